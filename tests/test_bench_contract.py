"""bench.py's output contract: one JSON line with the keys the driver reads."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e", "gpu_launches"}


def _run(args, timeout):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True,
                       timeout=timeout, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, "exactly one JSON line expected"
    return json.loads(lines[0])


def test_reference_arm_prints_the_contract_line():
    d = _run(["--impl", "reference", "--steps", "1", "--warmup", "1"], timeout=600)
    assert BASE_KEYS <= set(d) and d["impl"] == "reference"
    assert d["metric"] == "fused optimizer step HBM GB/s" and d["unit"] == "GB/s" and d["higher_is_better"] is True
    assert d["vs_baseline"] is None and "workload" in d["config"] and "model" not in d["config"]
    assert d["warmup"] == 3 and d["steps"] == 1          # W >= 3 like the B200 arm, K as asked
    assert set(d["config"]) == {"workload", "l2", "step", "value_counts", "parallelism", "bucket_elems", "collective", "collective_note"}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


@pytest.mark.gpu
def test_b200_arm_prints_the_contract_line():
    d = _run(["--steps", "20", "--warmup", "3", "--no-train", "--no-config5"], timeout=900)
    assert BASE_KEYS | {"roofline", "clocks", "parity"} <= set(d) and "impl" not in d
    par = d["parity"]
    assert par["world"] == 1 and par["state_bit_exact"] is True and par["skip_agreed"] is True and par["all_ranks_ok"] is True
    assert d["n_gpus"] == 1 and d["steps"] == 20 and d["warmup"] == 3 and d["scaling"] == "weak" and d["data"] == "synthetic"
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert r["algorithmic_bytes_per_launch"] == 28 * 124_439_808 and 0.5 < r["frac"] < 1.3
    assert d["gpu_launches"] == 3 * d["steps"]                       # check + scaler update + fused step, nothing else
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] == 2 * 124_439_808 and e["d2h_bytes_per_step"] > 0 and 0 < e["value"] < d["value"]
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["value"] < d["value"]
    assert d["clocks"]["sm_max_mhz"] and not {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(d["clocks"]["reasons"])
