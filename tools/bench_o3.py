#!/usr/bin/env python
"""Tile-size / groups-in-flight sweep for the all-16-bit (amp O3) fused Adam on the config-5 list
(34k ragged tensors, 1.3 B elements, 14 B/param).  A/B between library builds: APEX_B200_SO=<path>.

    python tools/bench_o3.py [--total 1300000000] > gpurun_out/o3_sweep.json
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

from apex_example_b200 import _lib, ops  # noqa: E402
from apex_example_b200.tables import TensorTable  # noqa: E402
from workloads import many_small  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--total", type=int, default=1_300_000_000)
    ap.add_argument("--tiles", default="4096,8192,16384")
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--uniform", action="store_true", help="GPT-2-shaped sizes scaled to the same total (all multiples of 8)")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    sizes = many_small(total=args.total, seed=0)
    if args.uniform:
        from workloads import gpt2_shapes
        sizes = gpt2_shapes(2048, 24)
    n = sum(sizes)
    offs, acc = [], 0
    for s_ in sizes:
        acc += (-acc) % 16
        offs.append(acc)
        acc += s_
    flat = [torch.zeros(acc, dtype=torch.float16, device=dev) for _ in range(4)]
    flat[0].normal_(0, 1e-3)
    flat[1].normal_(0, 0.02)
    cols = [[f[o:o + s_] for o, s_ in zip(offs, sizes)] for f in flat]
    res = {"so": _lib.SO_PATH, "params": n, "tensors": len(sizes), "runs": {}}
    for tile in [int(t) for t in args.tiles.split(",")]:
        tbl = TensorTable(cols, tile_elems=tile)
        grp = [ops.AdamGroup(1e-4, 0.9, 0.999, 1e-8, 0.0, 1, 1, 1)]
        for _ in range(3):
            ops.fused_adam(tbl, torch.float16, torch.float16, None, grp)
        torch.cuda.synchronize()
        ts = []
        for _ in range(args.reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops.fused_adam(tbl, torch.float16, torch.float16, None, grp)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        res["runs"][str(tile)] = {"n_tiles": tbl.n_tiles, "ms_median": ts[len(ts) // 2], "ms_min": ts[0],
                                  "gbs_at_14B": 14 * n / (ts[len(ts) // 2] * 1e-3) / 1e9}
        ctbl = TensorTable([cols[0]], tile_elems=max(tile, 32768))
        flag = torch.zeros(1, dtype=torch.int32, device=dev)
        ts = []
        for _ in range(args.reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops.multi_tensor_check(ctbl, 0, torch.float16, flag)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        res["runs"][str(tile)]["check_ms_median"] = ts[len(ts) // 2]
        del tbl, ctbl
    print(json.dumps(res))


if __name__ == "__main__":
    main()
