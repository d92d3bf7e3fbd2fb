#!/usr/bin/env python
"""bench.py — the mixed-precision step hot path on B200 (BASELINE.json metric
"fused optimizer step HBM GB/s; train samples/sec at 1/2/4/8 B200 (amp O2)").

    python bench.py --gpus 1 --steps 50 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference            # the CPU restatement on the host cores

Headline workload (BASELINE.json configs[1]): the 148-tensor /
124,439,808-element GPT-2-small parameter list at amp O2 — fp16 gradients carrying the loss scale,
fp32 master / m / v, fp16 model copy.  One *step* = everything between backward and the next forward:
  N = 1 : overflow check over the gradients -> loss-scaler update -> fused unscale+AdamW+copy
  N > 1 : gradient mean across ranks -> loss-scaler update -> fused unscale+AdamW+copy
          (weak scaling: every rank owns a full replica).  The mean runs over NVLink peer memory by
          default (--collective peer: one reduce-scatter+mean+check kernel, then the optimizer
          kernel pulls each tile's reduced gradient from its owner); --collective nccl selects
          ncclAllReduce per 10M-element bucket + the epilogue kernel.
`value` = 28 B/param x params x ranks / step time, inputs resident in HBM.  Before anything is timed
the same buffers run a PARITY block: one step and one injected-overflow step are compared bit for bit
with oracle/ on every rank (key `parity`; a mismatch ends the run with a non-zero exit code).
`e2e` = the same step through the public API (apex.amp + apex.optimizers.FusedAdam, at N > 1 with
apex.parallel.Reducer for the gradient mean) with the step's gradients copied in from pinned host
memory and the scaler state read back, every step.  Further keys, one per BASELINE.json config:
`train` (GPT-2 medium O2 + FusedAdam + apex DDP samples/s, with torch's DDP as the incumbent),
`train_resnet50` (ResNet-50 O2 + FusedSGD + apex DDP images/s), `config5` (1.3 B-parameter
many-small-tensor list at O3 with injected inf/NaN on one random rank), `allreduce` (NCCL bus GB/s on
GPT-2 medium's 709.6 MB for several message sizes).
"""
from __future__ import annotations

import argparse
import json
import os
import random
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

BYTES_PER_PARAM = 28            # 2 (fp16 g) + 12 (p,m,v in) + 12 (p,m,v out) + 2 (fp16 copy)
LOSS_SCALE = 65536.0
HYPER = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01)
BUCKET_ELEMS = 10_000_000       # apex DistributedDataParallel message_size default
METRIC = "fused optimizer step HBM GB/s"


def workload_sizes(name):
    from workloads import GPT2_MEDIUM, GPT2_SMALL
    if name == "gpt2_small":
        return GPT2_SMALL, "GPT-2-small-shaped list: 148 tensors, 124,439,808 params, amp O2 (fp16 grad, fp32 master/m/v, fp16 copy)"
    if name == "gpt2_medium":
        return GPT2_MEDIUM, "GPT-2-medium-shaped list: 292 tensors, 354,823,168 params, amp O2"
    raise SystemExit(f"unknown workload {name}")


def config_dict(desc, world, collective="peer", note=None, step_mode="spec"):
    """The `config` object both arms print (the driver compares them)."""
    spec = step_mode == "spec" and (world == 1 or collective == "peer")
    if step_mode == "sharded" and world == 1:
        spec = True
    if world == 1:
        coll = "none"
    elif collective == "peer" and step_mode == "sharded":
        coll = ("peer-memory: every rank reads all ranks' gradients for ITS shard and writes its shard's new fp16 weights "
                "into every rank's model buffer, inside the update kernel (ZeRO-1 style sharded optimizer state)")
    elif collective == "peer" and spec:
        coll = "peer-memory reduce-scatter fused with the speculative update in one kernel (chunk-wise progress flags)"
    elif collective == "peer":
        coll = "peer-memory reduce-scatter + pull-in-optimizer"
    else:
        coll = "nccl all-reduce"
    return {"workload": desc, "l2": "inputs larger than L2 (3.5 GB touched per step vs 126 MB L2)",
            "step": ("sharded: reduce-scatter + speculative AdamW on the own 1/N shard + all-gather of the fp16 weights in ONE "
                     "kernel per rank -> flag exchange -> scaler update + commit -> repair-if-skipped (every rank ends the step "
                     "with the whole updated fp16 model; fp32 master/m/v exist once per job, not once per rank)")
            if (step_mode == "sharded" and world > 1 and collective == "peer") else
                    ("speculative single pass (check + unscale + AdamW + fp16 copy into double-buffered state) -> "
                     "scaler update + commit -> repair-if-skipped") if spec else
                    "check|allreduce+epilogue -> scaler update -> fused unscale+AdamW+fp16 copy",
            "value_counts": "28 B/param x params x ranks / step time: the bytes a replicated-state step of this N-rank job moves "
                            "(the N = 1 unit); with the sharded step fewer bytes really move — see roofline.algorithmic_bytes_per_launch "
                            "and roofline.nvlink" if (step_mode == "sharded" and world > 1 and collective == "peer") else
                            "28 B/param x params x ranks / step time",
            "parallelism": f"dp{world}", "bucket_elems": BUCKET_ELEMS, "collective": coll, "collective_note": note}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.idx = device_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "100"], stdout=f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        try:
            for line in open(self.path):
                c = [x.strip() for x in line.split(",")]
                if len(c) < 9:
                    continue
                try:
                    sm.append(float(c[1])); mx.append(float(c[2]))
                except ValueError:
                    continue
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


def host_list(sizes, seed=0):
    """Seeded host tensors of the workload (also what the CPU baseline steps on)."""
    g = torch.Generator().manual_seed(seed)
    copy = [(torch.randn(n, generator=g) * 0.02).half() for n in sizes]
    p = [c.float() for c in copy]
    m = [torch.zeros(n) for n in sizes]
    v = [torch.zeros(n) for n in sizes]
    grad = [(torch.randn(n, generator=g) * 1e-3 * LOSS_SCALE).half() for n in sizes]
    return grad, p, m, v, copy


# ---------------------------------------------------------------------------------------------------
# reference arm: the CPU restatement (oracle/) on the host cores
# ---------------------------------------------------------------------------------------------------
def run_cpu(sizes, steps, warmup):
    import oracle
    torch.set_num_threads(os.cpu_count() or 1)
    grad, p, m, v, copy = host_list(sizes)
    times = []
    for s in range(warmup + steps):
        t0 = time.perf_counter()
        oracle.adam_list(grad, p, m, v, copy, inv_scale=1.0 / LOSS_SCALE, step=s + 1, adam_w_mode=True,
                         check_overflow=True, **HYPER)
        if s >= warmup:
            times.append(time.perf_counter() - t0)
    return times


def omp_threads():
    try:
        import ctypes
        return int(ctypes.CDLL("libgomp.so.1").omp_get_max_threads())
    except Exception:
        return os.cpu_count() or 1


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def reference_arm(args, sizes, desc):
    """The reference's hot path lives in NVIDIA/apex (CUDA only, not vendored, not installable offline),
    so this arm times its CPU restatement, oracle/amp_step_oracle.c, with OpenMP on every host core, on the
    same workload, step count and warm-up as the B200 arm (each step is the whole 124 M-parameter list:
    ~40 ms, so even the default 1000 steps end within a minute)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = sum(sizes)
    steps = max(1, min(args.steps, 2000))
    times = run_cpu(sizes, steps, args.warmup)
    t = sum(times) / len(times)
    gbs = BYTES_PER_PARAM * n / t / 1e9
    cores = omp_threads()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(desc, world, args.collective, None, args.step_mode),
        "reference_note": "reference's hot path lives in NVIDIA/apex (CUDA only, not installed); this arm times its CPU "
                          "restatement oracle/amp_step_oracle.c on the host cores (one replica, whatever N is)",
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "port", "cpu": cpu_model(),
                         "sample": f"full list, {steps} steps of check + unscale+AdamW+copy (OpenMP, {cores} threads)"},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------
# parity block: the timed buffers, one clean step and one overflow step, against oracle/
# ---------------------------------------------------------------------------------------------------
def _ulp_fp16(a, b):
    """max distance in fp16 ulps between two finite fp16 tensors (sign-magnitude to ordered ints)."""
    def key(t):
        i = t.view(torch.int16).to(torch.int32)
        return torch.where(i < 0, -(i & 0x7fff), i)
    return int((key(a) - key(b)).abs().max()) if a.numel() else 0


def parity_block(ctx, sabotage=False):
    """Runs on every rank, untimed, before the warm-up.  Clean step: the gradient mean this rank owns (peer
    path: its `reduced` shard; NCCL path: the whole all-reduced buffer) and p / m / v / fp16 copy after the
    fused step are compared with oracle.peer_reduce + oracle.adam_list computed from EVERY rank's inputs.
    Overflow step: one inf on the LAST rank only -> every rank must skip (found_inf = 1, state untouched,
    loss scale halved; SURVEY.md §3.4, reference skip path test_apex_distributed_spawn.py:160-164)."""
    import torch.distributed as dist
    import oracle
    from apex_example_b200 import ops
    world, rank, dev, n, sizes = ctx["world"], ctx["rank"], ctx["dev"], ctx["n"], ctx["sizes"]
    gflat, state, reducer = ctx["gflat"], ctx["state"], ctx["reducer"]
    cols = ctx["cols"]()                                # p, m, v, copy as current now (device lists)
    res = {"world": world, "path": ctx["path"], "params": n}

    gen = torch.Generator(device=dev).manual_seed(4242 + rank)
    gflat.copy_((torch.randn(n, device=dev, generator=gen) * 1e-3 * LOSS_SCALE).half())
    if world > 1:
        parts = [torch.empty_like(gflat) for _ in range(world)]
        dist.all_gather(parts, gflat)
        host_all = [t.cpu() for t in parts]
        del parts
    else:
        host_all = [gflat.cpu()]
    before = [[t.cpu() for t in col] for col in cols]
    torch.cuda.synchronize()

    factor = 1.0 / world
    ctx["step"](factor_override=(2.0 * factor if sabotage else None))
    torch.cuda.synchronize()
    found = int(state.cpu()[ops.IDX_FOUND_INF])

    # oracle: mean in rank order (fp32 accumulate, round to fp16, x 1/world, round), then the fused step
    want_mean = torch.zeros(n, dtype=torch.float16)
    wflag = oracle.peer_reduce(host_all, want_mean, 0, n, factor) if world > 1 else oracle.check(host_all[0])
    if world == 1:
        want_mean = host_all[0]
    if ctx["path"] == "peer-sharded":
        # the kernel leaves the mean of this rank's shard in the shard of its own gradient buffer
        b0, b1 = ctx["bounds"][rank], ctx["bounds"][rank + 1]
        got = gflat[b0:b1].cpu()
        exact = bool(torch.equal(got.view(torch.int16), want_mean[b0:b1].view(torch.int16)))
        res["mean_bit_exact"] = exact
        step_grad = want_mean
        if ctx.get("in_switch_sum"):
            # summed inside the switch: neither the addend order nor the precision of its partial sums is specified
            # (measured: up to 4 fp16 ulps from the rank-ordered fp32 sum at 4 ranks), so the mean gets the same bound as
            # the NCCL arm (<= world ulps) and the update is then checked bit for bit on the mean the switch produced
            res["mean_max_ulp_vs_oracle"] = 0 if exact else _ulp_fp16(got, want_mean[b0:b1])
            res["mean_ok"] = res["mean_max_ulp_vs_oracle"] <= world
            step_grad = want_mean.clone()
            step_grad[b0:b1] = got
    elif ctx["path"] in ("peer", "peer-spec"):
        b0, b1 = ctx["bounds"][rank], ctx["bounds"][rank + 1]
        got = ctx["pbufs"].local["reduced"][b0:b1].cpu()
        res["mean_bit_exact"] = bool(torch.equal(got.view(torch.int16), want_mean[b0:b1].view(torch.int16)))
        step_grad = want_mean
    elif world > 1:
        got = gflat.cpu()
        exact = bool(torch.equal(got.view(torch.int16), want_mean.view(torch.int16)))
        res["mean_bit_exact"] = exact
        res["mean_max_ulp_vs_oracle"] = 0 if exact else _ulp_fp16(got, want_mean)
        # NCCL does not specify its summation order: the mean must agree with the rank-ordered fp32 sum within
        # a few fp16 ulps and be IDENTICAL on every rank; the optimizer is then checked on the mean NCCL produced
        digest = torch.tensor([int(got.view(torch.int16).to(torch.int64).sum())], device=dev)
        lo, hi = digest.clone(), digest.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        res["mean_identical_on_all_ranks"] = bool(int(lo) == int(hi))
        res["mean_ok"] = res["mean_identical_on_all_ranks"] and res["mean_max_ulp_vs_oracle"] <= world
        step_grad = got
    else:
        res["mean_bit_exact"] = True
        step_grad = want_mean
    assert wflag == 0 and found == 0, "parity block: finite gradients flagged as overflow"
    if ctx.get("segments") is not None:
        hg = [step_grad[f:f + c].clone() for f, c in ctx["segments"]]       # this rank's shard only
    else:
        hg = list(torch.split(step_grad, sizes))
    hp, hm, hv, hc = before[:4]
    oracle.adam_list(hg, hp, hm, hv, hc, inv_scale=1.0 / LOSS_SCALE, step=ctx["step_no"][0], adam_w_mode=True,
                     check_overflow=False, **HYPER)
    ok = True
    cols = ctx["cols"]()
    for col, want in zip(cols, (hp, hm, hv, hc)):
        for d, h in zip(col, want):
            it = torch.int32 if h.element_size() == 4 else torch.int16
            if not torch.equal(d.cpu().view(it), h.view(it)):
                ok = False
                break
    res["state_bit_exact"] = ok
    if ctx["path"] == "peer-sharded":
        # every rank checked its own shard (state and weights) against the oracle; the whole model buffer must be the
        # same on every rank, so every rank holds the oracle's weights everywhere
        digest = torch.tensor([int(cols[4][0].view(torch.int16).to(torch.int64).sum())], device=dev)
        lo, hi = digest.clone(), digest.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        res["model_identical_on_all_ranks"] = bool(int(lo) == int(hi))
        res["state_bit_exact"] = ok and res["model_identical_on_all_ranks"]
    del host_all, before, hg

    # overflow on the last rank only
    snap = [[t.clone() for t in col] for col in cols]
    if rank == world - 1:
        gflat[n - 7] = float("inf")
    ctx["step"]()
    torch.cuda.synchronize()
    st = state.cpu()
    skipped = int(st[ops.IDX_FOUND_INF]) == 1
    cols = ctx["cols"]()
    untouched = all(torch.equal(a, b) for ca, cb in zip(cols, snap) for a, b in zip(ca, cb))
    halved = float(st.view(torch.float32)[ops.IDX_LOSS_SCALE]) == LOSS_SCALE / 2
    res["skip_agreed"] = bool(skipped and untouched and halved)
    del snap
    mean_good = res["mean_ok"] if (ctx["path"] == "nccl" or ctx.get("in_switch_sum")) else res["mean_bit_exact"]
    if ctx["path"] == "peer-spec":
        res["overflow_seen_by"] = "every rank's own update role (it inspects every reduced gradient it pulls)"
    good = res["state_bit_exact"] and res["skip_agreed"] and mean_good
    if world > 1:
        t = torch.tensor([1 if good else 0], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        res["all_ranks_ok"] = bool(int(t.item()))
        if "mean_max_ulp_vs_oracle" in res:                                 # MAX over ranks
            t = torch.tensor([res["mean_max_ulp_vs_oracle"]], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res["mean_max_ulp_vs_oracle"] = int(t.item())
        for k in ("mean_bit_exact", "state_bit_exact", "skip_agreed"):      # AND over ranks
            t = torch.tensor([1 if res.get(k, True) else 0], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            res[k] = bool(int(t.item()))
    else:
        res["all_ranks_ok"] = bool(good)
    # back to a clean slate for the timed run
    ops.scaler_init(state, LOSS_SCALE)
    torch.cuda.synchronize()
    return res


# ---------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="gpt2_small", choices=["gpt2_small", "gpt2_medium"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", dest="parity", action="store_false", help="skip the untimed parity block")
    ap.add_argument("--parity-sabotage", action="store_true",
                    help="self-test of the parity block: run its clean step with a wrong averaging factor (must fail)")
    ap.add_argument("--no-train", dest="train", action="store_false",
                    help="skip the training measurements (GPT-2 medium under 'train', ResNet-50 under 'train_resnet50')")
    ap.add_argument("--no-resnet", dest="resnet", action="store_false")
    ap.add_argument("--no-config5", dest="config5", action="store_false",
                    help="skip BASELINE configs[4] (1.3 B-parameter many-small-tensor list at O3 with injected overflows)")
    ap.add_argument("--config5-total", type=int, default=1_300_000_000)
    ap.add_argument("--config5-steps", type=int, default=30)
    ap.add_argument("--no-incumbent", dest="incumbent", action="store_false",
                    help="N>1: skip the torch.nn.parallel.DistributedDataParallel comparison run")
    ap.add_argument("--train-batch", type=int, default=4, help="per-GPU micro-batch of 1024-token sequences")
    ap.add_argument("--resnet-batch", type=int, default=256, help="per-GPU batch of 224x224 images")
    ap.add_argument("--ddp-backend", default="nccl", help="apex DDP bucket all-reduce for the training measurement: "
                    "nccl | peer | nccl,peer (measure both; the first is reported under 'train')")
    ap.add_argument("--message-size", type=int, default=0, help="apex DDP message_size for the training runs (0 = apex default)")
    ap.add_argument("--threads", type=int, default=0, help="tuning: threads per CTA (256|512)")
    ap.add_argument("--unroll", type=int, default=0, help="tuning: 8-element groups in flight per thread (1|2)")
    ap.add_argument("--ctas-per-sm", type=int, default=0)
    ap.add_argument("--peer-barrier", default="signal", choices=["signal", "nccl"],
                    help="peer path: in-kernel signal-pad hand-shake (default) or two 4-byte NCCL all-reduces")
    ap.add_argument("--step-mode", default="auto", choices=["auto", "spec", "checked", "sharded"],
                    help="'spec' (N = 1): speculative single pass — the overflow check rides on the update, p/m/v are "
                         "double-buffered (+12 B/param resident), commit flips a device-side parity; 'checked': "
                         "check pass -> scaler update -> update in place")
    ap.add_argument("--multicast", default="auto", choices=["auto", "none", "stores", "reduce", "both"],
                    help="sharded mode: NVSwitch multicast for the weight stores (multimem.st) and / or the gradient sum "
                         "(multimem.ld_reduce: in-switch, addend order unspecified)")
    ap.add_argument("--reducers", type=int, default=0, help="spec mode, N>1: CTAs of the fused kernel that take the reduce role")
    ap.add_argument("--chunk-elems", type=int, default=0, help="spec mode, N>1: elements per announced chunk of the fused kernel (0 = default)")
    ap.add_argument("--collective", default="peer", choices=["peer", "nccl"],
                    help="N>1: 'peer' = reduce-scatter kernel + optimizer pulling reduced shards over NVLink; "
                         "'nccl' = ncclAllReduce per bucket + epilogue kernel")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    n_ranks = int(os.environ.get("WORLD_SIZE", "1"))
    if args.step_mode == "auto":
        # N = 1: speculative single pass.  N > 1 over peer memory: the sharded step (fastest at every N measured,
        # profiles/r02_summary.md); `--step-mode spec` keeps the replicated-state fused step of §6a.
        args.step_mode = "sharded" if (n_ranks > 1 and args.collective == "peer") else "spec"
    if args.multicast == "auto":
        # the multicast paths pay from 4 ranks on (at 2 ranks the echo of the own stores costs more than it saves)
        args.multicast = "both" if n_ranks >= 4 else "none"
    sizes, desc = workload_sizes(args.workload)
    if args.impl == "reference":
        return reference_arm(args, sizes, desc)

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 arm has no CPU fallback (use --impl reference for the CPU restatement)")
    import torch.distributed as dist
    from apex_example_b200 import _lib
    if not os.path.exists(_lib.SO_PATH) and int(os.environ.get("LOCAL_RANK", "0")) == 0:
        _lib.build()                                 # normally built by __graft_entry__.build() and shipped in-tree
    from apex_example_b200 import ops
    from apex_example_b200.tables import TensorTable

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if args.threads or args.unroll or args.ctas_per_sm:
        ops.set_launch_config(args.ctas_per_sm, args.threads or 256, args.unroll or 2)

    n = sum(sizes)
    gen = torch.Generator(device=dev).manual_seed(0)
    # gradients live in flat buckets (what apex DDP's flatten produces; at N=1 just contiguous storage)
    use_peer = world > 1 and args.collective == "peer"
    peer_note = None
    if use_peer:
        # the peer-memory path needs every pair of GPUs to be peer-accessible (NVLink / NVSwitch); all ranks
        # must take the same decision, so the capability is agreed on with a MIN all-reduce
        ok = all(torch.cuda.can_device_access_peer(local, d) for d in range(torch.cuda.device_count()) if d != local)
        t = torch.tensor([1 if ok else 0], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        if int(t.item()) == 0:
            use_peer = False
            peer_note = "peer access not available between all GPUs: fell back to NCCL"
    reducer, bounds = None, None
    spec = args.step_mode == "spec" and (world == 1 or use_peer)
    sharded = args.step_mode == "sharded" and use_peer
    if args.step_mode == "sharded" and not sharded:
        spec = world == 1                          # one rank (or no peer access): the shard is the whole model
    if use_peer:
        from apex_example_b200.parallel.peer import PeerBuffers, PeerGradientReducer
        from apex_example_b200.tables import tile_aligned_shards
        offs, acc = [], 0
        for s_ in sizes:
            offs.append(acc)
            acc += s_
        pbufs = PeerBuffers(n, torch.float16, names=("grads", "model") if sharded else ("grads", "reduced"),
                            multicast={"none": (), "stores": ("model",), "reduce": ("grads",),
                                       "both": ("grads", "model")}[args.multicast] if sharded else ())
        bounds, first_tiles = tile_aligned_shards(offs, sizes, n, world, return_first_tiles=True)
        if not spec and not sharded:               # the fused steps bring their own hand-shake
            reducer = PeerGradientReducer(pbufs, bounds, first_tiles=first_tiles, barrier=args.peer_barrier)
        gflat = pbufs.local["grads"]
    else:
        gflat = torch.empty(n, dtype=torch.float16, device=dev)

    def bench_gradients():
        gflat.copy_((torch.randn(n, device=dev, generator=gen) * 1e-3 * LOSS_SCALE).half())

    grads = list(torch.split(gflat, sizes))
    shard_stepper, segs = None, None
    if sharded:
        # ZeRO-1 style: the fp16 model lives in a peer-addressable flat buffer (every rank writes its shard's new
        # weights into every rank's buffer); fp32 master / m / v exist for this rank's shard only
        from apex_example_b200.parallel.peer import PeerShardedStep
        wflat = pbufs.local["model"]
        wflat.copy_((torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(77)) * 0.02).half())
        copy = list(torch.split(wflat, sizes))
        shard_stepper = PeerShardedStep(pbufs, offs, sizes, use_multicast=args.multicast)
        shard_multicast, shard_mc_note = shard_stepper.multicast_stores, pbufs.multicast_note
        shard_mc_reduce = shard_stepper.multicast_reduce
        bounds = shard_stepper.bounds
        segs = [(flat, cnt) for _, _, cnt, flat in shard_stepper.segments]
        parity_word = shard_stepper.parity
        table = shard_stepper.table
    else:
        copy = [(torch.randn(s, device=dev, generator=gen) * 0.02).half() for s in sizes]
    if sharded:
        pass
    elif spec:
        # double-buffered state: every p / m / v tensor is followed by its shadow copy (b200_spec_shadow_delta)
        pb = [ops.spec_state_like(c.float(), fill=c.float()) for c in copy]
        mb = [ops.spec_state_like(c.float()) for c in copy]
        vb = [ops.spec_state_like(c.float()) for c in copy]
        p, m, v = [x for x, _ in pb], [x for x, _ in mb], [x for x, _ in vb]
        parity_word = torch.zeros(1, dtype=torch.int32, device=dev)
    else:
        p = [c.float() for c in copy]
        m = [torch.zeros(s, device=dev) for s in sizes]
        v = [torch.zeros(s, device=dev) for s in sizes]
    if not sharded:
        table = TensorTable([grads, p, m, v, copy])
    stepper = None
    if spec and use_peer:
        from apex_example_b200.parallel.peer import PeerSpecStep
        stepper = PeerSpecStep(pbufs, bounds, table, chunk_elems=args.chunk_elems or None, n_reducers=args.reducers or None)

    def current_cols():
        """p, m, v, copy as they are current right now (synchronises in spec mode to read the parity)."""
        if sharded:
            par = int(parity_word.item())
            return [shard_stepper.current("p", par), shard_stepper.current("m", par), shard_stepper.current("v", par),
                    [wflat[f:f + c] for f, c in segs], [wflat]]
        if not spec:
            return [p, m, v, copy]
        par = int(parity_word.item())
        return [[ops.spec_current(x, b, par) for x, b in pb], [ops.spec_current(x, b, par) for x, b in mb],
                [ops.spec_current(x, b, par) for x, b in vb], copy]
    check_table = TensorTable([grads])              # one column: the scheduler picks 32768-element tiles
    buckets = []
    off = 0
    while off < n:
        end = min(n, off + BUCKET_ELEMS)
        buckets.append(gflat[off:end])
        off = end
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device=dev)
    ops.scaler_init(state, LOSS_SCALE)
    fstate = state.view(torch.float32)
    inv_ref, found_ref, acc_ref = fstate[ops.IDX_INV_SCALE:], state[ops.IDX_FOUND_INF:], state[ops.IDX_OVERFLOW_ACC:]
    inv_loss_ref = fstate[ops.IDX_INV_LOSS_SCALE:]
    step_no = [0]

    def step(ev=None, factor_override=None):
        step_no[0] += 1
        if sharded:
            # ONE kernel per rank: gradients of all ranks for the own shard -> mean -> speculative AdamW -> new fp16
            # weights stored into every rank's model buffer; then flag exchange, commit, repair (exits at once)
            if ev is not None:
                ev[2].record()
                ev[3].record()
                ev[0].record()
            grp = [ops.AdamGroup(HYPER["lr"], HYPER["beta1"], HYPER["beta2"], HYPER["eps"], HYPER["weight_decay"],
                                 step_no[0], 1, 1)]
            if factor_override is not None:
                keep = shard_stepper.fan.factor
                shard_stepper.fan.factor = factor_override
            shard_stepper.step(grp, state, ev_after_kernel=ev[1] if ev is not None else None)
            if factor_override is not None:
                shard_stepper.fan.factor = keep
            return
        if spec:
            # ONE pass: check + unscale + AdamW + copy into the shadow state; commit = scaler update + parity flip;
            # repair exits at once unless the step was skipped
            if ev is not None:
                ev[2].record()
                ev[3].record()
                ev[0].record()
            grp = [ops.AdamGroup(HYPER["lr"], HYPER["beta1"], HYPER["beta2"], HYPER["eps"], HYPER["weight_decay"],
                                 step_no[0], 1, 1)]
            if stepper is not None:
                # N > 1: ONE kernel does the gradient mean (reducer CTAs, chunk-wise progress) AND the update
                if factor_override is not None:
                    keep, stepper.factor = stepper.factor, factor_override
                stepper.step(torch.float16, torch.float32, torch.float16, grp, acc_ref, parity_word, dev_inv_scale=inv_loss_ref)
                if factor_override is not None:
                    stepper.factor = keep
            else:
                ops.fused_adam_spec(table, torch.float16, torch.float32, torch.float16, grp, acc_ref, parity_word,
                                    dev_inv_scale=inv_loss_ref)
            if ev is not None:
                ev[1].record()
            ops.spec_commit(state, True, 2000, 2.0, 0.0, 2.0 ** 24, parity_word)
            ops.spec_repair(table, torch.float32, torch.float16, found_ref, parity_word)
            return
        if use_peer:
            # ONE kernel: (hand-shake) reduce-scatter + mean + overflow check over NVLink -> flag OR
            if ev is not None:
                ev[2].record()
            if factor_override is not None:
                keep, reducer.factor = reducer.factor, factor_override
            reducer.reduce(acc_ref)
            if factor_override is not None:
                reducer.factor = keep
            if ev is not None:
                ev[3].record()
        elif world > 1:
            # in-place SUM then x 1/N: after the first step every rank holds the mean, which the
            # following steps reproduce exactly (N equal addends / N), so values stay finite
            if ev is not None:
                ev[2].record()
            for b in buckets:
                dist.all_reduce(b)
                ops.bucket_epilogue(b, (1.0 / world) if factor_override is None else factor_override, acc_ref)
            if ev is not None:
                ev[3].record()
        else:
            if ev is not None:
                ev[2].record()
            ops.multi_tensor_check(check_table, 0, torch.float16, acc_ref)
            if ev is not None:
                ev[3].record()
        ops.scaler_update(state, True, 2000, 2.0, 0.0, 2.0 ** 24)
        if ev is not None:
            ev[0].record()
        grp = [ops.AdamGroup(HYPER["lr"], HYPER["beta1"], HYPER["beta2"], HYPER["eps"], HYPER["weight_decay"],
                             step_no[0], 1, 1)]
        ops.fused_adam(table, torch.float16, torch.float32, torch.float16, grp, dev_inv_scale=inv_ref,
                       dev_found_inf=found_ref, pull=reducer.pull if use_peer else None)
        if ev is not None:
            ev[1].record()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- parity (untimed) ----
    parity = None
    if args.parity:
        ctx = dict(world=world, rank=rank, dev=dev, n=n, sizes=sizes, gflat=gflat, state=state, reducer=reducer,
                   pbufs=pbufs if use_peer else None,
                   segments=segs, in_switch_sum=bool(sharded and shard_mc_reduce),
                   cols=current_cols, path=("peer-sharded" if sharded else "peer-spec" if spec else "peer") if use_peer else ("nccl" if world > 1 else ("single-spec" if spec else "single")),
                   bounds=bounds, step=step, step_no=step_no)
        parity = parity_block(ctx, sabotage=args.parity_sabotage)
        step_no[0] -= 1                               # the skipped step did not count
        if not parity["all_ranks_ok"]:
            if rank == 0:
                print(json.dumps({"metric": METRIC, "error": "parity block failed", "parity": parity}), flush=True)
            if use_peer:
                del reducer, stepper, table
                pbufs.close()
            if world > 1:
                dist.destroy_process_group()
            raise SystemExit(3)
    bench_gradients()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ops.launch_count()
    evs = [tuple(torch.cuda.Event(enable_timing=True) for _ in range(4)) for _ in range(args.steps)]
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_begin.record()
    for k in range(args.steps):
        step(evs[k])
    t_end.record()
    barrier()
    launches = ops.launch_count() - l0
    # nvidia-smi samples every 100 ms; a short timed region (K x 0.6 ms) would give it nothing to see,
    # so the SAME step keeps running (untimed) until the sampler has had ~0.7 s under this load
    region_s = t_begin.elapsed_time(t_end) * 1e-3
    extra_steps = 0
    if region_s < 0.7:
        extra_steps = int((0.7 - region_s) / max(region_s / args.steps, 1e-4)) + 1
    if world > 1:                                   # every rank must run the same number (collectives inside)
        t = torch.tensor([extra_steps], device=dev, dtype=torch.int64)
        dist.broadcast(t, 0)
        extra_steps = int(t.item())
    if extra_steps:
        for _ in range(extra_steps):
            step()
        barrier()
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None:
        clocks["sampled_over"] = f"timed region + {extra_steps} identical untimed steps"
    total_ms = t_begin.elapsed_time(t_end)
    adam_ms = [e[0].elapsed_time(e[1]) for e in evs]
    check_ms = [e[2].elapsed_time(e[3]) for e in evs]
    if world > 1:
        t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = BYTES_PER_PARAM * n * world / (ms_per_step * 1e-3) / 1e9
    skipped = int(state.cpu()[ops.IDX_FOUND_INF])
    assert skipped == 0, "synthetic gradients are finite: the step must not have been skipped"
    if use_peer:
        pbufs.check_signals()

    # ---- roofline of the dominant kernel (fused step) ----
    peak, peak_src = peaks()
    adam_avg_ms = sum(adam_ms) / len(adam_ms)
    achieved = BYTES_PER_PARAM * n / (adam_avg_ms * 1e-3) / 1e9
    roof = {"bound": "hbm", "kernel": "tile_kernel<AdamOp<f16,f32,f16>> (b200_fused_adam)", "achieved": achieved,
            "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
            "kernel_ms_avg": adam_avg_ms, "kernel_ms_min": min(adam_ms), "kernel_share_of_step": adam_avg_ms / ms_per_step,
            "algorithmic_bytes_per_launch": BYTES_PER_PARAM * n, "traffic": None}
    cm = sum(check_ms) / len(check_ms)
    if sharded:
        per = 2 * n // world                                # bytes of one shard of fp16 gradients / weights
        # NVLink bytes per GPU: gradients in + weights in; gradients out + weights out
        g_in = per if shard_mc_reduce else per * (world - 1)        # in-switch sum: only the sum comes in
        g_out = per * world if shard_mc_reduce else per * (world - 1)   # ... but every GPU's whole buffer feeds the switch
        w_out = per if shard_multicast else per * (world - 1)       # multicast store: leaves once
        w_in = per * world if shard_multicast else per * (world - 1)    # ... and the sender receives its own echo
        link_in, link_out = g_in + w_in, g_out + w_out
        hbm = BYTES_PER_PARAM * n // world + g_out + w_in    # own shard + gradients served + weights received
        t = adam_avg_ms * 1e-3
        roof.update({"kernel": "tile_kernel<ShardStepOp<f16,f32,f16,W>> (b200_peer_shard_step: all ranks' gradients of the own "
                               "shard -> mean -> check + unscale + AdamW (speculative) -> fp16 weights stored into "
                               "every rank's model buffer; one launch, no inter-tile dependency)",
                     "achieved": hbm / t / 1e9, "frac": hbm / t / 1e9 / peak, "algorithmic_bytes_per_launch": hbm,
                     "limiter": "NVLink, not HBM: per GPU the kernel receives and sends the bytes in `nvlink` while it touches "
                                "only 1/N of the optimizer state; NCCL's best all-reduce bus bandwidth on this box is in `allreduce`",
                     "nvlink": {"bytes_in_per_gpu": link_in, "bytes_out_per_gpu": link_out, "inbound_gbs": link_in / t / 1e9,
                                "outbound_gbs": link_out / t / 1e9, "nvlink_peak_gbs_per_dir": 900.0,
                                "frac_of_nvlink_peak": max(link_in, link_out) / t / 1e9 / 900.0},
                     "weights_pushed_by": "multimem.st through the NVSwitch multicast address (one store leaves the GPU per 16 bytes)"
                     if shard_multicast else "one store per rank", "multicast_note": shard_mc_note,
                     "gradients_summed_by": "multimem.ld_reduce inside the NVSwitch (fp32 accumulation, one rounding, addend order "
                     "unspecified)" if shard_mc_reduce else "one load per rank, fp32 sum in rank order (bit-exact vs the oracle)",
                     "replicated_equivalent_gbs_per_gpu": BYTES_PER_PARAM * n / t / 1e9,
                     "value_note": "`value` counts 28 B/param x N: the bytes a replicated-state step of the same N-rank job "
                                   "moves (the unit the N = 1 line is quoted in); this step moves fewer (see hbm / nvlink)"})
    elif spec and world > 1:
        roof["kernel"] = ("peer_spec_step_kernel<f16,f32,f16> (b200_peer_spec_step: a few reducer CTAs reduce-scatter this rank's shard "
                          "over NVLink through bulk-copy staged shared memory while all other CTAs run check + unscale + AdamW + "
                          "copy on tiles whose chunk has been announced; one launch)")
        roof["nvlink"] = {"bytes_in_per_gpu": 2 * n * 2 * (world - 1) // world, "note": "(N-1)/N of the gradient bytes twice: "
                          "peers' gradients into the reduce role, reduced shards into the update role",
                          "inbound_gbs": 2 * n * 2 * (world - 1) / world / (adam_avg_ms * 1e-3) / 1e9, "nvlink_peak_gbs_per_dir": 900.0}
    elif spec:
        roof["kernel"] = "tile_kernel<AdamSpecOp<f16,f32,f16>> (b200_fused_adam_spec: check + unscale + AdamW + copy in one pass)"
        roof["other_kernels"] = {"note": "no separate overflow-check pass: spec_commit (1 thread) and spec_repair (exits at once) follow"}
    elif world == 1:
        roof["other_kernels"] = {"multi_tensor_check": {"ms_avg": cm, "achieved_gbs": 2 * n / (cm * 1e-3) / 1e9,
                                                        "algorithmic_bytes_per_launch": 2 * n}}
    elif not spec:
        # gradient mean: every rank receives (N-1)/N of the buffer over NVLink in the reduce phase and
        # again while the optimizer pulls the reduced shards; bus bandwidth as NCCL defines it
        nbytes = n * 2
        roof["gradient_mean"] = {"path": "peer reduce-scatter kernel (in-kernel hand-shake) + flag wait" if use_peer
                                 else "ncclAllReduce per 10M-element bucket + epilogue kernel",
                                 "ms_avg": cm, "bytes": nbytes,
                                 "busbw_gbs": nbytes / (cm * 1e-3) * 2 * (world - 1) / world / 1e9 if not use_peer
                                 else nbytes / (cm * 1e-3) * (world - 1) / world / 1e9,
                                 "busbw_note": "nccl: bytes/t*2(n-1)/n; peer: reduce phase only, bytes/t*(n-1)/n "
                                               "(the gather half is inside the optimizer kernel)",
                                 "nvlink_peak_gbs_per_dir": 900.0}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            roof["traffic"] = json.load(open(tpath)).get(args.workload)
        except Exception:
            pass

    extra = {}
    del table, check_table, grads, copy, buckets
    if sharded:
        del shard_stepper, wflat
    else:
        del p, m, v
    if spec:
        del pb, mb, vb
    if use_peer:
        del reducer, gflat, stepper
        pbufs.close()
    else:
        del gflat
    torch.cuda.empty_cache()

    # ---- NCCL all-reduce bus bandwidth on config 4's gradient bytes (N > 1) ----
    if world > 1:
        extra["allreduce"] = run_allreduce_sweep(dev, world, barrier)

    # ---- e2e through the public API with host buffers ----
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(sizes, dev, world, max(3, min(args.steps, 20)), args.warmup, barrier)

    # ---- the second half of BASELINE.json's metric: train samples/s ----
    train = None
    msg = args.message_size or None
    if args.train:
        try:
            backends = [b for b in args.ddp_backend.split(",") if b] if world > 1 else ["nccl"]
            runs = [run_train(dev, world, rank, barrier, args.train_batch, ddp_backend=b, message_size=msg) for b in backends]
            train = runs[0]
            if len(runs) > 1:
                train["other_ddp_backends"] = runs[1:]
            if world > 1 and args.incumbent:
                train["incumbent_torch_ddp"] = run_train(dev, world, rank, barrier, args.train_batch, ddp_backend="torch")
        except Exception as exc:                    # reported, never fatal for the headline line
            train = {"error": f"{type(exc).__name__}: {exc}"[:300]}
        if args.resnet:
            try:
                extra["train_resnet50"] = run_train_resnet50(dev, world, rank, barrier, args.resnet_batch, message_size=msg)
            except Exception as exc:
                extra["train_resnet50"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
        torch.cuda.empty_cache()

    # ---- BASELINE configs[4]: many small tensors, O3 dynamic, inf/NaN on one random rank ----
    if args.config5:
        try:
            extra["config5"] = run_config5(dev, world, rank, barrier, args.config5_total, args.config5_steps)
        except Exception as exc:
            extra["config5"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
        torch.cuda.empty_cache()

    # ---- CPU baseline (rank 0, N = 1 only) ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        times = run_cpu(sizes, 3, 1)
        tc = sum(times) / len(times)
        cores = omp_threads()
        cpu = {"value": BYTES_PER_PARAM * n / tc / 1e9, "unit": "GB/s", "cores": cores, "kind": "port",
               "cpu": cpu_model(), "ms_per_step": tc * 1e3,
               "sample": f"full list, 3 steps of check + unscale+AdamW+copy via oracle/amp_step_oracle.c (OpenMP, {cores} threads)"}

    bad_trace = isinstance(extra.get("config5"), dict) and extra["config5"].get("scale_trace_equals_oracle") is False
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(desc, world, "peer" if use_peer else "nccl", peer_note, args.step_mode),
            "parity": parity, "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clocks, "train": train,
        }
        line.update(extra)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if bad_trace:
        raise SystemExit(4)


def run_allreduce_sweep(dev, world, barrier):
    """ncclAllReduce(SUM, fp16) in place on GPT-2 medium's gradient bytes (354,823,168 elements = 709.6 MB,
    BASELINE configs[3]) cut into messages of several sizes, back to back on one stream; bus GB/s as NCCL
    defines it (bytes/t x 2(n-1)/n) against NVLink 5's 900 GB/s per direction."""
    import torch.distributed as dist
    n = 354_823_168
    flat = torch.zeros(n, dtype=torch.float16, device=dev)
    out = {"bytes_per_step": n * 2, "nvlink_peak_gbs_per_dir": 900.0, "by_message_elems": {}}
    for msg in (10_000_000, 32_000_000, 64_000_000, n):
        chunks = [flat[o:min(n, o + msg)] for o in range(0, n, msg)]
        for c in chunks:                                # warm-up: NCCL sets its channels up per size class
            dist.all_reduce(c)
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        a0.record()
        for _ in range(reps):
            for c in chunks:
                dist.all_reduce(c)
        a1.record()
        barrier()
        ms = a0.elapsed_time(a1) / reps
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        out["by_message_elems"][str(msg)] = {"messages": len(chunks), "ms": ms,
                                             "busbw_gbs": n * 2 / (ms * 1e-3) * 2 * (world - 1) / world / 1e9}
    best = max(out["by_message_elems"].items(), key=lambda kv: kv[1]["busbw_gbs"])
    apex_default = out["by_message_elems"][str(10_000_000)]
    out.update(ms=apex_default["ms"], busbw_gbs=apex_default["busbw_gbs"], buckets=apex_default["messages"],
               best_message_elems=int(best[0]), best_busbw_gbs=best[1]["busbw_gbs"])
    del flat
    return out


def _ddp_wrap(model, world, ddp_backend, message_size, local):
    if world <= 1:
        return model, None
    if ddp_backend == "torch":
        ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local])
        return ddp, None
    from apex.parallel import DistributedDataParallel
    kw = {"message_size": message_size} if message_size else {}
    ddp = DistributedDataParallel(model, allreduce_backend=ddp_backend, **kw)
    return ddp, ddp


def run_train(dev, world, rank, barrier, micro_batch=4, seq=1024, steps=8, warmup=3, ddp_backend="nccl", message_size=None):
    """GPT-2 medium (24 x 1024, 354.8M parameters, random init) on synthetic tokens through the
    drop-in API exactly as the reference drives it: amp.initialize(O2) -> apex DDP -> scale_loss /
    backward / FusedAdam.step.  The transformer's forward/backward is torch/transformers library
    code; what is under test is everything between backward and the next forward.
    ddp_backend="torch" swaps apex DDP for torch.nn.parallel.DistributedDataParallel (the incumbent; the
    reference's other branch, test_apex_distributed_spawn.py:124) with everything else unchanged."""
    import torch.distributed as dist
    from transformers import GPT2Config, GPT2LMHeadModel
    from apex import amp
    from apex.optimizers import FusedAdam
    from apex_example_b200 import ops

    amp._amp_state.reset()
    torch.manual_seed(0)
    cfg = GPT2Config(n_embd=1024, n_layer=24, n_head=16)
    try:
        cfg._attn_implementation = "sdpa"
    except Exception:
        pass
    model = GPT2LMHeadModel(cfg).to(dev)
    n_params = sum(p.numel() for p in model.parameters())
    opt = FusedAdam(model.parameters(), lr=1e-4, weight_decay=0.01)
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    model, ddp = _ddp_wrap(model, world, ddp_backend, message_size, dev.index)
    host_tokens = torch.randint(0, 50257, (micro_batch, seq), generator=torch.Generator().manual_seed(rank)).pin_memory()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]

    def one(e=None):
        tokens = host_tokens.to(dev, non_blocking=True)
        out = model(input_ids=tokens, labels=tokens)
        loss = out["loss"] if isinstance(out, dict) else out.loss
        opt.zero_grad()
        with amp.scale_loss(loss, opt) as scaled:
            scaled.backward()
            if e is not None:
                e[0].record()
        if e is not None:
            e[1].record()
        opt.step()
        if e is not None:
            e[2].record()
        return loss

    for _ in range(warmup):
        one()
    barrier()
    l0 = ops.launch_count()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for k in range(steps):
        loss = one(ev[k])
    t1.record()
    barrier()
    ms = t0.elapsed_time(t1) / steps
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    post_bwd = sum(e[0].elapsed_time(e[1]) for e in ev) / steps      # scale_loss exit: (DDP join) + check + scaler update
    step_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / steps       # FusedAdam.step: one launch per dtype set
    scaler = amp._amp_state.loss_scalers[0]
    wrap = {"nccl": "apex DistributedDataParallel (bucket all-reduce: nccl)", "peer": "apex DistributedDataParallel (bucket all-reduce: peer)",
            "torch": "torch.nn.parallel.DistributedDataParallel (incumbent)"}[ddp_backend] if world > 1 else "single GPU"
    res = {"model": "GPT-2 medium (GPT2Config(n_embd=1024, n_layer=24, n_head=16)), amp O2, FusedAdam, " + wrap,
           "params": n_params, "micro_batch_per_gpu": micro_batch, "seq_len": seq, "n_gpus": world,
           "samples_per_s": micro_batch * world / (ms * 1e-3), "tokens_per_s": micro_batch * seq * world / (ms * 1e-3),
           "ms_per_step": ms, "ms_scale_loss_exit": post_bwd, "ms_optimizer_step": step_ms,
           "optimizer_step_gbs": BYTES_PER_PARAM * n_params / (step_ms * 1e-3) / 1e9,
           "b200_kernel_launches_per_step": (ops.launch_count() - l0) / steps,
           "message_size": (message_size or 10_000_000) if ddp_backend != "torch" else "25 MiB buckets",
           "final_loss": float(loss.item()), "loss_scale": scaler.loss_scale(), "steps": steps}
    if ddp is not None:
        ddp.close()
    del model, opt, ddp
    torch.cuda.empty_cache()
    return res


def run_train_resnet50(dev, world, rank, barrier, batch=256, steps=8, warmup=4, message_size=None):
    """BASELINE configs[2]: torchvision resnet50() (161 tensors, 25,557,032 parameters; the 106 BatchNorm
    tensors stay fp32 at O2 -> two dtype launch sets), amp O2, FusedSGD(lr 0.1, momentum 0.9, wd 1e-4),
    apex DistributedDataParallel, synthetic 224x224 images (channels_last), dynamic loss scale."""
    import torch.distributed as dist
    import torchvision
    from apex import amp
    from apex.optimizers import FusedSGD
    from apex_example_b200 import ops

    amp._amp_state.reset()
    torch.manual_seed(0)
    model = torchvision.models.resnet50().to(dev).to(memory_format=torch.channels_last)
    n_params = sum(p.numel() for p in model.parameters())
    n_tensors = sum(1 for _ in model.parameters())
    opt = FusedSGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4, materialize_master_grads=False)
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    model, ddp = _ddp_wrap(model, world, "nccl", message_size, dev.index)
    g = torch.Generator().manual_seed(100 + rank)
    images = torch.randn(batch, 3, 224, 224, generator=g).pin_memory()
    labels = torch.randint(0, 1000, (batch,), generator=g).pin_memory()
    crit = torch.nn.CrossEntropyLoss().to(dev)
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
    torch.backends.cudnn.benchmark = True

    def one(e=None):
        x = images.to(dev, non_blocking=True).contiguous(memory_format=torch.channels_last)
        y = labels.to(dev, non_blocking=True)
        loss = crit(model(x), y)
        opt.zero_grad()
        with amp.scale_loss(loss, opt) as scaled:
            scaled.backward()
            if e is not None:
                e[0].record()
        if e is not None:
            e[1].record()
        opt.step()
        if e is not None:
            e[2].record()
        return loss

    for _ in range(warmup):
        one()
    barrier()
    l0 = ops.launch_count()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for k in range(steps):
        loss = one(ev[k])
    t1.record()
    barrier()
    ms = t0.elapsed_time(t1) / steps
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    step_us = sum(e[1].elapsed_time(e[2]) for e in ev) / steps * 1e3
    exit_us = sum(e[0].elapsed_time(e[1]) for e in ev) / steps * 1e3
    n_bn = sum(p.numel() for p in opt._amp_stash.all_fp32_params)
    algo = 20 * n_params       # momentum SGD at O2: fp16 params 2+4+4 in, 4+4+2 out; fp32 BatchNorm params 4+4+4 in, 4+4 out
    res = {"model": "torchvision resnet50(), amp O2, FusedSGD(lr=0.1, momentum=0.9, weight_decay=1e-4), "
                    + ("apex DistributedDataParallel (nccl)" if world > 1 else "single GPU") + ", channels_last",
           "params": n_params, "tensors": n_tensors, "fp32_batchnorm_params": n_bn, "batch_per_gpu": batch, "n_gpus": world,
           "images_per_s": batch * world / (ms * 1e-3), "ms_per_step": ms,
           "us_scale_loss_exit": exit_us, "us_optimizer_step": step_us,
           "optimizer_step_gbs": algo / (step_us * 1e-6) / 1e9, "optimizer_dtype_launch_sets": 2,
           "b200_kernel_launches_per_step": (ops.launch_count() - l0) / steps,
           "final_loss": float(loss.item()), "loss_scale": amp._amp_state.loss_scalers[0].loss_scale(), "steps": steps}
    if ddp is not None:
        ddp.close()
    del model, opt, ddp
    torch.cuda.empty_cache()
    return res


def run_config5(dev, world, rank, barrier, total=1_300_000_000, steps=30):
    """BASELINE configs[4] / SURVEY.md §8d config 5: a 1.3 B-parameter list of many small tensors (sizes
    log-uniform in [256, 262144], most not divisible by 8) at O3 with dynamic loss scaling — all-fp16 p, m,
    v, g — stepped by amp + FusedAdam.  The gradient mean across the N ranks goes through
    apex.parallel.Reducer (in place on the flat gradient arena: ncclAllReduce per message + the epilogue
    kernel, which also raises amp's overflow flag).  At steps drawn Bernoulli(0.1) (seed 1; plus the first
    two steps) ONE random rank gets an inf or NaN in one element (first / last / random) of one random
    tensor.  Pass = the skip / loss-scale trace equals the oracle's state machine on EVERY rank and sampled
    p / m / v are untouched on skipped steps."""
    import torch.distributed as dist
    import oracle
    from workloads import many_small
    from apex import amp
    from apex.optimizers import FusedAdam
    from apex.parallel import Reducer
    from apex_example_b200 import ops

    amp._amp_state.reset()
    sizes = many_small(total=total, seed=0)
    n = sum(sizes)
    offs, acc = [], 0
    for s_ in sizes:                                   # every tensor starts 32-byte aligned inside the gradient arena
        acc += (-acc) % 16
        offs.append(acc)
        acc += s_
    t0 = time.time()
    torch.manual_seed(0)
    flat_p = torch.randn(n, device=dev) * 0.02
    flat_g = torch.zeros(acc, dtype=torch.float16, device=dev)

    class Holder(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.ps = torch.nn.ParameterList([torch.nn.Parameter(t) for t in torch.split(flat_p, sizes)])

    model = Holder()
    opt = FusedAdam(model.parameters(), lr=1e-4, set_grad_none=False)
    model, opt = amp.initialize(model, opt, opt_level="O3", loss_scale="dynamic", verbosity=0)
    del flat_p
    params = list(model.parameters())
    assert all(p.dtype == torch.float16 for p in params)
    for p, o, s_ in zip(params, offs, sizes):
        p.grad = flat_g[o:o + s_]
    reducer = Reducer([p.grad for p in params], message_size=64_000_000).covers(params) if world > 1 else None
    setup_s = time.time() - t0
    rng = random.Random(1)                             # the same draws on every rank
    sc = oracle.LossScaler()
    scaler = amp._amp_state.loss_scalers[0]
    dummy = torch.zeros((), device=dev)
    sample = sorted(set([0, len(sizes) - 1] + [rng.randrange(len(sizes)) for _ in range(62)]))
    trace_ok, untouched_ok = True, True
    normal_ms, skip_ms, host_ms, reduce_ms = [], [], [], []
    gen = torch.Generator(device=dev).manual_seed(rank)
    for it in range(steps):
        flat_g.copy_((torch.randn(flat_g.numel(), device=dev, generator=gen) * 1e-3 * sc.loss_scale).half())
        inject = rng.random() < 0.1 or it in (0, 1)
        t_idx = rng.randrange(len(sizes))
        e_idx = rng.choice([0, sizes[t_idx] - 1, rng.randrange(sizes[t_idx])])
        who = rng.randrange(world)
        bad = float("nan") if rng.random() < 0.3 else float("inf")
        if inject and rank == who:
            params[t_idx].grad[e_idx] = bad
        before = None
        if inject:
            before = [(params[i].detach().clone(), opt.state[params[i]]["exp_avg"].clone() if params[i] in opt.state else None)
                      for i in sample]
        torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        h0 = time.perf_counter()
        e[0].record()
        if reducer is not None:
            reducer.reduce()                           # gradient mean + overflow check (epilogue kernel)
        e[1].record()
        with amp.scale_loss(dummy, opt):
            pass
        opt.step()
        e[2].record()
        host_ms.append((time.perf_counter() - h0) * 1e3)
        torch.cuda.synchronize()
        reduce_ms.append(e[0].elapsed_time(e[1]))
        (skip_ms if inject else normal_ms).append(e[1].elapsed_time(e[2]))
        skipped = sc.update_scale(inject)
        trace_ok &= scaler.loss_scale() == sc.loss_scale and skipped == inject
        if inject:
            for i, (pb, mb) in zip(sample, before):
                untouched_ok &= torch.equal(params[i].detach(), pb)
                if mb is not None:
                    untouched_ok &= torch.equal(opt.state[params[i]]["exp_avg"], mb)
    if world > 1:
        t = torch.tensor([1 if trace_ok else 0, 1 if untouched_ok else 0], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        trace_ok, untouched_ok = bool(int(t[0])), bool(int(t[1]))
    med = lambda xs: sorted(xs)[len(xs) // 2] if xs else None  # noqa: E731
    res = {"config": "BASELINE configs[4]: many-small-tensor list, amp O3 + dynamic loss scale + FusedAdam, inf/NaN injected on "
                     "one random rank" + (", gradient mean through apex.parallel.Reducer (in place, NCCL + epilogue kernel)" if world > 1 else ""),
           "n_gpus": world, "tensors": len(sizes), "params": n, "not_divisible_by_8": sum(1 for s in sizes if s % 8),
           "steps": steps, "injected": len(skip_ms), "scale_trace_equals_oracle": bool(trace_ok),
           "scale_trace_checked_on": "every rank (AND over ranks)", "state_untouched_on_skips": bool(untouched_ok),
           "normal_step_ms_gpu": med(normal_ms),
           "normal_step_gbs_at_14B": 14 * n / (med(normal_ms) * 1e-3) / 1e9 if normal_ms else None,
           "skip_step_ms_gpu": med(skip_ms), "gradient_mean_ms": med(reduce_ms) if world > 1 else None,
           "host_ms_per_step_python": med(host_ms), "setup_s": setup_s,
           "step_definition": "scale_loss exit (check unless the Reducer's epilogue did it, scaler update) + FusedAdam.step",
           "final_loss_scale": sc.loss_scale}
    del model, opt, params, flat_g, reducer
    torch.cuda.empty_cache()
    return res


def run_e2e(sizes, dev, world, steps, warmup, barrier):
    """Same step through apex.amp + apex.optimizers.FusedAdam; per step the gradients come from
    pinned host memory (H2D) and the scaler state is read back (D2H).  At N > 1 the gradient mean is
    taken by apex.parallel.Reducer (in place on the flat gradient buffer: NCCL all-reduce per message +
    this library's epilogue kernel, which also does the overflow check)."""
    import torch.distributed as dist
    from apex import amp
    from apex.optimizers import FusedAdam
    from apex.parallel import Reducer

    amp._amp_state.reset()
    n = sum(sizes)
    gen = torch.Generator(device=dev).manual_seed(0)

    class ParamList(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.ps = torch.nn.ParameterList(
                [torch.nn.Parameter(torch.randn(s, device=dev, generator=gen) * 0.02) for s in sizes])

    model = ParamList()
    opt = FusedAdam(model.parameters(), lr=HYPER["lr"], betas=(HYPER["beta1"], HYPER["beta2"]), eps=HYPER["eps"],
                    weight_decay=HYPER["weight_decay"], set_grad_none=False)
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    host_g = (torch.randn(n) * 1e-3 * LOSS_SCALE).half().pin_memory()
    dev_g = torch.empty(n, dtype=torch.float16, device=dev)
    params = list(model.parameters())
    for prm, g in zip(params, torch.split(dev_g, sizes)):
        prm.grad = g
    reducer = Reducer([p_.grad for p_ in params]).covers(params) if world > 1 else None
    dummy = torch.zeros((), device=dev)
    scaler = amp._amp_state.loss_scalers[0]

    # A caller whose gradients live in host memory overlaps the transfer of step k+1 with the work of
    # step k: the H2D copy runs on its own stream into a staging buffer, the step's stream waits for it,
    # moves it into the gradient buffer (device-to-device) and steps.  Every step still pays one full
    # H2D of its gradients and one D2H read of its result inside the timed region.
    staging = torch.empty(n, dtype=torch.float16, device=dev)
    copy_stream = torch.cuda.Stream(device=dev)
    arrived, consumed = torch.cuda.Event(), torch.cuda.Event()

    def upload():
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed)                   # staging may be overwritten only after the D2D read it
            staging.copy_(host_g, non_blocking=True)           # H2D: the NEXT step's gradients
            arrived.record(copy_stream)

    def one():
        main = torch.cuda.current_stream()
        main.wait_event(arrived)
        dev_g.copy_(staging, non_blocking=True)
        consumed.record(main)
        upload()                                               # overlaps with everything below
        if reducer is not None:
            reducer.reduce()
        with amp.scale_loss(dummy, opt):
            pass                                               # backward already "happened": grads are in place
        opt.step()
        return scaler.loss_scale()                             # D2H: the step's result (scale / skip state)

    consumed.record(torch.cuda.current_stream())
    upload()
    for _ in range(warmup):
        one()
    barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(steps):
        one()
    t1.record()
    barrier()
    ms = t0.elapsed_time(t1) / steps
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    res = {"value": BYTES_PER_PARAM * n * world / (ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": ms,
           "h2d_bytes_per_step": n * 2, "d2h_bytes_per_step": 32, "steps": steps,
           "api": "apex.amp.scale_loss + apex.optimizers.FusedAdam.step (O2)" + (" + apex.parallel.Reducer.reduce" if world > 1 else ""),
           "note": "bound by the host link: 248.9 MB of fp16 gradients per step over PCIe, uploaded on a side stream "
                   "while the previous step runs"}
    del model, opt, params, dev_g, staging, reducer
    torch.cuda.empty_cache()
    return res


if __name__ == "__main__":
    main()
