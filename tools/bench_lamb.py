#!/usr/bin/env python
"""Time one FusedLAMB step (amp O2 layout: fp16 gradients, fp32 master/m/v, fp16 model copy) on the GPT-2-small-shaped
list of bench.py, kernel by kernel, with CUDA events.  Algorithmic bytes per parameter: gradient norm 2, parameter norm 4,
stage 1 26 (g 2, p 4, m 8, v 8, u 4), update norm 4, stage 2 14 (u 4, p 8, copy 2) = 50."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from apex_example_b200 import ops  # noqa: E402
from apex_example_b200.tables import TensorTable  # noqa: E402
from bench import workload_sizes  # noqa: E402


def main():
    sizes, desc = workload_sizes("gpt2_small")
    dev = torch.device("cuda", 0)
    n = sum(sizes)
    g = [torch.randn(s, device=dev).mul_(1e-3).half() for s in sizes]
    p = [torch.randn(s, device=dev).mul_(0.02) for s in sizes]
    m = [torch.zeros(s, device=dev) for s in sizes]
    v = [torch.zeros(s, device=dev) for s in sizes]
    u = [torch.zeros(s, device=dev) for s in sizes]
    c = [x.half() for x in p]
    t1 = TensorTable([g, p, m, v, u])
    t2 = TensorTable([u, p, c], tile_elems=t1.tile_elems)
    first = t1.first_tiles()
    tile_ss = torch.empty(t1.n_tiles, device=dev)
    ss, pn, un = (torch.empty(len(sizes), device=dev) for _ in range(3))
    gsum = torch.zeros(1, device=dev)
    names = ["grad_norm", "param_norm", "stage1", "update_norm", "stage2"]
    bytes_per = [2, 4, 26, 4, 14]

    def step(k, evs=None):
        grp = [ops.LambGroup(1e-3, 0.9, 0.999, 1e-6, 0.01, k, 1, 1, 1, 0)]
        fns = [lambda: (gsum.zero_(), ops.l2norm_per_tensor(t1, 0, torch.float16, first, tile_ss, ss, total=gsum)),
               lambda: ops.l2norm_per_tensor(t1, 1, torch.float32, first, tile_ss, ss, norms=pn),
               lambda: ops.lamb_stage1(t1, torch.float16, torch.float32, grp, gsum, 1.0),
               lambda: ops.l2norm_per_tensor(t1, 4, torch.float32, first, tile_ss, ss, norms=un),
               lambda: ops.lamb_stage2(t2, torch.float32, torch.float16, grp, pn, un)]
        for i, f in enumerate(fns):
            if evs is not None:
                evs[i][0].record()
            f()
            if evs is not None:
                evs[i][1].record()

    for k in range(1, 6):
        step(k)
    torch.cuda.synchronize()
    K = 20
    all_evs = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in names] for _ in range(K)]
    t0, t1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for k in range(K):
        step(6 + k, all_evs[k])
    t1e.record()
    torch.cuda.synchronize()
    out = {"workload": desc, "params": n, "ms_per_step": t0.elapsed_time(t1e) / K, "tile_elems": t1.tile_elems}
    for i, nm in enumerate(names):
        ms = sum(e[i][0].elapsed_time(e[i][1]) for e in all_evs) / K
        out[nm] = {"ms": ms, "gbs": bytes_per[i] * n / (ms * 1e-3) / 1e9}
    out["gbs_at_50B"] = 50 * n / (out["ms_per_step"] * 1e-3) / 1e9
    print(json.dumps(out))


if __name__ == "__main__":
    main()
