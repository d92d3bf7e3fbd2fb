#!/usr/bin/env python
"""The Blackwell-capable incumbents already on the box, on the same tensor list and the same B200
(SURVEY.md §2.4): torch 2.11's own sm_100 kernels arranged the way apex arranges its passes.

  apex-style 3 passes:  fp16 grad -> fp32 master grad (x 1/scale) + inf check   [_foreach copy + _amp_foreach_non_finite_check_and_unscale_]
                        Adam on fp32 master/m/v with found_inf skip              [torch._fused_adam_ / _fused_adamw_]
                        fp32 master -> fp16 model copy                            [torch._foreach_copy_]
  ours:                 check -> scaler update -> ONE fused kernel (28 B/param)

    python tools/incumbents.py > gpurun_out/incumbents.json
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

from apex_example_b200 import ops  # noqa: E402
from apex_example_b200.tables import TensorTable  # noqa: E402
from workloads import GPT2_MEDIUM, GPT2_SMALL  # noqa: E402


def timeit(fn, reps=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def run(sizes, name):
    dev = "cuda"
    n = sum(sizes)
    S = 65536.0
    g16 = [(torch.randn(s, device=dev) * 1e-3 * S).half() for s in sizes]
    c16 = [(torch.randn(s, device=dev) * 0.02).half() for s in sizes]
    p = [x.float() for x in c16]
    m = [torch.zeros(s, device=dev) for s in sizes]
    v = [torch.zeros(s, device=dev) for s in sizes]
    mg = [torch.empty(s, device=dev) for s in sizes]
    steps = [torch.zeros((), device=dev) for _ in sizes]
    found = torch.zeros(1, device=dev)
    inv = torch.full((1,), 1.0 / S, device=dev)

    def torch_three_pass():
        torch._foreach_copy_(mg, g16)                                                   # fp16 -> fp32 master grads
        found.zero_()
        torch._amp_foreach_non_finite_check_and_unscale_(mg, found, inv)               # x 1/scale + inf check
        torch._foreach_add_(steps, 1)
        torch._fused_adamw_(p, mg, m, v, [], steps, lr=1e-3, beta1=0.9, beta2=0.999, weight_decay=0.01, eps=1e-8,
                            amsgrad=False, maximize=False, grad_scale=None, found_inf=found)
        torch._foreach_copy_(c16, p)                                                    # master -> model

    def torch_fused_with_scale():
        # torch's fused Adam can unscale and skip by itself, but it has no fp16-grad / fp16-copy path:
        # the grads must already be fp32 and the model copy stays a separate pass
        torch._foreach_copy_(mg, g16)
        found.zero_()
        torch._amp_foreach_non_finite_check_and_unscale_(mg, found, torch.ones(1, device=dev))   # check only
        torch._foreach_add_(steps, 1)
        torch._fused_adamw_(p, mg, m, v, [], steps, lr=1e-3, beta1=0.9, beta2=0.999, weight_decay=0.01, eps=1e-8,
                            amsgrad=False, maximize=False, grad_scale=torch.full((1,), S, device=dev).squeeze(), found_inf=found)
        torch._foreach_copy_(c16, p)

    table = TensorTable([g16, p, m, v, c16])
    chk = TensorTable([g16])
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device=dev)
    ops.scaler_init(state, S)
    k = [0]

    def ours():
        k[0] += 1
        ops.multi_tensor_check(chk, 0, torch.float16, state[ops.IDX_OVERFLOW_ACC:])
        ops.scaler_update(state, True, 2000, 2.0, 0.0, 2.0 ** 24)
        ops.fused_adam(table, torch.float16, torch.float32, torch.float16,
                       [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.01, k[0], 1, 1)],
                       dev_inv_scale=state.view(torch.float32)[ops.IDX_INV_SCALE:], dev_found_inf=state[ops.IDX_FOUND_INF:])

    res = {"list": name, "tensors": len(sizes), "params": n}
    for label, fn in (("torch_apex_style_3_pass", torch_three_pass), ("torch_fused_adamw_grad_scale", torch_fused_with_scale),
                      ("ours_check_update_fused", ours)):
        try:
            ms = timeit(fn)
            res[label] = {"ms": ms, "gbs_at_28B": 28 * n / (ms * 1e-3) / 1e9}
        except Exception as e:                                       # an incumbent API that moved: report, do not hide
            res[label] = {"error": f"{type(e).__name__}: {e}"[:200]}
    if "ms" in res.get("torch_apex_style_3_pass", {}) and "ms" in res["ours_check_update_fused"]:
        res["speedup_vs_torch_3_pass"] = res["torch_apex_style_3_pass"]["ms"] / res["ours_check_update_fused"]["ms"]
    return res


if __name__ == "__main__":
    out = [run(GPT2_SMALL, "GPT-2 small (124.4M)"), run(GPT2_MEDIUM, "GPT-2 medium (354.8M)")]
    print(json.dumps(out, indent=1))
