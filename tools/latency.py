#!/usr/bin/env python
"""Launch-bound regime and skip path: step latency on the small lists of BASELINE configs 1 and 3
(reference ConvNet: 10 tensors / 29,034 params; ResNet-50: 161 tensors / 25.6M params) and the
cost of a SKIPPED step (overflow: only the 2 B/param check runs, the step kernel exits at once)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

from apex_example_b200 import ops  # noqa: E402
from apex_example_b200.tables import TensorTable  # noqa: E402
from workloads import CONVNET, GPT2_SMALL  # noqa: E402


def timeit(fn, reps=200, warm=20):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3      # microseconds


def resnet50_sizes():
    import torchvision
    return [p.numel() for p in torchvision.models.resnet50().parameters()]


def run(name, sizes, reps):
    dev = "cuda"
    g = [(torch.randn(s, device=dev) * 64).half() for s in sizes]
    c = [(torch.randn(s, device=dev) * 0.02).half() for s in sizes]
    p = [x.float() for x in c]
    m = [torch.zeros(s, device=dev) for s in sizes]
    v = [torch.zeros(s, device=dev) for s in sizes]
    t5, t1 = TensorTable([g, p, m, v, c]), TensorTable([g])
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device=dev)
    ops.scaler_init(state, 64.0)
    fs = state.view(torch.float32)
    k = [0]

    def step():
        k[0] += 1
        ops.multi_tensor_check(t1, 0, torch.float16, state[ops.IDX_OVERFLOW_ACC:])
        ops.scaler_update(state, True, 2000, 2.0, 0.0, 2.0 ** 24)
        ops.fused_adam(t5, torch.float16, torch.float32, torch.float16, [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.0, k[0], 1, 1)],
                       dev_inv_scale=fs[ops.IDX_INV_SCALE:], dev_found_inf=state[ops.IDX_FOUND_INF:])

    normal = timeit(step, reps)
    # capture the same three launches in a CUDA graph (hyper-parameters frozen: valid for SGD-like constants;
    # shown only as the launch-overhead floor)
    gr = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        step(); torch.cuda.synchronize()
        with torch.cuda.graph(gr, stream=s):
            step()
    graph = timeit(gr.replay, reps)
    g[-1][-1] = float("inf")                     # every step now overflows and is skipped on the device
    ops.scaler_init(state, 64.0)
    skipped = timeit(step, reps)
    n = sum(sizes)
    return {"list": name, "tensors": len(sizes), "params": n, "step_us": normal, "step_us_cuda_graph": graph,
            "skipped_step_us": skipped, "skip_gbs_at_2B": 2 * n / (skipped * 1e-6) / 1e9,
            "loss_scale_after_skips": float(state.cpu().view(torch.float32)[0])}


if __name__ == "__main__":
    out = [run("reference ConvNet", CONVNET, 500), run("ResNet-50", resnet50_sizes(), 300), run("GPT-2 small", GPT2_SMALL, 50)]
    print(json.dumps(out, indent=1))
