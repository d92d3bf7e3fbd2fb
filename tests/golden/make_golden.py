"""Generates tests/golden/*.npz.  Run HERE (the build container), never on the GPU box:

    python tests/golden/make_golden.py

The reference repository holds no golden vectors for this path, and its own
arithmetic lives in NVIDIA/apex, which is not installed.  So the fixtures are produced by the
numerical contract BASELINE.json names — torch.optim.Adam / AdamW / SGD (foreach=False, CPU,
torch 2.11.0) — on the parameter shapes of the reference's own model, obtained by importing
/root/reference/test_apex_distributed_spawn.py with `apex` stubbed out.

Files:
  convnet_inputs.npz
      p0_i            initial fp32 master of tensor i (10 tensors, 29,034 elements)
      g_s_i           fp16 gradient of tensor i at step s, already multiplied by the loss scale
      scale, names
  convnet_adam.npz / convnet_adamw.npz / convnet_sgd.npz
      p_i, m_i, v_i   (or p_i, buf_i) torch's result after 5 steps; hyper-parameters as 0-d arrays
  scale_trace.npz
      flags, window, scales, unskipped   torch._amp_update_scale_ trace for SURVEY.md §8c (ii)
      cap_scales                          the 2**24 cap trace (iii): apex rule, cross-checked
                                          against torch's uncapped rule up to the cap
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/test_apex_distributed_spawn.py"


def reference_shapes():
    for name in ("apex", "apex.parallel", "apex.amp"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["apex"].amp = sys.modules["apex.amp"]
    sys.modules["apex"].parallel = sys.modules["apex.parallel"]
    spec = importlib.util.spec_from_file_location("ref_script", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.set_seed(0)
    net = mod.ConvNet()
    return [(n, tuple(p.shape)) for n, p in net.named_parameters()], [p.detach().clone() for p in net.parameters()]


def main():
    names, params = reference_shapes()
    assert sum(p.numel() for p in params) == 29034
    S = 65536.0
    steps = 5
    gen = torch.Generator().manual_seed(1234)
    grads = [[((torch.randn(p.shape, generator=gen) * 1e-3) * S).half() for p in params] for _ in range(steps)]

    def run(cls, **kw):
        ps = [torch.nn.Parameter(p.clone()) for p in params]
        opt = cls(ps, foreach=False, **kw)
        for s in range(steps):
            for p, g in zip(ps, grads[s]):
                p.grad = g.float() * (1.0 / S)
            opt.step()
        return ps, opt

    common = {f"p0_{i}": p.numpy() for i, p in enumerate(params)}
    common.update({f"g_{s}_{i}": g.numpy() for s in range(steps) for i, g in enumerate(grads[s])})
    common["scale"] = np.float32(S)
    common["names"] = np.array([n for n, _ in names])
    np.savez_compressed(os.path.join(HERE, "convnet_inputs.npz"), **common)

    for tag, cls, kw in (("adam", torch.optim.Adam, dict(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0)),
                         ("adamw", torch.optim.AdamW, dict(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01))):
        ps, opt = run(cls, **kw)
        out = {}
        for i, p in enumerate(ps):
            out[f"p_{i}"] = p.detach().numpy()
            out[f"m_{i}"] = opt.state[p]["exp_avg"].numpy()
            out[f"v_{i}"] = opt.state[p]["exp_avg_sq"].numpy()
        out.update(lr=kw["lr"], beta1=kw["betas"][0], beta2=kw["betas"][1], eps=kw["eps"], weight_decay=kw["weight_decay"])
        np.savez_compressed(os.path.join(HERE, f"convnet_{tag}.npz"), **out)

    kw = dict(lr=0.1, momentum=0.9, nesterov=True, weight_decay=1e-4)
    ps, opt = run(torch.optim.SGD, **kw)
    out = {}
    for i, p in enumerate(ps):
        out[f"p_{i}"] = p.detach().numpy()
        out[f"buf_{i}"] = opt.state[p]["momentum_buffer"].numpy()
    out.update(lr=kw["lr"], momentum=kw["momentum"], weight_decay=kw["weight_decay"], nesterov=1)
    np.savez_compressed(os.path.join(HERE, "convnet_sgd.npz"), **out)

    flags = [0, 0, 0, 0, 1, 0, 0, 1, 1, 0, 0, 0, 0, 0, 0]
    s, t = torch.tensor(65536.0), torch.tensor(0, dtype=torch.int32)
    scales, unsk = [], []
    for f in flags:
        torch._amp_update_scale_(s, t, torch.tensor(float(f)), 2.0, 0.5, 3)
        scales.append(float(s)); unsk.append(int(t))
    assert scales == [65536, 65536, 131072, 131072, 65536, 65536, 65536, 32768, 16384, 16384, 16384, 32768, 32768,
                      32768, 65536]
    s, t = torch.tensor(2.0 ** 23), torch.tensor(0, dtype=torch.int32)
    cap = []
    for _ in range(9):
        torch._amp_update_scale_(s, t, torch.tensor(0.0), 2.0, 0.5, 3)
        cap.append(min(float(s), 2.0 ** 24))      # apex max_loss_scale
    np.savez_compressed(os.path.join(HERE, "scale_trace.npz"), flags=np.array(flags), window=3,
                        scales=np.array(scales, dtype=np.float32), unskipped=np.array(unsk),
                        cap_scales=np.array(cap, dtype=np.float32))
    print("wrote", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))


if __name__ == "__main__":
    main()
