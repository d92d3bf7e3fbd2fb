"""Pins oracle/amp_step_oracle.c (the CPU restatement) against torch 2.11's own CPU ops and the
committed golden fixtures.  No GPU.  See the oracle header for what "pinned" means here."""
import os

import numpy as np
import pytest
import torch

import oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden")
RTOL, ATOL = 1e-6, 1e-9          # the tolerance BASELINE.json's north_star / SURVEY.md §8c state


def _bits_equal_or_both_nan(a, b):
    ai = a.view(torch.int16 if a.element_size() == 2 else torch.int32)
    bi = b.view(torch.int16 if b.element_size() == 2 else torch.int32)
    return bool(((ai == bi) | (a.isnan() & b.isnan())).all())


@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16])
def test_conversions_match_tensor_copy(dt):
    torch.manual_seed(0)
    x = torch.randn(1 << 18) * torch.logspace(-9, 5, 1 << 18)
    x[:9] = torch.tensor([float("inf"), -float("inf"), float("nan"), 65520.0, 65519.0, 1 + 2 ** -11, 1 + 3 * 2 ** -11,
                          1e-8, -0.0])
    assert _bits_equal_or_both_nan(oracle.convert(x, dt), x.to(dt))
    allbits = torch.arange(65536, dtype=torch.int32).to(torch.int16).view(dt)
    assert _bits_equal_or_both_nan(oracle.convert(allbits, torch.float32), allbits.float())


def test_rne_known_answers():
    # SURVEY.md App. C probe 3
    x = torch.tensor([1 + 2 ** -11, 1 + 3 * 2 ** -11, 65520.0, 65519.0, 1e-8])
    assert oracle.convert(x, torch.float16).tolist() == [1.0, 1.001953125, float("inf"), 65504.0, 0.0]


def test_check_and_scale_match_torch_amp_op():
    torch.manual_seed(1)
    for bad in (None, float("inf"), float("nan"), -float("inf")):
        g = (torch.randn(10001) * 100).half()
        if bad is not None:
            g[7777] = bad
        master, found = oracle.torch_unscale_check([g], 1 / 1024.0)
        out = torch.empty(10001)
        flag = oracle.scale(g, out, 1 / 1024.0)
        assert flag == found == oracle.check(g) == int(bad is not None)
        assert _bits_equal_or_both_nan(out, master[0])


def test_axpby():
    torch.manual_seed(2)
    x, y = torch.randn(999).half(), torch.randn(999)
    out = torch.empty(999)
    assert oracle.axpby(x, y, out, 0.5, 2.0) == 0
    torch.testing.assert_close(out, 0.5 * x.float() + 2.0 * y, rtol=1e-6, atol=1e-7)
    y[5] = float("nan")
    assert oracle.axpby(x, y, out, 0.5, 2.0, arg_to_check=0) == 0
    assert oracle.axpby(x, y, out, 0.5, 2.0, arg_to_check=1) == 1
    assert oracle.axpby(x, y, out, 0.5, 2.0, arg_to_check=-1) == 1


def _adam_run(p0, grads, S, **kw):
    p, m, v = p0.clone(), torch.zeros_like(p0), torch.zeros_like(p0)
    c = torch.empty(p0.shape, dtype=torch.float16)
    for s, g in enumerate(grads, 1):
        oracle.adam(g, p, m, v, c, step=s, inv_scale=1.0 / S, **kw)
    return p, m, v, c


@pytest.mark.parametrize("adamw,wd", [(False, 0.0), (True, 0.0), (True, 0.01)])
def test_adam_matches_torch_optim(adamw, wd):
    torch.manual_seed(3)
    N, S = 1 << 18, 65536.0
    p0 = torch.randn(N) * 0.05
    grads = [((torch.randn(N) * 1e-3) * S).half() for _ in range(6)]
    pt, mt, vt = oracle.torch_adam_steps(p0, grads, adamw=adamw, inv_scale=1 / S, lr=1e-3, betas=(0.9, 0.999),
                                         eps=1e-8, weight_decay=wd)
    p, m, v, c = _adam_run(p0, grads, S, lr=1e-3, weight_decay=wd, adam_w_mode=adamw)
    torch.testing.assert_close(p, pt, rtol=RTOL, atol=ATOL)
    torch.testing.assert_close(m, mt, rtol=RTOL, atol=ATOL)
    torch.testing.assert_close(v, vt, rtol=RTOL, atol=ATOL)
    assert torch.equal(c, p.half())                       # model copy is a pure RNE cast of the master


def test_adam_l2_mode_as_accurate_as_torch_fp32():
    """L2 weight decay (g + wd*p) cancels catastrophically for a few elements, where torch's own
    fp32 result misses its fp64 result by more than rtol 1e-6.  The restatement
    must (a) agree with torch fp32 on all but a 1e-4 fraction and (b) be as close to fp64."""
    torch.manual_seed(4)
    N, S = 1 << 18, 65536.0
    p0 = torch.randn(N) * 0.05
    grads = [((torch.randn(N) * 1e-3) * S).half() for _ in range(5)]
    kw = dict(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01)
    pt, _, _ = oracle.torch_adam_steps(p0, grads, adamw=False, inv_scale=1 / S, **kw)
    p64 = torch.nn.Parameter(p0.double())
    o = torch.optim.Adam([p64], foreach=False, **kw)
    for g in grads:
        p64.grad = g.double() / S
        o.step()
    p, _, _, _ = _adam_run(p0, grads, S, lr=1e-3, weight_decay=0.01, adam_w_mode=False)
    frac_bad = (~torch.isclose(p, pt, rtol=RTOL, atol=ATOL)).float().mean().item()
    assert frac_bad < 1e-4
    err_mine = (p.double() - p64.detach()).abs().max().item()
    err_torch = (pt.double() - p64.detach()).abs().max().item()
    assert err_mine <= 1.5 * err_torch + 1e-9


def test_adam_beta1_below_half_uses_other_lerp_form():
    torch.manual_seed(5)
    N = 4096
    p0 = torch.randn(N) * 0.05
    grads = [torch.randn(N) * 1e-3 for _ in range(3)]
    pt, mt, vt = oracle.torch_adam_steps(p0, grads, adamw=False, lr=1e-3, betas=(0.3, 0.999), eps=1e-8)
    p, m, v = p0.clone(), torch.zeros(N), torch.zeros(N)
    for s, g in enumerate(grads, 1):
        oracle.adam(g, p, m, v, None, lr=1e-3, beta1=0.3, step=s, adam_w_mode=False)
    torch.testing.assert_close(p, pt, rtol=RTOL, atol=ATOL)
    torch.testing.assert_close(m, mt, rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("kw", [dict(lr=0.1), dict(lr=0.1, momentum=0.9),
                                dict(lr=0.1, momentum=0.9, nesterov=True, weight_decay=1e-4),
                                dict(lr=0.1, momentum=0.9, dampening=0.1, weight_decay=1e-4)])
def test_sgd_matches_torch_optim(kw):
    torch.manual_seed(6)
    N, S = 1 << 18, 1024.0
    p0 = torch.randn(N) * 0.05
    grads = [((torch.randn(N) * 1e-3) * S).half() for _ in range(4)]
    pt, bt = oracle.torch_sgd_steps(p0, grads, inv_scale=1 / S, **kw)
    p, b = p0.clone(), torch.zeros(N)
    c = torch.empty(N, dtype=torch.bfloat16)
    for s, g in enumerate(grads):
        oracle.sgd(g, p, b, c, first_run=(s == 0), inv_scale=1 / S, **kw)
    torch.testing.assert_close(p, pt, rtol=RTOL, atol=ATOL)
    if bt is not None:
        torch.testing.assert_close(b, bt, rtol=RTOL, atol=ATOL)
    assert torch.equal(c, p.bfloat16())


def test_sgd_wd_after_momentum_rule():
    # apex-only flag: g = mom; g += wd*w AFTER the momentum update
    p, b = torch.tensor([1.0, -2.0]), torch.tensor([0.5, 0.5])
    g = torch.tensor([0.25, 0.25])
    oracle.sgd(g, p, b, None, lr=0.1, momentum=0.5, weight_decay=0.1, wd_after_momentum=True)
    mom = 0.5 * 0.5 + 0.25
    assert b.tolist() == [mom, mom]
    torch.testing.assert_close(p, torch.tensor([1.0 - 0.1 * (mom + 0.1 * 1.0), -2.0 - 0.1 * (mom - 0.2)]))


def test_o3_half_state_close_to_fp32_math():
    torch.manual_seed(7)
    N = 8192
    p = (torch.randn(N) * 0.05).half()
    m, v = torch.zeros(N).half(), torch.zeros(N).half()
    g = (torch.randn(N) * 1e-2).half()
    pf, mf, vf = p.float(), m.float(), v.float()
    oracle.adam(g, pf, mf, vf, None, lr=1e-3, step=1)
    oracle.adam(g, p, m, v, None, lr=1e-3, step=1)
    assert torch.equal(p, pf.half()) and torch.equal(m, mf.half()) and torch.equal(v, vf.half())


def test_scaler_traces_golden_and_torch():
    z = np.load(os.path.join(GOLD, "scale_trace.npz"))
    sc = oracle.LossScaler(scale_window=int(z["window"]))
    s, t = 65536.0, 0
    for f, want_s, want_u in zip(z["flags"], z["scales"], z["unskipped"]):
        skipped = sc.update_scale(bool(f))
        s, t = oracle.torch_update_scale(s, t, int(f), int(z["window"]))
        assert skipped == bool(f)
        assert sc.loss_scale == float(want_s) == s and sc.unskipped == int(want_u) == t
    sc = oracle.LossScaler(loss_scale=2.0 ** 23, scale_window=3)
    got = []
    for _ in range(9):
        sc.update_scale(False)
        got.append(sc.loss_scale)
    assert got == z["cap_scales"].tolist() and got[-1] == 2.0 ** 24


def test_scaler_static_never_skips_and_min_scale():
    sc = oracle.LossScaler(loss_scale=128.0, dynamic=False)
    assert sc.update_scale(True) is False and sc.loss_scale == 128.0
    sc = oracle.LossScaler(loss_scale=4.0, min_loss_scale=2.0)
    assert sc.update_scale(True) and sc.loss_scale == 2.0
    assert sc.update_scale(True) and sc.loss_scale == 2.0
    assert sc.inv_scale == 0.5


def test_bucket_epilogue_matches_torch_mul():
    torch.manual_seed(8)
    for dt in (torch.float16, torch.bfloat16, torch.float32):
        x = (torch.randn(4099) * 300).to(dt)
        want = x.clone().mul_(1.0 / 8)
        assert oracle.bucket_epilogue(x, 1.0 / 8) == 0
        assert torch.equal(x, want)
    x = torch.tensor([1.0, float("inf"), 2.0]).half()
    assert oracle.bucket_epilogue(x, 0.5) == 1


def _load_inputs():
    z = np.load(os.path.join(GOLD, "convnet_inputs.npz"))
    n = len(z["names"])
    p0 = [torch.from_numpy(z[f"p0_{i}"]).flatten().clone() for i in range(n)]
    grads = [[torch.from_numpy(z[f"g_{s}_{i}"]).flatten().clone() for i in range(n)] for s in range(5)]
    return n, p0, grads, float(z["scale"])


@pytest.mark.parametrize("tag,adamw", [("adam", False), ("adamw", True)])
def test_golden_convnet_adam(tag, adamw):
    n, p0, grads, S = _load_inputs()
    z = np.load(os.path.join(GOLD, f"convnet_{tag}.npz"))
    assert sum(p.numel() for p in p0) == 29034
    for i in range(n):
        p, m, v = p0[i].clone(), torch.zeros_like(p0[i]), torch.zeros_like(p0[i])
        for s in range(5):
            oracle.adam(grads[s][i], p, m, v, None, lr=float(z["lr"]), beta1=float(z["beta1"]), beta2=float(z["beta2"]),
                        eps=float(z["eps"]), weight_decay=float(z["weight_decay"]), step=s + 1, adam_w_mode=adamw,
                        inv_scale=1.0 / S)
        torch.testing.assert_close(p, torch.from_numpy(z[f"p_{i}"]).flatten(), rtol=RTOL, atol=ATOL)
        torch.testing.assert_close(m, torch.from_numpy(z[f"m_{i}"]).flatten(), rtol=RTOL, atol=ATOL)
        torch.testing.assert_close(v, torch.from_numpy(z[f"v_{i}"]).flatten(), rtol=RTOL, atol=ATOL)


def test_golden_convnet_sgd():
    n, p0, grads, S = _load_inputs()
    z = np.load(os.path.join(GOLD, "convnet_sgd.npz"))
    for i in range(n):
        p, b = p0[i].clone(), torch.zeros_like(p0[i])
        for s in range(5):
            oracle.sgd(grads[s][i], p, b, None, lr=float(z["lr"]), momentum=float(z["momentum"]),
                       weight_decay=float(z["weight_decay"]), nesterov=True, first_run=(s == 0), inv_scale=1.0 / S)
        torch.testing.assert_close(p, torch.from_numpy(z[f"p_{i}"]).flatten(), rtol=RTOL, atol=ATOL)
        torch.testing.assert_close(b, torch.from_numpy(z[f"buf_{i}"]).flatten(), rtol=RTOL, atol=ATOL)


def test_list_step_skips_everything_on_overflow():
    torch.manual_seed(9)
    ps = [torch.randn(n) for n in (7, 1000, 65537)]
    gs = [(torch.randn_like(p) * 64).half() for p in ps]
    ms, vs = [torch.zeros_like(p) for p in ps], [torch.zeros_like(p) for p in ps]
    cs = [torch.zeros(p.shape, dtype=torch.float16) for p in ps]
    before = [p.clone() for p in ps]
    gs[2][-1] = float("inf")
    assert oracle.adam_list(gs, ps, ms, vs, cs, inv_scale=1 / 64, lr=1e-3, step=1) == 1
    assert all(torch.equal(a, b) for a, b in zip(ps, before)) and all(int(m.abs().sum()) == 0 for m in ms)
    gs[2][-1] = 1.0
    assert oracle.adam_list(gs, ps, ms, vs, cs, inv_scale=1 / 64, lr=1e-3, step=1) == 0
    assert not torch.equal(ps[0], before[0]) and torch.equal(cs[1], ps[1].half())


def test_vectorised_o2_path_is_bit_identical_to_the_generic_loop():
    """The timed CPU baseline uses an F16C/AVX2-vectorisable loop for the O2 layout; it must give
    exactly the generic scalar loop's bits (including inf / NaN / subnormal inputs)."""
    import ctypes
    force = ctypes.c_int.in_dll(oracle.lib(), "oracle_force_generic")
    N = 1 << 16
    for adamw, wd, b1 in ((True, 0.01, 0.9), (False, 0.01, 0.9), (True, 0.0, 0.3)):
        outs = []
        for f in (0, 1):
            force.value = f
            try:
                gen = torch.Generator().manual_seed(5)
                p = torch.randn(N, generator=gen) * 0.05
                m = torch.randn(N, generator=gen) * 1e-3
                v = (torch.randn(N, generator=gen) * 1e-3) ** 2
                c = torch.zeros(N, dtype=torch.float16)
                g = (torch.randn(N, generator=gen) * 64).half()
                g[5], g[6], g[7], g[8] = float("inf"), float("nan"), 6e-8, -0.0
                oracle.adam(g, p, m, v, c, lr=1e-3, beta1=b1, weight_decay=wd, step=3, adam_w_mode=adamw, inv_scale=1 / 64)
                outs.append((p, m, v, c))
            finally:
                force.value = 0
        for a, b in zip(*outs):
            assert _bits_equal_or_both_nan(a, b)


def test_lamb_oracle_follows_the_published_rule():
    """oracle.lamb_list against a double-precision restatement of LAMB (You et al. 2020, Alg. 2, with the global-norm
    clip and bias correction apex.optimizers.FusedLAMB documents).  apex itself is not in the image: UNPINNED against
    apex; this pins the oracle to the published rule within fp32 rounding (rtol 1e-5)."""
    gen = torch.Generator().manual_seed(9)
    sizes = [5, 4096, 4097, 10000]
    ps = [torch.randn(n, generator=gen) * 0.1 for n in sizes]
    ms = [torch.zeros(n) for n in sizes]
    vs = [torch.zeros(n) for n in sizes]
    dp, dm, dv = ([t.double() for t in lst] for lst in (ps, ms, vs))
    lr, wd, b1, b2, eps, max_norm = 1e-2, 0.01, 0.9, 0.999, 1e-6, 1.0
    for step in (1, 2, 3):
        gs = [torch.randn(n, generator=gen) * 0.05 for n in sizes]
        ss, total, flag = oracle.l2norm_per_tensor(gs, 4096)
        assert flag == 0
        torch.testing.assert_close(ss.double(), torch.stack([g.double().square().sum() for g in gs]), rtol=1e-5, atol=0)
        oracle.lamb_list(gs, ps, ms, vs, tile_elems=4096, grad_sumsq=total, max_grad_norm=max_norm, lr=lr,
                         weight_decay=wd, step=step, eps=eps)
        gn = torch.sqrt(sum(g.double().square().sum() for g in gs))
        clip = max(gn.item() / max_norm, 1.0)
        for p, g, m, v in zip(dp, gs, dm, dv):
            g = g.double() / clip
            m.mul_(b1).add_(g, alpha=1 - b1)
            v.mul_(b2).addcmul_(g, g, value=1 - b2)
            u = (m / (1 - b1 ** step)) / ((v / (1 - b2 ** step)).sqrt() + eps) + wd * p
            p.sub_(lr * p.norm() / u.norm() * u)
        for a, b in zip(ps, dp):
            torch.testing.assert_close(a.double(), b, rtol=1e-5, atol=1e-8)


def test_l2norm_oracle_flags_non_finite_values():
    xs = [torch.ones(10), torch.tensor([1.0, float("inf")])]
    _, _, flag = oracle.l2norm_per_tensor(xs, 4096)
    assert flag == 1


def test_novograd_oracle_follows_the_published_rule():
    """oracle.novograd_list against a double-precision restatement of NovoGrad with apex's documented conventions
    (layer-wise second moment kept as a norm, first step initialised from the gradient norm, bias corrections 1-b1^t and
    sqrt(1-b2^t)).  UNPINNED against apex itself; rtol 1e-5 against the published rule."""
    gen = torch.Generator().manual_seed(21)
    sizes = [5, 4096, 4097, 10000]
    ps = [torch.randn(n, generator=gen) * 0.1 for n in sizes]
    ms = [torch.zeros(n) for n in sizes]
    v = torch.zeros(len(sizes))
    dp, dm = ([t.double() for t in lst] for lst in (ps, ms))
    dv = [None] * len(sizes)
    lr, wd, b1, b2, eps = 1e-2, 0.01, 0.95, 0.98, 1e-8
    for step in (1, 2, 3):
        gs = [torch.randn(n, generator=gen) * 0.05 for n in sizes]
        ss, _, _ = oracle.l2norm_per_tensor(gs, 4096)
        oracle.novograd_list(gs, ps, ms, v, ss, lr=lr, weight_decay=wd, beta1=b1, beta2=b2, eps=eps, step=step)
        for i, (p, g, m) in enumerate(zip(dp, gs, dm)):
            g = g.double()
            n = g.norm()
            dv[i] = n if step == 1 else torch.sqrt(b2 * dv[i] ** 2 + (1 - b2) * n ** 2)
            m.mul_(b1).add_(g, alpha=1 - b1)
            denom = dv[i] / (1 - b2 ** step) ** 0.5 + eps
            p.sub_(lr * ((m / (1 - b1 ** step)) / denom + wd * p))
        torch.testing.assert_close(v.double(), torch.stack(dv), rtol=1e-5, atol=0)
        for a, b in zip(ps, dp):
            torch.testing.assert_close(a.double(), b, rtol=1e-5, atol=1e-8)
