"""-m gpu tests of the drop-in API: apex.optimizers.FusedAdam/FusedSGD, apex.amp (O0-O3),
amp_C / multi_tensor_applier, against the oracle (bit-exact where the inputs are identical) and
against a plain-torch emulation of apex's O2 recipe (tolerance: the forward/backward is torch's
own and identical on both sides; only the optimizer arithmetic differs, rtol 1e-6 / atol 1e-7)."""
import copy
import io
import contextlib

import pytest
import torch

import oracle
from gpu_util import assert_lists_bit_equal

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _fresh_amp_state():
    from apex_example_b200.amp._amp_state import _amp_state
    _amp_state.reset()
    yield
    _amp_state.reset()


def _mlp(seed=0, din=32, dh=64, dout=8):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(din, dh), torch.nn.ReLU(), torch.nn.Linear(dh, dh), torch.nn.ReLU(),
                               torch.nn.Linear(dh, dout)).cuda()


# ------------------------------------------------------------------------------------------------
# fused optimizers without amp: Python-side grouping / step logic, bit-exact vs the oracle
# ------------------------------------------------------------------------------------------------
def test_fused_adam_param_groups_bit_exact_vs_oracle():
    from apex.optimizers import FusedAdam
    gen = torch.Generator().manual_seed(0)
    sizes = [(33, 7), (129,), (64, 64), (5,), (4097,)]
    host = [torch.randn(s, generator=gen) * 0.1 for s in sizes]
    params = [torch.nn.Parameter(h.clone().cuda()) for h in host]
    opt = FusedAdam([{"params": params[:2], "lr": 1e-2, "weight_decay": 0.1},
                     {"params": params[2:4], "betas": (0.8, 0.95)},
                     {"params": params[4:], "bias_correction": False, "eps": 1e-6}], lr=1e-3, weight_decay=0.0)
    hyper = [dict(lr=1e-2, weight_decay=0.1), dict(lr=1e-2, weight_decay=0.1),
             dict(lr=1e-3, beta1=0.8, beta2=0.95), dict(lr=1e-3, beta1=0.8, beta2=0.95),
             dict(lr=1e-3, bias_correction=False, eps=1e-6)]
    hp = [h.flatten().clone() for h in host]
    hm = [torch.zeros_like(h) for h in hp]
    hv = [torch.zeros_like(h) for h in hp]
    for step in range(1, 5):
        grads = [torch.randn(s, generator=gen) for s in sizes]
        for p, g in zip(params, grads):
            p.grad = g.cuda()
        opt.step()
        for i, g in enumerate(grads):
            oracle.adam(g.flatten(), hp[i], hm[i], hv[i], None, step=step, adam_w_mode=True, **hyper[i])
    torch.cuda.synchronize()
    assert_lists_bit_equal([p.detach().flatten() for p in params], hp, "p")
    assert_lists_bit_equal([opt.state[p]["exp_avg"].flatten() for p in params], hm, "exp_avg")
    assert_lists_bit_equal([opt.state[p]["exp_avg_sq"].flatten() for p in params], hv, "exp_avg_sq")
    assert [g["step"] for g in opt.param_groups] == [4, 4, 4]


def test_fused_adam_matches_torch_adamw_cuda():
    from apex.optimizers import FusedAdam
    m1, m2 = _mlp(1), _mlp(1)
    o1 = FusedAdam(m1.parameters(), lr=1e-3, weight_decay=0.01)
    o2 = torch.optim.AdamW(m2.parameters(), lr=1e-3, weight_decay=0.01, foreach=False)
    x = torch.randn(64, 32, device="cuda")
    for _ in range(5):
        for m, o in ((m1, o1), (m2, o2)):
            o.zero_grad()
            m(x).pow(2).mean().backward()
            o.step()
    for a, b in zip(m1.parameters(), m2.parameters()):
        torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-7)


def test_fused_sgd_matches_torch_sgd_and_oracle():
    from apex.optimizers import FusedSGD
    gen = torch.Generator().manual_seed(3)
    host = [torch.randn(n, generator=gen) * 0.1 for n in (1000, 17, 4096)]
    params = [torch.nn.Parameter(h.clone().cuda()) for h in host]
    tparams = [torch.nn.Parameter(h.clone().cuda()) for h in host]
    opt = FusedSGD(params, lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True)
    topt = torch.optim.SGD(tparams, lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True, foreach=False)
    hp = [h.clone() for h in host]
    hb = [torch.zeros_like(h) for h in host]
    for step in range(4):
        grads = [torch.randn(h.shape, generator=gen) for h in host]
        for p, tp, g in zip(params, tparams, grads):
            p.grad = g.cuda()
            tp.grad = g.cuda()
        opt.step()
        topt.step()
        for i, g in enumerate(grads):
            oracle.sgd(g, hp[i], hb[i], None, lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True, first_run=(step == 0))
    torch.cuda.synchronize()
    assert_lists_bit_equal([p.detach() for p in params], hp, "w")
    assert_lists_bit_equal([opt.state[p]["momentum_buffer"] for p in params], hb, "momentum_buffer")
    for a, b in zip(params, tparams):
        torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-7)


def test_fused_optimizer_error_behaviour():
    from apex.optimizers import FusedAdam, FusedSGD
    p = torch.nn.Parameter(torch.zeros(4, device="cuda"))
    with pytest.raises(RuntimeError, match="AMSGrad"):
        FusedAdam([p], amsgrad=True)
    opt = FusedAdam([p])
    p.grad = torch.zeros(4, device="cuda")
    with pytest.raises(RuntimeError, match="has been updated"):
        opt.step(grads=[p.grad])
    p64 = torch.nn.Parameter(torch.zeros(4, device="cuda", dtype=torch.float64))
    p64.grad = torch.zeros(4, device="cuda", dtype=torch.float64)
    with pytest.raises(RuntimeError, match="only support fp16 and fp32"):
        FusedAdam([p64]).step()
    cpu = torch.nn.Parameter(torch.zeros(4))
    cpu.grad = torch.zeros(4)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        FusedSGD([cpu], lr=0.1).step()
    with pytest.raises(ValueError):
        FusedSGD([p], lr=0.1, nesterov=True)


# ------------------------------------------------------------------------------------------------
# the step window: skips decided on the device, exact step numbers, no host sync in between
# ------------------------------------------------------------------------------------------------
def test_device_side_skips_keep_bias_correction_exact():
    from apex import amp
    from apex.optimizers import FusedAdam
    gen = torch.Generator().manual_seed(4)
    sizes = [5000, 33, 4096]
    p16 = [torch.nn.Parameter((torch.randn(n, generator=gen) * 0.1).cuda()) for n in sizes]

    class Holder(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.ps = torch.nn.ParameterList(p16)
    model = Holder()
    opt = FusedAdam(model.parameters(), lr=1e-2, weight_decay=0.01)
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    hp = [p.detach().float().cpu().clone() for p in model.parameters()]      # masters start as the fp16 values
    hm = [torch.zeros_like(h) for h in hp]
    hv = [torch.zeros_like(h) for h in hp]
    hc = [h.half() for h in hp]
    sc = oracle.LossScaler()
    overflow_at = {0, 1, 2, 6, 7, 11}          # includes three in a row at the start (the usual warm-up pattern)
    good = 0
    dummy = torch.zeros((), device="cuda")
    for it in range(14):
        S = sc.loss_scale
        grads = [(torch.randn(n, generator=gen) * 1e-3 * S).half() for n in sizes]
        if it in overflow_at:
            grads[it % 3][it] = float("inf") if it % 2 else float("nan")
        for p, g in zip(model.parameters(), grads):
            p.grad = g.cuda()
        with amp.scale_loss(dummy, opt):
            pass
        opt.step()                              # never synchronises; the kernel picks the right step number
        skip = sc.update_scale(any(oracle.check(g) for g in grads))
        assert skip == (it in overflow_at)
        if not skip:
            good += 1
            for i, g in enumerate(grads):
                oracle.adam(g, hp[i], hm[i], hv[i], hc[i], lr=1e-2, weight_decay=0.01, step=good, inv_scale=sc.inv_scale)
    torch.cuda.synchronize()
    masters = list(amp.master_params(opt))
    assert_lists_bit_equal([m.detach() for m in masters], hp, "master")
    assert_lists_bit_equal([opt.state[m]["exp_avg"] for m in masters], hm, "exp_avg")
    assert_lists_bit_equal([opt.state[m]["exp_avg_sq"] for m in masters], hv, "exp_avg_sq")
    assert_lists_bit_equal([p.detach() for p in model.parameters()], hc, "model copy")
    scaler = amp._amp_state.loss_scalers[0]
    assert scaler.loss_scale() == sc.loss_scale and scaler._unskipped == sc.unskipped
    assert opt.state_dict()["param_groups"][0]["step"] == good == 8


# ------------------------------------------------------------------------------------------------
# amp end to end vs a plain-torch emulation of apex's O2 recipe
# ------------------------------------------------------------------------------------------------
def _torch_o2_emulation(model32, xs, ys, lr, wd, overflow_steps, make_opt):
    model16 = copy.deepcopy(model32).half()
    masters = [p.detach().clone().float().requires_grad_(True) for p in model16.parameters()]
    opt = make_opt(masters)
    scale, unskipped, trace = 65536.0, 0, []
    for it, (x, y) in enumerate(zip(xs, ys)):
        if it in overflow_steps:
            x = x * float("inf")
        for p in model16.parameters():
            p.grad = None
        loss = torch.nn.functional.mse_loss(model16(x.half()).float(), y)
        (loss.float() * scale).backward()
        grads = [p.grad for p in model16.parameters()]
        found = any(not torch.isfinite(g).all() for g in grads)
        if found:
            scale, unskipped = scale / 2, 0
        else:
            unskipped += 1
            for m, g in zip(masters, grads):
                m.grad = g.float() * (1.0 / (scale))
            opt.step()
            with torch.no_grad():
                for p, m in zip(model16.parameters(), masters):
                    p.copy_(m)
            if unskipped == 2000:
                scale, unskipped = min(scale * 2, 2.0 ** 24), 0
        trace.append((found, scale))
    return model16, masters, trace


@pytest.mark.parametrize("which", ["fused_adam", "fused_sgd_lazy", "fused_sgd_materialized", "torch_sgd"])
def test_amp_o2_training_matches_torch_emulation(which):
    from apex import amp
    from apex.optimizers import FusedAdam, FusedSGD
    model = _mlp(7)
    ref = copy.deepcopy(model)
    torch.manual_seed(11)
    xs = [torch.randn(32, 32, device="cuda") for _ in range(12)]
    ys = [torch.randn(32, 8, device="cuda") for _ in range(12)]
    overflow = {0, 1, 5}
    lr, wd = 1e-2, 0.01
    if which == "fused_adam":
        opt = FusedAdam(model.parameters(), lr=lr, weight_decay=wd)
        mk = lambda ps: torch.optim.AdamW(ps, lr=lr, weight_decay=wd, foreach=False)
    elif which.startswith("fused_sgd"):
        opt = FusedSGD(model.parameters(), lr=lr, momentum=0.9, weight_decay=wd,
                       materialize_master_grads=which.endswith("materialized"))
        mk = lambda ps: torch.optim.SGD(ps, lr=lr, momentum=0.9, weight_decay=wd, foreach=False)
    else:   # the reference script's own choice: a stock torch optimizer under amp O2
        opt = torch.optim.SGD(model.parameters(), lr=lr, momentum=0.9, weight_decay=wd, foreach=False)
        mk = lambda ps: torch.optim.SGD(ps, lr=lr, momentum=0.9, weight_decay=wd, foreach=False)
    out = io.StringIO()
    with contextlib.redirect_stdout(out):
        model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=1)
        trace = []
        for it, (x, y) in enumerate(zip(xs, ys)):
            if it in overflow:
                x = x * float("inf")
            loss = torch.nn.functional.mse_loss(model(x), y)          # forward patch casts x to fp16, output to fp32
            opt.zero_grad()
            with amp.scale_loss(loss, opt) as scaled:
                scaled.backward()
            opt.step()
            trace.append(amp._amp_state.loss_scalers[0].loss_scale())
        with amp.scale_loss(torch.zeros((), device="cuda"), opt, delay_unscale=True):
            pass                                                       # flushes any pending overflow message
    ref16, ref_masters, ref_trace = _torch_o2_emulation(ref, xs, ys, lr, wd, overflow, mk)
    assert trace == [s for _, s in ref_trace]                           # scale decisions: exact
    assert out.getvalue().count("Gradient overflow.  Skipping step, loss scaler 0 reducing loss scale to") == len(overflow)
    for p, q in zip(model.parameters(), ref16.parameters()):
        assert p.dtype == torch.float16
        torch.testing.assert_close(p.detach().float(), q.detach().float(), rtol=0, atol=2e-3 * float(q.detach().abs().max()))
    # Tolerance: both sides run torch's own fp16 forward/backward, so the first fp16 weight that rounds
    # differently (masters differ by ~1e-9) changes every later gradient at the 1e-3 relative level, and
    # Adam turns that into ~1e-3 * lr per step.  A wrong recipe (missing unscale, a skipped step counted
    # in the bias correction, a missed skip) is off by ~lr = 1e-2 per step, 50x the bound used here.
    for m, q in zip(amp.master_params(opt), ref_masters):
        assert m.dtype == torch.float32
        torch.testing.assert_close(m.detach(), q.detach(), rtol=0, atol=2e-4)


@pytest.mark.parametrize("level", ["O0", "O1", "O3"])
def test_other_opt_levels_run_and_keep_their_dtypes(level):
    from apex import amp
    from apex.optimizers import FusedAdam
    model = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.BatchNorm1d(32), torch.nn.ReLU(),
                                torch.nn.Linear(32, 4)).cuda()
    opt = FusedAdam(model.parameters(), lr=1e-3)
    kwargs = dict(loss_scale="dynamic") if level == "O3" else {}
    model, opt = amp.initialize(model, opt, opt_level=level, verbosity=0, **kwargs)
    want = {"O0": torch.float32, "O1": torch.float32, "O3": torch.float16}[level]
    assert all(p.dtype == want for p in model.parameters())
    x = torch.randn(8, 16, device="cuda")
    before = [p.detach().clone() for p in model.parameters()]
    for _ in range(2):
        out = model(x)
        assert out.dtype == (torch.float16 if level == "O1" else torch.float32)
        loss = out.float().pow(2).mean()
        opt.zero_grad()
        with amp.scale_loss(loss, opt) as scaled:
            scaled.backward()
        opt.step()
    torch.cuda.synchronize()
    assert any(not torch.equal(a, p.detach()) for a, p in zip(before, model.parameters()))
    assert list(amp.master_params(opt))[0].dtype == want


def test_o2_keeps_batchnorm_fp32_and_state_dict_is_fp32():
    from apex import amp
    from apex.optimizers import FusedSGD
    model = torch.nn.Sequential(torch.nn.Conv2d(1, 4, 3), torch.nn.BatchNorm2d(4), torch.nn.ReLU(), torch.nn.Flatten(),
                                torch.nn.Linear(4 * 26 * 26, 10)).cuda()
    opt = FusedSGD(model.parameters(), lr=0.1, momentum=0.9, materialize_master_grads=False)
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    assert model[0].weight.dtype == torch.float16 and model[1].weight.dtype == torch.float32
    assert all(v.dtype != torch.float16 for v in model.state_dict().values())
    x = torch.randn(4, 1, 28, 28, device="cuda")
    loss = model(x).pow(2).mean()
    with amp.scale_loss(loss, opt) as scaled:
        scaled.backward()
    opt.step()
    sd = amp.state_dict()
    assert sd == {"loss_scaler0": {"loss_scale": 65536.0, "unskipped": 1}}
    amp.load_state_dict({"loss_scaler0": {"loss_scale": 1024.0, "unskipped": 7}})
    assert amp.state_dict()["loss_scaler0"] == {"loss_scale": 1024.0, "unskipped": 7}


def test_master_params_materialises_unscaled_fp32_grads_for_clipping():
    from apex import amp
    from apex.optimizers import FusedAdam
    model = _mlp(9)
    opt = FusedAdam(model.parameters(), lr=1e-3)
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    x = torch.randn(16, 32, device="cuda")
    loss = model(x).pow(2).mean()
    with amp.scale_loss(loss, opt) as scaled:
        scaled.backward()
    model_grads = [p.grad.detach().clone() for p in model.parameters()]
    masters = list(amp.master_params(opt))
    for m, g in zip(masters, model_grads):
        assert m.grad is not None and m.grad.dtype == torch.float32
        assert torch.equal(m.grad, g.float() * (1.0 / 65536.0))
    torch.nn.utils.clip_grad_norm_(masters, 0.1)
    before = [m.detach().clone() for m in masters]
    opt.step()
    torch.cuda.synchronize()
    assert all(not torch.equal(a, m.detach()) for a, m in zip(before, masters))


def test_amp_api_misuse_raises():
    from apex import amp
    from apex.optimizers import FusedAdam
    model = _mlp(2)
    opt = FusedAdam(model.parameters())
    with pytest.raises(RuntimeError, match="has not been initialized"):
        with amp.scale_loss(torch.zeros((), device="cuda"), opt):
            pass
    with pytest.raises(RuntimeError, match="Unexpected optimization level"):
        amp.initialize(model, opt, opt_level="02")
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    with pytest.raises(RuntimeError, match="closure"):
        opt.step(closure=lambda: None)
    with pytest.raises(RuntimeError, match="only be passed through amp.initialize once"):
        amp.initialize(_mlp(3), opt, opt_level="O2", verbosity=0)
    m2, o2 = amp.initialize(_mlp(4), FusedAdam(_mlp(4).parameters()), enabled=False)
    assert next(m2.parameters()).dtype == torch.float32


def test_amp_c_compat_through_multi_tensor_applier():
    import amp_C
    from apex.multi_tensor_apply import multi_tensor_applier
    gen = torch.Generator().manual_seed(5)
    hs = [(torch.randn(n, generator=gen) * 50).half() for n in (100, 4097, 7)]
    src = [h.cuda() for h in hs]
    dst = [torch.empty(h.shape, device="cuda") for h in hs]
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    multi_tensor_applier(amp_C.multi_tensor_scale, flag, [src, dst], 1.0 / 128)
    assert int(flag.item()) == 0
    for d, h in zip(dst, hs):
        assert torch.equal(d.cpu(), h.float() * (1.0 / 128))
    src[1][-1] = float("nan")
    multi_tensor_applier(amp_C.multi_tensor_scale, flag, [src, dst], 1.0)
    assert int(flag.item()) == 1
    # adam / sgd early-out on a raised noop flag, like apex's `if(*noop) return`
    p = [torch.ones(64, device="cuda")]
    g = [torch.ones(64, device="cuda")]
    m, v = [torch.zeros(64, device="cuda")], [torch.zeros(64, device="cuda")]
    multi_tensor_applier(amp_C.multi_tensor_adam, flag, [g, p, m, v], 1e-3, 0.9, 0.999, 1e-8, 1, 1, 1, 0.0)
    multi_tensor_applier(amp_C.multi_tensor_sgd, flag, [g, p, m], 0.0, 0.9, 0.0, 0.1, False, True, False, 1.0)
    assert torch.equal(p[0], torch.ones(64, device="cuda"))
    flag.zero_()
    multi_tensor_applier(amp_C.multi_tensor_sgd, flag, [g, p, m], 0.0, 0.9, 0.0, 0.1, False, True, False, 1.0)
    assert torch.allclose(p[0], torch.full((64,), 0.9, device="cuda")) and torch.equal(m[0], g[0])
    out = [torch.zeros(64, device="cuda")]
    multi_tensor_applier(amp_C.multi_tensor_axpby, flag, [g, p, out], 2.0, 1.0, -1)
    assert torch.allclose(out[0], torch.full((64,), 2.9, device="cuda"))


def test_resnet50_o2_fused_sgd_two_dtype_sets():
    """BASELINE config 3 in miniature: ResNet-50 at O2 keeps its 106 BatchNorm tensors in fp32, so the
    optimizer runs two launch sets — (fp16 grad, fp32 master, fp16 copy) and (fp32, fp32, none)."""
    torchvision = pytest.importorskip("torchvision")
    from apex import amp
    from apex.optimizers import FusedSGD
    from apex_example_b200 import ops
    torch.manual_seed(0)
    model = torchvision.models.resnet50().cuda()
    opt = FusedSGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4, materialize_master_grads=False)
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    n16 = sum(p.numel() for p in model.parameters() if p.dtype == torch.float16)
    n32 = sum(p.numel() for p in model.parameters() if p.dtype == torch.float32)
    assert n16 + n32 == 25_557_032 and n32 == 53_120                     # SURVEY.md §8 sizes
    x = torch.randn(8, 3, 64, 64, device="cuda")
    y = torch.randint(0, 1000, (8,), device="cuda")
    bn_before = model.bn1.weight.detach().clone()
    conv_before = model.conv1.weight.detach().clone()
    launches = []
    for _ in range(14):                     # the first steps overflow at scale 2**16 and are skipped: the usual warm-up
        loss = torch.nn.functional.cross_entropy(model(x), y)
        opt.zero_grad()
        l0 = ops.launch_count()
        with amp.scale_loss(loss, opt) as scaled:
            scaled.backward()
        opt.step()
        launches.append(ops.launch_count() - l0)
    torch.cuda.synchronize()
    assert torch.isfinite(loss)
    # check fp16 + check fp32 + scaler update + one SGD launch per dtype set (+ the scaler's one-time init)
    assert launches[0] in (5, 6) and set(launches[1:]) == {5}
    assert opt.state_dict()["param_groups"][0]["step"] >= 1          # counts only the steps that really ran
    assert not torch.equal(model.bn1.weight, bn_before) and model.bn1.weight.dtype == torch.float32
    assert not torch.equal(model.conv1.weight, conv_before) and model.conv1.weight.dtype == torch.float16
    stash = opt._amp_stash
    for master, model_p in stash.model_of.items():
        assert master.dtype == torch.float32 and torch.equal(model_p, master.half())
        assert "momentum_buffer" in opt.state[master] and opt.state[master]["momentum_buffer"].dtype == torch.float32


# ------------------------------------------------------------------------------------------------
# checkpoint surface, gradient accumulation
# ------------------------------------------------------------------------------------------------
def _run_steps(model, opt, xs, start, stop, overflow=()):
    from apex import amp
    for it in range(start, stop):
        x = xs[it] * (float("inf") if it in overflow else 1.0)
        loss = model(x).float().pow(2).mean()
        opt.zero_grad()
        with amp.scale_loss(loss, opt) as scaled:
            scaled.backward()
        opt.step()


@pytest.mark.parametrize("which", ["adam", "sgd", "torch_adamw", "torch_sgd"])
def test_checkpoint_round_trip_continues_bit_exactly(which):
    """optimizer.state_dict() + amp.state_dict() + model.state_dict() (fp32 at O2) restore a run so that
    it continues exactly like the uninterrupted one — including the skipped steps' effect on the step
    counter and on the loss scale."""
    from apex import amp
    from apex.optimizers import FusedAdam, FusedSGD
    from apex_example_b200.amp._amp_state import _amp_state

    def build():
        _amp_state.reset()
        model = _mlp(21)
        if which == "adam":
            opt = FusedAdam(model.parameters(), lr=1e-2, weight_decay=0.01)
        elif which == "sgd":
            opt = FusedSGD(model.parameters(), lr=1e-2, momentum=0.9, materialize_master_grads=False)
        elif which == "torch_adamw":            # stock torch optimizers adopted by the fused kernels
            opt = torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.01)
        else:
            opt = torch.optim.SGD(model.parameters(), lr=1e-2, momentum=0.9)
        return amp.initialize(model, opt, opt_level="O2", verbosity=0)

    torch.manual_seed(5)
    xs = [torch.randn(16, 32, device="cuda") for _ in range(10)]
    overflow = {0, 3}
    model_a, opt_a = build()
    _run_steps(model_a, opt_a, xs, 0, 10, overflow)
    ref = [p.detach().clone() for p in amp.master_params(opt_a)]
    ref_scale = amp.state_dict()

    model_b, opt_b = build()
    _run_steps(model_b, opt_b, xs, 0, 6, overflow)
    ckpt = {"model": model_b.state_dict(), "opt": opt_b.state_dict(), "amp": amp.state_dict()}
    assert all(v.dtype != torch.float16 for v in ckpt["model"].values())
    assert ckpt["opt"]["param_groups"][0]["step"] == 4              # 6 iterations, 2 skipped
    model_c, opt_c = build()
    model_c.load_state_dict(ckpt["model"])
    opt_c.load_state_dict(ckpt["opt"])
    amp.load_state_dict(ckpt["amp"])
    # model.load_state_dict() handed the checkpoint's exact fp32 values to the masters (pre-hook installed by
    # amp.initialize): nothing else to restore
    for m_c, m_b in zip(amp.master_params(opt_c), amp.master_params(opt_b)):
        assert torch.equal(m_c.detach(), m_b.detach())
    _run_steps(model_c, opt_c, xs, 6, 10, overflow)
    torch.cuda.synchronize()
    for a, c in zip(ref, amp.master_params(opt_c)):
        assert torch.equal(a, c.detach())
    assert amp.state_dict() == ref_scale
    assert opt_c.state_dict()["param_groups"][0]["step"] == opt_a.state_dict()["param_groups"][0]["step"] == 8


@pytest.mark.parametrize("which", ["fused_adam", "torch_sgd"])
def test_gradient_accumulation_with_delay_unscale(which):
    """Two micro-batches: the first backward runs under delay_unscale=True (gradients stay scaled and
    accumulate), the second unscales and checks once.  Must equal one step on the summed gradient."""
    from apex import amp
    from apex.optimizers import FusedAdam
    from apex_example_b200.amp._amp_state import _amp_state
    torch.manual_seed(8)
    x1, x2 = torch.randn(16, 32, device="cuda"), torch.randn(16, 32, device="cuda")

    def build():
        _amp_state.reset()
        model = _mlp(33)
        opt = (FusedAdam(model.parameters(), lr=1e-2, set_grad_none=False) if which == "fused_adam"
               else torch.optim.SGD(model.parameters(), lr=1e-2, momentum=0.9, foreach=False))
        return amp.initialize(model, opt, opt_level="O2", loss_scale=1024.0, verbosity=0)

    model_a, opt_a = build()
    opt_a.zero_grad()
    with amp.scale_loss(model_a(x1).float().pow(2).mean(), opt_a, delay_unscale=True) as s1:
        s1.backward()
    with amp.scale_loss(model_a(x2).float().pow(2).mean(), opt_a) as s2:
        s2.backward()
    opt_a.step()

    model_b, opt_b = build()
    opt_b.zero_grad()
    loss = model_b(x1).float().pow(2).mean() + model_b(x2).float().pow(2).mean()
    with amp.scale_loss(loss, opt_b) as s:
        s.backward()
    opt_b.step()
    torch.cuda.synchronize()
    for a, b in zip(amp.master_params(opt_a), amp.master_params(opt_b)):
        # fp16 accumulation of two scaled gradients vs one backward of the sum: equal up to fp16 rounding of the
        # gradient (2^-11 relative), which Adam/SGD turn into at most lr * 2^-10 here
        torch.testing.assert_close(a.detach(), b.detach(), rtol=0, atol=2e-5)
    assert any(not torch.equal(a.detach(), p.detach().float()) or True for a, p in zip(amp.master_params(opt_a), model_a.parameters()))


# ------------------------------------------------------------------------------------------------
# l2norm + gradient clipping, fused
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16, torch.float32])
def test_multi_tensor_l2norm_matches_torch(dt):
    import amp_C
    from apex.multi_tensor_apply import multi_tensor_applier
    gen = torch.Generator().manual_seed(3)
    tensors = [(torch.randn(n, generator=gen) * 3).to(dt).cuda() for n in (1, 7, 4096, 70001, 123457, 9)]
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    norm, _ = multi_tensor_applier(amp_C.multi_tensor_l2norm, flag, [tensors], False)
    want = torch.linalg.vector_norm(torch.cat([t.double() for t in tensors]))
    # fp32 accumulation in a fixed tree vs an fp64 reference: 1e-6 relative
    assert abs(float(norm) - float(want)) <= 1e-6 * float(want) and int(flag.item()) == 0
    again, _ = multi_tensor_applier(amp_C.multi_tensor_l2norm, flag, [tensors], False)
    assert float(again) == float(norm)                      # deterministic: ordered partial sums, no float atomics
    tensors[3][-1] = float("inf")
    multi_tensor_applier(amp_C.multi_tensor_l2norm, flag, [tensors], False)
    assert int(flag.item()) == 1


@pytest.mark.parametrize("max_norm", [0.05, 1e6])
def test_fused_clip_grad_norm_matches_torch_clip_on_master_grads(max_norm):
    """opt.clip_grad_norm_(max_norm) + step() == clip_grad_norm_(amp.master_params(opt), max_norm) + step()."""
    from apex import amp
    from apex.optimizers import FusedAdam
    from apex_example_b200.amp._amp_state import _amp_state
    torch.manual_seed(12)
    x = torch.randn(32, 32, device="cuda")

    def run(fused_clip):
        _amp_state.reset()
        model = _mlp(44)
        opt = FusedAdam(model.parameters(), lr=1e-2, weight_decay=0.01)
        model, opt = amp.initialize(model, opt, opt_level="O2", loss_scale=4096.0, verbosity=0)
        norms = []
        for _ in range(3):
            loss = model(x).float().pow(2).mean()
            opt.zero_grad()
            with amp.scale_loss(loss, opt) as scaled:
                scaled.backward()
            if fused_clip:
                norms.append(float(opt.clip_grad_norm_(max_norm)))
            else:
                norms.append(float(torch.nn.utils.clip_grad_norm_(list(amp.master_params(opt)), max_norm)))
            opt.step()
        torch.cuda.synchronize()
        return [m.detach().clone() for m in amp.master_params(opt)], norms

    a, na = run(True)
    b, nb = run(False)
    for x_, y_ in zip(na, nb):
        assert abs(x_ - y_) <= 1e-5 * abs(y_)
    if max_norm < 1:
        assert na[0] > max_norm                          # the clip is actually active
    for p, q in zip(a, b):
        # same gradients, same clip rule; the fused path multiplies by (inv_scale*coef) once instead of twice
        torch.testing.assert_close(p, q, rtol=0, atol=2e-6)


def test_bf16_o2_training_runs_with_bf16_model_and_fp32_masters():
    """north_star's addition to apex: cast_model_type=torch.bfloat16 (static scale 1 is the usual choice)."""
    from apex import amp
    from apex.optimizers import FusedAdam
    model = _mlp(55)
    opt = FusedAdam(model.parameters(), lr=1e-2)
    model, opt = amp.initialize(model, opt, opt_level="O2", cast_model_type=torch.bfloat16, loss_scale=1.0, verbosity=0)
    assert all(p.dtype == torch.bfloat16 for p in model.parameters())
    x = torch.randn(32, 32, device="cuda")
    first = None
    for _ in range(20):
        loss = model(x).float().pow(2).mean()
        first = first if first is not None else float(loss)
        opt.zero_grad()
        with amp.scale_loss(loss, opt) as scaled:
            scaled.backward()
        opt.step()
    torch.cuda.synchronize()
    assert float(loss) < first
    for master, model_p in opt._amp_stash.model_of.items():
        assert master.dtype == torch.float32 and torch.equal(model_p, master.bfloat16())


# ------------------------------------------------------------------------------------------------
# less-travelled paths
# ------------------------------------------------------------------------------------------------
def test_step_window_stress_200_unsynchronised_steps_with_random_skips():
    """The host runs far ahead of the device (tiny tensors, no sync): the step counter must fall back to a
    sync when more than B200_STEP_WINDOW-1 launches are unobserved, and the result must still equal the
    oracle's sequential semantics bit for bit."""
    import random
    from apex import amp
    from apex.optimizers import FusedAdam
    rng = random.Random(3)
    gen = torch.Generator().manual_seed(6)
    p = torch.nn.Parameter((torch.randn(257, generator=gen) * 0.1).cuda())

    class Holder(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.p = p
    model = Holder()
    opt = FusedAdam(model.parameters(), lr=1e-2, set_grad_none=False)
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    hp = model.p.detach().float().cpu().clone()
    hm, hv, hc = torch.zeros(257), torch.zeros(257), hp.half()
    sc = oracle.LossScaler()
    dummy = torch.zeros((), device="cuda")
    gbuf = torch.zeros(257, dtype=torch.float16, device="cuda")
    model.p.grad = gbuf
    good = 0
    for it in range(200):
        bad = rng.random() < 0.15
        g = (torch.randn(257, generator=gen) * 1e-3 * sc.loss_scale).half()
        if bad:
            g[rng.randrange(257)] = float("inf")
        gbuf.copy_(g, non_blocking=True)
        with amp.scale_loss(dummy, opt):
            pass
        opt.step()
        if not sc.update_scale(bad):
            good += 1
            oracle.adam(g, hp, hm, hv, hc, lr=1e-2, step=good, inv_scale=sc.inv_scale)
    torch.cuda.synchronize()
    master = next(amp.master_params(opt))
    assert torch.equal(master.detach().cpu(), hp) and torch.equal(model.p.detach().cpu(), hc)
    assert opt.state_dict()["param_groups"][0]["step"] == good
    assert amp._amp_state.loss_scalers[0].loss_scale() == sc.loss_scale


def test_two_losses_two_optimizers_and_frozen_parameters():
    from apex import amp
    from apex.optimizers import FusedAdam, FusedSGD
    net_a, net_b = _mlp(61), _mlp(62)
    for q in list(net_b.parameters())[:2]:
        q.requires_grad_(False)                                  # frozen: never gets a gradient
    opt_a = FusedAdam(net_a.parameters(), lr=1e-2)
    opt_b = FusedSGD([q for q in net_b.parameters() if q.requires_grad], lr=1e-2, momentum=0.9,
                     materialize_master_grads=False)
    (net_a, net_b), (opt_a, opt_b) = amp.initialize([net_a, net_b], [opt_a, opt_b], opt_level="O2", num_losses=2,
                                                    verbosity=0)
    assert len(amp._amp_state.loss_scalers) == 2
    x = torch.randn(16, 32, device="cuda")
    frozen = [q.detach().clone() for q in list(net_b.parameters())[:2]]
    before_a = [q.detach().clone() for q in net_a.parameters()]
    before_b = [q.detach().clone() for q in list(net_b.parameters())[2:]]
    for it in range(4):
        la = net_a(x).float().pow(2).mean()
        lb = net_b(x * (float("inf") if it == 0 else 1.0)).float().pow(2).mean()      # only loss 1 overflows
        opt_a.zero_grad()
        opt_b.zero_grad()
        with amp.scale_loss(la, opt_a, loss_id=0) as s:
            s.backward()
        with amp.scale_loss(lb, opt_b, loss_id=1) as s:
            s.backward()
        opt_a.step()
        opt_b.step()
    torch.cuda.synchronize()
    assert [s.loss_scale() for s in amp._amp_state.loss_scalers] == [65536.0, 32768.0]
    assert all(torch.equal(a, b.detach()) for a, b in zip(frozen, list(net_b.parameters())[:2]))
    assert all(not torch.equal(a, b.detach()) for a, b in zip(before_a, net_a.parameters()))
    assert all(not torch.equal(a, b.detach()) for a, b in zip(before_b, list(net_b.parameters())[2:]))
    assert opt_a.state_dict()["param_groups"][0]["step"] == 4 and opt_b.state_dict()["param_groups"][0]["step"] == 3


def test_add_param_group_after_steps_starts_its_own_step_count():
    from apex.optimizers import FusedAdam
    gen = torch.Generator().manual_seed(9)
    a = torch.nn.Parameter((torch.randn(100, generator=gen)).cuda())
    b = torch.nn.Parameter((torch.randn(50, generator=gen)).cuda())
    opt = FusedAdam([a], lr=1e-2, weight_decay=0.0)
    ha, hb = a.detach().cpu().clone(), b.detach().cpu().clone()
    st = {k: torch.zeros(n) for k, n in (("ma", 100), ("va", 100), ("mb", 50), ("vb", 50))}
    for step in range(1, 6):
        ga = torch.randn(100, generator=gen)
        a.grad = ga.cuda()
        if step == 3:
            opt.add_param_group({"params": [b], "lr": 1e-3})
        if step >= 3:
            gb = torch.randn(50, generator=gen)
            b.grad = gb.cuda()
            oracle.adam(gb, hb, st["mb"], st["vb"], None, lr=1e-3, step=step - 2, adam_w_mode=True)
        opt.step()
        oracle.adam(ga, ha, st["ma"], st["va"], None, lr=1e-2, step=step, adam_w_mode=True)
    torch.cuda.synchronize()
    assert torch.equal(a.detach().cpu(), ha) and torch.equal(b.detach().cpu(), hb)
    assert [g["step"] for g in opt.param_groups] == [5, 3]


# ------------------------------------------------------------------------------------------------
# round 2: gradient accumulation without delay_unscale, sticky skips, adopted torch optimizers
# ------------------------------------------------------------------------------------------------
class _PlainSGD(torch.optim.SGD):
    """A subclass is never adopted by the fused kernels: it exercises amp's generic path."""


def _make_opt(which, params, lr=1e-2):
    from apex.optimizers import FusedAdam, FusedSGD
    if which == "fused_adam":
        return FusedAdam(params, lr=lr, set_grad_none=False)
    if which == "fused_sgd":
        return FusedSGD(params, lr=lr, momentum=0.9, materialize_master_grads=False)
    if which == "torch_sgd":
        return torch.optim.SGD(params, lr=lr, momentum=0.9)
    if which == "torch_adamw":
        return torch.optim.AdamW(params, lr=lr, weight_decay=0.01)
    if which == "generic_sgd":
        return _PlainSGD(params, lr=lr, momentum=0.9, foreach=False)
    raise ValueError(which)


def _bn_mlp(seed):
    torch.manual_seed(seed)           # BatchNorm stays fp32 at O2: covers the fp32-parameter half of the path
    return torch.nn.Sequential(torch.nn.Linear(32, 64), torch.nn.BatchNorm1d(64), torch.nn.ReLU(),
                               torch.nn.Linear(64, 8)).cuda()


@pytest.mark.parametrize("level", ["O2", "O1"])
@pytest.mark.parametrize("which", ["fused_adam", "fused_sgd", "torch_sgd", "torch_adamw", "generic_sgd"])
def test_gradient_accumulation_without_delay_unscale(which, level):
    """Two scale_loss blocks, no zero_grad between them, no delay_unscale: each block is
    unscaled on its own and merged, so the result must equal ONE backward of the summed loss."""
    from apex import amp
    from apex_example_b200.amp._amp_state import _amp_state
    torch.manual_seed(8)
    x1, x2 = torch.randn(16, 32, device="cuda"), torch.randn(16, 32, device="cuda")

    def build():
        _amp_state.reset()
        model = _bn_mlp(33)
        model.eval()                  # fixed BatchNorm statistics: both arrangements see the same function
        for p in model.parameters():
            p.requires_grad_(True)
        return amp.initialize(model, _make_opt(which, model.parameters()), opt_level=level, loss_scale=1024.0, verbosity=0)

    def fwd(model, x):
        return model(x).float().pow(2).mean()

    model_a, opt_a = build()
    opt_a.zero_grad()
    with amp.scale_loss(fwd(model_a, x1), opt_a) as s1:
        s1.backward()
    with amp.scale_loss(fwd(model_a, x2), opt_a) as s2:
        s2.backward()
    opt_a.step()

    model_b, opt_b = build()
    opt_b.zero_grad()
    with amp.scale_loss(fwd(model_b, x1) + fwd(model_b, x2), opt_b) as s:
        s.backward()
    opt_b.step()
    torch.cuda.synchronize()
    moved = False
    for a, b, p0 in zip(amp.master_params(opt_a), amp.master_params(opt_b), _bn_mlp(33).parameters()):
        # the wrong merge (2*g1 + g2, or a gradient left scaled) is off by ~lr*|g| = 1e-3..1e-2; rounding of the
        # fp16 gradient sum is 2^-11 relative -> <= 2e-5 after one lr=1e-2 step
        torch.testing.assert_close(a.detach().float(), b.detach().float(), rtol=0, atol=3e-5)
        moved |= not torch.equal(a.detach().float(), p0.detach().float())
    assert moved


@pytest.mark.parametrize("which", ["fused_adam", "torch_sgd", "generic_sgd"])
def test_overflow_in_first_micro_batch_skips_the_accumulated_step(which):
    """apex patches step() into a one-shot skip at the scale_loss exit that saw the overflow and leaves it
    there until step() is called; a clean second backward must not un-skip the step."""
    from apex import amp
    model = _bn_mlp(5)
    model, opt = amp.initialize(model, _make_opt(which, model.parameters()), opt_level="O2", verbosity=0)
    x = torch.randn(16, 32, device="cuda")
    before = [p.detach().clone() for p in model.parameters()]
    opt.zero_grad()
    with amp.scale_loss(model(x * float("inf")).float().pow(2).mean(), opt) as s:
        s.backward()
    with amp.scale_loss(model(x).float().pow(2).mean(), opt) as s:
        s.backward()
    opt.step()
    torch.cuda.synchronize()
    assert all(torch.equal(a, p.detach()) for a, p in zip(before, model.parameters())), "the poisoned step ran"
    assert amp._amp_state.loss_scalers[0].loss_scale() == 2.0 ** 15
    # and the next ordinary step runs again
    opt.zero_grad()
    with amp.scale_loss(model(x).float().pow(2).mean(), opt) as s:
        s.backward()
    opt.step()
    torch.cuda.synchronize()
    assert any(not torch.equal(a, p.detach()) for a, p in zip(before, model.parameters()))


@pytest.mark.parametrize("which", ["torch_sgd", "torch_adamw"])
def test_stock_torch_optimizer_takes_the_fused_sync_free_path(which):
    """The reference's own configuration (torch.optim.SGD at test_apex_distributed_spawn.py:119): same launch
    count as FusedSGD/FusedAdam, no host synchronisation anywhere in the step, torch-compatible state_dict."""
    from apex import amp
    from apex_example_b200 import ops
    model = _mlp(3)
    opt = _make_opt(which, model.parameters())
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    assert getattr(opt, "_b200_engine", None) is not None
    xs = [torch.randn(8, 32, device="cuda") for _ in range(6)]

    def one(x):
        loss = model(x).float().pow(2).mean()
        opt.zero_grad()
        with amp.scale_loss(loss, opt) as s:
            s.backward()
        opt.step()

    one(xs[0])
    one(xs[1])                         # tables and plans are built by now
    torch.cuda.synchronize()
    l0 = ops.launch_count()
    torch.cuda.set_sync_debug_mode("error")
    try:
        for x in xs[2:]:
            one(x)
    finally:
        torch.cuda.set_sync_debug_mode("default")
    torch.cuda.synchronize()
    assert (ops.launch_count() - l0) / 4 == 3          # check + scaler update + ONE fused step launch
    sd = opt.state_dict()
    fresh = type(opt)([torch.nn.Parameter(p.detach().clone()) for p in amp.master_params(opt)], lr=1e-2,
                      **({"momentum": 0.9} if which == "torch_sgd" else {"weight_decay": 0.01}))
    fresh.load_state_dict(sd)                          # an unpatched torch optimizer accepts it
    if which == "torch_adamw":
        steps = {float(st["step"]) for st in fresh.state.values()}
        assert steps == {6.0}
        assert all("exp_avg" in st and "exp_avg_sq" in st for st in fresh.state.values())
    else:
        assert all("momentum_buffer" in st for st in fresh.state.values())


def test_fused_sgd_equals_torch_sgd_on_fp32_params_nesterov():
    """O0-like use (no amp): FusedSGD engine vs torch.optim.SGD(foreach=False) on identical fp32 inputs."""
    from apex.optimizers import FusedSGD
    gen = torch.Generator().manual_seed(4)
    ps = [torch.randn(n, generator=gen).cuda() for n in (1000, 17, 4096)]
    a = [torch.nn.Parameter(p.clone()) for p in ps]
    b = [torch.nn.Parameter(p.clone()) for p in ps]
    oa = FusedSGD(a, lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True)
    ob = torch.optim.SGD(b, lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True, foreach=False)
    for _ in range(4):
        gs = [torch.randn(p.shape, generator=gen).cuda() for p in ps]
        for x, y, g in zip(a, b, gs):
            x.grad, y.grad = g.clone(), g.clone()
        oa.step()
        ob.step()
    for x, y in zip(a, b):            # vs torch's CUDA kernels (which may fuse multiply-adds): a few fp32 ulps at |p| ~ 1
        torch.testing.assert_close(x.detach(), y.detach(), rtol=1e-6, atol=1e-6)


def test_fused_sgd_loads_a_torch_sgd_checkpoint_without_resetting_momentum():
    from apex.optimizers import FusedSGD
    gen = torch.Generator().manual_seed(9)
    p0 = torch.randn(5000, generator=gen).cuda()
    a, b = torch.nn.Parameter(p0.clone()), torch.nn.Parameter(p0.clone())
    ta = torch.optim.SGD([a], lr=0.1, momentum=0.9, foreach=False)
    for _ in range(3):
        a.grad = torch.randn(5000, generator=gen).cuda()
        ta.step()
    sd = copy.deepcopy(ta.state_dict())       # load_state_dict keeps same-dtype tensors by reference: do not share buffers
    assert "step" not in sd["param_groups"][0]
    b.data.copy_(a.data)
    fb = FusedSGD([b], lr=0.1, momentum=0.9)
    fb.load_state_dict(sd)
    g = torch.randn(5000, generator=gen).cuda()
    a.grad, b.grad = g.clone(), g.clone()
    ta.step()
    fb.step()
    # a reset momentum buffer (mom = grad instead of 0.9*mom + grad) would be off by ~|mom| ~ 1
    torch.testing.assert_close(b.detach(), a.detach(), rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(fb.state[b]["momentum_buffer"], ta.state[a]["momentum_buffer"], rtol=1e-6, atol=1e-6)


def test_plan_notices_moved_parameter_storage_on_the_very_next_step():
    """model.half()/.to()/EMA swaps re-point p.data while the gradient object stays the same."""
    from apex.optimizers import FusedAdam
    p = torch.nn.Parameter(torch.randn(5000, device="cuda"))
    opt = FusedAdam([p], lr=1e-2, set_grad_none=False)
    p.grad = torch.randn(5000, device="cuda")
    opt.step()
    opt.step()
    old = p.data
    keep = old.clone()
    p.data = old.clone()               # new storage, same values; the grad object is untouched
    opt.step()
    torch.cuda.synchronize()
    assert torch.equal(old, keep), "the step wrote through a stale address"
    assert not torch.equal(p.data, keep)


@pytest.mark.parametrize("grad_layout", ["channels_last", "nchw"])
def test_channels_last_parameters_are_stepped_in_their_own_memory_order(grad_layout):
    """The kernels walk memory order; a channels_last convolution weight (and its state) is dense but not
    row-major.  Whatever layout the gradient arrives in, the result must equal torch.optim.SGD's."""
    from apex.optimizers import FusedSGD
    gen = torch.Generator().manual_seed(12)
    w = torch.randn(32, 16, 3, 3, generator=gen).cuda().contiguous(memory_format=torch.channels_last)
    a, b = torch.nn.Parameter(w.clone()), torch.nn.Parameter(w.clone())
    assert not a.is_contiguous() and a.stride() == w.stride()
    oa = FusedSGD([a], lr=0.1, momentum=0.9, weight_decay=1e-4)
    ob = torch.optim.SGD([b], lr=0.1, momentum=0.9, weight_decay=1e-4, foreach=False)
    for _ in range(3):
        g = torch.randn(32, 16, 3, 3, generator=gen).cuda()
        if grad_layout == "channels_last":
            g = g.contiguous(memory_format=torch.channels_last)
        a.grad, b.grad = g.clone(), g.clone()
        oa.step()
        ob.step()
    # a wrong element pairing between the layouts would be off by ~|g| ~ 1; fp32 rounding is ~1e-7 here
    torch.testing.assert_close(a.detach(), b.detach(), rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(oa.state[a]["momentum_buffer"], ob.state[b]["momentum_buffer"], rtol=1e-6, atol=1e-6)
