"""-m gpu tests of the per-tensor L2 norms and FusedLAMB: b200_l2norm_per_tensor,
b200_lamb_stage1, b200_lamb_stage2 through the C ABI, bit-exact against oracle/ (whose summation order is part of the
contract), and apex.optimizers.FusedLAMB against a straight torch restatement of the published rule (tolerance written
below).  Parity against apex itself is unpinned: apex is not vendored in the reference."""
import pytest
import torch

import oracle
from apex_example_b200 import ops
from apex_example_b200.tables import TensorTable
from gpu_util import ADVERSARIAL_SIZES, assert_lists_bit_equal

pytestmark = pytest.mark.gpu

SIZES = [s for s in ADVERSARIAL_SIZES if s] + [40000]


def _norm_bufs(table):
    dev = table.dev.device
    return (table.first_tiles(), torch.empty(max(table.n_tiles, 1), device=dev), torch.zeros(table.n_tensors, device=dev),
            torch.zeros(table.n_tensors, device=dev), torch.zeros(1, device=dev))


@pytest.mark.parametrize("dt", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("tile", [4096, 8192])
def test_per_tensor_norms_bit_exact(dt, tile):
    gen = torch.Generator().manual_seed(5)
    hx = [(torch.randn(n, generator=gen) * 3).to(dt) for n in SIZES]
    dx = [h.cuda() for h in hx]
    table = TensorTable([dx], tile_elems=tile)
    first, tile_ss, ss, norms, total = _norm_bufs(table)
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    total.fill_(1.5)                                          # the total ACCUMULATES
    ops.l2norm_per_tensor(table, 0, dt, first, tile_ss, ss, norms=norms, total=total, dev_flag=flag)
    want_ss, want_total, want_flag = oracle.l2norm_per_tensor(hx, tile)
    assert torch.equal(ss.cpu(), want_ss)
    assert torch.equal(norms.cpu(), want_ss.sqrt())
    assert total.item() == (torch.tensor(1.5) + torch.tensor(want_total)).item()
    assert flag.item() == 0 and want_flag == 0
    # sanity against double precision
    ref = torch.stack([h.double().square().sum() for h in hx])
    torch.testing.assert_close(ss.cpu().double(), ref, rtol=1e-5, atol=0)
    # inf/NaN raises the flag
    dx[3][-1] = float("nan")
    ops.l2norm_per_tensor(table, 0, dt, first, tile_ss, ss, dev_flag=flag)
    assert flag.item() == 1


def test_amp_c_l2norm_per_tensor():
    import amp_C
    from apex.multi_tensor_apply import multi_tensor_applier
    gen = torch.Generator().manual_seed(6)
    hx = [torch.randn(n, generator=gen) for n in (7, 5000, 1, 12345)]
    dx = [h.cuda() for h in hx]
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    norm, per = multi_tensor_applier(amp_C.multi_tensor_l2norm, flag, [dx], True)
    torch.testing.assert_close(per.cpu(), torch.stack([h.norm() for h in hx]), rtol=1e-5, atol=0)
    torch.testing.assert_close(norm.cpu()[0], torch.cat(hx).norm(), rtol=1e-5, atol=0)
    assert flag.item() == 0


def _lamb_case(gdt, sdt, cdt, hyper, max_norm, steps=3, tile=4096, inv_scale=1.0 / 256):
    gen = torch.Generator().manual_seed(11)
    hp = [(torch.randn(n, generator=gen) * 0.05).to(sdt) for n in SIZES]
    hm = [torch.zeros(n, dtype=sdt) for n in SIZES]
    hv = [torch.zeros(n, dtype=sdt) for n in SIZES]
    hc = [h.to(cdt) for h in hp] if cdt is not None else None
    dp, dm, dv = ([h.cuda() for h in lst] for lst in (hp, hm, hv))
    dc = [h.cuda() for h in hc] if hc is not None else None
    dg = [torch.zeros(n, dtype=gdt, device="cuda") for n in SIZES]
    du = [torch.zeros(n, device="cuda") for n in SIZES]
    t1 = TensorTable([dg, dp, dm, dv, du], tile_elems=tile)
    t2 = TensorTable([du, dp] + ([dc] if dc is not None else []), tile_elems=tile)
    first, tile_ss, ss, pn, _ = _norm_bufs(t1)
    un = torch.zeros_like(pn)
    gsum = torch.zeros(1, device="cuda")
    for step in range(1, steps + 1):
        hg = [(torch.randn(n, generator=gen) * 0.02 / inv_scale).to(gdt) for n in SIZES]
        for d, h in zip(dg, hg):
            d.copy_(h)
        grp = [ops.LambGroup(hyper["lr"], 0.9, 0.999, 1e-6, hyper["weight_decay"], step, 1, hyper["adam_w_mode"],
                             hyper.get("grad_averaging", 1), hyper.get("use_nvlamb", 0))]
        gsum.zero_()
        ops.l2norm_per_tensor(t1, 0, gdt, first, tile_ss, ss, total=gsum)
        ops.l2norm_per_tensor(t1, 1, sdt, first, tile_ss, ss, norms=pn)
        ops.lamb_stage1(t1, gdt, sdt, grp, gsum, max_norm, inv_scale=inv_scale)
        ops.l2norm_per_tensor(t1, 4, torch.float32, first, tile_ss, ss, norms=un)
        ops.lamb_stage2(t2, sdt, cdt, grp, pn, un)
        _, want_gsum, _ = oracle.l2norm_per_tensor(hg, tile)
        assert gsum.item() == want_gsum
        oracle.lamb_list(hg, hp, hm, hv, hc, tile_elems=tile, grad_sumsq=want_gsum, inv_scale=inv_scale,
                         max_grad_norm=max_norm, step=step, eps=1e-6, **hyper)
        assert_lists_bit_equal(dm, hm, f"m step {step}")
        assert_lists_bit_equal(dv, hv, f"v step {step}")
        assert_lists_bit_equal(dp, hp, f"p step {step}")
        if dc is not None:
            assert_lists_bit_equal(dc, hc, f"copy step {step}")


@pytest.mark.parametrize("gdt,sdt,cdt", [(torch.float16, torch.float32, torch.float16),
                                         (torch.bfloat16, torch.float32, torch.bfloat16),
                                         (torch.float32, torch.float32, None),
                                         (torch.float16, torch.float16, None)])
@pytest.mark.parametrize("adam_w_mode", [1, 0])
def test_lamb_stages_bit_exact(gdt, sdt, cdt, adam_w_mode):
    _lamb_case(gdt, sdt, cdt, dict(lr=2e-2, weight_decay=0.01, adam_w_mode=adam_w_mode), max_norm=1.0)


@pytest.mark.parametrize("hyper,max_norm", [(dict(lr=1e-2, weight_decay=0.0, adam_w_mode=1), 0.0),
                                            (dict(lr=1e-2, weight_decay=0.0, adam_w_mode=1, use_nvlamb=1), 0.5),
                                            (dict(lr=1e-2, weight_decay=0.1, adam_w_mode=0, grad_averaging=0), 1e9)])
def test_lamb_variants_bit_exact(hyper, max_norm):
    _lamb_case(torch.float32, torch.float32, None, hyper, max_norm, tile=8192)


def _torch_lamb_step(ps, gs, ms, vs, step, lr, wd, max_norm, eps=1e-6, b1=0.9, b2=0.999):
    gn = torch.sqrt(sum(g.double().square().sum() for g in gs)).float()
    clip = gn / max_norm if gn > max_norm else torch.tensor(1.0, device=gn.device)
    for p, g, m, v in zip(ps, gs, ms, vs):
        g = g / clip
        m.mul_(b1).add_(g, alpha=1 - b1)
        v.mul_(b2).addcmul_(g, g, value=1 - b2)
        u = (m / (1 - b1 ** step)) / ((v / (1 - b2 ** step)).sqrt() + eps) + wd * p
        pn, un = p.norm(), u.norm()
        ratio = lr * pn / un if (wd != 0 and pn != 0 and un != 0) else lr
        p.sub_(ratio * u)


def test_fused_lamb_follows_the_published_rule():
    from apex.optimizers import FusedLAMB
    torch.manual_seed(3)
    model = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.Tanh(), torch.nn.Linear(128, 10)).cuda()
    params = list(model.parameters())
    opt = FusedLAMB(params, lr=1e-2, weight_decay=0.01, max_grad_norm=1.0)
    rp = [p.detach().clone() for p in params]
    rm = [torch.zeros_like(p) for p in rp]
    rv = [torch.zeros_like(p) for p in rp]
    x = torch.randn(32, 64, device="cuda")
    y = torch.randint(0, 10, (32,), device="cuda")
    for step in range(1, 6):
        opt.zero_grad()
        torch.nn.functional.cross_entropy(model(x), y).backward()
        gs = [p.grad.detach().clone() for p in params]
        # the reference trajectory is re-seeded from the fused one each step: per-step agreement is what is pinned
        for r, p in zip(rp, params):
            r.copy_(p.detach())
        _torch_lamb_step(rp, gs, rm, rv, step, 1e-2, 0.01, 1.0)
        opt.step()
        for r, p in zip(rp, params):
            torch.testing.assert_close(p.detach(), r, rtol=2e-5, atol=1e-7)   # fp32 rounding of differently ordered sums
    assert opt.param_groups[0]["step"] == 5
    sd = opt.state_dict()
    assert set(sd["state"][0]) == {"exp_avg", "exp_avg_sq"}


def test_fused_lamb_under_amp_o2_trains_and_skips_on_overflow():
    from apex import amp
    from apex.optimizers import FusedLAMB
    torch.manual_seed(4)
    model = torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.ReLU(), torch.nn.Linear(256, 10)).cuda()
    opt = FusedLAMB(model.parameters(), lr=2e-2)
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    x = torch.randn(64, 64, device="cuda")
    y = torch.randint(0, 10, (64,), device="cuda")
    losses = []
    for it in range(30):
        opt.zero_grad()
        loss = torch.nn.functional.cross_entropy(model(x).float(), y)
        scale_before = amp._amp_state.loss_scalers[0].loss_scale()
        if it == 10:
            loss = loss * float("inf")                        # every gradient becomes inf/NaN
        with amp.scale_loss(loss, opt) as scaled:
            scaled.backward()
        if it == 10:
            before = [p.detach().clone() for p in amp.master_params(opt)]
            opt.step()
            for b, p in zip(before, amp.master_params(opt)):
                assert torch.equal(b, p.detach())
            assert opt.param_groups[0]["step"] == 10           # the skipped step did not count
            assert amp._amp_state.loss_scalers[0].loss_scale() < scale_before or scale_before == 1.0
            continue
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < 0.85 * losses[0]            # lr x trust ratio keeps LAMB's steps small: 29 steps on one batch
    for mp, p in zip(model.parameters(), amp.master_params(opt)):
        assert torch.equal(mp.detach(), p.detach().to(mp.dtype))
