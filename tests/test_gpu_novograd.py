"""-m gpu tests of FusedNovoGrad: b200_novograd_blend + b200_fused_novograd through the C ABI,
bit-exact against oracle/ (layer-wise second moments from the deterministic per-tensor norms), and
apex.optimizers.FusedNovoGrad against a torch restatement of the published rule.  Unpinned against apex itself."""
import pytest
import torch

import oracle
from apex_example_b200 import ops
from apex_example_b200.tables import TensorTable
from gpu_util import ADVERSARIAL_SIZES, assert_lists_bit_equal

pytestmark = pytest.mark.gpu

SIZES = [s for s in ADVERSARIAL_SIZES if s] + [40000]


@pytest.mark.parametrize("gdt,sdt,cdt", [(torch.float16, torch.float32, torch.float16),
                                         (torch.bfloat16, torch.float32, torch.bfloat16),
                                         (torch.float32, torch.float32, None),
                                         (torch.float16, torch.float16, None)])
@pytest.mark.parametrize("hyper", [dict(lr=1e-2, weight_decay=0.01),
                                   dict(lr=1e-2, weight_decay=0.01, reg_inside_moment=1, grad_averaging=0),
                                   dict(lr=1e-2, weight_decay=0.0, init_zero=1, bias_correction=0)])
def test_novograd_bit_exact(gdt, sdt, cdt, hyper):
    gen = torch.Generator().manual_seed(13)
    inv_scale = 1.0 / 128
    hp = [(torch.randn(n, generator=gen) * 0.05).to(sdt) for n in SIZES]
    hm = [torch.zeros(n, dtype=sdt) for n in SIZES]
    hc = [h.to(cdt) for h in hp] if cdt is not None else None
    hv = torch.zeros(len(SIZES))
    dp, dm = ([h.cuda() for h in lst] for lst in (hp, hm))
    dc = [h.cuda() for h in hc] if hc is not None else None
    dg = [torch.zeros(n, dtype=gdt, device="cuda") for n in SIZES]
    dv = torch.zeros(len(SIZES), device="cuda")
    table = TensorTable([dg, dp, dm] + ([dc] if dc is not None else []))
    first = table.first_tiles()
    tile_ss = torch.empty(table.n_tiles, device="cuda")
    ss = torch.zeros(len(SIZES), device="cuda")
    for step in (1, 2, 3):
        hg = [(torch.randn(n, generator=gen) * 0.02 / inv_scale).to(gdt) for n in SIZES]
        for d, h in zip(dg, hg):
            d.copy_(h)
        grp = [ops.NovoGradGroup(hyper["lr"], 0.9, 0.999, 1e-8, hyper["weight_decay"], step, hyper.get("bias_correction", 1),
                                 hyper.get("reg_inside_moment", 0), hyper.get("grad_averaging", 1), hyper.get("init_zero", 0))]
        ops.l2norm_per_tensor(table, 0, gdt, first, tile_ss, ss)
        ops.novograd_blend(table, grp, dv, ss, inv_scale=inv_scale)
        ops.fused_novograd(table, gdt, sdt, cdt, grp, dv, inv_scale=inv_scale)
        want_ss, _, _ = oracle.l2norm_per_tensor(hg, table.tile_elems)
        assert torch.equal(ss.cpu(), want_ss)
        oracle.novograd_list(hg, hp, hm, hv, want_ss, hc, inv_scale=inv_scale, step=step, **hyper)
        assert torch.equal(dv.cpu().view(torch.int32), hv.view(torch.int32)), f"layer-wise moments, step {step}"
        assert_lists_bit_equal(dm, hm, f"m step {step}")
        assert_lists_bit_equal(dp, hp, f"p step {step}")
        if dc is not None:
            assert_lists_bit_equal(dc, hc, f"copy step {step}")
    # a raised found_inf skips both kernels
    found = torch.ones(1, dtype=torch.int32, device="cuda")
    snap_v, snap_p = dv.clone(), [t.clone() for t in dp]
    ops.novograd_blend(table, grp, dv, ss, inv_scale=inv_scale, dev_found_inf=found)
    ops.fused_novograd(table, gdt, sdt, cdt, grp, dv, inv_scale=inv_scale, dev_found_inf=found)
    assert torch.equal(dv, snap_v) and all(torch.equal(a, b) for a, b in zip(dp, snap_p))


def test_fused_novograd_follows_the_published_rule_and_checkpoints():
    from apex.optimizers import FusedNovoGrad
    torch.manual_seed(5)
    model = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.Tanh(), torch.nn.Linear(128, 10)).cuda()
    params = list(model.parameters())
    lr, wd, b1, b2, eps = 1e-2, 0.01, 0.95, 0.98, 1e-8
    opt = FusedNovoGrad(params, lr=lr, weight_decay=wd, betas=(b1, b2))
    rm = [torch.zeros_like(p) for p in params]
    rv = [None] * len(params)
    x = torch.randn(32, 64, device="cuda")
    y = torch.randint(0, 10, (32,), device="cuda")
    for step in range(1, 6):
        opt.zero_grad()
        torch.nn.functional.cross_entropy(model(x), y).backward()
        want = []
        for i, p in enumerate(params):
            g = p.grad.detach()
            n = g.norm()
            rv[i] = n if step == 1 else torch.sqrt(b2 * rv[i] ** 2 + (1 - b2) * n ** 2)
            rm[i] = b1 * rm[i] + (1 - b1) * g
            denom = rv[i] / (1 - b2 ** step) ** 0.5 + eps
            want.append(p.detach() - lr * ((rm[i] / (1 - b1 ** step)) / denom + wd * p.detach()))
        opt.step()
        for w, p in zip(want, params):
            torch.testing.assert_close(p.detach(), w, rtol=2e-5, atol=1e-7)
    assert opt.param_groups[0]["step"] == 5
    sd = opt.state_dict()
    assert set(sd["state"][0]) == {"exp_avg", "exp_avg_sq"} and sd["state"][0]["exp_avg_sq"].dim() == 0
    # a fresh optimizer that loads the checkpoint continues identically
    import copy
    model2 = copy.deepcopy(model)
    opt2 = FusedNovoGrad(model2.parameters(), lr=lr, weight_decay=wd, betas=(b1, b2))
    opt2.load_state_dict(copy.deepcopy(sd))
    for mdl, o in ((model, opt), (model2, opt2)):
        o.zero_grad()
        torch.nn.functional.cross_entropy(mdl(x), y).backward()
        o.step()
    for a, b in zip(model.parameters(), model2.parameters()):
        assert torch.equal(a.detach(), b.detach())


def test_fused_novograd_under_amp_o2_skips_on_overflow():
    from apex import amp
    from apex.optimizers import FusedNovoGrad
    torch.manual_seed(6)
    model = torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.ReLU(), torch.nn.Linear(256, 10)).cuda()
    opt = FusedNovoGrad(model.parameters(), lr=1e-2, betas=(0.95, 0.98))
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    x = torch.randn(64, 64, device="cuda")
    y = torch.randint(0, 10, (64,), device="cuda")
    losses = []
    for it in range(30):
        opt.zero_grad()
        loss = torch.nn.functional.cross_entropy(model(x).float(), y)
        if it == 10:
            loss = loss * float("inf")
        with amp.scale_loss(loss, opt) as scaled:
            scaled.backward()
        if it == 10:
            before = [p.detach().clone() for p in amp.master_params(opt)]
            steps_before = opt.param_groups[0].get("step", 0)
            opt.step()
            assert all(torch.equal(b, p.detach()) for b, p in zip(before, amp.master_params(opt)))
            assert opt.param_groups[0].get("step", 0) == steps_before
            continue
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0]
    for mp, p in zip(model.parameters(), amp.master_params(opt)):
        assert torch.equal(mp.detach(), p.detach().to(mp.dtype))
