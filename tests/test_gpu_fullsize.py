"""-m gpu tests at BASELINE.json's full sizes (config 2: the 124M-parameter GPT-2-small list):
a whole-list comparison with the oracle plus size-independent properties of the path."""
import pytest
import torch

import oracle
from apex_example_b200 import ops
from apex_example_b200.tables import TensorTable
from gpu_util import assert_lists_bit_equal
from workloads import GPT2_SMALL, many_small

pytestmark = pytest.mark.gpu


def _alloc(sizes, seed=0, S=65536.0):
    torch.manual_seed(seed)
    g = torch.Generator(device="cuda").manual_seed(seed)
    n = sum(sizes)
    flat = {
        "c": (torch.randn(n, device="cuda", generator=g) * 0.02).half(),
        "g": (torch.randn(n, device="cuda", generator=g) * 1e-3 * S).half(),
    }
    flat["p"] = flat["c"].float()
    flat["m"] = torch.randn(n, device="cuda", generator=g) * 1e-4
    flat["v"] = (torch.randn(n, device="cuda", generator=g) * 1e-4).square_()
    lists = {k: list(torch.split(v, sizes)) for k, v in flat.items()}
    return flat, lists


def test_gpt2_small_list_one_step_equals_oracle_everywhere():
    S = 65536.0
    flat, L = _alloc(GPT2_SMALL)
    host = {k: [t.cpu() for t in v] for k, v in L.items()}
    table = TensorTable([L["g"], L["p"], L["m"], L["v"], L["c"]])
    assert table.n_tensors == 148 and table.total_elems == 124_439_808
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.multi_tensor_check(table, 0, torch.float16, flag)
    ops.fused_adam(table, torch.float16, torch.float32, torch.float16,
                   [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.01, 4, 1, 1)], inv_scale=1.0 / S, dev_found_inf=flag)
    skipped = oracle.adam_list(host["g"], host["p"], host["m"], host["v"], host["c"], inv_scale=1.0 / S, lr=1e-3,
                               weight_decay=0.01, step=4)
    torch.cuda.synchronize()
    assert skipped == 0 == int(flag.item())
    for k in ("p", "m", "v", "c"):
        assert_lists_bit_equal(L[k], host[k], k)
    # the model copy is exactly the RNE cast of the master (a8 in SURVEY.md §8a)
    assert torch.equal(flat["c"], flat["p"].half())


def test_one_inf_anywhere_skips_the_whole_step():
    """Checksum property: with a single inf in the LAST gradient the device-side flag must stop
    every tile of every tensor from being written."""
    S = 1024.0
    flat, L = _alloc(GPT2_SMALL, seed=1, S=S)
    sums = {k: flat[k].double().sum().item() for k in ("p", "m", "v")}
    csum = flat["c"].float().sum().item()
    L["g"][-1][-1] = float("inf")
    table = TensorTable([L["g"], L["p"], L["m"], L["v"], L["c"]])
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device="cuda")
    ops.scaler_init(state, S)
    ops.multi_tensor_check(table, 0, torch.float16, state[ops.IDX_OVERFLOW_ACC:])
    ops.scaler_update(state, True, 2000, 2.0, 0.0, 2.0 ** 24)
    ops.fused_adam(table, torch.float16, torch.float32, torch.float16, [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.0, 1, 1, 1)],
                   dev_inv_scale=state.view(torch.float32)[ops.IDX_INV_SCALE:], dev_found_inf=state[ops.IDX_FOUND_INF:])
    torch.cuda.synchronize()
    st = state.cpu()
    assert int(st[ops.IDX_FOUND_INF]) == 1 and float(st.view(torch.float32)[0]) == S / 2
    for k in ("p", "m", "v"):
        assert flat[k].double().sum().item() == sums[k]
    assert flat["c"].float().sum().item() == csum
    # the next, clean, step goes through with the halved scale's predecessor (inv = 1/S)
    L["g"][-1][-1] = 0.0
    ops.multi_tensor_check(table, 0, torch.float16, state[ops.IDX_OVERFLOW_ACC:])
    ops.scaler_update(state, True, 2000, 2.0, 0.0, 2.0 ** 24)
    ops.fused_adam(table, torch.float16, torch.float32, torch.float16, [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.0, 1, 1, 1)],
                   dev_inv_scale=state.view(torch.float32)[ops.IDX_INV_SCALE:], dev_found_inf=state[ops.IDX_FOUND_INF:])
    torch.cuda.synchronize()
    assert flat["p"].double().sum().item() != sums["p"]
    assert float(state.cpu().view(torch.float32)[ops.IDX_INV_SCALE]) == 2.0 / S


def test_sgd_without_momentum_is_linear_in_the_gradient():
    """w' = w - lr*s*g exactly representable here: power-of-two lr and scale, small-integer g."""
    sizes = many_small(total=20_000_000, seed=3)
    n = sum(sizes)
    gen = torch.Generator(device="cuda").manual_seed(3)
    g = torch.randint(-8, 9, (n,), device="cuda", generator=gen).half()
    w = torch.randint(-64, 65, (n,), device="cuda", generator=gen).float()
    w0 = w.clone()
    c = torch.zeros(n, dtype=torch.float16, device="cuda")
    table = TensorTable([list(torch.split(g, sizes)), list(torch.split(w, sizes)), [None] * len(sizes),
                         list(torch.split(c, sizes))])
    ops.fused_sgd(table, torch.float16, torch.float32, torch.float16, [ops.SgdGroup(0.25, 0.0, 0.0, 0.0, 1, 0, 0)],
                  inv_scale=0.5)
    torch.cuda.synchronize()
    assert torch.equal(w, w0 - 0.125 * g.float())
    assert torch.equal(c, w.half())


def test_many_small_ragged_list_o3_equals_oracle():
    """Config-5-shaped slice (sizes log-uniform in [256, 262144], many not divisible by 8), O3 layout:
    fp16 p/m/v/g, no master."""
    sizes = many_small(total=30_000_000, seed=5)
    assert any(s % 8 for s in sizes)
    gen = torch.Generator().manual_seed(5)
    hp = [(torch.randn(s, generator=gen) * 0.05).half() for s in sizes]
    hm = [(torch.randn(s, generator=gen) * 1e-3).half() for s in sizes]
    hv = [(torch.randn(s, generator=gen) * 1e-2).square().half() for s in sizes]
    hg = [(torch.randn(s, generator=gen) * 4).half() for s in sizes]
    # one flat device allocation per column: consecutive tensors are then NOT 16-byte aligned
    def flat(hs):
        f = torch.cat(hs).cuda()
        return list(torch.split(f, sizes))
    dp, dm, dv, dg = flat(hp), flat(hm), flat(hv), flat(hg)
    table = TensorTable([dg, dp, dm, dv])
    ops.fused_adam(table, torch.float16, torch.float16, None, [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.0, 3, 1, 1)],
                   inv_scale=1.0 / 128)
    oracle.adam_list(hg, hp, hm, hv, None, inv_scale=1.0 / 128, lr=1e-3, step=3, check_overflow=False)
    torch.cuda.synchronize()
    assert_lists_bit_equal(dp, hp, "p"); assert_lists_bit_equal(dm, hm, "m"); assert_lists_bit_equal(dv, hv, "v")
