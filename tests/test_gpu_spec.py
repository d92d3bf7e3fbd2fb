"""-m gpu tests of the speculative single-pass step (b200_fused_adam_spec / b200_spec_commit /
b200_spec_repair): the overflow check rides on the update pass, p / m / v are double-buffered and a
device-side parity word says which copy is current.  Same bar as the other kernels: bit-exact
against oracle/ step by step, including skipped steps (state untouched, model copy repaired)."""
import pytest
import torch

import oracle
from apex_example_b200 import ops
from apex_example_b200.tables import TensorTable
from gpu_util import ADVERSARIAL_SIZES, assert_lists_bit_equal

pytestmark = pytest.mark.gpu


def _build(sizes, gdt, sdt, cdt, gen):
    hp = [(torch.randn(n, generator=gen) * 0.05).to(sdt) for n in sizes]
    hm = [(torch.randn(n, generator=gen) * 1e-3).to(sdt) for n in sizes]
    hv = [(torch.randn(n, generator=gen) * 1e-3).square().to(sdt) for n in sizes]
    hc = [h.to(cdt) for h in hp] if cdt is not None else None
    dev = {}
    for name, lst in (("p", hp), ("m", hm), ("v", hv)):
        pairs = [ops.spec_state_like(h.cuda(), fill=h.cuda()) for h in lst]
        dev[name] = pairs
    dc = [h.cuda() for h in hc] if hc is not None else None
    dg = [torch.zeros(n, dtype=gdt, device="cuda") for n in sizes]
    cols = [dg] + [[pr for pr, _ in dev[k]] for k in "pmv"] + ([dc] if dc is not None else [])
    return hp, hm, hv, hc, dev, dc, dg, TensorTable(cols)


@pytest.mark.parametrize("gdt,sdt,cdt", [(torch.float16, torch.float32, torch.float16),
                                         (torch.bfloat16, torch.float32, torch.bfloat16),
                                         (torch.float32, torch.float32, None),
                                         (torch.float16, torch.float16, None)])
def test_speculative_step_matches_oracle_through_skips_and_parity_flips(gdt, sdt, cdt):
    gen = torch.Generator().manual_seed(21)
    sizes = [s for s in ADVERSARIAL_SIZES if s]
    hp, hm, hv, hc, dev, dc, dg, table = _build(sizes, gdt, sdt, cdt, gen)
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device="cuda")
    S = 1024.0
    ops.scaler_init(state, S)
    fstate = state.view(torch.float32)
    parity = torch.zeros(1, dtype=torch.int32, device="cuda")
    sc = oracle.LossScaler(loss_scale=S, scale_window=3)
    good, expect_parity = 0, 0
    overflow_at = {1, 4, 5}
    for it in range(9):
        hg = [(torch.randn(n, generator=gen) * 1e-2 * sc.loss_scale).to(gdt) for n in sizes]
        if it in overflow_at:
            hg[it % len(sizes)][-1] = float("inf") if it % 2 else float("nan")
        for d, h in zip(dg, hg):
            d.copy_(h)
        grp = [ops.AdamGroup(1e-2, 0.9, 0.999, 1e-8, 0.01, good + 1, 1, 1)]
        ops.fused_adam_spec(table, gdt, sdt, cdt, grp, state[ops.IDX_OVERFLOW_ACC:], parity,
                            dev_inv_scale=fstate[ops.IDX_INV_LOSS_SCALE:])
        ops.spec_commit(state, True, 3, 2.0, 0.0, 2.0 ** 24, parity)
        if cdt is not None:
            ops.spec_repair(table, sdt, cdt, state[ops.IDX_FOUND_INF:], parity)
        inv = 1.0 / sc.loss_scale
        skip = sc.update_scale(it in overflow_at)
        torch.cuda.synchronize()
        assert int(state[ops.IDX_FOUND_INF]) == int(skip) == int(it in overflow_at)
        assert float(fstate[ops.IDX_LOSS_SCALE]) == sc.loss_scale
        if not skip:
            good += 1
            expect_parity ^= 1
            for i, g in enumerate(hg):
                oracle.adam(g, hp[i], hm[i], hv[i], hc[i] if hc is not None else None, lr=1e-2, weight_decay=0.01,
                            step=good, inv_scale=inv)
        assert int(parity) == expect_parity
        for name, want in (("p", hp), ("m", hm), ("v", hv)):
            cur = [ops.spec_current(pr, buf, expect_parity) for pr, buf in dev[name]]
            assert_lists_bit_equal(cur, want, f"{name} (iteration {it}, parity {expect_parity})")
        if dc is not None:              # after a skip the repaired copy equals cast(current master) again
            assert_lists_bit_equal(dc, hc, f"model copy (iteration {it})")
    assert good == 6


def test_speculative_step_never_writes_outside_its_two_copies():
    """Sentinels right after every shadow copy (the buffers are exactly primary + padding + shadow)."""
    gen = torch.Generator().manual_seed(3)
    sizes = [1, 7, 4097, 65537]
    hp, hm, hv, hc, dev, dc, dg, table = _build(sizes, torch.float16, torch.float32, torch.float16, gen)
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device="cuda")
    ops.scaler_init(state, 1.0)
    parity = torch.zeros(1, dtype=torch.int32, device="cuda")
    pads = []
    for name in "pmv":
        for pr, buf in dev[name]:
            delta = ops.spec_shadow_delta(pr.numel(), 4) // 4
            if delta > pr.numel():
                buf[pr.numel():delta] = 777.0           # the gap between the copies
                pads.append(buf[pr.numel():delta])
    for step in (1, 2, 3):
        for d in dg:
            d.normal_(0, 1e-2)
        ops.fused_adam_spec(table, torch.float16, torch.float32, torch.float16,
                            [ops.AdamGroup(1e-2, 0.9, 0.999, 1e-8, 0.0, step, 1, 1)], state[ops.IDX_OVERFLOW_ACC:], parity)
        ops.spec_commit(state, True, 2000, 2.0, 0.0, 2.0 ** 24, parity)
    torch.cuda.synchronize()
    assert int(parity) == 1
    assert all(bool((x == 777.0).all()) for x in pads)


# ---- fused gradient mean + speculative step (b200_peer_spec_step) ---------------------------------------------
def _layout(sizes):
    offs, off = [], 0
    for n in sizes:
        off += (-off) % 16
        offs.append(off)
        off += n
    return offs, off


@pytest.mark.parametrize("world,chunk", [(2, 131072), (4, 262144), (4, 1 << 20), (8, 131072)])
def test_fused_peer_step_emulated_ranks_match_oracle(world, chunk):
    """All `world` ranks live on ONE GPU: each has its own buffers, state, signal pad, scratch and stream, and
    its kernel (grid capped to 12 CTAs so that all of them are resident together) really waits for the others'
    progress flags.  Two steps (the second flips parity back) + one overflow step injected on the last rank."""
    from apex_example_b200.tables import tile_aligned_shards
    ops.set_peer_grid(12)
    try:
        gen = torch.Generator().manual_seed(40 + world)
        sizes = [5000, 17, 4096, 70001, 8, 12345, 9000, 930000]
        offs, total = _layout(sizes)
        bounds = tile_aligned_shards(offs, sizes, total, world)
        S = 256.0
        pads = torch.zeros(world, ops.PEER_STEP_PAD_WORDS, dtype=torch.int32, device="cuda")   # pad + progress board
        pad_ptrs = [pads.data_ptr() + 4 * ops.PEER_STEP_PAD_WORDS * r for r in range(world)]
        grads = [torch.zeros(total, dtype=torch.float16, device="cuda") for _ in range(world)]
        reduced = [torch.full((total,), float("nan"), dtype=torch.float16, device="cuda") for _ in range(world)]
        hp = [torch.randn(n, generator=gen) * 0.05 for n in sizes]
        hm, hv = [torch.zeros(n) for n in sizes], [torch.zeros(n) for n in sizes]
        hc = [h.half() for h in hp]
        ranks = []
        for r in range(world):
            st = {k: [ops.spec_state_like(h.cuda(), fill=h.cuda()) for h in lst] for k, lst in (("p", hp), ("m", hm), ("v", hv))}
            dc = [h.cuda() for h in hc]
            cols = [[grads[r][o:o + n] for o, n in zip(offs, sizes)]] + [[a for a, _ in st[k]] for k in "pmv"] + [dc]
            table = TensorTable(cols).order_for_peer_step(grads[r].data_ptr(), 2, bounds, r, chunk)
            state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device="cuda")
            ops.scaler_init(state, S)
            ranks.append(dict(st=st, dc=dc, table=table, state=state, parity=torch.zeros(1, dtype=torch.int32, device="cuda"),
                              scratch=ops.new_peer_step_scratch("cuda"), stream=torch.cuda.Stream(),
                              pull=ops.make_peer_pull(grads[r].data_ptr(), bounds, [t.data_ptr() for t in reduced], 2)))
        torch.cuda.synchronize()
        scale, good, par = S, 0, 0
        for it in range(4):
            host_g = [(torch.randn(total, generator=gen) * 1e-2 * scale).half() for _ in range(world)]
            if it == 2:
                host_g[world - 1][total - 3] = float("inf")
            for r in range(world):
                grads[r].copy_(host_g[r])
            torch.cuda.synchronize()
            for r, rk in enumerate(ranks):
                with torch.cuda.stream(rk["stream"]):
                    fs = rk["state"].view(torch.float32)
                    ops.peer_spec_step(rk["table"], torch.float16, torch.float32, torch.float16,
                                       [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.01, good + 1, 1, 1)],
                                       rk["state"][ops.IDX_OVERFLOW_ACC:], rk["parity"], rk["pull"],
                                       [g.data_ptr() for g in grads], reduced[r], 1.0 / world,
                                       ops.make_peer_sync(pad_ptrs, r, it + 1), 3, chunk, rk["scratch"],
                                       dev_inv_scale=fs[ops.IDX_INV_LOSS_SCALE:])
                    ops.spec_commit(rk["state"], True, 2000, 2.0, 0.0, 2.0 ** 24, rk["parity"])
                    ops.spec_repair(rk["table"], torch.float32, torch.float16, rk["state"][ops.IDX_FOUND_INF:], rk["parity"])
            torch.cuda.synchronize()
            assert int(pads[:, 16].abs().sum()) == 0, "a bounded wait timed out"
            want = torch.zeros(total, dtype=torch.float16)
            wflag = oracle.peer_reduce(host_g, want, 0, total, 1.0 / world)
            assert wflag == (1 if it == 2 else 0)
            if not wflag:
                good += 1
                par ^= 1
                for i, (o, n) in enumerate(zip(offs, sizes)):
                    oracle.adam(want[o:o + n].clone(), hp[i], hm[i], hv[i], hc[i], lr=1e-3, weight_decay=0.01, step=good,
                                inv_scale=1.0 / scale)
            else:
                scale /= 2
            for r, rk in enumerate(ranks):
                assert int(rk["state"][ops.IDX_FOUND_INF]) == wflag, f"rank {r}, iteration {it}"
                assert int(rk["parity"]) == par
                for k, h in (("p", hp), ("m", hm), ("v", hv)):
                    assert_lists_bit_equal([ops.spec_current(a, b, par) for a, b in rk["st"][k]], h, f"{k} (rank {r}, it {it})")
                assert_lists_bit_equal(rk["dc"], hc, f"copy (rank {r}, it {it})")
        assert good == 3
    finally:
        ops.set_peer_grid(0)
