"""-m gpu tests of the sharded step (b200_peer_shard_step / b200_peer_shard_repair) with the ranks EMULATED on one GPU
(every "rank" is a set of buffers on cuda:0 and the kernels run one after the other, no hand-shake): the mean is
formed in rank order, the shard's p / m / v and EVERY rank's model buffer must equal oracle.peer_reduce +
oracle.adam bit for bit, a non-finite gradient on one rank skips the step everywhere and the repair restores the
weights.  The real-rank version (CUDA IPC, signal pads) is tests/test_gpu_peer.py::test_real_rank_sharded_step."""
import pytest
import torch

import oracle
from apex_example_b200 import ops
from apex_example_b200.parallel.peer import even_shards, shard_segments
from apex_example_b200.tables import TensorTable
from gpu_util import assert_lists_bit_equal

pytestmark = pytest.mark.gpu


def _layout(sizes, align=16):
    offs, o = [], 0
    for n in sizes:
        offs.append(o)
        o += (n + align - 1) // align * align
    return offs, o


@pytest.mark.parametrize("world,dt", [(2, torch.float16), (4, torch.float16), (8, torch.float16), (3, torch.bfloat16)])
def test_emulated_sharded_step_matches_oracle(world, dt):
    sizes = [5000, 17, 4096, 70001, 8, 12345, 9000, 300000]
    groups_of = [0, 1, 0, 0, 1, 0, 1, 0]
    offs, total = _layout(sizes)
    bounds = even_shards(total, world)
    gen = torch.Generator().manual_seed(3)
    hp = [torch.randn(n, generator=gen) * 0.05 for n in sizes]
    hm, hv = [torch.zeros(n) for n in sizes], [torch.zeros(n) for n in sizes]
    hc = [h.to(dt) for h in hp]
    flat_p = torch.zeros(total)
    for o, n, h in zip(offs, sizes, hp):
        flat_p[o:o + n] = h
    grads = [torch.zeros(total, dtype=dt, device="cuda") for _ in range(world)]
    models = [flat_p.to(dt).cuda() for _ in range(world)]
    ranks = []
    for r in range(world):
        segs = shard_segments(offs, sizes, bounds[r], bounds[r + 1])
        st = {k: [] for k in "pmv"}
        for t, first, n, flat in segs:
            init = flat_p[flat:flat + n].cuda()
            st["p"].append(ops.spec_state_like(init, fill=init))
            st["m"].append(ops.spec_state_like(torch.zeros(n, device="cuda")))
            st["v"].append(ops.spec_state_like(torch.zeros(n, device="cuda")))
        table = TensorTable([[grads[r][f:f + n] for _, _, n, f in segs]] + [[a for a, _ in st[k]] for k in "pmv"] +
                            [[models[r][f:f + n] for _, _, n, f in segs]], groups=[groups_of[t] for t, *_ in segs])
        fan = ops.make_peer_fan([g.data_ptr() for g in grads], [m.data_ptr() for m in models], r, 1.0 / world)
        ranks.append((segs, st, table, fan))
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device="cuda")
    S = 256.0
    ops.scaler_init(state, S)
    fs = state.view(torch.float32)
    parity = torch.zeros(1, dtype=torch.int32, device="cuda")
    hyper = [dict(lr=1e-3, weight_decay=0.01), dict(lr=2e-3, weight_decay=0.0)]
    scale, good, par = S, 0, 0
    for it in range(5):
        host_all = [(torch.randn(total, generator=torch.Generator().manual_seed(50 * it + r)) * 1e-2 * scale).to(dt)
                    for r in range(world)]
        if it in (1, 3):
            host_all[world - 1][offs[3] + 777] = float("inf") if it == 1 else float("nan")
        for g, h in zip(grads, host_all):
            g.copy_(h)
        grp = [ops.AdamGroup(h["lr"], 0.9, 0.999, 1e-8, h["weight_decay"], good + 1, 1, 1) for h in hyper]
        for segs, st, table, fan in ranks:
            ops.peer_shard_step(table, dt, torch.float32, dt, grp, parity, fan, dev_flag=state[ops.IDX_OVERFLOW_ACC:],
                                dev_inv_scale=fs[ops.IDX_INV_LOSS_SCALE:])
        ops.spec_commit(state, True, 2000, 2.0, 0.0, 2.0 ** 24, parity)
        for segs, st, table, fan in ranks:
            ops.peer_shard_repair(table, torch.float32, dt, state[ops.IDX_FOUND_INF:], parity, fan)
        torch.cuda.synchronize()
        want = torch.zeros(total, dtype=dt)
        wflag = oracle.peer_reduce(host_all, want, 0, total, 1.0 / world)
        assert wflag == (1 if it in (1, 3) else 0)
        assert int(state.cpu()[ops.IDX_FOUND_INF]) == wflag
        if not wflag:
            good += 1
            par ^= 1
            for i, (o, n) in enumerate(zip(offs, sizes)):
                oracle.adam(want[o:o + n].clone(), hp[i], hm[i], hv[i], hc[i], step=good, inv_scale=1.0 / scale,
                            **hyper[groups_of[i]])
        else:
            scale /= 2
        assert int(parity) == par
        for segs, st, table, fan in ranks:
            for k, h in (("p", hp), ("m", hm), ("v", hv)):
                got = [ops.spec_current(a, b, par) for a, b in st[k]]
                assert_lists_bit_equal(got, [h[t][first:first + n] for t, first, n, _ in segs], f"{k} it {it}")
        if not wflag:                       # every rank left the mean of its shard in its own gradient buffer
            for r, (segs, *_rest) in enumerate(ranks):
                for _, _, n, f in segs:
                    assert torch.equal(grads[r][f:f + n].cpu().view(torch.int16), want[f:f + n].view(torch.int16)), \
                        f"mean of rank {r}, it {it}"
        for r in range(world):              # every rank's model buffer holds the whole new model
            got = [models[r][o:o + n] for o, n in zip(offs, sizes)]
            assert_lists_bit_equal(got, hc, f"model of rank {r}, it {it}")


def test_repair_restores_weights_after_a_skipped_step():
    world, total, dt = 2, 40960, torch.float16
    bounds = even_shards(total, world)
    flat_p = torch.randn(total, generator=torch.Generator().manual_seed(1)) * 0.1
    grads = [torch.full((total,), float("nan"), dtype=dt, device="cuda") for _ in range(world)]
    models = [flat_p.to(dt).cuda() for _ in range(world)]
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device="cuda")
    ops.scaler_init(state, 128.0)
    parity = torch.zeros(1, dtype=torch.int32, device="cuda")
    built = []
    for r in range(world):
        a, b = bounds[r], bounds[r + 1]
        init = flat_p[a:b].cuda()
        p = ops.spec_state_like(init, fill=init)
        m = ops.spec_state_like(torch.zeros(b - a, device="cuda"))
        v = ops.spec_state_like(torch.zeros(b - a, device="cuda"))
        table = TensorTable([[grads[r][a:b]], [p[0]], [m[0]], [v[0]], [models[r][a:b]]])
        fan = ops.make_peer_fan([g.data_ptr() for g in grads], [w.data_ptr() for w in models], r, 0.5)
        built.append((table, fan, p))
    grp = [ops.AdamGroup(1e-2, 0.9, 0.999, 1e-8, 0.0, 1, 1, 1)]
    for table, fan, _ in built:
        ops.peer_shard_step(table, dt, torch.float32, dt, grp, parity, fan, dev_flag=state[ops.IDX_OVERFLOW_ACC:])
    torch.cuda.synchronize()
    assert all(bool(torch.isnan(w.float()).any()) for w in models)        # the speculative weights went out
    ops.spec_commit(state, True, 2000, 2.0, 0.0, 2.0 ** 24, parity)
    for table, fan, _ in built:
        ops.peer_shard_repair(table, torch.float32, dt, state[ops.IDX_FOUND_INF:], parity, fan)
    torch.cuda.synchronize()
    assert int(parity) == 0 and int(state.cpu()[ops.IDX_FOUND_INF]) == 1
    for w in models:
        assert torch.equal(w.cpu(), flat_p.to(dt))
    for _, _, p in built:
        assert not bool(torch.isnan(p[0]).any())                           # the current masters were never touched
