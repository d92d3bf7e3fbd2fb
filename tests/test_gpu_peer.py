"""-m gpu tests of the peer-memory gradient mean (b200_peer_reduce_shard + b200_fused_*_peer).

On one GPU the ranks are emulated as separate buffers on the same device (the kernels only see
pointers), which covers the arithmetic, the shard/tile ownership logic and the overflow flag.
The CUDA-IPC plumbing and real NVLink loads need `gpurun --gpus 2` (skipped otherwise)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from apex_example_b200 import ops
from apex_example_b200.tables import TensorTable, tile_aligned_shards
from gpu_util import assert_lists_bit_equal, bit_equal
from test_ddp_gloo import _free_port

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _layout(sizes):
    offs, off = [], 0
    for n in sizes:
        off += (-off) % 16                       # keep every tensor 32-byte aligned inside the flat buffer
        offs.append(off)
        off += n
    return offs, off


@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_peer_reduce_shard_emulated_ranks_match_oracle(dt, world):
    gen = torch.Generator().manual_seed(world)
    n = 100_003
    host = [(torch.randn(n, generator=gen) * 3000).to(dt) for _ in range(world)]
    dev = [h.cuda() for h in host]
    bounds = [0] + sorted(int(x) for x in torch.randint(1, n - 1, (world - 1,), generator=gen)) + [n]
    out = torch.zeros(n, dtype=dt, device="cuda")
    want = torch.zeros(n, dtype=dt)
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    wflag = 0
    for r in range(world):                       # each emulated rank reduces its own shard
        ops.peer_reduce_shard([d.data_ptr() for d in dev], out, bounds[r], bounds[r + 1], 1.0 / world, flag)
        wflag |= oracle.peer_reduce(host, want, bounds[r], bounds[r + 1], 1.0 / world)
    torch.cuda.synchronize()
    assert bit_equal(out, want)
    assert int(flag.item()) == wflag
    if dt == torch.float16 and world >= 2:       # the SUM overflows although the mean would not
        assert wflag == 1 or float(want.float().abs().max()) < 65504
    host[-1][bounds[-2]] = float("nan")
    dev[-1][bounds[-2]] = float("nan")
    flag.zero_()
    ops.peer_reduce_shard([d.data_ptr() for d in dev], out, bounds[-2], bounds[-1], 1.0 / world, flag)
    assert int(flag.item()) == 1


def test_fp16_sum_overflow_is_flagged_like_apex_sum_then_scale():
    a = torch.full((64,), 40000.0).half().cuda()
    b = torch.full((64,), 40000.0).half().cuda()
    out = torch.zeros(64, dtype=torch.float16, device="cuda")
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.peer_reduce_shard([a.data_ptr(), b.data_ptr()], out, 0, 64, 0.5, flag)
    assert int(flag.item()) == 1 and bool(torch.isinf(out).all())     # SUM = 80000 > fp16 max, as with NCCL fp16 SUM


@pytest.mark.parametrize("world", [2, 4])
def test_fused_adam_pulls_each_tile_from_its_owner_emulated(world):
    gen = torch.Generator().manual_seed(10 + world)
    sizes = [5000, 17, 4096, 70001, 8, 12345, 9000]
    offs, total = _layout(sizes)
    bounds, first = tile_aligned_shards(offs, sizes, total, world, return_first_tiles=True)
    S = 256.0
    host_g = [(torch.randn(total, generator=gen) * 1e-2 * S).half() for _ in range(world)]
    dev_g = [h.cuda() for h in host_g]
    reduced = [torch.full((total,), float("nan"), dtype=torch.float16, device="cuda") for _ in range(world)]
    want_red = torch.zeros(total, dtype=torch.float16)
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    for r in range(world):                       # rank r writes ONLY its shard of its own `reduced` buffer
        ops.peer_reduce_shard([d.data_ptr() for d in dev_g], reduced[r], bounds[r], bounds[r + 1], 1.0 / world, flag)
        oracle.peer_reduce(host_g, want_red, bounds[r], bounds[r + 1], 1.0 / world)
    hp = [torch.randn(n, generator=gen) * 0.05 for n in sizes]
    hm = [torch.zeros(n) for n in sizes]
    hv = [torch.zeros(n) for n in sizes]
    hc = [torch.zeros(n, dtype=torch.float16) for n in sizes]
    for rank in range(world):                    # every emulated rank must end with the same state
        dp, dm, dv, dc = ([t.clone().cuda() for t in lst] for lst in (hp, hm, hv, hc))
        gviews = [dev_g[rank][o:o + n] for o, n in zip(offs, sizes)]         # table column 0: the LOCAL flat buffer
        table = TensorTable([gviews, dp, dm, dv, dc])
        pull = ops.make_peer_pull(dev_g[rank].data_ptr(), bounds, [t.data_ptr() for t in reduced], 2,
                                  first_tile=first[rank])           # rotated sweep: same result, other order
        ops.fused_adam(table, torch.float16, torch.float32, torch.float16,
                       [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.01, 1, 1, 1)], inv_scale=1.0 / S, pull=pull)
        op, om, ov, oc = ([t.clone() for t in lst] for lst in (hp, hm, hv, hc))
        for i, (o, n) in enumerate(zip(offs, sizes)):
            oracle.adam(want_red[o:o + n].clone(), op[i], om[i], ov[i], oc[i], lr=1e-3, weight_decay=0.01, step=1,
                        inv_scale=1.0 / S)
        torch.cuda.synchronize()
        assert_lists_bit_equal(dp, op, f"p(rank {rank})")
        assert_lists_bit_equal(dm, om, f"m(rank {rank})")
        assert_lists_bit_equal(dc, oc, f"copy(rank {rank})")


@pytest.mark.parametrize("n,world", [(10, 4), (10, 2), (3, 8), (64, 4), (70, 8)])
def test_small_bucket_with_empty_trailing_shards_reduces_and_gathers(n, world):
    """DDP's peer backend shards a bucket in multiples of 8 elements, so a small bucket whose size
    is not a multiple of 8 (a lone 10-element bias at world 4: bounds [0, 8, 10, 10, 10]) has interior bounds
    equal to its size.  Both halves of the all-reduce must accept and handle them."""
    gen = torch.Generator().manual_seed(n * 10 + world)
    host = [(torch.randn(n, generator=gen) * 8).half() for _ in range(world)]
    dev = [h.cuda() for h in host]
    shard = -(-n // world)
    shard += (-shard) % 8
    bounds = [min(r * shard, n) for r in range(world)] + [n]
    reduced = [torch.full((n,), float("nan"), dtype=torch.float16, device="cuda") for _ in range(world)]
    want = torch.zeros(n, dtype=torch.float16)
    for r in range(world):
        ops.peer_reduce_shard([d.data_ptr() for d in dev], reduced[r], bounds[r], bounds[r + 1], 1.0 / world)
        oracle.peer_reduce(host, want, bounds[r], bounds[r + 1], 1.0 / world)
    out = torch.zeros(n, dtype=torch.float16, device="cuda")
    ops.peer_gather([t.data_ptr() for t in reduced], out, bounds)
    torch.cuda.synchronize()
    assert bit_equal(out, want)
    # a boundary INSIDE the buffer that cuts an 8-element group is still refused
    if n > 9:
        with pytest.raises(RuntimeError, match="multiple of 8"):
            ops.peer_gather([t.data_ptr() for t in reduced[:2]], out, [0, 5, n])


def test_tile_aligned_shards_never_cut_a_tile():
    sizes = [5000, 17, 4096, 70001, 8, 12345, 9000]
    offs, total = _layout(sizes)
    for world in (1, 2, 3, 8):
        b = tile_aligned_shards(offs, sizes, total, world)
        assert b[0] == 0 and b[-1] == total and len(b) == world + 1
        for o, n in zip(offs, sizes):
            for s in range(o, o + n, 4096):
                e = min(s + 4096, o + n)
                assert not any(s < x < e for x in b), (world, s, e, b)


# ---- real peers over CUDA IPC / NVLink ----------------------------------------------------------------
def _worker_ipc(rank, world, port, barrier="nccl", repeats=1):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from apex_example_b200.parallel.peer import PeerBuffers, PeerGradientReducer
    sizes = [5000, 17, 4096, 70001, 8, 12345, 9000]
    offs, total = _layout(sizes)
    bounds, first = tile_aligned_shards(offs, sizes, total, world, return_first_tiles=True)
    bufs = PeerBuffers(total, torch.float16)
    red = PeerGradientReducer(bufs, bounds, first_tiles=first, barrier=barrier)
    S = 256.0
    host_all = [(torch.randn(total, generator=torch.Generator().manual_seed(100 + r)) * 1e-2 * S).half() for r in range(world)]
    want_red = torch.zeros(total, dtype=torch.float16)
    oracle.peer_reduce(host_all, want_red, 0, total, 1.0 / world)
    gen = torch.Generator().manual_seed(7)
    hp = [torch.randn(n, generator=gen) * 0.05 for n in sizes]
    hm, hv = [torch.zeros(n) for n in sizes], [torch.zeros(n) for n in sizes]
    hc = [torch.zeros(n, dtype=torch.float16) for n in sizes]
    dp, dm, dv, dc = ([t.clone().cuda() for t in lst] for lst in (hp, hm, hv, hc))
    g_local = bufs.local["grads"]
    table = TensorTable([[g_local[o:o + n] for o, n in zip(offs, sizes)], dp, dm, dv, dc])
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device="cuda")
    ops.scaler_init(state, S)
    for step in (1, 2, 3):
        if step == 2:
            host_all[world - 1][total - 1] = float("inf")          # overflow on the LAST rank only
        else:
            host_all[world - 1][total - 1] = 1.0
        g_local.copy_(host_all[rank])
        red.reduce(state[ops.IDX_OVERFLOW_ACC:])
        ops.scaler_update(state, True, 2000, 2.0, 0.0, 2.0 ** 24)
        ops.fused_adam(table, torch.float16, torch.float32, torch.float16,
                       [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.01, 1 if step == 1 else 2, 1, 1)],
                       dev_inv_scale=state.view(torch.float32)[ops.IDX_INV_SCALE:], dev_found_inf=state[ops.IDX_FOUND_INF:],
                       pull=red.pull)
        torch.cuda.synchronize()
        found = int(state.cpu()[ops.IDX_FOUND_INF])
        assert found == (1 if step == 2 else 0), f"rank {rank}: skip decision {found} at step {step}"
        if step != 2:
            oracle.peer_reduce(host_all, want_red, 0, total, 1.0 / world)
            for i, (o, n) in enumerate(zip(offs, sizes)):
                # the overflow at step 2 halved the loss scale, so step 3 is unscaled by 1/(S/2)
                oracle.adam(want_red[o:o + n].clone(), hp[i], hm[i], hv[i], hc[i], lr=1e-3, weight_decay=0.01,
                            step=1 if step == 1 else 2, inv_scale=(1.0 / S) if step == 1 else (2.0 / S))
        assert_lists_bit_equal(dp, hp, f"p(rank {rank}, step {step})")
        assert_lists_bit_equal(dc, hc, f"copy(rank {rank}, step {step})")
    # hammer the hand-shake: many back-to-back steps without any host synchronisation in between
    for _ in range(repeats):
        red.reduce(state[ops.IDX_OVERFLOW_ACC:])
        ops.scaler_update(state, True, 2000, 2.0, 0.0, 2.0 ** 24)
        ops.fused_adam(table, torch.float16, torch.float32, torch.float16,
                       [ops.AdamGroup(0.0, 0.9, 0.999, 1e-8, 0.0, 3, 1, 1)],      # lr 0: state moves, weights do not
                       dev_inv_scale=state.view(torch.float32)[ops.IDX_INV_SCALE:], dev_found_inf=state[ops.IDX_FOUND_INF:],
                       pull=red.pull)
    torch.cuda.synchronize()
    bufs.check_signals()
    red_now = bufs.local["reduced"][bounds[rank]:bounds[rank + 1]].cpu()
    assert torch.equal(red_now, want_red[bounds[rank]:bounds[rank + 1]])
    del table
    bufs.close()
    dist.destroy_process_group()


def _worker_fused(rank, world, port, repeats):
    """Real ranks: b200_peer_spec_step over CUDA-IPC buffers, against the oracle, incl. a one-rank overflow."""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from apex_example_b200.parallel.peer import PeerBuffers, PeerSpecStep
    sizes = [5000, 17, 4096, 70001, 8, 12345, 9000, 1600000]
    offs, total = _layout(sizes)
    bounds = tile_aligned_shards(offs, sizes, total, world)
    bufs = PeerBuffers(total, torch.float16)
    S = 256.0
    gen = torch.Generator().manual_seed(7)
    hp = [torch.randn(n, generator=gen) * 0.05 for n in sizes]
    hm, hv = [torch.zeros(n) for n in sizes], [torch.zeros(n) for n in sizes]
    hc = [h.half() for h in hp]
    st = {k: [ops.spec_state_like(h.cuda(), fill=h.cuda()) for h in lst] for k, lst in (("p", hp), ("m", hm), ("v", hv))}
    dc = [h.cuda() for h in hc]
    g_local = bufs.local["grads"]
    table = TensorTable([[g_local[o:o + n] for o, n in zip(offs, sizes)]] + [[a for a, _ in st[k]] for k in "pmv"] + [dc])
    stepper = PeerSpecStep(bufs, bounds, table, chunk_elems=131072, n_reducers=24)
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device="cuda")
    ops.scaler_init(state, S)
    fs = state.view(torch.float32)
    parity = torch.zeros(1, dtype=torch.int32, device="cuda")

    def launch(step_no, lr=1e-3, wd=0.01):
        stepper.step(torch.float16, torch.float32, torch.float16, [ops.AdamGroup(lr, 0.9, 0.999, 1e-8, wd, step_no, 1, 1)],
                     state[ops.IDX_OVERFLOW_ACC:], parity, dev_inv_scale=fs[ops.IDX_INV_LOSS_SCALE:])
        ops.spec_commit(state, True, 2000, 2.0, 0.0, 2.0 ** 24, parity)
        ops.spec_repair(stepper.table, torch.float32, torch.float16, state[ops.IDX_FOUND_INF:], parity)

    scale, good, par = S, 0, 0
    for it in range(4):
        host_all = [(torch.randn(total, generator=torch.Generator().manual_seed(100 * it + r)) * 1e-2 * scale).half()
                    for r in range(world)]
        if it == 1:
            host_all[world - 1][total - 1] = float("inf")          # overflow on the LAST rank only
        g_local.copy_(host_all[rank])
        launch(good + 1)
        torch.cuda.synchronize()
        want = torch.zeros(total, dtype=torch.float16)
        wflag = oracle.peer_reduce(host_all, want, 0, total, 1.0 / world)
        assert wflag == (1 if it == 1 else 0)
        assert int(state.cpu()[ops.IDX_FOUND_INF]) == wflag, f"rank {rank}: skip decision at iteration {it}"
        if not wflag:
            good += 1
            par ^= 1
            for i, (o, n) in enumerate(zip(offs, sizes)):
                oracle.adam(want[o:o + n].clone(), hp[i], hm[i], hv[i], hc[i], lr=1e-3, weight_decay=0.01, step=good,
                            inv_scale=1.0 / scale)
        else:
            scale /= 2
        assert int(parity) == par
        for k, h in (("p", hp), ("m", hm), ("v", hv)):
            assert_lists_bit_equal([ops.spec_current(a, b, par) for a, b in st[k]], h, f"{k} (rank {rank}, it {it})")
        assert_lists_bit_equal(dc, hc, f"copy (rank {rank}, it {it})")
    for _ in range(repeats):                 # hammer the progress protocol: back-to-back steps, no host sync
        launch(good + 1, lr=0.0, wd=0.0)
    torch.cuda.synchronize()
    bufs.check_signals()
    del table, stepper
    bufs.close()
    dist.destroy_process_group()


def _worker_sharded(rank, world, port, repeats, multicast):
    """Real ranks: PeerShardedStep (b200_peer_shard_step + flag exchange + commit + repair) over CUDA-IPC buffers,
    against the oracle, incl. an overflow on the last rank only (every rank skips, weights repaired everywhere)."""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from apex_example_b200.parallel.peer import PeerBuffers, PeerShardedStep, shard_segments
    sizes = [5000, 17, 4096, 70001, 8, 12345, 9000, 1600000]
    offs, total = _layout(sizes)
    bufs = PeerBuffers(total, torch.float16, names=("grads", "model"), multicast=("grads", "model") if multicast != "none" else ())
    if multicast != "none" and rank == 0:
        print(f"[sharded step, world {world}] multicast address: {bool(bufs.mc_ptr('model'))} ({bufs.multicast_note})", flush=True)
    S = 256.0
    gen = torch.Generator().manual_seed(7)
    hp = [torch.randn(n, generator=gen) * 0.05 for n in sizes]
    hm, hv = [torch.zeros(n) for n in sizes], [torch.zeros(n) for n in sizes]
    hc = [h.half() for h in hp]
    flat_p = torch.zeros(total)
    for o, n, h in zip(offs, sizes, hp):
        flat_p[o:o + n] = h
    g_local, w_local = bufs.local["grads"], bufs.local["model"]
    w_local.copy_(flat_p.half())
    stepper = PeerShardedStep(bufs, offs, sizes, masters=flat_p, use_multicast=multicast)
    in_switch = stepper.multicast_reduce
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device="cuda")
    ops.scaler_init(state, S)

    def launch(step_no, lr=1e-3, wd=0.01):
        stepper.step([ops.AdamGroup(lr, 0.9, 0.999, 1e-8, wd, step_no, 1, 1)], state)

    scale, good, par = S, 0, 0
    for it in range(4):
        host_all = [(torch.randn(total, generator=torch.Generator().manual_seed(100 * it + r)) * 1e-2 * scale).half()
                    for r in range(world)]
        if it == 1:
            host_all[world - 1][total - 1] = float("inf")          # overflow on the LAST rank only
        g_local.copy_(host_all[rank])
        launch(good + 1)
        torch.cuda.synchronize()
        dist.barrier()
        want = torch.zeros(total, dtype=torch.float16)
        wflag = oracle.peer_reduce(host_all, want, 0, total, 1.0 / world)
        assert wflag == (1 if it == 1 else 0)
        assert int(state.cpu()[ops.IDX_FOUND_INF]) == wflag, f"rank {rank}: skip decision at iteration {it}"
        if not wflag:
            # every rank left the mean of ITS shard in its own gradient buffer
            parts = [torch.empty_like(g_local) for _ in range(world)]
            dist.all_gather(parts, g_local)
            b = stepper.bounds
            mean = want.clone()                  # (the alignment gaps between tensors are not part of any shard's table)
            for r in range(world):
                for _, _, n, f in shard_segments(offs, sizes, b[r], b[r + 1]):
                    mean[f:f + n] = parts[r][f:f + n].cpu()
            if in_switch:       # addend order and partial-sum precision inside the switch are unspecified: <= world ulps
                ka, kb = (t.view(torch.int16).to(torch.int32) for t in (mean, want))
                ka, kb = torch.where(ka < 0, -(ka & 0x7fff), ka), torch.where(kb < 0, -(kb & 0x7fff), kb)
                assert int((ka - kb).abs().max()) <= world, f"rank {rank}: in-switch mean differs by more than {world} ulps"
            else:
                assert torch.equal(mean.view(torch.int16), want.view(torch.int16)), f"rank {rank}: mean"
            good += 1
            par ^= 1
            for i, (o, n) in enumerate(zip(offs, sizes)):
                oracle.adam(mean[o:o + n].clone(), hp[i], hm[i], hv[i], hc[i], lr=1e-3, weight_decay=0.01, step=good,
                            inv_scale=1.0 / scale)
        else:
            scale /= 2
        assert int(stepper.parity) == par
        for k, h in (("p", hp), ("m", hm), ("v", hv)):
            assert_lists_bit_equal(stepper.current(k, par), [h[t][a:a + n] for t, a, n, _ in stepper.segments],
                                   f"{k} (rank {rank}, it {it})")
        assert_lists_bit_equal([w_local[o:o + n] for o, n in zip(offs, sizes)], hc, f"model (rank {rank}, it {it})")
        dist.barrier()
    for _ in range(repeats):                 # hammer the hand-shake: back-to-back steps, no host sync
        launch(good + 1, lr=0.0, wd=0.0)
    torch.cuda.synchronize()
    bufs.check_signals()
    del stepper
    bufs.close()
    dist.destroy_process_group()


def _worker_sharded_training(rank, world, port):
    """Real ranks: a small fp16 model whose weights live in the peer-addressable model buffer and whose gradients are
    produced straight into the flat gradient buffer, trained with PeerShardedStep."""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from apex_example_b200.parallel.peer import PeerBuffers, PeerShardedStep
    torch.manual_seed(0)                                             # the same initial weights on every rank
    model = torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.ReLU(), torch.nn.Linear(256, 10)).cuda().half()
    params = list(model.parameters())
    sizes = [p.numel() for p in params]
    offs, total = _layout(sizes)
    mc = ("grads", "model") if world >= 4 else ()
    bufs = PeerBuffers(total, torch.float16, names=("grads", "model"), multicast=mc)
    g_flat, w_flat = bufs.local["grads"], bufs.local["model"]
    for p, o, n in zip(params, offs, sizes):
        w_flat[o:o + n].copy_(p.data.view(-1))
        p.data = w_flat[o:o + n].view_as(p)
        p.grad = g_flat[o:o + n].view_as(p)
    stepper = PeerShardedStep(bufs, offs, sizes, use_multicast="both" if world >= 4 else "none")
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device="cuda")
    ops.scaler_init(state, 1024.0)
    fs = state.view(torch.float32)
    gen = torch.Generator().manual_seed(1)
    x_all = torch.randn(32 * world, 64, generator=gen)
    y_all = torch.randint(0, 10, (32 * world,), generator=gen)
    x, y = x_all[rank * 32:(rank + 1) * 32].cuda().half(), y_all[rank * 32:(rank + 1) * 32].cuda()
    losses, good = [], 0
    for it in range(14):
        g_flat.zero_()
        loss = torch.nn.functional.cross_entropy(model(x).float(), y)
        scaled = loss * fs[ops.IDX_LOSS_SCALE]
        if it == 5 and rank == world - 1:
            scaled = scaled * float("inf")                          # overflow on the last rank only
        scaled.backward()
        assert params[0].grad.data_ptr() == g_flat.data_ptr() + 2 * offs[0], "autograd replaced the gradient view"
        before = w_flat.clone()
        stepper.step([ops.AdamGroup(1e-2, 0.9, 0.999, 1e-8, 0.0, good + 1, 1, 1)], state)
        torch.cuda.synchronize()
        found = int(state.cpu()[ops.IDX_FOUND_INF])
        assert found == (1 if it == 5 else 0), f"rank {rank}: skip decision at iteration {it}"
        if found:
            assert torch.equal(w_flat, before), f"rank {rank}: weights changed on a skipped step"
        else:
            good += 1
            assert not torch.equal(w_flat, before)
            losses.append(float(loss))
        dist.barrier()
    assert all(l == l and abs(l) < 1e4 for l in losses)
    mean_loss = torch.tensor([losses[0], losses[-1]], device="cuda")
    dist.all_reduce(mean_loss)
    assert float(mean_loss[1]) < 0.8 * float(mean_loss[0]), f"loss did not fall: {mean_loss.tolist()}"
    parts = [torch.empty_like(w_flat) for _ in range(world)]
    dist.all_gather(parts, w_flat)
    assert all(torch.equal(parts[0], q) for q in parts[1:]), "ranks hold different weights"
    assert float(fs[ops.IDX_LOSS_SCALE]) == 512.0
    torch.cuda.synchronize()
    bufs.check_signals()
    for p in params:
        p.grad = None
    del stepper, params, model, g_flat, w_flat, parts, before
    bufs.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_real_rank_training_with_the_sharded_step(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs (run under gpurun --gpus {world})")
    mp.spawn(_worker_sharded_training, args=(world, _free_port()), nprocs=world, join=True)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("multicast", ["none", "stores", "both"])
def test_real_rank_sharded_step(world, multicast):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs (run under gpurun --gpus {world})")
    mp.spawn(_worker_sharded, args=(world, _free_port(), 100, multicast), nprocs=world, join=True)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_real_rank_fused_reduce_and_speculative_step(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs (run under gpurun --gpus {world})")
    mp.spawn(_worker_fused, args=(world, _free_port(), 100), nprocs=world, join=True)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("barrier,repeats", [("nccl", 1), ("signal", 200)])
def test_real_rank_ipc_reduce_and_fused_pull(barrier, repeats, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs (run under gpurun --gpus {world})")
    mp.spawn(_worker_ipc, args=(world, _free_port(), barrier, repeats), nprocs=world, join=True)
