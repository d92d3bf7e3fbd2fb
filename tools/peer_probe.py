#!/usr/bin/env python
"""NVLink probe for the peer-memory path (run under torchrun, one rank per GPU):
times the barrier all-reduces, the reduce-scatter kernel and plain peer copies separately."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from apex_example_b200 import ops  # noqa: E402
from apex_example_b200.parallel.peer import PeerBuffers, _wrap  # noqa: E402


def timeit(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / reps], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = 124_439_808
    bufs = PeerBuffers(n, torch.float16)
    bufs.local["grads"].copy_(torch.randn(n, device="cuda").half())
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    shard = n // world // 8 * 8
    b0, b1 = rank * shard, (rank + 1) * shard
    gp = bufs.ptrs("grads")
    res = {}
    res["barrier_x2_ms"] = timeit(lambda: (bufs.barrier(), bufs.barrier(flag)))
    res["reduce_kernel_ms"] = timeit(lambda: ops.peer_reduce_shard(gp, bufs.local["reduced"], b0, b1, 1.0 / world, flag))
    inbound = (world - 1) * shard * 2
    res["reduce_inbound_gbs"] = inbound / res["reduce_kernel_ms"] / 1e6
    nxt = (rank + 1) % world
    peer = _wrap(gp[nxt], n, torch.float16, torch.device("cuda", local))
    mine = bufs.local["reduced"]
    k = 32 * 1024 * 1024
    res["pull_copy_64MB_gbs"] = 2 * k / timeit(lambda: mine[:k].copy_(peer[:k])) / 1e6
    rp = _wrap(bufs.ptrs("reduced")[nxt], n, torch.float16, torch.device("cuda", local))
    res["push_copy_64MB_gbs"] = 2 * k / timeit(lambda: rp[:k].copy_(bufs.local["grads"][:k])) / 1e6
    if rank == 0:
        print(world, {k_: round(v, 4) for k_, v in res.items()}, flush=True)
    del peer, rp, mine
    bufs.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
