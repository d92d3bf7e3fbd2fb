#!/usr/bin/env python
"""Kernel-level timing sweep over launch configurations (CUDA events, inputs >> L2).
Usage on the GPU box:  python tools/tune.py [--workload gpt2_small] > gpurun_out/tune.txt"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

from apex_example_b200 import ops  # noqa: E402
from apex_example_b200.tables import TensorTable  # noqa: E402
from workloads import GPT2_MEDIUM, GPT2_SMALL, many_small  # noqa: E402


def timeit(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in evs:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="gpt2_small")
    ap.add_argument("--tiles", default="4096")
    ap.add_argument("--threads", default="256,512")
    ap.add_argument("--unrolls", default="1,2")
    ap.add_argument("--ctas", default="0,1,2,3,4")
    args = ap.parse_args()
    sizes = {"gpt2_small": GPT2_SMALL, "gpt2_medium": GPT2_MEDIUM,
             "many_small": many_small(total=300_000_000)}[args.workload]
    n = sum(sizes)
    dev = "cuda"
    g = [torch.randn(s, device=dev).mul_(64).half() for s in sizes]
    c = [torch.randn(s, device=dev).mul_(0.02).half() for s in sizes]
    p = [x.float() for x in c]
    m = [torch.zeros(s, device=dev) for s in sizes]
    v = [torch.zeros(s, device=dev) for s in sizes]
    mg = [torch.zeros(s, device=dev) for s in sizes]
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    flat = torch.randn(n, device=dev).half()
    print(f"workload {args.workload}: {len(sizes)} tensors, {n} elements")
    for tile in [int(t) for t in args.tiles.split(",")]:
        t5 = TensorTable([g, p, m, v, c], tile_elems=tile)
        t1 = TensorTable([g], tile_elems=tile)
        t2 = TensorTable([g, mg], tile_elems=tile)
        t4 = TensorTable([g, p, m, c], tile_elems=tile)
        for threads in [int(x) for x in args.threads.split(',')]:
            for unroll in [int(x) for x in args.unrolls.split(',')]:
                for ctas in [int(x) for x in args.ctas.split(',')]:
                    ops.set_launch_config(ctas, threads, unroll)
                    step = [0]

                    def adam():
                        step[0] += 1
                        ops.fused_adam(t5, torch.float16, torch.float32, torch.float16,
                                       [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.01, step[0], 1, 1)], inv_scale=1 / 64)

                    def sgd():
                        ops.fused_sgd(t4, torch.float16, torch.float32, torch.float16,
                                      [ops.SgdGroup(0.1, 0.9, 0.0, 1e-4, 2, 0, 0)], inv_scale=1 / 64)

                    res = {
                        "adam28": (timeit(adam), 28 * n),
                        "sgd20": (timeit(sgd), 20 * n),
                        "check2": (timeit(lambda: ops.multi_tensor_check(t1, 0, torch.float16, flag)), 2 * n),
                        "scale6": (timeit(lambda: ops.multi_tensor_scale(t2, torch.float16, torch.float32, 1 / 64, dev_flag=flag)), 6 * n),
                    }
                    line = f"tile {tile:6d} thr {threads} unroll {unroll} ctas/sm {ctas}: "
                    for k, ((med, best), nbytes) in res.items():
                        line += f"{k} {med*1e3:7.1f}us {nbytes/med/1e6:7.0f}GB/s | "
                    print(line, flush=True)
    ops.set_launch_config(0, 256, 2)
    # peer pull emulated on one GPU: "reduced" shards live in two other local buffers
    from apex_example_b200.tables import tile_aligned_shards
    gflat = torch.randn(n, device=dev).mul_(64).half()
    offs, acc = [], 0
    for s_ in sizes:
        offs.append(acc); acc += s_
    for world in (2, 8):
        red = [gflat.clone() for _ in range(world)]
        t5 = TensorTable([list(torch.split(gflat, sizes)), p, m, v, c])
        pull = ops.make_peer_pull(gflat.data_ptr(), tile_aligned_shards(offs, sizes, n, world), [r.data_ptr() for r in red], 2)
        med, best = timeit(lambda: ops.fused_adam(t5, torch.float16, torch.float32, torch.float16,
                                                  [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.01, 5, 1, 1)], inv_scale=1 / 64, pull=pull))
        print(f"adam28 with pull emulated locally, world {world}: {med*1e3:.1f}us {28*n/med/1e6:.0f}GB/s")
        out = torch.empty_like(gflat)
        b = tile_aligned_shards(offs, sizes, n, world)
        med, best = timeit(lambda: ops.peer_reduce_shard([r.data_ptr() for r in red], out, b[0], b[1], 1.0 / world, flag))
        print(f"peer_reduce_shard emulated locally, world {world}: shard {b[1]-b[0]} elems {med*1e3:.1f}us "
              f"{(world+1)*2*(b[1]-b[0])/med/1e6:.0f}GB/s (HBM bytes)")
        del red, out
    tn = TensorTable([g])
    sumsq, scratch = torch.zeros(1, device=dev), ops.new_norm_scratch(dev)
    med, best = timeit(lambda: ops.multi_tensor_l2norm(tn, 0, torch.float16, sumsq, scratch, flag))
    print(f"l2norm+check (2B/elem): {med*1e3:.1f}us {2*n/med/1e6:.0f}GB/s")
    med, best = timeit(lambda: ops.bucket_epilogue(flat, 0.125, flag))
    print(f"bucket_epilogue rw (4B/elem): {med*1e3:.1f}us {4*n/med/1e6:.0f}GB/s")
    med, best = timeit(lambda: ops.bucket_epilogue(flat, 1.0, flag))
    print(f"bucket_epilogue check-only (2B/elem): {med*1e3:.1f}us {2*n/med/1e6:.0f}GB/s")
    a, b = torch.empty(n, device=dev), torch.empty(n, device=dev)
    med, best = timeit(lambda: b.copy_(a))
    print(f"torch copy_ fp32 ({8*n/1e9:.2f} GB): {med*1e3:.1f}us {8*n/med/1e6:.0f}GB/s (roofline denominator style)")


if __name__ == "__main__":
    main()
