#!/usr/bin/env python
"""Timing probe for b200_peer_spec_step on the GPT-2-small list (run under torchrun, N >= 2).
With a library built with -DB200_STEP_PROFILE (APEX_B200_SO=...) it also prints where the CTAs' time goes."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from apex_example_b200 import _lib, ops  # noqa: E402
from apex_example_b200.parallel.peer import PeerBuffers, PeerSpecStep  # noqa: E402
from apex_example_b200.tables import TensorTable, tile_aligned_shards  # noqa: E402
from workloads import GPT2_SMALL  # noqa: E402


def main():
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    chunk = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    sizes = GPT2_SMALL
    n = sum(sizes)
    offs, acc = [], 0
    for s_ in sizes:
        offs.append(acc)
        acc += s_
    bufs = PeerBuffers(n, torch.float16)
    bounds = tile_aligned_shards(offs, sizes, n, world)
    g = bufs.local["grads"]
    g.copy_((torch.randn(n, device=dev) * 65.0).half())
    copy = [(torch.randn(s, device=dev) * 0.02).half() for s in sizes]
    pb = [ops.spec_state_like(c.float(), fill=c.float()) for c in copy]
    mb = [ops.spec_state_like(c.float()) for c in copy]
    vb = [ops.spec_state_like(c.float()) for c in copy]
    table = TensorTable([list(torch.split(g, sizes)), [x for x, _ in pb], [x for x, _ in mb], [x for x, _ in vb], copy])
    stepper = PeerSpecStep(bufs, bounds, table, chunk_elems=chunk or None, n_reducers=int(os.environ.get('REDUCERS', '0')) or None)
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device=dev)
    ops.scaler_init(state, 65536.0)
    fs = state.view(torch.float32)
    parity = torch.zeros(1, dtype=torch.int32, device=dev)

    def step(k):
        stepper.step(torch.float16, torch.float32, torch.float16, [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.01, k, 1, 1)],
                     state[ops.IDX_OVERFLOW_ACC:], parity, dev_inv_scale=fs[ops.IDX_INV_LOSS_SCALE:])
        ops.spec_commit(state, True, 2000, 2.0, 0.0, 2.0 ** 24, parity)
        ops.spec_repair(stepper.table, torch.float32, torch.float16, state[ops.IDX_FOUND_INF:], parity)

    for k in range(5):
        step(k + 1)
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(20):
        step(6 + k)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    res = {"rank": rank, "world": world, "ms_per_step": ms, "chunk_elems": stepper.chunk_elems, "so": _lib.SO_PATH}
    prof = stepper.scratch[2048:2048 + 20 * 512].view(torch.int64).view(512, 10).cpu()
    if int(prof[:, 8].sum()) > 0:
        live = prof[prof[:, 8] > 0].double()
        names = ["setup", "top(next slot, tickets)", "piece issue", "tile", "wait+fold", "barrier A", "count+ready", "barrier B"]
        tot = live[:, :8].sum(1).mean()
        res["profile"] = {"ctas": int(live.shape[0]), "iters_per_cta": float(live[:, 8].mean()), "tiles_per_cta": float(live[:, 9].mean()),
                          "kcycles_per_cta": float(tot / 1e3),
                          "share": {nm: round(float(live[:, i].mean() / tot), 3) for i, nm in enumerate(names)}}
    print(json.dumps(res), flush=True)
    del table, stepper
    bufs.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
