#!/usr/bin/env python
"""Single-GPU timing of b200_peer_spec_step's loop: world 2 where rank 1 is a GHOST whose hand-shake and progress
flags are pre-set, so rank 0's kernel runs alone at full grid with every buffer local (no NVLink).  Separates the
kernel's own efficiency from interconnect effects; can be run under ncu."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

from apex_example_b200 import _lib, ops  # noqa: E402
from apex_example_b200.tables import TensorTable, tile_aligned_shards  # noqa: E402
from workloads import GPT2_SMALL  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    sizes = GPT2_SMALL
    n = sum(sizes)
    offs, acc = [], 0
    for s_ in sizes:
        offs.append(acc)
        acc += s_
    world = 2
    bounds = tile_aligned_shards(offs, sizes, n, world)
    grads = [(torch.randn(n, device=dev) * 65.0).half() for _ in range(world)]
    reduced = [(torch.randn(n, device=dev) * 65.0).half() for _ in range(world)]      # the ghost's shard keeps these values
    pads = torch.zeros(world, ops.PEER_STEP_PAD_WORDS, dtype=torch.int32, device=dev)
    BIG = 1 << 20
    pads[0, 1] = BIG                                            # ready[1] on rank 0's pad
    pads[0, 32 + _lib.PEER_STEP_MAX_CHUNKS: 32 + 2 * _lib.PEER_STEP_MAX_CHUNKS] = BIG      # board[owner 1][*]
    pad_ptrs = [pads.data_ptr() + 4 * ops.PEER_STEP_PAD_WORDS * r for r in range(world)]
    copy = [(torch.randn(s, device=dev) * 0.02).half() for s in sizes]
    pb = [ops.spec_state_like(c.float(), fill=c.float()) for c in copy]
    mb = [ops.spec_state_like(c.float()) for c in copy]
    vb = [ops.spec_state_like(c.float()) for c in copy]
    chunk = 1 << 19
    table = TensorTable([list(torch.split(grads[0], sizes)), [x for x, _ in pb], [x for x, _ in mb], [x for x, _ in vb], copy])
    table.order_for_peer_step(grads[0].data_ptr(), 2, bounds, 0, chunk)
    plain = TensorTable([list(torch.split(grads[0], sizes)), [x for x, _ in pb], [x for x, _ in mb], [x for x, _ in vb], copy])
    pull = ops.make_peer_pull(grads[0].data_ptr(), bounds, [t.data_ptr() for t in reduced], 2)
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device=dev)
    ops.scaler_init(state, 65536.0)
    fs = state.view(torch.float32)
    parity = torch.zeros(1, dtype=torch.int32, device=dev)
    scratch = ops.new_peer_step_scratch(dev)
    epoch = [0]

    def fused(k):
        epoch[0] += 1
        ops.peer_spec_step(table, torch.float16, torch.float32, torch.float16, [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.01, k, 1, 1)],
                           state[ops.IDX_OVERFLOW_ACC:], parity, pull, [g.data_ptr() for g in grads], reduced[0], 0.5,
                           ops.make_peer_sync(pad_ptrs, 0, epoch[0]), int(os.environ.get('REDUCERS', '32')), chunk, scratch, dev_inv_scale=fs[ops.IDX_INV_LOSS_SCALE:])

    def plain_spec(k):
        ops.fused_adam_spec(plain, torch.float16, torch.float32, torch.float16, [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.01, k, 1, 1)],
                            state[ops.IDX_OVERFLOW_ACC:], parity, dev_inv_scale=fs[ops.IDX_INV_LOSS_SCALE:], pull=pull)

    res = {"so": _lib.SO_PATH}
    for name, fn in (("fused_ghost", fused), ("plain_spec_with_pull", plain_spec)):
        for k in range(3):
            fn(k + 1)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(10):
            fn(4 + k)
        e1.record()
        torch.cuda.synchronize()
        res[name + "_ms"] = e0.elapsed_time(e1) / 10
    res["timed_out"] = int(pads[:, 16].abs().sum())
    print(json.dumps(res))


if __name__ == "__main__":
    main()
