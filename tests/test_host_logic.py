"""CPU tests of host-side logic that needs no GPU: plan re-validation, the scaler's keep-skip mode."""
import torch

import oracle
from apex_example_b200.optimizers._fused_base import ListWatch


class _Holder:
    def __init__(self, g):
        self.grad = g


def test_listwatch_small_lists_are_checked_completely_every_step():
    gs = [torch.zeros(3) for _ in range(10)]
    hs = [_Holder(g) for g in gs]
    ts = [torch.zeros(4) for _ in range(10)]
    w = ListWatch(hs, list(gs), ts)
    assert w.ok() and w.ok()
    hs[7].grad = torch.zeros(3)                 # same values, other object
    assert not w.ok()
    hs[7].grad = gs[7]
    assert w.ok()
    ts[9].data = torch.zeros(4)                 # storage moved, tensor object unchanged
    assert not w.ok()


def test_listwatch_large_lists_rotate_and_catch_a_change_within_one_cycle():
    n = 5 * ListWatch.WINDOW + 17
    gs = [torch.zeros(1) for _ in range(n)]
    hs = [_Holder(g) for g in gs]
    ts = [torch.zeros(1) for _ in range(n)]
    w = ListWatch(hs, list(gs), ts)
    for _ in range(20):
        assert w.ok()
    hs[n - 3].grad = None
    cycle = -(-n // ListWatch.WINDOW)
    assert not all(w.ok() for _ in range(cycle))
    hs[n - 3].grad = gs[n - 3]
    ts[2 * ListWatch.WINDOW + 5].data = torch.zeros(1)
    assert not all(w.ok() for _ in range(cycle))


def test_scaler_keep_skip_mode_matches_apex_one_shot_skip_step():
    """Overflow in micro-batch 1, clean micro-batch 2, no step in between: the step stays skipped, the
    scale is halved once, the window counter sees one clean backward."""
    s = oracle.LossScaler(loss_scale=2.0 ** 16, scale_window=3)
    assert s.update_scale(True) is True
    assert s.loss_scale == 2.0 ** 15
    assert s.update_scale(False, keep_skip=True) is True          # still skipped
    assert s.loss_scale == 2.0 ** 15 and s.unskipped == 1
    assert s._s.skipped_total == 1
    assert s.update_scale(False) is False                         # after a step(): fresh decision
    # without the mode bit the old behaviour is unchanged
    t = oracle.LossScaler(loss_scale=2.0 ** 16, scale_window=3)
    t.update_scale(True)
    assert t.update_scale(False) is False


def test_even_shards_and_shard_segments():
    from apex_example_b200.parallel.peer import even_shards, shard_segments
    sizes = [5000, 17, 4096, 70001, 8]
    offs, o = [], 0
    for n in sizes:
        offs.append(o)
        o += (n + 15) // 16 * 16
    total = o
    for world in (1, 2, 3, 8):
        b = even_shards(total, world)
        assert b[0] == 0 and b[-1] == total and len(b) == world + 1 and b == sorted(b)
        assert all(x % 1024 == 0 for x in b[1:-1])
        covered = [0] * len(sizes)
        for r in range(world):
            for t, first, n, flat in shard_segments(offs, sizes, b[r], b[r + 1]):
                assert flat == offs[t] + first and n > 0 and b[r] <= flat and flat + n <= b[r + 1]
                covered[t] += n
        assert covered == sizes                                       # every element belongs to exactly one rank
    assert even_shards(100, 4) == [0, 100, 100, 100, 100]             # tiny buffers: trailing shards are empty
