"""-m gpu tests of apex.parallel.DistributedDataParallel over NCCL.  The single-rank cases run on
one GPU; the 2-rank cases need `gpurun --gpus 2` and skip themselves otherwise."""
import os
import subprocess
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from test_ddp_gloo import _free_port

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _init(rank, world, port):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))


def _mlp(seed):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.ReLU(), torch.nn.Linear(256, 256), torch.nn.ReLU(),
                               torch.nn.Linear(256, 16))


def _worker_train(rank, world, port, kwargs, out):
    _init(rank, world, port)
    from apex import amp
    from apex.optimizers import FusedAdam, FusedSGD
    from apex.parallel import DistributedDataParallel
    kwargs = dict(kwargs)
    which = kwargs.pop("optimizer", "fused_adam")
    model = _mlp(rank).cuda()                       # different per rank; DDP must broadcast rank 0's
    if which == "fused_adam":
        opt = FusedAdam(model.parameters(), lr=1e-3, weight_decay=0.01)
    elif which == "fused_sgd":                      # set_grad_none=False: gradients stay views of the buckets
        opt = FusedSGD(model.parameters(), lr=1e-2, momentum=0.9)
    else:                                           # the reference script's own choice (test_apex_distributed_spawn.py:119)
        opt = torch.optim.SGD(model.parameters(), 1e-4)
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    ddp = DistributedDataParallel(model, **kwargs)
    gen = torch.Generator().manual_seed(50 + rank)
    scales = []
    for it in range(8):
        x = torch.randn(32, 64, generator=gen).cuda()
        if it in (0, 4) and rank == world - 1:
            x = x * float("inf")                    # overflow on ONE rank only: every rank must skip
        loss = ddp(x).float().pow(2).mean()
        opt.zero_grad()
        if it >= 1 and which in ("fused_sgd", "torch_sgd"):
            # zero_grad without set-to-none: the gradients are still the bucket views and every bucket was
            # cleared as a whole (one memset per bucket)
            for b in ddp.buckets:
                assert all(p.grad is v for p, v in zip(b.params, b.views))
                assert float(b.flat.float().abs().max()) == 0.0
        with amp.scale_loss(loss, opt) as scaled:
            scaled.backward()
        # post-allreduce gradients are bit-identical on every rank
        for p in model.parameters():
            g = p.grad.detach().float().clone()
            lo, hi = g.clone(), g.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            same = (lo == hi) | (lo.isnan() & hi.isnan()) | (g.isnan())
            assert bool(same.all()), "gradients differ across ranks after the bucket all-reduce"
        opt.step()
        scales.append(amp._amp_state.loss_scalers[0].loss_scale())
    assert scales == [32768.0] * 4 + [16384.0] * 4, scales
    for p in list(model.parameters()) + list(amp.master_params(opt)):
        v = p.detach().float().clone()
        lo, hi = v.clone(), v.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        assert torch.equal(lo, hi), "parameters diverged across ranks"
    if rank == 0:
        out.put((scales, len(ddp.buckets)))
    ddp.close()
    dist.destroy_process_group()


def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs (run under gpurun --gpus {n})")


_TRAIN_CASES = [dict(), dict(message_size=20000), dict(delay_allreduce=True),
                dict(message_size=20000, num_allreduce_streams=2),
                dict(allreduce_always_fp32=True), dict(gradient_predivide_factor=2.0),
                dict(optimizer="torch_sgd"), dict(optimizer="fused_sgd", message_size=20000),
                dict(allreduce_backend="peer"), dict(allreduce_backend="peer", message_size=20000),
                dict(allreduce_backend="peer", delay_allreduce=True, num_allreduce_streams=2),
                dict(allreduce_backend="peer", message_size=10)]      # every bucket closes at once: tiny 16-element buckets


@pytest.mark.parametrize("world,kwargs", [(2, k) for k in _TRAIN_CASES] +
                         [(w, k) for w in (4, 8) for k in (dict(), dict(optimizer="torch_sgd"),
                                                            dict(allreduce_backend="peer", message_size=20000),
                                                            dict(allreduce_backend="peer", message_size=10))])
def test_multi_rank_training_identical_ranks_and_skips(world, kwargs):
    """One rank's inf must make EVERY rank skip; gradients after the all-reduce and parameters after the steps
    are bit-identical on all ranks (SURVEY.md §3.4, §4 "distributed GPU" tier) — at world 2, 4 and 8."""
    _need(world)
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    mp.spawn(_worker_train, args=(world, _free_port(), kwargs, out), nprocs=world, join=True)
    scales, n_buckets = out.get()
    ms = kwargs.get("message_size")
    assert n_buckets == (6 if ms == 10 else 2 if ms else 1)           # [16,4096,256,65536] | [256,16384]


def _convnet(seed):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Conv2d(8, 16, 3, padding=1), torch.nn.ReLU(), torch.nn.Conv2d(16, 8, 3, padding=1),
                               torch.nn.Flatten(), torch.nn.Linear(8 * 6 * 6, 4)).to(memory_format=torch.channels_last)


def _worker_mean(rank, world, port, backend="nccl", conv=False):
    _init(rank, world, port)
    from apex.parallel import DistributedDataParallel
    make = _convnet if conv else _mlp                # conv: channels_last weights -> bucket views in that layout
    model = make(0).cuda().half()
    ref = make(0).cuda().half()
    ddp = DistributedDataParallel(model, message_size=30000, allreduce_backend=backend)
    gen = torch.Generator().manual_seed(9 + rank)
    x = (torch.randn(16, 8, 6, 6, generator=gen).cuda().half().contiguous(memory_format=torch.channels_last) if conv
         else torch.randn(16, 64, generator=gen).cuda().half())
    ddp(x).float().pow(2).mean().backward()
    ref(x).float().pow(2).mean().backward()
    for p, q in zip(model.parameters(), ref.parameters()):
        want = q.grad.clone()
        dist.all_reduce(want)                       # fp16 SUM, then x 1/world in fp16: apex's order
        want.mul_(1.0 / world)
        # two addends: the fp16 sum is the correctly rounded exact sum whichever way it is accumulated, so the
        # NCCL path and the peer path (fp32 accumulation, one rounding) must both reproduce it exactly
        assert torch.equal(p.grad, want)
        assert p.grad.stride() == p.stride() or p.grad.is_contiguous() and p.is_contiguous()
    ddp.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("conv", [False, True])
@pytest.mark.parametrize("backend", ["nccl", "peer"])
def test_two_rank_bucket_mean_is_sum_then_scale_in_grad_dtype(backend, conv):
    _need(2)
    mp.spawn(_worker_mean, args=(2, _free_port(), backend, conv), nprocs=2, join=True)


def _worker_mean_oracle(rank, world, port):
    """Peer backend at any world size: the bucket mean must equal the oracle's rank-ordered fp32 sum, rounded
    to fp16, times 1/world — bit for bit, on every rank."""
    _init(rank, world, port)
    import oracle
    from apex.parallel import DistributedDataParallel
    model = _mlp(0).cuda().half()
    ddp = DistributedDataParallel(model, message_size=30000, allreduce_backend="peer")
    refs = []
    for r in range(world):                          # every rank replays every rank's local backward
        ref = _mlp(0).cuda().half()
        x = torch.randn(16, 64, generator=torch.Generator().manual_seed(9 + r)).cuda().half()
        ref(x).float().pow(2).mean().backward()
        refs.append([q.grad.detach().cpu().reshape(-1) for q in ref.parameters()])
    x = torch.randn(16, 64, generator=torch.Generator().manual_seed(9 + rank)).cuda().half()
    ddp(x).float().pow(2).mean().backward()
    torch.cuda.synchronize()
    for i, p in enumerate(model.parameters()):
        want = torch.zeros(p.numel(), dtype=torch.float16)
        oracle.peer_reduce([refs[r][i] for r in range(world)], want, 0, p.numel(), 1.0 / world)
        assert torch.equal(p.grad.detach().cpu().reshape(-1).view(torch.int16), want.view(torch.int16)), f"param {i}"
    ddp.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_peer_bucket_mean_equals_oracle_at_every_world_size(world):
    _need(world)
    mp.spawn(_worker_mean_oracle, args=(world, _free_port()), nprocs=world, join=True)


def _worker_single(rank, world, port):
    _init(rank, world, port)
    from apex import amp
    from apex.optimizers import FusedSGD
    from apex.parallel import DistributedDataParallel
    model = _mlp(3).cuda()
    opt = FusedSGD(model.parameters(), lr=0.1, momentum=0.9, materialize_master_grads=False)
    model, opt = amp.initialize(model, opt, opt_level="O2", verbosity=0)
    with pytest.raises(RuntimeError, match="AFTER"):
        amp.initialize(DistributedDataParallel(_mlp(4).cuda()), None, opt_level="O2", verbosity=0)
    ddp = DistributedDataParallel(model)
    x = torch.randn(8, 64, device="cuda")
    before = [p.detach().clone() for p in model.parameters()]
    for _ in range(3):
        loss = ddp(x).float().pow(2).mean()
        opt.zero_grad()
        with amp.scale_loss(loss, opt) as scaled:
            scaled.backward()
        opt.step()
    torch.cuda.synchronize()
    assert all(not torch.equal(a, p.detach()) for a, p in zip(before, model.parameters()))
    dist.destroy_process_group()


def test_single_rank_nccl_wrapper_is_a_pass_through():
    mp.spawn(_worker_single, args=(1, _free_port()), nprocs=1, join=True)


@pytest.mark.parametrize("optimizer", ["sgd", "fused_adam"])
def test_reference_workload_example_runs(optimizer):
    """The reference's flow (examples/convnet_ddp_spawn.py mirrors its call order) on 1 GPU, amp O2."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "convnet_ddp_spawn.py"), "--gpus", "1", "--steps",
                        "100", "--optimizer", optimizer, "--port", str(_free_port())], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Training complete in" in r.stdout and "identical across ranks: True" in r.stdout
    losses = [float(l.rsplit("Loss:", 1)[1]) for l in r.stdout.splitlines() if "Loss:" in l]
    assert len(losses) == 2 and all(l == l for l in losses)
    if optimizer == "fused_adam":
        assert losses[-1] < losses[0]


def test_the_reference_script_itself_runs_through_the_drop_in():
    """SURVEY.md §4 "drop-in API" tier: the reference file, unmodified, one short epoch at --gpus 1 with a synthetic
    MNIST stand-in (examples/run_reference.py).  The file is not part of this repository and must not be copied
    into it, so the test can only run where /root/reference (or $APEX_EXAMPLE_REF) exists."""
    ref = os.environ.get("APEX_EXAMPLE_REF", "/root/reference/test_apex_distributed_spawn.py")
    if not os.path.exists(ref):
        pytest.skip("the reference script is not on this machine (it does not travel with the repository)")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "run_reference.py"), "--ref", ref, "--gpus", "1",
                        "--epochs", "1", "--port", str(_free_port())], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    losses = [float(l.rsplit("Loss:", 1)[1]) for l in r.stdout.splitlines() if "Loss:" in l]
    assert len(losses) == 6 and all(l == l for l in losses) and "Training complete in" in r.stdout
