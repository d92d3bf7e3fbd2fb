"""CPU, world_size-2 (gloo) tests of the data-parallel host logic: bucket layout, hook-driven
reduction, averaging rule, structure check — with the bucket kernels replaced by the oracle
(injected through the wrapper's test seam; the product itself refuses CPU tensors).  Also
BASELINE.json config 1: the reference's own ConvNet under torch DDP/gloo with torch.optim.Adam."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/test_apex_distributed_spawn.py"


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


class OracleBucketOps:
    """CPU stand-in for the two bucket kernels, built on the oracle (tests only)."""

    def gather(self, srcs, dsts, scale, table_cache, key):
        import oracle
        for s, d in zip(srcs, dsts):
            oracle.scale(s.contiguous().view(-1), d.view(-1), scale)

    def epilogue(self, flat, factor, flag):
        import oracle
        oracle.bucket_epilogue(flat, factor)


def _init(rank, world, port):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)


def _model(seed):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(20, 50), torch.nn.ReLU(), torch.nn.Linear(50, 30), torch.nn.ReLU(),
                               torch.nn.Linear(30, 4))


def _worker_ddp(rank, world, port, kwargs, out):
    _init(rank, world, port)
    from apex_example_b200.parallel import DistributedDataParallel
    model = _model(100 + rank)                       # different init per rank: the ctor must broadcast rank 0's
    ddp = DistributedDataParallel(model, _bucket_ops=OracleBucketOps(), **kwargs)
    ref = _model(100)                                # == rank 0's parameters
    for p, q in zip(model.parameters(), ref.parameters()):
        assert torch.equal(p, q), "parameters were not broadcast from rank 0"
    g = torch.Generator().manual_seed(7 + rank)
    x = torch.randn(16, 20, generator=g)
    ddp(x).pow(2).sum().backward()
    ref(x).pow(2).sum().backward()
    pre = kwargs.get("gradient_predivide_factor", 1.0)
    for p, q in zip(model.parameters(), ref.parameters()):
        want = q.grad / pre
        dist.all_reduce(want)
        if kwargs.get("gradient_average", True):
            want = want * (pre / world)
        elif pre != 1.0:
            want = want * pre
        torch.testing.assert_close(p.grad, want, rtol=1e-6, atol=1e-7)
        # param.grad now lives inside a persistent bucket (no unflatten copy)
        b = ddp._bucket_of[id(p)]
        assert b.flat.data_ptr() <= p.grad.data_ptr() < b.flat.data_ptr() + b.flat.numel() * 4
    sizes = [sum(p.numel() for p in b.params) for b in ddp.buckets]
    if rank == 0:
        out.put((len(ddp.buckets), sizes))
    # second iteration accumulates into the bucket views in place and is reduced again
    ddp.zero_grad()
    ddp(x).pow(2).sum().backward()
    g2 = [p.grad.clone() for p in model.parameters()]
    for p, q, g1 in zip(model.parameters(), ref.parameters(), g2):
        want = q.grad / pre
        dist.all_reduce(want)
        if kwargs.get("gradient_average", True):
            want = want * (pre / world)
        elif pre != 1.0:
            want = want * pre
        torch.testing.assert_close(g1, want, rtol=1e-6, atol=1e-7)
    # disable_allreduce(): local gradients only
    ddp.disable_allreduce()
    for p in model.parameters():
        p.grad = None
    ddp(x).pow(2).sum().backward()
    for p, q in zip(model.parameters(), ref.parameters()):
        torch.testing.assert_close(p.grad, q.grad, rtol=1e-6, atol=1e-7)
    dist.destroy_process_group()


@pytest.mark.parametrize("kwargs,expect_buckets", [
    (dict(message_size=10_000_000), 1),
    (dict(message_size=100), 3),                     # reverse order: [4,120] | [30,1500] | [50,1000]
    (dict(message_size=1000), 2),                    # [4,120,30,1500] | [50,1000]
    (dict(message_size=100, delay_allreduce=True), 3),
    (dict(message_size=1000, gradient_predivide_factor=2.0), 2),
    (dict(gradient_average=False), 1),
    (dict(num_allreduce_streams=2, message_size=100), 3),
])
def test_ddp_two_ranks_gloo(kwargs, expect_buckets):
    port = _free_port()
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    mp.spawn(_worker_ddp, args=(2, port, kwargs, out), nprocs=2, join=True)
    n_buckets, sizes = out.get()
    assert n_buckets == expect_buckets and sum(sizes) == 20 * 50 + 50 + 50 * 30 + 30 + 30 * 4 + 4
    if kwargs.get("message_size", 10_000_000) < 10_000_000:
        assert all(s >= kwargs["message_size"] for s in sizes[:-1])


def _worker_options(rank, world, port):
    _init(rank, world, port)
    from apex_example_b200.parallel import DistributedDataParallel
    model = _model(5)
    params = list(model.parameters())
    # a bucket closes at each trigger parameter (reverse order: [4,120] | [30,1500,50] | [1000])
    ddp = DistributedDataParallel(model, allreduce_trigger_params=[params[4], params[1]], retain_allreduce_buffers=True,
                                  _bucket_ops=OracleBucketOps())
    assert [sum(p.numel() for p in b.params) for b in ddp.buckets] == [124, 1580, 1000]
    assert len(ddp.allreduce_buffers) == 3 and all(buf is b.flat for buf, b in zip(ddp.allreduce_buffers, ddp.buckets))
    x = torch.randn(8, 20, generator=torch.Generator().manual_seed(rank))
    ddp(x).pow(2).sum().backward()
    flat = torch.cat([b.flat for b in ddp.buckets])
    lo, hi = flat.clone(), flat.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert torch.equal(lo, hi)                              # the retained buffers hold the same mean on every rank
    with pytest.raises(ValueError, match="delay_allreduce"):
        DistributedDataParallel(_model(5), allreduce_trigger_params=[params[0]], delay_allreduce=True,
                                _bucket_ops=OracleBucketOps())
    dist.destroy_process_group()


def test_ddp_trigger_params_and_retained_buffers():
    mp.spawn(_worker_options, args=(2, _free_port()), nprocs=2, join=True)


def test_ddp_requires_process_group_and_cuda():
    sys.path.insert(0, ROOT)
    from apex_example_b200.parallel import DistributedDataParallel
    with pytest.raises(RuntimeError, match="torch.distributed to be initialized"):
        DistributedDataParallel(_model(0))


def _worker_cpu_refused(rank, world, port):
    _init(rank, world, port)
    from apex_example_b200.parallel import DistributedDataParallel
    with pytest.raises(TypeError, match="no CPU fallback"):
        DistributedDataParallel(_model(0))
    with pytest.raises(ValueError, match="shared_param"):
        DistributedDataParallel(_model(0), shared_param=True, _bucket_ops=OracleBucketOps())
    dist.destroy_process_group()


def test_ddp_refuses_cpu_parameters_without_the_test_seam():
    mp.spawn(_worker_cpu_refused, args=(1, _free_port()), nprocs=1, join=True)


# ---- BASELINE.json config 1 -------------------------------------------------------------------------
def _worker_config1(rank, world, port, out):
    _init(rank, world, port)
    import importlib.util
    import types
    for name in ("apex", "apex.parallel", "apex.amp"):       # the script imports apex at module scope
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["apex"].amp = sys.modules["apex.amp"]
    sys.modules["apex"].parallel = sys.modules["apex.parallel"]
    spec = importlib.util.spec_from_file_location("ref_script", REF)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    ref.set_seed(0)                                          # REF :113 precedes model construction :117
    model = ref.ConvNet()
    model = torch.nn.parallel.DistributedDataParallel(model)  # REF :124 with device_ids=None (CPU)
    opt = torch.optim.Adam(model.parameters(), 1e-4)
    crit = torch.nn.CrossEntropyLoss()
    g = torch.Generator().manual_seed(1234 + rank)
    for _ in range(5):
        x = torch.randn(100, 1, 28, 28, generator=g)
        y = torch.randint(0, 10, (100,), generator=g)
        loss = crit(model(x), y)
        opt.zero_grad()
        loss.backward()
        opt.step()
    out.put((rank, float(sum(p.double().sum() for p in model.parameters()))))
    dist.destroy_process_group()


@pytest.mark.skipif(not os.path.exists(REF), reason="reference script not present on this machine")
def test_config1_reference_convnet_torch_ddp_gloo_adam():
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    mp.spawn(_worker_config1, args=(2, _free_port(), out), nprocs=2, join=True)
    sums = dict(out.get() for _ in range(2))
    assert sums[0] == sums[1]


# ---- apex.parallel.Reducer (manual-trigger gradient mean) ---------------------------------------------
def _worker_reducer(rank, world, port, flat):
    _init(rank, world, port)
    from apex_example_b200.parallel import Reducer
    model = _model(200 + rank)
    red = Reducer(model, message_size=700)           # several chunks
    ref = _model(200)
    for p, q in zip(model.parameters(), ref.parameters()):
        assert torch.equal(p, q), "Reducer(module) must broadcast rank 0's parameters"
    g = torch.Generator().manual_seed(17 + rank)
    x = torch.randn(16, 20, generator=g)
    if flat:                                         # gradients as views of one flat arena: reduced in place
        n = sum(p.numel() for p in model.parameters())
        arena = torch.zeros(n)
        off = 0
        for p in model.parameters():
            p.grad = arena[off:off + p.numel()].view_as(p)
            off += p.numel()
    model(x).pow(2).sum().backward()
    ref(x).pow(2).sum().backward()
    ptrs = [p.grad.data_ptr() for p in model.parameters()]
    red.reduce()
    for p, q, ptr in zip(model.parameters(), ref.parameters(), ptrs):
        want = q.grad.clone()
        dist.all_reduce(want)
        torch.testing.assert_close(p.grad, want / world, rtol=1e-6, atol=1e-7)
        assert p.grad.data_ptr() == ptr              # gradients stay where they were
    # a list of gradient tensors instead of a module
    grads = [torch.full((5,), float(rank + 1)), torch.full((3, 2), float(10 * (rank + 1)))]
    Reducer(grads).reduce()
    assert torch.equal(grads[0], torch.full((5,), (1 + world) / 2.0))
    assert torch.equal(grads[1], torch.full((3, 2), 10 * (1 + world) / 2.0))
    dist.destroy_process_group()


@pytest.mark.parametrize("flat", [False, True])
def test_reducer_averages_gradients_on_demand_gloo(flat):
    mp.spawn(_worker_reducer, args=(2, _free_port(), flat), nprocs=2, join=True)
