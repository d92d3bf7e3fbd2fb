#!/usr/bin/env python
"""The reference workload end to end on this repository's apex drop-in.

Same flow, flags and call order as /root/reference/test_apex_distributed_spawn.py (spawn one
process per GPU -> NCCL -> amp.initialize(O2) -> apex DDP -> scale_loss/backward/step), with two
differences forced by the sandbox: MNIST cannot be downloaded, so the data are seeded synthetic
28x28 images, and ``--steps`` bounds the run.  The model is the same shape of network
(two conv/bn/relu/pool blocks and a linear head: 29,034 parameters in 10 tensors).

    python examples/convnet_ddp_spawn.py --gpus 1 --steps 200
"""
import argparse
import os
import random
import sys
import time

import torch
import torch.multiprocessing as mp
import torch.nn as nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import apex.parallel  # noqa: E402
from apex import amp  # noqa: E402


def build_net(num_classes=10):
    def block(cin, cout):
        return nn.Sequential(nn.Conv2d(cin, cout, kernel_size=5, stride=1, padding=2), nn.BatchNorm2d(cout), nn.ReLU(),
                             nn.MaxPool2d(kernel_size=2, stride=2))
    return nn.Sequential(block(1, 16), block(16, 32), nn.Flatten(), nn.Linear(7 * 7 * 32, num_classes))


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", default=1, type=int)
    ap.add_argument("--gpus", default=1, type=int)
    ap.add_argument("--nr", default=0, type=int)
    ap.add_argument("--apex_enabled", default=True, type=bool)       # same quirk as the reference: any non-empty string is True
    ap.add_argument("--apex_opt_level", default="O2", type=str)
    ap.add_argument("--epochs", default=1, type=int)
    ap.add_argument("--batch_size", default=100, type=int)
    ap.add_argument("--steps", default=100, type=int, help="iterations per epoch (synthetic data)")
    ap.add_argument("--optimizer", default="sgd", choices=["sgd", "fused_sgd", "fused_adam"])
    ap.add_argument("--port", default=8888, type=int)
    return ap.parse_args()


def seed_everything(seed):
    os.environ["PYTHONHASHSEED"] = str(seed)
    random.seed(seed)
    torch.manual_seed(seed)
    torch.cuda.manual_seed_all(seed)


def train(gpu, args):
    rank = args.nr * args.gpus + gpu
    torch.distributed.init_process_group(backend="nccl", init_method="env://", world_size=args.world_size, rank=rank)
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False
    seed_everything(0)
    torch.cuda.set_device(gpu)

    model = build_net().cuda(gpu)
    if args.optimizer == "sgd":
        optimizer = torch.optim.SGD(model.parameters(), 1e-4)
    elif args.optimizer == "fused_sgd":
        from apex.optimizers import FusedSGD
        optimizer = FusedSGD(model.parameters(), lr=1e-2, momentum=0.9, materialize_master_grads=False)
    else:
        from apex.optimizers import FusedAdam
        optimizer = FusedAdam(model.parameters(), lr=1e-3)
    if args.apex_enabled:
        model, optimizer = amp.initialize(model, optimizer, opt_level=args.apex_opt_level)
        model = apex.parallel.DistributedDataParallel(model)
    else:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[gpu])

    gen = torch.Generator().manual_seed(1234 + rank)
    criterion = nn.CrossEntropyLoss().cuda(gpu)
    # a learnable synthetic task: the label is a function of the image.  A pool of pinned host batches is
    # prepared once (this rank's shard of the "dataset"); the loop only does the reference's H2D copies.
    proto = torch.randn(10, 1, 28, 28, generator=torch.Generator().manual_seed(99))
    pool = []
    for _ in range(32):
        lab = torch.randint(0, 10, (args.batch_size,), generator=gen)
        img = proto[lab] + 0.5 * torch.randn(args.batch_size, 1, 28, 28, generator=gen)
        pool.append((img.pin_memory(), lab.pin_memory()))
    torch.cuda.synchronize()
    start = time.time()
    losses = []
    for epoch in range(args.epochs):
        for i in range(args.steps):
            images, labels = pool[i % len(pool)]
            images = images.cuda(non_blocking=True)
            labels = labels.cuda(non_blocking=True)

            outputs = model(images)
            loss = criterion(outputs, labels)

            optimizer.zero_grad()
            if args.apex_enabled:
                with amp.scale_loss(loss, optimizer) as scaled_loss:
                    scaled_loss.backward()
            else:
                loss.backward()
            optimizer.step()

            if (i + 1) % 50 == 0 and gpu == 0:
                losses.append(loss.item())
                print(f"Epoch [{epoch + 1}/{args.epochs}], Step [{i + 1}/{args.steps}], Loss: {losses[-1]:.4f}", flush=True)
    torch.cuda.synchronize()
    if gpu == 0:
        dt = time.time() - start
        n = args.epochs * args.steps * args.batch_size * args.world_size
        print(f"Training complete in: {dt:.2f}s  ({n / dt:.0f} samples/s over {args.world_size} GPU(s))", flush=True)
    checksum = float(sum(p.detach().double().sum() for p in model.parameters()))
    t = torch.tensor([checksum], dtype=torch.float64, device=f"cuda:{gpu}")
    lo, hi = t.clone(), t.clone()
    torch.distributed.all_reduce(lo, op=torch.distributed.ReduceOp.MIN)
    torch.distributed.all_reduce(hi, op=torch.distributed.ReduceOp.MAX)
    if gpu == 0:
        print(f"parameter checksum identical across ranks: {bool(lo.item() == hi.item())} ({checksum:.10f})", flush=True)
    torch.distributed.destroy_process_group()


def main():
    args = parse_args()
    args.world_size = args.gpus * args.nodes
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(args.port)
    mp.spawn(train, nprocs=args.gpus, args=(args,))


if __name__ == "__main__":
    main()
