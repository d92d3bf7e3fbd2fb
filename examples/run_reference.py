#!/usr/bin/env python
"""Run the reference file ITSELF (enijkamp/apex_example, test_apex_distributed_spawn.py) through this
repository's drop-in `apex` package — no line of it is changed or copied.

    python examples/run_reference.py --ref /root/reference/test_apex_distributed_spawn.py --gpus 1 --epochs 1

What is substituted, and why:
  * `torchvision.datasets.MNIST` -> a synthetic stand-in (60,000 seeded 1x28x28 images whose label is a fixed
    function of the image, so the loss can fall): the real set needs a download and there is no network.
  * the launcher (`spawn_workers`, reference :53-57): its `mp.spawn` children would re-import the module without the
    stand-in, so this script spawns the ranks itself and calls the reference's own `train(gpu, args)` (:106-170) in
    each of them.  Everything from `init_process_group` to the last `optimizer.step()` is the reference's code.
The reference lives outside this repository (it is not copied in), so the GPU test that uses this script skips itself
where the file is absent — e.g. on a GPU box that only receives this repository."""
import argparse
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402


class SyntheticMNIST(torch.utils.data.Dataset):
    """Signature of torchvision.datasets.MNIST as the reference calls it (:127-132)."""

    def __init__(self, root=None, train=True, transform=None, download=False, size=60000, **_):
        g = torch.Generator().manual_seed(1234)
        self.images = torch.rand(size, 1, 28, 28, generator=g)
        proj = torch.randn(28 * 28, 10, generator=g)
        self.labels = (self.images.view(size, -1) @ proj).argmax(1)

    def __len__(self):
        return self.images.shape[0]

    def __getitem__(self, i):
        return self.images[i], int(self.labels[i])


def load_reference(path):
    import torchvision
    torchvision.datasets.MNIST = SyntheticMNIST
    spec = importlib.util.spec_from_file_location("reference_script", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)                     # `import apex.parallel` / `from apex import amp` resolve to this repo
    return mod


def _rank(gpu, path, ns):
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    mod = load_reference(path)
    import apex
    assert os.path.realpath(os.path.dirname(os.path.dirname(apex.__file__))) == os.path.realpath(ROOT), "not this repository's apex"
    mod.train(gpu, ns)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default=os.environ.get("APEX_EXAMPLE_REF", "/root/reference/test_apex_distributed_spawn.py"))
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--epochs", type=int, default=1)
    ap.add_argument("--batch_size", type=int, default=100)
    ap.add_argument("--apex_opt_level", default="O2")
    ap.add_argument("--port", default="8888")
    a = ap.parse_args()
    if not os.path.exists(a.ref):
        raise SystemExit(f"reference script not found at {a.ref} (pass --ref or set APEX_EXAMPLE_REF)")
    # the namespace parse_args() + spawn_workers() of the reference would have produced (:39-56)
    ns = argparse.Namespace(nodes=1, gpus=a.gpus, nr=0, apex_enabled=True, apex_opt_level=a.apex_opt_level,
                            epochs=a.epochs, batch_size=a.batch_size, world_size=a.gpus)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = a.port
    if a.gpus == 1:
        _rank(0, a.ref, ns)
    else:
        mp.spawn(_rank, nprocs=a.gpus, args=(a.ref, ns))


if __name__ == "__main__":
    main()
