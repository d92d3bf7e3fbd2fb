"""Helpers shared by the -m gpu parity tests: build seeded tensor lists on the device, mirror
them on the host for the oracle, and compare bit for bit (NaNs compare equal to NaNs)."""
import torch

INT_VIEW = {2: torch.int16, 4: torch.int32}


def bit_equal(a: torch.Tensor, b: torch.Tensor) -> bool:
    a, b = a.detach().cpu().contiguous(), b.detach().cpu().contiguous()
    if a.dtype != b.dtype or a.shape != b.shape:
        return False
    if a.numel() == 0:
        return True
    ai, bi = a.view(INT_VIEW[a.element_size()]), b.view(INT_VIEW[b.element_size()])
    return bool(((ai == bi) | (a.isnan() & b.isnan())).all())


def assert_lists_bit_equal(dev_list, host_list, what=""):
    for i, (d, h) in enumerate(zip(dev_list, host_list)):
        if not bit_equal(d, h):
            dc = d.detach().cpu().float().flatten()
            hc = h.float().flatten()
            bad = (dc != hc) & ~(dc.isnan() & hc.isnan())
            idx = int(bad.nonzero()[0])
            raise AssertionError(f"{what}[{i}] differs from the oracle at {int(bad.sum())}/{dc.numel()} elements; "
                                 f"first at {idx}: gpu={dc[idx].item()!r} oracle={hc[idx].item()!r}")


def make_list(sizes, dtype, gen, scale=1.0, device="cuda", misalign=()):
    """Seeded host list + device copies.  Tensors whose index is in `misalign` are views that
    start one element into their storage (so they are NOT 16/32-byte aligned)."""
    host, dev = [], []
    for i, n in enumerate(sizes):
        h = (torch.randn(n, generator=gen) * scale).to(dtype)
        if i in misalign:
            store = torch.empty(n + 1, dtype=dtype, device=device)
            d = store[1:]
            d.copy_(h)
        else:
            d = h.to(device)
        host.append(h)
        dev.append(d)
    return host, dev


ADVERSARIAL_SIZES = [1, 7, 8, 9, 4095, 4096, 4097, 65535, 65537, 0, 12345, 2304]
