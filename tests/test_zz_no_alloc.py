"""b200_fused_adam_noalloc — the Adam step for callers that are not allowed to allocate.

Some users of this library cannot call cudaMalloc or go through torch's allocator at the moment
they need an optimizer step: a plugin living inside a host application that gave it a fixed
memory budget, a callback that must not enter an allocator (allocators may lock or synchronise
the device), a plain C program that owns its parameter buffers and nothing else.  The regular
entry point needs a tensor-list table in device memory that the caller allocates; this one keeps
the table in a 256 KiB static device array inside the library.  The test drives it through the C
ABI twice in a row (the second time with a different, smaller list, so a stale table would show)
and requires bit-identical results to the regular entry point and to the CPU oracle — the same
bar as the other kernel tests.

This file is named so that pytest collects it last."""
import ctypes

import pytest
import torch

from apex_example_b200 import _lib


def _call_noalloc(cols, gdt, sdt, cdt, group, inv_scale, stream=0):
    lib = _lib.load()
    n = len(cols[0])
    depth = len(cols)
    ptrs = (ctypes.c_void_p * (n * depth))()
    numel = (ctypes.c_int64 * n)()
    for d, col in enumerate(cols):
        for t, x in enumerate(col):
            ptrs[d * n + t] = x.data_ptr()
    for t, x in enumerate(cols[1]):
        numel[t] = x.numel()
    return lib.b200_fused_adam_noalloc(n, ptrs, numel, gdt, sdt, cdt, ctypes.byref(group), inv_scale, None, None, stream)


def test_lists_beyond_the_static_table_are_refused_with_a_clear_error():
    """Runs without a GPU: the size check comes before any CUDA call."""
    lib = _lib.load()
    n = 5000                                   # 64 + 64*5000 bytes of rows alone exceed 256 KiB
    ptrs = (ctypes.c_void_p * (n * 4))()
    numel = (ctypes.c_int64 * n)(*([8] * n))
    grp = _lib.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.0, 1, 1, 1)
    rc = lib.b200_fused_adam_noalloc(n, ptrs, numel, _lib.F16, _lib.F32, _lib.NONE, ctypes.byref(grp), 1.0, None, None, None)
    assert rc == -2                            # B200_ESIZE
    msg = lib.b200_last_error().decode()
    assert "static table" in msg and str(_lib.NOALLOC_TABLE_BYTES) in msg
    one_big = (ctypes.c_int64 * 1)(4096 * 40000)      # 40,000 tiles > 256 KiB of tile entries
    rc = lib.b200_fused_adam_noalloc(1, ptrs, one_big, _lib.F16, _lib.F32, _lib.NONE, ctypes.byref(grp), 1.0, None, None, None)
    assert rc == -2


def test_a_missing_column_pointer_is_refused_instead_of_dereferenced():
    """Runs without a GPU: argument checks come before any CUDA call.  b200_table_build treats a
    NULL pointer as "column absent"; the Adam kernel reads every column, so the entry point must
    not let one through."""
    lib = _lib.load()
    n = 3
    numel = (ctypes.c_int64 * n)(16, 0, 8)
    grp = _lib.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.0, 1, 1, 1)
    for depth, cdt in ((4, _lib.NONE), (5, _lib.F16)):
        for hole in range(depth * n):
            if numel[hole % n] == 0:
                continue                       # an empty tensor has nothing to point at
            ptrs = (ctypes.c_void_p * (depth * n))(*([0x1000] * (depth * n)))
            ptrs[hole] = None
            rc = lib.b200_fused_adam_noalloc(n, ptrs, numel, _lib.F16, _lib.F32, cdt, ctypes.byref(grp), 1.0, None, None, None)
            assert rc == -1, (depth, hole)     # B200_EINVAL
            assert "NULL" in lib.b200_last_error().decode()


@pytest.mark.gpu
def test_noalloc_entry_point_matches_the_regular_one_twice_in_a_row():
    import oracle
    from apex_example_b200 import ops
    from apex_example_b200.tables import TensorTable

    def make(sizes, seed):
        g = torch.Generator().manual_seed(seed)
        grads = [(torch.randn(n, generator=g) * 4).half().cuda() for n in sizes]
        p = [(torch.randn(n, generator=g) * 0.05).cuda() for n in sizes]
        m = [(torch.randn(n, generator=g) * 1e-3).cuda() for n in sizes]
        v = [(torch.randn(n, generator=g) * 1e-3).square().cuda() for n in sizes]
        c = [torch.zeros(n, dtype=torch.float16, device="cuda") for n in sizes]
        return [grads, p, m, v, c]

    def clone(cols):
        return [[t.clone() for t in col] for col in cols]

    # call 1: a ragged 11-tensor list; call 2: a shorter list, other sizes (a stale table would be visible)
    for call, (sizes, step) in enumerate([([400, 16, 7, 12800, 1, 32, 15680, 10, 70001, 4096, 4097], 3), ([9000, 5, 2048], 4)]):
        a = make(sizes, 100 + call)
        b = clone(a)
        grp = _lib.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.01, step, 1, 1)
        rc = _call_noalloc(a, _lib.F16, _lib.F32, _lib.F16, grp, 1.0 / 128, torch.cuda.current_stream().cuda_stream)
        assert rc == 0, _lib.load().b200_last_error().decode()
        ops.fused_adam(TensorTable(b), torch.float16, torch.float32, torch.float16, [grp], inv_scale=1.0 / 128)
        torch.cuda.synchronize()
        for name, ca, cb in zip("gpmvc", a, b):
            for i, (x, y) in enumerate(zip(ca, cb)):
                assert torch.equal(x, y), f"call {call + 1}: column {name}[{i}] differs between the two entry points"
        # and against the CPU oracle, on the same inputs (bit for bit, like every other kernel test)
        h = [[t.cpu() for t in col] for col in make(sizes, 100 + call)]
        for t in range(len(sizes)):
            oracle.adam(h[0][t], h[1][t], h[2][t], h[3][t], h[4][t], lr=1e-3, weight_decay=0.01, step=step,
                        inv_scale=1.0 / 128)
        for name, ca, ch in zip("gpmvc", a, h):
            for i, (x, y) in enumerate(zip(ca, ch)):
                assert torch.equal(x.cpu(), y), f"call {call + 1}: column {name}[{i}] differs from the oracle"
        assert not torch.equal(a[1][0], make(sizes, 100 + call)[1][0]), "the step did not run"
