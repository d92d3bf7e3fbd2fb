"""CPU checks of the C-ABI boundary: the library loads, exports every symbol the header
declares, and the host half of the chunk scheduler (table builder) covers every element of any
tensor list exactly once.  No compute calls — there is no GPU here."""
import ctypes
import os
import re

import pytest
import torch
from hypothesis import given, settings, strategies as st

from apex_example_b200 import _lib, tables

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    inc = os.path.join(ROOT, "include")
    for fn in os.listdir(inc):
        if fn.endswith(".h"):
            src = open(os.path.join(inc, fn)).read()
            src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
            names |= set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src))
    return names


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    declared = _declared_symbols()
    assert len(declared) >= 15
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/ but not exported"
    assert declared == set(_lib.SIGNATURES), "ctypes signatures out of sync with the header"
    assert lib.b200_abi_version() == 1


def test_struct_sizes_match_header():
    assert ctypes.sizeof(_lib.TableHeader) == 64
    assert ctypes.sizeof(_lib.TableRow) == 64
    assert ctypes.sizeof(_lib.TableTile) == 8
    assert ctypes.sizeof(_lib.AdamGroup) == 56
    assert ctypes.sizeof(_lib.SgdGroup) == 48
    assert ctypes.sizeof(_lib.StepSync) == 24


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "apex_example_b200")
    shim = os.path.join(ROOT, "apex")
    for base in (pkg, shim):
        for dirpath, _, files in os.walk(base):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                    text = open(os.path.join(dirpath, f)).read()
                    assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                    assert "liboracle" not in text, f


def _parse(image: torch.Tensor):
    buf = bytes(image.numpy())
    h = _lib.TableHeader.from_buffer_copy(buf[:64])
    rows = [_lib.TableRow.from_buffer_copy(buf[h.rows_off + 64 * i: h.rows_off + 64 * (i + 1)]) for i in range(h.n_tensors)]
    tiles = [_lib.TableTile.from_buffer_copy(buf[h.tiles_off + 8 * i: h.tiles_off + 8 * (i + 1)]) for i in range(h.n_tiles)]
    return h, rows, tiles


@settings(max_examples=60, deadline=None)
@given(sizes=st.lists(st.integers(min_value=0, max_value=70000), min_size=0, max_size=12),
       tile=st.sampled_from([8, 64, 2048, 4096, 65536]))
def test_table_covers_every_element_once(sizes, tile):
    cols = [[torch.empty(n, dtype=torch.float16) for n in sizes], [torch.empty(n, dtype=torch.float32) for n in sizes]]
    if not sizes:           # an empty list is legal: header only, and the launchers turn into no-ops
        image, n_tiles = tables.build_image(cols, tile_elems=tile)
        assert n_tiles == 0 and image.numel() == 64
        return
    image, n_tiles = tables.build_image(cols, tile_elems=tile)
    h, rows, tiles = _parse(image)
    assert h.magic == 0x42325442 and h.n_tensors == len(sizes) and h.depth == 2 and h.tile_elems == tile
    assert h.n_tiles == n_tiles == sum((n + tile - 1) // tile for n in sizes)
    assert h.total_elems == sum(sizes) and h.image_bytes == image.numel()
    covered = [0] * len(sizes)
    seen = set()
    for tl in tiles:
        assert (tl.tensor, tl.tile) not in seen
        seen.add((tl.tensor, tl.tile))
        start = tl.tile * tile
        assert start < sizes[tl.tensor]
        covered[tl.tensor] += min(tile, sizes[tl.tensor] - start)
    assert covered == sizes
    for t, r in enumerate(rows):
        assert r.numel == sizes[t]
        assert (r.ptr[0] or 0) == cols[0][t].data_ptr() and (r.ptr[1] or 0) == cols[1][t].data_ptr()
        assert r.ptr[2] is None


def test_alignment_flag_and_groups():
    base16 = torch.empty(64, dtype=torch.float16)
    base32 = torch.empty(64, dtype=torch.float32)
    a16, a32 = base16[:32], base32[:32]              # torch CPU allocations are 64-byte aligned
    off16, off32 = base16[1:33], base32[4:36]        # +2 bytes / +16 bytes
    assert a16.data_ptr() % 16 == 0 and a32.data_ptr() % 32 == 0
    image, _ = tables.build_image([[a16, off16, a16, a16], [a32, a32, off32, a32]], groups=[0, 1, 2, 7])
    _, rows, _ = _parse(image)
    assert [r.flags & 1 for r in rows] == [1, 0, 0, 1]   # fp32 needs 32 B (256-bit access), 16-bit types 16 B
    assert [r.group for r in rows] == [0, 1, 2, 7]
    with pytest.raises(RuntimeError):
        tables.build_image([[a16]], groups=[8])


def test_builder_rejects_bad_input():
    a = torch.empty(10, dtype=torch.float32)
    with pytest.raises(RuntimeError):
        tables.build_image([[a], [torch.empty(11)]])
    with pytest.raises(RuntimeError):
        tables.build_image([[torch.empty(4, 8)[:, ::2]]])          # holes in memory: not a dense tensor
    with pytest.raises(RuntimeError):
        tables.build_image([[torch.empty(4, 4).t()], [torch.empty(4, 4)]])   # dense, but the columns disagree on the layout
    img, tiles = tables.build_image([[torch.empty(4, 4).t()], [torch.empty(4, 4).t()]])  # same dense layout: memory order
    assert tiles == 1
    with pytest.raises(RuntimeError):
        tables.build_image([[a]], tile_elems=100)
    with pytest.raises(RuntimeError):
        tables.build_image([[a]] * 6)
    lib = _lib.load()
    assert lib.b200_table_bytes(1, (ctypes.c_int64 * 1)(-5), 4096) < 0
    assert b"negative" in lib.b200_last_error()


def test_absent_column_and_gpt2_small_tile_count():
    sizes = [768] * 74 + [768 * 2304] * 12 + [2304] * 12 + [768 * 768] * 12 + [768 * 3072] * 12 + [3072] * 12 + \
            [3072 * 768] * 12 + [50257 * 768, 1024 * 768]
    assert len(sizes) == 148 and sum(sizes) == 124439808           # SURVEY.md App. C shape table
    numel = (ctypes.c_int64 * len(sizes))(*sizes)
    nbytes = _lib.load().b200_table_bytes(len(sizes), numel, 4096)
    n_tiles = (nbytes - 64 - 64 * len(sizes)) // 8
    assert n_tiles == sum((n + 4095) // 4096 for n in sizes)
    g = torch.empty(100, dtype=torch.float16)
    image, _ = tables.build_image([[g], [None]])
    _, rows, _ = _parse(image)
    assert rows[0].ptr[1] is None and rows[0].flags & 1


def test_kernel_entry_points_validate_arguments_without_a_gpu():
    lib = _lib.load()
    grp = (_lib.AdamGroup * 1)(_lib.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.0, 0, 1, 1))
    assert lib.b200_fused_adam(None, 1, _lib.F16, _lib.F32, _lib.F16, grp, 1, 1.0, None, None, None, None) == -1   # step 0
    assert lib.b200_fused_adam(None, 1, _lib.F16, _lib.F32, _lib.F16, grp, 9, 1.0, None, None, None, None) == -1   # groups
    assert lib.b200_multi_tensor_scale(None, 1, 7, _lib.F32, 1.0, None, None, None) == -3
    assert lib.b200_bucket_epilogue(None, -1, _lib.F16, 1.0, None, None) == -1
    assert lib.b200_set_launch_config(0, 100, 2) == -1
    # round-2 entry points: everything that can be refused is refused before the first CUDA call
    ok = (_lib.AdamGroup * 1)(_lib.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.0, 1, 1, 1))
    dummy = ctypes.c_void_p(256)
    assert lib.b200_fused_adam_spec(None, 1, _lib.F16, _lib.F32, _lib.F16, ok, 1, 1.0, None, None, None, None, None, None) == -1  # no parity word
    assert lib.b200_peer_shard_step(None, 1, _lib.F16, _lib.F32, _lib.F16, ok, 1, 1.0, None, None, None, None, None, None, None) == -1
    assert lib.b200_peer_shard_step(None, 1, _lib.F16, _lib.F32, _lib.F16, ok, 1, 1.0, None, None, dummy, None, None, None, None) == -1  # no fan
    fan = _lib.PeerFan()
    fan.world, fan.rank = 9, 0
    assert lib.b200_peer_shard_step(None, 1, _lib.F16, _lib.F32, _lib.F16, ok, 1, 1.0, None, None, dummy, None, ctypes.byref(fan), None, None) == -1
    assert "world 9" in lib.b200_last_error().decode()
    fan.world = 2
    assert lib.b200_peer_shard_repair(None, 1, _lib.F32, _lib.F16, dummy, dummy, ctypes.byref(fan), None, None) == -1   # NULL buffers
    assert lib.b200_l2norm_per_tensor(None, 1, 0, 7, None, 1, None, None, None, None, None, None) == -3                 # dtype
    assert lib.b200_l2norm_per_tensor(None, 1, 9, _lib.F32, None, 1, None, None, None, None, None, None) == -1         # column
    lg = (_lib.LambGroup * 1)(_lib.LambGroup(1e-3, 0.9, 0.999, 1e-6, 0.01, 1, 1, 1, 1, 0))
    assert lib.b200_lamb_stage1(None, 1, _lib.F16, _lib.F32, lg, 1, 1.0, None, None, None, 1.0, None) == -1             # no gradient norm
    assert lib.b200_lamb_stage2(None, 1, _lib.F32, _lib.F16, lg, 1, None, None, None, None) == -1                        # no norm arrays
    ng = (_lib.NovoGradGroup * 1)(_lib.NovoGradGroup(1e-3, 0.9, 0.999, 1e-8, 0.0, 0, 1, 0, 1, 0))
    assert lib.b200_fused_novograd(None, 1, _lib.F16, _lib.F32, _lib.F16, ng, 1, 1.0, None, None, dummy, None) == -1     # step 0
    assert lib.b200_fused_novograd(None, 1, _lib.F16, _lib.F32, _lib.F16, ng, 1, 1.0, None, None, None, None) == -1      # no moments
    with pytest.raises(RuntimeError):
        _lib.check(lib.b200_set_launch_config(0, 256, 3), "set_launch_config")
    assert lib.b200_set_launch_config(0, 256, 2) == 0


def test_header_is_plain_c_and_the_library_links_from_c(tmp_path):
    """The boundary is a C ABI: the header must compile as C99 and a C program must link against the .so."""
    import shutil
    import subprocess
    cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else shutil.which("gcc")
    if cc is None:
        pytest.skip("no C compiler")
    src = tmp_path / "abi.c"
    src.write_text(
        '#include <stdio.h>\n#include "apex_b200.h"\n'
        "int main(void) {\n"
        "    int64_t numel[2] = {5000, 17};\n"
        "    int64_t bytes = b200_table_bytes(2, numel, B200_TILE_ELEMS_DEFAULT);\n"
        "    if (sizeof(b200_scaler_state) != 48 || sizeof(b200_peer_signal_pad) != 128 || sizeof(b200_table_row) != 64) return 2;\n"
        "    return (b200_abi_version() == B200_ABI_VERSION && bytes == 64 + 2 * 64 + 3 * 8) ? 0 : 1;\n"
        "}\n")
    exe = tmp_path / "abi"
    libdir = os.path.dirname(_lib.SO_PATH)
    subprocess.run([cc, "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe),
                    "-L", libdir, "-l:" + os.path.basename(_lib.SO_PATH), "-Wl,-rpath," + libdir], check=True)
    assert subprocess.run([str(exe)]).returncode == 0


def test_ctypes_structs_match_the_header_field_by_field(tmp_path):
    """Every struct that crosses the C ABI by value or pointer: size and the offset of every field as gcc lays the header
    out, against the ctypes mirror in apex_example_b200/_lib.py (guards against a field added on one side only)."""
    import subprocess
    pairs = {"b200_adam_group": _lib.AdamGroup, "b200_sgd_group": _lib.SgdGroup, "b200_lamb_group": _lib.LambGroup,
             "b200_novograd_group": _lib.NovoGradGroup, "b200_step_sync": _lib.StepSync, "b200_peer_sync": _lib.PeerSync,
             "b200_peer_pull": _lib.PeerPull, "b200_peer_fan": _lib.PeerFan, "b200_table_header": _lib.TableHeader}
    header_names = {"b200_step_sync": {"dev_counters": "dev_counters"}}
    lines = ["#include <stdio.h>", "#include <stddef.h>", '#include "apex_b200.h"', "int main(void) {"]
    for cname, cls in pairs.items():
        lines.append(f'    printf("{cname} size %zu\\n", sizeof({cname}));')
        for fname, _ in cls._fields_:
            hname = header_names.get(cname, {}).get(fname, fname)
            lines.append(f'    printf("{cname} {fname} %zu\\n", offsetof({cname}, {hname}));')
    lines += ["    return 0;", "}"]
    src = tmp_path / "probe.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "probe"
    inc = os.path.join(ROOT, "include")
    subprocess.run(["gcc", "-std=c99", "-I", inc, str(src), "-o", str(exe)], check=True, capture_output=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    got = {}
    for line in out.splitlines():
        cname, fname, val = line.split()
        got[(cname, fname)] = int(val)
    for cname, cls in pairs.items():
        assert got[(cname, "size")] == ctypes.sizeof(cls), cname
        for fname, _ in cls._fields_:
            assert got[(cname, fname)] == getattr(cls, fname).offset, f"{cname}.{fname}"
