"""-m gpu parity tests: every kernel of libapex_b200.so, called through the C ABI, against
oracle/amp_step_oracle.c on the same seeded inputs.  The bar is bit-exact (integer decisions
AND floating-point state): both sides do fp32 arithmetic in the same order without FMA."""
import pytest
import torch

import oracle
from apex_example_b200 import ops
from apex_example_b200.tables import TensorTable
from gpu_util import ADVERSARIAL_SIZES, assert_lists_bit_equal, bit_equal, make_list

pytestmark = pytest.mark.gpu


def _flag():
    return torch.zeros(1, dtype=torch.int32, device="cuda")


def _adam_case(sizes, gdt, sdt, cdt, *, steps=2, misalign=(), groups=None, hypers=None, inv_scale=1.0 / 1024,
               tile_elems=4096, seed=0):
    gen = torch.Generator().manual_seed(seed)
    hp, dp = make_list(sizes, sdt, gen, 0.05, misalign=misalign)
    hm, dm = make_list(sizes, sdt, gen, 1e-3, misalign=misalign)
    hv, dv = make_list(sizes, sdt, gen, 1e-3, misalign=misalign)
    for x, y in zip(hv, dv):
        x.abs_(); y.abs_()
    hc = dc = None
    if cdt is not None:
        hc = [torch.zeros(n, dtype=cdt) for n in sizes]
        dc = [torch.zeros(n, dtype=cdt, device="cuda") for n in sizes]
    hypers = hypers or [dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, bias_correction=True,
                             adam_w_mode=True)]
    groups = groups or [0] * len(sizes)
    for step in range(1, steps + 1):
        hg, dg = make_list(sizes, gdt, gen, 1e-3 / inv_scale, misalign=misalign)
        cols = [dg, dp, dm, dv] + ([dc] if dc else [])
        table = TensorTable(cols, groups=groups, tile_elems=tile_elems)
        grp = [ops.AdamGroup(h["lr"], h["beta1"], h["beta2"], h["eps"], h["weight_decay"], step,
                             int(h["bias_correction"]), int(h["adam_w_mode"])) for h in hypers]
        ops.fused_adam(table, gdt, sdt, cdt, grp, inv_scale=inv_scale)
        for i in range(len(sizes)):
            oracle.adam(hg[i], hp[i], hm[i], hv[i], hc[i] if hc else None, step=step, inv_scale=inv_scale,
                        **hypers[groups[i]])
        torch.cuda.synchronize()
        assert_lists_bit_equal(dp, hp, f"p(step {step})")
        assert_lists_bit_equal(dm, hm, f"m(step {step})")
        assert_lists_bit_equal(dv, hv, f"v(step {step})")
        if dc:
            assert_lists_bit_equal(dc, hc, f"copy(step {step})")


@pytest.mark.parametrize("gdt,sdt,cdt", [
    (torch.float16, torch.float32, torch.float16),       # O2
    (torch.bfloat16, torch.float32, torch.bfloat16),     # O2 bf16
    (torch.float32, torch.float32, None),                # O0 / O1 / fp32 BN params at O2
    (torch.float32, torch.float32, torch.float16),       # O2 with materialised fp32 master grads
    (torch.float32, torch.float32, torch.bfloat16),
    (torch.float16, torch.float16, None),                # O3
    (torch.bfloat16, torch.bfloat16, None),
    (torch.float16, torch.float32, None),
    (torch.bfloat16, torch.float32, None),
])
def test_fused_adam_matches_oracle_bit_exact(gdt, sdt, cdt):
    _adam_case(ADVERSARIAL_SIZES, gdt, sdt, cdt, misalign=(3, 10))


@pytest.mark.parametrize("threads,unroll,ctas", [(256, 1, 0), (256, 2, 0), (512, 1, 0), (512, 2, 0), (256, 2, 1), (256, 2, 16)])
def test_fused_adam_every_launch_config(threads, unroll, ctas):
    ops.set_launch_config(ctas, threads, unroll)
    try:
        _adam_case(ADVERSARIAL_SIZES + [300001], torch.float16, torch.float32, torch.float16, steps=1, misalign=(1,))
    finally:
        ops.set_launch_config(0, 256, 2)


@pytest.mark.parametrize("tile", [8, 2048, 65536])
def test_fused_adam_other_tile_sizes(tile):
    _adam_case([5, 70000, 4096, 131073], torch.float16, torch.float32, torch.float16, steps=1, tile_elems=tile)


def test_fused_adam_hyperparameter_variants_and_groups():
    hypers = [
        dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, bias_correction=True, adam_w_mode=True),
        dict(lr=3e-4, beta1=0.9, beta2=0.95, eps=1e-6, weight_decay=0.1, bias_correction=True, adam_w_mode=False),
        dict(lr=1e-2, beta1=0.3, beta2=0.999, eps=1e-8, weight_decay=0.01, bias_correction=False, adam_w_mode=True),
    ]
    sizes = [1000, 4097, 20000, 8, 33333, 64]
    _adam_case(sizes, torch.float16, torch.float32, torch.float16, steps=3, groups=[0, 1, 2, 2, 1, 0], hypers=hypers)


def test_fused_adam_skips_on_device_flag_and_reads_device_scale():
    gen = torch.Generator().manual_seed(5)
    sizes = [4096, 77, 10000]
    hp, dp = make_list(sizes, torch.float32, gen, 0.05)
    hm, dm = make_list(sizes, torch.float32, gen, 1e-3)
    hv, dv = make_list(sizes, torch.float32, gen, 1e-6)
    for x, y in zip(hv, dv):
        x.abs_(); y.abs_()
    hg, dg = make_list(sizes, torch.float16, gen, 1.0)
    dc = [torch.zeros(n, dtype=torch.float16, device="cuda") for n in sizes]
    hc = [torch.zeros(n, dtype=torch.float16) for n in sizes]
    table = TensorTable([dg, dp, dm, dv, dc])
    grp = [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.0, 1, 1, 1)]
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device="cuda")
    ops.scaler_init(state, 512.0)
    fstate = state.view(torch.float32)
    state[ops.IDX_FOUND_INF] = 1
    ops.fused_adam(table, torch.float16, torch.float32, torch.float16, grp,
                   dev_inv_scale=fstate[ops.IDX_INV_SCALE:], dev_found_inf=state[ops.IDX_FOUND_INF:])
    torch.cuda.synchronize()
    assert_lists_bit_equal(dp, hp, "p after skipped step")
    assert_lists_bit_equal(dm, hm, "m after skipped step")
    assert_lists_bit_equal(dv, hv, "v after skipped step")
    assert all(int(c.abs().sum()) == 0 for c in dc)
    state[ops.IDX_FOUND_INF] = 0
    ops.fused_adam(table, torch.float16, torch.float32, torch.float16, grp, inv_scale=123.0,   # ignored: device scale wins
                   dev_inv_scale=fstate[ops.IDX_INV_SCALE:], dev_found_inf=state[ops.IDX_FOUND_INF:])
    for i in range(len(sizes)):
        oracle.adam(hg[i], hp[i], hm[i], hv[i], hc[i], lr=1e-3, step=1, inv_scale=1.0 / 512)
    torch.cuda.synchronize()
    assert_lists_bit_equal(dp, hp, "p"); assert_lists_bit_equal(dc, hc, "copy")


SGD_VARIANTS = [
    dict(lr=0.1),
    dict(lr=0.1, momentum=0.9),
    dict(lr=0.1, momentum=0.9, nesterov=True, weight_decay=1e-4),
    dict(lr=0.05, momentum=0.9, dampening=0.1, weight_decay=1e-4),
    dict(lr=0.1, momentum=0.9, weight_decay=1e-4, wd_after_momentum=True),
]


@pytest.mark.parametrize("gdt,sdt,cdt", [
    (torch.float16, torch.float32, torch.float16), (torch.float32, torch.float32, None),
    (torch.float16, torch.float16, None), (torch.float32, torch.float32, torch.float16),
    (torch.bfloat16, torch.float32, torch.bfloat16), (torch.bfloat16, torch.bfloat16, None),
])
@pytest.mark.parametrize("variant", range(len(SGD_VARIANTS)))
def test_fused_sgd_matches_oracle_bit_exact(gdt, sdt, cdt, variant):
    kw = SGD_VARIANTS[variant]
    gen = torch.Generator().manual_seed(100 + variant)
    sizes = ADVERSARIAL_SIZES
    hp, dp = make_list(sizes, sdt, gen, 0.05, misalign=(2,))
    hb = [torch.zeros(n, dtype=sdt) for n in sizes]
    db = [torch.zeros(n, dtype=sdt, device="cuda") for n in sizes]
    has_mom = kw.get("momentum", 0.0) != 0.0
    hc = [torch.zeros(n, dtype=cdt) for n in sizes] if cdt else None
    dc = [torch.zeros(n, dtype=cdt, device="cuda") for n in sizes] if cdt else None
    inv = 1.0 / 256
    for step in range(3):
        hg, dg = make_list(sizes, gdt, gen, 1e-2 / inv, misalign=(2,))
        cols = [dg, dp, db if has_mom else [None] * len(sizes)] + ([dc] if dc else [])
        table = TensorTable(cols)
        grp = [ops.SgdGroup(kw["lr"], kw.get("momentum", 0.0), kw.get("dampening", 0.0), kw.get("weight_decay", 0.0),
                            step + 1, int(kw.get("nesterov", False)), int(kw.get("wd_after_momentum", False)))]
        ops.fused_sgd(table, gdt, sdt, cdt, grp, inv_scale=inv)
        for i in range(len(sizes)):
            oracle.sgd(hg[i], hp[i], hb[i] if has_mom else None, hc[i] if hc else None, first_run=(step == 0),
                       inv_scale=inv, **kw)
        torch.cuda.synchronize()
        assert_lists_bit_equal(dp, hp, f"w(step {step})")
        if has_mom:
            assert_lists_bit_equal(db, hb, f"mom(step {step})")
        if dc:
            assert_lists_bit_equal(dc, hc, f"copy(step {step})")


@pytest.mark.parametrize("idt", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("odt", [torch.float16, torch.bfloat16, torch.float32])
def test_multi_tensor_scale_matches_oracle(idt, odt):
    gen = torch.Generator().manual_seed(7)
    sizes = ADVERSARIAL_SIZES
    hi, di = make_list(sizes, idt, gen, 100.0, misalign=(4,))
    ho = [torch.zeros(n, dtype=odt) for n in sizes]
    do = [torch.zeros(n, dtype=odt, device="cuda") for n in sizes]
    flag = _flag()
    ops.multi_tensor_scale(TensorTable([di, do]), idt, odt, scale=1.0 / 128, dev_flag=flag)
    want = 0
    for a, b in zip(hi, ho):
        want |= oracle.scale(a, b, 1.0 / 128)
    assert int(flag.item()) == want == 0
    assert_lists_bit_equal(do, ho, "out")


@pytest.mark.parametrize("where", ["first_elem_first_tensor", "last_elem_last_tensor", "tail_elem", "misaligned_tensor", "nan"])
def test_overflow_flag_positions(where):
    gen = torch.Generator().manual_seed(8)
    sizes = [4099, 65537, 1, 7, 20001]
    hi, di = make_list(sizes, torch.float16, gen, 10.0, misalign=(3,))
    t, e, val = {"first_elem_first_tensor": (0, 0, float("inf")), "last_elem_last_tensor": (4, 20000, float("-inf")),
                 "tail_elem": (0, 4098, float("inf")), "misaligned_tensor": (3, 5, float("inf")),
                 "nan": (1, 40000, float("nan"))}[where]
    di[t][e] = val
    hi[t][e] = val
    outs = [torch.empty(n, dtype=torch.float32, device="cuda") for n in sizes]     # must outlive the launches
    table = TensorTable([di, outs])
    for fn in ("check", "scale"):
        flag = _flag()
        if fn == "check":
            ops.multi_tensor_check(table, 0, torch.float16, flag)
        else:
            ops.multi_tensor_scale(table, torch.float16, torch.float32, scale=0.5, dev_flag=flag)
        assert int(flag.item()) == 1 == max(oracle.check(h) for h in hi)
    clean = _flag()
    di[t][e] = 1.0
    ops.multi_tensor_check(table, 0, torch.float16, clean)
    assert int(clean.item()) == 0


def test_multi_tensor_axpby_matches_oracle():
    gen = torch.Generator().manual_seed(9)
    sizes = ADVERSARIAL_SIZES
    hx, dx = make_list(sizes, torch.float16, gen, 10.0)
    hy, dy = make_list(sizes, torch.float32, gen, 1.0)
    flag = _flag()
    ops.multi_tensor_axpby(TensorTable([dx, dy, dy]), torch.float16, torch.float32, torch.float32, 1.0 / 64, 1.0,
                           arg_to_check=0, dev_flag=flag)     # in-place accumulate into y, like apex's stashed-grad merge
    for a, b in zip(hx, hy):
        oracle.axpby(a, b, b, 1.0 / 64, 1.0, 0)
    assert int(flag.item()) == 0
    assert_lists_bit_equal(dy, hy, "y")
    dy[1][3] = float("nan")
    f0, f1 = _flag(), _flag()
    t = TensorTable([dx, dy, dy])
    ops.multi_tensor_axpby(t, torch.float16, torch.float32, torch.float32, 1.0, 1.0, arg_to_check=0, dev_flag=f0)
    ops.multi_tensor_axpby(t, torch.float16, torch.float32, torch.float32, 1.0, 1.0, arg_to_check=1, dev_flag=f1)
    assert (int(f0.item()), int(f1.item())) == (0, 1)


@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("n,off", [(0, 0), (1, 0), (5, 1), (4096, 0), (100003, 3), (1 << 22, 0), ((1 << 22) + 13, 1)])
def test_bucket_epilogue_matches_oracle(dt, n, off):
    gen = torch.Generator().manual_seed(n + off)
    h = (torch.randn(n, generator=gen) * 1000).to(dt)
    store = torch.empty(n + off, dtype=dt, device="cuda")
    d = store[off:]
    d.copy_(h)
    flag = _flag()
    ops.bucket_epilogue(d, 1.0 / 8, flag)
    want = oracle.bucket_epilogue(h, 1.0 / 8)
    assert int(flag.item()) == want == 0
    assert bit_equal(d, h)
    if n:
        before = d.clone()
        d[n - 1] = float("inf")
        ops.bucket_epilogue(d, 1.0, flag)            # factor 1: read-only check
        assert int(flag.item()) == 1
        d[n - 1] = before[n - 1]
        assert bit_equal(d, before)


def test_scaler_state_machine_matches_oracle_and_mirror():
    state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device="cuda")
    mirror = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32).pin_memory()
    ops.scaler_init(state, 65536.0)
    sc = oracle.LossScaler(scale_window=3)
    flags = [0, 0, 0, 0, 1, 0, 0, 1, 1, 0, 0, 0, 0, 0, 0]
    for i, f in enumerate(flags):
        if f:
            state[ops.IDX_OVERFLOW_ACC] = 1
        ops.scaler_update(state, True, 3, 2.0, 0.0, 2.0 ** 24, mirror)
        skip = sc.update_scale(bool(f))
        torch.cuda.synchronize()
        got = state.cpu()
        gf = got.view(torch.float32)
        assert float(gf[ops.IDX_LOSS_SCALE]) == sc.loss_scale and float(gf[ops.IDX_INV_SCALE]) == sc.inv_scale
        assert int(got[ops.IDX_FOUND_INF]) == int(skip) and int(got[ops.IDX_UNSKIPPED]) == sc.unskipped
        assert int(got[ops.IDX_OVERFLOW_ACC]) == 0 and int(got[ops.IDX_SEQ]) == i + 1
        assert torch.equal(mirror, got)
    # cap at 2**24, floor at min_loss_scale, static mode never skips
    ops.scaler_init(state, 2.0 ** 23)
    for _ in range(9):
        ops.scaler_update(state, True, 3, 2.0, 0.0, 2.0 ** 24)
    assert float(state.cpu().view(torch.float32)[0]) == 2.0 ** 24
    ops.scaler_init(state, 4.0)
    for _ in range(3):
        state[ops.IDX_OVERFLOW_ACC] = 1
        ops.scaler_update(state, True, 2000, 2.0, 2.0, 2.0 ** 24)
    assert float(state.cpu().view(torch.float32)[0]) == 2.0
    ops.scaler_init(state, 128.0)
    state[ops.IDX_OVERFLOW_ACC] = 1
    ops.scaler_update(state, False, 2000, 2.0, 0.0, 2.0 ** 24)
    got = state.cpu()
    assert float(got.view(torch.float32)[0]) == 128.0 and int(got[ops.IDX_FOUND_INF]) == 0 and int(got[ops.IDX_HAD_OVERFLOW]) == 1


def test_error_paths_raise():
    t = TensorTable([[torch.zeros(8, device="cuda")] for _ in range(4)])
    with pytest.raises(RuntimeError, match="not instantiated"):
        ops.fused_adam(t, torch.float16, torch.float16, torch.float32, [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0, 1, 1, 1)])
    with pytest.raises(RuntimeError, match="step"):
        ops.fused_adam(t, torch.float32, torch.float32, None, [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0, 0, 1, 1)])
    with pytest.raises(RuntimeError):
        ops.bucket_epilogue(torch.zeros(8), 1.0)
    assert ops.launch_count() > 0


def test_no_kernel_writes_outside_its_tensors_canaries():
    """compute-sanitizer is closed on this GPU pool, so out-of-bounds writes are hunted directly: every
    tensor sits between 64-element sentinel regions that must survive all kernels untouched."""
    gen = torch.Generator().manual_seed(77)
    sizes = [1, 7, 8, 9, 15, 4095, 4096, 4097, 65535, 70001]
    PAD = 64

    def guarded(n, dtype, fill):
        store = torch.full((n + 2 * PAD,), fill, dtype=dtype, device="cuda")
        return store, store[PAD:PAD + n]

    def intact(store, n, fill):
        return bool((store[:PAD] == fill).all()) and bool((store[PAD + n:] == fill).all())

    stores, cols = {k: [] for k in "gpmvc"}, {k: [] for k in "gpmvc"}
    fills = {"g": 3.0, "p": 5.0, "m": 7.0, "v": 9.0, "c": 11.0}
    dts = {"g": torch.float16, "p": torch.float32, "m": torch.float32, "v": torch.float32, "c": torch.float16}
    for n in sizes:
        for k in "gpmvc":
            s, view = guarded(n, dts[k], fills[k])
            view.copy_((torch.randn(n, generator=gen) * 0.01).abs().to(dts[k]))
            stores[k].append(s)
            cols[k].append(view)
    flag = _flag()
    table = TensorTable([cols[k] for k in "gpmvc"])
    ops.multi_tensor_check(TensorTable([cols["g"]]), 0, torch.float16, flag)
    ops.fused_adam(table, torch.float16, torch.float32, torch.float16, [ops.AdamGroup(1e-3, 0.9, 0.999, 1e-8, 0.01, 1, 1, 1)])
    ops.fused_sgd(TensorTable([cols["g"], cols["p"], cols["m"], cols["c"]]), torch.float16, torch.float32, torch.float16,
                  [ops.SgdGroup(0.1, 0.9, 0.0, 1e-4, 1, 1, 0)])
    ops.multi_tensor_scale(TensorTable([cols["g"], cols["v"]]), torch.float16, torch.float32, scale=0.5, dev_flag=flag)
    ops.multi_tensor_axpby(TensorTable([cols["g"], cols["m"], cols["m"]]), torch.float16, torch.float32, torch.float32,
                           1.0, 1.0, -1, flag)
    for view in cols["g"]:
        ops.bucket_epilogue(view, 0.5, flag)
    out = [guarded(n, torch.float16, 13.0) for n in sizes]
    for (s, view), g in zip(out, cols["g"]):
        ops.peer_reduce_shard([g.data_ptr(), g.data_ptr()], view, 0, g.numel(), 0.5, flag)
    torch.cuda.synchronize()
    for i, n in enumerate(sizes):
        for k in "gpmvc":
            assert intact(stores[k][i], n, fills[k]), f"column {k}, tensor {i} (numel {n}): sentinel overwritten"
        assert intact(out[i][0], n, 13.0), f"peer_reduce output {i}: sentinel overwritten"
