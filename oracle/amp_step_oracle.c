/*
 * amp_step_oracle.c — CPU restatement of the mixed-precision step hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under apex_example_b200/ may import, link or call this
 * file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs do, and there only as the checker or the timed CPU baseline.
 *
 * What it restates.  The reference script (/root/reference/test_apex_distributed_spawn.py)
 * reaches this arithmetic through NVIDIA/apex pinned at f3a960f80244cf9e80558ab30f7f7e8cbf03c0a0
 * (/root/reference/README.md:16-19).  Apex is NOT vendored and NOT installed here, so the
 * algorithm is restated from (1) the numerical contract BASELINE.json's north_star names —
 * torch.optim.Adam / torch.optim.SGD, whose source IS on disk — and (2) the published apex
 * semantics summarised in SURVEY.md Appendix B.
 *
 * PARITY STATUS: unpinned against apex itself (the reference holds no golden vectors, tests
 * or fixtures for this path and apex cannot be run here).  Pinned against torch 2.11's own
 * CPU ops: tests/test_oracle.py checks every function below against torch.optim.Adam/AdamW/SGD
 * (foreach=False), torch._amp_foreach_non_finite_check_and_unscale_, torch._amp_update_scale_
 * and Tensor.copy_ casts, and against the committed fixtures in tests/golden/ that
 * tests/golden/make_golden.py generated from those torch ops.
 *
 * Arithmetic rules (the CUDA kernels follow the same ones, which is why they match this file
 * bit for bit): fp32 scalar math, IEEE sqrtf / division, NO fused multiply-add (compile with
 * -ffp-contract=off), hyper-parameters derived in double and rounded to float once.
 *
 * TORCH = site-packages/torch (2.11.0).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define DT_F32  0
#define DT_F16  1
#define DT_BF16 2

/* ---- element conversions (software; round-to-nearest-even like Tensor.copy_) ------------ */
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

static inline float half_to_float(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu, man = h & 0x3ffu;
    if (exp == 0) {
        if (man == 0) return u2f(sign);
        /* subnormal: value = man * 2^-24 */
        float v = (float)man * 5.9604644775390625e-08f;
        return sign ? -v : v;
    }
    if (exp == 31) return u2f(sign | 0x7f800000u | (man << 13));
    return u2f(sign | ((exp + 112u) << 23) | (man << 13));
}

static inline uint16_t float_to_half(float f) {
    uint32_t x = f2u(f);
    uint16_t sign = (uint16_t)((x >> 16) & 0x8000u);
    uint32_t ax = x & 0x7fffffffu;
    if (ax >= 0x7f800000u)                       /* inf / NaN */
        return (uint16_t)(sign | 0x7c00u | (ax > 0x7f800000u ? 0x200u : 0u));
    if (ax >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u);   /* >= 65520 rounds to inf */
    if (ax < 0x38800000u) {                      /* below the smallest normal half (2^-14) */
        if (ax < 0x33000000u) return sign;       /* < 2^-25 rounds to zero */
        uint32_t exp = ax >> 23;
        uint32_t man = (ax & 0x7fffffu) | 0x800000u;
        uint32_t shift = 126u - exp;             /* 14..24 */
        uint32_t q = man >> shift;
        uint32_t rem = man & ((1u << shift) - 1u), halfway = 1u << (shift - 1u);
        if (rem > halfway || (rem == halfway && (q & 1u))) q++;
        return (uint16_t)(sign | q);
    }
    uint32_t q = (ax - 0x38000000u) >> 13;       /* rebias 127 -> 15 */
    uint32_t rem = ax & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (q & 1u))) q++;
    return (uint16_t)(sign | q);
}

static inline float bf16_to_float(uint16_t b) { return u2f((uint32_t)b << 16); }

static inline uint16_t float_to_bf16(float f) {
    uint32_t x = f2u(f);
    if ((x & 0x7fffffffu) > 0x7f800000u) return 0x7fffu;        /* NaN (cuda's canonical NaN) */
    uint32_t lsb = (x >> 16) & 1u;
    x += 0x7fffu + lsb;
    return (uint16_t)(x >> 16);
}

static inline float load_as_float(const void* p, int dt, int64_t i) {
    switch (dt) {
    case DT_F16:  return half_to_float(((const uint16_t*)p)[i]);
    case DT_BF16: return bf16_to_float(((const uint16_t*)p)[i]);
    default:      return ((const float*)p)[i];
    }
}

static inline void store_from_float(void* p, int dt, int64_t i, float v) {
    switch (dt) {
    case DT_F16:  ((uint16_t*)p)[i] = float_to_half(v); break;
    case DT_BF16: ((uint16_t*)p)[i] = float_to_bf16(v); break;
    default:      ((float*)p)[i] = v; break;
    }
}

static inline int finite_f(float v) { return (f2u(v) & 0x7f800000u) != 0x7f800000u; }

/* exported so the tests can pin the conversions against Tensor.copy_ / Tensor.float() */
void oracle_convert(const void* in, int in_dt, void* out, int out_dt, int64_t n) {
    for (int64_t i = 0; i < n; i++) store_from_float(out, out_dt, i, load_as_float(in, in_dt, i));
}

/* ---- overflow check --------------------------------------------------------------------- */
/* apex LossScaler.unscale raises the flag if any INPUT element is inf/NaN;
 * same rule as torch._amp_foreach_non_finite_check_and_unscale_ (TORCH/amp/grad_scaler.py:292-298). */
int oracle_check(const void* g, int dt, int64_t n) {
    int flag = 0;
    if (dt == DT_F16 || dt == DT_BF16) {            /* same predicate on the raw bits (exponent all ones) */
        const uint16_t* h = (const uint16_t*)g;
        const uint16_t mask = dt == DT_F16 ? 0x7c00u : 0x7f80u;
        #pragma omp parallel for reduction(|:flag) schedule(static)
        for (int64_t i = 0; i < n; i++) flag |= (h[i] & mask) == mask;
        return flag;
    }
    #pragma omp parallel for reduction(|:flag) schedule(static)
    for (int64_t i = 0; i < n; i++) flag |= !finite_f(load_as_float(g, dt, i));
    return flag;
}

/* amp_C.multi_tensor_scale: out = in * scale, flag on non-finite input. */
void oracle_scale(const void* in, int in_dt, void* out, int out_dt, int64_t n, float scale, int* flag) {
    int f = 0;
    #pragma omp parallel for reduction(|:f) schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float x = load_as_float(in, in_dt, i);
        f |= !finite_f(x);
        store_from_float(out, out_dt, i, x * scale);
    }
    if (f && flag) *flag = 1;
}

/* amp_C.multi_tensor_axpby: out = a*x + b*y. arg_to_check: 0 x, 1 y, -1 both. */
void oracle_axpby(const void* x, int x_dt, const void* y, int y_dt, void* out, int out_dt,
                  int64_t n, float a, float b, int arg_to_check, int* flag) {
    int f = 0;
    #pragma omp parallel for reduction(|:f) schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float xv = load_as_float(x, x_dt, i), yv = load_as_float(y, y_dt, i);
        if (arg_to_check == 0) f |= !finite_f(xv);
        else if (arg_to_check == 1) f |= !finite_f(yv);
        else f |= (!finite_f(xv)) | (!finite_f(yv));
        float ax = a * xv, by = b * yv;
        store_from_float(out, out_dt, i, ax + by);
    }
    if (f && flag) *flag = 1;
}

/* ---- fused unscale + Adam + master->model copy ------------------------------------------- */
typedef struct {
    double  lr, beta1, beta2, eps, weight_decay;
    int64_t step;
    int32_t bias_correction;
    int32_t adam_w_mode;
} oracle_adam_group;

/* Follows torch.optim.Adam's single-tensor path op for op:
 *   TORCH/optim/adam.py:416-419  AdamW: param.mul_(1 - lr*wd)
 *   TORCH/optim/adam.py:429      L2:    grad = grad.add(param, alpha=wd)
 *   TORCH/optim/adam.py:457      exp_avg.lerp_(grad, 1-beta1)
 *   TORCH/optim/adam.py:476      exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2)
 *   TORCH/optim/adam.py:531-547  bias corrections in double; denom = sqrt(v)/sqrt(bc2)+eps;
 *                                param.addcdiv_(exp_avg, denom, value=-lr/bc1)
 * with apex FusedAdam's surroundings: grad is first multiplied by
 * inv_scale (the unscale of LossScaler.unscale), and the fp16/bf16 model copy is written from
 * the new fp32 master (apex _master_params_to_model_params).  State may be fp16/bf16 (O3):
 * values are widened to float, updated, and rounded back. */
#if defined(__x86_64__) && defined(__GNUC__)
/* Same arithmetic as the generic loop below for the O2 layout (fp16 grad, fp32 state, fp16 copy),
 * written so that gcc can vectorise it: hardware fp16 conversions (F16C, round-to-nearest-even),
 * IEEE vsqrtps / vdivps, and no FMA contraction (-ffp-contract=off).  Used only so that the timed
 * CPU baseline is not handicapped by the software conversions; tests/test_oracle.py checks that it
 * gives bit-identical results to the generic path. */
__attribute__((target("avx2,f16c"), optimize("O3")))
static void adam_o2_f16c(const uint16_t* g16, float* p, float* m, float* v, uint16_t* c16, int64_t n,
                         float inv_scale, float w1, float b2, float w2, float neg_step, float bc2_sqrt, float eps,
                         float wd, float decay, int has_wd, int adamw) {
    const _Float16* g = (const _Float16*)g16;
    _Float16* c = (_Float16*)c16;
    const float om_w1 = 1.0f - w1;
    const int lerp_hi = !(w1 < 0.5f);
    #pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float gi = (float)g[i] * inv_scale;
        float pi = p[i], mi = m[i], vi = v[i];
        if (has_wd) {
            if (adamw) pi = pi * decay;
            else gi = gi + wd * pi;
        }
        float d = gi - mi;
        mi = lerp_hi ? gi - d * om_w1 : mi + w1 * d;
        vi = vi * b2 + (w2 * gi) * gi;
        float denom = sqrtf(vi) / bc2_sqrt + eps;
        pi = pi + neg_step * (mi / denom);
        p[i] = pi; m[i] = mi; v[i] = vi;
        if (c) c[i] = (_Float16)pi;
    }
}
static int have_f16c(void) { return __builtin_cpu_supports("avx2") && __builtin_cpu_supports("f16c"); }
#else
static int have_f16c(void) { return 0; }
#endif

int oracle_force_generic = 0;       /* tests flip this to compare the two paths */

void oracle_adam(const void* g, int g_dt, void* p, void* m, void* v, int s_dt,
                 void* copy, int c_dt, int64_t n,
                 const oracle_adam_group* grp, float inv_scale) {
    const float w1 = (float)(1.0 - grp->beta1);
    const float b2 = (float)grp->beta2;
    const float w2 = (float)(1.0 - grp->beta2);
    const double bc1 = grp->bias_correction ? 1.0 - pow(grp->beta1, (double)grp->step) : 1.0;
    const double bc2 = grp->bias_correction ? 1.0 - pow(grp->beta2, (double)grp->step) : 1.0;
    const float neg_step = (float)(-(grp->lr / bc1));
    const float bc2_sqrt = (float)pow(bc2, 0.5);
    const float eps = (float)grp->eps;
    const float wd = (float)grp->weight_decay;
    const float decay = (float)(1.0 - grp->lr * grp->weight_decay);
    const int has_wd = grp->weight_decay != 0.0;
    const int adamw = grp->adam_w_mode;

#if defined(__x86_64__) && defined(__GNUC__)
    if (!oracle_force_generic && g_dt == DT_F16 && s_dt == DT_F32 && (!copy || c_dt == DT_F16) && have_f16c()) {
        adam_o2_f16c((const uint16_t*)g, (float*)p, (float*)m, (float*)v, (uint16_t*)copy, n, inv_scale, w1, b2, w2,
                     neg_step, bc2_sqrt, eps, wd, decay, has_wd, adamw);
        return;
    }
#endif
    #pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float gi = load_as_float(g, g_dt, i) * inv_scale;
        float pi = load_as_float(p, s_dt, i);
        float mi = load_as_float(m, s_dt, i);
        float vi = load_as_float(v, s_dt, i);
        if (has_wd) {
            if (adamw) pi = pi * decay;
            else { float t = wd * pi; gi = gi + t; }
        }
        /* lerp_: ATen picks the form by |weight| < 0.5 */
        float d = gi - mi;
        if (w1 < 0.5f) { float t = w1 * d; mi = mi + t; }
        else { float t = d * (1.0f - w1); mi = gi - t; }
        float vb = vi * b2;
        float t1 = w2 * gi;
        float t2 = t1 * gi;
        vi = vb + t2;
        float denom = sqrtf(vi) / bc2_sqrt + eps;
        float q = mi / denom;
        float u = neg_step * q;
        pi = pi + u;
        store_from_float(p, s_dt, i, pi);
        store_from_float(m, s_dt, i, mi);
        store_from_float(v, s_dt, i, vi);
        if (copy) store_from_float(copy, c_dt, i, s_dt == DT_F32 ? pi : load_as_float(p, s_dt, i));
    }
}

/* ---- fused unscale + SGD + master->model copy -------------------------------------------- */
typedef struct {
    double  lr, momentum, dampening, weight_decay;
    int32_t nesterov;
    int32_t first_run;
    int32_t wd_after_momentum;
    int32_t reserved;
} oracle_sgd_group;

/* torch.optim.SGD single-tensor path (TORCH/optim/sgd.py:340-380: wd :356, first step
 * buf = grad.clone() :361-363, buf.mul_(momentum).add_(grad, alpha=1-dampening) :365,
 * nesterov :367-370, param.add_(grad, alpha=-lr) :380) plus apex FusedSGD's extras
 * (SURVEY.md App. B.6): in-kernel grad scale, wd_after_momentum, first_run, model copy. */
void oracle_sgd(const void* g, int g_dt, void* p, void* mom, int s_dt,
                void* copy, int c_dt, int64_t n,
                const oracle_sgd_group* grp, float inv_scale) {
    const float wd = (float)grp->weight_decay;
    const float mu = (float)grp->momentum;
    const float damp = (float)(1.0 - grp->dampening);
    const float neg_lr = (float)(-grp->lr);
    const int has_wd = grp->weight_decay != 0.0;
    const int has_mom = grp->momentum != 0.0;

    #pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float gi = load_as_float(g, g_dt, i) * inv_scale;
        float pi = load_as_float(p, s_dt, i);
        if (has_wd && !grp->wd_after_momentum) { float t = wd * pi; gi = gi + t; }
        if (has_mom) {
            float bi;
            if (grp->first_run) bi = gi;
            else {
                float a = load_as_float(mom, s_dt, i) * mu;
                float t = damp * gi;
                bi = a + t;
            }
            store_from_float(mom, s_dt, i, bi);
            if (s_dt != DT_F32) bi = load_as_float(mom, s_dt, i);   /* what later ops see is the stored value */
            if (grp->nesterov) { float t = mu * bi; gi = gi + t; }
            else gi = bi;
        }
        if (has_wd && grp->wd_after_momentum) { float t = wd * pi; gi = gi + t; }
        float u = neg_lr * gi;
        pi = pi + u;
        store_from_float(p, s_dt, i, pi);
        if (copy) store_from_float(copy, c_dt, i, s_dt == DT_F32 ? pi : load_as_float(p, s_dt, i));
    }
}

/* ---- loss-scaler state machine ------------------------------------------------------------ */
typedef struct {
    float   loss_scale;
    float   inv_scale;
    int32_t found_inf;
    int32_t unskipped;
    int32_t overflow_acc;
    int32_t seq;
    int32_t had_overflow;
    float   inv_loss_scale;
    int32_t skipped_total;
    int32_t reserved[3];
} oracle_scaler_state;

/* apex LossScaler.update_scale; equal to torch._amp_update_scale_
 * (TORCH/amp/grad_scaler.py:549-556) with backoff 0.5 / growth `scale_factor`, plus apex's
 * max_loss_scale (2**24) and optional min_loss_scale which torch lacks. */
void oracle_scaler_update(oracle_scaler_state* s, int mode, int scale_window,
                          float scale_factor, float min_scale, float max_scale) {
    /* mode bit 0: dynamic scaling; bit 1: "no optimizer step has run since the previous update" — apex's
     * one-shot skip_step stays patched in until step() is called, so a skip decided by
     * an earlier backward of a gradient-accumulation cycle survives later clean backwards. */
    int dynamic = mode & 1;
    int carried = (mode & 2) ? s->found_inf : 0;
    int flag = s->overflow_acc != 0;
    s->inv_scale = 1.0f / s->loss_scale;
    s->had_overflow = flag;
    s->found_inf = ((dynamic && flag) || carried) ? 1 : 0;
    if (dynamic) {
        if (flag) {
            float ns = s->loss_scale / 2.0f;
            if (min_scale > 0.0f && ns < min_scale) ns = min_scale;
            s->loss_scale = ns;
            s->unskipped = 0;
        } else {
            s->unskipped += 1;
        }
        if (s->unskipped == scale_window) {
            float ns = s->loss_scale * scale_factor;
            if (ns > max_scale) ns = max_scale;
            s->loss_scale = ns;
            s->unskipped = 0;
        }
    }
    s->overflow_acc = 0;
    s->seq += 1;
    s->skipped_total += (s->found_inf && !carried) ? 1 : 0;
    s->inv_loss_scale = 1.0f / s->loss_scale;
}

/* ---- DDP bucket epilogue -------------------------------------------------------------------- */
/* apex allreduce_bucket after the NCCL SUM: tensor.mul_(1/world_size) in the gradient dtype
 * (SURVEY.md App. B.7), with the overflow check of the following unscale folded in. */
void oracle_bucket_epilogue(void* x, int dt, int64_t n, float factor, int* flag) {
    int f = 0;
    #pragma omp parallel for reduction(|:f) schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float r = load_as_float(x, dt, i) * factor;
        store_from_float(x, dt, i, r);
        f |= !finite_f(load_as_float(x, dt, i));
    }
    if (f && flag) *flag = 1;
}

/* ---- gradient mean across ranks, restated for the peer-memory path ---------------------------------
 * apex allreduce_bucket: all_reduce(SUM) in the gradient dtype, then mul_(1/world) (SURVEY.md
 * App. B.7).  NCCL does not specify its summation order, so the restatement fixes one: fp32
 * accumulation over ranks 0..world-1, ONE rounding to the gradient dtype (the "sum in the
 * gradient dtype" result, where fp16 can overflow to inf although the mean would not), then
 * x factor with a second rounding.  Flag on a non-finite result. */
void oracle_peer_reduce(const void* const* bufs, int world, void* out, int dt, int64_t begin, int64_t end,
                        float factor, int* flag) {
    int f = 0;
    #pragma omp parallel for reduction(|:f) schedule(static)
    for (int64_t i = begin; i < end; i++) {
        float acc = load_as_float(bufs[0], dt, i);
        for (int r = 1; r < world; r++) acc = acc + load_as_float(bufs[r], dt, i);
        store_from_float(out, dt, i, acc);                       /* rounding 1 */
        float v = load_as_float(out, dt, i) * factor;
        store_from_float(out, dt, i, v);                         /* rounding 2 */
        f |= !finite_f(load_as_float(out, dt, i));
    }
    if (f && flag) *flag = 1;
}

/* ---- whole-list step, used as the timed CPU baseline ------------------------------------------
 * One O2 step over a tensor list the way the path runs it: check every gradient, and unless
 * an overflow was found, unscale + Adam + model copy.  Returns the flag. */
int oracle_adam_list(int n_tensors, const void* const* g, int g_dt,
                     void* const* p, void* const* m, void* const* v, int s_dt,
                     void* const* copy, int c_dt, const int64_t* numel,
                     const oracle_adam_group* grp, float inv_scale, int check) {
    int flag = 0;
    if (check)
        for (int t = 0; t < n_tensors; t++) flag |= oracle_check(g[t], g_dt, numel[t]);
    if (flag) return 1;
    for (int t = 0; t < n_tensors; t++)
        oracle_adam(g[t], g_dt, p[t], m[t], v[t], s_dt, copy ? copy[t] : 0, c_dt, numel[t], grp, inv_scale);
    return 0;
}

int oracle_sgd_list(int n_tensors, const void* const* g, int g_dt,
                    void* const* p, void* const* mom, int s_dt,
                    void* const* copy, int c_dt, const int64_t* numel,
                    const oracle_sgd_group* grp, float inv_scale, int check) {
    int flag = 0;
    if (check)
        for (int t = 0; t < n_tensors; t++) flag |= oracle_check(g[t], g_dt, numel[t]);
    if (flag) return 1;
    for (int t = 0; t < n_tensors; t++)
        oracle_sgd(g[t], g_dt, p[t], mom ? mom[t] : 0, s_dt, copy ? copy[t] : 0, c_dt, numel[t], grp, inv_scale);
    return 0;
}

/* ---- per-tensor L2 norms and LAMB -------------------------------------
 * PARITY UNPINNED against apex: apex is a pip dependency of the reference (test_apex_distributed_spawn.py:14-16),
 * not vendored, and cannot be built here.  The rule below restates the LAMB paper (You et al. 2020, Alg. 2) with the
 * conventions apex.optimizers.FusedLAMB documents: global-norm clip of the gradients before the moments
 * (max_grad_norm), beta3 = 1-beta1 when grad_averaging, bias correction, L2 or decoupled decay, trust ratio
 * ||p||/||u|| applied when weight_decay != 0 or use_nvlamb, lr otherwise.  tests/test_gpu_lamb.py additionally
 * compares with a straight torch restatement of the same published rule within a tolerance.
 * The summation ORDER is part of the contract (the trust ratio depends on it bit for bit): a tile of `tile_elems`
 * elements is summed by 256 strided accumulators (fused multiply-add), folded by a 32-wide butterfly and then over the
 * 8 groups in order; a tensor's tiles are summed by 32 strided accumulators + butterfly; tensors likewise. */
static float butterfly32(float* x) {
    for (int o = 16; o > 0; o >>= 1) {
        float y[32];
        for (int l = 0; l < 32; l++) y[l] = x[l] + x[l ^ o];
        memcpy(x, y, sizeof(y));
    }
    return x[0];
}

static float tile_sumsq(const void* x, int dt, int64_t start, int n, int* bad) {
    float acc[256];
    for (int t = 0; t < 256; t++) acc[t] = 0.0f;
    for (int i = 0; i < n; i++) {
        const float v = load_as_float(x, dt, start + i);
        if (!finite_f(v)) *bad = 1;
        acc[i & 255] = fmaf(v, v, acc[i & 255]);
    }
    float tot = 0.0f;
    for (int w = 0; w < 8; w++) tot += butterfly32(acc + 32 * w);
    return tot;
}

static float strided_sum32(const float* x, int64_t n) {
    float acc[32];
    for (int l = 0; l < 32; l++) acc[l] = 0.0f;
    for (int64_t i = 0; i < n; i++) acc[i & 31] += x[i];
    return butterfly32(acc);
}

/* sumsq[t] for every tensor; returns the inf/NaN flag; *total += sum over tensors (if total) */
int oracle_l2norm_per_tensor(int n_tensors, const void* const* x, int dt, const int64_t* numel, int tile_elems,
                             float* sumsq, float* total) {
    int bad = 0;
    for (int t = 0; t < n_tensors; t++) {
        const int64_t tiles = numel[t] <= 0 ? 0 : (numel[t] + tile_elems - 1) / tile_elems;
        float* ss = (float*)malloc(sizeof(float) * (size_t)(tiles > 0 ? tiles : 1));
        for (int64_t k = 0; k < tiles; k++) {
            const int64_t start = k * tile_elems;
            const int64_t left = numel[t] - start;
            ss[k] = tile_sumsq(x[t], dt, start, (int)(left < tile_elems ? left : tile_elems), &bad);
        }
        sumsq[t] = strided_sum32(ss, tiles);
        free(ss);
    }
    if (total) *total += strided_sum32(sumsq, n_tensors);
    return bad;
}

typedef struct {
    double  lr, beta1, beta2, eps, weight_decay;
    int64_t step;
    int32_t bias_correction;
    int32_t adam_w_mode;
    int32_t grad_averaging;
    int32_t use_nvlamb;
} oracle_lamb_group;

/* One LAMB step over a list that shares one hyper-parameter group.  grad_sumsq is the sum of squares of ALL (still
 * scaled) gradients of the optimizer (every list).  u is fp32 scratch per tensor. */
void oracle_lamb_list(int n_tensors, const void* const* g, int g_dt, void* const* p, void* const* m, void* const* v,
                      int s_dt, void* const* copy, int c_dt, float* const* u, const int64_t* numel, int tile_elems,
                      const oracle_lamb_group* grp, float inv_scale, float grad_sumsq, float max_grad_norm) {
    const float b1 = (float)grp->beta1, b2 = (float)grp->beta2, one_m_b2 = (float)(1.0 - grp->beta2);
    const float b3 = grp->grad_averaging ? (float)(1.0 - grp->beta1) : 1.0f;
    const float bc1 = grp->bias_correction ? (float)(1.0 - pow(grp->beta1, (double)grp->step)) : 1.0f;
    const float bc2 = grp->bias_correction ? (float)(1.0 - pow(grp->beta2, (double)grp->step)) : 1.0f;
    const float eps = (float)grp->eps, wd = (float)grp->weight_decay, lr = (float)grp->lr;
    const float gn = sqrtf(grad_sumsq) * inv_scale;
    const float clip = (max_grad_norm > 0.0f && gn > max_grad_norm) ? gn / max_grad_norm : 1.0f;
    for (int t = 0; t < n_tensors; t++)
        for (int64_t i = 0; i < numel[t]; i++) {
            float gi = load_as_float(g[t], g_dt, i) * inv_scale;
            gi = gi / clip;
            const float pi = load_as_float(p[t], s_dt, i);
            float mi = load_as_float(m[t], s_dt, i), vi = load_as_float(v[t], s_dt, i);
            if (!grp->adam_w_mode) { float d = wd * pi; gi = gi + d; }
            { float a = b1 * mi, b = b3 * gi; mi = a + b; }
            { float a = b2 * vi, b = one_m_b2 * gi; b = b * gi; vi = a + b; }
            const float mh = mi / bc1, vh = vi / bc2;
            float ui = mh / (sqrtf(vh) + eps);
            if (grp->adam_w_mode) { float d = wd * pi; ui = ui + d; }
            store_from_float(m[t], s_dt, i, mi);
            store_from_float(v[t], s_dt, i, vi);
            u[t][i] = ui;
        }
    float* pn = (float*)malloc(sizeof(float) * (size_t)(n_tensors + 1));
    float* un = (float*)malloc(sizeof(float) * (size_t)(n_tensors + 1));
    oracle_l2norm_per_tensor(n_tensors, (const void* const*)p, s_dt, numel, tile_elems, pn, 0);
    oracle_l2norm_per_tensor(n_tensors, (const void* const*)u, DT_F32, numel, tile_elems, un, 0);
    for (int t = 0; t < n_tensors; t++) {
        float ratio = lr;
        if (grp->use_nvlamb || grp->weight_decay != 0.0) {
            const float a = sqrtf(pn[t]), b = sqrtf(un[t]);
            if (a != 0.0f && b != 0.0f) ratio = lr * (a / b);
        }
        for (int64_t i = 0; i < numel[t]; i++) {
            float pi = load_as_float(p[t], s_dt, i);
            { float d = ratio * u[t][i]; pi = pi - d; }
            store_from_float(p[t], s_dt, i, pi);
            if (copy) store_from_float(copy[t], c_dt, i, s_dt == DT_F32 ? pi : load_as_float(p[t], s_dt, i));
        }
    }
    free(pn);
    free(un);
}

/* ---- NovoGrad ------------------------------------------------------------
 * PARITY UNPINNED against apex (not vendored, cannot be built here).  Restates NovoGrad (Ginsburg et al. 2019,
 * "Stochastic Gradient Methods with Layer-wise Adaptive Moments") with the conventions apex.optimizers.FusedNovoGrad
 * documents: the second moment is one number per tensor, kept as a NORM and blended as sqrt(b2 v^2 + (1-b2) n^2) with
 * n = ||g||_2 (norm_type 2); on the first step it starts from n unless init_zero; bias corrections 1-b1^t and
 * sqrt(1-b2^t); weight decay inside the moment (reg_inside_moment) or on the update (default). */
typedef struct {
    double  lr, beta1, beta2, eps, weight_decay;
    int64_t step;
    int32_t bias_correction;
    int32_t reg_inside_moment;
    int32_t grad_averaging;
    int32_t init_zero;
} oracle_novograd_group;

/* v: one float per tensor (state).  grad_sumsq[t]: sum of squares of the still scaled gradient of tensor t. */
void oracle_novograd_list(int n_tensors, const void* const* g, int g_dt, void* const* p, void* const* m, int s_dt,
                          void* const* copy, int c_dt, float* v, const float* grad_sumsq, const int64_t* numel,
                          const oracle_novograd_group* grp, float inv_scale) {
    const float b1 = (float)grp->beta1, b2 = (float)grp->beta2, one_m_b2 = (float)(1.0 - grp->beta2);
    const float b3 = grp->grad_averaging ? (float)(1.0 - grp->beta1) : 1.0f;
    const float bc1 = grp->bias_correction ? (float)(1.0 - pow(grp->beta1, (double)grp->step)) : 1.0f;
    const float bc2 = grp->bias_correction ? (float)sqrt(1.0 - pow(grp->beta2, (double)grp->step)) : 1.0f;
    const float eps = (float)grp->eps, wd = (float)grp->weight_decay, lr = (float)grp->lr;
    const int first = grp->step == 1 && !grp->init_zero;
    for (int t = 0; t < n_tensors; t++) {
        const float n = sqrtf(grad_sumsq[t]) * inv_scale;
        const float old = first ? n : v[t];
        { float a = old * old; a = b2 * a; float b = n * n; b = one_m_b2 * b; v[t] = sqrtf(a + b); }
        const float denom = v[t] / bc2 + eps;
        for (int64_t i = 0; i < numel[t]; i++) {
            float gi = load_as_float(g[t], g_dt, i) * inv_scale;
            float pi = load_as_float(p[t], s_dt, i), mi = load_as_float(m[t], s_dt, i);
            if (grp->reg_inside_moment) {
                { float a = gi / denom, b = wd * pi; gi = a + b; }
                { float a = b1 * mi, b = b3 * gi; mi = a + b; }
                { float a = mi / bc1; a = lr * a; pi = pi - a; }
            } else {
                { float a = b1 * mi, b = b3 * gi; mi = a + b; }
                { float a = mi / bc1; a = a / denom; float b = wd * pi; a = a + b; a = lr * a; pi = pi - a; }
            }
            store_from_float(p[t], s_dt, i, pi);
            store_from_float(m[t], s_dt, i, mi);
            if (copy) store_from_float(copy[t], c_dt, i, s_dt == DT_F32 ? pi : load_as_float(p[t], s_dt, i));
        }
    }
}
