// kernels.cu — the sm_100a kernels of libapex_b200.so and their extern "C" launchers.
//
// Every kernel here is a single streaming pass bounded by HBM bandwidth: no
// data is reused, so there is no shared-memory tiling and no tensor-core work; what matters is
// 128/256-bit coalesced access, enough bytes in flight per SM, a persistent grid sized in
// multiples of the SM count, and never touching a byte twice.
//
// Arithmetic contract: fp32, IEEE sqrt/div, explicit __f*_rn intrinsics so that nothing is
// contracted into an FMA — results are bit-identical to oracle/amp_step_oracle.c.
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <type_traits>

#include "b200_common.cuh"

namespace b200 {

// ============================================================================================
// host-side state
// ============================================================================================
static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};
static std::atomic<int> g_ctas_per_sm{0};     // 0: per-kernel default
static std::atomic<int> g_threads{256};
static std::atomic<int> g_unroll{2};
static std::atomic<int> g_peer_max_ctas{0};   // 0: fill the GPU; >0: run the peer kernels as a small "communication" grid

int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

static int cuda_fail(cudaError_t e, const char* what) {
    return set_error(static_cast<int>(e), "%s: %s", what, cudaGetErrorString(e));
}

static int sm_count() {
    static int cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

// index of the current device into the per-device caches below (occupancy depends on the device)
static int current_device_slot() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    return dev;
}

// Persistent grids must not exceed what is resident at once (a second wave would run at a
// fraction of the occupancy): CTAs/SM comes from the occupancy calculator, per kernel and device.
template <class Kernel>
static int resident_ctas_per_sm(Kernel kernel, int threads) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, 0) != cudaSuccess || n < 1) {
        cudaGetLastError();
        n = 1;
    }
    return n;
}

static unsigned grid_for(int64_t n_tiles, int resident_per_sm) {
    int per_sm = g_ctas_per_sm.load();
    if (per_sm <= 0 || per_sm > resident_per_sm) per_sm = resident_per_sm;
    int64_t g = static_cast<int64_t>(sm_count()) * per_sm;
    if (g > n_tiles) g = n_tiles;
    if (g < 1) g = 1;
    return static_cast<unsigned>(g);
}

static int after_launch(const char* name) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, name);
    g_launches.fetch_add(1);
    return 0;
}

// ============================================================================================
// generic tile kernel
// ============================================================================================
// per-thread accumulator carried through the walk: a bool ("saw inf/NaN") unless the Op declares its own
template <class Op, class = void> struct AccOf { using type = bool; };
template <class Op> struct AccOf<Op, std::void_t<typename Op::Acc>> { using type = typename Op::Acc; };
// optional per-Op cap on resident CTAs per SM (0 = whatever the occupancy calculator allows)
template <class Op, class = void> struct MaxCtasOf { static constexpr int value = 0; };
template <class Op> struct MaxCtasOf<Op, std::void_t<decltype(Op::kMaxCtasPerSm)>> { static constexpr int value = Op::kMaxCtasPerSm; };

template <class Op, int kThreads, int kUnroll>
__global__ void __launch_bounds__(kThreads)
tile_kernel(const void* __restrict__ table, const __grid_constant__ typename Op::Params prm) {
    int win = 0;
    if (Op::prologue(prm, win)) return;
    typename AccOf<Op>::type acc{};
    walk_tiles<Op, kThreads, kUnroll * Op::kUnrollMul>(table, prm, win, Op::rotation(prm), acc);
    Op::finish(prm, acc);
}

template <class Op, int kThreads, int kUnroll>
static int launch_one(const char* name, const void* dev_table, int64_t n_tiles, const typename Op::Params& prm,
                      cudaStream_t st, unsigned max_grid) {
    auto kernel = tile_kernel<Op, kThreads, kUnroll>;
    static int occ[64];                       // per device; one static per instantiation
    const int dev = current_device_slot();
    if (occ[dev] == 0) {
        occ[dev] = resident_ctas_per_sm(kernel, kThreads);
        if constexpr (MaxCtasOf<Op>::value > 0)
            if (occ[dev] > MaxCtasOf<Op>::value) occ[dev] = MaxCtasOf<Op>::value;
    }
    unsigned grid = grid_for(n_tiles, occ[dev]);
    if (max_grid && grid > max_grid) grid = max_grid;
    kernel<<<grid, kThreads, 0, st>>>(dev_table, prm);
    return after_launch(name);
}

template <class Op>
static int launch_tiles(const char* name, const void* dev_table, int64_t n_tiles,
                        const typename Op::Params& prm, void* stream, unsigned max_grid = 0) {
    if (n_tiles <= 0) return 0;
    if (dev_table == nullptr) return set_error(B200_EINVAL, "%s: dev_table is NULL", name);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int threads = g_threads.load(), unroll = g_unroll.load();
    if (threads == 512) {
        if (unroll == 1) return launch_one<Op, 512, 1>(name, dev_table, n_tiles, prm, st, max_grid);
        return launch_one<Op, 512, 2>(name, dev_table, n_tiles, prm, st, max_grid);
    }
    if (unroll == 1) return launch_one<Op, 256, 1>(name, dev_table, n_tiles, prm, st, max_grid);
    return launch_one<Op, 256, 2>(name, dev_table, n_tiles, prm, st, max_grid);
}

// ============================================================================================
// multi_tensor_scale / axpby / check
// ============================================================================================
struct ScaleParams {
    float scale;
    const float* dev_scale;
    int32_t* flag;
};

template <int I, int O>
struct ScaleOp {
    using Params = ScaleParams;
    static constexpr int kUnrollMul = 2;     // 1-2 light columns: keep more groups in flight
    struct Ctx { float s; };
    static __device__ __forceinline__ bool prologue(const Params&, int&) { return false; }
    static __device__ __forceinline__ uint32_t rotation(const Params&) { return 0u; }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row&, int, int64_t) {
        return Ctx{p.dev_scale ? __ldg(p.dev_scale) : p.scale};
    }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], bool& bad) {
        Raw8<I> in[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (act[u]) load8<I, true>(r.ptr[0], idx[u], in[u]);
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            float x[kVec];
            unpack8<I>(in[u], x);
#pragma unroll
            for (int i = 0; i < kVec; i++) {
                bad |= !is_finite(x[i]);
                x[i] = __fmul_rn(x[i], c.s);
            }
            store8<O>(r.ptr[1], idx[u], x);
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, bool& bad) {
        const float x = load1<I>(r.ptr[0], i);
        bad |= !is_finite(x);
        store1<O>(r.ptr[1], i, __fmul_rn(x, c.s));
    }
    static __device__ __forceinline__ void finish(const Params& p, bool bad) {
        if (p.flag) raise_flag(p.flag, bad);
    }
};

struct AxpbyParams {
    float a, b;
    int32_t arg_to_check;
    int32_t* flag;
};

template <int X, int Y, int O>
struct AxpbyOp {
    using Params = AxpbyParams;
    static constexpr int kUnrollMul = 1;
    struct Ctx { float a, b; int32_t chk; };
    static __device__ __forceinline__ bool prologue(const Params&, int&) { return false; }
    static __device__ __forceinline__ uint32_t rotation(const Params&) { return 0u; }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row&, int, int64_t) {
        return Ctx{p.a, p.b, p.arg_to_check};
    }
    static __device__ __forceinline__ float one(float x, float y, const Ctx& c, bool& bad) {
        if (c.chk == 0) bad |= !is_finite(x);
        else if (c.chk == 1) bad |= !is_finite(y);
        else bad |= (!is_finite(x)) | (!is_finite(y));
        return __fadd_rn(__fmul_rn(c.a, x), __fmul_rn(c.b, y));
    }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], bool& bad) {
        Raw8<X> xr[U];
        Raw8<Y> yr[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (act[u]) {
                load8<X, true>(r.ptr[0], idx[u], xr[u]);
                load8<Y, false>(r.ptr[1], idx[u], yr[u]);   // y may alias out (in-place accumulate)
            }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            float x[kVec], y[kVec];
            unpack8<X>(xr[u], x);
            unpack8<Y>(yr[u], y);
#pragma unroll
            for (int i = 0; i < kVec; i++) x[i] = one(x[i], y[i], c, bad);
            store8<O>(r.ptr[2], idx[u], x);
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, bool& bad) {
        store1<O>(r.ptr[2], i, one(load1<X>(r.ptr[0], i), load1<Y>(r.ptr[1], i), c, bad));
    }
    static __device__ __forceinline__ void finish(const Params& p, bool bad) {
        if (p.flag) raise_flag(p.flag, bad);
    }
};

struct CheckParams {
    int32_t col;
    int32_t* flag;
};

template <int I>
struct CheckOp {
    using Params = CheckParams;
    static constexpr int kUnrollMul = 4;     // read-only, one column: 4-8 loads in flight per thread
    struct Ctx { int32_t col; };
    static __device__ __forceinline__ bool prologue(const Params&, int&) { return false; }
    static __device__ __forceinline__ uint32_t rotation(const Params&) { return 0u; }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row&, int, int64_t) { return Ctx{p.col}; }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], bool& bad) {
        Raw8<I> in[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (act[u]) load8<I, true>(r.ptr[c.col], idx[u], in[u]);
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            float x[kVec];
            unpack8<I>(in[u], x);
#pragma unroll
            for (int i = 0; i < kVec; i++) bad |= !is_finite(x[i]);
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, bool& bad) {
        bad |= !is_finite(load1<I>(r.ptr[c.col], i));
    }
    static __device__ __forceinline__ void finish(const Params& p, bool bad) { raise_flag(p.flag, bad); }
};

// ============================================================================================
// multi_tensor_l2norm fused with the overflow check (one read-only pass over the gradients)
// ============================================================================================
struct NormParams {
    int32_t col;
    int32_t* flag;          // may be NULL
    float* scratch;         // [0] ticket (as uint32), [1] pad, [2 + b] partial sum of squares of block b
    float* out;             // out[0] += sum of squares of this list (the caller zeroes it once per step)
};

struct NormAcc {
    float ss;
    bool bad;
};

template <int I>
struct NormCheckOp {
    using Params = NormParams;
    using Acc = NormAcc;
    static constexpr int kUnrollMul = 4;
    struct Ctx { int32_t col; };
    static __device__ __forceinline__ bool prologue(const Params&, int&) { return false; }
    static __device__ __forceinline__ uint32_t rotation(const Params&) { return 0u; }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row&, int, int64_t) { return Ctx{p.col}; }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], Acc& acc) {
        Raw8<I> in[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (act[u]) load8<I, true>(r.ptr[c.col], idx[u], in[u]);
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            float x[kVec];
            unpack8<I>(in[u], x);
            float s = 0.0f;
#pragma unroll
            for (int i = 0; i < kVec; i++) {
                acc.bad |= !is_finite(x[i]);
                s = __fmaf_rn(x[i], x[i], s);
            }
            acc.ss += s;
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, Acc& acc) {
        const float x = load1<I>(r.ptr[c.col], i);
        acc.bad |= !is_finite(x);
        acc.ss = __fmaf_rn(x, x, acc.ss);
    }
    // Deterministic for a given grid: block partials are written to scratch and summed in index order
    // by the last block to arrive (ticket), never with floating-point atomics.
    static __device__ __forceinline__ void finish(const Params& p, const Acc& acc) {
        if (p.flag) raise_flag(p.flag, acc.bad);
        __shared__ float s_warp[32];
        __shared__ bool s_last;
        float v = acc.ss;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
        if (lane == 0) s_warp[warp] = v;
        __syncthreads();
        if (warp == 0) {
            float t = lane < n_warps ? s_warp[lane] : 0.0f;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
            if (lane == 0) {
                p.scratch[2 + blockIdx.x] = t;
                __threadfence();
                const unsigned ticket = atomicAdd(reinterpret_cast<unsigned*>(p.scratch), 1u);
                s_last = ticket == gridDim.x - 1;
            }
        }
        __syncthreads();
        if (s_last && warp == 0) {
            __threadfence();
            float t = 0.0f;
            for (unsigned b = lane; b < gridDim.x; b += 32) t += *(reinterpret_cast<volatile float*>(p.scratch) + 2 + b);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
            if (lane == 0) {
                p.out[0] += t;
                *reinterpret_cast<unsigned*>(p.scratch) = 0u;       // ready for the next launch
            }
        }
    }
};

// clip coefficient on the device: scale_out = inv_scale * min(1, max_norm / (||g|| * inv_scale + 1e-6))
// (torch.nn.utils.clip_grad_norm_'s rule), norm_out = ||g|| * inv_scale.  One thread.
__global__ void clip_scale_kernel(float* sumsq, const float* dev_inv_scale, float inv_scale, float max_norm,
                                  float* scale_out, float* norm_out) {
    const float inv = dev_inv_scale ? *dev_inv_scale : inv_scale;
    const float norm = __fmul_rn(__fsqrt_rn(*sumsq), inv);
    *sumsq = 0.0f;                                   // consumed: the accumulator is ready for the next step
    float coef = __fdiv_rn(max_norm, __fadd_rn(norm, 1e-6f));
    if (coef > 1.0f) coef = 1.0f;
    *scale_out = __fmul_rn(inv, coef);
    if (norm_out) *norm_out = norm;
}

// ============================================================================================
// per-tensor L2 norms (amp_C.multi_tensor_l2norm(per_tensor=True)) and FusedLAMB
// ============================================================================================
// Deterministic by construction: one CTA per tile sums its squares in a fixed order, one warp per tensor sums
// the tensor's tiles in a fixed order, one warp sums the tensors.  No floating-point atomics.
template <int DT, int kThreads>
__global__ void __launch_bounds__(kThreads)
tile_sumsq_kernel(const void* __restrict__ table, int col, const uint32_t* __restrict__ first_tile,
                  float* __restrict__ tile_ss, int32_t* flag) {
    const TableView tv = open_table(table);
    const uint32_t t = blockIdx.x;
    if (t >= tv.n_tiles) return;
    const b200_table_tile tl = tv.tiles[t];
    const b200_table_row& row = tv.rows[tl.tensor];
    const int64_t start = static_cast<int64_t>(tl.tile) * tv.tile_elems;
    const int64_t left = row.numel - start;
    const int n = left < static_cast<int64_t>(tv.tile_elems) ? static_cast<int>(left) : static_cast<int>(tv.tile_elems);
    float s = 0.0f;
    bool bad = false;
    for (int i = threadIdx.x; i < n; i += kThreads) {
        const float x = load1<DT>(row.ptr[col], start + i);
        bad |= !is_finite(x);
        s = __fmaf_rn(x, x, s);
    }
    __shared__ float s_warp[kThreads / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float tot = 0.0f;
        for (int w = 0; w < kThreads / 32; w++) tot += s_warp[w];
        tile_ss[first_tile[tl.tensor] + tl.tile] = tot;  // natural position: independent of the tile order
    }
    if (flag) raise_flag(flag, bad);
}

// first_tile[t] .. first_tile[t + 1]: the tiles of tensor t in tile_ss (exclusive prefix sum of the tile counts).  Block b handles tensor b (one warp).
__global__ void tensor_norm_kernel(const float* __restrict__ tile_ss, const uint32_t* __restrict__ first_tile, int n_tensors,
                                   float* __restrict__ sumsq_out, float* __restrict__ norm_out) {
    const int t = blockIdx.x;
    if (t >= n_tensors) return;
    float s = 0.0f;
    for (uint32_t i = first_tile[t] + threadIdx.x; i < first_tile[t + 1]; i += 32) s += tile_ss[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (threadIdx.x == 0) {
        if (sumsq_out) sumsq_out[t] = s;
        if (norm_out) norm_out[t] = __fsqrt_rn(s);
    }
}

__global__ void total_sumsq_kernel(const float* __restrict__ sumsq, int n, float* __restrict__ total) {
    float s = 0.0f;
    for (int i = threadIdx.x; i < n; i += 32) s += sumsq[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (threadIdx.x == 0) total[0] += s;                // several lists (dtypes) accumulate into the same total
}

struct LambConsts {
    float b1, b3, b2, one_m_b2, bc1, bc2, eps, decay, lr;
    uint32_t flags;                 // bit0: decoupled decay (adam_w_mode), bit1: trust ratio on even without decay (nvlamb)
};

struct Lamb1Params {
    LambConsts grp[B200_MAX_GROUPS];
    float inv_scale;
    const float* dev_inv_scale;
    const int32_t* dev_found_inf;
    const float* dev_grad_sumsq;    // sum of squares of the (still scaled) gradients of ALL lists
    float max_grad_norm;
};

// stage 1: moments and the un-trusted update u (column 4, fp32); p is only read
template <int G, int S>
struct Lamb1Op {
    using Params = Lamb1Params;
    static constexpr int kUnrollMul = 1;
    struct Ctx { LambConsts k; float s, clip; };
    static __device__ __forceinline__ uint32_t rotation(const Params&) { return 0u; }
    static __device__ __forceinline__ bool prologue(const Params& p, int&) {
        return p.dev_found_inf != nullptr && *reinterpret_cast<const volatile int32_t*>(p.dev_found_inf) != 0;
    }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row& r, int, int64_t) {
        const float s = p.dev_inv_scale ? __ldg(p.dev_inv_scale) : p.inv_scale;
        const float gn = __fmul_rn(__fsqrt_rn(__ldg(p.dev_grad_sumsq)), s);     // norm of the unscaled gradients
        const float clip = (p.max_grad_norm > 0.0f && gn > p.max_grad_norm) ? __fdiv_rn(gn, p.max_grad_norm) : 1.0f;
        return Ctx{p.grp[r.group], s, clip};
    }
    static __device__ __forceinline__ float one(float g_in, float p, float& m, float& v, const Ctx& c) {
        float g = __fdiv_rn(__fmul_rn(g_in, c.s), c.clip);
        if (!(c.k.flags & 1u)) g = __fadd_rn(g, __fmul_rn(c.k.decay, p));
        m = __fadd_rn(__fmul_rn(c.k.b1, m), __fmul_rn(c.k.b3, g));
        v = __fadd_rn(__fmul_rn(c.k.b2, v), __fmul_rn(__fmul_rn(c.k.one_m_b2, g), g));
        const float mh = __fdiv_rn(m, c.k.bc1), vh = __fdiv_rn(v, c.k.bc2);
        float u = __fdiv_rn(mh, __fadd_rn(__fsqrt_rn(vh), c.k.eps));
        if (c.k.flags & 1u) u = __fadd_rn(u, __fmul_rn(c.k.decay, p));
        return u;
    }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], bool&) {
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            Raw8<G> rg;
            Raw8<S> rp, rm, rv;
            load8<G, true>(r.ptr[0], idx[u], rg);
            load8<S, true>(r.ptr[1], idx[u], rp);
            load8<S, false>(r.ptr[2], idx[u], rm);
            load8<S, false>(r.ptr[3], idx[u], rv);
            float g[kVec], p[kVec], m[kVec], v[kVec], o[kVec];
            unpack8<G>(rg, g);
            unpack8<S>(rp, p);
            unpack8<S>(rm, m);
            unpack8<S>(rv, v);
#pragma unroll
            for (int i = 0; i < kVec; i++) o[i] = one(g[i], p[i], m[i], v[i], c);
            store8<S>(r.ptr[2], idx[u], m);
            store8<S>(r.ptr[3], idx[u], v);
            store8<B200_F32>(r.ptr[4], idx[u], o);
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, bool&) {
        float m = load1<S>(r.ptr[2], i), v = load1<S>(r.ptr[3], i);
        const float u = one(load1<G>(r.ptr[0], i), load1<S>(r.ptr[1], i), m, v, c);
        store1<S>(r.ptr[2], i, m);
        store1<S>(r.ptr[3], i, v);
        store1<B200_F32>(r.ptr[4], i, u);
    }
    static __device__ __forceinline__ void finish(const Params&, bool) {}
};

struct Lamb2Params {
    LambConsts grp[B200_MAX_GROUPS];
    const int32_t* dev_found_inf;
    const float* p_norm;            // per tensor (row index)
    const float* u_norm;
};

// stage 2: p -= trust_ratio * u, columns 0 = u (fp32), 1 = p, 2 = model copy (optional)
template <int S, int C>
struct Lamb2Op {
    using Params = Lamb2Params;
    static constexpr int kUnrollMul = 1;
    struct Ctx { float ratio; };
    static __device__ __forceinline__ uint32_t rotation(const Params&) { return 0u; }
    static __device__ __forceinline__ bool prologue(const Params& p, int&) {
        return p.dev_found_inf != nullptr && *reinterpret_cast<const volatile int32_t*>(p.dev_found_inf) != 0;
    }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row& r, int, int64_t) {
        const LambConsts& k = p.grp[r.group];
        float ratio = k.lr;
        if ((k.flags & 2u) || k.decay != 0.0f) {
            const float pn = __ldg(p.p_norm + r.reserved), un = __ldg(p.u_norm + r.reserved);
            if (pn != 0.0f && un != 0.0f) ratio = __fmul_rn(k.lr, __fdiv_rn(pn, un));
        }
        return Ctx{ratio};
    }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], bool&) {
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            Raw8<B200_F32> ru;
            Raw8<S> rp;
            load8<B200_F32, true>(r.ptr[0], idx[u], ru);
            load8<S, false>(r.ptr[1], idx[u], rp);
            float uu[kVec], p[kVec];
            unpack8<B200_F32>(ru, uu);
            unpack8<S>(rp, p);
#pragma unroll
            for (int i = 0; i < kVec; i++) p[i] = __fsub_rn(p[i], __fmul_rn(c.ratio, uu[i]));
            store8<S>(r.ptr[1], idx[u], p);
            if constexpr (C != B200_NONE) {
                if constexpr (S != B200_F32) {
#pragma unroll
                    for (int i = 0; i < kVec; i++) p[i] = round_through<S>(p[i]);
                }
                store8<C>(r.ptr[2], idx[u], p);
            }
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, bool&) {
        const float p = __fsub_rn(load1<S>(r.ptr[1], i), __fmul_rn(c.ratio, load1<B200_F32>(r.ptr[0], i)));
        store1<S>(r.ptr[1], i, p);
        if constexpr (C != B200_NONE) store1<C>(r.ptr[2], i, round_through<S>(p));
    }
    static __device__ __forceinline__ void finish(const Params&, bool) {}
};

// ============================================================================================
// NovoGrad: layer-wise second moments (one float per tensor) + momentum update
// ============================================================================================
struct NovoConsts {
    float b1, b3, b2, one_m_b2, bc1, bc2, eps, decay, lr;
    uint32_t flags;                 // bit0: regularise inside the moment (apex MOMENT_MODE_0), bit1: first step and !init_zero
};

struct NovoParams {
    NovoConsts grp[B200_MAX_GROUPS];
    float inv_scale;
    const float* dev_inv_scale;
    const int32_t* dev_found_inf;
    const float* v;                 // per tensor (row index): the blended gradient norm of this step
};

// v[t] <- sqrt(b2 * v[t]^2 + (1 - b2) * n^2), n = sqrt(sumsq[t]) * inv_scale (norm of the unscaled gradient of tensor t);
// on the first step (unless init_zero) v[t] starts from n.  One thread per tensor.
__global__ void novograd_blend_kernel(const void* __restrict__ table, const __grid_constant__ NovoParams prm,
                                      float* __restrict__ v, const float* __restrict__ grad_sumsq) {
    if (prm.dev_found_inf != nullptr && *reinterpret_cast<const volatile int32_t*>(prm.dev_found_inf) != 0) return;
    const TableView tv = open_table(table);
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= reinterpret_cast<const b200_table_header*>(table)->n_tensors) return;
    const NovoConsts& k = prm.grp[tv.rows[t].group];
    const float s = prm.dev_inv_scale ? __ldg(prm.dev_inv_scale) : prm.inv_scale;
    const float n = __fmul_rn(__fsqrt_rn(grad_sumsq[t]), s);
    float old = (k.flags & 2u) ? n : v[t];
    const float a = __fmul_rn(k.b2, __fmul_rn(old, old));
    const float b = __fmul_rn(k.one_m_b2, __fmul_rn(n, n));
    v[t] = __fsqrt_rn(__fadd_rn(a, b));
}

// columns: 0 g, 1 p, 2 m, 3 model copy (optional)
template <int G, int S, int C>
struct NovoGradOp {
    using Params = NovoParams;
    static constexpr int kUnrollMul = 1;
    struct Ctx { NovoConsts k; float s, denom; };
    static constexpr int kCopyCol = 3;
    static __device__ __forceinline__ uint32_t rotation(const Params&) { return 0u; }
    static __device__ __forceinline__ bool prologue(const Params& p, int&) {
        return p.dev_found_inf != nullptr && *reinterpret_cast<const volatile int32_t*>(p.dev_found_inf) != 0;
    }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row& r, int, int64_t) {
        const NovoConsts& k = p.grp[r.group];
        const float s = p.dev_inv_scale ? __ldg(p.dev_inv_scale) : p.inv_scale;
        return Ctx{k, s, __fadd_rn(__fdiv_rn(__ldg(p.v + r.reserved), k.bc2), k.eps)};
    }
    static __device__ __forceinline__ void one(float g_in, float& p, float& m, const Ctx& c) {
        float g = __fmul_rn(g_in, c.s);
        if (c.k.flags & 1u) {
            g = __fadd_rn(__fdiv_rn(g, c.denom), __fmul_rn(c.k.decay, p));
            m = __fadd_rn(__fmul_rn(c.k.b1, m), __fmul_rn(c.k.b3, g));
            p = __fsub_rn(p, __fmul_rn(c.k.lr, __fdiv_rn(m, c.k.bc1)));
        } else {
            m = __fadd_rn(__fmul_rn(c.k.b1, m), __fmul_rn(c.k.b3, g));
            const float u = __fadd_rn(__fdiv_rn(__fdiv_rn(m, c.k.bc1), c.denom), __fmul_rn(c.k.decay, p));
            p = __fsub_rn(p, __fmul_rn(c.k.lr, u));
        }
    }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], bool&) {
        Raw8<G> rg[U];
        Raw8<S> rp[U], rm[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (act[u]) {
                load8<G, true>(r.ptr[0], idx[u], rg[u]);
                load8<S, false>(r.ptr[1], idx[u], rp[u]);
                load8<S, false>(r.ptr[2], idx[u], rm[u]);
            }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            float g[kVec], p[kVec], m[kVec];
            unpack8<G>(rg[u], g);
            unpack8<S>(rp[u], p);
            unpack8<S>(rm[u], m);
#pragma unroll
            for (int i = 0; i < kVec; i++) one(g[i], p[i], m[i], c);
            store8<S>(r.ptr[1], idx[u], p);
            store8<S>(r.ptr[2], idx[u], m);
            if constexpr (C != B200_NONE) {
                if constexpr (S != B200_F32) {
#pragma unroll
                    for (int i = 0; i < kVec; i++) p[i] = round_through<S>(p[i]);
                }
                store8<C>(r.ptr[kCopyCol], idx[u], p);
            }
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, bool&) {
        float p = load1<S>(r.ptr[1], i), m = load1<S>(r.ptr[2], i);
        one(load1<G>(r.ptr[0], i), p, m, c);
        store1<S>(r.ptr[1], i, p);
        store1<S>(r.ptr[2], i, m);
        if constexpr (C != B200_NONE) store1<C>(r.ptr[kCopyCol], i, round_through<S>(p));
    }
    static __device__ __forceinline__ void finish(const Params&, bool) {}
};

// ============================================================================================
// fused unscale + Adam + master->model copy
// ============================================================================================
struct AdamConsts {
    float w1, one_m_w1, b2, w2, neg_step, bc2_sqrt, eps, wd, decay;
    uint32_t flags;     // bit0 weight decay on, bit1 decoupled (AdamW), bit2 lerp "high weight" form
};

struct AdamParams {
    AdamConsts grp[B200_MAX_GROUPS][B200_STEP_WINDOW];   // [group][step-window candidate]
    float inv_scale;
    const float* dev_inv_scale;
    const int32_t* dev_found_inf;
    StepSync sync;
    PeerPull pull;
};

// One element, torch.optim.Adam's operation order (see oracle_adam for the line references).
__device__ __forceinline__ void adam_elem(float g_in, float& p, float& m, float& v, const AdamConsts& k, float s) {
    float g = __fmul_rn(g_in, s);
    if (k.flags & 1u) {
        if (k.flags & 2u) p = __fmul_rn(p, k.decay);
        else g = __fadd_rn(g, __fmul_rn(k.wd, p));
    }
    const float d = __fsub_rn(g, m);
    if (!(k.flags & 4u)) m = __fadd_rn(m, __fmul_rn(k.w1, d));
    else m = __fsub_rn(g, __fmul_rn(d, k.one_m_w1));
    v = __fadd_rn(__fmul_rn(v, k.b2), __fmul_rn(__fmul_rn(k.w2, g), g));
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), k.bc2_sqrt), k.eps);
    p = __fadd_rn(p, __fmul_rn(k.neg_step, __fdiv_rn(m, denom)));
}

// 16-bit state (amp O3: p, m, v and g all fp16/bf16) moves half the bytes per element, so a thread keeps
// twice as many 8-element groups in flight and the table uses 8192-element tiles (tables.auto_tile_elems)
#ifndef B200_U16_MUL
#define B200_U16_MUL 2
#endif

template <int G, int S, int C>      // grad, state (p/m/v) and copy element types; C may be B200_NONE
struct AdamOp {
    using Params = AdamParams;
    static constexpr int kUnrollMul = (S != B200_F32 && G != B200_F32) ? B200_U16_MUL : 1;
    struct Ctx { AdamConsts k; float s; const void* g; };     // g: the tile's gradient column (local or a peer's)
    static constexpr int kCopyCol = 4;
    static __device__ __forceinline__ uint32_t rotation(const Params& p) { return p.pull.first_tile; }
    static __device__ __forceinline__ bool prologue(const Params& p, int& win) {
        const bool skip = step_prologue(p.dev_found_inf, p.sync, win);
        if (skip) step_epilogue(p.sync, true);
        return skip;
    }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row& r, int win, int64_t start) {
        using GT = typename Elem<G>::type;
        const char* g0 = reinterpret_cast<const char*>(r.ptr[0]);
        const long long d = peer_delta(p.pull, reinterpret_cast<const GT*>(g0) + start);
        return Ctx{p.grp[r.group][win], p.dev_inv_scale ? __ldg(p.dev_inv_scale) : p.inv_scale, g0 + d};
    }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], bool&) {
        Raw8<G> rg[U];
        Raw8<S> rp[U], rm[U], rv[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (act[u]) {
                load8<G, true>(c.g, idx[u], rg[u]);
                load8<S, false>(r.ptr[1], idx[u], rp[u]);
                load8<S, false>(r.ptr[2], idx[u], rm[u]);
                load8<S, false>(r.ptr[3], idx[u], rv[u]);
            }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            float g[kVec], p[kVec], m[kVec], v[kVec];
            unpack8<G>(rg[u], g);
            unpack8<S>(rp[u], p);
            unpack8<S>(rm[u], m);
            unpack8<S>(rv[u], v);
#pragma unroll
            for (int i = 0; i < kVec; i++) adam_elem(g[i], p[i], m[i], v[i], c.k, c.s);
            store8<S>(r.ptr[1], idx[u], p);
            store8<S>(r.ptr[2], idx[u], m);
            store8<S>(r.ptr[3], idx[u], v);
            if constexpr (C != B200_NONE) {
                if constexpr (S != B200_F32) {
#pragma unroll
                    for (int i = 0; i < kVec; i++) p[i] = round_through<S>(p[i]);
                }
                store8<C>(r.ptr[kCopyCol], idx[u], p);
            }
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, bool&) {
        float p = load1<S>(r.ptr[1], i), m = load1<S>(r.ptr[2], i), v = load1<S>(r.ptr[3], i);
        adam_elem(load1<G>(c.g, i), p, m, v, c.k, c.s);
        store1<S>(r.ptr[1], i, p);
        store1<S>(r.ptr[2], i, m);
        store1<S>(r.ptr[3], i, v);
        if constexpr (C != B200_NONE) store1<C>(r.ptr[kCopyCol], i, round_through<S>(p));
    }
    static __device__ __forceinline__ void finish(const Params& p, bool) { step_epilogue(p.sync, false); }
};


// ============================================================================================
// speculative single-pass step (no separate overflow-check pass)
// ============================================================================================
// The all-or-nothing skip normally costs a pass of its own: every gradient has to be inspected
// before the first state element may be overwritten.  Here p, m and v are double-buffered — every
// state tensor is followed by a shadow copy of itself (b200_spec_shadow_delta bytes further on) —
// and ONE pass reads the gradient and the CURRENT copies, inspects the gradient, and writes the
// new state into the OTHER copies and the new fp16 weights into the model.  A one-thread commit
// kernel then runs the loss-scaler update and flips the device-side parity word only if nothing
// overflowed; after an overflow (rare: the scaler backs off) the parity stays, the state is
// untouched by construction, and a repair pass re-derives the fp16 weights from the current masters.
struct SpecParams {
    AdamConsts grp[B200_MAX_GROUPS][B200_STEP_WINDOW];
    float inv_scale;
    const float* dev_inv_scale;
    int32_t* flag;                  // raised on a non-finite gradient (the loss scaler's overflow_acc)
    const int32_t* parity;          // 0: the primary copies are current, 1: the shadows are
    StepSync sync;
    PeerPull pull;
};

__device__ __forceinline__ int64_t spec_delta(int64_t numel, int elem_size) {
    return (numel * elem_size + 255) / 256 * 256;
}

template <int G, int S, int C>
struct AdamSpecOp {
    using Params = SpecParams;
    static constexpr int kUnrollMul = (S != B200_F32 && G != B200_F32) ? B200_U16_MUL : 1;
    struct Ctx { AdamConsts k; float s; const void* g; int64_t rd, wr; };
    static constexpr int kCopyCol = 4;
    static __device__ __forceinline__ uint32_t rotation(const Params& p) { return p.pull.first_tile; }
    static __device__ __forceinline__ bool prologue(const Params& p, int& win) {
        step_prologue(nullptr, p.sync, win);
        return false;                                   // never skipped up front: this pass IS the check
    }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row& r, int win, int64_t start) {
        using GT = typename Elem<G>::type;
        using ST = typename Elem<S>::type;
        const char* g0 = reinterpret_cast<const char*>(r.ptr[0]);
        const long long d = peer_delta(p.pull, reinterpret_cast<const GT*>(g0) + start);
        const int cur = *reinterpret_cast<const volatile int32_t*>(p.parity) & 1;
        const int64_t delta = spec_delta(r.numel, sizeof(ST)) / static_cast<int64_t>(sizeof(ST));   // in elements
        return Ctx{p.grp[r.group][win], p.dev_inv_scale ? __ldg(p.dev_inv_scale) : p.inv_scale, g0 + d,
                   cur ? delta : 0, cur ? 0 : delta};
    }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], bool& bad) {
        Raw8<G> rg[U];
        Raw8<S> rp[U], rm[U], rv[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (act[u]) {
                load8<G, true>(c.g, idx[u], rg[u]);
                load8<S, true>(r.ptr[1], idx[u] + c.rd, rp[u]);     // read-only here: the other copy is written
                load8<S, true>(r.ptr[2], idx[u] + c.rd, rm[u]);
                load8<S, true>(r.ptr[3], idx[u] + c.rd, rv[u]);
            }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            float g[kVec], p[kVec], m[kVec], v[kVec];
            unpack8<G>(rg[u], g);
            unpack8<S>(rp[u], p);
            unpack8<S>(rm[u], m);
            unpack8<S>(rv[u], v);
#pragma unroll
            for (int i = 0; i < kVec; i++) {
                bad |= !is_finite(g[i]);
                adam_elem(g[i], p[i], m[i], v[i], c.k, c.s);
            }
            store8<S>(r.ptr[1], idx[u] + c.wr, p);
            store8<S>(r.ptr[2], idx[u] + c.wr, m);
            store8<S>(r.ptr[3], idx[u] + c.wr, v);
            if constexpr (C != B200_NONE) {
                if constexpr (S != B200_F32) {
#pragma unroll
                    for (int i = 0; i < kVec; i++) p[i] = round_through<S>(p[i]);
                }
                store8<C>(r.ptr[kCopyCol], idx[u], p);
            }
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, bool& bad) {
        float p = load1<S>(r.ptr[1], i + c.rd), m = load1<S>(r.ptr[2], i + c.rd), v = load1<S>(r.ptr[3], i + c.rd);
        const float g = load1<G>(c.g, i);
        bad |= !is_finite(g);
        adam_elem(g, p, m, v, c.k, c.s);
        store1<S>(r.ptr[1], i + c.wr, p);
        store1<S>(r.ptr[2], i + c.wr, m);
        store1<S>(r.ptr[3], i + c.wr, v);
        if constexpr (C != B200_NONE) store1<C>(r.ptr[kCopyCol], i, round_through<S>(p));
    }
    static __device__ __forceinline__ void finish(const Params& p, bool bad) {
        if (p.flag) raise_flag(p.flag, bad);
    }
};

// repair after a skipped speculative step: model copy = cast(current master); exits at once otherwise
struct RepairParams {
    const int32_t* found_inf;
    const int32_t* parity;
};

template <int S, int C>
struct SpecRepairOp {
    using Params = RepairParams;
    static constexpr int kUnrollMul = 2;
    struct Ctx { int64_t rd; };
    static __device__ __forceinline__ uint32_t rotation(const Params&) { return 0u; }
    static __device__ __forceinline__ bool prologue(const Params& p, int&) {
        return *reinterpret_cast<const volatile int32_t*>(p.found_inf) == 0;
    }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row& r, int, int64_t) {
        using ST = typename Elem<S>::type;
        const int cur = *reinterpret_cast<const volatile int32_t*>(p.parity) & 1;
        return Ctx{cur ? spec_delta(r.numel, sizeof(ST)) / static_cast<int64_t>(sizeof(ST)) : 0};
    }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], bool&) {
        Raw8<S> in[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (act[u]) load8<S, true>(r.ptr[1], idx[u] + c.rd, in[u]);
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            float x[kVec];
            unpack8<S>(in[u], x);
            store8<C>(r.ptr[4], idx[u], x);
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, bool&) {
        store1<C>(r.ptr[4], i, load1<S>(r.ptr[1], i + c.rd));
    }
    static __device__ __forceinline__ void finish(const Params&, bool) {}
};

// ============================================================================================
// fused unscale + SGD + master->model copy
// ============================================================================================
struct SgdConsts {
    float wd, mu, damp, neg_lr;
    uint32_t flags;     // bit0 wd on, bit1 momentum on, bit2 nesterov, bit3 first_run, bit4 wd_after_momentum
};

struct SgdParams {
    SgdConsts grp[B200_MAX_GROUPS][B200_STEP_WINDOW];
    float inv_scale;
    const float* dev_inv_scale;
    const int32_t* dev_found_inf;
    StepSync sync;
    PeerPull pull;
};

// torch.optim.SGD's order plus apex FusedSGD's extras (see oracle_sgd for the line references).
template <int S>
__device__ __forceinline__ void sgd_elem(float g_in, float& p, float& mom, const SgdConsts& k, float s) {
    float g = __fmul_rn(g_in, s);
    const bool wd_on = k.flags & 1u, wd_after = k.flags & 16u;
    if (wd_on && !wd_after) g = __fadd_rn(g, __fmul_rn(k.wd, p));
    if (k.flags & 2u) {
        float b = (k.flags & 8u) ? g : __fadd_rn(__fmul_rn(mom, k.mu), __fmul_rn(k.damp, g));
        mom = b;
        b = round_through<S>(b);
        g = (k.flags & 4u) ? __fadd_rn(g, __fmul_rn(k.mu, b)) : b;
    }
    if (wd_on && wd_after) g = __fadd_rn(g, __fmul_rn(k.wd, p));
    p = __fadd_rn(p, __fmul_rn(k.neg_lr, g));
}

template <int G, int S, int C>
struct SgdOp {
    using Params = SgdParams;
    static constexpr int kUnrollMul = (S != B200_F32 && G != B200_F32) ? B200_U16_MUL : 1;
    // measured (profiles/r01_tune_sweep_*.txt): 2 resident CTAs/SM stream at 6.5-6.7 TB/s, the 3 the
    // register budget would allow only at 6.0 TB/s (more concurrent streams per DRAM page)
    static constexpr int kMaxCtasPerSm = 2;
    struct Ctx { SgdConsts k; float s; const void* g; };
    static constexpr int kCopyCol = 3;
    static __device__ __forceinline__ uint32_t rotation(const Params& p) { return p.pull.first_tile; }
    static __device__ __forceinline__ bool prologue(const Params& p, int& win) {
        const bool skip = step_prologue(p.dev_found_inf, p.sync, win);
        if (skip) step_epilogue(p.sync, true);
        return skip;
    }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row& r, int win, int64_t start) {
        using GT = typename Elem<G>::type;
        const char* g0 = reinterpret_cast<const char*>(r.ptr[0]);
        const long long d = peer_delta(p.pull, reinterpret_cast<const GT*>(g0) + start);
        return Ctx{p.grp[r.group][win], p.dev_inv_scale ? __ldg(p.dev_inv_scale) : p.inv_scale, g0 + d};
    }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], bool&) {
        const bool has_mom = c.k.flags & 2u;
        const bool read_mom = has_mom && !(c.k.flags & 8u);
        Raw8<G> rg[U];
        Raw8<S> rp[U], rm[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (act[u]) {
                load8<G, true>(c.g, idx[u], rg[u]);
                load8<S, false>(r.ptr[1], idx[u], rp[u]);
                if (read_mom) load8<S, false>(r.ptr[2], idx[u], rm[u]);
            }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            float g[kVec], p[kVec], m[kVec];
            unpack8<G>(rg[u], g);
            unpack8<S>(rp[u], p);
            if (read_mom) unpack8<S>(rm[u], m);
            else {
#pragma unroll
                for (int i = 0; i < kVec; i++) m[i] = 0.0f;
            }
#pragma unroll
            for (int i = 0; i < kVec; i++) sgd_elem<S>(g[i], p[i], m[i], c.k, c.s);
            store8<S>(r.ptr[1], idx[u], p);
            if (has_mom) store8<S>(r.ptr[2], idx[u], m);
            if constexpr (C != B200_NONE) {
                if constexpr (S != B200_F32) {
#pragma unroll
                    for (int i = 0; i < kVec; i++) p[i] = round_through<S>(p[i]);
                }
                store8<C>(r.ptr[kCopyCol], idx[u], p);
            }
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, bool&) {
        const bool has_mom = c.k.flags & 2u;
        float p = load1<S>(r.ptr[1], i);
        float m = (has_mom && !(c.k.flags & 8u)) ? load1<S>(r.ptr[2], i) : 0.0f;
        sgd_elem<S>(load1<G>(c.g, i), p, m, c.k, c.s);
        store1<S>(r.ptr[1], i, p);
        if (has_mom) store1<S>(r.ptr[2], i, m);
        if constexpr (C != B200_NONE) store1<C>(r.ptr[kCopyCol], i, round_through<S>(p));
    }
    static __device__ __forceinline__ void finish(const Params& p, bool) { step_epilogue(p.sync, false); }
};

// ============================================================================================
// DDP bucket epilogue (flat buffer, in place)
// ============================================================================================
template <int DT, int kThreads>
__global__ void __launch_bounds__(kThreads)
bucket_epilogue_kernel(void* __restrict__ buf, int64_t n, float factor, int read_only, int32_t* flag) {
    using T = typename Elem<DT>::type;
    constexpr int kU = 4;
    constexpr int kAlign = (DT == B200_F32) ? 32 : 16;
    bool bad = false;
    T* x = reinterpret_cast<T*>(buf);
    // ragged head up to the first aligned element
    int64_t head = ((kAlign - (reinterpret_cast<uintptr_t>(x) & (kAlign - 1))) & (kAlign - 1)) / sizeof(T);
    if (head > n) head = n;
    const int64_t gtid = static_cast<int64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (gtid < head) {
        const float r = __fmul_rn(Elem<DT>::to_float(x[gtid]), factor);
        const float rr = round_through<DT>(r);
        bad |= !is_finite(rr);
        if (!read_only) x[gtid] = Elem<DT>::from_float(r);
    }
    T* body = x + head;
    const int64_t nb = n - head;
    const int64_t n_groups = nb / kVec;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kThreads;
    for (int64_t g0 = gtid; g0 < n_groups; g0 += stride * kU) {
        Raw8<DT> raw[kU];
        bool act[kU];
#pragma unroll
        for (int u = 0; u < kU; u++) {
            const int64_t g = g0 + u * stride;
            act[u] = g < n_groups;
            if (act[u]) load8<DT, false>(body, g * kVec, raw[u]);
        }
#pragma unroll
        for (int u = 0; u < kU; u++) {
            if (!act[u]) continue;
            float f[kVec];
            unpack8<DT>(raw[u], f);
#pragma unroll
            for (int i = 0; i < kVec; i++) {
                f[i] = __fmul_rn(f[i], factor);
                bad |= !is_finite(round_through<DT>(f[i]));
            }
            if (!read_only) store8<DT>(body, (g0 + u * stride) * kVec, f);
        }
    }
    const int64_t tail0 = n_groups * kVec;
    if (gtid < nb - tail0) {
        const float r = __fmul_rn(Elem<DT>::to_float(body[tail0 + gtid]), factor);
        bad |= !is_finite(round_through<DT>(r));
        if (!read_only) body[tail0 + gtid] = Elem<DT>::from_float(r);
    }
    if (flag) raise_flag(flag, bad);
}

// ============================================================================================
// reduce-scatter + mean + overflow check over peer memory (one shard per rank)
// ============================================================================================
struct PeerBufs {
    const void* p[B200_MAX_PEERS];
};

// value of one reduced element: fp32 sum in rank order, one rounding to the gradient dtype, then
// x factor with a second rounding (oracle_peer_reduce in oracle/amp_step_oracle.c)
template <int DT, int kWorld>
__device__ __forceinline__ float peer_mean(const float (&x)[kWorld], int world, float factor) {
    float acc = x[0];
#pragma unroll
    for (int r = 1; r < kWorld; r++)
        if (r < world) acc = __fadd_rn(acc, x[r]);
    return __fmul_rn(round_through<DT>(acc), factor);
}

// kWorld = compile-time upper bound on the rank count (2, 4 or 8): the loads of every rank and of
// kU groups are issued before the first use, so a thread keeps kWorld*kU 128/256-bit requests in
// flight — NVLink round trips are 2-3x an HBM access, and this kernel is bound by them.
// Cross-GPU hand-shake through signal pads in peer memory (b200_peer_sync).  Every rank owns one
// pad; peers write into it with system-scope release stores and the owner polls it with acquire
// loads.  Epochs only grow, so nothing has to be cleared between steps.
struct PeerSync {
    b200_peer_signal_pad* pad[B200_MAX_PEERS];      // pad[r]: rank r's pad as addressable from this GPU; NULL = no sync
    int world, rank, epoch;
};

__device__ __forceinline__ void st_release_sys(int32_t* p, int32_t v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int32_t ld_acquire_sys(const int32_t* p) {
    int32_t v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Spin until *p >= want (as seen with acquire/system scope).  Bounded: after 6e10 cycles (~30 s; ranks can be
// seconds apart in a first iteration) the
// pad's error word is set and the wait gives up, so a missing peer shows up as an error, not a hung GPU.
__device__ __forceinline__ void wait_at_least(const int32_t* p, int32_t want, int32_t* error) {
    const long long t0 = clock64();
    while (ld_acquire_sys(p) < want) {
        __nanosleep(64);
        if (clock64() - t0 > 60000000000ll) {
            *reinterpret_cast<volatile int32_t*>(error) = 1;
            __threadfence();
            break;
        }
    }
}

// A wait that gave up poisons the step: the caller raises the overflow flag, so the optimizer kernel writes
// nothing (the same all-or-nothing path as a gradient overflow) instead of updating from incomplete sums.
__device__ __forceinline__ bool wait_failed(const int32_t* error) {
    return *reinterpret_cast<const volatile int32_t*>(error) != 0;
}

// Mean of the elements [begin, end) over all ranks' buffers, written to `reduced` at the same offsets; the
// participating threads are numbered gtid = 0 .. stride-1 (a whole grid, or the reducer CTAs of the fused step).
template <int DT, int kThreads, int kWorld>
__device__ __forceinline__ void reduce_span(const PeerBufs& bufs, int world, void* __restrict__ reduced, int64_t begin,
                                            int64_t end, float factor, int64_t gtid, int64_t stride, bool& bad) {
    constexpr int kU = (DT == B200_F32 ? 4 : 8) / kWorld > 0 ? (DT == B200_F32 ? 4 : 8) / kWorld : 1;
    using T = typename Elem<DT>::type;
    T* out = reinterpret_cast<T*>(reduced);
    // every buffer starts 32-byte aligned, so element offsets that are multiples of 8 are vector-safe
    int64_t head = (kVec - (begin % kVec)) % kVec;
    if (head > end - begin) head = end - begin;
    if (gtid < head) {
        float x[kWorld];
#pragma unroll
        for (int r = 0; r < kWorld; r++) x[r] = r < world ? load1<DT>(bufs.p[r], begin + gtid) : 0.0f;
        const float v = peer_mean<DT, kWorld>(x, world, factor);
        bad |= !is_finite(round_through<DT>(v));
        store1<DT>(out, begin + gtid, v);
    }
    const int64_t body = begin + head;
    const int64_t n_groups = (end - body) / kVec;
    for (int64_t g0 = gtid; g0 < n_groups; g0 += stride * kU) {
        Raw8<DT> raw[kU][kWorld];
        bool act[kU];
#pragma unroll
        for (int u = 0; u < kU; u++) {
            const int64_t g = g0 + u * stride;
            act[u] = g < n_groups;
#pragma unroll
            for (int r = 0; r < kWorld; r++)
                if (act[u] && r < world) load8<DT, false>(bufs.p[r], body + g * kVec, raw[u][r]);
        }
#pragma unroll
        for (int u = 0; u < kU; u++) {
            if (!act[u]) continue;
            float acc[kVec];
            unpack8<DT>(raw[u][0], acc);
#pragma unroll
            for (int r = 1; r < kWorld; r++) {
                if (r < world) {
                    float f[kVec];
                    unpack8<DT>(raw[u][r], f);
#pragma unroll
                    for (int i = 0; i < kVec; i++) acc[i] = __fadd_rn(acc[i], f[i]);
                }
            }
#pragma unroll
            for (int i = 0; i < kVec; i++) {
                acc[i] = __fmul_rn(round_through<DT>(acc[i]), factor);
                bad |= !is_finite(round_through<DT>(acc[i]));
            }
            store8<DT>(out, body + (g0 + u * stride) * kVec, acc);
        }
    }
    const int64_t tail0 = body + n_groups * kVec;
    if (gtid < end - tail0) {
        float x[kWorld];
#pragma unroll
        for (int r = 0; r < kWorld; r++) x[r] = r < world ? load1<DT>(bufs.p[r], tail0 + gtid) : 0.0f;
        const float v = peer_mean<DT, kWorld>(x, world, factor);
        bad |= !is_finite(round_through<DT>(v));
        store1<DT>(out, tail0 + gtid, v);
    }
}

template <int DT, int kThreads, int kWorld>
__global__ void __launch_bounds__(kThreads)
peer_reduce_kernel(const __grid_constant__ PeerBufs bufs, int world, void* __restrict__ reduced, int64_t begin,
                   int64_t end, float factor, int32_t* flag, const __grid_constant__ PeerSync sync) {
    bool bad = false;
    const bool synced = sync.pad[0] != nullptr;
    if (synced) {
        // head: tell every rank "my gradients of this epoch are in place (and I am done with last epoch's
        // reduced buffers)", then wait until every rank has said so before reading anybody's gradients
        b200_peer_signal_pad* mine = sync.pad[sync.rank];
        if (blockIdx.x == 0 && threadIdx.x < sync.world) st_release_sys(&sync.pad[threadIdx.x]->ready[sync.rank], sync.epoch);
        if (threadIdx.x < sync.world) wait_at_least(&mine->ready[threadIdx.x], sync.epoch, &mine->error);
        __syncthreads();
        if (wait_failed(&mine->error)) {      // a rank never arrived: read nobody's buffer, poison the step
            bad = true;
            end = begin;
        }
    }
    reduce_span<DT, kThreads, kWorld>(bufs, world, reduced, begin, end, factor,
                                      static_cast<int64_t>(blockIdx.x) * kThreads + threadIdx.x,
                                      static_cast<int64_t>(gridDim.x) * kThreads, bad);
    if (!synced) {
        if (flag) raise_flag(flag, bad);
        return;
    }
    // tail: the last block to finish publishes "my shard of this epoch is written" together with this
    // rank's overflow bit to every rank's pad; b200_peer_wait_flags (next on the stream) collects them
    b200_peer_signal_pad* mine = sync.pad[sync.rank];
    raise_flag(&mine->local_flag, bad);
    __shared__ bool s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        s_last = atomicAdd(reinterpret_cast<unsigned*>(&mine->ticket), 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
        __threadfence_system();
        const int32_t bit = *reinterpret_cast<volatile int32_t*>(&mine->local_flag) != 0 ? 1 : 0;
        if (threadIdx.x < sync.world) st_release_sys(&sync.pad[threadIdx.x]->done[sync.rank], (sync.epoch << 1) | bit);
        __syncthreads();
        if (threadIdx.x == 0) {
            mine->local_flag = 0;
            mine->ticket = 0;
        }
    }
}

// Waits until every rank has published its shard of `epoch`, then ORs the ranks' overflow bits into *flag.
__global__ void peer_wait_flags_kernel(const __grid_constant__ PeerSync sync, int32_t* flag) {
    b200_peer_signal_pad* mine = sync.pad[sync.rank];
    __shared__ int s_any;
    if (threadIdx.x == 0) s_any = 0;
    __syncthreads();
    if (threadIdx.x < sync.world) {
        wait_at_least(&mine->done[threadIdx.x], sync.epoch << 1, &mine->error);
        if ((ld_acquire_sys(&mine->done[threadIdx.x]) & 1) || wait_failed(&mine->error)) atomicOr(&s_any, 1);
    }
    __syncthreads();
    if (threadIdx.x == 0 && s_any && flag) *flag = 1;
}

// ============================================================================================
// sharded step (ZeRO-1 style): reduce-scatter + speculative Adam on the own shard + all-gather, ONE pass
// ============================================================================================
// Each rank owns the fp32 master / m / v of its shard only.  For every 8-element group of its shard a thread
// reads the gradients of ALL ranks (its own from HBM, the others over NVLink), averages them in rank order
// exactly like peer_reduce_kernel (fp32 sum -> round to the gradient type -> x factor -> round), inspects the
// mean, runs the speculative Adam update into the other copy of p / m / v, and STORES the new 16-bit weights
// into the model buffer of every rank (posted writes over NVLink).  Tiles are independent of each other —
// no chunk flags, no reducer/updater roles — and the inbound (gradients) and outbound (weights) directions of the
// link are busy at the same time.  Per GPU and step: HBM 28 B x P/N + 2 x 2 B x P(N-1)/N served to / received
// from peers; NVLink 2 B x P(N-1)/N each way.
// Hand-shake: head = "my gradients of this epoch are in place and I no longer read my model weights" (ready);
// tail = last CTA publishes "my reads of your gradients are over, my writes to your model have landed" + this
// rank's overflow bit (done).  b200_peer_wait_flags then ORs the bits into the scaler's overflow word and
// b200_spec_commit flips the parity unless any rank overflowed; b200_peer_shard_repair re-sends the old weights
// after an overflow (rare), and exits at once otherwise.
struct PeerFan {
    long long grad_delta[B200_MAX_PEERS];    // bytes from this rank's gradient buffer to rank r's (0 for itself)
    long long copy_delta[B200_MAX_PEERS];    // bytes from this rank's model buffer to rank r's
    int world;
    int store_world;                         // = world (a debug build of the host side may lower it)
    float factor;
    int use_mc;                              // 1: 16-byte groups go out as ONE multicast store (copy_mc_delta)
    long long copy_mc_delta;                 // bytes from this rank's model buffer to the multicast address of the model buffers
    int use_mc_grad;                         // 1: 16-byte groups of gradients are summed inside the switch (grad_mc_delta)
    long long grad_mc_delta;
};

struct ShardStepParams {
    AdamConsts grp[B200_MAX_GROUPS][B200_STEP_WINDOW];
    float inv_scale;
    const float* dev_inv_scale;
    const int32_t* parity;
    int32_t* flag;                           // used when there is no hand-shake (emulated ranks)
    StepSync sync;
    PeerFan fan;
    PeerSync psync;
};

// tail of a kernel that takes part in the hand-shake: the last CTA to finish publishes done[rank] everywhere
__device__ __forceinline__ void publish_done(const PeerSync& sync, bool bad) {
    b200_peer_signal_pad* mine = sync.pad[sync.rank];
    raise_flag(&mine->local_flag, bad);
    __shared__ bool s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        s_last = atomicAdd(reinterpret_cast<unsigned*>(&mine->ticket), 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
        __threadfence_system();
        const int32_t bit = *reinterpret_cast<volatile int32_t*>(&mine->local_flag) != 0 ? 1 : 0;
        if (threadIdx.x < sync.world) st_release_sys(&sync.pad[threadIdx.x]->done[sync.rank], (sync.epoch << 1) | bit);
        __syncthreads();
        if (threadIdx.x == 0) {
            mine->local_flag = 0;
            mine->ticket = 0;
        }
    }
}

template <int G, int S, int C, int kWorld>
struct ShardStepOp {
    using Params = ShardStepParams;
    static constexpr int kUnrollMul = 1;
    struct Ctx { AdamConsts k; float s; const char* g; char* c; int64_t rd, wr; const PeerFan* fan; };
    struct Acc { bool bad; };
    static __device__ __forceinline__ uint32_t rotation(const Params&) { return 0u; }
    static __device__ __forceinline__ bool prologue(const Params& p, int& win) {
        step_prologue(nullptr, p.sync, win);
        if (p.psync.pad[0] != nullptr) {
            b200_peer_signal_pad* mine = p.psync.pad[p.psync.rank];
            if (blockIdx.x == 0 && threadIdx.x < p.psync.world)
                st_release_sys(&p.psync.pad[threadIdx.x]->ready[p.psync.rank], p.psync.epoch);
            if (threadIdx.x < p.psync.world) wait_at_least(&mine->ready[threadIdx.x], p.psync.epoch, &mine->error);
            __syncthreads();
            if (wait_failed(&mine->error)) {             // a rank never arrived: touch nobody's memory, poison the step
                publish_done(p.psync, true);
                return true;
            }
        }
        return false;
    }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row& r, int win, int64_t) {
        using ST = typename Elem<S>::type;
        const int cur = *reinterpret_cast<const volatile int32_t*>(p.parity) & 1;
        const int64_t delta = spec_delta(r.numel, sizeof(ST)) / static_cast<int64_t>(sizeof(ST));
        return Ctx{p.grp[r.group][win], p.dev_inv_scale ? __ldg(p.dev_inv_scale) : p.inv_scale,
                   reinterpret_cast<const char*>(r.ptr[0]), reinterpret_cast<char*>(r.ptr[4]),
                   cur ? delta : 0, cur ? 0 : delta, &p.fan};
    }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], Acc& acc) {
        const int world = c.fan->world;
        Raw8<G> rg[U][kWorld];
        Raw8<S> rp[U], rm[U], rv[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (act[u]) {
                if (c.fan->use_mc_grad) {
                    multimem_ld_reduce128<G>(c.g + c.fan->grad_mc_delta + idx[u] * 2, rg[u][0].r);
                } else {
#pragma unroll
                    for (int w = 0; w < kWorld; w++)
                        if (w < world) load8<G, false>(c.g + c.fan->grad_delta[w], idx[u], rg[u][w]);
                }
                load8<S, true>(r.ptr[1], idx[u] + c.rd, rp[u]);
                load8<S, true>(r.ptr[2], idx[u] + c.rd, rm[u]);
                load8<S, true>(r.ptr[3], idx[u] + c.rd, rv[u]);
            }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            float g[kVec], p[kVec], m[kVec], v[kVec];
            unpack8<G>(rg[u][0], g);
#pragma unroll
            for (int w = 1; w < kWorld; w++)
                if (w < world && !c.fan->use_mc_grad) {
                    float f[kVec];
                    unpack8<G>(rg[u][w], f);
#pragma unroll
                    for (int i = 0; i < kVec; i++) g[i] = __fadd_rn(g[i], f[i]);
                }
            unpack8<S>(rp[u], p);
            unpack8<S>(rm[u], m);
            unpack8<S>(rv[u], v);
#pragma unroll
            for (int i = 0; i < kVec; i++) {
                g[i] = round_through<G>(__fmul_rn(round_through<G>(g[i]), c.fan->factor));
                acc.bad |= !is_finite(g[i]);
            }
            store8<G>(const_cast<char*>(c.g), idx[u], g);      // the mean replaces this rank's own gradient (its shard only)
#pragma unroll
            for (int i = 0; i < kVec; i++) adam_elem(g[i], p[i], m[i], v[i], c.k, c.s);
            store8<S>(r.ptr[1], idx[u] + c.wr, p);
            store8<S>(r.ptr[2], idx[u] + c.wr, m);
            store8<S>(r.ptr[3], idx[u] + c.wr, v);
            if (c.fan->use_mc) {
                uint32_t packed[4];
                pack8_16<C>(p, packed);
                multimem_st128(c.c + c.fan->copy_mc_delta + idx[u] * 2, packed);
            } else {
#pragma unroll
                for (int w = 0; w < kWorld; w++)
                    if (w < c.fan->store_world) store8<C>(c.c + c.fan->copy_delta[w], idx[u], p);
            }
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, Acc& acc) {
        const int world = c.fan->world;
        float g = load1<G>(c.g + c.fan->grad_delta[0], i);
        for (int w = 1; w < world; w++) g = __fadd_rn(g, load1<G>(c.g + c.fan->grad_delta[w], i));
        g = round_through<G>(__fmul_rn(round_through<G>(g), c.fan->factor));
        acc.bad |= !is_finite(g);
        store1<G>(const_cast<char*>(c.g), i, g);
        float p = load1<S>(r.ptr[1], i + c.rd), m = load1<S>(r.ptr[2], i + c.rd), v = load1<S>(r.ptr[3], i + c.rd);
        adam_elem(g, p, m, v, c.k, c.s);
        store1<S>(r.ptr[1], i + c.wr, p);
        store1<S>(r.ptr[2], i + c.wr, m);
        store1<S>(r.ptr[3], i + c.wr, v);
        for (int w = 0; w < c.fan->store_world; w++) store1<C>(c.c + c.fan->copy_delta[w], i, p);
    }
    static __device__ __forceinline__ void finish(const Params& p, const Acc& acc) {
        if (p.psync.pad[0] != nullptr) publish_done(p.psync, acc.bad);
        else if (p.flag) raise_flag(p.flag, acc.bad);
    }
};

// after a skipped sharded step: every rank re-sends cast(current master) of its shard to every rank's model buffer
struct ShardRepairParams {
    const int32_t* found_inf;
    const int32_t* parity;
    PeerFan fan;
    PeerSync psync;
};

template <int S, int C, int kWorld>
struct ShardRepairOp {
    using Params = ShardRepairParams;
    static constexpr int kUnrollMul = 2;
    struct Ctx { int64_t rd; char* c; const PeerFan* fan; };
    static __device__ __forceinline__ uint32_t rotation(const Params&) { return 0u; }
    static __device__ __forceinline__ bool prologue(const Params& p, int&) {
        return *reinterpret_cast<const volatile int32_t*>(p.found_inf) == 0;     // every rank sees the same word
    }
    static __device__ __forceinline__ Ctx make_ctx(const Params& p, const b200_table_row& r, int, int64_t) {
        using ST = typename Elem<S>::type;
        const int cur = *reinterpret_cast<const volatile int32_t*>(p.parity) & 1;
        return Ctx{cur ? spec_delta(r.numel, sizeof(ST)) / static_cast<int64_t>(sizeof(ST)) : 0,
                   reinterpret_cast<char*>(r.ptr[4]), &p.fan};
    }
    template <int U>
    static __device__ __forceinline__ void vec(const b200_table_row& r, const Ctx& c, const int64_t (&idx)[U],
                                               const bool (&act)[U], bool&) {
        Raw8<S> in[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (act[u]) load8<S, true>(r.ptr[1], idx[u] + c.rd, in[u]);
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!act[u]) continue;
            float x[kVec];
            unpack8<S>(in[u], x);
            if (c.fan->use_mc) {
                uint32_t packed[4];
                pack8_16<C>(x, packed);
                multimem_st128(c.c + c.fan->copy_mc_delta + idx[u] * 2, packed);
            } else {
#pragma unroll
                for (int w = 0; w < kWorld; w++)
                    if (w < c.fan->store_world) store8<C>(c.c + c.fan->copy_delta[w], idx[u], x);
            }
        }
    }
    static __device__ __forceinline__ void scalar(const b200_table_row& r, const Ctx& c, int64_t i, bool&) {
        const float x = load1<S>(r.ptr[1], i + c.rd);
        for (int w = 0; w < c.fan->store_world; w++) store1<C>(c.c + c.fan->copy_delta[w], i, x);
    }
    static __device__ __forceinline__ void finish(const Params& p, bool) {
        if (p.psync.pad[0] != nullptr) publish_done(p.psync, false);
    }
};

// waits for every rank's done[] of `epoch` — only when *cond != 0 (the ranks agree on *cond)
__global__ void peer_wait_done_if_kernel(const __grid_constant__ PeerSync sync, const int32_t* cond) {
    if (*reinterpret_cast<const volatile int32_t*>(cond) == 0) return;
    b200_peer_signal_pad* mine = sync.pad[sync.rank];
    if (threadIdx.x < sync.world) wait_at_least(&mine->done[threadIdx.x], sync.epoch << 1, &mine->error);
}

// All-gather half as a kernel of its own (used where the averaged gradient must be materialised, i.e.
// inside apex.parallel.DistributedDataParallel): wait for every rank's shard, OR the overflow bits,
// then copy each 8-element group from its owner's `reduced` buffer into the local bucket.
struct ShardBounds {
    long long b[B200_MAX_PEERS + 1];
};

template <int DT, int kThreads>
__global__ void __launch_bounds__(kThreads)
peer_gather_kernel(const __grid_constant__ PeerBufs reduced, int world, void* __restrict__ local_out,
                   const __grid_constant__ ShardBounds bounds, int32_t* flag, const __grid_constant__ PeerSync sync) {
    using T = typename Elem<DT>::type;
    constexpr int kU = 4;
    if (sync.pad[0] != nullptr) {
        b200_peer_signal_pad* mine = sync.pad[sync.rank];
        __shared__ int s_any;
        if (threadIdx.x == 0) s_any = 0;
        __syncthreads();
        if (threadIdx.x < sync.world) {
            wait_at_least(&mine->done[threadIdx.x], sync.epoch << 1, &mine->error);
            if ((ld_acquire_sys(&mine->done[threadIdx.x]) & 1) || wait_failed(&mine->error)) atomicOr(&s_any, 1);
        }
        __syncthreads();
        if (blockIdx.x == 0 && threadIdx.x == 0 && s_any && flag) *flag = 1;
    }
    T* out = reinterpret_cast<T*>(local_out);
    const long long n = bounds.b[world];
    const long long n_groups = n / kVec;
    const long long gtid = static_cast<long long>(blockIdx.x) * kThreads + threadIdx.x;
    const long long stride = static_cast<long long>(gridDim.x) * kThreads;
    auto owner_of = [&](long long elem) {
        int o = 0;
#pragma unroll
        for (int k = 1; k < B200_MAX_PEERS; k++) o += (k < world && elem >= bounds.b[k]) ? 1 : 0;
        return o;
    };
    for (long long g0 = gtid; g0 < n_groups; g0 += stride * kU) {
        Raw8<DT> raw[kU];
        bool act[kU];
#pragma unroll
        for (int u = 0; u < kU; u++) {
            const long long g = g0 + u * stride;
            act[u] = g < n_groups;
            if (act[u]) load8<DT, false>(reduced.p[owner_of(g * kVec)], g * kVec, raw[u]);
        }
#pragma unroll
        for (int u = 0; u < kU; u++) {
            if (!act[u]) continue;
            T* dst = out + (g0 + u * stride) * kVec;
            if constexpr (DT == B200_F32) stg256(dst, raw[u].r);
            else stg128(dst, raw[u].r);
        }
    }
    const long long tail0 = n_groups * kVec;
    if (gtid < n - tail0) {
        const long long i = tail0 + gtid;
        out[i] = reinterpret_cast<const T*>(reduced.p[owner_of(i)])[i];
    }
}

// ============================================================================================
// fused gradient mean + speculative step over peer memory: ONE kernel per rank, no serial phase
// ============================================================================================
// With an in-place update the reduce-scatter has to finish on every rank before the first state element may
// be written (an overflow anywhere must stop every write everywhere), so NVLink time and HBM time add up.
// The speculative update writes only shadow state, so nothing has to wait, and every CTA does both jobs at
// once: per iteration it (1) starts the loads of one reduce PIECE (2048 elements of this rank's shard from
// every rank's gradient buffer) as cp.async copies into shared memory — no registers are held while they
// cross NVLink —, (2) runs the speculative Adam pass on one update tile, whose HBM traffic hides the NVLink
// round trip, and (3) folds the piece (mean + roundings exactly as b200_peer_reduce_shard) into this rank's
// `reduced` buffer.  Pieces and tiles are drawn from two shared counters.  When the last piece of a chunk is
// folded the chunk is announced on every rank's progress board (an array of epoch flags that follows the
// signal pad in peer memory); a tile is processed only after the owner of its gradient has announced the
// chunk it lies in, and a CTA blocks on that only when there are no pieces left to hand out (so producers
// can never be starved by waiting consumers).  The table lists the tiles chunk-major
// (b200_table_order_for_peer_step): consumers trail producers by construction.  Every rank inspects every
// reduced gradient it pulls, so each rank's overflow flag is complete on its own: no flag exchange.
constexpr int kPieceElems = 256 * kVec;
constexpr int kRunPieces = 64;                              // pieces a reducer CTA folds between two fences
constexpr long long kRunElems = static_cast<long long>(kRunPieces) * kPieceElems;

__device__ __forceinline__ int32_t* peer_board(b200_peer_signal_pad* pad) {
    return reinterpret_cast<int32_t*>(pad) + sizeof(b200_peer_signal_pad) / 4;
}

__device__ __forceinline__ bool epoch_reached(const int32_t* p, int32_t epoch) {
    return static_cast<int32_t>(static_cast<uint32_t>(ld_acquire_sys(p)) - static_cast<uint32_t>(epoch)) >= 0;
}

__device__ __forceinline__ void wait_epoch(const int32_t* p, int32_t epoch, int32_t* error) {
    const long long t0 = clock64();
    while (!epoch_reached(p, epoch)) {
        __nanosleep(32);
        if (clock64() - t0 > 60000000000ll) {
            *reinterpret_cast<volatile int32_t*>(error) = 1;
            __threadfence();
            break;
        }
    }
}

struct PeerStepParams {
    SpecParams spec;                // update half; spec.pull says where each shard's reduced gradients live
    PeerBufs grads;                 // reduce half: every rank's flat gradient buffer
    void* reduced;                  // this rank's reduced buffer
    long long begin, end;           // this rank's shard
    long long chunk_elems;          // announcement granularity, a multiple of kPieceElems
    float factor;
    int world;
    int n_reducers;                 // CTAs [0, n_reducers) reduce, the others update
    int n_stages;                   // pieces a reducer CTA keeps in flight (shared-memory stages)
    PeerSync sync;
    int32_t* scratch;               // [2] exits [3] run tickets [8 + c] pieces of chunk c folded
};

template <int DT>
__device__ __forceinline__ void lds8(const void* smem, Raw8<DT>& raw) {
    const uint4* q = reinterpret_cast<const uint4*>(smem);
    const uint4 a = q[0];
    raw.r[0] = a.x; raw.r[1] = a.y; raw.r[2] = a.z; raw.r[3] = a.w;
    if constexpr (DT == B200_F32) {
        const uint4 b = q[1];
        raw.r[4] = b.x; raw.r[5] = b.y; raw.r[6] = b.z; raw.r[7] = b.w;
    }
}

// One update tile of the fused step (the body of walk_tiles for one tile).  Kept inlined: ptxas 12.9 crashes on a
// __noinline__ version; the fused kernel instead keeps its scheduling state in shared memory so that the Adam
// loop has the registers to itself.
template <class Op, int kThreads, int kU>
__device__ __forceinline__ bool step_one_tile(const typename Op::Params& sp, const b200_table_row* row_in, uint32_t tile_idx,
                                           uint32_t tile_elems, int win) {
    bool bad = false;
    const b200_table_row row = *row_in;
    const int64_t start = static_cast<int64_t>(tile_idx) * tile_elems;
    const int64_t left = row.numel - start;
    const int n = left < static_cast<int64_t>(tile_elems) ? static_cast<int>(left) : static_cast<int>(tile_elems);
    const typename Op::Ctx ctx = Op::make_ctx(sp, row, win, start);
    if (row.flags & 1u) {
        const int n8 = n & ~(kVec - 1);
        for (int base = threadIdx.x * kVec; base < n8; base += kThreads * kVec * kU) {
            int64_t idx[kU];
            bool act[kU];
#pragma unroll
            for (int u = 0; u < kU; u++) {
                const int off = base + u * kThreads * kVec;
                act[u] = off < n8;
                idx[u] = start + off;
            }
            Op::template vec<kU>(row, ctx, idx, act, bad);
        }
        const int tail = n - n8;
        if (static_cast<int>(threadIdx.x) < tail) Op::scalar(row, ctx, start + n8 + threadIdx.x, bad);
    } else {
        for (int i = threadIdx.x; i < n; i += kThreads) Op::scalar(row, ctx, start + i, bad);
    }
    return bad;
}

struct StepShared {                     // the fused step's scheduling state (shared memory, not registers)
    const b200_table_row* rows;
    const b200_table_tile* tiles;
    uint32_t n_tiles, tile_elems, n_pieces;
    int vec_ok, win;
    uint32_t idx[2];                    // tile (index into the table's list) of the iteration using the slot
    uint32_t tile[2];                   // ... its tile index inside its tensor
    uint32_t pc[2];                     // reduce piece that iteration starts and folds
    int go[2];                          // 1: the tile's gradient has been announced, process it
    uint32_t idx1;                      // list index of the next iteration's tile
    uint2 e1[2];                        // e1[s]: (tensor, tile) entry of the tile that goes into slot s next (8-byte cp.async)
    uint32_t piece_next;                // piece of the next iteration
    int seen_owner;
    long long seen_lo, seen_hi;         // chunks of seen_owner already seen announced
};

// Thread 0: may the tile in slot `sl` run?  Every chunk its gradient touches must have been announced by its owner.
template <class GT, int kWorld>
__device__ __forceinline__ int step_tile_ready(const PeerStepParams& prm, const b200_table_row* row, uint32_t tile_idx,
                                            uint32_t tile_elems, StepShared* sc, bool block) {
    b200_peer_signal_pad* mine = prm.sync.pad[prm.sync.rank];
    const char* g0 = reinterpret_cast<const char*>(row->ptr[0]);      // thread 0's own 16 bytes of the row have arrived
    const long long flat = (g0 - prm.spec.pull.local_base) / static_cast<long long>(sizeof(GT)) +
                           static_cast<long long>(tile_idx) * tile_elems;
    int owner = 0;
#pragma unroll
    for (int k = 1; k < kWorld; k++) owner += (k < prm.world && flat >= prm.spec.pull.bounds[k]) ? 1 : 0;
    long long last = flat + tile_elems;                                // the tile may be shorter (end of its tensor)
    if (last > prm.spec.pull.bounds[owner + 1]) last = prm.spec.pull.bounds[owner + 1];
    // pieces are folded by whichever CTA draws them, so chunks complete in any order: every chunk the tile touches
    // has to be announced (one or two: chunks are at least a tile long)
    const long long c0 = (flat - prm.spec.pull.bounds[owner]) / prm.chunk_elems;
    const long long c1 = (last - 1 - prm.spec.pull.bounds[owner]) / prm.chunk_elems;
    if (owner == sc->seen_owner && c0 >= sc->seen_lo && c1 <= sc->seen_hi) return 1;
    const int32_t* board = peer_board(mine);
    for (long long c = c0; c <= c1; c++) {
        const int32_t* f = board + owner * B200_PEER_STEP_MAX_CHUNKS + c;
        if (block) wait_epoch(f, prm.sync.epoch, &mine->error);
        if (!epoch_reached(f, prm.sync.epoch) && !wait_failed(&mine->error)) return 0;
    }
    sc->seen_owner = owner;
    sc->seen_lo = c0;
    sc->seen_hi = c1;
    return 1;
}

template <int G, int S, int C, int kThreads, int kWorld>
__global__ void __launch_bounds__(kThreads, 2)          // two CTAs per SM, like the plain step kernel: <= 128 registers
peer_spec_step_kernel(const void* __restrict__ table, const __grid_constant__ PeerStepParams prm) {
    static_assert(kThreads * kVec == kPieceElems, "one 8-element group per thread and piece");
    using Op = AdamSpecOp<G, S, C>;
    using GT = typename Elem<G>::type;
    constexpr int kU = 2 * Op::kUnrollMul;
    extern __shared__ __align__(128) unsigned char s_stage[];              // reducer stages: kWorld slices of one piece each
    __shared__ __align__(16) RowSlot s_slot[2];
    __shared__ StepShared s_sh;
    bool bad = false;

    // every rank's gradients of this epoch must be in place before anybody's buffer is read
    if (blockIdx.x == 0 && threadIdx.x < prm.sync.world)
        st_release_sys(&prm.sync.pad[threadIdx.x]->ready[prm.sync.rank], prm.sync.epoch);
    if (threadIdx.x < prm.sync.world)
        wait_at_least(&prm.sync.pad[prm.sync.rank]->ready[threadIdx.x], prm.sync.epoch, &prm.sync.pad[prm.sync.rank]->error);
    __syncthreads();
    const bool dead = wait_failed(&prm.sync.pad[prm.sync.rank]->error);     // a rank never arrived: read nobody's buffer
    bad |= dead;

    if (static_cast<int>(blockIdx.x) < prm.n_reducers) {
        // ================= REDUCE role: this rank's shard, one RUN of kRunPieces pieces at a time ==============
        // The slices of a piece (2048 elements from every rank) are staged in shared memory by cp.async, kStages
        // pieces deep, so a reducer CTA keeps kStages x world x 4 KB in flight without holding registers; a run is
        // counted (one fence for all its stores) when it is done, and the CTA that completes a chunk announces it.
        const int n_stage = prm.n_stages;
        const size_t stage_bytes = static_cast<size_t>(kWorld) * kPieceElems * sizeof(GT);      // one slice per rank
        const long long shard = prm.end - prm.begin;
        const long long n_runs = (shard + kRunElems - 1) / kRunElems;
        const bool vec_ok = (prm.begin % kVec) == 0;                  // cp.async needs 16-byte aligned sources
        __shared__ long long s_run;
        __shared__ __align__(8) uint64_t s_bar[12];                    // one mbarrier per stage
        if (threadIdx.x == 0) {
            for (int i = 0; i < n_stage; i++) mbar_init(&s_bar[i], 1);
            fence_barrier_init();
        }
        uint32_t phase = 0;                                            // bit s: parity the next wait on stage s expects
        constexpr unsigned kSliceBytes = kPieceElems * sizeof(GT);
        for (;;) {
            __syncthreads();
            if (threadIdx.x == 0) s_run = atomicAdd(reinterpret_cast<unsigned*>(prm.scratch + 3), 1u);
            __syncthreads();
            const long long run = s_run;
            if (run >= n_runs) break;
            const long long rb = prm.begin + run * kRunElems;
            const long long re = rb + kRunElems < prm.end ? rb + kRunElems : prm.end;
            const int n_pc = static_cast<int>((re - rb + kPieceElems - 1) / kPieceElems);
            // a piece travels as `world` bulk copies (one per rank, 4 or 8 KB each) issued by ONE thread; they
            // complete on the stage's mbarrier.  Ragged or unaligned pieces are not staged (element path below).
            auto staged = [&](int j) {
                return j < n_pc && !dead && vec_ok && rb + static_cast<long long>(j + 1) * kPieceElems <= re;
            };
            auto issue = [&](int j) {                                  // thread 0: piece j of the run -> stage j % n_stage
                if (!staged(j)) return;
                const long long pb = rb + static_cast<long long>(j) * kPieceElems;
                const int sidx = j % n_stage;
                unsigned char* st = s_stage + sidx * stage_bytes;
                mbar_expect_tx(&s_bar[sidx], kSliceBytes * prm.world);
                for (int r = 0; r < prm.world; r++)
                    bulk_load(st + static_cast<size_t>(r) * kSliceBytes, reinterpret_cast<const GT*>(prm.grads.p[r]) + pb,
                              kSliceBytes, &s_bar[sidx]);
            };
            if (threadIdx.x == 0)
                for (int j = 0; j < n_stage - 1; j++) issue(j);
            for (int j = 0; j < n_pc; j++) {
                __syncthreads();                                       // everybody is done with the stage issued next
                if (threadIdx.x == 0) issue(j + n_stage - 1);
                const long long pb = rb + static_cast<long long>(j) * kPieceElems;
                const long long pe = pb + kPieceElems < re ? pb + kPieceElems : re;
                if (!dead) {
                    if (staged(j)) {
                        const int sidx = j % n_stage;
                        mbar_wait(&s_bar[sidx], (phase >> sidx) & 1u); // piece j's slices have landed
                        phase ^= 1u << sidx;
                        const unsigned char* st = s_stage + sidx * stage_bytes;
                        float acc[kVec];
                        Raw8<G> raw;
                        lds8<G>(st + (threadIdx.x * kVec) * sizeof(GT), raw);
                        unpack8<G>(raw, acc);
#pragma unroll
                        for (int r = 1; r < kWorld; r++) {
                            if (r < prm.world) {
                                float f[kVec];
                                lds8<G>(st + (static_cast<size_t>(r) * kPieceElems + threadIdx.x * kVec) * sizeof(GT), raw);
                                unpack8<G>(raw, f);
#pragma unroll
                                for (int i = 0; i < kVec; i++) acc[i] = __fadd_rn(acc[i], f[i]);
                            }
                        }
#pragma unroll
                        for (int i = 0; i < kVec; i++) acc[i] = __fmul_rn(round_through<G>(acc[i]), prm.factor);
                        store8<G>(prm.reduced, pb + threadIdx.x * kVec, acc);
                    } else {                                           // ragged last piece / unaligned shard
                        for (int i = 0; i < kVec; i++) {
                            const long long e = pb + static_cast<long long>(threadIdx.x) * kVec + i;
                            if (e < pe) {
                                float x[kWorld];
#pragma unroll
                                for (int r = 0; r < kWorld; r++) x[r] = r < prm.world ? load1<G>(prm.grads.p[r], e) : 0.0f;
                                store1<G>(prm.reduced, e, peer_mean<G, kWorld>(x, prm.world, prm.factor));
                            }
                        }
                    }
                }
            }
            // the run is folded: one fence for all of its stores, then count its pieces towards its chunk
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) {
                const long long c = (rb - prm.begin) / prm.chunk_elems;
                const long long c_end = (c + 1) * prm.chunk_elems < shard ? (c + 1) * prm.chunk_elems : shard;
                const int in_chunk = static_cast<int>((c_end - c * prm.chunk_elems + kPieceElems - 1) / kPieceElems);
                if (atomicAdd(prm.scratch + 8 + c, n_pc) + n_pc == in_chunk) {
                    prm.scratch[8 + c] = 0;                            // ready for the next step
                    __threadfence_system();
                    for (int r = 0; r < prm.sync.world; r++)
                        st_release_sys(peer_board(prm.sync.pad[r]) + prm.sync.rank * B200_PEER_STEP_MAX_CHUNKS + c, prm.sync.epoch);
                }
            }
        }
    } else {
        // ================= UPDATE role: the speculative Adam pass over the (chunk-major) tile list ==============
        // Static striding over the update CTAs, next row prefetched with cp.async (as walk_tiles does); before the
        // barrier that hands a row over, thread 0 waits until the owner of that tile's gradient has announced the
        // chunk(s) it lies in.  Update CTAs never produce anything a reducer waits for, so blocking is safe.
        const TableView tv = open_table(table);
        int win = 0;
        Op::prologue(prm.spec, win);
        const uint32_t Gu = gridDim.x - prm.n_reducers;
        uint32_t t = blockIdx.x - prm.n_reducers;
        const bool loader = threadIdx.x < 4;
        if (t < tv.n_tiles) {
            uint2 tl_next = make_uint2(0u, 0u);
            if (loader) {
                const uint2 tl0 = __ldg(reinterpret_cast<const uint2*>(tv.tiles + t));
                cp_async16(reinterpret_cast<char*>(&s_slot[0].row) + threadIdx.x * 16,
                           reinterpret_cast<const char*>(tv.rows + tl0.x) + threadIdx.x * 16);
                if (threadIdx.x == 0) s_sh.tile[0] = tl0.y;
                if (t + Gu < tv.n_tiles) tl_next = __ldg(reinterpret_cast<const uint2*>(tv.tiles + t + Gu));
                cp_async_wait_all();
                if (threadIdx.x == 0) {
                    s_sh.seen_owner = -1;
                    s_sh.seen_lo = 0;
                    s_sh.seen_hi = -1;
                    step_tile_ready<GT, kWorld>(prm, &s_slot[0].row, s_sh.tile[0], tv.tile_elems, &s_sh, true);
                }
            }
            __syncthreads();
            int slot = 0;
            for (;;) {
                const bool has_next = t + Gu < tv.n_tiles;
                if (loader && has_next) {
                    cp_async16(reinterpret_cast<char*>(&s_slot[slot ^ 1].row) + threadIdx.x * 16,
                               reinterpret_cast<const char*>(tv.rows + tl_next.x) + threadIdx.x * 16);
                    if (threadIdx.x == 0) s_sh.tile[slot ^ 1] = tl_next.y;
                    const uint32_t t2 = t + 2 * Gu;
                    if (t2 < tv.n_tiles && t2 > t) tl_next = __ldg(reinterpret_cast<const uint2*>(tv.tiles + t2));
                }
                bad |= step_one_tile<Op, kThreads, kU>(prm.spec, &s_slot[slot].row, s_sh.tile[slot], tv.tile_elems, win);
                if (!has_next) break;
                if (loader) {
                    cp_async_wait_all();
                    if (threadIdx.x == 0)
                        step_tile_ready<GT, kWorld>(prm, &s_slot[slot ^ 1].row, s_sh.tile[slot ^ 1], tv.tile_elems, &s_sh, true);
                }
                __syncthreads();
                slot ^= 1;
                t += Gu;
            }
        }
    }
    bad |= wait_failed(&prm.sync.pad[prm.sync.rank]->error);
    Op::finish(prm.spec, bad);
    // the last CTA to leave resets the run tickets (the next launch is ordered after this one on the stream)
    __syncthreads();
    if (threadIdx.x == 0) {
        if (atomicAdd(prm.scratch + 2, 1) == static_cast<int>(gridDim.x) - 1) {
            prm.scratch[2] = 0;
            prm.scratch[3] = 0;
        }
    }
}

// ============================================================================================
// loss-scaler state machine (one thread; the whole point is that it runs on the device)
// ============================================================================================
__global__ void scaler_init_kernel(b200_scaler_state* s, float init_scale) {
    s->loss_scale = init_scale;
    s->inv_scale = __fdiv_rn(1.0f, init_scale);
    s->found_inf = 0;
    s->unskipped = 0;
    s->overflow_acc = 0;
    s->seq = 0;
    s->had_overflow = 0;
    s->inv_loss_scale = s->inv_scale;
    s->skipped_total = 0;
    s->reserved[0] = s->reserved[1] = s->reserved[2] = 0;
}

__device__ __forceinline__ int scaler_update_body(b200_scaler_state* s, int mode, int scale_window, float scale_factor,
                                                  float min_scale, float max_scale, b200_scaler_state* mirror) {
    b200_scaler_state st = *s;
    const int dynamic = mode & B200_SCALER_DYNAMIC;
    const int flag = st.overflow_acc != 0;
    st.inv_scale = __fdiv_rn(1.0f, st.loss_scale);
    st.had_overflow = flag;
    // B200_SCALER_KEEP_SKIP: no optimizer step has consumed the previous decision yet (a second backward of a
    // gradient-accumulation cycle), so an earlier overflow must keep the pending step skipped — apex's patched
    // skip_step stays in place until step() is called
    const int carried = (mode & B200_SCALER_KEEP_SKIP) ? st.found_inf : 0;
    st.found_inf = ((dynamic && flag) || carried) ? 1 : 0;
    if (dynamic) {
        if (flag) {
            float ns = __fdiv_rn(st.loss_scale, 2.0f);
            if (min_scale > 0.0f && ns < min_scale) ns = min_scale;
            st.loss_scale = ns;
            st.unskipped = 0;
        } else {
            st.unskipped += 1;
        }
        if (st.unskipped == scale_window) {
            float ns = __fmul_rn(st.loss_scale, scale_factor);
            if (ns > max_scale) ns = max_scale;
            st.loss_scale = ns;
            st.unskipped = 0;
        }
    }
    st.overflow_acc = 0;
    st.seq += 1;
    st.skipped_total += (st.found_inf && !carried) ? 1 : 0;
    st.inv_loss_scale = __fdiv_rn(1.0f, st.loss_scale);
    *s = st;
    if (mirror) {
        // payload first, sequence number last, so a host that polls `seq` sees a whole record
        volatile b200_scaler_state* mv = mirror;
        mv->loss_scale = st.loss_scale;
        mv->inv_scale = st.inv_scale;
        mv->found_inf = st.found_inf;
        mv->unskipped = st.unskipped;
        mv->overflow_acc = 0;
        mv->had_overflow = st.had_overflow;
        mv->inv_loss_scale = st.inv_loss_scale;
        mv->skipped_total = st.skipped_total;
        __threadfence_system();
        mv->seq = st.seq;
    }
    return st.found_inf;
}

__global__ void scaler_update_kernel(b200_scaler_state* s, int mode, int scale_window, float scale_factor,
                                     float min_scale, float max_scale, b200_scaler_state* mirror) {
    scaler_update_body(s, mode, scale_window, scale_factor, min_scale, max_scale, mirror);
}

// Commit of a speculative step: the loss-scaler update (which turns the accumulated overflow flag into
// found_inf and the new scale) followed by the parity flip that makes the freshly written copies current —
// only if nothing overflowed.  Also advances the {calls, skips} step counters.
__global__ void spec_commit_kernel(b200_scaler_state* s, int mode, int scale_window, float scale_factor, float min_scale,
                                   float max_scale, b200_scaler_state* mirror, int32_t* parity, StepSync sync) {
    const int found = scaler_update_body(s, mode, scale_window, scale_factor, min_scale, max_scale, mirror);
    if (!found) *parity = (*parity & 1) ^ 1;
    step_epilogue(sync, found != 0);
}

// ============================================================================================
// dtype dispatch helpers
// ============================================================================================
static bool valid_dt(int dt) { return dt == B200_F32 || dt == B200_F16 || dt == B200_BF16; }

#define B200_DISPATCH_DT(dt, NAME, ...)                                    \
    switch (dt) {                                                          \
    case B200_F32: { constexpr int NAME = B200_F32; __VA_ARGS__; } break;  \
    case B200_F16: { constexpr int NAME = B200_F16; __VA_ARGS__; } break;  \
    default: { constexpr int NAME = B200_BF16; __VA_ARGS__; } break;       \
    }

// The (grad, state, copy) triples the mixed-precision path produces:
//   O2        : (f16|bf16, f32, same 16-bit)     fp16 model grad, fp32 master, fp16 model copy
//   O2, BN    : (f32, f32, none)                 params kept in fp32
//   O2 master grads materialised: (f32, f32, f16|bf16)
//   O0 / O1   : (f32, f32, none)
//   O3        : (f16, f16, none) / (bf16, bf16, none)
//   O2 without copy (copy done elsewhere): (f16|bf16, f32, none)
template <template <int, int, int> class OpT, class Params>
static int dispatch_step(const char* name, int g, int s, int c, const void* tbl, int64_t n_tiles,
                         const Params& prm, void* stream) {
#define B200_TRY(G_, S_, C_) \
    if (g == G_ && s == S_ && c == C_) return launch_tiles<OpT<G_, S_, C_>>(name, tbl, n_tiles, prm, stream);
    B200_TRY(B200_F16, B200_F32, B200_F16)
    B200_TRY(B200_BF16, B200_F32, B200_BF16)
    B200_TRY(B200_F32, B200_F32, B200_NONE)
    B200_TRY(B200_F32, B200_F32, B200_F16)
    B200_TRY(B200_F32, B200_F32, B200_BF16)
    B200_TRY(B200_F16, B200_F16, B200_NONE)
    B200_TRY(B200_BF16, B200_BF16, B200_NONE)
    B200_TRY(B200_F16, B200_F32, B200_NONE)
    B200_TRY(B200_BF16, B200_F32, B200_NONE)
#undef B200_TRY
    return set_error(B200_EDTYPE, "%s: (grad,state,copy) dtypes (%d,%d,%d) not instantiated", name, g, s, c);
}

}  // namespace b200

// ================================================================================================
// extern "C" entry points
// ================================================================================================
using namespace b200;

extern "C" {

int32_t b200_abi_version(void) { return B200_ABI_VERSION; }
const char* b200_last_error(void) { return g_err; }
int64_t b200_launch_count(void) { return g_launches.load(); }

int32_t b200_set_launch_config(int32_t ctas_per_sm, int32_t threads, int32_t unroll) {
    if (threads != 256 && threads != 512) return set_error(B200_EINVAL, "threads must be 256 or 512");
    if (unroll != 1 && unroll != 2) return set_error(B200_EINVAL, "unroll must be 1 or 2");
    if (ctas_per_sm < 0 || ctas_per_sm > 32) return set_error(B200_EINVAL, "ctas_per_sm out of range");
    g_ctas_per_sm.store(ctas_per_sm);
    g_threads.store(threads);
    g_unroll.store(unroll);
    return 0;
}

int32_t b200_table_upload(void* dev_image, const void* host_image, int64_t image_bytes, void* stream) {
    if (!dev_image || !host_image || image_bytes <= 0) return set_error(B200_EINVAL, "table_upload: bad argument");
    cudaError_t e = cudaMemcpyAsync(dev_image, host_image, static_cast<size_t>(image_bytes), cudaMemcpyHostToDevice,
                                    static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? 0 : cuda_fail(e, "table_upload");
}

int32_t b200_scaler_init(b200_scaler_state* dev_state, float init_scale, void* stream) {
    if (!dev_state || !(init_scale > 0.0f)) return set_error(B200_EINVAL, "scaler_init: bad argument");
    scaler_init_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(dev_state, init_scale);
    return after_launch("scaler_init");
}

int32_t b200_scaler_update(b200_scaler_state* dev_state, int32_t dynamic, int32_t scale_window, float scale_factor,
                           float min_scale, float max_scale, b200_scaler_state* mirror, void* stream) {
    if (!dev_state) return set_error(B200_EINVAL, "scaler_update: dev_state is NULL");
    if (dynamic & ~(B200_SCALER_DYNAMIC | B200_SCALER_KEEP_SKIP)) return set_error(B200_EINVAL, "scaler_update: unknown mode bits");
    if ((dynamic & B200_SCALER_DYNAMIC) && scale_window <= 0) return set_error(B200_EINVAL, "scaler_update: scale_window must be > 0");
    scaler_update_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(dev_state, dynamic, scale_window, scale_factor,
                                                                          min_scale, max_scale, mirror);
    return after_launch("scaler_update");
}

int32_t b200_multi_tensor_scale(const void* dev_table, int64_t n_tiles, int32_t in_dtype, int32_t out_dtype, float scale,
                                const float* dev_scale, int32_t* dev_flag, void* stream) {
    if (!valid_dt(in_dtype) || !valid_dt(out_dtype)) return set_error(B200_EDTYPE, "multi_tensor_scale: bad dtype");
    ScaleParams prm{scale, dev_scale, dev_flag};
    B200_DISPATCH_DT(in_dtype, I,
        B200_DISPATCH_DT(out_dtype, O,
            return launch_tiles<ScaleOp<I, O>>("multi_tensor_scale", dev_table, n_tiles, prm, stream)))
    return 0;
}

int32_t b200_multi_tensor_axpby(const void* dev_table, int64_t n_tiles, int32_t x_dtype, int32_t y_dtype, int32_t out_dtype,
                                float a, float b, int32_t arg_to_check, int32_t* dev_flag, void* stream) {
    if (!valid_dt(x_dtype) || !valid_dt(y_dtype) || !valid_dt(out_dtype))
        return set_error(B200_EDTYPE, "multi_tensor_axpby: bad dtype");
    // apex uses axpby to merge freshly unscaled grads into stashed ones: out and y share a type.
    if (y_dtype != out_dtype) return set_error(B200_EDTYPE, "multi_tensor_axpby: y and out must share a dtype");
    AxpbyParams prm{a, b, arg_to_check, dev_flag};
    B200_DISPATCH_DT(x_dtype, X,
        B200_DISPATCH_DT(out_dtype, O,
            return launch_tiles<AxpbyOp<X, O, O>>("multi_tensor_axpby", dev_table, n_tiles, prm, stream)))
    return 0;
}

int32_t b200_multi_tensor_check(const void* dev_table, int64_t n_tiles, int32_t col, int32_t dtype, int32_t* dev_flag,
                                void* stream) {
    if (!valid_dt(dtype)) return set_error(B200_EDTYPE, "multi_tensor_check: bad dtype");
    if (col < 0 || col >= B200_MAX_DEPTH) return set_error(B200_EINVAL, "multi_tensor_check: bad column");
    if (!dev_flag) return set_error(B200_EINVAL, "multi_tensor_check: dev_flag is NULL");
    CheckParams prm{col, dev_flag};
    B200_DISPATCH_DT(dtype, I, return launch_tiles<CheckOp<I>>("multi_tensor_check", dev_table, n_tiles, prm, stream))
    return 0;
}

static StepSync to_sync(const b200_step_sync* s) {
    StepSync r{nullptr, nullptr, 0, 0};
    if (s) { r.counters = s->dev_counters; r.mirror = s->mirror; r.known_skips = s->known_skips; r.read_only = s->read_only; }
    return r;
}

static int to_pull(const b200_peer_pull* p, PeerPull& out, const char* name) {
    memset(&out, 0, sizeof(out));
    out.world = 1;
    out.elem_size = 2;
    if (!p) return 0;
    if (p->world < 1 || p->world > B200_MAX_PEERS) return set_error(B200_EINVAL, "%s: world must be 1..%d", name, B200_MAX_PEERS);
    if (p->elem_size != 2 && p->elem_size != 4) return set_error(B200_EINVAL, "%s: elem_size must be 2 or 4", name);
    if (p->world > 1 && !p->local_base) return set_error(B200_EINVAL, "%s: local_base is NULL", name);
    out.local_base = static_cast<const char*>(p->local_base);
    out.world = p->world;
    out.elem_size = p->elem_size;
    out.first_tile = p->first_tile;
    for (int r = 0; r <= p->world; r++) out.bounds[r] = p->bounds[r];
    for (int r = 0; r < p->world; r++) out.delta[r] = p->delta[r];
    for (int r = 0; r < p->world; r++)
        if (p->bounds[r] > p->bounds[r + 1]) return set_error(B200_EINVAL, "%s: shard bounds must be non-decreasing", name);
    return 0;
}

int32_t b200_multi_tensor_l2norm(const void* dev_table, int64_t n_tiles, int32_t col, int32_t dtype, float* dev_sumsq,
                                 void* dev_scratch, int64_t scratch_bytes, int32_t* dev_flag, void* stream) {
    if (!valid_dt(dtype)) return set_error(B200_EDTYPE, "multi_tensor_l2norm: bad dtype");
    if (col < 0 || col >= B200_MAX_DEPTH) return set_error(B200_EINVAL, "multi_tensor_l2norm: bad column");
    if (!dev_sumsq || !dev_scratch) return set_error(B200_EINVAL, "multi_tensor_l2norm: NULL output or scratch");
    if (scratch_bytes < B200_NORM_SCRATCH_BYTES)
        return set_error(B200_ESIZE, "multi_tensor_l2norm: scratch must be at least %d bytes (zeroed once)", B200_NORM_SCRATCH_BYTES);
    NormParams prm{col, dev_flag, static_cast<float*>(dev_scratch), dev_sumsq};
    const unsigned cap = B200_NORM_SCRATCH_BYTES / 4 - 2;
    B200_DISPATCH_DT(dtype, I, return launch_tiles<NormCheckOp<I>>("multi_tensor_l2norm", dev_table, n_tiles, prm, stream, cap))
    return 0;
}

int32_t b200_clip_scale(float* dev_sumsq, const float* dev_inv_scale, float inv_scale, float max_norm, float* dev_scale_out,
                        float* dev_norm_out, void* stream) {
    if (!dev_sumsq || !dev_scale_out) return set_error(B200_EINVAL, "clip_scale: NULL argument");
    if (!(max_norm > 0.0f)) return set_error(B200_EINVAL, "clip_scale: max_norm must be > 0");
    clip_scale_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(dev_sumsq, dev_inv_scale, inv_scale, max_norm,
                                                                     dev_scale_out, dev_norm_out);
    return after_launch("clip_scale");
}

// Per-group constants of one Adam launch for every step-window candidate.  Hyper-parameters are derived in
// double and rounded to float once, exactly like the Python scalars torch.optim.Adam hands to its fp32 tensor ops.
static int fill_adam_consts(const char* name, const b200_adam_group* groups, int32_t n_groups,
                            AdamConsts (*grp)[B200_STEP_WINDOW]) {
    if (!groups || n_groups < 1 || n_groups > B200_MAX_GROUPS)
        return set_error(B200_EINVAL, "%s: n_groups must be 1..%d", name, B200_MAX_GROUPS);
    for (int i = 0; i < n_groups; i++) {
        const b200_adam_group& h = groups[i];
        if (h.step < 1) return set_error(B200_EINVAL, "%s: step must be >= 1", name);
        for (int w = 0; w < B200_STEP_WINDOW; w++) {
            AdamConsts& k = grp[i][w];
            const int64_t step = h.step - w > 1 ? h.step - w : 1;
            k.w1 = static_cast<float>(1.0 - h.beta1);
            k.one_m_w1 = 1.0f - k.w1;
            k.b2 = static_cast<float>(h.beta2);
            k.w2 = static_cast<float>(1.0 - h.beta2);
            const double bc1 = h.bias_correction ? 1.0 - std::pow(h.beta1, static_cast<double>(step)) : 1.0;
            const double bc2 = h.bias_correction ? 1.0 - std::pow(h.beta2, static_cast<double>(step)) : 1.0;
            k.neg_step = static_cast<float>(-(h.lr / bc1));
            k.bc2_sqrt = static_cast<float>(std::pow(bc2, 0.5));
            k.eps = static_cast<float>(h.eps);
            k.wd = static_cast<float>(h.weight_decay);
            k.decay = static_cast<float>(1.0 - h.lr * h.weight_decay);
            k.flags = (h.weight_decay != 0.0 ? 1u : 0u) | (h.adam_w_mode ? 2u : 0u) | (k.w1 < 0.5f ? 0u : 4u);
        }
    }
    return 0;
}

int32_t b200_fused_adam(const void* dev_table, int64_t n_tiles, int32_t grad_dtype, int32_t state_dtype, int32_t copy_dtype,
                        const b200_adam_group* groups, int32_t n_groups, float inv_scale, const float* dev_inv_scale,
                        const int32_t* dev_found_inf, const b200_step_sync* sync, void* stream) {
    return b200_fused_adam_peer(dev_table, n_tiles, grad_dtype, state_dtype, copy_dtype, groups, n_groups, inv_scale,
                                dev_inv_scale, dev_found_inf, sync, nullptr, stream);
}

int32_t b200_fused_sgd(const void* dev_table, int64_t n_tiles, int32_t grad_dtype, int32_t state_dtype, int32_t copy_dtype,
                       const b200_sgd_group* groups, int32_t n_groups, float inv_scale, const float* dev_inv_scale,
                       const int32_t* dev_found_inf, const b200_step_sync* sync, void* stream) {
    return b200_fused_sgd_peer(dev_table, n_tiles, grad_dtype, state_dtype, copy_dtype, groups, n_groups, inv_scale,
                               dev_inv_scale, dev_found_inf, sync, nullptr, stream);
}

int32_t b200_fused_adam_peer(const void* dev_table, int64_t n_tiles, int32_t grad_dtype, int32_t state_dtype,
                             int32_t copy_dtype, const b200_adam_group* groups, int32_t n_groups, float inv_scale,
                             const float* dev_inv_scale, const int32_t* dev_found_inf, const b200_step_sync* sync,
                             const b200_peer_pull* pull, void* stream) {
    AdamParams prm;
    memset(&prm, 0, sizeof(prm));
    if (int rc = to_pull(pull, prm.pull, "fused_adam")) return rc;
    if (int rc = fill_adam_consts("fused_adam", groups, n_groups, prm.grp)) return rc;
    prm.inv_scale = inv_scale;
    prm.dev_inv_scale = dev_inv_scale;
    prm.dev_found_inf = dev_found_inf;
    prm.sync = to_sync(sync);
    return dispatch_step<AdamOp>("fused_adam", grad_dtype, state_dtype, copy_dtype, dev_table, n_tiles, prm, stream);
}

// ------------------------------------------------------------------------------------------------
// per-tensor norms and LAMB
// ------------------------------------------------------------------------------------------------
int32_t b200_l2norm_per_tensor(const void* dev_table, int64_t n_tiles, int32_t col, int32_t dtype,
                               const uint32_t* dev_first_tile, int32_t n_tensors, float* dev_tile_ss,
                               float* dev_sumsq_per_tensor, float* dev_norm_per_tensor, float* dev_total_sumsq,
                               int32_t* dev_flag, void* stream) {
    if (!valid_dt(dtype)) return set_error(B200_EDTYPE, "l2norm_per_tensor: bad dtype");
    if (col < 0 || col >= B200_MAX_DEPTH) return set_error(B200_EINVAL, "l2norm_per_tensor: bad column");
    if (n_tensors < 0 || n_tiles < 0 || n_tiles > 0x7fffffffll) return set_error(B200_EINVAL, "l2norm_per_tensor: bad size");
    if (n_tensors == 0) return 0;
    if (!dev_table || !dev_first_tile || !dev_tile_ss || !dev_sumsq_per_tensor)
        return set_error(B200_EINVAL, "l2norm_per_tensor: NULL argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n_tiles > 0) {
        B200_DISPATCH_DT(dtype, I, (tile_sumsq_kernel<I, 256><<<static_cast<unsigned>(n_tiles), 256, 0, st>>>(
                                       dev_table, col, dev_first_tile, dev_tile_ss, dev_flag)))
        g_launches.fetch_add(1);
    }
    tensor_norm_kernel<<<n_tensors, 32, 0, st>>>(dev_tile_ss, dev_first_tile, n_tensors, dev_sumsq_per_tensor, dev_norm_per_tensor);
    g_launches.fetch_add(1);
    if (dev_total_sumsq) {
        total_sumsq_kernel<<<1, 32, 0, st>>>(dev_sumsq_per_tensor, n_tensors, dev_total_sumsq);
        g_launches.fetch_add(1);
    }
    return after_launch("l2norm_per_tensor");
}

static int fill_lamb_consts(const char* name, const b200_lamb_group* groups, int32_t n_groups, LambConsts* grp) {
    if (!groups || n_groups < 1 || n_groups > B200_MAX_GROUPS)
        return set_error(B200_EINVAL, "%s: n_groups must be 1..%d", name, B200_MAX_GROUPS);
    for (int i = 0; i < n_groups; i++) {
        const b200_lamb_group& h = groups[i];
        if (h.step < 1) return set_error(B200_EINVAL, "%s: step must be >= 1", name);
        LambConsts& k = grp[i];
        k.b1 = static_cast<float>(h.beta1);
        k.b3 = h.grad_averaging ? static_cast<float>(1.0 - h.beta1) : 1.0f;
        k.b2 = static_cast<float>(h.beta2);
        k.one_m_b2 = static_cast<float>(1.0 - h.beta2);
        k.bc1 = h.bias_correction ? static_cast<float>(1.0 - std::pow(h.beta1, static_cast<double>(h.step))) : 1.0f;
        k.bc2 = h.bias_correction ? static_cast<float>(1.0 - std::pow(h.beta2, static_cast<double>(h.step))) : 1.0f;
        k.eps = static_cast<float>(h.eps);
        k.decay = static_cast<float>(h.weight_decay);
        k.lr = static_cast<float>(h.lr);
        k.flags = (h.adam_w_mode ? 1u : 0u) | (h.use_nvlamb ? 2u : 0u);
    }
    return 0;
}

int32_t b200_lamb_stage1(const void* dev_table, int64_t n_tiles, int32_t grad_dtype, int32_t state_dtype,
                         const b200_lamb_group* groups, int32_t n_groups, float inv_scale, const float* dev_inv_scale,
                         const int32_t* dev_found_inf, const float* dev_grad_sumsq, float max_grad_norm, void* stream) {
    if (!dev_grad_sumsq) return set_error(B200_EINVAL, "lamb_stage1: dev_grad_sumsq is NULL");
    Lamb1Params prm;
    memset(&prm, 0, sizeof(prm));
    if (int rc = fill_lamb_consts("lamb_stage1", groups, n_groups, prm.grp)) return rc;
    prm.inv_scale = inv_scale;
    prm.dev_inv_scale = dev_inv_scale;
    prm.dev_found_inf = dev_found_inf;
    prm.dev_grad_sumsq = dev_grad_sumsq;
    prm.max_grad_norm = max_grad_norm;
#define B200_TRY(G_, S_) \
    if (grad_dtype == G_ && state_dtype == S_) return launch_tiles<Lamb1Op<G_, S_>>("lamb_stage1", dev_table, n_tiles, prm, stream);
    B200_TRY(B200_F16, B200_F32)
    B200_TRY(B200_BF16, B200_F32)
    B200_TRY(B200_F32, B200_F32)
    B200_TRY(B200_F16, B200_F16)
    B200_TRY(B200_BF16, B200_BF16)
#undef B200_TRY
    return set_error(B200_EDTYPE, "lamb_stage1: (grad,state) dtypes (%d,%d) not instantiated", grad_dtype, state_dtype);
}

int32_t b200_lamb_stage2(const void* dev_table, int64_t n_tiles, int32_t state_dtype, int32_t copy_dtype,
                         const b200_lamb_group* groups, int32_t n_groups, const int32_t* dev_found_inf,
                         const float* dev_p_norm, const float* dev_u_norm, void* stream) {
    if (!dev_p_norm || !dev_u_norm) return set_error(B200_EINVAL, "lamb_stage2: NULL norm array");
    Lamb2Params prm;
    memset(&prm, 0, sizeof(prm));
    if (int rc = fill_lamb_consts("lamb_stage2", groups, n_groups, prm.grp)) return rc;
    prm.dev_found_inf = dev_found_inf;
    prm.p_norm = dev_p_norm;
    prm.u_norm = dev_u_norm;
#define B200_TRY(S_, C_) \
    if (state_dtype == S_ && copy_dtype == C_) return launch_tiles<Lamb2Op<S_, C_>>("lamb_stage2", dev_table, n_tiles, prm, stream);
    B200_TRY(B200_F32, B200_NONE)
    B200_TRY(B200_F32, B200_F16)
    B200_TRY(B200_F32, B200_BF16)
    B200_TRY(B200_F16, B200_NONE)
    B200_TRY(B200_BF16, B200_NONE)
#undef B200_TRY
    return set_error(B200_EDTYPE, "lamb_stage2: (state,copy) dtypes (%d,%d) not instantiated", state_dtype, copy_dtype);
}

// ------------------------------------------------------------------------------------------------
// NovoGrad
// ------------------------------------------------------------------------------------------------
static int fill_novo(const char* name, const b200_novograd_group* groups, int32_t n_groups, float inv_scale,
                     const float* dev_inv_scale, const int32_t* dev_found_inf, const float* dev_v, NovoParams& prm) {
    memset(&prm, 0, sizeof(prm));
    if (!groups || n_groups < 1 || n_groups > B200_MAX_GROUPS)
        return set_error(B200_EINVAL, "%s: n_groups must be 1..%d", name, B200_MAX_GROUPS);
    if (!dev_v) return set_error(B200_EINVAL, "%s: dev_v is NULL", name);
    for (int i = 0; i < n_groups; i++) {
        const b200_novograd_group& h = groups[i];
        if (h.step < 1) return set_error(B200_EINVAL, "%s: step must be >= 1", name);
        NovoConsts& k = prm.grp[i];
        k.b1 = static_cast<float>(h.beta1);
        k.b3 = h.grad_averaging ? static_cast<float>(1.0 - h.beta1) : 1.0f;
        k.b2 = static_cast<float>(h.beta2);
        k.one_m_b2 = static_cast<float>(1.0 - h.beta2);
        k.bc1 = h.bias_correction ? static_cast<float>(1.0 - std::pow(h.beta1, static_cast<double>(h.step))) : 1.0f;
        k.bc2 = h.bias_correction ? static_cast<float>(std::sqrt(1.0 - std::pow(h.beta2, static_cast<double>(h.step)))) : 1.0f;
        k.eps = static_cast<float>(h.eps);
        k.decay = static_cast<float>(h.weight_decay);
        k.lr = static_cast<float>(h.lr);
        k.flags = (h.reg_inside_moment ? 1u : 0u) | ((h.step == 1 && !h.init_zero) ? 2u : 0u);
    }
    prm.inv_scale = inv_scale;
    prm.dev_inv_scale = dev_inv_scale;
    prm.dev_found_inf = dev_found_inf;
    prm.v = dev_v;
    return 0;
}

int32_t b200_novograd_blend(const void* dev_table, int32_t n_tensors, const b200_novograd_group* groups, int32_t n_groups,
                            float inv_scale, const float* dev_inv_scale, const int32_t* dev_found_inf, float* dev_v,
                            const float* dev_grad_sumsq, void* stream) {
    if (n_tensors <= 0) return 0;
    if (!dev_table || !dev_grad_sumsq) return set_error(B200_EINVAL, "novograd_blend: NULL argument");
    NovoParams prm;
    if (int rc = fill_novo("novograd_blend", groups, n_groups, inv_scale, dev_inv_scale, dev_found_inf, dev_v, prm)) return rc;
    novograd_blend_kernel<<<(n_tensors + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(dev_table, prm, dev_v,
                                                                                                 dev_grad_sumsq);
    return after_launch("novograd_blend");
}

int32_t b200_fused_novograd(const void* dev_table, int64_t n_tiles, int32_t grad_dtype, int32_t state_dtype,
                            int32_t copy_dtype, const b200_novograd_group* groups, int32_t n_groups, float inv_scale,
                            const float* dev_inv_scale, const int32_t* dev_found_inf, const float* dev_v, void* stream) {
    NovoParams prm;
    if (int rc = fill_novo("fused_novograd", groups, n_groups, inv_scale, dev_inv_scale, dev_found_inf, dev_v, prm)) return rc;
    return dispatch_step<NovoGradOp>("fused_novograd", grad_dtype, state_dtype, copy_dtype, dev_table, n_tiles, prm, stream);
}

// ------------------------------------------------------------------------------------------------
// speculative single-pass step (see AdamSpecOp)
// ------------------------------------------------------------------------------------------------
int64_t b200_spec_shadow_delta(int64_t numel, int32_t elem_size) {
    if (numel < 0 || (elem_size != 2 && elem_size != 4)) return set_error(B200_EINVAL, "spec_shadow_delta: bad argument");
    return (numel * elem_size + 255) / 256 * 256;
}

int32_t b200_fused_adam_spec(const void* dev_table, int64_t n_tiles, int32_t grad_dtype, int32_t state_dtype,
                             int32_t copy_dtype, const b200_adam_group* groups, int32_t n_groups, float inv_scale,
                             const float* dev_inv_scale, int32_t* dev_flag, const int32_t* dev_parity,
                             const b200_step_sync* sync, const b200_peer_pull* pull, void* stream) {
    if (!dev_parity) return set_error(B200_EINVAL, "fused_adam_spec: dev_parity is NULL");
    SpecParams prm;
    memset(&prm, 0, sizeof(prm));
    if (int rc = to_pull(pull, prm.pull, "fused_adam_spec")) return rc;
    if (int rc = fill_adam_consts("fused_adam_spec", groups, n_groups, prm.grp)) return rc;
    prm.inv_scale = inv_scale;
    prm.dev_inv_scale = dev_inv_scale;
    prm.flag = dev_flag;
    prm.parity = dev_parity;
    prm.sync = to_sync(sync);
    prm.sync.read_only = 1;                              // the commit kernel advances the counters
    return dispatch_step<AdamSpecOp>("fused_adam_spec", grad_dtype, state_dtype, copy_dtype, dev_table, n_tiles, prm, stream);
}

int32_t b200_spec_commit(b200_scaler_state* dev_state, int32_t mode, int32_t scale_window, float scale_factor,
                         float min_scale, float max_scale, b200_scaler_state* mirror, int32_t* dev_parity,
                         const b200_step_sync* sync, void* stream) {
    if (!dev_state || !dev_parity) return set_error(B200_EINVAL, "spec_commit: NULL argument");
    if (mode & ~(B200_SCALER_DYNAMIC | B200_SCALER_KEEP_SKIP)) return set_error(B200_EINVAL, "spec_commit: unknown mode bits");
    if ((mode & B200_SCALER_DYNAMIC) && scale_window <= 0) return set_error(B200_EINVAL, "spec_commit: scale_window must be > 0");
    spec_commit_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(dev_state, mode, scale_window, scale_factor, min_scale,
                                                                       max_scale, mirror, dev_parity, to_sync(sync));
    return after_launch("spec_commit");
}

int32_t b200_spec_repair(const void* dev_table, int64_t n_tiles, int32_t state_dtype, int32_t copy_dtype,
                         const int32_t* dev_found_inf, const int32_t* dev_parity, void* stream) {
    if (!dev_found_inf || !dev_parity) return set_error(B200_EINVAL, "spec_repair: NULL argument");
    RepairParams prm{dev_found_inf, dev_parity};
    if (state_dtype == B200_F32 && copy_dtype == B200_F16)
        return launch_tiles<SpecRepairOp<B200_F32, B200_F16>>("spec_repair", dev_table, n_tiles, prm, stream);
    if (state_dtype == B200_F32 && copy_dtype == B200_BF16)
        return launch_tiles<SpecRepairOp<B200_F32, B200_BF16>>("spec_repair", dev_table, n_tiles, prm, stream);
    return set_error(B200_EDTYPE, "spec_repair: (state,copy) dtypes (%d,%d) not instantiated", state_dtype, copy_dtype);
}

int32_t b200_fused_sgd_peer(const void* dev_table, int64_t n_tiles, int32_t grad_dtype, int32_t state_dtype,
                            int32_t copy_dtype, const b200_sgd_group* groups, int32_t n_groups, float inv_scale,
                            const float* dev_inv_scale, const int32_t* dev_found_inf, const b200_step_sync* sync,
                            const b200_peer_pull* pull, void* stream) {
    if (!groups || n_groups < 1 || n_groups > B200_MAX_GROUPS)
        return set_error(B200_EINVAL, "fused_sgd: n_groups must be 1..%d", B200_MAX_GROUPS);
    SgdParams prm;
    memset(&prm, 0, sizeof(prm));
    if (int rc = to_pull(pull, prm.pull, "fused_sgd")) return rc;
    for (int i = 0; i < n_groups; i++) {
        const b200_sgd_group& h = groups[i];
        if (h.step < 1) return set_error(B200_EINVAL, "fused_sgd: step must be >= 1");
        for (int w = 0; w < B200_STEP_WINDOW; w++) {
            SgdConsts& k = prm.grp[i][w];
            const bool first_run = h.step - w <= 1;
            k.wd = static_cast<float>(h.weight_decay);
            k.mu = static_cast<float>(h.momentum);
            k.damp = static_cast<float>(1.0 - h.dampening);
            k.neg_lr = static_cast<float>(-h.lr);
            k.flags = (h.weight_decay != 0.0 ? 1u : 0u) | (h.momentum != 0.0 ? 2u : 0u) | (h.nesterov ? 4u : 0u) |
                      (first_run ? 8u : 0u) | (h.wd_after_momentum ? 16u : 0u);
        }
    }
    prm.inv_scale = inv_scale;
    prm.dev_inv_scale = dev_inv_scale;
    prm.dev_found_inf = dev_found_inf;
    prm.sync = to_sync(sync);
    return dispatch_step<SgdOp>("fused_sgd", grad_dtype, state_dtype, copy_dtype, dev_table, n_tiles, prm, stream);
}

// ------------------------------------------------------------------------------------------------
// peer buffer set-up (CUDA IPC).  Not part of the step: called once when the wrapper is built.
// ------------------------------------------------------------------------------------------------
static_assert(sizeof(cudaIpcMemHandle_t) == B200_IPC_HANDLE_BYTES, "IPC handle size is part of the ABI");

int32_t b200_peer_alloc(int64_t bytes, void** dev_ptr, void* handle_out) {
    if (bytes <= 0 || !dev_ptr || !handle_out) return set_error(B200_EINVAL, "peer_alloc: bad argument");
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, static_cast<size_t>(bytes));
    if (e != cudaSuccess) return cuda_fail(e, "peer_alloc: cudaMalloc");
    e = cudaMemset(p, 0, static_cast<size_t>(bytes));
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(static_cast<cudaIpcMemHandle_t*>(handle_out), p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return cuda_fail(e, "peer_alloc: cudaIpcGetMemHandle");
    }
    *dev_ptr = p;
    return 0;
}

int32_t b200_peer_open(const void* handle, void** dev_ptr) {
    if (!handle || !dev_ptr) return set_error(B200_EINVAL, "peer_open: bad argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return cuda_fail(e, "peer_open: cudaIpcOpenMemHandle (are the GPUs peer-accessible?)");
    *dev_ptr = p;
    return 0;
}

int32_t b200_peer_close(void* dev_ptr) {
    if (!dev_ptr) return 0;
    cudaError_t e = cudaIpcCloseMemHandle(dev_ptr);
    return e == cudaSuccess ? 0 : cuda_fail(e, "peer_close");
}

int32_t b200_peer_free(void* dev_ptr) {
    if (!dev_ptr) return 0;
    cudaError_t e = cudaFree(dev_ptr);
    return e == cudaSuccess ? 0 : cuda_fail(e, "peer_free");
}

// ------------------------------------------------------------------------------------------------
// b200_peer_reduce_shard: reduce-scatter + mean + overflow check over NVLink peer memory.
// ------------------------------------------------------------------------------------------------
static int to_peer_sync(const b200_peer_sync* s, PeerSync& out, const char* name) {
    memset(&out, 0, sizeof(out));
    if (!s) return 0;
    if (s->world < 1 || s->world > B200_MAX_PEERS || s->rank < 0 || s->rank >= s->world || !s->pads)
        return set_error(B200_EINVAL, "%s: bad b200_peer_sync (world %d, rank %d)", name, s->world, s->rank);
    if (s->epoch < 1 || s->epoch >= (1 << 30)) return set_error(B200_EINVAL, "%s: epoch must be in [1, 2^30)", name);
    for (int r = 0; r < s->world; r++) {
        if (!s->pads[r]) return set_error(B200_EINVAL, "%s: signal pad %d is NULL", name, r);
        out.pad[r] = s->pads[r];
    }
    out.world = s->world;
    out.rank = s->rank;
    out.epoch = s->epoch;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// sharded step (see ShardStepOp)
// ------------------------------------------------------------------------------------------------
static int to_fan(const b200_peer_fan* f, PeerFan& out, const char* name) {
    if (!f) return set_error(B200_EINVAL, "%s: fan is NULL", name);
    if (f->world < 1 || f->world > B200_MAX_PEERS || f->rank < 0 || f->rank >= f->world)
        return set_error(B200_EINVAL, "%s: bad b200_peer_fan (world %d, rank %d)", name, f->world, f->rank);
    memset(&out, 0, sizeof(out));
    for (int r = 0; r < f->world; r++) {
        if (!f->grads[r] || !f->models[r]) return set_error(B200_EINVAL, "%s: buffer of rank %d is NULL", name, r);
        out.grad_delta[r] = static_cast<const char*>(f->grads[r]) - static_cast<const char*>(f->grads[f->rank]);
        out.copy_delta[r] = static_cast<const char*>(f->models[r]) - static_cast<const char*>(f->models[f->rank]);
        if (out.grad_delta[r] % 32 || out.copy_delta[r] % 32)
            return set_error(B200_EINVAL, "%s: the ranks' buffers must be 32-byte aligned relative to each other", name);
    }
    out.world = f->world;
    out.store_world = f->world;
    out.factor = f->factor;
    if (f->grads_mc) {
        out.grad_mc_delta = static_cast<const char*>(f->grads_mc) - static_cast<const char*>(f->grads[f->rank]);
        if (out.grad_mc_delta % 32) return set_error(B200_EINVAL, "%s: grads_mc must be 32-byte aligned like the buffers", name);
        out.use_mc_grad = 1;
    }
    if (f->model_mc) {
        out.copy_mc_delta = static_cast<const char*>(f->model_mc) - static_cast<const char*>(f->models[f->rank]);
        if (out.copy_mc_delta % 32) return set_error(B200_EINVAL, "%s: model_mc must be 32-byte aligned like the buffers", name);
        out.use_mc = 1;
    }
    // measurement aid (tools/, never set by the package): B200_SHARD_DEBUG bit 0 = store to the local model buffer only,
    // bit 1 = read every rank's gradient from the local buffer.  Results are then wrong by construction.
    if (const char* e = std::getenv("B200_SHARD_DEBUG")) {
        const int dbg = std::atoi(e);
        if (dbg & 1) {
            out.store_world = 1;
            out.copy_delta[0] = 0;
        }
        if (dbg & 2)
            for (int r = 0; r < f->world; r++) out.grad_delta[r] = 0;
    }
    return 0;
}

int32_t b200_peer_shard_step(const void* dev_table, int64_t n_tiles, int32_t grad_dtype, int32_t state_dtype,
                             int32_t copy_dtype, const b200_adam_group* groups, int32_t n_groups, float inv_scale,
                             const float* dev_inv_scale, int32_t* dev_flag, const int32_t* dev_parity,
                             const b200_step_sync* sync, const b200_peer_fan* fan, const b200_peer_sync* psync,
                             void* stream) {
    if (!dev_parity) return set_error(B200_EINVAL, "peer_shard_step: dev_parity is NULL");
    ShardStepParams prm;
    memset(&prm, 0, sizeof(prm));
    if (int rc = to_fan(fan, prm.fan, "peer_shard_step")) return rc;
    if (int rc = to_peer_sync(psync, prm.psync, "peer_shard_step")) return rc;
    if (psync && (psync->world != fan->world || psync->rank != fan->rank))
        return set_error(B200_EINVAL, "peer_shard_step: fan and psync disagree on world / rank");
    if (int rc = fill_adam_consts("peer_shard_step", groups, n_groups, prm.grp)) return rc;
    prm.inv_scale = inv_scale;
    prm.dev_inv_scale = dev_inv_scale;
    prm.parity = dev_parity;
    prm.flag = dev_flag;
    prm.sync = to_sync(sync);
    prm.sync.read_only = 1;
    if (n_tiles <= 0) return set_error(B200_EINVAL, "peer_shard_step: the shard must hold at least one tile");
    const unsigned cap = static_cast<unsigned>(g_peer_max_ctas.load());
    // with both multicast paths on, a thread issues ONE load and ONE store per group whatever the rank count: the
    // 2-rank instantiation (fewest registers, 2 CTAs/SM) serves every world size (ragged elements loop over `world`)
    const int w = (prm.fan.use_mc && prm.fan.use_mc_grad) ? 2 : fan->world;
#define B200_TRY(G_, W_)                                                                                              \
    if (grad_dtype == G_ && state_dtype == B200_F32 && copy_dtype == G_ && w <= W_)                                    \
        return launch_tiles<ShardStepOp<G_, B200_F32, G_, W_>>("peer_shard_step", dev_table, n_tiles, prm, stream, cap);
    B200_TRY(B200_F16, 2)
    B200_TRY(B200_F16, 4)
    B200_TRY(B200_F16, 8)
    B200_TRY(B200_BF16, 2)
    B200_TRY(B200_BF16, 4)
    B200_TRY(B200_BF16, 8)
#undef B200_TRY
    return set_error(B200_EDTYPE, "peer_shard_step: (grad,state,copy) dtypes (%d,%d,%d) not instantiated (16-bit gradients "
                     "and model, fp32 state)", grad_dtype, state_dtype, copy_dtype);
}

int32_t b200_peer_shard_repair(const void* dev_table, int64_t n_tiles, int32_t state_dtype, int32_t copy_dtype,
                               const int32_t* dev_found_inf, const int32_t* dev_parity, const b200_peer_fan* fan,
                               const b200_peer_sync* psync, void* stream) {
    if (!dev_found_inf || !dev_parity) return set_error(B200_EINVAL, "peer_shard_repair: NULL argument");
    ShardRepairParams prm;
    memset(&prm, 0, sizeof(prm));
    if (int rc = to_fan(fan, prm.fan, "peer_shard_repair")) return rc;
    if (int rc = to_peer_sync(psync, prm.psync, "peer_shard_repair")) return rc;
    prm.found_inf = dev_found_inf;
    prm.parity = dev_parity;
    if (n_tiles <= 0) return set_error(B200_EINVAL, "peer_shard_repair: the shard must hold at least one tile");
    if (state_dtype != B200_F32 || (copy_dtype != B200_F16 && copy_dtype != B200_BF16))
        return set_error(B200_EDTYPE, "peer_shard_repair: fp32 state and a 16-bit model copy expected");
    const unsigned cap = static_cast<unsigned>(g_peer_max_ctas.load());
    int rc = 0;
    if (copy_dtype == B200_F16)
        rc = launch_tiles<ShardRepairOp<B200_F32, B200_F16, 8>>("peer_shard_repair", dev_table, n_tiles, prm, stream, cap);
    else
        rc = launch_tiles<ShardRepairOp<B200_F32, B200_BF16, 8>>("peer_shard_repair", dev_table, n_tiles, prm, stream, cap);
    if (rc || !psync) return rc;
    peer_wait_done_if_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(prm.psync, dev_found_inf);
    return after_launch("peer_shard_repair(wait)");
}

int32_t b200_peer_wait_flags(const b200_peer_sync* sync, int32_t* dev_flag, void* stream) {
    if (!sync) return set_error(B200_EINVAL, "peer_wait_flags: sync is NULL");
    PeerSync ps;
    if (int rc = to_peer_sync(sync, ps, "peer_wait_flags")) return rc;
    peer_wait_flags_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(ps, dev_flag);
    return after_launch("peer_wait_flags");
}

int32_t b200_peer_reduce_shard(const void* const* peer_bufs, int32_t world, void* reduced, int64_t begin, int64_t end,
                               int32_t dtype, float factor, int32_t* dev_flag, const b200_peer_sync* sync, void* stream) {
    PeerSync ps;
    if (int rc = to_peer_sync(sync, ps, "peer_reduce_shard")) return rc;
    if (sync && sync->world != world) return set_error(B200_EINVAL, "peer_reduce_shard: sync->world != world");
    if (world < 1 || world > B200_MAX_PEERS) return set_error(B200_EINVAL, "peer_reduce_shard: world must be 1..%d", B200_MAX_PEERS);
    if (!peer_bufs || !reduced) return set_error(B200_EINVAL, "peer_reduce_shard: NULL buffer");
    if (begin < 0 || end < begin) return set_error(B200_EINVAL, "peer_reduce_shard: bad shard [%lld, %lld)",
                                                   static_cast<long long>(begin), static_cast<long long>(end));
    if (!valid_dt(dtype)) return set_error(B200_EDTYPE, "peer_reduce_shard: bad dtype");
    if (end == begin && !sync) return 0;              // with sync even an empty shard must take part in the hand-shake
    PeerBufs pb;
    for (int r = 0; r < B200_MAX_PEERS; r++) pb.p[r] = r < world ? peer_bufs[r] : nullptr;
    for (int r = 0; r < world; r++)
        if (!pb.p[r]) return set_error(B200_EINVAL, "peer_reduce_shard: peer buffer %d is NULL", r);
    constexpr int kThreads = 256;
    const int64_t groups = (end - begin) / kVec + 1;
    const int64_t want = (groups + kThreads - 1) / kThreads;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
#define B200_LAUNCH_REDUCE(W_)                                                                                    \
    {                                                                                                             \
        auto kernel = peer_reduce_kernel<DT, kThreads, W_>;                                                       \
        static int occ_dev[64];                                                                                   \
        int& occ = occ_dev[current_device_slot()];                                                                \
        if (occ == 0) occ = resident_ctas_per_sm(kernel, kThreads);                                               \
        int64_t blocks = static_cast<int64_t>(sm_count()) * occ;                                                  \
        if (blocks > want) blocks = want;                                                                         \
        const int cap = g_peer_max_ctas.load();                                                                   \
        if (cap > 0 && blocks > cap) blocks = cap;                                                                \
        if (blocks < 1) blocks = 1;                                                                               \
        kernel<<<static_cast<unsigned>(blocks), kThreads, 0, st>>>(pb, world, reduced, begin, end, factor, dev_flag, ps); \
    }
    B200_DISPATCH_DT(dtype, DT, {
        if (world <= 2) B200_LAUNCH_REDUCE(2)
        else if (world <= 4) B200_LAUNCH_REDUCE(4)
        else B200_LAUNCH_REDUCE(8)
    })
#undef B200_LAUNCH_REDUCE
    return after_launch("peer_reduce_shard");
}

// ------------------------------------------------------------------------------------------------
// b200_peer_spec_step: gradient mean + speculative step in one kernel per rank (see peer_spec_step_kernel)
// ------------------------------------------------------------------------------------------------
}  // extern "C" (templates need C++ linkage)

template <int G, int S, int C>
static int launch_peer_spec_step(const void* dev_table, int64_t n_tiles, const PeerStepParams& in, int32_t n_reducers,
                                 cudaStream_t st) {
    constexpr int kThreads = 256;
    PeerStepParams prm = in;
    // a reducer CTA keeps n_stages pieces (world x 4 or 8 KB each) in flight; two CTAs per SM share 227 KB
    const size_t stage = static_cast<size_t>(prm.world <= 2 ? 2 : prm.world <= 4 ? 4 : 8) * kPieceElems * sizeof(typename Elem<G>::type);
    prm.n_stages = static_cast<int>(96 * 1024 / stage);
    static const int stage_cap = getenv("B200_STEP_STAGES") ? atoi(getenv("B200_STEP_STAGES")) : 6;   // tuning knob
    if (prm.n_stages > stage_cap) prm.n_stages = stage_cap;
    if (prm.n_stages < 2) prm.n_stages = 2;
    const size_t smem = stage * prm.n_stages;
#define B200_LAUNCH_STEP(W_)                                                                              \
    {                                                                                                     \
        auto kernel = peer_spec_step_kernel<G, S, C, kThreads, W_>;                                       \
        static int occ_dev[64];                                                                           \
        int& occ = occ_dev[current_device_slot()];                                                        \
        if (occ == 0) {                                                                                   \
            cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)); \
            if (e != cudaSuccess) return cuda_fail(e, "peer_spec_step: cudaFuncSetAttribute");            \
            int n = 0;                                                                                    \
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kThreads, smem) != cudaSuccess || n < 1) { \
                cudaGetLastError();                                                                       \
                n = 1;                                                                                    \
            }                                                                                             \
            occ = n;                                                                                      \
        }                                                                                                 \
        int64_t blocks = static_cast<int64_t>(sm_count()) * occ;                                          \
        const int cap = g_peer_max_ctas.load();                                                           \
        if (cap > 0 && blocks > cap) blocks = cap;                                                        \
        if (blocks > n_tiles + n_reducers) blocks = n_tiles + n_reducers;                                 \
        if (blocks < 2) blocks = 2;                                                                       \
        prm.n_reducers = n_reducers < blocks - 1 ? n_reducers : static_cast<int>(blocks - 1);            \
        if (prm.n_reducers < 1) prm.n_reducers = 1;                                                       \
        kernel<<<static_cast<unsigned>(blocks), kThreads, smem, st>>>(dev_table, prm);                    \
    }
    if (prm.world <= 2) B200_LAUNCH_STEP(2)
    else if (prm.world <= 4) B200_LAUNCH_STEP(4)
    else B200_LAUNCH_STEP(8)
#undef B200_LAUNCH_STEP
    return after_launch("peer_spec_step");
}

extern "C" {

int32_t b200_peer_spec_step(const void* dev_table, int64_t n_tiles, int32_t grad_dtype, int32_t state_dtype,
                            int32_t copy_dtype, const b200_adam_group* groups, int32_t n_groups, float inv_scale,
                            const float* dev_inv_scale, int32_t* dev_flag, const int32_t* dev_parity,
                            const b200_step_sync* sync, const b200_peer_pull* pull, const void* const* peer_grads,
                            void* reduced, float factor, const b200_peer_sync* psync, int32_t n_reducers,
                            int64_t chunk_elems, void* dev_scratch, int64_t scratch_bytes, void* stream) {
    if (!dev_table || n_tiles <= 0) return set_error(B200_EINVAL, "peer_spec_step: empty table");
    if (!dev_parity || !dev_flag || !pull || !peer_grads || !reduced || !psync || !dev_scratch)
        return set_error(B200_EINVAL, "peer_spec_step: NULL argument");
    PeerStepParams prm;
    memset(&prm, 0, sizeof(prm));
    if (int rc = to_pull(pull, prm.spec.pull, "peer_spec_step")) return rc;
    if (prm.spec.pull.world < 2) return set_error(B200_EINVAL, "peer_spec_step: needs world >= 2 (use b200_fused_adam_spec)");
    if (int rc = fill_adam_consts("peer_spec_step", groups, n_groups, prm.spec.grp)) return rc;
    if (int rc = to_peer_sync(psync, prm.sync, "peer_spec_step")) return rc;
    if (psync->world != prm.spec.pull.world) return set_error(B200_EINVAL, "peer_spec_step: sync->world != pull->world");
    if (chunk_elems < kRunElems || chunk_elems % kRunElems)
        return set_error(B200_EINVAL, "peer_spec_step: chunk_elems must be a multiple of %lld", kRunElems);
    if (n_reducers < 1) n_reducers = 1;
    prm.spec.pull.first_tile = 0;                        // the table is already ordered for this rank
    prm.spec.inv_scale = inv_scale;
    prm.spec.dev_inv_scale = dev_inv_scale;
    prm.spec.flag = dev_flag;
    prm.spec.parity = dev_parity;
    prm.spec.sync = to_sync(sync);
    prm.spec.sync.read_only = 1;
    prm.world = prm.spec.pull.world;
    prm.begin = prm.spec.pull.bounds[psync->rank];
    prm.end = prm.spec.pull.bounds[psync->rank + 1];
    prm.chunk_elems = chunk_elems;
    prm.factor = factor;
    prm.reduced = reduced;
    prm.scratch = static_cast<int32_t*>(dev_scratch);
    long long max_chunks = 0;
    for (int r = 0; r < prm.world; r++) {
        const long long c = (prm.spec.pull.bounds[r + 1] - prm.spec.pull.bounds[r] + chunk_elems - 1) / chunk_elems;
        if (c > max_chunks) max_chunks = c;
    }
    if (max_chunks > B200_PEER_STEP_MAX_CHUNKS)
        return set_error(B200_ESIZE, "peer_spec_step: %lld chunks per shard, the progress board holds %d (raise chunk_elems)",
                         max_chunks, B200_PEER_STEP_MAX_CHUNKS);
    if (scratch_bytes < static_cast<int64_t>(8 + max_chunks) * 4)
        return set_error(B200_ESIZE, "peer_spec_step: scratch must hold %lld bytes (zeroed once)", (8 + max_chunks) * 4);
    for (int r = 0; r < B200_MAX_PEERS; r++) prm.grads.p[r] = r < prm.world ? peer_grads[r] : nullptr;
    for (int r = 0; r < prm.world; r++)
        if (!prm.grads.p[r]) return set_error(B200_EINVAL, "peer_spec_step: gradient buffer %d is NULL", r);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (grad_dtype == B200_F16 && state_dtype == B200_F32 && copy_dtype == B200_F16)
        return launch_peer_spec_step<B200_F16, B200_F32, B200_F16>(dev_table, n_tiles, prm, n_reducers, st);
    if (grad_dtype == B200_BF16 && state_dtype == B200_F32 && copy_dtype == B200_BF16)
        return launch_peer_spec_step<B200_BF16, B200_F32, B200_BF16>(dev_table, n_tiles, prm, n_reducers, st);
    if (grad_dtype == B200_F32 && state_dtype == B200_F32 && copy_dtype == B200_NONE)
        return launch_peer_spec_step<B200_F32, B200_F32, B200_NONE>(dev_table, n_tiles, prm, n_reducers, st);
    return set_error(B200_EDTYPE, "peer_spec_step: (grad,state,copy) dtypes (%d,%d,%d) not instantiated", grad_dtype,
                     state_dtype, copy_dtype);
}

int32_t b200_set_peer_grid(int32_t max_ctas) {
    if (max_ctas < 0) return set_error(B200_EINVAL, "set_peer_grid: max_ctas must be >= 0");
    g_peer_max_ctas.store(max_ctas);
    return 0;
}

int32_t b200_peer_gather(const void* const* peer_reduced, int32_t world, void* local_out, const int64_t* bounds, int32_t dtype,
                         int32_t* dev_flag, const b200_peer_sync* sync, void* stream) {
    PeerSync ps;
    if (int rc = to_peer_sync(sync, ps, "peer_gather")) return rc;
    if (world < 1 || world > B200_MAX_PEERS || !peer_reduced || !local_out || !bounds)
        return set_error(B200_EINVAL, "peer_gather: bad argument");
    if (!valid_dt(dtype)) return set_error(B200_EDTYPE, "peer_gather: bad dtype");
    PeerBufs pb;
    ShardBounds sb;
    for (int r = 0; r < B200_MAX_PEERS; r++) pb.p[r] = r < world ? peer_reduced[r] : nullptr;
    for (int r = 0; r <= B200_MAX_PEERS; r++) sb.b[r] = r <= world ? bounds[r] : bounds[world];
    if (sb.b[0] != 0) return set_error(B200_EINVAL, "peer_gather: bounds[0] must be 0");
    for (int r = 0; r < world; r++) {
        if (!pb.p[r]) return set_error(B200_EINVAL, "peer_gather: peer buffer %d is NULL", r);
        if (sb.b[r] > sb.b[r + 1]) return set_error(B200_EINVAL, "peer_gather: bounds must be non-decreasing");
        // a shard boundary inside the buffer must not cut an 8-element group; boundaries equal to the total
        // (empty trailing shards of a small bucket) are fine
        if (r > 0 && sb.b[r] % kVec && sb.b[r] != sb.b[world])
            return set_error(B200_EINVAL, "peer_gather: bounds[%d] = %lld is inside the buffer and not a multiple of 8", r,
                             static_cast<long long>(sb.b[r]));
    }
    constexpr int kThreads = 256;
    const int64_t groups = sb.b[world] / kVec + 1;
    const int64_t want = (groups + kThreads * 4 - 1) / (kThreads * 4);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    B200_DISPATCH_DT(dtype, DT, {
        auto kernel = peer_gather_kernel<DT, kThreads>;
        static int occ_dev[64];
        int& occ = occ_dev[current_device_slot()];
        if (occ == 0) occ = resident_ctas_per_sm(kernel, kThreads);
        int64_t blocks = static_cast<int64_t>(sm_count()) * occ;
        if (blocks > want) blocks = want;
        const int cap = g_peer_max_ctas.load();
        if (cap > 0 && blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        kernel<<<static_cast<unsigned>(blocks), kThreads, 0, st>>>(pb, world, local_out, sb, dev_flag, ps);
    })
    return after_launch("peer_gather");
}

// ------------------------------------------------------------------------------------------------
// b200_fused_adam_noalloc: the Adam step for callers that cannot allocate (see the header).
// The tensor-list table — the one piece of global memory this path needs beyond the caller's own
// buffers — lives in a static device array instead of caller-allocated memory.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_table[B200_NOALLOC_TABLE_BYTES];

// In host code `g_noalloc_table` is only a handle for the runtime, not an address the GPU can
// use; the real device address is looked up once per device (with that device current).
static void* g_noalloc_table_addr[64];
static unsigned char g_noalloc_host_image[B200_NOALLOC_TABLE_BYTES];
static std::mutex g_noalloc_mutex;

static int noalloc_table_address(void** out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "fused_adam_noalloc: cudaGetDevice");
    if (dev < 0 || dev >= 64) return set_error(B200_ENODEV, "fused_adam_noalloc: device index %d not supported", dev);
    if (g_noalloc_table_addr[dev] == nullptr) {
        void* ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_noalloc_table);
        if (e != cudaSuccess) return cuda_fail(e, "fused_adam_noalloc: cudaGetSymbolAddress");
        if (ptr == nullptr) return set_error(B200_ENODEV, "fused_adam_noalloc: static table has no device address");
        g_noalloc_table_addr[dev] = ptr;
    }
    *out = g_noalloc_table_addr[dev];
    return 0;
}

int32_t b200_fused_adam_noalloc(int32_t n_tensors, const void* const* ptrs, const int64_t* numel, int32_t grad_dtype,
                                int32_t state_dtype, int32_t copy_dtype, const b200_adam_group* group, float inv_scale,
                                const float* dev_inv_scale, const int32_t* dev_found_inf, void* stream) {
    if (n_tensors < 0 || (n_tensors > 0 && (!ptrs || !numel)) || !group)
        return set_error(B200_EINVAL, "fused_adam_noalloc: bad argument");
    if (!valid_dt(grad_dtype) || !valid_dt(state_dtype) || (copy_dtype != B200_NONE && !valid_dt(copy_dtype)))
        return set_error(B200_EDTYPE, "fused_adam_noalloc: bad dtype");
    if (n_tensors == 0) return 0;
    const int64_t need = b200_table_bytes(n_tensors, numel, B200_TILE_ELEMS_DEFAULT);
    if (need < 0) return static_cast<int32_t>(need);
    if (need > B200_NOALLOC_TABLE_BYTES)
        return set_error(B200_ESIZE,
                         "fused_adam_noalloc: this tensor list needs a %lld-byte table but the built-in static table "
                         "holds %d bytes; use b200_fused_adam with a caller-allocated table instead",
                         static_cast<long long>(need), B200_NOALLOC_TABLE_BYTES);
    // every column the kernel dereferences must be there (b200_table_build itself accepts NULL as
    // "column absent", which the Adam kernel has no use for)
    const int n_cols = copy_dtype == B200_NONE ? 4 : 5;
    for (int d = 0; d < n_cols; d++)
        for (int t = 0; t < n_tensors; t++)
            if (numel[t] > 0 && ptrs[static_cast<size_t>(d) * n_tensors + t] == nullptr)
                return set_error(B200_EINVAL, "fused_adam_noalloc: ptrs[%d * n_tensors + %d] (column %d of tensor %d) is NULL",
                                 d, t, d, t);
    const int depth = copy_dtype == B200_NONE ? 4 : 5;
    const int g_sz = grad_dtype == B200_F32 ? 4 : 2, s_sz = state_dtype == B200_F32 ? 4 : 2;
    const int32_t elem_size[5] = {g_sz, s_sz, s_sz, s_sz, copy_dtype == B200_F32 ? 4 : 2};

    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    void* dev_table = nullptr;
    int rc = noalloc_table_address(&dev_table);
    if (rc != 0) return rc;
    const int64_t n_tiles = b200_table_build(g_noalloc_host_image, B200_NOALLOC_TABLE_BYTES, n_tensors, depth, ptrs, elem_size,
                                             numel, nullptr, B200_TILE_ELEMS_DEFAULT);
    if (n_tiles < 0) return static_cast<int32_t>(n_tiles);
    // The source is ordinary (pageable) host memory: in a normal, non-captured stream the runtime
    // stages it before cudaMemcpyAsync returns, so the next call (serialised by the mutex above) may
    // overwrite the static host buffer; the device-side copy is ordered on `stream` ahead of the kernel.
    cudaError_t e = cudaMemcpyAsync(dev_table, g_noalloc_host_image, static_cast<size_t>(need), cudaMemcpyHostToDevice,
                                    static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "fused_adam_noalloc: table upload");
    return b200_fused_adam(dev_table, n_tiles, grad_dtype, state_dtype, copy_dtype, group, 1, inv_scale, dev_inv_scale,
                           dev_found_inf, nullptr, stream);
}

int32_t b200_bucket_epilogue(void* dev_bucket, int64_t numel, int32_t dtype, float factor, int32_t* dev_flag, void* stream) {
    if (numel < 0) return set_error(B200_EINVAL, "bucket_epilogue: negative numel");
    if (numel == 0) return 0;
    if (!dev_bucket) return set_error(B200_EINVAL, "bucket_epilogue: dev_bucket is NULL");
    if (!valid_dt(dtype)) return set_error(B200_EDTYPE, "bucket_epilogue: bad dtype");
    constexpr int kThreads = 256;
    const int read_only = factor == 1.0f;
    const int64_t groups = numel / kVec + 1;
    const int64_t want = (groups + kThreads * 4 - 1) / (kThreads * 4);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    B200_DISPATCH_DT(dtype, DT, {
        auto kernel = bucket_epilogue_kernel<DT, kThreads>;
        static int occ_dev[64];
        int& occ = occ_dev[current_device_slot()];
        if (occ == 0) occ = resident_ctas_per_sm(kernel, kThreads);
        int64_t blocks = static_cast<int64_t>(sm_count()) * occ;      // one resident wave, grid-stride inside
        if (blocks > want) blocks = want;
        if (blocks < 1) blocks = 1;
        kernel<<<static_cast<unsigned>(blocks), kThreads, 0, st>>>(dev_bucket, numel, factor, read_only, dev_flag);
    })
    return after_launch("bucket_epilogue");
}

}  // extern "C"
