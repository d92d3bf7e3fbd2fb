// b200_common.cuh — element traits, 128/256-bit global memory access and the tile walker
// shared by every kernel of libapex_b200.so.  sm_100a only.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "apex_b200.h"

namespace b200 {

constexpr int kVec = 8;            // elements one thread moves per memory instruction group

// ------------------------------------------------------------------------------------------
// Element conversions.  All arithmetic is done in fp32; storage types are widened on load and
// rounded to nearest-even on store (same as Tensor.copy_ and oracle/amp_step_oracle.c).
// ------------------------------------------------------------------------------------------
template <int DT> struct Elem;
template <> struct Elem<B200_F32> {
    using type = float;
    static __device__ __forceinline__ float to_float(float x) { return x; }
    static __device__ __forceinline__ float from_float(float x) { return x; }
};
template <> struct Elem<B200_F16> {
    using type = __half;
    static __device__ __forceinline__ float to_float(__half x) { return __half2float(x); }
    static __device__ __forceinline__ __half from_float(float x) { return __float2half_rn(x); }
};
template <> struct Elem<B200_BF16> {
    using type = __nv_bfloat16;
    static __device__ __forceinline__ float to_float(__nv_bfloat16 x) { return __bfloat162float(x); }
    static __device__ __forceinline__ __nv_bfloat16 from_float(float x) { return __float2bfloat16_rn(x); }
};

// value a later reader of a DT-typed array would see for the fp32 value x
template <int DT> __device__ __forceinline__ float round_through(float x) {
    return Elem<DT>::to_float(Elem<DT>::from_float(x));
}

__device__ __forceinline__ bool is_finite(float x) {
    return (__float_as_uint(x) & 0x7f800000u) != 0x7f800000u;
}

// ------------------------------------------------------------------------------------------
// 8-element vector access.  fp32 columns use Blackwell's 256-bit LDG/STG (ld.global.v8.f32,
// new with sm_100); fp16/bf16 columns use 128-bit.  Everything here streams exactly once, so
// loads bypass L1 (L1::no_allocate); read-only columns additionally take the .nc path.
// Requirements: address aligned to 32 B (fp32) / 16 B (16-bit types) — the table's row flag
// guarantees it, otherwise the scalar path is taken.
// ------------------------------------------------------------------------------------------
template <bool kReadOnly>
__device__ __forceinline__ void ldg256(const void* p, uint32_t (&r)[8]) {
    if constexpr (kReadOnly) {
        asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                     : "l"(p));
    } else {
        asm volatile("ld.global.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                     : "l"(p));
    }
}

template <bool kReadOnly>
__device__ __forceinline__ void ldg128(const void* p, uint32_t (&r)[4]) {
    if constexpr (kReadOnly) {
        asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "l"(p));
    } else {
        asm volatile("ld.global.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "l"(p));
    }
}

__device__ __forceinline__ void stg256(void* p, const uint32_t (&r)[8]) {
    asm volatile("st.global.L1::no_allocate.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(p), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
}

__device__ __forceinline__ void stg128(void* p, const uint32_t (&r)[4]) {
    asm volatile("st.global.L1::no_allocate.v4.b32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]) : "memory");
}

// raw 8-element register image of a column slice
template <int DT> struct Raw8 { uint32_t r[DT == B200_F32 ? 8 : 4]; };

template <int DT, bool kReadOnly>
__device__ __forceinline__ void load8(const void* base, int64_t idx, Raw8<DT>& raw) {
    using T = typename Elem<DT>::type;
    const T* p = reinterpret_cast<const T*>(base) + idx;
    if constexpr (DT == B200_F32) ldg256<kReadOnly>(p, raw.r);
    else ldg128<kReadOnly>(p, raw.r);
}

template <int DT>
__device__ __forceinline__ void unpack8(const Raw8<DT>& raw, float (&f)[kVec]) {
    if constexpr (DT == B200_F32) {
#pragma unroll
        for (int i = 0; i < 8; i++) f[i] = __uint_as_float(raw.r[i]);
    } else if constexpr (DT == B200_F16) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            __half2 h = *reinterpret_cast<const __half2*>(&raw.r[i]);
            f[2 * i] = __low2float(h);
            f[2 * i + 1] = __high2float(h);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 4; i++) {   // bf16 -> f32 is a 16-bit shift
            f[2 * i] = __uint_as_float(raw.r[i] << 16);
            f[2 * i + 1] = __uint_as_float(raw.r[i] & 0xffff0000u);
        }
    }
}

template <int DT>
__device__ __forceinline__ void store8(void* base, int64_t idx, const float (&f)[kVec]) {
    using T = typename Elem<DT>::type;
    T* p = reinterpret_cast<T*>(base) + idx;
    if constexpr (DT == B200_F32) {
        uint32_t r[8];
#pragma unroll
        for (int i = 0; i < 8; i++) r[i] = __float_as_uint(f[i]);
        stg256(p, r);
    } else if constexpr (DT == B200_F16) {
        uint32_t r[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            __half2 h = __halves2half2(__float2half_rn(f[2 * i]), __float2half_rn(f[2 * i + 1]));
            r[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        stg128(p, r);
    } else {
        uint32_t r[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            __nv_bfloat162 h = __halves2bfloat162(__float2bfloat16_rn(f[2 * i]), __float2bfloat16_rn(f[2 * i + 1]));
            r[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        stg128(p, r);
    }
}

// 8 floats -> 8 sixteen-bit values (16 bytes), rounded like store8
template <int DT>
__device__ __forceinline__ void pack8_16(const float (&f)[kVec], uint32_t (&r)[4]) {
    static_assert(DT == B200_F16 || DT == B200_BF16, "16-bit types only");
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if constexpr (DT == B200_F16) {
            __half2 h = __halves2half2(__float2half_rn(f[2 * i]), __float2half_rn(f[2 * i + 1]));
            r[i] = *reinterpret_cast<uint32_t*>(&h);
        } else {
            __nv_bfloat162 h = __halves2bfloat162(__float2bfloat16_rn(f[2 * i]), __float2bfloat16_rn(f[2 * i + 1]));
            r[i] = *reinterpret_cast<uint32_t*>(&h);
        }
    }
}

// One 16-byte store through a MULTICAST address: the switch writes it into the same offset of every GPU bound to
// the multicast object (this GPU included).  Stores only — no load, and nothing TMA, ever goes through this address.
__device__ __forceinline__ void multimem_st128(void* mc_addr, const uint32_t (&r)[4]) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 :: "l"(mc_addr), "f"(__uint_as_float(r[0])), "f"(__uint_as_float(r[1])), "f"(__uint_as_float(r[2])),
                    "f"(__uint_as_float(r[3]))
                 : "memory");
}

// Eight 16-bit values summed over every GPU bound to the multicast object, inside the switch (.acc::f32; the order of the
// addends and the precision of the switch's partial sums are not specified: a few 16-bit ulps from a rank-ordered sum).  A plain (non-TMA) load instruction.
template <int DT>
__device__ __forceinline__ void multimem_ld_reduce128(const void* mc_addr, uint32_t (&r)[4]) {
    static_assert(DT == B200_F16 || DT == B200_BF16, "16-bit types only");
    if constexpr (DT == B200_F16)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "l"(mc_addr) : "memory");
    else
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "l"(mc_addr) : "memory");
}

template <int DT>
__device__ __forceinline__ float load1(const void* base, int64_t idx) {
    using T = typename Elem<DT>::type;
    return Elem<DT>::to_float(reinterpret_cast<const T*>(base)[idx]);
}

template <int DT>
__device__ __forceinline__ void store1(void* base, int64_t idx, float x) {
    using T = typename Elem<DT>::type;
    reinterpret_cast<T*>(base)[idx] = Elem<DT>::from_float(x);
}

// ------------------------------------------------------------------------------------------
// Table access
// ------------------------------------------------------------------------------------------
struct TableView {
    const b200_table_row* rows;
    const b200_table_tile* tiles;
    uint32_t n_tiles;
    uint32_t tile_elems;
};

__device__ __forceinline__ TableView open_table(const void* table) {
    const b200_table_header* h = reinterpret_cast<const b200_table_header*>(table);
    const char* base = reinterpret_cast<const char*>(table);
    TableView v;
    v.rows = reinterpret_cast<const b200_table_row*>(base + __ldg(&h->rows_off));
    v.tiles = reinterpret_cast<const b200_table_tile*>(base + __ldg(&h->tiles_off));
    v.n_tiles = __ldg(&h->n_tiles);
    v.tile_elems = __ldg(&h->tile_elems);
    return v;
}

// ------------------------------------------------------------------------------------------
// Step window (b200_step_sync): which of the host's candidate step numbers is the true one,
// and the {calls, skips} bookkeeping that lets the host catch up without synchronising.
// ------------------------------------------------------------------------------------------
struct StepSync {
    int32_t* counters;
    int32_t* mirror;
    int32_t known_skips;
    int32_t read_only;
};

// returns true when the whole step must be skipped; `win` = candidate index otherwise
__device__ __forceinline__ bool step_prologue(const int32_t* found_inf, const StepSync& ss, int& win) {
    win = 0;
    if (ss.counters) {
        // counters[1] only changes in launches that skip (where `win` is not used), so every CTA of
        // a stepping launch reads the same value no matter when it starts
        const int d = *(reinterpret_cast<const volatile int32_t*>(ss.counters) + 1) - ss.known_skips;
        win = d < 0 ? 0 : (d > B200_STEP_WINDOW - 1 ? B200_STEP_WINDOW - 1 : d);
    }
    return found_inf != nullptr && *reinterpret_cast<const volatile int32_t*>(found_inf) != 0;
}

__device__ __forceinline__ void step_epilogue(const StepSync& ss, bool skipped) {
    if (ss.counters && !ss.read_only && blockIdx.x == 0 && threadIdx.x == 0) {
        const int calls = ss.counters[0] + 1;
        const int skips = ss.counters[1] + (skipped ? 1 : 0);
        if (skipped) ss.counters[1] = skips;
        ss.counters[0] = calls;
        if (ss.mirror) {
            volatile int32_t* m = ss.mirror;
            m[1] = skips;
            __threadfence_system();
            m[0] = calls;
        }
    }
}

// ------------------------------------------------------------------------------------------
// Peer pull (b200_peer_pull): where a tile's gradient lives when the mean was reduce-scattered.
// ------------------------------------------------------------------------------------------
struct PeerPull {
    const char* local_base;
    long long bounds[B200_MAX_PEERS + 1];
    long long delta[B200_MAX_PEERS];
    int world;
    int elem_size;
    unsigned first_tile;
};

// byte offset to add to a column-0 pointer for the tile that starts at `tile_ptr`
__device__ __forceinline__ long long peer_delta(const PeerPull& pp, const void* tile_ptr) {
    if (pp.world <= 1) return 0;
    const long long flat = (reinterpret_cast<const char*>(tile_ptr) - pp.local_base) / pp.elem_size;
    int owner = 0;
#pragma unroll
    for (int k = 1; k < B200_MAX_PEERS; k++) owner += (k < pp.world && flat >= pp.bounds[k]) ? 1 : 0;
    return pp.delta[owner];
}

// Raise a 0/1 flag for the whole grid: one plain store per warp that saw something.  Every
// writer stores the same value, so no atomic is needed.
__device__ __forceinline__ void raise_flag(int32_t* flag, bool bad) {
    if (__any_sync(0xffffffffu, bad)) {
        if ((threadIdx.x & 31) == 0) *flag = 1;
    }
}

// ------------------------------------------------------------------------------------------
// The tile walker.  A persistent grid strides over the tile list; for each tile the CTA runs
// Op's vector body over the 8-element groups and its scalar body over the ragged tail (or
// over the whole tile when the row is not aligned for vector access).
//
// Op interface:
//   using Params;                                      // kernel-argument payload
//   struct Ctx;  static Ctx make_ctx(params, row, win, start) // per-tile constants (group lookup…);
//                                                      // win = step-window candidate (see b200_step_sync)
//   template <int U> static void vec(row, ctx, idx[U], active[U], bad&)   // U groups in flight
//   static void scalar(row, ctx, idx, bad&)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    const uint32_t s = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" :: "r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async8(volatile void* smem_dst, const void* gmem_src) {
    const uint32_t s = static_cast<uint32_t>(__cvta_generic_to_shared(const_cast<void*>(smem_dst)));
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"(s), "l"(gmem_src) : "memory");
}
// ---- mbarrier + bulk asynchronous copy (TMA without a tensor map, no multicast): one thread moves KBs ----
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
    const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(bar));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(a), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(bar));
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(a), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(bar));
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" :: "r"(a), "r"(parity) : "memory");
}
// global -> this CTA's shared memory, `bytes` a multiple of 16, both addresses 16-byte aligned; completion is
// signalled on `bar` as transaction bytes
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, unsigned bytes, uint64_t* bar) {
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
    const uint32_t b = static_cast<uint32_t>(__cvta_generic_to_shared(bar));
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(d), "l"(gmem_src), "r"(bytes), "r"(b) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
// wait until at most `pending` of the committed groups are still in flight (pending < 12)
__device__ __forceinline__ void cp_async_wait_pending(int pending, int /*max*/) {
    switch (pending) {
    case 0: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
    case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
    case 2: asm volatile("cp.async.wait_group 2;" ::: "memory"); break;
    case 3: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
    case 4: asm volatile("cp.async.wait_group 4;" ::: "memory"); break;
    case 5: asm volatile("cp.async.wait_group 5;" ::: "memory"); break;
    case 6: asm volatile("cp.async.wait_group 6;" ::: "memory"); break;
    case 7: asm volatile("cp.async.wait_group 7;" ::: "memory"); break;
    case 8: asm volatile("cp.async.wait_group 8;" ::: "memory"); break;
    case 9: asm volatile("cp.async.wait_group 9;" ::: "memory"); break;
    case 10: asm volatile("cp.async.wait_group 10;" ::: "memory"); break;
    default: asm volatile("cp.async.wait_group 11;" ::: "memory"); break;
    }
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// The descriptor chain (tile entry -> row -> data) is three dependent memory latencies.  To keep
// it off the critical path, the row of the NEXT tile is fetched with cp.async into a two-slot
// shared-memory ring by four threads while the current tile streams, and the tile entry after
// that is already sitting in a register; one __syncthreads per tile hands the slot over.
struct RowSlot {
    b200_table_row row;      // 64 B, filled by 4 x cp.async(16 B)
};

template <class Op, int kThreads, int kUnroll, class Acc>
__device__ __forceinline__ void walk_tiles(const void* table, const typename Op::Params& prm, int win, uint32_t rot,
                                           Acc& bad) {
    __shared__ __align__(16) RowSlot s_slot[2];
    __shared__ uint32_t s_tile[2];
    const TableView tv = open_table(table);
    const uint32_t G = gridDim.x;
    uint32_t t = blockIdx.x;
    if (t >= tv.n_tiles) return;
    const bool loader = threadIdx.x < 4;
    // `rot` rotates the traversal: logical tile i is list entry (i + rot) mod n_tiles.  The peer path
    // starts every rank at its OWN shard, so at any moment the ranks pull from different owners.
    if (rot >= tv.n_tiles) rot = 0;
    auto entry = [&](uint32_t i) {
        uint32_t j = i + rot;
        if (j >= tv.n_tiles) j -= tv.n_tiles;
        return __ldg(reinterpret_cast<const uint2*>(tv.tiles + j));
    };
    uint2 tl_next = make_uint2(0u, 0u);          // tile entry of iteration i+1 (loader threads only)
    if (loader) {
        const uint2 tl0 = entry(t);
        cp_async16(reinterpret_cast<char*>(&s_slot[0].row) + threadIdx.x * 16,
                   reinterpret_cast<const char*>(tv.rows + tl0.x) + threadIdx.x * 16);
        if (threadIdx.x == 0) s_tile[0] = tl0.y;
        if (t + G < tv.n_tiles) tl_next = entry(t + G);
        cp_async_wait_all();
    }
    __syncthreads();
    int slot = 0;
    for (;;) {
        const bool has_next = t + G < tv.n_tiles;
        if (loader && has_next) {
            cp_async16(reinterpret_cast<char*>(&s_slot[slot ^ 1].row) + threadIdx.x * 16,
                       reinterpret_cast<const char*>(tv.rows + tl_next.x) + threadIdx.x * 16);
            if (threadIdx.x == 0) s_tile[slot ^ 1] = tl_next.y;
            const uint32_t t2 = t + 2 * G;
            if (t2 < tv.n_tiles && t2 > t) tl_next = entry(t2);
        }
        const b200_table_row row = s_slot[slot].row;
        const uint32_t tile_idx = s_tile[slot];
        const int64_t start = static_cast<int64_t>(tile_idx) * tv.tile_elems;
        const int64_t left = row.numel - start;
        const int n = left < static_cast<int64_t>(tv.tile_elems) ? static_cast<int>(left)
                                                                   : static_cast<int>(tv.tile_elems);
        const typename Op::Ctx ctx = Op::make_ctx(prm, row, win, start);
        if (row.flags & 1u) {
            const int n8 = n & ~(kVec - 1);
            for (int base = threadIdx.x * kVec; base < n8; base += kThreads * kVec * kUnroll) {
                int64_t idx[kUnroll];
                bool act[kUnroll];
#pragma unroll
                for (int u = 0; u < kUnroll; u++) {
                    const int off = base + u * kThreads * kVec;
                    act[u] = off < n8;
                    idx[u] = start + off;
                }
                Op::template vec<kUnroll>(row, ctx, idx, act, bad);
            }
            const int tail = n - n8;
            if (static_cast<int>(threadIdx.x) < tail) Op::scalar(row, ctx, start + n8 + threadIdx.x, bad);
        } else {
            for (int i = threadIdx.x; i < n; i += kThreads) Op::scalar(row, ctx, start + i, bad);
        }
        if (!has_next) break;
        if (loader) cp_async_wait_all();
        __syncthreads();
        slot ^= 1;
        t += G;
    }
}

}  // namespace b200
