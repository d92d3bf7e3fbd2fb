// table.cc — S1, the host half of the chunk scheduler: turns a tensor list into the table image
// described in include/apex_b200.h.  Pure host code (no CUDA calls), so it is unit-tested on
// machines without a GPU.  Replaces the per-call packing loop of apex's
// csrc/multi_tensor_apply.cuh.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

#include "apex_b200.h"

namespace b200 {
int set_error(int code, const char* fmt, ...);
}

namespace {

constexpr uint32_t kMagic = 0x42325442u;
static_assert(sizeof(b200_table_header) == 64 && sizeof(b200_table_row) == 64 && sizeof(b200_table_tile) == 8 &&
                  sizeof(b200_scaler_state) == 48,
              "table / scaler layouts are part of the ABI");

inline int64_t tiles_of(int64_t numel, int32_t tile_elems) {
    return numel <= 0 ? 0 : (numel + tile_elems - 1) / tile_elems;
}

bool valid_tile(int32_t tile_elems) { return tile_elems >= 8 && tile_elems % 8 == 0 && tile_elems <= (1 << 24); }

}  // namespace

extern "C" {

int64_t b200_table_bytes(int32_t n_tensors, const int64_t* numel, int32_t tile_elems) {
    if (n_tensors < 0 || (n_tensors > 0 && !numel)) return b200::set_error(B200_EINVAL, "table_bytes: bad argument");
    if (!valid_tile(tile_elems)) return b200::set_error(B200_EINVAL, "table_bytes: tile_elems must be a multiple of 8");
    int64_t tiles = 0;
    for (int32_t t = 0; t < n_tensors; t++) {
        if (numel[t] < 0) return b200::set_error(B200_EINVAL, "table_bytes: negative numel at %d", t);
        tiles += tiles_of(numel[t], tile_elems);
    }
    if (tiles > 0xffffffffll) return b200::set_error(B200_ESIZE, "table_bytes: more than 2^32 tiles");
    return static_cast<int64_t>(sizeof(b200_table_header)) + static_cast<int64_t>(sizeof(b200_table_row)) * n_tensors +
           static_cast<int64_t>(sizeof(b200_table_tile)) * tiles;
}

int64_t b200_table_build(void* host_image, int64_t image_bytes, int32_t n_tensors, int32_t depth,
                         const void* const* ptrs, const int32_t* elem_size, const int64_t* numel,
                         const int32_t* group, int32_t tile_elems) {
    if (!host_image || depth < 1 || depth > B200_MAX_DEPTH || (n_tensors > 0 && (!ptrs || !numel || !elem_size)))
        return b200::set_error(B200_EINVAL, "table_build: bad argument");
    const int64_t need = b200_table_bytes(n_tensors, numel, tile_elems);
    if (need < 0) return need;
    if (image_bytes < need) return b200::set_error(B200_ESIZE, "table_build: image needs %lld bytes, got %lld",
                                                   static_cast<long long>(need), static_cast<long long>(image_bytes));
    for (int32_t d = 0; d < depth; d++)
        if (elem_size[d] != 2 && elem_size[d] != 4)
            return b200::set_error(B200_EINVAL, "table_build: elem_size[%d] must be 2 or 4", d);

    char* base = static_cast<char*>(host_image);
    b200_table_header* h = reinterpret_cast<b200_table_header*>(base);
    b200_table_row* rows = reinterpret_cast<b200_table_row*>(base + sizeof(b200_table_header));
    b200_table_tile* tiles = reinterpret_cast<b200_table_tile*>(base + sizeof(b200_table_header) +
                                                                sizeof(b200_table_row) * static_cast<size_t>(n_tensors));
    memset(h, 0, sizeof(*h));
    uint64_t total = 0;
    uint32_t n_tiles = 0;
    for (int32_t t = 0; t < n_tensors; t++) {
        b200_table_row& r = rows[t];
        memset(&r, 0, sizeof(r));
        bool aligned = true;
        for (int32_t d = 0; d < depth; d++) {
            const void* p = ptrs[static_cast<size_t>(d) * n_tensors + t];
            r.ptr[d] = const_cast<void*>(p);
            // 8 elements per access: 32 B for 4-byte types (256-bit LDG/STG), 16 B for 2-byte types
            const uintptr_t mask = elem_size[d] == 4 ? 31u : 15u;
            if (p && (reinterpret_cast<uintptr_t>(p) & mask)) aligned = false;
        }
        r.numel = numel[t];
        const int32_t g = group ? group[t] : 0;
        if (g < 0 || g >= B200_MAX_GROUPS) return b200::set_error(B200_EINVAL, "table_build: group %d out of range", g);
        r.group = static_cast<uint32_t>(g);
        r.flags = aligned ? 1u : 0u;
        r.reserved = static_cast<uint64_t>(t);          // the row's own index (per-tensor quantities: LAMB trust ratios)
        const int64_t k = tiles_of(numel[t], tile_elems);
        for (int64_t i = 0; i < k; i++) {
            tiles[n_tiles].tensor = static_cast<uint32_t>(t);
            tiles[n_tiles].tile = static_cast<uint32_t>(i);
            n_tiles++;
        }
        total += static_cast<uint64_t>(numel[t]);
    }
    h->magic = kMagic;
    h->version = B200_ABI_VERSION;
    h->n_tensors = static_cast<uint32_t>(n_tensors);
    h->depth = static_cast<uint32_t>(depth);
    h->tile_elems = static_cast<uint32_t>(tile_elems);
    h->n_tiles = n_tiles;
    h->rows_off = sizeof(b200_table_header);
    h->tiles_off = sizeof(b200_table_header) + sizeof(b200_table_row) * static_cast<uint64_t>(n_tensors);
    h->total_elems = total;
    h->image_bytes = static_cast<uint64_t>(need);
    return n_tiles;
}

// Tile order for b200_peer_spec_step.  The update role of that kernel consumes reduced gradient tiles while
// the ranks' reduce roles are still producing them, chunk by chunk, each over its own shard.  To never wait
// for data that is far from ready, the tiles are visited CHUNK-MAJOR: first the tiles whose gradients lie in
// chunk 0 of any shard, then chunk 1 of any shard, ...; inside a chunk index the shards are visited starting
// with this rank's own (so at any moment the ranks pull from different owners).  A tile that straddles two
// chunks is keyed on the later one.  Pure host code, stable, O(n_tiles log n_tiles).
int32_t b200_table_order_for_peer_step(void* host_image, const void* local_base, int32_t grad_elem_size,
                                       const int64_t* bounds, int32_t world, int32_t rank, int64_t chunk_elems) {
    if (!host_image || !local_base || !bounds || world < 1 || world > B200_MAX_PEERS || rank < 0 || rank >= world ||
        chunk_elems < 8 || (grad_elem_size != 2 && grad_elem_size != 4))
        return b200::set_error(B200_EINVAL, "table_order_for_peer_step: bad argument");
    char* base = static_cast<char*>(host_image);
    b200_table_header* h = reinterpret_cast<b200_table_header*>(base);
    if (h->magic != kMagic) return b200::set_error(B200_EINVAL, "table_order_for_peer_step: not a table image");
    const b200_table_row* rows = reinterpret_cast<const b200_table_row*>(base + h->rows_off);
    b200_table_tile* tiles = reinterpret_cast<b200_table_tile*>(base + h->tiles_off);
    struct Key {
        int64_t chunk, owner_rot, flat;
        b200_table_tile tile;
    };
    std::vector<Key> keys(h->n_tiles);
    for (uint32_t i = 0; i < h->n_tiles; i++) {
        const b200_table_row& r = rows[tiles[i].tensor];
        const int64_t start = static_cast<int64_t>(tiles[i].tile) * h->tile_elems;
        const int64_t n = std::min<int64_t>(h->tile_elems, r.numel - start);
        const int64_t flat = (static_cast<const char*>(r.ptr[0]) - static_cast<const char*>(local_base)) / grad_elem_size + start;
        if (flat < bounds[0] || flat + n > bounds[world])
            return b200::set_error(B200_EINVAL, "table_order_for_peer_step: tile %u lies outside the flat gradient buffer", i);
        int owner = 0;
        for (int k = 1; k < world; k++) owner += flat >= bounds[k] ? 1 : 0;
        if (flat + n > bounds[owner + 1])
            return b200::set_error(B200_EINVAL, "table_order_for_peer_step: tile %u straddles a shard boundary "
                                                "(use tile-aligned shards)", i);
        keys[i] = Key{(flat + n - 1 - bounds[owner]) / chunk_elems, (owner - rank + world) % world, flat, tiles[i]};
    }
    std::stable_sort(keys.begin(), keys.end(), [](const Key& a, const Key& b) {
        if (a.chunk != b.chunk) return a.chunk < b.chunk;
        if (a.owner_rot != b.owner_rot) return a.owner_rot < b.owner_rot;
        return a.flat < b.flat;
    });
    for (uint32_t i = 0; i < h->n_tiles; i++) tiles[i] = keys[i].tile;
    return 0;
}

}  // extern "C"
