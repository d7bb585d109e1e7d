// fb_layout.cuh -- private HBM layouts of the per-call state (geom / image / inst) and of the
// batched scoring workspace.  The reference keeps comparable state in three byte buffers
// (DGR/cuda_rasterizer/rasterizer_impl.h:30-70); those layouts are opaque to its callers
// (only round-tripped to backward), so this one is free to be B200-shaped:
//   * one 32-byte record per visible Gaussian (exactly one DRAM sector): {px,py,depth,opacity |
//     conic a,b,c, radius} on the drop-in path, {px,py,hx,hy | conic a,b,c, opacity} on the scoring
//     path (hx,hy = half extents of the alpha >= 1/255 footprint, used for warp-level culling)
//   * one 16-byte entry per visible Gaussian in a compact list {id, depth bits, packed tile rect}
//     that the key scatter reads coalesced
//   * instances are 8-byte keys (depth_bits<<32 | gaussian_id) binned per tile, so a tile's
//     whole work list is one contiguous segment that a CTA can pull into shared memory.
#pragma once
#include "fb_common.cuh"

namespace fb {

static inline __host__ __device__ size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// One projected Gaussian of one view: 32 bytes, 32-byte aligned.
#ifndef FB_TC_STRIDE
#define FB_TC_STRIDE 32   // uint32 slots per tile counter (128 B)
#endif

struct __align__(16) GRec {
    float4 a;  // px, py, depth, opacity
    float4 b;  // conic.x, conic.y, conic.z, footprint half extents as half2 (single view) / opacity (batched: a = px,py,hx,hy)
};

// View-batched binning state.  All arrays are [V][...] with the given strides.
// Tile lists of >= FB_HEAVY_KEYS keys (among the FB_HEAVY_RANKS heaviest tiles of a view) are swept in segments of
// FB_SEG_KEYS keys by different warps (fb_score.cuh: sweep_px2_segment).
#ifndef FB_HEAVY_KEYS
#define FB_HEAVY_KEYS 16384
#endif
#ifndef FB_HEAVY_RANKS
#define FB_HEAVY_RANKS 64
#endif
#ifndef FB_SEG_KEYS
#define FB_SEG_KEYS 8192
#endif
// work items the segment list needs for `instances` keys over at most `slots` heavy tiles (a hard bound)
static inline __host__ __device__ size_t seg_item_capacity(long long instances, long long slots) {
    return (size_t)(instances / FB_SEG_KEYS + slots + 1);
}

// Work order of the sort / sweep kernels: tile slots by descending instance count inside a view, the views of a GROUP of
// FB_ORDER_GROUP interleaved (rank-major inside the group), the groups one after the other.  A group of V views is the
// plain interleave tile_order[rank * V + v]; smaller groups keep the records of fewer views in flight (L2).
#ifndef FB_ORDER_GROUP
#define FB_ORDER_GROUP 64
#endif
static inline __host__ __device__ unsigned order_index(unsigned rank, unsigned v, unsigned V, unsigned tiles) {
    const unsigned g = v / FB_ORDER_GROUP, g0 = g * FB_ORDER_GROUP;
    const unsigned gs = (V - g0 < FB_ORDER_GROUP) ? V - g0 : FB_ORDER_GROUP;
    return g0 * tiles + rank * gs + (v - g0);
}
static inline __host__ __device__ unsigned order_rank(unsigned idx, unsigned V, unsigned tiles) {
    const unsigned g0 = (idx / (FB_ORDER_GROUP * tiles)) * FB_ORDER_GROUP;
    const unsigned gs = (V - g0 < FB_ORDER_GROUP) ? V - g0 : FB_ORDER_GROUP;
    return (idx - g0 * tiles) / gs;
}

struct Bins {
    GRec*     recs;        // [V][P]   drop-in and H_train paths: indexed by Gaussian id
    float4*   recsB;       // unused by the kernels (the H_train path keeps its moment rows in that region)
    // scoring path: records live in the view's VISIBLE SLOT (dense, coalesced writes, arrays sized by vis_stride, not P)
    float4*   recA;        // [V][vis_stride]     {px, py, hx, hy}: all the tile-list scan needs
    float4*   recB;        // [V][vis_stride][5]  {conic', opacity | r+g+b, coef0, -, - | coef1-4 | coef5-8 | coef9-12}
    uint4*    vis_list;    // [V][vis_stride]   visible Gaussians, unordered: {id, depth bits, x0|y0<<8|x1<<16|y1<<24, 0}
    size_t    vis_stride;  // entries per view in vis_list (P, or the scoring path's visible capacity)
    int       key_is_slot; // scoring path: the low word of a key is the visible slot, not the Gaussian id
    uint32_t* vis_count;   // [V]
    u64*      view_total;  // [V]  chained scan: (1 << 32 | instances of the view) once published (zeroed per chunk)
    uint32_t* tile_count;  // [V][tiles][FB_TC_STRIDE]  instances per tile (cursor of the scatter); one counter per
                           // 128-byte line: atomics on one L2 line serialise, 32 hot counters per line cost 10x
    uint2*    tile_range;  // [V][tiles]   [start,end) into keys
    uint32_t* tile_order;  // [V*tiles]    scoring only (else null): tile slots by descending instance count
    uint32_t* total;       // [4]: total instances (lo), overflow flag, prefiltered-trap flag, max tile count
    u64*      keys;        // [capacity]
    uint4*    seg_list;    // scoring path (else null): heavy tile lists {tile slot, segments, first item, -} (k_scan_tiles -> k_tile_score_mix);
                           // total[10] items, [11] ticket counter, [12] lists, [13] longest list in segments
    uint4*    seg_flags;   // [items][2]  chain flags and completion counters of the work items (fb_score.cuh: SegWs)
    uint4*    sub_list;    // scoring / H_train paths (else null): pieces of the long segments (split_segment in k_sort_tiles); count in total[8], reports in total[9]
    long long capacity;
    int       P;
    int       V;
    int       tiles;
};

struct GeomLayout {
    size_t recs, rgb, clamped, vis_list, vis_count, total;
};
static inline __host__ GeomLayout geom_layout(int P) {
    GeomLayout L; size_t o = 0; size_t Pp = (size_t)(P > 0 ? P : 1);
    L.recs = o;      o = align_up(o + Pp * sizeof(GRec), 256);
    L.rgb = o;       o = align_up(o + Pp * 3 * sizeof(float), 256);
    L.clamped = o;   o = align_up(o + Pp * 4, 256);
    L.vis_list = o;  o = align_up(o + Pp * 16, 256);
    L.vis_count = o; o = align_up(o + 256, 256);
    L.total = o;
    return L;
}
struct ImageLayout {
    size_t final_T, n_contrib, tile_range, tile_count, tile_order, totals, total;   // totals: uint32[16] counters, then u64 view_total[1]
};
static inline __host__ ImageLayout image_layout(int W, int H) {
    ImageLayout L; size_t o = 0; size_t HW = (size_t)W * H;
    size_t tiles = (size_t)((W + FB_TILE_X - 1) / FB_TILE_X) * ((H + FB_TILE_Y - 1) / FB_TILE_Y);
    L.final_T = o;    o = align_up(o + HW * 4, 256);
    L.n_contrib = o;  o = align_up(o + HW * 4, 256);
    L.tile_range = o; o = align_up(o + tiles * 8, 256);
    L.tile_count = o; o = align_up(o + tiles * 4 * FB_TC_STRIDE, 256);
    L.tile_order = o; o = align_up(o + tiles * 4, 256);
    L.totals = o;     o = align_up(o + 256, 256);
    L.total = o;
    return L;
}

}  // namespace fb
