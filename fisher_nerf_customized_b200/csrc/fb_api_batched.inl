// fb_api_batched.inl -- debug export, batched scoring / H_train entry points, stage 5 entry points.
// Included by fb_api.cu.

// FB_SCORE_PX2 (default): the pair sweep is k_tile_score2 (two pixels per lane, packed FP32); 0 selects k_tile_score1.
// FB_SORT_SPLIT: tile segments of >= FB_SPLIT_MIN keys (among the FB_SPLIT_RANKS heaviest tiles of a view) are partitioned by
// depth and sorted piecewise by FB_SORT_SUB_CTAS CTAs (fb_stage12.cuh: split_segment).
#ifndef FB_SORT_SPLIT
#define FB_SORT_SPLIT 1
#endif
#ifndef FB_SPLIT_RANKS
#define FB_SPLIT_RANKS 64
#endif
#ifndef FB_SORT_SUB_CTAS
#define FB_SORT_SUB_CTAS (148 * 6)
#endif
#ifndef FB_HESS_PX2_MIN_TILES
#define FB_HESS_PX2_MIN_TILES 8192     // (view, tile) pairs of a chunk from which k_tile_hessian2 is used
#endif
#ifndef FB_SCORE_PX2
#define FB_SCORE_PX2 1
#endif
// FB_SCORE_MIX (default, with FB_SCORE_PX2): k_tile_score_mix -- lists of >= FB_HEAVY_KEYS keys are swept in chained depth segments.
#ifndef FB_SCORE_MIX
#define FB_SCORE_MIX FB_SCORE_PX2
#endif
#if FB_SCORE_PX2
#define FB_SCORE_KERNEL k_tile_score2
#define FB_SCORE_BLOCKS FB_S2_BLOCKS
#define FB_SCORE_KWPC FB_S2_WPC
#define FB_SCORE_KSMEM FB_S2_SMEM_BYTES
#else
#define FB_SCORE_KERNEL k_tile_score1
#define FB_SCORE_BLOCKS FB_TILE_WARPS
#define FB_SCORE_KWPC FB_SCORE_WPC
#define FB_SCORE_KSMEM FB_SCORE_SMEM_BYTES
#endif

namespace fb {

__global__ void k_debug_geom(int P, Cam cam, const GRec* recs, const int* radii_or_null, const uint4* vis_list,
                             uint32_t nvis_cap, const uint32_t* vis_count, uint32_t* tiles_touched, float* depths,
                             float* means2D, float* conic_opacity) {
    // Gaussians that are not visible keep zeros (outputs are pre-zeroed by the host wrapper).
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    (void)radii_or_null; (void)nvis_cap;
    if (i >= vis_count[0]) return;
    unsigned gid = vis_list[i].x;
    if ((int)gid >= P) return;
    GRec r = recs[gid];
    const unsigned rect = vis_list[i].z;          // x0 | y0 << 8 | x1 << 16 | y1 << 24, as stage 1 computed it
    const int x0 = rect & 0xff, y0 = (rect >> 8) & 0xff, x1 = (rect >> 16) & 0xff, y1 = rect >> 24;
    (void)cam;
    if (tiles_touched) tiles_touched[gid] = (uint32_t)((x1 - x0) * (y1 - y0));
    if (depths) depths[gid] = r.a.z;
    if (means2D) { means2D[2 * gid] = r.a.x; means2D[2 * gid + 1] = r.a.y; }
    if (conic_opacity) {
        conic_opacity[4 * gid] = r.b.x; conic_opacity[4 * gid + 1] = r.b.y;
        conic_opacity[4 * gid + 2] = r.b.z; conic_opacity[4 * gid + 3] = r.a.w;
    }
}
__global__ void k_debug_inst(Bins b, uint32_t* point_list, u64* point_keys) {
    const int t = blockIdx.x;
    uint2 r = b.tile_range[t];
    for (unsigned i = r.x + threadIdx.x; i < r.y; i += blockDim.x) {
        u64 k = b.keys[i];
        if (point_list) point_list[i] = (uint32_t)(k & 0xffffffffull);
        if (point_keys) point_keys[i] = ((u64)t << 32) | (k >> 32);
    }
}

__global__ void k_status_update(const uint32_t* total, const uint32_t* vis_count, int V, int* status) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        // atomics: with two chunk lanes (fb_score_candidates) two of these may run at the same time
        int nvis = 0;
        for (int v = 0; v < V; ++v) nvis += (int)vis_count[v];
        atomicAdd(&status[4], nvis);
        atomicMax(&status[5], (int)total[4]);                     // largest visible count of a view
        if (total[1]) atomicExch(&status[0], FB_E_CAPACITY);
        atomicMax(&status[1], (int)total[0]);
        atomicMax(&status[2], (int)total[3]);
        atomicAdd(&status[3], (int)total[0]);
    }
}
__global__ void k_fill(float* p, size_t n, float v) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// Control words of one chunk, cleared with ONE memset per chunk: vis_count[64] | totals (uint32[64]) | view_total[64].
#define FB_CTRL_BYTES (256 + 256 + 512)
struct CtrlPtrs { uint32_t* vis_count; uint32_t* total; u64* view_total; };
static CtrlPtrs ctrl_ptrs(char* base) { return CtrlPtrs{(uint32_t*)base, (uint32_t*)(base + 256), (u64*)(base + 512)}; }

// Scoring workspace of one chunk lane.  Records live in the view's visible slot: [V][vis_cap] rows of 16 + 80 + 16
// bytes (vis_cap = the caller's visible capacity, <= P), not [V][P] rows.
struct ScoreLayout {
    size_t recA, recB, vis_list, ctrl, tile_count, tile_range, tile_order, partial, keys, keys_alt, sub_list, seg_list, seg_zero, seg_state, seg_out, seg_items, total;
};
static ScoreLayout score_layout(int P, int W, int H, int V, long long cap, long long vis_cap) {
    ScoreLayout L; size_t o = 0; size_t Vv = (size_t)(V > 0 ? V : 1);
    size_t vc = (size_t)(vis_cap > 0 && vis_cap < P ? vis_cap : (P > 0 ? P : 1));
    size_t tiles = (size_t)((W + FB_TILE_X - 1) / FB_TILE_X) * ((H + FB_TILE_Y - 1) / FB_TILE_Y);
    L.recA = o;       o = align_up(o + Vv * vc * 16, 256);
    L.recB = o;       o = align_up(o + Vv * vc * 80, 256);
    L.vis_list = o;   o = align_up(o + Vv * vc * 16, 256);
    L.ctrl = o;       o = align_up(o + FB_CTRL_BYTES, 256);
    L.tile_count = o; o = align_up(o + Vv * tiles * 4 * FB_TC_STRIDE, 256);
    L.tile_range = o; o = align_up(o + Vv * tiles * 8, 256);
    L.tile_order = o; o = align_up(o + Vv * tiles * 4, 256);
    L.partial = o;    o = align_up(o + Vv * tiles * 4 * FB_TILE_WARPS, 256);
    L.keys = o;       o = align_up(o + (size_t)(cap > 0 ? cap : 1) * 8, 256);
    L.keys_alt = o;   o = align_up(o + (size_t)(cap > 0 ? cap : 1) * 8, 256);
    L.sub_list = o;   o = align_up(o + sort_sub_capacity(cap > 0 ? cap : 1, (long long)Vv * FB_SPLIT_RANKS) * 16, 256);
    L.seg_items = seg_item_capacity(cap > 0 ? cap : 1, (long long)Vv * FB_HEAVY_RANKS);
    L.seg_list = o;   o = align_up(o + L.seg_items * 16, 256);
    L.seg_zero = o;   o = align_up(o + L.seg_items * 32, 256);          // chain flags and completion counters: [items][8]
    L.seg_state = o;  o = align_up(o + L.seg_items * 4 * 32 * 16, 256);
    L.seg_out = o;    o = align_up(o + L.seg_items * 4 * 64 * 16, 256);
    L.total = o;
    return L;
}
static Bins score_bins(const ScoreLayout& L, char* ws, int P, int V, const Cam& cam, long long cap, long long vis_cap) {
    Bins b;
    b.recs = nullptr; b.recsB = nullptr;
    b.recA = (float4*)(ws + L.recA); b.recB = (float4*)(ws + L.recB);
    b.vis_list = (uint4*)(ws + L.vis_list);
    b.vis_stride = (size_t)(vis_cap > 0 && vis_cap < P ? vis_cap : P); b.key_is_slot = 1;
    CtrlPtrs c = ctrl_ptrs(ws + L.ctrl);
    b.vis_count = c.vis_count; b.total = c.total; b.view_total = c.view_total;
    b.tile_count = (uint32_t*)(ws + L.tile_count); b.tile_range = (uint2*)(ws + L.tile_range);
    b.tile_order = (uint32_t*)(ws + L.tile_order);
    b.keys = (u64*)(ws + L.keys); b.sub_list = (uint4*)(ws + L.sub_list); b.seg_list = FB_SCORE_MIX ? (uint4*)(ws + L.seg_list) : nullptr; b.seg_flags = (uint4*)(ws + L.seg_zero);
    b.capacity = cap; b.P = P; b.V = V; b.tiles = cam.tiles_x * cam.tiles_y;
    return b;
}

// H_train workspace (fb_accumulate_hessian): records and moment rows are indexed by Gaussian id, [V][P].
struct HessLayout {
    size_t recs, mom, vis_list, ctrl, tile_count, tile_range, tile_order, keys, keys_alt, sub_list, total;
};
static HessLayout hess_layout(int P, int W, int H, int V, long long cap) {
    HessLayout L; size_t o = 0; size_t Pp = (size_t)(P > 0 ? P : 1), Vv = (size_t)(V > 0 ? V : 1);
    size_t tiles = (size_t)((W + FB_TILE_X - 1) / FB_TILE_X) * ((H + FB_TILE_Y - 1) / FB_TILE_Y);
    L.recs = o;       o = align_up(o + Vv * Pp * sizeof(GRec), 256);
    L.mom = o;        o = align_up(o + Vv * Pp * FB_NACC * 4, 256);
    L.vis_list = o;   o = align_up(o + Vv * Pp * 16, 256);
    L.ctrl = o;       o = align_up(o + FB_CTRL_BYTES, 256);
    L.tile_count = o; o = align_up(o + Vv * tiles * 4 * FB_TC_STRIDE, 256);
    L.tile_range = o; o = align_up(o + Vv * tiles * 8, 256);
    L.tile_order = o; o = align_up(o + Vv * tiles * 4, 256);
    L.keys = o;       o = align_up(o + (size_t)(cap > 0 ? cap : 1) * 8, 256);
    L.keys_alt = o;   o = align_up(o + (size_t)(cap > 0 ? cap : 1) * 8, 256);
    L.sub_list = o;   o = align_up(o + sort_sub_capacity(cap > 0 ? cap : 1, (long long)Vv * FB_SPLIT_RANKS) * 16, 256);
    L.total = o;
    return L;
}
static Bins hess_bins(const HessLayout& L, char* ws, int P, int V, const Cam& cam, long long cap) {
    Bins b;
    b.recs = (GRec*)(ws + L.recs); b.recsB = nullptr; b.recA = nullptr; b.recB = nullptr;
    b.vis_list = (uint4*)(ws + L.vis_list); b.vis_stride = (size_t)P; b.key_is_slot = 0;
    CtrlPtrs c = ctrl_ptrs(ws + L.ctrl);
    b.vis_count = c.vis_count; b.total = c.total; b.view_total = c.view_total;
    b.tile_count = (uint32_t*)(ws + L.tile_count); b.tile_range = (uint2*)(ws + L.tile_range);
    b.tile_order = (uint32_t*)(ws + L.tile_order);
    b.keys = (u64*)(ws + L.keys); b.sub_list = (uint4*)(ws + L.sub_list); b.seg_list = nullptr; b.seg_flags = nullptr;
    b.capacity = cap; b.P = P; b.V = V; b.tiles = cam.tiles_x * cam.tiles_y;
    return b;
}
static GaussIn map_in(const fb_map* m) {
    GaussIn i;
    i.means3D = m->means3D; i.opacities = m->opacities; i.shs = nullptr; i.colors = m->colors;
    i.scales = m->scales; i.rotations = m->rotations; i.cov3D = nullptr;
    return i;
}


}  // namespace fb

extern "C" {

int fb_debug_export(const fb_camera* c, int P, int num_rendered, const void* geom, const void* image, const void* inst,
                    uint32_t* tiles_touched, float* depths, float* means2D, float* conic_opacity, uint32_t* ranges,
                    float* final_T, uint32_t* n_contrib, uint32_t* point_list, uint64_t* point_keys, void* stream_) {
    cudaStream_t st = (cudaStream_t)stream_;
    if (!c || P <= 0 || !geom || !image) return FB_E_INVALID;
    Cam cam = make_cam(c);
    Bins b = single_bins(P, cam, (void*)geom, (void*)image, (void*)inst, (long long)num_rendered);
    ImageLayout IL = image_layout(cam.W, cam.H);
    size_t HW = (size_t)cam.W * cam.H;
    if (tiles_touched) FB_CUDA(cudaMemsetAsync(tiles_touched, 0, (size_t)P * 4, st));
    if (depths) FB_CUDA(cudaMemsetAsync(depths, 0, (size_t)P * 4, st));
    if (means2D) FB_CUDA(cudaMemsetAsync(means2D, 0, (size_t)P * 8, st));
    if (conic_opacity) FB_CUDA(cudaMemsetAsync(conic_opacity, 0, (size_t)P * 16, st));
    k_debug_geom<<<(P + 255) / 256, 256, 0, st>>>(P, cam, b.recs, nullptr, b.vis_list, (uint32_t)P, b.vis_count,
                                                  tiles_touched, depths, means2D, conic_opacity);
    FB_LAUNCH_CHECK();
    if (ranges) FB_CUDA(cudaMemcpyAsync(ranges, b.tile_range, (size_t)b.tiles * 8, cudaMemcpyDeviceToDevice, st));
    if (final_T) FB_CUDA(cudaMemcpyAsync(final_T, (const char*)image + IL.final_T, HW * 4, cudaMemcpyDeviceToDevice, st));
    if (n_contrib) FB_CUDA(cudaMemcpyAsync(n_contrib, (const char*)image + IL.n_contrib, HW * 4, cudaMemcpyDeviceToDevice, st));
    if ((point_list || point_keys) && inst && num_rendered > 0) {
        k_debug_inst<<<b.tiles, 256, 0, st>>>(b, point_list, (u64*)point_keys);
        FB_LAUNCH_CHECK();
    }
    return FB_OK;
}

size_t fb_score_workspace_bytes(int P, int width, int height, int views_per_chunk, long long inst_capacity,
                                long long vis_capacity) {
    return score_layout(P, width, height, views_per_chunk, inst_capacity, vis_capacity).total;
}
size_t fb_hessian_workspace_bytes(int P, int width, int height, int views_per_chunk, long long inst_capacity) {
    return hess_layout(P, width, height, views_per_chunk, inst_capacity).total;
}

// Per-device library state (function attributes are per device; the second chunk lane's stream and events belong to
// the device they were created on).  Indexed by the CURRENT device of the calling thread, which the host side makes
// the device of the tensors it passes (torch.cuda.device guard in _lib.py).
struct DevState {
    bool score_attr, hess_attr, sort_attr;
    cudaStream_t lane_stream; cudaEvent_t lane_fork, lane_join;
};
static DevState g_dev[64];
static DevState* dev_state() {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= 64) return nullptr;
    return &g_dev[d];
}

// Two chunk lanes.  The launch sequence of a chunk is a chain of kernels with very different limits (preprocess and the
// pair sweep issue-bound; scan, scatter and sort latency-bound).
// When the caller's workspace holds two chunks (>= 2 x fb_score_workspace_bytes), consecutive chunks alternate between
// the caller's stream and a library-owned stream, each in its own half of the workspace, so the latency-bound kernels
// of one chunk run under the issue-bound kernels of the other.  Results are unchanged (each view's arithmetic is its
// own); the caller's stream waits for the second lane before fb_score_candidates' work is complete on it.
// While per-kernel profiling is enabled the chunks stay on one stream so that the event pairs time kernels alone.
static int g_lanes_enabled = 1;
int fb_score_set_lanes(int on) { g_lanes_enabled = on != 0; return FB_OK; }

int fb_score_candidates(const fb_camera* c, const fb_map* map, const float* w2c, int K, const float* weight, int C_H,
                        float grad_scale, float* scores, int* status_dev, void* workspace, size_t workspace_bytes,
                        int views_per_chunk, long long inst_capacity, long long vis_capacity, void* stream_) {
    return fb_score_candidates_sets(c, map, w2c, K, weight, nullptr, C_H, grad_scale, scores, status_dev, workspace,
                                    workspace_bytes, views_per_chunk, inst_capacity, vis_capacity, stream_);
}

int fb_score_candidates_sets(const fb_camera* c, const fb_map* map, const float* w2c, int K, const float* weight,
                             const int* weight_set, int C_H, float grad_scale, float* scores, int* status_dev,
                             void* workspace, size_t workspace_bytes, int views_per_chunk, long long inst_capacity,
                             long long vis_capacity, void* stream_) {
    cudaStream_t st = (cudaStream_t)stream_;
    if (!c || !map || K < 0 || (C_H != 4 && C_H != 11) || views_per_chunk <= 0 || views_per_chunk > 64 ||
        c->image_width <= 0 || c->image_height <= 0 || c->image_width > 4080 || c->image_height > 4080 ||
        inst_capacity <= 0 || vis_capacity < 0 || !status_dev) return FB_E_INVALID;
    if (K == 0) return FB_OK;
    const int P = map->P;
    if (P <= 0 || !map->means3D || !map->opacities || !map->scales || !map->rotations || !map->colors || !w2c ||
        !weight || !scores || !workspace) return FB_E_INVALID;
    const long long vcap = (vis_capacity > 0 && vis_capacity < P) ? vis_capacity : P;
    if (vcap * 80ll >= (1ll << 32)) return FB_E_INVALID;       // 32-bit record offsets inside a view (k_tile_score*)
    Cam cam = make_cam(c);
    const int V = views_per_chunk;
    ScoreLayout L = score_layout(P, cam.W, cam.H, V, inst_capacity, vcap);
    if (workspace_bytes < L.total) return FB_E_WORKSPACE;
    DevState* ds = dev_state();
    if (!ds) return FB_E_CUDA;
    if (!ds->score_attr) {
        FB_CUDA(cudaFuncSetAttribute(k_sort_tiles, cudaFuncAttributeMaxDynamicSharedMemorySize, FB_SORT_SMEM_BYTES));
        FB_CUDA(cudaFuncSetAttribute(FB_SCORE_KERNEL, cudaFuncAttributeMaxDynamicSharedMemorySize, FB_SCORE_KSMEM));
        FB_CUDA(cudaFuncSetAttribute(k_tile_score_mix, cudaFuncAttributeMaxDynamicSharedMemorySize, FB_MIX_SMEM_BYTES));
        FB_CUDA(cudaFuncSetAttribute(k_preprocess_score<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PSmem<4>)));
        FB_CUDA(cudaFuncSetAttribute(k_preprocess_score<11>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PSmem<11>)));
        ds->score_attr = true; ds->sort_attr = true;
    }
    cudaStream_t st0 = st;
    // (splitting a single small chunk over the two lanes was measured and dropped: 4 keyframes + 32 candidates per
    //  planning step take 2.75 ms as two half chunks against 2.63 ms as one: launches that small are latency-bound)
    const int lanes = (g_lanes_enabled && !g_prof_on && K > V && workspace_bytes >= 2 * L.total) ? 2 : 1;
    const int Vstep = V;
    char* ws = (char*)workspace;
    FB_CUDA(cudaMemsetAsync(status_dev, 0, 8 * sizeof(int), st0));
    const size_t tc_bytes = (size_t)V * cam.tiles_x * cam.tiles_y * 4 * FB_TC_STRIDE;
    FB_CUDA(cudaMemsetAsync(ws + L.tile_count, 0, tc_bytes, st0));
    if (lanes == 2) {
        if (!ds->lane_stream) {
            FB_CUDA(cudaStreamCreateWithFlags(&ds->lane_stream, cudaStreamNonBlocking));
            FB_CUDA(cudaEventCreateWithFlags(&ds->lane_fork, cudaEventDisableTiming));
            FB_CUDA(cudaEventCreateWithFlags(&ds->lane_join, cudaEventDisableTiming));
        }
        FB_CUDA(cudaEventRecord(ds->lane_fork, st0));                 // after the status reset and everything queued before
        FB_CUDA(cudaStreamWaitEvent(ds->lane_stream, ds->lane_fork, 0));
        FB_CUDA(cudaMemsetAsync(ws + L.total + L.tile_count, 0, tc_bytes, ds->lane_stream));
    }
    GaussIn gi = map_in(map);
    char* const ws0 = ws;
    int chunk = 0;
    for (int k0 = 0; k0 < K; k0 += Vstep, ++chunk) {
        const int Vc = (K - k0 < Vstep) ? (K - k0) : Vstep;
        const int lane = (lanes == 2) ? (chunk & 1) : 0;
        st = lane ? ds->lane_stream : st0;
        ws = ws0 + (size_t)lane * L.total;
        Bins b = score_bins(L, ws, P, Vc, cam, inst_capacity, vcap);
        FB_CUDA(cudaMemsetAsync(ws + L.ctrl, 0, FB_CTRL_BYTES, st));
        const float* w2c_k = w2c + (size_t)k0 * 16;
        const dim3 ps_grid((P + FB_PS_THREADS - 1) / FB_PS_THREADS, min(FB_PS_SPLIT, (Vc + FB_PS_VG - 1) / FB_PS_VG));
        if (C_H == 4)
            { FB_PROF(FB_K_PREPROCESS); k_preprocess_score<4><<<ps_grid, FB_PS_THREADS, sizeof(PSmem<4>), st>>>(gi, cam, w2c_k, b, weight, weight_set ? weight_set + k0 : nullptr); }
        else
            { FB_PROF(FB_K_PREPROCESS); k_preprocess_score<11><<<ps_grid, FB_PS_THREADS, sizeof(PSmem<11>), st>>>(gi, cam, w2c_k, b, weight, weight_set ? weight_set + k0 : nullptr); }
        FB_LAUNCH_CHECK();
        { FB_PROF(FB_K_SCAN); k_scan_tiles<<<Vc, 1024, 0, st>>>(b); }
        FB_LAUNCH_CHECK();
        ++g_launches; k_status_update<<<1, 32, 0, st>>>(b.total, b.vis_count, b.V, status_dev);
#if FB_SCATTER_CTA
        if (b.tiles <= FB_SC_MAX_TILES) {
            int sblocks = (int)((b.vis_stride + 256 * FB_SC_EPT - 1) / (256 * FB_SC_EPT)); if (sblocks > 148 * FB_SC_GRID) sblocks = 148 * FB_SC_GRID;
            { FB_PROF(FB_K_SCATTER); k_scatter_keys_cta<<<dim3(sblocks, Vc), 256, 0, st>>>(b, cam); }
        } else
#endif
        {
            int sblocks = (int)((b.vis_stride + 255) / 256); if (sblocks > 148 * 2) sblocks = 148 * 2;
            { FB_PROF(FB_K_SCATTER); k_scatter_keys<<<dim3(sblocks, Vc), 256, 0, st>>>(b, cam); }
        }
        FB_LAUNCH_CHECK();
        {
            // long segments are cut into pieces by their CTA, the pieces sorted by the last FB_SORT_SUB_CTAS CTAs of the grid
            const int split_slots = FB_SORT_SPLIT ? (b.tiles < FB_SPLIT_RANKS ? b.tiles : FB_SPLIT_RANKS) * Vc : 0;
            const int sub_ctas = split_slots ? FB_SORT_SUB_CTAS : 0;
            { FB_PROF(FB_K_SORT); k_sort_tiles<<<b.tiles * Vc + sub_ctas, 256, FB_SORT_SMEM_BYTES, st>>>(b, (u64*)(ws + L.keys_alt), FB_SORT_SMEM_KEYS, b.tiles * Vc, split_slots); }
        }
        FB_LAUNCH_CHECK();
#if FB_SCORE_MIX
        {
            SegWs sw;
            sw.state = (float4*)(ws + L.seg_state); sw.out = (float4*)(ws + L.seg_out);
            sw.flag = (unsigned*)(ws + L.seg_zero);
            FB_PROF(FB_K_TILE_SCORE);
            k_tile_score_mix<<<FB_SEG_WORKERS + b.tiles * Vc, 128, FB_MIX_SMEM_BYTES, st>>>(b, cam, grad_scale, (float*)(ws + L.partial), sw, FB_SEG_WORKERS);
        }
        FB_LAUNCH_CHECK();
        { FB_PROF(FB_K_FINALIZE); k_score_finalize<<<Vc, 32, 0, st>>>((const float*)(ws + L.partial), b.tiles * FB_TILE_WARPS, scores + k0); }
#else
        { FB_PROF(FB_K_TILE_SCORE); FB_SCORE_KERNEL<<<dim3(b.tiles * (FB_SCORE_BLOCKS / FB_SCORE_KWPC), Vc), 32 * FB_SCORE_KWPC, FB_SCORE_KSMEM, st>>>(b, cam, grad_scale, (float*)(ws + L.partial)); }
        FB_LAUNCH_CHECK();
        { FB_PROF(FB_K_FINALIZE); k_score_finalize<<<Vc, 32, 0, st>>>((const float*)(ws + L.partial), b.tiles * FB_SCORE_BLOCKS, scores + k0); }
#endif
        FB_LAUNCH_CHECK();
    }
    if (lanes == 2) {
        FB_CUDA(cudaEventRecord(ds->lane_join, ds->lane_stream));
        FB_CUDA(cudaStreamWaitEvent(st0, ds->lane_join, 0));
    }
    return FB_OK;
}

int fb_accumulate_hessian(const fb_camera* c, const fb_map* map, const float* w2c, int K, int C_H, float grad_scale,
                          float* H_out, int* status_dev, void* workspace, size_t workspace_bytes, int views_per_chunk,
                          long long inst_capacity, void* stream_) {
    return fb_accumulate_hessian_sets(c, map, w2c, K, C_H, grad_scale, H_out, nullptr, status_dev, workspace,
                                      workspace_bytes, views_per_chunk, inst_capacity, stream_);
}

int fb_accumulate_hessian_sets(const fb_camera* c, const fb_map* map, const float* w2c, int K, int C_H, float grad_scale,
                               float* H_out, const int* out_set, int* status_dev, void* workspace, size_t workspace_bytes,
                               int views_per_chunk, long long inst_capacity, void* stream_) {
    cudaStream_t st = (cudaStream_t)stream_;
    if (!c || !map || K < 0 || (C_H != 4 && C_H != 11) || views_per_chunk <= 0 || views_per_chunk > 64 ||
        c->image_width <= 0 || c->image_height <= 0 || c->image_width > 4080 || c->image_height > 4080 ||
        inst_capacity <= 0 || !status_dev)
        return FB_E_INVALID;                                   // (tile coordinates are packed in 8 bits: 4080 px)
    if (K == 0) return FB_OK;
    const int P = map->P;
    if (P <= 0 || !map->means3D || !map->opacities || !map->scales || !map->rotations || !map->colors || !w2c ||
        !H_out || !workspace) return FB_E_INVALID;
    Cam cam = make_cam(c);
    HessLayout L = hess_layout(P, cam.W, cam.H, views_per_chunk, inst_capacity);
    if (workspace_bytes < L.total) return FB_E_WORKSPACE;
    char* ws = (char*)workspace;
    DevState* ds = dev_state();
    if (!ds) return FB_E_CUDA;
    if (!ds->hess_attr) {
        FB_CUDA(cudaFuncSetAttribute(k_sort_tiles, cudaFuncAttributeMaxDynamicSharedMemorySize, FB_SORT_SMEM_BYTES));
        FB_CUDA(cudaFuncSetAttribute(k_tile_hessian, cudaFuncAttributeMaxDynamicSharedMemorySize, FB_SCORE_SMEM_BYTES));
        FB_CUDA(cudaFuncSetAttribute(k_tile_hessian2, cudaFuncAttributeMaxDynamicSharedMemorySize, FB_MIX_SMEM_BYTES));
        ds->hess_attr = true; ds->sort_attr = true;
    }
    const int V = views_per_chunk;
    float* mom = (float*)(ws + L.mom);     // [V][P][16] floats
    FB_CUDA(cudaMemsetAsync(status_dev, 0, 8 * sizeof(int), st));
    FB_CUDA(cudaMemsetAsync(mom, 0, (size_t)(K < V ? K : V) * P * FB_NACC * 4, st));   // k_expand_H clears the rows it consumes
    FB_CUDA(cudaMemsetAsync(ws + L.tile_count, 0, (size_t)V * cam.tiles_x * cam.tiles_y * 4 * FB_TC_STRIDE, st));
    GaussIn gi = map_in(map);
    // V keyframes per launch sequence: a single 256x256 view is 256 CTAs, less than two per SM
    for (int k0 = 0; k0 < K; k0 += V) {
        const int Vc = (K - k0 < V) ? (K - k0) : V;
        Bins b = hess_bins(L, ws, P, Vc, cam, inst_capacity);
        const float* w2c_k = w2c + (size_t)k0 * 16;
        FB_CUDA(cudaMemsetAsync(ws + L.ctrl, 0, FB_CTRL_BYTES, st));
        { FB_PROF(FB_K_PREPROCESS); k_preprocess<true, 0><<<dim3((P + 255) / 256, min(FB_PS_SPLIT, Vc)), 256, Vc * 48, st>>>(gi, cam, w2c_k, b, nullptr, nullptr, nullptr, nullptr); }
        FB_LAUNCH_CHECK();
        { FB_PROF(FB_K_SCAN); k_scan_tiles<<<Vc, 1024, 0, st>>>(b); }
        ++g_launches; k_status_update<<<1, 32, 0, st>>>(b.total, b.vis_count, b.V, status_dev);
        int sblocks = (P + 255) / 256; if (sblocks > 148 * 2) sblocks = 148 * 2;
#if FB_SCATTER_CTA
        if (b.tiles <= FB_SC_MAX_TILES) { FB_PROF(FB_K_SCATTER); k_scatter_keys_cta<<<dim3(sblocks, Vc), 256, 0, st>>>(b, cam); }
        else
#endif
        { FB_PROF(FB_K_SCATTER); k_scatter_keys<<<dim3(sblocks, Vc), 256, 0, st>>>(b, cam); }
        {
            // long segments are cut into pieces by their CTA, the pieces sorted by the last FB_SORT_SUB_CTAS CTAs of the grid
            const int split_slots = FB_SORT_SPLIT ? (b.tiles < FB_SPLIT_RANKS ? b.tiles : FB_SPLIT_RANKS) * Vc : 0;
            const int sub_ctas = split_slots ? FB_SORT_SUB_CTAS : 0;
            { FB_PROF(FB_K_SORT); k_sort_tiles<<<b.tiles * Vc + sub_ctas, 256, FB_SORT_SMEM_BYTES, st>>>(b, (u64*)(ws + L.keys_alt), FB_SORT_SMEM_KEYS, b.tiles * Vc, split_slots); }
        }
        FB_LAUNCH_CHECK();
#if FB_HESS_PX2
        // the 8x8 two-pixel form issues fewer instructions; with few (view, tile) pairs the 8x4 form's twice as many warps win
        // (at 256 tiles, 8x4 against 8x8 form: 4 keyframes 369 / 584 us, 16: 981 / 1041, 32: 2041 / 1857, 64: 4068 / 3559)
        if ((long long)b.tiles * Vc >= FB_HESS_PX2_MIN_TILES)
        { FB_PROF(FB_K_BWD_TILES); k_tile_hessian2<<<b.tiles * Vc, 128, FB_MIX_SMEM_BYTES, st>>>(b, cam, map->colors, grad_scale, mom); }
        else
#endif
        { FB_PROF(FB_K_BWD_TILES); k_tile_hessian<<<dim3(b.tiles * (FB_TILE_WARPS / FB_SCORE_WPC), Vc), 32 * FB_SCORE_WPC, FB_SCORE_SMEM_BYTES, st>>>(
              b, cam, map->colors, grad_scale, mom); }
        FB_LAUNCH_CHECK();
        const int* set_k = out_set ? out_set + k0 : nullptr;
        if (C_H == 4) { FB_PROF(FB_K_EXPAND); k_expand_H<4><<<(P + 127) / 128, 128, 0, st>>>(b, gi, cam, w2c_k, mom, H_out, set_k); }
        else { FB_PROF(FB_K_EXPAND); k_expand_H<11><<<(P + 127) / 128, 128, 0, st>>>(b, gi, cam, w2c_k, mom, H_out, set_k); }
        FB_LAUNCH_CHECK();
    }
    return FB_OK;
}

// ---- stage 5 ---------------------------------------------------------------------------------
static fb::KnnWs knn_ws(int P, void* workspace, size_t* total) {
    using namespace fb;
    KnnWs w; char* ws = (char*)workspace; size_t o = 0; size_t Pp = (size_t)(P > 0 ? P : 1);
    unsigned cap = (unsigned)(4 * Pp + 64);
    unsigned nblocks = (cap + 1 + 1024 * FB_SCAN_ITEMS - 1) / (1024 * FB_SCAN_ITEMS);
    w.cap = cap;
    w.bbox = (unsigned*)(ws + o);     o = align_up(o + 32, 256);
    w.grid = (KnnGrid*)(ws + o);      o = align_up(o + sizeof(KnnGrid), 256);
    w.cell = (unsigned*)(ws + o);     o = align_up(o + ((size_t)cap + 1) * 4, 256);
    w.cursor = (unsigned*)(ws + o);   o = align_up(o + (size_t)cap * 4, 256);
    w.blocksum = (unsigned*)(ws + o); o = align_up(o + (size_t)nblocks * 4, 256);
    w.sorted = (float4*)(ws + o);     o = align_up(o + Pp * 16, 256);
    *total = o;
    return w;
}
size_t fb_knn_workspace_bytes(int P) { size_t t; knn_ws(P, nullptr, &t); return t; }

int fb_dist2_knn3(int P, const float* points, float* mean_dist2, void* workspace, size_t workspace_bytes, void* stream_) {
    using namespace fb;
    cudaStream_t st = (cudaStream_t)stream_;
    if (P < 0) return FB_E_INVALID;
    if (P == 0) return FB_OK;
    if (!points || !mean_dist2 || !workspace) return FB_E_INVALID;
    size_t need; KnnWs w = knn_ws(P, workspace, &need);
    if (workspace_bytes < need) return FB_E_WORKSPACE;
    const unsigned n = w.cap + 1, nb = (n + 1024 * FB_SCAN_ITEMS - 1) / (1024 * FB_SCAN_ITEMS);
    g_launches += 8;                                   // the eight helper kernels below; the query counts itself
    k_knn_init<<<148 * 4, 256, 0, st>>>(w);
    k_knn_bbox<<<148 * 4, 256, 0, st>>>(P, points, w);
    k_knn_setup<<<1, 32, 0, st>>>(P, w);
    k_knn_count<<<(P + 255) / 256, 256, 0, st>>>(P, points, w);
    FB_LAUNCH_CHECK();
    k_scan_block<<<nb, 1024, 0, st>>>(w.cell, n, w.blocksum);
    k_scan_sums<<<1, 1024, 0, st>>>(w.blocksum, nb);
    k_scan_add<<<nb, 1024, 0, st>>>(w.cell, n, w.blocksum);
    k_knn_scatter<<<(P + 255) / 256, 256, 0, st>>>(P, points, w);
    { FB_PROF(FB_K_KNN); k_knn_query<<<(P + 127) / 128, 128, 0, st>>>(P, w, mean_dist2); }
    FB_LAUNCH_CHECK();
    return FB_OK;
}

int fb_count_visible(int N, const float* points, int K, const double* w2c, const double* intr, int img_w, int img_h,
                     int* counts, uint8_t* mask, void* stream_) {
    using namespace fb;
    cudaStream_t st = (cudaStream_t)stream_;
    if (N < 0 || K < 0 || K > 256) return FB_E_INVALID;
    if (K == 0) return FB_OK;
    if (!counts || !w2c || !intr) return FB_E_INVALID;
    FB_CUDA(cudaMemsetAsync(counts, 0, (size_t)K * 4, st));
    if (N == 0) return FB_OK;
    if (!points) return FB_E_INVALID;
    int blocks = (N + 255) / 256; if (blocks > 148 * 8) blocks = 148 * 8;
    { FB_PROF(FB_K_VISIBLE); k_count_visible<<<blocks, 256, (size_t)K * (12 * 8 + 4), st>>>(N, points, K, w2c, intr, img_w, img_h, counts, mask); }
    FB_LAUNCH_CHECK();
    return FB_OK;
}

int fb_block_outer_accumulate(int Nv, const long long* rows, const float* g_mean, const float* g_op, const float* g_rot,
                              const float* g_scale, float divide_by, float* H_blocks, void* stream_) {
    using namespace fb;
    cudaStream_t st = (cudaStream_t)stream_;
    if (Nv < 0) return FB_E_INVALID;
    if (Nv == 0) return FB_OK;
    if (!rows || !g_mean || !H_blocks) return FB_E_INVALID;
    const int d = 3 + (g_op ? 1 : 0) + (g_rot ? 4 : 0) + (g_scale ? 3 : 0);
    const long long n = (long long)Nv * d * d;
    long long blocks = (n + 255) / 256; if (blocks > 148 * 16) blocks = 148 * 16;
    BlockCols g{g_mean, g_op, g_rot, g_scale};
    { FB_PROF(FB_K_BLOCKS); k_block_outer<<<(int)blocks, 256, 0, st>>>(n, d, rows, g, divide_by, H_blocks); }
    FB_LAUNCH_CHECK();
    return FB_OK;
}

int fb_block_opt_partials(int n_blocks) { return (n_blocks + FB_BLOCK_OPT_THREADS - 1) / FB_BLOCK_OPT_THREADS; }

int fb_block_opt(int n_blocks, int d, const float* Hm_blocks, const long long* idx_m, const float* J_blocks,
                 const long long* idx_j, float lam, int criterion, double* partial, float* per_block, void* stream_) {
    using namespace fb;
    cudaStream_t st = (cudaStream_t)stream_;
    if (n_blocks < 0 || d < 1 || d > 11 || (criterion != 0 && criterion != 1)) return FB_E_INVALID;
    if (n_blocks == 0) return FB_OK;
    if (!Hm_blocks || !J_blocks || !partial) return FB_E_INVALID;
    const int grid = fb_block_opt_partials(n_blocks);
    const size_t smem = (size_t)(d * d + d) * FB_BLOCK_OPT_THREADS * sizeof(double);
    FB_PROF(FB_K_BLOCKS);
    if (criterion == 0) k_block_opt<false><<<grid, FB_BLOCK_OPT_THREADS, smem, st>>>(n_blocks, d, Hm_blocks, idx_m, J_blocks, idx_j, lam, partial, per_block);
    else                k_block_opt<true><<<grid, FB_BLOCK_OPT_THREADS, smem, st>>>(n_blocks, d, Hm_blocks, idx_m, J_blocks, idx_j, lam, partial, per_block);
    FB_LAUNCH_CHECK();
    return FB_OK;
}

int fb_unproject_append(int width, int height, int downsample, const float* color, const float* depth,
                        const float* intr, const float* c2w, const long long* pix, long long n,
                        float* means3D, int mean_stride, float* rgb, int rgb_stride, float* unnorm_rotations,
                        float* logit_opacities, float* log_scales, int log_scale_cols, float* mean3_sq_dist,
                        void* stream_) {
    using namespace fb;
    cudaStream_t st = (cudaStream_t)stream_;
    if (width <= 0 || height <= 0 || downsample <= 0 || n < 0 || !intr || !c2w) return FB_E_INVALID;
    if (n == 0) return FB_OK;
    if (!color || !depth || !means3D || !rgb || mean_stride < 3 || rgb_stride < 3 ||
        (log_scales && log_scale_cols != 1 && log_scale_cols != 3)) return FB_E_INVALID;
    UnprojectArgs a;
    a.W = width; a.H = height; a.ds = downsample; a.Wd = (width + downsample - 1) / downsample;
    if (!pix && n > (long long)a.Wd * ((height + downsample - 1) / downsample)) return FB_E_INVALID;
    a.intr = intr; a.c2w = c2w;
    a.color = color; a.depth = depth; a.pix = pix; a.n = n;
    a.means3D = means3D; a.mean_stride = mean_stride; a.rgb = rgb; a.rgb_stride = rgb_stride;
    a.rots = unnorm_rotations; a.logit_op = logit_opacities; a.log_scales = log_scales; a.ls_cols = log_scale_cols;
    a.msd = mean3_sq_dist;
    { FB_PROF(FB_K_MAPUPDATE); k_unproject_append<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a); }
    FB_LAUNCH_CHECK();
    return FB_OK;
}

int fb_morton_codes(int P, const float* means3D, const float* lo, const float* ext, int* codes, void* stream_) {
    using namespace fb;
    cudaStream_t st = (cudaStream_t)stream_;
    if (P < 0) return FB_E_INVALID;
    if (P == 0) return FB_OK;
    if (!means3D || !lo || !ext || !codes) return FB_E_INVALID;
    { FB_PROF(FB_K_MAPUPDATE); k_morton30<<<(P + 255) / 256, 256, 0, st>>>(P, means3D, lo, ext, codes); }
    FB_LAUNCH_CHECK();
    return FB_OK;
}

size_t fb_occ_workspace_bytes(int gx, int gz) {
    const size_t n = (size_t)(gx > 0 ? gx : 1) * (gz > 0 ? gz : 1);
    return fb::align_up(n * 4, 256) * 2 + fb::align_up(n, 256) + 256;
}

int fb_occ_update(float* occ_map, int gx, int gz, const float* depth, int width, int height, int downsample,
                  const float* c2w, const float* intr, const float* planes, const float* map_center, float cell_size,
                  float height_lower, float height_upper, float far_distance, int* cam_cell, void* workspace,
                  size_t workspace_bytes, void* stream_) {
    using namespace fb;
    cudaStream_t st = (cudaStream_t)stream_;
    if (!occ_map || gx <= 0 || gz <= 0 || !depth || width <= 0 || height <= 0 || downsample <= 0 || !c2w || !intr ||
        !planes || !map_center || !(cell_size > 0.f) || !workspace) return FB_E_INVALID;
    if (workspace_bytes < fb_occ_workspace_bytes(gx, gz)) return FB_E_WORKSPACE;
    const size_t n = (size_t)gx * gz;
    char* ws = (char*)workspace;
    OccArgs a;
    a.occ_map = occ_map; a.gx = gx; a.gz = gz; a.depth = depth; a.W = width; a.H = height; a.ds = downsample;
    a.c2w = c2w; a.intr = intr; a.planes = planes; a.map_center = map_center;
    a.cell = cell_size; a.h_lo = height_lower; a.h_hi = height_upper; a.far = far_distance;
    a.cnt_free = (int*)ws; a.cnt_occ = (int*)(ws + align_up(n * 4, 256));
    a.canvas = (unsigned char*)(ws + 2 * align_up(n * 4, 256));
    a.cam_cell = (int*)(ws + 2 * align_up(n * 4, 256) + align_up(n, 256));
    FB_CUDA(cudaMemsetAsync(ws, 0, 2 * align_up(n * 4, 256) + align_up(n, 256), st));
    const long long pts = (long long)((width + downsample - 1) / downsample) * ((height + downsample - 1) / downsample) * 11;
    { FB_PROF(FB_K_MAPUPDATE); k_occ_points<<<(unsigned)((pts + 255) / 256), 256, 0, st>>>(a); }
    FB_LAUNCH_CHECK();
    { FB_PROF(FB_K_MAPUPDATE); k_occ_lines<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a); }
    { FB_PROF(FB_K_MAPUPDATE); k_occ_fuse<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a); }
    FB_LAUNCH_CHECK();
    if (cam_cell) FB_CUDA(cudaMemcpyAsync(cam_cell, a.cam_cell, 8, cudaMemcpyDeviceToDevice, st));
    return FB_OK;
}

// ---- fb_occ_update_static: the occupancy update with its scratch in a file-scope device array ----------------------
// For callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc): same kernels
// and arguments as fb_occ_update, but the counters / canvas / camera cell live in g_occ_static_scratch, which every GPU
// that loads this library holds until exit.  No counterpart in the reference.
// The array's name is only a host-side handle: its device address is looked up once per device (with that
// device current) through cudaGetSymbolAddress and handed to the kernels.  Calls share the array and it keeps the last
// call's contents: the used part is cleared on the stream first, and two calls that overlap on one GPU race.
#define FB_OCC_STATIC_BYTES (2 * (size_t)FB_OCC_STATIC_MAX_CELLS * 4 + (size_t)FB_OCC_STATIC_MAX_CELLS + 256)
static __device__ __align__(256) unsigned char g_occ_static_scratch[FB_OCC_STATIC_BYTES];
static void* g_occ_static_ptr[64] = {nullptr};

int fb_occ_update_static(float* occ_map, int gx, int gz, const float* depth, int width, int height, int downsample,
                         const float* c2w, const float* intr, const float* planes, const float* map_center,
                         float cell_size, float height_lower, float height_upper, float far_distance, int* cam_cell,
                         void* stream_) {
    using namespace fb;
    cudaStream_t st = (cudaStream_t)stream_;
    if (!occ_map || gx <= 0 || gz <= 0 || !depth || width <= 0 || height <= 0 || downsample <= 0 || !c2w || !intr ||
        !planes || !map_center || !(cell_size > 0.f)) return FB_E_INVALID;
    if ((long long)gx * gz > (long long)FB_OCC_STATIC_MAX_CELLS) return FB_E_STATIC_LIMIT;
    int dev = -1;
    FB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return FB_E_INVALID;
    if (!g_occ_static_ptr[dev]) {
        void* p = nullptr;
        FB_CUDA(cudaGetSymbolAddress(&p, g_occ_static_scratch));
        if (!p) return FB_E_CUDA;
        g_occ_static_ptr[dev] = p;
    }
    const size_t n = (size_t)gx * gz;
    char* ws = (char*)g_occ_static_ptr[dev];
    OccArgs a;
    a.occ_map = occ_map; a.gx = gx; a.gz = gz; a.depth = depth; a.W = width; a.H = height; a.ds = downsample;
    a.c2w = c2w; a.intr = intr; a.planes = planes; a.map_center = map_center;
    a.cell = cell_size; a.h_lo = height_lower; a.h_hi = height_upper; a.far = far_distance;
    a.cnt_free = (int*)ws; a.cnt_occ = (int*)(ws + align_up(n * 4, 256));
    a.canvas = (unsigned char*)(ws + 2 * align_up(n * 4, 256));
    a.cam_cell = (int*)(ws + 2 * align_up(n * 4, 256) + align_up(n, 256));
    FB_CUDA(cudaMemsetAsync(ws, 0, 2 * align_up(n * 4, 256) + align_up(n, 256), st));
    const long long pts = (long long)((width + downsample - 1) / downsample) * ((height + downsample - 1) / downsample) * 11;
    // (no FB_PROF here: the profiling scope creates events when enabled; the launches are counted by hand)
    k_occ_points<<<(unsigned)((pts + 255) / 256), 256, 0, st>>>(a);
    FB_LAUNCH_CHECK();
    k_occ_lines<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a);
    k_occ_fuse<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a);
    FB_LAUNCH_CHECK();
    g_launches += 3;
    if (cam_cell) FB_CUDA(cudaMemcpyAsync(cam_cell, a.cam_cell, 8, cudaMemcpyDeviceToDevice, st));
    return FB_OK;
}

int fb_count_vs_mask(int N, const float* means3D, const float* logit_opacities, const float* w2c, const float* intr,
                     const uint8_t* mask, int width, int height, float opacity_thresh, uint8_t* code, int* counts,
                     void* stream_) {
    using namespace fb;
    cudaStream_t st = (cudaStream_t)stream_;
    if (N < 0 || !counts || width <= 0 || height <= 0) return FB_E_INVALID;
    FB_CUDA(cudaMemsetAsync(counts, 0, 3 * sizeof(int), st));
    if (N == 0) return FB_OK;
    if (!means3D || !logit_opacities || !w2c || !intr || !mask || !code) return FB_E_INVALID;
    { FB_PROF(FB_K_VISIBLE); k_count_vs_mask<<<(N + 255) / 256, 256, 0, st>>>(N, means3D, logit_opacities, w2c, intr, mask,
                                                                            width, height, opacity_thresh, code, counts); }
    FB_LAUNCH_CHECK();
    return FB_OK;
}

}  // extern "C"
