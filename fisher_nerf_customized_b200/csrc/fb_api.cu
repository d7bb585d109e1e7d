// fb_api.cu -- the C ABI (include/fisher_b200.h) over the sm_100a kernels.  No torch types here.
#include "fb_score.cuh"
#include "fb_render_warp.cuh"
// FB_SCATTER_CTA (default): keys are scattered by k_scatter_keys_cta (shared-memory aggregation: one returning global
// atomic per (CTA, tile)) whenever the view has <= 1024 tiles; 0 keeps the warp-aggregated k_scatter_keys everywhere.
#ifndef FB_SCATTER_CTA
#define FB_SCATTER_CTA 1
#endif
#ifndef FB_SC_GRID
#define FB_SC_GRID 2              // CTAs per SM and view in the grid of k_scatter_keys_cta
#endif
#include "fb_backward.cuh"
#include "fb_stage5.cuh"
#include "fb_blocks.cuh"
#include "fb_mapupdate.cuh"
#include "fb_mapnbr.cuh"
#include <limits.h>
#include <stdlib.h>
#include <string.h>

using namespace fb;

static thread_local int g_last_cuda = 0;
#define FB_CUDA(call)                                                   \
    do {                                                                \
        cudaError_t e__ = (call);                                       \
        if (e__ != cudaSuccess) { g_last_cuda = (int)e__; return FB_E_CUDA; } \
    } while (0)
#define FB_LAUNCH_CHECK() FB_CUDA(cudaGetLastError())

// ---- optional per-kernel timing (fb_profile_*): CUDA events recorded on the launching stream around
// every kernel of a class while enabled; read back (with a sync) by fb_profile_read.  Off by default.
#include <vector>
enum { FB_K_PREPROCESS = 0, FB_K_SCAN, FB_K_SCATTER, FB_K_SORT, FB_K_RENDER, FB_K_BWD_TILES, FB_K_EXPAND,
       FB_K_TILE_SCORE, FB_K_FINALIZE, FB_K_KNN, FB_K_VISIBLE, FB_K_BLOCKS, FB_K_MAPUPDATE, FB_K_CLASSES };
static const char* const g_class_names[FB_K_CLASSES] = {"k_preprocess", "k_scan_tiles", "k_scatter_keys", "k_sort_tiles",
    "k_render_fwd", "k_bwd_tiles", "k_expand", "k_tile_score", "k_score_finalize", "k_knn", "k_count_visible", "k_block", "k_unproject_append"};
struct ProfRec { int cls; cudaEvent_t a, b; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof;
static std::vector<cudaEvent_t> g_ev_pool;
static cudaEvent_t prof_event() {
    cudaEvent_t e;
    if (!g_ev_pool.empty()) { e = g_ev_pool.back(); g_ev_pool.pop_back(); return e; }
    cudaEventCreate(&e);
    return e;
}
static unsigned long long g_launches = 0;      // kernels of this library launched so far (every launch site sits in a ProfScope)
struct ProfScope {
    int cls; cudaStream_t st; cudaEvent_t a;
    ProfScope(int c, cudaStream_t s) : cls(c), st(s), a(nullptr) {
        ++g_launches;
        if (g_prof_on) { a = prof_event(); cudaEventRecord(a, st); }
    }
    ~ProfScope() {
        if (a) { cudaEvent_t b = prof_event(); cudaEventRecord(b, st); g_prof.push_back({cls, a, b}); }
    }
};
#define FB_PROF(cls) ProfScope prof_scope__(cls, st)

extern "C" {

const char* fb_error_string(int s) {
    switch (s) {
        case FB_OK: return "ok";
        case FB_E_INVALID: return "invalid argument";
        case FB_E_WORKSPACE: return "workspace too small";
        case FB_E_CUDA: return "CUDA error";
        case FB_E_CAPACITY: return "instance capacity exceeded";
        case FB_E_PREFILTERED: return "Point is filtered although prefiltered is set";
        case FB_E_STATIC_LIMIT: return "size exceeds the static scratch of the no-allocation entry point (FB_OCC_STATIC_MAX_CELLS)";
        default: return "unknown";
    }
}
int fb_last_cuda_error(void) { return g_last_cuda; }
int fb_version(void) { return 200; }

int fb_profile_enable(int on) { g_prof_on = on != 0; return FB_OK; }
unsigned long long fb_launch_count(void) { return g_launches; }
int fb_profile_classes(void) { return FB_K_CLASSES; }
const char* fb_profile_class_name(int cls) { return (cls >= 0 && cls < FB_K_CLASSES) ? g_class_names[cls] : ""; }
int fb_profile_read(float* ms_total, int* launches) {
    for (int i = 0; i < FB_K_CLASSES; ++i) { if (ms_total) ms_total[i] = 0.f; if (launches) launches[i] = 0; }
    for (ProfRec& r : g_prof) {
        FB_CUDA(cudaEventSynchronize(r.b));
        float ms = 0.f;
        FB_CUDA(cudaEventElapsedTime(&ms, r.a, r.b));
        if (ms_total) ms_total[r.cls] += ms;
        if (launches) launches[r.cls] += 1;
        g_ev_pool.push_back(r.a); g_ev_pool.push_back(r.b);
    }
    g_prof.clear();
    return FB_OK;
}

size_t fb_geom_bytes(int P) { return geom_layout(P).total; }
size_t fb_image_bytes(int w, int h) { return image_layout(w, h).total; }
// sorted keys + an equally sized ping-pong area for tile lists longer than the shared-memory sort capacity
size_t fb_inst_bytes(int R) { return 2 * align_up((size_t)(R > 0 ? R : 1) * 8, 256); }
size_t fb_forward_scratch_bytes(int, int, int, int) { return 256; }
// moment rows (power 1 / 2, precomputed colours) or the per-pair fallback's chain rows (any other power, SH colours)
size_t fb_backward_scratch_bytes(int P, int, int power, int sh) {
    const size_t per = (power == 1 || power == 2) && !sh ? (size_t)FB_NACC * 4 : (size_t)FB_ROW_FLOATS * 4;
    return align_up((size_t)(P > 0 ? P : 1) * per, 256);
}

}  // extern "C"

static int check_gauss(const fb_gaussians* g) {
    if (!g || g->P < 0) return FB_E_INVALID;
    if (g->P == 0) return FB_OK;
    if (!g->means3D || !g->opacities) return FB_E_INVALID;
    if ((g->shs == nullptr) == (g->colors_precomp == nullptr)) return FB_E_INVALID;
    const bool sr = g->scales != nullptr && g->rotations != nullptr;
    if (sr == (g->cov3D_precomp != nullptr)) return FB_E_INVALID;
    if (!sr && (g->scales || g->rotations)) return FB_E_INVALID;
    return FB_OK;
}
static GaussIn to_in(const fb_gaussians* g) {
    GaussIn i;
    i.means3D = g->means3D; i.opacities = g->opacities; i.shs = g->shs; i.colors = g->colors_precomp;
    i.scales = g->scales; i.rotations = g->rotations; i.cov3D = g->cov3D_precomp;
    return i;
}
static Bins single_bins(int P, const Cam& cam, void* geom, void* image, void* inst, long long cap) {
    GeomLayout GL = geom_layout(P);
    ImageLayout IL = image_layout(cam.W, cam.H);
    char* gb = (char*)geom; char* ib = (char*)image;
    Bins b;
    b.recs = (GRec*)(gb + GL.recs); b.recsB = nullptr; b.recA = nullptr; b.recB = nullptr;
    b.vis_list = (uint4*)(gb + GL.vis_list); b.vis_count = (uint32_t*)(gb + GL.vis_count);
    b.vis_stride = (size_t)(P > 0 ? P : 1); b.key_is_slot = 0;
    b.view_total = (u64*)(ib + IL.totals + 64);            // zeroed with the totals (256 B) before every scan
    b.tile_count = (uint32_t*)(ib + IL.tile_count); b.tile_range = (uint2*)(ib + IL.tile_range);
    b.tile_order = (uint32_t*)(ib + IL.tile_order);
    b.total = (uint32_t*)(ib + IL.totals);
    b.keys = (u64*)inst; b.sub_list = nullptr; b.seg_list = nullptr; b.seg_flags = nullptr; b.capacity = cap; b.P = P; b.V = 1; b.tiles = cam.tiles_x * cam.tiles_y;
    return b;
}

// Keys of a tile segment that are staged in shared memory by k_sort_tiles; longer segments are sorted through L2 with
// the same shared memory as the bucket array.  Small on purpose: the kernel is latency-bound and the L2 path is as
// fast per key, so occupancy wins (4096 -> 1024: 930 -> 410 us per 64 views).
#ifndef FB_SORT_SMEM_KEYS
#define FB_SORT_SMEM_KEYS 1024
#endif
#define FB_SORT_SMEM_BYTES ((FB_SORT_STAGE ? 2 : 1) * FB_SORT_SMEM_KEYS * 8)
static_assert(FB_SORT_SMEM_BYTES >= FB_SPLIT_FINE * 4, "split_segment keeps its 4096 bucket counters in the sort's dynamic shared memory");
static_assert(FB_SPLIT_MAXJ <= (int)(sizeof(fb::SortShared::work) / 4), "the piece table of split_segment lives in SortShared::work");

extern "C" {

int fb_raster_forward_bin(const fb_camera* c, const fb_gaussians* g, void* geom, size_t geom_bytes, void* image,
                          size_t image_bytes, void* scratch, size_t scratch_bytes, int* radii, int* num_rendered_host,
                          void* stream_) {
    (void)scratch; (void)scratch_bytes;
    cudaStream_t st = (cudaStream_t)stream_;
    if (!c || !num_rendered_host || c->image_width <= 0 || c->image_height <= 0 || c->image_width > 4080 ||
        c->image_height > 4080) return FB_E_INVALID;   // tile coordinates are packed in 8 bits
    int rc = check_gauss(g);
    if (rc) return rc;
    *num_rendered_host = 0;
    const int P = g->P;
    if (P == 0) return FB_OK;
    if (!geom || !image || !radii) return FB_E_INVALID;
    if (g->shs && (c->sh_coeffs <= 0 || c->sh_degree < 0 || c->sh_degree > 3 ||
                   (c->sh_degree + 1) * (c->sh_degree + 1) > c->sh_coeffs)) return FB_E_INVALID;
    if (geom_bytes < fb_geom_bytes(P) || image_bytes < fb_image_bytes(c->image_width, c->image_height)) return FB_E_WORKSPACE;
    Cam cam = make_cam(c);
    Bins b = single_bins(P, cam, geom, image, nullptr, LLONG_MAX);
    GeomLayout GL = geom_layout(P);
    FB_CUDA(cudaMemsetAsync(b.vis_count, 0, 256, st));
    FB_CUDA(cudaMemsetAsync(b.tile_count, 0, (size_t)b.tiles * 4 * FB_TC_STRIDE, st));
    FB_CUDA(cudaMemsetAsync(b.total, 0, 256, st));
    { FB_PROF(FB_K_PREPROCESS); k_preprocess<false, 0><<<(P + 255) / 256, 256, 0, st>>>(to_in(g), cam, nullptr, b, radii,
                                                           (float*)((char*)geom + GL.rgb), (bool*)((char*)geom + GL.clamped), nullptr); }
    FB_LAUNCH_CHECK();
    { FB_PROF(FB_K_SCAN); k_scan_tiles<<<1, 1024, 0, st>>>(b); }
    FB_LAUNCH_CHECK();
    uint32_t tot[4];
    FB_CUDA(cudaMemcpyAsync(tot, b.total, sizeof(tot), cudaMemcpyDeviceToHost, st));
    FB_CUDA(cudaStreamSynchronize(st));   // the reference syncs here too (rasterizer_impl.cu:282)
    if (tot[2]) return FB_E_PREFILTERED;
    *num_rendered_host = (int)tot[0];
    return FB_OK;
}

int fb_raster_forward_render(const fb_camera* c, const fb_gaussians* g, void* geom, size_t geom_bytes, void* image,
                             size_t image_bytes, void* inst, size_t inst_bytes, void* scratch, size_t scratch_bytes,
                             int num_rendered, float* out_color, float* out_depth, void* stream_) {
    return fb_raster_forward_render2(c, g, nullptr, geom, geom_bytes, image, image_bytes, inst, inst_bytes, scratch,
                                     scratch_bytes, num_rendered, out_color, nullptr, out_depth, stream_);
}

int fb_raster_forward_render2(const fb_camera* c, const fb_gaussians* g, const float* extra_colors, void* geom,
                              size_t geom_bytes, void* image, size_t image_bytes, void* inst, size_t inst_bytes,
                              void* scratch, size_t scratch_bytes, int num_rendered, float* out_color, float* out_extra,
                              float* out_depth, void* stream_) {
    (void)scratch; (void)scratch_bytes; (void)geom_bytes; (void)image_bytes;
    cudaStream_t st = (cudaStream_t)stream_;
    if (!c || !g || !out_color || !out_depth || num_rendered < 0) return FB_E_INVALID;
    if ((extra_colors == nullptr) != (out_extra == nullptr)) return FB_E_INVALID;
    const int P = g->P;
    if (P == 0) return FB_OK;
    if (!geom || !image || !inst) return FB_E_INVALID;
    if (inst_bytes < fb_inst_bytes(num_rendered)) return FB_E_WORKSPACE;
    Cam cam = make_cam(c);
    Bins b = single_bins(P, cam, geom, image, inst, (long long)num_rendered);
    GeomLayout GL = geom_layout(P);
    ImageLayout IL = image_layout(cam.W, cam.H);
    if (num_rendered > 0) {
        int sblocks = (P + 255) / 256; if (sblocks > 148 * 8) sblocks = 148 * 8;
#if FB_SCATTER_CTA
        if (b.tiles <= FB_SC_MAX_TILES) { FB_PROF(FB_K_SCATTER); k_scatter_keys_cta<<<dim3(sblocks, 1), 256, 0, st>>>(b, cam); }
        else
#endif
        { FB_PROF(FB_K_SCATTER); k_scatter_keys<<<dim3(sblocks, 1), 256, 0, st>>>(b, cam); }
        FB_LAUNCH_CHECK();
        static bool attr_set = false;
        if (!attr_set) {
            FB_CUDA(cudaFuncSetAttribute(k_sort_tiles, cudaFuncAttributeMaxDynamicSharedMemorySize, FB_SORT_SMEM_BYTES));
            attr_set = true;
        }
        { FB_PROF(FB_K_SORT); k_sort_tiles<<<b.tiles, 256, FB_SORT_SMEM_BYTES, st>>>(
            b, (u64*)((char*)inst + align_up((size_t)num_rendered * 8, 256)), FB_SORT_SMEM_KEYS, b.tiles, 0); }
        FB_LAUNCH_CHECK();
    }
    const float* feat = g->colors_precomp ? g->colors_precomp : (const float*)((char*)geom + GL.rgb);
    if (extra_colors)
        { FB_PROF(FB_K_RENDER); k_render_warp<1><<<dim3(b.tiles * (FB_TILE_WARPS / FB_SCORE_WPC), 1), 32 * FB_SCORE_WPC, FB_SCORE_SMEM_BYTES, st>>>(
            b, cam, feat, extra_colors, (float*)((char*)image + IL.final_T), (uint32_t*)((char*)image + IL.n_contrib), out_color, out_extra, out_depth); }
    else
        { FB_PROF(FB_K_RENDER); k_render_warp<0><<<dim3(b.tiles * (FB_TILE_WARPS / FB_SCORE_WPC), 1), 32 * FB_SCORE_WPC, FB_SCORE_SMEM_BYTES, st>>>(
            b, cam, feat, nullptr, (float*)((char*)image + IL.final_T), (uint32_t*)((char*)image + IL.n_contrib), out_color, nullptr, out_depth); }
    FB_LAUNCH_CHECK();
    return FB_OK;
}

int fb_raster_backward(const fb_camera* c, const fb_gaussians* g, const int* radii, const void* geom, size_t geom_bytes,
                       const void* image, size_t image_bytes, const void* inst, size_t inst_bytes, void* scratch,
                       size_t scratch_bytes, int num_rendered, const float* dL_dpix, int power, float* dL_dmeans2D,
                       float* dL_dconic, float* dL_dopacity, float* dL_dcolors, float* dL_dmeans3D, float* dL_dcov3D,
                       float* dL_dsh, float* dL_dscales, float* dL_drotations, void* stream_) {
    return fb_raster_backward2(c, g, nullptr, radii, geom, geom_bytes, image, image_bytes, inst, inst_bytes, scratch,
                               scratch_bytes, num_rendered, dL_dpix, nullptr, power, dL_dmeans2D, dL_dconic, dL_dopacity,
                               dL_dcolors, nullptr, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations, stream_);
}

int fb_raster_backward2(const fb_camera* c, const fb_gaussians* g, const float* extra_colors, const int* radii,
                        const void* geom, size_t geom_bytes, const void* image, size_t image_bytes, const void* inst,
                        size_t inst_bytes, void* scratch, size_t scratch_bytes, int num_rendered, const float* dL_dpix,
                        const float* dL_dextra, int power, float* dL_dmeans2D, float* dL_dconic, float* dL_dopacity,
                        float* dL_dcolors, float* dL_dextra_colors, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                        float* dL_dscales, float* dL_drotations, void* stream_) {
    (void)geom_bytes; (void)image_bytes; (void)inst_bytes;
    cudaStream_t st = (cudaStream_t)stream_;
    if (!c || !dL_dpix) return FB_E_INVALID;
    if ((extra_colors == nullptr) != (dL_dextra == nullptr)) return FB_E_INVALID;
    if (extra_colors && power != 1) return FB_E_INVALID;   // the two renders' gradients add only for power 1
    int rc = check_gauss(g);
    if (rc) return rc;
    const int P = g->P;
    if (P == 0 || num_rendered == 0) return FB_OK;
    if (!geom || !image || !inst || !radii || !scratch) return FB_E_INVALID;
    const bool pairwise = (power != 1 && power != 2) || (g->shs && c->sh_degree > 0);
    if (pairwise && extra_colors) return FB_E_INVALID;
    if (scratch_bytes < fb_backward_scratch_bytes(P, num_rendered, power, pairwise ? 1 : 0)) return FB_E_WORKSPACE;
    Cam cam = make_cam(c);
    Bins b = single_bins(P, cam, (void*)geom, (void*)image, (void*)inst, (long long)num_rendered);
    GeomLayout GL = geom_layout(P);
    ImageLayout IL = image_layout(cam.W, cam.H);
    if (pairwise) {
        // any integer power / SH degree > 0: the reference's per-pair definition (fb_backward.cuh: k_bwd_pairs)
        float* rows = (float*)scratch;
        const float* featp = g->colors_precomp ? g->colors_precomp : (const float*)((const char*)geom + GL.rgb);
        ShArgs sh{g->shs, (const bool*)((const char*)geom + GL.clamped), c->sh_degree, c->sh_coeffs, nullptr};
        // campos lives in the camera struct; k_chain_rows reads it from device-visible kernel parameters
        Cam* cam_ptr = nullptr; (void)cam_ptr;
        static_assert(sizeof(Cam) < 4000, "Cam travels as a kernel parameter");
        { FB_PROF(FB_K_EXPAND); k_chain_rows<<<(P + 127) / 128, 128, 0, st>>>(P, to_in(g), cam, b.recs, radii, sh, rows); }
        FB_LAUNCH_CHECK();
        PairOut po{dL_dmeans2D, dL_dconic, dL_dopacity, dL_dcolors, dL_dmeans3D, dL_dcov3D, dL_dscales, dL_drotations,
                   (g->shs && c->sh_degree > 0) ? dL_dsh : nullptr};
        { FB_PROF(FB_K_BWD_TILES); k_bwd_pairs<<<dim3(b.tiles * (FB_TILE_WARPS / FB_SCORE_WPC), 1), 32 * FB_SCORE_WPC, FB_SCORE_SMEM_BYTES, st>>>(
            b, cam, featp, (const float*)((const char*)image + IL.final_T), (const uint32_t*)((const char*)image + IL.n_contrib),
            dL_dpix, rows, power, c->sh_coeffs, g->shs ? c->sh_degree : 0, (g->scales && g->rotations) ? 1 : 0, po); }
        FB_LAUNCH_CHECK();
        return FB_OK;
    }
    float* mom = (float*)scratch;
    FB_CUDA(cudaMemsetAsync(mom, 0, (size_t)P * FB_NACC * 4, st));
    const float* feat = g->colors_precomp ? g->colors_precomp : (const float*)((const char*)geom + GL.rgb);
    const float* fT = (const float*)((const char*)image + IL.final_T);
    const uint32_t* nc = (const uint32_t*)((const char*)image + IL.n_contrib);
    dim3 grid(b.tiles * (FB_TILE_WARPS / FB_SCORE_WPC), 1);
    GradOut o{dL_dmeans2D, dL_dconic, dL_dopacity, dL_dcolors, dL_dmeans3D, dL_dcov3D, dL_dscales, dL_drotations, dL_dextra_colors};
    if (power == 2) {
        { FB_PROF(FB_K_BWD_TILES); k_bwd_warp<true, 0><<<grid, 32 * FB_SCORE_WPC, FB_SCORE_SMEM_BYTES, st>>>(b, cam, feat, nullptr, fT, nc, dL_dpix, nullptr, mom); }
        FB_LAUNCH_CHECK();
        { FB_PROF(FB_K_EXPAND); k_expand_grads<true><<<(P + 127) / 128, 128, 0, st>>>(P, to_in(g), cam, b.recs, radii, mom, o); }
    } else {
        if (extra_colors) {
            { FB_PROF(FB_K_BWD_TILES); k_bwd_warp<false, 1><<<grid, 32 * FB_SCORE_WPC, FB_SCORE_SMEM_BYTES, st>>>(b, cam, feat, extra_colors, fT, nc, dL_dpix, dL_dextra, mom); }
            FB_LAUNCH_CHECK();
            { FB_PROF(FB_K_EXPAND); k_expand_grads<false, true><<<(P + 127) / 128, 128, 0, st>>>(P, to_in(g), cam, b.recs, radii, mom, o); }
            FB_LAUNCH_CHECK();
            return FB_OK;
        }
        { FB_PROF(FB_K_BWD_TILES); k_bwd_warp<false, 0><<<grid, 32 * FB_SCORE_WPC, FB_SCORE_SMEM_BYTES, st>>>(b, cam, feat, nullptr, fT, nc, dL_dpix, nullptr, mom); }
        FB_LAUNCH_CHECK();
        { FB_PROF(FB_K_EXPAND); k_expand_grads<false><<<(P + 127) / 128, 128, 0, st>>>(P, to_in(g), cam, b.recs, radii, mom, o); }
    }
    FB_LAUNCH_CHECK();
    return FB_OK;
}

int fb_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present,
                    void* stream_) {
    if (P < 0 || !viewmatrix || !projmatrix) return FB_E_INVALID;
    if (P == 0) return FB_OK;
    if (!means3D || !present) return FB_E_INVALID;
    Cam cam; memset(&cam, 0, sizeof(cam));
    for (int i = 0; i < 16; ++i) { cam.view[i] = viewmatrix[i]; cam.proj[i] = projmatrix[i]; }
    cudaStream_t st = (cudaStream_t)stream_;
    { FB_PROF(FB_K_PREPROCESS); k_mark_visible<<<(P + 255) / 256, 256, 0, st>>>(P, means3D, cam, present); }
    FB_LAUNCH_CHECK();
    return FB_OK;
}

}  // extern "C"

#include "fb_api_batched.inl"
