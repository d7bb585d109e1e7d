/*
 * fisher_b200.h -- C ABI of the B200-native FisherRF candidate-view scoring path.
 *
 * Every entry point takes raw DEVICE pointers (unless the name says host), plain
 * sizes, an explicit cudaStream_t (passed as void*) and returns an int status
 * (FB_OK == 0, negative on error; fb_error_string() gives the text).  No torch
 * types cross this boundary; all buffers are allocated by the caller.
 *
 * Reference interfaces replaced (paths relative to the upstream repo,
 * DGR = thirdparty/diff-gaussian-rasterization-modified, KNN = thirdparty/simple-knn):
 *
 *   fb_raster_forward_bin + fb_raster_forward_render
 *        <- CudaRasterizer::Rasterizer::forward      DGR/cuda_rasterizer/rasterizer.h:29-56,
 *           bound by RasterizeGaussiansCUDA          DGR/rasterize_points.cu:35-115
 *   fb_raster_backward
 *        <- CudaRasterizer::Rasterizer::backward     DGR/cuda_rasterizer/rasterizer.h:58-84,
 *           bound by RasterizeGaussiansBackwardCUDA  DGR/rasterize_points.cu:117-196
 *   fb_mark_visible
 *        <- CudaRasterizer::Rasterizer::markVisible  DGR/cuda_rasterizer/rasterizer.h:24-27
 *   fb_dist2_knn3
 *        <- SimpleKNN::knn(int P, float3*, float*)   KNN/simple_knn.h:18
 *   fb_count_visible
 *        <- count_visible_points (numpy)             tester_navigator.py:147-180
 *   fb_score_candidates / fb_accumulate_hessian   (additive, batched)
 *        <- the Python loops GaussianSLAM.pose_eval / compute_H_train / compute_Hessian
 *           models/SLAM/gaussian.py:1338-1375,1503-1570 (11-column variant
 *           models/SLAM/gaussian_object.py:1940-2045)
 *
 *   fb_block_outer_accumulate / fb_block_opt   (POp-GS block form)
 *        <- estimate_block_JtJ, t_opt_blocks, d_opt_blocks, pose_eval_popgs_blocks
 *           models/SLAM/gaussian_object.py:1660-1732, 2112-2176
 *   fb_unproject_append   (map update, densification)
 *        <- get_pointcloud + initialize_new_params + the torch.cat growth of add_new_gaussians
 *           models/SLAM/gaussian.py:75-142, 299-318, 397-401
 *
 * Threading: one host thread per process drives the library (one process per GPU,
 * as the reference runs).  Launches are asynchronous on the caller's stream; the
 * profiling switch, the chunk-lane stream and the kernel-attribute set-up are
 * process-wide state without locks.
 *
 * The reference hands the rasteriser three std::function<char*(size_t)> resize
 * callbacks (rasterizer.h:32-34).  That is not C-ABI safe; here the caller asks
 * for sizes (fb_*_bytes) and allocates.  The one size that is only known on the
 * device, num_rendered, is returned by fb_raster_forward_bin to the host exactly
 * where the reference reads it back (rasterizer_impl.cu:282).
 */
#ifndef FISHER_B200_H
#define FISHER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FB_OK                 0
#define FB_E_INVALID        (-1)  /* bad argument (null pointer, non-positive size, unsupported combination) */
#define FB_E_WORKSPACE      (-2)  /* caller-provided workspace too small */
#define FB_E_CUDA           (-3)  /* a CUDA runtime call or launch failed; see fb_last_cuda_error() */
#define FB_E_CAPACITY       (-4)  /* device-side instance capacity exceeded; *needed tells how much */
#define FB_E_PREFILTERED    (-5)  /* prefiltered=1 but a point was culled (reference: device __trap, auxiliary.h:156-160) */
#define FB_E_STATIC_LIMIT   (-6)  /* fb_occ_update_static: gx * gz exceeds FB_OCC_STATIC_MAX_CELLS */

/* Compile-time raster constants, identical to DGR/cuda_rasterizer/config.h:15-17. */
#define FB_NUM_CHANNELS 3
#define FB_TILE_X 16
#define FB_TILE_Y 16

/* Camera + raster constants: the numeric fields of GaussianRasterizationSettings
 * (DGR/diff_gaussian_rasterization/__init__.py:140-151).  Matrices are the 16
 * floats exactly as the reference passes them, i.e. the row-major memory of the
 * TRANSPOSED matrix (column-major for the kernels; auxiliary.h:58-77). */
typedef struct fb_camera {
    int   image_width;
    int   image_height;
    float tanfovx;
    float tanfovy;
    float scale_modifier;
    int   sh_degree;       /* D: active SH degree 0..3 */
    int   sh_coeffs;       /* M: coefficients per Gaussian in `shs` (0 if none) */
    int   prefiltered;
    float bg[3];
    float campos[3];
    float viewmatrix[16];
    float projmatrix[16];
} fb_camera;

/* Gaussian inputs of one rasterisation call.  Exactly one of shs / colors_precomp
 * and exactly one of (scales, rotations) / cov3D_precomp is non-null, as enforced
 * by GaussianRasterizer.forward (__init__.py:174-178). */
typedef struct fb_gaussians {
    int          P;
    const float* means3D;        /* [P,3] */
    const float* opacities;      /* [P]   */
    const float* shs;            /* [P,M,3] or NULL */
    const float* colors_precomp; /* [P,3]   or NULL */
    const float* scales;         /* [P,3]   or NULL */
    const float* rotations;      /* [P,4] (w,x,y,z; used un-normalised) or NULL */
    const float* cov3D_precomp;  /* [P,6]   or NULL */
} fb_gaussians;

const char* fb_error_string(int status);
int         fb_last_cuda_error(void);     /* cudaError_t of the last FB_E_CUDA, as int */
int         fb_version(void);
/* Number of kernels this library has launched in this process so far (every launch site counts itself). */
unsigned long long fb_launch_count(void);

/* Optional per-kernel timing for bench.py's roofline: while enabled, CUDA events are recorded on the
 * launching stream around every kernel launch; fb_profile_read synchronises on them, sums the elapsed
 * milliseconds and launch counts per kernel class (arrays of fb_profile_classes() entries) and resets. */
int         fb_profile_enable(int on);
int         fb_profile_classes(void);
const char* fb_profile_class_name(int cls);
int         fb_profile_read(float* ms_total, int* launches);

/* ---- single-view drop-in path --------------------------------------------- */

/* Per-call state sizes.  `geom` holds the per-Gaussian state kept for backward
 * (the role of the reference's geomBuffer, rasterizer_impl.h:30-46); `image` the
 * per-pixel + per-tile state (imgBuffer, :48-57); `inst` the sorted per-instance
 * records (binningBuffer, :59-70).  Layouts are private to this library. */
size_t fb_geom_bytes(int P);
size_t fb_image_bytes(int width, int height);
size_t fb_inst_bytes(int num_rendered);
/* Scratch needed only during fb_raster_forward_* (unsorted keys, scan temp). */
size_t fb_forward_scratch_bytes(int P, int width, int height, int num_rendered);
/* Scratch needed only during fb_raster_backward (per-instance moment rows). */
size_t fb_backward_scratch_bytes(int P, int num_rendered, int power, int has_sh);

/* Stage 1 + first half of stage 2: preprocess all Gaussians, count instances per
 * Gaussian and per tile, scan.  Writes radii[P] (int32).  Blocks until the total
 * instance count is on the host and stores it in *num_rendered_host. */
int fb_raster_forward_bin(const fb_camera* cam, const fb_gaussians* g,
                          void* geom, size_t geom_bytes,
                          void* image, size_t image_bytes,
                          void* scratch, size_t scratch_bytes,
                          int* radii,
                          int* num_rendered_host,
                          void* stream);

/* Second half of stage 2 + stage 3: emit (depth|gaussian) keys per tile, sort each
 * tile, gather the sorted per-instance records, alpha-blend.  Writes
 * out_color[3,H,W], out_depth[1,H,W]. */
int fb_raster_forward_render(const fb_camera* cam, const fb_gaussians* g,
                             void* geom, size_t geom_bytes,
                             void* image, size_t image_bytes,
                             void* inst, size_t inst_bytes,
                             void* scratch, size_t scratch_bytes,
                             int num_rendered,
                             float* out_color, float* out_depth,
                             void* stream);

/* Stage 4, drop-in form.  All nine outputs must be ZERO-FILLED by the caller
 * (the reference glue uses torch::zeros, rasterize_points.cu:151-159) and have
 * the reference's shapes: dL_dmeans2D[P,3], dL_dconic[P,2,2], dL_dopacity[P],
 * dL_dcolors[P,3], dL_dmeans3D[P,3], dL_dcov3D[P,6], dL_dsh[P,M,3],
 * dL_dscales[P,3], dL_drotations[P,4].  `power` is backward_power
 * (__init__.py:154): every per-(pixel,Gaussian) partial is raised to it before
 * accumulation (backward.cu:1093-1137). */
int fb_raster_backward(const fb_camera* cam, const fb_gaussians* g,
                       const int* radii,
                       const void* geom, size_t geom_bytes,
                       const void* image, size_t image_bytes,
                       const void* inst, size_t inst_bytes,
                       void* scratch, size_t scratch_bytes,
                       int num_rendered,
                       const float* dL_dpix,   /* [3,H,W] */
                       int power,
                       float* dL_dmeans2D, float* dL_dconic, float* dL_dopacity,
                       float* dL_dcolors, float* dL_dmeans3D, float* dL_dcov3D,
                       float* dL_dsh, float* dL_dscales, float* dL_drotations,
                       void* stream);

/* Fused second render.  SplaTAM-style mapping / tracking rasterises every iteration twice with the
 * same geometry: colours, then colors_precomp = [depth, 1, depth^2] for depth + silhouette (models/SLAM/gaussian.py:
 * 207,211; models/SLAM/utils/slam_helpers.py:235-252,268-279).  These take the second colour triple along:
 * preprocess, binning and sort run once (fb_raster_forward_bin), one sweep blends six channels, one backward sweep
 * (power 1 only: the two renders' gradients add) returns the summed gradients of the shared inputs, the gradients of
 * both colour triples, and dL_dmeans2D of the COLOUR render alone (the densification statistic, gaussian.py:208).
 * extra_colors [P,3], out_extra / dL_dextra [3,H,W]; NULL extra_colors = the plain calls above. */
int fb_raster_forward_render2(const fb_camera* cam, const fb_gaussians* g, const float* extra_colors,
                              void* geom, size_t geom_bytes, void* image, size_t image_bytes,
                              void* inst, size_t inst_bytes, void* scratch, size_t scratch_bytes,
                              int num_rendered, float* out_color, float* out_extra, float* out_depth, void* stream);
int fb_raster_backward2(const fb_camera* cam, const fb_gaussians* g, const float* extra_colors, const int* radii,
                        const void* geom, size_t geom_bytes, const void* image, size_t image_bytes,
                        const void* inst, size_t inst_bytes, void* scratch, size_t scratch_bytes,
                        int num_rendered, const float* dL_dpix, const float* dL_dextra, int power,
                        float* dL_dmeans2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolors,
                        float* dL_dextra_colors, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                        float* dL_dscales, float* dL_drotations, void* stream);

/* present[i] = (viewmatrix * p_i).z > 0.001  (rasterizer_impl.cu:54-66). */
int fb_mark_visible(int P, const float* means3D, const float* viewmatrix,
                    const float* projmatrix, uint8_t* present, void* stream);

/* Introspection for parity tests: copies of the private state in the
 * reference's vocabulary.  Any output pointer may be NULL.
 *   tiles_touched[P] u32, depths[P], means2D[P,2], conic_opacity[P,4] (geom)
 *   ranges[tiles,2] u32, final_T[HW], n_contrib[HW] u32            (image)
 *   point_list[R] u32 (sorted Gaussian ids), point_keys[R] u64 (tile<<32|depth bits) (inst) */
int fb_debug_export(const fb_camera* cam, int P, int num_rendered,
                    const void* geom, const void* image, const void* inst,
                    uint32_t* tiles_touched, float* depths, float* means2D, float* conic_opacity,
                    uint32_t* ranges, float* final_T, uint32_t* n_contrib,
                    uint32_t* point_list, uint64_t* point_keys,
                    void* stream);

/* ---- batched candidate scoring (resident map, many poses, no host sync) ---- */

/* The resident Gaussian map in the activated form compute_Hessian feeds the
 * rasteriser (gaussian.py:1523-1543): world-frame means, NORMALISED rotations,
 * sigmoid opacities, exp scales, rgb colours. */
typedef struct fb_map {
    int          P;
    const float* means3D;    /* [P,3] world frame */
    const float* opacities;  /* [P]   */
    const float* scales;     /* [P,3] */
    const float* rotations;  /* [P,4] */
    const float* colors;     /* [P,3] */
} fb_map;

/* Workspace for scoring `views_per_chunk` (1..64) poses at a time with room for
 * `inst_capacity` (Gaussian,tile) instances per chunk and `vis_capacity` visible Gaussians (radii > 0) per
 * view (0 or >= P: P).  Per-view records are stored in the view's visible slot, so the workspace is
 * 112 * views_per_chunk * vis_capacity bytes plus keys, not 112 * views_per_chunk * P.  The map may be given in
 * any order; a spatially coherent order (the Python GaussianMap stores it Morton-sorted) lets the kernels
 * aggregate their tile counters per warp and is ~20 % faster, with the same results. */
size_t fb_score_workspace_bytes(int P, int width, int height, int views_per_chunk,
                                long long inst_capacity, long long vis_capacity);
/* Workspace of fb_accumulate_hessian (per-view records and moment rows are indexed by Gaussian id). */
size_t fb_hessian_workspace_bytes(int P, int width, int height, int views_per_chunk,
                                  long long inst_capacity);

/* Chunk lanes.  If `workspace_bytes` >= 2 * fb_score_workspace_bytes(...) and K > views_per_chunk,
 * fb_score_candidates alternates consecutive chunks between the caller's stream and a library-owned
 * stream, each in its own half of the workspace (the latency-bound binning kernels of one chunk run
 * under the issue-bound kernels of the other); the
 * caller's stream waits for both before the call's work is complete on it.  Results do not change.
 * fb_score_set_lanes(0) keeps every chunk on the caller's stream; so does fb_profile_enable(1), so that
 * its event pairs time kernels alone.  Library state (the second stream, function attributes) is kept per
 * device and belongs to the device that is CURRENT when the call is made: make the device of the buffers
 * current before calling (one host thread per device). */
int fb_score_set_lanes(int on);

/* scores[k] = sum_{g,c} H_k[g,c] * weight[g,c]   for k in [0,K)
 * where H_k is the Fisher diagonal of pose k: the grads compute_Hessian
 * concatenates (C_H = 4: d/dmeans3D(3), d/dopacity(1), gaussian.py:1554-1560;
 * C_H = 11: + d/dscales(3) + d/drotations(4), gaussian_object.py:2024-2032) with
 * backward_power = 2 and dL_dpix = grad_scale (1e-3 in the reference, gaussian.py:1549),
 * and weight = 1/(H_train+lambda) (gaussian.py:1357,1367).
 * w2c: K row-major 4x4 world-to-camera matrices (the poses AFTER torch.linalg.inv,
 * gaussian.py:1364); means are moved to the camera frame with them while rotations
 * stay in the world frame, exactly as the reference does (gaussian.py:1527-1529).
 * status_dev (8 ints): [0] 0 or FB_E_CAPACITY (inst_capacity or vis_capacity too small: grow and call again),
 * [1] the instance count the largest chunk needed, [2] the longest per-tile list seen, [3] total
 * (Gaussian,tile) instances over all K poses (the reference's num_rendered summed), [4] total visible
 * Gaussians (radii > 0) over all K poses, [5] the largest visible count of a single pose (what vis_capacity
 * must hold), [6..7] reserved; read it whenever the scores are read.  No host sync inside. */
int fb_score_candidates(const fb_camera* cam, const fb_map* map,
                        const float* w2c, int K,
                        const float* weight, int C_H, float grad_scale,
                        float* scores, int* status_dev,
                        void* workspace, size_t workspace_bytes,
                        int views_per_chunk, long long inst_capacity, long long vis_capacity,
                        void* stream);

/* Path evaluation (plan_best_path / path_evaluation, tester_gaussians_navigation.py:1659-1727, 1886-1975): every
 * candidate path carries its own running Hessian, so the poses of one launch score against different weights.
 * `weight` is [sets][P][C_H]; pose k uses tensor weight_set[k] (i32 [K], device; NULL = tensor 0 for all). */
int fb_score_candidates_sets(const fb_camera* cam, const fb_map* map,
                             const float* w2c, int K,
                             const float* weight, const int* weight_set, int C_H, float grad_scale,
                             float* scores, int* status_dev,
                             void* workspace, size_t workspace_bytes,
                             int views_per_chunk, long long inst_capacity, long long vis_capacity,
                             void* stream);

/* H_out[g,c] += sum_k H_k[g,c]   (compute_H_train, gaussian.py:1338-1348).
 * H_out is [P,C_H], caller-initialised.  Keyframes are processed views_per_chunk (<= 64) per launch
 * sequence; the per-Gaussian sum over the views of a chunk is formed in a fixed order (no atomics on H_out).
 * Workspace: fb_hessian_workspace_bytes; same status convention as fb_score_candidates. */
int fb_accumulate_hessian(const fb_camera* cam, const fb_map* map,
                          const float* w2c, int K,
                          int C_H, float grad_scale,
                          float* H_out, int* status_dev,
                          void* workspace, size_t workspace_bytes,
                          int views_per_chunk, long long inst_capacity,
                          void* stream);

/* As fb_accumulate_hessian, with H_out = [sets][P][C_H] and the term of pose k added to tensor out_set[k]
 * (i32 [K], device; NULL = tensor 0): the per-path `H_train_path + cur_H` of path evaluation for all paths at once. */
int fb_accumulate_hessian_sets(const fb_camera* cam, const fb_map* map,
                               const float* w2c, int K,
                               int C_H, float grad_scale,
                               float* H_out, const int* out_set, int* status_dev,
                               void* workspace, size_t workspace_bytes,
                               int views_per_chunk, long long inst_capacity,
                               void* stream);

/* ---- stage 5 ------------------------------------------------------------- */

size_t fb_knn_workspace_bytes(int P);
/* mean_dist2[i] = mean of the 3 smallest squared distances from point i to any
 * other point (self excluded by index), FLT_MAX terms if P < 4
 * (KNN/simple_knn.cu:148-184). */
int fb_dist2_knn3(int P, const float* points, float* mean_dist2,
                  void* workspace, size_t workspace_bytes, void* stream);

/* counts[k] = #{ i : z_cam > 0 and 0 <= u < img_w and 0 <= v < img_h } for pose k,
 * evaluated in float64 like numpy does (tester_navigator.py:160-178).
 * points: f32 [N,3]; w2c: f64 K x 16 row-major (already inverted on the host);
 * intr: f64 {fx, fy, cx, cy}.  mask (optional, may be NULL): u8 [K,N]. */
int fb_count_visible(int N, const float* points, int K, const double* w2c,
                     const double* intr, int img_w, int img_h,
                     int* counts, uint8_t* mask, void* stream);

/* ---- POp-GS block form (models/SLAM/gaussian_object.py) ------------------- */

/* estimate_block_JtJ's accumulation (:2160-2174), one probe per call:
 *   H_blocks[i] (d x d row-major) += g g^T,  g = row rows[i] of
 *   [g_mean[P,3] | g_op[P,1]? | g_rot[P,4]? | g_scale[P,3]?]   (NULL = column group absent,
 *   the reference's use_opacity / use_rot / use_scale switches), d = 3 (+1) (+4) (+3).
 * The product is rounded before the add like the reference's tensor expression; if
 * divide_by != 0 every element is multiplied by fl(1/divide_by) after the add (`H_blocks / float(K)`
 * as torch evaluates it on a CUDA tensor; folded into the last probe).  rows: i64 [Nv] device. */
int fb_block_outer_accumulate(int Nv, const long long* rows,
                              const float* g_mean, const float* g_op,
                              const float* g_rot, const float* g_scale,
                              float divide_by, float* H_blocks, void* stream);

/* t_opt_blocks (:1720-1724, criterion 0):  -sum_b trace((Hm_b + J_b + lam I)^-1)
 * d_opt_blocks (:1726-1732, criterion 1):  sum_b logabsdet(Hm_b + lam I + J_b) - logabsdet(Hm_b + lam I)
 * over n_blocks pairs (Hm_blocks[idx_m[b]], J_blocks[idx_j[b]]); idx_* may be NULL (identity):
 * they are the reference's `Hm_blocks[idx_train]`, `Jb[idx_cur]` gathers (:1688-1689).
 * The FP32 matrix is formed as the reference forms it and factorised (LU, partial pivoting)
 * in FP64.  partial: f64 [fb_block_opt_partials(n_blocks)] device, the score is their sum;
 * per_block (optional, may be NULL): f32 [n_blocks] the per-block terms.  1 <= d <= 11. */
int fb_block_opt_partials(int n_blocks);
int fb_block_opt(int n_blocks, int d,
                 const float* Hm_blocks, const long long* idx_m,
                 const float* J_blocks, const long long* idx_j,
                 float lam, int criterion,
                 double* partial, float* per_block, void* stream);

/* ---- map update: densification from an RGB-D frame (models/SLAM/gaussian.py) ---- */

/* get_pointcloud (:75-142, transform_pts=True, "projective" mean_sq_dist) + initialize_new_params
 * (:299-318) for n selected pixels, written in place: row i of the outputs is pixel pix[i] of the
 * grid downsampled by `downsample` (row-major, row length ceil(width / downsample); pix == NULL
 * selects pixels 0..n-1).  The caller passes pointers to the FIRST NEW ROW of its parameter arrays
 * (the reference re-allocates every array with torch.cat, :399-401).
 *   means3D (stride mean_stride >= 3 floats), rgb (stride rgb_stride >= 3): required; with both
 *   pointing into one [n,6] array (strides 6) the result is get_pointcloud's point_cld.
 *   unnorm_rotations [n,4] = (1,0,0,0), logit_opacities [n] = 0, log_scales [n,log_scale_cols]
 *   = log(sqrt(mean_sq_dist)) (3 columns, or 1 if isotropic), mean3_sq_dist [n]: each optional (NULL).
 * color [3,H,W], depth [1,H,W], pix (i64), intr = {fx, fy, cx, cy}, c2w = 16 floats row-major (the
 * inverse of w2c, :105): all device pointers (no host round trip of the pose). */
int fb_unproject_append(int width, int height, int downsample,
                        const float* color, const float* depth,
                        const float* intr, const float* c2w,
                        const long long* pix, long long n,
                        float* means3D, int mean_stride, float* rgb, int rgb_stride,
                        float* unnorm_rotations, float* logit_opacities,
                        float* log_scales, int log_scale_cols, float* mean3_sq_dist,
                        void* stream);

/* codes[i] = 30-bit Morton code (10 bits per axis, x lowest) of means3D[i] quantised in the box [lo, lo + ext]
 * (lo, ext: 3 floats each, device): the order the resident map is kept in (see fb_score_workspace_bytes). */
int fb_morton_codes(int P, const float* means3D, const float* lo, const float* ext, int* codes, void* stream);

/* ---- occupancy grid and masked visibility (planning/astar.py, models/SLAM/gaussian_object.py) ---- */

/* AstarPlanner.update_occ_map (planning/astar.py:202-303) on the device, without its host loop:
 * occ_map [3,gz,gx] (unknown, occupied, free) is updated in place from one depth image [height,width]:
 * the 3x3 block around the camera cell is set to 1e3 on the free channel; 10 samples along every pixel's ray
 * (planes[0..9] * depth) are counted as free and the depth point (planes[10] = 1) as occupied, after the
 * height filter, per grid cell (what torch.unique(..., return_counts=True) returns); one 8-connected line per
 * occupied cell to the camera cell marks free space (cv2.line, thickness 1); the three channels are normalised
 * and added.  c2w (16), intr {fx,fy,cx,cy}, planes (11), map_center (2): device pointers, so the pose never
 * visits the host.  cam_cell (optional, 2 ints, device): the camera's (x, z) cell, the reference's cam_pos. */
size_t fb_occ_workspace_bytes(int gx, int gz);
int fb_occ_update(float* occ_map, int gx, int gz,
                  const float* depth, int width, int height, int downsample,
                  const float* c2w, const float* intr, const float* planes, const float* map_center,
                  float cell_size, float height_lower, float height_upper, float far_distance,
                  int* cam_cell, void* workspace, size_t workspace_bytes, void* stream);

/* fb_occ_update without any allocation or caller workspace.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, serving loops).  Same arguments and
 * results as fb_occ_update; the scratch is a file-scope device array inside the library (9.0 MiB on every GPU that loads
 * it, held until exit) sized for gx * gz <= FB_OCC_STATIC_MAX_CELLS cells (the reference's 768 x 768 grid fits);
 * larger grids return FB_E_STATIC_LIMIT.  The call itself uses only cudaMemsetAsync, kernel launches and one
 * device-to-device cudaMemcpyAsync on `stream`.  All calls on a GPU share the array: calls that overlap on the GPU
 * (different streams, or different host threads) race; serialise them. */
#define FB_OCC_STATIC_MAX_CELLS (1024 * 1024)
int fb_occ_update_static(float* occ_map, int gx, int gz,
                         const float* depth, int width, int height, int downsample,
                         const float* c2w, const float* intr, const float* planes, const float* map_center,
                         float cell_size, float height_lower, float height_upper, float far_distance,
                         int* cam_cell, void* stream);

/* count_gaussians_vs_mask (models/SLAM/gaussian_object.py:359-428): code[i] = 0 if sigmoid(logit_i) < thresh,
 * 1 if the Gaussian projects inside the mask, 2 if it projects into the image but outside the mask, 3 if it is
 * active but behind the camera or outside the image; counts = {in, out (codes 2 and 3), active}.
 * mask: u8 [height,width]; w2c: 16 floats row-major; intr {fx,fy,cx,cy}: device pointers. */
int fb_count_vs_mask(int N, const float* means3D, const float* logit_opacities,
                     const float* w2c, const float* intr, const uint8_t* mask, int width, int height,
                     float opacity_thresh, uint8_t* code, int* counts, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FISHER_B200_H */
