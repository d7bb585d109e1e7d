// fb_stage12.cuh -- stage 1 (Gaussian preprocess) and stage 2 (tile binning + depth sort).
//
// Replaces, for one view or a batch of V views sharing one resident map:
//   preprocessCUDA<3>          DGR/cuda_rasterizer/forward.cu:155-256
//   InclusiveSum + D2H sync    DGR/cuda_rasterizer/rasterizer_impl.cu:277-282
//   duplicateWithKeys          DGR/cuda_rasterizer/rasterizer_impl.cu:70-111
//   DeviceRadixSort::SortPairs DGR/cuda_rasterizer/rasterizer_impl.cu:304-309
//   identifyTileRanges         DGR/cuda_rasterizer/rasterizer_impl.cu:116-138
//
// B200 shape: instead of one global 41-bit radix sort (>= 6 passes over 12 B/instance) the
// instances are counted per tile while the Gaussian is still in registers, a single-CTA scan
// turns counts into segment offsets, the keys are scattered straight into their tile's
// segment, and each tile's segment (a few KB) is sorted once in shared memory by the CTA that
// owns the tile.  Sorting (depth_bits<<32 | gaussian_id) reproduces exactly the order of the
// reference's stable (tile|depth) radix sort: ascending depth, ties by ascending Gaussian id
// (the reference's emission order, rasterizer_impl.cu:92-108).
#pragma once
#include "fb_layout.cuh"
#include "fb_fisher.cuh"
#include <cuda_fp16.h>
#include <type_traits>

namespace fb {

// footprint half extents of the single-view record: two halves (rounded up) in GRec.b.w
__device__ __forceinline__ float pack_extents(float hx, float hy) {
    const __half2 h = __halves2half2(__float2half_ru(hx), __float2half_ru(hy));
    return __uint_as_float(*reinterpret_cast<const unsigned*>(&h));
}
__device__ __forceinline__ float2 unpack_extents(float w) {
    const unsigned u = __float_as_uint(w);
    return __half22float2(*reinterpret_cast<const __half2*>(&u));
}


// ---- SH -> RGB (forward.cu:20-71); coefficients are the standard real SH constants -------------
__device__ __forceinline__ float3 sh_to_rgb(int deg, int M, const float3& pos, const float* campos,
                                            const float* sh_g, bool* clamped3) {
    const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
    const float C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                         -1.0925484305920792f, 0.5462742152960396f};
    const float C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                         -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};
    float3 d = make_float3(pos.x - campos[0], pos.y - campos[1], pos.z - campos[2]);
    float len = sqrtf(d.x * d.x + d.y * d.y + d.z * d.z);
    d.x = d.x / len; d.y = d.y / len; d.z = d.z / len;
    float res[3];
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        const float* sh = sh_g + ch;   // sh[k] is sh_g[3*k + ch]
#define SHK(k) sh[3 * (k)]
        float r = C0 * SHK(0);
        if (deg > 0) {
            float x = d.x, y = d.y, z = d.z;
            r = r - C1 * y * SHK(1) + C1 * z * SHK(2) - C1 * x * SHK(3);
            if (deg > 1) {
                float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                r = r + C2[0] * xy * SHK(4) + C2[1] * yz * SHK(5) + C2[2] * (2.0f * zz - xx - yy) * SHK(6) +
                    C2[3] * xz * SHK(7) + C2[4] * (xx - yy) * SHK(8);
                if (deg > 2) {
                    r = r + C3[0] * y * (3.0f * xx - yy) * SHK(9) + C3[1] * xy * z * SHK(10) +
                        C3[2] * y * (4.0f * zz - xx - yy) * SHK(11) +
                        C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * SHK(12) +
                        C3[4] * x * (4.0f * zz - xx - yy) * SHK(13) + C3[5] * z * (xx - yy) * SHK(14) +
                        C3[6] * x * (xx - 3.0f * yy) * SHK(15);
                }
            }
        }
#undef SHK
        r += 0.5f;
        clamped3[ch] = (r < 0);
        res[ch] = fmaxf(r, 0.0f);
    }
    (void)M;
    return make_float3(res[0], res[1], res[2]);
}

// Gaussian inputs as device pointers (fb_gaussians / fb_map).
struct GaussIn {
    const float* means3D; const float* opacities; const float* shs; const float* colors;
    const float* scales; const float* rotations; const float* cov3D;
};

// ---- stage 1 -------------------------------------------------------------------------------
// One thread per Gaussian; the Gaussian (44 B: mean 12 + scale 12 + quat 16 + opacity 4) is
// loaded ONCE, its 3D covariance is built once, and it is then projected into every view of
// the batch while still in registers.  w2c (BATCHED only): V row-major 4x4 matrices; the mean
// is moved to the camera frame as torch's fp32 matmul does in the reference caller
// (models/SLAM/gaussian.py:1523-1527), the rotation stays in the world frame.
// CH > 0 (scoring): additionally emits the 64-byte scoring record {rgb, 13 polynomial coefficients}
// built from weight[g, 0..CH) = 1/(H_train+lambda)  (fb_fisher.cuh: score_coefs).
template <bool BATCHED, int CH>
__global__ void __launch_bounds__(256) k_preprocess(GaussIn g, Cam cam, const float* __restrict__ w2c, Bins bins,
                                                    int* __restrict__ radii, float* __restrict__ rgb,
                                                    bool* __restrict__ clamped, const float* __restrict__ weight) {
    extern __shared__ float s_w2c[];  // V*12
    const int P = bins.P, V = bins.V;
    if (BATCHED) {
        for (int i = threadIdx.x; i < V * 12; i += blockDim.x) s_w2c[i] = w2c[(i / 12) * 16 + (i % 12)];
        __syncthreads();
    }
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = idx < P;
    const int lane = threadIdx.x & 31;
    float3 p = make_float3(0.f, 0.f, 0.f);
    float cov6[6] = {0, 0, 0, 0, 0, 0};
    float opacity = 0.f;
    if (live) {
        p = make_float3(g.means3D[3 * idx], g.means3D[3 * idx + 1], g.means3D[3 * idx + 2]);
        opacity = g.opacities[idx];
        if (g.cov3D != nullptr) {
#pragma unroll
            for (int i = 0; i < 6; ++i) cov6[i] = g.cov3D[6 * (size_t)idx + i];
        } else {
            float3 sc = make_float3(g.scales[3 * idx], g.scales[3 * idx + 1], g.scales[3 * idx + 2]);
            float4 q = *reinterpret_cast<const float4*>(g.rotations + 4 * (size_t)idx);
            cov3d_from_scale_rot(sc, cam.scale_modifier, q, cov6);
        }
    }
    for (int v = blockIdx.y; v < V; v += gridDim.y) {      // batched launches deal the views over grid.y (shorter blocks)
        float3 pc = p;
        if (BATCHED) {
            const float* m = s_w2c + v * 12;
            pc.x = fmaf(m[2], p.z, fmaf(m[1], p.y, m[0] * p.x)) + m[3];
            pc.y = fmaf(m[6], p.z, fmaf(m[5], p.y, m[4] * p.x)) + m[7];
            pc.z = fmaf(m[10], p.z, fmaf(m[9], p.y, m[8] * p.x)) + m[11];
        }
        Splat s; int culled = 0; int vis = 0;
        if (live) vis = project_splat(pc, cov6, cam, s, culled);
        if (!BATCHED) {
            if (live) radii[idx] = vis ? s.radius : 0;
            if (live && culled && cam.prefiltered) bins.total[2] = 1u;
        }
        // warp-aggregated append to the view's visible list
        unsigned m = __ballot_sync(0xffffffffu, vis);
        if (m) {
            int leader = __ffs(m) - 1;
            unsigned base = 0;
            if (lane == leader) base = atomicAdd(&bins.vis_count[v], (unsigned)__popc(m));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (vis) {
                unsigned slot = base + __popc(m & ((1u << lane) - 1u));
                bins.vis_list[(size_t)v * bins.vis_stride + slot] =
                    make_uint4((unsigned)idx, __float_as_uint(s.depth),
                               (unsigned)s.x0 | ((unsigned)s.y0 << 8) | ((unsigned)s.x1 << 16) | ((unsigned)s.y1 << 24), 0u);
                GRec r;
                if (CH > 0 || BATCHED) {
                    // scoring / batched-Hessian layout: half extents of the region where alpha = o*exp(power) can reach 1/255
                    // (d^T Q d <= 2 ln(255 o)); a pair outside it is skipped by the reference's own
                    // alpha < 1/255 test (forward.cu:352, backward.cu:974), so culling with it is exact.
                    // Margin: 1e-3 relative + 0.05 px, far above the rounding of pair_power / expf.
                    float hx = -1.f, hy = -1.f;                       // o < 1/255: never contributes
                    if (opacity >= 1.0f / 255.0f) {
                        const float tau2 = 2.0f * logf(255.0f * opacity) + 1e-3f;
                        const float dq = s.ca * s.cc - s.cb * s.cb;
                        hx = sqrtf(tau2 * s.cc / dq) * 1.001f + 0.05f;
                        hy = sqrtf(tau2 * s.ca / dq) * 1.001f + 0.05f;
                        if (!(dq > 0.f) || !(hx == hx) || !(hy == hy)) { hx = 3.0e38f; hy = 3.0e38f; }   // never cull
                    }
                    r.a = make_float4(s.px, s.py, hx, hy);
                    r.b = make_float4(s.ca, s.cb, s.cc, opacity);
                } else {
                    // single-view layout: depth and opacity stay in the record (forward needs both), the footprint
                    // half extents (as above) travel as two halves rounded up
                    float hx = -1.f, hy = -1.f;
                    if (opacity >= 1.0f / 255.0f) {
                        const float tau2 = 2.0f * logf(255.0f * opacity) + 1e-3f;
                        const float dq = s.ca * s.cc - s.cb * s.cb;
                        hx = sqrtf(tau2 * s.cc / dq) * 1.001f + 0.05f;
                        hy = sqrtf(tau2 * s.ca / dq) * 1.001f + 0.05f;
                        if (!(dq > 0.f) || !(hx == hx) || !(hy == hy)) { hx = 3.0e38f; hy = 3.0e38f; }   // never cull
                    }
                    r.a = make_float4(s.px, s.py, s.depth, opacity);
                    r.b = make_float4(s.ca, s.cb, s.cc, pack_extents(hx, hy));
                }
                bins.recs[(size_t)v * P + idx] = r;
                if (CH > 0) {
                    float3 sc = make_float3(g.scales[3 * idx], g.scales[3 * idx + 1], g.scales[3 * idx + 2]);
                    float4 q = *reinterpret_cast<const float4*>(g.rotations + 4 * (size_t)idx);
                    SRCtx sr = make_sr(sc, cam.scale_modifier, q);
                    GCtx ctx = make_ctx(pc, cov6, cam);
                    float wv[CH > 0 ? CH : 1];
#pragma unroll
                    for (int k = 0; k < CH; ++k) wv[k] = weight[(size_t)idx * CH + k];
                    float coef[13];
                    score_coefs<CH>(ctx, cam, sr, s.ca, s.cb, s.cc, opacity, wv, coef);
                    float4* rb = bins.recsB + ((size_t)v * P + idx) * 4;
                    rb[0] = make_float4(g.colors[3 * idx], g.colors[3 * idx + 1], g.colors[3 * idx + 2], coef[0]);
                    rb[1] = make_float4(coef[1], coef[2], coef[3], coef[4]);
                    rb[2] = make_float4(coef[5], coef[6], coef[7], coef[8]);
                    rb[3] = make_float4(coef[9], coef[10], coef[11], coef[12]);
                }
                uint32_t* tc = bins.tile_count + (size_t)v * bins.tiles * FB_TC_STRIDE;
                for (int y = s.y0; y < s.y1; ++y)
                    for (int x = s.x0; x < s.x1; ++x) atomicAdd(&tc[(y * cam.tiles_x + x) * FB_TC_STRIDE], 1u);
                if (!BATCHED && g.colors == nullptr) {
                    bool cl[3];
                    float3 c = sh_to_rgb(cam.sh_degree, cam.sh_coeffs, p, cam.campos,
                                         g.shs + (size_t)idx * cam.sh_coeffs * 3, cl);
                    rgb[3 * idx] = c.x; rgb[3 * idx + 1] = c.y; rgb[3 * idx + 2] = c.z;
                    clamped[4 * idx] = cl[0]; clamped[4 * idx + 1] = cl[1]; clamped[4 * idx + 2] = cl[2];
                }
            }
        }
    }
}

// Warp-aggregated update of per-tile counters.  Every lane brings a tile rectangle (empty if x0 >= x1).  When the
// union of the warp's rectangles is small (a spatially ordered map: the 32 Gaussians of a warp land on a few
// tiles) the warp walks the union, one ballot per tile, and the first participating lane issues ONE atomic for
// all of them; otherwise every lane walks its own rectangle.  f(t, n, leader) performs the atomic on the leader
// and returns its result to every lane; the callback g(t, value, rank) runs on the participating lanes.
#define FB_AGG_MAX_UNION 24
#define FB_AGG_BIG_RECT 8
template <bool COOP, class Atomic, class Each>
__device__ __forceinline__ void warp_tiles(int x0, int y0, int x1, int y1, int tiles_x, u64 payload, Atomic atom, Each each) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const bool any = x0 < x1 && y0 < y1;
    const int ux0 = __reduce_min_sync(full, any ? x0 : 0x7fffffff), uy0 = __reduce_min_sync(full, any ? y0 : 0x7fffffff);
    const int ux1 = __reduce_max_sync(full, any ? x1 : -1), uy1 = __reduce_max_sync(full, any ? y1 : -1);
    if (ux1 <= ux0) return;                                   // nobody has a tile
    if ((ux1 - ux0) * (uy1 - uy0) <= FB_AGG_MAX_UNION) {
        for (int y = uy0; y < uy1; ++y)
            for (int x = ux0; x < ux1; ++x) {
                const bool in = any && x >= x0 && x < x1 && y >= y0 && y < y1;
                const unsigned m = __ballot_sync(full, in);
                if (m == 0u) continue;
                const int leader = __ffs(m) - 1;
                const int t = y * tiles_x + x;
                unsigned r = 0;
                if (lane == leader) r = atom(t, (unsigned)__popc(m));
                r = __shfl_sync(full, r, leader);
                if (in) each(t, r, (unsigned)__popc(m & ((1u << lane) - 1u)), payload);
            }
        return;
    }
    // scattered warp: small rectangles stay with their lane; a large one (a Gaussian close to the camera can
    // cover every tile) is spread over the 32 lanes so that no single thread walks hundreds of tiles
    // (COOP: only where the atomic returns a value the thread waits for; fire-and-forget counters do not stall)
    const bool big = COOP && any && (x1 - x0) * (y1 - y0) > FB_AGG_BIG_RECT;
    if (any && !big) {
        for (int y = y0; y < y1; ++y)
            for (int x = x0; x < x1; ++x) {
                const int t = y * tiles_x + x;
                each(t, atom(t, 1u), 0u, payload);
            }
    }
    unsigned bm = COOP ? __ballot_sync(full, big) : 0u;
    while (bm) {
        const int src = __ffs(bm) - 1;
        bm &= bm - 1u;
        const int sx0 = __shfl_sync(full, x0, src), sy0 = __shfl_sync(full, y0, src);
        const int sx1 = __shfl_sync(full, x1, src), sy1 = __shfl_sync(full, y1, src);
        const u64 pl = ((u64)__shfl_sync(full, (unsigned)(payload >> 32), src) << 32) |
                       (u64)__shfl_sync(full, (unsigned)payload, src);
        const int w = sx1 - sx0, cnt = w * (sy1 - sy0);
        for (int j = lane; j < cnt; j += 32) {
            const int t = (sy0 + j / w) * tiles_x + sx0 + j % w;
            each(t, atom(t, 1u), 0u, pl);
        }
    }
}

// ---- stage 1, scoring form -------------------------------------------------------------------
// Same outputs as k_preprocess<true, CH> but split in two phases per group of FB_PS_VG views, because
// only ~30 % of the (Gaussian, view) pairs are visible and the 13-coefficient build costs ~10x the projection:
//   phase 1  thread = Gaussian: project into each view, count tiles, append to the view's visible list, write
//            the 32-byte GRec, and push (local Gaussian, view) onto a queue in shared memory;
//   phase 2  thread = queue entry: rebuild the conic of that pair from the Gaussian's parameters (kept in
//            shared memory; same expressions, same bits) and run score_coefs with every lane busy.
#ifndef FB_PS_VG
#define FB_PS_VG 16
#endif
#define FB_PS_THREADS 256
#ifndef FB_PS_IDENT
#define FB_PS_IDENT 1         // written-out products for an identity view matrix in the coefficient phase
#endif
#ifndef FB_PS_FAST_COEF
#define FB_PS_FAST_COEF 1      // approximate divisions in the coefficient phase (phase 2) of k_preprocess_score
#endif
#ifndef FB_PS_EARLY_CULL
#define FB_PS_EARLY_CULL 1     // conservative per-warp visibility test before the projection (fb_common.cuh: maybe_on_screen)
#endif
#ifndef FB_PS_SPLIT
#define FB_PS_SPLIT 4          // view groups of one Gaussian group go to up to this many blocks (grid.y)
#endif
// conic (as project_splat forms it) and the 13 scoring coefficients of one visible (Gaussian, view) pair
template <int CH, bool FC, bool ID>
__device__ __forceinline__ void build_coefs(const float3& pc, const float* c6, const Cam& cam, const SRCtx& sr, float op,
                                            const float* wv, float* coef, float& ca, float& cb, float& cc) {
    const GCtx ctx = make_ctx<FC, ID>(pc, c6, cam);
    const float det = __fmaf_rn(ctx.c2.a, ctx.c2.c, -__fmul_rn(ctx.c2.b, ctx.c2.b));
    const float det_inv = fdiv<FC>(1.f, det);
    ca = __fmul_rn(ctx.c2.c, det_inv); cb = __fmul_rn(-ctx.c2.b, det_inv); cc = __fmul_rn(ctx.c2.a, det_inv);
    score_coefs<CH, FC, ID>(ctx, cam, sr, ca, cb, cc, op, wv, coef);
}

template <int CH> struct PSmem {
    float p[3][FB_PS_THREADS], cov6[6][FB_PS_THREADS], op[FB_PS_THREADS], sc[3][FB_PS_THREADS], q[4][FB_PS_THREADS],
        col[3][FB_PS_THREADS], w[CH][FB_PS_THREADS];
    float w2c[FB_PS_VG * 12];
    unsigned qslot[FB_PS_THREADS * FB_PS_VG];          // visible slot of the queued (Gaussian, view) pair
    unsigned short queue[FB_PS_THREADS * FB_PS_VG];
    unsigned qcount;
};

// Records are written in the view's VISIBLE SLOT (the position the warp-aggregated atomicAdd on vis_count hands out),
// not at the Gaussian's index: the arrays are dense ([V][vis_stride], vis_stride = the caller's visible capacity), a
// warp's visible lanes own consecutive slots (coalesced 16-B rows in phase 1, contiguous 80-B rows in phase 2), and the
// scoring kernel's tile-list scan touches two records per 32-B sector.  A view with more visible Gaussians than
// vis_stride raises the overflow flag in k_scan_tiles (status FB_E_CAPACITY, the caller grows and re-runs).
template <int CH>
#ifndef FB_PS_MINB
#define FB_PS_MINB 3
#endif
#ifndef FB_PS_MINB11
#define FB_PS_MINB11 3          // C_H = 11 (2: 128 registers, no spills, 907 us; 3: 80 registers, 184 B of spills, 890 us per 64 views)
#endif
__global__ void __launch_bounds__(FB_PS_THREADS, (CH > 4 ? FB_PS_MINB11 : FB_PS_MINB)) k_preprocess_score(GaussIn g, Cam cam, const float* __restrict__ w2c,
                                                                    Bins bins, const float* __restrict__ weight,
                                                                    const int* __restrict__ weight_set) {
    // weight_set (optional): view v scores against weight tensor weight_set[v] of a [sets][P][CH] array (path
    // evaluation: every path carries its own running Hessian); NULL: one tensor for all views, staged in shared memory
    extern __shared__ __align__(16) unsigned char ps_raw[];
    PSmem<CH>& S = *reinterpret_cast<PSmem<CH>*>(ps_raw);
    const int P = bins.P, V = bins.V;
    const int tid = threadIdx.x, lane = tid & 31;
    const int idx = blockIdx.x * FB_PS_THREADS + tid;
    const bool live = idx < P;
    const unsigned vcap = (unsigned)bins.vis_stride;
    float3 p = make_float3(0.f, 0.f, 0.f);
    float cov6[6] = {0, 0, 0, 0, 0, 0};
    float opacity = 0.f;
    if (live) {
        p = make_float3(g.means3D[3 * idx], g.means3D[3 * idx + 1], g.means3D[3 * idx + 2]);
        opacity = g.opacities[idx];
        const float3 sc = make_float3(g.scales[3 * idx], g.scales[3 * idx + 1], g.scales[3 * idx + 2]);
        const float4 q = *reinterpret_cast<const float4*>(g.rotations + 4 * (size_t)idx);
        cov3d_from_scale_rot(sc, cam.scale_modifier, q, cov6);
        S.p[0][tid] = p.x; S.p[1][tid] = p.y; S.p[2][tid] = p.z;
#pragma unroll
        for (int i = 0; i < 6; ++i) S.cov6[i][tid] = cov6[i];
        S.op[tid] = opacity;
        S.sc[0][tid] = sc.x; S.sc[1][tid] = sc.y; S.sc[2][tid] = sc.z;
        S.q[0][tid] = q.x; S.q[1][tid] = q.y; S.q[2][tid] = q.z; S.q[3][tid] = q.w;
#pragma unroll
        for (int k = 0; k < 3; ++k) S.col[k][tid] = g.colors[3 * idx + k];
#pragma unroll
        for (int k = 0; k < CH; ++k) S.w[k][tid] = weight[(size_t)idx * CH + k];
    }
    const float tr_sigma = (cov6[0] + cov6[3]) + cov6[5];
    // gridDim.y blocks share a Gaussian group and deal its view groups round-robin: 4x more, 4x shorter blocks,
    // so the last wave of the grid is a small fraction of the kernel
    for (int v0 = blockIdx.y * FB_PS_VG; v0 < V; v0 += gridDim.y * FB_PS_VG) {
        const int nv = min(FB_PS_VG, V - v0);
        __syncthreads();                                   // previous group's phase 2 is done with w2c / queue
        for (int i = tid; i < nv * 12; i += FB_PS_THREADS) S.w2c[i] = w2c[(size_t)(v0 + i / 12) * 16 + (i % 12)];
        if (tid == 0) S.qcount = 0u;
        __syncthreads();
        // ---- phase 1
        for (int vv = 0; vv < nv; ++vv) {
            const int v = v0 + vv;
            const float* m = S.w2c + vv * 12;
            float3 pc;
            pc.x = fmaf(m[2], p.z, fmaf(m[1], p.y, m[0] * p.x)) + m[3];
            pc.y = fmaf(m[6], p.z, fmaf(m[5], p.y, m[4] * p.x)) + m[7];
            pc.z = fmaf(m[10], p.z, fmaf(m[9], p.y, m[8] * p.x)) + m[11];
            Splat s; int culled = 0; int vis = 0;
#if FB_PS_EARLY_CULL
            // most (warp, view) pairs of a spatially ordered map see nothing: a ~60-instruction conservative test
            // instead of the ~350-instruction projection decides that for the whole warp
            const bool maybe = live && maybe_on_screen(pc, tr_sigma, cam);
            if (!__any_sync(0xffffffffu, maybe)) continue;
            if (maybe) vis = project_splat(pc, cov6, cam, s, culled);
#else
            if (live) vis = project_splat(pc, cov6, cam, s, culled);
#endif
            const unsigned bm = __ballot_sync(0xffffffffu, vis);
            if (bm == 0u) continue;
            {
                uint32_t* tc = bins.tile_count + (size_t)v * bins.tiles * FB_TC_STRIDE;
                warp_tiles<false>(vis ? s.x0 : 0, vis ? s.y0 : 0, vis ? s.x1 : 0, vis ? s.y1 : 0, cam.tiles_x, 0ull,
                           [&](int t, unsigned n) { atomicAdd(&tc[t * FB_TC_STRIDE], n); return 0u; },
                           [&](int, unsigned, unsigned, u64) {});
            }
            const int leader = __ffs(bm) - 1;
            unsigned base = 0, qbase = 0;
            if (lane == leader) {
                base = atomicAdd(&bins.vis_count[v], (unsigned)__popc(bm));
                qbase = atomicAdd(&S.qcount, (unsigned)__popc(bm));
            }
            base = __shfl_sync(0xffffffffu, base, leader);
            qbase = __shfl_sync(0xffffffffu, qbase, leader);
            if (!vis) continue;
            const unsigned rank = __popc(bm & ((1u << lane) - 1u));
            const unsigned slot = base + rank;
            S.queue[qbase + rank] = (unsigned short)(tid | (vv << 8));
            S.qslot[qbase + rank] = slot;
            if (slot >= vcap) continue;                        // overflow: the chunk is void (k_scan_tiles flags it)
            bins.vis_list[(size_t)v * vcap + slot] =
                make_uint4((unsigned)idx, __float_as_uint(s.depth),
                           (unsigned)s.x0 | ((unsigned)s.y0 << 8) | ((unsigned)s.x1 << 16) | ((unsigned)s.y1 << 24), 0u);
            // half extents of the alpha >= 1/255 footprint (see k_preprocess)
            float hx = -1.f, hy = -1.f;
            if (opacity >= 1.0f / 255.0f) {
                const float tau2 = 2.0f * logf(255.0f * opacity) + 1e-3f;
                const float dq = s.ca * s.cc - s.cb * s.cb;
                hx = sqrtf(tau2 * s.cc / dq) * 1.001f + 0.05f;
                hy = sqrtf(tau2 * s.ca / dq) * 1.001f + 0.05f;
                if (!(dq > 0.f) || !(hx == hx) || !(hy == hy)) { hx = 3.0e38f; hy = 3.0e38f; }
            }
            bins.recA[(size_t)v * vcap + slot] = make_float4(s.px, s.py, hx, hy);
        }
        __syncthreads();
        // ---- phase 2
        const unsigned qn = S.qcount;
        for (unsigned i = tid; i < qn; i += FB_PS_THREADS) {
            const unsigned e = S.queue[i];
            const unsigned slot = S.qslot[i];
            if (slot >= vcap) continue;
            const int gi = e & 255, vv = e >> 8;
            const float* m = S.w2c + vv * 12;
            const float3 pw = make_float3(S.p[0][gi], S.p[1][gi], S.p[2][gi]);
            float3 pc;
            pc.x = fmaf(m[2], pw.z, fmaf(m[1], pw.y, m[0] * pw.x)) + m[3];
            pc.y = fmaf(m[6], pw.z, fmaf(m[5], pw.y, m[4] * pw.x)) + m[7];
            pc.z = fmaf(m[10], pw.z, fmaf(m[9], pw.y, m[8] * pw.x)) + m[11];
            float c6[6];
#pragma unroll
            for (int k = 0; k < 6; ++k) c6[k] = S.cov6[k][gi];
            // nothing this phase produces decides an integer any more (its conic is read by the pair sweep only), so
            // its divisions are approximate (FB_PS_FAST_COEF; the visible list, radii and keys come from phase 1) and an
            // identity view matrix (the scoring path's usual camera) takes the written-out products
            const SRCtx sr = make_sr(make_float3(S.sc[0][gi], S.sc[1][gi], S.sc[2][gi]), cam.scale_modifier,
                                     make_float4(S.q[0][gi], S.q[1][gi], S.q[2][gi], S.q[3][gi]));
            float wv[CH];
            if (weight_set == nullptr) {
#pragma unroll
                for (int k = 0; k < CH; ++k) wv[k] = S.w[k][gi];
            } else {
                const float* wsrc = weight + ((size_t)weight_set[v0 + vv] * P + (size_t)blockIdx.x * FB_PS_THREADS + gi) * CH;
#pragma unroll
                for (int k = 0; k < CH; ++k) wv[k] = wsrc[k];
            }
            float coef[13], ca, cb, cc;
            constexpr bool FC = FB_PS_FAST_COEF != 0;
            if constexpr (CH == 4 && FC && FB_PS_IDENT) {
                if (cam.view_ident) build_coefs<CH, FC, true>(pc, c6, cam, sr, S.op[gi], wv, coef, ca, cb, cc);
                else build_coefs<CH, FC, false>(pc, c6, cam, sr, S.op[gi], wv, coef, ca, cb, cc);
            } else {
                build_coefs<CH, FC, false>(pc, c6, cam, sr, S.op[gi], wv, coef, ca, cb, cc);
            }
            float4* rb = bins.recB + ((size_t)(v0 + vv) * vcap + slot) * 5;
            // the scoring kernel evaluates G = ex2(dx (A dx + B dy) + C dy^2): conic pre-multiplied by
            // -log2(e)/2, -log2(e), -log2(e)/2 (the sign of the exponent is kept); r+g+b is the one colour
            rb[0] = make_float4(ca * -0.72134752044448170368f, cb * -1.4426950408889634f, cc * -0.72134752044448170368f, S.op[gi]);
            rb[1] = make_float4((S.col[0][gi] + S.col[1][gi]) + S.col[2][gi], coef[0], 0.f, 0.f);
            rb[2] = make_float4(coef[1], coef[2], coef[3], coef[4]);
            rb[3] = make_float4(coef[5], coef[6], coef[7], coef[8]);
            rb[4] = make_float4(coef[9], coef[10], coef[11], coef[12]);
        }
    }
}

// ---- stage 2a: counts -> segment offsets (one CTA per view, chained) ---------------------------
// Empty tiles get range (0,0) as in the reference (cudaMemset + identifyTileRanges,
// rasterizer_impl.cu:311-319).  total[0] = instance count, total[1] = overflow flag,
// total[3] = largest tile segment, total[4] = largest visible count of a view.
// CTA v scans the counters of view v, publishes the view's instance total, then adds the totals of the views
// before it: CTAs of a grid are dispatched in index order, so every CTA this one waits for is resident or done
// (the chained-scan argument).  V <= 64 CTAs replace the single CTA that walked all V * tiles counters
// (45 us per 64 views on the critical path of every chunk).  bins.view_total and bins.total are zeroed per chunk.
// Work order for the sort and pair-sweep kernels: each view's tile slots by descending instance count (counting sort,
// 1024 buckets over [0, max of the view]), interleaved over the views -- tile_order[rank * V + v] -- so the heaviest
// tiles of every view start first.  The order only decides when a tile is scheduled, never what it computes.
__global__ void __launch_bounds__(1024) k_scan_tiles(Bins bins) {
    __shared__ unsigned warp_sums[32];
    __shared__ unsigned carry_s, maxc_s, base_s;
    __shared__ unsigned bucket[1024];
    const int v = blockIdx.x, V = bins.V, n = bins.tiles;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t* tc = bins.tile_count + (size_t)v * n * FB_TC_STRIDE;
    uint2* tr = bins.tile_range + (size_t)v * n;
    if (tid == 0) { carry_s = 0; maxc_s = 0; base_s = 0; }
    __syncthreads();
    unsigned local_max = 0;
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const unsigned c = (i < n) ? tc[(size_t)i * FB_TC_STRIDE] : 0u;
        local_max = max(local_max, c);
        unsigned x = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) warp_sums[wid] = x;
        __syncthreads();
        if (wid == 0) {
            unsigned w = warp_sums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
            warp_sums[lane] = w;
        }
        __syncthreads();
        const unsigned incl = x + (wid > 0 ? warp_sums[wid - 1] : 0u) + carry_s;
        if (i < n) tr[i] = make_uint2(incl - c, incl);          // relative to the view's first instance, for now
        __syncthreads();
        if (tid == 1023) carry_s = incl;
        __syncthreads();
    }
    local_max = __reduce_max_sync(0xffffffffu, local_max);
    if (lane == 0) atomicMax(&maxc_s, local_max);
    const unsigned my_total = carry_s;
    if (tid == 0) {
        atomicExch(reinterpret_cast<unsigned long long*>(&bins.view_total[v]), (1ull << 32) | (u64)my_total);
    }
    // instances of the views before this one
    if (tid < v) {
        const volatile u64* vt = reinterpret_cast<const volatile u64*>(bins.view_total);
        u64 x;
        do { x = vt[tid]; } while ((x >> 32) == 0ull);
        atomicAdd(&base_s, (unsigned)(x & 0xffffffffull));      // (sums of unsigned wrap alike in any order)
    }
    __syncthreads();
    const unsigned base_inst = base_s;
    const unsigned maxc_view = maxc_s;
    for (int i = tid; i < n; i += 1024) {
        const uint2 r = tr[i];
        tr[i] = (r.y > r.x) ? make_uint2(r.x + base_inst, r.y + base_inst) : make_uint2(0u, 0u);
    }
    if (tid == 0) {
        atomicMax(&bins.total[3], maxc_view);
        atomicMax(&bins.total[4], bins.vis_count[v]);
        if (bins.vis_count[v] > (unsigned)bins.vis_stride) atomicExch(&bins.total[1], 1u);
        if (v == V - 1) {
            const u64 all = (u64)base_inst + (u64)my_total;
            bins.total[0] = (unsigned)min(all, (u64)0xffffffffu);
            if ((long long)all > bins.capacity) atomicExch(&bins.total[1], 1u);
        }
    }
    if (bins.tile_order == nullptr) return;
    const unsigned maxc = maxc_view > 0u ? maxc_view : 1u;
    bucket[tid] = 0u;
    __syncthreads();
    for (int i = tid; i < n; i += 1024) {
        const unsigned c = tc[(size_t)i * FB_TC_STRIDE];
        atomicAdd(&bucket[1023u - (unsigned)(((u64)c * 1023ull) / maxc)], 1u);
    }
    __syncthreads();
    {
        const unsigned c = bucket[tid];
        unsigned x = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) warp_sums[wid] = x;
        __syncthreads();
        if (wid == 0) {
            unsigned w = warp_sums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
            warp_sums[lane] = w;
        }
        __syncthreads();
        bucket[tid] = x - c + (wid > 0 ? warp_sums[wid - 1] : 0u);
    }
    __syncthreads();
    for (int i = tid; i < n; i += 1024) {
        const unsigned c = tc[(size_t)i * FB_TC_STRIDE];
        const unsigned pos = atomicAdd(&bucket[1023u - (unsigned)(((u64)c * 1023ull) / maxc)], 1u);
        bins.tile_order[order_index(pos, (unsigned)v, (unsigned)V, (unsigned)n)] = (unsigned)(v * n + i);
        // a heavy list: ceil(c / FB_SEG_KEYS) segments, their state / output rows reserved consecutively
        if (bins.seg_list && pos < FB_HEAVY_RANKS && c >= FB_HEAVY_KEYS) {
            const unsigned S = (c + FB_SEG_KEYS - 1) / FB_SEG_KEYS;
            const unsigned base = atomicAdd(&bins.total[10], S);           // items (state / output rows) of the chunk
            // (the rows are sized for `capacity` instances; a chunk that needs more raises the overflow flag below and
            //  none of its later kernels runs, so the list is simply not written)
            if ((size_t)base + S > seg_item_capacity(bins.capacity, (long long)V * FB_HEAVY_RANKS)) continue;
            const unsigned l = atomicAdd(&bins.total[12], 1u);             // lists
            atomicMax(&bins.total[13], S);
            bins.seg_list[l] = make_uint4((unsigned)(v * n + i), S, base, 0u);
            for (unsigned k = 0; k < S; ++k) {
                bins.seg_flags[2 * (base + k)] = make_uint4(0u, 0u, 0u, 0u);
                bins.seg_flags[2 * (base + k) + 1] = make_uint4(0u, 0u, 0u, 0u);
            }
        }
    }
}

// ---- stage 2b: scatter keys into tile segments ----------------------------------------------
// One thread per visible Gaussian of a view, reading the compact 16-byte entries stage 1 left
// (coalesced).  tile_count doubles as the cursor (counted down).
__global__ void __launch_bounds__(256) k_scatter_keys(Bins bins, Cam cam) {
    const int v = blockIdx.y;
    uint32_t* tc = bins.tile_count + (size_t)v * bins.tiles * FB_TC_STRIDE;
    if (bins.total[1]) {
        // capacity exceeded: the caller re-runs with a larger buffer.  Leave the cursors clean.
        if (blockIdx.x == 0)
            for (int i = threadIdx.x; i < bins.tiles; i += blockDim.x) tc[i * FB_TC_STRIDE] = 0u;
        return;
    }
    const unsigned nvis = bins.vis_count[v];
    const uint2* tr = bins.tile_range + (size_t)v * bins.tiles;
    // warp-uniform trip count: every lane of a warp stays in the loop while any lane has an entry
    for (unsigned i0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; i0 < nvis; i0 += gridDim.x * blockDim.x) {
        const unsigned i = i0 + (threadIdx.x & 31);
        uint4 e = make_uint4(0u, 0u, 0u, 0u);
        if (i < nvis) e = bins.vis_list[(size_t)v * bins.vis_stride + i];
        const int x0 = e.z & 0xff, y0 = (e.z >> 8) & 0xff, x1 = (e.z >> 16) & 0xff, y1 = e.z >> 24;
        // low word: the Gaussian id, or (scoring path) the visible slot its records live in
        const u64 key = ((u64)e.y << 32) | (u64)(bins.key_is_slot ? i : e.x);
        warp_tiles<true>(x0, y0, x1, y1, cam.tiles_x, key,
                   [&](int t, unsigned n) { return atomicSub(&tc[t * FB_TC_STRIDE], n); },
                   [&](int t, unsigned old, unsigned rank, u64 k) { bins.keys[tr[t].x + old - 1u - rank] = k; });
    }
}

// CTA-aggregated form of the scatter (tiles <= FB_SC_MAX_TILES): a CTA takes FB_SC_EPT * 256 consecutive entries of a
// view's visible list, counts their instances per tile with SHARED-memory atomics, claims each touched tile's room
// with ONE global atomicSub per (CTA, tile) -- the warp-aggregated form above issues one per (warp, tile), ~13x as many
// returning L2 atomics, and is bound by their latency -- and writes the keys at ranks handed out again in shared memory.
// The order inside a segment is arbitrary either way; k_sort_tiles establishes it.
#define FB_SC_MAX_TILES 1024
#ifndef FB_SC_EPT
#define FB_SC_EPT 1
#endif
__global__ void __launch_bounds__(256) k_scatter_keys_cta(Bins bins, Cam cam) {
    __shared__ unsigned cnt[FB_SC_MAX_TILES];
    __shared__ unsigned base[FB_SC_MAX_TILES];
    const int v = blockIdx.y, tid = threadIdx.x, tiles = bins.tiles;
    uint32_t* tc = bins.tile_count + (size_t)v * tiles * FB_TC_STRIDE;
    if (bins.total[1]) {
        if (blockIdx.x == 0)
            for (int i = tid; i < tiles; i += blockDim.x) tc[i * FB_TC_STRIDE] = 0u;
        return;
    }
    const unsigned nvis = bins.vis_count[v];
    const uint2* tr = bins.tile_range + (size_t)v * tiles;
    const unsigned per_cta = 256u * FB_SC_EPT;
    for (unsigned start = blockIdx.x * per_cta; start < nvis; start += gridDim.x * per_cta) {
        for (int t = tid; t < tiles; t += 256) cnt[t] = 0u;
        __syncthreads();
        uint4 e[FB_SC_EPT];
#pragma unroll
        for (int k = 0; k < FB_SC_EPT; ++k) {
            const unsigned i = start + k * 256u + tid;
            e[k] = make_uint4(0u, 0u, 0u, 0u);
            if (i < nvis) e[k] = bins.vis_list[(size_t)v * bins.vis_stride + i];
            const int x0 = e[k].z & 0xff, y0 = (e[k].z >> 8) & 0xff, x1 = (e[k].z >> 16) & 0xff, y1 = e[k].z >> 24;
            for (int y = y0; y < y1; ++y)
                for (int x = x0; x < x1; ++x) atomicAdd(&cnt[y * cam.tiles_x + x], 1u);
        }
        __syncthreads();
        for (int t = tid; t < tiles; t += 256) {
            const unsigned c = cnt[t];
            if (c) base[t] = tr[t].x + atomicSub(&tc[t * FB_TC_STRIDE], c) - 1u;      // position of this CTA's rank 0
            cnt[t] = 0u;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < FB_SC_EPT; ++k) {
            const unsigned i = start + k * 256u + tid;
            const int x0 = e[k].z & 0xff, y0 = (e[k].z >> 8) & 0xff, x1 = (e[k].z >> 16) & 0xff, y1 = e[k].z >> 24;
            const u64 key = ((u64)e[k].y << 32) | (u64)(bins.key_is_slot ? i : e[k].x);
            for (int y = y0; y < y1; ++y)
                for (int x = x0; x < x1; ++x) {
                    const int t = y * cam.tiles_x + x;
                    bins.keys[base[t] - atomicAdd(&cnt[t], 1u)] = key;
                }
        }
        __syncthreads();
    }
}

// ---- stage 2c: in-CTA sort of one tile segment ----------------------------------------------
// Stable LSD radix sort of the 8-byte keys (depth_bits<<32 | gaussian_id) on the 32 DEPTH bits only
// (four 8-bit passes; a pass whose digit is the same for every key is skipped), followed by a
// fix-up that orders the rare runs of bit-identical depths by Gaussian id.  The result is the total
// order by (depth, id), i.e. exactly the reference's stable (tile|depth) radix sort of keys emitted
// in ascending Gaussian index (rasterizer_impl.cu:92-108, :304-309).
// Each warp owns a contiguous chunk of the segment, so ranking keys with __match_any_sync inside
// the warp and per-(warp, digit) counters across warps keeps every pass stable without atomics.
// src/dst may be shared or global memory (tiles longer than the shared capacity ping-pong in L2).
#define FB_RADIX_WARPS (FB_TILE_PIX / 32)

__device__ __forceinline__ bool radix_pass_u64(const u64* __restrict__ src, u64* __restrict__ dst, int n, int shift,
                                               unsigned (*hist)[256]) {
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const unsigned lt = (1u << lane) - 1u;
    for (int i = tid; i < FB_RADIX_WARPS * 256; i += FB_TILE_PIX) (&hist[0][0])[i] = 0u;
    __syncthreads();
    const int chunk = (((n + FB_RADIX_WARPS - 1) / FB_RADIX_WARPS) + 31) & ~31;
    const int beg = min(n, w * chunk), end = min(n, beg + chunk);
    for (int i = beg; i < end; i += 32) {
        const int k = i + lane;
        const unsigned d = (k < end) ? (unsigned)((src[k] >> shift) & 255ull) : (256u + lane);
        const unsigned m = __match_any_sync(0xffffffffu, d);
        if (k < end && (m & lt) == 0u) hist[w][d] += __popc(m);
    }
    __syncthreads();
    // exclusive scan over (digit major, warp minor); thread d owns digit d
    __shared__ unsigned wsum[FB_RADIX_WARPS];
    __shared__ int same_digit;
    unsigned tot = 0;
    {
        unsigned run = 0;
#pragma unroll
        for (int ww = 0; ww < FB_RADIX_WARPS; ++ww) { unsigned c = hist[ww][tid]; hist[ww][tid] = run; run += c; }
        tot = run;
    }
    if (tid == 0) same_digit = 0;
    __syncthreads();
    if (tot == (unsigned)n) same_digit = 1;
    unsigned x = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) wsum[w] = x;
    __syncthreads();
    if (same_digit) return false;          // every key has the same digit: the pass is the identity
    unsigned base = x - tot;
#pragma unroll
    for (int ww = 0; ww < FB_RADIX_WARPS; ++ww) if (ww < w) base += wsum[ww];
#pragma unroll
    for (int ww = 0; ww < FB_RADIX_WARPS; ++ww) hist[ww][tid] += base;
    __syncthreads();
    for (int i = beg; i < end; i += 32) {
        const int k = i + lane;
        u64 key = 0;
        unsigned d = 256u + lane;
        if (k < end) { key = src[k]; d = (unsigned)((key >> shift) & 255ull); }
        const unsigned m = __match_any_sync(0xffffffffu, d);
        unsigned off = 0;
        if (k < end) off = hist[w][d];
        __syncwarp();
        if (k < end && (m & lt) == 0u) hist[w][d] = off + __popc(m);
        __syncwarp();
        if (k < end) dst[off + __popc(m & lt)] = key;
    }
    __syncthreads();
    return true;
}

// Total order of the keys: depth bits, then GAUSSIAN id (the reference's stable sort keeps emission order = ascending
// Gaussian index among equal depths, rasterizer_impl.cu:92-108).  On the scoring path the low word of a key is the
// visible slot (assigned by atomics, in no particular order), so equal depths -- rare -- look the id up in the view's
// visible list; elsewhere the low word is the id itself and this is the plain 64-bit compare.
__device__ __forceinline__ bool key_less(u64 a, u64 b, const uint4* __restrict__ vis) {
    const unsigned da = (unsigned)(a >> 32), db = (unsigned)(b >> 32);
    if (da != db) return da < db;
    if (vis == nullptr) return (unsigned)a < (unsigned)b;
    return vis[(unsigned)a].x < vis[(unsigned)b].x;
}

// Orders runs of equal depth bits by id (key_less).  Runs are disjoint, one thread each.
__device__ __forceinline__ void fix_equal_depth_runs(u64* a, int n, const uint4* __restrict__ vis) {
    for (int k = threadIdx.x; k + 1 < n; k += blockDim.x) {
        const unsigned d = (unsigned)(a[k] >> 32);
        if ((unsigned)(a[k + 1] >> 32) != d) continue;
        if (k > 0 && (unsigned)(a[k - 1] >> 32) == d) continue;   // not the start of the run
        int e = k + 1;
        while (e < n && (unsigned)(a[e] >> 32) == d) ++e;
        for (int i = k + 1; i < e; ++i) {                          // insertion sort (runs are short)
            const u64 x = a[i];
            int j = i - 1;
            while (j >= k && key_less(x, a[j], vis)) { a[j + 1] = a[j]; --j; }
            a[j + 1] = x;
        }
    }
    __syncthreads();
}

// Sort every tile segment of every view of the chunk.  Dynamic shared memory: smem_keys keys (x2 if FB_SORT_STAGE).
//
// The required order is the total order of the 64-bit key (depth_bits << 32 | gaussian_id), so ANY complete
// sort of the full key reproduces the reference's stable radix sort; stability is not needed.  Segments that
// fit in shared memory use a histogram-equalised bucket sort (5 light passes, no per-digit passes):
//   A  stage keys, min/max of the depth bits                 B  256 coarse buckets (linear in the depth bits)
//   C  each coarse bucket gets fine buckets in proportion to its population (<= 2048 in total), so the
//      monotone piecewise-linear map depth -> fine bucket leaves ~1 key per bucket even when the tile's
//      depths cluster on a few surfaces; fine histogram + scan
//   D  scatter into bucket order (arrival order inside a bucket)
//   E  every key counts the smaller keys of its own bucket and is written to its final place.
// Segments longer than the shared capacity, or with a pathologically full bucket, take the LSD radix path.
#ifndef FB_SORT_STAGE
#define FB_SORT_STAGE 1
#endif
#define FB_SORT_COARSE 256
#ifndef FB_SORT_FINE
#define FB_SORT_FINE 2048
#endif
#ifndef FB_SORT_MAX_BUCKET
#define FB_SORT_MAX_BUCKET 64
#endif

__device__ __forceinline__ unsigned block_excl_scan_256(unsigned x, unsigned* warp_sums, unsigned* total) {
    // exclusive scan of one value per thread over a 256-thread block (all threads call; contains barriers)
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    unsigned inc = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
    __syncthreads();                       // warp_sums may still be read from a previous call
    if (lane == 31) warp_sums[w] = inc;
    __syncthreads();
    unsigned base = 0, tot = 0;
#pragma unroll
    for (int ww = 0; ww < FB_RADIX_WARPS; ++ww) { const unsigned v = warp_sums[ww]; if (ww < w) base += v; tot += v; }
    if (total) *total = tot;
    return base + inc - x;
}

struct SortShared {
    unsigned work[FB_SORT_FINE > FB_RADIX_WARPS * 256 ? FB_SORT_FINE : FB_RADIX_WARPS * 256];   // fine-bucket cursors, or the radix histograms
    unsigned c_cnt[FB_SORT_COARSE], c_base[FB_SORT_COARSE], c_nf[FB_SORT_COARSE];
    unsigned wsums[FB_RADIX_WARPS];
    unsigned s_lo, s_hi, s_maxb;
};

// Sorts the n keys of one segment into `seg`.  `alt`: n keys of scratch in global memory; with in_alt the unsorted keys
// are there (k_sort_split left them in bucket order), else in seg.  All threads of the 256-thread CTA call; the caller
// puts a barrier between two calls.
__device__ __forceinline__ void sort_segment(u64* seg, u64* alt, const int n, const bool in_alt,
                                             const uint4* __restrict__ vis, const int smem_keys, u64* s_keys, SortShared& sh) {
    unsigned* work = sh.work;
    unsigned* c_cnt = sh.c_cnt; unsigned* c_base = sh.c_base; unsigned* c_nf = sh.c_nf;
    unsigned* wsums = sh.wsums;
    unsigned& s_lo = sh.s_lo; unsigned& s_hi = sh.s_hi; unsigned& s_maxb = sh.s_maxb;
    const int tid = threadIdx.x;
    if (n < 2) { if (in_alt && n == 1 && tid == 0) seg[0] = alt[0]; return; }
    if (in_alt && n > smem_keys) {                            // (a pathological sub-segment: rare)
        for (int i = tid; i < n; i += FB_TILE_PIX) seg[i] = alt[i];
        __syncthreads();
    }
    const u64* src = (in_alt && n <= smem_keys) ? alt : seg;
    bool radix = false;
    {
        // Segments that fit: keys staged and scattered in shared memory, 2048 fine buckets in `work`.
        // Longer segments (a view close to a wall puts > 10k instances on a tile): same algorithm with the keys
        // in global memory (seg -> alt -> seg, L2-resident) and the dynamic shared memory used as the
        // bucket array instead (16 KB per 4096 entries), so the buckets stay short.
        const bool big = n > smem_keys;
        const u64* s_in = big ? seg : s_keys;
        u64* s_out = big ? alt : s_keys + smem_keys;
        unsigned* fine = big ? reinterpret_cast<unsigned*>(s_keys) : work;
#define FB_FIDX(f) (f)
        const int NF = big ? min(smem_keys * 4, 32768) & ~255 : FB_SORT_FINE;   // u32 entries available; multiple of 256
        // A
        if (tid == 0) { s_lo = 0xffffffffu; s_hi = 0u; s_maxb = 0u; }
        c_cnt[tid] = 0u;
        __syncthreads();
        unsigned lo = 0xffffffffu, hi = 0u;
#pragma unroll 4
        for (int i = tid; i < n; i += FB_TILE_PIX) {
            const u64 k = src[i];
            if (!big) s_keys[i] = k;
            const unsigned d = (unsigned)(k >> 32);
            lo = min(lo, d); hi = max(hi, d);
        }
        lo = __reduce_min_sync(0xffffffffu, lo); hi = __reduce_max_sync(0xffffffffu, hi);
        if ((tid & 31) == 0) { atomicMin(&s_lo, lo); atomicMax(&s_hi, hi); }
        __syncthreads();
        lo = s_lo;
        const unsigned range = s_hi - lo;
        const int nb = 32 - __clz(range | 1u);                 // bits needed for (d - lo)
        const int shc = nb > 8 ? nb - 8 : 0;                   // coarse bucket = (d - lo) >> shc  (< 256)
        const unsigned wmask = (1u << shc) - 1u;
        // B
#pragma unroll 4
        for (int i = tid; i < n; i += FB_TILE_PIX)
            atomicAdd(&c_cnt[((unsigned)(s_in[i] >> 32) - lo) >> shc], 1u);
        __syncthreads();
        // C: fine buckets per coarse bucket, proportional to its population
        {
            const unsigned cnt = c_cnt[tid];
            const unsigned nf = (unsigned)(((u64)cnt * (unsigned)(NF - FB_SORT_COARSE) + (unsigned)n - 1u) / (unsigned)n);
            c_nf[tid] = nf;
            c_base[tid] = block_excl_scan_256(nf, wsums, nullptr);
        }
        for (int i = tid; i < NF; i += FB_TILE_PIX) fine[i] = 0u;
        __syncthreads();
#define FB_FINE_BUCKET(d) ({ const unsigned x_ = (d) - lo; const unsigned c_ = x_ >> shc; \
                             c_base[c_] + (unsigned)(((u64)(x_ & wmask) * c_nf[c_]) >> shc); })
#pragma unroll 4
        for (int i = tid; i < n; i += FB_TILE_PIX) atomicAdd(&fine[FB_FIDX(FB_FINE_BUCKET((unsigned)(s_in[i] >> 32)))], 1u);
        __syncthreads();
        {
            // exclusive scan of the fine counts.  Warp w owns the contiguous range [w, w+1) * NF/8 and walks it 32
            // buckets at a time (lane-contiguous: no bank conflicts, whatever NF is): totals first, then the scan.
            const int lane = tid & 31, w = tid >> 5;
            const int wlen = NF / FB_RADIX_WARPS, wbeg = w * wlen;
            unsigned sum = 0, mx = 0;
            for (int j = lane; j < wlen; j += 32) { const unsigned c = fine[wbeg + j]; sum += c; mx = max(mx, c); }
            sum = __reduce_add_sync(0xffffffffu, sum);
            mx = __reduce_max_sync(0xffffffffu, mx);
            if (lane == 0) { wsums[w] = sum; atomicMax(&s_maxb, mx); }
            __syncthreads();
            unsigned carry = 0;
#pragma unroll
            for (int ww = 0; ww < FB_RADIX_WARPS; ++ww) if (ww < w) carry += wsums[ww];
            for (int j = lane; j < wlen; j += 32) {
                const unsigned c = fine[wbeg + j];
                unsigned inc = c;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
                fine[wbeg + j] = carry + inc - c;
                carry += __shfl_sync(0xffffffffu, inc, 31);
            }
        }
        __syncthreads();
        if (s_maxb > FB_SORT_MAX_BUCKET) {
            radix = true;                                      // uniform: s_maxb is block-wide
        } else {
            // D
    #pragma unroll 4
        for (int i = tid; i < n; i += FB_TILE_PIX) {
                const u64 k = s_in[i];
                const unsigned pos = atomicAdd(&fine[FB_FIDX(FB_FINE_BUCKET((unsigned)(k >> 32)))], 1u);
                s_out[pos] = k;
            }
            __syncthreads();                                   // (global writes of this block are visible to it)
            // E: fine[f] is now the END of bucket f, i.e. the start of bucket f + 1
    #pragma unroll 4
        for (int i = tid; i < n; i += FB_TILE_PIX) {
                const u64 k = s_out[i];
                const unsigned f = FB_FINE_BUCKET((unsigned)(k >> 32));
                const unsigned e = fine[FB_FIDX(f)], b0 = f ? fine[FB_FIDX(f - 1)] : 0u;
                // rank inside the bucket by depth; keys of EQUAL depth (rare) are ranked among themselves by id
                const unsigned dk = (unsigned)(k >> 32);
                unsigned rank = 0, ties = 0;
                for (unsigned j = b0; j < e; ++j) {
                    const unsigned dj = (unsigned)(s_out[j] >> 32);
                    rank += (dj < dk) ? 1u : 0u;
                    ties += (dj == dk) ? 1u : 0u;
                }
                if (ties > 1u)
                    for (unsigned j = b0; j < e; ++j) {
                        const u64 o = s_out[j];
                        if ((unsigned)(o >> 32) == dk && o != k && key_less(o, k, vis)) ++rank;
                    }
                seg[b0 + rank] = k;
            }
            return;
        }
#undef FB_FINE_BUCKET
        __syncthreads();
    }
    // LSD radix on the depth bits through global memory (L2-resident ping-pong), then the equal-depth fix-up
    {
        unsigned (*hist)[256] = reinterpret_cast<unsigned (*)[256]>(work);
        u64* a = seg;
        u64* b = alt;
        if (in_alt && n <= smem_keys) { a = alt; b = seg; }     // a piece that was not copied: its keys are still in alt
#pragma unroll 1
        for (int pass = 0; pass < 4; ++pass) {
            if (radix_pass_u64(a, b, n, 32 + 8 * pass, hist)) { u64* tmp = a; a = b; b = tmp; }
        }
        fix_equal_depth_runs(a, n, vis);
        if (a != seg) {
            for (int i = tid; i < n; i += blockDim.x) seg[i] = a[i];
        }
    }
}

// ---- long segments: split by depth first, sort the pieces with every SM ------------------------------------
// One CTA per tile sorts a segment, and a segment longer than the shared staging goes through global memory at a
// fraction of the speed: a view along a 30 m floor has tiles with > 100 k keys (cfg 4: the longest 272 k) and the
// kernel waited milliseconds for the CTAs that own them.  A segment of >= FB_SPLIT_MIN keys (among the first
// `split_slots` entries of the tile order) is instead PARTITIONED by its CTA into depth buckets
// (keys -> keys_alt in bucket order), and the bucket sequence is cut into pieces of
// ~FB_SPLIT_T keys.  A piece never cuts a bucket, so equal depths stay together and the sorted pieces, concatenated,
// are the sorted segment.  The pieces go onto one list for the chunk (bins.sub_list, count in bins.total[8]); the LAST
// `sub_ctas` CTAs of the grid wait until the first split_slots CTAs have reported (bins.total[9]) and sort the pieces
// (keys_alt -> keys) in shared memory.  CTAs start in index order, so every CTA a waiting one depends on is resident
// or done (the argument of the chained scan).
#ifndef FB_SPLIT_RANKS
#define FB_SPLIT_RANKS 64
#endif
#define FB_SPLIT_RANKS_ FB_SPLIT_RANKS
#ifndef FB_SPLIT_MIN
#define FB_SPLIT_MIN 16384
#endif
#ifndef FB_SPLIT_T
#define FB_SPLIT_T 768
#endif
#define FB_SPLIT_FINE 4096        // fine buckets of the partition: the dynamic shared memory as unsigned[4096]
#define FB_SPLIT_MAXJ 2048        // piece table: SortShared::work
__host__ __device__ __forceinline__ bool sort_is_split(int n) { return n >= FB_SPLIT_MIN && n <= FB_SPLIT_MAXJ * FB_SPLIT_T; }
// entries the piece list needs for `instances` keys over at most `slots` split segments (a hard bound)
__host__ __device__ __forceinline__ size_t sort_sub_capacity(long long instances, long long slots) {
    return (size_t)(instances / FB_SPLIT_T + slots + 1);
}

// The partition uses a FIXED monotone map depth -> bucket (no min / max and no equalisation passes: two reads of the
// segment instead of four): 512 buckets per octave of depth over [0.125, 32) m, i.e. bits 14.. of the float above
// 0.125f.  A tile whose depths cluster in a few buckets yields pieces longer than the staging; sort_segment takes
// those through its global-memory path, as it would have taken the whole segment.
__device__ __forceinline__ unsigned split_bucket(unsigned depth_bits) {
    const unsigned x = depth_bits > 0x3E000000u ? depth_bits - 0x3E000000u : 0u;
    return min(x >> 14, (unsigned)(FB_SPLIT_FINE - 1));
}
__device__ __forceinline__ void split_segment(const Bins& bins, const u64* __restrict__ seg, u64* __restrict__ alt, const int n,
                                              const unsigned off, const unsigned view, unsigned* fine, SortShared& sh) {
    unsigned* tab = sh.work;
    const int tid = threadIdx.x, lane = tid & 31;
    constexpr int NF = FB_SPLIT_FINE, NT = FB_TILE_PIX, U = 8;
    for (int i = tid; i < NF; i += NT) fine[i] = 0u;
    const int J = (n + FB_SPLIT_T - 1) / FB_SPLIT_T;
    for (int i = tid; i < J; i += NT) tab[i] = 0xffffffffu;
    __syncthreads();
    // histogram: U loads in flight per thread
    for (int i0 = 0; i0 < n; i0 += NT * U) {
        unsigned d[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { const int i = i0 + u * NT + tid; d[u] = i < n ? (unsigned)(__ldcg(seg + i) >> 32) : 0u; }
#pragma unroll
        for (int u = 0; u < U; ++u) if (i0 + u * NT + tid < n) atomicAdd(&fine[split_bucket(d[u])], 1u);
    }
    __syncthreads();
    {
        // exclusive scan of the counts: thread t owns buckets [16 t, 16 t + 16)
        constexpr int PER = NF / NT;
        unsigned sum = 0;
#pragma unroll
        for (int k = 0; k < PER; ++k) sum += fine[tid * PER + k];
        unsigned start = block_excl_scan_256(sum, sh.wsums, nullptr);
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            const unsigned c = fine[tid * PER + k];
            fine[tid * PER + k] = start;
            // piece j = the buckets whose first key position lies in [j T, (j + 1) T): its start is the smallest of them
            if (c) atomicMin(&tab[start / FB_SPLIT_T], start);
            start += c;
        }
    }
    __syncthreads();
    for (int i0 = 0; i0 < n; i0 += NT * U) {
        u64 k[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { const int i = i0 + u * NT + tid; k[u] = i < n ? __ldcg(seg + i) : 0ull; }
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (i0 + u * NT + tid < n) alt[atomicAdd(&fine[split_bucket((unsigned)(k[u] >> 32))], 1u)] = k[u];
    }
    // the pieces: one list entry each {first key, keys, view, -}; a warp reserves its entries with one atomic
    for (int j0 = 0; j0 < J; j0 += NT) {
        const int j = j0 + tid;
        const unsigned st = j < J ? tab[j] : 0xffffffffu;
        const bool has = st != 0xffffffffu;
        unsigned en = (unsigned)n;
        if (has) for (int k = j + 1; k < J; ++k) { const unsigned t = tab[k]; if (t != 0xffffffffu) { en = t; break; } }
        const unsigned m = __ballot_sync(0xffffffffu, has);
        unsigned base = 0;
        if (lane == 0 && m) base = atomicAdd(&bins.total[8], (unsigned)__popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (has) bins.sub_list[base + __popc(m & ((1u << lane) - 1u))] = make_uint4(off + st, en - st, view, 0u);
    }
}

// Sort every tile segment of every view of the chunk.  Dynamic shared memory: smem_keys keys (x2 if FB_SORT_STAGE).
// Grid: [one CTA per (view, tile), heaviest first; the first split_slots may partition instead][sub_ctas piece sorters].
#ifndef FB_SORT_MINB
#define FB_SORT_MINB 6
#endif
__global__ void __launch_bounds__(FB_TILE_PIX, FB_SORT_MINB) k_sort_tiles(Bins bins, u64* __restrict__ keys_alt, int smem_keys, int tile_ctas,
                                                               int split_slots) {
    extern __shared__ u64 s_keys[];
    __shared__ SortShared sh;
    if (bins.total[1]) return;
    const int tid = threadIdx.x;
    if ((int)blockIdx.x >= tile_ctas) {
        // piece sorter: wait for the partitions, then walk the list
        if (tid == 0) {
            volatile unsigned* done = bins.total + 9;
            while (*done < (unsigned)split_slots) __nanosleep(200);
            __threadfence();
        }
        __syncthreads();
        const unsigned cnt = *(volatile unsigned*)(bins.total + 8);
        const unsigned nsub = gridDim.x - (unsigned)tile_ctas;
        for (unsigned i = blockIdx.x - (unsigned)tile_ctas; i < cnt; i += nsub) {
            const uint4 d = __ldcg(bins.sub_list + i);
            const uint4* vis = bins.key_is_slot ? bins.vis_list + (size_t)d.z * bins.vis_stride : nullptr;
            sort_segment(bins.keys + d.x, keys_alt + d.x, (int)d.y, true, vis, smem_keys, s_keys, sh);
            __syncthreads();
        }
        return;
    }
    const unsigned lin = blockIdx.x;
    const unsigned slot = bins.tile_order ? bins.tile_order[lin] : lin;    // heaviest segments first
    const uint2 r = bins.tile_range[slot];
    const int n = (int)(r.y - r.x);
    const bool reports = split_slots > 0 && order_rank(lin, (unsigned)bins.V, (unsigned)bins.tiles) < FB_SPLIT_RANKS_;
    if (reports && sort_is_split(n)) {
        split_segment(bins, bins.keys + r.x, keys_alt + r.x, n, r.x, slot / (unsigned)bins.tiles, reinterpret_cast<unsigned*>(s_keys), sh);
        __threadfence();
        __syncthreads();
        if (tid == 0) atomicAdd(&bins.total[9], 1u);
        return;
    }
    if (reports && tid == 0) atomicAdd(&bins.total[9], 1u);
    if (n < 2) return;
    const uint4* vis = bins.key_is_slot ? bins.vis_list + (size_t)(slot / (unsigned)bins.tiles) * bins.vis_stride : nullptr;
    sort_segment(bins.keys + r.x, keys_alt + r.x, n, false, vis, smem_keys, s_keys, sh);
}

}  // namespace fb
