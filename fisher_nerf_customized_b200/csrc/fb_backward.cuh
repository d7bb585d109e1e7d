// fb_backward.cuh -- drop-in backward: per-Gaussian moment rows (k_bwd_warp, fb_render_warp.cuh) -> expansion into the
// reference's nine gradient tensors; the per-pair fallback for other powers / SH colours (RasterizeGaussiansBackwardCUDA, DGR/rasterize_points.cu:117-196).
#pragma once
#include "fb_fisher.cuh"

namespace fb {

// Outputs in the reference's shapes; any pointer may be null (skipped).
struct GradOut {
    float* dmeans2D;   // [P,3]
    float* dconic;     // [P,2,2]  (x=[0][0], y=[0][1], w=[1][1])
    float* dopacity;   // [P]
    float* dcolors;    // [P,3]
    float* dmeans3D;   // [P,3]
    float* dcov3D;     // [P,6]
    float* dscales;    // [P,3]
    float* drots;      // [P,4]
    float* dextra;     // [P,3]  fused second render only: gradient of the extra colour triple
};

// One thread per Gaussian.  Outputs are caller-zeroed, so invisible Gaussians are skipped.
// MULTI (fused colour + extra render, power 1): dmeans2D is the colour render's own (moment slots 6-7), everything
// else is the sum of the two renders, dextra comes from slots 8-10 (see k_bwd_warp).
template <bool P2, bool MULTI = false>
__global__ void __launch_bounds__(128) k_expand_grads(int P, GaussIn g, Cam cam, const GRec* __restrict__ recs,
                                                     const int* __restrict__ radii, const float* __restrict__ mom,
                                                     GradOut out) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P || !(radii[idx] > 0)) return;
    float acc[FB_NACC];
#pragma unroll
    for (int k = 0; k < 4; ++k) *reinterpret_cast<float4*>(acc + 4 * k) = *reinterpret_cast<const float4*>(mom + (size_t)idx * FB_NACC + 4 * k);
    const GRec r = recs[idx];
    const float A = r.b.x, B = r.b.y, C = r.b.z, op = r.a.w;
    const float halfW = 0.5f * cam.W, halfH = 0.5f * cam.H;
    const float3 mean = make_float3(g.means3D[3 * idx], g.means3D[3 * idx + 1], g.means3D[3 * idx + 2]);
    float cov6[6];
    const bool has_sr = (g.cov3D == nullptr);
    SRCtx sr;
    if (has_sr) {
        float3 sc = make_float3(g.scales[3 * idx], g.scales[3 * idx + 1], g.scales[3 * idx + 2]);
        float4 q = *reinterpret_cast<const float4*>(g.rotations + 4 * (size_t)idx);
        sr = make_sr(sc, cam.scale_modifier, q);
        cov3d_from_scale_rot(sc, cam.scale_modifier, q, cov6);
    } else {
#pragma unroll
        for (int i = 0; i < 6; ++i) cov6[i] = g.cov3D[6 * (size_t)idx + i];
    }
    const GCtx ctx = make_ctx(mean, cov6, cam);
    // result rows: 0-5 cov3D, 6-8 means3D, 9-11 scales, 12-15 rotations
    float res[16];
    float m2d[2], con[3], dop;
    if (!P2) {
        float w[5] = {acc[1], acc[2], acc[3], acc[4], acc[5]};
        float u[5];
        phi_to_u(w, A, B, C, halfW, halfH, u);
        ChainOut co = chain_geom(ctx, cam, u[0], u[1], u[2], u[3], u[4]);
#pragma unroll
        for (int i = 0; i < 6; ++i) res[i] = co.cov3D[i];
#pragma unroll
        for (int i = 0; i < 3; ++i) res[6 + i] = co.mean[i];
        if (has_sr) chain_scale_rot(sr, co.cov3D, res + 9, res + 12);
        con[0] = u[0]; con[1] = u[1]; con[2] = u[2];
        m2d[0] = u[3]; m2d[1] = u[4];
        if (MULTI) {
            m2d[0] = -(A * acc[6] + B * acc[7]) * halfW;
            m2d[1] = -(C * acc[7] + B * acc[6]) * halfH;
        }
        dop = acc[0] / op;
    } else {
        float col[5][16];
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            float w[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
            w[i] = 1.f;
            float u[5];
            phi_to_u(w, A, B, C, halfW, halfH, u);
            ChainOut co = chain_geom(ctx, cam, u[0], u[1], u[2], u[3], u[4]);
#pragma unroll
            for (int k = 0; k < 6; ++k) col[i][k] = co.cov3D[k];
#pragma unroll
            for (int k = 0; k < 3; ++k) col[i][6 + k] = co.mean[k];
            if (has_sr) chain_scale_rot(sr, co.cov3D, &col[i][9], &col[i][12]);
            else {
#pragma unroll
                for (int k = 9; k < 16; ++k) col[i][k] = 0.f;
            }
        }
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            double sum = 0.0;
#pragma unroll
            for (int i = 0; i < 5; ++i)
#pragma unroll
                for (int j = 0; j < 5; ++j) sum += (double)col[i][k] * (double)col[j][k] * (double)moment_phi(acc, i, j);
            res[k] = (float)sum;
        }
        con[0] = 0.25f * acc[8]; con[1] = 0.25f * acc[10]; con[2] = 0.25f * acc[12];
        m2d[0] = halfW * halfW * (A * A * acc[1] + 2.f * A * B * acc[2] + B * B * acc[3]);
        m2d[1] = halfH * halfH * (C * C * acc[3] + 2.f * B * C * acc[2] + B * B * acc[1]);
        dop = acc[0] / (op * op);
    }
    if (out.dmeans2D) { out.dmeans2D[3 * idx] = m2d[0]; out.dmeans2D[3 * idx + 1] = m2d[1]; }
    if (out.dconic) { out.dconic[4 * idx] = con[0]; out.dconic[4 * idx + 1] = con[1]; out.dconic[4 * idx + 3] = con[2]; }
    if (out.dopacity) out.dopacity[idx] = dop;
    if (out.dcolors) { out.dcolors[3 * idx] = acc[13]; out.dcolors[3 * idx + 1] = acc[14]; out.dcolors[3 * idx + 2] = acc[15]; }
    if (MULTI && out.dextra) { out.dextra[3 * idx] = acc[8]; out.dextra[3 * idx + 1] = acc[9]; out.dextra[3 * idx + 2] = acc[10]; }
    if (out.dmeans3D) { out.dmeans3D[3 * idx] = res[6]; out.dmeans3D[3 * idx + 1] = res[7]; out.dmeans3D[3 * idx + 2] = res[8]; }
    if (out.dcov3D) {
#pragma unroll
        for (int i = 0; i < 6; ++i) out.dcov3D[6 * (size_t)idx + i] = res[i];
    }
    if (has_sr) {
        if (out.dscales) { out.dscales[3 * idx] = res[9]; out.dscales[3 * idx + 1] = res[10]; out.dscales[3 * idx + 2] = res[11]; }
        if (out.drots) {
#pragma unroll
            for (int i = 0; i < 4; ++i) out.drots[4 * (size_t)idx + i] = res[12 + i];
        }
    }
}

// ---- per-pair fallback: any integer backward_power, SH degree > 0 ---------------------------------------------
// The moment formulation above needs sum_pairs f(partial) to be a polynomial in the pair's monomials: true for
// backward_power 1 and 2 with precomputed colours.  For every other case the reference's definition is evaluated as
// written -- per (pixel, Gaussian) pair every partial is formed, raised with powf(x, (float)power) and summed
// (backward.cu:1093-1137) -- but the per-pair chain rule (computeCov2DCUDARelocated / preprocessCUDARelocated /
// computeCov3D, ~300 FLOPs and 192 B of local memory per pair in the reference) is replaced by a product with
// per-Gaussian rows: every geometric partial of a pair is LINEAR in u = (dL/dconic.{x,y,w}, dL/dmean2D.{x,y}) and, with
// SH colours, in the colour partials.  k_chain_rows builds the rows once per visible Gaussian; k_bwd_pairs forms
// u per pair, multiplies, applies powf, sums the 32 pixels of the block with a halving butterfly and issues one atomic
// per (Gaussian, pixel block, output) instead of one per pair.
// SH (degree > 0), reproduced as the reference compiles it, including its quirks: the per-pair SH
// pointer is `shs + M * global_id` on a FLOAT pointer (backward.cu:1067; the view-direction part of dL/dmeans reads the
// coefficients found there), and dL_dsh is accumulated only when D > 0 (:1117).
#define FB_ROW_FLOATS 112     // 16 x 5 geometry rows | 3 x 3 SH mean rows | 16 SH basis values | 3 clamp factors | pad
struct ShArgs { const float* shs; const bool* clamped; int D, M; const float* unused; };

__global__ void __launch_bounds__(128) k_chain_rows(int P, GaussIn g, Cam cam, const GRec* __restrict__ recs,
                                                   const int* __restrict__ radii, ShArgs sh, float* __restrict__ rows_all) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P || !(radii[idx] > 0)) return;
    float* rows = rows_all + (size_t)idx * FB_ROW_FLOATS;
    const GRec r = recs[idx];
    const float A = r.b.x, B = r.b.y, C = r.b.z;
    const float halfW = 0.5f * cam.W, halfH = 0.5f * cam.H;
    const float3 mean = make_float3(g.means3D[3 * idx], g.means3D[3 * idx + 1], g.means3D[3 * idx + 2]);
    float cov6[6];
    const bool has_sr = (g.cov3D == nullptr);
    SRCtx sr;
    if (has_sr) {
        float3 sc = make_float3(g.scales[3 * idx], g.scales[3 * idx + 1], g.scales[3 * idx + 2]);
        float4 q = *reinterpret_cast<const float4*>(g.rotations + 4 * (size_t)idx);
        sr = make_sr(sc, cam.scale_modifier, q);
        cov3d_from_scale_rot(sc, cam.scale_modifier, q, cov6);
    } else {
#pragma unroll
        for (int i = 0; i < 6; ++i) cov6[i] = g.cov3D[6 * (size_t)idx + i];
    }
    const GCtx ctx = make_ctx(mean, cov6, cam);
    // column j: the response of (cov3D[6], means3D[3], scales[3], rotations[4]) to the unit vector e_j of
    // (dL/dconic.x, .y, .w, dL/dmean2D.x, .y)
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        float e[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
        e[j] = 1.f;
        ChainOut co = chain_geom(ctx, cam, e[0], e[1], e[2], e[3], e[4]);
        float ds[3] = {0.f, 0.f, 0.f}, dr[4] = {0.f, 0.f, 0.f, 0.f};
        if (has_sr) chain_scale_rot(sr, co.cov3D, ds, dr);
#pragma unroll
        for (int k = 0; k < 6; ++k) rows[k * 5 + j] = co.cov3D[k];
#pragma unroll
        for (int k = 0; k < 3; ++k) rows[(6 + k) * 5 + j] = co.mean[k];
#pragma unroll
        for (int k = 0; k < 3; ++k) rows[(9 + k) * 5 + j] = ds[k];
#pragma unroll
        for (int k = 0; k < 4; ++k) rows[(12 + k) * 5 + j] = dr[k];
    }
    (void)A; (void)B; (void)C; (void)halfW; (void)halfH;
    float* S = rows + 80;      // S[i][c]: d(means3D[i]) / d(dL_dRGB[c]) through the view direction
    float* bas = rows + 89;    // dRGB/dsh_k
    float* clm = rows + 105;   // 1 - clamped
#pragma unroll
    for (int i = 0; i < 9; ++i) S[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) bas[i] = 0.f;
    clm[0] = clm[1] = clm[2] = 1.f;
    if (sh.shs == nullptr) return;
    clm[0] = sh.clamped[4 * idx] ? 0.f : 1.f; clm[1] = sh.clamped[4 * idx + 1] ? 0.f : 1.f; clm[2] = sh.clamped[4 * idx + 2] ? 0.f : 1.f;
    // backward.cu:20-139 with the pointer of :1067
    const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
    const float C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f};
    const float C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                         -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};
    const float3 dir_orig = make_float3(mean.x - cam.campos[0], mean.y - cam.campos[1], mean.z - cam.campos[2]);
    const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
    const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
    const float* shp = sh.shs + (size_t)sh.M * idx;              // (sic) float offset M * idx, not 3 * M * idx
#define SHV(k, c) shp[3 * (k) + (c)]
    float dx[3] = {0.f, 0.f, 0.f}, dy[3] = {0.f, 0.f, 0.f}, dz[3] = {0.f, 0.f, 0.f};
    bas[0] = C0;
    const int deg = sh.D;
    if (deg > 0) {
        bas[1] = -C1 * y; bas[2] = C1 * z; bas[3] = -C1 * x;
#pragma unroll
        for (int c = 0; c < 3; ++c) { dx[c] = -C1 * SHV(3, c); dy[c] = -C1 * SHV(1, c); dz[c] = C1 * SHV(2, c); }
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            bas[4] = C2[0] * xy; bas[5] = C2[1] * yz; bas[6] = C2[2] * (2.f * zz - xx - yy); bas[7] = C2[3] * xz; bas[8] = C2[4] * (xx - yy);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                dx[c] += C2[0] * y * SHV(4, c) + C2[2] * 2.f * -x * SHV(6, c) + C2[3] * z * SHV(7, c) + C2[4] * 2.f * x * SHV(8, c);
                dy[c] += C2[0] * x * SHV(4, c) + C2[1] * z * SHV(5, c) + C2[2] * 2.f * -y * SHV(6, c) + C2[4] * 2.f * -y * SHV(8, c);
                dz[c] += C2[1] * y * SHV(5, c) + C2[2] * 2.f * 2.f * z * SHV(6, c) + C2[3] * x * SHV(7, c);
            }
            if (deg > 2) {
                bas[9] = C3[0] * y * (3.f * xx - yy); bas[10] = C3[1] * xy * z; bas[11] = C3[2] * y * (4.f * zz - xx - yy);
                bas[12] = C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy); bas[13] = C3[4] * x * (4.f * zz - xx - yy);
                bas[14] = C3[5] * z * (xx - yy); bas[15] = C3[6] * x * (xx - 3.f * yy);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    dx[c] += C3[0] * SHV(9, c) * 3.f * 2.f * xy + C3[1] * SHV(10, c) * yz + C3[2] * SHV(11, c) * -2.f * xy +
                             C3[3] * SHV(12, c) * -3.f * 2.f * xz + C3[4] * SHV(13, c) * (-3.f * xx + 4.f * zz - yy) +
                             C3[5] * SHV(14, c) * 2.f * xz + C3[6] * SHV(15, c) * 3.f * (xx - yy);
                    dy[c] += C3[0] * SHV(9, c) * 3.f * (xx - yy) + C3[1] * SHV(10, c) * xz + C3[2] * SHV(11, c) * (-3.f * yy + 4.f * zz - xx) +
                             C3[3] * SHV(12, c) * -3.f * 2.f * yz + C3[4] * SHV(13, c) * -2.f * xy + C3[5] * SHV(14, c) * -2.f * yz +
                             C3[6] * SHV(15, c) * -3.f * 2.f * xy;
                    dz[c] += C3[1] * SHV(10, c) * xy + C3[2] * SHV(11, c) * 4.f * 2.f * yz + C3[3] * SHV(12, c) * 3.f * (2.f * zz - xx - yy) +
                             C3[4] * SHV(13, c) * 4.f * 2.f * xz + C3[5] * SHV(14, c) * (xx - yy);
                }
            }
        }
    }
#undef SHV
    // dL_dmean = dnormvdv(dir_orig, (dx . dRGB, dy . dRGB, dz . dRGB))  (auxiliary.h:107-117): linear in dRGB
    const float sum2 = dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z;
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    const float vx = dir_orig.x, vy = dir_orig.y, vz = dir_orig.z;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float ddx = dx[c], ddy = dy[c], ddz = dz[c];
        S[0 * 3 + c] = ((+sum2 - vx * vx) * ddx - vy * vx * ddy - vz * vx * ddz) * invsum32;
        S[1 * 3 + c] = (-vx * vy * ddx + (sum2 - vy * vy) * ddy - vz * vy * ddz) * invsum32;
        S[2 * 3 + c] = (-vx * vz * ddx - vy * vz * ddy + (sum2 - vz * vz) * ddz) * invsum32;
    }
}

// Halving butterfly over 32 values per lane: afterwards lane L holds the warp total of v[L] (returned).
__device__ __forceinline__ float warp_reduce32(float* v, int lane) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const bool up = lane & 16;
        const float send = up ? v[i] : v[i + 16], keep = up ? v[i + 16] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const bool up = lane & 8;
        const float send = up ? v[i] : v[i + 8], keep = up ? v[i + 8] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const bool up = lane & 4;
        const float send = up ? v[i] : v[i + 4], keep = up ? v[i + 4] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const bool up = lane & 2;
        const float send = up ? v[i] : v[i + 2], keep = up ? v[i + 2] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    {
        const bool up = lane & 1;
        const float send = up ? v[0] : v[1], keep = up ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
    }
    return v[0];
}

struct PairOut {
    float* dmeans2D; float* dconic; float* dopacity; float* dcolors; float* dmeans3D; float* dcov3D; float* dscales;
    float* drots; float* dsh;
};
// destination of output slot L (0..24) for Gaussian gid: the reference's nine gradient tensors
__device__ __forceinline__ float* pair_dest(const PairOut& o, int L, unsigned gid, bool has_sr) {
    if (L < 2) return o.dmeans2D ? o.dmeans2D + 3 * (size_t)gid + L : nullptr;
    if (L < 5) return o.dconic ? o.dconic + 4 * (size_t)gid + (L == 4 ? 3 : L - 2) : nullptr;
    if (L < 6) return o.dopacity ? o.dopacity + gid : nullptr;
    if (L < 9) return o.dcolors ? o.dcolors + 3 * (size_t)gid + (L - 6) : nullptr;
    if (L < 15) return o.dcov3D ? o.dcov3D + 6 * (size_t)gid + (L - 9) : nullptr;
    if (L < 18) return o.dmeans3D ? o.dmeans3D + 3 * (size_t)gid + (L - 15) : nullptr;
    if (L < 21) return (o.dscales && has_sr) ? o.dscales + 3 * (size_t)gid + (L - 18) : nullptr;
    if (L < 25) return (o.drots && has_sr) ? o.drots + 4 * (size_t)gid + (L - 21) : nullptr;
    return nullptr;
}

__global__ void __launch_bounds__(32 * FB_SCORE_WPC) k_bwd_pairs(Bins bins, Cam cam, const float* __restrict__ colors,
                                                                 const float* __restrict__ final_T,
                                                                 const uint32_t* __restrict__ n_contrib,
                                                                 const float* __restrict__ dL_dpix,
                                                                 const float* __restrict__ rows_all, int power, int sh_coeffs,
                                                                 int sh_degree, int has_sr, PairOut out) {
    extern __shared__ float4 smem[];
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const WarpGeom wg = warp_geom(bins, cam, wq);
    const float ex = 0.5f * (FB_WB_X - 1), ey = 0.5f * (FB_WB_Y - 1);
    if (bins.total[1]) return;
    const uint2 range = bins.tile_range[wg.tile_slot];
    const GRec* __restrict__ recs = bins.recs;
    const u64* __restrict__ sorted = bins.keys + range.x;
    WarpQ& q = reinterpret_cast<WarpQ*>(smem)[wq];
    const int pix_id = cam.W * wg.py + wg.px;
    const float T_final = wg.inside ? final_T[pix_id] : 0.f;
    const int last = wg.inside ? (int)n_contrib[pix_id] : 0;
    float dpix[3], bgdot = 0.f;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        dpix[ch] = wg.inside ? dL_dpix[(size_t)ch * cam.H * cam.W + pix_id] : 0.f;
        bgdot += cam.bg[ch] * dpix[ch];
    }
    const int warp_last = min(__reduce_max_sync(0xffffffffu, last), (int)(range.y - range.x));
    if (warp_last == 0) return;
    const float fpow = (float)power;
    const float halfW = 0.5f * cam.W, halfH = 0.5f * cam.H;
    const int nsh = (sh_degree > 0 && out.dsh) ? (sh_degree + 1) * (sh_degree + 1) : 0;
    float T = T_final;
    float accum_rec[3] = {0.f, 0.f, 0.f}, last_color[3] = {0.f, 0.f, 0.f};
    float last_alpha = 0.f;
    int qn = 0, qh = 0;
    const int nsteps = (warp_last + 31) >> 5;
    bool v_nxt = lane < warp_last;
    unsigned gid_nxt = v_nxt ? (unsigned)(sorted[warp_last - 1 - lane] & 0xffffffffull) : 0u;
    GRec r_nxt; r_nxt.a = make_float4(0.f, 0.f, 0.f, 0.f); r_nxt.b = r_nxt.a;
    if (v_nxt) { r_nxt.a = __ldg(&recs[gid_nxt].a); r_nxt.b = __ldg(&recs[gid_nxt].b); }
    u64 key2 = (32 + lane < warp_last) ? sorted[warp_last - 1 - 32 - lane] : 0ull;
    for (int step = 0; step <= nsteps; ++step) {
        if (step < nsteps) {
            const GRec r = r_nxt;
            const unsigned gid = gid_nxt;
            const bool valid = v_nxt;
            v_nxt = (step + 1) * 32 + lane < warp_last;
            gid_nxt = (unsigned)(key2 & 0xffffffffull);
            if (v_nxt) { r_nxt.a = __ldg(&recs[gid_nxt].a); r_nxt.b = __ldg(&recs[gid_nxt].b); }
            key2 = ((step + 2) * 32 + lane < warp_last) ? sorted[warp_last - 1 - (step + 2) * 32 - lane] : 0ull;
            const float2 h = unpack_extents(r.b.w);
            const bool hit = valid && (fabsf(r.a.x - wg.cx) <= h.x + ex) && (fabsf(r.a.y - wg.cy) <= h.y + ey);
            const unsigned mask = __ballot_sync(0xffffffffu, hit);
            if (hit) {
                const int slot = (qh + qn + __popc(mask & lt)) & (FB_WQ - 1);
                q.a[slot] = make_float4(r.a.x, r.a.y, __int_as_float(warp_last - 1 - (step * 32 + lane)), __uint_as_float(gid));
                q.b[slot] = make_float4(r.b.x, r.b.y, r.b.z, r.a.w);
                const float* c = colors + 3 * (size_t)gid;
                q.rb[0][slot] = make_float4(c[0], c[1], c[2], 0.f);
            }
            qn += __popc(mask);
        }
        const bool flush = (step == nsteps);
        if (qn >= 32 || (flush && qn > 0)) {
            const int cnt = min(qn, 32);
            __syncwarp();
            for (int hh = 0; hh < cnt; ++hh) {
                const int slot = qh + hh;
                const float4 a = q.a[slot];
                const float4 co = q.b[slot];
                const float dx = a.x - wg.pxf, dy = a.y - wg.pyf;
                const float pw = pair_power(co.x, co.y, co.z, dx, dy);
                const float G = expf(pw);
                const float alpha = fminf(0.99f, co.w * G);
                const bool ok = __float_as_int(a.z) < last && !(pw > 0.0f) && !(alpha < 1.0f / 255.0f);
                if (!__any_sync(0xffffffffu, ok)) continue;
                const unsigned gid = __float_as_uint(a.w);
                const float4 f = q.rb[0][slot];
                const float col[3] = {f.x, f.y, f.z};
                const float* __restrict__ rows = rows_all + (size_t)gid * FB_ROW_FLOATS;
                float v[32];
#pragma unroll
                for (int k = 0; k < 32; ++k) v[k] = 0.f;
                float dcol[3] = {0.f, 0.f, 0.f};
                if (ok) {
                    // backward.cu:978-1038
                    T = T / (1.f - alpha);
                    const float dchannel_dcolor = alpha * T;
                    float dL_dalpha = 0.0f;
#pragma unroll
                    for (int ch = 0; ch < 3; ++ch) {
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = col[ch];
                        dL_dalpha += (col[ch] - accum_rec[ch]) * dpix[ch];
                        dcol[ch] = dchannel_dcolor * dpix[ch];
                    }
                    dL_dalpha *= T;
                    last_alpha = alpha;
                    dL_dalpha += (-T_final / (1.f - alpha)) * bgdot;
                    const float s = G * (co.w * dL_dalpha);
                    const float w[5] = {s * dx * dx, s * dx * dy, s * dy * dy, s * dx, s * dy};
                    float u[5];
                    phi_to_u(w, co.x, co.y, co.z, halfW, halfH, u);
                    v[0] = u[3]; v[1] = u[4]; v[2] = u[0]; v[3] = u[1]; v[4] = u[2];
                    v[5] = G * dL_dalpha;
                    v[6] = dcol[0]; v[7] = dcol[1]; v[8] = dcol[2];
#pragma unroll
                    for (int k = 0; k < 16; ++k) {
                        float t = 0.f;
#pragma unroll
                        for (int j = 0; j < 5; ++j) t += rows[k * 5 + j] * u[j];
                        v[9 + k] = t;
                    }
                    if (nsh) {
                        // view-direction path of dL/dmeans (backward.cu:124-138), added before the power is applied
#pragma unroll
                        for (int i = 0; i < 3; ++i) {
                            float t = 0.f;
#pragma unroll
                            for (int c = 0; c < 3; ++c) t += rows[80 + i * 3 + c] * (rows[105 + c] * dcol[c]);
                            v[15 + i] += t;
                        }
                    }
#pragma unroll
                    for (int k = 0; k < 25; ++k) v[k] = powf(v[k], fpow);
                }
                const float tot = warp_reduce32(v, lane);
                float* dst = pair_dest(out, lane, gid, has_sr != 0);
                if (dst && tot != 0.f) atomicAdd(dst, tot);
                if (nsh) {
                    // dL_dsh[k][ch] = (dRGB/dsh_k) * dL_dRGB[ch]   (backward.cu:47-104), 3 * nsh values in groups of 32
                    for (int base = 0; base < 3 * nsh; base += 32) {
#pragma unroll
                        for (int k = 0; k < 32; ++k) {
                            const int e = base + k;
                            float t = 0.f;
                            if (ok && e < 3 * nsh) t = powf(rows[89 + e / 3] * (rows[105 + e % 3] * dcol[e % 3]), fpow);
                            v[k] = t;
                        }
                        const float ts = warp_reduce32(v, lane);
                        const int e = base + lane;
                        if (e < 3 * nsh && ts != 0.f) atomicAdd(out.dsh + (size_t)gid * sh_coeffs * 3 + e, ts);
                    }
                }
            }
            __syncwarp();
            qn -= cnt; qh = (qh + cnt) & (FB_WQ - 1);
        }
    }
}

// H_out[g, 0..CH) += sum over the V views of a batch of the view's Fisher diagonal (compute_H_train,
// models/SLAM/gaussian.py:1338-1348).  One thread per Gaussian walks the views in order (deterministic, no atomics
// on H); a view in which the Gaussian collected nothing has an all-zero moment row and is skipped.  Rows are cleared
// after use so the buffer is all-zero again for the next batch.  w2c: V row-major 4x4 matrices.
template <int CH>
__global__ void __launch_bounds__(128) k_expand_H(Bins bins, GaussIn g, Cam cam, const float* __restrict__ w2c_all,
                                                 float* __restrict__ mom_all, float* __restrict__ H_out,
                                                 const int* __restrict__ out_set) {
    // out_set (optional): the term of view v is added to H_out[out_set[v]] of a [sets][P][CH] array (path evaluation:
    // one running Hessian per path); NULL: every view goes to the one [P][CH] tensor.  Consecutive views of one set are
    // summed in registers in view order and added once, as before.
    const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (unsigned)bins.P) return;
    constexpr int NOUT = (CH == 11) ? 10 : 3;
    double Hk[NOUT];
    double Hop = 0.0;
#pragma unroll
    for (int k = 0; k < NOUT; ++k) Hk[k] = 0.0;
    bool any = false, pending = false;
    int cur_set = 0;
    float3 p, sc; float4 q; float cov6[6]; SRCtx sr;
    auto flush = [&]() {
        float* H = H_out + ((size_t)cur_set * bins.P + idx) * CH;
#pragma unroll
        for (int k = 0; k < NOUT; ++k) {
            const int c = (k < 3) ? k : k + 1;
            H[c] += (float)Hk[k];
            Hk[k] = 0.0;
        }
        H[3] += (float)Hop;
        Hop = 0.0;
        pending = false;
    };
    for (int v = 0; v < bins.V; ++v) {
        float acc[FB_NACC];
        float4* row = reinterpret_cast<float4*>(mom_all + ((size_t)v * bins.P + idx) * FB_NACC);
        bool nz = false;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float4 r4 = row[k];
            *reinterpret_cast<float4*>(acc + 4 * k) = r4;
            nz = nz || r4.x != 0.f || r4.y != 0.f || r4.z != 0.f || r4.w != 0.f;
        }
        if (!nz) continue;
#pragma unroll
        for (int k = 0; k < 4; ++k) row[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        const int set = out_set ? out_set[v] : 0;
        if (pending && set != cur_set) flush();
        cur_set = set;
        pending = true;
        if (!any) {
            any = true;
            p = make_float3(g.means3D[3 * idx], g.means3D[3 * idx + 1], g.means3D[3 * idx + 2]);
            sc = make_float3(g.scales[3 * idx], g.scales[3 * idx + 1], g.scales[3 * idx + 2]);
            q = *reinterpret_cast<const float4*>(g.rotations + 4 * (size_t)idx);
            cov3d_from_scale_rot(sc, cam.scale_modifier, q, cov6);
            sr = make_sr(sc, cam.scale_modifier, q);
        }
        const float* w2c = w2c_all + 16 * v;
        const GRec r = bins.recs[(size_t)v * bins.P + idx];
        float3 pc;
        pc.x = fmaf(w2c[2], p.z, fmaf(w2c[1], p.y, w2c[0] * p.x)) + w2c[3];
        pc.y = fmaf(w2c[6], p.z, fmaf(w2c[5], p.y, w2c[4] * p.x)) + w2c[7];
        pc.z = fmaf(w2c[10], p.z, fmaf(w2c[9], p.y, w2c[8] * p.x)) + w2c[11];
        const GCtx ctx = make_ctx(pc, cov6, cam);
        // columns of the chain: means through mean_chain_matrix, scale / rotation through scale_rot_matrix (they
        // depend on phi only through its conic part) -- one pass each instead of five through chain_geom
        const float halfW = 0.5f * cam.W, halfH = 0.5f * cam.H;
        float M[3][5];
        mean_chain_matrix(ctx, cam, r.b.x, r.b.y, r.b.z, halfW, halfH, M);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            double sum = 0.0;
#pragma unroll
            for (int a = 0; a < 5; ++a)
#pragma unroll
                for (int b = 0; b < 5; ++b) sum += (double)M[k][a] * (double)M[k][b] * (double)moment_phi(acc, a, b);
            Hk[k] += (double)(float)sum;          // each view's term is rounded to float as the per-view tensor is
        }
        if (CH == 11) {
            float Z[3][7];
            scale_rot_matrix(ctx, sr, Z);
#pragma unroll
            for (int k = 0; k < 7; ++k) {
                double sum = 0.0;
#pragma unroll
                for (int a = 0; a < 3; ++a)
#pragma unroll
                    for (int b = 0; b < 3; ++b) sum += (double)Z[a][k] * (double)Z[b][k] * (double)moment_phi(acc, a, b);
                Hk[(CH == 11) ? 3 + k : 0] += (double)(float)sum;
            }
        }
        Hop += (double)(acc[0] / (r.b.w * r.b.w));          // batched record layout: opacity in b.w
    }
    if (pending) flush();
}

}  // namespace fb
