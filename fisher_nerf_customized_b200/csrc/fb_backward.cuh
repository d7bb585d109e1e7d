// fb_backward.cuh -- drop-in backward: tile walk -> per-Gaussian moment rows -> expansion into the
// reference's nine gradient tensors (RasterizeGaussiansBackwardCUDA, DGR/rasterize_points.cu:117-196).
#pragma once
#include "fb_fisher_tile.cuh"

namespace fb {

struct MomFlush {
    float* mom;  // [P][16]
    __device__ __forceinline__ void operator()(int, int gid, const float* row) const {
        float4* dst = reinterpret_cast<float4*>(mom + (size_t)gid * FB_NACC);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float4 v = *reinterpret_cast<const float4*>(row + 4 * k);
            if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f) atomicAdd(dst + k, v);  // red.global.add.v4.f32
        }
    }
};

template <bool P2, bool COLOR>
__global__ void __launch_bounds__(FB_TILE_PIX) k_bwd_tiles(Bins bins, Cam cam, const float* __restrict__ colors,
                                                          const float* __restrict__ final_T,
                                                          const uint32_t* __restrict__ n_contrib,
                                                          const float* __restrict__ dL_dpix, float* __restrict__ mom) {
    __shared__ BwdSmem sm;
    // blockIdx.z = view of a batch: per-view records, ranges, pixel state and moment rows; dL_dpix is shared
    int tx = blockIdx.x, ty = blockIdx.y, v = blockIdx.z;
    if (bins.tile_order) {            // batched: heaviest (view, tile) slots first
        const unsigned slot = bins.tile_order[(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x];
        v = (int)(slot / (unsigned)bins.tiles);
        const int t = (int)(slot % (unsigned)bins.tiles);
        tx = t % cam.tiles_x; ty = t / cam.tiles_x;
    }
    {
        const size_t HWv = (size_t)cam.W * cam.H * v;
        final_T += HWv; n_contrib += HWv;
        bins.recs += (size_t)v * bins.P;
        bins.tile_range += (size_t)v * bins.tiles;
        mom += (size_t)v * bins.P * FB_NACC;
    }
    const int px = tx * FB_TILE_X + threadIdx.x % FB_TILE_X, py = ty * FB_TILE_Y + threadIdx.x / FB_TILE_X;
    const bool inside = px < cam.W && py < cam.H;
    const int pix_id = cam.W * py + px;
    if (bins.total[1]) return;
    uint2 range = bins.tile_range[ty * cam.tiles_x + tx];
    PixState ps;
    ps.T_final = inside ? final_T[pix_id] : 0.f;
    ps.last = inside ? (int)n_contrib[pix_id] : 0;
    ps.bgdot = 0.f;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        ps.dpix[ch] = inside ? dL_dpix[(size_t)ch * cam.H * cam.W + pix_id] : 0.f;
        ps.bgdot += cam.bg[ch] * ps.dpix[ch];
    }
    MomFlush fl{mom};
    tile_backward<P2, COLOR>(sm, bins.keys + range.x, (int)(range.y - range.x), bins.recs, colors, ps, (float)px, (float)py,
                            0.5f * cam.W, 0.5f * cam.H, fl);
}

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
        float col[5][NOUT];
        const float halfW = 0.5f * cam.W, halfH = 0.5f * cam.H;
#pragma unroll
        for (int b = 0; b < 5; ++b) {
            float e[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
            e[b] = 1.f;
            float u[5];
            phi_to_u(e, r.b.x, r.b.y, r.b.z, halfW, halfH, u);
            ChainOut co = chain_geom(ctx, cam, u[0], u[1], u[2], u[3], u[4]);
            col[b][0] = co.mean[0]; col[b][1] = co.mean[1]; col[b][2] = co.mean[2];
            if (CH == 11) {
                float ds[3], dr[4];
                chain_scale_rot(sr, co.cov3D, ds, dr);
                col[b][3] = ds[0]; col[b][4] = ds[1]; col[b][5] = ds[2];
                col[b][6] = dr[0]; col[b][7] = dr[1]; col[b][8] = dr[2]; col[b][9] = dr[3];
            }
        }
#pragma unroll
        for (int k = 0; k < NOUT; ++k) {
            double sum = 0.0;
#pragma unroll
            for (int a = 0; a < 5; ++a)
#pragma unroll
                for (int b = 0; b < 5; ++b) sum += (double)col[a][k] * (double)col[b][k] * (double)moment_phi(acc, a, b);
            Hk[k] += (double)(float)sum;          // each view's term is rounded to float as the per-view tensor is
        }
        Hop += (double)(acc[0] / (r.b.w * r.b.w));          // batched record layout: opacity in b.w
    }
    if (pending) flush();
}

}  // namespace fb
