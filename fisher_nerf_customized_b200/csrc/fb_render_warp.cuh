// fb_render_warp.cuh -- the drop-in single-view forward render and moment backward in warp-autonomous form.
//
// Same per-pixel arithmetic, in the same order, as renderCUDA<3> (DGR/cuda_rasterizer/forward.cu:261-393) and the
// pair loop of renderCUDAFused<3> (backward.cu:939-1038); what changes is who visits what.  The reference
// lets each of a tile's 256 pixels test every
// instance binned to the tile, with a block barrier per 256 instances.  Here a warp owns an 8x4 pixel block,
// walks the tile's sorted list on its own, 32 entries per step, drops the entries whose alpha >= 1/255 footprint
// box misses the block (pairs the reference's own alpha test skips; skipped pairs change no state, so every
// pixel's result is bit-identical) and runs the pair arithmetic only for the survivors, whose colour arrives by
// cp.async in a per-warp ring.  No block barriers; a warp whose 32 pixels are saturated stops.
// 200k Gaussians at 256x256 (BASELINE config 2): forward 156 -> ~40 us, backward 230 -> ~60 us.
#pragma once
#include "fb_score.cuh"

namespace fb {

struct WarpGeom {
    int tile_slot, warp, px, py;
    bool inside;
    float pxf, pyf, cx, cy;
};
__device__ __forceinline__ WarpGeom warp_geom(const Bins& bins, const Cam& cam, int wq) {
    constexpr int CPT = FB_TILE_WARPS / FB_SCORE_WPC;
    const unsigned lin = blockIdx.y * gridDim.x + blockIdx.x;
    WarpGeom g;
    g.tile_slot = bins.tile_order ? (int)bins.tile_order[lin / CPT] : (int)(lin / CPT);
    const int t = g.tile_slot % bins.tiles;
    g.warp = (int)(lin % CPT) * FB_SCORE_WPC + wq;
    const int lane = threadIdx.x & 31;
    const int tx = t % cam.tiles_x, ty = t / cam.tiles_x;
    const int bx = tx * FB_TILE_X + (g.warp & 1) * FB_WB_X, by = ty * FB_TILE_Y + (g.warp >> 1) * FB_WB_Y;
    g.px = bx + (lane & 7); g.py = by + (lane >> 3);
    g.inside = g.px < cam.W && g.py < cam.H;
    g.pxf = (float)g.px; g.pyf = (float)g.py;
    g.cx = (float)bx + 0.5f * (FB_WB_X - 1); g.cy = (float)by + 0.5f * (FB_WB_Y - 1);
    return g;
}

// ---- forward: out_color[3,H,W], out_depth[H,W], final_T, n_contrib ------------------------------------------
// NX = 1: a second colour triple per Gaussian (`extra`, e.g. SplaTAM's [depth, 1, depth^2], slam_helpers.py:235-252)
// is blended in the same sweep into out_extra[3,H,W]: the reference rasterises twice for it (gaussian.py:207,211),
// binning and sorting the same geometry twice.
template <int NX>
__global__ void __launch_bounds__(32 * FB_SCORE_WPC) k_render_warp(Bins bins, Cam cam, const float* __restrict__ features,
                                                                   const float* __restrict__ extra,
                                                                   float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                                                                   float* __restrict__ out_color, float* __restrict__ out_extra,
                                                                   float* __restrict__ out_depth) {
    extern __shared__ float4 smem[];
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const WarpGeom wg = warp_geom(bins, cam, wq);
    const float ex = 0.5f * (FB_WB_X - 1), ey = 0.5f * (FB_WB_Y - 1);
    const uint2 range = bins.total[1] ? make_uint2(0u, 0u) : bins.tile_range[wg.tile_slot];
    const int n = (int)(range.y - range.x);
    const GRec* __restrict__ recs = bins.recs;
    const u64* __restrict__ sorted = bins.keys + range.x;
    WarpQ& q = reinterpret_cast<WarpQ*>(smem)[wq];

    float T = 1.0f, D = 15.0f;
    float C[3] = {0.f, 0.f, 0.f};
    float X[3] = {0.f, 0.f, 0.f};
    int last = 0;
    bool done = !wg.inside;
    if (n > 0 && !__all_sync(0xffffffffu, done)) {
        int qn = 0, qh = 0;
        const int nsteps = (n + 31) >> 5;
        bool v_nxt = lane < n;
        unsigned gid_nxt = v_nxt ? (unsigned)(sorted[lane] & 0xffffffffull) : 0u;
        GRec r_nxt; r_nxt.a = make_float4(0.f, 0.f, 0.f, 0.f); r_nxt.b = r_nxt.a;
        if (v_nxt) { r_nxt.a = __ldg(&recs[gid_nxt].a); r_nxt.b = __ldg(&recs[gid_nxt].b); }
        u64 key2 = (32 + lane < n) ? sorted[32 + lane] : 0ull;
        for (int step = 0; step <= nsteps; ++step) {
            if (step < nsteps) {
                const GRec r = r_nxt;
                const unsigned gid = gid_nxt;
                const bool valid = v_nxt;
                v_nxt = (step + 1) * 32 + lane < n;
                gid_nxt = (unsigned)(key2 & 0xffffffffull);
                if (v_nxt) { r_nxt.a = __ldg(&recs[gid_nxt].a); r_nxt.b = __ldg(&recs[gid_nxt].b); }
                key2 = ((step + 2) * 32 + lane < n) ? sorted[(step + 2) * 32 + lane] : 0ull;
                const float2 h = unpack_extents(r.b.w);
                const bool hit = valid && (fabsf(r.a.x - wg.cx) <= h.x + ex) && (fabsf(r.a.y - wg.cy) <= h.y + ey);
                const unsigned mask = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    const int slot = (qh + qn + __popc(mask & lt)) & (FB_WQ - 1);
                    q.a[slot] = make_float4(r.a.x, r.a.y, __int_as_float(step * 32 + lane), r.a.z);     // px, py, position, depth
                    q.b[slot] = make_float4(r.b.x, r.b.y, r.b.z, r.a.w);                                // conic, opacity
                    const float* c = features + 3 * (size_t)gid;
                    float* dst = reinterpret_cast<float*>(&q.rb[0][slot]);
#pragma unroll
                    for (int k = 0; k < 3; ++k) __pipeline_memcpy_async(dst + k, c + k, 4);
                    if (NX) {
                        const float* c2 = extra + 3 * (size_t)gid;
                        float* dst2 = reinterpret_cast<float*>(&q.rb[1][slot]);
#pragma unroll
                        for (int k = 0; k < 3; ++k) __pipeline_memcpy_async(dst2 + k, c2 + k, 4);
                    }
                }
                __pipeline_commit();
                qn += __popc(mask);
            }
            const bool flush = (step == nsteps);
            if (qn >= 32 || (flush && qn > 0)) {
                const int cnt = min(qn, 32);
                __pipeline_wait_prior(0);
                __syncwarp();
                for (int hh = 0; hh < cnt; ++hh) {
                    const int slot = qh + hh;
                    const float4 a = q.a[slot];
                    const float4 co = q.b[slot];
                    const float dx = a.x - wg.pxf, dy = a.y - wg.pyf;
                    const float power = pair_power(co.x, co.y, co.z, dx, dy);
                    const float alpha = fminf(0.99f, co.w * expf(power));
                    const float test_T = __fmul_rn(T, __fsub_rn(1.f, alpha));
                    const bool ok = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                    const bool term = ok && test_T < 0.0001f;
                    done = done || term;
                    const bool upd = ok && !term;
                    if (!__any_sync(0xffffffffu, upd)) continue;
                    const float4 f = q.rb[0][slot];
                    const float c0 = __fmaf_rn(T, __fmul_rn(alpha, f.x), C[0]);     // as compiled (forward.cu:361)
                    const float c1 = __fmaf_rn(T, __fmul_rn(alpha, f.y), C[1]);
                    const float c2 = __fmaf_rn(T, __fmul_rn(alpha, f.z), C[2]);
                    C[0] = upd ? c0 : C[0]; C[1] = upd ? c1 : C[1]; C[2] = upd ? c2 : C[2];
                    if (NX) {
                        const float4 f2 = q.rb[1][slot];
                        const float x0 = __fmaf_rn(T, __fmul_rn(alpha, f2.x), X[0]);
                        const float x1 = __fmaf_rn(T, __fmul_rn(alpha, f2.y), X[1]);
                        const float x2 = __fmaf_rn(T, __fmul_rn(alpha, f2.z), X[2]);
                        X[0] = upd ? x0 : X[0]; X[1] = upd ? x1 : X[1]; X[2] = upd ? x2 : X[2];
                    }
                    D = (upd && T > 0.5f && (double)test_T < 0.5) ? a.w : D;       // median depth, forward.cu:367-371
                    T = upd ? test_T : T;
                    last = upd ? __float_as_int(a.z) + 1 : last;
                }
                __syncwarp();
                qn -= cnt; qh = (qh + cnt) & (FB_WQ - 1);
                if (__all_sync(0xffffffffu, done)) break;
            }
        }
        __pipeline_wait_prior(0);
        __syncwarp();
    }
    if (wg.inside) {
        const int W = cam.W, H = cam.H;
        const int pix_id = W * wg.py + wg.px;
        final_T[pix_id] = T;
        n_contrib[pix_id] = (uint32_t)last;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) out_color[(size_t)ch * H * W + pix_id] = __fmaf_rn(T, cam.bg[ch], C[ch]);
        if (NX) {
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) out_extra[(size_t)ch * H * W + pix_id] = __fmaf_rn(T, cam.bg[ch], X[ch]);
        }
        out_depth[pix_id] = D;
    }
}

// ---- backward: per-Gaussian moment rows mom[P][16] (consumed by k_expand_grads) ------------------------------
// P2: backward_power == 2 (13 monomial moments of s^2) or 1 (6 sums of s); colour sums in 13..15 either way.
// NX = 1 (power 1 only: gradients of the two renders ADD, squares would not): the second colour triple and its image
// gradient dL_dextra ride along.  dL/dalpha sums over the six channels, so the six geometry sums (slots 0-5) are those of
// both renders together; slots 6-7 keep sum(s_rgb dx), sum(s_rgb dy) of the colour render alone (the reference takes
// the densification statistic means2D.grad from the colour render only, gaussian.py:208); slots 8-10 are the extra
// colours' gradients, 13-15 the colours'.
template <bool P2, int NX>
__global__ void __launch_bounds__(32 * FB_SCORE_WPC) k_bwd_warp(Bins bins, Cam cam, const float* __restrict__ colors,
                                                                const float* __restrict__ extra,
                                                                const float* __restrict__ final_T,
                                                                const uint32_t* __restrict__ n_contrib,
                                                                const float* __restrict__ dL_dpix,
                                                                const float* __restrict__ dL_dextra, float* __restrict__ mom) {
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
    float dpix[3], dext[3] = {0.f, 0.f, 0.f}, bgdot = 0.f, bgdot_x = 0.f;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        dpix[ch] = wg.inside ? dL_dpix[(size_t)ch * cam.H * cam.W + pix_id] : 0.f;
        bgdot += cam.bg[ch] * dpix[ch];
        if (NX) {
            dext[ch] = wg.inside ? dL_dextra[(size_t)ch * cam.H * cam.W + pix_id] : 0.f;
            bgdot_x += cam.bg[ch] * dext[ch];
        }
    }
    float accum_x[3] = {0.f, 0.f, 0.f}, last_x[3] = {0.f, 0.f, 0.f};
    const int warp_last = min(__reduce_max_sync(0xffffffffu, last), (int)(range.y - range.x));
    if (warp_last == 0) return;
    float T = T_final;
    float accum_rec[3] = {0.f, 0.f, 0.f}, last_color[3] = {0.f, 0.f, 0.f};
    float last_alpha = 0.f;
    const int rslot = reduce16_slot(lane);
    const float halfW = 0.5f * cam.W, halfH = 0.5f * cam.H;
    (void)halfW; (void)halfH;
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
                float* dst = reinterpret_cast<float*>(&q.rb[0][slot]);
#pragma unroll
                for (int k = 0; k < 3; ++k) __pipeline_memcpy_async(dst + k, c + k, 4);
                if (NX) {
                    const float* c2 = extra + 3 * (size_t)gid;
                    float* dst2 = reinterpret_cast<float*>(&q.rb[1][slot]);
#pragma unroll
                    for (int k = 0; k < 3; ++k) __pipeline_memcpy_async(dst2 + k, c2 + k, 4);
                }
            }
            __pipeline_commit();
            qn += __popc(mask);
        }
        const bool flush = (step == nsteps);
        if (qn >= 32 || (flush && qn > 0)) {
            const int cnt = min(qn, 32);
            __pipeline_wait_prior(0);
            __syncwarp();
            for (int hh = 0; hh < cnt; ++hh) {
                const int slot = qh + hh;
                const float4 a = q.a[slot];
                const float4 co = q.b[slot];
                const float dx = a.x - wg.pxf, dy = a.y - wg.pyf;
                const float power = pair_power(co.x, co.y, co.z, dx, dy);
                const float G = expf(power);
                const float alpha = fminf(0.99f, co.w * G);
                const bool ok = __float_as_int(a.z) < last && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                if (!__any_sync(0xffffffffu, ok)) continue;
                const float4 f = q.rb[0][slot];
                const float col[3] = {f.x, f.y, f.z};
                float v[FB_NACC];
#pragma unroll
                for (int k = 0; k < FB_NACC; ++k) v[k] = 0.f;
                if (ok) {
                    // backward.cu:978-1038, the reference's statements in the same order
                    T = T / (1.f - alpha);
                    const float dchannel_dcolor = alpha * T;
                    float dL_dalpha = 0.0f;
#pragma unroll
                    for (int ch = 0; ch < 3; ++ch) {
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = col[ch];
                        dL_dalpha += (col[ch] - accum_rec[ch]) * dpix[ch];
                        const float cc = dchannel_dcolor * dpix[ch];
                        v[13 + ch] = P2 ? cc * cc : cc;
                    }
                    dL_dalpha *= T;
                    float dL_dalpha_x = 0.0f;
                    if (NX) {
                        const float4 f2 = q.rb[1][slot];
                        const float colx[3] = {f2.x, f2.y, f2.z};
#pragma unroll
                        for (int ch = 0; ch < 3; ++ch) {
                            accum_x[ch] = last_alpha * last_x[ch] + (1.f - last_alpha) * accum_x[ch];
                            last_x[ch] = colx[ch];
                            dL_dalpha_x += (colx[ch] - accum_x[ch]) * dext[ch];
                            v[8 + ch] = dchannel_dcolor * dext[ch];
                        }
                        dL_dalpha_x *= T;
                        dL_dalpha_x += (-T_final / (1.f - alpha)) * bgdot_x;
                    }
                    last_alpha = alpha;
                    dL_dalpha += (-T_final / (1.f - alpha)) * bgdot;
                    const float s_rgb = G * (co.w * dL_dalpha);
                    const float s = NX ? s_rgb + G * (co.w * dL_dalpha_x) : s_rgb;
                    if (NX) { v[6] = s_rgb * dx; v[7] = s_rgb * dy; }
                    if (P2) {
                        const float qq = s * s;
                        const float qx = qq * dx, qy = qq * dy;
                        const float m20 = qx * dx, m11 = qx * dy, m02 = qy * dy;
                        const float m30 = m20 * dx, m21 = m20 * dy, m12 = m02 * dx, m03 = m02 * dy;
                        v[0] = qq; v[1] = m20; v[2] = m11; v[3] = m02;
                        v[4] = m30; v[5] = m21; v[6] = m12; v[7] = m03;
                        v[8] = m30 * dx; v[9] = m30 * dy; v[10] = m21 * dy; v[11] = m03 * dx; v[12] = m03 * dy;
                    } else {
                        const float sx = s * dx, sy = s * dy;
                        v[0] = s; v[1] = sx * dx; v[2] = sx * dy; v[3] = sy * dy; v[4] = sx; v[5] = sy;
                    }
                }
                const float tot = warp_reduce16(v, lane);
                float4 quad;
                quad.x = tot;
                quad.y = __shfl_down_sync(0xffffffffu, tot, 2);
                quad.z = __shfl_down_sync(0xffffffffu, tot, 4);
                quad.w = __shfl_down_sync(0xffffffffu, tot, 6);
                if ((lane & 7) == 0 && (quad.x != 0.f || quad.y != 0.f || quad.z != 0.f || quad.w != 0.f))
                    atomicAdd(reinterpret_cast<float4*>(mom + (size_t)__float_as_uint(a.w) * FB_NACC) + (rslot >> 2), quad);
            }
            __syncwarp();
            qn -= cnt; qh = (qh + cnt) & (FB_WQ - 1);
        }
    }
}

}  // namespace fb
