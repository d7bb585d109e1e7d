// fb_score.cuh -- fused per-(view, tile) candidate-scoring kernel.
//
// One CTA owns one 16x16 tile of one candidate view; its depth-sorted key segment was produced by
// k_scatter_keys + k_sort_tiles.  Without leaving the SM it does
//   (1) front-to-back transmittance pass -> T_final, n_contrib per pixel (registers)      [stage 3]
//   (2) back-to-front pass: per contributing pair, s = G*o*dL/dalpha and
//       score += s^2 * poly_g(dx,dy)                                                      [stage 4]
//   (3) warp reduce -> one float per (view, tile, pixel block); k_score_finalize sums them in fixed order.
// No image, no final_T/n_contrib round trip, no per-Gaussian Hessian tensor, no atomics.
// Replaces the K-iteration Python loop  models/SLAM/gaussian.py:1362-1370  (+ compute_Hessian
// :1503-1570) around renderCUDA / renderCUDAFused (forward.cu:261-393, backward.cu:850-1140).
//
// Pair culling.  The reference tests every pixel of a tile against every Gaussian binned to it and
// rejects most pairs with `power > 0` / `alpha < 1/255`.  Here each warp owns an 8x4 pixel block and
// first tests 32 staged Gaussians at once (one per lane) against the block with the precomputed half
// extents of the alpha >= 1/255 footprint (GRec.a.zw, see k_preprocess); a ballot gives the short,
// still depth-ordered list of Gaussians that can touch the block, and only those run the pair
// arithmetic.  A culled pair is one the reference's own alpha test skips, and skipped pairs change no
// state (T, accum_rec, last_alpha are only updated after the tests), so every pixel's result is
// unchanged; the per-pair arithmetic itself is identical (pair_power / expf as compiled).
#pragma once
#include "fb_stage12.cuh"
#include <cuda_pipeline.h>

namespace fb {

#define FB_WB_X 8      // warp pixel block
#define FB_WB_Y 4
// pixel block of the single-sweep scoring kernel (32 pixels: 8x4 default; 4x8 and 16x2 measured)
#ifndef FB_S1_WX
#define FB_S1_WX 8
#endif
#define FB_S1_WY (32 / FB_S1_WX)
#ifndef FB_S1_UNROLL
#define FB_S1_UNROLL 8
#endif
constexpr int kS1Unroll = FB_S1_UNROLL;   // (#pragma unroll takes a constant expression, not a macro)

// Warp-autonomous traversal.  There is NO block-wide barrier in the pair loops: each warp walks the
// tile's sorted list on its own, 32 entries per step (one per lane: key, then the 16-byte {px,py,hx,hy}
// half of the record, both prefetched one / two steps ahead), culls them against its 8x4 pixel block
// and appends the survivors, in order, to a 64-entry ring in shared memory that is private to the
// warp.  The appending lane also starts a cp.async of the rest of the record (conic+opacity, and in the
// backward pass the 64-byte scoring record) into the ring slot, so by the time 32 survivors have
// queued up their data has landed; the warp then runs the pair arithmetic for those 32 entries with
// broadcast shared-memory reads.  Warps of a tile never wait for each other (pixel blocks differ a lot
// in how many Gaussians touch them), and a warp whose 32 pixels are saturated stops scanning.
#define FB_WQ 64   // ring entries per warp
struct WarpQ {
    float4 a[FB_WQ];       // px, py, list position (as int bits), -
    float4 b[FB_WQ];       // conic a, b, c, opacity
    float4 rb[4][FB_WQ];   // rgb+coef0 | coef1-4 | coef5-8 | coef9-12   (backward pass only)
};
// Warps per CTA.  A warp never talks to another warp, so the CTA is only a launch unit: small CTAs let the
// SM backfill as soon as a warp's pixel block is finished (with 8-warp CTAs 23 % of all warp samples sat in
// the final barrier waiting for the slowest block of the tile; profiles/r1_summary_v4.md).
#ifndef FB_SCORE_WPC
#define FB_SCORE_WPC 4
#endif
#define FB_TILE_WARPS (FB_TILE_PIX / 32)
// FB_SCORE_FAST (default): the scoring path evaluates exp with MUFU.EX2 (ex2.approx(x*log2e), |x| <= 5.6 here so
// <= 6 ulp), 1/(1-alpha) with MUFU.RCP, and runs the colour recurrence of backward.cu:997-1008 once on r+g+b
// (dL/dpix is the same scalar for the three channels in compute_Hessian, gaussian.py:1549, and the recurrence is
// linear).  These differ from the reference's compiled arithmetic by a few ulp per pair; scores stay far inside the
// stated rtol 1e-3 (tests/test_gpu_scoring.py).  The drop-in rasteriser kernels keep the as-compiled arithmetic.
#ifndef FB_SCORE_FAST
#define FB_SCORE_FAST 1
#endif
// FB_SCORE_SINGLE_PASS (default): k_tile_score1 below; 0 selects the two-sweep kernel k_tile_score.
#ifndef FB_SCORE_SINGLE_PASS
#define FB_SCORE_SINGLE_PASS 1
#endif
__device__ __forceinline__ float score_exp(float x) {
#if FB_SCORE_FAST
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x * 1.4426950408889634f));   // FMUL + MUFU.EX2
    return r;
#else
    return expf(x);
#endif
}
__device__ __forceinline__ float score_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));                          // MUFU.RCP
    return r;
}
// FB_SCORE_PRESCALE (default with the single-sweep fast path): k_preprocess_score stores the conic pre-multiplied by
// -log2(e)/2 | -log2(e) and r+g+b instead of rgb, so k_tile_score1's pair body needs no exponent scaling and no
// colour sum (61 -> 54 instructions per surviving entry).  The two-sweep kernel reads the unscaled records.
// FB_HESS_FAST: k_tile_hessian (H_train path) with the scoring kernel's MUFU exp / reciprocal instead of expf and the
// IEEE division of backward.cu:987 (both sweeps use the same exp, so their alpha decisions stay consistent); the H_train
// path is held to rtol 1e-3 (tests/test_gpu_scoring.py), the MUFU forms are ~1e-6.
#ifndef FB_HESS_FAST
#define FB_HESS_FAST 0            // build switch; the default keeps the reference's arithmetic
#endif
__device__ __forceinline__ float hess_exp(float x) {
#if FB_HESS_FAST
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x * 1.4426950408889634f));
    return r;
#else
    return expf(x);
#endif
}
__device__ __forceinline__ float hess_div(float a, float b) {
#if FB_HESS_FAST
    return a * score_rcp(b);
#else
    return a / b;
#endif
}
#define FB_SCORE_SMEM_BYTES ((int)sizeof(WarpQ) * FB_SCORE_WPC)

__global__ void __launch_bounds__(32 * FB_SCORE_WPC) k_tile_score(Bins bins, Cam cam, float grad_scale,
                                                           float* __restrict__ partial) {
    extern __shared__ float4 smem[];
    constexpr int CPT = FB_TILE_WARPS / FB_SCORE_WPC;      // CTAs per tile
    // CTAs are dispatched in linear block order; tile_order lists the (view, tile) slots of the chunk by
    // descending instance count, so the longest lists start first and the grid drains evenly.
    const unsigned lin = blockIdx.y * gridDim.x + blockIdx.x;
    const int tile_slot = (int)bins.tile_order[lin / CPT];
    const int t = tile_slot % bins.tiles, v = tile_slot / bins.tiles;
    const int tid = threadIdx.x, lane = tid & 31, wq = tid >> 5;
    const int warp = (int)(lin % CPT) * FB_SCORE_WPC + wq;   // pixel block of the tile, 0..7
    const unsigned lt = (1u << lane) - 1u;
    const int tx = t % cam.tiles_x, ty = t / cam.tiles_x;
    // warp -> 8x4 pixel block of the tile, lane -> pixel of the block
    const int bx = tx * FB_TILE_X + (warp & 1) * FB_WB_X, by = ty * FB_TILE_Y + (warp >> 1) * FB_WB_Y;
    const int px = bx + (lane & 7), py = by + (lane >> 3);
    const bool inside = px < cam.W && py < cam.H;
    const float pxf = (float)px, pyf = (float)py;
    const float cx = (float)bx + 0.5f * (FB_WB_X - 1), cy = (float)by + 0.5f * (FB_WB_Y - 1);
    const float ex = 0.5f * (FB_WB_X - 1), ey = 0.5f * (FB_WB_Y - 1);
    float* out = partial + (size_t)tile_slot * FB_TILE_WARPS + warp;
    if (bins.total[1]) { if (lane == 0) *out = 0.f; return; }
    const uint2 range = bins.tile_range[tile_slot];
    const int n = (int)(range.y - range.x);
    if (n == 0) { if (lane == 0) *out = 0.f; return; }
    const GRec* __restrict__ recs = bins.recs + (size_t)v * bins.P;
    const float4* __restrict__ recsB = bins.recsB + (size_t)v * bins.P * 4;
    const u64* __restrict__ sorted = bins.keys + range.x;
    WarpQ& q = reinterpret_cast<WarpQ*>(smem)[wq];

    // (1) transmittance pass (forward.cu:330-379 without colour / depth)
    float T = 1.0f;
    int last = 0;
    bool done = !inside;
    if (!__all_sync(0xffffffffu, done)) {
        int qn = 0, qh = 0;
        const int nsteps = (n + 31) >> 5;
        bool v_nxt = lane < n;
        unsigned gid_nxt = v_nxt ? (unsigned)(sorted[lane] & 0xffffffffull) : 0u;
        float4 a_nxt = make_float4(0.f, 0.f, 0.f, 0.f);
        if (v_nxt) a_nxt = __ldg(reinterpret_cast<const float4*>(recs + gid_nxt));
        u64 key2 = (32 + lane < n) ? sorted[32 + lane] : 0ull;
        for (int step = 0; step <= nsteps; ++step) {
            unsigned mask = 0u;
            if (step < nsteps) {
                const float4 a = a_nxt;
                const unsigned gid = gid_nxt;
                const bool valid = v_nxt;
                v_nxt = (step + 1) * 32 + lane < n;
                gid_nxt = (unsigned)(key2 & 0xffffffffull);
                if (v_nxt) a_nxt = __ldg(reinterpret_cast<const float4*>(recs + gid_nxt));
                key2 = ((step + 2) * 32 + lane < n) ? sorted[(step + 2) * 32 + lane] : 0ull;
                const bool hit = valid && (fabsf(a.x - cx) <= a.z + ex) && (fabsf(a.y - cy) <= a.w + ey);
                mask = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    const int slot = (qh + qn + __popc(mask & lt)) & (FB_WQ - 1);
                    q.a[slot] = make_float4(a.x, a.y, __int_as_float(step * 32 + lane), 0.f);
                    __pipeline_memcpy_async(&q.b[slot], reinterpret_cast<const float4*>(recs + gid) + 1, 16);
                }
                __pipeline_commit();
                qn += __popc(mask);
            }
            const bool flush = (step == nsteps);
            if (qn >= 32 || (flush && qn > 0)) {
                const int cnt = min(qn, 32);
                __pipeline_wait_prior(0);
                __syncwarp();
#if FB_SCORE_FAST
                // branch-free: nearly every entry of the ring has a contributing lane, so the skips of
                // forward.cu:343-358 become predicates and selects instead of divergent branches
#pragma unroll 4
                for (int h = 0; h < cnt; ++h) {
                    const int slot = (qh + h) & (FB_WQ - 1);
                    const float4 a = q.a[slot];
                    const float4 co = q.b[slot];
                    const float dx = a.x - pxf, dy = a.y - pyf;
                    const float power = pair_power(co.x, co.y, co.z, dx, dy);
                    const float alpha = fminf(0.99f, co.w * score_exp(power));
                    const float test_T = __fmul_rn(T, __fsub_rn(1.f, alpha));
                    const bool ok = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                    const bool term = ok && test_T < 0.0001f;
                    done = done || term;
                    const bool upd = ok && !term;
                    T = upd ? test_T : T;
                    last = upd ? __float_as_int(a.z) + 1 : last;
                }
#else
                for (int h = 0; h < cnt; ++h) {
                    if (done) continue;
                    const int slot = (qh + h) & (FB_WQ - 1);
                    const float4 a = q.a[slot];
                    const float4 co = q.b[slot];
                    const float dx = a.x - pxf, dy = a.y - pyf;
                    const float power = pair_power(co.x, co.y, co.z, dx, dy);
                    if (power > 0.0f) continue;
                    const float alpha = fminf(0.99f, co.w * score_exp(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float test_T = __fmul_rn(T, __fsub_rn(1.f, alpha));
                    if (test_T < 0.0001f) { done = true; continue; }
                    T = test_T;
                    last = __float_as_int(a.z) + 1;
                }
#endif
                __syncwarp();
                qn -= cnt; qh = (qh + cnt) & (FB_WQ - 1);
                if (__all_sync(0xffffffffu, done)) break;
            }
        }
        __pipeline_wait_prior(0);
        __syncwarp();
    }

    // (2) back-to-front scoring pass (backward.cu:939-1038)
    const int warp_last = __reduce_max_sync(0xffffffffu, last);
    const float T_final = T;
    const float dpix = inside ? grad_scale : 0.f;
    const float bgdot = (cam.bg[0] + cam.bg[1] + cam.bg[2]) * dpix;
    float accum_rec[3] = {0.f, 0.f, 0.f}, last_color[3] = {0.f, 0.f, 0.f};
    float last_alpha = 0.f;
    float score = 0.f;
    if (warp_last > 0) {
        // step s, lane l looks at list position warp_last-1 - (32 s + l): descending depth order
        int qn = 0, qh = 0;
        const int nsteps = (warp_last + 31) >> 5;
        bool v_nxt = lane < warp_last;
        unsigned gid_nxt = v_nxt ? (unsigned)(sorted[warp_last - 1 - lane] & 0xffffffffull) : 0u;
        float4 a_nxt = make_float4(0.f, 0.f, 0.f, 0.f);
        if (v_nxt) a_nxt = __ldg(reinterpret_cast<const float4*>(recs + gid_nxt));
        u64 key2 = (32 + lane < warp_last) ? sorted[warp_last - 1 - 32 - lane] : 0ull;
        for (int step = 0; step <= nsteps; ++step) {
            if (step < nsteps) {
                const float4 a = a_nxt;
                const unsigned gid = gid_nxt;
                const bool valid = v_nxt;
                v_nxt = (step + 1) * 32 + lane < warp_last;
                gid_nxt = (unsigned)(key2 & 0xffffffffull);
                if (v_nxt) a_nxt = __ldg(reinterpret_cast<const float4*>(recs + gid_nxt));
                key2 = ((step + 2) * 32 + lane < warp_last) ? sorted[warp_last - 1 - (step + 2) * 32 - lane] : 0ull;
                const bool hit = valid && (fabsf(a.x - cx) <= a.z + ex) && (fabsf(a.y - cy) <= a.w + ey);
                const unsigned mask = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    const int slot = (qh + qn + __popc(mask & lt)) & (FB_WQ - 1);
                    q.a[slot] = make_float4(a.x, a.y, __int_as_float(warp_last - 1 - (step * 32 + lane)), 0.f);
                    __pipeline_memcpy_async(&q.b[slot], reinterpret_cast<const float4*>(recs + gid) + 1, 16);
                    const float4* rb = recsB + (size_t)gid * 4;
#pragma unroll
                    for (int k = 0; k < 4; ++k) __pipeline_memcpy_async(&q.rb[k][slot], rb + k, 16);
                }
                __pipeline_commit();
                qn += __popc(mask);
            }
            const bool flush = (step == nsteps);
            if (qn >= 32 || (flush && qn > 0)) {
                const int cnt = min(qn, 32);
                __pipeline_wait_prior(0);
                __syncwarp();
#if FB_SCORE_FAST
                for (int h = 0; h < cnt; ++h) {
                    const int slot = (qh + h) & (FB_WQ - 1);
                    const float4 a = q.a[slot];
                    const float4 co = q.b[slot];
                    const float dx = a.x - pxf, dy = a.y - pyf;
                    const float power = pair_power(co.x, co.y, co.z, dx, dy);
                    const float G = score_exp(power);
                    const float alpha = fminf(0.99f, co.w * G);
                    const bool ok = __float_as_int(a.z) < last && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                    if (!__any_sync(0xffffffffu, ok)) continue;          // a quarter of the entries: nobody contributes
                    const float4 r0 = q.rb[0][slot], r1 = q.rb[1][slot], r2 = q.rb[2][slot], r3 = q.rb[3][slot];
                    const float inv1ma = score_rcp(1.f - alpha);
                    const float Tn = T * inv1ma;
                    const float csum = (r0.x + r0.y) + r0.z;
                    const float ar = last_alpha * last_color[0] + (1.f - last_alpha) * accum_rec[0];
                    float dL_dalpha = (csum - ar) * dpix * Tn;
                    if (bgdot != 0.f) dL_dalpha += (-T_final * inv1ma) * bgdot;
                    const float sv = ok ? G * (co.w * dL_dalpha) : 0.f;
                    const float c[13] = {r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w, r3.x, r3.y, r3.z, r3.w};
                    score += (sv * sv) * eval_poly13(c, dx, dy);
                    T = ok ? Tn : T;
                    accum_rec[0] = ok ? ar : accum_rec[0];
                    last_color[0] = ok ? csum : last_color[0];
                    last_alpha = ok ? alpha : last_alpha;
                }
#else
                for (int h = 0; h < cnt; ++h) {
                    const int slot = (qh + h) & (FB_WQ - 1);
                    const float4 a = q.a[slot];
                    if (__float_as_int(a.z) >= last) continue;
                    const float4 co = q.b[slot];
                    const float dx = a.x - pxf, dy = a.y - pyf;
                    const float power = pair_power(co.x, co.y, co.z, dx, dy);
                    if (power > 0.0f) continue;
                    const float G = score_exp(power);
                    const float alpha = fminf(0.99f, co.w * G);
                    if (alpha < 1.0f / 255.0f) continue;
                    const float4 r0 = q.rb[0][slot];
                    T = T / (1.f - alpha);
                    const float col[3] = {r0.x, r0.y, r0.z};
                    float dL_dalpha = 0.0f;
#pragma unroll
                    for (int ch = 0; ch < 3; ++ch) {
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = col[ch];
                        dL_dalpha += (col[ch] - accum_rec[ch]) * dpix;
                    }
                    dL_dalpha *= T;
                    last_alpha = alpha;
                    if (bgdot != 0.f) dL_dalpha += (-T_final / (1.f - alpha)) * bgdot;   // exact: x + (-0) == x
                    const float s = G * (co.w * dL_dalpha);
                    const float4 r1 = q.rb[1][slot], r2 = q.rb[2][slot], r3 = q.rb[3][slot];
                    const float c[13] = {r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w, r3.x, r3.y, r3.z, r3.w};
                    score += (s * s) * eval_poly13(c, dx, dy);
                }
#endif
                __syncwarp();
                qn -= cnt; qh = (qh + cnt) & (FB_WQ - 1);
            }
        }
    }

    // (3) warp reduce; k_score_finalize adds the (tile, block) partials in fixed order
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) score += __shfl_xor_sync(0xffffffffu, score, o);
    if (lane == 0) *out = score;
}

// ---- single-pass form -------------------------------------------------------------------------
// The reference needs two sweeps of a pixel's list: forward for T_final / n_contrib, then backward for
//   dL/dalpha_i = T_i (c_i - S_i) - T_final bg / (1 - alpha_i),   S_i = colour accumulated BEHIND Gaussian i
// (backward.cu:978-1017).  With C = sum_j c_j alpha_j T_j the colour of the pixel and Cp_i the same sum over j <= i,
//   S_i = (C - Cp_i) / (T_i (1 - alpha_i))   =>   dL/dalpha_i = A_i - (C + T_final bg) B_i,
//   A_i = T_i c_i + Cp_i / (1 - alpha_i),   B_i = 1 / (1 - alpha_i),
// and A_i, B_i are known when Gaussian i is visited FRONT TO BACK; only the scalar C' = C + T_final bg is not.  The
// score of the pixel, sum_i (G_i o_i dL/dalpha_i)^2 poly_i, is a quadratic in C':
//   Q_AA - 2 C' Q_AB + C'^2 Q_BB,   Q_XY = sum_i g_i X_i Y_i,   g_i = (G_i o_i)^2 poly_i >= 0,
// so one front-to-back sweep with three accumulators per pixel replaces both sweeps: the list is scanned and
// culled once, alpha is evaluated once per pair, T is the forward product T (1 - alpha) itself (no T / (1 - alpha)
// reconstruction), and n_contrib never exists.  c is r+g+b because dL/dpix is one scalar for the three channels
// (gaussian.py:1549).  The blend rules (skips, 0.99 clamp, stop BEFORE the Gaussian that would take T below 1e-4)
// are those of forward.cu:343-358, which are also the set backward.cu:962-975 revisits.
// Rounding: the three sums are of non-negative terms and the cancellation happens once per pixel; measured
// against the reference extension the scores agree to ~1e-5 (tests/test_gpu_scoring.py, rtol 1e-3 required).
__global__ void __launch_bounds__(32 * FB_SCORE_WPC) k_tile_score1(Bins bins, Cam cam, float grad_scale,
                                                                   float* __restrict__ partial) {
    extern __shared__ float4 smem[];
    constexpr int CPT = FB_TILE_WARPS / FB_SCORE_WPC;
    const unsigned lin = blockIdx.y * gridDim.x + blockIdx.x;
    const int tile_slot = (int)bins.tile_order[lin / CPT];
    const int t = tile_slot % bins.tiles, v = tile_slot / bins.tiles;
    const int tid = threadIdx.x, lane = tid & 31, wq = tid >> 5;
    const int warp = (int)(lin % CPT) * FB_SCORE_WPC + wq;
    const unsigned lt = (1u << lane) - 1u;
    const int tx = t % cam.tiles_x, ty = t / cam.tiles_x;
    constexpr int BPR = FB_TILE_X / FB_S1_WX;                       // pixel blocks per tile row
    const int bx = tx * FB_TILE_X + (warp % BPR) * FB_S1_WX, by = ty * FB_TILE_Y + (warp / BPR) * FB_S1_WY;
    const int px = bx + (lane % FB_S1_WX), py = by + (lane / FB_S1_WX);
    const bool inside = px < cam.W && py < cam.H;
    const float pxf = (float)px, pyf = (float)py;
    const float cx = (float)bx + 0.5f * (FB_S1_WX - 1), cy = (float)by + 0.5f * (FB_S1_WY - 1);
    const float ex = 0.5f * (FB_S1_WX - 1), ey = 0.5f * (FB_S1_WY - 1);
    float* out = partial + (size_t)tile_slot * FB_TILE_WARPS + warp;
    if (bins.total[1]) { if (lane == 0) *out = 0.f; return; }
    const uint2 range = bins.tile_range[tile_slot];
    const int n = (int)(range.y - range.x);
    if (n == 0) { if (lane == 0) *out = 0.f; return; }
    // warp-uniform 64-bit bases + 32-bit byte offsets (P <= 2^26 Gaussians): the gathers need no 64-bit address math
    const char* __restrict__ recs = reinterpret_cast<const char*>(bins.recs + (size_t)v * bins.P);
    const char* __restrict__ recsB = reinterpret_cast<const char*>(bins.recsB + (size_t)v * bins.P * 4);
    const unsigned* __restrict__ sorted = reinterpret_cast<const unsigned*>(bins.keys + range.x);   // low word = id
    WarpQ& q = reinterpret_cast<WarpQ*>(smem)[wq];

    float T = 1.0f, Cp = 0.f, Qaa = 0.f, Qab = 0.f, Qbb = 0.f;
#if FB_SCORE_STATS
    float stat = 0.f;                                                    // measurement builds only (scratch/lane_stats.py)
#endif
    bool done = !inside;
    if (!__all_sync(0xffffffffu, done)) {
        int qn = 0, qh = 0;
        const int nsteps = (n + 31) >> 5;
        bool v_nxt = lane < n;
        unsigned gid_nxt = v_nxt ? sorted[2 * lane] : 0u;
        float4 a_nxt = make_float4(0.f, 0.f, 0.f, 0.f);
        if (v_nxt) a_nxt = __ldg(reinterpret_cast<const float4*>(recs + (gid_nxt << 5)));
        unsigned key2 = (32 + lane < n) ? sorted[2 * (32 + lane)] : 0u;
        for (int step = 0; step <= nsteps; ++step) {
            if (step < nsteps) {
                const float4 a = a_nxt;
                const unsigned gid = gid_nxt;
                const bool valid = v_nxt;
                v_nxt = (step + 1) * 32 + lane < n;
                gid_nxt = key2;
                if (v_nxt) a_nxt = __ldg(reinterpret_cast<const float4*>(recs + (gid_nxt << 5)));
                key2 = ((step + 2) * 32 + lane < n) ? sorted[2 * ((step + 2) * 32 + lane)] : 0u;
                const bool hit = valid && (fabsf(a.x - cx) <= a.z + ex) && (fabsf(a.y - cy) <= a.w + ey);
                const unsigned mask = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    const int slot = (qh + qn + __popc(mask & lt)) & (FB_WQ - 1);
                    q.a[slot] = a;
                    __pipeline_memcpy_async(&q.b[slot], reinterpret_cast<const float4*>(recs + (gid << 5)) + 1, 16);
                    const float4* rb = reinterpret_cast<const float4*>(recsB + (gid << 6));
#pragma unroll
                    for (int k = 0; k < 4; ++k) __pipeline_memcpy_async(&q.rb[k][slot], rb + k, 16);
                }
                __pipeline_commit();
                qn += __popc(mask);
            }
            const bool flush = (step == nsteps);
            if (qn >= 32 || (flush && qn > 0)) {
                const int cnt = min(qn, 32);
                __pipeline_wait_prior(0);
                __syncwarp();
                // batches are 32 entries except the last one, so qh is 0 or 32 and qh + h never wraps
                auto hit = [&](int slot) {
                    const float4 a = q.a[slot];
                    const float4 co = q.b[slot];
                    const float dx = a.x - pxf, dy = a.y - pyf;
#if FB_SCORE_PRESCALE
                    // conic pre-multiplied in k_preprocess_score: power * log2(e) = dx (A dx + B dy) + C dy^2
                    const float power = fmaf(co.z, dy * dy, dx * fmaf(co.y, dy, co.x * dx));
                    float G;
                    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(G) : "f"(power));
#else
                    const float power = pair_power(co.x, co.y, co.z, dx, dy);
                    const float G = score_exp(power);
#endif
                    const float alpha = fminf(0.99f, co.w * G);
                    const float oma = __fsub_rn(1.f, alpha);
                    const float test_T = __fmul_rn(T, oma);
                    const bool ok = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                    const bool upd = ok && !(test_T < 0.0001f);
                    done = done || (ok && !upd);                         // stop BEFORE the Gaussian that takes T below 1e-4
#if FB_SCORE_STATS == 1
                    stat += 1.f;                                         // (lane, survivor) visits
#elif FB_SCORE_STATS == 3
                    stat += upd ? 1.f : 0.f;                             // lanes that take the survivor
#endif
                    if (!__any_sync(0xffffffffu, upd)) return;           // ~30 % of the entries: nobody contributes
#if FB_SCORE_STATS == 2
                    stat += 1.f;                                         // (lane, survivor) visits that run the full body
#endif
                    const float4 r0 = q.rb[0][slot], r1 = q.rb[1][slot], r2 = q.rb[2][slot], r3 = q.rb[3][slot];
                    const float inv = score_rcp(oma);
#if FB_SCORE_PRESCALE
                    const float csum = r0.x;
#else
                    const float csum = (r0.x + r0.y) + r0.z;
#endif
                    const float Cn = fmaf(csum, alpha * T, Cp);
                    const float A = fmaf(T, csum, Cn * inv);
#if FB_SCORE_PRESCALE
                    const float c[13] = {r0.y, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w, r3.x, r3.y, r3.z, r3.w};
#else
                    const float c[13] = {r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w, r3.x, r3.y, r3.z, r3.w};
#endif
                    const float go = upd ? G * co.w : 0.f;               // select BEFORE the polynomial: no branch around it
                    const float g = (go * go) * eval_poly13(c, dx, dy);
                    const float gA = g * A, gB = g * inv;
                    Qaa = fmaf(gA, A, Qaa);
                    Qab = fmaf(gA, inv, Qab);
                    Qbb = fmaf(gB, inv, Qbb);
                    Cp = upd ? Cn : Cp;
                    T = upd ? test_T : T;
                };
                if (cnt == 32) {
#pragma unroll kS1Unroll
                    for (int h = 0; h < 32; ++h) hit(qh + h);
                } else {
                    for (int h = 0; h < cnt; ++h) hit(qh + h);
                }
                __syncwarp();
                qn -= cnt; qh = (qh + cnt) & (FB_WQ - 1);
                if (__all_sync(0xffffffffu, done)) break;
            }
        }
        __pipeline_wait_prior(0);
        __syncwarp();
    }
    const float dpix = inside ? grad_scale : 0.f;
    const float Cf = fmaf(T, (cam.bg[0] + cam.bg[1]) + cam.bg[2], Cp);     // C + T_final * bg
    float score = (dpix * dpix) * fmaf(Cf, fmaf(Cf, Qbb, -2.f * Qab), Qaa);
#if FB_SCORE_STATS
    score = stat;
#endif
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) score += __shfl_xor_sync(0xffffffffu, score, o);
    if (lane == 0) *out = score;
}

// ---- per-Gaussian Hessian moments (compute_H_train), warp-autonomous form ----------------------------------
// Same traversal as k_tile_score (two sweeps, the reference's arithmetic: expf, IEEE division, the recurrences of
// backward.cu:978-1017 with r+g+b as the one colour channel), but what a contributing pair produces is the 13
// monomial moments  s^2 * dx^i dy^j  of fb_fisher_tile.cuh (power = 2).  They are summed over the 32 pixels of the
// block with the halving butterfly (16 SHFL), four neighbouring sums are gathered into one lane and added to the
// Gaussian's 16-float row with ONE red.global.add.v4.f32 per quad: 4 vector atomics per (Gaussian, pixel block).
// k_expand_H turns the rows into H columns.  Replaces k_render_fwd + k_bwd_tiles on the batched H_train path (every
// pixel tests every instance, block barriers per 256 instances): 170 -> ~60 us per keyframe at 500k Gaussians.
__global__ void __launch_bounds__(32 * FB_SCORE_WPC) k_tile_hessian(Bins bins, Cam cam, const float* __restrict__ colors,
                                                                    float grad_scale, float* __restrict__ mom_all) {
    extern __shared__ float4 smem[];
    constexpr int CPT = FB_TILE_WARPS / FB_SCORE_WPC;
    const unsigned lin = blockIdx.y * gridDim.x + blockIdx.x;
    const int tile_slot = (int)bins.tile_order[lin / CPT];
    const int t = tile_slot % bins.tiles, v = tile_slot / bins.tiles;
    const int tid = threadIdx.x, lane = tid & 31, wq = tid >> 5;
    const int warp = (int)(lin % CPT) * FB_SCORE_WPC + wq;
    const unsigned lt = (1u << lane) - 1u;
    const int tx = t % cam.tiles_x, ty = t / cam.tiles_x;
    const int bx = tx * FB_TILE_X + (warp & 1) * FB_WB_X, by = ty * FB_TILE_Y + (warp >> 1) * FB_WB_Y;
    const int px = bx + (lane & 7), py = by + (lane >> 3);
    const bool inside = px < cam.W && py < cam.H;
    const float pxf = (float)px, pyf = (float)py;
    const float cx = (float)bx + 0.5f * (FB_WB_X - 1), cy = (float)by + 0.5f * (FB_WB_Y - 1);
    const float ex = 0.5f * (FB_WB_X - 1), ey = 0.5f * (FB_WB_Y - 1);
    if (bins.total[1]) return;
    const uint2 range = bins.tile_range[tile_slot];
    const int n = (int)(range.y - range.x);
    if (n == 0) return;
    const GRec* __restrict__ recs = bins.recs + (size_t)v * bins.P;
    float* __restrict__ mom = mom_all + (size_t)v * bins.P * FB_NACC;
    const u64* __restrict__ sorted = bins.keys + range.x;
    WarpQ& q = reinterpret_cast<WarpQ*>(smem)[wq];

    // (1) transmittance sweep (forward.cu:330-379 without colour / depth)
    float T = 1.0f;
    int last = 0;
    bool done = !inside;
    if (!__all_sync(0xffffffffu, done)) {
        int qn = 0, qh = 0;
        const int nsteps = (n + 31) >> 5;
        bool v_nxt = lane < n;
        unsigned gid_nxt = v_nxt ? (unsigned)(sorted[lane] & 0xffffffffull) : 0u;
        float4 a_nxt = make_float4(0.f, 0.f, 0.f, 0.f);
        if (v_nxt) a_nxt = __ldg(reinterpret_cast<const float4*>(recs + gid_nxt));
        u64 key2 = (32 + lane < n) ? sorted[32 + lane] : 0ull;
        for (int step = 0; step <= nsteps; ++step) {
            if (step < nsteps) {
                const float4 a = a_nxt;
                const unsigned gid = gid_nxt;
                const bool valid = v_nxt;
                v_nxt = (step + 1) * 32 + lane < n;
                gid_nxt = (unsigned)(key2 & 0xffffffffull);
                if (v_nxt) a_nxt = __ldg(reinterpret_cast<const float4*>(recs + gid_nxt));
                key2 = ((step + 2) * 32 + lane < n) ? sorted[(step + 2) * 32 + lane] : 0ull;
                const bool hit = valid && (fabsf(a.x - cx) <= a.z + ex) && (fabsf(a.y - cy) <= a.w + ey);
                const unsigned mask = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    const int slot = (qh + qn + __popc(mask & lt)) & (FB_WQ - 1);
                    q.a[slot] = make_float4(a.x, a.y, __int_as_float(step * 32 + lane), 0.f);
                    __pipeline_memcpy_async(&q.b[slot], reinterpret_cast<const float4*>(recs + gid) + 1, 16);
                }
                __pipeline_commit();
                qn += __popc(mask);
            }
            const bool flush = (step == nsteps);
            if (qn >= 32 || (flush && qn > 0)) {
                const int cnt = min(qn, 32);
                __pipeline_wait_prior(0);
                __syncwarp();
                for (int h = 0; h < cnt; ++h) {
                    const int slot = qh + h;
                    const float4 a = q.a[slot];
                    const float4 co = q.b[slot];
                    const float dx = a.x - pxf, dy = a.y - pyf;
                    const float power = pair_power(co.x, co.y, co.z, dx, dy);
                    const float alpha = fminf(0.99f, co.w * hess_exp(power));
                    const float test_T = __fmul_rn(T, __fsub_rn(1.f, alpha));
                    const bool ok = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                    const bool term = ok && test_T < 0.0001f;
                    done = done || term;
                    const bool upd = ok && !term;
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

    // (2) back-to-front sweep (backward.cu:939-1038), moments instead of gradients
    const int warp_last = __reduce_max_sync(0xffffffffu, last);
    if (warp_last == 0) return;
    const float T_final = T;
    const float dpix = inside ? grad_scale : 0.f;
    const float bgdot = ((cam.bg[0] + cam.bg[1]) + cam.bg[2]) * dpix;
    float accum_rec = 0.f, last_color = 0.f, last_alpha = 0.f;
    const int rslot = reduce16_slot(lane);
    {
        int qn = 0, qh = 0;
        const int nsteps = (warp_last + 31) >> 5;
        bool v_nxt = lane < warp_last;
        unsigned gid_nxt = v_nxt ? (unsigned)(sorted[warp_last - 1 - lane] & 0xffffffffull) : 0u;
        float4 a_nxt = make_float4(0.f, 0.f, 0.f, 0.f);
        if (v_nxt) a_nxt = __ldg(reinterpret_cast<const float4*>(recs + gid_nxt));
        u64 key2 = (32 + lane < warp_last) ? sorted[warp_last - 1 - 32 - lane] : 0ull;
        for (int step = 0; step <= nsteps; ++step) {
            if (step < nsteps) {
                const float4 a = a_nxt;
                const unsigned gid = gid_nxt;
                const bool valid = v_nxt;
                v_nxt = (step + 1) * 32 + lane < warp_last;
                gid_nxt = (unsigned)(key2 & 0xffffffffull);
                if (v_nxt) a_nxt = __ldg(reinterpret_cast<const float4*>(recs + gid_nxt));
                key2 = ((step + 2) * 32 + lane < warp_last) ? sorted[warp_last - 1 - (step + 2) * 32 - lane] : 0ull;
                const bool hit = valid && (fabsf(a.x - cx) <= a.z + ex) && (fabsf(a.y - cy) <= a.w + ey);
                const unsigned mask = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    const int slot = (qh + qn + __popc(mask & lt)) & (FB_WQ - 1);
                    q.a[slot] = make_float4(a.x, a.y, __int_as_float(warp_last - 1 - (step * 32 + lane)), __uint_as_float(gid));
                    __pipeline_memcpy_async(&q.b[slot], reinterpret_cast<const float4*>(recs + gid) + 1, 16);
                    const float* c = colors + 3 * (size_t)gid;
                    float* dst = reinterpret_cast<float*>(&q.rb[0][slot]);
#pragma unroll
                    for (int k = 0; k < 3; ++k) __pipeline_memcpy_async(dst + k, c + k, 4);
                }
                __pipeline_commit();
                qn += __popc(mask);
            }
            const bool flush = (step == nsteps);
            if (qn >= 32 || (flush && qn > 0)) {
                const int cnt = min(qn, 32);
                __pipeline_wait_prior(0);
                __syncwarp();
                for (int h = 0; h < cnt; ++h) {
                    const int slot = qh + h;
                    const float4 a = q.a[slot];
                    const float4 co = q.b[slot];
                    const float dx = a.x - pxf, dy = a.y - pyf;
                    const float power = pair_power(co.x, co.y, co.z, dx, dy);
                    const float G = hess_exp(power);
                    const float alpha = fminf(0.99f, co.w * G);
                    const bool ok = __float_as_int(a.z) < last && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                    if (!__any_sync(0xffffffffu, ok)) continue;
                    const float4 c3 = q.rb[0][slot];
                    const float Tn = hess_div(T, 1.f - alpha);
                    const float csum = (c3.x + c3.y) + c3.z;
                    const float ar = last_alpha * last_color + (1.f - last_alpha) * accum_rec;
                    float dL_dalpha = (csum - ar) * dpix * Tn;
                    if (bgdot != 0.f) dL_dalpha += (-T_final / (1.f - alpha)) * bgdot;
                    const float sv = ok ? G * (co.w * dL_dalpha) : 0.f;
                    T = ok ? Tn : T;
                    accum_rec = ok ? ar : accum_rec;
                    last_color = ok ? csum : last_color;
                    last_alpha = ok ? alpha : last_alpha;
                    float m[FB_NACC];
                    {
                        const float qq = sv * sv;
                        const float qx = qq * dx, qy = qq * dy;
                        const float m20 = qx * dx, m11 = qx * dy, m02 = qy * dy;
                        const float m30 = m20 * dx, m21 = m20 * dy, m12 = m02 * dx, m03 = m02 * dy;
                        m[0] = qq; m[1] = m20; m[2] = m11; m[3] = m02;
                        m[4] = m30; m[5] = m21; m[6] = m12; m[7] = m03;
                        m[8] = m30 * dx; m[9] = m30 * dy; m[10] = m21 * dy; m[11] = m03 * dx; m[12] = m03 * dy;
                        m[13] = 0.f; m[14] = 0.f; m[15] = 0.f;
                    }
                    const float tot = warp_reduce16(m, lane);        // even lane L holds the block sum of moment rslot
                    // quad k = moments 4k..4k+3 sits on lanes 8k, 8k+2, 8k+4, 8k+6: gather it on lane 8k
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
}

// scores[k0 + v] = sum over (tile, pixel block) of partial[v][t][w], fixed order (deterministic given partials).
__global__ void k_score_finalize(const float* __restrict__ partial, int tiles, float* __restrict__ scores) {
    const int v = blockIdx.x;
    float s = 0.f;
    for (int t = threadIdx.x; t < tiles; t += 32) s += partial[v * tiles + t];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (threadIdx.x == 0) scores[v] = s;
}

}  // namespace fb
