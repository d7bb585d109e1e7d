// fb_score.cuh -- fused per-(view, tile) candidate-scoring kernels and the H_train moment kernel.
//
// A warp owns one pixel block of one 16x16 tile of one candidate view; the tile's depth-sorted key segment was
// produced by k_scatter_keys + k_sort_tiles.  Without leaving the SM it makes ONE front-to-back sweep that carries
// the transmittance and three accumulators per pixel (see "single sweep" below) and reduces them to one float per
// (view, tile, pixel block); k_score_finalize sums those in fixed order.  No image, no final_T / n_contrib round
// trip, no per-Gaussian Hessian tensor, no atomics.
// Replaces the K-iteration Python loop  models/SLAM/gaussian.py:1362-1370  (+ compute_Hessian
// :1503-1570) around renderCUDA / renderCUDAFused (forward.cu:261-393, backward.cu:850-1140).
//
// Pair culling.  The reference tests every pixel of a tile against every Gaussian binned to it and
// rejects most pairs with `power > 0` / `alpha < 1/255`.  Here each warp first tests 32 list entries at once (one per
// lane) against its pixel block with the precomputed half extents of the alpha >= 1/255 footprint (recA.zw, see
// k_preprocess_score); a ballot gives the short, still depth-ordered list of Gaussians that can touch the block, and
// only those run the pair arithmetic.  A culled pair is one the reference's own alpha test skips, and skipped pairs
// change no state, so every pixel's result is unchanged.
#pragma once
#include "fb_stage12.cuh"
#include <cuda_pipeline.h>

namespace fb {

#ifndef FB_HESS_PFD
#define FB_HESS_PFD 1  // the same for k_tile_hessian; 4 keyframes: 1 / 2 / 4 / 8 ahead -> 0.60 / 0.62 / 0.58 / 0.58 ms per H_train call, 16 keyframes: 1.52 / 1.57 / 1.61 / 1.89
#endif
#ifndef FB_S2_PFD
#define FB_S2_PFD 1    // scan steps the record gather of the pair sweeps runs ahead of its use
#endif
#define FB_WB_X 8      // warp pixel block of the H_train moment kernel
#define FB_WB_Y 4
// pixel block of k_tile_score1 (32 pixels, one per lane: 8x4 default; 4x8 and 16x2 measured)
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
// record, both prefetched one / two steps ahead), culls them against its pixel block
// and appends the survivors, in order, to a 64-entry ring in shared memory that is private to the
// warp.  The appending lane also starts a cp.async of the 80-byte scoring record (conic+opacity, colour, 13
// coefficients) into the ring slot, so by the time 32 survivors have
// queued up their data has landed; the warp then runs the pair arithmetic for those 32 entries with
// broadcast shared-memory reads.  Warps of a tile never wait for each other (pixel blocks differ a lot
// in how many Gaussians touch them), and a warp whose pixels are saturated stops scanning.
#define FB_WQ 64   // ring entries per warp
struct WarpQ {
    float4 a[FB_WQ];       // px, py, hx, hy
    float4 b[FB_WQ];       // conic a, b, c (pre-multiplied by -log2(e)/2, -log2(e), -log2(e)/2), opacity
    float4 rb[4][FB_WQ];   // r+g+b, coef0, -, - | coef1-4 | coef5-8 | coef9-12
};
// Warps per CTA.  A warp never talks to another warp, so the CTA is only a launch unit: small CTAs let the
// SM backfill as soon as a warp's pixel block is finished (with 8-warp CTAs 23 % of all warp samples sat in
// the final barrier waiting for the slowest block of the tile; profiles/r1_summary_v4.md).
#ifndef FB_SCORE_WPC
#define FB_SCORE_WPC 4
#endif
#define FB_TILE_WARPS (FB_TILE_PIX / 32)
// The scoring path evaluates exp with MUFU.EX2 on an exponent that arrives pre-scaled by log2(e) (|x| <= 8 here, <= 2
// ulp), 1/(1-alpha) with MUFU.RCP, and runs the colour recurrence of backward.cu:997-1008 once on r+g+b (dL/dpix is
// the same scalar for the three channels in compute_Hessian, gaussian.py:1549, and the recurrence is linear).  These
// differ from the reference's compiled arithmetic by a few ulp per pair; scores stay far inside the stated rtol 1e-3
// (tests/test_gpu_scoring.py).  The drop-in rasteriser kernels keep the as-compiled arithmetic.
__device__ __forceinline__ float score_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));                          // MUFU.RCP
    return r;
}
__device__ __forceinline__ float score_ex2(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));                          // MUFU.EX2
    return r;
}
// FB_HESS_FAST: k_tile_hessian (H_train path) with the scoring kernel's MUFU exp / reciprocal instead of expf and the
// IEEE division of backward.cu:987 (both sweeps use the same exp, so their alpha decisions stay consistent); the H_train
// path is held to rtol 1e-3 (tests/test_gpu_scoring.py), the MUFU forms are ~1e-6.
#ifndef FB_HESS_FAST
#define FB_HESS_FAST 1
#endif
__device__ __forceinline__ float hess_exp(float x) {
#if FB_HESS_FAST
    return score_ex2(x * 1.4426950408889634f);
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

// ---- single sweep ----------------------------------------------------------------------------------
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

// sweep_px1: one pixel per lane, 8x4 block per warp.  D = scan steps the record gather runs ahead of its use.
template <int D>
__device__ __forceinline__ void sweep_px1(const Bins& bins, const Cam& cam, float grad_scale, int tile_slot, int warp,
                                          WarpQ& q, float* __restrict__ out) {
    const int t = tile_slot % bins.tiles, v = tile_slot / bins.tiles;
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int tx = t % cam.tiles_x, ty = t / cam.tiles_x;
    constexpr int BPR = FB_TILE_X / FB_S1_WX;                       // pixel blocks per tile row
    const int bx = tx * FB_TILE_X + (warp % BPR) * FB_S1_WX, by = ty * FB_TILE_Y + (warp / BPR) * FB_S1_WY;
    const int px = bx + (lane % FB_S1_WX), py = by + (lane / FB_S1_WX);
    const bool inside = px < cam.W && py < cam.H;
    const float pxf = (float)px, pyf = (float)py;
    const float cx = (float)bx + 0.5f * (FB_S1_WX - 1), cy = (float)by + 0.5f * (FB_S1_WY - 1);
    const float ex = 0.5f * (FB_S1_WX - 1), ey = 0.5f * (FB_S1_WY - 1);
    if (bins.total[1]) { if (lane == 0) *out = 0.f; return; }
    const uint2 range = bins.tile_range[tile_slot];
    const int n = (int)(range.y - range.x);
    if (n == 0) { if (lane == 0) *out = 0.f; return; }
    // warp-uniform 64-bit bases + 32-bit byte offsets (80 B x visible capacity < 4 GB per view, checked by the host
    // entry point): the gathers need no 64-bit address math
    const char* __restrict__ recA = reinterpret_cast<const char*>(bins.recA + (size_t)v * bins.vis_stride);
    const char* __restrict__ recB = reinterpret_cast<const char*>(bins.recB + (size_t)v * bins.vis_stride * 5);
    const unsigned* __restrict__ sorted = reinterpret_cast<const unsigned*>(bins.keys + range.x);   // low word = visible slot

    float T = 1.0f, Cp = 0.f, Qaa = 0.f, Qab = 0.f, Qbb = 0.f;
    bool done = !inside;
    if (!__all_sync(0xffffffffu, done)) {
        int qn = 0, qh = 0;
        const int nsteps = (n + 31) >> 5;
        // the 16-byte {px, py, hx, hy} gathers run D scan steps ahead of their use, the keys one more
        bool vq[D]; unsigned gq[D]; float4 aq[D];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            vq[d] = d * 32 + lane < n;
            gq[d] = vq[d] ? sorted[2 * (d * 32 + lane)] : 0u;
            aq[d] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (vq[d]) aq[d] = __ldg(reinterpret_cast<const float4*>(recA + (gq[d] << 4)));
        }
        unsigned key2 = (D * 32 + lane < n) ? sorted[2 * (D * 32 + lane)] : 0u;
        for (int step = 0; step <= nsteps; ++step) {
            if (step < nsteps) {
                const float4 a = aq[0];
                const unsigned gid = gq[0];
                const bool valid = vq[0];
#pragma unroll
                for (int d = 0; d + 1 < D; ++d) { vq[d] = vq[d + 1]; gq[d] = gq[d + 1]; aq[d] = aq[d + 1]; }
                vq[D - 1] = (step + D) * 32 + lane < n;
                gq[D - 1] = key2;
                if (vq[D - 1]) aq[D - 1] = __ldg(reinterpret_cast<const float4*>(recA + (key2 << 4)));
                key2 = ((step + D + 1) * 32 + lane < n) ? sorted[2 * ((step + D + 1) * 32 + lane)] : 0u;
                const bool hit = valid && (fabsf(a.x - cx) <= a.z + ex) && (fabsf(a.y - cy) <= a.w + ey);
                const unsigned mask = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    const int slot = (qh + qn + __popc(mask & lt)) & (FB_WQ - 1);
                    q.a[slot] = a;
                    const float4* rb = reinterpret_cast<const float4*>(recB + gid * 80u);
                    __pipeline_memcpy_async(&q.b[slot], rb, 16);
#pragma unroll
                    for (int k = 0; k < 4; ++k) __pipeline_memcpy_async(&q.rb[k][slot], rb + 1 + k, 16);
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
                    // conic pre-multiplied in k_preprocess_score: power * log2(e) = dx (A dx + B dy) + C dy^2
                    const float power = fmaf(co.z, dy * dy, dx * fmaf(co.y, dy, co.x * dx));
                    const float G = score_ex2(power);
                    const float alpha = fminf(0.99f, co.w * G);
                    const float oma = __fsub_rn(1.f, alpha);
                    const float test_T = __fmul_rn(T, oma);
                    const bool ok = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                    const bool upd = ok && !(test_T < 0.0001f);
                    done = done || (ok && !upd);                         // stop BEFORE the Gaussian that takes T below 1e-4
                    if (!__any_sync(0xffffffffu, upd)) return;           // ~30 % of the entries: nobody contributes
                    const float4 r0 = q.rb[0][slot], r1 = q.rb[1][slot], r2 = q.rb[2][slot], r3 = q.rb[3][slot];
                    const float inv = score_rcp(oma);
                    const float csum = r0.x;
                    const float Cn = fmaf(csum, alpha * T, Cp);
                    const float A = fmaf(T, csum, Cn * inv);
                    const float c[13] = {r0.y, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w, r3.x, r3.y, r3.z, r3.w};
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
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) score += __shfl_xor_sync(0xffffffffu, score, o);
    if (lane == 0) *out = score;
}

// k_tile_score1: the 8x4 form alone (the round-1 kernel on the slot-indexed records; built with -DFB_SCORE_PX2=0 as
// the independent form the parity tests can be run against).
__global__ void __launch_bounds__(32 * FB_SCORE_WPC) k_tile_score1(Bins bins, Cam cam, float grad_scale,
                                                                   float* __restrict__ partial) {
    extern __shared__ float4 smem[];
    constexpr int CPT = FB_TILE_WARPS / FB_SCORE_WPC;
    const unsigned lin = blockIdx.y * gridDim.x + blockIdx.x;
    const int tile_slot = (int)bins.tile_order[lin / CPT];
    const int wq = threadIdx.x >> 5;
    const int warp = (int)(lin % CPT) * FB_SCORE_WPC + wq;
    sweep_px1<FB_S2_PFD>(bins, cam, grad_scale, tile_slot, warp, reinterpret_cast<WarpQ*>(smem)[wq],
                         partial + (size_t)tile_slot * FB_TILE_WARPS + warp);
}

// ---- k_tile_score2: two pixels per lane, packed FP32 ------------------------------------------------
// The pair loop is bound by instruction issue, not by a pipe or by memory (round 1: 79 % of the issue slots, fma 42 %,
// alu 42 %).  This form issues fewer instructions per (pixel, Gaussian) pair:
//   * a warp owns an 8x8 pixel block and a lane a COLUMN of two pixels, (x, y) and (x, y + 4).  Everything that
//     depends on dx alone is computed once per lane and shared by its two pixels: the exponent's A dx^2 and B dx, and
//     the quartic written as a polynomial in dy whose five coefficients are polynomials in dx
//         poly = P0(dx) + dy (P1(dx) + dy (P2(dx) + dy (P3(dx) + dy c12)));
//   * what differs between the two pixels runs as packed FP32 (fma.rn.f32x2 -> FFMA2: one issue slot for both pixels;
//     measured on B200: the same FLOP/s as FFMA at half the issue slots; per-Gaussian scalars enter as broadcast
//     operands, no register duplication);
//   * half as many warps scan a tile's list (4 blocks of 8x8 instead of 8 of 8x4);
//   * each batch of 32 ring entries is first tested EXACTLY, one entry per lane: the maximum of the (concave) exponent
//     over the block's pixel-centre rectangle is compared with the alpha >= 1/255 threshold, so entries whose bounding
//     box touches the block but whose ellipse does not (a third of them) never reach the pair loop.
// Per surviving entry: 64 pixel-visits in ~66 issue slots (k_tile_score1: 32 in 54).
typedef float2 f32x2;
__device__ __forceinline__ f32x2 f2_pack(float lo, float hi) { return make_float2(lo, hi); }
__device__ __forceinline__ f32x2 f2_dup(float s) { return make_float2(s, s); }
__device__ __forceinline__ float f2_lo(f32x2 v) { return v.x; }
__device__ __forceinline__ float f2_hi(f32x2 v) { return v.y; }
#ifndef FB_S2_PACKED
#define FB_S2_PACKED 1            // 0: the same kernel with scalar FP32 (A/B measurement of FFMA2)
#endif
#if FB_S2_PACKED
__device__ __forceinline__ f32x2 f2_fma(f32x2 a, f32x2 b, f32x2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ f32x2 f2_mul(f32x2 a, f32x2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ f32x2 f2_add(f32x2 a, f32x2 b) { return __fadd2_rn(a, b); }
#else
__device__ __forceinline__ f32x2 f2_fma(f32x2 a, f32x2 b, f32x2 c) { return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)); }
__device__ __forceinline__ f32x2 f2_mul(f32x2 a, f32x2 b) { return make_float2(a.x * b.x, a.y * b.y); }
__device__ __forceinline__ f32x2 f2_add(f32x2 a, f32x2 b) { return make_float2(a.x + b.x, a.y + b.y); }
#endif
__device__ __forceinline__ void f2_acc(f32x2& acc, f32x2 a, f32x2 b) { acc = f2_fma(a, b, acc); }
#ifndef FB_S2_EXACT
#define FB_S2_EXACT 1             // exact ellipse / block test of every ring batch
#endif
#ifndef FB_S2_ANY
#define FB_S2_ANY 1               // skip the rest of the body when no pixel of the block takes the entry
#endif
#ifndef FB_S2_STATS
#define FB_S2_STATS 0             // measurement builds (tools/lane_stats.py): the "score" of a view becomes a count --
#endif                            // 1 scan steps, 2 ring entries, 3 entries after the exact test, 4 full bodies, 5 contributing pixel pairs
#define FB_S2_BLOCKS 4            // 8x8 pixel blocks per 16x16 tile
#ifndef FB_S2_WPC
#define FB_S2_WPC 4               // warps per CTA
#endif
#ifndef FB_S2_MINB
#define FB_S2_MINB 8              // minimum resident CTAs per SM asked of ptxas (register budget)
#endif
#ifndef FB_S2_LIST
#define FB_S2_LIST 1              // 1: surviving ring slots are compacted into a byte list (no BREV/FLO in the loop)
#endif
struct WarpQ2 { WarpQ q; unsigned char list[32]; unsigned char pad[96]; };
#ifndef FB_S2_SMEM_EXTRA
#define FB_S2_SMEM_EXTRA 0        // extra dynamic shared memory per CTA: caps the resident CTAs per SM so that the other chunk
#endif                            // lane's latency-bound kernels (scatter, sort) find registers and warp slots beside the sweep
#define FB_S2_SMEM_BYTES ((int)sizeof(WarpQ2) * FB_S2_WPC + FB_S2_SMEM_EXTRA)

struct Px2State { f32x2 T, Cp, Qaa, Qab, Qbb; float athr0, athr1; };   // athr: 1/255 while the pixel is live, 2 once done / outside

// The pair sweep of one 8x8 pixel block (corner bx, by) over n depth-sorted keys, continuing from the state S.
// MODE 0: the scoring sweep (stop rule, three accumulators).  MODE 1: transmittance and colour only and WITHOUT the
// stop rule -- the plain products a later segment of the same list needs as its starting state (sweep_px2_segment).
template <int D, int MODE>
__device__ __forceinline__ void sweep_px2_core(const char* __restrict__ recA, const char* __restrict__ recB,
                                               const unsigned* __restrict__ sorted, const int n, const int bx, const int by,
                                               WarpQ2& q2, Px2State& S, float& stat) {
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const float pxf = (float)(bx + (lane & 7));
    const f32x2 npy = f2_pack(-(float)(by + (lane >> 3)), -(float)(by + (lane >> 3) + 4));
    const float cx = (float)bx + 3.5f, cy = (float)by + 3.5f;
    WarpQ& q = q2.q;
    f32x2 T = S.T, Cp = S.Cp, Qaa = S.Qaa, Qab = S.Qab, Qbb = S.Qbb;
    float athr0 = S.athr0, athr1 = S.athr1;
    (void)stat;
    {
        int qn = 0, qh = 0;
        const int nsteps = (n + 31) >> 5;
        // the 16-byte {px, py, hx, hy} gathers run D scan steps ahead of their use, the keys one more
        bool vq[D]; unsigned gq[D]; float4 aq[D];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            vq[d] = d * 32 + lane < n;
            gq[d] = vq[d] ? sorted[2 * (d * 32 + lane)] : 0u;
            aq[d] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (vq[d]) aq[d] = __ldg(reinterpret_cast<const float4*>(recA + (gq[d] << 4)));
        }
        unsigned key2 = (D * 32 + lane < n) ? sorted[2 * (D * 32 + lane)] : 0u;
        for (int step = 0; step <= nsteps; ++step) {
            if (step < nsteps) {
                const float4 a = aq[0];
                const unsigned gid = gq[0];
                const bool valid = vq[0];
#pragma unroll
                for (int d = 0; d + 1 < D; ++d) { vq[d] = vq[d + 1]; gq[d] = gq[d + 1]; aq[d] = aq[d + 1]; }
                vq[D - 1] = (step + D) * 32 + lane < n;
                gq[D - 1] = key2;
                if (vq[D - 1]) aq[D - 1] = __ldg(reinterpret_cast<const float4*>(recA + (key2 << 4)));
                key2 = ((step + D + 1) * 32 + lane < n) ? sorted[2 * ((step + D + 1) * 32 + lane)] : 0u;
                const bool hit = valid && (fabsf(a.x - cx) <= a.z + 3.5f) && (fabsf(a.y - cy) <= a.w + 3.5f);
                const unsigned mask = __ballot_sync(0xffffffffu, hit);
#if FB_S2_STATS == 1
                if (lane == 0) stat += 1.f;
#elif FB_S2_STATS == 2
                if (lane == 0) stat += (float)__popc(mask);
#endif
                if (hit) {
                    const int slot = (qh + qn + __popc(mask & lt)) & (FB_WQ - 1);
                    q.a[slot] = a;
                    const float4* rb = reinterpret_cast<const float4*>(recB + gid * 80u);
                    __pipeline_memcpy_async(&q.b[slot], rb, 16);
#pragma unroll
                    for (int k = 0; k < (MODE == 0 ? 4 : 1); ++k) __pipeline_memcpy_async(&q.rb[k][slot], rb + 1 + k, 16);
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
                unsigned todo = cnt == 32 ? 0xffffffffu : ((1u << cnt) - 1u);
#if FB_S2_EXACT
                {
                    // lane h tests ring entry qh + h: max over the block's pixel-centre rectangle of the exponent
                    // p(d) = A dx^2 + B dx dy + C dy^2 (A, C < 0, negative definite; d = mean - pixel) against
                    // alpha >= 1/255  <=>  p >= -log2(255 o).  0.02 in log2 units (1.4 % in alpha) of slack covers the
                    // approximate reciprocals and every rounding below.  The maximum sits at d = 0 when the mean is
                    // inside the rectangle, else on one of the four edges (a clamped 1-D parabola each).
                    const float4 a = q.a[qh + lane];
                    const float4 co = q.b[qh + lane];
                    const float X1 = a.x - (float)bx, X0 = X1 - 7.f, Y1 = a.y - (float)by, Y0 = Y1 - 7.f;
                    const float ia = score_rcp(co.x + co.x), ic = score_rcp(co.z + co.z);       // 1 / (2A), 1 / (2C)
                    auto pw = [&](float x, float y) { return fmaf(co.z, y * y, x * fmaf(co.y, y, co.x * x)); };
                    auto ex_ = [&](float Y) { return pw(fminf(fmaxf(-co.y * Y * ia, X0), X1), Y); };
                    auto ey_ = [&](float X) { return pw(X, fminf(fmaxf(-co.y * X * ic, Y0), Y1)); };
                    float m = fmaxf(fmaxf(ex_(Y0), ex_(Y1)), fmaxf(ey_(X0), ey_(X1)));
                    if (X0 <= 0.f && X1 >= 0.f && Y0 <= 0.f && Y1 >= 0.f) m = 0.f;
                    float lg;
                    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(255.0f * co.w));
                    const bool keep = !(m < -lg - 0.02f);                 // NaN (degenerate conic) keeps the entry
                    todo &= __ballot_sync(0xffffffffu, keep);
                }
#endif
#if FB_S2_STATS == 3
                if (lane == 0) stat += (float)__popc(todo);
#endif
#if FB_S2_LIST
                if ((todo >> lane) & 1u) q2.list[__popc(todo & lt)] = (unsigned char)(qh + lane);
                __syncwarp();
                const int nkeep = __popc(todo);
                for (int i = 0; i < nkeep; ++i) {
                    const int slot = q2.list[i];
#else
                while (todo) {
                    const int h = __ffs(todo) - 1;
                    todo &= todo - 1u;
                    const int slot = qh + h;
#endif
                    const float2 mu = *reinterpret_cast<const float2*>(&q.a[slot]);
                    const float4 co = q.b[slot];
                    const float dx = mu.x - pxf;
                    const f32x2 dy = f2_add(f2_dup(mu.y), npy);
                    const float Adx = co.x * dx;
                    const float q0 = Adx * dx, q1 = co.y * dx;
                    const f32x2 pw = f2_fma(f2_fma(f2_dup(co.z), dy, f2_dup(q1)), dy, f2_dup(q0));
                    const float G0 = score_ex2(f2_lo(pw)), G1 = score_ex2(f2_hi(pw));
                    const f32x2 oG = f2_mul(f2_dup(co.w), f2_pack(G0, G1));
                    const float al0 = fminf(0.99f, f2_lo(oG)), al1 = fminf(0.99f, f2_hi(oG));
                    const f32x2 oma = f2_fma(f2_pack(al0, al1), f2_dup(-1.f), f2_dup(1.f));   // 1 - alpha, one rounding
                    const f32x2 tT = f2_mul(T, oma);
                    // (the exponent of a negative-definite form is <= 0 up to rounding: the reference's `power > 0`
                    //  skip, forward.cu:343, can only fire within ~1e-7 px of the mean and is not tested here)
                    // athr = 1/255 while the pixel is live, 2 once it is done (alpha <= 0.99 never reaches it)
                    const bool ok0 = al0 >= athr0, ok1 = al1 >= athr1;
                    if constexpr (MODE == 1) {
                        if (!__any_sync(0xffffffffu, ok0 || ok1)) continue;
                        const f32x2 go1 = f2_pack(ok0 ? f2_lo(oG) : 0.f, ok1 ? f2_hi(oG) : 0.f);
                        const f32x2 aT1 = f2_mul(f2_pack(fminf(0.99f, f2_lo(go1)), fminf(0.99f, f2_hi(go1))), T);
                        f2_acc(Cp, f2_dup(q.rb[0][slot].x), aT1);
                        f2_acc(T, aT1, f2_dup(-1.f));
                        continue;
                    }
                    const bool upd0 = ok0 && !(f2_lo(tT) < 0.0001f), upd1 = ok1 && !(f2_hi(tT) < 0.0001f);
                    athr0 = (ok0 && !upd0) ? 2.0f : athr0;               // stop BEFORE the Gaussian that takes T below 1e-4
                    athr1 = (ok1 && !upd1) ? 2.0f : athr1;
#if FB_S2_ANY
                    if (!__any_sync(0xffffffffu, upd0 || upd1)) continue;
#endif
#if FB_S2_STATS == 4
                    if (lane == 0) stat += 1.f;
#elif FB_S2_STATS == 5
                    stat += (upd0 ? 1.f : 0.f) + (upd1 ? 1.f : 0.f);
#endif
                    const float4 r0 = q.rb[0][slot], r1 = q.rb[1][slot], r2 = q.rb[2][slot], r3 = q.rb[3][slot];
                    const f32x2 inv = f2_pack(score_rcp(f2_lo(oma)), score_rcp(f2_hi(oma)));
                    // a pixel that does not take the entry runs the same arithmetic with go = alpha = 0: nothing changes
                    const f32x2 go = f2_pack(upd0 ? f2_lo(oG) : 0.f, upd1 ? f2_hi(oG) : 0.f);
                    const f32x2 aT = f2_mul(f2_pack(fminf(0.99f, f2_lo(go)), fminf(0.99f, f2_hi(go))), T);
                    const float csum = r0.x;
                    f2_acc(Cp, f2_dup(csum), aT);                                          // colour up to and including this one
                    const f32x2 A = f2_fma(T, f2_dup(csum), f2_mul(Cp, inv));
                    f2_acc(T, aT, f2_dup(-1.f));                                           // T (1 - alpha) as T - alpha T
                    // coefficients in monomial order (1, dx2, dxdy, dy2, dx3, dx2dy, dxdy2, dy3, dx4, dx3dy, dx2dy2, dxdy3, dy4)
                    //                               = r0.y r1.x r1.y  r1.z r1.w r2.x   r2.y   r2.z r2.w r3.x   r3.y    r3.z   r3.w
                    const float dx2 = dx * dx;
                    const float P0 = fmaf(dx2, fmaf(dx, fmaf(dx, r2.w, r1.w), r1.x), r0.y);
                    const float P1 = dx * fmaf(dx, fmaf(dx, r3.x, r2.x), r1.y);
                    const float P2 = fmaf(dx, fmaf(dx, r3.y, r2.y), r1.z);
                    const float P3 = fmaf(dx, r3.z, r2.z);
                    f32x2 poly = f2_fma(f2_dup(r3.w), dy, f2_dup(P3));
                    poly = f2_fma(poly, dy, f2_dup(P2));
                    poly = f2_fma(poly, dy, f2_dup(P1));
                    poly = f2_fma(poly, dy, f2_dup(P0));
                    const f32x2 g = f2_mul(f2_mul(go, go), poly);
                    const f32x2 gA = f2_mul(g, A), gB = f2_mul(g, inv);
                    f2_acc(Qaa, gA, A);
                    f2_acc(Qab, gA, inv);
                    f2_acc(Qbb, gB, inv);
                }
                __syncwarp();
                qn -= cnt; qh = (qh + cnt) & (FB_WQ - 1);
                if constexpr (MODE == 1) {      // every pixel's product below the stop threshold: what follows cannot matter
                    if (__all_sync(0xffffffffu, (athr0 > 1.f || f2_lo(T) < 0.0001f) && (athr1 > 1.f || f2_hi(T) < 0.0001f))) break;
                } else if (__all_sync(0xffffffffu, athr0 > 1.f && athr1 > 1.f)) break;
            }
        }
        __pipeline_wait_prior(0);
        __syncwarp();
    }
    S.T = T; S.Cp = Cp; S.Qaa = Qaa; S.Qab = Qab; S.Qbb = Qbb; S.athr0 = athr0; S.athr1 = athr1;
}

template <int D>
__device__ __forceinline__ void sweep_px2(const Bins& bins, const Cam& cam, float grad_scale, int tile_slot, int warp,
                                          WarpQ2& q2, float* __restrict__ out) {
    const int t = tile_slot % bins.tiles, v = tile_slot / bins.tiles;
    const int lane = threadIdx.x & 31;
    const int tx = t % cam.tiles_x, ty = t / cam.tiles_x;
    const int bx = tx * FB_TILE_X + (warp & 1) * 8, by = ty * FB_TILE_Y + (warp >> 1) * 8;
    const int px = bx + (lane & 7), py0 = by + (lane >> 3), py1 = py0 + 4;
    const bool in0 = px < cam.W && py0 < cam.H, in1 = px < cam.W && py1 < cam.H;
    if (bins.total[1]) { if (lane == 0) *out = 0.f; return; }
    const uint2 range = bins.tile_range[tile_slot];
    const int n = (int)(range.y - range.x);
    if (n == 0) { if (lane == 0) *out = 0.f; return; }
    const char* __restrict__ recA = reinterpret_cast<const char*>(bins.recA + (size_t)v * bins.vis_stride);
    const char* __restrict__ recB = reinterpret_cast<const char*>(bins.recB + (size_t)v * bins.vis_stride * 5);
    const unsigned* __restrict__ sorted = reinterpret_cast<const unsigned*>(bins.keys + range.x);   // low word = visible slot
    Px2State S;
    S.T = f2_dup(1.0f); S.Cp = f2_dup(0.f); S.Qaa = f2_dup(0.f); S.Qab = f2_dup(0.f); S.Qbb = f2_dup(0.f);
    S.athr0 = in0 ? 1.0f / 255.0f : 2.0f; S.athr1 = in1 ? 1.0f / 255.0f : 2.0f;
    float stat = 0.f;
    if (__any_sync(0xffffffffu, in0 || in1)) sweep_px2_core<D, 0>(recA, recB, sorted, n, bx, by, q2, S, stat);
    const f32x2 T = S.T, Cp = S.Cp, Qaa = S.Qaa, Qab = S.Qab, Qbb = S.Qbb;
    const float bgs = (cam.bg[0] + cam.bg[1]) + cam.bg[2];
    const float Cf0 = fmaf(f2_lo(T), bgs, f2_lo(Cp)), Cf1 = fmaf(f2_hi(T), bgs, f2_hi(Cp));     // C + T_final * bg
    const float s0 = fmaf(Cf0, fmaf(Cf0, f2_lo(Qbb), -2.f * f2_lo(Qab)), f2_lo(Qaa));
    const float s1 = fmaf(Cf1, fmaf(Cf1, f2_hi(Qbb), -2.f * f2_hi(Qab)), f2_hi(Qaa));
    float score = (grad_scale * grad_scale) * ((in0 ? s0 : 0.f) + (in1 ? s1 : 0.f));
#if FB_S2_STATS
    score = stat;
#endif
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) score += __shfl_xor_sync(0xffffffffu, score, o);
    if (lane == 0) *out = score;
}

__global__ void __launch_bounds__(32 * FB_S2_WPC, FB_S2_MINB) k_tile_score2(Bins bins, Cam cam, float grad_scale,
                                                                 float* __restrict__ partial) {
    extern __shared__ float4 smem[];
    constexpr int CPT = FB_S2_BLOCKS / FB_S2_WPC;
    const unsigned lin = blockIdx.y * gridDim.x + blockIdx.x;
    const int tile_slot = (int)bins.tile_order[lin / CPT];
    const int wq = threadIdx.x >> 5;
    const int warp = (int)(lin % CPT) * FB_S2_WPC + wq;            // 8x8 block of the tile, 0..3
    sweep_px2<FB_S2_PFD>(bins, cam, grad_scale, tile_slot, warp, reinterpret_cast<WarpQ2*>(smem)[wq],
                         partial + (size_t)tile_slot * FB_S2_BLOCKS + warp);
}

// ---- heavy tile lists: segments swept by different warps -------------------------------------------------
// A warp walks its tile's list alone, so the longest list of the chunk bounds the kernel: a view along a 30 m floor
// puts > 100 k tiny splats on each tile of the horizon row (cfg 4: ~30 tiles of a view hold 80 % of its instances, the
// longest list 272 k keys) and a lone warp needs milliseconds for one of them while the rest of the GPU has drained
// (clock64 per warp at 3M Gaussians: the longest warp of a view runs as long as the whole kernel).
// The front-to-back recurrence is cut instead.  For a segment [k0, k1) of the list, the state of a pixel after it is
//   T_out = T_in * Tl,   Cp_out = Cp_in + T_in * Cpl,      Tl = prod (1 - alpha_i),  Cpl = sum c_i alpha_i prod_{j<i} (1 - alpha_j)
// (products over the pairs with alpha >= 1/255), and the stop rule -- stop BEFORE the Gaussian that takes T below 1e-4
// -- is a threshold on the same monotone product: a pixel is done on entry to a segment iff T_in < 1e-4.  So
//   pass 1  every segment but the last computes (Tl, Cpl) of its keys alone (exponent and alpha only: no polynomial,
//           no accumulators, 32 of the record's 80 bytes), all segments of a list in parallel;
//   chain   segment s waits for (T_in, Cp_in) of segment s - 1, publishes its own (T_out, Cp_out): a few us per link;
//   pass 2  the ordinary scoring sweep of the segment's keys from (T_in, Cp_in), stop rule included.
// The accumulators of the segments add (Q_AA, Q_AB, Q_BB are sums over the list), the final C + T bg is taken from
// the segment in which the pixel stopped (or the last one), and the warp that finishes a (tile, pixel block) last
// combines the segments in order: a fixed order of additions, no float atomics.  Against one sequential sweep only
// the association of the product T_in differs (ulps; a stop decision at the 1e-4 threshold may move by one entry).
// The heavy lists {tile slot, segments, first item} come from k_scan_tiles; the warps of the first `workers` CTAs of
// the grid take tickets in the order (segment index, list, pixel block), so the segment a warp waits for has always
// been started.  While there are more lists than warps the previous segment of a list has FINISHED by the time the next
// one is taken up: its exact end state is there, pass 1 is skipped, and the list is swept segment after segment at
// the cost of one sequential sweep (a list whose pixels are all done costs a flag read per remaining segment).  When
// the lists run out -- the tail this is for -- idle warps take segments whose predecessor is still running, and those
// go through pass 1 / chain / pass 2 in parallel.
struct SegWs {
    float4*   state;   // [items][4 blocks][32 lanes]      {T0, T1, Cp0, Cp1} after the segment (plain products)
    float4*   out;     // [items][4 blocks][2][32 lanes]   {Qaa0, Qaa1, Qab0, Qab1}, {Qbb0, Qbb1, Cf0, Cf1} (Cf = NaN unless the pixel's final state is here)
    unsigned* flag;    // [items][8]  0-3: state of the block published; 4-7: finished segments of the block, kept at the
};                     //             list's first item (bins.seg_flags: zeroed by k_scan_tiles when it emits the item)
#ifndef FB_SEG_AHEAD_WAVES
#define FB_SEG_AHEAD_WAVES 100000   // pass 1 may run ahead of an unfinished predecessor only in the last so many waves of worker warps (100000: always)
#endif
#ifndef FB_SEG_SLEEP_NS
#define FB_SEG_SLEEP_NS 100   // pause between two looks at the predecessor's flag
#endif
#define FB_SEG_WS_BYTES_PER_ITEM (4 * 32 * 16 + 4 * 64 * 16 + 16 + 16)

template <int D>
__device__ __forceinline__ void sweep_px2_segment(const Bins& bins, const Cam& cam, float grad_scale, const SegWs& ws,
                                                  const int tile_slot, const int seg, const int S, const unsigned base,
                                                  const int warp, const bool may_run_ahead, WarpQ2& q2,
                                                  float* __restrict__ partial) {
    const unsigned item = base + (unsigned)seg;
    const int t = tile_slot % bins.tiles, v = tile_slot / bins.tiles;
    const int lane = threadIdx.x & 31;
    const int tx = t % cam.tiles_x, ty = t / cam.tiles_x;
    const int bx = tx * FB_TILE_X + (warp & 1) * 8, by = ty * FB_TILE_Y + (warp >> 1) * 8;
    const int px = bx + (lane & 7), py0 = by + (lane >> 3), py1 = py0 + 4;
    const bool in0 = px < cam.W && py0 < cam.H, in1 = px < cam.W && py1 < cam.H;
    const uint2 range = bins.tile_range[tile_slot];
    const int k0 = seg * FB_SEG_KEYS, n = min((int)(range.y - range.x) - k0, FB_SEG_KEYS);
    const char* __restrict__ recA = reinterpret_cast<const char*>(bins.recA + (size_t)v * bins.vis_stride);
    const char* __restrict__ recB = reinterpret_cast<const char*>(bins.recB + (size_t)v * bins.vis_stride * 5);
    const unsigned* __restrict__ sorted = reinterpret_cast<const unsigned*>(bins.keys + range.x + k0);
    const bool last = seg + 1 == S;
    const size_t me = (size_t)item * 4 + warp;
    volatile unsigned* my_flag = ws.flag + (size_t)item * 8 + warp;
    volatile unsigned* prev_flag = my_flag - 8;
    float stat = 0.f;
    Px2State P;
    P.T = f2_dup(1.0f); P.Cp = f2_dup(0.f); P.Qaa = f2_dup(0.f); P.Qab = f2_dup(0.f); P.Qbb = f2_dup(0.f);
    P.athr0 = in0 ? 1.0f / 255.0f : 2.0f; P.athr1 = in1 ? 1.0f / 255.0f : 2.0f;
    // the state on entry: there already if the previous segment has finished (the usual case while the GPU is busy:
    // the list is then swept segment after segment at no extra cost); else pass 1 runs on this segment meanwhile
    f32x2 Tin = f2_dup(1.0f), Cin = f2_dup(0.f);
    bool have_in = seg == 0;
    if (seg > 0) {
        unsigned f = 0;
        if (lane == 0) f = *prev_flag;
        f = __shfl_sync(0xffffffffu, f, 0);
        if (f) {
            __threadfence();
            const float4 st = __ldcg(ws.state + (me - 4) * 32 + lane);
            Tin = f2_pack(st.x, st.y); Cin = f2_pack(st.z, st.w);
            have_in = true;
        }
    }
    if (!have_in && !may_run_ahead) {
        // plenty of other lists are waiting for a warp: pass 1 would spend issue slots they can use, so wait for the
        // predecessor (picked a whole row of tickets earlier: nearly done) and sweep from its exact end state
        if (lane == 0) { while (*prev_flag == 0u) __nanosleep(FB_SEG_SLEEP_NS); }
        __syncwarp();
        __threadfence();
        const float4 st = __ldcg(ws.state + (me - 4) * 32 + lane);
        Tin = f2_pack(st.x, st.y); Cin = f2_pack(st.z, st.w);
        have_in = true;
    }
    const bool direct = have_in;                       // no pass 1: the state published is the exact one after pass 2
    if (!direct && !last && __any_sync(0xffffffffu, in0 || in1)) sweep_px2_core<D, 1>(recA, recB, sorted, n, bx, by, q2, P, stat);
    if (!have_in) {
        if (lane == 0) { while (*prev_flag == 0u) __nanosleep(FB_SEG_SLEEP_NS); }
        __syncwarp();
        __threadfence();
        const float4 st = __ldcg(ws.state + (me - 4) * 32 + lane);
        Tin = f2_pack(st.x, st.y); Cin = f2_pack(st.z, st.w);
    }
    f32x2 Tout = f2_mul(Tin, P.T), Cout = f2_fma(Tin, P.Cp, Cin);
    if (!direct && !last) {
        __stcg(ws.state + me * 32 + lane, make_float4(f2_lo(Tout), f2_hi(Tout), f2_lo(Cout), f2_hi(Cout)));
        __threadfence();
        __syncwarp();
        if (lane == 0) *my_flag = 1u;
    }
    // pass 2: the scoring sweep of this segment's keys from (T_in, Cp_in); a pixel is done on entry iff T_in < 1e-4
    const bool live0 = in0 && !(f2_lo(Tin) < 0.0001f), live1 = in1 && !(f2_hi(Tin) < 0.0001f);
    P.T = Tin; P.Cp = Cin;
    P.athr0 = live0 ? 1.0f / 255.0f : 2.0f; P.athr1 = live1 ? 1.0f / 255.0f : 2.0f;
    if (__any_sync(0xffffffffu, live0 || live1)) sweep_px2_core<D, 0>(recA, recB, sorted, n, bx, by, q2, P, stat);
    if (direct) {
        // exact state after the segment; a pixel that stopped inside it is handed on as T = 0 (done)
        Tout = f2_pack(P.athr0 > 1.f ? 0.f : f2_lo(P.T), P.athr1 > 1.f ? 0.f : f2_hi(P.T));
        if (!last) {
            __stcg(ws.state + me * 32 + lane, make_float4(f2_lo(Tout), f2_hi(Tout), f2_lo(P.Cp), f2_hi(P.Cp)));
            __threadfence();
            __syncwarp();
            if (lane == 0) *my_flag = 1u;
        }
    }
    // the pixel's final state is here if it was live on entry and the next segment finds it done (or there is none)
    const float bgs = (cam.bg[0] + cam.bg[1]) + cam.bg[2];
    const float qnan = __int_as_float(0x7fc00000);
    const bool own0 = live0 && (last || f2_lo(Tout) < 0.0001f), own1 = live1 && (last || f2_hi(Tout) < 0.0001f);
    const float Cf0 = own0 ? fmaf(f2_lo(P.T), bgs, f2_lo(P.Cp)) : qnan, Cf1 = own1 ? fmaf(f2_hi(P.T), bgs, f2_hi(P.Cp)) : qnan;
    __stcg(ws.out + me * 64 + lane, make_float4(f2_lo(P.Qaa), f2_hi(P.Qaa), f2_lo(P.Qab), f2_hi(P.Qab)));
    __stcg(ws.out + me * 64 + 32 + lane, make_float4(f2_lo(P.Qbb), f2_hi(P.Qbb), Cf0, Cf1));
    __threadfence();
    __syncwarp();
    unsigned fin = 0;
    if (lane == 0) fin = atomicAdd(ws.flag + (size_t)base * 8 + 4 + warp, 1u);
    fin = __shfl_sync(0xffffffffu, fin, 0);
    if (fin + 1 != (unsigned)S) return;
    // last to finish this (tile, pixel block): combine the segments in order
    __threadfence();
    float Qaa0 = 0.f, Qaa1 = 0.f, Qab0 = 0.f, Qab1 = 0.f, Qbb0 = 0.f, Qbb1 = 0.f, C0 = 0.f, C1 = 0.f;
    for (int k = 0; k < S; ++k) {
        const size_t o = ((size_t)(base + k) * 4 + warp) * 64;
        const float4 u = __ldcg(ws.out + o + lane), w = __ldcg(ws.out + o + 32 + lane);
        Qaa0 += u.x; Qaa1 += u.y; Qab0 += u.z; Qab1 += u.w; Qbb0 += w.x; Qbb1 += w.y;
        if (w.z == w.z) C0 = w.z;
        if (w.w == w.w) C1 = w.w;
    }
    const float s0 = fmaf(C0, fmaf(C0, Qbb0, -2.f * Qab0), Qaa0), s1 = fmaf(C1, fmaf(C1, Qbb1, -2.f * Qab1), Qaa1);
    float score = (grad_scale * grad_scale) * ((in0 ? s0 : 0.f) + (in1 ? s1 : 0.f));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) score += __shfl_xor_sync(0xffffffffu, score, o);
    if (lane == 0) { partial[(size_t)tile_slot * FB_TILE_WARPS + warp] = score; partial[(size_t)tile_slot * FB_TILE_WARPS + warp + 4] = 0.f; }
}

// k_tile_score_mix: grid = [workers: CTAs whose warps take (heavy work item, pixel block) tickets][1 CTA per tile, the
// four 8x8 blocks of the tile; the CTA of a heavy tile exits].  partial: FB_TILE_WARPS floats per tile (four used).
#define FB_MIX_SMEM_BYTES ((int)sizeof(WarpQ2) * 4)
#ifndef FB_SEG_WORKERS
#define FB_SEG_WORKERS (148 * 8)
#endif
__global__ void __launch_bounds__(128, FB_S2_MINB) k_tile_score_mix(Bins bins, Cam cam, float grad_scale,
                                                                    float* __restrict__ partial, SegWs ws, int workers) {
    extern __shared__ float4 smem[];
    const int wq = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (bins.total[1]) {
        if ((int)blockIdx.x >= workers && lane == 0) {
            float* out = partial + (size_t)bins.tile_order[blockIdx.x - workers] * FB_TILE_WARPS + wq;
            out[0] = 0.f; out[4] = 0.f;
        }
        return;
    }
    if ((int)blockIdx.x < workers) {
        // tickets in the order (segment index, list, pixel block): segment s of every list before any segment s + 1
        const unsigned row = 4u * bins.total[12], tickets = row * bins.total[13];
        for (;;) {
            unsigned tk = 0;
            if (lane == 0) tk = atomicAdd(&bins.total[11], 1u);
            tk = __shfl_sync(0xffffffffu, tk, 0);
            if (tk >= tickets) return;
            const unsigned sgm = tk / row, r = tk - sgm * row;
            const uint4 it = bins.seg_list[r >> 2];                        // {tile slot, segments, first item, -}
            if (sgm < it.y)
                sweep_px2_segment<FB_S2_PFD>(bins, cam, grad_scale, ws, (int)it.x, (int)sgm, (int)it.y, it.z, (int)(r & 3u),
                                             tk + (unsigned)(FB_SEG_AHEAD_WAVES * 4 * workers) >= tickets,
                                             reinterpret_cast<WarpQ2*>(smem)[wq], partial);
            __syncwarp();
        }
    }
    const int idx = (int)blockIdx.x - workers;
    const int tile_slot = (int)bins.tile_order[idx];
    if (order_rank((unsigned)idx, (unsigned)bins.V, (unsigned)bins.tiles) < FB_HEAVY_RANKS) {
        const uint2 r = bins.tile_range[tile_slot];
        if ((int)(r.y - r.x) >= FB_HEAVY_KEYS) return;                     // its segments are work items
    }
    float* out = partial + (size_t)tile_slot * FB_TILE_WARPS + wq;
    if (lane == 0) out[4] = 0.f;
    sweep_px2<FB_S2_PFD>(bins, cam, grad_scale, tile_slot, wq, reinterpret_cast<WarpQ2*>(smem)[wq], out);
}

// ---- per-Gaussian Hessian moments (compute_H_train), warp-autonomous form ----------------------------------
// Same traversal as k_tile_score (two sweeps, the reference's arithmetic: expf, IEEE division, the recurrences of
// backward.cu:978-1017 with r+g+b as the one colour channel), but what a contributing pair produces is the 13
// monomial moments  s^2 * dx^i dy^j  of fb_fisher.cuh (power = 2).  They are summed over the 32 pixels of the
// block with the halving butterfly (16 SHFL), four neighbouring sums are gathered into one lane and added to the
// Gaussian's 16-float row with ONE red.global.add.v4.f32 per quad: 4 vector atomics per (Gaussian, pixel block).
// k_expand_H turns the rows into H columns.  Replaces a per-view render + block-synchronous tile backward on the batched H_train path (every
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
        // record gathers FB_HESS_PFD scan steps ahead of their use, keys one more (measured: no gain from more than one)
        constexpr int D = FB_HESS_PFD;
        bool vq[D]; unsigned gq[D]; float4 aq[D];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            vq[d] = d * 32 + lane < n;
            gq[d] = vq[d] ? (unsigned)(sorted[d * 32 + lane] & 0xffffffffull) : 0u;
            aq[d] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (vq[d]) aq[d] = __ldg(reinterpret_cast<const float4*>(recs + gq[d]));
        }
        u64 key2 = (D * 32 + lane < n) ? sorted[D * 32 + lane] : 0ull;
        for (int step = 0; step <= nsteps; ++step) {
            if (step < nsteps) {
                const float4 a = aq[0];
                const unsigned gid = gq[0];
                const bool valid = vq[0];
#pragma unroll
                for (int d = 0; d + 1 < D; ++d) { vq[d] = vq[d + 1]; gq[d] = gq[d + 1]; aq[d] = aq[d + 1]; }
                vq[D - 1] = (step + D) * 32 + lane < n;
                gq[D - 1] = (unsigned)(key2 & 0xffffffffull);
                if (vq[D - 1]) aq[D - 1] = __ldg(reinterpret_cast<const float4*>(recs + gq[D - 1]));
                key2 = ((step + D + 1) * 32 + lane < n) ? sorted[(step + D + 1) * 32 + lane] : 0ull;
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
        constexpr int D = FB_HESS_PFD;
        bool vq[D]; unsigned gq[D]; float4 aq[D];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            vq[d] = d * 32 + lane < warp_last;
            gq[d] = vq[d] ? (unsigned)(sorted[warp_last - 1 - d * 32 - lane] & 0xffffffffull) : 0u;
            aq[d] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (vq[d]) aq[d] = __ldg(reinterpret_cast<const float4*>(recs + gq[d]));
        }
        u64 key2 = (D * 32 + lane < warp_last) ? sorted[warp_last - 1 - D * 32 - lane] : 0ull;
        for (int step = 0; step <= nsteps; ++step) {
            if (step < nsteps) {
                const float4 a = aq[0];
                const unsigned gid = gq[0];
                const bool valid = vq[0];
#pragma unroll
                for (int d = 0; d + 1 < D; ++d) { vq[d] = vq[d + 1]; gq[d] = gq[d + 1]; aq[d] = aq[d + 1]; }
                vq[D - 1] = (step + D) * 32 + lane < warp_last;
                gq[D - 1] = (unsigned)(key2 & 0xffffffffull);
                if (vq[D - 1]) aq[D - 1] = __ldg(reinterpret_cast<const float4*>(recs + gq[D - 1]));
                key2 = ((step + D + 1) * 32 + lane < warp_last) ? sorted[warp_last - 1 - (step + D + 1) * 32 - lane] : 0ull;
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

// ---- k_tile_hessian2: the H_train moments on 8x8 blocks, two pixels per lane ---------------------------------
// Same quantities as k_tile_hessian in the form of the scoring sweep: a warp owns an 8x8 pixel block, a lane the pixels
// (x, y) and (x, y + 4), packed FP32 where the two differ, the exact ellipse test on every ring batch.  Both passes run
// FRONT TO BACK with the same recurrence (and therefore stop at the same entry): pass 0 leaves C' = C + T_final bg of
// every pixel, pass 1 forms dL/dalpha_i = A_i - C' B_i as the scoring sweep does ("single sweep" above) -- no
// back-to-front traversal, no T / (1 - alpha) reconstruction -- and adds  s^2 dx^i dy^j  of the 64 pixels to the
// Gaussian's row: two pixels summed in the lane, the halving butterfly, one red.global.add.v4.f32 per quad.
// Half as many (Gaussian, block) reductions and atomics as with 8x4 blocks, a sixth fewer bodies (exact test).
#ifndef FB_HESS_PX2
#define FB_HESS_PX2 1
#endif
template <int PASS>
__device__ __forceinline__ void hess_px2_pass(const GRec* __restrict__ recs, const float* __restrict__ colors,
                                              const u64* __restrict__ sorted, const int n, const int bx, const int by,
                                              WarpQ2& q2, Px2State& S, const f32x2 Cf, const float dpix, float* __restrict__ mom) {
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const float pxf = (float)(bx + (lane & 7));
    const f32x2 npy = f2_pack(-(float)(by + (lane >> 3)), -(float)(by + (lane >> 3) + 4));
    const float cx = (float)bx + 3.5f, cy = (float)by + 3.5f;
    WarpQ& q = q2.q;
    f32x2 T = S.T, Cp = S.Cp;
    float athr0 = S.athr0, athr1 = S.athr1;
    const int rslot = reduce16_slot(lane);
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
            const bool hit = valid && (fabsf(a.x - cx) <= a.z + 3.5f) && (fabsf(a.y - cy) <= a.w + 3.5f);
            const unsigned mask = __ballot_sync(0xffffffffu, hit);
            if (hit) {
                const int slot = (qh + qn + __popc(mask & lt)) & (FB_WQ - 1);
                q.a[slot] = make_float4(a.x, a.y, 0.f, __uint_as_float(gid));
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
            unsigned todo = cnt == 32 ? 0xffffffffu : ((1u << cnt) - 1u);
            {
                // lane h owns ring entry qh + h: the conic goes into the exponent's log2 units (as the scoring records
                // carry it), then the exact ellipse / block test of sweep_px2_core
                float4 co = q.b[qh + lane];
                co.x *= -0.72134752044448170f; co.y *= -1.4426950408889634f; co.z *= -0.72134752044448170f;
                q.b[qh + lane] = co;
                const float4 a = q.a[qh + lane];
                const float X1 = a.x - (float)bx, X0 = X1 - 7.f, Y1 = a.y - (float)by, Y0 = Y1 - 7.f;
                const float ia = score_rcp(co.x + co.x), ic = score_rcp(co.z + co.z);
                auto pw = [&](float x, float y) { return fmaf(co.z, y * y, x * fmaf(co.y, y, co.x * x)); };
                auto ex_ = [&](float Y) { return pw(fminf(fmaxf(-co.y * Y * ia, X0), X1), Y); };
                auto ey_ = [&](float X) { return pw(X, fminf(fmaxf(-co.y * X * ic, Y0), Y1)); };
                float m = fmaxf(fmaxf(ex_(Y0), ex_(Y1)), fmaxf(ey_(X0), ey_(X1)));
                if (X0 <= 0.f && X1 >= 0.f && Y0 <= 0.f && Y1 >= 0.f) m = 0.f;
                float lg;
                asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(255.0f * co.w));
                const bool keep = !(m < -lg - 0.02f);
                todo &= __ballot_sync(0xffffffffu, keep);
            }
            if ((todo >> lane) & 1u) q2.list[__popc(todo & lt)] = (unsigned char)(qh + lane);
            __syncwarp();
            const int nkeep = __popc(todo);
            for (int i = 0; i < nkeep; ++i) {
                const int slot = q2.list[i];
                const float4 mu = q.a[slot];
                const float4 co = q.b[slot];
                const float dx = mu.x - pxf;
                const f32x2 dy = f2_add(f2_dup(mu.y), npy);
                const float Adx = co.x * dx;
                const float q0 = Adx * dx, q1 = co.y * dx;
                const f32x2 pw = f2_fma(f2_fma(f2_dup(co.z), dy, f2_dup(q1)), dy, f2_dup(q0));
                const float G0 = score_ex2(f2_lo(pw)), G1 = score_ex2(f2_hi(pw));
                const f32x2 oG = f2_mul(f2_dup(co.w), f2_pack(G0, G1));
                const float al0 = fminf(0.99f, f2_lo(oG)), al1 = fminf(0.99f, f2_hi(oG));
                const f32x2 oma = f2_fma(f2_pack(al0, al1), f2_dup(-1.f), f2_dup(1.f));
                const f32x2 tT = f2_mul(T, oma);
                const bool ok0 = al0 >= athr0, ok1 = al1 >= athr1;
                const bool upd0 = ok0 && !(f2_lo(tT) < 0.0001f), upd1 = ok1 && !(f2_hi(tT) < 0.0001f);
                athr0 = (ok0 && !upd0) ? 2.0f : athr0;
                athr1 = (ok1 && !upd1) ? 2.0f : athr1;
                if (!__any_sync(0xffffffffu, upd0 || upd1)) continue;
                const float4 c3 = q.rb[0][slot];
                const float csum = (c3.x + c3.y) + c3.z;
                const f32x2 go = f2_pack(upd0 ? f2_lo(oG) : 0.f, upd1 ? f2_hi(oG) : 0.f);
                const f32x2 aT = f2_mul(f2_pack(fminf(0.99f, f2_lo(go)), fminf(0.99f, f2_hi(go))), T);
                if constexpr (PASS == 0) {
                    f2_acc(Cp, f2_dup(csum), aT);
                    f2_acc(T, aT, f2_dup(-1.f));
                } else {
                    const f32x2 inv = f2_pack(score_rcp(f2_lo(oma)), score_rcp(f2_hi(oma)));
                    f2_acc(Cp, f2_dup(csum), aT);
                    const f32x2 A = f2_fma(T, f2_dup(csum), f2_mul(Cp, inv));
                    f2_acc(T, aT, f2_dup(-1.f));
                    const f32x2 dLda = f2_fma(f2_mul(Cf, f2_dup(-1.f)), inv, A);               // A - C' B
                    const f32x2 sv = f2_mul(f2_mul(go, dLda), f2_dup(dpix));
                    const f32x2 qq = f2_mul(sv, sv);
                    const f32x2 dxx = f2_dup(dx);
                    const f32x2 qx = f2_mul(qq, dxx), qy = f2_mul(qq, dy);
                    const f32x2 m20 = f2_mul(qx, dxx), m11 = f2_mul(qx, dy), m02 = f2_mul(qy, dy);
                    const f32x2 m30 = f2_mul(m20, dxx), m21 = f2_mul(m20, dy), m12 = f2_mul(m02, dxx), m03 = f2_mul(m02, dy);
                    const f32x2 m40 = f2_mul(m30, dxx), m31 = f2_mul(m30, dy), m22 = f2_mul(m21, dy), m13 = f2_mul(m03, dxx), m04 = f2_mul(m03, dy);
                    float m[FB_NACC];
                    m[0] = f2_lo(qq) + f2_hi(qq);   m[1] = f2_lo(m20) + f2_hi(m20); m[2] = f2_lo(m11) + f2_hi(m11);
                    m[3] = f2_lo(m02) + f2_hi(m02); m[4] = f2_lo(m30) + f2_hi(m30); m[5] = f2_lo(m21) + f2_hi(m21);
                    m[6] = f2_lo(m12) + f2_hi(m12); m[7] = f2_lo(m03) + f2_hi(m03); m[8] = f2_lo(m40) + f2_hi(m40);
                    m[9] = f2_lo(m31) + f2_hi(m31); m[10] = f2_lo(m22) + f2_hi(m22); m[11] = f2_lo(m13) + f2_hi(m13);
                    m[12] = f2_lo(m04) + f2_hi(m04); m[13] = 0.f; m[14] = 0.f; m[15] = 0.f;
                    const float tot = warp_reduce16(m, lane);        // even lane L holds the block sum of moment rslot
                    float4 quad;
                    quad.x = tot;
                    quad.y = __shfl_down_sync(0xffffffffu, tot, 2);
                    quad.z = __shfl_down_sync(0xffffffffu, tot, 4);
                    quad.w = __shfl_down_sync(0xffffffffu, tot, 6);
                    if ((lane & 7) == 0 && (quad.x != 0.f || quad.y != 0.f || quad.z != 0.f || quad.w != 0.f))
                        atomicAdd(reinterpret_cast<float4*>(mom + (size_t)__float_as_uint(mu.w) * FB_NACC) + (rslot >> 2), quad);
                }
            }
            __syncwarp();
            qn -= cnt; qh = (qh + cnt) & (FB_WQ - 1);
            if (__all_sync(0xffffffffu, athr0 > 1.f && athr1 > 1.f)) break;
        }
    }
    __pipeline_wait_prior(0);
    __syncwarp();
    S.T = T; S.Cp = Cp; S.athr0 = athr0; S.athr1 = athr1;
}

__global__ void __launch_bounds__(128, FB_S2_MINB) k_tile_hessian2(Bins bins, Cam cam, const float* __restrict__ colors,
                                                                   float grad_scale, float* __restrict__ mom_all) {
    extern __shared__ float4 smem[];
    const int tile_slot = (int)bins.tile_order[blockIdx.x];
    const int t = tile_slot % bins.tiles, v = tile_slot / bins.tiles;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tx = t % cam.tiles_x, ty = t / cam.tiles_x;
    const int bx = tx * FB_TILE_X + (warp & 1) * 8, by = ty * FB_TILE_Y + (warp >> 1) * 8;
    const int px = bx + (lane & 7), py0 = by + (lane >> 3), py1 = py0 + 4;
    const bool in0 = px < cam.W && py0 < cam.H, in1 = px < cam.W && py1 < cam.H;
    if (bins.total[1]) return;
    const uint2 range = bins.tile_range[tile_slot];
    const int n = (int)(range.y - range.x);
    if (n == 0 || !__any_sync(0xffffffffu, in0 || in1)) return;
    const GRec* __restrict__ recs = bins.recs + (size_t)v * bins.P;
    float* __restrict__ mom = mom_all + (size_t)v * bins.P * FB_NACC;
    const u64* __restrict__ sorted = bins.keys + range.x;
    WarpQ2& q2 = reinterpret_cast<WarpQ2*>(smem)[warp];
    Px2State S;
    S.T = f2_dup(1.0f); S.Cp = f2_dup(0.f);
    S.athr0 = in0 ? 1.0f / 255.0f : 2.0f; S.athr1 = in1 ? 1.0f / 255.0f : 2.0f;
    hess_px2_pass<0>(recs, colors, sorted, n, bx, by, q2, S, f2_dup(0.f), 0.f, nullptr);
    const float bgs = (cam.bg[0] + cam.bg[1]) + cam.bg[2];
    const f32x2 Cf = f2_pack(fmaf(f2_lo(S.T), bgs, f2_lo(S.Cp)), fmaf(f2_hi(S.T), bgs, f2_hi(S.Cp)));
    S.T = f2_dup(1.0f); S.Cp = f2_dup(0.f);
    S.athr0 = in0 ? 1.0f / 255.0f : 2.0f; S.athr1 = in1 ? 1.0f / 255.0f : 2.0f;
    __syncwarp();
    hess_px2_pass<1>(recs, colors, sorted, n, bx, by, q2, S, Cf, grad_scale, mom);
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
