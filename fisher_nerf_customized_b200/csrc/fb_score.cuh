// fb_score.cuh -- fused per-(view, tile) candidate-scoring kernel.
//
// One CTA owns one 16x16 tile of one candidate view and does, without leaving the SM:
//   (1) pull the tile's key segment into shared memory and sort it by (depth, id)        [stage 2c]
//   (2) front-to-back transmittance pass -> T_final, n_contrib per pixel (registers)      [stage 3]
//   (3) back-to-front pass: per contributing pair, s = G*o*dL/dalpha and
//       score += s^2 * poly_g(dx,dy)                                                      [stage 4]
//   (4) block reduce -> one float per (view, tile); k_score_finalize sums the tiles in fixed order.
// No image, no final_T/n_contrib round trip, no per-Gaussian Hessian tensor, no atomics.
// Replaces the K-iteration Python loop  models/SLAM/gaussian.py:1362-1370  (+ compute_Hessian
// :1503-1570) around renderCUDA / renderCUDAFused (forward.cu:261-393, backward.cu:850-1140).
#pragma once
#include "fb_stage12.cuh"

namespace fb {

struct ScoreBatch {
    float4 xyo[FB_TILE_PIX];
    float4 con[FB_TILE_PIX];
    float4 rb[4][FB_TILE_PIX];   // rgb+coef0 | coef1-4 | coef5-8 | coef9-12
};

// dynamic smem: [u64 keys[smem_keys]] [ScoreBatch] [float red[8]] [int block_last]
__global__ void __launch_bounds__(FB_TILE_PIX) k_tile_score(Bins bins, Cam cam, float grad_scale, int smem_keys,
                                                           float* __restrict__ partial) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* s_keys = reinterpret_cast<u64*>(smem_raw);
    ScoreBatch& sb = *reinterpret_cast<ScoreBatch*>(smem_raw + (size_t)smem_keys * 8);
    float* red = reinterpret_cast<float*>(smem_raw + (size_t)smem_keys * 8 + sizeof(ScoreBatch));
    int* s_last = reinterpret_cast<int*>(red + 8);

    const int t = blockIdx.x, v = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31;
    const int tx = t % cam.tiles_x, ty = t / cam.tiles_x;
    const int px = tx * FB_TILE_X + tid % FB_TILE_X, py = ty * FB_TILE_Y + tid / FB_TILE_X;
    const bool inside = px < cam.W && py < cam.H;
    const float pxf = (float)px, pyf = (float)py;
    const int tile_slot = v * bins.tiles + t;
    if (bins.total[1]) { if (tid == 0) partial[tile_slot] = 0.f; return; }
    const uint2 range = bins.tile_range[tile_slot];
    const int n = (int)(range.y - range.x);
    if (n == 0) { if (tid == 0) partial[tile_slot] = 0.f; return; }
    const GRec* __restrict__ recs = bins.recs + (size_t)v * bins.P;
    const float4* __restrict__ recsB = bins.recsB + (size_t)v * bins.P * 4;

    // (1) sort
    u64* seg = bins.keys + range.x;
    const u64* sorted;
    if (n <= smem_keys) {
        for (int i = tid; i < n; i += FB_TILE_PIX) s_keys[i] = seg[i];
        __syncthreads();
        cta_sort_u64(s_keys, n);
        sorted = s_keys;
    } else {
        cta_sort_u64(seg, n);
        sorted = seg;
    }
    __syncthreads();

    // (2) transmittance pass (forward.cu:330-379 without colour / depth)
    float T = 1.0f;
    int last = 0;
    bool done = !inside;
    for (int base = 0; base < n; base += FB_TILE_PIX) {
        if (__syncthreads_count(done) == FB_TILE_PIX) break;
        int i = base + tid;
        if (i < n) {
            GRec r = recs[(unsigned)(sorted[i] & 0xffffffffull)];
            sb.xyo[tid] = r.a;
            sb.con[tid] = r.b;
        }
        __syncthreads();
        const int m = min(FB_TILE_PIX, n - base);
        for (int j = 0; !done && j < m; ++j) {
            const float4 xy = sb.xyo[j];
            const float4 co = sb.con[j];
            const float dx = xy.x - pxf, dy = xy.y - pyf;
            const float power = pair_power(co.x, co.y, co.z, dx, dy);
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, xy.w * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = __fmul_rn(T, __fsub_rn(1.f, alpha));
            if (test_T < 0.0001f) { done = true; continue; }
            T = test_T;
            last = base + j + 1;
        }
    }

    // (3) back-to-front scoring pass (backward.cu:939-1038)
    if (tid == 0) *s_last = 0;
    __syncthreads();
    const int warp_last = __reduce_max_sync(0xffffffffu, last);
    if (lane == 0 && warp_last > 0) atomicMax(s_last, warp_last);
    __syncthreads();
    const int block_last = *s_last;
    const float T_final = T;
    const float dpix = inside ? grad_scale : 0.f;
    const float bgdot = (cam.bg[0] + cam.bg[1] + cam.bg[2]) * dpix;
    float accum_rec[3] = {0.f, 0.f, 0.f}, last_color[3] = {0.f, 0.f, 0.f};
    float last_alpha = 0.f;
    float score = 0.f;
    for (int top = block_last; top > 0; top -= FB_TILE_PIX) {
        const int m = min(FB_TILE_PIX, top);
        __syncthreads();
        if (tid < m) {
            unsigned gid = (unsigned)(sorted[top - 1 - tid] & 0xffffffffull);
            GRec r = recs[gid];
            sb.xyo[tid] = r.a;
            sb.con[tid] = r.b;
            const float4* rb = recsB + (size_t)gid * 4;
#pragma unroll
            for (int k = 0; k < 4; ++k) sb.rb[k][tid] = rb[k];
        }
        __syncthreads();
        for (int j = max(0, top - warp_last); j < m; ++j) {
            const int pos = top - 1 - j;
            if (pos >= last) continue;
            const float4 xy = sb.xyo[j];
            const float4 co = sb.con[j];
            const float dx = xy.x - pxf, dy = xy.y - pyf;
            const float power = pair_power(co.x, co.y, co.z, dx, dy);
            if (power > 0.0f) continue;
            const float G = expf(power);
            const float alpha = fminf(0.99f, xy.w * G);
            if (alpha < 1.0f / 255.0f) continue;
            T = T / (1.f - alpha);
            const float4 r0 = sb.rb[0][j];
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
            dL_dalpha += (-T_final / (1.f - alpha)) * bgdot;
            const float s = G * (xy.w * dL_dalpha);
            const float4 r1 = sb.rb[1][j], r2 = sb.rb[2][j], r3 = sb.rb[3][j];
            const float c[13] = {r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w, r3.x, r3.y, r3.z, r3.w};
            score += (s * s) * eval_poly13(c, dx, dy);
        }
    }

    // (4) block reduce
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) score += __shfl_xor_sync(0xffffffffu, score, o);
    __syncthreads();
    if (lane == 0) red[tid >> 5] = score;
    __syncthreads();
    if (tid == 0) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < FB_TILE_PIX / 32; ++w) s += red[w];
        partial[tile_slot] = s;
    }
}

// scores[k0 + v] = sum over tiles of partial[v][t], fixed order (deterministic given partials).
__global__ void k_score_finalize(const float* __restrict__ partial, int tiles, float* __restrict__ scores) {
    const int v = blockIdx.x;
    float s = 0.f;
    for (int t = threadIdx.x; t < tiles; t += 32) s += partial[v * tiles + t];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (threadIdx.x == 0) scores[v] = s;
}

}  // namespace fb
