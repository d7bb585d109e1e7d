// fb_fisher_tile.cuh -- the per-tile backward walk (pixel-parallel, back to front) that turns
// every (pixel, Gaussian) pair into monomial moments, reduces them across the tile and hands one
// 16-float row per (Gaussian, tile) instance to a flush functor.
// Pair arithmetic: DGR/cuda_rasterizer/backward.cu:939-1038.
#pragma once
#include "fb_fisher.cuh"

namespace fb {

struct PixState {
    float T_final;
    int   last;       // n_contrib of this pixel (0 if outside / nothing blended)
    float dpix[3];
    float bgdot;
};

struct BwdSmem {
    TileBatch sb;
    float acc[FB_TILE_PIX][FB_NACC];
    int   block_last;
};

// sorted: tile's depth-sorted keys (shared or global), n entries.  recs: this view's records.
// colors: [P,3].  flush(i_in_batch, gid, acc_row) is called by thread i for every instance of the
// batch after the tile-wide reduction.
template <bool P2, bool COLOR, class Flush>
__device__ __forceinline__ void tile_backward(BwdSmem& sm, const u64* sorted, int n, const GRec* __restrict__ recs,
                                              const float* __restrict__ colors, const PixState& ps, float pxf, float pyf,
                                              float halfW, float halfH, Flush flush) {
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid == 0) sm.block_last = 0;
    __syncthreads();
    const int warp_last = __reduce_max_sync(0xffffffffu, ps.last);
    if (lane == 0 && warp_last > 0) atomicMax(&sm.block_last, warp_last);
    __syncthreads();
    const int block_last = min(sm.block_last, n);

    float T = ps.T_final;
    float accum_rec[3] = {0.f, 0.f, 0.f};
    float last_alpha = 0.f;
    float last_color[3] = {0.f, 0.f, 0.f};
    const int slot = reduce16_slot(lane);

    for (int top = block_last; top > 0; top -= FB_TILE_PIX) {
        const int m = min(FB_TILE_PIX, top);
        __syncthreads();  // previous epilogue done with sb/acc
        if (tid < m) {
            unsigned gid = (unsigned)(sorted[top - 1 - tid] & 0xffffffffull);
            GRec r = recs[gid];
            sm.sb.xyo[tid] = r.a;
            sm.sb.con[tid] = r.b;
            sm.sb.gid[tid] = (int)gid;
            const float* c = colors + 3 * (size_t)gid;
            sm.sb.col[0][tid] = c[0]; sm.sb.col[1][tid] = c[1]; sm.sb.col[2][tid] = c[2];
        }
#pragma unroll
        for (int k = 0; k < FB_NACC; k += 4) *reinterpret_cast<float4*>(&sm.acc[tid][k]) = make_float4(0.f, 0.f, 0.f, 0.f);
        __syncthreads();

        // warp-uniform start: skip instances behind every pixel of this warp
        int j0 = max(0, top - warp_last);
        for (int j = j0; j < m; ++j) {
            const int pos = top - 1 - j;
            float v[FB_NACC];
#pragma unroll
            for (int k = 0; k < FB_NACC; ++k) v[k] = 0.f;
            bool hit = false;
            if (pos < ps.last) {
                const float4 xy = sm.sb.xyo[j];
                const float4 co = sm.sb.con[j];
                const float dx = xy.x - pxf, dy = xy.y - pyf;
                const float power = pair_power(co.x, co.y, co.z, dx, dy);
                if (power <= 0.0f) {
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, xy.w * G);
                    if (alpha >= 1.0f / 255.0f) {
                        hit = true;
                        T = T / (1.f - alpha);
                        const float dchannel_dcolor = alpha * T;
                        float dL_dalpha = 0.0f;
#pragma unroll
                        for (int ch = 0; ch < 3; ++ch) {
                            const float c = sm.sb.col[ch][j];
                            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                            last_color[ch] = c;
                            dL_dalpha += (c - accum_rec[ch]) * ps.dpix[ch];
                            if (COLOR) {
                                float cc = dchannel_dcolor * ps.dpix[ch];
                                v[13 + ch] = P2 ? cc * cc : cc;
                            }
                        }
                        dL_dalpha *= T;
                        last_alpha = alpha;
                        dL_dalpha += (-ps.T_final / (1.f - alpha)) * ps.bgdot;
                        const float s = G * (xy.w * dL_dalpha);
                        if (P2) {
                            const float q = s * s;
                            const float qx = q * dx, qy = q * dy;
                            const float m20 = qx * dx, m11 = qx * dy, m02 = qy * dy;
                            const float m30 = m20 * dx, m21 = m20 * dy, m12 = m02 * dx, m03 = m02 * dy;
                            v[0] = q; v[1] = m20; v[2] = m11; v[3] = m02;
                            v[4] = m30; v[5] = m21; v[6] = m12; v[7] = m03;
                            v[8] = m30 * dx; v[9] = m30 * dy; v[10] = m21 * dy; v[11] = m03 * dx; v[12] = m03 * dy;
                        } else {
                            const float sx = s * dx, sy = s * dy;
                            v[0] = s; v[1] = sx * dx; v[2] = sx * dy; v[3] = sy * dy; v[4] = sx; v[5] = sy;
                        }
                    }
                }
            }
            if (__any_sync(0xffffffffu, hit)) {
                float tot = warp_reduce16(v, lane);
                if ((lane & 1) == 0 && (COLOR || slot < 13) && tot != 0.f) atomicAdd(&sm.acc[j][slot], tot);
            }
        }
        __syncthreads();
        if (tid < m) flush(tid, sm.sb.gid[tid], sm.acc[tid]);
    }
}

}  // namespace fb
