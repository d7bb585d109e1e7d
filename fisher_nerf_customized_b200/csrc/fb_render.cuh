// fb_render.cuh -- stage 3: per-tile front-to-back alpha blend (drop-in single-view path).
//
// Replaces renderCUDA<3>  DGR/cuda_rasterizer/forward.cu:261-393.  Blend rules reproduced
// (SURVEY 8.1 #9): skip power>0; alpha=min(0.99,o*exp(power)); skip alpha<1/255; stop when
// T*(1-alpha)<1e-4 BEFORE blending that Gaussian; n_contrib = list position of the last blended
// Gaussian; median depth (default 15.0) latched when T crosses 0.5.
#pragma once
#include "fb_layout.cuh"

namespace fb {

struct TileBatch {
    float4 xyo[FB_TILE_PIX];   // px, py, depth, opacity
    float4 con[FB_TILE_PIX];   // conic a,b,c, (unused)
    float  col[3][FB_TILE_PIX];
    int    gid[FB_TILE_PIX];
};

__global__ void __launch_bounds__(FB_TILE_PIX) k_render_fwd(Bins bins, Cam cam, const float* __restrict__ features,
                                                           float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                                                           float* __restrict__ out_color, float* __restrict__ out_depth) {
    __shared__ TileBatch sb;
    // blockIdx.z = view of a batch (per-view strides P, HW); the single-view drop-in launches with gridDim.z = 1
    int tx = blockIdx.x, ty = blockIdx.y, v = blockIdx.z;
    if (bins.tile_order) {            // batched: heaviest (view, tile) slots first
        const unsigned slot = bins.tile_order[(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x];
        v = (int)(slot / (unsigned)bins.tiles);
        const int t = (int)(slot % (unsigned)bins.tiles);
        tx = t % cam.tiles_x; ty = t / cam.tiles_x;
    }
    {
        const size_t HWv = (size_t)cam.W * cam.H * v;
        final_T += HWv; n_contrib += HWv; out_color += 3 * HWv; out_depth += HWv;
        bins.recs += (size_t)v * bins.P;
        bins.tile_range += (size_t)v * bins.tiles;
    }
    const int lx = threadIdx.x % FB_TILE_X, ly = threadIdx.x / FB_TILE_X;
    const int px = tx * FB_TILE_X + lx, py = ty * FB_TILE_Y + ly;
    const int W = cam.W, H = cam.H;
    const bool inside = px < W && py < H;
    const int pix_id = W * py + px;
    const float pxf = (float)px, pyf = (float)py;
    bool done = !inside;
    uint2 range = bins.total[1] ? make_uint2(0u, 0u) : bins.tile_range[ty * cam.tiles_x + tx];
    const int n = (int)(range.y - range.x);
    const u64* seg = bins.keys + range.x;

    float T = 1.0f;
    uint32_t contributor = 0, last_contributor = 0;
    float C[3] = {0.f, 0.f, 0.f};
    float D = 15.0f;
    for (int base = 0; base < n; base += FB_TILE_PIX) {
        if (__syncthreads_count(done) == FB_TILE_PIX) break;
        int i = base + threadIdx.x;
        if (i < n) {
            unsigned gid = (unsigned)(seg[i] & 0xffffffffull);
            GRec r = bins.recs[gid];
            sb.xyo[threadIdx.x] = r.a;
            sb.con[threadIdx.x] = r.b;
            sb.gid[threadIdx.x] = (int)gid;
        }
        __syncthreads();
        const int m = min(FB_TILE_PIX, n - base);
        for (int j = 0; !done && j < m; ++j) {
            contributor++;
            float4 xy = sb.xyo[j];
            float4 co = sb.con[j];
            float dx = xy.x - pxf, dy = xy.y - pyf;
            float power = pair_power(co.x, co.y, co.z, dx, dy);
            if (power > 0.0f) continue;
            float alpha = fminf(0.99f, xy.w * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            float test_T = __fmul_rn(T, __fsub_rn(1.f, alpha));
            if (test_T < 0.0001f) { done = true; continue; }
            const float* f = features + 3 * (size_t)sb.gid[j];
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) C[ch] = __fmaf_rn(T, __fmul_rn(alpha, f[ch]), C[ch]);   // as compiled
            if (T > 0.5f && (double)test_T < 0.5) D = xy.z;
            T = test_T;
            last_contributor = contributor;
        }
    }
    if (inside) {
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) out_color[(size_t)ch * H * W + pix_id] = __fmaf_rn(T, cam.bg[ch], C[ch]);
        out_depth[pix_id] = D;
    }
}

// present[i] = (view * p_i).z > 0.001   (checkFrustum, rasterizer_impl.cu:54-66; the projection
// is computed by the reference but does not influence the result).
__global__ void k_mark_visible(int P, const float* __restrict__ means, Cam cam, uint8_t* __restrict__ present) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    float3 p = make_float3(means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
    float3 pv = xform4x3(p, cam.view);
    present[idx] = pv.z <= 0.001f ? 0 : 1;
}

}  // namespace fb
