// fb_stage5.cuh -- stage 5: distCUDA2 (exact 3-NN mean squared distance) and the visibility counter.
//
// distCUDA2 replaces SimpleKNN::knn  KNN/simple_knn.cu:186-222 (Morton sort + 1024-point boxes +
// pruned brute force, :148-184).  Result definition kept bit-for-bit: for point i the three
// smallest values of  dx*dx + dy*dy + dz*dz  over all j != i (same expression shape, hence the
// same FMA contraction), then (b0 + b1 + b2) / 3.0f with b ascending (:183).  The three smallest
// values are a property of the point set, not of the search order, so any EXACT search yields the
// same floats.  B200 shape: counting-sort the points into a uniform grid (~2 cells per point),
// then each query walks outward ring by ring and stops as soon as its third-best distance is
// provably inside the searched block.  No host sync, no allocation, caller-provided workspace.
//
// fb_count_visible replaces count_visible_points  tester_navigator.py:147-180 (numpy, float64).
#pragma once
#include "fb_common.cuh"
#include <float.h>

namespace fb {

struct KnnGrid {
    float minx, miny, minz;
    float inv_h, h;
    int   nx, ny, nz;
    unsigned ncells;
};

struct KnnWs {
    unsigned* bbox;      // [6] ordered-int encoded min xyz, max xyz
    KnnGrid*  grid;
    unsigned* cell;      // [cap+1] counts -> starts
    unsigned* cursor;    // [cap]
    unsigned* blocksum;  // [scan blocks]
    float4*   sorted;    // [P] x,y,z,idx
    unsigned  cap;
};

__device__ __forceinline__ unsigned f2ord(float f) { unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float ord2f(unsigned u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

__global__ void k_knn_init(KnnWs ws) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 3) ws.bbox[i] = 0xffffffffu;
    else if (i < 6) ws.bbox[i] = 0u;
    for (unsigned c = i; c <= ws.cap; c += gridDim.x * blockDim.x) { ws.cell[c] = 0u; if (c < ws.cap) ws.cursor[c] = 0u; }
}

__global__ void __launch_bounds__(256) k_knn_bbox(int P, const float* __restrict__ pts, KnnWs ws) {
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            // non-finite coordinates (a diverged map) stay out of the bounding box; such points land in a border cell
            float c = pts[3 * (size_t)i + a];
            if (fabsf(c) <= FLT_MAX) { mn[a] = fminf(mn[a], c); mx[a] = fmaxf(mx[a], c); }
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { atomicMin(&ws.bbox[a], f2ord(mn[a])); atomicMax(&ws.bbox[3 + a], f2ord(mx[a])); }
    }
}

// Grid sizing on the device: cell edge from the bounding volume so that there are about two
// cells per point, then coarsened until the cell count fits the workspace.
__global__ void k_knn_setup(int P, KnnWs ws) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float mn[3], ex[3];
    float emax = 0.f;
    for (int a = 0; a < 3; ++a) {
        mn[a] = ord2f(ws.bbox[a]); ex[a] = ord2f(ws.bbox[3 + a]) - mn[a];
        if (!(ex[a] >= 0.f) || !(ex[a] <= FLT_MAX)) { ex[a] = 0.f; if (!(fabsf(mn[a]) <= FLT_MAX)) mn[a] = 0.f; }   // no finite point on this axis
        emax = fmaxf(emax, ex[a]);
    }
    if (!(emax > 0.f)) emax = 1.f;
    double vol = 1.0;
    for (int a = 0; a < 3; ++a) vol *= (double)fmaxf(ex[a], 1e-3f * emax);
    double h = cbrt(vol / (2.0 * (double)(P > 0 ? P : 1)));
    int n[3] = {1, 1, 1};
    bool fits = false;
    for (int it = 0; it < 96 && !fits; ++it) {         // bounded: h grows 2x every 3 rounds, 2^32 over 96
        double prod = 1.0;
        for (int a = 0; a < 3; ++a) { double d = floor((double)ex[a] / h) + 1.0; if (!(d <= 2048.0)) d = 2048.0; n[a] = (int)d; prod *= d; }
        fits = prod <= (double)ws.cap;
        if (!fits) h *= 1.2599210498948732;
    }
    if (!fits || !(h > 0.0) || !(h <= 3.0e38)) { n[0] = n[1] = n[2] = 1; h = (double)emax > 0.0 ? (double)emax : 1.0; }   // one cell: brute force, still exact
    KnnGrid g;
    g.minx = mn[0]; g.miny = mn[1]; g.minz = mn[2];
    g.h = (float)h; g.inv_h = (float)(1.0 / h);
    g.nx = n[0]; g.ny = n[1]; g.nz = n[2];
    g.ncells = (unsigned)n[0] * (unsigned)n[1] * (unsigned)n[2];
    *ws.grid = g;
}

__device__ __forceinline__ void knn_cell_of(const KnnGrid& g, float x, float y, float z, int& cx, int& cy, int& cz) {
    cx = min(g.nx - 1, max(0, (int)((x - g.minx) * g.inv_h)));
    cy = min(g.ny - 1, max(0, (int)((y - g.miny) * g.inv_h)));
    cz = min(g.nz - 1, max(0, (int)((z - g.minz) * g.inv_h)));
}

__global__ void __launch_bounds__(256) k_knn_count(int P, const float* __restrict__ pts, KnnWs ws) {
    const KnnGrid g = *ws.grid;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    int cx, cy, cz;
    knn_cell_of(g, pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2], cx, cy, cz);
    atomicAdd(&ws.cell[((size_t)cz * g.ny + cy) * g.nx + cx], 1u);
}

// Exclusive scan of cell[0..cap] in three kernels (8192 entries per CTA).
#define FB_SCAN_ITEMS 8
__global__ void __launch_bounds__(1024) k_scan_block(unsigned* data, unsigned n, unsigned* blocksum) {
    __shared__ unsigned ws_[32];
    const unsigned base = blockIdx.x * 1024u * FB_SCAN_ITEMS + threadIdx.x * FB_SCAN_ITEMS;
    unsigned v[FB_SCAN_ITEMS], sum = 0;
#pragma unroll
    for (int k = 0; k < FB_SCAN_ITEMS; ++k) { v[k] = (base + k < n) ? data[base + k] : 0u; sum += v[k]; }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned x = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) ws_[wid] = x;
    __syncthreads();
    if (wid == 0) {
        unsigned w = ws_[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
        ws_[lane] = w;
    }
    __syncthreads();
    unsigned excl = x - sum + (wid > 0 ? ws_[wid - 1] : 0u);
#pragma unroll
    for (int k = 0; k < FB_SCAN_ITEMS; ++k) { if (base + k < n) data[base + k] = excl; excl += v[k]; }
    if (threadIdx.x == 1023) blocksum[blockIdx.x] = excl;
}
__global__ void __launch_bounds__(1024) k_scan_sums(unsigned* blocksum, unsigned nb) {
    __shared__ unsigned ws_[32];
    __shared__ unsigned carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (unsigned base = 0; base < nb; base += 1024) {
        unsigned i = base + threadIdx.x;
        unsigned c = i < nb ? blocksum[i] : 0u, x = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) ws_[wid] = x;
        __syncthreads();
        if (wid == 0) {
            unsigned w = ws_[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
            ws_[lane] = w;
        }
        __syncthreads();
        unsigned incl = x + (wid > 0 ? ws_[wid - 1] : 0u) + carry;
        if (i < nb) blocksum[i] = incl - c;
        __syncthreads();
        if (threadIdx.x == 1023) carry = incl;
        __syncthreads();
    }
}
__global__ void __launch_bounds__(1024) k_scan_add(unsigned* data, unsigned n, const unsigned* blocksum) {
    const unsigned base = blockIdx.x * 1024u * FB_SCAN_ITEMS + threadIdx.x * FB_SCAN_ITEMS;
    const unsigned add = blocksum[blockIdx.x];
#pragma unroll
    for (int k = 0; k < FB_SCAN_ITEMS; ++k) if (base + k < n) data[base + k] += add;
}

__global__ void __launch_bounds__(256) k_knn_scatter(int P, const float* __restrict__ pts, KnnWs ws) {
    const KnnGrid g = *ws.grid;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    float x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
    int cx, cy, cz;
    knn_cell_of(g, x, y, z, cx, cy, cz);
    size_t c = ((size_t)cz * g.ny + cy) * g.nx + cx;
    unsigned slot = ws.cell[c] + atomicAdd(&ws.cursor[c], 1u);
    ws.sorted[slot] = make_float4(x, y, z, __int_as_float(i));
}

// updateKBest<3>  (simple_knn.cu:132-146)
__device__ __forceinline__ void k3_update(float dist, float* best) {
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        if (best[j] > dist) { float t = best[j]; best[j] = dist; dist = t; }
    }
}

__global__ void __launch_bounds__(128) k_knn_query(int P, KnnWs ws, float* __restrict__ out) {
    const KnnGrid g = *ws.grid;
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P) return;
    const float4 me = ws.sorted[s];
    const int my_idx = __float_as_int(me.w);
    int cx, cy, cz;
    knn_cell_of(g, me.x, me.y, me.z, cx, cy, cz);
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    const int rmax = max(g.nx, max(g.ny, g.nz));
    for (int r = 0; r <= rmax; ++r) {
        const int z0 = max(0, cz - r), z1 = min(g.nz - 1, cz + r);
        const int y0 = max(0, cy - r), y1 = min(g.ny - 1, cy + r);
        const int x0 = max(0, cx - r), x1 = min(g.nx - 1, cx + r);
        for (int z = z0; z <= z1; ++z) {
            const bool zshell = (z == cz - r) || (z == cz + r);
            for (int y = y0; y <= y1; ++y) {
                const bool yshell = zshell || (y == cy - r) || (y == cy + r);
                const size_t row = ((size_t)z * g.ny + y) * g.nx;
                if (yshell) {
                    // whole x-run of this row is on the shell: cells are contiguous in memory
                    const unsigned b = ws.cell[row + x0], e = ws.cell[row + x1 + 1];
                    for (unsigned k = b; k < e; ++k) {
                        const float4 o = ws.sorted[k];
                        if (__float_as_int(o.w) == my_idx) continue;
                        const float dx = o.x - me.x, dy = o.y - me.y, dz = o.z - me.z;
                        k3_update(dot3(dx, dx, dy, dy, dz, dz), best);
                    }
                } else {
                    // interior row: only the two end cells belong to the shell
                    for (int side = 0; side < 2; ++side) {
                        const int x = side ? cx + r : cx - r;
                        if (x < 0 || x >= g.nx || (side && r == 0)) continue;
                        const unsigned b = ws.cell[row + x], e = ws.cell[row + x + 1];
                        for (unsigned k = b; k < e; ++k) {
                            const float4 o = ws.sorted[k];
                            if (__float_as_int(o.w) == my_idx) continue;
                            const float dx = o.x - me.x, dy = o.y - me.y, dz = o.z - me.z;
                            k3_update(dot3(dx, dx, dy, dy, dz, dz), best);
                        }
                    }
                }
            }
        }
        // distance from the query to the nearest face of the searched block that has cells beyond it
        float face = FLT_MAX;
        const float margin = 1e-3f * g.h;
        if (cx - r > 0)        face = fminf(face, me.x - (g.minx + (float)(cx - r) * g.h));
        if (cx + r < g.nx - 1) face = fminf(face, (g.minx + (float)(cx + r + 1) * g.h) - me.x);
        if (cy - r > 0)        face = fminf(face, me.y - (g.miny + (float)(cy - r) * g.h));
        if (cy + r < g.ny - 1) face = fminf(face, (g.miny + (float)(cy + r + 1) * g.h) - me.y);
        if (cz - r > 0)        face = fminf(face, me.z - (g.minz + (float)(cz - r) * g.h));
        if (cz + r < g.nz - 1) face = fminf(face, (g.minz + (float)(cz + r + 1) * g.h) - me.z);
        if (face == FLT_MAX) break;  // block covers the whole grid
        face = fmaxf(0.f, face - margin);
        if (best[2] <= face * face) break;
    }
    out[my_idx] = (best[0] + best[1] + best[2]) / 3.0f;
}

// ---- visibility counter ----------------------------------------------------------------------
// counts[k] = #{i : z>0, 0<=u<W, 0<=v<H} in FP64, u = fx*x/z + cx  (tester_navigator.py:160-178).
// One pass over the points serves all K poses (pose constants in shared memory).
__global__ void __launch_bounds__(256) k_count_visible(int N, const float* __restrict__ pts, int K,
                                                      const double* __restrict__ w2c, const double* __restrict__ intr,
                                                      int img_w, int img_h, int* __restrict__ counts,
                                                      uint8_t* __restrict__ mask) {
    extern __shared__ double s_pose[];  // K*12 doubles, then K ints
    int* s_cnt = reinterpret_cast<int*>(s_pose + (size_t)K * 12);
    for (int i = threadIdx.x; i < K * 12; i += blockDim.x) s_pose[i] = w2c[(i / 12) * 16 + (i % 12)];
    for (int i = threadIdx.x; i < K; i += blockDim.x) s_cnt[i] = 0;
    __syncthreads();
    const double fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
    const double Wd = (double)img_w, Hd = (double)img_h;
    const int lane = threadIdx.x & 31;
    for (int base = blockIdx.x * blockDim.x; base < N; base += gridDim.x * blockDim.x) {
        const int i = base + threadIdx.x;
        const bool live = i < N;
        double x = 0, y = 0, z = 0;
        if (live) { x = (double)pts[3 * (size_t)i]; y = (double)pts[3 * (size_t)i + 1]; z = (double)pts[3 * (size_t)i + 2]; }
        for (int k = 0; k < K; ++k) {
            const double* m = s_pose + k * 12;
            const double xc = __dadd_rn(__fma_rn(m[2], z, __fma_rn(m[1], y, __dmul_rn(m[0], x))), m[3]);
            const double yc = __dadd_rn(__fma_rn(m[6], z, __fma_rn(m[5], y, __dmul_rn(m[4], x))), m[7]);
            const double zc = __dadd_rn(__fma_rn(m[10], z, __fma_rn(m[9], y, __dmul_rn(m[8], x))), m[11]);
            bool vis = false;
            if (live && zc > 0.0) {
                const double u = __dadd_rn(__ddiv_rn(__dmul_rn(fx, xc), zc), cx);
                const double v = __dadd_rn(__ddiv_rn(__dmul_rn(fy, yc), zc), cy);
                vis = (u >= 0.0) && (u < Wd) && (v >= 0.0) && (v < Hd);
            }
            if (mask != nullptr && live) mask[(size_t)k * N + i] = vis ? 1 : 0;
            const unsigned b = __ballot_sync(0xffffffffu, vis);
            if (lane == 0 && b) atomicAdd(&s_cnt[k], __popc(b));
        }
    }
    __syncthreads();
    for (int k = threadIdx.x; k < K; k += blockDim.x)
        if (s_cnt[k]) atomicAdd(&counts[k], s_cnt[k]);
}

}  // namespace fb
