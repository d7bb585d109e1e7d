// fb_mapupdate.cuh -- map update, densification half: un-project the masked pixels of an RGB-D frame into new
// Gaussians, written IN PLACE at the tail of the resident parameter arrays (first piece).
//
// Reference: models/SLAM/gaussian.py
//   get_pointcloud          :75-142    pixel grid -> camera points -> world (c2w @ pts4), colours, projective
//                                      mean_sq_dist = (downsample * z / ((FX + FY) / 2))^2, boolean-mask selection
//   initialize_new_params   :299-318   means3D, rgb_colors, unnorm_rotations = (1,0,0,0), logit_opacities = 0,
//                                      log_scales = log(sqrt(mean_sq_dist)) tiled to 3 (or 1 if isotropic)
//   add_new_gaussians       :399-401   params[k] = cat(params[k], new[k])  -- a copy of the WHOLE map per frame;
//                                      here the arrays have capacity and only the new rows are written.
// The reference builds ~12 full-frame temporaries (grids, pts4, ones, the transposed product, the [HW,6] cat) and
// then selects; this kernel reads depth + colour of the selected pixels once and writes 44-56 B per new Gaussian.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace fb {

struct UnprojectArgs {
    int W, H, ds, Wd;                 // frame size, downsample, downsampled row length
    const float* intr;                // device {fx, fy, cx, cy}
    const float* c2w;                 // device, 16 floats row-major (camera-to-world); rows 0-2 are used
    const float* color;               // [3,H,W]
    const float* depth;               // [1,H,W]
    const long long* pix;             // [n] indices into the downsampled grid (ascending), or null = all pixels
    long long n;
    float* means3D; int mean_stride;  // row stride in floats (3 for the map arrays, 6 for get_pointcloud's [n,6])
    float* rgb;     int rgb_stride;
    float* rots;                      // [n,4] or null
    float* logit_op;                  // [n]   or null
    float* log_scales; int ls_cols;   // [n,ls_cols] (3 or 1) or null
    float* msd;                       // [n] or null
};

__global__ void __launch_bounds__(256) k_unproject_append(UnprojectArgs a) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const float fx = __ldg(a.intr), fy = __ldg(a.intr + 1), cx = __ldg(a.intr + 2), cy = __ldg(a.intr + 3);
    const long long p = a.pix ? a.pix[i] : i;
    const int gy = (int)(p / a.Wd), gx = (int)(p - (long long)gy * a.Wd);
    const int x = gx * a.ds, y = gy * a.ds;
    const size_t src = (size_t)y * a.W + x, HW = (size_t)a.W * a.H;
    const float z = __ldg(a.depth + src);
    const float xx = __fdiv_rn(__fsub_rn((float)x, cx), fx);        // (x_grid - CX) / FX
    const float yy = __fdiv_rn(__fsub_rn((float)y, cy), fy);
    const float pcx = __fmul_rn(xx, z), pcy = __fmul_rn(yy, z);
    float* m = a.means3D + (size_t)i * a.mean_stride;
#pragma unroll
    for (int r = 0; r < 3; ++r)                                         // (c2w @ [pc, 1])[r]
        m[r] = fmaf(__ldg(a.c2w + 4 * r + 2), z, fmaf(__ldg(a.c2w + 4 * r + 1), pcy, __ldg(a.c2w + 4 * r) * pcx)) +
               __ldg(a.c2w + 4 * r + 3);
    float* c = a.rgb + (size_t)i * a.rgb_stride;
#pragma unroll
    for (int k = 0; k < 3; ++k) c[k] = __ldg(a.color + k * HW + src);
    if (a.rots) reinterpret_cast<float4*>(a.rots)[i] = make_float4(1.f, 0.f, 0.f, 0.f);
    if (a.logit_op) a.logit_op[i] = 0.f;
    if (a.msd || a.log_scales) {
        const float half_f = __fmul_rn(__fadd_rn(fx, fy), 0.5f);    // (FX + FY) / 2
        const float sg = __fdiv_rn(__fmul_rn((float)a.ds, z), half_f);  // downsample * depth_z / ((FX + FY) / 2)
        const float msd = __fmul_rn(sg, sg);
        if (a.msd) a.msd[i] = msd;
        if (a.log_scales) {
            const float ls = logf(sqrtf(msd));
            for (int k = 0; k < a.ls_cols; ++k) a.log_scales[(size_t)i * a.ls_cols + k] = ls;
        }
    }
}

// 30-bit Morton code of each mean inside the bounding box [lo, lo + ext] (10 bits per axis): the key the resident map
// is sorted by (one kernel instead of ~35 elementwise launches; the sort itself is a library radix sort).
__device__ __forceinline__ unsigned spread10(unsigned v) {
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
__global__ void __launch_bounds__(256) k_morton30(int P, const float* __restrict__ means, const float* __restrict__ lo,
                                                  const float* __restrict__ ext, int* __restrict__ codes) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    unsigned q[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float e = fmaxf(ext[a], 1e-12f);
        const float t = (means[3 * (size_t)i + a] - lo[a]) / e * 1023.0f;
        q[a] = (unsigned)min(1023, max(0, (int)t));                      // NaN -> 0
    }
    codes[i] = (int)(spread10(q[0]) | (spread10(q[1]) << 1) | (spread10(q[2]) << 2));
}

}  // namespace fb
