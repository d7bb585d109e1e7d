// fb_mapnbr.cuh -- the per-step scatter kernels either side of the scoring path:
//   AstarPlanner.update_occ_map      planning/astar.py:202-303   depth rays splatted into the 2-D occupancy grid
//   count_gaussians_vs_mask          models/SLAM/gaussian_object.py:359-428   masked variant of the visibility counter
//
// update_occ_map in the reference: 11 depth planes per pixel are unprojected (10 "free" samples along the ray and the
// depth point itself), moved to the world with a [4,4] @ [4,N] matmul, discretised, counted per cell with
// torch.unique(dim=0, return_counts=True) (a sort of up to 11 * H * W rows, twice), then one cv2.line per occupied cell
// is drawn ON THE HOST from that cell to the camera cell (a .cpu() sync and a Python loop), and the three channels are
// normalised and added to the persistent map.  Here: one pass counts both classes with integer atomics (the counts are
// what torch.unique returns), one thread per occupied cell walks the line on the device, one pass fuses.  No sort, no
// host round trip.
// Arithmetic choices that the reference leaves to its libraries are fixed here (the CPU checker restates them):
//   * c2w @ p is evaluated as fma(m2, z, fma(m1, y, m0 * x)) + m3 (k-sequential, like the scoring path's transform);
//   * cv2.line(..., thickness 1, LINE_8) is OpenCV's 8-connected Bresenham (imgproc/src/drawing.cpp, LineIterator with
//     leftToRight = true as Line() constructs it): OpenCV is not in this image, so that part is "parity unpinned".
#pragma once
#include "fb_common.cuh"

namespace fb {

struct OccArgs {
    float* occ_map;          // [3][gz][gx] persistent map, updated in place
    int gx, gz;
    const float* depth;      // [H][W]
    int W, H, ds;
    const float* c2w;        // 16 floats, device
    const float* intr;       // fx, fy, cx, cy, device
    const float* planes;     // 11 floats: torch.linspace(1e-3, 0.95, 11) with the last one set to 1 (astar.py:237-239)
    const float* map_center; // 2 floats, device
    float cell, h_lo, h_hi, far;
    int* cnt_free; int* cnt_occ; unsigned char* canvas; int* cam_cell;   // workspace
};

__device__ __forceinline__ int occ_cell(float w, float centre, float cell, int dim) {
    // discretize_coords (datasets/util/map_utils.py:106-124): floor((w - c) / cell) + (dim - 1) / 2.0, .int(), clamp
    const float b = __fadd_rn(floorf(__fdiv_rn(__fsub_rn(w, centre), cell)), (float)(dim - 1) / 2.0f);
    const int i = (int)b;
    return min(dim - 1, max(0, i));
}

// thread 0 of block 0 also fixes the camera cell (astar.py:211-212), kept on the device
__global__ void __launch_bounds__(256) k_occ_points(OccArgs a) {
    const int Wd = (a.W + a.ds - 1) / a.ds, Hd = (a.H + a.ds - 1) / a.ds;
    const long long n = (long long)Wd * Hd * 11;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        const float cx = __fadd_rn(__fdiv_rn(__fsub_rn(a.c2w[3], a.map_center[0]), a.cell), (float)(a.gx / 2));
        const float cz = __fadd_rn(__fdiv_rn(__fsub_rn(a.c2w[11], a.map_center[1]), a.cell), (float)(a.gz / 2));
        a.cam_cell[0] = (int)cx; a.cam_cell[1] = (int)cz;
    }
    if (i >= n) return;
    const int k = (int)(i / ((long long)Wd * Hd));
    const int r = (int)(i % ((long long)Wd * Hd));
    const int px = (r % Wd) * a.ds, py = (r / Wd) * a.ds;
    const float fx = a.intr[0], fy = a.intr[1], cx = a.intr[2], cy = a.intr[3];
    const float xx = __fdiv_rn(__fsub_rn((float)px, cx), fx), yy = __fdiv_rn(__fsub_rn((float)py, cy), fy);
    const float dz = __fmul_rn(a.planes[k], a.depth[(size_t)py * a.W + px]);
    if (!(dz > 0.f && dz < a.far)) return;                                      // astar.py:243
    const float x = __fmul_rn(xx, dz), y = __fmul_rn(yy, dz);
    const float* m = a.c2w;
    const float wx = __fadd_rn(__fmaf_rn(m[2], dz, __fmaf_rn(m[1], y, __fmul_rn(m[0], x))), m[3]);
    const float wy = __fadd_rn(__fmaf_rn(m[6], dz, __fmaf_rn(m[5], y, __fmul_rn(m[4], x))), m[7]);
    const float wz = __fadd_rn(__fmaf_rn(m[10], dz, __fmaf_rn(m[9], y, __fmul_rn(m[8], x))), m[11]);
    if (!(wy >= a.h_lo && wy <= a.h_hi)) return;                                // floor and ceiling removed, :264 / :281
    const int gx = occ_cell(wx, a.map_center[0], a.cell, a.gx), gz = occ_cell(wz, a.map_center[1], a.cell, a.gz);
    atomicAdd((k < 10 ? a.cnt_free : a.cnt_occ) + (size_t)gz * a.gx + gx, 1);
}

// One thread per occupied cell: cv2.line(canvas, (x, z), (cam_x, cam_z), 1, 1) (astar.py:296-299).
// OpenCV LineIterator, 8-connected, leftToRight: start from the endpoint with the smaller x; dx, dy made positive
// (the step signs keep the direction); the major axis is the longer one; err = dx - 2 dy, and at each of the
// dx + 1 points: mask = err < 0; err += -2 dy + (2 dx & mask); minor step taken when mask.
__global__ void __launch_bounds__(256) k_occ_lines(OccArgs a) {
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= a.gx * a.gz || a.cnt_occ[cell] == 0) return;
    int x1 = cell % a.gx, y1 = cell / a.gx, x2 = a.cam_cell[0], y2 = a.cam_cell[1];
    // clipLine to the canvas: the occupied cell is inside by construction; clamp the camera cell (a camera outside the
    // grid is outside the reference's use: its own 3x3 write at :214 would wrap)
    x2 = min(a.gx - 1, max(0, x2)); y2 = min(a.gz - 1, max(0, y2));
    if (x2 < x1) { int t = x1; x1 = x2; x2 = t; t = y1; y1 = y2; y2 = t; }       // leftToRight
    int dx = x2 - x1, dy = y2 - y1;
    const int ystep = dy < 0 ? -1 : 1;
    dy = dy < 0 ? -dy : dy;
    const bool steep = dy > dx;
    const int major = steep ? dy : dx, minor = steep ? dx : dy;
    int err = major - (minor + minor);
    const int plus = major + major, minus = -(minor + minor);
    int x = x1, y = y1;
    for (int i = 0; i <= major; ++i) {
        a.canvas[(size_t)y * a.gx + x] = 1;
        const bool m = err < 0;
        err += minus + (m ? plus : 0);
        if (steep) { y += ystep; if (m) x += 1; }
        else { x += 1; if (m) y += ystep; }
    }
}

// occ_map += tmp / (tmp.sum(0) + 1e-5) with tmp = [0, 100 (n_occ + 1e-5), 0.01 (n_free + 1e-5) or 1 on a ray]
// (astar.py:267-270, 284-303), after the 3x3 block around the camera was set to 1e3 on the free channel (:214).
__global__ void __launch_bounds__(256) k_occ_fuse(OccArgs a) {
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = a.gx * a.gz;
    if (cell >= n) return;
    const int x = cell % a.gx, z = cell / a.gx;
    const int nf = a.cnt_free[cell], no = a.cnt_occ[cell];
    float t2 = nf > 0 ? __fmul_rn(0.01f, __fadd_rn((float)nf, 1e-5f)) : 0.f;
    const float t1 = no > 0 ? __fmul_rn(__fadd_rn((float)no, 1e-5f), 100.f) : 0.f;
    if (a.canvas[cell]) t2 = 1.f;
    const float den = __fadd_rn(__fadd_rn(__fadd_rn(0.f, t1), t2), 1e-5f);
    float* p0 = a.occ_map + cell; float* p1 = p0 + n; float* p2 = p1 + n;
    const int cx = a.cam_cell[0], cz = a.cam_cell[1];
    float base2 = *p2;
    if (x >= cx - 1 && x <= cx + 1 && z >= cz - 1 && z <= cz + 1) base2 = 1e3f;
    *p0 = __fadd_rn(*p0, __fdiv_rn(0.f, den));
    *p1 = __fadd_rn(*p1, __fdiv_rn(t1, den));
    *p2 = __fadd_rn(base2, __fdiv_rn(t2, den));
}

// count_gaussians_vs_mask: code[i] = 0 inactive (sigmoid(logit) < thresh), 1 projects inside the mask, 2 visible in the
// image but outside the mask, 3 active but behind the camera / outside the image; counts = {in, out, active}.
__global__ void __launch_bounds__(256) k_count_vs_mask(int N, const float* __restrict__ means, const float* __restrict__ logit,
                                                      const float* __restrict__ w2c, const float* __restrict__ intr,
                                                      const unsigned char* __restrict__ mask, int W, int H, float thresh,
                                                      unsigned char* __restrict__ code, int* __restrict__ counts) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    int c = 0;
    if (i < N) {
        const float alpha = __fdiv_rn(1.f, __fadd_rn(1.f, expf(-logit[i])));                 // torch.sigmoid
        if (alpha >= thresh) {
            const float x = means[3 * (size_t)i], y = means[3 * (size_t)i + 1], z = means[3 * (size_t)i + 2];
            const float* m = w2c;
            const float xc = __fadd_rn(__fmaf_rn(m[2], z, __fmaf_rn(m[1], y, __fmul_rn(m[0], x))), m[3]);
            const float yc = __fadd_rn(__fmaf_rn(m[6], z, __fmaf_rn(m[5], y, __fmul_rn(m[4], x))), m[7]);
            const float zc = __fadd_rn(__fmaf_rn(m[10], z, __fmaf_rn(m[9], y, __fmul_rn(m[8], x))), m[11]);
            const float zz = fmaxf(zc, 1e-6f);                                                // clamp_min(1e-6), :405
            const float u = __fadd_rn(__fmul_rn(intr[0], __fdiv_rn(xc, zz)), intr[2]);
            const float v = __fadd_rn(__fmul_rn(intr[1], __fdiv_rn(yc, zz)), intr[3]);
            const bool ok = zc > 0.f && u >= 0.f && u < (float)W && v >= 0.f && v < (float)H;
            c = 3;
            if (ok) {
                const int iu = min(W - 1, max(0, (int)rintf(u))), iv = min(H - 1, max(0, (int)rintf(v)));   // round half to even
                c = mask[(size_t)iv * W + iu] ? 1 : 2;
            }
            code[i] = (unsigned char)c;
        } else {
            code[i] = 0;
        }
    }
    const unsigned lane = threadIdx.x & 31;
    const unsigned b_in = __ballot_sync(0xffffffffu, c == 1), b_act = __ballot_sync(0xffffffffu, c != 0);
    if (lane == 0) {
        if (b_in) atomicAdd(&counts[0], __popc(b_in));
        if (b_act) { atomicAdd(&counts[2], __popc(b_act)); atomicAdd(&counts[1], __popc(b_act) - __popc(b_in)); }
    }
}

}  // namespace fb
