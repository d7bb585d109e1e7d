// fb_blocks.cuh -- POp-GS block form: per-Gaussian d x d blocks of J^T J and their T-/D-optimality scores.
//
// Reference: models/SLAM/gaussian_object.py
//   estimate_block_JtJ   :2112-2176   H_blocks[i] += Gv[i] (x) Gv[i] per probe, / K at the end
//   t_opt_blocks         :1720-1724   -sum_b trace( (Hm_b + J_b + lam I)^-1 )
//   d_opt_blocks         :1726-1732   sum_b logabsdet(Hm_b + lam I + J_b) - logabsdet(Hm_b + lam I)
// The block row of a visible Gaussian is g = [d means3D (3) | d opacity (1)? | d rotations (4)? | d scales (3)?],
// d in {3,4,6,7,8,10,11}.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace fb {

struct BlockCols {
    const float* mean;    // [P,3]
    const float* op;      // [P,1] or null
    const float* rot;     // [P,4] or null
    const float* sc;      // [P,3] or null
};

// k_block_outer: thread = one element of one block, blocks laid out [Nv, d, d] row-major so that the read-modify-
// write of H is fully coalesced (2 * 4 * d^2 bytes per visible Gaussian and probe: the kernel is HBM-bound); the
// Gaussian's gradient row (<= 44 B) is gathered through L1.  The product is rounded before the add, as
// `H_blocks += Gv.unsqueeze(2) * Gv.unsqueeze(1)` does (:2170); with `div` != 0 the element is scaled by fl(1/div)
// after the add: `H_blocks / float(K)` (:2174) on a CUDA tensor is evaluated by torch as a multiplication with the
// rounded reciprocal of the scalar, and that is the rounding reproduced here (folded into the last probe's launch).
__device__ __forceinline__ float block_col(const BlockCols& g, long long row, int c, int o_op, int o_rot, int o_sc) {
    if (c < 3) return __ldg(g.mean + row * 3 + c);
    if (c < o_rot) return __ldg(g.op + row);                 // only reached when op is present (o_op == 3 < o_rot)
    if (c < o_sc) return __ldg(g.rot + row * 4 + (c - o_rot));
    return __ldg(g.sc + row * 3 + (c - o_sc));
}
__global__ void __launch_bounds__(256) k_block_outer(long long n_elem, int d, const long long* __restrict__ rows,
                                                     BlockCols g, float div, float* __restrict__ H) {
    const int o_op = 3, o_rot = 3 + (g.op ? 1 : 0), o_sc = o_rot + (g.rot ? 4 : 0);
    const int dd = d * d;
    const float inv_div = div != 0.f ? __fdiv_rn(1.f, div) : 1.f;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < n_elem;
         e += (long long)gridDim.x * blockDim.x) {
        const long long i = e / dd;
        const int k = (int)(e - i * dd), r = k / d, c = k - r * d;
        const long long row = rows[i];
        const float gr = block_col(g, row, r, o_op, o_rot, o_sc), gc = block_col(g, row, c, o_op, o_rot, o_sc);
        float h = __fadd_rn(H[e], __fmul_rn(gr, gc));
        if (div != 0.f) h = __fmul_rn(h, inv_div);
        H[e] = h;
    }
}

// k_block_opt: thread = one block.  The matrix M = (Hm + J) + lam I (T-opt, :1722) or Hm + lam I and (Hm + lam I) + J
// (D-opt, :1728-1729) is formed in FP32 exactly as the reference forms it, then factorised by LU with partial
// pivoting -- the algorithm behind torch.linalg.inv / slogdet -- in FP64, in the thread's own column of shared
// memory (element e of thread t at sm[e * blockDim.x + t]: conflict-free).  FP64 because the blocks are sums of a
// handful of outer products plus lam = 1e-6: condition numbers reach 1e8, beyond what an FP32 factorisation
// resolves (the reference's FP32 result is then rounding noise around the value this kernel returns).
//   T-opt: value = -trace(M^-1), from d solves with the factors.     D-opt: value = log|det M1| - log|det M0|.
// One double per CTA leaves (warp-shuffle sum in lane order, then warps in order): the host adds the CTA partials.
#define FB_BLOCK_OPT_THREADS 32
template <bool DOPT>
__global__ void __launch_bounds__(FB_BLOCK_OPT_THREADS) k_block_opt(int nb, int d, const float* __restrict__ Hm,
                                                                    const long long* __restrict__ idx_m,
                                                                    const float* __restrict__ J,
                                                                    const long long* __restrict__ idx_j, float lam,
                                                                    double* __restrict__ partial,
                                                                    float* __restrict__ per_block) {
    extern __shared__ double sm_blk[];
    const int T = FB_BLOCK_OPT_THREADS, t = threadIdx.x;
    const int b = blockIdx.x * T + t;
    const int dd = d * d;
    double* A = sm_blk + t;                 // A[e * T]
    double* x = sm_blk + (size_t)dd * T + t; // x[i * T]
    double val = 0.0;
    if (b < nb) {
        const float* hm = Hm + (idx_m ? idx_m[b] : (long long)b) * dd;
        const float* jj = J + (idx_j ? idx_j[b] : (long long)b) * dd;
        for (int pass = 0; pass < (DOPT ? 2 : 1); ++pass) {
            for (int e = 0; e < dd; ++e) {
                const float eye = (e / d == e % d) ? lam : 0.f;          // lam * I: zero off the diagonal
                float m;
                if (!DOPT) m = __fadd_rn(__fadd_rn(hm[e], jj[e]), eye);   // Hm + J + lam I
                else {
                    m = __fadd_rn(hm[e], eye);                            // Hm + lam I
                    if (pass == 1) m = __fadd_rn(m, jj[e]);               // (Hm + lam I) + J
                }
                A[e * T] = (double)m;
            }
            // LU, partial pivoting, in place (unit lower factor below the diagonal)
            double logabs = 0.0;
            bool singular = false;
            int piv[11];
            for (int k = 0; k < d; ++k) {
                int p = k; double best = fabs(A[(k * d + k) * T]);
                for (int i = k + 1; i < d; ++i) { const double a = fabs(A[(i * d + k) * T]); if (a > best) { best = a; p = i; } }
                piv[k] = p;
                if (p != k) for (int c = 0; c < d; ++c) { const double s = A[(k * d + c) * T]; A[(k * d + c) * T] = A[(p * d + c) * T]; A[(p * d + c) * T] = s; }
                const double akk = A[(k * d + k) * T];
                if (akk == 0.0) { singular = true; continue; }
                logabs += log(fabs(akk));
                const double inv = 1.0 / akk;
                for (int i = k + 1; i < d; ++i) {
                    const double l = A[(i * d + k) * T] * inv;
                    A[(i * d + k) * T] = l;
                    for (int c = k + 1; c < d; ++c) A[(i * d + c) * T] -= l * A[(k * d + c) * T];
                }
            }
            if (DOPT) {
                if (singular) logabs = -INFINITY;                         // slogdet of a singular matrix
                val += (pass == 1) ? logabs : -logabs;
            } else {
                if (singular) { val = NAN; break; }                       // torch.linalg.inv raises; flagged as NaN here
                double tr = 0.0;
                for (int j = 0; j < d; ++j) {                             // column j of the inverse, its j-th entry
                    for (int i = 0; i < d; ++i) x[i * T] = (i == j) ? 1.0 : 0.0;
                    for (int k = 0; k < d; ++k) if (piv[k] != k) { const double s = x[k * T]; x[k * T] = x[piv[k] * T]; x[piv[k] * T] = s; }
                    for (int i = 1; i < d; ++i) { double s = x[i * T]; for (int c = 0; c < i; ++c) s -= A[(i * d + c) * T] * x[c * T]; x[i * T] = s; }
                    for (int i = d - 1; i >= j; --i) {                     // rows below j are not needed
                        double s = x[i * T];
                        for (int c = i + 1; c < d; ++c) s -= A[(i * d + c) * T] * x[c * T];
                        x[i * T] = s / A[(i * d + i) * T];
                    }
                    tr += x[j * T];
                }
                val = -tr;
            }
        }
        if (per_block) per_block[b] = (float)val;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) val += __shfl_down_sync(0xffffffffu, val, o);
    if (t == 0) partial[blockIdx.x] = val;
}

}  // namespace fb
