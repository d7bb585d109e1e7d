// fb_common.cuh -- shared device math for the FisherRF candidate-view scoring path (sm_100a).
//
// Everything here is plain FP32 SIMT arithmetic: the path has no dense contraction,
// so tensor cores are deliberately unused.  The per-Gaussian
// projection follows the arithmetic ORDER of the reference so that integer
// outputs (radii, tile rectangles, keys) come out identical:
//   reference preprocess      DGR/cuda_rasterizer/forward.cu:155-256
//   reference helpers         DGR/cuda_rasterizer/auxiliary.h:41-97,139-164
//   GLM mat3*mat3 order       DGR/third_party/glm/glm/detail/type_mat3x3.inl:514-545
//
// Rounding.  The reference is compiled with nvcc's default -fmad=true, so which multiplies are
// fused into FMAs is part of its observable behaviour (it moves radii / tile rectangles / depth
// key bits by an ulp).  The fused sequence nvcc 12.9 -O3 emits for the reference source was read
// from its SASS (ptxas fuses further than the PTX shows); it is written out here with explicit __fmul_rn/__fmaf_rn/__fadd_rn so that
// it cannot drift with compiler heuristics.  Three-term dot products are
//     t = a1*b1;  t = fma(a0,b0,t);  t = fma(a2,b2,t)        (dot3)
// and a trailing "+ c" is a separate rounded add.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "../../include/fisher_b200.h"

#define FB_TILE_PIX (FB_TILE_X * FB_TILE_Y)

namespace fb {

typedef unsigned long long u64;

// Column-major 3x3 with the reference's (GLM) element naming: c[col][row].
struct M3 {
    float c[3][3];
};

__device__ __forceinline__ float dot3(float a0, float b0, float a1, float b1, float a2, float b2) {
    float t = __fmul_rn(a1, b1);
    t = __fmaf_rn(a0, b0, t);
    return __fmaf_rn(a2, b2, t);
}

// (A*B)[col][row] = A[0][row]*B[col][0] + A[1][row]*B[col][1] + A[2][row]*B[col][2], fused as dot3
__device__ __forceinline__ M3 m3_mul(const M3& A, const M3& B) {
    M3 R;
#pragma unroll
    for (int col = 0; col < 3; ++col) {
#pragma unroll
        for (int row = 0; row < 3; ++row)
            R.c[col][row] = dot3(A.c[0][row], B.c[col][0], A.c[1][row], B.c[col][1], A.c[2][row], B.c[col][2]);
    }
    return R;
}
__device__ __forceinline__ M3 m3_T(const M3& A) {
    M3 R;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) R.c[i][j] = A.c[j][i];
    return R;
}

// Camera constants resident in constant-like kernel parameters (one per launch).
struct Cam {
    int   W, H;
    int   tiles_x, tiles_y;
    float tanfovx, tanfovy;
    float focal_x, focal_y;
    float scale_modifier;
    int   sh_degree, sh_coeffs, prefiltered;
    float bg[3];
    float campos[3];
    float view[16];
    float proj[16];
    float cull_k;       // bound on ||W J||^2 z^2 (see maybe_on_screen), 0 = no early cull
    int   view_ident;   // the view matrix is exactly the identity
};

__host__ inline Cam make_cam(const fb_camera* c) {
    Cam k;
    k.W = c->image_width; k.H = c->image_height;
    k.tiles_x = (k.W + FB_TILE_X - 1) / FB_TILE_X;
    k.tiles_y = (k.H + FB_TILE_Y - 1) / FB_TILE_Y;
    k.tanfovx = c->tanfovx; k.tanfovy = c->tanfovy;
    // rasterizer_impl.cu:222-223
    k.focal_y = k.H / (2.0f * c->tanfovy);
    k.focal_x = k.W / (2.0f * c->tanfovx);
    k.scale_modifier = c->scale_modifier;
    k.sh_degree = c->sh_degree; k.sh_coeffs = c->sh_coeffs; k.prefiltered = c->prefiltered;
    for (int i = 0; i < 3; ++i) { k.bg[i] = c->bg[i]; k.campos[i] = c->campos[i]; }
    for (int i = 0; i < 16; ++i) { k.view[i] = c->viewmatrix[i]; k.proj[i] = c->projmatrix[i]; }
    // ||W||_2^2 of the view's 3x3 block: 1 for a rigid transform (checked), its Frobenius norm otherwise
    double wtw[3][3], fro = 0.0; bool rigid = true;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0.0;
            for (int r = 0; r < 3; ++r) s += (double)k.view[4 * i + r] * (double)k.view[4 * j + r];
            wtw[i][j] = s;
            if (fabs(s - (i == j ? 1.0 : 0.0)) > 1e-5) rigid = false;
            if (i == j) fro += s;
        }
    (void)wtw;
    const double lx = 1.3 * k.tanfovx, ly = 1.3 * k.tanfovy;
    const double jb = (double)k.focal_x * k.focal_x * (1.0 + lx * lx) + (double)k.focal_y * k.focal_y * (1.0 + ly * ly);
    const double ck = (rigid ? 1.0 : fro) * jb * 1.02;
    k.cull_k = (ck == ck && ck < 1e30) ? (float)ck : 0.f;
    k.view_ident = 1;
    for (int i = 0; i < 16; ++i) if (k.view[i] != ((i % 5 == 0) ? 1.0f : 0.0f)) k.view_ident = 0;
    return k;
}

// auxiliary.h:58-77 (matrices are consumed column-major: m[0],m[4],m[8],m[12] is row 0)
__device__ __forceinline__ float3 xform4x3(const float3& p, const float* m) {
    float3 r;
    r.x = __fadd_rn(m[12], dot3(m[0], p.x, m[4], p.y, m[8], p.z));
    r.y = __fadd_rn(m[13], dot3(m[1], p.x, m[5], p.y, m[9], p.z));
    r.z = __fadd_rn(m[14], dot3(m[2], p.x, m[6], p.y, m[10], p.z));
    return r;
}
__device__ __forceinline__ float4 xform4x4(const float3& p, const float* m) {
    float4 r;
    r.x = __fadd_rn(m[12], dot3(m[0], p.x, m[4], p.y, m[8], p.z));
    r.y = __fadd_rn(m[13], dot3(m[1], p.x, m[5], p.y, m[9], p.z));
    r.z = __fadd_rn(m[14], dot3(m[2], p.x, m[6], p.y, m[10], p.z));
    r.w = __fadd_rn(m[15], dot3(m[3], p.x, m[7], p.y, m[11], p.z));
    return r;
}

// auxiliary.h:41-44 -- the literals are double, so ((v + 1.0) * S - 1.0) * 0.5 is evaluated in
// FP64 and rounded to FP32 once.
__device__ __forceinline__ float ndc_to_pix_exact(float v, int S) {
    double d = __dmul_rn(__fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0), 0.5);   // one fma.rn.f64, as compiled
    return (float)d;
}

// auxiliary.h:46-56: float division, (int) truncation, clamp to [0, grid].
__device__ __forceinline__ void tile_rect(float px, float py, int max_radius, int tiles_x, int tiles_y,
                                          int& x0, int& y0, int& x1, int& y1) {
    x0 = min(tiles_x, max(0, (int)((px - max_radius) / FB_TILE_X)));
    y0 = min(tiles_y, max(0, (int)((py - max_radius) / FB_TILE_Y)));
    x1 = min(tiles_x, max(0, (int)((px + max_radius + FB_TILE_X - 1) / FB_TILE_X)));
    y1 = min(tiles_y, max(0, (int)((py + max_radius + FB_TILE_Y - 1) / FB_TILE_Y)));
}

// forward.cu:118-152.  Quaternion (w,x,y,z) is used UN-normalised.
__device__ __forceinline__ void rot_scale_M(const float3& scale, float mod, const float4& q, M3& R, float3& s, M3& M) {
    float r = q.x, x = q.y, y = q.z, z = q.w;
    // GLM constructor fills columns: (a,b,c) is column 0.
    // as compiled (SASS): of each +/- pair of products one is rounded and shared, the other is fused:
    //   yy+zz: mul,mul,add;  xx+zz, xx+yy: fma(x,x,zz), fma(x,x,yy);
    //   xy-/+rz: fma(x,y,-/+rz);  yz-/+rx: fma(y,z,-/+rx);  xz+ry: fma(r,y,xz);  xz-ry: fma(-r,y,xz)
    const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
    const float rz = __fmul_rn(r, z), xz = __fmul_rn(x, z), rx = __fmul_rn(r, x);
    R.c[0][0] = __fsub_rn(1.f, 2.f * __fadd_rn(yy, zz)); R.c[0][1] = 2.f * __fmaf_rn(x, y, -rz); R.c[0][2] = 2.f * __fmaf_rn(r, y, xz);
    R.c[1][0] = 2.f * __fmaf_rn(x, y, rz); R.c[1][1] = __fsub_rn(1.f, 2.f * __fmaf_rn(x, x, zz)); R.c[1][2] = 2.f * __fmaf_rn(y, z, -rx);
    R.c[2][0] = 2.f * __fmaf_rn(-r, y, xz); R.c[2][1] = 2.f * __fmaf_rn(y, z, rx); R.c[2][2] = __fsub_rn(1.f, 2.f * __fmaf_rn(x, x, yy));
    s.x = __fmul_rn(mod, scale.x); s.y = __fmul_rn(mod, scale.y); s.z = __fmul_rn(mod, scale.z);
    M3 S;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) S.c[i][j] = 0.f;
    S.c[0][0] = s.x; S.c[1][1] = s.y; S.c[2][2] = s.z;
    M = m3_mul(S, R);
}
__device__ __forceinline__ void cov3d_from_scale_rot(const float3& scale, float mod, const float4& q, float* cov6) {
    M3 R, M; float3 s;
    rot_scale_M(scale, mod, q, R, s, M);
    M3 Sigma = m3_mul(m3_T(M), M);
    cov6[0] = Sigma.c[0][0]; cov6[1] = Sigma.c[0][1]; cov6[2] = Sigma.c[0][2];
    cov6[3] = Sigma.c[1][1]; cov6[4] = Sigma.c[1][2]; cov6[5] = Sigma.c[2][2];
}

// Shared by forward (forward.cu:74-113) and backward (backward.cu:276-330): the clamped
// camera-space point t, T = W*J and the 2D covariance (a,b,c) including the +0.3 low-pass.
struct Cov2D {
    float3 t;       // clamped
    float  txtz, tytz;
    M3     T;
    float  a, b, c;
};
// FAST: approximate divisions (div.approx, 2 ulp).  Only for values that never decide an integer: the scoring path's
// coefficient phase (its conic and polynomial feed the pair sweep alone, held to rtol 1e-3 on the scores).
template <bool FAST> __device__ __forceinline__ float fdiv(float a, float b) { return FAST ? __fdividef(a, b) : a / b; }
// IDENT: the view matrix is exactly the identity (the scoring path: the caller's camera is set up with w2c = I and the
// candidate pose is applied to the means, gaussian.py:1523-1527), so t = mean and T = W J = J; the products with 0 and 1
// are written out.  Only used together with FAST (values that decide no integer).
template <bool FAST = false, bool IDENT = false>
__device__ __forceinline__ Cov2D cov2d(const float3& mean, const Cam& cam, const float* cov6) {
    Cov2D o;
    float3 t = IDENT ? mean : xform4x3(mean, cam.view);
    const float limx = 1.3f * cam.tanfovx;
    const float limy = 1.3f * cam.tanfovy;
    o.txtz = fdiv<FAST>(t.x, t.z);
    o.tytz = fdiv<FAST>(t.y, t.z);
    t.x = __fmul_rn(fminf(limx, fmaxf(-limx, o.txtz)), t.z);
    t.y = __fmul_rn(fminf(limy, fmaxf(-limy, o.tytz)), t.z);
    o.t = t;
    M3 J;
    const float tzz = __fmul_rn(t.z, t.z);
    J.c[0][0] = fdiv<FAST>(cam.focal_x, t.z); J.c[0][1] = 0.f; J.c[0][2] = fdiv<FAST>(-__fmul_rn(cam.focal_x, t.x), tzz);
    J.c[1][0] = 0.f; J.c[1][1] = fdiv<FAST>(cam.focal_y, t.z); J.c[1][2] = fdiv<FAST>(-__fmul_rn(cam.focal_y, t.y), tzz);
    J.c[2][0] = 0.f; J.c[2][1] = 0.f; J.c[2][2] = 0.f;
    if (IDENT) {
        o.T = J;
        const float j00 = J.c[0][0], j02 = J.c[0][2], j11 = J.c[1][1], j12 = J.c[1][2];
        const float u0 = j00 * cov6[0] + j02 * cov6[2], u2 = j00 * cov6[2] + j02 * cov6[5];       // (J row 0) V, x and z parts
        const float u1 = j00 * cov6[1] + j02 * cov6[4];
        const float w1 = j11 * cov6[3] + j12 * cov6[4], w2 = j11 * cov6[4] + j12 * cov6[5];
        o.a = (u0 * j00 + u2 * j02) + 0.3f;
        o.b = u1 * j11 + u2 * j12;
        o.c = (w1 * j11 + w2 * j12) + 0.3f;
        return o;
    }
    M3 Wm;
    Wm.c[0][0] = cam.view[0]; Wm.c[0][1] = cam.view[4]; Wm.c[0][2] = cam.view[8];
    Wm.c[1][0] = cam.view[1]; Wm.c[1][1] = cam.view[5]; Wm.c[1][2] = cam.view[9];
    Wm.c[2][0] = cam.view[2]; Wm.c[2][1] = cam.view[6]; Wm.c[2][2] = cam.view[10];
    o.T = m3_mul(Wm, J);
    M3 V;
    V.c[0][0] = cov6[0]; V.c[0][1] = cov6[1]; V.c[0][2] = cov6[2];
    V.c[1][0] = cov6[1]; V.c[1][1] = cov6[3]; V.c[1][2] = cov6[4];
    V.c[2][0] = cov6[2]; V.c[2][1] = cov6[4]; V.c[2][2] = cov6[5];
    M3 cov = m3_mul(m3_mul(m3_T(o.T), m3_T(V)), o.T);
    o.a = __fadd_rn(cov.c[0][0], 0.3f);
    o.b = cov.c[0][1];
    o.c = __fadd_rn(cov.c[1][1], 0.3f);
    return o;
}

// Result of projecting one Gaussian (stage 1).
struct Splat {
    float px, py;     // pixel-space centre
    float depth;      // p_view.z
    float ca, cb, cc; // conic (inverse 2D covariance): x, y, z of the reference's float3
    int   radius;
    int   x0, y0, x1, y1; // tile rectangle [x0,x1) x [y0,y1)
};

// Returns 0 if dropped; 1 if visible.  `culled_near` reports the near-plane cull for the
// prefiltered trap (auxiliary.h:154-160).
__device__ __forceinline__ int project_splat(const float3& p_orig, const float* cov6, const Cam& cam,
                                             Splat& s, int& culled_near) {
    culled_near = 0;
    float3 p_view = xform4x3(p_orig, cam.view);
    if (p_view.z <= 0.001f) { culled_near = 1; return 0; }
    float4 p_hom = xform4x4(p_orig, cam.proj);
    float p_w = 1.0f / __fadd_rn(p_hom.w, 0.0000001f);
    float projx = __fmul_rn(p_hom.x, p_w), projy = __fmul_rn(p_hom.y, p_w);
    Cov2D c2 = cov2d(p_orig, cam, cov6);
    // as compiled (SASS): det = fma(a, c, -(b*b)); mid*mid - det = fma(mid, mid, -det)
    float det = __fmaf_rn(c2.a, c2.c, -__fmul_rn(c2.b, c2.b));
    if (det == 0.0f) return 0;
    float det_inv = 1.f / det;
    s.ca = __fmul_rn(c2.c, det_inv); s.cb = __fmul_rn(-c2.b, det_inv); s.cc = __fmul_rn(c2.a, det_inv);
    float mid = __fmul_rn(0.5f, __fadd_rn(c2.a, c2.c));
    float root = sqrtf(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
    float lambda1 = __fadd_rn(mid, root);
    float lambda2 = __fsub_rn(mid, root);
    float my_radius = ceilf(__fmul_rn(3.f, sqrtf(fmaxf(lambda1, lambda2))));
    s.px = ndc_to_pix_exact(projx, cam.W);
    s.py = ndc_to_pix_exact(projy, cam.H);
    s.radius = (int)my_radius;
    tile_rect(s.px, s.py, s.radius, cam.tiles_x, cam.tiles_y, s.x0, s.y0, s.x1, s.y1);
    if ((s.x1 - s.x0) * (s.y1 - s.y0) == 0) return 0;
    s.depth = p_view.z;
    return 1;
}

// Conservative early test of stage 1: false only if project_splat would certainly return 0 (behind the near plane, or
// the 3-sigma tile rectangle certainly empty), so skipping the projection for such a Gaussian changes nothing.
// lambda_max(cov2D) <= ||W J||_F^2 tr(Sigma) + 0.3 with ||J||_F^2 <= (fx^2 (1 + lx^2) + fy^2 (1 + ly^2)) / z^2 (the
// Jacobian's t.x/t.z, t.y/t.z are clamped to lx, ly = 1.3 tan(fov/2), forward.cu:82-87); the reference's radius is
// ceil(3 sqrt(lambda)) with lambda <= lambda_max + sqrt(0.1) (its max(0.1, .) under the root, forward.cu:229-232).
// The rectangle is empty iff px + radius < 1 or (px - radius) / 16 >= tiles_x (auxiliary.h:46-56), same in y.
__device__ __forceinline__ bool maybe_on_screen(const float3& p_orig, float tr_sigma, const Cam& cam) {
    const float3 p_view = xform4x3(p_orig, cam.view);
    if (p_view.z <= 0.001f) return false;                                   // the reference's own near cull
    if (cam.cull_k == 0.f) return true;
    const float4 p_hom = xform4x4(p_orig, cam.proj);
    const float p_w = 1.0f / __fadd_rn(p_hom.w, 0.0000001f);
    const float px = ndc_to_pix_exact(__fmul_rn(p_hom.x, p_w), cam.W), py = ndc_to_pix_exact(__fmul_rn(p_hom.y, p_w), cam.H);
    const float lam = cam.cull_k * tr_sigma / (p_view.z * p_view.z) + 0.65f;
    const float rb = 3.0f * sqrtf(lam) * 1.001f + 1.5f;
    const bool off = (px + rb < 0.f) || (px - rb > (float)(cam.tiles_x * FB_TILE_X) + 1.f) ||
                     (py + rb < 0.f) || (py - rb > (float)(cam.tiles_y * FB_TILE_Y) + 1.f);
    return !off;                                                            // NaN anywhere: not culled
}

// Exponent of a (pixel, Gaussian) pair, forward.cu:341 / backward.cu:968:
//   -0.5f*(A*dx*dx + C*dy*dy) - B*dx*dy   as compiled (SASS):
//   q = fma(dx, dx*A, dy*(dy*C));  power = fma(q, -0.5f, -(dy*(dx*B)))
__device__ __forceinline__ float pair_power(float A, float B, float C, float dx, float dy) {
    const float q = __fmaf_rn(dx, __fmul_rn(dx, A), __fmul_rn(dy, __fmul_rn(dy, C)));
    return __fmaf_rn(q, -0.5f, -__fmul_rn(dy, __fmul_rn(dx, B)));
}

}  // namespace fb
