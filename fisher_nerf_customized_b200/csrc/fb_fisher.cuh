// fb_fisher.cuh -- stage 4: backward pass fused with FisherRF's diagonal-Hessian accumulation.
//
// Replaces renderCUDAFused<3> and its relocated per-pair chain
//   DGR/cuda_rasterizer/backward.cu:850-1140 (pixel loop), :276-408 (cov2D chain),
//   :412-475 (cov3D -> scale/rotation), :532-583 (projection chain).
//
// The reference runs the whole chain rule once per (pixel, Gaussian) pair and issues 25 float
// atomics per pair after raising every partial to `grad_power` (:1093-1137).  Here the pair loop
// only produces what actually varies per pair.  With d = mean2D - pixel, G = exp(power),
// s = G * opacity * dL/dalpha, every partial of the pair is  s * (per-Gaussian constant) *
// (a monomial of d):
//     dL/dconic   = -1/2 s (dx^2, dx dy, dy^2)                         (:1035-1037)
//     dL/dmean2D  = -1/2 s (W (A dx + B dy), H (C dy + B dx))          (:1032-1033)
//     dL/dopacity = s / opacity                                        (:1038)
// and everything downstream (cov2D -> cov3D -> scales/rotations, means3D) is a LINEAR map with
// per-Gaussian coefficients applied to those five numbers.  Therefore
//     power = 1:  sum_pairs (a.u)   = a . (sum_pairs u)          -> 6 accumulators (+3 colour)
//     power = 2:  sum_pairs (a.u)^2 = a^T (sum_pairs u u^T) a    -> 13 monomial moments
//                                                                   sum s^2 dx^i dy^j, i+j in {0,2,3,4}
// The moments are reduced across the 256 pixels of the tile with a halving warp butterfly
// (16 shuffles for 16 values) BEFORE any atomic, and the chain rule runs once per
// (Gaussian, tile) instance in an epilogue instead of once per pair.
#pragma once
#include "fb_layout.cuh"
#include "fb_render.cuh"

namespace fb {

#define FB_NACC 16

// ---- per-Gaussian context for the chain rule --------------------------------------------------
struct GCtx {
    Cov2D c2;
    float denom, d2i;
    float xmul, ymul;
    float V[6];
    float3 m;       // mean as given to the rasteriser
    float m_w, mul1, mul2;
};

template <bool FAST = false, bool IDENT = false>
__device__ __forceinline__ GCtx make_ctx(const float3& mean, const float* cov6, const Cam& cam) {
    GCtx g;
    g.c2 = cov2d<FAST, IDENT>(mean, cam, cov6);
    const float limx = 1.3f * cam.tanfovx, limy = 1.3f * cam.tanfovy;
    g.xmul = (g.c2.txtz < -limx || g.c2.txtz > limx) ? 0.f : 1.f;
    g.ymul = (g.c2.tytz < -limy || g.c2.tytz > limy) ? 0.f : 1.f;
    const float a = g.c2.a, b = g.c2.b, c = g.c2.c;
    g.denom = a * c - b * b;
    g.d2i = fdiv<FAST>(1.0f, (g.denom * g.denom) + 0.0000001f);
#pragma unroll
    for (int i = 0; i < 6; ++i) g.V[i] = cov6[i];
    g.m = mean;
    float4 mh = xform4x4(mean, cam.proj);
    g.m_w = fdiv<FAST>(1.0f, mh.w + 0.0000001f);
    g.mul1 = (cam.proj[0] * mean.x + cam.proj[4] * mean.y + cam.proj[8] * mean.z + cam.proj[12]) * g.m_w * g.m_w;
    g.mul2 = (cam.proj[1] * mean.x + cam.proj[5] * mean.y + cam.proj[9] * mean.z + cam.proj[13]) * g.m_w * g.m_w;
    return g;
}

struct ChainOut {
    float cov3D[6];
    float mean[3];
};

// Linear chain: (dL/dconic.{x,y,w}, dL/dmean2D.{x,y}) -> dL/dcov3D[6], dL/dmeans3D[3].
// backward.cu:333-407 and :566-574.
__device__ __forceinline__ ChainOut chain_geom(const GCtx& g, const Cam& cam, float gx, float gy, float gw,
                                               float hx, float hy) {
    ChainOut o;
    const float a = g.c2.a, b = g.c2.b, c = g.c2.c, denom = g.denom, d2i = g.d2i;
    const M3& T = g.c2.T;
    float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
    if (d2i != 0.f) {
        dL_da = d2i * (-c * c * gx + 2 * b * c * gy + (denom - a * c) * gw);
        dL_dc = d2i * (-a * a * gw + 2 * a * b * gy + (denom - a * c) * gx);
        dL_db = d2i * 2 * (b * c * gx - (denom + 2 * b * b) * gy + a * b * gw);
        o.cov3D[0] = (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
        o.cov3D[3] = (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
        o.cov3D[5] = (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
        o.cov3D[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][1] * dL_dc;
        o.cov3D[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][2] * dL_dc;
        o.cov3D[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db + 2 * T.c[1][1] * T.c[1][2] * dL_dc;
    } else {
#pragma unroll
        for (int i = 0; i < 6; ++i) o.cov3D[i] = 0.f;
    }
    // Vrk (symmetric): V00 V01 V02 / V01 V11 V12 / V02 V12 V22  = V[0] V[1] V[2] / . V[3] V[4] / . . V[5]
    const float V00 = g.V[0], V01 = g.V[1], V02 = g.V[2], V11 = g.V[3], V12 = g.V[4], V22 = g.V[5];
    const float r0x = T.c[0][0] * V00 + T.c[0][1] * V01 + T.c[0][2] * V02;
    const float r0y = T.c[0][0] * V01 + T.c[0][1] * V11 + T.c[0][2] * V12;
    const float r0z = T.c[0][0] * V02 + T.c[0][1] * V12 + T.c[0][2] * V22;
    const float r1x = T.c[1][0] * V00 + T.c[1][1] * V01 + T.c[1][2] * V02;
    const float r1y = T.c[1][0] * V01 + T.c[1][1] * V11 + T.c[1][2] * V12;
    const float r1z = T.c[1][0] * V02 + T.c[1][1] * V12 + T.c[1][2] * V22;
    const float dT00 = 2 * r0x * dL_da + r1x * dL_db;
    const float dT01 = 2 * r0y * dL_da + r1y * dL_db;
    const float dT02 = 2 * r0z * dL_da + r1z * dL_db;
    const float dT10 = 2 * r1x * dL_dc + r0x * dL_db;
    const float dT11 = 2 * r1y * dL_dc + r0y * dL_db;
    const float dT12 = 2 * r1z * dL_dc + r0z * dL_db;
    // W (GLM naming, W[col][row]) built from the view matrix as in cov2d().
    const float W00 = cam.view[0], W01 = cam.view[4], W02 = cam.view[8];
    const float W10 = cam.view[1], W11 = cam.view[5], W12 = cam.view[9];
    const float W20 = cam.view[2], W21 = cam.view[6], W22 = cam.view[10];
    const float dJ00 = W00 * dT00 + W01 * dT01 + W02 * dT02;
    const float dJ02 = W20 * dT00 + W21 * dT01 + W22 * dT02;
    const float dJ11 = W10 * dT10 + W11 * dT11 + W12 * dT12;
    const float dJ12 = W20 * dT10 + W21 * dT11 + W22 * dT12;
    const float3 t = g.c2.t;
    const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
    const float hxf = cam.focal_x, hyf = cam.focal_y;
    const float dtx = g.xmul * -hxf * tz2 * dJ02;
    const float dty = g.ymul * -hyf * tz2 * dJ12;
    const float dtz = -hxf * tz2 * dJ00 - hyf * tz2 * dJ11 + (2 * hxf * t.x) * tz3 * dJ02 + (2 * hyf * t.y) * tz3 * dJ12;
    // transformVec4x3Transpose (auxiliary.h:89-97)
    float mx = cam.view[0] * dtx + cam.view[1] * dty + cam.view[2] * dtz;
    float my = cam.view[4] * dtx + cam.view[5] * dty + cam.view[6] * dtz;
    float mz = cam.view[8] * dtx + cam.view[9] * dty + cam.view[10] * dtz;
    // projection path (backward.cu:566-574); the two paths are summed BEFORE the power is applied.
    const float* pr = cam.proj;
    const float m_w = g.m_w, mul1 = g.mul1, mul2 = g.mul2;
    mx += (pr[0] * m_w - pr[3] * mul1) * hx + (pr[1] * m_w - pr[3] * mul2) * hy;
    my += (pr[4] * m_w - pr[7] * mul1) * hx + (pr[5] * m_w - pr[7] * mul2) * hy;
    mz += (pr[8] * m_w - pr[11] * mul1) * hx + (pr[9] * m_w - pr[11] * mul2) * hy;
    o.mean[0] = mx; o.mean[1] = my; o.mean[2] = mz;
    return o;
}

// dL/dcov3D[6] -> dL/dscale[3], dL/drot[4]   (backward.cu:412-475; no quaternion-normalisation
// Jacobian).
struct SRCtx { M3 R, M; float3 s; float4 q; };
__device__ __forceinline__ SRCtx make_sr(const float3& scale, float mod, const float4& q) {
    SRCtx c; c.q = q;
    rot_scale_M(scale, mod, q, c.R, c.s, c.M);
    return c;
}
__device__ __forceinline__ void chain_scale_rot(const SRCtx& c, const float* d6, float* dscale3, float* drot4) {
    M3 dS;
    dS.c[0][0] = d6[0]; dS.c[0][1] = 0.5f * d6[1]; dS.c[0][2] = 0.5f * d6[2];
    dS.c[1][0] = 0.5f * d6[1]; dS.c[1][1] = d6[3]; dS.c[1][2] = 0.5f * d6[4];
    dS.c[2][0] = 0.5f * d6[2]; dS.c[2][1] = 0.5f * d6[4]; dS.c[2][2] = d6[5];
    M3 M2;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) M2.c[i][j] = 2.0f * c.M.c[i][j];
    M3 dM = m3_mul(M2, dS);
    M3 Rt = m3_T(c.R);
    M3 dMt = m3_T(dM);
    dscale3[0] = Rt.c[0][0] * dMt.c[0][0] + Rt.c[0][1] * dMt.c[0][1] + Rt.c[0][2] * dMt.c[0][2];
    dscale3[1] = Rt.c[1][0] * dMt.c[1][0] + Rt.c[1][1] * dMt.c[1][1] + Rt.c[1][2] * dMt.c[1][2];
    dscale3[2] = Rt.c[2][0] * dMt.c[2][0] + Rt.c[2][1] * dMt.c[2][1] + Rt.c[2][2] * dMt.c[2][2];
#pragma unroll
    for (int j = 0; j < 3; ++j) { dMt.c[0][j] *= c.s.x; dMt.c[1][j] *= c.s.y; dMt.c[2][j] *= c.s.z; }
    const float r = c.q.x, x = c.q.y, y = c.q.z, z = c.q.w;
    drot4[0] = 2 * z * (dMt.c[0][1] - dMt.c[1][0]) + 2 * y * (dMt.c[2][0] - dMt.c[0][2]) + 2 * x * (dMt.c[1][2] - dMt.c[2][1]);
    drot4[1] = 2 * y * (dMt.c[1][0] + dMt.c[0][1]) + 2 * z * (dMt.c[2][0] + dMt.c[0][2]) + 2 * r * (dMt.c[1][2] - dMt.c[2][1]) - 4 * x * (dMt.c[2][2] + dMt.c[1][1]);
    drot4[2] = 2 * x * (dMt.c[1][0] + dMt.c[0][1]) + 2 * r * (dMt.c[2][0] - dMt.c[0][2]) + 2 * z * (dMt.c[1][2] + dMt.c[2][1]) - 4 * y * (dMt.c[2][2] + dMt.c[0][0]);
    drot4[3] = 2 * r * (dMt.c[0][1] - dMt.c[1][0]) + 2 * x * (dMt.c[2][0] + dMt.c[0][2]) + 2 * y * (dMt.c[1][2] + dMt.c[2][1]) - 4 * z * (dMt.c[1][1] + dMt.c[0][0]);
}

// ---- moment bookkeeping -----------------------------------------------------------------------
// acc[] layout.  POWER 2: 0:q 1:q dx2 2:q dxdy 3:q dy2 4:q dx3 5:q dx2dy 6:q dxdy2 7:q dy3
//                         8:q dx4 9:q dx3dy 10:q dx2dy2 11:q dxdy3 12:q dy4   (q = s*s)
//                POWER 1: 0:s 1:s dx2 2:s dxdy 3:s dy2 4:s dx 5:s dy
//                both:    13..15: colour channels, sum (alpha*T*dL/dpix_ch)^power
// phi basis of the quadratic form: (dx2, dxdy, dy2, dx, dy).
__device__ __forceinline__ float moment_phi(const float* acc, int i, int j) {
    // index of sum q*phi_i*phi_j in acc[]
    const int tab[5][5] = {{8, 9, 10, 4, 5}, {9, 10, 11, 5, 6}, {10, 11, 12, 6, 7}, {4, 5, 6, 1, 2}, {5, 6, 7, 2, 3}};
    return acc[tab[i][j]];
}

// u = L * (s*phi):  conic.x=-1/2 w0, conic.y=-1/2 w1, conic.w=-1/2 w2,
//                   mean2D.x=-1/2 W (A w3 + B w4), mean2D.y=-1/2 H (C w4 + B w3)
__device__ __forceinline__ void phi_to_u(const float* w, float A, float B, float C, float halfW, float halfH, float* u) {
    u[0] = -0.5f * w[0];
    u[1] = -0.5f * w[1];
    u[2] = -0.5f * w[2];
    u[3] = -(A * w[3] + B * w[4]) * halfW;
    u[4] = -(C * w[4] + B * w[3]) * halfH;
}

// Halving butterfly: 16 values per lane in, afterwards lane L (L even) holds the warp total of
// value index (L>>1) in v[0].  16 SHFL + 16 FADD instead of 80 + 80.
__device__ __forceinline__ float warp_reduce16(float* v, int lane) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const bool up = lane & 16;
        float send = up ? v[i] : v[i + 8];
        float keep = up ? v[i + 8] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const bool up = lane & 8;
        float send = up ? v[i] : v[i + 4];
        float keep = up ? v[i + 4] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const bool up = lane & 4;
        float send = up ? v[i] : v[i + 2];
        float keep = up ? v[i + 2] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    {
        const bool up = lane & 2;
        float send = up ? v[0] : v[1];
        float keep = up ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
    return v[0];
}
// value index held by `lane` after warp_reduce16: bit4 -> +8, bit3 -> +4, bit2 -> +2, bit1 -> +1
__device__ __forceinline__ int reduce16_slot(int lane) { return ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1); }

// ---- scoring polynomial -----------------------------------------------------------------------
// For candidate scoring the Hessian diagonal is only ever consumed through
//     score = sum_g sum_c H[g,c] * weight[g,c]            (models/SLAM/gaussian.py:1367)
// and H[g,c] = sum_pairs (a_c . (s*phi))^2, so
//     score = sum_pairs s^2 * poly_g(dx,dy),   poly_g = sum_c weight[g,c] (a_c . phi)^2 + weight_op / o^2
// a bivariate quartic with 13 coefficients per (view, Gaussian), in the monomial order of the
// moment row (1, dx2, dxdy, dy2, dx3, dx2dy, dxdy2, dy3, dx4, dx3dy, dx2dy2, dxdy3, dy4).
// H column order: means3D(3), opacity(1) [C_H=4, gaussian.py:1554-1560] + scales(3), rotations(4)
// [C_H=11, gaussian_object.py:2024-2032].
// C_H = 4 only needs dL/dmeans3D = M (s*phi); M (3x5) written out as a product of the chain's small matrices
// instead of five passes through chain_geom with unit vectors (the compiler may not fold x*0, so each pass costs the
// whole chain): K = d(cov2D a,b,c)/d(conic) with the -1/2 of phi_to_u folded in, D = d(t)/d(a,b,c) through
// T = W J (backward.cu:333-407), the view rotation, and the projection path for the two mean2D columns (:566-574).
template <bool FAST = false, bool IDENT = false>
__device__ __forceinline__ void mean_chain_matrix(const GCtx& g, const Cam& cam, float A, float B, float C,
                                                  float halfW, float halfH, float M[3][5]) {
    const float a = g.c2.a, b = g.c2.b, c = g.c2.c, denom = g.denom;
    const M3& T = g.c2.T;
    const float h = -0.5f * g.d2i, dmac = denom - a * c;
    float K[3][3];                                   // rows (da, db, dc), columns (phi0, phi1, phi2)
    K[0][0] = h * (-c * c);        K[0][1] = h * (2.f * b * c);                 K[0][2] = h * dmac;
    K[1][0] = 2.f * h * (b * c);   K[1][1] = -2.f * h * (denom + 2.f * b * b);  K[1][2] = 2.f * h * (a * b);
    K[2][0] = h * dmac;            K[2][1] = h * (2.f * a * b);                 K[2][2] = h * (-a * a);
    const float V00 = g.V[0], V01 = g.V[1], V02 = g.V[2], V11 = g.V[3], V12 = g.V[4], V22 = g.V[5];
    float r0[3], r1[3];
    if (IDENT) {            // T = J: T.c[0][1] = T.c[1][0] = 0
        r0[0] = T.c[0][0] * V00 + T.c[0][2] * V02; r0[1] = T.c[0][0] * V01 + T.c[0][2] * V12; r0[2] = T.c[0][0] * V02 + T.c[0][2] * V22;
        r1[0] = T.c[1][1] * V01 + T.c[1][2] * V02; r1[1] = T.c[1][1] * V11 + T.c[1][2] * V12; r1[2] = T.c[1][1] * V12 + T.c[1][2] * V22;
    } else {
        r0[0] = T.c[0][0] * V00 + T.c[0][1] * V01 + T.c[0][2] * V02; r0[1] = T.c[0][0] * V01 + T.c[0][1] * V11 + T.c[0][2] * V12;
        r0[2] = T.c[0][0] * V02 + T.c[0][1] * V12 + T.c[0][2] * V22;
        r1[0] = T.c[1][0] * V00 + T.c[1][1] * V01 + T.c[1][2] * V02; r1[1] = T.c[1][0] * V01 + T.c[1][1] * V11 + T.c[1][2] * V12;
        r1[2] = T.c[1][0] * V02 + T.c[1][1] * V12 + T.c[1][2] * V22;
    }
    const float* vw = cam.view;
    const float p00 = IDENT ? r0[0] : vw[0] * r0[0] + vw[4] * r0[1] + vw[8] * r0[2], p01 = IDENT ? r1[0] : vw[0] * r1[0] + vw[4] * r1[1] + vw[8] * r1[2];
    const float p10 = IDENT ? r0[1] : vw[1] * r0[0] + vw[5] * r0[1] + vw[9] * r0[2], p11 = IDENT ? r1[1] : vw[1] * r1[0] + vw[5] * r1[1] + vw[9] * r1[2];
    const float p20 = IDENT ? r0[2] : vw[2] * r0[0] + vw[6] * r0[1] + vw[10] * r0[2], p21 = IDENT ? r1[2] : vw[2] * r1[0] + vw[6] * r1[1] + vw[10] * r1[2];
    const float3 t = g.c2.t;
    const float tz = fdiv<FAST>(1.f, t.z), tz2 = tz * tz, tz3 = tz2 * tz;
    const float fx2 = -cam.focal_x * tz2, fy2 = -cam.focal_y * tz2;
    const float kx = g.xmul * fx2, ky = g.ymul * fy2;
    const float cx3 = 2.f * cam.focal_x * t.x * tz3, cy3 = 2.f * cam.focal_y * t.y * tz3;
    float D[3][3];                                   // rows (dtx, dty, dtz), columns (da, db, dc)
    D[0][0] = 2.f * kx * p20;                D[0][1] = kx * p21;                                      D[0][2] = 0.f;
    D[1][0] = 0.f;                           D[1][1] = ky * p20;                                      D[1][2] = 2.f * ky * p21;
    D[2][0] = 2.f * (fx2 * p00 + cx3 * p20); D[2][1] = fx2 * p01 + fy2 * p10 + cx3 * p21 + cy3 * p20; D[2][2] = 2.f * (fy2 * p11 + cy3 * p21);
    float E[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) E[i][j] = D[i][0] * K[0][j] + D[i][1] * K[1][j] + D[i][2] * K[2][j];
    const float* pr = cam.proj;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) M[i][j] = IDENT ? E[i][j] : vw[4 * i] * E[0][j] + vw[4 * i + 1] * E[1][j] + vw[4 * i + 2] * E[2][j];
        const float px = pr[4 * i] * g.m_w - pr[4 * i + 3] * g.mul1, py = pr[4 * i + 1] * g.m_w - pr[4 * i + 3] * g.mul2;
        M[i][3] = -(px * (A * halfW) + py * (B * halfH));
        M[i][4] = -(px * (B * halfW) + py * (C * halfH));
    }
}

// Z[i][0..6] = d(scale 3, rotation 4) / d(phi_i) for the three conic components of phi (the mean2D components do not
// reach cov3D): dL/dcov3D = N K e_i with K as in mean_chain_matrix (the -1/2 of phi_to_u folded in) and N the
// T-products of backward.cu:370-390, then chain_scale_rot.  Equal to chain_geom + chain_scale_rot on unit vectors.
__device__ __forceinline__ void scale_rot_matrix(const GCtx& g, const SRCtx& sr, float Z[3][7]) {
    const float a = g.c2.a, b = g.c2.b, c = g.c2.c, denom = g.denom;
    const float h = -0.5f * g.d2i, dmac = denom - a * c;
    float K[3][3];
    K[0][0] = h * (-c * c);        K[0][1] = h * (2.f * b * c);                 K[0][2] = h * dmac;
    K[1][0] = 2.f * h * (b * c);   K[1][1] = -2.f * h * (denom + 2.f * b * b);  K[1][2] = 2.f * h * (a * b);
    K[2][0] = h * dmac;            K[2][1] = h * (2.f * a * b);                 K[2][2] = h * (-a * a);
    const M3& T = g.c2.T;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float da = K[0][i], db = K[1][i], dc = K[2][i];
        float d6[6];
        d6[0] = T.c[0][0] * T.c[0][0] * da + T.c[0][0] * T.c[1][0] * db + T.c[1][0] * T.c[1][0] * dc;
        d6[3] = T.c[0][1] * T.c[0][1] * da + T.c[0][1] * T.c[1][1] * db + T.c[1][1] * T.c[1][1] * dc;
        d6[5] = T.c[0][2] * T.c[0][2] * da + T.c[0][2] * T.c[1][2] * db + T.c[1][2] * T.c[1][2] * dc;
        d6[1] = 2 * T.c[0][0] * T.c[0][1] * da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * db + 2 * T.c[1][0] * T.c[1][1] * dc;
        d6[2] = 2 * T.c[0][0] * T.c[0][2] * da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * db + 2 * T.c[1][0] * T.c[1][2] * dc;
        d6[4] = 2 * T.c[0][2] * T.c[0][1] * da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * db + 2 * T.c[1][1] * T.c[1][2] * dc;
        chain_scale_rot(sr, d6, &Z[i][0], &Z[i][3]);
    }
}

#ifndef FB_COEF11_DIRECT
#define FB_COEF11_DIRECT 1
#endif
template <int CH, bool FAST = false, bool IDENT = false>
__device__ __forceinline__ void score_coefs(const GCtx& ctx, const Cam& cam, const SRCtx& sr, float A, float B, float C,
                                            float op, const float* __restrict__ w, float* coef) {
    constexpr int NOUT = (CH == 11) ? 10 : 3;
    const float halfW = 0.5f * cam.W, halfH = 0.5f * cam.H;
    if constexpr (CH == 11 && FB_COEF11_DIRECT) {
        // The mean columns as for C_H = 4; the scale / rotation columns depend on phi only through the conic
        // (phi0..2): dL/dcov3D = N K e_i with K as in mean_chain_matrix and N the T-products of backward.cu:370-390,
        // then chain_scale_rot -- three short passes instead of five through the whole chain, and no 5 x 10 array
        // (which cost 708 bytes of spills at 80 registers).
        float M[3][5];
        mean_chain_matrix<FAST, IDENT>(ctx, cam, A, B, C, halfW, halfH, M);
        float Z[3][7];
        scale_rot_matrix(ctx, sr, Z);
#pragma unroll
        for (int m = 0; m < 13; ++m) coef[m] = 0.f;
        coef[0] = fdiv<FAST>(w[3], op * op);
        const int tab[5][5] = {{8, 9, 10, 4, 5}, {9, 10, 11, 5, 6}, {10, 11, 12, 6, 7}, {4, 5, 6, 1, 2}, {5, 6, 7, 2, 3}};
#pragma unroll
        for (int i = 0; i < 5; ++i)
#pragma unroll
            for (int j = i; j < 5; ++j) {
                float bsum = 0.f;
#pragma unroll
                for (int k = 0; k < 3; ++k) bsum += w[k] * M[k][i] * M[k][j];
                if (i < 3 && j < 3) {
#pragma unroll
                    for (int k = 0; k < 7; ++k) bsum += w[4 + k] * Z[i][k] * Z[j][k];
                }
                coef[tab[i][j]] += (i == j) ? bsum : 2.f * bsum;
            }
        return;
    }
    float col[5][NOUT];
    if (CH == 4) {
        float M[3][5];
        mean_chain_matrix<FAST, IDENT>(ctx, cam, A, B, C, halfW, halfH, M);
#pragma unroll
        for (int i = 0; i < 5; ++i) { col[i][0] = M[0][i]; col[i][1] = M[1][i]; col[i][2] = M[2][i]; }
    } else
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        float e[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
        e[i] = 1.f;
        float u[5];
        phi_to_u(e, A, B, C, halfW, halfH, u);
        ChainOut co = chain_geom(ctx, cam, u[0], u[1], u[2], u[3], u[4]);
        col[i][0] = co.mean[0]; col[i][1] = co.mean[1]; col[i][2] = co.mean[2];
        if (CH == 11) {
            float ds[3], dr[4];
            chain_scale_rot(sr, co.cov3D, ds, dr);
            col[i][3] = ds[0]; col[i][4] = ds[1]; col[i][5] = ds[2];
            col[i][6] = dr[0]; col[i][7] = dr[1]; col[i][8] = dr[2]; col[i][9] = dr[3];
        }
    }
    float wk[NOUT];
    wk[0] = w[0]; wk[1] = w[1]; wk[2] = w[2];
    if (CH == 11) {
#pragma unroll
        for (int k = 3; k < 10; ++k) wk[k] = w[k + 1];
    }
#pragma unroll
    for (int m = 0; m < 13; ++m) coef[m] = 0.f;
    coef[0] = fdiv<FAST>(w[3], op * op);
    const int tab[5][5] = {{8, 9, 10, 4, 5}, {9, 10, 11, 5, 6}, {10, 11, 12, 6, 7}, {4, 5, 6, 1, 2}, {5, 6, 7, 2, 3}};
    // C_H = 4: the columns are pre-multiplied by the weights (45 FMAs for the 15 products instead of 135 operations);
    // C_H = 11 keeps the three-factor form (a second 5x10 array would not fit the register budget)
    float colw[5][CH == 4 ? NOUT : 1];
    if (CH == 4) {
#pragma unroll
        for (int i = 0; i < 5; ++i)
#pragma unroll
            for (int k = 0; k < NOUT; ++k) colw[i][k] = wk[k] * col[i][k];
    }
#pragma unroll
    for (int i = 0; i < 5; ++i)
#pragma unroll
        for (int j = i; j < 5; ++j) {
            float b = 0.f;
#pragma unroll
            for (int k = 0; k < NOUT; ++k) b += (CH == 4) ? colw[i][CH == 4 ? k : 0] * col[j][k] : wk[k] * col[i][k] * col[j][k];
            coef[tab[i][j]] += (i == j) ? b : 2.f * b;
        }
}

// q * poly(dx,dy) with the 13 coefficients above (12 FMA + 3 MUL).
__device__ __forceinline__ float eval_poly13(const float* c, float dx, float dy) {
    const float dx2 = dx * dx, dxy = dx * dy, dy2 = dy * dy;
    float r = c[0];
    r += dx2 * (c[1] + dx * (c[4] + dx * c[8] + dy * c[9]) + dy * (c[5] + dy * c[10]));
    r += dxy * (c[2] + dy * (c[6] + dy * c[11]));
    r += dy2 * (c[3] + dy * (c[7] + dy * c[12]));
    return r;
}

}  // namespace fb
