// fisher_oracle.cpp -- CPU restatement of the reference's candidate-view scoring path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing in the product package (fisher_nerf_customized_b200/)
// may import, link or call this file; only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / reference legs use it, and only as the checker.
//
// Parity status: the reference ships NO tests, golden vectors or fixtures for this path
// (SURVEY.md section 4), so this restatement is pinned against OUTPUTS OF THE REFERENCE ITSELF:
// the unmodified reference CUDA extension is rebuilt for sm_100a (oracle/build_ref.py ->
// oracle/_ref/), run on the GPU box on seeded inputs by tests/golden/make_golden.py, and the
// resulting fixtures under tests/golden/ are checked against this file in tests/test_oracle.py.
//
// Every function cites the reference lines it follows.  DGR = thirdparty/
// diff-gaussian-rasterization-modified, DGR/cr = DGR/cuda_rasterizer, KNN = thirdparty/simple-knn.
//
// Floating-point contraction.  The reference is CUDA compiled by nvcc with the default
// -fmad=true, so a*b+c*d+e*f is NOT three roundings: nvcc fuses some of the multiplies into
// FMAs.  For the arithmetic that decides INTEGER outputs (radii, tile rectangles, depth key
// bits, n_contrib) this file spells out the exact fused sequence that nvcc 12.9 -O3 emits for
// the reference source (read from the SASS of the reference's forward.cu built here -- cuobjdump -sass; ptxas
// fuses mul+add/sub pairs that the PTX still shows separately, so the PTX alone is not enough).
// It must therefore be compiled with -ffp-contract=off so that the host compiler adds none of
// its own.  The observed rule for a three-term dot product  a0*b0 + a1*b1 + a2*b2  is
//     t = a1*b1;  t = fma(a0, b0, t);  t = fma(a2, b2, t)                      ("dot3")
// and a trailing "+ c" is a separate rounded add.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <algorithm>
#include <vector>

namespace {

inline float fma_(float a, float b, float c) { return std::fmaf(a, b, c); }
inline float dot3(float a0, float b0, float a1, float b1, float a2, float b2) {
    float t = a1 * b1;
    t = fma_(a0, b0, t);
    t = fma_(a2, b2, t);
    return t;
}

struct Camera {   // same field order as fb_camera (include/fisher_b200.h)
    int   W, H;
    float tanfovx, tanfovy;
    float scale_modifier;
    int   D, M, prefiltered;
    float bg[3];
    float campos[3];
    float view[16];
    float proj[16];
};

const int BLOCK_X = 16, BLOCK_Y = 16, BLOCK_SIZE = 256;   // DGR/cr/config.h:15-17

struct f3 { float x, y, z; };
struct f4 { float x, y, z, w; };

// DGR/cr/auxiliary.h:58-66 (matrix consumed column-major).  Fusion as compiled: see header.
inline f3 transformPoint4x3(const f3& p, const float* m) {
    f3 r;
    r.x = m[12] + dot3(m[0], p.x, m[4], p.y, m[8], p.z);
    r.y = m[13] + dot3(m[1], p.x, m[5], p.y, m[9], p.z);
    r.z = m[14] + dot3(m[2], p.x, m[6], p.y, m[10], p.z);
    return r;
}
// DGR/cr/auxiliary.h:68-77
inline f4 transformPoint4x4(const f3& p, const float* m) {
    f4 r;
    r.x = m[12] + dot3(m[0], p.x, m[4], p.y, m[8], p.z);
    r.y = m[13] + dot3(m[1], p.x, m[5], p.y, m[9], p.z);
    r.z = m[14] + dot3(m[2], p.x, m[6], p.y, m[10], p.z);
    r.w = m[15] + dot3(m[3], p.x, m[7], p.y, m[11], p.z);
    return r;
}
// DGR/cr/auxiliary.h:89-97
inline f3 transformVec4x3Transpose(const f3& p, const float* m) {
    f3 r;
    r.x = m[0] * p.x + m[1] * p.y + m[2] * p.z;
    r.y = m[4] * p.x + m[5] * p.y + m[6] * p.z;
    r.z = m[8] * p.x + m[9] * p.y + m[10] * p.z;
    return r;
}
// DGR/cr/auxiliary.h:41-44: the literals are double, so the expression is evaluated in FP64
// ((v+1.0)*S-1.0 compiles to one fma.rn.f64) and rounded to FP32 once.
inline float ndc2Pix(float v, int S) {
    double d = std::fma((double)v + 1.0, (double)S, -1.0) * 0.5;
    return (float)d;
}
// DGR/cr/auxiliary.h:46-56
inline void getRect(float px, float py, int max_radius, int gx, int gy, int& x0, int& y0, int& x1, int& y1) {
    x0 = std::min(gx, std::max(0, (int)((px - max_radius) / BLOCK_X)));
    y0 = std::min(gy, std::max(0, (int)((py - max_radius) / BLOCK_Y)));
    x1 = std::min(gx, std::max(0, (int)((px + max_radius + BLOCK_X - 1) / BLOCK_X)));
    y1 = std::min(gy, std::max(0, (int)((py + max_radius + BLOCK_Y - 1) / BLOCK_Y)));
}

// Column-major 3x3 with GLM's indexing m[col][row].
struct mat3 { float m[3][3]; };
inline mat3 transpose(const mat3& a) {
    mat3 r;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r.m[i][j] = a.m[j][i];
    return r;
}
// GLM mat3*mat3 (DGR/third_party/glm/glm/detail/type_mat3x3.inl, operator*):
// R[col][row] = A[0][row]*B[col][0] + A[1][row]*B[col][1] + A[2][row]*B[col][2], fused as dot3.
inline mat3 mul(const mat3& A, const mat3& B) {
    mat3 R;
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r)
            R.m[c][r] = dot3(A.m[0][r], B.m[c][0], A.m[1][r], B.m[c][1], A.m[2][r], B.m[c][2]);
    return R;
}

// Rotation matrix of the UN-normalised quaternion (r,x,y,z); DGR/cr/forward.cu:126-138 (fwd) and
// DGR/cr/backward.cu:415-425 (bwd).  As compiled (SASS): of each +/- pair of products one is rounded and
// shared, the other fused: yy+zz is mul,mul,add; xx+zz = fma(x,x,zz); xx+yy = fma(x,x,yy);
// xy-/+rz = fma(x,y,-/+rz); yz-/+rx = fma(y,z,-/+rx); xz+ry = fma(r,y,xz); xz-ry = fma(-r,y,xz).
inline mat3 quat_R(float r, float x, float y, float z) {
    mat3 R;
    const float yy = y * y, zz = z * z, rz = r * z, xz = x * z, rx = r * x;
    R.m[0][0] = 1.f - 2.f * (yy + zz);
    R.m[0][1] = 2.f * fma_(x, y, -rz);
    R.m[0][2] = 2.f * fma_(r, y, xz);
    R.m[1][0] = 2.f * fma_(x, y, rz);
    R.m[1][1] = 1.f - 2.f * fma_(x, x, zz);
    R.m[1][2] = 2.f * fma_(y, z, -rx);
    R.m[2][0] = 2.f * fma_(-r, y, xz);
    R.m[2][1] = 2.f * fma_(y, z, rx);
    R.m[2][2] = 1.f - 2.f * fma_(x, x, yy);
    return R;
}

// DGR/cr/forward.cu:118-152
inline void computeCov3D(const float* scale, float mod, const float* rot, float* cov3D) {
    mat3 S; std::memset(&S, 0, sizeof(S));
    S.m[0][0] = mod * scale[0]; S.m[1][1] = mod * scale[1]; S.m[2][2] = mod * scale[2];
    mat3 R = quat_R(rot[0], rot[1], rot[2], rot[3]);
    mat3 M = mul(S, R);
    mat3 Sigma = mul(transpose(M), M);
    cov3D[0] = Sigma.m[0][0]; cov3D[1] = Sigma.m[0][1]; cov3D[2] = Sigma.m[0][2];
    cov3D[3] = Sigma.m[1][1]; cov3D[4] = Sigma.m[1][2]; cov3D[5] = Sigma.m[2][2];
}

// Shared by forward (DGR/cr/forward.cu:74-113) and backward (DGR/cr/backward.cu:300-330).
struct Cov2DState {
    f3 t; float txtz, tytz;
    mat3 J, Wm, T, Vrk, cov;
};
inline Cov2DState cov2d_state(const f3& mean, float fx, float fy, float tan_fovx, float tan_fovy, const float* cov3D,
                              const float* view) {
    Cov2DState s;
    f3 t = transformPoint4x3(mean, view);
    const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
    s.txtz = t.x / t.z; s.tytz = t.y / t.z;
    t.x = std::min(limx, std::max(-limx, s.txtz)) * t.z;
    t.y = std::min(limy, std::max(-limy, s.tytz)) * t.z;
    s.t = t;
    std::memset(&s.J, 0, sizeof(mat3));
    s.J.m[0][0] = fx / t.z; s.J.m[0][2] = -(fx * t.x) / (t.z * t.z);
    s.J.m[1][1] = fy / t.z; s.J.m[1][2] = -(fy * t.y) / (t.z * t.z);
    mat3& W = s.Wm;
    W.m[0][0] = view[0]; W.m[0][1] = view[4]; W.m[0][2] = view[8];
    W.m[1][0] = view[1]; W.m[1][1] = view[5]; W.m[1][2] = view[9];
    W.m[2][0] = view[2]; W.m[2][1] = view[6]; W.m[2][2] = view[10];
    s.T = mul(W, s.J);
    mat3& V = s.Vrk;
    V.m[0][0] = cov3D[0]; V.m[0][1] = cov3D[1]; V.m[0][2] = cov3D[2];
    V.m[1][0] = cov3D[1]; V.m[1][1] = cov3D[3]; V.m[1][2] = cov3D[4];
    V.m[2][0] = cov3D[2]; V.m[2][1] = cov3D[4]; V.m[2][2] = cov3D[5];
    s.cov = mul(mul(transpose(s.T), transpose(V)), s.T);
    s.cov.m[0][0] += 0.3f;
    s.cov.m[1][1] += 0.3f;
    return s;
}

const float SH_C0 = 0.28209479177387814f, SH_C1 = 0.4886025119029199f;
const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f,
                        0.5462742152960396f};
const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                        -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

// DGR/cr/forward.cu:20-71 (per channel; floating point only, plain rounding)
inline void computeColorFromSH(int idx, int deg, int M, const float* means, const float* campos, const float* shs,
                               uint8_t* clamped, float* out3) {
    float dx = means[3 * idx] - campos[0], dy = means[3 * idx + 1] - campos[1], dz = means[3 * idx + 2] - campos[2];
    float len = std::sqrt(dx * dx + dy * dy + dz * dz);
    float x = dx / len, y = dy / len, z = dz / len;
    const float* sh = shs + (size_t)idx * M * 3;
    for (int ch = 0; ch < 3; ++ch) {
        auto S = [&](int k) { return sh[3 * k + ch]; };
        float r = SH_C0 * S(0);
        if (deg > 0) {
            r = r - SH_C1 * y * S(1) + SH_C1 * z * S(2) - SH_C1 * x * S(3);
            if (deg > 1) {
                float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                r = r + SH_C2[0] * xy * S(4) + SH_C2[1] * yz * S(5) + SH_C2[2] * (2.0f * zz - xx - yy) * S(6) +
                    SH_C2[3] * xz * S(7) + SH_C2[4] * (xx - yy) * S(8);
                if (deg > 2) {
                    r = r + SH_C3[0] * y * (3.0f * xx - yy) * S(9) + SH_C3[1] * xy * z * S(10) +
                        SH_C3[2] * y * (4.0f * zz - xx - yy) * S(11) +
                        SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
                        SH_C3[4] * x * (4.0f * zz - xx - yy) * S(13) + SH_C3[5] * z * (xx - yy) * S(14) +
                        SH_C3[6] * x * (xx - 3.0f * yy) * S(15);
                }
            }
        }
        r += 0.5f;
        clamped[3 * idx + ch] = (r < 0);
        out3[ch] = std::max(r, 0.0f);
    }
}

}  // namespace

// ================================================================================================
// Forward: Rasterizer::forward  DGR/cr/rasterizer_impl.cu:198-339
// ================================================================================================
namespace {

struct State {
    Camera cam;
    int P = 0;
    float focal_x = 0, focal_y = 0;
    int grid_x = 0, grid_y = 0;
    // inputs kept for backward (the reference re-reads them from the caller's tensors)
    std::vector<float> means3D, opacities, shs, colors_in, scales, rotations, cov3D_in;
    bool has_sh = false, has_colors = false, has_sr = false, has_cov = false;
    // GeometryState (rasterizer_impl.h:30-46)
    std::vector<int> radii;
    std::vector<uint32_t> tiles_touched, point_offsets;
    std::vector<float> depths, means2D, cov3D, conic_opacity, rgb;
    std::vector<uint8_t> clamped;
    // BinningState (rasterizer_impl.h:59-70)
    std::vector<uint64_t> keys;      // sorted
    std::vector<uint32_t> point_list;
    // ImageState (rasterizer_impl.h:48-57)
    std::vector<uint32_t> ranges;    // [tiles][2]
    std::vector<float> final_T, out_color, out_depth;
    std::vector<uint32_t> n_contrib;
    long long num_rendered = 0;
    int prefiltered_trap = 0;
};

// preprocessCUDA<3>  DGR/cr/forward.cu:155-256
void preprocess(State& s) {
    const Camera& c = s.cam;
    const int P = s.P;
    for (int idx = 0; idx < P; ++idx) {
        s.radii[idx] = 0;
        s.tiles_touched[idx] = 0;
        // in_frustum  DGR/cr/auxiliary.h:139-164
        f3 p_orig = {s.means3D[3 * idx], s.means3D[3 * idx + 1], s.means3D[3 * idx + 2]};
        f3 p_view = transformPoint4x3(p_orig, c.view);
        if (p_view.z <= 0.001f) {
            if (c.prefiltered) s.prefiltered_trap = 1;   // device printf + __trap() in the reference
            continue;
        }
        f4 p_hom = transformPoint4x4(p_orig, c.proj);
        float p_w = 1.0f / (p_hom.w + 0.0000001f);
        float projx = p_hom.x * p_w, projy = p_hom.y * p_w;
        const float* cov3D;
        if (s.has_cov) {
            cov3D = &s.cov3D_in[6 * (size_t)idx];
        } else {
            computeCov3D(&s.scales[3 * (size_t)idx], c.scale_modifier, &s.rotations[4 * (size_t)idx], &s.cov3D[6 * (size_t)idx]);
            cov3D = &s.cov3D[6 * (size_t)idx];
        }
        Cov2DState c2 = cov2d_state(p_orig, s.focal_x, s.focal_y, c.tanfovx, c.tanfovy, cov3D, c.view);
        const float cx = c2.cov.m[0][0], cy = c2.cov.m[0][1], cz = c2.cov.m[1][1];
        // as compiled (SASS): det = fma(x, z, -(y*y)); mid*mid - det = fma(mid, mid, -det)
        const float cyy = cy * cy;
        float det = fma_(cx, cz, -cyy);
        if (det == 0.0f) continue;
        float det_inv = 1.f / det;
        float conic[3] = {cz * det_inv, -cy * det_inv, cx * det_inv};
        float mid = 0.5f * (cx + cz);
        float disc = std::max(0.1f, fma_(mid, mid, -det));
        float lambda1 = mid + std::sqrt(disc);
        float lambda2 = mid - std::sqrt(disc);
        float my_radius = std::ceil(3.f * std::sqrt(std::max(lambda1, lambda2)));
        float px = ndc2Pix(projx, c.W), py = ndc2Pix(projy, c.H);
        int x0, y0, x1, y1;
        getRect(px, py, (int)my_radius, s.grid_x, s.grid_y, x0, y0, x1, y1);
        if ((x1 - x0) * (y1 - y0) == 0) continue;
        if (!s.has_colors) computeColorFromSH(idx, c.D, c.M, s.means3D.data(), c.campos, s.shs.data(), s.clamped.data(), &s.rgb[3 * (size_t)idx]);
        s.depths[idx] = p_view.z;
        s.radii[idx] = (int)my_radius;
        s.means2D[2 * idx] = px; s.means2D[2 * idx + 1] = py;
        s.conic_opacity[4 * idx] = conic[0]; s.conic_opacity[4 * idx + 1] = conic[1];
        s.conic_opacity[4 * idx + 2] = conic[2]; s.conic_opacity[4 * idx + 3] = s.opacities[idx];
        s.tiles_touched[idx] = (uint32_t)((y1 - y0) * (x1 - x0));
    }
}

// InclusiveSum (rasterizer_impl.cu:277), duplicateWithKeys (:70-111), SortPairs (:304-309, stable,
// ascending, low 32+bit key bits), identifyTileRanges (:116-138).
void bin_and_sort(State& s) {
    const int P = s.P;
    uint32_t run = 0;
    for (int i = 0; i < P; ++i) { run += s.tiles_touched[i]; s.point_offsets[i] = run; }
    const size_t R = P ? run : 0;
    s.num_rendered = (long long)R;
    std::vector<uint64_t> keys_unsorted(R);
    std::vector<uint32_t> vals_unsorted(R);
    for (int idx = 0; idx < P; ++idx) {
        if (s.radii[idx] > 0) {
            uint32_t off = (idx == 0) ? 0 : s.point_offsets[idx - 1];
            int x0, y0, x1, y1;
            getRect(s.means2D[2 * idx], s.means2D[2 * idx + 1], s.radii[idx], s.grid_x, s.grid_y, x0, y0, x1, y1);
            for (int y = y0; y < y1; ++y)
                for (int x = x0; x < x1; ++x) {
                    uint64_t key = (uint64_t)(y * s.grid_x + x);
                    key <<= 32;
                    uint32_t db; std::memcpy(&db, &s.depths[idx], 4);
                    key |= db;
                    keys_unsorted[off] = key;
                    vals_unsorted[off] = (uint32_t)idx;
                    off++;
                }
        }
    }
    // getHigherMsb(tile_grid.x * tile_grid.y) (rasterizer_impl.cu:35-50): the sort covers bits
    // [0, 32+bit).  Tile ids are < tiles <= 2^bit, so the masked key equals the full key.
    std::vector<uint32_t> order(R);
    for (size_t i = 0; i < R; ++i) order[i] = (uint32_t)i;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return keys_unsorted[a] < keys_unsorted[b]; });
    s.keys.resize(R); s.point_list.resize(R);
    for (size_t i = 0; i < R; ++i) { s.keys[i] = keys_unsorted[order[i]]; s.point_list[i] = vals_unsorted[order[i]]; }
    const int tiles = s.grid_x * s.grid_y;
    std::fill(s.ranges.begin(), s.ranges.end(), 0u);   // cudaMemset, rasterizer_impl.cu:311
    for (size_t i = 0; i < R; ++i) {
        uint32_t tile = (uint32_t)(s.keys[i] >> 32);
        if ((int)tile >= tiles) continue;
        if (i == 0) s.ranges[2 * tile] = 0;
        else {
            uint32_t prev = (uint32_t)(s.keys[i - 1] >> 32);
            if (tile != prev) { s.ranges[2 * prev + 1] = (uint32_t)i; s.ranges[2 * tile] = (uint32_t)i; }
        }
        if (i == R - 1) s.ranges[2 * tile + 1] = (uint32_t)R;
    }
}

// The pair exponent of renderCUDA / renderCUDAFused (forward.cu:341, backward.cu:968), fused as
// nvcc compiles it:  -0.5f*(A*dx*dx + C*dy*dy) - B*dx*dy  ->
//   q = fma(dx, dx*A, dy*(dy*C));  power = fma(q, -0.5f, -(dy*(dx*B)))      (SASS)
inline float pair_power(float A, float B, float C, float dx, float dy) {
    const float t0 = dx * A;
    const float t1 = dy * (dy * C);
    const float q = fma_(dx, t0, t1);
    const float t2 = dy * (dx * B);
    return fma_(q, -0.5f, -t2);
}

// renderCUDA<3>  DGR/cr/forward.cu:261-393.  One pixel at a time; the batch/`done` voting of the
// CUDA kernel only decides when a whole tile stops fetching and does not change any pixel's result.
void render(State& s) {
    const Camera& c = s.cam;
    const int W = c.W, H = c.H;
    const float* features = s.has_colors ? s.colors_in.data() : s.rgb.data();
    for (int py = 0; py < H; ++py)
        for (int px = 0; px < W; ++px) {
            const int tile = (py / BLOCK_Y) * s.grid_x + (px / BLOCK_X);
            const uint32_t r0 = s.ranges[2 * tile], r1 = s.ranges[2 * tile + 1];
            const float pxf = (float)px, pyf = (float)py;
            float T = 1.0f;
            uint32_t contributor = 0, last_contributor = 0;
            float C[3] = {0, 0, 0};
            float D = 15.0f;
            for (uint32_t i = r0; i < r1; ++i) {
                contributor++;
                const uint32_t id = s.point_list[i];
                const float dx = s.means2D[2 * id] - pxf, dy = s.means2D[2 * id + 1] - pyf;
                const float* co = &s.conic_opacity[4 * (size_t)id];
                const float power = pair_power(co[0], co[1], co[2], dx, dy);
                if (power > 0.0f) continue;
                const float alpha = std::min(0.99f, co[3] * std::exp(power));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = T * (1 - alpha);
                if (test_T < 0.0001f) break;   // done = true
                for (int ch = 0; ch < 3; ++ch) C[ch] = fma_(T, alpha * features[3 * (size_t)id + ch], C[ch]);
                if (T > 0.5f && (double)test_T < 0.5) D = s.depths[id];
                T = test_T;
                last_contributor = contributor;
            }
            const int pix = W * py + px;
            s.final_T[pix] = T;
            s.n_contrib[pix] = last_contributor;
            for (int ch = 0; ch < 3; ++ch) s.out_color[(size_t)ch * H * W + pix] = fma_(T, c.bg[ch], C[ch]);
            s.out_depth[pix] = D;
        }
}

}  // namespace

// ================================================================================================
// Backward: Rasterizer::backward -> BACKWARD::renderFused -> renderCUDAFused<3>
//   DGR/cr/rasterizer_impl.cu:343-434, DGR/cr/backward.cu:850-1140
// Floating point only: plain IEEE FP32 per pair (contraction off), per-Gaussian sums kept in
// FP64 (the reference's float atomicAdd order is non-deterministic; FP64 is the order-free value
// every FP32 summation order approximates to ~1e-6 relative).
// ================================================================================================
namespace {

struct Grads {   // per Gaussian, FP64 accumulators, reference shapes
    std::vector<double> mean2D, conic, opacity, colors, mean3D, cov3D, sh, scale, rot;
};

// computeCov2DCUDARelocated  DGR/cr/backward.cu:276-408 (one pair; `idx = 0` in the reference)
void cov2d_backward(const State& s, int gid, const float* dL_dconic4, f3& dL_dmeans, float* dL_dcov6) {
    const Camera& c = s.cam;
    const float* cov3D = s.has_cov ? &s.cov3D_in[6 * (size_t)gid] : &s.cov3D[6 * (size_t)gid];
    f3 mean = {s.means3D[3 * gid], s.means3D[3 * gid + 1], s.means3D[3 * gid + 2]};
    const float dcx = dL_dconic4[0], dcy = dL_dconic4[1], dcz = dL_dconic4[3];
    Cov2DState st = cov2d_state(mean, s.focal_x, s.focal_y, c.tanfovx, c.tanfovy, cov3D, c.view);
    const float limx = 1.3f * c.tanfovx, limy = 1.3f * c.tanfovy;
    const float x_grad_mul = (st.txtz < -limx || st.txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (st.tytz < -limy || st.tytz > limy) ? 0.f : 1.f;
    const mat3& T = st.T; const mat3& Vrk = st.Vrk; const mat3& W = st.Wm;
    const f3 t = st.t;
    const float h_x = s.focal_x, h_y = s.focal_y;
    float a = st.cov.m[0][0], b = st.cov.m[0][1], cc = st.cov.m[1][1];
    float denom = a * cc - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    if (denom2inv != 0) {
        dL_da = denom2inv * (-cc * cc * dcx + 2 * b * cc * dcy + (denom - a * cc) * dcz);
        dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * cc) * dcx);
        dL_db = denom2inv * 2 * (b * cc * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
        dL_dcov6[0] = (T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db + T.m[1][0] * T.m[1][0] * dL_dc);
        dL_dcov6[3] = (T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db + T.m[1][1] * T.m[1][1] * dL_dc);
        dL_dcov6[5] = (T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db + T.m[1][2] * T.m[1][2] * dL_dc);
        dL_dcov6[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db + 2 * T.m[1][0] * T.m[1][1] * dL_dc;
        dL_dcov6[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db + 2 * T.m[1][0] * T.m[1][2] * dL_dc;
        dL_dcov6[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db + 2 * T.m[1][1] * T.m[1][2] * dL_dc;
    } else {
        for (int i = 0; i < 6; i++) dL_dcov6[i] = 0;
    }
    float dL_dT00 = 2 * (T.m[0][0] * Vrk.m[0][0] + T.m[0][1] * Vrk.m[0][1] + T.m[0][2] * Vrk.m[0][2]) * dL_da +
                    (T.m[1][0] * Vrk.m[0][0] + T.m[1][1] * Vrk.m[0][1] + T.m[1][2] * Vrk.m[0][2]) * dL_db;
    float dL_dT01 = 2 * (T.m[0][0] * Vrk.m[1][0] + T.m[0][1] * Vrk.m[1][1] + T.m[0][2] * Vrk.m[1][2]) * dL_da +
                    (T.m[1][0] * Vrk.m[1][0] + T.m[1][1] * Vrk.m[1][1] + T.m[1][2] * Vrk.m[1][2]) * dL_db;
    float dL_dT02 = 2 * (T.m[0][0] * Vrk.m[2][0] + T.m[0][1] * Vrk.m[2][1] + T.m[0][2] * Vrk.m[2][2]) * dL_da +
                    (T.m[1][0] * Vrk.m[2][0] + T.m[1][1] * Vrk.m[2][1] + T.m[1][2] * Vrk.m[2][2]) * dL_db;
    float dL_dT10 = 2 * (T.m[1][0] * Vrk.m[0][0] + T.m[1][1] * Vrk.m[0][1] + T.m[1][2] * Vrk.m[0][2]) * dL_dc +
                    (T.m[0][0] * Vrk.m[0][0] + T.m[0][1] * Vrk.m[0][1] + T.m[0][2] * Vrk.m[0][2]) * dL_db;
    float dL_dT11 = 2 * (T.m[1][0] * Vrk.m[1][0] + T.m[1][1] * Vrk.m[1][1] + T.m[1][2] * Vrk.m[1][2]) * dL_dc +
                    (T.m[0][0] * Vrk.m[1][0] + T.m[0][1] * Vrk.m[1][1] + T.m[0][2] * Vrk.m[1][2]) * dL_db;
    float dL_dT12 = 2 * (T.m[1][0] * Vrk.m[2][0] + T.m[1][1] * Vrk.m[2][1] + T.m[1][2] * Vrk.m[2][2]) * dL_dc +
                    (T.m[0][0] * Vrk.m[2][0] + T.m[0][1] * Vrk.m[2][1] + T.m[0][2] * Vrk.m[2][2]) * dL_db;
    float dL_dJ00 = W.m[0][0] * dL_dT00 + W.m[0][1] * dL_dT01 + W.m[0][2] * dL_dT02;
    float dL_dJ02 = W.m[2][0] * dL_dT00 + W.m[2][1] * dL_dT01 + W.m[2][2] * dL_dT02;
    float dL_dJ11 = W.m[1][0] * dL_dT10 + W.m[1][1] * dL_dT11 + W.m[1][2] * dL_dT12;
    float dL_dJ12 = W.m[2][0] * dL_dT10 + W.m[2][1] * dL_dT11 + W.m[2][2] * dL_dT12;
    float tz = 1.f / t.z;
    float tz2 = tz * tz;
    float tz3 = tz2 * tz;
    float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12;
    dL_dmeans = transformVec4x3Transpose({dL_dtx, dL_dty, dL_dtz}, c.view);   // ASSIGNED (backward.cu:407)
}

// computeCov3D (backward)  DGR/cr/backward.cu:412-475
void cov3d_backward(const float* scale, float mod, const float* rot, const float* dL_dcov3D, float* dL_dscale3, float* dL_drot4) {
    const float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
    mat3 R = quat_R(r, x, y, z);
    mat3 S; std::memset(&S, 0, sizeof(S));
    const float sx = mod * scale[0], sy = mod * scale[1], sz = mod * scale[2];
    S.m[0][0] = sx; S.m[1][1] = sy; S.m[2][2] = sz;
    mat3 M = mul(S, R);
    mat3 dS;
    dS.m[0][0] = dL_dcov3D[0]; dS.m[0][1] = 0.5f * dL_dcov3D[1]; dS.m[0][2] = 0.5f * dL_dcov3D[2];
    dS.m[1][0] = 0.5f * dL_dcov3D[1]; dS.m[1][1] = dL_dcov3D[3]; dS.m[1][2] = 0.5f * dL_dcov3D[4];
    dS.m[2][0] = 0.5f * dL_dcov3D[2]; dS.m[2][1] = 0.5f * dL_dcov3D[4]; dS.m[2][2] = dL_dcov3D[5];
    mat3 M2;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) M2.m[i][j] = 2.0f * M.m[i][j];
    // plain-rounded product here (floating-point output only)
    mat3 dL_dM;
    for (int c = 0; c < 3; ++c) for (int rr = 0; rr < 3; ++rr)
        dL_dM.m[c][rr] = M2.m[0][rr] * dS.m[c][0] + M2.m[1][rr] * dS.m[c][1] + M2.m[2][rr] * dS.m[c][2];
    mat3 Rt = transpose(R);
    mat3 dMt = transpose(dL_dM);
    for (int k = 0; k < 3; ++k)
        dL_dscale3[k] = Rt.m[k][0] * dMt.m[k][0] + Rt.m[k][1] * dMt.m[k][1] + Rt.m[k][2] * dMt.m[k][2];
    for (int j = 0; j < 3; ++j) { dMt.m[0][j] *= sx; dMt.m[1][j] *= sy; dMt.m[2][j] *= sz; }
    dL_drot4[0] = 2 * z * (dMt.m[0][1] - dMt.m[1][0]) + 2 * y * (dMt.m[2][0] - dMt.m[0][2]) + 2 * x * (dMt.m[1][2] - dMt.m[2][1]);
    dL_drot4[1] = 2 * y * (dMt.m[1][0] + dMt.m[0][1]) + 2 * z * (dMt.m[2][0] + dMt.m[0][2]) + 2 * r * (dMt.m[1][2] - dMt.m[2][1]) - 4 * x * (dMt.m[2][2] + dMt.m[1][1]);
    dL_drot4[2] = 2 * x * (dMt.m[1][0] + dMt.m[0][1]) + 2 * r * (dMt.m[2][0] - dMt.m[0][2]) + 2 * z * (dMt.m[1][2] + dMt.m[2][1]) - 4 * y * (dMt.m[2][2] + dMt.m[0][0]);
    dL_drot4[3] = 2 * r * (dMt.m[0][1] - dMt.m[1][0]) + 2 * x * (dMt.m[2][0] + dMt.m[0][2]) + 2 * y * (dMt.m[1][2] + dMt.m[2][1]) - 4 * z * (dMt.m[1][1] + dMt.m[0][0]);
}

}  // namespace

namespace {

inline double powd(float v, int power) {
    // powf(x, (float)grad_power) (backward.cu:1093-1137) for the integer powers the callers use.
    if (power == 1) return (double)v;
    if (power == 2) return (double)(v * v);
    return (double)std::pow(v, (float)power);
}

// renderCUDAFused<3>  DGR/cr/backward.cu:850-1140, pixel by pixel.
int backward(const State& s, const float* dL_dpixels, int power, Grads& g) {
    const Camera& c = s.cam;
    const int W = c.W, H = c.H, P = s.P;
    if (s.has_sh && c.D > 0) return -1;   // SH backward of degree > 0: not restated
    g.mean2D.assign((size_t)P * 3, 0.0); g.conic.assign((size_t)P * 4, 0.0); g.opacity.assign(P, 0.0);
    g.colors.assign((size_t)P * 3, 0.0); g.mean3D.assign((size_t)P * 3, 0.0); g.cov3D.assign((size_t)P * 6, 0.0);
    g.scale.assign((size_t)P * 3, 0.0); g.rot.assign((size_t)P * 4, 0.0);
    const float* colors = s.has_colors ? s.colors_in.data() : s.rgb.data();
    const float ddelx_dx = 0.5 * W, ddely_dy = 0.5 * H;
    for (int py = 0; py < H; ++py)
        for (int px = 0; px < W; ++px) {
            const int tile = (py / BLOCK_Y) * s.grid_x + (px / BLOCK_X);
            const uint32_t r0 = s.ranges[2 * tile], r1 = s.ranges[2 * tile + 1];
            const int pix = W * py + px;
            const float pxf = (float)px, pyf = (float)py;
            const float T_final = s.final_T[pix];
            float T = T_final;
            uint32_t contributor = r1 - r0;
            const uint32_t last_contributor = s.n_contrib[pix];
            float accum_rec[3] = {0, 0, 0}, dL_dpixel[3], last_color[3] = {0, 0, 0};
            for (int i = 0; i < 3; ++i) dL_dpixel[i] = dL_dpixels[(size_t)i * H * W + pix];
            float last_alpha = 0;
            for (uint32_t k = 0; k < r1 - r0; ++k) {
                contributor--;
                if (contributor >= last_contributor) continue;
                const uint32_t gid = s.point_list[r1 - k - 1];
                const float dx = s.means2D[2 * gid] - pxf, dy = s.means2D[2 * gid + 1] - pyf;
                const float* co = &s.conic_opacity[4 * (size_t)gid];
                const float pw = pair_power(co[0], co[1], co[2], dx, dy);
                if (pw > 0.0f) continue;
                const float G = std::exp(pw);
                const float alpha = std::min(0.99f, co[3] * G);
                if (alpha < 1.0f / 255.0f) continue;
                T = T / (1.f - alpha);
                const float dchannel_dcolor = alpha * T;
                float cur_dcolors[3];
                float dL_dalpha = 0.0f;
                for (int ch = 0; ch < 3; ++ch) {
                    const float col = colors[3 * (size_t)gid + ch];
                    accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                    last_color[ch] = col;
                    const float dL_dchannel = dL_dpixel[ch];
                    dL_dalpha += (col - accum_rec[ch]) * dL_dchannel;
                    cur_dcolors[ch] = dchannel_dcolor * dL_dchannel;
                }
                dL_dalpha *= T;
                last_alpha = alpha;
                float bg_dot_dpixel = 0;
                for (int i = 0; i < 3; ++i) bg_dot_dpixel += c.bg[i] * dL_dpixel[i];
                dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
                const float dL_dG = co[3] * dL_dalpha;
                const float gdx = G * dx, gdy = G * dy;
                const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                const float dG_ddely = -gdy * co[2] - gdx * co[1];
                float cur_mean2D[2] = {dL_dG * dG_ddelx * ddelx_dx, dL_dG * dG_ddely * ddely_dy};
                float cur_conic[4] = {-0.5f * gdx * dx * dL_dG, -0.5f * gdx * dy * dL_dG, 0.f, -0.5f * gdy * dy * dL_dG};
                const float cur_dopacity = G * dL_dalpha;
                f3 cur_means; float cur_cov3D[6];
                cov2d_backward(s, (int)gid, cur_conic, cur_means, cur_cov3D);
                // preprocessCUDARelocated  backward.cu:532-583
                {
                    f3 m = {s.means3D[3 * gid], s.means3D[3 * gid + 1], s.means3D[3 * gid + 2]};
                    const float* proj = c.proj;
                    f4 m_hom = transformPoint4x4(m, proj);
                    float m_w = 1.0f / (m_hom.w + 0.0000001f);
                    float mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
                    float mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
                    cur_means.x += (proj[0] * m_w - proj[3] * mul1) * cur_mean2D[0] + (proj[1] * m_w - proj[3] * mul2) * cur_mean2D[1];
                    cur_means.y += (proj[4] * m_w - proj[7] * mul1) * cur_mean2D[0] + (proj[5] * m_w - proj[7] * mul2) * cur_mean2D[1];
                    cur_means.z += (proj[8] * m_w - proj[11] * mul1) * cur_mean2D[0] + (proj[9] * m_w - proj[11] * mul2) * cur_mean2D[1];
                }
                float cur_scale[3] = {0, 0, 0}, cur_rot[4] = {0, 0, 0, 0};
                if (s.has_sr)
                    cov3d_backward(&s.scales[3 * (size_t)gid], c.scale_modifier, &s.rotations[4 * (size_t)gid], cur_cov3D, cur_scale, cur_rot);
                for (int ch = 0; ch < 3; ++ch) g.colors[3 * (size_t)gid + ch] += powd(cur_dcolors[ch], power);
                g.mean2D[3 * (size_t)gid] += powd(cur_mean2D[0], power);
                g.mean2D[3 * (size_t)gid + 1] += powd(cur_mean2D[1], power);
                g.conic[4 * (size_t)gid] += powd(cur_conic[0], power);
                g.conic[4 * (size_t)gid + 1] += powd(cur_conic[1], power);
                g.conic[4 * (size_t)gid + 3] += powd(cur_conic[3], power);
                g.mean3D[3 * (size_t)gid] += powd(cur_means.x, power);
                g.mean3D[3 * (size_t)gid + 1] += powd(cur_means.y, power);
                g.mean3D[3 * (size_t)gid + 2] += powd(cur_means.z, power);
                for (int ch = 0; ch < 6; ++ch) g.cov3D[6 * (size_t)gid + ch] += powd(cur_cov3D[ch], power);
                for (int ch = 0; ch < 3; ++ch) g.scale[3 * (size_t)gid + ch] += powd(cur_scale[ch], power);
                for (int ch = 0; ch < 4; ++ch) g.rot[4 * (size_t)gid + ch] += powd(cur_rot[ch], power);
                g.opacity[gid] += powd(cur_dopacity, power);
            }
        }
    return 0;
}

template <class T> void copy_out(const std::vector<T>& v, T* dst) { if (dst && !v.empty()) std::memcpy(dst, v.data(), v.size() * sizeof(T)); }
void copy_f(const std::vector<double>& v, float* dst) { if (dst) for (size_t i = 0; i < v.size(); ++i) dst[i] = (float)v[i]; }

}  // namespace

// ================================================================================================
// C entry points (loaded by oracle/oracle.py through ctypes)
// ================================================================================================
extern "C" {

// Runs preprocess + binning + sort + ranges + render and returns an opaque handle (NULL on bad
// arguments).  Pointers follow RasterizeGaussiansCUDA's argument meaning (DGR/rasterize_points.cu:35-115);
// absent optional inputs are NULL.
void* orc_forward(const void* cam_, int P, const float* means3D, const float* opacities, const float* shs,
                  const float* colors_precomp, const float* scales, const float* rotations, const float* cov3D_precomp) {
    const Camera* cam = (const Camera*)cam_;
    if (!cam || P < 0) return nullptr;
    if (P > 0 && (!means3D || !opacities)) return nullptr;
    if (P > 0 && ((shs == nullptr) == (colors_precomp == nullptr))) return nullptr;
    const bool sr = scales && rotations;
    if (P > 0 && (sr == (cov3D_precomp != nullptr))) return nullptr;
    State* s = new State();
    s->cam = *cam; s->P = P;
    // Rasterizer::forward  rasterizer_impl.cu:222-223
    s->focal_y = cam->H / (2.0f * cam->tanfovy);
    s->focal_x = cam->W / (2.0f * cam->tanfovx);
    s->grid_x = (cam->W + BLOCK_X - 1) / BLOCK_X;
    s->grid_y = (cam->H + BLOCK_Y - 1) / BLOCK_Y;
    s->has_sh = shs != nullptr; s->has_colors = colors_precomp != nullptr; s->has_sr = sr; s->has_cov = cov3D_precomp != nullptr;
    const size_t Pz = (size_t)P;
    s->means3D.assign(means3D, means3D + 3 * Pz);
    s->opacities.assign(opacities, opacities + Pz);
    if (shs) s->shs.assign(shs, shs + Pz * cam->M * 3);
    if (colors_precomp) s->colors_in.assign(colors_precomp, colors_precomp + 3 * Pz);
    if (sr) { s->scales.assign(scales, scales + 3 * Pz); s->rotations.assign(rotations, rotations + 4 * Pz); }
    if (cov3D_precomp) s->cov3D_in.assign(cov3D_precomp, cov3D_precomp + 6 * Pz);
    s->radii.assign(Pz, 0); s->tiles_touched.assign(Pz, 0); s->point_offsets.assign(Pz, 0);
    s->depths.assign(Pz, 0.f); s->means2D.assign(2 * Pz, 0.f); s->cov3D.assign(6 * Pz, 0.f);
    s->conic_opacity.assign(4 * Pz, 0.f); s->rgb.assign(3 * Pz, 0.f); s->clamped.assign(3 * Pz, 0);
    const size_t HW = (size_t)cam->W * cam->H;
    s->ranges.assign(2 * (size_t)s->grid_x * s->grid_y, 0u);
    s->final_T.assign(HW, 0.f); s->n_contrib.assign(HW, 0u); s->out_color.assign(3 * HW, 0.f); s->out_depth.assign(HW, 0.f);
    if (P == 0) return s;   // RasterizeGaussiansCUDA short-circuits P == 0 (rasterize_points.cu:81): outputs stay zero
    preprocess(*s);
    bin_and_sort(*s);
    render(*s);
    return s;
}

long long orc_num_rendered(const void* h) { return ((const State*)h)->num_rendered; }
int orc_prefiltered_trap(const void* h) { return ((const State*)h)->prefiltered_trap; }

// Any pointer may be NULL.  Shapes: radii[P] i32, tiles_touched[P] u32, depths[P], means2D[P,2],
// cov3D[P,6], conic_opacity[P,4], rgb[P,3], clamped[P,3] u8, keys[R] u64, point_list[R] u32,
// ranges[tiles,2] u32, final_T[HW], n_contrib[HW] u32, out_color[3,H,W], out_depth[1,H,W].
void orc_export(const void* h, int* radii, uint32_t* tiles_touched, float* depths, float* means2D, float* cov3D,
                float* conic_opacity, float* rgb, uint8_t* clamped, uint64_t* keys, uint32_t* point_list,
                uint32_t* ranges, float* final_T, uint32_t* n_contrib, float* out_color, float* out_depth) {
    const State& s = *(const State*)h;
    copy_out(s.radii, radii); copy_out(s.tiles_touched, tiles_touched); copy_out(s.depths, depths);
    copy_out(s.means2D, means2D); copy_out(s.cov3D, cov3D); copy_out(s.conic_opacity, conic_opacity);
    copy_out(s.rgb, rgb); copy_out(s.clamped, clamped); copy_out(s.keys, keys); copy_out(s.point_list, point_list);
    copy_out(s.ranges, ranges); copy_out(s.final_T, final_T); copy_out(s.n_contrib, n_contrib);
    copy_out(s.out_color, out_color); copy_out(s.out_depth, out_depth);
}

// RasterizeGaussiansBackwardCUDA  DGR/rasterize_points.cu:117-196.  Outputs in the reference's shapes:
// dL_dmeans2D[P,3], dL_dconic[P,2,2], dL_dopacity[P], dL_dcolors[P,3], dL_dmeans3D[P,3], dL_dcov3D[P,6],
// dL_dscales[P,3], dL_drotations[P,4].  The *_f64 variants (may be NULL) return the FP64 sums.
int orc_backward(const void* h, const float* dL_dpix, int power, float* dL_dmeans2D, float* dL_dconic, float* dL_dopacity,
                 float* dL_dcolors, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dscales, float* dL_drotations,
                 double* dL_dmeans3D_f64, double* dL_dopacity_f64, double* dL_dscales_f64, double* dL_drotations_f64) {
    const State& s = *(const State*)h;
    Grads g;
    if (s.P == 0) return 0;
    int rc = backward(s, dL_dpix, power, g);
    if (rc) return rc;
    copy_f(g.mean2D, dL_dmeans2D); copy_f(g.conic, dL_dconic); copy_f(g.opacity, dL_dopacity); copy_f(g.colors, dL_dcolors);
    copy_f(g.mean3D, dL_dmeans3D); copy_f(g.cov3D, dL_dcov3D); copy_f(g.scale, dL_dscales); copy_f(g.rot, dL_drotations);
    copy_out(g.mean3D, dL_dmeans3D_f64); copy_out(g.opacity, dL_dopacity_f64);
    copy_out(g.scale, dL_dscales_f64); copy_out(g.rot, dL_drotations_f64);
    return 0;
}

void orc_free(void* h) { delete (State*)h; }

// checkFrustum  DGR/cr/rasterizer_impl.cu:54-66
void orc_mark_visible(int P, const float* means3D, const float* view, uint8_t* present) {
    for (int i = 0; i < P; ++i) {
        f3 p = {means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]};
        f3 pv = transformPoint4x3(p, view);
        present[i] = pv.z <= 0.001f ? 0 : 1;
    }
}

// SimpleKNN::knn  KNN/simple_knn.cu:186-222 restated by its RESULT: boxMeanDist (:148-184) prunes
// conservatively, so out[i] is the mean of the three smallest squared distances from point i to
// every other point (self excluded by index; FLT_MAX terms when P < 4), best[] ascending,
// (b0+b1+b2)/3.0f.  Distance as updateKBest<3> computes it (:132-146), fused as dot3.
void orc_knn3(int P, const float* pts, float* out) {
    for (int i = 0; i < P; ++i) {
        float best[3] = {3.402823466e+38f, 3.402823466e+38f, 3.402823466e+38f};
        const float rx = pts[3 * i], ry = pts[3 * i + 1], rz = pts[3 * i + 2];
        for (int j = 0; j < P; ++j) {
            if (j == i) continue;
            const float dx = pts[3 * j] - rx, dy = pts[3 * j + 1] - ry, dz = pts[3 * j + 2] - rz;
            float dist = dot3(dx, dx, dy, dy, dz, dz);
            for (int k = 0; k < 3; ++k)
                if (best[k] > dist) { float t = best[k]; best[k] = dist; dist = t; }
        }
        out[i] = (best[0] + best[1] + best[2]) / 3.0f;
    }
}

// count_visible_points  tester_navigator.py:147-180, counts only, for K poses (FP64 like numpy).
// w2c: K x 16 row-major; the per-point transform is written as numpy's (3,3)@(3,N) + t evaluates
// it for one point: ((r0*x + r1*y) + r2*z) + t  with separate roundings.
void orc_count_visible(int N, const float* pts, int K, const double* w2c, double fx, double fy, double cx, double cy,
                       int img_w, int img_h, int* counts) {
    for (int k = 0; k < K; ++k) {
        const double* m = w2c + 16 * (size_t)k;
        int n = 0;
        for (int i = 0; i < N; ++i) {
            const double x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
            const double xc = ((m[0] * x + m[1] * y) + m[2] * z) + m[3];
            const double yc = ((m[4] * x + m[5] * y) + m[6] * z) + m[7];
            const double zc = ((m[8] * x + m[9] * y) + m[10] * z) + m[11];
            if (!(zc > 0.0)) continue;
            const double u = (fx * xc) / zc + cx, v = (fy * yc) / zc + cy;
            if (u >= 0.0 && u < (double)img_w && v >= 0.0 && v < (double)img_h) ++n;
        }
        counts[k] = n;
    }
}

}  // extern "C"
