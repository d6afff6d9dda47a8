/*
 * gsr_oracle.c -- CPU restatement of the reference rasterizer.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference-cpu legs may load this.
 * The product path (gaussian_splatting_with_eye_tracking_b200/) never imports, links or calls it.
 *
 * What it restates (all paths under /root/reference/submodules/):
 *   orc_preprocess        diff-gaussian-rasterization/cuda_rasterizer/forward.cu:20-256, auxiliary.h:41-164
 *   orc_scan/duplicate    cuda_rasterizer/rasterizer_impl.cu:70-111,277
 *   orc_sort              cub::DeviceRadixSort::SortPairs call site rasterizer_impl.cu:303-308 (CUB is an
 *                         un-vendored toolkit dependency; its contract -- stable ascending sort on a bit range --
 *                         is restated as a stable LSD byte radix sort)
 *   orc_ranges            rasterizer_impl.cu:116-138,310-317
 *   orc_blend_forward     forward.cu:261-374
 *   orc_blend_backward    backward.cu:399-557
 *   orc_preprocess_backward  backward.cu:20-396
 *   orc_amr_*             diff-gaussian-rasterization-amr/cuda_rasterizer/rasterizer_impl.cu:181-243,590-656,
 *                         forward.cu:261-648
 *
 * FP32 evaluation order: the reference is compiled by nvcc with -fmad=true, which contracts a*b+c*d+e*f
 * into fma(e,f, fma(a,b, c*d)) (read off the sm_100a SASS of preprocessCUDA/renderCUDA).  Integer parity
 * (depth bits in the keys, radii, tile rects) needs exactly that, so every contraction is written out with
 * fmaf() and this file must be compiled with -ffp-contract=off.  MUFU.EX2 inside CUDA's expf cannot be
 * reproduced bit-for-bit on a CPU; the blend therefore matches to ~2 ulp of exp(), not bitwise.
 *
 * Pinning: checked against outputs of the reference CUDA rasterizer itself (oracle/_ref, built from the
 * sources under /root/reference) captured on a B200 and committed under tests/golden/ (see
 * oracle/make_golden.py).  The gaze-driven level cap (orc_fovea_levels with gaze) is an extension without
 * reference code: "parity unpinned" for that term only.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                               -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

ORC_API int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* a*b + c*d + e*f as nvcc contracts it */
static inline float dot3c(float a, float b, float c, float d, float e, float f) {
    return fmaf(e, f, fmaf(a, b, c * d));
}

/* glm 3x3, column-major: m[c][r] */
typedef struct { float m[3][3]; } mat3;

/* glm operator* (third_party/glm/glm/detail/type_mat3x3.inl:486-518) with nvcc's contraction */
static mat3 mat3_mul(const mat3* A, const mat3* B) {
    mat3 R;
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r)
            R.m[c][r] = dot3c(A->m[0][r], B->m[c][0], A->m[1][r], B->m[c][1], A->m[2][r], B->m[c][2]);
    return R;
}
static mat3 mat3_transpose(const mat3* A) {
    mat3 R;
    for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) R.m[c][r] = A->m[r][c];
    return R;
}

/* auxiliary.h:58-77 */
static inline float xform_row(const float* m, int row, float x, float y, float z) {
    return dot3c(m[row] , x, m[4 + row], y, m[8 + row], z) + m[12 + row];
}

/* auxiliary.h:41-44: double arithmetic, (v+1.0)*S-1.0 is one DFMA in the compiled reference */
static inline float ndc2Pix(float v, int S) {
    return (float)(fma((double)v + 1.0, (double)S, -1.0) * 0.5);
}

/* auxiliary.h:46-56 */
static void getRect(float px, float py, int max_radius, int block, int gx, int gy, uint32_t rmin[2], uint32_t rmax[2]) {
    const float r = (float)max_radius, b = (float)block, inv = 1.0f / (float)block;
    int ax = (int)((px - r) * inv), ay = (int)((py - r) * inv);
    int bx = (int)((((px + r) + b) - 1.0f) * inv), by = (int)((((py + r) + b) - 1.0f) * inv);
    if (ax < 0) ax = 0; if (ay < 0) ay = 0; if (bx < 0) bx = 0; if (by < 0) by = 0;
    rmin[0] = (uint32_t)ax < (uint32_t)gx ? (uint32_t)ax : (uint32_t)gx;
    rmin[1] = (uint32_t)ay < (uint32_t)gy ? (uint32_t)ay : (uint32_t)gy;
    rmax[0] = (uint32_t)bx < (uint32_t)gx ? (uint32_t)bx : (uint32_t)gx;
    rmax[1] = (uint32_t)by < (uint32_t)gy ? (uint32_t)by : (uint32_t)gy;
}

/* forward.cu:118-152 */
static void computeCov3D(const float* scale, float mod, const float* rot, float* cov3D) {
    const float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
    const float xz = x * z, rx = r * x, rz = r * z, yy = y * y, zz = z * z;
    mat3 R;
    R.m[0][0] = 1.f - ((yy + zz) + (yy + zz));
    R.m[0][1] = fmaf(x, y, -rz) + fmaf(x, y, -rz);
    R.m[0][2] = fmaf(r, y, xz) + fmaf(r, y, xz);
    R.m[1][0] = fmaf(x, y, rz) + fmaf(x, y, rz);
    R.m[1][1] = 1.f - (fmaf(x, x, zz) + fmaf(x, x, zz));
    R.m[1][2] = fmaf(y, z, -rx) + fmaf(y, z, -rx);
    R.m[2][0] = fmaf(-r, y, xz) + fmaf(-r, y, xz);
    R.m[2][1] = fmaf(y, z, rx) + fmaf(y, z, rx);
    R.m[2][2] = 1.f - (fmaf(x, x, yy) + fmaf(x, x, yy));
    mat3 S; memset(&S, 0, sizeof(S));
    S.m[0][0] = mod * scale[0]; S.m[1][1] = mod * scale[1]; S.m[2][2] = mod * scale[2];
    mat3 M = mat3_mul(&S, &R);
    mat3 Mt = mat3_transpose(&M);
    mat3 Sigma = mat3_mul(&Mt, &M);
    cov3D[0] = Sigma.m[0][0]; cov3D[1] = Sigma.m[0][1]; cov3D[2] = Sigma.m[0][2];
    cov3D[3] = Sigma.m[1][1]; cov3D[4] = Sigma.m[1][2]; cov3D[5] = Sigma.m[2][2];
}

/* forward.cu:74-113 */
static void computeCov2D(const float* mean, float focal_x, float focal_y, float tan_fovx, float tan_fovy,
                         const float* cov3D, const float* V, float cov[3]) {
    float tx = xform_row(V, 0, mean[0], mean[1], mean[2]);
    float ty = xform_row(V, 1, mean[0], mean[1], mean[2]);
    const float tz = xform_row(V, 2, mean[0], mean[1], mean[2]);
    const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
    const float txtz = tx / tz, tytz = ty / tz;
    const float cx = fminf(limx, fmaxf(-limx, txtz)), cy = fminf(limy, fmaxf(-limy, tytz));
    mat3 J; memset(&J, 0, sizeof(J));
    J.m[0][0] = focal_x / tz; J.m[0][2] = ((tz * -cx) * focal_x) / (tz * tz);
    J.m[1][1] = focal_y / tz; J.m[1][2] = ((tz * -cy) * focal_y) / (tz * tz);
    mat3 W = {{{V[0], V[4], V[8]}, {V[1], V[5], V[9]}, {V[2], V[6], V[10]}}};
    mat3 T = mat3_mul(&W, &J);
    mat3 Vrk = {{{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}}};
    mat3 Tt = mat3_transpose(&T), Vt = mat3_transpose(&Vrk);
    mat3 X = mat3_mul(&Tt, &Vt);
    mat3 C = mat3_mul(&X, &T);
    cov[0] = C.m[0][0] + 0.3f; cov[1] = C.m[0][1]; cov[2] = C.m[1][1] + 0.3f;
}

/* forward.cu:20-71 */
static void computeColorFromSH(int deg, const float* pos, const float* campos, const float* sh, float rgb[3], uint8_t clamped[3]) {
    const float dx = pos[0] - campos[0], dy = pos[1] - campos[1], dz = pos[2] - campos[2];
    const float len = sqrtf(dot3c(dx, dx, dy, dy, dz, dz));
    const float x = dx / len, y = dy / len, z = dz / len;
    float res[3];
    for (int c = 0; c < 3; ++c) res[c] = SH_C0 * sh[c];
    if (deg > 0) {
        const float a = y * SH_C1, b = z * SH_C1, d = x * SH_C1;
        for (int c = 0; c < 3; ++c) {
            res[c] = fmaf(-a, sh[3 + c], res[c]);
            res[c] = fmaf(b, sh[6 + c], res[c]);
            res[c] = fmaf(-d, sh[9 + c], res[c]);
        }
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = y * x, yz = z * y, xz = z * x;
            const float zz2 = zz + zz, xx_yy = xx - yy;
            const float t[5] = {xy * SH_C2[0], yz * SH_C2[1], ((zz2 - xx) - yy) * SH_C2[2], xz * SH_C2[3], xx_yy * SH_C2[4]};
            for (int c = 0; c < 3; ++c)
                for (int k = 0; k < 5; ++k) res[c] = fmaf(t[k], sh[3 * (4 + k) + c], res[c]);
            if (deg > 2) {
                const float q = fmaf(zz, 4.f, -xx) - yy;
                const float u[7] = {
                    (y * SH_C3[0]) * fmaf(xx, 3.f, -yy),
                    (xy * SH_C3[1]) * z,
                    (y * SH_C3[2]) * q,
                    (z * SH_C3[3]) * fmaf(yy, -3.f, fmaf(xx, -3.f, zz2)),
                    q * (x * SH_C3[4]),
                    xx_yy * (z * SH_C3[5]),
                    (x * SH_C3[6]) * fmaf(yy, -3.f, xx)};
                for (int c = 0; c < 3; ++c)
                    for (int k = 0; k < 7; ++k) res[c] = fmaf(u[k], sh[3 * (9 + k) + c], res[c]);
            }
        }
    }
    for (int c = 0; c < 3; ++c) {
        res[c] += 0.5f;
        clamped[c] = res[c] < 0.f;
        rgb[c] = res[c] > 0.f ? res[c] : 0.f;
    }
}

/* forward.cu:155-256 (preprocessCUDA) + in_frustum auxiliary.h:139-164.  block = 16 (R/) or 32 (A/). */
ORC_API void orc_preprocess(int P, int D, int M, const float* means3D, const float* scales, float scale_modifier,
                            const float* rotations, const float* opacities, const float* shs, const float* cov3D_precomp,
                            const float* colors_precomp, const float* viewmatrix, const float* projmatrix,
                            const float* cam_pos, int W, int H, float tan_fovx, float tan_fovy, int block,
                            int* radii, float* means2D, float* depths, float* cov3Ds, float* rgb, float* conic_opacity,
                            uint32_t* tiles_touched, uint8_t* clamped) {
    const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);   /* rasterizer_impl.cu:222-223 */
    const int gx = (W + block - 1) / block, gy = (H + block - 1) / block;
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        radii[idx] = 0; tiles_touched[idx] = 0;
        const float* p = means3D + 3 * (size_t)idx;
        const float depth = xform_row(viewmatrix, 2, p[0], p[1], p[2]);
        if (depth <= 0.2f) continue;
        const float hx = xform_row(projmatrix, 0, p[0], p[1], p[2]);
        const float hy = xform_row(projmatrix, 1, p[0], p[1], p[2]);
        const float hw = xform_row(projmatrix, 3, p[0], p[1], p[2]);
        const float p_w = 1.0f / (hw + 0.0000001f);
        const float projx = hx * p_w, projy = hy * p_w;
        const float* cov3D;
        if (cov3D_precomp) cov3D = cov3D_precomp + 6 * (size_t)idx;
        else { computeCov3D(scales + 3 * (size_t)idx, scale_modifier, rotations + 4 * (size_t)idx, cov3Ds + 6 * (size_t)idx); cov3D = cov3Ds + 6 * (size_t)idx; }
        float cov[3];
        computeCov2D(p, focal_x, focal_y, tan_fovx, tan_fovy, cov3D, viewmatrix, cov);
        const float det = fmaf(cov[0], cov[2], -(cov[1] * cov[1]));
        if (det == 0.0f) continue;
        const float det_inv = 1.f / det;
        const float conic[3] = {cov[2] * det_inv, cov[1] * -det_inv, cov[0] * det_inv};
        const float mid = (cov[0] + cov[2]) * 0.5f;
        const float disc = sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
        const float lambda = fmaxf(mid + disc, mid - disc);
        const int my_radius = (int)ceilf(sqrtf(lambda) * 3.f);
        const float pix[2] = {ndc2Pix(projx, W), ndc2Pix(projy, H)};
        uint32_t rmin[2], rmax[2];
        getRect(pix[0], pix[1], my_radius, block, gx, gy, rmin, rmax);
        if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) continue;
        if (!colors_precomp)
            computeColorFromSH(D, p, cam_pos, shs + (size_t)idx * M * 3, rgb + 3 * (size_t)idx, clamped + 3 * (size_t)idx);
        depths[idx] = depth;
        radii[idx] = my_radius;
        means2D[2 * (size_t)idx] = pix[0]; means2D[2 * (size_t)idx + 1] = pix[1];
        float* co = conic_opacity + 4 * (size_t)idx;
        co[0] = conic[0]; co[1] = conic[1]; co[2] = conic[2]; co[3] = opacities[idx];
        tiles_touched[idx] = (rmax[1] - rmin[1]) * (rmax[0] - rmin[0]);
    }
}

/* rasterizer_impl.cu:54-66 (checkFrustum) */
ORC_API void orc_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present) {
    for (int i = 0; i < P; ++i)
        present[i] = xform_row(viewmatrix, 2, means3D[3 * (size_t)i], means3D[3 * (size_t)i + 1], means3D[3 * (size_t)i + 2]) > 0.2f;
}

/* ------------------------------------------------------------------ binning */
/* cub::DeviceScan::InclusiveSum call site rasterizer_impl.cu:277; returns num_rendered (:281) */
ORC_API long long orc_scan(int P, const uint32_t* tiles_touched, uint32_t* point_offsets) {
    uint32_t acc = 0;
    for (int i = 0; i < P; ++i) { acc += tiles_touched[i]; point_offsets[i] = acc; }
    return (long long)acc;
}

/* rasterizer_impl.cu:70-111 (duplicateWithKeys) */
ORC_API void orc_duplicate(int P, const float* means2D, const float* depths, const uint32_t* offsets, const int* radii,
                           int block, int gx, int gy, uint64_t* keys, uint32_t* vals) {
#pragma omp parallel for schedule(dynamic, 1024)
    for (int idx = 0; idx < P; ++idx) {
        if (radii[idx] <= 0) continue;
        uint32_t off = idx == 0 ? 0 : offsets[idx - 1];
        uint32_t rmin[2], rmax[2];
        getRect(means2D[2 * (size_t)idx], means2D[2 * (size_t)idx + 1], radii[idx], block, gx, gy, rmin, rmax);
        uint32_t dbits; memcpy(&dbits, depths + idx, 4);
        for (uint32_t y = rmin[1]; y < rmax[1]; ++y)
            for (uint32_t x = rmin[0]; x < rmax[0]; ++x) {
                uint64_t key = (uint64_t)(y * (uint32_t)gx + x);
                key <<= 32; key |= dbits;
                keys[off] = key; vals[off] = (uint32_t)idx; ++off;
            }
    }
}

/* rasterizer_impl.cu:35-50 */
ORC_API uint32_t orc_higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) { step /= 2; if (n >> msb) msb += step; else msb -= step; }
    if (n >> msb) msb++;
    return msb;
}

/* stable ascending sort of pairs on key bits [0,end_bit): the SortPairs contract (rasterizer_impl.cu:303-308) */
ORC_API void orc_sort(long long n, const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int end_bit) {
    if (n <= 0) return;
    uint64_t* ka = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)n); uint64_t* kb = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)n);
    uint32_t* va = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n); uint32_t* vb = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
    memcpy(ka, keys_in, sizeof(uint64_t) * (size_t)n); memcpy(va, vals_in, sizeof(uint32_t) * (size_t)n);
    for (int shift = 0; shift < end_bit; shift += 8) {
        int bits = end_bit - shift < 8 ? end_bit - shift : 8;
        uint32_t mask = (1u << bits) - 1u;
        size_t cnt[257]; memset(cnt, 0, sizeof(cnt));
        for (long long i = 0; i < n; ++i) cnt[((ka[i] >> shift) & mask) + 1]++;
        for (int b = 0; b < 256; ++b) cnt[b + 1] += cnt[b];
        for (long long i = 0; i < n; ++i) { size_t d = cnt[(ka[i] >> shift) & mask]++; kb[d] = ka[i]; vb[d] = va[i]; }
        uint64_t* tk = ka; ka = kb; kb = tk; uint32_t* tv = va; va = vb; vb = tv;
    }
    memcpy(keys_out, ka, sizeof(uint64_t) * (size_t)n); memcpy(vals_out, va, sizeof(uint32_t) * (size_t)n);
    free(ka); free(kb); free(va); free(vb);
}

/* rasterizer_impl.cu:116-138 + memset :310.  ranges = [num_tiles][2] */
ORC_API void orc_ranges(long long L, const uint64_t* keys, uint32_t* ranges, int num_tiles) {
    memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)num_tiles);
    for (long long idx = 0; idx < L; ++idx) {
        uint32_t cur = (uint32_t)(keys[idx] >> 32);
        if (idx == 0) ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(keys[idx - 1] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)idx; ranges[2 * cur] = (uint32_t)idx; }
        }
        if (idx == L - 1) ranges[2 * cur + 1] = (uint32_t)L;
    }
}

/* ------------------------------------------------------------------ blend forward */
static inline float blend_power(float cx, float cy, float cz, float dx, float dy) {
    const float s = fmaf(dx, dx * cx, dy * (dy * cz));
    const float m = dy * (dx * cy);
    return fmaf(s, -0.5f, -m);
}

/* one pixel of renderCUDA (forward.cu:306-372); the CTA-level early exit does not change per-pixel results */
static void blend_pixel(uint32_t px, uint32_t py, uint32_t r0, uint32_t r1, const uint32_t* point_list, const float* means2D,
                        const float* colors, const float* conic_opacity, const float* bg, float* T_out, uint32_t* ncontrib_out, float C_out[3]) {
    float T = 1.0f, C[3] = {0, 0, 0};
    uint32_t contributor = 0, last_contributor = 0;
    const float pxf = (float)px, pyf = (float)py;
    for (uint32_t i = r0; i < r1; ++i) {
        contributor++;
        const uint32_t id = point_list[i];
        const float dx = means2D[2 * (size_t)id] - pxf, dy = means2D[2 * (size_t)id + 1] - pyf;
        const float* co = conic_opacity + 4 * (size_t)id;
        const float power = blend_power(co[0], co[1], co[2], dx, dy);
        if (power > 0.0f) continue;
        const float alpha = fminf(0.99f, co[3] * expf(power));
        if (alpha < 1.0f / 255.0f) continue;
        const float test_T = T * (1 - alpha);
        if (test_T < 0.0001f) break;
        for (int ch = 0; ch < 3; ++ch) C[ch] = fmaf(T, alpha * colors[3 * (size_t)id + ch], C[ch]);
        T = test_T;
        last_contributor = contributor;
    }
    *T_out = T; *ncontrib_out = last_contributor;
    for (int ch = 0; ch < 3; ++ch) C_out[ch] = fmaf(T, bg[ch], C[ch]);
}

ORC_API void orc_blend_forward(int W, int H, int block, const uint32_t* ranges, const uint32_t* point_list, const float* means2D,
                               const float* colors, const float* conic_opacity, const float* bg, float* final_T,
                               uint32_t* n_contrib, float* out_color) {
    const int gx = (W + block - 1) / block;
    const size_t HW = (size_t)H * W;
#pragma omp parallel for schedule(dynamic, 4)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const int tile = (y / block) * gx + (x / block);
            float C[3];
            const size_t pix = (size_t)W * y + x;
            blend_pixel((uint32_t)x, (uint32_t)y, ranges[2 * tile], ranges[2 * tile + 1], point_list, means2D, colors, conic_opacity, bg,
                        final_T + pix, n_contrib + pix, C);
            for (int ch = 0; ch < 3; ++ch) out_color[ch * HW + pix] = C[ch];
        }
}

/* ------------------------------------------------------------------ blend backward (backward.cu:399-557)
 * Gradients are accumulated in double and rounded once (the reference uses non-deterministic float atomics).
 * dL_dmean2D [P,3] (z unused), dL_dconic [P,4] (x,y,-,w), dL_dopacity [P], dL_dcolors [P,3]: all += into zeroed arrays. */
ORC_API void orc_blend_backward(int P, int W, int H, int block, const uint32_t* ranges, const uint32_t* point_list, const float* bg,
                                const float* means2D, const float* conic_opacity, const float* colors, const float* final_Ts,
                                const uint32_t* n_contrib, const float* dL_dpixels, float* dL_dmean2D, float* dL_dconic,
                                float* dL_dopacity, float* dL_dcolors) {
    const int gx = (W + block - 1) / block;
    const size_t HW = (size_t)H * W;
    double* acc = (double*)calloc((size_t)P * 9, sizeof(double));
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const int tile = (y / block) * gx + (x / block);
            const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
            const size_t pix = (size_t)W * y + x;
            const float T_final = final_Ts[pix];
            float T = T_final;
            const uint32_t last_contributor = n_contrib[pix];
            float accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0}, last_alpha = 0, dL_dpixel[3];
            for (int ch = 0; ch < 3; ++ch) dL_dpixel[ch] = dL_dpixels[ch * HW + pix];
            const float pxf = (float)x, pyf = (float)y;
            uint32_t contributor = r1 - r0;
            for (uint32_t i = r1; i-- > r0;) {
                contributor--;
                if (contributor >= last_contributor) continue;
                const uint32_t id = point_list[i];
                const float dx = means2D[2 * (size_t)id] - pxf, dy = means2D[2 * (size_t)id + 1] - pyf;
                const float* co = conic_opacity + 4 * (size_t)id;
                const float power = blend_power(co[0], co[1], co[2], dx, dy);
                if (power > 0.0f) continue;
                const float G = expf(power);
                const float alpha = fminf(0.99f, co[3] * G);
                if (alpha < 1.0f / 255.0f) continue;
                T = T / (1.f - alpha);
                const float dchannel_dcolor = alpha * T;
                float dL_dalpha = 0.0f;
                double* a = acc + 9 * (size_t)id;
                for (int ch = 0; ch < 3; ++ch) {
                    const float c = colors[3 * (size_t)id + ch];
                    accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                    last_color[ch] = c;
                    dL_dalpha += (c - accum_rec[ch]) * dL_dpixel[ch];
                    a[6 + ch] += dchannel_dcolor * dL_dpixel[ch];
                }
                dL_dalpha *= T;
                last_alpha = alpha;
                float bg_dot = 0;
                for (int ch = 0; ch < 3; ++ch) bg_dot += bg[ch] * dL_dpixel[ch];
                dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                const float dL_dG = co[3] * dL_dalpha;
                const float gdx = G * dx, gdy = G * dy;
                const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                const float dG_ddely = -gdy * co[2] - gdx * co[1];
                a[0] += dL_dG * dG_ddelx * ddelx_dx;
                a[1] += dL_dG * dG_ddely * ddely_dy;
                a[2] += -0.5f * gdx * dx * dL_dG;
                a[3] += -0.5f * gdx * dy * dL_dG;
                a[4] += -0.5f * gdy * dy * dL_dG;
                a[5] += G * dL_dalpha;
            }
        }
    for (int i = 0; i < P; ++i) {
        const double* a = acc + 9 * (size_t)i;
        dL_dmean2D[3 * (size_t)i] += (float)a[0]; dL_dmean2D[3 * (size_t)i + 1] += (float)a[1];
        dL_dconic[4 * (size_t)i] += (float)a[2]; dL_dconic[4 * (size_t)i + 1] += (float)a[3]; dL_dconic[4 * (size_t)i + 3] += (float)a[4];
        dL_dopacity[i] += (float)a[5];
        for (int ch = 0; ch < 3; ++ch) dL_dcolors[3 * (size_t)i + ch] += (float)a[6 + ch];
    }
    free(acc);
}

/* ------------------------------------------------------------------ preprocess backward
 * computeCov2DCUDA (backward.cu:144-274), preprocessCUDA bwd (:346-396), computeColorFromSH bwd (:20-139),
 * computeCov3D bwd (:278-341).  Plain FP32 in the reference's expression order (no pinned contraction: the
 * gradient tolerance is 1e-3 relative).  Outputs must be zero-initialised by the caller (reference: torch::zeros). */
static void sh_backward(int idx, int deg, int M, const float* means, const float* campos, const float* shs, const uint8_t* clamped,
                        const float* dL_dcolor, float* dL_dmeans, float* dL_dshs) {
    const float* pos = means + 3 * (size_t)idx;
    const float ox = pos[0] - campos[0], oy = pos[1] - campos[1], oz = pos[2] - campos[2];
    const float len = sqrtf(ox * ox + oy * oy + oz * oz);
    const float x = ox / len, y = oy / len, z = oz / len;
    const float* sh = shs + (size_t)idx * M * 3;
    float g[3];
    for (int c = 0; c < 3; ++c) g[c] = dL_dcolor[3 * (size_t)idx + c] * (clamped[3 * (size_t)idx + c] ? 0.f : 1.f);
    float dRdx[3] = {0, 0, 0}, dRdy[3] = {0, 0, 0}, dRdz[3] = {0, 0, 0};
    float w[16]; for (int k = 0; k < 16; ++k) w[k] = 0.f;
    int ncoef = 1;
    w[0] = SH_C0;
#define S(k, c) sh[3 * (k) + (c)]
    if (deg > 0) {
        ncoef = 4;
        w[1] = -SH_C1 * y; w[2] = SH_C1 * z; w[3] = -SH_C1 * x;
        for (int c = 0; c < 3; ++c) { dRdx[c] = -SH_C1 * S(3, c); dRdy[c] = -SH_C1 * S(1, c); dRdz[c] = SH_C1 * S(2, c); }
        if (deg > 1) {
            ncoef = 9;
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            w[4] = SH_C2[0] * xy; w[5] = SH_C2[1] * yz; w[6] = SH_C2[2] * (2.f * zz - xx - yy); w[7] = SH_C2[3] * xz; w[8] = SH_C2[4] * (xx - yy);
            for (int c = 0; c < 3; ++c) {
                dRdx[c] += SH_C2[0] * y * S(4, c) + SH_C2[2] * 2.f * -x * S(6, c) + SH_C2[3] * z * S(7, c) + SH_C2[4] * 2.f * x * S(8, c);
                dRdy[c] += SH_C2[0] * x * S(4, c) + SH_C2[1] * z * S(5, c) + SH_C2[2] * 2.f * -y * S(6, c) + SH_C2[4] * 2.f * -y * S(8, c);
                dRdz[c] += SH_C2[1] * y * S(5, c) + SH_C2[2] * 2.f * 2.f * z * S(6, c) + SH_C2[3] * x * S(7, c);
            }
            if (deg > 2) {
                ncoef = 16;
                w[9] = SH_C3[0] * y * (3.f * xx - yy); w[10] = SH_C3[1] * xy * z; w[11] = SH_C3[2] * y * (4.f * zz - xx - yy);
                w[12] = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy); w[13] = SH_C3[4] * x * (4.f * zz - xx - yy);
                w[14] = SH_C3[5] * z * (xx - yy); w[15] = SH_C3[6] * x * (xx - 3.f * yy);
                for (int c = 0; c < 3; ++c) {
                    dRdx[c] += SH_C3[0] * S(9, c) * 3.f * 2.f * xy + SH_C3[1] * S(10, c) * yz + SH_C3[2] * S(11, c) * -2.f * xy +
                               SH_C3[3] * S(12, c) * -3.f * 2.f * xz + SH_C3[4] * S(13, c) * (-3.f * xx + 4.f * zz - yy) +
                               SH_C3[5] * S(14, c) * 2.f * xz + SH_C3[6] * S(15, c) * 3.f * (xx - yy);
                    dRdy[c] += SH_C3[0] * S(9, c) * 3.f * (xx - yy) + SH_C3[1] * S(10, c) * xz + SH_C3[2] * S(11, c) * (-3.f * yy + 4.f * zz - xx) +
                               SH_C3[3] * S(12, c) * -3.f * 2.f * yz + SH_C3[4] * S(13, c) * -2.f * xy + SH_C3[5] * S(14, c) * -2.f * yz +
                               SH_C3[6] * S(15, c) * -3.f * 2.f * xy;
                    dRdz[c] += SH_C3[1] * S(10, c) * xy + SH_C3[2] * S(11, c) * 4.f * 2.f * yz + SH_C3[3] * S(12, c) * 3.f * (2.f * zz - xx - yy) +
                               SH_C3[4] * S(13, c) * 4.f * 2.f * xz + SH_C3[5] * S(14, c) * (xx - yy);
                }
            }
        }
    }
#undef S
    float* out = dL_dshs + (size_t)idx * M * 3;
    for (int k = 0; k < ncoef; ++k) for (int c = 0; c < 3; ++c) out[3 * k + c] = w[k] * g[c];
    const float ddx = dRdx[0] * g[0] + dRdx[1] * g[1] + dRdx[2] * g[2];
    const float ddy = dRdy[0] * g[0] + dRdy[1] * g[1] + dRdy[2] * g[2];
    const float ddz = dRdz[0] * g[0] + dRdz[1] * g[1] + dRdz[2] * g[2];
    /* dnormvdv (auxiliary.h:107-118) */
    const float sum2 = ox * ox + oy * oy + oz * oz;
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    dL_dmeans[3 * (size_t)idx] += ((+sum2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * invsum32;
    dL_dmeans[3 * (size_t)idx + 1] += (-ox * oy * ddx + (sum2 - oy * oy) * ddy - oz * oy * ddz) * invsum32;
    dL_dmeans[3 * (size_t)idx + 2] += (-ox * oz * ddx - oy * oz * ddy + (sum2 - oz * oz) * ddz) * invsum32;
}

static void cov3d_backward(int idx, const float* scale, float mod, const float* rot, const float* dL_dcov3Ds, float* dL_dscales, float* dL_drots) {
    const float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
    /* R[c][r] glm */
    const float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                           {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                           {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    const float s[3] = {mod * scale[0], mod * scale[1], mod * scale[2]};
    float Mm[3][3];
    for (int c = 0; c < 3; ++c) for (int rr = 0; rr < 3; ++rr) Mm[c][rr] = s[rr] * R[c][rr];
    const float* d = dL_dcov3Ds + 6 * (size_t)idx;
    const float Sg[3][3] = {{d[0], 0.5f * d[1], 0.5f * d[2]}, {0.5f * d[1], d[3], 0.5f * d[4]}, {0.5f * d[2], 0.5f * d[4], d[5]}};
    float dM[3][3];   /* dL_dM = 2 * M * dL_dSigma : dM[c][r] = sum_k (2 M[k][r]) Sg[c][k] */
    for (int c = 0; c < 3; ++c) for (int rr = 0; rr < 3; ++rr)
        dM[c][rr] = 2.f * Mm[0][rr] * Sg[c][0] + 2.f * Mm[1][rr] * Sg[c][1] + 2.f * Mm[2][rr] * Sg[c][2];
    float dMt[3][3], Rt[3][3];
    for (int c = 0; c < 3; ++c) for (int rr = 0; rr < 3; ++rr) { dMt[c][rr] = dM[rr][c]; Rt[c][rr] = R[rr][c]; }
    for (int i = 0; i < 3; ++i) dL_dscales[3 * (size_t)idx + i] = Rt[i][0] * dMt[i][0] + Rt[i][1] * dMt[i][1] + Rt[i][2] * dMt[i][2];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) dMt[i][j] *= s[i];
    float* q = dL_drots + 4 * (size_t)idx;
    q[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
    q[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
    q[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
    q[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
}

ORC_API void orc_preprocess_backward(int P, int D, int M, const float* means3D, const int* radii, const float* shs, const uint8_t* clamped,
                                     const float* scales, const float* rotations, float scale_modifier, const float* cov3Ds,
                                     const float* view, const float* proj, int W, int H, float tan_fovx, float tan_fovy,
                                     const float* campos, const float* dL_dmean2D /*[P,3]*/, const float* dL_dconics /*[P,4]*/,
                                     float* dL_dmeans /*[P,3]*/, const float* dL_dcolor /*[P,3]*/, float* dL_dcov /*[P,6]*/,
                                     float* dL_dsh, float* dL_dscale, float* dL_drot) {
    const float h_y = H / (2.0f * tan_fovy), h_x = W / (2.0f * tan_fovx);
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        if (!(radii[idx] > 0)) continue;
        /* ---- computeCov2DCUDA */
        const float* cov3D = cov3Ds + 6 * (size_t)idx;
        const float* mean = means3D + 3 * (size_t)idx;
        const float dcx = dL_dconics[4 * (size_t)idx], dcy = dL_dconics[4 * (size_t)idx + 1], dcz = dL_dconics[4 * (size_t)idx + 3];
        float tx = view[0] * mean[0] + view[4] * mean[1] + view[8] * mean[2] + view[12];
        float ty = view[1] * mean[0] + view[5] * mean[1] + view[9] * mean[2] + view[13];
        const float tz = view[2] * mean[0] + view[6] * mean[1] + view[10] * mean[2] + view[14];
        const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
        const float txtz = tx / tz, tytz = ty / tz;
        tx = fminf(limx, fmaxf(-limx, txtz)) * tz; ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        /* glm [c][r] */
        const float J[3][3] = {{h_x / tz, 0.f, -(h_x * tx) / (tz * tz)}, {0.f, h_y / tz, -(h_y * ty) / (tz * tz)}, {0, 0, 0}};
        const float Wm[3][3] = {{view[0], view[4], view[8]}, {view[1], view[5], view[9]}, {view[2], view[6], view[10]}};
        const float Vrk[3][3] = {{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}};
        float T[3][3], X[3][3], c2[3][3];
        for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) T[c][r] = Wm[0][r] * J[c][0] + Wm[1][r] * J[c][1] + Wm[2][r] * J[c][2];
        for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) X[c][r] = T[r][0] * Vrk[0][c] + T[r][1] * Vrk[1][c] + T[r][2] * Vrk[2][c];
        for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) c2[c][r] = X[0][r] * T[c][0] + X[1][r] * T[c][1] + X[2][r] * T[c][2];
        const float a = c2[0][0] + 0.3f, b = c2[0][1], c = c2[1][1] + 0.3f;
        const float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        float* dcov = dL_dcov + 6 * (size_t)idx;
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
            dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
            dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
            dcov[0] = (T[0][0] * T[0][0] * dL_da + T[0][0] * T[1][0] * dL_db + T[1][0] * T[1][0] * dL_dc);
            dcov[3] = (T[0][1] * T[0][1] * dL_da + T[0][1] * T[1][1] * dL_db + T[1][1] * T[1][1] * dL_dc);
            dcov[5] = (T[0][2] * T[0][2] * dL_da + T[0][2] * T[1][2] * dL_db + T[1][2] * T[1][2] * dL_dc);
            dcov[1] = 2 * T[0][0] * T[0][1] * dL_da + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][1] * dL_dc;
            dcov[2] = 2 * T[0][0] * T[0][2] * dL_da + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][2] * dL_dc;
            dcov[4] = 2 * T[0][2] * T[0][1] * dL_da + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_db + 2 * T[1][1] * T[1][2] * dL_dc;
        } else { for (int i = 0; i < 6; ++i) dcov[i] = 0; }
        float dT0[3], dT1[3];
        for (int j = 0; j < 3; ++j) {
            const float u0 = T[0][0] * Vrk[j][0] + T[0][1] * Vrk[j][1] + T[0][2] * Vrk[j][2];
            const float u1 = T[1][0] * Vrk[j][0] + T[1][1] * Vrk[j][1] + T[1][2] * Vrk[j][2];
            dT0[j] = 2 * u0 * dL_da + u1 * dL_db;
            dT1[j] = 2 * u1 * dL_dc + u0 * dL_db;
        }
        const float dJ00 = Wm[0][0] * dT0[0] + Wm[0][1] * dT0[1] + Wm[0][2] * dT0[2];
        const float dJ02 = Wm[2][0] * dT0[0] + Wm[2][1] * dT0[1] + Wm[2][2] * dT0[2];
        const float dJ11 = Wm[1][0] * dT1[0] + Wm[1][1] * dT1[1] + Wm[1][2] * dT1[2];
        const float dJ12 = Wm[2][0] * dT1[0] + Wm[2][1] * dT1[1] + Wm[2][2] * dT1[2];
        const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
        const float dtx = x_grad_mul * -h_x * itz2 * dJ02;
        const float dty = y_grad_mul * -h_y * itz2 * dJ12;
        const float dtz = -h_x * itz2 * dJ00 - h_y * itz2 * dJ11 + (2 * h_x * tx) * itz3 * dJ02 + (2 * h_y * ty) * itz3 * dJ12;
        float* dm = dL_dmeans + 3 * (size_t)idx;
        dm[0] = view[0] * dtx + view[1] * dty + view[2] * dtz;      /* transformVec4x3Transpose; ASSIGNED (backward.cu:273) */
        dm[1] = view[4] * dtx + view[5] * dty + view[6] * dtz;
        dm[2] = view[8] * dtx + view[9] * dty + view[10] * dtz;
        /* ---- preprocessCUDA backward */
        const float m_w = 1.0f / ((proj[3] * mean[0] + proj[7] * mean[1] + proj[11] * mean[2] + proj[15]) + 0.0000001f);
        const float mul1 = (proj[0] * mean[0] + proj[4] * mean[1] + proj[8] * mean[2] + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * mean[0] + proj[5] * mean[1] + proj[9] * mean[2] + proj[13]) * m_w * m_w;
        const float gx2 = dL_dmean2D[3 * (size_t)idx], gy2 = dL_dmean2D[3 * (size_t)idx + 1];
        dm[0] += (proj[0] * m_w - proj[3] * mul1) * gx2 + (proj[1] * m_w - proj[3] * mul2) * gy2;
        dm[1] += (proj[4] * m_w - proj[7] * mul1) * gx2 + (proj[5] * m_w - proj[7] * mul2) * gy2;
        dm[2] += (proj[8] * m_w - proj[11] * mul1) * gx2 + (proj[9] * m_w - proj[11] * mul2) * gy2;
        if (shs) sh_backward(idx, D, M, means3D, campos, shs, clamped, dL_dcolor, dL_dmeans, dL_dsh);
        if (scales) cov3d_backward(idx, scales + 3 * (size_t)idx, scale_modifier, rotations + 4 * (size_t)idx, dL_dcov, dL_dscale, dL_drot);
    }
}

/* ------------------------------------------------------------------ AMR (diff-gaussian-rasterization-amr) */
static int cmp_u32(const void* a, const void* b) { uint32_t x = *(const uint32_t*)a, y = *(const uint32_t*)b; return (x > y) - (x < y); }

/* calculateIntersections (:181-188) + SortKeys + percentile pick (:609-637) + setAMRLevelsKernel (:190-205) */
ORC_API void orc_amr_levels(int num_tiles, const uint32_t* ranges, uint32_t* n_intersections, uint32_t* levels, uint32_t* thresholds) {
    uint32_t* sorted = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)num_tiles);
    for (int t = 0; t < num_tiles; ++t) { n_intersections[t] = ranges[2 * t + 1] - ranges[2 * t]; sorted[t] = n_intersections[t]; }
    qsort(sorted, (size_t)num_tiles, sizeof(uint32_t), cmp_u32);
    const float percentiles[3] = {0.25f, 0.5f, 0.9f};
    for (int i = 0; i < 3; ++i) {
        int k = (int)(percentiles[i] * num_tiles);      /* float multiply, as in the reference (:630) */
        if (k > num_tiles - 1) k = num_tiles - 1;
        thresholds[i] = sorted[k];
    }
    for (int t = 0; t < num_tiles; ++t) {
        const uint32_t v = n_intersections[t];
        levels[t] = v <= thresholds[0] ? 1 : v <= thresholds[1] ? 2 : v <= thresholds[2] ? 3 : 4;
    }
    free(sorted);
}

/* setFoveaAMRLevelsKernel (:208-243).  gaze = {cx, cy, r1..r4} or NULL.  The gaze cap is an EXTENSION:
 * lvl_eff = min(level, max(1, max{k : dist(tile centre, gaze) <= r_k})); single-pass mode writes it back to levels. */
ORC_API void orc_fovea_levels(int step, uint32_t* last, uint32_t* cur, uint32_t* levels, int num_tiles, int tiles_x, int tile_px, const float* gaze) {
    for (int t = 0; t < num_tiles; ++t) {
        uint32_t lvl = levels[t];
        if (gaze) {
            const float tx = ((float)(t % tiles_x) + 0.5f) * (float)tile_px, ty = ((float)(t / tiles_x) + 0.5f) * (float)tile_px;
            const float dx = tx - gaze[0], dy = ty - gaze[1];
            const float d2 = dx * dx + dy * dy;
            uint32_t cap = 1;
            for (int k = 1; k < 4; ++k) if (d2 <= gaze[2 + k] * gaze[2 + k]) cap = (uint32_t)k + 1;
            if (cap < lvl) lvl = cap;
            if (step < 0) levels[t] = lvl;
        }
        switch (step) {
            case 0: break;
            case 1: last[t] = 0; cur[t] = lvl >= 1 ? 1 : last[t]; break;
            case 2: case 3: case 4:
                last[t] = cur[t]; cur[t] = lvl >= (uint32_t)step ? (uint32_t)step : last[t]; break;
            default: last[t] = 0; cur[t] = lvl; break;
        }
    }
}

/* AMR renderCUDA (A/forward.cu:261-518): 32x32 tiles, 2x2 lattice rounds; writes only rendered lattice pixels */
ORC_API void orc_blend_amr(int W, int H, const uint32_t* ranges, const uint32_t* levels, const uint32_t* levels_last, int fovea_step,
                           const uint32_t* point_list, const float* means2D, const float* colors, const float* conic_opacity,
                           const float* bg, float* final_T, uint32_t* n_contrib, float* out_color) {
    const int gx = (W + 31) / 32, gy = (H + 31) / 32;
    const size_t HW = (size_t)H * W;
#pragma omp parallel for schedule(dynamic, 1)
    for (int by = 0; by < gy * 2; ++by)
        for (int bx = 0; bx < gx * 2; ++bx) {
            const int tile = (by / 2) * gx + (bx / 2);
            const uint32_t lvl_last = levels_last[tile];
            uint32_t lvl = levels[tile];
            if (lvl <= lvl_last) continue;
            const uint32_t ox = bx % 2, oy = by % 2;
            const uint32_t round = ox == 0 ? (oy == 0 ? 1 : 4) : (oy == 0 ? 3 : 2);
            if (lvl > 4) lvl = 4;
            if (fovea_step > 0 && round <= lvl_last) continue;
            if (round > lvl) continue;
            for (int ty = 0; ty < 16; ++ty)
                for (int tx = 0; tx < 16; ++tx) {
                    const uint32_t px = (bx / 2) * 32 + tx * 2 + ox, py = (by / 2) * 32 + ty * 2 + oy;
                    if (px >= (uint32_t)W || py >= (uint32_t)H) continue;
                    const size_t pix = (size_t)W * py + px;
                    float C[3];
                    blend_pixel(px, py, ranges[2 * tile], ranges[2 * tile + 1], point_list, means2D, colors, conic_opacity, bg,
                                final_T + pix, n_contrib + pix, C);
                    for (int ch = 0; ch < 3; ++ch) out_color[ch * HW + pix] = C[ch];
                }
        }
}

/* interpolateCUDA (A/forward.cu:520-648).  Where the reference reads a pixel that another CTA of the same
 * launch concurrently overwrites with out_color_precomp (level <= last in stepped mode), the copied-in
 * value is used (deterministic restatement of a racy read). */
ORC_API void orc_interpolate(int W, int H, const uint32_t* levels, const uint32_t* levels_last, int fovea_step, float* final_T,
                             uint32_t* n_contrib, float* out_color, const float* out_color_precomp) {
    const int gx = (W + 31) / 32;
    const size_t HW = (size_t)H * W;
    float* src_img = (float*)malloc(sizeof(float) * 3 * HW);
    float* src_T = (float*)malloc(sizeof(float) * HW);
    uint32_t* src_n = (uint32_t*)malloc(sizeof(uint32_t) * HW);
    memcpy(src_img, out_color, sizeof(float) * 3 * HW); memcpy(src_T, final_T, sizeof(float) * HW); memcpy(src_n, n_contrib, sizeof(uint32_t) * HW);
    if (!out_color_precomp) out_color_precomp = src_img;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const int tile = (y / 32) * gx + (x / 32);
            const uint32_t ox = x & 1, oy = y & 1;
            const int round = ox == 0 ? (oy == 0 ? 1 : 4) : (oy == 0 ? 3 : 2);
            uint32_t lvl = levels[tile]; if (lvl > 4) lvl = 4;
            const size_t pix = (size_t)W * y + x;
            int from_precomp = 0;
            if (fovea_step > 0) {
                const int last = (int)levels_last[tile];
                if ((int)lvl <= last) { for (int ch = 0; ch < 3; ++ch) out_color[ch * HW + pix] = out_color_precomp[ch * HW + pix]; from_precomp = 1; }
                if (round < last) { for (int ch = 0; ch < 3; ++ch) out_color[ch * HW + pix] = out_color_precomp[ch * HW + pix]; continue; }
            }
            if ((uint32_t)round <= lvl) continue;
            const uint32_t off = lvl >= 3 ? 1u : 0u;
            const uint32_t sx = ((uint32_t)x & ~1u) + off, sy = ((uint32_t)y & ~1u) + off;
            if (sx < (uint32_t)W && sy < (uint32_t)H) {
                const size_t sp = (size_t)W * sy + sx;
                final_T[pix] = src_T[sp]; n_contrib[pix] = src_n[sp];
                const float* s = from_precomp ? out_color_precomp : src_img;
                for (int ch = 0; ch < 3; ++ch) out_color[ch * HW + pix] = s[ch * HW + sp];
            }
        }
    free(src_img); free(src_T); free(src_n);
}
