// gsr_preprocess.cu -- per-Gaussian stages of the rasterizer for sm_100a.
//
//   preprocess_fwd   replaces preprocessCUDA<3> + computeCov3D + computeCov2D + computeColorFromSH
//                    (submodules/diff-gaussian-rasterization/cuda_rasterizer/forward.cu:20-256) and the
//                    first half of the scan: each 256-Gaussian CTA also emits its tiles_touched sum.
//   block_scan       replaces cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:277) at CTA granularity;
//                    the per-Gaussian prefix is re-derived inside the duplicate kernel (gsr_binning.cu).
//   mark_visible     replaces checkFrustum (rasterizer_impl.cu:54-66).
//   preprocess_bwd   replaces computeCov2DCUDA + preprocessCUDA(backward) + SH/cov3D backward
//                    (backward.cu:20-396), fused into ONE pass that also finalises every gradient row
//                    (so the host never zero-fills 300 B/Gaussian, rasterize_points.cu:151-159).
//
// HBM-bound stages: one thread per Gaussian, 128-bit loads where the AoS inputs allow it, culled
// Gaussians leave after reading 12 B.
#include <mutex>

#include "gsr_common.cuh"
#include "gsr_math.cuh"

namespace gsr {

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
struct PreFwdParams {
    int P, D, M;
    const float* __restrict__ means3D;
    const float* __restrict__ scales;
    const float* __restrict__ rotations;
    const float* __restrict__ opacities;
    const float* __restrict__ shs;
    const float* __restrict__ cov3D_precomp;
    const float* __restrict__ colors_precomp;
    const float* __restrict__ shs_rest;     // raw-parameter entry: shs = features_dc [P,1,3], shs_rest = features_rest [P,M-1,3]
    int raw;                                // 1: opacities / scales / rotations are raw (logit / log / un-normalised)
    float scale_modifier;
    const float* __restrict__ viewmatrix;
    const float* __restrict__ projmatrix;
    const float* __restrict__ cam_pos;
    int W, H;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    int tile, grid_x, grid_y;
    int* __restrict__ radii;
    float* __restrict__ depths;
    unsigned char* __restrict__ clamped;
    float2* __restrict__ means2D;
    float* __restrict__ cov3D;
    float4* __restrict__ conic_opacity;
    float* __restrict__ rgb;
    uint32_t* __restrict__ tiles_touched;
    uint2* __restrict__ rects;              // (x0 | x1 << 16, y0 | y1 << 16) tile rect, (0,0) when nothing is touched
    uint32_t* __restrict__ block_sums;
    uint32_t* __restrict__ block_sums_r;    // per-CTA sum of tile rows spanned (runs)
    uint32_t* __restrict__ block_sums_v;    // per-CTA number of visible Gaussians; bit 31 = "a Gaussian was culled although
                                            // prefiltered" (the reference would __trap)
    uint4* zero_ptr; size_t zero_n16;       // scratch cleared on the side (grid-stride), see PreprocessArgs
    uint32_t* __restrict__ ckeys;           // per-block compaction of the visible Gaussians: depth key (float bits) ...
    uint32_t* __restrict__ cids;            // ... index ...
    uint32_t* __restrict__ crows;           // ... (y0 << 16) | tile rows the rect spans (= runs the Gaussian emits) ...
    uint32_t* __restrict__ cxw;             // ... and (x0 << 10) | width in tiles: what the run emission needs, so that it never
                                            // gathers rects[] through the sorted order
    uint32_t* __restrict__ crank;           // ... and its slot inside its depth bin (return value of the counting atomic)
    FrameStatSlot* stat_slots;              // [kStatSlots] pair total, flags
    unsigned long long* bins64;             // [nbins] (runs << 32) | records per depth bin, cleared by the frame's memset node
    int nbins, depth_shift;
    GeomHeader* header;
    int prefiltered;
    int band_row0, band_row1;
};

// SH -> RGB (forward.cu:20-71).  `sh` points at this Gaussian's [M][3] coefficients.
__device__ __forceinline__ void sh_to_rgb(int deg, const float* __restrict__ sh, bool vec4_ok,
                                          float dx, float dy, float dz, float out[3], bool clamp_flag[3],
                                          const float* __restrict__ sh_rest = nullptr) {
    // direction = (pos - campos) / |pos - campos|
    const float len = __fsqrt_rn(sum3(dx, dx, dy, dy, dz, dz));
    const float x = __fdiv_rn(dx, len), y = __fdiv_rn(dy, len), z = __fdiv_rn(dz, len);

    float c[48];
    const int ncoef = (deg + 1) * (deg + 1);
    if (sh_rest != nullptr) {           // split layout: 3 floats of features_dc, then the row of features_rest
        c[0] = __ldg(sh); c[1] = __ldg(sh + 1); c[2] = __ldg(sh + 2);
#pragma unroll
        for (int i = 3; i < 48; ++i) c[i] = (i < 3 * ncoef) ? __ldg(sh_rest + i - 3) : 0.f;
    } else if (vec4_ok && deg == 3) {
        const float4* s4 = reinterpret_cast<const float4*>(sh);
#pragma unroll
        for (int i = 0; i < 12; ++i) {
            float4 v = __ldg(s4 + i);
            c[4 * i] = v.x; c[4 * i + 1] = v.y; c[4 * i + 2] = v.z; c[4 * i + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 48; ++i) c[i] = (i < 3 * ncoef) ? __ldg(sh + i) : 0.f;
    }

    float r[3];
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) r[ch] = __fmul_rn(kSH_C0, c[ch]);
    if (deg > 0) {
        const float a = __fmul_rn(y, kSH_C1), b = __fmul_rn(z, kSH_C1), d = __fmul_rn(x, kSH_C1);
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            r[ch] = __fmaf_rn(-a, c[3 + ch], r[ch]);
            r[ch] = __fmaf_rn(b, c[6 + ch], r[ch]);
            r[ch] = __fmaf_rn(-d, c[9 + ch], r[ch]);
        }
        if (deg > 1) {
            const float xx = __fmul_rn(x, x), yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
            const float xy = __fmul_rn(y, x), yz = __fmul_rn(z, y), xz = __fmul_rn(z, x);
            const float zz2 = __fadd_rn(zz, zz);
            const float t4 = __fmul_rn(xy, kSH_C2_0);
            const float t5 = __fmul_rn(yz, kSH_C2_1);
            const float t6 = __fmul_rn(__fadd_rn(__fadd_rn(zz2, -xx), -yy), kSH_C2_2);
            const float t7 = __fmul_rn(xz, kSH_C2_3);
            const float xx_yy = __fadd_rn(xx, -yy);
            const float t8 = __fmul_rn(xx_yy, kSH_C2_4);
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) {
                r[ch] = __fmaf_rn(t4, c[12 + ch], r[ch]);
                r[ch] = __fmaf_rn(t5, c[15 + ch], r[ch]);
                r[ch] = __fmaf_rn(t6, c[18 + ch], r[ch]);
                r[ch] = __fmaf_rn(t7, c[21 + ch], r[ch]);
                r[ch] = __fmaf_rn(t8, c[24 + ch], r[ch]);
            }
            if (deg > 2) {
                const float q = __fadd_rn(__fmaf_rn(zz, 4.f, -xx), -yy);     // 4zz - xx - yy
                const float t9 = __fmul_rn(__fmul_rn(y, kSH_C3_0), __fmaf_rn(xx, 3.f, -yy));
                const float t10 = __fmul_rn(__fmul_rn(xy, kSH_C3_1), z);
                const float t11 = __fmul_rn(__fmul_rn(y, kSH_C3_2), q);
                const float t12 = __fmul_rn(__fmul_rn(z, kSH_C3_3), __fmaf_rn(yy, -3.f, __fmaf_rn(xx, -3.f, zz2)));
                const float t13 = __fmul_rn(q, __fmul_rn(x, kSH_C3_4));
                const float t14 = __fmul_rn(xx_yy, __fmul_rn(z, kSH_C3_5));
                const float t15 = __fmul_rn(__fmul_rn(x, kSH_C3_6), __fmaf_rn(yy, -3.f, xx));
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                    r[ch] = __fmaf_rn(t9, c[27 + ch], r[ch]);
                    r[ch] = __fmaf_rn(t10, c[30 + ch], r[ch]);
                    r[ch] = __fmaf_rn(t11, c[33 + ch], r[ch]);
                    r[ch] = __fmaf_rn(t12, c[36 + ch], r[ch]);
                    r[ch] = __fmaf_rn(t13, c[39 + ch], r[ch]);
                    r[ch] = __fmaf_rn(t14, c[42 + ch], r[ch]);
                    r[ch] = __fmaf_rn(t15, c[45 + ch], r[ch]);
                }
            }
        }
    }
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        const float v = __fadd_rn(r[ch], 0.5f);
        clamp_flag[ch] = (v < 0.f);
        out[ch] = fmaxf(v, 0.f);
    }
}

// Sigma = (S R)^T (S R) (forward.cu:118-152).  Quaternion (r,x,y,z) is used un-normalised, as in the
// reference.  Multiplications by the literal zeros of S are exact no-ops for finite inputs and dropped.
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod, float4 q, float cov[6]) {
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    sx = __fmul_rn(sx, mod); sy = __fmul_rn(sy, mod); sz = __fmul_rn(sz, mod);
    const float xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
    const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
    const float C = __fmaf_rn(r, y, xz);
    const float G = __fmaf_rn(-r, y, xz);
    const float F = __fmaf_rn(y, z, -rx);
    const float Hh = __fmaf_rn(y, z, rx);
    const float B = __fmaf_rn(x, y, -rz);
    const float Dd = __fmaf_rn(x, y, rz);
    const float xx_yy = __fmaf_rn(x, x, yy);
    const float yy_zz = __fadd_rn(yy, zz);
    const float xx_zz = __fmaf_rn(x, x, zz);
    // R (column-major, glm): R[0]=(A,B2,C2) R[1]=(D2,E,F2) R[2]=(G2,H2,I)
    const float A = __fadd_rn(1.f, -__fadd_rn(yy_zz, yy_zz));
    const float E = __fadd_rn(1.f, -__fadd_rn(xx_zz, xx_zz));
    const float I = __fadd_rn(1.f, -__fadd_rn(xx_yy, xx_yy));
    const float B2 = __fadd_rn(B, B), C2 = __fadd_rn(C, C), D2 = __fadd_rn(Dd, Dd);
    const float F2 = __fadd_rn(F, F), G2 = __fadd_rn(G, G), H2 = __fadd_rn(Hh, Hh);
    // M = S * R : M[c] = (sx*R[c][0], sy*R[c][1], sz*R[c][2])
    const float M00 = __fmul_rn(sx, A),  M01 = __fmul_rn(sy, B2), M02 = __fmul_rn(sz, C2);
    const float M10 = __fmul_rn(sx, D2), M11 = __fmul_rn(sy, E),  M12 = __fmul_rn(sz, F2);
    const float M20 = __fmul_rn(sx, G2), M21 = __fmul_rn(sy, H2), M22 = __fmul_rn(sz, I);
    // Sigma = M^T M : Sigma[c][r] = M[r][0]M[c][0] + M[r][1]M[c][1] + M[r][2]M[c][2]
    cov[0] = sum3(M00, M00, M01, M01, M02, M02);
    cov[1] = sum3(M00, M10, M01, M11, M02, M12);
    cov[2] = sum3(M00, M20, M01, M21, M02, M22);
    cov[3] = sum3(M10, M10, M11, M11, M12, M12);
    cov[4] = sum3(M10, M20, M11, M21, M12, M22);
    cov[5] = sum3(M20, M20, M21, M21, M22, M22);
}

// EWA projection of the 3D covariance (forward.cu:74-113).  Returns (cov00, cov01, cov11) incl. the +0.3
// low-pass.  t = view-space mean (already computed by the caller with xform_row).
__device__ __forceinline__ float3 cov2d_from_cov3d(float tx, float ty, float tz, float focal_x, float focal_y,
                                                   float tan_fovx, float tan_fovy, const float c[6], const float* V) {
    const float limx = __fmul_rn(1.3f, tan_fovx), limy = __fmul_rn(1.3f, tan_fovy);
    const float txtz = __fdiv_rn(tx, tz), tytz = __fdiv_rn(ty, tz);
    const float cx = fminf(limx, fmaxf(-limx, txtz));
    const float cy = fminf(limy, fmaxf(-limy, tytz));
    const float tz2 = __fmul_rn(tz, tz);
    const float J00 = __fdiv_rn(focal_x, tz);
    const float J11 = __fdiv_rn(focal_y, tz);
    const float J02 = __fdiv_rn(__fmul_rn(__fmul_rn(tz, -cx), focal_x), tz2);   // -(fx * (cx*tz)) / tz^2
    const float J12 = __fdiv_rn(__fmul_rn(__fmul_rn(tz, -cy), focal_y), tz2);
    // T = W * J, rows 0/1 of J only: T0r = W[0][r]*J00 + W[2][r]*J02 ; T1r = W[1][r]*J11 + W[2][r]*J12
    // with W[0]=(V0,V4,V8) W[1]=(V1,V5,V9) W[2]=(V2,V6,V10)
    const float T00 = __fmaf_rn(V[2], J02, __fmul_rn(V[0], J00));
    const float T01 = __fmaf_rn(V[6], J02, __fmul_rn(V[4], J00));
    const float T02 = __fmaf_rn(V[10], J02, __fmul_rn(V[8], J00));
    const float T10 = __fmaf_rn(V[2], J12, __fmul_rn(V[1], J11));
    const float T11 = __fmaf_rn(V[6], J12, __fmul_rn(V[5], J11));
    const float T12 = __fmaf_rn(V[10], J12, __fmul_rn(V[9], J11));
    // X = T^T * Vrk^T :  X[c][r] = T[r][0]*Vrk[0][c] + T[r][1]*Vrk[1][c] + T[r][2]*Vrk[2][c]
    const float X00 = sum3(T00, c[0], T01, c[1], T02, c[2]);
    const float X10 = sum3(T00, c[1], T01, c[3], T02, c[4]);
    const float X20 = sum3(T00, c[2], T01, c[4], T02, c[5]);
    const float X01 = sum3(T10, c[0], T11, c[1], T12, c[2]);
    const float X11 = sum3(T10, c[1], T11, c[3], T12, c[4]);
    const float X21 = sum3(T10, c[2], T11, c[4], T12, c[5]);
    // cov = X * T
    const float cov00 = sum3(T00, X00, T01, X10, T02, X20);
    const float cov01 = sum3(T00, X01, T01, X11, T02, X21);
    const float cov11 = sum3(T10, X01, T11, X11, T12, X21);
    return make_float3(__fadd_rn(cov00, 0.3f), cov01, __fadd_rn(cov11, 0.3f));
}

__global__ void __launch_bounds__(kPreBlock, 4) preprocess_fwd_kernel(const PreFwdParams p) {
    __shared__ float sV[16], sPm[16], sCam[3];
    __shared__ uint32_t s_warp_sum[kPreBlock / 32], s_warp_sum_v[kPreBlock / 32];
    const int tid = threadIdx.x;
    if (tid < 16) { sV[tid] = p.viewmatrix[tid]; sPm[tid] = p.projmatrix[tid]; }
    if (tid < 3) sCam[tid] = p.cam_pos[tid];
    __syncthreads();

    const int idx = blockIdx.x * kPreBlock + tid;
    uint32_t touched = 0, rows = 0, key = 0, ry0 = 0, rxw = 0;
    int radius = 0, trapped = 0;
    for (size_t i = (size_t)idx; i < p.zero_n16; i += (size_t)gridDim.x * kPreBlock) p.zero_ptr[i] = make_uint4(0u, 0u, 0u, 0u);

    if (idx < p.P) {
        const float px = __ldg(p.means3D + 3 * (size_t)idx);
        const float py = __ldg(p.means3D + 3 * (size_t)idx + 1);
        const float pz = __ldg(p.means3D + 3 * (size_t)idx + 2);
        // near-plane cull (auxiliary.h:139-164): p_view.z <= 0.2 -> invisible
        const float depth = xform_row(sV[2], sV[6], sV[10], sV[14], px, py, pz);
        bool alive = depth > 0.2f;
        if (!alive && p.prefiltered) trapped = 1;   // reference: printf + __trap(); here: an error status from the call

        float2 pix = make_float2(0.f, 0.f);
        float3 conic = make_float3(0.f, 0.f, 0.f);
        uint32_t x0 = 0, y0 = 0, x1 = 0, y1 = 0;
        bool wrote_cov = false;
        if (alive) {
            const float hw = __fadd_rn(xform_row(sPm[3], sPm[7], sPm[11], sPm[15], px, py, pz), 0.0000001f);
            const float hx = xform_row(sPm[0], sPm[4], sPm[8], sPm[12], px, py, pz);
            const float hy = xform_row(sPm[1], sPm[5], sPm[9], sPm[13], px, py, pz);
            const float pw = __frcp_rn(hw);
            const float ndc_x = __fmul_rn(hx, pw), ndc_y = __fmul_rn(hy, pw);

            float c3[6];
            if (p.cov3D_precomp != nullptr) {
#pragma unroll
                for (int i = 0; i < 6; ++i) c3[i] = __ldg(p.cov3D_precomp + 6 * (size_t)idx + i);
            } else {
                float sx = __ldg(p.scales + 3 * (size_t)idx), sy = __ldg(p.scales + 3 * (size_t)idx + 1),
                      sz = __ldg(p.scales + 3 * (size_t)idx + 2);
                float4 q = __ldg(reinterpret_cast<const float4*>(p.rotations) + idx);
                if (p.raw) { sx = act_exp(sx); sy = act_exp(sy); sz = act_exp(sz); q = act_normalize(q); }
                cov3d_from_scale_rot(sx, sy, sz, p.scale_modifier, q, c3);
                wrote_cov = true;
#pragma unroll
                for (int i = 0; i < 6; ++i) p.cov3D[6 * (size_t)idx + i] = c3[i];
            }
            const float tx = xform_row(sV[0], sV[4], sV[8], sV[12], px, py, pz);
            const float ty = xform_row(sV[1], sV[5], sV[9], sV[13], px, py, pz);
            const float3 cov = cov2d_from_cov3d(tx, ty, depth, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy, c3, sV);

            const float det = __fmaf_rn(cov.x, cov.z, -__fmul_rn(cov.y, cov.y));
            alive = (det != 0.0f);
            if (alive) {
                const float det_inv = __frcp_rn(det);
                conic = make_float3(__fmul_rn(cov.z, det_inv), __fmul_rn(cov.y, -det_inv), __fmul_rn(cov.x, det_inv));
                const float mid = __fmul_rn(__fadd_rn(cov.x, cov.z), 0.5f);
                const float disc = __fsqrt_rn(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
                const float lam = fmaxf(__fadd_rn(mid, disc), __fadd_rn(mid, -disc));
                radius = (int)ceilf(__fmul_rn(__fsqrt_rn(lam), 3.f));
                pix = make_float2(ndc2pix(ndc_x, p.W), ndc2pix(ndc_y, p.H));
                get_rect(pix.x, pix.y, radius, p.tile, p.grid_x, p.grid_y, x0, y0, x1, y1);
                clip_band(y0, y1, p.band_row0, p.band_row1);
                touched = (x1 - x0) * (y1 - y0);
                alive = (touched != 0);
                if (!alive) radius = 0;
            }
        }
        // Every per-Gaussian output is written for EVERY Gaussian (zeros when culled): a warp then always stores whole
        // 32 B sectors.  Writing only the visible ~46 % left almost every touched sector partially written, and each
        // of those costs a DRAM read-modify-write (ncu: 553 -> 448 MB read per launch, 761 -> 670 MB in total).
        float rgb[3] = {0.f, 0.f, 0.f}; bool cl[3] = {false, false, false};
        if (alive && p.colors_precomp == nullptr) {
            const size_t shoff = (size_t)idx * p.M * 3;
            const bool vec_ok = ((p.M * 3) % 4 == 0) && ((reinterpret_cast<size_t>(p.shs) & 15) == 0);
            if (p.shs_rest != nullptr)
                sh_to_rgb(p.D, p.shs + 3 * (size_t)idx, false, __fadd_rn(px, -sCam[0]), __fadd_rn(py, -sCam[1]), __fadd_rn(pz, -sCam[2]), rgb, cl,
                          p.shs_rest + (size_t)idx * (p.M - 1) * 3);
            else
                sh_to_rgb(p.D, p.shs + shoff, vec_ok, __fadd_rn(px, -sCam[0]), __fadd_rn(py, -sCam[1]),
                          __fadd_rn(pz, -sCam[2]), rgb, cl);
        }
        if (p.colors_precomp == nullptr) {
            p.rgb[3 * (size_t)idx] = rgb[0]; p.rgb[3 * (size_t)idx + 1] = rgb[1]; p.rgb[3 * (size_t)idx + 2] = rgb[2];
            p.clamped[3 * (size_t)idx] = cl[0]; p.clamped[3 * (size_t)idx + 1] = cl[1]; p.clamped[3 * (size_t)idx + 2] = cl[2];
        }
        p.depths[idx] = alive ? depth : 0.f;
        key = __float_as_uint(depth);
        p.means2D[idx] = alive ? pix : make_float2(0.f, 0.f);
        float opac = 0.f;
        if (alive) { opac = __ldg(p.opacities + idx); if (p.raw) opac = act_sigmoid(opac); }
        p.conic_opacity[idx] = alive ? make_float4(conic.x, conic.y, conic.z, opac) : make_float4(0.f, 0.f, 0.f, 0.f);
        if (!wrote_cov && p.cov3D_precomp == nullptr) {
#pragma unroll
            for (int i = 0; i < 6; ++i) p.cov3D[6 * (size_t)idx + i] = 0.f;
        }
        p.radii[idx] = radius;
        p.tiles_touched[idx] = touched;
        if (!alive) { x0 = x1 = y0 = y1 = 0; }
        p.rects[idx] = make_uint2(x0 | (x1 << 16), y0 | (y1 << 16));
        rows = y1 - y0; ry0 = y0; rxw = (x0 << kRunX0Shift) | (x1 - x0);
    }

    // ---- block-local compaction, depth-bin count, frame statistics.
    // There is no scan kernel and no global cursor: the visible Gaussians of this 256-block go to the FRONT of the block's
    // own segment of the compaction arrays and the block's count to block_sums_v -- the depth sort (gsr_depthsort.cu) walks
    // blocks, not a dense array.  Every visible Gaussian also counts itself into its depth bin here (one 64-bit atomic:
    // +1 record, +rows runs; the returned count is its slot inside the bin), spread over the kernel's whole HBM-bound run
    // time instead of a pass of its own.  The pair total goes to one of kStatSlots statistics lines (same-address atomics
    // serialise at ~30 ns each: one hot word for 11 719 CTAs would cost more than the kernel itself).
    uint32_t v = touched;
    const uint32_t vis_mask = __ballot_sync(0xffffffffu, touched != 0);
    uint32_t rank_in_bin = 0, beyond = 0;
    if (touched) {
        uint32_t bin = (key - kDepthKey0) >> p.depth_shift;
        if (key < kDepthKey0 || bin >= (uint32_t)p.nbins) { bin = (uint32_t)p.nbins - 1u; beyond = 1u; }
        rank_in_bin = (uint32_t)atomicAdd(p.bins64 + bin, ((unsigned long long)rows << 32) | 1ull);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((tid & 31) == 0) { s_warp_sum[tid >> 5] = v; s_warp_sum_v[tid >> 5] = __popc(vis_mask); }
    const int flags_any = __syncthreads_or((trapped ? 1 : 0) | (beyond ? 2 : 0));
    if (touched) {
        uint32_t pos = __popc(vis_mask & ((1u << (tid & 31)) - 1u));
#pragma unroll
        for (int w = 0; w < kPreBlock / 32; ++w) pos += (w < (tid >> 5)) ? s_warp_sum_v[w] : 0u;
        const size_t at = (size_t)blockIdx.x * kPreBlock + pos;
        p.ckeys[at] = key; p.cids[at] = (uint32_t)idx; p.crows[at] = (ry0 << 16) | rows; p.cxw[at] = rxw; p.crank[at] = rank_in_bin;
    }
    if (tid == 0) {
        uint32_t s = 0, sv = 0;
#pragma unroll
        for (int w = 0; w < kPreBlock / 32; ++w) { s += s_warp_sum[w]; sv += s_warp_sum_v[w]; }
        p.block_sums_v[blockIdx.x] = sv;
        FrameStatSlot* slot = p.stat_slots + (blockIdx.x & (kStatSlots - 1));
        if (sv) atomicAdd(&slot->num_rendered, s);
        if (flags_any) atomicOr(&slot->flags, (uint32_t)flags_any);
    }
}

void launch_preprocess(const PreprocessArgs& a, const GeomPtrs& g, cudaStream_t s) {
    PreFwdParams p;
    p.P = a.P; p.D = a.D; p.M = a.M;
    p.means3D = a.means3D; p.scales = a.scales; p.rotations = a.rotations; p.opacities = a.opacities;
    p.shs = a.shs; p.cov3D_precomp = a.cov3D_precomp; p.colors_precomp = a.colors_precomp;
    p.shs_rest = a.shs_rest; p.raw = a.raw ? 1 : 0;
    p.scale_modifier = a.scale_modifier;
    p.viewmatrix = a.viewmatrix; p.projmatrix = a.projmatrix; p.cam_pos = a.cam_pos;
    p.W = a.W; p.H = a.H; p.tan_fovx = a.tan_fovx; p.tan_fovy = a.tan_fovy; p.focal_x = a.focal_x; p.focal_y = a.focal_y;
    p.tile = a.tile; p.grid_x = a.grid_x; p.grid_y = a.grid_y;
    p.radii = a.radii; p.depths = g.depths; p.clamped = g.clamped; p.means2D = g.means2D; p.cov3D = g.cov3D;
    p.conic_opacity = g.conic_opacity; p.rgb = g.rgb; p.tiles_touched = g.tiles_touched;
    p.rects = g.rects; p.block_sums_r = g.block_sums_r; p.block_sums_v = g.block_sums_v;
    p.block_sums = g.block_sums; p.header = g.header; p.prefiltered = a.prefiltered ? 1 : 0;
    p.band_row0 = a.band_row0; p.band_row1 = a.band_row1;
    p.zero_ptr = reinterpret_cast<uint4*>(a.zero_ptr); p.zero_n16 = a.zero_ptr ? a.zero_bytes / 16 : 0;
    p.ckeys = g.dkeys_a; p.cids = g.order_tmp; p.crows = g.dkeys_b; p.cxw = g.order; p.crank = g.crank; p.stat_slots = g.stat_slots;
    p.bins64 = g.bins64; p.nbins = g.nbins; p.depth_shift = depth_shift(a.P);
    preprocess_fwd_kernel<<<g.nblocks, kPreBlock, 0, s>>>(p);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// scan of the per-CTA sums: one CTA, exclusive offsets + grand total into the header.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) block_scan_kernel(const uint32_t* __restrict__ sums, uint32_t* __restrict__ offsets,
                                                          int n, GeomHeader* header, uint32_t P, const uint32_t* __restrict__ sums_t,
                                                          const uint32_t* __restrict__ sums_r, uint32_t* mailbox) {
    // offsets = exclusive scan of sums (+ grand total at [n]).  With a header: sums = visible counts, and the plain totals
    // of sums_t (tiles touched -> num_rendered) and sums_r (tile rows -> num_runs) are taken on the way.
    // One pass: every thread owns a contiguous slice of the (small, cache-resident) array.
    __shared__ uint32_t s_warp[32], s_warp_t[32], s_warp_r[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int per = (n + 1023) / 1024;
    const int begin = min(n, tid * per), end = min(n, begin + per);
    // with a header, bit 31 of a visible count is the prefiltered-trap flag of that block
    const uint32_t vmask = header != nullptr ? 0x7fffffffu : 0xffffffffu;
    uint32_t local = 0, t = 0, r = 0, trap = 0;
    for (int i = begin; i < end; ++i) { const uint32_t v = sums[i]; local += v & vmask; trap |= v & ~vmask; }
    if (header != nullptr) for (int i = begin; i < end; ++i) { t += sums_t[i]; r += sums_r[i]; }
    const int trap_any = __syncthreads_or(trap != 0u);
    uint32_t inc = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { t += __shfl_xor_sync(0xffffffffu, t, o); r += __shfl_xor_sync(0xffffffffu, r, o); }
    if (lane == 31) s_warp[wid] = inc;
    if (lane == 0) { s_warp_t[wid] = t; s_warp_r[wid] = r; }
    __syncthreads();
    uint32_t wbase = 0, total = 0, tt = 0, tr = 0;
#pragma unroll
    for (int k = 0; k < 32; ++k) { const uint32_t v = s_warp[k]; wbase += (k < wid) ? v : 0u; total += v; tt += s_warp_t[k]; tr += s_warp_r[k]; }
    uint32_t running = wbase + inc - local;
    for (int i = begin; i < end; ++i) { const uint32_t v = sums[i] & vmask; offsets[i] = running; running += v; }
    if (tid == 0) {
        offsets[n] = total;
        if (header) {   // every header field is (re)written here: no host memset of the header
            const uint32_t err = trap_any ? 0x100u : 0u;
            header->magic = kGeomMagic; header->P = P; header->num_rendered = tt; header->num_runs = tr; header->num_visible = total;
            header->sort_error = err;
            if (mailbox) {   // mapped pinned host memory: the host reads the totals without a DMA copy (a D2H memcpy would
                             // queue behind the caller's own frame read-back on the copy engine)
                volatile uint32_t* m = mailbox;
                m[0] = kGeomMagic; m[1] = P; m[2] = tt; m[3] = err; m[4] = tr; m[5] = total;
                // no __threadfence_system(): the host reads after the event recorded behind this kernel, and posted
                // PCIe writes stay ordered; a system fence here waits out whatever DMA traffic is queued (+50 us
                // while a 25 MB frame read-back is in flight)
            }
        }
    }
}

void launch_block_scan(const GeomPtrs& g, int P, uint32_t* mailbox_dev, cudaStream_t s) {
    block_scan_kernel<<<1, 1024, 0, s>>>(g.block_sums_v, g.block_offsets, g.nblocks, g.header, (uint32_t)P, g.block_sums, g.block_sums_r, mailbox_dev);
    count_launch();
}

void launch_scan_u32(const uint32_t* sums, uint32_t* offsets, int n, cudaStream_t s) {
    block_scan_kernel<<<1, 1024, 0, s>>>(sums, offsets, n, nullptr, 0u, nullptr, nullptr, nullptr);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// mark_visible
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) mark_visible_kernel(int P, const float* __restrict__ means3D,
                                                           const float* __restrict__ V, unsigned char* __restrict__ present) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= P) return;
    const float x = __ldg(means3D + 3 * (size_t)idx), y = __ldg(means3D + 3 * (size_t)idx + 1), z = __ldg(means3D + 3 * (size_t)idx + 2);
    const float depth = xform_row(__ldg(V + 2), __ldg(V + 6), __ldg(V + 10), __ldg(V + 14), x, y, z);
    present[idx] = depth > 0.2f ? 1 : 0;
}

void launch_mark_visible(int P, const float* means3D, const float* view, const float* proj, unsigned char* present, cudaStream_t s) {
    (void)proj;  // the reference computes p_proj but the side-plane test is commented out (auxiliary.h:154)
    if (P <= 0) return;
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, view, present);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// backward: computeCov2DCUDA (backward.cu:144-274) + preprocessCUDA bwd (:346-396) + SH bwd (:20-139)
// + cov3D bwd (:278-341), one pass.  Gradient tolerance is 1e-3 relative (the reference accumulates with
// non-deterministic float atomics), so arithmetic here is plain FP32 without pinned contraction.
// ------------------------------------------------------------------------------------------------
struct PreBwdParams {
    int P, D, M;                   // P = one past the last Gaussian this launch handles
    int block0;                    // first 256-Gaussian block of this launch (range launches, gsr_backward_preprocess_range)
    const float* __restrict__ means3D; const float* __restrict__ scales; const float* __restrict__ rotations;
    const float* __restrict__ shs; const float* __restrict__ cov3D;
    const float* __restrict__ shs_rest; int raw;      // raw-parameter entry (see PreFwdParams)
    int cov_from_scales, colors_from_sh;
    float scale_modifier;
    const float* __restrict__ viewmatrix; const float* __restrict__ projmatrix; const float* __restrict__ cam_pos;
    float focal_x, focal_y, tan_fovx, tan_fovy;
    const int* __restrict__ radii; const unsigned char* __restrict__ clamped;
    const float* __restrict__ scratch;
    float *dL_dmeans2D, *dL_dcolors, *dL_dopacity, *dL_dmeans3D, *dL_dcov3D, *dL_dsh, *dL_dscales, *dL_drot, *dL_dconic;
    float* dL_drgb_clamped;        // optional [P,3]; with dL_dsh == nullptr the SH rows are not written (gsr_grads)
    // fused view-parallel exchange (gsr_grads.push_*): rows of the 11 dense floats go to the owner rank's staging area
    char* const* push_bases; size_t push_offset; int push_rank, push_world, push_rows_per_rank;
    uint32_t push_epoch;
};

// SH rows are staged through shared memory (STAGED): a thread's 3M coefficients / gradients are 192 B apart from its
// neighbour's, so direct per-thread access costs one 32 B sector per 4 useful bytes.  Rows sit in smem with an odd
// stride (conflict-free per-thread access); a warp loads only its VISIBLE rows and stores all 32 gradient rows as one
// contiguous 6 KB run with 128-bit accesses.
constexpr int kShStride = 49;                 // floats per staged row (3 * 16 coefficients + 1 pad)

template <bool STAGED>
__global__ void __launch_bounds__(256) preprocess_bwd_kernel(const PreBwdParams p) {
    __shared__ float V[16], Pm[16], cam[3];
    extern __shared__ float s_sh[];           // STAGED: [256][kShStride]
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid < 16) { V[tid] = p.viewmatrix[tid]; Pm[tid] = p.projmatrix[tid]; }
    if (tid < 3) cam[tid] = p.cam_pos[tid];
    __syncthreads();
    const int idx = (blockIdx.x + p.block0) * 256 + tid;
    const bool valid = idx < p.P;
    const bool pushing = p.push_bases != nullptr;                   // whole CTAs stay: every row is pushed, then the block is published
    if (!STAGED && !valid && !pushing) return;
    const size_t i3 = 3 * (size_t)idx;

    const bool visible = valid && p.radii[idx] > 0;
    float4 sA = make_float4(0, 0, 0, 0), sB = sA, sC = sA;
    if (visible) {
        const float4* sp = reinterpret_cast<const float4*>(p.scratch) + 3 * (size_t)idx;
        sA = sp[0]; sB = sp[1]; sC = sp[2];
    }
    // blend-stage gradients leave in the reference's tensor shapes
    if (valid) {
        p.dL_dmeans2D[i3] = sA.x; p.dL_dmeans2D[i3 + 1] = sA.y; p.dL_dmeans2D[i3 + 2] = 0.f;
        if (p.push_bases == nullptr) p.dL_dopacity[idx] = sB.y;
        p.dL_dcolors[i3] = sB.z; p.dL_dcolors[i3 + 1] = sB.w; p.dL_dcolors[i3 + 2] = sC.x;
        if (p.dL_dconic) reinterpret_cast<float4*>(p.dL_dconic)[idx] = make_float4(sA.z, sA.w, 0.f, sB.x);
    }

    float dcov[6] = {0, 0, 0, 0, 0, 0};
    float dmx = 0.f, dmy = 0.f, dmz = 0.f;
    float dscale[3] = {0, 0, 0};
    float4 drot = make_float4(0, 0, 0, 0);

    const float mx = visible ? p.means3D[i3] : 0.f, my = visible ? p.means3D[i3 + 1] : 0.f, mz = visible ? p.means3D[i3 + 2] : 0.f;

    if (visible) {
        // ---------------- conic -> cov2D -> cov3D, mean (backward.cu:144-274)
        float c3[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) c3[i] = p.cov3D[6 * (size_t)idx + i];
        const float dcx = sA.z, dcy = sA.w, dcz = sB.x;        // dL/dconic (x, y, w)
        float tx = V[0] * mx + V[4] * my + V[8] * mz + V[12];
        float ty = V[1] * mx + V[5] * my + V[9] * mz + V[13];
        const float tz = V[2] * mx + V[6] * my + V[10] * mz + V[14];
        const float limx = 1.3f * p.tan_fovx, limy = 1.3f * p.tan_fovy;
        const float txtz = tx / tz, tytz = ty / tz;
        tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
        ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        const float hx = p.focal_x, hy = p.focal_y;
        const float J00 = hx / tz, J11 = hy / tz, J02 = -(hx * tx) / (tz * tz), J12 = -(hy * ty) / (tz * tz);
        const float T00 = V[0] * J00 + V[2] * J02, T01 = V[4] * J00 + V[6] * J02, T02 = V[8] * J00 + V[10] * J02;
        const float T10 = V[1] * J11 + V[2] * J12, T11 = V[5] * J11 + V[6] * J12, T12 = V[9] * J11 + V[10] * J12;
        // rows of T * Vrk (Vrk symmetric: c0 c1 c2 / c1 c3 c4 / c2 c4 c5)
        const float A0 = T00 * c3[0] + T01 * c3[1] + T02 * c3[2];
        const float A1 = T00 * c3[1] + T01 * c3[3] + T02 * c3[4];
        const float A2 = T00 * c3[2] + T01 * c3[4] + T02 * c3[5];
        const float B0 = T10 * c3[0] + T11 * c3[1] + T12 * c3[2];
        const float B1 = T10 * c3[1] + T11 * c3[3] + T12 * c3[4];
        const float B2 = T10 * c3[2] + T11 * c3[4] + T12 * c3[5];
        const float a = A0 * T00 + A1 * T01 + A2 * T02 + 0.3f;
        const float b = B0 * T00 + B1 * T01 + B2 * T02;
        const float c = B0 * T10 + B1 * T11 + B2 * T12 + 0.3f;
        const float denom = a * c - b * b;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
        if (denom2inv != 0.f) {
            dL_da = denom2inv * (-c * c * dcx + 2.f * b * c * dcy + (denom - a * c) * dcz);
            dL_dc = denom2inv * (-a * a * dcz + 2.f * a * b * dcy + (denom - a * c) * dcx);
            dL_db = denom2inv * 2.f * (b * c * dcx - (denom + 2.f * b * b) * dcy + a * b * dcz);
            dcov[0] = T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc;
            dcov[3] = T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc;
            dcov[5] = T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc;
            dcov[1] = 2.f * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2.f * T10 * T11 * dL_dc;
            dcov[2] = 2.f * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2.f * T10 * T12 * dL_dc;
            dcov[4] = 2.f * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2.f * T11 * T12 * dL_dc;
        }
        const float dT00 = 2.f * A0 * dL_da + B0 * dL_db, dT01 = 2.f * A1 * dL_da + B1 * dL_db, dT02 = 2.f * A2 * dL_da + B2 * dL_db;
        const float dT10 = 2.f * B0 * dL_dc + A0 * dL_db, dT11 = 2.f * B1 * dL_dc + A1 * dL_db, dT12 = 2.f * B2 * dL_dc + A2 * dL_db;
        const float dJ00 = V[0] * dT00 + V[4] * dT01 + V[8] * dT02;
        const float dJ02 = V[2] * dT00 + V[6] * dT01 + V[10] * dT02;
        const float dJ11 = V[1] * dT10 + V[5] * dT11 + V[9] * dT12;
        const float dJ12 = V[2] * dT10 + V[6] * dT11 + V[10] * dT12;
        const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
        const float dtx = x_grad_mul * -hx * itz2 * dJ02;
        const float dty = y_grad_mul * -hy * itz2 * dJ12;
        const float dtz = -hx * itz2 * dJ00 - hy * itz2 * dJ11 + (2.f * hx * tx) * itz3 * dJ02 + (2.f * hy * ty) * itz3 * dJ12;
        dmx = V[0] * dtx + V[1] * dty + V[2] * dtz;
        dmy = V[4] * dtx + V[5] * dty + V[6] * dtz;
        dmz = V[8] * dtx + V[9] * dty + V[10] * dtz;

        // ---------------- screen-space mean -> 3D mean (backward.cu:369-387)
        const float mw = 1.0f / (Pm[3] * mx + Pm[7] * my + Pm[11] * mz + Pm[15] + 0.0000001f);
        const float mul1 = (Pm[0] * mx + Pm[4] * my + Pm[8] * mz + Pm[12]) * mw * mw;
        const float mul2 = (Pm[1] * mx + Pm[5] * my + Pm[9] * mz + Pm[13]) * mw * mw;
        const float d2x = sA.x, d2y = sA.y;
        dmx += (Pm[0] * mw - Pm[3] * mul1) * d2x + (Pm[1] * mw - Pm[3] * mul2) * d2y;
        dmy += (Pm[4] * mw - Pm[7] * mul1) * d2x + (Pm[5] * mw - Pm[7] * mul2) * d2y;
        dmz += (Pm[8] * mw - Pm[11] * mul1) * d2x + (Pm[9] * mw - Pm[11] * mul2) * d2y;
    }

    // ---------------- SH (backward.cu:20-139)
    const bool sh_rows = p.dL_dsh != nullptr;                 // false: only the clamped colour gradient leaves
    if (p.dL_drgb_clamped != nullptr && valid && !(visible && p.colors_from_sh)) {
        p.dL_drgb_clamped[i3] = 0.f; p.dL_drgb_clamped[i3 + 1] = 0.f; p.dL_drgb_clamped[i3 + 2] = 0.f;
    }
    if (sh_rows || p.colors_from_sh) {
        const int L = 3 * p.M;
        const int warp_idx0 = (blockIdx.x + p.block0) * 256 + wid * 32;
        float* wbase = s_sh + wid * 32 * kShStride;
        float* out = STAGED ? (wbase + lane * kShStride) : (sh_rows ? p.dL_dsh + (size_t)idx * L : nullptr);
        const int ncoef = (p.D + 1) * (p.D + 1);
        const int f4_per_row = L >> 2;
        if (STAGED) {
            // cooperative load of the visible rows: 32 / f4_per_row rows per instruction, 128-bit each
            const uint32_t m = __ballot_sync(0xffffffffu, visible && p.colors_from_sh);
            const int nv = __popc(m), rpi = 32 / f4_per_row;
            const int sel = lane / f4_per_row, part = lane - sel * f4_per_row;
            for (int i = 0; i < nv; i += rpi) {
                const int k = i + sel;
                if (sel < rpi && k < nv) {
                    const int src_lane = __fns(m, 0, k + 1);
                    float4 v;
                    if (p.shs_rest != nullptr) {       // split layout: coefficient c of the row = dc[c] (c < 3) or rest[c - 3]
                        const float* dc = p.shs + 3 * (size_t)(warp_idx0 + src_lane);
                        const float* rs = p.shs_rest + (size_t)(warp_idx0 + src_lane) * (L - 3);
                        float t[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) { const int c = 4 * part + e; t[e] = c < 3 ? __ldg(dc + c) : __ldg(rs + c - 3); }
                        v = make_float4(t[0], t[1], t[2], t[3]);
                    } else v = __ldg(reinterpret_cast<const float4*>(p.shs + (size_t)(warp_idx0 + src_lane) * L) + part);
                    float* dst = wbase + src_lane * kShStride + 4 * part;
                    dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w;
                }
            }
            __syncwarp();
        }
        if (visible && p.colors_from_sh) {
            const float* sh = STAGED ? out : (p.shs + (size_t)idx * L);
            const float ox = mx - cam[0], oy = my - cam[1], oz = mz - cam[2];
            const float inv_len = 1.f / sqrtf(ox * ox + oy * oy + oz * oz);
            const float x = ox * inv_len, y = oy * inv_len, z = oz * inv_len;
            float g[3] = {sB.z, sB.w, sC.x};
            g[0] *= p.clamped[i3] ? 0.f : 1.f; g[1] *= p.clamped[i3 + 1] ? 0.f : 1.f; g[2] *= p.clamped[i3 + 2] ? 0.f : 1.f;
            if (p.dL_drgb_clamped != nullptr) { p.dL_drgb_clamped[i3] = g[0]; p.dL_drgb_clamped[i3 + 1] = g[1]; p.dL_drgb_clamped[i3 + 2] = g[2]; }
            float w[16];                       // dRGB/dsh_k (same for the three channels)
            float dx_[3] = {0, 0, 0}, dy_[3] = {0, 0, 0}, dz_[3] = {0, 0, 0};   // dRGB/d(dir)
            w[0] = kSH_C0;
#pragma unroll
            for (int k = 1; k < 16; ++k) w[k] = 0.f;
            auto S = [&](int k, int ch) { return sh[3 * k + ch]; };
            if (p.D > 0) {
                w[1] = -kSH_C1 * y; w[2] = kSH_C1 * z; w[3] = -kSH_C1 * x;
                for (int ch = 0; ch < 3; ++ch) { dx_[ch] = -kSH_C1 * S(3, ch); dy_[ch] = -kSH_C1 * S(1, ch); dz_[ch] = kSH_C1 * S(2, ch); }
                if (p.D > 1) {
                    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                    w[4] = kSH_C2_0 * xy; w[5] = kSH_C2_1 * yz; w[6] = kSH_C2_2 * (2.f * zz - xx - yy);
                    w[7] = kSH_C2_3 * xz; w[8] = kSH_C2_4 * (xx - yy);
                    for (int ch = 0; ch < 3; ++ch) {
                        dx_[ch] += kSH_C2_0 * y * S(4, ch) + kSH_C2_2 * 2.f * -x * S(6, ch) + kSH_C2_3 * z * S(7, ch) + kSH_C2_4 * 2.f * x * S(8, ch);
                        dy_[ch] += kSH_C2_0 * x * S(4, ch) + kSH_C2_1 * z * S(5, ch) + kSH_C2_2 * 2.f * -y * S(6, ch) + kSH_C2_4 * 2.f * -y * S(8, ch);
                        dz_[ch] += kSH_C2_1 * y * S(5, ch) + kSH_C2_2 * 2.f * 2.f * z * S(6, ch) + kSH_C2_3 * x * S(7, ch);
                    }
                    if (p.D > 2) {
                        w[9] = kSH_C3_0 * y * (3.f * xx - yy);
                        w[10] = kSH_C3_1 * xy * z;
                        w[11] = kSH_C3_2 * y * (4.f * zz - xx - yy);
                        w[12] = kSH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
                        w[13] = kSH_C3_4 * x * (4.f * zz - xx - yy);
                        w[14] = kSH_C3_5 * z * (xx - yy);
                        w[15] = kSH_C3_6 * x * (xx - 3.f * yy);
                        for (int ch = 0; ch < 3; ++ch) {
                            dx_[ch] += kSH_C3_0 * S(9, ch) * 3.f * 2.f * xy + kSH_C3_1 * S(10, ch) * yz + kSH_C3_2 * S(11, ch) * -2.f * xy +
                                       kSH_C3_3 * S(12, ch) * -3.f * 2.f * xz + kSH_C3_4 * S(13, ch) * (-3.f * xx + 4.f * zz - yy) +
                                       kSH_C3_5 * S(14, ch) * 2.f * xz + kSH_C3_6 * S(15, ch) * 3.f * (xx - yy);
                            dy_[ch] += kSH_C3_0 * S(9, ch) * 3.f * (xx - yy) + kSH_C3_1 * S(10, ch) * xz + kSH_C3_2 * S(11, ch) * (-3.f * yy + 4.f * zz - xx) +
                                       kSH_C3_3 * S(12, ch) * -3.f * 2.f * yz + kSH_C3_4 * S(13, ch) * -2.f * xy +
                                       kSH_C3_5 * S(14, ch) * -2.f * yz + kSH_C3_6 * S(15, ch) * -3.f * 2.f * xy;
                            dz_[ch] += kSH_C3_1 * S(10, ch) * xy + kSH_C3_2 * S(11, ch) * 4.f * 2.f * yz + kSH_C3_3 * S(12, ch) * 3.f * (2.f * zz - xx - yy) +
                                       kSH_C3_4 * S(13, ch) * 4.f * 2.f * xz + kSH_C3_5 * S(14, ch) * (xx - yy);
                        }
                    }
                }
            }
            if (sh_rows) {
                for (int k = 0; k < p.M; ++k) {
                    const float wk = (k < ncoef && k < 16) ? w[k] : 0.f;
                    out[3 * k] = wk * g[0]; out[3 * k + 1] = wk * g[1]; out[3 * k + 2] = wk * g[2];
                }
            }
            // direction gradient -> mean, through the normalisation (dnormvdv, auxiliary.h:107-118)
            const float ddx = dx_[0] * g[0] + dx_[1] * g[1] + dx_[2] * g[2];
            const float ddy = dy_[0] * g[0] + dy_[1] * g[1] + dy_[2] * g[2];
            const float ddz = dz_[0] * g[0] + dz_[1] * g[1] + dz_[2] * g[2];
            const float sum2 = ox * ox + oy * oy + oz * oz;
            const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
            dmx += ((sum2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * invsum32;
            dmy += (-ox * oy * ddx + (sum2 - oy * oy) * ddy - oz * oy * ddz) * invsum32;
            dmz += (-ox * oz * ddx - oy * oz * ddy + (sum2 - oz * oz) * ddz) * invsum32;
        } else if (sh_rows && (STAGED || valid)) {
            for (int k = 0; k < L; ++k) out[k] = 0.f;
        }
        if (STAGED && sh_rows) {
            __syncwarp();
            const int nrows = min(32, p.P - warp_idx0);
            const int total4 = nrows * f4_per_row;
            float4* gout = reinterpret_cast<float4*>(p.dL_dsh + (size_t)warp_idx0 * L);
            for (int e = lane; e < total4; e += 32) {
                const int r = e / f4_per_row, c = (e - r * f4_per_row) * 4;
                const float* src = wbase + r * kShStride + c;
                gout[e] = make_float4(src[0], src[1], src[2], src[3]);
            }
        }
    }
    if (!valid && !pushing) return;

    // ---------------- cov3D -> scale, rotation (backward.cu:278-341)
    if (visible && p.cov_from_scales) {
        float4 q = reinterpret_cast<const float4*>(p.rotations)[idx];
        float sc0 = p.scales[i3], sc1 = p.scales[i3 + 1], sc2 = p.scales[i3 + 2];
        if (p.raw) { q = act_normalize(q); sc0 = act_exp(sc0); sc1 = act_exp(sc1); sc2 = act_exp(sc2); }
        const float r = q.x, x = q.y, y = q.z, z = q.w;
        const float s0 = p.scale_modifier * sc0, s1 = p.scale_modifier * sc1, s2 = p.scale_modifier * sc2;
        // R[c][r] as in the forward (glm column-major)
        const float R00 = 1.f - 2.f * (y * y + z * z), R01 = 2.f * (x * y - r * z), R02 = 2.f * (x * z + r * y);
        const float R10 = 2.f * (x * y + r * z), R11 = 1.f - 2.f * (x * x + z * z), R12 = 2.f * (y * z - r * x);
        const float R20 = 2.f * (x * z - r * y), R21 = 2.f * (y * z + r * x), R22 = 1.f - 2.f * (x * x + y * y);
        // M = S*R : M[c][r] = s_r * R[c][r]
        const float M00 = s0 * R00, M01 = s1 * R01, M02 = s2 * R02;
        const float M10 = s0 * R10, M11 = s1 * R11, M12 = s2 * R12;
        const float M20 = s0 * R20, M21 = s1 * R21, M22 = s2 * R22;
        // dL_dSigma (symmetric, off-diagonals halved)
        const float S00 = dcov[0], S01 = 0.5f * dcov[1], S02 = 0.5f * dcov[2], S11 = dcov[3], S12 = 0.5f * dcov[4], S22 = dcov[5];
        // dL_dM = 2 * M * dL_dSigma : dM[c][r] = 2 * sum_k M[k][r] * Sg[c][k]
        const float dM00 = 2.f * (M00 * S00 + M10 * S01 + M20 * S02), dM01 = 2.f * (M01 * S00 + M11 * S01 + M21 * S02), dM02 = 2.f * (M02 * S00 + M12 * S01 + M22 * S02);
        const float dM10 = 2.f * (M00 * S01 + M10 * S11 + M20 * S12), dM11 = 2.f * (M01 * S01 + M11 * S11 + M21 * S12), dM12 = 2.f * (M02 * S01 + M12 * S11 + M22 * S12);
        const float dM20 = 2.f * (M00 * S02 + M10 * S12 + M20 * S22), dM21 = 2.f * (M01 * S02 + M11 * S12 + M21 * S22), dM22 = 2.f * (M02 * S02 + M12 * S12 + M22 * S22);
        // Rt = transpose(R), dMt = transpose(dM): Rt[c][r] = R[r][c], dMt[c][r] = dM[r][c]
        // dL_dscale_i = dot(Rt[i], dMt[i]) = sum_r R[r][i] * dM[r][i]
        dscale[0] = R00 * dM00 + R10 * dM10 + R20 * dM20;
        dscale[1] = R01 * dM01 + R11 * dM11 + R21 * dM21;
        dscale[2] = R02 * dM02 + R12 * dM12 + R22 * dM22;
        // dMt[i] *= s_i  => t[i][j] = dMt[i][j] * s_i = dM[j][i] * s_i
        const float t00 = dM00 * s0, t01 = dM10 * s0, t02 = dM20 * s0;
        const float t10 = dM01 * s1, t11 = dM11 * s1, t12 = dM21 * s1;
        const float t20 = dM02 * s2, t21 = dM12 * s2, t22 = dM22 * s2;
        drot.x = 2.f * z * (t01 - t10) + 2.f * y * (t20 - t02) + 2.f * x * (t12 - t21);
        drot.y = 2.f * y * (t10 + t01) + 2.f * z * (t20 + t02) + 2.f * r * (t12 - t21) - 4.f * x * (t22 + t11);
        drot.z = 2.f * x * (t10 + t01) + 2.f * r * (t20 - t02) + 2.f * z * (t12 + t21) - 4.f * y * (t22 + t00);
        drot.w = 2.f * r * (t01 - t10) + 2.f * x * (t20 + t02) + 2.f * y * (t12 + t21) - 4.f * z * (t11 + t00);
        // note: like the reference, dL_dscale is the gradient w.r.t. (modifier * scale); no extra factor is applied
    }

    if (valid) {
#pragma unroll
        for (int i = 0; i < 6; ++i) p.dL_dcov3D[6 * (size_t)idx + i] = dcov[i];
    }
    if (pushing) {
        // The CTA's 256 x 11 dense floats go to the rank that OWNS the block (peer memory over NVLink).  Ownership is block-cyclic
        // (256-Gaussian blocks dealt round-robin to the ranks): at any moment the CTAs in flight push to ALL owners at once;
        // contiguous slices would make every rank push to the same owner at the same time (7-to-1 incast on one GPU's ingress).
        // Wire format = 128-byte LINES of 30 data floats + a 2-word flag (the epoch), each written by 8 lanes of one store
        // instruction: a line arrives whole, so the owner validates every line by its own flag and no fence, ready word or kernel
        // boundary is needed -- gsr_peer_owner_sum consumes the block while this kernel is still running.  (A fence.sys + ready
        // word per CTA was measured first: the fence drains the SM's whole memory pipeline, 0.72 -> 1.53 ms for the kernel.)
        // Rows of invisible Gaussians are zeros -- the terms the dense per-view tensors hold for them -- so the owner adds all
        // ranks' lines unconditionally: no masks, no read-modify-write.  94 lines (2816 floats) per block in a 12 KB slot.
        const int blk = blockIdx.x + p.block0, owner = blk % p.push_world;
        float* brow = s_sh;                                                 // [256][11]; the SH staging rows are dead by now
        __syncthreads();
        float* r = brow + tid * 11;
        r[0] = visible ? dmx : 0.f; r[1] = visible ? dmy : 0.f; r[2] = visible ? dmz : 0.f; r[3] = visible ? sB.y : 0.f;
        r[4] = visible ? dscale[0] : 0.f; r[5] = visible ? dscale[1] : 0.f; r[6] = visible ? dscale[2] : 0.f;
        r[7] = visible ? drot.x : 0.f; r[8] = visible ? drot.y : 0.f; r[9] = visible ? drot.z : 0.f; r[10] = visible ? drot.w : 0.f;
        __syncthreads();
        float4* dst = reinterpret_cast<float4*>(p.push_bases[owner] + p.push_offset) +
                      ((size_t)p.push_rank * (p.push_rows_per_rank >> 8) + blk / p.push_world) * kPushSlotF4;
        const float flag = __uint_as_float(p.push_epoch);
        for (int e = tid; e < kPushLines * 8; e += 256) {
            const int f0 = (e >> 3) * 30 + (e & 7) * 4;                     // first data float of this 16-byte piece
            float4 v;
            v.x = f0 < 2816 ? brow[f0] : 0.f; v.y = f0 + 1 < 2816 ? brow[f0 + 1] : 0.f;
            if ((e & 7) == 7) { v.z = flag; v.w = flag; }
            else { v.z = f0 + 2 < 2816 ? brow[f0 + 2] : 0.f; v.w = f0 + 3 < 2816 ? brow[f0 + 3] : 0.f; }
            dst[e] = v;
        }
        return;
    }
    p.dL_dmeans3D[i3] = dmx; p.dL_dmeans3D[i3 + 1] = dmy; p.dL_dmeans3D[i3 + 2] = dmz;
    p.dL_dscales[i3] = dscale[0]; p.dL_dscales[i3 + 1] = dscale[1]; p.dL_dscales[i3 + 2] = dscale[2];
    reinterpret_cast<float4*>(p.dL_drot)[idx] = drot;
}

// resolve both instantiations now (lazy module loading, see gsr_peer.cu preload)
void preload_preprocess_backward() {
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, preprocess_bwd_kernel<true>);
    cudaFuncGetAttributes(&fa, preprocess_bwd_kernel<false>);
}

void launch_preprocess_backward(const PreBwdArgs& a, cudaStream_t s) {
    if (a.P <= 0) return;
    const int first = a.count >= 0 ? a.first : 0;
    const int end = a.count >= 0 ? min(a.P, a.first + a.count) : a.P;
    if (end <= first) return;
    PreBwdParams p;
    p.P = end; p.block0 = first / 256; p.D = a.D; p.M = a.M;
    p.means3D = a.means3D; p.scales = a.scales; p.rotations = a.rotations; p.shs = a.shs; p.cov3D = a.cov3D;
    p.shs_rest = a.shs_rest; p.raw = a.raw ? 1 : 0;
    p.cov_from_scales = a.cov_from_scales ? 1 : 0; p.colors_from_sh = a.colors_from_sh ? 1 : 0;
    p.scale_modifier = a.scale_modifier;
    p.viewmatrix = a.viewmatrix; p.projmatrix = a.projmatrix; p.cam_pos = a.cam_pos;
    p.focal_x = a.focal_x; p.focal_y = a.focal_y; p.tan_fovx = a.tan_fovx; p.tan_fovy = a.tan_fovy;
    p.radii = a.radii; p.clamped = a.clamped; p.scratch = a.scratch;
    p.dL_dmeans2D = a.out.dL_dmeans2D; p.dL_dcolors = a.out.dL_dcolors; p.dL_dopacity = a.out.dL_dopacity;
    p.dL_dmeans3D = a.out.dL_dmeans3D; p.dL_dcov3D = a.out.dL_dcov3D; p.dL_dsh = a.out.dL_dsh;
    p.dL_dscales = a.out.dL_dscales; p.dL_drot = a.out.dL_drotations; p.dL_dconic = a.out.dL_dconic;
    p.dL_drgb_clamped = a.out.dL_dcolors_clamped;
    p.push_bases = (char* const*)a.out.push_bases; p.push_offset = a.out.push_offset; p.push_rank = a.out.push_rank;
    p.push_rows_per_rank = a.out.push_rows_per_rank; p.push_world = a.out.push_world > 0 ? a.out.push_world : 1;
    p.push_epoch = a.out.push_epoch;
    // staged SH path: needs M % 4 == 0 (128-bit rows), M <= 16 (row fits the staged stride) and 16 B-aligned tensors
    const bool staged = (p.dL_dsh != nullptr || p.colors_from_sh) && p.M > 0 && p.M <= 16 && (p.M % 4) == 0 && ((size_t)p.dL_dsh & 15) == 0 &&
                        (!p.colors_from_sh || p.shs_rest != nullptr || ((size_t)p.shs & 15) == 0);
    if (staged) {
        const int smem = 256 * kShStride * (int)sizeof(float);
        static bool attr_done = false;
        if (!attr_done) { cudaFuncSetAttribute(preprocess_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); attr_done = true; }
        preprocess_bwd_kernel<true><<<(end - first + 255) / 256, 256, smem, s>>>(p);
    } else {
        preprocess_bwd_kernel<false><<<(end - first + 255) / 256, 256, p.push_bases ? 256 * 11 * (int)sizeof(float) : 0, s>>>(p);
    }
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// data-parallel helper: dL_dsh summed over views from the per-view clamped colour gradients (gsr_sh_grad_from_views).
// Per view the SH gradient of a Gaussian is rank 1: dL_dsh[k][c] = basis_k(dir) * g[c]  (backward.cu:20-139; the same
// w[] as preprocess_bwd_kernel).  Exchanging g (3 floats) instead of the row (3M floats) cuts the gradient payload of a
// view-parallel step from 59 to 14 floats per Gaussian; this kernel rebuilds the sum.  Rows leave through shared
// memory as whole 6 KB warp segments (same staging as the backward).
// ------------------------------------------------------------------------------------------------
// PACKED: g_all holds, per view, only the rows of the Gaussians whose bit is set in that view's mask ([K][cap][3], index
// order); meta holds per view `meta_stride` words: mask words [W32], then the number of set bits before each word [W32].
// One half of a Gaussian's row: coefficients [8 HALF, 8 HALF + 8) summed over the views into 24 registers and parked in the
// staged row.  Two passes of 24 accumulators instead of one of 48 keep the kernel at 64 registers = 4 CTAs per SM (the 48-wide
// version sat at 89 registers, 2 CTAs, 23 % of the warp slots, 32 % of DRAM peak: r2 ncu); a pass re-reads the view's mask word
// and 12-byte row, which hit L1.  Every accumulator still adds the views in index order, so the sums keep their bits.
template <bool PACKED, int HALF>
__device__ __forceinline__ void sh_views_half(int idx, int P, int D, int ncoef, float mx, float my, float mz, int K,
                                              const float* __restrict__ cams, const float* __restrict__ g_all,
                                              const uint32_t* __restrict__ meta, int meta_stride, int W32, int cap, float* row, int L) {
    float acc[24];
#pragma unroll
    for (int i = 0; i < 24; ++i) acc[i] = 0.f;
    for (int v = 0; v < K; ++v) {
        const float* g;
        if (PACKED) {
            const uint32_t* mv = meta + (size_t)v * meta_stride;
            const uint32_t word = __ldg(mv + (idx >> 5));
            if (!((word >> (idx & 31)) & 1u)) continue;
            const uint32_t slot = __ldg(mv + W32 + (idx >> 5)) + __popc(word & ((1u << (idx & 31)) - 1u));
            if (slot >= (uint32_t)cap) continue;                    // cannot happen with a consistent pack; never read out of bounds
            g = g_all + ((size_t)v * cap + slot) * 3;
        } else {
            g = g_all + ((size_t)v * P + idx) * 3;
        }
        const float g0 = __ldg(g), g1 = __ldg(g + 1), g2 = __ldg(g + 2);
        if (g0 == 0.f && g1 == 0.f && g2 == 0.f) continue;          // culled or fully clamped in this view
        const float ox = mx - __ldg(cams + 3 * v), oy = my - __ldg(cams + 3 * v + 1), oz = mz - __ldg(cams + 3 * v + 2);
        const float inv_len = 1.f / sqrtf(ox * ox + oy * oy + oz * oz);
        const float x = ox * inv_len, y = oy * inv_len, z = oz * inv_len;
        float w[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) w[k] = 0.f;
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        if (HALF == 0) {
            w[0] = kSH_C0;
            if (D > 0) {
                w[1] = -kSH_C1 * y; w[2] = kSH_C1 * z; w[3] = -kSH_C1 * x;
                if (D > 1) { w[4] = kSH_C2_0 * xy; w[5] = kSH_C2_1 * yz; w[6] = kSH_C2_2 * (2.f * zz - xx - yy); w[7] = kSH_C2_3 * xz; }
            }
        } else {
            if (D > 1) {
                w[0] = kSH_C2_4 * (xx - yy);
                if (D > 2) {
                    w[1] = kSH_C3_0 * y * (3.f * xx - yy);
                    w[2] = kSH_C3_1 * xy * z;
                    w[3] = kSH_C3_2 * y * (4.f * zz - xx - yy);
                    w[4] = kSH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
                    w[5] = kSH_C3_4 * x * (4.f * zz - xx - yy);
                    w[6] = kSH_C3_5 * z * (xx - yy);
                    w[7] = kSH_C3_6 * x * (xx - 3.f * yy);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float wk = (8 * HALF + k) < ncoef ? w[k] : 0.f;
            // product first, then the sum over views: one view reproduces preprocess_bwd_kernel's rows bit for bit
            acc[3 * k] += wk * g0; acc[3 * k + 1] += wk * g1; acc[3 * k + 2] += wk * g2;
        }
    }
#pragma unroll
    for (int i = 0; i < 24; ++i) if (24 * HALF + i < L) row[24 * HALF + i] = acc[i];
}

template <bool PACKED>
__global__ void __launch_bounds__(256, 4) sh_grad_from_views_kernel(int P, int D, int M, const float* __restrict__ means3D, int K,
                                                                    const float* __restrict__ cams, const float* __restrict__ g_all,
                                                                    const uint32_t* __restrict__ meta, int meta_stride, int W32, int cap,
                                                                    float* __restrict__ dL_dsh) {
    extern __shared__ float s_rows[];            // [256][kShStride]
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int idx = blockIdx.x * 256 + tid;
    const bool valid = idx < P;
    const int ncoef = min(16, min(M, (D + 1) * (D + 1)));
    const int L = 3 * M;
    float* row = s_rows + tid * kShStride;
    if (valid) {
        const float mx = __ldg(means3D + 3 * (size_t)idx), my = __ldg(means3D + 3 * (size_t)idx + 1), mz = __ldg(means3D + 3 * (size_t)idx + 2);
        sh_views_half<PACKED, 0>(idx, P, D, ncoef, mx, my, mz, K, cams, g_all, meta, meta_stride, W32, cap, row, L);
        if (L > 24) sh_views_half<PACKED, 1>(idx, P, D, ncoef, mx, my, mz, K, cams, g_all, meta, meta_stride, W32, cap, row, L);
    }
    __syncwarp();
    const int warp_idx0 = blockIdx.x * 256 + wid * 32;
    const int nrows = min(32, P - warp_idx0);
    if (nrows <= 0) return;
    const int f4_per_row = L >> 2;
    const float* wbase = s_rows + wid * 32 * kShStride;
    float4* gout = reinterpret_cast<float4*>(dL_dsh + (size_t)warp_idx0 * L);
    for (int e = lane; e < nrows * f4_per_row; e += 32) {
        const int r = e / f4_per_row, c = (e - r * f4_per_row) * 4;
        const float* src = wbase + r * kShStride + c;
        gout[e] = make_float4(src[0], src[1], src[2], src[3]);
    }
}

int launch_sh_grad_from_views(int P, int D, int M, const float* means3D, int K, const float* cams, const float* g_all,
                              const uint32_t* meta, int meta_stride, int cap, float* dL_dsh, cudaStream_t s) {
    if (P <= 0) return GSR_OK;
    if (M <= 0 || M > 16 || (M % 4) != 0 || ((size_t)dL_dsh & 15) != 0) {
        set_error("sh_grad_from_views: needs 0 < M <= 16, M %% 4 == 0 and a 16-byte aligned output (M=%d)", M); return GSR_ERR_INVALID;
    }
    const int smem = 256 * kShStride * (int)sizeof(float);
    static bool attr_done = false;
    if (!attr_done) {
        cudaFuncSetAttribute(sh_grad_from_views_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        cudaFuncSetAttribute(sh_grad_from_views_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        attr_done = true;
    }
    if (meta) sh_grad_from_views_kernel<true><<<(P + 255) / 256, 256, smem, s>>>(P, D, M, means3D, K, cams, g_all, meta, meta_stride, (P + 31) / 32, cap, dL_dsh);
    else sh_grad_from_views_kernel<false><<<(P + 255) / 256, 256, smem, s>>>(P, D, M, means3D, K, cams, g_all, nullptr, 0, 0, 0, dL_dsh);
    count_launch();
    return GSR_OK;
}

// ------------------------------------------------------------------------------------------------
// gsr_pack_visible_rows: rows of src [P,3] of the Gaussians with radii > 0, in index order, plus the bit mask and the
// per-word exclusive popcount prefix that let a reader find a Gaussian's row (see sh_grad_from_views_kernel<true>).
// Three small launches: per-1024 counts + mask words, one-CTA scan of the counts, scatter.
// ------------------------------------------------------------------------------------------------
constexpr int kPackBlock = 1024;

__global__ void __launch_bounds__(kPackBlock) pack_mask_kernel(int P, const int* __restrict__ radii, uint32_t* __restrict__ mask,
                                                                uint32_t* __restrict__ prefix, uint32_t* __restrict__ block_counts) {
    __shared__ uint32_t s_cnt[kPackBlock / 32];
    const int idx = blockIdx.x * kPackBlock + threadIdx.x, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const bool vis = idx < P && radii[idx] > 0;
    const uint32_t word = __ballot_sync(0xffffffffu, vis);
    if (lane == 0) s_cnt[wid] = __popc(word);
    __syncthreads();
    if (wid == 0) {
        const uint32_t c = s_cnt[lane];
        uint32_t inc = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
        s_cnt[lane] = inc - c;                                    // exclusive, local to the block
        if (lane == 31) block_counts[blockIdx.x] = inc;
    }
    __syncthreads();
    if (lane == 0 && (idx >> 5) < (P + 31) / 32) { mask[idx >> 5] = word; prefix[idx >> 5] = s_cnt[wid]; }
}

__global__ void __launch_bounds__(1024) pack_scan_kernel(uint32_t* __restrict__ block_counts, int n) {
    // in-place exclusive scan of n block counts by one CTA (n = P/1024: a few thousand)
    __shared__ uint32_t s_w[32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const uint32_t v = i < n ? block_counts[i] : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
        if (lane == 31) s_w[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            const uint32_t w = s_w[lane];
            uint32_t winc = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, winc, d); if (lane >= d) winc += t; }
            s_w[lane] = winc - w;
        }
        __syncthreads();
        const uint32_t carry = s_carry;
        if (i < n) block_counts[i] = carry + s_w[wid] + inc - v;
        __syncthreads();
        if (tid == 1023) s_carry = carry + s_w[31] + inc;
        __syncthreads();
    }
    if (tid == 0) block_counts[n] = s_carry;                      // total, for the caller's consistency check
}

__global__ void __launch_bounds__(kPackBlock) pack_scatter_kernel(int P, const float* __restrict__ src, const uint32_t* __restrict__ mask,
                                                                   uint32_t* __restrict__ prefix, const uint32_t* __restrict__ block_offsets,
                                                                   int cap, float* __restrict__ rows) {
    const int idx = blockIdx.x * kPackBlock + threadIdx.x, lane = threadIdx.x & 31;
    if (idx >= ((P + 31) / 32) * 32) return;                      // whole warps stay: the word's lane 0 finalises the prefix
    const uint32_t word = mask[idx >> 5];
    const uint32_t base = block_offsets[blockIdx.x] + prefix[idx >> 5];
    __syncwarp();
    if (lane == 0) prefix[idx >> 5] = base;                       // now global: set bits before this word
    if (idx < P && ((word >> lane) & 1u)) {
        const uint32_t slot = base + __popc(word & ((1u << lane) - 1u));
        if (slot < (uint32_t)cap) {
            const size_t i3 = 3 * (size_t)idx, o3 = 3 * (size_t)slot;
            rows[o3] = src[i3]; rows[o3 + 1] = src[i3 + 1]; rows[o3 + 2] = src[i3 + 2];
        }
    }
}

size_t pack_temp_bytes(int P) { return sizeof(uint32_t) * ((size_t)(P + kPackBlock - 1) / kPackBlock + 1); }

int launch_pack_visible_rows(int P, const int* radii, const float* src, int cap, float* rows, uint32_t* mask, uint32_t* prefix,
                             uint32_t* temp, cudaStream_t s) {
    if (P <= 0) return GSR_OK;
    const int nb = (P + kPackBlock - 1) / kPackBlock;
    pack_mask_kernel<<<nb, kPackBlock, 0, s>>>(P, radii, mask, prefix, temp);
    pack_scan_kernel<<<1, 1024, 0, s>>>(temp, nb);
    pack_scatter_kernel<<<nb, kPackBlock, 0, s>>>(P, src, mask, prefix, temp, cap, rows);
    count_launch(); count_launch(); count_launch();
    return GSR_OK;
}

// clamped colour gradient right after the blend stage (gsr_backward_phase GSR_BWD_BLEND): scratch rows hold dL_dcolor
// in (1.z, 1.w, 2.x); the SH clamp mask zeroes channels (backward.cu:56-58); culled Gaussians get zero rows.
__global__ void __launch_bounds__(256) clamped_rgb_kernel(int P, const int* __restrict__ radii, const float* __restrict__ scratch,
                                                          const unsigned char* __restrict__ clamped, int colors_from_sh,
                                                          float* __restrict__ out) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= P) return;
    const size_t i3 = 3 * (size_t)idx;
    float g0 = 0.f, g1 = 0.f, g2 = 0.f;
    if (colors_from_sh && radii[idx] > 0) {
        const float4* sp = reinterpret_cast<const float4*>(scratch) + 3 * (size_t)idx;
        const float4 sB = sp[1], sC = sp[2];
        g0 = sB.z * (clamped[i3] ? 0.f : 1.f); g1 = sB.w * (clamped[i3 + 1] ? 0.f : 1.f); g2 = sC.x * (clamped[i3 + 2] ? 0.f : 1.f);
    }
    out[i3] = g0; out[i3 + 1] = g1; out[i3 + 2] = g2;
}

void launch_clamped_rgb(int P, const int* radii, const float* scratch, const unsigned char* clamped, bool colors_from_sh, float* out, cudaStream_t s) {
    if (P <= 0) return;
    clamped_rgb_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, radii, scratch, clamped, colors_from_sh ? 1 : 0, out);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// gsr_pack_visible_rows_noalloc: gsr_pack_visible_rows with its scratch (the per-block counts / offsets of the scan) in a
// file-scope device array instead of caller memory, for callers that cannot allocate (CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The array costs kPackStaticBytes of device memory on every GPU that loads the library, until exit.  All calls on a device
// share it: two calls that overlap on the GPU (different streams) race.  Nothing in it has to start at zero -- every call writes
// all the words it reads (pack_mask_kernel writes counts [0, nb), pack_scan_kernel rewrites them and writes word nb) -- so the
// previous call's contents are harmless.
// ------------------------------------------------------------------------------------------------
constexpr int kPackStaticMaxP = 1 << 24;                                       // 16 777 216 Gaussians
constexpr int kPackStaticWords = kPackStaticMaxP / kPackBlock + 64;            // nb + 1 words are used; 65 792 bytes
static __device__ __align__(256) uint32_t g_pack_scratch[kPackStaticWords];

// The array's name is a host-side stand-in in host code, not a device address: resolve it once per device, with that device
// current (cudaGetSymbolAddress), and hand the kernels the resolved pointer.
static int pack_scratch_address(uint32_t** out) {
    static uint32_t* addr[64] = {};
    static std::mutex mu;
    int dev = 0;
    int rc = check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
    if (rc) return rc;
    if (dev < 0 || dev >= 64) { set_error("pack_visible_rows_noalloc: device index %d not supported", dev); return GSR_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        uint32_t* ptr = nullptr;
        if ((rc = check_cuda(cudaGetSymbolAddress((void**)&ptr, g_pack_scratch), "cudaGetSymbolAddress(g_pack_scratch)"))) return rc;
        if (!ptr) { set_error("pack_visible_rows_noalloc: no device address for the scratch array"); return GSR_ERR_CUDA; }
        addr[dev] = ptr;
    }
    *out = addr[dev];
    return GSR_OK;
}

}  // namespace gsr

extern "C" int gsr_pack_visible_rows_noalloc(int P, const int* radii, const float* rows_in, int cap, float* rows_out, uint32_t* meta,
                                             void* stream) {
    using namespace gsr;
    if (P < 0 || cap < 0) { set_error("pack_visible_rows_noalloc: bad sizes"); return GSR_ERR_INVALID; }
    if (P > kPackStaticMaxP) {
        set_error("pack_visible_rows_noalloc: P = %d exceeds the static scratch (at most %d Gaussians); use gsr_pack_visible_rows with caller scratch",
                  P, kPackStaticMaxP);
        return GSR_ERR_INVALID;
    }
    if (P == 0) return GSR_OK;
    if (!radii || !rows_in || !meta || (cap > 0 && !rows_out)) { set_error("pack_visible_rows_noalloc: null pointer"); return GSR_ERR_INVALID; }
    uint32_t* temp = nullptr;
    int rc = pack_scratch_address(&temp);
    if (rc) return rc;
    const int W32 = (P + 31) / 32;
    rc = launch_pack_visible_rows(P, radii, rows_in, cap, rows_out, meta, meta + W32, temp, (cudaStream_t)stream);
    if (rc) return rc;
    return check_last("pack_visible_rows_noalloc", false, (cudaStream_t)stream);
}

extern "C" int gsr_pack_noalloc_max_gaussians(void) { return gsr::kPackStaticMaxP; }
