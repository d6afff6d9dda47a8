// gsr_math.cuh -- device arithmetic helpers with PINNED rounding.
//
// Tile keys embed raw depth bits and tile rects depend on ceil(3*sqrt(lambda)), so integer parity with
// the reference requires bit-identical FP32 results in preprocess.  nvcc's default -fmad=true contracts
// the reference's expressions in a specific, deterministic way; that contraction was read off the
// reference's sm_100a SASS (preprocessCUDA / renderCUDA in cuda_rasterizer/forward.cu) and is pinned
// here with explicit round-to-nearest intrinsics, which the compiler never re-associates or fuses.
//
// The one rule that covers almost everything: a three-term sum of products written
//     a0*a1 + b0*b1 + c0*c1            (glm mat3*mat3, type_mat3x3.inl:509-517; auxiliary.h:58-77)
// is evaluated as   fma(c0,c1, fma(a0,a1, b0*b1)).
#pragma once
#include <cuda_runtime.h>

namespace gsr {

__device__ __forceinline__ float sum3(float a0, float a1, float b0, float b1, float c0, float c1) {
    return __fmaf_rn(c0, c1, __fmaf_rn(a0, a1, __fmul_rn(b0, b1)));
}

// one row of transformPoint4x3 / 4x4 (auxiliary.h:58-77): m0*x + m4*y + m8*z + m12
__device__ __forceinline__ float xform_row(float m0, float m4, float m8, float m12, float x, float y, float z) {
    return __fadd_rn(sum3(x, m0, y, m4, z, m8), m12);
}

// ndc2Pix (auxiliary.h:41-44): evaluated in double, (v+1.0)*S-1.0 is one DFMA, result rounded to float
__device__ __forceinline__ float ndc2pix(float v, int S) {
    double t = __fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0);
    return __double2float_rn(__dmul_rn(t, 0.5));
}

// SH basis constants (auxiliary.h:22-39)
constexpr float kSH_C0 = 0.28209479177387814f;
constexpr float kSH_C1 = 0.4886025119029199f;
constexpr float kSH_C2_0 = 1.0925484305920792f;
constexpr float kSH_C2_1 = -1.0925484305920792f;
constexpr float kSH_C2_2 = 0.31539156525252005f;
constexpr float kSH_C2_3 = -1.0925484305920792f;
constexpr float kSH_C2_4 = 0.5462742152960396f;
constexpr float kSH_C3_0 = -0.5900435899266435f;
constexpr float kSH_C3_1 = 2.890611442640554f;
constexpr float kSH_C3_2 = -0.4570457994644658f;
constexpr float kSH_C3_3 = 0.3731763325901154f;
constexpr float kSH_C3_4 = -0.4570457994644658f;
constexpr float kSH_C3_5 = 1.445305721320277f;
constexpr float kSH_C3_6 = -0.5900435899266435f;

// tile rect (auxiliary.h:46-56).  TILE is 16 (baseline) or 32 (AMR); division by a power of two is an
// exact multiply, (p + r + TILE - 1) is three separate float adds in that order.
__device__ __forceinline__ void get_rect(float px, float py, int radius, int tile, int grid_x, int grid_y,
                                         uint32_t& x0, uint32_t& y0, uint32_t& x1, uint32_t& y1) {
    const float r = (float)radius;
    const float inv = 1.0f / (float)tile;
    const float t = (float)tile;
    int ax = (int)__fmul_rn(__fadd_rn(px, -r), inv);
    int ay = (int)__fmul_rn(__fadd_rn(py, -r), inv);
    int bx = (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(px, r), t), -1.0f), inv);
    int by = (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(py, r), t), -1.0f), inv);
    x0 = min((uint32_t)grid_x, (uint32_t)max(0, ax));
    y0 = min((uint32_t)grid_y, (uint32_t)max(0, ay));
    x1 = min((uint32_t)grid_x, (uint32_t)max(0, bx));
    y1 = min((uint32_t)grid_y, (uint32_t)max(0, by));
}

// screen-tile band (multi-GPU tile bands): keep only rows [row0,row1) of a rect; row1 == 0 means "no band"
__device__ __forceinline__ void clip_band(uint32_t& y0, uint32_t& y1, int row0, int row1) {
    if (row1 > 0) {
        y0 = min(max(y0, (uint32_t)row0), (uint32_t)row1);
        y1 = min(max(y1, (uint32_t)row0), (uint32_t)row1);
    }
}

// the blend exponent (forward.cu:334): -0.5f*(cx*dx*dx + cz*dy*dy) - cy*dx*dy, contraction as compiled
__device__ __forceinline__ float blend_power(float cx, float cy, float cz, float dx, float dy) {
    float s = __fmaf_rn(dx, __fmul_rn(dx, cx), __fmul_rn(dy, __fmul_rn(dy, cz)));
    float m = __fmul_rn(dy, __fmul_rn(dx, cy));
    return __fmaf_rn(s, -0.5f, -m);
}

}  // namespace gsr
