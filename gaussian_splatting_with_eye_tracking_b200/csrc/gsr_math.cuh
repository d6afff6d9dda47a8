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

// activations of the raw-parameter entry (gsr_scene.raw_params; scene/gaussian_model.py:33-41): the expressions torch evaluates
__device__ __forceinline__ float act_sigmoid(float x) { return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x))); }
__device__ __forceinline__ float act_exp(float x) { return expf(x); }
__device__ __forceinline__ float4 act_normalize(float4 q) {      // F.normalize: x / max(||x||_2, 1e-12)
    const float n = __fsqrt_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(q.x, q.x), __fmul_rn(q.y, q.y)), __fmul_rn(q.z, q.z)), __fmul_rn(q.w, q.w)));
    const float d = fmaxf(n, 1e-12f);
    return make_float4(__fdiv_rn(q.x, d), __fdiv_rn(q.y, d), __fdiv_rn(q.z, d), __fdiv_rn(q.w, d));
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

// ---- packed FP32x2 (sm_100a FADD2 / FMUL2 / FFMA2): two IEEE round-to-nearest operations per issued instruction.
// Each half is bit-identical to the scalar __f*_rn intrinsic; pk(v, v) compiles to a broadcast operand (no move).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// blend_power for two horizontally adjacent pixels (dx differs, dy shared): same operations, same order, per half.
// NOTE: every mul2 below feeds an fma2 as a multiplicand or as the addend of a DIFFERENT product, so ptxas has no
// mul+add pair to contract (it does contract mul.rn.f32x2 + add.rn.f32x2, unlike the scalar .rn forms).
__device__ __forceinline__ f32x2 blend_power2(float cx, float cy, float cz, f32x2 dx2, float dy) {
    const float b = __fmul_rn(dy, __fmul_rn(dy, cz));
    const f32x2 s2 = fma2(dx2, mul2(dx2, pk(cx, cx)), pk(b, b));
    const f32x2 m2n = mul2(pk(-dy, -dy), mul2(dx2, pk(cy, cy)));     // -(dy * (dx * cy)), negation is exact
    return fma2(s2, pk(-0.5f, -0.5f), m2n);
}

}  // namespace gsr
