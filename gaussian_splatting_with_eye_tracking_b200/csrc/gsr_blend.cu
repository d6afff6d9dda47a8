// gsr_blend.cu -- per-tile front-to-back alpha blending for sm_100a, forward / AMR-forward / backward.
//
//   blend_fwd_kernel   replaces renderCUDA<3> forward (cuda_rasterizer/forward.cu:261-374)
//   blend_amr_kernel   replaces the AMR fork's renderCUDA<3> (A/cuda_rasterizer/forward.cu:261-518): ONE CTA per
//                      32x32 binning tile walks the tile's list once and blends every active 2x2-lattice round
//                      from the same staged batch (the reference launches 4 CTAs per tile that each re-stage it)
//   interpolate_kernel replaces interpolateCUDA<3> (A/cuda_rasterizer/forward.cu:520-648)
//   blend_bwd_kernel   replaces renderCUDA<3> backward (cuda_rasterizer/backward.cu:399-557): per-warp butterfly
//                      reduction of the 9 per-Gaussian gradient terms, then ONE packed atomic set per
//                      (warp, Gaussian) instead of 9 atomics per (pixel, Gaussian)
//
// Common structure: a batch of 256 list entries is gathered (id -> xy, conic, opacity, rgb) by the 256 threads of
// the CTA, each loader thread runs an exact-preserving tile-level cull (a Gaussian whose best-case alpha over the
// whole tile rectangle is below 1/255 can never pass the per-pixel test at forward.cu:343), survivors are
// compacted into shared memory in list order together with their 1-based list position, and the next batch's
// gathers are issued before the current batch is blended (register double buffering).  The per-pixel arithmetic
// (power, expf, alpha, T test, accumulation order) is the reference's, operation for operation.
#include "gsr_common.cuh"
#include "gsr_math.cuh"

namespace gsr {

constexpr int kBlendThreads = 256;
constexpr float kAlphaMin = 1.0f / 255.0f;

struct GaussRec {        // what a loader thread holds for one list entry
    float2 xy; float4 co; float3 col; uint32_t id;
};

// Exact-preserving cull: true  => for EVERY pixel centre in [x0,x1]x[y0,y1] the reference's per-pixel test yields
// power > 0 or alpha < 1/255, so the entry cannot change any pixel (it only advances `contributor`).
__device__ __forceinline__ bool cannot_touch(const float2 xy, const float4 co, float x0, float y0, float x1, float y1) {
    const float a = co.x, b = co.y, c = co.z, o = co.w;
    if (!(o >= kAlphaMin)) return true;                 // alpha = min(.99, o*exp(power<=0)) <= o
    if (!(a > 0.f && c > 0.f && a * c - b * b > 0.f)) return false;   // not positive definite: do not reason about it
    const float tau = __logf(255.0f * o) * 1.0001f + 0.02f;            // q > tau  <=>  o*exp(-q) < 1/255 (with margin)
    const float cx = fminf(fmaxf(xy.x, x0), x1), cy = fminf(fmaxf(xy.y, y0), y1);
    const float ex = xy.x - cx, ey = xy.y - cy;        // offset of the mean from the closest rect point
    if (ex == 0.f && ey == 0.f) return false;           // mean inside the tile
    // d = mean - pixel; over the rect d.x in [xy.x - x1, xy.x - x0], d.y likewise
    const float dxlo = xy.x - x1, dxhi = xy.x - x0, dylo = xy.y - y1, dyhi = xy.y - y0;
    float qmin = 3.0e38f;
    if (ex != 0.f) {   // facing vertical edge: d.x fixed = ex, minimise over d.y
        const float dy = fminf(fmaxf(-b * ex / c, dylo), dyhi);
        const float t0 = 0.5f * a * ex * ex, t1 = 0.5f * c * dy * dy, t2 = b * ex * dy;
        qmin = fminf(qmin, (t0 + t1 + t2) - 1e-5f * (t0 + t1 + fabsf(t2)));
    }
    if (ey != 0.f) {   // facing horizontal edge
        const float dx = fminf(fmaxf(-b * ey / a, dxlo), dxhi);
        const float t0 = 0.5f * a * dx * dx, t1 = 0.5f * c * ey * ey, t2 = b * dx * ey;
        qmin = fminf(qmin, (t0 + t1 + t2) - 1e-5f * (t0 + t1 + fabsf(t2)));
    }
    return qmin > tau;
}

// Per-pixel early reject, exact-preserving: power < reject_below(o)  =>  o * expf(power) < (1/255) * 0.98, so the
// reference's `alpha < 1/255` test (forward.cu:343) is certain to skip the pair and expf need not be evaluated.
// Same margin as cannot_touch; only called for o >= 1/255.
__device__ __forceinline__ float reject_below(float o) { return -(__logf(255.0f * o) * 1.0001f + 0.02f); }

__device__ __forceinline__ GaussRec gather(const uint32_t* __restrict__ point_list, const float2* __restrict__ means2D,
                                           const float4* __restrict__ conic_opacity, const float* __restrict__ colors, uint32_t li) {
    GaussRec r;
    r.id = __ldg(point_list + li);
    r.xy = __ldg(means2D + r.id);
    r.co = __ldg(conic_opacity + r.id);
    r.col = make_float3(__ldg(colors + 3 * (size_t)r.id), __ldg(colors + 3 * (size_t)r.id + 1), __ldg(colors + 3 * (size_t)r.id + 2));
    return r;
}

// CTA-wide stable compaction of `keep` flags; returns slot (or -1) and sets s_count.
template <int NTHREADS = kBlendThreads>
__device__ __forceinline__ int compact_slot(bool keep, uint32_t* s_wcnt, uint32_t* s_count) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t m = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_wcnt[wid] = __popc(m);
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < NTHREADS / 32; ++w) { const uint32_t c = s_wcnt[w]; base += (w < wid) ? c : 0u; tot += c; }
    if (threadIdx.x == 0) *s_count = tot;
    return keep ? (int)(base + __popc(m & ((1u << lane) - 1u))) : -1;
}

// ------------------------------------------------------------------------------------------------
// forward, baseline 16x16 tiles
// ------------------------------------------------------------------------------------------------
struct FwdParams {
    int W, H, tiles_x, row0;      // row0: first tile row of the launch (tile bands); blockIdx.y counts from it
    const uint2* __restrict__ ranges; const uint32_t* __restrict__ point_list;
    const float2* __restrict__ means2D; const float4* __restrict__ conic_opacity; const float* __restrict__ colors;
    const float* __restrict__ bg;
    float* __restrict__ final_T; uint32_t* __restrict__ n_contrib; float* __restrict__ out_color;
};

__global__ void __launch_bounds__(kBlendThreads) blend_fwd_kernel(const FwdParams p) {
    // staged entry = 3 x float4: (x, y, conic.x, conic.y) (conic.z, reject threshold, opacity, 1-based list position)
    // (r, g, b, -): one base pointer, immediate offsets, 128-bit broadcast loads; the colour vector is only read by
    // pixels the entry contributes to
    __shared__ float4 s_ent[3 * kBlendThreads];
    __shared__ uint32_t s_wcnt[kBlendThreads / 32];
    __shared__ uint32_t s_count;

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    // warp = 8x4 pixel block (results do not depend on the thread->pixel map)
    const uint32_t lx = (wid & 1) * 8 + (lane & 7), ly = (wid >> 1) * 4 + (lane >> 3);
    const uint32_t tile_y = blockIdx.y + (uint32_t)p.row0;
    const uint32_t pix_x = blockIdx.x * 16 + lx, pix_y = tile_y * 16 + ly;
    const bool inside = pix_x < (uint32_t)p.W && pix_y < (uint32_t)p.H;
    const float pxf = (float)pix_x, pyf = (float)pix_y;
    const float rx0 = (float)(blockIdx.x * 16), ry0 = (float)(tile_y * 16);
    const float rx1 = rx0 + 15.f, ry1 = ry0 + 15.f;

    const uint2 range = p.ranges[tile_y * p.tiles_x + blockIdx.x];
    const int todo = (int)(range.y - range.x);
    const int rounds = (todo + kBlendThreads - 1) / kBlendThreads;

    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
    uint32_t last_contributor = 0;
    bool done = !inside;

    GaussRec nxt; bool nxt_valid = false;
    if (tid < todo) { nxt = gather(p.point_list, p.means2D, p.conic_opacity, p.colors, range.x + tid); nxt_valid = true; }

    for (int r = 0; r < rounds; ++r) {
        // block vote doubles as the "everyone left the previous batch" barrier (forward.cu:309)
        if (__syncthreads_count(done) == kBlendThreads) break;
        const bool keep = nxt_valid && !cannot_touch(nxt.xy, nxt.co, rx0, ry0, rx1, ry1);
        const int slot = compact_slot(keep, s_wcnt, &s_count);
        if (slot >= 0) {
            s_ent[3 * slot] = make_float4(nxt.xy.x, nxt.xy.y, nxt.co.x, nxt.co.y);
            s_ent[3 * slot + 1] = make_float4(nxt.co.z, reject_below(nxt.co.w), nxt.co.w, __uint_as_float((uint32_t)(r * kBlendThreads + tid + 1)));
            s_ent[3 * slot + 2] = make_float4(nxt.col.x, nxt.col.y, nxt.col.z, 0.f);
        }
        __syncthreads();
        const int count = (int)s_count;
        // issue the next batch's gathers; they complete while this batch is blended
        const int nprog = (r + 1) * kBlendThreads + tid;
        nxt_valid = nprog < todo;
        if (nxt_valid) nxt = gather(p.point_list, p.means2D, p.conic_opacity, p.colors, range.x + nprog);

        if (!done) {
            const float4* e = s_ent;
            int last_j = -1;
            for (int j = 0; j < count; ++j, e += 3) {
                const float4 A = e[0], B = e[1];
                const float dx = __fadd_rn(A.x, -pxf), dy = __fadd_rn(A.y, -pyf);
                const float power = blend_power(A.z, A.w, B.x, dx, dy);
                if (power < B.y || power > 0.0f) continue;       // certain alpha < 1/255 (reject_below) | forward.cu:337
                const float alpha = fminf(0.99f, __fmul_rn(B.z, expf(power)));
                if (alpha < kAlphaMin) continue;
                const float test_T = __fmul_rn(T, __fadd_rn(1.f, -alpha));
                if (test_T < 0.0001f) { done = true; break; }
                const float4 Cc = e[2];
                C0 = __fmaf_rn(T, __fmul_rn(alpha, Cc.x), C0);
                C1 = __fmaf_rn(T, __fmul_rn(alpha, Cc.y), C1);
                C2 = __fmaf_rn(T, __fmul_rn(alpha, Cc.z), C2);
                T = test_T;
                last_j = j;
            }
            if (last_j >= 0) last_contributor = __float_as_uint(s_ent[3 * last_j + 1].w);
        }
    }
    if (inside) {
        const uint32_t pix_id = (uint32_t)p.W * pix_y + pix_x;
        const size_t HW = (size_t)p.H * p.W;
        p.final_T[pix_id] = T;
        p.n_contrib[pix_id] = last_contributor;
        p.out_color[pix_id] = __fmaf_rn(T, __ldg(p.bg), C0);
        p.out_color[HW + pix_id] = __fmaf_rn(T, __ldg(p.bg + 1), C1);
        p.out_color[2 * HW + pix_id] = __fmaf_rn(T, __ldg(p.bg + 2), C2);
    }
}

// ------------------------------------------------------------------------------------------------
// forward, two pixels per thread with packed FP32x2 arithmetic (default).  The kernel above is bound by instruction
// issue (ncu: 79 % issue-active, FMA pipe 49 %); here one FADD2/FMUL2/FFMA2 does the exponent arithmetic of two
// horizontally adjacent pixels, and one pair of 128-bit shared loads feeds both.  128 threads per 16x16 tile,
// warp = 8x8 pixels.  Per-pixel results are bit-identical (every half is an IEEE round-to-nearest operation).
// ------------------------------------------------------------------------------------------------
constexpr int kBlend2Threads = 128;

__global__ void __launch_bounds__(kBlend2Threads) blend_fwd2_kernel(const FwdParams p) {
    __shared__ float4 s_ent[3 * kBlend2Threads];
    __shared__ uint32_t s_wcnt[kBlend2Threads / 32];
    __shared__ uint32_t s_count;

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t lx = (wid & 1) * 8 + (lane & 3) * 2, ly = (wid >> 1) * 8 + (lane >> 2);
    const uint32_t tile_y = blockIdx.y + (uint32_t)p.row0;
    const uint32_t pix_x = blockIdx.x * 16 + lx, pix_y = tile_y * 16 + ly;
    const bool inside0 = pix_x < (uint32_t)p.W && pix_y < (uint32_t)p.H;
    const bool inside1 = pix_x + 1 < (uint32_t)p.W && pix_y < (uint32_t)p.H;
    const float pyf = (float)pix_y;
    const f32x2 npx2 = pk(-(float)pix_x, -(float)(pix_x + 1));
    const float rx0 = (float)(blockIdx.x * 16), ry0 = (float)(tile_y * 16);
    const float rx1 = rx0 + 15.f, ry1 = ry0 + 15.f;

    const uint2 range = p.ranges[tile_y * p.tiles_x + blockIdx.x];
    const int todo = (int)(range.y - range.x);
    const int rounds = (todo + kBlend2Threads - 1) / kBlend2Threads;

    float T0 = 1.0f, T1 = 1.0f, Ca0 = 0.f, Ca1 = 0.f, Ca2 = 0.f, Cb0 = 0.f, Cb1 = 0.f, Cb2 = 0.f;
    uint32_t last0 = 0, last1 = 0;
    bool done0 = !inside0, done1 = !inside1;

    GaussRec nxt; bool nxt_valid = false;
    if (tid < todo) { nxt = gather(p.point_list, p.means2D, p.conic_opacity, p.colors, range.x + tid); nxt_valid = true; }

    for (int r = 0; r < rounds; ++r) {
        if (__syncthreads_count(done0 && done1) == kBlend2Threads) break;
        const bool keep = nxt_valid && !cannot_touch(nxt.xy, nxt.co, rx0, ry0, rx1, ry1);
        const int slot = compact_slot<kBlend2Threads>(keep, s_wcnt, &s_count);
        if (slot >= 0) {
            s_ent[3 * slot] = make_float4(nxt.xy.x, nxt.xy.y, nxt.co.x, nxt.co.y);
            s_ent[3 * slot + 1] = make_float4(nxt.co.z, reject_below(nxt.co.w), nxt.co.w, __uint_as_float((uint32_t)(r * kBlend2Threads + tid + 1)));
            s_ent[3 * slot + 2] = make_float4(nxt.col.x, nxt.col.y, nxt.col.z, 0.f);
        }
        __syncthreads();
        const int count = (int)s_count;
        const int nprog = (r + 1) * kBlend2Threads + tid;
        nxt_valid = nprog < todo;
        if (nxt_valid) nxt = gather(p.point_list, p.means2D, p.conic_opacity, p.colors, range.x + nprog);

        if (!(done0 && done1)) {
            const float4* e = s_ent;
            int lj0 = -1, lj1 = -1;
            for (int j = 0; j < count; ++j, e += 3) {
                const float4 A = e[0], B = e[1];
                const float dy = __fadd_rn(A.y, -pyf);
                float pw0, pw1;
                upk(blend_power2(A.z, A.w, B.x, add2(pk(A.x, A.x), npx2), dy), pw0, pw1);
                // B.y: below it alpha < 1/255 is certain (reject_below); > 0: forward.cu:337
                const bool c0 = !done0 && !(pw0 < B.y) && !(pw0 > 0.0f);
                const bool c1 = !done1 && !(pw1 < B.y) && !(pw1 > 0.0f);
                if (!(c0 || c1)) continue;
                const float4 Cc = e[2];
                if (c0) {
                    const float alpha = fminf(0.99f, __fmul_rn(B.z, expf(pw0)));
                    if (!(alpha < kAlphaMin)) {
                        const float test_T = __fmul_rn(T0, __fadd_rn(1.f, -alpha));
                        if (test_T < 0.0001f) done0 = true;
                        else {
                            Ca0 = __fmaf_rn(T0, __fmul_rn(alpha, Cc.x), Ca0);
                            Ca1 = __fmaf_rn(T0, __fmul_rn(alpha, Cc.y), Ca1);
                            Ca2 = __fmaf_rn(T0, __fmul_rn(alpha, Cc.z), Ca2);
                            T0 = test_T; lj0 = j;
                        }
                    }
                }
                if (c1) {
                    const float alpha = fminf(0.99f, __fmul_rn(B.z, expf(pw1)));
                    if (!(alpha < kAlphaMin)) {
                        const float test_T = __fmul_rn(T1, __fadd_rn(1.f, -alpha));
                        if (test_T < 0.0001f) done1 = true;
                        else {
                            Cb0 = __fmaf_rn(T1, __fmul_rn(alpha, Cc.x), Cb0);
                            Cb1 = __fmaf_rn(T1, __fmul_rn(alpha, Cc.y), Cb1);
                            Cb2 = __fmaf_rn(T1, __fmul_rn(alpha, Cc.z), Cb2);
                            T1 = test_T; lj1 = j;
                        }
                    }
                }
                if (done0 && done1) break;
            }
            if (lj0 >= 0) last0 = __float_as_uint(s_ent[3 * lj0 + 1].w);
            if (lj1 >= 0) last1 = __float_as_uint(s_ent[3 * lj1 + 1].w);
        }
    }
    const size_t HW = (size_t)p.H * p.W;
    const float bg0 = __ldg(p.bg), bg1 = __ldg(p.bg + 1), bg2 = __ldg(p.bg + 2);
    const size_t pix_id = (size_t)p.W * pix_y + pix_x;
    if (inside0) {
        p.final_T[pix_id] = T0; p.n_contrib[pix_id] = last0;
        p.out_color[pix_id] = __fmaf_rn(T0, bg0, Ca0);
        p.out_color[HW + pix_id] = __fmaf_rn(T0, bg1, Ca1);
        p.out_color[2 * HW + pix_id] = __fmaf_rn(T0, bg2, Ca2);
    }
    if (inside1) {
        p.final_T[pix_id + 1] = T1; p.n_contrib[pix_id + 1] = last1;
        p.out_color[pix_id + 1] = __fmaf_rn(T1, bg0, Cb0);
        p.out_color[HW + pix_id + 1] = __fmaf_rn(T1, bg1, Cb1);
        p.out_color[2 * HW + pix_id + 1] = __fmaf_rn(T1, bg2, Cb2);
    }
}

// ------------------------------------------------------------------------------------------------
// measurement only: the ALGORITHMIC work of the blend stage ("pair-test" = one pixel x one list entry the
// reference's loop examines, forward.cu:328-361; "hit" = a pair-test that passes both skips and is accumulated).  One
// pixel per thread, no cull, no staging tricks: per pixel the number of entries examined before `done` (or the whole
// list), the number accumulated, and n_contrib (what the backward walks, backward.cu:486-488).  Never on the frame path.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) blend_count_kernel(const FwdParams p, unsigned long long* __restrict__ totals) {
    __shared__ float4 s_a[256];
    __shared__ float2 s_b[256];
    __shared__ unsigned long long s_tot[3];
    const int tid = threadIdx.x;
    if (tid < 3) s_tot[tid] = 0ull;
    const uint32_t tile_y = blockIdx.y + (uint32_t)p.row0;
    const uint32_t pix_x = blockIdx.x * 16 + (tid & 15), pix_y = tile_y * 16 + (tid >> 4);
    const bool inside = pix_x < (uint32_t)p.W && pix_y < (uint32_t)p.H;
    const float pxf = (float)pix_x, pyf = (float)pix_y;
    const uint2 range = p.ranges[tile_y * p.tiles_x + blockIdx.x];
    const int todo = (int)(range.y - range.x);
    float T = 1.0f;
    bool done = !inside;
    uint32_t examined = 0, hits = 0, last = 0;
    for (int base = 0; base < todo; base += 256) {
        if (__syncthreads_count(done) == 256) break;
        if (base + tid < todo) {
            const uint32_t id = __ldg(p.point_list + range.x + base + tid);
            const float2 xy = __ldg(p.means2D + id); const float4 co = __ldg(p.conic_opacity + id);
            s_a[tid] = make_float4(xy.x, xy.y, co.x, co.y); s_b[tid] = make_float2(co.z, co.w);
        }
        __syncthreads();
        const int cnt = min(256, todo - base);
        for (int j = 0; !done && j < cnt; ++j) {
            ++examined;
            const float4 A = s_a[j]; const float2 B = s_b[j];
            const float dx = __fadd_rn(A.x, -pxf), dy = __fadd_rn(A.y, -pyf);
            const float power = blend_power(A.z, A.w, B.x, dx, dy);
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, __fmul_rn(B.y, expf(power)));
            if (alpha < kAlphaMin) continue;
            const float test_T = __fmul_rn(T, __fadd_rn(1.f, -alpha));
            if (test_T < 0.0001f) { done = true; continue; }
            T = test_T; ++hits; last = (uint32_t)(base + j + 1);
        }
    }
    __syncthreads();
    if (inside) {
        atomicAdd(&s_tot[0], (unsigned long long)examined);
        atomicAdd(&s_tot[1], (unsigned long long)hits);
        atomicAdd(&s_tot[2], (unsigned long long)last);
    }
    __syncthreads();
    if (tid < 3 && s_tot[tid]) atomicAdd(&totals[tid], s_tot[tid]);
}

void launch_blend_count(const BlendArgs& a, unsigned long long* totals, cudaStream_t s) {
    FwdParams p;
    p.W = a.W; p.H = a.H; p.tiles_x = (a.W + 15) / 16;
    p.ranges = a.ranges; p.point_list = a.point_list; p.means2D = a.means2D; p.conic_opacity = a.conic_opacity;
    p.colors = a.colors; p.bg = a.bg; p.final_T = a.final_T; p.n_contrib = a.n_contrib; p.out_color = a.out_color;
    const int tiles_y = (a.H + 15) / 16;
    const int r0 = a.row1 > 0 ? a.row0 : 0, r1 = a.row1 > 0 ? a.row1 : tiles_y;
    if (r1 <= r0) return;
    p.row0 = r0;
    blend_count_kernel<<<dim3((a.W + 15) / 16, r1 - r0, 1), 256, 0, s>>>(p, totals);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// forward, ASYNC-STAGED variant (GSR_OPT_BLEND_VARIANT = 2; A/B only): the same kernel as blend_fwd2_kernel
// with the per-batch gather issued as cp.async (global -> shared without passing through registers, committed one batch ahead,
// waited with cp.async.wait_group) instead of __ldg into a register double buffer.  north_star (3) names "staging Gaussian
// batches into shared memory via TMA/cp.async": a list entry is a 4-byte id followed by an indexed gather of 8 + 16 + 12 bytes
// from three arrays, so there is no tile for a TMA descriptor; cp.async is the applicable form.  The exact tile cull and the
// compaction still need the values: they are read back from the raw staging area, culled, and compacted into s_ent.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async_4(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_8(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__global__ void __launch_bounds__(kBlend2Threads) blend_fwd3_kernel(const FwdParams p) {
    __shared__ float4 s_ent[3 * kBlend2Threads];
    __shared__ __align__(16) float4 s_co[2][kBlend2Threads];       // raw staging, double buffered
    __shared__ __align__(8) float2 s_xy[2][kBlend2Threads];
    __shared__ float s_col[2][3 * kBlend2Threads];
    __shared__ uint32_t s_wcnt[kBlend2Threads / 32];
    __shared__ uint32_t s_count;

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t lx = (wid & 1) * 8 + (lane & 3) * 2, ly = (wid >> 1) * 8 + (lane >> 2);
    const uint32_t tile_y = blockIdx.y + (uint32_t)p.row0;
    const uint32_t pix_x = blockIdx.x * 16 + lx, pix_y = tile_y * 16 + ly;
    const bool inside0 = pix_x < (uint32_t)p.W && pix_y < (uint32_t)p.H;
    const bool inside1 = pix_x + 1 < (uint32_t)p.W && pix_y < (uint32_t)p.H;
    const float pyf = (float)pix_y;
    const f32x2 npx2 = pk(-(float)pix_x, -(float)(pix_x + 1));
    const float rx0 = (float)(blockIdx.x * 16), ry0 = (float)(tile_y * 16);
    const float rx1 = rx0 + 15.f, ry1 = ry0 + 15.f;

    const uint2 range = p.ranges[tile_y * p.tiles_x + blockIdx.x];
    const int todo = (int)(range.y - range.x);
    const int rounds = (todo + kBlend2Threads - 1) / kBlend2Threads;

    float T0 = 1.0f, T1 = 1.0f, Ca0 = 0.f, Ca1 = 0.f, Ca2 = 0.f, Cb0 = 0.f, Cb1 = 0.f, Cb2 = 0.f;
    uint32_t last0 = 0, last1 = 0;
    bool done0 = !inside0, done1 = !inside1;

    auto issue = [&](int r, uint32_t id) {            // batch r's gather for this thread's entry, into buffer r & 1
        const int b = r & 1;
        cp_async_8(&s_xy[b][tid], p.means2D + id);
        cp_async_16(&s_co[b][tid], p.conic_opacity + id);
        cp_async_4(&s_col[b][3 * tid], p.colors + 3 * (size_t)id);
        cp_async_4(&s_col[b][3 * tid + 1], p.colors + 3 * (size_t)id + 1);
        cp_async_4(&s_col[b][3 * tid + 2], p.colors + 3 * (size_t)id + 2);
    };
    uint32_t id_next = 0;
    if (tid < todo) issue(0, __ldg(p.point_list + range.x + tid));
    cp_async_commit();
    if (kBlend2Threads + tid < todo) id_next = __ldg(p.point_list + range.x + kBlend2Threads + tid);

    for (int r = 0; r < rounds; ++r) {
        cp_async_wait_all();                            // this thread's copies of batch r have landed
        if (__syncthreads_count(done0 && done1) == kBlend2Threads) break;
        const bool valid = r * kBlend2Threads + tid < todo;
        const int b = r & 1;
        if ((r + 1) * kBlend2Threads + tid < todo) issue(r + 1, id_next);      // overlaps the blending of batch r
        cp_async_commit();
        if ((r + 2) * kBlend2Threads + tid < todo) id_next = __ldg(p.point_list + range.x + (r + 2) * kBlend2Threads + tid);
        GaussRec cur;
        if (valid) { cur.xy = s_xy[b][tid]; cur.co = s_co[b][tid]; cur.col = make_float3(s_col[b][3 * tid], s_col[b][3 * tid + 1], s_col[b][3 * tid + 2]); }
        const bool keep = valid && !cannot_touch(cur.xy, cur.co, rx0, ry0, rx1, ry1);
        const int slot = compact_slot<kBlend2Threads>(keep, s_wcnt, &s_count);
        if (slot >= 0) {
            s_ent[3 * slot] = make_float4(cur.xy.x, cur.xy.y, cur.co.x, cur.co.y);
            s_ent[3 * slot + 1] = make_float4(cur.co.z, reject_below(cur.co.w), cur.co.w, __uint_as_float((uint32_t)(r * kBlend2Threads + tid + 1)));
            s_ent[3 * slot + 2] = make_float4(cur.col.x, cur.col.y, cur.col.z, 0.f);
        }
        __syncthreads();
        const int count = (int)s_count;
        if (!(done0 && done1)) {
            const float4* e = s_ent;
            int lj0 = -1, lj1 = -1;
            for (int j = 0; j < count; ++j, e += 3) {
                const float4 A = e[0], B = e[1];
                const float dy = __fadd_rn(A.y, -pyf);
                float pw0, pw1;
                upk(blend_power2(A.z, A.w, B.x, add2(pk(A.x, A.x), npx2), dy), pw0, pw1);
                const bool c0 = !done0 && !(pw0 < B.y) && !(pw0 > 0.0f);
                const bool c1 = !done1 && !(pw1 < B.y) && !(pw1 > 0.0f);
                if (!(c0 || c1)) continue;
                const float4 Cc = e[2];
                if (c0) {
                    const float alpha = fminf(0.99f, __fmul_rn(B.z, expf(pw0)));
                    if (!(alpha < kAlphaMin)) {
                        const float test_T = __fmul_rn(T0, __fadd_rn(1.f, -alpha));
                        if (test_T < 0.0001f) done0 = true;
                        else {
                            Ca0 = __fmaf_rn(T0, __fmul_rn(alpha, Cc.x), Ca0); Ca1 = __fmaf_rn(T0, __fmul_rn(alpha, Cc.y), Ca1);
                            Ca2 = __fmaf_rn(T0, __fmul_rn(alpha, Cc.z), Ca2);
                            T0 = test_T; lj0 = j;
                        }
                    }
                }
                if (c1) {
                    const float alpha = fminf(0.99f, __fmul_rn(B.z, expf(pw1)));
                    if (!(alpha < kAlphaMin)) {
                        const float test_T = __fmul_rn(T1, __fadd_rn(1.f, -alpha));
                        if (test_T < 0.0001f) done1 = true;
                        else {
                            Cb0 = __fmaf_rn(T1, __fmul_rn(alpha, Cc.x), Cb0); Cb1 = __fmaf_rn(T1, __fmul_rn(alpha, Cc.y), Cb1);
                            Cb2 = __fmaf_rn(T1, __fmul_rn(alpha, Cc.z), Cb2);
                            T1 = test_T; lj1 = j;
                        }
                    }
                }
                if (done0 && done1) break;
            }
            if (lj0 >= 0) last0 = __float_as_uint(s_ent[3 * lj0 + 1].w);
            if (lj1 >= 0) last1 = __float_as_uint(s_ent[3 * lj1 + 1].w);
        }
    }
    cp_async_wait_all();
    const size_t HW = (size_t)p.H * p.W;
    const float bg0 = __ldg(p.bg), bg1 = __ldg(p.bg + 1), bg2 = __ldg(p.bg + 2);
    const size_t pix_id = (size_t)p.W * pix_y + pix_x;
    if (inside0) {
        p.final_T[pix_id] = T0; p.n_contrib[pix_id] = last0;
        p.out_color[pix_id] = __fmaf_rn(T0, bg0, Ca0); p.out_color[HW + pix_id] = __fmaf_rn(T0, bg1, Ca1); p.out_color[2 * HW + pix_id] = __fmaf_rn(T0, bg2, Ca2);
    }
    if (inside1) {
        p.final_T[pix_id + 1] = T1; p.n_contrib[pix_id + 1] = last1;
        p.out_color[pix_id + 1] = __fmaf_rn(T1, bg0, Cb0); p.out_color[HW + pix_id + 1] = __fmaf_rn(T1, bg1, Cb1); p.out_color[2 * HW + pix_id + 1] = __fmaf_rn(T1, bg2, Cb2);
    }
}

static int g_blend_variant = 0;
void set_blend_variant(int v) { g_blend_variant = v; }

void launch_blend_forward(const BlendArgs& a, cudaStream_t s) {
    FwdParams p;
    p.W = a.W; p.H = a.H; p.tiles_x = (a.W + 15) / 16;
    p.ranges = a.ranges; p.point_list = a.point_list; p.means2D = a.means2D; p.conic_opacity = a.conic_opacity;
    p.colors = a.colors; p.bg = a.bg; p.final_T = a.final_T; p.n_contrib = a.n_contrib; p.out_color = a.out_color;
    const int tiles_y = (a.H + 15) / 16;
    const int r0 = a.row1 > 0 ? a.row0 : 0, r1 = a.row1 > 0 ? a.row1 : tiles_y;
    if (r1 <= r0) return;
    p.row0 = r0;
    dim3 grid((a.W + 15) / 16, r1 - r0, 1);
    if (g_blend_variant == 1) blend_fwd_kernel<<<grid, kBlendThreads, 0, s>>>(p);      // one pixel per thread (A/B reference)
    else if (g_blend_variant == 2) blend_fwd3_kernel<<<grid, kBlend2Threads, 0, s>>>(p);   // cp.async staging (A/B, measured: DESIGN section 3)
    else blend_fwd2_kernel<<<grid, kBlend2Threads, 0, s>>>(p);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// forward, AMR: 32x32 binning tile, thread = one 2x2 quad, round k of the quad is pixel offset
// (0,0),(1,1),(1,0),(0,1) for k = 1..4 (A/forward.cu:313-338).  Rounds blended: last < k <= min(level,4).
// ------------------------------------------------------------------------------------------------
struct AmrParams {
    FwdParams f;
    const uint32_t* __restrict__ levels; const uint32_t* __restrict__ levels_last;
};

__global__ void __launch_bounds__(kBlendThreads) blend_amr_kernel(const AmrParams q) {
    const FwdParams& p = q.f;
    __shared__ float4 s_ent[3 * kBlendThreads];     // same staged-entry layout as blend_fwd_kernel
    __shared__ uint32_t s_wcnt[kBlendThreads / 32];
    __shared__ uint32_t s_count;

    const uint32_t tile_y = blockIdx.y + (uint32_t)p.row0;
    const int tile_id = tile_y * p.tiles_x + blockIdx.x;
    const uint32_t last = q.levels_last[tile_id];
    uint32_t level = q.levels[tile_id];
    if (level <= last) return;                  // A/forward.cu:285-288
    if (level > 4u) level = 4u;                 // A/forward.cu:341-342
    // active rounds: last < k <= level
    const int k_lo = (int)last + 1, k_hi = (int)level;
    if (k_lo > k_hi) return;

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t qx = (wid & 1) * 8 + (lane & 7), qy = (wid >> 1) * 4 + (lane >> 3);   // quad coords 0..15
    const uint32_t bx = blockIdx.x * 32 + 2 * qx, by = tile_y * 32 + 2 * qy;
    const float rx0 = (float)(blockIdx.x * 32), ry0 = (float)(tile_y * 32);
    const float rx1 = rx0 + 31.f, ry1 = ry0 + 31.f;

    // per-round pixel state; index k-1
    const int offx[4] = {0, 1, 1, 0}, offy[4] = {0, 1, 0, 1};
    float T[4], Cr[4], Cg[4], Cb[4], pxf[4], pyf[4];
    uint32_t lastc[4];
    bool done[4], ins[4];
    bool all_done = true;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t px = bx + offx[k], py = by + offy[k];
        const bool act = (k + 1 >= k_lo) && (k + 1 <= k_hi);
        ins[k] = act && px < (uint32_t)p.W && py < (uint32_t)p.H;
        done[k] = !ins[k];
        all_done = all_done && done[k];
        T[k] = 1.f; Cr[k] = Cg[k] = Cb[k] = 0.f; lastc[k] = 0;
        pxf[k] = (float)px; pyf[k] = (float)py;
    }

    const uint2 range = p.ranges[tile_id];
    const int todo = (int)(range.y - range.x);
    const int rounds = (todo + kBlendThreads - 1) / kBlendThreads;

    GaussRec nxt; bool nxt_valid = false;
    if (tid < todo) { nxt = gather(p.point_list, p.means2D, p.conic_opacity, p.colors, range.x + tid); nxt_valid = true; }

    for (int r = 0; r < rounds; ++r) {
        if (__syncthreads_count(all_done) == kBlendThreads) break;
        const bool keep = nxt_valid && !cannot_touch(nxt.xy, nxt.co, rx0, ry0, rx1, ry1);
        const int slot = compact_slot(keep, s_wcnt, &s_count);
        if (slot >= 0) {
            s_ent[3 * slot] = make_float4(nxt.xy.x, nxt.xy.y, nxt.co.x, nxt.co.y);
            s_ent[3 * slot + 1] = make_float4(nxt.co.z, reject_below(nxt.co.w), nxt.co.w, __uint_as_float((uint32_t)(r * kBlendThreads + tid + 1)));
            s_ent[3 * slot + 2] = make_float4(nxt.col.x, nxt.col.y, nxt.col.z, 0.f);
        }
        __syncthreads();
        const int count = (int)s_count;
        const int nprog = (r + 1) * kBlendThreads + tid;
        nxt_valid = nprog < todo;
        if (nxt_valid) nxt = gather(p.point_list, p.means2D, p.conic_opacity, p.colors, range.x + nprog);

        if (!all_done) {
            const float4* e = s_ent;
            for (int j = 0; !all_done && j < count; ++j, e += 3) {
                const float4 A = e[0], B = e[1];
                const float4 Cc = e[2];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (done[k]) continue;
                    const float dx = __fadd_rn(A.x, -pxf[k]), dy = __fadd_rn(A.y, -pyf[k]);
                    const float power = blend_power(A.z, A.w, B.x, dx, dy);
                    if (power < B.y || power > 0.0f) continue;   // certain alpha < 1/255 (reject_below) | forward.cu:337
                    const float alpha = fminf(0.99f, __fmul_rn(B.z, expf(power)));
                    if (alpha < kAlphaMin) continue;
                    const float test_T = __fmul_rn(T[k], __fadd_rn(1.f, -alpha));
                    if (test_T < 0.0001f) { done[k] = true; continue; }
                    Cr[k] = __fmaf_rn(T[k], __fmul_rn(alpha, Cc.x), Cr[k]);
                    Cg[k] = __fmaf_rn(T[k], __fmul_rn(alpha, Cc.y), Cg[k]);
                    Cb[k] = __fmaf_rn(T[k], __fmul_rn(alpha, Cc.z), Cb[k]);
                    T[k] = test_T;
                    lastc[k] = __float_as_uint(B.w);
                }
                all_done = done[0] && done[1] && done[2] && done[3];
            }
        }
    }
    const size_t HW = (size_t)p.H * p.W;
    const float bg0 = __ldg(p.bg), bg1 = __ldg(p.bg + 1), bg2 = __ldg(p.bg + 2);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (!ins[k]) continue;
        const uint32_t pix_id = (uint32_t)p.W * (by + offy[k]) + (bx + offx[k]);
        p.final_T[pix_id] = T[k];
        p.n_contrib[pix_id] = lastc[k];
        p.out_color[pix_id] = __fmaf_rn(T[k], bg0, Cr[k]);
        p.out_color[HW + pix_id] = __fmaf_rn(T[k], bg1, Cg[k]);
        p.out_color[2 * HW + pix_id] = __fmaf_rn(T[k], bg2, Cb[k]);
    }
}

// nearest-copy fill of the lattice pixels a tile's level does not render (A/forward.cu:520-648).
// One thread per pixel.  Deterministic restatement: where the reference reads a pixel that another
// CTA of the same launch is concurrently overwriting with out_color_precomp (level <= last in stepped
// mode), the copied-in value is used.
struct InterpParams {
    int W, H, tiles_x, fovea_step, y_begin, y_end;     // pixel rows [y_begin, y_end) of the launch (tile bands)
    const uint32_t* __restrict__ levels; const uint32_t* __restrict__ levels_last;
    float* final_T; uint32_t* n_contrib; float* out_color; const float* out_color_precomp;
};

__global__ void __launch_bounds__(256) interpolate_kernel(const InterpParams p) {
    const uint32_t x = blockIdx.x * 32 + (threadIdx.x & 31), y = (uint32_t)p.y_begin + blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= (uint32_t)p.W || y >= (uint32_t)p.y_end) return;
    const int tile_id = (y / 32) * p.tiles_x + (x / 32);
    const uint32_t ox = x & 1, oy = y & 1;
    const int round = (ox == 0) ? (oy == 0 ? 1 : 4) : (oy == 0 ? 3 : 2);
    uint32_t level = p.levels[tile_id];
    if (level > 4u) level = 4u;
    const size_t HW = (size_t)p.H * p.W;
    const uint32_t pix = (uint32_t)p.W * y + x;
    bool src_from_precomp = false;
    if (p.fovea_step > 0) {
        const int last = (int)p.levels_last[tile_id];
        if ((int)level <= last) {
            for (int ch = 0; ch < 3; ++ch) p.out_color[ch * HW + pix] = p.out_color_precomp[ch * HW + pix];
            src_from_precomp = true;
        }
        if (round < last) {
            for (int ch = 0; ch < 3; ++ch) p.out_color[ch * HW + pix] = p.out_color_precomp[ch * HW + pix];
            return;
        }
    }
    if ((uint32_t)round <= level) return;          // rendered accurately
    const uint32_t sx = (x & ~1u) + (level >= 3u ? 1u : 0u), sy = (y & ~1u) + (level >= 3u ? 1u : 0u);
    if (sx < (uint32_t)p.W && sy < (uint32_t)p.H) {
        const uint32_t spix = (uint32_t)p.W * sy + sx;
        p.final_T[pix] = p.final_T[spix];
        p.n_contrib[pix] = p.n_contrib[spix];
        const float* src = src_from_precomp ? p.out_color_precomp : p.out_color;
        for (int ch = 0; ch < 3; ++ch) p.out_color[ch * HW + pix] = src[ch * HW + spix];
    }
}

void launch_blend_amr(const AmrBlendArgs& a, cudaStream_t s) {
    AmrParams q;
    FwdParams& p = q.f;
    const BlendArgs& b = a.b;
    p.W = b.W; p.H = b.H; p.tiles_x = (b.W + 31) / 32;
    p.ranges = b.ranges; p.point_list = b.point_list; p.means2D = b.means2D; p.conic_opacity = b.conic_opacity;
    p.colors = b.colors; p.bg = b.bg; p.final_T = b.final_T; p.n_contrib = b.n_contrib; p.out_color = b.out_color;
    q.levels = a.levels; q.levels_last = a.levels_last;
    const int tiles_y = (b.H + 31) / 32;
    const int r0 = b.row1 > 0 ? b.row0 : 0, r1 = b.row1 > 0 ? b.row1 : tiles_y;
    if (r1 <= r0) return;
    p.row0 = r0;
    dim3 grid((b.W + 31) / 32, r1 - r0, 1);
    blend_amr_kernel<<<grid, kBlendThreads, 0, s>>>(q);
    count_launch();
    if (a.interpolate) {
        InterpParams ip;
        ip.W = b.W; ip.H = b.H; ip.tiles_x = p.tiles_x; ip.fovea_step = a.fovea_step;
        ip.levels = a.levels; ip.levels_last = a.levels_last;
        ip.final_T = b.final_T; ip.n_contrib = b.n_contrib; ip.out_color = b.out_color;
        ip.out_color_precomp = a.out_color_precomp ? a.out_color_precomp : b.out_color;
        ip.y_begin = r0 * 32; ip.y_end = min(b.H, r1 * 32);
        dim3 g2((b.W + 31) / 32, (ip.y_end - ip.y_begin + 7) / 8, 1);
        interpolate_kernel<<<g2, 256, 0, s>>>(ip);
        count_launch();
    }
}

// ------------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------------
struct BwdParams {
    int W, H, tiles_x, row0;
    const uint2* __restrict__ ranges; const uint32_t* __restrict__ point_list;
    const float2* __restrict__ means2D; const float4* __restrict__ conic_opacity; const float* __restrict__ colors;
    const float* __restrict__ bg; const float* __restrict__ final_T; const uint32_t* __restrict__ n_contrib;
    const float* __restrict__ dL_dpix;
    float* scratch;
};

__global__ void __launch_bounds__(kBlendThreads) blend_bwd_kernel(const BwdParams p) {
    __shared__ float4 s_a[kBlendThreads];
    __shared__ float4 s_b[kBlendThreads];
    __shared__ float4 s_c[kBlendThreads];     // b, list position (1-based), id, unused
    __shared__ uint32_t s_wcnt[kBlendThreads / 32];
    __shared__ uint32_t s_count;

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t lx = (wid & 1) * 8 + (lane & 7), ly = (wid >> 1) * 4 + (lane >> 3);
    const uint32_t tile_y = blockIdx.y + (uint32_t)p.row0;
    const uint32_t pix_x = blockIdx.x * 16 + lx, pix_y = tile_y * 16 + ly;
    const bool inside = pix_x < (uint32_t)p.W && pix_y < (uint32_t)p.H;
    const uint32_t pix_id = (uint32_t)p.W * pix_y + pix_x;
    const float pxf = (float)pix_x, pyf = (float)pix_y;
    const float rx0 = (float)(blockIdx.x * 16), ry0 = (float)(tile_y * 16);
    const float rx1 = rx0 + 15.f, ry1 = ry0 + 15.f;
    const size_t HW = (size_t)p.H * p.W;

    const uint2 range = p.ranges[tile_y * p.tiles_x + blockIdx.x];
    const int todo = (int)(range.y - range.x);
    const int rounds = (todo + kBlendThreads - 1) / kBlendThreads;

    const float T_final = inside ? p.final_T[pix_id] : 0.f;
    float T = T_final;
    const uint32_t last_contributor = inside ? p.n_contrib[pix_id] : 0u;
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f;          // accum_rec
    float lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, last_alpha = 0.f;
    float g0 = 0.f, g1 = 0.f, g2 = 0.f;
    if (inside) { g0 = p.dL_dpix[pix_id]; g1 = p.dL_dpix[HW + pix_id]; g2 = p.dL_dpix[2 * HW + pix_id]; }
    const float bg_dot = p.bg[0] * g0 + p.bg[1] * g1 + p.bg[2] * g2;
    const float ddelx_dx = 0.5f * (float)p.W, ddely_dy = 0.5f * (float)p.H;

    // the CTA only needs list entries up to the largest n_contrib of its pixels
    __shared__ uint32_t s_maxc;
    if (tid == 0) s_maxc = 0;
    __syncthreads();
    {
        uint32_t m = last_contributor;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) atomicMax(&s_maxc, m);
    }
    __syncthreads();
    const int max_c = (int)s_maxc;        // entries with 1-based position > max_c are skipped by every pixel

    GaussRec nxt; bool nxt_valid = false;
    // batch r covers 1-based positions todo - r*256 - tid   (back to front)
    auto pos_of = [&](int r) { return todo - r * kBlendThreads - tid; };
    const int r_first = (todo - max_c) / kBlendThreads;  // earlier batches hold only positions > max_c
    {
        const int pos = pos_of(r_first);
        nxt_valid = pos >= 1 && pos <= max_c;
        if (nxt_valid) nxt = gather(p.point_list, p.means2D, p.conic_opacity, p.colors, range.x + pos - 1);
    }

    for (int r = r_first; r < rounds; ++r) {
        __syncthreads();
        const bool keep = nxt_valid && !cannot_touch(nxt.xy, nxt.co, rx0, ry0, rx1, ry1);
        const int slot = compact_slot(keep, s_wcnt, &s_count);
        if (slot >= 0) {
            s_a[slot] = make_float4(nxt.xy.x, nxt.xy.y, nxt.co.x, nxt.co.y);
            s_b[slot] = make_float4(nxt.co.z, nxt.co.w, nxt.col.x, nxt.col.y);
            s_c[slot] = make_float4(nxt.col.z, __uint_as_float((uint32_t)pos_of(r)), __uint_as_float(nxt.id), reject_below(nxt.co.w));
        }
        __syncthreads();
        const int count = (int)s_count;
        {
            const int pos = pos_of(r + 1);
            nxt_valid = (r + 1 < rounds) && pos >= 1 && pos <= max_c;
            if (nxt_valid) nxt = gather(p.point_list, p.means2D, p.conic_opacity, p.colors, range.x + pos - 1);
        }

        for (int j = 0; j < count; ++j) {
            const float4 A = s_a[j], B = s_b[j], Cc = s_c[j];
            const uint32_t pos = __float_as_uint(Cc.y);
            const float dx = A.x - pxf, dy = A.y - pyf;
            const float power = blend_power(A.z, A.w, B.x, dx, dy);
            // (!inside pixels carry last_contributor = 0 and positions are 1-based, so no separate `inside` test)
            // candidates: not behind the pixel's last contributor, not certain to fail the alpha test (reject_below)
            const bool cand = pos <= last_contributor && !(power < Cc.w) && !(power > 0.0f);
            if (!__any_sync(0xffffffffu, cand)) continue;
            const float G = expf(power);
            const float alpha = fminf(0.99f, B.y * G);
            const bool act = cand && !(alpha < kAlphaMin);
            if (!__any_sync(0xffffffffu, act)) continue;

            float v[8], v8 = 0.f;
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = 0.f;
            if (act) {
                const float inv_1ma = __frcp_rn(1.f - alpha);      // one correctly rounded reciprocal serves both divisions
                T = T * inv_1ma;
                const float dchannel_dcolor = alpha * T;
                acc0 = last_alpha * lc0 + (1.f - last_alpha) * acc0;
                acc1 = last_alpha * lc1 + (1.f - last_alpha) * acc1;
                acc2 = last_alpha * lc2 + (1.f - last_alpha) * acc2;
                lc0 = B.z; lc1 = B.w; lc2 = Cc.x;
                float dL_dalpha = (lc0 - acc0) * g0 + (lc1 - acc1) * g1 + (lc2 - acc2) * g2;
                dL_dalpha *= T;
                last_alpha = alpha;
                dL_dalpha += (-T_final * inv_1ma) * bg_dot;
                const float dL_dG = B.y * dL_dalpha;
                const float gdx = G * dx, gdy = G * dy;
                const float dG_ddelx = -gdx * A.z - gdy * A.w;
                const float dG_ddely = -gdy * B.x - gdx * A.w;
                v[0] = dL_dG * dG_ddelx * ddelx_dx;
                v[1] = dL_dG * dG_ddely * ddely_dy;
                v[2] = -0.5f * gdx * dx * dL_dG;
                v[3] = -0.5f * gdx * dy * dL_dG;
                v[4] = -0.5f * gdy * dy * dL_dG;
                v[5] = G * dL_dalpha;
                v[6] = dchannel_dcolor * g0;
                v[7] = dchannel_dcolor * g1;
                v8 = dchannel_dcolor * g2;
            }
            // butterfly reduction: after three halving steps lane l holds value index (l>>2)&7 summed over
            // lane bits 4,3,2; two more steps sum over bits 1,0.  9 shuffles for 8 values instead of 40.
            float w4[4], w2[2], w1;
            {
                const bool hi = lane & 16;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float keepv = hi ? v[4 + i] : v[i], send = hi ? v[i] : v[4 + i];
                    w4[i] = keepv + __shfl_xor_sync(0xffffffffu, send, 16);
                }
            }
            {
                const bool hi = lane & 8;
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const float keepv = hi ? w4[2 + i] : w4[i], send = hi ? w4[i] : w4[2 + i];
                    w2[i] = keepv + __shfl_xor_sync(0xffffffffu, send, 8);
                }
            }
            {
                const bool hi = lane & 4;
                const float keepv = hi ? w2[1] : w2[0], send = hi ? w2[0] : w2[1];
                w1 = keepv + __shfl_xor_sync(0xffffffffu, send, 4);
            }
            w1 += __shfl_xor_sync(0xffffffffu, w1, 2);
            w1 += __shfl_xor_sync(0xffffffffu, w1, 1);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v8 += __shfl_xor_sync(0xffffffffu, v8, o);

            float* dst = p.scratch + 12 * (size_t)__float_as_uint(Cc.z);
            if ((lane & 3) == 0) atomicAdd(dst + ((lane >> 2) & 7), w1);
            if (lane == 1) atomicAdd(dst + 8, v8);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// backward, two pixels per thread (default).  blend_bwd_kernel is bound by instruction issue (ncu: 84 % issue-active);
// here a thread owns two horizontally adjacent pixels, the gradient chain runs in packed FP32x2, the two pixels'
// terms are added before the warp butterfly (one reduction and one atomic set per 64 pixels instead of 32), and 128
// threads cover the tile.  A pixel that an entry does not touch is carried through the same arithmetic with
// alpha = 0 and G = 0: T, the deferred colour accumulator and every gradient term then come out unchanged / zero
// exactly (acc' = a*lc + (1-a)*acc applied early followed by a zero-weight entry is the same recurrence), so no
// per-pixel selects are needed.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlend2Threads, 6) blend_bwd2_kernel(const BwdParams p) {
    __shared__ float4 s_a[kBlend2Threads];
    __shared__ float4 s_b[kBlend2Threads];
    __shared__ float4 s_c[kBlend2Threads];     // b, list position (1-based), id, reject threshold
    __shared__ uint32_t s_wcnt[kBlend2Threads / 32];
    __shared__ uint32_t s_count;
    __shared__ uint32_t s_maxc;

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t lx = (wid & 1) * 8 + (lane & 3) * 2, ly = (wid >> 1) * 8 + (lane >> 2);
    const uint32_t tile_y = blockIdx.y + (uint32_t)p.row0;
    const uint32_t pix_x = blockIdx.x * 16 + lx, pix_y = tile_y * 16 + ly;
    const bool in0 = pix_x < (uint32_t)p.W && pix_y < (uint32_t)p.H;
    const bool in1 = pix_x + 1 < (uint32_t)p.W && pix_y < (uint32_t)p.H;
    const size_t pix_id = (size_t)p.W * pix_y + pix_x;
    const float pyf = (float)pix_y;
    const f32x2 npx2 = pk(-(float)pix_x, -(float)(pix_x + 1));
    const float rx0 = (float)(blockIdx.x * 16), ry0 = (float)(tile_y * 16);
    const float rx1 = rx0 + 15.f, ry1 = ry0 + 15.f;
    const size_t HW = (size_t)p.H * p.W;

    const uint2 range = p.ranges[tile_y * p.tiles_x + blockIdx.x];
    const int todo = (int)(range.y - range.x);
    const int rounds = (todo + kBlend2Threads - 1) / kBlend2Threads;

    const float Tf0 = in0 ? p.final_T[pix_id] : 0.f, Tf1 = in1 ? p.final_T[pix_id + 1] : 0.f;
    const uint32_t lc0 = in0 ? p.n_contrib[pix_id] : 0u, lc1 = in1 ? p.n_contrib[pix_id + 1] : 0u;
    float ga0 = 0.f, ga1 = 0.f, ga2 = 0.f, gb0 = 0.f, gb1 = 0.f, gb2 = 0.f;
    if (in0) { ga0 = p.dL_dpix[pix_id]; ga1 = p.dL_dpix[HW + pix_id]; ga2 = p.dL_dpix[2 * HW + pix_id]; }
    if (in1) { gb0 = p.dL_dpix[pix_id + 1]; gb1 = p.dL_dpix[HW + pix_id + 1]; gb2 = p.dL_dpix[2 * HW + pix_id + 1]; }
    const float bgx = p.bg[0], bgy = p.bg[1], bgz = p.bg[2];
    const f32x2 g0_2 = pk(ga0, gb0), g1_2 = pk(ga1, gb1), g2_2 = pk(ga2, gb2);
    const f32x2 nTf_bg2 = pk(-Tf0 * (bgx * ga0 + bgy * ga1 + bgz * ga2), -Tf1 * (bgx * gb0 + bgy * gb1 + bgz * gb2));   // -T_final * bg_dot
    const float ddelx_dx = 0.5f * (float)p.W, ddely_dy = 0.5f * (float)p.H;
    f32x2 T2 = pk(Tf0, Tf1);
    f32x2 acc0_2 = pk(0.f, 0.f), acc1_2 = acc0_2, acc2_2 = acc0_2, la2 = acc0_2;      // accum_rec, last_alpha
    float lcol0 = 0.f, lcol1 = 0.f, lcol2 = 0.f;                                        // last colour (same for both pixels)

    // the CTA only needs list entries up to the largest n_contrib of its pixels
    if (tid == 0) s_maxc = 0;
    __syncthreads();
    {
        uint32_t m = max(lc0, lc1);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) atomicMax(&s_maxc, m);
    }
    __syncthreads();
    const int max_c = (int)s_maxc;

    GaussRec nxt; bool nxt_valid = false;
    auto pos_of = [&](int r) { return todo - r * kBlend2Threads - tid; };      // batch r, back to front, 1-based
    const int r_first = (todo - max_c) / kBlend2Threads;
    {
        const int pos = pos_of(r_first);
        nxt_valid = pos >= 1 && pos <= max_c;
        if (nxt_valid) nxt = gather(p.point_list, p.means2D, p.conic_opacity, p.colors, range.x + pos - 1);
    }

    for (int r = r_first; r < rounds; ++r) {
        __syncthreads();
        const bool keep = nxt_valid && !cannot_touch(nxt.xy, nxt.co, rx0, ry0, rx1, ry1);
        const int slot = compact_slot<kBlend2Threads>(keep, s_wcnt, &s_count);
        if (slot >= 0) {
            s_a[slot] = make_float4(nxt.xy.x, nxt.xy.y, nxt.co.x, nxt.co.y);
            s_b[slot] = make_float4(nxt.co.z, nxt.co.w, nxt.col.x, nxt.col.y);
            s_c[slot] = make_float4(nxt.col.z, __uint_as_float((uint32_t)pos_of(r)), __uint_as_float(nxt.id), reject_below(nxt.co.w));
        }
        __syncthreads();
        const int count = (int)s_count;
        {
            const int pos = pos_of(r + 1);
            nxt_valid = (r + 1 < rounds) && pos >= 1 && pos <= max_c;
            if (nxt_valid) nxt = gather(p.point_list, p.means2D, p.conic_opacity, p.colors, range.x + pos - 1);
        }

        for (int j = 0; j < count; ++j) {
            const float4 A = s_a[j], B = s_b[j], Cc = s_c[j];
            const uint32_t pos = __float_as_uint(Cc.y);
            const float dy = A.y - pyf;
            const f32x2 dx2 = add2(pk(A.x, A.x), npx2);
            float pw0, pw1;
            upk(blend_power2(A.z, A.w, B.x, dx2, dy), pw0, pw1);
            const bool c0 = pos <= lc0 && !(pw0 < Cc.w) && !(pw0 > 0.0f);
            const bool c1 = pos <= lc1 && !(pw1 < Cc.w) && !(pw1 > 0.0f);
            if (!__any_sync(0xffffffffu, c0 || c1)) continue;
            float G0 = expf(pw0), G1 = expf(pw1);
            float al0 = fminf(0.99f, B.y * G0), al1 = fminf(0.99f, B.y * G1);
            const bool a0 = c0 && !(al0 < kAlphaMin), a1 = c1 && !(al1 < kAlphaMin);
            if (!__any_sync(0xffffffffu, a0 || a1)) continue;
            if (!a0) { al0 = 0.f; G0 = 0.f; }
            if (!a1) { al1 = 0.f; G1 = 0.f; }
            const f32x2 al2 = pk(al0, al1), G2 = pk(G0, G1);
            const f32x2 inv2 = pk(__frcp_rn(1.f - al0), __frcp_rn(1.f - al1));       // 1 for an untouched pixel
            T2 = mul2(T2, inv2);
            const f32x2 dch2 = mul2(al2, T2);                                          // dchannel_dcolor
            const f32x2 oml2 = add2(pk(1.f, 1.f), la2 ^ 0x8000000080000000ull);         // 1 - last_alpha
            acc0_2 = fma2(la2, pk(lcol0, lcol0), mul2(oml2, acc0_2));
            acc1_2 = fma2(la2, pk(lcol1, lcol1), mul2(oml2, acc1_2));
            acc2_2 = fma2(la2, pk(lcol2, lcol2), mul2(oml2, acc2_2));
            lcol0 = B.z; lcol1 = B.w; lcol2 = Cc.x;
            f32x2 dLa2 = mul2(add2(pk(lcol0, lcol0), acc0_2 ^ 0x8000000080000000ull), g0_2);
            dLa2 = fma2(add2(pk(lcol1, lcol1), acc1_2 ^ 0x8000000080000000ull), g1_2, dLa2);
            dLa2 = fma2(add2(pk(lcol2, lcol2), acc2_2 ^ 0x8000000080000000ull), g2_2, dLa2);
            dLa2 = mul2(dLa2, T2);
            la2 = al2;
            dLa2 = fma2(nTf_bg2, inv2, dLa2);
            const f32x2 dLdG2 = mul2(pk(B.y, B.y), dLa2);
            const f32x2 gdx2 = mul2(G2, dx2), gdy2 = mul2(G2, pk(dy, dy));
            const f32x2 ncx = pk(-A.z, -A.z), ncy = pk(-A.w, -A.w), ncz = pk(-B.x, -B.x);
            const f32x2 dGx2 = fma2(gdx2, ncx, mul2(gdy2, ncy));                       // -gdx*cx - gdy*cy
            const f32x2 dGy2 = fma2(gdy2, ncz, mul2(gdx2, ncy));                       // -gdy*cz - gdx*cy
            const f32x2 h2 = mul2(pk(-0.5f, -0.5f), dLdG2);
            float v[8], v8, lo, hi;
            upk(mul2(mul2(dLdG2, dGx2), pk(ddelx_dx, ddelx_dx)), lo, hi); v[0] = lo + hi;
            upk(mul2(mul2(dLdG2, dGy2), pk(ddely_dy, ddely_dy)), lo, hi); v[1] = lo + hi;
            upk(mul2(mul2(gdx2, dx2), h2), lo, hi); v[2] = lo + hi;
            upk(mul2(mul2(gdx2, pk(dy, dy)), h2), lo, hi); v[3] = lo + hi;
            upk(mul2(mul2(gdy2, pk(dy, dy)), h2), lo, hi); v[4] = lo + hi;
            upk(mul2(G2, dLa2), lo, hi); v[5] = lo + hi;
            upk(mul2(dch2, g0_2), lo, hi); v[6] = lo + hi;
            upk(mul2(dch2, g1_2), lo, hi); v[7] = lo + hi;
            upk(mul2(dch2, g2_2), lo, hi); v8 = lo + hi;

            // butterfly reduction (as in blend_bwd_kernel): 9 shuffles for 8 values
            float w4[4], w2[2], w1;
            {
                const bool hi_ = lane & 16;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float keepv = hi_ ? v[4 + i] : v[i], send = hi_ ? v[i] : v[4 + i];
                    w4[i] = keepv + __shfl_xor_sync(0xffffffffu, send, 16);
                }
            }
            {
                const bool hi_ = lane & 8;
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const float keepv = hi_ ? w4[2 + i] : w4[i], send = hi_ ? w4[i] : w4[2 + i];
                    w2[i] = keepv + __shfl_xor_sync(0xffffffffu, send, 8);
                }
            }
            {
                const bool hi_ = lane & 4;
                const float keepv = hi_ ? w2[1] : w2[0], send = hi_ ? w2[0] : w2[1];
                w1 = keepv + __shfl_xor_sync(0xffffffffu, send, 4);
            }
            w1 += __shfl_xor_sync(0xffffffffu, w1, 2);
            w1 += __shfl_xor_sync(0xffffffffu, w1, 1);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v8 += __shfl_xor_sync(0xffffffffu, v8, o);

            float* dst = p.scratch + 12 * (size_t)__float_as_uint(Cc.z);
            if ((lane & 3) == 0) atomicAdd(dst + ((lane >> 2) & 7), w1);
            if (lane == 1) atomicAdd(dst + 8, v8);
        }
    }
}

void launch_blend_backward(const BlendBwdArgs& a, cudaStream_t s) {
    BwdParams p;
    p.W = a.W; p.H = a.H; p.tiles_x = (a.W + 15) / 16;
    p.ranges = a.ranges; p.point_list = a.point_list; p.means2D = a.means2D; p.conic_opacity = a.conic_opacity;
    p.colors = a.colors; p.bg = a.bg; p.final_T = a.final_T; p.n_contrib = a.n_contrib; p.dL_dpix = a.dL_dpix;
    p.scratch = a.scratch;
    const int tiles_y = (a.H + 15) / 16;
    const int r0 = a.row1 > 0 ? a.row0 : 0, r1 = a.row1 > 0 ? a.row1 : tiles_y;
    if (r1 <= r0) return;
    p.row0 = r0;
    dim3 grid((a.W + 15) / 16, r1 - r0, 1);
    if (g_blend_variant == 1) blend_bwd_kernel<<<grid, kBlendThreads, 0, s>>>(p);      // one pixel per thread (A/B reference)
    else blend_bwd2_kernel<<<grid, kBlend2Threads, 0, s>>>(p);
    count_launch();
}

}  // namespace gsr
