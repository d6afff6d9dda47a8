// gsr_train.cu -- the two HBM-bound kernels either side of the rasterizer in a training step:
//
//   l1_ssim_fwd_kernel / l1_ssim_bwd_kernel   <- utils/loss_utils.py:17-63 (l1_loss, create_window, ssim, _ssim) as used by
//                                                train.py:86-89: loss = (1-l)*L1 + l*(1 - ssim(image, gt))
//   adam_multi_kernel                         <- torch.optim.Adam(l, lr=0.0, eps=1e-15) over the six parameter groups
//                                                (scene/gaussian_model.py:149-175), train.py:119-121
//
// The reference evaluates SSIM as five grouped 11x11 conv2d calls plus ~15 elementwise kernels and lets autograd replay
// them (ten more convolutions); here one kernel per direction does the separable 11-tap Gaussian window in shared
// memory on a 32x32 pixel tile (+5 pixel halo), forward writes the three partial-derivative maps the backward
// convolves.  Zero padding (F.conv2d padding=5) is reproduced by zero-filling the halo outside the image.
#include "gsr_common.cuh"

namespace gsr {

namespace {

constexpr int kWin = 11;
constexpr int kHalo = kWin / 2;         // 5
constexpr int kTile = 32;               // output pixels per CTA edge
constexpr int kIn = kTile + 2 * kHalo;  // 42
constexpr int kLossThreads = 256;
constexpr float kC1 = 0.01f * 0.01f;
constexpr float kC2 = 0.03f * 0.03f;

// gaussian(11, 1.5) normalised (loss_utils.py:23-25), computed on the host in double and rounded to float once --
// torch builds it from exp() in Python floats and stores float32, the same values.
struct Window { float w[kWin]; };

Window make_window() {
    // torch.Tensor([exp(..) ...]) rounds each value to float32; gauss.sum() in torch equals the correctly rounded double
    // sum of those floats (a sequential float sum is one ulp lower) -- reproduced so the 1-D window is bit-identical.
    float gf[kWin]; double sd = 0.0;
    for (int i = 0; i < kWin; ++i) { double d = i - kWin / 2; gf[i] = (float)exp(-(d * d) / (2.0 * 1.5 * 1.5)); sd += (double)gf[i]; }
    const float sf = (float)sd;
    Window w;
    for (int i = 0; i < kWin; ++i) w.w[i] = gf[i] / sf;
    return w;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// One CTA = one 32x32 tile of one channel.  256 threads.
//   phase 0: stage x (image) and y (ground truth) with the halo, zero outside the image
//   phase 1: horizontal 11-tap pass of {x, y, xx, yy, xy} for 42 rows x 32 columns (each thread: 4 adjacent columns of
//            one row from a 14-value register window)
//   phase 2: vertical pass for the 32x32 outputs (each thread: 4 adjacent rows of one column), SSIM, the three maps
//   phase 3: block reduction of sum|x-y| and sum(ssim) into partials[cta]
template <bool kWriteMaps>
__global__ void __launch_bounds__(kLossThreads)
l1_ssim_fwd_kernel(int C, int H, int W, const float* __restrict__ img, const float* __restrict__ gt, Window win,
                   float* __restrict__ maps, float* __restrict__ partials) {
    __shared__ float sx[kIn][kIn + 1];
    __shared__ float sy[kIn][kIn + 1];
    __shared__ float hs[5][kIn][kTile + 1];
    __shared__ float red[2][kLossThreads / 32];

    const int c = blockIdx.z;
    const int x0 = blockIdx.x * kTile, y0 = blockIdx.y * kTile;
    const size_t plane = (size_t)H * W;
    const float* px = img + (size_t)c * plane;
    const float* py = gt + (size_t)c * plane;
    const int t = threadIdx.x;

    for (int i = t; i < kIn * kIn; i += kLossThreads) {
        const int r = i / kIn, q = i - r * kIn;
        const int gy = y0 + r - kHalo, gx = x0 + q - kHalo;
        float a = 0.f, b = 0.f;
        if (gy >= 0 && gy < H && gx >= 0 && gx < W) { a = __ldg(px + (size_t)gy * W + gx); b = __ldg(py + (size_t)gy * W + gx); }
        sx[r][q] = a; sy[r][q] = b;
    }
    __syncthreads();

    // horizontal: 42 rows x 8 column groups of 4 = 336 work items
    for (int i = t; i < kIn * (kTile / 4); i += kLossThreads) {
        const int r = i / (kTile / 4), cg = (i - r * (kTile / 4)) * 4;
        float a[kWin + 3], b[kWin + 3];
#pragma unroll
        for (int k = 0; k < kWin + 3; ++k) { a[k] = sx[r][cg + k]; b[k] = sy[r][cg + k]; }
#pragma unroll
        for (int o = 0; o < 4; ++o) {
            float m1 = 0.f, m2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
            for (int k = 0; k < kWin; ++k) {
                const float w = win.w[k], xa = a[o + k], yb = b[o + k];
                const float wx = w * xa, wy = w * yb;
                m1 += wx; m2 += wy; e11 = fmaf(wx, xa, e11); e22 = fmaf(wy, yb, e22); e12 = fmaf(wx, yb, e12);
            }
            hs[0][r][cg + o] = m1; hs[1][r][cg + o] = m2; hs[2][r][cg + o] = e11; hs[3][r][cg + o] = e22; hs[4][r][cg + o] = e12;
        }
    }
    __syncthreads();

    // vertical: thread -> column (t & 31), rows 4*(t >> 5) .. +3
    float l1_acc = 0.f, ss_acc = 0.f;
    {
        const int col = t & 31, rg = (t >> 5) * 4;
        float v[5][kWin + 3];
#pragma unroll
        for (int q = 0; q < 5; ++q)
#pragma unroll
            for (int k = 0; k < kWin + 3; ++k) v[q][k] = hs[q][rg + k][col];
#pragma unroll
        for (int o = 0; o < 4; ++o) {
            float s[5];
#pragma unroll
            for (int q = 0; q < 5; ++q) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < kWin; ++k) acc = fmaf(win.w[k], v[q][o + k], acc);
                s[q] = acc;
            }
            const int gy = y0 + rg + o, gx = x0 + col;
            if (gy < H && gx < W) {
                const float mu1 = s[0], mu2 = s[1];
                const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
                const float sig1 = s[2] - mu1_sq, sig2 = s[3] - mu2_sq, sig12 = s[4] - mu12;
                const float A = 2.f * mu12 + kC1, B = 2.f * sig12 + kC2;
                const float Cc = mu1_sq + mu2_sq + kC1, D = sig1 + sig2 + kC2;
                const float inv = 1.f / (Cc * D);
                const float val = (A * B) * inv;
                ss_acc += val;
                const float xv = sx[rg + o + kHalo][col + kHalo], yv = sy[rg + o + kHalo][col + kHalo];
                l1_acc += fabsf(xv - yv);
                if (kWriteMaps) {
                    // d ssim / d{mu1, E[xx], E[xy]} at this window centre (E[.] = windowed second moments)
                    const float d_mu1 = 2.f * mu2 * (B - A) * inv - val * 2.f * mu1 / Cc + val * 2.f * mu1 / D;
                    const float d_e11 = -val / D;
                    const float d_e12 = 2.f * A * inv;
                    const size_t o3 = (size_t)c * plane + (size_t)gy * W + gx;
                    const size_t CHW = (size_t)C * plane;
                    maps[o3] = d_mu1; maps[CHW + o3] = d_e11; maps[2 * CHW + o3] = d_e12;
                }
            }
        }
    }
    l1_acc = warp_sum(l1_acc); ss_acc = warp_sum(ss_acc);
    if ((t & 31) == 0) { red[0][t >> 5] = l1_acc; red[1][t >> 5] = ss_acc; }
    __syncthreads();
    if (t == 0) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int k = 0; k < kLossThreads / 32; ++k) { a += red[0][k]; b += red[1][k]; }
        const size_t cta = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        partials[2 * cta] = a; partials[2 * cta + 1] = b;
    }
}

// fixed-order final sum (deterministic): out = {w_l1*L1 + w_ssim*SSIM + w_const, L1, SSIM}
__global__ void __launch_bounds__(1024)
l1_ssim_finalize_kernel(const float* __restrict__ partials, int n, double inv_count, float w_l1, float w_ssim, float w_const,
                        float* __restrict__ out) {
    __shared__ double ra[32], rb[32];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < n; i += 1024) { a += (double)partials[2 * i]; b += (double)partials[2 * i + 1]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
    if ((threadIdx.x & 31) == 0) { ra[threadIdx.x >> 5] = a; rb[threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        a = 0.0; b = 0.0;
        for (int k = 0; k < 32; ++k) { a += ra[k]; b += rb[k]; }
        const float l1 = (float)(a * inv_count), ss = (float)(b * inv_count);
        out[0] = w_l1 * l1 + w_ssim * ss + w_const; out[1] = l1; out[2] = ss;
    }
}

// dL/dimg(p) = g * [ w_ssim/N * ( conv(d_mu1)(p) + 2 x(p) conv(d_e11)(p) + y(p) conv(d_e12)(p) ) + w_l1/N * sign(x(p) - y(p)) ]
// (the window is symmetric, so the transposed convolution of autograd is the same correlation; maps are zero outside
// the image because no window is centred there).
__global__ void __launch_bounds__(kLossThreads)
l1_ssim_bwd_kernel(int C, int H, int W, const float* __restrict__ img, const float* __restrict__ gt, Window win,
                   const float* __restrict__ maps, const float* __restrict__ dloss, float c_l1, float c_ssim,
                   float* __restrict__ dimg) {
    __shared__ float sm[3][kIn][kIn + 1];
    __shared__ float hs[3][kIn][kTile + 1];
    const int c = blockIdx.z;
    const int x0 = blockIdx.x * kTile, y0 = blockIdx.y * kTile;
    const size_t plane = (size_t)H * W, CHW = (size_t)C * plane;
    const int t = threadIdx.x;
    const float g = dloss ? __ldg(dloss) : 1.f;

    if (c_ssim != 0.f) {
        const float* pm = maps + (size_t)c * plane;
        for (int i = t; i < kIn * kIn; i += kLossThreads) {
            const int r = i / kIn, q = i - r * kIn;
            const int gy = y0 + r - kHalo, gx = x0 + q - kHalo;
            float a = 0.f, b = 0.f, d = 0.f;
            if (gy >= 0 && gy < H && gx >= 0 && gx < W) {
                const size_t o = (size_t)gy * W + gx;
                a = __ldg(pm + o); b = __ldg(pm + CHW + o); d = __ldg(pm + 2 * CHW + o);
            }
            sm[0][r][q] = a; sm[1][r][q] = b; sm[2][r][q] = d;
        }
        __syncthreads();
        for (int i = t; i < kIn * (kTile / 4); i += kLossThreads) {
            const int r = i / (kTile / 4), cg = (i - r * (kTile / 4)) * 4;
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                float a[kWin + 3];
#pragma unroll
                for (int k = 0; k < kWin + 3; ++k) a[k] = sm[q][r][cg + k];
#pragma unroll
                for (int o = 0; o < 4; ++o) {
                    float acc = 0.f;
#pragma unroll
                    for (int k = 0; k < kWin; ++k) acc = fmaf(win.w[k], a[o + k], acc);
                    hs[q][r][cg + o] = acc;
                }
            }
        }
        __syncthreads();
    }
    const int col = t & 31, rg = (t >> 5) * 4;
    float s[3][4];
#pragma unroll
    for (int q = 0; q < 3; ++q)
#pragma unroll
        for (int o = 0; o < 4; ++o) s[q][o] = 0.f;
    if (c_ssim != 0.f) {
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            float v[kWin + 3];
#pragma unroll
            for (int k = 0; k < kWin + 3; ++k) v[k] = hs[q][rg + k][col];
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < kWin; ++k) acc = fmaf(win.w[k], v[o + k], acc);
                s[q][o] = acc;
            }
        }
    }
#pragma unroll
    for (int o = 0; o < 4; ++o) {
        const int gy = y0 + rg + o, gx = x0 + col;
        if (gy < H && gx < W) {
            const size_t o3 = (size_t)c * plane + (size_t)gy * W + gx;
            const float xv = __ldg(img + o3), yv = __ldg(gt + o3);
            const float d = xv - yv;
            const float sgn = (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f);
            const float gs = s[0][o] + 2.f * xv * s[1][o] + yv * s[2][o];
            dimg[o3] = g * (c_ssim * gs + c_l1 * sgn);
        }
    }
}

// ---------------------------------------------------------------- multi-tensor Adam
constexpr int kAdamMaxGroups = 8;
constexpr int kAdamThreads = 256;
constexpr int kAdamVecPerThread = 4;                       // float4 per thread
constexpr int kAdamChunk = kAdamThreads * kAdamVecPerThread * 4;   // floats per CTA

struct AdamPack {
    float* p[kAdamMaxGroups]; const float* g[kAdamMaxGroups]; float* m[kAdamMaxGroups]; float* v[kAdamMaxGroups];
    long long n[kAdamMaxGroups];
    float step_size[kAdamMaxGroups];      // lr / (1 - beta1^t)
    float bc2_sqrt[kAdamMaxGroups];       // sqrt(1 - beta2^t)
    int block_end[kAdamMaxGroups];        // exclusive end of the tensor's CTA range
    int count;
    float beta1, beta2, omb1, omb2, eps;   // omb = 1 - beta, rounded from double (1.f - 0.999f is off by 4.7e-5 relative)
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float omb1, float b2, float omb2, float eps, float step_size, float bc2_sqrt) {
    // torch/optim/adam.py (_single_tensor_adam / _multi_tensor_adam, no amsgrad, no weight decay, maximize=False):
    //   exp_avg.lerp_(grad, 1-beta1); exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2)
    //   denom = exp_avg_sq.sqrt() / bias_correction2_sqrt + eps;  param.addcdiv_(exp_avg, denom, value=-step_size)
    m = fmaf(omb1, g - m, m);
    v = fmaf(omb2, g * g, v * b2);
    const float denom = __fdiv_rn(__fsqrt_rn(v), bc2_sqrt) + eps;
    p = p - step_size * __fdiv_rn(m, denom);
}

__global__ void __launch_bounds__(kAdamThreads) adam_multi_kernel(const AdamPack a) {
    int t = 0;
    while (t < a.count - 1 && (int)blockIdx.x >= a.block_end[t]) ++t;
    const int first_block = t ? a.block_end[t - 1] : 0;
    const long long base = (long long)(blockIdx.x - first_block) * kAdamChunk;
    const long long n = a.n[t];
    float* __restrict__ p = a.p[t]; const float* __restrict__ g = a.g[t];
    float* __restrict__ m = a.m[t]; float* __restrict__ v = a.v[t];
    const float ss = a.step_size[t], bs = a.bc2_sqrt[t];
    const bool vec = ((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0);
    if (vec && base + kAdamChunk <= n) {
        float4 P4[kAdamVecPerThread], G4[kAdamVecPerThread], M4[kAdamVecPerThread], V4[kAdamVecPerThread];
#pragma unroll
        for (int k = 0; k < kAdamVecPerThread; ++k) {
            const long long i = base / 4 + (long long)k * kAdamThreads + threadIdx.x;
            P4[k] = ((const float4*)p)[i]; G4[k] = __ldcs((const float4*)g + i); M4[k] = ((const float4*)m)[i]; V4[k] = ((const float4*)v)[i];
        }
#pragma unroll
        for (int k = 0; k < kAdamVecPerThread; ++k) {
            adam_one(P4[k].x, G4[k].x, M4[k].x, V4[k].x, a.omb1, a.beta2, a.omb2, a.eps, ss, bs);
            adam_one(P4[k].y, G4[k].y, M4[k].y, V4[k].y, a.omb1, a.beta2, a.omb2, a.eps, ss, bs);
            adam_one(P4[k].z, G4[k].z, M4[k].z, V4[k].z, a.omb1, a.beta2, a.omb2, a.eps, ss, bs);
            adam_one(P4[k].w, G4[k].w, M4[k].w, V4[k].w, a.omb1, a.beta2, a.omb2, a.eps, ss, bs);
        }
#pragma unroll
        for (int k = 0; k < kAdamVecPerThread; ++k) {
            const long long i = base / 4 + (long long)k * kAdamThreads + threadIdx.x;
            ((float4*)p)[i] = P4[k]; ((float4*)m)[i] = M4[k]; ((float4*)v)[i] = V4[k];
        }
    } else {
        for (long long i = base + threadIdx.x; i < min(n, base + (long long)kAdamChunk); i += kAdamThreads) {
            float pp = p[i], mm = m[i], vv = v[i];
            adam_one(pp, g[i], mm, vv, a.omb1, a.beta2, a.omb2, a.eps, ss, bs);
            p[i] = pp; m[i] = mm; v[i] = vv;
        }
    }
}

}  // namespace

}  // namespace gsr

using namespace gsr;

namespace gsr {
// 4 pixels per thread: three 16-byte-ish channel reads (scalar when unaligned / at the tail), one 12-byte store group
__global__ void __launch_bounds__(256) pack_rgb8_kernel(const float* __restrict__ img, long long px, unsigned char* __restrict__ out) {
    const long long g = (long long)blockIdx.x * 256 + threadIdx.x;
    const long long p0 = g * 4;
    if (p0 >= px) return;
    const int n = (int)min(4LL, px - p0);
    unsigned char v[12];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float* src = img + (size_t)c * px + p0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float x = k < n ? __ldg(src + k) : 0.f;
            const float y = fminf(fmaxf(x, 0.f), 1.f) * 255.f;          // clamp(min=0,max=1): NaN -> propagates in torch; here NaN -> 0
            v[3 * k + c] = (unsigned char)(int)y;                        // .byte(): truncation toward zero
        }
    }
    unsigned char* dst = out + p0 * 3;
    if (n == 4 && ((size_t)dst & 3) == 0) {
        uint32_t* d4 = reinterpret_cast<uint32_t*>(dst);
        d4[0] = v[0] | (v[1] << 8) | (v[2] << 16) | ((uint32_t)v[3] << 24);
        d4[1] = v[4] | (v[5] << 8) | (v[6] << 16) | ((uint32_t)v[7] << 24);
        d4[2] = v[8] | (v[9] << 8) | (v[10] << 16) | ((uint32_t)v[11] << 24);
    } else {
        for (int k = 0; k < 3 * n; ++k) dst[k] = v[k];
    }
}
}  // namespace gsr

extern "C" {

size_t gsr_l1_ssim_partials(int channels, int height, int width) {
    if (channels <= 0 || height <= 0 || width <= 0) return 0;
    return (size_t)channels * ((height + kTile - 1) / kTile) * ((width + kTile - 1) / kTile);
}

int gsr_l1_ssim_forward(int channels, int height, int width, const float* image, const float* gt, float w_l1, float w_ssim,
                        float w_const, float* maps, float* partials, float* out, void* stream) {
    if (channels <= 0 || height <= 0 || width <= 0) { set_error("l1_ssim_forward: bad sizes"); return GSR_ERR_INVALID; }
    if (!image || !gt || !partials || !out) { set_error("l1_ssim_forward: null pointer"); return GSR_ERR_INVALID; }
    static const Window win = make_window();
    cudaStream_t s = (cudaStream_t)stream;
    dim3 grid((width + kTile - 1) / kTile, (height + kTile - 1) / kTile, channels);
    if (grid.z > 65535 || grid.y > 65535) { set_error("l1_ssim_forward: image too large"); return GSR_ERR_INVALID; }
    if (maps) l1_ssim_fwd_kernel<true><<<grid, kLossThreads, 0, s>>>(channels, height, width, image, gt, win, maps, partials);
    else l1_ssim_fwd_kernel<false><<<grid, kLossThreads, 0, s>>>(channels, height, width, image, gt, win, nullptr, partials);
    const int n = (int)gsr_l1_ssim_partials(channels, height, width);
    const double inv = 1.0 / ((double)channels * height * width);
    l1_ssim_finalize_kernel<<<1, 1024, 0, s>>>(partials, n, inv, w_l1, w_ssim, w_const, out);
    count_launch(2);
    return check_last("l1_ssim_forward", false, s);
}

int gsr_l1_ssim_backward(int channels, int height, int width, const float* image, const float* gt, float w_l1, float w_ssim,
                         const float* maps, const float* dloss, float* dL_dimage, void* stream) {
    if (channels <= 0 || height <= 0 || width <= 0) { set_error("l1_ssim_backward: bad sizes"); return GSR_ERR_INVALID; }
    if (!image || !gt || !dL_dimage || (w_ssim != 0.f && !maps)) { set_error("l1_ssim_backward: null pointer"); return GSR_ERR_INVALID; }
    static const Window win = make_window();
    cudaStream_t s = (cudaStream_t)stream;
    dim3 grid((width + kTile - 1) / kTile, (height + kTile - 1) / kTile, channels);
    const double inv = 1.0 / ((double)channels * height * width);
    l1_ssim_bwd_kernel<<<grid, kLossThreads, 0, s>>>(channels, height, width, image, gt, win, maps, dloss, (float)(w_l1 * inv),
                                                     (float)(w_ssim * inv), dL_dimage);
    count_launch();
    return check_last("l1_ssim_backward", false, s);
}

int gsr_adam_step(const gsr_adam_tensor* tensors, int count, double beta1, double beta2, double eps, void* stream) {
    if (count < 0 || (count > 0 && !tensors)) { set_error("adam_step: bad arguments"); return GSR_ERR_INVALID; }
    cudaStream_t s = (cudaStream_t)stream;
    int i = 0, launched = 0;
    while (i < count) {
        AdamPack a; a.count = 0; a.beta1 = (float)beta1; a.beta2 = (float)beta2; a.omb1 = (float)(1.0 - beta1); a.omb2 = (float)(1.0 - beta2); a.eps = (float)eps;
        long long blocks = 0;
        for (; i < count && a.count < kAdamMaxGroups; ++i) {
            const gsr_adam_tensor& t = tensors[i];
            if (t.n < 0 || t.step < 1) { set_error("adam_step: tensor %d: bad n/step", i); return GSR_ERR_INVALID; }
            if (t.n == 0) continue;
            if (!t.param || !t.grad || !t.exp_avg || !t.exp_avg_sq) { set_error("adam_step: tensor %d: null pointer", i); return GSR_ERR_INVALID; }
            const long long nb = (t.n + kAdamChunk - 1) / kAdamChunk;
            if (blocks + nb > 0x7fffffffLL) break;
            const int k = a.count++;
            a.p[k] = t.param; a.g[k] = t.grad; a.m[k] = t.exp_avg; a.v[k] = t.exp_avg_sq; a.n[k] = t.n;
            // the bias corrections in double like Python floats (torch: 1 - beta ** step), then one rounding to float
            const double bc1 = 1.0 - pow(beta1, (double)t.step), bc2 = 1.0 - pow(beta2, (double)t.step);
            a.step_size[k] = (float)(t.lr / bc1);
            a.bc2_sqrt[k] = (float)sqrt(bc2);
            blocks += nb; a.block_end[k] = (int)blocks;
        }
        if (a.count == 0) { if (i < count && blocks == 0) { set_error("adam_step: tensor too large"); return GSR_ERR_INVALID; } continue; }
        adam_multi_kernel<<<(unsigned)blocks, kAdamThreads, 0, s>>>(a);
        count_launch(); ++launched;
    }
    return launched ? check_last("adam_step", false, s) : GSR_OK;
}

// ---- presenter: float [3,H,W] frame -> the viewer's uint8 [H,W,3] layout.
// Replaces `(torch.clamp(net_image, min=0, max=1.0) * 255).byte().permute(1, 2, 0).contiguous()` (train.py:60): four passes
// over the frame in the reference, one here (12 B read + 3 B written per pixel).  `.byte()` truncates toward zero.
int gsr_pack_rgb8(const float* image_chw, int width, int height, unsigned char* out_hwc, void* stream) {
    if (width <= 0 || height <= 0 || !image_chw || !out_hwc) { set_error("pack_rgb8: bad arguments"); return GSR_ERR_INVALID; }
    const long long px = (long long)width * height;
    const long long groups = (px + 3) / 4;
    pack_rgb8_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, (cudaStream_t)stream>>>(image_chw, px, out_hwc);
    count_launch();
    return check_last("pack_rgb8", false, (cudaStream_t)stream);
}

}  // extern "C"
