// gsr_peer.cu -- sum of the view-parallel gradients over NVLink peer memory, in place of the NCCL
// all-reduce of the 11 dense floats per Gaussian (means3D 3, opacity 1, scales 3, rotations 4).
//
// Every rank holds the same-sized buffer in memory its peers can address (the host maps them: torch symmetric memory,
// cudaIpc, or simply two buffers of one device in the tests) and passes the device array of the `world` base pointers.
// The sum is the two-shot exchange written as ONE data kernel between two flag barriers:
//     barrier     every rank's gradients are final (release/acquire flags in the peers' buffers, monotonic epochs)
//     reduce      rank r owns the r-th 1/world of the elements: it loads that slice from every peer (rank order, so the
//                 float sum is the same expression on every run and every rank receives the SAME bits), adds, and stores
//                 the result into every peer's buffer -- (world-1)/world of the payload in and out per GPU, full duplex
//     barrier     all stores have landed
// No staging copies, no ring: NVSwitch gives every pair full bandwidth, so the pull and the push of all slices run at once.
#include "gsr_common.cuh"

namespace gsr {

void preload_preprocess_backward();       // gsr_preprocess.cu

namespace {

constexpr int kMaxPeers = 8;
constexpr int kPeerThreads = 256;

struct PeerPtrs { float4* p[kMaxPeers]; };
int g_peer_ctas_per_sm = 0;
int g_peer_diag_mode = 0;   // diagnostics only (tools/peer_diag.py): 1 = remote loads, local stores; 2 = local loads, remote stores

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// One CTA, one thread per peer.  flags = `world` words inside every rank's buffer at flag_offset; word s of rank d's
// flags is written only by rank s.  Epochs only grow, so nothing is ever reset.
// A rank that has not arrived after the watchdog time (default 10 minutes, like a collective library's; a rank may be busy
// with a checkpoint, an evaluation pass or a cold start -- GSR_OPT_PEER_TIMEOUT_MS) must never lead to a silent sum of stale
// data: the barrier then (1) sets the local error word flags[world], which makes every data kernel behind it on this rank
// write NaN instead of sums, and (2) raises the process-wide host-visible error word that gsr_peer_error() reports -- the
// host shim checks it whenever it waits for an exchange.  The device is not hung and the context survives.
__global__ void peer_barrier_kernel(char* const* __restrict__ bases, size_t flag_offset, int world, int rank, uint32_t epoch,
                                    long long timeout_ns, uint32_t* host_err) {
    const int p = threadIdx.x;
    if (p >= world) return;
    __threadfence_system();
    uint32_t* theirs = (uint32_t*)(bases[p] + flag_offset);
    uint32_t* mine = (uint32_t*)(bases[rank] + flag_offset);
    st_release_sys(theirs + rank, epoch);
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    unsigned spins = 0;
    while ((int32_t)(ld_acquire_sys(mine + p) - epoch) < 0) {
        if ((++spins & 1023u) == 0u) {
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if ((long long)(t1 - t0) > timeout_ns) {
                mine[world] = 1u;
                if (host_err) { *((volatile uint32_t*)host_err) = 1u; __threadfence_system(); }
                break;
            }
        }
        __nanosleep(100);
    }
}

__device__ __forceinline__ bool peer_poisoned(char* const* bases, size_t flag_offset, int world, int rank) {
    return ((volatile const uint32_t*)(bases[rank] + flag_offset))[world] != 0u;
}

template <int WORLD>
__global__ void __launch_bounds__(kPeerThreads)
peer_reduce_kernel(char* const* __restrict__ bases, size_t data_offset, long long n4, int rank, int diag, size_t flag_offset) {
    const bool poisoned = peer_poisoned(bases, flag_offset, WORLD, rank);      // a barrier of this rank timed out
    const float nanv = __int_as_float(0x7fc00000);
    PeerPtrs P, Q;
#pragma unroll
    for (int k = 0; k < WORLD; ++k) {
        P.p[k] = (float4*)(bases[diag == 2 ? rank : k] + data_offset);       // load side
        Q.p[k] = (float4*)(bases[diag == 1 ? rank : k] + data_offset);       // store side
    }
    const long long per = (n4 + WORLD - 1) / WORLD;
    const long long lo = per * rank, hi = min(n4, lo + per);
    const long long stride = (long long)gridDim.x * kPeerThreads;
    for (long long i = lo + (long long)blockIdx.x * kPeerThreads + threadIdx.x; i < hi; i += 2 * stride) {
        const long long j = i + stride;
        const bool two = j < hi;
        float4 a[WORLD], b[WORLD];
#pragma unroll
        for (int k = 0; k < WORLD; ++k) {
            a[k] = __ldcg(P.p[k] + i);
            if (two) b[k] = __ldcg(P.p[k] + j);
        }
        float4 s = a[0], t = two ? b[0] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 1; k < WORLD; ++k) {
            s.x += a[k].x; s.y += a[k].y; s.z += a[k].z; s.w += a[k].w;
            if (two) { t.x += b[k].x; t.y += b[k].y; t.z += b[k].z; t.w += b[k].w; }
        }
        if (poisoned) { s = make_float4(nanv, nanv, nanv, nanv); t = s; }
#pragma unroll
        for (int k = 0; k < WORLD; ++k) {
            __stcg(Q.p[k] + i, s);
            if (two) __stcg(Q.p[k] + j, t);
        }
    }
    __threadfence_system();
}

// NVSwitch variant: the same slice ownership, but the sum is formed INSIDE the switch (multimem.ld_reduce pulls one reduced
// copy of the slice instead of `world` copies) and one multimem.st delivers it to every rank: (1 + 1/world) x payload per
// direction and GPU instead of 2 (world-1)/world.  `mc` is the multicast mapping of the same buffers (cuMulticast*, set up
// by the host).  The switch fixes the order of its adds, so all ranks still receive identical bits, but that order is not
// the rank order of the unicast kernel (last-bit differences between the two variants).
__global__ void __launch_bounds__(kPeerThreads)
peer_reduce_mc_kernel(float4* __restrict__ mc, long long n4, int world, int rank, char* const* __restrict__ bases, size_t flag_offset) {
    const bool poisoned = peer_poisoned(bases, flag_offset, world, rank);
    const float nanv = __int_as_float(0x7fc00000);
    const long long per = (n4 + world - 1) / world;
    const long long lo = per * rank, hi = min(n4, lo + per);
    const long long stride = (long long)gridDim.x * kPeerThreads;
    for (long long i = lo + (long long)blockIdx.x * kPeerThreads + threadIdx.x; i < hi; i += 4 * stride) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const long long j = i + u * stride;
            if (j < hi)
                asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                             : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(mc + j) : "memory");
            if (poisoned) v[u] = make_float4(nanv, nanv, nanv, nanv);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const long long j = i + u * stride;
            if (j < hi)
                asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                             :: "l"(mc + j), "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w) : "memory");
        }
    }
    __threadfence_system();
}

long long g_peer_timeout_ms = 600000;
long long peer_timeout_ns() { return g_peer_timeout_ms * 1000000LL; }
// process-wide error word in mapped pinned host memory (portable: valid on every device under unified addressing)
uint32_t* peer_error_word() {
    static uint32_t* w = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        if (cudaHostAlloc((void**)&w, 64, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) { w = nullptr; cudaGetLastError(); }
        else *w = 0u;
    }
    return w;
}

// ---- fused exchange, second half (gsr_peer_owner_sum): a STREAMING consumer.  It runs beside preprocess_bwd_kernel (its own
// stream, a few dozen resident CTAs) and follows it block by block.  A 256-Gaussian block travels as 94 self-validating 128-byte
// lines (30 floats + a 2-word epoch flag, written whole by 8 lanes of one store: see the producer): a CTA loads line e of the
// block from all `world` slots, re-loads the ones whose flag is not this epoch yet, adds them in rank order (the same float
// expression on every run, so every rank receives the SAME bits; rows of invisible Gaussians were staged as zeros, exactly the
// terms the dense per-view tensors hold), gathers the [256][11] totals in shared memory and writes every array of the dense
// block with 128-bit stores: means3D 3 x float4 per 4 Gaussians, opacity 1, scales 3, rotations 4 -- to all ranks at once.
// When the producers finish only the last few blocks and the closing barrier are exposed.
struct OwnerSumParams {
    char* const* bases; float* mc;            // mc != nullptr: multicast mapping of the buffers (one multimem.st per value)
    size_t staging_offset, final_offset, flag_offset;
    int P4, rows_per_rank, world, rank;
    uint32_t epoch; long long timeout_ns; uint32_t* host_err;
};

template <bool MC>
__device__ __forceinline__ void deliver(const OwnerSumParams& p, size_t float_index, float4 v) {
    if (MC) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                     :: "l"(p.mc + (p.final_offset >> 2) + float_index), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    } else {
        for (int k = 0; k < p.world; ++k) __stcg(reinterpret_cast<float4*>(p.bases[k] + p.final_offset) + (float_index >> 2), v);
    }
}

__device__ __forceinline__ float4 ld_line_piece(const float4* p) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

// MAXW = compile-time bound of world (2, 4 or 8), DEPTH = 256-piece chunks of a block whose loads are issued before the first is
// validated (3 = the whole block: three times fewer exposed load latencies per block, used at world 2 where an owner has the most
// blocks).  64 registers at most: a consumer CTA then displaces exactly one CTA of the backward beside it.
template <bool MC, int MAXW, int DEPTH>
__global__ void __launch_bounds__(kPeerThreads, 4) peer_owner_sum_kernel(const OwnerSumParams p) {
    __shared__ float s_f[2816 + 8];             // the block's totals, [256 Gaussians][11 floats]
    __shared__ int s_bad;
    const int tid = threadIdx.x, lane = tid & 31;
    const float nanv = __int_as_float(0x7fc00000);
    const float4* staging = reinterpret_cast<const float4*>(p.bases[p.rank] + p.staging_offset);
    uint32_t* mine = reinterpret_cast<uint32_t*>(p.bases[p.rank] + p.flag_offset);
    const size_t P4 = (size_t)p.P4;
    const int nblk = p.rows_per_rank >> 8;
    const size_t slot = (size_t)nblk * kPushSlotF4;                       // float4s between two ranks' copies of a block
    const bool flag_lane = (lane & 7) == 7;
    if (tid == 0) s_bad = 0;
    bool bad = false;                           // this warp's watchdog fired: it stops waiting for good
    // this rank owns the 256-Gaussian blocks rank, rank + world, rank + 2 world, ... (block-cyclic, see preprocess_bwd_kernel);
    // local block j <-> Gaussians ((j * world + rank) * 256 ...)
    for (int j = blockIdx.x; j < nblk; j += gridDim.x) {
        const long long idx0 = ((long long)j * p.world + p.rank) << 8;
        if (idx0 >= (long long)P4) break;
        __syncthreads();                                                  // s_f of the previous block has been read
        for (int e0 = 0; e0 < kPushLines * 8; e0 += kPeerThreads * DEPTH) {       // whole warps stay: the line votes are warp-wide
            float4 v[DEPTH][MAXW];
            const float4* src[DEPTH];
            bool act[DEPTH];
#pragma unroll
            for (int c = 0; c < DEPTH; ++c) {
                const int e = e0 + c * kPeerThreads + tid;
                act[c] = e < kPushLines * 8;
                src[c] = staging + (size_t)j * kPushSlotF4 + (act[c] ? e : 0);
#pragma unroll
                for (int k = 0; k < MAXW; ++k)
                    if (k < p.world) v[c][k] = ld_line_piece(src[c] + k * slot);
            }
#pragma unroll
            for (int c = 0; c < DEPTH; ++c) {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int k = 0; k < MAXW; ++k) {
                    if (k >= p.world) break;
                    // a line is valid when the flag its 8th piece carries is this epoch; the 8 pieces come from ONE load instruction
                    unsigned long long t0 = 0;
                    unsigned spins = 0;
                    for (;;) {
                        const bool ok = !act[c] || !flag_lane || (__float_as_uint(v[c][k].z) == p.epoch && __float_as_uint(v[c][k].w) == p.epoch);
                        if (__all_sync(0xffffffffu, ok) || bad) break;
                        if ((++spins & 63u) == 0u) {
                            unsigned long long t1;
                            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                            if (t0 == 0) t0 = t1;
                            else if ((long long)(t1 - t0) > p.timeout_ns) bad = true;      // never a silent sum of stale rows
                        }
                        bad = __any_sync(0xffffffffu, bad);
                        if (bad) break;
                        __nanosleep(500);
                        v[c][k] = ld_line_piece(src[c] + k * slot);        // the whole warp re-loads: pieces and flag stay one load
                    }
                    acc.x += v[c][k].x; acc.y += v[c][k].y; acc.z += v[c][k].z; acc.w += v[c][k].w;      // rank order
                }
                if (act[c]) {
                    const int e = e0 + c * kPeerThreads + tid;
                    const int f0 = (e >> 3) * 30 + (e & 7) * 4;
                    s_f[f0] = acc.x; s_f[f0 + 1] = acc.y;                 // (the last line's tail lands in the 8 pad floats)
                    if (!flag_lane) { s_f[f0 + 2] = acc.z; s_f[f0 + 3] = acc.w; }
                }
            }
            if (bad && lane == 0) {
                mine[p.world] = 1u;
                if (p.host_err) { *((volatile uint32_t*)p.host_err) = 1u; __threadfence_system(); }
                s_bad = 1;
            }
        }
        __syncthreads();
        const bool poisoned = s_bad != 0 || ((volatile const uint32_t*)mine)[p.world] != 0u;
        // dense block: [means3D 3 x P4 | opacity P4 | scales 3 x P4 | rotations 4 x P4]; this block's n Gaussians (multiple of 4)
        const int n = (int)min((long long)256, (long long)P4 - idx0);
        const int n34 = 3 * n / 4, n4 = n / 4;
        const size_t i0 = (size_t)idx0;
        for (int t = tid; t < 2 * n34 + n4 + n; t += kPeerThreads) {
            float4 o;
            size_t at;
            if (t < 2 * n34) {                                             // means3D (columns 0..2) or scales (columns 4..6)
                const bool sc = t >= n34;
                const int m = sc ? t - n34 : t, c0 = sc ? 4 : 0;
                float v[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) { const int el = 4 * m + q; v[q] = s_f[(el / 3) * 11 + c0 + el % 3]; }
                o = make_float4(v[0], v[1], v[2], v[3]); at = (sc ? 4 * P4 : 0) + 3 * i0 + 4 * (size_t)m;
            } else if (t < 2 * n34 + n4) {                                 // opacity (column 3)
                const int m = t - 2 * n34;
                o = make_float4(s_f[(4 * m) * 11 + 3], s_f[(4 * m + 1) * 11 + 3], s_f[(4 * m + 2) * 11 + 3], s_f[(4 * m + 3) * 11 + 3]);
                at = 3 * P4 + i0 + 4 * (size_t)m;
            } else {                                                       // rotations (columns 7..10)
                const int m = t - 2 * n34 - n4;
                o = make_float4(s_f[m * 11 + 7], s_f[m * 11 + 8], s_f[m * 11 + 9], s_f[m * 11 + 10]);
                at = 7 * P4 + 4 * (i0 + m);
            }
            if (poisoned) o = make_float4(nanv, nanv, nanv, nanv);
            deliver<MC>(p, at, o);
        }
    }
    __threadfence_system();
}

// Lazy module loading resolves a kernel at its first launch and may wait for the device to drain -- behind a barrier
// kernel that is itself waiting for a peer this host thread has yet to launch, that is a stall until the watchdog.  So
// every kernel of this file is resolved before the first barrier is queued.
template <int WORLD>
void preload_reduce() {
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, peer_reduce_kernel<WORLD>);
}
void preload(int world) {
    static bool done[kMaxPeers + 1] = {};
    if (done[world]) return;
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, peer_barrier_kernel);
    cudaFuncGetAttributes(&fa, peer_reduce_mc_kernel);
    cudaFuncGetAttributes(&fa, peer_owner_sum_kernel<true, 2, 3>);
    cudaFuncGetAttributes(&fa, peer_owner_sum_kernel<false, 2, 3>);
    cudaFuncGetAttributes(&fa, peer_owner_sum_kernel<true, 4, 1>);
    cudaFuncGetAttributes(&fa, peer_owner_sum_kernel<false, 4, 1>);
    cudaFuncGetAttributes(&fa, peer_owner_sum_kernel<true, 8, 1>);
    cudaFuncGetAttributes(&fa, peer_owner_sum_kernel<false, 8, 1>);
    preload_preprocess_backward();            // the producer the streaming owner sum waits for
    switch (world) {
        case 2: preload_reduce<2>(); break; case 3: preload_reduce<3>(); break; case 4: preload_reduce<4>(); break;
        case 5: preload_reduce<5>(); break; case 6: preload_reduce<6>(); break; case 7: preload_reduce<7>(); break;
        case 8: preload_reduce<8>(); break; default: break;
    }
    done[world] = true;
}

template <int WORLD>
void launch_reduce(char* const* bases, size_t off, long long n4, int rank, size_t flag_off, cudaStream_t s) {
    const long long per = (n4 + WORLD - 1) / WORLD;
    long long blocks = (per + 2LL * kPeerThreads - 1) / (2LL * kPeerThreads);
    blocks = std::max(1LL, std::min(blocks, 148LL * (g_peer_ctas_per_sm ? g_peer_ctas_per_sm : 2)));
    peer_reduce_kernel<WORLD><<<(unsigned)blocks, kPeerThreads, 0, s>>>(bases, off, n4, rank, g_peer_diag_mode, flag_off);
}

}  // namespace
void set_peer_ctas_per_sm(int v) { g_peer_ctas_per_sm = v & 0xff; g_peer_diag_mode = (v >> 8) & 3; }
void set_peer_timeout_ms(int v) { g_peer_timeout_ms = v > 0 ? v : 600000; }
}  // namespace gsr

using namespace gsr;

extern "C" {

size_t gsr_peer_flag_bytes(int world) { return world <= 0 ? 0 : (size_t)((world + 1) * 4 + 255) / 256 * 256; }

int gsr_peer_barrier(void* const* peer_bases, size_t flag_offset, int world, int rank, unsigned epoch, void* stream) {
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || !peer_bases || (flag_offset & 3)) {
        set_error("peer_barrier: bad arguments (world %d, rank %d)", world, rank); return GSR_ERR_INVALID;
    }
    preload(world);
    peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>((char* const*)peer_bases, flag_offset, world, rank, epoch, peer_timeout_ns(), peer_error_word());
    count_launch();
    return check_last("peer_barrier", false, (cudaStream_t)stream);
}

int gsr_peer_error(void) {
    uint32_t* w = peer_error_word();
    if (!w) return 0;
    const uint32_t v = *((volatile uint32_t*)w);
    if (v) *((volatile uint32_t*)w) = 0u;
    return v ? 1 : 0;
}

int gsr_peer_owner_sum(void* const* peer_bases, void* multicast_base, size_t staging_offset, size_t final_offset,
                       int P, int rows_per_rank, size_t flag_offset, int world, int rank, unsigned epoch,
                       int ctas, void* stream) {
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || !peer_bases) {
        set_error("peer_owner_sum: bad arguments (world %d, rank %d)", world, rank); return GSR_ERR_INVALID;
    }
    if (P < 0 || rows_per_rank <= 0 || (rows_per_rank % 256) != 0 || (long long)rows_per_rank * world < (long long)P ||
        (staging_offset & 127) || (final_offset & 15) || (flag_offset & 3) || ((uintptr_t)multicast_base & 15) || epoch == 0) {
        set_error("peer_owner_sum: bad sizes / alignment (P %d, rows per rank %d)", P, rows_per_rank); return GSR_ERR_INVALID;
    }
    cudaStream_t s = (cudaStream_t)stream;
    char* const* bases = (char* const*)peer_bases;
    preload(world);
    // opening barrier: every rank is past the last reader of the totals of the previous exchange (the broadcasts below overwrite them)
    peer_barrier_kernel<<<1, 32, 0, s>>>(bases, flag_offset, world, rank, epoch, peer_timeout_ns(), peer_error_word());
    OwnerSumParams p;
    p.bases = bases; p.mc = (float*)multicast_base; p.staging_offset = staging_offset; p.final_offset = final_offset;
    p.flag_offset = flag_offset;
    p.P4 = (P + 3) / 4 * 4; p.rows_per_rank = rows_per_rank; p.world = world; p.rank = rank;
    p.epoch = epoch; p.timeout_ns = peer_timeout_ns(); p.host_err = peer_error_word();
    const int mine = (int)((((long long)p.P4 + 255) / 256 + world - 1 - rank) / world);       // blocks this rank owns
    // resident consumer CTAs: every one of them takes an occupancy slot from the backward beside it (a register-file quarter of
    // its SM: measured +0.13 ms on the 0.72 ms kernel with 74, +0.3 ms with 148), so as few as keep up with the producers:
    // one per ~160 owned blocks, 37..74
    int blocks = ctas > 0 ? ctas : (g_peer_ctas_per_sm ? 148 * g_peer_ctas_per_sm : std::max(37, std::min(74, mine / 160)));
    blocks = std::max(1, std::min(blocks, std::max(1, mine)));
    if (world <= 2) {
        if (multicast_base) peer_owner_sum_kernel<true, 2, 3><<<(unsigned)blocks, kPeerThreads, 0, s>>>(p);
        else peer_owner_sum_kernel<false, 2, 3><<<(unsigned)blocks, kPeerThreads, 0, s>>>(p);
    } else if (world <= 4) {
        if (multicast_base) peer_owner_sum_kernel<true, 4, 1><<<(unsigned)blocks, kPeerThreads, 0, s>>>(p);
        else peer_owner_sum_kernel<false, 4, 1><<<(unsigned)blocks, kPeerThreads, 0, s>>>(p);
    } else {
        if (multicast_base) peer_owner_sum_kernel<true, 8, 1><<<(unsigned)blocks, kPeerThreads, 0, s>>>(p);
        else peer_owner_sum_kernel<false, 8, 1><<<(unsigned)blocks, kPeerThreads, 0, s>>>(p);
    }
    peer_barrier_kernel<<<1, 32, 0, s>>>(bases, flag_offset, world, rank, epoch + 1u, peer_timeout_ns(), peer_error_word());
    count_launch(3);
    return check_last("peer_owner_sum", false, s);
}

int gsr_peer_allreduce_sum_mc(void* const* peer_bases, void* multicast_base, size_t data_offset, long long n_floats,
                              size_t flag_offset, int world, int rank, unsigned epoch, void* stream) {
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || !peer_bases || !multicast_base) {
        set_error("peer_allreduce_sum_mc: bad arguments (world %d, rank %d)", world, rank); return GSR_ERR_INVALID;
    }
    if (n_floats < 0 || (n_floats & 3) || (data_offset & 15) || (flag_offset & 3) || ((uintptr_t)multicast_base & 15)) {
        set_error("peer_allreduce_sum_mc: n_floats must be a multiple of 4 and the offsets 16-byte aligned"); return GSR_ERR_INVALID;
    }
    if (n_floats == 0) return GSR_OK;
    cudaStream_t s = (cudaStream_t)stream;
    char* const* bases = (char* const*)peer_bases;
    preload(world);
    peer_barrier_kernel<<<1, 32, 0, s>>>(bases, flag_offset, world, rank, epoch, peer_timeout_ns(), peer_error_word());
    const long long n4 = n_floats / 4, per = (n4 + world - 1) / world;
    long long blocks = (per + 4LL * kPeerThreads - 1) / (4LL * kPeerThreads);
    blocks = std::max(1LL, std::min(blocks, 148LL * (g_peer_ctas_per_sm ? g_peer_ctas_per_sm : 1)));
    peer_reduce_mc_kernel<<<(unsigned)blocks, kPeerThreads, 0, s>>>((float4*)((char*)multicast_base + data_offset), n4, world, rank, bases, flag_offset);
    peer_barrier_kernel<<<1, 32, 0, s>>>(bases, flag_offset, world, rank, epoch + 1u, peer_timeout_ns(), peer_error_word());
    count_launch(3);
    return check_last("peer_allreduce_sum_mc", false, s);
}

int gsr_peer_allreduce_sum(void* const* peer_bases, size_t data_offset, long long n_floats, size_t flag_offset, int world,
                           int rank, unsigned epoch, void* stream) {
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || !peer_bases) {
        set_error("peer_allreduce_sum: bad arguments (world %d, rank %d)", world, rank); return GSR_ERR_INVALID;
    }
    if (n_floats < 0 || (n_floats & 3) || (data_offset & 15) || (flag_offset & 3)) {
        set_error("peer_allreduce_sum: n_floats must be a multiple of 4 and the offsets 16-byte aligned"); return GSR_ERR_INVALID;
    }
    if (n_floats == 0) return GSR_OK;
    cudaStream_t s = (cudaStream_t)stream;
    char* const* bases = (char* const*)peer_bases;
    preload(world);
    peer_barrier_kernel<<<1, 32, 0, s>>>(bases, flag_offset, world, rank, epoch, peer_timeout_ns(), peer_error_word());
    const long long n4 = n_floats / 4;
    switch (world) {
        case 1: break;                                   // a single rank's sum is its own buffer
        case 2: launch_reduce<2>(bases, data_offset, n4, rank, flag_offset, s); break;
        case 3: launch_reduce<3>(bases, data_offset, n4, rank, flag_offset, s); break;
        case 4: launch_reduce<4>(bases, data_offset, n4, rank, flag_offset, s); break;
        case 5: launch_reduce<5>(bases, data_offset, n4, rank, flag_offset, s); break;
        case 6: launch_reduce<6>(bases, data_offset, n4, rank, flag_offset, s); break;
        case 7: launch_reduce<7>(bases, data_offset, n4, rank, flag_offset, s); break;
        default: launch_reduce<8>(bases, data_offset, n4, rank, flag_offset, s); break;
    }
    peer_barrier_kernel<<<1, 32, 0, s>>>(bases, flag_offset, world, rank, epoch + 1u, peer_timeout_ns(), peer_error_word());
    count_launch(world > 1 ? 3 : 2);
    return check_last("peer_allreduce_sum", false, s);
}

}  // extern "C"
