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
// flags is written only by rank s.  Epochs only grow, so nothing is ever reset.  A rank that never arrives trips the
// watchdog (~2 s) instead of hanging the device: flags[world] (error word, local) is set and the kernel returns.
__global__ void peer_barrier_kernel(char* const* __restrict__ bases, size_t flag_offset, int world, int rank, uint32_t epoch) {
    const int p = threadIdx.x;
    if (p >= world) return;
    __threadfence_system();
    uint32_t* theirs = (uint32_t*)(bases[p] + flag_offset);
    uint32_t* mine = (uint32_t*)(bases[rank] + flag_offset);
    st_release_sys(theirs + rank, epoch);
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(mine + p) - epoch) < 0) {
        if (clock64() - t0 > 4000000000LL) { mine[world] = 1u; break; }
        __nanosleep(100);
    }
}

template <int WORLD>
__global__ void __launch_bounds__(kPeerThreads)
peer_reduce_kernel(char* const* __restrict__ bases, size_t data_offset, long long n4, int rank, int diag) {
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
peer_reduce_mc_kernel(float4* __restrict__ mc, long long n4, int world, int rank) {
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
    switch (world) {
        case 2: preload_reduce<2>(); break; case 3: preload_reduce<3>(); break; case 4: preload_reduce<4>(); break;
        case 5: preload_reduce<5>(); break; case 6: preload_reduce<6>(); break; case 7: preload_reduce<7>(); break;
        case 8: preload_reduce<8>(); break; default: break;
    }
    done[world] = true;
}

template <int WORLD>
void launch_reduce(char* const* bases, size_t off, long long n4, int rank, cudaStream_t s) {
    const long long per = (n4 + WORLD - 1) / WORLD;
    long long blocks = (per + 2LL * kPeerThreads - 1) / (2LL * kPeerThreads);
    blocks = std::max(1LL, std::min(blocks, 148LL * (g_peer_ctas_per_sm ? g_peer_ctas_per_sm : 2)));
    peer_reduce_kernel<WORLD><<<(unsigned)blocks, kPeerThreads, 0, s>>>(bases, off, n4, rank, g_peer_diag_mode);
}

}  // namespace
void set_peer_ctas_per_sm(int v) { g_peer_ctas_per_sm = v & 0xff; g_peer_diag_mode = (v >> 8) & 3; }
}  // namespace gsr

using namespace gsr;

extern "C" {

size_t gsr_peer_flag_bytes(int world) { return world <= 0 ? 0 : (size_t)((world + 1) * 4 + 255) / 256 * 256; }

int gsr_peer_barrier(void* const* peer_bases, size_t flag_offset, int world, int rank, unsigned epoch, void* stream) {
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || !peer_bases || (flag_offset & 3)) {
        set_error("peer_barrier: bad arguments (world %d, rank %d)", world, rank); return GSR_ERR_INVALID;
    }
    preload(world);
    peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>((char* const*)peer_bases, flag_offset, world, rank, epoch);
    count_launch();
    return check_last("peer_barrier", false, (cudaStream_t)stream);
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
    peer_barrier_kernel<<<1, 32, 0, s>>>(bases, flag_offset, world, rank, epoch);
    const long long n4 = n_floats / 4, per = (n4 + world - 1) / world;
    long long blocks = (per + 4LL * kPeerThreads - 1) / (4LL * kPeerThreads);
    blocks = std::max(1LL, std::min(blocks, 148LL * (g_peer_ctas_per_sm ? g_peer_ctas_per_sm : 1)));
    peer_reduce_mc_kernel<<<(unsigned)blocks, kPeerThreads, 0, s>>>((float4*)((char*)multicast_base + data_offset), n4, world, rank);
    peer_barrier_kernel<<<1, 32, 0, s>>>(bases, flag_offset, world, rank, epoch + 1u);
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
    peer_barrier_kernel<<<1, 32, 0, s>>>(bases, flag_offset, world, rank, epoch);
    const long long n4 = n_floats / 4;
    switch (world) {
        case 1: break;                                   // a single rank's sum is its own buffer
        case 2: launch_reduce<2>(bases, data_offset, n4, rank, s); break;
        case 3: launch_reduce<3>(bases, data_offset, n4, rank, s); break;
        case 4: launch_reduce<4>(bases, data_offset, n4, rank, s); break;
        case 5: launch_reduce<5>(bases, data_offset, n4, rank, s); break;
        case 6: launch_reduce<6>(bases, data_offset, n4, rank, s); break;
        case 7: launch_reduce<7>(bases, data_offset, n4, rank, s); break;
        default: launch_reduce<8>(bases, data_offset, n4, rank, s); break;
    }
    peer_barrier_kernel<<<1, 32, 0, s>>>(bases, flag_offset, world, rank, epoch + 1u);
    count_launch(world > 1 ? 3 : 2);
    return check_last("peer_allreduce_sum", false, s);
}

}  // extern "C"
