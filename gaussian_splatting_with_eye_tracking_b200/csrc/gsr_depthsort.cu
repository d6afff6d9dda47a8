// gsr_depthsort.cu -- depth ordering of the visible Gaussians for sm_100a: one counting sort into fine depth bins, then a
// tiny exact sort inside every bin.
//
// First half of the hierarchical replacement of cub::DeviceRadixSort::SortPairs<u64,u32> on (tile << 32 | depth) keys
// (cuda_rasterizer/rasterizer_impl.cu:303-308): the VISIBLE Gaussians are ordered by (depth bits, index) -- the order in
// which CUB's stable LSD passes leave equal-tile pairs (ties on equal depth keep the ascending index of duplicateWithKeys,
// rasterizer_impl.cu:70-111) -- and the run emission / row partition / expansion of gsr_binning.cu turn that order into the
// per-tile lists.
//
// Round 1 ran 4 global onesweep passes over the ~1.4 M keys of a 1080p frame (+ block scan, compaction, histogram, block
// sums, a second scan: ten dependent launches, 0.15 ms).  A radix pass costs 13-19 us at this size whether it runs in
// shared or global memory (the match-based ranking dominates), so the passes are replaced altogether:
//
//   preprocess_fwd_kernel  (gsr_preprocess.cu) leaves, per 256-Gaussian block, the (depth key, id, tile rows) of its visible
//                          Gaussians compacted to the front of the block's segment, the block's visible count, and the key
//                          range / pair and run totals in 128 spread statistics slots (no single hot atomic address:
//                          same-address global atomics serialise at ~30 ns each on this part)
//   bin_hist_kernel        bin = (key - key_min) >> shift over depth_bins(P) bins: two REDs per Gaussian (count, tile rows);
//                          the CTA that finishes last scans both histograms (bin start, first run of the bin), checks the
//                          largest bin and posts the frame totals to the host mailbox
//   bin_scatter_kernel     one 16-byte record per Gaussian into its bin's segment, slot = atomicAdd on the bin cursor
//                          (~20 Gaussians per address; order inside a bin: arbitrary)
//   bin_sort_kernel        a warp per bin: exact rank of every record by all-pairs comparison of the 64-bit composite
//                          (depth key, id) -- bins hold a few dozen records -- then the warp's prefix of the tile rows
//                          gives every Gaussian the absolute offset of its first run.  Replaces the remaining radix
//                          passes AND order_block_sums + scan.
//
// A bin with more than kBinCap records (thousands of Gaussians within 1/depth_bins of the depth range, e.g. all-equal
// depths) is not sorted by one warp: bin_hist_kernel then sets sort_mode = 2, the two kernels behind it leave at once, and
// the host runs the round-1 onesweep chain instead (gsr_api.cu), so that everything downstream sees `order` + `runoff`.
#include "gsr_common.cuh"

namespace gsr {

static int ds_num_sms() {
    static int n = 0;
    if (n == 0) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); if (n <= 0) n = 148; }
    return n;
}

// ------------------------------------------------------------------------------------------------
// histogram + scan
// ------------------------------------------------------------------------------------------------
constexpr int kHistThreads = 1024;

__global__ void __launch_bounds__(kHistThreads) bin_hist_kernel(GeomHeader* hdr, const FrameStatSlot* __restrict__ slots, int P, int nblocks,
                                                                const uint32_t* __restrict__ block_vis, const uint32_t* __restrict__ ckeys,
                                                                const uint32_t* __restrict__ crows, uint32_t* bin_cur, uint32_t* bin_runs,
                                                                int nbins, uint32_t* runoff, uint32_t* mailbox) {
    __shared__ uint32_t s_a[32], s_b[32], s_c[32], s_d[32];
    __shared__ uint32_t s_kmin, s_shift, s_last;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;

    // ---- key range and totals from the statistics slots (every CTA, redundantly: 128 loads)
    uint32_t kmx = 0, kmn = 0, npairs = 0, trap = 0;
    if (tid < kStatSlots) { const FrameStatSlot s = slots[tid]; kmx = s.key_max; kmn = s.key_min_inv; npairs = s.num_rendered; trap = s.flags; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        kmx = max(kmx, __shfl_xor_sync(0xffffffffu, kmx, o)); kmn = max(kmn, __shfl_xor_sync(0xffffffffu, kmn, o));
        npairs += __shfl_xor_sync(0xffffffffu, npairs, o); trap |= __shfl_xor_sync(0xffffffffu, trap, o);
    }
    if (lane == 0) { s_a[wid] = kmx; s_b[wid] = kmn; s_c[wid] = npairs; s_d[wid] = trap; }
    __syncthreads();
    if (tid == 0) {
        uint32_t a = 0, b = 0;
        for (int w = 0; w < kStatSlots / 32; ++w) { a = max(a, s_a[w]); b = max(b, s_b[w]); }
        const uint32_t kmin = ~b;
        const uint32_t range = (b != 0u && a >= kmin) ? a - kmin : 0u;        // b == 0: nothing visible
        int nb_log = 0; while ((1 << nb_log) < nbins) ++nb_log;
        const int bits = 32 - __clz(range);
        s_kmin = kmin; s_shift = bits > nb_log ? (uint32_t)(bits - nb_log) : 0u;
    }
    __syncthreads();
    const uint32_t kmin = s_kmin, shift = s_shift;
    uint32_t tot_pairs = 0, tot_trap = 0;
    for (int w = 0; w < kStatSlots / 32; ++w) { tot_pairs += s_c[w]; tot_trap |= s_d[w]; }

    // ---- two REDs per visible Gaussian.  Four 256-blocks per iteration; a block's records sit at the front of its segment.
    const int sub = tid >> 8, t = tid & 255;
    for (int b = (int)blockIdx.x * 4 + sub; b < nblocks; b += (int)gridDim.x * 4) {
        const uint32_t cnt = __ldg(block_vis + b);
        if ((uint32_t)t < cnt) {
            const size_t i = (size_t)b * kPreBlock + t;
            const uint32_t bin = (__ldg(ckeys + i) - kmin) >> shift;
            atomicAdd(&bin_cur[bin], 1u);
            atomicAdd(&bin_runs[bin], __ldg(crows + i));
        }
    }
    __syncthreads();
    if (tid == 0) { __threadfence(); s_last = (atomicAdd(&hdr->done_hist, 1u) == gridDim.x - 1) ? 1u : 0u; }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

    // ---- last CTA: exclusive scans of both histograms (in place), largest bin, totals.  Every thread owns nbins / 1024
    // consecutive bins: one pass for the sums, one block scan, one pass that writes the prefixes.
    const int per = nbins / kHistThreads;               // nbins is a power of two >= 1024
    const int i0 = tid * per;
    uint32_t sc = 0, sr = 0, mx = 0;
    for (int i = 0; i < per; ++i) { const uint32_t c = __ldcg(bin_cur + i0 + i); sc += c; sr += __ldcg(bin_runs + i0 + i); mx = max(mx, c); }
    uint32_t ic = sc, ir = sr;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, ic, o), v = __shfl_up_sync(0xffffffffu, ir, o);
        if (lane >= o) { ic += u; ir += v; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    __syncthreads();
    if (lane == 31) { s_a[wid] = ic; s_b[wid] = ir; }
    if (lane == 0) s_c[wid] = mx;
    __syncthreads();
    uint32_t bc = 0, br = 0, V = 0, R = 0, maxbin = 0;
#pragma unroll
    for (int w = 0; w < 32; ++w) {
        const uint32_t a = s_a[w], b = s_b[w];
        bc += (w < wid) ? a : 0u; br += (w < wid) ? b : 0u; V += a; R += b; maxbin = max(maxbin, s_c[w]);
    }
    uint32_t rc = bc + ic - sc, rr = br + ir - sr;
    for (int i = 0; i < per; ++i) {
        const uint32_t c = __ldcg(bin_cur + i0 + i), r = __ldcg(bin_runs + i0 + i);
        bin_cur[i0 + i] = rc; bin_runs[i0 + i] = rr;   // bin start (the scatter's cursor) and the bin's first run
        rc += c; rr += r;
    }
    if (tid == 0) {
        const bool ok = maxbin <= (uint32_t)kBinCap;
        const uint32_t err = tot_trap ? 0x100u : 0u;
        hdr->magic = kGeomMagic; hdr->P = (uint32_t)P; hdr->num_rendered = tot_pairs; hdr->num_runs = R; hdr->num_visible = V;
        hdr->sort_error = err; hdr->key_min = kmin; hdr->shift = shift; hdr->sort_mode = ok ? 1u : 2u; hdr->max_bin = maxbin;
        runoff[V] = R;                                 // read by the run emission of the last block of sorted positions
        if (mailbox) {   // mapped pinned host memory: the host reads the totals without a DMA copy; posted PCIe writes stay
                         // ordered and the host reads after an event recorded behind this kernel
            volatile uint32_t* m = mailbox;
            m[0] = kGeomMagic; m[1] = (uint32_t)P; m[2] = tot_pairs; m[3] = err; m[4] = R; m[5] = V; m[6] = ok ? 1u : 2u; m[7] = maxbin;
        }
    }
}

void launch_bin_hist(const GeomPtrs& g, int P, uint32_t* mailbox_dev, cudaStream_t s) {
    int grid = (g.nblocks + 3) / 4;
    if (grid > ds_num_sms()) grid = ds_num_sms();
    if (grid < 1) grid = 1;
    bin_hist_kernel<<<grid, kHistThreads, 0, s>>>(g.header, g.stat_slots, P, g.nblocks, g.block_sums_v, g.dkeys_a, g.dkeys_b, g.bin_cur, g.bin_runs,
                                                  g.nbins, g.runoff, mailbox_dev);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// scatter into the bins' segments
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kHistThreads) bin_scatter_kernel(const GeomHeader* __restrict__ hdr, int nblocks,
                                                                   const uint32_t* __restrict__ block_vis, const uint32_t* __restrict__ ckeys,
                                                                   const uint32_t* __restrict__ cids, const uint32_t* __restrict__ crows,
                                                                   uint32_t* bin_cur, uint4* __restrict__ brec) {
    if (hdr->sort_mode != 1u) return;
    const uint32_t kmin = hdr->key_min, shift = hdr->shift;
    const int sub = threadIdx.x >> 8, t = threadIdx.x & 255;
    for (int b = (int)blockIdx.x * 4 + sub; b < nblocks; b += (int)gridDim.x * 4) {
        const uint32_t cnt = __ldg(block_vis + b);
        if ((uint32_t)t < cnt) {
            const size_t i = (size_t)b * kPreBlock + t;
            const uint32_t key = __ldg(ckeys + i);
            const uint32_t pos = atomicAdd(&bin_cur[(key - kmin) >> shift], 1u);
            brec[pos] = make_uint4(key, __ldg(cids + i), __ldg(crows + i), 0u);
        }
    }
}

void launch_bin_scatter(const GeomPtrs& g, int P, cudaStream_t s) {
    int grid = (g.nblocks + 3) / 4;
    const int cap = ds_num_sms() * 2;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    bin_scatter_kernel<<<grid, kHistThreads, 0, s>>>(g.header, g.nblocks, g.block_sums_v, g.dkeys_a, g.order_tmp, g.dkeys_b, g.bin_cur, g.brec);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// exact order inside every bin + run offsets
// ------------------------------------------------------------------------------------------------
constexpr int kBinSortWarps = 8;

__global__ void __launch_bounds__(kBinSortWarps * 32) bin_sort_kernel(const GeomHeader* __restrict__ hdr, const uint32_t* __restrict__ bin_end,
                                                                      const uint32_t* __restrict__ bin_runs, int nbins, const uint4* __restrict__ brec,
                                                                      uint32_t* __restrict__ order, uint32_t* __restrict__ runoff) {
    // per warp: the bin's composites (unsorted) and its (id, rows) in sorted order
    extern __shared__ __align__(16) unsigned char s_raw[];
    if (hdr->sort_mode != 1u) return;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int warp_global = blockIdx.x * kBinSortWarps + wid, warps_total = gridDim.x * kBinSortWarps;
    unsigned long long* comp = reinterpret_cast<unsigned long long*>(s_raw) + (size_t)wid * kBinCap;
    uint2* sorted = reinterpret_cast<uint2*>(s_raw + sizeof(unsigned long long) * kBinSortWarps * kBinCap) + (size_t)wid * kBinCap;
    // bins are dealt round-robin: neighbouring (similarly loaded) bins go to different warps
    for (int bin = warp_global; bin < nbins; bin += warps_total) {
        const uint32_t end = __ldg(bin_end + bin), start = bin ? __ldg(bin_end + bin - 1) : 0u;   // after the scatter a cursor = its bin's end
        const int n = (int)(end - start);
        if (n <= 0) continue;
        uint32_t run0 = __ldg(bin_runs + bin);
        if (n <= 32) {
            // ---- register path: one record per lane, rank by n shuffled comparisons
            uint4 r = make_uint4(0xffffffffu, 0xffffffffu, 0u, 0u);
            if (lane < n) r = __ldg(brec + start + lane);
            const unsigned long long c = ((unsigned long long)r.x << 32) | r.y;
            int rank = 0;
            for (int j = 0; j < n; ++j) {
                const unsigned long long o = __shfl_sync(0xffffffffu, c, j);
                rank += (o < c) ? 1 : 0;
            }
            __syncwarp();
            if (lane < n) sorted[rank] = make_uint2(r.y, r.z);
            __syncwarp();
            uint2 v = make_uint2(0u, 0u);
            if (lane < n) v = sorted[lane];
            uint32_t inc = v.y;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
            if (lane < n) { order[start + lane] = v.x; runoff[start + lane] = run0 + inc - v.y; }
        } else {
            // ---- shared-memory path (33 .. kBinCap records): all-pairs rank, records strided over the lanes
            for (int i = lane; i < n; i += 32) { const uint4 r = __ldg(brec + start + i); comp[i] = ((unsigned long long)r.x << 32) | r.y; }
            __syncwarp();
            for (int i = lane; i < n; i += 32) {
                const unsigned long long c = comp[i];
                int rank = 0;
                for (int j = 0; j < n; ++j) rank += (comp[j] < c) ? 1 : 0;       // broadcast reads
                const uint4 r = __ldg(brec + start + i);
                sorted[rank] = make_uint2(r.y, r.z);
            }
            __syncwarp();
            for (int base = 0; base < n; base += 32) {
                const int i = base + lane;
                uint2 v = make_uint2(0u, 0u);
                if (i < n) v = sorted[i];
                uint32_t inc = v.y;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
                if (i < n) { order[start + i] = v.x; runoff[start + i] = run0 + inc - v.y; }
                run0 += __shfl_sync(0xffffffffu, inc, 31);
            }
        }
        __syncwarp();
    }
}

void launch_bin_sort(const GeomPtrs& g, int P, cudaStream_t s) {
    int grid = (g.nbins + kBinSortWarps * 8 - 1) / (kBinSortWarps * 8);        // ~8 bins per warp
    const int cap = ds_num_sms() * 3;                                           // 3 CTAs (3 x 64 KB) per SM
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    const size_t smem = (sizeof(unsigned long long) + sizeof(uint2)) * kBinSortWarps * kBinCap;
    static bool attr_done = false;
    if (!attr_done) { cudaFuncSetAttribute(bin_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); attr_done = true; }
    bin_sort_kernel<<<grid, kBinSortWarps * 32, smem, s>>>(g.header, g.bin_cur, g.bin_runs, g.nbins, g.brec, g.order, g.runoff);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// fallback: run offsets of an order produced by the onesweep chain (order_block_sums + scan give the offset of every
// 256-block of sorted positions; this finishes the scan inside the block)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kPreBlock) runoff_from_blocks_kernel(GeomHeader* hdr, const uint32_t* __restrict__ order, const uint2* __restrict__ rects,
                                                                       const uint32_t* __restrict__ block_offsets, uint32_t* __restrict__ runoff) {
    __shared__ uint32_t s_w[kPreBlock / 32];
    const uint32_t V = hdr->num_visible;
    const uint32_t j = blockIdx.x * kPreBlock + threadIdx.x;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t h = 0;
    if (j < V) { const uint32_t ry = rects[order[j]].y; h = (ry >> 16) - (ry & 0xffffu); }
    uint32_t inc = h;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (lane == 31) s_w[wid] = inc;
    __syncthreads();
    uint32_t base = __ldg(block_offsets + blockIdx.x);
#pragma unroll
    for (int w = 0; w < kPreBlock / 32; ++w) base += (w < wid) ? s_w[w] : 0u;
    if (j < V) runoff[j] = base + inc - h;
    if (j == V) runoff[j] = hdr->num_runs;
    if (blockIdx.x == 0 && threadIdx.x == 0) hdr->sort_mode = 3u;
}

void launch_runoff_from_blocks(const GeomPtrs& g, int P, cudaStream_t s) {
    runoff_from_blocks_kernel<<<P / kPreBlock + 1, kPreBlock, 0, s>>>(g.header, g.order, g.rects, g.oblock_offsets, g.runoff);
    count_launch();
}

}  // namespace gsr
