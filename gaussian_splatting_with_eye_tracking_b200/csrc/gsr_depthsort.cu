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
    // The next block's count is fetched before the current block's records are touched (the loop is a chain of dependent
    // L2 round trips otherwise).
    const int sub = tid >> 8, t = tid & 255;
    const int bstep = (int)gridDim.x * 4;
    int b = (int)blockIdx.x * 4 + sub;
    uint32_t cnt = b < nblocks ? __ldg(block_vis + b) : 0u;
    while (b < nblocks) {
        const int bn = b + bstep;
        const uint32_t cnt_next = bn < nblocks ? __ldg(block_vis + bn) : 0u;
        if ((uint32_t)t < cnt) {
            const size_t i = (size_t)b * kPreBlock + t;
            const uint32_t bin = (__ldg(ckeys + i) - kmin) >> shift;
            atomicAdd(&bin_cur[bin], 1u);
            atomicAdd(&bin_runs[bin], __ldg(crows + i) & 0xffffu);
        }
        b = bn; cnt = cnt_next;
    }
    __syncthreads();
    if (tid == 0) { __threadfence(); s_last = (atomicAdd(&hdr->done_hist, 1u) == gridDim.x - 1) ? 1u : 0u; }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

    // ---- last CTA: exclusive scans of both histograms (in place), largest bin, totals.  Every thread owns nbins / 1024
    // consecutive bins (a multiple of 4 from 4 Ki bins on: 128-bit loads, four in flight per array).
    const int per = nbins / kHistThreads;               // nbins is a power of two >= 1024
    const int i0 = tid * per;
    uint32_t sc = 0, sr = 0, mx = 0;
    if ((per & 15) == 0) {
        const uint4* c4 = reinterpret_cast<const uint4*>(bin_cur + i0);
        const uint4* r4 = reinterpret_cast<const uint4*>(bin_runs + i0);
        for (int i = 0; i < per / 4; i += 4) {
            uint4 c[4], r[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) { c[k] = __ldcg(c4 + i + k); r[k] = __ldcg(r4 + i + k); }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                sc += c[k].x + c[k].y + c[k].z + c[k].w; sr += r[k].x + r[k].y + r[k].z + r[k].w;
                mx = max(mx, max(max(c[k].x, c[k].y), max(c[k].z, c[k].w)));
            }
        }
    } else {
        for (int i = 0; i < per; ++i) { const uint32_t c = __ldcg(bin_cur + i0 + i); sc += c; sr += __ldcg(bin_runs + i0 + i); mx = max(mx, c); }
    }
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
        const uint32_t a = s_a[w], b2 = s_b[w];
        bc += (w < wid) ? a : 0u; br += (w < wid) ? b2 : 0u; V += a; R += b2; maxbin = max(maxbin, s_c[w]);
    }
    uint32_t rc = bc + ic - sc, rr = br + ir - sr;
    if ((per & 15) == 0) {
        uint4* c4 = reinterpret_cast<uint4*>(bin_cur + i0);
        uint4* r4 = reinterpret_cast<uint4*>(bin_runs + i0);
        for (int i = 0; i < per / 4; i += 4) {
            uint4 c[4], r[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) { c[k] = __ldcg(c4 + i + k); r[k] = __ldcg(r4 + i + k); }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                uint4 oc, orr;
                oc.x = rc; rc += c[k].x; oc.y = rc; rc += c[k].y; oc.z = rc; rc += c[k].z; oc.w = rc; rc += c[k].w;
                orr.x = rr; rr += r[k].x; orr.y = rr; rr += r[k].y; orr.z = rr; rr += r[k].z; orr.w = rr; rr += r[k].w;
                c4[i + k] = oc; r4[i + k] = orr;       // bin start (the scatter's cursor) and the bin's first run
            }
        }
    } else {
        for (int i = 0; i < per; ++i) {
            const uint32_t c = __ldcg(bin_cur + i0 + i), r = __ldcg(bin_runs + i0 + i);
            bin_cur[i0 + i] = rc; bin_runs[i0 + i] = rr;
            rc += c; rr += r;
        }
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
                                                                   const uint32_t* __restrict__ cxw, uint32_t* bin_cur, uint4* __restrict__ brec) {
    if (hdr->sort_mode != 1u) return;
    const uint32_t kmin = hdr->key_min, shift = hdr->shift;
    const int sub = threadIdx.x >> 8, t = threadIdx.x & 255;
    const int bstep = (int)gridDim.x * 4;
    int b = (int)blockIdx.x * 4 + sub;
    uint32_t cnt = b < nblocks ? __ldg(block_vis + b) : 0u;
    while (b < nblocks) {
        const int bn = b + bstep;
        const uint32_t cnt_next = bn < nblocks ? __ldg(block_vis + bn) : 0u;
        if ((uint32_t)t < cnt) {
            const size_t i = (size_t)b * kPreBlock + t;
            const uint32_t key = __ldg(ckeys + i);
            const uint4 rec = make_uint4(key, __ldg(cids + i), __ldg(crows + i), __ldg(cxw + i));
            const uint32_t pos = atomicAdd(&bin_cur[(key - kmin) >> shift], 1u);
            brec[pos] = rec;
        }
        b = bn; cnt = cnt_next;
    }
}

void launch_bin_scatter(const GeomPtrs& g, int P, cudaStream_t s) {
    int grid = (g.nblocks + 3) / 4;
    const int cap = ds_num_sms() * 2;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    bin_scatter_kernel<<<grid, kHistThreads, 0, s>>>(g.header, g.nblocks, g.block_sums_v, g.dkeys_a, g.order_tmp, g.dkeys_b, g.order, g.bin_cur, g.brec);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// exact order inside every bin + run offsets
// ------------------------------------------------------------------------------------------------
constexpr int kBinSortWarps = 8;

// No shared memory, so that 64 warps per SM are in flight: a bin costs three dependent L2 round trips (cursor words,
// records, stores) and ~20 of them queue behind each other in a warp -- the kernel is latency-bound, not work-bound.
// `scratch` (one uint4 per record, indexed like brec) is only touched by bins with more than 32 records.
__global__ void __launch_bounds__(kBinSortWarps * 32) bin_sort_kernel(const GeomHeader* __restrict__ hdr, const uint32_t* __restrict__ bin_end,
                                                                      const uint32_t* __restrict__ bin_runs, int nbins, const uint4* __restrict__ brec,
                                                                      uint4* scratch, uint32_t* __restrict__ order, uint32_t* __restrict__ runoff,
                                                                      uint2* __restrict__ srect) {
    if (hdr->sort_mode != 1u) return;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int warp_global = blockIdx.x * kBinSortWarps + wid, warps_total = gridDim.x * kBinSortWarps;
    // bins are dealt round-robin: neighbouring (similarly loaded) bins go to different warps
    for (int bin = warp_global; bin < nbins; bin += warps_total) {
        const uint32_t end = __ldg(bin_end + bin), start = bin ? __ldg(bin_end + bin - 1) : 0u;   // after the scatter a cursor = its bin's end
        const int n = (int)(end - start);
        if (n <= 0) continue;
        uint32_t run0 = __ldg(bin_runs + bin);
        if (n <= 32) {
            // ---- one record per lane: rank by n shuffled comparisons of the 64-bit composite (depth key, id), then pull the
            // record of rank `lane` from the lane that holds it
            uint4 r = make_uint4(0xffffffffu, 0xffffffffu, 0u, 0u);
            if (lane < n) r = __ldg(brec + start + lane);
            const unsigned long long c = ((unsigned long long)r.x << 32) | r.y;
            int rank = 0;
            for (int j = 0; j < n; ++j) rank += (__shfl_sync(0xffffffffu, c, j) < c) ? 1 : 0;
            int src = 0;
            for (int j = 0; j < n; ++j) src = (__shfl_sync(0xffffffffu, rank, j) == lane) ? j : src;
            const uint32_t id = __shfl_sync(0xffffffffu, r.y, src), yh = __shfl_sync(0xffffffffu, r.z, src), xw = __shfl_sync(0xffffffffu, r.w, src);
            const uint32_t h = lane < n ? (yh & 0xffffu) : 0u;
            uint32_t inc = h;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
            if (lane < n) { order[start + lane] = id; runoff[start + lane] = run0 + inc - h; srect[start + lane] = make_uint2(xw, yh >> 16); }
        } else {
            // ---- 33 .. kBinCap records: all-pairs rank with the records read through the L1, sorted records parked in
            // `scratch`, then the scan in sorted order
            for (int i = lane; i < n; i += 32) {
                const uint4 r = __ldg(brec + start + i);
                const unsigned long long c = ((unsigned long long)r.x << 32) | r.y;
                int rank = 0;
                for (int j = 0; j < n; ++j) { const uint4 o = __ldg(brec + start + j); rank += ((((unsigned long long)o.x << 32) | o.y) < c) ? 1 : 0; }
                scratch[start + rank] = r;
            }
            __syncwarp();
            for (int base = 0; base < n; base += 32) {
                const int i = base + lane;
                uint4 r = make_uint4(0u, 0u, 0u, 0u);
                if (i < n) r = __ldcg(scratch + start + i);
                const uint32_t h = r.z & 0xffffu;
                uint32_t inc = h;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
                if (i < n) { order[start + i] = r.y; runoff[start + i] = run0 + inc - h; srect[start + i] = make_uint2(r.w, r.z >> 16); }
                run0 += __shfl_sync(0xffffffffu, inc, 31);
            }
        }
    }
}

void launch_bin_sort(const GeomPtrs& g, int P, cudaStream_t s) {
    int grid = (g.nbins + kBinSortWarps * 4 - 1) / (kBinSortWarps * 4);        // >= 4 bins per warp
    const int cap = ds_num_sms() * 8;                                           // 8 CTAs x 8 warps per SM
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    // scratch for the rare large bins: the per-block compaction arrays are dead by now (dkeys_a .. order_tmp hold 2 x 4 B x P,
    // contiguous in the layout; a uint4 per record needs 16 B x V <= 16 B x P): use the cov3D-independent sort temp instead
    bin_sort_kernel<<<grid, kBinSortWarps * 32, 0, s>>>(g.header, g.bin_cur, g.bin_runs, g.nbins, g.brec, g.brec_sorted, g.order, g.runoff, g.srect);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// fallback: run offsets of an order produced by the onesweep chain (order_block_sums + scan give the offset of every
// 256-block of sorted positions; this finishes the scan inside the block)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kPreBlock) runoff_from_blocks_kernel(GeomHeader* hdr, const uint32_t* __restrict__ order, const uint2* __restrict__ rects,
                                                                       const uint32_t* __restrict__ block_offsets, uint32_t* __restrict__ runoff,
                                                                       uint2* __restrict__ srect) {
    __shared__ uint32_t s_w[kPreBlock / 32];
    const uint32_t V = hdr->num_visible;
    const uint32_t j = blockIdx.x * kPreBlock + threadIdx.x;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t h = 0;
    if (j < V) {
        const uint2 r = rects[order[j]];
        const uint32_t x0 = r.x & 0xffffu, y0 = r.y & 0xffffu;
        h = (r.y >> 16) - y0;
        srect[j] = make_uint2((x0 << kRunX0Shift) | ((r.x >> 16) - x0), y0);
    }
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
    runoff_from_blocks_kernel<<<P / kPreBlock + 1, kPreBlock, 0, s>>>(g.header, g.order, g.rects, g.oblock_offsets, g.runoff, g.srect);
    count_launch();
}

}  // namespace gsr
