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
//   preprocess_fwd_kernel  (gsr_preprocess.cu) leaves, per 256-Gaussian block, the (depth key, id, rect) of its visible
//                          Gaussians compacted to the front of the block's segment, and COUNTS every one of them into its
//                          depth bin -- bin = (key - bits(0.2f)) >> shift, a fixed map, so no key-range pass is needed --
//                          with one 64-bit atomic (+1 record, +rows runs) whose return value is the record's slot inside
//                          the bin.  The atomics ride under the kernel's HBM-bound run time.
//   bin_scan_kernel        one CTA per 1024 bins: exclusive prefix of records and runs inside the segment + segment totals
//                          (consumers add the prefix of the <= 1024 segment totals themselves); the CTA that finishes last
//                          posts the frame totals to the host mailbox
//   bin_scatter_kernel     one 16-byte record per Gaussian to bin start + slot: plain loads and stores, no atomics
//   bin_sort_kernel        a warp per bin (32 bins per coalesced look-up, next bin's records prefetched): exact rank of
//                          every record by all-pairs comparison of the composite (key position inside the bin, id) -- bins
//                          hold a dozen records -- and, in the same loop, the runs emitted by the records before it: every
//                          Gaussian gets its sorted position and the absolute offset of its first run.  Replaces the
//                          remaining radix passes AND order_block_sums + scan.
//
// A bin with more than kBinCap records (thousands of Gaussians within 1/depth_bins of the depth range, e.g. all-equal
// depths) is not sorted by one warp: bin_scan_kernel then sets sort_mode = 2, the two kernels behind it leave at once, and
// the host runs the round-1 onesweep chain instead (gsr_api.cu), so that everything downstream sees `order` + `runoff`.
#include "gsr_common.cuh"

namespace gsr {

static int ds_num_sms() {
    static int n = 0;
    if (n == 0) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); if (n <= 0) n = 148; }
    return n;
}

// ------------------------------------------------------------------------------------------------
// scan: one CTA per segment of 1024 bins -- exclusive prefix of records / runs INSIDE the segment + the segment's totals.
// Consumers add the exclusive prefix of the segment totals themselves (seg_base()).  The CTA that finishes last sums the
// totals and the statistics slots and posts the frame's counts where the host reads them.
// ------------------------------------------------------------------------------------------------
constexpr int kIlp = 4;                        // 256-blocks a thread of the scatter has in flight

__global__ void __launch_bounds__(kSegBins) bin_scan_kernel(GeomHeader* hdr, const FrameStatSlot* __restrict__ slots, int P, int nbins,
                                                            unsigned long long* bins64, uint32_t* __restrict__ bin_cnt, uint4* seg_tot,
                                                            uint32_t* runoff, int shift, int id_bits, uint32_t* mailbox) {
    __shared__ uint32_t s_a[32], s_b[32], s_c[32], s_d[32];
    __shared__ uint32_t s_last;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int i = blockIdx.x * kSegBins + tid;
    const unsigned long long packed = bins64[i];
    const uint32_t c = (uint32_t)packed, r = (uint32_t)(packed >> 32);
    uint32_t ic = c, ir = r, mx = c;
    uint32_t sq = c > 32u ? min(c, 0xffffu) * min(c, 0xffffu) : 0u;        // all-pairs cost of a bin the warp cannot rank in registers
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, ic, o), v = __shfl_up_sync(0xffffffffu, ir, o);
        if (lane >= o) { ic += u; ir += v; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o)); sq += __shfl_xor_sync(0xffffffffu, sq, o); }
    if (lane == 31) { s_a[wid] = ic; s_b[wid] = ir; }
    if (lane == 0) { s_c[wid] = mx; s_d[wid] = sq; }
    __syncthreads();
    uint32_t bc = 0, br = 0, tc = 0, tr = 0, maxbin = 0, tsq = 0;
#pragma unroll
    for (int w = 0; w < 32; ++w) {
        const uint32_t a = s_a[w], b = s_b[w];
        bc += (w < wid) ? a : 0u; br += (w < wid) ? b : 0u; tc += a; tr += b; maxbin = max(maxbin, s_c[w]);
        tsq = (tsq + s_d[w] < tsq) ? 0xffffffffu : tsq + s_d[w];           // saturating
    }
    if (tc != 0u) {                                    // (segments without records keep their zeros: nobody reads them)
        bins64[i] = ((unsigned long long)(br + ir - r) << 32) | (unsigned long long)(bc + ic - c);   // (first run, first slot), relative
        bin_cnt[i] = c;
    } else bin_cnt[i] = 0u;
    if (tid == 0) {
        seg_tot[blockIdx.x] = make_uint4(tc, tr, tsq, 0u);
        if (maxbin) atomicMax(&hdr->max_bin, maxbin);
        __threadfence();
        s_last = (atomicAdd(&hdr->done_hist, 1u) == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // ---- last CTA: frame totals
    uint32_t v = 0, rr = 0, pairs = 0, flags = 0;
    unsigned long long cost = 0;
    for (int k = tid; k < (int)gridDim.x; k += kSegBins) { const uint4 t2 = __ldcg(seg_tot + k); v += t2.x; rr += t2.y; cost += t2.z; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cost += __shfl_xor_sync(0xffffffffu, cost, o);
    __shared__ unsigned long long s_cost[32];
    if (lane == 0) s_cost[wid] = cost;                 // summed by thread 0 behind the block-wide OR below
    if (tid < kStatSlots) { pairs = slots[tid].num_rendered; flags = slots[tid].flags; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v += __shfl_xor_sync(0xffffffffu, v, o); rr += __shfl_xor_sync(0xffffffffu, rr, o);
        pairs += __shfl_xor_sync(0xffffffffu, pairs, o); flags |= __shfl_xor_sync(0xffffffffu, flags, o);
    }
    __syncthreads();
    if (lane == 0) { s_a[wid] = v; s_b[wid] = rr; s_c[wid] = pairs; }
    const int flags_any = __syncthreads_or((int)flags);
    if (tid == 0) {
        uint32_t V = 0, R = 0, N = 0;
        for (int w = 0; w < 32; ++w) { V += s_a[w]; R += s_b[w]; N += s_c[w]; }
        const uint32_t mb = atomicMax(&hdr->max_bin, 0u);
        unsigned long long pair_cost = 0;
        for (int w = 0; w < 32; ++w) pair_cost += s_cost[w];
        const bool ok = mb <= (uint32_t)kBinCap && !(flags_any & 2) && pair_cost <= kBinPairBudget;
        const uint32_t err = (flags_any & 1) ? 0x100u : 0u;
        hdr->magic = kGeomMagic; hdr->P = (uint32_t)P; hdr->num_rendered = N; hdr->num_runs = R; hdr->num_visible = V;
        hdr->sort_error = err; hdr->shift = (uint32_t)shift; hdr->sort_mode = ok ? 1u : 2u;
        hdr->comp32 = (shift + id_bits <= 32) ? 1u : 0u;
        runoff[V] = R;                                 // read by the run emission of the last block of sorted positions
        if (mailbox) {   // mapped pinned host memory: the host reads the totals without a DMA copy; posted PCIe writes stay
                         // ordered and the host reads after an event recorded behind this kernel
            volatile uint32_t* m = mailbox;
            m[0] = kGeomMagic; m[1] = (uint32_t)P; m[2] = N; m[3] = err; m[4] = R; m[5] = V; m[6] = ok ? 1u : 2u; m[7] = mb;
        }
    }
}

void launch_bin_scan(const GeomPtrs& g, int P, uint32_t* mailbox_dev, cudaStream_t s) {
    bin_scan_kernel<<<g.nbins / kSegBins, kSegBins, 0, s>>>(g.header, g.stat_slots, P, g.nbins, g.bins64, g.bin_cnt, g.seg_tot, g.runoff,
                                                           depth_shift(P), index_bits(P), mailbox_dev);
    count_launch();
}

// exclusive prefix of the segment totals, by every consumer CTA for itself (<= 1024 segments): s_seg[k] = (records, runs) before segment k
__device__ __forceinline__ void seg_base(const uint4* __restrict__ seg_tot, int nseg, uint2* s_seg) {
    __shared__ uint2 s_w[32];
    __shared__ uint2 s_run;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nthr = blockDim.x, nw = nthr / 32;
    if (tid == 0) s_run = make_uint2(0u, 0u);
    for (int base = 0; base < nseg; base += nthr) {
        const int k = base + tid;
        uint2 t = make_uint2(0u, 0u);
        if (k < nseg) { const uint4 t4 = __ldcg(seg_tot + k); t = make_uint2(t4.x, t4.y); }
        uint2 inc = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, inc.x, o), v = __shfl_up_sync(0xffffffffu, inc.y, o);
            if (lane >= o) { inc.x += u; inc.y += v; }
        }
        __syncthreads();
        if (lane == 31) s_w[wid] = inc;
        __syncthreads();
        uint2 b = s_run, tot = make_uint2(0u, 0u);
        for (int w = 0; w < nw; ++w) { const uint2 x = s_w[w]; if (w < wid) { b.x += x.x; b.y += x.y; } tot.x += x.x; tot.y += x.y; }
        if (k < nseg) s_seg[k] = make_uint2(b.x + inc.x - t.x, b.y + inc.y - t.y);
        __syncthreads();
        if (tid == 0) { s_run.x += tot.x; s_run.y += tot.y; }
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// scatter into the bins' segments
// ------------------------------------------------------------------------------------------------
// record = (composite, (y0 << 16) | rows, (x0 << 10) | width, id).  composite = position of the key inside its bin (low `shift`
// key bits) above the id bits when that fits 32 bits (hdr->comp32), so that one 32-bit compare orders (depth, index);
// otherwise the full key (ties then compare the id word).  Slot = first slot of the bin + the count the preprocess kernel's
// atomic returned: no atomics here.
__global__ void __launch_bounds__(1024) bin_scatter_kernel(const GeomHeader* __restrict__ hdr, int nblocks, int nseg, int id_bits,
                                                           const uint32_t* __restrict__ block_vis, const uint32_t* __restrict__ ckeys,
                                                           const uint32_t* __restrict__ cids, const uint32_t* __restrict__ crows,
                                                           const uint32_t* __restrict__ cxw, const uint32_t* __restrict__ crank,
                                                           const unsigned long long* __restrict__ bins64, const uint4* __restrict__ seg_tot,
                                                           uint4* __restrict__ brec) {
    extern __shared__ uint2 s_seg[];                  // [nseg]
    if (hdr->sort_mode != 1u) return;
    seg_base(seg_tot, nseg, s_seg);
    const uint32_t shift = hdr->shift, comp32 = hdr->comp32;
    const uint32_t low = (1u << shift) - 1u;
    const int sub = threadIdx.x >> 8, t = threadIdx.x & 255;
    const int bstep = (int)gridDim.x * 4;
    for (int b0 = (int)blockIdx.x * 4 + sub; b0 < nblocks; b0 += bstep * kIlp) {
        uint4 rec[kIlp]; uint32_t bin[kIlp], slot[kIlp]; bool on[kIlp];
#pragma unroll
        for (int k = 0; k < kIlp; ++k) { const int b = b0 + k * bstep; on[k] = b < nblocks && (uint32_t)t < __ldg(block_vis + b); }
#pragma unroll
        for (int k = 0; k < kIlp; ++k) if (on[k]) {
            const size_t i = (size_t)(b0 + k * bstep) * kPreBlock + t;
            const uint32_t key = __ldg(ckeys + i), id = __ldg(cids + i);
            const uint32_t rel = key - kDepthKey0;
            bin[k] = rel >> shift;
            slot[k] = __ldg(crank + i);
            rec[k] = make_uint4(comp32 ? (((rel & low) << id_bits) | id) : key, __ldg(crows + i), __ldg(cxw + i), id);
        }
#pragma unroll
        for (int k = 0; k < kIlp; ++k) if (on[k]) slot[k] += s_seg[bin[k] / kSegBins].x + (uint32_t)__ldg(bins64 + bin[k]);
#pragma unroll
        for (int k = 0; k < kIlp; ++k) if (on[k]) brec[slot[k]] = rec[k];
    }
}

void launch_bin_scatter(const GeomPtrs& g, int P, cudaStream_t s) {
    int grid = (g.nblocks + 4 * kIlp - 1) / (4 * kIlp);
    const int cap = ds_num_sms() * 2;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    const int nseg = g.nbins / kSegBins;
    bin_scatter_kernel<<<grid, 1024, sizeof(uint2) * nseg, s>>>(g.header, g.nblocks, nseg, index_bits(P), g.block_sums_v, g.dkeys_a, g.order_tmp,
                                                              g.dkeys_b, g.order, g.crank, g.bins64, g.seg_tot, g.brec);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// exact order inside every bin + run offsets
// ------------------------------------------------------------------------------------------------
constexpr int kBinSortWarps = 8;

// A warp takes 32 consecutive bins per step (one coalesced look-up of their counts) and then one non-empty bin at a time, the
// next bin's records already in flight.  Every lane ranks its record against the bin's others and PUSHES its outputs to slot
// start + rank (the bin's slots are one short contiguous stretch); the number of runs emitted by the records ordered before
// it -- the exclusive prefix of the tile rows in sorted order -- is accumulated in the same loop, so no inverse permutation
// and no scan is needed.
__global__ void __launch_bounds__(kBinSortWarps * 32) bin_sort_kernel(const GeomHeader* __restrict__ hdr, const unsigned long long* __restrict__ bins64,
                                                                      const uint32_t* __restrict__ bin_cnt, const uint4* __restrict__ seg_tot,
                                                                      int nbins, const uint4* __restrict__ brec, uint32_t* __restrict__ order,
                                                                      uint32_t* __restrict__ runoff, uint2* __restrict__ srect) {
    extern __shared__ uint2 s_seg[];                  // [nbins / kSegBins]
    if (hdr->sort_mode != 1u) return;
    seg_base(seg_tot, nbins / kSegBins, s_seg);
    const bool comp32 = hdr->comp32 != 0u;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int warp_global = blockIdx.x * kBinSortWarps + wid, warps_total = gridDim.x * kBinSortWarps;
    const int ngroups = nbins / 32;
    // groups of 32 bins are dealt round-robin: neighbouring (similarly loaded) groups go to different warps
    for (int grp = warp_global; grp < ngroups; grp += warps_total) {
        const int bin = grp * 32 + lane;
        const uint32_t cnt = __ldg(bin_cnt + bin);
        uint32_t mask = __ballot_sync(0xffffffffu, cnt != 0u);
        if (mask == 0u) continue;
        uint32_t my_start = 0, my_run0 = 0;
        if (cnt) {
            const unsigned long long sr = __ldg(bins64 + bin);
            const uint2 sb = s_seg[bin / kSegBins];
            my_start = sb.x + (uint32_t)sr; my_run0 = sb.y + (uint32_t)(sr >> 32);
        }
        int k = __ffs(mask) - 1;
        int n = (int)__shfl_sync(0xffffffffu, cnt, k);
        uint32_t start = __shfl_sync(0xffffffffu, my_start, k), run0 = __shfl_sync(0xffffffffu, my_run0, k);
        uint4 r = make_uint4(0xffffffffu, 0u, 0u, 0xffffffffu);
        if (lane < n && n <= 32) r = __ldg(brec + start + lane);
        while (true) {
            mask &= mask - 1u;
            // the next non-empty bin of the group: its records are requested before this bin's are ranked
            int k2 = 0, n2 = 0; uint32_t start2 = 0, run02 = 0;
            uint4 r2 = make_uint4(0xffffffffu, 0u, 0u, 0xffffffffu);
            if (mask) {
                k2 = __ffs(mask) - 1;
                n2 = (int)__shfl_sync(0xffffffffu, cnt, k2);
                start2 = __shfl_sync(0xffffffffu, my_start, k2); run02 = __shfl_sync(0xffffffffu, my_run0, k2);
                if (lane < n2 && n2 <= 32) r2 = __ldg(brec + start2 + lane);
            }
            if (n <= 32) {
                const uint32_t h = r.y & 0xffffu;
                uint32_t rank = 0, before = 0;
                if (comp32) {
                    for (int j = 0; j < n; ++j) {
                        const uint32_t oc = __shfl_sync(0xffffffffu, r.x, j), oh = __shfl_sync(0xffffffffu, h, j);
                        const bool lt = oc < r.x;
                        rank += lt ? 1u : 0u; before += lt ? oh : 0u;
                    }
                } else {
                    for (int j = 0; j < n; ++j) {
                        const uint32_t ok = __shfl_sync(0xffffffffu, r.x, j), oi = __shfl_sync(0xffffffffu, r.w, j), oh = __shfl_sync(0xffffffffu, h, j);
                        const bool lt = ok < r.x || (ok == r.x && oi < r.w);
                        rank += lt ? 1u : 0u; before += lt ? oh : 0u;
                    }
                }
                if (lane < n) { const uint32_t at = start + rank; order[at] = r.w; runoff[at] = run0 + before; srect[at] = make_uint2(r.z, r.y >> 16); }
            } else {
                // 33 .. kBinCap records: the same all-pairs loop with the bin's records re-read through the L1
                for (int i = lane; i < n; i += 32) {
                    const uint4 q = __ldg(brec + start + i);
                    uint32_t rank = 0, before = 0;
#pragma unroll 4
                    for (int j = 0; j < n; ++j) {
                        const uint4 o = __ldg(brec + start + j);
                        const bool lt = o.x < q.x || (!comp32 && o.x == q.x && o.w < q.w);
                        rank += lt ? 1u : 0u; before += lt ? (o.y & 0xffffu) : 0u;
                    }
                    const uint32_t at = start + rank;
                    order[at] = q.w; runoff[at] = run0 + before; srect[at] = make_uint2(q.z, q.y >> 16);
                }
            }
            if (!mask) break;
            k = k2; n = n2; start = start2; run0 = run02; r = r2;
        }
    }
}

void launch_bin_sort(const GeomPtrs& g, int P, cudaStream_t s) {
    const int ngroups = g.nbins / 32;
    int grid = (ngroups + kBinSortWarps * 2 - 1) / (kBinSortWarps * 2);
    const int cap = ds_num_sms() * 8;                                           // 8 CTAs x 8 warps per SM
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    bin_sort_kernel<<<grid, kBinSortWarps * 32, sizeof(uint2) * (g.nbins / kSegBins), s>>>(g.header, g.bins64, g.bin_cnt, g.seg_tot, g.nbins, g.brec,
                                                                                          g.order, g.runoff, g.srect);
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
