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
// Consumers add the exclusive prefix of the segment totals themselves (seg_base()).  No atomics, no "last CTA" ticket: a
// thousand CTAs bumping one counter would serialise for longer than the scan takes (same-address atomics: ~30 ns each); the
// frame totals are taken by the one-CTA bin_totals_kernel behind it.
// ------------------------------------------------------------------------------------------------
constexpr int kIlp = 4;                        // 256-blocks a thread of the scatter has in flight

__global__ void __launch_bounds__(kSegBins) bin_scan_kernel(unsigned long long* bins64, uint4* __restrict__ seg_tot) {
    __shared__ uint32_t s_a[32], s_b[32], s_c[32], s_d[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int i = blockIdx.x * kSegBins + tid;
    const unsigned long long packed = bins64[i];
    const uint32_t c = (uint32_t)packed, r = (uint32_t)(packed >> 32);
    uint32_t ic = c, ir = r, mx = c;
    uint32_t sq = c > 32u ? min(c, 0xffffu) * min(c, 0xffffu) : 0u;        // all-pairs cost of a bin too full for one thread
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
    if (tc != 0u) bins64[i] = ((unsigned long long)(br + ir - r) << 32) | (unsigned long long)(bc + ic - c);   // (first run, first slot), relative
    if (tid == 0) seg_tot[blockIdx.x] = make_uint4(tc, tr, tsq, maxbin);
}

// frame totals: one CTA sums the segment totals and the statistics slots and posts the counts where the host reads them
__global__ void __launch_bounds__(1024) bin_totals_kernel(GeomHeader* hdr, const FrameStatSlot* __restrict__ slots, int P, int nseg,
                                                          const uint4* __restrict__ seg_tot, uint2* __restrict__ seg_pre, uint32_t* runoff,
                                                          int shift, int id_bits, uint32_t* mailbox) {
    __shared__ uint32_t s_a[32], s_b[32], s_c[32], s_d[32];
    __shared__ unsigned long long s_cost[32];
    __shared__ uint2 s_w[32];
    __shared__ uint2 s_run;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    // exclusive prefix of the segment totals (records, runs): what the scatter and the per-bin sort add to segment-relative offsets
    if (tid == 0) s_run = make_uint2(0u, 0u);
    for (int base = 0; base < nseg; base += 1024) {
        const int k = base + tid;
        uint2 t = make_uint2(0u, 0u);
        if (k < nseg) { const uint4 t4 = seg_tot[k]; t = make_uint2(t4.x, t4.y); }
        uint2 inc = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, inc.x, o), v2 = __shfl_up_sync(0xffffffffu, inc.y, o);
            if (lane >= o) { inc.x += u; inc.y += v2; }
        }
        __syncthreads();
        if (lane == 31) s_w[wid] = inc;
        __syncthreads();
        uint2 b = s_run, tot = make_uint2(0u, 0u);
        for (int w = 0; w < 32; ++w) { const uint2 x = s_w[w]; if (w < wid) { b.x += x.x; b.y += x.y; } tot.x += x.x; tot.y += x.y; }
        if (k < nseg) seg_pre[k] = make_uint2(b.x + inc.x - t.x, b.y + inc.y - t.y);
        __syncthreads();
        if (tid == 0) { s_run.x += tot.x; s_run.y += tot.y; }
    }
    __syncthreads();
    uint32_t v = 0, rr = 0, pairs = 0, flags = 0, mb = 0;
    unsigned long long cost = 0;
    for (int k = tid; k < nseg; k += 1024) { const uint4 t = seg_tot[k]; v += t.x; rr += t.y; cost += t.z; mb = max(mb, t.w); }
    if (tid < kStatSlots) { pairs = slots[tid].num_rendered; flags = slots[tid].flags; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v += __shfl_xor_sync(0xffffffffu, v, o); rr += __shfl_xor_sync(0xffffffffu, rr, o); cost += __shfl_xor_sync(0xffffffffu, cost, o);
        pairs += __shfl_xor_sync(0xffffffffu, pairs, o); flags |= __shfl_xor_sync(0xffffffffu, flags, o);
        mb = max(mb, __shfl_xor_sync(0xffffffffu, mb, o));
    }
    if (lane == 0) { s_a[wid] = v; s_b[wid] = rr; s_c[wid] = pairs; s_d[wid] = mb; s_cost[wid] = cost; }
    const int flags_any = __syncthreads_or((int)flags);
    if (tid == 0) {
        uint32_t V = 0, R = 0, N = 0, maxbin = 0;
        unsigned long long pair_cost = 0;
        for (int w = 0; w < 32; ++w) { V += s_a[w]; R += s_b[w]; N += s_c[w]; maxbin = max(maxbin, s_d[w]); pair_cost += s_cost[w]; }
        const bool ok = maxbin <= (uint32_t)kBinCap && !(flags_any & 2) && pair_cost <= kBinPairBudget;
        const uint32_t err = (flags_any & 1) ? 0x100u : 0u;
        hdr->magic = kGeomMagic; hdr->P = (uint32_t)P; hdr->num_rendered = N; hdr->num_runs = R; hdr->num_visible = V;
        hdr->sort_error = err; hdr->shift = (uint32_t)shift; hdr->sort_mode = ok ? 1u : 2u; hdr->max_bin = maxbin;
        hdr->comp32 = (shift + id_bits <= 32) ? 1u : 0u;
        runoff[V] = R;                                 // read by the run emission of the last block of sorted positions
        if (mailbox) {   // mapped pinned host memory: the host reads the totals without a DMA copy; posted PCIe writes stay
                         // ordered and the host reads after an event recorded behind this kernel
            volatile uint32_t* m = mailbox;
            m[0] = kGeomMagic; m[1] = (uint32_t)P; m[2] = N; m[3] = err; m[4] = R; m[5] = V; m[6] = ok ? 1u : 2u; m[7] = maxbin;
        }
    }
}

void launch_bin_scan(const GeomPtrs& g, int P, uint32_t* mailbox_dev, cudaStream_t s) {
    const int nseg = g.nbins / kSegBins;
    bin_scan_kernel<<<nseg, kSegBins, 0, s>>>(g.bins64, g.seg_tot);
    bin_totals_kernel<<<1, 1024, 0, s>>>(g.header, g.stat_slots, P, nseg, g.seg_tot, g.seg_pre, g.runoff, depth_shift(P), index_bits(P), mailbox_dev);
    count_launch(2);
}

// ------------------------------------------------------------------------------------------------
// scatter into the bins' segments
// ------------------------------------------------------------------------------------------------
// record = (composite, (y0 << 16) | rows, (x0 << 10) | width, id).  composite = position of the key inside its bin (low `shift`
// key bits) above the id bits when that fits 32 bits (hdr->comp32), so that one 32-bit compare orders (depth, index);
// otherwise the full key (ties then compare the id word).  Slot = first slot of the bin + the count the preprocess kernel's
// atomic returned: no atomics here.  One 256-Gaussian block per CTA iteration, kIlp blocks in flight.
__global__ void __launch_bounds__(kPreBlock) bin_scatter_kernel(const GeomHeader* __restrict__ hdr, int nblocks, int nseg, int id_bits,
                                                                const uint32_t* __restrict__ block_vis, const uint32_t* __restrict__ ckeys,
                                                                const uint32_t* __restrict__ cids, const uint32_t* __restrict__ crows,
                                                                const uint32_t* __restrict__ cxw, const uint32_t* __restrict__ crank,
                                                                const unsigned long long* __restrict__ bins64, const uint2* __restrict__ seg_pre,
                                                                uint4* __restrict__ brec) {
    if (hdr->sort_mode != 1u) return;
    const uint32_t shift = hdr->shift, comp32 = hdr->comp32;
    const uint32_t low = (1u << shift) - 1u;
    const int t = threadIdx.x;
    const int bstep = (int)gridDim.x;
    for (int b0 = (int)blockIdx.x; b0 < nblocks; b0 += bstep * kIlp) {
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
        for (int k = 0; k < kIlp; ++k) if (on[k]) slot[k] += __ldg(seg_pre + bin[k] / kSegBins).x + (uint32_t)__ldg(bins64 + bin[k]);
#pragma unroll
        for (int k = 0; k < kIlp; ++k) if (on[k]) brec[slot[k]] = rec[k];
    }
}

void launch_bin_scatter(const GeomPtrs& g, int P, cudaStream_t s) {
    int grid = (g.nblocks + kIlp - 1) / kIlp;
    const int cap = ds_num_sms() * 8;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    const int nseg = g.nbins / kSegBins;
    bin_scatter_kernel<<<grid, kPreBlock, 0, s>>>(g.header, g.nblocks, nseg, index_bits(P), g.block_sums_v, g.dkeys_a, g.order_tmp,
                                                  g.dkeys_b, g.order, g.crank, g.bins64, g.seg_pre, g.brec);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// exact order inside every bin + run offsets
// ------------------------------------------------------------------------------------------------
constexpr int kBinSortThreads = 256;
constexpr int kBinsPerCta = 128;               // consecutive bins one CTA orders (their records are one contiguous stretch of brec)
constexpr int kBinChunk = 2048;                // records staged in shared memory at a time (whole bins; a bin holds <= kBinCap)

// A CTA per 128 consecutive bins.  Their records are contiguous in brec: the CTA stages them in shared memory with coalesced
// 128-bit loads (whole bins, up to 2048 records at a time) and then works a THREAD PER RECORD: the thread finds its bin in the
// CTA's table of bin starts (binary search), counts the records of that bin that order before its own -- (composite, rows) pairs
// read from shared memory, n compares -- and, in the same loop, the runs those records emit; the record's id, the absolute
// offset of its first run and its rect go to slot start + rank.  (Per-thread or per-warp bins reading global memory are bound by
// the L1's one-sector-per-clock rate on scattered accesses: two earlier forms of this kernel took 70 - 250 us.)
__global__ void __launch_bounds__(kBinSortThreads) bin_sort_kernel(const GeomHeader* __restrict__ hdr, const unsigned long long* __restrict__ bins64,
                                                                   const uint4* __restrict__ seg_tot, const uint2* __restrict__ seg_pre,
                                                                   int nbins, const uint4* __restrict__ brec, uint32_t* __restrict__ order,
                                                                   uint32_t* __restrict__ runoff, uint2* __restrict__ srect) {
    __shared__ uint32_t s_start[kBinsPerCta + 1];     // first record of every bin of this CTA (absolute), + end
    __shared__ uint32_t s_run0[kBinsPerCta];          // first run of every bin (absolute)
    __shared__ uint2 s_rec[kBinChunk];                // (composite, rows) of the staged records
    __shared__ uint32_t s_tie[kBinChunk];             // id word (tie break when the composite is the full key)
    __shared__ int s_b1;
    if (hdr->sort_mode != 1u) return;
    const bool comp32 = hdr->comp32 != 0u;
    const int tid = threadIdx.x;
    const int bin0 = blockIdx.x * kBinsPerCta;
    const int seg = bin0 / kSegBins;
    const uint4 st = __ldg(seg_tot + seg);
    if (st.x == 0u) return;                           // empty segment (its bins64 words were never rewritten)
    const uint2 sp = __ldg(seg_pre + seg);
    if (tid < kBinsPerCta) {
        const unsigned long long sr = __ldg(bins64 + bin0 + tid);
        s_start[tid] = sp.x + (uint32_t)sr; s_run0[tid] = sp.y + (uint32_t)(sr >> 32);
    }
    if (tid == kBinsPerCta) {                         // end of the CTA's last bin: start of the next bin, or the segment's end
        const int nb = bin0 + kBinsPerCta;
        s_start[kBinsPerCta] = (nb % kSegBins) ? sp.x + (uint32_t)__ldg(bins64 + nb) : sp.x + st.x;
    }
    __syncthreads();
    if (s_start[kBinsPerCta] == s_start[0]) return;   // no records in these bins
    int b0 = 0;
    while (b0 < kBinsPerCta) {
        // chunk = bins [b0, b1): as many whole bins as fit kBinChunk records (at least one: a bin holds <= kBinCap < kBinChunk)
        if (tid == 0) {
            const uint32_t lim = s_start[b0] + kBinChunk;
            int lo = b0 + 1, hi = kBinsPerCta;        // largest b1 in [b0 + 1, 128] with s_start[b1] <= lim
            while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (s_start[mid] <= lim) lo = mid; else hi = mid - 1; }
            s_b1 = lo;
        }
        __syncthreads();
        const int b1 = s_b1;
        const uint32_t c0 = s_start[b0], cn = s_start[b1] - c0;
        for (uint32_t i = tid; i < cn; i += kBinSortThreads) {
            const uint4 q = __ldg(brec + c0 + i);
            s_rec[i] = make_uint2(q.x, q.y & 0xffffu); s_tie[i] = q.w;
        }
        __syncthreads();
        for (uint32_t i = tid; i < cn; i += kBinSortThreads) {
            const uint32_t at0 = c0 + i;
            int lo = b0, hi = b1;                     // bin of record at0: s_start[lo] <= at0 < s_start[lo + 1]
            while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (s_start[mid] <= at0) lo = mid; else hi = mid; }
            const uint32_t s = s_start[lo] - c0, e = s_start[lo + 1] - c0;
            const uint2 me = s_rec[i];
            uint32_t rank = 0, before = 0;
            if (comp32) {
                for (uint32_t j = s; j < e; ++j) { const uint2 o = s_rec[j]; const bool lt = o.x < me.x; rank += lt ? 1u : 0u; before += lt ? o.y : 0u; }
            } else {
                const uint32_t myid = s_tie[i];
                for (uint32_t j = s; j < e; ++j) {
                    const uint2 o = s_rec[j];
                    const bool lt = o.x < me.x || (o.x == me.x && s_tie[j] < myid);
                    rank += lt ? 1u : 0u; before += lt ? o.y : 0u;
                }
            }
            const uint4 q = __ldg(brec + at0);        // L1 hit: id, rect words
            const uint32_t at = c0 + s + rank;
            order[at] = q.w; runoff[at] = s_run0[lo] + before; srect[at] = make_uint2(q.z, q.y >> 16);
        }
        __syncthreads();
        b0 = b1;
    }
}

void launch_bin_sort(const GeomPtrs& g, int P, cudaStream_t s) {
    bin_sort_kernel<<<g.nbins / kBinsPerCta, kBinSortThreads, 0, s>>>(g.header, g.bins64, g.seg_tot, g.seg_pre, g.nbins, g.brec, g.order, g.runoff, g.srect);
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
