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
// histogram
// ------------------------------------------------------------------------------------------------
constexpr int kHistThreads = 1024;
constexpr int kSegBins = 1024;                 // bins per scan segment (one CTA of bin_scan_kernel)
constexpr int kIlp = 4;                        // 256-blocks a thread has in flight (the loops are chains of dependent L2 round trips)

// key range from the statistics slots: every CTA reduces the 128 slots itself (128 loads)
struct KeyRange { uint32_t kmin, shift, pairs, trap; };
__device__ __forceinline__ KeyRange reduce_stat_slots(const FrameStatSlot* __restrict__ slots, int nbins, uint32_t* s_a, uint32_t* s_b,
                                                      uint32_t* s_c, uint32_t* s_d) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    uint32_t kmx = 0, kmn = 0, npairs = 0, trap = 0;
    if (tid < kStatSlots) { const FrameStatSlot s = slots[tid]; kmx = s.key_max; kmn = s.key_min_inv; npairs = s.num_rendered; trap = s.flags; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        kmx = max(kmx, __shfl_xor_sync(0xffffffffu, kmx, o)); kmn = max(kmn, __shfl_xor_sync(0xffffffffu, kmn, o));
        npairs += __shfl_xor_sync(0xffffffffu, npairs, o); trap |= __shfl_xor_sync(0xffffffffu, trap, o);
    }
    if (lane == 0 && wid < kStatSlots / 32) { s_a[wid] = kmx; s_b[wid] = kmn; s_c[wid] = npairs; s_d[wid] = trap; }
    __syncthreads();
    uint32_t a = 0, b = 0; KeyRange r; r.pairs = 0; r.trap = 0;
#pragma unroll
    for (int w = 0; w < kStatSlots / 32; ++w) { a = max(a, s_a[w]); b = max(b, s_b[w]); r.pairs += s_c[w]; r.trap |= s_d[w]; }
    r.kmin = ~b;
    const uint32_t range = (b != 0u && a >= r.kmin) ? a - r.kmin : 0u;        // b == 0: nothing visible
    int nb_log = 0; while ((1 << nb_log) < nbins) ++nb_log;
    const int bits = 32 - __clz(range);
    r.shift = bits > nb_log ? (uint32_t)(bits - nb_log) : 0u;
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(kHistThreads) bin_hist_kernel(const FrameStatSlot* __restrict__ slots, int nblocks,
                                                                const uint32_t* __restrict__ block_vis, const uint32_t* __restrict__ ckeys,
                                                                const uint32_t* __restrict__ crows, uint32_t* bin_cur, uint32_t* bin_runs, int nbins) {
    __shared__ uint32_t s_a[4], s_b[4], s_c[4], s_d[4];
    const KeyRange kr = reduce_stat_slots(slots, nbins, s_a, s_b, s_c, s_d);
    // two REDs per visible Gaussian; a block's records sit at the front of its 256-slot segment
    const int sub = threadIdx.x >> 8, t = threadIdx.x & 255;
    const int bstep = (int)gridDim.x * 4;
    for (int b0 = (int)blockIdx.x * 4 + sub; b0 < nblocks; b0 += bstep * kIlp) {
        uint32_t key[kIlp], rows[kIlp]; bool on[kIlp];
#pragma unroll
        for (int k = 0; k < kIlp; ++k) { const int b = b0 + k * bstep; on[k] = b < nblocks && (uint32_t)t < __ldg(block_vis + b); }
#pragma unroll
        for (int k = 0; k < kIlp; ++k) if (on[k]) { const size_t i = (size_t)(b0 + k * bstep) * kPreBlock + t; key[k] = __ldg(ckeys + i); rows[k] = __ldg(crows + i) & 0xffffu; }
#pragma unroll
        for (int k = 0; k < kIlp; ++k) if (on[k]) { const uint32_t bin = (key[k] - kr.kmin) >> kr.shift; atomicAdd(&bin_cur[bin], 1u); atomicAdd(&bin_runs[bin], rows[k]); }
    }
}

// ------------------------------------------------------------------------------------------------
// scan: one CTA per segment of 1024 bins -- exclusive prefix INSIDE the segment (in place) + the segment's totals.
// Consumers add the exclusive prefix of the segment totals themselves (seg_base()).  The CTA that finishes last sums the
// totals and posts the frame's counts where the host reads them.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kSegBins) bin_scan_kernel(GeomHeader* hdr, const FrameStatSlot* __restrict__ slots, int P, int nbins,
                                                            uint32_t* bin_cur, uint32_t* bin_runs, uint2* seg_tot, uint32_t* runoff,
                                                            int id_bits, uint32_t* mailbox) {
    __shared__ uint32_t s_a[32], s_b[32], s_c[32], s_d[4];
    __shared__ uint32_t s_last;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int i = blockIdx.x * kSegBins + tid;
    const uint32_t c = bin_cur[i], r = bin_runs[i];
    uint32_t ic = c, ir = r, mx = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, ic, o), v = __shfl_up_sync(0xffffffffu, ir, o);
        if (lane >= o) { ic += u; ir += v; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 31) { s_a[wid] = ic; s_b[wid] = ir; }
    if (lane == 0) s_c[wid] = mx;
    __syncthreads();
    uint32_t bc = 0, br = 0, tc = 0, tr = 0, maxbin = 0;
#pragma unroll
    for (int w = 0; w < 32; ++w) {
        const uint32_t a = s_a[w], b = s_b[w];
        bc += (w < wid) ? a : 0u; br += (w < wid) ? b : 0u; tc += a; tr += b; maxbin = max(maxbin, s_c[w]);
    }
    bin_cur[i] = bc + ic - c;                  // the scatter's cursor, relative to the segment
    bin_runs[i] = br + ir - r;                 // first run of the bin, relative to the segment
    if (tid == 0) {
        seg_tot[blockIdx.x] = make_uint2(tc, tr);
        if (maxbin) atomicMax(&hdr->max_bin, maxbin);
        __threadfence();
        s_last = (atomicAdd(&hdr->done_hist, 1u) == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // ---- last CTA: frame totals
    __syncthreads();
    const KeyRange kr = reduce_stat_slots(slots, nbins, s_a, s_b, s_c, s_d);
    uint32_t v = 0, rr = 0;
    for (int k = tid; k < (int)gridDim.x; k += kSegBins) { const uint2 t2 = __ldcg(seg_tot + k); v += t2.x; rr += t2.y; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { v += __shfl_xor_sync(0xffffffffu, v, o); rr += __shfl_xor_sync(0xffffffffu, rr, o); }
    __syncthreads();
    if (lane == 0) { s_a[wid] = v; s_b[wid] = rr; }
    __syncthreads();
    if (tid == 0) {
        uint32_t V = 0, R = 0;
        for (int w = 0; w < 32; ++w) { V += s_a[w]; R += s_b[w]; }
        const uint32_t mb = atomicMax(&hdr->max_bin, 0u);
        const bool ok = mb <= (uint32_t)kBinCap;
        const uint32_t err = kr.trap ? 0x100u : 0u;
        hdr->magic = kGeomMagic; hdr->P = (uint32_t)P; hdr->num_rendered = kr.pairs; hdr->num_runs = R; hdr->num_visible = V;
        hdr->sort_error = err; hdr->key_min = kr.kmin; hdr->shift = kr.shift; hdr->sort_mode = ok ? 1u : 2u;
        hdr->comp32 = (kr.shift + (uint32_t)id_bits <= 32u) ? 1u : 0u;
        runoff[V] = R;                                 // read by the run emission of the last block of sorted positions
        if (mailbox) {   // mapped pinned host memory: the host reads the totals without a DMA copy; posted PCIe writes stay
                         // ordered and the host reads after an event recorded behind this kernel
            volatile uint32_t* m = mailbox;
            m[0] = kGeomMagic; m[1] = (uint32_t)P; m[2] = kr.pairs; m[3] = err; m[4] = R; m[5] = V; m[6] = ok ? 1u : 2u; m[7] = mb;
        }
    }
}

static int id_bits_of(int P) { int b = 1; while (b < 32 && (1ll << b) < (long long)P) ++b; return b; }

void launch_bin_hist(const GeomPtrs& g, int P, uint32_t* mailbox_dev, cudaStream_t s) {
    int grid = (g.nblocks + 4 * kIlp - 1) / (4 * kIlp);
    if (grid > ds_num_sms()) grid = ds_num_sms();
    if (grid < 1) grid = 1;
    bin_hist_kernel<<<grid, kHistThreads, 0, s>>>(g.stat_slots, g.nblocks, g.block_sums_v, g.dkeys_a, g.dkeys_b, g.bin_cur, g.bin_runs, g.nbins);
    bin_scan_kernel<<<g.nbins / kSegBins, kSegBins, 0, s>>>(g.header, g.stat_slots, P, g.nbins, g.bin_cur, g.bin_runs, g.seg_tot, g.runoff,
                                                           id_bits_of(P), mailbox_dev);
    count_launch(2);
}

// exclusive prefix of the segment totals, by every consumer CTA for itself (<= 256 segments): s_seg[k] = (records, runs) before segment k
__device__ __forceinline__ void seg_base(const uint2* __restrict__ seg_tot, int nseg, uint2* s_seg) {
    __shared__ uint2 s_w[8];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    uint2 t = make_uint2(0u, 0u), inc = make_uint2(0u, 0u);
    if (tid < 256) {
        if (tid < nseg) t = __ldcg(seg_tot + tid);
        inc = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, inc.x, o), v = __shfl_up_sync(0xffffffffu, inc.y, o);
            if (lane >= o) { inc.x += u; inc.y += v; }
        }
        if (lane == 31) s_w[wid] = inc;
    }
    __syncthreads();
    if (tid < 256) {
        uint2 b = make_uint2(0u, 0u);
#pragma unroll
        for (int w = 0; w < 8; ++w) if (w < wid) { b.x += s_w[w].x; b.y += s_w[w].y; }
        s_seg[tid] = make_uint2(b.x + inc.x - t.x, b.y + inc.y - t.y);
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// scatter into the bins' segments
// ------------------------------------------------------------------------------------------------
// record = (composite, (y0 << 16) | rows, (x0 << 10) | width, id).  composite = position of the key inside its bin (low `shift`
// key bits) above the id bits when that fits 32 bits (hdr->comp32), so that one 32-bit compare orders (depth, index);
// otherwise the full key (ties then compare the id word).
__global__ void __launch_bounds__(kHistThreads) bin_scatter_kernel(const GeomHeader* __restrict__ hdr, int nblocks, int nseg, int id_bits,
                                                                   const uint32_t* __restrict__ block_vis, const uint32_t* __restrict__ ckeys,
                                                                   const uint32_t* __restrict__ cids, const uint32_t* __restrict__ crows,
                                                                   const uint32_t* __restrict__ cxw, uint32_t* bin_cur,
                                                                   const uint2* __restrict__ seg_tot, uint4* __restrict__ brec) {
    __shared__ uint2 s_seg[256];
    if (hdr->sort_mode != 1u) return;
    seg_base(seg_tot, nseg, s_seg);
    const uint32_t kmin = hdr->key_min, shift = hdr->shift, comp32 = hdr->comp32;
    const uint32_t low = shift ? ((1u << shift) - 1u) : 0u;
    const int sub = threadIdx.x >> 8, t = threadIdx.x & 255;
    const int bstep = (int)gridDim.x * 4;
    for (int b0 = (int)blockIdx.x * 4 + sub; b0 < nblocks; b0 += bstep * kIlp) {
        uint4 rec[kIlp]; uint32_t bin[kIlp], pos[kIlp]; bool on[kIlp];
#pragma unroll
        for (int k = 0; k < kIlp; ++k) { const int b = b0 + k * bstep; on[k] = b < nblocks && (uint32_t)t < __ldg(block_vis + b); }
#pragma unroll
        for (int k = 0; k < kIlp; ++k) if (on[k]) {
            const size_t i = (size_t)(b0 + k * bstep) * kPreBlock + t;
            const uint32_t key = __ldg(ckeys + i), id = __ldg(cids + i);
            const uint32_t rel = key - kmin;
            bin[k] = rel >> shift;
            rec[k] = make_uint4(comp32 ? (((rel & low) << id_bits) | id) : key, __ldg(crows + i), __ldg(cxw + i), id);
        }
#pragma unroll
        for (int k = 0; k < kIlp; ++k) if (on[k]) pos[k] = s_seg[bin[k] / kSegBins].x + atomicAdd(&bin_cur[bin[k]], 1u);
#pragma unroll
        for (int k = 0; k < kIlp; ++k) if (on[k]) brec[pos[k]] = rec[k];
    }
}

void launch_bin_scatter(const GeomPtrs& g, int P, cudaStream_t s) {
    int grid = (g.nblocks + 4 * kIlp - 1) / (4 * kIlp);
    const int cap = ds_num_sms() * 2;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    bin_scatter_kernel<<<grid, kHistThreads, 0, s>>>(g.header, g.nblocks, g.nbins / kSegBins, id_bits_of(P), g.block_sums_v, g.dkeys_a, g.order_tmp,
                                                     g.dkeys_b, g.order, g.bin_cur, g.seg_tot, g.brec);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// exact order inside every bin + run offsets
// ------------------------------------------------------------------------------------------------
constexpr int kBinSortWarps = 8;

// A warp per bin.  Every lane ranks its record against the bin's others and PUSHES its outputs to slot start + rank (the bin's
// slots are one short contiguous stretch); the number of runs emitted by the records ordered before it -- the exclusive
// prefix of the tile rows in sorted order -- is accumulated in the same loop, so no inverse permutation and no scan is needed.
// No shared memory beyond the segment table: 64 warps per SM hide the three dependent L2 round trips a bin costs.
__global__ void __launch_bounds__(kBinSortWarps * 32) bin_sort_kernel(const GeomHeader* __restrict__ hdr, const uint32_t* __restrict__ bin_end,
                                                                      const uint32_t* __restrict__ bin_runs, const uint2* __restrict__ seg_tot,
                                                                      int nbins, const uint4* __restrict__ brec, uint32_t* __restrict__ order,
                                                                      uint32_t* __restrict__ runoff, uint2* __restrict__ srect) {
    __shared__ uint2 s_seg[256];
    if (hdr->sort_mode != 1u) return;
    seg_base(seg_tot, nbins / kSegBins, s_seg);
    const bool comp32 = hdr->comp32 != 0u;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int warp_global = blockIdx.x * kBinSortWarps + wid, warps_total = gridDim.x * kBinSortWarps;
    // bins are dealt round-robin: neighbouring (similarly loaded) bins go to different warps
    for (int bin = warp_global; bin < nbins; bin += warps_total) {
        const uint2 sb = s_seg[bin / kSegBins];
        const uint32_t end_l = __ldg(bin_end + bin), start_l = (bin % kSegBins) ? __ldg(bin_end + bin - 1) : 0u;   // after the scatter a cursor = its bin's end
        const int n = (int)(end_l - start_l);
        if (n <= 0) continue;
        const uint32_t start = sb.x + start_l;
        const uint32_t run0 = sb.y + __ldg(bin_runs + bin);
        if (n <= 32) {
            uint4 r = make_uint4(0xffffffffu, 0u, 0u, 0xffffffffu);
            if (lane < n) r = __ldg(brec + start + lane);
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
            // 33 .. kBinCap records: the same all-pairs loop with the bin's (composite, rows, id) re-read through the L1
            for (int i = lane; i < n; i += 32) {
                const uint4 r = __ldg(brec + start + i);
                uint32_t rank = 0, before = 0;
                for (int j = 0; j < n; ++j) {
                    const uint4 o = __ldg(brec + start + j);
                    const bool lt = o.x < r.x || (!comp32 && o.x == r.x && o.w < r.w);
                    rank += lt ? 1u : 0u; before += lt ? (o.y & 0xffffu) : 0u;
                }
                const uint32_t at = start + rank;
                order[at] = r.w; runoff[at] = run0 + before; srect[at] = make_uint2(r.z, r.y >> 16);
            }
        }
    }
}

void launch_bin_sort(const GeomPtrs& g, int P, cudaStream_t s) {
    int grid = (g.nbins + kBinSortWarps * 4 - 1) / (kBinSortWarps * 4);        // >= 4 bins per warp
    const int cap = ds_num_sms() * 8;                                           // 8 CTAs x 8 warps per SM
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    bin_sort_kernel<<<grid, kBinSortWarps * 32, 0, s>>>(g.header, g.bin_cur, g.bin_runs, g.seg_tot, g.nbins, g.brec, g.order, g.runoff, g.srect);
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
