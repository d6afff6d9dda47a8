// gsr_depthsort.cu -- bucketed depth sort of the visible Gaussians for sm_100a.
//
// First half of the hierarchical replacement of cub::DeviceRadixSort::SortPairs<u64,u32> on (tile << 32 | depth) keys
// (cuda_rasterizer/rasterizer_impl.cu:303-308): the VISIBLE Gaussians are ordered by (depth bits, index) -- the order in
// which CUB's stable LSD passes leave equal-tile pairs (ties on equal depth keep the ascending index of duplicateWithKeys,
// rasterizer_impl.cu:70-111) -- and the run emission / row partition / expansion of gsr_binning.cu turn that order into the
// per-tile lists.
//
// Round 1 ran 4 global onesweep passes over the 1.4 M keys of a 1080p frame (+ compaction, histogram, block sums, two
// single-CTA scans: ten dependent launches, 0.14 ms, latency-bound at 0.11 of the HBM roofline).  Here the keys are
// range-partitioned ONCE into buckets of at most kBucketCap keys and every bucket is sorted by one CTA in shared memory:
//
//   preprocess_fwd_kernel   (gsr_preprocess.cu) writes an UNORDERED compaction (key, id) of the visible Gaussians (one
//                           atomic per CTA reserves its slots) and the key range [key_min, key_max]
//   bucket_hist_kernel      histogram of (key - key_min) >> shift over depth_bins(P) bins (global atomics); the last CTA
//                           scans it and cuts it into buckets: bucket b = bins whose exclusive prefix lies in
//                           [b*Cq, (b+1)*Cq), Cq = kBucketCap - largest bin, so no bucket exceeds kBucketCap
//   bucket_scatter_kernel   (key, id) -> its bucket's segment, slot = atomic cursor (order inside a bucket: arbitrary)
//   bucket_sort_kernel      one CTA per bucket: stable LSD radix passes in shared memory on the composite
//                           (key - bucket's first key, id) -- id bits first, so the result is the canonical
//                           (depth, index) order whatever the arrival order was; then the exclusive prefix of the tile
//                           rows each Gaussian spans (its run count) inside the bucket; the last CTA scans the buckets'
//                           run totals.  Replaces order_block_sums + scan.
//
// A bin holding more than kBucketCap - kBucketMinQuantum keys (thousands of Gaussians within 1/depth_bins of the depth
// range, or all-equal depths) cannot be split by range: bucket_hist_kernel then sets sort_mode = 2, the two kernels
// behind it leave at once, and the host runs the round-1 onesweep chain instead (gsr_api.cu), followed by
// bucket_sort_kernel in `presorted` mode on uniform buckets, so that everything downstream sees one format.
#include "gsr_common.cuh"

namespace gsr {

static int ds_num_sms() {
    static int n = 0;
    if (n == 0) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); if (n <= 0) n = 148; }
    return n;
}

// ------------------------------------------------------------------------------------------------
// histogram + bucket plan
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) bucket_hist_kernel(GeomHeader* hdr, const uint32_t* __restrict__ ckeys, uint32_t* bhist, int nbins,
                                                           uint32_t* __restrict__ bucket_start, int maxb, uint32_t* mailbox) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_last, s_carry, s_maxbin;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t V = hdr->num_visible;              // written by the last CTA of preprocess_fwd_kernel
    const uint32_t kmin = ~hdr->key_min_inv, kmax = hdr->key_max;
    const uint32_t range = (V && kmax >= kmin) ? kmax - kmin : 0u;
    int nb_log = 0; while ((1 << nb_log) < nbins) ++nb_log;
    const int bits = 32 - __clz(range);               // range == 0 -> 0
    const uint32_t shift = bits > nb_log ? (uint32_t)(bits - nb_log) : 0u;
    for (uint32_t i = blockIdx.x * 1024u + tid; i < V; i += gridDim.x * 1024u)
        atomicAdd(&bhist[(__ldg(ckeys + i) - kmin) >> shift], 1u);
    __syncthreads();
    if (tid == 0) { __threadfence(); s_last = (atomicAdd(&hdr->done_hist, 1u) == gridDim.x - 1) ? 1u : 0u; }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

    // ---- last CTA: largest bin, exclusive scan, bucket cut
    volatile uint32_t* h = bhist;
    uint32_t mx = 0;
    for (int i = tid; i < nbins; i += 1024) mx = max(mx, h[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) s_warp[wid] = mx;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    if (tid == 0) { uint32_t m = 0; for (int w = 0; w < 32; ++w) m = max(m, s_warp[w]); s_maxbin = m; }
    __syncthreads();
    const uint32_t maxbin = s_maxbin;
    const bool overflow = maxbin > (uint32_t)(kBucketCap - kBucketMinQuantum);
    const uint32_t Cq = overflow ? (uint32_t)kBucketMinQuantum : (uint32_t)kBucketCap - maxbin;
    const uint32_t nbk = V ? (V - 1) / Cq + 1 : 0u;
    if (!overflow && nbk <= (uint32_t)maxb) {
        // exclusive scan of the bins, 1024 at a time; bin i belongs to bucket prefix[i] / Cq, which replaces its count in
        // place.  Bucket ids skipped by a large bin get empty segments (start[b] = start[b+1]).
        uint32_t prev_b_carry = 0xffffffffu;          // bucket of the bin before this tile's first one (-1: none)
        for (int base = 0; base < nbins; base += 1024) {
            const int i = base + tid;
            const uint32_t c = i < nbins ? h[i] : 0u;
            uint32_t inc = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
            __syncthreads();                           // s_warp / s_carry of the previous tile are consumed
            if (lane == 31) s_warp[wid] = inc;
            __syncthreads();
            uint32_t wbase = 0, tot = 0;
#pragma unroll
            for (int w = 0; w < 32; ++w) { const uint32_t x = s_warp[w]; wbase += (w < wid) ? x : 0u; tot += x; }
            const uint32_t carry = s_carry;
            const uint32_t excl = carry + wbase + inc - c;
            const uint32_t b = min(excl / Cq, nbk ? nbk - 1 : 0u);     // (excl == V only for trailing empty bins)
            // bucket of the previous bin (previous lane / warp / tile)
            uint32_t pb = __shfl_up_sync(0xffffffffu, b, 1);
            __syncthreads();
            if (lane == 31) s_warp[wid] = b;
            __syncthreads();
            if (lane == 0) pb = wid ? s_warp[wid - 1] : prev_b_carry;
            if (i < nbins) {
                h[i] = b;
                for (uint32_t q = pb + 1u; q <= b; ++q) bucket_start[q] = excl;   // pb == -1 wraps to 0; an empty bin's
                                                                                  // prefix is its successor's, so it may open a bucket
            }
            prev_b_carry = s_warp[31];
            __syncthreads();
            if (tid == 0) s_carry = carry + tot;
        }
        __syncthreads();
        // bucket ids past the last bin's (skipped by a large last bin) are empty
        if (tid == 0) for (uint32_t q = prev_b_carry + 1u; q <= nbk; ++q) bucket_start[q] = V;
        if (tid == 0) bucket_start[nbk] = V;
    }
    if (tid == 0) {
        const bool ok = !overflow && nbk <= (uint32_t)maxb;
        hdr->shift = shift; hdr->num_buckets = ok ? nbk : 0u; hdr->bucket_quantum = Cq; hdr->sort_mode = ok ? 1u : 2u;
        if (mailbox) { volatile uint32_t* m = mailbox; m[6] = ok ? 1u : 2u; m[7] = nbk; }
    }
}

void launch_bucket_hist(const GeomPtrs& g, int P, uint32_t* mailbox_dev, cudaStream_t s) {
    const long long want = ((long long)P + 1024 * 8 - 1) / (1024 * 8);
    int grid = (int)(want < 1 ? 1 : (want > ds_num_sms() * 2 ? ds_num_sms() * 2 : want));
    bucket_hist_kernel<<<grid, 1024, 0, s>>>(g.header, g.dkeys_a, g.bhist, g.nbins, g.bucket_start, g.maxb, mailbox_dev);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// scatter into the buckets' segments
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bucket_scatter_kernel(const GeomHeader* __restrict__ hdr, const uint32_t* __restrict__ ckeys,
                                                             const uint32_t* __restrict__ cids, const uint32_t* __restrict__ bin2bucket,
                                                             const uint32_t* __restrict__ bucket_start, uint32_t* bucket_cursor,
                                                             uint32_t* __restrict__ bkeys, uint32_t* __restrict__ bids) {
    if (hdr->sort_mode != 1u) return;
    const uint32_t V = hdr->num_visible, kmin = ~hdr->key_min_inv, shift = hdr->shift;
    for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < V; i += gridDim.x * 256u) {
        const uint32_t key = __ldg(ckeys + i), id = __ldg(cids + i);
        const uint32_t b = __ldg(bin2bucket + ((key - kmin) >> shift));
        const uint32_t pos = __ldg(bucket_start + b) + atomicAdd(&bucket_cursor[b], 1u);
        bkeys[pos] = key; bids[pos] = id;
    }
}

void launch_bucket_scatter(const GeomPtrs& g, int P, cudaStream_t s) {
    const long long want = ((long long)P + 255) / 256;
    const long long cap = (long long)ds_num_sms() * 16;
    bucket_scatter_kernel<<<(unsigned)(want < cap ? (want < 1 ? 1 : want) : cap), 256, 0, s>>>(
        g.header, g.dkeys_a, g.order_tmp, g.bhist, g.bucket_start, g.bucket_cursor, g.dkeys_b, g.bids);
    count_launch();
}

// uniform buckets of kBucketCap sorted positions (fallback path: the onesweep chain has already produced `order`)
__global__ void __launch_bounds__(256) uniform_buckets_kernel(GeomHeader* hdr, uint32_t* __restrict__ bucket_start, int maxb) {
    const uint32_t V = hdr->num_visible;
    const uint32_t nbk = (V + kBucketCap - 1) / kBucketCap;
    for (uint32_t b = blockIdx.x * 256u + threadIdx.x; b <= nbk && b <= (uint32_t)maxb; b += gridDim.x * 256u)
        bucket_start[b] = min(V, b * (uint32_t)kBucketCap);
    if (blockIdx.x == 0 && threadIdx.x == 0) { hdr->num_buckets = nbk; hdr->sort_mode = 3u; hdr->done_sort = 0u; }
}

void launch_uniform_buckets(const GeomPtrs& g, int P, cudaStream_t s) {
    const int nbk = (P + kBucketCap - 1) / kBucketCap + 1;
    uniform_buckets_kernel<<<(nbk + 255) / 256, 256, 0, s>>>(g.header, g.bucket_start, g.maxb);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// per-bucket sort in shared memory
// ------------------------------------------------------------------------------------------------
// lanes whose digit equals mine: one ballot per digit bit (constant cost whatever the digit diversity)
__device__ __forceinline__ uint32_t match_bits(uint32_t d, int nbits) {
    uint32_t peers = 0xffffffffu;
    for (int b = 0; b < nbits; ++b) {
        const uint32_t bal = __ballot_sync(0xffffffffu, (d >> b) & 1u);
        peers &= ((d >> b) & 1u) ? bal : ~bal;
    }
    return peers;
}

constexpr int kBsWarps = kBucketThreads / 32;
constexpr int kBsRadix = 256;

struct BucketSortParams {
    GeomHeader* hdr;
    const uint32_t* __restrict__ bkeys; const uint32_t* __restrict__ bids;
    const uint32_t* __restrict__ bucket_start;
    const uint2* __restrict__ rects;
    uint32_t* __restrict__ order; uint32_t* __restrict__ runoff;
    uint32_t* bucket_runs; uint32_t* bucket_base;
    int id_bits;          // bits needed for a Gaussian index (32 - clz(P - 1))
    int presorted;        // 1: `order` already holds the sorted ids (fallback); only the run offsets are computed
};

// one stable LSD pass over the CTA's items (warp-striped in registers) on `nbits` bits of the 64-bit composite starting
// at bit `shift` (composite = rel key << 32 | id); the items come back warp-striped in the new order.
__device__ __forceinline__ void smem_radix_pass(uint32_t (&key)[kBucketItems], uint32_t (&id)[kBucketItems], int shift, int nbits,
                                                uint32_t* s_k, uint32_t* s_i, uint32_t* s_whist, uint32_t* s_bin, uint32_t* s_wtot) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t mask = (1u << nbits) - 1u;
    const int radix = 1 << nbits;
    for (int i = tid; i < kBsWarps * radix; i += kBucketThreads) s_whist[i] = 0;
    __syncthreads();
    uint32_t* wh = s_whist + wid * radix;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t rank[kBucketItems];
#pragma unroll
    for (int i = 0; i < kBucketItems; ++i) {
        const uint32_t d = (shift >= 32 ? (key[i] >> (shift - 32)) : (id[i] >> shift)) & mask;
        const uint32_t peers = match_bits(d, nbits);
        const uint32_t pre = wh[d];
        __syncwarp();
        if ((peers & lt) == 0) wh[d] = pre + __popc(peers);
        __syncwarp();
        rank[i] = pre + __popc(peers & lt);
    }
    __syncthreads();
    // per bin: exclusive scan across warps, then across bins
    uint32_t cnt = 0;
    if (tid < radix) {
        uint32_t acc = 0;
#pragma unroll
        for (int w = 0; w < kBsWarps; ++w) { const uint32_t t = s_whist[w * radix + tid]; s_whist[w * radix + tid] = acc; acc += t; }
        cnt = acc;
    }
    {
        uint32_t inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
        if (tid < radix && lane == 31) s_wtot[wid] = inc;
        __syncthreads();
        if (tid < radix) {
            uint32_t b = 0;
            for (int w = 0; w < (radix + 31) / 32; ++w) b += (w < wid) ? s_wtot[w] : 0u;
            s_bin[tid] = b + inc - cnt;
        }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kBucketItems; ++i) {
        const uint32_t d = (shift >= 32 ? (key[i] >> (shift - 32)) : (id[i] >> shift)) & mask;
        const uint32_t pos = s_bin[d] + wh[d] + rank[i];
        s_k[pos] = key[i]; s_i[pos] = id[i];
    }
    __syncthreads();
    const int wbase = wid * 32 * kBucketItems;
#pragma unroll
    for (int i = 0; i < kBucketItems; ++i) { key[i] = s_k[wbase + i * 32 + lane]; id[i] = s_i[wbase + i * 32 + lane]; }
    __syncthreads();
}

__global__ void __launch_bounds__(kBucketThreads, 2) bucket_sort_kernel(const BucketSortParams p) {
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_k = smem;                                // [kBucketCap]
    uint32_t* s_i = s_k + kBucketCap;                    // [kBucketCap]
    uint32_t* s_whist = s_i + kBucketCap;                // [kBsWarps][kBsRadix]
    uint32_t* s_bin = s_whist + kBsWarps * kBsRadix;     // [kBsRadix]
    __shared__ uint32_t s_wtot[32];
    __shared__ uint32_t s_red[kBsWarps];
    __shared__ uint32_t s_flag;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t mode = p.hdr->sort_mode;
    if (mode != (p.presorted ? 3u : 1u)) return;
    const uint32_t nbk = p.hdr->num_buckets;

    for (uint32_t b = blockIdx.x; b < nbk; b += gridDim.x) {
        const uint32_t start = __ldg(p.bucket_start + b), end = __ldg(p.bucket_start + b + 1);
        const int n = (int)(end - start);                // <= kBucketCap by construction
        uint32_t key[kBucketItems], id[kBucketItems];
        const int wbase = wid * 32 * kBucketItems;
        if (!p.presorted) {
            uint32_t kmin = 0xffffffffu, kmaxv = 0u;
#pragma unroll
            for (int i = 0; i < kBucketItems; ++i) {
                const int lp = wbase + i * 32 + lane;
                key[i] = lp < n ? __ldg(p.bkeys + start + lp) : 0xffffffffu;
                id[i] = lp < n ? __ldg(p.bids + start + lp) : 0xffffffffu;
                if (lp < n) { kmin = min(kmin, key[i]); kmaxv = max(kmaxv, key[i]); }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, o)); kmaxv = max(kmaxv, __shfl_xor_sync(0xffffffffu, kmaxv, o)); }
            if (lane == 0) { s_red[wid] = kmin; s_wtot[wid] = kmaxv; }
            __syncthreads();
            kmin = 0xffffffffu; kmaxv = 0u;
#pragma unroll
            for (int w = 0; w < kBsWarps; ++w) { kmin = min(kmin, s_red[w]); kmaxv = max(kmaxv, s_wtot[w]); }
            __syncthreads();
            // relative keys: padding items keep 0xffffffff (and id 0xffffffff) -> they sort last in every pass
            const uint32_t span = n ? kmaxv - kmin : 0u;
#pragma unroll
            for (int i = 0; i < kBucketItems; ++i) if (wbase + i * 32 + lane < n) key[i] -= kmin;
            const int key_bits = 32 - __clz(span);       // 0 when all keys are equal
            // id bits first (least significant part of the composite), then the relative key bits; 8 bits per pass at most,
            // spread evenly.  Padding items: all-ones digits in every pass -> stay behind the real ones.
            {
                const int np = (p.id_bits + 7) / 8;
                for (int ps = 0, sh = 0; ps < np; ++ps) {
                    const int nb = (p.id_bits - sh + (np - ps) - 1) / (np - ps);
                    smem_radix_pass(key, id, sh, nb, s_k, s_i, s_whist, s_bin, s_wtot);
                    sh += nb;
                }
            }
            {
                const int np = (key_bits + 7) / 8;
                for (int ps = 0, sh = 0; ps < np; ++ps) {
                    const int nb = (key_bits - sh + (np - ps) - 1) / (np - ps);
                    smem_radix_pass(key, id, 32 + sh, nb, s_k, s_i, s_whist, s_bin, s_wtot);
                    sh += nb;
                }
            }
            // blocked order for the scan below: through shared memory
#pragma unroll
            for (int i = 0; i < kBucketItems; ++i) s_i[wbase + i * 32 + lane] = id[i];
            __syncthreads();
        } else {
            for (int j = tid; j < kBucketCap; j += kBucketThreads) s_i[j] = j < n ? __ldg(p.order + start + j) : 0xffffffffu;
            __syncthreads();
        }
        // ---- run counts (tile rows spanned) in sorted order, exclusive prefix inside the bucket
        uint32_t h[kBucketItems], sum = 0;
#pragma unroll
        for (int i = 0; i < kBucketItems; ++i) {
            const int j = tid * kBucketItems + i;
            const uint32_t g = s_i[j];
            id[i] = g;
            uint32_t rows = 0;
            if (j < n) { const uint32_t ry = __ldg(&p.rects[g].y); rows = (ry >> 16) - (ry & 0xffffu); }
            h[i] = rows; sum += rows;
        }
        uint32_t inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
        if (lane == 31) s_red[wid] = inc;
        __syncthreads();
        uint32_t wb = 0, tot = 0;
#pragma unroll
        for (int w = 0; w < kBsWarps; ++w) { const uint32_t x = s_red[w]; wb += (w < wid) ? x : 0u; tot += x; }
        uint32_t run = wb + inc - sum;
#pragma unroll
        for (int i = 0; i < kBucketItems; ++i) {
            const int j = tid * kBucketItems + i;
            if (j < n) { if (!p.presorted) p.order[start + j] = id[i]; p.runoff[start + j] = run; }
            run += h[i];
        }
        if (tid == 0) p.bucket_runs[b] = tot;
        __syncthreads();
    }

    // ---- the CTA that finishes last turns the buckets' run totals into exclusive offsets
    if (tid == 0) { __threadfence(); s_flag = (atomicAdd(&p.hdr->done_sort, 1u) == gridDim.x - 1) ? 1u : 0u; }
    __syncthreads();
    if (!s_flag) return;
    __threadfence();
    volatile uint32_t* br = p.bucket_runs;
    uint32_t carry = 0;
    for (uint32_t base = 0; base < nbk; base += kBucketThreads) {
        const uint32_t i = base + tid;
        const uint32_t c = i < nbk ? br[i] : 0u;
        uint32_t inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
        __syncthreads();
        if (lane == 31) s_red[wid] = inc;
        __syncthreads();
        uint32_t wb = 0, tot = 0;
#pragma unroll
        for (int w = 0; w < kBsWarps; ++w) { const uint32_t x = s_red[w]; wb += (w < wid) ? x : 0u; tot += x; }
        if (i < nbk) p.bucket_base[i] = carry + wb + inc - c;
        carry += tot;
    }
    if (tid == 0) p.bucket_base[nbk] = carry;
}

void launch_bucket_sort(const GeomPtrs& g, int P, bool presorted, cudaStream_t s) {
    BucketSortParams p;
    p.hdr = g.header; p.bkeys = g.dkeys_b; p.bids = g.bids; p.bucket_start = g.bucket_start; p.rects = g.rects;
    p.order = g.order; p.runoff = g.runoff; p.bucket_runs = g.bucket_runs; p.bucket_base = g.bucket_base;
    int idb = 1; while (idb < 32 && (1ll << idb) < (long long)P) ++idb;
    p.id_bits = idb; p.presorted = presorted ? 1 : 0;
    const size_t smem = sizeof(uint32_t) * (2 * (size_t)kBucketCap + kBsWarps * kBsRadix + kBsRadix);
    static bool attr_done = false;
    if (!attr_done) { cudaFuncSetAttribute(bucket_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); attr_done = true; }
    // grid: the bucket count lives on the device; buckets are handed out round-robin to at most 2 CTAs per SM
    long long want = (long long)P / (kBucketCap / 2) + 1;
    const long long cap = (long long)ds_num_sms() * 2;
    bucket_sort_kernel<<<(unsigned)(want < cap ? want : cap), kBucketThreads, smem, s>>>(p);
    count_launch();
}

}  // namespace gsr
