// gsr_binning.cu -- tile binning for sm_100a: key duplication, onesweep radix sort, tile ranges, AMR levels.
//
//   emit_pairs_kernel  replaces duplicateWithKeys (rasterizer_impl.cu:70-111) AND the second half of
//                      cub::DeviceScan::InclusiveSum (:277) AND the histogram pass of the sort: a persistent grid
//                      walks 256-Gaussian blocks IN DEPTH ORDER, scans their tiles_touched in shared memory, emits
//                      (tile id, Gaussian id) with fully coalesced stores (one thread per PAIR, not per Gaussian ->
//                      no divergence on fat Gaussians) and accumulates the digit histograms of the tile ids.
//   onesweep_kernel    replaces cub::DeviceRadixSort::SortPairs<u64,u32> (rasterizer_impl.cu:303-308): stable LSD
//                      radix sort, one read + one write of every pair per pass, chained-scan ("decoupled
//                      look-back") prefix across tiles.  Used hierarchically: 4 passes over the P depth keys, then
//                      2 passes over the N (tile, id) pairs -- 40 B/pair of traffic instead of 132 B/pair.
//   tile_ranges_kernel replaces cudaMemset + identifyTileRanges (rasterizer_impl.cu:116-138,310-317) and the AMR
//                      fork's calculateIntersections (A/rasterizer_impl.cu:181-188): per-tile binary search on
//                      the sorted keys (8 B x 2 log2 N per tile instead of 16 B per pair).
//   amr_levels_kernel  replaces SortKeys + cudaMalloc/cudaFree + 3x D2H + setAMRLevelsKernel
//                      (A/rasterizer_impl.cu:590-648): single-CTA radix select of the three percentile
//                      thresholds, thresholds never leave the device.
//   fovea_levels_kernel replaces setFoveaAMRLevelsKernel (A/rasterizer_impl.cu:208-243) + gaze extension (A7).
#include "gsr_common.cuh"
#include "gsr_math.cuh"

namespace gsr {

// ------------------------------------------------------------------------------------------------
// sort plan
// ------------------------------------------------------------------------------------------------
SortPlan make_sort_plan(int end_bit) {
    SortPlan pl;
    if (end_bit < 1) end_bit = 1;
    if (end_bit > 64) end_bit = 64;
    int passes = (end_bit + 8) / 9;                 // ceil(end_bit / 9): 9-bit digits max
    pl.passes = passes;
    int base = end_bit / passes, extra = end_bit % passes, sh = 0;
    for (int i = 0; i < passes; ++i) {
        int b = base + (i < extra ? 1 : 0);
        pl.shift[i] = sh; pl.bits[i] = b; sh += b;
    }
    return pl;
}

SortPlan make_sort_plan_range(int begin_bit, int nbits) {
    SortPlan pl = make_sort_plan(nbits);
    for (int i = 0; i < pl.passes; ++i) pl.shift[i] += begin_bit;
    return pl;
}

// ------------------------------------------------------------------------------------------------
// pair emission in depth order
// ------------------------------------------------------------------------------------------------
// The reference sorts N (tile, depth) pairs on 45+ key bits.  The same permutation is obtained far more cheaply
// hierarchically: (1) stable-sort the VISIBLE Gaussians (compacted, ascending index) by their 32-bit depth key, (2) emit every Gaussian's pairs in THAT order, (3) stable-partition the pairs by tile id
// (13-15 bits -> two radix passes on (u32 tile, u32 id) instead of five on (u64 key, u32 id)).  Ties between equal
// depths inside a tile fall back to ascending Gaussian index in both schemes, so point_list is bit-identical.
__global__ void __launch_bounds__(kPreBlock) compact_keys_kernel(int P, const float* __restrict__ depths,
                                                                 const uint32_t* __restrict__ tiles_touched,
                                                                 const uint32_t* __restrict__ vis_offsets,   // per 256-block, exclusive
                                                                 uint32_t* __restrict__ keys, uint32_t* __restrict__ ids) {
    __shared__ uint32_t s_w[kPreBlock / 32];
    const int i = blockIdx.x * kPreBlock + threadIdx.x;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const bool vis = (i < P) && tiles_touched[i] != 0;
    const uint32_t m = __ballot_sync(0xffffffffu, vis);
    if (lane == 0) s_w[wid] = __popc(m);
    __syncthreads();
    if (vis) {
        uint32_t pos = vis_offsets[blockIdx.x] + __popc(m & ((1u << lane) - 1u));
#pragma unroll
        for (int k = 0; k < kPreBlock / 32; ++k) pos += (k < wid) ? s_w[k] : 0u;
        keys[pos] = __float_as_uint(depths[i]);
        ids[pos] = (uint32_t)i;
    }
}

// per-256 block sums, in depth order, of the number of tile rows each Gaussian's rect spans (= runs it emits)
__global__ void __launch_bounds__(kPreBlock) order_block_sums_kernel(const uint32_t* __restrict__ n_dev, const uint32_t* __restrict__ order,
                                                                     const uint2* __restrict__ rects, uint32_t* __restrict__ sums) {
    __shared__ uint32_t s_w[kPreBlock / 32];
    const uint32_t j = blockIdx.x * kPreBlock + threadIdx.x;
    uint32_t v = 0;
    if (j < __ldg(n_dev)) { const uint32_t ry = rects[order[j]].y; v = (ry >> 16) - (ry & 0xffffu); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) { uint32_t t = 0; for (int w = 0; w < kPreBlock / 32; ++w) t += s_w[w]; sums[blockIdx.x] = t; }
}

// ------------------------------------------------------------------------------------------------
// Run-length binning.  The reference materialises every (tile, Gaussian) pair with a 64-bit key and radix-sorts all
// N of them (duplicateWithKeys + SortPairs, rasterizer_impl.cu:70-111,303-308).  A Gaussian's tiles form a rectangle,
// so its pairs are h runs of w consecutive tiles: we emit one 8-byte RUN per (Gaussian, tile row) in depth order
// (N / w-bar of them), stable-sort the runs by row (one radix pass), and expand each row's runs straight into the
// per-tile lists with an exact rank computation -- the N pairs are written once (4 B each) and never sorted.
//   rank of Gaussian j in tile (row, x) = number of runs of that row, earlier in depth order, that cover column x.
// ------------------------------------------------------------------------------------------------
struct EmitRunsParams {
    const GeomHeader* __restrict__ hdr;          // num_visible
    const uint32_t* __restrict__ order;          // depth-sorted Gaussian ids
    const uint2* __restrict__ srect;             // [num_visible] ((x0 << 10) | width, y0) of every sorted position
    const uint32_t* __restrict__ runoff;         // [num_visible + 1] first run of every sorted position (exclusive prefix of the tile rows)
    uint32_t* __restrict__ rkeys;                // out: (row << 20) | (x0 << 10) | w
    uint32_t* __restrict__ rvals;                // out: Gaussian id
    uint32_t* hist;                              // [passes][kSortMaxRadix] global digit counts of the row field
    int passes, shift0, bits0, shift1, bits1;    // at most two passes (<= 4095 tile rows)
};

// One THREAD PER RUN: the runs of a block of 256 depth-sorted Gaussians are consecutive in the output, so thread r of the
// block writes run base + r -- fully coalesced stores (one thread per Gaussian looping over its rows wrote 32 different
// sectors per store instruction, 1 cycle each in the L1) and no divergence on tall Gaussians.  The Gaussian of a run comes
// from a marker + prefix-max pass over tiles of 1024 runs in shared memory: every Gaussian marks the slot of its first
// run with its index, an inclusive max-scan spreads the index over the Gaussian's runs (~30 instructions per run; a binary
// search per run in the block's offsets cost ~150 and made the kernel issue-bound).
constexpr int kEmitTile = 4 * kPreBlock;

__global__ void __launch_bounds__(kPreBlock) emit_runs_kernel(const EmitRunsParams p) {
    extern __shared__ uint32_t s_hist[];            // [passes][kSortMaxRadix]
    __shared__ uint32_t s_off[kPreBlock + 1];
    __shared__ uint32_t s_id[kPreBlock];
    __shared__ uint32_t s_x[kPreBlock], s_y0[kPreBlock];
    __shared__ __align__(16) uint32_t s_g[kEmitTile];
    __shared__ uint32_t s_wmax[kPreBlock / 32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t V = p.hdr->num_visible;
    const uint32_t j0 = blockIdx.x * kPreBlock;
    if (j0 >= V) return;                             // (CTA-uniform)
    const uint32_t j = j0 + tid;
    const int cnt = (int)min((uint32_t)kPreBlock, V - j0);
    for (int i = tid; i < p.passes * kSortMaxRadix; i += kPreBlock) s_hist[i] = 0;
    if (tid < cnt) {
        const uint2 r = __ldg(p.srect + j);
        s_id[tid] = __ldg(p.order + j); s_x[tid] = r.x; s_y0[tid] = r.y;
        s_off[tid] = __ldg(p.runoff + j);
    }
    if (tid == 0) s_off[cnt] = __ldg(p.runoff + j0 + cnt);
    __syncthreads();
    const uint32_t base = s_off[0], total = s_off[cnt] - base;
    const uint32_t my_st = tid < cnt ? s_off[tid] - base : 0u, my_en = tid < cnt ? s_off[tid + 1] - base : 0u;
    for (uint32_t tile0 = 0; tile0 < total; tile0 += kEmitTile) {
        reinterpret_cast<uint4*>(s_g)[tid] = make_uint4(0u, 0u, 0u, 0u);
        if (tid == 0) s_carry = 0;
        __syncthreads();
        if (my_en > my_st) {                                          // Gaussians with at least one run
            if (my_st >= tile0 && my_st < tile0 + kEmitTile) s_g[my_st - tile0] = (uint32_t)tid + 1u;
            else if (my_st < tile0 && my_en > tile0) s_carry = (uint32_t)tid + 1u;      // the one whose runs straddle the tile start
        }
        __syncthreads();
        // inclusive prefix-max over the tile: 4 consecutive slots per thread, warp scan, 8 warp totals
        uint4 v = reinterpret_cast<const uint4*>(s_g)[tid];
        v.y = max(v.y, v.x); v.z = max(v.z, v.y); v.w = max(v.w, v.z);
        uint32_t m = v.w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, m, o); if (lane >= o) m = max(m, u); }
        if (lane == 31) s_wmax[wid] = m;
        uint32_t ex = __shfl_up_sync(0xffffffffu, m, 1);
        if (lane == 0) ex = 0;
        __syncthreads();
        uint32_t pre = s_carry;
#pragma unroll
        for (int w = 0; w < kPreBlock / 32; ++w) pre = max(pre, (w < wid) ? s_wmax[w] : 0u);
        pre = max(pre, ex);
        reinterpret_cast<uint4*>(s_g)[tid] = make_uint4(max(v.x, pre), max(v.y, pre), max(v.z, pre), max(v.w, pre));   // own slots only
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t r = tile0 + k * kPreBlock + tid;
            if (r < total) {
                const uint32_t lo = s_g[k * kPreBlock + tid] - 1u;
                const uint32_t at = base + r;
                const uint32_t key = ((s_y0[lo] + (at - s_off[lo])) << kRunRowShift) | s_x[lo];
                p.rkeys[at] = key;
                p.rvals[at] = s_id[lo];
                atomicAdd(&s_hist[(key >> p.shift0) & ((1u << p.bits0) - 1u)], 1u);
                if (p.passes > 1) atomicAdd(&s_hist[kSortMaxRadix + ((key >> p.shift1) & ((1u << p.bits1) - 1u))], 1u);
            }
        }
        __syncthreads();
    }
    for (int i = tid; i < p.passes * kSortMaxRadix; i += kPreBlock) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&p.hist[i], c);
    }
}

static int g_num_sms = 0;
static int num_sms() {
    if (g_num_sms == 0) {
        int dev = 0; cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (g_num_sms <= 0) g_num_sms = 148;
    }
    return g_num_sms;
}

void launch_compact_keys(const GeomPtrs& g, int P, uint32_t* keys, uint32_t* ids, cudaStream_t s) {
    compact_keys_kernel<<<(P + kPreBlock - 1) / kPreBlock, kPreBlock, 0, s>>>(P, g.depths, g.tiles_touched, g.block_offsets, keys, ids);
    count_launch();
}

void launch_order_block_sums(int P, const uint32_t* n_dev, const uint32_t* order, const uint2* rects, uint32_t* sums, cudaStream_t s) {
    order_block_sums_kernel<<<(P + kPreBlock - 1) / kPreBlock, kPreBlock, 0, s>>>(n_dev, order, rects, sums);
    count_launch();
}

void launch_emit_runs(const GeomPtrs& g, long long V, uint32_t* rkeys, uint32_t* rvals, uint32_t* hist, const SortPlan& plan, cudaStream_t s) {
    EmitRunsParams p;
    p.hdr = g.header; p.order = g.order; p.srect = g.srect; p.runoff = g.runoff;
    p.rkeys = rkeys; p.rvals = rvals; p.hist = hist;
    p.passes = plan.passes;
    p.shift0 = plan.shift[0]; p.bits0 = plan.bits[0];
    p.shift1 = plan.passes > 1 ? plan.shift[1] : 0; p.bits1 = plan.passes > 1 ? plan.bits[1] : 1;
    if (V < 1) return;                 // the host knows the visible count by now: the grid covers exactly the sorted positions
    emit_runs_kernel<<<(unsigned)((V + kPreBlock - 1) / kPreBlock), kPreBlock, sizeof(uint32_t) * plan.passes * kSortMaxRadix, s>>>(p);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// standalone digit histogram (keys that do not come out of emit_pairs_kernel: the depth sort, gsr_sort_pairs)
// ------------------------------------------------------------------------------------------------
template <typename KeyT>
struct HistParams { const KeyT* keys; long long n; const uint32_t* n_dev; uint32_t* hist; int passes; int shift[kSortMaxPasses]; int bits[kSortMaxPasses]; };

template <typename KeyT>
__global__ void __launch_bounds__(256) sort_hist_kernel(const HistParams<KeyT> p) {
    extern __shared__ uint32_t s_hist[];
    const int tid = threadIdx.x, lane = tid & 31;
    for (int i = tid; i < p.passes * kSortMaxRadix; i += 256) s_hist[i] = 0;
    __syncthreads();
    const long long n = p.n_dev ? (long long)__ldg(p.n_dev) : p.n;
    const long long stride = (long long)gridDim.x * 256;
    const long long n_round = (n + 31) / 32 * 32;
    for (long long i = (long long)blockIdx.x * 256 + tid; i < n_round; i += stride) {
        const bool valid = i < n;
        const KeyT key = valid ? p.keys[i] : (KeyT)0;
        const uint32_t act = __ballot_sync(0xffffffffu, valid);
        if (valid) {
            // low digits are close to uniform: plain shared atomics (few conflicts).  The most significant digit has a
            // handful of distinct values (depth exponents, tile rows): aggregate equal digits in the warp first.
            for (int ps = 0; ps + 1 < p.passes; ++ps)
                atomicAdd(&s_hist[ps * kSortMaxRadix + ((uint32_t)(key >> p.shift[ps]) & ((1u << p.bits[ps]) - 1u))], 1u);
            const int ps = p.passes - 1;
            const uint32_t d = (uint32_t)(key >> p.shift[ps]) & ((1u << p.bits[ps]) - 1u);
            const uint32_t peers = __match_any_sync(act, d);
            if (lane == (31 - __clz(peers))) atomicAdd(&s_hist[ps * kSortMaxRadix + d], (uint32_t)__popc(peers));
        }
    }
    __syncthreads();
    for (int i = tid; i < p.passes * kSortMaxRadix; i += 256) { const uint32_t h = s_hist[i]; if (h) atomicAdd(&p.hist[i], h); }
}

// counts -> exclusive offsets, one CTA per pass
__global__ void __launch_bounds__(kSortMaxRadix) hist_scan_kernel(uint32_t* hist) {
    __shared__ uint32_t s_w[kSortMaxRadix / 32];
    uint32_t* h = hist + blockIdx.x * kSortMaxRadix;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t v = h[tid];
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (lane == 31) s_w[wid] = inc;
    __syncthreads();
    uint32_t base = 0;
#pragma unroll
    for (int w = 0; w < kSortMaxRadix / 32; ++w) base += (w < wid) ? s_w[w] : 0u;
    h[tid] = base + inc - v;
}

// ------------------------------------------------------------------------------------------------
// onesweep pass: one read and one write of every (key, value) pair, chained-scan prefix across tiles.
//   KeyT = uint32_t : depth keys of Gaussians, tile ids of (tile, Gaussian) pairs   (the rasterizer's own sorts)
//   KeyT = uint64_t : gsr_sort_pairs (the cub::DeviceRadixSort::SortPairs<u64,u32> replacement as a stand-alone op)
// Tile = 512 threads x ITEMS keys, warp-striped loads (coalesced, order-preserving), match-any multi-split ranks,
// keys AND values staged in shared memory so that each digit bin leaves as one contiguous run.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t ST_AGG = 1u << 30, ST_INC = 2u << 30, ST_MASK = (1u << 30) - 1u;

template <typename KeyT>
struct SweepParams {
    const KeyT* __restrict__ keys_in; KeyT* __restrict__ keys_out;
    const uint32_t* __restrict__ vals_in;    // nullptr: value = global index (first pass of an index sort)
    uint32_t* __restrict__ vals_out;
    long long n; const uint32_t* __restrict__ n_dev;   // n_dev != nullptr: the actual count lives on the device (<= n)
    int shift; int bits;
    const uint32_t* __restrict__ bin_base;   // [RADIX] exclusive global offsets of this pass
    uint32_t* status;                        // [tiles][RADIX]
    uint32_t* tile_counter;
    uint32_t* err_flag;
};

// lanes whose digit equals mine, built from one ballot per digit bit: constant cost whatever the number of distinct
// digits in the warp (the hardware MATCH.ANY degrades with digit diversity, which is the common case here)
template <int BITS>
__device__ __forceinline__ uint32_t match_digit(uint32_t d) {
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < BITS; ++b) {
        const uint32_t bal = __ballot_sync(0xffffffffu, (d >> b) & 1u);
        peers &= ((d >> b) & 1u) ? bal : ~bal;
    }
    return peers;
}

template <int R> struct Log2 { static constexpr int v = 1 + Log2<R / 2>::v; };
template <> struct Log2<1> { static constexpr int v = 0; };

// THREADS x ITEMS keys per tile; THREADS >= RADIX.  MINB = CTAs per SM the register budget is planned for.
template <typename KeyT, int RADIX, int THREADS, int ITEMS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) onesweep_kernel(const SweepParams<KeyT> p) {
    static_assert(THREADS >= RADIX, "one thread per bin in the scan / look-back phases");
    constexpr int WARPS = THREADS / 32;
    constexpr int TILE = THREADS * ITEMS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    KeyT* s_keys = reinterpret_cast<KeyT*>(smem_raw);                                         // [TILE]
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(smem_raw + sizeof(KeyT) * TILE);           // [TILE]
    uint32_t* s_whist = s_vals + TILE;                                                        // [WARPS][RADIX]
    uint32_t* s_bin_start = s_whist + WARPS * RADIX;                                          // [RADIX]
    uint32_t* s_gofs = s_bin_start + RADIX;                                                   // [RADIX]
    __shared__ uint32_t s_wtot[RADIX / 32];

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    // grid sized for the upper bound n: tiles past the device-side count leave at once (nobody waits on them)
    if (p.n_dev != nullptr && (long long)blockIdx.x * (THREADS * ITEMS) >= (long long)__ldg(p.n_dev)) return;
    for (int i = tid; i < WARPS * RADIX; i += THREADS) s_whist[i] = 0;
    __syncthreads();
    // The tile's place in the chained scan is its block index: blocks are dispatched in increasing linear index (the
    // assumption cub::DeviceScan makes too), so every predecessor is resident or done.  A global ticket counter costs a
    // serialised same-address atomic round trip per CTA; the watchdog below still bounds any wait.
    const uint32_t tile = blockIdx.x;
    const long long tile_base = (long long)tile * TILE;
    const long long n_items = p.n_dev ? (long long)__ldg(p.n_dev) : p.n;
    if (tile_base >= n_items) return;          // grid was sized for the upper bound (CTA-uniform exit; nobody waits on this tile)
    const int count = (int)min((long long)TILE, n_items - tile_base);
    const uint32_t mask = (1u << p.bits) - 1u;

    // ---- load keys and values, warp-striped: item i of lane l sits at warp_base + i*32 + l
    KeyT key[ITEMS];
    uint32_t val[ITEMS];
    const int wbase = wid * 32 * ITEMS;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const int lp = wbase + i * 32 + lane;
        key[i] = (lp < count) ? p.keys_in[tile_base + lp] : (KeyT)~(KeyT)0;
    }
    if (p.vals_in != nullptr) {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const int lp = wbase + i * 32 + lane;
            val[i] = (lp < count) ? p.vals_in[tile_base + lp] : 0u;
        }
    } else {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) val[i] = (uint32_t)(tile_base + wbase + i * 32 + lane);
    }

    // ---- stable ranks inside the warp (match-any multi-split)
    uint32_t rank[ITEMS];
    uint32_t* wh = s_whist + wid * RADIX;
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const uint32_t d = (uint32_t)(key[i] >> p.shift) & mask;
        const uint32_t peers = match_digit<Log2<RADIX>::v>(d);
        const uint32_t pre = wh[d];
        __syncwarp();
        if ((peers & lt) == 0) wh[d] = pre + __popc(peers);
        __syncwarp();
        rank[i] = pre + __popc(peers & lt);
    }
    __syncthreads();

    // ---- per-bin: exclusive scan across warps, publish the tile aggregate as early as possible
    uint32_t bin_count = 0;
    if (tid < RADIX) {
        uint32_t acc = 0;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) { const uint32_t t = s_whist[w * RADIX + tid]; s_whist[w * RADIX + tid] = acc; acc += t; }
        bin_count = acc;
        *((volatile uint32_t*)(p.status + (size_t)tile * RADIX + tid)) = (tile == 0 ? ST_INC : ST_AGG) | acc;
    }
    {   // tile-local exclusive scan of the bin counts (warps 0 .. RADIX/32-1)
        uint32_t inc = bin_count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
        if (tid < RADIX && lane == 31) s_wtot[wid] = inc;
        __syncthreads();
        if (tid < RADIX) {
            uint32_t b = 0;
#pragma unroll
            for (int w = 0; w < RADIX / 32; ++w) b += (w < wid) ? s_wtot[w] : 0u;
            s_bin_start[tid] = b + inc - bin_count;
        }
    }
    __syncthreads();

    // ---- stage keys and values in shared memory in bin order (overlaps the look-back below)
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const uint32_t d = (uint32_t)(key[i] >> p.shift) & mask;
        const uint32_t lp = s_bin_start[d] + wh[d] + rank[i];
        s_keys[lp] = key[i];
        s_vals[lp] = val[i];
    }

    // ---- decoupled look-back, one thread per bin
    if (tid < RADIX) {
        uint32_t excl = 0;
        if (tile > 0) {
            long long t = (long long)tile - 1;
            uint32_t spins = 0;
            while (true) {
                const uint32_t v = *((volatile uint32_t*)(p.status + (size_t)t * RADIX + tid));
                const uint32_t f = v & ~ST_MASK;
                if (f == 0) {
                    if (++spins > 64u) __nanosleep(64);
                    if (spins > (1u << 22)) { atomicOr(p.err_flag, 1u); break; }   // watchdog: never hang the GPU
                    continue;
                }
                excl += v & ST_MASK;
                if (f == ST_INC) break;
                --t;
            }
            *((volatile uint32_t*)(p.status + (size_t)tile * RADIX + tid)) = ST_INC | ((excl + bin_count) & ST_MASK);
        }
        s_gofs[tid] = p.bin_base[tid] + excl - s_bin_start[tid];   // global position of smem slot j of this bin = s_gofs + j
    }
    __syncthreads();

    // ---- scatter: consecutive smem slots of one bin are consecutive global positions
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const int j = tid + i * THREADS;
        if (j < count) {
            const KeyT k = s_keys[j];
            const uint32_t d = (uint32_t)(k >> p.shift) & mask;
            const uint32_t g = s_gofs[d] + (uint32_t)j;
            p.keys_out[g] = k;
            p.vals_out[g] = s_vals[j];
        }
    }
}

template <typename KeyT, int RADIX, int THREADS, int ITEMS>
static size_t sweep_smem() {
    return (sizeof(KeyT) + sizeof(uint32_t)) * THREADS * ITEMS + sizeof(uint32_t) * ((THREADS / 32) * RADIX + 2 * RADIX);
}

struct SortTemp { uint32_t* hist; uint32_t* counters; uint32_t* status; };
static SortTemp carve_sort_temp(void* temp) {
    char* b = aligned_base(temp);
    SortTemp t;
    t.hist = (uint32_t*)b; b += align_up(sizeof(uint32_t) * kSortMaxPasses * kSortMaxRadix);
    t.counters = (uint32_t*)b; b += align_up(256);
    t.status = (uint32_t*)b;
    return t;
}
uint32_t* sort_hist_ptr(void* temp) { return (uint32_t*)aligned_base(temp); }

// tile shapes (threads x items per thread, CTAs per SM).  Variant 0 is the default; the others exist for tuning
// sweeps through gsr_set_option(GSR_OPT_SORT_VARIANT, v) and are all bit-identical in result.
static int g_sort_variant = 0;
void set_sort_variant(int v) { g_sort_variant = v; }

template <typename KeyT, int RADIX, int THREADS, int ITEMS, int MINB>
static void launch_sweep_cfg(const SweepParams<KeyT>& sp, long long n, cudaStream_t s) {
    constexpr int TILE = THREADS * ITEMS;
    static_assert(TILE >= kSortTileItems, "temp sizing assumes no tile is smaller than kSortTileItems");
    static bool attr_done = false;
    const size_t smem = sweep_smem<KeyT, RADIX, THREADS, ITEMS>();
    if (!attr_done) {
        cudaFuncSetAttribute(onesweep_kernel<KeyT, RADIX, THREADS, ITEMS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr_done = true;
    }
    const long long tiles = (n + TILE - 1) / TILE;
    onesweep_kernel<KeyT, RADIX, THREADS, ITEMS, MINB><<<(unsigned)tiles, THREADS, smem, s>>>(sp);
}

template <typename KeyT, int RADIX> struct SweepDispatch;
template <int RADIX> struct SweepDispatch<uint64_t, RADIX> {
    static int tile_items() { return 512 * 9; }
    static void launch(const SweepParams<uint64_t>& sp, long long n, cudaStream_t s) { launch_sweep_cfg<uint64_t, RADIX, 512, 9, 2>(sp, n, s); }
};
template <int RADIX> struct SweepDispatch<uint32_t, RADIX> {
    static int tile_items() {
        switch (g_sort_variant) { case 1: return 256 * 12; case 2: return 256 * 16; case 3: return 384 * 12; case 4: return 256 * 8;
                                  case 5: return 512 * 8; default: return 512 * 12; }
    }
    static void launch(const SweepParams<uint32_t>& sp, long long n, cudaStream_t s) {
        switch (g_sort_variant) {
            case 1: launch_sweep_cfg<uint32_t, RADIX, 256, 12, 4>(sp, n, s); break;
            case 2: launch_sweep_cfg<uint32_t, RADIX, 256, 16, 3>(sp, n, s); break;
            case 3: launch_sweep_cfg<uint32_t, RADIX, 384, 12, 2>(sp, n, s); break;
            case 4: launch_sweep_cfg<uint32_t, RADIX, 256, 8, 5>(sp, n, s); break;
            case 5: launch_sweep_cfg<uint32_t, RADIX, 512, 8, 2>(sp, n, s); break;
            default: launch_sweep_cfg<uint32_t, RADIX, 512, 12, 2>(sp, n, s); break;
        }
    }
};

// look-back words one pass needs: tiles x radix of the widest pass of the plan, for the active tile shape
template <typename KeyT>
static size_t pass_status_words(long long n, const SortPlan& plan) {
    size_t words = 0;
    for (int ps = 0; ps < plan.passes; ++ps) {
        const int radix = plan.bits[ps] <= 7 ? 128 : (plan.bits[ps] <= 8 ? 256 : 512);
        const int tile_items = radix == 128 ? SweepDispatch<KeyT, 128>::tile_items()
                             : radix == 256 ? SweepDispatch<KeyT, 256>::tile_items() : 512 * 9;
        const size_t w = (size_t)((n + tile_items - 1) / tile_items) * radix;
        if (w > words) words = w;
    }
    return (words + 3) / 4 * 4;
}

size_t sort_prezero_bytes(long long n, const SortPlan& plan, size_t temp_bytes, bool keys64) {
    if (n <= 0) return 0;
    const size_t head = align_up(sizeof(uint32_t) * kSortMaxPasses * kSortMaxRadix) + align_up(256);
    const size_t words = keys64 ? pass_status_words<uint64_t>(n, plan) : pass_status_words<uint32_t>(n, plan);
    const size_t need = head + sizeof(uint32_t) * words * plan.passes;
    return need + kAlign <= temp_bytes ? need : 0;      // kAlign: slack of aligned_base()
}

__global__ void __launch_bounds__(256) zero_kernel(uint4* __restrict__ p, size_t n16) {
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n16; i += (size_t)gridDim.x * 256) p[i] = make_uint4(0u, 0u, 0u, 0u);
}
void launch_zero(void* ptr, size_t bytes, cudaStream_t s) {
    if (bytes == 0) return;
    const size_t n16 = bytes / 16;
    const int grid = (int)min((size_t)num_sms() * 8, (n16 + 255) / 256);
    zero_kernel<<<grid, 256, 0, s>>>(reinterpret_cast<uint4*>(aligned_base(ptr)), n16);
    count_launch();
}

template <typename KeyT>
int launch_onesweep_t(const KeyT* kin0, const uint32_t* vin0, KeyT* kout, uint32_t* vout, KeyT* ktmp, uint32_t* vtmp,
                      long long n, const SortPlan& plan, void* temp, size_t temp_bytes, bool hist_ready, uint32_t* err_flag,
                      int prof_slot0, cudaStream_t s, const uint32_t* n_dev = nullptr, bool prezeroed = false) {
    if (n <= 0) return GSR_OK;
    if (n >= (1ll << 30)) { set_error("radix sort: n=%lld exceeds 2^30 pairs", n); return GSR_ERR_INVALID; }
    if (temp_bytes < sort_temp_bytes(n)) { set_error("radix sort: temp too small"); return GSR_ERR_ALLOC; }
    SortTemp t = carve_sort_temp(temp);
    const size_t pass_words = pass_status_words<KeyT>(n, plan);
    if (prezeroed && sort_prezero_bytes(n, plan, temp_bytes, sizeof(KeyT) == 8) == 0) { set_error("radix sort: prezeroed temp does not fit"); return GSR_ERR_ALLOC; }
    if (!hist_ready) {
        if (!prezeroed) cudaMemsetAsync(t.hist, 0, sizeof(uint32_t) * kSortMaxPasses * kSortMaxRadix, s);
        HistParams<KeyT> hp; hp.keys = kin0; hp.n = n; hp.n_dev = n_dev; hp.hist = t.hist; hp.passes = plan.passes;
        for (int i = 0; i < kSortMaxPasses; ++i) { hp.shift[i] = i < plan.passes ? plan.shift[i] : 0; hp.bits[i] = i < plan.passes ? plan.bits[i] : 1; }
        long long want = (n + 256 * 16 - 1) / (256 * 16);
        int grid = (int)min((long long)num_sms() * 4, want); if (grid < 1) grid = 1;
        sort_hist_kernel<KeyT><<<grid, 256, sizeof(uint32_t) * plan.passes * kSortMaxRadix, s>>>(hp);
        count_launch();
    }
    hist_scan_kernel<<<plan.passes, kSortMaxRadix, 0, s>>>(t.hist);
    count_launch();
    if (!prezeroed) cudaMemsetAsync(t.counters, 0, 256, s);
    const KeyT* kin = kin0; const uint32_t* vin = vin0;
    KeyT* kb[2]; uint32_t* vb[2];
    if (plan.passes & 1) { kb[0] = kout; vb[0] = vout; kb[1] = ktmp; vb[1] = vtmp; }
    else                 { kb[0] = ktmp; vb[0] = vtmp; kb[1] = kout; vb[1] = vout; }
    for (int ps = 0; ps < plan.passes; ++ps) {
        SweepParams<KeyT> sp;
        sp.keys_in = kin; sp.keys_out = kb[ps & 1]; sp.vals_in = vin; sp.vals_out = vb[ps & 1];
        sp.n = n; sp.n_dev = n_dev; sp.shift = plan.shift[ps]; sp.bits = plan.bits[ps];
        sp.bin_base = t.hist + ps * kSortMaxRadix; sp.tile_counter = t.counters + ps; sp.err_flag = err_flag;
        const int radix = plan.bits[ps] <= 7 ? 128 : (plan.bits[ps] <= 8 ? 256 : 512);
        // prezeroed: every pass has its own look-back region, cleared by the caller in one go; otherwise one region,
        // re-zeroed before each pass
        sp.status = prezeroed ? t.status + (size_t)ps * pass_words : t.status;
        if (!prezeroed) cudaMemsetAsync(t.status, 0, sizeof(uint32_t) * pass_words, s);
        {
            ProfScope pr(prof_slot0 >= 0 ? prof_slot0 + ps : -1, s);
            if (radix == 128) SweepDispatch<KeyT, 128>::launch(sp, n, s);
            else if (radix == 256) SweepDispatch<KeyT, 256>::launch(sp, n, s);
            else launch_sweep_cfg<KeyT, 512, 512, 9, 2>(sp, n, s);
        }
        count_launch();
        kin = kb[ps & 1]; vin = vb[ps & 1];
    }
    return GSR_OK;
}

int launch_onesweep(const uint64_t* kin, const uint32_t* vin, uint64_t* kout, uint32_t* vout, uint64_t* ktmp, uint32_t* vtmp,
                    long long n, const SortPlan& plan, void* temp, size_t temp_bytes, bool hist_ready, uint32_t* err_flag, cudaStream_t s) {
    return launch_onesweep_t<uint64_t>(kin, vin, kout, vout, ktmp, vtmp, n, plan, temp, temp_bytes, hist_ready, err_flag, -1, s);
}
int launch_onesweep32(const uint32_t* kin, const uint32_t* vin, uint32_t* kout, uint32_t* vout, uint32_t* ktmp, uint32_t* vtmp,
                      long long n, const SortPlan& plan, void* temp, size_t temp_bytes, bool hist_ready, uint32_t* err_flag,
                      int prof_slot0, cudaStream_t s, const uint32_t* n_dev, bool prezeroed) {
    return launch_onesweep_t<uint32_t>(kin, vin, kout, vout, ktmp, vtmp, n, plan, temp, temp_bytes, hist_ready, err_flag, prof_slot0, s, n_dev, prezeroed);
}

// ------------------------------------------------------------------------------------------------
// expansion of the row-sorted runs into per-tile lists (+ tile ranges, + AMR intersection counts)
//   replaces cudaMemset + identifyTileRanges (rasterizer_impl.cu:116-138,310-317) and the AMR fork's
//   calculateIntersections (A/rasterizer_impl.cu:181-188)
// ------------------------------------------------------------------------------------------------
// exclusive scan over a 1024-thread CTA; returns the exclusive prefix, *total = CTA sum.  s_warp: 32 words.
__device__ __forceinline__ uint32_t cta1024_exclusive_scan(uint32_t v, uint32_t* s_warp, uint32_t* total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    __syncthreads();                         // s_warp may still be read from a previous call
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    uint32_t wsum = s_warp[lane], winc = wsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += t; }
    const uint32_t tot = __shfl_sync(0xffffffffu, winc, 31);
    const uint32_t wexcl = __shfl_sync(0xffffffffu, winc - wsum, wid);
    *total = tot;
    return wexcl + inc - v;
}

// row_start[r] = first sorted run of tile row r (binary search on the row field); chunk_row_start[r] = first chunk of row r
__global__ void __launch_bounds__(1024) chunk_table_kernel(const uint32_t* __restrict__ rkeys, uint32_t R, int tiles_y, int chunk,
                                                           const uint32_t* __restrict__ row_offsets,   // sort's digit offsets or nullptr
                                                           uint32_t* __restrict__ row_start, uint32_t* __restrict__ chunk_row_start) {
    __shared__ uint32_t s_warp[32];
    const int tid = threadIdx.x;
    for (int r = tid; r <= tiles_y; r += 1024) {
        uint32_t lo = 0, hi = R;
        if (row_offsets != nullptr) lo = (r < tiles_y) ? row_offsets[r] : R;     // one-pass row sort: the digit IS the row
        else while (lo < hi) { const uint32_t mid = lo + ((hi - lo) >> 1); if ((__ldg(rkeys + mid) >> kRunRowShift) < (uint32_t)r) lo = mid + 1; else hi = mid; }
        row_start[r] = lo;
    }
    __syncthreads();
    uint32_t carry = 0;
    for (int base = 0; base <= tiles_y; base += 1024) {
        const int r = base + tid;
        uint32_t nch = 0;
        if (r < tiles_y) { const uint32_t len = row_start[r + 1] - row_start[r]; nch = (len + chunk - 1) / chunk; }
        uint32_t tot;
        const uint32_t ex = cta1024_exclusive_scan(nch, s_warp, &tot);
        if (r <= tiles_y) chunk_row_start[r] = carry + ex;      // entry [tiles_y] = total number of chunks
        carry += tot;
    }
}

struct ChunkCtx { int row; uint32_t r0, r1; };
// which rows' runs does chunk c cover?  (binary search over the <= 4096 rows; CTA-uniform)
__device__ __forceinline__ bool chunk_lookup(uint32_t c, int tiles_y, int chunk, const uint32_t* __restrict__ chunk_row_start,
                                             const uint32_t* __restrict__ row_start, ChunkCtx& cx) {
    if (c >= __ldg(chunk_row_start + tiles_y)) return false;
    int lo = 0, hi = tiles_y;                  // invariant: chunk_row_start[lo] <= c < chunk_row_start[hi]
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (__ldg(chunk_row_start + mid) <= c) lo = mid; else hi = mid; }
    cx.row = lo;
    const uint32_t k = c - __ldg(chunk_row_start + lo);
    cx.r0 = __ldg(row_start + lo) + k * (uint32_t)chunk;
    cx.r1 = min(cx.r0 + (uint32_t)chunk, __ldg(row_start + lo + 1));
    return true;
}

// per chunk: how many of its runs cover each tile column (difference array + scan)
__global__ void __launch_bounds__(256) chunk_count_kernel(const uint32_t* __restrict__ rkeys, int tiles_x, int tiles_y, int chunk,
                                                          const uint32_t* __restrict__ chunk_row_start, const uint32_t* __restrict__ row_start,
                                                          uint32_t* __restrict__ chunk_counts, uint4* __restrict__ chunk_info) {
    extern __shared__ int s_diff[];               // [tiles_x + 1], then reused for the scanned counts
    __shared__ int s_wsum[8];
    ChunkCtx cx;
    const bool live = chunk_lookup(blockIdx.x, tiles_y, chunk, chunk_row_start, row_start, cx);
    // the chunk's place, for expand_runs_kernel (one 16-byte load there instead of a second binary search)
    if (threadIdx.x == 0) chunk_info[blockIdx.x] = live ? make_uint4((uint32_t)cx.row, cx.r0, cx.r1, 1u) : make_uint4(0u, 0u, 0u, 0u);
    if (!live) return;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int i = tid; i <= tiles_x; i += 256) s_diff[i] = 0;
    __syncthreads();
    for (uint32_t r = cx.r0 + tid; r < cx.r1; r += 256) {
        const uint32_t key = __ldg(rkeys + r);
        const uint32_t x0 = (key >> kRunX0Shift) & kRunFieldMask, w = key & kRunFieldMask;
        atomicAdd(&s_diff[x0], 1);
        atomicAdd(&s_diff[x0 + w], -1);
    }
    __syncthreads();
    // inclusive scan over the columns: 4 consecutive columns per thread (tiles_x <= 1023)
    int v[4], sum = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) { const int x = 4 * tid + i; v[i] = (x < tiles_x) ? s_diff[x] : 0; sum += v[i]; }
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) s_wsum[wid] = inc;
    __syncthreads();
    int run = inc - sum;
#pragma unroll
    for (int k = 0; k < 8; ++k) run += (k < wid) ? s_wsum[k] : 0;
    uint32_t* out = chunk_counts + (size_t)blockIdx.x * tiles_x;
#pragma unroll
    for (int i = 0; i < 4; ++i) { const int x = 4 * tid + i; run += v[i]; if (x < tiles_x) out[x] = (uint32_t)run; }
}

// per tile row: exclusive prefix over the row's chunks for every column (in place) + per-tile totals.
// CTA = 32 chunk-lanes x 32 columns; a lane owns a contiguous slice of the row's chunks (two passes: slice sums, then
// the running prefix), so the serial chain is chunks/32 long and every load is a coalesced 128 B row segment.
__global__ void __launch_bounds__(1024) row_scan_kernel(int tiles_x, const uint32_t* __restrict__ chunk_row_start,
                                                        uint32_t* __restrict__ chunk_counts, uint32_t* __restrict__ tile_count) {
    __shared__ uint32_t s_part[32][33];
    const int row = blockIdx.y;
    const int xl = threadIdx.x & 31, l = threadIdx.x >> 5;
    const int x = blockIdx.x * 32 + xl;
    const uint32_t c0 = chunk_row_start[row], c1 = chunk_row_start[row + 1];
    const uint32_t per = (c1 - c0 + 31) / 32;
    const uint32_t a = min(c1, c0 + l * per), b = min(c1, a + per);
    uint32_t sum = 0;
    if (x < tiles_x)
        for (uint32_t c = a; c < b; ++c) sum += chunk_counts[(size_t)c * tiles_x + x];
    s_part[l][xl] = sum;
    __syncthreads();
    uint32_t base = 0, total = 0;
#pragma unroll
    for (int k = 0; k < 32; ++k) { const uint32_t v = s_part[k][xl]; base += (k < l) ? v : 0u; total += v; }
    if (x < tiles_x) {
        for (uint32_t c = a; c < b; ++c) {
            uint32_t* q = chunk_counts + (size_t)c * tiles_x + x;
            const uint32_t v = *q; *q = base; base += v;
        }
        if (l == 0) tile_count[(size_t)row * tiles_x + x] = total;
    }
}

// exclusive scan of the per-tile totals -> tile_start (in place), ranges, AMR intersection counts
__global__ void __launch_bounds__(1024) tile_scan_kernel(uint32_t* __restrict__ tile_start, int num_tiles, uint2* __restrict__ ranges,
                                                         uint32_t* __restrict__ n_intersections, const GeomHeader* hdr, uint32_t* mailbox,
                                                         uint32_t fake_error) {
    __shared__ uint32_t s_warp[32];
    const int tid = threadIdx.x;
    // Every radix pass of the frame (depth-sort fallback, row partition of the runs) is behind us in stream order: post the
    // watchdog word where the host finds it at its next call on this thread (gsr_api.cu: check_sort_mailbox).
    if (tid == 0 && mailbox != nullptr) *((volatile uint32_t*)mailbox + 8) = 0x80000000u | ((hdr->sort_error | fake_error) & 1u);
    const int per = (num_tiles + 1023) / 1024;
    const int begin = min(num_tiles, tid * per), end = min(num_tiles, begin + per);
    uint32_t local = 0;
    for (int t = begin; t < end; ++t) local += tile_start[t];
    uint32_t tot;
    uint32_t st = cta1024_exclusive_scan(local, s_warp, &tot);
    for (int t = begin; t < end; ++t) {
        const uint32_t cnt = tile_start[t];
        tile_start[t] = st;
        ranges[t] = cnt ? make_uint2(st, st + cnt) : make_uint2(0u, 0u);   // empty tiles keep the reference's memset value
        if (n_intersections) n_intersections[t] = cnt;
        st += cnt;
    }
}

// per chunk: write every run's Gaussian id into the lists of the tiles it covers, at its exact depth-order rank.
// A chunk's 1024 runs form 32 warp-sized groups in run order; the coverage of column x is 32 words (word q = group q, bit =
// lane = run q*32 + lane), stored COLUMN-MAJOR.  Phase 1: every run sets its bit in the words of the columns it covers
// (shared atomicOr).  Phase 2: a WARP PER COLUMN decodes that column's 1024-bit map: lane q takes word q, a warp scan of
// the popcounts gives its first slot, and the set bits are enumerated into consecutive list slots -- the lanes of a warp
// write one contiguous stretch of one tile's list (a few sectors per store), where the round-1 kernel had every lane
// scatter into a different tile (32 sectors per store instruction, 1 cycle each in the L1: that, not DRAM, was its limit).
// No per-(group, column) offset table, no rank look-ups.
constexpr int kExpandThreads = 256;
// CHUNK = 1024 or 2048 runs per CTA.  The per-column decode has a fixed cost per column and chunk (scan, look-ups), so the larger
// chunk halves it per run; it needs [CHUNK] ids + tiles_x x (CHUNK/32 + pad) coverage words of shared memory (40 KB at 1080p for
// 2048) and is used up to 600 tile columns.
template <int CHUNK>
__global__ void __launch_bounds__(kExpandThreads) expand_runs_kernel(const uint32_t* __restrict__ rkeys, const uint32_t* __restrict__ rvals,
                                                                     int tiles_x, const uint4* __restrict__ chunk_info,
                                                                     const uint32_t* __restrict__ chunk_counts,
                                                                     const uint32_t* __restrict__ tile_start, uint32_t* __restrict__ point_list) {
    constexpr int BATCHES = CHUNK / kExpandThreads;                     // runs per thread
    constexpr int PLANES = CHUNK / 32;                                  // coverage words per column
    constexpr int WPL = PLANES / 32;                                    // words per lane in the decode
    constexpr int STRIDE = PLANES + (WPL == 1 ? 1 : 2);                 // 33: odd; 66: lane's two words 8-byte aligned, columns 2 banks apart
    extern __shared__ __align__(16) uint32_t s_mem[];
    uint32_t* s_ids = s_mem;                                            // [CHUNK] Gaussian id of run r
    uint32_t* s_bits = s_ids + CHUNK;                                   // [tiles_x][STRIDE] coverage words
    const uint4 info = __ldg(chunk_info + blockIdx.x);                  // (row, first run, end run, valid)
    if (info.w == 0u) return;
    const uint32_t row = info.x, r0 = info.y, r1 = info.z;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    uint32_t keys[BATCHES];
#pragma unroll
    for (int b = 0; b < BATCHES; ++b) {
        const uint32_t r = r0 + b * kExpandThreads + tid;
        keys[b] = (r < r1) ? __ldg(rkeys + r) : 0u;                     // width field 0: covers nothing
        s_ids[b * kExpandThreads + tid] = (r < r1) ? __ldg(rvals + r) : 0u;
    }
    for (int i = tid; i < tiles_x * STRIDE; i += kExpandThreads) s_bits[i] = 0;
    __syncthreads();
#pragma unroll
    for (int b = 0; b < BATCHES; ++b) {
        const int q = b * (kExpandThreads / 32) + wid;                  // group (coverage word) of run b*256 + tid
        const uint32_t x0 = (keys[b] >> kRunX0Shift) & kRunFieldMask, w = keys[b] & kRunFieldMask;
        for (uint32_t i = 0; i < w; ++i) atomicOr(&s_bits[(x0 + i) * STRIDE + q], 1u << lane);
    }
    __syncthreads();
    const uint32_t* cc = chunk_counts + (size_t)blockIdx.x * tiles_x;
    const uint32_t* ts = tile_start + (size_t)row * tiles_x;
    for (int x = wid; x < tiles_x; x += kExpandThreads / 32) {
        uint32_t word[WPL];
        uint32_t c = 0;
#pragma unroll
        for (int k = 0; k < WPL; ++k) { word[k] = s_bits[x * STRIDE + lane * WPL + k]; c += __popc(word[k]); }
        uint32_t inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
        if (__shfl_sync(0xffffffffu, inc, 31) == 0u) continue;          // nothing of this chunk in column x (warp-uniform)
        uint32_t slot = __ldg(ts + x) + __ldg(cc + x) + inc - c;        // first list slot of this lane's runs
#pragma unroll
        for (int k = 0; k < WPL; ++k) {
            const uint32_t* ids = s_ids + (lane * WPL + k) * 32;
            uint32_t wd = word[k];
            while (wd) {
                const int bpos = __ffs(wd) - 1;
                wd &= wd - 1u;
                point_list[slot++] = ids[bpos];
            }
        }
    }
}

static uint32_t g_fake_sort_error = 0;
void set_fake_sort_error(int v) { g_fake_sort_error = v ? 1u : 0u; }

void launch_expand_runs(const BinPtrs& b, const uint32_t* rkeys_sorted, const uint32_t* rvals_sorted, long long R, int tiles_x, int tiles_y,
                        const uint32_t* row_offsets, uint2* ranges, uint32_t* n_intersections, const GeomHeader* hdr, uint32_t* mailbox_dev,
                        int prof_ranges, int prof_expand, cudaStream_t s) {
    const int num_tiles = tiles_x * tiles_y;
    // runs per expansion chunk.  2048 (half the per-column decode overhead per run, 40 KB of shared memory) measured SLOWER than
    // 1024 at 1080p (110-120 us against 94-104 us: 5 instead of 8 CTAs per SM); kept selectable for wide images only by hand.
    const int chunk = kRunChunk;
    const unsigned grid = (unsigned)(R / chunk + tiles_y + 1);
    {
        ProfScope ps(prof_ranges, s);
        chunk_table_kernel<<<1, 1024, 0, s>>>(rkeys_sorted, (uint32_t)R, tiles_y, chunk, row_offsets, b.row_start, b.chunk_row_start);
        chunk_count_kernel<<<grid, 256, sizeof(int) * (tiles_x + 1), s>>>(rkeys_sorted, tiles_x, tiles_y, chunk, b.chunk_row_start, b.row_start, b.chunk_counts, b.chunk_info);
        row_scan_kernel<<<dim3((tiles_x + 31) / 32, tiles_y), 1024, 0, s>>>(tiles_x, b.chunk_row_start, b.chunk_counts, b.tile_start);
        tile_scan_kernel<<<1, 1024, 0, s>>>(b.tile_start, num_tiles, ranges, n_intersections, hdr, mailbox_dev, g_fake_sort_error);
        count_launch(4);
    }
    {
        ProfScope ps(prof_expand, s);
        static bool attr_done = false;
        if (!attr_done) {
            cudaFuncSetAttribute(expand_runs_kernel<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(uint32_t) * (1024 + 1023 * 33)));
            cudaFuncSetAttribute(expand_runs_kernel<2048>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(uint32_t) * (2048 + 600 * 66)));
            attr_done = true;
        }
        if (chunk == 2048) {
            const size_t smem = sizeof(uint32_t) * (2048 + (size_t)tiles_x * 66);      // 40 KB at 1080p
            expand_runs_kernel<2048><<<grid, kExpandThreads, smem, s>>>(rkeys_sorted, rvals_sorted, tiles_x, b.chunk_info, b.chunk_counts, b.tile_start, b.vals);
        } else {
            const size_t smem = sizeof(uint32_t) * (1024 + (size_t)tiles_x * 33);      // 139 KB at the 1023-column limit
            expand_runs_kernel<1024><<<grid, kExpandThreads, smem, s>>>(rkeys_sorted, rvals_sorted, tiles_x, b.chunk_info, b.chunk_counts, b.tile_start, b.vals);
        }
        count_launch();
    }
}

// ------------------------------------------------------------------------------------------------
// AMR levels: thresholds v_i = sorted(n_intersections)[int(p_i * num_tiles)], p = {.25,.5,.9}
// (A/rasterizer_impl.cu:609-631: the index is computed in FLOAT), then
// level = 1 + (n > v0) + (n > v1) + (n > v2)   (setAMRLevelsKernel, :190-205)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) amr_levels_kernel(const uint32_t* __restrict__ counts, int n, int k0, int k1, int k2,
                                                          uint32_t* __restrict__ levels, uint32_t* __restrict__ thresholds) {
    __shared__ uint32_t s_hist[3][256];
    __shared__ uint32_t s_prefix[3], s_rank[3];
    const int tid = threadIdx.x;
    if (tid == 0) { s_rank[0] = k0; s_rank[1] = k1; s_rank[2] = k2; s_prefix[0] = s_prefix[1] = s_prefix[2] = 0; }
    // MSB-first radix select, 4 x 8 bits, three ranks at once
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        const uint32_t pmask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
        for (int i = tid; i < 3 * 256; i += 1024) (&s_hist[0][0])[i] = 0;
        __syncthreads();
        for (int i = tid; i < n; i += 1024) {
            const uint32_t v = counts[i];
#pragma unroll
            for (int q = 0; q < 3; ++q)
                if ((v & pmask) == s_prefix[q]) atomicAdd(&s_hist[q][(v >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (tid < 3) {
            uint32_t r = s_rank[tid], acc = 0; int b = 0;
            for (; b < 256; ++b) { const uint32_t h = s_hist[tid][b]; if (acc + h > r) break; acc += h; }
            if (b > 255) b = 255;
            s_rank[tid] = r - acc;
            s_prefix[tid] |= (uint32_t)b << shift;
        }
        __syncthreads();
    }
    const uint32_t v0 = s_prefix[0], v1 = s_prefix[1], v2 = s_prefix[2];
    if (tid < 3) thresholds[tid] = s_prefix[tid];
    for (int i = tid; i < n; i += 1024) {
        const uint32_t c = counts[i];
        levels[i] = 1u + (c > v0) + (c > v1) + (c > v2);
    }
}

void launch_amr_levels(const ImgPtrs& im, int num_tiles, cudaStream_t s) {
    const float pct[3] = {0.25f, 0.5f, 0.9f};
    int k[3];
    for (int i = 0; i < 3; ++i) { k[i] = static_cast<int>(pct[i] * num_tiles); if (k[i] > num_tiles - 1) k[i] = num_tiles - 1; }
    amr_levels_kernel<<<1, 1024, 0, s>>>(im.n_intersections, num_tiles, k[0], k[1], k[2], im.levels, im.thresholds);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// per-step (last, current) level state machine + gaze-driven cap
// ------------------------------------------------------------------------------------------------
struct GazeP { int on; float cx, cy, r[4]; };

__global__ void __launch_bounds__(256) fovea_levels_kernel(int step, uint32_t* __restrict__ last, uint32_t* __restrict__ cur,
                                                           uint32_t* __restrict__ levels, int num_tiles, int tiles_x, int tile_px, GazeP gz) {
    const int t = blockIdx.x * 256 + threadIdx.x;
    if (t >= num_tiles) return;
    uint32_t lvl = levels[t];
    if (gz.on) {
        // extension A7: the finest step a tile may reach is cap = max(1, max{k : dist(tile centre, gaze) <= r_k});
        // round 1 (quarter density) is always rendered, so r_1 only documents the coarse ring.
        const float tx = ((float)(t % tiles_x) + 0.5f) * (float)tile_px, ty = ((float)(t / tiles_x) + 0.5f) * (float)tile_px;
        const float dx = tx - gz.cx, dy = ty - gz.cy;
        const float d2 = dx * dx + dy * dy;
        uint32_t cap = 1;
#pragma unroll
        for (int k = 1; k < 4; ++k) if (d2 <= gz.r[k] * gz.r[k]) cap = k + 1;
        lvl = min(lvl, cap);
        if (step < 0) levels[t] = lvl;             // single-pass mode renders with `levels` directly
    }
    if (step == 0) return;
    if (step >= 1 && step <= 4) {
        const uint32_t l = (step == 1) ? 0u : cur[t];
        last[t] = l;
        cur[t] = (lvl >= (uint32_t)step) ? (uint32_t)step : l;
    } else {          // step < 0 (or > 4): full single-pass render
        last[t] = 0u;
        cur[t] = lvl;
    }
}

void launch_fovea_levels(const ImgPtrs& im, int num_tiles, int fovea_step, const gsr_gaze* gaze, cudaStream_t s) {
    GazeP gz; gz.on = gaze ? 1 : 0; gz.cx = gz.cy = 0.f; gz.r[0] = gz.r[1] = gz.r[2] = gz.r[3] = 0.f;
    if (gaze) { gz.cx = gaze->cx; gz.cy = gaze->cy; for (int i = 0; i < 4; ++i) gz.r[i] = gaze->radii[i]; }
    fovea_levels_kernel<<<(num_tiles + 255) / 256, 256, 0, s>>>(fovea_step, im.levels_last, im.levels_current, im.levels,
                                                              num_tiles, im.tiles_x, im.tile, gz);
    count_launch();
}

}  // namespace gsr
