// gsr_binning.cu -- tile binning for sm_100a: key duplication, onesweep radix sort, tile ranges, AMR levels.
//
//   duplicate_kernel   replaces duplicateWithKeys (rasterizer_impl.cu:70-111) AND the second half of
//                      cub::DeviceScan::InclusiveSum (:277) AND the histogram pass of the sort: a persistent grid
//                      walks 256-Gaussian blocks, scans their tiles_touched in shared memory, emits
//                      key = (tile << 32) | depth_bits / value = Gaussian id with fully coalesced stores
//                      (one thread per PAIR, not per Gaussian -> no divergence on fat Gaussians), and accumulates
//                      the per-pass digit histograms of the keys it just produced.
//   onesweep_kernel    replaces cub::DeviceRadixSort::SortPairs<u64,u32> (rasterizer_impl.cu:303-308): stable LSD
//                      radix sort, one read + one write of every pair per pass, chained-scan ("decoupled
//                      look-back") prefix across tiles, 8- or 9-bit digits (45 key bits -> 5 passes).
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

// ------------------------------------------------------------------------------------------------
// duplicate + per-pass histograms
// ------------------------------------------------------------------------------------------------
struct DupParams {
    int P, nblocks, tile, grid_x, grid_y;
    const int* __restrict__ radii;
    const float2* __restrict__ means2D;
    const float* __restrict__ depths;
    const uint32_t* __restrict__ tiles_touched;
    const uint32_t* __restrict__ block_offsets;
    uint64_t* __restrict__ keys;
    uint32_t* __restrict__ vals;
    uint32_t* hist;                 // [passes][kSortMaxRadix] global counts, or nullptr
    int passes; int shift[kSortMaxPasses]; int bits[kSortMaxPasses];
};

constexpr int kDupThreads = 256;

__global__ void __launch_bounds__(kDupThreads) duplicate_kernel(const DupParams p) {
    extern __shared__ uint32_t s_hist[];            // [passes][kSortMaxRadix]
    __shared__ uint32_t s_off[kPreBlock + 1];
    __shared__ uint32_t s_x0[kPreBlock], s_y0[kPreBlock], s_w[kPreBlock], s_depth[kPreBlock];
    __shared__ uint32_t s_wsum[kDupThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const bool do_hist = (p.hist != nullptr);
    if (do_hist) for (int i = tid; i < p.passes * kSortMaxRadix; i += kDupThreads) s_hist[i] = 0;

    for (int blk = blockIdx.x; blk < p.nblocks; blk += gridDim.x) {
        __syncthreads();                            // previous iteration's smem readers are done
        const int idx = blk * kPreBlock + tid;
        uint32_t t = 0;
        if (idx < p.P) {
            t = p.tiles_touched[idx];
            if (t > 0) {
                const float2 xy = p.means2D[idx];
                uint32_t x0, y0, x1, y1;
                get_rect(xy.x, xy.y, p.radii[idx], p.tile, p.grid_x, p.grid_y, x0, y0, x1, y1);
                s_x0[tid] = x0; s_y0[tid] = y0; s_w[tid] = x1 - x0;
                s_depth[tid] = __float_as_uint(p.depths[idx]);
            }
        }
        // exclusive scan of t over the CTA
        uint32_t inc = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
        if (lane == 31) s_wsum[wid] = inc;
        __syncthreads();
        uint32_t wbase = 0;
#pragma unroll
        for (int w = 0; w < kDupThreads / 32; ++w) wbase += (w < wid) ? s_wsum[w] : 0u;
        s_off[tid] = wbase + inc - t;
        if (tid == kDupThreads - 1) s_off[kPreBlock] = wbase + inc;
        __syncthreads();

        const uint32_t total = s_off[kPreBlock];
        const uint32_t gbase = p.block_offsets[blk];
        for (uint32_t q0 = 0; q0 < total; q0 += kDupThreads) {
            const uint32_t q = q0 + tid;
            const bool valid = q < total;
            uint64_t key = 0;
            if (valid) {
                // largest g with s_off[g] <= q  (s_off non-decreasing, s_off[256] = total > q)
                int lo = 0, hi = kPreBlock;
#pragma unroll
                for (int it = 0; it < 8; ++it) { int mid = (lo + hi) >> 1; if (s_off[mid] <= q) lo = mid; else hi = mid; }
                const int g = lo;
                const uint32_t k = q - s_off[g];
                const uint32_t w = s_w[g];
                const uint32_t ty = s_y0[g] + k / w, tx = s_x0[g] + k % w;
                key = ((uint64_t)(ty * (uint32_t)p.grid_x + tx) << 32) | (uint64_t)s_depth[g];
                p.keys[(size_t)gbase + q] = key;
                p.vals[(size_t)gbase + q] = (uint32_t)(blk * kPreBlock + g);
            }
            if (do_hist) {
                const uint32_t act = __ballot_sync(0xffffffffu, valid);
                if (valid) {
                    for (int ps = 0; ps < p.passes; ++ps) {
                        const uint32_t d = (uint32_t)(key >> p.shift[ps]) & ((1u << p.bits[ps]) - 1u);
                        const uint32_t peers = __match_any_sync(act, d);
                        if (lane == (31 - __clz(peers))) atomicAdd(&s_hist[ps * kSortMaxRadix + d], (uint32_t)__popc(peers));
                    }
                }
            }
        }
    }
    if (do_hist) {
        __syncthreads();
        for (int i = tid; i < p.passes * kSortMaxRadix; i += kDupThreads) {
            const uint32_t h = s_hist[i];
            if (h) atomicAdd(&p.hist[i], h);
        }
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

void launch_duplicate(const GeomPtrs& g, int P, const int* radii, int tile, int grid_x, int grid_y,
                      uint64_t* keys, uint32_t* vals, uint32_t* hist, const SortPlan& plan, cudaStream_t s) {
    DupParams p;
    p.P = P; p.nblocks = g.nblocks; p.tile = tile; p.grid_x = grid_x; p.grid_y = grid_y;
    p.radii = radii; p.means2D = g.means2D; p.depths = g.depths; p.tiles_touched = g.tiles_touched;
    p.block_offsets = g.block_offsets; p.keys = keys; p.vals = vals; p.hist = hist;
    p.passes = plan.passes;
    for (int i = 0; i < kSortMaxPasses; ++i) { p.shift[i] = i < plan.passes ? plan.shift[i] : 0; p.bits[i] = i < plan.passes ? plan.bits[i] : 1; }
    int grid = num_sms() * 4;
    if (grid > g.nblocks) grid = g.nblocks;
    if (grid < 1) return;
    size_t smem = hist ? sizeof(uint32_t) * plan.passes * kSortMaxRadix : 0;
    duplicate_kernel<<<grid, kDupThreads, smem, s>>>(p);
    count_launch();
}

// ------------------------------------------------------------------------------------------------
// standalone histogram (used by gsr_sort_pairs when keys do not come from duplicate_kernel)
// ------------------------------------------------------------------------------------------------
struct HistParams { const uint64_t* keys; long long n; uint32_t* hist; int passes; int shift[kSortMaxPasses]; int bits[kSortMaxPasses]; };

__global__ void __launch_bounds__(256) sort_hist_kernel(const HistParams p) {
    extern __shared__ uint32_t s_hist[];
    const int tid = threadIdx.x, lane = tid & 31;
    for (int i = tid; i < p.passes * kSortMaxRadix; i += 256) s_hist[i] = 0;
    __syncthreads();
    const long long stride = (long long)gridDim.x * 256;
    const long long n_round = (p.n + 31) / 32 * 32;
    for (long long i = (long long)blockIdx.x * 256 + tid; i < n_round; i += stride) {
        const bool valid = i < p.n;
        const uint64_t key = valid ? p.keys[i] : 0;
        const uint32_t act = __ballot_sync(0xffffffffu, valid);
        if (valid) {
            for (int ps = 0; ps < p.passes; ++ps) {
                const uint32_t d = (uint32_t)(key >> p.shift[ps]) & ((1u << p.bits[ps]) - 1u);
                const uint32_t peers = __match_any_sync(act, d);
                if (lane == (31 - __clz(peers))) atomicAdd(&s_hist[ps * kSortMaxRadix + d], (uint32_t)__popc(peers));
            }
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
// onesweep pass
// ------------------------------------------------------------------------------------------------
constexpr int SORT_THREADS = 512;
constexpr int SORT_ITEMS = 9;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;
constexpr int SORT_WARPS = SORT_THREADS / 32;
static_assert(SORT_TILE == kSortTileItems, "tile size mismatch with gsr_common.cuh");

constexpr uint32_t ST_AGG = 1u << 30, ST_INC = 2u << 30, ST_MASK = (1u << 30) - 1u;

struct SweepParams {
    const uint64_t* __restrict__ keys_in; uint64_t* __restrict__ keys_out;
    const uint32_t* __restrict__ vals_in; uint32_t* __restrict__ vals_out;
    long long n; int shift; int bits;
    const uint32_t* __restrict__ bin_base;   // [RADIX] exclusive global offsets of this pass
    uint32_t* status;                        // [tiles][RADIX]
    uint32_t* tile_counter;
    uint32_t* err_flag;
};

template <int RADIX>
__global__ void __launch_bounds__(SORT_THREADS, 2) onesweep_kernel(const SweepParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* s_keys = reinterpret_cast<uint64_t*>(smem_raw);                       // [SORT_TILE]
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(smem_raw);                       // aliases s_keys (after keys are out)
    uint32_t* s_whist = reinterpret_cast<uint32_t*>(smem_raw + sizeof(uint64_t) * SORT_TILE);   // [SORT_WARPS][RADIX]
    uint32_t* s_bin_start = s_whist + SORT_WARPS * RADIX;                            // [RADIX]
    uint32_t* s_gofs = s_bin_start + RADIX;                                          // [RADIX]
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_wtot[RADIX / 32];

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(p.tile_counter, 1u);
    for (int i = tid; i < SORT_WARPS * RADIX; i += SORT_THREADS) s_whist[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const long long tile_base = (long long)tile * SORT_TILE;
    const int count = (int)min((long long)SORT_TILE, p.n - tile_base);
    const uint32_t mask = (1u << p.bits) - 1u;

    // ---- load keys, warp-striped: item i of lane l sits at warp_base + i*32 + l
    uint64_t key[SORT_ITEMS];
    const int wbase = wid * 32 * SORT_ITEMS;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const int lp = wbase + i * 32 + lane;
        key[i] = (lp < count) ? p.keys_in[tile_base + lp] : ~0ull;
    }

    // ---- stable ranks inside the warp (match-any multi-split)
    uint32_t rank[SORT_ITEMS];
    uint32_t* wh = s_whist + wid * RADIX;
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const uint32_t d = (uint32_t)(key[i] >> p.shift) & mask;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const uint32_t pre = wh[d];
        __syncwarp();
        if ((peers & lt) == 0) wh[d] = pre + __popc(peers);
        __syncwarp();
        rank[i] = pre + __popc(peers & lt);
    }
    __syncthreads();

    // ---- per-bin: exclusive scan across warps, publish tile aggregate, look back
    uint32_t bin_count = 0;
    if (tid < RADIX) {
        uint32_t acc = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w) { const uint32_t t = s_whist[w * RADIX + tid]; s_whist[w * RADIX + tid] = acc; acc += t; }
        bin_count = acc;
        volatile uint32_t* st = p.status + (size_t)tile * RADIX + tid;
        *st = (tile == 0 ? ST_INC : ST_AGG) | acc;
    }
    // tile-local exclusive scan of bin counts (warps 0 .. RADIX/32-1)
    {
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
    if (tid < RADIX) {
        uint32_t excl = 0;
        if (tile > 0) {
            long long t = (long long)tile - 1;
            uint32_t spins = 0;
            while (true) {
                const uint32_t v = *((volatile uint32_t*)(p.status + (size_t)t * RADIX + tid));
                const uint32_t f = v & ~ST_MASK;
                if (f == 0) {
                    __nanosleep(32);
                    if (++spins > (1u << 24)) { atomicOr(p.err_flag, 1u); break; }   // watchdog: never hang the GPU
                    continue;
                }
                excl += v & ST_MASK;
                if (f == ST_INC) break;
                --t;
            }
            *((volatile uint32_t*)(p.status + (size_t)tile * RADIX + tid)) = ST_INC | ((excl + bin_count) & ST_MASK);
        }
        s_gofs[tid] = p.bin_base[tid] + excl;
    }
    __syncthreads();

    // ---- reorder keys through shared memory so that each bin is one contiguous run
    uint32_t lpos[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const uint32_t d = (uint32_t)(key[i] >> p.shift) & mask;
        lpos[i] = s_bin_start[d] + wh[d] + rank[i];
        s_keys[lpos[i]] = key[i];
    }
    __syncthreads();
    uint32_t gpos[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const int j = tid + i * SORT_THREADS;
        gpos[i] = 0xffffffffu;
        if (j < count) {
            const uint64_t k = s_keys[j];
            const uint32_t d = (uint32_t)(k >> p.shift) & mask;
            gpos[i] = s_gofs[d] + ((uint32_t)j - s_bin_start[d]);
            p.keys_out[gpos[i]] = k;
        }
    }
    __syncthreads();

    // ---- values follow the same permutation
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const int lp = wbase + i * 32 + lane;
        if (lp < count) s_vals[lpos[i]] = p.vals_in[tile_base + lp];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const int j = tid + i * SORT_THREADS;
        if (j < count) p.vals_out[gpos[i]] = s_vals[j];
    }
}

template <int RADIX>
static size_t sweep_smem() { return sizeof(uint64_t) * SORT_TILE + sizeof(uint32_t) * (SORT_WARPS * RADIX + 2 * RADIX); }

struct SortTemp { uint32_t* hist; uint32_t* counters; uint32_t* status; };
static SortTemp carve_sort_temp(void* temp, long long n) {
    long long tiles = (n + SORT_TILE - 1) / SORT_TILE; if (tiles < 1) tiles = 1;
    char* b = aligned_base(temp);
    SortTemp t;
    t.hist = (uint32_t*)b; b += align_up(sizeof(uint32_t) * kSortMaxPasses * kSortMaxRadix);
    t.counters = (uint32_t*)b; b += align_up(256);
    t.status = (uint32_t*)b;
    return t;
}
uint32_t* sort_hist_ptr(void* temp) { return (uint32_t*)aligned_base(temp); }

int launch_onesweep(const uint64_t* kin0, const uint32_t* vin0, uint64_t* kout, uint32_t* vout, uint64_t* ktmp, uint32_t* vtmp,
                    long long n, const SortPlan& plan, void* temp, size_t temp_bytes, bool hist_ready, uint32_t* err_flag, cudaStream_t s) {
    if (n <= 0) return GSR_OK;
    if (n >= (1ll << 30)) { set_error("radix sort: n=%lld exceeds 2^30 pairs", n); return GSR_ERR_INVALID; }
    if (temp_bytes < sort_temp_bytes(n)) { set_error("radix sort: temp too small"); return GSR_ERR_ALLOC; }
    SortTemp t = carve_sort_temp(temp, n);
    const long long tiles = (n + SORT_TILE - 1) / SORT_TILE;
    static bool attr_done = false;
    if (!attr_done) {
        cudaFuncSetAttribute(onesweep_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sweep_smem<256>());
        cudaFuncSetAttribute(onesweep_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sweep_smem<512>());
        attr_done = true;
    }
    if (!hist_ready) {
        cudaMemsetAsync(t.hist, 0, sizeof(uint32_t) * kSortMaxPasses * kSortMaxRadix, s);
        HistParams hp; hp.keys = kin0; hp.n = n; hp.hist = t.hist; hp.passes = plan.passes;
        for (int i = 0; i < kSortMaxPasses; ++i) { hp.shift[i] = i < plan.passes ? plan.shift[i] : 0; hp.bits[i] = i < plan.passes ? plan.bits[i] : 1; }
        long long want = (n + 256 * 16 - 1) / (256 * 16);
        int grid = (int)min((long long)num_sms() * 4, want); if (grid < 1) grid = 1;
        sort_hist_kernel<<<grid, 256, sizeof(uint32_t) * plan.passes * kSortMaxRadix, s>>>(hp);
        count_launch();
    }
    hist_scan_kernel<<<plan.passes, kSortMaxRadix, 0, s>>>(t.hist);
    count_launch();
    cudaMemsetAsync(t.counters, 0, 256, s);
    const uint64_t* kin = kin0; const uint32_t* vin = vin0;
    uint64_t* kb[2]; uint32_t* vb[2];
    if (plan.passes & 1) { kb[0] = kout; vb[0] = vout; kb[1] = ktmp; vb[1] = vtmp; }
    else                 { kb[0] = ktmp; vb[0] = vtmp; kb[1] = kout; vb[1] = vout; }
    for (int ps = 0; ps < plan.passes; ++ps) {
        SweepParams sp;
        sp.keys_in = kin; sp.keys_out = kb[ps & 1]; sp.vals_in = vin; sp.vals_out = vb[ps & 1];
        sp.n = n; sp.shift = plan.shift[ps]; sp.bits = plan.bits[ps];
        sp.bin_base = t.hist + ps * kSortMaxRadix; sp.status = t.status; sp.tile_counter = t.counters + ps; sp.err_flag = err_flag;
        if (plan.bits[ps] <= 8) {
            cudaMemsetAsync(t.status, 0, sizeof(uint32_t) * (size_t)tiles * 256, s);
            ProfScope pr(GSR_PROF_SORT_PASS0 + ps, s);
            onesweep_kernel<256><<<(unsigned)tiles, SORT_THREADS, sweep_smem<256>(), s>>>(sp);
        } else {
            cudaMemsetAsync(t.status, 0, sizeof(uint32_t) * (size_t)tiles * 512, s);
            ProfScope pr(GSR_PROF_SORT_PASS0 + ps, s);
            onesweep_kernel<512><<<(unsigned)tiles, SORT_THREADS, sweep_smem<512>(), s>>>(sp);
        }
        count_launch();
        kin = kb[ps & 1]; vin = vb[ps & 1];
    }
    return GSR_OK;
}

// ------------------------------------------------------------------------------------------------
// tile ranges by binary search on the sorted keys (+ AMR intersection counts)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t lower_bound_tile(const uint64_t* __restrict__ keys, uint32_t n, uint32_t tile) {
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if ((uint32_t)(__ldg(keys + mid) >> 32) < tile) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(128) tile_ranges_kernel(const uint64_t* __restrict__ keys, uint32_t n, uint2* __restrict__ ranges,
                                                          uint32_t* __restrict__ n_intersections, int num_tiles) {
    const int t = blockIdx.x * 128 + threadIdx.x;
    if (t >= num_tiles) return;
    uint32_t lo = 0, hi = 0;
    if (n > 0) { lo = lower_bound_tile(keys, n, (uint32_t)t); hi = lower_bound_tile(keys, n, (uint32_t)t + 1u); }
    // empty tiles keep the reference's memset value (0,0)
    ranges[t] = (hi > lo) ? make_uint2(lo, hi) : make_uint2(0u, 0u);
    if (n_intersections) n_intersections[t] = hi - lo;
}

void launch_tile_ranges(const uint64_t* keys_sorted, long long n, uint2* ranges, uint32_t* n_intersections, int num_tiles, cudaStream_t s) {
    tile_ranges_kernel<<<(num_tiles + 127) / 128, 128, 0, s>>>(keys_sorted, (uint32_t)n, ranges, n_intersections, num_tiles);
    count_launch();
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
