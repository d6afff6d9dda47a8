// gsr_common.cuh -- shared definitions for the sm_100a rasterizer kernels.
//
// Work-buffer layouts replace the reference's GeometryState / BinningState / ImageState
// (submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.h:31-63 and
// rasterizer_impl.cu:155-194).  They are NOT byte-compatible with the reference (callers only ever
// thread the opaque byte tensors back into this library); they are self-describing through a
// device-side header so that backward / AMR steps can find num_rendered without host state.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/gsr_b200.h"

namespace gsr {

constexpr int kAlign = 256;
constexpr int kPreBlock = 256;          // Gaussians per preprocess / duplicate CTA
constexpr uint32_t kGeomMagic = 0x47535231u;  // "GSR1"

__host__ __device__ inline size_t align_up(size_t x, size_t a = kAlign) { return (x + a - 1) / a * a; }

// ---------------------------------------------------------------- radix sort temp sizing
constexpr int kSortTileItems = 256 * 8;    // smallest tile of any onesweep variant (gsr_binning.cu); sizes the look-back words
constexpr int kSortMaxPasses = 8;
constexpr int kSortMaxRadix = 512;

__host__ __device__ inline size_t sort_temp_bytes(long long n) {
    long long tiles = (n + kSortTileItems - 1) / kSortTileItems;
    if (tiles < 1) tiles = 1;
    size_t hist = sizeof(uint32_t) * kSortMaxPasses * kSortMaxRadix;        // global digit histograms
    size_t counters = 256;                                                  // per-pass tile counters + flags
    size_t status = sizeof(uint32_t) * (size_t)tiles * kSortMaxRadix;       // look-back words, ONE pass (re-zeroed per pass)
    return ((hist + 255) / 256 * 256) + 256 + ((status + 255) / 256 * 256) + 256;
}

// ---------------------------------------------------------------- geometry buffer
// fused gradient exchange (gsr_grads.push_*): a 256-Gaussian block travels as 94 lines of 128 bytes (30 floats + 2 flag words)
// in a slot of 96 lines = 768 float4 = 12 KB
constexpr int kPushLines = 94;
constexpr int kPushSlotF4 = 768;

struct GeomHeader {      // lives at the start of the (aligned) geometry buffer, device memory
    uint32_t magic;
    uint32_t P;
    uint32_t num_rendered;   // = sum(tiles_touched); reference reads point_offsets[P-1] (rasterizer_impl.cu:281)
    uint32_t sort_error;     // bit 0: radix-sort watchdog; bit 8: a Gaussian was culled although prefiltered
    uint32_t num_runs;       // = sum over visible Gaussians of the tile rows their rect spans ((Gaussian, row) runs)
    uint32_t num_visible;    // Gaussians with tiles_touched > 0: only these are depth-sorted
    // ---- depth sort (gsr_depthsort.cu)
    uint32_t done_hist;      // CTAs of bin_scan_kernel that have finished (the last one posts the frame totals)
    uint32_t shift;          // bin = (key - kDepthKey0) >> shift
    uint32_t sort_mode;      // 1 = bins valid; 2 = a bin exceeds kBinCap / a key beyond the binned range -> the host runs the
                             // onesweep fallback; 3 = fallback done
    uint32_t max_bin;        // records in the fullest bin
    uint32_t comp32;         // 1: (position of the key inside its bin, id) fits one 32-bit composite
    uint32_t pad[53];
};
static_assert(sizeof(GeomHeader) == 256, "header is one alignment unit");

// Frame statistics accumulated by preprocess_fwd_kernel with atomics.  Same-address global atomics serialise (~30 ns each
// on B200), so the 11 719 CTAs of a 3 M-Gaussian frame are spread over kStatSlots slots, one 128-byte line each.
constexpr int kStatSlots = 128;
struct FrameStatSlot { uint32_t num_rendered, flags /* 1: culled although prefiltered, 2: depth key beyond the binned range */, pad[30]; };
static_assert(sizeof(FrameStatSlot) == 128, "one line per slot");

// ---------------------------------------------------------------- depth sort sizing
// Depth keys are the float bits of the view-space depth, > 0.2 by the near-plane cull (auxiliary.h:139-164).  The bins cover
// keys [kDepthKey0, kDepthKey0 + 2^28) -- depths up to ~1e6 -- at a FIXED resolution, so that the preprocess kernel itself can
// count them (no key-range pass): 2^shift keys per bin, shift chosen so that (key position inside the bin, Gaussian index)
// fits one 32-bit word for the scene size.
constexpr uint32_t kDepthKey0 = 0x3E4CCCCDu;                  // bits of 0.2f
constexpr int kDepthSpanLog2 = 28;
constexpr int kBinCap = 512;                                  // records one warp orders by all-pairs ranking; fuller bin -> fallback
constexpr int kSegBins = 1024;                                // bins per scan segment
__host__ __device__ inline int index_bits(int P) { int b = 1; while (b < 31 && (1ll << b) < (long long)P) ++b; return b; }
// ~10 records per populated bin at the densities of trained scenes (measured on the 3 M-Gaussian bench scene: shift 8 -> mean 10,
// max 83; shift 10 -> mean 33, max 270 and the all-pairs ranking of the full bins dominates the frame)
__host__ __device__ inline int depth_shift(int P) { int sh = 30 - index_bits(P); return sh > 10 ? 10 : (sh < 6 ? 6 : sh); }
__host__ __device__ inline int depth_bins(int P) { return 1 << (kDepthSpanLog2 - depth_shift(P)); }       // 256 Ki .. 4 Mi
constexpr unsigned long long kBinPairBudget = 48ull << 20;    // sum of n^2 over the bins with more than 32 records above which the
                                                              // frame takes the onesweep fallback (the all-pairs ranking would cost more)

struct GeomLayout {
    size_t header, depths, clamped, radii, means2D, cov3D, conic_opacity, rgb, tiles_touched, rects,
           block_sums, block_sums_r, block_sums_v, block_offsets, dkeys_a, dkeys_b, order_tmp, order, oblock_sums, oblock_offsets, sort_temp,
           stat_slots, brec, crank, seg_tot, seg_pre, runoff, srect, bins64, bin_cnt, total;
    int nblocks, nbins;
    __host__ __device__ explicit GeomLayout(int P) {
        nblocks = (P + kPreBlock - 1) / kPreBlock;
        size_t o = 0;
        header = o;        o = align_up(o + sizeof(GeomHeader));
        depths = o;        o = align_up(o + sizeof(float) * (size_t)P);
        clamped = o;       o = align_up(o + 3 * (size_t)P);
        radii = o;         o = align_up(o + sizeof(int) * (size_t)P);
        means2D = o;       o = align_up(o + sizeof(float2) * (size_t)P);
        cov3D = o;         o = align_up(o + sizeof(float) * 6 * (size_t)P);
        conic_opacity = o; o = align_up(o + sizeof(float4) * (size_t)P);
        rgb = o;           o = align_up(o + sizeof(float) * 3 * (size_t)P);
        tiles_touched = o; o = align_up(o + sizeof(uint32_t) * (size_t)P);
        rects = o;         o = align_up(o + sizeof(uint2) * (size_t)P);
        block_sums = o;    o = align_up(o + sizeof(uint32_t) * (size_t)(nblocks + 1));
        block_sums_r = o;  o = align_up(o + sizeof(uint32_t) * (size_t)(nblocks + 1));
        block_sums_v = o;  o = align_up(o + sizeof(uint32_t) * (size_t)(nblocks + 1));
        block_offsets = o; o = align_up(o + sizeof(uint32_t) * (size_t)(nblocks + 1));
        // depth sort of the Gaussians (keys ping-pong, index ping-pong) and the scan in sorted order
        dkeys_a = o;       o = align_up(o + sizeof(uint32_t) * (size_t)P);
        dkeys_b = o;       o = align_up(o + sizeof(uint32_t) * (size_t)P);
        order_tmp = o;     o = align_up(o + sizeof(uint32_t) * (size_t)P);
        order = o;         o = align_up(o + sizeof(uint32_t) * (size_t)P);
        oblock_sums = o;   o = align_up(o + sizeof(uint32_t) * (size_t)(nblocks + 1));
        oblock_offsets = o; o = align_up(o + sizeof(uint32_t) * (size_t)(nblocks + 1));
        sort_temp = o;     o = align_up(o + sort_temp_bytes(P));
        // depth sort: per-block compaction (dkeys_a = keys, order_tmp = ids, dkeys_b = rows, order = x-pack, crank = rank inside the
        // bin) -> brec (binned records) -> order, runoff, srect.  stat_slots + bins64 are cleared by ONE memset node per frame.
        nbins = depth_bins(P);
        brec = o;          o = align_up(o + sizeof(uint4) * (size_t)P);
        crank = o;         o = align_up(o + sizeof(uint32_t) * (size_t)P);
        seg_tot = o;       o = align_up(o + sizeof(uint4) * (size_t)(nbins / kSegBins));   // (records, runs, n^2 of the full bins, fullest bin) per scan segment
        seg_pre = o;       o = align_up(o + sizeof(uint2) * (size_t)(nbins / kSegBins));   // exclusive prefix of (records, runs) over the segments
        runoff = o;        o = align_up(o + sizeof(uint32_t) * ((size_t)P + 1));
        srect = o;         o = align_up(o + sizeof(uint2) * (size_t)P);
        bin_cnt = o;       o = align_up(o + sizeof(uint32_t) * (size_t)nbins);
        stat_slots = o;    o = align_up(o + sizeof(FrameStatSlot) * kStatSlots);
        bins64 = o;        o = align_up(o + sizeof(unsigned long long) * (size_t)nbins);      // directly behind stat_slots
        total = o;
    }
};

struct GeomPtrs {
    GeomHeader* header;
    float* depths; unsigned char* clamped; int* radii; float2* means2D; float* cov3D;
    float4* conic_opacity; float* rgb; uint32_t* tiles_touched; uint2* rects; uint32_t* block_sums; uint32_t* block_sums_r; uint32_t* block_sums_v;
    uint32_t* block_offsets;
    uint32_t *dkeys_a, *dkeys_b, *order_tmp, *order, *oblock_sums, *oblock_offsets; void* sort_temp; size_t sort_temp_bytes;
    FrameStatSlot* stat_slots; uint4* brec; uint4* seg_tot; uint2* seg_pre; uint32_t *runoff, *crank, *bin_cnt; unsigned long long* bins64; uint2* srect;
    size_t frame_zero_bytes;      // stat_slots .. end of bins64: what the per-frame memset node clears
    int nblocks, nbins;
};

inline char* aligned_base(const void* p) { return (char*)align_up((size_t)p); }

inline GeomPtrs geom_ptrs(const void* buf, int P) {
    GeomLayout L(P);
    char* b = aligned_base(buf);
    GeomPtrs g;
    g.header = (GeomHeader*)(b + L.header);
    g.depths = (float*)(b + L.depths);
    g.clamped = (unsigned char*)(b + L.clamped);
    g.radii = (int*)(b + L.radii);
    g.means2D = (float2*)(b + L.means2D);
    g.cov3D = (float*)(b + L.cov3D);
    g.conic_opacity = (float4*)(b + L.conic_opacity);
    g.rgb = (float*)(b + L.rgb);
    g.tiles_touched = (uint32_t*)(b + L.tiles_touched);
    g.rects = (uint2*)(b + L.rects);
    g.block_sums = (uint32_t*)(b + L.block_sums);
    g.block_sums_r = (uint32_t*)(b + L.block_sums_r);
    g.block_sums_v = (uint32_t*)(b + L.block_sums_v);
    g.block_offsets = (uint32_t*)(b + L.block_offsets);
    g.dkeys_a = (uint32_t*)(b + L.dkeys_a); g.dkeys_b = (uint32_t*)(b + L.dkeys_b);
    g.order_tmp = (uint32_t*)(b + L.order_tmp); g.order = (uint32_t*)(b + L.order);
    g.oblock_sums = (uint32_t*)(b + L.oblock_sums); g.oblock_offsets = (uint32_t*)(b + L.oblock_offsets);
    g.sort_temp = (void*)(b + L.sort_temp); g.sort_temp_bytes = sort_temp_bytes(P);
    g.stat_slots = (FrameStatSlot*)(b + L.stat_slots); g.brec = (uint4*)(b + L.brec); g.seg_tot = (uint4*)(b + L.seg_tot); g.seg_pre = (uint2*)(b + L.seg_pre); g.runoff = (uint32_t*)(b + L.runoff);
    g.crank = (uint32_t*)(b + L.crank); g.bin_cnt = (uint32_t*)(b + L.bin_cnt); g.bins64 = (unsigned long long*)(b + L.bins64); g.srect = (uint2*)(b + L.srect);
    g.frame_zero_bytes = L.total - L.stat_slots;
    g.nblocks = L.nblocks; g.nbins = L.nbins;
    return g;
}

// ---------------------------------------------------------------- binning buffer
// point_list[N] (the only part backward needs, at offset 0), then the run-length binning scratch: a "run" is one
// (Gaussian, tile row) span of consecutive tiles, key = (row << 20) | (x0 << 10) | width, value = Gaussian id.
constexpr int kRunChunk = 1024;          // runs per expansion chunk (a chunk never straddles two tile rows)
constexpr int kRunRowShift = 20, kRunX0Shift = 10, kRunFieldMask = 1023, kRunMaxRows = 4095;

struct BinLayout {
    size_t vals, rkeyA, rkeyB, rkeyC, rvalA, rvalB, sort_temp, row_start, chunk_row_start, chunk_counts, chunk_info, tile_start, total;
    long long max_chunks;
    __host__ __device__ BinLayout(long long N, long long R, int tiles_x, int tiles_y) {
        size_t n = (size_t)(N > 0 ? N : 0), r = (size_t)(R > 0 ? R : 0);
        max_chunks = (long long)(r / kRunChunk) + tiles_y + 1;
        size_t o = 0;
        vals = o;  o = align_up(o + 4 * n);
        rkeyA = o; o = align_up(o + 4 * r);
        rkeyB = o; o = align_up(o + 4 * r);
        rkeyC = o; o = align_up(o + (tiles_y > kSortMaxRadix ? 4 * r : 0));   // ping-pong partner of a two-pass row sort (> 512 rows)
        rvalA = o; o = align_up(o + 4 * r);
        rvalB = o; o = align_up(o + 4 * r);
        sort_temp = o; o = align_up(o + sort_temp_bytes((long long)r));
        row_start = o; o = align_up(o + 4 * (size_t)(tiles_y + 2));
        chunk_row_start = o; o = align_up(o + 4 * (size_t)(tiles_y + 2));
        chunk_counts = o; o = align_up(o + 4 * (size_t)max_chunks * (size_t)tiles_x);
        chunk_info = o; o = align_up(o + 16 * (size_t)max_chunks);     // (row, first run, end run, valid) per chunk
        tile_start = o; o = align_up(o + 4 * ((size_t)tiles_x * tiles_y + 1));
        total = o;
    }
};

struct BinPtrs {
    uint32_t *vals, *rkeyA, *rkeyB, *rkeyC, *rvalA, *rvalB; void* sort_temp; size_t sort_temp_bytes;
    uint32_t *row_start, *chunk_row_start, *chunk_counts, *tile_start; uint4* chunk_info; long long max_chunks;
};

inline BinPtrs bin_ptrs(const void* buf, long long N, long long R, int tiles_x, int tiles_y) {
    BinLayout L(N, R, tiles_x, tiles_y);
    char* b = aligned_base(buf);
    BinPtrs p;
    p.vals = (uint32_t*)(b + L.vals);
    p.rkeyA = (uint32_t*)(b + L.rkeyA); p.rkeyB = (uint32_t*)(b + L.rkeyB); p.rkeyC = (uint32_t*)(b + L.rkeyC);
    p.rvalA = (uint32_t*)(b + L.rvalA); p.rvalB = (uint32_t*)(b + L.rvalB);
    p.sort_temp = (void*)(b + L.sort_temp); p.sort_temp_bytes = sort_temp_bytes(R > 0 ? R : 0);
    p.row_start = (uint32_t*)(b + L.row_start); p.chunk_row_start = (uint32_t*)(b + L.chunk_row_start);
    p.chunk_counts = (uint32_t*)(b + L.chunk_counts); p.tile_start = (uint32_t*)(b + L.tile_start);
    p.chunk_info = (uint4*)(b + L.chunk_info);
    p.max_chunks = L.max_chunks;
    return p;
}
// the sorted Gaussian-id list (reference: point_list) sits at the start of the buffer
inline uint32_t* bin_point_list(const void* buf) { return (uint32_t*)aligned_base(buf); }

// ---------------------------------------------------------------- image buffer
struct ImgLayout {
    size_t accum_alpha, n_contrib, ranges, n_intersections, levels, levels_last, levels_current, thresholds, total;
    int tiles_x, tiles_y, tile;
    __host__ __device__ ImgLayout(int W, int H, int mode) {
        tile = (mode == GSR_MODE_AMR) ? 32 : 16;
        tiles_x = (W + tile - 1) / tile; tiles_y = (H + tile - 1) / tile;
        size_t px = (size_t)W * H, nt = (size_t)tiles_x * tiles_y;
        size_t o = 0;
        accum_alpha = o; o = align_up(o + 4 * px);
        n_contrib = o;   o = align_up(o + 4 * px);
        ranges = o;      o = align_up(o + 8 * nt);
        n_intersections = levels = levels_last = levels_current = thresholds = 0;
        if (mode == GSR_MODE_AMR) {
            n_intersections = o; o = align_up(o + 4 * nt);
            levels = o;          o = align_up(o + 4 * nt);
            levels_last = o;     o = align_up(o + 4 * nt);
            levels_current = o;  o = align_up(o + 4 * nt);
            thresholds = o;      o = align_up(o + 16);
        }
        total = o;
    }
};

struct ImgPtrs {
    float* accum_alpha; uint32_t* n_contrib; uint2* ranges;
    uint32_t *n_intersections, *levels, *levels_last, *levels_current, *thresholds;
    int tiles_x, tiles_y, tile;
};

inline ImgPtrs img_ptrs(const void* buf, int W, int H, int mode) {
    ImgLayout L(W, H, mode);
    char* b = aligned_base(buf);
    ImgPtrs p;
    p.accum_alpha = (float*)(b + L.accum_alpha);
    p.n_contrib = (uint32_t*)(b + L.n_contrib);
    p.ranges = (uint2*)(b + L.ranges);
    bool amr = (mode == GSR_MODE_AMR);
    p.n_intersections = amr ? (uint32_t*)(b + L.n_intersections) : nullptr;
    p.levels = amr ? (uint32_t*)(b + L.levels) : nullptr;
    p.levels_last = amr ? (uint32_t*)(b + L.levels_last) : nullptr;
    p.levels_current = amr ? (uint32_t*)(b + L.levels_current) : nullptr;
    p.thresholds = amr ? (uint32_t*)(b + L.thresholds) : nullptr;
    p.tiles_x = L.tiles_x; p.tiles_y = L.tiles_y; p.tile = L.tile;
    return p;
}

// ---------------------------------------------------------------- launch bookkeeping / errors
void count_launch(int n = 1);
void set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
int check_last(const char* what, bool debug, cudaStream_t s);

// ---------------------------------------------------------------- optional per-stage CUDA-event timing
// (gsr_profile_enable / gsr_profile_read in the C ABI).  No-ops unless enabled; events are recorded on the
// caller's stream, so the figures are device time of the stage inside a real step, not a separate micro-run.
void prof_begin(int slot, cudaStream_t s);
void prof_end(int slot, cudaStream_t s);
struct ProfScope {
    int slot; cudaStream_t s;
    ProfScope(int slot_, cudaStream_t s_) : slot(slot_), s(s_) { prof_begin(slot, s); }
    ~ProfScope() { prof_end(slot, s); }
};

// ---------------------------------------------------------------- stage entry points (host)
struct PreprocessArgs {
    int P, D, M;
    const float *means3D, *scales, *rotations, *opacities, *shs, *cov3D_precomp, *colors_precomp;
    const float* shs_rest = nullptr; bool raw = false;       // raw-parameter entry (gsr_scene.raw_params)
    float scale_modifier;
    const float *viewmatrix, *projmatrix, *cam_pos;
    int W, H; float tan_fovx, tan_fovy, focal_x, focal_y;
    int tile;              // 16 or 32
    int grid_x, grid_y;    // tile grid
    int* radii;            // output (caller's or internal)
    bool prefiltered;
    int band_row0, band_row1;   // tile-row band [row0,row1); 0,0 = whole frame
    void* zero_ptr; size_t zero_bytes;   // 16 B aligned scratch the kernel clears on the side (the depth sort's histograms and
                                         // look-back words), so that the frame needs no memset operation; may be null / 0
};
void launch_preprocess(const PreprocessArgs& a, const GeomPtrs& g, cudaStream_t s);
void launch_block_scan(const GeomPtrs& g, int P, uint32_t* mailbox_dev, cudaStream_t s);
// depth sort (gsr_depthsort.cu): scan of the bins counted by preprocess, scatter into the bins, exact order inside every bin (+ run offsets)
void launch_bin_scan(const GeomPtrs& g, int P, uint32_t* mailbox_dev, cudaStream_t s);
void launch_bin_scatter(const GeomPtrs& g, int P, cudaStream_t s);
void launch_bin_sort(const GeomPtrs& g, int P, cudaStream_t s);
void launch_runoff_from_blocks(const GeomPtrs& g, int P, cudaStream_t s);      // fallback: run offsets of an onesweep-sorted order
void launch_clamped_rgb(int P, const int* radii, const float* scratch, const unsigned char* clamped, bool colors_from_sh, float* out, cudaStream_t s);
int launch_sh_grad_from_views(int P, int D, int M, const float* means3D, int K, const float* cams, const float* g_all,
                              const uint32_t* meta, int meta_stride, int cap, float* dL_dsh, cudaStream_t s);   // meta == nullptr: dense [K,P,3]
size_t pack_temp_bytes(int P);
int launch_pack_visible_rows(int P, const int* radii, const float* src, int cap, float* rows, uint32_t* mask, uint32_t* prefix,
                             uint32_t* temp, cudaStream_t s);
void launch_mark_visible(int P, const float* means3D, const float* view, const float* proj, unsigned char* present, cudaStream_t s);

struct SortPlan {
    int passes; int shift[kSortMaxPasses]; int bits[kSortMaxPasses];
};
SortPlan make_sort_plan(int end_bit);

SortPlan make_sort_plan_range(int begin_bit, int nbits);
// compaction of the visible Gaussians: keys[k] = depth bits, ids[k] = index, k = rank among the visible ones
void launch_compact_keys(const GeomPtrs& g, int P, uint32_t* keys, uint32_t* ids, cudaStream_t s);
void launch_order_block_sums(int P, const uint32_t* n_dev, const uint32_t* order, const uint2* rects, uint32_t* sums, cudaStream_t s);
void launch_scan_u32(const uint32_t* sums, uint32_t* offsets, int n, cudaStream_t s);
// run-length binning (gsr_binning.cu): emit (Gaussian,row) runs in depth order -> [sort by row] -> expand into per-tile lists
void launch_emit_runs(const GeomPtrs& g, long long V, uint32_t* rkeys, uint32_t* rvals, uint32_t* hist, const SortPlan& plan, cudaStream_t s);
void launch_expand_runs(const BinPtrs& b, const uint32_t* rkeys_sorted, const uint32_t* rvals_sorted, long long R, int tiles_x, int tiles_y,
                        const uint32_t* row_offsets /* digit offsets of a one-pass row sort, or nullptr */, uint2* ranges, uint32_t* n_intersections,
                        const GeomHeader* hdr, uint32_t* mailbox_dev, int prof_ranges, int prof_expand, cudaStream_t s);
void set_fake_sort_error(int v);
// stable LSD sort of n pairs.  Reads (kin,vin), ends in (kout,vout), uses (ktmp,vtmp) as the ping-pong partner.
// Pass order is in -> b1 -> b2 -> b1 ... with b1 = out when the pass count is odd, tmp when it is even, so the last
// pass always writes `out`.  In-place use: kin may alias ktmp when the count is odd and kout when it is even (never
// b1).  vin == nullptr: values = 0..n-1.
int launch_onesweep(const uint64_t* kin, const uint32_t* vin, uint64_t* kout, uint32_t* vout, uint64_t* ktmp, uint32_t* vtmp,
                    long long n, const SortPlan& plan, void* temp, size_t temp_bytes, bool hist_ready, uint32_t* err_flag, cudaStream_t s);
// prezeroed: the caller cleared sort_prezero_bytes() of `temp` (no memset is issued here)
int launch_onesweep32(const uint32_t* kin, const uint32_t* vin, uint32_t* kout, uint32_t* vout, uint32_t* ktmp, uint32_t* vtmp,
                      long long n, const SortPlan& plan, void* temp, size_t temp_bytes, bool hist_ready, uint32_t* err_flag,
                      int prof_slot0, cudaStream_t s, const uint32_t* n_dev = nullptr,   // n_dev: actual count on the device (<= n)
                      bool prezeroed = false);
uint32_t* sort_hist_ptr(void* temp);
// Bytes at the start of `temp` that launch_onesweep* expects to be zero when called with prezeroed = true (histograms,
// tile counters and one look-back region PER PASS); 0 if the per-pass regions do not fit the temp area.
size_t sort_prezero_bytes(long long n, const SortPlan& plan, size_t temp_bytes, bool keys64);
void launch_zero(void* ptr, size_t bytes, cudaStream_t s);   // bytes % 16 == 0, ptr 16 B aligned
void set_sort_variant(int v);
void set_blend_variant(int v);
void set_peer_ctas_per_sm(int v);   // gsr_peer.cu: grid cap of the reduce kernels in CTAs per SM; 0 = default
void set_peer_timeout_ms(int v);    // gsr_peer.cu: barrier watchdog

void launch_amr_levels(const ImgPtrs& im, int num_tiles, cudaStream_t s);
void launch_fovea_levels(const ImgPtrs& im, int num_tiles, int fovea_step, const gsr_gaze* gaze, cudaStream_t s);

struct BlendArgs {
    int W, H;
    int row0, row1;        // tile rows to blend (in units of the kernel's CTA tile: 16 px baseline, 32 px AMR)
    const uint2* ranges; const uint32_t* point_list;
    const float2* means2D; const float4* conic_opacity; const float* colors;
    const float* bg;
    float* final_T; uint32_t* n_contrib; float* out_color;
};
void launch_blend_forward(const BlendArgs& a, cudaStream_t s);
void launch_blend_count(const BlendArgs& a, unsigned long long* totals /* device, [3], caller-zeroed */, cudaStream_t s);

struct AmrBlendArgs {
    BlendArgs b;
    const uint32_t* levels;        // "tile_AMR_levels" argument of the reference kernel
    const uint32_t* levels_last;
    int fovea_step; const float* out_color_precomp; bool interpolate;
};
void launch_blend_amr(const AmrBlendArgs& a, cudaStream_t s);

struct BlendBwdArgs {
    int W, H;
    int row0, row1;
    const uint2* ranges; const uint32_t* point_list;
    const float2* means2D; const float4* conic_opacity; const float* colors;
    const float* bg; const float* final_T; const uint32_t* n_contrib; const float* dL_dpix;
    float* scratch;   // [P][12] packed accumulator: {dmean.x, dmean.y, dconic.x, dconic.y | dconic.w, dopacity, dcol.r, dcol.g | dcol.b, 0,0,0}
};
void launch_blend_backward(const BlendBwdArgs& a, cudaStream_t s);

struct PreBwdArgs {
    int P, D, M;
    const float *means3D, *scales, *rotations, *shs, *cov3D /*precomp or internal*/;
    const float* shs_rest = nullptr; bool raw = false;
    bool cov_from_scales, colors_from_sh;
    float scale_modifier;
    const float *viewmatrix, *projmatrix, *cam_pos;
    float focal_x, focal_y, tan_fovx, tan_fovy;
    const int* radii; const unsigned char* clamped;
    const float* scratch;
    gsr_grads out;
    int first = 0, count = -1;     // count >= 0: only Gaussians [first, first + count), first a multiple of 256
};
void launch_preprocess_backward(const PreBwdArgs& a, cudaStream_t s);

}  // namespace gsr
