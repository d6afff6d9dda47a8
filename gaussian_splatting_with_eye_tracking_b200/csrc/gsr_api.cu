// gsr_api.cu -- extern "C" entry points (include/gsr_b200.h): orchestration only, every kernel lives in
// gsr_preprocess.cu / gsr_binning.cu / gsr_blend.cu.  Replaces CudaRasterizer::Rasterizer::{forward,backward,
// markVisible} (cuda_rasterizer/rasterizer_impl.cu:141-153,198-434; AMR fork :296-694).
#include "gsr_common.cuh"
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>
#include <cstdlib>
#include <nvtx3/nvToolsExt.h>     // header-only NVTX 3: a no-op unless a tool (Nsight Systems / Compute) is attached

namespace gsr {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

void set_error(const char* fmt, ...) {
    va_list ap; va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return GSR_OK;
    set_error("%s: %s", what, cudaGetErrorString(e));
    return GSR_ERR_CUDA;
}

// launch-error check; with debug also synchronises (reference CHECK_CUDA, auxiliary.h:166-173)
int check_last(const char* what, bool debug, cudaStream_t s) {
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess && debug) e = cudaStreamSynchronize(s);
    return check_cuda(e, what);
}

// ---- optional stage timing
struct ProfRec { int slot; cudaEvent_t a, b; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof_recs;
static std::vector<cudaEvent_t> g_prof_pool;
static cudaEvent_t g_prof_open[GSR_PROF_SLOTS];
static std::mutex g_prof_mu;

static cudaEvent_t prof_event() {
    if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
    cudaEvent_t e = nullptr; cudaEventCreate(&e); return e;
}
// NVTX ranges per stage: off by default, on with GSR_NVTX=1 in the environment or
// gsr_set_option(GSR_OPT_NVTX, 1).  Range names are the profile slot names.
static int g_nvtx_on = -1;
static const char* const kSlotNames[GSR_PROF_SLOTS] = {
    "gsr:preprocess", "gsr:depth_histogram", "gsr:emit_runs", "gsr:row_partition", "gsr:tile_ranges", "gsr:blend_fwd", "gsr:amr_levels",
    "gsr:amr_blend", "gsr:blend_bwd", "gsr:preprocess_bwd", "gsr:depth_sort", "gsr:expand_runs", "gsr:sort_pass0", "gsr:bin_scatter",
    "gsr:bin_sort", "gsr:sort_pass3", "gsr:sort_pass4", "gsr:sort_pass5", "gsr:sort_pass6", "gsr:sort_pass7"};
static bool nvtx_on() {
    if (g_nvtx_on < 0) { const char* e = getenv("GSR_NVTX"); g_nvtx_on = (e && e[0] && e[0] != '0') ? 1 : 0; }
    return g_nvtx_on == 1;
}
void set_nvtx(int v) { g_nvtx_on = v ? 1 : 0; }

void prof_begin(int slot, cudaStream_t s) {
    if (slot >= 0 && slot < GSR_PROF_SLOTS && nvtx_on()) nvtxRangePushA(kSlotNames[slot]);
    if (!g_prof_on || slot < 0) return;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    cudaEvent_t e = prof_event(); cudaEventRecord(e, s); g_prof_open[slot] = e;
}
void prof_end(int slot, cudaStream_t s) {
    if (slot >= 0 && slot < GSR_PROF_SLOTS && nvtx_on()) nvtxRangePop();
    if (!g_prof_on || slot < 0) return;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (!g_prof_open[slot]) return;
    cudaEvent_t e = prof_event(); cudaEventRecord(e, s);
    g_prof_recs.push_back({slot, g_prof_open[slot], e}); g_prof_open[slot] = nullptr;
}

static uint32_t* pinned_mailbox(uint32_t** dev_alias = nullptr) {
    // 64 B of mapped pinned host memory per host thread: target of the small D2H header copies, and -- through its
    // device alias -- written directly by block_scan_kernel (no DMA; see build_lists)
    static thread_local uint32_t* box = nullptr;
    static thread_local uint32_t* box_dev = nullptr;
    if (!box) {
        if (cudaHostAlloc((void**)&box, 64, cudaHostAllocMapped) != cudaSuccess) { box = nullptr; cudaGetLastError(); }
        else {
            memset(box, 0, 64);
            if (cudaHostGetDevicePointer((void**)&box_dev, box, 0) != cudaSuccess) { box_dev = nullptr; cudaGetLastError(); }
        }
    }
    if (dev_alias) *dev_alias = box_dev;
    return box;
}

static int read_u32(const uint32_t* dev, uint32_t* host, int count, cudaStream_t s) {
    uint32_t* box = pinned_mailbox();
    if (box) {
        int rc = check_cuda(cudaMemcpyAsync(box, dev, sizeof(uint32_t) * count, cudaMemcpyDeviceToHost, s), "D2H header");
        if (rc) return rc;
        rc = check_cuda(cudaStreamSynchronize(s), "stream sync (header)");
        if (rc) return rc;
        memcpy(host, box, sizeof(uint32_t) * count);
        return GSR_OK;
    }
    int rc = check_cuda(cudaMemcpyAsync(host, dev, sizeof(uint32_t) * count, cudaMemcpyDeviceToHost, s), "D2H header");
    if (rc) return rc;
    return check_cuda(cudaStreamSynchronize(s), "stream sync (header)");
}

// Header of the geometry buffer this host thread built last: lets the follow-up calls of the 5-step fovea protocol and
// gsr_amr_render_prebuilt skip the blocking header read-back (and the pipeline bubble behind it) when they are handed
// that same buffer, which is the only way the reference's Python ever calls them (gaussian_renderer_amr/__init__.py).
struct LastBuild { const void* geom; int P; int N; long long R; long long V; };
static thread_local LastBuild g_last_build = {nullptr, 0, 0, 0, 0};
static int g_force_sort_fallback = 0;      // GSR_OPT_FORCE_SORT_FALLBACK: run the onesweep depth sort even when the buckets fit (tests)

// The radix-sort watchdog (a look-back that never saw its predecessor) cannot stop a frame that is already queued; the
// last binning kernel posts the flag to the mailbox and the NEXT call of this host thread into the library -- the
// backward of the same frame, the next forward, gsr_last_counts -- returns GSR_ERR_SORT instead of working on lists that
// may be corrupt.  With debug != 0 the forward itself reports it.
static int check_sort_mailbox(const char* where) {
    uint32_t* box = pinned_mailbox();
    if (!box) return GSR_OK;
    const uint32_t w = ((volatile uint32_t*)box)[8];
    if ((w & 0x80000001u) == 0x80000001u) {
        box[8] = 0;
        set_error("%s: the radix-sort watchdog fired in the previous frame of this thread (tile lists may be corrupt)", where);
        return GSR_ERR_SORT;
    }
    return GSR_OK;
}

static int geom_header(const void* geom, const GeomHeader* header, int P, int* N, cudaStream_t s) {
    if (geom == g_last_build.geom && P == g_last_build.P) { *N = g_last_build.N; return GSR_OK; }
    uint32_t hdr[4];
    int rc = read_u32(&header->magic, hdr, 4, s);
    if (rc) return rc;
    if (hdr[0] != kGeomMagic || hdr[1] != (uint32_t)P) { set_error("geometry buffer does not belong to this scene (magic %08x, P %u)", hdr[0], hdr[1]); return GSR_ERR_WORKSPACE; }
    *N = (int)hdr[2];
    return GSR_OK;
}

static unsigned higher_msb(unsigned n) {   // reference getHigherMsb (rasterizer_impl.cu:35-50)
    unsigned msb = sizeof(n) * 4, step = msb;
    while (step > 1) { step /= 2; if (n >> msb) msb += step; else msb -= step; }
    if (n >> msb) msb++;
    return msb;
}

static int validate(const gsr_scene* sc, const gsr_camera* cam, bool forward = true) {
    if (!sc || !cam) { set_error("null scene/camera"); return GSR_ERR_INVALID; }
    if (sc->P < 0 || cam->width <= 0 || cam->height <= 0) { set_error("bad sizes P=%d W=%d H=%d", sc->P, cam->width, cam->height); return GSR_ERR_INVALID; }
    if (sc->P > 0) {
        if (!sc->means3D || (forward && !sc->opacities)) { set_error("means3D/opacities are required"); return GSR_ERR_INVALID; }
        if (!sc->colors_precomp && !sc->shs) { set_error("provide SHs or precomputed colors"); return GSR_ERR_INVALID; }
        if (!sc->cov3D_precomp && (!sc->scales || !sc->rotations)) { set_error("provide scale/rotation or precomputed 3D covariance"); return GSR_ERR_INVALID; }
        if (!sc->colors_precomp && (sc->sh_degree < 0 || sc->sh_degree > 3 || sc->sh_coeffs < (sc->sh_degree + 1) * (sc->sh_degree + 1))) {
            set_error("sh_degree=%d needs %d coefficients, got M=%d", sc->sh_degree, (sc->sh_degree + 1) * (sc->sh_degree + 1), sc->sh_coeffs);
            return GSR_ERR_INVALID;
        }
    }
    if (sc->raw_params && sc->P > 0) {
        if (sc->cov3D_precomp || !sc->scales || !sc->rotations) { set_error("raw_params needs scales and rotations (no precomputed covariance)"); return GSR_ERR_INVALID; }
        if (!sc->colors_precomp && (!sc->shs_rest && sc->sh_coeffs > 1)) { set_error("raw_params needs shs (features_dc) and shs_rest (features_rest)"); return GSR_ERR_INVALID; }
    }
    if (!cam->background || !cam->viewmatrix || !cam->projmatrix || !cam->cam_pos) { set_error("camera pointers are required"); return GSR_ERR_INVALID; }
    if (cam->band_row0 < 0 || cam->band_row1 < 0 || (cam->band_row1 > 0 && cam->band_row0 >= cam->band_row1) || (cam->band_row1 == 0 && cam->band_row0 != 0)) {
        set_error("bad tile band [%d,%d)", cam->band_row0, cam->band_row1); return GSR_ERR_INVALID;
    }
    return GSR_OK;
}

// preprocess .. tile ranges; shared by the baseline and AMR forward.  On success *N_out = num_rendered.
static int build_lists(const gsr_scene* sc, const gsr_camera* cam, int mode, gsr_alloc_fn alloc, void* user,
                       int* radii_user, void** geom_out, void** bin_out, void** img_out, int* N_out, bool debug, cudaStream_t s) {
    const int P = sc->P, W = cam->width, H = cam->height;
    const size_t gbytes = gsr_geom_bytes(P), ibytes = gsr_image_bytes(W, H, mode);
    g_last_build = {nullptr, 0, 0, 0, 0};
    { int rc0 = check_sort_mailbox("forward"); if (rc0) return rc0; }
    void* geom = alloc(user, GSR_BUF_GEOM, gbytes);
    void* img = alloc(user, GSR_BUF_IMAGE, ibytes);
    if (!geom || !img) { set_error("allocation callback returned NULL"); return GSR_ERR_ALLOC; }
    GeomPtrs g = geom_ptrs(geom, P);
    ImgPtrs im = img_ptrs(img, W, H, mode);
    const int num_tiles = im.tiles_x * im.tiles_y;
    if (cam->band_row1 > im.tiles_y) { set_error("tile band [%d,%d) exceeds the %d tile rows", cam->band_row0, cam->band_row1, im.tiles_y); return GSR_ERR_INVALID; }
    if (im.tiles_x > kRunFieldMask || im.tiles_y > kRunMaxRows) { set_error("image of %dx%d tiles exceeds the %dx%d tile grid limit", im.tiles_x, im.tiles_y, kRunFieldMask, kRunMaxRows); return GSR_ERR_INVALID; }
    int rc = GSR_OK;
    // Two memset nodes per frame: the header (256 B) and the statistics slots + depth bins (2 - 8 MB, contiguous), all
    // accumulated with atomics by the preprocess kernel.
    if ((rc = check_cuda(cudaMemsetAsync(g.header, 0, sizeof(GeomHeader), s), "memset header"))) return rc;
    if ((rc = check_cuda(cudaMemsetAsync(g.stat_slots, 0, g.frame_zero_bytes, s), "memset bins"))) return rc;
    uint32_t* box_dev = nullptr;
    uint32_t* box = pinned_mailbox(&box_dev);
    if (box) { box[0] = 0; box[6] = 0; }

    PreprocessArgs a;
    a.P = P; a.D = sc->sh_degree; a.M = sc->sh_coeffs;
    a.means3D = sc->means3D; a.scales = sc->scales; a.rotations = sc->rotations; a.opacities = sc->opacities;
    a.shs = sc->shs; a.cov3D_precomp = sc->cov3D_precomp; a.colors_precomp = sc->colors_precomp;
    a.shs_rest = sc->raw_params ? sc->shs_rest : nullptr; a.raw = sc->raw_params != 0;
    a.scale_modifier = sc->scale_modifier;
    a.viewmatrix = cam->viewmatrix; a.projmatrix = cam->projmatrix; a.cam_pos = cam->cam_pos;
    a.W = W; a.H = H; a.tan_fovx = cam->tan_fovx; a.tan_fovy = cam->tan_fovy;
    a.focal_y = H / (2.0f * cam->tan_fovy);          // rasterizer_impl.cu:222-223
    a.focal_x = W / (2.0f * cam->tan_fovx);
    a.tile = im.tile; a.grid_x = im.tiles_x; a.grid_y = im.tiles_y;
    a.radii = radii_user ? radii_user : g.radii;
    a.prefiltered = cam->prefiltered != 0;
    a.band_row0 = cam->band_row0; a.band_row1 = cam->band_row1;
    a.zero_ptr = nullptr; a.zero_bytes = 0;
    { ProfScope ps(GSR_PROF_PREPROCESS, s); launch_preprocess(a, g, s); }
    if ((rc = check_last("preprocess", debug, s))) return rc;
    // histogram of the depth keys + bucket plan (replaces the scan kernel's slot in the profile)
    { ProfScope ps(GSR_PROF_SCAN, s); launch_bin_scan(g, P, box_dev, s); }
    if ((rc = check_last("depth-bin scan", debug, s))) return rc;

    // The one host round-trip of the forward (reference: rasterizer_impl.cu:281): the binning buffer is sized by N.
    // The last CTAs of the two kernels above store the totals straight into mapped host memory and the host waits on an
    // event recorded BEFORE the depth sort, so the host allocates the binning buffer while the GPU is already sorting.
    // No cudaMemcpy: a D2H copy, however small, queues behind whatever the application has in flight on the copy
    // engine (a 25 MB frame read-back delayed this read by up to 0.45 ms).
    static thread_local cudaEvent_t ev_hdr_by_dev[64] = {};
    int dev_now = 0; cudaGetDevice(&dev_now);
    cudaEvent_t* ev_slot = (dev_now >= 0 && dev_now < 64) ? &ev_hdr_by_dev[dev_now] : nullptr;   // events belong to a device
    if (box && box_dev && ev_slot && !*ev_slot) cudaEventCreateWithFlags(ev_slot, cudaEventDisableTiming);
    cudaEvent_t ev_hdr = ev_slot ? *ev_slot : nullptr;
    const bool mailed = box && box_dev && ev_hdr;
    if (mailed && (rc = check_cuda(cudaEventRecord(ev_hdr, s), "event record"))) return rc;

    // (1) order the visible Gaussians by (depth bits, index): counting sort into fine depth bins, exact order inside each
    {
        ProfScope ps(GSR_PROF_DEPTH_SORT, s);
        { ProfScope p1(GSR_PROF_SORT_PASS0 + 1, s); launch_bin_scatter(g, P, s); }
        { ProfScope p2(GSR_PROF_SORT_PASS0 + 2, s); launch_bin_sort(g, P, s); }
    }
    if ((rc = check_last("depth sort", debug, s))) return rc;

    uint32_t hdr[8];
    if (mailed) {
        if ((rc = check_cuda(cudaEventSynchronize(ev_hdr), "event sync (header)"))) return rc;
        memcpy(hdr, box, 32);
    } else {
        if ((rc = read_u32(&g.header->magic, hdr, 6, s))) return rc;
        uint32_t mode = 0;
        if ((rc = read_u32(&g.header->sort_mode, &mode, 1, s))) return rc;
        hdr[6] = mode >= 2 ? 2u : mode; hdr[7] = 0;
    }
    if (hdr[0] != kGeomMagic || hdr[1] != (uint32_t)P || hdr[6] == 0) { set_error("frame totals did not arrive (magic %08x, P %u, mode %u)", hdr[0], hdr[1], hdr[6]); return GSR_ERR_WORKSPACE; }
    if (hdr[3] & 0x100u) { set_error("a Gaussian was culled although prefiltered=true (reference would __trap)"); return GSR_ERR_INVALID; }
    const long long N = hdr[2], R = hdr[4], V = hdr[5];
    if (hdr[6] == 2u || g_force_sort_fallback) {
        // A depth bin holds more records than one warp orders (or the test hook asks for it): the round-1 chain instead --
        // ordered compaction of the visible Gaussians, 4 stable onesweep passes over the 32 depth bits, block sums + scan --
        // then the run offsets per sorted position.  The bin kernels queued above have left at once (sort_mode != 1).
        ProfScope ps(GSR_PROF_DEPTH_SORT, s);
        const SortPlan dplan = make_sort_plan(32);
        if (dplan.passes & 1) { set_error("depth sort plan must have an even pass count"); return GSR_ERR_INVALID; }
        launch_scan_u32(g.block_sums_v, g.block_offsets, g.nblocks, s);
        launch_compact_keys(g, P, g.dkeys_a, g.order, s);
        rc = launch_onesweep32(g.dkeys_a, g.order, g.dkeys_a, g.order, g.dkeys_b, g.order_tmp, P, dplan, g.sort_temp, g.sort_temp_bytes,
                               /*hist_ready=*/false, &g.header->sort_error, -1, s, &g.header->num_visible, /*prezeroed=*/false);
        if (rc) return rc;
        launch_order_block_sums(P, &g.header->num_visible, g.order, g.rects, g.oblock_sums, s);
        launch_scan_u32(g.oblock_sums, g.oblock_offsets, g.nblocks, s);
        launch_runoff_from_blocks(g, P, s);
        if ((rc = check_last("depth sort (fallback)", debug, s))) return rc;
    }
    void* bin = alloc(user, GSR_BUF_BINNING, gsr_binning_bytes(N, R, W, H, mode));
    if (!bin) { set_error("allocation callback returned NULL (binning)"); return GSR_ERR_ALLOC; }
    BinPtrs b = bin_ptrs(bin, N, R, im.tiles_x, im.tiles_y);

    const uint32_t* rk_sorted = b.rkeyB; const uint32_t* rv_sorted = b.rvalB;
    const uint32_t* row_offsets = nullptr;
    if (R > 0) {
        // (2) emit one run per (Gaussian, tile row) in depth order, (3) stable partition of the runs by row
        const SortPlan plan = make_sort_plan_range(kRunRowShift, (int)higher_msb((unsigned)im.tiles_y));
        uint32_t* hist = sort_hist_ptr(b.sort_temp);
        const size_t rzero = sort_prezero_bytes(R, plan, b.sort_temp_bytes, false);
        if (rzero) launch_zero(hist, rzero, s);       // one clearing kernel for histogram + counters + look-back words
        else if ((rc = check_cuda(cudaMemsetAsync(hist, 0, sizeof(uint32_t) * kSortMaxPasses * kSortMaxRadix, s), "memset hist"))) return rc;
        { ProfScope ps(GSR_PROF_DUPLICATE, s);
          launch_emit_runs(g, V, b.rkeyA, b.rvalA, hist, plan, s); }
        if ((rc = check_last("emit runs", debug, s))) return rc;
        {
            ProfScope ps(GSR_PROF_SORT, s);
            // one pass (<= 512 tile rows): A -> B.  Two passes: A -> C -> B, the value partner being the not-yet-written
            // point list (N >= R words).
            if (plan.passes > 2 || (plan.passes == 2 && im.tiles_y <= kSortMaxRadix)) { set_error("unexpected row-sort plan"); return GSR_ERR_INVALID; }
            rc = launch_onesweep32(b.rkeyA, b.rvalA, b.rkeyB, b.rvalB, (plan.passes & 1) ? b.rkeyA : b.rkeyC,
                                   (plan.passes & 1) ? b.rvalA : b.vals, R, plan, b.sort_temp, b.sort_temp_bytes,
                                   /*hist_ready=*/true, &g.header->sort_error, GSR_PROF_SORT_PASS0, s, nullptr, /*prezeroed=*/rzero != 0);
        }
        if (rc) return rc;
        if ((rc = check_last("radix sort", debug, s))) return rc;
        if (plan.passes == 1 && plan.shift[0] == kRunRowShift) row_offsets = hist;   // exclusive digit offsets == first run of every row
    }
    // (4) expand the row-sorted runs into the per-tile lists; also yields tile ranges and AMR intersection counts
    launch_expand_runs(b, rk_sorted, rv_sorted, R, im.tiles_x, im.tiles_y, row_offsets, im.ranges, im.n_intersections, g.header, box_dev,
                       GSR_PROF_RANGES, GSR_PROF_EXPAND, s);
    if ((rc = check_last("expand runs", debug, s))) return rc;
    if (debug && (rc = check_sort_mailbox("forward (debug)"))) return rc;      // debug synchronised: the word has arrived
    *geom_out = geom; *bin_out = bin; *img_out = img; *N_out = (int)N;
    g_last_build = {geom, P, (int)N, R, (long long)hdr[5]};
    return GSR_OK;
}

}  // namespace gsr

using namespace gsr;

extern "C" {

size_t gsr_geom_bytes(int P) { return GeomLayout(P > 0 ? P : 0).total + kAlign; }
size_t gsr_binning_bytes(long long N, long long R, int W, int H, int mode) {
    ImgLayout il(W > 0 ? W : 1, H > 0 ? H : 1, mode);
    return BinLayout(N, R, il.tiles_x, il.tiles_y).total + kAlign;
}
size_t gsr_image_bytes(int W, int H, int mode) { return ImgLayout(W, H, mode).total + kAlign; }
size_t gsr_sort_temp_bytes(long long n) { return sort_temp_bytes(n) + kAlign; }
unsigned gsr_higher_msb(unsigned n) { return higher_msb(n); }
unsigned long long gsr_launch_count(void) { return g_launches.load(); }

void gsr_profile_enable(int on) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    g_prof_on = on != 0;
    if (g_prof_on) while (g_prof_pool.size() < 2048) { cudaEvent_t e = nullptr; if (cudaEventCreate(&e) != cudaSuccess) break; g_prof_pool.push_back(e); }
}

int gsr_profile_read(float* ms, int* counts) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    int rc = GSR_OK;
    for (const ProfRec& r : g_prof_recs) {
        float t = 0.f;
        cudaError_t e = cudaEventSynchronize(r.b);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&t, r.a, r.b);
        if (e != cudaSuccess) { rc = check_cuda(e, "profile read"); }
        else if (r.slot >= 0 && r.slot < GSR_PROF_SLOTS) { if (ms) ms[r.slot] += t; if (counts) counts[r.slot] += 1; }
        g_prof_pool.push_back(r.a); g_prof_pool.push_back(r.b);
    }
    g_prof_recs.clear();
    return rc;
}
const char* gsr_last_error(void) { return g_err; }
int gsr_set_option(int option, int value) {
    if (option == GSR_OPT_SORT_VARIANT && value >= 0 && value <= 5) { set_sort_variant(value); return GSR_OK; }
    if (option == GSR_OPT_BLEND_VARIANT && value >= 0 && value <= 2) { set_blend_variant(value); return GSR_OK; }
    if (option == GSR_OPT_PEER_TIMEOUT_MS && value >= 0) { set_peer_timeout_ms(value); return GSR_OK; }
    if (option == GSR_OPT_NVTX && (value == 0 || value == 1)) { set_nvtx(value); return GSR_OK; }
    if (option == GSR_OPT_FAKE_SORT_ERROR && (value == 0 || value == 1)) { set_fake_sort_error(value); return GSR_OK; }
    if (option == GSR_OPT_FORCE_SORT_FALLBACK && (value == 0 || value == 1)) { g_force_sort_fallback = value; return GSR_OK; }
    if (option == GSR_OPT_PEER_CTAS_PER_SM && value >= 0 && (value & 0xff) <= 32 && value < 0x300) { set_peer_ctas_per_sm(value); return GSR_OK; }
    set_error("unknown option %d / value %d", option, value);
    return GSR_ERR_INVALID;
}
const char* gsr_version(void) { return "gsr_b200 0.1 (sm_100a)"; }

int gsr_forward(const gsr_scene* sc, const gsr_camera* cam, gsr_alloc_fn alloc, void* user,
                float* out_color, int* radii, int* num_rendered, int debug, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    int rc = validate(sc, cam);
    if (rc) return rc;
    if (!alloc || !out_color || !num_rendered) { set_error("alloc/out_color/num_rendered are required"); return GSR_ERR_INVALID; }
    *num_rendered = 0;
    const int W = cam->width, H = cam->height;
    if (sc->P == 0) {   // reference: nothing is launched, out_color keeps its zero fill (rasterize_points.cu:68,80)
        return check_cuda(cudaMemsetAsync(out_color, 0, sizeof(float) * 3 * (size_t)W * H, s), "memset out_color");
    }
    void *geom, *bin, *img; int N = 0;
    rc = build_lists(sc, cam, GSR_MODE_BASELINE, alloc, user, radii, &geom, &bin, &img, &N, debug != 0, s);
    if (rc) return rc;
    GeomPtrs g = geom_ptrs(geom, sc->P);
    ImgPtrs im = img_ptrs(img, W, H, GSR_MODE_BASELINE);
    BlendArgs ba;
    ba.W = W; ba.H = H; ba.row0 = cam->band_row0; ba.row1 = cam->band_row1; ba.ranges = im.ranges; ba.point_list = bin_point_list(bin); ba.means2D = g.means2D;
    ba.conic_opacity = g.conic_opacity; ba.colors = sc->colors_precomp ? sc->colors_precomp : g.rgb;
    ba.bg = cam->background; ba.final_T = im.accum_alpha; ba.n_contrib = im.n_contrib; ba.out_color = out_color;
    { ProfScope ps(GSR_PROF_BLEND, s); launch_blend_forward(ba, s); }
    if ((rc = check_last("blend forward", debug != 0, s))) return rc;
    *num_rendered = N;
    return GSR_OK;
}

int gsr_backward(const gsr_scene* sc, const gsr_camera* cam, const int* radii, const void* geom_buffer,
                 const void* binning_buffer, const void* image_buffer, int num_rendered, const float* dL_dpix,
                 const gsr_grads* gr, int debug, void* stream) {
    return gsr_backward_phase(sc, cam, radii, geom_buffer, binning_buffer, image_buffer, num_rendered, dL_dpix, gr,
                              GSR_BWD_ALL, debug, stream);
}

static int backward_impl(const gsr_scene* sc, const gsr_camera* cam, const int* radii, const void* geom_buffer,
                         const void* binning_buffer, const void* image_buffer, int num_rendered, const float* dL_dpix,
                         const gsr_grads* gr, int phase, int first, int count, int debug, void* stream);

int gsr_backward_phase(const gsr_scene* sc, const gsr_camera* cam, const int* radii, const void* geom_buffer,
                       const void* binning_buffer, const void* image_buffer, int num_rendered, const float* dL_dpix,
                       const gsr_grads* gr, int phase, int debug, void* stream) {
    return backward_impl(sc, cam, radii, geom_buffer, binning_buffer, image_buffer, num_rendered, dL_dpix, gr, phase, 0, -1, debug, stream);
}

int gsr_backward_preprocess_range(const gsr_scene* sc, const gsr_camera* cam, const int* radii, const void* geom_buffer,
                                  const void* binning_buffer, const void* image_buffer, int num_rendered,
                                  const gsr_grads* gr, int first, int count, int debug, void* stream) {
    if (first < 0 || count < 0 || (first % 256) != 0) { set_error("backward_preprocess_range: first must be a non-negative multiple of 256, count >= 0"); return GSR_ERR_INVALID; }
    return backward_impl(sc, cam, radii, geom_buffer, binning_buffer, image_buffer, num_rendered, nullptr, gr, GSR_BWD_PREPROCESS, first, count, debug, stream);
}

static int backward_impl(const gsr_scene* sc, const gsr_camera* cam, const int* radii, const void* geom_buffer,
                         const void* binning_buffer, const void* image_buffer, int num_rendered, const float* dL_dpix,
                         const gsr_grads* gr, int phase, int first, int count, int debug, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    if (phase != GSR_BWD_ALL && phase != GSR_BWD_BLEND && phase != GSR_BWD_PREPROCESS) { set_error("bad backward phase %d", phase); return GSR_ERR_INVALID; }
    (void)num_rendered;
    int rc = validate(sc, cam, /*forward=*/false);
    if (rc) return rc;
    if (!gr || (!dL_dpix && phase != GSR_BWD_PREPROCESS)) { set_error("grads/dL_dpix are required"); return GSR_ERR_INVALID; }
    const int P = sc->P, W = cam->width, H = cam->height;
    if (P == 0) return GSR_OK;
    if ((rc = check_sort_mailbox("backward"))) return rc;
    if (!geom_buffer || !binning_buffer || !image_buffer) { set_error("work buffers are required"); return GSR_ERR_WORKSPACE; }
    if (!gr->dL_dmeans2D || !gr->dL_dcolors || !gr->dL_dopacity || !gr->dL_dmeans3D || !gr->dL_dcov3D || !gr->dL_dscales ||
        !gr->dL_drotations || !gr->scratch || (phase != GSR_BWD_BLEND && sc->sh_coeffs > 0 && sc->shs && !gr->dL_dsh && !gr->dL_dcolors_clamped)) {
        set_error("gradient output pointers are required"); return GSR_ERR_INVALID;
    }
    if (((size_t)gr->scratch & 15) != 0) { set_error("scratch must be 16-byte aligned"); return GSR_ERR_INVALID; }
    if (sc->raw_params && sc->shs_rest && phase != GSR_BWD_BLEND &&
        (sc->sh_coeffs > 16 || (sc->sh_coeffs % 4) != 0 || (gr->dL_dsh && ((size_t)gr->dL_dsh & 15) != 0))) {
        set_error("raw-parameter backward needs M in {4, 8, 12, 16} and a 16-byte aligned dL_dsh (M = %d)", sc->sh_coeffs); return GSR_ERR_INVALID;
    }
    if (gr->push_bases && (gr->push_world < 1 || gr->push_rank < 0 || gr->push_rank >= gr->push_world || gr->push_rows_per_rank <= 0 ||
                           (gr->push_rows_per_rank % 256) != 0 || (gr->push_offset & 127) != 0 || gr->push_epoch == 0 ||
                           (long long)gr->push_rows_per_rank * gr->push_world < (long long)P)) {
        set_error("bad push_* fields (world %d, rank %d, rows per rank %d for P = %d)", gr->push_world, gr->push_rank, gr->push_rows_per_rank, P);
        return GSR_ERR_INVALID;
    }
    GeomPtrs g = geom_ptrs(geom_buffer, P);
    ImgPtrs im = img_ptrs(image_buffer, W, H, GSR_MODE_BASELINE);
    if (!radii) radii = g.radii;

    const bool colors_from_sh = (sc->shs != nullptr) && (sc->colors_precomp == nullptr);          // backward.cu:390 `if (shs)`
    if (phase != GSR_BWD_PREPROCESS) {
    if ((rc = check_cuda(cudaMemsetAsync(gr->scratch, 0, sizeof(float) * 12 * (size_t)P, s), "memset scratch"))) return rc;
    BlendBwdArgs bb;
    bb.W = W; bb.H = H; bb.row0 = cam->band_row0; bb.row1 = cam->band_row1; bb.ranges = im.ranges; bb.point_list = bin_point_list(binning_buffer); bb.means2D = g.means2D;
    bb.conic_opacity = g.conic_opacity; bb.colors = sc->colors_precomp ? sc->colors_precomp : g.rgb;
    bb.bg = cam->background; bb.final_T = im.accum_alpha; bb.n_contrib = im.n_contrib; bb.dL_dpix = dL_dpix;
    bb.scratch = gr->scratch;
    { ProfScope ps(GSR_PROF_BWD_BLEND, s); launch_blend_backward(bb, s); }
    if ((rc = check_last("blend backward", debug != 0, s))) return rc;
    if (phase == GSR_BWD_BLEND && gr->dL_dcolors_clamped) {
        // the clamped colour gradient is final once the blend stage is done: emit it now so that a view-parallel
        // caller can start exchanging it while the preprocess stage still runs
        launch_clamped_rgb(P, radii, gr->scratch, g.clamped, colors_from_sh, gr->dL_dcolors_clamped, s);
        if ((rc = check_last("clamped rgb", debug != 0, s))) return rc;
    }
    }
    if (phase == GSR_BWD_BLEND) return GSR_OK;

    PreBwdArgs pb;
    pb.P = P; pb.D = sc->sh_degree; pb.M = sc->sh_coeffs;
    pb.means3D = sc->means3D; pb.scales = sc->scales; pb.rotations = sc->rotations; pb.shs = sc->shs;
    pb.shs_rest = sc->raw_params ? sc->shs_rest : nullptr; pb.raw = sc->raw_params != 0;
    pb.cov3D = sc->cov3D_precomp ? sc->cov3D_precomp : g.cov3D;
    pb.cov_from_scales = (sc->cov3D_precomp == nullptr) && sc->scales && sc->rotations;   // backward.cu:394 `if (scales)`
    pb.colors_from_sh = colors_from_sh;
    pb.scale_modifier = sc->scale_modifier;
    pb.viewmatrix = cam->viewmatrix; pb.projmatrix = cam->projmatrix; pb.cam_pos = cam->cam_pos;
    pb.focal_y = H / (2.0f * cam->tan_fovy); pb.focal_x = W / (2.0f * cam->tan_fovx);
    pb.tan_fovx = cam->tan_fovx; pb.tan_fovy = cam->tan_fovy;
    pb.radii = radii; pb.clamped = g.clamped; pb.scratch = gr->scratch; pb.out = *gr;
    if (sc->sh_coeffs == 0) pb.out.dL_dsh = nullptr;
    if (phase == GSR_BWD_PREPROCESS) pb.out.dL_dcolors_clamped = nullptr;      // already emitted by the blend phase
    pb.first = first; pb.count = count;
    { ProfScope ps(GSR_PROF_BWD_PREPROCESS, s); launch_preprocess_backward(pb, s); }
    return check_last("preprocess backward", debug != 0, s);
}

int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     unsigned char* present, void* stream) {
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) { set_error("mark_visible: bad arguments"); return GSR_ERR_INVALID; }
    launch_mark_visible(P, means3D, viewmatrix, projmatrix, present, (cudaStream_t)stream);
    return check_last("mark_visible", false, (cudaStream_t)stream);
}

// The AMR kernels write only the lattice pixels they render: everything else of the frame (or of this call's tile band) must read 0
// (reference: out_color = torch::full(0), A/rasterize_points.cu:116).  With a tile band only the band's pixel rows are cleared --
// rows outside it are not this call's to write (a rank of a banded 4K frame would otherwise clear 100 MB to use 12).
static int zero_out_color(float* out_color, int W, int H, int band_row0, int band_row1, cudaStream_t s) {
    if (band_row1 <= 0) return check_cuda(cudaMemsetAsync(out_color, 0, sizeof(float) * 3 * (size_t)W * H, s), "memset out_color");
    const int y0 = min(H, band_row0 * 32), y1 = min(H, band_row1 * 32);
    if (y1 <= y0) return GSR_OK;
    return check_cuda(cudaMemset2DAsync(out_color + (size_t)y0 * W, sizeof(float) * (size_t)W * H, 0, sizeof(float) * (size_t)W * (y1 - y0), 3, s),
                      "memset out_color (band)");
}

int gsr_amr_forward(const gsr_scene* sc, const gsr_camera* cam, int fovea_step, const float* out_color_precomp,
                    void* geom_precomp, void* binning_precomp, void* image_precomp, int interpolate_image,
                    const gsr_gaze* gaze, gsr_alloc_fn alloc, void* user, float* out_color, int* radii,
                    int* num_rendered, int debug, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    int rc = validate(sc, cam);
    if (rc) return rc;
    if (!out_color || !num_rendered) { set_error("out_color/num_rendered are required"); return GSR_ERR_INVALID; }
    *num_rendered = 0;
    const int P = sc->P, W = cam->width, H = cam->height;
    // the AMR kernels write only the lattice pixels they render: everything else must read 0
    // (reference: out_color = torch::full(0), A/rasterize_points.cu:116)
    if ((rc = zero_out_color(out_color, W, H, cam->band_row0, cam->band_row1, s))) return rc;
    if (P == 0) return GSR_OK;

    void *geom = geom_precomp, *bin = binning_precomp, *img = image_precomp;
    int N = 0;
    if (fovea_step >= 1) {
        // progressive round k: everything was built by step 0 (A/rasterizer_impl.cu:334-462)
        if (!geom || !bin || !img) { set_error("fovea_step >= 1 needs the buffers built by step 0"); return GSR_ERR_WORKSPACE; }
        GeomPtrs g0 = geom_ptrs(geom, P);
        if ((rc = geom_header(geom, g0.header, P, &N, s))) return rc;
    } else {
        if (!alloc) { set_error("alloc callback is required"); return GSR_ERR_INVALID; }
        rc = build_lists(sc, cam, GSR_MODE_AMR, alloc, user, radii, &geom, &bin, &img, &N, debug != 0, s);
        if (rc) return rc;
    }
    GeomPtrs g = geom_ptrs(geom, P);
    ImgPtrs im = img_ptrs(img, W, H, GSR_MODE_AMR);
    const int num_tiles = im.tiles_x * im.tiles_y;
    *num_rendered = N;

    if (fovea_step < 1) {
        { ProfScope ps(GSR_PROF_AMR_LEVELS, s); launch_amr_levels(im, num_tiles, s); }
        if ((rc = check_last("amr levels", debug != 0, s))) return rc;
        if (fovea_step == 0) return GSR_OK;             // step 0: build only (A/rasterizer_impl.cu:650-653)
    }
    launch_fovea_levels(im, num_tiles, fovea_step, gaze, s);
    if ((rc = check_last("fovea levels", debug != 0, s))) return rc;

    AmrBlendArgs ab;
    ab.b.W = W; ab.b.H = H; ab.b.row0 = cam->band_row0; ab.b.row1 = cam->band_row1; ab.b.ranges = im.ranges; ab.b.point_list = bin_point_list(bin); ab.b.means2D = g.means2D;
    ab.b.conic_opacity = g.conic_opacity; ab.b.colors = sc->colors_precomp ? sc->colors_precomp : g.rgb;
    ab.b.bg = cam->background; ab.b.final_T = im.accum_alpha; ab.b.n_contrib = im.n_contrib; ab.b.out_color = out_color;
    // reference: steps >= 1 render with tile_AMR_levels_current, single pass with tile_AMR_levels (:401 vs :667)
    ab.levels = (fovea_step >= 1) ? im.levels_current : im.levels;
    ab.levels_last = im.levels_last;
    ab.fovea_step = fovea_step; ab.out_color_precomp = out_color_precomp; ab.interpolate = interpolate_image != 0;
    { ProfScope ps(GSR_PROF_AMR_BLEND, s); launch_blend_amr(ab, s); }
    return check_last("blend amr", debug != 0, s);
}

int gsr_amr_relevel(void* image_buffer, int W, int H, void* stream) {
    if (!image_buffer || W <= 0 || H <= 0) { set_error("amr_relevel: bad arguments"); return GSR_ERR_INVALID; }
    ImgPtrs im = img_ptrs(image_buffer, W, H, GSR_MODE_AMR);
    launch_amr_levels(im, im.tiles_x * im.tiles_y, (cudaStream_t)stream);
    return check_last("amr relevel", false, (cudaStream_t)stream);
}

int gsr_amr_render_prebuilt(const gsr_scene* sc, const gsr_camera* cam, const void* geom_buffer, const void* binning_buffer,
                            void* image_buffer, int interpolate_image, const gsr_gaze* gaze, float* out_color, int debug, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    int rc = validate(sc, cam);
    if (rc) return rc;
    if (!out_color) { set_error("out_color is required"); return GSR_ERR_INVALID; }
    const int P = sc->P, W = cam->width, H = cam->height;
    if ((rc = zero_out_color(out_color, W, H, cam->band_row0, cam->band_row1, s))) return rc;
    if (P == 0) return GSR_OK;
    if (!geom_buffer || !binning_buffer || !image_buffer) { set_error("prebuilt work buffers are required"); return GSR_ERR_WORKSPACE; }
    GeomPtrs g = geom_ptrs(geom_buffer, P);
    int N_unused = 0;
    if ((rc = geom_header(geom_buffer, g.header, P, &N_unused, s))) return rc;
    ImgPtrs im = img_ptrs(image_buffer, W, H, GSR_MODE_AMR);
    if (cam->band_row1 > im.tiles_y) { set_error("tile band exceeds the tile rows"); return GSR_ERR_INVALID; }
    launch_fovea_levels(im, im.tiles_x * im.tiles_y, -1, gaze, s);
    if ((rc = check_last("fovea levels", debug != 0, s))) return rc;
    AmrBlendArgs ab;
    ab.b.W = W; ab.b.H = H; ab.b.row0 = cam->band_row0; ab.b.row1 = cam->band_row1; ab.b.ranges = im.ranges; ab.b.point_list = bin_point_list(binning_buffer);
    ab.b.means2D = g.means2D; ab.b.conic_opacity = g.conic_opacity; ab.b.colors = sc->colors_precomp ? sc->colors_precomp : g.rgb;
    ab.b.bg = cam->background; ab.b.final_T = im.accum_alpha; ab.b.n_contrib = im.n_contrib; ab.b.out_color = out_color;
    ab.levels = im.levels; ab.levels_last = im.levels_last; ab.fovea_step = -1; ab.out_color_precomp = nullptr;
    ab.interpolate = interpolate_image != 0;
    { ProfScope ps(GSR_PROF_AMR_BLEND, s); launch_blend_amr(ab, s); }
    return check_last("blend amr (prebuilt)", debug != 0, s);
}

int gsr_sh_grad_from_views(int P, int sh_degree, int sh_coeffs, const float* means3D, int num_views,
                           const float* cam_positions, const float* dL_dcolors_clamped, float* dL_dsh, void* stream) {
    if (P < 0 || num_views < 0 || sh_degree < 0 || sh_degree > 3) { set_error("sh_grad_from_views: bad sizes"); return GSR_ERR_INVALID; }
    if (P > 0 && (!means3D || !dL_dsh || (num_views > 0 && (!cam_positions || !dL_dcolors_clamped)))) { set_error("sh_grad_from_views: null pointer"); return GSR_ERR_INVALID; }
    int rc = launch_sh_grad_from_views(P, sh_degree, sh_coeffs, means3D, num_views, cam_positions, dL_dcolors_clamped, nullptr, 0, 0, dL_dsh, (cudaStream_t)stream);
    if (rc) return rc;
    return check_last("sh_grad_from_views", false, (cudaStream_t)stream);
}

size_t gsr_pack_meta_words(int P) { return P < 0 ? 0 : 2 * (size_t)((P + 31) / 32); }
size_t gsr_pack_temp_bytes(int P) { return P < 0 ? 0 : pack_temp_bytes(P); }

int gsr_pack_visible_rows(int P, const int* radii, const float* rows_in, int cap, float* rows_out, uint32_t* meta,
                          void* temp, size_t temp_bytes, void* stream) {
    if (P < 0 || cap < 0) { set_error("pack_visible_rows: bad sizes"); return GSR_ERR_INVALID; }
    if (P == 0) return GSR_OK;
    if (!radii || !rows_in || !meta || !temp || (cap > 0 && !rows_out)) { set_error("pack_visible_rows: null pointer"); return GSR_ERR_INVALID; }
    if (temp_bytes < pack_temp_bytes(P)) { set_error("pack_visible_rows: temp too small (%zu < %zu)", temp_bytes, pack_temp_bytes(P)); return GSR_ERR_INVALID; }
    const int W32 = (P + 31) / 32;
    int rc = launch_pack_visible_rows(P, radii, rows_in, cap, rows_out, meta, meta + W32, (uint32_t*)temp, (cudaStream_t)stream);
    if (rc) return rc;
    return check_last("pack_visible_rows", false, (cudaStream_t)stream);
}

int gsr_sh_grad_from_packed_views(int P, int sh_degree, int sh_coeffs, const float* means3D, int num_views, const float* cam_positions,
                                  const uint32_t* meta, size_t meta_stride_words, const float* rows, int cap, float* dL_dsh, void* stream) {
    if (P < 0 || num_views < 0 || cap < 0 || sh_degree < 0 || sh_degree > 3) { set_error("sh_grad_from_packed_views: bad sizes"); return GSR_ERR_INVALID; }
    if (P > 0 && (!means3D || !dL_dsh || (num_views > 0 && (!cam_positions || !meta || (cap > 0 && !rows))))) { set_error("sh_grad_from_packed_views: null pointer"); return GSR_ERR_INVALID; }
    if (meta_stride_words < gsr_pack_meta_words(P) || meta_stride_words > 0x7fffffffull) { set_error("sh_grad_from_packed_views: meta stride too small"); return GSR_ERR_INVALID; }
    if (num_views == 0) return gsr_sh_grad_from_views(P, sh_degree, sh_coeffs, means3D, 0, nullptr, nullptr, dL_dsh, stream);
    int rc = launch_sh_grad_from_views(P, sh_degree, sh_coeffs, means3D, num_views, cam_positions, rows, meta, (int)meta_stride_words, cap, dL_dsh, (cudaStream_t)stream);
    if (rc) return rc;
    return check_last("sh_grad_from_packed_views", false, (cudaStream_t)stream);
}

int gsr_last_counts(const void* geom_buffer, int P, long long* num_rendered, long long* num_runs, long long* num_visible) {
    if (!geom_buffer || geom_buffer != g_last_build.geom || P != g_last_build.P) {
        set_error("last_counts: not the geometry buffer this thread built last"); return GSR_ERR_INVALID;
    }
    { int rc0 = check_sort_mailbox("last_counts"); if (rc0) return rc0; }
    if (num_rendered) *num_rendered = g_last_build.N;
    if (num_runs) *num_runs = g_last_build.R;
    if (num_visible) *num_visible = g_last_build.V;
    return GSR_OK;
}

int gsr_blend_stats(int P, int W, int H, const void* geom_buffer, const void* binning_buffer, const void* image_buffer,
                    unsigned long long* totals_dev, void* stream) {
    if (P <= 0 || W <= 0 || H <= 0 || !geom_buffer || !binning_buffer || !image_buffer || !totals_dev) { set_error("blend_stats: bad arguments"); return GSR_ERR_INVALID; }
    cudaStream_t s = (cudaStream_t)stream;
    GeomPtrs g = geom_ptrs(geom_buffer, P);
    ImgPtrs im = img_ptrs(image_buffer, W, H, GSR_MODE_BASELINE);
    int rc = check_cuda(cudaMemsetAsync(totals_dev, 0, 3 * sizeof(unsigned long long), s), "memset totals");
    if (rc) return rc;
    BlendArgs ba;
    ba.W = W; ba.H = H; ba.row0 = 0; ba.row1 = 0; ba.ranges = im.ranges; ba.point_list = bin_point_list(binning_buffer); ba.means2D = g.means2D;
    ba.conic_opacity = g.conic_opacity; ba.colors = nullptr; ba.bg = nullptr; ba.final_T = nullptr; ba.n_contrib = nullptr; ba.out_color = nullptr;
    launch_blend_count(ba, totals_dev, s);
    return check_last("blend_stats", false, s);
}

int gsr_view_geom(const void* geom_buffer, int P, gsr_geom_view* out) {
    if (!geom_buffer || !out || P < 0) { set_error("view_geom: bad arguments"); return GSR_ERR_INVALID; }
    GeomPtrs g = geom_ptrs(geom_buffer, P);
    out->P = P; out->depths = g.depths; out->clamped = g.clamped; out->radii = g.radii;
    out->means2D = (const float*)g.means2D; out->cov3D = g.cov3D; out->conic_opacity = (const float*)g.conic_opacity;
    out->rgb = g.rgb; out->tiles_touched = g.tiles_touched; out->block_offsets = g.block_offsets; out->depth_order = g.order;
    out->num_rendered = &g.header->num_rendered;
    return GSR_OK;
}

int gsr_view_binning(const void* binning_buffer, long long N, gsr_binning_view* out) {
    if (!binning_buffer || !out || N < 0) { set_error("view_binning: bad arguments"); return GSR_ERR_INVALID; }
    out->N = N; out->keys_unsorted = nullptr; out->values_unsorted = nullptr;   // pairs are never materialised unsorted
    out->tiles_sorted = nullptr;              // nor are their tile ids: they follow from the tile ranges
    out->point_list = bin_point_list(binning_buffer);
    return GSR_OK;
}

int gsr_view_image(const void* image_buffer, int W, int H, int mode, gsr_image_view* out) {
    if (!image_buffer || !out || W <= 0 || H <= 0) { set_error("view_image: bad arguments"); return GSR_ERR_INVALID; }
    ImgPtrs im = img_ptrs(image_buffer, W, H, mode);
    out->width = W; out->height = H; out->mode = mode; out->tiles_x = im.tiles_x; out->tiles_y = im.tiles_y;
    out->accum_alpha = im.accum_alpha; out->n_contrib = im.n_contrib; out->ranges = (const uint32_t*)im.ranges;
    out->n_intersections = im.n_intersections; out->levels = im.levels; out->levels_last = im.levels_last;
    out->levels_current = im.levels_current; out->thresholds = im.thresholds;
    return GSR_OK;
}

int gsr_sort_pairs(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                   uint64_t* keys_tmp, uint32_t* vals_tmp, long long n, int end_bit, void* temp, size_t temp_bytes, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    if (n < 0 || end_bit < 1 || end_bit > 64) { set_error("sort_pairs: bad n/end_bit"); return GSR_ERR_INVALID; }
    if (n == 0) return GSR_OK;
    if (!keys_in || !keys_out || !vals_in || !vals_out || !keys_tmp || !vals_tmp || !temp) { set_error("sort_pairs: null buffer"); return GSR_ERR_INVALID; }
    const SortPlan plan = make_sort_plan(end_bit);
    // the watchdog flag lives in the last word of the counters block
    uint32_t* err = (uint32_t*)(aligned_base(temp) + align_up(sizeof(uint32_t) * kSortMaxPasses * kSortMaxRadix)) + 32;
    int rc = check_cuda(cudaMemsetAsync(err, 0, sizeof(uint32_t), s), "memset sort flag");
    if (rc) return rc;
    rc = launch_onesweep(keys_in, vals_in, keys_out, vals_out, keys_tmp, vals_tmp, n, plan, temp, temp_bytes, false, err, s);
    if (rc) return rc;
    if ((rc = check_last("sort_pairs", false, s))) return rc;
    // stand-alone operator: report the watchdog synchronously (a caller of a sort wants sorted data or an error)
    uint32_t flag = 0;
    if ((rc = read_u32(err, &flag, 1, s))) return rc;
    if (flag & 1u) { set_error("sort_pairs: the radix-sort watchdog fired (a look-back never saw its predecessor)"); return GSR_ERR_SORT; }
    return GSR_OK;
}

}  // extern "C"
