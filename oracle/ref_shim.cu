/*
 * ref_shim.cu -- C-ABI wrapper around the UNMODIFIED reference CUDA rasterizer.  TEST / BASELINE INFRASTRUCTURE.
 *
 * Compiled by oracle/Makefile together with the reference's own sources, taken where they lie under
 * /root/reference/submodules/diff-gaussian-rasterization{,-amr}/cuda_rasterizer/{forward,backward,rasterizer_impl}.cu
 * into oracle/_ref/libref_raster{,_amr}.so (git-ignored; travels to the GPU box).  No reference source is copied
 * into this repository: this file only #includes the reference headers at build time and calls
 * CudaRasterizer::Rasterizer::{forward,backward,markVisible} (cuda_rasterizer/rasterizer.h:20-86) plus the
 * reference's own fromChunk() decoders (rasterizer_impl.h:31-63) to expose its intermediate buffers.
 *
 * Used (a) to generate golden vectors and check parity on the GPU box, (b) as the timed reference baseline.
 *
 * -DREF_AMR selects the AMR fork's signature.  The AMR fork passes a host stack array to a kernel
 * (rasterizer_impl.cu:610,644); to run it unmodified on a machine without pageable-memory access the call is
 * executed on a worker thread whose stack has been cudaHostRegister'ed (UVA makes it device-addressable).
 */
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <string>
#include <pthread.h>

#include "rasterizer.h"
#include "rasterizer_impl.h"

#define REF_API extern "C" __attribute__((visibility("default")))

typedef char* (*ref_alloc_fn)(void* user, int which, size_t bytes);   // which: 0 geom, 1 binning, 2 image

static thread_local std::string g_err;
static std::string g_err_shared;
REF_API const char* ref_last_error(void) { return g_err_shared.c_str(); }
REF_API int ref_is_amr(void) {
#ifdef REF_AMR
    return 1;
#else
    return 0;
#endif
}

struct FwdArgs {
    int P, D, M; const float* bg; int W, H;
    const float *means3D, *shs, *colors_precomp, *opacities, *scales; float scale_modifier;
    const float *rotations, *cov3D_precomp, *view, *proj, *campos; float tan_fovx, tan_fovy; int prefiltered;
    int fovea_step; const float* out_color_precomp; char *geom_pre, *bin_pre, *img_pre;
    float* out_color; int* radii; int interpolate; int debug;
    ref_alloc_fn alloc; void* user;
    int result;
};

static void run_forward(FwdArgs* a) {
    try {
        ref_alloc_fn alloc = a->alloc; void* user = a->user;
        std::function<char*(size_t)> gf = [=](size_t n) { return alloc(user, 0, n); };
        std::function<char*(size_t)> bf = [=](size_t n) { return alloc(user, 1, n); };
        std::function<char*(size_t)> imf = [=](size_t n) { return alloc(user, 2, n); };
        a->result = CudaRasterizer::Rasterizer::forward(gf, bf, imf, a->P, a->D, a->M, a->bg, a->W, a->H, a->means3D, a->shs,
            a->colors_precomp, a->opacities, a->scales, a->scale_modifier, a->rotations, a->cov3D_precomp, a->view, a->proj,
            a->campos, a->tan_fovx, a->tan_fovy, a->prefiltered != 0,
#ifdef REF_AMR
            a->fovea_step, a->out_color_precomp, a->geom_pre, a->bin_pre, a->img_pre,
#endif
            a->out_color, a->radii,
#ifdef REF_AMR
            a->interpolate != 0,
#endif
            a->debug != 0);
    } catch (const std::exception& e) {
        g_err_shared = e.what();
        a->result = -1;
    }
}

#ifdef REF_AMR
// ---- persistent worker thread with a device-addressable stack
static pthread_t g_thr; static bool g_thr_up = false; static int g_thr_ok = 0;
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t g_cv_req = PTHREAD_COND_INITIALIZER, g_cv_done = PTHREAD_COND_INITIALIZER;
static FwdArgs* g_req = nullptr; static bool g_done = false; static int g_dev = 0;

static void* worker(void*) {
    cudaSetDevice(g_dev);
    pthread_mutex_lock(&g_mu);
    for (;;) {
        while (!g_req) pthread_cond_wait(&g_cv_req, &g_mu);
        FwdArgs* a = g_req;
        pthread_mutex_unlock(&g_mu);
        run_forward(a);
        pthread_mutex_lock(&g_mu);
        g_req = nullptr; g_done = true;
        pthread_cond_signal(&g_cv_done);
    }
    return nullptr;
}

static int ensure_worker() {
    if (g_thr_up) return g_thr_ok;
    g_thr_up = true;
    cudaGetDevice(&g_dev);
    int can = 0;
    cudaDeviceGetAttribute(&can, cudaDevAttrCanUseHostPointerForRegisteredMem, g_dev);
    int pageable = 0;
    cudaDeviceGetAttribute(&pageable, cudaDevAttrPageableMemoryAccess, g_dev);
    const size_t stack_bytes = 16u << 20;
    void* stack = nullptr;
    if (posix_memalign(&stack, 1 << 16, stack_bytes) != 0) { g_err_shared = "posix_memalign failed"; return g_thr_ok = 0; }
    memset(stack, 0, stack_bytes);
    if (!pageable) {
        if (!can) { g_err_shared = "device cannot address registered host memory by host pointer; AMR reference cannot run unmodified"; return g_thr_ok = 0; }
        cudaError_t e = cudaHostRegister(stack, stack_bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
        if (e != cudaSuccess) { g_err_shared = std::string("cudaHostRegister(stack): ") + cudaGetErrorString(e); cudaGetLastError(); return g_thr_ok = 0; }
    }
    pthread_attr_t at; pthread_attr_init(&at);
    pthread_attr_setstack(&at, stack, stack_bytes);
    if (pthread_create(&g_thr, &at, worker, nullptr) != 0) { g_err_shared = "pthread_create failed"; return g_thr_ok = 0; }
    return g_thr_ok = 1;
}

static void dispatch(FwdArgs* a) {
    if (!ensure_worker()) { a->result = -2; return; }
    pthread_mutex_lock(&g_mu);
    g_req = a; g_done = false;
    pthread_cond_signal(&g_cv_req);
    while (!g_done) pthread_cond_wait(&g_cv_done, &g_mu);
    pthread_mutex_unlock(&g_mu);
}
#else
static void dispatch(FwdArgs* a) { run_forward(a); }
#endif

REF_API int ref_forward(int P, int D, int M, const float* bg, int W, int H, const float* means3D, const float* shs,
                        const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                        const float* rotations, const float* cov3D_precomp, const float* view, const float* proj,
                        const float* campos, float tan_fovx, float tan_fovy, int prefiltered,
                        int fovea_step, const float* out_color_precomp, char* geom_pre, char* bin_pre, char* img_pre,
                        float* out_color, int* radii, int interpolate, int debug, ref_alloc_fn alloc, void* user) {
    FwdArgs a;
    a.P = P; a.D = D; a.M = M; a.bg = bg; a.W = W; a.H = H; a.means3D = means3D; a.shs = shs; a.colors_precomp = colors_precomp;
    a.opacities = opacities; a.scales = scales; a.scale_modifier = scale_modifier; a.rotations = rotations;
    a.cov3D_precomp = cov3D_precomp; a.view = view; a.proj = proj; a.campos = campos; a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
    a.prefiltered = prefiltered; a.fovea_step = fovea_step; a.out_color_precomp = out_color_precomp;
    a.geom_pre = geom_pre; a.bin_pre = bin_pre; a.img_pre = img_pre; a.out_color = out_color; a.radii = radii;
    a.interpolate = interpolate; a.debug = debug; a.alloc = alloc; a.user = user; a.result = -1;
    dispatch(&a);
    return a.result;
}

REF_API int ref_backward(int P, int D, int M, int R, const float* bg, int W, int H, const float* means3D, const float* shs,
                         const float* colors_precomp, const float* scales, float scale_modifier, const float* rotations,
                         const float* cov3D_precomp, const float* view, const float* proj, const float* campos,
                         float tan_fovx, float tan_fovy, const int* radii, char* geom, char* binning, char* image,
                         const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                         float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, int debug) {
    try {
        CudaRasterizer::Rasterizer::backward(P, D, M, R, bg, W, H, means3D, shs, colors_precomp, scales, scale_modifier, rotations,
            cov3D_precomp, view, proj, campos, tan_fovx, tan_fovy, radii, geom, binning, image, dL_dpix, dL_dmean2D, dL_dconic,
            dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug != 0);
        return 0;
    } catch (const std::exception& e) { g_err_shared = e.what(); return -1; }
}

REF_API void ref_mark_visible(int P, float* means3D, float* view, float* proj, bool* present) {
    CudaRasterizer::Rasterizer::markVisible(P, means3D, view, proj, present);
}

// ---- views into the reference's buffers, decoded by the reference's own fromChunk()
struct RefGeomView { const float* depths; const bool* clamped; const int* internal_radii; const float* means2D; const float* cov3D;
                     const float* conic_opacity; const float* rgb; const uint32_t* point_offsets; const uint32_t* tiles_touched; };
struct RefBinView { const uint64_t* keys_unsorted; const uint64_t* keys; const uint32_t* list_unsorted; const uint32_t* list; };
struct RefImgView { const uint32_t* ranges; const uint32_t* n_contrib; const float* accum_alpha;
                    const uint32_t* n_intersections; const uint32_t* levels; const uint32_t* levels_last; const uint32_t* levels_current; };

REF_API void ref_view_geom(char* chunk, int P, RefGeomView* o) {
    CudaRasterizer::GeometryState g = CudaRasterizer::GeometryState::fromChunk(chunk, P);
    o->depths = g.depths; o->clamped = g.clamped; o->internal_radii = g.internal_radii; o->means2D = (const float*)g.means2D;
    o->cov3D = g.cov3D; o->conic_opacity = (const float*)g.conic_opacity; o->rgb = g.rgb; o->point_offsets = g.point_offsets;
    o->tiles_touched = g.tiles_touched;
}
REF_API void ref_view_binning(char* chunk, int N, RefBinView* o) {
    CudaRasterizer::BinningState b = CudaRasterizer::BinningState::fromChunk(chunk, N);
    o->keys_unsorted = b.point_list_keys_unsorted; o->keys = b.point_list_keys; o->list_unsorted = b.point_list_unsorted; o->list = b.point_list;
}
REF_API void ref_view_image(char* chunk, int W, int H, RefImgView* o) {
    CudaRasterizer::ImageState im = CudaRasterizer::ImageState::fromChunk(chunk, (size_t)W * H);
    o->ranges = (const uint32_t*)im.ranges; o->n_contrib = im.n_contrib; o->accum_alpha = im.accum_alpha;
#ifdef REF_AMR
    o->n_intersections = im.n_intersections; o->levels = im.tile_AMR_levels; o->levels_last = im.tile_AMR_levels_last;
    o->levels_current = im.tile_AMR_levels_current;
#else
    o->n_intersections = nullptr; o->levels = nullptr; o->levels_last = nullptr; o->levels_current = nullptr;
#endif
}
