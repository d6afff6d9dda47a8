/*
 * gsr_b200.h -- C ABI of the B200-native differentiable Gaussian-splatting rasterizer.
 *
 * This is the drop-in boundary for the hot path of
 * XinShuo-ph/gaussian_splatting_with_eye_tracking.  Every entry point replaces one method of
 * the reference's C++ core class CudaRasterizer::Rasterizer (the only thing its torch binding,
 * and the SIBR viewer, ever call):
 *
 *   gsr_forward        <- Rasterizer::forward   submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:30-53
 *   gsr_backward       <- Rasterizer::backward  .../cuda_rasterizer/rasterizer.h:55-84
 *   gsr_mark_visible   <- Rasterizer::markVisible .../cuda_rasterizer/rasterizer.h:23-28
 *   gsr_amr_forward    <- Rasterizer::forward (AMR fork)
 *                         submodules/diff-gaussian-rasterization-amr/cuda_rasterizer/rasterizer.h:30-66
 *   gsr_*_bytes        <- required<GeometryState|BinningState|ImageState>()
 *                         .../cuda_rasterizer/rasterizer_impl.h:66-72
 *
 * Conventions (same as the reference unless stated):
 *   - all data pointers are DEVICE pointers; NULL means "input absent" (reference: nullptr checks at
 *     forward.cu:205,241, rasterizer_impl.cu:229,321,389,411).
 *   - matrices are 16 floats read column-major: element (row r, col c) = m[c*4 + r]
 *     (auxiliary.h:58-77); the Python side stores them transposed (scene/cameras.py:54-56).
 *   - no exceptions, no std::function: status codes + gsr_last_error(); the three growable work
 *     buffers are obtained through a C callback (the analogue of resizeFunctional,
 *     rasterize_points.cu:27-33).
 *   - every launch goes to the caller's stream; the library owns no device memory.
 *
 * No torch types appear here; the PyTorch host shim (gaussian_splatting_with_eye_tracking_b200/_capi.py)
 * binds these with ctypes.
 */
#ifndef GSR_B200_H_INCLUDED
#define GSR_B200_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define GSR_API __attribute__((visibility("default")))
#else
#define GSR_API
#endif

/* status codes */
enum {
    GSR_OK = 0,
    GSR_ERR_INVALID = 1,   /* bad argument (the reference would AT_ERROR / throw)        */
    GSR_ERR_CUDA = 2,      /* a CUDA runtime call failed; see gsr_last_error()            */
    GSR_ERR_ALLOC = 3,     /* the allocation callback returned NULL / too small a buffer  */
    GSR_ERR_WORKSPACE = 4, /* a passed-in work buffer does not carry a valid header       */
    GSR_ERR_SORT = 5       /* device-side watchdog tripped inside the radix sort          */
};

/* which work buffer an allocation request is for (reference: geomFunc/binningFunc/imgFunc) */
enum { GSR_BUF_GEOM = 0, GSR_BUF_BINNING = 1, GSR_BUF_IMAGE = 2 };

/* tile modes */
enum {
    GSR_MODE_BASELINE = 0, /* 16x16 tiles, reference R/ (config.h:16-17)                          */
    GSR_MODE_AMR = 1       /* 32x32 binning tiles rendered as 2x2-strided 16x16 lattices, ref A/  */
};

/* Allocation callback: return a DEVICE pointer to at least `bytes` bytes, 256-byte aligned,
 * valid until the caller of gsr_* releases it.  Called at most once per buffer per gsr_* call. */
typedef void* (*gsr_alloc_fn)(void* user, int which, size_t bytes);

typedef struct gsr_scene {
    int P;                       /* number of Gaussians                                    */
    int sh_degree;               /* D: active SH degree 0..3                               */
    int sh_coeffs;               /* M: coefficients per Gaussian in `shs` (16 for deg 3)    */
    const float* means3D;        /* [P,3]                                                  */
    const float* shs;            /* [P,M,3] or NULL                                        */
    const float* colors_precomp; /* [P,3]   or NULL                                        */
    const float* opacities;      /* [P]  (activated)                                       */
    const float* scales;         /* [P,3] (activated) or NULL                              */
    const float* rotations;      /* [P,4] (r,x,y,z), not normalised by the kernel, or NULL */
    const float* cov3D_precomp;  /* [P,6]  or NULL                                         */
    float scale_modifier;
    /* ---- raw-parameter entry (optional, raw_params == 0: off).  The reference's caller materialises activated
     * copies of every parameter per frame (scene/gaussian_model.py:95-115: sigmoid, exp, normalize and torch.cat of the two SH
     * tensors = 236 B/Gaussian read + 236 B written before the rasterizer reads them again).  With raw_params = 1 the kernels
     * apply the activations on load:
     *     opacities  = raw logits            -> 1 / (1 + exp(-x))
     *     scales     = log-scales            -> exp(x)
     *     rotations  = raw quaternions       -> x / max(|x|, 1e-12)
     *     shs        = features_dc   [P,1,3],   shs_rest = features_rest [P,M-1,3]   (sh_coeffs = M as before)
     * Gradients are still returned w.r.t. the ACTIVATED quantities and the concatenated [P,M,3] SH layout; the host chains them
     * through the activations (a few floats per Gaussian). */
    int raw_params;
    const float* shs_rest;
} gsr_scene;

typedef struct gsr_camera {
    int width, height;
    float tan_fovx, tan_fovy;
    const float* background;     /* [3]  device */
    const float* viewmatrix;     /* [16] device */
    const float* projmatrix;     /* [16] device */
    const float* cam_pos;        /* [3]  device */
    int prefiltered;
    /* Screen-tile band (no reference counterpart): this call bins and blends only tile rows
     * [band_row0, band_row1) of the mode's tile grid (16-px rows baseline, 32-px rows AMR).  Tile ids stay global.
     * radii / num_rendered / out_color then describe the band only.  0,0 = whole frame (reference behaviour).  */
    int band_row0, band_row1;
} gsr_camera;

/* Gaze extension (no reference implementation exists -- the dead tensors at
 * gaussian_renderer_amr/__init__.py:98-106 are its only spec).  NULL => reference behaviour. */
typedef struct gsr_gaze {
    float cx, cy;        /* gaze point in pixels                          */
    float radii[4];      /* step k (1..4) renders tiles within radii[k-1] */
} gsr_gaze;

typedef struct gsr_grads {
    float* dL_dmeans2D;   /* [P,3] (z column written as 0)  */
    float* dL_dcolors;    /* [P,3]                          */
    float* dL_dopacity;   /* [P]                            */
    float* dL_dmeans3D;   /* [P,3]                          */
    float* dL_dcov3D;     /* [P,6]                          */
    float* dL_dsh;        /* [P,M,3] or NULL when M == 0    */
    float* dL_dscales;    /* [P,3]                          */
    float* dL_drotations; /* [P,4]                          */
    float* dL_dconic;     /* [P,4] scratch (reference: internal dL_dconic, rasterize_points.cu:154) */
    float* scratch;       /* [P,12] floats, packed blend-gradient accumulator; must be 16B aligned   */
    float* dL_dcolors_clamped; /* optional [P,3]: dL_dcolors after the SH clamp mask (backward.cu:56-58), i.e. the vector
                                  that multiplies the SH basis; zero rows for culled Gaussians.  When it is given and
                                  dL_dsh is NULL, the [P,M,3] SH gradient is not written at all: data-parallel training
                                  exchanges these 3 floats per Gaussian and view and rebuilds the summed SH gradient with
                                  gsr_sh_grad_from_views (the per-view SH gradient is the rank-1 product basis x this). */
    /* ---- view-parallel training, fused gradient exchange (optional; push_bases == NULL: off).  The preprocess-stage
     * backward then does NOT write dL_dmeans3D / dL_dopacity / dL_dscales / dL_drotations locally: the 256 x 11 floats of every
     * 256-Gaussian block (zeros where a Gaussian is not visible in this view) go straight into the staging area of the rank
     * that OWNS the block (peer memory over NVLink, under the kernel's own run time):
     *     block = idx / 256,  owner = block % push_world  (blocks dealt round-robin, so that all owners receive at once),
     *     slot address = push_bases[owner] + push_offset + (push_rank * (push_rows_per_rank / 256) + block / push_world) * 12288
     * A slot holds the block's 2816 floats ([256][11], row-major) as 94 lines of 128 bytes: 30 data floats followed by two words
     * that both hold push_epoch.  A line is written by 8 lanes of one store instruction and arrives whole, so it validates
     * itself: gsr_peer_owner_sum -- queued on another stream BEFORE this backward, so that it runs beside it -- re-reads a
     * line until its flag is the epoch, sums every block in rank order and delivers the totals to all ranks.  No fence, no
     * kernel boundary between producer and consumer.  Relies on an aligned 128-byte line written by one warp store instruction
     * becoming visible whole -- true within a GPU and over NVLink (what NCCL's LL128 protocol relies on), NOT over PCIe: use it
     * between NVLink peers only.  push_rows_per_rank must be a multiple of 256, push_offset of 128;
     * epochs only grow and are never 0. */
    void* const* push_bases;   /* DEVICE array of `push_world` peer-mapped base pointers (as for gsr_peer_allreduce_sum) */
    size_t push_offset;        /* byte offset of the staging area inside every rank's buffer, 128-byte aligned */
    int push_rank, push_world, push_rows_per_rank;
    unsigned push_epoch;       /* the epoch the matching gsr_peer_owner_sum call waits for */
} gsr_grads;

/* ---- sizes of the three work buffers (bytes, including alignment slack) ---- */
GSR_API size_t gsr_geom_bytes(int P);
GSR_API size_t gsr_binning_bytes(long long num_rendered, long long num_runs, int width, int height, int mode);
GSR_API size_t gsr_image_bytes(int width, int height, int mode);

/* ---- baseline rasterizer ---- */
GSR_API int gsr_forward(const gsr_scene* scene, const gsr_camera* cam,
                        gsr_alloc_fn alloc, void* alloc_user,
                        float* out_color /* [3,H,W] */, int* radii /* [P] or NULL */,
                        int* num_rendered /* host out */, int debug, void* stream);

GSR_API int gsr_backward(const gsr_scene* scene, const gsr_camera* cam,
                         const int* radii, const void* geom_buffer, const void* binning_buffer,
                         const void* image_buffer, int num_rendered,
                         const float* dL_dpix /* [3,H,W] */, const gsr_grads* grads,
                         int debug, void* stream);

/* The same in two calls, so that a view-parallel caller can overlap its gradient exchange with the second stage:
 * GSR_BWD_BLEND      blend-stage backward into `scratch`; with grads->dL_dcolors_clamped given, that tensor is final on
 *                    return of this phase (in stream order)
 * GSR_BWD_PREPROCESS preprocess-stage backward from `scratch` (dL_dpix may be NULL); dL_dsh == NULL with
 *                    dL_dcolors_clamped given = SH rows not written, dL_dcolors_clamped not rewritten
 * GSR_BWD_ALL        both (== gsr_backward) */
enum { GSR_BWD_ALL = 0, GSR_BWD_BLEND = 1, GSR_BWD_PREPROCESS = 2 };
GSR_API int gsr_backward_phase(const gsr_scene* scene, const gsr_camera* cam, const int* radii, const void* geom_buffer,
                               const void* binning_buffer, const void* image_buffer, int num_rendered, const float* dL_dpix,
                               const gsr_grads* grads, int phase, int debug, void* stream);
/* The GSR_BWD_PREPROCESS phase for the Gaussians [first, first + count) only (first a multiple of 256): a view-parallel
 * caller runs the stage as a few consecutive ranges and starts reducing a range's rows of dL_dmeans3D / dL_dopacity /
 * dL_dscales / dL_drotations while the next range is computed.  The ranges of one backward must cover [0, P) once. */
GSR_API int gsr_backward_preprocess_range(const gsr_scene* scene, const gsr_camera* cam, const int* radii, const void* geom_buffer,
                                          const void* binning_buffer, const void* image_buffer, int num_rendered,
                                          const gsr_grads* grads, int first, int count, int debug, void* stream);

GSR_API int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix,
                             const float* projmatrix, unsigned char* present, void* stream);

/* ---- AMR / foveated rasterizer (forward only; the reference AMR backward is non-functional,
 *      SURVEY 2.3 K17) ----
 * fovea_step < 0 : single-pass AMR render;  0 : build buffers + levels only (out_color untouched);
 * 1..4 : progressive round k using the *_precomp buffers built by step 0.                        */
GSR_API int gsr_amr_forward(const gsr_scene* scene, const gsr_camera* cam,
                            int fovea_step, const float* out_color_precomp /* [3,H,W] or NULL */,
                            void* geom_precomp, void* binning_precomp, void* image_precomp,
                            int interpolate_image, const gsr_gaze* gaze /* NULL = reference */,
                            gsr_alloc_fn alloc, void* alloc_user,
                            float* out_color, int* radii, int* num_rendered,
                            int debug, void* stream);

/* ---- tile-band helpers for multi-GPU foveated frames ----
 * After every rank has built its band with gsr_amr_forward(fovea_step = 0) and the per-tile pair counts
 * (gsr_image_view.n_intersections) have been summed over ranks in place, gsr_amr_relevel recomputes the percentile
 * thresholds and tile levels from those global counts (bit-identical to a single-GPU build), and
 * gsr_amr_render_prebuilt renders the band in a single pass (all lattice rounds) from the prebuilt buffers. */
GSR_API int gsr_amr_relevel(void* image_buffer, int width, int height, void* stream);
GSR_API int gsr_amr_render_prebuilt(const gsr_scene* scene, const gsr_camera* cam,
                                    const void* geom_buffer, const void* binning_buffer, void* image_buffer,
                                    int interpolate_image, const gsr_gaze* gaze /* NULL = reference */,
                                    float* out_color, int debug, void* stream);

/* ---- read-only views into the work buffers (tests, band-parallel driver, debugging) ----
 * Fills `out` with device pointers; fields that do not exist for the buffer's mode are NULL. */
typedef struct gsr_geom_view {
    int P;
    const float* depths;            /* [P]   */
    const unsigned char* clamped;   /* [3P]  */
    const int* radii;               /* [P]   */
    const float* means2D;           /* [P,2] */
    const float* cov3D;             /* [P,6] */
    const float* conic_opacity;     /* [P,4] */
    const float* rgb;               /* [P,3] */
    const uint32_t* tiles_touched;  /* [P]   */
    const uint32_t* block_offsets;  /* [ceil(P/256)+1] exclusive counts of VISIBLE Gaussians per 256-block */
    const uint32_t* depth_order;    /* [num_visible] visible Gaussian ids, stable-sorted by depth bits     */
    const uint32_t* num_rendered;   /* [1]   */
} gsr_geom_view;

typedef struct gsr_binning_view {
    long long N;
    const uint64_t* keys_unsorted;  /* always NULL: pairs are never materialised unsorted (run-length binning)  */
    const uint32_t* values_unsorted;/* always NULL                                                              */
    const uint32_t* tiles_sorted;   /* always NULL: the tile of entry i follows from gsr_image_view.ranges; the */
                                    /*     reference's 64-bit key is (tile << 32) | depth_bits[point_list[i]]   */
    const uint32_t* point_list;     /* [N] Gaussian id of every sorted pair (reference: point_list)         */
} gsr_binning_view;

typedef struct gsr_image_view {
    int width, height, mode, tiles_x, tiles_y;
    const float* accum_alpha;        /* [H*W]              */
    const uint32_t* n_contrib;       /* [H*W]              */
    const uint32_t* ranges;          /* [tiles,2]          */
    const uint32_t* n_intersections; /* [tiles] (AMR)      */
    const uint32_t* levels;          /* [tiles] (AMR)      */
    const uint32_t* levels_last;     /* [tiles] (AMR)      */
    const uint32_t* levels_current;  /* [tiles] (AMR)      */
    const uint32_t* thresholds;      /* [3]     (AMR)      */
} gsr_image_view;

GSR_API int gsr_view_geom(const void* geom_buffer, int P, gsr_geom_view* out);
GSR_API int gsr_view_binning(const void* binning_buffer, long long N, gsr_binning_view* out);
GSR_API int gsr_view_image(const void* image_buffer, int width, int height, int mode, gsr_image_view* out);

/* ---- standalone stages (unit-level parity + roofline benches) ---- */
/* stable LSD radix sort of (u64 key, u32 value) pairs on bits [0, end_bit); replaces
 * cub::DeviceRadixSort::SortPairs (rasterizer_impl.cu:303-308).  temp >= gsr_sort_temp_bytes(n). */
GSR_API size_t gsr_sort_temp_bytes(long long n);
GSR_API int gsr_sort_pairs(const uint64_t* keys_in, uint64_t* keys_out,
                           const uint32_t* vals_in, uint32_t* vals_out,
                           uint64_t* keys_tmp, uint32_t* vals_tmp,
                           long long n, int end_bit, void* temp, size_t temp_bytes, void* stream);

/* number of bits needed for tile ids (reference getHigherMsb, rasterizer_impl.cu:35-50) */
GSR_API unsigned gsr_higher_msb(unsigned n);

/* ---- per-stage device timing (CUDA events on the caller's stream, recorded inside real calls) ---- */
enum {
    GSR_PROF_PREPROCESS = 0, GSR_PROF_SCAN, GSR_PROF_DUPLICATE, GSR_PROF_SORT, GSR_PROF_RANGES, GSR_PROF_BLEND,
    GSR_PROF_AMR_LEVELS, GSR_PROF_AMR_BLEND, GSR_PROF_BWD_BLEND, GSR_PROF_BWD_PREPROCESS, GSR_PROF_DEPTH_SORT,
    GSR_PROF_EXPAND,
    GSR_PROF_SORT_PASS0,                      /* .. GSR_PROF_SORT_PASS0 + 7: individual onesweep launches */
    GSR_PROF_SLOTS = GSR_PROF_SORT_PASS0 + 8
};
GSR_API void gsr_profile_enable(int on);
/* waits for the recorded intervals, ADDS their milliseconds into ms[slot] and their number into counts[slot]
 * (both [GSR_PROF_SLOTS], caller-zeroed), then forgets them.  Returns GSR_OK or GSR_ERR_CUDA. */
GSR_API int gsr_profile_read(float* ms, int* counts);

/* ---- data-parallel helper: summed SH gradient of `num_views` views from their clamped colour gradients ----
 * dL_dsh[g][k][c] = sum_v basis_k(normalize(means3D[g] - cam_positions[v])) * dL_dcolors_clamped[v][g][c]
 * (computeColorFromSH backward, backward.cu:20-139: every coefficient's gradient is its basis value times dL_dRGB).
 * Views are summed in index order, so every rank that holds the same gathered inputs gets bit-identical sums.
 * cam_positions [num_views,3], dL_dcolors_clamped [num_views,P,3], dL_dsh [P,sh_coeffs,3] (16-byte aligned), all device. */
GSR_API int gsr_sh_grad_from_views(int P, int sh_degree, int sh_coeffs, const float* means3D, int num_views,
                                   const float* cam_positions, const float* dL_dcolors_clamped, float* dL_dsh, void* stream);

/* ---- the same exchange with only the visible Gaussians' rows on the wire ----
 * gsr_pack_visible_rows: rows_out[slot] = rows_in[g] ([.,3] floats) for the Gaussians with radii[g] > 0, slots in index
 *   order; meta [gsr_pack_meta_words(P)] = ceil(P/32) mask words (bit g%32 of word g/32) followed by ceil(P/32) words
 *   "set bits before this word".  Rows whose slot >= cap are dropped (size cap from num_visible, gsr_last_counts).
 *   temp: gsr_pack_temp_bytes(P) bytes of device scratch.  Every view's (rows, meta) then travels instead of [P,3].
 * gsr_sh_grad_from_packed_views: gsr_sh_grad_from_views reading view v's rows at rows + v*cap*3 and its meta at
 *   meta + v*meta_stride_words; a Gaussian whose bit is clear in a view contributes nothing from it.
 * gsr_last_counts: totals of the forward this host thread ran last on `geom_buffer` (from the header mailbox, no
 *   device round trip): pairs, (Gaussian, tile-row) runs, visible Gaussians.  GSR_ERR_INVALID for any other buffer. */
GSR_API size_t gsr_pack_meta_words(int P);
GSR_API size_t gsr_pack_temp_bytes(int P);
GSR_API int gsr_pack_visible_rows(int P, const int* radii, const float* rows_in, int cap, float* rows_out, uint32_t* meta,
                                  void* temp, size_t temp_bytes, void* stream);
/* gsr_pack_visible_rows_noalloc: gsr_pack_visible_rows for callers that cannot allocate device memory (code under CUDA graph
 *   capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 *   Same inputs, outputs and results; the scan scratch lives in a file-scope device array inside the
 *   library (65 792 bytes on every GPU that loads the library, held until exit) instead of `temp`.  The entry point calls no
 *   cudaMalloc / cudaFree / allocating routine.  Limit: P <= gsr_pack_noalloc_max_gaussians() = 16 777 216, larger P is
 *   rejected with GSR_ERR_INVALID (use gsr_pack_visible_rows with caller scratch).  All calls on a device share the array:
 *   calls that overlap on the GPU (different streams without an ordering between them) RACE -- keep them on one stream or order
 *   them with events.  Nothing has to be cleared between calls: every call writes each scratch word before it reads it. */
GSR_API int gsr_pack_visible_rows_noalloc(int P, const int* radii, const float* rows_in, int cap, float* rows_out, uint32_t* meta,
                                          void* stream);
GSR_API int gsr_pack_noalloc_max_gaussians(void);
GSR_API int gsr_sh_grad_from_packed_views(int P, int sh_degree, int sh_coeffs, const float* means3D, int num_views,
                                          const float* cam_positions, const uint32_t* meta, size_t meta_stride_words,
                                          const float* rows, int cap, float* dL_dsh, void* stream);
GSR_API int gsr_last_counts(const void* geom_buffer, int P, long long* num_rendered, long long* num_runs, long long* num_visible);

/* ---- training-step surround: the loss and the optimizer either side of the rasterizer ----
 * gsr_l1_ssim_forward  <- utils/loss_utils.py:17-18 (l1_loss) + :39-63 (ssim, window 11, sigma 1.5, zero padding, mean),
 *                         combined as train.py:87-88.  image, gt: [channels,height,width] floats.
 *                         out[3] = { w_l1*L1 + w_ssim*SSIM + w_const, L1, SSIM }   (train.py: w_l1 = 1-l, w_ssim = -l, w_const = l)
 *                         maps: [3,channels,height,width] scratch the backward reads (d ssim / d{mu1, E[xx], E[xy]} per window
 *                         centre), or NULL when no gradient is wanted.  partials: 2*gsr_l1_ssim_partials(...) floats of scratch;
 *                         the final sum runs in a fixed order in double (deterministic).
 * gsr_l1_ssim_backward <- autograd of the above w.r.t. image: dL_dimage [channels,height,width];
 *                         dloss = device scalar upstream gradient of out[0], or NULL for 1.
 * gsr_adam_step        <- torch.optim.Adam.step() as configured by scene/gaussian_model.py:149-175 (betas, eps=1e-15, no weight
 *                         decay, no amsgrad), every tensor of every group in ONE launch (per 8 tensors); `step` is the 1-based
 *                         step count of that tensor AFTER this update, lr its group's learning rate. */
typedef struct gsr_adam_tensor {
    float* param;            /* [n] updated in place */
    const float* grad;       /* [n]                  */
    float* exp_avg;          /* [n] updated in place */
    float* exp_avg_sq;       /* [n] updated in place */
    long long n;
    double lr;               /* doubles like the Python floats torch derives its step size from */
    int step;
} gsr_adam_tensor;
GSR_API size_t gsr_l1_ssim_partials(int channels, int height, int width);
GSR_API int gsr_l1_ssim_forward(int channels, int height, int width, const float* image, const float* gt, float w_l1, float w_ssim,
                                float w_const, float* maps, float* partials, float* out, void* stream);
GSR_API int gsr_l1_ssim_backward(int channels, int height, int width, const float* image, const float* gt, float w_l1, float w_ssim,
                                 const float* maps, const float* dloss, float* dL_dimage, void* stream);
GSR_API int gsr_adam_step(const gsr_adam_tensor* tensors, int count, double beta1, double beta2, double eps, void* stream);

/* ---- presenter (replaces train.py:60): float32 [3,H,W] frame -> uint8 [H,W,3], clamp to [0,1], x255, truncate.
 * The 8-bit interleaved layout is what the reference's network viewer protocol carries (gaussian_renderer/network_gui.py:50-55). */
GSR_API int gsr_pack_rgb8(const float* image_chw, int width, int height, unsigned char* out_hwc, void* stream);

/* ---- view-parallel gradient sum over NVLink peer memory (replaces the NCCL all-reduce of the 11 dense
 * floats per Gaussian that follows the preprocess-stage backward) ----
 * peer_bases: DEVICE array of `world` base pointers, entry k = rank k's buffer as addressable from THIS device (peer-mapped
 * by the host: torch symmetric memory / cudaIpc; all on one device in the single-GPU tests).  Every buffer holds the data
 * ([n_floats] at data_offset, 16-byte aligned) and gsr_peer_flag_bytes(world) of flag words at flag_offset, zeroed once
 * before first use.  gsr_peer_allreduce_sum = barrier(epoch) + one kernel in which rank r loads the r-th 1/world slice
 * from every peer in rank order, adds, and stores the sum into every peer's buffer + barrier(epoch + 1): in place, the same
 * bits on every rank.  `epoch` must grow by 2 per call (by 1 per gsr_peer_barrier), identically on all ranks, from 1. */
GSR_API size_t gsr_peer_flag_bytes(int world);
GSR_API int gsr_peer_barrier(void* const* peer_bases, size_t flag_offset, int world, int rank, unsigned epoch, void* stream);
/* the same with the sum formed inside the NVSwitch: multicast_base = the multicast (cuMulticast) mapping of the same buffers;
 * multimem.ld_reduce + multimem.st, (1 + 1/world) x payload per direction instead of 2 (world-1)/world.  Identical bits on
 * all ranks; the switch's add order is not the rank order of the unicast variant. */
GSR_API int gsr_peer_allreduce_sum_mc(void* const* peer_bases, void* multicast_base, size_t data_offset, long long n_floats,
                                      size_t flag_offset, int world, int rank, unsigned epoch, void* stream);
/* Second half of the fused exchange (see gsr_grads.push_*), a streaming consumer: barrier(epoch) -- every rank is past the last
 * reader of the previous totals --, then rank r takes the blocks it owns (block % world == r) in order; every 128-byte line of
 * a block is re-read from each rank's slot until its flag words hold `epoch`, the lines are added in rank order (so every run
 * gives the same bits and every rank the same totals) and the totals are written into EVERY rank's dense gradient block at
 * final_offset -- [means3D 3 x P4 | opacity P4 | scales 3 x P4 | rotations 4 x P4] floats, P4 = P rounded up to 4, the layout
 * the torch host carves the four tensors from -- through the multicast mapping when multicast_base != NULL (one store per
 * value, replicated by the NVSwitch) or with one store per peer; barrier(epoch + 1).  Queue it on its own stream before the
 * backward that pushes: it then runs beside that kernel (`ctas` resident CTAs of 256 threads, 0 = one per ~160 owned blocks, 37..74) and only the
 * last blocks and the closing barrier remain when the backward ends.  A line that does not arrive within the watchdog time
 * poisons the totals (NaN) and raises gsr_peer_error.  epoch must not be 0. */
GSR_API int gsr_peer_owner_sum(void* const* peer_bases, void* multicast_base, size_t staging_offset, size_t final_offset,
                               int P, int rows_per_rank, size_t flag_offset, int world, int rank, unsigned epoch, int ctas,
                               void* stream);
/* 1 if a peer barrier of this process timed out since the last call (the word is cleared); the kernels that follow a timed-out
 * barrier write NaN instead of sums, so a rank that ran ahead can never train on stale gradients silently. */
GSR_API int gsr_peer_error(void);
GSR_API int gsr_peer_allreduce_sum(void* const* peer_bases, size_t data_offset, long long n_floats, size_t flag_offset, int world,
                                   int rank, unsigned epoch, void* stream);

/* ---- measurement: algorithmic work of the blend stage of a frame built by gsr_forward.
 * totals_dev[0] = pair-tests of the forward  (sum over pixels of the list entries the reference's loop examines before the
 *                 pixel is done, cuda_rasterizer/forward.cu:328-361),
 * totals_dev[1] = hits (pair-tests that pass both skips and are accumulated; what the backward differentiates),
 * totals_dev[2] = sum over pixels of n_contrib (entries the backward walks, backward.cu:486-488).
 * One extra kernel, one pixel per thread, no culling; never part of a frame.  totals_dev: 3 x u64 in device memory. */
GSR_API int gsr_blend_stats(int P, int W, int H, const void* geom_buffer, const void* binning_buffer, const void* image_buffer,
                            unsigned long long* totals_dev, void* stream);

/* ---- tuning knobs (results never depend on them, with ONE exception: GSR_OPT_PEER_CTAS_PER_SM values >= 0x100 put the
 * unicast peer kernel into a link-diagnostic mode -- 0x1xx = remote loads only, 0x2xx = remote stores only -- that does NOT
 * compute the sum; tools/peer_diag.py uses it to time the two directions and nothing else may) ---- */
enum { GSR_OPT_SORT_VARIANT = 0,  /* onesweep tile shape for u32 keys: 0 = default, 1..5 see gsr_binning.cu */
       GSR_OPT_BLEND_VARIANT = 1, /* forward blend: 0 = two pixels per thread, packed FP32x2 (default); 1 = one pixel per thread; 2 = as 0 with cp.async staging (A/B) */
       GSR_OPT_PEER_CTAS_PER_SM = 2, /* gsr_peer_allreduce_sum*: grid cap in CTAs per SM, 0 = default */
       GSR_OPT_PEER_TIMEOUT_MS = 6, /* peer barrier watchdog in milliseconds (default 600 000 = 10 min, NCCL-like); on expiry the
                                       barrier posts gsr_peer_error and the data kernel behind it writes NaN */
       GSR_OPT_NVTX = 5,            /* 1: NVTX range per stage (names "gsr:<stage>"); also enabled by GSR_NVTX=1 in the environment */
       GSR_OPT_FAKE_SORT_ERROR = 4, /* 1: the next frames post a tripped radix-sort watchdog (test hook for GSR_ERR_SORT) */
       GSR_OPT_FORCE_SORT_FALLBACK = 3 /* 1: depth-sort with the global onesweep chain even when the range buckets fit (the
                                          path taken when a depth-histogram bin overflows a bucket); results are identical */ };
GSR_API int gsr_set_option(int option, int value);

/* counters: how many kernels this library launched since load (bench.py "gpu_launches") */
GSR_API unsigned long long gsr_launch_count(void);

GSR_API const char* gsr_last_error(void);
GSR_API const char* gsr_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GSR_B200_H_INCLUDED */
