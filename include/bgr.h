/*
 * bgr.h — C ABI of the Blackwell Gaussian rasterizer (libbgr.so, sm_100a only).
 *
 * This is the drop-in boundary for the reference's rasterizer core. Every entry point names the
 * reference interface it replaces (paths relative to
 * /root/reference/submodules/diff-gaussian-rasterization/):
 *
 *   bgr_forward       <- CudaRasterizer::Rasterizer::forward      cuda_rasterizer/rasterizer.h:31-54
 *                        (called from RasterizeGaussiansCUDA,      rasterize_points.cu:35-115)
 *   bgr_backward      <- CudaRasterizer::Rasterizer::backward     cuda_rasterizer/rasterizer.h:56-85
 *                        (called from RasterizeGaussiansBackwardCUDA, rasterize_points.cu:117-196)
 *   bgr_mark_visible  <- CudaRasterizer::Rasterizer::markVisible  cuda_rasterizer/rasterizer.h:24-29
 *                        (called from markVisible,                 rasterize_points.cu:198-217)
 *   bgr_alloc_fn      <- std::function<char*(size_t)>             cuda_rasterizer/rasterizer.h:32-34
 *                        (resizeFunctional,                        rasterize_points.cu:27-33)
 *
 * Plain pointers and sizes only; no C++ types, no exceptions cross this boundary. All data
 * pointers are DEVICE pointers to contiguous float32/int32 arrays with the reference's shapes; an
 * absent optional input is NULL (the reference's "empty tensor => nullptr" convention,
 * rasterize_points.cu:94-111). All work is enqueued on `stream` (a cudaStream_t passed as void*).
 * bgr_forward synchronises that stream once to read back the number of rendered tile instances
 * and the longest tile list (the reference does a blocking cudaMemcpy at rasterizer_impl.cu:281).
 *
 * Every function returns BGR_OK (0) or a negative status; bgr_last_error() returns a
 * thread-local human-readable message for the last failure on the calling thread.
 */
#ifndef BGR_H_INCLUDED
#define BGR_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define BGR_API __attribute__((visibility("default")))
#else
#define BGR_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define BGR_OK                 0
#define BGR_ERR_INVALID_ARG   -1   /* bad shape / NULL where data is required              */
#define BGR_ERR_CUDA          -2   /* a CUDA runtime call or (debug mode) a kernel failed   */
#define BGR_ERR_ALLOC         -3   /* an allocator callback returned NULL                   */
#define BGR_ERR_OVERFLOW      -4   /* num_rendered >= 2^30 tile instances, counted in 64    */
                                   /* bits (the reference overflows its int silently,       */
                                   /* rasterizer_impl.cu:280-283)                           */

/* Grows (or allocates) a scratch buffer to `bytes` and returns its device address. */
typedef void* (*bgr_alloc_fn)(size_t bytes, void* user);

typedef struct BgrForwardArgs {
    int P;                       /* number of Gaussians                                     */
    int D;                       /* active SH degree (0..3)                                 */
    int M;                       /* SH coefficients per channel in `shs` (0 when shs==NULL) */
    int width, height;
    const float* background;     /* [3]                                                     */
    const float* means3D;        /* [P,3]                                                   */
    const float* shs;            /* [P,M,3] or NULL                                         */
    const float* colors_precomp; /* [P,3]   or NULL                                         */
    const float* opacities;      /* [P]                                                     */
    const float* scales;         /* [P,3]   or NULL                                         */
    float        scale_modifier;
    const float* rotations;      /* [P,4]   or NULL                                         */
    const float* cov3D_precomp;  /* [P,6]   or NULL                                         */
    const float* viewmatrix;     /* [16]                                                    */
    const float* projmatrix;     /* [16]                                                    */
    const float* cam_pos;        /* [3]                                                     */
    float        tan_fovx, tan_fovy;
    int          prefiltered;
    int          debug;          /* sync + check after every stage (auxiliary.h:166-173)    */
    float*       out_color;      /* [3,H,W]  every element is written                       */
    int*         radii;          /* [P]      every element is written                       */
} BgrForwardArgs;

typedef struct BgrBackwardArgs {
    int P, D, M;
    int R;                       /* num_rendered returned by bgr_forward                    */
    int width, height;
    const float* background;
    const float* means3D;
    const float* shs;            /* or NULL */
    const float* colors_precomp; /* or NULL */
    const float* scales;         /* or NULL */
    float        scale_modifier;
    const float* rotations;      /* or NULL */
    const float* cov3D_precomp;  /* or NULL */
    const float* viewmatrix;
    const float* projmatrix;
    const float* cam_pos;
    float        tan_fovx, tan_fovy;
    const int*   radii;          /* [P] from forward                                        */
    const void*  geom_buffer;    /* the three scratch buffers filled by bgr_forward         */
    const void*  binning_buffer;
    const void*  image_buffer;
    const float* dL_dpix;        /* [3,H,W]                                                 */
    void*        scratch;        /* bgr_backward_scratch_bytes(P) bytes, contents ignored   */
    /* outputs: every element of every array is written (no pre-zeroing required)          */
    float* dL_dmean2D;           /* [P,3] (z column = 0)                                    */
    float* dL_dcolor;            /* [P,3]                                                   */
    float* dL_dopacity;          /* [P]                                                     */
    float* dL_dmean3D;           /* [P,3]                                                   */
    float* dL_dcov3D;            /* [P,6]                                                   */
    float* dL_dsh;               /* [P,M,3]  (may be NULL when M == 0)                      */
    float* dL_dscale;            /* [P,3]                                                   */
    float* dL_drot;              /* [P,4]                                                   */
    int    debug;
    /* View-parallel path only (both NULL otherwise). dL/dsh = basis(dir) x dL/dRGB is rank one, so
     * instead of writing dL_dsh the backward MULTICASTS (NVLink multimem.st) the clamp-masked
     * dL/dRGB [P,3] into this rank's slot of a symmetric buffer on every GPU, and its camera
     * position into its [4]-float slot; every rank then rebuilds the view-summed dL_dsh locally with
     * bgr_sh_grad_from_factors. Needs P % 4 == 0. dL_dsh is not written by bgr_backward then.   */
    float* sh_factor_multicast;  /* multicast address of slot [rank][P][3]                  */
    float* campos_multicast;     /* multicast address of slot [rank][4]                     */
    /* Optional split of the call (0 = do everything, the drop-in behaviour). A view-parallel
     * caller runs BGR_BACKWARD_ACCUMULATE once (clear + render backward, reference
     * backward.cu:399-557) and then BGR_BACKWARD_PREPROCESS (backward.cu:140-396 fused) over
     * consecutive Gaussian ranges [gaussian_begin, gaussian_end), exchanging the finished range
     * while the next one is computed. Range bounds must be multiples of 4 on the multicast path. */
    int    stages;
    int    gaussian_begin;
    int    gaussian_end;
} BgrBackwardArgs;
#define BGR_BACKWARD_ACCUMULATE 1
#define BGR_BACKWARD_PREPROCESS 2

/* Forward rasterization. Writes *num_rendered (total tile instances, the reference's return
 * value). geom/binning/image are grown through the callbacks exactly once each. */
BGR_API int bgr_forward(const BgrForwardArgs* args,
                bgr_alloc_fn geom_alloc, void* geom_user,
                bgr_alloc_fn binning_alloc, void* binning_user,
                bgr_alloc_fn image_alloc, void* image_user,
                void* stream, int* num_rendered);

BGR_API int bgr_backward(const BgrBackwardArgs* args, void* stream);

/* ---- extended entry: the trainer's activations fused in (new) -------------
 * The trainer keeps RAW leaves and activates them on every step before it calls the rasterizer
 * (scene/gaussian_model.py:95-115 via gaussian_renderer/__init__.py:53-82):
 *     scales    = exp(_scaling)              rotations = normalize(_rotation)   (x / max(|x|, 1e-12))
 *     opacities = sigmoid(_opacity)          shs       = cat(_features_dc, _features_rest, dim=1)
 * each of which materialises a P-sized tensor forward and another one backward (the SH cat alone
 * copies 192 B per Gaussian each way). These two entry points take the raw leaves and do the
 * activations inside the preprocess kernels, and the backward returns the gradients w.r.t. the RAW
 * leaves. The argument blocks are the same, read as follows:
 *     scales = _scaling [P,3]   rotations = _rotation [P,4]   opacities = _opacity [P]
 *     shs = _features_dc [P,1,3]      features_rest = _features_rest [P,M-1,3] (NULL when M == 1)
 *     M   = coefficients per channel of the two SH leaves together (1..16)
 *     dL_dscale / dL_drot / dL_dopacity = gradients of the raw leaves
 *     dL_dsh = dL/d_features_dc [P,1,3]     dL_dfeatures_rest = dL/d_features_rest [P,M-1,3]
 * colors_precomp and cov3D_precomp must be NULL (use the drop-in entry for those modes), and so must
 * the view-parallel multicast slots. Everything else - buffers, num_rendered, the image, radii and
 * the remaining gradients - is exactly what bgr_forward / bgr_backward produce for the activated
 * values. */
BGR_API int bgr_forward_raw(const BgrForwardArgs* args, const float* features_rest,
                bgr_alloc_fn geom_alloc, void* geom_user,
                bgr_alloc_fn binning_alloc, void* binning_user,
                bgr_alloc_fn image_alloc, void* image_user,
                void* stream, int* num_rendered);
BGR_API int bgr_backward_raw(const BgrBackwardArgs* args, const float* features_rest,
                float* dL_dfeatures_rest, void* stream);

/* ---- the step after the path (new) ---------------------------------------
 * The trainer's loss, train.py:91-92 with utils/loss_utils.py:16-63:
 *     loss = (1 - lambda_dssim) * mean|image - gt| + lambda_dssim * (1 - ssim(image, gt))
 * (11x11 Gaussian window, sigma 1.5, zero padding, C1 = 0.01^2, C2 = 0.03^2, mean over 3 H W) and
 * its gradient w.r.t. the rendered image, in three launches instead of PyTorch's ~50:
 * image, gt: [3,H,W]. scratch: bgr_loss_scratch_bytes(W,H) bytes. loss3 (device): {loss, l1, ssim}.
 * dL_dimage: [3,H,W] = d loss / d image (pass NULL to skip the backward kernel). */
BGR_API size_t bgr_loss_scratch_bytes(int width, int height);
BGR_API int bgr_l1_ssim_loss(const float* image, const float* gt, int width, int height,
                             float lambda_dssim, void* scratch, float* loss3, float* dL_dimage,
                             void* stream);
/* (row f3) Densification statistics of one view, train.py:114-116 + scene/gaussian_model.py:405-407,
 * in one pass: for every Gaussian with radii > 0
 *     max_radii2D = max(max_radii2D, radii);  xyz_gradient_accum += |dL_dmean2D[:, :2]|;  denom += 1
 * radii: [P] int32 and dL_dmean2D: [P,3] as bgr_forward / bgr_backward return them; the three
 * accumulators are [P] (or [P,1]) float32, updated in place. */
BGR_API int bgr_densification_stats(int P, const int* radii, const float* dL_dmean2D, float* max_radii2D,
                                    float* xyz_gradient_accum, float* denom, void* stream);
/* torch.optim.Adam's update (scene/gaussian_model.py:149-160: six parameter groups, one learning
 * rate each, eps 1e-15) for up to 8 groups in ONE launch. step is the 1-based update count. The
 * hyper-parameters are doubles, as in torch: 1 - beta, the bias corrections and lr / (1 - beta1^t)
 * are formed in double on the host and rounded to float32 once. */
BGR_API int bgr_adam_step(int n_groups, float* const* params, const float* const* grads,
                          float* const* exp_avg, float* const* exp_avg_sq, const size_t* numel,
                          const double* lr, double beta1, double beta2, double eps, int step, void* stream);

/* present[i] = 1 iff Gaussian i passes the near-plane test (auxiliary.h:139-164). */
BGR_API int bgr_mark_visible(int P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, uint8_t* present, void* stream);

/* Scratch sizes (bytes). */
BGR_API size_t bgr_geom_bytes(int P);
BGR_API size_t bgr_image_bytes(int width, int height);
BGR_API size_t bgr_binning_bytes(int R);
BGR_API size_t bgr_backward_scratch_bytes(int P);

/* Introspection for the parity tests: byte offsets of the sub-arrays inside the scratch buffers.
 * geom   : [0] records (P x 48 B: x,y,conicA,conicB | conicC,opacity,r,g | b,thr,depth,clamp bits)
 *          [1] tiles_touched u32[P]
 * image  : [0] final_T f32[W*H]  [1] n_contrib u32[W*H]  [2] ranges uint2[tiles]
 *          [3] control block {u32 num_rendered, u32 longest tile list, u32 visible, ...}
 * binning: [0] point_list u32[R] (sorted Gaussian ids)  [1] per-tile key buckets u64[R]
 *          [2] alternate key array u64[R] (ping-pong partner of the per-tile sort's general path) */
BGR_API void bgr_geom_layout(int P, size_t* offsets2);
BGR_API void bgr_image_layout(int width, int height, size_t* offsets4);
BGR_API void bgr_binning_layout(int R, size_t* offsets3);

/* ---- view-parallel exchange (new; the reference has no multi-GPU path) ---------- */
/* In-switch (NVLink multicast / NVLS) sum of floats [offset, offset+count) of a buffer that is
 * symmetrically mapped on `world` GPUs and bound to the multicast address `multicast_ptr`. Rank
 * `rank` reduces and re-broadcasts its 1/world share with multimem.ld_reduce / multimem.st. The
 * caller orders it against the other ranks with a cross-GPU barrier before and after (the kernel
 * never waits on a peer). offset_floats must be a multiple of 4. */
BGR_API int bgr_nvls_allreduce_f32(void* multicast_ptr, size_t offset_floats, size_t count_floats,
                                   int rank, int world, int max_blocks, void* stream);

/* dL_dsh[i][k][:] = sum over the `world` views v of basis_k(dir_v(i)) * factors_all[v][i][:]
 * (k < (D+1)^2, zero above), summed in view order on every rank => bit-identical replicas.
 * campos_all: [world][4], factors_all: [world][P][3] (local addresses of the symmetric buffer). */
BGR_API int bgr_sh_grad_from_factors(int P, int D, int M, int world, const float* means3D,
                                     const float* campos_all, const float* factors_all, float* dL_dsh,
                                     void* stream);
/* The same for the Gaussians [begin, end) only (pipelined exchange). */
BGR_API int bgr_sh_grad_from_factors_range(int P, int begin, int end, int D, int M, int world,
                                           const float* means3D, const float* campos_all,
                                           const float* factors_all, float* dL_dsh, void* stream);

/* bgr_nvls_allreduce_f32 over up to four disjoint ranges in one launch (the Gaussian range
 * [p0, p1) of the structure-of-arrays leaf block is four short ranges). Offsets and counts are in
 * floats from multicast_ptr and must be multiples of 4. */
BGR_API int bgr_nvls_allreduce_segments_f32(void* multicast_ptr, int nseg, const size_t* offsets_floats,
                                            const size_t* counts_floats, int rank, int world,
                                            int max_blocks, void* stream);

/* ---- diagnostics (bench.py / tests; nothing here is on the product path) --------------------- */
/* Total number of CUDA kernels this library has launched in this process. */
BGR_API unsigned long long bgr_launch_count(void);
/* What the last bgr_forward on the calling thread chose for its binning (only when
 * num_rendered > 0): [0] 1 = some tile list exceeded the shared-memory sort capacity and went
 * through the global-memory path, [1] index bits of the bucket keys, [2] depth bits (32),
 * [4] longest tile list, [5] shared-memory sort capacity in keys, [7] R; [3],[6] unused. */
BGR_API void bgr_last_forward_info(int* out8);
/* Per-stage CUDA-event timing on the launch stream: enable, run bgr_forward / bgr_backward, then
 * read the durations (ms; -1 for a stage that did not run). Stage i is named bgr_stage_name(i). */
BGR_API void bgr_stage_timing(int enable);
BGR_API int bgr_stage_count(void);
BGR_API const char* bgr_stage_name(int i);
BGR_API int bgr_stage_times(float* ms, int cap);
/* Re-traverses the forward compositing of a finished bgr_forward (its three buffers) and counts
 * stats4_host[0] staged (tile, splat) entries, [1] (pixel, splat) pairs examined, [2] pairs whose
 * exponential was evaluated, [3] pairs blended. Synchronises the stream. */
BGR_API int bgr_render_stats(int P, int R, int width, int height, const void* geom_buffer,
                             const void* binning_buffer, const void* image_buffer,
                             unsigned long long* stats4_host, void* stream);

BGR_API const char* bgr_last_error(void);
BGR_API const char* bgr_version(void);

#ifdef __cplusplus
}
#endif
#endif /* BGR_H_INCLUDED */
