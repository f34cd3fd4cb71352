// extern "C" entry points of libbgr.so (include/bgr.h) and the host-side orchestration of the
// forward / backward passes (replaces reference rasterizer_impl.cu:198-434).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/bgr.h"
#include "bgr_common.cuh"

namespace bgr {
std::atomic<unsigned long long> g_kernel_launches{0};
}

namespace {

thread_local std::string g_last_error;
// what the last bgr_forward on this thread decided (bgr_last_forward_info)
thread_local int g_forward_info[8] = {0, 0, 0, 0, 0, 0, 0, 0};

int fail(int code, const std::string& msg)
{
    g_last_error = msg;
    return code;
}

int check_cuda(cudaError_t e, const char* what)
{
    if (e == cudaSuccess) return BGR_OK;
    return fail(BGR_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

// debug mode: synchronise and surface kernel faults with the stage name (auxiliary.h:166-173)
int stage_check(bool debug, cudaStream_t s, const char* stage)
{
    cudaError_t e = cudaPeekAtLastError();
    if (e == cudaSuccess && debug) e = cudaStreamSynchronize(s);
    if (e == cudaSuccess && debug) e = cudaGetLastError();
    if (e != cudaSuccess) {
        std::fprintf(stderr, "\n[CUDA ERROR] in bgr stage '%s': %s\n", stage, cudaGetErrorString(e));
        return fail(BGR_ERR_CUDA, std::string("stage '") + stage + "': " + cudaGetErrorString(e));
    }
    return BGR_OK;
}

bgr::GeomControl* pinned_control()
{
    thread_local bgr::GeomControl* host = nullptr;
    if (host == nullptr) {
        if (cudaMallocHost(reinterpret_cast<void**>(&host), sizeof(bgr::GeomControl)) != cudaSuccess)
            host = nullptr;
    }
    return host;
}

// ---- diagnostics: per-stage CUDA-event timing (off by default) and a kernel-launch counter ------
enum Stage {
    ST_PREPROCESS = 0, ST_SCAN, ST_EMIT, ST_SORT, ST_RENDER_FWD, ST_ACC_CLEAR,
    ST_RENDER_BWD, ST_PREPROCESS_BWD, ST_COUNT
};
const char* const kStageNames[ST_COUNT] = {
    "preprocess", "tile_scan", "scatter_keys", "tile_sort", "render_forward",
    "clear_accumulators", "render_backward", "preprocess_backward"};

struct StageTimer {
    bool on = false;
    bool have_events = false;
    cudaEvent_t ev[2 * ST_COUNT];
    bool used[ST_COUNT] = {};
    void begin(int st, cudaStream_t s)
    {
        if (!on) return;
        if (!have_events) {
            for (auto& e : ev) cudaEventCreate(&e);
            have_events = true;
        }
        cudaEventRecord(ev[2 * st], s);
    }
    void end(int st, cudaStream_t s)
    {
        if (!on) return;
        cudaEventRecord(ev[2 * st + 1], s);
        used[st] = true;
    }
};
StageTimer g_timer;

bool env_flag(const char* name)
{
    const char* v = std::getenv(name);
    return v != nullptr && v[0] != '\0' && v[0] != '0';
}

}  // namespace

extern "C" {

BGR_API const char* bgr_last_error(void) { return g_last_error.c_str(); }
BGR_API const char* bgr_version(void) { return "bgr 0.1 (sm_100a)"; }

BGR_API unsigned long long bgr_launch_count(void) { return bgr::g_kernel_launches.load(std::memory_order_relaxed); }
BGR_API void bgr_last_forward_info(int* out8) { std::copy(g_forward_info, g_forward_info + 8, out8); }

BGR_API void bgr_stage_timing(int enable)
{
    g_timer.on = enable != 0;
    for (bool& u : g_timer.used) u = false;
}
BGR_API int bgr_stage_count(void) { return ST_COUNT; }
BGR_API const char* bgr_stage_name(int i) { return (i >= 0 && i < ST_COUNT) ? kStageNames[i] : ""; }
BGR_API int bgr_stage_times(float* ms, int cap)
{
    int n = 0;
    for (int i = 0; i < ST_COUNT && i < cap; ++i, ++n) {
        ms[i] = -1.0f;
        if (!g_timer.used[i]) continue;
        if (cudaEventSynchronize(g_timer.ev[2 * i + 1]) != cudaSuccess) continue;
        float t = 0.f;
        if (cudaEventElapsedTime(&t, g_timer.ev[2 * i], g_timer.ev[2 * i + 1]) == cudaSuccess) ms[i] = t;
    }
    return n;
}

BGR_API size_t bgr_geom_bytes(int P) { return bgr::geom_layout((size_t)std::max(P, 0)).total; }
BGR_API size_t bgr_image_bytes(int width, int height)
{
    return bgr::image_layout((size_t)std::max(width, 0), (size_t)std::max(height, 0)).total;
}
BGR_API size_t bgr_binning_bytes(int R) { return bgr::binning_layout((size_t)std::max(R, 0), true).total; }
BGR_API size_t bgr_backward_scratch_bytes(int P)
{
    return bgr::align_up((size_t)std::max(P, 0) * bgr::ACC_STRIDE * sizeof(float), 256);
}

BGR_API void bgr_geom_layout(int P, size_t* o)
{
    const bgr::GeomLayout L = bgr::geom_layout((size_t)P);
    o[0] = L.rec; o[1] = L.tiles;
}
BGR_API void bgr_image_layout(int width, int height, size_t* o)
{
    const bgr::ImageLayout L = bgr::image_layout((size_t)width, (size_t)height);
    o[0] = L.final_T; o[1] = L.n_contrib; o[2] = L.ranges; o[3] = L.control;
}
BGR_API void bgr_binning_layout(int R, size_t* o)
{
    const bgr::BinningLayout L = bgr::binning_layout((size_t)R, true);
    o[0] = L.point_list; o[1] = L.keys; o[2] = L.keys_alt;
}

// raw: the argument block holds the trainer's raw leaves (bgr_forward_raw, include/bgr.h)
static int forward_impl(const BgrForwardArgs* a, bool raw, const float* features_rest,
                        bgr_alloc_fn geom_alloc, void* geom_user,
                        bgr_alloc_fn binning_alloc, void* binning_user, bgr_alloc_fn image_alloc,
                        void* image_user, void* stream_v, int* num_rendered)
{
    using namespace bgr;
    cudaStream_t s = static_cast<cudaStream_t>(stream_v);
    if (a == nullptr || num_rendered == nullptr) return fail(BGR_ERR_INVALID_ARG, "null argument");
    *num_rendered = 0;
    if (a->P < 0 || a->width <= 0 || a->height <= 0)
        return fail(BGR_ERR_INVALID_ARG, "P must be >= 0 and the image non-empty");
    const bool debug = a->debug != 0;
    const size_t npix = (size_t)a->width * a->height;

    if (a->P == 0) {
        // the reference skips the core entirely: the image stays zero, NOT background
        // (rasterize_points.cu:81)
        int rc = check_cuda(cudaMemsetAsync(a->out_color, 0, 3 * npix * sizeof(float), s), "memset");
        return rc;
    }
    if (a->means3D == nullptr || a->opacities == nullptr || a->viewmatrix == nullptr ||
        a->projmatrix == nullptr || a->cam_pos == nullptr || a->background == nullptr ||
        a->out_color == nullptr || a->radii == nullptr)
        return fail(BGR_ERR_INVALID_ARG, "required pointer is NULL");
    if (a->shs == nullptr && a->colors_precomp == nullptr)
        return fail(BGR_ERR_INVALID_ARG, "provide SHs or precomputed colours");
    if (a->cov3D_precomp == nullptr && (a->scales == nullptr || a->rotations == nullptr))
        return fail(BGR_ERR_INVALID_ARG, "provide scales+rotations or a precomputed 3D covariance");
    if (a->shs != nullptr && a->colors_precomp == nullptr &&
        (a->D < 0 || a->D > 3 || (a->D + 1) * (a->D + 1) > a->M))
        return fail(BGR_ERR_INVALID_ARG, "SH degree does not fit the coefficient count");
    if (raw && (a->shs == nullptr || a->colors_precomp != nullptr || a->cov3D_precomp != nullptr ||
                a->M < 1 || a->M > 16 || (a->M > 1 && features_rest == nullptr)))
        return fail(BGR_ERR_INVALID_ARG,
                    "raw-leaf entry: needs _features_dc (+ _features_rest when M > 1, M <= 16), _scaling, "
                    "_rotation, _opacity and no precomputed colours / covariance");

    const int grid_x = (a->width + TILE_X - 1) / TILE_X;
    const int grid_y = (a->height + TILE_Y - 1) / TILE_Y;
    const uint32_t num_tiles = (uint32_t)grid_x * (uint32_t)grid_y;
    const float focal_y = a->height / (2.0f * a->tan_fovy);
    const float focal_x = a->width / (2.0f * a->tan_fovx);

    const GeomLayout GL = geom_layout((size_t)a->P);
    const ImageLayout IL = image_layout((size_t)a->width, (size_t)a->height);
    unsigned char* geom = static_cast<unsigned char*>(geom_alloc(GL.total, geom_user));
    unsigned char* img = static_cast<unsigned char*>(image_alloc(IL.total, image_user));
    if (geom == nullptr || img == nullptr) return fail(BGR_ERR_ALLOC, "scratch allocation failed");

    GaussRec* rec = reinterpret_cast<GaussRec*>(geom + GL.rec);
    uint32_t* tiles = reinterpret_cast<uint32_t*>(geom + GL.tiles);
    float* final_T = reinterpret_cast<float*>(img + IL.final_T);
    uint32_t* n_contrib = reinterpret_cast<uint32_t*>(img + IL.n_contrib);
    uint2* ranges = reinterpret_cast<uint2*>(img + IL.ranges);
    uint32_t* cursor = reinterpret_cast<uint32_t*>(img + IL.cursor);
    GeomControl* ctl = reinterpret_cast<GeomControl*>(img + IL.control);
    int* coverage = reinterpret_cast<int*>(img + IL.grid);

    int rc;
    // control block (counters) + coverage difference grid: adjacent, one memset
    if ((rc = check_cuda(cudaMemsetAsync(ctl, 0, IL.total - IL.control, s), "memset control")))
        return rc;

    PreprocessParams pp;
    pp.P = a->P; pp.D = a->D; pp.M = a->M; pp.W = a->width; pp.H = a->height;
    pp.means3D = a->means3D; pp.scales = a->scales; pp.scale_modifier = a->scale_modifier;
    pp.rotations = a->rotations; pp.opacities = a->opacities; pp.shs = a->shs;
    pp.cov3D_precomp = a->cov3D_precomp; pp.colors_precomp = a->colors_precomp;
    pp.viewmatrix = a->viewmatrix; pp.projmatrix = a->projmatrix; pp.cam_pos = a->cam_pos;
    pp.tan_fovx = a->tan_fovx; pp.tan_fovy = a->tan_fovy; pp.focal_x = focal_x; pp.focal_y = focal_y;
    pp.raw = raw ? 1 : 0; pp.shs_rest = features_rest;
    pp.grid_x = grid_x; pp.grid_y = grid_y; pp.prefiltered = a->prefiltered;
    pp.radii = a->radii; pp.rec = rec; pp.tiles_touched = tiles; pp.control = ctl;
    pp.coverage_grid = coverage;
    g_timer.begin(ST_PREPROCESS, s);
    launch_preprocess(pp, s);
    g_timer.end(ST_PREPROCESS, s);
    if ((rc = stage_check(debug, s, "preprocess"))) return rc;

    g_timer.begin(ST_SCAN, s);
    if (launch_tile_scan(grid_x, grid_y, coverage, ranges, cursor, ctl, s) != 0)
        return fail(BGR_ERR_CUDA, "tile scan launch configuration rejected");
    g_timer.end(ST_SCAN, s);
    if ((rc = stage_check(debug, s, "tile scan"))) return rc;

    GeomControl* host = pinned_control();
    if (host == nullptr) return fail(BGR_ERR_CUDA, "pinned host allocation failed");
    if ((rc = check_cuda(cudaMemcpyAsync(host, ctl, sizeof(GeomControl), cudaMemcpyDeviceToHost, s),
                         "read back num_rendered")))
        return rc;
    if ((rc = check_cuda(cudaStreamSynchronize(s), "sync after the tile scan"))) return rc;

    // the 64-bit total decides: a 32-bit count wraps at 2^32 (the reference overflows silently,
    // rasterizer_impl.cu:280-283)
    if (host->tiles_total >= (1ull << 30))
        return fail(BGR_ERR_OVERFLOW, "num_rendered = " + std::to_string(host->tiles_total) +
                                          " >= 2^30 tile instances is not supported");
    const uint32_t R = host->num_rendered;
    if ((unsigned long long)R != host->tiles_total)
        return fail(BGR_ERR_CUDA, "internal error: tile scan total differs from the sum of tiles_touched");

    // shared-memory capacity of the per-tile sort, from the longest list of this view
    uint32_t smem_keys = std::min<uint32_t>(std::max<uint32_t>(host->max_tile_list, 128u),
                                            TILE_SORT_MAX_SMEM_KEYS);
    if (const char* cap = std::getenv("BGR_TILE_SORT_SMEM_KEYS")) {   // tests: force the global path
        const long v = std::atol(cap);
        if (v >= 64) smem_keys = std::min<uint32_t>((uint32_t)v, TILE_SORT_MAX_SMEM_KEYS);
    }
    smem_keys = (smem_keys + 1u) & ~1u;
    const bool long_lists = host->max_tile_list > smem_keys;

    // the alternate key array backs the general (global-memory) path of the per-tile sort, which
    // a tile takes when its list does not fit shared memory or its depths are badly clustered
    const BinningLayout BL = binning_layout((size_t)R, true);
    unsigned char* bin = nullptr;
    uint32_t* point_list = nullptr;

    if (R > 0) {
        bin = static_cast<unsigned char*>(binning_alloc(BL.total, binning_user));
        if (bin == nullptr) return fail(BGR_ERR_ALLOC, "binning allocation failed");
        point_list = reinterpret_cast<uint32_t*>(bin + BL.point_list);
        const int nb_idx = bits_for((uint64_t)(a->P - 1));
        uint64_t* keys = reinterpret_cast<uint64_t*>(bin + BL.keys);
        uint64_t* keys_alt = reinterpret_cast<uint64_t*>(bin + BL.keys_alt);
        const int info[8] = {long_lists ? 1 : 0, nb_idx, 32, 0, (int)host->max_tile_list, (int)smem_keys,
                             0, (int)R};
        std::copy(info, info + 8, g_forward_info);

        g_timer.begin(ST_EMIT, s);
        launch_scatter_keys(a->P, rec, a->radii, grid_x, grid_y, nb_idx, cursor, keys, s);
        g_timer.end(ST_EMIT, s);
        if ((rc = stage_check(debug, s, "scatter keys"))) return rc;

        g_timer.begin(ST_SORT, s);
        if (launch_tile_sort(num_tiles, ranges, keys, keys_alt, nb_idx, smem_keys, point_list, s) != 0)
            return fail(BGR_ERR_CUDA, "tile sort launch configuration rejected");
        g_timer.end(ST_SORT, s);
        if ((rc = stage_check(debug, s, "tile sort"))) return rc;
    }

    g_timer.begin(ST_RENDER_FWD, s);
    launch_render_forward(a->width, a->height, ranges, point_list, rec, nullptr, a->background,
                          final_T, n_contrib, a->out_color, s);
    g_timer.end(ST_RENDER_FWD, s);
    if ((rc = stage_check(debug, s, "render forward"))) return rc;

    *num_rendered = (int)R;
    return BGR_OK;
}

BGR_API int bgr_forward(const BgrForwardArgs* a, bgr_alloc_fn geom_alloc, void* geom_user,
                bgr_alloc_fn binning_alloc, void* binning_user, bgr_alloc_fn image_alloc,
                void* image_user, void* stream_v, int* num_rendered)
{
    return forward_impl(a, false, nullptr, geom_alloc, geom_user, binning_alloc, binning_user, image_alloc,
                        image_user, stream_v, num_rendered);
}

BGR_API int bgr_forward_raw(const BgrForwardArgs* a, const float* features_rest, bgr_alloc_fn geom_alloc,
                void* geom_user, bgr_alloc_fn binning_alloc, void* binning_user,
                bgr_alloc_fn image_alloc, void* image_user, void* stream_v, int* num_rendered)
{
    return forward_impl(a, true, features_rest, geom_alloc, geom_user, binning_alloc, binning_user,
                        image_alloc, image_user, stream_v, num_rendered);
}

static int backward_impl(const BgrBackwardArgs* a, bool raw, const float* features_rest,
                         float* dL_dfeatures_rest, void* stream_v)
{
    using namespace bgr;
    cudaStream_t s = static_cast<cudaStream_t>(stream_v);
    if (a == nullptr) return fail(BGR_ERR_INVALID_ARG, "null argument");
    if (a->P < 0 || a->width <= 0 || a->height <= 0 || a->R < 0)
        return fail(BGR_ERR_INVALID_ARG, "bad sizes");
    if (a->P == 0) return BGR_OK;
    if (a->geom_buffer == nullptr || a->image_buffer == nullptr || a->scratch == nullptr ||
        a->dL_dpix == nullptr || a->radii == nullptr || a->means3D == nullptr)
        return fail(BGR_ERR_INVALID_ARG, "required pointer is NULL");
    if (a->R > 0 && a->binning_buffer == nullptr)
        return fail(BGR_ERR_INVALID_ARG, "binning buffer is NULL");
    if (raw && (a->shs == nullptr || a->dL_dsh == nullptr || a->scales == nullptr || a->rotations == nullptr ||
                a->colors_precomp != nullptr || a->cov3D_precomp != nullptr || a->M < 1 || a->M > 16 ||
                (a->M > 1 && (features_rest == nullptr || dL_dfeatures_rest == nullptr)) ||
                a->sh_factor_multicast != nullptr))
        return fail(BGR_ERR_INVALID_ARG,
                    "raw-leaf entry: needs _features_dc (+ _features_rest when M > 1, M <= 16) and their gradient "
                    "arrays, _scaling, _rotation, no precomputed colours / covariance and no multicast slots");
    const bool debug = a->debug != 0;

    const GeomLayout GL = geom_layout((size_t)a->P);
    const ImageLayout IL = image_layout((size_t)a->width, (size_t)a->height);
    const BinningLayout BL = binning_layout((size_t)a->R);
    const unsigned char* geom = static_cast<const unsigned char*>(a->geom_buffer);
    const unsigned char* img = static_cast<const unsigned char*>(a->image_buffer);
    const unsigned char* bin = static_cast<const unsigned char*>(a->binning_buffer);
    const GaussRec* rec = reinterpret_cast<const GaussRec*>(geom + GL.rec);
    const float* final_T = reinterpret_cast<const float*>(img + IL.final_T);
    const uint32_t* n_contrib = reinterpret_cast<const uint32_t*>(img + IL.n_contrib);
    const uint2* ranges = reinterpret_cast<const uint2*>(img + IL.ranges);
    float* acc = static_cast<float*>(a->scratch);

    // stages == 0: everything. Otherwise the caller splits the call: ACCUMULATE once, then
    // PREPROCESS over consecutive Gaussian ranges, so that it can exchange the gradients of one
    // range while the next is being computed.
    const bool do_accumulate = a->stages == 0 || (a->stages & BGR_BACKWARD_ACCUMULATE) != 0;
    const bool do_preprocess = a->stages == 0 || (a->stages & BGR_BACKWARD_PREPROCESS) != 0;
    int begin = 0, end = a->P;
    if (a->stages != 0 && do_preprocess) {
        begin = a->gaussian_begin;
        end = a->gaussian_end;
        if (begin < 0 || end > a->P || begin > end)
            return fail(BGR_ERR_INVALID_ARG, "bad Gaussian range");
    }

    int rc;
    if (do_accumulate) {
    g_timer.begin(ST_ACC_CLEAR, s);
    if ((rc = check_cuda(cudaMemsetAsync(acc, 0, (size_t)a->P * ACC_STRIDE * sizeof(float), s),
                         "memset accumulators")))
        return rc;
    g_timer.end(ST_ACC_CLEAR, s);
    }

    if (do_accumulate && a->R > 0) {
        const uint32_t* point_list = reinterpret_cast<const uint32_t*>(bin + BL.point_list);
        g_timer.begin(ST_RENDER_BWD, s);
        launch_render_backward(a->width, a->height, ranges, point_list, rec, nullptr, a->background,
                               final_T, n_contrib, a->dL_dpix, acc, s);
        g_timer.end(ST_RENDER_BWD, s);
        if ((rc = stage_check(debug, s, "render backward"))) return rc;
    }

    if (!do_preprocess) return BGR_OK;
    PreprocessBackwardParams pb;
    pb.begin = begin; pb.end = end;
    pb.P = a->P; pb.D = a->D; pb.M = a->M; pb.W = a->width; pb.H = a->height;
    pb.means3D = a->means3D; pb.radii = a->radii; pb.shs = a->shs; pb.scales = a->scales;
    pb.rotations = a->rotations; pb.scale_modifier = a->scale_modifier;
    pb.cov3D_precomp = a->cov3D_precomp; pb.colors_precomp = a->colors_precomp;
    pb.viewmatrix = a->viewmatrix; pb.projmatrix = a->projmatrix; pb.cam_pos = a->cam_pos;
    pb.focal_y = a->height / (2.0f * a->tan_fovy);
    pb.focal_x = a->width / (2.0f * a->tan_fovx);
    pb.tan_fovx = a->tan_fovx; pb.tan_fovy = a->tan_fovy;
    pb.rec = rec; pb.acc = acc;
    pb.dL_dmean2D = a->dL_dmean2D; pb.dL_dcolor = a->dL_dcolor; pb.dL_dopacity = a->dL_dopacity;
    pb.dL_dmean3D = a->dL_dmean3D; pb.dL_dcov3D = a->dL_dcov3D; pb.dL_dsh = a->dL_dsh;
    pb.dL_dscale = a->dL_dscale; pb.dL_drot = a->dL_drot;
    pb.factor_mc = a->sh_factor_multicast; pb.campos_mc = a->campos_multicast;
    pb.raw = raw ? 1 : 0; pb.shs_rest = features_rest; pb.dL_dsh_rest = dL_dfeatures_rest;
    if (pb.factor_mc != nullptr) {
        if (pb.shs == nullptr || pb.campos_mc == nullptr || (a->P & 3) != 0 || (begin & 3) != 0 ||
            (end & 3) != 0 || (reinterpret_cast<uintptr_t>(pb.factor_mc) & 15) != 0)
            return fail(BGR_ERR_INVALID_ARG,
                        "SH factor multicast needs SHs, a camera slot, P and the range bounds % 4 == 0 "
                        "and a 16-byte aligned slot");
    } else if (pb.shs != nullptr && pb.dL_dsh == nullptr) {
        return fail(BGR_ERR_INVALID_ARG, "dL_dsh is NULL although SHs were given");
    }
    g_timer.begin(ST_PREPROCESS_BWD, s);
    launch_preprocess_backward(pb, s);
    g_timer.end(ST_PREPROCESS_BWD, s);
    if ((rc = stage_check(debug, s, "preprocess backward"))) return rc;
    return BGR_OK;
}

BGR_API int bgr_backward(const BgrBackwardArgs* a, void* stream_v)
{
    return backward_impl(a, false, nullptr, nullptr, stream_v);
}

BGR_API int bgr_backward_raw(const BgrBackwardArgs* a, const float* features_rest, float* dL_dfeatures_rest,
                             void* stream_v)
{
    return backward_impl(a, true, features_rest, dL_dfeatures_rest, stream_v);
}

BGR_API size_t bgr_loss_scratch_bytes(int width, int height)
{
    return (width > 0 && height > 0) ? bgr::loss_scratch_bytes(width, height) : 0;
}

BGR_API int bgr_l1_ssim_loss(const float* image, const float* gt, int width, int height, float lambda_dssim,
                             void* scratch, float* loss3, float* dL_dimage, void* stream_v)
{
    cudaStream_t s = static_cast<cudaStream_t>(stream_v);
    if (bgr::launch_l1_ssim_loss(image, gt, width, height, lambda_dssim, scratch, loss3, dL_dimage, s) != 0)
        return fail(BGR_ERR_INVALID_ARG, "loss: NULL pointer or empty image");
    return stage_check(false, s, "l1 + ssim loss");
}

BGR_API int bgr_densification_stats(int P, const int* radii, const float* dL_dmean2D, float* max_radii2D,
                                    float* xyz_gradient_accum, float* denom, void* stream_v)
{
    cudaStream_t s = static_cast<cudaStream_t>(stream_v);
    if (bgr::launch_densification_stats(P, radii, dL_dmean2D, max_radii2D, xyz_gradient_accum, denom, s) != 0)
        return fail(BGR_ERR_INVALID_ARG, "densification stats: P < 0 or NULL array");
    return stage_check(false, s, "densification stats");
}

BGR_API int bgr_adam_step(int n_groups, float* const* params, const float* const* grads, float* const* exp_avg,
                          float* const* exp_avg_sq, const size_t* numel, const double* lr, double beta1,
                          double beta2, double eps, int step, void* stream_v)
{
    cudaStream_t s = static_cast<cudaStream_t>(stream_v);
    if (params == nullptr || grads == nullptr || exp_avg == nullptr || exp_avg_sq == nullptr || numel == nullptr ||
        lr == nullptr ||
        bgr::launch_adam(n_groups, params, grads, exp_avg, exp_avg_sq, numel, lr, beta1, beta2, eps, step, s) != 0)
        return fail(BGR_ERR_INVALID_ARG, "adam: 1..8 groups with non-NULL arrays and step >= 1");
    return stage_check(false, s, "adam step");
}

BGR_API int bgr_mark_visible(int P, const float* means3D, const float* viewmatrix,
                     const float* /*projmatrix*/, uint8_t* present, void* stream_v)
{
    cudaStream_t s = static_cast<cudaStream_t>(stream_v);
    if (P < 0) return fail(BGR_ERR_INVALID_ARG, "P < 0");
    if (P == 0) return BGR_OK;
    if (means3D == nullptr || viewmatrix == nullptr || present == nullptr)
        return fail(BGR_ERR_INVALID_ARG, "required pointer is NULL");
    bgr::launch_mark_visible(P, means3D, viewmatrix, present, s);
    return stage_check(false, s, "mark visible");
}

BGR_API int bgr_render_stats(int P, int R, int width, int height, const void* geom_buffer,
                             const void* binning_buffer, const void* image_buffer,
                             unsigned long long* stats4_host, void* stream_v)
{
    using namespace bgr;
    cudaStream_t s = static_cast<cudaStream_t>(stream_v);
    if (stats4_host == nullptr || P < 0 || R < 0 || width <= 0 || height <= 0)
        return fail(BGR_ERR_INVALID_ARG, "bad argument");
    for (int i = 0; i < 4; ++i) stats4_host[i] = 0;
    if (P == 0 || R == 0) return BGR_OK;
    if (geom_buffer == nullptr || binning_buffer == nullptr || image_buffer == nullptr)
        return fail(BGR_ERR_INVALID_ARG, "required pointer is NULL");
    const GeomLayout GL = geom_layout((size_t)P);
    const ImageLayout IL = image_layout((size_t)width, (size_t)height);
    const BinningLayout BL = binning_layout((size_t)R);
    const unsigned char* geom = static_cast<const unsigned char*>(geom_buffer);
    const unsigned char* img = static_cast<const unsigned char*>(image_buffer);
    const unsigned char* bin = static_cast<const unsigned char*>(binning_buffer);
    unsigned long long* dev = nullptr;
    int rc;
    if ((rc = check_cuda(cudaMalloc(reinterpret_cast<void**>(&dev), 4 * sizeof(unsigned long long)), "cudaMalloc"))) return rc;
    if ((rc = check_cuda(cudaMemsetAsync(dev, 0, 4 * sizeof(unsigned long long), s), "memset stats"))) {
        cudaFree(dev);
        return rc;
    }
    launch_render_stats(width, height, reinterpret_cast<const uint2*>(img + IL.ranges),
                        reinterpret_cast<const uint32_t*>(bin + BL.point_list),
                        reinterpret_cast<const GaussRec*>(geom + GL.rec), dev, s);
    rc = check_cuda(cudaMemcpyAsync(stats4_host, dev, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s),
                    "copy stats");
    if (rc == 0) rc = check_cuda(cudaStreamSynchronize(s), "render stats");
    cudaFree(dev);
    return rc;
}

BGR_API int bgr_nvls_allreduce_f32(void* multicast_ptr, size_t offset_floats, size_t count_floats,
                                   int rank, int world, int max_blocks, void* stream_v)
{
    cudaStream_t s = static_cast<cudaStream_t>(stream_v);
    if (bgr::launch_nvls_allreduce(multicast_ptr, offset_floats, count_floats, rank, world,
                                   max_blocks, s) != 0)
        return fail(BGR_ERR_INVALID_ARG, "bad NVLS all-reduce arguments (NULL multicast pointer, rank, or unaligned offset)");
    return stage_check(false, s, "nvls all-reduce");
}

BGR_API int bgr_nvls_allreduce_segments_f32(void* multicast_ptr, int nseg, const size_t* offsets_floats,
                                            const size_t* counts_floats, int rank, int world,
                                            int max_blocks, void* stream_v)
{
    cudaStream_t s = static_cast<cudaStream_t>(stream_v);
    if (nseg > 0 && (offsets_floats == nullptr || counts_floats == nullptr))
        return fail(BGR_ERR_INVALID_ARG, "NULL segment table");
    if (bgr::launch_nvls_allreduce_segments(multicast_ptr, nseg, offsets_floats, counts_floats, rank,
                                            world, max_blocks, s) != 0)
        return fail(BGR_ERR_INVALID_ARG, "bad NVLS segment all-reduce arguments (NULL multicast pointer, rank, "
                                         "more than 4 segments, or offsets/counts not multiples of 4 floats)");
    return stage_check(false, s, "nvls segment all-reduce");
}

BGR_API int bgr_sh_grad_from_factors_range(int P, int begin, int end, int D, int M, int world,
                                           const float* means3D, const float* campos_all,
                                           const float* factors_all, float* dL_dsh, void* stream_v)
{
    cudaStream_t s = static_cast<cudaStream_t>(stream_v);
    if (P < 0 || D < 0 || D > 3 || M < (D + 1) * (D + 1) || world <= 0 || begin < 0 || end > P ||
        begin > end)
        return fail(BGR_ERR_INVALID_ARG, "bad sizes");
    if (begin == end) return BGR_OK;
    if (means3D == nullptr || campos_all == nullptr || factors_all == nullptr || dL_dsh == nullptr)
        return fail(BGR_ERR_INVALID_ARG, "required pointer is NULL");
    bgr::launch_sh_grad_from_factors(P, begin, end, D, M, world, means3D, campos_all, factors_all,
                                     dL_dsh, s);
    return stage_check(false, s, "sh gradient from factors");
}

BGR_API int bgr_sh_grad_from_factors(int P, int D, int M, int world, const float* means3D,
                                     const float* campos_all, const float* factors_all, float* dL_dsh,
                                     void* stream_v)
{
    return bgr_sh_grad_from_factors_range(P, 0, P, D, M, world, means3D, campos_all, factors_all,
                                          dL_dsh, stream_v);
}

}  // extern "C"
