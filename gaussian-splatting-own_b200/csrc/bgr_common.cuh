// Shared declarations for the sm_100a rasterizer kernels (internal; the public ABI is include/bgr.h).
#pragma once
#include <atomic>
#include <cstdint>
#include <cstddef>
#include <cuda_runtime.h>

namespace bgr {

constexpr int TILE_X = 16;                 // config.h:16-17 — part of the bit-exact integer contract
constexpr int TILE_Y = 16;
constexpr int TILE_PIXELS = TILE_X * TILE_Y;

// ---------------------------------------------------------------------------------------------
// Per-Gaussian record written by the forward preprocess and gathered by the render kernels.
// 48 bytes = three float4, so one gather touches 1.5–2 sectors instead of 3+ for separate arrays.
//   q0 = {mean2D.x, mean2D.y, conic.x, conic.y}
//   q1 = {conic.z, opacity, r, g}
//   q2 = {b, thr, depth, clamp bits}
// thr = -(ln(255*opacity) + 1e-4): power < thr  =>  opacity*exp(power) < 1/255 with a safe margin.
// ---------------------------------------------------------------------------------------------
struct __align__(16) GaussRec {
    float4 q0, q1, q2;
};
static_assert(sizeof(GaussRec) == 48, "record layout");

// Control block of one forward call (device side; lives in the image buffer next to the tile
// counters so that ONE memset clears it together with the coverage grid).
struct GeomControl {
    uint32_t num_rendered;           // total tile instances R (written by the tile scan)
    uint32_t max_tile_list;          // longest per-tile list (sizes the sort's shared memory)
    uint32_t visible;                // number of Gaussians with radii > 0
    uint32_t pad;
    unsigned long long tiles_total;  // 64-bit sum of tiles_touched: the overflow guard
    unsigned long long pad2;
};
static_assert(sizeof(GeomControl) == 32, "control block layout");

struct GeomLayout {
    size_t rec, tiles, total;
};
struct ImageLayout {
    size_t final_T, n_contrib, ranges, cursor, control, grid, total;
};
struct BinningLayout {
    size_t point_list, keys, keys_alt, total;
};

// Number of kernels this library has launched (diagnostic; bumped by every launch_* wrapper).
extern std::atomic<unsigned long long> g_kernel_launches;   // several host threads may launch

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

inline GeomLayout geom_layout(size_t P)
{
    GeomLayout L;
    size_t o = 0;
    L.rec = o;        o = align_up(o + P * sizeof(GaussRec), 256);
    L.tiles = o;      o = align_up(o + P * 4, 256);
    L.total = o;
    return L;
}

// Image buffer. `grid` is the (gx+1) x (gy+1) coverage difference grid the preprocess adds the
// four corners of every tile rectangle to; its 2-D prefix sum is the per-tile instance count.
inline ImageLayout image_layout(size_t W, size_t H)
{
    ImageLayout L;
    const size_t gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    const size_t tiles = gx * gy;
    size_t o = 0;
    L.final_T = o;   o = align_up(o + W * H * 4, 256);
    L.n_contrib = o; o = align_up(o + W * H * 4, 256);
    L.ranges = o;    o = align_up(o + tiles * 8, 256);
    L.cursor = o;    o = align_up(o + tiles * 4, 256);
    L.control = o;   o = align_up(o + sizeof(GeomControl), 256);
    L.grid = o;      o = align_up(o + (gx + 1) * (gy + 1) * 4, 256);
    L.total = o;
    return L;
}

// Binning buffer: the sorted Gaussian-id list the render kernels read, the per-tile key buckets
// the scatter fills, and the second key array the per-tile sort's general (global-memory) path
// ping-pongs with.
inline BinningLayout binning_layout(size_t R, bool with_alt = false)
{
    BinningLayout L;
    size_t o = 0;
    L.point_list = o; o = align_up(o + R * 4, 256);
    L.keys = o;       o = align_up(o + R * 8, 256);
    L.keys_alt = o;   o = align_up(o + (with_alt ? R * 8 : 0), 256);
    L.total = o;
    return L;
}

// Per-tile sort geometry (binning.cu): tiles whose list fits TILE_SORT_MAX_SMEM_KEYS keys are
// sorted in shared memory, longer ones through global memory.
constexpr uint32_t TILE_SORT_MAX_SMEM_KEYS = 12288;   // 2 x 96 KB of keys + 34 KB of counters

// Per-Gaussian accumulators of the backward render (AoS, 12 floats), with u = G * dL/dalpha and
// (dx, dy) = splat centre - pixel, summed over the pixels the splat was blended into:
//   [0] sum u (cx dx + cy dy)   [1] sum u (cz dy + cy dx)   (cx, cy, cz = conic)
//   [2] sum u dx dx   [3] sum u dx dy   [4] sum u dy dy
//   [5] sum u (= dL/dopacity)     [6..8] dL/dcolor  [9..11] unused
constexpr int ACC_STRIDE = 12;

// ---------------------------------------------------------------------------------------------
// Instance key inside a tile bucket: depth_bits << nb_idx | index. The tile is implied by the
// bucket, so sorting a bucket by this key is the order of a stable sort on the reference key
// tile << 32 | depth_bits (ties in depth keep emission order = ascending index,
// rasterizer_impl.cu:98-108). 32 depth bits + at most 31 index bits always fit 64.
// ---------------------------------------------------------------------------------------------
inline int bits_for(uint64_t max_value)
{
    int b = 0;
    while (max_value) { ++b; max_value >>= 1; }
    return b;
}

// ---------------------------------------------------------------------------------------------
// Kernel launch prototypes (definitions in the .cu files)
// ---------------------------------------------------------------------------------------------
struct PreprocessParams {
    int P, D, M, W, H;
    const float* means3D; const float* scales; float scale_modifier; const float* rotations;
    const float* opacities; const float* shs; const float* cov3D_precomp;
    const float* colors_precomp; const float* viewmatrix; const float* projmatrix;
    const float* cam_pos; float tan_fovx, tan_fovy, focal_x, focal_y;
    int grid_x, grid_y; int prefiltered;
    int* radii; GaussRec* rec; uint32_t* tiles_touched; GeomControl* control;
    int* coverage_grid;   // (grid_x + 1) x (grid_y + 1) difference grid, zero on entry
    // raw-leaf entry (bgr_forward_raw): scales / rotations / opacities are pre-activation, shs is
    // _features_dc [P,1,3] and shs_rest is _features_rest [P,M-1,3]
    int raw; const float* shs_rest;
};

void launch_preprocess(const PreprocessParams& p, cudaStream_t s);
void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present,
                         cudaStream_t s);
int  launch_tile_scan(int grid_x, int grid_y, int* coverage_grid, uint2* ranges, uint32_t* cursor,
                      GeomControl* control, cudaStream_t s);
void launch_scatter_keys(int P, const GaussRec* rec, const int* radii, int grid_x, int grid_y,
                         int nb_idx, uint32_t* cursor, uint64_t* keys, cudaStream_t s);
// smem_keys: capacity (keys) of the shared-memory path chosen by the caller from max_tile_list;
// keys_alt may be NULL when no tile exceeds it.
int  launch_tile_sort(uint32_t num_tiles, const uint2* ranges, uint64_t* keys, uint64_t* keys_alt,
                      int nb_idx, uint32_t smem_keys, uint32_t* point_list, cudaStream_t s);

void launch_render_forward(int W, int H, const uint2* ranges, const uint32_t* point_list,
                           const GaussRec* rec, const float* colors_precomp, const float* bg,
                           float* final_T, uint32_t* n_contrib, float* out_color, cudaStream_t s);
void launch_render_stats(int W, int H, const uint2* ranges, const uint32_t* point_list,
                         const GaussRec* rec, unsigned long long* stats, cudaStream_t s);
void launch_render_backward(int W, int H, const uint2* ranges, const uint32_t* point_list,
                            const GaussRec* rec, const float* colors_precomp, const float* bg,
                            const float* final_T, const uint32_t* n_contrib, const float* dL_dpix,
                            float* acc, cudaStream_t s);

int launch_nvls_allreduce(void* multicast_ptr, size_t offset, size_t count, int rank, int world,
                          int max_blocks, cudaStream_t s);

int launch_nvls_allreduce_segments(void* multicast_ptr, int nseg, const size_t* offsets,
                                   const size_t* counts, int rank, int world, int max_blocks,
                                   cudaStream_t s);

struct PreprocessBackwardParams {
    int P, D, M, W, H;
    const float* means3D; const int* radii; const float* shs; const float* scales;
    const float* rotations; float scale_modifier; const float* cov3D_precomp;
    const float* colors_precomp;
    const float* viewmatrix; const float* projmatrix; const float* cam_pos;
    float focal_x, focal_y, tan_fovx, tan_fovy;
    const GaussRec* rec; const float* acc;
    float* dL_dmean2D; float* dL_dcolor; float* dL_dopacity; float* dL_dmean3D; float* dL_dcov3D;
    float* dL_dsh; float* dL_dscale; float* dL_drot;
    // view-parallel path: multicast address of this rank's factor slot [P][3] and of its camera
    // position slot [4] (nullptr => write dL_dsh normally)
    float* factor_mc; float* campos_mc;
    // Gaussian range [begin, end) handled by this launch (the whole scene unless the caller
    // pipelines the kernel against the gradient exchange chunk by chunk)
    int begin, end;
    // raw-leaf entry (bgr_backward_raw): see PreprocessParams; dL_dsh is then [P,1,3] and
    // dL_dsh_rest [P,M-1,3], and dL_dscale / dL_drot / dL_dopacity are w.r.t. the raw leaves
    int raw; const float* shs_rest; float* dL_dsh_rest;
};
void launch_preprocess_backward(const PreprocessBackwardParams& p, cudaStream_t s);
void launch_sh_grad_from_factors(int P, int begin, int end, int D, int M, int world,
                                 const float* means3D, const float* campos_all,
                                 const float* factors_all, float* dL_dsh, cudaStream_t s);


// ---- the step after the path (train_step.cu) ------------------------------
size_t loss_scratch_bytes(int W, int H);
int launch_l1_ssim_loss(const float* image, const float* gt, int W, int H, float lambda, void* scratch,
                        float* loss3, float* dL_dimage, cudaStream_t s);
int launch_densification_stats(int P, const int* radii, const float* dL_dmean2D, float* max_radii2D,
                               float* grad_accum, float* denom, cudaStream_t s);
int launch_adam(int count, float* const* params, const float* const* grads, float* const* exp_avg,
                float* const* exp_avg_sq, const size_t* numel, const double* lr, double beta1, double beta2,
                double eps, int step, cudaStream_t s);

}  // namespace bgr
