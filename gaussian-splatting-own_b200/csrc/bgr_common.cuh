// Shared declarations for the sm_100a rasterizer kernels (internal; the public ABI is include/bgr.h).
#pragma once
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

// Control block at the tail of the geometry buffer (device side).
struct GeomControl {
    uint32_t num_rendered;   // total tile instances (written by the scan)
    uint32_t depth_min_bits; // max over visible Gaussians of ~float_as_uint(depth) (inverted so 0-init works)
    uint32_t depth_max_bits; // max over visible Gaussians of float_as_uint(depth)
    uint32_t visible;        // number of Gaussians with radii > 0
    uint32_t scan_ticket;    // dynamic block id for the look-back scan
    uint32_t pad;
    unsigned long long tiles_total;  // 64-bit sum of tiles_touched: the u32 scan total may wrap
};
static_assert(sizeof(GeomControl) == 32, "control block layout");

struct GeomLayout {
    size_t rec, tiles, offsets, control, scan_state, total;
};
struct ImageLayout {
    size_t final_T, n_contrib, ranges, total;
};
struct BinningLayout {
    size_t point_list, keys_a, keys_b, vals_a, vals_b, sort_temp, total;
};

// Number of kernels this library has launched (diagnostic; bumped by every launch_* wrapper).
extern unsigned long long g_kernel_launches;

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

inline GeomLayout geom_layout(size_t P)
{
    GeomLayout L;
    size_t o = 0;
    L.rec = o;        o = align_up(o + P * sizeof(GaussRec), 256);
    L.tiles = o;      o = align_up(o + P * 4, 256);
    L.offsets = o;    o = align_up(o + P * 4, 256);
    L.control = o;    o = align_up(o + sizeof(GeomControl), 256);
    L.scan_state = o; o = align_up(o + ((P + SCAN_TILE - 1) / SCAN_TILE + 1) * 8, 256);
    L.total = o;
    return L;
}

inline ImageLayout image_layout(size_t W, size_t H)
{
    ImageLayout L;
    size_t tiles = ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    size_t o = 0;
    L.final_T = o;   o = align_up(o + W * H * 4, 256);
    L.n_contrib = o; o = align_up(o + W * H * 4, 256);
    L.ranges = o;    o = align_up(o + tiles * 8, 256);
    L.total = o;
    return L;
}

// ---------------------------------------------------------------------------------------------
// Radix sort geometry
// ---------------------------------------------------------------------------------------------
constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 16;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;
constexpr int SORT_MAX_PASSES = 8;

struct SortTempLayout {
    size_t hist;      // u32[SORT_MAX_PASSES][RADIX]   global digit histograms -> exclusive bases
    size_t tickets;   // u32[SORT_MAX_PASSES]
    size_t status;    // u32[SORT_MAX_PASSES][nblocks][RADIX] decoupled look-back state
    size_t total;
    size_t nblocks;
};

inline SortTempLayout sort_temp_layout(size_t n)
{
    SortTempLayout L;
    L.nblocks = (n + SORT_TILE - 1) / SORT_TILE;
    if (L.nblocks == 0) L.nblocks = 1;
    size_t o = 0;
    L.hist = o;    o = align_up(o + SORT_MAX_PASSES * RADIX * 4, 256);
    L.tickets = o; o = align_up(o + SORT_MAX_PASSES * 4, 256);
    L.status = o;  o = align_up(o + (size_t)SORT_MAX_PASSES * L.nblocks * RADIX * 4, 256);
    L.total = o;
    return L;
}

inline BinningLayout binning_layout(size_t R)
{
    BinningLayout L;
    size_t o = 0;
    L.point_list = o; o = align_up(o + R * 4, 256);
    L.keys_a = o;     o = align_up(o + R * 8, 256);
    L.keys_b = o;     o = align_up(o + R * 8, 256);
    L.vals_a = o;     o = align_up(o + R * 4, 256);
    L.vals_b = o;     o = align_up(o + R * 4, 256);
    L.sort_temp = o;  o = align_up(o + sort_temp_layout(R).total, 256);
    L.total = o;
    return L;
}

// Per-Gaussian accumulators of the backward render (AoS, 12 floats):
//   [0] dL/dmean2D.x  [1] dL/dmean2D.y  [2] dL/dconic.x  [3] dL/dconic.y
//   [4] dL/dconic.w   [5] dL/dopacity   [6..8] dL/dcolor  [9..11] unused
constexpr int ACC_STRIDE = 12;

// ---------------------------------------------------------------------------------------------
// Key packing chosen per call from the data (host side, after the num_rendered read-back).
//   packed mode : key = tile << (nb_depth+nb_idx) | (depth_bits - depth_min) << nb_idx | idx
//                 (keys only, 8 B per instance; sorted on bits [nb_idx, nb_idx+nb_depth+nb_tile))
//   pairs mode  : key = tile << nb_depth | (depth_bits - depth_min), value = idx
// Both orders are identical to a stable sort on the reference key tile<<32|depth_bits because
// depth_bits - depth_min is monotone in depth_bits and ties keep emission (= index) order.
// ---------------------------------------------------------------------------------------------
struct KeyFormat {
    int nb_idx, nb_depth, nb_tile;
    int pairs;            // 1 = separate u32 values
    uint32_t depth_min;
    int sort_lo, sort_hi; // sorted bit range
    int tile_shift;       // key >> tile_shift = tile id
};

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
};

void launch_preprocess(const PreprocessParams& p, cudaStream_t s);
void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present,
                         cudaStream_t s);
void launch_inclusive_scan(const uint32_t* in, uint32_t* out, size_t n, uint64_t* state,
                           uint32_t* ticket, uint32_t* total_out, cudaStream_t s);
void launch_emit_keys(int P, const GaussRec* rec, const uint32_t* offsets, const int* radii,
                      int grid_x, int grid_y, KeyFormat kf, uint64_t* keys, uint32_t* vals,
                      cudaStream_t s);
void launch_tile_ranges(size_t R, const uint64_t* keys_sorted, const uint32_t* vals_sorted,
                        KeyFormat kf, uint32_t num_tiles, uint2* ranges, uint32_t* point_list,
                        cudaStream_t s);
int  radix_sort(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, size_t n,
                int bit_lo, int bit_hi, void* temp, cudaStream_t s, int* result_in_b,
                 int scattered_from_bit = -1);
int  radix_sort_passes(int bit_lo, int bit_hi);

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
};
void launch_preprocess_backward(const PreprocessBackwardParams& p, cudaStream_t s);
void launch_sh_grad_from_factors(int P, int begin, int end, int D, int M, int world,
                                 const float* means3D, const float* campos_all,
                                 const float* factors_all, float* dL_dsh, cudaStream_t s);

}  // namespace bgr
