// Tile binning pipeline around the sort: inclusive scan of tiles_touched (L1), key emission (K2)
// and tile-range identification (K3, fused with the zero fill of empty tiles and the extraction of
// the sorted Gaussian-id list).
//
// Spec: SURVEY.md App. A.4 (reference rasterizer_impl.cu:70-138, 277, 310-318).
#include "bgr_common.cuh"
#include "gs_math.cuh"

namespace bgr {

namespace {

// --------------------------------------------------------------------------------------------
// Single-pass inclusive scan with decoupled look-back.
// state[b]: bits 63..32 = flag (0 = not ready, 1 = block aggregate, 2 = inclusive prefix),
//           bits 31..0  = value. Block ids come from an atomic ticket so that a block only ever
//           waits on blocks that have already started.
// --------------------------------------------------------------------------------------------
constexpr uint64_t FLAG_AGG = 1ull << 32;
constexpr uint64_t FLAG_PREFIX = 2ull << 32;

__device__ __forceinline__ uint64_t ld_volatile_u64(const uint64_t* p)
{
    return *reinterpret_cast<const volatile uint64_t*>(p);
}
__device__ __forceinline__ void st_volatile_u64(uint64_t* p, uint64_t v)
{
    *reinterpret_cast<volatile uint64_t*>(p) = v;
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, size_t n, uint64_t* state,
            uint32_t* ticket, uint32_t* total_out)
{
    __shared__ uint32_t s_bid;
    __shared__ uint32_t s_warp[SCAN_THREADS / 32];
    __shared__ uint32_t s_excl;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_bid = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t bid = s_bid;
    const size_t base = (size_t)bid * SCAN_TILE + (size_t)tid * SCAN_ITEMS;

    uint32_t v[SCAN_ITEMS];
    if (base + SCAN_ITEMS <= n && (reinterpret_cast<uintptr_t>(in) & 15) == 0) {
        const uint4 a = *reinterpret_cast<const uint4*>(in + base);
        const uint4 b = *reinterpret_cast<const uint4*>(in + base + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
        v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; ++i) v[i] = (base + i < n) ? in[base + i] : 0u;
    }
    uint32_t tsum = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) { tsum += v[i]; v[i] = tsum; }  // thread-inclusive

    // warp-inclusive scan of the thread totals
    uint32_t incl = tsum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, block_total = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; ++w) {
        const uint32_t t = s_warp[w];
        if (w < warp) warp_off += t;
        block_total += t;
    }
    const uint32_t thread_excl = warp_off + incl - tsum;

    if (warp == 0) {
        if (lane == 0)
            st_volatile_u64(&state[bid], (bid == 0 ? FLAG_PREFIX : FLAG_AGG) | block_total);
        uint32_t excl = 0;
        if (bid > 0) {
            int pred = (int)bid - 1;
            while (true) {
                const int j = pred - lane;
                uint64_t s = (j >= 0) ? ld_volatile_u64(&state[j]) : FLAG_PREFIX;
                while (__any_sync(0xffffffffu, (s >> 32) == 0))
                    s = (j >= 0) ? ld_volatile_u64(&state[j]) : FLAG_PREFIX;
                const unsigned pmask = __ballot_sync(0xffffffffu, (s >> 32) == 2);
                if (pmask) {
                    const int first = __ffs(pmask) - 1;
                    excl += __reduce_add_sync(0xffffffffu, lane <= first ? (uint32_t)s : 0u);
                    break;
                }
                excl += __reduce_add_sync(0xffffffffu, (uint32_t)s);
                pred -= 32;
            }
            if (lane == 0) st_volatile_u64(&state[bid], FLAG_PREFIX | (uint64_t)(excl + block_total));
        }
        if (lane == 0) {
            s_excl = excl;
            if ((size_t)(bid + 1) * SCAN_TILE >= n && total_out != nullptr)
                *total_out = excl + block_total;
        }
    }
    __syncthreads();
    const uint32_t add = s_excl + thread_excl;
    if (base + SCAN_ITEMS <= n && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
        *reinterpret_cast<uint4*>(out + base) =
            make_uint4(v[0] + add, v[1] + add, v[2] + add, v[3] + add);
        *reinterpret_cast<uint4*>(out + base + 4) =
            make_uint4(v[4] + add, v[5] + add, v[6] + add, v[7] + add);
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; ++i)
            if (base + i < n) out[base + i] = v[i] + add;
    }
}

// --------------------------------------------------------------------------------------------
// K2: one key per (Gaussian, touched tile), written at the Gaussian's scan offset in y-major,
// x-minor order (rasterizer_impl.cu:98-108). Key layout: bgr_common.cuh KeyFormat.
// --------------------------------------------------------------------------------------------
// The 32 Gaussians of a warp own one contiguous run of the output (their scan offsets are
// adjacent), so the warp writes that run cooperatively: output slot s is mapped back to its
// source lane by a 5-step search over the warp's exclusive prefix, which makes every store
// instruction cover 32 consecutive keys (the reference writes 12 B per instance at addresses
// strided by each thread's own tile count).
__global__ void __launch_bounds__(256)
emit_keys_kernel(int P, const GaussRec* __restrict__ rec, const uint32_t* __restrict__ offsets,
                 const int* __restrict__ radii, int grid_x, int grid_y, KeyFormat kf,
                 uint64_t* __restrict__ keys, uint32_t* __restrict__ vals)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t count = 0, minx = 0, miny = 0, width = 1;
    uint64_t low = 0;
    if (idx < P) {
        const int radius = radii[idx];
        if (radius > 0) {
            const float4 q0 = __ldg(&rec[idx].q0);
            const float depth = __ldg(&rec[idx].q2).z;
            uint2 rmin, rmax;
            tile_rect(make_float2(q0.x, q0.y), radius, grid_x, grid_y, rmin, rmax);
            minx = rmin.x; miny = rmin.y;
            width = rmax.x - rmin.x;
            count = width * (rmax.y - rmin.y);
            const uint64_t drel = (uint64_t)(__float_as_uint(depth) - kf.depth_min);
            low = kf.pairs ? drel : ((drel << kf.nb_idx) | (uint64_t)(uint32_t)idx);
            if (width == 0) width = 1;
        }
    }
    // exclusive prefix of the counts inside the warp
    uint32_t incl = count;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    const uint32_t excl = incl - count;
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    if (total == 0) return;
    // first output slot of the warp = scan offset preceding its first Gaussian
    const int first_idx = idx - lane;
    uint32_t warp_off = 0;
    if (lane == 0) warp_off = (first_idx == 0) ? 0u : offsets[first_idx - 1];
    warp_off = __shfl_sync(0xffffffffu, warp_off, 0);

    for (uint32_t base = 0; base < total; base += 32) {
        const uint32_t s = base + lane;
        // largest lane l with excl[l] <= s (zero-count lanes share a prefix with their successor;
        // taking the largest such lane lands on the one that actually owns the slot)
        int src = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
            const int cand = src + step;
            const uint32_t e = __shfl_sync(0xffffffffu, excl, cand & 31);
            if (cand < 32 && e <= s) src = cand;
        }
        const uint32_t e_src = __shfl_sync(0xffffffffu, excl, src);
        const uint32_t w_src = __shfl_sync(0xffffffffu, width, src);
        const uint32_t x_src = __shfl_sync(0xffffffffu, minx, src);
        const uint32_t y_src = __shfl_sync(0xffffffffu, miny, src);
        const uint64_t low_src = __shfl_sync(0xffffffffu, low, src);
        if (s < total) {
            const uint32_t local = s - e_src;
            const uint32_t ty = local / w_src, tx = local - ty * w_src;
            const uint64_t tile = (uint64_t)((y_src + ty) * (uint32_t)grid_x + x_src + tx);
            keys[warp_off + s] = (tile << kf.tile_shift) | low_src;
            if (kf.pairs) vals[warp_off + s] = (uint32_t)(first_idx + src);
        }
    }
}

// --------------------------------------------------------------------------------------------
// K3: ranges[t] = [first, last+1) of tile t in the sorted list; tiles without instances get
// (0,0) (the reference memsets, rasterizer_impl.cu:310). Also writes the sorted id list.
// --------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
tile_ranges_kernel(uint32_t R, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                   KeyFormat kf, uint32_t num_tiles, uint2* __restrict__ ranges,
                   uint32_t* __restrict__ point_list)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const uint64_t key = keys[i];
    const uint32_t cur = (uint32_t)(key >> kf.tile_shift);
    point_list[i] = kf.pairs ? vals[i] : (uint32_t)(key & ((1ull << kf.nb_idx) - 1ull));
    if (i == 0) {
        for (uint32_t t = 0; t < cur; ++t) ranges[t] = make_uint2(0u, 0u);
        ranges[cur].x = 0u;
    } else {
        const uint32_t prev = (uint32_t)(keys[i - 1] >> kf.tile_shift);
        if (cur != prev) {
            ranges[prev].y = i;
            ranges[cur].x = i;
            for (uint32_t t = prev + 1; t < cur; ++t) ranges[t] = make_uint2(0u, 0u);
        }
    }
    if (i == R - 1) {
        ranges[cur].y = R;
        for (uint32_t t = cur + 1; t < num_tiles; ++t) ranges[t] = make_uint2(0u, 0u);
    }
}

}  // namespace

void launch_inclusive_scan(const uint32_t* in, uint32_t* out, size_t n, uint64_t* state,
                           uint32_t* ticket, uint32_t* total_out, cudaStream_t s)
{
    if (n == 0) return;
    const unsigned blocks = (unsigned)((n + SCAN_TILE - 1) / SCAN_TILE);
    ++g_kernel_launches;
    scan_kernel<<<blocks, SCAN_THREADS, 0, s>>>(in, out, n, state, ticket, total_out);
}

void launch_emit_keys(int P, const GaussRec* rec, const uint32_t* offsets, const int* radii,
                      int grid_x, int grid_y, KeyFormat kf, uint64_t* keys, uint32_t* vals,
                      cudaStream_t s)
{
    if (P <= 0) return;
    ++g_kernel_launches;
    emit_keys_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, rec, offsets, radii, grid_x, grid_y, kf,
                                                    keys, vals);
}

void launch_tile_ranges(size_t R, const uint64_t* keys_sorted, const uint32_t* vals_sorted,
                        KeyFormat kf, uint32_t num_tiles, uint2* ranges, uint32_t* point_list,
                        cudaStream_t s)
{
    if (R == 0) return;
    ++g_kernel_launches;
    tile_ranges_kernel<<<(unsigned)((R + 255) / 256), 256, 0, s>>>(
        (uint32_t)R, keys_sorted, vals_sorted, kf, num_tiles, ranges, point_list);
}

}  // namespace bgr
