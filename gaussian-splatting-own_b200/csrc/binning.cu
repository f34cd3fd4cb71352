// Tile binning pipeline: per-tile instance counts from the coverage grid (replaces the prefix sum
// over tiles_touched, L1), scatter of one key per (Gaussian, touched tile) into its tile's bucket
// (K2), and a per-tile sort of every bucket (replaces the global sort L2 and, together with the
// tile scan, identifyTileRanges K3).
//
// Spec: SURVEY.md App. A.4 (reference rasterizer_impl.cu:70-138, 277, 303-318): the list of tile t
// holds every Gaussian whose tile rectangle covers t, ordered by (depth bits, Gaussian index);
// ranges[t] = [first, last+1) in the concatenation of the lists in tile order, (0,0) for an empty
// tile. The reference gets there by sorting all R instances on a 64-bit tile|depth key (6 radix
// passes of 12 bytes per instance at 1080p). Here the tile is never part of a key:
//   1. tile_scan_kernel   2-D prefix sum of the coverage grid -> count per tile; exclusive scan in
//                         tile order -> ranges, bucket cursors, R, the longest list;
//   2. scatter_keys_kernel  key = depth_bits << nb_idx | index, written at atomicAdd(cursor[tile]);
//   3. tile_sort_kernel   one block per tile sorts its bucket in shared memory (one distribution
//                         pass on the top bits of the depth relative to the tile's nearest
//                         Gaussian, then an exact ordering inside the small bins) and writes the
//                         Gaussian ids: 8 B written + 2 x 8 B read + 4 B written per instance
//                         instead of ~124.
// Within a bucket all keys are distinct (one per Gaussian), so the sorted order is unique whatever
// order the atomics filled the bucket in: point_list and ranges are bit-identical to the
// reference's.
#include <atomic>
#include "bgr_common.cuh"
#include "gs_math.cuh"

namespace bgr {

namespace {

// --------------------------------------------------------------------------------------------
// Tile scan. grid is the (gx+1) x (gy+1) difference grid: the preprocess added +1 at
// (min.x, min.y) and (max.x, max.y) and -1 at (max.x, min.y) and (min.x, max.y) of every visible
// Gaussian's tile rectangle, so the inclusive 2-D prefix sum at (x, y) is the number of rectangles
// covering tile (x, y). One block: the grid has 8 349 cells at 1080p, 32 776 at 4K.
// --------------------------------------------------------------------------------------------
constexpr int TSCAN_THREADS = 1024;
constexpr size_t TSCAN_MAX_SMEM = 200 * 1024;   // grids up to 51 200 cells are scanned in shared memory

// kSmem: the grid is first copied into shared memory (coalesced) and scanned there; larger grids
// are scanned in place in global memory.
template <bool kSmem>
__global__ void __launch_bounds__(TSCAN_THREADS)
tile_scan_kernel(int gx, int gy, int* __restrict__ grid_global, uint2* __restrict__ ranges,
                 uint32_t* __restrict__ cursor, GeomControl* __restrict__ ctl)
{
    extern __shared__ int s_grid[];
    __shared__ uint32_t s_part[TSCAN_THREADS / 32];
    __shared__ uint32_t s_max[TSCAN_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int gw = gx + 1;
    int* grid = kSmem ? s_grid : grid_global;
    if (kSmem) {
        const int cells = gw * gy;     // the last row (y == gy) only holds closing corners
        for (int i = tid; i < cells; i += TSCAN_THREADS) s_grid[i] = grid_global[i];
        __syncthreads();
    }
    // down the columns
    for (int x = tid; x < gx; x += TSCAN_THREADS) {
        int run = 0;
#pragma unroll 4
        for (int y = 0; y < gy; ++y) {
            run += grid[y * gw + x];
            grid[y * gw + x] = run;
        }
    }
    __syncthreads();
    // along the rows, one warp per row
    for (int y = warp; y < gy; y += TSCAN_THREADS / 32) {
        int carry = 0;
        for (int x0 = 0; x0 < gx; x0 += 32) {
            const int x = x0 + lane;
            int incl = x < gx ? grid[y * gw + x] : 0;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            if (x < gx) grid[y * gw + x] = incl + carry;
            carry += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
    __syncthreads();
    // Exclusive scan of the counts in tile order. The grid rows are contiguous except for one
    // padding cell per row (x == gx), so the scan runs over the padded array with that cell
    // counted as zero: each warp takes a contiguous span, 32 cells at a time (conflict-free
    // shared loads, coalesced stores); cell e of row y is tile e - y.
    const uint32_t cells = (uint32_t)gw * (uint32_t)gy;
    const uint32_t span = ((cells + TSCAN_THREADS - 1) / TSCAN_THREADS) * 32u;
    const uint32_t e0 = min(cells, (uint32_t)warp * span), e1 = min(cells, e0 + span);
    uint32_t sum = 0, longest = 0;
    {
        uint32_t e = e0 + lane;
        uint32_t y = e / (uint32_t)gw, x = e - y * (uint32_t)gw;
        for (; e < e1; e += 32) {
            const uint32_t c = x < (uint32_t)gx ? (uint32_t)grid[e] : 0u;
            sum += c;
            longest = max(longest, c);
            x += 32;
            while (x >= (uint32_t)gw) { x -= gw; ++y; }
        }
    }
    sum = __reduce_add_sync(0xffffffffu, sum);
    longest = __reduce_max_sync(0xffffffffu, longest);
    if (lane == 0) { s_part[warp] = sum; s_max[warp] = longest; }
    __syncthreads();
    uint32_t carry = 0, total = 0, block_longest = 0;
#pragma unroll
    for (int w = 0; w < TSCAN_THREADS / 32; ++w) {
        const uint32_t v = s_part[w];
        if (w < warp) carry += v;
        total += v;
        block_longest = max(block_longest, s_max[w]);
    }
    {
        uint32_t eb = e0;
        uint32_t e = e0 + lane;
        uint32_t y = e / (uint32_t)gw, x = e - y * (uint32_t)gw;
        for (; eb < e1; eb += 32, e += 32) {
            const bool tile = e < e1 && x < (uint32_t)gx;
            const uint32_t c = tile ? (uint32_t)grid[e] : 0u;
            uint32_t incl = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            const uint32_t base = carry + incl - c;
            if (tile) {
                const uint32_t t = e - y;
                // tiles without instances keep (0,0): the reference memsets (rasterizer_impl.cu:310)
                ranges[t] = c ? make_uint2(base, base + c) : make_uint2(0u, 0u);
                cursor[t] = base;
            }
            carry += __shfl_sync(0xffffffffu, incl, 31);
            x += 32;
            while (x >= (uint32_t)gw) { x -= gw; ++y; }
        }
    }
    if (tid == 0) {
        ctl->num_rendered = total;
        ctl->max_tile_list = block_longest;
    }
}

// --------------------------------------------------------------------------------------------
// K2: one key per (Gaussian, touched tile) (rasterizer_impl.cu:70-111), appended to the tile's
// bucket. The 32 Gaussians of a warp emit cooperatively: output slot s of the warp's run is mapped
// back to its source lane by a 5-step search over the warp's exclusive prefix, so a Gaussian that
// covers thousands of tiles does not serialise its warp.
// --------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
scatter_keys_kernel(int P, const GaussRec* __restrict__ rec, const int* __restrict__ radii, int grid_x,
                    int grid_y, int nb_idx, uint32_t* __restrict__ cursor, uint64_t* __restrict__ keys)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t count = 0, minx = 0, miny = 0, width = 1;
    uint64_t key = 0;
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
            key = ((uint64_t)__float_as_uint(depth) << nb_idx) | (uint64_t)(uint32_t)idx;
            if (width == 0) width = 1;
        }
    }
    uint32_t incl = count;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    const uint32_t excl = incl - count;
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    // Four slots per lane and round: the four cursor atomics are issued back to back, so their
    // L2 round trips overlap (the kernel is bound by the latency of these atomics). Giving every
    // cursor its own 32-byte sector was measured: 0.100 -> 0.095 ms at 3M / 1080p, but 1.37 -> 1.46 ms
    // at 10M / 4K / mod 2 and a slower tile scan (strided cursor stores) - not kept.
    for (uint32_t base = 0; base < total; base += 128) {
        uint32_t tile[4];
        uint64_t kv[4];
        bool live[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t s = base + 32u * u + lane;
            // largest lane l with excl[l] <= s (zero-count lanes share a prefix with their
            // successor; taking the largest such lane lands on the one that owns the slot)
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
            kv[u] = __shfl_sync(0xffffffffu, key, src);
            live[u] = s < total;
            const uint32_t local = live[u] ? s - e_src : 0u;
            const uint32_t ty = local / w_src, tx = local - ty * w_src;
            tile[u] = (y_src + ty) * (uint32_t)grid_x + x_src + tx;
        }
        uint32_t pos[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) pos[u] = live[u] ? atomicAdd(cursor + tile[u], 1u) : 0u;
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (live[u]) keys[pos[u]] = kv[u];
    }
}

// --------------------------------------------------------------------------------------------
// Per-tile sort: one block per tile (4 to 32 warps, by the longest list of the view).
//
// All keys of a bucket are distinct (one per Gaussian), so no stability is needed and ONE
// distribution pass does almost all the work:
//   1. the tile's depth range [dmin, dmax] is reduced over the bucket;
//   2. every key is counted into the bin of the top bits of (depth - dmin), with one to two bins
//      per key (512 .. 8192 bins); the shared-memory atomic also tells the key its arrival number
//      inside its bin, which stays in a register;
//   3. an exclusive scan of the bins gives every key its place: bin base + arrival number;
//   4. a bin holds a handful of keys in arrival order: each key counts the keys of its own bin
//      that are smaller (whole 64-bit key: depth bits, then index) and writes its Gaussian id to
//      out[bin base + that count].
// Per key: one shared atomic, one 8-byte shared store, a few compares - no radix passes and no
// warp votes. A bin that ends up with more than TS_MAX_BIN keys (many Gaussians at the same or
// nearly the same depth in one tile) would make step 4 quadratic: such a bucket - like one that
// does not fit in shared memory at all - takes the general path instead, a stable LSD radix sort
// of the bucket in place through global memory (9-bit digits; the index digits only if two equal
// depths arrived out of index order).
// --------------------------------------------------------------------------------------------
constexpr int TS_KEYS_PER_THREAD = 16;                 // fast path: n <= 16 x block threads
constexpr uint32_t TS_MAX_BIN = 48;
constexpr int TS_BINS_PER_KEY_LOG2 = 0;               // bins = 2^(bits(n) + this): 1..2 per key (measured
                                                       // against 2..4 and 0.5..1: 0.070 / 0.072 / 0.078 ms at 3M)
constexpr int TS_MIN_BIN_BITS = 9, TS_MAX_BIN_BITS = 13;

constexpr int TS_BITS = 9;                             // general path: LSD digit
constexpr int TS_BINS = 1 << TS_BITS;
constexpr int TS_ROW = TS_BINS + TS_BINS / 16;          // padded: digit d lives at d + d/16
constexpr int TS_LSD_WARPS = 4;                         // warps that work in the general path
__device__ __forceinline__ uint32_t ts_slot(uint32_t d) { return d + (d >> 4); }

__device__ __forceinline__ unsigned same_digit_lanes(uint32_t d, bool valid)
{
    unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
    for (int b = 0; b < TS_BITS; ++b) {
        const bool bit = (d >> b) & 1u;
        const unsigned m = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? m : ~m;
    }
    return peers;
}

// General path, one stable counting pass on key bits [shift, shift + 9) & mask: src[0, n) ->
// dst[0, n) in global memory. Warp w < TS_LSD_WARPS owns a contiguous segment, takes it 32 keys at
// a time in memory order and ranks a step's keys by lane order. cnt / off: [TS_LSD_WARPS][TS_ROW]
// counter and offset rows; every working warp derives the offsets of ITS row from all counter
// rows, so the pass needs two block barriers. All warps of the block must call it.
__device__ __forceinline__ void tile_lsd_pass(const uint64_t* src, uint64_t* dst, uint32_t n, int shift,
                                              uint32_t mask, uint32_t* cnt, uint32_t* off)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool working = warp < TS_LSD_WARPS;
    const uint32_t seg = ((n + TS_LSD_WARPS - 1) / TS_LSD_WARPS + 31u) & ~31u;
    const uint32_t begin = working ? min(n, (uint32_t)warp * seg) : n, end = min(n, begin + seg);
    uint32_t* my_cnt = cnt + (working ? warp : 0) * TS_ROW;
    uint32_t* my_off = off + (working ? warp : 0) * TS_ROW;
    if (working) {
        for (int k = lane; k < TS_ROW; k += 32) my_cnt[k] = 0;
        __syncwarp();
        for (uint32_t i = begin + lane; i < end; i += 32)
            atomicAdd(&my_cnt[ts_slot((uint32_t)(src[i] >> shift) & mask)], 1u);
    }
    __syncthreads();
    if (working) {
        // lane owns the 16 consecutive digits [16 lane, 16 lane + 16): total over the warps, and
        // the part counted by the warps before this one
        uint32_t below[16], lane_total = 0;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const uint32_t slot = (uint32_t)lane * 17u + k;      // == ts_slot(16 lane + k)
            uint32_t t = 0, bsum = 0;
#pragma unroll
            for (int w = 0; w < TS_LSD_WARPS; ++w) {
                const uint32_t c = cnt[w * TS_ROW + slot];
                bsum += w < warp ? c : 0u;
                t += c;
            }
            below[k] = lane_total + bsum;                         // exclusive inside the lane's digits
            lane_total += t;
        }
        uint32_t incl = lane_total;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        const uint32_t lane_base = incl - lane_total;
#pragma unroll
        for (int k = 0; k < 16; ++k) my_off[(uint32_t)lane * 17u + k] = lane_base + below[k];
        __syncwarp();
        for (uint32_t i0 = begin; i0 < end; i0 += 32) {
            const uint32_t i = i0 + lane;
            const bool valid = i < end;
            const uint64_t key = valid ? src[i] : 0ull;
            const uint32_t d = (uint32_t)(key >> shift) & mask;
            const unsigned peers = same_digit_lanes(d, valid);
            const uint32_t slot = ts_slot(d);
            uint32_t pos = 0;
            if (valid) pos = my_off[slot];
            __syncwarp();
            if (valid) {
                if (lane == __ffs(peers) - 1) my_off[slot] = pos + (uint32_t)__popc(peers);
                dst[pos + (uint32_t)__popc(peers & ((1u << lane) - 1u))] = key;
            }
            __syncwarp();
        }
    }
    __syncthreads();
}

// General path: sorts bucket[0, n) (global memory) by the whole key, ping-ponging with alt[0, n);
// returns the array holding the result.
__device__ __forceinline__ const uint64_t* tile_lsd_sort(uint64_t* bucket, uint64_t* alt, uint32_t n, int nb_idx,
                                                         uint32_t* cnt, uint32_t* off, uint32_t* s_red)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    // key bits that differ inside the bucket
    uint64_t diff = 0;
    const uint64_t first = bucket[0];
    for (uint32_t i = tid; i < n; i += blockDim.x) diff |= bucket[i] ^ first;
    uint32_t lo = __reduce_or_sync(0xffffffffu, (uint32_t)diff);
    uint32_t hi = __reduce_or_sync(0xffffffffu, (uint32_t)(diff >> 32));
    __syncthreads();
    if (lane == 0) { s_red[warp] = lo; s_red[32 + warp] = hi; }
    __syncthreads();
    lo = hi = 0;
    for (int w = 0; w < nwarps; ++w) { lo |= s_red[w]; hi |= s_red[32 + w]; }
    diff = ((uint64_t)hi << 32) | lo;
    const int top = 64 - __clzll((long long)(diff | 1ull));
    uint64_t* src = bucket;
    uint64_t* dst = alt;
    for (int attempt = 0; attempt < 2; ++attempt) {
        // attempt 0: depth digits only. attempt 1: index digits, then the depth digits again
        int shift = attempt == 0 ? nb_idx : 0;
        while (shift < top) {
            // a digit never straddles the index/depth boundary
            const int stop = (shift < nb_idx) ? min(nb_idx, top) : top;
            const int nb = min(TS_BITS, stop - shift);
            tile_lsd_pass(src, dst, n, shift, (1u << nb) - 1u, cnt, off);
            uint64_t* t = src; src = dst; dst = t;
            shift += nb;
        }
        if (attempt == 1 || nb_idx == 0) break;
        // in order already? (equal depth bits + descending index somewhere => redo on all digits)
        bool bad = false;
        for (uint32_t i = tid; i + 1 < n; i += blockDim.x) bad |= src[i] > src[i + 1];
        if (!__syncthreads_or(bad)) break;
    }
    return src;
}

template <int kWarps>
__global__ void __launch_bounds__(kWarps * 32)
tile_sort_kernel(const uint2* __restrict__ ranges, uint64_t* __restrict__ keys, uint64_t* __restrict__ keys_alt,
                 int nb_idx, uint32_t smem_keys, int max_bin_bits, uint32_t* __restrict__ point_list)
{
    constexpr int kThreads = kWarps * 32;
    extern __shared__ __align__(16) unsigned char tile_sort_smem[];
    uint64_t* s_keys = reinterpret_cast<uint64_t*>(tile_sort_smem);          // [smem_keys]
    uint32_t* s_bins = reinterpret_cast<uint32_t*>(s_keys + smem_keys);      // [2^max_bin_bits + 1] | LSD rows
    __shared__ uint32_t s_red[64];
    __shared__ uint32_t s_part[32];

    const uint2 range = ranges[blockIdx.x];
    const uint32_t n = range.y - range.x;
    if (n == 0) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t idx_mask = (1ull << nb_idx) - 1ull;
    uint64_t* bucket = keys + range.x;
    uint32_t* out = point_list + range.x;

    if (n <= 64) {
        // short list: every key counts the keys below it
        if ((uint32_t)tid < n) s_keys[tid] = bucket[tid];
        __syncthreads();
        if ((uint32_t)tid < n) {
            const uint64_t mine = s_keys[tid];
            uint32_t below = 0;
            for (uint32_t j = 0; j < n; ++j) below += s_keys[j] < mine ? 1u : 0u;
            out[below] = (uint32_t)(mine & idx_mask);
        }
        return;
    }

    bool general = n > smem_keys || n > (uint32_t)(kThreads * TS_KEYS_PER_THREAD);
    if (!general) {
        // 1. depth range of the tile
        uint32_t dmin = 0xffffffffu, dmax = 0u;
        for (uint32_t i = tid; i < n; i += kThreads) {
            const uint32_t d = (uint32_t)(bucket[i] >> nb_idx);
            dmin = min(dmin, d);
            dmax = max(dmax, d);
        }
        dmin = __reduce_min_sync(0xffffffffu, dmin);
        dmax = __reduce_max_sync(0xffffffffu, dmax);
        if (lane == 0) { s_red[warp] = dmin; s_red[32 + warp] = dmax; }
        __syncthreads();
#pragma unroll
        for (int w = 0; w < kWarps; ++w) { dmin = min(dmin, s_red[w]); dmax = max(dmax, s_red[32 + w]); }
        // one to two bins per key; the bin is the top `bin_bits` of the depth relative to dmin
        const int span_bits = 32 - __clz((int)(dmax - dmin));                  // 0 when all depths are equal
        const int want_bits = min(max(32 + TS_BINS_PER_KEY_LOG2 - __clz((int)n), TS_MIN_BIN_BITS), max_bin_bits);
        const int bin_bits = min(want_bits, max(span_bits, 1));
        const int bin_shift = max(span_bits - bin_bits, 0);
        const uint32_t bins = 1u << bin_bits;
        for (uint32_t k = tid; k <= bins; k += kThreads) s_bins[k] = 0;
        __syncthreads();
        // 2. count; the atomic's return value is the key's arrival number inside its bin
        uint32_t place[TS_KEYS_PER_THREAD];
#pragma unroll
        for (int k = 0; k < TS_KEYS_PER_THREAD; ++k) {
            if ((uint32_t)k * kThreads >= n) break;      // block-uniform: the list ends here
            const uint32_t i = (uint32_t)tid + (uint32_t)k * kThreads;
            place[k] = 0;
            if (i < n) {
                const uint32_t d = ((uint32_t)(bucket[i] >> nb_idx) - dmin) >> bin_shift;
                place[k] = atomicAdd(&s_bins[d], 1u);
            }
        }
        __syncthreads();
        // 3. exclusive scan of the bins (each warp a contiguous span, 32 bins at a time) and the
        //    fullest bin
        {
            const uint32_t span = ((bins + kThreads - 1) / kThreads) * 32u;
            const uint32_t b0 = min(bins, (uint32_t)warp * span), b1 = min(bins, b0 + span);
            uint32_t sum = 0, fullest = 0;
            for (uint32_t bb = b0 + lane; bb < b1; bb += 32) {
                const uint32_t c = s_bins[bb];
                sum += c;
                fullest = max(fullest, c);
            }
            sum = __reduce_add_sync(0xffffffffu, sum);
            fullest = __reduce_max_sync(0xffffffffu, fullest);
            if (lane == 0) { s_part[warp] = sum; s_red[warp] = fullest; }
            __syncthreads();
            uint32_t carry = 0;
            fullest = 0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
                if (w < warp) carry += s_part[w];
                fullest = max(fullest, s_red[w]);
            }
            general = fullest > TS_MAX_BIN;                                    // block-uniform
            if (!general) {
                for (uint32_t bb = b0; bb < b1; bb += 32) {
                    const uint32_t c = bb + lane < b1 ? s_bins[bb + lane] : 0u;
                    uint32_t incl = c;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
                        if (lane >= d) incl += t;
                    }
                    if (bb + lane < b1) s_bins[bb + lane] = carry + incl - c;
                    carry += __shfl_sync(0xffffffffu, incl, 31);
                }
                if (tid == 0) s_bins[bins] = n;
            }
            __syncthreads();
        }
        if (!general) {
            // place = bin base + arrival number; the bucket lands in shared memory bin by bin
#pragma unroll
            for (int k = 0; k < TS_KEYS_PER_THREAD; ++k) {
                if ((uint32_t)k * kThreads >= n) break;
                const uint32_t i = (uint32_t)tid + (uint32_t)k * kThreads;
                if (i < n) {
                    const uint64_t key = bucket[i];
                    place[k] += s_bins[((uint32_t)(key >> nb_idx) - dmin) >> bin_shift];
                    s_keys[place[k]] = key;
                }
            }
            __syncthreads();
            // 4. order inside the bin: count the smaller keys of the own bin
#pragma unroll
            for (int k = 0; k < TS_KEYS_PER_THREAD; ++k) {
                if ((uint32_t)k * kThreads >= n) break;
                const uint32_t i = (uint32_t)tid + (uint32_t)k * kThreads;
                if (i < n) {
                    const uint64_t key = s_keys[place[k]];
                    const uint32_t d = ((uint32_t)(key >> nb_idx) - dmin) >> bin_shift;
                    const uint32_t first = s_bins[d], last = s_bins[d + 1];
                    uint32_t below = 0;
                    for (uint32_t j = first; j < last; ++j) below += s_keys[j] < key ? 1u : 0u;
                    out[first + below] = (uint32_t)(key & idx_mask);
                }
            }
            return;
        }
    }
    // general path (block-uniform): stable LSD radix sort of the bucket through global memory
    __syncthreads();
    uint32_t* cnt = s_bins;
    uint32_t* off = cnt + TS_LSD_WARPS * TS_ROW;
    const uint64_t* sorted = tile_lsd_sort(bucket, keys_alt + range.x, n, nb_idx, cnt, off, s_red);
    for (uint32_t i = tid; i < n; i += kThreads) out[i] = (uint32_t)(sorted[i] & idx_mask);
}

constexpr size_t tile_sort_smem_bytes(uint32_t smem_keys, int max_bin_bits)
{
    const size_t bins = ((size_t)1 << max_bin_bits) + 1;
    const size_t rows = (size_t)TS_LSD_WARPS * TS_ROW * 2;
    return (size_t)smem_keys * 8 + (bins > rows ? bins : rows) * 4;
}

// function attributes are per device: raise the dynamic shared-memory limits once on each
int prepare_binning_kernels()
{
    static std::atomic<unsigned long long> attr_done_mask{0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    const unsigned long long dev_bit = 1ull << (dev & 63);
    if ((attr_done_mask.load(std::memory_order_relaxed) & dev_bit) != 0) return 0;
    const int most = (int)tile_sort_smem_bytes(TILE_SORT_MAX_SMEM_KEYS, TS_MAX_BIN_BITS);
    if (cudaFuncSetAttribute(tile_sort_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, most) != cudaSuccess ||
        cudaFuncSetAttribute(tile_sort_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, most) != cudaSuccess ||
        cudaFuncSetAttribute(tile_sort_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, most) != cudaSuccess ||
        cudaFuncSetAttribute(tile_sort_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, most) != cudaSuccess ||
        cudaFuncSetAttribute(tile_scan_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)TSCAN_MAX_SMEM) != cudaSuccess)
        return -1;
    attr_done_mask.fetch_or(dev_bit, std::memory_order_relaxed);
    return 0;
}

}  // namespace

int launch_tile_scan(int grid_x, int grid_y, int* coverage_grid, uint2* ranges, uint32_t* cursor,
                     GeomControl* control, cudaStream_t s)
{
    if (prepare_binning_kernels() != 0) return -1;
    const size_t bytes = (size_t)(grid_x + 1) * (size_t)grid_y * sizeof(int);
    ++g_kernel_launches;
    if (bytes <= TSCAN_MAX_SMEM)
        tile_scan_kernel<true><<<1, TSCAN_THREADS, bytes, s>>>(grid_x, grid_y, coverage_grid, ranges, cursor, control);
    else
        tile_scan_kernel<false><<<1, TSCAN_THREADS, 0, s>>>(grid_x, grid_y, coverage_grid, ranges, cursor, control);
    return 0;
}

void launch_scatter_keys(int P, const GaussRec* rec, const int* radii, int grid_x, int grid_y,
                         int nb_idx, uint32_t* cursor, uint64_t* keys, cudaStream_t s)
{
    if (P <= 0) return;
    ++g_kernel_launches;
    scatter_keys_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, rec, radii, grid_x, grid_y, nb_idx, cursor, keys);
}

int launch_tile_sort(uint32_t num_tiles, const uint2* ranges, uint64_t* keys, uint64_t* keys_alt,
                     int nb_idx, uint32_t smem_keys, uint32_t* point_list, cudaStream_t s)
{
    if (num_tiles == 0) return 0;
    if (smem_keys < 64 || smem_keys > TILE_SORT_MAX_SMEM_KEYS || keys_alt == nullptr ||
        prepare_binning_kernels() != 0)
        return -1;
    // bins for the longest list this launch sorts in shared memory
    int max_bin_bits = TS_MIN_BIN_BITS;
    while (max_bin_bits < TS_MAX_BIN_BITS && (1u << max_bin_bits) < (2u << TS_BINS_PER_KEY_LOG2) * smem_keys) ++max_bin_bits;
    const size_t smem = tile_sort_smem_bytes(smem_keys, max_bin_bits);
    ++g_kernel_launches;
    // Block size by the longest list (16 keys per thread at most). Measured on 1M-10M Gaussians at
    // 4K and 3M / 6M at 1080p: four warps are best up to 1 024 keys, eight from there to 4 096
    // (e.g. longest list 1 802: 0.68 / 0.56 / 0.76 ms with 4 / 8 / 16 warps); more warps only when
    // the keys-per-thread limit demands them (4 950: 1.62 ms with 16 warps, 2.35 with 32).
    if (smem_keys <= 1024)
        tile_sort_kernel<4><<<num_tiles, 128, smem, s>>>(ranges, keys, keys_alt, nb_idx, smem_keys, max_bin_bits, point_list);
    else if (smem_keys <= 8 * 32 * TS_KEYS_PER_THREAD)
        tile_sort_kernel<8><<<num_tiles, 256, smem, s>>>(ranges, keys, keys_alt, nb_idx, smem_keys, max_bin_bits, point_list);
    else if (smem_keys <= 16 * 32 * TS_KEYS_PER_THREAD)
        tile_sort_kernel<16><<<num_tiles, 512, smem, s>>>(ranges, keys, keys_alt, nb_idx, smem_keys, max_bin_bits, point_list);
    else
        tile_sort_kernel<32><<<num_tiles, 1024, smem, s>>>(ranges, keys, keys_alt, nb_idx, smem_keys, max_bin_bits, point_list);
    return 0;
}

}  // namespace bgr
