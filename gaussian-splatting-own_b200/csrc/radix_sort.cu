// Stable least-significant-digit "onesweep" radix sort of u64 keys (optionally with u32 values),
// restricted to the key bits [bit_lo, bit_hi). Replaces cub::DeviceRadixSort::SortPairs at
// reference rasterizer_impl.cu:303-308.
//
// Structure (one read + one write of every element per 8-bit digit):
//   1. sort_hist_kernel   : one pass over the keys builds the digit histograms of ALL passes.
//   2. sort_bases_kernel  : exclusive scan of each 256-bin histogram -> global digit bases.
//   3. onesweep_kernel x passes: each block ranks its 4096-key tile stably (warp match_any
//      multi-split), publishes its per-digit counts, resolves its global offsets with a decoupled
//      look-back over the preceding blocks (no second global pass), and scatters through shared
//      memory so that global writes are contiguous runs per digit.
// Stability: keys are taken in memory order warp by warp (warp-striped), ranks inside a warp step
// follow lane order, steps follow item order, warps follow warp order, blocks follow block order.
#include <atomic>
#include "bgr_common.cuh"

namespace bgr {

namespace {

constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr uint32_t ST_AGG = 1u << 30;
constexpr uint32_t ST_PREFIX = 2u << 30;
constexpr uint32_t ST_VALUE_MASK = (1u << 30) - 1u;

// Lanes holding my digit. kBallot: eight ballots (constant cost); otherwise MATCH.ANY, whose cost
// grows with the number of distinct digits in the warp (cheap on the depth digits, where the
// ~4 keys of one Gaussian are adjacent and equal; 4-5x slower on the tile digits).
template <bool kBallot>
__device__ __forceinline__ unsigned digit_peers(uint32_t d, bool valid)
{
    if (!kBallot) return __match_any_sync(0xffffffffu, d);
    unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
    for (int b = 0; b < RADIX_BITS; ++b) {
        const bool bit = (d >> b) & 1u;
        const unsigned m = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? m : ~m;
    }
    return valid ? peers : (1u << (threadIdx.x & 31));
}

__global__ void __launch_bounds__(256)
sort_hist_kernel(const uint64_t* __restrict__ keys, uint32_t n, int bit_lo, int bit_hi, int npasses,
                 uint32_t* __restrict__ hist)
{
    __shared__ uint32_t s_hist[SORT_MAX_PASSES * RADIX];
    for (int i = threadIdx.x; i < npasses * RADIX; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t n_round = (n + 31u) & ~31u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool valid = i < n;
        const uint64_t key = valid ? keys[i] : 0ull;
        for (int p = 0; p < npasses; ++p) {
            const int shift = bit_lo + p * RADIX_BITS;
            const int nb = min(RADIX_BITS, bit_hi - shift);
            const uint32_t d = valid ? ((uint32_t)(key >> shift) & ((1u << nb) - 1u)) : 0xffffffffu;
            if (valid) atomicAdd(&s_hist[p * RADIX + d], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < npasses * RADIX; i += blockDim.x) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// One block per pass: exclusive scan of the 256 bins, in place.
__global__ void __launch_bounds__(RADIX) sort_bases_kernel(uint32_t* __restrict__ hist)
{
    __shared__ uint32_t s_w[RADIX / 32];
    uint32_t* h = hist + blockIdx.x * RADIX;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t c = h[t];
    uint32_t incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
    for (int w = 0; w < warp; ++w) off += s_w[w];
    h[t] = off + incl - c;
}

template <bool PAIRS, bool kBallot>
__global__ void __launch_bounds__(SORT_THREADS, 4)
onesweep_kernel(const uint64_t* __restrict__ keys_in, uint64_t* __restrict__ keys_out,
                const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out, uint32_t n,
                int shift, uint32_t digit_mask, const uint32_t* __restrict__ bases,
                uint32_t* status, uint32_t* ticket)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* s_keys = reinterpret_cast<uint64_t*>(smem_raw);                        // SORT_TILE
    uint32_t* s_warp_hist = reinterpret_cast<uint32_t*>(s_keys + SORT_TILE);         // WARPS*RADIX
    uint32_t* s_digit_start = s_warp_hist + SORT_WARPS * RADIX;                      // RADIX
    uint32_t* s_goff = s_digit_start + RADIX;                                        // RADIX
    uint32_t* s_misc = s_goff + RADIX;                                               // 16
    uint32_t* s_vals = s_misc + 16;                                                  // SORT_TILE (PAIRS)

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_misc[0] = atomicAdd(ticket, 1u);
    for (int k = lane; k < RADIX; k += 32) s_warp_hist[warp * RADIX + k] = 0;
    __syncthreads();
    const uint32_t bid = s_misc[0];
    const uint32_t tile_base = bid * (uint32_t)SORT_TILE;
    const uint32_t valid_count = min((uint32_t)SORT_TILE, n - tile_base);

    uint64_t key[SORT_ITEMS];
    uint32_t val[SORT_ITEMS];
    uint32_t rank[SORT_ITEMS];
    const uint32_t warp_base = warp * (SORT_ITEMS * 32) + lane;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const uint32_t local = warp_base + i * 32;
        key[i] = (local < valid_count) ? keys_in[tile_base + local] : ~0ull;
        if (PAIRS) val[i] = (local < valid_count) ? vals_in[tile_base + local] : 0u;
    }

    const unsigned lt_mask = (1u << lane) - 1u;
    uint32_t* my_hist = s_warp_hist + warp * RADIX;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const bool valid = (warp_base + i * 32) < valid_count;
        const uint32_t d = valid ? ((uint32_t)(key[i] >> shift) & digit_mask) : 0xffffffffu;
        const unsigned peers = digit_peers<kBallot>(d, valid);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = my_hist[d];
            my_hist[d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[i] = old + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();

    // One thread per digit: warp counts -> exclusive warp offsets, block count, look-back.
    {
        const int t = tid;  // digit (SORT_THREADS == RADIX)
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w) {
            const uint32_t c = s_warp_hist[w * RADIX + t];
            s_warp_hist[w * RADIX + t] = run;
            run += c;
        }
        volatile uint32_t* st = status;
        st[(size_t)bid * RADIX + t] = (bid == 0 ? ST_PREFIX : ST_AGG) | run;

        // block-exclusive scan of the digit counts
        uint32_t incl = run;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
        }
        if (lane == 31) s_misc[1 + warp] = incl;
        __syncthreads();
        uint32_t off = 0;
        for (int w = 0; w < warp; ++w) off += s_misc[1 + w];
        const uint32_t dstart = off + incl - run;

        uint32_t excl = 0;
        if (bid > 0) {
            for (int b = (int)bid - 1;; --b) {
                uint32_t v;
                do {
                    v = st[(size_t)b * RADIX + t];
                } while ((v >> 30) == 0);
                excl += v & ST_VALUE_MASK;
                if ((v >> 30) == 2) break;
            }
            st[(size_t)bid * RADIX + t] = ST_PREFIX | (excl + run);
        }
        s_digit_start[t] = dstart;
        s_goff[t] = bases[t] + excl - dstart;  // wraps; only ever used as s_goff[d] + pos
    }
    __syncthreads();

#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        if ((warp_base + i * 32) < valid_count) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & digit_mask;
            const uint32_t pos = s_digit_start[d] + s_warp_hist[warp * RADIX + d] + rank[i];
            s_keys[pos] = key[i];
            if (PAIRS) s_vals[pos] = val[i];
        }
    }
    __syncthreads();

#pragma unroll
    for (int k = 0; k < SORT_ITEMS; ++k) {
        const uint32_t pos = k * SORT_THREADS + tid;
        if (pos < valid_count) {
            const uint64_t kk = s_keys[pos];
            const uint32_t d = (uint32_t)(kk >> shift) & digit_mask;
            const uint32_t dst = s_goff[d] + pos;
            keys_out[dst] = kk;
            if (PAIRS) vals_out[dst] = s_vals[pos];
        }
    }
}

constexpr size_t onesweep_smem(bool pairs)
{
    return (size_t)SORT_TILE * 8 + (size_t)SORT_WARPS * RADIX * 4 + RADIX * 4 * 2 + 16 * 4 +
           (pairs ? (size_t)SORT_TILE * 4 : 0);
}

}  // namespace

int radix_sort_passes(int bit_lo, int bit_hi)
{
    if (bit_hi <= bit_lo) return 0;
    return (bit_hi - bit_lo + RADIX_BITS - 1) / RADIX_BITS;
}

int radix_sort(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, size_t n,
               int bit_lo, int bit_hi, void* temp, cudaStream_t s, int* result_in_b,
               int scattered_from_bit)
{
    static_assert(SORT_THREADS == RADIX, "one thread per digit in the look-back stage");
    *result_in_b = 0;
    const int npasses = radix_sort_passes(bit_lo, bit_hi);
    if (n == 0 || npasses == 0) return 0;
    if (npasses > SORT_MAX_PASSES || n >= (1ull << 30)) return -1;
    const bool pairs = vals_a != nullptr;

    const SortTempLayout L = sort_temp_layout(n);
    unsigned char* tb = static_cast<unsigned char*>(temp);
    uint32_t* hist = reinterpret_cast<uint32_t*>(tb + L.hist);
    uint32_t* tickets = reinterpret_cast<uint32_t*>(tb + L.tickets);
    uint32_t* status = reinterpret_cast<uint32_t*>(tb + L.status);
    if (cudaMemsetAsync(tb, 0, L.status + (size_t)npasses * L.nblocks * RADIX * 4, s) != cudaSuccess) return -1;

    // function attributes are per device (context): opt in once on each device that sorts
    static std::atomic<unsigned long long> attr_done_mask{0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    const unsigned long long dev_bit = 1ull << (dev & 63);
    if ((attr_done_mask.load(std::memory_order_relaxed) & dev_bit) == 0) {
        const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
        if (cudaFuncSetAttribute(onesweep_kernel<false, false>, attr, (int)onesweep_smem(false)) != cudaSuccess ||
            cudaFuncSetAttribute(onesweep_kernel<false, true>, attr, (int)onesweep_smem(false)) != cudaSuccess ||
            cudaFuncSetAttribute(onesweep_kernel<true, false>, attr, (int)onesweep_smem(true)) != cudaSuccess ||
            cudaFuncSetAttribute(onesweep_kernel<true, true>, attr, (int)onesweep_smem(true)) != cudaSuccess)
            return -1;
        attr_done_mask.fetch_or(dev_bit, std::memory_order_relaxed);
    }

    const unsigned hist_blocks = (unsigned)std::min<size_t>((n + 255) / 256, 148 * 8);
    ++g_kernel_launches;
    sort_hist_kernel<<<hist_blocks, 256, 0, s>>>(keys_a, (uint32_t)n, bit_lo, bit_hi, npasses, hist);
    ++g_kernel_launches;
    sort_bases_kernel<<<npasses, RADIX, 0, s>>>(hist);

    uint64_t* kin = keys_a;
    uint64_t* kout = keys_b;
    uint32_t* vin = vals_a;
    uint32_t* vout = vals_b;
    for (int p = 0; p < npasses; ++p) {
        const int shift = bit_lo + p * RADIX_BITS;
        const int nb = std::min(RADIX_BITS, bit_hi - shift);
        const uint32_t mask = (1u << nb) - 1u;
        uint32_t* st = status + (size_t)p * L.nblocks * RADIX;
        ++g_kernel_launches;
        // digits that reach into [scattered_from_bit, ...) are high-entropy inside a warp
        const bool ballot = scattered_from_bit >= 0 && shift + nb > scattered_from_bit;
        const unsigned nblk = (unsigned)L.nblocks;
        const size_t smem = onesweep_smem(pairs);
        const uint32_t* bases = hist + p * RADIX;
        if (pairs && ballot)
            onesweep_kernel<true, true><<<nblk, SORT_THREADS, smem, s>>>(
                kin, kout, vin, vout, (uint32_t)n, shift, mask, bases, st, tickets + p);
        else if (pairs)
            onesweep_kernel<true, false><<<nblk, SORT_THREADS, smem, s>>>(
                kin, kout, vin, vout, (uint32_t)n, shift, mask, bases, st, tickets + p);
        else if (ballot)
            onesweep_kernel<false, true><<<nblk, SORT_THREADS, smem, s>>>(
                kin, kout, nullptr, nullptr, (uint32_t)n, shift, mask, bases, st, tickets + p);
        else
            onesweep_kernel<false, false><<<nblk, SORT_THREADS, smem, s>>>(
                kin, kout, nullptr, nullptr, (uint32_t)n, shift, mask, bases, st, tickets + p);
        std::swap(kin, kout);
        std::swap(vin, vout);
    }
    *result_in_b = npasses & 1;
    return 0;
}

}  // namespace bgr
