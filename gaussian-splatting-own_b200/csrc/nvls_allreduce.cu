// In-switch (NVLS) sum of the per-Gaussian gradient arena across the GPUs of one NVSwitch domain.
//
// The reference has no collective; this is the ONE exchange step of the
// view-parallel path (§8e): after the backward, every rank holds 59 floats per Gaussian of leaf
// gradients in a buffer that is mapped symmetrically on all ranks and bound to an NVLink multicast
// object. Rank r owns the r-th share of the range: for each 16-byte vector of its share it issues
//     multimem.ld_reduce.add.v4.f32  -> the switch reads the vector from all ranks and adds them,
//     multimem.st.v4.f32             -> the switch writes the sum back to all ranks,
// so every byte crosses each GPU's links once per direction (~size / link bandwidth in total)
// and all replicas receive bit-identical sums. Cross-rank ordering (all producers finished before
// the loads, all stores landed before consumers read) is provided by the caller with the symmetric
// memory barrier before and after this kernel; the kernel itself never waits on another GPU.
#include <algorithm>
#include "bgr_common.cuh"

namespace bgr {

namespace {

__device__ __forceinline__ float4 mm_ld_reduce(const float4* p)
{
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}
__device__ __forceinline__ void mm_st(float4* p, float4 v)
{
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
                 "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ float mm_ld_reduce1(const float* p)
{
    float v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st1(float* p, float v)
{
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}

constexpr int NVLS_THREADS = 512;
constexpr int NVLS_UNROLL = 4;

__global__ void __launch_bounds__(NVLS_THREADS)
nvls_allreduce_kernel(float* __restrict__ mc, size_t vec_begin, size_t vec_end, size_t tail_begin,
                      size_t tail_end)
{
    float4* mc4 = reinterpret_cast<float4*>(mc);
    const size_t stride = (size_t)gridDim.x * NVLS_THREADS;
    size_t i = vec_begin + (size_t)blockIdx.x * NVLS_THREADS + threadIdx.x;
    // NVLS_UNROLL independent in-switch reductions in flight per thread
    for (; i + (NVLS_UNROLL - 1) * stride < vec_end; i += NVLS_UNROLL * stride) {
        float4 v[NVLS_UNROLL];
#pragma unroll
        for (int u = 0; u < NVLS_UNROLL; ++u) v[u] = mm_ld_reduce(mc4 + i + u * stride);
#pragma unroll
        for (int u = 0; u < NVLS_UNROLL; ++u) mm_st(mc4 + i + u * stride, v[u]);
    }
    for (; i < vec_end; i += stride) mm_st(mc4 + i, mm_ld_reduce(mc4 + i));
    // scalar tail (only the last rank gets one, when the range is not a multiple of 4 floats)
    for (size_t t = tail_begin + (size_t)blockIdx.x * NVLS_THREADS + threadIdx.x; t < tail_end; t += stride)
        mm_st1(mc + t, mm_ld_reduce1(mc + t));
    __threadfence_system();
}

// Up to four disjoint ranges in ONE launch: the Gaussian range [p0, p1) of the structure-of-arrays
// leaf block (means3D | opacity | scales | rotations) is four short ranges, reduced as soon as the
// backward preprocess has produced that chunk on every rank (view_parallel.py pipelines them).
struct NvlsSegments {
    size_t begin[4];   // this rank's share of each range, in float4 units from the multicast base
    size_t end[4];
};

__global__ void __launch_bounds__(NVLS_THREADS)
nvls_allreduce_segments_kernel(float* __restrict__ mc, NvlsSegments seg)
{
    float4* mc4 = reinterpret_cast<float4*>(mc);
    const size_t stride = (size_t)gridDim.x * NVLS_THREADS;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const size_t vec_end = seg.end[k];
        size_t i = seg.begin[k] + (size_t)blockIdx.x * NVLS_THREADS + threadIdx.x;
        for (; i + (NVLS_UNROLL - 1) * stride < vec_end; i += NVLS_UNROLL * stride) {
            float4 v[NVLS_UNROLL];
#pragma unroll
            for (int u = 0; u < NVLS_UNROLL; ++u) v[u] = mm_ld_reduce(mc4 + i + u * stride);
#pragma unroll
            for (int u = 0; u < NVLS_UNROLL; ++u) mm_st(mc4 + i + u * stride, v[u]);
        }
        for (; i < vec_end; i += stride) mm_st(mc4 + i, mm_ld_reduce(mc4 + i));
    }
    __threadfence_system();
}

// ---- row-sparse variant -----------------------------------------------------------------------
// A Gaussian that no rank sees (radii == 0 everywhere) has an all-zero gradient row on every rank,
// so its row need not cross the fabric at all. vis_words holds one bit per Gaussian: the OR over
// ranks of (radii > 0), produced by pack_visibility_kernel + nvls_or_kernel below. Only the rows
// whose bit is set are summed; the others already hold the correct value (zero) everywhere.
__global__ void pack_visibility_kernel(const int* __restrict__ radii, int P, uint32_t* __restrict__ words)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool vis = i < P && radii[i] > 0;
    const unsigned w = __ballot_sync(0xffffffffu, vis);
    if ((threadIdx.x & 31) == 0 && i < P + 31) words[i >> 5] = w;
}

__global__ void nvls_or_kernel(uint32_t* __restrict__ mc_words, size_t begin, size_t end)
{
    for (size_t i = begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < end;
         i += (size_t)gridDim.x * blockDim.x) {
        uint32_t v;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.or.b32 %0, [%1];" : "=r"(v) : "l"(mc_words + i) : "memory");
        asm volatile("multimem.st.relaxed.sys.global.b32 [%0], %1;" ::"l"(mc_words + i), "r"(v) : "memory");
    }
    __threadfence_system();
}

// Rows [32*chunk_begin, 32*chunk_end) of two row-major blocks that share the row index:
//   block A: vecs_a float4 per row (rotations: 1), block B: vecs_b float4 per row (sh: 3M/4).
// A half-warp sums one visible row per step (lane k < vecs_a + vecs_b takes vector k).
__global__ void __launch_bounds__(NVLS_THREADS)
nvls_sparse_rows_kernel(float4* __restrict__ mc_a, int vecs_a, float4* __restrict__ mc_b, int vecs_b,
                        const uint32_t* __restrict__ vis_words, size_t chunk_begin, size_t chunk_end)
{
    const int lane = threadIdx.x & 31, half = lane >> 4, hl = lane & 15;
    const size_t warp = ((size_t)blockIdx.x * NVLS_THREADS + threadIdx.x) >> 5;
    const size_t nwarps = ((size_t)gridDim.x * NVLS_THREADS) >> 5;
    const int vecs = vecs_a + vecs_b;
    for (size_t c = chunk_begin + warp; c < chunk_end; c += nwarps) {
        unsigned bits = __ldg(vis_words + c);   // warp-uniform
        while (bits) {
            // up to four visible rows in flight per warp: two per half-warp
            float4 v[2];
            float4* addr[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                int b0 = -1, b1 = -1;
                if (bits) { b0 = __ffs(bits) - 1; bits &= bits - 1; }
                if (bits) { b1 = __ffs(bits) - 1; bits &= bits - 1; }
                const int row = half == 0 ? b0 : b1;
                addr[u] = nullptr;
                if (row >= 0 && hl < vecs) {
                    const size_t r = c * 32 + (size_t)row;
                    addr[u] = hl < vecs_a ? mc_a + r * vecs_a + hl : mc_b + r * vecs_b + (hl - vecs_a);
                    v[u] = mm_ld_reduce(addr[u]);
                }
            }
#pragma unroll
            for (int u = 0; u < 2; ++u)
                if (addr[u] != nullptr) mm_st(addr[u], v[u]);
        }
    }
    __threadfence_system();
}

}  // namespace

void launch_pack_visibility(const int* radii, int P, uint32_t* words, cudaStream_t s)
{
    if (P <= 0) return;
    const int n = (P + 31) / 32 * 32;
    ++g_kernel_launches;
    pack_visibility_kernel<<<(n + 255) / 256, 256, 0, s>>>(radii, P, words);
}

int launch_nvls_or(void* multicast_words, size_t nwords, int rank, int world, cudaStream_t s)
{
    if (multicast_words == nullptr || world <= 0 || rank < 0 || rank >= world) return -1;
    const size_t per = (nwords + world - 1) / world;
    const size_t b = std::min(nwords, per * (size_t)rank), e = std::min(nwords, b + per);
    if (e == b) return 0;
    ++g_kernel_launches;
    nvls_or_kernel<<<(unsigned)std::min<size_t>((e - b + 255) / 256, 296), 256, 0, s>>>(
        static_cast<uint32_t*>(multicast_words), b, e);
    return 0;
}

int launch_nvls_sparse_rows(void* mc_a, int vecs_a, void* mc_b, int vecs_b, const uint32_t* vis_words,
                            size_t rows, int rank, int world, int max_blocks, cudaStream_t s)
{
    if (mc_a == nullptr || mc_b == nullptr || vis_words == nullptr || world <= 0 || rank < 0 ||
        rank >= world || vecs_a < 0 || vecs_b < 0 || vecs_a + vecs_b <= 0 || vecs_a + vecs_b > 16)
        return -1;
    if (((reinterpret_cast<uintptr_t>(mc_a) | reinterpret_cast<uintptr_t>(mc_b)) & 15) != 0) return -1;
    const size_t chunks = (rows + 31) / 32;
    const size_t per = (chunks + world - 1) / world;
    const size_t b = std::min(chunks, per * (size_t)rank), e = std::min(chunks, b + per);
    if (e == b) return 0;
    const size_t warps_per_block = NVLS_THREADS / 32;
    size_t blocks = std::min<size_t>((e - b + warps_per_block - 1) / warps_per_block, (size_t)std::max(1, max_blocks));
    ++g_kernel_launches;
    nvls_sparse_rows_kernel<<<(unsigned)blocks, NVLS_THREADS, 0, s>>>(
        static_cast<float4*>(mc_a), vecs_a, static_cast<float4*>(mc_b), vecs_b, vis_words, b, e);
    return 0;
}

// Sums floats [offset, offset + count) of the multicast-mapped buffer across `world` ranks; this
// rank processes share `rank`. offset must be a multiple of 4 floats (16-byte alignment).
int launch_nvls_allreduce(void* multicast_ptr, size_t offset, size_t count, int rank, int world,
                          int max_blocks, cudaStream_t s)
{
    if (multicast_ptr == nullptr || world <= 0 || rank < 0 || rank >= world || (offset & 3) != 0)
        return -1;
    if (count == 0) return 0;
    float* mc = static_cast<float*>(multicast_ptr) + offset;
    const size_t nvec = count / 4;
    const size_t per = (nvec + world - 1) / world;
    const size_t vb = std::min(nvec, per * (size_t)rank);
    const size_t ve = std::min(nvec, vb + per);
    const bool last = rank == world - 1;
    const size_t tb = last ? nvec * 4 : 0, te = last ? count : 0;
    const size_t work = (ve - vb) + (te - tb);
    if (work == 0) return 0;
    size_t blocks = (work + (size_t)NVLS_THREADS * NVLS_UNROLL - 1) / ((size_t)NVLS_THREADS * NVLS_UNROLL);
    blocks = std::max<size_t>(1, std::min<size_t>(blocks, (size_t)std::max(1, max_blocks)));
    ++g_kernel_launches;
    nvls_allreduce_kernel<<<(unsigned)blocks, NVLS_THREADS, 0, s>>>(mc, vb, ve, tb, te);
    return 0;
}

int launch_nvls_allreduce_segments(void* multicast_ptr, int nseg, const size_t* offsets,
                                   const size_t* counts, int rank, int world, int max_blocks,
                                   cudaStream_t s)
{
    if (multicast_ptr == nullptr || world <= 0 || rank < 0 || rank >= world || nseg < 0 || nseg > 4)
        return -1;
    NvlsSegments seg;
    size_t work = 0;
    for (int k = 0; k < 4; ++k) {
        seg.begin[k] = seg.end[k] = 0;
        if (k >= nseg) continue;
        if ((offsets[k] & 3) != 0 || (counts[k] & 3) != 0) return -1;
        const size_t nvec = counts[k] / 4, first = offsets[k] / 4;
        const size_t per = (nvec + world - 1) / world;
        const size_t vb = std::min(nvec, per * (size_t)rank);
        const size_t ve = std::min(nvec, vb + per);
        seg.begin[k] = first + vb;
        seg.end[k] = first + ve;
        work = std::max(work, ve - vb);
    }
    if (work == 0) return 0;
    size_t blocks = (work + (size_t)NVLS_THREADS * NVLS_UNROLL - 1) / ((size_t)NVLS_THREADS * NVLS_UNROLL);
    blocks = std::max<size_t>(1, std::min<size_t>(blocks, (size_t)std::max(1, max_blocks)));
    ++g_kernel_launches;
    nvls_allreduce_segments_kernel<<<(unsigned)blocks, NVLS_THREADS, 0, s>>>(
        static_cast<float*>(multicast_ptr), seg);
    return 0;
}

}  // namespace bgr
