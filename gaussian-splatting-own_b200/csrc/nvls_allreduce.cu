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

}  // namespace

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
