// Forward tile renderer (K4): front-to-back alpha compositing of the depth-sorted per-tile lists.
//
// Spec: SURVEY.md App. A.4 (reference forward.cu:261-374). Per-pixel arithmetic (power, expf,
// alpha clamp, 1/255 skip, T < 1e-4 stop, colour FMA order) follows the reference instruction for
// instruction so that n_contrib is bit-identical and the image matches to rounding.
//
// What is different from the reference:
//  * the cooperative fetch gathers ONE 48-byte record per list entry (position, conic, opacity,
//    colour) instead of three arrays + three dependent global colour loads inside the pixel loop;
//  * the fetching thread runs an exact-conservative splat/tile rejection test (render_common.cuh)
//    and the batch is order-preservingly compacted in shared memory (ballot + prefix), so the
//    256 pixel threads only iterate over splats that can reach alpha >= 1/255 somewhere in the tile;
//  * a second, per-warp culling level: warps own 8x4 pixel blocks; 32 staged splats at a time are
//    tested (one per lane) against the warp's block and only the survivors are visited, so a pair
//    that cannot contribute costs ~2 issued instructions instead of ~18;
//  * per pair, a single compare against the precomputed log-threshold skips the expf for pairs
//    whose alpha is provably < 1/255.
#include <cstdlib>
#include "render_common.cuh"

namespace bgr {

namespace {

// kStats: diagnostic instantiation used by bgr_render_stats (never by bgr_forward). It runs the
// identical traversal and additionally counts, over the whole image,
//   stats[0] staged (tile, splat) entries that survived the tile rejection,
//   stats[1] (pixel, splat) pairs examined (power evaluated),
//   stats[2] pairs whose expf was evaluated,  stats[3] pairs blended into the pixel.
template <bool kStats>
__global__ void __launch_bounds__(RENDER_THREADS, 5)
render_forward_kernel(int W, int H, const uint2* __restrict__ ranges,
                      const uint32_t* __restrict__ point_list, const GaussRec* __restrict__ rec,
                      const float* __restrict__ bg, float* __restrict__ final_T,
                      uint32_t* __restrict__ n_contrib, float* __restrict__ out_color,
                      unsigned long long* __restrict__ stats)
{
    // Staged splats of the current round, split by when a field is needed:
    //   s_geo : x, y, conic.x, conic.y            (culling test + every pair)
    //   s_opa : conic.z, thr, opacity, list pos+1 (culling test + every pair)
    //   s_col : r, g, b                            (only pairs that blend)
    __shared__ float4 s_rec[3 * RENDER_THREADS];   // planes: geo | opa | col (see lds128)
    float4* const s_geo = s_rec;
    float4* const s_opa = s_rec + RENDER_THREADS;
    float4* const s_col = s_rec + 2 * RENDER_THREADS;
    uint32_t s_base;   // opaque to the optimiser so that it stays in a register across the loops
    asm volatile("mov.u32 %0, %1;" : "=r"(s_base) : "r"((uint32_t)__cvta_generic_to_shared(s_rec)));
    __shared__ uint32_t s_warp_count[RENDER_WARPS];
    unsigned long long st_kept = 0, st_exam = 0, st_exp = 0, st_blend = 0;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tiles_x = (W + TILE_X - 1) / TILE_X;
    const uint32_t tile_px = blockIdx.x * TILE_X, tile_py = blockIdx.y * TILE_Y;
    int lx, ly;
    tile_pixel(tid, lx, ly);
    const uint32_t px = tile_px + lx, py = tile_py + ly;
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const float2 neg_pix = make_float2(-(float)px, -(float)py);
    // pixel-centre rectangle of the part of the tile that is inside the image
    const float rx0 = (float)tile_px, ry0 = (float)tile_py;
    const float rx1 = (float)min(tile_px + TILE_X - 1, (uint32_t)W - 1);
    const float ry1 = (float)min(tile_py + TILE_Y - 1, (uint32_t)H - 1);
    // ... and of this warp's 8x4 pixel block
    const uint32_t wpx = tile_px + ((warp & 1) << 3), wpy = tile_py + ((warp >> 1) << 2);
    const float wx0 = (float)wpx, wy0 = (float)wpy;
    const float wx1 = (float)min(wpx + 7, (uint32_t)W - 1);
    const float wy1 = (float)min(wpy + 3, (uint32_t)H - 1);

    const uint2 range = ranges[blockIdx.y * tiles_x + blockIdx.x];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + RENDER_THREADS - 1) / RENDER_THREADS;

    bool done = !inside;
    float T = 1.0f;
    uint32_t last_contributor = 0;
    float2 C01 = make_float2(0.f, 0.f);
    float C2 = 0.f;

    for (int r = 0; r < rounds; ++r) {
        // all pixels finished -> stop fetching (also orders the previous round's smem reads
        // before this round's writes)
        if (__syncthreads_count(done) == RENDER_THREADS) break;

        const int progress = r * RENDER_THREADS + tid;
        bool keep = false;
        float4 q0, q1, q2;
        if (progress < total) {
            const uint32_t id = point_list[range.x + progress];
            const GaussRec* g = rec + id;
            q0 = __ldg(&g->q0);
            q1 = __ldg(&g->q1);
            q2 = __ldg(&g->q2);
            keep = !tile_rejects(q0.x, q0.y, q0.z, q0.w, q1.x, q2.y, rx0, ry0, rx1, ry1);
        }
        const unsigned kmask = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) s_warp_count[warp] = __popc(kmask);
        __syncthreads();
        int before = 0, kept = 0;
#pragma unroll
        for (int w = 0; w < RENDER_WARPS; ++w) {
            const int c = (int)s_warp_count[w];
            if (w < warp) before += c;
            kept += c;
        }
        if (keep) {
            const int pos = before + __popc(kmask & ((1u << lane) - 1u));
            s_geo[pos] = make_float4(q0.x, q0.y, q1.x, q0.w);   // x, y, cz, cy
            s_opa[pos] = make_float4(q0.z, q2.y, q1.y, __uint_as_float((uint32_t)progress + 1u));   // cx, thr, opacity
            s_col[pos] = make_float4(q1.z, q1.w, q2.x, 0.f);
        }
        __syncthreads();
        if (kStats && tid == 0) st_kept += (unsigned long long)kept;

        // Per-warp culling: 32 staged splats at a time, each lane tests one of them against the
        // warp's 8x4 pixel block; the warp then visits only the survivors, in list order.
        for (int base = 0; base < kept; base += 32) {
            if (__all_sync(0xffffffffu, done)) break;
            const int jt = base + lane;
            bool hit = false;
            if (jt < kept) {
                const float4 a = s_geo[jt];
                const float4 b = s_opa[jt];
                hit = !tile_rejects(a.x, a.y, b.x, a.w, a.z, b.y, wx0, wy0, wx1, wy1);
            }
            unsigned hits = __ballot_sync(0xffffffffu, hit);
            while (hits) {
                const int j = base + __ffs(hits) - 1;
                hits &= hits - 1;
                // Straight-line body: in 94 % of the iterations some lane of the warp blends, so
                // per-lane branches only add reconvergence overhead; every lane computes, the
                // updates are predicated. Values and tests are those of forward.cu:330-356.
                const uint32_t ja = s_base + 16u * (uint32_t)j;
                const float4 a = lds128(ja);
                const float4 b = lds128(ja + PLANE_BYTES);
                float cxdx, czdy;
                const float power = gauss_power2(b.x, make_float2(a.z, a.w),
                                                 add2(make_float2(a.x, a.y), neg_pix), cxdx, czdy);
                bool ok = !done && !(power > 0.0f || power < b.y);
                if (kStats) st_exam += done ? 0 : 1;
                const float alpha = fminf(0.99f, b.z * expf(power));
                if (kStats) st_exp += ok ? 1 : 0;
                ok = ok && !(alpha < 1.0f / 255.0f);
                const float test_T = T * (1 - alpha);
                const bool stop = ok && test_T < 0.0001f;
                done = done || stop;
                ok = ok && !stop;
                const float4 c = lds128(ja + 2 * PLANE_BYTES);
                if (ok) {
                    C01 = fma2(mul2(make_float2(c.x, c.y), make_float2(alpha, alpha)), make_float2(T, T), C01);
                    C2 = __fmaf_rn(__fmul_rn(c.z, alpha), T, C2);
                    T = test_T;
                    last_contributor = __float_as_uint(b.w);
                }
                if (kStats) st_blend += ok ? 1 : 0;
            }
        }
    }

    if (kStats) {
        atomicAdd(&stats[0], st_kept);
        atomicAdd(&stats[1], st_exam);
        atomicAdd(&stats[2], st_exp);
        atomicAdd(&stats[3], st_blend);
        return;  // statistics only: the image buffers are not written
    }

    if (inside) {
        const size_t pix_id = (size_t)py * W + px;
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
        const size_t plane = (size_t)H * W;
        out_color[pix_id] = __fmaf_rn(T, bg[0], C01.x);
        out_color[plane + pix_id] = __fmaf_rn(T, bg[1], C01.y);
        out_color[2 * plane + pix_id] = __fmaf_rn(T, bg[2], C2);
    }
}

}  // namespace

void launch_render_forward(int W, int H, const uint2* ranges, const uint32_t* point_list,
                           const GaussRec* rec, const float* /*colors_precomp*/, const float* bg,
                           float* final_T, uint32_t* n_contrib, float* out_color, cudaStream_t s)
{
    const dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y, 1);
    ++g_kernel_launches;
    render_forward_kernel<false><<<grid, RENDER_THREADS, 0, s>>>(W, H, ranges, point_list, rec, bg,
                                                                final_T, n_contrib, out_color,
                                                                nullptr);
}

void launch_render_stats(int W, int H, const uint2* ranges, const uint32_t* point_list,
                         const GaussRec* rec, unsigned long long* stats, cudaStream_t s)
{
    const dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y, 1);
    render_forward_kernel<true><<<grid, RENDER_THREADS, 0, s>>>(W, H, ranges, point_list, rec, nullptr,
                                                               nullptr, nullptr, nullptr, stats);
}

}  // namespace bgr
