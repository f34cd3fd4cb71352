// Forward tile renderer, variant with TWO pixels per lane (opt-in: BGR_RENDER_2PX=1; the backward
// counterpart is in render_bwd.cu). Included by render_fwd.cu inside namespace bgr { namespace {.
//
// A warp owns an 8x8 pixel block, lane (lx, ly) the pixels (lx, ly) and (lx, ly + 4): the two 8x4
// halves are exactly the blocks two warps of render_forward_kernel own. The per-splat fixed cost
// (list walk, shared-memory loads, dx) is paid once for both halves; each half is skipped as a
// whole (warp-uniform) when the splat provably misses it or all of its pixels are finished.
// Per-pixel arithmetic, tests and traversal order are those of render_forward_kernel
// (reference forward.cu:330-356), so every output is bit-identical to it.
#pragma once

constexpr int F2_THREADS = 128;
constexpr int F2_WARPS = F2_THREADS / 32;
constexpr uint32_t F2_PLANE_BYTES = F2_THREADS * 16;

__global__ void __launch_bounds__(F2_THREADS)
render_forward_2px_kernel(int W, int H, const uint2* __restrict__ ranges,
                          const uint32_t* __restrict__ point_list, const GaussRec* __restrict__ rec,
                          const float* __restrict__ bg, float* __restrict__ final_T,
                          uint32_t* __restrict__ n_contrib, float* __restrict__ out_color)
{
    __shared__ float4 s_rec[3 * F2_THREADS];   // planes: geo | opa | col
    float4* const s_geo = s_rec;
    float4* const s_opa = s_rec + F2_THREADS;
    float4* const s_col = s_rec + 2 * F2_THREADS;
    uint32_t s_base;
    asm volatile("mov.u32 %0, %1;" : "=r"(s_base) : "r"((uint32_t)__cvta_generic_to_shared(s_rec)));
    __shared__ uint32_t s_warp_count[F2_WARPS];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tiles_x = (W + TILE_X - 1) / TILE_X;
    const uint32_t tile_px = blockIdx.x * TILE_X, tile_py = blockIdx.y * TILE_Y;
    const uint32_t wpx = tile_px + ((warp & 1) << 3), wpy = tile_py + ((warp >> 1) << 3);
    const uint32_t px = wpx + (lane & 7);
    const uint32_t pyA = wpy + (lane >> 3), pyB = pyA + 4;
    const bool insideA = px < (uint32_t)W && pyA < (uint32_t)H;
    const bool insideB = px < (uint32_t)W && pyB < (uint32_t)H;
    const float pixfx = (float)px, pixfyA = (float)pyA, pixfyB = (float)pyB;
    // tile rectangle (staging) and the two 8x4 halves of this warp (culling)
    const float rx0 = (float)tile_px, ry0 = (float)tile_py;
    const float rx1 = (float)min(tile_px + TILE_X - 1, (uint32_t)W - 1);
    const float ry1 = (float)min(tile_py + TILE_Y - 1, (uint32_t)H - 1);
    const float wx0 = (float)wpx, wx1 = (float)min(wpx + 7, (uint32_t)W - 1);
    const float wyA0 = (float)wpy, wyA1 = (float)min(wpy + 3, (uint32_t)H - 1);
    const float wyB0 = (float)(wpy + 4), wyB1 = (float)min(wpy + 7, (uint32_t)H - 1);

    const uint2 range = ranges[blockIdx.y * tiles_x + blockIdx.x];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + F2_THREADS - 1) / F2_THREADS;

    bool doneA = !insideA, doneB = !insideB;
    float TA = 1.0f, TB = 1.0f;
    uint32_t lastA = 0, lastB = 0;
    float A0 = 0.f, A1 = 0.f, A2 = 0.f, B0 = 0.f, B1 = 0.f, B2 = 0.f;

    for (int r = 0; r < rounds; ++r) {
        // all pixels finished -> stop fetching (also orders the previous round's smem reads
        // before this round's writes)
        if (__syncthreads_count(doneA && doneB) == F2_THREADS) break;

        const int progress = r * F2_THREADS + tid;
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
        for (int w = 0; w < F2_WARPS; ++w) {
            const int c = (int)s_warp_count[w];
            if (w < warp) before += c;
            kept += c;
        }
        if (keep) {
            const int pos = before + __popc(kmask & ((1u << lane) - 1u));
            s_geo[pos] = q0;
            s_opa[pos] = make_float4(q1.x, q2.y, q1.y, __uint_as_float((uint32_t)progress + 1u));
            s_col[pos] = make_float4(q1.z, q1.w, q2.x, 0.f);
        }
        __syncthreads();

        for (int base = 0; base < kept; base += 32) {
            // a half whose pixels are all finished (or outside the image) takes no more splats
            const bool liveA = !__all_sync(0xffffffffu, doneA);
            const bool liveB = !__all_sync(0xffffffffu, doneB);
            if (!liveA && !liveB) break;
            const int jt = base + lane;
            bool hitA = false, hitB = false;
            if (jt < kept) {
                const float4 a = s_geo[jt];
                const float4 b = s_opa[jt];
                hitA = liveA && !tile_rejects(a.x, a.y, a.z, a.w, b.x, b.y, wx0, wyA0, wx1, wyA1);
                hitB = liveB && !tile_rejects(a.x, a.y, a.z, a.w, b.x, b.y, wx0, wyB0, wx1, wyB1);
            }
            const unsigned hitsA = __ballot_sync(0xffffffffu, hitA);
            const unsigned hitsB = __ballot_sync(0xffffffffu, hitB);
            unsigned hits = hitsA | hitsB;
            while (hits) {
                const int bit = __ffs(hits) - 1;
                hits &= hits - 1;
                const bool inA = (hitsA >> bit) & 1u;   // warp-uniform
                const bool inB = (hitsB >> bit) & 1u;
                const uint32_t ja = s_base + 16u * (uint32_t)(base + bit);
                const float4 a = lds128(ja);
                const float4 b = lds128(ja + F2_PLANE_BYTES);
                const float4 c = lds128(ja + 2 * F2_PLANE_BYTES);
                const float dx = a.x - pixfx;
                if (inA) {
                    const float dy = a.y - pixfyA;
                    const float power = gauss_power(a.z, a.w, b.x, dx, dy);
                    bool ok = !doneA && !(power > 0.0f || power < b.y);
                    const float alpha = fminf(0.99f, b.z * expf(power));
                    ok = ok && !(alpha < 1.0f / 255.0f);
                    const float test_T = TA * (1 - alpha);
                    const bool stop = ok && test_T < 0.0001f;
                    doneA = doneA || stop;
                    ok = ok && !stop;
                    if (ok) {
                        A0 = __fmaf_rn(__fmul_rn(c.x, alpha), TA, A0);
                        A1 = __fmaf_rn(__fmul_rn(c.y, alpha), TA, A1);
                        A2 = __fmaf_rn(__fmul_rn(c.z, alpha), TA, A2);
                        TA = test_T;
                        lastA = __float_as_uint(b.w);
                    }
                }
                if (inB) {
                    const float dy = a.y - pixfyB;
                    const float power = gauss_power(a.z, a.w, b.x, dx, dy);
                    bool ok = !doneB && !(power > 0.0f || power < b.y);
                    const float alpha = fminf(0.99f, b.z * expf(power));
                    ok = ok && !(alpha < 1.0f / 255.0f);
                    const float test_T = TB * (1 - alpha);
                    const bool stop = ok && test_T < 0.0001f;
                    doneB = doneB || stop;
                    ok = ok && !stop;
                    if (ok) {
                        B0 = __fmaf_rn(__fmul_rn(c.x, alpha), TB, B0);
                        B1 = __fmaf_rn(__fmul_rn(c.y, alpha), TB, B1);
                        B2 = __fmaf_rn(__fmul_rn(c.z, alpha), TB, B2);
                        TB = test_T;
                        lastB = __float_as_uint(b.w);
                    }
                }
            }
        }
    }

    const size_t plane = (size_t)H * W;
    const float bg0 = bg[0], bg1 = bg[1], bg2 = bg[2];
    if (insideA) {
        const size_t pix_id = (size_t)pyA * W + px;
        final_T[pix_id] = TA;
        n_contrib[pix_id] = lastA;
        out_color[pix_id] = __fmaf_rn(TA, bg0, A0);
        out_color[plane + pix_id] = __fmaf_rn(TA, bg1, A1);
        out_color[2 * plane + pix_id] = __fmaf_rn(TA, bg2, A2);
    }
    if (insideB) {
        const size_t pix_id = (size_t)pyB * W + px;
        final_T[pix_id] = TB;
        n_contrib[pix_id] = lastB;
        out_color[pix_id] = __fmaf_rn(TB, bg0, B0);
        out_color[plane + pix_id] = __fmaf_rn(TB, bg1, B1);
        out_color[2 * plane + pix_id] = __fmaf_rn(TB, bg2, B2);
    }
}
