// Backward tile renderer (K5): back-to-front re-traversal producing per-Gaussian screen-space
// gradients (dL/dmean2D, dL/dconic, dL/dopacity, dL/dcolour).
//
// Spec: SURVEY.md App. A.8 (reference backward.cu:399-557). The set of (pixel, splat) pairs that
// contribute is decided by the same bit-exact tests as the forward pass; gradient values are
// tolerance-level (the reference accumulates them with unordered float atomics).
//
// Differences from the reference:
//  * only list entries below the tile's maximum n_contrib are visited at all;
//  * same record gather + conservative tile rejection + ordered compaction as the forward;
//  * the inner loop is warp-uniform: a splat that no lane of the warp blends is skipped by one
//    vote; otherwise the nine per-pixel gradient terms are reduced across the warp with a
//    transposing butterfly (12 shuffles instead of 45) and leave the SM as ONE reduction per
//    component per (warp, splat) instead of one per (pixel, splat): 9 REDs instead of up to 288;
//  * what is accumulated per Gaussian are six moments of G * dL/dalpha about the splat centre and
//    three colour sums (one 48-byte AoS row, bgr_common.cuh ACC_STRIDE); the per-splat constant
//    factors of backward.cu:534-552 are applied once per Gaussian by the fused backward preprocess.
// The per-pair arithmetic follows backward.cu:500-552 (a transcription of a mathematically fixed
// expression, regrouped as described at the accumulation).
#include <cstdlib>
#include "render_common.cuh"

namespace bgr {

namespace {

template <int kMinBlocks>
__global__ void __launch_bounds__(RENDER_THREADS, kMinBlocks)
render_backward_kernel(int W, int H, const uint2* __restrict__ ranges,
                       const uint32_t* __restrict__ point_list, const GaussRec* __restrict__ rec,
                       const float* __restrict__ bg, const float* __restrict__ final_T,
                       const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix,
                       float* __restrict__ acc)
{
    __shared__ float4 s_rec[3 * RENDER_THREADS];   // planes: geo | opa | col (see lds128)
    float4* const s_geo = s_rec;
    float4* const s_opa = s_rec + RENDER_THREADS;
    float4* const s_col = s_rec + 2 * RENDER_THREADS;
    uint32_t s_base;   // opaque to the optimiser so that it stays in a register across the loops
    asm volatile("mov.u32 %0, %1;" : "=r"(s_base) : "r"((uint32_t)__cvta_generic_to_shared(s_rec)));
    __shared__ uint32_t s_warp_count[RENDER_WARPS];
    __shared__ uint32_t s_warp_max[RENDER_WARPS];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tiles_x = (W + TILE_X - 1) / TILE_X;
    const uint32_t tile_px = blockIdx.x * TILE_X, tile_py = blockIdx.y * TILE_Y;
    int lx, ly;
    tile_pixel(tid, lx, ly);
    const uint32_t px = tile_px + lx, py = tile_py + ly;
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const size_t pix_id = (size_t)py * W + px;
    const float2 neg_pix = make_float2(-(float)px, -(float)py);
    const float rx0 = (float)tile_px, ry0 = (float)tile_py;
    const float rx1 = (float)min(tile_px + TILE_X - 1, (uint32_t)W - 1);
    const float ry1 = (float)min(tile_py + TILE_Y - 1, (uint32_t)H - 1);
    const uint32_t wpx = tile_px + ((warp & 1) << 3), wpy = tile_py + ((warp >> 1) << 2);
    const float wx0 = (float)wpx, wy0 = (float)wpy;
    const float wx1 = (float)min(wpx + 7, (uint32_t)W - 1);
    const float wy1 = (float)min(wpy + 3, (uint32_t)H - 1);

    const uint2 range = ranges[blockIdx.y * tiles_x + blockIdx.x];

    const float T_final = inside ? final_T[pix_id] : 0.0f;
    float T = T_final;
    const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0u;

    // Entries at or beyond the largest n_contrib of the tile cannot contribute to any pixel.
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, last_contributor);
    if (lane == 0) s_warp_max[warp] = wmax;
    __syncthreads();
    uint32_t total_u = 0;
#pragma unroll
    for (int w = 0; w < RENDER_WARPS; ++w) total_u = max(total_u, s_warp_max[w]);
    total_u = min(total_u, range.y - range.x);
    const int total = (int)total_u;
    if (total == 0) return;
    const int rounds = (total + RENDER_THREADS - 1) / RENDER_THREADS;

    // Per-pixel state of the back-to-front recursion (backward.cu:433-437, 509-521): T before the
    // current splat and the colour accumulated behind it. The reference keeps (last_alpha,
    // last_color) and folds them into accum_rec lazily at the next contributing splat; folding them
    // in right away gives the same value with four registers instead of eight.
    float accum0 = 0.f, accum1 = 0.f, accum2 = 0.f;
    float g0 = 0.f, g1 = 0.f, g2 = 0.f;
    if (inside) {
        const size_t plane = (size_t)H * W;
        g0 = dL_dpix[pix_id];
        g1 = dL_dpix[plane + pix_id];
        g2 = dL_dpix[2 * plane + pix_id];
    }
    // background term of dL/dalpha: -T_final / (1 - alpha) * (bg . dL/dpixel)  (backward.cu:529-532)
    const float bg_term = -T_final * (bg[0] * g0 + bg[1] * g1 + bg[2] * g2);
    // which of the nine reduced sums this lane ends up holding (see the butterfly below): lane 4k
    // holds sum k = bit2 + 2*bit3 + 4*bit4 of the lane id, lane 2 the ninth
    const bool red_writer = (lane & 3) == 0 || lane == 2;
    const int red_slot = lane == 2 ? 8 : (((lane >> 2) & 1) | (((lane >> 3) & 1) << 1) | (((lane >> 4) & 1) << 2));

    for (int r = 0; r < rounds; ++r) {
        __syncthreads();
        const int progress = r * RENDER_THREADS + tid;
        bool keep = false;
        float4 q0, q1, q2;
        uint32_t id = 0, li = 0;
        if (progress < total) {
            li = (uint32_t)(total - 1 - progress);  // back to front
            id = point_list[range.x + li];
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
            s_opa[pos] = make_float4(q0.z, q2.y, q1.y, __uint_as_float(li));   // cx, thr, opacity, list position
            s_col[pos] = make_float4(q1.z, q1.w, q2.x, __uint_as_float(id));
        }
        __syncthreads();

        // Per-warp culling (see render_fwd.cu): each lane tests one staged splat against the
        // warp's 8x4 pixel block and against the warp's largest n_contrib; the warp visits only
        // the survivors, in (reverse) list order.
        for (int base = 0; base < kept; base += 32) {
        const int jt = base + lane;
        bool hit = false;
        if (jt < kept) {
            const float4 ta = s_geo[jt];
            const float4 tb = s_opa[jt];
            hit = __float_as_uint(tb.w) < wmax &&
                  !tile_rejects(ta.x, ta.y, tb.x, ta.w, ta.z, tb.y, wx0, wy0, wx1, wy1);
        }
        unsigned hits = __ballot_sync(0xffffffffu, hit);
        while (hits) {
            const int j = base + __ffs(hits) - 1;
            hits &= hits - 1;
            const uint32_t ja = s_base + 16u * (uint32_t)j;
            const float4 a = lds128(ja);
            const float4 o = lds128(ja + PLANE_BYTES);
            const uint32_t pos_in_list = __float_as_uint(o.w);
            bool valid = pos_in_list < last_contributor;  // false for pixels outside the image
            const float2 dxy = add2(make_float2(a.x, a.y), neg_pix);   // mean - pixel
            const float dx = dxy.x, dy = dxy.y;
            float cxdx, czdy;
            const float power = gauss_power2(o.x, make_float2(a.z, a.w), dxy, cxdx, czdy);
            valid = valid && !(power > 0.0f || power < o.y);
            if (!__any_sync(0xffffffffu, valid)) continue;
            const float G = expf(power);
            const float alpha = fminf(0.99f, o.z * G);
            valid = valid && !(alpha < 1.0f / 255.0f);
            // (no second vote here: after the first one practically every iteration has a lane
            // that blends - 8.08 M of 8.08 M on the bench scene)
            const float4 c = lds128(ja + 2 * PLANE_BYTES);
            // What leaves the pixel are MOMENTS of u = G * dL/dalpha about the splat centre and the
            // colour weights w = alpha * T; the factors that are constant per splat (opacity, conic,
            // 0.5 W / 0.5 H) are applied once per Gaussian by the backward preprocess:
            //   dL/dmean2D.x = -0.5 W o Mx,  Mx = sum u (cx dx + cy dy)     dL/dconic.x = -0.5 o Sxx
            //   dL/dmean2D.y = -0.5 H o My,  My = sum u (cz dy + cy dx)     dL/dconic.y = -0.5 o Sxy
            //   dL/dopacity  = S1                                           dL/dconic.w = -0.5 o Syy
            // (backward.cu:534-552 with dL_dG = o dL/dalpha, gdx = G dx, gdy = G dy factored out).
            // cx dx + cy dy is formed per pixel, as the reference does: for an elongated, rotated
            // splat the two products nearly cancel and separate sums would lose the difference.
            //
            // Straight-line, like the forward: every lane computes, the state updates and the two
            // weights are predicated. 1/(1-alpha) is MUFU.RCP + one Newton step (1-alpha is in
            // [0.01, 1]) instead of the reference's two IEEE divisions. Without the Newton step
            // (1 ulp per factor of the T chain) a small scene put 2.5e-5 of the dL_dcov3D elements
            // outside the bar where two runs of the reference agree exactly.
            const float one_m_alpha = 1.f - alpha;
            float rcp_oma;
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rcp_oma) : "f"(one_m_alpha));
            rcp_oma = __fmaf_rn(rcp_oma, __fmaf_rn(-one_m_alpha, rcp_oma, 1.0f), rcp_oma);
            const float Tn = T * rcp_oma;
            const float d0 = c.x - accum0, d1 = c.y - accum1, d2 = c.z - accum2;
            float dL_dalpha = d0 * g0;
            dL_dalpha = __fmaf_rn(d1, g1, dL_dalpha);
            dL_dalpha = __fmaf_rn(d2, g2, dL_dalpha);
            dL_dalpha = __fmaf_rn(bg_term, rcp_oma, dL_dalpha * Tn);
            const float u = valid ? G * dL_dalpha : 0.f;
            const float w = valid ? alpha * Tn : 0.f;
            if (valid) {
                T = Tn;
                accum0 = __fmaf_rn(alpha, d0, accum0);   // = alpha c + (1 - alpha) accum
                accum1 = __fmaf_rn(alpha, d1, accum1);
                accum2 = __fmaf_rn(alpha, d2, accum2);
            }
            // moments, two per instruction where the operations pair up (FMUL2 / FFMA2)
            const float2 uu = make_float2(u, u);
            const float2 uxy = mul2(uu, dxy);                                   // u dx, u dy
            const float2 s24 = mul2(uxy, dxy);                                  // u dx^2, u dy^2
            const float2 lin = make_float2(__fmaf_rn(a.w, dy, cxdx), __fmaf_rn(a.w, dx, czdy));   // cx dx + cy dy, cz dy + cy dx
            const float2 v01 = mul2(uu, lin);
            const float2 v67 = mul2(make_float2(w, w), make_float2(g0, g1));
            const float v3 = uxy.x * dy;
            float v8 = w * g2;

            // Transposing butterfly: 8 values x 32 lanes -> each value summed over the warp.
            // pairs: (v0, v1) | (v2 = Sxx, v3 = Sxy) | (v4 = Syy, v5 = u) | (v6, v7)
            const bool h1 = (lane & 16) != 0;
            // (selects written per component: a float2 ?: becomes predicated moves)
            const float2 k01 = make_float2(h1 ? s24.y : v01.x, h1 ? u : v01.y);         // kept: v4,v5 | v0,v1
            const float2 k23 = make_float2(h1 ? v67.x : s24.x, h1 ? v67.y : v3);        //       v6,v7 | v2,v3
            const float2 t01 = make_float2(h1 ? v01.x : s24.y, h1 ? v01.y : u);         // sent
            const float2 t23 = make_float2(h1 ? s24.x : v67.x, h1 ? v3 : v67.y);
            const float2 w01 = add2(k01, make_float2(__shfl_xor_sync(0xffffffffu, t01.x, 16),
                                                     __shfl_xor_sync(0xffffffffu, t01.y, 16)));
            const float2 w23 = add2(k23, make_float2(__shfl_xor_sync(0xffffffffu, t23.x, 16),
                                                     __shfl_xor_sync(0xffffffffu, t23.y, 16)));
            const bool h2 = (lane & 8) != 0;
            const float2 k2 = make_float2(h2 ? w23.x : w01.x, h2 ? w23.y : w01.y);
            const float2 t2 = make_float2(h2 ? w01.x : w23.x, h2 ? w01.y : w23.y);
            const float2 u01 = add2(k2, make_float2(__shfl_xor_sync(0xffffffffu, t2.x, 8),
                                                    __shfl_xor_sync(0xffffffffu, t2.y, 8)));
            const bool h3 = (lane & 4) != 0;
            float sacc = (h3 ? u01.y : u01.x) + __shfl_xor_sync(0xffffffffu, h3 ? u01.x : u01.y, 4);
            // ninth value: plain butterfly over lane bits 4..2, then it shares the last two
            // steps with the other eight (lanes with bit 1 set carry it from here on)
            v8 += __shfl_xor_sync(0xffffffffu, v8, 16);
            v8 += __shfl_xor_sync(0xffffffffu, v8, 8);
            v8 += __shfl_xor_sync(0xffffffffu, v8, 4);
            const bool h4 = (lane & 2) != 0;
            float fin = (h4 ? v8 : sacc) + __shfl_xor_sync(0xffffffffu, h4 ? sacc : v8, 2);
            fin += __shfl_xor_sync(0xffffffffu, fin, 1);

            if (red_writer) atomicAdd(acc + (size_t)__float_as_uint(c.w) * ACC_STRIDE + red_slot, fin);
        }
        }
    }
}

}  // namespace

void launch_render_backward(int W, int H, const uint2* ranges, const uint32_t* point_list,
                            const GaussRec* rec, const float* /*colors_precomp*/, const float* bg,
                            const float* final_T, const uint32_t* n_contrib, const float* dL_dpix,
                            float* acc, cudaStream_t s)
{
    const dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y, 1);
    ++g_kernel_launches;
    // four resident blocks per SM (60 registers): five (48 registers, 8 bytes spilled) measured the
    // same within run-to-run noise (1.18-1.30 ms either way on the 3M / 1080p view)
    render_backward_kernel<4><<<grid, RENDER_THREADS, 0, s>>>(W, H, ranges, point_list, rec, bg, final_T,
                                                             n_contrib, dL_dpix, acc);
}

}  // namespace bgr
