// Shared pieces of the forward and backward tile renderers.
#pragma once
#include "bgr_common.cuh"

namespace bgr {

constexpr int RENDER_THREADS = TILE_PIXELS;  // one thread per pixel of the 16x16 tile
constexpr int RENDER_WARPS = RENDER_THREADS / 32;

// Thread -> pixel mapping inside the tile: each warp owns an 8x4 pixel block (instead of a 16x2
// strip) so that the per-warp skip of non-overlapping splats is tighter. The mapping does not
// change any result; every pixel is still computed by exactly one thread.
__device__ __forceinline__ void tile_pixel(int tid, int& lx, int& ly)
{
    const int warp = tid >> 5, lane = tid & 31;
    lx = ((warp & 1) << 3) | (lane & 7);
    ly = ((warp >> 1) << 2) | (lane >> 3);
}

// 128-bit shared-memory load from a 32-bit shared-window address. The staged records of both render
// kernels live in ONE float4 array [3][RENDER_THREADS] (geo | opa | col), so the address of entry j is
// base + 16*j and the three planes are immediate offsets of the same register: the hot loops then
// issue one address computation per splat instead of recomputing the shared window base per access
// (which is what nvcc emitted for three separate __shared__ arrays).
__device__ __forceinline__ float4 lds128(uint32_t addr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "r"(addr));
    return v;
}
constexpr uint32_t PLANE_BYTES = RENDER_THREADS * 16;

// Packed FP32 pairs (sm_100: FMUL2 / FADD2 / FFMA2, one issue slot for two IEEE-rn operations).
// Both render kernels are issue-bound, not pipe-bound, so pairing independent operations of the
// same kind buys issue slots. Each half is rounded exactly like the scalar instruction.
__device__ __forceinline__ float2 mul2(float2 a, float2 b)
{
    float2 r;
    asm("{ .reg .b64 ra, rb, rr; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; mul.rn.f32x2 rr, ra, rb; "
        "mov.b64 {%0,%1}, rr; }"
        : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return r;
}
__device__ __forceinline__ float2 add2(float2 a, float2 b)
{
    float2 r;
    asm("{ .reg .b64 ra, rb, rr; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; add.rn.f32x2 rr, ra, rb; "
        "mov.b64 {%0,%1}, rr; }"
        : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return r;
}
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c)
{
    float2 r;
    asm("{ .reg .b64 ra, rb, rc, rr; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; mov.b64 rc, {%6,%7}; "
        "fma.rn.f32x2 rr, ra, rb, rc; mov.b64 {%0,%1}, rr; }"
        : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return r;
}

// Exactly the reference's evaluation order for the Gaussian exponent (forward.cu:335, SASS shape in
// SURVEY.md A.5): power = -0.5*(cx*dx*dx + cz*dy*dy) - cy*dx*dy with
//   t = (cz*dy)*dy, u = cx*dx, q = fma(u,dx,t), v = (cy*dx)*dy, power = fma(-0.5,q,-v).
__device__ __forceinline__ float gauss_power(float cx, float cy, float cz, float dx, float dy)
{
    const float t = __fmul_rn(__fmul_rn(cz, dy), dy);
    const float u = __fmul_rn(cx, dx);
    const float q = __fmaf_rn(u, dx, t);
    const float v = __fmul_rn(__fmul_rn(cy, dx), dy);
    return __fmaf_rn(-0.5f, q, -v);
}

// The same operations, the independent ones issued in pairs. The staged record keeps (cz, cy)
// next to each other (s_geo = x, y, cz, cy; s_opa = cx, ...) so that the pair needs no moves:
//   (czdy, cydx) = (cz, cy) * (dy, dx);  cxdx = cx*dx;  (t, v) = (czdy, cydx) * dy
// czdy and cxdx are handed back: the backward's first moments reuse them.
__device__ __forceinline__ float gauss_power2(float cx, float2 czcy, float2 d, float& cxdx, float& czdy)
{
    const float2 m = mul2(czcy, make_float2(d.y, d.x));
    cxdx = __fmul_rn(cx, d.x);
    czdy = m.x;
    const float2 tv = mul2(m, make_float2(d.y, d.y));
    const float q = __fmaf_rn(cxdx, d.x, tv.x);
    return __fmaf_rn(-0.5f, q, -tv.y);
}

// Conservative tile-level rejection. Returns true when NO pixel centre of the rectangle
// [px0,px1]x[py0,py1] can pass the reference's per-pixel tests (power <= 0 and
// opacity*exp(power) >= 1/255), so dropping the splat from this tile's working list cannot change
// any pixel. q(d) = -power is a convex quadratic when the conic is positive definite; its minimum
// over the rectangle is 0 if the centre is inside, otherwise it lies on one of the (at most two)
// edges facing the centre, where it is a 1-D quadratic with a clamped closed-form minimiser.
// Everything non-finite or not positive definite is kept (all comparisons are written so that NaN
// evaluates to "keep"). The tolerance covers the rounding difference between this bound and the
// exact per-pixel evaluation.
__device__ __forceinline__ bool tile_rejects(float mx, float my, float cx, float cy, float cz,
                                             float thr, float px0, float py0, float px1, float py1)
{
    // d = mean - pixel: intervals of dx, dy over the rectangle
    const float dx_lo = mx - px1, dx_hi = mx - px0;
    const float dy_lo = my - py1, dy_hi = my - py0;
    const bool in_x = (dx_lo <= 0.0f) && (dx_hi >= 0.0f);
    const bool in_y = (dy_lo <= 0.0f) && (dy_hi >= 0.0f);
    const float det = cx * cz - cy * cy;
    if (!(cx > 0.0f) || !(cz > 0.0f) || !(det > 0.0f)) return false;
    if (in_x && in_y) return false;
    const float tau = -thr;  // ln(255*opacity) + margin
    if (!(tau == tau)) return false;
    float qmin = 3.0e38f;
    float mag = 0.0f;
    if (!in_x) {
        // vertical edge facing the centre: dx = e fixed, minimise over dy
        const float e = (dx_lo > 0.0f) ? dx_lo : dx_hi;
        float dy = __fdividef(-cy * e, cz);  // minimiser position: a 2-ulp error moves q by O(ulp^2)
        dy = fminf(fmaxf(dy, dy_lo), dy_hi);
        const float a = 0.5f * cx * e * e, b = 0.5f * cz * dy * dy, c = cy * e * dy;
        const float q = a + b + c;
        if (q < qmin) { qmin = q; mag = fabsf(a) + fabsf(b) + fabsf(c); }
    }
    if (!in_y) {
        const float e = (dy_lo > 0.0f) ? dy_lo : dy_hi;
        float dx = __fdividef(-cy * e, cx);
        dx = fminf(fmaxf(dx, dx_lo), dx_hi);
        const float a = 0.5f * cz * e * e, b = 0.5f * cx * dx * dx, c = cy * e * dx;
        const float q = a + b + c;
        if (q < qmin) { qmin = q; mag = fabsf(a) + fabsf(b) + fabsf(c); }
    }
    return qmin > tau + 1e-3f + 8e-6f * mag;
}

}  // namespace bgr
