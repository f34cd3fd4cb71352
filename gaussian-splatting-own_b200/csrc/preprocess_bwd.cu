// Backward of the per-Gaussian preprocess: K6 (conic -> 3D covariance / mean) and K7 (screen
// mean -> 3D mean, colour -> SH, 3D covariance -> scale / rotation) fused into ONE pass that also
// writes the rows of culled Gaussians as zeros, so no output needs a prior memset (the reference
// zero-fills 300 B/Gaussian per backward, rasterize_points.cu:151-159).
//
// Spec: SURVEY.md App. A.9 / A.10 (reference backward.cu:144-396, auxiliary.h:107-117).
// Tolerance-level parity (gradients); the forward quantities it recomputes (view-space mean,
// Jacobian, 3D covariance) use the same helpers as the forward preprocess.
#include "bgr_common.cuh"
#include "gs_math.cuh"

namespace bgr {

namespace {

constexpr int PB_THREADS = 128;   // 26 KB SH tile per block, ~5 blocks per SM

__device__ __forceinline__ void store3(float* p, float a, float b, float c)
{
    p[0] = a; p[1] = b; p[2] = c;
}

// kVecSH: a Gaussian's SH row is 192 contiguous bytes, so per-thread 128-bit accesses of a warp
// touch 32 half sectors 192 bytes apart. Rows therefore move through a shared-memory tile:
//  * loads: the warp fetches the rows of its VISIBLE lanes cooperatively (twelve lanes per row,
//    whole sectors, culled rows not read at all) and every lane then reads its own row from the
//    tile, four coefficients at a time;
//  * stores (!kFactor): every lane leaves its 16 basis values and dL/dRGB (zeros when culled) in
//    the same tile row and the warp writes its 32 x 48 block of products with 512-byte stores.
constexpr int SH_TILE_STRIDE = 52;   // 48 SH floats (then 16 basis values + dL/dRGB); 52 = 4 * 13
                                     // keeps rows 16-byte aligned and quarter-warps conflict free

// Per-Gaussian body. Returns the clamp-masked colour gradient dL/dRGB (zero for culled
// Gaussians): dL/dsh[k] = basis_k(dir) * dL/dRGB is rank one, which the view-parallel path
// exploits (kFactor: dL_dsh is NOT written here, the caller multicasts the returned factor).
// kRaw (bgr_backward_raw): the inputs are the trainer's raw leaves (activated here exactly as
// the forward did) and dL_dscale / dL_drot / dL_dopacity are chained through exp / normalize /
// sigmoid; the SH row and the SH gradients move through the tile as with kVecSH, the kernel splits
// them between the _features_dc and _features_rest layouts.
template <bool kVecSH, bool kFactor, bool kRaw>
__device__ __forceinline__ V3 preprocess_backward_body(const PreprocessBackwardParams& p, int idx,
                                                       bool visible, const float* s_view,
                                                       const float* s_proj, const float* s_cam,
                                                       float* tile_row)
{
    constexpr bool kCoop = kVecSH && !kFactor;
    const size_t i = (size_t)idx;
    const int ncoef = (p.D + 1) * (p.D + 1);

    if (!visible) {
        if (kCoop) {
#pragma unroll
            for (int k = 0; k < 19; ++k) tile_row[k] = 0.f;
        }
        if (kRaw) store3(p.dL_dsh + 3 * i, 0.f, 0.f, 0.f);
        store3(p.dL_dmean2D + 3 * i, 0.f, 0.f, 0.f);
        store3(p.dL_dcolor + 3 * i, 0.f, 0.f, 0.f);
        p.dL_dopacity[i] = 0.f;
        store3(p.dL_dmean3D + 3 * i, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < 6; ++k) p.dL_dcov3D[6 * i + k] = 0.f;
        store3(p.dL_dscale + 3 * i, 0.f, 0.f, 0.f);
        if ((reinterpret_cast<uintptr_t>(p.dL_drot) & 15) == 0) {
            reinterpret_cast<float4*>(p.dL_drot)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) p.dL_drot[4 * i + k] = 0.f;
        }
        if (p.M > 0 && !kFactor && !kCoop) {
            if (kVecSH) {
                float4* dst = reinterpret_cast<float4*>(p.dL_dsh) + i * 12;
#pragma unroll
                for (int k = 0; k < 12; ++k) dst[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            } else {
                float* dst = p.dL_dsh + i * p.M * 3;
                for (int k = 0; k < p.M * 3; ++k) dst[k] = 0.f;
            }
        }
        return v3(0.f, 0.f, 0.f);
    }
    V3 factor = v3(0.f, 0.f, 0.f);

    // ---- screen-space gradients accumulated by the backward render ---------------------------
    // The render kernel leaves per-Gaussian MOMENTS of u = G * dL/dalpha about the splat centre
    // (render_bwd.cu); the factors of backward.cu:534-552 that are constant per splat are applied
    // here, once:  dL_dG = o * dL/dalpha,  dG/ddelx = -gdx*cx - gdy*cy,  ddelx/dx = 0.5 W, ...
    const float4* arow = reinterpret_cast<const float4*>(p.acc + i * ACC_STRIDE);
    const float4 a0 = arow[0], a1 = arow[1], a2 = arow[2];
    const float opac = __ldg(&p.rec[idx].q1).y;
    const float Mx = a0.x, My = a0.y, Sxx = a0.z, Sxy = a0.w, Syy = a1.x;
    const float mh_o = -0.5f * opac;
    const float gmx = mh_o * Mx * (float)p.W;                             // dL/dmean2D
    const float gmy = mh_o * My * (float)p.H;
    const float gcx = mh_o * Sxx, gcy = mh_o * Sxy, gcw = mh_o * Syy;     // dL/dconic (x, y, w)
    const float gop = a1.y;                                               // dL/dopacity = sum of u
    const V3 gcol = v3(a1.z, a1.w, a2.x);
    store3(p.dL_dmean2D + 3 * i, gmx, gmy, 0.f);
    store3(p.dL_dcolor + 3 * i, gcol.x, gcol.y, gcol.z);
    // raw leaf: d sigmoid(x) = o (1 - o)
    p.dL_dopacity[i] = kRaw ? gop * ((1.0f - opac) * opac) : gop;

    const float3 mean = {p.means3D[3 * i], p.means3D[3 * i + 1], p.means3D[3 * i + 2]};

    // ---- 3D covariance exactly as the forward saw it ------------------------------------------
    float cov3D[6];
    float3 sc = {0.f, 0.f, 0.f};
    float4 q = {0.f, 0.f, 0.f, 0.f};
    if (p.scales != nullptr) {
        sc = float3{p.scales[3 * i], p.scales[3 * i + 1], p.scales[3 * i + 2]};
        q = float4{p.rotations[4 * i], p.rotations[4 * i + 1], p.rotations[4 * i + 2],
                   p.rotations[4 * i + 3]};
    }
    float q_inv_norm = 1.0f;
    if (kRaw) {
        sc = float3{expf(sc.x), expf(sc.y), expf(sc.z)};
        q_inv_norm = 1.0f / fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
        q = float4{q.x * q_inv_norm, q.y * q_inv_norm, q.z * q_inv_norm, q.w * q_inv_norm};
    }
    if (p.cov3D_precomp != nullptr) {
#pragma unroll
        for (int k = 0; k < 6; ++k) cov3D[k] = p.cov3D_precomp[6 * i + k];
    } else {
        cov3d_from_scale_rot(sc, p.scale_modifier, q, cov3D);
    }

    // ---- K6: conic -> cov2D -> (cov3D, T) -> J -> view-space mean -----------------------------
    // Several terms here cancel catastrophically (denom - a*c == -b*b up to the rounding of denom,
    // the four-term dL/dtz sum), so their value is dominated by WHICH products the reference
    // build fuses into FMAs. To stay within 1e-4 relative of the reference's gradients the
    // roundings below are pinned with intrinsics to the contraction observed in the reference's
    // sm_100a SASS for computeCov2DCUDA (backward.cu:144-274); rn(.) = one IEEE rounding.
    const ProjFrame f = make_proj_frame(mean, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy, s_view);
    const float limx = 1.3f * p.tan_fovx, limy = 1.3f * p.tan_fovy;
    const float x_grad_mul = (f.txtz < -limx || f.txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (f.tytz < -limy || f.tytz > limy) ? 0.f : 1.f;

    // T = W * J with J = [J00 0 J02; 0 J11 J12; 0 0 0]:
    //   T[0][r] = fma(W2r, J02, fma(W0r, J00, rn(0 * W1r))),  T[1][r] = fma(W2r, J12, fma(0, W0r, rn(W1r * J11)))
    const float J00 = f.J.m[0][0], J02 = f.J.m[0][2], J11 = f.J.m[1][1], J12 = f.J.m[1][2];
    float T0[3], T1[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const float w0 = s_view[4 * r], w1 = s_view[4 * r + 1], w2 = s_view[4 * r + 2];
        T0[r] = __fmaf_rn(w2, J02, __fmaf_rn(w0, J00, __fmul_rn(0.0f, w1)));
        T1[r] = __fmaf_rn(w2, J12, __fmaf_rn(0.0f, w0, __fmul_rn(w1, J11)));
    }
    // S0[k] = sum_j T0[j] V[j][k], S1[k] likewise: fma(V2k, T[2], fma(V0k, T[0], rn(V1k * T[1])))
    const float Vm[3][3] = {{cov3D[0], cov3D[1], cov3D[2]},
                            {cov3D[1], cov3D[3], cov3D[4]},
                            {cov3D[2], cov3D[4], cov3D[5]}};
    float S0[3], S1[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        S0[k] = __fmaf_rn(Vm[2][k], T0[2], __fmaf_rn(Vm[0][k], T0[0], __fmul_rn(Vm[1][k], T0[1])));
        S1[k] = __fmaf_rn(Vm[2][k], T1[2], __fmaf_rn(Vm[0][k], T1[0], __fmul_rn(Vm[1][k], T1[1])));
    }
    const float a = __fadd_rn(__fmaf_rn(T0[2], S0[2], __fmaf_rn(T0[1], S0[1], __fmul_rn(T0[0], S0[0]))), 0.3f);
    const float b = __fmaf_rn(T0[2], S1[2], __fmaf_rn(T0[1], S1[1], __fmul_rn(T0[0], S1[0])));
    const float c = __fadd_rn(__fmaf_rn(T1[2], S1[2], __fmaf_rn(T1[1], S1[1], __fmul_rn(T1[0], S1[0]))), 0.3f);

    const float ac = __fmul_rn(a, c);
    const float denom = __fmaf_rn(-b, b, ac);
    const float denom2inv = 1.0f / __fmaf_rn(denom, denom, 0.0000001f);
    float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
    float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (denom2inv != 0) {
        const float two_b = __fadd_rn(b, b), two_a = __fadd_rn(a, a);
        const float d_m_ac = __fadd_rn(denom, -ac);                 // == -b*b up to rounding
        const float d_p_2bb = __fmaf_rn(b, two_b, denom);
        const float two_bc = __fmul_rn(c, two_b), two_ab = __fmul_rn(b, two_a);
        const float bc = __fmul_rn(b, c), ab = __fmul_rn(b, a);
        const float na = __fmaf_rn(gcw, d_m_ac, __fmaf_rn(gcy, two_bc, -__fmul_rn(gcx, __fmul_rn(c, c))));
        const float nc = __fmaf_rn(gcx, d_m_ac, __fmaf_rn(gcy, two_ab, -__fmul_rn(gcw, __fmul_rn(a, a))));
        const float nb = __fmaf_rn(gcw, ab, __fmaf_rn(gcx, bc, -__fmul_rn(gcy, d_p_2bb)));
        dL_da = __fmul_rn(na, denom2inv);
        dL_dc = __fmul_rn(nc, denom2inv);
        dL_db = __fmul_rn(__fadd_rn(denom2inv, denom2inv), nb);
        // diagonal: fma(dc, rn(T1k^2), fma(da, rn(T0k^2), rn(db * rn(T0k*T1k))))
        dcov[0] = __fmaf_rn(dL_dc, __fmul_rn(T1[0], T1[0]), __fmaf_rn(dL_da, __fmul_rn(T0[0], T0[0]), __fmul_rn(dL_db, __fmul_rn(T0[0], T1[0]))));
        dcov[3] = __fmaf_rn(dL_dc, __fmul_rn(T1[1], T1[1]), __fmaf_rn(dL_da, __fmul_rn(T0[1], T0[1]), __fmul_rn(dL_db, __fmul_rn(T0[1], T1[1]))));
        dcov[5] = __fmaf_rn(dL_dc, __fmul_rn(T1[2], T1[2]), __fmaf_rn(dL_da, __fmul_rn(T0[2], T0[2]), __fmul_rn(dL_db, __fmul_rn(T0[2], T1[2]))));
        // off-diagonal (doubled): fma(dc, rn(2T1i*T1j), fma(da, rn(2T0i*T0j), rn(db * fma(T0i, T1j, rn(T0j*T1i)))))
        dcov[1] = __fmaf_rn(dL_dc, __fmul_rn(T1[1], __fadd_rn(T1[0], T1[0])),
                            __fmaf_rn(dL_da, __fmul_rn(T0[1], __fadd_rn(T0[0], T0[0])),
                                      __fmul_rn(dL_db, __fmaf_rn(T0[0], T1[1], __fmul_rn(T0[1], T1[0])))));
        dcov[2] = __fmaf_rn(dL_dc, __fmul_rn(T1[2], __fadd_rn(T1[0], T1[0])),
                            __fmaf_rn(dL_da, __fmul_rn(T0[2], __fadd_rn(T0[0], T0[0])),
                                      __fmul_rn(dL_db, __fmaf_rn(T0[0], T1[2], __fmul_rn(T0[2], T1[0])))));
        dcov[4] = __fmaf_rn(dL_dc, __fmul_rn(T1[2], __fadd_rn(T1[1], T1[1])),
                            __fmaf_rn(dL_da, __fmul_rn(T0[1], __fadd_rn(T0[2], T0[2])),
                                      __fmul_rn(dL_db, __fmaf_rn(T0[1], T1[2], __fmul_rn(T0[2], T1[1])))));
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) p.dL_dcov3D[6 * i + k] = dcov[k];

    // dL/dT0k = fma(2*S0k, da, rn(S1k*db)),  dL/dT1k = fma(S0k, db, rn(2*S1k*dc))
    float dT0[3], dT1[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        dT0[k] = __fmaf_rn(__fadd_rn(S0[k], S0[k]), dL_da, __fmul_rn(S1[k], dL_db));
        dT1[k] = __fmaf_rn(S0[k], dL_db, __fmul_rn(__fadd_rn(S1[k], S1[k]), dL_dc));
    }
    // T = W * J  =>  dL/dJ[c][k] = sum_r W[k][r] dL/dT[c][r] = fma(Wk2, d2, fma(Wk0, d0, rn(Wk1 * d1)))
    const float dL_dJ00 = __fmaf_rn(s_view[8], dT0[2], __fmaf_rn(s_view[0], dT0[0], __fmul_rn(s_view[4], dT0[1])));
    const float dL_dJ02 = __fmaf_rn(s_view[10], dT0[2], __fmaf_rn(s_view[2], dT0[0], __fmul_rn(s_view[6], dT0[1])));
    const float dL_dJ11 = __fmaf_rn(s_view[9], dT1[2], __fmaf_rn(s_view[1], dT1[0], __fmul_rn(s_view[5], dT1[1])));
    const float dL_dJ12 = __fmaf_rn(s_view[10], dT1[2], __fmaf_rn(s_view[2], dT1[0], __fmul_rn(s_view[6], dT1[1])));

    const float tz = 1.f / f.t.z;
    const float tz2 = __fmul_rn(tz, tz);
    const float tz3 = __fmul_rn(tz2, tz);
    const float hx = p.focal_x, hy = p.focal_y;
    const float dL_dtx = __fmul_rn(dL_dJ02, __fmul_rn(tz2, __fmul_rn(x_grad_mul, -hx)));
    const float dL_dty = __fmul_rn(dL_dJ12, __fmul_rn(tz2, __fmul_rn(y_grad_mul, -hy)));
    float dL_dtz = __fmaf_rn(dL_dJ00, __fmul_rn(tz2, -hx), -__fmul_rn(dL_dJ11, __fmul_rn(tz2, hy)));
    dL_dtz = __fmaf_rn(dL_dJ02, __fmul_rn(tz3, __fmul_rn(f.t.x, __fadd_rn(hx, hx))), dL_dtz);
    dL_dtz = __fmaf_rn(dL_dJ12, __fmul_rn(tz3, __fmul_rn(f.t.y, __fadd_rn(hy, hy))), dL_dtz);
    // back through t = view * mean (rotation part transposed): fma(dtz, v2, fma(dtx, v0, rn(dty * v1)))
    V3 dmean = v3(__fmaf_rn(dL_dtz, s_view[2], __fmaf_rn(dL_dtx, s_view[0], __fmul_rn(dL_dty, s_view[1]))),
                  __fmaf_rn(dL_dtz, s_view[6], __fmaf_rn(dL_dtx, s_view[4], __fmul_rn(dL_dty, s_view[5]))),
                  __fmaf_rn(dL_dtz, s_view[10], __fmaf_rn(dL_dtx, s_view[8], __fmul_rn(dL_dty, s_view[9]))));

    // ---- K7a: screen-space mean -> 3D mean through the perspective divide ---------------------
    {
        const float4 m_hom = xform_4x4(mean, s_proj);
        const float m_w = 1.0f / (m_hom.w + 0.0000001f);
        const float mul1 = (s_proj[0] * mean.x + s_proj[4] * mean.y + s_proj[8] * mean.z + s_proj[12]) * m_w * m_w;
        const float mul2 = (s_proj[1] * mean.x + s_proj[5] * mean.y + s_proj[9] * mean.z + s_proj[13]) * m_w * m_w;
        V3 d;
        d.x = (s_proj[0] * m_w - s_proj[3] * mul1) * gmx + (s_proj[1] * m_w - s_proj[3] * mul2) * gmy;
        d.y = (s_proj[4] * m_w - s_proj[7] * mul1) * gmx + (s_proj[5] * m_w - s_proj[7] * mul2) * gmy;
        d.z = (s_proj[8] * m_w - s_proj[11] * mul1) * gmx + (s_proj[9] * m_w - s_proj[11] * mul2) * gmy;
        dmean = dmean + d;
    }

    // ---- K7b: colour -> SH coefficients and view direction ------------------------------------
    if (p.shs != nullptr) {
        const uint32_t clampbits = __float_as_uint(__ldg(&p.rec[idx].q2).w);
        V3 dRGB = gcol;
        if (clampbits & 1u) dRGB.x = 0.f;
        if (clampbits & 2u) dRGB.y = 0.f;
        if (clampbits & 4u) dRGB.z = 0.f;
        factor = dRGB;

        const V3 dir_orig = v3(mean.x, mean.y, mean.z) - v3(s_cam[0], s_cam[1], s_cam[2]);
        const V3 dir = dir_orig / length3(dir_orig);
        const float x = dir.x, y = dir.y, z = dir.z;
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;

        // basis values and their partial derivatives w.r.t. the unit direction
        float bas[16], bx[16], by[16], bz[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) { bas[k] = 0.f; bx[k] = 0.f; by[k] = 0.f; bz[k] = 0.f; }
        bas[0] = kC0;
        bas[1] = -kC1 * y;                 by[1] = -kC1;
        bas[2] = kC1 * z;                  bz[2] = kC1;
        bas[3] = -kC1 * x;                 bx[3] = -kC1;
        bas[4] = kC2_0 * xy;               bx[4] = kC2_0 * y;          by[4] = kC2_0 * x;
        bas[5] = kC2_1 * yz;               by[5] = kC2_1 * z;          bz[5] = kC2_1 * y;
        bas[6] = kC2_2 * (2.f * zz - xx - yy);
        bx[6] = kC2_2 * 2.f * -x;          by[6] = kC2_2 * 2.f * -y;   bz[6] = kC2_2 * 2.f * 2.f * z;
        bas[7] = kC2_3 * xz;               bx[7] = kC2_3 * z;          bz[7] = kC2_3 * x;
        bas[8] = kC2_4 * (xx - yy);        bx[8] = kC2_4 * 2.f * x;    by[8] = kC2_4 * 2.f * -y;
        bas[9] = kC3_0 * y * (3.f * xx - yy);
        bx[9] = kC3_0 * 3.f * 2.f * xy;    by[9] = kC3_0 * 3.f * (xx - yy);
        bas[10] = kC3_1 * xy * z;
        bx[10] = kC3_1 * yz;               by[10] = kC3_1 * xz;        bz[10] = kC3_1 * xy;
        bas[11] = kC3_2 * y * (4.f * zz - xx - yy);
        bx[11] = kC3_2 * -2.f * xy;        by[11] = kC3_2 * (-3.f * yy + 4.f * zz - xx);
        bz[11] = kC3_2 * 4.f * 2.f * yz;
        bas[12] = kC3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
        bx[12] = kC3_3 * -3.f * 2.f * xz;  by[12] = kC3_3 * -3.f * 2.f * yz;
        bz[12] = kC3_3 * 3.f * (2.f * zz - xx - yy);
        bas[13] = kC3_4 * x * (4.f * zz - xx - yy);
        bx[13] = kC3_4 * (-3.f * xx + 4.f * zz - yy);  by[13] = kC3_4 * -2.f * xy;
        bz[13] = kC3_4 * 4.f * 2.f * xz;
        bas[14] = kC3_5 * z * (xx - yy);
        bx[14] = kC3_5 * 2.f * xz;         by[14] = kC3_5 * -2.f * yz; bz[14] = kC3_5 * (xx - yy);
        bas[15] = kC3_6 * x * (xx - 3.f * yy);
        bx[15] = kC3_6 * 3.f * (xx - yy);  by[15] = kC3_6 * -3.f * 2.f * xy;

        V3 dRGBdx = v3(0.f, 0.f, 0.f), dRGBdy = v3(0.f, 0.f, 0.f), dRGBdz = v3(0.f, 0.f, 0.f);
        if (kVecSH) {
            // the lane's SH row is in its tile row (cooperative load in the kernel); it is consumed
            // four coefficients (three 128-bit shared loads) at a time, then the row is reused
            // for the basis values and dL/dRGB of the cooperative store
            float4* dst = reinterpret_cast<float4*>(p.dL_dsh) + i * 12;
            const float4* row4 = reinterpret_cast<const float4*>(tile_row);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                float sh[12];
                if (4 * c < ncoef) {
#pragma unroll
                    for (int q = 0; q < 3; ++q) {
                        const float4 v = row4[3 * c + q];
                        sh[4 * q] = v.x; sh[4 * q + 1] = v.y; sh[4 * q + 2] = v.z; sh[4 * q + 3] = v.w;
                    }
                } else {
#pragma unroll
                    for (int q = 0; q < 12; ++q) sh[q] = 0.f;
                }
                float out[12];
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const int k = 4 * c + kk;
                    const bool on = k < ncoef;
                    const float bk = on ? bas[k] : 0.f;
                    if (!kFactor && !kCoop) {
                        out[3 * kk] = bk * dRGB.x; out[3 * kk + 1] = bk * dRGB.y; out[3 * kk + 2] = bk * dRGB.z;
                    }
                    if (on && k > 0) {
                        const V3 s = v3(sh[3 * kk], sh[3 * kk + 1], sh[3 * kk + 2]);
                        dRGBdx = dRGBdx + bx[k] * s;
                        dRGBdy = dRGBdy + by[k] * s;
                        dRGBdz = dRGBdz + bz[k] * s;
                    }
                }
                if (!kFactor && !kCoop) {
#pragma unroll
                    for (int q = 0; q < 3; ++q)
                        dst[3 * c + q] = make_float4(out[4 * q], out[4 * q + 1], out[4 * q + 2], out[4 * q + 3]);
                }
            }
            if (kCoop) {
#pragma unroll
                for (int k = 0; k < 16; ++k) tile_row[k] = k < ncoef ? bas[k] : 0.f;
                tile_row[16] = dRGB.x; tile_row[17] = dRGB.y; tile_row[18] = dRGB.z;
            }
            if (kRaw) store3(p.dL_dsh + 3 * i, bas[0] * dRGB.x, bas[0] * dRGB.y, bas[0] * dRGB.z);
        } else {
            const float* src = p.shs + i * p.M * 3;
            float* dst = p.dL_dsh + i * p.M * 3;
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                if (k < p.M) {
                    const bool on = k < ncoef;
                    const float bk = on ? bas[k] : 0.f;
                    if (!kFactor) {
                        dst[3 * k] = bk * dRGB.x; dst[3 * k + 1] = bk * dRGB.y; dst[3 * k + 2] = bk * dRGB.z;
                    }
                    if (on && k > 0) {
                        const V3 s = v3(src[3 * k], src[3 * k + 1], src[3 * k + 2]);
                        dRGBdx = dRGBdx + bx[k] * s;
                        dRGBdy = dRGBdy + by[k] * s;
                        dRGBdz = dRGBdz + bz[k] * s;
                    }
                }
            }
            if (!kFactor)
                for (int k = 16; k < p.M; ++k) store3(dst + 3 * k, 0.f, 0.f, 0.f);
        }

        const V3 dL_ddir = v3(dot3(dRGBdx, dRGB), dot3(dRGBdy, dRGB), dot3(dRGBdz, dRGB));
        // derivative of v/|v|
        const V3 v = dir_orig;
        const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
        const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
        V3 dn;
        dn.x = ((+sum2 - v.x * v.x) * dL_ddir.x - v.y * v.x * dL_ddir.y - v.z * v.x * dL_ddir.z) * invsum32;
        dn.y = (-v.x * v.y * dL_ddir.x + (sum2 - v.y * v.y) * dL_ddir.y - v.z * v.y * dL_ddir.z) * invsum32;
        dn.z = (-v.x * v.z * dL_ddir.x - v.y * v.z * dL_ddir.y + (sum2 - v.z * v.z) * dL_ddir.z) * invsum32;
        dmean = dmean + dn;
    }
    store3(p.dL_dmean3D + 3 * i, dmean.x, dmean.y, dmean.z);

    // ---- K7c: 3D covariance -> scale and (unnormalised) quaternion ----------------------------
    if (p.scales != nullptr) {
        const float r = q.x, x = q.y, y = q.z, z = q.w;
        const M3 Rg = quat_to_m3(r, x, y, z);
        // Rs[a][b]: row a, column b of the matrix the reference calls R (Rg is column-major)
        float Rs[3][3], Ms[3][3];
        const float s[3] = {p.scale_modifier * sc.x, p.scale_modifier * sc.y, p.scale_modifier * sc.z};
#pragma unroll
        for (int aa = 0; aa < 3; ++aa)
#pragma unroll
            for (int bb = 0; bb < 3; ++bb) {
                Rs[aa][bb] = Rg.m[bb][aa];
                Ms[aa][bb] = s[aa] * Rs[aa][bb];
            }
        // symmetric dL/dSigma with halved off-diagonals
        const float dS[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                                {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                                {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
        float dM[3][3];  // dL/dM = 2 M dSigma
#pragma unroll
        for (int aa = 0; aa < 3; ++aa)
#pragma unroll
            for (int bb = 0; bb < 3; ++bb)
                dM[aa][bb] = 2.0f * Ms[aa][0] * dS[0][bb] + 2.0f * Ms[aa][1] * dS[1][bb] +
                             2.0f * Ms[aa][2] * dS[2][bb];
        float dscale[3];
        float Gm[3][3];  // dL/dR
#pragma unroll
        for (int aa = 0; aa < 3; ++aa) {
            dscale[aa] = Rs[aa][0] * dM[aa][0] + Rs[aa][1] * dM[aa][1] + Rs[aa][2] * dM[aa][2];
#pragma unroll
            for (int bb = 0; bb < 3; ++bb) Gm[aa][bb] = s[aa] * dM[aa][bb];
        }
        if (kRaw) {   // d exp(x) = exp(x)
            dscale[0] *= sc.x; dscale[1] *= sc.y; dscale[2] *= sc.z;
        }
        store3(p.dL_dscale + 3 * i, dscale[0], dscale[1], dscale[2]);

        float4 dq;
        dq.x = 2 * z * (Gm[0][1] - Gm[1][0]) + 2 * y * (Gm[2][0] - Gm[0][2]) + 2 * x * (Gm[1][2] - Gm[2][1]);
        dq.y = 2 * y * (Gm[1][0] + Gm[0][1]) + 2 * z * (Gm[2][0] + Gm[0][2]) + 2 * r * (Gm[1][2] - Gm[2][1]) - 4 * x * (Gm[2][2] + Gm[1][1]);
        dq.z = 2 * x * (Gm[1][0] + Gm[0][1]) + 2 * r * (Gm[2][0] - Gm[0][2]) + 2 * z * (Gm[1][2] + Gm[2][1]) - 4 * y * (Gm[2][2] + Gm[0][0]);
        dq.w = 2 * r * (Gm[0][1] - Gm[1][0]) + 2 * x * (Gm[2][0] + Gm[0][2]) + 2 * y * (Gm[1][2] + Gm[2][1]) - 4 * z * (Gm[1][1] + Gm[0][0]);
        if (kRaw) {   // through q / |q|: (g - q (q . g)) / |q| with q already unit length
            const float qg = r * dq.x + x * dq.y + y * dq.z + z * dq.w;
            dq.x = (dq.x - r * qg) * q_inv_norm; dq.y = (dq.y - x * qg) * q_inv_norm;
            dq.z = (dq.z - y * qg) * q_inv_norm; dq.w = (dq.w - z * qg) * q_inv_norm;
        }
        if ((reinterpret_cast<uintptr_t>(p.dL_drot) & 15) == 0) {
            reinterpret_cast<float4*>(p.dL_drot)[i] = dq;
        } else {
            p.dL_drot[4 * i] = dq.x; p.dL_drot[4 * i + 1] = dq.y;
            p.dL_drot[4 * i + 2] = dq.z; p.dL_drot[4 * i + 3] = dq.w;
        }
    } else {
        store3(p.dL_dscale + 3 * i, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < 4; ++k) p.dL_drot[4 * i + k] = 0.f;
    }
    return factor;
}

__device__ __forceinline__ void mm_st4(float4* p, float4 v)
{
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
                 "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// kFactor (view-parallel path): instead of writing 3M floats of dL/dsh per Gaussian, the block
// stages its 256 x 3 colour-gradient factors in shared memory and MULTICASTS them (multimem.st:
// one store, the NVSwitch replicates it) into slot `dp_rank` of a symmetric buffer on every GPU —
// the all-gather is done by the kernel that produces the data, while it is still computing. Each
// rank later rebuilds sum_v basis(dir_v) x factor_v locally (sh_grad_from_factors_kernel).
template <bool kVecSH, bool kFactor, bool kRaw>
__global__ void __launch_bounds__(PB_THREADS) preprocess_backward_kernel(PreprocessBackwardParams p)
{
    static_assert(!kRaw || (kVecSH && !kFactor), "the raw-leaf path uses the tile and writes its SH gradients");
    __shared__ float s_view[16], s_proj[16], s_cam[3];
    __shared__ __align__(16) float s_factor[kFactor ? 3 * PB_THREADS : 4];
    if (threadIdx.x < 16) {
        s_view[threadIdx.x] = p.viewmatrix[threadIdx.x];
        s_proj[threadIdx.x] = p.projmatrix[threadIdx.x];
    }
    if (threadIdx.x < 3) s_cam[threadIdx.x] = p.cam_pos[threadIdx.x];
    __syncthreads();

    constexpr bool kCoop = kVecSH && !kFactor;
    __shared__ __align__(16) float s_tile[kVecSH ? PB_THREADS * SH_TILE_STRIDE : 4];
    float* tile_row = kVecSH ? s_tile + threadIdx.x * SH_TILE_STRIDE : nullptr;

    const int block_first = p.begin + blockIdx.x * PB_THREADS;
    const int idx = block_first + threadIdx.x;
    const bool visible = idx < p.end && p.radii[idx] > 0;
    // raw leaves: the warp's 32 rows of _features_rest are one contiguous block of 32 x rw floats
    // (rw = 3 (M - 1), any alignment); element e of it belongs to row e / rw. raw_div = ceil(2^16 /
    // rw) turns the division into a multiply for e < 32 x 48.
    const int rw = kRaw ? 3 * (p.M - 1) : 0;
    const uint32_t raw_div = kRaw && rw > 0 ? (65536u + rw - 1) / rw : 0u;
    if (kRaw) {
        const int lane = threadIdx.x & 31;
        const unsigned vis = __ballot_sync(0xffffffffu, visible);
        const int row0 = idx - lane;
        const int live = 3 * ((p.D + 1) * (p.D + 1) - 1);     // floats of a rest row that are used
        float* tile = s_tile + (threadIdx.x - lane) * SH_TILE_STRIDE;
        if (visible) {
            const float* dc = p.shs + 3 * (size_t)idx;
            tile_row[0] = dc[0]; tile_row[1] = dc[1]; tile_row[2] = dc[2];
        }
        const float* src = p.shs_rest + (size_t)row0 * rw;
        for (int e = lane; e < 32 * rw; e += 32) {
            const int r = (int)(((uint32_t)e * raw_div) >> 16), c = e - r * rw;
            if (((vis >> r) & 1u) && c < live) tile[r * SH_TILE_STRIDE + 3 + c] = __ldg(src + e);
        }
        __syncwarp();
    } else if (kVecSH) {
        // cooperative load of the visible lanes' SH rows: float4 q of the warp's 32 x 12 block
        // belongs to row q / 12; a row is fetched by twelve neighbouring lanes (six whole sectors)
        const int lane = threadIdx.x & 31;
        const unsigned vis = __ballot_sync(0xffffffffu, visible);
        const int row0 = idx - lane;
        const int nvec = (3 * (p.D + 1) * (p.D + 1) + 3) >> 2;   // float4 per row that hold live coefficients
        const float4* src = reinterpret_cast<const float4*>(p.shs) + (size_t)row0 * 12;
        float* tile = s_tile + (threadIdx.x - lane) * SH_TILE_STRIDE;
        float4 v[12];
#pragma unroll
        for (int j = 0; j < 12; ++j) {
            const int q = j * 32 + lane, r = q / 12, c = q - r * 12;
            v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (((vis >> r) & 1u) && c < nvec) v[j] = __ldg(src + q);
        }
#pragma unroll
        for (int j = 0; j < 12; ++j) {
            const int q = j * 32 + lane, r = q / 12, c = q - r * 12;
            if ((vis >> r) & 1u) *reinterpret_cast<float4*>(tile + r * SH_TILE_STRIDE + 4 * c) = v[j];
        }
        __syncwarp();
    }
    V3 f = v3(0.f, 0.f, 0.f);
    if (idx < p.end)
        f = preprocess_backward_body<kVecSH, kFactor, kRaw>(p, idx, visible, s_view, s_proj, s_cam, tile_row);
    if (kRaw) {
        // dL/d_features_rest[r][k-1][c] = basis_k(r) * dL/dRGB_c(r), k >= 1: the warp's block of
        // 32 x rw floats, 128 contiguous bytes per store instruction
        __syncwarp();
        const int lane = threadIdx.x & 31;
        const int row0 = idx - lane;
        const int rows = min(32, p.end - row0);
        const float* tile = s_tile + (threadIdx.x - lane) * SH_TILE_STRIDE;
        float* dst = p.dL_dsh_rest + (size_t)row0 * rw;
        for (int e = lane; e < rows * rw; e += 32) {
            const int r = (int)(((uint32_t)e * raw_div) >> 16), c = e - r * rw;
            const int k = (c * 21846) >> 16;                  // c / 3 for c < 48
            const float* row = tile + r * SH_TILE_STRIDE;
            dst[e] = row[k + 1] * row[16 + c - 3 * k];
        }
    } else if (kCoop) {
        // dL_dsh[r][k][c] = basis_k(r) * dL/dRGB_c(r): the warp's 32 rows, 384 float4, written in
        // twelve instructions of 512 contiguous bytes each
        __syncwarp();
        const int lane = threadIdx.x & 31;
        const int row0 = idx - lane;
        const int rows = min(32, p.end - row0);
        const float* tile = s_tile + (threadIdx.x - lane) * SH_TILE_STRIDE;
        float4* dst = reinterpret_cast<float4*>(p.dL_dsh) + (size_t)row0 * 12;
#pragma unroll
        for (int j = 0; j < 12; ++j) {
            const int q = j * 32 + lane;          // float4 index inside the warp's block
            const int r = q / 12;
            if (r < rows) {
                const float* row = tile + r * SH_TILE_STRIDE;
                const int f0 = (q - r * 12) * 4;  // first of the four floats: k = f / 3, c = f % 3
                float o[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const int fl = f0 + t, k = fl / 3;
                    o[t] = row[k] * row[16 + fl - 3 * k];
                }
                dst[q] = make_float4(o[0], o[1], o[2], o[3]);
            }
        }
    }
    if (kFactor) {
        s_factor[3 * threadIdx.x] = f.x;
        s_factor[3 * threadIdx.x + 1] = f.y;
        s_factor[3 * threadIdx.x + 2] = f.z;
        __syncthreads();
        // 3 floats per thread; begin and end are multiples of 4 (checked by the host), so
        // the block's first float is 16-byte aligned and the valid part of the last block is a
        // whole number of float4 too
        const size_t first = (size_t)block_first * 3;                        // float index in the slot
        const size_t valid = (size_t)min(PB_THREADS, p.end - block_first) * 3;
        float4* dst = reinterpret_cast<float4*>(p.factor_mc + first);
        if (threadIdx.x * 4 < valid)
            mm_st4(dst + threadIdx.x, reinterpret_cast<const float4*>(s_factor)[threadIdx.x]);
        if (block_first == 0 && threadIdx.x < 3)
            asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p.campos_mc + threadIdx.x),
                         "f"(s_cam[threadIdx.x])
                         : "memory");
    }
}

// dL/dsh[i][k][c] = sum over views v of basis_k(dir_v(i)) * factor_v[i][c]   (k < (D+1)^2, else 0),
// summed in rank order, i.e. deterministically and identically on every rank.
constexpr int SHG_THREADS = 128;   // 4 warps x 6.1 KB of transpose tile

__global__ void __launch_bounds__(SHG_THREADS)
sh_grad_from_factors_kernel(int P, int begin, int end, int D, int M, int world,
                            const float* __restrict__ means3D, const float* __restrict__ campos_all,
                            const float* __restrict__ factors_all, float* __restrict__ dL_dsh)
{
    const int idx = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (idx - (int)(threadIdx.x & 31) >= end) return;          // whole warp past the range
    const bool live = idx < end;                               // tail lanes help with the stores
    const size_t i = (size_t)(live ? idx : end - 1);
    const int ncoef = (D + 1) * (D + 1);
    float acc[48];
#pragma unroll
    for (int k = 0; k < 48; ++k) acc[k] = 0.f;
    const V3 mean = v3(means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]);
    if (!live) world = 0;
    // Views in groups of four: the group's factor loads are issued together (the loop is bound
    // by the latency of these loads, not by the arithmetic), then consumed in view order.
    constexpr int kGroup = 4;
    for (int v0 = 0; v0 < world; v0 += kGroup) {
        V3 gv[kGroup];
#pragma unroll
        for (int u = 0; u < kGroup; ++u) {
            gv[u] = v3(0.f, 0.f, 0.f);
            if (v0 + u < world) {
                const float* fsrc = factors_all + ((size_t)(v0 + u) * P + i) * 3;
                gv[u] = v3(__ldcs(fsrc), __ldcs(fsrc + 1), __ldcs(fsrc + 2));
            }
        }
#pragma unroll
        for (int u = 0; u < kGroup; ++u) {
            const V3 g = gv[u];
            // not seen (or fully clamped) in this view; also skips the padding of the last group
            if (g.x == 0.f && g.y == 0.f && g.z == 0.f) continue;
            const int v = v0 + u;
            const V3 cam = v3(campos_all[4 * v], campos_all[4 * v + 1], campos_all[4 * v + 2]);
            const V3 dir_orig = mean - cam;
            const V3 dir = dir_orig / length3(dir_orig);
            const float x = dir.x, y = dir.y, z = dir.z;
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            float bas[16];
            bas[0] = kC0;
            bas[1] = -kC1 * y; bas[2] = kC1 * z; bas[3] = -kC1 * x;
            bas[4] = kC2_0 * xy; bas[5] = kC2_1 * yz; bas[6] = kC2_2 * (2.f * zz - xx - yy);
            bas[7] = kC2_3 * xz; bas[8] = kC2_4 * (xx - yy);
            bas[9] = kC3_0 * y * (3.f * xx - yy); bas[10] = kC3_1 * xy * z;
            bas[11] = kC3_2 * y * (4.f * zz - xx - yy);
            bas[12] = kC3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
            bas[13] = kC3_4 * x * (4.f * zz - xx - yy); bas[14] = kC3_5 * z * (xx - yy);
            bas[15] = kC3_6 * x * (xx - 3.f * yy);
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                if (k < ncoef) {
                    // product rounded first, then added: the same value a plain sum of the
                    // per-view dL/dsh (bas*g rounded by K7) would accumulate
                    acc[3 * k] = __fadd_rn(acc[3 * k], __fmul_rn(bas[k], g.x));
                    acc[3 * k + 1] = __fadd_rn(acc[3 * k + 1], __fmul_rn(bas[k], g.y));
                    acc[3 * k + 2] = __fadd_rn(acc[3 * k + 2], __fmul_rn(bas[k], g.z));
                }
            }
        }
    }
    if (M == 16 && (reinterpret_cast<uintptr_t>(dL_dsh) & 15) == 0) {
        // A thread's row is 192 contiguous bytes; written directly, every store instruction of a
        // warp would touch 32 half sectors 192 bytes apart. Transpose the warp's 32 x 48 tile
        // through shared memory (row stride 49 words: conflict free) so that each instruction
        // writes 512 contiguous bytes.
        __shared__ float s_tile[SHG_THREADS / 32][32 * 49];
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        float* tile = s_tile[warp];
#pragma unroll
        for (int k = 0; k < 48; ++k) tile[lane * 49 + k] = acc[k];
        __syncwarp();
        const int row0 = idx - lane;                           // first Gaussian of this warp
        const int rows = min(32, end - row0);
        float4* dst = reinterpret_cast<float4*>(dL_dsh) + (size_t)row0 * 12;
#pragma unroll
        for (int j = 0; j < 12; ++j) {
            const int q = j * 32 + lane;                       // float4 index inside the tile
            const int r = q / 12, c = (q - r * 12) * 4;
            if (r < rows) {
                const float* src = tile + r * 49 + c;
                __stcs(dst + q, make_float4(src[0], src[1], src[2], src[3]));
            }
        }
    } else if (live) {
        float* dst = dL_dsh + i * M * 3;
#pragma unroll
        for (int k = 0; k < 48; ++k)
            if (k < M * 3) dst[k] = acc[k];
        for (int k = 48; k < M * 3; ++k) dst[k] = 0.f;
    }
}

}  // namespace

void launch_sh_grad_from_factors(int P, int begin, int end, int D, int M, int world,
                                 const float* means3D, const float* campos_all,
                                 const float* factors_all, float* dL_dsh, cudaStream_t s)
{
    if (end <= begin) return;
    ++g_kernel_launches;
    sh_grad_from_factors_kernel<<<(end - begin + SHG_THREADS - 1) / SHG_THREADS, SHG_THREADS, 0, s>>>(
        P, begin, end, D, M, world, means3D, campos_all, factors_all, dL_dsh);
}

void launch_preprocess_backward(const PreprocessBackwardParams& p, cudaStream_t s)
{
    if (p.end <= p.begin) return;
    const int blocks = (p.end - p.begin + PB_THREADS - 1) / PB_THREADS;
    const bool factor = p.factor_mc != nullptr && p.shs != nullptr;
    const bool vec = p.shs != nullptr && p.M == 16 &&
                     (reinterpret_cast<uintptr_t>(p.shs) & 15) == 0 &&
                     (factor || (reinterpret_cast<uintptr_t>(p.dL_dsh) & 15) == 0);
    ++g_kernel_launches;
    if (p.raw) {
        preprocess_backward_kernel<true, false, true><<<blocks, PB_THREADS, 0, s>>>(p);
    } else if (factor) {
        if (vec)
            preprocess_backward_kernel<true, true, false><<<blocks, PB_THREADS, 0, s>>>(p);
        else
            preprocess_backward_kernel<false, true, false><<<blocks, PB_THREADS, 0, s>>>(p);
    } else if (vec) {
        preprocess_backward_kernel<true, false, false><<<blocks, PB_THREADS, 0, s>>>(p);
    } else {
        preprocess_backward_kernel<false, false, false><<<blocks, PB_THREADS, 0, s>>>(p);
    }
}

}  // namespace bgr
