// Small column-major 3x3 / vec3 helpers and the shared projection math of the preprocess kernels.
//
// The integer outputs of the rasterizer (radii, tiles_touched, sort keys, n_contrib) are compared
// bit-for-bit with the reference, so the floating-point association order of every expression
// that feeds them matters. The reference evaluates these through glm:
//   mat3*mat3 : R[c][r] = A[0][r]*B[c][0] + A[1][r]*B[c][1] + A[2][r]*B[c][2], left to right,
//               including the products with structural zeros;
//   dot(vec3) : x*x' + y*y' + z*z' left to right;  length = sqrt(dot(v,v));  v/s = 3 divisions.
// The helpers below state that arithmetic directly (no glm), so that nvcc's default FMA
// contraction sees the same expression trees.
#pragma once
#include <cuda_runtime.h>

namespace bgr {

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 v3(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(float s, V3 v) { return V3{s * v.x, s * v.y, s * v.z}; }
__device__ __forceinline__ V3 operator*(V3 v, float s) { return V3{v.x * s, v.y * s, v.z * s}; }
__device__ __forceinline__ V3 operator/(V3 v, float s) { return V3{v.x / s, v.y / s, v.z / s}; }
__device__ __forceinline__ float dot3(V3 a, V3 b)
{
    float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z;
    return tx + ty + tz;
}
__device__ __forceinline__ float length3(V3 v) { return sqrtf(dot3(v, v)); }

// Column-major: m[c][r].
struct M3 {
    float m[3][3];
};
// Arguments are listed column by column (first three = column 0), like the reference's constructor.
__device__ __forceinline__ M3 m3_cols(float a, float b, float c, float d, float e, float f, float g,
                                      float h, float i)
{
    M3 r;
    r.m[0][0] = a; r.m[0][1] = b; r.m[0][2] = c;
    r.m[1][0] = d; r.m[1][1] = e; r.m[1][2] = f;
    r.m[2][0] = g; r.m[2][1] = h; r.m[2][2] = i;
    return r;
}
__device__ __forceinline__ M3 m3_mul(const M3& A, const M3& B)
{
    M3 R;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int r = 0; r < 3; ++r)
            R.m[c][r] = A.m[0][r] * B.m[c][0] + A.m[1][r] * B.m[c][1] + A.m[2][r] * B.m[c][2];
    return R;
}
__device__ __forceinline__ M3 m3_transpose(const M3& A)
{
    M3 R;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int r = 0; r < 3; ++r) R.m[c][r] = A.m[r][c];
    return R;
}

// SH basis constants (the standard real SH normalisation constants, auxiliary.h:22-39).
__device__ constexpr float kC0 = 0.28209479177387814f;
__device__ constexpr float kC1 = 0.4886025119029199f;
__device__ constexpr float kC2_0 = 1.0925484305920792f;
__device__ constexpr float kC2_1 = -1.0925484305920792f;
__device__ constexpr float kC2_2 = 0.31539156525252005f;
__device__ constexpr float kC2_3 = -1.0925484305920792f;
__device__ constexpr float kC2_4 = 0.5462742152960396f;
__device__ constexpr float kC3_0 = -0.5900435899266435f;
__device__ constexpr float kC3_1 = 2.890611442640554f;
__device__ constexpr float kC3_2 = -0.4570457994644658f;
__device__ constexpr float kC3_3 = 0.3731763325901154f;
__device__ constexpr float kC3_4 = -0.4570457994644658f;
__device__ constexpr float kC3_5 = 1.445305721320277f;
__device__ constexpr float kC3_6 = -0.5900435899266435f;

// out_i = m[i]*x + m[4+i]*y + m[8+i]*z + m[12+i]   (auxiliary.h:58-77)
__device__ __forceinline__ float3 xform_4x3(float3 p, const float* m)
{
    return float3{
        m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
        m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
        m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
    };
}
__device__ __forceinline__ float4 xform_4x4(float3 p, const float* m)
{
    return float4{
        m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
        m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
        m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
        m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15],
    };
}

// NDC -> pixel: the reference writes the literals without an f suffix, so the arithmetic is done
// in double precision and rounded once to float (auxiliary.h:41-44).
__device__ __forceinline__ float ndc_to_pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// Tile rectangle of a splat with integer radius (auxiliary.h:46-56).
__device__ __forceinline__ void tile_rect(float2 p, int max_radius, int grid_x, int grid_y,
                                          uint2& rmin, uint2& rmax)
{
    const unsigned gx = (unsigned)grid_x, gy = (unsigned)grid_y;
    rmin.x = min(gx, (unsigned)max((int)0, (int)((p.x - max_radius) / TILE_X)));
    rmin.y = min(gy, (unsigned)max((int)0, (int)((p.y - max_radius) / TILE_Y)));
    rmax.x = min(gx, (unsigned)max((int)0, (int)((p.x + max_radius + TILE_X - 1) / TILE_X)));
    rmax.y = min(gy, (unsigned)max((int)0, (int)((p.y + max_radius + TILE_Y - 1) / TILE_Y)));
}

// Rotation matrix of the (unnormalised) quaternion (r,x,y,z), listed column by column.
__device__ __forceinline__ M3 quat_to_m3(float r, float x, float y, float z)
{
    // The nine entries are sums/differences of two products; which of the two products nvcc
    // fuses into the FMA decides the last bit of the conic. Pinned to the contraction the
    // reference build shows on sm_100a (SASS of preprocessCUDA: e.g. x*z + r*y = fma(r, y, rn(x*z)),
    // y*y + z*z = rn(y*y) + rn(z*z)), checked bit-for-bit by the GPU parity tests.
    const float rz = __fmul_rn(r, z), rx = __fmul_rn(r, x), xz = __fmul_rn(x, z);
    const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
    const float xy_m_rz = __fmaf_rn(x, y, -rz), xy_p_rz = __fmaf_rn(x, y, rz);
    const float xz_p_ry = __fmaf_rn(r, y, xz), xz_m_ry = __fmaf_rn(-r, y, xz);
    const float yz_m_rx = __fmaf_rn(y, z, -rx), yz_p_rx = __fmaf_rn(y, z, rx);
    const float yy_zz = __fadd_rn(yy, zz);
    const float xx_zz = __fmaf_rn(x, x, zz), xx_yy = __fmaf_rn(x, x, yy);
    return m3_cols(
        1.f - 2.f * yy_zz, 2.f * xy_m_rz, 2.f * xz_p_ry,
        2.f * xy_p_rz, 1.f - 2.f * xx_zz, 2.f * yz_m_rx,
        2.f * xz_m_ry, 2.f * yz_p_rx, 1.f - 2.f * xx_yy);
}

// Sigma = (S R)^T (S R); six unique entries [00,01,02,11,12,22]   (forward.cu:118-152)
__device__ __forceinline__ void cov3d_from_scale_rot(float3 scale, float mod, float4 rot,
                                                     float* cov3D)
{
    M3 S = m3_cols(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
    S.m[0][0] = mod * scale.x;
    S.m[1][1] = mod * scale.y;
    S.m[2][2] = mod * scale.z;
    const M3 R = quat_to_m3(rot.x, rot.y, rot.z, rot.w);
    const M3 M = m3_mul(S, R);
    const M3 Sigma = m3_mul(m3_transpose(M), M);
    cov3D[0] = Sigma.m[0][0];
    cov3D[1] = Sigma.m[0][1];
    cov3D[2] = Sigma.m[0][2];
    cov3D[3] = Sigma.m[1][1];
    cov3D[4] = Sigma.m[1][2];
    cov3D[5] = Sigma.m[2][2];
}

// View-space mean with the 1.3*tan(fov) guard band, Jacobian J, W, T = W*J   (forward.cu:80-99)
struct ProjFrame {
    float3 t;          // clamped view-space position
    float txtz, tytz;  // unclamped ratios (the backward needs them for the clamp mask)
    M3 J, W, T;
};
__device__ __forceinline__ ProjFrame make_proj_frame(float3 mean, float focal_x, float focal_y,
                                                     float tan_fovx, float tan_fovy,
                                                     const float* view)
{
    ProjFrame f;
    f.t = xform_4x3(mean, view);
    const float limx = 1.3f * tan_fovx;
    const float limy = 1.3f * tan_fovy;
    f.txtz = f.t.x / f.t.z;
    f.tytz = f.t.y / f.t.z;
    f.t.x = fminf(limx, fmaxf(-limx, f.txtz)) * f.t.z;
    f.t.y = fminf(limy, fmaxf(-limy, f.tytz)) * f.t.z;
    f.J = m3_cols(focal_x / f.t.z, 0.0f, -(focal_x * f.t.x) / (f.t.z * f.t.z),
                  0.0f, focal_y / f.t.z, -(focal_y * f.t.y) / (f.t.z * f.t.z),
                  0.0f, 0.0f, 0.0f);
    f.W = m3_cols(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
    f.T = m3_mul(f.W, f.J);
    return f;
}

// cov2D = T^T * Vrk^T * T (left-associated), entries (00, 01, 11) before the low-pass.
__device__ __forceinline__ M3 cov2d_full(const M3& T, const float* cov3D)
{
    const M3 Vrk = m3_cols(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2],
                           cov3D[4], cov3D[5]);
    return m3_mul(m3_mul(m3_transpose(T), m3_transpose(Vrk)), T);
}

}  // namespace bgr
