// Forward per-Gaussian preprocess (K1) and the near-plane visibility test (K8).
//
// Behavioural spec: SURVEY.md App. A.2/A.3 (reference forward.cu:155-256, auxiliary.h:139-164).
// One thread per Gaussian. Instead of the reference's eight separate scratch arrays the kernel
// writes one 48-byte record per visible Gaussian (bgr_common.cuh: GaussRec) plus radii and
// tiles_touched for every Gaussian. It also adds the four corners of the Gaussian's tile rectangle
// to a coverage difference grid, whose 2-D prefix sum (binning.cu: tile_scan_kernel) is the number
// of instances per tile - the per-Gaussian prefix sum of the reference (rasterizer_impl.cu:277)
// is not needed. cov3D is not stored: the backward recomputes it.
#include <cstdio>
#include "bgr_common.cuh"
#include "gs_math.cuh"

namespace bgr {

namespace {

template <typename LoadCoeff>
__device__ __forceinline__ V3 sh_to_rgb(int deg, V3 dir, LoadCoeff sh)
{
    V3 result = kC0 * sh(0);
    if (deg > 0) {
        const float x = dir.x, y = dir.y, z = dir.z;
        result = result - kC1 * y * sh(1) + kC1 * z * sh(2) - kC1 * x * sh(3);
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z;
            const float xy = x * y, yz = y * z, xz = x * z;
            result = result + kC2_0 * xy * sh(4) + kC2_1 * yz * sh(5) +
                     kC2_2 * (2.0f * zz - xx - yy) * sh(6) + kC2_3 * xz * sh(7) +
                     kC2_4 * (xx - yy) * sh(8);
            if (deg > 2) {
                result = result + kC3_0 * y * (3.0f * xx - yy) * sh(9) + kC3_1 * xy * z * sh(10) +
                         kC3_2 * y * (4.0f * zz - xx - yy) * sh(11) +
                         kC3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh(12) +
                         kC3_4 * x * (4.0f * zz - xx - yy) * sh(13) + kC3_5 * z * (xx - yy) * sh(14) +
                         kC3_6 * x * (xx - 3.0f * yy) * sh(15);
            }
        }
    }
    result.x += 0.5f;
    result.y += 0.5f;
    result.z += 0.5f;
    return result;
}

constexpr int PRE_THREADS = 256;

// kVecSH: shs has M == 16 coefficients and a 16-byte aligned base, so a Gaussian's 48 floats are
// fetched as twelve 128-bit loads.
// kRaw (bgr_forward_raw): the arrays are the trainer's raw leaves; exp / normalize / sigmoid
// (scene/gaussian_model.py:95-115, the formulas of torch's CUDA kernels) are applied as the values
// are loaded, and coefficient 0 comes from _features_dc, the others from _features_rest.
template <bool kVecSH, bool kRaw>
__global__ void __launch_bounds__(PRE_THREADS, 4) preprocess_kernel(PreprocessParams p)
{
    __shared__ float s_view[16], s_proj[16], s_cam[3];
    if (threadIdx.x < 16) {
        s_view[threadIdx.x] = p.viewmatrix[threadIdx.x];
        s_proj[threadIdx.x] = p.projmatrix[threadIdx.x];
    }
    if (threadIdx.x < 3) s_cam[threadIdx.x] = p.cam_pos[threadIdx.x];
    __syncthreads();

    const int idx = blockIdx.x * PRE_THREADS + threadIdx.x;
    uint32_t tiles = 0;
    int radius_out = 0;

    if (idx < p.P) {
        const float3 p_orig = {p.means3D[3 * idx], p.means3D[3 * idx + 1], p.means3D[3 * idx + 2]};
        // scale and rotation are fetched before the cull decision: 28 coalesced bytes that a
        // culled Gaussian wastes, one dependent DRAM latency that a visible one saves
        float3 sc = {0.f, 0.f, 0.f};
        float4 q = {0.f, 0.f, 0.f, 0.f};
        if (p.cov3D_precomp == nullptr) {
            sc = {p.scales[3 * idx], p.scales[3 * idx + 1], p.scales[3 * idx + 2]};
            if ((reinterpret_cast<uintptr_t>(p.rotations) & 15) == 0) {
                q = __ldg(reinterpret_cast<const float4*>(p.rotations) + idx);
            } else {
                q = make_float4(p.rotations[4 * (size_t)idx], p.rotations[4 * (size_t)idx + 1],
                                p.rotations[4 * (size_t)idx + 2], p.rotations[4 * (size_t)idx + 3]);
            }
            if (kRaw) {
                sc = {expf(sc.x), expf(sc.y), expf(sc.z)};
                const float inv = 1.0f / fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
                q = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
            }
        }
        const float3 p_view = xform_4x3(p_orig, s_view);
        bool alive = p_view.z > 0.2f;
        if (!alive && p.prefiltered) {
            printf("Point is filtered although prefiltered is set. This shouldn't happen!");
            __trap();
        }
        if (alive) {
            const float4 p_hom = xform_4x4(p_orig, s_proj);
            const float p_w = 1.0f / (p_hom.w + 0.0000001f);
            const float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};

            float cov3D[6];
            if (p.cov3D_precomp != nullptr) {
#pragma unroll
                for (int k = 0; k < 6; ++k) cov3D[k] = p.cov3D_precomp[6 * (size_t)idx + k];
            } else {
                cov3d_from_scale_rot(sc, p.scale_modifier, q, cov3D);
            }

            const ProjFrame f =
                make_proj_frame(p_orig, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy, s_view);
            M3 cov = cov2d_full(f.T, cov3D);
            cov.m[0][0] += 0.3f;
            cov.m[1][1] += 0.3f;
            const float cx = cov.m[0][0], cy = cov.m[0][1], cz = cov.m[1][1];

            const float det = (cx * cz - cy * cy);
            if (det != 0.0f) {
                const float det_inv = 1.f / det;
                const float3 conic = {cz * det_inv, -cy * det_inv, cx * det_inv};

                const float mid = 0.5f * (cx + cz);
                const float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
                const float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
                const float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
                const float2 point_image = {ndc_to_pix(p_proj.x, p.W), ndc_to_pix(p_proj.y, p.H)};
                uint2 rmin, rmax;
                tile_rect(point_image, (int)my_radius, p.grid_x, p.grid_y, rmin, rmax);
                const uint32_t area = (rmax.x - rmin.x) * (rmax.y - rmin.y);
                if (area != 0) {
                    float3 rgb;
                    uint32_t clampbits = 0;
                    if (p.colors_precomp == nullptr) {
                        V3 dir = v3(p_orig.x, p_orig.y, p_orig.z) - v3(s_cam[0], s_cam[1], s_cam[2]);
                        dir = dir / length3(dir);
                        V3 c;
                        if (kVecSH) {
                            float sh[48];
                            const float4* src =
                                reinterpret_cast<const float4*>(p.shs) + (size_t)idx * 12;
                            const int nvec = (3 * (p.D + 1) * (p.D + 1) + 3) >> 2;
#pragma unroll
                            for (int k = 0; k < 12; ++k) {
                                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                                if (k < nvec) v = __ldg(src + k);
                                sh[4 * k + 0] = v.x; sh[4 * k + 1] = v.y;
                                sh[4 * k + 2] = v.z; sh[4 * k + 3] = v.w;
                            }
                            c = sh_to_rgb(p.D, dir, [&](int k) {
                                return v3(sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]);
                            });
                        } else if (kRaw) {
                            const float* dc = p.shs + (size_t)idx * 3;
                            const float* rest = p.shs_rest + (size_t)idx * (p.M - 1) * 3 - 3;
                            c = sh_to_rgb(p.D, dir, [&](int k) {
                                const float* src = k == 0 ? dc : rest + 3 * k;
                                return v3(src[0], src[1], src[2]);
                            });
                        } else {
                            const float* src = p.shs + (size_t)idx * p.M * 3;
                            c = sh_to_rgb(p.D, dir, [&](int k) {
                                return v3(src[3 * k], src[3 * k + 1], src[3 * k + 2]);
                            });
                        }
                        clampbits = (c.x < 0 ? 1u : 0u) | (c.y < 0 ? 2u : 0u) | (c.z < 0 ? 4u : 0u);
                        rgb = {fmaxf(c.x, 0.0f), fmaxf(c.y, 0.0f), fmaxf(c.z, 0.0f)};
                    } else {
                        rgb = {p.colors_precomp[3 * (size_t)idx], p.colors_precomp[3 * (size_t)idx + 1],
                               p.colors_precomp[3 * (size_t)idx + 2]};
                    }
                    float opacity = p.opacities[idx];
                    if (kRaw) opacity = 1.0f / (1.0f + expf(-opacity));
                    // power < thr implies opacity * exp(power) < 1/255 (margin covers logf/expf
                    // rounding); opacity <= 0 or NaN keeps the exact per-pixel path.
                    const float thr =
                        (opacity > 0.0f) ? -(logf(255.0f * opacity) + 1e-4f) : -3.0e38f;

                    GaussRec r;
                    r.q0 = make_float4(point_image.x, point_image.y, conic.x, conic.y);
                    r.q1 = make_float4(conic.z, opacity, rgb.x, rgb.y);
                    r.q2 = make_float4(rgb.z, thr, p_view.z, __uint_as_float(clampbits));
                    p.rec[idx] = r;

                    tiles = area;
                    radius_out = (int)my_radius;
                    // coverage of [rmin, rmax): +1 / -1 at the corners, summed by tile_scan_kernel
                    const int gw = p.grid_x + 1;
                    atomicAdd(p.coverage_grid + rmin.y * gw + rmin.x, 1);
                    atomicAdd(p.coverage_grid + rmin.y * gw + rmax.x, -1);
                    atomicAdd(p.coverage_grid + rmax.y * gw + rmin.x, -1);
                    atomicAdd(p.coverage_grid + rmax.y * gw + rmax.x, 1);
                }
            }
        }
        p.radii[idx] = radius_out;
        p.tiles_touched[idx] = tiles;
    }

    // Visible count and the 64-bit instance total (overflow guard): reduced over the block first,
    // so that the counters receive one pair of atomics per block.
    __shared__ uint32_t s_cnt[PRE_THREADS / 32];
    __shared__ uint32_t s_sum[PRE_THREADS / 32];   // <= 32 x tiles of one image: fits 32 bits
    const unsigned vis_mask = __ballot_sync(0xffffffffu, tiles != 0);
    const uint32_t wsum = __reduce_add_sync(0xffffffffu, tiles);
    if ((threadIdx.x & 31) == 0) {
        s_cnt[threadIdx.x >> 5] = (uint32_t)__popc(vis_mask);
        s_sum[threadIdx.x >> 5] = wsum;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t bcnt = 0u;
        unsigned long long bsum = 0ull;
#pragma unroll
        for (int w = 0; w < PRE_THREADS / 32; ++w) {
            bcnt += s_cnt[w];
            bsum += s_sum[w];
        }
        if (bcnt != 0) {
            atomicAdd(&p.control->visible, bcnt);
            atomicAdd(&p.control->tiles_total, bsum);
        }
    }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D,
                                    const float* __restrict__ view, uint8_t* __restrict__ present)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float3 p_orig = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    const float3 p_view = xform_4x3(p_orig, view);
    present[idx] = p_view.z > 0.2f ? 1 : 0;
}

}  // namespace

void launch_preprocess(const PreprocessParams& p, cudaStream_t s)
{
    if (p.P <= 0) return;
    const int blocks = (p.P + PRE_THREADS - 1) / PRE_THREADS;
    const bool vec = p.shs != nullptr && p.M == 16 && (reinterpret_cast<uintptr_t>(p.shs) & 15) == 0;
    ++g_kernel_launches;
    if (p.raw)
        preprocess_kernel<false, true><<<blocks, PRE_THREADS, 0, s>>>(p);
    else if (vec)
        preprocess_kernel<true, false><<<blocks, PRE_THREADS, 0, s>>>(p);
    else
        preprocess_kernel<false, false><<<blocks, PRE_THREADS, 0, s>>>(p);
}

void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present,
                         cudaStream_t s)
{
    if (P <= 0) return;
    ++g_kernel_launches;
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, viewmatrix, present);
}

}  // namespace bgr
