// The step AFTER the rasterizer: the trainer's loss and its optimizer update.
//
//   loss = (1 - lambda) * mean|x - y| + lambda * (1 - mean(ssim_map(x, y)))        train.py:91-92
//   ssim_map: utils/loss_utils.py:43-63 - five depthwise 11x11 Gaussian (sigma 1.5) convolutions
//             with zero padding (mu1, mu2, E[x^2], E[y^2], E[xy]) and the usual quotient
//   Adam:     torch.optim.Adam(l, lr=0.0, eps=1e-15) over the six parameter groups
//             (scene/gaussian_model.py:149-160), one learning rate per group
//
// In PyTorch the loss is ~25 elementwise / convolution kernels forward and as many backward over
// (3,H,W) tensors; here it is ONE kernel that produces the loss partial sums and the three
// derivative maps, a tiny deterministic finalisation, and ONE kernel for dL/dimage. The window is
// separable (the reference builds the 2-D window as the outer product of the 1-D one), so every
// convolution is an 11-tap row pass into shared memory followed by an 11-tap column pass.
//
//   S   = (2 m1 m2 + C1)(2 s12 + C2) / ((m1^2 + m2^2 + C1)(s1 + s2 + C2)),
//         s1 = e1 - m1^2, s2 = e2 - m2^2, s12 = e12 - m1 m2,  m = conv(.), e1 = conv(x^2), ...
//   dL/dx = gS * (conv(A) + 2 x conv(B) + y conv(C)) + gL * sign(x - y)
//         A = dS/dm1 (through m1, s1 and s12), B = dS/de1, C = dS/de12,
//         gS = -lambda / N, gL = (1 - lambda) / N, N = 3 H W
#include <algorithm>
#include <cmath>
#include "bgr_common.cuh"

namespace bgr {

namespace {

constexpr int LT = 32;             // output tile edge
constexpr int LR = 5;              // window radius (11 taps)
constexpr int LIN = LT + 2 * LR;   // input tile edge with halo
constexpr int LTHREADS = 256;

struct Window { float w[2 * LR + 1]; };

// Deterministic block sum (fixed tree): warp shuffles, then the warps' partials in order.
__device__ __forceinline__ float block_sum(float v, float* s_red)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) s_red[warp] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < LTHREADS / 32; ++w) t += s_red[w];
    return t;
}

// Loads the LIN x LIN input tile of one channel around output tile (tx0, ty0); zero outside the
// image (conv2d padding).
__device__ __forceinline__ void load_tile(const float* __restrict__ img, int W, int H, int tx0, int ty0,
                                          float (*dst)[LIN + 1])
{
    for (int i = threadIdx.x; i < LIN * LIN; i += LTHREADS) {
        const int r = i / LIN, c = i - r * LIN;
        const int gy = ty0 + r - LR, gx = tx0 + c - LR;
        dst[r][c] = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? __ldg(img + (size_t)gy * W + gx) : 0.f;
    }
}

// Forward: per output pixel the SSIM value, |x - y| and the three derivative maps A, B, C.
// grid (ceil(W/32), ceil(H/32), 3); partial[block] = (sum of ssim, sum of |x - y|).
__global__ void __launch_bounds__(LTHREADS)
ssim_forward_kernel(const float* __restrict__ x, const float* __restrict__ y, int W, int H, Window win,
                    float* __restrict__ maps, float2* __restrict__ partial)
{
    __shared__ float sx[LIN][LIN + 1], sy[LIN][LIN + 1];
    __shared__ float sh[5][LIN][LT + 1];
    __shared__ float s_red[LTHREADS / 32];
    const int tx0 = blockIdx.x * LT, ty0 = blockIdx.y * LT, ch = blockIdx.z;
    const size_t plane = (size_t)W * H;
    load_tile(x + ch * plane, W, H, tx0, ty0, sx);
    load_tile(y + ch * plane, W, H, tx0, ty0, sy);
    __syncthreads();
    // rows
    for (int i = threadIdx.x; i < LIN * LT; i += LTHREADS) {
        const int r = i / LT, c = i - r * LT;
        float h0 = 0.f, h1 = 0.f, h2 = 0.f, h3 = 0.f, h4 = 0.f;
#pragma unroll
        for (int k = 0; k <= 2 * LR; ++k) {
            const float a = sx[r][c + k], b = sy[r][c + k], w = win.w[k];
            h0 = fmaf(w, a, h0);
            h1 = fmaf(w, b, h1);
            h2 = fmaf(w, a * a, h2);
            h3 = fmaf(w, b * b, h3);
            h4 = fmaf(w, a * b, h4);
        }
        sh[0][r][c] = h0; sh[1][r][c] = h1; sh[2][r][c] = h2; sh[3][r][c] = h3; sh[4][r][c] = h4;
    }
    __syncthreads();
    // columns + the quotient
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    float sum_s = 0.f, sum_l = 0.f;
    for (int i = threadIdx.x; i < LT * LT; i += LTHREADS) {
        const int r = i / LT, c = i - r * LT;
        const int gx = tx0 + c, gy = ty0 + r;
        if (gx >= W || gy >= H) continue;
        float m1 = 0.f, m2 = 0.f, e1 = 0.f, e2 = 0.f, e12 = 0.f;
#pragma unroll
        for (int k = 0; k <= 2 * LR; ++k) {
            const float w = win.w[k];
            m1 = fmaf(w, sh[0][r + k][c], m1);
            m2 = fmaf(w, sh[1][r + k][c], m2);
            e1 = fmaf(w, sh[2][r + k][c], e1);
            e2 = fmaf(w, sh[3][r + k][c], e2);
            e12 = fmaf(w, sh[4][r + k][c], e12);
        }
        const float s1 = e1 - m1 * m1, s2 = e2 - m2 * m2, s12 = e12 - m1 * m2;
        const float num1 = 2.f * m1 * m2 + C1, num2 = 2.f * s12 + C2;
        const float den1 = m1 * m1 + m2 * m2 + C1, den2 = s1 + s2 + C2;
        const float inv = 1.f / (den1 * den2);
        const float S = num1 * num2 * inv;
        // dS/de12, dS/de1, and dS/dm1 through m1 itself, s1 = e1 - m1^2 and s12 = e12 - m1 m2
        const float Cc = 2.f * num1 * inv;
        const float Bb = -S / den2;
        const float Aa = 2.f * m2 * num2 * inv - 2.f * m1 * S / den1 - m2 * Cc - 2.f * m1 * Bb;
        const size_t o = ch * plane + (size_t)gy * W + gx;
        maps[o] = Aa;
        maps[3 * plane + o] = Bb;
        maps[6 * plane + o] = Cc;
        sum_s += S;
        sum_l += fabsf(sx[r + LR][c + LR] - sy[r + LR][c + LR]);
    }
    sum_s = block_sum(sum_s, s_red);
    sum_l = block_sum(sum_l, s_red);
    if (threadIdx.x == 0)
        partial[(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = make_float2(sum_s, sum_l);
}

// One block: the partial sums in a fixed order -> loss (and its two terms, for logging).
__global__ void __launch_bounds__(LTHREADS)
loss_finalize_kernel(const float2* __restrict__ partial, int n, float inv_count, float lambda,
                     float* __restrict__ out3)
{
    __shared__ float s_red[LTHREADS / 32];
    float a = 0.f, b = 0.f;
    for (int i = threadIdx.x; i < n; i += LTHREADS) {
        const float2 v = partial[i];
        a += v.x;
        b += v.y;
    }
    a = block_sum(a, s_red);
    b = block_sum(b, s_red);
    if (threadIdx.x == 0) {
        const float ssim = a * inv_count, l1 = b * inv_count;
        out3[0] = (1.f - lambda) * l1 + lambda * (1.f - ssim);
        out3[1] = l1;
        out3[2] = ssim;
    }
}

// Backward: dL/dx = gS * (conv(A) + 2 x conv(B) + y conv(C)) + gL * sign(x - y). The maps are zero
// outside the image (the transposed convolution sums over output positions inside the image only;
// the window is symmetric, so it is the same 11-tap pass).
__global__ void __launch_bounds__(LTHREADS)
ssim_backward_kernel(const float* __restrict__ x, const float* __restrict__ y, int W, int H, Window win,
                     const float* __restrict__ maps, float gS, float gL, float* __restrict__ dL_dx)
{
    __shared__ float sm[3][LIN][LIN + 1];
    __shared__ float sh[3][LIN][LT + 1];
    const int tx0 = blockIdx.x * LT, ty0 = blockIdx.y * LT, ch = blockIdx.z;
    const size_t plane = (size_t)W * H;
#pragma unroll
    for (int q = 0; q < 3; ++q) load_tile(maps + (3 * q + ch) * plane, W, H, tx0, ty0, sm[q]);
    __syncthreads();
    for (int i = threadIdx.x; i < LIN * LT; i += LTHREADS) {
        const int r = i / LT, c = i - r * LT;
        float h0 = 0.f, h1 = 0.f, h2 = 0.f;
#pragma unroll
        for (int k = 0; k <= 2 * LR; ++k) {
            const float w = win.w[k];
            h0 = fmaf(w, sm[0][r][c + k], h0);
            h1 = fmaf(w, sm[1][r][c + k], h1);
            h2 = fmaf(w, sm[2][r][c + k], h2);
        }
        sh[0][r][c] = h0; sh[1][r][c] = h1; sh[2][r][c] = h2;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < LT * LT; i += LTHREADS) {
        const int r = i / LT, c = i - r * LT;
        const int gx = tx0 + c, gy = ty0 + r;
        if (gx >= W || gy >= H) continue;
        float cA = 0.f, cB = 0.f, cC = 0.f;
#pragma unroll
        for (int k = 0; k <= 2 * LR; ++k) {
            const float w = win.w[k];
            cA = fmaf(w, sh[0][r + k][c], cA);
            cB = fmaf(w, sh[1][r + k][c], cB);
            cC = fmaf(w, sh[2][r + k][c], cC);
        }
        const size_t o = ch * plane + (size_t)gy * W + gx;
        const float xv = __ldg(x + o), yv = __ldg(y + o);
        const float d = xv - yv;
        const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
        dL_dx[o] = gS * (cA + 2.f * xv * cB + yv * cC) + gL * sgn;
    }
}

// Adam over up to eight parameter groups in one launch (torch.optim.Adam's update, one learning
// rate per group): m += (1-b1)(g-m); v = b2 v + (1-b2) g^2; p -= step_size * m / (sqrt(v)/sqrt(bc2) + eps)
// HBM-bound: 16 B read + 12 B written per element. A work item is four consecutive elements
// (128-bit accesses) while the group's arrays are 16-byte aligned, then up to three single ones.
struct AdamLaunch {
    float* p[8]; const float* g[8]; float* m[8]; float* v[8];
    unsigned long long end[8];     // running end of the groups' work items
    unsigned long long vec[8];     // 4-element items of the group (the rest are single elements)
    float step_size[8];            // lr / (1 - beta1^t)
    int count;
    // all derived on the host in double from the caller's doubles, as torch does with its Python
    // floats (1 - 0.999f in float32 is off by 1.3e-5 relative)
    float beta2, one_m_beta1, one_m_beta2, eps, bc2_sqrt;
};

__device__ __forceinline__ void adam_one(const AdamLaunch& a, float ss, float g, float& m, float& v, float& p)
{
    m = m + a.one_m_beta1 * (g - m);
    v = a.beta2 * v + a.one_m_beta2 * g * g;
    const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
    p = p - ss * (m / denom);
}

__global__ void __launch_bounds__(256) adam_kernel(AdamLaunch a)
{
    const unsigned long long total = a.end[a.count - 1];
    for (unsigned long long i = (unsigned long long)blockIdx.x * 256 + threadIdx.x; i < total;
         i += (unsigned long long)gridDim.x * 256) {
        int k = 0;
        unsigned long long base = 0;
#pragma unroll
        for (int q = 0; q < 7; ++q)
            if (q < a.count - 1 && i >= a.end[q]) { k = q + 1; base = a.end[q]; }
        const unsigned long long j = i - base;
        const float ss = a.step_size[k];
        if (j < a.vec[k]) {
            const float4 g = __ldcs(reinterpret_cast<const float4*>(a.g[k]) + j);
            float4 m = reinterpret_cast<float4*>(a.m[k])[j], v = reinterpret_cast<float4*>(a.v[k])[j];
            float4 p = reinterpret_cast<float4*>(a.p[k])[j];
            adam_one(a, ss, g.x, m.x, v.x, p.x);
            adam_one(a, ss, g.y, m.y, v.y, p.y);
            adam_one(a, ss, g.z, m.z, v.z, p.z);
            adam_one(a, ss, g.w, m.w, v.w, p.w);
            reinterpret_cast<float4*>(a.m[k])[j] = m;
            reinterpret_cast<float4*>(a.v[k])[j] = v;
            reinterpret_cast<float4*>(a.p[k])[j] = p;
        } else {
            const unsigned long long e = 4ull * a.vec[k] + (j - a.vec[k]);
            float m = a.m[k][e], v = a.v[k][e], p = a.p[k][e];
            adam_one(a, ss, a.g[k][e], m, v, p);
            a.m[k][e] = m;
            a.v[k][e] = v;
            a.p[k][e] = p;
        }
    }
}

// Densification statistics of one view (train.py:114-116, scene/gaussian_model.py:405-407), one
// pass instead of a boolean-mask gather / scatter chain: for Gaussians with radii > 0
//   max_radii2D = max(max_radii2D, radii); xyz_gradient_accum += |dL/dmean2D.xy|; denom += 1
__global__ void __launch_bounds__(256)
densification_stats_kernel(int P, const int* __restrict__ radii, const float* __restrict__ dL_dmean2D,
                           float* __restrict__ max_radii2D, float* __restrict__ grad_accum,
                           float* __restrict__ denom)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (r <= 0) return;
    const float gx = dL_dmean2D[3 * (size_t)i], gy = dL_dmean2D[3 * (size_t)i + 1];
    max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
    grad_accum[i] += sqrtf(gx * gx + gy * gy);
    denom[i] += 1.0f;
}

Window make_window()
{
    // utils/loss_utils.py:23-25: exp(-(x - 5)^2 / (2 * 1.5^2)) in double, stored as float32,
    // normalised by the float32 sum
    Window w;
    float g[2 * LR + 1], sum = 0.f;
    for (int k = 0; k <= 2 * LR; ++k) {
        g[k] = (float)exp(-(double)((k - LR) * (k - LR)) / (2.0 * 1.5 * 1.5));
        sum += g[k];
    }
    for (int k = 0; k <= 2 * LR; ++k) w.w[k] = g[k] / sum;
    return w;
}

}  // namespace

static size_t maps_bytes(int W, int H)   // the three derivative maps, padded to 256 bytes
{
    return ((size_t)9 * W * H * sizeof(float) + 255) & ~(size_t)255;
}

size_t loss_scratch_bytes(int W, int H)
{
    const size_t blocks = (size_t)((W + LT - 1) / LT) * ((H + LT - 1) / LT) * 3;
    return maps_bytes(W, H) + ((blocks * sizeof(float2) + 255) & ~(size_t)255);
}

// scratch: [maps 9 W H floats | partial sums]. loss3 (device): loss, l1, ssim.
int launch_l1_ssim_loss(const float* image, const float* gt, int W, int H, float lambda, void* scratch,
                        float* loss3, float* dL_dimage, cudaStream_t s)
{
    if (image == nullptr || gt == nullptr || scratch == nullptr || loss3 == nullptr || W <= 0 || H <= 0) return -1;
    const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, 3);
    const int blocks = (int)(grid.x * grid.y * grid.z);
    float* maps = static_cast<float*>(scratch);
    float2* partial = reinterpret_cast<float2*>(static_cast<unsigned char*>(scratch) + maps_bytes(W, H));
    static const Window win = make_window();
    const float inv_count = 1.0f / (3.0f * (float)W * (float)H);
    ++g_kernel_launches;
    ssim_forward_kernel<<<grid, LTHREADS, 0, s>>>(image, gt, W, H, win, maps, partial);
    ++g_kernel_launches;
    loss_finalize_kernel<<<1, LTHREADS, 0, s>>>(partial, blocks, inv_count, lambda, loss3);
    if (dL_dimage != nullptr) {
        ++g_kernel_launches;
        ssim_backward_kernel<<<grid, LTHREADS, 0, s>>>(image, gt, W, H, win, maps, -lambda * inv_count,
                                                      (1.0f - lambda) * inv_count, dL_dimage);
    }
    return 0;
}

int launch_densification_stats(int P, const int* radii, const float* dL_dmean2D, float* max_radii2D,
                               float* grad_accum, float* denom, cudaStream_t s)
{
    if (P < 0 || (P > 0 && (radii == nullptr || dL_dmean2D == nullptr || max_radii2D == nullptr ||
                            grad_accum == nullptr || denom == nullptr)))
        return -1;
    if (P == 0) return 0;
    ++g_kernel_launches;
    densification_stats_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, radii, dL_dmean2D, max_radii2D, grad_accum, denom);
    return 0;
}

int launch_adam(int count, float* const* params, const float* const* grads, float* const* exp_avg,
                float* const* exp_avg_sq, const size_t* numel, const double* lr, double beta1, double beta2,
                double eps, int step, cudaStream_t s)
{
    if (count <= 0 || count > 8 || step <= 0) return -1;
    AdamLaunch a;
    unsigned long long end = 0;
    const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
    for (int k = 0; k < 8; ++k) {
        const bool on = k < count;
        if (on && (params[k] == nullptr || grads[k] == nullptr || exp_avg[k] == nullptr || exp_avg_sq[k] == nullptr))
            return -1;
        a.p[k] = on ? params[k] : nullptr; a.g[k] = on ? grads[k] : nullptr;
        a.m[k] = on ? exp_avg[k] : nullptr; a.v[k] = on ? exp_avg_sq[k] : nullptr;
        unsigned long long vec = 0;
        if (on) {
            const uintptr_t bits = reinterpret_cast<uintptr_t>(params[k]) | reinterpret_cast<uintptr_t>(grads[k]) |
                                   reinterpret_cast<uintptr_t>(exp_avg[k]) | reinterpret_cast<uintptr_t>(exp_avg_sq[k]);
            vec = (bits & 15) == 0 ? numel[k] / 4 : 0;
            end += vec + (numel[k] - 4 * vec);
        }
        a.vec[k] = vec;
        a.end[k] = end;
        a.step_size[k] = on ? (float)(lr[k] / bc1) : 0.f;
    }
    if (end == 0) return 0;
    a.count = count; a.beta2 = (float)beta2; a.one_m_beta1 = (float)(1.0 - beta1);
    a.one_m_beta2 = (float)(1.0 - beta2); a.eps = (float)eps; a.bc2_sqrt = (float)sqrt(bc2);
    const unsigned blocks = (unsigned)std::min<unsigned long long>((end + 255) / 256, 148ull * 32);
    ++g_kernel_launches;
    adam_kernel<<<blocks, 256, 0, s>>>(a);
    return 0;
}

}  // namespace bgr
