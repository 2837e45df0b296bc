// gsr_loss.cu -- fused L1 + SSIM training loss, forward and backward, for sm_100a.
//
// Replaces, on the other side of the rasterizer, what the reference composes from torch ops every iteration
// (utils/loss_utils.py:17-64, train.py:126-127): five grouped 11x11 convolutions (mu1, mu2, E[x^2], E[y^2], E[xy]),
// a dozen full-image elementwise maps, two means, and autograd's backward through all of it.  Here:
//   forward  = ONE kernel: a 16x16 pixel tile per block and channel, both images' 26x26 halo tiles in shared
//              memory, the Gaussian window applied separably (11 horizontal taps into shared memory, 11 vertical
//              taps per pixel) for the five moments at once, the SSIM value and its three partial derivatives
//              (w.r.t. conv(x), conv(x^2), conv(xy)) per pixel, per-block partial sums of SSIM and |x - y|;
//              a one-block kernel adds the partials in a fixed order (deterministic loss).
//   backward = ONE kernel: d loss / d x = c_l1 sign(x - y) + c_ssim (conv(dm_dmu1) + 2 x conv(dm_dq) + y conv(dm_dr)),
//              the same separable scheme on the three derivative maps (the window is symmetric).
// Zero padding as F.conv2d(padding = 5): pixels outside the image contribute zeros in both directions.
#include "gsr_internal.cuh"
#include <cmath>

namespace gsr {

constexpr int LT = 16;              // tile edge
constexpr int LR = 5;               // window radius (window_size 11)
constexpr int LH = LT + 2 * LR;     // halo tile edge

struct SsimWindow { float w[2 * LR + 1]; };

// gaussian(11, 1.5) of utils/loss_utils.py:23-25, evaluated as there: float32 values of exp(), float32 sum and division
static SsimWindow make_window() {
    SsimWindow win;
    float sum = 0.f;
    for (int i = 0; i < 2 * LR + 1; i++) {
        win.w[i] = (float)std::exp(-(double)((i - LR) * (i - LR)) / (2.0 * 1.5 * 1.5));
        sum += win.w[i];
    }
    for (int i = 0; i < 2 * LR + 1; i++) win.w[i] /= sum;
    return win;
}

__device__ __forceinline__ void load_halo(float (*s)[LH + 1], const float* __restrict__ img, int H, int W, int x0, int y0) {
    for (int i = threadIdx.x; i < LH * LH; i += 256) {
        const int r = i / LH, c = i - r * LH;
        const int gy = y0 + r - LR, gx = x0 + c - LR;
        s[r][c] = (gy >= 0 && gy < H && gx >= 0 && gx < W) ? __ldg(img + (size_t)gy * W + gx) : 0.f;
    }
}

__global__ void __launch_bounds__(256) l1_ssim_fwd_kernel(const float* __restrict__ img1, const float* __restrict__ img2, int H, int W,
                                                          SsimWindow win, float C1, float C2, float* __restrict__ dm_dmu1,
                                                          float* __restrict__ dm_dq, float* __restrict__ dm_dr,
                                                          float* __restrict__ partial, int nblocks) {
    __shared__ float s_x[LH][LH + 1], s_y[LH][LH + 1];
    __shared__ float s_h[5][LH][LT];
    __shared__ float s_red[2][8];
    const int ch = blockIdx.z, x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
    const size_t plane = (size_t)ch * H * W;
    load_halo(s_x, img1 + plane, H, W, x0, y0);
    load_halo(s_y, img2 + plane, H, W, x0, y0);
    __syncthreads();
    // horizontal taps: 26 rows x 16 columns, five moments each
    for (int i = threadIdx.x; i < LH * LT; i += 256) {
        const int r = i / LT, c = i - r * LT;
        float a = 0.f, b = 0.f, aa = 0.f, bb = 0.f, ab = 0.f;
#pragma unroll
        for (int k = 0; k < 2 * LR + 1; k++) {
            const float x = s_x[r][c + k], y = s_y[r][c + k], w = win.w[k];
            const float wx = w * x, wy = w * y;
            a += wx; b += wy; aa += wx * x; bb += wy * y; ab += wx * y;
        }
        s_h[0][r][c] = a; s_h[1][r][c] = b; s_h[2][r][c] = aa; s_h[3][r][c] = bb; s_h[4][r][c] = ab;
    }
    __syncthreads();
    const int tx = threadIdx.x & (LT - 1), ty = threadIdx.x / LT;
    float mu1 = 0.f, mu2 = 0.f, q = 0.f, q2 = 0.f, r12 = 0.f;
#pragma unroll
    for (int k = 0; k < 2 * LR + 1; k++) {
        const float w = win.w[k];
        mu1 += w * s_h[0][ty + k][tx]; mu2 += w * s_h[1][ty + k][tx];
        q += w * s_h[2][ty + k][tx]; q2 += w * s_h[3][ty + k][tx]; r12 += w * s_h[4][ty + k][tx];
    }
    const int px = x0 + tx, py = y0 + ty;
    const bool inside = px < W && py < H;
    float m = 0.f, l1 = 0.f;
    if (inside) {
        // utils/loss_utils.py:47-61
        const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu1_mu2 = mu1 * mu2;
        const float s1 = q - mu1_sq, s2 = q2 - mu2_sq, s12 = r12 - mu1_mu2;
        const float A1 = 2.f * mu1_mu2 + C1, A2 = 2.f * s12 + C2, B1 = mu1_sq + mu2_sq + C1, B2 = s1 + s2 + C2;
        const float inv = 1.f / (B1 * B2);
        m = A1 * A2 * inv;
        if (dm_dmu1 != nullptr) {
            // m as a function of p = conv(x) (= mu1), q = conv(x^2), r = conv(x y):  s1 = q - p^2, s12 = r - p mu2
            const float dm_ds1 = -m / B2;                      // = -A1 A2 / (B1 B2^2)
            const float dm_ds12 = 2.f * A1 * inv;
            const float dm_dp = 2.f * mu2 * A2 * inv - 2.f * mu1 * m / B1 + dm_ds1 * (-2.f * mu1) + dm_ds12 * (-mu2);
            const size_t o = plane + (size_t)py * W + px;
            dm_dmu1[o] = dm_dp; dm_dq[o] = dm_ds1; dm_dr[o] = dm_ds12;
        }
        l1 = fabsf(s_x[ty + LR][tx + LR] - s_y[ty + LR][tx + LR]);
    }
    // per-block partial sums, fixed order
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { m += __shfl_xor_sync(0xffffffffu, m, d); l1 += __shfl_xor_sync(0xffffffffu, l1, d); }
    if ((threadIdx.x & 31) == 0) { s_red[0][threadIdx.x >> 5] = m; s_red[1][threadIdx.x >> 5] = l1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float sm = 0.f, sl = 0.f;
#pragma unroll
        for (int w = 0; w < 8; w++) { sm += s_red[0][w]; sl += s_red[1][w]; }
        const int b = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        partial[b] = sm; partial[nblocks + b] = sl;
    }
}

// out[0] = (1 - lambda) * mean|x - y| + lambda * (1 - mean ssim)  (train.py:127), out[1] = mean|x - y|, out[2] = mean ssim
__global__ void __launch_bounds__(1024) l1_ssim_finalize_kernel(const float* __restrict__ partial, int nblocks, double inv_n, float lambda,
                                                                float* __restrict__ out) {
    __shared__ double s[2][32];
    double sm = 0.0, sl = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 1024) { sm += (double)partial[i]; sl += (double)partial[nblocks + i]; }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { sm += __shfl_xor_sync(0xffffffffu, sm, d); sl += __shfl_xor_sync(0xffffffffu, sl, d); }
    if ((threadIdx.x & 31) == 0) { s[0][threadIdx.x >> 5] = sm; s[1][threadIdx.x >> 5] = sl; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < 32; w++) { a += s[0][w]; b += s[1][w]; }
        const double ssim = a * inv_n, l1 = b * inv_n;
        out[0] = (float)((1.0 - (double)lambda) * l1 + (double)lambda * (1.0 - ssim));
        out[1] = (float)l1;
        out[2] = (float)ssim;
    }
}

__global__ void __launch_bounds__(256) l1_ssim_bwd_kernel(const float* __restrict__ img1, const float* __restrict__ img2,
                                                          const float* __restrict__ dm_dmu1, const float* __restrict__ dm_dq,
                                                          const float* __restrict__ dm_dr, int H, int W, SsimWindow win, float coef_l1,
                                                          float coef_ssim, const float* __restrict__ grad_out, float* __restrict__ dimg) {
    __shared__ float s_a[LH][LH + 1], s_b[LH][LH + 1], s_c[LH][LH + 1];
    __shared__ float s_h[3][LH][LT];
    const int ch = blockIdx.z, x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
    const size_t plane = (size_t)ch * H * W;
    load_halo(s_a, dm_dmu1 + plane, H, W, x0, y0);
    load_halo(s_b, dm_dq + plane, H, W, x0, y0);
    load_halo(s_c, dm_dr + plane, H, W, x0, y0);
    __syncthreads();
    for (int i = threadIdx.x; i < LH * LT; i += 256) {
        const int r = i / LT, c = i - r * LT;
        float a = 0.f, b = 0.f, cc = 0.f;
#pragma unroll
        for (int k = 0; k < 2 * LR + 1; k++) {
            const float w = win.w[k];
            a += w * s_a[r][c + k]; b += w * s_b[r][c + k]; cc += w * s_c[r][c + k];
        }
        s_h[0][r][c] = a; s_h[1][r][c] = b; s_h[2][r][c] = cc;
    }
    __syncthreads();
    const int tx = threadIdx.x & (LT - 1), ty = threadIdx.x / LT;
    const int px = x0 + tx, py = y0 + ty;
    if (px >= W || py >= H) return;
    float A = 0.f, B = 0.f, Cc = 0.f;
#pragma unroll
    for (int k = 0; k < 2 * LR + 1; k++) {
        const float w = win.w[k];
        A += w * s_h[0][ty + k][tx]; B += w * s_h[1][ty + k][tx]; Cc += w * s_h[2][ty + k][tx];
    }
    const size_t o = plane + (size_t)py * W + px;
    const float x = __ldg(img1 + o), y = __ldg(img2 + o);
    const float d = x - y;
    const float sgn = (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f);      // torch.abs backward: 0 at 0
    const float go = grad_out ? __ldg(grad_out) : 1.f;
    dimg[o] = go * (coef_l1 * sgn + coef_ssim * (A + 2.f * x * B + y * Cc));
}

size_t l1_ssim_blocks(int planes, int H, int W) {
    return (size_t)planes * ((H + LT - 1) / LT) * ((W + LT - 1) / LT);
}

void launch_l1_ssim_forward(const float* img, const float* gt, int planes, int H, int W, float lambda, float* dm_maps, float* partial,
                            float* out3, cudaStream_t st) {
    const SsimWindow win = make_window();
    const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, planes);
    const int nb = (int)l1_ssim_blocks(planes, H, W);
    const size_t n = (size_t)planes * H * W;
    float* m0 = dm_maps; float* m1 = dm_maps ? dm_maps + n : nullptr; float* m2 = dm_maps ? dm_maps + 2 * n : nullptr;
    l1_ssim_fwd_kernel<<<grid, 256, 0, st>>>(img, gt, H, W, win, 0.01f * 0.01f, 0.03f * 0.03f, m0, m1, m2, partial, nb);
    l1_ssim_finalize_kernel<<<1, 1024, 0, st>>>(partial, nb, 1.0 / (double)n, lambda, out3);
    count_launch(2);
}

void launch_l1_ssim_backward(const float* img, const float* gt, const float* dm_maps, int planes, int H, int W, float coef_l1,
                             float coef_ssim, const float* grad_out, float* dimg, cudaStream_t st) {
    const SsimWindow win = make_window();
    const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, planes);
    const size_t n = (size_t)planes * H * W;
    l1_ssim_bwd_kernel<<<grid, 256, 0, st>>>(img, gt, dm_maps, dm_maps + n, dm_maps + 2 * n, H, W, win, coef_l1, coef_ssim, grad_out, dimg);
    count_launch(1);
}

}  // namespace gsr

// ---- gsr_l1_ssim_forward_noalloc: the forward above with its per-block partial sums in a file-scope device array ----
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart
// in the reference (include/gsr.h).
//
// The array lives in every GPU context that loads this library until the process exits (1 MiB per GPU).  In HOST code
// its name is only a stand-in for the symbol, not a device address: the address of this device's copy is fetched once per
// device, with that device current, by cudaGetSymbolAddress and that pointer is what the kernels get.  The forward kernel
// overwrites partial[0 .. 2*nblocks) completely and the finalize kernel reads exactly that range, so nothing has to be
// cleared between calls; the array keeps the last call's partial sums.  Every call on a device shares it: two calls that
// overlap on the GPU (different streams, no ordering between them) race on it.
#include "../../include/gsr.h"
#include <mutex>
#include <string>

namespace gsr {
constexpr size_t L1_SSIM_NOALLOC_MAX_BLOCKS = GSR_L1_SSIM_NOALLOC_MAX_BLOCKS;
static __device__ __align__(256) float g_l1_ssim_static_partial[2 * L1_SSIM_NOALLOC_MAX_BLOCKS];

static std::mutex g_static_partial_mutex;
static float* g_static_partial_ptr[GSR_MAX_DEVICES];      // device address of the array, per device ordinal (nullptr = not fetched)
}  // namespace gsr

extern "C" size_t gsr_l1_ssim_noalloc_max_blocks(void) { return gsr::L1_SSIM_NOALLOC_MAX_BLOCKS; }

extern "C" int gsr_l1_ssim_forward_noalloc(const float* img, const float* gt, int planes, int height, int width, float lambda_dssim,
                                           float* dm_maps, float* out3, void* cuda_stream) {
    using namespace gsr;
    if (planes <= 0 || height <= 0 || width <= 0) return set_error(GSR_ERR_INVALID_ARGUMENT, "bad image sizes");
    if (!img || !gt || !out3) return set_error(GSR_ERR_INVALID_ARGUMENT, "null pointer");
    if (planes > 65535) return set_error(GSR_ERR_INVALID_ARGUMENT, "image too large");
    const size_t nb = l1_ssim_blocks(planes, height, width);
    if (nb > L1_SSIM_NOALLOC_MAX_BLOCKS)
        return set_error(GSR_ERR_INVALID_ARGUMENT,
                         "gsr_l1_ssim_forward_noalloc: planes * ceil(H/16) * ceil(W/16) = " + std::to_string(nb) + " tiles exceed the static scratch of " +
                             std::to_string(L1_SSIM_NOALLOC_MAX_BLOCKS) + " tiles; use gsr_l1_ssim_forward with caller-provided scratch");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return set_error(GSR_ERR_CUDA, std::string("cudaGetDevice: ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= GSR_MAX_DEVICES) return set_error(GSR_ERR_INVALID_ARGUMENT, "device ordinal out of range");
    float* partial = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_static_partial_mutex);
        if (g_static_partial_ptr[dev] == nullptr) {      // once per device, with that device current
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_l1_ssim_static_partial);
            if (e != cudaSuccess || p == nullptr)
                return set_error(GSR_ERR_CUDA, std::string("cudaGetSymbolAddress(g_l1_ssim_static_partial): ") + cudaGetErrorString(e));
            g_static_partial_ptr[dev] = (float*)p;
        }
        partial = g_static_partial_ptr[dev];
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    launch_l1_ssim_forward(img, gt, planes, height, width, lambda_dssim, dm_maps, partial, out3, st);
    e = cudaGetLastError();
    if (e != cudaSuccess) return set_error(GSR_ERR_CUDA, std::string("stage l1+ssim forward (noalloc): ") + cudaGetErrorString(e));
    return GSR_OK;
}
