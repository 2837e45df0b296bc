// gsr_adam.cu -- one-launch Adam step over all parameter groups of the Gaussian model.
//
// Replaces the optimizer step of the reference's training loop: `torch.optim.Adam(l, lr=0.0, eps=1e-15)` over the six
// groups xyz / f_dc / f_rest / opacity / scaling / rotation (scene/gaussian_model.py:243-252), stepped once per
// iteration (train.py:173-175).  torch runs that as its multi-tensor ("foreach") path: lerp, mul, addcmul, sqrt, div,
// add, addcdiv -- seven passes over 59 floats per Gaussian and their two moments.  Here the whole update is ONE pass:
// every element is read once (param, grad, exp_avg, exp_avg_sq) and written once (param and both moments), 28 bytes
// per element, all groups in one grid.  The arithmetic is torch's, in torch's order (torch/optim/adam.py,
// _multi_tensor_adam, the non-capturable branch):
//
//     m  <- m + (1 - beta1) * (g - m)                      exp_avg.lerp_(grad, 1 - beta1)
//     v  <- v * beta2 + (1 - beta2) * g * g                exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value = 1 - beta2)
//     d  <- sqrt(v) / sqrt(1 - beta2^t) + eps              (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)
//     p  <- p + (-lr / (1 - beta1^t)) * (m / d)            param.addcdiv_(exp_avg, denom, value = -step_size)
//
// HBM-bound: 28 B per element, nothing else.  Blocks of 256 threads take 2048 consecutive elements of one group as
// 16-byte vectors, all loads of a thread issued before the first use.  An optional per-Gaussian `visible` byte mask
// skips rows entirely (parameters AND moments untouched) -- not the reference's semantics (torch decays the moments of
// Gaussians without gradient too), off unless asked for.
#include "gsr_internal.cuh"

namespace gsr {

constexpr int ADAM_THREADS = 256;
constexpr int ADAM_ITERS = 2;                                   // 16-byte vectors per thread and array
constexpr int ADAM_TILE = ADAM_THREADS * ADAM_ITERS * 4;        // 2048 elements per block

struct AdamConsts { float beta2, w1, w2, eps; };                // w1 = 1 - beta1, w2 = 1 - beta2 (rounded from double, as torch does)

// Every rounding is pinned (no compiler-chosen contraction): the FMAs sit where torch's own CUDA functors have them
// (lerp: fma(w, end - self, self); addcmul: fma(value * t1, t2, self); addcdiv: fma(value, t1 / t2, self)); sqrt and
// the divisions are IEEE round-to-nearest as in torch's build (no fast-math).
__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, const AdamConsts& c, float bc2_sqrt, float neg_step) {
    m = __fmaf_rn(c.w1, __fsub_rn(g, m), m);
    v = __fmul_rn(v, c.beta2);
    v = __fmaf_rn(__fmul_rn(c.w2, g), g, v);
    const float d = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), bc2_sqrt), c.eps);
    p = __fmaf_rn(neg_step, __fdiv_rn(m, d), p);
}

template <bool VEC, bool MASKED>
__global__ void __launch_bounds__(ADAM_THREADS, 4) adam_kernel(const __grid_constant__ AdamLaunch L) {
    // which group does this block belong to (at most ADAM_MAX_GROUPS entries, block_start is ascending)
    int gi = 0;
#pragma unroll
    for (int k = 1; k < ADAM_MAX_GROUPS; k++)
        if (k < L.n_groups && blockIdx.x >= L.block_start[k]) gi = k;
    const AdamGroupDev G = L.g[gi];
    const AdamConsts c = {L.beta2, L.w1, L.w2, L.eps};
    const size_t base = (size_t)(blockIdx.x - L.block_start[gi]) * ADAM_TILE;

    if (VEC) {
        float4 p[ADAM_ITERS], g[ADAM_ITERS], m[ADAM_ITERS], v[ADAM_ITERS];
        size_t e[ADAM_ITERS];
        bool full[ADAM_ITERS];
#pragma unroll
        for (int k = 0; k < ADAM_ITERS; k++) {
            e[k] = base + ((size_t)k * ADAM_THREADS + threadIdx.x) * 4;
            full[k] = e[k] + 4 <= G.n;
            if (full[k]) {
                p[k] = *reinterpret_cast<const float4*>(G.p + e[k]);
                g[k] = __ldg(reinterpret_cast<const float4*>(G.g + e[k]));
                m[k] = *reinterpret_cast<const float4*>(G.m + e[k]);
                v[k] = *reinterpret_cast<const float4*>(G.v + e[k]);
            }
        }
#pragma unroll
        for (int k = 0; k < ADAM_ITERS; k++) {
            if (full[k]) {
                bool on[4] = {true, true, true, true};
                if (MASKED) {
#pragma unroll
                    for (int j = 0; j < 4; j++) on[j] = L.visible[(e[k] + j) / G.row_width] != 0;
                    if (!(on[0] | on[1] | on[2] | on[3])) continue;
                }
                float4 pn = p[k], mn = m[k], vn = v[k];
                adam_update(pn.x, g[k].x, mn.x, vn.x, c, G.bc2_sqrt, G.neg_step);
                adam_update(pn.y, g[k].y, mn.y, vn.y, c, G.bc2_sqrt, G.neg_step);
                adam_update(pn.z, g[k].z, mn.z, vn.z, c, G.bc2_sqrt, G.neg_step);
                adam_update(pn.w, g[k].w, mn.w, vn.w, c, G.bc2_sqrt, G.neg_step);
                if (MASKED) {
                    if (!on[0]) { pn.x = p[k].x; mn.x = m[k].x; vn.x = v[k].x; }
                    if (!on[1]) { pn.y = p[k].y; mn.y = m[k].y; vn.y = v[k].y; }
                    if (!on[2]) { pn.z = p[k].z; mn.z = m[k].z; vn.z = v[k].z; }
                    if (!on[3]) { pn.w = p[k].w; mn.w = m[k].w; vn.w = v[k].w; }
                }
                *reinterpret_cast<float4*>(G.p + e[k]) = pn;
                *reinterpret_cast<float4*>(G.m + e[k]) = mn;
                *reinterpret_cast<float4*>(G.v + e[k]) = vn;
            } else {
                // ragged end of the group: at most 3 elements, one thread
                for (size_t i = e[k]; i < G.n; i++) {
                    if (MASKED && L.visible[i / G.row_width] == 0) continue;
                    float pp = G.p[i], mm = G.m[i], vv = G.v[i];
                    adam_update(pp, G.g[i], mm, vv, c, G.bc2_sqrt, G.neg_step);
                    G.p[i] = pp; G.m[i] = mm; G.v[i] = vv;
                }
            }
        }
    } else {
        // some tensor of this launch is not 16-byte aligned: element-wise, still coalesced
#pragma unroll 4
        for (int k = 0; k < ADAM_ITERS * 4; k++) {
            const size_t i = base + (size_t)k * ADAM_THREADS + threadIdx.x;
            if (i >= G.n) break;
            if (MASKED && L.visible[i / G.row_width] == 0) continue;
            float pp = G.p[i], mm = G.m[i], vv = G.v[i];
            adam_update(pp, G.g[i], mm, vv, c, G.bc2_sqrt, G.neg_step);
            G.p[i] = pp; G.m[i] = mm; G.v[i] = vv;
        }
    }
}

size_t adam_blocks(size_t n) { return (n + ADAM_TILE - 1) / ADAM_TILE; }

void launch_adam(const AdamLaunch& L, unsigned total_blocks, bool vec, cudaStream_t st) {
    if (total_blocks == 0) return;
    const bool masked = L.visible != nullptr;
    if (vec) {
        if (masked) adam_kernel<true, true><<<total_blocks, ADAM_THREADS, 0, st>>>(L);
        else adam_kernel<true, false><<<total_blocks, ADAM_THREADS, 0, st>>>(L);
    } else {
        if (masked) adam_kernel<false, true><<<total_blocks, ADAM_THREADS, 0, st>>>(L);
        else adam_kernel<false, false><<<total_blocks, ADAM_THREADS, 0, st>>>(L);
    }
    count_launch(1);
}

}  // namespace gsr
