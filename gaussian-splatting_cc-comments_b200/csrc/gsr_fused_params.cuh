// gsr_fused_params.cuh -- the model's RAW parameters inside the per-Gaussian kernels.
//
// The reference's render() (gaussian_renderer/__init__.py:69-101) hands the rasterizer ACTIVATED tensors, each made by
// separate torch kernels every iteration (scene/gaussian_model.py:114-138):
//     scales    = torch.exp(_scaling)                          opacity = torch.sigmoid(_opacity)
//     rotations = torch.nn.functional.normalize(_rotation)     shs     = torch.cat((_features_dc, _features_rest), dim=1)
// and autograd runs their backward after the rasterizer's.  The helpers here let preprocess_fwd_kernel and the
// per-Gaussian backward take `_scaling`, `_rotation`, `_opacity`, `_features_dc` and `_features_rest` as they are stored:
//   * the three activations and their derivatives, in torch's arithmetic (same expressions as ATen's CUDA functors);
//   * SH colour from SPLIT storage: the (P,1,3) DC block and the (P,M-1,3) rest block, whose 12*(M-1)-byte rows are only
//     4-byte aligned -- read as aligned 16-byte vectors starting up to 3 floats early and consumed with a compile-time
//     shift (4 variants), so the traffic is the same 12 vectors per Gaussian as the concatenated layout;
//   * the backward's staging of rest rows into its 16-byte aligned shared-memory rows and the split write-out of dL_dsh
//     into dL_dfeatures_dc / dL_dfeatures_rest (what torch.cat's backward would otherwise copy out of dL_dsh).
// Pure index / arithmetic code: compiled for the host by tests/ (GSR_EMU) to check every mapping against numpy.
#pragma once
#include "gsr_math.cuh"

namespace gsr {

// ---- activations (forward) -------------------------------------------------------------------------------------------
GSR_DI float act_sigmoid(float x) { return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x))); }     // ATen sigmoid: 1 / (1 + exp(-x))
GSR_DI float act_exp(float x) { return expf(x); }

struct QuatNorm { float4 q; float norm, denom; };     // q = r / denom, denom = max(|r|, 1e-12)  (F.normalize, p = 2, eps = 1e-12)
GSR_DI QuatNorm act_normalize(float4 r) {
    QuatNorm o;
    o.norm = __fsqrt_rn(fmar(r.w, r.w, fmar(r.z, r.z, fmar(r.y, r.y, mulr(r.x, r.x)))));
    o.denom = fmaxf(o.norm, 1e-12f);
    o.q = make_float4(__fdiv_rn(r.x, o.denom), __fdiv_rn(r.y, o.denom), __fdiv_rn(r.z, o.denom), __fdiv_rn(r.w, o.denom));
    return o;
}

// ---- activations (backward): gradient w.r.t. the raw value from the gradient w.r.t. the activated one ------------------
GSR_DI float act_sigmoid_bwd(float y, float g) { return g * (1.0f - y) * y; }                 // sigmoid_backward: grad * (1 - y) * y
GSR_DI float act_exp_bwd(float y, float g) { return g * y; }                                  // exp backward: grad * result
GSR_DI float4 act_normalize_bwd(const QuatNorm& f, float4 g) {
    const float inv = 1.0f / f.denom;
    if (!(f.norm >= 1e-12f)) return make_float4(g.x * inv, g.y * inv, g.z * inv, g.w * inv);  // clamped: denom is a constant
    const float d = f.q.x * g.x + f.q.y * g.y + f.q.z * g.z + f.q.w * g.w;                    // (g - q (q.g)) / |r|
    return make_float4((g.x - f.q.x * d) * inv, (g.y - f.q.y * d) * inv, (g.z - f.q.z * d) * inv, (g.w - f.q.w * d) * inv);
}

// ---- SH colour from split storage (forward) ------------------------------------------------------------------------
// res[ch] = k[0]*dc[ch] + sum_{kk>=1} k[kk]*rest[3(kk-1)+ch], accumulated in the same order as the concatenated path
// (sh_color_vec in gsr_preprocess.cu), so both layouts give bit-identical colours.
template <int DEG, int SHIFT>
GSR_DI void sh_rest_accum(const float4* __restrict__ base, const float* k, float* res) {
    constexpr int K = (DEG + 1) * (DEG + 1);
    constexpr int NF = 3 * (K - 1);
    constexpr int NV = (SHIFT + NF + 3) / 4;
#pragma unroll
    for (int i = 0; i < NV; i++) {
        const float4 v = __ldg(base + i);
        const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int f = 4 * i + j - SHIFT;          // index into this Gaussian's rest row
            if (f >= 0 && f < NF) res[f % 3] = fmar(k[f / 3 + 1], e[j], res[f % 3]);
        }
    }
}

template <int DEG>
GSR_DI void sh_color_split(const float* __restrict__ dc, const float* __restrict__ rest_row, bool vec_ok, const float* k, float* res) {
    constexpr int K = (DEG + 1) * (DEG + 1);
#pragma unroll
    for (int ch = 0; ch < 3; ch++) res[ch] = mulr(__ldg(dc + ch), k[0]);
    if (DEG == 0) return;
    if (vec_ok) {
        const int shift = (int)((reinterpret_cast<uintptr_t>(rest_row) >> 2) & 3);
        const float4* base = reinterpret_cast<const float4*>(rest_row - shift);
        switch (shift) {
            case 0: sh_rest_accum<DEG, 0>(base, k, res); break;
            case 1: sh_rest_accum<DEG, 1>(base, k, res); break;
            case 2: sh_rest_accum<DEG, 2>(base, k, res); break;
            default: sh_rest_accum<DEG, 3>(base, k, res); break;
        }
    } else {
        for (int kk = 1; kk < K; kk++)
            for (int ch = 0; ch < 3; ch++) res[ch] = fmar(k[kk], __ldg(rest_row + 3 * (kk - 1) + ch), res[ch]);
    }
}

// ---- backward staging: rest rows -> shared rows [dc(3) | rest] of stride S floats -------------------------------------------
GSR_DI void copy4_to_shared(float* dst_shared, const float* src_global) {
#ifdef GSR_EMU
    *dst_shared = *src_global;
#else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"((uint32_t)__cvta_generic_to_shared(dst_shared)), "l"(src_global) : "memory");
#endif
}

// src: the block's rest rows, nrows * Lr contiguous floats.  Only rows with s_vis set and columns < live are fetched.
// Thread t of NT takes floats t, t + NT, ... (coalesced); row / column advance without a division per element.
template <int NT>
GSR_DI void stage_rest_rows(float* s_rows, int S, const float* __restrict__ src, int nrows, int Lr, int live,
                            const uint8_t* s_vis, int t) {
    if (Lr <= 0) return;
    const int n = nrows * Lr;
    const int qstep = NT / Lr, rstep = NT % Lr;
    int row = t / Lr, col = t - row * Lr;
    for (int f = t; f < n; f += NT) {
        if (s_vis[row] && col < live) copy4_to_shared(s_rows + row * S + 3 + col, src + f);
        row += qstep; col += rstep;
        if (col >= Lr) { col -= Lr; row++; }
    }
}

// Shared rows hold dL_dsh[0 .. live) (zero beyond is NOT assumed) -> dc_out (nrows,3) and rest_out (nrows,Lr), zeros for
// culled Gaussians and for coefficients beyond the active degree; 16-byte stores into rest_out when it is aligned.
template <int NT>
GSR_DI void store_split_rows(const float* s_rows, int S, float* __restrict__ dc_out, float* __restrict__ rest_out, int nrows,
                             int Lr, int live /* 3K */, const uint8_t* s_vis, int t) {
    for (int f = t; f < 3 * nrows; f += NT) {
        const int row = f / 3, c = f - 3 * row;
        dc_out[f] = s_vis[row] ? s_rows[row * S + c] : 0.f;
    }
    if (Lr <= 0) return;
    const int n = nrows * Lr;
    const int n4 = ((reinterpret_cast<uintptr_t>(rest_out) & 15) == 0) ? (n >> 2) : 0;
    {
        const int qstep = (4 * NT) / Lr, rstep = (4 * NT) % Lr;
        int row = (4 * t) / Lr, col = 4 * t - row * Lr;
        for (int e4 = t; e4 < n4; e4 += NT) {
            float v[4];
            int r = row, c = col;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                v[j] = (s_vis[r] && 3 + c < live) ? s_rows[r * S + 3 + c] : 0.f;
                if (++c >= Lr) { c = 0; r++; }
            }
            reinterpret_cast<float4*>(rest_out)[e4] = make_float4(v[0], v[1], v[2], v[3]);
            row += qstep; col += rstep;
            if (col >= Lr) { col -= Lr; row++; }
        }
    }
    for (int f = 4 * n4 + t; f < n; f += NT) {
        const int row = f / Lr, c = f - row * Lr;
        rest_out[f] = (s_vis[row] && 3 + c < live) ? s_rows[row * S + 3 + c] : 0.f;
    }
}

}  // namespace gsr
