// gsr_preprocess.cu -- forward per-Gaussian stage for sm_100a.
//
// Replaces preprocessCUDA<3> (DGR/cuda_rasterizer/forward.cu:192-324) and checkFrustum
// (DGR/cuda_rasterizer/rasterizer_impl.cu:56-69).  One thread per Gaussian; HBM-bound
// (P*(44+8) + P_vis*(12K+40) algorithmic bytes).  Differences from the
// reference in structure (not in results):
//   * camera matrices are staged once per block in shared memory,
//   * SH coefficients are fetched as 16-byte vectors through the read-only path,
//   * everything the blending kernels need is written as ONE 48-byte record per Gaussian
//     (3 x st.v4) instead of five separate arrays, plus conservative cull extents,
//   * cov3D is not stored (the backward recomputes it from scale/rotation),
//   * the depth-sort key/value pair is emitted here.
#include "gsr_internal.cuh"
#include "gsr_math.cuh"
#include "gsr_smem.cuh"
#include "gsr_fused_params.cuh"
#include <stdio.h>

namespace gsr {

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

// colour from SH (forward.cu:21-81); vector path
template <int DEG>
__device__ __forceinline__ void sh_color_vec(const float* __restrict__ sh, const float* k, float* res) {
    constexpr int K = (DEG + 1) * (DEG + 1);
    constexpr int NF = 3 * K;
    constexpr int NV = (NF + 3) / 4;
#pragma unroll
    for (int i = 0; i < NV; i++) {
        const float4 v = ldg4(sh + 4 * i);
        const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int f = 4 * i + j;
            if (f < NF) {
                const int kk = f / 3, ch = f % 3;
                res[ch] = (kk == 0) ? mulr(e[j], k[0]) : fmar(k[kk], e[j], res[ch]);
            }
        }
    }
}

__device__ __forceinline__ void sh_color_scalar(int K, const float* __restrict__ sh, const float* k, float* res) {
    for (int ch = 0; ch < 3; ch++) res[ch] = mulr(__ldg(sh + ch), k[0]);
    for (int kk = 1; kk < K; kk++)
        for (int ch = 0; ch < 3; ch++) res[ch] = fmar(k[kk], __ldg(sh + 3 * kk + ch), res[ch]);
}

// 4 blocks per SM (64 registers): measured 0.194 -> 0.170 ms at C2 against the 85-register build
// RAW  : opacities / scales / rotations hold the model's raw parameters; sigmoid / exp / normalize are applied here.
// SPLIT: SH coefficients come from a.f_dc (P,1,3) and a.f_rest (P,M-1,3) instead of a.shs (P,M,3).
// <false,false> is the reference-shaped path; the other instantiations are separate kernels, it is not touched by them.
template <bool RAW, bool SPLIT>
__global__ void __launch_bounds__(256, 4) preprocess_fwd_kernel(PreprocessArgs a) {
    __shared__ float s_cam[36];
    if (threadIdx.x < 16) s_cam[threadIdx.x] = a.cam.view[threadIdx.x];
    else if (threadIdx.x < 32) s_cam[threadIdx.x] = a.cam.proj[threadIdx.x - 16];
    else if (threadIdx.x < 35) s_cam[threadIdx.x] = a.cam.campos[threadIdx.x - 32];
    __syncthreads();
    const float* V = s_cam;
    const float* Pm = s_cam + 16;

    // threads past the end redo the last Gaussian without storing anything: the whole block reaches the reduction below
    const int idx_raw = blockIdx.x * blockDim.x + threadIdx.x;
    const bool in_range = idx_raw < a.P;
    const int idx = in_range ? idx_raw : a.P - 1;

    int out_radius = 0;
    uint32_t out_tiles = 0;
    uint32_t out_key = 0xFFFFFFFFu;   // sorts behind every visible Gaussian; emits nothing

    const float mx = __ldg(a.means3D + 3 * idx), my = __ldg(a.means3D + 3 * idx + 1), mz = __ldg(a.means3D + 3 * idx + 2);
    // every per-Gaussian input except the SH row is fetched up front (one memory round trip instead of three)
    float4 q_in = make_float4(0.f, 0.f, 0.f, 1.f);
    float sx_in = 0.f, sy_in = 0.f, sz_in = 0.f;
    if (a.cov3D_precomp == nullptr) {
        if ((reinterpret_cast<uintptr_t>(a.rotations) & 15) == 0) q_in = ldg4(a.rotations + 4 * (size_t)idx);
        else q_in = make_float4(__ldg(a.rotations + 4 * (size_t)idx), __ldg(a.rotations + 4 * (size_t)idx + 1),
                                __ldg(a.rotations + 4 * (size_t)idx + 2), __ldg(a.rotations + 4 * (size_t)idx + 3));
        sx_in = __ldg(a.scales + 3 * idx); sy_in = __ldg(a.scales + 3 * idx + 1); sz_in = __ldg(a.scales + 3 * idx + 2);
    }
    float opacity = __ldg(a.opacities + idx);
    if (RAW) {
        opacity = act_sigmoid(opacity);
        if (a.cov3D_precomp == nullptr) {
            sx_in = act_exp(sx_in); sy_in = act_exp(sy_in); sz_in = act_exp(sz_in);
            q_in = act_normalize(q_in).q;
        }
    }
    // in_frustum (auxiliary.h:144-175): only the near-plane test is live
    const float depth = dot4(V[2], V[6], V[10], V[14], mx, my, mz);
    bool alive = in_range && !(depth <= 0.2f);
    // The SH row is needed only after the whole projection has been computed (a second dependent memory round trip): ask L2
    // for its two 128-byte lines now.  A fifth of these rows belong to Gaussians that turn out to be off-screen.
    if (alive && !SPLIT && a.colors_precomp == nullptr) {
        const char* row = reinterpret_cast<const char*>(a.shs + (size_t)idx * a.M * 3);
        prefetch_l2(row);
        if (a.M * 12 > 128) prefetch_l2(row + 128);
    }
    if (in_range && !alive && a.prefiltered) {
        printf("Point is filtered although prefiltered is set. This shouldn't happen!");
        __trap();
    }
    if (alive) {
        const float hx = dot4(Pm[0], Pm[4], Pm[8], Pm[12], mx, my, mz);
        const float hy = dot4(Pm[1], Pm[5], Pm[9], Pm[13], mx, my, mz);
        const float hw = dot4(Pm[3], Pm[7], Pm[11], Pm[15], mx, my, mz);
        const float p_w = __frcp_rn(addr(hw, 0.0000001f));
        const float projx = mulr(hx, p_w), projy = mulr(hy, p_w);

        float c3[6];
        if (a.cov3D_precomp != nullptr) {
#pragma unroll
            for (int i = 0; i < 6; i++) c3[i] = __ldg(a.cov3D_precomp + 6 * (size_t)idx + i);
        } else {
            cov3d_from_scale_rot(sx_in, sy_in, sz_in, a.scale_modifier, q_in, c3);
        }
        const EwaT e = ewa_T(mx, my, mz, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, V);
        const float3 cov = ewa_cov2d(e, c3);
        const float det = fmar(cov.x, cov.z, -mulr(cov.y, cov.y));
        if (det != 0.0f) {
            const float det_inv = __frcp_rn(det);
            const float conx = mulr(cov.z, det_inv), cony = mulr(cov.y, -det_inv), conz = mulr(cov.x, det_inv);
            const float mid = mulr(addr(cov.x, cov.z), 0.5f);
            const float sq = __fsqrt_rn(fmaxf(fmar(mid, mid, -det), 0.1f));
            const float lambda1 = addr(mid, sq), lambda2 = subr(mid, sq);
            const float my_radius = ceilf(mulr(__fsqrt_rn(fmaxf(lambda1, lambda2)), 3.0f));
            const int irad = __float2int_rz(my_radius);
            const float px = ndc2pix(projx, a.W), py = ndc2pix(projy, a.H);
            uint32_t x0, y0, x1, y1;
            get_rect(px, py, irad, (uint32_t)a.grid_x, (uint32_t)a.grid_y, x0, y0, x1, y1);
            // band-clipped tile count (tile-row sharded rendering); the full grid by default, then band_tiles is the whole rect
            const uint32_t by0 = max(y0, (uint32_t)a.row_begin), by1 = min(y1, (uint32_t)a.row_end);
            const uint32_t band_tiles = (by1 > by0) ? (by1 - by0) * (x1 - x0) : 0u;
            // radii are those of the full image on every rank; a Gaussian that misses this rank's band emits nothing, so its
            // SH row is not even read and no record is written (render-only sharding: 7/8 of the SH traffic at 8 ranks)
            if ((x1 - x0) * (y1 - y0) != 0) out_radius = irad;
            if (band_tiles != 0) {
                // ---- colour -------------------------------------------------------------------
                float rgb[3];
                uint32_t clampbits = 0;
                if (a.colors_precomp == nullptr) {
                    const float dx = subr(mx, V[32]), dy = subr(my, V[33]), dz = subr(mz, V[34]);
                    const float len = __fsqrt_rn(fmar(dz, dz, fmar(dx, dx, mulr(dy, dy))));
                    const float x = __fdiv_rn(dx, len), y = __fdiv_rn(dy, len), z = __fdiv_rn(dz, len);
                    float k[16];
                    sh_basis(a.D, x, y, z, k);
                    float res[3] = {0.f, 0.f, 0.f};
                    if (SPLIT) {
                        const float* dc = a.f_dc + 3 * (size_t)idx;
                        const float* rest = a.f_rest + (size_t)idx * (a.M - 1) * 3;
                        // aligned 16-byte reads start up to 3 floats before the row and end up to 3 floats after it:
                        // inside the tensor for every Gaussian but the last one
                        const bool vec_ok = ((reinterpret_cast<uintptr_t>(a.f_rest) & 15) == 0) && idx < a.P - 1;
                        if (a.D == 3) sh_color_split<3>(dc, rest, vec_ok, k, res);
                        else if (a.D == 2) sh_color_split<2>(dc, rest, vec_ok, k, res);
                        else if (a.D == 1) sh_color_split<1>(dc, rest, vec_ok, k, res);
                        else sh_color_split<0>(dc, rest, vec_ok, k, res);
                    } else {
                    const float* sh = a.shs + (size_t)idx * a.M * 3;
                    const bool vec_ok = ((a.M * 3) % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.shs) & 15) == 0);
                    if (vec_ok && a.D == 3) sh_color_vec<3>(sh, k, res);
                    else if (vec_ok && a.D == 2 && a.M * 3 >= 28) sh_color_vec<2>(sh, k, res);
                    else if (vec_ok && a.D == 1) sh_color_vec<1>(sh, k, res);
                    else if (vec_ok && a.D == 0) sh_color_vec<0>(sh, k, res);
                    else sh_color_scalar((a.D + 1) * (a.D + 1), sh, k, res);
                    }
#pragma unroll
                    for (int ch = 0; ch < 3; ch++) {
                        const float v = addr(res[ch], 0.5f);
                        if (v < 0.0f) clampbits |= 1u << ch;
                        rgb[ch] = (v < 0.0f) ? 0.0f : v;
                    }
                } else {
#pragma unroll
                    for (int ch = 0; ch < 3; ch++) rgb[ch] = __ldg(a.colors_precomp + 3 * (size_t)idx + ch);
                }

                // ---- conservative cull data for the blending kernels ------------------------------
                // alpha = min(.99, o*exp(power)) >= 1/255 needs power >= -ln(255 o).  The ellipse
                // {q <= q0} has axis-aligned half extents sqrt(q0*cov.x), sqrt(q0*cov.z).  The extents
                // are inflated for the rounding of `power` (cancellation grows with the anisotropy
                // lambda1/lambda2), so a culled pair is always one the exact test would skip.
                float ext_x = -1.0f, ext_y = -1.0f, thr = 1.0f;
                const float o255 = opacity * 255.0f;
                if (o255 > 1.0f) {   // NaN opacity -> never contributes in the reference either
                    const float L = logf(o255);
                    thr = -L - 1e-4f * L - 1e-4f;
                    const float q0 = 2.0f * L * 1.0002f + 2e-4f;
                    const float kappa = lambda1 / fmaxf(lambda2, 1e-12f);
                    const float infl = 1.0f + 1e-3f + 2e-6f * kappa;
                    ext_x = sqrtf(q0 * cov.x) * infl + 1e-2f;
                    ext_y = sqrtf(q0 * cov.z) * infl + 1e-2f;
                    if (!(ext_x == ext_x) || !(ext_y == ext_y)) { ext_x = 3.0e38f; ext_y = 3.0e38f; }
                }
                float4* rec = a.packed + 3 * (size_t)idx;
                rec[0] = make_float4(px, py, conx, cony);
                rec[1] = make_float4(conz, thr, opacity, rgb[0]);
                rec[2] = make_float4(rgb[1], rgb[2], ext_x, ext_y);
                a.clamped[idx] = (uint8_t)clampbits;

                out_tiles = band_tiles;
                out_key = __float_as_uint(depth);
                // the duplicate stage reads this back in depth order (8 sequential bytes instead of gathering record and radius)
                if (a.rects) a.rects[idx] = make_uint2(x0 | (by0 << 16), x1 - x0);
            } else if (out_radius != 0) {
                // visible, but not in this rank's band: a neutral record (a banded backward reads it beside all-zero accumulators)
                float4* rec = a.packed + 3 * (size_t)idx;
                rec[0] = make_float4(px, py, 0.f, 0.f);
                rec[1] = make_float4(0.f, 1.0f, 0.f, 0.f);
                rec[2] = make_float4(0.f, 0.f, -1.0f, -1.0f);
                a.clamped[idx] = 0;
            }
        }
    }
    if (in_range) {
        if (a.radii) a.radii[idx] = out_radius;
        a.tiles_touched[idx] = out_tiles;
        a.depth_keys[idx] = out_key;
        a.ids[idx] = (uint32_t)idx;
    }
    // num_rendered = sum of tiles_touched is known here, long before the sorted prefix sum: the host reads it back
    // while the depth sort runs (gsr_api.cu)
    if (a.total != nullptr) {
        __shared__ uint32_t s_sum[8];
        const uint32_t wsum = __reduce_add_sync(0xffffffffu, out_tiles);
        if ((threadIdx.x & 31) == 0) s_sum[threadIdx.x >> 5] = wsum;
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t bsum = 0;
#pragma unroll
            for (int w = 0; w < 8; w++) bsum += s_sum[w];
            if (bsum) atomicAdd(a.total, bsum);
        }
    }
}

void launch_preprocess(const PreprocessArgs& a, cudaStream_t st) {
    if (a.P <= 0) return;
    const unsigned nb = (unsigned)((a.P + 255) / 256);
    const bool split = a.f_dc != nullptr;
    if (!a.raw_act && !split) preprocess_fwd_kernel<false, false><<<nb, 256, 0, st>>>(a);
    else if (a.raw_act && !split) preprocess_fwd_kernel<true, false><<<nb, 256, 0, st>>>(a);
    else if (!a.raw_act && split) preprocess_fwd_kernel<false, true><<<nb, 256, 0, st>>>(a);
    else preprocess_fwd_kernel<true, true><<<nb, 256, 0, st>>>(a);
    count_launch(1);
}

// rasterizer_impl.cu:56-69 (checkFrustum) -> present[idx] = p_view.z > 0.2 (auxiliary.h:165)
__global__ void __launch_bounds__(256) mark_visible_kernel(int P, const float* __restrict__ means,
                                                           const float* __restrict__ view,
                                                           unsigned char* __restrict__ present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float v2 = __ldg(view + 2), v6 = __ldg(view + 6), v10 = __ldg(view + 10), v14 = __ldg(view + 14);
    const float depth = dot4(v2, v6, v10, v14, __ldg(means + 3 * idx), __ldg(means + 3 * idx + 1), __ldg(means + 3 * idx + 2));
    present[idx] = !(depth <= 0.2f);
}

void launch_mark_visible(int P, const float* means, const float* view, unsigned char* present, cudaStream_t st) {
    if (P <= 0) return;
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, means, view, present);
    count_launch(1);
}

}  // namespace gsr
