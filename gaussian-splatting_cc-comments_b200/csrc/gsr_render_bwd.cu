// gsr_render_bwd.cu -- backward of the per-tile alpha blending for sm_100a.
//
// Replaces renderCUDA<3> (backward) (DGR/cuda_rasterizer/backward.cu:408-601), which issues nine
// scalar float atomicAdd per (pixel, Gaussian) pair.  Here:
//   * the tile is replayed back to front starting at the tile's LAST contributor (block max of
//     n_contrib) instead of at the end of the tile's list;
//   * each warp owns an 8x4 pixel block and culls staged Gaussians against it exactly as the
//     forward does (same extents, same `power < thr` skip, so the same pairs are blended);
//   * the nine per-pair gradient terms are reduced across the warp with a shuffle
//     reduce-scatter (8 values in 9 shuffles + 1 value in 5) and leave as ONE 8-lane
//     red.global.add.f32 on consecutive addresses + one scalar red, into a (P,12) float scratch
//     that the preprocess backward consumes;
// Per-pixel replay order (the T division chain, backward.cu:534) is unchanged; only the
// cross-pixel summation order differs (gradient tolerance 1e-4 relative L2).
#include "gsr_internal.cuh"
#include "gsr_math.cuh"
#include "gsr_smem.cuh"

namespace gsr {

constexpr int RBB = GSR_TILE_PIX;
constexpr int BOFF_B = RBB * 16, BOFF_C = 2 * RBB * 16, BOFF_ID = 3 * RBB * 16;   // byte offsets of the b / c / id planes
constexpr int BOFF_MB = BOFF_ID + RBB * 4;                                         // per-instance warp masks (1 byte each)

template <int N>
__device__ __forceinline__ void reduce_scatter_step(float* v, int bit, bool upper) {
#pragma unroll
    for (int k = 0; k < N / 2; k++) {
        const float send = upper ? v[k] : v[k + N / 2];
        const float keep = upper ? v[k + N / 2] : v[k];
        v[k] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
    }
}

// 5 blocks per SM: 6 and 8 measured slower (1.408 / 1.395 vs 1.360 ms at C2)
__global__ void __launch_bounds__(256, 5) render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                                                         int W, int H, int grid_x, int row_begin,
                                                         const float4* __restrict__ packed, const float* __restrict__ bg,
                                                         const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                                                         const float* __restrict__ dL_dpixels, float* __restrict__ grad_acc,
                                                         const uint8_t* __restrict__ warp_mask) {
    __shared__ __align__(16) float4 s_rec[3 * RBB + RBB / 4 + RBB / 16];    // planes a | b | c of the staged records, ids, warp masks
    __shared__ uint32_t s_max[8];
    const uint32_t sb = smem_addr(s_rec);

    const int tile_x = blockIdx.x, tile_y = blockIdx.y + row_begin;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int bx0 = tile_x * GSR_TILE_X + (warp & 1) * 8;
    const int by0 = tile_y * GSR_TILE_Y + (warp >> 1) * 4;
    const int pix_x = bx0 + (lane & 7), pix_y = by0 + (lane >> 3);
    const bool inside = pix_x < W && pix_y < H;
    const float pixfx = (float)pix_x, pixfy = (float)pix_y;
    const size_t pix_id = (size_t)pix_y * W + pix_x;
    const size_t HW = (size_t)H * W;

    const uint2 range = ranges[tile_y * grid_x + tile_x];

    const float T_final = inside ? final_Ts[pix_id] : 0.0f;
    float T = T_final;
    const int last_contributor = inside ? (int)n_contrib[pix_id] : 0;
    float dpix0 = 0.f, dpix1 = 0.f, dpix2 = 0.f;
    if (inside) {
        dpix0 = dL_dpixels[pix_id];
        dpix1 = dL_dpixels[HW + pix_id];
        dpix2 = dL_dpixels[2 * HW + pix_id];
    }
    const float bg_dot_dpixel = __ldg(bg) * dpix0 + __ldg(bg + 1) * dpix1 + __ldg(bg + 2) * dpix2;
    float ar0 = 0.f, ar1 = 0.f, ar2 = 0.f;       // accum_rec
    float lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;       // last_color
    float last_alpha = 0.f;

    // block / warp maxima of the last contributor: nothing behind them has a gradient
    const int wmax = (int)__reduce_max_sync(0xffffffffu, (uint32_t)last_contributor);
    if (lane == 0) s_max[warp] = (uint32_t)wmax;
    __syncthreads();
    int bmax = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) bmax = max(bmax, (int)s_max[w]);
    const int rounds = (bmax + RBB - 1) / RBB;

    const uint32_t my_slot = sb + threadIdx.x * 16;
    const bool red_lane = (lane & 3) == 0 || lane == 1;     // lanes that hold a finished sum after the reduce-scatter
    const int red_idx = lane == 1 ? 8 : (lane >> 2);
    for (int i = 0; i < rounds; i++) {
        __syncthreads();
        const int first_pos = bmax - 1 - i * RBB;          // position staged by thread 0
        const int mypos = first_pos - (int)threadIdx.x;
        if (mypos >= 0) {
            const uint32_t id = point_list[range.x + mypos];
            const float4* rec = packed + 3 * (size_t)id;
            const float4 ra = __ldg(rec), rb = __ldg(rec + 1), rc = __ldg(rec + 2);
            sts128<0>(my_slot, ra);
            sts128<BOFF_B>(my_slot, rb);
            sts128<BOFF_C>(my_slot, rc);
            sts32u<BOFF_ID>(sb + threadIdx.x * 4, id);
            sts8u<BOFF_MB>(sb + threadIdx.x, warp_mask[range.x + mypos]);
        }
        __syncthreads();
        const int count = min(RBB, first_pos + 1);
        if (first_pos - (count - 1) >= wmax) continue;      // whole batch is behind this warp's last contributor
        for (int cbase = 0; cbase < count; cbase += 32) {
            // lane l tests staged Gaussian cbase + 31 - l: the next one to replay is the highest set ballot bit
            const int jj = cbase + 31 - lane;
            // the forward recorded, per instance, which warps of the tile blended it (gsr_render.cu): exactly the
            // (Gaussian, warp) pairs with a gradient, so no geometric test and no empty replays here
            bool hit = false;
            if (jj < count && first_pos - jj < wmax) hit = (lds8u<BOFF_MB>(sb + (uint32_t)jj) >> warp) & 1u;
            uint32_t m = __ballot_sync(0xffffffffu, hit);
            while (m) {
                const int top = 31 - __clz(m);
                m ^= 1u << top;
                const int j = cbase + 31 - top;
                const int pos = first_pos - j;
                const uint32_t ja = sb + (uint32_t)j * 16;
                float v[8];
                float v8;
                // Branch-free replay: the forward's warp mask guarantees at least one contributing lane, so every
                // instruction below issues anyway; lanes that do not contribute run with alpha = G = 0, which
                // leaves T (x 1/(1-0)), the accum_rec recurrence (0*c + 1*rec) and all nine sums unchanged.
                const float4 a = lds128<0>(ja);
                const float4 b = lds128<BOFF_B>(ja);
                const float2 gb = lds64<BOFF_C>(ja);
                const float dx = subr(a.x, pixfx), dy = subr(a.y, pixfy);
                const float s = fmar(dx, mulr(dx, a.z), mulr(dy, mulr(dy, b.x)));
                const float power = fmar(s, -0.5f, -mulr(dy, mulr(dx, a.w)));
                const float G0 = expf(power);
                const float alpha0 = fminf(0.99f, mulr(b.z, G0));
                const bool c = (pos < last_contributor) && !(power > 0.0f) && !(power < b.y) && !(alpha0 < 1.0f / 255.0f);
                const float alpha = c ? alpha0 : 0.0f;
                const float G = c ? G0 : 0.0f;
                const float inv = rcp_approx_ftz(1.0f - alpha);
                T = T * inv;
                const float dchannel_dcolor = alpha * T;
                ar0 = last_alpha * lc0 + (1.f - last_alpha) * ar0;
                ar1 = last_alpha * lc1 + (1.f - last_alpha) * ar1;
                ar2 = last_alpha * lc2 + (1.f - last_alpha) * ar2;
                lc0 = b.w; lc1 = gb.x; lc2 = gb.y;
                float dL_dalpha = (lc0 - ar0) * dpix0 + (lc1 - ar1) * dpix1 + (lc2 - ar2) * dpix2;
                v[6] = dchannel_dcolor * dpix0;
                v[7] = dchannel_dcolor * dpix1;
                v8 = dchannel_dcolor * dpix2;
                dL_dalpha *= T;
                last_alpha = alpha;
                dL_dalpha += (-T_final * inv) * bg_dot_dpixel;
                const float w = (b.z * dL_dalpha) * G;
                const float wx = w * dx, wy = w * dy;
                v[0] = wx; v[1] = wy; v[2] = wx * dx; v[3] = wx * dy; v[4] = wy * dy;
                v[5] = G * dL_dalpha;
                reduce_scatter_step<8>(v, 16, (lane & 16) != 0);
                reduce_scatter_step<4>(v, 8, (lane & 8) != 0);
                reduce_scatter_step<2>(v, 4, (lane & 4) != 0);
                float r = v[0];
                r += __shfl_xor_sync(0xffffffffu, r, 2);
                r += __shfl_xor_sync(0xffffffffu, r, 1);
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) v8 += __shfl_xor_sync(0xffffffffu, v8, d);
                float* dst = grad_acc + 12 * (size_t)lds32u<BOFF_ID>(sb + (uint32_t)j * 4);
                if (red_lane) atomicAdd(dst + red_idx, lane == 1 ? v8 : r);
            }
        }
    }
}

void launch_render_backward(int grid_x, int row_begin, int row_end, const uint2* ranges, const uint32_t* point_list,
                            int W, int H, const float4* packed, const float* bg, const float* final_T,
                            const uint32_t* n_contrib, const float* dL_dpix, float* grad_acc, const uint8_t* warp_mask,
                            cudaStream_t st) {
    if (row_end <= row_begin || grid_x <= 0) return;
    dim3 grid(grid_x, row_end - row_begin, 1);
    render_bwd_kernel<<<grid, 256, 0, st>>>(ranges, point_list, W, H, grid_x, row_begin, packed, bg, final_T, n_contrib,
                                             dL_dpix, grad_acc, warp_mask);
    count_launch(1);
}

}  // namespace gsr
