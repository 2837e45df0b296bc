// gsr_render_bwd.cu -- backward of the per-tile alpha blending for sm_100a.
//
// Replaces renderCUDA<3> (backward) (DGR/cuda_rasterizer/backward.cu:408-601), which issues nine
// scalar float atomicAdd per (pixel, Gaussian) pair.  The kernel is bound by warp-instruction issue
// slots (ncu: 83 % issue active, DRAM 4 %), so its design is about instructions per replayed pair:
//
//   * 128 threads per 16x16 tile: each warp owns an 8x8 pixel block and each lane TWO pixels (rows ly
//     and ly + 4 of the block, same column).  Everything that depends only on the Gaussian or on the
//     column (record loads, dx, dx * conic terms, loop control) is paid once per pixel pair, and the
//     per-pixel arithmetic runs as packed pairs (FADD2 / FMUL2 / FFMA2: two IEEE operations per issue slot).
//   * the tile is replayed back to front from its LAST contributor (block max of n_contrib); which
//     (Gaussian, block) pairs carry a gradient comes from the warp masks the forward recorded
//     (an 8x8 block = two of the forward's 8x4 blocks), so there is no geometric test and no empty replay.
//   * Phase 1 (lanes = pixels, sequential over the Gaussians of the block): only what needs the per-pixel
//     recurrences (T, accumulated colour) is computed -- per pixel the two weights
//         w   = opacity * G * dL/dalpha        (all positional / conic / opacity gradients are moments of w)
//         dch = alpha * T                      (dL/dcolour = sum dch * dL/dpixel)
//     and parked in shared memory, one 16-byte store per lane and replay, in a slot of a 16-slot batch.
//   * Phase 2 (lanes = Gaussians, once per 16 replays): lane (slot g, half h) walks 32 of the block's 64
//     pixels, accumulates the nine sums of its Gaussian in registers with packed FMAs, the two halves are
//     combined with one shuffle per sum, and the row leaves as TWO red.global.add.v4.f32 + one scalar red
//     into the (P,12) accumulator the preprocess backward consumes.  This replaces the per-replay
//     warp-shuffle reduce-scatter (47 of 126 instructions per replay before) by about 19.
//   * records are staged by cp.async (LDGSTS) into two buffers: the copies of round r+1 fly while round r
//     is replayed, one block barrier per round; ids and masks of round r+2 are prefetched into registers.
// Per-pixel replay order (the T division chain, backward.cu:534) is unchanged; only the cross-pixel
// summation order differs (gradient tolerance 1e-4 relative L2).
#include "gsr_internal.cuh"
#include "gsr_math.cuh"
#include "gsr_smem.cuh"

namespace gsr {

constexpr int BW_T = 128;                 // threads per CTA (4 warps x 8x8 pixels)
constexpr int BW_R = 64;                  // records staged per round and buffer
constexpr int BW_SLOTS = 16;              // replays per phase-2 batch
// one staging buffer: planes a | b | c of the 48-byte records, ids, warp masks
constexpr int SG_B = BW_R * 16, SG_C = 2 * BW_R * 16, SG_ID = 3 * BW_R * 16, SG_MB = SG_ID + BW_R * 4;
constexpr int SG_BYTES = (SG_MB + BW_R + 15) & ~15;
// per warp: 32 rows (one per pixel lane) of 16 slots x 16 bytes {w0, w1, dch0, dch1} + 16 bytes (row k < 16: the
// Gaussian of slot k: x, y, opacity, id), then 32 rows of 24 bytes {dL/dpixel: c0p0 c0p1 c1p0 c1p1 c2p0 c2p1}
constexpr int SL_ROW = BW_SLOTS * 16 + 16;
constexpr int SL_INFO = BW_SLOTS * 16;
constexpr int DP_ROW = 24;
constexpr int WP_DP = 32 * SL_ROW;
constexpr int WP_BYTES = WP_DP + 32 * DP_ROW;
constexpr int BW_SMEM = 2 * SG_BYTES + 4 * WP_BYTES;

// Phase 2: the nine sums of the n <= 16 Gaussians parked in this warp's slots.  Warp-uniform call.
__device__ __forceinline__ void flush_slots(const uint32_t wbuf, const int n, const int lane, const float fx0, const float fy0,
                                            float* __restrict__ grad_acc) {
    __syncwarp();
    const int g = lane & 15, h = lane >> 4;
    const float4 info = lds128<SL_INFO>(wbuf + (uint32_t)g * SL_ROW);          // x, y, opacity, id
    // lane (g, h) covers pixel lanes p = 16 h + i, i = 0..15: column i & 7, rows 2h + (i >> 3) and that + 4
    const float fyh = fy0 + (float)(2 * h);
    const float2 dyA = make_float2(subr(info.y, fyh), subr(info.y, fyh + 4.0f));
    const float2 dyB = make_float2(subr(info.y, fyh + 1.0f), subr(info.y, fyh + 5.0f));
    float dxs[8];
#pragma unroll
    for (int i = 0; i < 8; i++) dxs[i] = subr(info.x, fx0 + (float)i);
    const uint32_t base = wbuf + (uint32_t)(h * 16 * SL_ROW + g * 16);
    const uint32_t dbase = wbuf + (uint32_t)(WP_DP + h * 16 * DP_ROW);
    float2 S = make_float2(0.f, 0.f), V0 = S, V1 = S, V2 = S, V3 = S, V4 = S, V6 = S, V7 = S, V8 = S;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        float4 e;
        float2 d0, d1, d2;
        switch (i) {   // immediate offsets (the loop is fully unrolled; the switch folds away)
#define GSR_ROW(I) case I: e = lds128<I * SL_ROW>(base); d0 = lds64<I * DP_ROW>(dbase); d1 = lds64<I * DP_ROW + 8>(dbase); d2 = lds64<I * DP_ROW + 16>(dbase); break;
            GSR_ROW(0) GSR_ROW(1) GSR_ROW(2) GSR_ROW(3) GSR_ROW(4) GSR_ROW(5) GSR_ROW(6) GSR_ROW(7)
            GSR_ROW(8) GSR_ROW(9) GSR_ROW(10) GSR_ROW(11) GSR_ROW(12) GSR_ROW(13) GSR_ROW(14) default: GSR_ROW(15)
#undef GSR_ROW
        }
        const float2 Wp = make_float2(e.x, e.y), Dc = make_float2(e.z, e.w);
        const float2 DY = i < 8 ? dyA : dyB;
        const float2 DX = bc2(dxs[i & 7]);
        const float2 WX = mul2(Wp, DX), WY = mul2(Wp, DY);
        S = add2(S, Wp);
        V0 = add2(V0, WX);
        V1 = add2(V1, WY);
        V2 = fma2(WX, DX, V2);
        V3 = fma2(WX, DY, V3);
        V4 = fma2(WY, DY, V4);
        V6 = fma2(Dc, d0, V6);
        V7 = fma2(Dc, d1, V7);
        V8 = fma2(Dc, d2, V8);
    }
    float v0 = V0.x + V0.y, v1 = V1.x + V1.y, v2 = V2.x + V2.y, v3 = V3.x + V3.y, v4 = V4.x + V4.y, s = S.x + S.y;
    float v6 = V6.x + V6.y, v7 = V7.x + V7.y, v8 = V8.x + V8.y;
    v0 += __shfl_xor_sync(0xffffffffu, v0, 16);
    v1 += __shfl_xor_sync(0xffffffffu, v1, 16);
    v2 += __shfl_xor_sync(0xffffffffu, v2, 16);
    v3 += __shfl_xor_sync(0xffffffffu, v3, 16);
    v4 += __shfl_xor_sync(0xffffffffu, v4, 16);
    s += __shfl_xor_sync(0xffffffffu, s, 16);
    v6 += __shfl_xor_sync(0xffffffffu, v6, 16);
    v7 += __shfl_xor_sync(0xffffffffu, v7, 16);
    v8 += __shfl_xor_sync(0xffffffffu, v8, 16);
    if (lane < n) {        // h == 0 lanes of the filled slots
        // sum G dL/dalpha = (sum w) / opacity; opacity 0 never contributes (alpha = 0 < 1/255), then s = 0 as well
        const float v5 = info.z != 0.0f ? __fdividef(s, info.z) : 0.0f;
        float* dst = grad_acc + 12 * (size_t)__float_as_uint(info.w);
        red_add_v4(dst, v0, v1, v2, v3);
        red_add_v4(dst + 4, v4, v5, v6, v7);
        red_add_f32(dst + 8, v8);
    }
    __syncwarp();          // the slots are free again
}

__global__ void __launch_bounds__(BW_T, 5) render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                                                            int W, int H, int grid_x, int row_begin,
                                                            const float4* __restrict__ packed, const float* __restrict__ bg,
                                                            const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                                                            const float* __restrict__ dL_dpixels, float* __restrict__ grad_acc,
                                                            const uint8_t* __restrict__ warp_mask) {
    __shared__ __align__(16) unsigned char s_raw[BW_SMEM];
    __shared__ uint32_t s_max[4];
    const uint32_t sb = smem_addr(s_raw);

    const int tile_x = blockIdx.x, tile_y = blockIdx.y + row_begin;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int bxi = warp & 1, byi = warp >> 1;
    const int bx0 = tile_x * GSR_TILE_X + bxi * 8;
    const int by0 = tile_y * GSR_TILE_Y + byi * 8;
    const int pix_x = bx0 + (lane & 7), pix_y0 = by0 + (lane >> 3), pix_y1 = pix_y0 + 4;
    const bool in0 = pix_x < W && pix_y0 < H, in1 = pix_x < W && pix_y1 < H;
    const float pixfx = (float)pix_x;
    const float2 npixfy = make_float2(-(float)pix_y0, -(float)pix_y1);         // a.y + (-py) == a.y - py
    const size_t id0 = (size_t)pix_y0 * W + pix_x, id1 = (size_t)pix_y1 * W + pix_x;
    const size_t HW = (size_t)H * W;
    // the forward's 8x4 blocks: warp w' covers columns 8 (w' & 1), rows 4 (w' >> 1); this block is w' = bxi + 4 byi and that + 2
    const uint32_t mbits = 0x5u << (bxi + 4 * byi);

    const uint2 range = ranges[tile_y * grid_x + tile_x];

    const float Tf0 = in0 ? final_Ts[id0] : 0.0f, Tf1 = in1 ? final_Ts[id1] : 0.0f;
    float2 T = make_float2(Tf0, Tf1);
    const int lastc0 = in0 ? (int)n_contrib[id0] : 0, lastc1 = in1 ? (int)n_contrib[id1] : 0;
    float2 dp0 = make_float2(0.f, 0.f), dp1 = dp0, dp2 = dp0;
    if (in0) { dp0.x = dL_dpixels[id0]; dp1.x = dL_dpixels[HW + id0]; dp2.x = dL_dpixels[2 * HW + id0]; }
    if (in1) { dp0.y = dL_dpixels[id1]; dp1.y = dL_dpixels[HW + id1]; dp2.y = dL_dpixels[2 * HW + id1]; }
    const float bg0 = __ldg(bg), bg1 = __ldg(bg + 1), bg2 = __ldg(bg + 2);
    // -T_final * (bg . dL/dpixel): the background term of dL/dalpha is this times 1 / (1 - alpha)
    const float2 nTfbg = make_float2(-Tf0 * (bg0 * dp0.x + bg1 * dp1.x + bg2 * dp2.x), -Tf1 * (bg0 * dp0.y + bg1 * dp1.y + bg2 * dp2.y));
    float2 ar0 = make_float2(0.f, 0.f), ar1 = ar0, ar2 = ar0;      // accum_rec per channel
    float lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;                          // last_color (the Gaussian's: same for both pixels)
    float2 last_alpha = make_float2(0.f, 0.f);

    const uint32_t wbuf = sb + 2 * SG_BYTES + (uint32_t)warp * WP_BYTES;
    const uint32_t myrow = wbuf + (uint32_t)lane * SL_ROW;
    {
        const uint32_t dpr = wbuf + WP_DP + (uint32_t)lane * DP_ROW;
        sts64<0>(dpr, dp0);
        sts64<8>(dpr, dp1);
        sts64<16>(dpr, dp2);
    }
    const float fx0 = (float)bx0, fy0 = (float)by0;

    // block / warp maxima of the last contributor: nothing behind them has a gradient
    const int wmax = (int)__reduce_max_sync(0xffffffffu, (uint32_t)max(lastc0, lastc1));
    if (lane == 0) s_max[warp] = (uint32_t)wmax;
    __syncthreads();
    const int bmax = max(max((int)s_max[0], (int)s_max[1]), max((int)s_max[2], (int)s_max[3]));
    const int rounds = (bmax + BW_R - 1) / BW_R;

    // ---- staging pipeline: thread t < BW_R owns position first_pos(r) - t of round r ----
    const bool stager = threadIdx.x < BW_R;
    const int t = (int)threadIdx.x;
    auto fetch = [&](int r, uint32_t& id, uint32_t& mk) -> bool {
        const int pos = bmax - 1 - r * BW_R - t;
        const bool ok = stager && pos >= 0;
        if (ok) { id = point_list[range.x + pos]; mk = warp_mask[range.x + pos]; }
        return ok;
    };
    auto issue = [&](int r, bool ok, uint32_t id, uint32_t mk) {
        if (ok) {
            const uint32_t buf = sb + (uint32_t)(r & 1) * SG_BYTES;
            const float4* rec = packed + 3 * (size_t)id;
            cp_async16(buf + t * 16, rec);
            cp_async16(buf + SG_B + t * 16, rec + 1);
            cp_async16(buf + SG_C + t * 16, rec + 2);
            sts32u<SG_ID>(buf + t * 4, id);
            sts8u<SG_MB>(buf + t, mk);
        }
        cp_async_commit();
    };
    uint32_t nid = 0, nmk = 0;
    bool nok = fetch(0, nid, nmk);
    issue(0, nok, nid, nmk);
    nok = fetch(1, nid, nmk);

    int ns = 0;                                                     // filled slots of this warp's batch
    for (int r = 0; r < rounds; r++) {
        cp_async_wait_all();
        __syncthreads();                                            // round r is staged; everyone is done with round r - 1
        issue(r + 1, nok, nid, nmk);                                // copies of round r + 1 fly while round r is replayed
        nok = fetch(r + 2, nid, nmk);
        const uint32_t buf = sb + (uint32_t)(r & 1) * SG_BYTES;
        const int first_pos = bmax - 1 - r * BW_R;                  // position staged by thread 0
        const int count = min(BW_R, first_pos + 1);
        if (first_pos - (count - 1) >= wmax) continue;              // whole batch is behind this warp's last contributor
        for (int cbase = 0; cbase < count; cbase += 32) {
            // lane l tests staged Gaussian cbase + 31 - l: the next one to replay is the highest set ballot bit
            const int jj = cbase + 31 - lane;
            bool hit = false;
            if (jj < count && first_pos - jj < wmax) hit = (lds8u<SG_MB>(buf + (uint32_t)jj) & mbits) != 0u;
            uint32_t m = __ballot_sync(0xffffffffu, hit);
            while (m) {
                const int top = 31 - __clz(m);
                m ^= 1u << top;
                const int j = cbase + 31 - top;
                const int pos = first_pos - j;
                const uint32_t ja = buf + (uint32_t)j * 16;
                // Branch-free replay: the forward's masks guarantee a contributing pixel in the block, so every
                // instruction below issues anyway; pixels that do not contribute run with alpha = G = 0, which
                // leaves T (x 1/(1-0)), the accum_rec recurrence (0*c + 1*rec) and both weights (0) unchanged.
                const float4 a = lds128<0>(ja);           // x, y, conic.x, conic.y
                const float4 b = lds128<SG_B>(ja);        // conic.z, power_thr, opacity, r
                const float2 gb = lds64<SG_C>(ja);        // g, b
                const uint32_t gid = lds32u<SG_ID>(buf + (uint32_t)j * 4);
                // forward.cu:433-437 in the reference binary's contraction, the y-dependent half as packed pairs
                const float dx = subr(a.x, pixfx);
                const float2 dy = add2(bc2(a.y), npixfy);
                const float t1 = mulr(dx, a.z);
                const float un = mulr(dx, -a.w);
                const float2 s2 = fma2(bc2(dx), bc2(t1), mul2(dy, mul2(dy, bc2(b.x))));
                const float2 power = fma2(s2, bc2(-0.5f), mul2(dy, bc2(un)));
                const float2 G0 = make_float2(expf(power.x), expf(power.y));
                float2 al0 = mul2(bc2(b.z), G0);
                al0.x = fminf(0.99f, al0.x);
                al0.y = fminf(0.99f, al0.y);
                // power < power_thr implies alpha < 1/255 (the preprocess writes a conservative threshold), so the
                // forward's three tests collapse into two here
                const bool c0 = (pos < lastc0) && !(power.x > 0.0f) && !(al0.x < 1.0f / 255.0f);
                const bool c1 = (pos < lastc1) && !(power.y > 0.0f) && !(al0.y < 1.0f / 255.0f);
                const float2 alpha = make_float2(c0 ? al0.x : 0.0f, c1 ? al0.y : 0.0f);
                const float2 G = make_float2(c0 ? G0.x : 0.0f, c1 ? G0.y : 0.0f);
                const float2 oma = sub2(bc2(1.0f), alpha);
                const float2 inv = make_float2(rcp_approx_ftz(oma.x), rcp_approx_ftz(oma.y));
                T = mul2(T, inv);
                const float2 dch = mul2(alpha, T);
                const float2 oml = sub2(bc2(1.0f), last_alpha);
                ar0 = fma2(last_alpha, bc2(lc0), mul2(oml, ar0));
                ar1 = fma2(last_alpha, bc2(lc1), mul2(oml, ar1));
                ar2 = fma2(last_alpha, bc2(lc2), mul2(oml, ar2));
                lc0 = b.w; lc1 = gb.x; lc2 = gb.y;
                float2 dLa = mul2(sub2(bc2(lc0), ar0), dp0);
                dLa = fma2(sub2(bc2(lc1), ar1), dp1, dLa);
                dLa = fma2(sub2(bc2(lc2), ar2), dp2, dLa);
                last_alpha = alpha;
                dLa = fma2(inv, nTfbg, mul2(dLa, T));
                const float2 w = mul2(mul2(bc2(b.z), dLa), G);
                sts128<0>(myrow + (uint32_t)ns * 16, make_float4(w.x, w.y, dch.x, dch.y));
                if (lane == 0) sts128<SL_INFO>(wbuf + (uint32_t)ns * SL_ROW, make_float4(a.x, a.y, b.z, __uint_as_float(gid)));
                if (++ns == BW_SLOTS) {
                    flush_slots(wbuf, BW_SLOTS, lane, fx0, fy0, grad_acc);
                    ns = 0;
                }
            }
        }
    }
    if (ns) flush_slots(wbuf, ns, lane, fx0, fy0, grad_acc);
}

void launch_render_backward(int grid_x, int row_begin, int row_end, const uint2* ranges, const uint32_t* point_list,
                            int W, int H, const float4* packed, const float* bg, const float* final_T,
                            const uint32_t* n_contrib, const float* dL_dpix, float* grad_acc, const uint8_t* warp_mask,
                            cudaStream_t st) {
    if (row_end <= row_begin || grid_x <= 0) return;
    static bool configured[GSR_MAX_DEVICES] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < GSR_MAX_DEVICES && !configured[dev]) {
        cudaFuncSetAttribute(render_bwd_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        configured[dev] = true;
    }
    dim3 grid(grid_x, row_end - row_begin, 1);
    render_bwd_kernel<<<grid, BW_T, 0, st>>>(ranges, point_list, W, H, grid_x, row_begin, packed, bg, final_T, n_contrib,
                                              dL_dpix, grad_acc, warp_mask);
    count_launch(1);
}

}  // namespace gsr
