// gsr_render_bwd.cu -- backward of the per-tile alpha blending for sm_100a.
//
// Replaces renderCUDA<3> (backward) (DGR/cuda_rasterizer/backward.cu:408-601), which issues nine
// scalar float atomicAdd per (pixel, Gaussian) pair.  The kernel is bound by warp-instruction issue
// slots (ncu: 83 % issue active, DRAM 4 %), so its design is about instructions per replayed pair:
//
//   * one warp per 8x8 pixel block (four per 16x16 tile), each lane TWO pixels (rows ly and ly + 4 of the
//     block, same column).  Everything that depends only on the Gaussian or on the column (record loads,
//     dx, dx * conic terms, loop control) is paid once per pixel pair, and the per-pixel arithmetic runs
//     as packed pairs (FADD2 / FMUL2 / FFMA2: two IEEE operations per issue slot).
//   * warps share nothing: no block barrier anywhere (a barrier per staging round cost 26 % of the warp
//     time when four warps of a tile shared their staging, ncu r02_ncu_rbwd_v13a).  Each warp walks its
//     tile's list back to front from its OWN last contributor in chunks of 32 positions; which (Gaussian,
//     block) pairs carry a gradient comes from the warp masks the forward recorded (one bit per 8x8 block
//     of the tile), so there is no geometric test and no empty replay, and only records that
//     are replayed are fetched at all.
//   * the walk is software-pipelined three deep: masks of chunk c+3 and ids of the hit positions of chunk
//     c+2 are in flight in registers, the records of chunk c+1 travel by cp.async (LDGSTS) into the other of
//     two per-warp rings, while chunk c is replayed.
//   * Phase 1 (lanes = pixels, sequential over the Gaussians of the block): only what needs the per-pixel
//     recurrences (T, accumulated colour) is computed -- per pixel the two weights
//         w   = G * dL/dalpha                  (all positional / conic / opacity gradients are moments of w)
//         dch = alpha * T                      (dL/dcolour = sum dch * dL/dpixel)
//     and parked in shared memory, one 16-byte store per lane and replay, in a slot of an NS-slot batch.
//   * Phase 2 (lanes = Gaussians, once per NS replays): lane (slot g, part h) walks its share of the
//     block's 64 pixels, accumulates the nine sums of its Gaussian in registers with packed FMAs, the parts
//     are combined with shuffles, and the row leaves as TWO red.global.add.v4.f32 + one scalar red into the
//     (P,12) accumulator the preprocess backward consumes.  This replaces the per-replay warp-shuffle
//     reduce-scatter (47 of 126 instructions per 32-pixel replay before).
// Per-pixel replay order (the T division chain, backward.cu:534) is unchanged; only the cross-pixel
// summation order differs (gradient tolerance 1e-4 relative L2).
#include "gsr_internal.cuh"
#include "gsr_math.cuh"
#include "gsr_smem.cuh"

namespace gsr {

// Shared-memory layout of ONE warp, by template parameter NS (replays per phase-2 batch):
//   two staging rings of up to 32 records (planes a | b | c of the 48-byte records, then the 32 ids),
//   32 rows (one per pixel lane) of NS slots x 16 bytes {w0, w1, dch0, dch1} + 16 bytes (row k < NS: the Gaussian of
//   slot k: x, y, opacity, id), then 32 rows of 24 bytes {dL/dpixel: c0p0 c0p1 c1p0 c1p1 c2p0 c2p1}
template <int NS>
struct BwLayout {
    static constexpr int SG_B = 32 * 16, SG_C = 2 * 32 * 16, SG_ID = 3 * 32 * 16;
    static constexpr int SG_BYTES = SG_ID + 32 * 4;
    static constexpr int SL_ROW = NS * 16 + 16, SL_INFO = NS * 16;
    static constexpr int DP_ROW = 24;
    static constexpr int WP_SLOTS = 2 * SG_BYTES;
    static constexpr int WP_DP = WP_SLOTS + 32 * SL_ROW;
    static constexpr int WP_BYTES = WP_DP + 32 * DP_ROW;
};

// Phase 2, one pixel column of the lane's two rows: raw moments about the column c0 nearest to the Gaussian.  X = column - c0 (exact).
template <int OFF_E, int OFF_D, int COL>
__device__ __forceinline__ void phase2_col(const uint32_t base, const uint32_t dbase, const float c0, float2& S, float2& Mx, float2& Mxx, float2& V6,
                                           float2& V7, float2& V8) {
    const float4 e = lds128<OFF_E>(base);
    const float2 d0 = lds64<OFF_D>(dbase), d1 = lds64<OFF_D + 8>(dbase), d2 = lds64<OFF_D + 16>(dbase);
    const float2 Wp = make_float2(e.x, e.y), Dc = make_float2(e.z, e.w);
    const float X = (float)COL - c0;
    S = add2(S, Wp);
    Mx = fma2(Wp, bc2(X), Mx);
    Mxx = fma2(Wp, bc2(X * X), Mxx);
    V6 = fma2(Dc, d0, V6);
    V7 = fma2(Dc, d1, V7);
    V8 = fma2(Dc, d2, V8);
}

// Phase 2: the nine sums of the n <= NS Gaussians parked in this warp's slots.  Warp-uniform call.  NS = 8:
// lane (slot g = lane % 8, part h = lane / 8) walks the 8 pixel lanes of rows h and h + 4 of the block.  Per pixel only the
// raw moments sum w, sum w x, sum w x^2 about a block column c0 are accumulated (x = column - c0 is one subtraction per column
// and lane, y is constant along a row); the moments about the Gaussian's centre the gradients need,
//     sum w dx = ax S - Mx,   sum w dx^2 = ax^2 S - 2 ax Mx + Mxx,   sum w dy = dy S,   sum w dx dy = dy (ax S - Mx),   sum w dy^2 = dy^2 S
// with ax = x_gaussian - x_c0, follow once per lane.  c0 is the block column nearest to the Gaussian, so the shift does not
// cancel (|ax| <= 1/2 when the centre is over the block; otherwise dx keeps one sign).  6 packed FP32 operations per pixel pair
// instead of 11: the kernel is co-bound by the FP32 pipes.
template <int NS>
__device__ __forceinline__ void flush_slots(const uint32_t wbuf, const int n, const int lane, const float fx0, const float fy0,
                                            float* __restrict__ grad_acc) {
    static_assert(NS == 8, "phase 2 is written for 8 slots: 8 x 4 lanes = (slot, row pair)");
    using L = BwLayout<NS>;
    __syncwarp();
    const int g = lane & 7, h = lane >> 3;
    const float4 info = lds128<L::WP_SLOTS + L::SL_INFO>(wbuf + (uint32_t)g * L::SL_ROW);          // x, y, opacity, id
    const uint32_t base = wbuf + (uint32_t)(L::WP_SLOTS + h * 8 * L::SL_ROW + g * 16);
    const uint32_t dbase = wbuf + (uint32_t)(L::WP_DP + h * 8 * L::DP_ROW);
    // x origin of the raw moments: the block column nearest to the Gaussian's centre.  With the centre inside the block's columns
    // the shift below is at most half a pixel; outside, dx has one sign over the whole block: no cancellation either way.
    const float c0 = rintf(fminf(fmaxf(subr(info.x, fx0), 0.0f), 7.0f));
    float2 S = make_float2(0.f, 0.f), Mx = S, Mxx = S, V6 = S, V7 = S, V8 = S;
    phase2_col<0 * L::SL_ROW, 0 * L::DP_ROW, 0>(base, dbase, c0, S, Mx, Mxx, V6, V7, V8);
    phase2_col<1 * L::SL_ROW, 1 * L::DP_ROW, 1>(base, dbase, c0, S, Mx, Mxx, V6, V7, V8);
    phase2_col<2 * L::SL_ROW, 2 * L::DP_ROW, 2>(base, dbase, c0, S, Mx, Mxx, V6, V7, V8);
    phase2_col<3 * L::SL_ROW, 3 * L::DP_ROW, 3>(base, dbase, c0, S, Mx, Mxx, V6, V7, V8);
    phase2_col<4 * L::SL_ROW, 4 * L::DP_ROW, 4>(base, dbase, c0, S, Mx, Mxx, V6, V7, V8);
    phase2_col<5 * L::SL_ROW, 5 * L::DP_ROW, 5>(base, dbase, c0, S, Mx, Mxx, V6, V7, V8);
    phase2_col<6 * L::SL_ROW, 6 * L::DP_ROW, 6>(base, dbase, c0, S, Mx, Mxx, V6, V7, V8);
    phase2_col<7 * L::SL_ROW, 7 * L::DP_ROW, 7>(base, dbase, c0, S, Mx, Mxx, V6, V7, V8);
    // shift to the Gaussian's centre (dx = x_g - x_pixel = ax - X, dy = y_g - y_row)
    const float ax = subr(info.x, fx0 + c0);
    const float fyh = fy0 + (float)h;
    const float2 dy = make_float2(subr(info.y, fyh), subr(info.y, fyh + 4.0f));
    const float2 sdx = sub2(mul2(bc2(ax), S), Mx);                                   // sum w dx per row
    const float2 sdxx = fma2(bc2(ax), sdx, fma2(bc2(-ax), Mx, Mxx));                 // sum w dx^2 per row
    const float2 sdy = mul2(dy, S), sdxy = mul2(dy, sdx), sdyy = mul2(dy, sdy);
    float v0 = sdx.x + sdx.y, v1 = sdy.x + sdy.y, v2 = sdxx.x + sdxx.y, v3 = sdxy.x + sdxy.y, v4 = sdyy.x + sdyy.y, s = S.x + S.y;
    float v6 = V6.x + V6.y, v7 = V7.x + V7.y, v8 = V8.x + V8.y;
#pragma unroll
    for (int d = 16; d >= 8; d >>= 1) {
        v0 += __shfl_xor_sync(0xffffffffu, v0, d);
        v1 += __shfl_xor_sync(0xffffffffu, v1, d);
        v2 += __shfl_xor_sync(0xffffffffu, v2, d);
        v3 += __shfl_xor_sync(0xffffffffu, v3, d);
        v4 += __shfl_xor_sync(0xffffffffu, v4, d);
        s += __shfl_xor_sync(0xffffffffu, s, d);
        v6 += __shfl_xor_sync(0xffffffffu, v6, d);
        v7 += __shfl_xor_sync(0xffffffffu, v7, d);
        v8 += __shfl_xor_sync(0xffffffffu, v8, d);
    }
    if (lane < n) {        // part 0 lanes of the filled slots
        // phase 1 parks w = G dL/dalpha; the moments carry the opacity factor (backward.cu:568: dL/dG = opacity dL/dalpha),
        // the opacity gradient is the plain sum
        const float op = info.z;
        float* dst = grad_acc + 12 * (size_t)__float_as_uint(info.w);
        red_add_v4(dst, op * v0, op * v1, op * v2, op * v3);
        red_add_v4(dst + 4, op * v4, s, v6, v7);
        red_add_f32(dst + 8, v8);
    }
    __syncwarp();          // the slots are free again
}

// One warp = one 8x8 pixel block of a tile; warps share nothing (no block barrier anywhere), so a block is just WPC
// consecutive pixel blocks.  Per warp, the tile's list is walked back to front from the warp's own last contributor in
// chunks of 32 positions, software-pipelined three deep: masks of chunk c+3, ids of the hit positions of chunk c+2 and
// the records of chunk c+1 (cp.async into the other ring) are in flight while chunk c is replayed.
template <int NS, int WPC, int MINB>
__global__ void __launch_bounds__(32 * WPC, MINB) render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                                                                   int W, int H, int grid_x, int row_begin, int n_blocks,
                                                                   const float4* __restrict__ packed, const float* __restrict__ bg,
                                                                   const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                                                                   const float* __restrict__ dL_dpixels, float* __restrict__ grad_acc,
                                                                   const uint8_t* __restrict__ warp_mask) {
    using L = BwLayout<NS>;
    constexpr int SG_B = L::SG_B, SG_C = L::SG_C, SG_ID = L::SG_ID, SG_BYTES = L::SG_BYTES;
    constexpr int SL_ROW = L::SL_ROW, SL_INFO = L::SL_INFO, DP_ROW = L::DP_ROW, WP_SLOTS = L::WP_SLOTS, WP_DP = L::WP_DP, WP_BYTES = L::WP_BYTES;
    GSR_DYN_SMEM(s_dyn);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int gw = blockIdx.x * WPC + warp;                         // pixel block: 4 per tile, tiles row-major inside the band
    if (gw >= n_blocks) return;
    const uint32_t wbuf = smem_addr(s_dyn) + (uint32_t)warp * WP_BYTES;

    const int tile = gw >> 2, sub = gw & 3;
    const int tile_y = tile / grid_x + row_begin, tile_x = tile - (tile / grid_x) * grid_x;
    const int bxi = sub & 1, byi = sub >> 1;
    const int bx0 = tile_x * GSR_TILE_X + bxi * 8;
    const int by0 = tile_y * GSR_TILE_Y + byi * 8;
    const int pix_x = bx0 + (lane & 7), pix_y0 = by0 + (lane >> 3), pix_y1 = pix_y0 + 4;
    const bool in0 = pix_x < W && pix_y0 < H, in1 = pix_x < W && pix_y1 < H;
    const float pixfx = (float)pix_x;
    const float2 npixfy = make_float2(-(float)pix_y0, -(float)pix_y1);         // a.y + (-py) == a.y - py
    const size_t id0 = (size_t)pix_y0 * W + pix_x, id1 = (size_t)pix_y1 * W + pix_x;
    const size_t HW = (size_t)H * W;
    // the forward records one bit per 8x8 block of the tile (its warp w' covers columns 8 (w' & 1), rows 8 (w' >> 1))
    const uint32_t mbits = 1u << sub;

    const uint32_t list0 = ranges[tile_y * grid_x + tile_x].x;

    const float Tf0 = in0 ? final_Ts[id0] : 0.0f, Tf1 = in1 ? final_Ts[id1] : 0.0f;
    float2 T = make_float2(Tf0, Tf1);
    const int lastc0 = in0 ? (int)n_contrib[id0] : 0, lastc1 = in1 ? (int)n_contrib[id1] : 0;
    float2 dp0 = make_float2(0.f, 0.f), dp1 = dp0, dp2 = dp0;
    if (in0) { dp0.x = dL_dpixels[id0]; dp1.x = dL_dpixels[HW + id0]; dp2.x = dL_dpixels[2 * HW + id0]; }
    if (in1) { dp0.y = dL_dpixels[id1]; dp1.y = dL_dpixels[HW + id1]; dp2.y = dL_dpixels[2 * HW + id1]; }
    const float bg0 = __ldg(bg), bg1 = __ldg(bg + 1), bg2 = __ldg(bg + 2);
    // The background is the colour "behind everything": the reference's separate term -T_final / (1 - alpha) * (bg . dL/dpixel)
    // of dL/dalpha (backward.cu:560-563) obeys the same recurrence as accum_rec, so it is folded into A's START value
    // (B_j = -T_final (bg . dL/dpix) / T_after(j) satisfies B_j = (1 - alpha_{j+1}) B_{j+1}; A - B starts at bg . dL/dpix).
    float2 A = make_float2(bg0 * dp0.x + bg1 * dp1.x + bg2 * dp2.x, bg0 * dp0.y + bg1 * dp1.y + bg2 * dp2.y);   // (accum_rec . dL/dpixel), background included
    float2 Dl = make_float2(0.f, 0.f);                              // last_color . dL/dpixel
    float2 last_alpha = make_float2(0.f, 0.f), last_oma = make_float2(1.f, 1.f);

    const uint32_t myrow = wbuf + WP_SLOTS + (uint32_t)lane * SL_ROW;
    {
        const uint32_t dpr = wbuf + WP_DP + (uint32_t)lane * DP_ROW;
        sts64<0>(dpr, dp0);
        sts64<8>(dpr, dp1);
        sts64<16>(dpr, dp2);
    }
    const float fx0 = (float)bx0, fy0 = (float)by0;

    // nothing behind the block's last contributor has a gradient
    const int wmax = (int)__reduce_max_sync(0xffffffffu, (uint32_t)max(lastc0, lastc1));
    const int nchunks = (wmax + 31) >> 5;
    // chunk c covers positions top(c) - 31 .. top(c), lane l <-> position top(c) - 31 + l: the next Gaussian to replay is the
    // highest set bit of the chunk's hit mask
    auto load_mask = [&](int c) -> uint32_t {
        const int p = wmax - 32 - 32 * c + lane;
        return (c < nchunks && p >= 0) ? (uint32_t)warp_mask[list0 + p] : 0u;
    };
    auto load_id = [&](int c, bool hit) -> uint32_t {
        return hit ? point_list[list0 + (wmax - 32 - 32 * c + lane)] : 0u;
    };
    auto issue = [&](int c, uint32_t hits, uint32_t id) {          // records of the hit positions of chunk c -> ring c & 1, in replay order
        if ((hits >> lane) & 1u) {
            const uint32_t k = (uint32_t)__popc((hits >> lane) >> 1);
            const uint32_t stg = wbuf + (uint32_t)(c & 1) * SG_BYTES;
            const float4* rec = packed + 3 * (size_t)id;
            cp_async16(stg + k * 16, rec);
            cp_async16(stg + SG_B + k * 16, rec + 1);
            cp_async16(stg + SG_C + k * 16, rec + 2);
            sts32u<SG_ID>(stg + k * 4, id);
        }
        cp_async_commit();
    };
    uint32_t hb0, hb1, nid1, mk2;
    {
        const uint32_t mkA = load_mask(0), mkB = load_mask(1);
        mk2 = load_mask(2);
        hb0 = __ballot_sync(0xffffffffu, (mkA & mbits) != 0u);
        const uint32_t idA = load_id(0, (mkA & mbits) != 0u);
        hb1 = __ballot_sync(0xffffffffu, (mkB & mbits) != 0u);
        nid1 = load_id(1, (mkB & mbits) != 0u);
        issue(0, hb0, idA);
    }

    int ns = 0;                                                     // filled slots of this warp's batch
    for (int c = 0; c < nchunks; c++) {
        issue(c + 1, hb1, nid1);                                    // copies of chunk c + 1 fly while chunk c is replayed
        const bool hit2 = (mk2 & mbits) != 0u;
        const uint32_t hb2 = __ballot_sync(0xffffffffu, hit2);
        const uint32_t nid2 = load_id(c + 2, hit2);
        mk2 = load_mask(c + 3);
        cp_async_wait_1();
        __syncwarp();                                               // chunk c's records (every lane's copies) are in ring c & 1
        const uint32_t stg = wbuf + (uint32_t)(c & 1) * SG_BYTES;
        const int pos_base = wmax - 32 - 32 * c;
        uint32_t m = hb0;
        uint32_t ja = stg;
        while (m) {
            const int top = 31 - __clz(m);
            m ^= 1u << top;
            const int pos = pos_base + top;
            // Branch-free replay: the forward's masks guarantee a contributing pixel in the block, so every
            // instruction below issues anyway; pixels that do not contribute run with alpha = G = 0, which
            // leaves T (x 1/(1-0)), the accum_rec recurrence (0*c + 1*rec) and both weights (0) unchanged.
            const float4 a = lds128<0>(ja);           // x, y, conic.x, conic.y
            const float4 b = lds128<SG_B>(ja);        // conic.z, power_thr, opacity, r
            const float2 gb = lds64<SG_C>(ja);        // g, b
            const uint32_t gid = lds32u<SG_ID>(stg + ((ja - stg) >> 2));
            ja += 16;
            // forward.cu:433-437 in the reference binary's contraction, the y-dependent half as packed pairs
            const float dx = subr(a.x, pixfx);
            const float2 dy = add2(bc2(a.y), npixfy);
            const float t1 = mulr(dx, a.z);
            const float un = mulr(dx, -a.w);
            const float2 s2 = fma2(bc2(dx), bc2(t1), mul2(dy, mul2(dy, bc2(b.x))));
            const float2 power = fma2(s2, bc2(-0.5f), mul2(dy, bc2(un)));
            const float2 G0 = expf2(power);
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
            // backward.cu:545-556 with the three colour channels collapsed BEFORE the recurrence: the reference keeps
            // accum_rec[c] (colour accumulated behind this Gaussian) per channel and forms sum_c (c_c - accum_rec[c]) dL/dpix_c;
            // the recurrence is linear, so A = sum_c accum_rec[c] dL/dpix_c obeys the same one with D = colour . dL/dpix in
            // place of the colour: 8 FP32 operations per pixel instead of 14 (the kernel is co-bound by the FP32 pipes).
            A = fma2(last_alpha, Dl, mul2(last_oma, A));
            Dl = fma2(bc2(gb.y), dp2, fma2(bc2(gb.x), dp1, mul2(bc2(b.w), dp0)));
            last_alpha = alpha;
            last_oma = oma;
            const float2 dLa = mul2(sub2(Dl, A), T);
            // w = G dL/dalpha WITHOUT the opacity factor: phase 2 scales the five moments by it once per Gaussian
            const float2 w = mul2(dLa, G);
            sts128<0>(myrow + (uint32_t)ns * 16, make_float4(w.x, w.y, dch.x, dch.y));
            if (lane == 0) sts128<WP_SLOTS + SL_INFO>(wbuf + (uint32_t)ns * SL_ROW, make_float4(a.x, a.y, b.z, __uint_as_float(gid)));
            if (++ns == NS) {
                flush_slots<NS>(wbuf, NS, lane, fx0, fy0, grad_acc);
                ns = 0;
            }
        }
        __syncwarp();                                               // ring c & 1 is free for chunk c + 2
        hb0 = hb1; hb1 = hb2; nid1 = nid2;
    }
    if (ns) flush_slots<NS>(wbuf, ns, lane, fx0, fy0, grad_acc);
}

template <int NS, int WPC, int MINB>
static void launch_variant(int n_blocks, cudaStream_t st, int dev, const uint2* ranges, const uint32_t* point_list, int W, int H, int grid_x,
                           int row_begin, const float4* packed, const float* bg, const float* final_T, const uint32_t* n_contrib,
                           const float* dL_dpix, float* grad_acc, const uint8_t* warp_mask) {
    constexpr int SMEM = BwLayout<NS>::WP_BYTES * WPC;
    static bool configured[GSR_MAX_DEVICES] = {};
    if (dev >= 0 && dev < GSR_MAX_DEVICES && !configured[dev]) {     // as much of the SM's memory as shared memory as the blocks need
        cudaFuncSetAttribute(render_bwd_kernel<NS, WPC, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        cudaFuncSetAttribute(render_bwd_kernel<NS, WPC, MINB>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        configured[dev] = true;
    }
    render_bwd_kernel<NS, WPC, MINB><<<(n_blocks + WPC - 1) / WPC, 32 * WPC, SMEM, st>>>(ranges, point_list, W, H, grid_x, row_begin, n_blocks, packed, bg,
                                                                                        final_T, n_contrib, dL_dpix, grad_acc, warp_mask);
}

void launch_render_backward(int grid_x, int row_begin, int row_end, const uint2* ranges, const uint32_t* point_list,
                            int W, int H, const float4* packed, const float* bg, const float* final_T,
                            const uint32_t* n_contrib, const float* dL_dpix, float* grad_acc, const uint8_t* warp_mask,
                            cudaStream_t st) {
    if (row_end <= row_begin || grid_x <= 0) return;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) dev = -1;
    const int n_blocks = 4 * grid_x * (row_end - row_begin);         // 8x8 pixel blocks in the band
    // GSR_RB_VARIANT (measurement only): other batch / block-size / occupancy choices of the same kernel
    static const int variant = [] { const char* e = getenv("GSR_RB_VARIANT"); return e ? atoi(e) : 0; }();
#define GSR_RB_ARGS n_blocks, st, dev, ranges, point_list, W, H, grid_x, row_begin, packed, bg, final_T, n_contrib, dL_dpix, grad_acc, warp_mask
    switch (variant) {
        case 1: launch_variant<8, 4, 6>(GSR_RB_ARGS); break;
        case 2: launch_variant<8, 4, 7>(GSR_RB_ARGS); break;
        case 3: launch_variant<8, 1, 20>(GSR_RB_ARGS); break;
        case 4: launch_variant<8, 1, 28>(GSR_RB_ARGS); break;
        case 5: launch_variant<8, 2, 12>(GSR_RB_ARGS); break;
        case 6: launch_variant<8, 1, 32>(GSR_RB_ARGS); break;
        default: launch_variant<8, 1, 24>(GSR_RB_ARGS); break;      // measured best at C2 (0.809 ms), see profiles/r02_sweep_rb_*.log
    }
#undef GSR_RB_ARGS
    count_launch(1);
}

}  // namespace gsr
