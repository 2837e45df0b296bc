// gsr_binning.cu -- scan, stable radix sort, instance duplication and tile ranges for sm_100a.
//
// Replaces cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:323-325), duplicateWithKeys
// (:78-126), cub::DeviceRadixSort::SortPairs on (tile<<32 | depth_bits) (:357-374) and
// identifyTileRanges (:133-159) of DGR/cuda_rasterizer/rasterizer_impl.cu.
//
// The reference sorts R = num_rendered 12-byte pairs on 32+bit key bits (6 passes at 1080p).
// The same total order -- ascending (tile, depth bits), ties in ascending Gaussian index -- is
// produced here in two cheaper steps, both stable LSD radix sorts on 32-bit keys:
//   (1) sort the P Gaussians by depth bits (4 passes over P 8-byte pairs);
//   (2) emit instances in that order and sort them by tile id only (2 passes over R 8-byte
//       pairs for <= 65536 tiles).
// Stability of (2) keeps depth order (and index order among equal depths) inside each tile.
// All of it is HBM-bound integer work: coalesced key loads, warp-level match_any ranking in
// shared memory, no tensor cores.
#include <cstdlib>
#include "gsr_internal.cuh"
#include "gsr_smem.cuh"
#include "gsr_math.cuh"

namespace gsr {

// ------------------------------------------------------------------------------------------------
// exclusive scan (three small kernels; sizes here are a few MB, launch-latency bound)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}

// block-wide exclusive scan of one value per thread; returns exclusive prefix, *total = block sum
template <int THREADS>
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* s_warp /* THREADS/32 + 1 */, uint32_t* total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t inc = warp_incl_scan(v, lane);
    if (lane == 31) s_warp[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint32_t x = (lane < THREADS / 32) ? s_warp[lane] : 0u;
        const uint32_t xi = warp_incl_scan(x, lane);
        if (lane < THREADS / 32) s_warp[lane] = xi - x;
        if (lane == 31) s_warp[THREADS / 32] = xi;
    }
    __syncthreads();
    const uint32_t r = s_warp[w] + inc - v;
    *total = s_warp[THREADS / 32];
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_block_sums_kernel(const uint32_t* __restrict__ in, size_t n, uint32_t* __restrict__ sums) {
    __shared__ uint32_t s_warp[SCAN_THREADS / 32];
    const size_t base = (size_t)blockIdx.x * SCAN_TILE;
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const size_t i = base + (size_t)k * SCAN_THREADS + threadIdx.x;
        if (i < n) acc += in[i];
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t x = (threadIdx.x < SCAN_THREADS / 32) ? s_warp[threadIdx.x] : 0u;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) x += __shfl_xor_sync(0xffffffffu, x, d);
        if (threadIdx.x == 0) sums[blockIdx.x] = x;
    }
}

// one block: exclusive scan of sums[0..nb) in place, sums[nb] = grand total
__global__ void __launch_bounds__(1024) scan_sums_kernel(uint32_t* __restrict__ sums, size_t nb, uint32_t* __restrict__ total_out) {
    __shared__ uint32_t s_warp[33];
    uint32_t carry = 0;
    for (size_t base = 0; base < nb; base += 1024) {
        const size_t i = base + threadIdx.x;
        const uint32_t v = (i < nb) ? sums[i] : 0u;
        uint32_t tot;
        const uint32_t ex = block_excl_scan<1024>(v, s_warp, &tot);
        if (i < nb) sums[i] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0) {
        sums[nb] = carry;
        if (total_out) *total_out = carry;
    }
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, size_t n,
                                                                  const uint32_t* __restrict__ sums) {
    __shared__ uint32_t s_warp[SCAN_THREADS / 32 + 1];
    const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    const bool full = ((size_t)blockIdx.x + 1) * SCAN_TILE <= n;
    if (full) {
        const uint4* p = reinterpret_cast<const uint4*>(in + base);
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS / 4; k++) {
            const uint4 q = p[k];
            v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
        }
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) v[k] = (base + k < n) ? in[base + k] : 0u;
    }
    uint32_t tsum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) tsum += v[k];
    uint32_t tot;
    uint32_t run = block_excl_scan<SCAN_THREADS>(tsum, s_warp, &tot) + sums[blockIdx.x];
    uint32_t o[SCAN_ITEMS];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) { o[k] = run; run += v[k]; }
    if (full) {
        uint4* p = reinterpret_cast<uint4*>(out + base);
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS / 4; k++) p[k] = make_uint4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]);
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) if (base + k < n) out[base + k] = o[k];
    }
}

void scan_exclusive_u32(const uint32_t* in, uint32_t* out, size_t n, uint32_t* tmp, uint32_t* total, cudaStream_t st) {
    if (n == 0) {
        if (total) cudaMemsetAsync(total, 0, sizeof(uint32_t), st);
        return;
    }
    const size_t nb = scan_blocks(n);
    scan_block_sums_kernel<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, n, tmp);
    scan_sums_kernel<<<1, 1024, 0, st>>>(tmp, nb, total);
    scan_apply_kernel<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, out, n, tmp);
    count_launch(3);
}

// ------------------------------------------------------------------------------------------------
// stable LSD radix sort pass on (u32 key, u32 value) pairs, <= 8 key bits per pass
// ------------------------------------------------------------------------------------------------
// One pass = a histogram kernel (global digit counts) + ONE single-sweep kernel: each block takes a ticket
// (blocks therefore start in tile order), ranks its 4096 items, publishes its per-digit counts and obtains
// the counts of all earlier tiles with a decoupled look-back over a status-tagged table, reorders its
// items by digit in shared memory and writes them out as contiguous runs.  Per item the pass moves
// 4 B (histogram read) + 8 B read + 8 B write.  Deterministic and stable: a tile's items keep their input
// order within each digit, tiles are concatenated in order.
//
// Ranking: warp w owns the w-th 512-item slice of the tile and walks it in 16 rounds of 32 consecutive
// items, so "earlier in the input" == (lower warp, lower round, lower lane).  Within a round lanes with
// equal digits find each other with match.any; the lowest such lane bumps the warp's private digit counter.
constexpr uint32_t LB_AGG = 1u << 30, LB_PREFIX = 2u << 30, LB_MASK = (1u << 30) - 1u;

struct RsPasses { int n; int shift[4]; int bits[4]; };

// Digit histograms of ALL passes of one sort in a single read of the keys (the key set does not change between
// passes, only its order).  ghist[p*256 + d] += count of keys whose pass-p digit is d.
template <int NP>
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const uint32_t* __restrict__ keys, size_t n, RsPasses ps,
                                                             uint32_t* __restrict__ ghist) {
    __shared__ uint32_t s_hist[8 * NP * RS_BINS];          // [8 warps][NP][RS_BINS]
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t* wh = s_hist + w * NP * RS_BINS;
    for (int i = lane; i < NP * RS_BINS; i += 32) wh[i] = 0;
    __syncwarp();
    const size_t n4 = n >> 2;
    const uint4* k4 = reinterpret_cast<const uint4*>(keys);
    // warp-uniform trip count (the loop body votes): lanes past the end carry an impossible digit
    for (size_t base = (size_t)blockIdx.x * RS_THREADS + (size_t)w * 32; base < n4; base += (size_t)gridDim.x * RS_THREADS) {
        const size_t i = base + lane;
        const bool ok = i < n4;
        const uint4 q = ok ? __ldg(k4 + i) : make_uint4(0, 0, 0, 0);
        const uint32_t kk[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int p = 0; p < NP; p++) {
            const uint32_t mask = (1u << ps.bits[p]) - 1u;
            uint32_t* hp = wh + p * RS_BINS;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                // whole warp on one digit (high bytes of depth keys): one add instead of a 32-way same-address atomic
                const uint32_t dj = ok ? ((kk[j] >> ps.shift[p]) & mask) : 0xFFFFFFFFu;
                const uint32_t d0 = __shfl_sync(0xffffffffu, dj, 0);
                if (__all_sync(0xffffffffu, dj == d0)) { if (lane == 0) hp[d0] += 32; }
                else if (ok) atomicAdd(&hp[dj], 1u);
                __syncwarp();
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        const uint32_t key = keys[4 * n4 + threadIdx.x];
#pragma unroll
        for (int p = 0; p < NP; p++) atomicAdd(&wh[p * RS_BINS + ((key >> ps.shift[p]) & ((1u << ps.bits[p]) - 1u))], 1u);
    }
    __syncthreads();
    const int d = threadIdx.x;   // RS_THREADS == RS_BINS
#pragma unroll
    for (int p = 0; p < NP; p++) {
        uint32_t c = 0;
#pragma unroll
        for (int ww = 0; ww < 8; ww++) c += s_hist[(ww * NP + p) * RS_BINS + d];
        if (c) atomicAdd(&ghist[p * RS_BINS + d], c);
    }
}

#ifndef RS_MINB
#define RS_MINB 5
#endif
template <bool BALLOT, int ITEMS>
__global__ void __launch_bounds__(RS_THREADS, RS_MINB) rs_sweep_kernel(const uint32_t* __restrict__ kin, const uint32_t* __restrict__ vin,
                                                              uint32_t* __restrict__ kout, uint32_t* __restrict__ vout, size_t n,
                                                              int shift, int bits, const uint32_t* __restrict__ ghist,
                                                              uint32_t* ticket, uint32_t* lookback) {
    __shared__ uint32_t s_hist[8 * RS_BINS];
    __shared__ uint32_t s_keys[(RS_THREADS * ITEMS)];
    __shared__ uint32_t s_vals[(RS_THREADS * ITEMS)];
    __shared__ uint32_t s_dstart[RS_BINS];     // first local (digit-sorted) position of each digit
    __shared__ uint32_t s_gbase[RS_BINS];      // global position of local position 0 of each digit's run
    __shared__ uint32_t s_warp[RS_THREADS / 32 + 1];
    __shared__ uint32_t s_bid;
    if (threadIdx.x == 0) s_bid = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t bid = s_bid;
    const uint32_t mask = (1u << bits) - 1u;
    const int radix = 1 << bits;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const size_t tile_base = (size_t)bid * (RS_THREADS * ITEMS);
    const int tile_n = (int)min((size_t)(RS_THREADS * ITEMS), n - tile_base);

    // ---- rank ----
    uint32_t* wh = s_hist + w * RS_BINS;
    for (int i = lane; i < RS_BINS; i += 32) wh[i] = 0;
    const int wofs = w * (32 * ITEMS);
    uint32_t key[ITEMS];
    uint16_t rank[ITEMS];
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        const int li = wofs + r * 32 + lane;
        key[r] = (li < tile_n) ? kin[tile_base + li] : 0xFFFFFFFFu;
    }
    __syncwarp();
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        const bool valid = wofs + r * 32 + lane < tile_n;
        const uint32_t d = valid ? ((key[r] >> shift) & mask) : 0xFFFFFFFFu;
        uint32_t peers;
        if (BALLOT) {
            peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll 8
            for (int b = 0; b < bits; b++) {
                const bool one = (d >> b) & 1u;
                const uint32_t mb = __ballot_sync(0xffffffffu, one);
                peers &= one ? mb : ~mb;
            }
        } else {
            peers = __match_any_sync(0xffffffffu, d);
        }
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) { old = wh[d]; wh[d] = old + __popc(peers); }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[r] = (uint16_t)(old + __popc(peers & lt));
    }
    __syncthreads();

    // ---- per digit: block count, exclusive prefix over the 8 warps, look-back over earlier tiles ----
    const int d = threadIdx.x;
    uint32_t cnt = 0;
#pragma unroll
    for (int ww = 0; ww < 8; ww++) {
        const uint32_t c = s_hist[ww * RS_BINS + d];
        s_hist[ww * RS_BINS + d] = cnt;
        cnt += c;
    }
    volatile uint32_t* lb = lookback;
    if (d < radix) lb[(size_t)bid * RS_BINS + d] = (bid == 0 ? LB_PREFIX : LB_AGG) | cnt;
    uint32_t tot;
    const uint32_t dstart = block_excl_scan<RS_THREADS>(cnt, s_warp, &tot);
    const uint32_t gh = (d < radix) ? ghist[d] : 0u;
    const uint32_t gstart = block_excl_scan<RS_THREADS>(gh, s_warp, &tot);
    uint32_t ex = 0;
    if (d < radix && bid > 0) {
        int b = (int)bid - 1;
        while (true) {
            const uint32_t v = lb[(size_t)b * RS_BINS + d];
            const uint32_t flag = v & ~LB_MASK;
            if (flag == 0) continue;                 // not published yet: earlier tickets are resident, spin
            ex += v & LB_MASK;
            if (flag == LB_PREFIX) break;
            b--;
        }
        lb[(size_t)bid * RS_BINS + d] = LB_PREFIX | (ex + cnt);
    }
    s_dstart[d] = dstart;
    s_gbase[d] = gstart + ex - dstart;
    __syncthreads();

    // ---- reorder by digit in shared memory ----
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        const int li = wofs + r * 32 + lane;
        if (li < tile_n) {
            const uint32_t dg = (key[r] >> shift) & mask;
            const uint32_t lp = s_dstart[dg] + wh[dg] + rank[r];
            s_keys[lp] = key[r];
            s_vals[lp] = vin[tile_base + li];
        }
    }
    __syncthreads();
    // ---- contiguous runs out ----
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        const int j = k * RS_THREADS + threadIdx.x;
        if (j < tile_n) {
            const uint32_t kk = s_keys[j];
            const uint32_t pos = s_gbase[(kk >> shift) & mask] + (uint32_t)j;
            kout[pos] = kk;
            vout[pos] = s_vals[j];
        }
    }
}

// Stable LSD sort of n (key, value) pairs on the given digit passes; input in keys[s0]/vals[s0], the result ends
// in keys[s0 ^ (npass & 1)].  `table` (rs_table_words(n) words) holds the digit histograms of all passes
// [p*256, p*256+256), the ticket counters [1024 + p] and one look-back table per pass from word 1280.
void radix_sort_u32(uint32_t* const keys[2], uint32_t* const vals[2], int s0, size_t n, int npass, const int* shifts,
                    const int* bits, uint32_t* table, cudaStream_t st) {
    if (n == 0 || npass <= 0) return;
    const unsigned nb = (unsigned)rs_blocks(n);
    cudaMemsetAsync(table, 0, (1280 + (size_t)npass * nb * RS_BINS) * sizeof(uint32_t), st);
    RsPasses ps;
    ps.n = npass;
    for (int p = 0; p < 4; p++) { ps.shift[p] = p < npass ? shifts[p] : 0; ps.bits[p] = p < npass ? bits[p] : 1; }
    const unsigned hb = (unsigned)min((size_t)(148 * 8), (n / 4 + RS_THREADS - 1) / RS_THREADS + 1);
    switch (npass) {
        case 1: rs_hist_kernel<1><<<hb, RS_THREADS, 0, st>>>(keys[s0], n, ps, table); break;
        case 2: rs_hist_kernel<2><<<hb, RS_THREADS, 0, st>>>(keys[s0], n, ps, table); break;
        case 3: rs_hist_kernel<3><<<hb, RS_THREADS, 0, st>>>(keys[s0], n, ps, table); break;
        default: rs_hist_kernel<4><<<hb, RS_THREADS, 0, st>>>(keys[s0], n, ps, table); break;
    }
    int s = s0;
    for (int p = 0; p < npass; p++) {
        // one ballot per digit bit beats match.any (one trip per distinct value among the 32 lanes); measured on B200 for the
        // 6- and 7-bit tile digits (0.30 -> 0.25 ms) and for the 8-bit depth digits (0.1765 -> 0.1667 ms)
        // (tiles of 8 or 12 items per thread measured slower than 16 for both sorts)
        if (bits[p] <= 8)
            rs_sweep_kernel<true, RS_ITEMS><<<nb, RS_THREADS, 0, st>>>(keys[s], vals[s], keys[s ^ 1], vals[s ^ 1], n, shifts[p], bits[p],
                                                                       table + p * RS_BINS, table + 1024 + p, table + 1280 + (size_t)p * nb * RS_BINS);
        else
            rs_sweep_kernel<false, RS_ITEMS><<<nb, RS_THREADS, 0, st>>>(keys[s], vals[s], keys[s ^ 1], vals[s ^ 1], n, shifts[p], bits[p],
                                                                        table + p * RS_BINS, table + 1024 + p, table + 1280 + (size_t)p * nb * RS_BINS);
        s ^= 1;
    }
    count_launch(1 + npass);
}


// ------------------------------------------------------------------------------------------------
// Counting tile sort.  The instance list leaves duplicate in depth order; what the blending kernels need is that list
// transposed -- grouped by tile, depth order kept inside each tile (the reference gets it from the 64-bit radix sort,
// rasterizer_impl.cu:366-389).  With at most TC_MAX_TILES tiles the write cursor of EVERY tile fits in shared memory, so one
// pass does it:  (1) each chunk of the list counts its instances per tile (tc_count_kernel), (2) a column scan over the
// chunks (tc_colscan_kernel) and a scan over the tiles (prologue of tc_scatter_kernel) turn the counts into each chunk's
// first slot in every tile's run -- and into the tile ranges, which need no pass over sorted keys any more --, (3) each
// chunk walks its instances in order and hands out slots from its cursors (tc_scatter_kernel).
// Slots of one tile must be handed out in list order: inside 32 instances by lane order (fixed up in the rare case that two
// of them share a tile), between the warps of a block by taking turns, between blocks by the bases from (2).
// Measured on B200 at C2 (R = 12.0 M, 8160 tiles): 0.153 ms against 0.274 ms for the two radix passes + the ranges pass.
// What bounds it now: 12 M scattered 4-byte stores (one L2 transaction each); the turn-taking costs 0.012 ms of it.
__global__ void __launch_bounds__(TC_THREADS) tc_count_kernel(const uint32_t* __restrict__ keys, size_t R, size_t chunk, int Tpad,
                                                              uint32_t* __restrict__ table) {
    GSR_DYN_SMEM(s_dyn);
    uint32_t* s_cnt = reinterpret_cast<uint32_t*>(s_dyn);
    for (int t = threadIdx.x; t < Tpad; t += TC_THREADS) s_cnt[t] = 0;
    __syncthreads();
    const size_t b = (size_t)blockIdx.x * chunk, e = min(R, b + chunk);
    // chunk is a multiple of TC_ROUND and the key array is 16-byte aligned: whole 16-byte words up to the last few keys
    const size_t n4 = (e - b) >> 2;
    const uint4* k4 = reinterpret_cast<const uint4*>(keys + b);
#pragma unroll 4
    for (size_t i = threadIdx.x; i < n4; i += TC_THREADS) {               // unrolled: several 16-byte loads in flight per thread
        const uint4 q = __ldg(k4 + i);
        atomicAdd(&s_cnt[q.x], 1u); atomicAdd(&s_cnt[q.y], 1u); atomicAdd(&s_cnt[q.z], 1u); atomicAdd(&s_cnt[q.w], 1u);
    }
    for (size_t i = b + 4 * n4 + threadIdx.x; i < e; i += TC_THREADS) atomicAdd(&s_cnt[keys[i]], 1u);
    __syncthreads();
    uint32_t* row = table + TC_ROWS_AT + (size_t)blockIdx.x * Tpad;
    for (int t = threadIdx.x; t < Tpad; t += TC_THREADS) row[t] = s_cnt[t];
}

// counts[chunk][tile] -> exclusive prefix over the chunks (in place) and the per-tile totals.  One block per 32 tiles;
// 32 segments of chunks per block, each thread walks its segment twice (the second time out of L1 / L2).
__global__ void __launch_bounds__(1024) tc_colscan_kernel(uint32_t* __restrict__ table, int C, int Tpad) {
    __shared__ uint32_t s_part[32][33];
    const int lane = threadIdx.x & 31, seg = threadIdx.x >> 5;
    const int bin = blockIdx.x * 32 + lane;
    const int Cs = (C + 31) / 32;
    const int c0 = min(C, seg * Cs), c1 = min(C, c0 + Cs);
    uint32_t* col = table + TC_ROWS_AT + bin;
    uint32_t sum = 0;
#pragma unroll 4
    for (int c = c0; c < c1; c++) sum += col[(size_t)c * Tpad];
    s_part[seg][lane] = sum;
    __syncthreads();
    uint32_t run = 0, total = 0;
    for (int sgm = 0; sgm < 32; sgm++) {
        const uint32_t v = s_part[sgm][lane];
        if (sgm < seg) run += v;
        total += v;
    }
#pragma unroll 4
    for (int c = c0; c < c1; c++) {
        const uint32_t v = col[(size_t)c * Tpad];
        col[(size_t)c * Tpad] = run;
        run += v;
    }
    if (seg == 0) table[bin] = total;
}

// Each warp owns EPT * 32 consecutive instances of a round.  The warps take turns, in list order, to advance the cursors of
// their tiles: a token passed round the block's 8 warps over named barriers (a waiting warp sleeps, no polling).  Only the
// cursor updates -- one shared-memory atomic and one read-back per lane and 32 instances, performed in order -- are inside
// the turn; loads, the rare fix-up and the stores overlap with the other warps' turns and with the other blocks of the SM.
template <int EPT>
__global__ void __launch_bounds__(TC_THREADS, 3) tc_scatter_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                                                                              size_t R, size_t chunk, int T, int Tpad, int tile_bits,
                                                                              const uint32_t* __restrict__ table, uint32_t* __restrict__ out, uint2* __restrict__ ranges) {
    GSR_DYN_SMEM(s_dyn);
    uint32_t* s_ctr = reinterpret_cast<uint32_t*>(s_dyn);                 // next free slot of every tile, for this chunk
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    constexpr int NW = TC_THREADS / 32;
    {
        // first slot of every tile = exclusive scan of the per-tile totals (every block redoes it: 32 KB out of L2, a few
        // microseconds, instead of one more launch), plus this chunk's base inside the tile's run.  Block 0 also writes the
        // tile ranges: tiles without instances keep (0, 0) as in the reference (ranges are zeroed and only tiles that occur
        // in the list are written, rasterizer_impl.cu:133-159,377).
        __shared__ uint32_t s_wsum[NW];
        const int Tw = ((T + NW - 1) / NW + 31) & ~31;                     // tiles per warp, whole 32s
        const int t0 = w * Tw, t1 = min(T, t0 + Tw);
#pragma unroll 8
        for (int t = tid; t < T; t += TC_THREADS) s_ctr[t] = table[t];     // independent loads first, the scan runs out of shared memory
        __syncthreads();
        uint32_t carry = 0;
        for (int t = t0 + lane; t < t0 + Tw; t += 32) {
            const uint32_t x = t < t1 ? s_ctr[t] : 0u;
            const uint32_t inc = warp_incl_scan(x, lane);
            if (t < t1) s_ctr[t] = carry + inc - x;
            carry += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane == 0) s_wsum[w] = carry;
        __syncthreads();
        uint32_t before = 0;
        for (int ww = 0; ww < w; ww++) before += s_wsum[ww];
        const uint32_t* base = table + TC_ROWS_AT + (size_t)blockIdx.x * Tpad;
        for (int t = t0 + lane; t < t1; t += 32) {
            const uint32_t at = s_ctr[t] + before;
            if (blockIdx.x == 0) {
                const uint32_t c = table[t];
                ranges[t] = c ? make_uint2(at, at + c) : make_uint2(0u, 0u);
            }
            s_ctr[t] = at + base[t];
        }
    }
    const uint32_t ctr_addr = smem_addr(s_ctr);
    const size_t b = (size_t)blockIdx.x * chunk, e = min(R, b + chunk);
    const size_t mine = (size_t)w * (32 * EPT) + lane;                  // my first instance inside a round
    uint32_t nk[EPT], nv[EPT];
#pragma unroll
    for (int j = 0; j < EPT; j++) {
        const size_t i = b + mine + 32 * j;
        nk[j] = i < e ? keys[i] : 0xFFFFFFFFu;
        nv[j] = i < e ? vals[i] : 0u;
    }
    __syncthreads();
    const uint32_t lt = (1u << lane) - 1u;
    const int wait_id = 1 + w, next_id = 1 + ((w + 1) % NW);
    bool first = w == 0;                                                   // the block's first turn has no token to wait for
    for (size_t sub = b; sub < e; sub += (TC_THREADS * EPT)) {
        uint32_t k[EPT], v[EPT], old[EPT], now[EPT];
        const bool last_round = sub + (TC_THREADS * EPT) >= e;
#pragma unroll
        for (int j = 0; j < EPT; j++) { k[j] = nk[j]; v[j] = nv[j]; }
#pragma unroll
        for (int j = 0; j < EPT; j++) {                                 // the next round's loads fly during this one
            const size_t i = sub + (TC_THREADS * EPT) + mine + 32 * j;
            nk[j] = i < e ? keys[i] : 0xFFFFFFFFu;
            nv[j] = i < e ? vals[i] : 0u;
        }
        // ---- my turn: (round, warp) in list order ----
        if (!first) bar_sync(wait_id, 64);
        first = false;
#pragma unroll
        for (int j = 0; j < EPT; j++) {
            // Every lane takes a slot of its tile and reads the cursor back.  Shared-memory operations of a warp are performed
            // in order, so the read-back sees exactly this instruction's additions (nobody else touches the cursors during my
            // turn) and the token can move on as soon as these are ISSUED: bar.arrive / bar.sync is PTX's producer / consumer
            // pattern (store, arrive | sync, load); a fence would also wait for the warp's global accesses in flight.
            old[j] = now[j] = 0;
            if (k[j] != 0xFFFFFFFFu) old[j] = atoms_add_u32(ctr_addr + 4u * k[j], 1u);
            __syncwarp();                                                  // all of the warp's additions, then the read-backs
            if (k[j] != 0xFFFFFFFFu) now[j] = lds_u32_ordered(ctr_addr + 4u * k[j]);
            __syncwarp();                                                  // ... before the next 32 instances add theirs
        }
        if (!(last_round && w == NW - 1)) bar_arrive(next_id, 64);
#pragma unroll
        for (int j = 0; j < EPT; j++) {
            const bool valid = k[j] != 0xFFFFFFFFu;
            uint32_t pos = old[j];
            // Lanes of one tile inside these 32 instances (rare: they are different Gaussians) got their slots in an order the
            // hardware chooses; list order = lane order.  At least one of them sees the cursor more than one past its slot.
            if (__any_sync(0xffffffffu, valid && now[j] - old[j] != 1u)) {
                uint32_t peers = __ballot_sync(0xffffffffu, valid);
                for (int bit = 0; bit < tile_bits; bit++) {                // one ballot per bit of the tile id (match.any: a trip per distinct value)
                    const uint32_t mb = __ballot_sync(0xffffffffu, (k[j] >> bit) & 1u);
                    peers &= mb ^ (((k[j] >> bit) & 1u) - 1u);             // bit set: mb, clear: ~mb
                }
                pos = now[j] - (uint32_t)__popc(peers) + (uint32_t)__popc(peers & lt);
            }
            if (valid) out[pos] = v[j];
        }
    }
}

bool tile_count_sort_applies(size_t tiles) {
#ifdef GSR_EMU
    if (tiles * 4 > 96 * 1024) return false;          // host builds of this file (GSR_EMU) have a 96 KiB dynamic shared-memory window
#endif
    return tiles >= 1 && tiles <= (size_t)TC_MAX_TILES;
}

// (tile key, id) in list order -> ids grouped by tile in `out`, tile ranges in `ranges`.  `table`: tc_table_words(R) words.
void tile_count_sort(const uint32_t* keys, const uint32_t* vals, size_t R, size_t tiles, uint32_t* table, uint32_t* out, uint2* ranges,
                     cudaStream_t st) {
    const int T = (int)tiles, Tpad = (T + 31) & ~31;
    if (R == 0) { cudaMemsetAsync(ranges, 0, tiles * sizeof(uint2), st); return; }
    const size_t chunk = tc_chunk(R, tiles);
    const int C = (int)tc_chunks(R, tiles);
    const size_t smem = (size_t)Tpad * 4;                                  // the chunk's counters / cursors
    static bool configured[GSR_MAX_DEVICES] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < GSR_MAX_DEVICES && !configured[dev]) {
        cudaFuncSetAttribute(tc_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_MAX_TILES * 4);
        cudaFuncSetAttribute(tc_scatter_kernel<TC_EPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_MAX_TILES * 4);
        cudaFuncSetAttribute(tc_scatter_kernel<TC_EPT>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        configured[dev] = true;
    }
    int tile_bits = 0;
    while ((1u << tile_bits) < (unsigned)T) tile_bits++;
    tc_count_kernel<<<C, TC_THREADS, smem, st>>>(keys, R, chunk, Tpad, table);
    tc_colscan_kernel<<<Tpad / 32, 1024, 0, st>>>(table, C, Tpad);
    tc_scatter_kernel<TC_EPT><<<C, TC_THREADS, smem, st>>>(keys, vals, R, chunk, T, Tpad, tile_bits, table, out, ranges);
    count_launch(3);
}

// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) gather_u32_kernel(const uint32_t* __restrict__ src, const uint32_t* __restrict__ idx,
                                                         uint32_t* __restrict__ dst, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = __ldg(src + idx[i]);
}
void launch_gather_u32(const uint32_t* src, const uint32_t* idx, uint32_t* dst, size_t n, cudaStream_t st) {
    if (n == 0) return;
    gather_u32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(src, idx, dst, n);
    count_launch(1);
}

// tiles_touched and the tile rectangle into depth order in one pass (the rectangle only for Gaussians that emit instances).
// One Gaussian per thread: four per thread (more gathers in flight per thread) measured slower, 0.057 against 0.051 ms for the
// stage at P = 3 M -- the strided 8-byte stores of the rectangles cost more than the extra loads in flight gain.
__global__ void __launch_bounds__(256) gather_tt_rect_kernel(const uint32_t* __restrict__ tt, const uint2* __restrict__ rects,
                                                             const uint32_t* __restrict__ idx, uint32_t* __restrict__ tt_sorted,
                                                             uint2* __restrict__ rect_sorted, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t g = idx[i];
    const uint32_t c = __ldg(tt + g);
    tt_sorted[i] = c;
    if (c) rect_sorted[i] = __ldg(rects + g);
}
void launch_gather_tt_rect(const uint32_t* tt, const uint2* rects, const uint32_t* idx, uint32_t* tt_sorted, uint2* rect_sorted, size_t n,
                           cudaStream_t st) {
    if (n == 0) return;
    gather_tt_rect_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(tt, rects, idx, tt_sorted, rect_sorted, n);
    count_launch(1);
}

// duplicateWithKeys (rasterizer_impl.cu:78-126), walked in depth order; key = tile id only.
__global__ void __launch_bounds__(256) duplicate_kernel(int P, const uint32_t* __restrict__ order, const uint32_t* __restrict__ offsets,
                                                        const uint32_t* __restrict__ tt_sorted, const float4* __restrict__ packed,
                                                        const int* __restrict__ radii, int grid_x, int grid_y, int row_begin, int row_end,
                                                        uint32_t* __restrict__ tile_keys, uint32_t* __restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    if (tt_sorted[i] == 0) return;
    const uint32_t g = order[i];
    const float4 a = __ldg(packed + 3 * (size_t)g);
    uint32_t x0, y0, x1, y1;
    get_rect(a.x, a.y, radii[g], (uint32_t)grid_x, (uint32_t)grid_y, x0, y0, x1, y1);
    y0 = max(y0, (uint32_t)row_begin);
    y1 = min(y1, (uint32_t)row_end);
    uint32_t off = offsets[i];
    for (uint32_t y = y0; y < y1; y++)
        for (uint32_t x = x0; x < x1; x++) {
            tile_keys[off] = y * (uint32_t)grid_x + x;
            vals[off] = g;
            off++;
        }
}

// Warp-cooperative form: the 32 Gaussians of a warp own one contiguous span of the instance list (offsets are
// the prefix sum in this same order).  Lanes walk that span 32 entries at a time, find each entry's owner with
// a 5-step shuffle search over the lanes' span starts, and write keys and ids fully coalesced; a Gaussian that
// covers hundreds of tiles is spread over the warp instead of serialising one lane.
// Everything it reads is sequential in depth order (id, offset, count, rectangle: 20 bytes per Gaussian); round 1 gathered the
// 16-byte record and the radius per Gaussian here (361 MB of sector traffic for 156 MB of algorithmic bytes at C2).
__global__ void __launch_bounds__(256) duplicate_coop_kernel(int P, const uint32_t* __restrict__ order, const uint32_t* __restrict__ offsets,
                                                             const uint32_t* __restrict__ tt_sorted, const uint2* __restrict__ rect_sorted,
                                                             int grid_x, uint32_t* __restrict__ tile_keys, uint32_t* __restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t cnt = 0, g = 0, off = 0, xy0 = 0, w = 1;
    if (i < P) {
        cnt = tt_sorted[i];
        off = offsets[i];
        if (cnt != 0) {
            g = order[i];
            const uint2 rc = rect_sorted[i];
            xy0 = rc.x;
            w = rc.y;
        }
    }
    const uint32_t base = __shfl_sync(0xffffffffu, off, 0);
    const uint32_t total = __reduce_max_sync(0xffffffffu, (i < P) ? off + cnt - base : 0u);
    const uint32_t rel = (i < P) ? off - base : 0xFFFFFFFFu;      // non-decreasing over the lanes
    for (uint32_t k = lane; k < ((total + 31u) & ~31u); k += 32) {
        int lo = 0;
        uint32_t rlo = 0;                                          // lane 0 starts at 0
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
            const uint32_t r = __shfl_sync(0xffffffffu, rel, lo + step);
            if (r <= k) { lo += step; rlo = r; }
        }
        const uint32_t og = __shfl_sync(0xffffffffu, g, lo);
        const uint32_t oxy = __shfl_sync(0xffffffffu, xy0, lo);
        const uint32_t ow = __shfl_sync(0xffffffffu, w, lo);
        if (k < total) {
            const uint32_t t = k - rlo;
            const uint32_t dy = t / ow, dx = t - dy * ow;
            tile_keys[base + k] = ((oxy >> 16) + dy) * (uint32_t)grid_x + (oxy & 0xFFFFu) + dx;
            vals[base + k] = og;
        }
    }
}

void launch_duplicate(int P, const uint32_t* order, const uint32_t* offsets, const uint32_t* tt_sorted, const uint2* rect_sorted,
                      const float4* packed, const int* radii, int grid_x, int grid_y, int row_begin, int row_end,
                      uint32_t* tile_keys, uint32_t* vals, cudaStream_t st) {
    if (P <= 0) return;
    if (grid_x < 65536 && grid_y < 65536 && rect_sorted != nullptr)   // measured 0.111 -> 0.075 ms at C2 against the one-thread-per-Gaussian walk
        duplicate_coop_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, order, offsets, tt_sorted, rect_sorted, grid_x, tile_keys, vals);
    else
        duplicate_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, order, offsets, tt_sorted, packed, radii, grid_x, grid_y,
                                                          row_begin, row_end, tile_keys, vals);
    count_launch(1);
}

// identifyTileRanges (rasterizer_impl.cu:133-159); ranges are zeroed first (:377)
__global__ void __launch_bounds__(256) tile_ranges_kernel(size_t R, const uint32_t* __restrict__ keys, uint2* __restrict__ ranges) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const uint32_t cur = keys[i];
    if (i == 0) ranges[cur].x = 0;
    else {
        const uint32_t prev = keys[i - 1];
        if (cur != prev) {
            ranges[prev].y = (uint32_t)i;
            ranges[cur].x = (uint32_t)i;
        }
    }
    if (i == R - 1) ranges[cur].y = (uint32_t)R;
}

// four keys per thread (one 16-byte load + the key before them)
__global__ void __launch_bounds__(256) tile_ranges4_kernel(size_t R, const uint32_t* __restrict__ keys, uint2* __restrict__ ranges) {
    const size_t i0 = 4 * ((size_t)blockIdx.x * blockDim.x + threadIdx.x);
    if (i0 >= R) return;
    uint32_t k[5];
    k[0] = i0 ? keys[i0 - 1] : 0u;
    if (i0 + 4 <= R) {
        const uint4 q = *reinterpret_cast<const uint4*>(keys + i0);
        k[1] = q.x; k[2] = q.y; k[3] = q.z; k[4] = q.w;
    } else {
#pragma unroll
        for (int j = 0; j < 4; j++) k[1 + j] = (i0 + j < R) ? keys[i0 + j] : 0u;
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const size_t i = i0 + j;
        if (i >= R) break;
        const uint32_t cur = k[1 + j];
        if (i == 0) ranges[cur].x = 0;
        else if (cur != k[j]) {
            ranges[k[j]].y = (uint32_t)i;
            ranges[cur].x = (uint32_t)i;
        }
        if (i == R - 1) ranges[cur].y = (uint32_t)R;
    }
}

void launch_tile_ranges(size_t R, const uint32_t* sorted_tile_keys, uint2* ranges, size_t tiles, cudaStream_t st) {
    cudaMemsetAsync(ranges, 0, tiles * sizeof(uint2), st);
    if (R == 0) return;
    if ((reinterpret_cast<uintptr_t>(sorted_tile_keys) & 15) == 0)
        tile_ranges4_kernel<<<(unsigned)((R + 1023) / 1024), 256, 0, st>>>(R, sorted_tile_keys, ranges);
    else
        tile_ranges_kernel<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(R, sorted_tile_keys, ranges);
    count_launch(1);
}

}  // namespace gsr
