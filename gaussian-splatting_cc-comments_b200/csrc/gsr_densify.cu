// gsr_densify.cu -- adaptive density control of the Gaussian model (second half).
//
// Two things the reference's training loop does to the model besides the optimizer step:
//   * every iteration until densify_until_iter (train.py:153-158): max_radii2D, xyz_gradient_accum and denom are
//     updated for the visible Gaussians through boolean-mask indexing -- eight torch kernels and two nonzero()
//     synchronisations.  densify_stats_kernel is the same update as one pass, 32 bytes per Gaussian, no host round trip;
//   * every densification_interval iterations: densify_and_prune (scene/gaussian_model.py:558-597) clones small
//     high-gradient Gaussians, splits large ones into N children, then prunes transparent / oversized ones -- three
//     rounds of mask-indexing and torch.cat over 6 parameter tensors and 12 Adam moments, every row copied up to three
//     times.  Here the three rounds collapse into decisions per ORIGINAL Gaussian (a clone has its parent's opacity and
//     scale, so it survives the final prune exactly when the parent's values pass; a child's scale is the parent's
//     shrunk by 1/(0.8N), recomputed here in the reference's arithmetic), a prefix sum of four counters, and one
//     kernel that moves every surviving row once: HBM-bound, 12 bytes per kept float moved (param + two moments read and
//     written), reads coalesced over the block's 256 consecutive rows, writes in coalesced runs because the ranks are
//     monotone.
#include "gsr_internal.cuh"
#include "gsr_fused_params.cuh"

namespace gsr {

// ---- per-iteration statistics ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) densify_stats_kernel(int P, const float* __restrict__ vs_grad, const int* __restrict__ radii,
                                                            float* __restrict__ accum, float* __restrict__ denom,
                                                            float* __restrict__ max_radii) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int r = __ldg(radii + i);
    if (r <= 0) return;                                              // visibility_filter = radii > 0
    const float gx = __ldg(vs_grad + 3 * (size_t)i), gy = __ldg(vs_grad + 3 * (size_t)i + 1);
    accum[i] = addr(accum[i], __fsqrt_rn(fmar(gy, gy, mulr(gx, gx))));   // torch.norm(grad[:, :2], dim=-1)
    denom[i] = addr(denom[i], 1.0f);
    if (max_radii) max_radii[i] = fmaxf(max_radii[i], (float)r);
}

void launch_densify_stats(int P, const float* vs_grad, const int* radii, float* accum, float* denom, float* max_radii, cudaStream_t st) {
    if (P <= 0) return;
    densify_stats_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, vs_grad, radii, accum, denom, max_radii);
    count_launch(1);
}

// ---- classify --------------------------------------------------------------------------------------------------------
DensifyState DensifyState::carve(char* base, size_t P) {
    DensifyState s;
    const size_t nb = (P + DENSIFY_THREADS - 1) / DENSIFY_THREADS;
    s.block_offsets = reinterpret_cast<uint32_t*>(base);
    s.bytes = ((4 * nb * sizeof(uint32_t)) + 255) & ~(size_t)255;
    return s;
}

// Decisions of densify_and_clone, densify_and_split and the final prune for ONE original Gaussian, from its own values.
GSR_DI unsigned densify_decide(float accum, float denom, float s0, float s1, float s2, float opacity_raw, const DensifyThresholds& t) {
    float g = __fdiv_rn(accum, denom);                               // grads = xyz_gradient_accum / denom
    if (g != g) g = 0.f;                                             // grads[grads.isnan()] = 0.0
    const float e0 = act_exp(s0), e1 = act_exp(s1), e2 = act_exp(s2);
    const float smax = fmaxf(fmaxf(e0, e1), e2);                     // torch.max(self.get_scaling, dim=1).values
    const bool clone_hot = __fsqrt_rn(mulr(g, g)) >= t.max_grad;     // torch.norm(grads, dim=-1) over the size-1 dimension
    const bool split_hot = g >= t.max_grad;                          // padded_grad >= grad_threshold
    const bool clone = clone_hot && smax <= t.dense_extent;
    const bool split = split_hot && smax > t.dense_extent;
    const bool low = act_sigmoid(opacity_raw) < t.min_opacity;       // self.get_opacity < min_opacity
    const bool big = t.use_screen && smax > t.big_extent;            // big_points_ws (big_points_vs is always False, see gsr.h)
    // a child's activated scale: exp(log(exp(s) * inv)), evaluated as the reference stores and re-activates it
    const float c0 = act_exp(logf(mulr(e0, t.inv_shrink))), c1 = act_exp(logf(mulr(e1, t.inv_shrink))), c2 = act_exp(logf(mulr(e2, t.inv_shrink)));
    const bool big_child = t.use_screen && fmaxf(fmaxf(c0, c1), c2) > t.big_extent;
    unsigned f = 0;
    if (!split && !(low || big)) f |= 1u;                            // the original stays
    if (clone) f |= 8u;
    if (clone && !(low || big)) f |= 2u;                             // its clone stays (same opacity and scale)
    if (split) f |= 16u;
    if (split && !(low || big_child)) f |= 4u;                       // its children stay
    return f;
}

__global__ void __launch_bounds__(DENSIFY_THREADS) densify_classify_kernel(int P, const float* __restrict__ accum, const float* __restrict__ denom,
                                                                           const float* __restrict__ scaling, const float* __restrict__ opacity,
                                                                           const __grid_constant__ DensifyThresholds t, uint8_t* __restrict__ flags,
                                                                           uint32_t* __restrict__ block_counts) {
    const int i = blockIdx.x * DENSIFY_THREADS + threadIdx.x;
    unsigned f = 0;
    if (i < P) {
        const float* s = scaling + 3 * (size_t)i;
        f = densify_decide(__ldg(accum + i), __ldg(denom + i), __ldg(s), __ldg(s + 1), __ldg(s + 2), __ldg(opacity + i), t);
        flags[i] = (uint8_t)f;
    }
    const int n_ok = __syncthreads_count(f & 1u), n_ck = __syncthreads_count(f & 2u);
    const int n_ss = __syncthreads_count(f & 16u), n_chk = __syncthreads_count(f & 4u);
    if (threadIdx.x == 0) {
        const size_t nb = gridDim.x;
        block_counts[blockIdx.x] = n_ok;
        block_counts[nb + blockIdx.x] = n_ck;
        block_counts[2 * nb + blockIdx.x] = n_ss;
        block_counts[3 * nb + blockIdx.x] = n_chk;
    }
}

// One block turns the four rows of per-block counts into exclusive offsets and writes the totals.
__global__ void __launch_bounds__(1024) densify_scan_kernel(int nb, uint32_t* __restrict__ block_counts, int copies, int* __restrict__ totals) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_total[4];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int per = (nb + 1023) / 1024;
    for (int c = 0; c < 4; c++) {
        uint32_t* row = block_counts + (size_t)c * nb;
        const int lo = min(t * per, nb), hi = min(lo + per, nb);
        uint32_t sum = 0;
        for (int j = lo; j < hi; j++) sum += row[j];
        uint32_t inc = sum;                                          // inclusive scan of the 1024 thread sums
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t o = __shfl_up_sync(0xffffffffu, w, d);
                if (lane >= d) w += o;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        uint32_t run = inc - sum + (warp ? s_warp[warp - 1] : 0u);   // exclusive offset of this thread's first block
        for (int j = lo; j < hi; j++) { const uint32_t n = row[j]; row[j] = run; run += n; }
        if (t == 1023) s_total[c] = run;
        __syncthreads();
    }
    if (t == 0) {
        totals[0] = (int)(s_total[0] + s_total[1] + (uint32_t)copies * s_total[3]);
        totals[1] = (int)s_total[2];
        totals[2] = (int)s_total[0];
        totals[3] = (int)s_total[1];
        totals[4] = (int)s_total[3];
    }
}

void launch_densify_classify(int P, const float* accum, const float* denom, const float* scaling, const float* opacity,
                             const DensifyThresholds& t, uint8_t* flags, int* totals, const DensifyState& s, cudaStream_t st) {
    const int nb = (P + DENSIFY_THREADS - 1) / DENSIFY_THREADS;
    densify_classify_kernel<<<nb, DENSIFY_THREADS, 0, st>>>(P, accum, denom, scaling, opacity, t, flags, s.block_offsets);
    densify_scan_kernel<<<1, 1024, 0, st>>>(nb, s.block_offsets, t.copies, totals);
    count_launch(2);
}

// ---- apply -------------------------------------------------------------------------------------------------------------
// build_rotation (utils/general_utils.py:90-114) of a RAW quaternion, each torch operation rounded on its own, applied to
// one sample: R s + xyz.  (The reference's torch.bmm leaves the order of the three products to cuBLAS; compared to 1e-6.)
GSR_DI void split_child_xyz(const float* rot, const float* xyz, const float* smp, float* out) {
    const float n = __fsqrt_rn(addr(addr(addr(mulr(rot[0], rot[0]), mulr(rot[1], rot[1])), mulr(rot[2], rot[2])), mulr(rot[3], rot[3])));
    const float r = __fdiv_rn(rot[0], n), x = __fdiv_rn(rot[1], n), y = __fdiv_rn(rot[2], n), z = __fdiv_rn(rot[3], n);
    const float R00 = subr(1.f, mulr(2.f, addr(mulr(y, y), mulr(z, z)))), R01 = mulr(2.f, subr(mulr(x, y), mulr(r, z))), R02 = mulr(2.f, addr(mulr(x, z), mulr(r, y)));
    const float R10 = mulr(2.f, addr(mulr(x, y), mulr(r, z))), R11 = subr(1.f, mulr(2.f, addr(mulr(x, x), mulr(z, z)))), R12 = mulr(2.f, subr(mulr(y, z), mulr(r, x)));
    const float R20 = mulr(2.f, subr(mulr(x, z), mulr(r, y))), R21 = mulr(2.f, addr(mulr(y, z), mulr(r, x))), R22 = subr(1.f, mulr(2.f, addr(mulr(x, x), mulr(y, y))));
    out[0] = addr(fmar(R02, smp[2], fmar(R01, smp[1], mulr(R00, smp[0]))), xyz[0]);
    out[1] = addr(fmar(R12, smp[2], fmar(R11, smp[1], mulr(R10, smp[0]))), xyz[1]);
    out[2] = addr(fmar(R22, smp[2], fmar(R21, smp[1], mulr(R20, smp[0]))), xyz[2]);
}

// Moves the block's rows of one tensor group.  s_dst[c][row]: destination row for category c (0 original, 1 clone,
// 2 first child; further children follow at a stride of n_chk rows), or -1.  Thread t takes floats t, t + NT, ... of the
// block's nrows * w contiguous floats; row and column advance without a division per element.  Four elements are in
// flight per thread (up to twelve loads issued before the first store): the kernel is a gather-free copy whose only
// limit should be HBM, and one load per thread at a time does not cover the memory latency.
constexpr int DENSIFY_UNROLL = 4;
template <int NT>
GSR_DI void densify_move_rows(const DensifyGroupDev& G, size_t first_row, int nrows, const int (*s_dst)[DENSIFY_THREADS], int copies, int n_chk,
                              bool children_computed, int t) {
    const int w = G.w;
    const size_t src0 = first_row * (size_t)w;
    const int n = nrows * w;
    const int qstep = NT / w, rstep = NT % w;
    const bool state = G.mo != nullptr;
    int row = t / w, col = t - row * w;
    for (int f0 = t; f0 < n; f0 += NT * DENSIFY_UNROLL) {
        int d_ok[DENSIFY_UNROLL], d_ck[DENSIFY_UNROLL], d_ch[DENSIFY_UNROLL], cl[DENSIFY_UNROLL];
        float val[DENSIFY_UNROLL], mv[DENSIFY_UNROLL], vv[DENSIFY_UNROLL];
#pragma unroll
        for (int u = 0; u < DENSIFY_UNROLL; u++) {
            const int f = f0 + u * NT;
            d_ok[u] = d_ck[u] = d_ch[u] = -1;
            cl[u] = col;
            if (f < n) {
                d_ok[u] = s_dst[0][row]; d_ck[u] = s_dst[1][row]; d_ch[u] = s_dst[2][row];
                if ((d_ok[u] & d_ck[u] & d_ch[u]) >= 0) val[u] = __ldg(G.p + src0 + f);           // at least one destination (not all -1)
                if (state && d_ok[u] >= 0) { mv[u] = __ldg(G.m + src0 + f); vv[u] = __ldg(G.v + src0 + f); }
            }
            row += qstep; col += rstep;
            if (col >= w) { col -= w; row++; }
        }
#pragma unroll
        for (int u = 0; u < DENSIFY_UNROLL; u++) {
            if (d_ok[u] >= 0) {
                const size_t d = (size_t)d_ok[u] * w + cl[u];
                G.po[d] = val[u];
                if (state) { G.mo[d] = mv[u]; G.vo[d] = vv[u]; }
            }
            if (d_ck[u] >= 0) {
                const size_t d = (size_t)d_ck[u] * w + cl[u];
                G.po[d] = val[u];
                if (state) { G.mo[d] = 0.f; G.vo[d] = 0.f; }
            }
            if (d_ch[u] >= 0) {
                for (int k = 0; k < copies; k++) {
                    const size_t d = ((size_t)d_ch[u] + (size_t)k * n_chk) * w + cl[u];
                    if (!children_computed) G.po[d] = val[u];
                    if (state) { G.mo[d] = 0.f; G.vo[d] = 0.f; }
                }
            }
        }
    }
}

__global__ void __launch_bounds__(DENSIFY_THREADS) densify_apply_kernel(const __grid_constant__ DensifyApply A) {
    __shared__ int s_dst[3][DENSIFY_THREADS];
    __shared__ int s_warp[4][DENSIFY_THREADS / 32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const size_t first_row = (size_t)blockIdx.x * DENSIFY_THREADS;
    const int nrows = (int)min((size_t)DENSIFY_THREADS, (size_t)A.P - first_row);
    const unsigned f = t < nrows ? A.flags[first_row + t] : 0u;

    // rank of this row among the block's rows of each category, then the block's offsets from the scan
    const unsigned lt = (1u << lane) - 1u;
    const unsigned b_ok = __ballot_sync(0xffffffffu, f & 1u), b_ck = __ballot_sync(0xffffffffu, f & 2u);
    const unsigned b_ss = __ballot_sync(0xffffffffu, f & 16u), b_chk = __ballot_sync(0xffffffffu, f & 4u);
    if (lane == 0) { s_warp[0][warp] = __popc(b_ok); s_warp[1][warp] = __popc(b_ck); s_warp[2][warp] = __popc(b_ss); s_warp[3][warp] = __popc(b_chk); }
    __syncthreads();
    int r_ok = __popc(b_ok & lt), r_ck = __popc(b_ck & lt), r_ss = __popc(b_ss & lt), r_chk = __popc(b_chk & lt);
    for (int w = 0; w < warp; w++) { r_ok += s_warp[0][w]; r_ck += s_warp[1][w]; r_ss += s_warp[2][w]; r_chk += s_warp[3][w]; }
    const size_t nb = gridDim.x;
    r_ok += (int)A.block_offsets[blockIdx.x];
    r_ck += (int)A.block_offsets[nb + blockIdx.x];
    r_ss += (int)A.block_offsets[2 * nb + blockIdx.x];
    r_chk += (int)A.block_offsets[3 * nb + blockIdx.x];
    const int child0 = A.n_ok + A.n_ck + r_chk;                      // first child's row; child k at + k * n_chk
    s_dst[0][t] = (f & 1u) ? r_ok : -1;
    s_dst[1][t] = (f & 2u) ? A.n_ok + r_ck : -1;
    s_dst[2][t] = (f & 4u) ? child0 : -1;
    __syncthreads();

#pragma unroll 1
    for (int gi = 0; gi < A.n_groups; gi++)
        densify_move_rows<DENSIFY_THREADS>(A.g[gi], first_row, nrows, s_dst, A.t.copies, A.n_chk,
                                           gi == A.xyz_group || gi == A.scaling_group, t);

    // the two computed tensors of the children: position sampled inside the parent, scale shrunk by 1 / (0.8 N)
    if (f & 4u) {
        const size_t i = first_row + t;
        const float* sc = A.g[A.scaling_group].p + 3 * i;
        const float* rq = A.g[A.rotation_group].p + 4 * i;
        const float* xp = A.g[A.xyz_group].p + 3 * i;
        const float rot[4] = {__ldg(rq), __ldg(rq + 1), __ldg(rq + 2), __ldg(rq + 3)};
        const float xyz[3] = {__ldg(xp), __ldg(xp + 1), __ldg(xp + 2)};
        float ns[3];
#pragma unroll
        for (int c = 0; c < 3; c++) ns[c] = logf(mulr(act_exp(__ldg(sc + c)), A.t.inv_shrink));   // scaling_inverse_activation(get_scaling / (0.8 N))
        for (int k = 0; k < A.t.copies; k++) {
            const size_t d = 3 * ((size_t)child0 + (size_t)k * A.n_chk);
            const float* sp = A.samples + 3 * ((size_t)k * A.n_split + r_ss);
            const float smp[3] = {__ldg(sp), __ldg(sp + 1), __ldg(sp + 2)};
            float o[3];
            split_child_xyz(rot, xyz, smp, o);
#pragma unroll
            for (int c = 0; c < 3; c++) { A.g[A.xyz_group].po[d + c] = o[c]; A.g[A.scaling_group].po[d + c] = ns[c]; }
        }
    }
}

void launch_densify_apply(const DensifyApply& a, cudaStream_t st) {
    if (a.P <= 0) return;
    densify_apply_kernel<<<(a.P + DENSIFY_THREADS - 1) / DENSIFY_THREADS, DENSIFY_THREADS, 0, st>>>(a);
    count_launch(1);
}

}  // namespace gsr
