// emu_next.cpp -- the Adam and k-NN kernels of the product, compiled for the HOST through tests/cuda_emu/emu.h and
// driven by plain C entry points for tests/test_kernels_emulated.py.  TEST INFRASTRUCTURE ONLY (see emu.h).
#define GSR_EMU 1
#include "emu.h"

namespace emu { BlockCtx g_ctx; thread_local int t_lane, t_warp; }
thread_local uint3 threadIdx, blockIdx;
thread_local dim3 blockDim, gridDim;

#include <algorithm>

#include "../../gaussian-splatting_cc-comments_b200/csrc/gsr_adam.cu"
#include "../../gaussian-splatting_cc-comments_b200/csrc/gsr_knn.cu"

using namespace gsr;

extern "C" {

// the host half of gsr_adam_step (csrc/gsr_api.cu), restated for one launch of up to ADAM_MAX_GROUPS groups
int emu_adam_step(int n_groups, float** param, const float** grad, float** exp_avg, float** exp_avg_sq, const size_t* n,
                  const double* lr, const long long* step, const int* row_width, double beta1, double beta2, double eps,
                  const unsigned char* visible, int force_scalar) {
    if (n_groups > ADAM_MAX_GROUPS) return -1;
    AdamLaunch L;
    std::memset(&L, 0, sizeof(L));
    L.beta2 = (float)beta2; L.w1 = (float)(1.0 - beta1); L.w2 = (float)(1.0 - beta2); L.eps = (float)eps;
    L.visible = visible;
    unsigned blocks = 0;
    bool vec = !force_scalar;
    for (int i = 0; i < n_groups; i++) {
        if (n[i] == 0) continue;
        AdamGroupDev& d = L.g[L.n_groups];
        d.p = param[i]; d.g = grad[i]; d.m = exp_avg[i]; d.v = exp_avg_sq[i]; d.n = n[i];
        d.neg_step = (float)(-(lr[i] / (1.0 - std::pow(beta1, (double)step[i]))));
        d.bc2_sqrt = (float)std::sqrt(1.0 - std::pow(beta2, (double)step[i]));
        d.row_width = (unsigned)(row_width[i] > 0 ? row_width[i] : 1);
        if (((uintptr_t)d.p | (uintptr_t)d.g | (uintptr_t)d.m | (uintptr_t)d.v) & 15) vec = false;
        L.block_start[L.n_groups++] = blocks;
        blocks += (unsigned)adam_blocks(n[i]);
    }
    for (int k = L.n_groups; k <= ADAM_MAX_GROUPS; k++) L.block_start[k] = blocks;
    if (blocks == 0) return 0;
    const bool masked = visible != nullptr;
    emu::launch(dim3(blocks), dim3(ADAM_THREADS), [&] {
        if (vec) { if (masked) adam_kernel<true, true>(L); else adam_kernel<true, false>(L); }
        else { if (masked) adam_kernel<false, true>(L); else adam_kernel<false, false>(L); }
    });
    return vec ? 1 : 2;
}

// launch_knn (csrc/gsr_knn.cu) with the radix sort replaced by std::stable_sort (the sort kernels are covered on the GPU)
int emu_knn_mean_dist2(int P, const float* points, float* mean_d2) {
    if (P <= 0) return 0;
    KnnState sizes = KnnState::carve(nullptr, (size_t)P);
    std::vector<char> raw(sizes.bytes + 256);
    char* base = raw.data() + (256 - ((uintptr_t)raw.data() & 255)) % 256;
    KnnState s = KnnState::carve(base, (size_t)P);
    std::memset(s.bounds, 0, 8 * sizeof(int));
    const unsigned rb = (unsigned)min((size_t)4, ((size_t)P + 255) / 256);      // few blocks: exercises the grid stride
    emu::launch(dim3(rb), dim3(256), [&] { knn_bounds_kernel(P, points, s.bounds); });
    emu::launch(dim3((unsigned)(((size_t)P + 255) / 256)), dim3(256), [&] { knn_morton_kernel(P, points, s.bounds, s.keys[0], s.vals[0]); });
    std::vector<uint32_t> order(P);
    for (int i = 0; i < P; i++) order[i] = (uint32_t)i;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return (s.keys[0][a] & 0x3FFFFFFFu) < (s.keys[0][b] & 0x3FFFFFFFu); });
    for (int i = 0; i < P; i++) { s.keys[1][i] = s.keys[0][order[i]]; s.vals[1][i] = s.vals[0][order[i]]; }
    const unsigned nb = (unsigned)knn_boxes((size_t)P);
    emu::launch(dim3(nb), dim3(KNN_BOX), [&] { knn_gather_kernel(P, points, s.vals[1], s.sorted, s.boxes); });
    emu::launch(dim3(nb), dim3(KNN_BOX), [&] { knn_search_kernel(P, s.sorted, s.boxes, (int)nb, mean_d2); });
    return (int)nb;
}

}  // extern "C"
