// emu_next.cpp -- the Adam and k-NN kernels of the product, compiled for the HOST through tests/cuda_emu/emu.h and
// driven by plain C entry points for tests/test_kernels_emulated.py.  TEST INFRASTRUCTURE ONLY (see emu.h).
#define GSR_EMU 1
#include "emu.h"

GSR_EMU_DEFINE_GLOBALS
GSR_EMU_DEFINE_RUNTIME      // count_launch + the few CUDA runtime calls the host code makes


#include "_gen/gsr_adam.cpp"
#include "_gen/gsr_binning.cpp"
#include "_gen/gsr_knn.cpp"

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

// gsr_knn_mean_dist2 as csrc/gsr_api.cu runs it: the product's own launch_knn, radix sort included
int emu_knn_mean_dist2(int P, const float* points, float* mean_d2) {
    if (P <= 0) return 0;
    KnnState sizes = KnnState::carve(nullptr, (size_t)P);
    std::vector<char> raw(sizes.bytes + 256);
    char* base = raw.data() + (256 - ((uintptr_t)raw.data() & 255)) % 256;
    launch_knn(P, points, mean_d2, KnnState::carve(base, (size_t)P), nullptr);
    return (int)knn_boxes((size_t)P);
}

}  // extern "C"
