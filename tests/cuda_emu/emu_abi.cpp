// emu_abi.cpp -- the WHOLE C ABI of the product (include/gsr.h) compiled for the HOST: csrc/gsr_api.cu and every kernel
// file it drives, launch syntax rewritten by cu2host.py, on tests/cuda_emu/emu.h and a no-op CUDA runtime.  The library
// exports the same symbols as libgsr_b200.so (gsr_forward, gsr_backward, gsr_adam_step, gsr_knn_mean_dist2, ...) and
// takes HOST pointers, so tests/test_abi_emulated.py drives the product's own host code and kernels end to end on the
// CPU against the fixtures from the reference's CUDA build.  TEST INFRASTRUCTURE ONLY (see emu.h); slow by design.
#define GSR_EMU 1
#include "emu.h"

GSR_EMU_DEFINE_GLOBALS
GSR_EMU_DEFINE_CUDART

#include "_gen/gsr_preprocess.cpp"
#include "_gen/gsr_binning.cpp"
#include "_gen/gsr_render.cpp"
#include "_gen/gsr_render_bwd.cpp"
#include "_gen/gsr_preprocess_bwd.cpp"
#include "_gen/gsr_loss.cpp"
#include "_gen/gsr_adam.cpp"
#include "_gen/gsr_knn.cpp"
#include "_gen/gsr_densify.cpp"

// csrc/gsr_exchange.cu is NVLS multimem PTX (multi-GPU only): not emulated
namespace gsr {
void launch_multimem_allreduce(float*, size_t, size_t, int, int, int, cudaStream_t) { std::abort(); }
void launch_multimem_broadcast(float*, const float*, size_t, int, cudaStream_t) { std::abort(); }
}

#include "_gen/gsr_api.cpp"
