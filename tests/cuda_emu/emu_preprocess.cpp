// emu_preprocess.cpp -- the per-Gaussian kernels of the product (csrc/gsr_preprocess.cu, csrc/gsr_preprocess_bwd.cu)
// compiled for the HOST through tests/cuda_emu/emu.h, with plain C entry points for tests/test_kernels_emulated.py.
// The reference-shaped instantiations are checked against the fixtures the reference's CUDA build produced on a B200
// (tests/golden/ref_cuda_*.npz); the raw-parameter instantiations against the reference-shaped ones.
// TEST INFRASTRUCTURE ONLY (see emu.h).
#define GSR_EMU 1
#include "emu.h"

GSR_EMU_DEFINE_GLOBALS
GSR_EMU_DEFINE_RUNTIME      // count_launch + the few CUDA runtime calls the host code makes

#include "_gen/gsr_preprocess.cpp"
#include "_gen/gsr_preprocess_bwd.cpp"

using namespace gsr;

extern "C" {

// preprocess_fwd_kernel<RAW, SPLIT>; NULL = not provided, as in include/gsr.h.  packed: (P,12) floats.
void emu_preprocess_fwd(int P, int D, int M, const float* means3D, const float* scales, float scale_modifier, const float* rotations,
                        const float* opacities, const float* shs, const float* f_dc, const float* f_rest, const float* view,
                        const float* proj, const float* campos, int W, int H, float tan_fovx, float tan_fovy, int raw_act, int* radii,
                        float* packed, unsigned* tiles_touched, unsigned char* clamped, unsigned* depth_keys, unsigned* total) {
    PreprocessArgs a;
    a.P = P; a.D = D; a.M = M;
    a.means3D = means3D; a.scales = scales; a.scale_modifier = scale_modifier; a.rotations = rotations;
    a.opacities = opacities; a.shs = shs; a.cov3D_precomp = nullptr; a.colors_precomp = nullptr;
    a.cam.view = view; a.cam.proj = proj; a.cam.campos = campos;
    a.W = W; a.H = H; a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
    a.focal_y = H / (2.0f * tan_fovy); a.focal_x = W / (2.0f * tan_fovx);
    a.grid_x = (W + 15) / 16; a.grid_y = (H + 15) / 16; a.row_begin = 0; a.row_end = a.grid_y;
    a.prefiltered = 0;
    a.radii = radii; a.packed = reinterpret_cast<float4*>(packed); a.tiles_touched = tiles_touched; a.clamped = clamped;
    a.depth_keys = depth_keys;
    std::vector<uint32_t> ids(P);
    a.ids = ids.data();
    *total = 0; a.total = total;
    a.raw_act = raw_act; a.f_dc = f_dc; a.f_rest = f_rest; a.rects = nullptr;
    const bool split = f_dc != nullptr;
    emu::launch(dim3((P + 255) / 256), dim3(256), [&] {
        if (!raw_act && !split) preprocess_fwd_kernel<false, false>(a);
        else if (raw_act && !split) preprocess_fwd_kernel<true, false>(a);
        else if (!raw_act && split) preprocess_fwd_kernel<false, true>(a);
        else preprocess_fwd_kernel<true, true>(a);
    });
}

// the per-Gaussian backward, staged kernel (or the generic one with generic != 0); same kernel selection and shared-memory
// row stride as launch_preprocess_backward.  grad_acc (P,12) are the blending backward's accumulators.
int emu_preprocess_bwd(int P, int D, int M, const float* means3D, const int* radii, const float* shs, const unsigned char* clamped,
                       const float* scales, const float* rotations, float scale_modifier, const float* view, const float* proj,
                       const float* campos, int W, int H, float tan_fovx, float tan_fovy, const float* grad_acc, const float* packed,
                       int raw_act, const float* f_dc, const float* f_rest, int generic, float* dL_dmean2D, float* dL_dconic,
                       float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale,
                       float* dL_drot, float* dL_df_dc, float* dL_df_rest) {
    PreprocessBwdArgs a;
    a.P = P; a.D = D; a.M = M;
    a.means3D = means3D; a.radii = radii; a.shs = shs; a.clamped = clamped;
    a.scales = scales; a.rotations = rotations; a.scale_modifier = scale_modifier; a.cov3D_precomp = nullptr;
    a.cam.view = view; a.cam.proj = proj; a.cam.campos = campos;
    a.W = W; a.H = H; a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
    a.focal_y = H / (2.0f * tan_fovy); a.focal_x = W / (2.0f * tan_fovx);
    a.grad_acc = grad_acc; a.packed = reinterpret_cast<const float4*>(packed);
    a.dL_dmean2D = dL_dmean2D; a.dL_dconic = dL_dconic; a.dL_dopacity = dL_dopacity; a.dL_dcolor = dL_dcolor;
    a.dL_dmean3D = dL_dmean3D; a.dL_dcov3D = dL_dcov3D; a.dL_dsh = dL_dsh; a.dL_dscale = dL_dscale; a.dL_drot = dL_drot;
    a.dL_drgb = nullptr; a.skip_dsh = 0; a.begin = 0; a.chunked = 0; a.chunk_events = nullptr; a.n_chunk_events = 0; a.small_pack = nullptr;
    a.raw_act = raw_act; a.f_dc = f_dc; a.f_rest = f_rest; a.dL_df_dc = dL_df_dc; a.dL_df_rest = dL_df_rest;
    const int L = 3 * M;
    if (f_dc != nullptr) {
        const int S4 = (L + 3) / 4, Ss = 4 * (S4 % 2 == 1 ? S4 : S4 + 1);
        emu::launch(dim3((P + PB - 1) / PB), dim3(PB), [&] {
            if (raw_act) preprocess_bwd_kernel_staged<8, false, true, true>(a, Ss);
            else preprocess_bwd_kernel_staged<8, false, false, true>(a, Ss);
        });
        return 3;
    }
    const int S = ((L / 4) % 2 == 1 || L == 0) ? L : L + 4;
    const bool aligned = (reinterpret_cast<uintptr_t>(shs) & 15) == 0 && (reinterpret_cast<uintptr_t>(dL_dsh) & 15) == 0 &&
                         (reinterpret_cast<uintptr_t>(rotations) & 15) == 0;
    const bool staged = !generic && (L % 4 == 0) && aligned;
    if (staged) {
        emu::launch(dim3((P + PB - 1) / PB), dim3(PB), [&] {
            if (raw_act) preprocess_bwd_kernel_staged<8, false, true, false>(a, S);
            else preprocess_bwd_kernel_staged<8, false, false, false>(a, S);
        });
        return 1;
    }
    emu::launch(dim3((P + 255) / 256), dim3(256), [&] {
        if (raw_act) preprocess_bwd_kernel_generic<true>(a);
        else preprocess_bwd_kernel_generic<false>(a);
    });
    return 2;
}

}  // extern "C"
