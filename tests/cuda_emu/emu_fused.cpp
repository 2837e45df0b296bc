// emu_fused.cpp -- the raw-parameter helpers of the per-Gaussian kernels (csrc/gsr_fused_params.cuh) compiled for the
// HOST through tests/cuda_emu/emu.h, with plain C entry points for tests/test_kernels_emulated.py.
// TEST INFRASTRUCTURE ONLY (see emu.h).
#define GSR_EMU 1
#include "emu.h"

GSR_EMU_DEFINE_GLOBALS

#include "gsr_fused_params.cuh"

using namespace gsr;

extern "C" {

// colour of every Gaussian from split SH storage, as preprocess_fwd_kernel<.., SPLIT = true> calls it
void emu_sh_color_split(int P, int M, int D, const float* dc, const float* rest, const float* k /* (P,16) */, float* out /* (P,3) */,
                        int allow_vec) {
    for (int idx = 0; idx < P; idx++) {
        const float* d = dc + 3 * (size_t)idx;
        const float* r = rest + (size_t)idx * (M - 1) * 3;
        const bool vec_ok = allow_vec && ((reinterpret_cast<uintptr_t>(rest) & 15) == 0) && idx < P - 1;
        float res[3] = {0.f, 0.f, 0.f};
        const float* kk = k + 16 * (size_t)idx;
        if (D == 3) sh_color_split<3>(d, r, vec_ok, kk, res);
        else if (D == 2) sh_color_split<2>(d, r, vec_ok, kk, res);
        else if (D == 1) sh_color_split<1>(d, r, vec_ok, kk, res);
        else sh_color_split<0>(d, r, vec_ok, kk, res);
        out[3 * idx] = res[0]; out[3 * idx + 1] = res[1]; out[3 * idx + 2] = res[2];
    }
}

// one block of the backward's staging: rest rows -> shared rows (returned), as preprocess_bwd_kernel_staged<.., SPLIT> does
void emu_stage_rest_rows(int nrows, int M, int live_rest, int S, const unsigned char* vis, const float* rest_block, float* rows /* (128,S) */) {
    emu::launch(dim3(1), dim3(128), [&] { stage_rest_rows<128>(rows, S, rest_block, nrows, 3 * (M - 1), live_rest, vis, (int)threadIdx.x); });
}

// ... and the split write-out of the rows after the per-Gaussian backward
void emu_store_split_rows(int nrows, int M, int live, int S, const unsigned char* vis, const float* rows, float* dc_out, float* rest_out) {
    emu::launch(dim3(1), dim3(128), [&] { store_split_rows<128>(rows, S, dc_out, rest_out, nrows, 3 * (M - 1), live, vis, (int)threadIdx.x); });
}

// activations forward + backward for n Gaussians: y = f(raw), g_raw = f'(raw) applied to g_y
void emu_activations(int n, const float* raw_op, const float* raw_s /* (n,3) */, const float* raw_q /* (n,4) */, const float* g_op,
                     const float* g_s, const float* g_q, float* op, float* s, float* q, float* d_op, float* d_s, float* d_q) {
    for (int i = 0; i < n; i++) {
        op[i] = act_sigmoid(raw_op[i]);
        d_op[i] = act_sigmoid_bwd(op[i], g_op[i]);
        for (int a = 0; a < 3; a++) { s[3 * i + a] = act_exp(raw_s[3 * i + a]); d_s[3 * i + a] = act_exp_bwd(s[3 * i + a], g_s[3 * i + a]); }
        const QuatNorm f = act_normalize(make_float4(raw_q[4 * i], raw_q[4 * i + 1], raw_q[4 * i + 2], raw_q[4 * i + 3]));
        q[4 * i] = f.q.x; q[4 * i + 1] = f.q.y; q[4 * i + 2] = f.q.z; q[4 * i + 3] = f.q.w;
        const float4 d = act_normalize_bwd(f, make_float4(g_q[4 * i], g_q[4 * i + 1], g_q[4 * i + 2], g_q[4 * i + 3]));
        d_q[4 * i] = d.x; d_q[4 * i + 1] = d.y; d_q[4 * i + 2] = d.z; d_q[4 * i + 3] = d.w;
    }
}

}  // extern "C"
