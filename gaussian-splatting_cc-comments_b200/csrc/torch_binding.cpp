// torch_binding.cpp -- the `_C` extension module: a thin torch layer over the C ABI (include/gsr.h).
//
// Same three entry points, argument order and return tuples as the reference's pybind module
// (DGR/ext.cpp:18-22, DGR/rasterize_points.h:18-67, DGR/rasterize_points.cu:38-237).  Differences:
// a CUDAGuard on means3D's device and the CURRENT torch stream are used (the reference runs on
// the legacy default stream of the current device); gradient outputs are views into ONE flat
// buffer (handy for a single NCCL all-reduce) that the kernels fully overwrite, so no zero-fill.
#include <torch/extension.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <tuple>
#include <stdexcept>
#include "../../include/gsr.h"

namespace {

char* resize_cb(void* user, size_t bytes) {
    auto* t = reinterpret_cast<torch::Tensor*>(user);
    t->resize_({(long long)bytes});
    return reinterpret_cast<char*>(t->data_ptr());
}

const float* fptr(const torch::Tensor& t) {      // empty tensor == "not provided" == NULL
    return t.numel() == 0 ? nullptr : t.data_ptr<float>();
}

torch::Tensor prep(const torch::Tensor& t, const torch::Device& dev, const char* name) {
    if (t.numel() == 0) return t;
    TORCH_CHECK(t.scalar_type() == torch::kFloat32, name, " must be float32");
    TORCH_CHECK(t.device() == dev, name, " must be on ", dev);
    return t.contiguous();
}

void check(int rc) {
    if (rc != GSR_OK) throw std::runtime_error(gsr_last_error());
}

}  // namespace

std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
RasterizeGaussiansCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                       const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                       const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                       const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy, const int image_height,
                       const int image_width, const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                       const bool prefiltered, const bool debug, const int tile_row_begin, const int tile_row_end) {
    if (means3D.ndimension() != 2 || means3D.size(1) != 3) AT_ERROR("means3D must have dimensions (num_points, 3)");
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (there is no CPU path)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const auto dev = means3D.device();
    const int P = means3D.size(0), H = image_height, W = image_width;
    auto fopt = means3D.options().dtype(torch::kFloat32);
    const bool banded = !(tile_row_begin == 0 && tile_row_end == 0);
    // the kernel writes every pixel of the rows it renders; a banded call leaves other rows at 0
    torch::Tensor out_color = (banded || P == 0) ? torch::zeros({3, H, W}, fopt) : torch::empty({3, H, W}, fopt);
    torch::Tensor radii = torch::empty({P}, means3D.options().dtype(torch::kInt32));
    auto bopt = torch::TensorOptions().dtype(torch::kByte).device(dev);
    torch::Tensor geomBuffer = torch::empty({0}, bopt), binningBuffer = torch::empty({0}, bopt), imgBuffer = torch::empty({0}, bopt);
    int rendered = 0;
    if (P != 0) {
        int M = 0;
        if (sh.numel() != 0) M = sh.size(1);
        auto bg = prep(background, dev, "bg"), m3 = prep(means3D, dev, "means3D"), col = prep(colors, dev, "colors_precomp"),
             op = prep(opacity, dev, "opacities"), sc = prep(scales, dev, "scales"), rot = prep(rotations, dev, "rotations"),
             cov = prep(cov3D_precomp, dev, "cov3D_precomp"), vm = prep(viewmatrix, dev, "viewmatrix"),
             pm = prep(projmatrix, dev, "projmatrix"), shc = prep(sh, dev, "shs"), cp = prep(campos, dev, "campos");
        gsr_forward_args a{};
        a.P = P; a.D = degree; a.M = M;
        a.background = fptr(bg); a.width = W; a.height = H;
        a.means3D = fptr(m3); a.shs = fptr(shc); a.colors_precomp = fptr(col); a.opacities = fptr(op);
        a.scales = fptr(sc); a.scale_modifier = scale_modifier; a.rotations = fptr(rot); a.cov3D_precomp = fptr(cov);
        a.viewmatrix = fptr(vm); a.projmatrix = fptr(pm); a.cam_pos = fptr(cp);
        a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy; a.prefiltered = prefiltered;
        a.out_color = out_color.data_ptr<float>(); a.radii = radii.data_ptr<int>();
        a.debug = debug; a.tile_row_begin = tile_row_begin; a.tile_row_end = tile_row_end;
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
        check(gsr_forward(&a, resize_cb, &geomBuffer, resize_cb, &binningBuffer, resize_cb, &imgBuffer, (void*)st, &rendered));
    }
    return std::make_tuple(rendered, out_color, radii, geomBuffer, binningBuffer, imgBuffer);
}

std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
RasterizeGaussiansBackwardCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                               const torch::Tensor& colors, const torch::Tensor& scales, const torch::Tensor& rotations,
                               const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                               const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy,
                               const torch::Tensor& dL_dout_color, const torch::Tensor& sh, const int degree,
                               const torch::Tensor& campos, const torch::Tensor& geomBuffer, const int R,
                               const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, const bool debug,
                               const int tile_row_begin, const int tile_row_end,
                               const c10::optional<torch::Tensor>& drgb_block, const bool skip_dsh, const int64_t drgb_event,
                               const std::vector<int64_t>& chunk_events, const c10::optional<torch::Tensor>& small_pack) {
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (there is no CPU path)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const auto dev = means3D.device();
    const int64_t P = means3D.size(0);
    const int H = dL_dout_color.size(1), W = dL_dout_color.size(2);
    int64_t M = 0;
    if (sh.numel() != 0) M = sh.size(1);
    auto fopt = means3D.options().dtype(torch::kFloat32);
    // one flat gradient arena; per-Gaussian widths: mean2D 3 (never exchanged between ranks: densification takes per-view
    // norms), then mean3D 3, opacity 1, scale 3, rot 4 (the small live blocks, contiguous so that ONE all-reduce covers
    // them in multi-view training), sh 3M, then the blocks training never reads: colour 3, cov3D 6, conic 4
    // Gradients of inputs that were not given (precomputed colours / covariances on the usual SH + scale/rotation path) and
    // dL_dconic (never returned) are neither allocated nor written: 52 bytes per Gaussian less for the per-Gaussian backward.
    const bool want_col = colors.numel() != 0, want_cov = cov3D_precomp.numel() != 0;
    const int64_t widths[8] = {3, 3, 1, 3, 4, 3 * M, want_col ? 3 : 0, want_cov ? 6 : 0};
    int64_t offs[9]; offs[0] = 0;
    for (int i = 0; i < 8; i++) offs[i + 1] = offs[i] + ((widths[i] * P + 3) / 4) * 4;   // keep 16-byte alignment
    torch::Tensor flat = torch::empty({offs[8]}, fopt);
    auto view = [&](int i, std::vector<int64_t> shape) { return flat.narrow(0, offs[i], widths[i] * P).view(shape); };
    torch::Tensor dL_dmeans2D = view(0, {P, 3}), dL_dmeans3D = view(1, {P, 3}), dL_dopacity = view(2, {P, 1}),
                  dL_dscales = view(3, {P, 3}), dL_drotations = view(4, {P, 4}), dL_dsh = view(5, {P, M, 3}),
                  dL_dcolors = want_col ? view(6, {P, 3}) : torch::empty({0}, fopt), dL_dcov3D = want_cov ? view(7, {P, 6}) : torch::empty({0}, fopt);
    if (P != 0) {
        auto bg = prep(background, dev, "bg"), m3 = prep(means3D, dev, "means3D"), col = prep(colors, dev, "colors_precomp"),
             sc = prep(scales, dev, "scales"), rot = prep(rotations, dev, "rotations"), cov = prep(cov3D_precomp, dev, "cov3D_precomp"),
             vm = prep(viewmatrix, dev, "viewmatrix"), pm = prep(projmatrix, dev, "projmatrix"), shc = prep(sh, dev, "shs"),
             cp = prep(campos, dev, "campos"), dpix = prep(dL_dout_color, dev, "dL_dout_color");
        auto rad = radii.contiguous();
        torch::Tensor scratch = torch::empty({(int64_t)gsr_backward_scratch_bytes((int)P)}, torch::TensorOptions().dtype(torch::kByte).device(dev));
        gsr_backward_args a{};
        a.P = (int)P; a.D = degree; a.M = (int)M; a.R = R;
        a.background = fptr(bg); a.width = W; a.height = H;
        a.means3D = fptr(m3); a.shs = fptr(shc); a.colors_precomp = fptr(col); a.scales = fptr(sc);
        a.scale_modifier = scale_modifier; a.rotations = fptr(rot); a.cov3D_precomp = fptr(cov);
        a.viewmatrix = fptr(vm); a.projmatrix = fptr(pm); a.cam_pos = fptr(cp);
        a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
        a.radii = rad.data_ptr<int>();
        a.geom_buffer = reinterpret_cast<const char*>(geomBuffer.data_ptr());
        a.binning_buffer = reinterpret_cast<const char*>(binningBuffer.data_ptr());
        a.image_buffer = reinterpret_cast<const char*>(imageBuffer.data_ptr());
        a.dL_dpix = fptr(dpix);
        a.dL_dmean2D = dL_dmeans2D.data_ptr<float>(); a.dL_dconic = nullptr;
        a.dL_dopacity = dL_dopacity.data_ptr<float>(); a.dL_dcolor = want_col ? dL_dcolors.data_ptr<float>() : nullptr;
        a.dL_dmean3D = dL_dmeans3D.data_ptr<float>(); a.dL_dcov3D = want_cov ? dL_dcov3D.data_ptr<float>() : nullptr;
        a.dL_dsh = M > 0 ? dL_dsh.data_ptr<float>() : nullptr;
        a.dL_dscale = dL_dscales.data_ptr<float>(); a.dL_drot = dL_drotations.data_ptr<float>();
        a.grad_scratch = reinterpret_cast<float*>(scratch.data_ptr());
        a.debug = debug; a.tile_row_begin = tile_row_begin; a.tile_row_end = tile_row_end;
        if (drgb_block.has_value() && drgb_block->numel() > 0) {      // multi-view exchange block: campos (4) | dL_drgb (P,3)
            TORCH_CHECK(drgb_block->scalar_type() == torch::kFloat32 && drgb_block->is_contiguous() && drgb_block->device() == dev &&
                        drgb_block->numel() >= 4 + 3 * P, "drgb_block must be a contiguous float32 tensor of >= 4 + 3P elements on ", dev);
            a.dL_drgb = drgb_block->data_ptr<float>() + 4;
            a.skip_dL_dsh = skip_dsh ? 1 : 0;
            a.drgb_ready_event = reinterpret_cast<void*>(drgb_event);      // cudaEvent_t (torch.cuda.Event.cuda_event) or 0
        }
        void* chunk_ev[GSR_MAX_CHUNKS];
        if (!chunk_events.empty()) {
            TORCH_CHECK((int)chunk_events.size() <= GSR_MAX_CHUNKS, "at most ", GSR_MAX_CHUNKS, " chunk events");
            for (size_t i = 0; i < chunk_events.size(); i++) chunk_ev[i] = reinterpret_cast<void*>(chunk_events[i]);
            a.chunk_done_events = chunk_ev;
            a.n_chunk_events = (int)chunk_events.size();
        }
        if (small_pack.has_value() && small_pack->numel() > 0) {      // (P,14) rows instead of the five small tensors
            TORCH_CHECK(small_pack->scalar_type() == torch::kFloat32 && small_pack->is_contiguous() && small_pack->device() == dev &&
                        small_pack->numel() >= GSR_SMALL_PACK_FLOATS * P, "small_pack must be a contiguous float32 tensor of >= 14P elements on ", dev);
            a.small_pack = small_pack->data_ptr<float>();
        }
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
        check(gsr_backward(&a, (void*)st));
    }
    return std::make_tuple(dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations);
}

// ---- raw-parameter path (include/gsr.h: raw_activations / features_dc / features_rest) ----
// The model's tensors as GaussianModel stores them (scene/gaussian_model.py:44-50): _xyz, _features_dc (P,1,3),
// _features_rest (P,M-1,3), _opacity (logits), _scaling (log), _rotation (unnormalised).  Same outputs as
// rasterize_gaussians; the backward returns the gradients w.r.t. these raw tensors.
std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
RasterizeGaussiansRawCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& features_dc,
                          const torch::Tensor& features_rest, const torch::Tensor& raw_opacity, const torch::Tensor& raw_scaling,
                          const torch::Tensor& raw_rotation, const float scale_modifier, const torch::Tensor& viewmatrix,
                          const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy, const int image_height,
                          const int image_width, const int degree, const torch::Tensor& campos, const bool prefiltered, const bool debug) {
    if (means3D.ndimension() != 2 || means3D.size(1) != 3) AT_ERROR("means3D must have dimensions (num_points, 3)");
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (there is no CPU path)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const auto dev = means3D.device();
    const int64_t P = means3D.size(0);
    const int H = image_height, W = image_width;
    TORCH_CHECK(features_dc.dim() == 3 && features_dc.size(0) == P && features_dc.size(1) == 1 && features_dc.size(2) == 3, "features_dc must be (P,1,3)");
    TORCH_CHECK(features_rest.dim() == 3 && features_rest.size(0) == P && features_rest.size(2) == 3, "features_rest must be (P,M-1,3)");
    TORCH_CHECK(raw_opacity.numel() == P && raw_scaling.numel() == 3 * P && raw_rotation.numel() == 4 * P, "opacity (P,1), scaling (P,3), rotation (P,4)");
    auto fopt = means3D.options().dtype(torch::kFloat32);
    torch::Tensor out_color = P == 0 ? torch::zeros({3, H, W}, fopt) : torch::empty({3, H, W}, fopt);
    torch::Tensor radii = torch::empty({P}, means3D.options().dtype(torch::kInt32));
    auto bopt = torch::TensorOptions().dtype(torch::kByte).device(dev);
    torch::Tensor geomBuffer = torch::empty({0}, bopt), binningBuffer = torch::empty({0}, bopt), imgBuffer = torch::empty({0}, bopt);
    int rendered = 0;
    if (P != 0) {
        auto bg = prep(background, dev, "bg"), m3 = prep(means3D, dev, "means3D"), dc = prep(features_dc, dev, "features_dc"),
             rest = prep(features_rest, dev, "features_rest"), op = prep(raw_opacity, dev, "opacity"), sc = prep(raw_scaling, dev, "scaling"),
             rot = prep(raw_rotation, dev, "rotation"), vm = prep(viewmatrix, dev, "viewmatrix"), pm = prep(projmatrix, dev, "projmatrix"),
             cp = prep(campos, dev, "campos");
        gsr_forward_args a{};
        a.P = (int)P; a.D = degree; a.M = 1 + (int)features_rest.size(1);
        a.background = fptr(bg); a.width = W; a.height = H;
        a.means3D = fptr(m3); a.opacities = fptr(op); a.scales = fptr(sc); a.scale_modifier = scale_modifier; a.rotations = fptr(rot);
        a.viewmatrix = fptr(vm); a.projmatrix = fptr(pm); a.cam_pos = fptr(cp);
        a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy; a.prefiltered = prefiltered;
        a.out_color = out_color.data_ptr<float>(); a.radii = radii.data_ptr<int>(); a.debug = debug;
        a.raw_activations = 1; a.features_dc = fptr(dc); a.features_rest = fptr(rest);
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
        check(gsr_forward(&a, resize_cb, &geomBuffer, resize_cb, &binningBuffer, resize_cb, &imgBuffer, (void*)st, &rendered));
    }
    return std::make_tuple(rendered, out_color, radii, geomBuffer, binningBuffer, imgBuffer);
}

// -> (dL_dmeans2D, dL_dopacity, dL_dmeans3D, dL_dfeatures_dc, dL_dfeatures_rest, dL_dscaling, dL_drotation), all w.r.t. the raw tensors
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
RasterizeGaussiansRawBackwardCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                                  const torch::Tensor& features_dc, const torch::Tensor& features_rest, const torch::Tensor& raw_scaling,
                                  const torch::Tensor& raw_rotation, const float scale_modifier, const torch::Tensor& viewmatrix,
                                  const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy,
                                  const torch::Tensor& dL_dout_color, const int degree, const torch::Tensor& campos,
                                  const torch::Tensor& geomBuffer, const int R, const torch::Tensor& binningBuffer,
                                  const torch::Tensor& imageBuffer, const bool debug) {
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (there is no CPU path)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const auto dev = means3D.device();
    const int64_t P = means3D.size(0);
    const int H = dL_dout_color.size(1), W = dL_dout_color.size(2);
    const int64_t Mr = features_rest.size(1);
    auto fopt = means3D.options().dtype(torch::kFloat32);
    // one flat arena as in rasterize_gaussians_backward: the live blocks first, then the ones training never reads
    // (the gradients of precomputed colours / covariances and dL_dconic have no reader on this path: not allocated, not written)
    const int64_t widths[7] = {3, 3, 1, 3, 4, 3, 3 * Mr};
    int64_t offs[8]; offs[0] = 0;
    for (int i = 0; i < 7; i++) offs[i + 1] = offs[i] + ((widths[i] * P + 3) / 4) * 4;
    torch::Tensor flat = torch::empty({offs[7]}, fopt);
    auto view = [&](int i, std::vector<int64_t> shape) { return flat.narrow(0, offs[i], widths[i] * P).view(shape); };
    torch::Tensor dL_dmeans3D = view(0, {P, 3}), dL_dmeans2D = view(1, {P, 3}), dL_dopacity = view(2, {P, 1}),
                  dL_dscales = view(3, {P, 3}), dL_drotations = view(4, {P, 4}), dL_ddc = view(5, {P, 1, 3}),
                  dL_drest = view(6, {P, Mr, 3});
    if (P != 0) {
        auto bg = prep(background, dev, "bg"), m3 = prep(means3D, dev, "means3D"), dc = prep(features_dc, dev, "features_dc"),
             rest = prep(features_rest, dev, "features_rest"), sc = prep(raw_scaling, dev, "scaling"), rot = prep(raw_rotation, dev, "rotation"),
             vm = prep(viewmatrix, dev, "viewmatrix"), pm = prep(projmatrix, dev, "projmatrix"), cp = prep(campos, dev, "campos"),
             dpix = prep(dL_dout_color, dev, "dL_dout_color");
        auto rad = radii.contiguous();
        torch::Tensor scratch = torch::empty({(int64_t)gsr_backward_scratch_bytes((int)P)}, torch::TensorOptions().dtype(torch::kByte).device(dev));
        gsr_backward_args a{};
        a.P = (int)P; a.D = degree; a.M = 1 + (int)Mr; a.R = R;
        a.background = fptr(bg); a.width = W; a.height = H;
        a.means3D = fptr(m3); a.scales = fptr(sc); a.scale_modifier = scale_modifier; a.rotations = fptr(rot);
        a.viewmatrix = fptr(vm); a.projmatrix = fptr(pm); a.cam_pos = fptr(cp);
        a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
        a.radii = rad.data_ptr<int>();
        a.geom_buffer = reinterpret_cast<const char*>(geomBuffer.data_ptr());
        a.binning_buffer = reinterpret_cast<const char*>(binningBuffer.data_ptr());
        a.image_buffer = reinterpret_cast<const char*>(imageBuffer.data_ptr());
        a.dL_dpix = fptr(dpix);
        a.dL_dmean2D = dL_dmeans2D.data_ptr<float>(); a.dL_dconic = nullptr;
        a.dL_dopacity = dL_dopacity.data_ptr<float>(); a.dL_dcolor = nullptr;
        a.dL_dmean3D = dL_dmeans3D.data_ptr<float>(); a.dL_dcov3D = nullptr;
        a.dL_dscale = dL_dscales.data_ptr<float>(); a.dL_drot = dL_drotations.data_ptr<float>();
        a.grad_scratch = reinterpret_cast<float*>(scratch.data_ptr());
        a.debug = debug;
        a.raw_activations = 1; a.features_dc = fptr(dc); a.features_rest = fptr(rest);
        a.dL_dfeatures_dc = dL_ddc.data_ptr<float>(); a.dL_dfeatures_rest = Mr > 0 ? dL_drest.data_ptr<float>() : nullptr;
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
        check(gsr_backward(&a, (void*)st));
    }
    return std::make_tuple(dL_dmeans2D, dL_dopacity, dL_dmeans3D, dL_ddc, dL_drest, dL_dscales, dL_drotations);
}

torch::Tensor markVisible(torch::Tensor& means3D, torch::Tensor& viewmatrix, torch::Tensor& projmatrix) {
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (there is no CPU path)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = means3D.size(0);
    torch::Tensor present = torch::full({P}, false, means3D.options().dtype(at::kBool));
    if (P != 0) {
        auto m3 = means3D.contiguous(), vm = viewmatrix.contiguous(), pm = projmatrix.contiguous();
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(means3D.device().index()).stream();
        check(gsr_mark_visible(P, m3.data_ptr<float>(), vm.data_ptr<float>(), pm.data_ptr<float>(),
                               reinterpret_cast<unsigned char*>(present.data_ptr<bool>()), (void*)st));
    }
    return present;
}

// Fused SH-gradient exchange for multi-view training (gsr_multiview_sh_grad, include/gsr.h).  peer_ptrs are the device
// addresses of every rank's exchange block (campos (4) | dL_drgb (P,3)) as mapped into THIS process (symmetric memory).
void multiviewShGrad(const std::vector<int64_t>& peer_ptrs, const torch::Tensor& means3D, torch::Tensor& dL_dsh, const int degree) {
    TORCH_CHECK(means3D.is_cuda() && dL_dsh.is_cuda() && dL_dsh.is_contiguous() && dL_dsh.dim() == 3, "bad tensors");
    const c10::cuda::CUDAGuard guard(means3D.device());
    std::vector<const float*> ptrs;
    for (auto p : peer_ptrs) ptrs.push_back(reinterpret_cast<const float*>(p));
    auto m3 = means3D.contiguous();
    cudaStream_t st = c10::cuda::getCurrentCUDAStream(means3D.device().index()).stream();
    check(gsr_multiview_sh_grad((int)means3D.size(0), degree, (int)dL_dsh.size(1), (int)ptrs.size(), ptrs.data(), m3.data_ptr<float>(),
                                dL_dsh.data_ptr<float>(), (void*)st));
}

// One range of Gaussians of the same exchange (gsr_multiview_sh_grad_range): block pointers may be virtual bases.
void multiviewShGradRange(const std::vector<int64_t>& peer_ptrs, const torch::Tensor& campos, const torch::Tensor& means3D, torch::Tensor& dL_dsh,
                          const int degree, const int64_t begin, const int64_t end) {
    TORCH_CHECK(means3D.is_cuda() && dL_dsh.is_cuda() && dL_dsh.is_contiguous() && dL_dsh.dim() == 3, "bad tensors");
    TORCH_CHECK(campos.is_cuda() && campos.is_contiguous() && campos.scalar_type() == torch::kFloat32 && campos.numel() >= 4 * (int64_t)peer_ptrs.size(),
                "campos must be a contiguous CUDA float32 (n_views,4) tensor");
    const c10::cuda::CUDAGuard guard(means3D.device());
    std::vector<const float*> ptrs;
    for (auto p : peer_ptrs) ptrs.push_back(reinterpret_cast<const float*>(p));
    auto m3 = means3D.contiguous();
    cudaStream_t st = c10::cuda::getCurrentCUDAStream(means3D.device().index()).stream();
    check(gsr_multiview_sh_grad_range((int)begin, (int)end, degree, (int)dL_dsh.size(1), (int)ptrs.size(), ptrs.data(), campos.data_ptr<float>(),
                                      m3.data_ptr<float>(), dL_dsh.data_ptr<float>(), (void*)st));
}

// introspection for the parity tests: (point_list, ranges, n_contrib, final_T, packed, tiles_touched, clamped) as tensors
std::vector<torch::Tensor> viewBuffers(int P, int R, int W, int H, const torch::Tensor& geomBuffer,
                                       const torch::Tensor& binningBuffer, const torch::Tensor& imgBuffer) {
    gsr_buffer_view v{};
    check(gsr_view_buffers(P, R, W, H, reinterpret_cast<const char*>(geomBuffer.data_ptr()),
                           reinterpret_cast<const char*>(binningBuffer.data_ptr()),
                           reinterpret_cast<const char*>(imgBuffer.data_ptr()), &v));
    const int tiles = ((W + 15) / 16) * ((H + 15) / 16);
    auto dev = geomBuffer.device();
    auto mk = [&](const void* p, std::vector<int64_t> shape, torch::ScalarType dt) {
        return torch::from_blob(const_cast<void*>(p), shape, torch::TensorOptions().dtype(dt).device(dev)).clone();
    };
    return {mk(v.point_list, {R}, torch::kInt32), mk(v.ranges, {tiles, 2}, torch::kInt32), mk(v.n_contrib, {H, W}, torch::kInt32),
            mk(v.final_T, {H, W}, torch::kFloat32), mk(v.packed, {P, 12}, torch::kFloat32), mk(v.tiles_touched, {P}, torch::kInt32),
            mk(v.clamped, {P}, torch::kUInt8), mk(v.warp_mask, {R}, torch::kUInt8)};
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    namespace py = pybind11;
    m.def("rasterize_gaussians", &RasterizeGaussiansCUDA, py::arg("background"), py::arg("means3D"), py::arg("colors"),
          py::arg("opacity"), py::arg("scales"), py::arg("rotations"), py::arg("scale_modifier"), py::arg("cov3D_precomp"),
          py::arg("viewmatrix"), py::arg("projmatrix"), py::arg("tan_fovx"), py::arg("tan_fovy"), py::arg("image_height"),
          py::arg("image_width"), py::arg("sh"), py::arg("degree"), py::arg("campos"), py::arg("prefiltered"), py::arg("debug"),
          py::arg("tile_row_begin") = 0, py::arg("tile_row_end") = 0);
    m.def("rasterize_gaussians_backward", &RasterizeGaussiansBackwardCUDA, py::arg("background"), py::arg("means3D"),
          py::arg("radii"), py::arg("colors"), py::arg("scales"), py::arg("rotations"), py::arg("scale_modifier"),
          py::arg("cov3D_precomp"), py::arg("viewmatrix"), py::arg("projmatrix"), py::arg("tan_fovx"), py::arg("tan_fovy"),
          py::arg("dL_dout_color"), py::arg("sh"), py::arg("degree"), py::arg("campos"), py::arg("geomBuffer"), py::arg("R"),
          py::arg("binningBuffer"), py::arg("imageBuffer"), py::arg("debug"), py::arg("tile_row_begin") = 0,
          py::arg("tile_row_end") = 0, py::arg("drgb_block") = py::none(), py::arg("skip_dsh") = false,
          py::arg("drgb_event") = 0, py::arg("chunk_events") = std::vector<int64_t>(),
          py::arg("small_pack") = py::none());
    m.def("rasterize_gaussians_raw", &RasterizeGaussiansRawCUDA, py::arg("background"), py::arg("means3D"), py::arg("features_dc"),
          py::arg("features_rest"), py::arg("opacity"), py::arg("scaling"), py::arg("rotation"), py::arg("scale_modifier"), py::arg("viewmatrix"),
          py::arg("projmatrix"), py::arg("tan_fovx"), py::arg("tan_fovy"), py::arg("image_height"), py::arg("image_width"), py::arg("degree"),
          py::arg("campos"), py::arg("prefiltered"), py::arg("debug"));
    m.def("rasterize_gaussians_raw_backward", &RasterizeGaussiansRawBackwardCUDA, py::arg("background"), py::arg("means3D"), py::arg("radii"),
          py::arg("features_dc"), py::arg("features_rest"), py::arg("scaling"), py::arg("rotation"), py::arg("scale_modifier"),
          py::arg("viewmatrix"), py::arg("projmatrix"), py::arg("tan_fovx"), py::arg("tan_fovy"), py::arg("dL_dout_color"), py::arg("degree"),
          py::arg("campos"), py::arg("geomBuffer"), py::arg("R"), py::arg("binningBuffer"), py::arg("imageBuffer"), py::arg("debug"));
    m.def("unpack_small_grads", [](int64_t begin, int64_t end, const torch::Tensor& pack, torch::Tensor& dmean3D, torch::Tensor& dmean2D,
                                   torch::Tensor& dop, torch::Tensor& dscale, torch::Tensor& drot) {
        TORCH_CHECK(pack.is_cuda() && pack.is_contiguous() && pack.scalar_type() == torch::kFloat32, "pack: contiguous float32 CUDA tensor");
        for (const torch::Tensor* t : {&dmean3D, &dmean2D, &dop, &dscale, &drot})
            TORCH_CHECK(t->is_cuda() && t->is_contiguous() && t->scalar_type() == torch::kFloat32 && t->device() == pack.device(),
                        "gradient tensors: contiguous float32 on the pack's device");
        TORCH_CHECK(end <= dmean3D.size(0) && GSR_SMALL_PACK_FLOATS * end <= pack.numel(), "range exceeds the tensors");
        const c10::cuda::CUDAGuard guard(pack.device());
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(pack.device().index()).stream();
        check(gsr_unpack_small_grads((int)begin, (int)end, pack.data_ptr<float>(), dmean3D.data_ptr<float>(), dmean2D.data_ptr<float>(),
                                     dop.data_ptr<float>(), dscale.data_ptr<float>(), drot.data_ptr<float>(), (void*)st));
    }, "rows [begin,end) of a (P,14) small_pack block -> the five gradient tensors (include/gsr.h)");
    m.def("multimem_allreduce", [](int64_t mc_ptr, int64_t begin, int64_t end, int64_t rank, int64_t world, int64_t device) {
        const c10::cuda::CUDAGuard guard(c10::Device(c10::kCUDA, (c10::DeviceIndex)device));
        cudaStream_t st = c10::cuda::getCurrentCUDAStream((c10::DeviceIndex)device).stream();
        check(gsr_multimem_allreduce(reinterpret_cast<void*>(mc_ptr), (size_t)begin, (size_t)end, (int)rank, (int)world, (void*)st));
    }, "two-shot all-reduce of floats [begin,end) through an NVLS multicast address (include/gsr.h)");
    m.def("multimem_broadcast", [](int64_t mc_dst_ptr, const torch::Tensor& src, int64_t n_floats) {
        TORCH_CHECK(src.is_cuda() && src.is_contiguous() && src.scalar_type() == torch::kFloat32 && src.numel() >= n_floats, "src: contiguous CUDA float32");
        const c10::cuda::CUDAGuard guard(src.device());
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(src.device().index()).stream();
        check(gsr_multimem_broadcast(reinterpret_cast<void*>(mc_dst_ptr), src.data_ptr<float>(), (size_t)n_floats, (void*)st));
    }, "store n floats through an NVLS multicast address into every rank's copy (include/gsr.h)");
    m.def("l1_ssim_forward", [](const torch::Tensor& img, const torch::Tensor& gt, double lambda_dssim, bool need_grad) {
        TORCH_CHECK(img.is_cuda() && gt.is_cuda(), "l1_ssim_forward: CUDA tensors only (there is no CPU path)");
        TORCH_CHECK(img.scalar_type() == torch::kFloat32 && gt.scalar_type() == torch::kFloat32, "float32 images");
        TORCH_CHECK(img.sizes() == gt.sizes() && img.dim() >= 3 && img.dim() <= 4 && img.device() == gt.device(), "images must have the same (C,H,W) or (B,C,H,W) shape and device");
        const c10::cuda::CUDAGuard guard(img.device());
        auto a = img.contiguous(), b = gt.contiguous();
        const int H = (int)a.size(-2), W = (int)a.size(-1);
        const int planes = (int)(a.numel() / ((int64_t)H * W));
        auto opt = a.options();
        torch::Tensor maps = need_grad ? torch::empty({3, (int64_t)planes, H, W}, opt) : torch::empty({0}, opt);
        torch::Tensor scratch = torch::empty({(int64_t)gsr_l1_ssim_scratch_floats(planes, H, W)}, opt);
        torch::Tensor out3 = torch::empty({3}, opt);
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(img.device().index()).stream();
        check(gsr_l1_ssim_forward(a.data_ptr<float>(), b.data_ptr<float>(), planes, H, W, (float)lambda_dssim,
                                  need_grad ? maps.data_ptr<float>() : nullptr, scratch.data_ptr<float>(), out3.data_ptr<float>(), (void*)st));
        return std::make_tuple(out3, maps);
    }, "fused L1 + SSIM forward: (loss, l1, ssim) and the derivative maps (include/gsr.h)");
    m.def("l1_ssim_backward", [](const torch::Tensor& img, const torch::Tensor& gt, const torch::Tensor& maps, double coef_l1, double coef_ssim,
                                 const torch::Tensor& grad_out) {
        TORCH_CHECK(img.is_cuda() && gt.is_cuda() && maps.is_cuda() && grad_out.is_cuda(), "l1_ssim_backward: CUDA tensors only");
        TORCH_CHECK(grad_out.numel() == 1 && grad_out.scalar_type() == torch::kFloat32, "grad_out: one float32 element");
        const c10::cuda::CUDAGuard guard(img.device());
        auto a = img.contiguous(), b = gt.contiguous(), g = grad_out.contiguous();
        const int H = (int)a.size(-2), W = (int)a.size(-1);
        const int planes = (int)(a.numel() / ((int64_t)H * W));
        TORCH_CHECK(maps.is_contiguous() && maps.numel() == 3 * a.numel(), "maps: the tensor l1_ssim_forward returned for these images");
        torch::Tensor dimg = torch::empty_like(a);
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(img.device().index()).stream();
        check(gsr_l1_ssim_backward(a.data_ptr<float>(), b.data_ptr<float>(), maps.data_ptr<float>(), planes, H, W, (float)coef_l1, (float)coef_ssim,
                                   g.data_ptr<float>(), dimg.data_ptr<float>(), (void*)st));
        return dimg;
    }, "fused L1 + SSIM backward (include/gsr.h)");
    m.def("adam_step", [](std::vector<torch::Tensor> params, std::vector<torch::Tensor> grads, std::vector<torch::Tensor> exp_avgs,
                          std::vector<torch::Tensor> exp_avg_sqs, std::vector<double> lrs, std::vector<int64_t> steps, double beta1, double beta2,
                          double eps, const c10::optional<torch::Tensor>& visible) {
        const size_t n = params.size();
        TORCH_CHECK(grads.size() == n && exp_avgs.size() == n && exp_avg_sqs.size() == n && lrs.size() == n && steps.size() == n,
                    "adam_step: one grad, exp_avg, exp_avg_sq, lr and step per parameter");
        if (n == 0) return;
        const auto dev = params[0].device();
        TORCH_CHECK(dev.is_cuda(), "adam_step: CUDA tensors only (there is no CPU path)");
        std::vector<gsr_adam_group> groups(n);
        int64_t rows = -1;
        if (visible.has_value()) {
            TORCH_CHECK(visible->is_cuda() && visible->device() == dev && visible->is_contiguous() && visible->dim() == 1 &&
                        (visible->scalar_type() == torch::kBool || visible->scalar_type() == torch::kUInt8),
                        "adam_step: visible must be a contiguous (P,) bool / uint8 tensor on the parameters' device");
            rows = visible->numel();
        }
        for (size_t i = 0; i < n; i++) {
            for (const torch::Tensor* t : {&params[i], &grads[i], &exp_avgs[i], &exp_avg_sqs[i]})
                TORCH_CHECK(t->is_cuda() && t->device() == dev && t->scalar_type() == torch::kFloat32 && t->is_contiguous() &&
                            t->numel() == params[i].numel(), "adam_step: contiguous float32 tensors of the parameter's size on one CUDA device");
            groups[i].param = params[i].data_ptr<float>();
            groups[i].grad = grads[i].data_ptr<float>();
            groups[i].exp_avg = exp_avgs[i].data_ptr<float>();
            groups[i].exp_avg_sq = exp_avg_sqs[i].data_ptr<float>();
            groups[i].n = (size_t)params[i].numel();
            groups[i].lr = lrs[i];
            groups[i].step = (long long)steps[i];
            groups[i].row_width = 1;
            if (rows >= 0) {
                TORCH_CHECK(params[i].dim() >= 1 && params[i].size(0) == rows, "adam_step: with `visible`, every parameter's first dimension is P");
                groups[i].row_width = rows > 0 ? (int)(params[i].numel() / rows) : 1;
            }
        }
        const c10::cuda::CUDAGuard guard(dev);
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
        check(gsr_adam_step(groups.data(), (int)n, beta1, beta2, eps,
                            visible.has_value() ? reinterpret_cast<const unsigned char*>(visible->data_ptr()) : nullptr, (void*)st));
    }, "one-launch Adam step over all parameter groups, in place (include/gsr.h: gsr_adam_step)",
          py::arg("params"), py::arg("grads"), py::arg("exp_avgs"), py::arg("exp_avg_sqs"), py::arg("lrs"), py::arg("steps"), py::arg("beta1"),
          py::arg("beta2"), py::arg("eps"), py::arg("visible") = py::none());
    m.def("dist_cuda2", [](const torch::Tensor& points) {
        TORCH_CHECK(points.is_cuda(), "distCUDA2: CUDA tensor only (there is no CPU path)");
        TORCH_CHECK(points.scalar_type() == torch::kFloat32 && points.dim() == 2 && points.size(1) == 3, "distCUDA2: points must be (P,3) float32");
        const c10::cuda::CUDAGuard guard(points.device());
        auto pts = points.contiguous();
        const int64_t P = pts.size(0);
        TORCH_CHECK(P <= 0x7fffffffll, "distCUDA2: too many points");
        torch::Tensor means = torch::full({P}, 0.0, pts.options());                       // spatial.cu:21
        if (P == 0) return means;
        torch::Tensor scratch = torch::empty({(int64_t)gsr_knn_scratch_bytes((int)P)}, pts.options().dtype(torch::kUInt8));
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(points.device().index()).stream();
        check(gsr_knn_mean_dist2((int)P, pts.data_ptr<float>(), means.data_ptr<float>(), reinterpret_cast<char*>(scratch.data_ptr()), (void*)st));
        return means;
    }, "mean squared distance to the 3 nearest neighbours (include/gsr.h: gsr_knn_mean_dist2; simple-knn's distCUDA2)");
    m.def("densify_stats", [](const torch::Tensor& viewspace_grad, const torch::Tensor& radii, torch::Tensor& accum, torch::Tensor& denom,
                              const c10::optional<torch::Tensor>& max_radii2D) {
        TORCH_CHECK(viewspace_grad.is_cuda(), "densify_stats: CUDA tensors only (there is no CPU path)");
        const auto dev = viewspace_grad.device();
        const int64_t P = radii.numel();
        TORCH_CHECK(viewspace_grad.scalar_type() == torch::kFloat32 && viewspace_grad.is_contiguous() && viewspace_grad.dim() == 2 &&
                    viewspace_grad.size(0) == P && viewspace_grad.size(1) == 3, "densify_stats: viewspace gradient must be contiguous (P,3) float32");
        TORCH_CHECK(radii.device() == dev && radii.scalar_type() == torch::kInt32 && radii.is_contiguous(), "densify_stats: radii must be contiguous (P,) int32");
        for (const torch::Tensor* t : {&accum, &denom})
            TORCH_CHECK(t->device() == dev && t->scalar_type() == torch::kFloat32 && t->is_contiguous() && t->numel() == P,
                        "densify_stats: xyz_gradient_accum and denom must be contiguous float32 with P elements");
        float* mr = nullptr;
        if (max_radii2D.has_value()) {
            TORCH_CHECK(max_radii2D->device() == dev && max_radii2D->scalar_type() == torch::kFloat32 && max_radii2D->is_contiguous() &&
                        max_radii2D->numel() == P, "densify_stats: max_radii2D must be contiguous (P,) float32");
            mr = max_radii2D->data_ptr<float>();
        }
        const c10::cuda::CUDAGuard guard(dev);
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
        check(gsr_densify_stats((int)P, viewspace_grad.data_ptr<float>(), radii.data_ptr<int>(), accum.data_ptr<float>(), denom.data_ptr<float>(), mr, (void*)st));
    }, "per-iteration densification statistics, in place (include/gsr.h: gsr_densify_stats)",
          py::arg("viewspace_grad"), py::arg("radii"), py::arg("xyz_gradient_accum"), py::arg("denom"), py::arg("max_radii2D") = py::none());
    m.def("densify_classify", [](const torch::Tensor& accum, const torch::Tensor& denom, const torch::Tensor& scaling, const torch::Tensor& opacity,
                                 double max_grad, double min_opacity, double extent, double percent_dense, double max_screen_size, int64_t copies) {
        TORCH_CHECK(scaling.is_cuda(), "densify_classify: CUDA tensors only (there is no CPU path)");
        const auto dev = scaling.device();
        const int64_t P = scaling.size(0);
        TORCH_CHECK(scaling.dim() == 2 && scaling.size(1) == 3, "densify_classify: scaling must be (P,3)");
        for (const torch::Tensor* t : {&accum, &denom, &scaling, &opacity})
            TORCH_CHECK(t->device() == dev && t->scalar_type() == torch::kFloat32 && t->is_contiguous(), "densify_classify: contiguous float32 tensors on one CUDA device");
        TORCH_CHECK(accum.numel() == P && denom.numel() == P && opacity.numel() == P, "densify_classify: statistics and opacity must have P elements");
        const c10::cuda::CUDAGuard guard(dev);
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
        auto bytes = scaling.options().dtype(torch::kUInt8);
        torch::Tensor flags = torch::empty({P}, bytes);
        torch::Tensor scratch = torch::empty({(int64_t)gsr_densify_scratch_bytes((int)P)}, bytes);
        torch::Tensor totals = torch::empty({5}, scaling.options().dtype(torch::kInt32));
        gsr_densify_params prm{max_grad, min_opacity, extent, percent_dense, max_screen_size, (int)copies};
        check(gsr_densify_classify((int)P, accum.data_ptr<float>(), denom.data_ptr<float>(), scaling.data_ptr<float>(), opacity.data_ptr<float>(), &prm,
                                   flags.data_ptr<uint8_t>(), totals.data_ptr<int>(), reinterpret_cast<char*>(scratch.data_ptr()), (void*)st));
        return std::make_tuple(flags, totals, scratch);
    }, "densify_and_prune, first call: (flags, totals, scratch) (include/gsr.h: gsr_densify_classify)");
    m.def("densify_apply", [](const torch::Tensor& flags, const torch::Tensor& scratch, std::vector<int64_t> totals, const torch::Tensor& samples,
                              std::vector<torch::Tensor> params, std::vector<c10::optional<torch::Tensor>> exp_avgs,
                              std::vector<c10::optional<torch::Tensor>> exp_avg_sqs, int64_t xyz_group, int64_t scaling_group, int64_t rotation_group,
                              double max_grad, double min_opacity, double extent, double percent_dense, double max_screen_size, int64_t copies) {
        const size_t n = params.size();
        TORCH_CHECK(n >= 1 && exp_avgs.size() == n && exp_avg_sqs.size() == n && totals.size() == 5, "densify_apply: one exp_avg / exp_avg_sq entry per parameter, five totals");
        TORCH_CHECK(flags.is_cuda(), "densify_apply: CUDA tensors only (there is no CPU path)");
        const auto dev = flags.device();
        const int64_t P = flags.numel(), P_out = totals[0];
        TORCH_CHECK(samples.device() == dev && samples.scalar_type() == torch::kFloat32 && samples.is_contiguous() && samples.numel() == 3 * copies * totals[1],
                    "densify_apply: samples must be contiguous (N * n_split, 3) float32");
        const c10::cuda::CUDAGuard guard(dev);
        std::vector<gsr_densify_group> groups(n);
        std::vector<torch::Tensor> out_p(n);
        std::vector<c10::optional<torch::Tensor>> out_m(n), out_v(n);
        for (size_t i = 0; i < n; i++) {
            const torch::Tensor& p = params[i];
            TORCH_CHECK(p.device() == dev && p.scalar_type() == torch::kFloat32 && p.is_contiguous() && p.dim() >= 1 && p.size(0) == P,
                        "densify_apply: parameters must be contiguous float32 with P rows");
            auto shape = p.sizes().vec();
            shape[0] = P_out;
            out_p[i] = torch::empty(shape, p.options());
            groups[i] = gsr_densify_group{p.data_ptr<float>(), nullptr, nullptr, out_p[i].data_ptr<float>(), nullptr, nullptr, (int)(P > 0 ? p.numel() / P : 1)};
            if (P == 0) { int w = 1; for (size_t d = 1; d < shape.size(); d++) w *= (int)shape[d]; groups[i].row_width = w; }
            TORCH_CHECK(exp_avgs[i].has_value() == exp_avg_sqs[i].has_value(), "densify_apply: moments come in pairs");
            if (exp_avgs[i].has_value()) {
                for (const torch::Tensor* t : {&*exp_avgs[i], &*exp_avg_sqs[i]})
                    TORCH_CHECK(t->device() == dev && t->scalar_type() == torch::kFloat32 && t->is_contiguous() && t->sizes() == p.sizes(),
                                "densify_apply: moments must match their parameter");
                out_m[i] = torch::empty(shape, p.options());
                out_v[i] = torch::empty(shape, p.options());
                groups[i].exp_avg = exp_avgs[i]->data_ptr<float>();
                groups[i].exp_avg_sq = exp_avg_sqs[i]->data_ptr<float>();
                groups[i].exp_avg_out = out_m[i]->data_ptr<float>();
                groups[i].exp_avg_sq_out = out_v[i]->data_ptr<float>();
            }
        }
        int th[5];
        for (int i = 0; i < 5; i++) th[i] = (int)totals[i];
        gsr_densify_params prm{max_grad, min_opacity, extent, percent_dense, max_screen_size, (int)copies};
        cudaStream_t st = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
        check(gsr_densify_apply((int)P, flags.data_ptr<uint8_t>(), reinterpret_cast<const char*>(scratch.data_ptr()), th, &prm,
                                samples.numel() ? samples.data_ptr<float>() : nullptr, groups.data(), (int)n, (int)xyz_group, (int)scaling_group,
                                (int)rotation_group, (void*)st));
        return std::make_tuple(out_p, out_m, out_v);
    }, "densify_and_prune, second call: new (params, exp_avgs, exp_avg_sqs) (include/gsr.h: gsr_densify_apply)");
    m.def("backward_chunk_plan", [](int64_t P) {
        int b[GSR_MAX_CHUNKS + 1];
        const int n = gsr_backward_chunk_plan((int)P, b);
        return std::vector<int64_t>(b, b + n + 1);
    }, "Gaussian ranges of the chunked per-Gaussian backward (include/gsr.h: gsr_backward_chunk_plan)");
    m.def("multiview_sh_grad", &multiviewShGrad);
    m.def("multiview_sh_grad_range", &multiviewShGradRange);
    m.def("mark_visible", &markVisible);
    m.def("view_buffers", &viewBuffers);
}
