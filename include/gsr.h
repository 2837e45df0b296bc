/*
 * gsr.h -- C ABI of the B200-native differentiable Gaussian-splatting rasterizer.
 *
 * This is the drop-in boundary underneath the reference's torch binding.  Each entry point
 * replaces one static method of CudaRasterizer::Rasterizer
 * (DGR = /root/reference/submodules/diff-gaussian-rasterization):
 *
 *   gsr_forward        <- Rasterizer::forward      DGR/cuda_rasterizer/rasterizer.h:31-55,
 *                                                  impl DGR/cuda_rasterizer/rasterizer_impl.cu:227-411
 *   gsr_backward       <- Rasterizer::backward     DGR/cuda_rasterizer/rasterizer.h:57-84,
 *                                                  impl DGR/cuda_rasterizer/rasterizer_impl.cu:416-518
 *   gsr_mark_visible   <- Rasterizer::markVisible  DGR/cuda_rasterizer/rasterizer.h:24-29,
 *                                                  impl DGR/cuda_rasterizer/rasterizer_impl.cu:162-174
 *   gsr_resize_fn      <- std::function<char*(size_t)> geometryBuffer/binningBuffer/imageBuffer
 *                                                  DGR/cuda_rasterizer/rasterizer.h:32-34,
 *                                                  made by resizeFunctional, DGR/rasterize_points.cu:28-36
 *
 * Plain C types only: device pointers, sizes, a cudaStream_t passed as void*.  All float
 * arrays are fp32, contiguous, on the current CUDA device.  "Not provided" optional inputs are
 * NULL, exactly as the reference's kernels test them (forward.cu:252,306; backward.cu:393,397).
 * Every function returns 0 on success and a non-zero gsr_status otherwise; gsr_last_error()
 * gives the text (thread-local).  Nothing here falls back to the CPU.
 */
#ifndef GSR_H_INCLUDED
#define GSR_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_VERSION 1

typedef enum gsr_status {
    GSR_OK = 0,
    GSR_ERR_INVALID_ARGUMENT = 1, /* bad sizes / missing required pointer */
    GSR_ERR_CUDA = 2,             /* a CUDA runtime call or (debug mode) a kernel failed */
    GSR_ERR_ALLOC = 3             /* a resize callback returned NULL */
} gsr_status;

/* Growable scratch buffer: called once per forward with the number of bytes needed; must return
 * a device pointer (>= 16-byte aligned) that stays valid until the matching backward is done. */
typedef char* (*gsr_resize_fn)(void* user, size_t bytes);

typedef struct gsr_forward_args {
    int P;                       /* number of Gaussians */
    int D;                       /* active SH degree 0..3 */
    int M;                       /* SH coefficients per Gaussian in `shs` (0 if shs == NULL) */
    const float* background;     /* (3) */
    int width, height;
    const float* means3D;        /* (P,3) */
    const float* shs;            /* (P,M,3) or NULL */
    const float* colors_precomp; /* (P,3) or NULL */
    const float* opacities;      /* (P,1) */
    const float* scales;         /* (P,3) or NULL */
    float scale_modifier;
    const float* rotations;      /* (P,4) or NULL */
    const float* cov3D_precomp;  /* (P,6) or NULL */
    const float* viewmatrix;     /* (4,4), flat index 4*col+row as auxiliary.h:60-79 reads it */
    const float* projmatrix;     /* (4,4) */
    const float* cam_pos;        /* (3) */
    float tan_fovx, tan_fovy;
    int prefiltered;
    float* out_color;            /* (3,H,W) out; rows outside the tile-row band are left untouched */
    int* radii;                  /* (P) out, may be NULL */
    int debug;                   /* sync + check after every stage (CHECK_CUDA, auxiliary.h:177-184) */
    /* Extension (not in the reference): restrict binning + blending to tile rows
     * [tile_row_begin, tile_row_end) for tile-row sharded multi-GPU rendering.  0,0 = all rows.
     * radii are computed on the full image either way. */
    int tile_row_begin, tile_row_end;
    /* Extension (not in the reference): the model's RAW parameters, as GaussianModel stores them,
     * instead of the activated tensors gaussian_renderer/__init__.py:69-101 builds with separate torch kernels.
     * raw_activations != 0: `opacities` holds logits (sigmoid applied here, scene/gaussian_model.py:128-130), `scales`
     *   log-scales (exp, :114-117), `rotations` unnormalised quaternions (F.normalize, eps 1e-12, :119-122).
     * features_dc != NULL: SH coefficients in split storage -- features_dc (P,1,3) and features_rest (P,M-1,3), the two
     *   tensors `get_features` concatenates (:132-138); then shs and colors_precomp must be NULL and M is the total
     *   coefficient count.  The two options are independent.  Zero / NULL = the reference's interface. */
    int raw_activations;
    const float* features_dc;
    const float* features_rest;
} gsr_forward_args;

typedef struct gsr_backward_args {
    int P, D, M, R;              /* R = num_rendered returned by the forward */
    const float* background;
    int width, height;
    const float* means3D;
    const float* shs;
    const float* colors_precomp;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* cov3D_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* cam_pos;
    float tan_fovx, tan_fovy;
    const int* radii;            /* (P) as written by the forward */
    const char* geom_buffer;     /* the three buffers the forward filled, unchanged */
    const char* binning_buffer;
    const char* image_buffer;
    const float* dL_dpix;        /* (3,H,W) */
    /* Outputs.  Every element is written (no pre-zeroing needed), so callers may pass
     * uninitialised memory -- the reference needs torch::zeros here (rasterize_points.cu:168-178). */
    float* dL_dmean2D;           /* (P,3) */
    float* dL_dconic;            /* (P,4) = (P,2,2), may be NULL (the reference does not return it) */
    float* dL_dopacity;          /* (P,1) */
    float* dL_dcolor;            /* (P,3); may be NULL when colors_precomp is NULL (then nobody reads it) */
    float* dL_dmean3D;           /* (P,3) */
    float* dL_dcov3D;            /* (P,6); may be NULL when cov3D_precomp is NULL */
    float* dL_dsh;               /* (P,M,3), NULL iff M == 0 */
    float* dL_dscale;            /* (P,3) */
    float* dL_drot;              /* (P,4) */
    /* Scratch: (P,12) floats, 16-byte aligned, used for the per-Gaussian gradient accumulators
     * the blending backward reduces into.  Zeroed internally. */
    float* grad_scratch;
    int debug;
    int tile_row_begin, tile_row_end;
    /* Extension for multi-view training (one camera per GPU): when dL_drgb is not NULL the colour gradient
     * with the SH clamp mask applied -- the (P,3) vector dL_dsh is an outer product of, see
     * gsr_multiview_sh_grad -- is written there; with skip_dL_dsh != 0 the (P,M,3) dL_dsh block is then left
     * untouched (it is produced by gsr_multiview_sh_grad from every rank's dL_drgb instead). */
    float* dL_drgb;
    int skip_dL_dsh;
    /* Optional cudaEvent_t recorded on the stream as soon as dL_drgb is complete -- right after the blending
     * backward, BEFORE the per-Gaussian backward runs -- so that the exchange of dL_drgb between ranks can
     * overlap with the rest of this call.  NULL = no event. */
    void* drgb_ready_event;
    /* Optional cudaEvent_t handles (n_chunk_events of them, at most GSR_MAX_CHUNKS).  With dL_drgb set the
     * per-Gaussian backward runs as gsr_backward_chunk_plan(P) consecutive launches over ranges of Gaussians and
     * chunk_done_events[c] is recorded on the stream after range c (handles beyond the number of ranges are
     * recorded after the last one): the caller can reduce a range's gradients across ranks while the later
     * ranges are still being computed.  NULL / 0 = no events. */
    void* const* chunk_done_events;
    int n_chunk_events;
    /* Optional (P,14) row-major block.  When set, the five small live gradients of a Gaussian -- dL_dmean3D (3),
     * dL_dmean2D (3), dL_dopacity (1), dL_dscale (3), dL_drot (4), in this order -- are written there as ONE
     * 56-byte row INSTEAD of into their five tensors, so that a range of Gaussians is one contiguous span for
     * the cross-rank reduction; gsr_unpack_small_grads then scatters reduced rows into the five tensors. */
    float* small_pack;
    /* Raw-parameter path, as in gsr_forward_args (same values as the forward call).  With raw_activations the gradients
     * dL_dopacity / dL_dscale / dL_drot are w.r.t. the RAW parameters (sigmoid / exp / normalize derivatives applied here).
     * With features_dc set, the SH gradient is written as dL_dfeatures_dc (P,1,3) and dL_dfeatures_rest (P,M-1,3) -- what
     * torch.cat's backward would copy out of dL_dsh -- and shs / dL_dsh must be NULL; M <= 16.  Not combined with dL_drgb /
     * small_pack. */
    int raw_activations;
    const float* features_dc;
    const float* features_rest;
    float* dL_dfeatures_dc;
    float* dL_dfeatures_rest;
} gsr_backward_args;
#define GSR_SMALL_PACK_FLOATS 14

/* The ranges of Gaussians the chunked per-Gaussian backward uses for P Gaussians: bounds[0] = 0 <= ... <=
 * bounds[n] = P, n returned (1..GSR_MAX_CHUNKS); bounds needs GSR_MAX_CHUNKS + 1 ints. */
#define GSR_MAX_CHUNKS 4
int gsr_backward_chunk_plan(int P, int* bounds);

/* Bytes needed for gsr_backward_args.grad_scratch */
size_t gsr_backward_scratch_bytes(int P);

/* Sizes of the three opaque buffers (what the resize callbacks will be asked for).
 * kind: 0 = geometry (n = P), 1 = binning (n = R), 2 = image (n = width*height, aux = #tiles). */
size_t gsr_required_bytes(int kind, size_t n, size_t aux);

int gsr_forward(const gsr_forward_args* args,
                gsr_resize_fn geometry_buffer, void* geometry_user,
                gsr_resize_fn binning_buffer, void* binning_user,
                gsr_resize_fn image_buffer, void* image_user,
                void* cuda_stream, int* num_rendered);

int gsr_backward(const gsr_backward_args* args, void* cuda_stream);

/* Fused gradient exchange for the SH block in multi-view training (extension, no reference counterpart).
 * Per view v the SH gradient is an outer product  dL_dsh_v[k] = basis_k(dir_v) * dL_drgb_v  (backward.cu:60-128),
 * so instead of all-reducing 12*M bytes per Gaussian each rank publishes 12 bytes (dL_drgb_v) plus its camera
 * position in peer-mapped memory and this ONE kernel reads every rank's block over NVLink (plain peer loads),
 * rebuilds the basis per view and writes the summed dL_dsh (P,M,3), coefficients >= (D+1)^2 zeroed.
 * peer_blocks[v] points to rank v's block: 4 floats (campos x,y,z,pad) followed by (P,3) floats dL_drgb.
 * All pointers must be 16-byte aligned.  The caller orders it after every rank's gsr_backward (device barrier). */
#define GSR_MAX_VIEWS 16
int gsr_multiview_sh_grad(int P, int D, int M, int n_views, const float* const* peer_blocks, const float* means3D,
                          float* dL_dsh, void* cuda_stream);

/* The same for Gaussians [begin, end) only (begin a multiple of 128), so that a caller can pipeline the exchange: gather the
 * dL_drgb of one range of Gaussians from every rank while the previous range's dL_dsh rows are being written.  campos is a
 * separate (n_views,4) array here; peer_blocks[v] + 4 + 3 i must be the address of view v's dL_drgb of Gaussian i for
 * i in [begin, end) -- the pointer itself may be a VIRTUAL base in front of a buffer that holds only that range. */
int gsr_multiview_sh_grad_range(int begin, int end, int D, int M, int n_views, const float* const* peer_blocks, const float* campos,
                                const float* means3D, float* dL_dsh, void* cuda_stream);

/* Rows [begin, end) of a (P,14) small_pack block -> the five gradient tensors (see gsr_backward_args.small_pack). */
int gsr_unpack_small_grads(int begin, int end, const float* small_pack, float* dL_dmean3D, float* dL_dmean2D,
                           float* dL_dopacity, float* dL_dscale, float* dL_drot, void* cuda_stream);

/* Sum over `world` ranks of floats [begin, end) of a buffer that every rank allocated symmetrically and bound to one
 * NVLS multicast address (`multicast_base`, 16-byte aligned; begin a multiple of 4): this rank reduces its 1/world
 * slice through the switch (multimem.ld_reduce) and broadcasts it into every rank's copy (multimem.st).  Collective:
 * every rank launches it with the same range; the caller places a device barrier before (all ranks' data written) and
 * after (all slices broadcast).  No reference counterpart (multi-view training extension). */
int gsr_multimem_allreduce(void* multicast_base, size_t begin, size_t end, int rank, int world, void* cuda_stream);

/* One rank's part of an all-gather done by the NVSwitch: n_floats (a multiple of 4) from local memory are stored once through
 * the multicast address `multicast_dst` (16-byte aligned; this rank's slot of a buffer every rank allocated symmetrically and
 * bound to one NVLS multicast object) and land in every rank's copy.  The caller places a device barrier after it on every
 * rank before reading.  Used for the dL_drgb blocks of multi-view training instead of an NCCL all-gather.  No reference
 * counterpart. */
int gsr_multimem_broadcast(void* multicast_dst, const float* src, size_t n_floats, void* cuda_stream);

/* ---- Fused L1 + SSIM training loss (replaces utils/loss_utils.py:17-64 + train.py:126-127) ----
 * Images are `planes` x H x W float32 (planes = channels, or batch*channels: the window is applied per plane and every
 * mean is over all planes*H*W elements, as F.conv2d(groups = channels) + .mean() do).  Window 11, sigma 1.5, zero padding.
 *
 * forward : out3[0] = (1 - lambda_dssim) * mean|img - gt| + lambda_dssim * (1 - mean ssim(img, gt))   (train.py:127)
 *           out3[1] = mean|img - gt| (l1_loss), out3[2] = mean ssim(img, gt) (ssim, size_average = True);
 *           dm_maps (3 * planes*H*W floats, or NULL when no gradient is needed) receives the per-pixel partial
 *           derivatives of the SSIM map the backward convolves; scratch holds gsr_l1_ssim_scratch_floats floats.
 * backward: dL_dimg = *grad_out * (coef_l1 * sign(img - gt) + coef_ssim * d(sum of the ssim map)/d img);
 *           for the train.py loss coef_l1 = (1 - lambda)/n, coef_ssim = -lambda/n with n = planes*H*W;
 *           grad_out is a DEVICE scalar (NULL = 1), so no host synchronisation is needed. */
size_t gsr_l1_ssim_scratch_floats(int planes, int height, int width);
int gsr_l1_ssim_forward(const float* img, const float* gt, int planes, int height, int width, float lambda_dssim,
                        float* dm_maps, float* scratch, float* out3, void* cuda_stream);
int gsr_l1_ssim_backward(const float* img, const float* gt, const float* dm_maps, int planes, int height, int width,
                         float coef_l1, float coef_ssim, const float* grad_out, float* dL_dimg, void* cuda_stream);

/* gsr_l1_ssim_forward without caller scratch and WITHOUT ANY ALLOCATION (no cudaMalloc / cudaFree, no allocating library
 * call): for callers that cannot allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc.  It has NO counterpart in the reference.
 * Same kernels, same arithmetic and the same outputs as gsr_l1_ssim_forward; the per-tile partial sums live in a
 * file-scope `static __device__` array of this library (csrc/gsr_loss.cu): 2 * GSR_L1_SSIM_NOALLOC_MAX_BLOCKS floats =
 * 1 MiB, held by every GPU context that loads the library until the process exits.
 * Limit: planes * ceil(H/16) * ceil(W/16) <= GSR_L1_SSIM_NOALLOC_MAX_BLOCKS (131072 tiles: 3 x 3840 x 2160 needs 97200);
 * larger images are rejected with GSR_ERR_INVALID_ARGUMENT and a text that names both numbers.
 * The array is SHARED by every call on a device and keeps the last call's partial sums (each call overwrites the range it
 * then reads, so nothing needs clearing): calls that OVERLAP on the GPU -- different streams without an ordering between
 * them -- race on it.  The first call on each device fetches the array's device address (cudaGetSymbolAddress, that
 * device current); call it with the device current on which img / gt / out3 live. */
#define GSR_L1_SSIM_NOALLOC_MAX_BLOCKS 131072
size_t gsr_l1_ssim_noalloc_max_blocks(void);
int gsr_l1_ssim_forward_noalloc(const float* img, const float* gt, int planes, int height, int width, float lambda_dssim,
                                float* dm_maps, float* out3, void* cuda_stream);

/* ---- Fused Adam step over all parameter groups ----
 * Replaces `self.optimizer.step()` (train.py:173-175) for the optimizer scene/gaussian_model.py:243-252 builds:
 * torch.optim.Adam(l, lr=0.0, eps=1e-15), six groups, no weight decay, no amsgrad.  One kernel launch updates every group
 * in place (28 bytes of traffic per element) with torch's arithmetic in torch's order:
 *     exp_avg    <- exp_avg + (1 - beta1) * (grad - exp_avg)
 *     exp_avg_sq <- exp_avg_sq * beta2 + (1 - beta2) * grad * grad
 *     param      <- param - lr / (1 - beta1^step) * exp_avg / (sqrt(exp_avg_sq) / sqrt(1 - beta2^step) + eps)
 * `step` is the group's step count AFTER this update (torch increments state['step'] before using it), so >= 1.
 * `visible` (optional, NULL = off; NOT the reference's semantics): one byte per Gaussian; rows with 0 are left
 * untouched -- parameters and both moments -- where row = element / row_width.  Groups with n == 0 are skipped. */
typedef struct gsr_adam_group {
    float* param;            /* (n) in/out */
    const float* grad;       /* (n) */
    float* exp_avg;          /* (n) in/out, first moment  (state['exp_avg'])    */
    float* exp_avg_sq;       /* (n) in/out, second moment (state['exp_avg_sq']) */
    size_t n;                /* elements */
    double lr;               /* the group's learning rate for this step (param_group['lr']) */
    long long step;          /* >= 1 */
    int row_width;           /* elements per Gaussian: 3, 3, 45, 1, 3, 4 for the six groups; used only with `visible` */
} gsr_adam_group;
int gsr_adam_step(const gsr_adam_group* groups, int n_groups, double beta1, double beta2, double eps,
                  const unsigned char* visible, void* cuda_stream);

/* ---- Densification statistics, every iteration (train.py:153-158) ----
 * Replaces, in one pass and without the host synchronisations of boolean-mask indexing,
 *     max_radii2D[vis] = max(max_radii2D[vis], radii[vis])                                   (train.py:156)
 *     xyz_gradient_accum[vis] += norm(viewspace_point_tensor.grad[vis, :2]); denom[vis] += 1 (gaussian_model.py:599-602)
 * with vis = radii > 0 (train.py:131, gaussian_renderer/__init__.py:108).  viewspace_grad is (P,3) float32 (dL_dmeans2D),
 * radii (P) int32; the three statistics are (P) float32, updated in place; max_radii2D may be NULL. */
int gsr_densify_stats(int P, const float* viewspace_grad, const int* radii, float* xyz_gradient_accum, float* denom,
                      float* max_radii2D, void* cuda_stream);

/* ---- densify_and_prune as two calls (scene/gaussian_model.py:558-597 with :366-556) ----
 * The reference clones, splits and prunes in three rounds of boolean-mask indexing and torch.cat over the six parameter
 * tensors and their twelve Adam moments (about 60 kernels and a dozen host synchronisations).  Here:
 *   gsr_densify_classify  reads what the decisions depend on (gradient statistics, raw scaling, raw opacity), writes one
 *                         flag byte per Gaussian, and leaves the per-block ranks and the five totals on the device;
 *   gsr_densify_apply     moves every surviving row of every tensor ONCE into its final place in new tensors.
 * The final order is the reference's: surviving originals, surviving clones, then the N rounds of split children
 * (round k of parent j uses samples[k * n_split + j], as .repeat(N, 1) lays them out).  Moments of new Gaussians are
 * zero (cat_tensors_to_optimizer); the three statistics are all zero afterwards (densification_postfix), which also
 * makes the reference's `max_radii2D > max_screen_size` test always false -- restated as is.
 * Thresholds are compared in float32 as torch compares a float tensor with a Python scalar:
 *   clone:  grad >= max_grad  and  max(exp(scaling)) <= percent_dense * extent
 *   split:  grad >= max_grad  and  max(exp(scaling)) >  percent_dense * extent,   grad = accum / denom, NaN -> 0
 *   prune:  sigmoid(opacity) < min_opacity  or  (max_screen_size != 0 and max(exp(scaling)) > 0.1 * extent)
 * children: scaling = log(exp(scaling) * (1 / (0.8 N)))  (torch's CUDA division by a scalar multiplies by the
 * reciprocal), xyz = R(rotation) samples + xyz with R = build_rotation (utils/general_utils.py:90-114).
 * totals (5 int32, device): [0] P_out, [1] n_split (rows of `samples` are n_split_copies * n_split), [2] originals kept,
 * [3] clones kept, [4] split parents whose children are kept.  The caller reads totals back (the one host
 * synchronisation), draws `samples` = torch.normal(0, exp(scaling)[split].repeat(N,1)) as the reference does, allocates the
 * outputs with P_out rows and calls gsr_densify_apply with the same scratch. */
#define GSR_DENSIFY_ORIG_KEEP 1u   /* flag bits */
#define GSR_DENSIFY_CLONE_KEEP 2u
#define GSR_DENSIFY_CHILD_KEEP 4u
#define GSR_DENSIFY_CLONE_SEL 8u
#define GSR_DENSIFY_SPLIT_SEL 16u
typedef struct gsr_densify_params {
    double max_grad, min_opacity, extent, percent_dense;   /* opt.densify_grad_threshold, 0.005, scene.cameras_extent, opt.percent_dense */
    double max_screen_size;                                /* size_threshold of train.py:161 (0 = None) */
    int n_split_copies;                                    /* N of densify_and_split (2) */
} gsr_densify_params;
typedef struct gsr_densify_group {
    const float* param; const float* exp_avg; const float* exp_avg_sq;      /* (P, row_width) in; moments may both be NULL */
    float* param_out; float* exp_avg_out; float* exp_avg_sq_out;            /* (P_out, row_width) out, distinct from the inputs */
    int row_width;
} gsr_densify_group;
size_t gsr_densify_scratch_bytes(int P);
int gsr_densify_classify(int P, const float* xyz_gradient_accum, const float* denom, const float* scaling_raw,
                         const float* opacity_raw, const gsr_densify_params* params, unsigned char* flags /* (P) out */,
                         int* totals /* (5) device out */, char* scratch, void* cuda_stream);
/* groups: n_groups <= 8 tensors sharing the row index; xyz_group / scaling_group / rotation_group say which of them
 * are _xyz (3), _scaling (3), _rotation (4).  samples: (n_split_copies * n_split, 3).  totals_host: the five values
 * read back from the device.  Nothing is written beyond row totals_host[0] of any output. */
int gsr_densify_apply(int P, const unsigned char* flags, const char* scratch, const int* totals_host,
                      const gsr_densify_params* params, const float* samples, const gsr_densify_group* groups, int n_groups,
                      int xyz_group, int scaling_group, int rotation_group, void* cuda_stream);

/* ---- Mean squared distance to the 3 nearest neighbours ----
 * Replaces distCUDA2 of the reference's simple-knn extension (submodules/simple-knn/spatial.cu:15-26, SimpleKNN::knn,
 * simple_knn.cu:185-221), used by GaussianModel.create_from_pcd to initialise the scales (scene/gaussian_model.py:207).
 * mean_dist2[i] = (d0 + d1 + d2) / 3 with d0 <= d1 <= d2 the three smallest squared distances from point i to the
 * other points (other = different index: coincident points count with distance 0); with fewer than 4 points the missing
 * neighbours keep the reference's FLT_MAX placeholders (simple_knn.cu:159): +inf for 1 or 2 points, about FLT_MAX/3 for 3.  points is (P,3) float32, scratch holds gsr_knn_scratch_bytes(P)
 * bytes (256-byte aligned) and is free again when the call's work on the stream is done.  No host synchronisation. */
size_t gsr_knn_scratch_bytes(int P);
int gsr_knn_mean_dist2(int P, const float* points, float* mean_dist2, char* scratch, void* cuda_stream);

int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     unsigned char* present /* (P) bool out */, void* cuda_stream);

/* Introspection for tests (device pointers into the opaque buffers; layout is private otherwise) */
typedef struct gsr_buffer_view {
    const uint32_t* point_list;   /* (R)   sorted Gaussian ids, tile-major then depth            */
    const uint32_t* ranges;       /* (T,2) start/end into point_list per tile                   */
    const uint32_t* n_contrib;    /* (H*W)                                                      */
    const float* final_T;         /* (H*W)                                                      */
    const float* packed;          /* (P,12) x,y,conic.xyz,power_thr,opacity,r,g,b,ext_x,ext_y   */
    const uint32_t* tiles_touched;/* (P)                                                        */
    const uint8_t* clamped;       /* (P) bit c set = colour channel c was clamped at 0          */
    const uint8_t* warp_mask;     /* (R) bit w (0..3) set = 8x8 pixel block w of the tile blended the instance (written by the forward) */
} gsr_buffer_view;

int gsr_view_buffers(int P, int R, int width, int height, const char* geom_buffer,
                     const char* binning_buffer, const char* image_buffer, gsr_buffer_view* out);

const char* gsr_last_error(void);
int gsr_version(void);

/* Measurement hooks (bench.py): number of kernels this library has launched since load, and optional
 * per-stage timing.  With profiling on, CUDA events are recorded on the launching stream around each stage
 * of gsr_forward / gsr_backward (preprocess, depth_sort, scan, duplicate, tile_sort, tile_ranges, render,
 * bwd_zero, render_bwd, preprocess_bwd) and folded into a running mean per stage without blocking.
 * gsr_set_profiling(1) clears the means; gsr_get_stage_ms waits for the last events and returns the mean
 * milliseconds (-1 = never ran) and sample counts.  Off by default (no events recorded). */
unsigned long long gsr_launch_count(void);
void gsr_set_profiling(int on);
int gsr_num_stages(void);
const char* gsr_stage_name(int i);
int gsr_get_stage_ms(float* mean_ms, int* samples, int n);

#ifdef __cplusplus
}
#endif
#endif /* GSR_H_INCLUDED */
