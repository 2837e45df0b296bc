// gsr_api.cu -- C-ABI orchestration (include/gsr.h).
//
// Replaces CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (DGR/cuda_rasterizer/rasterizer_impl.cu:162-174,227-411,416-518).  All work is enqueued on the
// caller's stream (the reference uses the legacy default stream).  The one host sync per forward
// -- reading num_rendered to size the binning buffer, rasterizer_impl.cu:331 -- is kept because
// the API returns it as an int; it is a 4-byte async copy into pinned memory + stream sync.
#include "../../include/gsr.h"
#include <cstdlib>
#include <cmath>
#include "gsr_internal.cuh"
#include <cstring>
#include <string>
#include <cstdio>
#include <atomic>
#include <mutex>

using namespace gsr;

// ---- launch counter + optional per-stage CUDA-event timing (bench.py's roofline leg) -----------
static std::atomic<unsigned long long> g_launches{0};
namespace gsr { void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); } }

enum Stage { ST_PREPROCESS = 0, ST_DEPTH_SORT, ST_SCAN, ST_DUPLICATE, ST_TILE_SORT, ST_RANGES, ST_RENDER,
             ST_BWD_ZERO, ST_RENDER_BWD, ST_PREPROCESS_BWD, ST_COUNT };
static const char* const k_stage_names[ST_COUNT] = {"preprocess", "depth_sort", "scan", "duplicate", "tile_sort", "tile_ranges",
                                                    "render", "bwd_zero", "render_bwd", "preprocess_bwd"};
// Per-device profiling state (one process may drive several devices, from several threads): events belong to the device
// they were created on, and the means are kept per device.  gsr_get_stage_ms reports the calling thread's current device.
static std::atomic<int> g_profiling{0};
struct StageProf {
    std::mutex mu;
    cudaEvent_t ev[ST_COUNT][2];
    bool made = false, pending[ST_COUNT] = {};
    double sum_ms[ST_COUNT] = {};
    int n[ST_COUNT] = {}, dropped[ST_COUNT] = {};
};
static StageProf g_prof[GSR_MAX_DEVICES];
static StageProf* current_prof() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= GSR_MAX_DEVICES) return nullptr;
    return &g_prof[dev];
}

// Fold a finished (start, stop) pair of stage s into its running mean.  Never blocks unless `wait`.  Caller holds p.mu.
static void harvest_stage(StageProf& p, int s, bool wait) {
    if (!p.made || !p.pending[s]) return;
    cudaError_t q = wait ? cudaEventSynchronize(p.ev[s][1]) : cudaEventQuery(p.ev[s][1]);
    if (q == cudaSuccess) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, p.ev[s][0], p.ev[s][1]) == cudaSuccess) { p.sum_ms[s] += ms; p.n[s]++; }
    } else {
        if (q == cudaErrorNotReady) (void)cudaGetLastError();
        p.dropped[s]++;
    }
    p.pending[s] = false;
}

struct StageTimer {        // records start/stop events around one stage on the launching stream
    int s; cudaStream_t st; StageProf* p;
    StageTimer(int stage, cudaStream_t stream) : s(stage), st(stream), p(g_profiling.load(std::memory_order_relaxed) ? current_prof() : nullptr) {
        if (!p) return;
        std::lock_guard<std::mutex> lock(p->mu);
        if (!p->made) {
            for (int i = 0; i < ST_COUNT; i++) { cudaEventCreate(&p->ev[i][0]); cudaEventCreate(&p->ev[i][1]); p->pending[i] = false; }
            p->made = true;
        }
        harvest_stage(*p, s, false);     // the previous step's pair for this stage has long finished (the forward syncs once)
        cudaEventRecord(p->ev[s][0], st);
    }
    ~StageTimer() {
        if (!p) return;
        std::lock_guard<std::mutex> lock(p->mu);
        cudaEventRecord(p->ev[s][1], st);
        p->pending[s] = true;
    }
};

static thread_local std::string g_err;
static thread_local uint32_t* g_pinned = nullptr;

// Side stream + events for reading num_rendered back while the depth sort runs (one set per device and host thread).
struct Readback { cudaStream_t side = nullptr; cudaEvent_t produced = nullptr, copied = nullptr; };
static thread_local Readback g_readback[GSR_MAX_DEVICES];

static int fail(int code, const std::string& msg) { g_err = msg; return code; }
namespace gsr { int set_error(int code, const std::string& msg) { return fail(code, msg); } }   // for entry points in other files

#define GSR_CUDA(call)                                                                         \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess)                                                                 \
            return fail(GSR_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));   \
    } while (0)

// CHECK_CUDA (auxiliary.h:177-184): with debug, sync and report the stage that failed
#define GSR_STAGE(name, st, debug)                                                             \
    do {                                                                                       \
        cudaError_t e_ = cudaGetLastError();                                                   \
        if (e_ == cudaSuccess && (debug)) e_ = cudaStreamSynchronize(st);                      \
        if (e_ != cudaSuccess)                                                                 \
            return fail(GSR_ERR_CUDA, std::string("stage ") + name + ": " + cudaGetErrorString(e_)); \
    } while (0)

static int bits_for(uint32_t n_values) {   // bits needed to represent ids 0..n_values-1
    int b = 0;
    while (b < 32 && (1ull << b) < (unsigned long long)n_values) b++;
    return b;
}

extern "C" {

const char* gsr_last_error(void) { return g_err.c_str(); }
int gsr_version(void) { return GSR_VERSION; }

unsigned long long gsr_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
void gsr_set_profiling(int on) {
    g_profiling.store(on, std::memory_order_relaxed);
    if (on)
        for (int d = 0; d < GSR_MAX_DEVICES; d++) {
            std::lock_guard<std::mutex> lock(g_prof[d].mu);
            for (int i = 0; i < ST_COUNT; i++) { g_prof[d].sum_ms[i] = 0.0; g_prof[d].n[i] = 0; g_prof[d].dropped[i] = 0; }
        }
}
int gsr_num_stages(void) { return ST_COUNT; }
const char* gsr_stage_name(int i) { return (i >= 0 && i < ST_COUNT) ? k_stage_names[i] : ""; }
int gsr_get_stage_ms(float* mean_ms, int* samples, int n) {
    if (!mean_ms || n < ST_COUNT) return fail(GSR_ERR_INVALID_ARGUMENT, "need room for gsr_num_stages() entries");
    StageProf* p = current_prof();
    for (int i = 0; i < ST_COUNT; i++) {
        if (p) {
            std::lock_guard<std::mutex> lock(p->mu);
            harvest_stage(*p, i, true);
            mean_ms[i] = p->n[i] ? (float)(p->sum_ms[i] / p->n[i]) : -1.0f;
            if (samples) samples[i] = p->n[i];
        } else {
            mean_ms[i] = -1.0f;
            if (samples) samples[i] = 0;
        }
    }
    return GSR_OK;
}

size_t gsr_backward_scratch_bytes(int P) { return (size_t)(P > 0 ? P : 0) * 12 * sizeof(float) + 256; }

size_t gsr_required_bytes(int kind, size_t n, size_t aux) {
    switch (kind) {
        case 0: return GeomState::carve(nullptr, n).bytes;
        case 1: return BinState::carve(nullptr, n).bytes;
        case 2: return ImageState::carve(nullptr, n, aux).bytes;
        default: return 0;
    }
}

int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     unsigned char* present, void* cuda_stream) {
    (void)projmatrix;   // the reference computes p_proj but only tests p_view.z (auxiliary.h:165)
    if (P < 0) return fail(GSR_ERR_INVALID_ARGUMENT, "P < 0");
    if (P == 0) return GSR_OK;
    if (!means3D || !viewmatrix || !present) return fail(GSR_ERR_INVALID_ARGUMENT, "null pointer");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    launch_mark_visible(P, means3D, viewmatrix, present, st);
    GSR_STAGE("mark_visible", st, 0);
    return GSR_OK;
}

int gsr_forward(const gsr_forward_args* a, gsr_resize_fn geom_fn, void* geom_user, gsr_resize_fn bin_fn, void* bin_user,
                gsr_resize_fn img_fn, void* img_user, void* cuda_stream, int* num_rendered) {
    if (!a || !geom_fn || !bin_fn || !img_fn || !num_rendered) return fail(GSR_ERR_INVALID_ARGUMENT, "null argument");
    *num_rendered = 0;
    if (a->P < 0 || a->width <= 0 || a->height <= 0) return fail(GSR_ERR_INVALID_ARGUMENT, "bad sizes");
    if (a->P == 0) return GSR_OK;     // rasterize_points.cu:94
    if (!a->means3D || !a->opacities || !a->viewmatrix || !a->projmatrix || !a->cam_pos || !a->background || !a->out_color)
        return fail(GSR_ERR_INVALID_ARGUMENT, "missing required pointer");
    const bool split_sh = a->features_dc != nullptr;
    if (split_sh && (a->shs || a->colors_precomp || a->M <= 0 || (a->M > 1 && !a->features_rest)))
        return fail(GSR_ERR_INVALID_ARGUMENT, "split SH storage: features_dc (P,1,3) + features_rest (P,M-1,3), no shs / colors_precomp");
    if (!a->colors_precomp && !split_sh && (!a->shs || a->M <= 0))
        return fail(GSR_ERR_INVALID_ARGUMENT, "provide SHs or precomputed colours");
    if (!a->cov3D_precomp && (!a->scales || !a->rotations))
        return fail(GSR_ERR_INVALID_ARGUMENT, "provide scales+rotations or a precomputed 3D covariance");
    if ((a->shs || split_sh) && (a->D < 0 || a->D > 3 || (a->D + 1) * (a->D + 1) > a->M))
        return fail(GSR_ERR_INVALID_ARGUMENT, "SH degree does not fit M coefficients");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int P = a->P, W = a->width, H = a->height;
    const int grid_x = (W + GSR_TILE_X - 1) / GSR_TILE_X, grid_y = (H + GSR_TILE_Y - 1) / GSR_TILE_Y;
    int row_begin = 0, row_end = grid_y;
    if (!(a->tile_row_begin == 0 && a->tile_row_end == 0)) {
        row_begin = a->tile_row_begin; row_end = a->tile_row_end;
        if (row_begin < 0 || row_end > grid_y || row_begin > row_end) return fail(GSR_ERR_INVALID_ARGUMENT, "bad tile row band");
    }
    const size_t tiles = (size_t)grid_x * grid_y;

    char* gptr = geom_fn(geom_user, GeomState::carve(nullptr, P).bytes);
    char* iptr = img_fn(img_user, ImageState::carve(nullptr, (size_t)W * H, tiles).bytes);
    if (!gptr || !iptr) return fail(GSR_ERR_ALLOC, "resize callback returned NULL");
    GeomState geom = GeomState::carve((char*)align_up((size_t)gptr, 256), P);
    ImageState img = ImageState::carve((char*)align_up((size_t)iptr, 256), (size_t)W * H, tiles);
    int* radii = a->radii ? a->radii : geom.radii_internal;

    PreprocessArgs pa;
    pa.P = P; pa.D = a->D; pa.M = a->M;
    pa.means3D = a->means3D; pa.scales = a->scales; pa.scale_modifier = a->scale_modifier; pa.rotations = a->rotations;
    pa.opacities = a->opacities; pa.shs = a->shs; pa.cov3D_precomp = a->cov3D_precomp; pa.colors_precomp = a->colors_precomp;
    pa.cam.view = a->viewmatrix; pa.cam.proj = a->projmatrix; pa.cam.campos = a->cam_pos;
    pa.W = W; pa.H = H; pa.tan_fovx = a->tan_fovx; pa.tan_fovy = a->tan_fovy;
    pa.focal_y = H / (2.0f * a->tan_fovy);     // rasterizer_impl.cu:251-252
    pa.focal_x = W / (2.0f * a->tan_fovx);
    pa.grid_x = grid_x; pa.grid_y = grid_y; pa.row_begin = row_begin; pa.row_end = row_end;
    pa.prefiltered = a->prefiltered;
    pa.radii = radii; pa.packed = geom.packed; pa.tiles_touched = geom.tiles_touched; pa.clamped = geom.clamped;
    pa.depth_keys = geom.keys[0]; pa.ids = geom.vals[0];
    pa.total = geom.total;
    pa.raw_act = a->raw_activations ? 1 : 0; pa.f_dc = a->features_dc; pa.f_rest = a->features_rest;
    const bool use_rects = grid_x < 65536 && grid_y < 65536;      // 16-bit tile coordinates in the packed rectangle
    pa.rects = use_rects ? geom.rects : nullptr;
    GSR_CUDA(cudaMemsetAsync(geom.total, 0, sizeof(uint32_t), st));
    { StageTimer t(ST_PREPROCESS, st); launch_preprocess(pa, st); }
    GSR_STAGE("preprocess", st, a->debug);

    // num_rendered is final once preprocess is done (a sum does not care about the depth order): copy it to the host
    // on a side stream now, so the round trip and the binning-buffer resize hide behind the depth sort
    int dev = 0;
    GSR_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= GSR_MAX_DEVICES) return fail(GSR_ERR_INVALID_ARGUMENT, "device ordinal out of range");
    Readback& rbk = g_readback[dev];
    if (!rbk.side) {
        GSR_CUDA(cudaStreamCreateWithFlags(&rbk.side, cudaStreamNonBlocking));
        GSR_CUDA(cudaEventCreateWithFlags(&rbk.produced, cudaEventDisableTiming));
        GSR_CUDA(cudaEventCreateWithFlags(&rbk.copied, cudaEventDisableTiming));
    }
    if (!g_pinned) GSR_CUDA(cudaHostAlloc((void**)&g_pinned, 64, cudaHostAllocDefault));
    GSR_CUDA(cudaEventRecord(rbk.produced, st));
    GSR_CUDA(cudaStreamWaitEvent(rbk.side, rbk.produced, 0));
    GSR_CUDA(cudaMemcpyAsync(g_pinned, geom.total, sizeof(uint32_t), cudaMemcpyDeviceToHost, rbk.side));
    GSR_CUDA(cudaEventRecord(rbk.copied, rbk.side));

    // depth order of the Gaussians: 4 stable 8-bit passes, result back in keys[0]/vals[0]
    {
        StageTimer t(ST_DEPTH_SORT, st);
        const int shifts[4] = {0, 8, 16, 24}, bits[4] = {8, 8, 8, 8};
        radix_sort_u32(geom.keys, geom.vals, 0, P, 4, shifts, bits, geom.table, st);
    }
    GSR_STAGE("depth sort", st, a->debug);
    {
        StageTimer t(ST_SCAN, st);
        // (a single-pass gather + scan with decoupled look-back measured slower here: 0.044 vs 0.036 ms at P = 3 M)
        if (use_rects) launch_gather_tt_rect(geom.tiles_touched, geom.rects, geom.vals[0], geom.tt_sorted, geom.rect_sorted, P, st);
        else launch_gather_u32(geom.tiles_touched, geom.vals[0], geom.tt_sorted, P, st);
        scan_exclusive_u32(geom.tt_sorted, geom.offsets, P, geom.scan_tmp, nullptr, st);
    }
    GSR_STAGE("scan", st, a->debug);

    GSR_CUDA(cudaEventSynchronize(rbk.copied));
    const size_t R = *g_pinned;
    if (R > 0x7fffffffull) return fail(GSR_ERR_INVALID_ARGUMENT, "num_rendered exceeds int range");

    char* bptr = bin_fn(bin_user, BinState::carve(nullptr, R).bytes);
    if (!bptr) return fail(GSR_ERR_ALLOC, "resize callback returned NULL");
    BinState bin = BinState::carve((char*)align_up((size_t)bptr, 256), R);

    // Instance list -> grouped by tile.  Grids of at most TC_MAX_TILES tiles (1080p: 8160, 4K: 32 400): one counting pass with
    // the tile cursors in shared memory, which also yields the ranges; larger grids: LSD radix passes on the tile id and
    // a pass over the sorted keys for the ranges.  GSR_TILE_SORT=radix forces the second (tests compare the two).
    const char* forced = getenv("GSR_TILE_SORT");           // read per call: the tests switch it inside one process
    const bool counting = tile_count_sort_applies(tiles) && !(forced && !strcmp(forced, "radix"));
    const int tile_bits = bits_for((uint32_t)tiles);
    const int npass = (tile_bits + 7) / 8;
    int s = counting ? 1 : (npass & 1);       // start so that the last pass lands in vals[0] == point_list
    {
        StageTimer t(ST_DUPLICATE, st);
        launch_duplicate(P, geom.vals[0], geom.offsets, geom.tt_sorted, use_rects ? geom.rect_sorted : nullptr, geom.packed, radii, grid_x, grid_y,
                         row_begin, row_end,
                         bin.keys[s], bin.vals[s], st);
    }
    GSR_STAGE("duplicate", st, a->debug);
    if (counting) {
        StageTimer t(ST_TILE_SORT, st);
        tile_count_sort(bin.keys[1], bin.vals[1], R, tiles, bin.table, bin.vals[0], img.ranges, st);
    } else {
        StageTimer t(ST_TILE_SORT, st);
        // digits of equal width (13 tile bits -> 7 + 6): fewer bins per pass = longer contiguous runs per block
        const int per = npass ? (tile_bits + npass - 1) / npass : 0;
        int shifts[4] = {0, 0, 0, 0}, bits[4] = {0, 0, 0, 0};
        for (int pass = 0, shift = 0; pass < npass; pass++) {
            bits[pass] = (tile_bits - shift) < per ? (tile_bits - shift) : per;
            shifts[pass] = shift;
            shift += bits[pass];
        }
        // one histogram launch per pass here: the fused histogram of both tile digits measured slower at R = 12 M
        // (0.075 ms vs 2 x 0.019 ms; the runs of equal high digits serialise on one shared-memory counter)
        for (int pass = 0; pass < npass; pass++) radix_sort_u32(bin.keys, bin.vals, s ^ (pass & 1), R, 1, shifts + pass, bits + pass, bin.table, st);
    }
    GSR_STAGE("tile sort", st, a->debug);
    if (!counting) { StageTimer t(ST_RANGES, st); launch_tile_ranges(R, bin.keys[0], img.ranges, tiles, st); }
    GSR_STAGE("tile ranges", st, a->debug);
    {
        StageTimer t(ST_RENDER, st);
        launch_render_forward(grid_x, row_begin, row_end, img.ranges, bin.vals[0], W, H, geom.packed, a->background,
                              a->out_color, img.final_T, img.n_contrib, bin.warp_mask, st);
    }
    GSR_STAGE("render", st, a->debug);
    *num_rendered = (int)R;
    return GSR_OK;
}

int gsr_backward(const gsr_backward_args* a, void* cuda_stream) {
    if (!a) return fail(GSR_ERR_INVALID_ARGUMENT, "null argument");
    if (a->P < 0 || a->R < 0 || a->width <= 0 || a->height <= 0) return fail(GSR_ERR_INVALID_ARGUMENT, "bad sizes");
    if (a->P == 0) return GSR_OK;     // rasterize_points.cu:180
    if (!a->means3D || !a->radii || !a->geom_buffer || !a->binning_buffer || !a->image_buffer || !a->dL_dpix ||
        !a->viewmatrix || !a->projmatrix || !a->cam_pos || !a->background || !a->grad_scratch)
        return fail(GSR_ERR_INVALID_ARGUMENT, "missing required pointer");
    const bool split_sh = a->features_dc != nullptr;
    if (!a->dL_dmean2D || !a->dL_dopacity || !a->dL_dmean3D || !a->dL_dscale || !a->dL_drot || (a->M > 0 && !split_sh && !a->dL_dsh))
        return fail(GSR_ERR_INVALID_ARGUMENT, "missing output pointer");
    // the gradient of an input that was not given may be left out (NULL): 36 bytes per Gaussian less to write on the SH / scale+rotation path
    if ((a->colors_precomp && !a->dL_dcolor) || (a->cov3D_precomp && !a->dL_dcov3D))
        return fail(GSR_ERR_INVALID_ARGUMENT, "dL_dcolor / dL_dcov3D are required when colors_precomp / cov3D_precomp are given");
    if (split_sh) {
        if (a->shs || a->dL_dsh || a->colors_precomp || a->M <= 0 || a->M > 16 || !a->dL_dfeatures_dc || (a->M > 1 && (!a->features_rest || !a->dL_dfeatures_rest)))
            return fail(GSR_ERR_INVALID_ARGUMENT, "split SH storage: features_dc/_rest with dL_dfeatures_dc/_rest, M <= 16, no shs / dL_dsh / colors_precomp");
        if (a->rotations && ((uintptr_t)a->rotations & 15)) return fail(GSR_ERR_INVALID_ARGUMENT, "split SH storage: rotations must be 16-byte aligned");
    }
    if ((split_sh || a->raw_activations) && (a->dL_drgb || a->small_pack))
        return fail(GSR_ERR_INVALID_ARGUMENT, "the raw-parameter path is not combined with the multi-view exchange options");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int P = a->P, W = a->width, H = a->height;
    const int grid_x = (W + GSR_TILE_X - 1) / GSR_TILE_X, grid_y = (H + GSR_TILE_Y - 1) / GSR_TILE_Y;
    int row_begin = 0, row_end = grid_y;
    if (!(a->tile_row_begin == 0 && a->tile_row_end == 0)) { row_begin = a->tile_row_begin; row_end = a->tile_row_end; }
    if (row_begin < 0 || row_end > grid_y || row_begin > row_end) return fail(GSR_ERR_INVALID_ARGUMENT, "bad tile row band");
    const size_t tiles = (size_t)grid_x * grid_y;
    GeomState geom = GeomState::carve((char*)align_up((size_t)a->geom_buffer, 256), P);
    BinState bin = BinState::carve((char*)align_up((size_t)a->binning_buffer, 256), (size_t)a->R);
    ImageState img = ImageState::carve((char*)align_up((size_t)a->image_buffer, 256), (size_t)W * H, tiles);

    float* acc = (float*)align_up((size_t)a->grad_scratch, 256);
    {
        StageTimer t(ST_BWD_ZERO, st);
        GSR_CUDA(cudaMemsetAsync(acc, 0, (size_t)P * 12 * sizeof(float), st));
    }
    if (a->R > 0) {
        StageTimer t(ST_RENDER_BWD, st);
        launch_render_backward(grid_x, row_begin, row_end, img.ranges, bin.vals[0], W, H, geom.packed, a->background,
                               img.final_T, img.n_contrib, a->dL_dpix, acc, bin.warp_mask, st);
    }
    GSR_STAGE("render backward", st, a->debug);

    const bool early_drgb = a->dL_drgb != nullptr && a->drgb_ready_event != nullptr;
    if (early_drgb) {     // publish the 12 B/Gaussian the other ranks need first, so their copies overlap the rest of this call
        launch_publish_drgb(P, acc, a->radii, a->shs ? geom.clamped : nullptr, a->dL_drgb, st);
        GSR_CUDA(cudaEventRecord((cudaEvent_t)a->drgb_ready_event, st));
    }
    PreprocessBwdArgs pb;
    pb.P = P; pb.D = a->D; pb.M = a->M;
    pb.means3D = a->means3D; pb.radii = a->radii; pb.shs = a->shs; pb.clamped = geom.clamped;
    pb.scales = a->scales; pb.rotations = a->rotations; pb.scale_modifier = a->scale_modifier; pb.cov3D_precomp = a->cov3D_precomp;
    pb.cam.view = a->viewmatrix; pb.cam.proj = a->projmatrix; pb.cam.campos = a->cam_pos;
    pb.W = W; pb.H = H; pb.tan_fovx = a->tan_fovx; pb.tan_fovy = a->tan_fovy;
    pb.focal_y = H / (2.0f * a->tan_fovy);     // rasterizer_impl.cu:459-460
    pb.focal_x = W / (2.0f * a->tan_fovx);
    pb.grad_acc = acc; pb.packed = geom.packed;
    pb.dL_dmean2D = a->dL_dmean2D; pb.dL_dconic = a->dL_dconic; pb.dL_dopacity = a->dL_dopacity; pb.dL_dcolor = a->dL_dcolor;
    pb.dL_dmean3D = a->dL_dmean3D; pb.dL_dcov3D = a->dL_dcov3D; pb.dL_dsh = a->dL_dsh; pb.dL_dscale = a->dL_dscale; pb.dL_drot = a->dL_drot;
    pb.dL_drgb = early_drgb ? nullptr : a->dL_drgb; pb.skip_dsh = (a->dL_drgb != nullptr && a->skip_dL_dsh) ? 1 : 0;
    pb.begin = 0; pb.chunked = a->dL_drgb != nullptr;
    pb.small_pack = a->small_pack;
    pb.raw_act = a->raw_activations ? 1 : 0; pb.f_dc = a->features_dc; pb.f_rest = a->features_rest;
    pb.dL_df_dc = a->dL_dfeatures_dc; pb.dL_df_rest = a->dL_dfeatures_rest;
    pb.chunk_events = a->chunk_done_events;
    pb.n_chunk_events = a->chunk_done_events ? (a->n_chunk_events < GSR_MAX_CHUNKS ? a->n_chunk_events : GSR_MAX_CHUNKS) : 0;
    { StageTimer t(ST_PREPROCESS_BWD, st); launch_preprocess_backward(pb, st); }
    GSR_STAGE("preprocess backward", st, a->debug);
    return GSR_OK;
}

int gsr_unpack_small_grads(int begin, int end, const float* small_pack, float* dL_dmean3D, float* dL_dmean2D,
                           float* dL_dopacity, float* dL_dscale, float* dL_drot, void* cuda_stream) {
    if (begin < 0 || end < begin) return fail(GSR_ERR_INVALID_ARGUMENT, "bad range");
    if (end == begin) return GSR_OK;
    if (!small_pack || !dL_dmean3D || !dL_dmean2D || !dL_dopacity || !dL_dscale || !dL_drot)
        return fail(GSR_ERR_INVALID_ARGUMENT, "null pointer");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    launch_unpack_small(begin, end, small_pack, dL_dmean3D, dL_dmean2D, dL_dopacity, dL_dscale, dL_drot, st);
    GSR_STAGE("unpack small gradients", st, 0);
    return GSR_OK;
}

int gsr_multimem_allreduce(void* multicast_base, size_t begin, size_t end, int rank, int world, void* cuda_stream) {
    if (!multicast_base || ((size_t)multicast_base & 15) || (begin & 3) || end < begin || world < 1 || rank < 0 || rank >= world)
        return fail(GSR_ERR_INVALID_ARGUMENT, "bad multimem all-reduce arguments");
    static const int max_blocks = getenv("GSR_MM_BLOCKS") ? atoi(getenv("GSR_MM_BLOCKS")) : 148;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    launch_multimem_allreduce((float*)multicast_base, begin, end, rank, world, max_blocks > 0 ? max_blocks : 148, st);
    GSR_STAGE("multimem all-reduce", st, 0);
    return GSR_OK;
}

int gsr_multimem_broadcast(void* multicast_dst, const float* src, size_t n_floats, void* cuda_stream) {
    if (n_floats == 0) return GSR_OK;
    if (!multicast_dst || !src || ((size_t)multicast_dst & 15) || ((size_t)src & 15) || (n_floats & 3))
        return fail(GSR_ERR_INVALID_ARGUMENT, "multicast destination / source must be 16-byte aligned and the length a multiple of 4 floats");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    launch_multimem_broadcast((float*)multicast_dst, src, n_floats, 148 * 8, st);
    GSR_STAGE("multimem broadcast", st, 0);
    return GSR_OK;
}

size_t gsr_l1_ssim_scratch_floats(int planes, int height, int width) {
    if (planes <= 0 || height <= 0 || width <= 0) return 0;
    return 2 * l1_ssim_blocks(planes, height, width);
}

int gsr_l1_ssim_forward(const float* img, const float* gt, int planes, int height, int width, float lambda_dssim,
                        float* dm_maps, float* scratch, float* out3, void* cuda_stream) {
    if (planes <= 0 || height <= 0 || width <= 0) return fail(GSR_ERR_INVALID_ARGUMENT, "bad image sizes");
    if (!img || !gt || !scratch || !out3) return fail(GSR_ERR_INVALID_ARGUMENT, "null pointer");
    if (l1_ssim_blocks(planes, height, width) > 0x7fffffffull || planes > 65535) return fail(GSR_ERR_INVALID_ARGUMENT, "image too large");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    launch_l1_ssim_forward(img, gt, planes, height, width, lambda_dssim, dm_maps, scratch, out3, st);
    GSR_STAGE("l1+ssim forward", st, 0);
    return GSR_OK;
}

int gsr_l1_ssim_backward(const float* img, const float* gt, const float* dm_maps, int planes, int height, int width,
                         float coef_l1, float coef_ssim, const float* grad_out, float* dL_dimg, void* cuda_stream) {
    if (planes <= 0 || height <= 0 || width <= 0) return fail(GSR_ERR_INVALID_ARGUMENT, "bad image sizes");
    if (!img || !gt || !dm_maps || !dL_dimg) return fail(GSR_ERR_INVALID_ARGUMENT, "null pointer");
    if (planes > 65535) return fail(GSR_ERR_INVALID_ARGUMENT, "image too large");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    launch_l1_ssim_backward(img, gt, dm_maps, planes, height, width, coef_l1, coef_ssim, grad_out, dL_dimg, st);
    GSR_STAGE("l1+ssim backward", st, 0);
    return GSR_OK;
}

// ---- fused Adam step (torch.optim.Adam as scene/gaussian_model.py:252 configures it) ----
int gsr_adam_step(const gsr_adam_group* groups, int n_groups, double beta1, double beta2, double eps,
                  const unsigned char* visible, void* cuda_stream) {
    if (n_groups < 0 || (n_groups > 0 && !groups)) return fail(GSR_ERR_INVALID_ARGUMENT, "adam: bad group list");
    if (!(beta1 >= 0.0 && beta1 < 1.0) || !(beta2 >= 0.0 && beta2 < 1.0) || !(eps >= 0.0))
        return fail(GSR_ERR_INVALID_ARGUMENT, "adam: betas must be in [0,1) and eps >= 0");     // torch/optim/adam.py's own checks
    for (int i = 0; i < n_groups; i++) {
        const gsr_adam_group& g = groups[i];
        if (g.n == 0) continue;
        if (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq) return fail(GSR_ERR_INVALID_ARGUMENT, "adam: null tensor in group " + std::to_string(i));
        if (g.step < 1) return fail(GSR_ERR_INVALID_ARGUMENT, "adam: step must be >= 1 (the count after this update)");
        if (!(g.lr >= 0.0)) return fail(GSR_ERR_INVALID_ARGUMENT, "adam: negative learning rate");
        if (visible && g.row_width < 1) return fail(GSR_ERR_INVALID_ARGUMENT, "adam: row_width must be >= 1 with a visibility mask");
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    int i = 0;
    while (i < n_groups) {
        AdamLaunch L;
        L.n_groups = 0;
        L.beta2 = (float)beta2; L.w1 = (float)(1.0 - beta1); L.w2 = (float)(1.0 - beta2); L.eps = (float)eps;
        L.visible = visible;
        unsigned long long blocks = 0;
        bool vec = true;
        for (; i < n_groups && L.n_groups < ADAM_MAX_GROUPS; i++) {
            const gsr_adam_group& g = groups[i];
            if (g.n == 0) continue;
            const unsigned long long nb = adam_blocks(g.n);
            if (blocks + nb > 0x7fffffffull) {
                if (L.n_groups == 0) return fail(GSR_ERR_INVALID_ARGUMENT, "adam: group too large for one grid");
                break;                                           // goes into the next launch
            }
            AdamGroupDev& d = L.g[L.n_groups];
            d.p = g.param; d.g = g.grad; d.m = g.exp_avg; d.v = g.exp_avg_sq; d.n = g.n;
            // host scalars in double, as torch computes them in Python floats (torch/optim/adam.py, _multi_tensor_adam)
            const double bc1 = 1.0 - std::pow(beta1, (double)g.step);
            const double bc2 = 1.0 - std::pow(beta2, (double)g.step);
            d.neg_step = (float)(-(g.lr / bc1));
            d.bc2_sqrt = (float)std::sqrt(bc2);
            d.row_width = (unsigned)(g.row_width > 0 ? g.row_width : 1);
            const uintptr_t a = (uintptr_t)g.param | (uintptr_t)g.grad | (uintptr_t)g.exp_avg | (uintptr_t)g.exp_avg_sq;
            if (a & 15) vec = false;
            L.block_start[L.n_groups] = (unsigned)blocks;
            blocks += nb;
            L.n_groups++;
        }
        for (int k = L.n_groups; k <= ADAM_MAX_GROUPS; k++) L.block_start[k] = (unsigned)blocks;
        for (int k = L.n_groups; k < ADAM_MAX_GROUPS; k++) L.g[k] = AdamGroupDev{nullptr, nullptr, nullptr, nullptr, 0, 0.f, 1.f, 1};
        if (blocks == 0) continue;                               // only empty groups in this batch
        launch_adam(L, (unsigned)blocks, vec, st);
        GSR_STAGE("adam step", st, 0);
    }
    return GSR_OK;
}

// ---- densification (train.py:153-158, scene/gaussian_model.py:558-602) ----
int gsr_densify_stats(int P, const float* viewspace_grad, const int* radii, float* xyz_gradient_accum, float* denom,
                      float* max_radii2D, void* cuda_stream) {
    if (P < 0) return fail(GSR_ERR_INVALID_ARGUMENT, "densify_stats: negative Gaussian count");
    if (P == 0) return GSR_OK;
    if (!viewspace_grad || !radii || !xyz_gradient_accum || !denom) return fail(GSR_ERR_INVALID_ARGUMENT, "densify_stats: null pointer");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    launch_densify_stats(P, viewspace_grad, radii, xyz_gradient_accum, denom, max_radii2D, st);
    GSR_STAGE("densify stats", st, 0);
    return GSR_OK;
}

static int densify_thresholds(const gsr_densify_params* p, DensifyThresholds& t) {
    if (!p) return fail(GSR_ERR_INVALID_ARGUMENT, "densify: null params");
    if (p->n_split_copies < 1 || p->n_split_copies > DENSIFY_MAX_COPIES)
        return fail(GSR_ERR_INVALID_ARGUMENT, "densify: n_split_copies must be in [1, " + std::to_string(DENSIFY_MAX_COPIES) + "]");
    // Python evaluates percent_dense * extent, 0.1 * extent and 0.8 * N in double; torch rounds the scalar to the tensor's float32
    t.max_grad = (float)p->max_grad;
    t.dense_extent = (float)(p->percent_dense * p->extent);
    t.big_extent = (float)(0.1 * p->extent);
    t.min_opacity = (float)p->min_opacity;
    t.inv_shrink = 1.0f / (float)(0.8 * p->n_split_copies);
    t.use_screen = p->max_screen_size != 0.0;                        // `if max_screen_size:`
    t.copies = p->n_split_copies;
    return GSR_OK;
}

size_t gsr_densify_scratch_bytes(int P) { return P > 0 ? DensifyState::carve(nullptr, (size_t)P).bytes : 0; }

int gsr_densify_classify(int P, const float* xyz_gradient_accum, const float* denom, const float* scaling_raw,
                         const float* opacity_raw, const gsr_densify_params* params, unsigned char* flags, int* totals,
                         char* scratch, void* cuda_stream) {
    if (P < 0) return fail(GSR_ERR_INVALID_ARGUMENT, "densify_classify: negative Gaussian count");
    DensifyThresholds t;
    if (int e = densify_thresholds(params, t)) return e;
    if (!totals) return fail(GSR_ERR_INVALID_ARGUMENT, "densify_classify: null totals");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (P == 0) { GSR_CUDA(cudaMemsetAsync(totals, 0, 5 * sizeof(int), st)); return GSR_OK; }
    if (!xyz_gradient_accum || !denom || !scaling_raw || !opacity_raw || !flags || !scratch)
        return fail(GSR_ERR_INVALID_ARGUMENT, "densify_classify: null pointer");
    if (((uintptr_t)scratch & 255) != 0) return fail(GSR_ERR_INVALID_ARGUMENT, "densify_classify: scratch must be 256-byte aligned");
    launch_densify_classify(P, xyz_gradient_accum, denom, scaling_raw, opacity_raw, t, flags, totals, DensifyState::carve(scratch, (size_t)P), st);
    GSR_STAGE("densify classify", st, 0);
    return GSR_OK;
}

int gsr_densify_apply(int P, const unsigned char* flags, const char* scratch, const int* totals_host,
                      const gsr_densify_params* params, const float* samples, const gsr_densify_group* groups, int n_groups,
                      int xyz_group, int scaling_group, int rotation_group, void* cuda_stream) {
    if (P < 0) return fail(GSR_ERR_INVALID_ARGUMENT, "densify_apply: negative Gaussian count");
    DensifyApply A;
    if (int e = densify_thresholds(params, A.t)) return e;
    if (P == 0) return GSR_OK;
    if (!flags || !scratch || !totals_host || !groups) return fail(GSR_ERR_INVALID_ARGUMENT, "densify_apply: null pointer");
    if (n_groups < 1 || n_groups > DENSIFY_MAX_GROUPS) return fail(GSR_ERR_INVALID_ARGUMENT, "densify_apply: 1 to 8 tensor groups");
    const long long P_out = totals_host[0], n_split = totals_host[1], n_ok = totals_host[2], n_ck = totals_host[3], n_chk = totals_host[4];
    if (n_ok < 0 || n_ck < 0 || n_chk < 0 || n_split < n_chk || n_ok + n_split > P || n_ck > P ||
        P_out != n_ok + n_ck + (long long)A.t.copies * n_chk)
        return fail(GSR_ERR_INVALID_ARGUMENT, "densify_apply: totals are not the ones gsr_densify_classify wrote for this model");
    if (P_out > 2147483647LL) return fail(GSR_ERR_INVALID_ARGUMENT, "densify_apply: result too large");
    auto role_ok = [&](int gi, int w) { return gi >= 0 && gi < n_groups && groups[gi].row_width == w; };
    if (!role_ok(xyz_group, 3) || !role_ok(scaling_group, 3) || !role_ok(rotation_group, 4) || xyz_group == scaling_group)
        return fail(GSR_ERR_INVALID_ARGUMENT, "densify_apply: xyz / scaling / rotation group indices or row widths are wrong");
    if (n_split > 0 && !samples) return fail(GSR_ERR_INVALID_ARGUMENT, "densify_apply: samples needed for the split");
    for (int i = 0; i < n_groups; i++) {
        const gsr_densify_group& g = groups[i];
        if (g.row_width < 1) return fail(GSR_ERR_INVALID_ARGUMENT, "densify_apply: row_width must be >= 1 in group " + std::to_string(i));
        if (!g.param || (P_out > 0 && !g.param_out)) return fail(GSR_ERR_INVALID_ARGUMENT, "densify_apply: null tensor in group " + std::to_string(i));
        const bool has_state = g.exp_avg || g.exp_avg_sq;
        if (has_state && (!g.exp_avg || !g.exp_avg_sq || (P_out > 0 && (!g.exp_avg_out || !g.exp_avg_sq_out))))
            return fail(GSR_ERR_INVALID_ARGUMENT, "densify_apply: moments must come in pairs, with outputs, in group " + std::to_string(i));
        if (g.param_out == g.param || (has_state && (g.exp_avg_out == g.exp_avg || g.exp_avg_sq_out == g.exp_avg_sq)))
            return fail(GSR_ERR_INVALID_ARGUMENT, "densify_apply: outputs must not alias the inputs (group " + std::to_string(i) + ")");
        A.g[i] = DensifyGroupDev{g.param, g.exp_avg, g.exp_avg_sq, g.param_out, has_state ? g.exp_avg_out : nullptr,
                                 has_state ? g.exp_avg_sq_out : nullptr, g.row_width};
    }
    for (int i = n_groups; i < DENSIFY_MAX_GROUPS; i++) A.g[i] = DensifyGroupDev{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 1};
    A.n_groups = n_groups; A.xyz_group = xyz_group; A.scaling_group = scaling_group; A.rotation_group = rotation_group;
    A.P = P; A.n_split = (int)n_split; A.n_ok = (int)n_ok; A.n_ck = (int)n_ck; A.n_chk = (int)n_chk;
    A.flags = flags;
    A.block_offsets = DensifyState::carve(const_cast<char*>(scratch), (size_t)P).block_offsets;
    A.samples = samples;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    launch_densify_apply(A, st);
    GSR_STAGE("densify apply", st, 0);
    return GSR_OK;
}

// ---- distCUDA2 of simple-knn ----
size_t gsr_knn_scratch_bytes(int P) { return P > 0 ? KnnState::carve(nullptr, (size_t)P).bytes : 0; }

int gsr_knn_mean_dist2(int P, const float* points, float* mean_dist2, char* scratch, void* cuda_stream) {
    if (P < 0) return fail(GSR_ERR_INVALID_ARGUMENT, "knn: negative point count");
    if (P == 0) return GSR_OK;
    if (!points || !mean_dist2 || !scratch) return fail(GSR_ERR_INVALID_ARGUMENT, "knn: null pointer");
    if (((uintptr_t)scratch & 255) != 0) return fail(GSR_ERR_INVALID_ARGUMENT, "knn: scratch must be 256-byte aligned");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    launch_knn(P, points, mean_dist2, KnnState::carve(scratch, (size_t)P), st);
    GSR_STAGE("knn", st, 0);
    return GSR_OK;
}

int gsr_backward_chunk_plan(int P, int* bounds) {
    if (!bounds || P < 0) return 0;
    int b[5];
    const int n = chunk_plan(P, 256, b);
    for (int i = 0; i <= n; i++) bounds[i] = b[i];
    return n;
}

int gsr_multiview_sh_grad(int P, int D, int M, int n_views, const float* const* peer_blocks, const float* means3D,
                          float* dL_dsh, void* cuda_stream) {
    if (P < 0 || D < 0 || D > 3 || M < (D + 1) * (D + 1) || n_views < 1 || n_views > GSR_MAX_VIEWS)
        return fail(GSR_ERR_INVALID_ARGUMENT, "bad sizes");
    if (P == 0) return GSR_OK;
    if (!peer_blocks || !means3D || !dL_dsh) return fail(GSR_ERR_INVALID_ARGUMENT, "null pointer");
    MultiviewShArgs a;
    a.P = P; a.D = D; a.M = M; a.n_views = n_views; a.means3D = means3D; a.dL_dsh = dL_dsh; a.begin = 0;
    a.campos = nullptr; a.no_split = 0;
    for (int v = 0; v < 16; v++) a.blocks[v] = nullptr;
    for (int v = 0; v < n_views; v++) {
        if (!peer_blocks[v] || ((size_t)peer_blocks[v] & 15)) return fail(GSR_ERR_INVALID_ARGUMENT, "peer block missing or not 16-byte aligned");
        a.blocks[v] = peer_blocks[v];
    }
    if ((size_t)dL_dsh & 15) return fail(GSR_ERR_INVALID_ARGUMENT, "dL_dsh not 16-byte aligned");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    launch_multiview_sh_grad(a, st);
    GSR_STAGE("multiview sh grad", st, 0);
    return GSR_OK;
}

int gsr_multiview_sh_grad_range(int begin, int end, int D, int M, int n_views, const float* const* peer_blocks, const float* campos,
                                const float* means3D, float* dL_dsh, void* cuda_stream) {
    if (begin < 0 || end < begin || (begin & 127) || D < 0 || D > 3 || M < (D + 1) * (D + 1) || n_views < 1 || n_views > GSR_MAX_VIEWS)
        return fail(GSR_ERR_INVALID_ARGUMENT, "bad sizes (begin must be a multiple of 128)");
    if (end == begin) return GSR_OK;
    if (!peer_blocks || !campos || !means3D || !dL_dsh) return fail(GSR_ERR_INVALID_ARGUMENT, "null pointer");
    MultiviewShArgs a;
    a.P = end; a.D = D; a.M = M; a.n_views = n_views; a.means3D = means3D; a.dL_dsh = dL_dsh; a.begin = begin;
    a.campos = campos; a.no_split = 1;
    for (int v = 0; v < 16; v++) a.blocks[v] = nullptr;
    for (int v = 0; v < n_views; v++) {
        if (!peer_blocks[v] || ((size_t)peer_blocks[v] & 15)) return fail(GSR_ERR_INVALID_ARGUMENT, "peer block missing or not 16-byte aligned");
        a.blocks[v] = peer_blocks[v];
    }
    if ((size_t)dL_dsh & 15) return fail(GSR_ERR_INVALID_ARGUMENT, "dL_dsh not 16-byte aligned");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    launch_multiview_sh_grad(a, st);
    GSR_STAGE("multiview sh grad (range)", st, 0);
    return GSR_OK;
}

int gsr_view_buffers(int P, int R, int width, int height, const char* geom_buffer, const char* binning_buffer,
                     const char* image_buffer, gsr_buffer_view* out) {
    if (!out || P < 0 || R < 0 || width <= 0 || height <= 0) return fail(GSR_ERR_INVALID_ARGUMENT, "bad argument");
    const int grid_x = (width + GSR_TILE_X - 1) / GSR_TILE_X, grid_y = (height + GSR_TILE_Y - 1) / GSR_TILE_Y;
    GeomState geom = GeomState::carve(geom_buffer ? (char*)align_up((size_t)geom_buffer, 256) : nullptr, P);
    BinState bin = BinState::carve(binning_buffer ? (char*)align_up((size_t)binning_buffer, 256) : nullptr, (size_t)R);
    ImageState img = ImageState::carve(image_buffer ? (char*)align_up((size_t)image_buffer, 256) : nullptr, (size_t)width * height,
                                       (size_t)grid_x * grid_y);
    out->point_list = bin.vals[0];
    out->ranges = reinterpret_cast<const uint32_t*>(img.ranges);
    out->n_contrib = img.n_contrib;
    out->final_T = img.final_T;
    out->packed = reinterpret_cast<const float*>(geom.packed);
    out->tiles_touched = geom.tiles_touched;
    out->clamped = geom.clamped;
    out->warp_mask = bin.warp_mask;
    return GSR_OK;
}

}  // extern "C"
