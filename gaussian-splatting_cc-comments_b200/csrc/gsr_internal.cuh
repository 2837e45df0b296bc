// gsr_internal.cuh -- private declarations shared by the sm_100a kernels and the C-ABI layer.
//
// Buffer layouts here are private to this implementation (the three byte buffers are opaque
// to callers, as in the reference: DGR/cuda_rasterizer/rasterizer_impl.h:33-81).
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <stdint.h>
#include <stddef.h>

#define GSR_TILE_X 16   // tile size is part of the observable contract (DGR config.h:16-17)
#define GSR_TILE_Y 16
#define GSR_TILE_PIX 256

namespace gsr {

// ---- per-Gaussian packed record consumed by both blending kernels (48 B, 3 x float4) ---------
//   a = (mean2D.x, mean2D.y, conic.x, conic.y)
//   b = (conic.z, power_thr, opacity, r)      a + b: everything up to the alpha test
//   c = (g, b, ext_x, ext_y)                  c.xy: only for contributing pairs; c.zw: only for the warp cull
// ext_x/ext_y: conservative half-extents (pixels) of the region where alpha can reach 1/255
//              (negative => can never contribute); power_thr: `power < power_thr` implies
//              alpha < 1/255 without evaluating exp.
struct alignas(16) PackedRec { float4 a, b, c; };

// ---- carve helpers ---------------------------------------------------------------------------
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct Carver {
    char* base; size_t off;
    explicit Carver(char* b) : base(b), off(0) {}
    template <typename T> T* take(size_t count) {
        off = align_up(off, 256);
        T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
        off += count * sizeof(T);
        return p;
    }
};

constexpr int GSR_MAX_DEVICES = 64;
constexpr int RS_THREADS = 256;          // radix sort: threads per block
constexpr int RS_ITEMS = 16;             // items per thread
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;   // 4096 items per block
constexpr int RS_BINS = 256;             // 8-bit digits
constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;   // 4096

static inline size_t rs_blocks(size_t n) { return (n + RS_TILE - 1) / RS_TILE; }
static inline size_t rs_table_words(size_t n) { return 1280 + 4 * rs_blocks(n) * RS_BINS; }   // see radix_sort_u32
static inline size_t scan_blocks(size_t n) { return (n + SCAN_TILE - 1) / SCAN_TILE; }
// Counting tile sort (tile_count_sort, gsr_binning.cu): one pass over the instance list with the per-tile write cursors in
// shared memory, for grids of at most TC_MAX_TILES tiles.  The list is cut into contiguous chunks (one
// block each, tc_blocks_per_sm per SM); table = [totals | per-chunk counts, then per-chunk bases].
constexpr int TC_MAX_TILES = 32768;                      // 128 KB of cursors: 3840x2160 is 32 400 tiles
constexpr int TC_THREADS = 256;
constexpr int TC_EPT = 4;                                // instances per thread and round of the scatter pass
constexpr int TC_ROUND = TC_THREADS * TC_EPT;
// chunks (blocks) per SM: what the cursors leave room for.  Measured on B200 at R = 12 M, 8160 tiles: 2-4 per SM 0.153-0.155 ms,
// 6: 0.169 ms (count table, column scan and prologue traffic grow with the number of chunks)
static inline int tc_blocks_per_sm(size_t tiles) { return tiles <= 8192 ? 3 : (tiles <= 16384 ? 2 : 1); }
static inline size_t tc_chunk(size_t R, size_t tiles) {
    const size_t max_chunks = 148 * (size_t)tc_blocks_per_sm(tiles);
    size_t c = (R + max_chunks - 1) / max_chunks;
    if (c < 4 * (size_t)TC_ROUND) c = 4 * (size_t)TC_ROUND;
    return (c + TC_ROUND - 1) / TC_ROUND * TC_ROUND;
}
static inline size_t tc_chunks(size_t R, size_t tiles) { return (R + tc_chunk(R, tiles) - 1) / tc_chunk(R, tiles); }
constexpr size_t TC_ROWS_AT = (size_t)TC_MAX_TILES;               // [per-tile totals | one row of counts, then bases, per chunk]
// upper bound over all grids: chunks x padded tiles is at most 444 x 8192, 296 x 16384 or 148 x 32768 words (19 MB)
static inline size_t tc_table_words(size_t R) {
    size_t rows = 0;
    const size_t t[3] = {8192, 16384, 32768};
    for (int i = 0; i < 3; i++) rows = rows > tc_chunks(R, t[i]) * t[i] ? rows : tc_chunks(R, t[i]) * t[i];
    return TC_ROWS_AT + rows;
}

struct GeomState {
    float4* packed;            // 3P float4
    uint32_t* tiles_touched;   // P (band-clipped count)
    uint8_t* clamped;          // P
    uint32_t* keys[2];         // P each: depth bits (0xFFFFFFFF = emits nothing), ping-pong
    uint32_t* vals[2];         // P each: Gaussian ids, ping-pong
    uint2* rects;              // P: (x0 | y0 << 16, width) of the band-clipped tile rectangle, for Gaussians that emit instances
    uint2* rect_sorted;        // P: the same in depth order
    uint32_t* tt_sorted;       // P: tiles_touched in depth order
    uint32_t* offsets;         // P: exclusive scan of tt_sorted
    uint32_t* table;           // rs_blocks(P) * RS_BINS + 512 (digit histogram, ticket, look-back table)
    uint32_t* scan_tmp;        // block sums for the scans over P-sized / table-sized arrays
    uint32_t* total;           // 1: num_rendered on the device
    int* radii_internal;       // P: used when the caller passes radii == NULL (rasterizer_impl.cu:262-265)
    size_t bytes;
    static GeomState carve(char* base, size_t P) {
        GeomState g; Carver c(base);
        g.packed = c.take<float4>(3 * P);
        g.tiles_touched = c.take<uint32_t>(P);
        g.clamped = c.take<uint8_t>(P);
        g.keys[0] = c.take<uint32_t>(P); g.keys[1] = c.take<uint32_t>(P);
        g.vals[0] = c.take<uint32_t>(P); g.vals[1] = c.take<uint32_t>(P);
        g.rects = c.take<uint2>(P); g.rect_sorted = c.take<uint2>(P);
        g.tt_sorted = c.take<uint32_t>(P);
        g.offsets = c.take<uint32_t>(P);
        size_t tbl = rs_table_words(P);
        g.table = c.take<uint32_t>(tbl);
        size_t big = tbl > P ? tbl : P;
        g.scan_tmp = c.take<uint32_t>(scan_blocks(big) + 1);
        g.total = c.take<uint32_t>(1);
        g.radii_internal = c.take<int>(P);
        g.bytes = align_up(c.off, 256) + 256;
        return g;
    }
};

struct BinState {
    uint32_t* vals[2];         // R each; vals[0] is the final point_list
    uint32_t* keys[2];         // R each: tile ids, ping-pong
    uint32_t* table;           // max(rs_table_words(R), tc_table_words(R)): radix look-back tables or the counting sort's chunk counts
    uint32_t* scan_tmp;
    uint8_t* warp_mask;        // R: bit w (0..3) = 8x8 block w of the instance's tile blended it in the forward (backward's cull)
    size_t bytes;
    static BinState carve(char* base, size_t R) {
        BinState b; Carver c(base);
        b.vals[0] = c.take<uint32_t>(R); b.vals[1] = c.take<uint32_t>(R);
        b.keys[0] = c.take<uint32_t>(R); b.keys[1] = c.take<uint32_t>(R);
        size_t tbl = rs_table_words(R);
        if (tc_table_words(R) > tbl) tbl = tc_table_words(R);
        b.table = c.take<uint32_t>(tbl);
        b.scan_tmp = c.take<uint32_t>(scan_blocks(tbl) + 1);
        b.warp_mask = c.take<uint8_t>(R);
        b.bytes = align_up(c.off, 256) + 256;
        return b;
    }
};

struct ImageState {
    float* final_T;            // N
    uint32_t* n_contrib;       // N
    uint2* ranges;             // tiles
    size_t bytes;
    static ImageState carve(char* base, size_t N, size_t tiles) {
        ImageState s; Carver c(base);
        s.final_T = c.take<float>(N);
        s.n_contrib = c.take<uint32_t>(N);
        s.ranges = c.take<uint2>(tiles);
        s.bytes = align_up(c.off, 256) + 256;
        return s;
    }
};

void count_launch(int n);   // bumps the library-wide kernel launch counter (gsr_launch_count)
int set_error(int code, const std::string& msg);   // stores the text gsr_last_error() returns (thread-local), returns `code`

// ---- stage launchers (each enqueues on `st`, none synchronises) --------------------------------
struct Camera {             // device pointers, read once per block
    const float* view; const float* proj; const float* campos;
};

struct PreprocessArgs {
    int P, D, M;
    const float* means3D; const float* scales; float scale_modifier; const float* rotations;
    const float* opacities; const float* shs; const float* cov3D_precomp; const float* colors_precomp;
    Camera cam;
    int W, H; float tan_fovx, tan_fovy, focal_x, focal_y;
    int grid_x, grid_y, row_begin, row_end;
    int prefiltered;
    int* radii;                // may alias an internal array
    float4* packed; uint32_t* tiles_touched; uint8_t* clamped; uint32_t* depth_keys; uint32_t* ids;
    uint32_t* total;           // zeroed by the caller; receives the sum of tiles_touched (num_rendered)
    // raw-parameter path (gsr_fused_params.cuh); appended so that the reference-shaped kernel's parameter offsets stay put
    int raw_act;               // opacities / scales / rotations are logits / log-scales / unnormalised quaternions
    const float* f_dc;         // (P,1,3) and
    const float* f_rest;       // (P,M-1,3): SH coefficients in split storage (then shs == nullptr)
    uint2* rects;              // (x0 | y0 << 16, width) of the band-clipped tile rectangle of a Gaussian that emits instances
};
void launch_preprocess(const PreprocessArgs& a, cudaStream_t st);
void launch_mark_visible(int P, const float* means, const float* view, unsigned char* present, cudaStream_t st);

// exclusive prefix sum of n uint32 (in may equal out); tmp holds scan_blocks(n)+1 words;
// if total != nullptr the grand total is written there.
void scan_exclusive_u32(const uint32_t* in, uint32_t* out, size_t n, uint32_t* tmp, uint32_t* total, cudaStream_t st);

// stable LSD radix sort on up to 4 digit passes (bits[p] <= 8 key bits starting at shifts[p]); see gsr_binning.cu
void radix_sort_u32(uint32_t* const keys[2], uint32_t* const vals[2], int s0, size_t n, int npass, const int* shifts,
                    const int* bits, uint32_t* table, cudaStream_t st);

void launch_gather_u32(const uint32_t* src, const uint32_t* idx, uint32_t* dst, size_t n, cudaStream_t st);
void launch_gather_tt_rect(const uint32_t* tt, const uint2* rects, const uint32_t* idx, uint32_t* tt_sorted, uint2* rect_sorted, size_t n,
                           cudaStream_t st);

void launch_duplicate(int P, const uint32_t* order, const uint32_t* offsets, const uint32_t* tt_sorted, const uint2* rect_sorted,
                      const float4* packed, const int* radii, int grid_x, int grid_y, int row_begin, int row_end,
                      uint32_t* tile_keys, uint32_t* vals, cudaStream_t st);

void launch_tile_ranges(size_t R, const uint32_t* sorted_tile_keys, uint2* ranges, size_t tiles, cudaStream_t st);
// one-pass counting sort of the instance list by tile + the tile ranges (grids of at most TC_MAX_TILES tiles)
bool tile_count_sort_applies(size_t tiles);
void tile_count_sort(const uint32_t* keys, const uint32_t* vals, size_t R, size_t tiles, uint32_t* table, uint32_t* out, uint2* ranges,
                     cudaStream_t st);

void launch_render_forward(int grid_x, int row_begin, int row_end, const uint2* ranges, const uint32_t* point_list,
                           int W, int H, const float4* packed, const float* bg, float* out_color,
                           float* final_T, uint32_t* n_contrib, uint8_t* warp_mask, cudaStream_t st);

void launch_render_backward(int grid_x, int row_begin, int row_end, const uint2* ranges, const uint32_t* point_list,
                            int W, int H, const float4* packed, const float* bg, const float* final_T,
                            const uint32_t* n_contrib, const float* dL_dpix, float* grad_acc /* (P,12) */,
                            const uint8_t* warp_mask, cudaStream_t st);

struct PreprocessBwdArgs {
    int P, D, M;
    const float* means3D; const int* radii; const float* shs; const uint8_t* clamped;
    const float* scales; const float* rotations; float scale_modifier; const float* cov3D_precomp;
    Camera cam;
    int W, H; float tan_fovx, tan_fovy, focal_x, focal_y;
    const float* grad_acc;     // (P,12): sum w dx, w dy, w dx dx, w dx dy, w dy dy, dL_dopacity, dL_dcolour[3]
    const float4* packed;      // the forward's records (conic)
    float* dL_dmean2D; float* dL_dconic; float* dL_dopacity; float* dL_dcolor;
    float* dL_dmean3D; float* dL_dcov3D; float* dL_dsh; float* dL_dscale; float* dL_drot;
    float* dL_drgb;            // optional (P,3): clamp-masked colour gradient (multi-view exchange)
    int skip_dsh;              // leave dL_dsh untouched
    int begin;                 // first Gaussian of this launch (chunked launches); P is the end bound
    int chunked;               // a collective runs beside this kernel: launch it in chunks (chunk_plan)
    void* const* chunk_events; int n_chunk_events;   // optional cudaEvent_t per chunk (see include/gsr.h)
    float* small_pack;         // optional (P,14): mean3D 3 | mean2D 3 | opacity 1 | scale 3 | rot 4 instead of the five tensors
    // raw-parameter path (gsr_fused_params.cuh), appended
    int raw_act;               // scales / rotations hold raw parameters; dL_dscale / dL_drot / dL_dopacity are w.r.t. the raw values
    const float* f_dc; const float* f_rest;      // split SH storage (then shs == nullptr) ...
    float* dL_df_dc; float* dL_df_rest;          // ... and its gradients (P,1,3), (P,M-1,3) (then dL_dsh == nullptr)
};
size_t l1_ssim_blocks(int planes, int H, int W);
void launch_l1_ssim_forward(const float* img, const float* gt, int planes, int H, int W, float lambda, float* dm_maps, float* partial,
                            float* out3, cudaStream_t st);
void launch_l1_ssim_backward(const float* img, const float* gt, const float* dm_maps, int planes, int H, int W, float coef_l1,
                             float coef_ssim, const float* grad_out, float* dimg, cudaStream_t st);
void launch_multimem_allreduce(float* mc_base, size_t begin, size_t end, int rank, int world, int max_blocks, cudaStream_t st);
void launch_multimem_broadcast(float* mc_dst, const float* src, size_t n_floats, int max_blocks, cudaStream_t st);
void launch_unpack_small(int begin, int end, const float* pack, float* dmean3D, float* dmean2D, float* dop, float* dscale,
                         float* drot, cudaStream_t st);
void launch_preprocess_backward(const PreprocessBwdArgs& a, cudaStream_t st);
// masked colour gradient (P,3) from the blending accumulators: out = radii > 0 ? acc[6..8] with SH-clamped channels zeroed : 0
void launch_publish_drgb(int P, const float* grad_acc, const int* radii, const uint8_t* clamped /* NULL: no mask */, float* out,
                         cudaStream_t st);

struct MultiviewShArgs {
    int P, D, M, n_views;
    const float* blocks[16];   // per view: campos (4 floats) then (P,3) masked colour gradients, peer-mapped
    const float* means3D;
    float* dL_dsh;
    int begin;                 // first Gaussian of this launch (chunked launches); P is the end bound
    const float* campos;       // optional (n_views,4): camera positions; nullptr = the first 4 floats of each block
    int no_split;              // 1 = one launch for [begin, P) (the caller pipelines ranges itself)
};
void launch_multiview_sh_grad(const MultiviewShArgs& a, cudaStream_t st);

// ---- fused Adam step (gsr_adam.cu) -------------------------------------------------------------------
constexpr int ADAM_MAX_GROUPS = 8;         // parameter groups per launch (the model has six)
struct AdamGroupDev {
    float* p; const float* g; float* m; float* v;   // param, grad, exp_avg, exp_avg_sq
    unsigned long long n;                           // elements
    float neg_step;                                 // -lr / (1 - beta1^t)
    float bc2_sqrt;                                 // sqrt(1 - beta2^t)
    unsigned row_width;                             // elements per Gaussian (visible-mask mode)
};
struct AdamLaunch {
    AdamGroupDev g[ADAM_MAX_GROUPS];
    unsigned block_start[ADAM_MAX_GROUPS + 1];      // first block of each group; [n_groups] = grid size
    int n_groups;
    float beta2, w1, w2, eps;                       // w1 = 1 - beta1, w2 = 1 - beta2
    const uint8_t* visible;                         // optional (rows) mask: 0 = leave the row untouched
};
size_t adam_blocks(size_t n);
void launch_adam(const AdamLaunch& L, unsigned total_blocks, bool vec, cudaStream_t st);

// ---- densification: statistics, classify, apply (gsr_densify.cu) -----------------------------------------
constexpr int DENSIFY_THREADS = 256;       // Gaussians per block in classify and apply (the ranks are per block)
constexpr int DENSIFY_MAX_GROUPS = 8;
constexpr int DENSIFY_MAX_COPIES = 8;
struct DensifyThresholds {                 // the Python scalars of densify_and_prune, rounded to float32 as torch does
    float max_grad, dense_extent, big_extent, min_opacity, inv_shrink;
    int use_screen, copies;
};
struct DensifyState {
    uint32_t* block_offsets;               // 4 x nblocks: per-block counts, then exclusive offsets (kept originals, kept
                                           // clones, split parents, split parents with kept children)
    size_t bytes;
    static DensifyState carve(char* base, size_t P);
};
struct DensifyGroupDev {
    const float* p; const float* m; const float* v;
    float* po; float* mo; float* vo;
    int w;
};
struct DensifyApply {
    DensifyGroupDev g[DENSIFY_MAX_GROUPS];
    int n_groups, xyz_group, scaling_group, rotation_group;
    int P, n_split, n_ok, n_ck, n_chk;     // totals read back by the caller
    const uint8_t* flags;
    const uint32_t* block_offsets;
    const float* samples;
    DensifyThresholds t;
};
void launch_densify_stats(int P, const float* vs_grad, const int* radii, float* accum, float* denom, float* max_radii, cudaStream_t st);
void launch_densify_classify(int P, const float* accum, const float* denom, const float* scaling, const float* opacity,
                             const DensifyThresholds& t, uint8_t* flags, int* totals, const DensifyState& s, cudaStream_t st);
void launch_densify_apply(const DensifyApply& a, cudaStream_t st);

// ---- 3-nearest-neighbour mean squared distance (gsr_knn.cu) ------------------------------------------
struct KnnState {
    uint32_t* keys[2];         // P each: Morton codes, ping-pong
    uint32_t* vals[2];         // P each: point indices, ping-pong
    uint32_t* table;           // radix sort tables, rs_table_words(P)
    float4* sorted;            // P: (x, y, z, original index) in Morton order
    float4* boxes;             // 2 per box: componentwise min, max
    int* bounds;               // 6 (+2): order-preserving integer images of the scene bounds
    size_t bytes;
    static KnnState carve(char* base, size_t P);
};
size_t knn_boxes(size_t P);
void launch_knn(int P, const float* points, float* mean_d2, const KnnState& s, cudaStream_t st);

// Chunked launches for kernels that run beside a collective: a grid of tens of thousands of short blocks keeps every
// SM full, and the collective's few large CTAs (launched a moment later on another stream) are not placed until that
// grid drains, whatever the stream priorities.  A small head chunk followed by three equal ones gives the block
// scheduler drain points where the waiting CTAs get in.  Returns the number of chunks; bounds[i]..bounds[i+1].
static inline int chunk_plan(int P, int unit, int bounds[5]) {
    bounds[0] = 0;
    if (P < 64 * 1024) { bounds[1] = P; return 1; }
    const int head = ((P / 16 + unit - 1) / unit) * unit;
    const int part = (((P - head) / 3 + unit - 1) / unit) * unit;
    bounds[1] = head;
    bounds[2] = head + part < P ? head + part : P;
    bounds[3] = head + 2 * part < P ? head + 2 * part : P;
    bounds[4] = P;
    return 4;
}

}  // namespace gsr
