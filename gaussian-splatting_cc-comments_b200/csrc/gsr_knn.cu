// gsr_knn.cu -- mean squared distance to the 3 nearest neighbours of every point.
//
// Replaces `distCUDA2` of the reference's simple-knn extension (submodules/simple-knn/spatial.cu:15-26 ->
// SimpleKNN::knn, simple_knn.cu:185-221), which GaussianModel.create_from_pcd uses once to initialise the scales
// (scene/gaussian_model.py:207).  The result is defined exactly -- for every point the three smallest squared distances
// to OTHER points (by index, so duplicates count with distance 0), averaged -- and does not depend on how the search is
// organised, so the organisation here is this library's own:
//
//   1. bounds       : one grid-stride reduction, atomics on order-preserving integer images of the floats; the bounds
//                     stay on the device (the reference reads both back to the host with two blocking copies).
//                     Like the reference's reductions (init = {0,0,0}, simple_knn.cu:190-199) they include the origin.
//   2. Morton codes : 3 x 10 bits (prepMorton / coord2Morton, simple_knn.cu:45-63).
//   3. sort         : this library's stable radix sort (gsr_binning.cu), 4 passes over (code, index).
//   4. gather       : points copied into Morton order as 16-byte records (x, y, z, original index), one bounding box per
//                     KNN_BOX consecutive records.  Everything after this reads coalesced 16-byte records -- the
//                     reference gathers `points[indices[i]]` inside its innermost loop.
//   5. search       : one block per box.  The box's points first meet each other and the two boxes next to them in
//                     Morton order; the largest third-best distance in the block then bounds which other boxes can
//                     matter for ANY of its points (box-to-box test, one candidate box per thread, ballot-compacted);
//                     surviving boxes are staged in shared memory once per block and each thread tests its own
//                     point-to-box distance against its current third-best before walking the 256 staged points.
//                     The reference walks all P/1024 boxes per point from global memory.
//
// Pruning is exact: box distances are evaluated with the same monotone expression as point distances
// (dx*dx, then two FMAs), so a box bound can never exceed the computed distance of a point inside it.
#include "gsr_internal.cuh"
#include <cfloat>

namespace gsr {

constexpr int KNN_BOX = 256;            // points per box = threads per search block

// ---- order-preserving float <-> int (signed compare of the image == float compare; image(0.0f) == 0) ----
__device__ __forceinline__ int f2ord(float f) { const int b = __float_as_int(f); return b >= 0 ? b : b ^ 0x7FFFFFFF; }
__device__ __forceinline__ float ord2f(int o) { return __int_as_float(o >= 0 ? o : o ^ 0x7FFFFFFF); }

// bounds[0..2] = min(0, x/y/z), bounds[3..5] = max(0, x/y/z); zeroed by the caller (image of 0.0f is 0)
__global__ void __launch_bounds__(256) knn_bounds_kernel(int P, const float* __restrict__ pts, int* __restrict__ bounds) {
    float lo[3] = {0.f, 0.f, 0.f}, hi[3] = {0.f, 0.f, 0.f};
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)P; i += (size_t)gridDim.x * blockDim.x) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float v = __ldg(pts + 3 * i + a);
            lo[a] = fminf(lo[a], v);      // NaN coordinates are ignored by fminf/fmaxf
            hi[a] = fmaxf(hi[a], v);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            if (lo[a] < 0.f) atomicMin(bounds + a, f2ord(lo[a]));
            if (hi[a] > 0.f) atomicMax(bounds + 3 + a, f2ord(hi[a]));
        }
    }
}

__device__ __forceinline__ uint32_t spread10(uint32_t x) {      // prepMorton, simple_knn.cu:45-52
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

__global__ void __launch_bounds__(256) knn_morton_kernel(int P, const float* __restrict__ pts, const int* __restrict__ bounds,
                                                         uint32_t* __restrict__ codes, uint32_t* __restrict__ ids) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)P) return;
    uint32_t q[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const float lo = ord2f(bounds[a]), hi = ord2f(bounds[3 + a]);
        const float ext = hi - lo;
        const float t = ext > 0.f ? (__ldg(pts + 3 * i + a) - lo) / ext * 1023.f : 0.f;
        // float -> unsigned conversion saturates and maps NaN to 0; the clamp keeps the code inside 10 bits
        q[a] = min(__float2uint_rz(t), 1023u);
    }
    codes[i] = spread10(q[0]) | (spread10(q[1]) << 1) | (spread10(q[2]) << 2);
    ids[i] = (uint32_t)i;
}

// sorted record = (x, y, z, original index); box b = componentwise min / max of records [b*KNN_BOX, (b+1)*KNN_BOX)
__global__ void __launch_bounds__(KNN_BOX) knn_gather_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ order,
                                                             float4* __restrict__ sorted, float4* __restrict__ boxes) {
    const size_t i = (size_t)blockIdx.x * KNN_BOX + threadIdx.x;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < (size_t)P) {
        const uint32_t id = order[i];
        const float x = __ldg(pts + 3 * (size_t)id), y = __ldg(pts + 3 * (size_t)id + 1), z = __ldg(pts + 3 * (size_t)id + 2);
        sorted[i] = make_float4(x, y, z, __uint_as_float(id));
        lo[0] = hi[0] = x; lo[1] = hi[1] = y; lo[2] = hi[2] = z;
    }
    __shared__ float s_lo[KNN_BOX / 32][3], s_hi[KNN_BOX / 32][3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
        if ((threadIdx.x & 31) == 0) { s_lo[threadIdx.x >> 5][a] = lo[a]; s_hi[threadIdx.x >> 5][a] = hi[a]; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            float l = s_lo[0][a], h = s_hi[0][a];
            for (int w = 1; w < KNN_BOX / 32; w++) { l = fminf(l, s_lo[w][a]); h = fmaxf(h, s_hi[w][a]); }
            lo[a] = l; hi[a] = h;
        }
        boxes[2 * (size_t)blockIdx.x] = make_float4(lo[0], lo[1], lo[2], 0.f);
        boxes[2 * (size_t)blockIdx.x + 1] = make_float4(hi[0], hi[1], hi[2], 0.f);
    }
}

// squared length of (dx, dy, dz), rounded as the point distances are: one product, two FMAs
// The contraction of the reference's sm_100 binary (boxMeanDist: FMUL on the y difference, then FFMA x, then FFMA z --
// nvcc evaluates d.x*d.x + d.y*d.y + d.z*d.z as fma(dz, dz, fma(dx, dx, dy*dy)); same order in distBoxPoint).  Pinned by the
// fixture the reference build wrote on a B200 (tests/golden/ref_knn_cuda.npz).
__device__ __forceinline__ float sq3(float dx, float dy, float dz) { return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy))); }

// updateKBest<3> (simple_knn.cu:136-150): keep the three smallest, ascending; min/max network == its compare-and-swap chain
__device__ __forceinline__ void keep3(float d, float& b0, float& b1, float& b2) {
    float t = fmaxf(b0, d); b0 = fminf(b0, d);
    d = fmaxf(b1, t); b1 = fminf(b1, t);
    b2 = fminf(b2, d);
}

// SELF: the staged box is the thread's own, entry `skip` is the thread's point (exclusion is by index, as the reference's `i == idx`)
template <bool SELF>
__device__ __forceinline__ void scan_staged(const float4* s_pts, int n, int skip, const float4& me, float& b0, float& b1, float& b2) {
#pragma unroll 4
    for (int j = 0; j < n; j++) {
        if (SELF && j == skip) continue;
        const float4 c = s_pts[j];                                            // same address for the whole warp: broadcast
        // point - ref, as updateKBest; the clamp turns NaN / inf distances (non-finite input points) into "never inserted"
        keep3(fminf(sq3(c.x - me.x, c.y - me.y, c.z - me.z), FLT_MAX), b0, b1, b2);
    }
}

// distBoxPoint (simple_knn.cu:121-131) in monotone form: per-axis gap to the box, 0 inside
__device__ __forceinline__ float box_point_d2(const float4& lo, const float4& hi, const float4& p) {
    const float dx = fmaxf(0.f, fmaxf(lo.x - p.x, p.x - hi.x));
    const float dy = fmaxf(0.f, fmaxf(lo.y - p.y, p.y - hi.y));
    const float dz = fmaxf(0.f, fmaxf(lo.z - p.z, p.z - hi.z));
    return sq3(dx, dy, dz);
}
__device__ __forceinline__ float box_box_d2(const float4& alo, const float4& ahi, const float4& blo, const float4& bhi) {
    const float dx = fmaxf(0.f, fmaxf(alo.x - bhi.x, blo.x - ahi.x));
    const float dy = fmaxf(0.f, fmaxf(alo.y - bhi.y, blo.y - ahi.y));
    const float dz = fmaxf(0.f, fmaxf(alo.z - bhi.z, blo.z - ahi.z));
    return sq3(dx, dy, dz);
}

__global__ void __launch_bounds__(KNN_BOX) knn_search_kernel(int P, const float4* __restrict__ sorted, const float4* __restrict__ boxes,
                                                             int n_boxes, float* __restrict__ mean_d2) {
    __shared__ float4 s_pts[KNN_BOX];
    __shared__ float4 s_box[2];
    __shared__ int s_cand[KNN_BOX];
    __shared__ int s_warp_n[KNN_BOX / 32];
    __shared__ float s_warp_r[KNN_BOX / 32];
    __shared__ float s_reject;

    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t first = (size_t)b * KNN_BOX;
    const bool valid = first + tid < (size_t)P;
    const float4 me = valid ? sorted[first + tid] : make_float4(0.f, 0.f, 0.f, 0.f);
    float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;

    // ---- seed: own box, then the boxes before and after it in Morton order ----
    for (int s = 0; s < 3; s++) {
        const int cb = s == 0 ? b : (s == 1 ? b - 1 : b + 1);
        if (cb < 0 || cb >= n_boxes) continue;                                // uniform over the block
        const size_t cfirst = (size_t)cb * KNN_BOX;
        const int cn = (int)min((size_t)KNN_BOX, (size_t)P - cfirst);
        __syncthreads();                                                       // previous contents no longer read
        if (tid < cn) s_pts[tid] = sorted[cfirst + tid];
        __syncthreads();
        if (valid) { if (s == 0) scan_staged<true>(s_pts, cn, tid, me, b0, b1, b2); else scan_staged<false>(s_pts, cn, -1, me, b0, b1, b2); }
    }

    // ---- the block's rejection radius: the largest third-best among its points ----
    float r = valid ? b2 : 0.f;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) r = fmaxf(r, __shfl_xor_sync(0xffffffffu, r, o));
    if (lane == 0) s_warp_r[warp] = r;
    __syncthreads();
    if (tid == 0) {
        float m = s_warp_r[0];
        for (int w = 1; w < KNN_BOX / 32; w++) m = fmaxf(m, s_warp_r[w]);
        s_reject = m;
    }
    __syncthreads();
    const float reject = s_reject;
    const float4 own_lo = boxes[2 * (size_t)b], own_hi = boxes[2 * (size_t)b + 1];

    // ---- every other box: box-to-box test (one box per thread), survivors staged and walked ----
    for (int base = 0; base < n_boxes; base += KNN_BOX) {
        const int c = base + tid;
        bool pass = false;
        if (c < n_boxes && (c < b - 1 || c > b + 1)) {
            const float4 lo = boxes[2 * (size_t)c], hi = boxes[2 * (size_t)c + 1];
            pass = box_box_d2(own_lo, own_hi, lo, hi) <= reject;
        }
        const unsigned bal = __ballot_sync(0xffffffffu, pass);
        if (lane == 0) s_warp_n[warp] = __popc(bal);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < KNN_BOX / 32; w++) { const int nw = s_warp_n[w]; if (w < warp) before += nw; total += nw; }
        if (pass) s_cand[before + __popc(bal & ((1u << lane) - 1u))] = c;
        __syncthreads();                                                       // s_cand complete; s_warp_n reads done
        for (int k = 0; k < total; k++) {
            const int cb = s_cand[k];
            const size_t cfirst = (size_t)cb * KNN_BOX;
            const int cn = (int)min((size_t)KNN_BOX, (size_t)P - cfirst);
            __syncthreads();                                                   // staged points of the previous box no longer read
            if (tid < cn) s_pts[tid] = sorted[cfirst + tid];
            if (tid < 2) s_box[tid] = boxes[2 * (size_t)cb + tid];
            __syncthreads();
            // the reference's test (simple_knn.cu:174-176): skip boxes farther than the current third-best
            if (valid && !(box_point_d2(s_box[0], s_box[1], me) > b2)) scan_staged<false>(s_pts, cn, -1, me, b0, b1, b2);
        }
        __syncthreads();                                                       // before s_cand / s_warp_n are rewritten
    }

    if (valid) mean_d2[__float_as_uint(me.w)] = (b0 + b1 + b2) / 3.0f;        // simple_knn.cu:182
}

size_t knn_boxes(size_t P) { return (P + KNN_BOX - 1) / KNN_BOX; }

KnnState KnnState::carve(char* base, size_t P) {
    KnnState s; Carver c(base);
    s.keys[0] = c.take<uint32_t>(P); s.keys[1] = c.take<uint32_t>(P);
    s.vals[0] = c.take<uint32_t>(P); s.vals[1] = c.take<uint32_t>(P);
    s.table = c.take<uint32_t>(rs_table_words(P));
    s.sorted = c.take<float4>(P);
    s.boxes = c.take<float4>(2 * knn_boxes(P));
    s.bounds = c.take<int>(8);
    s.bytes = align_up(c.off, 256) + 256;
    return s;
}

void launch_knn(int P, const float* points, float* mean_d2, const KnnState& s, cudaStream_t st) {
    if (P <= 0) return;
    cudaMemsetAsync(s.bounds, 0, 8 * sizeof(int), st);
    const unsigned rb = (unsigned)min((size_t)(148 * 8), ((size_t)P + 255) / 256);
    knn_bounds_kernel<<<rb, 256, 0, st>>>(P, points, s.bounds);
    knn_morton_kernel<<<(unsigned)(((size_t)P + 255) / 256), 256, 0, st>>>(P, points, s.bounds, s.keys[0], s.vals[0]);
    count_launch(2);
    const int shifts[4] = {0, 8, 16, 24}, bits[4] = {8, 8, 8, 6};            // 30-bit codes
    radix_sort_u32(s.keys, s.vals, 0, (size_t)P, 4, shifts, bits, s.table, st);   // 4 passes: result back in keys[0] / vals[0]
    const unsigned nb = (unsigned)knn_boxes((size_t)P);
    knn_gather_kernel<<<nb, KNN_BOX, 0, st>>>(P, points, s.vals[0], s.sorted, s.boxes);
    knn_search_kernel<<<nb, KNN_BOX, 0, st>>>(P, s.sorted, s.boxes, (int)nb, mean_d2);
    count_launch(2);
}

}  // namespace gsr
