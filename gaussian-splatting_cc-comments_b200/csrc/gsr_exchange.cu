// gsr_exchange.cu -- cross-GPU sums for multi-view training through the NVSwitch (NVLS multicast), sm_100a.
//
// No reference counterpart: the reference trains one view per step on one GPU.  With one camera per GPU the 14
// small gradient floats per Gaussian (gsr_backward_args.small_pack rows) have to be summed over the ranks.  The
// rows live in a buffer that every rank allocated symmetrically and that is bound to ONE multicast address:
//   * multimem.ld_reduce on that address returns the SUM of all ranks' copies of a 16-byte word -- the switch
//     pulls the word from every GPU and adds in flight;
//   * multimem.st on that address writes a word into EVERY rank's copy.
// Each rank does both for its 1/world slice of the range (two-shot all-reduce): per GPU about 1/world of the
// range comes in reduced and (world-1)/world comes in as broadcast, instead of the 2*(world-1)/world a ring
// moves, and the kernel is a handful of small CTAs that fit beside the per-Gaussian backward's blocks.
// The caller brackets the launch with device barriers (every rank's rows written before; every slice broadcast
// after).
#include "gsr_internal.cuh"
#include <cstdlib>

namespace gsr {

__device__ __forceinline__ float4 mm_ld_reduce_v4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st_v4(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float mm_ld_reduce_f32(const float* mc) {
    float v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st_f32(float* mc, float v) {
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" :: "l"(mc), "f"(v) : "memory");
}

// words [w0, w1) of 16 bytes belong to this rank; up to 3 trailing floats [t0, t1) go to the last rank
constexpr int MM_THREADS = 128;
constexpr int MM_UNROLL = 8;

__global__ void __launch_bounds__(MM_THREADS) multimem_allreduce_kernel(float* __restrict__ mc, size_t w0, size_t w1, size_t t0, size_t t1) {
    const size_t stride = (size_t)gridDim.x * MM_THREADS;
    size_t w = w0 + (size_t)blockIdx.x * MM_THREADS + threadIdx.x;
    // several independent reduced loads in flight per thread: the round trip goes through the switch to every GPU
    for (; w + (MM_UNROLL - 1) * stride < w1; w += MM_UNROLL * stride) {
        float4 v[MM_UNROLL];
#pragma unroll
        for (int u = 0; u < MM_UNROLL; u++) v[u] = mm_ld_reduce_v4(mc + 4 * (w + u * stride));
#pragma unroll
        for (int u = 0; u < MM_UNROLL; u++) mm_st_v4(mc + 4 * (w + u * stride), v[u]);
    }
    for (; w < w1; w += stride) mm_st_v4(mc + 4 * w, mm_ld_reduce_v4(mc + 4 * w));
    if (blockIdx.x == 0) {
        const size_t t = t0 + threadIdx.x;
        if (t < t1) mm_st_f32(mc + t, mm_ld_reduce_f32(mc + t));
    }
}

// One rank's contribution to an all-gather done by the switch: n4 16-byte words from local memory are stored once to the
// multicast address and land in EVERY rank's copy of the symmetric buffer (this rank's included).  Stores are fire-and-forget;
// four independent loads per thread keep the local read side busy.
__global__ void __launch_bounds__(256) multimem_broadcast_kernel(float* __restrict__ mc, const float4* __restrict__ src, size_t n4) {
    const size_t stride = (size_t)gridDim.x * 256;
    size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    for (; i + 3 * stride < n4; i += 4 * stride) {
        const float4 a = __ldg(src + i), b = __ldg(src + i + stride), c = __ldg(src + i + 2 * stride), d = __ldg(src + i + 3 * stride);
        mm_st_v4(mc + 4 * i, a);
        mm_st_v4(mc + 4 * (i + stride), b);
        mm_st_v4(mc + 4 * (i + 2 * stride), c);
        mm_st_v4(mc + 4 * (i + 3 * stride), d);
    }
    for (; i < n4; i += stride) mm_st_v4(mc + 4 * i, __ldg(src + i));
    __threadfence_system();      // the barrier kernel that follows on the stream signals the peers: order the stores before it
}

void launch_multimem_broadcast(float* mc_dst, const float* src, size_t n_floats, int max_blocks, cudaStream_t st) {
    const size_t n4 = n_floats / 4;
    if (n4 == 0) return;
    // Few CTAs: the stores are bounded by the NVLink port, not by SMs, and this kernel runs beside the per-Gaussian backward
    // (GSR_MC_BLOCKS overrides for sweeps).
    static const int env_blocks = [] { const char* e = getenv("GSR_MC_BLOCKS"); return e ? atoi(e) : 0; }();
    if (env_blocks > 0) max_blocks = env_blocks;
    size_t need = (n4 + 1023) / 1024;
    if (need < 1) need = 1;
    const unsigned grid = (unsigned)(need < (size_t)max_blocks ? need : (size_t)max_blocks);
    multimem_broadcast_kernel<<<grid, 256, 0, st>>>(mc_dst, reinterpret_cast<const float4*>(src), n4);
    count_launch(1);
}

void launch_multimem_allreduce(float* mc_base, size_t begin, size_t end, int rank, int world, int max_blocks, cudaStream_t st) {
    if (end <= begin) return;
    // 16-byte words of the range (begin is 16-byte aligned by contract), split evenly over the ranks
    const size_t words = (end - begin) / 4;
    const size_t per = (words + world - 1) / world;
    const size_t w0 = begin / 4 + (size_t)rank * per < begin / 4 + words ? begin / 4 + (size_t)rank * per : begin / 4 + words;
    const size_t w1 = w0 + per < begin / 4 + words ? w0 + per : begin / 4 + words;
    const bool last = rank == world - 1;
    const size_t t0 = begin + 4 * words, t1 = last ? end : t0;
    size_t need = (w1 - w0 + MM_THREADS * MM_UNROLL - 1) / (MM_THREADS * MM_UNROLL);
    if (need < 1) need = 1;
    const unsigned grid = (unsigned)(need < (size_t)max_blocks ? need : (size_t)max_blocks);
    multimem_allreduce_kernel<<<grid, MM_THREADS, 0, st>>>(mc_base, w0, w1, t0, t1);
    count_launch(1);
}

}  // namespace gsr
