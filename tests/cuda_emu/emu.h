// emu.h -- a tiny host-side stand-in for the CUDA execution model, TEST INFRASTRUCTURE ONLY.
//
// Lets tests/cuda_emu/emu_next.cpp compile the product's kernel sources (gsr_adam.cu, gsr_knn.cu) with plain g++ and
// run them on the CPU: one OS thread per CUDA thread, blocks one after another, __syncthreads() as a real barrier,
// warp shuffles / ballots as rendezvous between the 32 threads of a warp.  It checks kernel LOGIC (indexing, barriers,
// reductions, pruning) in the container that has no GPU; it says nothing about performance and it is sequentially
// consistent, so it cannot show memory-ordering bugs.  The GPU parity tests remain the gate.
#pragma once
#include <cuda_runtime.h>      // vector types, cudaStream_t (host-only view of the header)

#include <atomic>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>
#include <cstdio>
#include <cstdlib>

#undef __shared__
#define __shared__ static           // blocks run one at a time, so one static instance == the block's shared memory
#undef __launch_bounds__
#define __launch_bounds__(...)
#undef __grid_constant__
#define __grid_constant__
#undef __global__
#define __global__
#undef __device__
#define __device__
#undef __forceinline__
#define __forceinline__ inline
#undef __align__
#define __align__(n) __attribute__((aligned(n)))
// a file-scope device array is a host array here, and its "device address" is its address
template <class T> inline cudaError_t cudaGetSymbolAddress(void** p, T& symbol) { *p = (void*)&symbol; return cudaSuccess; }

namespace emu {

struct Barrier {                   // reusable barrier over the LIVE threads of a group (threads that returned drop out)
    std::mutex mu; std::condition_variable cv; int live = 0, arrived = 0; unsigned long gen = 0;
    void reset(int n) { live = n; arrived = 0; gen = 0; }
    void wait() {
        std::unique_lock<std::mutex> lk(mu);
        const unsigned long g = gen;
        if (++arrived == live) { arrived = 0; gen++; cv.notify_all(); }
        else cv.wait(lk, [&] { return gen != g; });
    }
    void leave() {
        std::unique_lock<std::mutex> lk(mu);
        live--;
        if (live > 0 && arrived == live) { arrived = 0; gen++; cv.notify_all(); }
    }
};

struct NamedBarrier {              // bar.sync / bar.arrive id, count: completes when `count` threads have arrived or synced
    std::mutex mu; std::condition_variable cv; int arrived = 0; unsigned long gen = 0;
    void arrive(int count) {
        std::unique_lock<std::mutex> lk(mu);
        if (++arrived == count) { arrived = 0; gen++; cv.notify_all(); }
    }
    void sync(int count) {
        std::unique_lock<std::mutex> lk(mu);
        const unsigned long g = gen;
        if (++arrived == count) { arrived = 0; gen++; cv.notify_all(); }
        else cv.wait(lk, [&] { return gen != g; });
    }
};

struct BlockCtx {
    Barrier block;
    NamedBarrier named[16];
    Barrier warp[32];
    uint32_t xchg[32][32];         // per warp, per lane exchange slots
    std::atomic<int> count{0};     // __syncthreads_count
};

extern BlockCtx g_ctx;
extern thread_local int t_lane, t_warp;

// "Shared memory" addressing: the kernels address shared memory through 32-bit shared addresses (csrc/gsr_smem.cuh:
// smem_addr / lds* / sts*).  On the host a shared address is the signed 32-bit distance from an anchor object of the same
// library to the byte -- static `__shared__` arrays and the dynamic window below all live in that library's data segment.
extern char g_smem_anchor;
inline uint32_t to_shared(const void* p) { return (uint32_t)(int32_t)(reinterpret_cast<const char*>(p) - &g_smem_anchor); }
inline char* from_shared(uint32_t a) { return &g_smem_anchor + (int32_t)a; }
extern float g_dyn_smem[24 * 1024];            // 96 KiB dynamic shared memory of the running block
inline float* dyn_smem() { return g_dyn_smem; }

template <typename F> void launch(dim3 grid, dim3 block, F body);

}  // namespace emu

extern thread_local uint3 threadIdx, blockIdx;
extern thread_local dim3 blockDim, gridDim;

// ---- intrinsics used by the kernels under test -----------------------------------------------------------------
inline void __syncthreads() { emu::g_ctx.block.wait(); }
inline void __threadfence_block() { std::atomic_thread_fence(std::memory_order_seq_cst); }

inline uint32_t emu_exchange(uint32_t v, int src_lane) {
    emu::BlockCtx& c = emu::g_ctx;
    c.xchg[emu::t_warp][emu::t_lane] = v;
    c.warp[emu::t_warp].wait();
    const uint32_t r = c.xchg[emu::t_warp][src_lane & 31];
    c.warp[emu::t_warp].wait();
    return r;
}
inline float __shfl_xor_sync(unsigned, float v, int mask) {
    uint32_t b; std::memcpy(&b, &v, 4);
    b = emu_exchange(b, emu::t_lane ^ mask);
    float r; std::memcpy(&r, &b, 4);
    return r;
}
inline unsigned __ballot_sync(unsigned, bool pred) {
    emu::BlockCtx& c = emu::g_ctx;
    c.xchg[emu::t_warp][emu::t_lane] = pred ? 1u : 0u;
    c.warp[emu::t_warp].wait();
    unsigned r = 0;
    for (int l = 0; l < 32; l++) r |= (c.xchg[emu::t_warp][l] & 1u) << l;
    c.warp[emu::t_warp].wait();
    return r;
}
inline void __syncwarp(unsigned = 0xffffffffu) { emu::g_ctx.warp[emu::t_warp].wait(); }
inline int __syncthreads_count(int pred) {
    emu::BlockCtx& c = emu::g_ctx;
    if (pred) c.count.fetch_add(1);
    c.block.wait();
    const int r = c.count.load();
    c.block.wait();
    if (threadIdx.x == 0 && threadIdx.y == 0 && threadIdx.z == 0) c.count.store(0);
    c.block.wait();
    return r;
}
inline uint32_t __shfl_sync(unsigned, uint32_t v, int src) { return emu_exchange(v, src); }
inline int __shfl_sync(unsigned, int v, int src) { return (int)emu_exchange((uint32_t)v, src); }
inline float __shfl_sync(unsigned, float v, int src) { uint32_t b; std::memcpy(&b, &v, 4); b = emu_exchange(b, src); float r; std::memcpy(&r, &b, 4); return r; }
inline uint32_t __shfl_xor_sync(unsigned, uint32_t v, int mask) { return emu_exchange(v, emu::t_lane ^ mask); }
inline int __shfl_xor_sync(unsigned, int v, int mask) { return (int)emu_exchange((uint32_t)v, emu::t_lane ^ mask); }
inline uint32_t __shfl_up_sync(unsigned, uint32_t v, unsigned d) { const uint32_t r = emu_exchange(v, emu::t_lane - (int)d); return emu::t_lane >= (int)d ? r : v; }
inline int __shfl_up_sync(unsigned, int v, unsigned d) { return (int)__shfl_up_sync(0xffffffffu, (uint32_t)v, d); }
inline double __shfl_xor_sync(unsigned, double v, int mask) {      // two 32-bit halves, as the hardware does it
    uint64_t b; std::memcpy(&b, &v, 8);
    const uint32_t lo = emu_exchange((uint32_t)b, emu::t_lane ^ mask), hi = emu_exchange((uint32_t)(b >> 32), emu::t_lane ^ mask);
    b = ((uint64_t)hi << 32) | lo;
    double r; std::memcpy(&r, &b, 8);
    return r;
}
inline int __all_sync(unsigned m, bool pred) { return __ballot_sync(m, pred) == 0xffffffffu; }
inline int __any_sync(unsigned m, bool pred) { return __ballot_sync(m, pred) != 0u; }
inline unsigned __reduce_or_sync(unsigned, unsigned v) { unsigned s = v; for (int o = 16; o >= 1; o >>= 1) s |= emu_exchange(s, emu::t_lane ^ o); return s; }
inline unsigned __reduce_max_sync(unsigned, unsigned v) { unsigned s = v; for (int o = 16; o >= 1; o >>= 1) { const unsigned t = emu_exchange(s, emu::t_lane ^ o); s = t > s ? t : s; } return s; }
inline int __reduce_max_sync(unsigned, int v) { int s = v; for (int o = 16; o >= 1; o >>= 1) { const int t = (int)emu_exchange((uint32_t)s, emu::t_lane ^ o); s = t > s ? t : s; } return s; }
inline unsigned __match_any_sync(unsigned, uint32_t v) {
    emu::BlockCtx& c = emu::g_ctx;
    c.xchg[emu::t_warp][emu::t_lane] = v;
    c.warp[emu::t_warp].wait();
    unsigned r = 0;
    for (int l = 0; l < 32; l++) r |= (unsigned)(c.xchg[emu::t_warp][l] == v) << l;
    c.warp[emu::t_warp].wait();
    return r;
}
inline int __ffs(int x) { return __builtin_ffs(x); }
inline int __clz(int x) { return x == 0 ? 32 : __builtin_clz((unsigned)x); }
inline unsigned __brev(unsigned x) { unsigned r = 0; for (int i = 0; i < 32; i++) r |= ((x >> i) & 1u) << (31 - i); return r; }
inline float atomicAdd(float* p, float v) {
    std::atomic_ref<float> a(*p); float old = a.load();
    while (!a.compare_exchange_weak(old, old + v)) {}
    return old;
}
inline unsigned __reduce_add_sync(unsigned, unsigned v) {
    unsigned s = v;
    for (int o = 16; o >= 1; o >>= 1) s += emu_exchange(s, emu::t_lane ^ o);
    return s;
}
inline unsigned atomicAdd(unsigned* p, unsigned v) { return std::atomic_ref<unsigned>(*p).fetch_add(v); }
inline void __trap() { std::abort(); }
inline float __frcp_rn(float a) { volatile float r = 1.0f / a; return r; }
inline int __popc(unsigned x) { return __builtin_popcount(x); }
template <typename T> inline T __ldg(const T* p) { return *p; }
inline int __float_as_int(float f) { int i; std::memcpy(&i, &f, 4); return i; }
inline float __int_as_float(int i) { float f; std::memcpy(&f, &i, 4); return f; }
inline float __fdividef(float a, float b) { return a / b; }
template <class F> inline cudaError_t cudaFuncSetAttribute(F*, cudaFuncAttribute, int) { return cudaSuccess; }   // launch hints mean nothing here
inline unsigned __float_as_uint(float f) { unsigned i; std::memcpy(&i, &f, 4); return i; }
inline float __uint_as_float(unsigned i) { float f; std::memcpy(&f, &i, 4); return f; }
inline float __fmaf_rn(float a, float b, float c) { return std::fmaf(a, b, c); }
inline float __fmul_rn(float a, float b) { volatile float r = a * b; return r; }
inline float __fadd_rn(float a, float b) { volatile float r = a + b; return r; }
inline float __fsub_rn(float a, float b) { volatile float r = a - b; return r; }
inline float __fdiv_rn(float a, float b) { volatile float r = a / b; return r; }
inline float __fsqrt_rn(float a) { return std::sqrt(a); }
inline double __dadd_rn(double a, double b) { volatile double r = a + b; return r; }
inline double __dmul_rn(double a, double b) { volatile double r = a * b; return r; }
inline double __fma_rn(double a, double b, double c) { return std::fma(a, b, c); }
inline float __double2float_rn(double a) { return (float)a; }
inline float __int2float_rn(int a) { return (float)a; }
inline int __float2int_rz(float f) {               // cvt.rzi.s32.f32: NaN -> 0, saturating
    if (!(f == f)) return 0;
    if (f >= 2147483648.f) return 2147483647;
    if (f <= -2147483648.f) return -2147483647 - 1;
    return (int)f;
}
inline unsigned __float2uint_rz(float f) {          // cvt.rzi.u32.f32: NaN -> 0, saturating
    if (!(f == f) || f <= 0.f) return 0u;
    if (f >= 4294967296.f) return 0xFFFFFFFFu;
    return (unsigned)f;
}
inline int atomicMin(int* p, int v) {
    std::atomic_ref<int> a(*p); int old = a.load();
    while (old > v && !a.compare_exchange_weak(old, v)) {}
    return old;
}
inline int atomicMax(int* p, int v) {
    std::atomic_ref<int> a(*p); int old = a.load();
    while (old < v && !a.compare_exchange_weak(old, v)) {}
    return old;
}
template <typename T> inline T min(T a, T b) { return a < b ? a : b; }
template <typename T> inline T max(T a, T b) { return a > b ? a : b; }

namespace emu {
template <typename F> void launch(dim3 grid, dim3 block, F body) {
    const int nt = (int)(block.x * block.y * block.z);
    for (unsigned bz = 0; bz < grid.z; bz++)
    for (unsigned by = 0; by < grid.y; by++)
    for (unsigned bx = 0; bx < grid.x; bx++) {
        g_ctx.block.reset(nt);
        g_ctx.count = 0;
        for (int w = 0; w < (nt + 31) / 32; w++) g_ctx.warp[w].reset(min(32, nt - 32 * w));
        std::vector<std::thread> th;
        th.reserve(nt);
        for (int t = 0; t < nt; t++)
            th.emplace_back([&, t] {
                threadIdx = {(unsigned)t % block.x, ((unsigned)t / block.x) % block.y, (unsigned)t / (block.x * block.y)};
                blockIdx = {bx, by, bz}; blockDim = block; gridDim = grid;
                t_lane = t & 31; t_warp = t >> 5;
                body();
                g_ctx.warp[t_warp].leave();
                g_ctx.block.leave();
            });
        for (auto& x : th) x.join();
    }
}
}  // namespace emu

// every emulation library defines these once
#define GSR_EMU_DEFINE_GLOBALS                                                                         \
    namespace emu { BlockCtx g_ctx; thread_local int t_lane, t_warp; char g_smem_anchor; float g_dyn_smem[24 * 1024]; } \
    thread_local uint3 threadIdx, blockIdx;                                                            \
    thread_local dim3 blockDim, gridDim;

// The handful of CUDA runtime calls the product's host code makes, on host memory (streams and events are no-ops: the
// emulation is synchronous), and the launch counter of gsr_internal.cuh.  Defined once per emulation library.
#define GSR_EMU_DEFINE_RUNTIME                                                                                                   \
    namespace gsr { static unsigned long long g_emu_launches = 0; void count_launch(int n) { g_emu_launches += (unsigned long long)n; } } \
    GSR_EMU_DEFINE_CUDART
#define GSR_EMU_DEFINE_CUDART                                                                                                    \
    extern "C" {                                                                                                                 \
    cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t) { std::memset(p, v, n); return cudaSuccess; }            \
    cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t) { std::memcpy(d, s, n); return cudaSuccess; } \
    cudaError_t cudaGetLastError(void) { return cudaSuccess; }                                                                   \
    cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }                                                      \
    cudaError_t cudaGetDevice(int* d) { *d = 0; return cudaSuccess; }                                                            \
    cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { *s = reinterpret_cast<cudaStream_t>(1); return cudaSuccess; } \
    cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { *e = reinterpret_cast<cudaEvent_t>(1); return cudaSuccess; } \
    cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = reinterpret_cast<cudaEvent_t>(1); return cudaSuccess; }                   \
    cudaError_t cudaHostAlloc(void** p, size_t n, unsigned) { *p = std::malloc(n); return *p ? cudaSuccess : cudaErrorMemoryAllocation; } \
    cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t) { return cudaSuccess; }                                               \
    cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned) { return cudaSuccess; }                                 \
    cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }                                                        \
    cudaError_t cudaEventQuery(cudaEvent_t) { return cudaSuccess; }                                                              \
    cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0.f; return cudaSuccess; }                     \
    const char* cudaGetErrorString(cudaError_t) { return "emulated CUDA runtime"; }                                              \
    }
