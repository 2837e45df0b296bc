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
#include <sys/mman.h>

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

struct BlockCtx {
    Barrier block;
    Barrier warp[32];
    uint32_t xchg[32][32];         // per warp, per lane exchange slots
};

extern BlockCtx g_ctx;
extern thread_local int t_lane, t_warp;

// Dynamic shared memory of the running block: one 96 KiB mapping BELOW 4 GiB, because the kernels address it through
// 32-bit shared addresses (csrc/gsr_smem.cuh: smem_addr / lds128 / sts128).
inline float* dyn_smem() {
    static float* p = [] {
        void* m = mmap(nullptr, 96 * 1024, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_32BIT, -1, 0);
        if (m == MAP_FAILED) { std::perror("emu: mmap(MAP_32BIT)"); std::abort(); }
        return static_cast<float*>(m);
    }();
    return p;
}

template <typename F> void launch(dim3 grid, dim3 block, F body);

}  // namespace emu

extern thread_local uint3 threadIdx, blockIdx;
extern thread_local dim3 blockDim, gridDim;

// ---- intrinsics used by the kernels under test -----------------------------------------------------------------
inline void __syncthreads() { emu::g_ctx.block.wait(); }

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
    const int nt = (int)block.x;
    for (unsigned bx = 0; bx < grid.x; bx++) {
        g_ctx.block.reset(nt);
        for (int w = 0; w < (nt + 31) / 32; w++) g_ctx.warp[w].reset(min(32, nt - 32 * w));
        std::vector<std::thread> th;
        th.reserve(nt);
        for (int t = 0; t < nt; t++)
            th.emplace_back([&, t] {
                threadIdx = {(unsigned)t, 0, 0}; blockIdx = {bx, 0, 0}; blockDim = block; gridDim = grid;
                t_lane = t & 31; t_warp = t >> 5;
                body();
                g_ctx.warp[t_warp].leave();
                g_ctx.block.leave();
            });
        for (auto& x : th) x.join();
    }
}
}  // namespace emu
