// common.cuh -- shared helpers for libmsr.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include <string>

#include "../../include/msr.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libmsr is written for sm_100a (B200) only"
#endif

namespace msr {

void set_error(const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;
inline void count_launch(uint64_t n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define MSR_CUDA_CHECK(expr)                                                                      \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess) {                                                                  \
            ::msr::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,                 \
                             cudaGetErrorString(_e));                                             \
            return MSR_ERR_CUDA;                                                                  \
        }                                                                                         \
    } while (0)

#define MSR_REQUIRE(cond, ...)                                                                    \
    do {                                                                                          \
        if (!(cond)) {                                                                            \
            ::msr::set_error(__VA_ARGS__);                                                        \
            return MSR_ERR_INVALID;                                                               \
        }                                                                                         \
    } while (0)

int use_device(int device);   // cudaSetDevice + sm_100 check; returns msr_status

// cudaFuncSetAttribute is per DEVICE, not per process: launch helpers remember, per call site, on which devices the
// attribute has been set (bit = device index), so a process that drives several GPUs sets it on each.
inline bool attr_needed(const std::atomic<uint64_t>& done) {
    int dev = 0;
    cudaGetDevice(&dev);
    return dev > 63 || !((done.load(std::memory_order_relaxed) >> dev) & 1ull);
}
inline void attr_mark(std::atomic<uint64_t>& done) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev <= 63) done.fetch_or(1ull << dev, std::memory_order_relaxed);
}

// RAII device buffer used for temporaries inside API calls
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) {
        if (p) { cudaFree(p); p = nullptr; }
        bytes = n;
        return cudaMalloc(&p, n ? n : 1);
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// grow-only workspace owned by a handle
struct Workspace {
    void* p = nullptr;
    size_t bytes = 0;
    cudaError_t reserve(size_t n, cudaStream_t s) {
        if (n <= bytes) return cudaSuccess;
        if (p) {
            cudaError_t e = cudaStreamSynchronize(s);
            if (e != cudaSuccess) return e;
            cudaFree(p);
            p = nullptr; bytes = 0;
        }
        cudaError_t e = cudaMalloc(&p, n);
        if (e == cudaSuccess) bytes = n;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
};

// ---------------------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------------------
#ifdef __CUDACC__

// order-preserving float32 <-> uint32 (larger float -> larger uint)
__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t b = __float_as_uint(f);
    return b ^ ((b >> 31) ? 0xFFFFFFFFu : 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
    uint32_t b = u ^ ((u >> 31) ? 0x80000000u : 0xFFFFFFFFu);
    return __uint_as_float(b);
}
// ranking key: larger key ranks first; ties in score -> smaller doc id first
__device__ __forceinline__ uint64_t make_key(float score, uint32_t doc) {
    return ((uint64_t)f2ord(score) << 32) | (uint64_t)(0xFFFFFFFFu - doc);
}
__device__ __forceinline__ float key_score(uint64_t k) { return ord2f((uint32_t)(k >> 32)); }
__device__ __forceinline__ uint32_t key_doc(uint64_t k) { return 0xFFFFFFFFu - (uint32_t)k; }

__device__ __forceinline__ uint32_t warp_reduce_max(uint32_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ uint32_t warp_reduce_add(uint32_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// In-place bitonic sort of n (power of two) u64 keys in shared memory, DESCENDING.  All threads of the CTA call.
__device__ __forceinline__ void bitonic_sort_desc(uint64_t* a, uint32_t n) {
    for (uint32_t size = 2; size <= n; size <<= 1) {
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (uint32_t t = threadIdx.x; t < (n >> 1); t += blockDim.x) {
                uint32_t lo = 2 * t - (t & (stride - 1));
                uint32_t hi = lo + stride;
                bool desc = ((lo & size) == 0);
                uint64_t x = a[lo], y = a[hi];
                if ((x < y) == desc) { a[lo] = y; a[hi] = x; }
            }
        }
    }
    __syncthreads();
}

#endif  // __CUDACC__

}  // namespace msr
