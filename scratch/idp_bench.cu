// micro-benchmark: issue rate of IDP.2A vs IMAD vs LOP3+SHF+IMAD on sm_100a (per SMSP, cycles per warp instruction)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(uint32_t* out, uint32_t w, uint32_t x0, long long* cyc) {
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x + i;
    uint32_t x = x0 + threadIdx.x;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < 4096; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) a[i] = __dp2a_lo(a[i], w, x);
            if (MODE == 1) a[i] = a[i] * w + x;
            if (MODE == 2) { a[i] = (a[i] & 0xFFFFu) * w + x; }
            if (MODE == 3) { a[i] = (a[i] & 0xFFFFu) * w + x; a[i] = (a[i] >> 16) * w + x; }
            if (MODE == 4) { a[i] = __dp2a_lo(a[i], w, x); a[i] = __dp2a_hi(a[i], w, x); }
        }
        
    }
    long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 1 << 24); cudaMallocManaged(&cyc, 8);
    for (int warps = 4; warps <= 32; warps *= 2) {
        for (int m = 0; m < 5; ++m) {
            if (m == 0) k<0><<<148, warps * 32>>>(out, 77, 12345, cyc);
            if (m == 1) k<1><<<148, warps * 32>>>(out, 77, 12345, cyc);
            if (m == 2) k<2><<<148, warps * 32>>>(out, 77, 12345, cyc);
            if (m == 3) k<3><<<148, warps * 32>>>(out, 77, 12345, cyc);
            if (m == 4) k<4><<<148, warps * 32>>>(out, 77, 12345, cyc);
            cudaDeviceSynchronize();
            // per SMSP: warps/4 warps each issuing 4096*8 "ops"
            double per = (double)*cyc / (4096.0 * 8 * (warps / 4.0));
            printf("warps/SM=%2d mode=%d cycles=%lld  cycles per op per SMSP=%.3f\n", warps, m, *cyc, per);
        }
    }
    return 0;
}
