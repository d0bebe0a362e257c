// microbench.cu -- measured ceilings of the resources the sparse scorer runs on (msr_microbench).
//
// The scorer shares every posting between thousands of queries through L2, so HBM does not bound it;
// its `roofline.binding` is reported against ceilings MEASURED here, on the same GPU, in the same bench run:
//   MSR_MB_L2_READ    L2 -> SM read bandwidth of an L2-resident buffer, with the dense phase's own access pattern:
//                     8-byte ld.global.nc.L1::no_allocate, 8 loads in flight per thread, 512-thread CTAs, 2 per SM
//   MSR_MB_HBM_READ   the same loop over a buffer far larger than L2 (HBM stream; cross-check of MEASURED_PEAKS.json)
//   MSR_MB_RED_SHARED red.shared.add.u32 warp instructions/s, one bank per lane (the conflict-free bucket layout)
//   MSR_MB_ISSUE      warp instructions/s of independent dp4a chains (fma pipe: one every 2 cycles per scheduler)
//   MSR_MB_ALU        warp instructions/s of independent PRMT chains (alu pipe)
//   MSR_MB_LDS        shared-memory wavefronts/s of conflict-free 16-byte loads (l1tex data pipe)
//   MSR_MB_IMMA       mma.sync m16n8k32 u8 instructions/s (legacy warp-level integer tensor op), alone / with 8 PRMT each
// Each call times `iters` launches with CUDA events after one warm-up launch and returns units/s.
#include "common.cuh"

namespace msr {

constexpr int kMbThreads = 512;

__global__ void __launch_bounds__(kMbThreads, 2)
mb_read_kernel(const uint2* __restrict__ buf, size_t n_vec, uint32_t passes, uint32_t* __restrict__ sink) {
    // grid-stride over the buffer, 8 independent 8-byte loads per thread and step
    const size_t stride = (size_t)gridDim.x * kMbThreads;
    uint32_t acc = 0;
    for (uint32_t p = 0; p < passes; ++p) {
        size_t i = (size_t)blockIdx.x * kMbThreads + threadIdx.x;
        for (; i + 7 * stride < n_vec; i += 8 * stride) {
            uint2 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u)
                asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(v[u].x), "=r"(v[u].y) : "l"(buf + i + u * stride));
#pragma unroll
            for (int u = 0; u < 8; ++u) acc += v[u].x ^ v[u].y;
        }
    }
    if (acc == 0x12345u) *sink = acc;
}

__global__ void __launch_bounds__(kMbThreads, 2)
mb_red_shared_kernel(uint32_t rounds, uint32_t* __restrict__ sink) {
    __shared__ uint32_t acc[8192];
    for (uint32_t i = threadIdx.x; i < 8192; i += kMbThreads) acc[i] = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(acc) + 4u * lane;
    uint32_t row = warp * 37u;
    for (uint32_t r = 0; r < rounds; ++r) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            // one wavefront: lane l hits bank l of a varying row
            asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(base + 128u * ((row + u * 11u) & 255u)), "r"(r + u) : "memory");
        }
        row += 97u;
    }
    __syncthreads();
    if (acc[threadIdx.x] == 0xFFFFFFFFu) *sink = 1;
}

__global__ void __launch_bounds__(kMbThreads, 2)
mb_issue_kernel(uint32_t rounds, uint32_t seed, uint32_t* __restrict__ sink) {
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x + i;
    const uint32_t v = seed | threadIdx.x, b = seed >> 3;
    for (uint32_t r = 0; r < rounds; ++r) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = __dp4a(v + r, b, a[i]);
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s ^= a[i];
    if (s == 0x7u) *sink = s;
}

// independent LOP3 / IADD3 chains: the alu pipe (the fma pipe's dp4a / IMAD repeat every 2 cycles per scheduler)
__global__ void __launch_bounds__(kMbThreads, 2)
mb_alu_kernel(uint32_t rounds, uint32_t seed, uint32_t* __restrict__ sink) {
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x + i;
    const uint32_t v = seed | threadIdx.x;
    for (uint32_t r = 0; r < rounds; ++r) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = __byte_perm(a[i], v + r, 0x5140);
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s ^= a[i];
    if (s == 0x7u) *sink = s;
}

// conflict-free 16-byte shared loads: 4 wavefronts per warp instruction (the l1tex data pipe)
__global__ void __launch_bounds__(kMbThreads, 2)
mb_lds_kernel(uint32_t rounds, uint32_t* __restrict__ sink) {
    __shared__ uint4 buf[2048];
    for (uint32_t i = threadIdx.x; i < 2048; i += kMbThreads) buf[i] = make_uint4(i, i, i, i);
    __syncthreads();
    uint32_t acc = 0, idx = threadIdx.x;
    for (uint32_t r = 0; r < rounds; ++r) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const uint4 v = buf[(idx + u * 512) & 2047];
            acc += v.x ^ v.w;
        }
        idx += 32;
    }
    if (acc == 0x12345u) *sink = acc;
}

// legacy warp-level integer tensor op: mma.sync m16n8k32 u8 x u8 -> s32, independent accumulator chains;
// with `prmt` > 0 that many byte permutes (alu pipe) are issued per mma (the 4x4 byte transposes of the dense phase)
template <int PRMT>
__global__ void __launch_bounds__(kMbThreads, 2)
mb_imma_kernel(uint32_t rounds, uint32_t seed, uint32_t* __restrict__ sink) {
    int32_t c[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) c[i][j] = 0;
    uint32_t a0 = seed | threadIdx.x, a1 = seed >> 1, a2 = seed >> 2, a3 = seed >> 3;
    const uint32_t b0 = seed >> 4, b1 = seed >> 5;
    for (uint32_t r = 0; r < rounds; ++r) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (PRMT) {
#pragma unroll
                for (int p = 0; p < PRMT / 4; ++p) {
                    a0 = __byte_perm(a0, a1, 0x5140); a1 = __byte_perm(a1, a2, 0x7362);
                    a2 = __byte_perm(a2, a3, 0x5410); a3 = __byte_perm(a3, a0, 0x7632);
                }
            }
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c[i][0]), "+r"(c[i][1]), "+r"(c[i][2]), "+r"(c[i][3])
                         : "r"(a0 + i), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
        }
    }
    int32_t s = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) s ^= c[i][0] ^ c[i][1] ^ c[i][2] ^ c[i][3];
    if (s == 0x7654321) *sink = (uint32_t)s;
}

}  // namespace msr

using namespace msr;

extern "C" int msr_microbench(int device, int what, uint64_t bytes, uint32_t iters, double* units_per_s, void* stream) {
    MSR_REQUIRE(units_per_s, "msr_microbench: NULL result");
    MSR_REQUIRE(iters >= 1 && iters <= 1000, "msr_microbench: iters outside [1,1000]");
    int rc = use_device(device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    int sms = 0;
    MSR_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const int grid = sms * 2;
    DevBuf sink, buf;
    MSR_CUDA_CHECK(sink.alloc(4));
    cudaEvent_t e0, e1;
    MSR_CUDA_CHECK(cudaEventCreate(&e0));
    MSR_CUDA_CHECK(cudaEventCreate(&e1));
    double units = 0;
    auto run = [&](int n) -> cudaError_t {
        for (int i = 0; i < n; ++i) {
            switch (what) {
                case MSR_MB_L2_READ: {
                    const size_t n_vec = bytes / 8;
                    const uint32_t passes = (uint32_t)std::max<uint64_t>(1, (2ull << 30) / std::max<uint64_t>(bytes, 1));
                    mb_read_kernel<<<grid, kMbThreads, 0, s>>>(buf.as<uint2>(), n_vec, passes, sink.as<uint32_t>());
                    const size_t stride = (size_t)grid * kMbThreads;
                    units = (double)(n_vec / (8 * stride)) * 8 * stride * 8.0 * passes;
                    break;
                }
                case MSR_MB_HBM_READ: {
                    const size_t n_vec = bytes / 8;
                    mb_read_kernel<<<grid, kMbThreads, 0, s>>>(buf.as<uint2>(), n_vec, 1, sink.as<uint32_t>());
                    const size_t stride = (size_t)grid * kMbThreads;
                    units = (double)(n_vec / (8 * stride)) * 8 * stride * 8.0;
                    break;
                }
                case MSR_MB_RED_SHARED: {
                    const uint32_t rounds = 4096;
                    mb_red_shared_kernel<<<grid, kMbThreads, 0, s>>>(rounds, sink.as<uint32_t>());
                    units = (double)grid * (kMbThreads / 32) * rounds * 8.0;          // warp-wide red instructions
                    break;
                }
                case MSR_MB_ISSUE: {
                    const uint32_t rounds = 8192;
                    mb_issue_kernel<<<grid, kMbThreads, 0, s>>>(rounds, 0x01010101u, sink.as<uint32_t>());
                    units = (double)grid * (kMbThreads / 32) * rounds * 8.0;          // warp-wide dp4a instructions
                    break;
                }
                case MSR_MB_ALU: {
                    const uint32_t rounds = 8192;
                    mb_alu_kernel<<<grid, kMbThreads, 0, s>>>(rounds, 0x01010101u, sink.as<uint32_t>());
                    units = (double)grid * (kMbThreads / 32) * rounds * 8.0;          // warp-wide PRMT instructions
                    break;
                }
                case MSR_MB_LDS: {
                    const uint32_t rounds = 4096;
                    mb_lds_kernel<<<grid, kMbThreads, 0, s>>>(rounds, sink.as<uint32_t>());
                    units = (double)grid * (kMbThreads / 32) * rounds * 8.0 * 4.0;    // shared-memory wavefronts
                    break;
                }
                case MSR_MB_IMMA: {
                    const uint32_t rounds = 4096;
                    mb_imma_kernel<0><<<grid, kMbThreads, 0, s>>>(rounds, 0x01020304u, sink.as<uint32_t>());
                    units = (double)grid * (kMbThreads / 32) * rounds * 4.0;          // warp-wide mma.sync instructions
                    break;
                }
                default: {
                    const uint32_t rounds = 4096;
                    mb_imma_kernel<8><<<grid, kMbThreads, 0, s>>>(rounds, 0x01020304u, sink.as<uint32_t>());
                    units = (double)grid * (kMbThreads / 32) * rounds * 4.0;          // mma.sync instructions, 8 PRMT each
                    break;
                }
            }
            count_launch();
        }
        return cudaGetLastError();
    };
    MSR_REQUIRE(what >= MSR_MB_L2_READ && what <= MSR_MB_IMMA_PRMT, "msr_microbench: unknown benchmark %d", what);
    if (what == MSR_MB_L2_READ || what == MSR_MB_HBM_READ) {
        MSR_REQUIRE(bytes >= (1u << 20) && bytes <= (16ull << 30), "msr_microbench: buffer size outside [1 MiB, 16 GiB]");
        MSR_CUDA_CHECK(buf.alloc(bytes));
        MSR_CUDA_CHECK(cudaMemsetAsync(buf.p, 1, bytes, s));
    }
    MSR_CUDA_CHECK(run(1));                                       // warm-up (also pulls an L2-sized buffer into L2)
    MSR_CUDA_CHECK(cudaEventRecord(e0, s));
    MSR_CUDA_CHECK(run((int)iters));
    MSR_CUDA_CHECK(cudaEventRecord(e1, s));
    MSR_CUDA_CHECK(cudaEventSynchronize(e1));
    float ms = 0;
    MSR_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *units_per_s = ms > 0 ? units * iters / (ms * 1e-3) : 0.0;
    return MSR_OK;
}
