// dense_tc.cuh -- kernel (c), tensor-core pass: approximate inner products of a query block against the whole
// flat index with tcgen05.mma (fp16 x fp16 -> fp32 in TMEM), operands staged by TMA.  sm_100a only.
//
// Replaces the cuBLAS HGEMM that faiss-gpu runs behind `--use_gpu` (reference src/search.py:254-270,
// GpuClonerOptions.useFloat16 -- fp16 storage is the reference's own choice) -- here as one persistent,
// warp-specialised kernel:
//
//   scores[q, d] = sum_h fp16(Q[q,h]) * fp16(X[d,h])            (fp32 accumulate)
//
// Two epilogues (template FILTER):
//   false  the 128 x 256 accumulator is written out as score rows (sample pass, tests, msr_dense_tc_scores);
//   true   FUSED SELECTION: each epilogue thread owns one query row of the tile, compares its 256 scores with the
//          query's threshold thr[q] and appends the few survivors (score bits, doc) to the query's candidate buffer
//          (one atomicAdd per 32-column chunk that has any).  The nq x N score matrix never exists in HBM.
//
//   tile      = 128 queries (UMMA M, TMEM lanes) x 256 docs (UMMA N, TMEM columns), K walked in 64-element slabs
//   warp 0    = TMA producer: one elected lane issues cp.async.bulk.tensor loads of the Q slab [128 x 64] and the
//               X slab [256 x 64] (128-byte swizzle) into a 4-stage shared-memory ring, completion on mbarriers
//   warp 1    = MMA issuer: one elected lane issues 4 x tcgen05.mma (K = 16 each) per slab; tcgen05.commit releases the
//               ring slot and, after the last slab, hands the accumulator to the epilogue; allocates / frees TMEM
//   warps 2-5 = epilogue: tcgen05.ld of the 128 x 256 fp32 accumulator (each warp its own 32-lane quarter), masked
//               128-bit stores of the score rows.  Two accumulators (2 x 256 TMEM columns) so the epilogue of tile i
//               overlaps the MMAs of tile i+1.
//   schedule  = persistent, one CTA per SM, tile id -> (doc block, query block) with the query block fastest, so
//               the CTAs running together share one X slab through L2 and X leaves HBM once per pass.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"

namespace msr {
namespace tc {

constexpr int BM = 128;            // queries per tile
constexpr int BN = 256;            // docs per tile
constexpr int BK = 64;             // fp16 elements per K slab = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int STAGES = 4;
constexpr int THREADS = 192;
constexpr uint32_t A_BYTES = BM * BK * 2;      // 16 KB
constexpr uint32_t B_BYTES = BN * BK * 2;      // 32 KB
constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
constexpr uint32_t TMEM_COLS = 512;            // 2 accumulators x 256 columns
constexpr size_t SMEM_BYTES = 1024 + (size_t)STAGES * STAGE_BYTES + 256;

// instruction descriptor (cute::UMMA::InstrDescriptor layout): fp32 accumulate (bit 4), A/B format 0 = fp16 (bits 7-9,
// 10-12), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int32_t c0, int32_t c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// K-major, 128-byte swizzle shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): rows of 128 bytes,
// 8-row swizzle atoms 1024 bytes apart (SBO), LBO unused (=1), descriptor version 1 (Blackwell), layout SWIZZLE_128B.
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(IDESC), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {    // arrives on `bar` when all MMAs issued so far are done
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct TcParams {
    uint32_t nq, n_docs, n_qblk, n_dblk, n_kslab;
    float* scores;                 // FILTER = false: [nq][ld] approximate inner products
    size_t ld;                     //                 row pitch in floats (multiple of 4)
    const float* thr;              // FILTER = true:  [nq] keep docs with score >= thr[q]
    uint32_t* cnt;                 //                 [nq] survivors appended so far (may exceed cap = overflow)
    uint2* cand;                   //                 [nq][cap] (score bits, local doc row)
    uint32_t cap;
};

template <bool FILTER>
__global__ void __launch_bounds__(THREADS, 1)
dense_tc_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_x, const TcParams P) {
    extern __shared__ unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;              // 128-byte swizzle atoms need 1024-byte alignment
    unsigned char* gen = smem_raw + (base - raw);
    const uint32_t bars = base + STAGES * STAGE_BYTES;          // full[STAGES] empty[STAGES] tfull[2] tempty[2]
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gen + STAGES * STAGE_BYTES + 8 * (2 * STAGES + 4));
    auto full = [&](uint32_t s) { return bars + 8 * s; };
    auto empty = [&](uint32_t s) { return bars + 8 * (STAGES + s); };
    auto tfull = [&](uint32_t a) { return bars + 8 * (2 * STAGES + a); };
    auto tempty = [&](uint32_t a) { return bars + 8 * (2 * STAGES + 2 + a); };

    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_q);
        tma_prefetch_desc(&map_x);
        for (uint32_t s = 0; s < STAGES; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 1); }
        for (uint32_t a = 0; a < 2; ++a) { mbar_init(tfull(a), 1); mbar_init(tempty(a), 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {                                            // one warp owns TMEM allocation and release
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32((const void*)tmem_slot)), "n"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t total = P.n_qblk * P.n_dblk;

    if (warp == 0) {
        if (lane == 0) {                                        // ===== TMA producer =====
            uint32_t stage = 0, phase = 0;
            for (uint32_t t = blockIdx.x; t < total; t += gridDim.x) {
                const int32_t q0 = (int32_t)((t % P.n_qblk) * BM), d0 = (int32_t)((t / P.n_qblk) * BN);
                for (uint32_t ks = 0; ks < P.n_kslab; ++ks) {
                    mbar_wait(empty(stage), phase ^ 1);
                    mbar_expect_tx(full(stage), STAGE_BYTES);
                    tma_load_2d(base + stage * STAGE_BYTES, &map_q, full(stage), (int32_t)(ks * BK), q0);
                    tma_load_2d(base + stage * STAGE_BYTES + A_BYTES, &map_x, full(stage), (int32_t)(ks * BK), d0);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {                                        // ===== MMA issuer =====
            uint32_t stage = 0, phase = 0, it = 0;
            for (uint32_t t = blockIdx.x; t < total; t += gridDim.x, ++it) {
                const uint32_t acc = it & 1, acc_phase = (it >> 1) & 1;
                mbar_wait(tempty(acc), acc_phase ^ 1);          // epilogue has drained this accumulator
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (uint32_t ks = 0; ks < P.n_kslab; ++ks) {
                    mbar_wait(full(stage), phase);              // TMA landed this slab
                    tcgen05_fence_after();
                    const uint64_t a_desc = umma_desc_k_sw128(base + stage * STAGE_BYTES);
                    const uint64_t b_desc = umma_desc_k_sw128(base + stage * STAGE_BYTES + A_BYTES);
#pragma unroll
                    for (uint32_t k = 0; k < BK / UMMA_K; ++k)  // +32 bytes along K inside the swizzle atom
                        umma_f16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, (ks | k) != 0 ? 1u : 0u);
                    umma_commit(empty(stage));                  // slot free once these MMAs retire
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(tfull(acc));                        // accumulator complete
            }
        }
        __syncwarp();
    } else {                                                    // ===== epilogue: TMEM -> score rows =====
        const uint32_t quarter = warp & 3;                      // a warp may only touch TMEM lanes 32*(warp%4)..+31
        const uint32_t row = quarter * 32 + lane;
        uint32_t it = 0;
        for (uint32_t t = blockIdx.x; t < total; t += gridDim.x, ++it) {
            const uint32_t acc = it & 1, acc_phase = (it >> 1) & 1;
            const uint32_t q = (t % P.n_qblk) * BM + row, d0 = (t / P.n_qblk) * BN;
            mbar_wait(tfull(acc), acc_phase);
            tcgen05_fence_after();
            if constexpr (FILTER) {
                // NaN never passes (>= is false); rows past nq and zero-filled doc rows past n_docs are masked out
                const float th = q < P.nq ? __ldg(P.thr + q) : INFINITY;
                uint2* mine = P.cand + (size_t)q * P.cap;
#pragma unroll 1
                for (uint32_t c = 0; c < BN / 32; ++c) {
                    uint32_t v[32];
                    tmem_ld32(tmem_base + ((quarter * 32u) << 16) + acc * BN + c * 32, v);
                    const uint32_t dc = d0 + c * 32;
                    uint32_t mask = 0;
#pragma unroll
                    for (uint32_t j = 0; j < 32; ++j) mask |= (__uint_as_float(v[j]) >= th ? 1u : 0u) << j;
                    if (dc + 32 > P.n_docs) mask &= dc < P.n_docs ? ((1u << (P.n_docs - dc)) - 1u) : 0u;
                    if (mask && q < P.nq) {
                        const uint32_t at = atomicAdd(P.cnt + q, (uint32_t)__popc(mask));
#pragma unroll
                        for (uint32_t j = 0; j < 32; ++j) {
                            if ((mask >> j) & 1u) {
                                const uint32_t pos = at + __popc(mask & ((1u << j) - 1u));
                                if (pos < P.cap) mine[pos] = make_uint2(v[j], dc + j);
                            }
                        }
                    }
                }
            } else {
            float* dst = P.scores + (size_t)q * P.ld + d0;
#pragma unroll 1
            for (uint32_t c = 0; c < BN / 32; ++c) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((quarter * 32u) << 16) + acc * BN + c * 32, v);
                if (q < P.nq) {
#pragma unroll
                    for (uint32_t j = 0; j < 8; ++j) {
                        const uint32_t d = d0 + c * 32 + 4 * j;
                        if (d + 3 < P.n_docs) {
                            *reinterpret_cast<float4*>(dst + c * 32 + 4 * j) =
                                make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                            __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
                        } else {
#pragma unroll
                            for (uint32_t e = 0; e < 4; ++e)
                                if (d + e < P.n_docs) dst[c * 32 + 4 * j + e] = __uint_as_float(v[4 * j + e]);
                        }
                    }
                }
            }
            }
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty(acc));
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
    }
}

// fp32 -> fp16 copy of the query block (row-major, same pitch)
__global__ void to_f16_kernel(const float* __restrict__ src, __half* __restrict__ dst, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        dst[i] = __float2half_rn(src[i]);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// fp16 [rows, dim] with a row pitch of `pitch` elements -> tensor map with a [box_rows x 64] box, 128-byte swizzle, zero
// fill out of bounds.  pitch = dim is the plain matrix; pitch = s * dim addresses every s-th row (the strided sample).
static int make_map_f16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t dim, uint64_t pitch, uint32_t box_rows) {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !p) {
            set_error("dense tensor-core pass: cuTensorMapEncodeTiled is not available from this driver");
            return MSR_ERR_CUDA;
        }
        fn = (EncodeTiledFn)p;
    }
    const cuuint64_t gdim[2] = {dim, rows};
    const cuuint64_t gstride[1] = {pitch * 2};
    const cuuint32_t box[2] = {(cuuint32_t)BK, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("dense tensor-core pass: cuTensorMapEncodeTiled failed with CUresult %d (rows=%llu dim=%llu)", (int)r,
                  (unsigned long long)rows, (unsigned long long)dim);
        return MSR_ERR_CUDA;
    }
    return MSR_OK;
}

static int tc_launch_common(const CUtensorMap& mq, const CUtensorMap& mx, TcParams& P, uint32_t nq, uint32_t n_docs,
                            uint32_t dim, bool filter, cudaStream_t s) {
    static std::atomic<uint64_t> attr_set{0};
    if (attr_needed(attr_set)) {
        MSR_CUDA_CHECK(cudaFuncSetAttribute(dense_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        MSR_CUDA_CHECK(cudaFuncSetAttribute(dense_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        attr_mark(attr_set);
    }
    P.nq = nq; P.n_docs = n_docs; P.n_qblk = (nq + BM - 1) / BM; P.n_dblk = (n_docs + BN - 1) / BN;
    P.n_kslab = (dim + BK - 1) / BK;
    int sms = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const uint64_t total = (uint64_t)P.n_qblk * P.n_dblk;
    if (total >= (1ull << 32)) { set_error("dense tensor-core pass: %llu tiles", (unsigned long long)total); return MSR_ERR_INVALID; }
    const uint32_t grid = (uint32_t)std::min<uint64_t>(total, (uint64_t)sms);
    if (filter) dense_tc_kernel<true><<<grid, THREADS, SMEM_BYTES, s>>>(mq, mx, P);
    else dense_tc_kernel<false><<<grid, THREADS, SMEM_BYTES, s>>>(mq, mx, P);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    return MSR_OK;
}

// scores[nq][ld] = fp16(Q) . fp16(X)^T on the tensor cores.  q16 [nq,dim] fp16; x16 = n_docs rows of fp16 at a pitch of
// x_pitch elements (dim = every row, s * dim = every s-th row); dim % 8 == 0.
static int dense_tc_scores_launch(const __half* q16, uint32_t nq, const __half* x16, uint32_t n_docs, uint32_t dim,
                                  uint64_t x_pitch, float* scores, size_t ld, cudaStream_t s) {
    CUtensorMap mq, mx;
    int rc = make_map_f16(&mq, q16, nq, dim, dim, BM);
    if (rc) return rc;
    rc = make_map_f16(&mx, x16, n_docs, dim, x_pitch, BN);
    if (rc) return rc;
    TcParams P{};
    P.scores = scores; P.ld = ld;
    return tc_launch_common(mq, mx, P, nq, n_docs, dim, false, s);
}

// Fused selection: survivors {d : fp16(Q[q]) . fp16(X[d]) >= thr[q]} appended to cand[q][0..cap), cnt[q] += survivors.
// cnt must be zeroed by the caller.
static int dense_tc_filter_launch(const __half* q16, uint32_t nq, const __half* x16, uint32_t n_docs, uint32_t dim,
                                  const float* thr, uint32_t* cnt, uint2* cand, uint32_t cap, cudaStream_t s) {
    CUtensorMap mq, mx;
    int rc = make_map_f16(&mq, q16, nq, dim, dim, BM);
    if (rc) return rc;
    rc = make_map_f16(&mx, x16, n_docs, dim, dim, BN);
    if (rc) return rc;
    TcParams P{};
    P.thr = thr; P.cnt = cnt; P.cand = cand; P.cap = cap;
    return tc_launch_common(mq, mx, P, nq, n_docs, dim, true, s);
}

}  // namespace tc
}  // namespace msr
