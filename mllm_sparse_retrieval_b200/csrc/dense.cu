// dense.cu -- kernel (c): dense flat inner-product top-k.
//
// Replaces tevatron FaissFlatSearcher / faiss.IndexFlatIP (reference call sites src/search.py:232,237,57,59 and the
// fp16 GPU clone :254-270) under rules D1-D6 of SURVEY.md 8(a).
//
// The index keeps two copies of the vectors in HBM: fp32 [n, dim] (exact scores) and bf16 [n, dim] (tensor-core
// pass).  Search modes:
//   1  exact: fp32 CUDA-core inner products in a fixed accumulation order -> score rows -> radix-select top-k;
//   2  tcgen05: bf16 x bf16 -> fp32 tensor-core pass (dense_tc.cuh) producing approximate score rows, top-k' of
//      those, exact fp32 re-score of the k' candidates, and a CERTIFICATE per query that no doc outside the
//      candidates can reach the exact k-th score; queries that fail it are re-run in mode 1.
//   0  auto: mode 2 when the batch is large enough to be tensor-bound, else mode 1 (HBM-bound GEMV regime).
#include <cuda_bf16.h>

#include "common.cuh"
#include "dense_tc.cuh"

namespace msr {
int topk_rows_launch(uint32_t nq, uint32_t n, const float* scores, size_t row_stride, uint32_t k, int32_t col_base,
                     float* out_scores, int32_t* out_docs, uint32_t* out_counts, cudaStream_t s);
}

struct msr_dense_index {
    int device = 0;
    uint32_t dim = 0, n_docs = 0, capacity = 0;
    int32_t doc_base = 0;
    float* v32 = nullptr;
    __nv_bfloat16* v16 = nullptr;
    float* err16 = nullptr;          // [capacity] || v32 - float(v16) ||_2 per row (certificate of mode 2)
    msr::Workspace ws;
};

namespace msr {

constexpr int kIpThreads = 256;
constexpr int kIpWarps = kIpThreads / 32;
constexpr int kIpQ = 4;              // queries held in shared memory per pass

// fp32 -> bf16 copy + per-row quantisation error norm
__global__ void dense_convert_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                     float* __restrict__ err, uint32_t n_rows, uint32_t dim) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n_rows) return;
    const float* s = src + (size_t)warp * dim;
    __nv_bfloat16* d = dst + (size_t)warp * dim;
    float e2 = 0.f;
    for (uint32_t i = lane; i < dim; i += 32) {
        float x = s[i];
        __nv_bfloat16 b = __float2bfloat16_rn(x);
        d[i] = b;
        float r = x - __bfloat162float(b);
        e2 = fmaf(r, r, e2);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e2 += __shfl_xor_sync(0xffffffffu, e2, o);
    if (lane == 0) err[warp] = sqrtf(e2) * 1.0001f;
}

// Exact pass: one warp per doc row, kIpQ queries per pass resident in shared memory.  Each lane accumulates its
// strided elements in ascending order with one fmaf chain per query, then a fixed xor-shuffle tree: the result is
// deterministic and independent of nq / batch composition.
__global__ void __launch_bounds__(kIpThreads)
dense_ip_exact_kernel(const float* __restrict__ docs, uint32_t n_docs, uint32_t dim, const float* __restrict__ queries,
                      uint32_t nq, float* __restrict__ scores, size_t row_stride) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* qs = reinterpret_cast<float*>(smem_raw);           // [kIpQ][dim]
    const uint32_t q0 = blockIdx.y * kIpQ;
    const uint32_t nqb = min((uint32_t)kIpQ, nq - q0);
    for (uint32_t i = threadIdx.x; i < kIpQ * dim; i += kIpThreads) {
        uint32_t qq = i / dim;
        qs[i] = qq < nqb ? queries[(size_t)(q0 + qq) * dim + (i - qq * dim)] : 0.f;
    }
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool vec = (dim & 3) == 0;
    for (uint32_t row = blockIdx.x * kIpWarps + warp; row < n_docs; row += gridDim.x * kIpWarps) {
        const float* d = docs + (size_t)row * dim;
        float acc[kIpQ];
#pragma unroll
        for (int j = 0; j < kIpQ; ++j) acc[j] = 0.f;
        if (vec) {
            const float4* d4 = reinterpret_cast<const float4*>(d);
            for (uint32_t i = lane; i < dim / 4; i += 32) {
                float4 x = __ldg(d4 + i);
#pragma unroll
                for (int j = 0; j < kIpQ; ++j) {
                    float4 y = reinterpret_cast<const float4*>(qs + (size_t)j * dim)[i];
                    acc[j] = fmaf(x.x, y.x, acc[j]); acc[j] = fmaf(x.y, y.y, acc[j]);
                    acc[j] = fmaf(x.z, y.z, acc[j]); acc[j] = fmaf(x.w, y.w, acc[j]);
                }
            }
        } else {
            for (uint32_t i = lane; i < dim; i += 32) {
                float x = __ldg(d + i);
#pragma unroll
                for (int j = 0; j < kIpQ; ++j) acc[j] = fmaf(x, qs[(size_t)j * dim + i], acc[j]);
            }
        }
#pragma unroll
        for (int j = 0; j < kIpQ; ++j) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
        }
        if (lane < nqb) {
            float v = acc[0];
#pragma unroll
            for (int j = 1; j < kIpQ; ++j) if ((int)lane == j) v = acc[j];
            scores[(size_t)(q0 + lane) * row_stride + row] = v;
        }
    }
}

// staging rows [nq,k_in] (int32 docs, counts) -> FAISS convention [nq,k_out]: int64 ids, -1 / -FLT_MAX padding (D5)
__global__ void dense_finish_kernel(const float* __restrict__ s_in, const int32_t* __restrict__ d_in,
                                    const uint32_t* __restrict__ cnt, uint32_t nq, uint32_t k_in, uint32_t k_out,
                                    float* __restrict__ s_out, int64_t* __restrict__ id_out) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= (size_t)nq * k_out) return;
    uint32_t q = (uint32_t)(i / k_out), r = (uint32_t)(i - (size_t)q * k_out);
    bool ok = r < min(cnt[q], k_in);
    s_out[i] = ok ? s_in[(size_t)q * k_in + r] : -3.4028234663852886e38f;
    id_out[i] = ok ? (int64_t)d_in[(size_t)q * k_in + r] : -1;
}

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace msr

using namespace msr;

extern "C" int msr_dense_index_create(int device, uint32_t dim, int32_t doc_base, msr_dense_index** out) {
    MSR_REQUIRE(out, "msr_dense_index_create: out is NULL");
    *out = nullptr;
    MSR_REQUIRE(dim >= 1 && dim <= 8192, "msr_dense_index_create: dim %u outside [1,8192]", dim);
    int rc = use_device(device);
    if (rc) return rc;
    msr_dense_index* h = new (std::nothrow) msr_dense_index();
    if (!h) { set_error("msr_dense_index_create: out of host memory"); return MSR_ERR_NOMEM; }
    h->device = device; h->dim = dim; h->doc_base = doc_base;
    *out = h;
    return MSR_OK;
}

extern "C" int msr_dense_index_destroy(msr_dense_index* h) {
    if (!h) return MSR_OK;
    cudaSetDevice(h->device);
    cudaFree(h->v32); cudaFree(h->v16); cudaFree(h->err16);
    h->ws.release();
    delete h;
    return MSR_OK;
}

extern "C" int msr_dense_index_info(const msr_dense_index* h, uint32_t info[5]) {
    MSR_REQUIRE(h && info, "msr_dense_index_info: NULL argument");
    uint64_t bytes = (uint64_t)h->capacity * h->dim * 6 + (uint64_t)h->capacity * 4;
    info[0] = h->n_docs; info[1] = h->dim; info[2] = (uint32_t)h->device;
    info[3] = (uint32_t)bytes; info[4] = (uint32_t)(bytes >> 32);
    return MSR_OK;
}

extern "C" int msr_dense_index_add(msr_dense_index* h, uint32_t n_rows, const float* vecs, int where, void* stream) {
    MSR_REQUIRE(h, "msr_dense_index_add: NULL handle");
    MSR_REQUIRE(where == MSR_HOST || where == MSR_DEVICE, "msr_dense_index_add: bad `where`");
    if (n_rows == 0) return MSR_OK;
    MSR_REQUIRE(vecs, "msr_dense_index_add: vecs is NULL");
    MSR_REQUIRE((uint64_t)h->n_docs + n_rows < (1ull << 31), "msr_dense_index_add: more than 2^31 rows");
    int rc = use_device(h->device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t need = h->n_docs + n_rows;
    if (need > h->capacity) {
        uint32_t cap = h->capacity ? (uint32_t)std::min<uint64_t>((uint64_t)h->capacity * 3 / 2, 0x7FFFFFFFull) : need;
        cap = std::max(cap, need);
        float* n32 = nullptr; __nv_bfloat16* n16 = nullptr; float* nerr = nullptr;
        MSR_CUDA_CHECK(cudaMalloc(&n32, (size_t)cap * h->dim * 4));
        MSR_CUDA_CHECK(cudaMalloc(&n16, (size_t)cap * h->dim * 2));
        MSR_CUDA_CHECK(cudaMalloc(&nerr, (size_t)cap * 4));
        if (h->n_docs) {
            MSR_CUDA_CHECK(cudaMemcpyAsync(n32, h->v32, (size_t)h->n_docs * h->dim * 4, cudaMemcpyDeviceToDevice, s));
            MSR_CUDA_CHECK(cudaMemcpyAsync(n16, h->v16, (size_t)h->n_docs * h->dim * 2, cudaMemcpyDeviceToDevice, s));
            MSR_CUDA_CHECK(cudaMemcpyAsync(nerr, h->err16, (size_t)h->n_docs * 4, cudaMemcpyDeviceToDevice, s));
            MSR_CUDA_CHECK(cudaStreamSynchronize(s));
        }
        cudaFree(h->v32); cudaFree(h->v16); cudaFree(h->err16);
        h->v32 = n32; h->v16 = n16; h->err16 = nerr; h->capacity = cap;
    }
    float* dst = h->v32 + (size_t)h->n_docs * h->dim;
    MSR_CUDA_CHECK(cudaMemcpyAsync(dst, vecs, (size_t)n_rows * h->dim * 4,
                                   where == MSR_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, s));
    const uint32_t blocks = (n_rows + 7) / 8;
    dense_convert_kernel<<<blocks, 256, 0, s>>>(dst, h->v16 + (size_t)h->n_docs * h->dim, h->err16 + h->n_docs, n_rows, h->dim);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    if (where == MSR_HOST) MSR_CUDA_CHECK(cudaStreamSynchronize(s));
    h->n_docs = need;
    return MSR_OK;
}

namespace msr {

// Exact pass over queries that are already on the device; writes FAISS-convention outputs [nq,k].
// scratch holds score rows [qc, n_docs] + top-k staging [qc,kk] (f32, i32) + counts [qc].
static int dense_search_exact(msr_dense_index* h, uint32_t nq, const float* d_q, uint32_t k, float* d_os, int64_t* d_oi,
                              unsigned char* scratch, size_t scratch_bytes, cudaStream_t s) {
    const uint32_t kk = std::min(k, h->n_docs);
    const size_t row_stride = align_up(h->n_docs, 4);
    const size_t per_q = row_stride * 4 + (size_t)kk * 8 + 4;
    uint32_t qc = (uint32_t)std::min<size_t>(nq, std::max<size_t>(1, (scratch_bytes - 1024) / per_q));
    if (qc >= (uint32_t)kIpQ) qc = qc / kIpQ * kIpQ;
    float* rows = (float*)scratch;
    float* ts = (float*)(scratch + align_up((size_t)qc * row_stride * 4, 256));
    int32_t* td = (int32_t*)((unsigned char*)ts + align_up((size_t)qc * kk * 4, 256));
    uint32_t* tc = (uint32_t*)((unsigned char*)td + align_up((size_t)qc * kk * 4, 256));
    const size_t smem = (size_t)kIpQ * h->dim * 4;
    static bool attr_set = false;
    if (!attr_set) {
        MSR_CUDA_CHECK(cudaFuncSetAttribute(dense_ip_exact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kIpQ * 8192 * 4));
        attr_set = true;
    }
    for (uint32_t q0 = 0; q0 < nq; q0 += qc) {
        const uint32_t n = std::min(qc, nq - q0);
        dim3 grid(std::min<uint32_t>((h->n_docs + kIpWarps - 1) / kIpWarps, 148 * 8), (n + kIpQ - 1) / kIpQ);
        dense_ip_exact_kernel<<<grid, kIpThreads, smem, s>>>(h->v32, h->n_docs, h->dim, d_q + (size_t)q0 * h->dim, n, rows, row_stride);
        count_launch();
        MSR_CUDA_CHECK(cudaGetLastError());
        int rc = topk_rows_launch(n, h->n_docs, rows, row_stride, kk, h->doc_base, ts, td, tc, s);
        if (rc) return rc;
        dense_finish_kernel<<<(unsigned)(((size_t)n * k + 255) / 256), 256, 0, s>>>(ts, td, tc, n, kk, k, d_os + (size_t)q0 * k,
                                                                                   d_oi + (size_t)q0 * k);
        count_launch();
        MSR_CUDA_CHECK(cudaGetLastError());
    }
    return MSR_OK;
}

static size_t dense_scratch_bytes(const msr_dense_index* h, uint32_t nq, uint32_t k) {
    const uint32_t kk = std::min(k, h->n_docs);
    const size_t per_q = align_up(h->n_docs, 4) * 4 + (size_t)kk * 8 + 4;
    const size_t want = per_q * nq + 4096;
    const size_t floor_b = per_q * std::min<uint32_t>(nq, kIpQ) + 4096;
    return std::max(floor_b, std::min<size_t>(want, 2ull << 30));
}

}  // namespace msr

extern "C" int msr_dense_search(msr_dense_index* h, uint32_t nq, const float* queries, uint32_t k, float* out_scores,
                                int64_t* out_ids, int mode, int where, void* stream) {
    MSR_REQUIRE(h, "msr_dense_search: NULL handle");
    MSR_REQUIRE(where == MSR_HOST || where == MSR_DEVICE, "msr_dense_search: bad `where`");
    MSR_REQUIRE(k >= 1 && k <= MSR_MAX_K, "msr_dense_search: k=%u outside [1,%u]", k, MSR_MAX_K);
    MSR_REQUIRE(mode >= 0 && mode <= 2, "msr_dense_search: mode %d outside {0,1,2}", mode);
    MSR_REQUIRE(h->n_docs > 0, "msr_dense_search: the index is empty");
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(queries && out_scores && out_ids, "msr_dense_search: NULL buffer");
    int rc = use_device(h->device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t sb = align_up(dense_scratch_bytes(h, nq, k), 256);
    size_t off_q = sb, off_os = 0, off_oi = 0, total = sb;
    if (where == MSR_HOST) {
        off_os = align_up(off_q + (size_t)nq * h->dim * 4, 256);
        off_oi = align_up(off_os + (size_t)nq * k * 4, 256);
        total = align_up(off_oi + (size_t)nq * k * 8, 256);
    }
    MSR_CUDA_CHECK(h->ws.reserve(total, s));
    unsigned char* base = (unsigned char*)h->ws.p;
    const float* d_q = queries;
    float* d_os = out_scores;
    int64_t* d_oi = out_ids;
    if (where == MSR_HOST) {
        MSR_CUDA_CHECK(cudaMemcpyAsync(base + off_q, queries, (size_t)nq * h->dim * 4, cudaMemcpyHostToDevice, s));
        d_q = (const float*)(base + off_q); d_os = (float*)(base + off_os); d_oi = (int64_t*)(base + off_oi);
    }
    rc = dense_search_exact(h, nq, d_q, k, d_os, d_oi, base, sb, s);
    if (rc) return rc;
    if (where == MSR_HOST) {
        MSR_CUDA_CHECK(cudaMemcpyAsync(out_scores, d_os, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, s));
        MSR_CUDA_CHECK(cudaMemcpyAsync(out_ids, d_oi, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, s));
        MSR_CUDA_CHECK(cudaStreamSynchronize(s));
    }
    return MSR_OK;
}


// ---------------------------------------------------------------------------------------------------------
// tensor-core pass on its own (tests, bench.py --dense, ncu): scores[nq, n_docs] = bf16(Q) . bf16(X)^T, fp32 out
// ---------------------------------------------------------------------------------------------------------
extern "C" int msr_dense_tc_scores(msr_dense_index* h, uint32_t nq, const float* queries, float* out_scores, size_t ld,
                                   void* stream) {
    MSR_REQUIRE(h && queries && out_scores, "msr_dense_tc_scores: NULL argument");
    MSR_REQUIRE(h->n_docs > 0 && nq > 0, "msr_dense_tc_scores: empty index or query block");
    MSR_REQUIRE(h->dim % 8 == 0, "msr_dense_tc_scores: dim %u is not a multiple of 8 (TMA row pitch)", h->dim);
    MSR_REQUIRE(ld >= h->n_docs && ld % 4 == 0, "msr_dense_tc_scores: ld must be >= n_docs and a multiple of 4");
    int rc = use_device(h->device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    MSR_CUDA_CHECK(h->ws.reserve(align_up((size_t)nq * h->dim * 2, 256), s));
    __nv_bfloat16* q16 = (__nv_bfloat16*)h->ws.p;
    const size_t n = (size_t)nq * h->dim;
    tc::to_bf16_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 148 * 16), 256, 0, s>>>(queries, q16, n);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    return tc::dense_tc_scores_launch(q16, nq, h->v16, h->n_docs, h->dim, out_scores, ld, s);
}
