// dense.cu -- kernel (c): dense flat inner-product top-k.
//
// Replaces tevatron FaissFlatSearcher / faiss.IndexFlatIP (reference call sites src/search.py:232,237,57,59 and the
// fp16 GPU clone :254-270) under rules D1-D6 of SURVEY.md 8(a).
//
// The index keeps two copies of the vectors in HBM: fp32 [n, dim] (exact scores) and fp16 [n, dim] (tensor-core
// pass; fp16 is what the reference's own GPU path stores, GpuClonerOptions.useFloat16).  Search modes:
//   1  exact: fp32 CUDA-core inner products in a fixed accumulation order -> score rows -> radix-select top-k;
//   2  tcgen05: fp16 x fp16 -> fp32 tensor-core pass (dense_tc.cuh) whose epilogue keeps only the docs above a
//      per-query threshold (selection fused into the GEMM, no score matrix in HBM), exact fp32 re-score of a
//      candidate set that is CERTIFIED to hold the exact top-k; queries that fail the certificate are re-run in mode 1.
//   0  auto: mode 2 when the batch is large enough to be tensor-bound, else mode 1 (HBM-bound GEMV regime).
#include <cuda_fp16.h>

#include <cmath>

#include <vector>

#include "common.cuh"
#include "dense_tc.cuh"

namespace msr {
int topk_rows_launch(uint32_t nq, uint32_t n, const float* scores, size_t row_stride, uint32_t k, int32_t col_base,
                     float* out_scores, int32_t* out_docs, uint32_t* out_counts, cudaStream_t s, const uint32_t* nq_dev = nullptr,
                     uint32_t nq_off = 0);
}

struct msr_dense_index {
    int device = 0;
    uint32_t dim = 0, n_docs = 0, capacity = 0;
    int32_t doc_base = 0;
    float* v32 = nullptr;
    __half* v16 = nullptr;          // fp16 copy for the tensor-core pass
    float* err16 = nullptr;          // [capacity] || v32 - float(v16) ||_2 per row (certificate of mode 2)
    float* stats = nullptr;          // [2] max row norm, max row fp16 error over the whole index (certificate of mode 2)
    msr::Workspace ws;
    uint32_t last_fallback = 0;      // queries of the last mode-2 search that needed the exact pass (host-side view)
    uint32_t* d_fail = nullptr;      // device: [0] uncertified queries of the last search, [1] != 0: more than one chunk's
                                     // kDenseFallbackMax of them (device path: the rest kept their uncertified rows)
    // optional per-kernel timing (msr_dense_profile): events bracket the fused GEMM pass and the re-score of every block
    bool profiling = false;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    double prof_tc_ms = 0, prof_rescore_ms = 0;
    uint64_t prof_launches = 0;
};

namespace msr {

constexpr int kIpThreads = 256;
constexpr int kIpWarps = kIpThreads / 32;
constexpr int kIpQ = 4;              // queries held in shared memory per pass
constexpr uint32_t kDenseFallbackMax = 128;  // uncertified queries per query block that the device re-runs on its own

// fp32 -> fp16 copy + per-row quantisation error norm (overflow to inf makes the error inf: mode 2 then certifies nothing)
__global__ void dense_convert_kernel(const float* __restrict__ src, __half* __restrict__ dst,
                                     float* __restrict__ err, float* __restrict__ stats, uint32_t n_rows, uint32_t dim) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n_rows) return;
    const float* s = src + (size_t)warp * dim;
    __half* d = dst + (size_t)warp * dim;
    float e2 = 0.f, n2 = 0.f;
    for (uint32_t i = lane; i < dim; i += 32) {
        float x = s[i];
        __half b = __float2half_rn(x);
        d[i] = b;
        float r = x - __half2float(b);
        e2 = fmaf(r, r, e2);
        n2 = fmaf(x, x, n2);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        e2 += __shfl_xor_sync(0xffffffffu, e2, o);
        n2 += __shfl_xor_sync(0xffffffffu, n2, o);
    }
    if (lane == 0) {
        const float e = sqrtf(e2) * 1.0001f, nrm = sqrtf(n2) * 1.0001f;
        err[warp] = e;
        // non-negative floats order like their bit patterns
        atomicMax(reinterpret_cast<int*>(stats), __float_as_int(nrm));
        atomicMax(reinterpret_cast<int*>(stats) + 1, __float_as_int(e));
    }
}

// Exact pass: one warp per doc row, NQ (1, 2 or kIpQ) queries per pass resident in shared memory.  Each lane accumulates its
// strided elements in ascending order with one fmaf chain per query, then a fixed xor-shuffle tree: the result is
// deterministic and independent of nq / batch composition / NQ.  A lane issues kIpLoads independent 16-byte loads of the
// row before it uses the first (the pass is an HBM stream: 410 MB at the reference's own batch sizes of 1-4 queries).
template <int NQ>
__global__ void __launch_bounds__(kIpThreads)
dense_ip_exact_kernel(const float* __restrict__ docs, uint32_t n_docs, uint32_t dim, const float* __restrict__ queries,
                      uint32_t nq, float* __restrict__ scores, size_t row_stride, const uint32_t* __restrict__ nq_dev,
                      uint32_t nq_off) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* qs = reinterpret_cast<float*>(smem_raw);           // [NQ][dim]
    if (nq_dev) {                                             // the count lives on the device (mode-2 fallback list)
        const uint32_t have = *nq_dev;
        nq = min(nq, have > nq_off ? have - nq_off : 0u);
    }
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool vec = (dim & 3) == 0;
    constexpr int kIpLoads = NQ > 2 ? 4 : 8;
    // groups of NQ queries: one per blockIdx.y, or (device-counted launches: gridDim.y == 1) all of them in turn, so that an
    // empty list costs one wave of no-op CTAs and a listed query still gets the whole chip's bandwidth
    for (uint32_t q0 = blockIdx.y * NQ; q0 < nq; q0 += gridDim.y * NQ) {
    const uint32_t nqb = min((uint32_t)NQ, nq - q0);
    __syncthreads();                                          // the previous group's queries are no longer read
    for (uint32_t i = threadIdx.x; i < NQ * dim; i += kIpThreads) {
        uint32_t qq = i / dim;
        qs[i] = qq < nqb ? queries[(size_t)(q0 + qq) * dim + (i - qq * dim)] : 0.f;
    }
    __syncthreads();
    for (uint32_t row = blockIdx.x * kIpWarps + warp; row < n_docs; row += gridDim.x * kIpWarps) {
        const float* d = docs + (size_t)row * dim;
        float acc[NQ];
#pragma unroll
        for (int j = 0; j < NQ; ++j) acc[j] = 0.f;
        if (vec) {
            const float4* d4 = reinterpret_cast<const float4*>(d);
            const uint32_t d4n = dim / 4;
            for (uint32_t i0 = lane; i0 < d4n; i0 += 32 * kIpLoads) {
                float4 x[kIpLoads];
#pragma unroll
                for (int u = 0; u < kIpLoads; ++u) {
                    const uint32_t i = i0 + 32 * u;
                    x[u] = i < d4n ? __ldg(d4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < kIpLoads; ++u) {
                    const uint32_t i = i0 + 32 * u;
                    if (i < d4n) {
#pragma unroll
                        for (int j = 0; j < NQ; ++j) {
                            const float4 y = reinterpret_cast<const float4*>(qs + (size_t)j * dim)[i];
                            acc[j] = fmaf(x[u].x, y.x, acc[j]); acc[j] = fmaf(x[u].y, y.y, acc[j]);
                            acc[j] = fmaf(x[u].z, y.z, acc[j]); acc[j] = fmaf(x[u].w, y.w, acc[j]);
                        }
                    }
                }
            }
        } else {
            for (uint32_t i = lane; i < dim; i += 32) {
                float x = __ldg(d + i);
#pragma unroll
                for (int j = 0; j < NQ; ++j) acc[j] = fmaf(x, qs[(size_t)j * dim + i], acc[j]);
            }
        }
#pragma unroll
        for (int j = 0; j < NQ; ++j) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
        }
        if (lane < nqb) {
            float v = acc[0];
#pragma unroll
            for (int j = 1; j < NQ; ++j) if ((int)lane == j) v = acc[j];
            scores[(size_t)(q0 + lane) * row_stride + row] = v;
        }
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
    if (cudaMalloc(&h->d_fail, 8) != cudaSuccess || cudaMemset(h->d_fail, 0, 8) != cudaSuccess) {
        set_error("msr_dense_index_create: device allocation failed");
        delete h;
        return MSR_ERR_CUDA;
    }
    *out = h;
    return MSR_OK;
}

extern "C" int msr_dense_index_destroy(msr_dense_index* h) {
    if (!h) return MSR_OK;
    cudaSetDevice(h->device);
    cudaFree(h->v32); cudaFree(h->v16); cudaFree(h->err16); cudaFree(h->stats); cudaFree(h->d_fail);
    for (cudaEvent_t e : h->ev) if (e) cudaEventDestroy(e);
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
        float* n32 = nullptr; __half* n16 = nullptr; float* nerr = nullptr;
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
    if (!h->stats) {
        MSR_CUDA_CHECK(cudaMalloc(&h->stats, 8));
        MSR_CUDA_CHECK(cudaMemsetAsync(h->stats, 0, 8, s));
    }
    float* dst = h->v32 + (size_t)h->n_docs * h->dim;
    MSR_CUDA_CHECK(cudaMemcpyAsync(dst, vecs, (size_t)n_rows * h->dim * 4,
                                   where == MSR_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, s));
    const uint32_t blocks = (n_rows + 7) / 8;
    dense_convert_kernel<<<blocks, 256, 0, s>>>(dst, h->v16 + (size_t)h->n_docs * h->dim, h->err16 + h->n_docs, h->stats, n_rows, h->dim);
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
                              unsigned char* scratch, size_t scratch_bytes, cudaStream_t s, const uint32_t* nq_dev = nullptr) {
    const uint32_t kk = std::min(k, h->n_docs);
    const size_t row_stride = align_up(h->n_docs, 4);
    const size_t per_q = row_stride * 4 + (size_t)kk * 8 + 4;
    uint32_t qc = (uint32_t)std::min<size_t>(nq, std::max<size_t>(1, (scratch_bytes - 1024) / per_q));
    if (qc >= (uint32_t)kIpQ) qc = qc / kIpQ * kIpQ;
    float* rows = (float*)scratch;
    float* ts = (float*)(scratch + align_up((size_t)qc * row_stride * 4, 256));
    int32_t* td = (int32_t*)((unsigned char*)ts + align_up((size_t)qc * kk * 4, 256));
    uint32_t* tc = (uint32_t*)((unsigned char*)td + align_up((size_t)qc * kk * 4, 256));
    // queries per pass of the kernel: 1 or 2 for the reference's own batch sizes (less shared memory, more CTAs per SM, no
    // multiplications by zero rows), else kIpQ
    const uint32_t gq = nq_dev ? (uint32_t)kIpQ : (nq == 1 ? 1u : nq == 2 ? 2u : (uint32_t)kIpQ);
    const size_t smem = (size_t)gq * h->dim * 4;
    static std::atomic<uint64_t> attr_set{0};
    if (attr_needed(attr_set)) {
        MSR_CUDA_CHECK(cudaFuncSetAttribute(dense_ip_exact_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 1 * 8192 * 4));
        MSR_CUDA_CHECK(cudaFuncSetAttribute(dense_ip_exact_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 8192 * 4));
        MSR_CUDA_CHECK(cudaFuncSetAttribute(dense_ip_exact_kernel<kIpQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, kIpQ * 8192 * 4));
        attr_mark(attr_set);
    }
    for (uint32_t q0 = 0; q0 < nq; q0 += qc) {
        const uint32_t n = std::min(qc, nq - q0);
        // device-counted launches (mode-2 fallback, usually empty): ONE grid row that walks the listed query groups itself --
        // an empty list is one wave of no-op CTAs, a listed query is answered at the full width of the chip
        dim3 grid(std::min<uint32_t>((h->n_docs + kIpWarps - 1) / kIpWarps, 148 * 8), nq_dev ? 1 : (n + gq - 1) / gq);
        const float* qp = d_q + (size_t)q0 * h->dim;
        if (gq == 1) dense_ip_exact_kernel<1><<<grid, kIpThreads, smem, s>>>(h->v32, h->n_docs, h->dim, qp, n, rows, row_stride, nq_dev, q0);
        else if (gq == 2) dense_ip_exact_kernel<2><<<grid, kIpThreads, smem, s>>>(h->v32, h->n_docs, h->dim, qp, n, rows, row_stride, nq_dev, q0);
        else dense_ip_exact_kernel<kIpQ><<<grid, kIpThreads, smem, s>>>(h->v32, h->n_docs, h->dim, qp, n, rows, row_stride, nq_dev, q0);
        count_launch();
        MSR_CUDA_CHECK(cudaGetLastError());
        int rc = topk_rows_launch(n, h->n_docs, rows, row_stride, kk, h->doc_base, ts, td, tc, s, nq_dev, q0);
        if (rc) return rc;
        dense_finish_kernel<<<(unsigned)(((size_t)n * k + 255) / 256), 256, 0, s>>>(ts, td, tc, n, kk, k, d_os + (size_t)q0 * k,
                                                                                   d_oi + (size_t)q0 * k);
        count_launch();
        MSR_CUDA_CHECK(cudaGetLastError());
    }
    return MSR_OK;
}

// ---- mode 2: tensor-core pass with fused selection -> exact fp32 re-score of a certified candidate set ----
//
// Notation: a(d) = fp16(q).fp16(x_d) accumulated in fp32 by tcgen05 (approximate), e(d) = the fp32 score of mode 1 (exact,
// fixed order).  |e(d) - a(d)| <= E_q for every doc, with
//     E_q = ||q - q^||*max||x|| + ||q^||*max||x - x^|| + (fp32 accumulation slack of both passes)        (Cauchy-Schwarz).
// Let a_k be the k-th largest a(d).  Every doc outside C = {d : a(d) >= a_k - 2 E_q} has e(d) < a_k - E_q, while the k docs
// with a(d) >= a_k have e(d) >= a_k - E_q: the exact top-k is a subset of C.  So:
//   P1  theta_q: an ESTIMATE of a_k - 2 E_q from a strided sample of m docs (rank r of the sample, r chosen so that the
//       sample holds r of the true top-k with probability < 1e-4) -- full-score GEMM on m docs + radix select;
//   P2  GEMM over all docs with the fused filter epilogue: survivors S = {a(d) >= theta_q} land in per-query buffers;
//   P3  per query: sort S, read a_k, keep C, and CERTIFY: |S| >= k, no buffer overflow and theta_q <= a_k - 2 E_q
//       (then S held every member of C).  The certificate is unconditional; only its success rate is statistical.
//   P4  exact re-score of C (same accumulation order as mode 1), final order (score desc, id asc), FAISS-convention rows.
// Uncertified queries are re-run by mode 1, so both modes return identical ids and bit-identical scores.

__device__ __forceinline__ float cert_E(float qn, float qe, const float* __restrict__ stats, uint32_t dim) {
    const float xn = stats[0], ex = stats[1], qh = qn + qe;
    // accumulation slack of BOTH passes: fp32 round-to-nearest on the exact side (dim * 2^-24), and the tensor pass, whose
    // fp32 accumulation is not documented as round-to-nearest (truncation: up to 2^-23 per add) -- 3 * dim * 2^-24 in all
    const float slack = ((float)dim * 1.7881393e-7f + 9.5367432e-7f) * qh * (xn + ex);
    return (qe * xn + qh * ex + slack) * 1.0001f;
}

// per query: ||q||_2 and ||q - fp16(q)||_2 (both rounded up a little)
__global__ void dense_query_norms_kernel(const float* __restrict__ q, uint32_t nq, uint32_t dim, float* __restrict__ qn,
                                         float* __restrict__ qe) {
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= nq) return;
    const float* s = q + (size_t)w * dim;
    float n2 = 0.f, e2 = 0.f;
    for (uint32_t i = lane; i < dim; i += 32) {
        const float x = s[i], r = x - __half2float(__float2half_rn(x));
        n2 = fmaf(x, x, n2);
        e2 = fmaf(r, r, e2);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        n2 += __shfl_xor_sync(0xffffffffu, n2, o);
        e2 += __shfl_xor_sync(0xffffffffu, e2, o);
    }
    if (lane == 0) { qn[w] = sqrtf(n2) * 1.0001f; qe[w] = sqrtf(e2) * 1.0001f; }
}

// P1 tail: theta_q = (r-th best sampled score) - 2 E_q, or -inf when the sample has fewer than r finite scores
__global__ void dense_thr_kernel(const float* __restrict__ top, const uint32_t* __restrict__ top_cnt, uint32_t r,
                                 const float* __restrict__ qn, const float* __restrict__ qe, const float* __restrict__ stats,
                                 uint32_t nq, uint32_t dim, float* __restrict__ thr) {
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    float t = -INFINITY;
    if (r > 0 && top_cnt[q] >= r) t = top[(size_t)q * r + r - 1] - 2.f * cert_E(qn[q], qe[q], stats, dim);
    thr[q] = (t == t) ? t : -INFINITY;
}
__global__ void fill_f32_kernel(float* __restrict__ p, float v, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

constexpr int kSelThreads = 256;

// P3: one CTA per query.  Sorts the survivors by (a desc, doc asc), cuts the candidate set C and certifies it.
__global__ void __launch_bounds__(kSelThreads)
dense_select_kernel(const uint2* __restrict__ cand, const uint32_t* __restrict__ cnt, uint32_t cap, const float* __restrict__ thr,
                    const float* __restrict__ qn, const float* __restrict__ qe, const float* __restrict__ stats, uint32_t kk,
                    uint32_t n_docs, uint32_t dim, int32_t doc_base, int32_t* __restrict__ out_docs,
                    uint32_t* __restrict__ out_cnt, uint32_t* __restrict__ flags) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* buf = reinterpret_cast<uint64_t*>(smem_raw);
    __shared__ uint32_t s_nc;
    const uint32_t q = blockIdx.x, tid = threadIdx.x;
    const uint32_t n_raw = cnt[q], n = min(n_raw, cap);
    uint32_t n2 = 2;
    while (n2 < n) n2 <<= 1;
    const uint2* mine = cand + (size_t)q * cap;
    for (uint32_t i = tid; i < n2; i += kSelThreads) {
        uint64_t key = 0;
        if (i < n) { const uint2 c = mine[i]; key = make_key(__uint_as_float(c.x), c.y); }
        buf[i] = key;
    }
    if (tid == 0) s_nc = 0;
    bitonic_sort_desc(buf, n2);
    bool ok = false;
    uint32_t nc = 0;
    if (n_raw <= cap) {
        if (n_raw >= n_docs) {                                  // every doc survived: C = all docs
            ok = true; nc = n;
        } else if (n >= kk) {
            const float a_k = key_score(buf[kk - 1]);
            const float E = cert_E(qn[q], qe[q], stats, dim);
            const float lim = a_k - 2.f * E;
            ok = isfinite(E) && isfinite(a_k) && thr[q] <= lim;  // NaN / inf anywhere (fp16 overflow of q or x) -> not certified
            if (ok) {
                uint32_t c = 0;
                for (uint32_t i = tid; i < n; i += kSelThreads) c += key_score(buf[i]) >= lim ? 1u : 0u;
                c = warp_reduce_add(c);
                if ((tid & 31) == 0 && c) atomicAdd(&s_nc, c);
                __syncthreads();
                nc = s_nc;                                      // the list is sorted: C is its first nc entries
            }
        }
    }
    for (uint32_t i = tid; i < nc; i += kSelThreads) out_docs[(size_t)q * cap + i] = doc_base + (int32_t)key_doc(buf[i]);
    if (tid == 0) { out_cnt[q] = nc; flags[q] = ok ? 1u : 0u; }
}

constexpr int kRsThreads = 256;
constexpr int kRsWarps = kRsThreads / 32;
constexpr int kRsPer = 4;            // candidates a warp scores together (independent fmaf chains)

// P4a: exact scores of the candidates of one query (CTA = query, query row in shared memory).  Each candidate is scored
// by one warp in the SAME accumulation order as dense_ip_exact_kernel (lane-strided fmaf chain over float4 elements,
// xor-shuffle tree), so both modes return bit-identical scores.  dim % 4 == 0.
__global__ void __launch_bounds__(kRsThreads)
dense_rescore_kernel(const float* __restrict__ docs, uint32_t dim, int32_t doc_base, const float* __restrict__ queries,
                     const int32_t* __restrict__ cand, const uint32_t* __restrict__ cnt, uint32_t cap, float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* qs = reinterpret_cast<float4*>(smem_raw);
    const uint32_t q = blockIdx.x, n = min(cnt[q], cap), d4n = dim / 4;
    if (n == 0) return;
    const float4* y4 = reinterpret_cast<const float4*>(queries + (size_t)q * dim);
    for (uint32_t i = threadIdx.x; i < d4n; i += kRsThreads) qs[i] = __ldg(y4 + i);
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int32_t* mine = cand + (size_t)q * cap;
    // gridDim.y CTAs share a query's candidates (small batches: one CTA per query left most SMs idle -- 16 queries at depth
    // 1000 took 1.0 ms here); every candidate is still scored by one warp in the same order
    for (uint32_t r0 = (blockIdx.y * kRsWarps + warp) * kRsPer; r0 < n; r0 += gridDim.y * kRsWarps * kRsPer) {
        const float4* d4[kRsPer];
        float acc[kRsPer];
#pragma unroll
        for (int j = 0; j < kRsPer; ++j) {
            const uint32_t r = min(r0 + j, n - 1);              // tail: recompute the last candidate, never stored
            d4[j] = reinterpret_cast<const float4*>(docs + (size_t)(mine[r] - doc_base) * dim);
            acc[j] = 0.f;
        }
#pragma unroll 2
        for (uint32_t i = lane; i < d4n; i += 32) {
            const float4 b = qs[i];
            float4 a[kRsPer];
#pragma unroll
            for (int j = 0; j < kRsPer; ++j) a[j] = __ldg(d4[j] + i);
#pragma unroll
            for (int j = 0; j < kRsPer; ++j) {
                acc[j] = fmaf(a[j].x, b.x, acc[j]); acc[j] = fmaf(a[j].y, b.y, acc[j]);
                acc[j] = fmaf(a[j].z, b.z, acc[j]); acc[j] = fmaf(a[j].w, b.w, acc[j]);
            }
        }
#pragma unroll
        for (int j = 0; j < kRsPer; ++j) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
            if (lane == 0 && r0 + j < n) out[(size_t)q * cap + r0 + j] = acc[j];
        }
    }
}

// P4b: final order by (exact score desc, id asc) and FAISS-convention output rows (int64 ids, -1 / -FLT_MAX padding, D5)
__global__ void __launch_bounds__(kSelThreads)
dense_final_kernel(const float* __restrict__ es, const int32_t* __restrict__ cand, const uint32_t* __restrict__ cnt, uint32_t cap,
                   uint32_t k, float* __restrict__ s_out, int64_t* __restrict__ id_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* buf = reinterpret_cast<uint64_t*>(smem_raw);
    const uint32_t q = blockIdx.x, tid = threadIdx.x, n = min(cnt[q], cap);
    uint32_t n2 = 2;
    while (n2 < n) n2 <<= 1;
    for (uint32_t i = tid; i < n2; i += kSelThreads)
        buf[i] = i < n ? make_key(es[(size_t)q * cap + i], (uint32_t)cand[(size_t)q * cap + i]) : 0ull;
    bitonic_sort_desc(buf, n2);
    for (uint32_t r = tid; r < k; r += kSelThreads) {
        const bool ok = r < n;
        s_out[(size_t)q * k + r] = ok ? key_score(buf[r]) : -3.4028234663852886e38f;
        id_out[(size_t)q * k + r] = ok ? (int64_t)(int32_t)key_doc(buf[r]) : -1;
    }
}

// P5 (device side): list the uncertified queries of a block (at most kDenseFallbackMax are re-run here), count them
__global__ void dense_fail_list_kernel(const uint32_t* __restrict__ flags, uint32_t n, uint32_t* __restrict__ gidx,
                                       uint32_t* __restrict__ n_list, uint32_t* __restrict__ fail, uint32_t cap) {
    __shared__ uint32_t s_n;
    if (threadIdx.x == 0) s_n = 0;
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x)
        if (!flags[i]) { const uint32_t p = atomicAdd(&s_n, 1u); if (p < cap) gidx[p] = i; }
    __syncthreads();
    if (threadIdx.x == 0) {
        *n_list = min(s_n, cap);
        if (s_n) atomicAdd(&fail[0], s_n);
        if (s_n > cap) atomicOr(&fail[1], 1u);
    }
}

__global__ void gather_rows_kernel(const float* __restrict__ src, const uint32_t* __restrict__ idx, uint32_t n, uint32_t dim,
                                   float* __restrict__ dst, const uint32_t* __restrict__ n_dev) {
    const uint32_t r = blockIdx.x;
    if (n_dev) n = min(n, *n_dev);
    if (r >= n) return;
    for (uint32_t i = threadIdx.x; i < dim; i += blockDim.x) dst[(size_t)r * dim + i] = src[(size_t)idx[r] * dim + i];
}
__global__ void scatter_results_kernel(const float* __restrict__ s_in, const int64_t* __restrict__ i_in,
                                       const uint32_t* __restrict__ idx, uint32_t n, uint32_t k, float* __restrict__ s_out,
                                       int64_t* __restrict__ i_out, const uint32_t* __restrict__ n_dev) {
    const uint32_t r = blockIdx.x;
    if (n_dev) n = min(n, *n_dev);
    if (r >= n) return;
    for (uint32_t j = threadIdx.x; j < k; j += blockDim.x) {
        s_out[(size_t)idx[r] * k + j] = s_in[(size_t)r * k + j];
        i_out[(size_t)idx[r] * k + j] = i_in[(size_t)r * k + j];
    }
}

static size_t dense_scratch_bytes(const msr_dense_index* h, uint32_t nq, uint32_t k) {
    const uint32_t kk = std::min(k, h->n_docs);
    const size_t per_q = align_up(h->n_docs, 4) * 4 + (size_t)kk * 8 + 4;
    const size_t want = per_q * nq + 4096;
    const size_t floor_b = per_q * std::min<uint32_t>(nq, kIpQ) + 4096;
    return std::max(floor_b, std::min<size_t>(want, 2ull << 30));
}

// workspace of the tensor-core path for a block of nqb queries
struct TcLayout {
    uint32_t nqb, kk, cap, m, stride, r;
    size_t ld_m, exact_bytes;
    size_t o_q16, o_samp, o_ts, o_td, o_tc, o_thr, o_cnt, o_cand, o_cd, o_es, o_ccnt, o_flags, o_qn, o_qe, o_gq, o_gos, o_goi,
        o_gidx, o_nlist, o_exact, total;
};

static TcLayout make_tc_layout(const msr_dense_index* h, uint32_t nq, uint32_t k) {
    TcLayout L{};
    const uint32_t N = h->n_docs;
    L.kk = std::min(k, N);
    // candidate buffer per query: the survivor count is ~ Gamma(r) * N / m, so its tail must stay below cap (a query that
    // overflows is re-run by the exact pass, which reads the whole fp32 index again)
    uint32_t cap = 4096;
    while (cap < 3 * L.kk && cap < 8192) cap <<= 1;
    L.cap = cap;
    if (N <= cap) {                                              // every doc fits the candidate buffer: no threshold needed
        L.m = 0; L.stride = 1; L.r = 0;
    } else {
        L.m = std::min<uint32_t>(N, std::min<uint32_t>(65536, std::max<uint32_t>(2048, N / 16)));
        L.stride = N / L.m;
        const double lam = (double)L.kk * L.m / N;
        // rank of the sample that estimates the threshold: the sample holds Poisson(lam) of the true top-k, and the estimate
        // is too high (query re-run by the exact pass, which reads the whole fp32 index) when it holds r or more.  4 sigma
        // leaves ~2e-5 per query -- one re-run per step over 8 GPUs x 4096 queries at config 5, 3 ms each, and every GPU
        // waits for it -- so take 6 sigma (~1e-9) where the expected survivor count r * N / m stays below a quarter of cap.
        const double r4 = std::ceil(lam + 4.0 * std::sqrt(lam) + 2.0), r6 = std::ceil(lam + 6.0 * std::sqrt(lam) + 4.0);
        const double r_cap = std::floor((double)cap / 4.0 * L.m / N);
        L.r = (uint32_t)std::min<double>(std::max(r4, std::min(r6, r_cap)), (double)std::min<uint32_t>(L.m, MSR_MAX_K));
    }
    L.ld_m = align_up(std::max<uint32_t>(L.m, 4), 4);
    const size_t per_q = (size_t)h->dim * 2 + L.ld_m * 4 + (size_t)L.r * 8 + (size_t)cap * 16 + (size_t)h->dim * 4 + (size_t)k * 12 + 64;
    uint32_t nqb = (uint32_t)std::min<size_t>(nq, std::max<size_t>(128, ((size_t)3 << 30) / per_q / 128 * 128));
    L.nqb = nqb;
    L.exact_bytes = dense_scratch_bytes(h, std::min<uint32_t>(nqb, kDenseFallbackMax), k);   // one pass for a full fallback list
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 256); return at; };
    L.o_q16 = take((size_t)nqb * h->dim * 2);
    L.o_samp = take((size_t)nqb * L.ld_m * 4);
    L.o_ts = take((size_t)nqb * std::max<uint32_t>(L.r, 1) * 4); L.o_td = take((size_t)nqb * std::max<uint32_t>(L.r, 1) * 4);
    L.o_tc = take((size_t)nqb * 4);
    L.o_thr = take((size_t)nqb * 4); L.o_cnt = take((size_t)nqb * 4);
    L.o_cand = take((size_t)nqb * cap * 8); L.o_cd = take((size_t)nqb * cap * 4); L.o_es = take((size_t)nqb * cap * 4);
    L.o_ccnt = take((size_t)nqb * 4); L.o_flags = take((size_t)nqb * 4);
    L.o_qn = take((size_t)nqb * 4); L.o_qe = take((size_t)nqb * 4);
    L.o_gq = take((size_t)nqb * h->dim * 4); L.o_gos = take((size_t)nqb * k * 4); L.o_goi = take((size_t)nqb * k * 8);
    L.o_gidx = take((size_t)nqb * 4);
    L.o_nlist = take(4);
    L.o_exact = take(L.exact_bytes);
    L.total = o;
    return L;
}

// Mode 2 over queries already on the device.  Synchronises the stream once per query block (to read the
// certificate flags); *n_fallback counts the queries that had to be re-run by the exact pass.
static int dense_search_tc(msr_dense_index* h, uint32_t nq, const float* d_q, uint32_t k, float* d_os, int64_t* d_oi,
                           unsigned char* base, const TcLayout& L, cudaStream_t s, uint32_t* n_fallback, bool on_device) {
    std::vector<uint32_t> flags, fail;
    static std::atomic<uint64_t> attr_set{0};
    if (attr_needed(attr_set)) {
        MSR_CUDA_CHECK(cudaFuncSetAttribute(dense_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 8));
        MSR_CUDA_CHECK(cudaFuncSetAttribute(dense_final_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 8));
        attr_mark(attr_set);
    }
    for (uint32_t q0 = 0; q0 < nq; q0 += L.nqb) {
        const uint32_t n = std::min(L.nqb, nq - q0);
        const float* q = d_q + (size_t)q0 * h->dim;
        __half* q16 = (__half*)(base + L.o_q16);
        float* samp = (float*)(base + L.o_samp);
        float* ts = (float*)(base + L.o_ts); int32_t* td = (int32_t*)(base + L.o_td); uint32_t* tcn = (uint32_t*)(base + L.o_tc);
        float* thr = (float*)(base + L.o_thr); uint32_t* cnt = (uint32_t*)(base + L.o_cnt);
        uint2* cand = (uint2*)(base + L.o_cand); int32_t* cd = (int32_t*)(base + L.o_cd); float* es = (float*)(base + L.o_es);
        uint32_t* ccnt = (uint32_t*)(base + L.o_ccnt); uint32_t* d_flags = (uint32_t*)(base + L.o_flags);
        float* qn = (float*)(base + L.o_qn); float* qe = (float*)(base + L.o_qe);
        const size_t ne = (size_t)n * h->dim;
        tc::to_f16_kernel<<<(unsigned)std::min<size_t>((ne + 255) / 256, 148 * 16), 256, 0, s>>>(q, q16, ne);
        dense_query_norms_kernel<<<(n + 7) / 8, 256, 0, s>>>(q, n, h->dim, qn, qe);
        count_launch(2);
        MSR_CUDA_CHECK(cudaMemsetAsync(cnt, 0, (size_t)n * 4, s));
        int rc;
        // P1: threshold from the strided sample
        if (L.m) {
            rc = tc::dense_tc_scores_launch(q16, n, h->v16, L.m, h->dim, (uint64_t)L.stride * h->dim, samp, L.ld_m, s);
            if (rc) return rc;
            rc = topk_rows_launch(n, L.m, samp, L.ld_m, L.r, 0, ts, td, tcn, s);
            if (rc) return rc;
            dense_thr_kernel<<<(n + 255) / 256, 256, 0, s>>>(ts, tcn, L.r, qn, qe, h->stats, n, h->dim, thr);
        } else {
            fill_f32_kernel<<<(n + 255) / 256, 256, 0, s>>>(thr, -INFINITY, n);
        }
        count_launch();
        // P2: all docs, fused filter epilogue
        if (h->profiling) MSR_CUDA_CHECK(cudaEventRecord(h->ev[0], s));
        rc = tc::dense_tc_filter_launch(q16, n, h->v16, h->n_docs, h->dim, thr, cnt, cand, L.cap, s);
        if (rc) return rc;
        if (h->profiling) MSR_CUDA_CHECK(cudaEventRecord(h->ev[1], s));
        // P3: candidate set + certificate, P4: exact re-score + final order
        dense_select_kernel<<<n, kSelThreads, (size_t)L.cap * 8, s>>>(cand, cnt, L.cap, thr, qn, qe, h->stats, L.kk, h->n_docs,
                                                                      h->dim, h->doc_base, cd, ccnt, d_flags);
        if (h->profiling) MSR_CUDA_CHECK(cudaEventRecord(h->ev[2], s));
        const uint32_t rs_split = std::max<uint32_t>(1, std::min<uint32_t>(16, (2 * 148 + n - 1) / n));
        dense_rescore_kernel<<<dim3(n, rs_split), kRsThreads, (size_t)h->dim * 4, s>>>(h->v32, h->dim, h->doc_base, q, cd, ccnt, L.cap, es);
        if (h->profiling) MSR_CUDA_CHECK(cudaEventRecord(h->ev[3], s));
        dense_final_kernel<<<n, kSelThreads, (size_t)L.cap * 8, s>>>(es, cd, ccnt, L.cap, k, d_os + (size_t)q0 * k,
                                                                     d_oi + (size_t)q0 * k);
        count_launch(3);
        MSR_CUDA_CHECK(cudaGetLastError());
        if (on_device) {
            // MSR_DEVICE stays asynchronous: the uncertified queries are listed, re-run by the exact pass and scattered over
            // the tensor-core rows by launches that take their count from device memory (an empty list costs four short
            // launches); at most kDenseFallbackMax per query block -- more sets the sticky flag msr_dense_last_fallback reports
            uint32_t* gidx = (uint32_t*)(base + L.o_gidx);
            uint32_t* n_list = (uint32_t*)(base + L.o_nlist);
            float* gq = (float*)(base + L.o_gq); float* gos = (float*)(base + L.o_gos); int64_t* goi = (int64_t*)(base + L.o_goi);
            const uint32_t F = std::min(kDenseFallbackMax, n);   // the scratch rows are sized for the block's own queries
            dense_fail_list_kernel<<<1, 256, 0, s>>>(d_flags, n, gidx, n_list, h->d_fail, F);
            gather_rows_kernel<<<F, 256, 0, s>>>(q, gidx, F, h->dim, gq, n_list);
            count_launch(2);
            rc = dense_search_exact(h, F, gq, k, gos, goi, base + L.o_exact, L.exact_bytes, s, n_list);
            if (rc) return rc;
            scatter_results_kernel<<<F, 128, 0, s>>>(gos, goi, gidx, F, k, d_os + (size_t)q0 * k, d_oi + (size_t)q0 * k, n_list);
            count_launch();
            MSR_CUDA_CHECK(cudaGetLastError());
            if (h->profiling) {
                MSR_CUDA_CHECK(cudaStreamSynchronize(s));
                float a = 0, b = 0;
                MSR_CUDA_CHECK(cudaEventElapsedTime(&a, h->ev[0], h->ev[1]));
                MSR_CUDA_CHECK(cudaEventElapsedTime(&b, h->ev[2], h->ev[3]));
                h->prof_tc_ms += a; h->prof_rescore_ms += b; h->prof_launches += 1;
            }
            continue;
        }
        flags.resize(n);
        MSR_CUDA_CHECK(cudaMemcpyAsync(flags.data(), d_flags, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
        MSR_CUDA_CHECK(cudaStreamSynchronize(s));
        if (h->profiling) {
            float a = 0, b = 0;
            MSR_CUDA_CHECK(cudaEventElapsedTime(&a, h->ev[0], h->ev[1]));
            MSR_CUDA_CHECK(cudaEventElapsedTime(&b, h->ev[2], h->ev[3]));
            h->prof_tc_ms += a; h->prof_rescore_ms += b; h->prof_launches += 1;
        }
        fail.clear();
        for (uint32_t i = 0; i < n; ++i) if (!flags[i]) fail.push_back(i);
        if (fail.empty()) continue;
        // uncertified queries: exact pass (still on the GPU), results scattered over the tensor-core ones
        *n_fallback += (uint32_t)fail.size();
        uint32_t* gidx = (uint32_t*)(base + L.o_gidx);
        float* gq = (float*)(base + L.o_gq); float* gos = (float*)(base + L.o_gos); int64_t* goi = (int64_t*)(base + L.o_goi);
        MSR_CUDA_CHECK(cudaMemcpyAsync(gidx, fail.data(), fail.size() * 4, cudaMemcpyHostToDevice, s));
        const uint32_t nf = (uint32_t)fail.size();
        gather_rows_kernel<<<nf, 256, 0, s>>>(q, gidx, nf, h->dim, gq, nullptr);
        count_launch();
        rc = dense_search_exact(h, nf, gq, k, gos, goi, base + L.o_exact, L.exact_bytes, s);
        if (rc) return rc;
        scatter_results_kernel<<<nf, 128, 0, s>>>(gos, goi, gidx, nf, k, d_os + (size_t)q0 * k, d_oi + (size_t)q0 * k, nullptr);
        count_launch();
        MSR_CUDA_CHECK(cudaGetLastError());
        MSR_CUDA_CHECK(cudaStreamSynchronize(s));               // `fail` (pageable) must outlive the copy
    }
    return MSR_OK;
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
    // mode 0: both paths return the same bits, so this is a pure cost rule (profiles/r2_dense_modes.txt, cold, top-100):
    //   exact pass   ~ 50 us + 4 bytes per element per group of 4 queries   (25 000 x 4096: 112 / 138 / 224 us at nq 1 / 4 / 8)
    //   tensor path  ~170 us + 2 bytes per element, any nq up to 128         (25 000 x 4096: 182-195 us; 1.25 M x 4096: 1.9 ms
    //                                                                         against 3.4 ms exact at nq = 1)
    const size_t bytes16 = (size_t)h->n_docs * h->dim * 2;
    const bool tc_pays = nq >= 8 || (nq >= 3 && bytes16 >= ((size_t)500 << 20)) || bytes16 >= ((size_t)800 << 20);
    bool use_tc = mode == 2 || (mode == 0 && h->n_docs >= 4096 && tc_pays);
    if (use_tc && h->dim % 8 != 0) {
        MSR_REQUIRE(mode != 2, "msr_dense_search: mode 2 needs dim %% 8 == 0 (TMA row pitch), dim is %u", h->dim);
        use_tc = false;
    }
    TcLayout L{};
    if (use_tc) L = make_tc_layout(h, nq, k);
    const size_t sb = use_tc ? L.total : align_up(dense_scratch_bytes(h, nq, k), 256);
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
    if (use_tc) {
        h->last_fallback = 0;
        MSR_CUDA_CHECK(cudaMemsetAsync(h->d_fail, 0, 8, s));
        rc = dense_search_tc(h, nq, d_q, k, d_os, d_oi, base, L, s, &h->last_fallback, where == MSR_DEVICE);
    } else {
        rc = dense_search_exact(h, nq, d_q, k, d_os, d_oi, base, sb, s);
    }
    if (rc) return rc;
    if (where == MSR_HOST) {
        MSR_CUDA_CHECK(cudaMemcpyAsync(out_scores, d_os, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, s));
        MSR_CUDA_CHECK(cudaMemcpyAsync(out_ids, d_oi, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, s));
        MSR_CUDA_CHECK(cudaStreamSynchronize(s));
    }
    return MSR_OK;
}


// ---------------------------------------------------------------------------------------------------------
// tensor-core pass on its own (tests, bench.py --dense, ncu): scores[nq, n_docs] = fp16(Q) . fp16(X)^T, fp32 out
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
    __half* q16 = (__half*)h->ws.p;
    const size_t n = (size_t)nq * h->dim;
    tc::to_f16_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 148 * 16), 256, 0, s>>>(queries, q16, n);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    return tc::dense_tc_scores_launch(q16, nq, h->v16, h->n_docs, h->dim, h->dim, out_scores, ld, s);
}

extern "C" int msr_dense_profile(msr_dense_index* h, int enable, double* tc_ms, double* rescore_ms, uint64_t* launches) {
    MSR_REQUIRE(h, "msr_dense_profile: NULL handle");
    int rc = use_device(h->device);
    if (rc) return rc;
    if (tc_ms) *tc_ms = h->prof_tc_ms;
    if (rescore_ms) *rescore_ms = h->prof_rescore_ms;
    if (launches) *launches = h->prof_launches;
    h->prof_tc_ms = h->prof_rescore_ms = 0; h->prof_launches = 0;
    if (enable && !h->ev[0])
        for (cudaEvent_t& e : h->ev) MSR_CUDA_CHECK(cudaEventCreate(&e));
    h->profiling = enable != 0;
    return MSR_OK;
}

extern "C" int msr_dense_last_fallback(const msr_dense_index* h, uint32_t* n) {
    MSR_REQUIRE(h && n, "msr_dense_last_fallback: NULL argument");
    int rc = use_device(h->device);
    if (rc) return rc;
    MSR_CUDA_CHECK(cudaDeviceSynchronize());                    // the device path counts on the device
    uint32_t f[2] = {0, 0};
    MSR_CUDA_CHECK(cudaMemcpy(f, h->d_fail, 8, cudaMemcpyDeviceToHost));
    *n = h->last_fallback + f[0];
    if (f[1]) {
        set_error("msr_dense_search (device path): more than %u uncertified queries in one query block; the first %u were re-run "
                  "exactly, the others kept their tensor-core rows -- re-run the batch with mode 1 or through the host path",
                  kDenseFallbackMax, kDenseFallbackMax);
        return MSR_ERR_RANGE;
    }
    return MSR_OK;
}

// ---------------------------------------------------------------------------------------------------------
// msr_dense_search_static: the exact pass (mode 1, device pointers) for callers that may not allocate
// ---------------------------------------------------------------------------------------------------------
// No counterpart in the reference.  Callers under CUDA graph capture, in a
// tight serving loop or in a harness that forbids cudaMalloc cannot use msr_dense_search, whose score rows and top-k
// staging live in the handle's grow-only workspace (Workspace::reserve = cudaMalloc on first use and on growth).  Here the
// same scratch is a file-scope device array of MSR_STATIC_SCRATCH_BYTES: every GPU that loads libmsr.so holds it until
// the process exits, every call on that GPU shares it, and it keeps the last call's contents.
//
// In host code `g_static_scratch` is only a host-side stand-in for the symbol, NOT a device address: its device address is
// asked for once per device, with that device current (cudaGetSymbolAddress), and that pointer is what the kernels get.
static __device__ __align__(256) unsigned char g_static_scratch[MSR_STATIC_SCRATCH_BYTES];

namespace msr {

constexpr int kStaticMaxDevices = 64;
static std::atomic<unsigned char*> g_static_ptr[kStaticMaxDevices];

// device address of g_static_scratch on the CURRENT device (the caller has made h->device current)
static int static_scratch_ptr(int device, unsigned char** out) {
    MSR_REQUIRE(device >= 0 && device < kStaticMaxDevices, "msr_dense_search_static: device %d outside [0,%d)", device, kStaticMaxDevices);
    unsigned char* p = g_static_ptr[device].load(std::memory_order_acquire);
    if (!p) {
        void* sym = nullptr;
        MSR_CUDA_CHECK(cudaGetSymbolAddress(&sym, g_static_scratch));
        MSR_REQUIRE(sym != nullptr, "msr_dense_search_static: cudaGetSymbolAddress returned NULL on device %d", device);
        p = (unsigned char*)sym;
        g_static_ptr[device].store(p, std::memory_order_release);   // racing threads store the same address
    }
    *out = p;
    return MSR_OK;
}

}  // namespace msr

extern "C" int msr_dense_search_static(msr_dense_index* h, uint32_t nq, const float* queries, uint32_t k, float* out_scores,
                                       int64_t* out_ids, void* stream) {
    MSR_REQUIRE(h, "msr_dense_search_static: NULL handle");
    MSR_REQUIRE(k >= 1 && k <= MSR_MAX_K, "msr_dense_search_static: k=%u outside [1,%u]", k, MSR_MAX_K);
    MSR_REQUIRE(h->n_docs > 0, "msr_dense_search_static: the index is empty");
    // one group of kIpQ score rows + its top-k staging must fit the static array (dense_search_exact then walks the
    // queries in passes of as many rows as fit)
    const size_t per_q = align_up(h->n_docs, 4) * 4 + (size_t)std::min(k, h->n_docs) * 8 + 4;
    const size_t need = per_q * kIpQ + 4096;
    MSR_REQUIRE(need <= (size_t)MSR_STATIC_SCRATCH_BYTES,
                "msr_dense_search_static: n_docs=%u, k=%u need %zu bytes of scratch (%d score rows of n_docs fp32 + top-k staging), "
                "the static array holds MSR_STATIC_SCRATCH_BYTES=%zu (n_docs <= ~%zu at this k); use msr_dense_search",
                h->n_docs, k, need, kIpQ, (size_t)MSR_STATIC_SCRATCH_BYTES,
                ((size_t)MSR_STATIC_SCRATCH_BYTES - 4096) / kIpQ / 4 - (size_t)std::min(k, h->n_docs) * 2 - 4);
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(queries && out_scores && out_ids, "msr_dense_search_static: NULL buffer");
    int rc = use_device(h->device);
    if (rc) return rc;
    unsigned char* scratch = nullptr;
    rc = static_scratch_ptr(h->device, &scratch);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    // Nothing in the scratch is read before this call has written it (score rows: every doc column of every row of a pass;
    // staging lists and counts: written by the row top-k before the finishing kernel reads them), so the last call's
    // contents are harmless.  The per-row count words are the one thing a reader could take for carried-over state: they
    // start at zero anyway (same layout arithmetic as dense_search_exact).
    {
        const size_t row_stride = align_up(h->n_docs, 4), kk = std::min(k, h->n_docs);
        size_t qc = std::min<size_t>(nq, std::max<size_t>(1, ((size_t)MSR_STATIC_SCRATCH_BYTES - 1024) / per_q));
        if (qc >= (size_t)kIpQ) qc = qc / kIpQ * kIpQ;
        const size_t off_counts = align_up(qc * row_stride * 4, 256) + 2 * align_up(qc * kk * 4, 256);
        MSR_CUDA_CHECK(cudaMemsetAsync(scratch + off_counts, 0, qc * 4, s));
    }
    return dense_search_exact(h, nq, queries, k, out_scores, out_ids, scratch, (size_t)MSR_STATIC_SCRATCH_BYTES, s);
}
