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

#include <vector>

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
    float* stats = nullptr;          // [2] max row norm, max row bf16 error over the whole index (certificate of mode 2)
    msr::Workspace ws;
    uint32_t last_fallback = 0;      // queries of the last mode-2 search that needed the exact pass
};

namespace msr {

constexpr int kIpThreads = 256;
constexpr int kIpWarps = kIpThreads / 32;
constexpr int kIpQ = 4;              // queries held in shared memory per pass

// fp32 -> bf16 copy + per-row quantisation error norm
__global__ void dense_convert_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                     float* __restrict__ err, float* __restrict__ stats, uint32_t n_rows, uint32_t dim) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n_rows) return;
    const float* s = src + (size_t)warp * dim;
    __nv_bfloat16* d = dst + (size_t)warp * dim;
    float e2 = 0.f, n2 = 0.f;
    for (uint32_t i = lane; i < dim; i += 32) {
        float x = s[i];
        __nv_bfloat16 b = __float2bfloat16_rn(x);
        d[i] = b;
        float r = x - __bfloat162float(b);
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
    cudaFree(h->v32); cudaFree(h->v16); cudaFree(h->err16); cudaFree(h->stats);
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

// ---- mode 2 helpers: tensor-core candidates -> exact fp32 re-score -> certificate ----

// per query: ||q||_2 and ||q - bf16(q)||_2 (both rounded up a little)
__global__ void dense_query_norms_kernel(const float* __restrict__ q, uint32_t nq, uint32_t dim, float* __restrict__ qn,
                                         float* __restrict__ qe) {
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= nq) return;
    const float* s = q + (size_t)w * dim;
    float n2 = 0.f, e2 = 0.f;
    for (uint32_t i = lane; i < dim; i += 32) {
        const float x = s[i], r = x - __bfloat162float(__float2bfloat16_rn(x));
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

// Exact score of candidate r of query q: one warp per (q, r), the SAME accumulation order as dense_ip_exact_kernel
// (lane-strided fmaf chain, xor-shuffle tree), so both modes return bit-identical scores.
__global__ void __launch_bounds__(256)
dense_rescore_kernel(const float* __restrict__ docs, uint32_t dim, int32_t doc_base, const float* __restrict__ queries,
                     const int32_t* __restrict__ cand, const uint32_t* __restrict__ cnt, uint32_t nq, uint32_t kp,
                     float* __restrict__ out) {
    const size_t w = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31;
    if (w >= (size_t)nq * kp) return;
    const uint32_t q = (uint32_t)(w / kp), r = (uint32_t)(w - (size_t)q * kp);
    if (r >= cnt[q]) return;
    const float* d = docs + (size_t)(cand[w] - doc_base) * dim;
    const float* y = queries + (size_t)q * dim;
    float acc = 0.f;
    if ((dim & 3) == 0) {
        const float4* d4 = reinterpret_cast<const float4*>(d);
        const float4* y4 = reinterpret_cast<const float4*>(y);
        for (uint32_t i = lane; i < dim / 4; i += 32) {
            const float4 a = __ldg(d4 + i), b = __ldg(y4 + i);
            acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc); acc = fmaf(a.z, b.z, acc); acc = fmaf(a.w, b.w, acc);
        }
    } else {
        for (uint32_t i = lane; i < dim; i += 32) acc = fmaf(__ldg(d + i), __ldg(y + i), acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) out[w] = acc;
}

// flags[q] = 1 when no doc outside the candidate list can reach the exact k-th score:
//   exact(d) <= approx(d) + E_q  for every doc, with
//   E_q = ||q - q^||*max||x|| + ||q^||*max||x - x^|| + (fp32 accumulation slack of both passes),
// and approx(d) <= a_min (the smallest approximate score among the candidates) for every d outside the list.
__global__ void dense_certify_kernel(const float* __restrict__ ts, const uint32_t* __restrict__ tcnt, const float* __restrict__ fs,
                                     const float* __restrict__ qn, const float* __restrict__ qe, const float* __restrict__ stats,
                                     uint32_t nq, uint32_t kp, uint32_t kk, uint32_t n_docs, uint32_t dim,
                                     uint32_t* __restrict__ flags) {
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const uint32_t c = tcnt[q];
    bool ok = c >= n_docs;                                       // every doc was re-scored
    if (!ok && c == kp && kp >= kk) {
        const float a_min = ts[(size_t)q * kp + kp - 1], e_k = fs[(size_t)q * kk + kk - 1];
        const float xn = stats[0], ex = stats[1], qh = qn[q] + qe[q];
        const float slack = ((float)dim * 1.1920929e-7f + 9.5367432e-7f) * qh * (xn + ex);
        const float E = qe[q] * xn + qh * ex + slack;
        ok = a_min + E * 1.0001f < e_k;                          // NaN anywhere -> not certified
    }
    flags[q] = ok ? 1u : 0u;
}

__global__ void gather_rows_kernel(const float* __restrict__ src, const uint32_t* __restrict__ idx, uint32_t n, uint32_t dim,
                                   float* __restrict__ dst) {
    const uint32_t r = blockIdx.x;
    if (r >= n) return;
    for (uint32_t i = threadIdx.x; i < dim; i += blockDim.x) dst[(size_t)r * dim + i] = src[(size_t)idx[r] * dim + i];
}
__global__ void scatter_results_kernel(const float* __restrict__ s_in, const int64_t* __restrict__ i_in,
                                       const uint32_t* __restrict__ idx, uint32_t n, uint32_t k, float* __restrict__ s_out,
                                       int64_t* __restrict__ i_out) {
    const uint32_t r = blockIdx.x;
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
    uint32_t nqb, kp, kk;
    size_t ld, scores_bytes;
    size_t o_scores, o_q16, o_ts, o_td, o_tc, o_es, o_fs, o_fd, o_fc, o_qn, o_qe, o_flags, o_gq, o_gos, o_goi, o_gidx, total;
};

static TcLayout make_tc_layout(const msr_dense_index* h, uint32_t nq, uint32_t k) {
    TcLayout L{};
    L.kk = std::min(k, h->n_docs);
    L.kp = std::min<uint32_t>(h->n_docs, std::min<uint32_t>(MSR_MAX_K, std::max<uint32_t>(4 * k, 256u)));
    L.kp = std::max(L.kp, L.kk);
    L.ld = align_up(h->n_docs, 4);
    const size_t row = L.ld * 4;
    uint32_t nqb = (uint32_t)std::min<size_t>(nq, std::max<size_t>(128, ((size_t)2 << 30) / row / 128 * 128));
    L.nqb = nqb;
    // the score block doubles as the scratch of the exact pass that re-runs uncertified queries
    L.scores_bytes = std::max((size_t)nqb * row, dense_scratch_bytes(h, std::min<uint32_t>(nqb, 64), k));
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 256); return at; };
    L.o_scores = take(L.scores_bytes);
    L.o_q16 = take((size_t)nqb * h->dim * 2);
    L.o_ts = take((size_t)nqb * L.kp * 4); L.o_td = take((size_t)nqb * L.kp * 4); L.o_tc = take((size_t)nqb * 4);
    L.o_es = take((size_t)nqb * L.kp * 4);
    L.o_fs = take((size_t)nqb * L.kk * 4); L.o_fd = take((size_t)nqb * L.kk * 4); L.o_fc = take((size_t)nqb * 4);
    L.o_qn = take((size_t)nqb * 4); L.o_qe = take((size_t)nqb * 4); L.o_flags = take((size_t)nqb * 4);
    L.o_gq = take((size_t)nqb * h->dim * 4); L.o_gos = take((size_t)nqb * k * 4); L.o_goi = take((size_t)nqb * k * 8);
    L.o_gidx = take((size_t)nqb * 4);
    L.total = o;
    return L;
}

// Mode 2 over queries already on the device.  Synchronises the stream once per query block (to read the
// certificate flags); *n_fallback counts the queries that had to be re-run by the exact pass.
static int dense_search_tc(msr_dense_index* h, uint32_t nq, const float* d_q, uint32_t k, float* d_os, int64_t* d_oi,
                           unsigned char* base, const TcLayout& L, cudaStream_t s, uint32_t* n_fallback) {
    std::vector<uint32_t> flags, fail;
    for (uint32_t q0 = 0; q0 < nq; q0 += L.nqb) {
        const uint32_t n = std::min(L.nqb, nq - q0);
        const float* q = d_q + (size_t)q0 * h->dim;
        float* scores = (float*)(base + L.o_scores);
        __nv_bfloat16* q16 = (__nv_bfloat16*)(base + L.o_q16);
        float* ts = (float*)(base + L.o_ts); int32_t* td = (int32_t*)(base + L.o_td); uint32_t* tcn = (uint32_t*)(base + L.o_tc);
        float* es = (float*)(base + L.o_es);
        float* fs = (float*)(base + L.o_fs); int32_t* fd = (int32_t*)(base + L.o_fd); uint32_t* fc = (uint32_t*)(base + L.o_fc);
        float* qn = (float*)(base + L.o_qn); float* qe = (float*)(base + L.o_qe); uint32_t* d_flags = (uint32_t*)(base + L.o_flags);
        // 1. approximate scores on the tensor cores
        const size_t ne = (size_t)n * h->dim;
        tc::to_bf16_kernel<<<(unsigned)std::min<size_t>((ne + 255) / 256, 148 * 16), 256, 0, s>>>(q, q16, ne);
        dense_query_norms_kernel<<<(n + 7) / 8, 256, 0, s>>>(q, n, h->dim, qn, qe);
        count_launch(2);
        int rc = tc::dense_tc_scores_launch(q16, n, h->v16, h->n_docs, h->dim, scores, L.ld, s);
        if (rc) return rc;
        // 2. k' best by approximate score, 3. exact fp32 re-score, 4. final order (score desc, id asc)
        rc = topk_rows_launch(n, h->n_docs, scores, L.ld, L.kp, h->doc_base, ts, td, tcn, s);
        if (rc) return rc;
        const size_t warps = (size_t)n * L.kp;
        dense_rescore_kernel<<<(unsigned)((warps + 7) / 8), 256, 0, s>>>(h->v32, h->dim, h->doc_base, q, td, tcn, n, L.kp, es);
        count_launch();
        MSR_CUDA_CHECK(cudaGetLastError());
        rc = msr_topk_merge(h->device, 1, n, L.kp, es, td, tcn, L.kk, fs, fd, fc, (void*)s);
        if (rc) return rc;
        // 5. certificate, FAISS-convention output
        dense_certify_kernel<<<(n + 255) / 256, 256, 0, s>>>(ts, tcn, fs, qn, qe, h->stats, n, L.kp, L.kk, h->n_docs, h->dim, d_flags);
        dense_finish_kernel<<<(unsigned)(((size_t)n * k + 255) / 256), 256, 0, s>>>(fs, fd, fc, n, L.kk, k, d_os + (size_t)q0 * k,
                                                                                   d_oi + (size_t)q0 * k);
        count_launch(2);
        MSR_CUDA_CHECK(cudaGetLastError());
        flags.resize(n);
        MSR_CUDA_CHECK(cudaMemcpyAsync(flags.data(), d_flags, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
        MSR_CUDA_CHECK(cudaStreamSynchronize(s));
        fail.clear();
        for (uint32_t i = 0; i < n; ++i) if (!flags[i]) fail.push_back(i);
        if (fail.empty()) continue;
        // 6. uncertified queries: exact pass (still on the GPU), results scattered over the tensor-core ones
        *n_fallback += (uint32_t)fail.size();
        uint32_t* gidx = (uint32_t*)(base + L.o_gidx);
        float* gq = (float*)(base + L.o_gq); float* gos = (float*)(base + L.o_gos); int64_t* goi = (int64_t*)(base + L.o_goi);
        MSR_CUDA_CHECK(cudaMemcpyAsync(gidx, fail.data(), fail.size() * 4, cudaMemcpyHostToDevice, s));
        const uint32_t nf = (uint32_t)fail.size();
        gather_rows_kernel<<<nf, 256, 0, s>>>(q, gidx, nf, h->dim, gq);
        count_launch();
        rc = dense_search_exact(h, nf, gq, k, gos, goi, base + L.o_scores, L.scores_bytes, s);
        if (rc) return rc;
        scatter_results_kernel<<<nf, 128, 0, s>>>(gos, goi, gidx, nf, k, d_os + (size_t)q0 * k, d_oi + (size_t)q0 * k);
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
    bool use_tc = mode == 2 || (mode == 0 && nq >= 16 && h->n_docs >= 256);
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
        rc = dense_search_tc(h, nq, d_q, k, d_os, d_oi, base, L, s, &h->last_fallback);
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

extern "C" int msr_dense_last_fallback(const msr_dense_index* h, uint32_t* n) {
    MSR_REQUIRE(h && n, "msr_dense_last_fallback: NULL argument");
    *n = h->last_fallback;
    return MSR_OK;
}
