// quantise.cu -- on-device query / document vector construction ("next" row f1 of SURVEY.md 8).
//
// Replaces, for the full-vocabulary branch, the host round trip of the reference's quantiser and query-string builder:
//   get_img_valid_tokens_values            /root/reference/src/encode.py:56-93   (logits.topk -> np.rint(v*100).astype(int))
//   get_text_valid_tokens_values (manual)  /root/reference/src/encode.py:114-124
//   query string  query += (' '+token)*v   /root/reference/src/search.py:419-422, 439-442  (duplicate tokens SUM, v<=0 drops)
//   doc vector    vector[token] = int(v)   /root/reference/src/encode.py:351-352, 369-370  (duplicate tokens: LAST wins)
//   pyserini TokFreqQueryEncoder + dictionary lookup (rule S4)
// and leaves the result as a query CSR in HBM that msr_sparse_search consumes without touching the host.
//
// Per activation row: radix-select top-`top` by (value desc, token id asc) [topk.cu], then one warp per row:
//   * token ids >= vocab_len are skipped the way the reference skips them: the token list shrinks but the value list
//     does not, so the j-th surviving token is paired with the j-th largest value (encode.py:75-83);
//   * weight = rint(float32(value) * 100) in float32, round-half-even, as numpy does;
//   * token id -> term id through `token_to_term` (host-built: lower-cased, optionally filter_token'ed vocabulary string
//     looked up in the index dictionary; -1 = not in the index, contributes nothing);
//   * tokens mapping to one term are merged: query mode sums their weights, doc mode keeps the last (smallest) one.
// Rows are written at a fixed pitch of `top` entries; unused slots carry weight 0, which the scorer ignores.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "common.cuh"

namespace msr {

int topk_rows_launch(uint32_t nq, uint32_t n, const float* scores, size_t row_stride, uint32_t k, int32_t col_base,
                     float* out_scores, int32_t* out_docs, uint32_t* out_counts, cudaStream_t s);

constexpr int kQuantMaxTop = 1024;

template <class T> __device__ __forceinline__ float to_f32(T x);
template <> __device__ __forceinline__ float to_f32<float>(float x) { return x; }
template <> __device__ __forceinline__ float to_f32<__half>(__half x) { return __half2float(x); }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 x) { return __bfloat162float(x); }

template <class T>
__global__ void rows_to_f32_kernel(const T* __restrict__ src, size_t src_stride, uint32_t n_rows, uint32_t n_cols,
                                   float* __restrict__ dst) {
    const size_t total = (size_t)n_rows * n_cols;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / n_cols, c = i - r * n_cols;
        dst[i] = to_f32<T>(src[r * src_stride + c]);
    }
}

// one warp per row; ts/td = the row's top entries sorted by (value desc, id asc), cnt = how many
__global__ void __launch_bounds__(128)
quantise_map_kernel(const float* __restrict__ ts, const int32_t* __restrict__ td, const uint32_t* __restrict__ cnt,
                    uint32_t n_rows, uint32_t top, uint32_t vocab_len, const int32_t* __restrict__ token_to_term,
                    int doc_mode, unsigned long long* __restrict__ out_indptr, uint32_t* __restrict__ out_terms,
                    uint32_t* __restrict__ out_weights) {
    extern __shared__ int32_t sm[];
    const uint32_t wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t row = blockIdx.x * (blockDim.x >> 5) + wib;
    int32_t* term = sm + (size_t)wib * 2 * top;          // term id of the j-th surviving token (-1 = not indexed)
    int32_t* wt = term + top;                             // its weight
    if (row == 0 && threadIdx.x == 0) out_indptr[n_rows] = (unsigned long long)n_rows * top;
    if (row >= n_rows) return;
    const uint32_t n = min(cnt[row], top);
    const float* v = ts + (size_t)row * top;
    const int32_t* id = td + (size_t)row * top;
    // 1. survivors in rank order, paired with the values of the leading ranks
    uint32_t kept = 0;
    for (uint32_t base = 0; base < n; base += 32) {
        const uint32_t r = base + lane;
        const int32_t t = r < n ? id[r] : -1;
        const bool ok = r < n && t >= 0 && (uint32_t)t < vocab_len;
        const uint32_t b = __ballot_sync(0xffffffffu, ok);
        if (ok) {
            const uint32_t j = kept + __popc(b & ((1u << lane) - 1u));
            term[j] = token_to_term ? token_to_term[t] : t;
            wt[j] = __float2int_rn(v[j] * 100.0f);        // np.rint(float32 * 100): round-half-even
        }
        kept += __popc(b);
    }
    __syncwarp();
    // 2. merge tokens of one term; the first occurrence carries the result
    uint32_t* o_t = out_terms + (size_t)row * top;
    uint32_t* o_w = out_weights + (size_t)row * top;
    uint32_t written = 0;
    for (uint32_t base = 0; base < kept; base += 32) {
        const uint32_t j = base + lane;
        int32_t t = j < kept ? term[j] : -1;
        long long w = 0;
        bool lead = t >= 0;
        if (lead) {
            for (uint32_t i = 0; i < j; ++i) if (term[i] == t) { lead = false; break; }
        }
        if (lead) {
            if (doc_mode) {                               // dict assignment: the last pair wins
                w = wt[j];
                for (uint32_t i = j + 1; i < kept; ++i) if (term[i] == t) w = wt[i];
            } else {                                      // string repetition: positive counts add up
                for (uint32_t i = j; i < kept; ++i) if (term[i] == t && wt[i] > 0) w += wt[i];
            }
        }
        const bool emit = lead && w > 0;
        const uint32_t b = __ballot_sync(0xffffffffu, emit);
        if (emit) {
            const uint32_t p = written + __popc(b & ((1u << lane) - 1u));
            o_t[p] = (uint32_t)t;
            o_w[p] = (uint32_t)min(w, (long long)0xFFFFFFFFu);
        }
        written += __popc(b);
    }
    for (uint32_t p = written + lane; p < top; p += 32) { o_t[p] = 0; o_w[p] = 0; }
    if (lane == 0) out_indptr[row] = (unsigned long long)row * top;
}

}  // namespace msr

using namespace msr;

extern "C" int msr_quantise_topk(int device, uint32_t n_rows, uint32_t vocab, const void* logits, int dtype,
                                 size_t row_stride, uint32_t top, uint32_t vocab_len, const int32_t* token_to_term,
                                 int doc_mode, uint64_t* out_indptr, uint32_t* out_terms, uint32_t* out_weights,
                                 void* stream) {
    MSR_REQUIRE(dtype >= 0 && dtype <= 2, "msr_quantise_topk: dtype %d outside {0 f32, 1 f16, 2 bf16}", dtype);
    MSR_REQUIRE(top >= 1 && top <= (uint32_t)kQuantMaxTop, "msr_quantise_topk: top=%u outside [1,%d]", top, kQuantMaxTop);
    MSR_REQUIRE(vocab >= top && row_stride >= vocab, "msr_quantise_topk: vocab (%u) must be >= top and row_stride >= vocab", vocab);
    MSR_REQUIRE(out_indptr, "msr_quantise_topk: NULL out_indptr");
    int rc = use_device(device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (n_rows == 0) {
        MSR_CUDA_CHECK(cudaMemsetAsync(out_indptr, 0, 8, s));
        return MSR_OK;
    }
    MSR_REQUIRE(logits && out_terms && out_weights, "msr_quantise_topk: NULL buffer");
    const uint32_t kk = top;
    // stream-ordered scratch: fp32 rows (when converting), top values / ids / counts
    float* f32 = nullptr; float* ts = nullptr; int32_t* td = nullptr; uint32_t* tc = nullptr;
    if (dtype != 0) MSR_CUDA_CHECK(cudaMallocAsync(&f32, (size_t)n_rows * vocab * 4, s));
    MSR_CUDA_CHECK(cudaMallocAsync(&ts, (size_t)n_rows * top * 4, s));
    MSR_CUDA_CHECK(cudaMallocAsync(&td, (size_t)n_rows * top * 4, s));
    MSR_CUDA_CHECK(cudaMallocAsync(&tc, (size_t)n_rows * 4, s));
    const float* rows = (const float*)logits;
    size_t stride = row_stride;
    if (dtype != 0) {
        const unsigned g = (unsigned)std::min<size_t>(((size_t)n_rows * vocab + 255) / 256, 148 * 32);
        if (dtype == 1) rows_to_f32_kernel<__half><<<g, 256, 0, s>>>((const __half*)logits, row_stride, n_rows, vocab, f32);
        else rows_to_f32_kernel<__nv_bfloat16><<<g, 256, 0, s>>>((const __nv_bfloat16*)logits, row_stride, n_rows, vocab, f32);
        count_launch();
        rows = f32; stride = vocab;
    }
    rc = topk_rows_launch(n_rows, vocab, rows, stride, kk, 0, ts, td, tc, s);
    if (rc == MSR_OK) {
        const uint32_t wpb = 4;
        quantise_map_kernel<<<(n_rows + wpb - 1) / wpb, wpb * 32, (size_t)wpb * 2 * kk * 4, s>>>(
            ts, td, tc, n_rows, kk, vocab_len, token_to_term, doc_mode, (unsigned long long*)out_indptr, out_terms, out_weights);
        count_launch();
        if (cudaGetLastError() != cudaSuccess) { set_error("msr_quantise_topk: launch failed"); rc = MSR_ERR_CUDA; }
    }
    if (f32) cudaFreeAsync(f32, s);
    cudaFreeAsync(ts, s); cudaFreeAsync(td, s); cudaFreeAsync(tc, s);
    return rc;
}
