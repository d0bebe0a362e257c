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

#include <algorithm>
#include <cmath>
#include <vector>

#include "common.cuh"

namespace msr {

int topk_rows_launch_typed(int dtype, uint32_t nq, uint32_t n, const void* scores, size_t row_stride, uint32_t k,
                           int32_t col_base, float* out_scores, int32_t* out_docs, uint32_t* out_counts, cudaStream_t s,
                           const uint32_t* nq_dev, uint32_t nq_off);
int topk_rows_launch(uint32_t nq, uint32_t n, const float* scores, size_t row_stride, uint32_t k, int32_t col_base,
                     float* out_scores, int32_t* out_docs, uint32_t* out_counts, cudaStream_t s, const uint32_t* nq_dev = nullptr,
                     uint32_t nq_off = 0);

constexpr int kQuantMaxTop = 1024;

// stream-ordered scratch buffers of one API call: freed (cudaFreeAsync on the call's stream) on every exit path
struct AsyncScratch {
    cudaStream_t s;
    std::vector<void*> ptrs;
    explicit AsyncScratch(cudaStream_t stream) : s(stream) {}
    ~AsyncScratch() { for (void* p : ptrs) cudaFreeAsync(p, s); }
    template <class T> cudaError_t alloc(T** out, size_t bytes) {
        void* p = nullptr;
        cudaError_t e = cudaMallocAsync(&p, bytes ? bytes : 1, s);
        if (e == cudaSuccess) ptrs.push_back(p);
        *out = reinterpret_cast<T*>(p);
        return e;
    }
};

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


// Candidate branch of get_text_valid_tokens_values (/root/reference/src/encode.py:96-135): the row's candidates are
// the token ids of the caption's content words (host tokenisation); keep the top-min(n_c, cap) of logits[cands] by
// (value desc, candidate position asc).  One CTA per row; ranks by counting in shared memory (n_c is ~10 for a caption).
// Writes the same (ts, td, cnt) triple the full-vocabulary selection produces, so quantise_map_kernel finishes the job.
constexpr int kCandMax = 4096;
template <class T>
__global__ void __launch_bounds__(128)
cand_select_kernel(const T* __restrict__ logits, size_t row_stride, const unsigned long long* __restrict__ c_indptr,
                   const int32_t* __restrict__ c_tok, uint32_t cap, uint32_t pitch, float* __restrict__ ts,
                   int32_t* __restrict__ td, uint32_t* __restrict__ cnt) {
    __shared__ float v[kCandMax];
    const uint32_t row = blockIdx.x;
    const unsigned long long lo = c_indptr[row];
    const uint32_t n = (uint32_t)(c_indptr[row + 1] - lo);
    if (n == 0) return;                                   // empty rows take the top-`fallback` path (host-driven)
    const T* x = logits + (size_t)row * row_stride;
    for (uint32_t j = threadIdx.x; j < n; j += blockDim.x) v[j] = to_f32<T>(x[c_tok[lo + j]]);
    __syncthreads();
    const uint32_t keep = min(n, cap);
    for (uint32_t j = threadIdx.x; j < n; j += blockDim.x) {
        const float a = v[j];
        uint32_t r = 0;
        for (uint32_t i = 0; i < n; ++i) {
            const float b = v[i];
            r += (b > a) || (b == a && i < j);
        }
        if (r < keep) { ts[(size_t)row * pitch + r] = a; td[(size_t)row * pitch + r] = c_tok[lo + j]; }
    }
    if (threadIdx.x == 0) cnt[row] = keep;
}


// Expansion terms of get_text_valid_tokens_values (/root/reference/src/encode.py:137-152): the top-n of the logits over
// `filtered_ids - token_ids`.  masked rows: -inf outside the expansion vocabulary, then -inf at the row's own candidates.
template <class T>
__global__ void expand_mask_rows_kernel(const T* __restrict__ logits, size_t row_stride, uint32_t row0, uint32_t n_rows,
                                        uint32_t vocab, const uint8_t* __restrict__ mask, float* __restrict__ dst) {
    const size_t total = (size_t)n_rows * vocab;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / vocab, c = i - r * vocab;
        dst[i] = mask[c] ? to_f32<T>(logits[(row0 + r) * row_stride + c]) : -INFINITY;
    }
}
__global__ void expand_exclude_kernel(const unsigned long long* __restrict__ c_indptr, const int32_t* __restrict__ c_tok,
                                      uint32_t row0, uint32_t vocab, float* __restrict__ dst) {
    const uint32_t r = blockIdx.x;
    const unsigned long long lo = c_indptr[row0 + r], hi = c_indptr[row0 + r + 1];
    for (unsigned long long j = lo + threadIdx.x; j < hi; j += blockDim.x) dst[(size_t)r * vocab + c_tok[j]] = -INFINITY;
}
// appends the finite expansion entries (already in descending order) behind the row's main entries
__global__ void expand_append_kernel(const float* __restrict__ es, const int32_t* __restrict__ ed, const uint32_t* __restrict__ ec,
                                     uint32_t n_expand, uint32_t row0, uint32_t n_rows, uint32_t pitch, float* __restrict__ ts,
                                     int32_t* __restrict__ td, uint32_t* __restrict__ cnt) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    const size_t row = (size_t)row0 + r;
    uint32_t n = cnt[row];
    const uint32_t have = min(ec[r], n_expand);                 // entries the select really returned (the rest is padding)
    for (uint32_t j = 0; j < have && n < pitch; ++j) {
        const float v = es[(size_t)r * n_expand + j];
        if (v == -INFINITY) break;
        ts[row * pitch + n] = v;
        td[row * pitch + n] = ed[(size_t)r * n_expand + j];
        ++n;
    }
    cnt[row] = n;
}

}  // namespace msr

using namespace msr;

// quantise_map_kernel keeps 2 * pitch words per warp in dynamic shared memory: above 48 KB that needs the opt-in attribute
static int map_kernel_smem(size_t bytes) {
    MSR_REQUIRE(bytes <= 200 * 1024, "quantiser: %zu bytes of shared memory per block (entries per row too large)", bytes);
    static std::atomic<uint64_t> attr_set{0};
    if (bytes > 48 * 1024 && attr_needed(attr_set)) {
        MSR_CUDA_CHECK(cudaFuncSetAttribute(quantise_map_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_mark(attr_set);
    }
    return MSR_OK;
}

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
    // stream-ordered scratch: top values / ids / counts
    AsyncScratch scratch(s);
    float* ts = nullptr; int32_t* td = nullptr; uint32_t* tc = nullptr;
    MSR_CUDA_CHECK(scratch.alloc(&ts, (size_t)n_rows * top * 4));
    MSR_CUDA_CHECK(scratch.alloc(&td, (size_t)n_rows * top * 4));
    MSR_CUDA_CHECK(scratch.alloc(&tc, (size_t)n_rows * 4));
    // rows are selected in place in their own type (no fp32 staging copy)
    rc = topk_rows_launch_typed(dtype, n_rows, vocab, logits, row_stride, kk, 0, ts, td, tc, s, nullptr, 0);
    if (rc == MSR_OK) rc = map_kernel_smem((size_t)4 * 2 * kk * 4);
    if (rc == MSR_OK) {
        const uint32_t wpb = 4;
        quantise_map_kernel<<<(n_rows + wpb - 1) / wpb, wpb * 32, (size_t)wpb * 2 * kk * 4, s>>>(
            ts, td, tc, n_rows, kk, vocab_len, token_to_term, doc_mode, (unsigned long long*)out_indptr, out_terms, out_weights);
        count_launch();
        if (cudaGetLastError() != cudaSuccess) { set_error("msr_quantise_topk: launch failed"); rc = MSR_ERR_CUDA; }
    }
    return rc;
}

extern "C" int msr_quantise_candidates(int device, uint32_t n_rows, uint32_t vocab, const void* logits, int dtype,
                                       size_t row_stride, const uint64_t* cand_indptr_host, const int32_t* cand_tokens_host,
                                       uint32_t cap, uint32_t fallback_top, uint32_t vocab_len, const int32_t* token_to_term,
                                       int doc_mode, const uint8_t* expand_mask, uint32_t n_expand, uint64_t* out_indptr,
                                       uint32_t* out_terms, uint32_t* out_weights, void* stream) {
    MSR_REQUIRE(dtype >= 0 && dtype <= 2, "msr_quantise_candidates: dtype %d outside {0 f32, 1 f16, 2 bf16}", dtype);
    MSR_REQUIRE(cap >= 1 && cap <= (uint32_t)kQuantMaxTop && fallback_top >= 1 && fallback_top <= (uint32_t)kQuantMaxTop,
                "msr_quantise_candidates: cap=%u / fallback_top=%u outside [1,%d]", cap, fallback_top, kQuantMaxTop);
    MSR_REQUIRE(vocab >= fallback_top && row_stride >= vocab, "msr_quantise_candidates: vocab (%u) must be >= fallback_top and row_stride >= vocab", vocab);
    MSR_REQUIRE(out_indptr, "msr_quantise_candidates: NULL out_indptr");
    int rc = use_device(device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (n_rows == 0) {
        MSR_CUDA_CHECK(cudaMemsetAsync(out_indptr, 0, 8, s));
        return MSR_OK;
    }
    MSR_REQUIRE(logits && out_terms && out_weights && cand_indptr_host, "msr_quantise_candidates: NULL buffer");
    MSR_REQUIRE(cand_indptr_host[0] == 0, "msr_quantise_candidates: cand_indptr[0] != 0");
    const uint64_t n_c = cand_indptr_host[n_rows];
    MSR_REQUIRE(n_c == 0 || cand_tokens_host, "msr_quantise_candidates: NULL cand_tokens");
    std::vector<uint32_t> empty_rows;
    for (uint32_t r = 0; r < n_rows; ++r) {
        MSR_REQUIRE(cand_indptr_host[r + 1] >= cand_indptr_host[r], "msr_quantise_candidates: cand_indptr not monotone at row %u", r);
        const uint64_t m = cand_indptr_host[r + 1] - cand_indptr_host[r];
        MSR_REQUIRE(m <= (uint64_t)kCandMax, "msr_quantise_candidates: row %u has %llu candidates (max %d)", r, (unsigned long long)m, kCandMax);
        if (m == 0) empty_rows.push_back(r);
    }
    for (uint64_t i = 0; i < n_c; ++i)
        MSR_REQUIRE(cand_tokens_host[i] >= 0 && (uint32_t)cand_tokens_host[i] < vocab,
                    "msr_quantise_candidates: candidate token id %d outside [0,%u)", cand_tokens_host[i], vocab);
    MSR_REQUIRE(n_expand == 0 || expand_mask, "msr_quantise_candidates: n_expand > 0 needs expand_mask");
    MSR_REQUIRE(n_expand <= (uint32_t)kQuantMaxTop && n_expand <= vocab, "msr_quantise_candidates: n_expand=%u above min(%d, vocab)", n_expand, kQuantMaxTop);
    const uint32_t pitch = std::max(cap, fallback_top) + n_expand;
    AsyncScratch scratch(s);
    float* ts = nullptr; int32_t* td = nullptr; uint32_t* tc = nullptr; unsigned long long* ci = nullptr; int32_t* ct = nullptr;
    float* f32 = nullptr;
    MSR_CUDA_CHECK(scratch.alloc(&ts, (size_t)n_rows * pitch * 4));
    MSR_CUDA_CHECK(scratch.alloc(&td, (size_t)n_rows * pitch * 4));
    MSR_CUDA_CHECK(scratch.alloc(&tc, (size_t)n_rows * 4));
    MSR_CUDA_CHECK(scratch.alloc(&ci, (size_t)(n_rows + 1) * 8));
    MSR_CUDA_CHECK(scratch.alloc(&ct, (size_t)std::max<uint64_t>(n_c, 1) * 4));
    MSR_CUDA_CHECK(cudaMemcpyAsync(ci, cand_indptr_host, (size_t)(n_rows + 1) * 8, cudaMemcpyHostToDevice, s));
    if (n_c) MSR_CUDA_CHECK(cudaMemcpyAsync(ct, cand_tokens_host, (size_t)n_c * 4, cudaMemcpyHostToDevice, s));
    if (dtype == 0) cand_select_kernel<float><<<n_rows, 128, 0, s>>>((const float*)logits, row_stride, ci, ct, cap, pitch, ts, td, tc);
    else if (dtype == 1) cand_select_kernel<__half><<<n_rows, 128, 0, s>>>((const __half*)logits, row_stride, ci, ct, cap, pitch, ts, td, tc);
    else cand_select_kernel<__nv_bfloat16><<<n_rows, 128, 0, s>>>((const __nv_bfloat16*)logits, row_stride, ci, ct, cap, pitch, ts, td, tc);
    count_launch();
    // rows without a candidate (encode.py:103-112): top-`fallback_top` of the whole vocabulary
    if (!empty_rows.empty() && dtype != 0) MSR_CUDA_CHECK(scratch.alloc(&f32, (size_t)vocab * 4));
    for (uint32_t r : empty_rows) {
        const float* row = (const float*)logits + (size_t)r * row_stride;
        if (dtype == 1) { rows_to_f32_kernel<__half><<<148, 256, 0, s>>>((const __half*)logits + (size_t)r * row_stride, row_stride, 1, vocab, f32); row = f32; count_launch(); }
        if (dtype == 2) { rows_to_f32_kernel<__nv_bfloat16><<<148, 256, 0, s>>>((const __nv_bfloat16*)logits + (size_t)r * row_stride, row_stride, 1, vocab, f32); row = f32; count_launch(); }
        rc = topk_rows_launch(1, vocab, row, vocab, fallback_top, 0, ts + (size_t)r * pitch, td + (size_t)r * pitch, tc + r, s);
        if (rc) break;
    }
    if (rc == MSR_OK && n_expand) {
        // expansion terms, a block of rows at a time (scratch: masked fp32 rows, <= 256 MB)
        const uint32_t rows_per = (uint32_t)std::max<size_t>(1, std::min<size_t>(n_rows, ((size_t)256 << 20) / ((size_t)vocab * 4)));
        float* masked = nullptr; float* es = nullptr; int32_t* ed = nullptr; uint32_t* ec = nullptr;
        MSR_CUDA_CHECK(scratch.alloc(&masked, (size_t)rows_per * vocab * 4));
        MSR_CUDA_CHECK(scratch.alloc(&es, (size_t)rows_per * n_expand * 4));
        MSR_CUDA_CHECK(scratch.alloc(&ed, (size_t)rows_per * n_expand * 4));
        MSR_CUDA_CHECK(scratch.alloc(&ec, (size_t)rows_per * 4));
        for (uint32_t r0 = 0; r0 < n_rows && rc == MSR_OK; r0 += rows_per) {
            const uint32_t nr = std::min(rows_per, n_rows - r0);
            const unsigned g = (unsigned)std::min<size_t>(((size_t)nr * vocab + 255) / 256, 148 * 32);
            if (dtype == 0) expand_mask_rows_kernel<float><<<g, 256, 0, s>>>((const float*)logits, row_stride, r0, nr, vocab, expand_mask, masked);
            else if (dtype == 1) expand_mask_rows_kernel<__half><<<g, 256, 0, s>>>((const __half*)logits, row_stride, r0, nr, vocab, expand_mask, masked);
            else expand_mask_rows_kernel<__nv_bfloat16><<<g, 256, 0, s>>>((const __nv_bfloat16*)logits, row_stride, r0, nr, vocab, expand_mask, masked);
            expand_exclude_kernel<<<nr, 128, 0, s>>>(ci, ct, r0, vocab, masked);
            count_launch(2);
            rc = topk_rows_launch(nr, vocab, masked, vocab, n_expand, 0, es, ed, ec, s);
            if (rc) break;
            expand_append_kernel<<<(nr + 127) / 128, 128, 0, s>>>(es, ed, ec, n_expand, r0, nr, pitch, ts, td, tc);
            count_launch();
        }
    }
    if (rc == MSR_OK) rc = map_kernel_smem((size_t)4 * 2 * pitch * 4);
    if (rc == MSR_OK) {
        const uint32_t wpb = 4;
        quantise_map_kernel<<<(n_rows + wpb - 1) / wpb, wpb * 32, (size_t)wpb * 2 * pitch * 4, s>>>(
            ts, td, tc, n_rows, pitch, vocab_len, token_to_term, doc_mode, (unsigned long long*)out_indptr, out_terms, out_weights);
        count_launch();
        if (cudaGetLastError() != cudaSuccess) { set_error("msr_quantise_candidates: launch failed"); rc = MSR_ERR_CUDA; }
    }
    // the host arrays were read by an async copy from pageable memory; make the call safe to return from
    cudaStreamSynchronize(s);
    return rc;
}
