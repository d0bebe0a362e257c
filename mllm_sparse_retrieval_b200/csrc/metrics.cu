// metrics.cu -- Recall@k hit counting on the device.
//
// Restates RecallMetrics._sort / _count of the reference (src/metrices.py:29-35, 76-84): for every query whose run is
// not empty, the first k entries of its ranked list are tested against the query's target set with torch.isin, and
// counts[k] is incremented when ANY target is among them (a hit metric, not |top-k ∩ targets| / |targets|); the
// reference then divides by the number of queries (metrices.py:86-100).  Here one warp handles one query: it finds
// the best rank at which a target appears and bumps every cutoff above it.
#include "common.cuh"

namespace msr {

constexpr int kRecallMaxCut = 32;

__global__ void __launch_bounds__(256)
recall_hits_kernel(const int32_t* __restrict__ docs, size_t row_stride, const uint32_t* __restrict__ counts, uint32_t nq,
                   const uint64_t* __restrict__ t_indptr, const int32_t* __restrict__ t_docs,
                   const uint32_t* __restrict__ cutoffs, uint32_t n_cut, unsigned long long* __restrict__ hits,
                   unsigned long long* __restrict__ n_eval) {
    const uint32_t q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (q >= nq) return;
    const uint32_t n = counts[q];
    if (n == 0) return;                                         // `if len(v['docs']) == 0: continue`
    const uint64_t t0 = t_indptr[q], t1 = t_indptr[q + 1];
    const int32_t* row = docs + (size_t)q * row_stride;
    uint32_t best = 0xFFFFFFFFu;
    for (uint32_t j0 = 0; j0 < n && best == 0xFFFFFFFFu; j0 += 32) {
        const uint32_t j = j0 + lane;
        bool hit = false;
        if (j < n) {
            const int32_t d = row[j];
            for (uint64_t t = t0; t < t1; ++t) hit |= (t_docs[t] == d);
        }
        const uint32_t b = __ballot_sync(0xffffffffu, hit);
        if (b) best = j0 + (uint32_t)__ffs(b) - 1;
    }
    if (lane == 0) atomicAdd(n_eval, 1ull);
    if (lane < n_cut && best < cutoffs[lane]) atomicAdd(hits + lane, 1ull);
}


// Provenance histogram of score_statistic.py (/root/reference/src/score_statistic.py:406-417): per query the fused
// list is ordered by score descending (Python's stable sort: equal scores keep insertion order), cut at `depth`, and
// the 1-based rank of every entry is appended to the list of its tag (dense / sparse / fuse, hybrid.py:76-85).  The
// three rank lists are only ever binned, so the device result is hist[tag-1][rank-1] summed over the batch.
// One CTA per query; rank by counting over the row (<= 2*MSR_MAX_K entries, read through L1).
__global__ void __launch_bounds__(256)
provenance_hist_kernel(const double* __restrict__ scores, const uint8_t* __restrict__ src, size_t row_stride,
                       const uint32_t* __restrict__ counts, uint32_t depth, unsigned long long* __restrict__ hist) {
    const uint32_t q = blockIdx.x;
    const uint32_t n = counts[q];
    const double* s = scores + (size_t)q * row_stride;
    const uint8_t* t = src + (size_t)q * row_stride;
    for (uint32_t j = threadIdx.x; j < n; j += blockDim.x) {
        const double a = s[j];
        uint32_t r = 0;
        for (uint32_t i = 0; i < n; ++i) {
            const double b = __ldg(s + i);
            r += (b > a) || (b == a && i < j);
        }
        const uint32_t tag = t[j];
        if (r < depth && tag >= 1 && tag <= 3) atomicAdd(hist + (size_t)(tag - 1) * depth + r, 1ull);
    }
}

// Final ordering of run rows (the CLI's last step): what Python's stable `sorted(..., key=score, reverse=True)` does in
// metrices.py:30 / hybrid.write_trec_run, and what get_run_dict does when it drops the query's own id (search.py:66-82),
// as ONE kernel.  One CTA per row: an entry is valid when it lies inside the row's count and is not the row's `self` doc;
// its rank is the number of valid entries that sort before it -- by (score desc, position asc) when by_score is set, by
// position otherwise -- counted against the row held in shared memory (rows have at most 2 * MSR_MAX_K entries).
template <typename T>
__device__ __forceinline__ uint64_t order_key(T x);
template <>
__device__ __forceinline__ uint64_t order_key<float>(float x) { return (uint64_t)f2ord(x + 0.0f) + 1ull; }   // -0.0 == 0.0, as in Python
template <>
__device__ __forceinline__ uint64_t order_key<double>(double x) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(x + 0.0);   // -0.0 == 0.0, as in Python
    const unsigned long long o = b ^ ((b >> 63) ? ~0ull : 0x8000000000000000ull);
    return o == ~0ull ? o : o + 1ull;                           // 0 is reserved for "not valid"
}

template <typename T>
__global__ void __launch_bounds__(256)
order_rows_kernel(const T* __restrict__ scores, const int32_t* __restrict__ docs, const uint8_t* __restrict__ tags, size_t row_stride,
                  const uint32_t* __restrict__ counts, const int32_t* __restrict__ self_ids, uint32_t width, int by_score,
                  T* __restrict__ out_scores, int32_t* __restrict__ out_docs, uint8_t* __restrict__ out_tags,
                  uint32_t* __restrict__ out_counts) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* key = reinterpret_cast<uint64_t*>(smem_raw);
    __shared__ uint32_t s_n;
    const uint32_t q = blockIdx.x, tid = threadIdx.x;
    const uint32_t n = min(counts[q], width);
    const int32_t self = self_ids ? self_ids[q] : -1;
    const T* srow = scores + (size_t)q * row_stride;
    const int32_t* drow = docs + (size_t)q * row_stride;
    if (tid == 0) s_n = 0;
    __syncthreads();
    uint32_t mine = 0;
    for (uint32_t i = tid; i < n; i += 256) {
        const bool valid = !(self >= 0 && drow[i] == self);
        key[i] = valid ? (by_score ? order_key<T>(srow[i]) : 1ull) : 0ull;
        mine += valid;
    }
    mine = warp_reduce_add(mine);
    if ((tid & 31) == 0 && mine) atomicAdd(&s_n, mine);
    __syncthreads();
    for (uint32_t i = tid; i < n; i += 256) {
        const uint64_t k = key[i];
        if (!k) continue;
        uint32_t rank = 0;
        for (uint32_t j = 0; j < n; ++j) {
            const uint64_t o = key[j];
            rank += (o > k) || (o == k && j < i);
        }
        out_scores[(size_t)q * row_stride + rank] = srow[i];
        out_docs[(size_t)q * row_stride + rank] = drow[i];
        if (tags) out_tags[(size_t)q * row_stride + rank] = tags[(size_t)q * row_stride + i];
    }
    if (tid == 0) out_counts[q] = s_n;
}

}  // namespace msr

using namespace msr;

extern "C" int msr_recall_hits(int device, uint32_t nq, const int32_t* docs, size_t row_stride, const uint32_t* counts,
                               const uint64_t* t_indptr, const int32_t* t_docs, uint32_t n_cut, const uint32_t* cutoffs,
                               uint64_t* hits, uint64_t* n_eval, void* stream) {
    MSR_REQUIRE(n_cut >= 1 && n_cut <= (uint32_t)kRecallMaxCut, "msr_recall_hits: n_cut=%u outside [1,%d]", n_cut, kRecallMaxCut);
    MSR_REQUIRE(hits && n_eval && cutoffs, "msr_recall_hits: NULL output / cutoffs");
    int rc = use_device(device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    MSR_CUDA_CHECK(cudaMemsetAsync(hits, 0, (size_t)n_cut * 8, s));
    MSR_CUDA_CHECK(cudaMemsetAsync(n_eval, 0, 8, s));
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(docs && counts && t_indptr, "msr_recall_hits: NULL input");
    recall_hits_kernel<<<(nq + 7) / 8, 256, 0, s>>>(docs, row_stride, counts, nq, t_indptr, t_docs, cutoffs, n_cut,
                                                    (unsigned long long*)hits, (unsigned long long*)n_eval);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    return MSR_OK;
}

extern "C" int msr_provenance_hist(int device, uint32_t nq, const double* scores, const uint8_t* src, size_t row_stride,
                                   const uint32_t* counts, uint32_t depth, uint64_t* hist, void* stream) {
    MSR_REQUIRE(depth >= 1 && depth <= 2 * MSR_MAX_K, "msr_provenance_hist: depth=%u outside [1,%u]", depth, 2 * MSR_MAX_K);
    MSR_REQUIRE(hist, "msr_provenance_hist: NULL hist");
    int rc = use_device(device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    MSR_CUDA_CHECK(cudaMemsetAsync(hist, 0, (size_t)3 * depth * 8, s));
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(scores && src && counts, "msr_provenance_hist: NULL input");
    provenance_hist_kernel<<<nq, 256, 0, s>>>(scores, src, row_stride, counts, depth, (unsigned long long*)hist);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    return MSR_OK;
}

extern "C" int msr_order_rows(int device, uint32_t nq, uint32_t width, const void* scores, int scores_f64, const int32_t* docs,
                              const uint8_t* tags, size_t row_stride, const uint32_t* counts, const int32_t* self_ids,
                              int by_score, void* out_scores, int32_t* out_docs, uint8_t* out_tags, uint32_t* out_counts,
                              void* stream) {
    MSR_REQUIRE(width >= 1 && width <= 2 * MSR_MAX_K, "msr_order_rows: width %u outside [1,%u]", width, 2 * MSR_MAX_K);
    MSR_REQUIRE(row_stride >= width, "msr_order_rows: row_stride < width");
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(scores && docs && counts && out_scores && out_docs && out_counts, "msr_order_rows: NULL buffer");
    MSR_REQUIRE(scores != out_scores && docs != out_docs, "msr_order_rows: the permutation is not done in place");
    MSR_REQUIRE(!tags || out_tags, "msr_order_rows: tags without out_tags");
    int rc = use_device(device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t smem = (size_t)width * 8;
    static std::atomic<uint64_t> attr_set{0};
    if (attr_needed(attr_set)) {
        MSR_CUDA_CHECK(cudaFuncSetAttribute(order_rows_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * MSR_MAX_K * 8));
        MSR_CUDA_CHECK(cudaFuncSetAttribute(order_rows_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * MSR_MAX_K * 8));
        attr_mark(attr_set);
    }
    if (scores_f64)
        order_rows_kernel<double><<<nq, 256, smem, s>>>((const double*)scores, docs, tags, row_stride, counts, self_ids, width, by_score,
                                                        (double*)out_scores, out_docs, out_tags, out_counts);
    else
        order_rows_kernel<float><<<nq, 256, smem, s>>>((const float*)scores, docs, tags, row_stride, counts, self_ids, width, by_score,
                                                       (float*)out_scores, out_docs, out_tags, out_counts);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    return MSR_OK;
}
