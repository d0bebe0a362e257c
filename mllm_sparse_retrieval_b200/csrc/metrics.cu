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
