/*
 * oracle/impact_oracle.c -- TEST INFRASTRUCTURE ONLY (CPU restatement; never on the product path).
 *
 * Plain-C restatement of the reference's sparse search arithmetic: the Lucene impact search that
 * /root/reference/src/search.py:85-99 (`sparse_search`) reaches through
 * pyserini.LuceneImpactSearcher.batch_search (call sites search.py:273-275, 86-87).  The engine itself
 * (pyserini -> Anserini SimpleImpactSearcher -> Lucene) is a third-party dependency that is NOT under
 * /root/reference and whose version the reference does not pin (no requirements file); its published
 * behaviour is restated here as rules S1-S10 of SURVEY.md section 8(a):
 *
 *   S1  tf(t,d) = vector[t] (integer impact >= 1); norms disabled          (JsonVectorCollection + --impact)
 *   S5  a query term is kept iff it is in the index dictionary and ln(N/(df+1)) > 0, i.e. df <= N-2
 *   S6  score = float32( sum_t float32(q_t * tf_t,d) ), term products rounded to float32 individually
 *       (ImpactSimilarity returns boost*freq as float), summed in double, cast to float32
 *   S7  only documents matching >=1 kept term are hits (score > 0); at most k hits
 *   S8  order: score descending, ties by external docid string ascending.  Internal doc ids handed to this
 *       file are already the rank of the docid string in UTF-8 byte order, so "smaller id wins".
 *   S10 one query per worker thread (Anserini thread pool) -> OpenMP parallel-for over queries.
 *
 * PARITY UNPINNED: the reference holds no test, golden vector or fixture for this path and
 * the real engine cannot run in the build container (no JVM).  Known-answer vectors in tests/golden are
 * hand-derived from S1-S10.
 *
 * Besides being the checker, this file is the "port" CPU baseline timed by bench.py (cpu_baseline.kind="port").
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { float score; uint32_t doc; } hit_t;

/* "a ranks before b": score desc, doc asc (S8) */
static inline int hit_before(hit_t a, hit_t b) {
    return (a.score > b.score) || (a.score == b.score && a.doc < b.doc);
}

/* min-heap on ranking order: root = the WORST of the kept hits */
static void heap_sift_down(hit_t* h, uint32_t n, uint32_t i) {
    for (;;) {
        uint32_t l = 2 * i + 1, r = l + 1, w = i;
        if (l < n && hit_before(h[w], h[l])) w = l;
        if (r < n && hit_before(h[w], h[r])) w = r;
        if (w == i) return;
        hit_t t = h[i]; h[i] = h[w]; h[w] = t;
        i = w;
    }
}

static int cmp_hit(const void* pa, const void* pb) {
    hit_t a = *(const hit_t*)pa, b = *(const hit_t*)pb;
    if (hit_before(a, b)) return -1;
    if (hit_before(b, a)) return 1;
    return 0;
}

int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/*
 * Canonical term-major CSR: postings of term t are docids[indptr[t]..indptr[t+1]) (ascending) with impacts[].
 * n_docs_global / df_global implement S5 with collection-wide statistics when the CSR is one shard of a
 * doc-sharded collection (df_global == NULL -> df from indptr, N = n_docs).
 * Query q: terms q_terms[q_indptr[q]..q_indptr[q+1]), weights q_weights[] (>=1; duplicates allowed, they add).
 * Outputs are [nq,k] row-major; row q holds out_counts[q] hits then padding (score 0, doc -1).
 * doc ids written are local_doc + doc_base.
 * Returns 0, or -1 on allocation failure.
 */
int oracle_impact_search(uint32_t n_docs, uint32_t n_terms,
                         const uint64_t* indptr, const uint32_t* docids, const uint16_t* impacts,
                         uint32_t n_docs_global, const uint32_t* df_global, int32_t doc_base,
                         uint32_t nq, const uint64_t* q_indptr, const uint32_t* q_terms, const uint32_t* q_weights,
                         uint32_t k, float* out_scores, int32_t* out_docs, uint32_t* out_counts,
                         int threads) {
    int fail = 0;
    if (threads < 1) threads = 1;
#ifdef _OPENMP
#pragma omp parallel num_threads(threads)
#endif
    {
        double* acc = (double*)calloc(n_docs ? n_docs : 1, sizeof(double));
        uint8_t* touched = (uint8_t*)calloc(n_docs ? n_docs : 1, 1);
        hit_t* heap = (hit_t*)malloc(sizeof(hit_t) * (k ? k : 1));
        if (!acc || !touched || !heap) {
#ifdef _OPENMP
#pragma omp atomic write
#endif
            fail = 1;
        } else {
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
            for (int64_t q = 0; q < (int64_t)nq; ++q) {
                /* term-at-a-time accumulation (S1, S5, S6) */
                for (uint64_t j = q_indptr[q]; j < q_indptr[q + 1]; ++j) {
                    uint32_t t = q_terms[j];
                    uint32_t w = q_weights[j];
                    if (t >= n_terms || w == 0) continue;
                    uint64_t df = df_global ? df_global[t] : (indptr[t + 1] - indptr[t]);
                    if (df == 0) continue;                              /* not in the dictionary */
                    if (!(log((double)n_docs_global / (double)(df + 1)) > 0.0)) continue; /* S5 */
                    for (uint64_t p = indptr[t]; p < indptr[t + 1]; ++p) {
                        float term = (float)w * (float)impacts[p];      /* S6: boost*freq in float32 */
                        acc[docids[p]] += (double)term;
                        touched[docids[p]] = 1;
                    }
                }
                /* top-k by (score desc, doc asc) (S7, S8) */
                uint32_t n = 0;
                for (uint32_t d = 0; d < n_docs; ++d) {
                    if (!touched[d]) continue;
                    hit_t h; h.score = (float)acc[d]; h.doc = d;
                    acc[d] = 0.0; touched[d] = 0;
                    if (k == 0 || !(h.score > 0.0f)) continue;
                    if (n < k) {
                        heap[n++] = h;
                        if (n == k) for (int64_t i = (int64_t)k / 2 - 1; i >= 0; --i) heap_sift_down(heap, k, (uint32_t)i);
                    } else if (hit_before(h, heap[0])) {
                        heap[0] = h;
                        heap_sift_down(heap, k, 0);
                    }
                }
                qsort(heap, n, sizeof(hit_t), cmp_hit);
                for (uint32_t i = 0; i < k; ++i) {
                    out_scores[(size_t)q * k + i] = i < n ? heap[i].score : 0.0f;
                    out_docs[(size_t)q * k + i] = i < n ? (int32_t)heap[i].doc + doc_base : -1;
                }
                out_counts[q] = n;
            }
        }
        free(acc); free(touched); free(heap);
    }
    return fail ? -1 : 0;
}

/*
 * Flat inner-product restatement (D3/D4): tevatron FaissFlatSearcher -> faiss.IndexFlatIP.search
 * (call sites /root/reference/src/search.py:55-59).  Exact scores accumulated in double, rounded once to
 * float32; order = score desc, row id asc (FAISS's own tie order is implementation-defined; Appendix A).
 * k is clamped by the caller to n_docs (D5).
 */
int oracle_flat_ip_search(uint32_t n_docs, uint32_t dim, const float* docs,
                          uint32_t nq, const float* queries, uint32_t k,
                          float* out_scores, int64_t* out_ids, int threads) {
    int fail = 0;
    if (threads < 1) threads = 1;
#ifdef _OPENMP
#pragma omp parallel num_threads(threads)
#endif
    {
        hit_t* heap = (hit_t*)malloc(sizeof(hit_t) * (k ? k : 1));
        if (!heap) {
#ifdef _OPENMP
#pragma omp atomic write
#endif
            fail = 1;
        } else {
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
            for (int64_t q = 0; q < (int64_t)nq; ++q) {
                const float* qv = queries + (size_t)q * dim;
                uint32_t n = 0;
                for (uint32_t d = 0; d < n_docs; ++d) {
                    const float* dv = docs + (size_t)d * dim;
                    double s = 0.0;
                    for (uint32_t i = 0; i < dim; ++i) s += (double)qv[i] * (double)dv[i];
                    hit_t h; h.score = (float)s; h.doc = d;
                    if (k == 0) continue;
                    if (n < k) {
                        heap[n++] = h;
                        if (n == k) for (int64_t i = (int64_t)k / 2 - 1; i >= 0; --i) heap_sift_down(heap, k, (uint32_t)i);
                    } else if (hit_before(h, heap[0])) {
                        heap[0] = h;
                        heap_sift_down(heap, k, 0);
                    }
                }
                qsort(heap, n, sizeof(hit_t), cmp_hit);
                for (uint32_t i = 0; i < k; ++i) {
                    out_scores[(size_t)q * k + i] = i < n ? heap[i].score : -3.4028234663852886e38f;
                    out_ids[(size_t)q * k + i] = i < n ? (int64_t)heap[i].doc : -1;
                }
            }
        }
        free(heap);
    }
    return fail ? -1 : 0;
}
