/*
 * oracle/impact_tuned.c -- TEST INFRASTRUCTURE ONLY (timed CPU baselines; never on the product path).
 *
 * Two TUNED CPU implementations of the same Lucene-impact search that oracle/impact_oracle.c restates naively
 * (rules S1-S10 of SURVEY.md 8(a); reference call sites /root/reference/src/search.py:85-99, 273-275).  They exist
 * because the exhaustive double-accumulator port is DRAM-latency-bound and does not scale with cores, which makes it
 * a soft target for a GPU/CPU ratio.  Both must return the oracle's lists bit for bit (tests/test_oracle_cpu.py).
 *
 *   mode 0  "port-exhaustive-blocked": int32 accumulators (what BASELINE.md section 3 promised), the doc-id space
 *           walked in blocks of TUNED_BLOCK docs so that the accumulator tile stays in the core's L2/L1; every
 *           posting of every kept term is applied (term-at-a-time inside a block), heap top-k per block.
 *   mode 1  "port-pruned" (MaxScore): the closest thing to what Lucene >= 8 does for this BooleanQuery of SHOULD
 *           term clauses under a top-k collector.  Terms are ordered by their score upper bound w_t * max_impact(t);
 *           the longest prefix whose bounds sum to <= the current k-th score is NON-ESSENTIAL: a document that
 *           matches only such terms cannot enter the top k (later docs lose score ties, S8), so those posting lists
 *           are not traversed -- they are only probed (galloping search) for the documents the essential terms
 *           produced, most promising term first, stopping as soon as the remaining bounds cannot lift the doc over
 *           the threshold.  Exact: no approximation, same (score desc, doc asc) order.
 *
 * Integer arithmetic is exact, and equal to S6's float32(sum of float32 products), while every product is below
 * 2^24 and the sum below 2^31; a query outside that range is answered by the wide (double) path below, which is the
 * oracle's loop restated.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TUNED_BLOCK 32768u     /* default accumulator tile: 128 KB of int32 */

typedef struct { float score; uint32_t doc; } thit_t;

static inline int thit_before(thit_t a, thit_t b) {
    return (a.score > b.score) || (a.score == b.score && a.doc < b.doc);
}
static void theap_down(thit_t* h, uint32_t n, uint32_t i) {
    for (;;) {
        uint32_t l = 2 * i + 1, r = l + 1, w = i;
        if (l < n && thit_before(h[w], h[l])) w = l;
        if (r < n && thit_before(h[w], h[r])) w = r;
        if (w == i) return;
        thit_t t = h[i]; h[i] = h[w]; h[w] = t;
        i = w;
    }
}
static int tcmp(const void* pa, const void* pb) {
    thit_t a = *(const thit_t*)pa, b = *(const thit_t*)pb;
    return thit_before(a, b) ? -1 : (thit_before(b, a) ? 1 : 0);
}
/* offer a hit (docs arrive in ascending id order, so a tie with the root loses) */
static inline void offer(thit_t* heap, uint32_t* n, uint32_t k, float score, uint32_t doc) {
    if (*n < k) {
        heap[*n].score = score; heap[*n].doc = doc;
        if (++*n == k) for (int64_t i = (int64_t)k / 2 - 1; i >= 0; --i) theap_down(heap, k, (uint32_t)i);
    } else if (score > heap[0].score) {
        heap[0].score = score; heap[0].doc = doc;
        theap_down(heap, k, 0);
    }
}

typedef struct {
    uint64_t cur, end;      /* posting cursor / end of the list */
    uint32_t w;             /* query weight (duplicates already merged) */
    uint64_t ub;            /* w * max_impact(t) */
} qterm_t;

static int cmp_ub(const void* a, const void* b) {
    uint64_t x = ((const qterm_t*)a)->ub, y = ((const qterm_t*)b)->ub;
    return x < y ? -1 : (x > y ? 1 : 0);
}

/* first position p in [lo, end) with docids[p] >= d (galloping from lo) */
static inline uint64_t gallop(const uint32_t* docids, uint64_t lo, uint64_t end, uint32_t d) {
    if (lo >= end || docids[lo] >= d) return lo;
    uint64_t step = 1, hi = lo + 1;
    while (hi < end && docids[hi] < d) { lo = hi; step <<= 1; hi = lo + step; }
    if (hi > end) hi = end;
    /* docids[lo] < d, (hi == end or docids[hi] >= d) */
    while (hi - lo > 1) {
        uint64_t mid = lo + ((hi - lo) >> 1);
        if (docids[mid] < d) lo = mid; else hi = mid;
    }
    return hi;
}

/* wide path: one query exactly as impact_oracle.c does it (double accumulators, float32 per-term products) */
static void wide_query(uint32_t n_docs, const uint32_t* docids, const uint16_t* impacts, const qterm_t* qt, uint32_t nt,
                       double* wacc, thit_t* heap, uint32_t* n, uint32_t k) {
    for (uint32_t i = 0; i < nt; ++i)
        for (uint64_t p = qt[i].cur; p < qt[i].end; ++p) wacc[docids[p]] += (double)((float)qt[i].w * (float)impacts[p]);
    for (uint32_t d = 0; d < n_docs; ++d) {
        if (wacc[d] != 0.0) {
            float s = (float)wacc[d];
            wacc[d] = 0.0;
            if (s > 0.0f) offer(heap, n, k, s, d);
        }
    }
}

int tuned_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/*
 * Same contract as oracle_impact_search (impact_oracle.c) plus:
 *   max_impact[n_terms]  per-term maximum impact (index-time statistic, as Lucene stores impacts per block)
 *   mode                 0 = exhaustive blocked, 1 = MaxScore-pruned
 *   block_docs           accumulator tile in docs (0 = TUNED_BLOCK)
 *   postings_touched     (optional) number of postings read, summed over the queries: shows what the pruning skips
 */
int tuned_impact_search(uint32_t n_docs, uint32_t n_terms,
                        const uint64_t* indptr, const uint32_t* docids, const uint16_t* impacts,
                        const uint16_t* max_impact,
                        uint32_t n_docs_global, const uint32_t* df_global, int32_t doc_base,
                        uint32_t nq, const uint64_t* q_indptr, const uint32_t* q_terms, const uint32_t* q_weights,
                        uint32_t k, float* out_scores, int32_t* out_docs, uint32_t* out_counts,
                        int mode, int threads, uint32_t block_docs, uint64_t* postings_touched) {
    int fail = 0;
    const uint32_t B = block_docs ? block_docs : TUNED_BLOCK;   /* tests use tiny blocks to exercise the pruning */
    uint64_t touched_total = 0;
    if (threads < 1) threads = 1;
#ifdef _OPENMP
#pragma omp parallel num_threads(threads) reduction(+ : touched_total)
#endif
    {
        int32_t* acc = (int32_t*)calloc(B, sizeof(int32_t));
        uint32_t* cand = (uint32_t*)malloc(sizeof(uint32_t) * B);
        thit_t* heap = (thit_t*)malloc(sizeof(thit_t) * (k ? k : 1));
        uint32_t qcap = 64;
        qterm_t* qt = (qterm_t*)malloc(sizeof(qterm_t) * qcap);
        double* wacc = NULL;
        if (!acc || !cand || !heap || !qt) {
#ifdef _OPENMP
#pragma omp atomic write
#endif
            fail = 1;
        } else {
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
            for (int64_t q = 0; q < (int64_t)nq; ++q) {
                if (fail) continue;
                const uint64_t qs = q_indptr[q], qe = q_indptr[q + 1];
                if (qe - qs > qcap) {
                    qcap = (uint32_t)(qe - qs) * 2;
                    free(qt);
                    qt = (qterm_t*)malloc(sizeof(qterm_t) * qcap);
                    if (!qt) { fail = 1; continue; }
                }
                /* kept terms (S5) */
                uint32_t nt = 0;
                uint64_t ub_sum = 0, ub_max = 0;
                for (uint64_t j = qs; j < qe; ++j) {
                    const uint32_t t = q_terms[j], w = q_weights[j];
                    if (t >= n_terms || w == 0) continue;
                    const uint64_t df = df_global ? df_global[t] : (indptr[t + 1] - indptr[t]);
                    if (df == 0 || !(log((double)n_docs_global / (double)(df + 1)) > 0.0)) continue;
                    if (indptr[t] == indptr[t + 1]) continue;
                    qt[nt].cur = indptr[t]; qt[nt].end = indptr[t + 1]; qt[nt].w = w;
                    qt[nt].ub = (uint64_t)w * max_impact[t];
                    if (qt[nt].ub > ub_max) ub_max = qt[nt].ub;
                    ub_sum += qt[nt].ub;
                    ++nt;
                }
                uint32_t n = 0;
                if (k == 0) { out_counts[q] = 0; continue; }
                if (ub_max >= (1ull << 24) || ub_sum >= (1ull << 31)) {
                    /* outside the exact integer range (S6 rounds per term): wide path */
                    if (!wacc) wacc = (double*)calloc(n_docs ? n_docs : 1, sizeof(double));
                    if (!wacc) { fail = 1; continue; }
                    for (uint32_t i = 0; i < nt; ++i) touched_total += qt[i].end - qt[i].cur;
                    wide_query(n_docs, docids, impacts, qt, nt, wacc, heap, &n, k);
                } else {
                    if (mode == 1) qsort(qt, nt, sizeof(qterm_t), cmp_ub);   /* ascending upper bound */
                    uint32_t n_ne = 0;                                       /* qt[0..n_ne) are non-essential */
                    uint64_t ne_ub = 0;                                      /* sum of their bounds */
                    for (uint32_t lo = 0; lo < n_docs; lo += B) {
                        const uint32_t hi = (n_docs - lo < B) ? n_docs : lo + B;
                        const uint32_t bl = hi - lo;
                        if (mode == 1 && n == k) {
                            /* grow the non-essential prefix: bounds summing to <= the k-th score cannot lift a doc in */
                            const uint64_t theta = (uint64_t)heap[0].score;  /* integer-valued float32 below 2^31 */
                            while (n_ne < nt && ne_ub + qt[n_ne].ub <= theta) ne_ub += qt[n_ne++].ub;
                        }
                        /* essential terms: apply every posting of the block */
                        for (uint32_t i = n_ne; i < nt; ++i) {
                            uint64_t p = qt[i].cur;
                            const uint64_t e = qt[i].end;
                            const int32_t w = (int32_t)qt[i].w;
                            const uint64_t p0 = p;
                            while (p < e && docids[p] < hi) { acc[docids[p] - lo] += w * (int32_t)impacts[p]; ++p; }
                            touched_total += p - p0;
                            qt[i].cur = p;
                        }
                        if (n_ne == 0) {
                            for (uint32_t d = 0; d < bl; ++d) {
                                const int32_t a = acc[d];
                                if (a) { acc[d] = 0; offer(heap, &n, k, (float)a, lo + d); }
                            }
                            continue;
                        }
                        /* candidates: docs the essential terms reached that could still beat the threshold */
                        uint32_t nc = 0;
                        {
                            const int64_t need = (int64_t)heap[0].score - (int64_t)ne_ub;   /* a + ne_ub > theta */
                            for (uint32_t d = 0; d < bl; ++d) {
                                const int32_t a = acc[d];
                                if (a) { if ((int64_t)a > need) cand[nc++] = d; else acc[d] = 0; }
                            }
                        }
                        /* complete them from the non-essential lists, largest bound first, with early exit */
                        for (uint32_t i = n_ne; i-- > 0 && nc;) {
                            uint64_t rest = 0;                               /* bounds of the terms still to come */
                            for (uint32_t j = 0; j < i; ++j) rest += qt[j].ub;
                            const int32_t w = (int32_t)qt[i].w;
                            const int64_t thr = (int64_t)heap[0].score;
                            uint64_t p = qt[i].cur;
                            const uint64_t e = qt[i].end;
                            const uint64_t pe = gallop(docids, p, e, hi);        /* end of the list inside this block */
                            uint32_t keep = 0;
                            for (uint32_t c = 0; c < nc; ++c) {                   /* early exit: bounds cannot lift it */
                                const uint32_t d = cand[c];
                                if ((int64_t)acc[d] + (int64_t)qt[i].ub + (int64_t)rest <= thr) { acc[d] = 0; continue; }
                                cand[keep++] = d;
                            }
                            nc = keep;
                            if ((uint64_t)nc * 8 >= pe - p) {
                                /* many survivors: walking the block's postings is cheaper than probing for each doc;
                                   only live candidates (acc != 0) take the contribution */
                                for (uint64_t x = p; x < pe; ++x) {
                                    const uint32_t d = docids[x] - lo;
                                    if (acc[d]) acc[d] += w * (int32_t)impacts[x];
                                }
                                touched_total += pe - p;
                            } else {
                                for (uint32_t c = 0; c < nc; ++c) {
                                    const uint32_t d = cand[c];
                                    p = gallop(docids, p, pe, lo + d);
                                    ++touched_total;
                                    if (p < pe && docids[p] == lo + d) acc[d] += w * (int32_t)impacts[p];
                                }
                            }
                            qt[i].cur = pe;
                        }
                        for (uint32_t c = 0; c < nc; ++c) {
                            const uint32_t d = cand[c];
                            const int32_t a = acc[d];
                            acc[d] = 0;
                            offer(heap, &n, k, (float)a, lo + d);
                        }
                        /* cursors of non-essential terms that the loop above did not visit (nc hit 0) */
                        for (uint32_t i = 0; i < n_ne; ++i) qt[i].cur = gallop(docids, qt[i].cur, qt[i].end, hi);
                    }
                }
                qsort(heap, n, sizeof(thit_t), tcmp);
                for (uint32_t i = 0; i < k; ++i) {
                    out_scores[(size_t)q * k + i] = i < n ? heap[i].score : 0.0f;
                    out_docs[(size_t)q * k + i] = i < n ? (int32_t)heap[i].doc + doc_base : -1;
                }
                out_counts[q] = n;
            }
        }
        free(acc); free(cand); free(heap); free(qt); free(wacc);
    }
    if (postings_touched) *postings_touched = touched_total;
    return fail ? -1 : 0;
}
