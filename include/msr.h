/*
 * msr.h -- C ABI of libmsr.so, the B200 (sm_100a) retrieval-scoring library.
 *
 * The reference (cjc20000323/mllm_sparse_retrieval) has no FFI layer of its own: its search step calls two
 * third-party engines and one in-tree pure-Python function.  Each entry point below names the reference call
 * site it replaces; INTEGRATION.md shows the ctypes stub a maintainer would add.
 *
 *   msr_sparse_*   replaces  pyserini LuceneImpactSearcher(<sparse_index>/index).batch_search(...)
 *                            called at /root/reference/src/search.py:273-275 and :86-87 (sparse_search, :85-99);
 *                            index built offline by scripts/sparse_index.sh:12-18.
 *   msr_dense_*    replaces  tevatron FaissFlatSearcher(...).add/search/batch_search
 *                            called at /root/reference/src/search.py:232,237 and :57,59 (search_queries, :55-63),
 *                            incl. the fp16 GPU clone at :254-270.
 *   msr_fuse       replaces  hybrid.fuse  /root/reference/src/hybrid.py:32-53 fed by get_run_dict search.py:66-82
 *                            (call site search.py:455-461).
 *   msr_topk_merge           the per-GPU list merge of the doc-sharded layout (FAISS IndexShards precedent,
 *                            search.py:265-270); inputs are what one NCCL all-gather delivers.
 *
 * Conventions: every function returns 0 (MSR_OK) or a negative msr_status; msr_last_error() returns a
 * thread-local message.  No exceptions cross the boundary.  Handles are opaque, owned by the library, freed
 * by *_destroy, and immutable after build except for an internal workspace: concurrent searches on ONE
 * handle must be serialised by the caller (one handle per GPU per stream).  `stream` is a cudaStream_t passed
 * as void* (NULL = legacy default stream).  Pointers are plain; `where` says whether a call's data pointers
 * are host (MSR_HOST) or device (MSR_DEVICE) memory.  With MSR_DEVICE the call is asynchronous on `stream`;
 * with MSR_HOST the library stages through device memory and returns after the results are in the host buffers.
 * There is NO CPU fallback: every entry point fails with MSR_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef MSR_H_
#define MSR_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    MSR_OK = 0,
    MSR_ERR_INVALID = -1,   /* bad argument */
    MSR_ERR_CUDA = -2,      /* CUDA runtime error / no usable device */
    MSR_ERR_RANGE = -3,     /* value outside the supported exact-arithmetic range (see msr_sparse_search) */
    MSR_ERR_NOMEM = -4
} msr_status;

enum { MSR_HOST = 0, MSR_DEVICE = 1 };

#define MSR_MAX_K 4096u            /* largest k / depth any search or merge accepts */

typedef struct msr_sparse_index msr_sparse_index;
typedef struct msr_dense_index msr_dense_index;

int msr_version(void);
const char* msr_last_error(void);
/* number of kernels launched by this library in this process so far (bench.py's gpu_launches) */
uint64_t msr_kernel_launches(void);

/* ---------------------------------------------------------------------------------------------------------
 * Sparse impact index (kernels (a)+(b)).
 *
 * Input = canonical term-major CSR of ONE shard of the collection: postings of term t are
 * docids[indptr[t] .. indptr[t+1]) strictly ascending LOCAL doc ids in [0,n_docs) with impacts[] >= 1
 * (tf(t,d) = vector[t], rule S1).  Local doc id d stands for global internal id doc_base + d; internal ids are
 * ranks of the external docid strings in UTF-8 byte order, so "smaller id wins" is Anserini's tie-break (S8).
 * n_docs_global / df_global[n_terms] are the collection-wide N and document frequencies used by the query
 * term filter S5 (keep t iff 0 < df and ln(N/(df+1)) > 0); df_global == NULL means "this shard is the
 * whole collection".  tile_docs = 0 picks the default doc-range tile (accumulators of one tile live in shared
 * memory); otherwise 1..53248.
 * The library re-buckets the postings on the device into its tiled layout and keeps no reference to the
 * caller's arrays.  nnz must be < 2^32.
 */
int msr_sparse_index_build(int device, uint32_t n_docs, uint32_t n_terms,
                           const uint64_t* indptr, const uint32_t* docids, const uint16_t* impacts,
                           uint32_t n_docs_global, const uint32_t* df_global, int32_t doc_base,
                           uint32_t tile_docs, int where, void* stream, msr_sparse_index** out);
int msr_sparse_index_destroy(msr_sparse_index* h);
/* n_docs, n_terms, n_tiles, tile_docs, nnz(lo), nnz(hi), device, bytes-resident(lo), bytes-resident(hi) */
int msr_sparse_index_info(const msr_sparse_index* h, uint32_t info[9]);

/*
 * Batched impact search: for query q with terms q_terms[q_indptr[q]..q_indptr[q+1]) and integer weights
 * q_weights[] (duplicates add, weight 0 ignored, term ids >= n_terms ignored):
 *     score(q,d) = float32( sum_t float32(w_t * tf(t,d)) )                       (rule S6)
 * over kept terms (S5); hits are docs with score > 0 (S7), ordered by score desc, internal id asc (S8).
 * Outputs are [nq,k] row-major: row q = out_counts[q] hits then padding (score 0, doc -1).  out_docs are
 * GLOBAL internal ids (doc_base added).
 * Range: the integer kernel is exact while every weight <= 65535 and sum_t w_t*max_impact(t) < 2^31.  A query beyond that
 * is scored by the WIDE path instead (u64 accumulators; rule S6 reproduced exactly, scores may exceed 2^31 as Lucene's
 * float32 scores would), on launches the batch makes anyway.  MSR_HOST handles any number of such queries; MSR_DEVICE is
 * asynchronous and takes one wide pass of 8 queries per call -- if a batch holds more, the first 8 are scored, the others'
 * rows stay empty and the sticky status (msr_sparse_index_status) reports MSR_ERR_RANGE.  k <= MSR_MAX_K.
 */
int msr_sparse_search(msr_sparse_index* h, uint32_t nq,
                      const uint64_t* q_indptr, const uint32_t* q_terms, const uint32_t* q_weights,
                      uint32_t k, float* out_scores, int32_t* out_docs, uint32_t* out_counts,
                      int where, void* stream);
/* Synchronises `stream`, returns and clears the sticky device-side status of earlier MSR_DEVICE searches. */
int msr_sparse_index_status(msr_sparse_index* h, void* stream);
/*
 * Algorithmic work of a query batch: *postings = sum_q sum_{t kept} df_shard(t),
 * *kept_terms = number of kept (query,term) pairs.  Used by bench.py for the roofline numerator.
 */
int msr_sparse_query_work(msr_sparse_index* h, uint32_t nq, const uint64_t* q_indptr, const uint32_t* q_terms,
                          const uint32_t* q_weights, uint64_t* postings, uint64_t* kept_terms,
                          int where, void* stream);

/*
 * PHYSICAL index bytes a query batch touches (device-path pointers): every distinct kept term of the batch counted
 * once -- a dense term as its u8 impact row (n_docs bytes; the u16 row too when some query weighs it >= 256) plus its
 * remainder postings, any other term as 4 bytes per posting slot over all tiles.  This is the DRAM floor of a batch
 * that starts with a cold L2 (bench.py's small-batch lines divide it by the measured time); it replaces no
 * reference call.
 */
int msr_sparse_query_traffic(msr_sparse_index* h, uint32_t nq, const uint64_t* q_indptr, const uint32_t* q_terms,
                             const uint32_t* q_weights, uint64_t* bytes, void* stream);

/*
 * Per-kernel timing for bench.py's roofline: synchronises `stream`, returns the summed CUDA-event durations (ms)
 * of the score kernel and of the merge kernel over all launches since the previous call plus their count, then
 * switches event recording on (enable != 0) or off for later searches.
 */
int msr_sparse_profile(msr_sparse_index* h, int enable, double* score_ms, double* merge_ms, uint64_t* launches,
                       void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Top-k list merge (doc-sharded multi-GPU).
 * scores/docs are [n_lists, nq, k_in], counts [n_lists, nq] (exactly the layout of an all-gather of per-GPU
 * results).  Output [nq,k] with the same (score desc, doc asc) order and padding.  Device pointers only.
 */
int msr_topk_merge(int device, uint32_t n_lists, uint32_t nq, uint32_t k_in,
                   const float* scores, const int32_t* docs, const uint32_t* counts,
                   uint32_t k, float* out_scores, int32_t* out_docs, uint32_t* out_counts, void* stream);

/*
 * msr_topk_merge over lists that are not packed back to back: list j starts list_stride elements after list j-1
 * (scores and docs) and count_list_stride elements (counts).  This is the layout of a symmetric-memory result buffer
 * [world][scores nq*k | docs nq*k | counts nq] that the fused exchange below fills.
 * The lists must be sorted best-first, as every libmsr search returns its rows (the kernel drops whatever lies below the
 * smallest of the lists' ceil(k / n_lists)-th keys before it sorts); msr_topk_merge accepts unsorted lists.
 */
int msr_topk_merge_strided(int device, uint32_t n_lists, uint32_t nq, uint32_t k_in, const float* scores,
                           const int32_t* docs, const uint32_t* counts, size_t list_stride, size_t count_list_stride,
                           uint32_t k, float* out_scores, int32_t* out_docs, uint32_t* out_counts, void* stream);

/*
 * Fused exchange of the doc-sharded search: after this call every MSR_DEVICE search on the handle stores
 * its merged result rows not only into out_* but ALSO into peer_scores[p] / peer_docs[p] / peer_counts[p] for p < n_peers
 * -- [nq,k] / [nq,k] / [nq] slots in the memory of other GPUs (peer pointers, e.g. from torch symmetric memory),
 * written by the epilogue of the tile-merge kernel over NVLink.  It replaces the all-gather that would follow the
 * search; the caller still has to synchronise the GPUs (a signal-pad barrier) before it merges the slots with
 * msr_topk_merge_strided.  n_peers = 0 switches it off.  At most 16 peers.
 */
int msr_sparse_set_peer_outputs(msr_sparse_index* h, uint32_t n_peers, float* const* peer_scores,
                                int32_t* const* peer_docs, uint32_t* const* peer_counts);

/*
 * Cross-GPU threshold sharing for the doc-sharded search (no reference counterpart -- FAISS IndexShards and
 * Lucene share nothing between shards).  `local_theta` u64[capacity] lives in symmetric memory of this GPU, `peer_theta[p]`
 * is the same array on peer p.  Device-path searches of at most `capacity` queries then keep the per-query pruning
 * threshold there and push every raise to the peers (red.max at system scope over NVLink): each shard prunes with the
 * k-th best key of the whole collection.  A shard's own list may then hold FEWER than k docs (never a wrong one); the merge
 * of the shards' lists is exact.  The caller clears `local_theta` before each search and must not reuse an array before
 * every GPU has finished the search that wrote into it (sharded.PeerExchange rotates three).  rot_num / rot_den: fraction
 * of the batch at which this GPU starts its walk over the queries (rank / world), so the GPUs publish different queries
 * first.  local_theta = NULL switches the sharing off.
 * Quantile sharing (local_quant != NULL; needs n_peers + 1 == rot_den, i.e. every GPU of the group): local_quant
 * u64[rot_den][capacity] holds, per GPU g and query, g's local r-th best key with r = ceil(k / rot_den); peer_quant[p] is
 * the same array on peer p.  Every GPU pushes its own row to all copies.  If every GPU holds r docs at or above t the
 * collection holds >= k, so min over g is a valid threshold -- for balanced shards close to the collection's true k-th key,
 * which a single shard's own k-th key is not.  The tile merge drops keys below max(theta, that minimum) as well.
 */
int msr_sparse_set_peer_theta(msr_sparse_index* h, uint32_t n_peers, uint64_t* local_theta, uint64_t* const* peer_theta,
                              uint32_t capacity, uint32_t rot_num, uint32_t rot_den, uint64_t* local_quant,
                              uint64_t* const* peer_quant);

/*
 * Publishing half of the fused exchange for searches whose last kernel does not store into peer slots itself (the dense
 * flat-IP search; the precedent is FAISS IndexShards behind GpuMultipleClonerOptions.shard, /root/reference/src/search.py:
 * 265-270): rows [nq,k] (scores f32, docs i32, counts) are copied into dst_*[p] for p < n_dst -- this rank's slot in its own
 * and in every peer's symmetric buffer -- with 16-byte stores when k % 4 == 0 and the bases are aligned.  Follow it with
 * msr_topk_merge_exchange.  Replaces three NCCL all-gathers.
 */
int msr_exchange_publish(int device, uint32_t nq, uint32_t k, const float* scores, const int32_t* docs, const uint32_t* counts,
                         uint32_t n_dst, float* const* dst_scores, int32_t* const* dst_docs, uint32_t* const* dst_counts,
                         void* stream);

/*
 * msr_topk_merge_strided behind the exchange's own one-way barrier (replaces a generic two-phase signal-pad barrier):
 * a one-warp kernel first stores `flag_value` into flag_peers[p][my_rank] on every peer (release, system scope; the kernel
 * that produced this GPU's rows precedes it on `stream`) and waits until flag_local[peer_ranks[p]] >= flag_value for every
 * peer (wrap-safe compare); the merge kernel follows on the stream.  flag_value must grow by one per exchange step; flags are u32 in symmetric memory,
 * zero-initialised.
 */
int msr_topk_merge_exchange(int device, uint32_t n_lists, uint32_t nq, uint32_t k_in, const float* scores,
                            const int32_t* docs, const uint32_t* counts, size_t list_stride, size_t count_list_stride,
                            uint32_t k, float* out_scores, int32_t* out_docs, uint32_t* out_counts, uint32_t n_peers,
                            uint32_t my_rank, uint32_t* flag_local, uint32_t* const* flag_peers, const uint32_t* peer_ranks,
                            uint32_t flag_value, void* stream);

/*
 * Stand-alone radix-select top-k over dense score rows (kernel (b) on its own): scores [nq, n] row-major
 * fp32 on the device; output k best per row by (score desc, column asc), columns offset by col_base.
 * Entries equal to -inf are never returned (counts < k possible).
 */
int msr_topk_rows(int device, uint32_t nq, uint32_t n, const float* scores, size_t row_stride,
                  uint32_t k, int32_t col_base, float* out_scores, int32_t* out_docs, uint32_t* out_counts,
                  void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Dense flat inner-product index (kernel (c)).
 * vecs = [n_docs, dim] row-major fp32 (the (reps, lookup) pickle of src/encode.py:386-457 after pickle_load,
 * search.py:49-52).  The index stores fp16 for the tensor-core pass (the reference's own
 * GPU storage type, GpuClonerOptions.useFloat16) AND fp32 for the exact re-score.
 * msr_dense_index_add appends rows (FaissFlatSearcher.add).  Row ids are doc_base + row.
 */
int msr_dense_index_create(int device, uint32_t dim, int32_t doc_base, msr_dense_index** out);
int msr_dense_index_add(msr_dense_index* h, uint32_t n_rows, const float* vecs, int where, void* stream);
int msr_dense_index_destroy(msr_dense_index* h);
/* n_docs, dim, device, bytes-resident(lo), bytes-resident(hi) */
int msr_dense_index_info(const msr_dense_index* h, uint32_t info[5]);
/*
 * Exact top-k by inner product: queries [nq, dim] fp32; out_scores fp32 [nq,k] are fp32 dot products
 * accumulated in a fixed order (|err| <= 1e-6 abs on unit vectors), out_ids int64 [nq,k] (FAISS convention:
 * -1 / -3.4028235e38 padding when k > n_docs).  Order = score desc, id asc.
 * mode: 0 = auto, 1 = exact fp32 CUDA-core pass only, 2 = tcgen05 fp16 pass + certified fp32 re-score.
 * Auto is a cost rule only (the results are the same bits): mode 2 for nq >= 8, or when the fp16 copy of the index is
 * large enough that reading half the bytes outweighs mode 2's fixed cost (>= 800 MB; >= 500 MB for nq >= 3); else mode 1.
 * Mode 2: approximate scores a(d) of all docs on the tensor cores with the selection FUSED into the GEMM epilogue (docs
 * with a(d) >= a per-query threshold, estimated from a strided sample of the index, are appended to a candidate buffer; no
 * score matrix is written).  With E = a Cauchy-Schwarz bound on |exact - a| from stored norms (fp16 rounding of both
 * sides + fp32 accumulation), the docs with a(d) >= a_k - 2E (a_k = k-th best a) provably contain the exact top-k; they
 * are re-scored in fp32 with the accumulation order of mode 1.  A query is certified when its buffer did not overflow,
 * holds >= k docs and the threshold was <= a_k - 2E; uncertified queries are re-run by mode 1.  Both modes therefore
 * return identical ids and bit-identical scores.  Needs dim % 8 == 0 (auto falls back to mode 1).  With MSR_DEVICE the
 * call is asynchronous in every mode: the uncertified queries of a query block are listed, re-run and scattered by launches
 * that take their count from device memory (at most 128 per block; more sets a sticky flag that msr_dense_last_fallback
 * reports as MSR_ERR_RANGE -- those queries kept their tensor-core rows).  MSR_HOST re-runs any number of them.
 */
int msr_dense_search(msr_dense_index* h, uint32_t nq, const float* queries, uint32_t k,
                     float* out_scores, int64_t* out_ids, int mode, int where, void* stream);

/*
 * Per-kernel timing of mode 2 for bench.py's roofline: returns the summed CUDA-event durations (ms) of the fused
 * tcgen05 pass (GEMM + selection epilogue) and of the exact re-score over all query blocks since the previous call
 * plus their count, then switches event recording on (enable != 0) or off.  Searches synchronise internally.
 */
int msr_dense_profile(msr_dense_index* h, int enable, double* tc_ms, double* rescore_ms, uint64_t* launches);
/*
 * Queries of the last mode-2 (or auto) search on this handle whose certificate failed and that were re-run exactly.
 * Synchronises the device (the device path counts there).  MSR_ERR_RANGE: see msr_dense_search.
 */
int msr_dense_last_fallback(const msr_dense_index* h, uint32_t* n);
/*
 * The tensor-core pass of mode 2 on its own (tests, benchmarks, profiling): out_scores[q*ld + d] =
 * sum_h fp16(queries[q,h]) * fp16(vecs[d,h]) accumulated in fp32 by tcgen05.mma.  Device pointers only; ld >= n_docs,
 * ld % 4 == 0, dim % 8 == 0.  Asynchronous on `stream`.
 */
int msr_dense_tc_scores(msr_dense_index* h, uint32_t nq, const float* queries, float* out_scores, size_t ld,
                        void* stream);
/*
 * msr_dense_search(mode 1, MSR_DEVICE) for callers that may not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid
 * cudaMalloc.  msr_dense_search keeps its score rows and top-k staging in the handle's grow-only workspace (cudaMalloc on
 * first use and on growth); this entry point calls no cudaMalloc / cudaFree / allocating library routine: its scratch is
 * a file-scope `static __device__` array of MSR_STATIC_SCRATCH_BYTES inside libmsr.so, which every GPU that loads the
 * library holds until the process exits (its device address is taken once per device with cudaGetSymbolAddress).
 * Same arguments, same kernels and therefore the same ids and bit-identical scores as mode 1: queries [nq,dim] fp32,
 * out_scores fp32 [nq,k], out_ids int64 [nq,k], all DEVICE pointers; asynchronous on `stream`.
 * Limit: four score rows + their staging must fit, (4*ceil4(n_docs) + 8*min(k,n_docs) + 4) * 4 + 4096 <=
 * MSR_STATIC_SCRATCH_BYTES, i.e. n_docs <= ~2.09 M at k = 100; beyond it the call returns MSR_ERR_INVALID with the sizes
 * in msr_last_error() and launches nothing.  Any nq (queries are taken in passes of as many rows as fit).
 * Every call on a GPU shares the one array and finds the last call's contents in it: nothing is read before the call has
 * written it and the count words are cleared first, so calls in stream order are independent -- but two calls that
 * OVERLAP on the same GPU (different streams, no ordering between them) race on it and return garbage.
 */
#define MSR_STATIC_SCRATCH_BYTES (32u << 20)
int msr_dense_search_static(msr_dense_index* h, uint32_t nq, const float* queries, uint32_t k, float* out_scores,
                            int64_t* out_ids, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * On-device query / document vector construction.
 * Replaces get_img_valid_tokens_values (/root/reference/src/encode.py:56-93; the full-vocabulary branch of
 * get_text_valid_tokens_values :114-124), the query-string builder (src/search.py:419-422) / doc-vector builder
 * (src/encode.py:351-352) and pyserini's TokFreqQueryEncoder for a batch of activation rows that are already on the GPU.
 * logits [n_rows, row_stride] of dtype 0 = f32, 1 = f16, 2 = bf16 (first `vocab` columns used).  Per row: top-`top`
 * by (value desc, token id asc); token ids >= vocab_len are skipped with the reference's value shift; weight =
 * rint(float32(value)*100) half-even; token id -> term id via token_to_term[vocab_len] (-1 = not in the index; NULL =
 * identity); duplicates merged (doc_mode 0: sum = query string semantics, 1: last wins = doc vector semantics).
 * Output = a query CSR at fixed pitch: out_indptr[r] = r*top, unused slots have weight 0 (ignored by msr_sparse_search).
 * Device pointers only, asynchronous on `stream`.  top <= 1024, vocab >= top.
 */
int msr_quantise_topk(int device, uint32_t n_rows, uint32_t vocab, const void* logits, int dtype, size_t row_stride,
                      uint32_t top, uint32_t vocab_len, const int32_t* token_to_term, int doc_mode,
                      uint64_t* out_indptr, uint32_t* out_terms, uint32_t* out_weights, void* stream);

/*
 * Candidate branch of the text quantiser: get_text_valid_tokens_values without --sparse_manual
 * (/root/reference/src/encode.py:96-112,125-135), the default of scripts/search_sparse.sh.  The caller tokenises the
 * caption on the host as the reference does (nltk words minus stop-words/punctuation -> tokenizer.encode -> set of token
 * ids, encode.py:97-100) and passes the per-row candidate lists as a HOST CSR (cand_indptr u64[n_rows+1], cand_tokens
 * i32[], in the order of the reference's list(token_ids)).  Per row: top-min(n_candidates, cap) of logits[candidates]
 * by (value desc, candidate position asc) -- the reference uses cap = 128; a row WITHOUT candidates falls back to the
 * top-`fallback_top` of the whole vocabulary (the reference uses 10).  Quantisation, the `id < vocab_len` skip with
 * its value shift, the token -> term map and the duplicate rule are those of msr_quantise_topk.  Output pitch =
 * max(cap, fallback_top) (+ n_expand) entries per row.  logits / token_to_term / out_* are device pointers; at most 4096 candidates
 * per row; candidate ids must be < vocab.  The call synchronises `stream` before returning (it reads pageable host
 * arrays).
 * Expansion terms (num_expended_tokens > 0, encode.py:137-152): with n_expand > 0 every row also gets the top-n_expand of
 * its logits over `filtered_ids - token_ids` = the tokens with expand_mask[v] != 0 (device u8[vocab], the reference's
 * get_filtered_ids, encode.py:38-47) minus the row's own candidates, appended behind the main entries exactly as the
 * reference appends them (so the value shift of skipped ids carries over); rows with fewer eligible tokens get fewer.
 * Output pitch = max(cap, fallback_top) + n_expand.  expand_mask NULL / n_expand 0 = off (what every script of the
 * reference passes).  The --sparse_manual and image variants of the expansion are not provided.
 */
int msr_quantise_candidates(int device, uint32_t n_rows, uint32_t vocab, const void* logits, int dtype, size_t row_stride,
                            const uint64_t* cand_indptr_host, const int32_t* cand_tokens_host, uint32_t cap,
                            uint32_t fallback_top, uint32_t vocab_len, const int32_t* token_to_term, int doc_mode,
                            const uint8_t* expand_mask, uint32_t n_expand,
                            uint64_t* out_indptr, uint32_t* out_terms, uint32_t* out_weights, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Hybrid fusion (kernel (d)) = hybrid.fuse([dense_run, sparse_run], [alpha, 1-alpha]) for a query batch.
 * Inputs per modality: scores f32 [nq,k_*], docs i32 [nq,k_*] (global internal ids), counts u32 [nq] and
 * self ids i32 [nq] (the query's own id in the doc id space, or -1): with remove_query != 0 a doc equal to
 * the self id is dropped AFTER min/max were taken over the full list (get_run_dict, search.py:66-82).
 * Output row q: the union in the reference's insertion order (dense list order, then sparse-only docs in sparse
 * list order): out_docs i32 [nq,k_d+k_s], out_scores f64 [nq,k_d+k_s], out_src u8 (1 dense only, 2 sparse only,
 * 3 both = fuse_statistic's tag, hybrid.py:56-90), out_counts u32 [nq].  w_dense / w_sparse are the reference's
 * weights [alpha, 1-alpha].  Arithmetic reproduces the reference under NumPy >= 2 bit for bit: float32 for the
 * dense term and for sums that contain one, float64 for sparse-only docs (rule H3).  dense_f32 = 0 switches the
 * dense term to float64 too (both runs hold Python floats, e.g. runs re-read with read_trec_run).  Doc ids must
 * be unique within a list.  Device pointers only.
 */
int msr_fuse(int device, uint32_t nq,
             uint32_t k_d, const float* d_scores, const int32_t* d_docs, const uint32_t* d_counts,
             uint32_t k_s, const float* s_scores, const int32_t* s_docs, const uint32_t* s_counts,
             const int32_t* self_ids, int remove_query, int dense_f32, double w_dense, double w_sparse,
             double* out_scores, int32_t* out_docs, uint8_t* out_src, uint32_t* out_counts, void* stream);
/*
 * msr_fuse for runs whose scores are arbitrary Python floats -- e.g. runs re-read with hybrid.read_trec_run
 * (/root/reference/src/hybrid.py:8-17), where float('0.1') is not float32(0.1): both score arrays are double and the whole
 * combination is float64, as the reference computes it on such runs (the dense_f32 = 0 typing of msr_fuse).
 */
int msr_fuse_f64(int device, uint32_t nq, uint32_t k_d, const double* d_scores, const int32_t* d_docs,
                 const uint32_t* d_counts, uint32_t k_s, const double* s_scores, const int32_t* s_docs,
                 const uint32_t* s_counts, const int32_t* self_ids, int remove_query, double w_dense, double w_sparse,
                 double* out_scores, int32_t* out_docs, uint8_t* out_src, uint32_t* out_counts, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Index builder: replaces the offline Lucene indexing step of the reference,
 * `python -m pyserini.index.lucene --collection JsonVectorCollection --impact --pretokenized` (scripts/sparse_index.sh:12-18),
 * for the JSONL files /root/reference/src/encode.py:353-359,426 writes ({"id", "content", "vector": {term: int}} per line).
 *
 * msr_jsonl_parse (host C++, no GPU needed): reads the files in the given order and builds the term dictionary, the
 * docid table and DOC-MAJOR postings under rules S1 (tf = vector[t], non-positive weights dropped, keys split on
 * whitespace, pieces add up; impacts above 65535 -> MSR_ERR_RANGE), S2 (terms verbatim) and S8 (doc_rank = rank of the
 * docid string in UTF-8 byte order; duplicate ids -> MSR_ERR_INVALID).  Term ids follow first appearance.
 * msr_collection_info: n_docs, n_terms, nnz, total bytes of docid strings, total bytes of term strings.
 * msr_collection_export copies into caller buffers (any may be NULL): doc_indptr u64[n_docs+1], terms u32[nnz],
 * impacts u16[nnz], doc_rank u32[n_docs] (file order -> internal id), and the two string tables as a byte blob plus
 * offsets u64[n+1] each (docids in FILE order).
 *
 * msr_invert_docs (device): doc-major postings -> the canonical term-major CSR msr_sparse_index_build takes (postings
 * grouped by term, internal doc ids strictly ascending): 64-bit keys (term, internal doc) from one warp per document,
 * a radix sort over the used key bits, a histogram scan for indptr.  doc_rank NULL = documents are already in internal
 * order.  Outputs: out_indptr u64[n_terms+1], out_docids u32[nnz], out_impacts u16[nnz]; host or device pointers per
 * `where` (the call synchronises `stream`).  A term occurring twice in one document is reported as MSR_ERR_INVALID.
 */
typedef struct msr_collection msr_collection;
int msr_jsonl_parse(const char* const* paths, uint32_t n_paths, msr_collection** out);
int msr_collection_info(const msr_collection* c, uint64_t info[5]);
int msr_collection_export(const msr_collection* c, uint64_t* doc_indptr, uint32_t* terms, uint16_t* impacts,
                          uint32_t* doc_rank, char* docid_blob, uint64_t* docid_off, char* term_blob, uint64_t* term_off);
int msr_collection_destroy(msr_collection* c);
int msr_invert_docs(int device, uint32_t n_docs, uint32_t n_terms, const uint64_t* doc_indptr, const uint32_t* terms,
                    const uint16_t* impacts, const uint32_t* doc_rank, uint64_t* out_indptr, uint32_t* out_docids,
                    uint16_t* out_impacts, int where, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Recall@k hit counts on the device.
 * Replaces RecallMetrics._sort/_count, /root/reference/src/metrices.py:29-35 and :76-84: for every query whose ranked
 * list is not empty (counts[q] > 0), hits[c] += 1 when ANY of the query's targets t_docs[t_indptr[q]..t_indptr[q+1])
 * appears among the first cutoffs[c] entries of docs[q*row_stride ...] (rows already in rank order: score desc, the
 * reference's stable sort).  *n_eval = number of non-empty lists.  The reference's recall is hits / (number of queries),
 * metrices.py:86-100.  Device pointers only, asynchronous on `stream`; n_cut <= 32.
 */
int msr_recall_hits(int device, uint32_t nq, const int32_t* docs, size_t row_stride, const uint32_t* counts,
                    const uint64_t* t_indptr, const int32_t* t_docs, uint32_t n_cut, const uint32_t* cutoffs,
                    uint64_t* hits, uint64_t* n_eval, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Provenance histogram: the loop of /root/reference/src/score_statistic.py:406-417 over the
 * output of fuse_statistic (hybrid.py:56-90) = msr_fuse's (out_scores, out_src, out_counts).  Per query the fused list
 * is ordered by score descending with equal scores in insertion order (Python's stable sort), cut at `depth` (the
 * reference uses 200), and hist[(tag-1)*depth + rank-1] += 1 for the entry at 1-based `rank` with tag 1 = dense only,
 * 2 = sparse only, 3 = both ("fuse").  hist u64[3*depth] is zeroed by the call.  Device pointers, asynchronous on `stream`.
 */
int msr_provenance_hist(int device, uint32_t nq, const double* scores, const uint8_t* src, size_t row_stride,
                        const uint32_t* counts, uint32_t depth, uint64_t* hist, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Final ordering of run rows: Python's stable `sorted(run.items(), key=score, reverse=True)` of
 * /root/reference/src/metrices.py:30 and hybrid.write_trec_run (hybrid.py:20-29), and the removal of the query's own id that
 * get_run_dict performs (search.py:66-82; order kept), in one kernel.  Row q has counts[q] entries (scores f32 or f64 per
 * `scores_f64`, docs i32, optional provenance tags u8); an entry whose doc equals self_ids[q] (>= 0) is dropped; the rest
 * are written to out_* by (score desc, original position asc) when by_score != 0, in their original order otherwise, and
 * out_counts[q] is what remains.  width <= 2 * MSR_MAX_K.  Not in place.  Device pointers, asynchronous on `stream`.
 */
int msr_order_rows(int device, uint32_t nq, uint32_t width, const void* scores, int scores_f64, const int32_t* docs,
                   const uint8_t* tags, size_t row_stride, const uint32_t* counts, const int32_t* self_ids, int by_score,
                   void* out_scores, int32_t* out_docs, uint8_t* out_tags, uint32_t* out_counts, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Measured ceilings for `roofline.binding` (csrc/microbench.cu; replaces no reference call).  The sparse scorer shares
 * postings across the query batch through L2, so HBM does not bound it; bench.py reports it against these numbers,
 * measured on the same GPU in the same run.  *units_per_s: bytes/s for the two read benchmarks (`bytes` = buffer size;
 * L2: choose <= 64 MiB so it stays L2-resident, HBM: >= 1 GiB), instructions (wavefronts for MSR_MB_LDS) per second for the others (`bytes` unused).
 * Synchronous.
 */
#define MSR_MB_L2_READ 0      /* ld.global.nc.L1::no_allocate.v2.u32, 8 in flight per thread, L2-resident buffer */
#define MSR_MB_HBM_READ 1     /* the same loop over a buffer far larger than L2 */
#define MSR_MB_RED_SHARED 2   /* red.shared.add.u32, one bank per lane */
#define MSR_MB_ISSUE 3        /* independent dp4a chains (fma pipe) */
#define MSR_MB_ALU 4          /* independent PRMT chains (alu pipe) */
#define MSR_MB_LDS 5          /* conflict-free 16-byte ld.shared: wavefronts/s of the l1tex data pipe */
#define MSR_MB_IMMA 6         /* mma.sync.m16n8k32 u8 x u8 -> s32 instructions/s */
#define MSR_MB_IMMA_PRMT 7    /* the same with 8 PRMT per mma (the dense phase's 4x4 byte transposes) */
int msr_microbench(int device, int what, uint64_t bytes, uint32_t iters, double* units_per_s, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MSR_H_ */
