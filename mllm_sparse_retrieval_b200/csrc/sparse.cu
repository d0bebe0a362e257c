// sparse.cu -- kernels (a)+(b): tiled inverted-index impact scorer with the top-k fused into its epilogue.
//
// Replaces pyserini LuceneImpactSearcher.batch_search (reference call sites src/search.py:86-87, 273-275) under
// the rules S1-S10 of SURVEY.md 8(a).
//
// HBM layout (built once by msr_sparse_index_build):
//   * the doc id space of the shard is cut into n_tiles contiguous tiles of tile_docs docs (<= 65536);
//   * postings are bucketed by (tile, term) and packed into ONE u32 each: (impact << 16) | (doc - tile_lo).
//     4 B/posting physical against the canonical 6 B (u32 docid + u16 impact) that SURVEY.md 8(d) fixes as the
//     algorithmic figure;
//   * buckets of at most kSmallBucket postings keep ascending doc order.  Larger buckets are laid out
//     BANK-CONFLICT-FREE: the bucket is cut into G = max(ceil(n/32), max postings per shared-memory bank)
//     groups of 32 slots, postings are dealt round-robin (bank-major) over the groups so that no group holds
//     two docs of the same bank (doc & 31), and the free slots of a group are filled with impact-0 postings on
//     the group's unused banks.  One warp-wide shared-memory atomic per group then costs one wavefront
//     instead of the ~3.5 a random 32-doc group costs.  Integer adds commute, so the order inside a bucket
//     never changes a score;
//   * seg_ptr[tile * n_terms + term] = first posting of the bucket (u32, one flat array, +1 sentinel);
//   * term_keep[term] = rule S5 with collection-wide df / N;  term_maximp[term] = max impact (range checks).
//
// Search kernel: one CTA per (query, tile); grid is tile-major so the CTAs resident at any moment share one
// tile's slice of the index through the 126 MB L2.  The CTA stages the query's (bucket start, length, weight)
// list in shared memory, streams the buckets with coalesced 32-bit no-allocate loads (8 in flight per lane)
// and accumulates integer impacts with shared-memory atomics into a u32 accumulator per doc of the tile.
// The accumulators never leave shared memory: the epilogue radix-selects the tile's top-k by
// (float32(score) desc, doc asc) and writes k packed 64-bit candidates per (query, tile); a small merge kernel
// (topk.cu) produces the final sorted list.
#include <cub/device/device_scan.cuh>

#include <math.h>
#include <stdlib.h>

#include <vector>

#include "common.cuh"
#include "sparse_kernel.cuh"

namespace msr {

constexpr uint32_t kDefaultTileDocs = 24576;   // 96 KB of accumulators -> 2 CTAs per SM
constexpr uint32_t kMaxTileDocs = 53248;       // 208 KB of accumulators + staging = the 227 KB opt-in limit
constexpr uint32_t kSmallBucket = 96;          // buckets above this size get the bank-conflict-free layout
constexpr float kDenseFraction = 0.15f;        // terms present in >= this fraction of the shard's docs are stored as rows
                                               // (measured at 1 M docs: 0.10 465 K, 0.15 469 K, 0.20 466 K, 0.25 460 K, 0.30 453 K QPS)
constexpr int kBuildWarps = 8;                 // warps per CTA of the bucket layout kernels

int topk_merge_keys_launch(uint32_t nq, uint32_t n_lists, uint32_t stride, const uint64_t* keys,
                           const uint32_t* cnts, uint32_t k, float* out_scores, int32_t* out_docs,
                           uint32_t* out_counts, cudaStream_t s, uint32_t n_peers, float* const* peer_scores,
                           int32_t* const* peer_docs, uint32_t* const* peer_counts, size_t peer_row0,
                           const unsigned long long* theta, const unsigned long long* quant, uint32_t quant_world,
                           uint32_t quant_cap, uint32_t wide_ctas, uint32_t wide_first, const uint32_t* wide_q,
                           const uint32_t* n_wide, const uint64_t* wide_keys, const uint32_t* wide_cnt, uint32_t wide_n_lists,
                           uint32_t wide_stride);

}  // namespace msr

struct msr_sparse_index {
    int device = 0;
    uint32_t n_docs = 0, n_terms = 0, tile_docs = 0, n_tiles = 0, n_docs_global = 0;
    int32_t doc_base = 0;
    uint64_t nnz = 0, n_slots = 0;    // real postings / posting slots incl. the padding of conflict-free buckets
    uint32_t* seg_ptr = nullptr;      // [n_tiles*n_terms + 1]
    uint32_t* postings = nullptr;     // [n_slots]
    uint8_t* term_keep = nullptr;     // [n_terms]
    uint32_t* term_maximp = nullptr;  // [n_terms]
    uint32_t* dense_id = nullptr;     // [n_terms] row of a dense term in dense_rows, kNotDense otherwise
    uint16_t* dense_rows = nullptr;   // [n_dense][dense_stride] impact of the term in every doc of the shard (0 = absent)
    uint8_t* dense_rows8 = nullptr;   // [n_dense][dense_stride] min(impact, 255); the rest (impact - 255) is an ordinary
                                      // posting of the term's (tile, term) bucket, so row + bucket add up to the impact
    uint32_t* dense_cursor = nullptr; // build-time only
    uint32_t n_dense = 0;
    size_t dense_stride = 0;
    uint32_t* df_local = nullptr;     // [n_terms]
    uint32_t* status = nullptr;       // sticky device status word
    uint64_t* work = nullptr;         // [2] scratch for msr_sparse_query_work
    msr::Workspace ws;
    size_t resident_bytes = 0;
    // optional per-kernel timing (msr_sparse_profile): events bracket every score / merge launch on the stream
    bool use_theta = true;            // cross-tile threshold sharing (MSR_SPARSE_NO_THETA=1 in the environment disables)
    bool fast_select = true;          // exact select bounded by the per-thread maxima first (MSR_SPARSE_FAST_SELECT=0 disables)
    bool profiling = false;
    std::vector<cudaEvent_t> ev;      // triples: before score, after score, after merge
    size_t ev_used = 0;
    // fused exchange (msr_sparse_set_peer_outputs): result slots on other GPUs that device-path searches also fill
    uint32_t n_peers = 0;
    float* peer_scores[16] = {};
    int32_t* peer_docs[16] = {};
    uint32_t* peer_counts[16] = {};
    // cross-GPU threshold sharing (msr_sparse_set_peer_theta)
    unsigned long long* ext_theta = nullptr;      // this GPU's theta[] in symmetric memory (capacity ext_theta_cap queries)
    unsigned long long* peer_theta[15] = {};
    unsigned long long* ext_quant = nullptr;      // [world][ext_theta_cap] r-th best keys of every GPU (see sparse_kernel.cuh)
    unsigned long long* peer_quant[15] = {};
    uint32_t n_peer_theta = 0, ext_theta_cap = 0, q_rot_num = 0, q_rot_den = 1;
};

namespace msr {

// ---------------------------------------------------------------------------------------------------------
// index build
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t term_of_posting(const uint64_t* __restrict__ indptr, uint32_t n_terms, uint64_t p) {
    // largest t with indptr[t] <= p
    uint32_t lo = 0, hi = n_terms;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (indptr[mid] <= p) lo = mid; else hi = mid;
    }
    return lo;
}

// Validates the CSR, records max impact per term, writes the postings of DENSE terms into their impact rows and
// counts the postings of the other terms per (tile, term) bucket.
__global__ void build_count_kernel(const uint64_t* __restrict__ indptr, const uint32_t* __restrict__ docids,
                                   const uint16_t* __restrict__ impacts, uint64_t nnz, uint32_t n_terms,
                                   uint32_t n_docs, uint32_t tile_docs, const uint32_t* __restrict__ dense_id,
                                   uint16_t* __restrict__ dense_rows, uint8_t* __restrict__ dense_rows8, size_t dense_stride,
                                   uint32_t* __restrict__ cnt, uint32_t* __restrict__ maximp, uint32_t* __restrict__ bad) {
    for (uint64_t p = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; p < nnz; p += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t t = term_of_posting(indptr, n_terms, p);
        uint32_t d = docids[p];
        uint32_t imp = impacts[p];
        if (d >= n_docs || imp == 0) { atomicOr(bad, 1u); continue; }
        if (p > indptr[t] && docids[p - 1] >= d) atomicOr(bad, 2u);
        const uint32_t di = dense_id[t];
        if (di != kNotDense) {
            dense_rows[(size_t)di * dense_stride + d] = (uint16_t)imp;
            dense_rows8[(size_t)di * dense_stride + d] = (uint8_t)min(imp, 255u);
            if (imp > 255u) atomicAdd(&cnt[(size_t)(d / tile_docs) * n_terms + t], 1u);
        } else {
            atomicAdd(&cnt[(size_t)(d / tile_docs) * n_terms + t], 1u);
        }
        atomicMax(&maximp[t], imp);
    }
}

__global__ void build_scatter_kernel(const uint64_t* __restrict__ indptr, const uint32_t* __restrict__ docids,
                                     const uint16_t* __restrict__ impacts, uint64_t nnz, uint32_t n_terms,
                                     uint32_t tile_docs, const uint32_t* __restrict__ dense_id,
                                     const uint32_t* __restrict__ seg_ptr, uint32_t* __restrict__ postings,
                                     uint32_t n_dense, uint32_t* __restrict__ dense_cursor) {
    for (uint64_t p = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; p < nnz; p += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t t = term_of_posting(indptr, n_terms, p);
        uint32_t d = docids[p];
        uint32_t tile = d / tile_docs;
        uint32_t tile_lo = tile * tile_docs;
        const uint32_t di = dense_id[t];
        if (di != kNotDense) {
            // dense term: only the part of an impact above 255 is a posting; slot order inside the bucket is
            // arbitrary (integer adds commute, the layout pass re-orders by bank anyway)
            const uint32_t imp = impacts[p];
            if (imp > 255u) {
                const uint32_t r = atomicAdd(&dense_cursor[(size_t)tile * n_dense + di], 1u);
                postings[seg_ptr[(size_t)tile * n_terms + t] + r] = ((imp - 255u) << 16) | perm32(d - tile_lo);
            }
            continue;
        }
        // first posting of term t whose doc >= tile_lo
        uint64_t lo = indptr[t], hi = p;   // docids[p] >= tile_lo, so the answer is in [lo, p]
        while (lo < hi) {
            uint64_t mid = (lo + hi) >> 1;
            if (docids[mid] < tile_lo) lo = mid + 1; else hi = mid;
        }
        uint32_t r = (uint32_t)(p - lo);
        postings[seg_ptr[(size_t)tile * n_terms + t] + r] = ((uint32_t)impacts[p] << 16) | perm32(d - tile_lo);
    }
}

__global__ void sum_u32_kernel(const uint32_t* __restrict__ x, size_t n, unsigned long long* __restrict__ out) {
    unsigned long long a = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) a += x[i];
    for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if ((threadIdx.x & 31) == 0 && a) atomicAdd(out, a);
}

// Per-bank posting counts of one large bucket (lane b returns the count of bank b); hist = 32 u32 of this warp.
__device__ __forceinline__ uint32_t bucket_bank_counts(const uint32_t* __restrict__ stage, uint32_t n, uint32_t* hist,
                                                       uint32_t lane) {
    hist[lane] = 0;
    __syncwarp();
    for (uint32_t i = lane; i < n; i += 32) atomicAdd(&hist[stage[i] & 31u], 1u);
    __syncwarp();
    return hist[lane];
}

// Pass 1 of the bucket layout: slots each bucket occupies (n for small buckets, 32*G for large ones).
__global__ void __launch_bounds__(kBuildWarps * 32)
build_plan_kernel(const uint32_t* __restrict__ seg0, const uint32_t* __restrict__ stage, size_t n_buckets,
                  uint32_t* __restrict__ slots) {
    __shared__ uint32_t hist[kBuildWarps][32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (size_t b = (size_t)blockIdx.x * kBuildWarps + warp; b < n_buckets; b += (size_t)gridDim.x * kBuildWarps) {
        const uint32_t lo = seg0[b], n = seg0[b + 1] - lo;
        if (n <= kSmallBucket) {
            if (lane == 0) slots[b] = n;
            continue;
        }
        uint32_t c = bucket_bank_counts(stage + lo, n, hist[warp], lane);
        const uint32_t g0 = (n + 31) / 32, G = max(g0, warp_reduce_max(c));
        if (lane == 0) slots[b] = G <= 2 * g0 ? G * 32 : n;     // skewed banks: keep doc order, no padding
        __syncwarp();
    }
}

// Pass 2: write the final posting array.  Large buckets: posting number s (bank-major order) goes to group s % G,
// slot s / G; the unused slots of a group get impact-0 postings on the group's free banks.
__global__ void __launch_bounds__(kBuildWarps * 32)
build_layout_kernel(const uint32_t* __restrict__ seg0, const uint32_t* __restrict__ stage, size_t n_buckets,
                    const uint32_t* __restrict__ seg_ptr, uint32_t* __restrict__ postings) {
    __shared__ uint32_t hist[kBuildWarps][32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    for (size_t b = (size_t)blockIdx.x * kBuildWarps + warp; b < n_buckets; b += (size_t)gridDim.x * kBuildWarps) {
        const uint32_t lo = seg0[b], n = seg0[b + 1] - lo;
        uint32_t* out = postings + seg_ptr[b];
        if (n <= kSmallBucket) {
            for (uint32_t i = lane; i < n; i += 32) out[i] = stage[lo + i];
            continue;
        }
        uint32_t c = bucket_bank_counts(stage + lo, n, hist[warp], lane);
        const uint32_t g0 = (n + 31) / 32, G = max(g0, warp_reduce_max(c));
        __syncwarp();
        if (G > 2 * g0) {
            for (uint32_t i = lane; i < n; i += 32) out[i] = stage[lo + i];
            continue;
        }
        hist[warp][lane] = warp_incl_scan(c, lane) - c;          // first sequence number of each bank
        __syncwarp();
        for (uint32_t i = 0; i < n; i += 32) {
            const uint32_t idx = i + lane;
            const bool act = idx < n;
            const uint32_t v = act ? stage[lo + idx] : 0u;
            const uint32_t bank = v & 31u;
            const uint32_t m = __match_any_sync(0xffffffffu, act ? bank : 32u + lane);
            const uint32_t r = __popc(m & lt);
            uint32_t s = 0;
            if (act) s = hist[warp][bank] + r;
            __syncwarp();
            if (act && r == 0) hist[warp][bank] += __popc(m);
            __syncwarp();
            if (act) out[(s % G) * 32 + s / G] = v;
        }
        __syncwarp();
        for (uint32_t g = 0; g < G; ++g) {
            const bool real = lane * G + g < n;
            const uint32_t v = real ? out[g * 32 + lane] : 0u;
            const uint32_t used = __reduce_or_sync(0xffffffffu, real ? (1u << (v & 31u)) : 0u);
            const uint32_t pads = __ballot_sync(0xffffffffu, !real);
            if (!real) out[g * 32 + lane] = __fns(~used, 0, __popc(pads & lt) + 1);
        }
        __syncwarp();
    }
}

// keep (rule S5), local df and the dense flag of every term (dense: df_local >= dense_min postings in this shard)
__global__ void build_term_meta_kernel(const uint64_t* __restrict__ indptr, const uint32_t* __restrict__ df_global,
                                       uint32_t n_terms, uint32_t n_docs_global, uint32_t dense_min,
                                       uint8_t* __restrict__ keep, uint32_t* __restrict__ df_local,
                                       uint32_t* __restrict__ dense_flag) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > n_terms) return;
    if (t == n_terms) { dense_flag[t] = 0; return; }              // sentinel slot of the scan
    uint32_t dfl = (uint32_t)(indptr[t + 1] - indptr[t]);
    uint64_t df = df_global ? df_global[t] : dfl;
    // S5: in the dictionary and ln(N/(df+1)) > 0  <=>  0 < df and df + 1 < N
    keep[t] = (df > 0 && df + 1 < (uint64_t)n_docs_global) ? 1 : 0;
    df_local[t] = dfl;
    dense_flag[t] = dfl >= dense_min ? 1u : 0u;
}

// dense_id[t] = row number of a dense term (exclusive scan of the flags), kNotDense otherwise
__global__ void build_dense_id_kernel(const uint32_t* __restrict__ dense_flag, uint32_t n_terms, uint32_t* __restrict__ dense_id) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n_terms) dense_id[t] = dense_flag[t] ? dense_id[t] : kNotDense;
}

__global__ void query_work_kernel(const uint8_t* __restrict__ keep, const uint32_t* __restrict__ df_local,
                                  uint32_t n_terms, uint64_t nnz_q, const uint32_t* __restrict__ q_terms,
                                  const uint32_t* __restrict__ q_weights, unsigned long long* __restrict__ out) {
    unsigned long long p = 0, c = 0;
    for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < nnz_q; j += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t t = q_terms[j];
        if (t < n_terms && q_weights[j] && keep[t]) { p += df_local[t]; c += 1; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        p += __shfl_xor_sync(0xffffffffu, p, o);
        c += __shfl_xor_sync(0xffffffffu, c, o);
    }
    if ((threadIdx.x & 31) == 0 && c) { atomicAdd(&out[0], p); atomicAdd(&out[1], c); }
}

// physical traffic of a batch: mark the distinct kept terms (flag 1 = used, 2 = used with a weight the u8 rows cannot take)
__global__ void traffic_mark_kernel(const uint8_t* __restrict__ keep, const uint32_t* __restrict__ maximp, uint32_t n_terms,
                                    uint64_t nnz_q, const uint32_t* __restrict__ q_terms, const uint32_t* __restrict__ q_weights,
                                    uint32_t* __restrict__ flags) {
    for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < nnz_q; j += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t t = q_terms[j], w = q_weights[j];
        if (t < n_terms && w && keep[t]) atomicOr(&flags[t], (w >= 256u || (uint64_t)w * maximp[t] >= (1u << 24)) ? 3u : 1u);
    }
}
__global__ void traffic_sum_kernel(const uint32_t* __restrict__ flags, uint32_t n_terms, uint32_t n_tiles, uint32_t n_docs,
                                   const uint32_t* __restrict__ seg_ptr, const uint32_t* __restrict__ dense_id,
                                   unsigned long long* __restrict__ out) {
    unsigned long long b = 0;
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n_terms; t += gridDim.x * blockDim.x) {
        const uint32_t f = flags[t];
        if (!f) continue;
        const bool dense = dense_id[t] != kNotDense;
        if (dense) b += (f & 2u) ? 3ull * n_docs : (unsigned long long)n_docs;   // u8 row (+ the u16 row for wide weights)
        {                                                        // bucket postings (a dense term's: the remainder above 255)
            for (uint32_t tile = 0; tile < n_tiles; ++tile) {
                const size_t i = (size_t)tile * n_terms + t;
                b += 4ull * (seg_ptr[i + 1] - seg_ptr[i]) + 8ull;   // slots + the two seg_ptr words
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) b += __shfl_xor_sync(0xffffffffu, b, o);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(out, b);
}

static size_t search_smem_bytes(uint32_t tile_docs) {
    return sizeof(StageUnion) + sizeof(SmallState) + (size_t)((tile_docs + 255) & ~255u) * 4;   // whole 256-doc blocks (one per warp and step)
}

}  // namespace msr

using namespace msr;

// ---------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------
extern "C" int msr_sparse_index_build(int device, uint32_t n_docs, uint32_t n_terms, const uint64_t* indptr,
                                      const uint32_t* docids, const uint16_t* impacts, uint32_t n_docs_global,
                                      const uint32_t* df_global, int32_t doc_base, uint32_t tile_docs, int where,
                                      void* stream, msr_sparse_index** out) {
    MSR_REQUIRE(out != nullptr, "msr_sparse_index_build: out is NULL");
    *out = nullptr;
    MSR_REQUIRE(n_docs > 0 && n_terms > 0, "msr_sparse_index_build: empty collection (n_docs=%u n_terms=%u)", n_docs, n_terms);
    MSR_REQUIRE(indptr != nullptr, "msr_sparse_index_build: indptr is NULL");
    MSR_REQUIRE(where == MSR_HOST || where == MSR_DEVICE, "msr_sparse_index_build: bad `where`");
    MSR_REQUIRE(tile_docs <= kMaxTileDocs, "msr_sparse_index_build: tile_docs %u > %u", tile_docs, kMaxTileDocs);
    MSR_REQUIRE(n_docs_global >= n_docs, "msr_sparse_index_build: n_docs_global < n_docs");
    int rc = use_device(device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;

    if (tile_docs == 0) {
        // balanced tiles no larger than the default
        uint32_t nt = (n_docs + kDefaultTileDocs - 1) / kDefaultTileDocs;
        tile_docs = (n_docs + nt - 1) / nt;
        tile_docs = (tile_docs + 255) & ~255u;
    }
    tile_docs = (tile_docs + 31) & ~31u;      // accumulator positions are a permutation inside 32-doc groups (perm32)
    const uint32_t n_tiles = (n_docs + tile_docs - 1) / tile_docs;

    // canonical CSR on the device
    DevBuf b_indptr, b_docids, b_impacts, b_df;
    const uint64_t* d_indptr = indptr;
    const uint32_t* d_docids = docids;
    const uint16_t* d_impacts = impacts;
    const uint32_t* d_df = df_global;
    uint64_t nnz = 0;
    if (where == MSR_HOST) {
        nnz = indptr[n_terms];
        MSR_CUDA_CHECK(b_indptr.alloc((size_t)(n_terms + 1) * 8));
        MSR_CUDA_CHECK(cudaMemcpyAsync(b_indptr.p, indptr, (size_t)(n_terms + 1) * 8, cudaMemcpyHostToDevice, s));
        d_indptr = b_indptr.as<uint64_t>();
    } else {
        MSR_CUDA_CHECK(cudaMemcpyAsync(&nnz, indptr + n_terms, 8, cudaMemcpyDeviceToHost, s));
        MSR_CUDA_CHECK(cudaStreamSynchronize(s));
    }
    MSR_REQUIRE(nnz < (1ull << 32), "msr_sparse_index_build: nnz %llu >= 2^32 (shard the collection)", (unsigned long long)nnz);
    MSR_REQUIRE(nnz == 0 || (docids && impacts), "msr_sparse_index_build: docids/impacts NULL");
    if (where == MSR_HOST) {
        MSR_CUDA_CHECK(b_docids.alloc(nnz * 4));
        MSR_CUDA_CHECK(b_impacts.alloc(nnz * 2));
        MSR_CUDA_CHECK(cudaMemcpyAsync(b_docids.p, docids, nnz * 4, cudaMemcpyHostToDevice, s));
        MSR_CUDA_CHECK(cudaMemcpyAsync(b_impacts.p, impacts, nnz * 2, cudaMemcpyHostToDevice, s));
        d_docids = b_docids.as<uint32_t>();
        d_impacts = b_impacts.as<uint16_t>();
        if (df_global) {
            MSR_CUDA_CHECK(b_df.alloc((size_t)n_terms * 4));
            MSR_CUDA_CHECK(cudaMemcpyAsync(b_df.p, df_global, (size_t)n_terms * 4, cudaMemcpyHostToDevice, s));
            d_df = b_df.as<uint32_t>();
        }
    }

    msr_sparse_index* h = new (std::nothrow) msr_sparse_index();
    if (!h) { set_error("msr_sparse_index_build: out of host memory"); return MSR_ERR_NOMEM; }
    h->device = device; h->n_docs = n_docs; h->n_terms = n_terms; h->tile_docs = tile_docs; h->n_tiles = n_tiles;
    h->n_docs_global = n_docs_global; h->doc_base = doc_base; h->nnz = nnz;
    { const char* e = getenv("MSR_SPARSE_NO_THETA"); h->use_theta = !(e && e[0] == '1'); }
    { const char* e = getenv("MSR_SPARSE_FAST_SELECT"); h->fast_select = !(e && e[0] == '0'); }
    const size_t n_seg = (size_t)n_tiles * n_terms + 1;
    DevBuf b_cnt, b_bad, b_tmp;
    cudaError_t e = cudaSuccess;
    auto fail = [&](const char* what) {
        set_error("msr_sparse_index_build: %s: %s", what, cudaGetErrorString(e));
        msr_sparse_index_destroy(h);
        return (int)MSR_ERR_CUDA;
    };
    if ((e = cudaMalloc(&h->seg_ptr, n_seg * 4)) != cudaSuccess) return fail("alloc seg_ptr");
    if ((e = cudaMalloc(&h->term_keep, n_terms)) != cudaSuccess) return fail("alloc term_keep");
    if ((e = cudaMalloc(&h->term_maximp, (size_t)n_terms * 4)) != cudaSuccess) return fail("alloc term_maximp");
    if ((e = cudaMalloc(&h->df_local, (size_t)n_terms * 4)) != cudaSuccess) return fail("alloc df_local");
    if ((e = cudaMalloc(&h->dense_id, (size_t)(n_terms + 1) * 4)) != cudaSuccess) return fail("alloc dense_id");
    if ((e = cudaMalloc(&h->status, 4)) != cudaSuccess) return fail("alloc status");
    if ((e = cudaMalloc(&h->work, 16)) != cudaSuccess) return fail("alloc work");
    DevBuf b_seg0, b_stage, b_flag;
    if ((e = b_cnt.alloc(n_seg * 4)) != cudaSuccess) return fail("alloc cnt");
    if ((e = b_seg0.alloc(n_seg * 4)) != cudaSuccess) return fail("alloc seg0");
    if ((e = b_flag.alloc((size_t)(n_terms + 1) * 4)) != cudaSuccess) return fail("alloc dense flags");
    if ((e = b_bad.alloc(4)) != cudaSuccess) return fail("alloc bad");
    cudaMemsetAsync(b_cnt.p, 0, n_seg * 4, s);
    cudaMemsetAsync(h->term_maximp, 0, (size_t)n_terms * 4, s);
    cudaMemsetAsync(b_bad.p, 0, 4, s);
    cudaMemsetAsync(h->status, 0, 4, s);

    // 0. per-term metadata; terms present in >= kDenseFraction of the shard's docs become dense impact rows
    const int bt = 256;
    size_t tmp_bytes = 0, tmp2 = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, b_cnt.as<uint32_t>(), b_seg0.as<uint32_t>(), n_seg, s);
    cub::DeviceScan::ExclusiveSum(nullptr, tmp2, b_flag.as<uint32_t>(), h->dense_id, n_terms + 1, s);
    tmp_bytes = std::max(tmp_bytes, tmp2);
    if ((e = b_tmp.alloc(tmp_bytes)) != cudaSuccess) return fail("alloc scan temp");
    uint32_t dense_min = 0xFFFFFFFFu;                               // dense rows need 16-byte aligned tile starts
    { const char* ev = getenv("MSR_SPARSE_NO_DENSE");
      if (tile_docs % 8 == 0 && !(ev && ev[0] == '1'))
      {
          double frac = kDenseFraction;                              // MSR_SPARSE_DENSE_FRACTION: tuning knob (0 < f <= 1)
          if (const char* ef = getenv("MSR_SPARSE_DENSE_FRACTION")) { const double f = atof(ef); if (f > 0.0 && f <= 1.0) frac = f; }
          dense_min = std::max<uint32_t>(32u, (uint32_t)std::ceil(frac * n_docs));
      } }
    build_term_meta_kernel<<<(n_terms + 1 + bt - 1) / bt, bt, 0, s>>>(d_indptr, d_df, n_terms, n_docs_global, dense_min,
                                                                      h->term_keep, h->df_local, b_flag.as<uint32_t>());
    count_launch();
    if ((e = cub::DeviceScan::ExclusiveSum(b_tmp.p, tmp_bytes, b_flag.as<uint32_t>(), h->dense_id, n_terms + 1, s)) != cudaSuccess)
        return fail("dense scan");
    if ((e = cudaMemcpyAsync(&h->n_dense, h->dense_id + n_terms, 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail("read n_dense");
    build_dense_id_kernel<<<(n_terms + bt - 1) / bt, bt, 0, s>>>(b_flag.as<uint32_t>(), n_terms, h->dense_id);
    count_launch();
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail("term metadata");
    h->dense_stride = ((size_t)n_tiles * tile_docs + 7) / 8 * 8;
    if ((uint64_t)h->n_dense * h->dense_stride * 2 >= (1ull << 32)) {
        // the kernel addresses dense rows with 32-bit byte offsets: keep every term as posting buckets instead
        build_term_meta_kernel<<<(n_terms + 1 + bt - 1) / bt, bt, 0, s>>>(d_indptr, d_df, n_terms, n_docs_global, 0xFFFFFFFFu,
                                                                          h->term_keep, h->df_local, b_flag.as<uint32_t>());
        build_dense_id_kernel<<<(n_terms + bt - 1) / bt, bt, 0, s>>>(b_flag.as<uint32_t>(), n_terms, h->dense_id);
        count_launch(2);
        h->n_dense = 0;
    }
    if (h->n_dense) {
        if ((e = cudaMalloc(&h->dense_rows, (size_t)h->n_dense * h->dense_stride * 2 + (size_t)kThreads * 16 + 1024)) != cudaSuccess) return fail("alloc dense rows");
        cudaMemsetAsync(h->dense_rows, 0, (size_t)h->n_dense * h->dense_stride * 2 + (size_t)kThreads * 16 + 1024, s);
        // the kernel loads two slots per thread, the second kThreads slots further on, without a bounds test (the
        // values are dropped): keep those loads inside the allocations
        const size_t pad8 = (size_t)kThreads * 8 + 512;
        if ((e = cudaMalloc(&h->dense_rows8, (size_t)h->n_dense * h->dense_stride + pad8)) != cudaSuccess) return fail("alloc dense rows (u8)");
        cudaMemsetAsync(h->dense_rows8, 0, (size_t)h->n_dense * h->dense_stride + pad8, s);
        if ((e = cudaMalloc(&h->dense_cursor, (size_t)h->n_dense * n_tiles * 4)) != cudaSuccess) return fail("alloc dense cursor");
        cudaMemsetAsync(h->dense_cursor, 0, (size_t)h->n_dense * n_tiles * 4, s);
    }
    if ((e = b_stage.alloc((nnz ? nnz : 1) * 4)) != cudaSuccess) return fail("alloc stage");

    // 1. validate, fill the dense rows, count the other postings per (tile, term) bucket
    const int grid = (int)std::min<uint64_t>((nnz + bt - 1) / bt, 148ull * 32);
    if (nnz) {
        build_count_kernel<<<grid, bt, 0, s>>>(d_indptr, d_docids, d_impacts, nnz, n_terms, n_docs, tile_docs, h->dense_id,
                                               h->dense_rows, h->dense_rows8, h->dense_stride, b_cnt.as<uint32_t>(), h->term_maximp,
                                               b_bad.as<uint32_t>());
        count_launch();
    }
    uint32_t bad = 0;
    if ((e = cudaMemcpyAsync(&bad, b_bad.p, 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail("read flags");
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail("count kernel");
    if (bad) {
        set_error("msr_sparse_index_build: malformed CSR (%s%s)", (bad & 1) ? "docid out of range or impact 0; " : "",
                  (bad & 2) ? "docids not strictly ascending within a term" : "");
        msr_sparse_index_destroy(h);
        return MSR_ERR_INVALID;
    }
    // 2. bucket the postings in doc order into a staging array
    uint32_t* seg0 = b_seg0.as<uint32_t>();
    uint32_t* stage = b_stage.as<uint32_t>();
    if ((e = cub::DeviceScan::ExclusiveSum(b_tmp.p, tmp_bytes, b_cnt.as<uint32_t>(), seg0, n_seg, s)) != cudaSuccess)
        return fail("bucket scan");
    if (nnz) {
        build_scatter_kernel<<<grid, bt, 0, s>>>(d_indptr, d_docids, d_impacts, nnz, n_terms, tile_docs, h->dense_id, seg0, stage,
                                                 h->n_dense, h->dense_cursor);
        count_launch();
    }
    // 3. final layout: slots per bucket (large buckets are padded to whole conflict-free groups), scan, write
    const size_t n_buckets = n_seg - 1;
    const int lgrid = (int)std::min<size_t>((n_buckets + kBuildWarps - 1) / kBuildWarps, 148 * 16);
    uint32_t* slots = b_cnt.as<uint32_t>();          // reuse: counts are no longer needed
    build_plan_kernel<<<lgrid, kBuildWarps * 32, 0, s>>>(seg0, stage, n_buckets, slots);
    count_launch();
    cudaMemsetAsync(slots + n_buckets, 0, 4, s);
    if ((e = cub::DeviceScan::ExclusiveSum(b_tmp.p, tmp_bytes, slots, h->seg_ptr, n_seg, s)) != cudaSuccess)
        return fail("slot scan");
    // total slots in 64 bits: the u32 prefix sum above must not have wrapped
    unsigned long long n_slots64 = 0;
    cudaMemsetAsync(h->work, 0, 16, s);
    sum_u32_kernel<<<148 * 4, 256, 0, s>>>(slots, n_buckets, (unsigned long long*)h->work);
    count_launch();
    if ((e = cudaMemcpyAsync(&n_slots64, h->work, 8, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail("read slots");
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail("plan kernel");
    if (n_slots64 >= (1ull << 32)) {
        set_error("msr_sparse_index_build: %llu posting slots after padding >= 2^32 (shard the collection)", n_slots64);
        msr_sparse_index_destroy(h);
        return MSR_ERR_INVALID;
    }
    h->n_slots = n_slots64;
    if ((e = cudaMalloc(&h->postings, (h->n_slots ? h->n_slots : 1) * 4)) != cudaSuccess) return fail("alloc postings");
    h->resident_bytes = n_seg * 4 + (h->n_slots ? h->n_slots : 1) * 4 + n_terms * 13ull + 20 +
                        (size_t)h->n_dense * h->dense_stride * 3;
    build_layout_kernel<<<lgrid, kBuildWarps * 32, 0, s>>>(seg0, stage, n_buckets, h->seg_ptr, h->postings);
    count_launch();
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail("build kernels");
    if ((e = cudaGetLastError()) != cudaSuccess) return fail("build kernels (launch)");
    cudaFree(h->dense_cursor); h->dense_cursor = nullptr;
    e = cudaFuncSetAttribute(sparse_score_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)search_smem_bytes(kMaxTileDocs));
    if (e != cudaSuccess) return fail("opt-in shared memory");
    *out = h;
    return MSR_OK;
}

extern "C" int msr_sparse_index_destroy(msr_sparse_index* h) {
    if (!h) return MSR_OK;
    cudaSetDevice(h->device);
    cudaFree(h->seg_ptr); cudaFree(h->postings); cudaFree(h->term_keep); cudaFree(h->term_maximp);
    cudaFree(h->df_local); cudaFree(h->status); cudaFree(h->work); cudaFree(h->dense_id); cudaFree(h->dense_rows);
    cudaFree(h->dense_rows8); cudaFree(h->dense_cursor);
    h->ws.release();
    for (cudaEvent_t e : h->ev) cudaEventDestroy(e);
    delete h;
    return MSR_OK;
}

extern "C" int msr_sparse_profile(msr_sparse_index* h, int enable, double* score_ms, double* merge_ms, uint64_t* launches,
                                  void* stream) {
    MSR_REQUIRE(h, "msr_sparse_profile: NULL handle");
    int rc = use_device(h->device);
    if (rc) return rc;
    MSR_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    double a = 0, b = 0;
    for (size_t i = 0; i + 2 < h->ev_used + 0 && i + 2 < h->ev.size() + 0; i += 3) {
        float x = 0, y = 0;
        MSR_CUDA_CHECK(cudaEventElapsedTime(&x, h->ev[i], h->ev[i + 1]));
        MSR_CUDA_CHECK(cudaEventElapsedTime(&y, h->ev[i + 1], h->ev[i + 2]));
        a += x; b += y;
    }
    if (score_ms) *score_ms = a;
    if (merge_ms) *merge_ms = b;
    if (launches) *launches = h->ev_used / 3;
    h->ev_used = 0;
    h->profiling = enable != 0;
    return MSR_OK;
}

extern "C" int msr_sparse_index_info(const msr_sparse_index* h, uint32_t info[9]) {
    MSR_REQUIRE(h && info, "msr_sparse_index_info: NULL argument");
    info[0] = h->n_docs; info[1] = h->n_terms; info[2] = h->n_tiles; info[3] = h->tile_docs;
    info[4] = (uint32_t)h->nnz; info[5] = (uint32_t)(h->nnz >> 32); info[6] = (uint32_t)h->device;
    info[7] = (uint32_t)h->resident_bytes; info[8] = (uint32_t)((uint64_t)h->resident_bytes >> 32);
    return MSR_OK;
}

extern "C" int msr_sparse_index_status(msr_sparse_index* h, void* stream) {
    MSR_REQUIRE(h, "msr_sparse_index_status: NULL handle");
    int rc = use_device(h->device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    uint32_t st = 0;
    MSR_CUDA_CHECK(cudaMemcpyAsync(&st, h->status, 4, cudaMemcpyDeviceToHost, s));
    MSR_CUDA_CHECK(cudaMemsetAsync(h->status, 0, 4, s));
    MSR_CUDA_CHECK(cudaStreamSynchronize(s));
    if (st) {
        set_error("msr_sparse_search (device path): more than %u queries of one batch lie outside the integer kernel's exact range "
                  "(weight > 65535 or sum_t w_t*max_impact(t) >= 2^31) and need the wide path; the first %u were scored, the rows of "
                  "the others are empty -- split the batch or use the host path, which runs as many passes as it takes",
                  kWideMax, kWideMax);
        return MSR_ERR_RANGE;
    }
    return MSR_OK;
}

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

extern "C" int msr_sparse_search(msr_sparse_index* h, uint32_t nq, const uint64_t* q_indptr, const uint32_t* q_terms,
                                 const uint32_t* q_weights, uint32_t k, float* out_scores, int32_t* out_docs,
                                 uint32_t* out_counts, int where, void* stream) {
    MSR_REQUIRE(h, "msr_sparse_search: NULL handle");
    MSR_REQUIRE(where == MSR_HOST || where == MSR_DEVICE, "msr_sparse_search: bad `where`");
    MSR_REQUIRE(k >= 1 && k <= MSR_MAX_K, "msr_sparse_search: k=%u outside [1,%u]", k, MSR_MAX_K);
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(q_indptr && out_scores && out_docs && out_counts, "msr_sparse_search: NULL buffer");
    int rc = use_device(h->device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;

    const uint32_t k_stride = std::min(k, h->tile_docs);
    // queries per launch: bound the candidate buffer (<= 1 GiB) and the 1-D grid
    const size_t per_q = (size_t)h->n_tiles * k_stride * 8 + (size_t)h->n_tiles * 4;
    uint32_t qc = (uint32_t)std::max<size_t>(1, std::min<size_t>(nq, (1ull << 30) / per_q));
    qc = (uint32_t)std::min<uint64_t>(qc, 0x7FFFFFFFull / h->n_tiles);

    // workspace layout: [cand keys | cand cnt | (host mode) q_indptr, q_terms, q_weights, out_scores, out_docs, out_counts]
    uint64_t nnz_q = 0;
    if (where == MSR_HOST) nnz_q = q_indptr[nq];
    size_t off_keys = 0;
    size_t off_cnt = align_up(off_keys + (size_t)qc * h->n_tiles * k_stride * 8, 256);
    size_t off_ok = align_up(off_cnt + (size_t)qc * h->n_tiles * 4, 256);
    size_t off_theta = align_up(off_ok + nq, 256);
    size_t off_tlist = align_up(off_theta + (size_t)qc * 8, 256);
    // wide path (queries outside the exact integer range): list of such queries, their count, kWideMax key arrays
    size_t off_wq = align_up(off_tlist + (size_t)qc * k * 8, 256);
    size_t off_nw = align_up(off_wq + (size_t)nq * 4, 256);
    size_t off_wkeys = align_up(off_nw + 4, 256);
    size_t off_wcnt = align_up(off_wkeys + (size_t)kWideMax * h->n_tiles * h->tile_docs * 8, 256);
    size_t off_end = align_up(off_wcnt + (size_t)kWideMax * h->n_tiles * 4, 256);
    size_t off_qi = off_end, off_qt = 0, off_qw = 0, off_os = 0, off_od = 0, off_oc = 0;
    if (where == MSR_HOST) {
        MSR_REQUIRE(nnz_q == 0 || (q_terms && q_weights), "msr_sparse_search: NULL query arrays");
        off_qt = align_up(off_qi + (size_t)(nq + 1) * 8, 256);
        off_qw = align_up(off_qt + nnz_q * 4, 256);
        off_os = align_up(off_qw + nnz_q * 4, 256);
        off_od = align_up(off_os + (size_t)nq * k * 4, 256);
        off_oc = align_up(off_od + (size_t)nq * k * 4, 256);
        off_end = align_up(off_oc + (size_t)nq * 4, 256);
    }
    MSR_CUDA_CHECK(h->ws.reserve(off_end, s));
    unsigned char* base = (unsigned char*)h->ws.p;
    const uint64_t* d_qi = q_indptr;
    const uint32_t* d_qt = q_terms;
    const uint32_t* d_qw = q_weights;
    float* d_os = out_scores;
    int32_t* d_od = out_docs;
    uint32_t* d_oc = out_counts;
    if (where == MSR_HOST) {
        MSR_CUDA_CHECK(cudaMemcpyAsync(base + off_qi, q_indptr, (size_t)(nq + 1) * 8, cudaMemcpyHostToDevice, s));
        if (nnz_q) {
            MSR_CUDA_CHECK(cudaMemcpyAsync(base + off_qt, q_terms, nnz_q * 4, cudaMemcpyHostToDevice, s));
            MSR_CUDA_CHECK(cudaMemcpyAsync(base + off_qw, q_weights, nnz_q * 4, cudaMemcpyHostToDevice, s));
        }
        d_qi = (const uint64_t*)(base + off_qi); d_qt = (const uint32_t*)(base + off_qt); d_qw = (const uint32_t*)(base + off_qw);
        d_os = (float*)(base + off_os); d_od = (int32_t*)(base + off_od); d_oc = (uint32_t*)(base + off_oc);
    }

    uint8_t* d_ok = (uint8_t*)(base + off_ok);
    uint32_t* d_wq = (uint32_t*)(base + off_wq);
    uint32_t* d_nw = (uint32_t*)(base + off_nw);
    MSR_CUDA_CHECK(cudaMemsetAsync(d_nw, 0, 4, s));
    query_prep_kernel<<<(nq + 7) / 8, 256, 0, s>>>(h->term_keep, h->term_maximp, h->n_terms, nq, d_qi, d_qt, d_qw, d_ok, d_wq, d_nw);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());

    SearchParams P;
    P.seg_ptr = h->seg_ptr; P.postings = h->postings; P.term_keep = h->term_keep; P.term_maximp = h->term_maximp;
    P.n_docs = h->n_docs; P.n_terms = h->n_terms; P.tile_docs = h->tile_docs; P.n_tiles = h->n_tiles;
    P.doc_base = h->doc_base; P.q_indptr = d_qi; P.q_terms = d_qt; P.q_weights = d_qw; P.q_ok = d_ok;
    P.dense_id = h->dense_id; P.dense_rows = h->dense_rows; P.dense_stride = h->dense_stride;
    P.dense_row_bytes = (uint32_t)(h->dense_stride * 2);
    P.dense_rows8 = h->dense_rows8;
    P.k = k; P.k_stride = k_stride;
    P.cand_keys = (uint64_t*)(base + off_keys); P.cand_cnt = (uint32_t*)(base + off_cnt);
    P.theta = (unsigned long long*)(base + off_theta); P.tlist = (uint64_t*)(base + off_tlist);
    const bool shared_theta = where == MSR_DEVICE && h->ext_theta && h->use_theta && nq <= h->ext_theta_cap;
    P.use_theta = ((h->n_tiles > 1 || shared_theta) && h->use_theta) ? 1u : 0u;
    P.n_peer_theta = shared_theta ? h->n_peer_theta : 0u;
    for (uint32_t p = 0; p < P.n_peer_theta; ++p) { P.peer_theta[p] = h->peer_theta[p]; P.peer_quant[p] = h->peer_quant[p]; }
    P.quant = shared_theta ? h->ext_quant : nullptr;
    P.quant_world = h->q_rot_den; P.quant_me = h->q_rot_num; P.quant_cap = h->ext_theta_cap;
    P.quant_rank = P.quant ? (k + h->q_rot_den - 1) / h->q_rot_den : 0u;
    // Re-publishing costs a rank count over c + k keys; it pays while the list is still loose.  Measured (k = 100):
    // 6 tiles k/2 3.47 ms, k/4 3.63, k/8 3.68; 21 tiles 10.36 / 10.45 / 10.70; 41 tiles 19.84 (+0.5 ms final merge) / 19.78 / 19.96.
    P.merge_min = std::max(1u, k >> (h->n_tiles <= 24 ? 1 : 2));
    if (const char* e = getenv("MSR_SPARSE_MERGE_SHIFT")) P.merge_min = std::max(1u, k >> (uint32_t)std::min(10, std::max(0, atoi(e))));
    P.fast_select = h->fast_select ? 1u : 0u;
    P.wide_q = d_wq; P.n_wide = d_nw; P.wide_first = 0; P.wide_ctas = 0; P.status = h->status;
    P.wide_acc = (unsigned long long*)(base + off_wkeys); P.wide_cnt = (uint32_t*)(base + off_wcnt);
    const size_t smem = search_smem_bytes(h->tile_docs);
    for (uint32_t q0 = 0; q0 < nq; q0 += qc) {
        const uint32_t n = std::min(qc, nq - q0);
        P.q_lo = q0; P.nq = n;
        P.q_rot = shared_theta ? (uint32_t)((uint64_t)n * h->q_rot_num / h->q_rot_den) : 0u;
        if (shared_theta) P.theta = h->ext_theta + q0;           // cleared by the caller (it rotates three buffers)
        cudaEvent_t* evs = nullptr;
        if (h->profiling) {
            if (h->ev_used + 3 > h->ev.size()) {
                for (int i = 0; i < 3; ++i) { cudaEvent_t e; MSR_CUDA_CHECK(cudaEventCreate(&e)); h->ev.push_back(e); }
            }
            evs = &h->ev[h->ev_used];
            h->ev_used += 3;
            MSR_CUDA_CHECK(cudaEventRecord(evs[0], s));
        }
        if (P.use_theta)      // theta[] and tlist[] are contiguous up to alignment: one clear
            MSR_CUDA_CHECK(cudaMemsetAsync(base + off_theta, 0, off_tlist - off_theta + (size_t)n * k * 8, s));
        static_assert(sizeof(SearchParams) <= 4000, "kernel parameter space");
        // the batch's last launch carries the wide path as extra CTAs at the end of the grid (they leave at once when the
        // prep kernel listed no query); the merge launch below carries the matching extra CTAs
        const bool last = q0 + n >= nq;
        P.wide_ctas = last ? kWideMax : 0u;
        sparse_score_topk_kernel<<<(n + P.wide_ctas) * h->n_tiles, kThreads, smem, s>>>(P);
        count_launch();
        MSR_CUDA_CHECK(cudaGetLastError());
        if (evs) MSR_CUDA_CHECK(cudaEventRecord(evs[1], s));
        rc = topk_merge_keys_launch(n, h->n_tiles, k_stride, P.cand_keys, P.cand_cnt, k, d_os + (size_t)q0 * k,
                                    d_od + (size_t)q0 * k, d_oc + q0, s, where == MSR_DEVICE ? h->n_peers : 0u,
                                    h->peer_scores, h->peer_docs, h->peer_counts, q0, P.use_theta ? P.theta : nullptr,
                                    P.quant ? P.quant + q0 : nullptr, P.quant_world, P.quant_cap,
                                    P.wide_ctas, 0u, d_wq, d_nw, (const uint64_t*)P.wide_acc, P.wide_cnt, h->n_tiles, h->tile_docs);
        if (rc) return rc;
        if (evs) MSR_CUDA_CHECK(cudaEventRecord(evs[2], s));
    }
    if (where == MSR_HOST) {
        // more than kWideMax queries needed the wide path: the host path knows (it synchronises anyway) and runs the rest
        uint32_t n_wide = 0;
        MSR_CUDA_CHECK(cudaMemcpyAsync(&n_wide, d_nw, 4, cudaMemcpyDeviceToHost, s));
        MSR_CUDA_CHECK(cudaStreamSynchronize(s));
        for (uint32_t first = kWideMax; first < n_wide; first += kWideMax) {
            P.q_lo = 0; P.nq = 0; P.wide_first = first; P.wide_ctas = kWideMax;
            sparse_score_topk_kernel<<<kWideMax * h->n_tiles, kThreads, smem, s>>>(P);
            count_launch();
            MSR_CUDA_CHECK(cudaGetLastError());
            rc = topk_merge_keys_launch(0, h->n_tiles, k_stride, P.cand_keys, P.cand_cnt, k, d_os, d_od, d_oc, s, 0u, h->peer_scores,
                                        h->peer_docs, h->peer_counts, 0, nullptr, nullptr, 0, 0, kWideMax, first, d_wq, d_nw,
                                        (const uint64_t*)P.wide_acc, P.wide_cnt, h->n_tiles, h->tile_docs);
            if (rc) return rc;
        }
        if (n_wide > kWideMax) MSR_CUDA_CHECK(cudaMemsetAsync(h->status, 0, 4, s));     // all of them were scored after all
        MSR_CUDA_CHECK(cudaMemcpyAsync(out_scores, d_os, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, s));
        MSR_CUDA_CHECK(cudaMemcpyAsync(out_docs, d_od, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, s));
        MSR_CUDA_CHECK(cudaMemcpyAsync(out_counts, d_oc, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
        return msr_sparse_index_status(h, stream);   // synchronises
    }
    return MSR_OK;
}

extern "C" int msr_sparse_query_work(msr_sparse_index* h, uint32_t nq, const uint64_t* q_indptr, const uint32_t* q_terms,
                                     const uint32_t* q_weights, uint64_t* postings, uint64_t* kept_terms, int where,
                                     void* stream) {
    MSR_REQUIRE(h && postings && kept_terms, "msr_sparse_query_work: NULL argument");
    MSR_REQUIRE(where == MSR_HOST || where == MSR_DEVICE, "msr_sparse_query_work: bad `where`");
    *postings = 0; *kept_terms = 0;
    if (nq == 0) return MSR_OK;
    int rc = use_device(h->device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    uint64_t nnz_q = 0;
    DevBuf bt, bw;
    const uint32_t* d_qt = q_terms;
    const uint32_t* d_qw = q_weights;
    if (where == MSR_HOST) {
        nnz_q = q_indptr[nq];
        if (nnz_q == 0) return MSR_OK;
        MSR_CUDA_CHECK(bt.alloc(nnz_q * 4));
        MSR_CUDA_CHECK(bw.alloc(nnz_q * 4));
        MSR_CUDA_CHECK(cudaMemcpyAsync(bt.p, q_terms, nnz_q * 4, cudaMemcpyHostToDevice, s));
        MSR_CUDA_CHECK(cudaMemcpyAsync(bw.p, q_weights, nnz_q * 4, cudaMemcpyHostToDevice, s));
        d_qt = bt.as<uint32_t>(); d_qw = bw.as<uint32_t>();
    } else {
        MSR_CUDA_CHECK(cudaMemcpyAsync(&nnz_q, q_indptr + nq, 8, cudaMemcpyDeviceToHost, s));
        MSR_CUDA_CHECK(cudaStreamSynchronize(s));
        if (nnz_q == 0) return MSR_OK;
    }
    MSR_CUDA_CHECK(cudaMemsetAsync(h->work, 0, 16, s));
    const int grid = (int)std::min<uint64_t>((nnz_q + 255) / 256, 148ull * 8);
    query_work_kernel<<<grid, 256, 0, s>>>(h->term_keep, h->df_local, h->n_terms, nnz_q, d_qt, d_qw,
                                           (unsigned long long*)h->work);
    count_launch();
    uint64_t res[2] = {0, 0};
    MSR_CUDA_CHECK(cudaMemcpyAsync(res, h->work, 16, cudaMemcpyDeviceToHost, s));
    MSR_CUDA_CHECK(cudaStreamSynchronize(s));
    *postings = res[0]; *kept_terms = res[1];
    return MSR_OK;
}

extern "C" int msr_sparse_query_traffic(msr_sparse_index* h, uint32_t nq, const uint64_t* q_indptr, const uint32_t* q_terms,
                                        const uint32_t* q_weights, uint64_t* bytes, void* stream) {
    MSR_REQUIRE(h && bytes, "msr_sparse_query_traffic: NULL argument");
    *bytes = 0;
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(q_indptr && q_terms && q_weights, "msr_sparse_query_traffic: NULL query arrays");
    int rc = use_device(h->device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    uint64_t nnz_q = 0;
    MSR_CUDA_CHECK(cudaMemcpyAsync(&nnz_q, q_indptr + nq, 8, cudaMemcpyDeviceToHost, s));
    MSR_CUDA_CHECK(cudaStreamSynchronize(s));
    if (nnz_q == 0) return MSR_OK;
    DevBuf flags;
    MSR_CUDA_CHECK(flags.alloc((size_t)h->n_terms * 4));
    MSR_CUDA_CHECK(cudaMemsetAsync(flags.p, 0, (size_t)h->n_terms * 4, s));
    MSR_CUDA_CHECK(cudaMemsetAsync(h->work, 0, 16, s));
    traffic_mark_kernel<<<(int)std::min<uint64_t>((nnz_q + 255) / 256, 148ull * 8), 256, 0, s>>>(
        h->term_keep, h->term_maximp, h->n_terms, nnz_q, q_terms, q_weights, flags.as<uint32_t>());
    traffic_sum_kernel<<<(h->n_terms + 255) / 256, 256, 0, s>>>(flags.as<uint32_t>(), h->n_terms, h->n_tiles, h->n_docs,
                                                                h->seg_ptr, h->dense_id, (unsigned long long*)h->work);
    count_launch(2);
    uint64_t res = 0;
    MSR_CUDA_CHECK(cudaMemcpyAsync(&res, h->work, 8, cudaMemcpyDeviceToHost, s));
    MSR_CUDA_CHECK(cudaStreamSynchronize(s));
    MSR_CUDA_CHECK(cudaGetLastError());
    *bytes = res + (uint64_t)nnz_q * 8 + (uint64_t)(nq + 1) * 8;   // + the query CSR itself
    return MSR_OK;
}

extern "C" int msr_sparse_set_peer_theta(msr_sparse_index* h, uint32_t n_peers, uint64_t* local_theta,
                                         uint64_t* const* peer_theta, uint32_t capacity, uint32_t rot_num, uint32_t rot_den,
                                         uint64_t* local_quant, uint64_t* const* peer_quant) {
    MSR_REQUIRE(h, "msr_sparse_set_peer_theta: NULL handle");
    MSR_REQUIRE(n_peers <= 15, "msr_sparse_set_peer_theta: at most 15 peers");
    if (!local_theta) {                                          // switch the sharing off
        h->ext_theta = nullptr; h->ext_quant = nullptr; h->n_peer_theta = 0; h->ext_theta_cap = 0; h->q_rot_num = 0; h->q_rot_den = 1;
        return MSR_OK;
    }
    MSR_REQUIRE(capacity > 0 && rot_den > 0 && rot_num < rot_den, "msr_sparse_set_peer_theta: bad capacity / rotation");
    MSR_REQUIRE(n_peers == 0 || peer_theta, "msr_sparse_set_peer_theta: NULL pointer table");
    for (uint32_t p = 0; p < n_peers; ++p) {
        MSR_REQUIRE(peer_theta[p], "msr_sparse_set_peer_theta: NULL peer pointer %u", p);
        h->peer_theta[p] = (unsigned long long*)peer_theta[p];
        MSR_REQUIRE(!local_quant || (peer_quant && peer_quant[p]), "msr_sparse_set_peer_theta: NULL peer quantile pointer %u", p);
        h->peer_quant[p] = local_quant ? (unsigned long long*)peer_quant[p] : nullptr;
    }
    MSR_REQUIRE(!local_quant || n_peers + 1 == rot_den, "msr_sparse_set_peer_theta: quantile sharing needs every GPU of the group as a peer");
    h->ext_quant = (unsigned long long*)local_quant;
    h->ext_theta = (unsigned long long*)local_theta; h->n_peer_theta = n_peers; h->ext_theta_cap = capacity;
    h->q_rot_num = rot_num; h->q_rot_den = rot_den;
    return MSR_OK;
}

extern "C" int msr_sparse_set_peer_outputs(msr_sparse_index* h, uint32_t n_peers, float* const* peer_scores,
                                           int32_t* const* peer_docs, uint32_t* const* peer_counts) {
    MSR_REQUIRE(h, "msr_sparse_set_peer_outputs: NULL handle");
    MSR_REQUIRE(n_peers <= 16, "msr_sparse_set_peer_outputs: at most 16 peers");
    MSR_REQUIRE(n_peers == 0 || (peer_scores && peer_docs && peer_counts), "msr_sparse_set_peer_outputs: NULL pointer table");
    for (uint32_t p = 0; p < n_peers; ++p) {
        MSR_REQUIRE(peer_scores[p] && peer_docs[p] && peer_counts[p], "msr_sparse_set_peer_outputs: NULL peer pointer %u", p);
        h->peer_scores[p] = peer_scores[p]; h->peer_docs[p] = peer_docs[p]; h->peer_counts[p] = peer_counts[p];
    }
    h->n_peers = n_peers;
    return MSR_OK;
}
