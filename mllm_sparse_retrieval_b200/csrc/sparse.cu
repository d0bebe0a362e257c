// sparse.cu -- kernels (a)+(b): tiled inverted-index impact scorer with the top-k fused into its epilogue.
//
// Replaces pyserini LuceneImpactSearcher.batch_search (reference call sites src/search.py:86-87, 273-275) under
// the rules S1-S10 of SURVEY.md 8(a).
//
// HBM layout (built once by msr_sparse_index_build):
//   * the doc id space of the shard is cut into n_tiles contiguous tiles of tile_docs docs (<= 65536);
//   * postings are bucketed by (tile, term) and packed into ONE u32 each: (impact << 16) | (doc - tile_lo),
//     ascending doc inside a bucket.  4 B/posting physical against the canonical 6 B (u32 docid + u16 impact)
//     that SURVEY.md 8(d) fixes as the algorithmic figure;
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

#include <vector>

#include "common.cuh"

namespace msr {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kSegMax = 256;      // query terms staged per pass
constexpr int kUnroll = 8;        // loads in flight per lane
constexpr int kChunk = 32 * kUnroll;
constexpr int kBins = 2048;
constexpr int kBinBits = 11;
constexpr int kCandMax = 1024;
constexpr uint32_t kDefaultTileDocs = 24576;   // 96 KB of accumulators -> 2 CTAs per SM
constexpr uint32_t kMaxTileDocs = 53248;       // 208 KB of accumulators + staging = the 227 KB opt-in limit

int topk_merge_keys_launch(uint32_t nq, uint32_t n_lists, uint32_t stride, const uint64_t* keys,
                           const uint32_t* cnts, uint32_t k, float* out_scores, int32_t* out_docs,
                           uint32_t* out_counts, cudaStream_t s);

}  // namespace msr

struct msr_sparse_index {
    int device = 0;
    uint32_t n_docs = 0, n_terms = 0, tile_docs = 0, n_tiles = 0, n_docs_global = 0;
    int32_t doc_base = 0;
    uint64_t nnz = 0;
    uint32_t* seg_ptr = nullptr;      // [n_tiles*n_terms + 1]
    uint32_t* postings = nullptr;     // [nnz]
    uint8_t* term_keep = nullptr;     // [n_terms]
    uint16_t* term_maximp = nullptr;  // [n_terms]
    uint32_t* df_local = nullptr;     // [n_terms]
    uint32_t* status = nullptr;       // sticky device status word
    uint64_t* work = nullptr;         // [2] scratch for msr_sparse_query_work
    msr::Workspace ws;
    size_t resident_bytes = 0;
    // optional per-kernel timing (msr_sparse_profile): events bracket every score / merge launch on the stream
    bool profiling = false;
    std::vector<cudaEvent_t> ev;      // triples: before score, after score, after merge
    size_t ev_used = 0;
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

__global__ void build_count_kernel(const uint64_t* __restrict__ indptr, const uint32_t* __restrict__ docids,
                                   const uint16_t* __restrict__ impacts, uint64_t nnz, uint32_t n_terms,
                                   uint32_t n_docs, uint32_t tile_docs, uint32_t* __restrict__ cnt,
                                   uint32_t* __restrict__ maximp32, uint32_t* __restrict__ bad) {
    for (uint64_t p = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; p < nnz; p += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t t = term_of_posting(indptr, n_terms, p);
        uint32_t d = docids[p];
        uint32_t imp = impacts[p];
        if (d >= n_docs || imp == 0) { atomicOr(bad, 1u); continue; }
        if (p > indptr[t] && docids[p - 1] >= d) atomicOr(bad, 2u);
        atomicAdd(&cnt[(size_t)(d / tile_docs) * n_terms + t], 1u);
        atomicMax(&maximp32[t], imp);
    }
}

__global__ void build_scatter_kernel(const uint64_t* __restrict__ indptr, const uint32_t* __restrict__ docids,
                                     const uint16_t* __restrict__ impacts, uint64_t nnz, uint32_t n_terms,
                                     uint32_t tile_docs, const uint32_t* __restrict__ seg_ptr,
                                     uint32_t* __restrict__ postings) {
    for (uint64_t p = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; p < nnz; p += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t t = term_of_posting(indptr, n_terms, p);
        uint32_t d = docids[p];
        uint32_t tile = d / tile_docs;
        uint32_t tile_lo = tile * tile_docs;
        // first posting of term t whose doc >= tile_lo
        uint64_t lo = indptr[t], hi = p;   // docids[p] >= tile_lo, so the answer is in [lo, p]
        while (lo < hi) {
            uint64_t mid = (lo + hi) >> 1;
            if (docids[mid] < tile_lo) lo = mid + 1; else hi = mid;
        }
        uint32_t r = (uint32_t)(p - lo);
        postings[seg_ptr[(size_t)tile * n_terms + t] + r] = ((uint32_t)impacts[p] << 16) | (d - tile_lo);
    }
}

__global__ void build_term_meta_kernel(const uint64_t* __restrict__ indptr, const uint32_t* __restrict__ df_global,
                                       const uint32_t* __restrict__ maximp32, uint32_t n_terms, uint32_t n_docs_global,
                                       uint8_t* __restrict__ keep, uint16_t* __restrict__ maximp, uint32_t* __restrict__ df_local) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_terms) return;
    uint32_t dfl = (uint32_t)(indptr[t + 1] - indptr[t]);
    uint64_t df = df_global ? df_global[t] : dfl;
    // S5: in the dictionary and ln(N/(df+1)) > 0  <=>  0 < df and df + 1 < N
    keep[t] = (df > 0 && df + 1 < (uint64_t)n_docs_global) ? 1 : 0;
    maximp[t] = (uint16_t)maximp32[t];
    df_local[t] = dfl;
}

// ---------------------------------------------------------------------------------------------------------
// search kernel
// ---------------------------------------------------------------------------------------------------------
struct SearchParams {
    const uint32_t* seg_ptr;
    const uint32_t* postings;
    const uint8_t* term_keep;
    const uint16_t* term_maximp;
    uint32_t n_docs, n_terms, tile_docs, n_tiles;
    int32_t doc_base;
    const uint64_t* q_indptr;
    const uint32_t* q_terms;
    const uint32_t* q_weights;
    uint32_t q_lo, nq, k, k_stride;
    uint64_t* cand_keys;   // [nq][n_tiles][k_stride]
    uint32_t* cand_cnt;    // [nq][n_tiles]
    uint32_t* status;
};

__device__ __forceinline__ uint32_t ld_stream_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

struct SegStage {                      // streaming phase view of the union region
    uint32_t start[kSegMax];
    uint32_t len[kSegMax];
    uint32_t w[kSegMax];               // bit 31 = "product may reach 2^24: round per term" (S6)
    uint32_t pref[kSegMax + 1];        // exclusive prefix of chunk counts
};
struct SelStage {                      // epilogue view of the union region
    uint32_t hist[kBins];
    uint64_t cand[kCandMax];
};
union StageUnion {
    SegStage seg;
    SelStage sel;
};
struct SmallState {
    unsigned long long bound;
    uint32_t wmax, vmax, hits, out_cnt, cand_cnt, sel_b, sel_clt, sel_bcount;
    uint32_t wsum[kWarps];
};

// exclusive scan of one value per thread over the CTA (kThreads threads); returns exclusive prefix, *total = sum
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* wsum, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = warp_incl_scan(v, lane);
    __syncthreads();                       // wsum may still be read from a previous call
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
        uint32_t x = wsum[w];
        if (w < warp) base += x;
        tot += x;
    }
    *total = tot;
    return base + inc - v;
}

__global__ void __launch_bounds__(kThreads, 2)
sparse_score_topk_kernel(const SearchParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    StageUnion* U = reinterpret_cast<StageUnion*>(smem_raw);
    SmallState* S = reinterpret_cast<SmallState*>(smem_raw + sizeof(StageUnion));
    uint32_t* acc = reinterpret_cast<uint32_t*>(smem_raw + sizeof(StageUnion) + sizeof(SmallState));

    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t q = blockIdx.x % P.nq;
    const uint32_t tile = blockIdx.x / P.nq;
    const uint32_t tile_lo = tile * P.tile_docs;
    const uint32_t tl = min(P.tile_docs, P.n_docs - tile_lo);
    const uint64_t qs = P.q_indptr[P.q_lo + q], qe = P.q_indptr[P.q_lo + q + 1];
    const uint32_t* seg_row = P.seg_ptr + (size_t)tile * P.n_terms;
    uint64_t* keys_out = P.cand_keys + ((size_t)q * P.n_tiles + tile) * P.k_stride;
    uint32_t* cnt_out = P.cand_cnt + (size_t)q * P.n_tiles + tile;

    for (uint32_t i = tid; i < tl; i += kThreads) acc[i] = 0;
    if (tid == 0) {
        S->bound = 0; S->wmax = 0; S->vmax = 0; S->hits = 0; S->out_cnt = 0; S->cand_cnt = 0;
    }
    __syncthreads();

    // ---- exact-arithmetic range check over the whole query (see msr.h) ----
    {
        unsigned long long b = 0;
        uint32_t wm = 0;
        for (uint64_t j = qs + tid; j < qe; j += kThreads) {
            uint32_t t = P.q_terms[j], w = P.q_weights[j];
            if (t < P.n_terms && w && P.term_keep[t]) {
                b += (unsigned long long)w * P.term_maximp[t];
                wm = max(wm, w);
            }
        }
        if (b) atomicAdd(&S->bound, b);
        if (wm) atomicMax(&S->wmax, wm);
    }
    __syncthreads();
    if (S->bound >= (1ull << 31) || S->wmax > 65535u) {
        if (tid == 0) { atomicOr(P.status, 1u); *cnt_out = 0; }
        return;
    }

    // ---- phase 1: stream the query's buckets of this tile, accumulate integer impacts ----
    for (uint64_t batch = qs; batch < qe; batch += kSegMax) {
        __syncthreads();   // previous batch fully consumed
        uint32_t nchunks = 0;
        if (tid < kSegMax) {
            uint32_t len = 0, start = 0, w = 0;
            uint64_t j = batch + tid;
            if (j < qe) {
                uint32_t t = P.q_terms[j];
                w = P.q_weights[j];
                if (t < P.n_terms && w && P.term_keep[t]) {
                    start = seg_row[t];
                    len = seg_row[t + 1] - start;
                    if ((uint64_t)w * P.term_maximp[t] >= (1u << 24)) w |= 0x80000000u;
                }
            }
            U->seg.start[tid] = start;
            U->seg.len[tid] = len;
            U->seg.w[tid] = w;
            nchunks = (len + kChunk - 1) / kChunk;
        }
        uint32_t total;
        uint32_t ex = block_excl_scan(nchunks, S->wsum, &total);
        if (tid < kSegMax) U->seg.pref[tid] = ex;
        if (tid == 0) U->seg.pref[kSegMax] = total;
        __syncthreads();

        uint32_t seg = 0;
        for (uint32_t g = warp; g < total; g += kWarps) {
            while (U->seg.pref[seg + 1] <= g) ++seg;          // monotone in g; empty buckets are skipped
            const uint32_t off = (g - U->seg.pref[seg]) * kChunk;
            const uint32_t n = min((uint32_t)kChunk, U->seg.len[seg] - off);
            const uint32_t* src = P.postings + U->seg.start[seg] + off + lane;
            const uint32_t wraw = U->seg.w[seg];
            const uint32_t w = wraw & 0x7FFFFFFFu;
            uint32_t v[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) v[u] = (u * 32 + lane < n) ? ld_stream_u32(src + u * 32) : 0u;
            if (!(wraw >> 31)) {
#pragma unroll
                for (int u = 0; u < kUnroll; ++u)
                    if (u * 32 + lane < n) atomicAdd(&acc[v[u] & 0xFFFFu], (v[u] >> 16) * w);
            } else {                                            // S6: float32(boost*freq) per term
#pragma unroll
                for (int u = 0; u < kUnroll; ++u)
                    if (u * 32 + lane < n) atomicAdd(&acc[v[u] & 0xFFFFu], (uint32_t)__uint2float_rn((v[u] >> 16) * w));
            }
        }
    }
    __syncthreads();

    // ---- phase 2: top-k of the tile by (float32(score) desc, doc asc) ----
    {
        uint32_t m = 0, h = 0;
        for (uint32_t i = tid; i < tl; i += kThreads) {
            uint32_t a = acc[i];
            m = max(m, a);
            h += (a != 0);
        }
        m = warp_reduce_max(m);
        h = warp_reduce_add(h);
        if (lane == 0) {
            if (m) atomicMax(&S->vmax, m);
            if (h) atomicAdd(&S->hits, h);
        }
    }
    __syncthreads();
    const uint32_t vmax = S->vmax, hits = S->hits;
    const uint32_t kq = min(P.k, tl);
    const uint32_t gdoc0 = (uint32_t)P.doc_base + tile_lo;
    if (hits == 0) {
        if (tid == 0) *cnt_out = 0;
        return;
    }
    if (hits <= kq) {                                           // S7: fewer hits than k -> all of them
        for (uint32_t i = tid; i < tl; i += kThreads) {
            uint32_t a = acc[i];
            if (a) keys_out[atomicAdd(&S->out_cnt, 1u)] = make_key(__uint2float_rn(a), gdoc0 + i);
        }
        if (tid == 0) *cnt_out = hits;
        return;
    }
    // selection key: integers below 2^24 are their own float32 image; above, rank by the rounded float (S6)
    const bool use_f = vmax >= (1u << 24);
#define MSR_SELKEY(a) (use_f ? __float_as_uint(__uint2float_rn(a)) : (a))
    const uint32_t kmax = MSR_SELKEY(vmax);
    uint32_t remaining = kq, sh_prev = 32 - __clz(kmax), pref = 0, s = 0, cand_count = 0;
    for (;;) {
        s = sh_prev > kBinBits ? sh_prev - kBinBits : 0;
        const uint32_t width = sh_prev - s;
        for (uint32_t i = tid; i < kBins; i += kThreads) U->sel.hist[i] = 0;
        __syncthreads();
        for (uint32_t i = tid; i < tl; i += kThreads) {
            uint32_t a = acc[i];
            if (a) {
                uint32_t d = kmax - MSR_SELKEY(a);
                uint32_t hi = sh_prev >= 32 ? 0u : (d >> sh_prev);
                if (hi == pref) atomicAdd(&U->sel.hist[(d >> s) & ((1u << width) - 1u)], 1u);
            }
        }
        __syncthreads();
        {
            uint32_t h4[4], ls = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) { h4[j] = U->sel.hist[tid * 4 + j]; ls += h4[j]; }
            uint32_t total;
            uint32_t ex = block_excl_scan(ls, S->wsum, &total);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (ex < remaining && ex + h4[j] >= remaining) {
                    S->sel_b = tid * 4 + j; S->sel_clt = ex; S->sel_bcount = h4[j];
                }
                ex += h4[j];
            }
        }
        __syncthreads();
        remaining -= S->sel_clt;
        pref = (width >= 32 ? 0u : (pref << width)) | S->sel_b;
        sh_prev = s;
        cand_count = S->sel_bcount;
        if (cand_count <= (uint32_t)kCandMax || s == 0) break;
    }
    const uint32_t n_def = kq - remaining;                      // docs strictly better than the candidate bin
    const bool collect = cand_count <= (uint32_t)kCandMax;
    for (uint32_t i = tid; i < tl; i += kThreads) {
        uint32_t a = acc[i];
        if (a) {
            uint32_t d = kmax - MSR_SELKEY(a);
            uint32_t x = d >> s;
            if (x < pref) {
                keys_out[atomicAdd(&S->out_cnt, 1u)] = make_key(__uint2float_rn(a), gdoc0 + i);
            } else if (x == pref && collect) {
                // descending sort of ~((d<<32)|i)  ==  ascending (d, i)  ==  best first
                U->sel.cand[atomicAdd(&S->cand_cnt, 1u)] = ~(((uint64_t)d << 32) | i);
            }
        }
    }
    __syncthreads();
    if (collect) {
        uint32_t n2 = 2;
        while (n2 < cand_count) n2 <<= 1;
        for (uint32_t i = cand_count + tid; i < n2; i += kThreads) U->sel.cand[i] = 0;
        bitonic_sort_desc(U->sel.cand, n2);
        for (uint32_t r = tid; r < remaining; r += kThreads) {
            uint32_t i = (uint32_t)(~U->sel.cand[r]);
            keys_out[n_def + r] = make_key(__uint2float_rn(acc[i]), gdoc0 + i);
        }
    } else {
        // more than kCandMax docs tie exactly at the threshold key: take the `remaining` smallest doc ids (S8)
        const uint32_t per = ((tl + kWarps - 1) / kWarps + 31) & ~31u;
        const uint32_t lo = warp * per, hi = min(tl, lo + per);
        uint32_t c = 0;
        for (uint32_t base = lo; base < hi; base += 32) {
            uint32_t i = base + lane;
            uint32_t a = i < hi ? acc[i] : 0;
            bool f = a && (kmax - MSR_SELKEY(a)) == pref;
            c += __popc(__ballot_sync(0xffffffffu, f));
        }
        __syncthreads();
        if (lane == 0) S->wsum[warp] = c;
        __syncthreads();
        uint32_t run = 0;
        for (uint32_t w2 = 0; w2 < warp; ++w2) run += S->wsum[w2];
        for (uint32_t base = lo; base < hi; base += 32) {
            uint32_t i = base + lane;
            uint32_t a = i < hi ? acc[i] : 0;
            bool f = a && (kmax - MSR_SELKEY(a)) == pref;
            uint32_t b = __ballot_sync(0xffffffffu, f);
            uint32_t r = run + __popc(b & ((1u << lane) - 1u));
            if (f && r < remaining) keys_out[n_def + r] = make_key(__uint2float_rn(a), gdoc0 + i);
            run += __popc(b);
        }
    }
#undef MSR_SELKEY
    if (tid == 0) *cnt_out = kq;
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

static size_t search_smem_bytes(uint32_t tile_docs) {
    return sizeof(StageUnion) + sizeof(SmallState) + (size_t)tile_docs * 4;
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
        tile_docs = (tile_docs + 31) & ~31u;
    }
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
    const size_t n_seg = (size_t)n_tiles * n_terms + 1;
    DevBuf b_cnt, b_max32, b_bad, b_tmp;
    cudaError_t e = cudaSuccess;
    auto fail = [&](const char* what) {
        set_error("msr_sparse_index_build: %s: %s", what, cudaGetErrorString(e));
        msr_sparse_index_destroy(h);
        return (int)MSR_ERR_CUDA;
    };
    if ((e = cudaMalloc(&h->seg_ptr, n_seg * 4)) != cudaSuccess) return fail("alloc seg_ptr");
    if ((e = cudaMalloc(&h->postings, (nnz ? nnz : 1) * 4)) != cudaSuccess) return fail("alloc postings");
    if ((e = cudaMalloc(&h->term_keep, n_terms)) != cudaSuccess) return fail("alloc term_keep");
    if ((e = cudaMalloc(&h->term_maximp, (size_t)n_terms * 2)) != cudaSuccess) return fail("alloc term_maximp");
    if ((e = cudaMalloc(&h->df_local, (size_t)n_terms * 4)) != cudaSuccess) return fail("alloc df_local");
    if ((e = cudaMalloc(&h->status, 4)) != cudaSuccess) return fail("alloc status");
    if ((e = cudaMalloc(&h->work, 16)) != cudaSuccess) return fail("alloc work");
    h->resident_bytes = n_seg * 4 + (nnz ? nnz : 1) * 4 + n_terms * 7ull + 20;
    if ((e = b_cnt.alloc(n_seg * 4)) != cudaSuccess) return fail("alloc cnt");
    if ((e = b_max32.alloc((size_t)n_terms * 4)) != cudaSuccess) return fail("alloc maximp32");
    if ((e = b_bad.alloc(4)) != cudaSuccess) return fail("alloc bad");
    cudaMemsetAsync(b_cnt.p, 0, n_seg * 4, s);
    cudaMemsetAsync(b_max32.p, 0, (size_t)n_terms * 4, s);
    cudaMemsetAsync(b_bad.p, 0, 4, s);
    cudaMemsetAsync(h->status, 0, 4, s);

    const int bt = 256;
    const int grid = (int)std::min<uint64_t>((nnz + bt - 1) / bt, 148ull * 32);
    if (nnz) {
        build_count_kernel<<<grid, bt, 0, s>>>(d_indptr, d_docids, d_impacts, nnz, n_terms, n_docs, tile_docs,
                                               b_cnt.as<uint32_t>(), b_max32.as<uint32_t>(), b_bad.as<uint32_t>());
        count_launch();
    }
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, b_cnt.as<uint32_t>(), h->seg_ptr, n_seg, s);
    if ((e = b_tmp.alloc(tmp_bytes)) != cudaSuccess) return fail("alloc scan temp");
    if ((e = cub::DeviceScan::ExclusiveSum(b_tmp.p, tmp_bytes, b_cnt.as<uint32_t>(), h->seg_ptr, n_seg, s)) != cudaSuccess)
        return fail("bucket scan");
    if (nnz) {
        build_scatter_kernel<<<grid, bt, 0, s>>>(d_indptr, d_docids, d_impacts, nnz, n_terms, tile_docs, h->seg_ptr, h->postings);
        count_launch();
    }
    build_term_meta_kernel<<<(n_terms + bt - 1) / bt, bt, 0, s>>>(d_indptr, d_df, b_max32.as<uint32_t>(), n_terms,
                                                                  n_docs_global, h->term_keep, h->term_maximp, h->df_local);
    count_launch();
    uint32_t bad = 0;
    if ((e = cudaMemcpyAsync(&bad, b_bad.p, 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail("read flags");
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail("build kernels");
    if ((e = cudaGetLastError()) != cudaSuccess) return fail("build kernels (launch)");
    if (bad) {
        set_error("msr_sparse_index_build: malformed CSR (%s%s)", (bad & 1) ? "docid out of range or impact 0; " : "",
                  (bad & 2) ? "docids not strictly ascending within a term" : "");
        msr_sparse_index_destroy(h);
        return MSR_ERR_INVALID;
    }
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
    cudaFree(h->df_local); cudaFree(h->status); cudaFree(h->work);
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
        set_error("msr_sparse_search: a query left the exact-arithmetic range (weight > 65535 or "
                  "sum_t w_t*max_impact(t) >= 2^31); its row was returned empty");
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
    size_t off_end = align_up(off_cnt + (size_t)qc * h->n_tiles * 4, 256);
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

    SearchParams P;
    P.seg_ptr = h->seg_ptr; P.postings = h->postings; P.term_keep = h->term_keep; P.term_maximp = h->term_maximp;
    P.n_docs = h->n_docs; P.n_terms = h->n_terms; P.tile_docs = h->tile_docs; P.n_tiles = h->n_tiles;
    P.doc_base = h->doc_base; P.q_indptr = d_qi; P.q_terms = d_qt; P.q_weights = d_qw;
    P.k = k; P.k_stride = k_stride;
    P.cand_keys = (uint64_t*)(base + off_keys); P.cand_cnt = (uint32_t*)(base + off_cnt); P.status = h->status;
    const size_t smem = search_smem_bytes(h->tile_docs);
    for (uint32_t q0 = 0; q0 < nq; q0 += qc) {
        const uint32_t n = std::min(qc, nq - q0);
        P.q_lo = q0; P.nq = n;
        cudaEvent_t* evs = nullptr;
        if (h->profiling) {
            if (h->ev_used + 3 > h->ev.size()) {
                for (int i = 0; i < 3; ++i) { cudaEvent_t e; MSR_CUDA_CHECK(cudaEventCreate(&e)); h->ev.push_back(e); }
            }
            evs = &h->ev[h->ev_used];
            h->ev_used += 3;
            MSR_CUDA_CHECK(cudaEventRecord(evs[0], s));
        }
        sparse_score_topk_kernel<<<n * h->n_tiles, kThreads, smem, s>>>(P);
        count_launch();
        MSR_CUDA_CHECK(cudaGetLastError());
        if (evs) MSR_CUDA_CHECK(cudaEventRecord(evs[1], s));
        rc = topk_merge_keys_launch(n, h->n_tiles, k_stride, P.cand_keys, P.cand_cnt, k, d_os + (size_t)q0 * k,
                                    d_od + (size_t)q0 * k, d_oc + q0, s);
        if (rc) return rc;
        if (evs) MSR_CUDA_CHECK(cudaEventRecord(evs[2], s));
    }
    if (where == MSR_HOST) {
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
