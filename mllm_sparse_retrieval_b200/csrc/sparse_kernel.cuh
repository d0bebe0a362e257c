// sparse_kernel.cuh -- device code of the sparse impact search (included by sparse.cu only).
//
//   query_prep_kernel         once per query batch: S5 term filter, S6 per-term rounding flag, exact-range check.
//   sparse_score_topk_kernel  one CTA per (query, doc tile): stream the query's posting buckets of the tile,
//                             accumulate integer impacts in shared memory, select the tile's candidates.
//
// Candidate selection uses a per-query THRESHOLD shared between the tiles of a query through global memory:
//   theta[q]    64-bit ranking key; invariant: at least k distinct docs of the shard have key >= theta[q].
//   tlist[q][k] keys of k such docs (any mix of real doc keys; readers de-duplicate).
// The first tile of a query (grid order is tile-major, so it finishes long before the others start) radix-selects
// its exact top-k and publishes it.  Every later tile emits only the docs whose key reaches theta -- one pass over
// the accumulators -- then merges them with tlist, re-publishes the best k and raises theta (atomicMax).  Any doc of
// the final top-k has key >= the true k-th key >= theta, so it is always emitted: the result is exact and
// independent of timing; only the amount of surplus candidates varies.  A tile that finds no theta yet, or more
// candidates than its output slots, falls back to the exact radix select.  8-byte stores are single transactions,
// so a reader racing a writer sees old or new keys, each a real doc: no locks, nothing waits on another CTA.
#pragma once
#include "common.cuh"
#include "sparse_dense_mma.cuh"

namespace msr {

#ifndef MSR_SPARSE_THREADS
#define MSR_SPARSE_THREADS 512
#endif
#ifndef MSR_SPARSE_MIN_CTAS
#define MSR_SPARSE_MIN_CTAS 2
#endif
constexpr int kThreads = MSR_SPARSE_THREADS;
constexpr int kWarps = kThreads / 32;
constexpr int kSegMax = 256;      // query terms staged per pass
constexpr int kUnroll = 8;        // loads in flight per lane
constexpr int kChunk = 32 * kUnroll;
constexpr int kBins = 2048;
constexpr int kBinBits = 11;
constexpr int kCandMax = 1024;
constexpr uint32_t kNotDense = 0xFFFFFFFFu;   // dense_id of a term that is stored as posting buckets
constexpr int kUnionMax = 2048;   // keys the threshold merge can sort (own candidates + k)
constexpr int kPer = 2048 / kThreads;   // bins / union keys per thread in the CTA-wide scans

struct SearchParams {
    const uint32_t* seg_ptr;
    const uint32_t* postings;
    uint32_t n_docs, n_terms, tile_docs, n_tiles;
    int32_t doc_base;
    const uint64_t* q_indptr;
    const uint32_t* q_terms;
    const uint32_t* q_weights;
    const uint8_t* term_keep;
    const uint32_t* term_maximp;
    const uint32_t* dense_id;     // [n_terms]
    const uint16_t* dense_rows;   // [n_dense][dense_stride]
    const uint8_t* dense_rows8;   // [n_dense][dense_stride] min(impact, 255); the remainder is a bucket posting
    size_t dense_stride;
    uint32_t dense_row_bytes;     // dense_stride * 2 (all rows together stay below 4 GiB)
    const uint8_t* q_ok;          // [nq_total] 1 = inside the exact-arithmetic range
    uint32_t q_lo, nq, k, k_stride;
    uint64_t* cand_keys;          // [nq][n_tiles][k_stride]
    uint32_t* cand_cnt;           // [nq][n_tiles]
    unsigned long long* theta;    // [nq]
    uint64_t* tlist;              // [nq][k]
    uint32_t use_theta;
    uint32_t merge_min;           // a tile re-publishes the query's threshold list when it adds >= this many candidates
    uint32_t fast_select;         // exact select through the per-thread-maxima bound first (MSR_SPARSE_FAST_SELECT, default 1)
    // Doc-sharded search over several GPUs (msr_sparse_set_peer_theta): theta[] then lives in symmetric memory and every
    // raise is ALSO pushed into the same slot of the other GPUs' arrays (fire-and-forget red.max over NVLink).  A key is
    // (float32 score, GLOBAL doc id), so "at least k docs of the COLLECTION reach theta" holds on every GPU and each shard
    // prunes as if it had all the tiles.  A shard's own list may then hold fewer than k docs; the merged list is exact.
    uint32_t n_peer_theta;
    uint32_t q_rot;               // the grid walks the queries from this offset (a different one per GPU), so the GPUs
                                  // publish different queries first and find each other's thresholds already there
    unsigned long long* peer_theta[15];
    // Quantile sharing: a GPU's local k-th key says little about the collection's k-th (the top k are spread over the
    // shards).  Every GPU therefore also publishes, per query, its local r-th best key with r = ceil(k / world)
    // (quant[g][q], pushed to all GPUs like theta).  If every GPU holds r docs at or above t, the collection holds
    // world * r >= k: min over g of quant[g][q] is a valid -- and for balanced shards nearly exact -- threshold.
    // wide path (see wide_score_cta): the last wide_ctas CTAs of the grid score the queries the prep kernel listed
    const uint32_t* wide_q;       // [nq_total] queries outside the exact integer range, in any order
    const uint32_t* n_wide;       // how many
    uint32_t wide_first, wide_ctas;
    unsigned long long* wide_acc; // [kWideMax][n_tiles][tile_docs] accumulators, then ranking keys
    uint32_t* wide_cnt;           // [kWideMax][n_tiles] keys per (query, tile) for the merge kernel
    uint32_t* status;
    unsigned long long* quant;    // [quant_world][quant_cap] on this GPU (NULL = off)
    unsigned long long* peer_quant[15];
    uint32_t quant_world, quant_me, quant_cap, quant_rank;
};

// this GPU's r-th best key of query q went up: tell everyone (quant[me][q] on every GPU)
__device__ __forceinline__ void quant_raise(const SearchParams& P, uint32_t q, unsigned long long key) {
    if (!P.quant) return;
    const size_t i = (size_t)P.quant_me * P.quant_cap + P.q_lo + q;
    atomicMax(P.quant + i, key);
    for (uint32_t p = 0; p < P.n_peer_theta; ++p)
        asm volatile("red.relaxed.sys.global.max.u64 [%0], %1;" ::"l"(P.peer_quant[p] + i), "l"(key) : "memory");
}
// the threshold a CTA starts from: the shared theta, or the smallest of the GPUs' r-th keys when that is larger
__device__ __forceinline__ unsigned long long theta_read(const SearchParams& P, uint32_t q) {
    unsigned long long t = *(volatile unsigned long long*)(P.theta + q);
    if (P.quant) {
        unsigned long long m = ~0ull;
        for (uint32_t g = 0; g < P.quant_world; ++g)
            m = min(m, *(volatile unsigned long long*)(P.quant + (size_t)g * P.quant_cap + P.q_lo + q));
        t = max(t, m);
    }
    return t;
}

// raise the query's threshold here and on every peer GPU
__device__ __forceinline__ void theta_raise(const SearchParams& P, unsigned long long* theta_p, uint32_t q, unsigned long long key) {
    atomicMax(theta_p, key);
    for (uint32_t p = 0; p < P.n_peer_theta; ++p)
        asm volatile("red.relaxed.sys.global.max.u64 [%0], %1;" ::"l"(P.peer_theta[p] + P.q_lo + q), "l"(key) : "memory");
}

// One warp per query: exact-arithmetic range check over the kept terms (see msr.h), done once instead of per tile.
__global__ void query_prep_kernel(const uint8_t* __restrict__ keep, const uint32_t* __restrict__ maximp, uint32_t n_terms,
                                  uint32_t nq, const uint64_t* __restrict__ q_indptr, const uint32_t* __restrict__ q_terms,
                                  const uint32_t* __restrict__ q_weights, uint8_t* __restrict__ q_ok,
                                  uint32_t* __restrict__ wide_q, uint32_t* __restrict__ n_wide) {
    const uint32_t q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (q >= nq) return;
    const uint64_t qs = q_indptr[q], qe = q_indptr[q + 1];
    unsigned long long bound = 0;
    uint32_t wmax = 0;
    for (uint64_t j = qs + lane; j < qe; j += 32) {
        const uint32_t t = q_terms[j];
        const uint32_t w = q_weights[j];
        if (t < n_terms && w && keep[t]) {
            bound += (unsigned long long)w * maximp[t];
            wmax = max(wmax, w);
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        bound += __shfl_xor_sync(0xffffffffu, bound, o);
        wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
    }
    if (lane == 0) {
        const bool ok = bound < (1ull << 31) && wmax <= 65535u;
        q_ok[q] = ok ? 1 : 0;
        if (!ok) wide_q[atomicAdd(n_wide, 1u)] = q;            // scored by wide_score_kernel instead (any order)
    }
}

// Wide path: a query outside the integer kernel's exact range (sum_t w_t * max_impact(t) >= 2^31, or a weight above
// 65535) is scored here instead, one CTA per (such query, tile), with u64 accumulators in global memory.  Rule S6 --
// float32(sum of float32(w * tf)), the sum taken in double -- is reproduced exactly: every product is rounded to float32
// first and is an integer below 2^48, so the u64 sum equals the double sum and one round-to-nearest cast ends it.
// Terms are applied one after the other (postings of one term hit distinct docs: no atomics).  The accumulators are then
// turned into ranking keys in place (0 = no hit) and extra CTAs of the merge kernel select the top k out of them and store
// the row where the query belongs.  Both ride on launches that happen anyway: a batch without such queries pays nothing.
constexpr uint32_t kWideMax = 8;          // queries per wide pass (the key buffer is kWideMax * n_docs * 8 bytes)

__device__ __forceinline__ void wide_score_cta(const SearchParams& P, uint32_t w, uint32_t tile) {
    const uint32_t tid = threadIdx.x;
    const uint32_t nw = *P.n_wide;
    if (P.wide_first + w >= nw) {
        if (tid == 0) P.wide_cnt[w * P.n_tiles + tile] = 0;
        return;
    }
    // more flagged queries than one pass takes: the host path runs further passes, the device path reports it (status)
    if (w == 0 && tile == 0 && tid == 0 && P.wide_first == 0 && nw > kWideMax) atomicOr(P.status, 2u);
    const uint32_t q = P.wide_q[P.wide_first + w];
    const uint32_t tile_lo = tile * P.tile_docs, tl = min(P.tile_docs, P.n_docs - tile_lo);
    unsigned long long* acc = P.wide_acc + ((size_t)w * P.n_tiles + tile) * P.tile_docs;
    const uint32_t tlp = (tl + 31) & ~31u;                      // accumulator positions (perm32 permutes inside 32-doc groups)
    for (uint32_t d = tid; d < tlp; d += blockDim.x) acc[d] = 0ull;
    __syncthreads();
    const uint32_t* seg_row = P.seg_ptr + (size_t)tile * P.n_terms;
    for (uint64_t j = P.q_indptr[q]; j < P.q_indptr[q + 1]; ++j) {
        const uint32_t t = P.q_terms[j], wt = P.q_weights[j];
        if (t >= P.n_terms || wt == 0 || !P.term_keep[t]) continue;         // uniform over the CTA
        const float wf = __uint2float_rn(wt);
        const uint32_t di = P.dense_id[t];
        if (di != kNotDense) {
            const uint16_t* row = P.dense_rows + (size_t)di * P.dense_stride + tile_lo;     // the whole impact
            for (uint32_t d = tid; d < tl; d += blockDim.x) {
                const uint32_t imp = row[d];
                if (imp) acc[perm32(d)] += (unsigned long long)__fmul_rn(wf, __uint2float_rn(imp));
            }
        } else {
            const uint32_t s0 = seg_row[t], s1 = seg_row[t + 1];
            for (uint32_t p = s0 + tid; p < s1; p += blockDim.x) {
                const uint32_t v = P.postings[p], imp = v >> 16;               // padding slots carry impact 0; low half = position
                if (imp) acc[v & 0xFFFFu] += (unsigned long long)__fmul_rn(wf, __uint2float_rn(imp));
            }
        }
        __syncthreads();
    }
    const uint32_t gdoc0 = (uint32_t)P.doc_base + tile_lo;
    for (uint32_t d = tid; d < tlp; d += blockDim.x) {
        const unsigned long long a = acc[d];
        acc[d] = a ? make_key(__ull2float_rn(a), gdoc0 + unperm32(d)) : 0ull;
    }
    if (tid == 0) P.wide_cnt[w * P.n_tiles + tile] = tlp;
}

__device__ __forceinline__ uint32_t ld_stream_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void red_shared_add(uint32_t saddr, uint32_t val) {
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(saddr), "r"(val) : "memory");
}
// One posting slot v = impact << 16 | doc offset, weight below 256 given as wb = w << 24:
// address = acc_s + 4 * offset and value = impact * w, one dp2a each (bytes (4,0,0,w) of `wb | 4`).
__device__ __forceinline__ void scatter_fast(uint32_t acc_s, uint32_t v, uint32_t wb4) {
    red_shared_add(__dp2a_lo(v, wb4, acc_s), __dp2a_hi(v, wb4, 0u));
}

struct SegStage {                      // streaming phase view of the union region
    uint32_t start[kSegMax];
    uint32_t len[kSegMax];
    uint32_t w[kSegMax];
    uint32_t pref[kSegMax + 1];        // exclusive prefix of chunk counts
    // dense terms of the batch: row start in 16-byte units and weight.  Plain terms fill the arrays from the front,
    // terms that need the per-term float32 rounding of S6 from the back.
    alignas(16) uint32_t drow[kSegMax];
    alignas(16) uint32_t dw[kSegMax];
#if MSR_DENSE_MMA
    uint32_t qw[kSegMax / 4];          // front list: the four u8 weights of quad i (terms 4i .. 4i+3 of the list), the mma's B operand
#else
    alignas(16) uint4 db[kSegMax];     // front list: (w, w << 8, w << 16, w << 24), the dp4a selectors of the weight
#endif
};
struct SelStage {                      // radix-select view
    uint32_t hist[kBins];
    uint64_t cand[kCandMax];
};
union StageUnion {
    SegStage seg;
    SelStage sel;
    uint64_t keys[kUnionMax];          // threshold-merge view
};
struct SmallState {
    uint32_t vmax, hits, out_cnt, cand_cnt, sel_b, sel_clt, sel_bcount, n_dense;
    uint32_t n_dense_f, pad;
    unsigned long long theta;          // the query's threshold as ONE thread read it (see phase 2a)
    uint32_t wsum[kWarps];
};
static_assert(sizeof(SmallState) % 16 == 0 && sizeof(StageUnion) % 16 == 0, "accumulators must stay 16-byte aligned");


// a[0..7] += w * (8 u16 impacts packed in v), any w below 2^16: unpack (ALU pipe) + IMAD (FMA pipe) per impact
__device__ __forceinline__ void dense_fma8_wide(uint32_t (&a)[8], const uint4 v, const uint32_t w) {
    a[0] += (v.x & 0xFFFFu) * w; a[1] += (v.x >> 16) * w;
    a[2] += (v.y & 0xFFFFu) * w; a[3] += (v.y >> 16) * w;
    a[4] += (v.z & 0xFFFFu) * w; a[5] += (v.z >> 16) * w;
    a[6] += (v.w & 0xFFFFu) * w; a[7] += (v.w >> 16) * w;
}
// Weight below 256 on the u8 rows: ONE instruction per impact.  dp4a multiplies the four u8 of its first operand by
// the four u8 of the second and adds a 32-bit accumulator (SASS IDP.4A.U8.U8, same issue rate as IMAD on sm_100a --
// scratch/idp_bench.cu); with the weight in byte i of the selector and zeros elsewhere it yields a + impact_i * w,
// without any mask / shift instruction on the ALU pipe.  v = impacts of 8 consecutive docs.
__device__ __forceinline__ void dense_dp4a8(uint32_t (&a)[8], const uint2 v, const uint4 b) {
    a[0] = __dp4a(v.x, b.x, a[0]); a[1] = __dp4a(v.x, b.y, a[1]); a[2] = __dp4a(v.x, b.z, a[2]); a[3] = __dp4a(v.x, b.w, a[3]);
    a[4] = __dp4a(v.y, b.x, a[4]); a[5] = __dp4a(v.y, b.y, a[5]); a[6] = __dp4a(v.y, b.z, a[6]); a[7] = __dp4a(v.y, b.w, a[7]);
}

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

// One chunk = kChunk consecutive posting slots of a bucket (the bucket's last chunk holds n < kChunk).  A warp loads
// chunk i+1 -- of the same bucket or of the next one -- before it applies chunk i, so the L2 round trip of a small
// bucket hides behind the previous chunk instead of serialising.
__device__ __forceinline__ void chunk_load(uint32_t (&v)[kUnroll], const uint32_t* src, uint32_t n, uint32_t lane) {
    if (n == (uint32_t)kChunk) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) v[u] = ld_stream_u32(src + u * 32);
    } else {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) v[u] = (u * 32 + lane < n) ? ld_stream_u32(src + u * 32) : 0u;
    }
}
// KIND 0: weight below 256, dp2a form (w = weight << 24 | 4); 1: weight below 2^16, mask / shift / IMAD form;
// 2: S6 term, float32(boost*freq) per posting.
template <int KIND>
__device__ __forceinline__ void posting_apply(uint32_t acc_s, uint32_t v, uint32_t w) {
    if (KIND == 0) scatter_fast(acc_s, v, w);
    else if (KIND == 1) red_shared_add(acc_s + (v & 0xFFFFu) * 4u, (v >> 16) * w);
    else red_shared_add(acc_s + (v & 0xFFFFu) * 4u, (uint32_t)__uint2float_rn((v >> 16) * w));
}
template <int KIND>
__device__ __forceinline__ void chunk_apply_kind(const uint32_t (&v)[kUnroll], uint32_t n, uint32_t w, uint32_t lane,
                                                 uint32_t acc_s) {
    if (n == (uint32_t)kChunk) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) posting_apply<KIND>(acc_s, v[u], w);
    } else {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
            if (u * 32 + lane < n) posting_apply<KIND>(acc_s, v[u], w);
    }
}
__device__ __forceinline__ void chunk_apply(const uint32_t (&v)[kUnroll], uint32_t n, uint32_t wraw, uint32_t lane,
                                            uint32_t acc_s) {
    if (wraw < 256u) chunk_apply_kind<0>(v, n, (wraw << 24) | 4u, lane, acc_s);
    else if (!(wraw >> 31)) chunk_apply_kind<1>(v, n, wraw, lane, acc_s);
    else chunk_apply_kind<2>(v, n, wraw & 0x7FFFFFFFu, lane, acc_s);
}

// Rank by counting with all threads busy: key i = tid >> pl is shared by 2^pl adjacent lanes, each counting the keys
// greater than it in its own part of [0, n); the partial counts are summed with shuffles.  Every thread of the CTA calls
// (n << pl <= kThreads); `mine` returns the key (0 = none), the result is the number of keys greater than it.
__device__ __forceinline__ uint32_t count_greater(const uint64_t* keys, uint32_t n, uint32_t pl, uint32_t tid, uint64_t& mine) {
    const uint32_t i = tid >> pl, parts = 1u << pl, part = tid & (parts - 1u);
    mine = i < n ? keys[i] : 0ull;
    uint32_t gt = 0;
    if (mine) {
        const uint32_t per = (n + parts - 1u) >> pl, lo = part * per, hi = min(n, lo + per);
        for (uint32_t j = lo; j < hi; ++j) gt += keys[j] > mine;
    }
    for (uint32_t o = 1; o < parts; o <<= 1) gt += __shfl_xor_sync(0xffffffffu, gt, o);
    return gt;
}
__device__ __forceinline__ uint32_t parts_log2_for(uint32_t n) {        // largest pl <= 2 with (n << pl) <= kThreads
    return (n << 2) <= (uint32_t)kThreads ? 2u : ((n << 1) <= (uint32_t)kThreads ? 1u : 0u);
}

__global__ void __launch_bounds__(kThreads, MSR_SPARSE_MIN_CTAS)
sparse_score_topk_kernel(const SearchParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    StageUnion* U = reinterpret_cast<StageUnion*>(smem_raw);
    SmallState* S = reinterpret_cast<SmallState*>(smem_raw + sizeof(StageUnion));
    uint32_t* acc = reinterpret_cast<uint32_t*>(smem_raw + sizeof(StageUnion) + sizeof(SmallState));
    const uint32_t acc_s = (uint32_t)__cvta_generic_to_shared(acc);

    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (blockIdx.x >= P.nq * P.n_tiles) {                       // the grid's tail: wide path
        const uint32_t x = blockIdx.x - P.nq * P.n_tiles;
        wide_score_cta(P, x / P.n_tiles, x % P.n_tiles);
        return;
    }
    // Grid order: tile-major (all CTAs in flight read ONE tile's slice of the index).  Interleaving groups of 2 / 4 / 8 tiles
    // was measured: 490 K -> 455 K / 408 K / 356 K q/s (looser shared threshold, worse L2 locality) -- and the extra index
    // arithmetic alone cost an 8-byte register spill and ~3 %.
    uint32_t q = blockIdx.x % P.nq;
    const uint32_t tile = blockIdx.x / P.nq;
    q += P.q_rot;
    if (q >= P.nq) q -= P.nq;
    const uint32_t tile_lo = tile * P.tile_docs;
    const uint32_t tl = min(P.tile_docs, P.n_docs - tile_lo);
    const uint64_t qs = P.q_indptr[P.q_lo + q], qe = P.q_indptr[P.q_lo + q + 1];
    const uint32_t* seg_row = P.seg_ptr + (size_t)tile * P.n_terms;
    uint64_t* keys_out = P.cand_keys + ((size_t)q * P.n_tiles + tile) * P.k_stride;
    uint32_t* cnt_out = P.cand_cnt + (size_t)q * P.n_tiles + tile;

    if (!P.q_ok[P.q_lo + q]) {                                  // outside the exact range: wide_score_kernel fills the row
        if (tid == 0) *cnt_out = 0;
        return;
    }
    if (tid == 0) { S->vmax = 0; S->hits = 0; S->out_cnt = 0; S->cand_cnt = 0; }

    // ---- phase 1: accumulate integer impacts of the tile's docs in shared memory ----
    // Dense terms (impact rows) are added doc-parallel with plain stores -- the first batch thereby also zeroes the
    // accumulators -- then the other terms' posting buckets are scattered with shared-memory atomics.
    uint64_t batch = qs;
    do {
        __syncthreads();   // previous batch fully consumed
        if (tid == 0) { S->n_dense = 0; S->n_dense_f = 0; }
#if MSR_DENSE_MMA
        if (tid < kSegMax / 4) U->seg.qw[tid] = 0;
#endif
        __syncthreads();
        uint32_t nchunks = 0;
        if (tid < kSegMax) {
            uint32_t len = 0, start = 0, w = 0;
            const uint64_t j = batch + tid;
            if (j < qe) {
                const uint32_t t = P.q_terms[j];
                w = P.q_weights[j];
                if (t < P.n_terms && w) {
                    // independent gathers first (one round trip), decisions after
                    const uint32_t keep = P.term_keep[t], mi = P.term_maximp[t], di = P.dense_id[t];
                    const uint32_t s0 = seg_row[t], s1 = seg_row[t + 1];
                    if (keep) {
                        // S6: float32(boost*freq) rounds once a product can reach 2^24 -> flag the term (bit 31)
                        if ((uint64_t)w * mi >= (1u << 24)) w |= 0x80000000u;
                        if (di != kNotDense) {
                            // front list: weights below 256 (dp2a form, stored as w | w << 24); back list: wider
                            // weights (IMAD form) and S6 terms (bit 31 kept)
                            // back list: u16 rows hold the whole impact.  Front list: u8 rows hold min(impact, 255)
                            // and the term's bucket the rest, so the bucket is scattered as well.
                            const bool fast = w < 256u;
                            const uint32_t pos = fast ? atomicAdd(&S->n_dense, 1u) : kSegMax - 1 - atomicAdd(&S->n_dense_f, 1u);
                            if (fast) {
                                U->seg.drow[pos] = di * (P.dense_row_bytes >> 1);
#if MSR_DENSE_MMA
                                atomicOr(&U->seg.qw[pos >> 2], w << (8 * (pos & 3)));
#else
                                U->seg.db[pos] = make_uint4(w, w << 8, w << 16, w << 24);
#endif
                                start = s0;
                                len = s1 - s0;
                            } else {
                                U->seg.drow[pos] = di * P.dense_row_bytes;
                                U->seg.dw[pos] = w;
                            }
                        } else {
                            start = s0;
                            len = s1 - s0;
                        }
                    } else {
                        w = 0;
                    }
                } else {
                    w = 0;
                }
            }
            U->seg.start[tid] = start;
            U->seg.len[tid] = len;
            U->seg.w[tid] = w;
            nchunks = (len + kChunk - 1) / kChunk;
        }
        __syncthreads();
#if MSR_DENSE_MMA
        {
            // dense phase on the integer tensor pipe (sparse_dense_mma.cuh): lane L of a warp owns accumulator positions
            // 8L .. 8L+7 of the warp's 256-doc block; two slots per thread and step as before
            const uint32_t nd = S->n_dense, ndf = S->n_dense_f;
            const bool first = batch == qs;
            if (first || nd || ndf) {
                // whole warps only (mma.sync is warp-wide): 32 slots = one 256-doc block per warp and step; slots past the tile's
                // end read the next tile's rows (or the row padding) into accumulators nobody looks at (positions >= tlp)
                const uint32_t nslots = ((tl + 255) & ~255u) >> 3;
                const char* rows = reinterpret_cast<const char*>(P.dense_rows + tile_lo);
                const char* rows8 = reinterpret_cast<const char*>(P.dense_rows8 + tile_lo);
                const uint32_t g = lane >> 2, t = lane & 3;
                const uint32_t m0 = g == t ? 0xFFFFFFFFu : 0u, m1 = g == t + 4 ? 0xFFFFFFFFu : 0u;   // block-diagonal B
                for (uint32_t slot = tid; slot < nslots; slot += 2 * kThreads) {
                    const bool two = slot + kThreads < nslots;  // else the second slot's loads are dummies (row padding)
                    uint32_t pa[8], pb[8];
                    uint4* ap = reinterpret_cast<uint4*>(acc + 8 * slot);
                    uint4* cp = ap + 2 * kThreads;
                    if (first) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) { pa[i] = 0; pb[i] = 0; }
                    } else {
                        const uint4 x = ap[0], y = ap[1];
                        pa[0] = x.x; pa[1] = x.y; pa[2] = x.z; pa[3] = x.w; pa[4] = y.x; pa[5] = y.y; pa[6] = y.z; pa[7] = y.w;
                        if (two) {
                            const uint4 x2 = cp[0], y2 = cp[1];
                            pb[0] = x2.x; pb[1] = x2.y; pb[2] = x2.z; pb[3] = x2.w; pb[4] = y2.x; pb[5] = y2.y; pb[6] = y2.z; pb[7] = y2.w;
                        }
                    }
                    const char* r8 = rows8 + slot * 8u;         // u8 rows: 8 docs = 8 bytes
                    constexpr uint32_t kB8 = kThreads * 8u;     // byte distance of the second slot in a u8 row
                    uint32_t d = 0;
                    for (; d + 4 <= nd; d += 4) {
                        const uint4 r = *reinterpret_cast<const uint4*>(&U->seg.drow[d]);
                        const uint32_t W = U->seg.qw[d >> 2];
                        dense_quad<4>(pa, pb, r8, kB8, r, W & m0, W & m1);
                    }
                    if (nd & 3) {
                        const uint4 r = *reinterpret_cast<const uint4*>(&U->seg.drow[d]);
                        const uint32_t W = U->seg.qw[d >> 2];
                        if ((nd & 3) == 3) dense_quad<3>(pa, pb, r8, kB8, r, W & m0, W & m1);
                        else if ((nd & 3) == 2) dense_quad<2>(pa, pb, r8, kB8, r, W & m0, W & m1);
                        else dense_quad<1>(pa, pb, r8, kB8, r, W & m0, W & m1);
                    }
                    if (ndf) {                                  // weights of 256 and more; S6 terms (u16 rows)
                        const char* c16 = rows + 2u * (8u * (slot & ~31u) + 32u * g + 16u * (t & 1u));
                        constexpr uint32_t kB16 = kThreads * 16u;
                        for (uint32_t f = 0; f < ndf; ++f) {
                            const uint32_t r = U->seg.drow[kSegMax - 1 - f], wr = U->seg.dw[kSegMax - 1 - f];
                            dense_wide_term(pa, c16 + r, t >> 1, wr);
                            dense_wide_term(pb, c16 + r + kB16, t >> 1, wr);
                        }
                    }
                    ap[0] = make_uint4(pa[0], pa[1], pa[2], pa[3]);
                    ap[1] = make_uint4(pa[4], pa[5], pa[6], pa[7]);
                    if (two) {
                        cp[0] = make_uint4(pb[0], pb[1], pb[2], pb[3]);
                        cp[1] = make_uint4(pb[4], pb[5], pb[6], pb[7]);
                    }
                }
            }
        }
#else
        {
            const uint32_t nd = S->n_dense, ndf = S->n_dense_f;
            const bool first = batch == qs;
            if (first || nd || ndf) {
                const uint32_t nslots = ((tl + 31) & ~31u) >> 3; // 8 docs per thread and step; whole 32-doc groups (the selection passes read them)
                const char* rows = reinterpret_cast<const char*>(P.dense_rows + tile_lo);
                const char* rows8 = reinterpret_cast<const char*>(P.dense_rows8 + tile_lo);
                // Two slots per thread and step (slot, slot + kThreads): the second one is reached with an immediate
                // offset, so one address computation serves two loads and 8 loads are in flight per thread.
                for (uint32_t slot = tid; slot < nslots; slot += 2 * kThreads) {
                    const bool two = slot + kThreads < nslots;  // else the second slot's loads are dummies (row padding)
                    uint32_t a[8], c[8];
                    uint4* ap = reinterpret_cast<uint4*>(acc + 8 * slot);
                    uint4* cp = ap + 2 * kThreads;
                    if (first) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) { a[i] = 0; c[i] = 0; }
                    } else {
                        const uint4 x = ap[0], y = ap[1];
                        a[0] = x.x; a[1] = x.y; a[2] = x.z; a[3] = x.w; a[4] = y.x; a[5] = y.y; a[6] = y.z; a[7] = y.w;
                        if (two) {
                            const uint4 x2 = cp[0], y2 = cp[1];
                            c[0] = x2.x; c[1] = x2.y; c[2] = x2.z; c[3] = x2.w; c[4] = y2.x; c[5] = y2.y; c[6] = y2.z; c[7] = y2.w;
                        }
                    }
                    const char* r8 = rows8 + slot * 8u;         // u8 rows: 8 docs = 8 bytes
                    constexpr uint32_t kB8 = kThreads * 8u;     // byte distance of the second slot in a u8 row
                    uint32_t d = 0;
                    for (; d + 4 <= nd; d += 4) {
                        const uint4 r = *reinterpret_cast<const uint4*>(&U->seg.drow[d]);
                        const uint2 v0 = ld_stream_v2(r8 + r.x), v1 = ld_stream_v2(r8 + r.y);
                        const uint2 v2 = ld_stream_v2(r8 + r.z), v3 = ld_stream_v2(r8 + r.w);
                        const uint2 u0 = ld_stream_v2(r8 + r.x + kB8), u1 = ld_stream_v2(r8 + r.y + kB8);
                        const uint2 u2 = ld_stream_v2(r8 + r.z + kB8), u3 = ld_stream_v2(r8 + r.w + kB8);
                        uint4 b = U->seg.db[d];
                        dense_dp4a8(a, v0, b); dense_dp4a8(c, u0, b);
                        b = U->seg.db[d + 1];
                        dense_dp4a8(a, v1, b); dense_dp4a8(c, u1, b);
                        b = U->seg.db[d + 2];
                        dense_dp4a8(a, v2, b); dense_dp4a8(c, u2, b);
                        b = U->seg.db[d + 3];
                        dense_dp4a8(a, v3, b); dense_dp4a8(c, u3, b);
                    }
                    if (nd & 2) {
                        const uint2 r = *reinterpret_cast<const uint2*>(&U->seg.drow[d]);
                        const uint2 v0 = ld_stream_v2(r8 + r.x), v1 = ld_stream_v2(r8 + r.y);
                        const uint2 u0 = ld_stream_v2(r8 + r.x + kB8), u1 = ld_stream_v2(r8 + r.y + kB8);
                        uint4 b = U->seg.db[d];
                        dense_dp4a8(a, v0, b); dense_dp4a8(c, u0, b);
                        b = U->seg.db[d + 1];
                        dense_dp4a8(a, v1, b); dense_dp4a8(c, u1, b);
                        d += 2;
                    }
                    if (nd & 1) {
                        const uint32_t r = U->seg.drow[d];
                        const uint2 v0 = ld_stream_v2(r8 + r), u0 = ld_stream_v2(r8 + r + kB8);
                        const uint4 b = U->seg.db[d];
                        dense_dp4a8(a, v0, b); dense_dp4a8(c, u0, b);
                    }
                    const char* r16 = rows + slot * 16u;        // u16 rows (row offsets are 32-bit byte offsets)
                    constexpr uint32_t kB16 = kThreads * 16u;
                    for (uint32_t f = 0; f < ndf; ++f) {        // weights of 256 and more; S6 terms
                        const uint32_t r = U->seg.drow[kSegMax - 1 - f];
                        const uint4 v = ld_stream_v4(r16 + r), u = ld_stream_v4(r16 + r + kB16);
                        const uint32_t wr = U->seg.dw[kSegMax - 1 - f], w = wr & 0x7FFFFFFFu;
                        if (!(wr >> 31)) {
                            dense_fma8_wide(a, v, w); dense_fma8_wide(c, u, w);
                        } else {                                // S6: float32(boost*freq) per term
                            const uint32_t e[4] = {v.x, v.y, v.z, v.w}, g[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                a[2 * i] += (uint32_t)__uint2float_rn((e[i] & 0xFFFFu) * w);
                                a[2 * i + 1] += (uint32_t)__uint2float_rn((e[i] >> 16) * w);
                                c[2 * i] += (uint32_t)__uint2float_rn((g[i] & 0xFFFFu) * w);
                                c[2 * i + 1] += (uint32_t)__uint2float_rn((g[i] >> 16) * w);
                            }
                        }
                    }
                    ap[0] = make_uint4(a[0], a[1], a[2], a[3]);
                    ap[1] = make_uint4(a[4], a[5], a[6], a[7]);
                    if (two) {
                        cp[0] = make_uint4(c[0], c[1], c[2], c[3]);
                        cp[1] = make_uint4(c[4], c[5], c[6], c[7]);
                    }
                }
            }
        }
#endif
        uint32_t total;
        const uint32_t ex = block_excl_scan(nchunks, S->wsum, &total);
        if (tid < kSegMax) U->seg.pref[tid] = ex;
        if (tid == 0) U->seg.pref[kSegMax] = total;
        __syncthreads();

        // each warp takes a contiguous range of chunk numbers [g, g1) of the batch, so it mostly walks inside one bucket
        uint32_t g = (uint32_t)(((uint64_t)total * warp) / kWarps);
        const uint32_t g1 = (uint32_t)(((uint64_t)total * (warp + 1)) / kWarps);
        if (g < g1) {
            uint32_t seg;
            {                                                   // last seg with pref[seg] <= g
                uint32_t lo = 0, hi = kSegMax;
                while (hi - lo > 1) {
                    const uint32_t mid = (lo + hi) >> 1;
                    if (U->seg.pref[mid] <= g) lo = mid; else hi = mid;
                }
                seg = lo;
            }
            // chunk stream: (source, valid slots, weight) of chunk number g, then ++g
            uint32_t b_sb = 0, b_se = 0, b_len = 0, b_w = 0;
            const uint32_t* b_src = nullptr;
            auto next_chunk = [&](const uint32_t*& src, uint32_t& n, uint32_t& w) {
                while (g >= b_se) {                             // enter the bucket that holds chunk g
                    b_sb = U->seg.pref[seg];
                    b_se = U->seg.pref[seg + 1];
                    if (b_se > g) {
                        b_len = U->seg.len[seg];
                        b_w = U->seg.w[seg];
                        b_src = P.postings + U->seg.start[seg] + lane;
                    }
                    ++seg;
                }
                const uint32_t c = g - b_sb;
                src = b_src + (size_t)c * kChunk;
                n = min((uint32_t)kChunk, b_len - c * kChunk);
                w = b_w;
                ++g;
            };
            const uint32_t* src;
            uint32_t n, w, v[kUnroll];
            next_chunk(src, n, w);
            chunk_load(v, src, n, lane);
            while (g < g1) {
                const uint32_t* src2;
                uint32_t n2, w2, nx[kUnroll];
                next_chunk(src2, n2, w2);
                chunk_load(nx, src2, n2, lane);                 // in flight while chunk g-1 is applied
                chunk_apply(v, n, w, lane, acc_s);
                n = n2; w = w2;
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) v[u] = nx[u];
            }
            chunk_apply(v, n, w, lane, acc_s);
        }
    } while ((batch += kSegMax) < qe);
    // The threshold is read ONCE per CTA and broadcast: theta rises while other tiles of the query publish, and
    // threads reading it on their own could disagree -- some still seeing 0 (exact select, with its barriers)
    // while others already take the threshold path.
    if (tid == 0) S->theta = P.use_theta ? theta_read(P, q) : 0ull;
    __syncthreads();

    // ---- phase 2: candidates of the tile by (float32(score) desc, doc asc) ----
    const uint32_t kq = min(P.k, tl);
    const uint32_t tlp = (tl + 31) & ~31u;                      // accumulator POSITIONS: position i holds doc unperm32(i); past-the-end docs hold 0
    const uint32_t gdoc0 = (uint32_t)P.doc_base + tile_lo;
    unsigned long long* theta_p = P.theta + q;
    uint64_t* tlist = P.tlist + (size_t)q * P.k;

    // 2a. threshold mode: one pass, emit the docs whose key reaches theta
    const unsigned long long theta = S->theta;
    if (theta) {
        const float tf = key_score(theta);                      // an integer-valued float32 in (0, 2^31)
        uint32_t amin = (uint32_t)tf;                           // smallest integer score whose float32 image is >= tf
        while (amin > 1 && __uint2float_rn(amin - 1) >= tf) --amin;
        const uint32_t cap = min(P.k_stride, (uint32_t)kUnionMax);
        const uint32_t nvec = tlp >> 2;                          // 4 positions per thread and step
#ifndef MSR_THETA_UNROLL
#define MSR_THETA_UNROLL 2
#endif
        // MSR_THETA_UNROLL independent 16-byte accumulator loads per step before any is tested (the pass is shared-memory
        // latency-bound: 10 % of the samples for 5 % of the instructions).  Measured (profiles/r2_theta_unroll_ab.txt):
        // 1 -> 19.278 ms, 2 -> 19.236 ms, 4 -> 19.283 ms per scorer launch on config 4, repeatable to 0.003 ms.
        for (uint32_t j0 = tid; j0 < nvec; j0 += MSR_THETA_UNROLL * kThreads) {
            uint4 xs[MSR_THETA_UNROLL];
#pragma unroll
            for (int v = 0; v < MSR_THETA_UNROLL; ++v) {
                const uint32_t j = j0 + v * kThreads;
                xs[v] = j < nvec ? reinterpret_cast<const uint4*>(acc)[j] : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int v = 0; v < MSR_THETA_UNROLL; ++v) {
                const uint4 x = xs[v];
                const uint32_t j = j0 + v * kThreads;
                if (max(max(x.x, x.y), max(x.z, x.w)) >= amin) {
                    const uint32_t e[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const uint32_t i = 4 * j + u;
                        if (e[u] >= amin) {
                            const uint64_t key = make_key(__uint2float_rn(e[u]), gdoc0 + unperm32(i));
                            if (key >= theta) {
                                const uint32_t pos = atomicAdd(&S->cand_cnt, 1u);
                                if (pos < cap) U->keys[pos] = key;
                            }
                        }
                    }
                }
            }
        }
        __syncthreads();
        const uint32_t c = S->cand_cnt;
        if (c <= cap) {
            for (uint32_t i = tid; i < c; i += kThreads) keys_out[i] = U->keys[i];
            if (tid == 0) *cnt_out = c;
            // merge with the published list: best k distinct keys of (own candidates U tlist) -> new list, new theta.
            // Skipped while the tile adds little (theta would hardly move) or the union does not fit.
            if (c < P.merge_min || c + P.k > (uint32_t)kUnionMax) return;
            for (uint32_t r = tid; r < P.k; r += kThreads) U->keys[c + r] = ((volatile const uint64_t*)tlist)[r];
            const uint32_t n = c + P.k;
            if (n <= (uint32_t)kThreads) {
                // rank by counting: one key per thread against all n (broadcast reads).  Own candidates and list
                // entries come from different tiles, so equal keys only appear when two tiles of the query raced
                // on the list; that case takes the sorting path below.
                // Equal keys get equal ranks, so a rank claimed twice (bitmap in wsum) reveals them.
                if (tid < (uint32_t)kWarps) S->wsum[tid] = 0;
                __syncthreads();
                const uint32_t pl = parts_log2_for(n);
                uint64_t mine;
                const uint32_t gt = count_greater(U->keys, n, pl, tid, mine);
                uint32_t dup = 0;
                if ((tid & ((1u << pl) - 1u)) != 0) mine = 0ull;        // one writer per key
                if (mine) dup = (atomicOr(&S->wsum[gt >> 5], 1u << (gt & 31)) >> (gt & 31)) & 1u;
                if (!__syncthreads_or((int)dup)) {
                    if (mine && gt < P.k) {
                        tlist[gt] = mine;
                        if (gt == P.k - 1) theta_raise(P, theta_p, q, (unsigned long long)mine);
                        if (gt == P.quant_rank - 1) quant_raise(P, q, (unsigned long long)mine);
                    }
                    return;
                }
            }
            uint32_t n2 = 2;
            while (n2 < n) n2 <<= 1;
            for (uint32_t i = n + tid; i < n2; i += kThreads) U->keys[i] = 0;
            bitonic_sort_desc(U->keys, n2);
            uint64_t e[kPer];
            uint32_t f[kPer], ls = 0;
            uint64_t prev = (tid > 0 && kPer * tid - 1 < n2) ? U->keys[kPer * tid - 1] : ~0ull;
#pragma unroll
            for (int j = 0; j < kPer; ++j) {
                const uint32_t idx = kPer * tid + j;
                e[j] = idx < n2 ? U->keys[idx] : 0ull;
                f[j] = (e[j] != 0ull && e[j] != prev) ? 1u : 0u;
                prev = e[j];
                ls += f[j];
            }
            uint32_t total;
            uint32_t ex = block_excl_scan(ls, S->wsum, &total);
#pragma unroll
            for (int j = 0; j < kPer; ++j) {
                if (f[j]) {
                    if (ex < P.k) {
                        tlist[ex] = e[j];
                        if (ex == P.k - 1) theta_raise(P, theta_p, q, (unsigned long long)e[j]);
                        if (ex == P.quant_rank - 1) quant_raise(P, q, (unsigned long long)e[j]);
                    }
                    ++ex;
                }
            }
            return;
        }
        __syncthreads();                                        // more candidates than slots: exact select below
        if (tid == 0) S->cand_cnt = 0;
    }

    // 2b. exact radix select of the tile's top-k
    const uint4* acc4 = reinterpret_cast<const uint4*>(acc);
    const uint32_t nvec4 = tlp >> 2;
    uint32_t tmax = 0;
    {
        // (accumulator slots past tl hold 0, so whole 16-byte vectors are read in all passes of the select)
        uint32_t m = 0, h = 0;
        for (uint32_t j = tid; j < nvec4; j += kThreads) {
            const uint4 x = acc4[j];
            m = max(max(m, max(x.x, x.y)), max(x.z, x.w));
            h += (x.x != 0) + (x.y != 0) + (x.z != 0) + (x.w != 0);
        }
        tmax = m;                                               // best score among this thread's own docs
        m = warp_reduce_max(m);
        h = warp_reduce_add(h);
        if (lane == 0) {
            if (m) atomicMax(&S->vmax, m);
            if (h) atomicAdd(&S->hits, h);
        }
    }
    __syncthreads();
    const uint32_t vmax = S->vmax, hits = S->hits;
    if (hits == 0) {
        if (tid == 0) *cnt_out = 0;
        return;
    }
    if (hits <= kq) {                                           // S7: fewer hits than k -> all of them
        for (uint32_t i = tid; i < tlp; i += kThreads) {
            uint32_t a = acc[i];
            if (a) keys_out[atomicAdd(&S->out_cnt, 1u)] = make_key(__uint2float_rn(a), gdoc0 + unperm32(i));
        }
        if (tid == 0) *cnt_out = hits;
        return;
    }
    // Fast path (k <= threads, scores below 2^24): the threads own disjoint docs, so the k-th largest of the per-thread
    // maxima is a score g that at least k docs of the tile reach -- and, the maxima being order statistics of ~tl/threads
    // docs each, only a few more than k do (about 112 of 20 K for k = 100).  g is found by a bitwise search with one
    // __syncthreads_count per bit; ONE compare pass then collects the docs with score >= g (all of the top k are among
    // them) and rank counting orders them.  No histogram, no shared atomics on the bulk of the accumulators.
    if (P.fast_select && P.k <= (uint32_t)kThreads && vmax < (1u << 24)) {
        uint32_t g = 0;
        for (int b = 31 - __clz(vmax); b >= 0; --b) {
            const uint32_t t = g | (1u << b);
            if ((uint32_t)__syncthreads_count(tmax >= t) >= P.k) g = t;
        }
        if (g) {
            for (uint32_t j = tid; j < nvec4; j += kThreads) {
                const uint4 x = acc4[j];
                if (max(max(x.x, x.y), max(x.z, x.w)) >= g) {
                    const uint32_t e[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const uint32_t i = 4 * j + u;
                        if (e[u] >= g) {
                            const uint32_t pos = atomicAdd(&S->cand_cnt, 1u);
                            if (pos < (uint32_t)kThreads) U->keys[pos] = make_key(__uint2float_rn(e[u]), gdoc0 + unperm32(i));
                        }
                    }
                }
            }
            __syncthreads();
            const uint32_t c = S->cand_cnt;                     // >= k by construction
            if (c <= (uint32_t)kThreads) {
                const uint32_t pl = parts_log2_for(c);
                uint64_t mine;
                const uint32_t gt = count_greater(U->keys, c, pl, tid, mine);      // distinct docs: distinct keys
                if (mine && (tid & ((1u << pl) - 1u)) == 0) {
                    if (gt < P.k) {
                        keys_out[gt] = mine;
                        if (P.use_theta) {
                            tlist[gt] = mine;
                            if (gt == P.k - 1) theta_raise(P, theta_p, q, (unsigned long long)mine);
                            if (gt == P.quant_rank - 1) quant_raise(P, q, (unsigned long long)mine);
                        }
                    }
                }
                if (tid == 0) *cnt_out = P.k;
                return;
            }
            __syncthreads();                                    // too many ties at g: the radix select below
            if (tid == 0) S->cand_cnt = 0;
        }
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
        for (uint32_t j = tid; j < nvec4; j += kThreads) {
            const uint4 x = acc4[j];
            const uint32_t e[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint32_t a = e[u];
                if (a) {
                    uint32_t d = kmax - MSR_SELKEY(a);
                    uint32_t hi = sh_prev >= 32 ? 0u : (d >> sh_prev);
                    if (hi == pref) atomicAdd(&U->sel.hist[(d >> s) & ((1u << width) - 1u)], 1u);
                }
            }
        }
        __syncthreads();
        {
            uint32_t h4[kPer], ls = 0;
#pragma unroll
            for (int j = 0; j < kPer; ++j) { h4[j] = U->sel.hist[tid * kPer + j]; ls += h4[j]; }
            uint32_t total;
            uint32_t ex = block_excl_scan(ls, S->wsum, &total);
#pragma unroll
            for (int j = 0; j < kPer; ++j) {
                if (ex < remaining && ex + h4[j] >= remaining) {
                    S->sel_b = tid * kPer + j; S->sel_clt = ex; S->sel_bcount = h4[j];
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
    for (uint32_t j = tid; j < nvec4; j += kThreads) {
        const uint4 v = acc4[j];
        const uint32_t e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t a = e[u], i = unperm32(4 * j + u);      // i = doc offset of this position
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
    }
    __syncthreads();
    if (collect) {
        uint32_t n2 = 2;
        while (n2 < cand_count) n2 <<= 1;
        for (uint32_t i = cand_count + tid; i < n2; i += kThreads) U->sel.cand[i] = 0;
        bitonic_sort_desc(U->sel.cand, n2);
        for (uint32_t r = tid; r < remaining; r += kThreads) {
            uint32_t i = (uint32_t)(~U->sel.cand[r]);
            keys_out[n_def + r] = make_key(__uint2float_rn(acc[perm32(i)]), gdoc0 + i);
        }
    } else {
        // more than kCandMax docs tie exactly at the threshold key: take the `remaining` smallest doc ids (S8)
        const uint32_t per = ((tl + kWarps - 1) / kWarps + 31) & ~31u;
        const uint32_t lo = warp * per, hi = min(tl, lo + per);
        uint32_t c = 0;
        for (uint32_t base = lo; base < hi; base += 32) {
            uint32_t i = base + lane;
            uint32_t a = i < hi ? acc[perm32(i)] : 0;
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
            uint32_t a = i < hi ? acc[perm32(i)] : 0;
            bool f = a && (kmax - MSR_SELKEY(a)) == pref;
            uint32_t b = __ballot_sync(0xffffffffu, f);
            uint32_t r = run + __popc(b & ((1u << lane) - 1u));
            if (f && r < remaining) keys_out[n_def + r] = make_key(__uint2float_rn(a), gdoc0 + i);
            run += __popc(b);
        }
    }
#undef MSR_SELKEY
    if (tid == 0) *cnt_out = kq;
    // publish this tile's exact top-k as the query's threshold list (keys_out[k-1] is its smallest key)
    if (P.use_theta && kq == P.k) {
        __syncthreads();
        for (uint32_t r = tid; r < P.k; r += kThreads) tlist[r] = keys_out[r];
        if (tid == 0) theta_raise(P, theta_p, q, (unsigned long long)keys_out[P.k - 1]);
    }
}

}  // namespace msr
