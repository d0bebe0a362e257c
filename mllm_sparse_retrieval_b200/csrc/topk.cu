// topk.cu -- kernel (b) stand-alone + list merge.
//
//  * topk_merge_kernel: one CTA per query merges n_lists candidate lists (per-tile lists of the sparse scorer,
//    or the per-GPU lists an NCCL all-gather delivered -- SURVEY.md 8(e)) into the final k by the 64-bit key
//    (float32 score desc, doc id asc): rounds of "append lists, bitonic-sort in shared memory, keep k".
//  * topk_rows_kernel: radix-select (11-bit digits on the distance to the row maximum) over a dense fp32 score
//    row with the same deterministic tie-break; used by the dense path's exact pass and exposed as msr_topk_rows.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "common.cuh"

namespace msr {

constexpr int kMergeThreads = 256;
constexpr int kMaxPeers = 16;

struct MergeParams {
    uint32_t nq, n_lists, stride, k, mode;   // mode 0: packed keys [nq][n_lists][stride]; 1: arrays [n_lists][nq][stride]
    const uint64_t* keys;
    const float* scores;
    const int32_t* docs;
    const uint32_t* cnts;
    float* out_scores;
    int32_t* out_docs;
    uint32_t* out_counts;
    uint32_t cap;                            // shared-memory capacity in keys (power of two, >= k + stride)
    size_t list_stride;                      // mode 1: elements between consecutive lists (0 = nq * stride, packed)
    size_t cnt_list_stride;                  // mode 1: elements between the count rows of consecutive lists (0 = nq)
    // Fused exchange (doc-sharded search): the merged rows are ALSO stored into the result slots of up to kMaxPeers
    // other GPUs over NVLink (peer pointers from symmetric memory), so no separate all-gather follows.
    uint32_t n_peers;
    float* peer_scores[kMaxPeers];
    int32_t* peer_docs[kMaxPeers];
    uint32_t* peer_counts[kMaxPeers];
    // Fused barrier (msr_topk_merge_exchange): before it reads the lists, the kernel tells every peer GPU that THIS GPU's
    // rows of step `flag_value` have been stored (flag_peers[p][my_rank] = flag_value, release at system scope, after the
    // producing kernel of the same stream has completed) and waits until every peer has said the same (flag_local[rank]).
    uint32_t n_flag_peers, my_rank, flag_value;
    // doc-sharded search with shared thresholds: keys below the collection-wide lower bound of the k-th key cannot be in
    // the final top k and are dropped here already (shorter rows to exchange, a much smaller final merge)
    const unsigned long long* theta;         // [nq] or NULL
    const unsigned long long* quant;         // [quant_world][quant_cap] (+ first query of this launch) or NULL
    uint32_t quant_world, quant_cap;
    // mode 1, lists sorted best-first (the rows every libmsr search returns): with r = ceil(k / n_lists), the smallest of the
    // lists' r-th keys has at least k keys at or above it, so everything below is dropped before the sort.  For doc-sharded
    // lists of balanced shards that leaves little more than k of the n_lists * k keys.
    uint32_t sorted_lists;
    // wide path of the sparse search (sparse_kernel.cuh, wide_score_cta): wide_ctas extra CTAs at the end of the grid merge
    // the per-tile KEY ARRAYS of the queries scored there (lists of tile_docs keys, 0 = no hit) and store each row at its
    // query's place in the batch (row = wide_q[...] - wide_row0 relative to out_* / peer_*).
    uint32_t wide_ctas, wide_first, wide_n_lists, wide_stride;
    const uint32_t* wide_q;
    const uint32_t* n_wide;
    const uint64_t* wide_keys;
    const uint32_t* wide_cnt;
    long long wide_row0;
    uint32_t vec_ok;                         // set by merge_launch: k % 4 == 0 and every output row base is 16-byte aligned
    uint32_t* flag_local;                    // [world] written by the peers
    uint32_t* flag_peers[kMaxPeers];         // the peers' flag arrays
    uint32_t flag_peer_rank[kMaxPeers];      // rank of peer p (index into flag_local)
};

// One-way barrier of the fused exchange (msr_topk_merge_exchange), one warp: lane p tells peer p that THIS GPU's rows of
// step `value` have been stored (the kernel that stored them precedes this one on the stream) and waits until peer p has
// said the same.  Flags only grow, so nothing is ever reset.  (Measured: the same wait done by every CTA of the merge
// kernel cost 0.05 ms per peer at 10 K queries -- 70 K system-scope acquires -- against ~3 us for this launch.)
struct FlagParams {
    uint32_t n_peers, my_rank, value;
    uint32_t* local;                         // [world] written by the peers
    uint32_t* peers[kMaxPeers];              // the peers' flag arrays
    uint32_t peer_rank[kMaxPeers];           // rank of peer p (index into local)
};
__global__ void exchange_flags_kernel(const FlagParams F) {
    const uint32_t p = threadIdx.x;
    if (p >= F.n_peers) return;
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(F.peers[p] + F.my_rank), "r"(F.value) : "memory");
    const uint32_t* f = F.local + F.peer_rank[p];
    uint32_t v;
    do {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
    } while ((int32_t)(v - F.value) < 0);
}

__device__ __forceinline__ uint32_t pow2_ceil(uint32_t x) {
    uint32_t n = 2;
    while (n < x) n <<= 1;
    return n;
}

// key number idx (0 <= idx < n_lists * stride) of query q, 0 when the slot is beyond its list's count
__device__ __forceinline__ uint64_t merge_key_at(const MergeParams& P, uint32_t q, uint32_t idx) {
    const uint32_t j = idx / P.stride, i = idx - j * P.stride;
    const size_t li = P.mode == 0 ? (size_t)q * P.n_lists + j : (size_t)j * P.nq + q;
    const size_t ci = (P.mode == 1 && P.cnt_list_stride) ? (size_t)j * P.cnt_list_stride + q : li;
    if (i >= min(P.cnts[ci], P.stride)) return 0ull;
    if (P.mode == 0) return P.keys[li * P.stride + i];
    const size_t ro = (P.mode == 1 && P.list_stride) ? (size_t)j * P.list_stride + (size_t)q * P.stride : li * P.stride;
    return make_key(P.scores[ro + i], (uint32_t)P.docs[ro + i]);
}

// One CTA per query.  Common case (large batches: the shared threshold keeps the per-tile lists short): all candidates fit
// the shared buffer, one bitonic sort.  Otherwise (small batches, deep k: every tile delivers a full list) the k-th largest
// key is located by an MSB-first radix select over the lists where they lie (8-bit digits; the lists stay in L2), and only
// the keys at or above it are gathered and sorted -- 41 lists x 1000 keys took 600 us by repeated sorting, ~30 us this way.
__device__ __forceinline__ void merge_body(const MergeParams& P, const uint32_t q, const long long orow, uint64_t* buf) {
    const uint32_t tid = threadIdx.x, nthr = blockDim.x;
    uint32_t fill = 0;
    __shared__ uint32_t s_total, s_fill, s_hist[256], s_sel[4];
    __shared__ unsigned long long s_floor, s_rmin;
    if (tid == 0) {
        s_total = 0; s_fill = 0;
        unsigned long long t = 0;
        if (P.theta) {
            t = *(volatile const unsigned long long*)(P.theta + q);
            if (P.quant) {
                unsigned long long m = ~0ull;
                for (uint32_t g = 0; g < P.quant_world; ++g) m = min(m, *(volatile const unsigned long long*)(P.quant + (size_t)g * P.quant_cap + q));
                t = max(t, m);
            }
        }
        s_floor = t;
        s_rmin = ~0ull;
    }
    if (P.mode == 1 && P.sorted_lists && P.n_lists > 1) {
        // one thread per list (independent loads), smallest r-th key by a shared-memory atomicMin
        __syncthreads();
        const uint32_t r = (P.k + P.n_lists - 1) / P.n_lists;
        for (uint32_t j = tid; j < P.n_lists; j += nthr) {
            const size_t ci = P.cnt_list_stride ? (size_t)j * P.cnt_list_stride + q : (size_t)j * P.nq + q;
            const size_t ro = P.list_stride ? (size_t)j * P.list_stride + (size_t)q * P.stride : ((size_t)j * P.nq + q) * P.stride;
            const unsigned long long x = (min(P.cnts[ci], P.stride) >= r)
                                             ? (unsigned long long)make_key(P.scores[ro + r - 1], (uint32_t)P.docs[ro + r - 1]) : 0ull;
            atomicMin(&s_rmin, x);
        }
        __syncthreads();
        if (tid == 0) s_floor = max(s_floor, s_rmin);
    }
    __syncthreads();
    {
        uint32_t mine = 0;
        for (uint32_t j = tid; j < P.n_lists; j += nthr) {
            const size_t li = P.mode == 0 ? (size_t)q * P.n_lists + j : (size_t)j * P.nq + q;
            const size_t ci = (P.mode == 1 && P.cnt_list_stride) ? (size_t)j * P.cnt_list_stride + q : li;
            mine += min(P.cnts[ci], P.stride);
        }
        mine = warp_reduce_add(mine);
        if ((tid & 31) == 0 && mine) atomicAdd(&s_total, mine);
    }
    __syncthreads();
    const uint32_t total = s_total;
    if (total <= P.cap) {
        // every warp reserves room for its lists with one shared atomic and copies them (the order is irrelevant, it is sorted next)
        const uint32_t lane = tid & 31, warp = tid >> 5;
        for (uint32_t j = warp; j < P.n_lists; j += nthr / 32) {
            const size_t li = P.mode == 0 ? (size_t)q * P.n_lists + j : (size_t)j * P.nq + q;
            const size_t ci = (P.mode == 1 && P.cnt_list_stride) ? (size_t)j * P.cnt_list_stride + q : li;
            const size_t ro = (P.mode == 1 && P.list_stride) ? (size_t)j * P.list_stride + (size_t)q * P.stride : li * P.stride;
            const uint32_t c = min(P.cnts[ci], P.stride);
            if (c == 0) continue;
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(&s_fill, c);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (P.mode == 0) {
                const uint64_t* src = P.keys + li * P.stride;
                const unsigned long long floor_key = total <= 8 * nthr ? s_floor : 0ull;    // pruned only when the compaction below runs
                for (uint32_t i = lane; i < c; i += 32) { const uint64_t x = src[i]; buf[base + i] = x >= floor_key ? x : 0ull; }
            } else if (P.sorted_lists && s_floor && total <= 8 * nthr) {
                // sorted list: the keys at or above the floor are a prefix; the rest is never copied
                const float* ss = P.scores + ro;
                const int32_t* dd = P.docs + ro;
                const unsigned long long floor_key = s_floor;
                uint32_t kept = 0;
                for (uint32_t i0 = 0; i0 < c; i0 += 32) {
                    const uint32_t i = i0 + lane;
                    const uint64_t x = i < c ? make_key(ss[i], (uint32_t)dd[i]) : 0ull;
                    const uint32_t m = __ballot_sync(0xffffffffu, i < c && x >= floor_key);
                    if (x >= floor_key && i < c) buf[base + kept + __popc(m & ((1u << lane) - 1u))] = x;
                    kept += __popc(m);
                    if (m != 0xffffffffu) break;
                }
                // the slots reserved beyond `kept` stay unused: zero them so they sort to the end
                for (uint32_t i = kept + lane; i < c; i += 32) buf[base + i] = 0ull;
            } else {
                const float* ss = P.scores + ro;
                const int32_t* dd = P.docs + ro;
                for (uint32_t i = lane; i < c; i += 32) buf[base + i] = make_key(ss[i], (uint32_t)dd[i]);
            }
        }
        __syncthreads();
        fill = total;
        if ((P.mode == 0 || P.sorted_lists) && s_floor && total <= 8 * nthr) {
            // compact the survivors (keys below the floor were zeroed): a much shorter sort
            uint64_t mine[8];
            uint32_t cnt = 0;
            for (uint32_t i = tid, j = 0; i < total && j < 8; i += nthr, ++j) { mine[j] = buf[i]; cnt += mine[j] != 0ull; }
            __syncthreads();
            if (tid == 0) s_fill = 0;
            __syncthreads();
            uint32_t pos = cnt ? atomicAdd(&s_fill, cnt) : 0u;
            for (uint32_t i = tid, j = 0; i < total && j < 8; i += nthr, ++j) if (mine[j]) buf[pos++] = mine[j];
            __syncthreads();
            fill = s_fill;
        }
    } else {
        // radix select of the k-th largest key (keys of real hits are distinct and non-zero)
        const uint32_t n_slots = P.n_lists * P.stride;
        uint32_t need = P.k, k_sel = P.k;                       // total > cap >= 2k slots, but some may hold no key (see pass 0)
        uint64_t prefix = 0;                                    // the digits fixed so far (top 8 * pass bits)
        uint32_t pass = 0;
        for (;; ++pass) {
            for (uint32_t i = tid; i < 256; i += nthr) s_hist[i] = 0;
            __syncthreads();
            const uint32_t shift = 56 - 8 * pass;
            for (uint32_t idx = tid; idx < n_slots; idx += nthr) {
                const uint64_t key = merge_key_at(P, q, idx);
                if (key && (pass == 0 || (key >> (shift + 8)) == prefix)) atomicAdd(&s_hist[(key >> shift) & 255u], 1u);
            }
            __syncthreads();
            if (tid < 32) {                                     // lane l owns bins 255-8l .. 248-8l (descending)
                uint32_t h[8], sum = 0;
#pragma unroll
                for (int i = 0; i < 8; ++i) { h[i] = s_hist[255 - 8 * tid - i]; sum += h[i]; }
                const uint32_t incl = warp_incl_scan(sum, tid);
                uint32_t before = incl - sum;                            // keys in higher bins
                // fewer real keys than asked for (key arrays of the wide path hold a 0 for every doc without a hit): take all
                const uint32_t have = __shfl_sync(0xffffffffu, incl, 31);
                if (pass == 0) need = min(need, have);
                if (tid == 0) { s_sel[3] = need; if (need == 0) { s_sel[0] = 0; s_sel[1] = 0; s_sel[2] = 0; } }
                if (need && before < need && before + sum >= need) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        if (before < need && before + h[i] >= need) { s_sel[0] = 255 - 8 * tid - i; s_sel[1] = before; s_sel[2] = h[i]; }
                        before += h[i];
                    }
                }
            }
            __syncthreads();
            if (pass == 0) { need = s_sel[3]; k_sel = need; }
            if (need == 0) break;
            const uint32_t bin = s_sel[0], greater = s_sel[1], cnt = s_sel[2];
            prefix = (prefix << 8) | bin;
            // keys above the bin are certain; stop as soon as they and the bin's keys fit the buffer together
            if ((k_sel - need) + greater + cnt <= P.cap || pass == 7) break;
            need -= greater;
            __syncthreads();
        }
        const uint32_t shift = 56 - 8 * pass;
        for (uint32_t idx = tid; idx < n_slots; idx += nthr) {
            const uint64_t key = merge_key_at(P, q, idx);
            if (key && (key >> shift) >= prefix) {
                const uint32_t pos = atomicAdd(&s_fill, 1u);
                if (pos < P.cap) buf[pos] = key;
            }
        }
        __syncthreads();
        fill = min(s_fill, P.cap);
    }
    if (fill > 1) {
        const uint32_t n2 = pow2_ceil(fill);
        for (uint32_t i = fill + tid; i < n2; i += nthr) buf[i] = 0;
        bitonic_sort_desc(buf, n2);
    } else {
        __syncthreads();
    }
    // keys are sorted best-first; zero keys (pruned slots, docs without a hit in a wide-path key array) sort to the end
    uint32_t n = min(fill, P.k);
    {
        uint32_t lo = 0, hi = n;                               // first zero in buf[0, n)
        while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (buf[mid]) lo = mid + 1; else hi = mid; }
        n = lo;
    }
    if (P.n_peers && P.vec_ok) {
        // rows leave as 16-byte stores: keys -> (score, doc) in place, then one uint4 per 4 ranks and destination
        __syncthreads();
        uint32_t* pair = reinterpret_cast<uint32_t*>(buf);
        for (uint32_t r = tid; r < P.k; r += nthr) {
            const bool ok = r < n;
            const uint64_t key = ok ? buf[r] : 0;
            pair[2 * r] = ok ? __float_as_uint(key_score(key)) : 0u;
            pair[2 * r + 1] = ok ? key_doc(key) : 0xFFFFFFFFu;
        }
        __syncthreads();
        const uint32_t nv = P.k >> 2, n_dst = P.n_peers + 1;
        for (uint32_t i = tid; i < 2 * nv * n_dst; i += nthr) {
            const uint32_t v = i % nv, what = (i / nv) & 1u, dst = i / (2 * nv);
            const uint32_t* src = pair + 8 * v + what;
            const uint4 x = make_uint4(src[0], src[2], src[4], src[6]);
            void* base = what ? (dst ? (void*)P.peer_docs[dst - 1] : (void*)P.out_docs)
                              : (dst ? (void*)P.peer_scores[dst - 1] : (void*)P.out_scores);
            reinterpret_cast<uint4*>(reinterpret_cast<uint32_t*>(base) + orow * P.k)[v] = x;
        }
    } else {
        for (uint32_t r = tid; r < P.k; r += nthr) {
            const bool ok = r < n;
            const uint64_t key = ok ? buf[r] : 0;
            const float sc = ok ? key_score(key) : 0.0f;
            const int32_t dc = ok ? (int32_t)key_doc(key) : -1;
            P.out_scores[orow * P.k + r] = sc;
            P.out_docs[orow * P.k + r] = dc;
            for (uint32_t p = 0; p < P.n_peers; ++p) {           // the exchange: same row into every peer's slot
                P.peer_scores[p][orow * P.k + r] = sc;
                P.peer_docs[p][orow * P.k + r] = dc;
            }
        }
    }
    if (tid == 0) {
        P.out_counts[orow] = n;
        for (uint32_t p = 0; p < P.n_peers; ++p) P.peer_counts[p][orow] = n;
    }
}

__global__ void __launch_bounds__(1024)
topk_merge_kernel(const MergeParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* buf = reinterpret_cast<uint64_t*>(smem_raw);
    if (blockIdx.x < P.nq) {
        merge_body(P, blockIdx.x, (long long)blockIdx.x, buf);
        return;
    }
    // wide path: this CTA's "lists" are the key arrays of one flagged query, one per tile
    const uint32_t w = blockIdx.x - P.nq;
    if (P.wide_first + w >= *P.n_wide) return;
    MergeParams W = P;
    W.mode = 0; W.nq = P.wide_ctas; W.n_lists = P.wide_n_lists; W.stride = P.wide_stride;
    W.keys = P.wide_keys; W.cnts = P.wide_cnt; W.theta = nullptr; W.quant = nullptr; W.sorted_lists = 0;
    merge_body(W, w, (long long)P.wide_q[P.wide_first + w] - P.wide_row0, buf);
}

static int merge_launch(MergeParams P, cudaStream_t s) {
    // buffer: at least 2k keys (the select path gathers k + the k-th key's bin), 2048 at least, a power of two
    uint32_t cap = 2048;
    while (cap < 2 * P.k) cap <<= 1;
    if (cap > 8192) { set_error("topk merge: k = %u exceeds %u", P.k, MSR_MAX_K); return MSR_ERR_INVALID; }
    P.cap = cap;
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
    P.vec_ok = (P.k & 3u) == 0 && al16(P.out_scores) && al16(P.out_docs);
    for (uint32_t p = 0; p < P.n_peers; ++p) P.vec_ok = P.vec_ok && al16(P.peer_scores[p]) && al16(P.peer_docs[p]);
    const size_t smem = (size_t)cap * 8;
    static std::atomic<uint64_t> attr_set{0};
    if (attr_needed(attr_set)) {
        MSR_CUDA_CHECK(cudaFuncSetAttribute(topk_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 8));
        attr_mark(attr_set);
    }
    // few queries cannot fill the GPU with CTAs: give each one more threads (the sort and the select scale with them)
    const int threads = P.nq <= 148 ? 1024 : (P.nq <= 2048 ? 512 : kMergeThreads);
    topk_merge_kernel<<<P.nq + P.wide_ctas, threads, smem, s>>>(P);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    return MSR_OK;
}

int topk_merge_keys_launch(uint32_t nq, uint32_t n_lists, uint32_t stride, const uint64_t* keys, const uint32_t* cnts,
                           uint32_t k, float* out_scores, int32_t* out_docs, uint32_t* out_counts, cudaStream_t s,
                           uint32_t n_peers, float* const* peer_scores, int32_t* const* peer_docs,
                           uint32_t* const* peer_counts, size_t peer_row0, const unsigned long long* theta,
                           const unsigned long long* quant, uint32_t quant_world, uint32_t quant_cap, uint32_t wide_ctas,
                           uint32_t wide_first, const uint32_t* wide_q, const uint32_t* n_wide, const uint64_t* wide_keys,
                           const uint32_t* wide_cnt, uint32_t wide_n_lists, uint32_t wide_stride) {
    MergeParams P{};
    P.theta = theta; P.quant = quant; P.quant_world = quant_world; P.quant_cap = quant_cap;
    P.wide_ctas = wide_ctas; P.wide_first = wide_first; P.wide_q = wide_q; P.n_wide = n_wide; P.wide_keys = wide_keys;
    P.wide_cnt = wide_cnt; P.wide_n_lists = wide_n_lists; P.wide_stride = wide_stride; P.wide_row0 = (long long)peer_row0;
    P.nq = nq; P.n_lists = n_lists; P.stride = stride; P.k = k; P.mode = 0;
    P.keys = keys; P.cnts = cnts; P.out_scores = out_scores; P.out_docs = out_docs; P.out_counts = out_counts;
    P.n_peers = n_peers;
    for (uint32_t p = 0; p < n_peers; ++p) {                      // peer_row0: first query of this launch in the batch
        P.peer_scores[p] = peer_scores[p] + peer_row0 * k;
        P.peer_docs[p] = peer_docs[p] + peer_row0 * k;
        P.peer_counts[p] = peer_counts[p] + peer_row0;
    }
    return merge_launch(P, s);
}

// ---------------------------------------------------------------------------------------------------------
// stand-alone row top-k
// ---------------------------------------------------------------------------------------------------------
constexpr int kRowThreads = 512;
constexpr int kRowWarps = kRowThreads / 32;
constexpr int kRowBins = 2048;
constexpr int kRowBinBits = 11;
constexpr int kRowCandMax = 1024;
constexpr size_t kRowSumBytes = 32 * 1024 + 16;   // per-vector maxima kept by pass 1 (u16 each)

struct RowsParams {
    uint32_t nq, n, k;
    int32_t col_base;
    const void* scores;          // rows of float, __half or __nv_bfloat16 (kernel template); row_stride in elements
    size_t row_stride;
    float* out_scores;
    int32_t* out_docs;
    uint32_t* out_counts;
    const uint32_t* nq_dev;      // optional: rows [nq_off, *nq_dev) exist (count kept on the device); others are skipped
    uint32_t nq_off;
    uint32_t sum_off, sum_cap;   // per-vector maxima of pass 1 (u16 key prefixes) at this byte offset of the dynamic shared
                                 // memory, room for sum_cap vectors (0: none)
};

struct RowsSmem {
    uint32_t hist[kRowBins];
    uint64_t cand[kRowCandMax];
    uint32_t wsum[kRowWarps];
    uint32_t vmax, hits, out_cnt, cand_cnt, sel_b, sel_clt, sel_bcount;
};

__device__ __forceinline__ uint32_t rows_block_excl_scan(uint32_t v, uint32_t* wsum, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = warp_incl_scan(v, lane);
    __syncthreads();
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kRowWarps; ++w) {
        uint32_t x = wsum[w];
        if (w < warp) base += x;
        tot += x;
    }
    *total = tot;
    return base + inc - v;
}

// key 0 is reserved for "not a candidate" (-inf / NaN are never returned)
__device__ __forceinline__ uint32_t row_key(float x) { return (x > -INFINITY) ? f2ord(x) : 0u; }

// rows are read in place in their own type: fp16 / bf16 activation rows of the quantiser need no fp32 staging copy
// (2.6 GB read + 5.2 GB written + 5.2 GB per select pass at 10 K x 131 072), and a 256 KB row stays in L2 between the passes
__device__ __forceinline__ float row_load(const float* r, uint32_t i) { return r[i]; }
__device__ __forceinline__ float row_load(const __half* r, uint32_t i) { return __half2float(r[i]); }
__device__ __forceinline__ float row_load(const __nv_bfloat16* r, uint32_t i) { return __bfloat162float(r[i]); }

#ifndef MSR_ROW_INFLIGHT
#define MSR_ROW_INFLIGHT 2
#endif
#ifndef MSR_ROW_MINBLOCKS
#define MSR_ROW_MINBLOCKS 4      // CTAs of 512 threads per SM the register allocation leaves room for (A/B: profiles/r2_rows_ab2.txt)
#endif
// 16 bytes of a row -> floats
__device__ __forceinline__ void row_unpack(const float*, const uint4 x, float (&v)[4]) {
    v[0] = __uint_as_float(x.x); v[1] = __uint_as_float(x.y); v[2] = __uint_as_float(x.z); v[3] = __uint_as_float(x.w);
}
__device__ __forceinline__ void row_unpack(const __half*, const uint4 x, float (&v)[8]) {
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        v[2 * j] = __half2float(__ushort_as_half((unsigned short)(w[j] & 0xFFFFu)));
        v[2 * j + 1] = __half2float(__ushort_as_half((unsigned short)(w[j] >> 16)));
    }
}
__device__ __forceinline__ void row_unpack(const __nv_bfloat16*, const uint4 x, float (&v)[8]) {
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) { v[2 * j] = __uint_as_float(w[j] << 16); v[2 * j + 1] = __uint_as_float(w[j] & 0xFFFF0000u); }
}

// f(i, row[i]) for every i < n, each element once, over the whole CTA: 16-byte loads (MSR_ROW_INFLIGHT in flight per thread) over the
// aligned body of the row, scalar loads for the few elements before and after it.  A pass over a 131 072-wide fp16 row
// with one 2-byte load per thread and step moved 64 bytes per warp instruction; this moves 512.
// pre(j): load and visit 16-byte vector j of the body at all?   post(j, m): m = fmaxf over the vector's elements (NaN skipped).
template <typename T, typename F, typename Pre, typename Post>
__device__ __forceinline__ void row_for_each(const T* __restrict__ row, const uint32_t n, const uint32_t tid, F&& f, Pre&& pre,
                                             Post&& post) {
    constexpr uint32_t VEC = 16 / sizeof(T);
    const uint32_t mis = (uint32_t)((reinterpret_cast<uintptr_t>(row) & 15u) / sizeof(T));
    const uint32_t head = mis ? min(n, VEC - mis) : 0u;
    if (tid < head) f(tid, row_load(row, tid));
    const uint32_t nv = (n - head) / VEC;
    const uint4* __restrict__ v = reinterpret_cast<const uint4*>(row + head);
    constexpr int kInFlight = MSR_ROW_INFLIGHT;                    // independent 16-byte loads per thread and step
    for (uint32_t j = tid; j < nv; j += kInFlight * kRowThreads) {
        uint4 x[kInFlight];
        bool on[kInFlight];
#pragma unroll
        for (int t = 0; t < kInFlight; ++t) {
            const uint32_t jt = j + t * kRowThreads;
            on[t] = jt < nv && pre(jt);
            x[t] = on[t] ? __ldg(v + jt) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int t = 0; t < kInFlight; ++t) {
            const uint32_t jt = j + t * kRowThreads;
            if (on[t]) {
                float e[VEC];
                row_unpack(row, x[t], e);
                float m = -INFINITY;
#pragma unroll
                for (uint32_t u = 0; u < VEC; ++u) { f(head + jt * VEC + u, e[u]); m = fmaxf(m, e[u]); }
                post(jt, m);
            }
        }
    }
    const uint32_t done = head + nv * VEC;
    if (done + tid < n) f(done + tid, row_load(row, done + tid));
}
template <typename T, typename F>
__device__ __forceinline__ void row_for_each(const T* __restrict__ row, const uint32_t n, const uint32_t tid, F&& f) {
    row_for_each(row, n, tid, f, [](uint32_t) { return true; }, [](uint32_t, float) {});
}
// number of 16-byte body vectors row_for_each visits
template <typename T>
__device__ __forceinline__ uint32_t row_body_vectors(const T* row, const uint32_t n) {
    constexpr uint32_t VEC = 16 / sizeof(T);
    const uint32_t mis = (uint32_t)((reinterpret_cast<uintptr_t>(row) & 15u) / sizeof(T));
    const uint32_t head = mis ? min(n, VEC - mis) : 0u;
    return (n - head) / VEC;
}

template <typename T>
__global__ void __launch_bounds__(kRowThreads, MSR_ROW_MINBLOCKS)
topk_rows_kernel(const RowsParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    RowsSmem* S = reinterpret_cast<RowsSmem*>(smem_raw);
    uint64_t* outk = reinterpret_cast<uint64_t*>(smem_raw + sizeof(RowsSmem));   // pow2_ceil(k) keys
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t q = blockIdx.x;
    if (P.nq_dev && P.nq_off + q >= *P.nq_dev) return;
    const T* row = reinterpret_cast<const T*>(P.scores) + (size_t)q * P.row_stride;
    const uint32_t n = P.n;
    // Pass 1 leaves the top 16 key bits of every 16-byte vector's maximum in shared memory (32 KB for a 131 072-wide fp16
    // row); the collecting pass then loads only the vectors that can hold a candidate -- for activation rows a handful, so
    // the row is read from HBM once instead of twice (the resident rows of all CTAs do not fit L2).
    uint16_t* vsum = reinterpret_cast<uint16_t*>(smem_raw + P.sum_off);
    const bool use_sum = P.sum_cap != 0 && row_body_vectors(row, n) <= P.sum_cap;
    if (tid == 0) { S->vmax = 0; S->hits = 0; S->out_cnt = 0; S->cand_cnt = 0; }
    for (uint32_t i = tid; i < kRowBins; i += kRowThreads) S->hist[i] = 0;
    __syncthreads();
    constexpr uint32_t kCoarseShift = 32 - kRowBinBits;          // coarse bin = top 11 bits of the key: sign, exponent, 2 mantissa bits
    {
        // One pass finds the maximum, the count of the rankable values (x > -inf; NaN fails the test) AND a first histogram,
        // over the ABSOLUTE key bits (index 0 = largest values): when the bin of the k-th value holds few enough candidates
        // the collecting pass follows at once -- two reads of the row instead of three.  Exact +0.0 (relu'd activation rows
        // are mostly zeros) is counted in a register: one shared atomic per warp instead of one per element on one address.
        float mf = -INFINITY;
        uint32_t h = 0, zc = 0;
        row_for_each(row, n, tid, [&](uint32_t, float x) {
            if (__float_as_uint(x) == 0u) { ++zc; return; }
            if (x > -INFINITY) {
                mf = fmaxf(mf, x);                              // max.f32 orders +0 above -0, as the keys do
                ++h;
                atomicAdd(&S->hist[(kRowBins - 1) - (f2ord(x) >> kCoarseShift)], 1u);
            }
        }, [](uint32_t) { return true; },
        [&](uint32_t j, float m) { if (use_sum) vsum[j] = (uint16_t)(row_key(m) >> 16); });
        if (zc) mf = fmaxf(mf, 0.0f);
        h += zc;
        uint32_t m = row_key(mf);
        m = warp_reduce_max(m);
        h = warp_reduce_add(h);
        zc = warp_reduce_add(zc);
        if (lane == 0) {
            if (m) atomicMax(&S->vmax, m);
            if (h) atomicAdd(&S->hits, h);
            if (zc) atomicAdd(&S->hist[(kRowBins - 1) - (f2ord(0.0f) >> kCoarseShift)], zc);
        }
    }
    __syncthreads();
    const uint32_t kmax = S->vmax, hits = S->hits;
    const uint32_t kq = min(P.k, hits);
    const uint32_t n2out = pow2_ceil(max(kq, 2u));
    for (uint32_t i = tid; i < n2out; i += kRowThreads) outk[i] = 0;
    __syncthreads();
    if (hits > 0 && hits <= P.k) {
        row_for_each(row, n, tid, [&](uint32_t i, float x) {
            if (row_key(x)) outk[atomicAdd(&S->out_cnt, 1u)] = make_key(x, (uint32_t)(P.col_base + (int32_t)i));
        });
    } else if (hits > 0) {
        uint32_t remaining = kq, sh_prev = 32 - __clz(kmax), pref = 0, s = 0, cand_count = 0;
        // keys >= k_def are certainly among the top kq, keys in [k_cand, k_def) are the candidates for the remaining places
        uint64_t k_def = 0;
        uint32_t k_cand = 0;
        bool coarse;
        {
            // the coarse histogram of pass 1: bin of the kq-th largest value
            uint32_t h4[4], ls = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) { h4[j] = S->hist[tid * 4 + j]; ls += h4[j]; }
            uint32_t total;
            uint32_t ex = rows_block_excl_scan(ls, S->wsum, &total);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (ex < remaining && ex + h4[j] >= remaining) {
                    S->sel_b = tid * 4 + j; S->sel_clt = ex; S->sel_bcount = h4[j];
                }
                ex += h4[j];
            }
            __syncthreads();
            coarse = S->sel_bcount <= (uint32_t)kRowCandMax;
            if (coarse) {
                const uint32_t bin = (kRowBins - 1) - S->sel_b;
                remaining -= S->sel_clt;
                cand_count = S->sel_bcount;
                k_def = ((uint64_t)bin + 1) << kCoarseShift;
                k_cand = max(bin << kCoarseShift, 1u);
            }
            __syncthreads();                                    // sel_* and hist are rewritten below
        }
        while (!coarse) {
            s = sh_prev > kRowBinBits ? sh_prev - kRowBinBits : 0;
            const uint32_t width = sh_prev - s;
            for (uint32_t i = tid; i < kRowBins; i += kRowThreads) S->hist[i] = 0;
            __syncthreads();
            // exact zeros (relu'd activation rows are mostly zeros) are counted in a register: one shared atomic per warp
            // instead of one per element on a single address
            const uint32_t kzero = f2ord(0.0f);
            uint32_t zc = 0;
            row_for_each(row, n, tid, [&](uint32_t, float x) {
                if (__float_as_uint(x) == 0u) { ++zc; return; }       // +0.0 (its key is kzero)
                const uint32_t kx = row_key(x);
                if (kx) {
                    const uint32_t d = kmax - kx;
                    const uint32_t hi = sh_prev >= 32 ? 0u : (d >> sh_prev);
                    if (hi == pref) atomicAdd(&S->hist[(d >> s) & ((1u << width) - 1u)], 1u);
                }
            });
            zc = warp_reduce_add(zc);
            if (lane == 0 && zc && kzero <= kmax) {
                const uint32_t d = kmax - kzero;
                const uint32_t hi = sh_prev >= 32 ? 0u : (d >> sh_prev);
                if (hi == pref) atomicAdd(&S->hist[(d >> s) & ((1u << width) - 1u)], zc);
            }
            __syncthreads();
            {
                uint32_t h4[4], ls = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) { h4[j] = S->hist[tid * 4 + j]; ls += h4[j]; }
                uint32_t total;
                uint32_t ex = rows_block_excl_scan(ls, S->wsum, &total);
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
            if (cand_count <= (uint32_t)kRowCandMax || s == 0) {
                // d >> s < pref  <=>  key > kmax - (pref << s);   d >> s == pref  <=>  key >= kmax - (((pref + 1) << s) - 1)
                const uint64_t dlo = (uint64_t)pref << s, dhi = (((uint64_t)pref + 1) << s) - 1;
                k_def = dlo > (uint64_t)kmax ? 0ull : (uint64_t)kmax - dlo + 1;
                k_cand = dhi >= (uint64_t)kmax ? 1u : max(kmax - (uint32_t)dhi, 1u);
                break;
            }
        }
        const uint32_t n_def = kq - remaining;
        const bool collect = cand_count <= (uint32_t)kRowCandMax;
        // one float compare rejects almost every element: keys are monotone in the value, NaN fails the compare, and the
        // bound is at least the key of -FLT_MAX, so -inf (key 0, never returned) fails it too
        k_cand = max(k_cand, 0x00800000u);
        const float xlow = ord2f(k_cand);
        const uint32_t kc16 = k_cand >> 16;                       // a vector whose maximum's key prefix is below holds no candidate
        row_for_each(row, n, tid, [&](uint32_t i, float xv) {
            if (!(xv >= xlow)) return;
            const uint32_t kx = row_key(xv);
            if ((uint64_t)kx >= k_def) outk[atomicAdd(&S->out_cnt, 1u)] = make_key(xv, (uint32_t)(P.col_base + (int32_t)i));
            else if (kx >= k_cand && collect) S->cand[atomicAdd(&S->cand_cnt, 1u)] = ~(((uint64_t)(kmax - kx) << 32) | i);
        }, [&](uint32_t j) { return !use_sum || (uint32_t)vsum[j] >= kc16; }, [](uint32_t, float) {});
        __syncthreads();
        if (collect) {
            uint32_t n2 = pow2_ceil(cand_count);
            for (uint32_t i = cand_count + tid; i < n2; i += kRowThreads) S->cand[i] = 0;
            bitonic_sort_desc(S->cand, n2);
            for (uint32_t r = tid; r < remaining; r += kRowThreads) {
                uint32_t i = (uint32_t)(~S->cand[r]);
                outk[n_def + r] = make_key(row_load(row, i), (uint32_t)(P.col_base + (int32_t)i));
            }
        } else {
            const uint32_t per = ((n + kRowWarps - 1) / kRowWarps + 31) & ~31u;
            const uint32_t lo = min(n, warp * per), hi = min(n, lo + per);
            uint32_t c = 0;
            for (uint32_t base = lo; base < hi; base += 32) {
                uint32_t i = base + lane;
                uint32_t kx = i < hi ? row_key(row_load(row, i)) : 0;
                bool f = kx && (kmax - kx) == pref;
                c += __popc(__ballot_sync(0xffffffffu, f));
            }
            __syncthreads();
            if (lane == 0) S->wsum[warp] = c;
            __syncthreads();
            uint32_t run = 0;
            for (uint32_t w2 = 0; w2 < warp; ++w2) run += S->wsum[w2];
            for (uint32_t base = lo; base < hi; base += 32) {
                uint32_t i = base + lane;
                float xv = i < hi ? row_load(row, i) : 0.f;
                uint32_t kx = i < hi ? row_key(xv) : 0;
                bool f = kx && (kmax - kx) == pref;
                uint32_t b = __ballot_sync(0xffffffffu, f);
                uint32_t r = run + __popc(b & ((1u << lane) - 1u));
                if (f && r < remaining) outk[n_def + r] = make_key(xv, (uint32_t)(P.col_base + (int32_t)i));
                run += __popc(b);
            }
        }
    }
    __syncthreads();
    bitonic_sort_desc(outk, n2out);
    for (uint32_t r = tid; r < P.k; r += kRowThreads) {
        const bool ok = r < kq;
        const uint64_t key = ok ? outk[r] : 0;
        P.out_scores[(size_t)q * P.k + r] = ok ? key_score(key) : 0.0f;
        P.out_docs[(size_t)q * P.k + r] = ok ? (int32_t)key_doc(key) : -1;
    }
    if (tid == 0) P.out_counts[q] = kq;
}

int topk_rows_launch_typed(int dtype, uint32_t nq, uint32_t n, const void* scores, size_t row_stride, uint32_t k,
                           int32_t col_base, float* out_scores, int32_t* out_docs, uint32_t* out_counts, cudaStream_t s,
                           const uint32_t* nq_dev, uint32_t nq_off);

int topk_rows_launch(uint32_t nq, uint32_t n, const float* scores, size_t row_stride, uint32_t k, int32_t col_base,
                     float* out_scores, int32_t* out_docs, uint32_t* out_counts, cudaStream_t s, const uint32_t* nq_dev,
                     uint32_t nq_off) {
    return topk_rows_launch_typed(0, nq, n, scores, row_stride, k, col_base, out_scores, out_docs, out_counts, s, nq_dev, nq_off);
}

// dtype: 0 = f32, 1 = f16, 2 = bf16 (the numbering of msr_quantise_topk)
int topk_rows_launch_typed(int dtype, uint32_t nq, uint32_t n, const void* scores, size_t row_stride, uint32_t k,
                           int32_t col_base, float* out_scores, int32_t* out_docs, uint32_t* out_counts, cudaStream_t s,
                           const uint32_t* nq_dev, uint32_t nq_off) {
    RowsParams P{};
    P.nq_dev = nq_dev; P.nq_off = nq_off;
    P.nq = nq; P.n = n; P.k = k; P.col_base = col_base; P.scores = scores; P.row_stride = row_stride;
    P.out_scores = out_scores; P.out_docs = out_docs; P.out_counts = out_counts;
    uint32_t n2 = 2;
    while (n2 < k) n2 <<= 1;
    // per-vector maxima of pass 1 behind the output keys when they fit 32 KB (rows up to 131 072 fp16 / 65 536 fp32 elements)
    const size_t esz = dtype == 0 ? 4 : 2;
    const size_t n_vec = (size_t)n * esz / 16 + 1;
    const size_t off = sizeof(RowsSmem) + (size_t)n2 * 8;
    P.sum_off = (uint32_t)off;
    P.sum_cap = n_vec * 2 <= kRowSumBytes ? (uint32_t)n_vec : 0u;
    const size_t smem = off + (P.sum_cap ? (size_t)P.sum_cap * 2 : 0);
    static std::atomic<uint64_t> attr_set{0};
    if (attr_needed(attr_set)) {
        const int mx = (int)(sizeof(RowsSmem) + MSR_MAX_K * 8 + kRowSumBytes);
        MSR_CUDA_CHECK(cudaFuncSetAttribute(topk_rows_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
        MSR_CUDA_CHECK(cudaFuncSetAttribute(topk_rows_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
        MSR_CUDA_CHECK(cudaFuncSetAttribute(topk_rows_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
        attr_mark(attr_set);
    }
    if (dtype == 1) topk_rows_kernel<__half><<<nq, kRowThreads, smem, s>>>(P);
    else if (dtype == 2) topk_rows_kernel<__nv_bfloat16><<<nq, kRowThreads, smem, s>>>(P);
    else topk_rows_kernel<float><<<nq, kRowThreads, smem, s>>>(P);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    return MSR_OK;
}

}  // namespace msr

using namespace msr;

extern "C" int msr_topk_merge(int device, uint32_t n_lists, uint32_t nq, uint32_t k_in, const float* scores,
                              const int32_t* docs, const uint32_t* counts, uint32_t k, float* out_scores,
                              int32_t* out_docs, uint32_t* out_counts, void* stream) {
    MSR_REQUIRE(n_lists >= 1 && k >= 1 && k_in >= 1, "msr_topk_merge: n_lists, k, k_in must be >= 1");
    MSR_REQUIRE(k <= MSR_MAX_K && k_in <= MSR_MAX_K, "msr_topk_merge: k/k_in above %u", MSR_MAX_K);
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(scores && docs && counts && out_scores && out_docs && out_counts, "msr_topk_merge: NULL buffer");
    int rc = use_device(device);
    if (rc) return rc;
    MergeParams P{};
    P.nq = nq; P.n_lists = n_lists; P.stride = k_in; P.k = k; P.mode = 1;
    P.scores = scores; P.docs = docs; P.cnts = counts;
    P.out_scores = out_scores; P.out_docs = out_docs; P.out_counts = out_counts;
    return merge_launch(P, (cudaStream_t)stream);
}

extern "C" int msr_topk_merge_strided(int device, uint32_t n_lists, uint32_t nq, uint32_t k_in, const float* scores,
                                      const int32_t* docs, const uint32_t* counts, size_t list_stride,
                                      size_t count_list_stride, uint32_t k, float* out_scores, int32_t* out_docs,
                                      uint32_t* out_counts, void* stream) {
    MSR_REQUIRE(n_lists >= 1 && k >= 1 && k_in >= 1, "msr_topk_merge_strided: n_lists, k, k_in must be >= 1");
    MSR_REQUIRE(k <= MSR_MAX_K && k_in <= MSR_MAX_K, "msr_topk_merge_strided: k/k_in above %u", MSR_MAX_K);
    MSR_REQUIRE(list_stride >= (size_t)nq * k_in && count_list_stride >= nq, "msr_topk_merge_strided: list strides too small");
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(scores && docs && counts && out_scores && out_docs && out_counts, "msr_topk_merge_strided: NULL buffer");
    int rc = use_device(device);
    if (rc) return rc;
    MergeParams P{};
    P.nq = nq; P.n_lists = n_lists; P.stride = k_in; P.k = k; P.mode = 1;
    P.scores = scores; P.docs = docs; P.cnts = counts; P.list_stride = list_stride; P.cnt_list_stride = count_list_stride;
    P.out_scores = out_scores; P.out_docs = out_docs; P.out_counts = out_counts;
    P.sorted_lists = 1;
    return merge_launch(P, (cudaStream_t)stream);
}

// rows of a finished local search -> this rank's slot on every GPU of the exchange (16-byte stores when aligned)
struct PublishParams {
    uint32_t nq, k, n_dst, vec_ok;
    const float* scores; const int32_t* docs; const uint32_t* counts;
    float* dst_scores[kMaxPeers]; int32_t* dst_docs[kMaxPeers]; uint32_t* dst_counts[kMaxPeers];
};
__global__ void __launch_bounds__(128)
exchange_publish_kernel(const PublishParams P) {
    const uint32_t q = blockIdx.x, tid = threadIdx.x;
    const size_t row = (size_t)q * P.k;
    if (P.vec_ok) {
        const uint32_t nv = P.k >> 2;
        for (uint32_t i = tid; i < 2 * nv * P.n_dst; i += 128) {
            const uint32_t v = i % nv, what = (i / nv) & 1u, dst = i / (2 * nv);
            const uint4 x = reinterpret_cast<const uint4*>(what ? (const void*)(P.docs + row) : (const void*)(P.scores + row))[v];
            reinterpret_cast<uint4*>(what ? (void*)(P.dst_docs[dst] + row) : (void*)(P.dst_scores[dst] + row))[v] = x;
        }
    } else {
        for (uint32_t r = tid; r < P.k; r += 128) {
            const float s = P.scores[row + r];
            const int32_t d = P.docs[row + r];
            for (uint32_t p = 0; p < P.n_dst; ++p) { P.dst_scores[p][row + r] = s; P.dst_docs[p][row + r] = d; }
        }
    }
    if (tid == 0) {
        const uint32_t c = P.counts[q];
        for (uint32_t p = 0; p < P.n_dst; ++p) P.dst_counts[p][q] = c;
    }
}

extern "C" int msr_exchange_publish(int device, uint32_t nq, uint32_t k, const float* scores, const int32_t* docs,
                                    const uint32_t* counts, uint32_t n_dst, float* const* dst_scores, int32_t* const* dst_docs,
                                    uint32_t* const* dst_counts, void* stream) {
    MSR_REQUIRE(k >= 1 && k <= MSR_MAX_K, "msr_exchange_publish: k=%u outside [1,%u]", k, MSR_MAX_K);
    MSR_REQUIRE(n_dst >= 1 && n_dst <= (uint32_t)kMaxPeers, "msr_exchange_publish: 1..%d destinations", kMaxPeers);
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(scores && docs && counts && dst_scores && dst_docs && dst_counts, "msr_exchange_publish: NULL buffer");
    int rc = use_device(device);
    if (rc) return rc;
    PublishParams P{};
    P.nq = nq; P.k = k; P.n_dst = n_dst; P.scores = scores; P.docs = docs; P.counts = counts;
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
    P.vec_ok = (k & 3u) == 0 && al16(scores) && al16(docs);
    for (uint32_t p = 0; p < n_dst; ++p) {
        MSR_REQUIRE(dst_scores[p] && dst_docs[p] && dst_counts[p], "msr_exchange_publish: NULL destination %u", p);
        P.dst_scores[p] = dst_scores[p]; P.dst_docs[p] = dst_docs[p]; P.dst_counts[p] = dst_counts[p];
        P.vec_ok = P.vec_ok && al16(dst_scores[p]) && al16(dst_docs[p]);
    }
    exchange_publish_kernel<<<nq, 128, 0, (cudaStream_t)stream>>>(P);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    return MSR_OK;
}

extern "C" int msr_topk_merge_exchange(int device, uint32_t n_lists, uint32_t nq, uint32_t k_in, const float* scores,
                                       const int32_t* docs, const uint32_t* counts, size_t list_stride,
                                       size_t count_list_stride, uint32_t k, float* out_scores, int32_t* out_docs,
                                       uint32_t* out_counts, uint32_t n_peers, uint32_t my_rank, uint32_t* flag_local,
                                       uint32_t* const* flag_peers, const uint32_t* peer_ranks, uint32_t flag_value,
                                       void* stream) {
    MSR_REQUIRE(n_lists >= 1 && k >= 1 && k_in >= 1, "msr_topk_merge_exchange: n_lists, k, k_in must be >= 1");
    MSR_REQUIRE(k <= MSR_MAX_K && k_in <= MSR_MAX_K, "msr_topk_merge_exchange: k/k_in above %u", MSR_MAX_K);
    MSR_REQUIRE(list_stride >= (size_t)nq * k_in && count_list_stride >= nq, "msr_topk_merge_exchange: list strides too small");
    MSR_REQUIRE(n_peers >= 1 && n_peers <= (uint32_t)kMaxPeers - 1 && flag_local && flag_peers && peer_ranks,
                "msr_topk_merge_exchange: bad peer arguments");
    MSR_REQUIRE(nq >= 1, "msr_topk_merge_exchange: nq = 0 (the flags must still be exchanged: pass at least one row)");
    MSR_REQUIRE(scores && docs && counts && out_scores && out_docs && out_counts, "msr_topk_merge_exchange: NULL buffer");
    int rc = use_device(device);
    if (rc) return rc;
    MergeParams P{};
    P.nq = nq; P.n_lists = n_lists; P.stride = k_in; P.k = k; P.mode = 1;
    P.scores = scores; P.docs = docs; P.cnts = counts; P.list_stride = list_stride; P.cnt_list_stride = count_list_stride;
    P.out_scores = out_scores; P.out_docs = out_docs; P.out_counts = out_counts;
    P.sorted_lists = 1;
    FlagParams F{};
    F.n_peers = n_peers; F.my_rank = my_rank; F.value = flag_value; F.local = flag_local;
    for (uint32_t p = 0; p < n_peers; ++p) {
        MSR_REQUIRE(flag_peers[p], "msr_topk_merge_exchange: NULL flag pointer %u", p);
        F.peers[p] = flag_peers[p]; F.peer_rank[p] = peer_ranks[p];
    }
    exchange_flags_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(F);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    return merge_launch(P, (cudaStream_t)stream);
}

extern "C" int msr_topk_rows(int device, uint32_t nq, uint32_t n, const float* scores, size_t row_stride, uint32_t k,
                             int32_t col_base, float* out_scores, int32_t* out_docs, uint32_t* out_counts, void* stream) {
    MSR_REQUIRE(k >= 1 && k <= MSR_MAX_K, "msr_topk_rows: k=%u outside [1,%u]", k, MSR_MAX_K);
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(scores && out_scores && out_docs && out_counts, "msr_topk_rows: NULL buffer");
    MSR_REQUIRE(row_stride >= n, "msr_topk_rows: row_stride < n");
    int rc = use_device(device);
    if (rc) return rc;
    return topk_rows_launch(nq, n, scores, row_stride, k, col_base, out_scores, out_docs, out_counts, (cudaStream_t)stream, nullptr, 0);
}
