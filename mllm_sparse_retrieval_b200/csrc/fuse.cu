// fuse.cu -- kernel (d): hybrid min-max normalise / interpolate / union for a batch of queries.
//
// Replaces hybrid.fuse (reference src/hybrid.py:32-53) applied to runs made by get_run_dict
// (src/search.py:66-82); call site src/search.py:455-461.  One CTA per query:
//   1. min/max of each list over the FULL returned list (get_run_dict takes them before remove_query drops
//      the query's own id -- rule H2);
//   2. the sparse list goes into a shared-memory hash table (doc -> position);
//   3. every dense entry is scored (and marks its sparse twin), then the sparse-only entries are appended in
//      sparse order by a block-wide ordered compaction -- the reference's dict insertion order (H1/H4).
// Arithmetic follows the reference under NumPy >= 2 (rule H3): the dense term is float32 arithmetic (np.float32
// scalars with weak Python floats), the sparse term float64, and their sum is float32 when a dense term exists,
// float64 otherwise.  Every operation is a separate IEEE op (no FMA contraction) so results are bit-identical
// to the reference on the same inputs.  Output scores are double (they hold either type exactly).
#include "common.cuh"

namespace msr {

constexpr int kFuseThreads = 256;
constexpr int kFuseWarps = kFuseThreads / 32;

struct FuseParams {
    uint32_t nq, k_d, k_s, table;            // table = power of two >= 2*k_s
    const void* d_scores; const int32_t* d_docs; const uint32_t* d_counts;     // scores: float, or double (msr_fuse_f64)
    const void* s_scores; const int32_t* s_docs; const uint32_t* s_counts;
    const int32_t* self_ids;
    int remove_query, dense_f32;
    double w_dense, w_sparse;
    double* out_scores; int32_t* out_docs; uint8_t* out_src; uint32_t* out_counts;
};

struct FuseSmall {
    float dmin, dmax, smin, smax;
    uint32_t wsum[kFuseWarps];
};

__device__ __forceinline__ uint32_t hash_doc(uint32_t d) { return d * 2654435761u; }

template <typename T>
__device__ __forceinline__ T block_minmax(T v, bool is_max, T* scratch) {
    // scratch: kFuseWarps values
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        T t = __shfl_xor_sync(0xffffffffu, v, o);
        v = is_max ? (t > v ? t : v) : (t < v ? t : v);
    }
    __syncthreads();
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    T r = scratch[0];
#pragma unroll
    for (int w = 1; w < kFuseWarps; ++w) r = is_max ? (scratch[w] > r ? scratch[w] : r) : (scratch[w] < r ? scratch[w] : r);
    return r;
}

// T = float: scores as the searchers return them (dense np.float32 or, with dense_f32 = 0, Python floats that are float32
// values).  T = double (msr_fuse_f64, dense_f32 = 0 only): runs of arbitrary Python floats, e.g. re-read from a TREC file --
// float('0.1') is not float32(0.1), and the reference combines the doubles it parsed.
template <typename T>
__global__ void __launch_bounds__(kFuseThreads)
fuse_kernel(const FuseParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    FuseSmall* S = reinterpret_cast<FuseSmall*>(smem_raw);
    T* scratch = reinterpret_cast<T*>(smem_raw + sizeof(FuseSmall));                      // kFuseWarps (room for doubles)
    uint32_t* tkeys = reinterpret_cast<uint32_t*>(smem_raw + sizeof(FuseSmall) + kFuseWarps * 8);   // [table]
    uint32_t* tvals = tkeys + P.table;                                                    // [table]
    uint8_t* matched = reinterpret_cast<uint8_t*>(tvals + P.table);                       // [k_s]

    const uint32_t q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t nd = P.d_counts ? min(P.d_counts[q], P.k_d) : 0;
    const uint32_t ns = P.s_counts ? min(P.s_counts[q], P.k_s) : 0;
    const T* ds = reinterpret_cast<const T*>(P.d_scores) + (size_t)q * P.k_d;
    const int32_t* dd = P.d_docs + (size_t)q * P.k_d;
    const T* ss = reinterpret_cast<const T*>(P.s_scores) + (size_t)q * P.k_s;
    const int32_t* sd = P.s_docs + (size_t)q * P.k_s;
    const int32_t self = (P.remove_query && P.self_ids) ? P.self_ids[q] : -1;
    const uint32_t kout = P.k_d + P.k_s;
    double* os = P.out_scores + (size_t)q * kout;
    int32_t* od = P.out_docs + (size_t)q * kout;
    uint8_t* osrc = P.out_src + (size_t)q * kout;

    // 1. min / max over the full lists (0/0 for an empty list)
    T mn = (T)INFINITY, mx = (T)-INFINITY;
    for (uint32_t i = tid; i < nd; i += kFuseThreads) { T v = ds[i]; mn = v < mn ? v : mn; mx = v > mx ? v : mx; }
    T dmin = block_minmax<T>(mn, false, scratch), dmax = block_minmax<T>(mx, true, scratch);
    mn = (T)INFINITY; mx = (T)-INFINITY;
    for (uint32_t i = tid; i < ns; i += kFuseThreads) { T v = ss[i]; mn = v < mn ? v : mn; mx = v > mx ? v : mx; }
    T smin = block_minmax<T>(mn, false, scratch), smax = block_minmax<T>(mx, true, scratch);
    if (nd == 0) { dmin = 0; dmax = 0; }
    if (ns == 0) { smin = 0; smax = 0; }

    // 2. sparse list -> hash table
    for (uint32_t i = tid; i < P.table; i += kFuseThreads) tkeys[i] = 0xFFFFFFFFu;
    for (uint32_t i = tid; i < ns; i += kFuseThreads) matched[i] = 0;
    __syncthreads();
    const uint32_t mask = P.table - 1;
    for (uint32_t i = tid; i < ns; i += kFuseThreads) {
        int32_t doc = sd[i];
        if (doc == self || doc < 0) continue;                 // dropped by remove_query / padding
        uint32_t slot = hash_doc((uint32_t)doc) & mask;
        for (;;) {
            uint32_t prev = atomicCAS(&tkeys[slot], 0xFFFFFFFFu, (uint32_t)doc);
            if (prev == 0xFFFFFFFFu) { tvals[slot] = i; break; }
            if (prev == (uint32_t)doc) { atomicMax(&tvals[slot], i); break; }   // dict: last duplicate wins
            slot = (slot + 1) & mask;
        }
    }
    __syncthreads();

    // normalisation constants, typed as the reference computes them
    const float d_den32 = __fsub_rn((float)dmax, (float)dmin);
    const bool d_den_small = 1e-9 > (double)d_den32;          // Python max(np.float32, 1e-9)
    const float d_den = d_den_small ? (float)1e-9 : d_den32;
    const float wd32 = (float)P.w_dense;                      // weak Python float * np.float32 -> float32
    const double s_den = fmax(__dsub_rn((double)smax, (double)smin), 1e-9);
    const double d_den64 = fmax(__dsub_rn((double)dmax, (double)dmin), 1e-9);   // both runs Python floats

    // 3a. dense entries, in dense order (ordered compaction: entries equal to the self id are dropped)
    uint32_t run = 0;
    for (uint32_t base = 0; base < nd; base += kFuseThreads) {
        const uint32_t i = base + tid;
        int32_t doc = -1;
        bool f = false;
        if (i < nd) { doc = dd[i]; f = doc >= 0 && doc != self; }
        const uint32_t b = __ballot_sync(0xffffffffu, f);
        __syncthreads();
        if (lane == 0) S->wsum[warp] = __popc(b);
        __syncthreads();
        uint32_t before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kFuseWarps; ++w) { uint32_t x = S->wsum[w]; if (w < (int)warp) before += x; total += x; }
        if (f) {
            float sc = __fmul_rn(wd32, __fdiv_rn(__fsub_rn((float)ds[i], (float)dmin), d_den));
            double sc64 = __dmul_rn(P.w_dense, __ddiv_rn(__dsub_rn((double)ds[i], (double)dmin), d_den64));
            uint8_t src = 1;
            uint32_t slot = hash_doc((uint32_t)doc) & mask;
            for (;;) {
                uint32_t kx = tkeys[slot];
                if (kx == 0xFFFFFFFFu) break;
                if (kx == (uint32_t)doc) {
                    uint32_t j = tvals[slot];
                    matched[j] = 1;
                    double st = __dmul_rn(P.w_sparse, __ddiv_rn(__dsub_rn((double)ss[j], (double)smin), s_den));
                    sc = __fadd_rn(sc, (float)st);            // np.float32 += weak Python float
                    sc64 = __dadd_rn(sc64, st);
                    src = 3;
                    break;
                }
                slot = (slot + 1) & mask;
            }
            const uint32_t pos = run + before + __popc(b & ((1u << lane) - 1u));
            os[pos] = P.dense_f32 ? (double)sc : sc64; od[pos] = doc; osrc[pos] = src;
        }
        run += total;
    }
    __syncthreads();

    // 3b. sparse-only entries, in sparse order (ordered compaction).  Doc ids are unique within a list.
    for (uint32_t base = 0; base < ns; base += kFuseThreads) {
        uint32_t i = base + tid;
        bool f = false;
        int32_t doc = -1;
        if (i < ns) {
            doc = sd[i];
            f = doc >= 0 && doc != self && !matched[i];
        }
        uint32_t b = __ballot_sync(0xffffffffu, f);
        uint32_t wcount = __popc(b);
        __syncthreads();
        if (lane == 0) S->wsum[warp] = wcount;
        __syncthreads();
        uint32_t before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kFuseWarps; ++w) { uint32_t x = S->wsum[w]; if (w < (int)warp) before += x; total += x; }
        if (f) {
            double st = __dmul_rn(P.w_sparse, __ddiv_rn(__dsub_rn((double)ss[i], (double)smin), s_den));
            uint32_t pos = run + before + __popc(b & ((1u << lane) - 1u));
            os[pos] = st; od[pos] = doc; osrc[pos] = 2;       // int 0 + Python float -> float64
        }
        run += total;
    }
    for (uint32_t i = run + tid; i < kout; i += kFuseThreads) { os[i] = 0.0; od[i] = -1; osrc[i] = 0; }
    if (tid == 0) P.out_counts[q] = run;
}

}  // namespace msr

using namespace msr;

static int fuse_launch(const FuseParams& P, bool f64, void* stream) {
    const size_t smem = sizeof(FuseSmall) + kFuseWarps * 8 + (size_t)P.table * 8 + ((P.k_s + 15) & ~15u);
    static std::atomic<uint64_t> attr_set{0};
    if (attr_needed(attr_set)) {
        MSR_CUDA_CHECK(cudaFuncSetAttribute(fuse_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        MSR_CUDA_CHECK(cudaFuncSetAttribute(fuse_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        attr_mark(attr_set);
    }
    if (f64) fuse_kernel<double><<<P.nq, kFuseThreads, smem, (cudaStream_t)stream>>>(P);
    else fuse_kernel<float><<<P.nq, kFuseThreads, smem, (cudaStream_t)stream>>>(P);
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    return MSR_OK;
}

extern "C" int msr_fuse(int device, uint32_t nq, uint32_t k_d, const float* d_scores, const int32_t* d_docs,
                        const uint32_t* d_counts, uint32_t k_s, const float* s_scores, const int32_t* s_docs,
                        const uint32_t* s_counts, const int32_t* self_ids, int remove_query, int dense_f32,
                        double w_dense, double w_sparse, double* out_scores, int32_t* out_docs, uint8_t* out_src,
                        uint32_t* out_counts, void* stream) {
    MSR_REQUIRE(k_d >= 1 && k_s >= 1 && k_d <= MSR_MAX_K && k_s <= MSR_MAX_K, "msr_fuse: k_d/k_s outside [1,%u]", MSR_MAX_K);
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(d_scores && d_docs && d_counts && s_scores && s_docs && s_counts, "msr_fuse: NULL input");
    MSR_REQUIRE(out_scores && out_docs && out_src && out_counts, "msr_fuse: NULL output");
    int rc = use_device(device);
    if (rc) return rc;
    FuseParams P{};
    P.nq = nq; P.k_d = k_d; P.k_s = k_s;
    P.table = 2;
    while (P.table < 2 * k_s) P.table <<= 1;
    P.d_scores = d_scores; P.d_docs = d_docs; P.d_counts = d_counts;
    P.s_scores = s_scores; P.s_docs = s_docs; P.s_counts = s_counts;
    P.self_ids = self_ids; P.remove_query = remove_query; P.dense_f32 = dense_f32; P.w_dense = w_dense; P.w_sparse = w_sparse;
    P.out_scores = out_scores; P.out_docs = out_docs; P.out_src = out_src; P.out_counts = out_counts;
    return fuse_launch(P, false, stream);
}

extern "C" int msr_fuse_f64(int device, uint32_t nq, uint32_t k_d, const double* d_scores, const int32_t* d_docs,
                            const uint32_t* d_counts, uint32_t k_s, const double* s_scores, const int32_t* s_docs,
                            const uint32_t* s_counts, const int32_t* self_ids, int remove_query, double w_dense,
                            double w_sparse, double* out_scores, int32_t* out_docs, uint8_t* out_src, uint32_t* out_counts,
                            void* stream) {
    MSR_REQUIRE(k_d >= 1 && k_s >= 1 && k_d <= MSR_MAX_K && k_s <= MSR_MAX_K, "msr_fuse_f64: k_d/k_s outside [1,%u]", MSR_MAX_K);
    if (nq == 0) return MSR_OK;
    MSR_REQUIRE(d_scores && d_docs && d_counts && s_scores && s_docs && s_counts, "msr_fuse_f64: NULL input");
    MSR_REQUIRE(out_scores && out_docs && out_src && out_counts, "msr_fuse_f64: NULL output");
    int rc = use_device(device);
    if (rc) return rc;
    FuseParams P{};
    P.nq = nq; P.k_d = k_d; P.k_s = k_s;
    P.table = 2;
    while (P.table < 2 * k_s) P.table <<= 1;
    P.d_scores = d_scores; P.d_docs = d_docs; P.d_counts = d_counts;
    P.s_scores = s_scores; P.s_docs = s_docs; P.s_counts = s_counts;
    P.self_ids = self_ids; P.remove_query = remove_query; P.dense_f32 = 0; P.w_dense = w_dense; P.w_sparse = w_sparse;
    P.out_scores = out_scores; P.out_docs = out_docs; P.out_src = out_src; P.out_counts = out_counts;
    return fuse_launch(P, true, stream);
}
