// invert.cu -- index builder, device half: doc-major vectors -> canonical term-major CSR.
//
// Replaces the inversion that Lucene performs inside `python -m pyserini.index.lucene --collection JsonVectorCollection
// --impact --pretokenized` (reference scripts/sparse_index.sh:12-18) for the vectors src/encode.py:353-359 writes.
// Input is what a JSONL reader produces: per document (in file order) a list of (term id, impact), plus the internal
// id of each document (its rank among the external docid strings in UTF-8 byte order, rule S8).  Output is the
// canonical CSR msr_sparse_index_build consumes: postings grouped by term, internal doc ids strictly ascending.
//
//   1. invert_keys_kernel   one warp per document: key = (term << 32) | internal doc, value = impact; term histogram
//   2. cub::DeviceRadixSort over the key bits that are actually used (log2(n_terms) + log2(n_docs) bits)
//   3. exclusive scan of the histogram -> indptr;  invert_unpack_kernel splits the sorted keys
// A duplicate (term, doc) pair -- impossible for a JSON object, possible for hand-made input -- is reported.
#include <cub/cub.cuh>

#include "common.cuh"

namespace msr {

__global__ void __launch_bounds__(256)
invert_keys_kernel(const uint64_t* __restrict__ doc_indptr, const uint32_t* __restrict__ terms,
                   const uint16_t* __restrict__ impacts, const uint32_t* __restrict__ doc_rank, uint32_t n_docs,
                   uint32_t n_terms, uint64_t* __restrict__ keys, uint16_t* __restrict__ vals,
                   unsigned long long* __restrict__ hist, uint32_t* __restrict__ bad) {
    const uint32_t lane = threadIdx.x & 31;
    for (uint32_t d = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; d < n_docs; d += (gridDim.x * blockDim.x) >> 5) {
        const uint64_t lo = doc_indptr[d], hi = doc_indptr[d + 1];
        const uint32_t id = doc_rank ? doc_rank[d] : d;
        if (id >= n_docs) { if (lane == 0) atomicOr(bad, 4u); continue; }
        for (uint64_t p = lo + lane; p < hi; p += 32) {
            const uint32_t t = terms[p];
            const uint16_t w = impacts[p];
            if (t >= n_terms || w == 0) { atomicOr(bad, 1u); keys[p] = ~0ull; vals[p] = 0; continue; }
            keys[p] = ((uint64_t)t << 32) | id;
            vals[p] = w;
            atomicAdd(hist + t, 1ull);
        }
    }
}

__global__ void invert_unpack_kernel(const uint64_t* __restrict__ keys, uint64_t nnz, uint32_t* __restrict__ docids,
                                     uint32_t* __restrict__ bad) {
    for (uint64_t p = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; p < nnz; p += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t k = keys[p];
        docids[p] = (uint32_t)k;
        if (p && keys[p - 1] == k) atomicOr(bad, 2u);
    }
}

static int bits_for(uint64_t n) {           // bits needed to represent values 0..n-1
    int b = 1;
    while (b < 64 && (1ull << b) < n) ++b;
    return b;
}

}  // namespace msr

using namespace msr;

extern "C" int msr_invert_docs(int device, uint32_t n_docs, uint32_t n_terms, const uint64_t* doc_indptr,
                               const uint32_t* terms, const uint16_t* impacts, const uint32_t* doc_rank,
                               uint64_t* out_indptr, uint32_t* out_docids, uint16_t* out_impacts, int where, void* stream) {
    MSR_REQUIRE(n_docs > 0 && n_terms > 0, "msr_invert_docs: empty collection (n_docs=%u n_terms=%u)", n_docs, n_terms);
    MSR_REQUIRE(doc_indptr && out_indptr, "msr_invert_docs: NULL indptr");
    MSR_REQUIRE(where == MSR_HOST || where == MSR_DEVICE, "msr_invert_docs: bad `where`");
    int rc = use_device(device);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    uint64_t nnz = 0;
    if (where == MSR_HOST) nnz = doc_indptr[n_docs];
    else {
        MSR_CUDA_CHECK(cudaMemcpyAsync(&nnz, doc_indptr + n_docs, 8, cudaMemcpyDeviceToHost, s));
        MSR_CUDA_CHECK(cudaStreamSynchronize(s));
    }
    MSR_REQUIRE(nnz < (1ull << 32), "msr_invert_docs: nnz %llu >= 2^32 (shard the collection)", (unsigned long long)nnz);
    MSR_REQUIRE(nnz == 0 || (terms && impacts && out_docids && out_impacts), "msr_invert_docs: NULL posting arrays");

    DevBuf b_ip, b_t, b_w, b_r, b_oip, b_od, b_ow, b_k0, b_k1, b_v0, b_hist, b_bad, b_tmp;
    const uint64_t* d_ip = doc_indptr; const uint32_t* d_t = terms; const uint16_t* d_w = impacts; const uint32_t* d_r = doc_rank;
    uint64_t* d_oip = out_indptr; uint32_t* d_od = out_docids; uint16_t* d_ow = out_impacts;
    const size_t n1 = nnz ? nnz : 1;
    if (where == MSR_HOST) {
        MSR_CUDA_CHECK(b_ip.alloc((size_t)(n_docs + 1) * 8)); MSR_CUDA_CHECK(b_t.alloc(n1 * 4)); MSR_CUDA_CHECK(b_w.alloc(n1 * 2));
        MSR_CUDA_CHECK(b_oip.alloc((size_t)(n_terms + 1) * 8)); MSR_CUDA_CHECK(b_od.alloc(n1 * 4)); MSR_CUDA_CHECK(b_ow.alloc(n1 * 2));
        MSR_CUDA_CHECK(cudaMemcpyAsync(b_ip.p, doc_indptr, (size_t)(n_docs + 1) * 8, cudaMemcpyHostToDevice, s));
        if (nnz) {
            MSR_CUDA_CHECK(cudaMemcpyAsync(b_t.p, terms, nnz * 4, cudaMemcpyHostToDevice, s));
            MSR_CUDA_CHECK(cudaMemcpyAsync(b_w.p, impacts, nnz * 2, cudaMemcpyHostToDevice, s));
        }
        if (doc_rank) {
            MSR_CUDA_CHECK(b_r.alloc((size_t)n_docs * 4));
            MSR_CUDA_CHECK(cudaMemcpyAsync(b_r.p, doc_rank, (size_t)n_docs * 4, cudaMemcpyHostToDevice, s));
            d_r = b_r.as<uint32_t>();
        }
        d_ip = b_ip.as<uint64_t>(); d_t = b_t.as<uint32_t>(); d_w = b_w.as<uint16_t>();
        d_oip = b_oip.as<uint64_t>(); d_od = b_od.as<uint32_t>(); d_ow = b_ow.as<uint16_t>();
    }
    MSR_CUDA_CHECK(b_k0.alloc(n1 * 8)); MSR_CUDA_CHECK(b_k1.alloc(n1 * 8)); MSR_CUDA_CHECK(b_v0.alloc(n1 * 2));
    MSR_CUDA_CHECK(b_hist.alloc((size_t)(n_terms + 1) * 8)); MSR_CUDA_CHECK(b_bad.alloc(4));
    MSR_CUDA_CHECK(cudaMemsetAsync(b_hist.p, 0, (size_t)(n_terms + 1) * 8, s));
    MSR_CUDA_CHECK(cudaMemsetAsync(b_bad.p, 0, 4, s));
    const int end_bit = std::min(64, 32 + bits_for(n_terms));
    size_t t_sort = 0, t_scan = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, t_sort, b_k0.as<uint64_t>(), b_k1.as<uint64_t>(), b_v0.as<uint16_t>(), d_ow,
                                    (uint64_t)nnz, 0, end_bit, s);
    cub::DeviceScan::ExclusiveSum(nullptr, t_scan, b_hist.as<uint64_t>(), d_oip, n_terms + 1, s);
    MSR_CUDA_CHECK(b_tmp.alloc(std::max(t_sort, t_scan)));
    size_t tmp_bytes = std::max(t_sort, t_scan);

    const uint32_t warps_per_block = 8;
    const uint32_t grid = std::min<uint32_t>((n_docs + warps_per_block - 1) / warps_per_block, 148 * 16);
    invert_keys_kernel<<<grid, 256, 0, s>>>(d_ip, d_t, d_w, d_r, n_docs, n_terms, b_k0.as<uint64_t>(), b_v0.as<uint16_t>(),
                                            (unsigned long long*)b_hist.p, b_bad.as<uint32_t>());
    count_launch();
    MSR_CUDA_CHECK(cudaGetLastError());
    if (nnz) {
        MSR_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(b_tmp.p, tmp_bytes, b_k0.as<uint64_t>(), b_k1.as<uint64_t>(),
                                                       b_v0.as<uint16_t>(), d_ow, (uint64_t)nnz, 0, end_bit, s));
        invert_unpack_kernel<<<(unsigned)std::min<uint64_t>((nnz + 255) / 256, 148ull * 32), 256, 0, s>>>(
            b_k1.as<uint64_t>(), nnz, d_od, b_bad.as<uint32_t>());
        count_launch();
    }
    MSR_CUDA_CHECK(cub::DeviceScan::ExclusiveSum(b_tmp.p, tmp_bytes, b_hist.as<uint64_t>(), d_oip, n_terms + 1, s));
    uint32_t bad = 0;
    MSR_CUDA_CHECK(cudaMemcpyAsync(&bad, b_bad.p, 4, cudaMemcpyDeviceToHost, s));
    if (where == MSR_HOST) {
        MSR_CUDA_CHECK(cudaMemcpyAsync(out_indptr, d_oip, (size_t)(n_terms + 1) * 8, cudaMemcpyDeviceToHost, s));
        if (nnz) {
            MSR_CUDA_CHECK(cudaMemcpyAsync(out_docids, d_od, nnz * 4, cudaMemcpyDeviceToHost, s));
            MSR_CUDA_CHECK(cudaMemcpyAsync(out_impacts, d_ow, nnz * 2, cudaMemcpyDeviceToHost, s));
        }
    }
    MSR_CUDA_CHECK(cudaStreamSynchronize(s));
    if (bad) {
        set_error("msr_invert_docs: malformed input (%s%s%s)", (bad & 1) ? "term id out of range or impact 0; " : "",
                  (bad & 2) ? "a term occurs twice in one document; " : "", (bad & 4) ? "doc_rank out of range" : "");
        return MSR_ERR_INVALID;
    }
    return MSR_OK;
}
