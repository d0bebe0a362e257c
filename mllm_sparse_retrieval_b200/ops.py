"""Thin tensor-level wrappers over the C ABI (include/msr.h).  torch is used for device memory and streams only.

Every function here ends in a libmsr.so kernel; nothing falls back to torch math or to the CPU.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import MSR_DEVICE, MSR_HOST, MSR_MAX_K, check, current_stream_ptr, ptr  # noqa: F401


def _dev_index(device) -> int:
    d = torch.device(device)
    if d.type != "cuda":
        raise RuntimeError(f"libmsr needs a CUDA (sm_100a) device, got {d}; there is no CPU fallback")
    return d.index if d.index is not None else torch.cuda.current_device()


def kernel_launches() -> int:
    return int(_lib.load().msr_kernel_launches())


class SparseIndex:
    """Device-resident tiled impact index of one shard (msr_sparse_index_*)."""

    def __init__(self, indptr, docids, impacts, n_docs, *, device="cuda", n_docs_global=None, df_global=None,
                 doc_base=0, tile_docs=0):
        """indptr u64/i64 [V+1], docids u32/i32 [nnz], impacts u16/i16 [nnz]: numpy (host) or CUDA tensors."""
        lib = _lib.load()
        self.device = _dev_index(device)
        on_device = isinstance(indptr, torch.Tensor) and indptr.is_cuda
        if on_device:
            assert indptr.dtype == torch.int64 and docids.dtype == torch.int32 and impacts.dtype == torch.int16
            indptr, docids, impacts = indptr.contiguous(), docids.contiguous(), impacts.contiguous()
            if df_global is not None:
                df_global = df_global.to(device=indptr.device, dtype=torch.int32).contiguous()
        else:
            indptr = np.ascontiguousarray(np.asarray(indptr), dtype=np.uint64)
            docids = np.ascontiguousarray(np.asarray(docids), dtype=np.uint32)
            impacts = np.ascontiguousarray(np.asarray(impacts), dtype=np.uint16)
            if df_global is not None:
                df_global = np.ascontiguousarray(np.asarray(df_global), dtype=np.uint32)
        n_terms = int(indptr.shape[0]) - 1
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            check(lib.msr_sparse_index_build(self.device, int(n_docs), n_terms, ptr(indptr), ptr(docids), ptr(impacts),
                                             int(n_docs if n_docs_global is None else n_docs_global), ptr(df_global),
                                             int(doc_base), int(tile_docs), MSR_DEVICE if on_device else MSR_HOST,
                                             current_stream_ptr(self.device), ctypes.byref(self._h)))
        info = (ctypes.c_uint32 * 9)()
        check(lib.msr_sparse_index_info(self._h, info))
        self.n_docs, self.n_terms, self.n_tiles, self.tile_docs = info[0], info[1], info[2], info[3]
        self.nnz = info[4] | (info[5] << 32)
        self.resident_bytes = info[7] | (info[8] << 32)
        self.doc_base = int(doc_base)

    def close(self):
        if getattr(self, "_h", None):
            _lib.load().msr_sparse_index_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:                                       # interpreter shutdown: module globals may be gone already
            pass

    def search(self, q_indptr, q_terms, q_weights, k, out=None):
        """Device path: CUDA tensors (int64 indptr, int32 terms/weights) -> (scores f32, docs i32, counts i32) [nq,k].
        Asynchronous on the current stream."""
        nq = q_indptr.numel() - 1
        dev = torch.device("cuda", self.device)
        assert q_indptr.is_cuda and q_indptr.dtype == torch.int64
        assert q_terms.dtype == torch.int32 and q_weights.dtype == torch.int32
        if out is None:
            out = (torch.empty((nq, k), dtype=torch.float32, device=dev),
                   torch.empty((nq, k), dtype=torch.int32, device=dev),
                   torch.empty((nq,), dtype=torch.int32, device=dev))
        with torch.cuda.device(self.device):
            check(_lib.load().msr_sparse_search(self._h, nq, ptr(q_indptr), ptr(q_terms), ptr(q_weights), int(k),
                                                ptr(out[0]), ptr(out[1]), ptr(out[2]), MSR_DEVICE,
                                                current_stream_ptr(self.device)))
        return out

    def set_peer_outputs(self, peer_scores=(), peer_docs=(), peer_counts=()):
        """Fused exchange (msr_sparse_set_peer_outputs): device-path searches also store their merged rows into these
        raw device addresses (ints: [nq,k] f32 / [nq,k] i32 / [nq] i32 slots on other GPUs).  Empty = off."""
        n = len(peer_scores)
        assert n == len(peer_docs) == len(peer_counts)
        arr = lambda xs: (ctypes.c_void_p * max(1, n))(*[ctypes.c_void_p(int(x)) for x in xs])
        check(_lib.load().msr_sparse_set_peer_outputs(self._h, n, arr(peer_scores), arr(peer_docs), arr(peer_counts)))

    def set_peer_theta(self, local_theta=0, peer_theta=(), capacity=0, rank=0, world=1, local_quant=0, peer_quant=()):
        """Cross-GPU threshold sharing (msr_sparse_set_peer_theta): raw device addresses of this GPU's u64[capacity]
        threshold array in symmetric memory and of the peers' copies, optionally with the u64[world][capacity] quantile
        arrays; ``local_theta=0`` switches it off."""
        n = len(peer_theta)
        arr = (ctypes.c_void_p * max(1, n))(*[ctypes.c_void_p(int(x)) for x in peer_theta])
        qarr = (ctypes.c_void_p * max(1, n))(*[ctypes.c_void_p(int(x)) for x in peer_quant]) if local_quant else None
        check(_lib.load().msr_sparse_set_peer_theta(self._h, n, ctypes.c_void_p(int(local_theta)) if local_theta else None, arr,
                                                    int(capacity), int(rank), int(max(1, world)),
                                                    ctypes.c_void_p(int(local_quant)) if local_quant else None, qarr))

    def status(self):
        """Synchronise and raise MsrError(MSR_ERR_RANGE) if an earlier device-path search held more than 8 queries outside the
        integer kernel's exact range (one wide pass per call; the host path takes any number)."""
        with torch.cuda.device(self.device):
            check(_lib.load().msr_sparse_index_status(self._h, current_stream_ptr(self.device)))

    def search_host(self, q_indptr, q_terms, q_weights, k, out=None):
        """Host path: numpy arrays or pinned CPU tensors in, host arrays out; H2D/D2H inside the call."""
        if isinstance(q_indptr, torch.Tensor):
            nq = q_indptr.numel() - 1
            assert not q_indptr.is_cuda and q_indptr.dtype == torch.int64
            assert q_terms.dtype == torch.int32 and q_weights.dtype == torch.int32
        else:
            q_indptr = np.ascontiguousarray(q_indptr, dtype=np.uint64)
            q_terms = np.ascontiguousarray(q_terms, dtype=np.uint32)
            q_weights = np.ascontiguousarray(q_weights, dtype=np.uint32)
            nq = len(q_indptr) - 1
        if out is None:
            out = (np.empty((nq, k), dtype=np.float32), np.empty((nq, k), dtype=np.int32),
                   np.empty((nq,), dtype=np.uint32))
        with torch.cuda.device(self.device):
            check(_lib.load().msr_sparse_search(self._h, nq, ptr(q_indptr), ptr(q_terms), ptr(q_weights), int(k),
                                                ptr(out[0]), ptr(out[1]), ptr(out[2]), MSR_HOST,
                                                current_stream_ptr(self.device)))
        return out

    def profile(self, enable=True):
        """(score_ms, merge_ms, launches) accumulated since the last call; then switch event timing on/off."""
        a, b, n = ctypes.c_double(), ctypes.c_double(), ctypes.c_uint64()
        with torch.cuda.device(self.device):
            check(_lib.load().msr_sparse_profile(self._h, int(bool(enable)), ctypes.byref(a), ctypes.byref(b),
                                                 ctypes.byref(n), current_stream_ptr(self.device)))
        return a.value, b.value, int(n.value)

    def query_work(self, q_indptr, q_terms, q_weights):
        """(postings touched, kept (query,term) pairs) of a device-resident query batch -- roofline numerator."""
        nq = q_indptr.numel() - 1
        p, c = ctypes.c_uint64(), ctypes.c_uint64()
        with torch.cuda.device(self.device):
            check(_lib.load().msr_sparse_query_work(self._h, nq, ptr(q_indptr), ptr(q_terms), ptr(q_weights),
                                                    ctypes.byref(p), ctypes.byref(c), MSR_DEVICE,
                                                    current_stream_ptr(self.device)))
        return int(p.value), int(c.value)

    def query_traffic(self, q_indptr, q_terms, q_weights):
        """Physical index bytes the batch touches, every distinct kept term counted once (msr_sparse_query_traffic):
        the DRAM floor of a cold-L2 search, used by bench.py's small-batch lines."""
        nq = q_indptr.numel() - 1
        b = ctypes.c_uint64()
        with torch.cuda.device(self.device):
            check(_lib.load().msr_sparse_query_traffic(self._h, nq, ptr(q_indptr), ptr(q_terms), ptr(q_weights),
                                                       ctypes.byref(b), current_stream_ptr(self.device)))
        return int(b.value)


MB_L2_READ, MB_HBM_READ, MB_RED_SHARED, MB_ISSUE, MB_ALU, MB_LDS, MB_IMMA, MB_IMMA_PRMT = range(8)


def microbench(what, nbytes=0, iters=5, device=0):
    """Measured ceiling of one resource the sparse scorer uses (msr_microbench): bytes/s or warp instructions/s."""
    v = ctypes.c_double()
    dev = torch.device(device)
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    with torch.cuda.device(idx):
        check(_lib.load().msr_microbench(idx, int(what), int(nbytes), int(iters), ctypes.byref(v), current_stream_ptr(dev)))
    return float(v.value)


def topk_merge(scores, docs, counts, k, out=None):
    """[n_lists,nq,k_in] lists (e.g. an all-gather of per-GPU results) -> merged [nq,k] (into ``out`` when given)."""
    n_lists, nq, k_in = scores.shape
    dev = scores.device
    if out is None:
        out = (torch.empty((nq, k), dtype=torch.float32, device=dev), torch.empty((nq, k), dtype=torch.int32, device=dev),
               torch.empty((nq,), dtype=torch.int32, device=dev))
    assert all(t.is_contiguous() for t in out)
    di = _dev_index(dev)
    with torch.cuda.device(di):
        check(_lib.load().msr_topk_merge(di, n_lists, nq, k_in, ptr(scores.contiguous()), ptr(docs.contiguous()),
                                         ptr(counts.contiguous()), int(k), ptr(out[0]), ptr(out[1]), ptr(out[2]),
                                         current_stream_ptr(di)))
    return out


def topk_merge_strided(scores_ptr, docs_ptr, counts_ptr, n_lists, nq, k_in, list_stride, count_list_stride, k, out, device):
    """msr_topk_merge_strided over raw device addresses (a symmetric-memory result buffer); ``out`` = (f32, i32, i32)."""
    di = _dev_index(device)
    with torch.cuda.device(di):
        check(_lib.load().msr_topk_merge_strided(di, n_lists, nq, k_in, ctypes.c_void_p(scores_ptr), ctypes.c_void_p(docs_ptr),
                                                 ctypes.c_void_p(counts_ptr), list_stride, count_list_stride, int(k),
                                                 ptr(out[0]), ptr(out[1]), ptr(out[2]), current_stream_ptr(di)))
    return out


def topk_merge_exchange(scores_ptr, docs_ptr, counts_ptr, n_lists, nq, k_in, list_stride, count_list_stride, k, out, device,
                        rank, flag_local, flag_peers, peer_ranks, flag_value):
    """msr_topk_merge_exchange: ``topk_merge_strided`` with the exchange barrier fused in (signal the peers, wait for
    them, merge -- one kernel).  flag_local / flag_peers are raw device addresses of u32 flag arrays in symmetric memory."""
    di = _dev_index(device)
    n = len(flag_peers)
    fp = (ctypes.c_void_p * max(1, n))(*[ctypes.c_void_p(int(x)) for x in flag_peers])
    pr = (ctypes.c_uint32 * max(1, n))(*[int(x) for x in peer_ranks])
    with torch.cuda.device(di):
        check(_lib.load().msr_topk_merge_exchange(di, n_lists, nq, k_in, ctypes.c_void_p(scores_ptr), ctypes.c_void_p(docs_ptr),
                                                  ctypes.c_void_p(counts_ptr), list_stride, count_list_stride, int(k),
                                                  ptr(out[0]), ptr(out[1]), ptr(out[2]), n, int(rank),
                                                  ctypes.c_void_p(int(flag_local)), fp, pr, int(flag_value) & 0xFFFFFFFF,
                                                  current_stream_ptr(di)))
    return out


def exchange_publish(scores, docs, counts, dst_scores, dst_docs, dst_counts):
    """msr_exchange_publish: rows of a finished local search (CUDA tensors f32 / i32 / i32) -> raw device addresses of this
    rank's slot on every GPU of the exchange."""
    nq, k = scores.shape
    di = _dev_index(scores.device)
    n = len(dst_scores)
    arr = lambda xs: (ctypes.c_void_p * max(1, n))(*[ctypes.c_void_p(int(x)) for x in xs])
    with torch.cuda.device(di):
        check(_lib.load().msr_exchange_publish(di, nq, k, ptr(scores.contiguous()), ptr(docs.contiguous()), ptr(counts.contiguous()), n,
                                               arr(dst_scores), arr(dst_docs), arr(dst_counts), current_stream_ptr(di)))


def topk_rows(scores, k, col_base=0):
    """Radix-select top-k of each row of a CUDA fp32 matrix by (score desc, column asc)."""
    assert scores.is_cuda and scores.dtype == torch.float32 and scores.stride(1) == 1
    nq, n = scores.shape
    dev = scores.device
    out = (torch.empty((nq, k), dtype=torch.float32, device=dev), torch.empty((nq, k), dtype=torch.int32, device=dev),
           torch.empty((nq,), dtype=torch.int32, device=dev))
    di = _dev_index(dev)
    with torch.cuda.device(di):
        check(_lib.load().msr_topk_rows(di, nq, n, ptr(scores), scores.stride(0), int(k), int(col_base), ptr(out[0]),
                                        ptr(out[1]), ptr(out[2]), current_stream_ptr(di)))
    return out


class DenseIndex:
    """Device-resident flat inner-product index (msr_dense_index_*)."""

    def __init__(self, dim, device="cuda", doc_base=0):
        self.device = _dev_index(device)
        self.dim = int(dim)
        self._h = ctypes.c_void_p()
        check(_lib.load().msr_dense_index_create(self.device, self.dim, int(doc_base), ctypes.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None):
            _lib.load().msr_dense_index_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:                                       # interpreter shutdown: module globals may be gone already
            pass

    @property
    def ntotal(self):
        info = (ctypes.c_uint32 * 5)()
        check(_lib.load().msr_dense_index_info(self._h, info))
        return int(info[0])

    def add(self, vecs):
        if isinstance(vecs, torch.Tensor) and vecs.is_cuda:
            vecs = vecs.to(torch.float32).contiguous()
            where = MSR_DEVICE
        else:
            vecs = np.ascontiguousarray(np.asarray(vecs), dtype=np.float32)
            where = MSR_HOST
        assert vecs.shape[1] == self.dim
        with torch.cuda.device(self.device):
            check(_lib.load().msr_dense_index_add(self._h, int(vecs.shape[0]), ptr(vecs), where,
                                                  current_stream_ptr(self.device)))

    @property
    def last_fallback(self):
        """Queries of the last tensor-core search whose certificate failed (re-run by the exact pass)."""
        n = ctypes.c_uint32()
        check(_lib.load().msr_dense_last_fallback(self._h, ctypes.byref(n)))
        return int(n.value)

    def profile(self, enable=True):
        """(fused tcgen05 pass ms, re-score ms, query blocks) accumulated since the last call; then switch timing on/off."""
        a, b, n = ctypes.c_double(), ctypes.c_double(), ctypes.c_uint64()
        with torch.cuda.device(self.device):
            check(_lib.load().msr_dense_profile(self._h, int(bool(enable)), ctypes.byref(a), ctypes.byref(b), ctypes.byref(n)))
        return a.value, b.value, int(n.value)

    def tc_scores(self, queries):
        """Tensor-core pass alone: CUDA fp32 queries [nq,dim] -> approximate scores fp32 [nq, n_docs] (fp16 inputs)."""
        q = queries.to(torch.float32).contiguous()
        n = self.ntotal
        ld = (n + 3) // 4 * 4
        out = torch.empty((q.shape[0], ld), dtype=torch.float32, device=q.device)
        with torch.cuda.device(self.device):
            check(_lib.load().msr_dense_tc_scores(self._h, int(q.shape[0]), ptr(q), ptr(out), ld,
                                                  current_stream_ptr(self.device)))
        return out[:, :n]

    def search_static(self, queries, k, out):
        """msr_dense_search_static: the exact pass (mode 1) with its scratch in libmsr.so's static device array -- the call
        itself allocates nothing, so the CALLER passes everything: ``queries`` a contiguous CUDA fp32 tensor [nq,dim],
        ``out`` = (scores f32 [nq,k], ids i64 [nq,k]) contiguous CUDA tensors.  No counterpart in the reference.
        Raises when the index is too large for MSR_STATIC_SCRATCH_BYTES."""
        assert queries.is_cuda and queries.dtype == torch.float32 and queries.is_contiguous() and queries.shape[1] == self.dim
        nq = int(queries.shape[0])
        s, i = out
        assert s.is_cuda and i.is_cuda and s.dtype == torch.float32 and i.dtype == torch.int64
        assert s.is_contiguous() and i.is_contiguous() and tuple(s.shape) == (nq, k) and tuple(i.shape) == (nq, k)
        check(_lib.load().msr_dense_search_static(self._h, nq, ptr(queries), int(k), ptr(s), ptr(i),
                                                  current_stream_ptr(self.device)))
        return out

    def search(self, queries, k, mode=0, out=None):
        """numpy / CPU tensor in -> host arrays out (host path, H2D + D2H inside the call); CUDA tensor in -> CUDA
        tensors out (device path).  ``out`` = preallocated (scores, ids) of the matching kind."""
        if isinstance(queries, torch.Tensor) and not queries.is_cuda:
            q = queries.to(torch.float32).contiguous()          # e.g. pinned host memory
            nq = q.shape[0]
            if out is None:
                out = (torch.empty((nq, k), dtype=torch.float32), torch.empty((nq, k), dtype=torch.int64))
            where = MSR_HOST
        elif isinstance(queries, torch.Tensor) and queries.is_cuda:
            q = queries.to(torch.float32).contiguous()
            nq = q.shape[0]
            if out is None:
                out = (torch.empty((nq, k), dtype=torch.float32, device=q.device),
                       torch.empty((nq, k), dtype=torch.int64, device=q.device))
            where = MSR_DEVICE
        else:
            q = np.ascontiguousarray(np.asarray(queries), dtype=np.float32)
            nq = q.shape[0]
            if out is None:
                out = (np.empty((nq, k), dtype=np.float32), np.empty((nq, k), dtype=np.int64))
            where = MSR_HOST
        assert q.shape[1] == self.dim
        with torch.cuda.device(self.device):
            check(_lib.load().msr_dense_search(self._h, nq, ptr(q), int(k), ptr(out[0]), ptr(out[1]), int(mode), where,
                                               current_stream_ptr(self.device)))
        return out


def fuse_lists(d_scores, d_docs, d_counts, s_scores, s_docs, s_counts, self_ids, remove_query, w_dense, w_sparse,
               dense_f32=True):
    """Kernel (d) on CUDA tensors: returns (scores f64 [nq,kd+ks], docs i32, src u8, counts i32).  float32 score tensors go
    through msr_fuse (rule H3 typing per ``dense_f32``); float64 score tensors -- runs of arbitrary Python floats -- through
    msr_fuse_f64."""
    nq, k_d = d_scores.shape
    k_s = s_scores.shape[1]
    dev = d_scores.device
    di = _dev_index(dev)
    out = (torch.empty((nq, k_d + k_s), dtype=torch.float64, device=dev),
           torch.empty((nq, k_d + k_s), dtype=torch.int32, device=dev),
           torch.empty((nq, k_d + k_s), dtype=torch.uint8, device=dev),
           torch.empty((nq,), dtype=torch.int32, device=dev))
    if d_scores.dtype == torch.float64 or s_scores.dtype == torch.float64:
        assert d_scores.dtype == s_scores.dtype == torch.float64 and not dense_f32, "float64 runs: both lists, dense_f32=False"
        with torch.cuda.device(di):
            check(_lib.load().msr_fuse_f64(di, nq, k_d, ptr(d_scores.contiguous()), ptr(d_docs.contiguous()),
                                           ptr(d_counts.contiguous()), k_s, ptr(s_scores.contiguous()), ptr(s_docs.contiguous()),
                                           ptr(s_counts.contiguous()), ptr(self_ids), int(bool(remove_query)), float(w_dense),
                                           float(w_sparse), ptr(out[0]), ptr(out[1]), ptr(out[2]), ptr(out[3]),
                                           current_stream_ptr(di)))
        return out
    with torch.cuda.device(di):
        check(_lib.load().msr_fuse(di, nq, k_d, ptr(d_scores.contiguous()), ptr(d_docs.contiguous()),
                                   ptr(d_counts.contiguous()), k_s, ptr(s_scores.contiguous()), ptr(s_docs.contiguous()),
                                   ptr(s_counts.contiguous()), ptr(self_ids), int(bool(remove_query)), int(bool(dense_f32)), float(w_dense),
                                   float(w_sparse), ptr(out[0]), ptr(out[1]), ptr(out[2]), ptr(out[3]),
                                   current_stream_ptr(di)))
    return out


def order_rows(scores, docs, counts, self_ids=None, by_score=True, tags=None):
    """msr_order_rows: stable descending order of each row's valid prefix (Python's ``sorted(..., reverse=True)``), and/or
    removal of the row's own doc id with the order kept (``get_run_dict(remove_query=True)``).  scores f32/f64 [nq,w] CUDA,
    docs i32, counts i32; returns (scores, docs, counts[, tags])."""
    assert scores.is_cuda and scores.dtype in (torch.float32, torch.float64) and docs.dtype == torch.int32
    scores, docs = scores.contiguous(), docs.contiguous()
    nq, w = scores.shape
    dev = scores.device
    di = _dev_index(dev)
    o_s, o_d = torch.zeros_like(scores), torch.full_like(docs, -1)
    o_c = torch.empty((nq,), dtype=torch.int32, device=dev)
    o_t = torch.zeros_like(tags) if tags is not None else None
    if self_ids is not None:
        self_ids = self_ids.to(device=dev, dtype=torch.int32).contiguous()
    with torch.cuda.device(di):
        check(_lib.load().msr_order_rows(di, nq, w, ptr(scores), int(scores.dtype == torch.float64), ptr(docs),
                                         ptr(tags.contiguous()) if tags is not None else None, scores.stride(0),
                                         ptr(counts.to(torch.int32).contiguous()), ptr(self_ids), int(bool(by_score)), ptr(o_s), ptr(o_d),
                                         ptr(o_t), ptr(o_c), current_stream_ptr(di)))
    return (o_s, o_d, o_c) if tags is None else (o_s, o_d, o_c, o_t)


RECALL_CUTOFFS = (1, 5, 10, 100, 200)           # metrices.py:9


def recall_hits(docs, counts, t_indptr, t_docs, cutoffs=RECALL_CUTOFFS):
    """Kernel of SURVEY 8(f) row 3.  docs i32 [nq, w] ranked lists (CUDA), counts i32 [nq], targets as a CSR of internal
    ids (t_indptr i64 [nq+1], t_docs i32).  Returns (hits per cutoff as a CUDA int64 tensor, n_eval CUDA int64 [1])."""
    assert docs.is_cuda and docs.dtype == torch.int32 and docs.stride(1) == 1
    nq = docs.shape[0]
    dev = docs.device
    di = _dev_index(dev)
    cuts = torch.tensor(list(cutoffs), dtype=torch.int32, device=dev)
    hits = torch.empty((len(cutoffs),), dtype=torch.int64, device=dev)
    n_eval = torch.empty((1,), dtype=torch.int64, device=dev)
    t_indptr = t_indptr.to(device=dev, dtype=torch.int64).contiguous()
    t_docs = t_docs.to(device=dev, dtype=torch.int32).contiguous()
    counts = counts.to(torch.int32).contiguous()
    with torch.cuda.device(di):
        check(_lib.load().msr_recall_hits(di, nq, ptr(docs), docs.stride(0) if nq else 0, ptr(counts), ptr(t_indptr), ptr(t_docs),
                                          len(cutoffs), ptr(cuts), ptr(hits), ptr(n_eval), current_stream_ptr(di)))
    return hits, n_eval


_DTYPES = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}


def quantise_topk(logits, top, token_to_term=None, vocab_len=None, doc_mode=False):
    """Activation rows on the GPU -> query CSR on the GPU (msr_quantise_topk): (indptr i64 [n+1], terms i32, weights i32).

    logits: CUDA tensor [n_rows, V] (f32 / f16 / bf16, last dim contiguous); token_to_term: CUDA int32 [vocab_len] or None."""
    assert logits.is_cuda and logits.dim() == 2 and logits.stride(1) == 1 and logits.dtype in _DTYPES
    n, V = logits.shape
    dev = logits.device
    vocab_len = V if vocab_len is None else int(vocab_len)
    if token_to_term is not None:
        token_to_term = token_to_term.to(device=dev, dtype=torch.int32).contiguous()
        assert token_to_term.numel() >= min(vocab_len, V)
    indptr = torch.empty((n + 1,), dtype=torch.int64, device=dev)
    terms = torch.empty((n * top,), dtype=torch.int32, device=dev)
    weights = torch.empty((n * top,), dtype=torch.int32, device=dev)
    di = _dev_index(dev)
    with torch.cuda.device(di):
        check(_lib.load().msr_quantise_topk(di, n, V, ptr(logits), _DTYPES[logits.dtype], logits.stride(0), int(top),
                                            vocab_len, ptr(token_to_term), int(bool(doc_mode)), ptr(indptr), ptr(terms),
                                            ptr(weights), current_stream_ptr(di)))
    return indptr, terms, weights


def text_token_ids(text, tokenizer, word_tokenize=None, stop_words=None):
    """Candidate token ids of one caption for ``quantise_candidates`` -- the host half of the reference's text quantiser
    (rule Q2; /root/reference/src/encode.py:97-100,103): the caption is lower-cased and split into words, English
    stop-words and punctuation marks are dropped, every remaining word is tokenised on its own (no special tokens), and
    the distinct ids are returned (as a list; the kernel ranks candidates by value, then position).

    ``word_tokenize`` / ``stop_words`` default to nltk's, as in the reference; pass them explicitly where nltk is absent."""
    import string
    if word_tokenize is None or stop_words is None:
        import nltk                                                # fails loudly when nltk is missing
        word_tokenize = word_tokenize or nltk.word_tokenize
        if stop_words is None:
            stop_words = nltk.corpus.stopwords.words("english")
    ignored = frozenset(stop_words) | frozenset(string.punctuation)
    # The ORDER of the returned list breaks value ties between candidates, and the reference's order is whatever CPython's
    # set iteration yields after these very insertions -- so the ids must go through one set, word by word, id by id.
    ids = set()
    for word in word_tokenize(text.lower()):
        if word not in ignored:
            ids.update(tokenizer.encode(word, add_special_tokens=False))
    return list(ids)


def expansion_vocab_mask(tokenizer, vocab_size=None):
    """u8 [V] mask of the expansion vocabulary (rule behind /root/reference/src/encode.py:38-47), in the form the kernel
    takes it (msr_quantise_candidates' ``expand_mask``): 1 for a vocabulary entry that -- once one leading word-boundary
    marker ('\u2581' or a blank) is removed -- is purely alphabetic or purely digits AND starts with an ASCII lower-case
    letter."""
    vocab = tokenizer.get_vocab()
    size = int(vocab_size) if vocab_size is not None else (max(vocab.values()) + 1 if vocab else 0)
    mask = np.zeros(size, dtype=np.uint8)
    for tok, tid in vocab.items():
        body = tok[1:] if tok[:1] in ("\u2581", " ") else tok
        if 0 <= tid < size and (body.isalpha() or body.isdigit()) and "a" <= body[:1] <= "z":
            mask[tid] = 1
    return mask


def get_filtered_ids(tokenizer):
    """The same vocabulary as a set of ids (the reference's container for it, encode.py:38-47)."""
    return set(np.flatnonzero(expansion_vocab_mask(tokenizer)).tolist())


def quantise_candidates(logits, cand_lists, token_to_term=None, vocab_len=None, doc_mode=False, cap=128, fallback_top=10,
                        filtered_ids=None, num_expended_tokens=0):
    """Candidate branch of the text quantiser on the GPU (msr_quantise_candidates): activation rows + per-row candidate
    token ids (``text_token_ids``) -> query CSR at pitch max(cap, fallback_top) + num_expended_tokens:
    (indptr i64 [n+1], terms i32, weights i32).  ``filtered_ids`` (``get_filtered_ids``) + ``num_expended_tokens`` > 0 add
    the reference's expansion terms (encode.py:137-152)."""
    assert logits.is_cuda and logits.dim() == 2 and logits.stride(1) == 1 and logits.dtype in _DTYPES
    n, V = logits.shape
    assert len(cand_lists) == n
    dev = logits.device
    vocab_len = V if vocab_len is None else int(vocab_len)
    if token_to_term is not None:
        token_to_term = token_to_term.to(device=dev, dtype=torch.int32).contiguous()
        assert token_to_term.numel() >= min(vocab_len, V)
    c_indptr = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum([len(c) for c in cand_lists], out=c_indptr[1:])
    c_tok = np.ascontiguousarray(np.concatenate([np.asarray(c, dtype=np.int64).reshape(-1) for c in cand_lists])
                                 if n else np.zeros(0), dtype=np.int32)
    n_expand = min(int(num_expended_tokens), V, 1024) if filtered_ids is not None else 0      # top_k = min(n, eligible)
    mask = None
    if n_expand > 0:
        if isinstance(filtered_ids, np.ndarray) and filtered_ids.dtype == np.uint8:          # expansion_vocab_mask()
            m = np.zeros(V, dtype=np.uint8)
            m[:min(V, len(filtered_ids))] = filtered_ids[:V]
        else:
            m = np.zeros(V, dtype=np.uint8)
            ids = np.fromiter((i for i in filtered_ids if 0 <= i < V), dtype=np.int64)
            m[ids] = 1
        mask = torch.from_numpy(m).to(dev)
    pitch = max(int(cap), int(fallback_top)) + n_expand
    indptr = torch.empty((n + 1,), dtype=torch.int64, device=dev)
    terms = torch.empty((n * pitch,), dtype=torch.int32, device=dev)
    weights = torch.empty((n * pitch,), dtype=torch.int32, device=dev)
    di = _dev_index(dev)
    with torch.cuda.device(di):
        check(_lib.load().msr_quantise_candidates(
            di, n, V, ptr(logits), _DTYPES[logits.dtype], logits.stride(0), c_indptr.ctypes.data, c_tok.ctypes.data,
            int(cap), int(fallback_top), vocab_len, ptr(token_to_term), int(bool(doc_mode)), ptr(mask), n_expand,
            ptr(indptr), ptr(terms), ptr(weights), current_stream_ptr(di)))
    return indptr, terms, weights


def provenance_hist(scores, src, counts, depth=200):
    """score_statistic.py:406-417 on the device (msr_provenance_hist): fused (scores f64, src u8, counts) from
    fuse_lists -> int64 [3, depth]; row 0/1/2 = dense-only / sparse-only / both, column r = rank r+1."""
    nq = scores.shape[0]
    dev = scores.device
    assert scores.dtype == torch.float64 and src.dtype == torch.uint8 and scores.stride(1) == 1 and src.stride() == scores.stride()
    hist = torch.empty((3, int(depth)), dtype=torch.int64, device=dev)
    di = _dev_index(dev)
    with torch.cuda.device(di):
        check(_lib.load().msr_provenance_hist(di, nq, ptr(scores), ptr(src), scores.stride(0),
                                              ptr(counts.to(torch.int32).contiguous()), int(depth), ptr(hist),
                                              current_stream_ptr(di)))
    return hist
