"""Drop-in searcher classes with the call surface the reference uses.

  LuceneImpactSearcher  <- pyserini.search.lucene.LuceneImpactSearcher as used at
                           /root/reference/src/search.py:273-275 (ctor + set_analyzer) and :86-87 (batch_search)
  FaissFlatSearcher     <- tevatron.retriever.searcher.FaissFlatSearcher as used at
                           /root/reference/src/search.py:232,237 (ctor + add), :57,59 (batch_search / search) and
                           :254-270 (``.index`` re-assignment by the faiss GPU cloner)

Both run on libmsr.so (CUDA, sm_100a) only; constructing either without a usable B200 raises.
With ``shard=True`` under an initialised torch.distributed process group the collection is doc-sharded over the
ranks and the per-rank top-k lists are exchanged and merged: by peer stores fused into the scorer's
last kernel when symmetric memory is available (``sharded.PeerExchange``), else by one NCCL all-gather
(``MSR_EXCHANGE`` = auto | p2p | nccl).  Every rank then holds the full result for the queries it passed (all ranks
must pass the SAME queries in that mode).
"""
from __future__ import annotations

import glob
import os

import numpy as np
import torch

from . import ops
from .index import ImpactCollection
from .sharded import PipelinedExchange, make_peer_exchange


class Hit:
    """pyserini hit: the reference reads ``hit.docid`` and ``hit.score`` (search.py:94-97)."""
    __slots__ = ("docid", "score")

    def __init__(self, docid, score):
        self.docid = docid
        self.score = score

    def __repr__(self):
        return f"Hit(docid={self.docid!r}, score={self.score!r})"


class JWhiteSpaceAnalyzer:
    """Placeholder for pyserini.analysis.JWhiteSpaceAnalyzer (search.py:274): queries are always split on whitespace."""


def _dist_info(shard):
    if shard and torch.distributed.is_available() and torch.distributed.is_initialized():
        return torch.distributed.get_rank(), torch.distributed.get_world_size()
    return 0, 1


def _find_collection_dir(index_dir):
    """The reference passes ``<sparse_index>/index`` (Lucene's directory); the JSONL lives one level up."""
    for cand in (index_dir, os.path.dirname(os.path.normpath(index_dir))):
        if glob.glob(os.path.join(cand, "corpus*.jsonl")):
            return cand
    raise FileNotFoundError(f"no corpus*.jsonl in {index_dir!r} or its parent; "
                            "point --sparse_index at the directory src/encode.py wrote")


class LuceneImpactSearcher:
    def __init__(self, index_dir, query_encoder=None, min_idf=0, *, device=None, shard=False, tile_docs=0,
                 collection: ImpactCollection | None = None):
        if query_encoder is not None:
            raise NotImplementedError("only query_encoder=None (token-frequency queries) is used by the reference")
        if min_idf != 0:
            raise NotImplementedError("min_idf is fixed at 0 (rule S5), as in the reference")
        self.collection = collection if collection is not None else ImpactCollection.from_jsonl_dir(
            _find_collection_dir(index_dir))
        c = self.collection
        self.num_docs = c.n_docs
        self.rank, self.world = _dist_info(shard)
        self._px, self._px_failed = {}, False
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.device = dev
        if self.world > 1:
            lo, hi = c.n_docs * self.rank // self.world, c.n_docs * (self.rank + 1) // self.world
            indptr, docs, imps = c.shard(lo, hi)
            self.index = ops.SparseIndex(indptr, docs, imps, hi - lo, device=dev, n_docs_global=c.n_docs,
                                         df_global=c.df, doc_base=lo, tile_docs=tile_docs)
        else:
            self.index = ops.SparseIndex(c.indptr, c.post_docs, c.post_impacts, c.n_docs, device=dev, tile_docs=tile_docs)

    def set_analyzer(self, analyzer):
        return None

    def _peer_exchange(self, nq, k):
        """One ``PeerExchange`` per k, grown when a larger batch arrives (every rank sees the same batch sizes, so the
        ranks take the same decisions).  ``MSR_EXCHANGE`` = auto (default) | p2p | nccl."""
        kind = os.environ.get("MSR_EXCHANGE", "auto")
        if kind == "nccl" or self._px_failed:
            return None
        px = self._px.get(k)
        if px is None or px.nq < nq:
            px = make_peer_exchange(max(nq, 64), k, self.device, self.world, self.rank, kind=kind)
            if px is None:
                self._px_failed = True
                return None
            self._px[k] = px
        return px

    # -- array-level API (used by the CLI, the fusion path and the benchmarks) --------------------------------
    def search_csr(self, d_qi, d_qt, d_qw, k):
        """Query CSR already on the device (i64 indptr, i32 terms, i32 weights -- e.g. straight from ops.quantise_topk):
        returns device tensors (scores f32 [nq,k], docs i32 internal ids, counts i32).  Doc-sharded searchers exchange and
        merge; every rank must pass the same batch."""
        nq = int(d_qi.numel()) - 1
        dev = self.device
        if nq == 0:
            return (torch.empty((0, k), dtype=torch.float32, device=dev), torch.empty((0, k), dtype=torch.int32, device=dev),
                    torch.empty((0,), dtype=torch.int32, device=dev))
        if self.world == 1:
            out = self.index.search(d_qi, d_qt, d_qw, k)
            self.index.status()
            return out
        px = self._peer_exchange(nq, k)
        if px is not None:          # exchange fused into the scorer's last kernel (peer stores), no collective call
            ms, md, mc = px.run(self.index, lambda o: self.index.search(d_qi, d_qt, d_qw, k, out=o), nq=nq)
        else:                       # NCCL all-gather + merge
            ex = PipelinedExchange(nq, k, dev, self.world)
            ms, md, mc = ex.run(lambda i, lo, hi, o: self.index.search(d_qi, d_qt, d_qw, k, out=o))
        self.index.status()
        return ms, md, mc

    def search_arrays(self, queries, k):
        """queries: list of strings / token->count maps.  Returns host arrays (scores f32, docs i32 internal ids, counts)."""
        qi, qt, qw = self.collection.encode_queries(queries)
        if self.world == 1:
            return self.index.search_host(qi, qt, qw, k)
        if len(queries) == 0:
            return (np.empty((0, k), np.float32), np.empty((0, k), np.int32), np.empty((0,), np.uint32))
        t = lambda a, dt: torch.from_numpy(a.astype(dt)).to(self.device)
        ms, md, mc = self.search_csr(t(qi, np.int64), t(qt, np.int32), t(qw, np.int32), k)
        return ms.cpu().numpy(), md.cpu().numpy(), mc.cpu().numpy().astype(np.uint32)

    # -- pyserini surface ---------------------------------------------------------------------------------------
    def batch_search(self, queries, qids, k=10, threads=1, fields=None):
        """{qid: [hit, ...]} with hit.docid (str) / hit.score (float).  ``threads`` sized the JVM pool (S10); the
        whole batch is one kernel launch here."""
        if len(queries) != len(qids):
            raise ValueError("queries and qids differ in length")
        s, d, c = self.search_arrays(list(queries), int(k))
        docids = self.collection.docids
        out = {}
        for i, qid in enumerate(qids):
            n = int(c[i])
            out[qid] = [Hit(docids[d[i, j]], float(s[i, j])) for j in range(n)]
        return out

    def search(self, q, k=10, fields=None):
        return self.batch_search([q], ["q"], k)["q"]

    def close(self):
        self.index.close()


class _FlatIPIndex:
    """Stands in for the ``faiss.IndexFlatIP`` object behind ``FaissFlatSearcher.index``."""

    def __init__(self, dim, device, doc_base=0):
        self.d = int(dim)
        self._ix = ops.DenseIndex(dim, device=device, doc_base=doc_base)

    @property
    def ntotal(self):
        return self._ix.ntotal

    def add(self, x):
        self._ix.add(x)

    def search(self, x, k, mode=0):
        return self._ix.search(x, k, mode=mode)


class FaissFlatSearcher:
    def __init__(self, init_reps: np.ndarray, *, device=None, shard=False):
        """Like tevatron: allocates an EMPTY inner-product index of init_reps.shape[1] dimensions (the reference
        therefore calls ``add(p_reps_0)`` right after, search.py:232-237)."""
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.device = dev
        self.rank, self.world = _dist_info(shard)
        self._n_global = 0
        self._segments = []                       # doc-sharded: (first global id, rows) of every block this rank holds
        self._row_global = None                   # local row -> global id (device, built on first search)
        self._px, self._px_failed = {}, False
        self.index = _FlatIPIndex(init_reps.shape[1], dev)

    def add(self, p_reps: np.ndarray):
        p_reps = np.asarray(p_reps)
        if self.world > 1:
            # doc-sharded (faiss IndexShards behind GpuMultipleClonerOptions.shard, search.py:265-270): of every block
            # added, this rank keeps rows [lo,hi); local row numbers grow with the global ids, so "smaller id wins"
            # means the same thing before and after the mapping back
            n = p_reps.shape[0]
            lo, hi = n * self.rank // self.world, n * (self.rank + 1) // self.world
            self.index.add(p_reps[lo:hi])
            self._segments.append((self._n_global + lo, hi - lo))
            self._row_global = None
        else:
            self.index.add(p_reps)
        self._n_global += p_reps.shape[0]

    def _peer_exchange(self, nq, k):
        kind = os.environ.get("MSR_EXCHANGE", "auto")
        if kind == "nccl" or self._px_failed:
            return None
        px = self._px.get(k)
        if px is None or px.nq < nq:
            px = make_peer_exchange(max(nq, 64), k, self.device, self.world, self.rank, kind=kind)
            if px is None:
                self._px_failed = True
                return None
            self._px[k] = px
        return px

    def search(self, q_reps: np.ndarray, k: int):
        q_reps = np.asarray(q_reps, dtype=np.float32)
        if self.world == 1:
            return self.index.search(q_reps, k)
        dev = self.device
        kk = min(int(k), ops.MSR_MAX_K)
        nq = q_reps.shape[0]
        if nq == 0:
            return np.empty((0, kk), np.float32), np.empty((0, kk), np.int64)
        if self._row_global is None:
            self._row_global = torch.cat([torch.arange(g0, g0 + n, dtype=torch.int32, device=dev) for g0, n in self._segments]
                                         or [torch.zeros(0, dtype=torch.int32, device=dev)])
        if self.index.ntotal:
            s, ids = self.index.search(torch.from_numpy(q_reps).to(dev), kk)
            self.index._ix.last_fallback                         # device path: raises if uncertified queries overflowed
            cnt = (ids >= 0).sum(1).to(torch.int32)
            gids = self._row_global[ids.clamp(min=0)]           # local row -> global id (indexing only)
        else:                                                   # this rank holds no rows (fewer docs than ranks)
            s = torch.zeros((nq, kk), dtype=torch.float32, device=dev)
            gids = torch.full((nq, kk), -1, dtype=torch.int32, device=dev)
            cnt = torch.zeros((nq,), dtype=torch.int32, device=dev)
        px = self._peer_exchange(nq, kk)
        if px is not None:          # rows stored into every GPU's slot over NVLink, one-warp barrier, merge: no collective call
            ms, md, mc = px.exchange_rows(s.contiguous(), gids.contiguous(), cnt)
        else:
            g_s = torch.empty((self.world, nq, kk), dtype=torch.float32, device=dev)
            g_d = torch.empty((self.world, nq, kk), dtype=torch.int32, device=dev)
            g_c = torch.empty((self.world, nq), dtype=torch.int32, device=dev)
            torch.distributed.all_gather_into_tensor(g_s, s.contiguous())
            torch.distributed.all_gather_into_tensor(g_d, gids.contiguous())
            torch.distributed.all_gather_into_tensor(g_c, cnt)
            ms, md, mc = ops.topk_merge(g_s, g_d, g_c, kk)
        ms, md, mc = ms.cpu().numpy(), md.cpu().numpy().astype(np.int64), mc.cpu().numpy()
        pad = np.arange(kk)[None, :] >= mc[:, None]
        ms[pad] = -3.4028234663852886e38
        md[pad] = -1
        return ms, md

    def batch_search(self, q_reps: np.ndarray, k: int, batch_size: int, quiet: bool = False):
        """tevatron loops over ``batch_size`` chunks and concatenates; chunking does not change any result."""
        q_reps = np.asarray(q_reps, dtype=np.float32)
        n = q_reps.shape[0]
        if batch_size <= 0 or n <= batch_size:
            return self.search(q_reps, k)
        scores, ids = [], []
        # chunk at a multiple of batch_size that keeps the launches large
        step = max(batch_size, (4096 // batch_size) * batch_size)
        for lo in range(0, n, step):
            s, i = self.search(q_reps[lo:lo + step], k)
            scores.append(s); ids.append(i)
        return np.concatenate(scores, 0), np.concatenate(ids, 0)
