"""GPU tests of the index builder's device half: msr_invert_docs (doc-major postings -> canonical term-major CSR)."""
import ctypes
import json

import numpy as np
import pytest

from helpers import python_jsonl_collection
import torch

pytestmark = pytest.mark.gpu


def _invert_numpy(n_terms, doc_indptr, terms, imps, rank):
    d = np.repeat(rank.astype(np.int64), np.diff(doc_indptr.astype(np.int64)))
    o = np.lexsort((d, terms.astype(np.int64)))
    indptr = np.zeros(n_terms + 1, dtype=np.uint64)
    np.cumsum(np.bincount(terms, minlength=n_terms), out=indptr[1:])
    return indptr, d[o].astype(np.uint32), imps[o]


@pytest.mark.parametrize("n_docs,n_terms,nnz_hi,where", [(1, 1, 2, 0), (1000, 50, 30, 0), (30000, 5000, 60, 0), (30000, 5000, 60, 1),
                                                         (200000, 131072, 16, 1)])
def test_invert_docs_matches_numpy(lib_built, n_docs, n_terms, nnz_hi, where):
    from mllm_sparse_retrieval_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(n_docs + n_terms)
    lens = rng.integers(0, min(nnz_hi, n_terms) + 1, size=n_docs)
    doc_indptr = np.zeros(n_docs + 1, dtype=np.uint64); np.cumsum(lens, out=doc_indptr[1:])
    nnz = int(doc_indptr[-1])
    # distinct terms inside a document: offsets of a random start, step coprime with n_terms
    start = rng.integers(0, n_terms, size=n_docs)
    pos = np.arange(nnz) - np.repeat(doc_indptr[:-1].astype(np.int64), lens)
    terms = ((np.repeat(start, lens) + pos) % n_terms).astype(np.uint32)
    imps = rng.integers(1, 65536, size=nnz).astype(np.uint16)
    rank = rng.permutation(n_docs).astype(np.uint32)
    e_ip, e_d, e_w = _invert_numpy(n_terms, doc_indptr, terms, imps, rank)
    if where == 0:
        ip = np.empty(n_terms + 1, np.uint64); d = np.empty(nnz, np.uint32); w = np.empty(nnz, np.uint16)
        _lib.check(lib.msr_invert_docs(0, n_docs, n_terms, _lib.ptr(doc_indptr), _lib.ptr(terms), _lib.ptr(imps), _lib.ptr(rank),
                                       _lib.ptr(ip), _lib.ptr(d), _lib.ptr(w), 0, _lib.current_stream_ptr(0)))
    else:
        t = lambda a, dt: torch.from_numpy(a.view(dt) if a.dtype != dt else a).cuda()
        g = (t(doc_indptr, np.int64), t(terms, np.int32), t(imps, np.int16), t(rank, np.int32))
        o = (torch.empty(n_terms + 1, dtype=torch.int64, device="cuda"), torch.empty(nnz, dtype=torch.int32, device="cuda"),
             torch.empty(nnz, dtype=torch.int16, device="cuda"))
        _lib.check(lib.msr_invert_docs(0, n_docs, n_terms, *[_lib.ptr(x) for x in g], *[_lib.ptr(x) for x in o], 1,
                                       _lib.current_stream_ptr(0)))
        ip, d, w = o[0].cpu().numpy().view(np.uint64), o[1].cpu().numpy().view(np.uint32), o[2].cpu().numpy().view(np.uint16)
    np.testing.assert_array_equal(ip, e_ip)
    np.testing.assert_array_equal(d, e_d)
    np.testing.assert_array_equal(w, e_w)


def test_invert_docs_rejects_duplicates_and_bad_ids(lib_built):
    from mllm_sparse_retrieval_b200 import _lib
    lib = _lib.load()
    doc_indptr = np.array([0, 2, 3], dtype=np.uint64)
    imps = np.array([1, 2, 3], dtype=np.uint16)
    out = (np.empty(4, np.uint64), np.empty(3, np.uint32), np.empty(3, np.uint16))
    for terms, what in ((np.array([1, 1, 0], np.uint32), "twice"), (np.array([1, 7, 0], np.uint32), "out of range")):
        rc = lib.msr_invert_docs(0, 2, 3, _lib.ptr(doc_indptr), _lib.ptr(terms), _lib.ptr(imps), None, *[_lib.ptr(x) for x in out], 0,
                                 _lib.current_stream_ptr(0))
        assert rc == -1 and what in lib.msr_last_error().decode()


def test_jsonl_dir_native_gpu_build_searches_like_python_build(oracle, lib_built, tmp_path):
    """End to end: JSONL -> C++ reader -> GPU inversion -> device index; same collection and same search results as the
    pure-Python reader."""
    from mllm_sparse_retrieval_b200.index import ImpactCollection
    from mllm_sparse_retrieval_b200.searchers import LuceneImpactSearcher
    rng = np.random.default_rng(21)
    ids = rng.permutation(9000)[:3000]
    with open(tmp_path / "corpus_0.jsonl", "w") as f:
        for i in ids:
            vec = {f"w{j}": int(rng.integers(1, 300)) for j in rng.choice(400, 20, replace=False)}
            f.write(json.dumps({"id": str(i), "content": "", "vector": vec}) + "\n")
    a = python_jsonl_collection(str(tmp_path))
    b = ImpactCollection.from_jsonl_dir(str(tmp_path), use_cache=False)
    assert a.terms == b.terms and a.docids == b.docids
    np.testing.assert_array_equal(a.indptr, b.indptr); np.testing.assert_array_equal(a.post_docs, b.post_docs)
    np.testing.assert_array_equal(a.post_impacts, b.post_impacts)
    s = LuceneImpactSearcher(str(tmp_path / "index"), None)
    topics = [" ".join(f"w{j}" for j in rng.choice(400, 6)) for _ in range(50)]
    hits = s.batch_search(topics, [str(i) for i in range(50)], k=10)
    qi, qt, qw = a.encode_queries(topics)
    es, ed, ec = oracle.impact_search_c(a.indptr, a.post_docs, a.post_impacts, a.n_docs, qi, qt, qw, 10)
    for r in range(50):
        assert [h.docid for h in hits[str(r)]] == [a.docids[j] for j in ed[r, :ec[r]]]
        assert [h.score for h in hits[str(r)]] == [float(x) for x in es[r, :ec[r]]]
