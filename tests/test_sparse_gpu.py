"""GPU parity tests of kernels (a)+(b) against the oracle, through the C ABI.  Bit-exact (integer work)."""
import numpy as np
import pytest
import torch

from helpers import random_csr, random_queries

pytestmark = pytest.mark.gpu


def _dev_queries(qi, qt, qw):
    dev = torch.device("cuda")
    return (torch.from_numpy(qi.astype(np.int64)).to(dev), torch.from_numpy(qt.astype(np.int32)).to(dev),
            torch.from_numpy(qw.astype(np.int32)).to(dev))


def _search_both(idx, qi, qt, qw, k):
    """device path and host path must agree with each other."""
    s, d, c = idx.search(*_dev_queries(qi, qt, qw), k)
    idx.status()
    hs, hd, hc = idx.search_host(qi, qt, qw, k)
    np.testing.assert_array_equal(s.cpu().numpy(), hs)
    np.testing.assert_array_equal(d.cpu().numpy(), hd)
    np.testing.assert_array_equal(c.cpu().numpy().astype(np.uint32), hc)
    return hs, hd, hc


def _check(oracle, indptr, docids, impacts, n_docs, qi, qt, qw, k, tile_docs=0, **kw):
    from mllm_sparse_retrieval_b200.ops import SparseIndex
    idx = SparseIndex(indptr, docids, impacts, n_docs, tile_docs=tile_docs, **kw)
    got = _search_both(idx, qi, qt, qw, k)
    exp = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw, k, threads=4,
                                 n_docs_global=kw.get("n_docs_global"), df_global=kw.get("df_global"),
                                 doc_base=kw.get("doc_base", 0))
    np.testing.assert_array_equal(got[2], exp[2])
    np.testing.assert_array_equal(got[1], exp[1])
    np.testing.assert_array_equal(got[0], exp[0])
    idx.close()
    return got


def test_known_answer_tiny(oracle, lib_built):
    indptr = np.array([0, 3, 5, 10], dtype=np.uint64)
    docids = np.array([0, 1, 2, 1, 3, 0, 1, 2, 3, 4], dtype=np.uint32)
    impacts = np.array([2, 1, 1, 3, 1, 9, 9, 9, 9, 9], dtype=np.uint16)
    qi = np.array([0, 4, 4, 5], dtype=np.uint64)          # second query is empty
    qt = np.array([0, 1, 2, 7, 2], dtype=np.uint32)
    qw = np.array([5, 2, 100, 3, 9], dtype=np.uint32)
    s, d, c = _check(oracle, indptr, docids, impacts, 6, qi, qt, qw, 10)
    assert c.tolist() == [4, 0, 0]
    assert d[0, :4].tolist() == [1, 0, 2, 3] and s[0, :4].tolist() == [11.0, 10.0, 5.0, 2.0]


@pytest.mark.parametrize("seed,n_docs,n_terms,nnz,nq,qlen,k,tile", [
    (1, 200, 64, 16, 20, 8, 10, 0),
    (2, 1000, 512, 32, 40, (1, 40), 100, 0),          # Flickr-1K shape (reduced vocab)
    (3, 50, 16, (0, 10), 10, 6, 1000, 0),             # k > hits
    (4, 3000, 2000, 24, 16, 12, 100, 256),            # 12 tiles
    (5, 5000, 4096, 128, 32, 128, 100, 0),            # COCO-5K t->i "manual" shape
    (6, 5000, 4096, 128, 8, 300, 1000, 1024),         # > 256 query terms (two staging passes), depth 1000
    (7, 26000, 8192, 10, 24, 128, 100, 0),            # COCO i->t shape: 2 default tiles
    (8, 700, 40, 30, 12, 20, 64, 32),                 # tiny tiles, k > tile size
])
def test_random_parity(oracle, lib_built, seed, n_docs, n_terms, nnz, nq, qlen, k, tile):
    rng = np.random.default_rng(seed)
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, nnz)
    qi, qt, qw = random_queries(rng, nq, n_terms, qlen, dup=(seed % 2 == 0))
    _check(oracle, indptr, docids, impacts, n_docs, qi, qt, qw, k, tile_docs=tile)


def test_massive_ties(oracle, lib_built):
    """every doc has the same score -> top-k is the k smallest doc ids (S8), through the ordered tie path."""
    n_docs = 6000
    indptr = np.array([0, n_docs - 2, n_docs - 2 + 3000], dtype=np.uint64)
    docids = np.concatenate([np.arange(n_docs - 2), np.arange(0, 6000, 2)]).astype(np.uint32)
    impacts = np.concatenate([np.full(n_docs - 2, 7), np.full(3000, 1)]).astype(np.uint16)
    qi = np.array([0, 1, 3], dtype=np.uint64)
    qt = np.array([0, 0, 1], dtype=np.uint32)
    qw = np.array([3, 3, 1], dtype=np.uint32)
    for tile in (0, 2048):
        s, d, c = _check(oracle, indptr, docids, impacts, n_docs, qi, qt, qw, 100, tile_docs=tile)
        assert d[0].tolist() == list(range(100))
        assert d[1].tolist() == list(range(0, 200, 2))


def test_score_above_2_24(oracle, lib_built):
    """S6: per-term float32 rounding and float32 ranking key once sums pass 2^24."""
    rng = np.random.default_rng(9)
    n_docs, n_terms = 4000, 50
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, 20, max_impact=65535, zipf=False)
    qi, qt, qw = random_queries(rng, 10, n_terms, 8, max_w=4000, zipf=False)
    s, d, c = _check(oracle, indptr, docids, impacts, n_docs, qi, qt, qw, 100, tile_docs=512)
    assert s.max() >= 2 ** 24


def test_out_of_range_queries_take_the_wide_path(oracle, lib_built):
    """a query outside the integer kernel's exact range (weight > 65535, or sum_t w_t*max_impact(t) >= 2^31) used
    to fail the whole batch.  It is now scored by the wide path (u64 accumulators, rule S6 reproduced exactly) and every
    query of the batch gets its row: one such query among normal ones, all of them, and more than one wide pass holds
    (host path); bit-exact against the oracle, which rounds every product to float32 and sums in double."""
    rng = np.random.default_rng(31)
    n_docs, n_terms = 5000, 60
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, 16, max_impact=65535, zipf=False)
    qi, qt, qw = random_queries(rng, 24, n_terms, 10, max_w=3000, zipf=False)
    qw = qw.copy()
    big = [int(qi[5]), int(qi[11]) + 2]                       # queries 5 and 11: one huge weight each
    qw[big[0]] = 70000                                        # weight above 65535
    qw[big[1]] = 60000                                        # 60000 * 65535 > 2^31
    for tile in (0, 512):
        s, d, c = _check(oracle, indptr, docids, impacts, n_docs, qi, qt, qw, 50, tile_docs=tile)
        assert s[5, 0] >= 2.0 ** 31 and s[11, 0] >= 2.0 ** 31
    # every query out of range, more than the 8 a wide pass takes: the host path runs further passes
    qw2 = (qw.astype(np.int64) * 0 + 66000).astype(np.uint32)
    _check_host = None
    from mllm_sparse_retrieval_b200.ops import SparseIndex
    from mllm_sparse_retrieval_b200._lib import MsrError, MSR_ERR_RANGE
    idx = SparseIndex(indptr, docids, impacts, n_docs, tile_docs=512)
    hs, hd, hc = idx.search_host(qi, qt, qw2, 50)
    es, ed, ec = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw2, 50, threads=4)
    np.testing.assert_array_equal(hd, ed); np.testing.assert_array_equal(hs, es); np.testing.assert_array_equal(hc, ec)
    # the device path is asynchronous and takes one wide pass: 24 such queries -> the first 8 listed are scored, status says so
    idx.search(*_dev_queries(qi, qt, qw2), 50)
    with pytest.raises(MsrError) as e:
        idx.status()
    assert e.value.code == MSR_ERR_RANGE
    s, d, c = idx.search(*_dev_queries(qi, qt, qw), 50)       # handle still usable; 2 wide queries fit one pass
    idx.status()
    es, ed, ec = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw, 50, threads=4)
    np.testing.assert_array_equal(d.cpu().numpy(), ed); np.testing.assert_array_equal(s.cpu().numpy(), es)
    idx.close()


def test_malformed_csr_rejected(lib_built):
    from mllm_sparse_retrieval_b200._lib import MsrError
    from mllm_sparse_retrieval_b200.ops import SparseIndex
    with pytest.raises(MsrError):
        SparseIndex(np.array([0, 2], np.uint64), np.array([3, 1], np.uint32), np.array([1, 1], np.uint16), 8)
    with pytest.raises(MsrError):
        SparseIndex(np.array([0, 1], np.uint64), np.array([9], np.uint32), np.array([1], np.uint16), 8)


def test_doc_sharded_merge_invariant(oracle, lib_built):
    """per-shard search (global df/N) + msr_topk_merge == unsharded, for G = 2, 3, 8."""
    from mllm_sparse_retrieval_b200.ops import SparseIndex, topk_merge
    rng = np.random.default_rng(21)
    n_docs, n_terms, k = 5000, 1000, 100
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, 40)
    qi, qt, qw = random_queries(rng, 25, n_terms, 16)
    ref = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw, k, threads=4)
    df = np.diff(indptr.astype(np.int64)).astype(np.uint32)
    term_of = np.repeat(np.arange(n_terms), df)
    dq = _dev_queries(qi, qt, qw)
    for G in (2, 3, 8):
        bounds = np.linspace(0, n_docs, G + 1).astype(int)
        outs = []
        for g in range(G):
            lo, hi = bounds[g], bounds[g + 1]
            m = (docids >= lo) & (docids < hi)
            ip = np.zeros(n_terms + 1, dtype=np.uint64)
            np.cumsum(np.bincount(term_of[m], minlength=n_terms), out=ip[1:])
            idx = SparseIndex(ip, docids[m] - lo, impacts[m], hi - lo, n_docs_global=n_docs, df_global=df,
                              doc_base=int(lo), tile_docs=300)
            outs.append(idx.search(*dq, k))
            idx.close()
        ms, md, mc = topk_merge(torch.stack([o[0] for o in outs]), torch.stack([o[1] for o in outs]),
                                torch.stack([o[2] for o in outs]), k)
        np.testing.assert_array_equal(mc.cpu().numpy().astype(np.uint32), ref[2])
        np.testing.assert_array_equal(md.cpu().numpy(), ref[1])
        np.testing.assert_array_equal(ms.cpu().numpy(), ref[0])


def test_query_work_counts(oracle, lib_built):
    from mllm_sparse_retrieval_b200.ops import SparseIndex
    rng = np.random.default_rng(31)
    n_docs, n_terms = 2000, 300
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, 30)
    qi, qt, qw = random_queries(rng, 17, n_terms, 9)
    idx = SparseIndex(indptr, docids, impacts, n_docs)
    p, c = idx.query_work(*_dev_queries(qi, qt, qw))
    df = np.diff(indptr.astype(np.int64))
    keep = (df > 0) & (df + 1 < n_docs)
    assert p == int(df[qt][keep[qt]].sum()) and c == int(keep[qt].sum())


def test_synthetic_generator_config_shapes(oracle, lib_built):
    """Flickr-1K stand-in of SURVEY.md 8(d) config 1 (V reduced for test time), generated on the GPU."""
    from mllm_sparse_retrieval_b200 import synth
    from mllm_sparse_retrieval_b200.ops import SparseIndex
    coll = synth.make_collection(1000, 128, V=8192, seed=1001, device="cuda")
    qs = synth.make_queries(500, 12, V=8192, seed=1001, device="cuda")
    idx = SparseIndex(coll.indptr, coll.docids, coll.impacts, coll.n_docs)
    s, d, c = idx.search(qs.indptr, qs.terms, qs.weights, 100)
    idx.status()
    exp = oracle.impact_search_c(*coll.numpy(), coll.n_docs, *qs.numpy(), 100, threads=4)
    np.testing.assert_array_equal(d.cpu().numpy(), exp[1])
    np.testing.assert_array_equal(s.cpu().numpy(), exp[0])


@pytest.mark.parametrize("k,tile,qlen", [(10, 512, 16), (100, 512, 16), (100, 1000, 40), (700, 1024, 24)])
def test_threshold_sharing_large_batch(oracle, lib_built, k, tile, qlen):
    """Enough queries that the tiles of one query run far apart in time: later tiles then select by the per-query
    threshold published by earlier ones (sparse_kernel.cuh) and merge their candidates into the shared list."""
    rng = np.random.default_rng(77 + k)
    n_docs, n_terms = 6000, 1500
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, 40)
    qi, qt, qw = random_queries(rng, 1200, n_terms, qlen)
    _check(oracle, indptr, docids, impacts, n_docs, qi, qt, qw, k, tile_docs=tile)


def test_threshold_overflow_falls_back(oracle, lib_built):
    """All docs tie: every later tile finds more candidates at the threshold than it has slots for."""
    n_docs = 8192
    indptr = np.array([0, n_docs - 2], dtype=np.uint64)
    docids = np.arange(n_docs - 2, dtype=np.uint32)
    impacts = np.full(n_docs - 2, 5, dtype=np.uint16)
    nq = 900
    qi = np.arange(nq + 1, dtype=np.uint64)
    qt = np.zeros(nq, dtype=np.uint32)
    qw = (1 + np.arange(nq) % 7).astype(np.uint32)
    s, d, c = _check(oracle, indptr, docids, impacts, n_docs, qi, qt, qw, 50, tile_docs=512)
    assert d[0].tolist() == list(range(50))


@pytest.mark.parametrize("env,val", [("MSR_SPARSE_NO_THETA", "1"), ("MSR_SPARSE_NO_DENSE", "1"), ("MSR_SPARSE_FAST_SELECT", "0")])
def test_layout_switches_do_not_change_results(oracle, lib_built, monkeypatch, env, val):
    """Dense impact rows, threshold sharing and the bounded select are pure optimisations: switching any off changes nothing."""
    monkeypatch.setenv(env, val)
    rng = np.random.default_rng(5)
    n_docs, n_terms = 5000, 800
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, 48, max_impact=65535)
    qi, qt, qw = random_queries(rng, 600, n_terms, 12, max_w=1500)
    s, d, c = _check(oracle, indptr, docids, impacts, n_docs, qi, qt, qw, 100, tile_docs=640)
    assert s.max() >= 2 ** 24          # S6 rounding exercised on the posting path too


def test_tile_size_not_multiple_of_8(oracle, lib_built):
    rng = np.random.default_rng(6)
    indptr, docids, impacts = random_csr(rng, 3001, 300, 30)
    qi, qt, qw = random_queries(rng, 400, 300, 10)
    _check(oracle, indptr, docids, impacts, 3001, qi, qt, qw, 20, tile_docs=501)


def test_threshold_read_is_uniform_per_cta_under_cold_timing(oracle, lib_built):
    """Regression: every thread used to read the query's shared threshold on its own; with the tiles of a query running
    skewed in time (cold L2 / TLB: first search after a build, or an L2 flush) some threads of a CTA could still see 0
    (exact select) while others took the threshold path.  All CTAs of this case are co-resident, k is close to the
    tile size and L2 is flushed before every search; 1 in ~30 searches failed before the fix."""
    from mllm_sparse_retrieval_b200.ops import SparseIndex
    rng = np.random.default_rng(6)
    n_docs, n_terms, k = 5000, 4096, 1000
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, 128)
    qi, qt, qw = random_queries(rng, 8, n_terms, 300, dup=True)
    exp = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw, k, threads=4)
    dq = _dev_queries(qi, qt, qw)
    for b in range(6):
        idx = SparseIndex(indptr, docids, impacts, n_docs, tile_docs=1024)
        for r in range(25):
            flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
            flush.fill_(r)
            del flush
            s, d, c = idx.search(*dq, k)
            idx.status()
            np.testing.assert_array_equal(d.cpu().numpy(), exp[1])
            np.testing.assert_array_equal(s.cpu().numpy(), exp[0])
        idx.close()


def test_fused_exchange_slots_and_strided_merge(oracle, lib_built):
    """SURVEY.md 8(e), fused exchange: with msr_sparse_set_peer_outputs every shard's tile-merge epilogue also stores its
    rows into its slot of the other "GPUs'" exchange buffers ([world][scores nq*k | docs nq*k | counts nq], here all on one
    device), and msr_topk_merge_strided over any of those buffers equals the unsharded search.  Bit-exact."""
    from mllm_sparse_retrieval_b200.ops import SparseIndex, topk_merge_strided
    rng = np.random.default_rng(23)
    n_docs, n_terms, k, nq, G = 6000, 1200, 100, 37, 3
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, 40)
    qi, qt, qw = random_queries(rng, nq, n_terms, 16)
    ref = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw, k, threads=4)
    df = np.diff(indptr.astype(np.int64)).astype(np.uint32)
    term_of = np.repeat(np.arange(n_terms), df)
    dq = _dev_queries(qi, qt, qw)
    dev = torch.device("cuda")
    slot = 2 * nq * k + nq + 5                                            # padded: strides need not be tight
    bufs = [torch.full((G * slot,), -7, dtype=torch.int32, device=dev) for _ in range(G)]   # one per emulated GPU

    def slot_ptrs(b, who):
        p = b.data_ptr() + 4 * who * slot
        return p, p + 4 * nq * k, p + 8 * nq * k

    bounds = np.linspace(0, n_docs, G + 1).astype(int)
    for g in range(G):
        lo, hi = bounds[g], bounds[g + 1]
        m = (docids >= lo) & (docids < hi)
        ip = np.zeros(n_terms + 1, dtype=np.uint64)
        np.cumsum(np.bincount(term_of[m], minlength=n_terms), out=ip[1:])
        idx = SparseIndex(ip, docids[m] - lo, impacts[m], hi - lo, n_docs_global=n_docs, df_global=df,
                          doc_base=int(lo), tile_docs=500)
        peers = [slot_ptrs(bufs[r], g) for r in range(G) if r != g]
        idx.set_peer_outputs([p[0] for p in peers], [p[1] for p in peers], [p[2] for p in peers])
        o = g * slot
        own = (bufs[g][o:o + nq * k].view(torch.float32).view(nq, k), bufs[g][o + nq * k:o + 2 * nq * k].view(nq, k),
               bufs[g][o + 2 * nq * k:o + 2 * nq * k + nq])
        idx.search(*dq, k, out=own)
        idx.set_peer_outputs()                                            # off again: the next search must not touch the slots
        before = [b.clone() for b in bufs]
        idx.search(*dq, k)
        torch.cuda.synchronize()
        for b, b0 in zip(bufs, before):
            assert torch.equal(b, b0)
        idx.close()
    for r in range(G):                                                    # every "GPU" holds the same G lists
        assert torch.equal(bufs[r], bufs[0])
        out = (torch.empty((nq, k), dtype=torch.float32, device=dev), torch.empty((nq, k), dtype=torch.int32, device=dev),
               torch.empty((nq,), dtype=torch.int32, device=dev))
        s0, d0, c0 = slot_ptrs(bufs[r], 0)
        topk_merge_strided(s0, d0, c0, G, nq, k, slot, slot, k, out, dev)
        np.testing.assert_array_equal(out[2].cpu().numpy().astype(np.uint32), ref[2])
        np.testing.assert_array_equal(out[1].cpu().numpy(), ref[1])
        np.testing.assert_array_equal(out[0].cpu().numpy(), ref[0])
    pad = bufs[0].view(G, slot)[:, 2 * nq * k + nq:]
    assert bool((pad == -7).all())                                        # nothing written past a slot


def test_strided_merge_rejects_short_strides(lib_built):
    from mllm_sparse_retrieval_b200._lib import MsrError
    from mllm_sparse_retrieval_b200.ops import topk_merge_strided
    dev = torch.device("cuda")
    b = torch.zeros(4096, dtype=torch.int32, device=dev)
    out = (torch.empty((4, 10), dtype=torch.float32, device=dev), torch.empty((4, 10), dtype=torch.int32, device=dev),
           torch.empty((4,), dtype=torch.int32, device=dev))
    with pytest.raises(MsrError):
        topk_merge_strided(b.data_ptr(), b.data_ptr(), b.data_ptr(), 2, 4, 10, 39, 4, 10, out, dev)


def _one_term_index(scores):
    """Index whose only term gives doc d the impact scores[d] (0 = no posting): with query weight 1 the score IS scores[d]."""
    docs = np.nonzero(scores)[0].astype(np.uint32)
    return (np.array([0, len(docs)], dtype=np.uint64), docs, scores[docs].astype(np.uint16))


@pytest.mark.parametrize("tile", [8192, 2048])
def test_bounded_select_edges(oracle, lib_built, tile):
    """The exact select first bounds the top-k by the k-th largest per-thread maximum (g) and collects score >= g.
    Edges: ties at g that fit the 512-key buffer (order by doc id), ties that do not (radix fallback), hits owned by
    fewer than k threads (g = 0, radix fallback), k = 1 and k = 512 (the largest k the bound handles)."""
    rng = np.random.default_rng(41)
    n = 8192
    qi = np.array([0, 1], dtype=np.uint64)
    qt = np.array([0], dtype=np.uint32)
    qw = np.array([1], dtype=np.uint32)
    # (a) 50 docs at 9, 300 at 7, the rest at 5: the bound lands on 7, 350 candidates, the last 50 places are ties
    sc = np.full(n, 5, dtype=np.int64)
    perm = rng.permutation(n)
    sc[perm[:50]] = 9
    sc[perm[50:350]] = 7
    sc[perm[-2:]] = 0                                       # df = N - 2: the term survives the S5 filter
    s, d, c = _check(oracle, *_one_term_index(sc), n, qi, qt, qw, 100, tile_docs=tile)
    assert (s[0, :50] == 9).all() and (s[0, 50:] == 7).all() and (np.diff(d[0, 50:]) > 0).all()
    # (b) same, but every other doc ties at the k-th score: more than 512 candidates
    sc[:] = 5
    sc[perm[:50]] = 9
    sc[perm[-2:]] = 0
    s, d, c = _check(oracle, *_one_term_index(sc), n, qi, qt, qw, 100, tile_docs=tile)
    assert (s[0, 50:] == 5).all() and (np.diff(d[0, 50:]) > 0).all()
    # (c) hits only on docs that 40 of the 512 threads own (thread t reads 16-byte vectors t, t + 512, ...): more than
    #     k hits, fewer than k non-zero thread maxima
    docs = np.arange(n)
    sc = np.where((docs // 4) % 512 < 40, rng.integers(1, 400, n), 0)
    assert np.count_nonzero(sc[:tile]) > 100
    _check(oracle, *_one_term_index(sc), n, qi, qt, qw, 100, tile_docs=tile)
    # (d) distinct-ish random scores, k = 1 / 512 / 513
    sc = rng.integers(1, 60000, n)
    sc[perm[-2:]] = 0
    for k in (1, 512, 513):
        _check(oracle, *_one_term_index(sc), n, qi, qt, qw, k, tile_docs=tile)


@pytest.mark.parametrize("n_lists,k_in,k,nq,fill", [
    (41, 100, 100, 3, "full"),          # 4100 keys > the 2048-key buffer: radix-select path (small batch at 1 M docs)
    (41, 1000, 1000, 2, "full"),        # the reference's depth (scripts/search.sh:25): 41 000 keys per query
    (8, 1000, 1000, 5, "ragged"),       # cross-GPU merge shape, ragged counts
    (300, 100, 37, 4, "ragged"),        # many lists, k below k_in
    (64, 64, 4096, 2, "full"),          # k above the number of keys: everything, sorted
    (5, 100, 100, 300, "ragged"),       # buffer path (<= 2048 keys), many queries
])
def test_topk_merge_select_path_vs_numpy(lib_built, n_lists, k_in, k, nq, fill):
    """msr_topk_merge on lists too long for one shared-memory sort (the MSB-first radix select over the lists in place)
    and on short ones, against a NumPy lexsort with the same (score desc, id asc) key.  Heavy score ties."""
    from mllm_sparse_retrieval_b200.ops import topk_merge
    rng = np.random.default_rng(n_lists * 1000 + k)
    dev = torch.device("cuda")
    scores = rng.integers(1, 400, size=(n_lists, nq, k_in)).astype(np.float32) * 64.0       # many equal scores
    docs = np.stack([rng.permutation(n_lists * k_in * 4)[:n_lists * k_in].reshape(n_lists, k_in) for _ in range(nq)], 1).astype(np.int32)
    counts = np.full((n_lists, nq), k_in, dtype=np.int32) if fill == "full" else rng.integers(0, k_in + 1, size=(n_lists, nq)).astype(np.int32)
    ms, md, mc = topk_merge(torch.from_numpy(scores).to(dev), torch.from_numpy(docs).to(dev), torch.from_numpy(counts).to(dev), k)
    ms, md, mc = ms.cpu().numpy(), md.cpu().numpy(), mc.cpu().numpy()
    for q in range(nq):
        s = np.concatenate([scores[j, q, :counts[j, q]] for j in range(n_lists)])
        d = np.concatenate([docs[j, q, :counts[j, q]] for j in range(n_lists)])
        order = np.lexsort((d, -s.astype(np.float64)))[:k]
        n = len(order)
        assert mc[q] == n
        np.testing.assert_array_equal(md[q, :n], d[order])
        np.testing.assert_array_equal(ms[q, :n], s[order])
        assert (md[q, n:] == -1).all() and (ms[q, n:] == 0).all()


@pytest.mark.parametrize("nq,k", [(1, 100), (4, 1000), (2, 10), (3, 4096)])
def test_small_batch_deep_k_matches_oracle(oracle, lib_built, nq, k):
    """The reference's operating point (scripts/search.sh: batch 2, depth 1000; search_sparse.sh: batch 4, depth 10): few
    queries, many tiles, every tile delivers a full list -> the merge's select path.  Bit-exact against the oracle."""
    from mllm_sparse_retrieval_b200 import synth
    coll = synth.make_collection(60000, 48, V=8192, seed=77, device="cuda")
    qs = synth.make_queries(nq, 24, V=8192, seed=78, device="cuda")
    _check(oracle, *coll.numpy(), coll.n_docs, *qs.numpy(), k, tile_docs=2048)          # 30 tiles
