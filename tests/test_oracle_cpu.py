"""CPU tests of the oracle itself: the C restatement against the independent scipy restatement and against
hand-derived known answers for rules S1-S10.  No GPU."""
import numpy as np
import pytest

from helpers import random_csr, random_queries


def test_known_answer_tiny(oracle):
    # 6 docs, 3 terms.  term0: d0:2 d1:1 d2:1 ; term1: d1:3 d3:1 ; term2: d0..d4 (df = N-1 -> dropped by S5)
    indptr = np.array([0, 3, 5, 10], dtype=np.uint64)
    docids = np.array([0, 1, 2, 1, 3, 0, 1, 2, 3, 4], dtype=np.uint32)
    impacts = np.array([2, 1, 1, 3, 1, 9, 9, 9, 9, 9], dtype=np.uint16)
    # query: term0 x5, term1 x2, term2 x100 (dropped), term 7 (unknown)
    qi = np.array([0, 4], dtype=np.uint64)
    qt = np.array([0, 1, 2, 7], dtype=np.uint32)
    qw = np.array([5, 2, 100, 3], dtype=np.uint32)
    for fn in (oracle.impact_search_c, oracle.impact_search_numpy):
        s, d, c = fn(indptr, docids, impacts, 6, qi, qt, qw, 10)
        assert c[0] == 4
        # scores: d0=10, d1=5+6=11, d2=5, d3=2
        assert d[0, :4].tolist() == [1, 0, 2, 3]
        assert s[0, :4].tolist() == [11.0, 10.0, 5.0, 2.0]
        assert d[0, 4:].tolist() == [-1] * 6 and s[0, 4:].tolist() == [0.0] * 6


def test_s5_threshold_df_n_minus_1(oracle):
    # N=5: df=4 -> ln(5/5)=0 -> dropped; df=3 -> ln(5/4)>0 -> kept
    indptr = np.array([0, 4, 7], dtype=np.uint64)
    docids = np.array([0, 1, 2, 3, 0, 1, 2], dtype=np.uint32)
    impacts = np.ones(7, dtype=np.uint16)
    qi = np.array([0, 1, 2], dtype=np.uint64)
    qt = np.array([0, 1], dtype=np.uint32)
    qw = np.array([1, 1], dtype=np.uint32)
    s, d, c = oracle.impact_search_c(indptr, docids, impacts, 5, qi, qt, qw, 5)
    assert c.tolist() == [0, 3]


def test_ties_break_by_doc_id(oracle):
    indptr = np.array([0, 6, 6], dtype=np.uint64)
    docids = np.arange(6, dtype=np.uint32)
    impacts = np.array([3, 7, 3, 7, 3, 7], dtype=np.uint16)
    qi = np.array([0, 1], dtype=np.uint64)
    s, d, c = oracle.impact_search_c(indptr, docids, impacts, 8, qi, np.array([0], np.uint32), np.array([2], np.uint32), 4)
    assert d[0].tolist() == [1, 3, 5, 0] and s[0].tolist() == [14.0, 14.0, 14.0, 6.0]


def test_duplicate_query_terms_add(oracle):
    indptr = np.array([0, 2], dtype=np.uint64)
    docids = np.array([0, 2], dtype=np.uint32)
    impacts = np.array([5, 1], dtype=np.uint16)
    qi = np.array([0, 2], dtype=np.uint64)
    s, d, c = oracle.impact_search_c(indptr, docids, impacts, 4, qi, np.array([0, 0], np.uint32), np.array([123, 100], np.uint32), 3)
    assert s[0, :2].tolist() == [1115.0, 223.0]


def test_float32_rounding_above_2_24(oracle):
    # S6: term product rounded to float32, then the sum rounded to float32
    indptr = np.array([0, 2, 4], dtype=np.uint64)
    docids = np.array([0, 1, 0, 1], dtype=np.uint32)
    impacts = np.array([65535, 65535, 1, 3], dtype=np.uint16)
    qi = np.array([0, 2], dtype=np.uint64)
    qt = np.array([0, 1], dtype=np.uint32)
    qw = np.array([65535, 1], dtype=np.uint32)
    s, d, c = oracle.impact_search_c(indptr, docids, impacts, 8, qi, qt, qw, 2)
    prod = np.float32(65535.0) * np.float32(65535.0)
    exp = [np.float32(np.float64(prod) + 3.0), np.float32(np.float64(prod) + 1.0)]
    assert s[0].tolist() == [float(e) for e in exp]
    assert d[0].tolist() == ([1, 0] if exp[0] != exp[1] else [0, 1])


@pytest.mark.parametrize("seed,n_docs,n_terms,nnz,nq,qlen,k", [
    (1, 200, 64, 16, 20, 8, 10),
    (2, 1000, 512, 32, 30, (1, 40), 100),
    (3, 50, 16, (0, 10), 10, 6, 1000),
    (4, 3000, 2000, 24, 16, 12, 100),
])
def test_c_matches_scipy(oracle, seed, n_docs, n_terms, nnz, nq, qlen, k):
    rng = np.random.default_rng(seed)
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, nnz)
    qi, qt, qw = random_queries(rng, nq, n_terms, qlen, dup=(seed % 2 == 0))
    a = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw, k, threads=2)
    b = oracle.impact_search_numpy(indptr, docids, impacts, n_docs, qi, qt, qw, k)
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)


def test_shard_merge_is_shard_count_invariant(oracle):
    """doc-sharded search + merge == unsharded, bit for bit (global df / N)."""
    rng = np.random.default_rng(11)
    n_docs, n_terms, k = 900, 300, 50
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, 20)
    qi, qt, qw = random_queries(rng, 12, n_terms, 10)
    ref = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw, k)
    df = np.diff(indptr.astype(np.int64)).astype(np.uint32)
    term_of = np.repeat(np.arange(n_terms), df)
    for G in (2, 3, 8):
        bounds = np.linspace(0, n_docs, G + 1).astype(int)
        parts = []
        for g in range(G):
            lo, hi = bounds[g], bounds[g + 1]
            m = (docids >= lo) & (docids < hi)
            ip = np.zeros(n_terms + 1, dtype=np.uint64)
            np.cumsum(np.bincount(term_of[m], minlength=n_terms), out=ip[1:])
            parts.append(oracle.impact_search_c(ip, docids[m] - lo, impacts[m], hi - lo, qi, qt, qw, k,
                                                n_docs_global=n_docs, df_global=df, doc_base=lo))
        merged = oracle.merge_topk([p[0] for p in parts], [p[1] for p in parts], [p[2] for p in parts], k)
        for x, y in zip(ref, merged):
            np.testing.assert_array_equal(x, y)


def test_flat_ip_c_matches_numpy(oracle):
    rng = np.random.default_rng(5)
    docs = rng.standard_normal((300, 64)).astype(np.float32)
    docs[7] = docs[3]                       # exact tie -> id order
    q = rng.standard_normal((9, 64)).astype(np.float32)
    a = oracle.flat_ip_search_c(docs, q, 20, threads=2)
    b = oracle.flat_ip_search_numpy(docs, q, 20)
    np.testing.assert_array_equal(a[1], b[1])
    np.testing.assert_array_equal(a[0], b[0])


def test_c_matches_scipy_property(oracle):
    """Property form of the cross-check: hypothesis draws ragged collections (empty docs, empty queries, out-of-vocabulary
    and zero-weight query terms, k above the hit count, heavy ties from a tiny impact range) and the two independent
    restatements must agree bit for bit."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=40, deadline=None)
    @given(st.integers(0, 2 ** 32 - 1), st.integers(3, 120), st.integers(1, 40), st.integers(0, 12),
           st.integers(0, 9), st.integers(1, 3), st.sampled_from([1, 7, 130]))
    def prop(seed, n_docs, n_terms, nnz, nq, max_imp, k):
        rng = np.random.default_rng(seed)
        indptr, docids, impacts = random_csr(rng, n_docs, n_terms, (0, min(nnz, n_terms) + 1), max_impact=max_imp, zipf=False)
        qi, qt, qw = random_queries(rng, nq, n_terms, (0, 8), max_w=3, dup=True, zipf=False)
        if len(qt):
            qt = qt.copy(); qw = qw.copy()
            qt[rng.random(len(qt)) < 0.1] = n_terms + 5          # a term the index has never seen
            qw[rng.random(len(qw)) < 0.1] = 0
        a = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw, k, threads=2)
        b = oracle.impact_search_numpy(indptr, docids, impacts, n_docs, qi, qt, qw, k)
        for x, y in zip(a, b):
            np.testing.assert_array_equal(x, y)

    prop()


# ---- the timed CPU baselines (oracle/impact_tuned.c) must be the oracle, bit for bit ----
@pytest.mark.parametrize("kind", ["exhaustive", "pruned"])
@pytest.mark.parametrize("seed,n_docs,n_terms,nnz,nq,qlen,k", [
    (1, 200, 64, 16, 20, 8, 10),
    (2, 1000, 512, 32, 30, (1, 40), 100),
    (3, 50, 16, (0, 10), 10, 6, 1000),
    (4, 3000, 2000, 24, 16, 12, 100),
    (5, 70000, 300, 12, 12, 10, 50),             # several 32 K-doc blocks: cursors and thresholds carry over
    (6, 70000, 300, 12, 12, (1, 30), 1),
])
def test_tuned_baselines_match_oracle(oracle, kind, seed, n_docs, n_terms, nnz, nq, qlen, k):
    rng = np.random.default_rng(seed)
    if n_docs > 10000:                           # vectorised generator for the multi-block cases
        t = rng.integers(0, n_terms, size=(n_docs, nnz))
        t.sort(axis=1)
        keep = np.ones_like(t, dtype=bool); keep[:, 1:] = t[:, 1:] != t[:, :-1]
        d = np.repeat(np.arange(n_docs), nnz).reshape(n_docs, nnz)[keep]; t = t[keep]
        w = rng.integers(1, 401, size=len(t))
        order = np.lexsort((d, t))
        indptr = np.zeros(n_terms + 1, dtype=np.uint64)
        np.cumsum(np.bincount(t, minlength=n_terms), out=indptr[1:])
        docids, impacts = d[order].astype(np.uint32), w[order].astype(np.uint16)
    else:
        indptr, docids, impacts = random_csr(rng, n_docs, n_terms, nnz)
    qi, qt, qw = random_queries(rng, nq, n_terms, qlen, dup=(seed % 2 == 0))
    a = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw, k, threads=2)
    b = oracle.impact_search_tuned(indptr, docids, impacts, n_docs, qi, qt, qw, k, kind=kind, threads=2)
    for x, y in zip(a, b[:3]):
        np.testing.assert_array_equal(x, y)
    if kind == "exhaustive":
        df = np.diff(indptr.astype(np.int64))
        kept = (df > 0) & (df + 1 < n_docs)
        valid = (qt < n_terms) & (qw > 0)
        expect = int(df[qt[valid]][kept[qt[valid]]].sum())
        assert b[3] == expect                    # every posting of every kept term was applied


def test_tuned_baselines_wide_path_and_shards(oracle):
    # S6: a per-term product >= 2^24 sends the query down the wide (double) path
    indptr = np.array([0, 2, 4], dtype=np.uint64)
    docids = np.array([0, 1, 0, 1], dtype=np.uint32)
    impacts = np.array([65535, 65535, 1, 3], dtype=np.uint16)
    qi = np.array([0, 2], dtype=np.uint64)
    qt = np.array([0, 1], dtype=np.uint32)
    qw = np.array([65535, 1], dtype=np.uint32)
    a = oracle.impact_search_c(indptr, docids, impacts, 8, qi, qt, qw, 2)
    for kind in ("exhaustive", "pruned"):
        b = oracle.impact_search_tuned(indptr, docids, impacts, 8, qi, qt, qw, 2, kind=kind)
        for x, y in zip(a, b[:3]):
            np.testing.assert_array_equal(x, y)
    # doc shard with collection-wide df / N and a doc base
    rng = np.random.default_rng(21)
    n_docs, n_terms, k = 900, 300, 50
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, 20)
    qi, qt, qw = random_queries(rng, 12, n_terms, 10)
    df = np.diff(indptr.astype(np.int64)).astype(np.uint32)
    term_of = np.repeat(np.arange(n_terms), df)
    lo, hi = 300, 650
    m = (docids >= lo) & (docids < hi)
    ip = np.zeros(n_terms + 1, dtype=np.uint64)
    np.cumsum(np.bincount(term_of[m], minlength=n_terms), out=ip[1:])
    kw = dict(n_docs_global=n_docs, df_global=df, doc_base=lo)
    a = oracle.impact_search_c(ip, docids[m] - lo, impacts[m], hi - lo, qi, qt, qw, k, **kw)
    for kind in ("exhaustive", "pruned"):
        b = oracle.impact_search_tuned(ip, docids[m] - lo, impacts[m], hi - lo, qi, qt, qw, k, kind=kind, **kw)
        for x, y in zip(a, b[:3]):
            np.testing.assert_array_equal(x, y)


def test_tuned_pruned_property(oracle):
    """MaxScore pruning against the exhaustive oracle on hypothesis-drawn ragged collections with heavy score ties
    (tiny impact range): the tie rule (a later doc never displaces an equal score) is what the pruning leans on."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=60, deadline=None)
    @given(st.integers(0, 2 ** 32 - 1), st.integers(3, 400), st.integers(1, 40), st.integers(0, 12),
           st.integers(0, 9), st.integers(1, 4), st.sampled_from([1, 3, 7, 130]))
    def prop(seed, n_docs, n_terms, nnz, nq, max_imp, k):
        rng = np.random.default_rng(seed)
        indptr, docids, impacts = random_csr(rng, n_docs, n_terms, (0, min(nnz, n_terms) + 1), max_impact=max_imp, zipf=False)
        qi, qt, qw = random_queries(rng, nq, n_terms, (0, 8), max_w=3, dup=True, zipf=False)
        if len(qt):
            qt = qt.copy(); qw = qw.copy()
            qt[rng.random(len(qt)) < 0.1] = n_terms + 5
            qw[rng.random(len(qw)) < 0.1] = 0
        a = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw, k, threads=2)
        for kind in ("exhaustive", "pruned"):
            for block in (0, 8, 37):                     # tiny accumulator tiles: the threshold is live from doc 8 on
                b = oracle.impact_search_tuned(indptr, docids, impacts, n_docs, qi, qt, qw, k, kind=kind, threads=2,
                                               block_docs=block)
                for x, y in zip(a, b[:3]):
                    np.testing.assert_array_equal(x, y)

    prop()
