"""GPU tests of the reference-facing Python surface: LuceneImpactSearcher / FaissFlatSearcher / hybrid.fuse."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "reference_glue.json")


@pytest.fixture(scope="module")
def golden():
    with open(GOLDEN) as f:
        return json.load(f)


def untyped(tv):
    tag, v = tv
    return np.float32(v) if tag == "f32" else (float(v) if tag == "f64" else int(v))


def test_fuse_dict_api_bit_exact_vs_reference(golden, lib_built):
    from mllm_sparse_retrieval_b200.hybrid import fuse, fuse_statistic, get_run_dict
    for case in golden["fuse"]:
        qids = case["qids"]
        dense = get_run_dict(qids, [[np.float32(x) for x in d["scores"]] for d in case["dense"]],
                             [d["docs"] for d in case["dense"]], case["remove_query"])
        sparse = get_run_dict(qids, [[float(x) for x in s["scores"]] for s in case["sparse"]],
                              [s["docs"] for s in case["sparse"]], case["remove_query"])
        w = [case["alpha"], 1 - case["alpha"]]
        fused = fuse([dense, sparse], w)
        for qid in qids:
            exp = case["fused"][qid]
            got = list(fused[qid].items())
            assert [d for d, _ in got] == [d for d, _ in exp]
            for (_, gv), (_, ev) in zip(got, exp):
                e = untyped(ev)
                assert type(gv) is type(e)
                assert gv == e, (qid, gv, e)
        stat = fuse_statistic([dense, sparse], w)
        for qid in qids:
            assert [(d, r.type) for d, r in stat[qid].items()] == [tuple(x) for x in case["fused_tags"][qid]]


def _write_collection(tmp_path, rng, n_docs, vocab, nnz):
    d = tmp_path / "sparse_output"
    (d / "index").mkdir(parents=True)                      # Lucene's directory: present but ignored
    ids = [str(x) for x in rng.permutation(n_docs * 3)[:n_docs]]
    vecs = []
    with open(d / "corpus_0.jsonl", "w") as f:
        for i in ids:
            toks = rng.choice(len(vocab), size=nnz, replace=False)
            vec = {vocab[t]: int(rng.integers(0, 300)) for t in toks}      # some zero weights (never indexed)
            vecs.append(vec)
            f.write(json.dumps({"id": i, "content": "", "vector": vec}) + "\n")
    return d, ids, vecs


def test_lucene_impact_searcher_dropin(oracle, lib_built, tmp_path):
    from mllm_sparse_retrieval_b200.searchers import JWhiteSpaceAnalyzer, LuceneImpactSearcher
    rng = np.random.default_rng(3)
    vocab = [f"tok{i}" for i in range(400)] + ["ġcat", "▁tree", "dog"]
    d, ids, vecs = _write_collection(tmp_path, rng, 700, vocab, 24)
    s = LuceneImpactSearcher(os.path.join(str(d), "index"), None)          # as search.py:273 passes it
    s.set_analyzer(JWhiteSpaceAnalyzer())
    assert s.num_docs == 700
    topics, qids = [], []
    for q in range(20):
        toks = rng.choice(len(vocab), size=10, replace=False)
        text = ""
        for t in toks:
            text += (" " + vocab[t]) * int(rng.integers(0, 40))
        topics.append((text + " unknown_token unknown_token").strip())
        qids.append(f"q{q}")
    topics.append(""); qids.append("empty")
    res = s.batch_search(topics, qids, 50, threads=16)
    # oracle on the same collection, through its own restatement of S1/S4/S8
    order = sorted(range(len(ids)), key=lambda i: ids[i].encode())
    internal = {ids[i]: r for r, i in enumerate(order)}
    terms = {}
    trip = []
    for i, vec in zip(ids, vecs):
        for tok, w in vec.items():
            if w > 0:
                trip.append((terms.setdefault(tok, len(terms)), internal[i], w))
    trip.sort()
    a = np.array(trip, dtype=np.int64)
    indptr = np.zeros(len(terms) + 1, dtype=np.uint64)
    np.cumsum(np.bincount(a[:, 0], minlength=len(terms)), out=indptr[1:])
    qi, qt, qw = [0], [], []
    for text in topics:
        for tok, c in oracle.query_string_to_counts(text).items():
            if tok in terms:
                qt.append(terms[tok]); qw.append(c)
        qi.append(len(qt))
    es, ed, ec = oracle.impact_search_c(indptr, a[:, 1], a[:, 2], 700, np.array(qi, np.uint64), np.array(qt, np.uint32),
                                        np.array(qw, np.uint32), 50)
    sorted_ids = [ids[i] for i in order]
    for r, qid in enumerate(qids):
        hits = res[qid]
        assert len(hits) == ec[r]
        assert [h.docid for h in hits] == [sorted_ids[x] for x in ed[r, :ec[r]]]
        assert [h.score for h in hits] == [float(x) for x in es[r, :ec[r]]]
        assert all(isinstance(h.score, float) and isinstance(h.docid, str) for h in hits)
    assert res["empty"] == []


def test_faiss_flat_searcher_dropin(oracle, lib_built):
    from mllm_sparse_retrieval_b200 import synth
    from mllm_sparse_retrieval_b200.searchers import FaissFlatSearcher
    docs, q = synth.make_dense(1500, 9, dim=256, seed=5)
    docs, q = docs.numpy(), q.numpy()
    r = FaissFlatSearcher(docs)
    assert r.index.ntotal == 0                              # tevatron allocates only; the reference adds next
    r.add(docs)
    assert r.index.ntotal == 1500
    s1, i1 = r.search(q, 20)
    s2, i2 = r.batch_search(q, 20, 2, True)
    np.testing.assert_array_equal(i1, i2)
    np.testing.assert_array_equal(s1, s2)
    assert s1.dtype == np.float32 and i1.dtype == np.int64
    es, ei = oracle.flat_ip_search_numpy(docs, q, 20)
    np.testing.assert_array_equal(i1, ei)
    np.testing.assert_allclose(s1, es, rtol=1e-3, atol=1e-6)


def test_end_to_end_hybrid_like_search_py(oracle, lib_built, tmp_path):
    """The reference's own loop shape: sparse_search + search_queries + get_run_dict + fuse (search.py:55-99,455-461)."""
    from mllm_sparse_retrieval_b200.hybrid import fuse, get_run_dict
    from mllm_sparse_retrieval_b200.searchers import FaissFlatSearcher, LuceneImpactSearcher
    from mllm_sparse_retrieval_b200 import synth
    rng = np.random.default_rng(8)
    vocab = [f"w{i}" for i in range(300)]
    d, ids, vecs = _write_collection(tmp_path, rng, 400, vocab, 30)
    sparse = LuceneImpactSearcher(os.path.join(str(d), "index"), None)
    reps, qreps = synth.make_dense(400, 6, dim=128, seed=9)
    dense = FaissFlatSearcher(reps.numpy()); dense.add(reps.numpy())
    look_up = ids
    batch_ids = [ids[3], ids[10], "nohit", ids[7], "x1", "x2"]
    topics = [" ".join(rng.choice(vocab, size=15)) for _ in batch_ids]
    depth = 100
    # sparse_search (search.py:85-99)
    results = sparse.batch_search(topics, batch_ids, depth, threads=16)
    sparse_scores = [[h.score for h in results[i]] for i in batch_ids]
    sparse_rank = [[h.docid for h in results[i]] for i in batch_ids]
    # search_queries (search.py:55-63)
    all_scores, all_idx = dense.batch_search(qreps.numpy(), depth, 2, True)
    psg = np.array([[str(look_up[x]) for x in row] for row in all_idx])
    dense_run = get_run_dict(batch_ids, all_scores, psg, True)
    sparse_run = get_run_dict(batch_ids, sparse_scores, sparse_rank, True)
    fused = fuse([dense_run, sparse_run], [0.5, 0.5])
    exp = oracle.fuse([oracle.get_run_dict(batch_ids, all_scores, psg, True),
                       oracle.get_run_dict(batch_ids, sparse_scores, sparse_rank, True)], [0.5, 0.5])
    for qid in batch_ids:
        assert list(fused[qid].keys()) == list(exp[qid].keys())
        for doc in exp[qid]:
            assert fused[qid][doc] == exp[qid][doc] and type(fused[qid][doc]) is type(exp[qid][doc])
        assert qid not in fused[qid]


def test_search_cli_writes_reference_format_runs(oracle, lib_built, tmp_path):
    """python -m mllm_sparse_retrieval_b200.search on the files src/encode.py writes (corpus_0.jsonl / corpus_0.pkl /
    query.tsv / query.pkl): hybrid run file equals oracle searchers -> get_run_dict -> fuse -> stable sort."""
    import json, pickle
    from mllm_sparse_retrieval_b200 import search as cli
    from mllm_sparse_retrieval_b200.index import ImpactCollection
    rng = np.random.default_rng(8)
    n_docs, nq, dim, V, depth = 300, 40, 64, 200, 50
    doc_ids = [str(x) for x in rng.permutation(1000)[:n_docs]]
    vecs = [{f"w{j}": int(rng.integers(1, 200)) for j in rng.choice(V, 12, replace=False)} for _ in range(n_docs)]
    sp = tmp_path / "sparse"; sp.mkdir()
    with open(sp / "corpus_0.jsonl", "w") as f:
        for i, v in zip(doc_ids, vecs):
            f.write(json.dumps({"id": i, "content": "", "vector": v}) + "\n")
    reps = rng.standard_normal((n_docs, dim)).astype(np.float32); reps /= np.linalg.norm(reps, axis=1, keepdims=True)
    dn = tmp_path / "dense"; dn.mkdir()
    with open(dn / "corpus_0.pkl", "wb") as f:
        pickle.dump((reps, doc_ids), f)
    qids = [doc_ids[i] if i % 3 == 0 else f"q{i}" for i in range(nq)]            # some query ids equal doc ids
    q_reps = rng.standard_normal((nq, dim)).astype(np.float32); q_reps /= np.linalg.norm(q_reps, axis=1, keepdims=True)
    with open(dn / "query.pkl", "wb") as f:
        pickle.dump((q_reps, qids), f)
    topics = [" ".join(f"w{j}" for j in rng.choice(V, 8) for _ in range(int(rng.integers(1, 5)))) for _ in range(nq)]
    with open(sp / "query.tsv", "w") as f:
        for q, t in zip(qids, topics):
            f.write(f"{q}\t{t}\n")
    out = tmp_path / "runs"
    outputs, names = cli.main(["--passage_reps", str(dn), "--sparse_index", str(sp), "--query_reps", str(dn), "--query_sparse", str(sp),
              "--depth", str(depth), "--alpha", "0.3", "--remove_query", "--save_dir", str(out), "--batch_size", "16",
              "--model_name_or_path", "ignored", "--threads", "16"])
    # expectation through the oracle and the reference-shaped glue
    coll = ImpactCollection.from_vectors(doc_ids, vecs)
    qi, qt, qw = coll.encode_queries(topics)
    ss, sd, sc = oracle.impact_search_c(coll.indptr, coll.post_docs, coll.post_impacts, coll.n_docs, qi, qt, qw, depth)
    ds, di = oracle.flat_ip_search_numpy(reps, q_reps, depth)
    dense_run = oracle.get_run_dict(qids, [[np.float32(x) for x in r] for r in ds], [[doc_ids[j] for j in r] for r in di], True)
    sparse_run = oracle.get_run_dict(qids, [[float(x) for x in ss[r, :sc[r]]] for r in range(nq)],
                                     [[coll.docids[j] for j in sd[r, :sc[r]]] for r in range(nq)], True)
    fused = oracle.fuse([dense_run, sparse_run], [0.3, 0.7])
    lines = (out / "run.fusion.trec").read_text().splitlines()
    got = {}
    for ln in lines:
        q, q0, d, rank, score, name = ln.split()
        assert q0 == "Q0" and name == "fusion"
        got.setdefault(q, []).append((d, int(rank), float(score)))
    for q in qids:
        exp = sorted(fused[q].items(), key=lambda kv: kv[1], reverse=True)
        g = got[q]
        assert [r for _, r, _ in g] == list(range(1, len(g) + 1))
        assert len(g) == len(exp)
        np.testing.assert_allclose([s for _, _, s in g], [float(v) for _, v in exp], rtol=1e-5, atol=1e-7)
        # identical docs wherever neighbouring fused scores are not within float32 noise of each other
        ev = np.array([float(v) for _, v in exp])
        for j, ((gd, _, _), (ed, _)) in enumerate(zip(g, exp)):
            if gd != ed:
                assert np.sum(np.abs(ev - ev[j]) < 1e-6) > 1, (q, j, gd, ed)
        assert q not in [d for d, _, _ in g]                                   # remove_query
    assert (out / "run.dense.trec").exists() and (out / "run.sparse.trec").exists()
    srun = (out / "run.sparse.trec").read_text().splitlines()
    first_q = qids[0]
    exp_docs = [d for d in sparse_run[first_q]["docs"]]
    assert [ln.split()[2] for ln in srun if ln.split()[0] == first_q] == exp_docs
    # Recall@k of the fused run, counted on the device, against the restatement of RecallMetrics (hit metric)
    qrels = {q: {doc_ids[int(j)] for j in rng.choice(n_docs, 3, replace=False)} for q in qids}
    f_qids, _, f_docs, f_cnt = outputs["fusion"]
    rec, n_eval = cli.recall_at(f_qids, {n: i for i, n in enumerate(names)}, f_docs, f_cnt, qrels)
    ranked = {q: {d: -r for d, r, _ in got[q]} for q in qids}                    # run-file order as scores
    exp_counts = oracle.recall_counts(ranked, {q: sorted(qrels[q]) for q in qids})
    assert {k: round(v * len(qids)) for k, v in rec.items()} == exp_counts
    assert n_eval == len(qids)


def test_fuse_on_trec_read_runs_is_float64_exact(lib_built, tmp_path):
    """runs of arbitrary Python floats (re-read with read_trec_run: float('0.1') is not float32(0.1)) are combined
    from the doubles themselves (msr_fuse_f64) -- bit-exact against the reference's own fuse() on such runs
    (tests/golden/make_fuse_f64_golden.py)."""
    from mllm_sparse_retrieval_b200.hybrid import fuse, read_trec_run
    with open(os.path.join(os.path.dirname(GOLDEN), "reference_fuse_f64.json")) as f:
        fx = json.load(f)
    for ci, case in enumerate(fx["cases"]):
        runs = []
        for ri, text in enumerate(case["trec"]):
            p = tmp_path / f"c{ci}_r{ri}.trec"
            p.write_text(text)
            runs.append(read_trec_run(str(p)))
        fused = fuse(runs, [case["alpha"], 1 - case["alpha"]])
        for qid, exp in case["fused"].items():
            got = list(fused[qid].items())
            assert [d for d, _ in got] == [d for d, _ in exp]
            for (_, gv), (_, ev) in zip(got, exp):
                assert type(gv) is float and gv == float.fromhex(ev), (ci, qid, gv, float.fromhex(ev))
