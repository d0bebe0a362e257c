"""CPU tests: the oracle's restatement of the reference's in-tree glue and the product's host-side mirrors against
fixtures generated from the reference itself (tests/golden/make_golden.py).  No GPU, no /root/reference at run time."""
import json
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "reference_glue.json")


@pytest.fixture(scope="module")
def golden():
    with open(GOLDEN) as f:
        return json.load(f)


def untyped(tv):
    tag, v = tv
    return np.float32(v) if tag == "f32" else (float(v) if tag == "f64" else int(v))


def runs_of(case, get_run_dict):
    qids = case["qids"]
    dense = get_run_dict(qids, [[np.float32(x) for x in d["scores"]] for d in case["dense"]],
                         [d["docs"] for d in case["dense"]], case["remove_query"])
    sparse = get_run_dict(qids, [[float(x) for x in s["scores"]] for s in case["sparse"]],
                          [s["docs"] for s in case["sparse"]], case["remove_query"])
    return dense, sparse


def check_fused(case, fused):
    for qid in case["qids"]:
        exp = case["fused"][qid]
        got = list(fused[qid].items())
        assert [d for d, _ in got] == [d for d, _ in exp]                         # insertion order (H1/H4)
        for (_, gv), (_, ev) in zip(got, exp):
            e = untyped(ev)
            assert type(gv) is type(e), (type(gv), type(e))                         # H3 typing
            assert gv == e, (gv, e)                                                  # bit-exact


@pytest.mark.parametrize("which", ["oracle", "product"])
def test_get_run_dict_matches_reference(golden, which):
    if which == "oracle":
        from oracle.oracle import get_run_dict
    else:
        from mllm_sparse_retrieval_b200.hybrid import get_run_dict
    for case in golden["fuse"]:
        dense, sparse = runs_of(case, get_run_dict)
        for run, mm in ((dense, case["dense_run_minmax"]), (sparse, case["sparse_run_minmax"])):
            for qid in case["qids"]:
                lo, hi = untyped(mm[qid][0]), untyped(mm[qid][1])
                assert run[qid]["min_score"] == lo and run[qid]["max_score"] == hi
                assert type(run[qid]["min_score"]) is type(lo)
        for qid in case["qids"]:
            assert list(dense[qid]["docs"].keys()) == case["dense_run_docs"][qid]
    rd = get_run_dict(["5", "x"], [[30., 20.], []], [["7", "5"], []], True)
    assert rd == golden["glue"]["run_dict_remove"]


def test_oracle_fuse_matches_reference(golden):
    from oracle.oracle import fuse, fuse_tags, get_run_dict
    for case in golden["fuse"]:
        dense, sparse = runs_of(case, get_run_dict)
        check_fused(case, fuse([dense, sparse], [case["alpha"], 1 - case["alpha"]]))
        tags = fuse_tags([dense, sparse])
        for qid in case["qids"]:
            assert list(tags[qid].items()) == [tuple(x) for x in case["fused_tags"][qid]]


def test_trec_io_matches_reference(golden, tmp_path):
    from mllm_sparse_retrieval_b200.hybrid import read_trec_run, write_trec_run
    g = golden["trec"]
    run = {"7": g["run"]["7"], "8": {"30": np.float32(0.75), "2": 1.5}}
    p = tmp_path / "run.trec"
    write_trec_run(run, str(p), name="fusion")
    assert p.read_text() == g["text"]
    assert read_trec_run(str(p)) == g["read_back"]


def test_quantiser_and_query_string_match_reference(golden):
    from oracle.oracle import build_doc_vector, build_query_string, quantise_topk
    for c in golden["quantiser"]:
        logits = np.asarray(c["logits"], dtype=np.float32)
        k = c["sparse_length"] if c["sparse_manual"] else 128
        ids, vals = quantise_topk(logits, k)
        assert vals.tolist() == c["values"]
        if not c["is_filtered"] and len(c["tokens"]) == len(ids):
            # torch.topk leaves the order of exactly tied logits unspecified: compare tokens at untied ranks only
            uniq, cnt = np.unique(logits, return_counts=True)
            tied = set(uniq[cnt > 1].tolist())
            for r, i in enumerate(ids):
                if float(logits[i]) not in tied:
                    assert c["vocab"][i].lower() == c["tokens"][r]
        assert build_query_string(c["tokens"], c["values"]) == c["query_string"]
        assert build_doc_vector(c["tokens"], c["values"]) == c["doc_vector"]


def test_query_string_to_csr_sums_duplicates(golden):
    """Q3 + S4: a query string with a repeated lower-cased token encodes to one term with the summed count."""
    from mllm_sparse_retrieval_b200.index import ImpactCollection
    c = golden["quantiser"][1]
    coll = ImpactCollection.from_vectors(["d0", "d1"], [c["doc_vector"], {"zebra": 3, "unused": 0, "neg": -2}])
    qi, qt, qw = coll.encode_queries([c["query_string"], "", "nosuchtoken zebra zebra"])
    from collections import Counter
    cnt = Counter(c["query_string"].split())
    got = {coll.terms[t]: int(w) for t, w in zip(qt[int(qi[0]):int(qi[1])], qw[int(qi[0]):int(qi[1])])}
    assert got == {t: n for t, n in cnt.items() if t in coll.term_to_id}
    assert int(qi[2]) == int(qi[1])                                   # empty query
    assert [coll.terms[t] for t in qt[int(qi[2]):]] == ["zebra"] and qw[int(qi[2]):].tolist() == [2]
    assert "unused" not in coll.term_to_id and "neg" not in coll.term_to_id     # S1: non-positive weights not indexed


def test_collection_doc_order_and_cache(tmp_path):
    """S8 internal ids = UTF-8 byte order of docid strings; JSONL loader + cache round trip."""
    from mllm_sparse_retrieval_b200.index import ImpactCollection
    d = tmp_path / "sparse"
    d.mkdir()
    rows = [{"id": "99", "content": "", "vector": {"a": 3, "b": 1}}, {"id": "100", "content": "", "vector": {"a": 2}},
            {"id": "9", "content": "", "vector": {"b c": 4, "": 7}}]
    (d / "corpus_0.jsonl").write_text("\n".join(json.dumps(r) for r in rows) + "\n")
    coll = ImpactCollection.from_jsonl_dir(str(d), host_inversion=True)      # no GPU in the CPU suite: explicit opt-in
    assert coll.docids == ["100", "9", "99"]
    a = coll.term_to_id["a"]
    lo, hi = int(coll.indptr[a]), int(coll.indptr[a + 1])
    assert coll.post_docs[lo:hi].tolist() == [0, 2] and coll.post_impacts[lo:hi].tolist() == [2, 3]
    assert set(coll.terms) == {"a", "b", "c"}                          # whitespace-analysed key, empty key dropped
    again = ImpactCollection.from_jsonl_dir(str(d))                    # served from <dir>/msr_index (raw arrays, no GPU needed)
    assert (d / "msr_index" / "meta.json").exists()
    assert again.docids == coll.docids and again.terms == coll.terms
    np.testing.assert_array_equal(again.post_docs, coll.post_docs)


# ---- Recall@k (metrices.py) ----------------------------------------------------------------------------------
METRICS = os.path.join(os.path.dirname(__file__), "golden", "reference_metrics.json")


def metrics_cases():
    with open(METRICS) as f:
        return json.load(f)["metrics"]


def test_oracle_recall_counts_match_reference():
    """oracle.recall_counts against RecallMetrics.sort_and_count of the reference (hit metric, stable order, empty
    lists skipped) on the committed fixtures."""
    from oracle import oracle
    for case in metrics_cases():
        for name in ("dense", "sparse", "fusion"):
            run = case["runs"][name]
            docs = {q: (v if name == "fusion" else v["docs"]) for q, v in run.items()}
            got = oracle.recall_counts(docs, case["targets"])
            assert {str(k): v for k, v in got.items()} == case["counts"][name], (case["nq"], name)


def test_text_quantiser_matches_reference(golden):
    """Q2: oracle.quantise_text_row against encode.get_text_valid_tokens_values itself (all three branches, the
    `i < len(vocab_dict)` quirk, > 128 candidates), and ops.text_token_ids against the reference's candidate set."""
    from oracle.oracle import build_doc_vector, build_query_string, quantise_text_row
    from mllm_sparse_retrieval_b200.ops import text_token_ids
    sys_path_hack = os.path.join(os.path.dirname(__file__), "golden")
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(sys_path_hack, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec); spec.loader.exec_module(mg)
    assert len(golden["text_quantiser"]) >= 12 and sum(c["num_expended_tokens"] > 0 for c in golden["text_quantiser"]) >= 4
    for c in golden["text_quantiser"]:
        logits = np.asarray(c["logits"], dtype=np.float32)
        cands = text_token_ids(c["text"], mg.FakeTokenizer(len(logits)), word_tokenize=lambda s: s.split(),
                               stop_words=c["stop_words"])
        assert cands == c["token_ids"]
        toks, vals = quantise_text_row(logits, cands, c["vocab"], c["sparse_manual"], c["sparse_length"], c["is_filtered"],
                                       filtered_ids=c["filtered_ids"], num_expended_tokens=c["num_expended_tokens"])
        assert toks == c["tokens"] and vals.tolist() == c["values"]
        assert build_query_string(toks, vals) == c["query_string"]
        assert build_doc_vector(toks, vals) == c["doc_vector"]
