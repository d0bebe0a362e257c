"""Generates tests/golden/*.json by IMPORTING the reference's own in-tree functions from /root/reference/src
(read-only; nothing is copied).  Run in the build container only -- /root/reference does not exist on the GPU box.

    python tests/golden/make_golden.py

Pinned functions: hybrid.fuse / hybrid.fuse_statistic / hybrid.write_trec_run / hybrid.read_trec_run (hybrid.py),
search.get_run_dict / search.sparse_search / search.search_queries (search.py:55-99) and
encode.get_img_valid_tokens_values / get_text_valid_tokens_values (encode.py:56-153); with ``--metrics`` RecallMetrics.sort_and_count
(metrices.py:29-84) -> reference_metrics.json.  Third-party modules that are absent here (faiss, tevatron,
pyserini, nltk, peft) are stubbed in sys.modules exactly as SURVEY.md Appendix B describes; none of the pinned
functions executes stub code except through the fake searcher objects handed to them below.
"""
import importlib.machinery
import json
import os
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/src"


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__spec__ = importlib.machinery.ModuleSpec(name, None)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def import_reference():
    import transformers  # noqa: F401  (must resolve before peft is stubbed)
    from transformers import TrainingArguments, Trainer  # noqa: F401
    from dataclasses import dataclass, field

    @dataclass
    class ModelArguments:
        model_name_or_path: str = field(default=None)
        lora: bool = field(default=False)
        lora_r: int = field(default=8)
        lora_alpha: int = field(default=64)
        lora_dropout: float = field(default=0.1)

    @dataclass
    class DataArguments:
        dataset_name: str = field(default=None)
        dataset_split: str = field(default="test")
        dataset_shard_index: int = field(default=0)
        encode_is_query: bool = field(default=False)

    class _Any:
        def __init__(self, *a, **k):
            pass

    _stub("faiss")
    tv = _stub("tevatron"); tr = _stub("tevatron.retriever"); tv.retriever = tr
    tr.arguments = _stub("tevatron.retriever.arguments", ModelArguments=ModelArguments, DataArguments=DataArguments)
    tr.searcher = _stub("tevatron.retriever.searcher", FaissFlatSearcher=_Any)
    tr.collator = _stub("tevatron.retriever.collator", TrainCollator=_Any)
    ps = _stub("pyserini"); ps.search = _stub("pyserini.search")
    ps.search.lucene = _stub("pyserini.search.lucene", LuceneImpactSearcher=_Any, LuceneSearcher=_Any)
    ps.analysis = _stub("pyserini.analysis", JWhiteSpaceAnalyzer=_Any)
    corpus = _stub("nltk.corpus", stopwords=types.SimpleNamespace(words=lambda lang: ["a", "the", "of"]))
    _stub("nltk", word_tokenize=lambda s: s.split(), corpus=corpus)
    _stub("peft", PeftModel=_Any, PeftConfig=_Any, LoraConfig=_Any, TaskType=_Any, get_peft_model=_Any)
    sys.path.insert(0, REF)
    import hybrid, search, encode  # noqa: E401
    return hybrid, search, encode


def typed(v):
    """JSON-able exact value + type tag."""
    if isinstance(v, np.float32):
        return ["f32", float(v)]
    if isinstance(v, (float, np.float64)):
        return ["f64", float(v)]
    return ["int", int(v)]


def make_fuse_cases(hybrid, search, rng):
    cases = []
    for case_id, (nq, kd, ks, n_docs, remove_query, alpha) in enumerate([
            (4, 10, 10, 40, False, 0.5), (5, 50, 37, 120, True, 0.3), (3, 100, 100, 150, True, 0.5),
            (2, 5, 8, 12, True, 0.7), (3, 1000, 1000, 3000, True, 0.5)]):
        qids = [str(3 * i + 1) for i in range(nq)]
        dense_lists, sparse_lists = [], []
        for q, qid in enumerate(qids):
            dd = rng.permutation(n_docs)[:kd]
            sd = np.concatenate([dd[: ks // 2], rng.permutation(np.setdiff1d(np.arange(n_docs), dd))[: ks - ks // 2]])
            if q % 2 == 0:
                dd[rng.integers(0, kd)] = int(qid)
                _, ui = np.unique(dd, return_index=True); dd = dd[np.sort(ui)]
            if q % 3 == 0 and int(qid) not in sd:
                sd[0] = int(qid)                                  # the removed doc carries the sparse MAX (H2)
            ds = np.sort(rng.random(len(dd)).astype(np.float32) * np.float32(0.6) + np.float32(0.2))[::-1]
            ssc = np.sort(rng.integers(1, 60000, len(sd)))[::-1].astype(np.float32)
            if case_id == 3 and q == 1:
                ds[:] = ds[0]                                     # zero dense range -> 1e-9 denominator
            dense_lists.append((ds, [str(x) for x in dd]))
            sparse_lists.append(([float(x) for x in ssc], [str(x) for x in sd]))
        dense_run = search.get_run_dict(qids, [d[0] for d in dense_lists], [d[1] for d in dense_lists], remove_query)
        sparse_run = search.get_run_dict(qids, [s[0] for s in sparse_lists], [s[1] for s in sparse_lists], remove_query)
        fused = hybrid.fuse([dense_run, sparse_run], [alpha, 1 - alpha])
        stat = hybrid.fuse_statistic([dense_run, sparse_run], [alpha, 1 - alpha])
        cases.append({
            "qids": qids, "remove_query": remove_query, "alpha": alpha,
            "dense": [{"scores": [float(x) for x in d[0]], "docs": d[1]} for d in dense_lists],
            "sparse": [{"scores": s[0], "docs": s[1]} for s in sparse_lists],
            "dense_run_minmax": {q: [typed(dense_run[q]["min_score"]), typed(dense_run[q]["max_score"])] for q in qids},
            "sparse_run_minmax": {q: [typed(sparse_run[q]["min_score"]), typed(sparse_run[q]["max_score"])] for q in qids},
            "dense_run_docs": {q: list(dense_run[q]["docs"].keys()) for q in qids},
            "fused": {q: [[doc, typed(v)] for doc, v in fused[q].items()] for q in qids},
            "fused_tags": {q: [[doc, rec.type] for doc, rec in stat[q].items()] for q in qids},
        })
    return cases


def make_trec_case(hybrid):
    run = {"7": {"docs": {"12": 3.5, "100": 9.25, "99": 3.5, "5": 0.125}, "min_score": 0.125, "max_score": 9.25},
           "8": {"30": np.float32(0.75), "2": 1.5}}
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "run.trec")
        hybrid.write_trec_run(run, path, name="fusion")
        text = open(path).read()
        back = hybrid.read_trec_run(path)
    return {"run": {"7": run["7"], "8": {"30": 0.75, "2": 1.5}}, "text": text, "read_back": back}


def make_glue_case(search):
    class FakeHit:
        def __init__(self, docid, score):
            self.docid, self.score = docid, score

    class FakeSparse:
        def batch_search(self, topics, ids, depth, threads=1):
            return {i: [FakeHit(str(10 * n + j), float(100 - j)) for j in range(min(depth, 3))] for n, i in enumerate(ids)}

    class FakeDense:
        def batch_search(self, q, depth, batch_size, quiet):
            n = q.shape[0]
            return np.tile(np.arange(depth, 0, -1, dtype=np.float32), (n, 1)), np.tile(np.arange(depth), (n, 1))

        def search(self, q, depth):
            return self.batch_search(q, depth, 0, True)

    args = types.SimpleNamespace(depth=4, threads=2, batch_size=2, quiet=True)
    ss, sr = search.sparse_search(FakeSparse(), ["a b", "c"], ["q1", "q0"], args)
    ds, dr = search.search_queries(FakeDense(), np.zeros((2, 8), np.float32), ["d%d" % i for i in range(10)], args)
    rd = search.get_run_dict(["5", "x"], [[30., 20.], []], [["7", "5"], []], True)
    return {"sparse_scores": ss, "sparse_rankings": sr, "dense_scores": ds.tolist(), "dense_rankings": dr.tolist(),
            "run_dict_remove": rd}


def make_quantiser_cases(encode, rng):
    import torch
    out = []
    vocab = ["Dog", "dog", "Ġcat", "zebra", "▁Tree", "tree", "sky", "Sky", "x1", "##ing", "blue", "Blue"] + [f"tok{i}" for i in range(52)]
    vocab_dict = {i: t for i, t in enumerate(vocab)}
    for sparse_manual, sparse_length, is_filtered in [(False, 128, False), (True, 8, False), (True, 16, True)]:
        V = len(vocab) if not sparse_manual else len(vocab) + 6          # ids >= len(vocab) exercise the `i < len` filter
        if not sparse_manual:
            vocab_full = {i: f"w{i}" for i in range(200)}
            vocab_full.update(vocab_dict)
            vd, V = vocab_full, 200
        else:
            vd = vocab_dict
        logits = torch.tensor(np.log1p(np.maximum(rng.standard_normal(V) * 2.0, 0)).astype(np.float32))
        logits[:4] = torch.tensor([1.23, 1.0, 0.565, 0.3])
        logits[4] = 0.125; logits[5] = 0.375                             # *100 -> 12.5 / 37.5: round-half-even
        data_args = types.SimpleNamespace(sparse_manual=sparse_manual, sparse_length=sparse_length, is_filtered=is_filtered,
                                          num_expended_tokens=0)
        tokens, values = encode.get_img_valid_tokens_values(None, logits, vd, data_args, set())
        query = ""
        for token, v in zip(tokens, values):
            query += (" " + token) * v
        vector = {}
        for token, v in zip(tokens, values):
            vector[token] = int(v)
        out.append({"sparse_manual": sparse_manual, "sparse_length": sparse_length, "is_filtered": is_filtered,
                    "vocab": [vd[i] for i in range(len(vd))], "logits": [float(x) for x in logits.numpy()],
                    "tokens": tokens, "values": [int(v) for v in values], "query_string": query.strip(), "doc_vector": vector})
    return out



class FakeTokenizer:
    """word -> 0..3 token ids, deterministic; stands in for the HF tokenizer handed to get_text_valid_tokens_values."""
    def __init__(self, V):
        self.V = V

    def encode(self, word, add_special_tokens=False):
        h = sum((i + 1) * ord(c) for i, c in enumerate(word))
        if word.startswith("oov"):
            return []
        n = 1 + h % 3
        return [(h * 7 + j * 31) % self.V for j in range(n)]


def make_text_quantiser_cases(encode):
    """encode.get_text_valid_tokens_values (encode.py:96-153) with the stubbed nltk (word_tokenize = str.split,
    stop-words = a/the/of) and FakeTokenizer; logits strictly positive and distinct so torch.topk has no ties."""
    import string
    import torch
    rng = np.random.default_rng(20261106)
    out = []
    base = ["Dog", "dog", "Ġcat", "zebra", "▁Tree", "tree", "sky", "Sky", "x1", "##ing", "blue", "Blue"]
    many = " ".join(f"w{i}" for i in range(140))
    for text, sparse_manual, sparse_length, is_filtered, V, vlen, n_exp in [
            ("A dog of the blue Sky , runs !", False, 128, False, 200, 200, 0),
            ("the of a , .", False, 128, False, 200, 200, 0),                   # no content word -> top-10 fallback
            ("oovword the", False, 128, True, 200, 200, 0),                    # words that tokenise to nothing
            ("two dogs chase a zebra", True, 8, False, 200, 200, 0),           # --sparse_manual
            ("Trees and sky : tree Tree", False, 128, True, 200, 200, 0),      # is_filtered
            (many, False, 128, False, 400, 400, 0),                            # > 128 candidates
            ("a quick brown fox", False, 128, False, 200, 190, 0),             # ids >= len(vocab_dict): value-shift quirk
            ("the", False, 128, False, 200, 195, 0),
            # expansion terms (num_expended_tokens > 0, encode.py:137-152); filtered_ids = every third id below 180
            ("A dog of the blue Sky , runs !", False, 128, False, 200, 200, 5),
            ("the of a , .", False, 128, False, 200, 200, 7),                   # fallback top-10 + expansion (overlap sums)
            ("a quick brown fox", False, 128, False, 200, 190, 4),             # value shift carries into the expansion
            ("Trees and sky : tree Tree", False, 128, True, 200, 200, 300)]:   # more requested than eligible
        vd = {i: (base[i] if i < len(base) else f"tok{i}") for i in range(vlen)}
        logits = torch.tensor((np.log1p(np.abs(rng.standard_normal(V)) * 2.0) + 1e-3).astype(np.float32))
        assert len(np.unique(logits.numpy())) == V
        tok = FakeTokenizer(V)
        data_args = types.SimpleNamespace(sparse_manual=sparse_manual, sparse_length=sparse_length, is_filtered=is_filtered,
                                          num_expended_tokens=n_exp)
        filtered_ids = set(range(0, 180, 3)) if n_exp else set()
        tokens, values = encode.get_text_valid_tokens_values(text, tok, logits, vd, data_args, set(filtered_ids))
        # the candidate list the reference built internally, rebuilt the same way (same set -> same iteration order)
        words = [i for i in text.lower().split() if i not in set(["a", "the", "of"] + list(string.punctuation))]
        token_ids = set()
        for w in words:
            token_ids.update(tok.encode(w, add_special_tokens=False))
        query = ""
        for token, v in zip(tokens, values):
            query += (" " + token) * v
        vector = {}
        for token, v in zip(tokens, values):
            vector[token] = int(v)
        out.append({"text": text, "sparse_manual": sparse_manual, "sparse_length": sparse_length, "is_filtered": is_filtered,
                    "vocab": [vd[i] for i in range(vlen)], "logits": [float(x) for x in logits.numpy()],
                    "token_ids": [int(t) for t in token_ids], "stop_words": ["a", "the", "of"],
                    "num_expended_tokens": n_exp, "filtered_ids": sorted(filtered_ids),
                    "tokens": tokens, "values": [int(v) for v in values], "query_string": query.strip(), "doc_vector": vector})
    return out


def make_metrics_cases():
    """RecallMetrics.sort_and_count (src/metrices.py:29-84) on random runs.  The class calls ``.cuda()`` on tiny tensors
    and ``dist.get_world_size()`` in its constructor; both are patched to CPU / 1 here, nothing else is touched."""
    import torch
    import torch.distributed as dist
    import metrices
    torch.Tensor.cuda = lambda self, *a, **k: self
    dist.get_world_size = lambda *a, **k: 1
    rng = np.random.default_rng(20261105)
    cases = []
    for nq, n_docs, depth, multi in [(40, 300, 250, False), (25, 60, 40, True), (30, 1000, 12, True), (8, 50, 0, False)]:
        targets, runs = {}, {}
        for name in ("dense", "sparse", "fusion"):
            run = {}
            for q in range(nq):
                n = 0 if depth == 0 or (q % 11 == 7) else int(rng.integers(1, depth + 1))
                docs = rng.choice(n_docs, size=min(n, n_docs), replace=False)
                # scores with ties: the stable sort keeps insertion order among equals
                sc = np.round(rng.random(len(docs)) * 8) / 8
                d = {str(int(x)): float(v) for x, v in zip(docs, sc)}
                run[str(q)] = d if name == "fusion" else {"docs": d, "min_score": 0.0, "max_score": 1.0}
            runs[name] = run
        for q in range(nq):
            tg = [str(int(x)) for x in rng.choice(n_docs, size=5 if multi else 1, replace=False)]
            targets[str(q)] = tg if multi else tg[0]

        class _DS:
            def get_target(self, qid, query_type):
                return targets[qid]
        rm = metrices.RecallMetrics(_DS(), runs["dense"], runs["sparse"], runs["fusion"], [], list(range(nq)),
                                    types.SimpleNamespace(query_type="text"))
        rm.sort_and_count()
        cases.append({"nq": nq, "n_docs": n_docs, "targets": targets, "runs": runs,
                      "counts": {"dense": rm.dense_counts, "sparse": rm.sparse_counts, "fusion": rm.fusion_counts}})
    return cases


def main():
    hybrid, search, encode = import_reference()
    if "--metrics" in sys.argv:
        with open(os.path.join(HERE, "reference_metrics.json"), "w") as f:
            json.dump({"metrics": make_metrics_cases()}, f)
        print("wrote reference_metrics.json", os.path.getsize(os.path.join(HERE, "reference_metrics.json")), "bytes")
        return
    rng = np.random.default_rng(20261104)
    golden = {
        "numpy": np.__version__,
        "fuse": make_fuse_cases(hybrid, search, rng),
        "trec": make_trec_case(hybrid),
        "glue": make_glue_case(search),
        "quantiser": make_quantiser_cases(encode, rng),
        "text_quantiser": make_text_quantiser_cases(encode),
    }
    with open(os.path.join(HERE, "reference_glue.json"), "w") as f:
        json.dump(golden, f)
    print("wrote", os.path.join(HERE, "reference_glue.json"), os.path.getsize(os.path.join(HERE, "reference_glue.json")), "bytes")


if __name__ == "__main__":
    main()
