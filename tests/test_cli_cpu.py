"""Host-only pieces of the search CLI (no GPU): argument conventions of the reference's HfArgumentParser dataclass
(src/arguments.py:56-68), the files src/encode.py writes (query.tsv, query logits + vocabulary), and the run-file writer of the
array path against hybrid.write_trec_run (src/hybrid.py:20-29)."""
import json

import numpy as np
import torch


def test_parse_args_follows_the_reference_dataclass():
    from mllm_sparse_retrieval_b200 import search as cli
    sa, ns = cli.parse_args([])
    # defaults of PromptRepsLLMSearchArguments (arguments.py:56-68)
    assert (sa.depth, sa.alpha, sa.batch_size, sa.threads, sa.query_type) == (1000, 0.5, 128, 1, "text")
    assert sa.quiet is True and sa.use_gpu is False and sa.remove_query is False and sa.save_dir is None
    assert ns.dist == "auto" and ns.sparse_length == 128 and ns.is_filtered is False
    # HfArgumentParser booleans: bare flag, or flag + value; encoder-side arguments are accepted and ignored
    sa, ns = cli.parse_args(["--remove_query", "--quiet", "false", "--depth", "10", "--alpha", "0.3", "--model_name_or_path", "x",
                             "--bf16", "--per_device_batch_size", "2", "--dist", "shard", "--is_filtered", "true"])
    assert sa.remove_query is True and sa.quiet is False and sa.depth == 10 and sa.alpha == 0.3
    assert ns.dist == "shard" and ns.is_filtered is True
    sa, _ = cli.parse_args(["--remove_query", "no", "--use_gpu", "1"])
    assert sa.remove_query is False and sa.use_gpu is True


def test_read_query_tsv_keeps_the_text_verbatim(tmp_path):
    from mllm_sparse_retrieval_b200 import search as cli
    p = tmp_path / "query.tsv"
    p.write_text("q1\tdog dog cat\n\n17\t\nq3\ta\tb c\n", encoding="utf-8")
    ids, texts = cli.read_query_tsv(str(p))
    assert ids == ["q1", "17", "q3"] and texts == ["dog dog cat", "", "a\tb c"]      # only the FIRST tab separates id and text
    assert cli._resolve(str(tmp_path), "query.tsv") == str(p) and cli._resolve(str(p), "query.tsv") == str(p)


def test_load_query_logits_accepts_both_layouts(tmp_path):
    from mllm_sparse_retrieval_b200 import search as cli
    logits = torch.arange(12, dtype=torch.float16).reshape(3, 4)
    torch.save({"ids": [5, "b", 7], "logits": logits}, str(tmp_path / "a.pt"))
    torch.save((logits, ["x", "y", "z"]), str(tmp_path / "b.pt"))
    (tmp_path / "list.json").write_text(json.dumps(["A", "b", "c", "d"]))
    (tmp_path / "map.json").write_text(json.dumps({"A": 0, "d": 3, "b": 1}))        # id 2 missing: empty string
    ids, lg, vocab = cli.load_query_logits(str(tmp_path / "a.pt"), str(tmp_path / "list.json"))
    assert ids == ["5", "b", "7"] and torch.equal(lg, logits) and vocab == ["A", "b", "c", "d"]
    ids, lg, vocab = cli.load_query_logits(str(tmp_path / "b.pt"), str(tmp_path / "map.json"))
    assert ids == ["x", "y", "z"] and vocab == ["A", "b", "", "d"]


def test_write_trec_arrays_equals_write_trec_run(tmp_path):
    """The array-form writer emits what hybrid.write_trec_run emits for the same run (rows already in rank order)."""
    from mllm_sparse_retrieval_b200 import hybrid, search as cli
    names = ["100", "99", "7", "abc"]
    qids = ["q1", "q2", "q3"]
    scores = np.array([[3.5, 2.25, 1.0], [9.0, 0.5, 0.0], [0.0, 0.0, 0.0]], dtype=np.float32)
    docs = np.array([[2, 0, 3], [1, 2, -1], [-1, -1, -1]], dtype=np.int32)
    counts = np.array([3, 2, 0], dtype=np.int32)
    cli.write_trec_arrays(str(tmp_path / "a.trec"), qids, names, scores, docs, counts, "fusion")
    run = {q: {names[docs[r, j]]: scores[r, j] for j in range(counts[r])} for r, q in enumerate(qids)}
    hybrid.write_trec_run(run, str(tmp_path / "b.trec"), name="fusion")
    assert (tmp_path / "a.trec").read_text() == (tmp_path / "b.trec").read_text()
    back = hybrid.read_trec_run(str(tmp_path / "a.trec"))
    assert list(back["q1"]["docs"]) == ["7", "100", "abc"] and back["q1"]["max_score"] == 3.5 and back["q1"]["min_score"] == 1.0
    assert "q3" not in back
