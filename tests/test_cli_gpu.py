"""GPU tests of the search CLI beyond the single-process file test in test_host_layer_gpu.py: the final-ordering kernel
(msr_order_rows), the string-free query path (--query_logits: encoder output -> msr_quantise_topk -> scorer), and the
reference's launch mode (scripts/search.sh:14: `deepspeed --num_gpus=N src/search.py`) under torchrun in both --dist modes."""
import json
import os
import pickle
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_order_rows_is_pythons_stable_sort(lib_built):
    """msr_order_rows == sorted(row, key=score, reverse=True) (stable: ties keep their position; metrices.py:30) on f32 and
    f64 scores with many ties, with and without dropping the row's own doc id, and the order-keeping removal of
    get_run_dict(remove_query=True)."""
    from mllm_sparse_retrieval_b200.ops import order_rows
    rng = np.random.default_rng(12)
    dev = torch.device("cuda")
    nq, w = 37, 300
    for dt in (np.float32, np.float64):
        scores = rng.integers(0, 40, size=(nq, w)).astype(dt) / 8            # heavy ties
        scores[3, :5] = [np.inf, -np.inf, 0.0, -0.0, 1e-30]
        docs = np.stack([rng.permutation(1000)[:w] for _ in range(nq)]).astype(np.int32)
        counts = rng.integers(0, w + 1, size=nq).astype(np.int32)
        counts[0] = 0; counts[1] = w
        self_ids = np.where(rng.random(nq) < 0.7, docs[np.arange(nq), rng.integers(0, w, nq)], -1).astype(np.int32)
        tags = rng.integers(1, 4, size=(nq, w)).astype(np.uint8)
        T = lambda a: torch.from_numpy(a).to(dev)
        for by_score in (True, False):
            for me in (None, self_ids):
                s, d, c, t = order_rows(T(scores), T(docs), T(counts), self_ids=None if me is None else T(me), by_score=by_score,
                                        tags=T(tags))
                s, d, c, t = s.cpu().numpy(), d.cpu().numpy(), c.cpu().numpy(), t.cpu().numpy()
                for q in range(nq):
                    rows = [(scores[q, i], docs[q, i], tags[q, i]) for i in range(counts[q]) if me is None or me[q] < 0 or docs[q, i] != me[q]]
                    if by_score:
                        rows = sorted(rows, key=lambda r: r[0], reverse=True)
                    assert c[q] == len(rows)
                    assert d[q, :c[q]].tolist() == [r[1] for r in rows]
                    assert s[q, :c[q]].tolist() == [r[0] for r in rows]
                    assert t[q, :c[q]].tolist() == [r[2] for r in rows]


def _write_problem(tmp_path, seed=8, n_docs=600, nq=48, dim=64, V=300):
    """The files src/encode.py writes (corpus_0.jsonl / corpus_0.pkl / query.tsv / query.pkl) + encoder-output logits whose
    quantisation IS query.tsv (oracle.quantise_row: top-k, rint(x*100), vocabulary strings, token repeated v times)."""
    from oracle import oracle
    rng = np.random.default_rng(seed)
    vocab = [f"W{j}" if j % 7 == 0 else f"w{j}" for j in range(V)]             # some upper-case entries: lower() collides
    vocab[5] = "w6"                                                            # two ids, one term: duplicates must add (Q3)
    doc_ids = [str(x) for x in rng.permutation(2000)[:n_docs]]
    vecs = [{vocab[j].lower(): int(rng.integers(1, 200)) for j in rng.choice(V, 16, replace=False)} for _ in range(n_docs)]
    sp = tmp_path / "sparse"; sp.mkdir()
    with open(sp / "corpus_0.jsonl", "w") as f:
        for i, v in zip(doc_ids, vecs):
            f.write(json.dumps({"id": i, "content": "", "vector": v}) + "\n")
    reps = rng.standard_normal((n_docs, dim)).astype(np.float32); reps /= np.linalg.norm(reps, axis=1, keepdims=True)
    dn = tmp_path / "dense"; dn.mkdir()
    with open(dn / "corpus_0.pkl", "wb") as f:
        pickle.dump((reps, doc_ids), f)
    qids = [doc_ids[i] if i % 3 == 0 else f"q{i}" for i in range(nq)]
    q_reps = rng.standard_normal((nq, dim)).astype(np.float32); q_reps /= np.linalg.norm(q_reps, axis=1, keepdims=True)
    with open(dn / "query.pkl", "wb") as f:
        pickle.dump((q_reps, qids), f)
    logits = np.log1p(np.maximum(rng.standard_normal((nq, V)).astype(np.float32) * 1.5, 0)).astype(np.float32)
    top = 24
    with open(sp / "query.tsv", "w") as f:
        for q, row in zip(qids, logits):
            counts = oracle.quantise_row(row, top, vocab)                       # {token: weight}, duplicates summed
            f.write(q + "\t" + " ".join(" ".join([tok] * int(v)) for tok, v in counts.items()) + "\n")
    torch.save({"ids": qids, "logits": torch.from_numpy(logits)}, str(sp / "query_logits.pt"))
    with open(sp / "vocab.json", "w") as f:
        json.dump(vocab, f)
    qrels = tmp_path / "qrels.txt"
    with open(qrels, "w") as f:
        for q in qids:
            f.write(q + "\t" + " ".join(doc_ids[int(j)] for j in rng.choice(n_docs, 3, replace=False)) + "\n")
    return sp, dn, qrels, top


def _args(sp, dn, qrels, out, extra=()):
    return ["--passage_reps", str(dn), "--sparse_index", str(sp), "--query_reps", str(dn), "--query_sparse", str(sp), "--depth", "40",
            "--alpha", "0.4", "--remove_query", "--save_dir", str(out), "--batch_size", "16", "--qrels", str(qrels)] + list(extra)


def test_cli_logits_path_equals_string_path(oracle, lib_built, tmp_path):
    """--query_logits (+ --vocab): activation rows -> msr_quantise_topk -> scorer, no query string in between; the run files
    are byte-identical to those of the query.tsv path built from the same encoder output."""
    from mllm_sparse_retrieval_b200 import search as cli
    sp, dn, qrels, top = _write_problem(tmp_path)
    cli.main(_args(sp, dn, qrels, tmp_path / "runs_tsv"))
    a = _args(sp, dn, qrels, tmp_path / "runs_logits", ["--query_logits", str(sp / "query_logits.pt"), "--vocab", str(sp / "vocab.json"),
                                                        "--sparse_length", str(top)])
    a[a.index("--query_sparse"):a.index("--query_sparse") + 2] = []
    cli.main(a)
    for name in ("run.fusion.trec", "run.sparse.trec", "run.dense.trec"):
        assert (tmp_path / "runs_tsv" / name).read_bytes() == (tmp_path / "runs_logits" / name).read_bytes(), name
    assert len((tmp_path / "runs_tsv" / "run.sparse.trec").read_text().splitlines()) > 48 * 10


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("dist_mode", ["replicas", "shard"])
def test_cli_under_torchrun_matches_single_process(oracle, lib_built, tmp_path, dist_mode):
    """`torchrun --nproc-per-node 2 -m mllm_sparse_retrieval_b200.search --dist replicas|shard` (the reference's own launch
    mode, scripts/search.sh:14 / search.py:112-131,180): run files byte for byte those of the single-process run, Recall@k
    lines identical (hit counters summed over the ranks with one all_reduce)."""
    sp, dn, qrels, top = _write_problem(tmp_path, nq=50)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    single = subprocess.run([sys.executable, "-m", "mllm_sparse_retrieval_b200.search"] + _args(sp, dn, qrels, tmp_path / "one"),
                            capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)
    assert single.returncode == 0, single.stderr[-2000:]
    multi = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                            "127.0.0.1", "--master-port", "29671" if dist_mode == "shard" else "29672", "-m",
                            "mllm_sparse_retrieval_b200.search"] + _args(sp, dn, qrels, tmp_path / "two", ["--dist", dist_mode]),
                           capture_output=True, text=True, env=env, cwd=ROOT, timeout=900)
    assert multi.returncode == 0, multi.stderr[-3000:]
    for name in ("run.fusion.trec", "run.sparse.trec", "run.dense.trec"):
        assert (tmp_path / "one" / name).read_bytes() == (tmp_path / "two" / name).read_bytes(), name
    rec = lambda out: [ln for ln in out.splitlines() if ln.startswith("Recall@") or ln.startswith("[")]
    assert rec(single.stdout) == rec(multi.stdout) and len(rec(single.stdout)) == 6
