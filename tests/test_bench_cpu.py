"""Host-side pieces of bench.py that need no GPU: the query rotations of the timed steps, and the rule that ncu counters
are only quoted for the exact configuration AND kernel source they were captured on."""
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_traffic_counters_are_bound_to_config_and_source(monkeypatch):
    import bench
    with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
        entry = json.load(f)["sparse_score_topk_kernel"]
    cfg = dict(entry["match"])
    if entry.get("source_sha") != bench.source_sha():
        # legitimate between a kernel change and its recapture: the bench line then carries no counters
        assert bench.ncu_counters("sparse_score_topk_kernel", **cfg) is None
        pytest.skip("profiles/traffic.json was captured on another version of the scorer: bench.py drops the counters")
    got = bench.ncu_counters("sparse_score_topk_kernel", **cfg)
    assert got is not None and got["dram_bytes_per_launch"] == entry["dram_bytes_per_launch"]
    other = dict(cfg, nq=cfg["nq"] + 1)                             # another configuration: nothing quoted
    assert bench.ncu_counters("sparse_score_topk_kernel", **other) is None
    monkeypatch.setattr(bench, "source_sha", lambda: "0" * 16)       # another kernel source: nothing quoted
    assert bench.ncu_counters("sparse_score_topk_kernel", **cfg) is None
    assert bench.ncu_counters("no_such_kernel", **cfg) is None


def test_source_sha_tracks_the_scorer_sources(tmp_path):
    import bench
    a = bench.source_sha()
    assert isinstance(a, str) and len(a) == 16 and a == bench.source_sha()


@pytest.mark.parametrize("off", [0, 1, 7, 13])
def test_rotate_queries_permutes_whole_queries(off):
    """Timed steps cycle through rotations of the query order so that a stale slot of the double-buffered exchange can never
    equal a fresh one: rotation i must hold exactly the old query (i + off) % nq -- terms and weights in their order."""
    import torch
    import bench
    from mllm_sparse_retrieval_b200.synth import QueryCSR
    rng = np.random.default_rng(off)
    nq = 13
    lens = rng.integers(0, 6, size=nq)
    indptr = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int64))
    terms = torch.from_numpy(rng.integers(0, 100, size=int(lens.sum())).astype(np.int32))
    weights = torch.from_numpy(rng.integers(1, 50, size=int(lens.sum())).astype(np.int32))
    qs = QueryCSR(indptr, terms, weights)
    rot, order = bench.rotate_queries(qs, off)
    assert order.tolist() == [(i + off) % nq for i in range(nq)]
    for i in range(nq):
        j = (i + off) % nq
        a0, a1 = int(rot.indptr[i]), int(rot.indptr[i + 1])
        b0, b1 = int(qs.indptr[j]), int(qs.indptr[j + 1])
        assert rot.terms[a0:a1].tolist() == qs.terms[b0:b1].tolist()
        assert rot.weights[a0:a1].tolist() == qs.weights[b0:b1].tolist()
