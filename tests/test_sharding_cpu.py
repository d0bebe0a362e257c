"""CPU tests of the doc-sharded multi-process path with world_size-2 gloo.

The product's per-shard search runs on the GPU; what is exercised here is everything around it that must agree
between ranks: ImpactCollection.shard (doc ranges, rebased ids, GLOBAL df/N for rule S5), the all-gather layout
[world, nq, k] that msr_topk_merge consumes, and the merge order.  The per-shard scorer is the oracle."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _make_collection(seed=11, n_docs=700, n_terms=120, nnz=14):
    from mllm_sparse_retrieval_b200.index import ImpactCollection
    rng = np.random.default_rng(seed)
    ids = [str(x) for x in rng.permutation(n_docs * 3)[:n_docs]]          # string order != numeric order
    vecs = []
    for _ in range(n_docs):
        t = rng.choice(n_terms, size=nnz, replace=False)
        vecs.append({f"t{j}": int(rng.integers(1, 300)) for j in t})
    vecs[0]["everywhere"] = 5
    for v in vecs[1:]:
        v["everywhere"] = 2                                               # df = N: dropped by S5 on every shard
    return ImpactCollection.from_vectors(ids, vecs)


def _queries(coll, seed=12, nq=40):
    rng = np.random.default_rng(seed)
    qs = []
    for _ in range(nq):
        toks = [coll.terms[i] for i in rng.choice(coll.n_terms, size=6, replace=False)]
        qs.append(" ".join(t for t in toks for _ in range(int(rng.integers(1, 4)))) + " everywhere")
    return qs


def _worker(rank, world, port, k, out_path):
    sys.path.insert(0, ROOT)
    from oracle import oracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        coll = _make_collection()
        qi, qt, qw = coll.encode_queries(_queries(coll))
        lo, hi = coll.n_docs * rank // world, coll.n_docs * (rank + 1) // world
        indptr, docs, imps = coll.shard(lo, hi)
        s, d, c = oracle.impact_search_c(indptr, docs, imps, hi - lo, qi, qt, qw, k, n_docs_global=coll.n_docs,
                                         df_global=coll.df, doc_base=lo)
        nq = len(qi) - 1
        # the same memory layout the NCCL path hands to msr_topk_merge: [world, nq, k] (gloo wants it flat)
        g_s = torch.empty((world * nq, k), dtype=torch.float32)
        g_d = torch.empty((world * nq, k), dtype=torch.int32)
        g_c = torch.empty((world * nq,), dtype=torch.int32)
        dist.all_gather_into_tensor(g_s, torch.from_numpy(s))
        dist.all_gather_into_tensor(g_d, torch.from_numpy(d))
        dist.all_gather_into_tensor(g_c, torch.from_numpy(c.astype(np.int32)))
        g_s, g_d, g_c = g_s.view(world, nq, k), g_d.view(world, nq, k), g_c.view(world, nq)
        ms, md, mc = oracle.merge_topk(list(g_s.numpy()), list(g_d.numpy()), list(g_c.numpy()), k)
        if rank == 0:
            np.savez(out_path, s=ms, d=md, c=mc)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("k", [10, 1000])
def test_two_rank_gloo_matches_unsharded(oracle, tmp_path, k):
    out = str(tmp_path / "merged.npz")
    mp.spawn(_worker, args=(2, _free_port(), k, out), nprocs=2, join=True)
    got = np.load(out)
    coll = _make_collection()
    qi, qt, qw = coll.encode_queries(_queries(coll))
    es, ed, ec = oracle.impact_search_c(coll.indptr, coll.post_docs, coll.post_impacts, coll.n_docs, qi, qt, qw, k)
    np.testing.assert_array_equal(got["c"], ec)
    np.testing.assert_array_equal(got["d"], ed)
    np.testing.assert_array_equal(got["s"], es)
    assert ec.max() <= coll.n_docs


def test_shard_keeps_global_term_filter():
    """S5 is decided on collection-wide df/N: a term present in every doc is dropped on every shard even though a
    shard's own df would keep it (df_shard + 1 < N)."""
    coll = _make_collection()
    t = coll.term_to_id["everywhere"]
    assert int(coll.df[t]) == coll.n_docs
    indptr, docs, imps = coll.shard(0, coll.n_docs // 2)
    assert int(indptr[t + 1] - indptr[t]) == coll.n_docs // 2          # present in the shard ...
    assert docs.max() < coll.n_docs // 2 and len(docs) == int(indptr[-1])
    from oracle import oracle
    qi, qt, qw = coll.encode_queries(["everywhere"])
    s, d, c = oracle.impact_search_c(indptr, docs, imps, coll.n_docs // 2, qi, qt, qw, 5, n_docs_global=coll.n_docs,
                                     df_global=coll.df, doc_base=0)
    assert int(c[0]) == 0                                                # ... but filtered by the global df


def test_peer_exchange_slot_layout_tiles_the_buffer():
    """Host arithmetic of the fused exchange (no GPU): the (step buffer, rank) slots with their scores / docs / counts
    sections, the three rotating threshold arrays and the flag words are disjoint, 16-byte aligned (the kernels store rows
    as uint4) and inside the symmetric buffer; slot strides are what msr_topk_merge_exchange is given."""
    from mllm_sparse_retrieval_b200.sharded import PeerExchange, peer_layout
    for nq, k, world in ((50, 7, 4), (1, 1, 2), (64, 100, 8), (10000, 100, 8), (33, 12, 3)):
        px = object.__new__(PeerExchange)                               # layout only: no symmetric memory on this box
        px.nq, px.k, px.world = nq, k, world
        px.slot, px.theta_off, px.flag_off, total = peer_layout(nq, k, world)
        base = 4096
        regions = []
        for par in (0, 1):
            for who in range(world):
                s, d, c = px._slot_ptrs(base, par, who)
                assert s % 16 == 0 and d % 16 == 0 and c % 16 == 0
                assert d - s >= 4 * nq * k and c - d >= 4 * nq * k
                regions += [(s, s + 4 * nq * k), (d, d + 4 * nq * k), (c, c + 4 * nq)]
                if who:
                    assert s - px._slot_ptrs(base, par, who - 1)[0] == 4 * px.slot      # the list stride of the merge
        for t in range(3):
            th = px._theta_ptr(base, t)
            assert th % 16 == 0
            regions.append((th, th + 8 * nq * (1 + world)))
        regions.append((base + 4 * px.flag_off, base + 4 * px.flag_off + 4 * world))
        regions.sort()
        assert regions[0][0] == base and regions[-1][1] <= base + 4 * total
        for (a0, a1), (b0, b1) in zip(regions[:-1], regions[1:]):
            assert a1 <= b0                                             # no overlap
