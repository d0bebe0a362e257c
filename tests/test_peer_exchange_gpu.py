"""Two-GPU test of the fused exchange: PeerExchange (peer stores from the tile-merge epilogue into
symmetric memory + signal-pad barrier + msr_topk_merge_strided) must equal the NCCL all-gather exchange and the unsharded
oracle, bit for bit, over several steps (the two step buffers alternate).  Skipped on a one-GPU box."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import random_csr, random_queries

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _problem():
    rng = np.random.default_rng(77)
    n_docs, n_terms = 9000, 1500
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, 48)
    qs = [random_queries(rng, 64 if i % 3 else 40, n_terms, 16) for i in range(8)]   # a different batch per step (some
    # smaller than the exchange's capacity): 8 steps walk the 2 step buffers and the 3 threshold arrays several times
    return n_docs, n_terms, indptr, docids, impacts, qs


def _worker(rank, world, port, k, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from mllm_sparse_retrieval_b200.ops import SparseIndex
        from mllm_sparse_retrieval_b200.sharded import PeerExchange, PipelinedExchange
        n_docs, n_terms, indptr, docids, impacts, qs = _problem()
        df = np.diff(indptr.astype(np.int64)).astype(np.uint32)
        lo, hi = n_docs * rank // world, n_docs * (rank + 1) // world
        m = (docids >= lo) & (docids < hi)
        ip = np.zeros(n_terms + 1, dtype=np.uint64)
        np.cumsum(np.bincount(np.repeat(np.arange(n_terms), df)[m], minlength=n_terms), out=ip[1:])
        idx = SparseIndex(ip, docids[m] - lo, impacts[m], hi - lo, n_docs_global=n_docs, df_global=df, doc_base=lo,
                          tile_docs=700, device=rank)
        nq = 64
        px = PeerExchange(nq, k, dev, world, rank)
        ex = PipelinedExchange(nq, k, dev, world)
        res = {}
        for step, (qi, qt, qw) in enumerate(qs):
            dq = (torch.from_numpy(qi.astype(np.int64)).to(dev), torch.from_numpy(qt.astype(np.int32)).to(dev),
                  torch.from_numpy(qw.astype(np.int32)).to(dev))
            n_step = dq[0].numel() - 1
            a = [t.clone() for t in px.run(idx, lambda o: idx.search(*dq, k, out=o), nq=n_step)]
            b = [t[:n_step].clone() for t in ex.run(lambda i, l, h, o: idx.search(*dq, k, out=tuple(x[:n_step] for x in o)))]
            torch.cuda.synchronize()
            for x, y in zip(a, b):
                assert torch.equal(x, y), f"rank {rank} step {step}: p2p exchange != NCCL exchange"
            res[f"s{step}"], res[f"d{step}"], res[f"c{step}"] = (t.cpu().numpy() for t in a)
        np.savez(out_path + f".{rank}.npz", **res)
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs with peer access")
@pytest.mark.parametrize("k", [100, 10])
def test_peer_exchange_two_gpus(oracle, lib_built, tmp_path, k):
    out = str(tmp_path / "px")
    mp.spawn(_worker, args=(2, _free_port(), k, out), nprocs=2, join=True)
    n_docs, n_terms, indptr, docids, impacts, qs = _problem()
    for r in range(2):                                                    # every rank ends with the full result
        got = np.load(out + f".{r}.npz")
        for step, (qi, qt, qw) in enumerate(qs):
            es, ed, ec = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw, k, threads=4)
            np.testing.assert_array_equal(got[f"c{step}"].astype(np.uint32), ec)
            np.testing.assert_array_equal(got[f"d{step}"], ed)
            np.testing.assert_array_equal(got[f"s{step}"], es)


def _searcher_worker(rank, world, port, k, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from test_sharding_cpu import _make_collection, _queries
        from mllm_sparse_retrieval_b200.searchers import LuceneImpactSearcher
        coll = _make_collection()
        qs = _queries(coll, nq=90)
        res = {}
        for kind in ("p2p", "nccl"):
            os.environ["MSR_EXCHANGE"] = kind
            S = LuceneImpactSearcher(None, collection=coll, shard=True, device=dev)
            for n in (40, 25, 90):                                        # a smaller batch reuses the slots, a larger one regrows them
                s, d, c = S.search_arrays(qs[:n], k)
                res[f"{kind}_s{n}"], res[f"{kind}_d{n}"], res[f"{kind}_c{n}"] = s, d, c
            assert (kind == "p2p") == bool(S._px), "the fused exchange was not the path taken"
            S.close()
        np.savez(out_path + f".{rank}.npz", **res)
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs with peer access")
def test_sharded_searcher_uses_fused_exchange(oracle, lib_built, tmp_path):
    """LuceneImpactSearcher(shard=True) on two GPUs: fused peer exchange == NCCL exchange == unsharded oracle."""
    from test_sharding_cpu import _make_collection, _queries
    k = 20
    out = str(tmp_path / "sx")
    mp.spawn(_searcher_worker, args=(2, _free_port(), k, out), nprocs=2, join=True)
    coll = _make_collection()
    qs = _queries(coll, nq=90)
    for n in (40, 25, 90):
        qi, qt, qw = coll.encode_queries(qs[:n])
        es, ed, ec = oracle.impact_search_c(coll.indptr, coll.post_docs, coll.post_impacts, coll.n_docs, qi, qt, qw, k)
        for r in range(2):
            got = np.load(out + f".{r}.npz")
            for kind in ("p2p", "nccl"):
                np.testing.assert_array_equal(got[f"{kind}_c{n}"], ec)
                np.testing.assert_array_equal(got[f"{kind}_d{n}"], ed)
                np.testing.assert_array_equal(got[f"{kind}_s{n}"], es)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_one_process_drives_two_devices(oracle, lib_built):
    """Opt-in shared memory (cudaFuncSetAttribute) is per device: a process that searched on cuda:0 must still be able
    to launch the > 48 KB tile merge, the row top-k and the dense kernels on cuda:1."""
    from mllm_sparse_retrieval_b200.ops import DenseIndex, SparseIndex, topk_rows
    rng = np.random.default_rng(88)
    n_docs, n_terms, k = 30000, 2000, 200
    indptr, docids, impacts = random_csr(rng, n_docs, n_terms, 32)
    qi, qt, qw = random_queries(rng, 16, n_terms, 12)
    exp = oracle.impact_search_c(indptr, docids, impacts, n_docs, qi, qt, qw, k, threads=4)
    x = rng.standard_normal((3000, 256)).astype(np.float32)
    xq = rng.standard_normal((8, 256)).astype(np.float32)
    for d in (0, 1):
        dev = torch.device("cuda", d)
        idx = SparseIndex(indptr, docids, impacts, n_docs, tile_docs=1024, device=dev)   # 30 tiles x 200: 64 KB merge buffer
        s, dd, c = idx.search(torch.from_numpy(qi.astype(np.int64)).to(dev), torch.from_numpy(qt.astype(np.int32)).to(dev),
                              torch.from_numpy(qw.astype(np.int32)).to(dev), k)
        idx.status()
        np.testing.assert_array_equal(dd.cpu().numpy(), exp[1])
        np.testing.assert_array_equal(s.cpu().numpy(), exp[0])
        idx.close()
        rows = torch.from_numpy(rng.standard_normal((4, 50000)).astype(np.float32)).to(dev)
        ts, td, _ = topk_rows(rows, 100)
        ref = torch.topk(rows, 100, dim=1)
        assert torch.equal(ts, ref.values)
        D = DenseIndex(256, device=dev)
        D.add(x)
        ds, di = D.search(xq, 10)
        ref_i = np.argsort(-(xq.astype(np.float64) @ x.T.astype(np.float64)), axis=1, kind="stable")[:, :10]
        np.testing.assert_array_equal(np.asarray(di), ref_i)
        D.close()


def _dense_worker(rank, world, port, out_path):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from mllm_sparse_retrieval_b200.searchers import FaissFlatSearcher
        rng = np.random.default_rng(5)
        x1 = rng.standard_normal((700, 128)).astype(np.float32)
        x2 = rng.standard_normal((301, 128)).astype(np.float32)
        x2[7] = x1[3]                                                      # an exact tie across the two blocks AND the two shards
        q = rng.standard_normal((90, 128)).astype(np.float32)
        res = {}
        for kind in ("p2p", "nccl"):
            os.environ["MSR_EXCHANGE"] = kind
            S = FaissFlatSearcher(x1, shard=True, device=dev)
            S.add(x1)
            S.add(x2)                                                      # a second block: the reference's own loop over shards
            for n, k in ((40, 10), (90, 100), (25, 100)):
                s, i = S.search(q[:n], k)
                res[f"{kind}_s{n}_{k}"], res[f"{kind}_i{n}_{k}"] = s, i
            assert (kind == "p2p") == bool(S._px), "the fused exchange was not the path taken"
        U = FaissFlatSearcher(x1, device=dev)                              # the same collection, unsharded, on this GPU
        U.add(x1); U.add(x2)
        for n, k in ((40, 10), (90, 100), (25, 100)):
            res[f"one_s{n}_{k}"], res[f"one_i{n}_{k}"] = U.search(q[:n], k)
        np.savez(out_path + f".{rank}.npz", **res)
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs with peer access")
def test_sharded_dense_searcher_uses_fused_exchange(oracle, lib_built, tmp_path):
    """FaissFlatSearcher(shard=True) on two GPUs, two add() blocks: rows published into peer slots + one-warp barrier + merge
    == NCCL all-gather + merge == the unsharded searcher, bit for bit (a row's fp32 score does not depend on the shard);
    and the float64 oracle: scores within 1e-3 relative, ids equal except where the oracle's own neighbouring scores tie
    within fp32 rounding."""
    out = str(tmp_path / "dx")
    mp.spawn(_dense_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    rng = np.random.default_rng(5)
    x1 = rng.standard_normal((700, 128)).astype(np.float32)
    x2 = rng.standard_normal((301, 128)).astype(np.float32)
    x2[7] = x1[3]
    q = rng.standard_normal((90, 128)).astype(np.float32)
    x = np.concatenate([x1, x2])
    for n, k in ((40, 10), (90, 100), (25, 100)):
        es, ei = oracle.flat_ip_search_numpy(x, q[:n], k)
        for r in range(2):
            got = np.load(out + f".{r}.npz")
            for kind in ("p2p", "nccl"):
                np.testing.assert_array_equal(got[f"{kind}_i{n}_{k}"], got[f"one_i{n}_{k}"])
                np.testing.assert_array_equal(got[f"{kind}_s{n}_{k}"], got[f"one_s{n}_{k}"])
            gi, gs = got[f"one_i{n}_{k}"], got[f"one_s{n}_{k}"]
            np.testing.assert_allclose(gs, es, rtol=1e-3, atol=1e-6)
            for row, col in zip(*np.nonzero(gi != ei)):
                near = np.abs(es[row] - es[row, col]) <= 4e-6 * max(1.0, abs(float(es[row, col])))
                assert near.sum() > 1, (n, k, row, col)                     # a genuine fp32 near-tie in the oracle's list
