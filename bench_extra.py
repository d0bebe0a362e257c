"""Secondary bench legs of bench.py (`--workload dense|hybrid`): the same JSON contract on the other kernels of the path.

dense  : BASELINE.json configs[1] (COCO-5K shape): dense flat-IP top-k, [nq, 4096] x [N, 4096], kernel (c).
         roofline.bound = "tensor": 2*nq*N*H flops over the tcgen05 pass's own CUDA-event time against
         MEASURED_PEAKS.json's bf16 peak.  cpu_baseline = torch-CPU (MKL sgemm) q @ X^T + topk, the shape of FAISS's own
         CPU path for nq >= 20 (real FAISS is unavailable on this box).
hybrid : configs[2]: fusion of a dense and a sparse run at depth 1000, kernel (d); QPS only (latency-bound).
"""
from __future__ import annotations

import json
import os
import time

import numpy as np


def _peaks(root):
    p = os.path.join(root, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            j = json.load(f)
        return float(j["bf16_tflops"]), float(j.get("bf16_tflops_sustained", j["bf16_tflops"])), "measured (MEASURED_PEAKS.json)"
    return 1590.0, 1400.0, "fallback (B200_PROFILING.md)"


def _traffic(root, kernel, **cfg):
    try:
        with open(os.path.join(root, "profiles", "traffic.json")) as f:
            e = json.load(f).get(kernel)
        if e and all(cfg.get(k) == v for k, v in e["match"].items()):
            return e["dram_bytes_per_launch"]
    except (OSError, ValueError, KeyError):
        pass
    return None


def _timed(fn, steps, warmup, torch):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


def run_dense(a, root, ClockSampler, emit=True):
    import torch
    from mllm_sparse_retrieval_b200 import synth
    from mllm_sparse_retrieval_b200.ops import DenseIndex, kernel_launches
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    N, nq, H, k = a.dense_docs, a.dense_nq, a.dense_dim, a.k
    docs, q = synth.make_dense(N, nq, dim=H, seed=1002, device=dev)
    idx = DenseIndex(H, device=dev)
    idx.add(docs)
    out = (torch.empty((nq, k), dtype=torch.float32, device=dev), torch.empty((nq, k), dtype=torch.int64, device=dev))
    h_q = q.cpu().pin_memory()
    h_out = (torch.empty((nq, k), dtype=torch.float32).pin_memory(), torch.empty((nq, k), dtype=torch.int64).pin_memory())

    sampler = ClockSampler(0)
    sampler.start()
    l0 = kernel_launches()
    t0 = time.time()
    ms_dev = _timed(lambda: idx.search(q, k, mode=0, out=out), a.steps, a.warmup, torch)
    t1 = time.time()
    launches = kernel_launches() - l0 - 0
    clocks = sampler.stop(t0, t1)
    fallback = idx.last_fallback
    ms_e2e = _timed(lambda: idx.search(h_q, k, mode=0, out=h_out), a.steps, max(1, a.warmup), torch)
    # the fused tcgen05 pass (GEMM + selection epilogue) timed by the library's own CUDA events inside real searches
    idx.profile(True)
    for _ in range(a.steps):
        idx.search(q, k, mode=0, out=out)
    tc_ms, rs_ms, n_blk = idx.profile(False)
    ms_tc, ms_rs = tc_ms / a.steps, rs_ms / a.steps
    ms_gemm_only = _timed(lambda: idx.tc_scores(q), a.steps, a.warmup, torch) / a.steps
    ms_exact = _timed(lambda: idx.search(q[:256], k, mode=1), 2, 1, torch) / 2

    # parity: ids of the tensor-core path == exact path on a sample, scores bit-identical
    s1, i1 = idx.search(q[:256], k, mode=1)
    ok = bool((i1 == out[1][:256]).all().item()) and bool((s1 == out[0][:256]).all().item())
    if not ok:
        print(json.dumps({"error": "dense: tensor-core path differs from the exact path"}), flush=True)
        raise SystemExit(3)

    burst, sustained, src = _peaks(root)
    flops = 2.0 * nq * N * H
    achieved = flops / (ms_tc * 1e-3) / 1e12
    # CPU stand-in for FAISS IndexFlatIP on the host cores, bounded sample
    cb = None
    if not a.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        xc, n_s = docs.cpu(), min(nq, 512)
        qc = q[:n_s].cpu()
        torch.topk(qc[:32] @ xc.T, k, dim=1)
        t = time.perf_counter()
        reps = 0
        while time.perf_counter() - t < max(2.0, a.cpu_seconds / 2):
            torch.topk(qc @ xc.T, k, dim=1)
            reps += 1
        dt = time.perf_counter() - t
        cb = {"value": n_s * reps / dt, "unit": "queries/s", "cores": os.cpu_count(), "kind": "port",
              "sample": f"{reps} x {n_s} of {nq} queries against all {N} docs, {dt:.1f} s; torch-CPU (MKL sgemm) q @ X^T + "
                        f"topk-{k}, fp32, all {os.cpu_count()} logical cores -- the shape of FAISS IndexFlatIP's CPU path "
                        f"for nq >= 20; real FAISS unavailable on this box (no wheel)"}
    res = {"metric": "dense flat-IP top-%d queries/sec" % k, "value": nq * a.steps / (ms_dev * 1e-3), "unit": "queries/s",
           "n_gpus": 1, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_dev / a.steps, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "fp16 tensor pass (fp32 accumulate) + fp32 exact re-score", "data": "synthetic",
           "config": {"workload": f"COCO-5K-shaped dense search: {nq} queries x {N} docs x {H} dims, top-{k} "
                                  f"(BASELINE.json configs[1], SURVEY 8(d) config 2)", "n_docs": N, "nq": nq, "dim": H, "k": k,
                      "l2": "fp32+fp16 index = %d MB > 126 MB L2" % (N * H * 6 // 2 ** 20)},
           "e2e": {"value": nq * a.steps / (ms_e2e * 1e-3), "unit": "queries/s", "h2d_bytes_per_step": nq * H * 4,
                   "d2h_bytes_per_step": nq * k * 12, "api": "msr_dense_search(MSR_HOST), pinned host buffers"},
           "gpu_launches": int(launches),
           "roofline": {"bound": "tensor", "kernel": "dense_tc_kernel<FILTER> (tcgen05.mma fp16, fused threshold selection)", "achieved": achieved, "peak": burst,
                        "unit": "TFLOP/s", "frac": achieved / burst, "frac_of_sustained": achieved / sustained,
                        "traffic": _traffic(root, "dense_tc_kernel<1>", n_docs=N, nq=nq, dim=H, k=k),
                        "peak_source": src + " dense bf16 figure (fp16 and bf16 share the kind::f16 tensor rate); burst: the kernel is "
                                      "bracketed by its own events", "flops_per_launch": flops,
                        "kernel_ms_per_launch": ms_tc, "kernel_share_of_step": ms_tc / (ms_dev / a.steps),
                        "rescore_ms_per_step": ms_rs, "unfused_gemm_ms": ms_gemm_only, "query_blocks_per_step": n_blk / a.steps},
           "certificate_fallback_queries": fallback, "exact_mode_qps": 256 / (ms_exact * 1e-3),
           "clocks": clocks, "parity_check": "ids and score bits equal to the exact fp32 pass on 256 queries"}
    if cb:
        res["cpu_baseline"] = cb
    if not emit:
        return res
    print(json.dumps(res), flush=True)


def run_hybrid(a, root, ClockSampler, emit=True):
    import torch
    from mllm_sparse_retrieval_b200.ops import fuse_lists, kernel_launches
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev)
    g.manual_seed(1003)
    nq, kd, ks, N = a.hybrid_nq, a.depth, a.depth, 5000
    kd, ks = min(kd, N), min(ks, N)
    # two runs over the same 5000-doc collection: random permutations give ~kd*ks/N shared docs per query
    d_docs = torch.argsort(torch.rand((nq, N), generator=g, device=dev), dim=1)[:, :kd].to(torch.int32).contiguous()
    s_docs = torch.argsort(torch.rand((nq, N), generator=g, device=dev), dim=1)[:, :ks].to(torch.int32).contiguous()
    d_s = torch.sort(torch.rand((nq, kd), generator=g, device=dev) * 0.6 + 0.2, dim=1, descending=True).values.contiguous()
    s_s = torch.sort(torch.randint(1, 60000, (nq, ks), generator=g, device=dev).float(), dim=1, descending=True).values.contiguous()
    cnt_d = torch.full((nq,), kd, dtype=torch.int32, device=dev)
    cnt_s = torch.full((nq,), ks, dtype=torch.int32, device=dev)
    self_ids = torch.arange(nq, device=dev, dtype=torch.int32) % N

    def step():
        return fuse_lists(d_s, d_docs, cnt_d, s_s, s_docs, cnt_s, self_ids, True, a.alpha, 1 - a.alpha)

    sampler = ClockSampler(0)
    sampler.start()
    l0 = kernel_launches()
    t0 = time.time()
    ms = _timed(step, a.steps, a.warmup, torch)
    t1 = time.time()
    launches = kernel_launches() - l0
    clocks = sampler.stop(t0, t1)
    # e2e: lists in pinned host memory -> device -> fuse -> fused lists back on the host
    hp = [t.cpu().pin_memory() for t in (d_s, d_docs, cnt_d, s_s, s_docs, cnt_s, self_ids)]

    ho = None

    def step_host():
        nonlocal ho
        dv = [t.to(dev, non_blocking=True) for t in hp]
        o = fuse_lists(dv[0], dv[1], dv[2], dv[3], dv[4], dv[5], dv[6], True, a.alpha, 1 - a.alpha)
        if ho is None:                                           # pinned result buffers, allocated once
            ho = [torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in o]
        for dst, src in zip(ho, o):
            dst.copy_(src, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return ho

    ms_e2e = _timed(step_host, a.steps, max(1, a.warmup), torch)
    fs, fd, fsrc, fc = step()
    torch.cuda.synchronize()
    # check + CPU baseline: the oracle's restatement of hybrid.fuse (pure Python, 1 core) on a bounded sample
    from oracle import oracle
    n_s = min(nq, 64)
    names = [str(i) for i in range(N)]
    h = [t[:n_s].cpu().numpy() for t in (d_s, d_docs, s_s, s_docs)]
    qids = [names[int(x)] for x in self_ids[:n_s].cpu().numpy()]
    qids = [f"{q}" for q in qids]
    dr = oracle.get_run_dict(qids, [[np.float32(x) for x in h[0][i]] for i in range(n_s)],
                             [[names[j] for j in h[1][i]] for i in range(n_s)], True)
    sr = oracle.get_run_dict(qids, [[float(x) for x in h[2][i]] for i in range(n_s)],
                             [[names[j] for j in h[3][i]] for i in range(n_s)], True)
    t = time.perf_counter()
    exp = oracle.fuse([dr, sr], [a.alpha, 1 - a.alpha])
    dt = time.perf_counter() - t
    ok = True
    if len(set(qids)) == n_s:
        for i, qid in enumerate(qids):
            n = int(fc[i])
            got = [(names[int(d)], float(s)) for d, s in zip(fd[i, :n].cpu().numpy(), fs[i, :n].cpu().numpy())]
            e = [(d, float(v)) for d, v in exp[qid].items()]
            ok = ok and [g_[0] for g_ in got] == [x[0] for x in e] and np.allclose([g_[1] for g_ in got], [x[1] for x in e], rtol=1e-6)
    if not ok:
        print(json.dumps({"error": "hybrid: fused lists differ from the oracle"}), flush=True)
        raise SystemExit(3)
    res = {"metric": "hybrid fusion queries/sec", "value": nq * a.steps / (ms * 1e-3), "unit": "queries/s", "n_gpus": 1,
           "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "f32/f64 (rule H3)", "data": "synthetic",
           "config": {"workload": f"COCO-5K hybrid fusion: {nq} queries, dense depth {kd} + sparse depth {ks}, alpha {a.alpha}, "
                                  f"remove_query (BASELINE.json configs[2])", "nq": nq, "k_dense": kd, "k_sparse": ks},
           "e2e": {"value": nq * a.steps / (ms_e2e * 1e-3), "unit": "queries/s",
                   "h2d_bytes_per_step": int(sum(t.numel() * t.element_size() for t in hp)),
                   "d2h_bytes_per_step": int(nq * (kd + ks) * 13 + nq * 4), "api": "pinned lists -> msr_fuse -> host"},
           "gpu_launches": int(launches),
           "roofline": {"bound": "hbm", "kernel": "fuse_kernel", "achieved": 16.0 * (kd + ks) * nq / (ms / a.steps * 1e-3) / 1e9,
                        "peak": None, "unit": "GB/s", "frac": None, "traffic": None,
                        "note": "latency-bound by design (<= 32 KB per query); QPS is the figure of merit (SURVEY 8(d))"},
           "clocks": clocks, "parity_check": f"doc order identical, scores within 1e-6 of the oracle on {n_s} queries",
           "cpu_baseline": {"value": n_s / dt, "unit": "queries/s", "cores": 1, "kind": "port",
                            "sample": f"{n_s} of {nq} queries, {dt:.2f} s; oracle.fuse = restatement of the reference's pure-Python "
                                      f"hybrid.fuse (src/hybrid.py:32-53), single core as in the reference"}}
    if not emit:
        return res
    print(json.dumps(res), flush=True)


# -------------------------------------------------------------------------------------------------------------
# config 5 (BASELINE.json configs[4]): 10 M-doc sparse index + 4096-d dense flat-IP, doc-sharded over the GPUs of one
# box with an NCCL all-gather merge.  Per-GPU work is FIXED (1.25 M docs of each kind per GPU): at 8 GPUs the collection
# is the named 10 M docs; at fewer GPUs it is the same shards, fewer of them ("scaling": "weak").
# -------------------------------------------------------------------------------------------------------------
def _dense_chunks(n_rows, dim, seed, dev, torch, chunk=65536, groups=4096):
    """Seeded generator of unit-norm fp32 doc blocks (same law as synth.make_dense), never materialised as a whole."""
    gen = torch.Generator(device=dev)
    gen.manual_seed(seed)
    cent = torch.randn((groups, dim), generator=gen, device=dev)
    cent = cent / cent.norm(dim=1, keepdim=True)
    scale = 0.3 * dim ** 0.5
    for lo in range(0, n_rows, chunk):
        n = min(chunk, n_rows - lo)
        g = torch.randint(0, groups, (n,), generator=gen, device=dev)
        x = torch.randn((n, dim), generator=gen, device=dev) + scale * cent[g]
        yield lo, x / x.norm(dim=1, keepdim=True)


def run_config5(a, rank, world, local_rank, root, ClockSampler):
    import torch
    import torch.distributed as dist
    from mllm_sparse_retrieval_b200 import synth
    from mllm_sparse_retrieval_b200.ops import DenseIndex, SparseIndex, fuse_lists, kernel_launches
    from mllm_sparse_retrieval_b200.sharded import PipelinedExchange, make_peer_exchange
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    per_gpu, V, k, H = a.c5_docs_per_gpu, a.vocab, a.k, a.dense_dim
    n_total = per_gpu * world
    base = rank * per_gpu

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- sparse shard: drawn independently per rank from the shared term law; df is summed over the ranks ----
    t_build = time.time()
    coll = synth.make_collection(per_gpu, a.nnz_per_doc, V=V, seed=1005 + 101 * (rank + 1), perm_seed=1005, device=dev)
    df = torch.diff(coll.indptr)
    if world > 1:
        dist.all_reduce(df, op=dist.ReduceOp.SUM)
    sidx = SparseIndex(coll.indptr, coll.docids, coll.impacts, per_gpu, device=dev, n_docs_global=n_total,
                       df_global=df.to(torch.int32), doc_base=base, tile_docs=a.tile_docs)
    coll_np = coll.numpy() if rank == 0 else None
    del coll
    torch.cuda.empty_cache()
    qs = synth.make_queries(a.nq, a.q_len, V=V, seed=1005, device=dev)
    postings, kept = sidx.query_work(qs.indptr, qs.terms, qs.weights)
    ex_s = PipelinedExchange(qs.nq, k, dev, world, chunks=a.chunks if world > 1 else 1)
    q_chunks = [qs.slice(lo, hi) for lo, hi in ex_s.chunks()]

    # fused exchange over peer memory (shared thresholds, peer stores, one-warp barrier) when it can be set up
    px = make_peer_exchange(qs.nq, k, dev, world, rank, kind=a.exchange) if (world > 1 and a.chunks == 1) else None

    def sparse_step():
        if px is not None:
            return px.run(sidx, lambda o: sidx.search(qs.indptr, qs.terms, qs.weights, k, out=o))
        return ex_s.run(lambda i, lo, hi, o: sidx.search(q_chunks[i].indptr, q_chunks[i].terms, q_chunks[i].weights, k, out=o))

    # ---- dense shard: generated on the device block by block ----
    didx = DenseIndex(H, device=dev, doc_base=base)
    for lo, x in _dense_chunks(per_gpu, H, 2005 + 101 * (rank + 1), dev, torch):
        didx.add(x)
    del x
    torch.cuda.empty_cache()
    gq = torch.Generator(device=dev)
    gq.manual_seed(3005)
    dq = torch.randn((a.c5_dense_nq, H), generator=gq, device=dev)
    # queries near docs of rank 0's first block so that true neighbours exist
    dq = dq / dq.norm(dim=1, keepdim=True)
    dbatch = a.c5_dense_batch
    ex_d = PipelinedExchange(a.c5_dense_nq, k, dev, world, chunks=max(1, a.c5_dense_nq // dbatch), min_chunk=dbatch,
                             id_dtype=torch.int64)
    ex_d.local[2].fill_(k)

    def dense_chunk(i, lo, hi, o):
        didx.search(dq[lo:hi], k, mode=0, out=(o[0], o[1]))

    # dense exchange: rows published into peer slots after each batch (msr_exchange_publish + one-warp barrier + merge) on
    # the search's own stream.  (Round 1 overlapped an NCCL all-gather of batch i with the GEMM of batch i+1: the NCCL
    # kernels took SMs from the persistent one-CTA-per-SM GEMM, whose statically scheduled late CTAs then stretched every
    # pass from 7.5 to 10.3 ms.)
    px_d = make_peer_exchange(dbatch, k, dev, world, rank, kind=a.exchange) if world > 1 else None
    d_final = (torch.empty((a.c5_dense_nq, k), dtype=torch.float32, device=dev), torch.empty((a.c5_dense_nq, k), dtype=torch.int32, device=dev),
               torch.empty((a.c5_dense_nq,), dtype=torch.int32, device=dev))
    d_loc = (torch.empty((dbatch, k), dtype=torch.float32, device=dev), torch.empty((dbatch, k), dtype=torch.int64, device=dev))
    d_cnt = torch.full((dbatch,), k, dtype=torch.int32, device=dev)

    def dense_step():
        if px_d is None:
            return ex_d.run(dense_chunk)
        for lo in range(0, a.c5_dense_nq, dbatch):
            hi = min(a.c5_dense_nq, lo + dbatch)
            n = hi - lo
            didx.search(dq[lo:hi], k, mode=0, out=(d_loc[0][:n], d_loc[1][:n]))
            ms_, md_, mc_ = px_d.exchange_rows(d_loc[0][:n], d_loc[1][:n].to(torch.int32), d_cnt[:n])
            d_final[0][lo:hi].copy_(ms_); d_final[1][lo:hi].copy_(md_); d_final[2][lo:hi].copy_(mc_)
        return d_final
    build_s = time.time() - t_build

    # ---- timed regions (device-resident inputs; max over ranks) ----
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(a.warmup):
        sparse_step()
    barrier()
    sidx.profile(True)
    l0 = kernel_launches()
    t0 = time.time()
    ms_sparse = timed(sparse_step, a.steps, 0)
    t1 = time.time()
    launches = kernel_launches() - l0
    score_ms, merge_ms, n_launch = sidx.profile(False)
    clocks = sampler.stop(t0, t1) if rank == 0 else None
    dense_step()
    dense_step()
    didx.profile(True)
    ms_dense = timed(dense_step, a.steps, 0)
    tc_ms, rs_ms, n_blk = didx.profile(False)
    # uncertified queries of one whole step, summed over every batch and every rank (outside the timed region: reading the
    # device-side counter synchronises).  One of them re-reads a rank's whole fp32 shard and every rank waits for it.
    n_fb = 0
    for lo in range(0, a.c5_dense_nq, dbatch):
        didx.search(dq[lo:min(a.c5_dense_nq, lo + dbatch)], k, mode=0)
        n_fb += didx.last_fallback
    if world > 1:
        t_fb = torch.tensor([n_fb], dtype=torch.int64, device=dev)
        dist.all_reduce(t_fb, op=dist.ReduceOp.SUM)
        n_fb = int(t_fb.item())

    # ---- e2e: pinned host query buffers in, merged lists back to pinned host memory on rank 0 ----
    h_q = tuple(t.cpu().pin_memory() for t in (qs.indptr, qs.terms, qs.weights))
    d_q = tuple(torch.empty_like(t) for t in (qs.indptr, qs.terms, qs.weights))
    h_out = (torch.empty((qs.nq, k), dtype=torch.float32).pin_memory(), torch.empty((qs.nq, k), dtype=torch.int32).pin_memory(),
             torch.empty((qs.nq,), dtype=torch.int32).pin_memory())
    cuts = [(int(qs.indptr[lo]), int(qs.indptr[hi])) for lo, hi in ex_s.chunks()]

    def sparse_host_step():
        for d, h in zip(d_q, h_q):
            d.copy_(h, non_blocking=True)

        def chunk(i, lo, hi, o):
            a0, a1 = cuts[i]
            sidx.search(d_q[0][lo:hi + 1] - a0, d_q[1][a0:a1], d_q[2][a0:a1], k, out=o)
        m = ex_s.run(chunk)
        if rank == 0:
            for dst, src in zip(h_out, m):
                dst.copy_(src, non_blocking=True)
        torch.cuda.current_stream().synchronize()
    ms_e2e = timed(sparse_host_step, a.steps, 1)

    s_fin = tuple(t.clone() for t in sparse_step())
    d_fin = tuple(t.clone() for t in dense_step())
    torch.cuda.synchronize()
    sidx.status()

    # ---- hybrid fusion of the merged lists (first c5_dense_nq sparse queries x the dense queries) on every rank ----
    nf = min(a.c5_dense_nq, qs.nq)
    self_ids = torch.full((nf,), -1, dtype=torch.int32, device=dev)
    f_args = (d_fin[0][:nf].contiguous(), d_fin[1][:nf].to(torch.int32).contiguous(), d_fin[2][:nf].contiguous(),
              s_fin[0][:nf].contiguous(), s_fin[1][:nf].contiguous(), s_fin[2][:nf].contiguous(), self_ids, False, a.alpha, 1 - a.alpha)
    ms_fuse = timed(lambda: fuse_lists(*f_args), a.steps, 1)

    # ---- parity at full size: size-independent properties + bounded oracle checks on rank 0's shard ----
    problems = []
    sc, dc = s_fin[2].long(), d_fin[2].long()
    for name, (S, D, C) in (("sparse", s_fin), ("dense", d_fin)):
        valid = torch.arange(k, device=dev)[None, :] < C.long()[:, None]
        pair_ok = (S[:, :-1] > S[:, 1:]) | ((S[:, :-1] == S[:, 1:]) & (D[:, :-1] < D[:, 1:]))
        if not bool((pair_ok | ~valid[:, 1:]).all()):
            problems.append(f"{name}: merged lists not ordered by (score desc, id asc)")
        if not bool(((D >= 0) & (D < n_total) | ~valid).all()):
            problems.append(f"{name}: doc id outside [0, {n_total})")
    if world > 1:                                               # every rank must hold the same merged lists
        for name, D in (("sparse", s_fin[1]), ("dense", d_fin[1])):
            h = (D.long() * torch.arange(1, D.numel() + 1, device=dev).view_as(D) % 1000003).sum().view(1)
            lo_, hi_ = h.clone(), h.clone()
            dist.all_reduce(lo_, op=dist.ReduceOp.MIN); dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
            if int(lo_) != int(hi_):
                problems.append(f"{name}: ranks disagree on the merged lists")
    if rank == 0 and a.check > 0:
        from oracle import oracle
        oracle.build()
        n = min(8, a.check)
        sub = qs.slice(0, n)
        ls, ld, lc = sidx.search(sub.indptr, sub.terms, sub.weights, k)
        es, ed, ec = oracle.impact_search_c(*coll_np, per_gpu, *sub.numpy(), k, threads=os.cpu_count() or 1,
                                            n_docs_global=n_total, df_global=df.cpu().numpy().astype(np.uint32), doc_base=base)
        if not (np.array_equal(ld.cpu().numpy(), ed) and np.array_equal(ls.cpu().numpy(), es)):
            problems.append("sparse: rank 0 shard differs from the oracle")
        # dense: float64 scores of 16 queries against the regenerated shard
        nd = 16
        best_s = torch.full((nd, 0), 0.0, dtype=torch.float64, device=dev)
        best_i = torch.zeros((nd, 0), dtype=torch.int64, device=dev)
        for lo, x in _dense_chunks(per_gpu, H, 2005 + 101 * (rank + 1), dev, torch):
            sc64 = dq[:nd].double() @ x.double().T
            best_s = torch.cat([best_s, sc64], 1); best_i = torch.cat([best_i, (base + lo + torch.arange(x.shape[0], device=dev))[None, :].expand(nd, -1)], 1)
            top = torch.topk(best_s, k, dim=1)
            best_s, best_i = top.values, torch.gather(best_i, 1, top.indices)
        gs, gi = didx.search(dq[:nd], k, mode=0)
        same = float((gi == best_i).float().mean())
        if same < 0.999 or not torch.allclose(gs.double(), best_s, rtol=1e-3, atol=2e-6):
            problems.append(f"dense: rank 0 shard differs from the float64 scores (id agreement {same:.4f})")
    if rank != 0:
        return
    if problems:
        print(json.dumps({"error": "; ".join(problems), "metric": "sparse top-%d queries/sec" % k}), flush=True)
        raise SystemExit(3)

    hbm, hbm_src = 6650.0, "fallback (B200_PROFILING.md)"
    p = os.path.join(root, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            hbm, hbm_src = float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    burst, sustained, tsrc = _peaks(root)
    alg = 6 * postings + 8 * kept + 8 * k * qs.nq
    per_launch = score_ms / max(1, n_launch)
    lps = max(1, n_launch // a.steps)
    achieved = alg / lps / (per_launch * 1e-3) / 1e9 if per_launch > 0 else 0.0
    d_flops = 2.0 * a.c5_dense_nq * per_gpu * H
    d_ach = d_flops / (tc_ms / a.steps * 1e-3) / 1e12 if tc_ms > 0 else 0.0
    res = {"metric": "sparse top-%d queries/sec" % k, "value": qs.nq * a.steps / (ms_sparse * 1e-3), "unit": "queries/s",
           "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_sparse / a.steps, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
           "config": {"workload": f"synthetic {n_total}-doc sparse index ({per_gpu} docs/GPU, Zipf s=1, V={V}, {a.nnz_per_doc} nnz/doc, "
                                  f"{qs.nq} queries x {a.q_len} terms) + {H}-d dense flat-IP over the same {n_total} docs "
                                  f"({a.c5_dense_nq} queries in batches of {dbatch}), top-{k}, doc-sharded x{world}, lists exchanged by "
                                  f"{'peer stores + one-warp barrier + merge' if px is not None else 'NCCL all-gather + merge'} "
                                  f"(BASELINE.json configs[4]; the named 10 M docs at 8 GPUs)",
                      "docs_per_gpu": per_gpu, "n_docs": n_total, "nq": qs.nq, "k": k, "dense_nq": a.c5_dense_nq, "dim": H,
                      "l2": "per-GPU sparse index %d MB and dense index %d GB are larger than the 126 MB L2" %
                            (sidx.resident_bytes >> 20, per_gpu * H * 6 >> 30)},
           "e2e": {"value": qs.nq * a.steps / (ms_e2e * 1e-3), "unit": "queries/s", "h2d_bytes_per_step": sum(t.numel() * t.element_size() for t in h_q),
                   "d2h_bytes_per_step": qs.nq * k * 8 + qs.nq * 4,
                   "api": "pinned query CSR -> msr_sparse_search(MSR_DEVICE) per shard -> NCCL all-gather -> msr_topk_merge -> pinned host lists on rank 0"},
           "gpu_launches": int(launches),
           "roofline": {"bound": "hbm", "kernel": "sparse_score_topk_kernel", "achieved": achieved, "peak": hbm, "unit": "GB/s",
                        "frac": achieved / hbm, "traffic": None, "peak_source": hbm_src, "algorithmic_bytes_per_launch": alg / lps,
                        "kernel_ms_per_launch": per_launch, "kernel_share_of_step": score_ms / ms_sparse, "rank": 0},
           "dense": {"value": a.c5_dense_nq * a.steps / (ms_dense * 1e-3), "unit": "queries/s", "ms_per_step": ms_dense / a.steps,
                     "fused_tc_ms_per_step": tc_ms / a.steps, "rescore_ms_per_step": rs_ms / a.steps, "achieved_tflops_rank0": d_ach,
                     "frac_of_peak": d_ach / burst, "frac_of_sustained": d_ach / sustained, "peak_source": tsrc,
                     "certificate_fallback_queries_per_step": n_fb, "query_blocks_per_step": n_blk / a.steps},
           "hybrid": {"value": nf * a.steps / (ms_fuse * 1e-3), "unit": "queries/s", "nq": nf},
           "clocks": clocks, "index_build_s": build_s,
           "parity_check": "merged lists ordered and identical on all ranks; rank-0 shard: sparse bit-exact vs the C oracle on 8 "
                           "queries, dense ids identical to float64 scores on 16 queries"}
    print(json.dumps(res), flush=True)
