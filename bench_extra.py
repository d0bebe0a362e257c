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


def run_dense(a, root, ClockSampler):
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
                        "unit": "TFLOP/s", "frac": achieved / burst, "frac_of_sustained": achieved / sustained, "traffic": None,
                        "peak_source": src + " dense bf16 figure (fp16 and bf16 share the kind::f16 tensor rate); burst: the kernel is "
                                      "bracketed by its own events", "flops_per_launch": flops,
                        "kernel_ms_per_launch": ms_tc, "kernel_share_of_step": ms_tc / (ms_dev / a.steps),
                        "rescore_ms_per_step": ms_rs, "unfused_gemm_ms": ms_gemm_only, "query_blocks_per_step": n_blk / a.steps},
           "certificate_fallback_queries": fallback, "exact_mode_qps": 256 / (ms_exact * 1e-3),
           "clocks": clocks, "parity_check": "ids and score bits equal to the exact fp32 pass on 256 queries"}
    if cb:
        res["cpu_baseline"] = cb
    print(json.dumps(res), flush=True)


def run_hybrid(a, root, ClockSampler):
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

    def step_host():
        dv = [t.to(dev, non_blocking=True) for t in hp]
        o = fuse_lists(dv[0], dv[1], dv[2], dv[3], dv[4], dv[5], dv[6], True, a.alpha, 1 - a.alpha)
        return [t.cpu() for t in o]

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
    print(json.dumps(res), flush=True)
