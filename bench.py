#!/usr/bin/env python
"""bench.py -- headline benchmark: sparse top-100 queries/sec on the synthetic 1M-doc vocab-sparse index
(BASELINE.json configs[3]), doc-sharded over N GPUs with an NCCL all-gather merge.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

One step = ALL nq queries scored against the whole collection.
  value : queries/sec with the query CSR already resident in HBM (device path of the C ABI), max over ranks.
  e2e   : the same through msr_sparse_search(MSR_HOST): queries in pinned host memory, H2D + D2H inside the timed step.
  roofline : sparse_score_topk_kernel, algorithmic bytes (6 B/posting + 8 B/term + 8 B/result) over the
             kernel's own CUDA-event time (msr_sparse_profile), against MEASURED_PEAKS.json's HBM copy bandwidth.
  cpu_baseline : oracle/impact_oracle.c (Lucene-impact restatement, one query per thread, all host cores) on a
             bounded sample of the same queries.  Real Lucene/pyserini is unavailable on this box (no JVM).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "sparse top-100 queries/sec"
UNIT = "queries/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n-docs", type=int, default=1_000_000)
    ap.add_argument("--nnz-per-doc", type=int, default=128)
    ap.add_argument("--vocab", type=int, default=131072)
    ap.add_argument("--nq", type=int, default=10_000)
    ap.add_argument("--q-len", type=int, default=32)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--seed", type=int, default=1004)
    ap.add_argument("--tile-docs", type=int, default=0)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--chunks", type=int, default=1, help="query chunks of the pipelined multi-GPU exchange (1 = score -> gather -> merge; more was not faster)")
    ap.add_argument("--exchange", default="auto", choices=["auto", "p2p", "nccl"],
                    help="multi-GPU exchange: p2p = fused into the tile-merge epilogue over symmetric memory (PeerExchange); "
                         "nccl = all-gather (PipelinedExchange); auto = p2p when it can be set up")
    ap.add_argument("--check", type=int, default=32, help="queries verified against the oracle after the run")
    ap.add_argument("--workload", default="sparse", choices=["sparse", "dense", "hybrid", "config5"],
                    help="sparse = the headline metric (default); dense / hybrid = the other kernels of the path (bench_extra.py)")
    ap.add_argument("--dense-docs", type=int, default=25000)
    ap.add_argument("--dense-nq", type=int, default=5000)
    ap.add_argument("--dense-dim", type=int, default=4096)
    ap.add_argument("--hybrid-nq", type=int, default=25000)
    ap.add_argument("--c5-docs-per-gpu", type=int, default=1_250_000, help="config5: docs of each index per GPU (x8 = 10 M)")
    ap.add_argument("--c5-dense-nq", type=int, default=4096)
    ap.add_argument("--c5-dense-batch", type=int, default=1024)
    ap.add_argument("--depth", type=int, default=1000)
    ap.add_argument("--alpha", type=float, default=0.5)
    return ap.parse_args()


def workload_config(a, world):
    named = {(1_000_000, 10_000, 32): "BASELINE.json configs[3]", (1000, 5000, 12): "configs[0], Flickr-1K shape",
             (1000, 5000, 128): "configs[0], Flickr-1K shape, --sparse_manual", (5000, 25000, 12): "configs[1], COCO-5K t->i shape",
             (5000, 25000, 128): "configs[1], COCO-5K t->i shape, --sparse_manual",
             (25000, 5000, 128): "configs[1], COCO-5K i->t shape"}.get((a.n_docs, a.nq, a.q_len), "parity-test size")
    return {"workload": f"synthetic {a.n_docs}-doc vocab-sparse index (Zipf s=1, V={a.vocab}, {a.nnz_per_doc} nnz/doc), "
                        f"{a.nq} queries x {a.q_len} terms, top-{a.k} ({named})",
            "n_docs": a.n_docs, "nq": a.nq, "q_len": a.q_len, "k": a.k, "vocab": a.vocab, "seed": a.seed,
            "sharding": f"doc-range x{world}, per-GPU top-k lists exchanged over NVLink and merged" if world > 1 else "single GPU",
            "l2": "index (4 B/posting physical, >= 512 MB) is larger than the 126 MB L2; no explicit flush"}


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for (t, r) in self.rows if t0 <= t <= t1 + 0.2] or [r for (_, r) in self.rows]
        sm, mx, reasons = [], None, set()
        for r in rows:
            try:
                sm.append(float(r[1])); mx = float(r[2])
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel, **cfg):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture of this exact
    configuration (profiles/traffic.json), else None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            e = json.load(f).get(kernel)
        if e and all(cfg.get(k) == v for k, v in e["match"].items()):
            return e["dram_bytes_per_launch"]
    except (OSError, ValueError, KeyError):
        pass
    return None


def shard_collection(coll, lo, hi):
    """Doc-range shard [lo,hi) of a term-major CSR held as torch tensors (docs rebased to 0)."""
    import torch
    from mllm_sparse_retrieval_b200.synth import SparseCSR
    m = (coll.docids >= lo) & (coll.docids < hi)
    df = torch.diff(coll.indptr)
    term_of = torch.repeat_interleave(torch.arange(coll.n_terms, device=coll.indptr.device), df)
    cnt = torch.bincount(term_of[m], minlength=coll.n_terms)
    indptr = torch.zeros(coll.n_terms + 1, dtype=torch.int64, device=coll.indptr.device)
    indptr[1:] = torch.cumsum(cnt, 0)
    return SparseCSR(hi - lo, coll.n_terms, indptr, (coll.docids[m] - lo).to(torch.int32), coll.impacts[m])


def cpu_baseline(a, coll_np, n_docs, q_np, budget_s, threads=None):
    """Oracle (C, OpenMP, one query per thread) on a bounded sample of the workload's queries."""
    from oracle import oracle
    oracle.build()
    threads = threads or os.cpu_count() or 1
    qi, qt, qw = q_np
    nq = len(qi) - 1
    batch = max(threads * 2, 8)
    done, t_used = 0, 0.0
    # untimed warm-up batch (page-in, thread pool)
    sub = slice_queries(qi, qt, qw, 0, min(batch, nq))
    oracle.impact_search_c(*coll_np, n_docs, *sub, a.k, threads=threads)
    while done < nq and t_used < budget_s:
        hi = min(nq, done + batch)
        sub = slice_queries(qi, qt, qw, done, hi)
        t0 = time.perf_counter()
        oracle.impact_search_c(*coll_np, n_docs, *sub, a.k, threads=threads)
        t_used += time.perf_counter() - t0
        done = hi
        if t_used < budget_s / 8:
            batch *= 2
    return {"value": done / t_used, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"first {done} of {nq} queries against the full {n_docs}-doc index, {t_used:.1f} s, "
                      f"oracle/impact_oracle.c term-at-a-time + heap top-{a.k}, one query per OpenMP thread; "
                      f"host has {os.cpu_count()} logical cores; real Lucene/pyserini unavailable on this box (no JVM)"}, done, t_used


def slice_queries(qi, qt, qw, lo, hi):
    a, b = int(qi[lo]), int(qi[hi])
    return (qi[lo:hi + 1] - qi[lo]).astype(np.uint64), qt[a:b], qw[a:b]


def run_reference(a, rank, world):
    """--impl reference: the CPU search path (oracle port of Lucene impact search) on the host cores, same config."""
    if rank != 0:
        return
    import torch
    from mllm_sparse_retrieval_b200 import synth
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    coll = synth.make_collection(a.n_docs, a.nnz_per_doc, V=a.vocab, seed=a.seed, device=dev)
    qs = synth.make_queries(a.nq, a.q_len, V=a.vocab, seed=a.seed, device=dev)
    coll_np, q_np = coll.numpy(), qs.numpy()
    del coll
    from oracle import oracle
    oracle.build()
    threads = os.cpu_count() or 1
    qi, qt, qw = q_np
    # step = a bounded sample of the workload: sized from one probe so K+W steps end within a few minutes
    probe = slice_queries(qi, qt, qw, 0, min(a.nq, threads * 2))
    t0 = time.perf_counter()
    oracle.impact_search_c(*coll_np, a.n_docs, *probe, a.k, threads=threads)
    per_q = (time.perf_counter() - t0) / (len(probe[0]) - 1)
    budget = 150.0 / max(1, a.steps + a.warmup)
    n_step = int(max(threads, min(a.nq, budget / max(per_q, 1e-9))))
    times = []
    for it in range(a.warmup + a.steps):
        lo = (it * n_step) % max(1, a.nq - n_step + 1)
        sub = slice_queries(qi, qt, qw, lo, lo + n_step)
        t0 = time.perf_counter()
        oracle.impact_search_c(*coll_np, a.n_docs, *sub, a.k, threads=threads)
        if it >= a.warmup:
            times.append(time.perf_counter() - t0)
    total = sum(times)
    v = n_step * len(times) / total
    sample = (f"{n_step} of {a.nq} queries per step against the full {a.n_docs}-doc index; oracle/impact_oracle.c "
              f"(restatement of pyserini LuceneImpactSearcher.batch_search; real Lucene unavailable: no JVM), "
              f"one query per OpenMP thread, {threads} threads")
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
                      "warmup": a.warmup, "ms_per_step": 1e3 * total / len(times), "step_queries": n_step,
                      "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u32",
                      "data": "synthetic", "config": workload_config(a, world),
                      "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                      "gpu_launches": 0}), flush=True)


def run_ours(a, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from mllm_sparse_retrieval_b200 import synth
    from mllm_sparse_retrieval_b200.ops import SparseIndex, kernel_launches

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)

    # ---- workload: every rank generates the same seeded collection on its GPU and keeps its doc range ----
    coll = synth.make_collection(a.n_docs, a.nnz_per_doc, V=a.vocab, seed=a.seed, device=dev)
    qs = synth.make_queries(a.nq, a.q_len, V=a.vocab, seed=a.seed, device=dev)
    df_global = torch.diff(coll.indptr).to(torch.int32)
    bounds = [a.n_docs * g // world for g in range(world + 1)]
    lo, hi = bounds[rank], bounds[rank + 1]
    shard = coll if world == 1 else shard_collection(coll, lo, hi)
    coll_np = coll.numpy() if rank == 0 else None          # host copy for the oracle legs (rank 0 only)
    if world > 1:
        del coll
    idx = SparseIndex(shard.indptr, shard.docids, shard.impacts, shard.n_docs, device=dev, n_docs_global=a.n_docs,
                      df_global=df_global, doc_base=lo, tile_docs=a.tile_docs)
    del shard
    torch.cuda.empty_cache()
    postings, kept = idx.query_work(qs.indptr, qs.terms, qs.weights)
    nq, k = qs.nq, a.k

    # doc-sharded exchange: chunk i's all-gather + merge run on a side stream while chunk i+1 is scored
    from mllm_sparse_retrieval_b200.sharded import PipelinedExchange
    ex = PipelinedExchange(nq, k, dev, world, chunks=a.chunks if world > 1 else 1)
    q_chunks = [qs.slice(lo, hi) for lo, hi in ex.chunks()]

    # fused exchange over peer memory when available (one chunk): no NCCL call on the path
    px = None
    if world > 1 and a.exchange != "nccl" and a.chunks == 1:
        try:
            from mllm_sparse_retrieval_b200.sharded import PeerExchange
            px = PeerExchange(nq, k, dev, world, rank)
        except Exception as e:                                   # symmetric memory unavailable: NCCL all-gather instead
            if a.exchange == "p2p":
                raise
            sys.stderr.write(f"[bench] rank {rank}: PeerExchange unavailable ({type(e).__name__}: {e}); using NCCL\n")
            px = None
        # all ranks must agree (a rendezvous that failed on one rank fails on all, but be explicit)
        flag = torch.tensor([1 if px is not None else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            px = None

    def step_device():
        if px is not None:
            return px.run(idx, lambda o: idx.search(qs.indptr, qs.terms, qs.weights, k, out=o))
        return ex.run(lambda i, lo, hi, o: idx.search(q_chunks[i].indptr, q_chunks[i].terms, q_chunks[i].weights, k, out=o))

    # e2e: pinned host buffers.  1 GPU: the MSR_HOST path of the C ABI.  N GPUs: every rank copies the pinned query CSR
    # to its GPU, scores its shard, exchanges on the device, and rank 0 reads the merged lists back to pinned memory.
    h_qi = qs.indptr.cpu().pin_memory(); h_qt = qs.terms.cpu().pin_memory(); h_qw = qs.weights.cpu().pin_memory()
    h_out = (torch.empty((nq, k), dtype=torch.float32).pin_memory(), torch.empty((nq, k), dtype=torch.int32).pin_memory(),
             torch.empty((nq,), dtype=torch.int32).pin_memory())
    h2d = h_qi.numel() * 8 + h_qt.numel() * 4 + h_qw.numel() * 4
    d2h = nq * k * 8 + nq * 4
    d_q = (torch.empty_like(qs.indptr), torch.empty_like(qs.terms), torch.empty_like(qs.weights))
    cuts = [(int(qs.indptr[lo]), int(qs.indptr[hi])) for lo, hi in ex.chunks()]

    def step_host():
        if world == 1:
            idx.search_host(h_qi, h_qt, h_qw, k, out=h_out)
            return h_out
        d_q[0].copy_(h_qi, non_blocking=True); d_q[1].copy_(h_qt, non_blocking=True); d_q[2].copy_(h_qw, non_blocking=True)

        def chunk(i, lo, hi, o):
            a0, a1 = cuts[i]
            idx.search(d_q[0][lo:hi + 1] - a0, d_q[1][a0:a1], d_q[2][a0:a1], k, out=o)
        m = px.run(idx, lambda o: idx.search(d_q[0], d_q[1], d_q[2], k, out=o)) if px is not None else ex.run(chunk)
        if rank == 0:
            for dst, src in zip(h_out, m):
                dst.copy_(src, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return h_out

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

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(a.warmup):
        step_device()
    barrier()
    idx.profile(True)                       # start per-kernel event timing for the timed region
    launches0 = kernel_launches()
    t_wall0 = time.time()
    ms_dev = timed(step_device, a.steps, 0)
    t_wall1 = time.time()
    launches = kernel_launches() - launches0
    score_ms, merge_ms, n_launch = idx.profile(False)
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    ms_e2e = timed(step_host, a.steps, max(1, a.warmup))

    # ---- correctness spot-check against the oracle (rank 0, outside the timed region) ----
    final = step_device()
    torch.cuda.synchronize()
    idx.status()
    if rank != 0:
        return
    check = "skipped"
    if a.check > 0:
        from oracle import oracle
        oracle.build()
        n = min(a.check, nq)
        sub = slice_queries(*qs.numpy(), 0, n)
        es, ed, ec = oracle.impact_search_c(*coll_np, a.n_docs, *sub, k, threads=os.cpu_count() or 1)
        ok = np.array_equal(final[1][:n].cpu().numpy(), ed) and np.array_equal(final[0][:n].cpu().numpy(), es)
        ok = ok and np.array_equal(h_out[1][:n].numpy(), ed) and np.array_equal(h_out[0][:n].numpy(), es)
        check = f"bit-exact on {n} queries" if ok else "MISMATCH"
        if not ok:
            print(json.dumps({"error": "GPU result differs from the oracle", "metric": METRIC}), flush=True)
            raise SystemExit(3)

    peak, peak_src = load_peaks()
    alg_bytes_step = 6 * postings + 8 * kept + 8 * k * nq              # this rank's shard
    per_launch_ms = score_ms / max(1, n_launch)
    launches_per_step = max(1, n_launch // a.steps)
    achieved = alg_bytes_step / launches_per_step / (per_launch_ms * 1e-3) / 1e9 if per_launch_ms > 0 else 0.0
    roof = {"bound": "hbm", "kernel": "sparse_score_topk_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak,
            "traffic": ncu_traffic("sparse_score_topk_kernel", n_docs=a.n_docs, nq=a.nq, q_len=a.q_len, k=a.k, n_gpus=world,
                                   tile_docs=a.tile_docs) if launches_per_step == 1 else None,
            "peak_source": peak_src,
            "algorithmic_bytes_per_launch": alg_bytes_step / launches_per_step,
            "physical_posting_bytes_per_launch": 4 * postings / launches_per_step,
            "kernel_ms_per_launch": per_launch_ms, "kernel_share_of_step": score_ms / ms_dev,
            "merge_ms_per_step": merge_ms / a.steps, "postings_per_query": postings / nq, "rank": 0}
    res = {"metric": METRIC, "value": nq * a.steps / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world, "steps": a.steps,
           "warmup": a.warmup, "ms_per_step": ms_dev / a.steps, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "u32", "data": "synthetic", "config": workload_config(a, world),
           "e2e": {"value": nq * a.steps / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                   "api": "msr_sparse_search(MSR_HOST) with pinned host buffers" if world == 1 else
                          ("pinned query CSR -> msr_sparse_search(MSR_DEVICE) per shard with the exchange fused into the tile-merge "
                           "epilogue (peer stores over NVLink, symmetric memory) -> barrier -> msr_topk_merge_strided -> pinned host "
                           "lists on rank 0" if px is not None else
                           "pinned query CSR -> msr_sparse_search(MSR_DEVICE) per shard -> NCCL all-gather -> msr_topk_merge -> "
                           "pinned host lists on rank 0")},
           "gpu_launches": int(launches), "roofline": roof, "clocks": clocks, "parity_check": check,
           "exchange": "single GPU" if world == 1 else ("p2p: fused peer stores + signal-pad barrier" if px is not None else "nccl all-gather"),
           "index": {"n_tiles": idx.n_tiles, "tile_docs": idx.tile_docs, "nnz_shard": idx.nnz, "resident_bytes": idx.resident_bytes}}
    if idx.resident_bytes * world < (126 << 20):
        res["config"]["l2"] = (f"index ({idx.resident_bytes * world >> 20} MB) fits the 126 MB L2: launch / latency-bound, "
                               "not a roofline configuration (SURVEY 8(d)); QPS and exactness only")
    if world == 1 and not a.no_cpu_baseline:
        cb, _, _ = cpu_baseline(a, coll_np, a.n_docs, qs.numpy(), a.cpu_seconds)
        res["cpu_baseline"] = cb
    print(json.dumps(res), flush=True)


def main():
    a = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference":
        run_reference(a, rank, world)
        return
    if a.workload in ("dense", "hybrid"):
        if rank == 0:
            import bench_extra
            (bench_extra.run_dense if a.workload == "dense" else bench_extra.run_hybrid)(a, ROOT, ClockSampler)
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        # NCCL may print its version banner on STDOUT at communicator creation; stdout carries the one JSON line only,
        # so fd 1 points at stderr until the first collective has run
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    try:
        if a.workload == "config5":
            import bench_extra
            bench_extra.run_config5(a, rank, world, local_rank, ROOT, ClockSampler)
        else:
            run_ours(a, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
