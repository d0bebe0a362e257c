"""Hybrid fusion and TREC run I/O with the reference's call surface (/root/reference/src/hybrid.py).

  fuse(runs, weights)            <- hybrid.fuse            hybrid.py:32-53   (caller search.py:455-461)
  fuse_statistic(runs, weights)  <- hybrid.fuse_statistic  hybrid.py:56-90   (caller score_statistic.py:393)
  get_run_dict(...)              <- search.get_run_dict    search.py:66-82
  read_trec_run / write_trec_run <- hybrid.py:8-29 (the run-file format; the reference never calls the writer)

``fuse`` materialises the two run dicts as padded arrays, runs kernel (d) (csrc/fuse.cu) and rebuilds the dict in
the reference's insertion order.  The arithmetic type follows the reference under NumPy >= 2: a run whose scores are
np.float32 (what FAISS returns) is combined in float32, a run of Python floats (Lucene hits, or runs re-read from
a TREC file) in float64.  ``fuse_arrays`` is the same kernel without the dict round trip.
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops


class ResultRecord:
    def __init__(self, score, type):
        self.score = score
        self.type = type


def get_run_dict(batch_ids, batch_scores, batch_rankings, remove_query):
    run_dict = {}
    for qid, scores, rankings in zip(batch_ids, batch_scores, batch_rankings):
        docs = {}
        for score, doc in zip(scores, rankings):
            if remove_query and doc == qid:
                continue
            docs[doc] = score
        n = len(scores)
        run_dict[qid] = {"docs": docs, "min_score": min(scores) if n else 0, "max_score": max(scores) if n else 0}
    return run_dict


def read_trec_run(file):
    run = {}
    with open(file, "r") as f:
        for line in f:
            qid, _, docid, _rank, score, _name = line.strip().split()
            score = float(score)
            entry = run.setdefault(qid, {"docs": {}, "max_score": score, "min_score": score})
            entry["docs"][docid] = score
            entry["min_score"] = score
    return run


def write_trec_run(run, file, name="fusion"):
    with open(file, "w") as f:
        for qid, entry in run.items():
            doc_score = entry["docs"] if "docs" in entry else entry
            ranked = sorted(doc_score.items(), key=lambda kv: kv[1], reverse=True)     # stable: ties keep insertion order
            for i, (doc, score) in enumerate(ranked):
                f.write(f"{qid} Q0 {doc} {i + 1} {score} {name}\n")


def _run_is_f32(run):
    for entry in run.values():
        for v in entry["docs"].values():
            return isinstance(v, np.float32)
    return False


def _fuse_two(runs, weights, device, dtype=np.float32):
    dense, sparse = runs
    qids = list(dict.fromkeys(list(dense.keys()) + list(sparse.keys())))
    for qid in qids:
        if qid not in dense or qid not in sparse:
            raise KeyError(qid)                                  # the reference indexes run[qid] unconditionally
    nq = len(qids)
    k_d = max(1, max((len(dense[q]["docs"]) for q in qids), default=1))
    k_s = max(1, max((len(sparse[q]["docs"]) for q in qids), default=1))
    if k_d > ops.MSR_MAX_K or k_s > ops.MSR_MAX_K:
        raise ValueError(f"run depth above {ops.MSR_MAX_K} is not supported by msr_fuse")
    d_s = np.zeros((nq, k_d), dtype); d_d = np.full((nq, k_d), -1, np.int32); d_c = np.zeros(nq, np.int32)
    s_s = np.zeros((nq, k_s), dtype); s_d = np.full((nq, k_s), -1, np.int32); s_c = np.zeros(nq, np.int32)
    names, name_id = [], {}

    def intern(doc):
        i = name_id.get(doc)
        if i is None:
            i = name_id[doc] = len(names)
            names.append(doc)
        return i

    # min/max of a run dict were taken over the unfiltered list (H2); they travel as two sentinel entries when they
    # are not attained by a surviving doc -- simpler: pass them through explicit arrays
    d_mm = np.zeros((nq, 2), np.float64); s_mm = np.zeros((nq, 2), np.float64)
    for r, qid in enumerate(qids):
        for (run, S, D, C, MM) in ((dense, d_s, d_d, d_c, d_mm), (sparse, s_s, s_d, s_c, s_mm)):
            docs = run[qid]["docs"]
            n = len(docs)
            if n:
                D[r, :n] = [intern(x) for x in docs.keys()]
                S[r, :n] = np.fromiter(docs.values(), dtype=np.float64, count=n)
            C[r] = n
            MM[r] = (run[qid]["min_score"], run[qid]["max_score"])
    return qids, names, (d_s, d_d, d_c, d_mm), (s_s, s_d, s_c, s_mm)


def fuse_arrays(d_scores, d_docs, d_counts, s_scores, s_docs, s_counts, weights, self_ids=None, remove_query=False,
                dense_f32=True):
    """Kernel (d) on CUDA tensors; see ops.fuse_lists."""
    return ops.fuse_lists(d_scores, d_docs, d_counts, s_scores, s_docs, s_counts, self_ids, remove_query,
                          weights[0], weights[1], dense_f32=dense_f32)


def _fuse_impl(runs, weights, device=None):
    if len(runs) != 2 or len(weights) != 2:
        raise NotImplementedError("fuse() covers the reference's use: runs=[dense_run, sparse_run], weights=[alpha, 1-alpha]")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    dense_f32 = _run_is_f32(runs[0])
    if _run_is_f32(runs[1]):
        raise NotImplementedError("the second run must hold Python floats (Lucene hit scores), as in the reference")
    # runs of Python floats (e.g. re-read with read_trec_run) are combined from the doubles themselves (msr_fuse_f64):
    # float('0.1') is not float32(0.1).  np.float32 dense runs keep the float32 / float64 typing of rule H3 (msr_fuse).
    dt = np.float32 if dense_f32 else np.float64
    qids, names, (d_s, d_d, d_c, d_mm), (s_s, s_d, s_c, s_mm) = _fuse_two(runs, weights, dev, dt)
    # min/max over the unfiltered lists: re-attach the removed entry (if any) as a trailing list element whose doc
    # id equals the self id, so the kernel's own min/max (taken before removal) reproduces the run dict's values.
    nq = len(qids)
    self_ids = np.full(nq, -1, np.int32)
    sentinel = len(names)                                        # an id no document uses

    def attach(S, D, C, MM):
        k = S.shape[1]
        S2 = np.zeros((nq, k + 2), dt); D2 = np.full((nq, k + 2), -1, np.int32)
        S2[:, :k] = S; D2[:, :k] = D
        C2 = C.copy()
        for r in range(nq):
            n = int(C[r])
            lo = S[r, :n].min() if n else None
            hi = S[r, :n].max() if n else None
            for v, cur, better in ((MM[r, 0], lo, lambda a, b: a < b), (MM[r, 1], hi, lambda a, b: a > b)):
                if cur is None or better(dt(v), dt(cur)):
                    if n == 0 and v == 0:
                        continue
                    S2[r, C2[r]] = v; D2[r, C2[r]] = sentinel; C2[r] += 1
        return S2, D2, C2

    d_s, d_d, d_c = attach(d_s, d_d, d_c, d_mm)
    s_s, s_d, s_c = attach(s_s, s_d, s_c, s_mm)
    self_ids[:] = sentinel
    t = lambda a: torch.from_numpy(a).to(dev)
    fs, fd, fsrc, fc = ops.fuse_lists(t(d_s), t(d_d), t(d_c), t(s_s), t(s_d), t(s_c), t(self_ids), True,
                                      weights[0], weights[1], dense_f32=dense_f32)
    return qids, names, fs.cpu().numpy(), fd.cpu().numpy(), fsrc.cpu().numpy(), fc.cpu().numpy(), dense_f32


def fuse(runs, weights, device=None):
    qids, names, fs, fd, fsrc, fc, dense_f32 = _fuse_impl(runs, weights, device)
    fused = {}
    for r, qid in enumerate(qids):
        n = int(fc[r])
        if dense_f32:
            fused[qid] = {names[fd[r, j]]: (np.float32(fs[r, j]) if fsrc[r, j] & 1 else float(fs[r, j])) for j in range(n)}
        else:
            fused[qid] = {names[fd[r, j]]: float(fs[r, j]) for j in range(n)}
    return fused


_TAGS = {1: "dense", 2: "sparse", 3: "fuse"}


def fuse_statistic(runs, weights, device=None):
    qids, names, fs, fd, fsrc, fc, dense_f32 = _fuse_impl(runs, weights, device)
    fused = {}
    for r, qid in enumerate(qids):
        n = int(fc[r])
        fused[qid] = {names[fd[r, j]]: ResultRecord((np.float32(fs[r, j]) if (dense_f32 and fsrc[r, j] & 1) else float(fs[r, j])),
                                                    _TAGS[int(fsrc[r, j])]) for j in range(n)}
    return fused


def provenance_histogram(runs, weights, depth=200, device=None):
    """The rank-by-provenance statistic of /root/reference/src/score_statistic.py:393-417 (fuse_statistic, stable sort by
    score, top-``depth``, 1-based ranks appended per tag) computed on the device as a by-product of kernel (d):
    returns {'dense' | 'sparse' | 'fuse': int64 [depth]}, entry r = how many queries have a doc of that tag at rank r+1.
    Expanding entry r into r+1 repeated that many times gives the reference's dense_list / sparse_list / fusion_list
    up to order (they are only ever binned)."""
    if len(runs) != 2 or len(weights) != 2:
        raise NotImplementedError("provenance_histogram covers runs=[dense_run, sparse_run]")
    qids, names, fs, fd, fsrc, fc, dense_f32 = _fuse_impl(runs, weights, device)
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    hist = ops.provenance_hist(torch.from_numpy(fs).to(dev), torch.from_numpy(fsrc).to(dev), torch.from_numpy(fc).to(dev),
                               depth).cpu().numpy()
    return {"dense": hist[0], "sparse": hist[1], "fuse": hist[2]}
