"""Shared generators for tests (small seeded collections)."""
import numpy as np


def random_csr(rng, n_docs, n_terms, nnz_per_doc, max_impact=400, zipf=True):
    """Canonical term-major CSR (indptr u64, docids u32, impacts u16) of a random collection."""
    rows = []
    p = 1.0 / np.arange(1, n_terms + 1)
    p /= p.sum()
    for d in range(n_docs):
        m = min(n_terms, int(nnz_per_doc if np.isscalar(nnz_per_doc) else rng.integers(*nnz_per_doc)))
        if m == 0:
            continue
        t = rng.choice(n_terms, size=m, replace=False, p=p if zipf else None)
        w = rng.integers(1, max_impact + 1, size=m)
        rows.append(np.stack([t, np.full(m, d), w], 1))
    if rows:
        a = np.concatenate(rows)
        order = np.lexsort((a[:, 1], a[:, 0]))
        a = a[order]
    else:
        a = np.zeros((0, 3), dtype=np.int64)
    indptr = np.zeros(n_terms + 1, dtype=np.uint64)
    np.cumsum(np.bincount(a[:, 0], minlength=n_terms), out=indptr[1:])
    return indptr, a[:, 1].astype(np.uint32), a[:, 2].astype(np.uint16)


def random_queries(rng, nq, n_terms, q_len, max_w=400, dup=False, zipf=True):
    p = 1.0 / np.arange(1, n_terms + 1)
    p /= p.sum()
    indptr = [0]
    terms, weights = [], []
    for _ in range(nq):
        m = int(q_len if np.isscalar(q_len) else rng.integers(*q_len))
        m = min(m, n_terms) if not dup else m
        t = rng.choice(n_terms, size=m, replace=dup, p=p if zipf else None)
        terms.append(t)
        weights.append(rng.integers(1, max_w + 1, size=m))
        indptr.append(indptr[-1] + m)
    cat = lambda xs, dt: (np.concatenate(xs) if xs else np.zeros(0)).astype(dt)
    return np.asarray(indptr, dtype=np.uint64), cat(terms, np.uint32), cat(weights, np.uint32)


def python_jsonl_collection(path):
    """The pure-Python reading of a --sparse_index directory (json.loads + ImpactCollection.from_vectors): the yardstick the
    native builder (C++ reader + GPU inversion) is compared with."""
    import glob
    import json
    import os
    from mllm_sparse_retrieval_b200.index import ImpactCollection
    ids, vecs = [], []
    for fn in sorted(glob.glob(os.path.join(path, "corpus*.jsonl"))):
        with open(fn, "r", encoding="utf-8") as f:
            for line in f:
                line = line.strip()
                if line:
                    obj = json.loads(line)
                    ids.append(str(obj["id"]))
                    vecs.append(obj.get("vector") or {})
    return ImpactCollection.from_vectors(ids, vecs)
