#!/usr/bin/env python
"""Small-batch probe (the reference's operating point: batch 1-4, depth 10/1000 -- scripts/search*.sh): a few searches of
nq queries at depth k against the config-4 index, for an ncu launch list / full capture.
    python profiles/probe_small.py [nq] [k] [reps] [n_docs]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mllm_sparse_retrieval_b200 import synth
from mllm_sparse_retrieval_b200.ops import SparseIndex

nq = int(sys.argv[1]) if len(sys.argv) > 1 else 4
k = int(sys.argv[2]) if len(sys.argv) > 2 else 100
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
n_docs = int(sys.argv[4]) if len(sys.argv) > 4 else 1_000_000
dev = torch.device("cuda", 0)
coll = synth.make_collection(n_docs, 128, V=131072, seed=1004, device=dev)
qs = synth.make_queries(max(nq, 8), 32, V=131072, seed=1004, device=dev).slice(0, nq)
idx = SparseIndex(coll.indptr, coll.docids, coll.impacts, coll.n_docs, device=dev)
for _ in range(reps):
    s, d, c = idx.search(qs.indptr, qs.terms, qs.weights, k)
torch.cuda.synchronize()
idx.status()
print("ok", nq, k, int(c.sum()))
