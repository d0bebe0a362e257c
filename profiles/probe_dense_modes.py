#!/usr/bin/env python
"""Exact pass (mode 1) against the tensor-core path (mode 2) at small batches over a large shard: which one should mode 0
pick?    python profiles/probe_dense_modes.py [n_docs] [dim]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mllm_sparse_retrieval_b200.ops import DenseIndex

n_docs = int(sys.argv[1]) if len(sys.argv) > 1 else 1_250_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(1005)
idx = DenseIndex(dim, device=dev)
for lo in range(0, n_docs, 65536):
    n = min(65536, n_docs - lo)
    idx.add(torch.nn.functional.normalize(torch.randn((n, dim), generator=g, device=dev), dim=1))
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for nq in (1, 2, 4, 8):
    q = torch.nn.functional.normalize(torch.randn((nq, dim), generator=g, device=dev), dim=1)
    res = {}
    for mode in (1, 2, 0):
        out = (torch.empty((nq, 100), dtype=torch.float32, device=dev), torch.empty((nq, 100), dtype=torch.int64, device=dev))
        for _ in range(2):
            idx.search(q, 100, mode=mode, out=out)
        ts = []
        for _ in range(5):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); idx.search(q, 100, mode=mode, out=out); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        res[mode] = (sorted(ts)[2], out[1].clone(), out[0].clone())
    same = torch.equal(res[1][1], res[2][1]) and torch.equal(res[1][2], res[2][2])
    print(f"n_docs={n_docs} dim={dim} nq={nq}: exact {res[1][0]:8.1f} us, tensor-core {res[2][0]:8.1f} us, auto {res[0][0]:8.1f} us, fallback {idx.last_fallback}, identical={same}")
