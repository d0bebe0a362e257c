#!/usr/bin/env python
"""Dense flat-IP at the reference's own operating point (scripts/search.sh: batch 2-4, depth 100 / 1000) over the COCO-5K
index shape (25 000 x 4096 fp32 = 410 MB), L2 flushed before every search.    python profiles/probe_dense_small.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mllm_sparse_retrieval_b200.ops import DenseIndex

dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(1002)
n_docs, dim = 25000, 4096
x = torch.nn.functional.normalize(torch.randn((n_docs, dim), generator=g, device=dev), dim=1)
idx = DenseIndex(dim, device=dev); idx.add(x)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for nq in (1, 2, 4, 16, 64):
    q = torch.nn.functional.normalize(torch.randn((nq, dim), generator=g, device=dev), dim=1)
    for k in (100, 1000):
        out = (torch.empty((nq, k), dtype=torch.float32, device=dev), torch.empty((nq, k), dtype=torch.int64, device=dev))
        for _ in range(3):
            idx.search(q, k, mode=0, out=out)
        ts = []
        for _ in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); idx.search(q, k, mode=0, out=out); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        ts.sort()
        us = ts[len(ts) // 2]
        print(f"nq={nq:3d} k={k:4d}: {us:8.1f} us cold  = {n_docs * dim * 4 / us / 1e3:7.0f} GB/s of fp32 index bytes (mode {'tc' if nq >= 16 else 'exact'})")
