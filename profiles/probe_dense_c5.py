#!/usr/bin/env python
"""Dense flat-IP at the per-GPU shape of BASELINE.json configs[4] (1.25 M docs x 4096, batches of 1024 queries) for an ncu
capture of the fused tcgen05 pass.    python profiles/probe_dense_c5.py [n_docs] [nq] [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mllm_sparse_retrieval_b200.ops import DenseIndex

n_docs = int(sys.argv[1]) if len(sys.argv) > 1 else 1_250_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(1005)
idx = DenseIndex(4096, device=dev)
cent = torch.randn((4096, 4096), generator=g, device=dev)
for lo in range(0, n_docs, 65536):
    n = min(65536, n_docs - lo)
    x = torch.randn((n, 4096), generator=g, device=dev) + 0.3 * cent[torch.randint(0, 4096, (n,), generator=g, device=dev)]
    idx.add(torch.nn.functional.normalize(x, dim=1))
q = torch.nn.functional.normalize(torch.randn((nq, 4096), generator=g, device=dev) + 0.3 * cent[:nq % 4096 or 4096][:nq], dim=1)
idx.profile(True)
for _ in range(reps):
    s, i = idx.search(q, 100, mode=2)
torch.cuda.synchronize()
tc_ms, rs_ms, n_blk = idx.profile(False)
flops = 2.0 * nq * n_docs * 4096 * reps
print(f"ok n_docs={n_docs} nq={nq} fused tc pass {tc_ms / reps:.3f} ms/search = {flops / (tc_ms * 1e-3) / 1e12:.1f} TFLOP/s, rescore {rs_ms / reps:.3f} ms, fallback {idx.last_fallback}")
