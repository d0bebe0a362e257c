#!/usr/bin/env python
"""msr_quantise_topk over the bench's activation rows (fp16 [10000, 131072], 32 non-zeros per row) for an ncu capture of
topk_rows_kernel<__half>.    python profiles/probe_quantise.py [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mllm_sparse_retrieval_b200 import ops

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(7)
nq, V = 10000, 131072
x = torch.zeros((nq, V), dtype=torch.float16, device=dev)
cols = torch.randint(0, V, (nq, 32), generator=g, device=dev)
x.scatter_(1, cols, (torch.randint(1, 256, (nq, 32), generator=g, device=dev).float() / 100).half())
for _ in range(reps):
    ops.quantise_topk(x, 32)
torch.cuda.synchronize()
print("ok")
