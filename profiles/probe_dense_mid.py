#!/usr/bin/env python
"""One mid-size dense search (64 queries x 25 000 docs x 4096, top-100) for a launch list."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mllm_sparse_retrieval_b200.ops import DenseIndex

nq = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(1002)
idx = DenseIndex(4096, device=dev)
idx.add(torch.nn.functional.normalize(torch.randn((25000, 4096), generator=g, device=dev), dim=1))
q = torch.nn.functional.normalize(torch.randn((nq, 4096), generator=g, device=dev), dim=1)
for _ in range(3):
    s, i = idx.search(q, 100, mode=2)
torch.cuda.synchronize()
print("ok")
