#!/usr/bin/env python
"""Row top-k (topk_rows_kernel<T>) on its own: the quantiser over fp16 activation rows [10000, 131072] (the bench's rows: 32
non-zeros each; and relu'd rows with ~16 % non-zeros), the dense search's threshold sample (1024 x 65536 fp32, rank 23) and
an exact-pass block (64 x 1.25 M fp32, top-100).    python profiles/probe_rows.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mllm_sparse_retrieval_b200 import ops

dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(7)


def timed(fn, reps=5):
    fn(); fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


nq, V = 10000, 131072
x = torch.zeros((nq, V), dtype=torch.float16, device=dev)
cols = torch.randint(0, V, (nq, 32), generator=g, device=dev)
x.scatter_(1, cols, (torch.randint(1, 256, (nq, 32), generator=g, device=dev).float() / 100).half())
ms = timed(lambda: ops.quantise_topk(x, 32))
print(f"quantise_topk fp16 [{nq},{V}] top-32, 32 non-zeros per row: {ms:.3f} ms = {nq * V * 2 / ms / 1e6:.0f} GB/s of row bytes")
for lo in range(0, nq, 1000):
    x[lo:lo + 1000] = torch.log1p(torch.relu(torch.randn((1000, V), generator=g, device=dev) * 2 - 2)).half()
ms = timed(lambda: ops.quantise_topk(x, 128))
print(f"quantise_topk fp16 [{nq},{V}] top-128, relu'd rows (16 % non-zero): {ms:.3f} ms = {nq * V * 2 / ms / 1e6:.0f} GB/s of row bytes")
del x
y = torch.randn((1024, 65536), generator=g, device=dev)
ms = timed(lambda: ops.topk_rows(y, 23))
print(f"topk_rows fp32 [1024,65536] top-23: {ms:.3f} ms = {y.numel() * 4 / ms / 1e6:.0f} GB/s")
del y
z = torch.randn((64, 1_250_000), generator=g, device=dev)
ms = timed(lambda: ops.topk_rows(z, 100))
print(f"topk_rows fp32 [64,1250000] top-100: {ms:.3f} ms = {z.numel() * 4 / ms / 1e6:.0f} GB/s")
