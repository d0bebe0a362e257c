import sys, torch, numpy as np
sys.path.insert(0, '.')
from mllm_sparse_retrieval_b200.ops import DenseIndex
torch.manual_seed(0)
for (n, nq, dim) in [(256, 128, 64), (1000, 37, 128), (5000, 300, 4096), (25000, 1024, 4096)]:
    x = torch.randn(n, dim, device='cuda'); x = x / x.norm(dim=1, keepdim=True)
    q = torch.randn(nq, dim, device='cuda'); q = q / q.norm(dim=1, keepdim=True)
    di = DenseIndex(dim); di.add(x)
    s = di.tc_scores(q); torch.cuda.synchronize()
    ref = (q.half().float() @ x.half().float().T)
    err = (s - ref).abs().max().item()
    print(n, nq, dim, 'max abs err vs fp16-input fp32 matmul', err, 'ref absmax', ref.abs().max().item(), flush=True)
    assert err < 2e-4, err
    # timing
    if n >= 25000:
        for _ in range(3): di.tc_scores(q)
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(10): di.tc_scores(q)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print('ms', ms, 'TFLOP/s', 2 * n * nq * dim / ms / 1e9)
print('ok')
