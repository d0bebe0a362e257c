"""msr_dense_search_static -- the exact dense search with its scratch in libmsr.so's file-scope device array, for callers
that may not call cudaMalloc (CUDA graph capture, tight serving loops).  No counterpart in the reference.
Checked against the existing entry point (msr_dense_search, mode 1, device path) through
the C ABI: same kernels, so ids AND score bit patterns must be equal -- on two calls in a row, because every call shares
the one array and finds the previous call's contents in it.

Named test_zz_* so that pytest collects it last (`pytest --collect-only -q`)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _raw_static(idx, q, k, s, i):
    """The call a C host makes: plain device pointers and sizes, no wrapper in between."""
    from mllm_sparse_retrieval_b200 import _lib
    lib = _lib.load()
    stream = ctypes.c_void_p(torch.cuda.current_stream(q.device).cuda_stream)
    return lib.msr_dense_search_static(idx._h, ctypes.c_uint32(q.shape[0]), ctypes.c_void_p(q.data_ptr()), ctypes.c_uint32(k),
                                       ctypes.c_void_p(s.data_ptr()), ctypes.c_void_p(i.data_ptr()), stream)


@pytest.mark.parametrize("n_docs,dim,k", [(30011, 96, 10), (5000, 50, 100), (257, 64, 300)])
def test_static_scratch_search_equals_existing_entry_point_twice(lib_built, n_docs, dim, k):
    from mllm_sparse_retrieval_b200.ops import DenseIndex
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(77 + n_docs)
    docs = torch.randn((n_docs, dim), generator=g, device=dev)
    docs = docs / docs.norm(dim=1, keepdim=True)
    docs[11] = docs[4]                                              # an exact tie: order by id
    idx = DenseIndex(dim, device=dev, doc_base=1000)
    idx.add(docs)
    # two different batches (the second smaller: rows of the first stay behind in the array), then the first again
    batches = [torch.randn((37, dim), generator=g, device=dev), torch.randn((5, dim), generator=g, device=dev)]
    batches.append(batches[0].clone())
    want = [idx.search(q, k, mode=1) for q in batches]              # existing entry point: msr_dense_search, MSR_DEVICE
    torch.cuda.synchronize()
    free_before = None
    for call, (q, (ws, wi)) in enumerate(zip(batches, want)):
        s = torch.full((q.shape[0], k), 7.0, dtype=torch.float32, device=dev)
        i = torch.full((q.shape[0], k), -7, dtype=torch.int64, device=dev)
        torch.cuda.synchronize()
        if call == 1:                                               # kernels are loaded by now: free memory must not drop
            free_before = torch.cuda.mem_get_info(dev)[0]
        rc = _raw_static(idx, q, k, s, i)
        assert rc == 0
        torch.cuda.synchronize()
        if call == 1:
            assert torch.cuda.mem_get_info(dev)[0] >= free_before, "the call allocated device memory"
        torch.testing.assert_close(i, wi, rtol=0, atol=0)
        assert torch.equal(s.view(torch.int32), ws.view(torch.int32)), f"call {call}: score bits differ"
    kk = min(k, n_docs)
    assert (want[0][1][:, :kk] >= 1000).all() and (want[0][1][:, kk:] == -1).all()      # doc_base applied, D5 padding
    # the wrapper over the same symbol
    s, i = idx.search_static(batches[1], k, (torch.empty((5, k), dtype=torch.float32, device=dev),
                                             torch.empty((5, k), dtype=torch.int64, device=dev)))
    torch.testing.assert_close(i, want[1][1], rtol=0, atol=0)
    torch.testing.assert_close(s, want[1][0], rtol=0, atol=0)


def test_static_scratch_search_rejects_what_does_not_fit(lib_built):
    """Four score rows of the index must fit MSR_STATIC_SCRATCH_BYTES (32 MiB): an index of 2.2 M docs is refused with a
    message that names the sizes, and nothing is launched (the outputs stay untouched)."""
    from mllm_sparse_retrieval_b200 import _lib
    from mllm_sparse_retrieval_b200.ops import DenseIndex
    dev = torch.device("cuda", 0)
    n_docs, dim, k = 2_200_000, 4, 10
    idx = DenseIndex(dim, device=dev)
    idx.add(torch.ones((n_docs, dim), device=dev))
    q = torch.ones((3, dim), device=dev)
    s = torch.full((3, k), 7.0, dtype=torch.float32, device=dev)
    i = torch.full((3, k), -7, dtype=torch.int64, device=dev)
    rc = _raw_static(idx, q, k, s, i)
    torch.cuda.synchronize()
    msg = _lib.load().msr_last_error().decode()
    assert rc != 0 and "MSR_STATIC_SCRATCH_BYTES" in msg and str(n_docs) in msg, msg
    assert (s == 7.0).all() and (i == -7).all()
    with pytest.raises(_lib.MsrError):
        idx.search_static(q, k, (s, i))
    # the allocating entry point still answers
    ws, wi = idx.search(q, k, mode=1)
    assert wi.shape == (3, k) and np.isfinite(ws.cpu().numpy()).all()
