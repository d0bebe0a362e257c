"""GPU parity of the stand-alone top-k (b), fusion (d) and the exact dense pass (c) against the oracle."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,k,kind", [(1000, 10, "normal"), (50000, 100, "normal"), (300000, 1000, "normal"),
                                      (5000, 100, "ties"), (20000, 64, "const"), (70, 100, "normal"),
                                      (4096, 4096, "normal"), (9000, 50, "neg")])
def test_topk_rows(lib_built, n, k, kind):
    from mllm_sparse_retrieval_b200.ops import topk_rows
    rng = np.random.default_rng(n + k)
    nq = 5
    if kind == "normal":
        x = rng.standard_normal((nq, n)).astype(np.float32)
    elif kind == "neg":
        x = -np.abs(rng.standard_normal((nq, n))).astype(np.float32) - 1
    elif kind == "ties":
        x = rng.integers(0, 20, size=(nq, n)).astype(np.float32)
    else:
        x = np.full((nq, n), 3.25, dtype=np.float32)
    x[0, : min(n, 5)] = -np.inf
    s, d, c = topk_rows(torch.from_numpy(x).cuda(), k, col_base=7)
    s, d, c = s.cpu().numpy(), d.cpu().numpy(), c.cpu().numpy()
    for q in range(nq):
        valid = np.isfinite(x[q]) | (x[q] == np.inf)
        order = np.lexsort((np.arange(n), -x[q].astype(np.float64)))
        order = order[valid[order]][:k]
        assert c[q] == len(order)
        np.testing.assert_array_equal(d[q, :len(order)], order + 7)
        np.testing.assert_array_equal(s[q, :len(order)], x[q, order])
        assert (d[q, len(order):] == -1).all()


@pytest.mark.parametrize("n,k,off", [(1001, 10, 1), (40003, 100, 3), (131072, 32, 2), (77, 100, 1), (5, 3, 3)])
def test_topk_rows_unaligned_zero_heavy_rows(lib_built, n, k, off):
    """Rows that do not start on a 16-byte boundary (a column slice of a wider matrix, odd row pitch): the 16-byte loads of
    the row passes must cover head, body and tail exactly once.  90 % exact zeros (relu'd activations): counted in
    registers by the histogram pass, they must still rank (0 > negative values, ties by column)."""
    from mllm_sparse_retrieval_b200.ops import topk_rows
    rng = np.random.default_rng(n * 7 + off)
    nq = 4
    big = rng.standard_normal((nq, n + 5)).astype(np.float32)
    big[rng.random(big.shape) < 0.9] = 0.0
    big[1] = np.where(big[1] > 0, 0.0, big[1])                              # row 1: zeros are the maximum
    big[2, off + 3 % n] = -0.0
    xt = torch.from_numpy(big).cuda()[:, off:off + n]
    assert xt.data_ptr() % 16 != 0 or xt.stride(0) * 4 % 16 != 0
    s, d, c = topk_rows(xt, k)
    s, d, c = s.cpu().numpy(), d.cpu().numpy(), c.cpu().numpy()
    x = big[:, off:off + n]
    for q in range(nq):
        order = np.lexsort((np.arange(n), -x[q].astype(np.float64)))[:k]
        assert c[q] == len(order)
        np.testing.assert_array_equal(s[q, :len(order)], x[q, order])
        # +0.0 and -0.0 compare equal in numpy but not in the kernel's total order (+0 first): compare ids where scores differ
        nz = (x[q, order] != 0) | (q != 2)
        np.testing.assert_array_equal(d[q, :len(order)][nz], order[nz])


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
def test_quantise_topk_unaligned_rows_equal_aligned_rows(lib_built, dtype):
    """msr_quantise_topk on a column slice of a wider activation matrix (rows off the 16-byte grid) == on its contiguous copy."""
    from mllm_sparse_retrieval_b200.ops import quantise_topk
    g = torch.Generator(device="cuda"); g.manual_seed(9)
    nq, V, top = 33, 30001, 48
    big = torch.log1p(torch.relu(torch.randn((nq, V + 7), generator=g, device="cuda") * 2 - 1)).to(dtype)
    for off in (1, 3, 6):
        view = big[:, off:off + V]
        a = quantise_topk(view, top)
        b = quantise_topk(view.contiguous(), top)
        for x, y in zip(a, b):
            n = int(a[0][-1]) if x.dim() == 1 and x.numel() != nq + 1 else x.numel()
            assert torch.equal(x[:n], y[:n])
        assert int(a[0][-1]) > nq * top // 2


def _lists_to_runs(oracle, qids, docnames, scores, docs, counts, remove_query, as_f32):
    batch_scores, batch_rank = [], []
    for q in range(len(qids)):
        n = int(counts[q])
        sc = scores[q, :n]
        batch_scores.append([np.float32(v) for v in sc] if as_f32 else [float(v) for v in sc])
        batch_rank.append([docnames[i] for i in docs[q, :n]])
    return oracle.get_run_dict(qids, batch_scores, batch_rank, remove_query)


@pytest.mark.parametrize("remove_query,alpha,kd,ks", [(False, 0.5, 100, 100), (True, 0.3, 1000, 1000), (True, 0.5, 10, 37)])
def test_fuse_matches_reference_semantics(oracle, lib_built, remove_query, alpha, kd, ks):
    from mllm_sparse_retrieval_b200.ops import fuse_lists
    rng = np.random.default_rng(kd + ks)
    nq, n_docs = 12, 3000
    docnames = [str(i) for i in range(n_docs)]
    qids = [str(i * 7) for i in range(nq)]                 # collide with doc ids on purpose
    d_s = np.zeros((nq, kd), np.float32); d_d = np.full((nq, kd), -1, np.int32); d_c = np.zeros(nq, np.int32)
    s_s = np.zeros((nq, ks), np.float32); s_d = np.full((nq, ks), -1, np.int32); s_c = np.zeros(nq, np.int32)
    for q in range(nq):
        nd = kd if q != 3 else 0                           # one query with an empty dense list
        ns = ks if q != 5 else max(1, ks // 3)             # one short sparse list
        pool = rng.permutation(n_docs)[: kd + ks]
        dd = pool[:nd].copy()
        sd = np.concatenate([pool[: ns // 2], pool[kd: kd + ns - ns // 2]])[:ns]
        if q % 2 == 0 and nd:
            dd[rng.integers(0, nd)] = q * 7                # the query's own id is in the dense list
        if q % 3 == 0 and ns and (q * 7) not in sd:
            sd[rng.integers(0, ns)] = q * 7
        _, ui = np.unique(dd, return_index=True); dd = dd[np.sort(ui)]
        _, ui = np.unique(sd, return_index=True); sd = sd[np.sort(ui)]
        nd, ns = len(dd), len(sd)
        d_d[q, :nd] = dd; d_s[q, :nd] = np.sort(rng.random(nd).astype(np.float32))[::-1]; d_c[q] = nd
        s_d[q, :ns] = sd; s_s[q, :ns] = np.sort(rng.integers(1, 90000, ns).astype(np.float32))[::-1]; s_c[q] = ns
    self_ids = np.array([int(x) for x in qids], np.int32)
    t = lambda a: torch.from_numpy(a).cuda()
    fs, fd, fsrc, fc = fuse_lists(t(d_s), t(d_d), t(d_c), t(s_s), t(s_d), t(s_c), t(self_ids), remove_query, alpha, 1 - alpha)
    fs, fd, fsrc, fc = fs.cpu().numpy(), fd.cpu().numpy(), fsrc.cpu().numpy(), fc.cpu().numpy()
    dense_run = _lists_to_runs(oracle, qids, docnames, d_s, d_d, d_c, remove_query, True)
    sparse_run = _lists_to_runs(oracle, qids, docnames, s_s, s_d, s_c, remove_query, False)
    exp = oracle.fuse([dense_run, sparse_run], [alpha, 1 - alpha])
    for q, qid in enumerate(qids):
        e = exp[qid]
        assert fc[q] == len(e)
        assert [docnames[i] for i in fd[q, :fc[q]]] == list(e.keys())          # insertion order (H1/H4)
        np.testing.assert_allclose(fs[q, :fc[q]], np.array(list(e.values()), dtype=np.float64), rtol=1e-3, atol=1e-6)
        for j, doc in enumerate(e.keys()):
            tag = (1 if doc in dense_run[qid]["docs"] else 0) | (2 if doc in sparse_run[qid]["docs"] else 0)
            assert fsrc[q, j] == tag
    # provenance histogram (score_statistic.py:406-417) of the same fused lists, quantised ties included
    from collections import Counter
    from mllm_sparse_retrieval_b200.ops import fuse_lists as _fl, provenance_hist
    tags = oracle.fuse_tags([dense_run, sparse_run])
    for depth in (200, 7):
        fs_d, fd_d, fsrc_d, fc_d = _fl(t(d_s), t(d_d), t(d_c), t(s_s), t(s_d), t(s_c), t(self_ids), remove_query, alpha, 1 - alpha)
        hist = provenance_hist(fs_d, fsrc_d, fc_d, depth).cpu().numpy()
        lists = oracle.provenance_lists({qid: {doc: (exp[qid][doc], tags[qid][doc]) for doc in exp[qid]} for qid in qids}, depth)
        for row, lst in enumerate(lists):
            c = Counter(lst)
            assert hist[row].tolist() == [c.get(r + 1, 0) for r in range(depth)]


@pytest.mark.parametrize("n_docs,dim,nq,k", [(1000, 4096, 7, 10), (5000, 4096, 33, 100), (300, 100, 5, 1000), (2000, 3584, 4, 100)])
def test_dense_exact_matches_oracle(oracle, lib_built, n_docs, dim, nq, k):
    from mllm_sparse_retrieval_b200 import synth
    from mllm_sparse_retrieval_b200.ops import DenseIndex
    docs, q = synth.make_dense(n_docs, nq, dim=dim, seed=1002)
    docs, q = docs.numpy(), q.numpy()
    docs[11] = docs[4]                                      # exact duplicate -> id tie-break
    idx = DenseIndex(dim)
    idx.add(docs[: n_docs // 2]); idx.add(docs[n_docs // 2:])   # FaissFlatSearcher.add appends
    assert idx.ntotal == n_docs
    s, ids = idx.search(q, k, mode=1)
    es, eids = oracle.flat_ip_search_numpy(docs, q, k)
    kk = min(k, n_docs)
    # ids identical except where the oracle's own neighbours are closer than fp32 resolution
    S64 = q.astype(np.float64) @ docs.astype(np.float64).T
    for r in range(nq):
        if not np.array_equal(ids[r, :kk], eids[r]):
            bad = np.nonzero(ids[r, :kk] != eids[r])[0]
            for b in bad:
                assert abs(S64[r, ids[r, b]] - S64[r, eids[r, b]]) < 2e-6, (r, b)
    np.testing.assert_allclose(s[:, :kk], np.take_along_axis(S64, ids[:, :kk], 1), rtol=1e-3, atol=1e-6)
    if k > n_docs:                                          # D5 padding convention
        assert (ids[:, kk:] == -1).all() and (s[:, kk:] < -3e38).all()
    # device path agrees with host path
    s2, ids2 = idx.search(torch.from_numpy(q).cuda(), k, mode=1)
    np.testing.assert_array_equal(ids2.cpu().numpy(), ids)
    np.testing.assert_array_equal(s2.cpu().numpy(), s)


@pytest.mark.parametrize("n_docs,dim,nq,k", [(5000, 4096, 200, 100), (25000, 1024, 300, 100), (777, 256, 130, 1000),
                                              (3000, 3584, 64, 10), (2048, 64, 129, 5), (30000, 512, 256, 1000),
                                              (70000, 256, 140, 10), (9000, 128, 33, 4096)])
def test_dense_tensor_core_mode_equals_exact_mode(oracle, lib_built, n_docs, dim, nq, k):
    """Mode 2 (tcgen05 fp16 pass with fused threshold selection + certified fp32 re-score) returns the ids AND the bit pattern of the scores of
    mode 1; both match the float64 oracle (ids identical up to fp32-resolution ties, scores within 1e-3 relative)."""
    from mllm_sparse_retrieval_b200 import synth
    from mllm_sparse_retrieval_b200.ops import DenseIndex
    docs, q = synth.make_dense(n_docs, nq, dim=dim, seed=1002, device="cuda")
    docs[11] = docs[4]
    idx = DenseIndex(dim)
    idx.add(docs)
    s1, i1 = idx.search(q, k, mode=1)
    s2, i2 = idx.search(q, k, mode=2)
    n_fb = idx.last_fallback
    s0, i0 = idx.search(q, k, mode=0)                        # auto picks the tensor-core path at this batch size
    torch.testing.assert_close(i2, i1, rtol=0, atol=0)
    torch.testing.assert_close(s2, s1, rtol=0, atol=0)
    torch.testing.assert_close(i0, i1, rtol=0, atol=0)
    assert n_fb <= nq // 4, f"{n_fb} of {nq} queries failed the certificate"
    kk = min(k, n_docs)
    es, eids = oracle.flat_ip_search_numpy(docs.cpu().numpy(), q.cpu().numpy(), kk)
    assert (i2[:, :kk].cpu().numpy() == eids).mean() > 0.999
    np.testing.assert_allclose(s2[:, :kk].cpu().numpy(), es, rtol=1e-3, atol=2e-6)
    # host path of the C ABI
    hs, hi = idx.search(q.cpu().numpy(), k, mode=2)
    np.testing.assert_array_equal(hi, i1.cpu().numpy())


def test_dense_certificate_falls_back_on_near_ties(oracle, lib_built):
    """Thousands of near-duplicate docs: fp16 cannot separate them, the certificate must fail and the exact pass
    must take over -- results still equal mode 1."""
    from mllm_sparse_retrieval_b200.ops import DenseIndex
    g = torch.Generator(device="cuda"); g.manual_seed(5)
    dim, n_docs, nq, k = 512, 6000, 48, 50
    base = torch.randn(dim, generator=g, device="cuda")
    docs = base[None, :] + 1e-4 * torch.randn(n_docs, dim, generator=g, device="cuda")
    docs = docs / docs.norm(dim=1, keepdim=True)
    q = base[None, :] + 0.05 * torch.randn(nq, dim, generator=g, device="cuda")
    q = q / q.norm(dim=1, keepdim=True)
    idx = DenseIndex(dim)
    idx.add(docs)
    s1, i1 = idx.search(q, k, mode=1)
    s2, i2 = idx.search(q, k, mode=2)
    assert idx.last_fallback == nq
    torch.testing.assert_close(i2, i1, rtol=0, atol=0)
    torch.testing.assert_close(s2, s1, rtol=0, atol=0)


def test_dense_clustered_docs_partial_fallback(oracle, lib_built):
    """Caption-group structure (5 near-identical docs per group, groups contiguous, so the strided sample is biased)
    plus a few queries that sit on a dense cluster: whatever mix of certified / fallback queries results, mode 2 must
    equal mode 1 bit for bit, and the bulk of the batch must stay on the tensor-core path."""
    from mllm_sparse_retrieval_b200.ops import DenseIndex
    g = torch.Generator(device="cuda"); g.manual_seed(11)
    dim, groups, nq, k = 1024, 6000, 300, 100
    cent = torch.randn(groups, dim, generator=g, device="cuda")
    docs = cent.repeat_interleave(5, dim=0) + 0.3 * torch.randn(groups * 5, dim, generator=g, device="cuda")
    docs[:5000] = cent[0][None, :] + 1e-5 * torch.randn(5000, dim, generator=g, device="cuda")      # one dense cluster > cap
    docs = docs / docs.norm(dim=1, keepdim=True)
    q = cent[torch.randint(0, groups, (nq,), generator=g, device="cuda")] + 0.5 * torch.randn(nq, dim, generator=g, device="cuda")
    q[:7] = cent[0][None, :] + 0.2 * torch.randn(7, dim, generator=g, device="cuda")               # land on the cluster
    q = q / q.norm(dim=1, keepdim=True)
    idx = DenseIndex(dim); idx.add(docs)
    s1, i1 = idx.search(q, k, mode=1)
    s2, i2 = idx.search(q, k, mode=2)
    assert 7 <= idx.last_fallback <= nq // 4, idx.last_fallback
    torch.testing.assert_close(i2, i1, rtol=0, atol=0)
    torch.testing.assert_close(s2, s1, rtol=0, atol=0)


def test_dense_fp16_overflow_is_never_certified(lib_built):
    """Vectors outside the fp16 range become inf in the tensor-core copy: the error bound is inf, nothing is certified
    and the exact pass answers every query."""
    from mllm_sparse_retrieval_b200.ops import DenseIndex
    g = torch.Generator(device="cuda"); g.manual_seed(12)
    docs = torch.randn(5000, 256, generator=g, device="cuda") * 1e5
    q = torch.randn(40, 256, generator=g, device="cuda")
    idx = DenseIndex(256); idx.add(docs)
    s1, i1 = idx.search(q, 10, mode=1)
    s2, i2 = idx.search(q, 10, mode=2)
    assert idx.last_fallback == 40
    torch.testing.assert_close(i2, i1, rtol=0, atol=0)
    torch.testing.assert_close(s2, s1, rtol=0, atol=0)


def test_tensor_core_scores_match_fp16_matmul(lib_built):
    """The tcgen05 pass alone against a torch fp32 matmul of the fp16-rounded operands (tolerance: fp32 accumulation)."""
    from mllm_sparse_retrieval_b200.ops import DenseIndex
    g = torch.Generator(device="cuda"); g.manual_seed(3)
    for n, nq, dim in [(256, 128, 64), (1000, 37, 128), (4099, 257, 4096)]:
        x = torch.randn(n, dim, generator=g, device="cuda"); x = x / x.norm(dim=1, keepdim=True)
        q = torch.randn(nq, dim, generator=g, device="cuda"); q = q / q.norm(dim=1, keepdim=True)
        idx = DenseIndex(dim); idx.add(x)
        got = idx.tc_scores(q)
        ref = q.half().float() @ x.half().float().T
        torch.testing.assert_close(got, ref, rtol=0, atol=2e-6)


def _csr_row_to_dict(coll, ip, t, w, r):
    lo, hi = int(ip[r]), int(ip[r + 1])
    return {coll.terms[int(a)]: int(b) for a, b in zip(t[lo:hi], w[lo:hi]) if int(b) > 0}


def test_quantise_topk_matches_reference_fixtures(oracle, lib_built):
    """msr_quantise_topk against the fixtures generated from the reference's own get_img_valid_tokens_values and
    query-string / doc-vector builders (tests/golden/reference_glue.json), and against the oracle restatement."""
    import json, os
    from collections import Counter
    from mllm_sparse_retrieval_b200.index import ImpactCollection
    from mllm_sparse_retrieval_b200.ops import quantise_topk
    with open(os.path.join(os.path.dirname(__file__), "golden", "reference_glue.json")) as f:
        golden = json.load(f)
    for c in golden["quantiser"]:
        vocab = c["vocab"]
        logits = np.asarray(c["logits"], dtype=np.float32)
        uniq, cnt = np.unique(logits, return_counts=True)
        top = c["sparse_length"] if c["sparse_manual"] else 128
        top = min(top, len(logits))
        if (cnt > 1).any() and logits[np.argsort(-logits)[top - 1]] in uniq[cnt > 1]:
            continue                                         # torch.topk tie order at the cut is unspecified
        # an index whose dictionary holds every lower-cased token, so nothing is dropped by the dictionary
        names = sorted({t for v in vocab for t in v.lower().split()} | {t[1:] for v in vocab for t in [v.lower()] if len(t) > 1})
        coll = ImpactCollection.from_vectors(["d"], [{n: 1 for n in names if n.split() == [n]}])
        vocab_len = len(vocab)
        tmap = torch.from_numpy(coll.token_to_term_map(vocab, is_filtered=c["is_filtered"])).cuda()
        x = torch.from_numpy(logits)[None, :].cuda()
        for dt in (torch.float32,):
            ip, t, w = quantise_topk(x.to(dt), top, tmap, vocab_len=vocab_len, doc_mode=False)
            got = _csr_row_to_dict(coll, ip.cpu().numpy(), t.cpu().numpy(), w.cpu().numpy(), 0)
            assert got == dict(Counter(c["query_string"].split())), c["sparse_length"]
            ip, t, w = quantise_topk(x.to(dt), top, tmap, vocab_len=vocab_len, doc_mode=True)
            got = _csr_row_to_dict(coll, ip.cpu().numpy(), t.cpu().numpy(), w.cpu().numpy(), 0)
            assert got == {k: v for k, v in c["doc_vector"].items() if v > 0}


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
def test_quantise_topk_feeds_sparse_search(oracle, lib_built, dtype):
    """logits -> query CSR on the device -> msr_sparse_search, equal to the host path (oracle quantiser -> query
    strings -> ImpactCollection.encode_queries -> search)."""
    from mllm_sparse_retrieval_b200.index import ImpactCollection
    from mllm_sparse_retrieval_b200.ops import SparseIndex, quantise_topk
    rng = np.random.default_rng(4)
    V, n_docs, nq, top = 3000, 400, 50, 64
    vocab = [f"Tok{i}" if i % 7 else f"tok{i - 1}" for i in range(V)]        # case-folding collisions: Tok8 / tok8 ...
    vocab_len = V - 5                                                        # ids beyond the dictionary (value-shift quirk)
    vecs = []
    for _ in range(n_docs):
        ids = rng.choice(V - 200, size=30, replace=False)                    # 200 tokens never indexed
        vecs.append({vocab[i].lower(): int(rng.integers(1, 300)) for i in ids})
    coll = ImpactCollection.from_vectors([str(i) for i in range(n_docs)], vecs)
    logits = torch.from_numpy(np.log1p(np.maximum(rng.standard_normal((nq, V)) * 2, 0)).astype(np.float32)).cuda().to(dtype)
    tmap = torch.from_numpy(coll.token_to_term_map(vocab)).cuda()
    ip, t, w = quantise_topk(logits, top, tmap, vocab_len=vocab_len)
    idx = SparseIndex(coll.indptr, coll.post_docs, coll.post_impacts, coll.n_docs)
    s, d, c = idx.search(ip, t, w, 20)
    idx.status()
    # host path through the oracle's restatement of the reference quantiser
    lf = logits.float().cpu().numpy()
    strings = []
    for r in range(nq):
        counts = oracle.quantise_row(lf[r], top, vocab, vocab_len=vocab_len)
        strings.append(" ".join(tok for tok, n in counts.items() for _ in range(n)))
        got = _csr_row_to_dict(coll, ip.cpu().numpy(), t.cpu().numpy(), w.cpu().numpy(), r)
        assert got == {k: v for k, v in counts.items() if k in coll.term_to_id}, r
    qi, qt, qw = coll.encode_queries(strings)
    es, ed, ec = oracle.impact_search_c(coll.indptr, coll.post_docs, coll.post_impacts, coll.n_docs, qi, qt, qw, 20)
    np.testing.assert_array_equal(d.cpu().numpy(), ed)
    np.testing.assert_array_equal(s.cpu().numpy(), es)


@pytest.mark.gpu
def test_recall_hits_kernel_matches_reference_fixtures(lib_built):
    """msr_recall_hits against the counts RecallMetrics.sort_and_count of the reference produced
    (tests/golden/reference_metrics.json): rows are put in the reference's stable score-descending order first."""
    import json, os
    from mllm_sparse_retrieval_b200.ops import recall_hits, RECALL_CUTOFFS
    with open(os.path.join(os.path.dirname(__file__), "golden", "reference_metrics.json")) as f:
        cases = json.load(f)["metrics"]
    for case in cases:
        nq = case["nq"]
        for name in ("dense", "sparse", "fusion"):
            run = case["runs"][name]
            lists = []
            for q in range(nq):
                v = run[str(q)]
                d = v if name == "fusion" else v["docs"]
                lists.append([int(x) for x, _ in sorted(d.items(), key=lambda kv: kv[1], reverse=True)])
            w = max(1, max(len(x) for x in lists))
            docs = torch.full((nq, w + 3), -1, dtype=torch.int32)           # pitch > width: padding must be ignored
            for q, x in enumerate(lists):
                docs[q, :len(x)] = torch.tensor(x, dtype=torch.int32)
            counts = torch.tensor([len(x) for x in lists], dtype=torch.int32)
            indptr, flat = [0], []
            for q in range(nq):
                t = case["targets"][str(q)]
                flat += [int(x) for x in (t if isinstance(t, list) else [t])]
                indptr.append(len(flat))
            hits, n_eval = recall_hits(docs.cuda(), counts.cuda(), torch.tensor(indptr), torch.tensor(flat, dtype=torch.int32))
            assert {str(k): int(h) for k, h in zip(RECALL_CUTOFFS, hits.cpu())} == case["counts"][name], (nq, name)
            assert int(n_eval) == sum(1 for x in lists if x)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_quantise_candidates_matches_reference_fixtures(oracle, lib_built, dtype):
    """msr_quantise_candidates (Q2, candidate branch + empty-caption fallback) against fixtures generated from the
    reference's own get_text_valid_tokens_values + query-string / doc-vector builders; all captions in ONE batch."""
    import json, os
    from collections import Counter
    from mllm_sparse_retrieval_b200.index import ImpactCollection
    from mllm_sparse_retrieval_b200.ops import quantise_candidates
    with open(os.path.join(os.path.dirname(__file__), "golden", "reference_glue.json")) as f:
        cases = [c for c in json.load(f)["text_quantiser"] if not c["sparse_manual"]]
    assert len(cases) >= 7
    by_shape = {}
    for c in cases:
        by_shape.setdefault((len(c["logits"]), len(c["vocab"]), c["is_filtered"], c["num_expended_tokens"]), []).append(c)
    assert sum(k[3] > 0 for k in by_shape) >= 3                 # expansion terms (encode.py:137-152) are covered
    for (V, vlen, filt, n_exp), group in by_shape.items():
        f_ids = group[0]["filtered_ids"]
        vocab = group[0]["vocab"]
        names = sorted({v.lower() for v in vocab} | {v.lower()[1:] for v in vocab if len(v) > 1})
        coll = ImpactCollection.from_vectors(["d"], [{n: 1 for n in names}])
        tmap = torch.from_numpy(coll.token_to_term_map(vocab, is_filtered=filt)).cuda()
        x = torch.tensor([c["logits"] for c in group], dtype=torch.float32).cuda().to(dtype)
        if dtype != torch.float32:                          # the fixture is fp32: compare with the oracle on the rounded row
            exp = []
            for r, c in enumerate(group):
                toks, vals = oracle.quantise_text_row(x[r].float().cpu().numpy(), c["token_ids"], vocab, False, 128, filt,
                                                      filtered_ids=f_ids, num_expended_tokens=n_exp)
                exp.append((oracle.build_query_string(toks, vals), oracle.build_doc_vector(toks, vals)))
        else:
            exp = [(c["query_string"], c["doc_vector"]) for c in group]
        for doc_mode in (False, True):
            ip, t, w = quantise_candidates(x, [c["token_ids"] for c in group], tmap, vocab_len=vlen, doc_mode=doc_mode,
                                           filtered_ids=f_ids if n_exp else None, num_expended_tokens=n_exp)
            ip, t, w = ip.cpu().numpy(), t.cpu().numpy(), w.cpu().numpy()
            assert ip.tolist() == [(128 + min(n_exp, V)) * r for r in range(len(group) + 1)]
            for r, (qs, dv) in enumerate(exp):
                got = _csr_row_to_dict(coll, ip, t, w, r)
                want = {k: v for k, v in dv.items() if v > 0} if doc_mode else dict(Counter(qs.split()))
                assert got == want, (group[r]["text"], doc_mode)


@pytest.mark.gpu
def test_quantise_candidates_rejects_bad_input(lib_built):
    from mllm_sparse_retrieval_b200._lib import MsrError
    from mllm_sparse_retrieval_b200.ops import quantise_candidates
    x = torch.rand(2, 64, device="cuda")
    with pytest.raises(MsrError):
        quantise_candidates(x, [[1, 2], [64]])                       # token id outside the vocabulary
    with pytest.raises(MsrError):
        quantise_candidates(x, [[1], [2]], fallback_top=100)          # fallback larger than the vocabulary
    ip, t, w = quantise_candidates(x[:0], [])
    assert ip.tolist() == [0]
