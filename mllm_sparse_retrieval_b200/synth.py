"""Synthetic vocab-sparse collections of the shapes named in BASELINE.json / SURVEY.md 8(d).

There is no network, no checkpoint and no COCO/Flickr test CSV in this environment, so every config is a
seeded stand-in with the statistics of the reference's encoder output (src/encode.py:56-153 writes at most
``sparse_length``=128 (token, int weight) pairs per vector, weights = rint(100*log1p(relu(logit)))):

  * term of each slot drawn WITHOUT replacement per vector from Zipf(s=1) over ranks 1..V (successive sampling:
    draw with replacement, keep the first ``nnz`` distinct values), rank -> term id by a fixed permutation;
  * impacts / query weights = clip(round(100*log1p(Exp(mean 1.5))), 1, 400);
  * external docids are decimal strings; internal doc id = rank of the string in UTF-8 byte order (S8), so
    string order != numeric order ("100" < "99").

Everything is torch so the 1M-doc configs are generated on the GPU in the bench (plumbing, not the product).
"""
from __future__ import annotations

import dataclasses

import numpy as np
import torch

V_DEFAULT = 131072


@dataclasses.dataclass
class SparseCSR:
    """Canonical term-major CSR: postings of term t = docids[indptr[t]:indptr[t+1]] (ascending) + impacts."""
    n_docs: int
    n_terms: int
    indptr: torch.Tensor    # int64 [n_terms+1]
    docids: torch.Tensor    # int32 [nnz]
    impacts: torch.Tensor   # int16 (bit pattern of u16) [nnz]

    def numpy(self):
        return (self.indptr.cpu().numpy().astype(np.uint64),
                self.docids.cpu().numpy().view(np.uint32) if self.docids.dtype == torch.int32
                else self.docids.cpu().numpy().astype(np.uint32),
                self.impacts.cpu().numpy().view(np.uint16))


@dataclasses.dataclass
class QueryCSR:
    indptr: torch.Tensor    # int64 [nq+1]
    terms: torch.Tensor     # int32 [nnz]
    weights: torch.Tensor   # int32 [nnz]

    @property
    def nq(self):
        return self.indptr.numel() - 1

    def numpy(self):
        return (self.indptr.cpu().numpy().astype(np.uint64), self.terms.cpu().numpy().astype(np.uint32),
                self.weights.cpu().numpy().astype(np.uint32))

    def slice(self, lo, hi):
        a, b = int(self.indptr[lo]), int(self.indptr[hi])
        return QueryCSR(self.indptr[lo:hi + 1] - a, self.terms[a:b], self.weights[a:b])


def _zipf_cdf(V, device):
    p = 1.0 / torch.arange(1, V + 1, dtype=torch.float64, device=device)
    return torch.cumsum(p / p.sum(), 0)


def _draw_distinct(n_rows, nnz, V, cdf, perm, gen, device, oversample=2):
    """[n_rows, nnz] term ids, -1 where fewer than nnz distinct values were drawn."""
    M = max(nnz * oversample, nnz + 16)
    u = torch.rand((n_rows, M), generator=gen, device=device, dtype=torch.float64)
    ranks = torch.searchsorted(cdf, u.reshape(-1)).clamp_(max=V - 1).reshape(n_rows, M)
    # first occurrence in draw order: stable sort by value, flag run heads, scatter back
    sv, si = torch.sort(ranks, dim=1, stable=True)
    head = torch.ones_like(sv, dtype=torch.bool)
    head[:, 1:] = sv[:, 1:] != sv[:, :-1]
    first = torch.zeros_like(head)
    first.scatter_(1, si, head)
    order = torch.cumsum(first.to(torch.int32), 1)
    keep = first & (order <= nnz)
    out = torch.full((n_rows, nnz), -1, dtype=torch.int64, device=device)
    rows = torch.arange(n_rows, device=device).unsqueeze(1).expand(n_rows, M)
    out[rows[keep], (order[keep] - 1).long()] = perm[ranks[keep]]
    return out


def _draw_weights(shape, gen, device):
    e = -1.5 * torch.log1p(-torch.rand(shape, generator=gen, device=device, dtype=torch.float32))
    return torch.clamp(torch.round(100.0 * torch.log1p(e)), 1, 400).to(torch.int32)


def make_collection(n_docs, nnz_per_doc=128, V=V_DEFAULT, seed=1004, device="cpu", chunk_docs=65536, perm_seed=None) -> SparseCSR:
    """perm_seed: seed of the rank -> term-id permutation when it must be shared between independently drawn shards of
    one collection (make_queries(seed=perm_seed) then uses the same permutation); default = ``seed``."""
    device = torch.device(device)
    gen = torch.Generator(device=device)
    gen.manual_seed(seed if perm_seed is None else perm_seed)
    cdf = _zipf_cdf(V, device)
    perm = torch.randperm(V, generator=gen, device=device)
    if perm_seed is not None:
        gen.manual_seed(seed)
    keys, imps = [], []
    for lo in range(0, n_docs, chunk_docs):
        n = min(chunk_docs, n_docs - lo)
        terms = _draw_distinct(n, nnz_per_doc, V, cdf, perm, gen, device)
        w = _draw_weights(terms.shape, gen, device)
        docs = torch.arange(lo, lo + n, device=device, dtype=torch.int64).unsqueeze(1).expand_as(terms)
        ok = terms >= 0
        keys.append(terms[ok] * n_docs + docs[ok])       # sort key (term, doc)
        imps.append(w[ok].to(torch.int16))
    key = torch.cat(keys)
    imp = torch.cat(imps)
    del keys, imps
    key, order = torch.sort(key)
    imp = imp[order]
    del order
    term = torch.div(key, n_docs, rounding_mode="floor")
    doc = (key - term * n_docs).to(torch.int32)
    del key
    df = torch.bincount(term, minlength=V)
    indptr = torch.zeros(V + 1, dtype=torch.int64, device=device)
    indptr[1:] = torch.cumsum(df, 0)
    return SparseCSR(n_docs, V, indptr, doc, imp)


def make_queries(nq, q_len=32, V=V_DEFAULT, seed=1004, device="cpu", chunk=65536) -> QueryCSR:
    device = torch.device(device)
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    cdf = _zipf_cdf(V, device)
    perm = torch.randperm(V, generator=gen, device=device)   # same permutation as make_collection(seed)
    gen.manual_seed(seed + 7919)
    t_all, w_all, n_all = [], [], []
    for lo in range(0, nq, chunk):
        n = min(chunk, nq - lo)
        terms = _draw_distinct(n, q_len, V, cdf, perm, gen, device, oversample=4)
        w = _draw_weights(terms.shape, gen, device)
        ok = terms >= 0
        t_all.append(terms[ok].to(torch.int32))
        w_all.append(w[ok])
        n_all.append(ok.sum(1))
    cnt = torch.cat(n_all)
    indptr = torch.zeros(nq + 1, dtype=torch.int64, device=device)
    indptr[1:] = torch.cumsum(cnt, 0)
    return QueryCSR(indptr, torch.cat(t_all), torch.cat(w_all))


def docid_strings(n_docs):
    """External docids for internal ids 0..n_docs-1: decimal strings of 0..n_docs-1 in string (UTF-8) order."""
    return sorted(str(i) for i in range(n_docs))


def make_dense(n_docs, nq, dim=4096, seed=1002, device="cpu", groups=None, dtype=torch.float32):
    """Unit-norm vectors = normalise(N(0,1) + 0.3*sqrt(dim)*centroid[group]) so near-ties exist (SURVEY 8(d) cfg 2)."""
    device = torch.device(device)
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    groups = groups or max(1, n_docs // 5)
    cent = torch.randn((groups, dim), generator=gen, device=device)
    cent = cent / cent.norm(dim=1, keepdim=True)
    scale = 0.3 * dim ** 0.5

    def draw(n):
        g = torch.randint(0, groups, (n,), generator=gen, device=device)
        x = torch.randn((n, dim), generator=gen, device=device) + scale * cent[g]
        return (x / x.norm(dim=1, keepdim=True)).to(dtype)

    return draw(n_docs), draw(nq)
