"""Host side of the sparse index: JSONL collection -> term dictionary + canonical CSR (+ on-disk cache).

Replaces the offline Lucene indexing step of the reference (scripts/sparse_index.sh:12-18:
``python -m pyserini.index.lucene --collection JsonVectorCollection --impact --pretokenized``) for the files
that src/encode.py:353-359,426 writes: one JSON object per line, ``{"id": ..., "content": "", "vector": {term: int}}``.

Rules implemented here:
  S1  tf(t,d) = vector[t]; non-positive weights are never indexed (Anserini repeats the key ``value`` times).
      The pretokenized pseudo-document is whitespace-analysed, so a key holding whitespace contributes each of its
      pieces (and an empty key nothing).
  S2  terms are taken verbatim (no stemming / case folding).
  S8  internal doc id = rank of the external docid string in UTF-8 byte order, so the kernels' "smaller id wins"
      equals Anserini's BREAK_SCORE_TIES_BY_DOCID.
"""
from __future__ import annotations

import glob
import json
import os
from collections import Counter

import numpy as np

CACHE_DIR = "msr_index"
CACHE_VERSION = 2


class StringTable:
    """Read-only list of strings kept as ONE utf-8 blob + offsets (+ an optional order): item i is decoded when it is asked
    for.  A 10 M-entry docid table is two NumPy arrays instead of 10 M Python objects; the cache stores it as is."""

    def __init__(self, blob, offsets, order=None):
        self.blob = blob if isinstance(blob, (bytes, bytearray, memoryview)) else np.ascontiguousarray(blob, dtype=np.uint8).tobytes()
        self.offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        self.order = None if order is None else np.ascontiguousarray(order, dtype=np.uint32)

    @classmethod
    def from_list(cls, items):
        enc = [str(x).encode("utf-8") for x in items]
        off = np.zeros(len(enc) + 1, dtype=np.uint64)
        np.cumsum([len(e) for e in enc], out=off[1:])
        return cls(b"".join(enc), off)

    def __len__(self):
        return len(self.order) if self.order is not None else len(self.offsets) - 1

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        i = int(i)
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError(i)
        j = int(self.order[i]) if self.order is not None else i
        return self.blob[int(self.offsets[j]):int(self.offsets[j + 1])].decode("utf-8")

    def __iter__(self):
        return (self[i] for i in range(len(self)))

    def __eq__(self, other):
        return len(self) == len(other) and all(a == b for a, b in zip(self, other))

    def tolist(self):
        return list(self)


class ImpactCollection:
    """Term dictionary + docid table + canonical term-major CSR on the host.  ``terms`` / ``docids`` are lists or
    ``StringTable``s; the string -> id dictionaries are built on first use (a search over pre-encoded queries needs the
    term dictionary only, a run file the id -> string direction only)."""

    def __init__(self, terms, docids, indptr, post_docs, post_impacts):
        self.terms = terms if isinstance(terms, StringTable) else list(terms)            # term id -> string
        self.docids = docids if isinstance(docids, StringTable) else list(docids)        # internal id -> external docid (sorted)
        self._term_to_id = None
        self._docid_to_internal = None
        self.indptr = np.ascontiguousarray(indptr, dtype=np.uint64)
        self.post_docs = np.ascontiguousarray(post_docs, dtype=np.uint32)
        self.post_impacts = np.ascontiguousarray(post_impacts, dtype=np.uint16)

    @property
    def term_to_id(self):
        if self._term_to_id is None:
            self._term_to_id = {t: i for i, t in enumerate(self.terms)}
        return self._term_to_id

    @property
    def docid_to_internal(self):
        if self._docid_to_internal is None:
            self._docid_to_internal = {d: i for i, d in enumerate(self.docids)}
        return self._docid_to_internal

    @property
    def n_docs(self):
        return len(self.docids)

    @property
    def n_terms(self):
        return len(self.terms)

    @property
    def df(self):
        return np.diff(self.indptr.astype(np.int64)).astype(np.uint32)

    # ------------------------------------------------------------------------------------------------------
    @classmethod
    def from_vectors(cls, doc_ids, vectors):
        """In-memory collections (tests, a handful of vectors produced on the fly): doc_ids: iterable of str; vectors:
        iterable of {term: int} (the JSONL ``vector`` objects).  Pure Python and meant for small inputs -- collections
        on disk go through ``from_jsonl_dir`` (C++ reader + GPU inversion)."""
        doc_ids = [str(d) for d in doc_ids]
        if len(set(doc_ids)) != len(doc_ids):
            raise ValueError("duplicate document ids in the collection")
        order = sorted(range(len(doc_ids)), key=lambda i: doc_ids[i].encode("utf-8"))       # S8
        internal = np.empty(len(doc_ids), dtype=np.int64)
        internal[order] = np.arange(len(doc_ids))
        term_to_id = {}
        d_list, t_list, w_list = [], [], []
        for i, vec in enumerate(vectors):
            tf = Counter()
            for key, v in vec.items():
                v = int(v)
                if v <= 0:
                    continue                                                               # S1
                for piece in str(key).split():
                    tf[piece] += v
            for term, v in tf.items():
                if v > 65535:
                    raise ValueError(f"impact {v} of term {term!r} in doc {doc_ids[i]!r} exceeds 65535")
                tid = term_to_id.setdefault(term, len(term_to_id))
                d_list.append(internal[i]); t_list.append(tid); w_list.append(v)
        n_terms = max(1, len(term_to_id))
        d = np.asarray(d_list, dtype=np.int64); t = np.asarray(t_list, dtype=np.int64); w = np.asarray(w_list, dtype=np.int64)
        o = np.lexsort((d, t))
        d, t, w = d[o], t[o], w[o]
        indptr = np.zeros(n_terms + 1, dtype=np.uint64)
        np.cumsum(np.bincount(t, minlength=n_terms), out=indptr[1:])
        terms = [None] * len(term_to_id)
        for s, i in term_to_id.items():
            terms[i] = s
        if not terms:
            terms = [""]
        return cls(terms, [doc_ids[i] for i in order], indptr, d, w)

    @classmethod
    def from_jsonl_files(cls, files, device=None, host_inversion=False):
        """Native builder: libmsr's C++ JSONL reader (msr_jsonl_parse: dictionary, docid table,
        doc-major postings under rules S1/S2/S8) followed by the inversion to the canonical term-major CSR -- on the GPU
        (msr_invert_docs: key build + radix sort + histogram scan).  Same numbering as ``from_vectors`` on the parsed
        lines.  Needs a CUDA device, like everything else here; ``host_inversion=True`` is the explicit opt-in for tooling
        on a box without one (a NumPy lexsort instead of msr_invert_docs -- index building only, never the search path)."""
        import ctypes
        from . import _lib
        lib = _lib.load()
        files = [os.fspath(f) for f in files]
        arr = (ctypes.c_char_p * len(files))(*[f.encode() for f in files])
        h = ctypes.c_void_p()
        _lib.check(lib.msr_jsonl_parse(arr, len(files), ctypes.byref(h)))
        try:
            info = (ctypes.c_uint64 * 5)()
            _lib.check(lib.msr_collection_info(h, info))
            n_docs, n_terms, nnz, db, tb = (int(x) for x in info)
            doc_indptr = np.empty(n_docs + 1, dtype=np.uint64)
            p_terms = np.empty(nnz, dtype=np.uint32); p_imps = np.empty(nnz, dtype=np.uint16)
            rank = np.empty(n_docs, dtype=np.uint32)
            d_blob = np.empty(max(db, 1), dtype=np.uint8); d_off = np.empty(n_docs + 1, dtype=np.uint64)
            t_blob = np.empty(max(tb, 1), dtype=np.uint8); t_off = np.empty(n_terms + 1, dtype=np.uint64)
            _lib.check(lib.msr_collection_export(h, _lib.ptr(doc_indptr), _lib.ptr(p_terms), _lib.ptr(p_imps), _lib.ptr(rank),
                                                 _lib.ptr(d_blob), _lib.ptr(d_off), _lib.ptr(t_blob), _lib.ptr(t_off)))
        finally:
            lib.msr_collection_destroy(h)
        # string tables stay blobs: docids in internal (sorted) order through `order` = inverse of the parser's rank[]
        order = np.empty(n_docs, dtype=np.uint32)
        order[rank] = np.arange(n_docs, dtype=np.uint32)
        docids = StringTable(d_blob[:db], d_off, order)
        terms = StringTable(t_blob[:tb], t_off)
        if n_terms == 0 or n_docs == 0:
            return cls(terms if n_terms else [""], docids, np.zeros(max(n_terms, 1) + 1, dtype=np.uint64), np.zeros(0, np.uint32),
                       np.zeros(0, np.uint16))
        import torch
        if host_inversion:
            d = np.repeat(rank.astype(np.int64), np.diff(doc_indptr.astype(np.int64)))
            t = p_terms.astype(np.int64)
            o = np.lexsort((d, t))
            indptr = np.zeros(n_terms + 1, dtype=np.uint64)
            np.cumsum(np.bincount(t, minlength=n_terms), out=indptr[1:])
            return cls(terms, docids, indptr, d[o], p_imps[o])
        if not torch.cuda.is_available():
            raise RuntimeError("ImpactCollection.from_jsonl_files inverts the collection on the GPU (msr_invert_docs); no CUDA "
                               "device is visible (host_inversion=True opts into the NumPy inversion for tooling)")
        dev = torch.cuda.current_device() if device is None else (torch.device(device).index or 0)
        indptr = np.empty(n_terms + 1, dtype=np.uint64)
        o_docs = np.empty(nnz, dtype=np.uint32); o_imps = np.empty(nnz, dtype=np.uint16)
        with torch.cuda.device(dev):
            _lib.check(lib.msr_invert_docs(dev, n_docs, n_terms, _lib.ptr(doc_indptr), _lib.ptr(p_terms), _lib.ptr(p_imps),
                                           _lib.ptr(rank), _lib.ptr(indptr), _lib.ptr(o_docs), _lib.ptr(o_imps), 0,
                                           _lib.current_stream_ptr(dev)))
        return cls(terms, docids, indptr, o_docs, o_imps)

    @classmethod
    def from_jsonl_dir(cls, path, use_cache=True, host_inversion=False):
        """path = the reference's ``--sparse_index`` directory (holds corpus_*.jsonl; Lucene's own ``index/`` is ignored).
        Built through libmsr (C++ reader + GPU inversion) and cached under ``<path>/msr_index/`` as raw arrays."""
        files = sorted(glob.glob(os.path.join(path, "corpus*.jsonl")))
        if not files:
            raise FileNotFoundError(f"no corpus*.jsonl under {path}")
        stamp = [[os.path.basename(f), os.path.getsize(f), int(os.path.getmtime(f))] for f in files]
        cdir = os.path.join(path, CACHE_DIR)
        if use_cache:
            got = cls._load_cache(cdir, stamp)
            if got is not None:
                return got
        coll = cls.from_jsonl_files(files, host_inversion=host_inversion)
        if use_cache:
            try:
                coll._save_cache(cdir, stamp)
            except OSError:
                pass                                     # read-only index directory: rebuild next time
        return coll

    # On-disk cache: raw little-endian arrays, loaded with np.fromfile -- no JSON string lists, no
    # Python dictionaries:  indptr.u64  docids.u32  impacts.u16  df.u32  terms.blob + terms.off.u64
    #                       docnames.blob + docnames.off.u64 (+ docnames.order.u32)  meta.json
    def _save_cache(self, cdir, stamp):
        os.makedirs(cdir, exist_ok=True)
        w = lambda name, a: np.ascontiguousarray(a).tofile(os.path.join(cdir, name))
        w("indptr.u64", self.indptr); w("docids.u32", self.post_docs); w("impacts.u16", self.post_impacts); w("df.u32", self.df)
        for name, tab in (("terms", self.terms), ("docnames", self.docids)):
            tab = tab if isinstance(tab, StringTable) else StringTable.from_list(tab)
            with open(os.path.join(cdir, name + ".blob"), "wb") as f:
                f.write(tab.blob)
            w(name + ".off.u64", tab.offsets)
            op = os.path.join(cdir, name + ".order.u32")
            if tab.order is not None:
                w(name + ".order.u32", tab.order)
            elif os.path.exists(op):
                os.remove(op)
        with open(os.path.join(cdir, "meta.json"), "w") as f:
            json.dump({"version": CACHE_VERSION, "source": stamp, "n_docs": self.n_docs, "n_terms": self.n_terms,
                       "nnz": int(self.indptr[-1])}, f)

    @classmethod
    def _load_cache(cls, cdir, stamp):
        try:
            with open(os.path.join(cdir, "meta.json")) as f:
                meta = json.load(f)
            if meta.get("version") != CACHE_VERSION or meta.get("source") != stamp:
                return None
            r = lambda name, dt: np.fromfile(os.path.join(cdir, name), dtype=dt)

            def table(name):
                with open(os.path.join(cdir, name + ".blob"), "rb") as f:
                    blob = f.read()
                op = os.path.join(cdir, name + ".order.u32")
                return StringTable(blob, r(name + ".off.u64", np.uint64), r(name + ".order.u32", np.uint32) if os.path.exists(op) else None)
            coll = cls(table("terms"), table("docnames"), r("indptr.u64", np.uint64), r("docids.u32", np.uint32), r("impacts.u16", np.uint16))
            if coll.n_docs != meta["n_docs"] or coll.n_terms != meta["n_terms"] or int(coll.indptr[-1]) != meta["nnz"]:
                return None
            return coll
        except (OSError, ValueError, KeyError):
            return None

    # ------------------------------------------------------------------------------------------------------
    def shard(self, lo, hi):
        """Canonical CSR of the internal-id range [lo,hi) with doc ids rebased to 0 (doc-sharded multi-GPU)."""
        m = (self.post_docs >= lo) & (self.post_docs < hi)
        term_of = np.repeat(np.arange(self.n_terms), self.df)
        indptr = np.zeros(self.n_terms + 1, dtype=np.uint64)
        np.cumsum(np.bincount(term_of[m], minlength=self.n_terms), out=indptr[1:])
        return indptr, (self.post_docs[m] - lo).astype(np.uint32), self.post_impacts[m]

    def token_to_term_map(self, vocab_dict, is_filtered=False):
        """int32 [len(vocab_dict)]: tokenizer id -> term id of this index, -1 when the token is not in the dictionary.

        The string is built as the reference builds it (src/encode.py:78-83): ``vocab_dict[i].lower()``, through
        ``filter_token`` when ``is_filtered`` -- then split on whitespace as the query encoder would (S4).  Used by
        ops.quantise_topk to go from activation rows to a query CSR without leaving the GPU."""
        n = len(vocab_dict)
        out = np.full(n, -1, dtype=np.int32)
        for i in range(n):
            tok = vocab_dict[i].lower()
            if is_filtered and tok and (ord(tok[0]) < ord('a') or ord(tok[0]) > ord('z')):
                tok = tok[1:]
            pieces = tok.split()
            if len(pieces) > 1:
                raise ValueError(f"vocabulary entry {i} ({tok!r}) contains whitespace: it is several index terms")
            if pieces:
                out[i] = self.term_to_id.get(pieces[0], -1)
        return out

    def encode_queries(self, queries):
        """S4 + dictionary lookup: list of query strings (or {token: weight} maps) -> query CSR (u64, u32, u32).

        A string is encoded as pyserini's TokFreqQueryEncoder does: Counter(text.strip().split()).  Tokens outside
        the dictionary are dropped here; the df-based half of S5 is applied on the device."""
        indptr = np.zeros(len(queries) + 1, dtype=np.uint64)
        terms, weights = [], []
        for i, q in enumerate(queries):
            counts = Counter(q.strip().split()) if isinstance(q, str) else q
            for tok, w in counts.items():
                tid = self.term_to_id.get(tok)
                w = int(w)
                if tid is not None and w > 0:
                    terms.append(tid); weights.append(w)
            indptr[i + 1] = len(terms)
        return indptr, np.asarray(terms, dtype=np.uint32), np.asarray(weights, dtype=np.uint32)
