"""Search CLI with the reference's own arguments (the search half of /root/reference/src/search.py:main, :102-468).

    python -m mllm_sparse_retrieval_b200.search --passage_reps <dir> --sparse_index <dir> \\
        --query_reps <dir|query.pkl> --query_sparse <dir|query.tsv> --depth 1000 --alpha 0.5 --save_dir runs/ [--remove_query]

The reference encodes every query with the MLLM inside its search loop; that half is out of scope here, so queries
come PRE-ENCODED in the files ``src/encode.py --encode_is_query`` writes (encode.py:407-424): ``query.pkl`` =
pickle of (float32 [nq,H], ids) for the dense side, ``query.tsv`` = ``id<TAB>token token ...`` for the sparse side.
Everything after the encoder follows the reference:

  dense   pickle_load(first corpus*.pkl) -> FaissFlatSearcher(reps); add(reps)        search.py:227-237
          search_queries: batch_search(q, depth, batch_size, quiet), ids -> lookup   search.py:55-63
  sparse  LuceneImpactSearcher(<sparse_index>/index); batch_search(topics, ids, depth, threads)   search.py:272-275, 85-99
  runs    get_run_dict(..., remove_query)                                             search.py:66-82
  fusion  fuse([dense, sparse], [alpha, 1-alpha])                                     search.py:455-461, hybrid.py:32-53
  output  write_trec_run under --save_dir (hybrid.py:20-29; the reference parses save_dir and never writes)

Argument names and defaults are PromptRepsLLMSearchArguments verbatim (src/arguments.py:56-68).  The searches, the
fusion and the final ordering run on the GPU over arrays; only the run-file text is produced on the host.
"""
from __future__ import annotations

import argparse
import dataclasses
import glob
import os
import pickle
import sys
from dataclasses import dataclass, field

import numpy as np
import torch

from . import ops
from .searchers import FaissFlatSearcher, LuceneImpactSearcher


@dataclass
class PromptRepsLLMSearchArguments:
    passage_reps: str = field(default=None, metadata={"help": "Path to passage dense representations"})
    sparse_index: str = field(default=None, metadata={"help": "Path to passage sparse representations"})
    depth: int = field(default=1000)
    save_dir: str = field(default=None, metadata={"help": "Where to save the run files"})
    quiet: bool = field(default=True, metadata={"help": "Whether to print the progress"})
    use_gpu: bool = field(default=False, metadata={"help": "Whether to use GPU"})
    alpha: float = field(default=0.5, metadata={"help": "The weight for dense retrieval"})
    batch_size: int = field(default=128, metadata={"help": "Batch size for retrieval"})
    remove_query: bool = field(default=False, metadata={"help": "Whether to remove query id from the ranking"})
    threads: int = field(default=1, metadata={"help": "Number of threads for sparse retrieval"})
    query_type: str = field(default='text')


def _str2bool(v):
    if isinstance(v, bool):
        return v
    if v.lower() in ("1", "true", "yes", "y"):
        return True
    if v.lower() in ("0", "false", "no", "n"):
        return False
    raise argparse.ArgumentTypeError(f"expected a boolean, got {v!r}")


def parse_args(argv=None):
    """HfArgumentParser conventions for the dataclass (``--flag`` or ``--flag true``), plus the pre-encoded query paths.
    Unknown arguments (the reference's model / data / training arguments) are accepted and ignored."""
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    for f in dataclasses.fields(PromptRepsLLMSearchArguments):
        if f.type in (bool, "bool"):
            ap.add_argument(f"--{f.name}", type=_str2bool, nargs="?", const=True, default=f.default)
        else:
            typ = {"int": int, "float": float, "str": str}.get(f.type if isinstance(f.type, str) else f.type.__name__, str)
            ap.add_argument(f"--{f.name}", type=typ, default=f.default, help=f.metadata.get("help"))
    ap.add_argument("--query_reps", default=None, help="query.pkl (or its directory) written by encode.py --encode_is_query")
    ap.add_argument("--query_sparse", default=None, help="query.tsv (or its directory) written by encode.py --encode_is_query")
    ap.add_argument("--qrels", default=None, help="optional: lines `qid<TAB>docid [docid ...]` -> Recall@{1,5,10,100,200}")
    ap.add_argument("--run_name", default="fusion")
    ns, unknown = ap.parse_known_args(argv)
    if unknown and not ns.quiet:
        print(f"[search] ignoring encoder-side arguments: {' '.join(unknown)}", file=sys.stderr)
    sa = PromptRepsLLMSearchArguments(**{f.name: getattr(ns, f.name) for f in dataclasses.fields(PromptRepsLLMSearchArguments)})
    return sa, ns


def pickle_load(path):
    """search.py:49-52"""
    with open(path, "rb") as f:
        reps, lookup = pickle.load(f)
    return np.array(reps), lookup


def _resolve(path, name):
    return os.path.join(path, name) if os.path.isdir(path) else path


def read_query_tsv(path):
    ids, texts = [], []
    with open(path, "r", encoding="utf-8") as f:
        for line in f:
            line = line.rstrip("\n")
            if not line:
                continue
            qid, _, text = line.partition("\t")
            ids.append(qid)
            texts.append(text)
    return ids, texts


def _sorted_run(scores, docs, counts):
    """Stable descending sort of each row's valid prefix (metrices.py:30 / write_trec_run order); torch plumbing."""
    k = scores.shape[1]
    pad = torch.arange(k, device=scores.device)[None, :] >= counts[:, None]
    key = scores.masked_fill(pad, float("-inf"))
    order = torch.sort(key, dim=1, descending=True, stable=True).indices
    return torch.gather(scores, 1, order), torch.gather(docs, 1, order)


def write_trec_arrays(path, qids, names, scores, docs, counts, run_name):
    """hybrid.write_trec_run for array-form runs (rows already in rank order)."""
    with open(path, "w") as f:
        for r, qid in enumerate(qids):
            n = int(counts[r])
            f.writelines(f"{qid} Q0 {names[docs[r, j]]} {j + 1} {scores[r, j]} {run_name}\n" for j in range(n))


def recall_at(qids, name_id, docs, counts, qrels, ks=ops.RECALL_CUTOFFS):
    """Recall@k as the reference computes it (metrices.py:29-100): a query scores a hit at k when ANY of its targets is
    among its first k ranked docs; queries with an empty list never hit; recall = hits / number of queries.
    docs i32 [nq, w] CUDA ranked lists of internal ids, counts [nq]; qrels = {qid: set of external docids}.
    The hit test runs on the device (msr_recall_hits); only the 5 counters come back."""
    indptr, flat = [0], []
    for qid in qids:
        for t in sorted(qrels.get(qid, ())):
            i = name_id.get(t)
            if i is not None:
                flat.append(i)
        indptr.append(len(flat))
    hits, n_eval = ops.recall_hits(docs.contiguous(), counts, torch.tensor(indptr, dtype=torch.int64),
                                   torch.tensor(flat, dtype=torch.int32), ks)
    hits = hits.cpu().tolist()
    n = max(1, len(qids))
    return {k: h / n for k, h in zip(ks, hits)}, int(n_eval.item())


def main(argv=None):
    sa, ns = parse_args(argv)
    if not torch.cuda.is_available():
        raise SystemExit("mllm_sparse_retrieval_b200.search needs a CUDA (sm_100a) device: there is no CPU fallback")
    if sa.passage_reps is None and sa.sparse_index is None:
        raise SystemExit("give --passage_reps and/or --sparse_index")
    dev = torch.device("cuda", torch.cuda.current_device())
    depth = int(sa.depth)
    dense = sparse = None
    names, name_id = [], {}

    # ---- sparse side --------------------------------------------------------------------------------------
    if sa.sparse_index is not None:
        if ns.query_sparse is None:
            raise SystemExit("--sparse_index needs --query_sparse (query.tsv)")
        searcher = LuceneImpactSearcher(os.path.join(sa.sparse_index, "index"), None)
        searcher.set_analyzer(None)
        s_qids, topics = read_query_tsv(_resolve(ns.query_sparse, "query.tsv"))
        names = list(searcher.collection.docids)                # internal id -> external docid string
        name_id = dict(searcher.collection.docid_to_internal)
        k_s = min(depth, ops.MSR_MAX_K)
        ss, sd, sc = searcher.search_arrays(topics, k_s)
        sparse = (s_qids, torch.from_numpy(ss).to(dev), torch.from_numpy(sd).to(dev), torch.from_numpy(sc.astype(np.int32)).to(dev))

    # ---- dense side ---------------------------------------------------------------------------------------
    if sa.passage_reps is not None:
        if ns.query_reps is None:
            raise SystemExit("--passage_reps needs --query_reps (query.pkl)")
        index_files = sorted(glob.glob(os.path.join(sa.passage_reps, "corpus*.pkl")))
        if not index_files:
            raise SystemExit(f"no corpus*.pkl under {sa.passage_reps}")
        p_reps, p_lookup = pickle_load(index_files[0])          # the reference loads the first shard only (search.py:228-237)
        retriever = FaissFlatSearcher(p_reps)
        retriever.add(p_reps)
        q_reps, d_qids = pickle_load(_resolve(ns.query_reps, "query.pkl"))
        d_qids = [str(x) for x in d_qids]
        k_d = min(depth, ops.MSR_MAX_K, len(p_lookup))
        if sa.batch_size > 0:
            d_scores, d_idx = retriever.batch_search(q_reps.astype(np.float32), k_d, sa.batch_size, sa.quiet)
        else:
            d_scores, d_idx = retriever.search(q_reps.astype(np.float32), k_d)
        # dense row number -> shared doc numbering (the sparse index's internal ids when both sides exist)
        if not names:
            names = [str(x) for x in p_lookup]
            row_to_id = np.arange(len(p_lookup), dtype=np.int32)
        else:
            row_to_id = np.empty(len(p_lookup), dtype=np.int32)
            for r, x in enumerate(p_lookup):
                x = str(x)
                i = name_id.get(x)
                if i is None:                                   # a doc only the dense side knows
                    i = name_id[x] = len(names)
                    names.append(x)
                row_to_id[r] = i
        dd = torch.from_numpy(row_to_id).to(dev)[torch.from_numpy(d_idx).to(dev).clamp_(min=0)]
        dc = torch.from_numpy((d_idx >= 0).sum(1).astype(np.int32)).to(dev)
        dense = (d_qids, torch.from_numpy(d_scores).to(dev), dd.to(torch.int32), dc)

    # ---- runs, fusion, order ------------------------------------------------------------------------------
    lookup = name_id if name_id else {n: i for i, n in enumerate(names)}

    def self_ids(qids):
        return torch.tensor([lookup.get(q, -1) for q in qids], dtype=torch.int32, device=dev)

    outputs = {}
    if dense is not None and sparse is not None:
        if dense[0] != sparse[0]:
            # the reference indexes run[qid] for every qid of the dense run: the query sets must agree
            pos = {q: i for i, q in enumerate(sparse[0])}
            missing = [q for q in dense[0] if q not in pos]
            if missing:
                raise KeyError(missing[0])
            sel = torch.tensor([pos[q] for q in dense[0]], device=dev)
            sparse = (dense[0], sparse[1][sel], sparse[2][sel], sparse[3][sel])
        qids = dense[0]
        fs, fd, fsrc, fc = ops.fuse_lists(dense[1].contiguous(), dense[2].contiguous(), dense[3], sparse[1].contiguous(),
                                          sparse[2].contiguous(), sparse[3], self_ids(qids), sa.remove_query, sa.alpha,
                                          1 - sa.alpha)
        outputs[ns.run_name] = (qids,) + _sorted_run(fs, fd, fc) + (fc,)
    for tag, side in (("dense", dense), ("sparse", sparse)):
        if side is None:
            continue
        qids, s, d, c = side
        if sa.remove_query:                                      # get_run_dict: drop doc == qid, keep the order
            me = self_ids(qids)[:, None]
            keep = (d != me) & (torch.arange(s.shape[1], device=dev)[None, :] < c[:, None])
            order = torch.sort((~keep).to(torch.int8), dim=1, stable=True).indices
            s, d, c = torch.gather(s, 1, order), torch.gather(d, 1, order), keep.sum(1).to(torch.int32)
        outputs[tag] = (qids, s, d, c)

    main_tag = ns.run_name if ns.run_name in outputs else ("dense" if dense is not None else "sparse")
    if sa.save_dir:
        os.makedirs(sa.save_dir, exist_ok=True)
        for tag, (qids, s, d, c) in outputs.items():
            write_trec_arrays(os.path.join(sa.save_dir, f"run.{tag}.trec"), qids, names, s.cpu().numpy(), d.cpu().numpy(),
                              c.cpu().numpy(), tag)
    if ns.qrels:
        qrels = {}
        with open(ns.qrels) as f:
            for line in f:
                parts = line.split()
                if parts:
                    qrels.setdefault(parts[0], set()).update(parts[1:])
        qids, s, d, c = outputs[main_tag]
        rec, n_eval = recall_at(qids, lookup, d, c, qrels)
        print(f"[{main_tag}] queries: {len(qids)} (non-empty lists: {n_eval})")
        for k, v in rec.items():
            print(f"Recall@{k}: {v:.4f}")
    return outputs, names


if __name__ == "__main__":
    main()
