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
    ap.add_argument("--dist", default="auto", choices=["auto", "none", "replicas", "shard"],
                    help="under torchrun / deepspeed: replicas = query slices per rank, whole index everywhere (the reference's "
                         "scheme, search.py:180); shard = doc ranges per rank, fused exchange; auto = replicas")
    ap.add_argument("--query_logits", default=None, help="torch.save({'ids', 'logits' [nq,V]}): encoder output, quantised on the device")
    ap.add_argument("--vocab", default=None, help="tokenizer vocabulary as JSON (list id -> string, or {string: id})")
    ap.add_argument("--sparse_length", type=int, default=128, help="terms kept per query vector (arguments.py:48)")
    ap.add_argument("--is_filtered", type=_str2bool, nargs="?", const=True, default=False, help="filter_token on vocabulary strings (arguments.py:45)")
    ap.add_argument("--return_all", type=_str2bool, nargs="?", const=True, default=False,
                    help="replicas mode: gather every rank's rows even without --save_dir (main() then returns the full runs)")
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


def write_trec_arrays(path, qids, names, scores, docs, counts, run_name):
    """hybrid.write_trec_run for array-form runs (rows already in rank order)."""
    with open(path, "w") as f:
        for r, qid in enumerate(qids):
            n = int(counts[r])
            f.writelines(f"{qid} Q0 {names[docs[r, j]]} {j + 1} {scores[r, j]} {run_name}\n" for j in range(n))


def recall_counters(qids, name_id, docs, counts, qrels, ks=ops.RECALL_CUTOFFS):
    """Hit counters of Recall@k as the reference computes it (metrices.py:29-100): a query scores a hit at k when ANY of its
    targets is among its first k ranked docs; queries with an empty list never hit.  docs i32 [nq, w] CUDA ranked lists of
    internal ids, counts [nq]; qrels = {qid: set of external docids}.  The hit test runs on the device (msr_recall_hits);
    returns (hits i64 [len(ks)], evaluated queries) as device tensors -- recall = hits / number of queries."""
    indptr, flat = [0], []
    for qid in qids:
        for t in sorted(qrels.get(qid, ())):
            i = name_id.get(t)
            if i is not None:
                flat.append(i)
        indptr.append(len(flat))
    return ops.recall_hits(docs.contiguous(), counts, torch.tensor(indptr, dtype=torch.int64),
                           torch.tensor(flat, dtype=torch.int32), ks)


def recall_at(qids, name_id, docs, counts, qrels, ks=ops.RECALL_CUTOFFS):
    hits, n_eval = recall_counters(qids, name_id, docs, counts, qrels, ks)
    n = max(1, len(qids))
    return {k: h / n for k, h in zip(ks, hits.cpu().tolist())}, int(n_eval.item())


def init_distributed(mode):
    """(rank, world, mode).  The reference is launched as `deepspeed --num_gpus=N src/search.py` and initialises NCCL itself
    (scripts/search.sh:14, search.py:112-131); under torchrun / deepspeed (WORLD_SIZE > 1) this does the same.
    mode: replicas = every rank holds the whole index and takes a slice of the queries (what search.py:180 does with a
    DistributedSampler, without its padding duplicates); shard = every rank holds a doc range, all ranks pass all queries
    (LuceneImpactSearcher / FaissFlatSearcher with shard=True); auto = replicas."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1 or mode == "none":
        return 0, 1, "none"
    import torch.distributed as dist
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return dist.get_rank(), world, ("replicas" if mode == "auto" else mode)


def load_query_logits(path, vocab_path):
    """Encoder output for the string-free sparse path: ``torch.save({"ids": [...], "logits": float tensor [nq, V]})`` (or
    a (logits, ids) pair) -- the rows the reference's quantiser receives (log(1 + relu(x)), encode.py:56-93) -- and the
    tokenizer vocabulary as JSON, a list (id -> string) or a {string: id} map."""
    import json
    obj = torch.load(path, map_location="cpu")
    logits, ids = (obj["logits"], obj["ids"]) if isinstance(obj, dict) else obj
    with open(vocab_path, "r", encoding="utf-8") as f:
        v = json.load(f)
    if isinstance(v, dict):
        size = max(v.values()) + 1
        vocab = [""] * size
        for tok, i in v.items():
            vocab[i] = tok
    else:
        vocab = list(v)
    return [str(x) for x in ids], logits, vocab


def gather_rows(rank, world, dev, tensors, counts_total):
    """replicas mode: rows of every rank (contiguous query slices, rank order) -> the full arrays, on every rank."""
    import torch.distributed as dist
    n_max = max(counts_total)
    out = []
    for t in tensors:
        pad = torch.zeros((n_max,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev)
        pad[:t.shape[0]] = t
        g = torch.empty((world,) + tuple(pad.shape), dtype=t.dtype, device=dev)
        dist.all_gather_into_tensor(g, pad)
        out.append(torch.cat([g[r, :counts_total[r]] for r in range(world)], 0))
    return out


def main(argv=None):
    sa, ns = parse_args(argv)
    if not torch.cuda.is_available():
        raise SystemExit("mllm_sparse_retrieval_b200.search needs a CUDA (sm_100a) device: there is no CPU fallback")
    if sa.passage_reps is None and sa.sparse_index is None:
        raise SystemExit("give --passage_reps and/or --sparse_index")
    rank, world, mode = init_distributed(ns.dist)
    dev = torch.device("cuda", torch.cuda.current_device())
    shard = mode == "shard"
    depth = int(sa.depth)
    dense = sparse = None
    names, name_id = [], None

    def my_slice(n):
        """queries this rank answers: all of them, or its contiguous share in replicas mode"""
        return (n * rank // world, n * (rank + 1) // world) if mode == "replicas" else (0, n)

    # ---- sparse side --------------------------------------------------------------------------------------
    if sa.sparse_index is not None:
        if ns.query_sparse is None and ns.query_logits is None:
            raise SystemExit("--sparse_index needs --query_sparse (query.tsv) or --query_logits (+ --vocab)")
        searcher = LuceneImpactSearcher(os.path.join(sa.sparse_index, "index"), None, shard=shard)
        searcher.set_analyzer(None)
        names = searcher.collection.docids                      # internal id -> external docid string
        k_s = min(depth, ops.MSR_MAX_K)
        if ns.query_logits is not None:
            # encoder output -> quantiser -> scorer without a string in between (search.py:410-447 builds `' tok' * v`
            # strings that pyserini counts again; here msr_quantise_topk writes the query CSR the scorer reads)
            if ns.vocab is None:
                raise SystemExit("--query_logits needs --vocab (tokenizer vocabulary as JSON)")
            s_qids, logits, vocab = load_query_logits(ns.query_logits, ns.vocab)
            lo, hi = my_slice(len(s_qids))
            t2t = torch.from_numpy(searcher.collection.token_to_term_map(vocab, is_filtered=ns.is_filtered)).to(dev)
            qi, qt, qw = ops.quantise_topk(logits[lo:hi].to(dev).contiguous(), int(ns.sparse_length), token_to_term=t2t,
                                           vocab_len=len(vocab))
            ss, sd, sc = searcher.search_csr(qi, qt, qw, k_s)
        else:
            s_qids, topics = read_query_tsv(_resolve(ns.query_sparse, "query.tsv"))
            lo, hi = my_slice(len(s_qids))
            ss, sd, sc = searcher.search_arrays(topics[lo:hi], k_s)
            ss, sd, sc = torch.from_numpy(ss).to(dev), torch.from_numpy(sd).to(dev), torch.from_numpy(sc.astype(np.int32)).to(dev)
        sparse = (s_qids, ss, sd, sc.to(torch.int32))

    # ---- dense side ---------------------------------------------------------------------------------------
    if sa.passage_reps is not None:
        if ns.query_reps is None:
            raise SystemExit("--passage_reps needs --query_reps (query.pkl)")
        index_files = sorted(glob.glob(os.path.join(sa.passage_reps, "corpus*.pkl")))
        if not index_files:
            raise SystemExit(f"no corpus*.pkl under {sa.passage_reps}")
        p_reps, p_lookup = pickle_load(index_files[0])          # the reference loads the first shard only (search.py:228-237)
        retriever = FaissFlatSearcher(p_reps, shard=shard)
        retriever.add(p_reps)
        q_reps, d_qids = pickle_load(_resolve(ns.query_reps, "query.pkl"))
        d_qids = [str(x) for x in d_qids]
        lo, hi = my_slice(len(d_qids))
        q_mine = q_reps[lo:hi].astype(np.float32)
        k_d = min(depth, ops.MSR_MAX_K, len(p_lookup))
        if sa.batch_size > 0:
            d_scores, d_idx = retriever.batch_search(q_mine, k_d, sa.batch_size, sa.quiet)
        else:
            d_scores, d_idx = retriever.search(q_mine, k_d)
        # dense row number -> shared doc numbering (the sparse index's internal ids when both sides exist)
        if len(names) == 0:
            names = [str(x) for x in p_lookup]
            row_to_id = np.arange(len(p_lookup), dtype=np.int32)
        else:
            name_id = dict(searcher.collection.docid_to_internal)
            names = list(names)
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
    all_qids = (dense or sparse)[0]
    lo, hi = my_slice(len(all_qids))
    qids = all_qids[lo:hi]
    need_lookup = sa.remove_query or ns.qrels
    if name_id is None and need_lookup:
        name_id = (dict(searcher.collection.docid_to_internal) if sparse is not None and dense is None
                   else {n: i for i, n in enumerate(names)})

    def self_ids(qs):
        return torch.tensor([name_id.get(q, -1) for q in qs] if name_id is not None else [-1] * len(qs), dtype=torch.int32, device=dev)

    outputs = {}
    if dense is not None and sparse is not None:
        if dense[0] != sparse[0]:
            # the reference indexes run[qid] for every qid of the dense run: the query sets must agree
            pos = {q: i for i, q in enumerate(sparse[0])}
            missing = [q for q in dense[0] if q not in pos]
            if missing:
                raise KeyError(missing[0])
            if mode == "replicas":
                raise SystemExit("--dist replicas needs query.tsv and query.pkl in the same query order")
            sel = torch.tensor([pos[q] for q in dense[0]], device=dev)
            sparse = (dense[0], sparse[1][sel], sparse[2][sel], sparse[3][sel])
        fs, fd, fsrc, fc = ops.fuse_lists(dense[1].contiguous(), dense[2].contiguous(), dense[3], sparse[1].contiguous(),
                                          sparse[2].contiguous(), sparse[3], self_ids(qids), sa.remove_query, sa.alpha,
                                          1 - sa.alpha)
        outputs[ns.run_name] = ops.order_rows(fs, fd, fc)       # stable, score descending (metrices.py:30)
    for tag, side in (("dense", dense), ("sparse", sparse)):
        if side is None:
            continue
        _, s, d, c = side
        if sa.remove_query:                                      # get_run_dict: drop doc == qid, keep the order
            s, d, c = ops.order_rows(s, d, c, self_ids=self_ids(qids), by_score=False)
        outputs[tag] = (s, d, c)

    main_tag = ns.run_name if ns.run_name in outputs else ("dense" if dense is not None else "sparse")
    # ---- Recall@k: hit counters per rank, ONE all_reduce (the reference gathers 15 Python objects, metrices.py:86-100) --
    recall = None
    if ns.qrels:
        qrels = {}
        with open(ns.qrels) as f:
            for line in f:
                parts = line.split()
                if parts:
                    qrels.setdefault(parts[0], set()).update(parts[1:])
        s, d, c = outputs[main_tag]
        hits, n_eval = recall_counters(qids, name_id, d, c, qrels)
        if mode == "replicas":
            import torch.distributed as dist
            acc = torch.cat([hits.to(torch.int64), n_eval.reshape(1).to(torch.int64)])
            dist.all_reduce(acc)
            hits, n_eval = acc[:-1], acc[-1]
        recall = ({k: h / max(1, len(all_qids)) for k, h in zip(ops.RECALL_CUTOFFS, hits.cpu().tolist())}, int(n_eval.item()))
    # ---- run files: written once, by rank 0 ----------------------------------------------------------------------
    if mode == "replicas" and (sa.save_dir or ns.return_all):
        per_rank = [len(all_qids) * (r + 1) // world - len(all_qids) * r // world for r in range(world)]
        outputs = {tag: tuple(gather_rows(rank, world, dev, list(v), per_rank)) for tag, v in outputs.items()}
        qids = all_qids
    if sa.save_dir and rank == 0:
        os.makedirs(sa.save_dir, exist_ok=True)
        for tag, (s, d, c) in outputs.items():
            write_trec_arrays(os.path.join(sa.save_dir, f"run.{tag}.trec"), qids, names, s.cpu().numpy(), d.cpu().numpy(),
                              c.cpu().numpy(), tag)
    if recall is not None and rank == 0:
        rec, n_eval = recall
        print(f"[{main_tag}] queries: {len(all_qids)} (non-empty lists: {n_eval})")
        for k, v in rec.items():
            print(f"Recall@{k}: {v:.4f}")
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    return {tag: (qids,) + tuple(v) for tag, v in outputs.items()}, names


if __name__ == "__main__":
    main()
