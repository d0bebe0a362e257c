#!/usr/bin/env python
"""Engine fixtures: the PIN that is still missing for the two third-party engines ("parity unpinned").

The reference's sparse and dense arithmetic lives in pyserini -> Anserini -> Lucene and tevatron -> faiss.IndexFlatIP,
neither of which exists in the build container (no JVM, no wheels, no network).  This script is the recipe that
produces real-engine answers on ANY machine that has them:

    pip install pyserini faiss-cpu            # + a JDK 21 on PATH (pyserini's requirement)
    python tests/golden/make_engine_golden.py --cases-only      # (re)writes tests/golden/engine_cases.json  [no engine needed]
    python tests/golden/make_engine_golden.py                   # writes engine_lucene.json and engine_faiss.json

It drives the engines exactly as the reference does:
  * index:  `python -m pyserini.index.lucene --collection JsonVectorCollection --generator DefaultLuceneDocumentGenerator
             --impact --pretokenized`                                        (/root/reference/scripts/sparse_index.sh:12-18)
  * search: `LuceneImpactSearcher(index_dir, None)`, `set_analyzer(JWhiteSpaceAnalyzer())`,
            `batch_search(topics, qids, k, threads=...)`, hits read as (hit.docid, hit.score)
                                                                              (/root/reference/src/search.py:273-275, 85-99)
  * dense:  `faiss.IndexFlatIP(dim)`; `add`; `search(q, k)`                   (tevatron FaissFlatSearcher; search.py:232-237, 55-59)

The CASES are committed (engine_cases.json) and are built to SEPARATE the rules the oracle restates from memory
 -- each case's "pins" field says which:
  S5   idf filter: a term with df = N-1 (ln(N/(df+1)) = 0, dropped) beside one with df = N-2 (kept); N counts documents
       whose vector is empty or holds only non-positive weights (random_ties_*, q3_*)
  S6   float32 per-term products and sums at and above 2^24 (odd products that float32 cannot hold)
  S7   fewer hits than k; a query whose terms are all unknown / all dropped
  S8   equal scores on docids "100" vs "99" vs "1000" (string order, not numeric)
  S8b  many exact ties (would the score-ties adjuster perturb the scores?)
  Q3   duplicate query tokens add; zero / negative vector weights are never indexed
  D3-5 exact fp32 inner products, exact ties (duplicate rows), k > N padding (-1 / -3.4028235e38)

Output schema (consumed by tests/test_engine_golden_cpu.py and tests/test_engine_golden_gpu.py):
  engine_cases.json   {"lucene": [{"name", "pins", "docs": [{"id": str, "vector": {tok: int}}], "queries": [{"qid", "text"}], "k"}],
                       "faiss":  [{"name", "pins", "dim", "docs": hex, "queries": hex, "k"}]}   (row-major little-endian f32 bytes as hex;
                       a query is {"qid", "text"} or {"qid", "rle": [[token, count], ...]} -- see query_text())
  engine_lucene.json  {"versions": {"pyserini", "java", "lucene"?}, "results": {case name: {qid: [[docid, score], ...]}}}
  engine_faiss.json   {"versions": {"faiss"}, "results": {case name: {"scores": [[hex f32]], "ids": [[int]]}}}
Scores are stored as the hex of their float32 bits, so the comparison is bit-exact.
"""
from __future__ import annotations

import argparse
import json
import os
import struct
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))


def query_text(q) -> str:
    """The query string of a case entry: "text" verbatim, or its run-length form "rle" expanded."""
    if "text" in q:
        return q["text"]
    return " ".join(" ".join([tok] * int(n)) for tok, n in q["rle"])


def rows_hex(a) -> str:
    """float32 matrix -> hex of its little-endian bytes (bit-exact, compact)."""
    import numpy as np
    return np.ascontiguousarray(a, dtype="<f4").tobytes().hex()


def hex_rows(h, dim):
    import numpy as np
    return np.frombuffer(bytes.fromhex(h), dtype="<f4").reshape(-1, dim).copy()


def f32hex(x) -> str:
    return struct.pack(">f", float(x)).hex()


def hexf32(h: str) -> float:
    return struct.unpack(">f", bytes.fromhex(h))[0]


# ----------------------------------------------------------------------------------------------------------------
# cases
# ----------------------------------------------------------------------------------------------------------------
def lucene_cases():
    import numpy as np
    cases = []

    # S5: N = 6 docs; "all" in 5 docs (df = N-1 -> idf 0 -> dropped), "most" in 4 (df = N-2 -> kept), "one" in 1
    docs = [{"id": str(i), "vector": {}} for i in range(6)]
    for i in range(5):
        docs[i]["vector"]["all"] = 3 + i
    for i in range(4):
        docs[i]["vector"]["most"] = 2 + i
    docs[5]["vector"]["one"] = 7
    docs[2]["vector"]["one2"] = 1
    cases.append({"name": "s5_idf_filter", "pins": ["S5", "S7"], "docs": docs, "k": 10,
                  "queries": [{"qid": "q_all", "text": "all all all"},                 # only the dropped term -> no hits
                              {"qid": "q_most", "text": "most most"},
                              {"qid": "q_mix", "text": "all most one one unknownterm"},
                              {"qid": "q_unknown", "text": "neverseen"}]})

    # S8 / S8b: equal scores, docids whose string order differs from numeric order
    ids = ["100", "99", "1000", "9", "10", "098", "a1", "B2"]
    docs = [{"id": d, "vector": {"tie": 5, "x": 1 + (i % 2)}} for i, d in enumerate(ids)]
    docs.append({"id": "zz", "vector": {"other": 1}})
    docs.append({"id": "zy", "vector": {"other": 2}})
    cases.append({"name": "s8_docid_ties", "pins": ["S8", "S8b"], "docs": docs, "k": 6,
                  "queries": [{"qid": "t1", "text": "tie tie"}, {"qid": "t2", "text": "tie x"}, {"qid": "t3", "text": "x"}]})

    # Q3 / S1: duplicate query tokens add; zero and negative vector weights are never indexed
    docs = [{"id": "d0", "vector": {"dog": 100, "cat": 0, "emu": -3}},
            {"id": "d1", "vector": {"dog": 1, "cat": 56}},
            {"id": "d2", "vector": {"cat": 1, "emu": 2}},
            {"id": "d3", "vector": {"fox": 9}},
            {"id": "d4", "vector": {"fox": 1, "gnu": 1}}]
    cases.append({"name": "q3_duplicates_and_nonpositive", "pins": ["Q3", "S1", "S4"], "docs": docs, "k": 5,
                  "queries": [{"qid": "a", "text": " ".join(["dog"] * 223)},
                              {"qid": "b", "text": "cat emu cat  dog\tcat"},
                              {"qid": "c", "text": "  emu  "}]})

    # S6: per-term products and sums at / above 2^24.  4001*5001 = 20009001 (odd, not a float32); sums of two such
    docs = [{"id": "p0", "vector": {"big": 4001, "small": 1}},
            {"id": "p1", "vector": {"big": 4001, "small": 2}},
            {"id": "p2", "vector": {"big": 4000, "small": 3, "big2": 4099}},
            {"id": "p3", "vector": {"big2": 4099, "small": 1}},
            {"id": "p4", "vector": {"small": 1}},
            {"id": "p5", "vector": {"pad": 1}},
            {"id": "p6", "vector": {"pad": 1}}]
    cases.append({"name": "s6_float32_rounding", "pins": ["S6"], "docs": docs, "k": 7,
                  # "rle" = run-length form of the query text (query_text() expands it): 5001 x "big" then 1 x "small", ...
                  "queries": [{"qid": "r1", "rle": [["big", 5001], ["small", 1]]},
                              {"qid": "r2", "rle": [["big", 5001], ["big2", 4097], ["small", 3]]},
                              {"qid": "r3", "rle": [["big", 4194], ["small", 1]]}]})   # 4001*4194 = 16780194 just above 2^24

    # a seeded ragged collection with heavy ties (impacts 1..3), k above and below the hit counts
    rng = np.random.default_rng(20261105)
    vocab = [f"t{i}" for i in range(40)]
    docs = []
    for i in range(120):
        m = int(rng.integers(0, 9))
        toks = rng.choice(40, size=m, replace=False)
        docs.append({"id": str(int(rng.integers(0, 10 ** 6))) + "_" + str(i), "vector": {vocab[int(t)]: int(rng.integers(1, 4)) for t in toks}})
    queries = []
    for qn in range(24):
        m = int(rng.integers(1, 7))
        toks = rng.choice(40, size=m, replace=True)
        text = " ".join(" ".join([vocab[int(t)]] * int(rng.integers(1, 4))) for t in toks)
        queries.append({"qid": f"r{qn}", "text": text})
    cases.append({"name": "random_ties_k10", "pins": ["S7", "S8", "S8b"], "docs": docs, "k": 10, "queries": queries})
    cases.append({"name": "random_ties_k1000", "pins": ["S7", "S8"], "docs": docs, "k": 1000, "queries": queries[:8]})
    return cases


def faiss_cases():
    import numpy as np
    rng = np.random.default_rng(4096)
    cases = []
    dim = 64
    docs = rng.standard_normal((50, dim)).astype(np.float32)
    docs /= np.linalg.norm(docs, axis=1, keepdims=True)
    docs[7] = docs[3]                               # exact ties: which id does FAISS return first?
    docs[21] = docs[3]
    q = rng.standard_normal((6, dim)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    q[1] = docs[3]
    enc = rows_hex
    cases.append({"name": "flat_ip_ties", "pins": ["D3", "D4"], "dim": dim, "docs": enc(docs), "queries": enc(q), "k": 10})
    cases.append({"name": "flat_ip_k_above_n", "pins": ["D5"], "dim": dim, "docs": enc(docs[:5]), "queries": enc(q[:2]), "k": 8})
    big = rng.standard_normal((200, 128)).astype(np.float32)
    qb = rng.standard_normal((4, 128)).astype(np.float32)
    cases.append({"name": "flat_ip_unnormalised", "pins": ["D3"], "dim": 128, "docs": enc(big), "queries": enc(qb), "k": 100})
    return cases


# ----------------------------------------------------------------------------------------------------------------
# engines
# ----------------------------------------------------------------------------------------------------------------
def run_lucene(cases, threads=4):
    from pyserini.analysis import JWhiteSpaceAnalyzer
    from pyserini.search.lucene import LuceneImpactSearcher
    import pyserini
    results = {}
    for case in cases:
        with tempfile.TemporaryDirectory() as tmp:
            coll = os.path.join(tmp, "coll")
            os.makedirs(coll)
            with open(os.path.join(coll, "corpus_0.jsonl"), "w") as f:
                for d in case["docs"]:
                    # the layout src/encode.py:353-359 writes: id, content, vector
                    f.write(json.dumps({"id": d["id"], "content": "", "vector": d["vector"]}) + "\n")
            index = os.path.join(coll, "index")
            subprocess.check_call([sys.executable, "-m", "pyserini.index.lucene", "--collection", "JsonVectorCollection",
                                   "--input", coll, "--index", index, "--generator", "DefaultLuceneDocumentGenerator",
                                   "--threads", "1", "--impact", "--pretokenized"])
            searcher = LuceneImpactSearcher(index, None)
            searcher.set_analyzer(JWhiteSpaceAnalyzer())
            topics = [query_text(q) for q in case["queries"]]
            qids = [q["qid"] for q in case["queries"]]
            hits = searcher.batch_search(topics, qids, case["k"], threads=threads)
            results[case["name"]] = {qid: [[h.docid, f32hex(h.score)] for h in hits[qid]] for qid in qids}
    java = subprocess.run(["java", "-version"], capture_output=True, text=True).stderr.strip().splitlines()
    return {"versions": {"pyserini": getattr(pyserini, "__version__", "unknown"), "java": java[0] if java else "unknown"},
            "results": results}


def run_faiss(cases):
    import faiss
    import numpy as np
    results = {}
    for case in cases:
        docs = hex_rows(case["docs"], case["dim"])
        q = hex_rows(case["queries"], case["dim"])
        index = faiss.IndexFlatIP(case["dim"])
        index.add(docs)
        s, i = index.search(q, case["k"])
        results[case["name"]] = {"scores": [[f32hex(x) for x in row] for row in s], "ids": [[int(x) for x in row] for row in i]}
    return {"versions": {"faiss": getattr(faiss, "__version__", "unknown")}, "results": results}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases-only", action="store_true", help="write engine_cases.json only (no engine needed)")
    ap.add_argument("--threads", type=int, default=4)
    a = ap.parse_args()
    cases = {"lucene": lucene_cases(), "faiss": faiss_cases()}
    with open(os.path.join(HERE, "engine_cases.json"), "w") as f:
        json.dump(cases, f)
    print("wrote engine_cases.json:", [c["name"] for c in cases["lucene"]], [c["name"] for c in cases["faiss"]])
    if a.cases_only:
        return
    with open(os.path.join(HERE, "engine_lucene.json"), "w") as f:
        json.dump(run_lucene(cases["lucene"], a.threads), f)
    with open(os.path.join(HERE, "engine_faiss.json"), "w") as f:
        json.dump(run_faiss(cases["faiss"]), f)
    print("wrote engine_lucene.json, engine_faiss.json")


if __name__ == "__main__":
    main()
