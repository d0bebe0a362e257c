#!/usr/bin/env python
"""Fixture for the float64 fusion path: the reference's own hybrid.fuse (/root/reference/src/hybrid.py:32-53) on
runs re-read with hybrid.read_trec_run (:8-17) from TREC text whose scores are NOT float32-representable (0.1, 12.3456789,
...).  Run in the build container (imports /root/reference/src/hybrid.py, which needs nothing else):
    python tests/golden/make_fuse_f64_golden.py   ->  tests/golden/reference_fuse_f64.json
Scores are stored as float.hex() so the comparison is bit-exact."""
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    sys.path.insert(0, "/root/reference/src")
    import hybrid
    rng = np.random.default_rng(64)
    cases = []
    for case in range(3):
        nq, n_docs = 6, 40
        texts = []
        for run in range(2):
            lines = []
            for q in range(nq):
                k = int(rng.integers(1, 15))
                docs = rng.choice(n_docs, size=k, replace=False)
                sc = np.sort(rng.random(k) * (30.0 if run else 1.0))[::-1]
                if case == 1 and q == 2:
                    sc[:] = 0.3                                  # max == min: the 1e-9 floor of the denominator
                for r, (d, s) in enumerate(zip(docs, sc)):
                    lines.append(f"q{q} Q0 d{d} {r + 1} {round(float(s), 7)} run{run}")
            texts.append("\n".join(lines) + "\n")
        runs = []
        for t in texts:
            with tempfile.NamedTemporaryFile("w", suffix=".trec", delete=False) as f:
                f.write(t)
            runs.append(hybrid.read_trec_run(f.name))
            os.unlink(f.name)
        alpha = [0.5, 0.3, 0.7][case]
        fused = hybrid.fuse(runs, [alpha, 1 - alpha])
        cases.append({"trec": texts, "alpha": alpha,
                      "fused": {q: [[d, float(v).hex()] for d, v in docs.items()] for q, docs in fused.items()}})
    with open(os.path.join(HERE, "reference_fuse_f64.json"), "w") as f:
        json.dump({"numpy": np.__version__, "cases": cases}, f)
    print("wrote reference_fuse_f64.json:", [len(c["fused"]) for c in cases])


if __name__ == "__main__":
    main()
