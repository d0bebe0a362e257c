"""CPU tests of the native index builder: libmsr's C++ JSONL reader against the pure-Python
reader (json.loads + ImpactCollection.from_vectors) on the same files.  No GPU: the inversion runs through NumPy here (host_inversion=True, the explicit opt-in);
tests/test_builder_gpu.py covers msr_invert_docs."""
import json
import os

import numpy as np
import pytest

from helpers import python_jsonl_collection


def _same(a, b):
    assert a.terms == b.terms and a.docids == b.docids
    np.testing.assert_array_equal(a.indptr, b.indptr)
    np.testing.assert_array_equal(a.post_docs, b.post_docs)
    np.testing.assert_array_equal(a.post_impacts, b.post_impacts)


def write_tricky(d):
    lines = [
        json.dumps({"id": "10", "content": "", "vector": {"dog": 5, "▁cat": 3, "a b": 2, "zero": 0, "neg": -4}}),
        json.dumps({"id": 9, "content": "x", "vector": {"dog": 1, "b": 7, "t u": 4, "é f": 1}}, ensure_ascii=False),
        "",
        '{"id": "abc", "extra": {"k": [1, 2, {"z": "q\\"}"}]}, "vector": {"dog": 2, "dog": 9, "q\\u00e9\\ud83d\\ude00": 3, '
        '"fl": 2.9, "s": "12", "tab\\tsep": 1}, "content": ""}',
        '  {"vector": null, "id": "n"}  ',
        json.dumps({"id": "100", "content": "", "vector": {}}),
    ]
    with open(os.path.join(d, "corpus_0.jsonl"), "w", encoding="utf-8") as f:
        f.write("\n".join(lines) + "\n")


def test_native_reader_equals_python_reader_on_tricky_lines(lib_built, tmp_path):
    """Escapes, surrogate pairs, raw UTF-8, keys with (Unicode) whitespace, duplicate keys (last wins), zero / negative /
    float / string weights, numeric ids, null and empty vectors, nested extra keys, blank lines."""
    from mllm_sparse_retrieval_b200.index import ImpactCollection
    write_tricky(str(tmp_path))
    a = python_jsonl_collection(str(tmp_path))
    b = ImpactCollection.from_jsonl_dir(str(tmp_path), use_cache=False, host_inversion=True)
    _same(a, b)
    assert b.docids == ["10", "100", "9", "abc", "n"]                       # UTF-8 byte order (S8)
    assert dict(zip(b.terms, b.df))["dog"] == 3


def test_native_reader_equals_python_reader_random(lib_built, tmp_path):
    from mllm_sparse_retrieval_b200.index import ImpactCollection
    rng = np.random.default_rng(5)
    vocab = [f"w{j}" for j in range(300)] + ["Ġtok", "▁x", "üß", "a/b", 'q"uote', "back\\slash"]
    ids = rng.permutation(5000)[:1200]
    for shard in range(2):
        with open(tmp_path / f"corpus_{shard}.jsonl", "w", encoding="utf-8") as f:
            for i in ids[shard::2]:
                vec = {vocab[j]: int(rng.integers(-2, 400)) for j in rng.choice(len(vocab), int(rng.integers(0, 40)), replace=False)}
                f.write(json.dumps({"id": str(i), "content": "", "vector": vec}, ensure_ascii=bool(i % 2)) + "\n")
    a = python_jsonl_collection(str(tmp_path))
    b = ImpactCollection.from_jsonl_dir(str(tmp_path), use_cache=False, host_inversion=True)
    _same(a, b)
    # the cache written by the native path reloads to the same collection
    c = ImpactCollection.from_jsonl_dir(str(tmp_path), use_cache=True, host_inversion=True)
    d = ImpactCollection.from_jsonl_dir(str(tmp_path), use_cache=True, host_inversion=True)
    _same(b, c); _same(c, d)


@pytest.mark.parametrize("line,what", [
    ('{"id": "1", "vector": {"a": 1}}\n{"id": "1", "vector": {"b": 1}}', "duplicate document id"),
    ('{"id": "1", "vector": {"a": 70000}}', "exceeds 65535"),
    ('{"id": "1", "vector": {"a": 1}', "expected"),
    ('{"vector": {"a": 1}}', 'no "id"'),
    ('{"id": "1", "vector": [1, 2]}', "must be an object"),
    ('{"id": "1", "vector": {"a": 1}} trailing', "extra data"),
])
def test_native_reader_errors(lib_built, tmp_path, line, what):
    from mllm_sparse_retrieval_b200._lib import MsrError
    from mllm_sparse_retrieval_b200.index import ImpactCollection
    (tmp_path / "corpus_0.jsonl").write_text(line + "\n")
    with pytest.raises(MsrError) as e:
        ImpactCollection.from_jsonl_files([str(tmp_path / "corpus_0.jsonl")])
    assert what in str(e.value)
