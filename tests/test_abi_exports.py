"""CPU-only checks of the drop-in boundary: the library loads, exports every function include/quorum_b200.h declares,
validates tag tables on the host, and refuses to compute without a GPU (no silent CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    with open(os.path.join(ROOT, "include", "quorum_b200.h")) as f:
        src = f.read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(qb200_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from quorum_b200 import _lib, build
    build.build()
    L = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared_functions()
    assert len(names) >= 9
    for n in names:
        assert hasattr(L, n), n
    assert sorted(_lib.EXPORTS) == names


def test_tag_table_validation_on_host():
    from quorum_b200.batch import TagTable
    TagTable(["think", "reason", "reasoning", "thought", "Thought"])
    TagTable([])
    TagTable([""])
    for bad in (["a.b"], ["a|b"], ["x" * 23], ["<b>"], ["a/b"], ["kkkkkkkk"], ["é"], ["sississippi"]):
        with pytest.raises(ValueError):
            TagTable(bad)
    with pytest.raises(TypeError):
        TagTable([None])


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import quorum_b200 as qb
    with pytest.raises(qb.QuorumB200Error):
        qb.strip_thinking_tags("a <think>b</think> c", ["think"])
    with pytest.raises(qb.QuorumB200Error):
        qb.ThinkingTagFilter(["think"])
    # flag False is the identity in the reference too (oai_proxy.py:134-135): no device needed
    assert qb.strip_thinking_tags("x", ["think"], False) == "x"


def test_product_does_not_import_oracle():
    """oracle/ is the checker: nothing under quorum_b200/ may import, load or link it."""
    pkg = os.path.join(ROOT, "quorum_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh")):
                with open(os.path.join(dirpath, fn)) as f:
                    txt = f.read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), fn
                assert "libquorum_oracle" not in txt and "quorum_oracle" not in txt, fn


def test_sharding_partition():
    import numpy as np
    from quorum_b200.sharding import partition_by_bytes, shard_offsets
    sizes = [10, 0, 5, 100, 1, 1, 50, 7]
    for w in (1, 2, 3, 8, 11):
        parts = partition_by_bytes(sizes, w)
        assert len(parts) == w and parts[0][0] == 0 and parts[-1][1] == len(sizes)
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
    off = np.array([0, 3, 3, 9, 20, 21], dtype=np.int64)
    grp = np.array([0, 2, 2, 5], dtype=np.int64)
    b0, b1, o, g = shard_offsets(off, grp, 1, 3)
    assert (b0, b1) == (3, 21) and o.tolist() == [0, 6, 17, 18] and g.tolist() == [0, 0, 3]
