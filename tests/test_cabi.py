"""The C-ABI library loads and exports every symbol include/asset_b200.h declares (no compute calls without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "asset_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ab_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported():
    lib = ctypes.CDLL(os.path.join(ROOT, "asset_asrl_b200", "libasset_b200.so"))
    names = declared_symbols()
    assert len(names) > 40
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_binding_covers_header():
    from asset_asrl_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()


def test_no_cpu_fallback():
    """Creating a device plan without a GPU must fail loudly (north_star: no CPU fallback on this path)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from asset_asrl_b200 import workloads as W
    from asset_asrl_b200._lib import AssetB200Error
    nlp = W.make_phase("hypersens", "LGL3", 4).transcribe()
    with pytest.raises(AssetB200Error):
        nlp.to_device(0)


def test_error_reporting():
    from asset_asrl_b200 import _lib
    g = _lib.lib.ab_graph_create()
    rc = _lib.lib.ab_graph_node(g, 9999, 1, 1, None, 0, None, 0, None, 0, -1, 0)
    assert rc < 0 and "unknown node kind" in _lib.last_error()
    _lib.lib.ab_graph_destroy(g)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "asset_asrl_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".hpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "liboracle" not in txt, f
