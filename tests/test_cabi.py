"""The C-ABI library loads and exports every symbol ``include/arborist_b200.h`` declares
(no compute calls: this file runs without a GPU), and fails loudly instead of falling back."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "arborist_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(arb_[a-z_0-9]+)\s*\(", src)))


def test_header_declares_the_entry_points():
    names = _declared()
    for need in ("arb_ctx_create", "arb_ingest_coo", "arb_ingest_csr", "arb_set_prior_labels",
                 "arb_fit_trees", "arb_fit_presence", "arb_get_fit", "arb_get_map", "arb_qz_update",
                 "arb_qy_update", "arb_elbo_terms", "arb_last_fit_stats", "arb_last_fit_profile"):
        assert need in names


def test_library_exports_every_declared_symbol():
    from arborist_b200 import engine
    lib = engine.load_library()
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    # and the binding covers all of them
    assert sorted(engine._SIGNATURES) == _declared()
    assert lib.arb_abi_version() == 2


def test_no_cpu_fallback_without_a_device():
    from arborist_b200 import engine
    if engine.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(engine.ArboristError, match="no usable CUDA device"):
        engine.Engine(0)
    import pandas as pd
    from arborist_b200 import arborist
    rc = pd.DataFrame({"snv": [1, 2], "cell": [1, 1], "alt": [1, 0], "total": [1, 1]})
    cl = pd.DataFrame({"snv": [1, 2], "cluster": [1, 2]})
    with pytest.raises(engine.ArboristError):
        arborist([[(0, 1), (1, 2)]], rc, cl)


def test_null_arguments_are_errors_not_crashes():
    from arborist_b200 import engine
    lib = engine.load_library()
    raw = ctypes.CDLL(engine.library_path())
    raw.arb_ctx_create.argtypes = [ctypes.c_int, ctypes.c_void_p]
    assert raw.arb_ctx_create(0, None) == -1
    assert b"null" in lib.arb_last_error()
    raw.arb_set_option.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int64]
    assert raw.arb_set_option(None, b"slots", 1) == -1
    assert raw.arb_ctx_destroy(None) == 0


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under arborist_b200/ or arborist/ may reference it."""
    for pkg in ("arborist_b200", "arborist"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, pkg)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h")):
                    text = open(os.path.join(dirpath, f)).read()
                    assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), (dirpath, f)
                    assert "liborc" not in text, (dirpath, f)
