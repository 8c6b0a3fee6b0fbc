"""The C-ABI library loads and exports every symbol include/gpk.h declares (no compute calls)."""
import os
import re

from conftest import ROOT
from gaussian_process_b200 import _lib


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "gpk.h")).read()
    return sorted(set(re.findall(r"GPK_API[^;(]*?\b(gpk_\w+)\s*\(", text)))


def test_header_and_binding_agree():
    names = declared_symbols()
    assert len(names) >= 14
    assert sorted(_lib.SIGNATURES) == names


def test_library_exports_every_symbol():
    lib = _lib.load()
    for name in declared_symbols():
        assert hasattr(lib, name), name
    assert lib.gpk_version() == 200
    assert lib.gpk_padded_n(1) == 128 and lib.gpk_padded_n(128) == 128 and lib.gpk_padded_n(4097) == 4224


def test_workspace_queries():
    lib = _lib.load()
    assert lib.gpk_loss_workspace_bytes(40, 100, 2, 0) <= 4096         # small path needs no scratch
    one = lib.gpk_loss_workspace_bytes(1, 4096, 3, 0)
    assert one >= 4096 * 4096 * 8
    # precision is a per-call argument (no library state): the emulated modes add the digit records
    assert lib.gpk_loss_workspace_bytes(1, 4096, 3, 1) > one + 4096 * 4096 * 7 // 2
    assert lib.gpk_loss_workspace_bytes(1, 4096, 3, 4) == lib.gpk_loss_workspace_bytes(1, 4096, 3, 1)
    assert lib.gpk_loss_workspace_bytes(256, 4096, 3, 0) >= 256 * 4096 * 4096 * 8
    assert lib.gpk_predict_workspace_bytes(4096, 4096) >= 4096 * 4096 * 8
    assert 0 < lib.gpk_fit_workspace_bytes(1000, 2, 0) < lib.gpk_fit_workspace_bytes(1000, 2, 1)
    assert lib.gpk_potrf_workspace_bytes(2, 256) == 2 * 2 * 2 * 128 * 128 * 8


def test_no_oracle_import_in_product():
    """The product path must not route through the oracle (or any CPU fallback)."""
    pkg = os.path.join(ROOT, "gaussian-process_b200")
    for dirpath, _dirs, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("no CPU fallback", ""), f
