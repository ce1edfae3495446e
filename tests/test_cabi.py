"""The C-ABI library loads on a machine without a GPU and exports every symbol include/fedicra_sm100.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "fedicra_sm100.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fedicra_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    names = declared_symbols()
    for must in ("fedicra_gcrf_fwd_bwd", "fedicra_mst_forward", "fedicra_mst_grid", "fedicra_bfs_forward",
                 "fedicra_refine_forward", "fedicra_refine_backward_feature", "fedicra_refine_backward_weight",
                 "fedicra_edge_weight_fwd", "fedicra_edge_weight_bwd", "fedicra_ala_update_multi",
                 "fedicra_ala_update_flat", "fedicra_scale_inplace", "fedicra_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from fedicra_b200 import _lib
    for name in declared_symbols():
        assert hasattr(_lib.lib, name), f"{name} declared in the header but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature"
    assert _lib.lib.fedicra_abi_version() == 4
    assert _lib.lib.fedicra_last_error() == b""


def test_argument_validation_needs_no_gpu():
    from fedicra_b200 import _lib
    assert _lib.lib.fedicra_gcrf_workspace_bytes(12, 256, 256) >= 64 + 4 * 12 * 8 * 16
    rc = _lib.lib.fedicra_mst_forward(None, None, 1, 3, 4, None, None, 0, None)
    assert rc < 0 and b"null pointer" in _lib.lib.fedicra_last_error()
    rc = _lib.lib.fedicra_ala_update_flat(None, None, None, None, None, 0, 1.0, None)
    assert rc < 0


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "fedicra_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "oracle/" not in src, f
