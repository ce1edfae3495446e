"""The reference itself on the GPU box -- TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench.py's reference legs).

`/root/reference` does not exist on the GPU box; `oracle/build_ref_ltf.py` (run where it does) leaves under the
git-ignored `oracle/_ref/`

  * tree_filter_cuda_ref.so -- lib_tree_filter/src rebuilt for sm_100a (only `#include <THC/THC.h>` removed), and
  * py/                     -- verbatim copies of the reference's Python on the loss path, in its own package layout.

This module loads them.  `install()` registers ONE proxy module as ``tree_filter_cuda`` whose backend can be switched
at run time (`use("reference" | "fedicra" | "oracle")`), so the reference's unmodified functions/*.py, modules/
tree_filter.py and flower_common.py -- which bind `tree_filter_cuda` once, at import -- can be driven on the
reference's own kernels, on this repository's kernels, or on the CPU restatement, inside one process.
"""
import importlib
import importlib.util
import os
import sys
import types

from oracle import ref_harness

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
PY_DIR = os.path.join(REF_DIR, "py")
EXT_SO = os.path.join(REF_DIR, "tree_filter_cuda_ref.so")

_EXT = None
_BACKENDS = {}
_STATE = {"backend": None, "installed": False}


def available():
    return os.path.exists(EXT_SO) and os.path.isdir(PY_DIR)


def wheel_load_status():
    """What happens when one tries the shipped wheel: recorded, never silently skipped."""
    return ("tree_filter-0.1-cp39-cp39-linux_x86_64.whl: not loadable (cp39 ABI, libcudart.so.11.0, torch 1.10 "
            "symbols, sm_75 SASS/PTX only); substitute = reference .cu rebuilt for sm_100a, THC include removed")


def ext():
    """The reference's pybind module (mst_forward, bfs_forward, refine_forward, refine_backward_{feature,weight})."""
    global _EXT
    if _EXT is None:
        if not os.path.exists(EXT_SO):
            raise RuntimeError("oracle/_ref/tree_filter_cuda_ref.so is missing: run `python oracle/build_ref_ltf.py` "
                               "where /root/reference is mounted")
        import torch  # noqa: F401  (libtorch must be loaded first)
        spec = importlib.util.spec_from_file_location("tree_filter_cuda_ref", EXT_SO)
        _EXT = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_EXT)
    return _EXT


def use(backend):
    """Route the reference glue's `_C.*` calls: "reference" (its own kernels), "fedicra" (this repository's
    tree_filter_cuda stand-in) or "oracle" (the CPU restatement)."""
    if backend not in _BACKENDS:
        if backend == "reference":
            _BACKENDS[backend] = ext()
        elif backend == "fedicra":
            from fedicra_b200 import tree_filter_cuda
            _BACKENDS[backend] = tree_filter_cuda
        elif backend == "oracle":
            from oracle import tree_cpu
            _BACKENDS[backend] = tree_cpu
        else:
            raise ValueError(backend)
    _STATE["backend"] = _BACKENDS[backend]


def install(backend="reference"):
    """Stub the absent third-party imports, put oracle/_ref/py first on sys.path and register the proxy."""
    if not os.path.isdir(PY_DIR):
        raise RuntimeError("oracle/_ref/py is missing: run `python oracle/build_ref_ltf.py` where /root/reference is")
    if not _STATE["installed"]:
        ref_harness.install_stubs()
        for name in ("matplotlib", "matplotlib.pyplot", "skimage", "skimage.measure"):
            try:
                importlib.import_module(name)
            except Exception:
                ref_harness._stub(name)
        proxy = types.ModuleType("tree_filter_cuda")
        proxy.__getattr__ = lambda name: getattr(_STATE["backend"], name)
        sys.modules["tree_filter_cuda"] = proxy
        # drop modules a previous harness may have imported from /root/reference so both never mix
        for name in list(sys.modules):
            if name == "utils" or name.startswith("utils.") or name in ("flower_common", "val_2D") \
                    or name == "networks" or name.startswith("networks.") \
                    or name == "dataloaders" or name.startswith("dataloaders."):
                del sys.modules[name]
        if ref_harness.REF_CODE in sys.path:
            sys.path.remove(ref_harness.REF_CODE)
        sys.path.insert(0, PY_DIR)
        _STATE["installed"] = True
    use(backend)


def gate_crf_loss():
    install(_current())
    return importlib.import_module("utils.gate_crf_loss")


def tree_filter():
    install(_current())
    return importlib.import_module("utils.TreeEnergyLoss.kernels.lib_tree_filter.modules.tree_filter")


def flower_common():
    install(_current())
    return importlib.import_module("flower_common")


def ours_script():
    """flower_pCE_2D_GateCRFMsacleTreeEnergyLoss_Ours.py (MyClient._train), imported as a module, unmodified."""
    install(_current())
    return importlib.import_module("flower_pCE_2D_GateCRFMsacleTreeEnergyLoss_Ours")


def _current():
    for k, v in _BACKENDS.items():
        if v is _STATE["backend"]:
            return k
    return "reference"
