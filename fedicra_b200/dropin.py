"""Make the reference's own scripts import this package instead of its PyTorch/C++ implementations.

``install()`` registers, in ``sys.modules``, the module names the reference imports on the hot path:

  tree_filter_cuda                                                        (functions/{mst,bfs,refine}.py:7)
  utils.gate_crf_loss                                                     (OURS.py:36)
  utils.TreeEnergyLoss.kernels.lib_tree_filter.modules.tree_filter        (flower_common.py:32-33)

``level="ops"`` only swaps the native extension (the reference's Python glue keeps running, on our kernels);
``level="modules"`` (default) also swaps the two Python modules for the fused implementations.
"""
import sys
import types


def install(level="modules"):
    from . import gate_crf_loss, tree_filter, tree_filter_cuda

    assert level in ("ops", "modules")
    sys.modules["tree_filter_cuda"] = tree_filter_cuda
    if level == "modules":
        for name in ("utils", "utils.TreeEnergyLoss", "utils.TreeEnergyLoss.kernels",
                     "utils.TreeEnergyLoss.kernels.lib_tree_filter",
                     "utils.TreeEnergyLoss.kernels.lib_tree_filter.modules"):
            if name not in sys.modules:
                try:
                    __import__(name)
                except Exception:
                    pkg = types.ModuleType(name)
                    pkg.__path__ = []
                    sys.modules[name] = pkg
        sys.modules["utils.gate_crf_loss"] = gate_crf_loss
        sys.modules["utils.TreeEnergyLoss.kernels.lib_tree_filter.modules.tree_filter"] = tree_filter
