"""Import the UNMODIFIED reference in this container -- TEST INFRASTRUCTURE ONLY, used to pin the oracle and to
generate the fixtures under tests/golden/.  It needs /root/reference, which does not exist on the GPU box, so nothing
in the `-m gpu` tests, smoke() or bench.py may call it.

The reference's third-party imports that are absent here (flwr, medpy, efficientnet_pytorch, tensorboardX, h5py) are
replaced by empty stand-ins; `tree_filter_cuda` (the pybind module the reference's functions/*.py import) is served by
oracle.tree_cpu.  See SURVEY.md §8(c).
"""
import importlib
import os
import sys
import types

REF_CODE = "/root/reference/code"


def available():
    return os.path.isdir(REF_CODE)


class _Anything:
    """Class usable as a base class, a callable and an attribute bag."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything()


class _AnyDict(dict):
    def __missing__(self, key):
        return ""


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)

    def _missing(attr):  # any other public name resolves to a fresh class
        if attr.startswith("__"):
            raise AttributeError(attr)
        return type(attr, (_Anything,), {})

    m.__getattr__ = _missing
    sys.modules[name] = m
    return m


def install_stubs():
    """Empty stand-ins for the third-party modules the reference imports and this image lacks.  Idempotent."""
    if "flwr" not in sys.modules:
        for name in ("flwr", "flwr.common", "flwr.common.logger", "flwr.client", "flwr.server", "flwr.server.server",
                     "flwr.server.history", "flwr.server.client_manager", "flwr.server.strategy",
                     "flwr.server.strategy.strategy", "flwr.server.strategy.fedavg", "flwr.server.strategy.fedadagrad",
                     "flwr.server.strategy.fedadam", "flwr.server.strategy.fedyogi", "medpy", "medpy.metric",
                     "tensorboardX", "h5py", "efficientnet_pytorch"):
            _stub(name)
        sys.modules["flwr"].client = sys.modules["flwr.client"]
        sys.modules["flwr"].common = sys.modules["flwr.common"]
        sys.modules["flwr"].server = sys.modules["flwr.server"]
        sys.modules["flwr.common.logger"].log = lambda *a, **k: None
        _stub("efficientnet_pytorch.utils", url_map=_AnyDict(), url_map_advprop=_AnyDict())


def install(tree_module=None):
    """Register the stand-ins and put the reference's code/ directory on sys.path.  Idempotent."""
    if not available():
        raise RuntimeError("the reference tree is not mounted at /root/reference")
    install_stubs()
    if tree_module is None:
        from oracle import tree_cpu as tree_module
    sys.modules["tree_filter_cuda"] = tree_module
    if REF_CODE not in sys.path:
        sys.path.insert(0, REF_CODE)


def gate_crf_loss():
    """The reference's code/utils/gate_crf_loss.py, imported as-is."""
    install()
    return importlib.import_module("utils.gate_crf_loss")


def tree_filter():
    """The reference's lib_tree_filter/modules/tree_filter.py (MinimumSpanningTree, TreeFilter2D), imported as-is."""
    install()
    return importlib.import_module("utils.TreeEnergyLoss.kernels.lib_tree_filter.modules.tree_filter")


def flower_common():
    """The reference's code/flower_common.py (MScaleRecurveTreeEnergyLoss, MyModel), imported as-is."""
    install()
    return importlib.import_module("flower_common")
