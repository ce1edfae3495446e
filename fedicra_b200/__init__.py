"""fedicra_b200 -- B200 (sm_100a) implementation of FedICRA's weak-supervision loss hot path.

Same Python names and call signatures as the reference (llmir/FedICRA):

    ModelLossSemsegGatedCRF                       code/utils/gate_crf_loss.py:5
    MinimumSpanningTree, TreeFilter2D             code/utils/TreeEnergyLoss/kernels/lib_tree_filter/modules/tree_filter.py:9,65
    tree_filter_cuda (5 operators)                .../lib_tree_filter/src/tree_filter.cpp:7-13
    TreeEnergyLoss, MScaleAddTreeEnergyLoss,
    MScaleRecurveTreeEnergyLoss                   code/flower_common.py:646-818
    adaptive aggregation / FedAvg                 code/flower_common.py:491-624, flwr FedAvg

All arithmetic runs in hand-written CUDA kernels behind the C ABI of ``libfedicra_sm100.so`` (include/fedicra_sm100.h).
There is no CPU path: importing the operators without the built library raises.
"""
from . import _lib  # noqa: F401  (fails loudly when the library is missing)
from .gate_crf_loss import ModelLossSemsegGatedCRF
from .tree_filter import MinimumSpanningTree, TreeFilter2D
from .tree_energy import MScaleAddTreeEnergyLoss, MScaleRecurveTreeEnergyLoss, TreeEnergyLoss
from .aggregation import AdaptiveAggregation, FlatModelState, fedavg_allreduce, fused_ala_update
from .runner import GraphedLossStack, LossStackStep, PipelinedLossStack
from .adaptive import adaptive_set_weights, install_set_weights
from . import tree_filter_cuda
from .dropin import install as install_dropin

__all__ = [
    "ModelLossSemsegGatedCRF", "MinimumSpanningTree", "TreeFilter2D", "TreeEnergyLoss", "MScaleAddTreeEnergyLoss",
    "MScaleRecurveTreeEnergyLoss", "AdaptiveAggregation", "FlatModelState", "fedavg_allreduce", "fused_ala_update",
    "tree_filter_cuda", "install_dropin", "LossStackStep", "GraphedLossStack", "PipelinedLossStack",
    "adaptive_set_weights", "install_set_weights",
]
