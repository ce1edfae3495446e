"""Stand-in for the reference's pybind11 extension module ``tree_filter_cuda``
(/root/reference/code/utils/TreeEnergyLoss/kernels/lib_tree_filter/src/tree_filter.cpp:7-13).

Same five functions, same argument lists and tensor layouts, so the reference's own
``functions/{mst,bfs,refine}.py`` and ``modules/tree_filter.py`` run on it unmodified once it is registered as
``sys.modules['tree_filter_cuda']`` (see fedicra_b200.dropin).  Everything executes in libfedicra_sm100.so.

Differences a caller can observe:
  * ``mst_forward`` returns the tree edges in increasing edge index (the reference: in Boruvka round order); the edge
    SET is identical;
  * ``bfs_forward`` returns the canonical level order (the reference: an atomic-dependent topological order) and only
    supports ``max_adj_per_vertex == 4`` (the only value the reference ever passes, tree_filter.py:135);
  * the refine operators need the sorted_* arrays of this module's ``bfs_forward`` (they rely on its level structure,
    remembered per ``sorted_index`` address, the tensor kept alive) and raise on foreign arrays.
"""
import collections

import torch

from . import ops

_LEVELS = collections.OrderedDict()   # sorted_index.data_ptr() -> (sorted_index itself, level_start, n_levels)
_LEVELS_MAX = 32


def _remember_levels(si, ls, nl):
    """The reference's autograd glue hands the refine operators re-wrapped tensors (ctx.saved_tensors), so the level
    structure cannot ride on the tensor object; it is looked up by address.  The entry holds a reference to the
    sorted_index tensor itself: while it is listed its memory cannot be freed and handed to another tree, so an address
    hit always is the same tree.  The last 32 orderings are kept; an older one raises instead of guessing."""
    _LEVELS[si.data_ptr()] = (si, ls, nl)
    _LEVELS.move_to_end(si.data_ptr())
    while len(_LEVELS) > _LEVELS_MAX:
        _LEVELS.popitem(last=False)


def _levels_of(si):
    hit = _LEVELS.get(si.data_ptr())
    if hit is None or hit[0].shape != si.shape or hit[0].device != si.device:
        raise RuntimeError("tree_filter_cuda.refine_*: sorted_index was not produced by one of this module's last "
                           f"{_LEVELS_MAX} bfs_forward calls (the kernels need its level structure)")
    return hit[1], hit[2]


def mst_forward(edge_index, edge_weight, vertex_count):
    """mst.hpp:4-7.  [B,E,2] int32, [B,E] float32, V -> [B,V-1,2] int32 (stays on the device, no host round trip)."""
    return ops.mst_forward(edge_index, edge_weight, vertex_count)


def bfs_forward(edge_index, max_adj_per_vertex):
    """bfs.hpp:4-8.  [B,V-1,2] int32 -> (sorted_index [B,V], sorted_parent [B,V], sorted_child [B,V,4])."""
    if max_adj_per_vertex != 4:
        raise NotImplementedError("bfs_forward: only max_adj_per_vertex == 4 (4-connected grid trees) is implemented")
    si, sp, sc, ls, nl = ops.bfs_forward(edge_index)
    _remember_levels(si, ls, nl)
    return si, sp, sc


def refine_forward(feature_in, edge_weight, sorted_index, sorted_parent, sorted_child):
    """refine.hpp:4-11.  Returns (feature_out, feature_aggr, feature_aggr_up, weight_sum, weight_sum_up)."""
    ls, nl = _levels_of(sorted_index)
    return ops.refine_forward(feature_in, edge_weight, sorted_index, sorted_parent, sorted_child, ls, nl)


def refine_backward_feature(feature_in, edge_weight, sorted_index, sorted_parent, sorted_child, feature_out,
                            feature_aggr, feature_aggr_up, weight_sum, weight_sum_up, grad_out):
    """refine.hpp:13-25."""
    ls, nl = _levels_of(sorted_index)
    return ops.refine_backward_feature(edge_weight, sorted_index, sorted_parent, sorted_child, ls, nl, weight_sum,
                                       grad_out)


def refine_backward_weight(feature_in, edge_weight, sorted_index, sorted_parent, sorted_child, feature_out,
                           feature_aggr, feature_aggr_up, weight_sum, weight_sum_up, grad_out):
    """refine.hpp:27-39."""
    ls, nl = _levels_of(sorted_index)
    return ops.refine_backward_weight(edge_weight, sorted_index, sorted_parent, sorted_child, ls, nl, feature_out,
                                      feature_aggr, feature_aggr_up, weight_sum, weight_sum_up, grad_out)
