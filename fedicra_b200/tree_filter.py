"""MinimumSpanningTree and TreeFilter2D with the reference's constructors and ``forward`` signatures
(/root/reference/code/utils/TreeEnergyLoss/kernels/lib_tree_filter/modules/tree_filter.py:9-62, 65-151), on the CUDA
kernels of libfedicra_sm100.so.

``MinimumSpanningTree.forward`` returns the same ``[B, V-1, 2]`` int32 edge tensor as the reference; the BFS ordering
is computed once per tree and remembered on that tensor object, so the four filters of the multiscale loss do not
re-derive it.
"""
import torch
from torch import nn

from . import ops


def _is_norm2(fn):
    return fn is None or getattr(fn, "__name__", "") == "norm2_distance"


class _TreeOrder:
    """BFS arrays of a batch of trees (sorted_index, sorted_parent, sorted_child, level_start, n_levels)."""

    __slots__ = ("si", "sp", "sc", "ls", "nl", "bands")

    def __init__(self, tree=None, arrays=None):
        if arrays is not None:
            self.si, self.sp, self.sc, self.ls, self.nl, self.bands = arrays
            return
        mask, hw = getattr(tree, "_fedicra_mask", None), getattr(tree, "_fedicra_hw", None)
        arrays = ops.bfs_from_mask(mask, *hw, root=ops.default_root(*hw)) if mask is not None else None
        self.si, self.sp, self.sc, self.ls, self.nl = arrays if arrays is not None else ops.bfs_forward(tree)
        self.bands = ops.band_plan(self.ls, self.nl)


def batched_tree_orders(embeds):
    """MST + BFS of several embeddings of one batch in ONE pass: the embeddings ([B,D_k,H,W], same B/H/W) are stacked
    along the batch axis -- by a pointer table, not by a copy -- so that each kernel of the topology stage is launched
    once for all of them.  Returns one _TreeOrder per embedding."""
    B, _, H, W = embeds[0].shape
    with torch.no_grad():
        fits = H * W <= max(ops.BFS_GRID_MAX_V, ops.BFS_EULER_MAX_V)
        if len(embeds) <= 8:   # one launch straight from the K tensors (pointer table), nothing stacked
            tree, mask = ops.mst_grid_multi([e.detach().float() for e in embeds], want_mask=True, want_edges=not fits)
        else:
            D = max(e.shape[1] for e in embeds)
            stacked = torch.cat([torch.nn.functional.pad(e.detach().float(), (0, 0, 0, 0, 0, D - e.shape[1])) for e in embeds], 0)
            tree, mask = ops.mst_grid(stacked, want_mask=True, want_edges=not fits)  # the mask alone feeds the grid BFS
        arrays = ops.bfs_from_mask(mask, H, W, root=ops.default_root(H, W)) if fits else ops.bfs_forward(tree)
        bands = ops.band_plan(arrays[3], arrays[4]).view(len(embeds) * B, -1)
    return [_TreeOrder(arrays=tuple(a[k * B:(k + 1) * B] for a in (*arrays, bands))) for k in range(len(embeds))]


def tree_order(tree):
    if isinstance(tree, _TreeOrder):
        return tree
    order = getattr(tree, "_fedicra_order", None)
    if order is None:
        order = _TreeOrder(tree)
        try:
            tree._fedicra_order = order
        except AttributeError:
            pass
    return order


class MinimumSpanningTree(nn.Module):
    def __init__(self, distance_func):
        super(MinimumSpanningTree, self).__init__()
        self.distance_func = distance_func

    @staticmethod
    def _build_matrix_index(fm):
        """[B,E,2] int32 grid edge list, vertical edges first (tree_filter.py:15-25); only used by the generic path."""
        batch, height, width = (fm.shape[0], *fm.shape[2:])
        ids = torch.arange(height * width, dtype=torch.int32, device=fm.device).view(height, width)
        vert = torch.stack([ids[:-1, :], ids[1:, :]], 2).reshape(1, -1, 2)
        horz = torch.stack([ids[:, :-1], ids[:, 1:]], 2).reshape(1, -1, 2)
        return torch.cat([vert, horz], 1).expand(batch, -1, -1)

    def _build_feature_weight(self, fm):
        batch = fm.shape[0]
        w_v = self.distance_func(fm[:, :, :-1, :], fm[:, :, 1:, :]).reshape(batch, -1)
        w_h = self.distance_func(fm[:, :, :, :-1], fm[:, :, :, 1:]).reshape(batch, -1)
        return torch.cat([w_v, w_h], dim=1) + 1

    def forward(self, guide_in, label=None):
        if label is not None:
            raise NotImplementedError("fedicra_b200 MinimumSpanningTree: the `label` branch is not implemented")
        with torch.no_grad():
            if _is_norm2(self.distance_func):
                # weights + MST fused, edge list never materialised; the adjacency mask rides along for the BFS
                tree, mask = ops.mst_grid(guide_in.detach().float(), want_mask=True)
                tree._fedicra_mask, tree._fedicra_hw = mask, tuple(guide_in.shape[2:])
            else:
                index = self._build_matrix_index(guide_in).contiguous()
                weight = self._build_feature_weight(guide_in).contiguous()
                tree = ops.mst_forward(index, weight, guide_in.shape[2] * guide_in.shape[3])
        return tree


class _EdgeWeight(torch.autograd.Function):
    """exp(-|e_v - e_parent|^2 / sigma) in BFS order, with its gradient w.r.t. the embedding."""

    @staticmethod
    def forward(ctx, embed, order, inv_sigma):
        ew = ops.edge_weight_fwd(embed, order.si, order.sp, inv_sigma)
        ctx.order, ctx.inv_sigma = order, inv_sigma
        ctx.save_for_backward(embed, ew)
        return ew

    @staticmethod
    def backward(ctx, grad_ew):
        embed, ew = ctx.saved_tensors
        o = ctx.order
        return ops.edge_weight_bwd(embed, o.si, o.sp, o.sc, ew, grad_ew, ctx.inv_sigma), None, None


class _Refine(torch.autograd.Function):
    """Two-pass tree filter (depth-banded kernels, csrc/bands.cu); same gradient policy as functions/refine.py:23-40
    (weight gradient for the high-level trees only, and only when the edge weights need one)."""

    @staticmethod
    def forward(ctx, feature_in, edge_weight, order, low_tree):
        out, saved = ops.tree_filter_forward(feature_in, edge_weight, order.si, order.sp, order.sc, order.ls, order.nl,
                                             order.bands)
        ctx.order, ctx.low_tree = order, low_tree
        ctx.save_for_backward(edge_weight, saved)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        edge_weight, saved = ctx.saved_tensors
        o = ctx.order
        want_w = (not ctx.low_tree) and ctx.needs_input_grad[1]
        grad_feature, grad_weight = ops.tree_filter_backward(grad_out.contiguous(), edge_weight, o.si, o.sp, o.sc, o.ls,
                                                             o.nl, o.bands, saved, want_w)
        return grad_feature, grad_weight, None, None


class TreeFilter2D(nn.Module):
    def __init__(self, groups=1, sigma=0.02, distance_func=None, enable_log=False):
        super(TreeFilter2D, self).__init__()
        self.groups = groups
        self.enable_log = enable_log
        self.distance_func = self.norm2_distance if distance_func is None else distance_func
        self.sigma = sigma

    @staticmethod
    def norm2_distance(fm_ref, fm_tar):
        diff = fm_ref - fm_tar
        return (diff * diff).sum(dim=1)

    def build_edge_weight(self, fm, order, low_tree):
        """tree_filter.py:91-108; `order` is this library's BFS record instead of the two raw index tensors."""
        batch, channel = fm.shape[0], fm.shape[1]
        fm = fm.reshape(batch, channel, -1)
        if _is_norm2(self.distance_func):
            return _EdgeWeight.apply(fm.float(), order, 1.0 / self.sigma if low_tree else 1.0)
        idx = order.si.long().unsqueeze(1).expand(-1, channel, -1)
        src = torch.gather(fm, 2, idx)
        dst = torch.gather(src, 2, order.sp.long().unsqueeze(1).expand(-1, channel, -1))
        dist = self.distance_func(src, dst)
        return torch.exp(-dist / self.sigma) if low_tree else torch.exp(-dist)

    def forward(self, feature_in, embed_in, tree, low_tree=True):
        if self.groups != 1:
            raise NotImplementedError("fedicra_b200 TreeFilter2D: groups != 1 is not implemented (FedICRA uses 1)")
        ori_shape = feature_in.shape
        order = tree_order(tree)
        edge_weight = self.build_edge_weight(embed_in, order, low_tree)
        self.enable_log = False
        feature = feature_in.reshape(ori_shape[0], ori_shape[1], -1).float()
        out = _Refine.apply(feature, edge_weight.contiguous(), order, low_tree)
        return out.reshape(ori_shape)
