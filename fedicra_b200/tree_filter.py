"""MinimumSpanningTree and TreeFilter2D with the reference's constructors and ``forward`` signatures
(/root/reference/code/utils/TreeEnergyLoss/kernels/lib_tree_filter/modules/tree_filter.py:9-62, 65-151), on the CUDA
kernels of libfedicra_sm100.so.

``MinimumSpanningTree.forward`` returns the same ``[B, V-1, 2]`` int32 edge tensor as the reference; the BFS ordering
is computed once per tree and remembered on that tensor object, so the four filters of the multiscale loss do not
re-derive it.
"""
import os

import torch
from torch import nn

from . import ops


_SIDE = {}


def side_stream(tag="crf"):
    """Per-device helper stream for work that is off the critical path of the step (the Gated CRF term beside the tree
    term, runner.py).  FEDICRA_OVERLAP=0 keeps everything on the caller's stream.  Measured and dropped: the edge-weight
    gradients of the high-level trees on a helper stream (+7 us per step: the chain filter's persistent CTAs hold the
    registers of every SM, so nothing runs beside them)."""
    want = os.environ.get("FEDICRA_OVERLAP", "1")
    if want == "0" or (want != "1" and tag not in want.split(",")):
        return None
    key = (torch.cuda.current_device(), tag)
    if key not in _SIDE:
        _SIDE[key] = torch.cuda.Stream()
    return _SIDE[key]


def _is_norm2(fn):
    """True when `fn` is this module's squared-L2 distance (or absent): only then may the fused weight + MST and
    edge-weight kernels stand in for it.  Identity, not name: a user function that happens to be called
    norm2_distance takes the generic path."""
    return fn is None or fn is TreeFilter2D.norm2_distance


class _Arrays:
    """One valid ordering of a batch of rooted trees in the reference's format (sorted_index, sorted_parent,
    sorted_child; bfs.cu:46-98) plus what this library's filter kernels need on top of it:
      kind "chain": group_start (heavy-chain order, csrc/chains.cu);  kind "bfs": level_start, n_levels, band plan."""

    __slots__ = ("kind", "si", "sp", "sc", "gs", "ls", "nl", "bands", "ew", "ew_key", "armed", "ew_out")

    def __init__(self, kind, si, sp, sc, gs=None, ls=None, nl=None, bands=None, ew=None, ew_key=None):
        self.kind, self.si, self.sp, self.sc, self.gs, self.ls, self.nl, self.bands = kind, si, sp, sc, gs, ls, nl, bands
        # edge weights that came out of the ordering pass, and what they were computed from: (embedding address,
        # shape, 1/sigma) -- build_edge_weight takes them only for exactly that request
        self.ew, self.ew_key = ew, ew_key
        # chain kind: buffers armed ahead for the filter launches on these trees (ops.chain_arm), each used once
        self.armed = []
        # the same edge weights as an output of the step's joint autograd node (edge_weights_multi), if there is one
        self.ew_out = None


def _want_chain(V):
    """The scan-based chain filter is the default; FEDICRA_FILTER=resident|multi selects the depth-banded one."""
    return os.environ.get("FEDICRA_FILTER", "chain") == "chain" and V <= ops.CHAIN_MAX_V


class _TreeOrder:
    """Topology record of a batch of trees: the adjacency mask (or edge list) and, built on first use, the orderings the
    filters run on.  `for_channels(C)` picks the one the filter for C feature channels takes."""

    def __init__(self, tree=None, mask=None, hw=None, chain=None, bfs=None):
        self.tree, self.mask, self.hw = tree, mask, hw
        if tree is not None and mask is None:
            self.mask, self.hw = getattr(tree, "_fedicra_mask", None), getattr(tree, "_fedicra_hw", None)
        self._chain, self._bfs = chain, bfs

    def chain(self):
        if self._chain is None:
            self._chain = _Arrays("chain", *ops.chain_order(self.mask, *self.hw))
        return self._chain

    def bfs(self):
        if self._bfs is None:
            arrays = ops.bfs_from_mask(self.mask, *self.hw, root=ops.default_root(*self.hw)) if self.mask is not None else None
            si, sp, sc, ls, nl = arrays if arrays is not None else ops.bfs_forward(self.tree)
            self._bfs = _Arrays("bfs", si, sp, sc, ls=ls, nl=nl, bands=ops.band_plan(ls, nl))
        return self._bfs

    def for_channels(self, C):
        if self.mask is not None and C <= ops.CHAIN_MAX_C and _want_chain(self.hw[0] * self.hw[1]):
            return self.chain()
        return self.bfs()


def _ew_key(embed, inv_sigma):
    return (embed.data_ptr(), embed.shape[0], embed.shape[1], int(embed.numel()), float(inv_sigma))


def batched_tree_orders(embeds, inv_sigmas=None):
    """MST + ordering of several embeddings of one batch in ONE pass: the embeddings ([B,D_k,H,W], same B/H/W) are
    stacked along the batch axis -- by a pointer table, not by a copy -- so that each kernel of the topology stage is
    launched once for all of them.  With `inv_sigmas` (one per embedding) the filters' edge weights come out of the same
    pass.  Returns one _TreeOrder per embedding."""
    B, _, H, W = embeds[0].shape
    K = len(embeds)
    with torch.no_grad():
        fits = H * W <= max(ops.BFS_GRID_MAX_V, ops.BFS_EULER_MAX_V)
        if K <= 8:   # one launch straight from the K tensors (pointer table), nothing stacked
            tree, mask = ops.mst_grid_multi([e.detach().float() for e in embeds], want_mask=True, want_edges=not fits)
        else:
            D = max(e.shape[1] for e in embeds)
            stacked = torch.cat([torch.nn.functional.pad(e.detach().float(), (0, 0, 0, 0, 0, D - e.shape[1])) for e in embeds], 0)
            tree, mask = ops.mst_grid(stacked, want_mask=True, want_edges=not fits)  # the mask alone feeds the ordering
        cut = lambda a, k: a[k * B:(k + 1) * B]
        if _want_chain(H * W):
            if inv_sigmas is not None and K <= 8:
                srcs = [e.detach().float().contiguous() for e in embeds]
                *arrays, ew = ops.chain_order_weights(mask, srcs, inv_sigmas, H, W)
                return [_TreeOrder(mask=cut(mask, k), hw=(H, W),
                                   chain=_Arrays("chain", *(cut(a, k) for a in arrays), ew=cut(ew, k),
                                                 ew_key=_ew_key(srcs[k], inv_sigmas[k]))) for k in range(K)]
            arrays = ops.chain_order(mask, H, W)
            return [_TreeOrder(mask=cut(mask, k), hw=(H, W), chain=_Arrays("chain", *(cut(a, k) for a in arrays)))
                    for k in range(K)]
        si, sp, sc, ls, nl = ops.bfs_from_mask(mask, H, W, root=ops.default_root(H, W)) if fits else ops.bfs_forward(tree)
        bands = ops.band_plan(ls, nl).view(K * B, -1)
    return [_TreeOrder(tree=None if tree is None else cut(tree, k), mask=cut(mask, k), hw=(H, W),
                       bfs=_Arrays("bfs", cut(si, k), cut(sp, k), cut(sc, k), ls=cut(ls, k), nl=cut(nl, k),
                                   bands=cut(bands, k))) for k in range(K)]


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
            if _is_norm2(self.distance_func) and guide_in.shape[1] <= 4:
                # weights + MST fused, edge list never materialised; the adjacency mask rides along for the ordering.
                # Up to 4 channels the kernel's in-order channel sum has the bits of torch's reduction (the MST is
                # decided on those bits); wider embeddings take the path below, on torch's own weights.
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
        if order.ew is not None and order.ew_key == _ew_key(embed, inv_sigma):
            ew, order.ew = order.ew, None      # computed by the ordering pass from this very tensor; used once (the
        else:                                  # filter writes w_0 = 0 into it)
            ew = ops.edge_weight_fwd(embed, order.si, order.sp, inv_sigma)
        ctx.order, ctx.inv_sigma = order, inv_sigma
        ctx.save_for_backward(embed, ew)
        return ew

    @staticmethod
    def backward(ctx, grad_ew):
        embed, ew = ctx.saved_tensors
        o = ctx.order
        return ops.edge_weight_bwd(embed, o.si, o.sp, o.sc, ew, grad_ew, ctx.inv_sigma), None, None


class _EdgeWeightMulti(torch.autograd.Function):
    """Edge weights of the trees of SEVERAL embeddings of one batch as one autograd node: the weights came out of the
    ordering pass, and the backward -- the gradients w.r.t. all the embeddings -- is one launch, issued when the last
    of their filters has delivered its weight gradient, instead of one launch between every two filter backwards."""

    @staticmethod
    def forward(ctx, orders, inv_sigmas, *embeds):
        ews = []
        for o in orders:
            ews.append(o.ew)
            o.ew = None
        ctx.orders, ctx.inv_sigmas = orders, inv_sigmas
        ctx.save_for_backward(*embeds, *ews)
        ctx.set_materialize_grads(False)
        return tuple(e.view_as(e) for e in ews)

    @staticmethod
    def backward(ctx, *grad_ews):
        K = len(ctx.orders)
        embeds, ews = ctx.saved_tensors[:K], ctx.saved_tensors[K:]
        live = [k for k in range(K) if grad_ews[k] is not None and ctx.needs_input_grad[2 + k]]
        out = [None] * K
        if live:
            grads = ops.edge_weight_bwd_multi([embeds[k] for k in live],
                                              [(ctx.orders[k].si, ctx.orders[k].sp, ctx.orders[k].sc) for k in live],
                                              [ews[k] for k in live], [grad_ews[k] for k in live],
                                              [ctx.inv_sigmas[k] for k in live])
            for k, g in zip(live, grads):
                out[k] = g
        return (None, None, *out)


def edge_weights_multi(orders, embeds, inv_sigmas):
    """Joins the edge weights that batched_tree_orders computed for `embeds` ([B,D,H,W], requiring grad) into one autograd
    node; `TreeFilter2D.build_edge_weight` then hands out its outputs.  orders: the chain `_Arrays` of those embeddings."""
    flat = [e.reshape(e.shape[0], e.shape[1], -1).float() for e in embeds]
    if any(o.ew is None or o.ew_key != _ew_key(f, s) for o, f, s in zip(orders, flat, inv_sigmas)):
        return
    outs = _EdgeWeightMulti.apply(tuple(orders), tuple(float(s) for s in inv_sigmas), *flat)
    for o, w in zip(orders, outs):
        o.ew_out = w


class _Refine(torch.autograd.Function):
    """Two-pass tree filter (affine scans along heavy chains, csrc/chains.cu; or the depth-banded kernels,
    csrc/bands.cu); same gradient policy as functions/refine.py:23-40 (weight gradient for the high-level trees only,
    and only when the edge weights need one)."""

    @staticmethod
    def forward(ctx, feature_in, edge_weight, order, low_tree):
        if order.kind == "chain":
            out, saved = ops.chain_filter_forward(feature_in, edge_weight, order.si, order.sp, order.sc, order.gs,
                                                  saved=order.armed.pop() if order.armed else None)
        else:
            out, saved = ops.tree_filter_forward(feature_in, edge_weight, order.si, order.sp, order.sc, order.ls,
                                                 order.nl, order.bands)
        ctx.order, ctx.low_tree = order, low_tree
        ctx.save_for_backward(edge_weight, saved)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        edge_weight, saved = ctx.saved_tensors
        o = ctx.order
        want_w = (not ctx.low_tree) and ctx.needs_input_grad[1]
        if o.kind == "chain":
            grad_feature, grad_weight = ops.chain_filter_backward(grad_out.contiguous(), edge_weight, o.si, o.sp, o.sc,
                                                                  o.gs, saved, want_w,
                                                                  armed_workspace=o.armed.pop() if o.armed else None)
        else:
            grad_feature, grad_weight = ops.tree_filter_backward(grad_out.contiguous(), edge_weight, o.si, o.sp, o.sc,
                                                                 o.ls, o.nl, o.bands, saved, want_w)
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
            inv_sigma = 1.0 / self.sigma if low_tree else 1.0
            if order.ew_out is not None and order.ew_key == _ew_key(fm.float(), inv_sigma):
                out, order.ew_out = order.ew_out, None      # used once: the filter writes w_0 = 0 into it
                return out
            return _EdgeWeight.apply(fm.float(), order, inv_sigma)
        idx = order.si.long().unsqueeze(1).expand(-1, channel, -1)
        src = torch.gather(fm, 2, idx)
        dst = torch.gather(src, 2, order.sp.long().unsqueeze(1).expand(-1, channel, -1))
        dist = self.distance_func(src, dst)
        return torch.exp(-dist / self.sigma) if low_tree else torch.exp(-dist)

    def forward(self, feature_in, embed_in, tree, low_tree=True):
        if self.groups != 1:
            raise NotImplementedError("fedicra_b200 TreeFilter2D: groups != 1 is not implemented (FedICRA uses 1)")
        ori_shape = feature_in.shape
        order = tree_order(tree).for_channels(ori_shape[1])
        edge_weight = self.build_edge_weight(embed_in, order, low_tree)
        self.enable_log = False
        feature = feature_in.reshape(ori_shape[0], ori_shape[1], -1).float()
        out = _Refine.apply(feature, edge_weight.contiguous(), order, low_tree)
        return out.reshape(ori_shape)
