"""Tree-energy losses with the reference's class names and ``forward`` signatures
(/root/reference/code/flower_common.py:646-689 TreeEnergyLoss, :692-753 MScaleAddTreeEnergyLoss,
:756-818 MScaleRecurveTreeEnergyLoss -- the one FedICRA's training loop uses, OURS.py:64,142).

The reference's own classes also run unchanged on this package (they only need MinimumSpanningTree / TreeFilter2D,
flower_common.py:32-33); these mirrors exist so the loss stack is usable without importing flower_common.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .tree_filter import MinimumSpanningTree, TreeFilter2D, batched_tree_orders, edge_weights_multi


def _prepare(preds, low_feats, unlabeled_ROIs):
    """flower_common.py:770-777: bring the image embedding and the ROI mask to the prediction size.  Returns the
    embedding, the float ROI + its sum for the generic path, and (when the mask already has the prediction size and the
    tensors are on the GPU) the mask as bytes for the fused masked-L1 kernel, in which case roi / N are not computed."""
    with torch.no_grad():
        h, w = preds.shape[2:]
        if low_feats.shape[2:] != (h, w):
            low_feats = F.interpolate(low_feats, size=(h, w), mode='bilinear', align_corners=False)
        if preds.is_cuda and unlabeled_ROIs.shape[-2:] == (h, w) and unlabeled_ROIs.dtype in (torch.bool, torch.uint8):
            roi_u8 = unlabeled_ROIs.contiguous().view(torch.uint8) if unlabeled_ROIs.dtype == torch.bool \
                else unlabeled_ROIs.contiguous()
            return low_feats, None, None, roi_u8
        roi = unlabeled_ROIs.unsqueeze(1).float()
        if roi.shape[2:] != (h, w):
            roi = F.interpolate(roi, size=(h, w), mode='nearest')
        return low_feats, roi, roi.sum(), None


class _BilinearHeads(torch.autograd.Function):
    """F.interpolate(head, size=(h, w), mode='bilinear', align_corners=False) of several heads in ONE launch per
    direction (csrc/glue.cu); the forward has torch's bits (the trees are built on them)."""

    @staticmethod
    def forward(ctx, h, w, *heads):
        ctx.hw, ctx.sizes = (h, w), [tuple(x.shape[2:]) for x in heads]
        return tuple(ops.bilinear_resize([x.contiguous() for x in heads], h, w))

    @staticmethod
    def backward(ctx, *grads):
        h, w = ctx.hw
        todo = [(k, g) for k, g in enumerate(grads) if g is not None and ctx.needs_input_grad[2 + k]]
        out = [None] * len(grads)
        if todo:
            res = ops.bilinear_resize([(g.contiguous(), ctx.sizes[k]) for k, g in todo], h, w, backward=True)
            for (k, _), r in zip(todo, res):
                out[k] = r
        return (None, None, *out)


def _resize_heads(heads, h, w):
    """All decoder heads brought to the prediction size (identity for those already there)."""
    out = list(heads)
    todo = [k for k, x in enumerate(heads) if x.shape[2:] != (h, w)]
    fused = [k for k in todo if heads[k].is_cuda and heads[k].dtype == torch.float32]
    if len(fused) == len(todo) and 1 <= len(fused) <= 4 and len({heads[k].shape[:2] for k in fused}) == 1:
        for k, r in zip(fused, _BilinearHeads.apply(h, w, *[heads[k] for k in fused])):
            out[k] = r
        return out
    return [_resize(x, h, w) for x in heads]


def _resize(feats, h, w):
    """Bilinear resize of a decoder head to the prediction size (identity when already there)."""
    if feats.shape[2:] == (h, w):
        return feats
    return F.interpolate(feats, size=(h, w), mode='bilinear', align_corners=False)


def _normalise(loss_sum, n, distributed):
    """`if N > 0: loss /= N` (flower_common.py:815-816) without the reference's device->host sync: N is a count, so
    dividing by max(N, 1) is the same value in both branches (the sum is 0 when N is 0).  With `distributed` the two
    scalars are summed over the ranks first (image sharding, fedicra_b200/sharding.py)."""
    if distributed:
        from .sharding import sharded_masked_l1
        return sharded_masked_l1(loss_sum, n)
    return loss_sum / torch.clamp(n, min=1.0)


class _ChannelSoftmax(torch.autograd.Function):
    """torch.softmax(preds, dim=1) (flower_common.py:781), one pass per direction (csrc/glue.cu)."""

    @staticmethod
    def forward(ctx, preds):
        prob = ops.channel_softmax_fwd(preds)
        ctx.save_for_backward(prob)
        return prob

    @staticmethod
    def backward(ctx, grad_prob):
        (prob,) = ctx.saved_tensors
        return ops.channel_softmax_bwd(prob, grad_prob.contiguous())


def _softmax(preds):
    if preds.is_cuda and preds.dtype == torch.float32 and preds.shape[1] <= 16:
        return _ChannelSoftmax.apply(preds.contiguous())
    return torch.softmax(preds, dim=1)


class _MaskedL1(torch.autograd.Function):
    """(ROI * |prob - filtered|).sum(), N and weight * sum / max(N, 1) in one pass (flower_common.py:777, 813-818).
    Returns (loss, sums): `loss` (0-d) is for the single-GPU path, `sums` = (sum, N) feeds the sharded normalisation
    (only sums[0] carries a gradient)."""

    @staticmethod
    def forward(ctx, prob, filtered, roi_u8, weight):
        sums, loss = ops.masked_l1_fwd(prob, filtered, roi_u8, weight)
        ctx.save_for_backward(prob, filtered, roi_u8, sums)
        ctx.weight = float(weight)
        ctx.set_materialize_grads(False)
        return loss.reshape(()), sums

    @staticmethod
    def backward(ctx, grad_loss, grad_sums):
        prob, filtered, roi_u8, sums = ctx.saved_tensors
        want_p, want_f = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        gp = gf = None
        if want_p or want_f:
            # whichever output the caller used: d loss (scaled by weight / max(N, 1) inside the kernel) or d sums[0]
            for g, n in ((grad_loss, sums[1:2]), (None if grad_sums is None else grad_sums[0:1], None)):
                if g is None:
                    continue
                a, b = ops.masked_l1_bwd(prob, filtered, roi_u8, g.contiguous(), n, ctx.weight, want_p, want_f)
                gp = a if gp is None else (gp + a if a is not None else gp)
                gf = b if gf is None else (gf + b if b is not None else gf)
        return gp, gf, None, None


def _masked_l1(roi, prob, filtered, n, distributed=False, weight=1.0, roi_u8=None, n_global=None):
    """weight * (roi * |prob - filtered|).sum() / N   (N over the global batch when `distributed`).  With `n_global` (a
    sharding.GlobalCount started before the step) the value is this rank's SHARE of the global loss -- local sum over
    global N; the shares of all ranks add up to the reference's value, the gradients are the reference's -- and no
    collective sits between forward and backward."""
    if isinstance(weight, torch.Tensor):   # keep a tensor weight on the device: scale outside the kernel
        return weight * _masked_l1(roi, prob, filtered, n, distributed, 1.0, roi_u8, n_global)
    if roi_u8 is not None and prob.is_cuda and prob.dtype == torch.float32 and filtered.dtype == torch.float32:
        loss, sums = _MaskedL1.apply(prob.contiguous(), filtered.contiguous(), roi_u8, weight)
        if n_global is not None:
            return weight * sums[0] / torch.clamp(n_global.get()[0], min=1.0)
        if not distributed:
            return loss
        return weight * _normalise(sums[0], sums[1].detach(), True)
    return weight * _normalise((roi * torch.abs(prob - filtered)).sum(), n, distributed)


class _TreeLossBase(nn.Module):
    def __init__(self, distributed=False, local_share=False):
        """distributed=True: the batch is one shard of a global batch split by image over torch.distributed ranks; the
        returned loss is the GLOBAL value (one 2-scalar all-reduce between forward and backward).  With
        local_share=True it is this rank's share of it (the shares add up to the global value, gradients identical) and
        the only collective, the count of unlabelled pixels, runs on a side stream under the step."""
        super().__init__()
        self.distributed = distributed
        self.local_share = bool(distributed and local_share)
        self.mst_layers = MinimumSpanningTree(TreeFilter2D.norm2_distance)
        self.tree_filter_layers = TreeFilter2D(groups=1, sigma=0.02)
        # called (no arguments) when the topology kernels of a forward have been enqueued and the filters have not: the
        # point where a caller with independent work for a helper stream (runner.py: the Gated CRF term) forks it
        self.on_topology_enqueued = None

    def _count(self, roi_u8):
        if not (self.local_share and roi_u8 is not None):
            return None
        from .sharding import GlobalCount
        return GlobalCount().start(roi_u8)

    def _trees(self, embeds, channels=None):
        """Topology of all the embeddings of this step at once (they do not depend on the filter chain), and with it the
        filters' edge weights: embedding 0 is the image (low-level tree, 1/sigma), the others are decoder heads (1)."""
        if all(e.is_cuda and e.shape[1] <= 4 for e in embeds) and len({(e.shape[0],) + tuple(e.shape[2:]) for e in embeds}) == 1:
            inv = [1.0 / self.tree_filter_layers.sigma] + [1.0] * (len(embeds) - 1)
            orders = batched_tree_orders(embeds, inv)
            B, V = embeds[0].shape[0], embeds[0].shape[2] * embeds[0].shape[3]
            if channels is not None and channels <= ops.CHAIN_MAX_C and V <= ops.CHAIN_MAX_V and torch.is_grad_enabled():
                # ONE memset arms the buffers of every filter launch of this step (forward and backward of each tree)
                # instead of one in front of each launch.  (Measured: on a helper stream beside the topology kernels it
                # gains nothing more in graph replay and costs 80 us of host time per eager step.)
                chains = [a for a in (o.for_channels(channels) for o in orders) if a.kind == "chain"]
                if chains:
                    bufs = ops.chain_arm(B, channels, V, 2 * len(chains), embeds[0].device)
                    for a in chains:
                        a.armed = [bufs.pop(), bufs.pop()]
            # the heads' edge weights as ONE autograd node: one backward launch for all of them
            if channels is not None and torch.is_grad_enabled():
                joint = [(o.for_channels(channels), e, s) for o, e, s in zip(orders[1:], embeds[1:], inv[1:]) if e.requires_grad]
                joint = [t for t in joint if t[0].kind == "chain" and t[0].ew is not None]
                if len(joint) >= 2:
                    edge_weights_multi([t[0] for t in joint], [t[1] for t in joint], [t[2] for t in joint])
            if self.on_topology_enqueued is not None:
                self.on_topology_enqueued()
            return orders
        return [self.mst_layers(e) for e in embeds]

    def _low(self, prob, low_feats, tree=None):
        tree = self.mst_layers(low_feats) if tree is None else tree
        return self.tree_filter_layers(feature_in=prob, embed_in=low_feats, tree=tree)

    def _high(self, feature, high_feats, h, w, tree=None):
        high_feats = _resize(high_feats, h, w)
        tree = self.mst_layers(high_feats) if tree is None else tree
        return self.tree_filter_layers(feature_in=feature, embed_in=high_feats, tree=tree, low_tree=False)


class TreeEnergyLoss(_TreeLossBase):
    def forward(self, preds, low_feats, high_feats, unlabeled_ROIs, weight):
        low_feats, roi, n, roi_u8 = _prepare(preds, low_feats, unlabeled_ROIs)
        h, w = preds.shape[2:]
        prob = _softmax(preds)
        highs = [] if high_feats is None else _resize_heads([high_feats], h, w)
        trees = self._trees([low_feats] + highs, prob.shape[1])
        AS = self._low(prob, low_feats, trees[0])
        if highs:
            AS = self._high(AS, highs[0], h, w, trees[1])
        return _masked_l1(roi, prob, AS, n, self.distributed, weight, roi_u8), AS


class MScaleAddTreeEnergyLoss(_TreeLossBase):
    def forward(self, preds, low_feats, high_feats_1, high_feats_2, high_feats_3, unlabeled_ROIs, weight):
        low_feats, roi, n, roi_u8 = _prepare(preds, low_feats, unlabeled_ROIs)
        h, w = preds.shape[2:]
        prob = _softmax(preds)
        highs = _resize_heads([high_feats_1, high_feats_2, high_feats_3], h, w)
        trees = self._trees([low_feats] + highs, prob.shape[1])
        AS = self._low(prob, low_feats, trees[0])
        outs = [self._high(AS, hf, h, w, t) for hf, t in zip(highs, trees[1:])]
        # flower_common.py:748-751: the three masked-L1 sums share one normaliser
        loss = sum(_masked_l1(roi, prob, o, n, self.distributed, weight, roi_u8) for o in outs)
        return (loss, *outs)


class MScaleRecurveTreeEnergyLoss(_TreeLossBase):
    def forward(self, preds, low_feats, high_feats_1, high_feats_2, high_feats_3, unlabeled_ROIs, weight):
        low_feats, roi, n, roi_u8 = _prepare(preds, low_feats, unlabeled_ROIs)
        n_global = self._count(roi_u8)
        h, w = preds.shape[2:]
        prob = _softmax(preds)
        highs = _resize_heads([high_feats_1, high_feats_2, high_feats_3], h, w)
        trees = self._trees([low_feats] + highs, prob.shape[1])
        cur = self._low(prob, low_feats, trees[0])
        outs = []
        for hf, t in zip(highs, trees[1:]):
            cur = self._high(cur, hf, h, w, t)
            outs.append(cur)
        return (_masked_l1(roi, prob, cur, n, self.distributed, weight, roi_u8, n_global), *outs)
