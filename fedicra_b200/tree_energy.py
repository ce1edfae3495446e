"""Tree-energy losses with the reference's class names and ``forward`` signatures
(/root/reference/code/flower_common.py:646-689 TreeEnergyLoss, :692-753 MScaleAddTreeEnergyLoss,
:756-818 MScaleRecurveTreeEnergyLoss -- the one FedICRA's training loop uses, OURS.py:64,142).

The reference's own classes also run unchanged on this package (they only need MinimumSpanningTree / TreeFilter2D,
flower_common.py:32-33); these mirrors exist so the loss stack is usable without importing flower_common.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .tree_filter import MinimumSpanningTree, TreeFilter2D, batched_tree_orders


def _prepare(preds, low_feats, unlabeled_ROIs):
    """flower_common.py:770-777: bring the image embedding and the ROI mask to the prediction size."""
    with torch.no_grad():
        h, w = preds.shape[2:]
        if low_feats.shape[2:] != (h, w):
            low_feats = F.interpolate(low_feats, size=(h, w), mode='bilinear', align_corners=False)
        roi = unlabeled_ROIs.unsqueeze(1).float()
        if roi.shape[2:] != (h, w):
            roi = F.interpolate(roi, size=(h, w), mode='nearest')
        return low_feats, roi, roi.sum()


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


def _masked_l1(roi, prob, filtered, n, distributed=False):
    return _normalise((roi * torch.abs(prob - filtered)).sum(), n, distributed)


class _TreeLossBase(nn.Module):
    def __init__(self, distributed=False):
        """distributed=True: the batch is one shard of a global batch split by image over torch.distributed ranks."""
        super().__init__()
        self.distributed = distributed
        self.mst_layers = MinimumSpanningTree(TreeFilter2D.norm2_distance)
        self.tree_filter_layers = TreeFilter2D(groups=1, sigma=0.02)

    def _trees(self, embeds):
        """Topology of all the embeddings of this step at once (they do not depend on the filter chain)."""
        if all(e.is_cuda for e in embeds) and len({(e.shape[0],) + tuple(e.shape[2:]) for e in embeds}) == 1:
            return batched_tree_orders(embeds)
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
        low_feats, roi, n = _prepare(preds, low_feats, unlabeled_ROIs)
        h, w = preds.shape[2:]
        prob = torch.softmax(preds, dim=1)
        highs = [] if high_feats is None else [_resize(high_feats, h, w)]
        trees = self._trees([low_feats] + highs)
        AS = self._low(prob, low_feats, trees[0])
        if highs:
            AS = self._high(AS, highs[0], h, w, trees[1])
        return weight * _masked_l1(roi, prob, AS, n, self.distributed), AS


class MScaleAddTreeEnergyLoss(_TreeLossBase):
    def forward(self, preds, low_feats, high_feats_1, high_feats_2, high_feats_3, unlabeled_ROIs, weight):
        low_feats, roi, n = _prepare(preds, low_feats, unlabeled_ROIs)
        h, w = preds.shape[2:]
        prob = torch.softmax(preds, dim=1)
        highs = [_resize(hf, h, w) for hf in (high_feats_1, high_feats_2, high_feats_3)]
        trees = self._trees([low_feats] + highs)
        AS = self._low(prob, low_feats, trees[0])
        outs = [self._high(AS, hf, h, w, t) for hf, t in zip(highs, trees[1:])]
        loss = _normalise(sum((roi * torch.abs(prob - o)).sum() for o in outs), n, self.distributed)
        return (weight * loss, *outs)


class MScaleRecurveTreeEnergyLoss(_TreeLossBase):
    def forward(self, preds, low_feats, high_feats_1, high_feats_2, high_feats_3, unlabeled_ROIs, weight):
        low_feats, roi, n = _prepare(preds, low_feats, unlabeled_ROIs)
        h, w = preds.shape[2:]
        prob = torch.softmax(preds, dim=1)
        highs = [_resize(hf, h, w) for hf in (high_feats_1, high_feats_2, high_feats_3)]
        trees = self._trees([low_feats] + highs)
        cur = self._low(prob, low_feats, trees[0])
        outs = []
        for hf, t in zip(highs, trees[1:]):
            cur = self._high(cur, hf, h, w, t)
            outs.append(cur)
        return (weight * _masked_l1(roi, prob, cur, n, self.distributed), *outs)
