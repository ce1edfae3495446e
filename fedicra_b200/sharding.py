"""Image sharding of the loss stack across the GPUs of one node.

Every image owns its trees, BFS orders, filters and CRF windows, so a contiguous split of the batch needs no data-path
collective.  The only cross-image coupling is the tree-energy normaliser N = ROI.sum() over the WHOLE batch
(flower_common.py:777, 813-816): ranks all-reduce the two scalars (masked-L1 sum, N) before dividing, which makes the
sharded loss and gradients equal the unsharded ones.  The Gated CRF denominator N*H*W is static.
"""
import torch
import torch.distributed as dist


def shard_range(n_items, rank, world):
    """Contiguous [lo, hi) of `n_items` owned by `rank`; earlier ranks take the remainder."""
    base, extra = divmod(n_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def sharded_masked_l1(local_sum, local_n, group=None):
    """(sum over ranks of local_sum) / (sum over ranks of local_n), differentiable w.r.t. local_sum."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local_sum / torch.clamp(local_n, min=1.0)
    pair = torch.stack([local_sum.detach().float(), local_n.detach().float()])
    dist.all_reduce(pair, op=dist.ReduceOp.SUM, group=group)
    # value = global sum / max(global N, 1); the gradient flows through this rank's share only (no host sync)
    return (local_sum + (pair[0] - local_sum.detach())) / torch.clamp(pair[1], min=1.0)


class GlobalCount:
    """The global normaliser N = ROI.sum() over all ranks, fetched AHEAD of the step: N depends on the labels alone, so
    its all-reduce can run on a side stream while the trees and filters of the step are computed, and nothing is left
    between the forward and the backward pass.  `start(roi)` launches it, `get()` joins and returns the [1] tensor."""

    def __init__(self, group=None):
        self.group, self.side, self.value = group, None, None

    def start(self, roi):
        main = torch.cuda.current_stream()
        if self.side is None:
            self.side = torch.cuda.Stream()
        self.side.wait_stream(main)
        with torch.cuda.stream(self.side):
            self.value = roi.sum(dtype=torch.float32).reshape(1)
            if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
                dist.all_reduce(self.value, op=dist.ReduceOp.SUM, group=self.group)
        return self

    def get(self):
        torch.cuda.current_stream().wait_stream(self.side)
        return self.value
