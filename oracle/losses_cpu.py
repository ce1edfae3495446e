"""CPU restatement of the tree-energy loss stack on top of oracle.tree_cpu -- TEST INFRASTRUCTURE ONLY.

Follows the reference (paths relative to /root/reference/code):
  grid edge list + weights  utils/TreeEnergyLoss/kernels/lib_tree_filter/modules/tree_filter.py:15-34, 53-62
  tree filter               .../modules/tree_filter.py:84-108, 133-151 and functions/refine.py:9-40
  multiscale loss           flower_common.py:756-818 (MScaleRecurveTreeEnergyLoss)
  full loss of one step     flower_pCE_2D_GateCRFMsacleTreeEnergyLoss_Ours.py:132-151

It travels to the GPU box (the reference's Python does not), where it is the parity checker for the `-m gpu` tests and
the timed CPU baseline of bench.py.  Pinned against the unmodified reference modules in tests/test_oracle_ref.py.
"""
import torch
import torch.nn.functional as F

from . import gcrf_cpu, tree_cpu


def grid_edges(H, W):
    """[E,2] int32: vertical edges (h,w)-(h+1,w) row-major first, then horizontal (h,w)-(h,w+1)  (tree_filter.py:15-25)."""
    ids = torch.arange(H * W, dtype=torch.int32).view(H, W)
    vert = torch.stack([ids[:-1, :], ids[1:, :]], -1).reshape(-1, 2)
    horz = torch.stack([ids[:, :-1], ids[:, 1:]], -1).reshape(-1, 2)
    return torch.cat([vert, horz], 0)


def grid_weights(embed):
    """[B,E] fp32: squared channel distance of the two end points, plus one  (tree_filter.py:27-34, 78-81)."""
    B = embed.shape[0]
    dv = embed[:, :, :-1, :] - embed[:, :, 1:, :]
    dh = embed[:, :, :, :-1] - embed[:, :, :, 1:]
    wv = (dv * dv).sum(1).reshape(B, -1)
    wh = (dh * dh).sum(1).reshape(B, -1)
    return torch.cat([wv, wh], 1) + 1


def minimum_spanning_tree(embed):
    """[B,D,H,W] -> [B,V-1,2] int32 tree edges.  Always on the fp32 embedding: tree parity is defined on fp32 weight bits,
    also when the rest of the stack runs in fp64 as the high-precision judge."""
    B, _, H, W = embed.shape
    with torch.no_grad():
        idx = grid_edges(H, W).unsqueeze(0).expand(B, -1, -1).contiguous()
        return tree_cpu.mst_forward(idx, grid_weights(embed.detach().float()).contiguous(), H * W)


class _Refine(torch.autograd.Function):
    """refine fwd / bwd-feature / bwd-weight, weight gradient only for the high-level trees (refine.py:23-40)."""

    @staticmethod
    def forward(ctx, feature, edge_weight, si, sp, sc, low_tree):
        ew = edge_weight.detach().clone()
        out, aggr, aggr_up, ws, ws_up = tree_cpu.refine_forward(feature.detach(), ew, si, sp, sc)
        ctx.save_for_backward(feature, ew, si, sp, sc, out, aggr, aggr_up, ws, ws_up)
        ctx.low_tree = low_tree
        return out

    @staticmethod
    def backward(ctx, grad_out):
        saved = ctx.saved_tensors
        g = grad_out.contiguous()
        gf = tree_cpu.refine_backward_feature(*saved, g)
        gw = None if ctx.low_tree else tree_cpu.refine_backward_weight(*saved, g)
        return gf, gw, None, None, None, None


def tree_filter(feature, embed, tree, low_tree=True, sigma=0.02):
    """TreeFilter2D.forward with groups=1  (tree_filter.py:133-151)."""
    B, C, H, W = feature.shape
    si, sp, sc = tree_cpu.bfs_forward(tree, 4)
    e = embed.reshape(B, embed.shape[1], -1)
    src = torch.gather(e, 2, si.long().unsqueeze(1).expand(-1, e.shape[1], -1))
    dst = torch.gather(src, 2, sp.long().unsqueeze(1).expand(-1, e.shape[1], -1))
    dist = ((src - dst) ** 2).sum(1)
    ew = torch.exp(-dist / sigma) if low_tree else torch.exp(-dist)
    out = _Refine.apply(feature.reshape(B, C, -1).contiguous(), ew, si, sp, sc, low_tree)
    return out.reshape(B, C, H, W)


def mscale_recurve_tree_energy_loss(preds, low_feats, high_1, high_2, high_3, unlabeled_rois, weight, sigma=0.02,
                                    dtype=torch.float32):
    """flower_common.py:768-818.  Returns (weight * loss, AS_1, AS_2, AS_3).

    dtype=torch.float64 evaluates everything after the (fp32) resize and the (fp32-weight) MST in double precision."""
    h, w = preds.shape[2:]
    with torch.no_grad():
        low = F.interpolate(low_feats, size=(h, w), mode="bilinear", align_corners=False)
        roi = F.interpolate(unlabeled_rois.unsqueeze(1).float(), size=(h, w), mode="nearest").to(dtype)
        n = roi.sum()
    prob = torch.softmax(preds.to(dtype), 1)
    cur = tree_filter(prob, low.to(dtype), minimum_spanning_tree(low), True, sigma)
    outs = []
    for hf in (high_1, high_2, high_3):
        hf = F.interpolate(hf, size=(h, w), mode="bilinear", align_corners=False)
        cur = tree_filter(cur, hf.to(dtype), minimum_spanning_tree(hf), False, sigma)
        outs.append(cur)
    loss = (roi * torch.abs(prob - cur)).sum()
    if n > 0:
        loss = loss / n
    return (weight * loss, *outs)


KERNELS_DESC = [{"weight": 1, "xy": 6, "rgb": 0.1}]  # OURS.py:68
RADIUS = 5  # OURS.py:69


def loss_stack(logits, image, aux1, aux2, aux3, label, num_classes, tree_weight=0.1, crf_weight=0.1,
               with_ce=False, gcrf="unfold", dtype=torch.float32):
    """One step's weak-supervision loss (OURS.py:132-151): w_tree*TEL + 0.1*GatedCRF (+ pCE when asked).

    image: [B,Cin,H,W]; the tree embedding is the image repeated to 3 channels when Cin == 1 (OURS.py:137-141), the
    CRF modality is the image itself (OURS.py:147).  label == num_classes marks unlabelled pixels.
    """
    rois = label == num_classes
    three = image.repeat(1, 3, 1, 1) if image.shape[1] == 1 else image
    tel, _, _, as3 = mscale_recurve_tree_energy_loss(logits, three, aux1, aux2, aux3, rois, tree_weight, dtype=dtype)
    soft = torch.softmax(logits.to(dtype), 1)
    if gcrf == "unfold":
        crf = gcrf_cpu.gated_crf_loss_unfold(soft, KERNELS_DESC, RADIUS, image)
    else:
        crf = gcrf_cpu.gated_crf_loss(soft, KERNELS_DESC, RADIUS, image, dtype=dtype)
    total = tel + crf_weight * crf
    if with_ce:
        total = total + F.cross_entropy(logits, label.long(), ignore_index=num_classes)
    return total, tel, crf, as3
