"""Fused element-wise ends of the tree-energy loss (csrc/glue.cu) and the multi-embedding MST launch, against PyTorch
expressions of the reference lines they replace (flower_common.py:777-781, 813-818)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-5


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("B,C,H,W", [(12, 2, 256, 256), (3, 3, 37, 41), (1, 1, 5, 7), (2, 16, 16, 16)])
def test_channel_softmax_forward_backward(B, C, H, W):
    from fedicra_b200.tree_energy import _ChannelSoftmax
    g = torch.Generator().manual_seed(1)
    x = (torch.randn(B, C, H, W, generator=g) * 3).cuda()
    gy = torch.randn(B, C, H, W, generator=g).cuda()
    x1, x2 = x.clone().requires_grad_(), x.clone().requires_grad_()
    y1 = _ChannelSoftmax.apply(x1)
    y2 = torch.softmax(x2, dim=1)
    assert rel(y1, y2) < TOL
    (y1 * gy).sum().backward()
    (y2 * gy).sum().backward()
    assert rel(x1.grad, x2.grad) < TOL


@pytest.mark.parametrize("B,C,H,W,frac", [(12, 2, 256, 256, 0.97), (2, 3, 33, 65, 0.5), (1, 2, 8, 8, 0.0)])
def test_masked_l1_matches_reference_expression(B, C, H, W, frac):
    from fedicra_b200.tree_energy import _masked_l1
    g = torch.Generator().manual_seed(2)
    prob = torch.rand(B, C, H, W, generator=g).cuda()
    filt = torch.rand(B, C, H, W, generator=g).cuda()
    filt[:, :, 0, :] = prob[:, :, 0, :]                       # exact ties: sign(0) = 0
    roi_b = (torch.rand(B, H, W, generator=g) < frac).cuda()
    weight = 0.1
    p1, f1 = prob.clone().requires_grad_(), filt.clone().requires_grad_()
    p2, f2 = prob.clone().requires_grad_(), filt.clone().requires_grad_()
    for _ in range(2):                                         # the scratch counter must be left reusable
        got = _masked_l1(None, p1, f1, None, False, weight, roi_b.view(torch.uint8))
    roi = roi_b.unsqueeze(1).float()                           # flower_common.py:776-777, 813-818
    n = roi.sum()
    want = (roi * torch.abs(p2 - f2)).sum()
    want = weight * (want / n if float(n) > 0 else want)
    assert abs(float(got) - float(want)) <= 1e-5 * max(1.0, abs(float(want)))
    p1.grad = f1.grad = None
    got.backward()
    want.backward()
    if float(n) > 0:
        assert rel(p1.grad, p2.grad) < TOL and rel(f1.grad, f2.grad) < TOL
    else:
        assert float(p1.grad.abs().max()) == 0.0


def test_masked_l1_sums_output_feeds_the_sharded_path():
    from fedicra_b200.tree_energy import _MaskedL1
    g = torch.Generator().manual_seed(3)
    prob = torch.rand(2, 2, 16, 16, generator=g).cuda().requires_grad_()
    filt = torch.rand(2, 2, 16, 16, generator=g).cuda().requires_grad_()
    roi_b = (torch.rand(2, 16, 16, generator=g) < 0.5).cuda()
    _, sums = _MaskedL1.apply(prob, filt, roi_b.view(torch.uint8), 1.0)
    (3.0 * sums[0]).backward()
    roi = roi_b.unsqueeze(1).float()
    assert abs(float(sums[1]) - float(roi.sum())) < 0.5
    want = 3.0 * roi * torch.sign(prob.detach() - filt.detach())
    assert rel(prob.grad, want) < TOL and rel(filt.grad, -want) < TOL


def test_mst_grid_multi_equals_separate_launches():
    from fedicra_b200 import ops
    g = torch.Generator().manual_seed(4)
    B, H, W = 3, 40, 56
    embs = [((torch.rand(B, 3, H, W, generator=g) * 255).round() / 255).cuda(), torch.randn(B, 2, H, W, generator=g).cuda(),
            torch.randn(B, 1, H, W, generator=g).cuda()]
    edges, mask = ops.mst_grid_multi(embs, want_mask=True, want_edges=True)
    for k, e in enumerate(embs):
        e1, m1 = ops.mst_grid(e, want_mask=True)
        assert torch.equal(edges[k * B:(k + 1) * B], e1) and torch.equal(mask[k * B:(k + 1) * B], m1)


@pytest.mark.parametrize("B,C,sizes,H,W", [(2, 2, [(64, 64), (128, 128)], 256, 256), (12, 3, [(64, 64), (128, 128)], 256, 256),
                                           (1, 2, [(16, 20)], 48, 64), (2, 3, [(13, 17), (30, 7)], 40, 52),
                                           (1, 1, [(128, 128), (256, 256)], 512, 512)])
def test_bilinear_resize_has_torchs_bits_and_its_gradient(B, C, sizes, H, W):
    """fedicra_bilinear_resize_fwd == F.interpolate(mode='bilinear', align_corners=False) BIT FOR BIT (the resized heads
    are the embeddings the trees are built on); the gather backward == autograd of F.interpolate within rounding."""
    import torch.nn.functional as F
    from fedicra_b200 import ops
    g = torch.Generator().manual_seed(5)
    heads = [torch.randn(B, C, h, w, generator=g).cuda() for h, w in sizes]
    outs = ops.bilinear_resize(heads, H, W)
    gouts = [torch.randn(B, C, H, W, generator=g).cuda() for _ in sizes]
    back = ops.bilinear_resize([(go, hw) for go, hw in zip(gouts, sizes)], H, W, backward=True)
    for x, o, go, gi in zip(heads, outs, gouts, back):
        xr = x.clone().requires_grad_()
        want = F.interpolate(xr, size=(H, W), mode="bilinear", align_corners=False)
        assert torch.equal(o, want.detach()), "resized head differs from torch's bits"
        want.backward(go)
        assert rel(gi, xr.grad) < 1e-5


def test_edge_weight_backward_of_several_embeddings_in_one_launch():
    """fedicra_edge_weight_bwd_multi == fedicra_edge_weight_bwd per embedding, bit for bit (different channel counts,
    different trees, different 1/sigma), and the joint autograd node of the loss modules hands out exactly the weights
    the ordering pass computed."""
    from fedicra_b200 import ops
    from fedicra_b200.tree_filter import TreeFilter2D, batched_tree_orders, edge_weights_multi
    g = torch.Generator().manual_seed(5)
    B, H, W = 3, 40, 56
    V = H * W
    embeds = [torch.randn(B, D, H, W, generator=g).cuda().requires_grad_() for D in (2, 3, 1)]
    inv = [1.0, 0.5, 50.0]
    orders = batched_tree_orders(embeds, inv)
    arrays = [o.for_channels(2) for o in orders]
    assert all(a.kind == "chain" for a in arrays)
    ews = [a.ew.clone() for a in arrays]
    gews = [torch.randn(B, V, generator=g).cuda() for _ in embeds]
    flat = [e.detach().reshape(B, e.shape[1], V) for e in embeds]
    single = [ops.edge_weight_bwd(f, a.si, a.sp, a.sc, w, gw, s) for f, a, w, gw, s in zip(flat, arrays, ews, gews, inv)]
    multi = ops.edge_weight_bwd_multi(flat, [(a.si, a.sp, a.sc) for a in arrays], ews, gews, inv)
    for x, y in zip(single, multi):
        assert torch.equal(x, y)
    edge_weights_multi(arrays, embeds, inv)
    tf = TreeFilter2D()
    outs = []
    for a, e, s, w in zip(arrays, embeds, inv, ews):
        assert a.ew is None and a.ew_out is not None
        tf.sigma = 1.0 / s
        out = tf.build_edge_weight(e, a, low_tree=True)       # low_tree: 1/sigma as given
        assert a.ew_out is None and torch.equal(out.detach(), w)
        outs.append(out)
    sum((o * gw).sum() for o, gw in zip(outs, gews)).backward()
    for e, want in zip(embeds, single):
        assert torch.equal(e.grad.reshape(B, -1, V), want)
