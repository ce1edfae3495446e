"""Module-level parity: the reference-signature classes on CUDA against the oracle loss stack and the golden fixtures."""
import os

import numpy as np
import pytest
import torch

from oracle import losses_cpu
from tests.helpers import loss_inputs, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-4


def run_gpu(logits, image3, a1, a2, a3, roi, weight=0.1):
    from fedicra_b200 import MScaleRecurveTreeEnergyLoss
    xs = [t.cuda().requires_grad_() for t in (logits, a1, a2, a3)]
    out = MScaleRecurveTreeEnergyLoss()(xs[0], image3.cuda(), xs[1], xs[2], xs[3], roi.cuda(), weight)
    out[0].backward()
    return out[0].detach().cpu(), [o.detach().cpu() for o in out[1:]], [x.grad.cpu() for x in xs]


@pytest.mark.parametrize("name", ["tel_faz.npz", "tel_odoc.npz"])
def test_tree_energy_golden(golden_dir, name):
    z = np.load(os.path.join(golden_dir, name))
    t = {k: torch.from_numpy(z[k]) for k in z.files}
    loss, outs, grads = run_gpu(t["logits"], t["image"], t["aux1"], t["aux2"], t["aux3"], t["roi"])
    assert rel_err(loss, t["loss"]) < TOL
    for o, k in zip(outs, ("as1", "as2", "as3")):
        assert rel_err(o, t[k]) < TOL, k
    for gr, k in zip(grads, ("g_logits", "g_aux1", "g_aux2", "g_aux3")):
        assert rel_err(gr, t[k]) < TOL, k


@pytest.mark.parametrize("B,C,Cin,H,W,kind", [(2, 2, 1, 64, 64, "quant8"), (1, 3, 3, 48, 80, "uniform"),
                                              (12, 2, 1, 256, 256, "quant8")])
def test_full_loss_stack_vs_oracle(B, C, Cin, H, W, kind):
    """TEL + 0.1 * GatedCRF as in OURS.py:142-151, gradients w.r.t. logits and the three aux heads."""
    from fedicra_b200 import ModelLossSemsegGatedCRF, MScaleRecurveTreeEnergyLoss
    image, logits, a1, a2, a3, label = loss_inputs(B, C, Cin, H, W, kind)
    xs_ref = [t.clone().requires_grad_() for t in (logits, a1, a2, a3)]
    total_ref, tel_ref, crf_ref, as3_ref = losses_cpu.loss_stack(xs_ref[0], image, xs_ref[1], xs_ref[2], xs_ref[3],
                                                                  label, C, gcrf="shift")
    total_ref.backward()

    xs = [t.cuda().requires_grad_() for t in (logits, a1, a2, a3)]
    img = image.cuda()
    three = img.repeat(1, 3, 1, 1) if Cin == 1 else img
    roi = (label == C).cuda()
    tel, _, _, as3 = MScaleRecurveTreeEnergyLoss()(xs[0], three, xs[1], xs[2], xs[3], roi, 0.1)
    crf = ModelLossSemsegGatedCRF()(torch.softmax(xs[0], 1), losses_cpu.KERNELS_DESC, 5, img, H, W)["loss"]
    total = tel + 0.1 * crf
    total.backward()
    assert rel_err(tel.detach().cpu(), tel_ref.detach()) < TOL
    assert rel_err(crf.detach().cpu(), crf_ref.detach()) < TOL
    assert rel_err(as3.detach().cpu(), as3_ref.detach()) < TOL
    for a, b in zip(xs, xs_ref):
        assert rel_err(a.grad.cpu(), b.grad) < TOL


def test_reference_glue_runs_on_the_operator_shim():
    """The op-level drop-in: a TreeFilter2D written the reference's way (gather-based edge weights, refine autograd
    Function over the five tree_filter_cuda operators) gives the same result as the fused module."""
    from fedicra_b200 import MinimumSpanningTree, TreeFilter2D, tree_filter_cuda as tfc
    g = torch.Generator().manual_seed(0)
    B, C, H, W = 2, 2, 32, 40
    feat = torch.rand(B, C, H, W, generator=g).cuda().requires_grad_()
    emb = torch.rand(B, 3, H, W, generator=g).cuda().requires_grad_()
    tree = MinimumSpanningTree(TreeFilter2D.norm2_distance)(emb)
    out = TreeFilter2D()(feat, emb, tree, low_tree=False)
    out.sum().backward()
    g_feat, g_emb = feat.grad.clone(), emb.grad.clone()
    feat.grad = emb.grad = None

    class Refine(torch.autograd.Function):  # functions/refine.py of the reference, restated
        @staticmethod
        def forward(ctx, f, w, si, sp, sc):
            res = tfc.refine_forward(f, w, si, sp, sc)
            ctx.save_for_backward(f, w, si, sp, sc, *res)
            return res[0]

        @staticmethod
        def backward(ctx, go):
            s = ctx.saved_tensors
            return (tfc.refine_backward_feature(*s, go.contiguous()), tfc.refine_backward_weight(*s, go.contiguous()),
                    None, None, None)

    si, sp, sc = tfc.bfs_forward(tree, 4)
    e = emb.reshape(B, 3, -1)
    src = torch.gather(e, 2, si.long().unsqueeze(1).expand(-1, 3, -1))
    dst = torch.gather(src, 2, sp.long().unsqueeze(1).expand(-1, 3, -1))
    ew = torch.exp(-((src - dst) ** 2).sum(1))
    out2 = Refine.apply(feat.reshape(B, C, -1), ew, si, sp, sc).reshape(B, C, H, W)
    out2.sum().backward()
    assert rel_err(out2.detach().cpu(), out.detach().cpu()) < 1e-5
    assert rel_err(feat.grad.cpu(), g_feat.cpu()) < 1e-5
    assert rel_err(emb.grad.cpu(), g_emb.cpu()) < TOL
