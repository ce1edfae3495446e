"""Pin the oracle against the reference itself.  Runs only where /root/reference is mounted (this container)."""
import numpy as np
import pytest
import torch

from oracle import gcrf_cpu, losses_cpu, ref_harness, tree_cpu
from tests.helpers import edge_set, loss_inputs, rel_err

pytestmark = pytest.mark.skipif(not ref_harness.available(), reason="reference tree not mounted")


@pytest.mark.parametrize("kind,H,W", [("quant8", 24, 31), ("uniform", 16, 16), ("constant", 9, 7), ("quant8", 40, 40)])
def test_port_boruvka_equals_reference_boruvka(kind, H, W):
    """Same edges AND same output order as boruvka.cpp (compiled from /root/reference into oracle/_ref)."""
    g = torch.Generator().manual_seed(H + W)
    emb = torch.rand(3, 3, H, W, generator=g)
    if kind == "quant8":
        emb = (emb * 7).round() / 7
    if kind == "constant":
        emb = torch.zeros_like(emb)
    idx = losses_cpu.grid_edges(H, W).unsqueeze(0).expand(3, -1, -1).contiguous()
    wgt = losses_cpu.grid_weights(emb).contiguous()
    tree_cpu.set_mst_impl("reference")
    try:
        ref = tree_cpu.mst_forward(idx, wgt, H * W)
    finally:
        tree_cpu.set_mst_impl("port")
    port = tree_cpu.mst_forward(idx, wgt, H * W)
    assert torch.equal(ref, port)


@pytest.mark.parametrize("N,C,Cin,H,W", [(2, 2, 1, 16, 20), (1, 3, 3, 13, 24)])
def test_gcrf_ports_equal_reference_module(N, C, Cin, H, W):
    mod = ref_harness.gate_crf_loss().ModelLossSemsegGatedCRF()
    g = torch.Generator().manual_seed(1)
    y = torch.softmax(torch.randn(N, C, H, W, generator=g), 1).requires_grad_()
    img = torch.rand(N, Cin, H, W, generator=g)
    kd = [{"weight": 1, "xy": 6, "rgb": 0.1}]
    ref = mod(y, kd, 5, img, H, W)["loss"]
    ref.backward()
    gref = y.grad.clone()
    y.grad = None
    a = gcrf_cpu.gated_crf_loss_unfold(y, kd, 5, img)
    a.backward()
    assert rel_err(a.detach(), ref.detach()) < 1e-6 and rel_err(y.grad, gref) < 1e-6
    l64, g64 = gcrf_cpu.gated_crf_loss_and_grad(y, kd, 5, img)
    assert rel_err(l64, ref.detach()) < 1e-5 and rel_err(g64, gref) < 1e-5


def test_loss_stack_port_equals_reference_glue():
    """oracle.losses_cpu vs the unmodified flower_common.MScaleRecurveTreeEnergyLoss (same native stand-in below both)."""
    fc = ref_harness.flower_common()
    image, logits, a1, a2, a3, label = loss_inputs(2, 3, 3, 20, 16, "quant8", seed=4)
    roi = label == 3

    def run(fn):
        xs = [t.clone().requires_grad_() for t in (logits, a1, a2, a3)]
        out = fn(xs[0], image, xs[1], xs[2], xs[3], roi, 0.1)
        out[0].backward()
        return out[0].detach(), out[3].detach(), [x.grad for x in xs]

    l_ref, as_ref, g_ref = run(fc.MScaleRecurveTreeEnergyLoss())
    l_port, as_port, g_port = run(losses_cpu.mscale_recurve_tree_energy_loss)
    assert torch.equal(l_ref, l_port) and torch.equal(as_ref, as_port)
    for a, b in zip(g_ref, g_port):
        assert torch.equal(a, b)
