"""The travelling oracle against the committed fixtures that the unmodified reference produced
(tests/golden/make_golden.py).  This is what keeps the oracle honest on the GPU box, where /root/reference is absent."""
import os

import numpy as np
import pytest
import torch

from oracle import gcrf_cpu, losses_cpu, tree_cpu
from tests.helpers import edge_set, rel_err


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


@pytest.mark.parametrize("name", ["gcrf_faz.npz", "gcrf_odoc.npz"])
def test_gcrf_oracle_vs_golden(golden_dir, name):
    z = _load(golden_dir, name)
    y, img = torch.from_numpy(z["y"]).requires_grad_(), torch.from_numpy(z["image"])
    kd = [{"weight": 1, "xy": float(z["sigma_xy"]), "rgb": float(z["sigma_rgb"])}]
    for fn in (gcrf_cpu.gated_crf_loss_unfold, gcrf_cpu.gated_crf_loss):
        y.grad = None
        loss = fn(y, kd, int(z["radius"]), img)
        loss.backward()
        assert rel_err(loss.detach(), z["loss"]) < 1e-5
        assert rel_err(y.grad, z["grad_y"]) < 1e-5


def test_gcrf_oracle_two_descriptors(golden_dir):
    z = _load(golden_dir, "gcrf_two_desc.npz")
    y, img = torch.from_numpy(z["y"]), torch.from_numpy(z["image"])
    kd = [{"weight": 0.9, "xy": 6, "rgb": 0.1}, {"weight": 0.1, "xy": 6}]
    loss, grad = gcrf_cpu.gated_crf_loss_and_grad(y, kd, int(z["radius"]), img)
    assert rel_err(loss, z["loss"]) < 1e-5 and rel_err(grad, z["grad_y"]) < 1e-5


@pytest.mark.parametrize("name", ["mst_real_faz.npz", "mst_quant8.npz", "mst_constant.npz", "mst_uniform.npz"])
def test_mst_oracle_vs_reference_boruvka_golden(golden_dir, name):
    z = _load(golden_dir, name)
    emb = torch.from_numpy(z["embed"])
    B, _, H, W = emb.shape
    wgt = losses_cpu.grid_weights(emb)
    assert np.array_equal(wgt.numpy(), z["weight"])  # torch reproduces the fixture's fp32 weight bits
    tree = losses_cpu.minimum_spanning_tree(emb)
    assert np.array_equal(tree.numpy(), z["tree"])   # same edges in the same order as boruvka.cpp


@pytest.mark.parametrize("name", ["tel_faz.npz", "tel_odoc.npz"])
def test_tree_energy_oracle_vs_golden(golden_dir, name):
    z = _load(golden_dir, name)
    t = {k: torch.from_numpy(z[k]) for k in z.files}
    xs = [t[k].clone().requires_grad_() for k in ("logits", "aux1", "aux2", "aux3")]
    loss, as1, as2, as3 = losses_cpu.mscale_recurve_tree_energy_loss(xs[0], t["image"], xs[1], xs[2], xs[3], t["roi"], 0.1)
    loss.backward()
    assert rel_err(loss.detach(), t["loss"]) < 1e-6
    for got, key in ((as1, "as1"), (as2, "as2"), (as3, "as3")):
        assert rel_err(got.detach(), t[key]) < 1e-6
    for x, key in zip(xs, ("g_logits", "g_aux1", "g_aux2", "g_aux3")):
        assert rel_err(x.grad, t[key]) < 1e-5
