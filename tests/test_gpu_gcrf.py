"""Gated CRF kernel (through the C ABI) against the golden fixtures and the CPU oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import gcrf_cpu
from tests.helpers import rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-4  # north_star: losses and gradients within 1e-4 relative error in fp32
KD = [{"weight": 1, "xy": 6, "rgb": 0.1}]


def run_module(y, img, kd=KD, radius=5):
    from fedicra_b200 import ModelLossSemsegGatedCRF
    yc = y.cuda().requires_grad_()
    out = ModelLossSemsegGatedCRF()(yc, kd, radius, img.cuda(), y.shape[2], y.shape[3])["loss"]
    assert out.dim() == 0
    out.backward()
    return out.detach().cpu(), yc.grad.cpu()


@pytest.mark.parametrize("name", ["gcrf_faz.npz", "gcrf_odoc.npz"])
def test_golden(golden_dir, name):
    z = np.load(os.path.join(golden_dir, name))
    loss, grad = run_module(torch.from_numpy(z["y"]), torch.from_numpy(z["image"]))
    assert rel_err(loss, z["loss"]) < TOL and rel_err(grad, z["grad_y"]) < TOL


def test_golden_two_descriptors_generic_path(golden_dir):
    z = np.load(os.path.join(golden_dir, "gcrf_two_desc.npz"))
    kd = [{"weight": 0.9, "xy": 6, "rgb": 0.1}, {"weight": 0.1, "xy": 6}]
    loss, grad = run_module(torch.from_numpy(z["y"]), torch.from_numpy(z["image"]), kd, int(z["radius"]))
    assert rel_err(loss, z["loss"]) < TOL and rel_err(grad, z["grad_y"]) < TOL


@pytest.mark.parametrize("N,C,Cin,H,W", [
    (2, 2, 1, 64, 64), (1, 3, 3, 37, 45), (1, 4, 3, 16, 130), (3, 2, 3, 5, 7), (1, 3, 1, 1, 1), (1, 2, 1, 11, 300),
    (1, 5, 2, 20, 20),   # channel counts outside the specialised kernels -> generic path
])
def test_against_fp64_oracle(N, C, Cin, H, W):
    g = torch.Generator().manual_seed(N * 1000 + H * 10 + W)
    y = torch.softmax(torch.randn(N, C, H, W, generator=g), 1)
    img = torch.rand(N, Cin, H, W, generator=g)
    l64, g64 = gcrf_cpu.gated_crf_loss_and_grad(y, KD, 5, img)
    loss, grad = run_module(y, img)
    assert rel_err(loss, l64) < TOL and rel_err(grad, g64) < TOL


@pytest.mark.parametrize("radius,kd", [(1, KD), (3, [{"weight": 2.0, "xy": 3}]), (7, [{"weight": 0.5, "rgb": 0.2}])])
def test_other_radii_and_modalities(radius, kd):
    g = torch.Generator().manual_seed(radius)
    y = torch.softmax(torch.randn(2, 3, 19, 23, generator=g), 1)
    img = torch.rand(2, 3, 19, 23, generator=g)
    l64, g64 = gcrf_cpu.gated_crf_loss_and_grad(y, kd, radius, img)
    loss, grad = run_module(y, img, kd, radius)
    assert rel_err(loss, l64) < TOL and rel_err(grad, g64) < TOL


@pytest.mark.parametrize("N,C,Cin,H,W", [(12, 2, 1, 256, 256), (4, 3, 3, 256, 256)])
def test_baseline_config_sizes(N, C, Cin, H, W):
    """BASELINE.json configs 1-3 shapes against the CPU oracle (fp32 shifted formulation, seconds on CPU)."""
    g = torch.Generator().manual_seed(2022)
    y = torch.softmax(torch.randn(N, C, H, W, generator=g), 1)
    img = (torch.rand(N, Cin, H, W, generator=g) * 255).round() / 255
    l_ref, g_ref = gcrf_cpu.gated_crf_loss_and_grad(y, KD, 5, img, dtype=torch.float64)
    loss, grad = run_module(y, img)
    assert rel_err(loss, l_ref) < TOL and rel_err(grad, g_ref) < TOL


def test_properties_at_512():
    """Size-independent properties on the 512x512 sweep shape: repeatability (bit-exact), linearity of K*y in y, and the
    identity loss = sum K + 0.5 * <y, grad> * denom / ... evaluated from the kernel's own outputs."""
    from fedicra_b200 import ops
    g = torch.Generator().manual_seed(5)
    N, C, H, W = 2, 3, 512, 512
    y1 = torch.softmax(torch.randn(N, C, H, W, generator=g), 1).cuda()
    y2 = torch.softmax(torch.randn(N, C, H, W, generator=g), 1).cuda()
    img = torch.rand(N, 3, H, W, generator=g).cuda()
    d = [(1.0, 6.0, 0.1)]
    la, ga = ops.gcrf_fwd_bwd(y1, img, 5, d)
    lb, gb = ops.gcrf_fwd_bwd(y1, img, 5, d)
    assert torch.equal(la, lb) and torch.equal(ga, gb)
    _, g2 = ops.gcrf_fwd_bwd(y2, img, 5, d)
    _, g12 = ops.gcrf_fwd_bwd(0.25 * y1 + 0.75 * y2, img, 5, d)
    assert rel_err(g12.cpu(), (0.25 * ga + 0.75 * g2).cpu()) < 1e-5
    # loss(y) - loss(0) = 0.5 * <y, grad(y)>   (quadratic form), loss(0) = sum K / denom
    l0, _ = ops.gcrf_fwd_bwd(torch.zeros_like(y1), img, 5, d)
    quad = 0.5 * (y1.double() * ga.double()).sum()
    assert abs((la.double() - l0.double()).item() - quad.item()) < 1e-4 * abs(la.item())


def test_forward_only_and_no_grad():
    from fedicra_b200 import ModelLossSemsegGatedCRF
    y = torch.softmax(torch.randn(1, 2, 32, 32), 1).cuda()
    img = torch.rand(1, 1, 32, 32).cuda()
    with torch.no_grad():
        a = ModelLossSemsegGatedCRF()(y, KD, 5, img, 32, 32)["loss"]
    b, _ = run_module(y.cpu(), img.cpu())
    assert abs(a.item() - b.item()) < 1e-6 * abs(b.item())
