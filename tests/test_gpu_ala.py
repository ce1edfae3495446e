"""Adaptive-aggregation kernels: bit-exact against the numpy oracle (pure element-wise, no fused multiply-add)."""
import numpy as np
import pytest
import torch

from oracle import ala_cpu

pytestmark = pytest.mark.gpu


def _tensors(sizes, seed):
    g = torch.Generator().manual_seed(seed)
    mk = lambda: [torch.randn(n, generator=g) for n in sizes]
    return mk(), mk(), mk(), [torch.rand(n, generator=g) for n in sizes]


def test_multi_tensor_update_bit_exact_with_missing_gradients():
    from fedicra_b200 import fused_ala_update
    sizes = [476032 // 8, 290, 1, 7, 4096, 4097, 12345]   # decoder-like mix incl. sizes not divisible by 4
    p, s, g, w = _tensors(sizes, 0)
    cp, cs, cw = ([x.cuda() for x in grp] for grp in (p, s, w))
    ct = [torch.nn.Parameter(torch.zeros(n, device="cuda")) for n in sizes]
    fused_ala_update(cp, cs, ct, cw, eta=1.0, use_grad=False)        # flower_common.py:553-555
    for k in range(len(sizes)):
        assert np.array_equal(ct[k].detach().cpu().numpy(), ala_cpu.ala_blend(p[k].numpy(), s[k].numpy(), w[k].numpy()))
    for k in range(len(sizes)):
        ct[k].grad = None if k == 2 else g[k].cuda()                 # tensor 2 has no gradient (:593-594)
    fused_ala_update(cp, cs, ct, cw, eta=1.0)
    for k in range(len(sizes)):
        if k == 2:
            w_ref, t_ref = w[k].numpy(), ala_cpu.ala_blend(p[k].numpy(), s[k].numpy(), w[k].numpy())
        else:
            w_ref, t_ref = ala_cpu.ala_update(p[k].numpy(), s[k].numpy(), g[k].numpy(), w[k].numpy(), 1.0)
        assert np.array_equal(cw[k].cpu().numpy(), w_ref), k
        assert np.array_equal(ct[k].detach().cpu().numpy(), t_ref), k


@pytest.mark.parametrize("n", [1, 3, 633010, 633155])
def test_flat_update_bit_exact_at_decoder_sizes(n):
    """633 010 / 633 155 = the 42 decoder tensors of UNet_LC_MultiHead, FAZ / ODOC."""
    from fedicra_b200 import AdaptiveAggregation
    (p,), (s,), (g,), _ = _tensors([n], n)
    agg = AdaptiveAggregation(p.cuda(), s.cuda(), eta=1.0)
    assert np.array_equal(agg.result().cpu().numpy(), ala_cpu.ala_blend(p.numpy(), s.numpy(), np.ones(n, np.float32)))
    w_ref = np.ones(n, np.float32)
    for it in range(3):
        gi = g * (it + 1)
        agg.step(gi.cuda())
        w_ref, t_ref = ala_cpu.ala_update(p.numpy(), s.numpy(), gi.numpy(), w_ref, 1.0)
    assert np.array_equal(agg.weights.cpu().numpy(), w_ref) and np.array_equal(agg.result().cpu().numpy(), t_ref)
    assert float(agg.weights.min()) >= 0 and float(agg.weights.max()) <= 1


def test_scale_inplace():
    from fedicra_b200 import ops
    x = torch.randn(2343287)
    y = x.cuda()
    ops.scale_inplace(y, 21.0 / 113.0)
    assert np.array_equal(y.cpu().numpy(), (x * np.float32(21.0 / 113.0)).numpy())
