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


def test_adaptive_set_weights_graph_mode_on_a_capture_safe_network():
    """adaptive_set_weights with use_graph=True against its own eager mode on a small network whose forward can be
    captured (the reference's UNet_LC_MultiHead cannot: unet.py:182 copies a host tensor per call)."""
    import copy
    import types
    from fedicra_b200.adaptive import adaptive_set_weights

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.down = torch.nn.Conv2d(1, 8, 3, padding=1)
            self.up1 = torch.nn.Conv2d(8, 8, 3, padding=1)
            self.out_conv = torch.nn.Conv2d(8, 2, 1)

        def forward(self, x):
            return (self.out_conv(torch.relu(self.up1(torch.relu(self.down(x))))),)

    torch.manual_seed(0)
    g = torch.Generator().manual_seed(1)
    loader = [{"image": torch.rand(4, 32, 32, generator=g), "label": torch.randint(0, 3, (4, 32, 32), generator=g, dtype=torch.uint8)}
              for _ in range(3)]
    args = types.SimpleNamespace(img_class="faz", num_classes=2, cid=0, strategy="FedICRA")
    base = types.SimpleNamespace(model=Net().cuda(), args=args, trainloader=loader, start_phase=False)
    weights = [(v.detach().cpu() + 0.1 * torch.randn(v.shape, generator=g)).numpy() for v in base.model.state_dict().values()]
    outs = []
    for use_graph in (False, True):
        w = types.SimpleNamespace(model=copy.deepcopy(base.model), args=args, trainloader=loader, start_phase=False)
        adaptive_set_weights(w, weights, {"iter_global": 51}, use_graph=use_graph, log=lambda *a: None)
        torch.cuda.synchronize()
        outs.append(([p.detach().clone() for p in w.model.parameters()], [x.clone() for x in w.fedaa_weights]))
    assert len(outs[0][1]) == 4                                    # up1 and out_conv, weight + bias
    # graph mode sees the first batch three extra times during its warm-up: same direction, not the same bits
    for a, b in zip(outs[0][1], outs[1][1]):
        assert float((a - b).abs().max()) < 0.05 and float(b.min()) >= 0.0 and float(b.max()) <= 1.0
    for a, b in zip(outs[0][0], outs[1][0]):
        assert float((a - b).abs().max()) < 0.05 * float(a.abs().max())
