"""CPU restatement of the Gated CRF loss -- TEST INFRASTRUCTURE ONLY.

Follows /root/reference/code/utils/gate_crf_loss.py:
  forward (live branch: no masks, Potts compatibility)        :20-124
  _create_kernels / _create_kernels_from_features             :134-172
  _get_mesh (absolute pixel coordinates x then y)             :174-181
  _unfold (zero padding of the FEATURES, so out-of-image taps have f_q = 0 and still add exp(-|f_p|^2/2)
           to kernels.sum())                                   :183-188

Two independent formulations:
  * ``gated_crf_loss_unfold``  -- the reference's algorithm (materialise the [N,*,d,d,H,W] window tensors); this is
    what the bench times as the CPU baseline because it has the reference's memory behaviour;
  * ``gated_crf_loss``         -- one shifted view per tap, O(N*H*W) memory, any dtype (fp64 for gradient checks).
Both are pinned against the unmodified reference module in tests/test_oracle_ref.py and tests/golden/.
"""
import torch
import torch.nn.functional as F


def _feature_stack(desc, sample, N, H, W, dtype):
    """Per-descriptor feature tensor [N,D,H,W]: dict order, each modality divided by its sigma (:142-155)."""
    feats = []
    for name, sigma in desc.items():
        if name == "weight":
            continue
        if name == "xy":
            xs = torch.arange(W, dtype=torch.float32).view(1, 1, 1, W).expand(N, 1, H, W)
            ys = torch.arange(H, dtype=torch.float32).view(1, 1, H, 1).expand(N, 1, H, W)
            f = torch.cat([xs, ys], 1)
        else:
            f = F.adaptive_avg_pool2d(sample.float(), (H, W))
        feats.append((f / sigma).to(dtype))
    return torch.cat(feats, 1)


def gated_crf_loss(y, kernels_desc, radius, sample, dtype=None):
    """loss = [sum K - sum_c y_p (K*y)_p] / (N*H*W); differentiable w.r.t. ``y`` through autograd."""
    N, C, H, W = y.shape
    dtype = dtype or y.dtype
    yy = y.to(dtype)
    ypad = F.pad(yy, [radius] * 4)
    ksum = torch.zeros((), dtype=dtype)
    prod = torch.zeros_like(yy)
    for desc in kernels_desc:
        f = _feature_stack(desc, sample, N, H, W, dtype)
        fpad = F.pad(f, [radius] * 4)
        for dy in range(-radius, radius + 1):
            for dx in range(-radius, radius + 1):
                if dy == 0 and dx == 0:
                    continue
                fq = fpad[:, :, radius + dy:radius + dy + H, radius + dx:radius + dx + W]
                k = desc["weight"] * torch.exp(-0.5 * ((fq - f) ** 2).sum(1, keepdim=True))
                ksum = ksum + k.sum()
                prod = prod + k * ypad[:, :, radius + dy:radius + dy + H, radius + dx:radius + dx + W]
    return (ksum - (prod * yy).sum()) / (N * H * W)


def gated_crf_loss_and_grad(y, kernels_desc, radius, sample, dtype=torch.float64):
    """Loss and d loss / d y without autograd: grad = -2 (K*y) / (N*H*W) (K is symmetric)."""
    N, C, H, W = y.shape
    yy = y.detach().to(dtype)
    ypad = F.pad(yy, [radius] * 4)
    ksum = torch.zeros((), dtype=dtype)
    prod = torch.zeros_like(yy)
    for desc in kernels_desc:
        f = _feature_stack(desc, sample, N, H, W, dtype)
        fpad = F.pad(f, [radius] * 4)
        for dy in range(-radius, radius + 1):
            for dx in range(-radius, radius + 1):
                if dy == 0 and dx == 0:
                    continue
                fq = fpad[:, :, radius + dy:radius + dy + H, radius + dx:radius + dx + W]
                k = desc["weight"] * torch.exp(-0.5 * ((fq - f) ** 2).sum(1, keepdim=True))
                ksum += k.sum()
                prod += k * ypad[:, :, radius + dy:radius + dy + H, radius + dx:radius + dx + W]
    denom = N * H * W
    return (ksum - (prod * yy).sum()) / denom, -2.0 * prod / denom


def _windows(t, radius):
    N, C, H, W = t.shape
    d = 2 * radius + 1
    return F.unfold(t, d, 1, radius).view(N, C, d, d, H, W)


def gated_crf_loss_unfold(y, kernels_desc, radius, sample):
    """Same value, computed the way the reference does it (window tensors); fp32; autograd for the gradient."""
    N, C, H, W = y.shape
    kern = None
    for desc in kernels_desc:
        f = _feature_stack(desc, sample, N, H, W, torch.float32)
        win = _windows(f, radius)
        centre = win[:, :, radius, radius].view(N, -1, 1, 1, H, W)
        k = torch.exp((-0.5 * (win - centre) ** 2).sum(1, keepdim=True))
        k[:, :, radius, radius] = 0
        kern = desc["weight"] * k if kern is None else kern + desc["weight"] * k
    agg = (kern * _windows(y, radius)).view(N, C, -1, H, W).sum(2)
    return (kern.sum() - (agg * y).sum()) / (N * H * W)
