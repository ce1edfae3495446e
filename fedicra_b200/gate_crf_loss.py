"""ModelLossSemsegGatedCRF with the reference's constructor and ``forward`` signature
(/root/reference/code/utils/gate_crf_loss.py:5-124), computed by one fused CUDA kernel (csrc/gcrf.cu) that produces the
loss and d loss / d y_hat_softmax together; no unfold tensor exists at any point.
"""
import torch
import torch.nn.functional as F

from . import ops


class _GatedCRF(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y, image, radius, descs):
        loss, grad = ops.gcrf_fwd_bwd(y, image, radius, descs, want_grad=ctx.needs_input_grad[0])
        ctx.save_for_backward(grad)
        return loss.reshape(())

    @staticmethod
    def backward(ctx, grad_loss):
        (grad,) = ctx.saved_tensors
        return grad * grad_loss, None, None, None


class ModelLossSemsegGatedCRF(torch.nn.Module):
    """Gated CRF loss (Obukhov et al. 2019), reference-compatible module.

    Supported: the branch FedICRA uses (gate_crf_loss.py:95-99) -- no source/destination masks, Potts compatibility.
    ``mask_src``, ``mask_dst``, ``compatibility`` and ``out_kernels_vis`` raise NotImplementedError (no silent fallback).
    As in the reference every non-'xy' modality of a kernel descriptor reads the one ``sample`` tensor (:148-152).
    """

    def forward(self, y_hat_softmax, kernels_desc, kernels_radius, sample, height_input, width_input,
                mask_src=None, mask_dst=None, compatibility=None, custom_modality_downsamplers=None,
                out_kernels_vis=False):
        assert y_hat_softmax.dim() == 4, 'Prediction must be a NCHW batch'
        N, C, height_pred, width_pred = y_hat_softmax.shape
        assert width_input % width_pred == 0 and height_input % height_pred == 0 and \
            width_input * height_pred == height_input * width_pred, \
            f'[{width_input}x{height_input}] !~= [{width_pred}x{height_pred}]'
        if mask_src is not None or mask_dst is not None:
            raise NotImplementedError("fedicra_b200 GatedCRF: source/destination masks are not implemented")
        if compatibility is not None:
            raise NotImplementedError("fedicra_b200 GatedCRF: only the Potts compatibility model is implemented")
        if out_kernels_vis:
            raise NotImplementedError("fedicra_b200 GatedCRF: kernel visualisation is not implemented")

        descs, image = [], None
        for desc in kernels_desc:
            inv_xy, inv_rgb = 0.0, 0.0
            for modality, sigma in desc.items():
                if modality == 'weight':
                    continue
                if modality == 'xy':
                    inv_xy += 1.0 / float(sigma) ** 2
                else:
                    inv_rgb += 1.0 / float(sigma) ** 2
                    if image is None:
                        image = self._downsample(sample, modality, height_pred, width_pred,
                                                 custom_modality_downsamplers)
            descs.append((float(desc['weight']), inv_xy ** -0.5 if inv_xy > 0 else 0.0,
                          inv_rgb ** -0.5 if inv_rgb > 0 else 0.0))
        if image is None:  # xy-only kernels: the image is not read, pass a 1-channel dummy
            image = torch.zeros(N, 1, height_pred, width_pred, dtype=torch.float32, device=y_hat_softmax.device)
        loss = _GatedCRF.apply(y_hat_softmax, image.detach().float(), int(kernels_radius), tuple(descs))
        return {'loss': loss}

    @staticmethod
    def _downsample(img, modality, height_dst, width_dst, custom_modality_downsamplers):
        """gate_crf_loss.py:126-132; at equal size (FedICRA's case) this is a no-op and is skipped."""
        if custom_modality_downsamplers is not None and modality in custom_modality_downsamplers:
            return custom_modality_downsamplers[modality](img, (height_dst, width_dst))
        if img.shape[2:] == (height_dst, width_dst):
            return img
        return F.adaptive_avg_pool2d(img, (height_dst, width_dst))
