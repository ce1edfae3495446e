"""Generate tests/golden/*.npz by running the UNMODIFIED reference in this container.

    python tests/golden/make_golden.py

Needs /root/reference (absent on the GPU box -- which is why the vectors are committed).  What runs:
  * gcrf_*.npz   /root/reference/code/utils/gate_crf_loss.py  ModelLossSemsegGatedCRF.forward + autograd
  * mst_*.npz    the reference's own boruvka.cpp (compiled into oracle/_ref by `make -C oracle ref`) on the edge list and
                 weights that the reference's MinimumSpanningTree._build_matrix_index/_build_feature_weight produce
  * tel_*.npz    /root/reference/code/flower_common.py  MScaleRecurveTreeEnergyLoss forward + backward; its native
                 extension `tree_filter_cuda` cannot be built here (THC headers, needs a GPU), so it runs on
                 oracle.tree_cpu with mst_forward routed to the reference's boruvka.cpp.  Everything else (tree_filter.py,
                 functions/*.py, the loss glue, autograd) is the reference's code.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
OUT = os.path.dirname(os.path.abspath(__file__))

from oracle import ref_harness, tree_cpu  # noqa: E402


def real_faz_crop(size=48):
    from PIL import Image
    d = "/root/reference/data/FAZ/Domain1/train/imgs"
    name = sorted(os.listdir(d))[0]
    im = np.asarray(Image.open(os.path.join(d, name)).convert("L"), dtype=np.float32) / 255.0
    h0, w0 = (im.shape[0] - size) // 2, (im.shape[1] - size) // 2
    return im[h0:h0 + size, w0:w0 + size].copy(), name


def make_gcrf():
    mod = ref_harness.gate_crf_loss().ModelLossSemsegGatedCRF()
    kd = [{"weight": 1, "xy": 6, "rgb": 0.1}]
    for tag, (N, C, Cin, H, W) in {"faz": (2, 2, 1, 20, 24), "odoc": (1, 3, 3, 18, 37)}.items():
        g = torch.Generator().manual_seed(2022)
        y = torch.softmax(torch.randn(N, C, H, W, generator=g), 1).requires_grad_()
        img = torch.rand(N, Cin, H, W, generator=g)
        loss = mod(y, kd, 5, img, H, W)["loss"]
        loss.backward()
        np.savez_compressed(os.path.join(OUT, f"gcrf_{tag}.npz"), y=y.detach().numpy(), image=img.numpy(),
                            loss=loss.detach().numpy(), grad_y=y.grad.numpy(), radius=5, sigma_xy=6.0, sigma_rgb=0.1)
    # two-descriptor kernel (RGBXY + XY), radius 3: exercises the generic path
    g = torch.Generator().manual_seed(7)
    N, C, Cin, H, W = 1, 2, 3, 14, 15
    y = torch.softmax(torch.randn(N, C, H, W, generator=g), 1).requires_grad_()
    img = torch.rand(N, Cin, H, W, generator=g)
    kd2 = [{"weight": 0.9, "xy": 6, "rgb": 0.1}, {"weight": 0.1, "xy": 6}]
    loss = mod(y, kd2, 3, img, H, W)["loss"]
    loss.backward()
    np.savez_compressed(os.path.join(OUT, "gcrf_two_desc.npz"), y=y.detach().numpy(), image=img.numpy(),
                        loss=loss.detach().numpy(), grad_y=y.grad.numpy(), radius=3)


def make_mst():
    tf = ref_harness.tree_filter()
    tree_cpu.set_mst_impl("reference")
    mst = tf.MinimumSpanningTree(tf.TreeFilter2D.norm2_distance)
    crop, name = real_faz_crop(48)
    g = torch.Generator().manual_seed(2022)
    cases = {
        "real_faz": torch.from_numpy(crop)[None, None].repeat(1, 3, 1, 1),
        "quant8": (torch.rand(2, 3, 24, 31, generator=g) * 255).round() / 255,
        "constant": torch.zeros(1, 1, 9, 7),
        "uniform": torch.rand(1, 2, 17, 16, generator=g),
    }
    for tag, emb in cases.items():
        idx = mst._build_matrix_index(emb).contiguous()
        wgt = mst._build_feature_weight(emb).contiguous()
        tree = tree_cpu.mst_forward(idx, wgt, emb.shape[2] * emb.shape[3])
        np.savez_compressed(os.path.join(OUT, f"mst_{tag}.npz"), embed=emb.numpy(), weight=wgt.numpy(),
                            tree=tree.numpy())
    tree_cpu.set_mst_impl("port")
    return name


def make_tel():
    fc = ref_harness.flower_common()
    tree_cpu.set_mst_impl("reference")
    loss_mod = fc.MScaleRecurveTreeEnergyLoss()
    for tag, (B, C, H, W, quant) in {"faz": (2, 2, 16, 20, True), "odoc": (1, 3, 12, 12, False)}.items():
        g = torch.Generator().manual_seed(2022)
        logits = torch.randn(B, C, H, W, generator=g).requires_grad_()
        img = torch.rand(B, 3, H, W, generator=g)
        if quant:
            img = (img[:, :1] * 255).round().div(255).repeat(1, 3, 1, 1)
        a1 = torch.randn(B, C, H // 4, W // 4, generator=g).requires_grad_()
        a2 = torch.randn(B, C, H // 2, W // 2, generator=g).requires_grad_()
        a3 = torch.randn(B, C, H, W, generator=g).requires_grad_()
        roi = torch.rand(B, H, W, generator=g) > 0.03
        loss, as1, as2, as3 = loss_mod(logits, img, a1, a2, a3, roi, 0.1)
        loss.backward()
        np.savez_compressed(
            os.path.join(OUT, f"tel_{tag}.npz"), logits=logits.detach().numpy(), image=img.numpy(),
            aux1=a1.detach().numpy(), aux2=a2.detach().numpy(), aux3=a3.detach().numpy(), roi=roi.numpy(),
            loss=loss.detach().numpy(), as1=as1.detach().numpy(), as2=as2.detach().numpy(), as3=as3.detach().numpy(),
            g_logits=logits.grad.numpy(), g_aux1=a1.grad.numpy(), g_aux2=a2.grad.numpy(), g_aux3=a3.grad.numpy())
    tree_cpu.set_mst_impl("port")


if __name__ == "__main__":
    tree_cpu.build(ref=True)
    make_gcrf()
    name = make_mst()
    make_tel()
    print("golden fixtures written to", OUT, "(real crop from", name + ")")
