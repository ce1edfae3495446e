"""Flower-less driver for the reference's own training loop -- TEST INFRASTRUCTURE ONLY (tests/, bench.py reference legs).

Runs `MyClient._train` (flower_pCE_2D_GateCRFMsacleTreeEnergyLoss_Ours.py:52-206) and `MyModel.set_weights`
(flower_common.py:491-633) EXACTLY as the reference ships them (the copies under oracle/_ref/py), with
`UNet_LC_MultiHead` from the reference's net_factory, on synthetic batches in the data loader's format
(SURVEY.md Appendix C).  What is swapped is only what the drop-in is meant to swap:

  backend = "reference" : the reference's Python modules on the reference's own CUDA extension (rebuilt for sm_100a)
  backend = "ops"       : the reference's Python modules on fedicra_b200.tree_filter_cuda   (install_dropin("ops"))
  backend = "modules"   : fedicra_b200's GatedCRF / MinimumSpanningTree / TreeFilter2D      (install_dropin("modules"))

One back end per process (the reference binds its imports at import time):

    python -m oracle.train_harness --backend modules --img-class faz --iters 4 --json out.json
"""
import argparse
import json
import os
import sys
import time
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def make_args(img_class, num_clients=5, cid=0, iters=4, rep_iters=3, img_size=256):
    """Namespace with the fields `_train`, BaseClient, MyModel and net_factory read; values are
    the script's own defaults (OURS.py:215-270)."""
    return types.SimpleNamespace(
        alpha=0.5, base_lr=0.01, cid=cid, img_class=img_class, img_size=img_size, iters=iters, max_iterations=30000,
        min_num_clients=num_clients, model="unet_lc_multihead", num_classes=2 if img_class == "faz" else 3,
        rep_iters=rep_iters, strategy="FedICRA", tree_loss_weight=0.1, amp=0, snapshot_path="/tmp",
        in_chns=1 if img_class == "faz" else 3, batch_size=12, seed=2022)


def make_loader(args, n_batches, batch_size, seed):
    """List of batches shaped like dataset.py:176-183,246-251 yields them: image float32 [B,H,W] (faz) or [B,3,H,W],
    label uint8 [B,H,W] with `num_classes` = unlabelled and ~3 % scribble pixels."""
    import torch
    g = torch.Generator().manual_seed(seed)
    H = W = args.img_size
    C = args.num_classes
    out = []
    for _ in range(n_batches):
        shape = (batch_size, H, W) if args.img_class == "faz" else (batch_size, 3, H, W)
        image = (torch.rand(*shape, generator=g) * 255).round() / 255
        label = torch.full((batch_size, H, W), C, dtype=torch.uint8)
        scribble = torch.rand(batch_size, H, W, generator=g) < 0.03
        label[scribble] = torch.randint(0, C, (int(scribble.sum()),), generator=g, dtype=torch.uint8)
        out.append({"image": image, "label": label, "idx": 0})
    return out


def install(backend):
    """Returns (OURS script module, flower_common module) with the requested back end behind them."""
    from oracle import ref_gpu
    if backend == "modules":
        ref_gpu.install("fedicra")                 # stubs + sys.path; tree_filter_cuda -> ours as well
        import fedicra_b200
        fedicra_b200.install_dropin("modules")     # utils.gate_crf_loss and ...modules.tree_filter -> ours
    elif backend == "ops":
        ref_gpu.install("fedicra")
    elif backend == "reference":
        ref_gpu.install("reference")
    else:
        raise ValueError(backend)
    return ref_gpu.ours_script(), ref_gpu.flower_common()


def build_client(backend, img_class, iters, rep_iters, n_batches=2, batch_size=12, img_size=256, seed=2022, cid=0):
    import torch
    ours, fc = install(backend)
    args = make_args(img_class, iters=iters, rep_iters=rep_iters, img_size=img_size, cid=cid)
    loader = make_loader(args, n_batches, batch_size, seed + cid)
    torch.manual_seed(seed)
    torch.cuda.manual_seed_all(seed)
    from networks.net_factory import net_factory

    class MyModelV2(fc.MyModel):                   # OURS.py:337-341, verbatim
        def forward(self, x, emb_idx=None):
            return self.model(x, emb_idx)

    model = MyModelV2(args, net_factory(args, net_type=args.model, in_chns=args.in_chns, class_num=args.num_classes),
                      loader, loader)
    model.cuda()
    client = ours.MyClient(args, model, loader, loader, amp=False)
    return ours, fc, args, client


def run(backend, img_class, iters, rep_iters, n_batches=2, batch_size=12, img_size=256, warmup_iters=0):
    """-> dict with the per-iteration losses the reference logs, timing, and which modules served the loss classes."""
    import torch
    ours, fc, args, client = build_client(backend, img_class, iters, rep_iters, n_batches, batch_size, img_size)
    lines = []
    ours.log = lambda level, msg, *a: lines.append(str(msg))
    if warmup_iters:
        client._train({"iters": warmup_iters})
        lines.clear()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    loss, metrics = client._train({"iters": iters})
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    per_iter = []
    for ln in lines:
        if "loss :" in ln:
            f = ln.replace(",", " ").split()
            per_iter.append({"loss": float(f[f.index("loss") + 2]), "loss_ce": float(f[f.index("loss_ce:") + 1]),
                             "loss_tree": float(f[f.index("loss_tree:") + 1])})
    tel = fc.MScaleRecurveTreeEnergyLoss()
    return {
        "backend": backend, "img_class": img_class, "iters": iters, "rep_iters": rep_iters, "batch_size": batch_size,
        "img_size": img_size, "final_loss": float(loss), "per_iter": per_iter, "seconds": dt,
        "images_per_s": iters * batch_size / dt,
        "gated_crf_module": ours.ModelLossSemsegGatedCRF.__module__,
        "tree_filter_module": type(tel.tree_filter_layers).__module__,
        "flower_common_file": fc.__file__, "script_file": ours.__file__,
    }


def check_set_weights(img_class="faz", n_batches=2, batch_size=6, img_size=128, use_graph=False):
    """MyModel.set_weights of the reference (flower_common.py:491-633) and fedicra_b200.adaptive.adaptive_set_weights on
    two copies of one client, same global weights, same batches: returns the largest relative difference of the
    initialised decoder parameters and of the learned aggregation weights, plus timings."""
    import copy
    import torch
    from fedicra_b200.adaptive import adaptive_set_weights
    ours, fc, args, client = build_client("modules", img_class, 4, 3, n_batches, batch_size, img_size)
    wrapper = client.model
    g = torch.Generator().manual_seed(7)
    weights = []
    trainable = {n for n, _ in wrapper.model.named_parameters()}
    for k, v in wrapper.model.state_dict().items():        # a "global model": the local one, parameters perturbed
        a = v.detach().cpu()
        weights.append((a + 0.02 * torch.randn(a.shape, generator=g)).numpy() if k in trainable else a.numpy())
    config = {"iter_global": 51}
    a_w, b_w = copy.deepcopy(wrapper), copy.deepcopy(wrapper)
    out = {}
    for tag, w, fn in (("reference", a_w, lambda: fc.MyModel.set_weights(a_w, weights, config)),
                       ("ours", b_w, lambda: adaptive_set_weights(b_w, weights, config, use_graph=use_graph,
                                                                  log=lambda *x: None))):
        torch.manual_seed(3)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        out[tag + "_seconds"] = time.perf_counter() - t0
    keys = [n for n, _ in a_w.model.named_parameters() if any(k in n for k in ("out_conv", "up4", "up3", "up2", "up1"))]
    pa, pb, p0 = (dict(m.model.named_parameters()) for m in (a_w, b_w, wrapper))
    scale = max(float(pa[k].data.abs().max()) for k in keys)          # tensors that are all (near) zero share the global scale
    rel = lambda x, y: float((x.double() - y.double()).abs().max() / max(float(y.abs().max()), 1e-2 * scale))
    out["param_rel"] = max(rel(pb[k].data, pa[k].data) for k in keys)
    out["weight_rel"] = max(rel(y, x) for x, y in zip(a_w.fedaa_weights, b_w.fedaa_weights))
    out["other_params_equal"] = all(torch.equal(pa[k].data, pb[k].data) for k in pa if k not in keys)
    out["moved"] = max(rel(pa[k].data, p0[k].data) for k in keys)
    out["n_tensors"], out["use_graph"] = len(keys), use_graph
    out["start_phase_after"] = [bool(a_w.start_phase), bool(b_w.start_phase)]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="modules", choices=["reference", "ops", "modules"])
    ap.add_argument("--mode", default="train", choices=["train", "set_weights", "set_weights_graph"])
    ap.add_argument("--img-class", default="faz", choices=["faz", "odoc"])
    ap.add_argument("--iters", type=int, default=4)
    ap.add_argument("--rep-iters", type=int, default=3)
    ap.add_argument("--batches", type=int, default=2)
    ap.add_argument("--batch-size", type=int, default=12)
    ap.add_argument("--img-size", type=int, default=256)
    ap.add_argument("--warmup-iters", type=int, default=0)
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    if a.mode != "train":
        res = check_set_weights(a.img_class, a.batches, min(a.batch_size, 6), min(a.img_size, 128), a.mode.endswith("graph"))
    else:
        res = run(a.backend, a.img_class, a.iters, a.rep_iters, a.batches, a.batch_size, a.img_size, a.warmup_iters)
    text = json.dumps(res)
    if a.json:
        with open(a.json, "w") as f:
            f.write(text)
    print(text)


if __name__ == "__main__":
    main()
