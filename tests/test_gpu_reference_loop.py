"""The reference's training loop, unchanged, on the drop-in (BASELINE.json configs[2]).

`MyClient._train` of flower_pCE_2D_GateCRFMsacleTreeEnergyLoss_Ours.py (copy under oracle/_ref/py) drives
`UNet_LC_MultiHead` for a few iterations on synthetic batches, once per back end, one process each (the reference binds
its imports at import time): the reference's own CUDA extension, `install_dropin("ops")`, `install_dropin("modules")`.
"""
import json
import os
import subprocess
import sys

import pytest

from oracle import ref_gpu

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_gpu.available(), reason="oracle/_ref was not built (needs /root/reference)")]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_backend(backend, img_class, iters, batch_size, img_size=256):
    cmd = [sys.executable, "-m", "oracle.train_harness", "--backend", backend, "--img-class", img_class, "--iters",
           str(iters), "--rep-iters", "3", "--batch-size", str(batch_size), "--img-size", str(img_size)]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (backend, r.stdout[-2000:], r.stderr[-4000:])
    return json.loads(r.stdout.strip().splitlines()[-1])


@pytest.mark.parametrize("img_class,batch_size", [("faz", 12), ("odoc", 6)])
def test_reference_train_loop_runs_unchanged_on_the_dropin(img_class, batch_size):
    iters = 4          # 1 iteration of phase 1 (only out_conv trains) + rep_iters = 3 of phase 2 (OURS.py:90-107)
    ref = run_backend("reference", img_class, iters, batch_size)
    ref2 = run_backend("reference", img_class, iters, batch_size)      # the reference against itself: run-to-run noise
    mods = run_backend("modules", img_class, iters, batch_size)
    ops_level = run_backend("ops", img_class, iters, batch_size)
    # who served the loss classes
    assert ref["gated_crf_module"] == "utils.gate_crf_loss" and ref["tree_filter_module"].endswith("modules.tree_filter")
    assert mods["gated_crf_module"] == "fedicra_b200.gate_crf_loss" and mods["tree_filter_module"] == "fedicra_b200.tree_filter"
    assert ops_level["gated_crf_module"] == "utils.gate_crf_loss"
    for r in (ref, mods, ops_level):
        assert r["script_file"].startswith(ref_gpu.PY_DIR) and r["flower_common_file"].startswith(ref_gpu.PY_DIR)
        assert len(r["per_iter"]) == iters
    rel = lambda a, b: abs(a - b) / max(abs(b), 1e-12)
    # iteration 1: same weights, same data -> the loss terms agree to fp32 accuracy
    for got in (mods, ops_level):
        for key in ("loss", "loss_ce", "loss_tree"):
            assert rel(got["per_iter"][0][key], ref["per_iter"][0][key]) < 2e-4, (got["backend"], key, got["per_iter"][0], ref["per_iter"][0])
    # later iterations: AdamW's sign-like first steps amplify last-bit gradient differences, for the reference against
    # itself as well (its scatter_add is atomic); the drop-in must stay inside a small multiple of that noise
    for k in range(1, iters):
        noise = max(rel(ref2["per_iter"][k]["loss"], ref["per_iter"][k]["loss"]), 1e-3)
        for got in (mods, ops_level):
            assert rel(got["per_iter"][k]["loss"], ref["per_iter"][k]["loss"]) < 20 * noise, (got["backend"], k, got["per_iter"], ref["per_iter"], ref2["per_iter"])


def test_adaptive_set_weights_equals_the_reference_set_weights():
    """fedicra_b200.adaptive.adaptive_set_weights (fused element-wise update, no dummy optimizer step) against the
    unmodified MyModel.set_weights (flower_common.py:491-633) on two copies of one UNet_LC_MultiHead client."""
    cmd = [sys.executable, "-m", "oracle.train_harness", "--mode", "set_weights", "--img-class", "faz"]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-4000:])
    res = json.loads(r.stdout.strip().splitlines()[-1])
    assert res["n_tensors"] == 42 and res["moved"] > 1e-3            # the call did initialise the 42 decoder tensors
    assert res["param_rel"] < 1e-4 and res["weight_rel"] < 1e-4, res
    assert res["other_params_equal"] and res["start_phase_after"] == [False, False]
