"""Print the per-tensor errors (vs the fp64 oracle) of the three back ends of the unmodified reference loss module."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import ref_gpu
from tests.helpers import loss_inputs, rel_err
from tests.test_gpu_reference_kernels import _smooth_run, _fp64_judge
import fedicra_b200 as fb

for (B, C, Cin, H, W, kind) in [(2, 2, 1, 64, 64, "quant8"), (2, 3, 3, 48, 80, "uniform"), (4, 2, 1, 256, 256, "quant8")]:
    image, logits, a1, a2, a3, label = loss_inputs(B, C, Cin, H, W, kind)
    three = image.repeat(1, 3, 1, 1) if Cin == 1 else image
    roi = label == C
    R = torch.randn(B, C, H, W, generator=torch.Generator().manual_seed(1))
    truth = _fp64_judge(logits, three, a1, a2, a3, roi, R)
    ref_gpu.install("reference")
    fc = ref_gpu.flower_common()
    runs = {}
    runs["reference"] = _smooth_run(fc.MScaleRecurveTreeEnergyLoss(), logits, three, a1, a2, a3, roi, R)
    ref_gpu.use("fedicra")
    runs["ops"] = _smooth_run(fc.MScaleRecurveTreeEnergyLoss(), logits, three, a1, a2, a3, roi, R)
    ref_gpu.use("reference")
    runs["fused"] = _smooth_run(fb.MScaleRecurveTreeEnergyLoss(), logits, three, a1, a2, a3, roi, R)
    for k, got in runs.items():
        e = [rel_err(got[0], truth[0])] + [rel_err(a, b) for a, b in zip(got[1], truth[1])] + [rel_err(a, b) for a, b in zip(got[2], truth[2])]
        print((B, C, H, W), k, " ".join(f"{x:.2e}" for x in e), flush=True)
