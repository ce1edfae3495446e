"""Where the step goes: sub-sequences of the loss-stack step captured into CUDA graphs and replayed (device time, L2
flushed between replays).  faz256 shape by default."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from fedicra_b200 import ops
from fedicra_b200.runner import LossStackStep, KERNELS_DESC, RADIUS
from fedicra_b200.tree_energy import MScaleRecurveTreeEnergyLoss, _resize_heads, _softmax
from fedicra_b200.gate_crf_loss import ModelLossSemsegGatedCRF
from fedicra_b200.tree_filter import batched_tree_orders

name = sys.argv[1] if len(sys.argv) > 1 else "faz256"
B, Cin, C, H, W, _ = bench.WORKLOADS[name]
dev = torch.device("cuda:0")
d = {k: v.to(dev) for k, v in bench.make_inputs(B, Cin, C, H, W, 2022).items()}
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timeit(fn, n=15):
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3): fn()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):   # same stream as the warm-up: scratch buffers are cached per stream
        keep = fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1000)
    ts.sort(); return ts[len(ts) // 2]

tel, crf = MScaleRecurveTreeEnergyLoss(), ModelLossSemsegGatedCRF()
three = d["image"].repeat(1, 3, 1, 1) if Cin == 1 else d["image"]
roi = d["label"] == C

def topo():
    highs = _resize_heads([d["aux1"], d["aux2"], d["aux3"]], H, W)
    return batched_tree_orders([three] + highs, [50.0, 1.0, 1.0, 1.0])

def tel_fwd():
    with torch.no_grad():
        return tel(d["logits"], three, d["aux1"], d["aux2"], d["aux3"], roi, 0.1)[0]

def tel_fb():
    leaves = [d[k].detach().requires_grad_() for k in ("logits", "aux1", "aux2", "aux3")]
    loss = tel(leaves[0], three, leaves[1], leaves[2], leaves[3], roi, 0.1)[0]
    loss.backward()
    return loss, [x.grad for x in leaves]

def crf_fb():
    x = d["logits"].detach().requires_grad_()
    loss = crf(torch.softmax(x, dim=1), KERNELS_DESC, RADIUS, d["image"], H, W)["loss"]
    loss.backward()
    return loss, x.grad

step = LossStackStep(C, H, W, 0.1, 0.1)
print(f"{name}: topology (resize + MST + order + weights) {timeit(topo):.0f} us | TEL forward {timeit(tel_fwd):.0f} us | "
      f"TEL forward+backward {timeit(tel_fb):.0f} us | CRF forward+backward {timeit(crf_fb):.0f} us | whole step {timeit(lambda: step(d)):.0f} us")
os.environ["FEDICRA_OVERLAP"] = "0"
print(f"   whole step without the helper stream {timeit(lambda: step(d)):.0f} us")
