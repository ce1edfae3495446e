"""Would two half batches on two streams beat one full batch on one?  (Kernels of the step are latency-bound; do two
independent chains fill each other's gaps?)  Times graph replays: one B=12 graph, one B=6 graph, two B=6 graphs on two
streams launched together."""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
from fedicra_b200.runner import LossStackStep, GraphedLossStack

_, Cin, C, H, W, _ = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "faz256"]
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
step = LossStackStep(C, H, W, 0.1, 0.1)
def graph(B, seed):
    d = {k: v.to(dev) for k, v in bench.make_inputs(B, Cin, C, H, W, seed).items()}
    return GraphedLossStack(step, d)
g12, g6a, g6b = graph(12, 1), graph(6, 2), graph(6, 3)
sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
def both():
    cur = torch.cuda.current_stream()
    sa.wait_stream(cur); sb.wait_stream(cur)
    with torch.cuda.stream(sa): g6a.graph.replay()
    with torch.cuda.stream(sb): g6b.graph.replay()
    cur.wait_stream(sa); cur.wait_stream(sb)
med = lambda f: statistics.median(bench.time_steps(f, 15, flush)) * 1000
print(f"one graph of 12 images {med(g12.graph.replay):.0f} us | one graph of 6 images {med(g6a.graph.replay):.0f} us | two graphs of 6 on two streams {med(both):.0f} us")
