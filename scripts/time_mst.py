"""Time fedicra_mst_grid_multi (48 trees of a faz256 / rgb512 step) from a CUDA graph, L2 flushed between replays."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedicra_b200 import ops

def timeit_graph(fn, reps=10, n=10):
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3): fn()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):   # same stream as the warm-up: scratch buffers are cached per stream
        for _ in range(reps): fn()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1000 / reps)
    ts.sort(); return ts[len(ts) // 2]

for (B, C, H, W) in ((12, 2, 256, 256), (12, 3, 512, 512)):
    g = torch.Generator().manual_seed(1)
    embs = [((torch.rand(B, 1, H, W, generator=g) * 255).round() / 255).repeat(1, 3, 1, 1).cuda()] + \
           [torch.nn.functional.interpolate(torch.randn(B, C, H // s, W // s, generator=g), size=(H, W), mode="bilinear").cuda() for s in (4, 2, 1)]
    t = timeit_graph(lambda: ops.mst_grid_multi(embs, want_mask=True, want_edges=False))
    print(f"{H}x{W}: mst_grid_multi of {4*B} trees: {t:.0f} us  (FEDICRA_MST_FINISH={os.environ.get('FEDICRA_MST_FINISH','contract')}, tile rounds {os.environ.get('FEDICRA_MST_TILE_ROUNDS','4')})")
