"""Time mst_grid and bfs_euler alone for the 48 trees of a faz256 step."""
import os, sys; sys.path.insert(0, '.')
import torch
from fedicra_b200 import ops
g = torch.Generator().manual_seed(0)
B, H, W = 48, 256, 256
emb = ((torch.rand(B, 3, H, W, generator=g) * 255).round() / 255).cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
def run(tag, fn, reps=10):
    for _ in range(3): fn()
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    ts.sort()
    print(f"{tag}: median {ts[len(ts)//2]*1e3:.1f} us  min {ts[0]*1e3:.1f} us")
tree, mask = ops.mst_grid(emb, want_mask=True, want_edges=False)
run("mst_grid CL=" + os.environ.get("FEDICRA_MST_CL", "auto"), lambda: ops.mst_grid(emb, want_mask=True, want_edges=False))
run("bfs_euler", lambda: ops.bfs_euler(mask, H, W)); sys.stdout.flush()
