"""Time tree_filter forward/backward alone (CUDA events) and print the resident kernel's phase stamps.
   FEDICRA_FILTER=multi|resident  FEDICRA_PROF=1"""
import ctypes, os, sys; sys.path.insert(0, '.')
import torch
from fedicra_b200 import ops
from fedicra_b200._lib import lib
g = torch.Generator().manual_seed(0)
B, C, H, W = 12, int(os.environ.get("C", 2)), int(os.environ.get("H", 256)), int(os.environ.get("H", 256))
emb = ((torch.rand(B, 3, H, W, generator=g) * 255).round() / 255).cuda()
feat = torch.rand(B, C, H * W, generator=g).cuda(); gout = torch.randn(B, C, H * W, generator=g).cuda()
tree, mask = ops.mst_grid(emb, want_mask=True)
order = ops.bfs_from_mask(mask, H, W)
bands = ops.band_plan(order[3], order[4])
pl = bands.view(B, -1).cpu()
print("nb per image:", pl[:, 0].tolist(), "n_levels:", order[4].cpu().tolist())
ew = ops.edge_weight_fwd(emb.reshape(B, 3, -1), order[0], order[1], 50.0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
NAMES_BAND = ["start", "gathered", "U1 done+arrive", "regathered", "flag_up seen", "U3 done", "D1 done+arrive", "flag_down seen", "D3 done", "epilogue done"]
NAMES_ST = ["tables", "preloaded", "cnt_up seen", "X loaded", "U2 chain done", "flag_up set", "cnt_down seen", "B loaded", "D2 chain done", "flag_down set"]

def prof(tag):
    if not os.environ.get("FEDICRA_PROF"): return
    n = lib.fedicra_debug_filter_profile(None, 0)
    buf = (ctypes.c_longlong * n)()
    lib.fedicra_debug_filter_profile(buf, n)
    import numpy as np
    a = np.array(buf[:]).reshape(-1, 12)
    nb = int(pl[0, 0])
    grid = lib_nbcap + 1
    st = a[grid - 1]; t0 = a[:nb, 0].min()
    print(f"--- {tag}: stitcher (us from first band start)")
    print("   " + "  ".join(f"{nm}={(st[i]-t0)/1e3:.1f}" for i, nm in enumerate(NAMES_ST)))
    bands_t = (a[:nb, :10] - t0) / 1e3
    print("   bands: mean / max per stamp")
    for i, nm in enumerate(NAMES_BAND):
        print(f"     {nm:18s} mean {bands_t[:, i].mean():7.1f}  min {bands_t[:, i].min():7.1f}  max {bands_t[:, i].max():7.1f}")

lib_nbcap = int(os.environ.get("FEDICRA_PLAN_NBCAP", 36 if H * W <= (1 << 17) else 72))
for name, fn in (("fwd", lambda: ops.tree_filter_forward(feat, ew, *order, bands)),):
    pass
out, sv = ops.tree_filter_forward(feat, ew, *order, bands)
def run(tag, fn, reps=20):
    for _ in range(3): fn()
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    ts.sort()
    print(f"{tag}: median {ts[len(ts)//2]*1e3:.1f} us  min {ts[0]*1e3:.1f} us")
run("forward ", lambda: ops.tree_filter_forward(feat, ew, *order, bands)); prof("forward")
run("backward+gw", lambda: ops.tree_filter_backward(gout, ew, *order, bands, sv, True)); prof("backward+gw")
run("backward   ", lambda: ops.tree_filter_backward(gout, ew, *order, bands, sv, False)); prof("backward")
