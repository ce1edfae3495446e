"""Per-item timeline of one chain-filter forward launch (FEDICRA_PROF=1): when each group's items were claimed, ready,
unblocked and published."""
import os, sys, ctypes
os.environ["FEDICRA_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from fedicra_b200 import ops
from fedicra_b200._lib import lib

B, C, H, W = 12, 2, int(os.environ.get("HW", 256)), int(os.environ.get("HW", 256))
TILE = 512
g = torch.Generator().manual_seed(1)
V = H * W
emb = torch.nn.functional.interpolate(torch.randn(B, C, H // 4, W // 4, generator=g), size=(H, W), mode="bilinear").cuda()
_, mask = ops.mst_grid(emb, want_mask=True)
feat = torch.rand(B, C, V, generator=g).cuda()
si, sp, sc, gs = ops.chain_order(mask, H, W)
cew = ops.edge_weight_fwd(emb.reshape(B, C, V), si, sp, 1.0)
for _ in range(3):
    out, sv = ops.chain_filter_forward(feat, cew, si, sp, sc, gs)
torch.cuda.synchronize()
n = 2 * 2 * 768 * 4
buf = (ctypes.c_longlong * n)()
assert lib.fedicra_debug_chain_profile(buf, n) == n
t = np.frombuffer(buf, dtype=np.int64).reshape(2, 2, 768, 4).astype(np.float64)
gsl = gs[0].tolist()
ntile = [((gsl[r + 1] - 1) // TILE - gsl[r] // TILE + 1) if gsl[r + 1] > gsl[r] else 0 for r in range(32)]
cum = np.concatenate([[0], np.cumsum(ntile)])
total = int(cum[-1])
t0 = t[0, 0, :total][t[0, 0, :total] > 0].min()
print("image 0: groups", [x for x in gsl if x < V], "items per group", [x for x in ntile if x], "total", total)
for phase, name in ((0, "leaf->root"), (1, "root->leaf")):
    rng = range(31, -1, -1) if phase == 0 else range(32)
    for r in rng:
        if ntile[r] == 0:
            continue
        x = (t[phase, 0, cum[r]:cum[r + 1]] - t0) / 1000.0
        print(f"  {name} group {r}: {ntile[r]:3d} items | claimed {x[:,0].min():7.1f}..{x[:,0].max():7.1f} | ready {x[:,1].min():7.1f}..{x[:,1].max():7.1f}"
              f" | unblocked {x[:,2].min():7.1f}..{x[:,2].max():7.1f} | published {x[:,3].min():7.1f}..{x[:,3].max():7.1f} us"
              f" | mean wait {np.mean(x[:,2]-x[:,1]):5.1f} mean work {np.mean(x[:,3]-x[:,2]):5.1f} mean load {np.mean(x[:,1]-x[:,0]):5.1f}")
