"""Phase timeline of the MST contraction kernel for tree 0 (FEDICRA_PROF=1)."""
import os, sys, ctypes
os.environ["FEDICRA_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedicra_b200 import ops
from fedicra_b200._lib import lib

B, C = 12, 2
H = W = int(os.environ.get("HW", 256))
g = torch.Generator().manual_seed(1)
embs = [((torch.rand(B, 1, H, W, generator=g) * 255).round() / 255).repeat(1, 3, 1, 1).cuda()] + \
       [torch.nn.functional.interpolate(torch.randn(B, C, H // s, W // s, generator=g), size=(H, W), mode="bilinear").cuda() for s in (4, 2, 1)]
for _ in range(3):
    ops.mst_grid_multi(embs, want_mask=True, want_edges=False)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 128)()
assert lib.fedicra_debug_mst_profile(buf, 128) == 128
v = list(buf)
mhz = 1965.0
t0, n, R, t1 = v[0], v[1], v[2], v[3]
print(f"tree 0: {n} records, {R} components; load {(t1 - t0) / mhz:.1f} us")
i, prev, rnd = 4, t1, 0
while i + 5 < 128 and v[i] != -1:
    if v[i + 1] == -1 or v[i + 2] == -1:
        break
    f, l, hk = v[i], v[i + 1], v[i + 2]
    if v[i + 3] == -1:
        print(f"round {rnd}: find {(f-prev)/mhz:.1f} locate {(l-f)/mhz:.1f}, nothing hooked"); break
    ch, cp, left = v[i + 3], v[i + 4], v[i + 5]
    print(f"round {rnd}: find {(f-prev)/mhz:5.1f} locate {(l-f)/mhz:5.1f} hook {(hk-l)/mhz:5.1f} chase {(ch-hk)/mhz:5.1f} compact {(cp-ch)/mhz:5.1f} us -> {left} records")
    prev, i, rnd = cp, i + 6, rnd + 1
print(f"total {(prev - t0) / mhz:.1f} us")
t = v[64:]
if t[0] > 0:
    names = {-10: "find", -11: "min weight", -12: "min code", -13: "hook"}
    out, prev, k, rnd = [f"weights {(t[1] - t[0]) / mhz:.1f}"], t[1], 2, 0
    while k < 64 and t[k] not in (0, -1, -2):
        if t[k] in names:
            out.append(f"{names[t[k]]} {(t[k + 1] - prev) / mhz:.1f}")
            prev, k = t[k + 1], k + 2
        else:
            out.append(f"rest {(t[k] - prev) / mhz:.1f} (round {rnd} done) |")
            prev, k, rnd = t[k], k + 1, rnd + 1
    if k < 61 and t[k] == -2:
        for j, lab in enumerate(["ranks+classify", "pair table", "emit"]):
            cur = t[k + 1 + j]
            out.append(f"{lab} {(cur - prev) / mhz:.1f}")
            prev = cur
        out.append(f"| total {(prev - t[0]) / mhz:.1f}")
    print("tile 0 of tree 0 (us):", " ".join(out))
