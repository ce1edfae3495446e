"""Adaptive-aggregation kernels against the HBM roofline: flat update at the decoder size and at a size
that does not fit L2, the 42-tensor single launch, and the FedAvg pre-scale."""
import json, sys; sys.path.insert(0, '.')
import torch
from fedicra_b200 import ops, fused_ala_update
peak = json.load(open('MEASURED_PEAKS.json'))['hbm_gbs']
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
def run(fn, reps=20, cold=True):
    for _ in range(3): fn()
    ts = []
    for _ in range(reps):
        if cold: flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    ts.sort(); return ts[len(ts) // 2] * 1e-3
g = torch.Generator().manual_seed(0)
for n in (633010, 64 << 20):
    p, s, gr = (torch.randn(n, generator=g).cuda() for _ in range(3))
    w, t = torch.ones(n, device='cuda'), torch.empty(n, device='cuda')
    for cold in (True, False):
        dt = run(lambda: ops.ala_update_flat(p, s, gr, w, t, 1.0), cold=cold)
        print(f"ala_update_flat n={n} {'cold' if cold else 'warm'}: {dt*1e6:.1f} us, {6*4*n/dt/1e9:.0f} GB/s = {6*4*n/dt/1e9/peak:.2f} of measured HBM peak")
    dt = run(lambda: ops.scale_inplace(p, 0.5), cold=True)
    print(f"scale_inplace   n={n} cold: {dt*1e6:.1f} us, {2*4*n/dt/1e9:.0f} GB/s = {2*4*n/dt/1e9/peak:.2f}")
# the 42 decoder tensors of UNet_LC_MultiHead: sizes summing to 633 010
sizes = [476032 // 8] * 8 + [119232 // 8] * 8 + [29920 // 8] * 8 + [7536 // 8] * 8 + [29] * 10
ps = [torch.randn(k, generator=g).cuda() for k in sizes]; ss = [torch.randn(k, generator=g).cuda() for k in sizes]
ts_ = [torch.zeros(k, device='cuda').requires_grad_() for k in sizes]; ws = [torch.ones(k, device='cuda') for k in sizes]
for t_ in ts_: t_.grad = torch.randn_like(t_)
dt = run(lambda: fused_ala_update(ps, ss, ts_, ws, 1.0), cold=True)
print(f"fused_ala_update {len(sizes)} tensors / {sum(sizes)} elements cold: {dt*1e6:.1f} us (one launch + pointer-table upload)")
