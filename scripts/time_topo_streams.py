"""Topology stage (MST + chain order) of K*B trees: one batch on one stream vs groups of trees on concurrent streams."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedicra_b200 import ops


def replay_time(fn, reps=5, n=10):
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):   # same stream as the warm-up: scratch buffers are cached per stream
        for _ in range(reps):
            fn()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1000 / reps)
    ts.sort()
    return ts[len(ts) // 2]


STREAMS = [torch.cuda.Stream() for _ in range(4)]


def topo(embs, H, W, groups):
    """groups: list of lists of embedding indices; each group runs on its own stream."""
    main = torch.cuda.current_stream()
    outs = []
    if len(groups) == 1:
        _, mask = ops.mst_grid_multi([embs[i] for i in groups[0]], want_mask=True, want_edges=False)
        return [ops.chain_order(mask, H, W)]
    for s, grp in zip(STREAMS, groups):
        s.wait_stream(main)
        with torch.cuda.stream(s):
            _, mask = ops.mst_grid_multi([embs[i] for i in grp], want_mask=True, want_edges=False)
            outs.append(ops.chain_order(mask, H, W))
    for s in STREAMS[:len(groups)]:
        main.wait_stream(s)
    return outs


for (B, C, H, W) in ((12, 2, 256, 256), (12, 3, 512, 512)):
    g = torch.Generator().manual_seed(1)
    embs = [((torch.rand(B, 3, H, W, generator=g) * 255).round() / 255).cuda()] + \
           [torch.nn.functional.interpolate(torch.randn(B, C, H // s, W // s, generator=g), size=(H, W), mode="bilinear").cuda() for s in (4, 2, 1)]
    for name, groups in (("1 stream x 4 embeddings", [[0, 1, 2, 3]]), ("2 streams x 2", [[0, 1], [2, 3]]),
                         ("4 streams x 1", [[0], [1], [2], [3]]), ("2 streams 1+3", [[0], [1, 2, 3]])):
        t = replay_time(lambda: topo(embs, H, W, groups))
        print(f"{H}x{W} B={B}: {name}: {t:.0f} us", flush=True)
    for k in (1, 2):
        t1 = replay_time(lambda: ops.mst_grid_multi(embs[:k], want_mask=True, want_edges=False))
        _, mask = ops.mst_grid_multi(embs[:k], want_mask=True, want_edges=False)
        t2 = replay_time(lambda: ops.chain_order(mask, H, W))
        print(f"   {k} embedding(s) alone: mst {t1:.0f} us, chain_order {t2:.0f} us", flush=True)
