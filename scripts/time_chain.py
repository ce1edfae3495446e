"""Time the chain-order builder and the chain filter next to the BFS order / depth-banded filter (one B200)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedicra_b200 import ops


def timeit_graph(fn, reps=20, n=10):
    """Device time per call with the host out of the picture: `reps` calls captured into one CUDA graph."""
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


def timeit(fn, n=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1000)
    ts.sort()
    return ts[len(ts) // 2]


for (B, C, H, W, K) in ((12, 2, 256, 256, 4), (12, 3, 256, 256, 4), (12, 3, 512, 512, 4)):
    g = torch.Generator().manual_seed(1)
    V = H * W
    embs = [((torch.rand(B, 3, H, W, generator=g) * 255).round() / 255).cuda()] + \
           [torch.nn.functional.interpolate(torch.randn(B, C, H // s, W // s, generator=g), size=(H, W), mode="bilinear").cuda() for s in (4, 2, 1)]
    _, mask = ops.mst_grid_multi(embs, want_mask=True, want_edges=False)
    t_bfs = timeit(lambda: ops.bfs_euler(mask, H, W))
    t_chain = timeit(lambda: ops.chain_order(mask, H, W))
    print(f"B={B} C={C} {H}x{W}: order of {K*B} trees: bfs_euler {t_bfs:.0f} us, chain_order {t_chain:.0f} us")
    feat = torch.rand(B, C, V, generator=g).cuda()
    gout = torch.randn(B, C, V, generator=g).cuda()
    m1 = mask[B:2 * B].contiguous()
    emb = embs[1].reshape(B, C, V)
    order = ops.bfs_euler(m1, H, W)
    bew = ops.edge_weight_fwd(emb, order[0], order[1], 1.0)
    bands = ops.band_plan(order[3], order[4])
    _, bsv = ops.tree_filter_forward(feat, bew, *order, bands)
    t_bf = timeit(lambda: ops.tree_filter_forward(feat, bew, *order, bands))
    t_bb = timeit(lambda: ops.tree_filter_backward(gout, bew, *order, bands, bsv, True))
    si, sp, sc, gs = ops.chain_order(m1, H, W)
    cew = ops.edge_weight_fwd(emb, si, sp, 1.0)
    _, csv = ops.chain_filter_forward(feat, cew, si, sp, sc, gs)
    t_cf = timeit_graph(lambda: ops.chain_filter_forward(feat, cew, si, sp, sc, gs))
    t_cb = timeit_graph(lambda: ops.chain_filter_backward(gout, cew, si, sp, sc, gs, csv, True))
    t_bf = timeit_graph(lambda: ops.tree_filter_forward(feat, bew, *order, bands))
    t_bb = timeit_graph(lambda: ops.tree_filter_backward(gout, bew, *order, bands, bsv, True))
    t_chain = timeit_graph(lambda: ops.chain_order(mask, H, W), reps=5)
    print(f"    (graph replay) chain_order {t_chain:.0f} us")
    print(f"    filter forward: bands {t_bf:.0f} us, chains {t_cf:.0f} us;  backward: bands {t_bb:.0f} us, chains {t_cb:.0f} us"
          f"   groups {[int(x) for x in gs[0].tolist() if x < V]}")
