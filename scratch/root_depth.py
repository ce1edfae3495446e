"""Tree depth as a function of the root: corner (reference) vs image centre vs true tree centre."""
import sys; sys.path.insert(0, '.')
import numpy as np, torch
from collections import deque
from oracle import tree_cpu
tree_cpu.build(ref=False)
from bench import make_inputs
from scratch.tree_stats_cpu_lib import grid_edges
def depths(adj, root, V):
    d = np.full(V, -1, np.int64); d[root] = 0; q = deque([root])
    while q:
        u = q.popleft()
        for v in adj[u]:
            if d[v] < 0: d[v] = d[u] + 1; q.append(v)
    return d
d = make_inputs(2, 1, 2, 256, 256, 2022)
embs = {'img': d['image'].repeat(1, 3, 1, 1), 'aux1': torch.nn.functional.interpolate(d['aux1'], size=(256, 256), mode='bilinear'),
        'aux3': d['aux3']}
V = 65536
for name, emb in embs.items():
    ei, ew = grid_edges(emb)
    tree = tree_cpu.mst_forward(ei, ew, V)
    for b in range(2):
        adj = [[] for _ in range(V)]
        for u, v in tree[b].numpy(): adj[u].append(v); adj[v].append(u)
        d0 = depths(adj, 0, V); a = int(d0.argmax()); da = depths(adj, a, V); bb = int(da.argmax()); diam = int(da.max())
        dc = depths(adj, 128 * 256 + 128, V)
        widths = np.bincount(dc)
        print(f"{name}{b}: depth from corner {d0.max()}  from image centre {dc.max()}  diameter {diam} (best possible {(diam + 1) // 2}); centre-root level width mean {widths.mean():.0f} max {widths.max()}")
