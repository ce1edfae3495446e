"""Heavy-path decomposition statistics of the bench trees (CPU oracle): light depth, group sizes, path lengths."""
import sys; sys.path.insert(0, '.')
import numpy as np, torch
from oracle import tree_cpu, losses_cpu
tree_cpu.build(ref=False)
from bench import make_inputs

def hld(si, sp, sc):
    V = si.shape[0]
    par = sp.astype(np.int64)
    size = np.ones(V, np.int64)
    for i in range(V - 1, 0, -1):
        size[par[i]] += size[i]
    heavy = np.full(V, -1, np.int64)
    best = np.zeros(V, np.int64)
    for i in range(1, V):        # positions increase => first child with max size wins
        p = par[i]
        if size[i] > best[p]:
            best[p] = size[i]; heavy[p] = i
    ld = np.zeros(V, np.int64)
    head = np.zeros(V, np.int64)
    for i in range(1, V):
        p = par[i]
        if heavy[p] == i:
            ld[i] = ld[p]; head[i] = head[p]
        else:
            ld[i] = ld[p] + 1; head[i] = i
    return size, heavy, ld, head

def stats(name, emb):
    B = emb.shape[0]
    tree = losses_cpu.minimum_spanning_tree(emb)
    si, sp, sc = tree_cpu.bfs_forward(tree, 4)
    for b in range(min(B, 2)):
        size, heavy, ld, head = hld(si[b].numpy(), sp[b].numpy(), sc[b].numpy())
        V = len(ld)
        groups = np.bincount(ld)
        plen = np.bincount(head)  # length of path per head
        plen = plen[plen > 0]
        npaths = len(plen)
        print(f"{name} img{b}: V={V} max light depth={ld.max()} groups={groups.tolist()}")
        print(f"    paths={npaths} len1={np.mean(plen==1):.3f} len<=4={np.mean(plen<=4):.3f} len<=32={np.mean(plen<=32):.3f} max={plen.max()} "
              f"vertices in paths>32: {plen[plen>32].sum()/V:.3f}  >256: {plen[plen>256].sum()/V:.3f} >1024: {plen[plen>1024].sum()/V:.3f}")

for (H, W) in ((256, 256), (512, 512)):
    d = make_inputs(2, 1 if H == 256 else 3, 2 if H == 256 else 3, H, W, 2022)
    img3 = d['image'].repeat(1, 3, 1, 1) if d['image'].shape[1] == 1 else d['image']
    stats(f'{H} image', img3)
    F = torch.nn.functional.interpolate
    stats(f'{H} aux1', F(d['aux1'], size=(H, W), mode='bilinear'))
    stats(f'{H} aux2', F(d['aux2'], size=(H, W), mode='bilinear'))
    stats(f'{H} aux3', d['aux3'])
