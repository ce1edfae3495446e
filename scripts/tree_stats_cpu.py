"""Tree statistics (CPU oracle) for the bench inputs: depth, level widths, band plan, leaves, chain lengths."""
import sys; sys.path.insert(0, '.')
import numpy as np, torch
from oracle import tree_cpu
tree_cpu.build(ref=False)
sys.path.insert(0, '.')
from bench import make_inputs

def grid_edges(emb):
    B, D, H, W = emb.shape
    idx = torch.arange(H*W, dtype=torch.int32).view(H, W)
    ev = torch.stack([idx[:-1].reshape(-1), idx[1:].reshape(-1)], 1)
    eh = torch.stack([idx[:, :-1].reshape(-1), idx[:, 1:].reshape(-1)], 1)
    ei = torch.cat([ev, eh], 0).unsqueeze(0).expand(B, -1, -1).contiguous()
    wv = ((emb[:, :, :-1] - emb[:, :, 1:])**2).sum(1).reshape(B, -1)
    wh = ((emb[:, :, :, :-1] - emb[:, :, :, 1:])**2).sum(1).reshape(B, -1)
    return ei, torch.cat([wv, wh], 1) + 1

def stats(name, emb, pmax=2560, kdiv=32):
    B, D, H, W = emb.shape; V = H*W
    ei, ew = grid_edges(emb)
    tree = tree_cpu.mst_forward(ei, ew, V)
    si, sp, sc = tree_cpu.bfs_forward(tree, 4)
    for b in range(min(B, 2)):
        par = sp[b].numpy().astype(np.int64)
        depth = np.zeros(V, np.int64)
        for i in range(1, V): depth[i] = depth[par[i]] + 1
        nl = depth.max() + 1
        widths = np.bincount(depth)
        nchild = (sc[b].numpy() > 0).sum(1)
        ls = np.concatenate([[0], np.cumsum(widths)])
        kmax = max(4, -(-nl // kdiv))
        cuts = [0]; L0 = 0
        while L0 < nl:
            L1 = L0 + 1
            while L1 < min(L0 + kmax, nl) and ls[L1 + 1] - ls[L0] <= pmax: L1 += 1
            cuts.append(L1); L0 = L1
        nb = len(cuts) - 1
        lastw = [widths[c - 1] for c in cuts[1:]]
        print(f"{name} img{b}: V={V} depth={nl} width mean={widths.mean():.1f} max={widths.max()} leaves={np.mean(nchild==0):.3f} unary={np.mean(nchild==1):.3f} binary={np.mean(nchild==2):.3f} tern={np.mean(nchild==3):.3f} | nb={nb} kmax={kmax} lastlevel width mean={np.mean(lastw):.1f} max={np.max(lastw)}")

d = make_inputs(2, 1, 2, 256, 256, 2022)
img3 = d['image'].repeat(1, 3, 1, 1)
stats('faz image', img3)
a1 = torch.nn.functional.interpolate(d['aux1'], size=(256, 256), mode='bilinear')
stats('faz aux1', a1)
a2 = torch.nn.functional.interpolate(d['aux2'], size=(256, 256), mode='bilinear')
stats('faz aux2', a2)
stats('faz aux3', d['aux3'])
