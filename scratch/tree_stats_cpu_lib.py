import sys; sys.path.insert(0, '.')
import numpy as np, torch
from oracle import tree_cpu
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

def level_starts(B=3):
    d = make_inputs(B, 1, 2, 256, 256, 2022)
    embs = {'img': d['image'].repeat(1, 3, 1, 1),
            'aux1': torch.nn.functional.interpolate(d['aux1'], size=(256, 256), mode='bilinear'),
            'aux2': torch.nn.functional.interpolate(d['aux2'], size=(256, 256), mode='bilinear'),
            'aux3': d['aux3']}
    out = []
    for name, emb in embs.items():
        V = 65536
        ei, ew = grid_edges(emb)
        tree = tree_cpu.mst_forward(ei, ew, V)
        si, sp, sc = tree_cpu.bfs_forward(tree, 4)
        for b in range(B):
            par = sp[b].numpy().astype(np.int64)
            depth = np.zeros(V, np.int64)
            for i in range(1, V): depth[i] = depth[par[i]] + 1
            widths = np.bincount(depth)
            out.append((f"{name}{b}", np.concatenate([[0], np.cumsum(widths)]).tolist()))
    return out
