"""How much of a grid MST is left after the tile-local Boruvka rounds (CPU simulation of csrc/mst.cu's tile kernel):
components and cross edges (edges between different components) per tree, for the bench's embeddings."""
import sys
import numpy as np
import torch

def stats(emb, TS=64, rounds=(2, 3, 4, 5, 6, 8)):
    D, H, W = emb.shape
    V = H * W
    idx = np.arange(V).reshape(H, W)
    eu = np.concatenate([idx[:-1].ravel(), idx[:, :-1].ravel()])
    ev = np.concatenate([idx[1:].ravel(), idx[:, 1:].ravel()])
    f = emb.reshape(D, V).astype(np.float32)
    w = ((f[:, eu] - f[:, ev]) ** 2).sum(0).astype(np.float32) + 1
    key = (w.view(np.uint32).astype(np.uint64) << np.uint64(32)) | np.arange(len(eu), dtype=np.uint64)
    tile = (np.arange(V) // W // TS) * ((W + TS - 1) // TS) + (np.arange(V) % W) // TS
    same_tile = tile[eu] == tile[ev]
    comp = np.arange(V)
    out = {}
    for r in range(1, max(rounds) + 1):
        cu, cv = comp[eu], comp[ev]
        cross = cu != cv
        best = np.full(V, np.iinfo(np.uint64).max, dtype=np.uint64)
        np.minimum.at(best, cu[cross], key[cross])
        np.minimum.at(best, cv[cross], key[cross])
        roots = np.unique(comp)
        parent = np.arange(V)
        for c in roots:
            k = best[c]
            if k == np.iinfo(np.uint64).max:
                continue
            e = int(k & np.uint64(0xffffffff))
            if not same_tile[e]:
                continue
            a, b = comp[eu[e]], comp[ev[e]]
            o = b if a == c else a
            if best[o] == k and c < o:
                continue
            parent[c] = o
        while True:
            p2 = parent[parent]
            if (p2 == parent).all():
                break
            parent = p2
        comp = parent[comp]
        if r in rounds:
            cu, cv = comp[eu], comp[ev]
            cross = cu != cv
            pairs = np.unique(np.stack([np.minimum(cu[cross], cv[cross]), np.maximum(cu[cross], cv[cross])]), axis=1).shape[1]
            out[r] = (len(np.unique(comp)), int(cross.sum()), pairs)
    return out

g = torch.Generator().manual_seed(1)
H = W = int(sys.argv[1]) if len(sys.argv) > 1 else 256
img = ((torch.rand(1, 1, H, W, generator=g) * 255).round() / 255).repeat(1, 3, 1, 1)
embs = {"image": img[0].numpy()}
for s in (4, 2, 1):
    embs[f"aux/{s}"] = torch.nn.functional.interpolate(torch.randn(1, 2, H // s, W // s, generator=g), size=(H, W), mode="bilinear")[0].numpy()
for ts in (64, 128):
    for name, e in embs.items():
        print(f"TS={ts} {name:8s}", {r: v for r, v in stats(e, ts).items()}, "  (components, cross edges, distinct component pairs)")
