"""Shared input builders and comparison helpers for the test-suite."""
import numpy as np
import torch


def rel_err(a, b):
    """max|a-b| / max|b|  (per-tensor relative error)."""
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    den = b.abs().max().item()
    return (a - b).abs().max().item() / (den if den > 0 else 1.0)


def edge_set(tree):
    """Canonical form of an undirected edge list [(V-1),2]: sorted array of (min,max) pairs."""
    t = np.asarray(tree).reshape(-1, 2)
    t = np.stack([t.min(1), t.max(1)], 1)
    return t[np.lexsort((t[:, 1], t[:, 0]))]


def parents_and_depth(tree, V):
    """Parent-of-vertex (root 0) and depth arrays of a spanning tree given as an edge list (host BFS)."""
    adj = [[] for _ in range(V)]
    for a, b in np.asarray(tree).reshape(-1, 2):
        adj[a].append(b)
        adj[b].append(a)
    parent = np.full(V, -1, np.int64)
    depth = np.full(V, -1, np.int64)
    parent[0], depth[0] = 0, 0
    queue = [0]
    for u in queue:
        for v in adj[u]:
            if depth[v] < 0:
                depth[v] = depth[u] + 1
                parent[v] = u
                queue.append(v)
    return parent, depth


def make_image(kind, B, Cin, H, W, gen):
    img = torch.rand(B, Cin, H, W, generator=gen)
    if kind == "quant8":
        img = (img * 255).round() / 255
    elif kind == "constant":
        img = torch.zeros(B, Cin, H, W)
    return img


def loss_inputs(B, C, Cin, H, W, kind="uniform", seed=2022):
    """Synthetic step inputs as in SURVEY.md §8(d)."""
    g = torch.Generator().manual_seed(seed)
    image = make_image(kind, B, Cin, H, W, g)
    logits = torch.randn(B, C, H, W, generator=g)
    aux1 = torch.randn(B, C, max(H // 4, 1), max(W // 4, 1), generator=g)
    aux2 = torch.randn(B, C, max(H // 2, 1), max(W // 2, 1), generator=g)
    aux3 = torch.randn(B, C, H, W, generator=g)
    label = torch.full((B, H, W), C, dtype=torch.uint8)
    scribble = torch.rand(B, H, W, generator=g) < 0.03
    label[scribble] = torch.randint(0, C, (int(scribble.sum()),), generator=g, dtype=torch.uint8)
    return image, logits, aux1, aux2, aux3, label
