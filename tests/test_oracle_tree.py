"""The C oracle against first principles: brute-force path products, fp64 finite differences, Kruskal."""
import numpy as np
import pytest
import torch

from oracle import losses_cpu, tree_cpu
from tests.helpers import edge_set, parents_and_depth


def kruskal(index, weight, V):
    """Unique MST under the total order (weight, edge index)."""
    order = np.lexsort((np.arange(len(weight)), weight))
    root = list(range(V))

    def find(x):
        while root[x] != x:
            root[x] = root[root[x]]
            x = root[x]
        return x

    out = []
    for e in order:
        a, b = find(index[e, 0]), find(index[e, 1])
        if a != b:
            root[a] = b
            out.append(index[e])
    return np.array(out)


@pytest.mark.parametrize("H,W,levels", [(5, 6, 4), (9, 7, 1), (12, 16, 256), (1, 9, 3), (8, 1, 3)])
def test_boruvka_equals_kruskal_under_ties(H, W, levels):
    g = torch.Generator().manual_seed(H * 100 + W)
    emb = (torch.rand(3, 2, H, W, generator=g) * levels).floor() / levels
    idx = losses_cpu.grid_edges(H, W)
    wgt = losses_cpu.grid_weights(emb)
    tree = tree_cpu.mst_forward(idx.unsqueeze(0).expand(3, -1, -1).contiguous(), wgt.contiguous(), H * W)
    for b in range(3):
        ref = kruskal(idx.numpy(), wgt[b].numpy(), H * W)
        assert np.array_equal(edge_set(tree[b]), edge_set(ref))


def test_bfs_semantics():
    H, W = 7, 9
    emb = torch.rand(2, 3, H, W, generator=torch.Generator().manual_seed(3))
    tree = losses_cpu.minimum_spanning_tree(emb)
    si, sp, sc = tree_cpu.bfs_forward(tree, 4)
    V = H * W
    for b in range(2):
        parent, depth = parents_and_depth(tree[b].numpy(), V)
        s, p, c = si[b].numpy(), sp[b].numpy(), sc[b].numpy()
        assert sorted(s.tolist()) == list(range(V)) and s[0] == 0 and p[0] == 0
        assert np.array_equal(parent[s[1:]], s[p[1:]])            # sorted_parent points at the parent's position
        assert np.all(np.diff(depth[s]) >= 0)                     # level order
        assert np.all(np.diff(p[1:]) >= 0)                        # FIFO: parents in position order
        for i in range(V):                                        # children lists
            kids = [k for k in c[i] if k > 0]
            assert all(p[k] == i for k in kids) and kids == sorted(kids)
            assert len(kids) == int(np.sum(p[1:] == i))
        _, _, _, ls = tree_cpu.bfs_levels(tree[b])
        assert ls[0] == 0 and ls[-1] == V and np.array_equal(np.searchsorted(depth[s], np.arange(depth.max() + 1)), ls[:-1].numpy())


def path_product_matrix(tree, w_of_vertex, V):
    """S[i,j] = product of edge weights on the tree path i..j (vertex order), brute force in fp64."""
    parent, depth = parents_and_depth(tree, V)
    S = np.eye(V)
    for i in range(V):
        for j in range(V):
            a, b, prod = i, j, 1.0
            while a != b:
                if depth[a] >= depth[b]:
                    prod *= w_of_vertex[a]
                    a = parent[a]
                else:
                    prod *= w_of_vertex[b]
                    b = parent[b]
            S[i, j] = prod
    return S


def test_refine_forward_is_the_path_product_filter():
    H, W, C = 5, 6, 2
    V = H * W
    g = torch.Generator().manual_seed(11)
    emb = torch.rand(1, 3, H, W, generator=g)
    feat = torch.rand(1, C, V, generator=g)
    tree = losses_cpu.minimum_spanning_tree(emb)
    si, sp, sc = tree_cpu.bfs_forward(tree, 4)
    ew = tree_cpu.edge_weight(emb.reshape(1, 3, V), si, sp, 1.0)
    w_vertex = np.zeros(V)
    w_vertex[si[0].numpy()] = ew[0].numpy()
    out, aggr, aggr_up, ws, ws_up = tree_cpu.refine_forward(feat, ew, si, sp, sc)
    assert ew[0, 0] == 0  # in-place write, refine.cu:43-46
    S = path_product_matrix(tree[0].numpy(), w_vertex, V)
    f = feat[0].double().numpy()
    np.testing.assert_allclose(aggr[0].numpy(), f @ S.T, rtol=2e-5)
    np.testing.assert_allclose(ws[0].numpy(), S.sum(1), rtol=2e-5)
    np.testing.assert_allclose(out[0].numpy(), (f @ S.T) / S.sum(1), rtol=2e-5)


def test_refine_backward_matches_fp64_autograd_of_the_path_product_filter():
    H, W, C = 4, 5, 2
    V = H * W
    g = torch.Generator().manual_seed(5)
    emb = torch.rand(1, 2, H, W, generator=g)
    feat = torch.rand(1, C, V, generator=g)
    gout = torch.randn(1, C, V, generator=g)
    tree = losses_cpu.minimum_spanning_tree(emb)
    si, sp, sc = tree_cpu.bfs_forward(tree, 4)
    ew = tree_cpu.edge_weight(emb.reshape(1, 2, V), si, sp, 1.0)
    saved = tree_cpu.refine_forward(feat, ew, si, sp, sc)
    gf = tree_cpu.refine_backward_feature(feat, ew, si, sp, sc, *saved, gout)
    gw = tree_cpu.refine_backward_weight(feat, ew, si, sp, sc, *saved, gout)

    # fp64 autograd through an explicit path-product construction
    parent, depth = parents_and_depth(tree[0].numpy(), V)
    order = si[0].long()
    w64 = ew[0].double().clone().requires_grad_()
    f64 = feat[0].double().clone().requires_grad_()
    w_vertex = torch.zeros(V, dtype=torch.float64).index_put((order,), w64)
    rows = []
    for i in range(V):
        row = []
        for j in range(V):
            a, b, prod = i, j, torch.ones((), dtype=torch.float64)
            while a != b:
                if depth[a] >= depth[b]:
                    prod = prod * w_vertex[a]
                    a = parent[a]
                else:
                    prod = prod * w_vertex[b]
                    b = parent[b]
            row.append(prod)
        rows.append(torch.stack(row))
    S = torch.stack(rows)
    out = (f64 @ S.T) / S.sum(1)
    (out * gout[0].double()).sum().backward()
    np.testing.assert_allclose(gf[0].numpy(), f64.grad.numpy(), rtol=1e-4, atol=1e-6)
    ref_gw = w64.grad.numpy().copy()
    ref_gw[0] = 0
    np.testing.assert_allclose(gw[0].numpy(), ref_gw, rtol=1e-4, atol=1e-6)


def test_edge_weight_backward_matches_autograd():
    H, W, D = 6, 5, 3
    V = H * W
    g = torch.Generator().manual_seed(2)
    emb = torch.rand(2, D, V, generator=g)
    tree = losses_cpu.minimum_spanning_tree(emb.reshape(2, D, H, W))
    si, sp, _ = tree_cpu.bfs_forward(tree, 4)
    gew = torch.randn(2, V, generator=g)
    e = emb.clone().requires_grad_()
    src = torch.gather(e, 2, si.long().unsqueeze(1).expand(-1, D, -1))
    dst = torch.gather(src, 2, sp.long().unsqueeze(1).expand(-1, D, -1))
    ew_t = torch.exp(-((src - dst) ** 2).sum(1) / 0.5)
    (ew_t * gew).sum().backward()
    ew = tree_cpu.edge_weight(emb, si, sp, 2.0)
    np.testing.assert_allclose(ew.numpy(), ew_t.detach().numpy(), rtol=1e-6)
    ge = tree_cpu.edge_weight_bwd(emb, si, sp, 2.0, ew, gew)
    np.testing.assert_allclose(ge.numpy(), e.grad.numpy(), rtol=1e-5, atol=1e-7)
