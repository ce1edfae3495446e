"""Chain order (heavy-path decomposition, csrc/euler.cu) and the scan-based tree filter on it (csrc/chains.cu), through
the C ABI, against a host restatement of the decomposition and against the sequential oracle (which accepts any valid
order, like the reference's kernels)."""
import numpy as np
import pytest
import torch

from oracle import losses_cpu, tree_cpu
from tests.helpers import make_image, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-4


def host_chain_order(tree, V):
    """Heavy-path order of one spanning tree [(V-1),2] rooted at 0: heavy child = largest subtree, smallest vertex id on
    ties; positions sorted by (light depth, preorder that visits the heavy child first, then light children by id)."""
    adj = [[] for _ in range(V)]
    for a, b in np.asarray(tree).reshape(-1, 2):
        adj[a].append(b); adj[b].append(a)
    parent = np.full(V, -1, np.int64); parent[0] = 0
    order = [0]
    for u in order:
        for v in sorted(adj[u]):
            if parent[v] < 0 and v != 0:
                parent[v] = u; order.append(v)
    size = np.ones(V, np.int64)
    for u in reversed(order[1:]):
        size[parent[u]] += size[u]
    kids = [[] for _ in range(V)]
    for u in order[1:]:
        kids[parent[u]].append(u)
    heavy = np.full(V, -1, np.int64)
    for u in range(V):
        best = 0
        for c in sorted(kids[u]):
            if size[c] > best:
                best = size[c]; heavy[u] = c
    ld = np.zeros(V, np.int64); pre = np.zeros(V, np.int64)
    stack = [0]; k = 0
    while stack:
        u = stack.pop()
        pre[u] = k; k += 1
        lights = [c for c in sorted(kids[u]) if c != heavy[u]]
        for c in reversed(lights):
            ld[c] = ld[u] + 1; stack.append(c)
        if heavy[u] >= 0:
            ld[heavy[u]] = ld[u]; stack.append(heavy[u])
    idx = np.lexsort((pre, ld))               # position -> vertex
    pos = np.empty(V, np.int64); pos[idx] = np.arange(V)
    sp = pos[parent[idx]]
    sc = np.zeros((V, 4), np.int64)
    for i, u in enumerate(idx):
        if heavy[u] >= 0:
            sc[i, 0] = pos[heavy[u]]
        for s, c in enumerate([c for c in sorted(kids[u]) if c != heavy[u]]):
            sc[i, 1 + s] = pos[c]
    gs = np.full(33, V, np.int64)
    counts = np.bincount(ld, minlength=1)
    gs[:len(counts)] = np.concatenate([[0], np.cumsum(counts)[:-1]])
    return idx, sp, sc, gs


@pytest.mark.parametrize("kind,B,D,H,W", [("quant8", 3, 3, 33, 47), ("constant", 1, 1, 40, 24), ("uniform", 2, 2, 1, 17),
                                          ("uniform", 2, 2, 19, 1), ("uniform", 1, 1, 1, 1), ("uniform", 1, 1, 1, 2),
                                          ("quant8", 2, 3, 128, 96), ("quant8", 2, 3, 256, 256)])
def test_chain_order_equals_host_decomposition(kind, B, D, H, W):
    from fedicra_b200 import ops
    emb = make_image(kind, B, D, H, W, torch.Generator().manual_seed(2))
    V = H * W
    tree = losses_cpu.minimum_spanning_tree(emb) if V > 1 else torch.zeros(B, 0, 2, dtype=torch.int32)
    _, mask = ops.mst_grid(emb.cuda(), want_mask=True)
    si, sp, sc, gs = [t.cpu().numpy() for t in ops.chain_order(mask, H, W)]
    for b in range(B):
        hi, hp, hc, hg = host_chain_order(tree[b].numpy(), V)
        assert np.array_equal(si[b], hi) and np.array_equal(sp[b], hp), b
        assert np.array_equal(sc[b], hc) and np.array_equal(gs[b], hg), b
        heavy = sc[b][:, 0]
        assert np.all((heavy == 0) | (heavy == np.arange(V) + 1))        # the heavy child is the next position


def _case(kind, B, C, D, H, W, seed=5):
    g = torch.Generator().manual_seed(seed)
    emb = make_image(kind, B, D, H, W, g)
    feat = torch.rand(B, C, H * W, generator=g)
    gout = torch.randn(B, C, H * W, generator=g)
    return emb, feat, gout


@pytest.mark.parametrize("kind,B,C,D,H,W,inv_sigma", [
    ("uniform", 2, 2, 3, 24, 31, 1.0), ("quant8", 1, 3, 3, 40, 40, 50.0), ("constant", 1, 2, 1, 9, 7, 1.0),
    ("uniform", 1, 1, 2, 1, 13, 1.0), ("uniform", 2, 2, 2, 1, 700, 0.05), ("constant", 2, 3, 1, 300, 2, 1.0),
    ("uniform", 1, 2, 1, 1, 1, 1.0), ("uniform", 1, 2, 1, 1, 2, 1.0), ("uniform", 2, 4, 3, 64, 64, 1.0),
    ("quant8", 12, 2, 3, 256, 256, 50.0), ("uniform", 3, 3, 3, 256, 256, 0.3), ("quant8", 2, 3, 3, 512, 512, 1.0),
])
def test_chain_filter_vs_oracle(kind, B, C, D, H, W, inv_sigma):
    """fedicra_chain_filter_forward/backward against the sequential oracle run on the SAME chain-ordered arrays (the
    oracle, like refine.cu, follows parent / child links of whatever valid order it is given), including paths
    (one chain), combs and single vertices."""
    from fedicra_b200 import ops
    emb, feat, gout = _case(kind, B, C, D, H, W)
    V = H * W
    _, mask = ops.mst_grid(emb.cuda(), want_mask=True)
    si, sp, sc, gs = ops.chain_order(mask, H, W)
    hsi, hsp, hsc = si.cpu(), sp.cpu(), sc.cpu()
    ew = tree_cpu.edge_weight(emb.reshape(B, D, V), hsi, hsp, inv_sigma)
    saved = tree_cpu.refine_forward(feat, ew, hsi, hsp, hsc)
    gf = tree_cpu.refine_backward_feature(feat, ew, hsi, hsp, hsc, *saved, gout)
    gw = tree_cpu.refine_backward_weight(feat, ew, hsi, hsp, hsc, *saved, gout)

    cew = ops.edge_weight_fwd(emb.reshape(B, D, V).cuda(), si, sp, inv_sigma)
    out, sv = ops.chain_filter_forward(feat.cuda(), cew, si, sp, sc, gs)
    torch.cuda.synchronize()
    assert ops.chain_watchdog_trips(sv) == 0, "a wait timed out"
    assert float(cew[:, 0].abs().max()) == 0.0
    assert rel_err(out.cpu(), saved[0]) < TOL
    up, agg = [t.cpu() for t in ops.chain_saved_arrays(sv, B, C, V)]
    assert rel_err(up[:, :C], saved[2]) < TOL and rel_err(up[:, C], saved[4]) < TOL      # feature_aggr_up, weight_sum_up
    to_vertex = lambda x: torch.zeros_like(x).scatter_(-1, hsi.long().unsqueeze(1).expand_as(x), x)
    assert rel_err(to_vertex(agg[:, :C]), saved[1]) < TOL                                 # feature_aggr
    assert rel_err(to_vertex(agg[:, C:])[:, 0], saved[3]) < TOL                           # weight_sum
    cgf, cgw = ops.chain_filter_backward(gout.cuda(), cew, si, sp, sc, gs, sv, True)
    torch.cuda.synchronize()
    ws = [v for k, v in ops._WORKSPACES.items() if k[0] == "chain_filter"][0]
    assert ops.chain_watchdog_trips(ws) == 0, "a wait timed out"
    assert rel_err(cgf.cpu(), gf) < TOL
    assert rel_err(cgw.cpu(), gw) < TOL
    cgf2, none = ops.chain_filter_backward(gout.cuda(), cew, si, sp, sc, gs, sv, False)
    assert none is None and rel_err(cgf2.cpu(), gf) < TOL


@pytest.mark.parametrize("kind,B,C,D,H,W", [("quant8", 3, 2, 3, 96, 80), ("uniform", 2, 3, 2, 33, 47), ("uniform", 1, 1, 1, 1, 1)])
def test_armed_buffers_give_the_same_bits(kind, B, C, D, H, W):
    """fedicra_chain_filter_arm + _forward_armed / _backward_armed (one memset for several launches) against the calls
    that arm their own buffer: the same bits; an armed buffer serves one launch, a used one must not be passed again
    (here: re-armed by hand and used again)."""
    from fedicra_b200 import ops
    emb, feat, gout = _case(kind, B, C, D, H, W)
    V = H * W
    _, mask = ops.mst_grid(emb.cuda(), want_mask=True)
    si, sp, sc, gs = ops.chain_order(mask, H, W)
    cew = ops.edge_weight_fwd(emb.reshape(B, D, V).cuda(), si, sp, 1.0)
    out, sv = ops.chain_filter_forward(feat.cuda(), cew, si, sp, sc, gs)
    gf, gw = ops.chain_filter_backward(gout.cuda(), cew, si, sp, sc, gs, sv, True)
    bufs = ops.chain_arm(B, C, V, 3, feat.cuda().device)
    assert all(b.data_ptr() % 256 == 0 for b in bufs)
    out2, sv2 = ops.chain_filter_forward(feat.cuda(), cew, si, sp, sc, gs, saved=bufs[0])
    assert sv2.data_ptr() == bufs[0].data_ptr()
    gf2, gw2 = ops.chain_filter_backward(gout.cuda(), cew, si, sp, sc, gs, sv2, True, armed_workspace=bufs[1])
    torch.cuda.synchronize()
    assert ops.chain_watchdog_trips(sv2) == 0 and ops.chain_watchdog_trips(bufs[1]) == 0
    assert torch.equal(out, out2) and torch.equal(gf, gf2) and torch.equal(gw, gw2)
    assert torch.equal(ops.chain_saved_arrays(sv, B, C, V), ops.chain_saved_arrays(sv2, B, C, V))
    bufs[1].fill_(255)                     # what fedicra_chain_filter_arm does
    gf3, gw3 = ops.chain_filter_backward(gout.cuda(), cew, si, sp, sc, gs, sv2, True, armed_workspace=bufs[1])
    assert torch.equal(gf, gf3) and torch.equal(gw, gw3)


def test_chain_and_band_filters_agree_at_full_size():
    """Two independent implementations (depth bands on the BFS order, affine scans on the chain order) of the same filter."""
    from fedicra_b200 import ops
    B, C, H, W = 12, 2, 256, 256
    emb, feat, gout = _case("quant8", B, C, 3, H, W, seed=9)
    V = H * W
    emb, feat, gout = emb.cuda(), feat.cuda(), gout.cuda()
    _, mask = ops.mst_grid(emb, want_mask=True)
    order = ops.bfs_euler(mask, H, W)
    bew = ops.edge_weight_fwd(emb.reshape(B, 3, V), order[0], order[1], 1.0)
    bands = ops.band_plan(order[3], order[4])
    bout, bsv = ops.tree_filter_forward(feat, bew, *order, bands)
    bgf, bgw = ops.tree_filter_backward(gout, bew, *order, bands, bsv, True)
    si, sp, sc, gs = ops.chain_order(mask, H, W)
    cew = ops.edge_weight_fwd(emb.reshape(B, 3, V), si, sp, 1.0)
    cout, csv = ops.chain_filter_forward(feat, cew, si, sp, sc, gs)
    cgf, cgw = ops.chain_filter_backward(gout, cew, si, sp, sc, gs, csv, True)
    assert rel_err(cout.cpu(), bout.cpu()) < 2e-5 and rel_err(cgf.cpu(), bgf.cpu()) < 2e-5
    by_vertex = lambda g, idx: torch.zeros(B, V, device="cuda").scatter_(1, idx.long(), g)
    assert rel_err(by_vertex(cgw, si).cpu(), by_vertex(bgw, order[0]).cpu()) < TOL
