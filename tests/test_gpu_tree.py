"""MST / BFS / edge weights / refine kernels (through the C ABI) against the CPU oracle and the golden fixtures."""
import os

import numpy as np
import pytest
import torch

from oracle import losses_cpu, tree_cpu
from tests.helpers import edge_set, make_image, parents_and_depth, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-4


def gpu_tree(emb):
    from fedicra_b200 import ops
    return ops.mst_grid(emb.cuda())


@pytest.mark.parametrize("name", ["mst_real_faz.npz", "mst_quant8.npz", "mst_constant.npz", "mst_uniform.npz"])
def test_mst_golden_edge_sets_bit_exact(golden_dir, name):
    z = np.load(os.path.join(golden_dir, name))
    emb = torch.from_numpy(z["embed"])
    tree = gpu_tree(emb).cpu().numpy()
    for b in range(emb.shape[0]):
        assert np.array_equal(edge_set(tree[b]), edge_set(z["tree"][b]))


@pytest.mark.parametrize("kind,B,D,H,W", [
    ("quant8", 3, 3, 33, 47), ("uniform", 2, 2, 64, 64), ("constant", 1, 1, 40, 24), ("quant8", 2, 1, 1, 50),
    ("quant8", 2, 1, 50, 1), ("uniform", 1, 3, 2, 2), ("quant8", 12, 3, 256, 256), ("quant8", 1, 3, 512, 512),
])
def test_mst_grid_and_explicit_equal_oracle(kind, B, D, H, W):
    from fedicra_b200 import tree_filter_cuda
    g = torch.Generator().manual_seed(H * 7 + W)
    emb = make_image(kind, B, D, H, W, g)
    ref = losses_cpu.minimum_spanning_tree(emb).numpy()
    got = gpu_tree(emb).cpu().numpy()
    idx = losses_cpu.grid_edges(H, W).unsqueeze(0).expand(B, -1, -1).contiguous()
    wgt = losses_cpu.grid_weights(emb).contiguous()
    got2 = tree_filter_cuda.mst_forward(idx.cuda(), wgt.cuda(), H * W).cpu().numpy()
    V = H * W
    for b in range(B):
        want = edge_set(ref[b])
        assert np.array_equal(edge_set(got[b]), want)       # fused grid weights + MST
        assert np.array_equal(edge_set(got2[b]), want)      # reference-style explicit edge list
        if V <= 4096:                                        # parent-of-vertex and depth arrays, rooted at 0
            p_ref, d_ref = parents_and_depth(ref[b], V)
            p_got, d_got = parents_and_depth(got[b], V)
            assert np.array_equal(p_ref, p_got) and np.array_equal(d_ref, d_got)


@pytest.mark.parametrize("kind,B,H,W", [("quant8", 3, 33, 47), ("constant", 1, 40, 24), ("uniform", 2, 1, 17),
                                        ("quant8", 2, 256, 256)])
def test_bfs_arrays_bit_exact(kind, B, H, W):
    from fedicra_b200 import ops
    emb = make_image(kind, B, 3, H, W, torch.Generator().manual_seed(1))
    tree = losses_cpu.minimum_spanning_tree(emb)
    si, sp, sc = tree_cpu.bfs_forward(tree, 4)
    gsi, gsp, gsc, gls, gnl = ops.bfs_forward(tree.cuda())
    assert torch.equal(gsi.cpu(), si) and torch.equal(gsp.cpu(), sp) and torch.equal(gsc.cpu(), sc)
    for b in range(B):
        _, _, _, ls = tree_cpu.bfs_levels(tree[b])
        n = int(gnl[b])
        assert n == len(ls) - 1 and torch.equal(gls[b, :n + 1].cpu(), ls)


def test_bfs_single_vertex():
    from fedicra_b200 import ops
    si, sp, sc, ls, nl = ops.bfs_forward(torch.zeros(2, 0, 2, dtype=torch.int32, device="cuda"))
    assert si.tolist() == [[0], [0]] and nl.tolist() == [1, 1] and ls.tolist() == [[0, 1], [0, 1]]


def _tree_case(kind, B, C, D, H, W, seed=3):
    g = torch.Generator().manual_seed(seed)
    emb = make_image(kind, B, D, H, W, g)
    feat = torch.rand(B, C, H * W, generator=g)
    gout = torch.randn(B, C, H * W, generator=g)
    tree = losses_cpu.minimum_spanning_tree(emb)
    return emb, feat, gout, tree


@pytest.mark.parametrize("kind,B,C,D,H,W,inv_sigma", [
    ("uniform", 2, 2, 3, 24, 31, 1.0), ("quant8", 1, 3, 3, 40, 40, 50.0), ("constant", 1, 2, 1, 9, 7, 1.0),
    ("uniform", 1, 1, 2, 1, 13, 1.0), ("uniform", 12, 2, 3, 256, 256, 50.0), ("quant8", 2, 3, 3, 512, 512, 1.0),
])
def test_reference_operator_surface_vs_oracle(kind, B, C, D, H, W, inv_sigma):
    """tree_filter_cuda.{bfs_forward,refine_forward,refine_backward_feature,refine_backward_weight} + edge weights."""
    from fedicra_b200 import ops, tree_filter_cuda as tfc
    emb, feat, gout, tree = _tree_case(kind, B, C, D, H, W)
    V = H * W
    si, sp, sc = tree_cpu.bfs_forward(tree, 4)
    ew = tree_cpu.edge_weight(emb.reshape(B, D, V), si, sp, inv_sigma)
    saved = tree_cpu.refine_forward(feat, ew, si, sp, sc)
    gf = tree_cpu.refine_backward_feature(feat, ew, si, sp, sc, *saved, gout)
    gw = tree_cpu.refine_backward_weight(feat, ew, si, sp, sc, *saved, gout)
    ge = tree_cpu.edge_weight_bwd(emb.reshape(B, D, V), si, sp, inv_sigma, ew, gw)

    csi, csp, csc = tfc.bfs_forward(tree.cuda(), 4)
    order = ops.bfs_forward(tree.cuda())
    cew = ops.edge_weight_fwd(emb.reshape(B, D, V).cuda(), csi, csp, inv_sigma)
    assert rel_err(cew.cpu()[:, 1:], tree_cpu.edge_weight(emb.reshape(B, D, V), si, sp, inv_sigma)[:, 1:]) < 1e-5
    cf = feat.cuda()
    csaved = tfc.refine_forward(cf, cew, csi, csp, csc)
    assert float(cew[:, 0].abs().max()) == 0.0  # refine.cu:43-46 side effect
    names = ("feature_out", "feature_aggr", "feature_aggr_up", "weight_sum", "weight_sum_up")
    for n, a, b in zip(names, csaved, saved):
        assert rel_err(a.cpu(), b) < TOL, n
    cgf = tfc.refine_backward_feature(cf, cew, csi, csp, csc, *csaved, gout.cuda())
    cgw = tfc.refine_backward_weight(cf, cew, csi, csp, csc, *csaved, gout.cuda())
    assert rel_err(cgf.cpu(), gf) < TOL
    assert rel_err(cgw.cpu(), gw) < TOL
    cge = ops.edge_weight_bwd(emb.reshape(B, D, V).cuda(), csi, csp, csc, cew, gw.cuda(), inv_sigma)
    assert rel_err(cge.cpu(), ge) < TOL


def test_filter_properties_at_512():
    """No oracle needed: a constant feature is a fixed point of the normalised filter, weight_sum >= 1, and the feature
    gradient is the adjoint of the (symmetric) un-normalised filter: <aggr(f), g> == <f, aggr(g)>."""
    from fedicra_b200 import ops
    B, C, H, W = 2, 3, 512, 512
    g = torch.Generator().manual_seed(9)
    emb = ((torch.rand(B, 3, H, W, generator=g) * 255).round() / 255).cuda()
    tree = ops.mst_grid(emb)
    assert tree.shape == (B, H * W - 1, 2) and int(tree.min()) >= 0
    si, sp, sc, ls, nl = ops.bfs_forward(tree)
    assert torch.equal(torch.sort(si, 1).values, torch.arange(H * W, device="cuda", dtype=torch.int32).expand(B, -1))
    ew = ops.edge_weight_fwd(emb.reshape(B, 3, -1), si, sp, 1.0)
    ones = torch.full((B, C, H * W), 0.37, device="cuda")
    out, aggr, _, wsum, _ = ops.refine_forward(ones, ew, si, sp, sc, ls, nl)
    assert rel_err(out.cpu(), ones.cpu()) < 1e-5 and float(wsum.min()) >= 1.0
    f = torch.rand(B, C, H * W, generator=g).cuda()
    h = torch.rand(B, C, H * W, generator=g).cuda()
    _, af, _, _, _ = ops.refine_forward(f, ew, si, sp, sc, ls, nl)
    _, ah, _, _, _ = ops.refine_forward(h, ew, si, sp, sc, ls, nl)
    lhs, rhs = (af.double() * h.double()).sum().item(), (f.double() * ah.double()).sum().item()
    assert abs(lhs - rhs) < 1e-5 * abs(lhs)


@pytest.mark.parametrize("kind,B,C,D,H,W,inv_sigma", [
    ("uniform", 2, 2, 3, 24, 31, 1.0), ("quant8", 1, 3, 3, 40, 40, 50.0), ("constant", 1, 2, 1, 9, 7, 1.0),
    ("uniform", 1, 1, 2, 1, 13, 1.0), ("uniform", 2, 2, 2, 1, 700, 0.05), ("constant", 2, 3, 1, 300, 2, 1.0),
    ("uniform", 1, 2, 1, 1, 1, 1.0), ("uniform", 1, 2, 1, 1, 2, 1.0), ("quant8", 12, 2, 3, 256, 256, 50.0),
    ("uniform", 3, 3, 3, 256, 256, 0.3), ("quant8", 2, 3, 3, 512, 512, 1.0),
])
@pytest.mark.parametrize("impl", ["resident", "multi", "resident-global"])
def test_depth_banded_filter_vs_oracle(kind, B, C, D, H, W, inv_sigma, impl, monkeypatch):
    """fedicra_tree_filter_forward/backward (csrc/bands.cu, csrc/resident.cuh) against the sequential oracle, including
    degenerate trees: paths (depth = V), combs, single vertices.  `resident` = one cooperative launch per filter,
    `multi` = the phase-per-kernel version, `resident-global` = the resident kernel with a shared-memory budget so
    small that every band and the stitcher take their global-memory code."""
    from fedicra_b200 import ops
    monkeypatch.setenv("FEDICRA_FILTER", "multi" if impl == "multi" else "resident")
    if impl == "resident-global":
        if H * W > 70000:
            pytest.skip("global-memory mode is exercised at the smaller sizes")
        monkeypatch.setenv("FEDICRA_PLAN_PMAX", "48")
    emb, feat, gout, tree = _tree_case(kind, B, C, D, H, W, seed=5)
    V = H * W
    if V == 1:
        tree = torch.zeros(B, 0, 2, dtype=torch.int32)
    si, sp, sc = tree_cpu.bfs_forward(tree, 4)
    ew = tree_cpu.edge_weight(emb.reshape(B, D, V), si, sp, inv_sigma)
    saved = tree_cpu.refine_forward(feat, ew, si, sp, sc)
    gf = tree_cpu.refine_backward_feature(feat, ew, si, sp, sc, *saved, gout)
    gw = tree_cpu.refine_backward_weight(feat, ew, si, sp, sc, *saved, gout)

    order = ops.bfs_forward(tree.cuda())
    cew = ops.edge_weight_fwd(emb.reshape(B, D, V).cuda(), order[0], order[1], inv_sigma)
    bands = ops.band_plan(order[3], order[4])
    plan = bands.view(B, -1).cpu()
    for b in range(B):                                  # the plan partitions the depths into non-empty bands
        nb = int(plan[b, 0]); cuts = plan[b, 1:nb + 2]
        assert cuts[0] == 0 and cuts[-1] == int(order[4][b]) and bool((cuts[1:] > cuts[:-1]).all())
    out, sv, aggr, wsum = ops.tree_filter_forward(feat.cuda(), cew, *order, bands, want_aggr=True)
    assert float(cew[:, 0].abs().max()) == 0.0
    assert rel_err(out.cpu(), saved[0]) < TOL
    assert rel_err(aggr.cpu(), saved[1]) < TOL
    assert rel_err(wsum.cpu(), saved[3]) < TOL
    cgf, cgw = ops.tree_filter_backward(gout.cuda(), cew, *order, bands, sv, True)
    assert rel_err(cgf.cpu(), gf) < TOL
    assert rel_err(cgw.cpu(), gw) < TOL
    cgf2, none = ops.tree_filter_backward(gout.cuda(), cew, *order, bands, sv, False)
    assert none is None and rel_err(cgf2.cpu(), gf) < TOL

@pytest.mark.parametrize("finish,pool_kb", [("contract", None), ("contract", 24), ("contract", 1), ("cluster", None)])
@pytest.mark.parametrize("kind,B,D,H,W", [("quant8", 4, 3, 256, 256), ("uniform", 2, 2, 130, 200), ("constant", 1, 1, 70, 65),
                                          ("quant8", 1, 3, 512, 512)])
def test_mst_finish_paths_bit_exact(kind, B, D, H, W, finish, pool_kb, monkeypatch):
    """Rounds after the tile kernel: on the contracted graph in shared memory (default), with the records streamed from L2
    (component arrays in shared memory: 24 KB hold ~1 500 components, not the records), wholly on the workspace (1 KB),
    and round 1's cluster kernel -- same edge set as the reference's Boruvka (oracle) and the same adjacency mask."""
    from fedicra_b200 import ops
    monkeypatch.setenv("FEDICRA_MST_FINISH", finish)
    if pool_kb is not None:
        monkeypatch.setenv("FEDICRA_MST_POOL_KB", str(pool_kb))
    emb = make_image(kind, B, D, H, W, torch.Generator().manual_seed(H + W))
    ref = losses_cpu.minimum_spanning_tree(emb).numpy()
    tree, mask = ops.mst_grid(emb.cuda(), want_mask=True)
    tree, mask = tree.cpu().numpy(), mask.cpu().long()
    only_mask = ops.mst_grid_multi([emb.cuda()], want_mask=True, want_edges=False)[1].cpu().long()
    assert torch.equal(mask, only_mask)
    V = H * W
    for b in range(B):
        assert np.array_equal(edge_set(tree[b]), edge_set(ref[b]))
        deg = torch.bincount(torch.from_numpy(ref[b]).reshape(-1).long(), minlength=V)
        assert torch.equal(sum(((mask[b] >> k) & 1) for k in range(4)), deg)


@pytest.mark.parametrize("kind,B,D,H,W", [("quant8", 3, 3, 33, 47), ("constant", 1, 1, 40, 24), ("uniform", 2, 2, 1, 17),
                                          ("uniform", 2, 2, 19, 1), ("uniform", 1, 1, 1, 1), ("quant8", 12, 3, 256, 256),
                                          ("constant", 1, 1, 300, 300), ("quant8", 2, 3, 512, 512)])
def test_bfs_from_adjacency_mask_bit_exact(kind, B, D, H, W):
    """fedicra_mst_grid's 4-bit mask + fedicra_bfs_grid (warp walk in shared memory) == oracle BFS of the oracle MST."""
    from fedicra_b200 import ops
    emb = make_image(kind, B, D, H, W, torch.Generator().manual_seed(2))
    V = H * W
    tree = losses_cpu.minimum_spanning_tree(emb) if V > 1 else torch.zeros(B, 0, 2, dtype=torch.int32)
    si, sp, sc = tree_cpu.bfs_forward(tree, 4)
    gtree, mask = ops.mst_grid(emb.cuda(), want_mask=True)
    deg = torch.zeros(B, V, dtype=torch.int64)
    for b in range(B):
        deg[b] = torch.bincount(tree[b].reshape(-1).long(), minlength=V)
    pop = sum(((mask.cpu().long() >> k) & 1) for k in range(4))
    assert torch.equal(pop, deg)
    for fn in (ops.bfs_grid, ops.bfs_euler):   # level-by-level walk in shared memory; Euler tour + prefix sums
        gsi, gsp, gsc, gls, gnl = fn(mask, H, W)
        assert torch.equal(gsi.cpu(), si) and torch.equal(gsp.cpu(), sp) and torch.equal(gsc.cpu(), sc), fn.__name__
        for b in range(B):
            _, _, _, ls = tree_cpu.bfs_levels(tree[b])
            n = int(gnl[b])
            assert n == len(ls) - 1 and torch.equal(gls[b, :n + 1].cpu(), ls), fn.__name__


@pytest.mark.parametrize("kind,B,C,H,W", [("quant8", 2, 2, 96, 80), ("uniform", 1, 3, 33, 47), ("constant", 1, 1, 40, 24)])
def test_rooted_order_is_a_level_order_and_the_filter_does_not_depend_on_the_root(kind, B, C, H, W):
    """fedicra_bfs_euler_rooted: started from the centre pixel the arrays are still a breadth-first order (levels,
    parents one level up, children contiguous and increasing), and the filter -- forward, feature gradient, edge-weight
    gradient mapped back to the embedding -- equals the one on the reference's vertex-0 order."""
    from fedicra_b200 import ops
    g = torch.Generator().manual_seed(11)
    emb = make_image(kind, B, 3, H, W, g).cuda()
    feat = torch.rand(B, C, H * W, generator=g).cuda()
    gout = torch.randn(B, C, H * W, generator=g).cuda()
    V = H * W
    _, mask = ops.mst_grid(emb, want_mask=True)
    results = []
    for root in (0, ops.centre_root(H, W)):
        si, sp, sc, ls, nl = ops.bfs_euler(mask, H, W, root)
        for b in range(B):
            s, p, c = si[b].cpu().long(), sp[b].cpu().long(), sc[b].cpu().long()
            n = int(nl[b]); starts = ls[b, :n + 1].cpu().long()
            assert int(s[0]) == root and int(p[0]) == 0 and sorted(s.tolist()) == list(range(V))
            level = torch.bucketize(torch.arange(V), starts[1:], right=True)
            assert int(starts[0]) == 0 and int(starts[n]) == V and int(starts[1]) == 1
            assert bool((level[p[1:]] == level[1:] - 1).all())                     # parents sit one level up
            kids = c[c > 0]
            assert sorted(kids.tolist()) == list(range(1, V))                      # every non-root is someone's child once
            for k in range(4):
                has = c[:, k] > 0
                assert bool((p[c[has, k]] == torch.arange(V)[has]).all())
            two = (c[:, 0] > 0) & (c[:, 1] > 0)
            assert bool((c[two, 1] == c[two, 0] + 1).all())                        # contiguous, increasing
        ew = ops.edge_weight_fwd(emb.reshape(B, 3, V), si, sp, 1.0)
        ew0 = ew.clone()
        bands = ops.band_plan(ls, nl)
        out, saved = ops.tree_filter_forward(feat, ew, si, sp, sc, ls, nl, bands)
        gf, gw = ops.tree_filter_backward(gout, ew, si, sp, sc, ls, nl, bands, saved, True)
        ge = ops.edge_weight_bwd(emb.reshape(B, 3, V), si, sp, sc, ew0, gw, 1.0)
        results.append((out, gf, ge))
    for a, b_ in zip(results[0], results[1]):
        assert rel_err(b_.cpu(), a.cpu()) < 2e-5


def test_corrupted_sync_words_fail_the_launch_instead_of_hanging():
    """The resident filter's CTAs wait on counters in the workspace head.  Garbage there (a workspace that was not
    zeroed) must end in a CUDA error within seconds -- the kernel traps once a wait outlasts FEDICRA_WATCHDOG_MS -- not in a hung GPU.
    Runs in a child process: the trap poisons the CUDA context."""
    import subprocess
    import sys
    if os.environ.get("FEDICRA_FILTER") == "multi":
        pytest.skip("the phase kernels have no inter-CTA waits")
    code = r"""
import sys, torch
sys.path.insert(0, %r)
from fedicra_b200 import ops
g = torch.Generator().manual_seed(0)
B, C, H, W = 2, 2, 64, 64
emb = torch.rand(B, 3, H, W, generator=g).cuda()
feat = torch.rand(B, C, H * W, generator=g).cuda()
_, mask = ops.mst_grid(emb, want_mask=True)
order = ops.bfs_euler(mask, H, W)
bands = ops.band_plan(order[3], order[4])
ew = ops.edge_weight_fwd(emb.reshape(B, 3, -1), order[0], order[1], 1.0)
ops.tree_filter_forward(feat, ew, *order, bands)          # creates (and zeroes) the cached workspace
torch.cuda.synchronize()
ws = [v for k, v in ops._WORKSPACES.items() if k[0] == "tree_filter"][0]
ws[:64].fill_(0x7f)                                       # corrupt the sync words of image 0 and 1
try:
    ops.tree_filter_forward(feat, ew, *order, bands)
    torch.cuda.synchronize()
except RuntimeError as exc:
    print("CUDA-ERROR", type(exc).__name__)
    sys.exit(0)
print("NO-ERROR")
""" % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, FEDICRA_WATCHDOG_MS="1500", FEDICRA_FILTER="resident")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120, env=env)
    assert "CUDA-ERROR" in r.stdout, (r.stdout[-500:], r.stderr[-500:])
