"""Thin tensor-level wrappers over the C ABI: argument checks, output allocation, workspace caching, stream hand-off.

PyTorch is only the allocator and the stream owner here; every function launches kernels of libfedicra_sm100.so.
"""
import torch

from ._lib import GcrfDesc, check, lib

_WORKSPACES = {}


def _stream():
    return torch.cuda.current_stream().cuda_stream


def workspace(tag, nbytes, device):
    """Zero-initialised scratch buffer, cached per (purpose, device, stream) and grown on demand."""
    key = (tag, device.index if device.index is not None else torch.cuda.current_device(), _stream())
    buf = _WORKSPACES.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.zeros(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        _WORKSPACES[key] = buf
    return buf


class _Here:
    """No-op context: the tensor's device is already the current one (the usual case; entering torch.cuda.device costs
    ~10 us of host time per call, and an eager step makes ~30 calls)."""

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


_HERE = _Here()


def _on(device):
    idx = device.index if isinstance(device, torch.device) else device
    return _HERE if (idx is None or idx == torch.cuda.current_device()) else torch.cuda.device(device)


def _cuda(t, dtype, name):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: fedicra_b200 has no CPU path")
    if t.dtype != dtype:
        raise RuntimeError(f"{name} must be {dtype}, got {t.dtype}")
    return t.contiguous()


def _p(t):
    return None if t is None else t.data_ptr()


# ------------------------------------------------------------------ Gated CRF


def gcrf_fwd_bwd(y, image, radius, descs, want_grad=True):
    """descs: list of (weight, sigma_xy or 0, sigma_rgb or 0).  Returns (loss[1], grad_y or None)."""
    y = _cuda(y, torch.float32, "y_hat_softmax")
    image = _cuda(image, torch.float32, "sample")
    N, C, H, W = y.shape
    assert image.shape[0] == N and image.shape[2:] == (H, W)
    with _on(y.device):
        loss = torch.empty(1, dtype=torch.float32, device=y.device)
        grad = torch.empty_like(y) if want_grad else None
        nbytes = lib.fedicra_gcrf_workspace_bytes(N, H, W)
        ws = workspace("gcrf", nbytes, y.device)
        arr = (GcrfDesc * len(descs))(*[GcrfDesc(float(w), float(sx), float(sr)) for (w, sx, sr) in descs])
        check(lib.fedicra_gcrf_fwd_bwd(_p(y), _p(image), N, C, image.shape[1], H, W, int(radius), arr, len(descs),
                                       _p(loss), _p(grad), _p(ws), ws.numel(), _stream()), "gcrf_fwd_bwd")
    return loss, grad


# ------------------------------------------------------------------ MST


def mst_forward(edge_index, edge_weight, vertex_count):
    ei = _cuda(edge_index, torch.int32, "edge_index")
    ew = _cuda(edge_weight, torch.float32, "edge_weight")
    B, E = ei.shape[0], ei.shape[1]
    V = int(vertex_count)
    with _on(ei.device):
        out = torch.empty(B, V - 1, 2, dtype=torch.int32, device=ei.device)
        ws = workspace("mst", lib.fedicra_mst_workspace_bytes(B, V, E), ei.device)
        check(lib.fedicra_mst_forward(_p(ei), _p(ew), B, E, V, _p(out), _p(ws), ws.numel(), _stream()), "mst_forward")
    return out


def mst_grid(embed, want_mask=False, want_edges=True):
    """MST of the 4-connected grid.  Returns the edge list [B,V-1,2] (and the 4-bit adjacency mask [B,V] when asked)."""
    e = _cuda(embed, torch.float32, "guide_in")
    B, D, H, W = e.shape
    V, E = H * W, 2 * H * W - H - W
    with _on(e.device):
        out = torch.empty(B, V - 1, 2, dtype=torch.int32, device=e.device) if want_edges else None
        mask = torch.empty(B, V, dtype=torch.uint8, device=e.device) if want_mask else None
        ws = workspace("mst", lib.fedicra_mst_workspace_bytes(B, V, E), e.device)
        check(lib.fedicra_mst_grid(_p(e), B, D, H, W, _p(out), _p(mask), _p(ws), ws.numel(), _stream()), "mst_grid")
    return (out, mask) if want_mask else out


def mst_grid_multi(embeds, want_mask=True, want_edges=False):
    """MST of the grids of several embeddings of one batch ([B,D_k,H,W], same B/H/W) in one launch, no stacked copy.
    Returns (edges [K*B,V-1,2] or None, mask [K*B,V] or None)."""
    import ctypes
    es = [_cuda(e, torch.float32, "guide_in") for e in embeds]
    B, _, H, W = es[0].shape
    K = len(es)
    assert all(e.shape[0] == B and tuple(e.shape[2:]) == (H, W) for e in es)
    V, E = H * W, 2 * H * W - H - W
    dev = es[0].device
    with _on(dev):
        out = torch.empty(K * B, V - 1, 2, dtype=torch.int32, device=dev) if want_edges else None
        mask = torch.empty(K * B, V, dtype=torch.uint8, device=dev) if want_mask else None
        ws = workspace("mst", lib.fedicra_mst_workspace_bytes(K * B, V, E), dev)
        ptrs = (ctypes.c_void_p * K)(*[e.data_ptr() for e in es])
        dims = (ctypes.c_int * K)(*[e.shape[1] for e in es])
        check(lib.fedicra_mst_grid_multi(ptrs, dims, K, B, H, W, _p(out), _p(mask), _p(ws), ws.numel(), _stream()),
              "mst_grid_multi")
    return out, mask


BFS_GRID_MAX_V = 2 * (227 * 1024 - 4 * 4096 * 4 - 256)
BFS_EULER_MAX_V = 1 << 18


def bfs_from_mask(mask, H, W, root=0):
    """Fastest BFS ordering available for an H x W grid tree given as adjacency mask (all three give identical arrays):
    Euler tour + prefix sums up to 262 144 pixels, else the shared-memory depth walk; None when neither applies.
    `root`: vertex the order starts from (Euler path only; the others always start from vertex 0)."""
    if H * W <= BFS_EULER_MAX_V:
        return bfs_euler(mask, H, W, root)
    if H * W <= BFS_GRID_MAX_V:
        return bfs_grid(mask, H, W)
    return None


def bfs_grid(mask, H, W):
    """BFS arrays from the adjacency mask of mst_grid (shared-memory walk); same outputs as bfs_forward."""
    m = _cuda(mask, torch.uint8, "tree_mask")
    B, V = m.shape
    dev = m.device
    with _on(dev):
        si = torch.empty(B, V, dtype=torch.int32, device=dev)
        sp = torch.empty(B, V, dtype=torch.int32, device=dev)
        sc = torch.empty(B, V, 4, dtype=torch.int32, device=dev)
        ls = torch.empty(B, V + 1, dtype=torch.int32, device=dev)
        nl = torch.empty(B, dtype=torch.int32, device=dev)
        check(lib.fedicra_bfs_grid(_p(m), B, H, W, _p(si), _p(sp), _p(sc), _p(ls), _p(nl), _stream()), "bfs_grid")
    return si, sp, sc, ls, nl


# ------------------------------------------------------------------ BFS


def bfs_forward(tree):
    """-> sorted_index [B,V], sorted_parent [B,V], sorted_child [B,V,4], level_start [B,V+1], n_levels [B]."""
    t = _cuda(tree, torch.int32, "tree")
    B, V = t.shape[0], t.shape[1] + 1
    dev = t.device
    with _on(dev):
        si = torch.zeros(B, V, dtype=torch.int32, device=dev)
        sp = torch.zeros(B, V, dtype=torch.int32, device=dev)
        sc = torch.zeros(B, V, 4, dtype=torch.int32, device=dev)
        ls = torch.empty(B, V + 1, dtype=torch.int32, device=dev)
        nl = torch.empty(B, dtype=torch.int32, device=dev)
        ws = workspace("bfs", lib.fedicra_bfs_workspace_bytes(B, V), dev)
        check(lib.fedicra_bfs_forward(_p(t), B, V, _p(si), _p(sp), _p(sc), _p(ls), _p(nl), _p(ws), ws.numel(),
                                      _stream()), "bfs_forward")
    return si, sp, sc, ls, nl


def centre_root(H, W):
    """The centre pixel as BFS root.  The filter's result does not depend on the root, the tree's shape does:
    centre-rooted trees are 15-40 % shallower than corner-rooted ones -- and as much wider."""
    return (H // 2) * W + W // 2 if H * W <= BFS_EULER_MAX_V else 0


def default_root(H, W):
    """Root TreeFilter2D orders its trees from.  Vertex 0, like the reference: measured on B200 (faz256) the centre root
    is a loss -- levels of ~80 instead of ~50 vertices need two passes of the 64 slots per channel and overflow the
    stitcher's shared memory, which costs more than the 30 % fewer levels save (FEDICRA_ROOT=centre to try it)."""
    import os
    return centre_root(H, W) if os.environ.get("FEDICRA_ROOT") == "centre" else 0


def bfs_euler(mask, H, W, root=0):
    """BFS arrays from the adjacency mask via Euler tour + prefix sums (no per-depth step); same outputs as bfs_grid
    (for root 0)."""
    m = _cuda(mask, torch.uint8, "tree_mask")
    B, V = m.shape
    dev = m.device
    with _on(dev):
        si = torch.empty(B, V, dtype=torch.int32, device=dev)
        sp = torch.empty(B, V, dtype=torch.int32, device=dev)
        sc = torch.empty(B, V, 4, dtype=torch.int32, device=dev)
        ls = torch.empty(B, V + 1, dtype=torch.int32, device=dev)
        nl = torch.empty(B, dtype=torch.int32, device=dev)
        ws = workspace("bfs_euler", lib.fedicra_bfs_euler_workspace_bytes(B, H, W), dev)
        check(lib.fedicra_bfs_euler_rooted(_p(m), B, H, W, int(root), _p(si), _p(sp), _p(sc), _p(ls), _p(nl), _p(ws),
                                           ws.numel(), _stream()), "bfs_euler")
    return si, sp, sc, ls, nl


def chain_order(mask, H, W):
    """Chain (heavy-path) order of grid trees given as adjacency masks: sorted_index, sorted_parent, sorted_child in the
    reference's sense plus group_start [B,33] (first position of every light depth).  Feeds chain_filter_*."""
    m = _cuda(mask, torch.uint8, "tree_mask")
    B, V = m.shape
    dev = m.device
    with _on(dev):
        si = torch.empty(B, V, dtype=torch.int32, device=dev)
        sp = torch.empty(B, V, dtype=torch.int32, device=dev)
        sc = torch.empty(B, V, 4, dtype=torch.int32, device=dev)
        gs = torch.empty(B, 33, dtype=torch.int32, device=dev)
        ws = workspace("bfs_euler", lib.fedicra_bfs_euler_workspace_bytes(B, H, W), dev)
        check(lib.fedicra_chain_order(_p(m), B, H, W, _p(si), _p(sp), _p(sc), _p(gs), _p(ws), ws.numel(), _stream()),
              "chain_order")
    return si, sp, sc, gs


def chain_order_weights(mask, embeds, inv_sigmas, H, W):
    """chain_order of the K*B trees of K embeddings ([B,D_k,H,W] each) plus, from the same pass, their filter edge
    weights [K*B,V] in chain order (inv_sigmas[k] = 1/sigma of embedding k's filter).  Returns (si, sp, sc, gs, ew)."""
    import ctypes
    m = _cuda(mask, torch.uint8, "tree_mask")
    es = [_cuda(e, torch.float32, "embed_in") for e in embeds]
    K, B = len(es), es[0].shape[0]
    KB, V = m.shape
    assert KB == K * B and V == H * W and all(e.shape[0] == B and tuple(e.shape[2:]) == (H, W) for e in es)
    dev = m.device
    with _on(dev):
        si = torch.empty(KB, V, dtype=torch.int32, device=dev)
        sp = torch.empty(KB, V, dtype=torch.int32, device=dev)
        sc = torch.empty(KB, V, 4, dtype=torch.int32, device=dev)
        gs = torch.empty(KB, 33, dtype=torch.int32, device=dev)
        ew = torch.empty(KB, V, dtype=torch.float32, device=dev)
        ws = workspace("bfs_euler", lib.fedicra_bfs_euler_workspace_bytes(KB, H, W), dev)
        ptrs = (ctypes.c_void_p * K)(*[e.data_ptr() for e in es])
        dims = (ctypes.c_int * K)(*[e.shape[1] for e in es])
        sig = (ctypes.c_float * K)(*[float(x) for x in inv_sigmas])
        check(lib.fedicra_chain_order_weights(_p(m), ptrs, dims, sig, K, B, H, W, _p(si), _p(sp), _p(sc), _p(gs), _p(ew),
                                              _p(ws), ws.numel(), _stream()), "chain_order_weights")
    return si, sp, sc, gs, ew


CHAIN_MAX_V = 1 << 18
CHAIN_MAX_C = lib.fedicra_chain_filter_max_channels()


def chain_arm(B, C, V, n, device):
    """n buffers for chain-filter launches (forward `saved` or backward workspace: same size), armed by ONE memset so
    that none sits between the filters of a step.  Each serves one launch (chain_filter_forward(..., saved=buf) /
    chain_filter_backward(..., armed_workspace=buf))."""
    size = (int(lib.fedicra_chain_filter_saved_bytes(B, C, V)) + 255) // 256 * 256
    with _on(device):
        block = torch.empty(n * size, dtype=torch.uint8, device=device)
        check(lib.fedicra_chain_filter_arm(_p(block), block.numel(), _stream()), "chain_filter_arm")
    return [block[i * size:(i + 1) * size] for i in range(n)]


def chain_filter_forward(feature_in, edge_weight, si, sp, sc, gs, saved=None):
    """Tree filter on chain-ordered trees.  Returns (feature_out [B,C,V] vertex order, saved); `saved` is an opaque
    byte buffer for chain_filter_backward, chain_saved_arrays(saved, B, C, V) views its two [B,C+1,V] arrays.
    `saved`: a buffer from chain_arm (not used by any launch yet)."""
    f = _cuda(feature_in, torch.float32, "feature_in")
    if not (edge_weight.is_cuda and edge_weight.dtype == torch.float32 and edge_weight.is_contiguous()):
        raise RuntimeError("edge_weight must be a contiguous float32 CUDA tensor (it is written in place)")
    B, C, V = f.shape
    dev = f.device
    with _on(dev):
        out = torch.empty_like(f)
        need = lib.fedicra_chain_filter_saved_bytes(B, C, V)
        if saved is not None and saved.numel() >= need and saved.device == dev:
            fn = lib.fedicra_chain_filter_forward_armed
        else:
            fn, saved = lib.fedicra_chain_filter_forward, torch.empty(need, dtype=torch.uint8, device=dev)
        check(fn(_p(f), _p(edge_weight), _p(si), _p(sp), _p(sc), _p(gs), B, C, V, _p(out), _p(saved), saved.numel(),
                 _stream()), "chain_filter_forward")
    return out, saved


def chain_saved_arrays(saved, B, C, V):
    """(leaf->root values, root->leaf aggregates), each [B,C+1,V] float32 in chain order (channel C = weight sums)."""
    off = lib.fedicra_chain_filter_saved_offset(B, C, V)
    return saved[off:off + 2 * B * (C + 1) * V * 4].view(torch.float32).view(2, B, C + 1, V)   # (an armed buffer is padded)


def chain_watchdog_trips(saved):
    """How many waits inside the last chain-filter launch that used this buffer ran into their time bound (0 = none)."""
    return int(saved[:8].view(torch.int32)[1]) + 1


def chain_filter_backward(grad_out, edge_weight, si, sp, sc, gs, saved, want_weight_grad, armed_workspace=None):
    """Returns (grad_feature [B,C,V] vertex order, grad_weight [B,V] chain order or None)."""
    g = _cuda(grad_out, torch.float32, "grad_output")
    B, C, V = g.shape
    dev = g.device
    with _on(dev):
        gf = torch.empty_like(g)
        gw = torch.empty(B, V, dtype=torch.float32, device=dev) if want_weight_grad else None
        need = lib.fedicra_chain_filter_workspace_bytes(B, C, V)
        if armed_workspace is not None and armed_workspace.numel() >= need and armed_workspace.device == dev:
            fn, ws = lib.fedicra_chain_filter_backward_armed, armed_workspace
        else:
            fn, ws = lib.fedicra_chain_filter_backward, workspace("chain_filter", need, dev)
        check(fn(_p(g), _p(edge_weight), _p(si), _p(sp), _p(sc), _p(gs), B, C, V, _p(saved), _p(gf), _p(gw), _p(ws),
                 ws.numel(), _stream()), "chain_filter_backward")
    return gf, gw


# ------------------------------------------------------------------ edge weights


def edge_weight_fwd(embed, si, sp, inv_sigma):
    e = _cuda(embed, torch.float32, "embed_in")
    B, D, V = e.shape
    with _on(e.device):
        ew = torch.empty(B, V, dtype=torch.float32, device=e.device)
        check(lib.fedicra_edge_weight_fwd(_p(e), _p(si), _p(sp), B, D, V, float(inv_sigma), _p(ew), _stream()),
              "edge_weight_fwd")
    return ew


def edge_weight_bwd(embed, si, sp, sc, ew, grad_ew, inv_sigma):
    e = _cuda(embed, torch.float32, "embed_in")
    g = _cuda(grad_ew, torch.float32, "grad_edge_weight")
    B, D, V = e.shape
    with _on(e.device):
        out = torch.empty_like(e)
        check(lib.fedicra_edge_weight_bwd(_p(e), _p(si), _p(sp), _p(sc), _p(ew), _p(g), B, D, V, float(inv_sigma),
                                          _p(out), _stream()), "edge_weight_bwd")
    return out


def edge_weight_bwd_multi(embeds, orders, ews, grad_ews, inv_sigmas):
    """edge_weight_bwd of K embeddings of one batch ([B,D_k,V]; orders[k] = (si, sp, sc)) in one launch.  Returns the
    K gradients."""
    import ctypes
    es = [_cuda(e, torch.float32, "embed_in") for e in embeds]
    gs = [_cuda(g, torch.float32, "grad_edge_weight") for g in grad_ews]
    K = len(es)
    B, _, V = es[0].shape
    assert all(e.shape[0] == B and e.shape[2] == V for e in es)
    with _on(es[0].device):
        outs = [torch.empty_like(e) for e in es]
        arr = lambda ts: (ctypes.c_void_p * K)(*[t.data_ptr() for t in ts])
        dims = (ctypes.c_int * K)(*[e.shape[1] for e in es])
        sig = (ctypes.c_float * K)(*[float(x) for x in inv_sigmas])
        check(lib.fedicra_edge_weight_bwd_multi(arr(es), arr([o[0] for o in orders]), arr([o[1] for o in orders]),
                                                arr([o[2] for o in orders]), arr(ews), arr(gs), dims, sig, K, B, V,
                                                arr(outs), _stream()), "edge_weight_bwd_multi")
    return outs


# ------------------------------------------------------------------ refine


def refine_forward(feature_in, edge_weight, si, sp, sc, ls, nl):
    f = _cuda(feature_in, torch.float32, "feature_in")
    if not (edge_weight.is_cuda and edge_weight.dtype == torch.float32 and edge_weight.is_contiguous()):
        raise RuntimeError("edge_weight must be a contiguous float32 CUDA tensor (it is written in place)")
    B, C, V = f.shape
    dev = f.device
    with _on(dev):
        out, aggr, aggr_up = torch.empty_like(f), torch.empty_like(f), torch.empty_like(f)
        wsum = torch.empty(B, V, dtype=torch.float32, device=dev)
        wsum_up = torch.empty(B, V, dtype=torch.float32, device=dev)
        check(lib.fedicra_refine_forward(_p(f), _p(edge_weight), _p(si), _p(sp), _p(sc), _p(ls), _p(nl), B, C, V,
                                         _p(out), _p(aggr), _p(aggr_up), _p(wsum), _p(wsum_up), _stream()),
              "refine_forward")
    return out, aggr, aggr_up, wsum, wsum_up


def refine_backward_feature(edge_weight, si, sp, sc, ls, nl, weight_sum, grad_out):
    g = _cuda(grad_out, torch.float32, "grad_output")
    B, C, V = g.shape
    with _on(g.device):
        out = torch.empty_like(g)
        ws = workspace("refine_bwd", lib.fedicra_refine_backward_workspace_bytes(B, C, V), g.device)
        check(lib.fedicra_refine_backward_feature(_p(edge_weight), _p(si), _p(sp), _p(sc), _p(ls), _p(nl),
                                                  _p(weight_sum), _p(g), B, C, V, _p(out), _p(ws), ws.numel(),
                                                  _stream()), "refine_backward_feature")
    return out


def refine_backward_weight(edge_weight, si, sp, sc, ls, nl, feature_out, feature_aggr, feature_aggr_up, weight_sum,
                           weight_sum_up, grad_out):
    g = _cuda(grad_out, torch.float32, "grad_output")
    B, C, V = g.shape
    with _on(g.device):
        out = torch.empty(B, V, dtype=torch.float32, device=g.device)
        ws = workspace("refine_bwd", lib.fedicra_refine_backward_workspace_bytes(B, C, V), g.device)
        check(lib.fedicra_refine_backward_weight(_p(edge_weight), _p(si), _p(sp), _p(sc), _p(ls), _p(nl),
                                                 _p(feature_out), _p(feature_aggr), _p(feature_aggr_up),
                                                 _p(weight_sum), _p(weight_sum_up), _p(g), B, C, V, _p(out), _p(ws),
                                                 ws.numel(), _stream()), "refine_backward_weight")
    return out


def band_plan(ls, nl):
    """Band plan of a batch of trees from their BFS level structure (one per topology, shared by all filter calls)."""
    B, V = ls.shape[0], ls.shape[1] - 1
    with _on(ls.device):
        bands = torch.empty(lib.fedicra_band_plan_ints(B), dtype=torch.int32, device=ls.device)
        check(lib.fedicra_band_plan(_p(ls), _p(nl), B, V, _p(bands), _stream()), "band_plan")
    return bands


def tree_filter_forward(feature_in, edge_weight, si, sp, sc, ls, nl, bands, want_aggr=False):
    """Depth-banded filter forward.  Returns (feature_out, saved[, feature_aggr, weight_sum])."""
    f = _cuda(feature_in, torch.float32, "feature_in")
    if not (edge_weight.is_cuda and edge_weight.dtype == torch.float32 and edge_weight.is_contiguous()):
        raise RuntimeError("edge_weight must be a contiguous float32 CUDA tensor (it is written in place)")
    B, C, V = f.shape
    dev = f.device
    with _on(dev):
        out = torch.empty_like(f)
        saved = torch.empty(lib.fedicra_tree_filter_saved_bytes(B, C, V), dtype=torch.uint8, device=dev)
        aggr = torch.empty_like(f) if want_aggr else None
        wsum = torch.empty(B, V, dtype=torch.float32, device=dev) if want_aggr else None
        ws = workspace("tree_filter", lib.fedicra_tree_filter_workspace_bytes(B, C, V), dev)
        check(lib.fedicra_tree_filter_forward(_p(f), _p(edge_weight), _p(si), _p(sp), _p(sc), _p(ls), _p(nl), _p(bands), B, C, V,
                                              _p(out), _p(aggr), _p(wsum), _p(saved), saved.numel(), _p(ws), ws.numel(),
                                              _stream()), "tree_filter_forward")
    return (out, saved, aggr, wsum) if want_aggr else (out, saved)


def tree_filter_backward(grad_out, edge_weight, si, sp, sc, ls, nl, bands, saved, want_weight_grad):
    """Depth-banded filter backward.  Returns (grad_feature, grad_weight or None)."""
    g = _cuda(grad_out, torch.float32, "grad_output")
    B, C, V = g.shape
    dev = g.device
    with _on(dev):
        gf = torch.empty_like(g)
        gw = torch.empty(B, V, dtype=torch.float32, device=dev) if want_weight_grad else None
        ws = workspace("tree_filter", lib.fedicra_tree_filter_workspace_bytes(B, C, V), dev)
        check(lib.fedicra_tree_filter_backward(_p(g), _p(edge_weight), _p(si), _p(sp), _p(sc), _p(ls), _p(nl), _p(bands), B, C, V,
                                               _p(saved), _p(gf), _p(gw), _p(ws), ws.numel(), _stream()),
              "tree_filter_backward")
    return gf, gw


# ------------------------------------------------------------------ loss glue


def channel_softmax_fwd(x):
    x = _cuda(x, torch.float32, "preds")
    B, C = x.shape[0], x.shape[1]
    P = x.numel() // (B * C)
    with _on(x.device):
        y = torch.empty_like(x)
        check(lib.fedicra_channel_softmax_fwd(_p(x), B, C, P, _p(y), _stream()), "channel_softmax_fwd")
    return y


def channel_softmax_bwd(y, grad_y):
    g = _cuda(grad_y, torch.float32, "grad_prob")
    B, C = y.shape[0], y.shape[1]
    P = y.numel() // (B * C)
    with _on(y.device):
        gx = torch.empty_like(y)
        check(lib.fedicra_channel_softmax_bwd(_p(y), _p(g), B, C, P, _p(gx), _stream()), "channel_softmax_bwd")
    return gx


def masked_l1_fwd(prob, filtered, roi_u8, weight):
    """-> sums [2] = (sum roi*|prob-filtered|, N), loss [1] = weight * sums[0] / max(N, 1)."""
    prob = _cuda(prob, torch.float32, "prob")
    filtered = _cuda(filtered, torch.float32, "filtered")
    roi = _cuda(roi_u8, torch.uint8, "roi")
    B, C = prob.shape[0], prob.shape[1]
    P = prob.numel() // (B * C)
    assert roi.numel() == B * P and filtered.shape == prob.shape
    with _on(prob.device):
        sums = torch.empty(2, dtype=torch.float32, device=prob.device)
        loss = torch.empty(1, dtype=torch.float32, device=prob.device)
        ws = workspace("masked_l1", lib.fedicra_masked_l1_workspace_bytes(), prob.device)
        check(lib.fedicra_masked_l1_fwd(_p(prob), _p(filtered), _p(roi), B, C, P, float(weight), _p(sums), _p(loss),
                                        _p(ws), ws.numel(), _stream()), "masked_l1_fwd")
    return sums, loss


def masked_l1_bwd(prob, filtered, roi_u8, grad_out, n, weight, want_prob=True, want_filtered=True):
    B, C = prob.shape[0], prob.shape[1]
    P = prob.numel() // (B * C)
    g = _cuda(grad_out.reshape(-1), torch.float32, "grad_loss")
    with _on(prob.device):
        gp = torch.empty_like(prob) if want_prob else None
        gf = torch.empty_like(prob) if want_filtered else None
        check(lib.fedicra_masked_l1_bwd(_p(prob), _p(filtered), _p(roi_u8), _p(g), _p(n), float(weight), B, C, P, _p(gp),
                                        _p(gf), _stream()), "masked_l1_bwd")
    return gp, gf


def bilinear_resize(heads, H, W, backward=False):
    """forward: heads = list of [B,C,h_k,w_k] -> list of [B,C,H,W] (torch's bilinear, align_corners=False, bit for bit);
    backward: heads = list of (grad [B,C,H,W], (h_k, w_k)) -> list of [B,C,h_k,w_k].  One launch for all heads."""
    import ctypes
    n = len(heads)
    if backward:
        srcs = [_cuda(g, torch.float32, "grad_resized") for g, _ in heads]
        sizes = [hw for _, hw in heads]
    else:
        srcs = [_cuda(x, torch.float32, "head") for x in heads]
        sizes = [tuple(x.shape[2:]) for x in srcs]
    B, C = srcs[0].shape[:2]
    dev = srcs[0].device
    with _on(dev):
        outs = [torch.empty(B, C, *(hw if backward else (H, W)), dtype=torch.float32, device=dev) for hw in sizes]
        sp = (ctypes.c_void_p * n)(*[x.data_ptr() for x in srcs])
        dp = (ctypes.c_void_p * n)(*[x.data_ptr() for x in outs])
        hh = (ctypes.c_int * n)(*[hw[0] for hw in sizes])
        ww = (ctypes.c_int * n)(*[hw[1] for hw in sizes])
        fn = lib.fedicra_bilinear_resize_bwd if backward else lib.fedicra_bilinear_resize_fwd
        check(fn(sp, dp, hh, ww, n, B * C, H, W, _stream()), "bilinear_resize")
    return outs


# ------------------------------------------------------------------ aggregation


def ala_update_flat(p, s, g, w, t, eta):
    n = p.numel()
    for name, x in (("p", p), ("s", s), ("w", w), ("t", t)):
        if not (x.is_cuda and x.dtype == torch.float32 and x.is_contiguous() and x.numel() == n):
            raise RuntimeError(f"ala_update_flat: {name} must be a contiguous float32 CUDA tensor of {n} elements")
    with _on(p.device):
        check(lib.fedicra_ala_update_flat(_p(p), _p(s), _p(g), _p(w), _p(t), n, float(eta), _stream()),
              "ala_update_flat")


def scale_inplace(x, alpha):
    if not (x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()):
        raise RuntimeError("scale_inplace: x must be a contiguous float32 CUDA tensor")
    with _on(x.device):
        check(lib.fedicra_scale_inplace(_p(x), x.numel(), float(alpha), _stream()), "scale_inplace")
