"""ctypes door onto ``liboracle_tree.so`` (oracle/tree_oracle.c) -- TEST INFRASTRUCTURE ONLY.

Exposes the five operators of the reference's pybind module ``tree_filter_cuda``
(lib_tree_filter/src/tree_filter.cpp:7-13) with the same argument lists, working on CPU torch tensors, so the
reference's unmodified Python glue (functions/{mst,bfs,refine}.py, modules/tree_filter.py) can run on top of it.

``set_mst_impl("reference")`` routes ``mst_forward`` through ``oracle/_ref/libboruvka_ref.so`` -- the reference's own
boruvka.cpp compiled by ``make -C oracle ref`` -- instead of the C restatement.
"""
import ctypes
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_LIB64 = None
_REF = None
_MST_IMPL = "port"

_I = ctypes.c_int
_P = ctypes.c_void_p
_F = ctypes.c_float


def build(ref=True):
    """Compile the C restatement (always) and the reference's boruvka.cpp (only where /root/reference exists)."""
    subprocess.run(["make", "-s", "-C", _HERE, "all"], check=True)
    if ref and os.path.isdir("/root/reference/code"):
        subprocess.run(["make", "-s", "-C", _HERE, "ref"], check=True)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle_tree.so")
        if not os.path.exists(path):
            build(ref=False)
        _LIB = ctypes.CDLL(path)
        _LIB.oracle_boruvka.restype = _I
        _LIB.oracle_mst_forward.restype = _I
        _LIB.oracle_bfs.restype = _I
        _LIB.oracle_bfs_forward.restype = _I
        _LIB.oracle_num_threads.restype = _I
    return _LIB


def lib64():
    """The same restatement compiled with REAL = double (liboracle_tree64.so): the high-precision judge."""
    global _LIB64
    if _LIB64 is None:
        path = os.path.join(_HERE, "liboracle_tree64.so")
        if not os.path.exists(path):
            build(ref=False)
        _LIB64 = ctypes.CDLL(path)
    return _LIB64


def _for(t):
    """(library, scalar ctype) matching the floating dtype of tensor `t`."""
    return (lib64(), ctypes.c_double) if t.dtype == torch.float64 else (lib(), _F)


def ref_lib():
    """The reference's boruvka.cpp as a shared object, or None when it has not been built."""
    global _REF
    if _REF is None:
        path = os.path.join(_HERE, "_ref", "libboruvka_ref.so")
        if not os.path.exists(path):
            return None
        _REF = ctypes.CDLL(path)
        _REF.ref_mst_forward.restype = _I
    return _REF


def set_mst_impl(kind):
    global _MST_IMPL
    assert kind in ("port", "reference")
    if kind == "reference" and ref_lib() is None:
        raise RuntimeError("oracle/_ref/libboruvka_ref.so is not built (run `make -C oracle ref` where /root/reference exists)")
    _MST_IMPL = kind


def num_threads():
    return lib().oracle_num_threads()


def set_num_threads(n):
    lib().oracle_set_num_threads(_I(int(n)))


def _i32(t):
    assert t.dtype == torch.int32 and t.device.type == "cpu"
    return t.contiguous()


def _f32(t, like=None):
    want = torch.float32 if like is None else like.dtype
    assert t.dtype == want and t.device.type == "cpu", (t.dtype, want)
    return t.contiguous()


def _fl(t):
    assert t.dtype in (torch.float32, torch.float64) and t.device.type == "cpu"
    return t.contiguous()


def _p(t):
    return _P(t.data_ptr())


# ---------------------------------------------------------------- the five tree_filter_cuda operators


def mst_forward(edge_index, edge_weight, vertex_count):
    """mst.hpp:4-7 -- [B,E,2] i32, [B,E] f32, V -> [B,V-1,2] i32."""
    ei, ew = _i32(edge_index), _f32(edge_weight)
    B, E = ei.shape[0], ei.shape[1]
    out = torch.empty(B, vertex_count - 1, 2, dtype=torch.int32)
    if _MST_IMPL == "reference":
        rc = ref_lib().ref_mst_forward(_I(B), _I(E), _I(vertex_count), _p(ei), _p(ew), _p(out))
    else:
        rc = lib().oracle_mst_forward(_I(B), _I(E), _I(vertex_count), _p(ei), _p(ew), _p(out))
    if rc != 0:
        raise RuntimeError("oracle mst_forward: input graph is not connected")
    return out


def bfs_forward(tree, max_adj):
    """bfs.hpp:4-8 -- [B,V-1,2] i32 -> sorted_index [B,V], sorted_parent [B,V], sorted_child [B,V,max_adj]."""
    t = _i32(tree)
    B, V = t.shape[0], t.shape[1] + 1
    si = torch.zeros(B, V, dtype=torch.int32)
    sp = torch.zeros(B, V, dtype=torch.int32)
    sc = torch.zeros(B, V, max_adj, dtype=torch.int32)
    rc = lib().oracle_bfs_forward(_I(B), _I(V), _p(t), _I(max_adj), _p(si), _p(sp), _p(sc))
    if rc != 0:
        raise RuntimeError("oracle bfs_forward: not a spanning tree with degree <= max_adj")
    return si, sp, sc


def bfs_levels(tree_one, max_adj=4):
    """One image: also return the level_start array (first BFS position of every depth)."""
    t = _i32(tree_one)
    V = t.shape[0] + 1
    si = torch.zeros(V, dtype=torch.int32)
    sp = torch.zeros(V, dtype=torch.int32)
    sc = torch.zeros(V, max_adj, dtype=torch.int32)
    ls = torch.zeros(V + 1, dtype=torch.int32)
    nl = ctypes.c_int(0)
    rc = lib().oracle_bfs(_I(V), _p(t), _I(max_adj), _p(si), _p(sp), _p(sc), _p(ls), ctypes.byref(nl))
    if rc != 0:
        raise RuntimeError("oracle bfs: not a spanning tree with degree <= max_adj")
    return si, sp, sc, ls[: nl.value + 1].clone()


def refine_forward(feature_in, edge_weight, sorted_index, sorted_parent, sorted_child):
    """refine.hpp:4-11.  Like the reference, writes 0 into edge_weight[:,0] through the raw pointer."""
    f = _fl(feature_in)
    assert edge_weight.is_contiguous() and edge_weight.dtype == f.dtype
    B, C, V = f.shape
    A = sorted_child.shape[2]
    out, aggr, aggr_up = torch.empty_like(f), torch.empty_like(f), torch.empty_like(f)
    ws, ws_up = torch.empty(B, V, dtype=f.dtype), torch.empty(B, V, dtype=f.dtype)
    _for(f)[0].oracle_refine_forward(_I(B), _I(C), _I(V), _I(A), _p(f), _p(edge_weight), _p(_i32(sorted_index)),
                                _p(_i32(sorted_parent)), _p(_i32(sorted_child)), _p(out), _p(aggr), _p(aggr_up),
                                _p(ws), _p(ws_up))
    return out, aggr, aggr_up, ws, ws_up


def refine_backward_feature(feature_in, edge_weight, sorted_index, sorted_parent, sorted_child, feature_out,
                            feature_aggr, feature_aggr_up, weight_sum, weight_sum_up, grad_out):
    """refine.hpp:13-25."""
    B, C, V = feature_in.shape
    A = sorted_child.shape[2]
    go = _fl(grad_out)
    g = torch.empty(B, C, V, dtype=go.dtype)
    _for(go)[0].oracle_refine_backward_feature(_I(B), _I(C), _I(V), _I(A), _p(_f32(edge_weight, go)),
                                               _p(_i32(sorted_index)), _p(_i32(sorted_parent)), _p(_i32(sorted_child)),
                                               _p(_f32(weight_sum, go)), _p(go), _p(g))
    return g


def refine_backward_weight(feature_in, edge_weight, sorted_index, sorted_parent, sorted_child, feature_out,
                           feature_aggr, feature_aggr_up, weight_sum, weight_sum_up, grad_out):
    """refine.hpp:27-39."""
    B, C, V = feature_in.shape
    A = sorted_child.shape[2]
    go = _fl(grad_out)
    g = torch.empty(B, V, dtype=go.dtype)
    _for(go)[0].oracle_refine_backward_weight(_I(B), _I(C), _I(V), _I(A), _p(_f32(edge_weight, go)),
                                              _p(_i32(sorted_index)), _p(_i32(sorted_parent)), _p(_i32(sorted_child)),
                                              _p(_f32(feature_out, go)), _p(_f32(feature_aggr, go)),
                                              _p(_f32(feature_aggr_up, go)), _p(_f32(weight_sum, go)),
                                              _p(_f32(weight_sum_up, go)), _p(go), _p(g))
    return g


# ---------------------------------------------------------------- edge weights (tree_filter.py:84-108)


def edge_weight(embed, sorted_index, sorted_parent, inv_sigma):
    e = _fl(embed)
    B, D, V = e.shape
    ew = torch.empty(B, V, dtype=e.dtype)
    L, S = _for(e)
    L.oracle_edge_weight(_I(B), _I(D), _I(V), _p(e), _p(_i32(sorted_index)), _p(_i32(sorted_parent)), S(inv_sigma), _p(ew))
    return ew


def edge_weight_bwd(embed, sorted_index, sorted_parent, inv_sigma, ew, grad_ew):
    e = _fl(embed)
    B, D, V = e.shape
    g = torch.empty(B, D, V, dtype=e.dtype)
    L, S = _for(e)
    L.oracle_edge_weight_bwd(_I(B), _I(D), _I(V), _p(e), _p(_i32(sorted_index)), _p(_i32(sorted_parent)), S(inv_sigma),
                             _p(_f32(ew, e)), _p(_f32(grad_ew, e)), _p(g))
    return g
