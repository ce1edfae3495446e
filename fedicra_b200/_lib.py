"""ctypes binding of libfedicra_sm100.so (the C ABI declared in include/fedicra_sm100.h).

The library is built in-tree by ``__graft_entry__.build()`` / ``make -C fedicra_b200/csrc``.  If it is missing this module
raises at import: there is deliberately no fallback implementation.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfedicra_sm100.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "or `make -C fedicra_b200/csrc` (nvcc, sm_100a). fedicra_b200 has no CPU or PyTorch fallback."
    )

lib = ctypes.CDLL(LIB_PATH)

c_int, c_size_t, c_void_p, c_float, c_int64 = ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_float, ctypes.c_int64


class GcrfDesc(ctypes.Structure):
    _fields_ = [("weight", c_float), ("sigma_xy", c_float), ("sigma_rgb", c_float)]


# name -> (restype, argtypes); every symbol include/fedicra_sm100.h declares
SIGNATURES = {
    "fedicra_last_error": (ctypes.c_char_p, []),
    "fedicra_abi_version": (c_int, []),
    "fedicra_launch_count": (ctypes.c_longlong, []),
    "fedicra_gcrf_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "fedicra_gcrf_fwd_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                     ctypes.POINTER(GcrfDesc), c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "fedicra_gcrf_noalloc_max_bytes": (c_size_t, []),
    "fedicra_gcrf_fwd_bwd_noalloc": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                             ctypes.POINTER(GcrfDesc), c_int, c_void_p, c_void_p, c_void_p]),
    "fedicra_mst_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "fedicra_mst_forward": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_size_t, c_void_p]),
    "fedicra_mst_grid": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "fedicra_mst_grid_multi": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                       c_size_t, c_void_p]),
    "fedicra_bfs_grid": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "fedicra_bfs_workspace_bytes": (c_size_t, [c_int, c_int]),
    "fedicra_bfs_forward": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                    c_size_t, c_void_p]),
    "fedicra_bfs_euler_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "fedicra_bfs_euler": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_size_t, c_void_p]),
    "fedicra_bfs_euler_rooted": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                         c_void_p, c_void_p, c_size_t, c_void_p]),
    "fedicra_chain_order": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                    c_void_p]),
    "fedicra_chain_order_weights": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                            c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "fedicra_chain_filter_max_channels": (c_int, []),
    "fedicra_chain_filter_saved_bytes": (c_size_t, [c_int, c_int, c_int]),
    "fedicra_chain_filter_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "fedicra_chain_filter_saved_offset": (c_size_t, [c_int, c_int, c_int]),
    "fedicra_chain_filter_forward": (c_int, [c_void_p] * 6 + [c_int, c_int, c_int, c_void_p, c_void_p, c_size_t, c_void_p]),
    "fedicra_chain_filter_backward": (c_int, [c_void_p] * 6 + [c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                                                c_size_t, c_void_p]),
    "fedicra_chain_filter_forward_armed": (c_int, [c_void_p] * 6 + [c_int, c_int, c_int, c_void_p, c_void_p, c_size_t, c_void_p]),
    "fedicra_chain_filter_backward_armed": (c_int, [c_void_p] * 6 + [c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                                                c_size_t, c_void_p]),
    "fedicra_chain_filter_arm": (c_int, [c_void_p, c_size_t, c_void_p]),
    "fedicra_edge_weight_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p, c_void_p]),
    "fedicra_edge_weight_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                        c_float, c_void_p, c_void_p]),
    "fedicra_edge_weight_bwd_multi": (c_int, [c_void_p] * 8 + [c_int, c_int, c_int, c_void_p, c_void_p]),
    "fedicra_refine_forward": (c_int, [c_void_p] * 7 + [c_int, c_int, c_int] + [c_void_p] * 5 + [c_void_p]),
    "fedicra_refine_backward_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "fedicra_refine_backward_feature": (c_int, [c_void_p] * 8 + [c_int, c_int, c_int, c_void_p, c_void_p, c_size_t,
                                                                 c_void_p]),
    "fedicra_refine_backward_weight": (c_int, [c_void_p] * 12 + [c_int, c_int, c_int, c_void_p, c_void_p, c_size_t,
                                                                 c_void_p]),
    "fedicra_tree_filter_saved_bytes": (c_size_t, [c_int, c_int, c_int]),
    "fedicra_tree_filter_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "fedicra_band_plan_ints": (c_size_t, [c_int]),
    "fedicra_band_plan": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "fedicra_tree_filter_forward": (c_int, [c_void_p] * 8 + [c_int, c_int, c_int] + [c_void_p] * 4 + [c_size_t, c_void_p,
                                                                                                    c_size_t, c_void_p]),
    "fedicra_tree_filter_backward": (c_int, [c_void_p] * 8 + [c_int, c_int, c_int] + [c_void_p] * 4 + [c_size_t, c_void_p]),
    "fedicra_channel_softmax_fwd": (c_int, [c_void_p, c_int, c_int, ctypes.c_longlong, c_void_p, c_void_p]),
    "fedicra_channel_softmax_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, ctypes.c_longlong, c_void_p, c_void_p]),
    "fedicra_masked_l1_workspace_bytes": (c_size_t, []),
    "fedicra_masked_l1_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, ctypes.c_longlong, c_float, c_void_p,
                                      c_void_p, c_void_p, c_size_t, c_void_p]),
    "fedicra_masked_l1_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_int, c_int,
                                      ctypes.c_longlong, c_void_p, c_void_p, c_void_p]),
    "fedicra_bilinear_resize_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "fedicra_bilinear_resize_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "fedicra_debug_filter_profile": (c_int, [c_void_p, c_int]),
    "fedicra_debug_chain_profile": (c_int, [c_void_p, c_int]),
    "fedicra_debug_mst_profile": (c_int, [c_void_p, c_int]),
    "fedicra_ala_update_flat": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_void_p]),
    "fedicra_ala_update_multi": (c_int, [c_void_p] * 8 + [c_int, c_int, c_float, c_void_p]),
    "fedicra_scale_inplace": (c_int, [c_void_p, c_int64, c_float, c_void_p]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here = library out of date with the header
    _fn.restype = _res
    _fn.argtypes = _args


def check(rc, what):
    """Turn a negative return code into RuntimeError, the way the reference's C++ exceptions surface in Python."""
    if rc != 0:
        raise RuntimeError(f"{what}: {lib.fedicra_last_error().decode()} (code {rc})")
