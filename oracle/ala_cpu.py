"""CPU restatement of the adaptive-aggregation element-wise update and the FedAvg reduce -- TEST INFRASTRUCTURE ONLY.

  ala_update   <- /root/reference/code/flower_common.py:590-602  (w <- clamp(w - eta*g*(s-p), 0, 1); t <- p + (s-p)*w)
  ala_blend    <- flower_common.py:553-555
  fedavg       <- flwr==1.0.0 (fed39v2.yaml:74) flwr/server/strategy/aggregate.py::aggregate, not vendored in the
                  reference: sum_i n_i*W_i / sum_i n_i per array.  Parity unpinned for this one (no reference source,
                  tests or vectors under /root/reference); restated from its published definition.
"""
import numpy as np


def ala_update(p, s, g, w, eta=1.0):
    """numpy fp32, evaluated operation by operation like the torch expression (no fused multiply-add)."""
    p, s, g, w = (np.asarray(a, dtype=np.float32) for a in (p, s, g, w))
    diff = s - p
    w_new = np.clip(w - np.float32(eta) * (g * diff), np.float32(0), np.float32(1)).astype(np.float32)
    t_new = (p + diff * w_new).astype(np.float32)
    return w_new, t_new


def ala_blend(p, s, w):
    p, s, w = (np.asarray(a, dtype=np.float32) for a in (p, s, w))
    return (p + (s - p) * w).astype(np.float32)


def fedavg(client_arrays, num_examples):
    """client_arrays: list (clients) of lists (arrays).  Returns the weighted mean per array."""
    total = float(sum(num_examples))
    out = []
    for arrays in zip(*client_arrays):
        acc = sum(np.asarray(a, dtype=np.float64) * n for a, n in zip(arrays, num_examples))
        out.append((acc / total).astype(arrays[0].dtype))
    return out
