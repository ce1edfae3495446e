"""Adaptive local aggregation (element-wise part) and FedAvg, B200-native.

Reference (/root/reference/code/flower_common.py):
  :553-555   temp = p + (s - p) * w                          initial blend
  :590-597   w = clamp(w - eta * grad(temp) * (s - p), 0, 1) per tensor, skipped when the tensor has no gradient
  :600-602   temp = p + (s - p) * w
  :623-624   p <- temp
The reference runs these as three Python loops over 42 tensors (~600 tiny launches per batch); here each step is ONE
kernel launch over all tensors (csrc/ala.cu) and is bit-identical to the torch expressions.

FedAvg (flwr 1.0.0 ``aggregate``: sum_i n_i W_i / sum_i n_i, fed by flower_common.py:488-489, 70-72) becomes a pre-scale
kernel plus one NCCL all-reduce over the flat state.
"""
import torch

from . import ops
from ._lib import check, lib

_CHUNK = 4096


class _Tables:
    """Device-side pointer / chunk tables for one fixed list of (p, s, t, w) tensors."""

    def __init__(self, params, server_params, temp_params, weights):
        dev = params[0].device
        self.n = len(params)
        sizes = [p.numel() for p in params]
        chunk_tensor, chunk_begin = [], []
        for k, n in enumerate(sizes):
            for b in range(0, n, _CHUNK):
                chunk_tensor.append(k)
                chunk_begin.append(b)
        self.n_chunks = len(chunk_tensor)
        i64 = dict(dtype=torch.int64, device=dev)
        self.p = torch.tensor([x.data_ptr() for x in params], **i64)
        self.s = torch.tensor([x.data_ptr() for x in server_params], **i64)
        self.t = torch.tensor([x.data_ptr() for x in temp_params], **i64)
        self.w = torch.tensor([x.data_ptr() for x in weights], **i64)
        self.sizes = torch.tensor(sizes, **i64)
        self.chunk_tensor = torch.tensor(chunk_tensor, dtype=torch.int32, device=dev)
        self.chunk_begin = torch.tensor(chunk_begin, **i64)
        self.key = tuple(x.data_ptr() for group in (params, server_params, temp_params, weights) for x in group)


_TABLE_CACHE = {}


def fused_ala_update(params_p, server_params_p, temp_params_p, fedaa_weights, eta=1.0, use_grad=True):
    """One launch for flower_common.py:590-602 (``use_grad=True``) or :553-555 (``use_grad=False``).

    params_p / server_params_p / temp_params_p / fedaa_weights: equally long lists of float32 CUDA tensors (parameters or
    their ``.data``); ``temp_params_p[k].grad`` supplies the gradient, tensors without one keep their weight (:593-594).
    Updates ``fedaa_weights`` and ``temp_params_p`` in place.
    """
    datas = [[getattr(x, "data", x) for x in group] for group in (params_p, server_params_p, temp_params_p, fedaa_weights)]
    for group in datas:
        for x in group:
            if not (x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()):
                raise RuntimeError("fused_ala_update: all tensors must be contiguous float32 CUDA tensors")
    # addresses AND sizes: the caching allocator may hand the same addresses to tensors of another shape
    key = tuple((x.data_ptr(), x.numel()) for group in datas for x in group)
    tab = _TABLE_CACHE.get(key)
    if tab is None:
        if len(_TABLE_CACHE) > 16:
            _TABLE_CACHE.clear()
        tab = _Tables(*datas)
        _TABLE_CACHE[key] = tab
    dev = datas[0][0].device
    if use_grad:
        grads = [getattr(t, "grad", None) for t in temp_params_p]
        gptrs = [0 if g is None else g.contiguous().data_ptr() for g in grads]
        for g in grads:
            if g is not None and not g.is_contiguous():
                raise RuntimeError("fused_ala_update: gradients must be contiguous")
    else:
        gptrs = [0] * tab.n
    # gradient pointers usually do not move between batches (optimizer.zero_grad(set_to_none=False), or the caching
    # allocator handing back the same blocks): upload the table only when they did
    gkey = tuple(gptrs)
    if getattr(tab, "g_key", None) != gkey:
        tab.g_table = torch.tensor(gptrs, dtype=torch.int64, device=dev)
        tab.g_key = gkey
    g_table = tab.g_table
    with torch.cuda.device(dev):
        check(lib.fedicra_ala_update_multi(tab.p.data_ptr(), tab.s.data_ptr(), g_table.data_ptr(), tab.w.data_ptr(),
                                           tab.t.data_ptr(), tab.sizes.data_ptr(), tab.chunk_tensor.data_ptr(),
                                           tab.chunk_begin.data_ptr(), tab.n_chunks, _CHUNK, float(eta),
                                           torch.cuda.current_stream().cuda_stream), "ala_update_multi")


class AdaptiveAggregation:
    """The element-wise state of MyModel.set_weights (flower_common.py:549-555, 590-602, 623-624) on flat buffers.

    ``global_p`` (p, the freshly loaded global values), ``local_s`` (s, the client's previous values) and the learned
    ``weights`` live in three flat float32 buffers; ``temp`` is the blended model the weight-learning loop evaluates.
    """

    def __init__(self, global_p, local_s, eta=1.0):
        assert global_p.is_cuda and global_p.dtype == torch.float32 and global_p.dim() == 1
        self.p, self.s, self.eta = global_p.contiguous(), local_s.contiguous(), float(eta)
        self.weights = torch.ones_like(self.p)  # re-created on every call in the reference (:549-550)
        self.temp = torch.empty_like(self.p)
        ops.ala_update_flat(self.p, self.s, None, self.weights, self.temp, self.eta)

    def step(self, grad_temp):
        """One batch of weight learning: w <- clamp(w - eta*g*(s-p), 0, 1); temp <- p + (s-p)*w."""
        ops.ala_update_flat(self.p, self.s, grad_temp.contiguous(), self.weights, self.temp, self.eta)
        return self.temp

    def result(self):
        return self.temp


def fedavg_allreduce(flat_state, num_examples, group=None):
    """FedAvg over the ranks of ``group``: flat_state <- sum_i n_i * flat_state_i / sum_i n_i, in place, in flwr's order
    of operations (aggregate.py::aggregate: every client's arrays times its n_i, summed, divided by the total).

    One scale kernel, one all-reduce of the flat fp32 state (NCCL over NVLink on the box), one tiny all-reduce of the
    example counts, one scale by the reciprocal total -- no host synchronisation.  Mirrors flwr's aggregate() fed with
    (get_weights(), len(trainloader)) pairs (flower_common.py:70-72, 488-489).
    """
    import torch.distributed as dist

    n = torch.tensor([float(num_examples)], dtype=torch.float32, device=flat_state.device)
    dist.all_reduce(n, op=dist.ReduceOp.SUM, group=group)
    if flat_state.is_cuda:
        ops.scale_inplace(flat_state, float(num_examples))
    else:  # gloo tests of the host logic only; the product path is CUDA + NCCL
        flat_state.mul_(float(num_examples))
    dist.all_reduce(flat_state, op=dist.ReduceOp.SUM, group=group)
    flat_state.div_(n)
    return flat_state


class FlatModelState:
    """The whole state_dict of a module in two persistent flat buffers, the module's own tensors re-pointed to views of
    them -- what `get_weights` / `set_weights` / `load_state_dict` move tensor by tensor through numpy in the reference
    (flower_common.py:488-489, 499-502):

      flat  float32: every floating entry (parameters and BatchNorm running statistics), in state_dict order
      ints  int64:   every integer entry (BatchNorm ``num_batches_tracked``)

    so that FedAvg is ONE in-place all-reduce of ``flat`` (plus a 21-element one for the counters) and loading the
    global model is no copy at all: the module already looks at the reduced buffer.
    """

    def __init__(self, module):
        entries = list(module.state_dict(keep_vars=True).items())
        self.names = [k for k, _ in entries]
        dev = entries[0][1].device
        fl = [(k, t) for k, t in entries if t.is_floating_point()]
        it = [(k, t) for k, t in entries if not t.is_floating_point()]
        self.flat = torch.empty(sum(t.numel() for _, t in fl), dtype=torch.float32, device=dev)
        self.ints = torch.empty(sum(t.numel() for _, t in it), dtype=torch.int64, device=dev)
        self.views, self.slices = {}, {}
        for buf, group in ((self.flat, fl), (self.ints, it)):
            off = 0
            for k, t in group:
                n = t.numel()
                view = buf[off:off + n].view(t.shape)
                view.copy_(t.detach())
                t.data = view                 # the module's parameter / buffer now lives inside the flat buffer
                self.views[k], self.slices[k] = view, (buf is self.flat, off, n)
                off += n

    def numel(self):
        return self.flat.numel() + self.ints.numel()

    def get_weights(self):
        """The reference's wire format (flower_common.py:488-489): one numpy array per state_dict entry."""
        return [self.views[k].detach().cpu().numpy() for k in self.names]

    def set_weights(self, arrays):
        """Inverse of get_weights: flwr's aggregate turns integer counters into floats, torch casts them back."""
        for k, a in zip(self.names, arrays):
            self.views[k].copy_(torch.as_tensor(a).to(self.views[k].dtype))

    def fedavg_(self, num_examples, group=None):
        """In-place FedAvg of the whole state over the ranks (one client per rank).  The integer counters go through
        the same weighted mean in float64 and are truncated on the way back, as numpy + load_state_dict do."""
        import torch.distributed as dist
        fedavg_allreduce(self.flat, num_examples, group)
        if self.ints.numel():
            c = torch.cat([self.ints.double() * float(num_examples),
                           torch.tensor([float(num_examples)], dtype=torch.float64, device=self.ints.device)])
            dist.all_reduce(c, op=dist.ReduceOp.SUM, group=group)
            self.ints.copy_((c[:-1] / c[-1]).to(torch.int64))
        return self
