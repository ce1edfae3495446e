"""Loss-stack step runner: the weak-supervision loss of one training iteration as ONE call, optionally captured into a
CUDA graph and fed from host memory through double-buffered copies.

The step is what the reference's loop computes between the network's forward and `optimizer.step()`
(/root/reference/code/flower_pCE_2D_GateCRFMsacleTreeEnergyLoss_Ours.py:132-151,172), minus the network and the
cross-entropy term: from (logits, aux1..3, image, label) to (loss, d/dlogits, d/daux1..3):

    tel  = MScaleRecurveTreeEnergyLoss()(logits, three_channel(image), aux1, aux2, aux3, label == C, tree_weight)[0]
    crf  = ModelLossSemsegGatedCRF()(softmax(logits), kernels_desc, radius, image, H, W)["loss"]
    loss = tel + crf_weight * crf ;  loss.backward()

`LossStackStep` is the eager call; `GraphedLossStack` replays it from static buffers; `PipelinedLossStack` keeps two
graphed copies and overlaps the host->device copy of step k+1 and the device->host copy of step k-1 with the replay of
step k on separate streams.
"""
import os

import torch

from .gate_crf_loss import ModelLossSemsegGatedCRF
from .tree_energy import MScaleRecurveTreeEnergyLoss
from .tree_filter import side_stream

KERNELS_DESC = [{"weight": 1, "xy": 6, "rgb": 0.1}]  # OURS.py:68
RADIUS = 5                                           # OURS.py:69
INPUT_KEYS = ("image", "logits", "aux1", "aux2", "aux3", "label")
GRAD_KEYS = ("logits", "aux1", "aux2", "aux3")


class LossStackStep:
    """One eager step.  `tel` / `crf` may be any modules with the reference signatures (the reference's own classes
    included); by default this package's."""

    def __init__(self, num_classes, height, width, tree_weight=0.1, crf_weight=0.1, distributed=False, world=1,
                 tel=None, crf=None, local_share=True):
        self.C, self.H, self.W = num_classes, height, width
        self.tree_weight = tree_weight
        # sharded: the returned loss is this rank's share of the global loss (gradients are the global ones); the count
        # of unlabelled pixels is all-reduced on a side stream under the step
        self.tel = tel if tel is not None else MScaleRecurveTreeEnergyLoss(distributed=distributed, local_share=local_share)
        self.crf = crf if crf is not None else ModelLossSemsegGatedCRF()
        # Gated CRF normalises by the LOCAL N*H*W (gate_crf_loss.py:63); the global batch is `world` times larger
        self.crf_scale = crf_weight / world

    def __call__(self, d):
        leaves = [d[k].detach().requires_grad_() for k in GRAD_KEYS]
        image = d["image"]
        three = image.repeat(1, 3, 1, 1) if image.shape[1] == 1 else image          # OURS.py:137-141
        roi = d["label"] == self.C                                                    # OURS.py:135
        side = side_stream("crf")
        if side is None:
            tel = self.tel(leaves[0], three, leaves[1], leaves[2], leaves[3], roi, self.tree_weight)[0]
            soft = torch.softmax(leaves[0], dim=1)                                        # OURS.py:132
            crf = self.crf(soft, KERNELS_DESC, RADIUS, image, self.H, self.W)["loss"]     # OURS.py:143-150
        else:
            # the Gated CRF term shares only `logits` and `image` with the tree term: it runs (forward, and through
            # autograd its backward) on a helper stream under the latency-bound tree kernels
            main = torch.cuda.current_stream()
            box = {}

            def crf_branch():
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    soft = torch.softmax(leaves[0], dim=1)
                    box["crf"] = self.crf_scale * self.crf(soft, KERNELS_DESC, RADIUS, image, self.H, self.W)["loss"]

            # where the branch forks: measured at faz256, per step: at the start 1.373 ms (it then competes with the MST tile
            # kernel, which fills every SM), once the topology is enqueued 1.347 ms, after the tree term 1.386 ms
            at = os.environ.get("FEDICRA_CRF_AT", "topology")
            hook = at == "topology" and hasattr(self.tel, "on_topology_enqueued")
            if at == "start" or (at == "topology" and not hook):
                crf_branch()
            if hook:
                self.tel.on_topology_enqueued = crf_branch
            try:
                tel = self.tel(leaves[0], three, leaves[1], leaves[2], leaves[3], roi, self.tree_weight)[0]
            finally:
                if hook:
                    self.tel.on_topology_enqueued = None
            if "crf" not in box:
                crf_branch()
            crf = box["crf"]
            main.wait_stream(side)
            crf.record_stream(main)
        loss = tel + (crf if side is not None else self.crf_scale * crf)              # OURS.py:151 (without pCE)
        loss.backward()                                                               # OURS.py:172
        return loss.detach(), [x.grad for x in leaves]


class GraphedLossStack:
    """The step captured once into a CUDA graph and replayed: static input buffers, outputs in the graph's pool.  Every
    kernel of the library launches on the capturing stream and none synchronises or allocates outside torch's allocator,
    so the step is capture-safe (at N > 1 the 2-scalar all-reduce of the normaliser is captured with it)."""

    def __init__(self, step, example):
        self.static_in = {k: v.clone() for k, v in example.items()}
        # Warm-up and capture on ONE stream: the library's scratch buffers are cached per (purpose, device, stream), and a
        # buffer first asked for during capture would put its zero-fill (hundreds of MB for the topology stage) into the
        # graph -- 0.13 ms on every replay of the 1.4 ms step.
        stream = torch.cuda.Stream()
        stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(stream):
            for _ in range(3):
                step(self.static_in)
        torch.cuda.current_stream().wait_stream(stream)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=stream):
            self.loss, self.grads = step(self.static_in)
        torch.cuda.synchronize()

    def load(self, src, non_blocking=True):
        for k, v in src.items():
            self.static_in[k].copy_(v, non_blocking=non_blocking)

    def __call__(self, d=None):
        if d is not None and d is not self.static_in:
            self.load(d)
        self.graph.replay()
        return self.loss, self.grads

    def reset(self):
        self.graph.reset()


class PipelinedLossStack:
    """Host-to-host step with the copies hidden behind the computation.

    `submit(pinned_inputs)` enqueues, on three streams, the host->device copy of the inputs into one of `depth` graphed
    copies of the step, its replay, and the device->host copy of loss + gradients into pinned buffers; `result()` waits
    for the OLDEST outstanding step and returns (loss float, list of pinned gradient tensors -- valid until that slot is
    used again).  With depth 2 the copies of the neighbouring steps overlap the replay of the current one; every step
    still moves all its inputs in and all its outputs out."""

    def __init__(self, step, example, depth=2, copy_gradients=True):
        """copy_gradients=False: only the loss travels back to the host every step (in a training loop the gradients
        are consumed on the device by the network's backward); the inputs still come from the host every step."""
        self.copy_gradients = copy_gradients
        self.slots = []
        self.h2d, self.d2h = torch.cuda.Stream(), torch.cuda.Stream()
        self.compute = torch.cuda.current_stream()
        for _ in range(depth):
            g = GraphedLossStack(step, example)
            host_grads = [torch.empty(x.shape, dtype=x.dtype).pin_memory() for x in g.grads]
            host_loss = torch.empty(1, dtype=torch.float32).pin_memory()
            self.slots.append({"g": g, "host_grads": host_grads, "host_loss": host_loss,
                               "in_ready": torch.cuda.Event(), "computed": torch.cuda.Event(),
                               "out_ready": torch.cuda.Event(), "busy": False})
        self.next, self.pending = 0, []
        self.h2d_bytes = sum(v.numel() * v.element_size() for v in example.values())
        self.d2h_bytes = 4 + (sum(x.numel() * x.element_size() for x in self.slots[0]["g"].grads) if copy_gradients else 0)

    def submit(self, pinned):
        s = self.slots[self.next % len(self.slots)]
        if s["busy"]:
            raise RuntimeError("PipelinedLossStack: all slots in flight; call result() first")
        g = s["g"]
        self.h2d.wait_event(s["computed"])            # the previous replay of this slot has consumed its inputs
        with torch.cuda.stream(self.h2d):
            for k, v in pinned.items():
                g.static_in[k].copy_(v, non_blocking=True)
            s["in_ready"].record(self.h2d)
        self.compute.wait_event(s["in_ready"])
        self.compute.wait_event(s["out_ready"])       # the previous outputs of this slot have left the device
        g.graph.replay()
        s["computed"].record(self.compute)
        self.d2h.wait_event(s["computed"])
        with torch.cuda.stream(self.d2h):
            s["host_loss"].copy_(g.loss.reshape(1), non_blocking=True)
            if self.copy_gradients:
                for h, x in zip(s["host_grads"], g.grads):
                    h.copy_(x, non_blocking=True)
            s["out_ready"].record(self.d2h)
        s["busy"] = True
        self.pending.append(s)
        self.next += 1

    def result(self):
        s = self.pending.pop(0)
        s["out_ready"].synchronize()
        s["busy"] = False
        return float(s["host_loss"][0]), s["host_grads"]

    def run(self, batches):
        """Feed an iterable of pinned input dicts; yields (loss, host gradients) per batch, in order."""
        depth = len(self.slots)
        for b in batches:
            if len(self.pending) == depth:
                yield self.result()
            self.submit(b)
        while self.pending:
            yield self.result()

    def reset(self):
        torch.cuda.synchronize()
        for s in self.slots:
            s["g"].reset()
