"""`MyModel.set_weights` of the FedICRA strategy (/root/reference/code/flower_common.py:491-633) with the element-wise
adaptive aggregation on the fused kernels and, optionally, the per-batch work of its weight-learning loop replayed from
a CUDA graph.

The reference's loop costs ~600 tiny launches per batch around a U-Net forward/backward: three Python loops over the 42
decoder tensors (:590-602), an AdamW step with lr = 0 that changes nothing (:560, :587-588), and a host
round trip per batch.  Here a batch is: copy the batch into static buffers, replay one graph holding forward ->
cross-entropy -> backward -> `fused_ala_update` (ONE launch for :590-602), nothing else.

Semantics kept, line by line: the deep copies and the load of the global weights (:496-502), the early returns (:521-526),
the `local_keys` selection (:504-511), the blend before the loop (:553-555), the loss bookkeeping and the convergence
rule (:609-621), `start_phase`, and the final copy (:623-624).  Also kept by default: the learned weights restart from
ones on every call, because the reference's guard tests an attribute that is never set (`hasattr(self, 'weights')`,
:549); `keep_weights=True` keeps them across rounds as the guard evidently intended.

    adaptive_set_weights(wrapper, weights, config)      # wrapper: the reference's MyModel / MyModelV2 instance
    install_set_weights(flower_common_module)           # patches MyModel.set_weights for strategy == 'FedICRA'
"""
import copy
from collections import OrderedDict

import numpy as np
import torch
import torch.nn as nn

from .aggregation import fused_ala_update

IGNORE_KEYS = ("out_conv", "up4", "up3", "up2", "up1")   # flower_common.py:507


def _batch_to_device(wrapper, sampled_batch):
    """flower_common.py:569-575."""
    if wrapper.args.img_class == "faz":
        volume, label = sampled_batch["image"].unsqueeze(1), sampled_batch["label"]
    else:
        volume, label = sampled_batch["image"], sampled_batch["label"]
    return volume.cuda(non_blocking=True), label.cuda(non_blocking=True)


class _GraphedBatch:
    """forward -> CE -> backward -> fused update of one batch, captured once; gradients live in persistent buffers."""

    def __init__(self, temp_model, ce_loss, update, volume, label):
        self.x, self.y = volume.clone(), label.clone()
        self.temp_params = [p for p in temp_model.parameters() if p.requires_grad]

        def one():
            for p in self.temp_params:
                if p.grad is not None:
                    p.grad.zero_()
            loss = ce_loss(temp_model(self.x)[0], self.y[:].long())
            loss.backward()
            update()
            return loss.detach()

        side = torch.cuda.Stream()         # warm-up and capture on the same stream (scratch buffers are cached per stream)
        side.wait_stream(torch.cuda.current_stream())
        self.warm = 3
        with torch.cuda.stream(side):
            for _ in range(self.warm):     # also creates the gradient buffers and the update's pointer tables
                one()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=side):
            self.loss = one()

    def __call__(self, volume, label):
        self.x.copy_(volume, non_blocking=True)
        self.y.copy_(label, non_blocking=True)
        self.graph.replay()
        return self.loss


def adaptive_set_weights(wrapper, weights, config, use_graph=False, keep_weights=False, eta=1.0, num_pre_loss=10,
                         threshold=0.1, log=print):
    """Drop-in body of MyModel.set_weights for strategy 'FedICRA'.

    `use_graph=True` replays the per-batch work from a CUDA graph.  It needs equally shaped batches and a network whose
    forward is capture-safe; the reference's UNet_LC_MultiHead is not (LCEncoder.get_emb builds its index tensor on the
    host and copies it over on every call, networks/unet.py:182), so with the reference network leave it False.  The
    warm-up before capture runs the first batch three extra times (BatchNorm statistics of the throw-away temp model
    see them; the learned weights do too), so graph mode is close to, not bit-for-bit, the eager loop."""
    model = wrapper.model
    server_model = copy.deepcopy(model)                                                  # :496
    state = OrderedDict((k, torch.as_tensor(v)) for k, v in zip(model.state_dict().keys(), weights))
    model.load_state_dict(state, strict=False)                                           # :497-500
    temp_model = copy.deepcopy(model)                                                    # :501
    local_keys = [n for n, _ in model.named_parameters() if any(k in n for k in IGNORE_KEYS)]
    params, server_params = list(model.parameters()), list(server_model.parameters())
    if torch.sum(server_params[0] - params[0]) == 0:                                     # :521-523
        return
    if config["iter_global"] <= 50:                                                      # :524-526
        log("skip", config)
        return
    pick = lambda m: [p for n, p in m.named_parameters() if n in local_keys]
    params_p, server_params_p, temp_params_p = pick(model), pick(server_model), pick(temp_model)
    for n, p in temp_model.named_parameters():                                           # :540-545
        p.requires_grad = n in local_keys
    if not (keep_weights and getattr(wrapper, "fedaa_weights", None) is not None):       # :548-550 (see module docstring)
        wrapper.fedaa_weights = [torch.ones_like(p.data) for p in params_p]
    w = wrapper.fedaa_weights
    fused_ala_update(params_p, server_params_p, temp_params_p, w, eta, use_grad=False)   # :553-555
    ce_loss = nn.CrossEntropyLoss(ignore_index=wrapper.args.num_classes)
    update = lambda: fused_ala_update(params_p, server_params_p, temp_params_p, w, eta, use_grad=True)   # :590-602
    graphed, losses, count = None, [], 0
    while True:
        loss = None
        for sampled_batch in wrapper.trainloader:
            volume, label = _batch_to_device(wrapper, sampled_batch)
            if use_graph:
                if graphed is None:
                    graphed = _GraphedBatch(temp_model, ce_loss, update, volume, label)
                loss = graphed(volume, label)
            else:
                for p in temp_params_p:
                    p.grad = None                                                         # optimizer.zero_grad(), :581
                loss = ce_loss(temp_model(volume)[0], label[:].long())                   # :577-579
                loss.backward()                                                           # :586 (the lr = 0 step: no-op)
                update()
        losses.append(loss.item())                                                        # :604
        count += 1
        log("Client:", wrapper.args.cid, "\tStd:", np.std(losses[-num_pre_loss:]), "\tALA epochs:", count,
            wrapper.start_phase)
        if not wrapper.start_phase:                                                       # :611-613
            break
        if len(losses) > num_pre_loss and np.std(losses[-num_pre_loss:]) < threshold:    # :615-619
            break
    wrapper.start_phase = False
    for p, t in zip(params_p, temp_params_p):                                            # :623-624
        p.data = t.data.clone()


def install_set_weights(flower_common, use_graph=False, keep_weights=False):
    """Replace flower_common.MyModel.set_weights for the FedICRA strategy (other strategies keep the original)."""
    original = flower_common.MyModel.set_weights

    def set_weights(self, weights, config):
        if self.args.strategy in ["FedICRA"]:
            return adaptive_set_weights(self, weights, config, use_graph=use_graph, keep_weights=keep_weights)
        return original(self, weights, config)

    flower_common.MyModel.set_weights = set_weights
    return original
