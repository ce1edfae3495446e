"""Image sharding on GPU tensors: the loss stack evaluated on two shards with the cross-shard normaliser equals the
unsharded evaluation -- loss and every gradient (the collective itself is covered over gloo in
tests/test_host_logic.py and over NCCL by bench.py --gpus N)."""
import pytest
import torch

from tests.helpers import loss_inputs, rel_err

pytestmark = pytest.mark.gpu


def test_two_shards_with_a_shared_normaliser_equal_the_whole_batch(monkeypatch):
    import fedicra_b200 as fb
    from fedicra_b200 import sharding, tree_energy
    B, C, H, W = 4, 2, 96, 128
    image, logits, a1, a2, a3, label = loss_inputs(B, C, 1, H, W, "quant8", seed=21)
    three = image.repeat(1, 3, 1, 1).cuda()
    roi = (label == C).cuda()

    def run(sl, distributed):
        xs = [t[sl].clone().cuda().requires_grad_() for t in (logits, a1, a2, a3)]
        out = fb.MScaleRecurveTreeEnergyLoss(distributed=distributed)(xs[0], three[sl], xs[1], xs[2], xs[3], roi[sl], 0.1)
        return xs, out

    xs_all, out_all = run(slice(0, B), False)
    out_all[0].backward()
    # a two-rank group faked in one process: the "all-reduce" adds the other shard's (sum, N) pair
    shards = [slice(0, 2), slice(2, 4)]
    pairs = []
    for sl in shards:   # pass 1: every shard's local (sum, N)
        captured = {}
        monkeypatch.setattr(tree_energy, "_normalise",
                            lambda s, n, d: captured.update(s=s.detach().clone(), n=n.detach().clone()) or s / n.clamp(min=1))
        run(sl, True)
        pairs.append((captured["s"], captured["n"]))
    monkeypatch.undo()
    tot_s, tot_n = pairs[0][0] + pairs[1][0], pairs[0][1] + pairs[1][1]
    grads, losses = [], []
    for k, sl in enumerate(shards):   # pass 2: the sharded formula with the summed pair (sharding.sharded_masked_l1)
        monkeypatch.setattr(sharding, "sharded_masked_l1",
                            lambda s, n, group=None: (s + (tot_s - s.detach())) / tot_n.clamp(min=1))
        xs, out = run(sl, True)
        out[0].backward()
        grads.append([x.grad for x in xs])
        losses.append(out[0].detach())
        monkeypatch.undo()
    assert rel_err(losses[0].cpu(), out_all[0].detach().cpu()) < 1e-5 and rel_err(losses[1].cpu(), losses[0].cpu()) < 1e-6
    for i, name in enumerate(("logits", "aux1", "aux2", "aux3")):
        whole = xs_all[i].grad
        got = torch.cat([grads[0][i], grads[1][i]], 0)
        assert rel_err(got.cpu(), whole.cpu()) < 1e-4, name


def test_local_share_mode_with_the_count_fetched_on_a_side_stream_equals_the_plain_loss():
    """distributed + local_share: N comes from sharding.GlobalCount (side stream, started before the trees are built);
    with one rank the share is the whole loss, gradients included."""
    import fedicra_b200 as fb
    B, C, H, W = 3, 2, 64, 96
    image, logits, a1, a2, a3, label = loss_inputs(B, C, 1, H, W, "quant8", seed=4)
    three = image.repeat(1, 3, 1, 1).cuda()
    roi = (label == C).cuda()
    res = []
    for kw in ({}, {"distributed": True, "local_share": True}):
        xs = [t.clone().cuda().requires_grad_() for t in (logits, a1, a2, a3)]
        out = fb.MScaleRecurveTreeEnergyLoss(**kw)(xs[0], three, xs[1], xs[2], xs[3], roi, 0.1)
        out[0].backward()
        torch.cuda.synchronize()
        res.append((out[0].detach().cpu(), [x.grad.cpu() for x in xs]))
    assert rel_err(res[1][0], res[0][0]) < 1e-6
    for a, b in zip(res[1][1], res[0][1]):
        assert rel_err(a, b) < 1e-5
