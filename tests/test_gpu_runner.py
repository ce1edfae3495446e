"""fedicra_b200/runner.py: the eager step, its CUDA-graph replay and the pipelined host-to-host runner agree."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(seed, B=3, C=2, H=64, W=80, Cin=1):
    g = torch.Generator().manual_seed(seed)
    label = torch.full((B, H, W), C, dtype=torch.uint8)
    scribble = torch.rand(B, H, W, generator=g) < 0.05
    label[scribble] = torch.randint(0, C, (int(scribble.sum()),), generator=g, dtype=torch.uint8)
    return {"image": (torch.rand(B, Cin, H, W, generator=g) * 255).round() / 255,
            "logits": torch.randn(B, C, H, W, generator=g),
            "aux1": torch.randn(B, C, H // 4, W // 4, generator=g),
            "aux2": torch.randn(B, C, H // 2, W // 2, generator=g),
            "aux3": torch.randn(B, C, H, W, generator=g), "label": label}


def _close(a, b, tol=1e-5):
    scale = float(b.abs().max().clamp_min(1e-30))
    return float((a.double().cpu() - b.double().cpu()).abs().max()) <= tol * scale


@pytest.mark.parametrize("overlap", ["1", "0"])
def test_graph_replay_equals_eager_step(overlap, monkeypatch):
    from fedicra_b200 import ops
    from fedicra_b200.runner import GraphedLossStack, LossStackStep
    monkeypatch.setenv("FEDICRA_OVERLAP", overlap)
    step = LossStackStep(2, 64, 80)
    a, b = ({k: v.cuda() for k, v in _inputs(s).items()} for s in (1, 2))
    loss_a, grads_a = step(a)
    loss_b, grads_b = step(b)
    graphed = GraphedLossStack(step, a)
    # the library's scratch buffers are cached per stream: warm-up and capture share one, so none was created (and
    # zero-filled) inside the capture
    streams = {}
    for (tag, dev, stream) in ops._WORKSPACES:
        streams.setdefault((tag, dev), set()).add(stream)
    for src, (loss, grads) in ((b, (loss_b, grads_b)), (a, (loss_a, grads_a)), (b, (loss_b, grads_b))):
        got_loss, got_grads = graphed(src)
        assert _close(got_loss, loss)
        for g, w in zip(got_grads, grads):
            assert _close(g, w)
    graphed.reset()


@pytest.mark.parametrize("copy_gradients", [True, False])
def test_pipelined_runner_returns_every_batch_in_order(copy_gradients):
    from fedicra_b200.runner import LossStackStep, PipelinedLossStack
    step = LossStackStep(2, 64, 80)
    batches = [_inputs(10 + s) for s in range(5)]
    want = [step({k: v.cuda() for k, v in b.items()}) for b in batches]
    pinned = [{k: v.pin_memory() for k, v in b.items()} for b in batches]
    pipe = PipelinedLossStack(step, {k: v.cuda() for k, v in batches[0].items()}, depth=2, copy_gradients=copy_gradients)
    assert pipe.h2d_bytes == sum(v.numel() * v.element_size() for v in batches[0].values())
    n = 0
    for (loss, grads), (w_loss, w_grads) in zip(pipe.run(pinned), want):
        assert abs(loss - float(w_loss)) <= 1e-5 * abs(float(w_loss))
        if copy_gradients:
            for g, w in zip(grads, w_grads):
                assert not g.is_cuda and _close(g, w)
        n += 1
    assert n == len(batches)
    with pytest.raises(RuntimeError):
        for b in pinned[:3]:
            pipe.submit(b)      # a third submit without a result(): both slots are in flight
    while pipe.pending:
        pipe.result()
    pipe.reset()


def test_scratch_buffers_are_not_created_inside_a_capture():
    from fedicra_b200 import ops
    from fedicra_b200.runner import GraphedLossStack, LossStackStep
    step = LossStackStep(2, 64, 80)
    data = {k: v.cuda() for k, v in _inputs(3).items()}
    before = set(ops._WORKSPACES)
    graphed = GraphedLossStack(step, data)
    new = set(ops._WORKSPACES) - before
    per_tag = {}
    for (tag, dev, stream) in new:
        per_tag.setdefault((tag, dev), set()).add(stream)
    assert all(len(s) == 1 for s in per_tag.values()), per_tag   # one stream: warm-up and capture
    graphed.reset()
