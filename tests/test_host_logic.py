"""Host-side logic that needs no GPU: ALA oracle, FedAvg over gloo (world size 2), drop-in registration, sharding."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import ala_cpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_ala_oracle_equals_torch_expressions():
    g = torch.Generator().manual_seed(0)
    p, s, grad = (torch.randn(1000, generator=g) for _ in range(3))
    w = torch.rand(1000, generator=g)
    w_ref = torch.clamp(w - 1.0 * (grad * (s - p)), 0, 1)  # flower_common.py:596-597
    t_ref = p + (s - p) * w_ref                            # flower_common.py:601-602
    w_o, t_o = ala_cpu.ala_update(p.numpy(), s.numpy(), grad.numpy(), w.numpy(), 1.0)
    assert np.array_equal(w_o, w_ref.numpy()) and np.array_equal(t_o, t_ref.numpy())


def test_fedavg_oracle_weights_by_batch_count():
    a = [np.ones(4, np.float32), np.full(2, 2.0, np.float32)]
    b = [np.full(4, 3.0, np.float32), np.full(2, 6.0, np.float32)]
    out = ala_cpu.fedavg([a, b], [1, 3])
    assert np.allclose(out[0], 2.5) and np.allclose(out[1], 5.0)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fedavg_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fedicra_b200.aggregation import fedavg_allreduce
    from fedicra_b200.sharding import shard_range, sharded_masked_l1
    state = torch.full((10,), float(rank + 1))
    out = fedavg_allreduce(state, num_examples=[21, 13][rank])
    # image-sharded tree-energy normaliser: sum over ranks of (masked L1 sum, N) before the division
    lo, hi = shard_range(5, rank, world)
    num = torch.tensor(float(hi - lo))
    n = torch.tensor(2.0 * (hi - lo))
    loss = sharded_masked_l1(num, n)
    q.put((rank, out.tolist(), (lo, hi), float(loss)))
    dist.destroy_process_group()


def test_fedavg_and_sharded_normaliser_over_gloo_world_size_2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_fedavg_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    expect = (1 * 21 + 2 * 13) / 34.0
    for rank, out, rng, loss in res:
        assert np.allclose(out, expect)
        assert abs(loss - 0.5) < 1e-6
    assert [r[2] for r in res] == [(0, 3), (3, 5)]


def _flat_state_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fedicra_b200.aggregation import FlatModelState
    torch.manual_seed(rank)
    net = torch.nn.Sequential(torch.nn.Conv2d(2, 3, 3), torch.nn.BatchNorm2d(3), torch.nn.Conv2d(3, 1, 1))
    net[1].num_batches_tracked.fill_(10 * (rank + 1) + 1)
    before = [v.clone() for v in net.state_dict().values()]
    st = FlatModelState(net)
    same = all(torch.equal(a, b) for a, b in zip(before, net.state_dict().values()))   # re-pointing keeps the values
    shared = net[0].weight.data_ptr() == st.flat.data_ptr()                             # ... and the module lives in `flat`
    wire = st.get_weights()
    st.fedavg_(num_examples=[21, 13][rank])
    x = torch.ones(1, 2, 5, 5)
    q.put((rank, [a.tolist() for a in wire], [v.tolist() for v in net.state_dict().values()], same, shared,
           net.eval()(x).flatten().tolist()))
    dist.destroy_process_group()


def test_flat_model_state_fedavg_equals_the_oracle_over_gloo_world_size_2():
    """FlatModelState: 160-tensor state_dicts as one buffer (flower_common.py:488-489, 499-502); FedAvg in place, integer
    counters averaged in float64 and truncated like numpy + load_state_dict; result == oracle.ala_cpu.fedavg."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_flat_state_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    wires = [[np.asarray(a) for a in r[1]] for r in res]
    want = ala_cpu.fedavg([[a.astype(np.float32) if a.dtype.kind == "f" else a for a in w] for w in wires], [21, 13])
    for rank, _, after, same, shared, out in res:
        assert same and shared
        for got, w, src in zip(after, want, wires[0]):
            if np.asarray(src).dtype.kind == "f":
                assert np.allclose(np.asarray(got, dtype=np.float32), np.asarray(w, dtype=np.float32), rtol=1e-6, atol=1e-7)
            else:
                assert int(np.asarray(got)) == int((11 * 21 + 21 * 13) / 34.0)      # truncation of the float mean
    assert res[0][5] == res[1][5]                                                    # both clients hold the same model now


def test_dropin_registers_reference_module_names():
    import fedicra_b200
    saved = {k: sys.modules.get(k) for k in list(sys.modules) if k.startswith("utils") or k == "tree_filter_cuda"}
    try:
        fedicra_b200.install_dropin("modules")
        import importlib
        m = importlib.import_module("utils.gate_crf_loss")
        assert m.ModelLossSemsegGatedCRF is fedicra_b200.ModelLossSemsegGatedCRF
        t = importlib.import_module("utils.TreeEnergyLoss.kernels.lib_tree_filter.modules.tree_filter")
        assert t.TreeFilter2D is fedicra_b200.TreeFilter2D
        assert sys.modules["tree_filter_cuda"] is fedicra_b200.tree_filter_cuda
    finally:
        for k in [k for k in sys.modules if k.startswith("utils") or k == "tree_filter_cuda"]:
            del sys.modules[k]
        sys.modules.update({k: v for k, v in saved.items() if v is not None})


def test_unsupported_branches_raise_instead_of_falling_back():
    from fedicra_b200 import MinimumSpanningTree, ModelLossSemsegGatedCRF, TreeFilter2D
    crf = ModelLossSemsegGatedCRF()
    y = torch.rand(1, 2, 8, 8)
    kd = [{"weight": 1, "xy": 6, "rgb": 0.1}]
    with pytest.raises(NotImplementedError):
        crf(y, kd, 5, torch.rand(1, 1, 8, 8), 8, 8, mask_src=torch.ones(1, 1, 8, 8))
    with pytest.raises(NotImplementedError):
        crf(y, kd, 5, torch.rand(1, 1, 8, 8), 8, 8, compatibility=torch.ones(2, 2))
    with pytest.raises(AssertionError):
        crf(y, kd, 5, torch.rand(1, 1, 8, 8), 9, 8)
    with pytest.raises(RuntimeError, match="no CPU path"):
        crf(y, kd, 5, torch.rand(1, 1, 8, 8), 8, 8)
    with pytest.raises(NotImplementedError):
        MinimumSpanningTree(TreeFilter2D.norm2_distance)(torch.rand(1, 3, 4, 4), label=torch.rand(1, 1, 4, 4))
    with pytest.raises(NotImplementedError):
        TreeFilter2D(groups=2)(torch.rand(1, 2, 4, 4), torch.rand(1, 2, 4, 4), torch.zeros(1, 15, 2, dtype=torch.int32))
