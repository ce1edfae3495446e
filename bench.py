#!/usr/bin/env python
"""bench.py -- GatedCRF + multiscale TreeEnergy loss, forward + backward, images/s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload faz256|odoc256|rgb512]

A step = one pass of the loss hot path over one batch of synthetic input: from (logits, aux1..3, image, label) to
(loss, d/dlogits, d/daux1..3), exactly the calls of the reference training loop
(/root/reference/code/flower_pCE_2D_GateCRFMsacleTreeEnergyLoss_Ours.py:132-151,167) minus the network itself.

One JSON line on stdout (rank 0).  `value` = images/s with the inputs resident in HBM (CUDA-graph replay; `eager` = the
same step launched eagerly, what a loop that is not captured gets); `e2e` = the same step called with HOST (pinned)
buffers through fedicra_b200.PipelinedLossStack -- host->device copies of every input and device->host copies of the
loss and all gradients inside the timed region, double-buffered (`e2e.serial_ms_per_step` = the same without overlap).
`roofline` describes the dominant kernel group, `cpu_baseline` the reference's CPU path timed on this box's host cores,
`reference_gpu` the reference's own modules on its own CUDA kernels (rebuilt for sm_100a, oracle/_ref) on the same GPU,
`extra_workloads` the other BASELINE shapes (odoc256, rgb512) measured the same way.  `--impl reference` times the
reference's CPU path alone.  `--workload odoc256_unet` drives the reference's unmodified training loop
(MyClient._train, UNet_LC_MultiHead) on the drop-in; `--workload fedround` one federated round per rank.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch
import torch.distributed as dist
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (per-GPU batch, Cin, classes, H, W, BASELINE.json config it stands for)
    "faz256": (12, 1, 2, 256, 256, "configs[1]: MScaleTreeEnergy (4 scales) + GatedCRF, FAZ-shape batch 12x1x256x256, 2 classes"),
    "odoc256": (12, 3, 3, 256, 256, "configs[2] loss stack: ODOC-shape RGB 12x3x256x256, 3 classes"),
    "rgb512": (12, 3, 3, 512, 512, "configs[4]: 512x512 RGB, 3 classes, per-GPU batch 12"),
}
KERNELS_DESC = [{"weight": 1, "xy": 6, "rgb": 0.1}]  # OURS.py:68
RADIUS = 5                                           # OURS.py:69
TREE_WEIGHT, CRF_WEIGHT = 0.1, 0.1                   # OURS.py:270, :151


def make_inputs(B, Cin, C, H, W, seed):
    """Synthetic step inputs: 8-bit-quantised image (tie-heavy like the PNGs), N(0,1) logits/heads,
    3 % scribble labels."""
    g = torch.Generator().manual_seed(seed)
    image = (torch.rand(B, Cin, H, W, generator=g) * 255).round() / 255
    logits = torch.randn(B, C, H, W, generator=g)
    aux1 = torch.randn(B, C, H // 4, W // 4, generator=g)
    aux2 = torch.randn(B, C, H // 2, W // 2, generator=g)
    aux3 = torch.randn(B, C, H, W, generator=g)
    label = torch.full((B, H, W), C, dtype=torch.uint8)
    scribble = torch.rand(B, H, W, generator=g) < 0.03
    label[scribble] = torch.randint(0, C, (int(scribble.sum()),), generator=g, dtype=torch.uint8)
    return {"image": image, "logits": logits, "aux1": aux1, "aux2": aux2, "aux3": aux3, "label": label}


# ----------------------------------------------------------------------------------------------- algorithmic bytes
def algorithmic_bytes(Cin, C, H, W):
    """Per-image algorithmic HBM bytes, SURVEY.md §8(d): u = 4*H*W; tree stage = (6C + 4De + 8)u (low tree 6C+3De+8),
    Gated CRF = (2C + Cin)*4 bytes per pixel."""
    u = 4 * H * W
    de = [3, C, C, C]
    stages = []
    for k, d in enumerate(de):
        stages.append({"mst_bfs": (d + 2) * u, "filter_fwd": (2 * C + d + 3) * u,
                       "filter_bwd": (4 * C + (d if k == 0 else 2 * d) + 3) * u})
    tree = sum(sum(s.values()) for s in stages)
    return {"u": u, "tree": tree, "gcrf": (2 * C + Cin) * u, "stages": stages}


# ----------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.lines, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        for line in self.lines:
            p = [x.strip() for x in line.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                smax.append(float(p[1]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------------- our arm
# The step itself lives in the package (fedicra_b200/runner.py): LossStackStep (eager), GraphedLossStack (CUDA-graph
# replay), PipelinedLossStack (host buffers in and out, copies overlapped with the replay).


def flush_l2(buf):
    buf.zero_()


def time_steps(fn, steps, flush_buf):
    """Per-step CUDA-event timing on the launching (current) stream; the L2 flush between steps is not timed."""
    times = []
    for _ in range(steps):
        flush_l2(flush_buf)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        times.append(a.elapsed_time(b))
    return times


def profile_ops(step, data, steps, flush_buf):
    """Second, instrumented pass: CUDA events around every C-ABI call, to attribute the step time to kernel groups."""
    from fedicra_b200 import ops
    names = ["gcrf_fwd_bwd", "mst_grid", "mst_grid_multi", "bfs_forward", "bfs_grid", "bfs_euler", "edge_weight_fwd",
             "edge_weight_bwd", "refine_forward", "refine_backward_feature", "refine_backward_weight", "tree_filter_forward",
             "tree_filter_backward", "band_plan", "chain_order", "chain_order_weights", "chain_filter_forward", "chain_filter_backward", "channel_softmax_fwd", "channel_softmax_bwd", "masked_l1_fwd",
             "masked_l1_bwd"]
    records, originals = {n: [] for n in names}, {}

    def wrap(name, fn):
        def inner(*a, **k):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn(*a, **k)
            e1.record()
            records[name].append((e0, e1))
            return out
        return inner

    for n in names:
        originals[n] = getattr(ops, n)
        setattr(ops, n, wrap(n, originals[n]))
    try:
        for _ in range(steps):
            flush_l2(flush_buf)
            step(data)
        torch.cuda.synchronize()
    finally:
        for n in names:
            setattr(ops, n, originals[n])
    out = {}
    for n, evs in records.items():
        if evs:
            ms = [a.elapsed_time(b) for a, b in evs]
            out[n] = {"calls_per_step": len(ms) / steps, "ms_per_call": sum(ms) / len(ms), "ms_per_step": sum(ms) / steps}
    return out


# ----------------------------------------------------------------------------------------------- reference arms
def isolated_call_times(data, C, H, W, flush_buf, reps=10, n=5):
    """Device time per CALL of the four heavy C-ABI calls of a step, each replayed alone from a CUDA graph (`reps` calls
    back to back, L2 flushed before every replay, CUDA events on the launching stream): no host launch gaps, no other
    kernels in between.  Inputs are the step's own image and decoder heads."""
    from fedicra_b200 import ops
    image = data["image"]
    three = image.repeat(1, 3, 1, 1) if image.shape[1] == 1 else image
    embs = [three.contiguous()] + [F.interpolate(data[k], size=(H, W), mode="bilinear", align_corners=False).contiguous()
                                   for k in ("aux1", "aux2", "aux3")]
    B, V = image.shape[0], H * W
    inv = [50.0, 1.0, 1.0, 1.0]
    _, mask = ops.mst_grid_multi(embs, want_mask=True, want_edges=False)
    si, sp, sc, gs, ew = ops.chain_order_weights(mask, embs, inv, H, W)
    cut = lambda a: a[B:2 * B]                                  # the first decoder-head tree
    feat = torch.softmax(data["logits"], 1).reshape(B, C, V).contiguous()
    gout = torch.randn_like(feat)
    o = [cut(a) for a in (si, sp, sc, gs)]
    w1 = cut(ew).clone()
    _, saved = ops.chain_filter_forward(feat, w1, *o)
    calls = {
        "mst_grid_multi": lambda: ops.mst_grid_multi(embs, want_mask=True, want_edges=False),
        "chain_order_weights": lambda: ops.chain_order_weights(mask, embs, inv, H, W),
        "chain_filter_forward": lambda: ops.chain_filter_forward(feat, w1, *o),
        "chain_filter_backward": lambda: ops.chain_filter_backward(gout, w1, *o, saved, True),
    }
    out = {}
    for name, fn in calls.items():
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):   # same stream as the warm-up: scratch buffers are cached per stream
            for _ in range(reps):
                fn()
        ts = time_steps(g.replay, n, flush_buf)
        out[name] = sorted(ts)[len(ts) // 2] / reps
        g.reset()
    return out


def cpu_reference_setup():
    """The reference's CPU path.  Where oracle/_ref was built (it travels to the GPU box) this is the reference's OWN
    Python, unmodified -- utils/gate_crf_loss.py, flower_common.MScaleRecurveTreeEnergyLoss, modules/tree_filter.py,
    functions/*.py -- on CPU tensors, with `tree_filter_cuda` served by the reference's boruvka.cpp (compiled as is)
    and the C restatement of its bfs / refine kernels (those exist only as CUDA in the reference).  Otherwise the oracle
    ports of the same files."""
    from oracle import ref_gpu, tree_cpu
    tree_cpu.build(ref=True)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    tree_cpu.set_num_threads(cores)
    have_ref_mst = tree_cpu.ref_lib() is not None
    tree_cpu.set_mst_impl("reference" if have_ref_mst else "port")
    if os.path.isdir(ref_gpu.PY_DIR) and have_ref_mst:
        ref_gpu.install("oracle")
        fc, gcrf = ref_gpu.flower_common(), ref_gpu.gate_crf_loss()
        tel_mod, crf_mod = fc.MScaleRecurveTreeEnergyLoss(), gcrf.ModelLossSemsegGatedCRF()

        def step(d, C):
            xs = [d[k].clone().requires_grad_() for k in ("logits", "aux1", "aux2", "aux3")]
            image = d["image"]
            three = image.repeat(1, 3, 1, 1) if image.shape[1] == 1 else image
            tel = tel_mod(xs[0], three, xs[1], xs[2], xs[3], d["label"] == C, TREE_WEIGHT)[0]
            crf = crf_mod(torch.softmax(xs[0], 1), KERNELS_DESC, RADIUS, image, image.shape[2], image.shape[3])["loss"]
            total = tel + CRF_WEIGHT * crf
            total.backward()
            return float(total.detach())

        what = ("reference Python unmodified (gate_crf_loss.py, flower_common.py, modules/tree_filter.py, functions/*.py) on "
                "CPU tensors; tree_filter_cuda = reference boruvka.cpp + C restatement of the CUDA-only bfs/refine kernels")
        return step, cores, "reference", what
    from oracle import losses_cpu

    def step(d, C):
        xs = [d[k].clone().requires_grad_() for k in ("logits", "aux1", "aux2", "aux3")]
        total, _, _, _ = losses_cpu.loss_stack(xs[0], d["image"], xs[1], xs[2], xs[3], d["label"], C, TREE_WEIGHT,
                                               CRF_WEIGHT, gcrf="unfold")
        total.backward()
        return float(total.detach())

    return step, cores, "port", "oracle ports of gate_crf_loss.py / flower_common.py glue; C port of boruvka.cpp, bfs, refine"


def run_reference(args, wl):
    """--impl reference: the reference's CPU implementation of the path on all host threads, the full batch per step."""
    B, Cin, C, H, W, desc = wl
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    step, cores, kind, what = cpu_reference_setup()
    sample_b = B if H * W <= 256 * 256 else min(B, 3)     # bounded: a 512^2 image costs ~0.5 s of CPU time
    data = make_inputs(sample_b, Cin, C, H, W, 2022)
    for _ in range(min(args.warmup, 2)):
        step(data, C)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(data, C)
    dt = time.perf_counter() - t0
    value = sample_b * args.steps / dt
    sample = f"{sample_b} of the {B} images of one batch per step, {cores} threads; {what}"
    emit(json.dumps({
        "impl": "reference", "metric": "GatedCRF+TreeEnergy loss fwd+bwd images/s", "value": value, "unit": "images/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": min(args.warmup, 2), "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "describes": desc, "batch_per_gpu": sample_b, "global_batch": sample_b,
                   "H": H, "W": W, "classes": C, "image_channels": Cin},
        "cpu_baseline": {"value": value, "unit": "images/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def reference_gpu_block(wl, data, steps):
    """The reference's own modules on its own CUDA kernels, same GPU, same inputs: ModelLossSemsegGatedCRF (torch-CUDA
    unfold algorithm) + MScaleRecurveTreeEnergyLoss on lib_tree_filter rebuilt for sm_100a (oracle/_ref; the shipped
    wheel cannot load).  Eager, like the reference's loop."""
    from oracle import ref_gpu
    B, Cin, C, H, W, _ = wl
    block = {"wheel_load": ref_gpu.wheel_load_status()}
    if not ref_gpu.available():
        block["unavailable"] = "oracle/_ref was not built (needs /root/reference at build time)"
        return block
    try:
        from fedicra_b200.runner import LossStackStep
        ref_gpu.install("reference")
        fc, gcrf = ref_gpu.flower_common(), ref_gpu.gate_crf_loss()
        step = LossStackStep(C, H, W, TREE_WEIGHT, CRF_WEIGHT, tel=fc.MScaleRecurveTreeEnergyLoss(),
                             crf=gcrf.ModelLossSemsegGatedCRF())
        for _ in range(2):
            step(data)
        torch.cuda.synchronize()
        ts = []
        for _ in range(steps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            step(data)
            b.record()
            b.synchronize()
            ts.append(a.elapsed_time(b))
        ms = sum(ts) / len(ts)
        block.update({"value": B / (ms * 1e-3), "unit": "images/s", "ms_per_step": ms, "steps": steps,
                      "what": "reference .cu rebuilt for sm_100a, THC include removed; reference Python unmodified; eager; "
                              "its mst_forward runs boruvka.cpp on host threads (mst.cu:93-112)"})
    except Exception as exc:  # a reference failure must not take the product's line with it
        block["unavailable"] = f"{type(exc).__name__}: {exc}"[:300]
    return block


# ----------------------------------------------------------------------------------------------- our arm, one workload
def group_bytes_table(B, C, ab):
    """Algorithmic bytes per CALL of each kernel group (all B images of the batch)."""
    all_stages = lambda key: sum(st[key] for st in ab["stages"])
    u = ab["u"]
    return {
        "mst_grid": B * (ab["stages"][0]["mst_bfs"] - u),
        "mst_grid_multi": B * (all_stages("mst_bfs") - 4 * u),          # one launch for the 4 embeddings
        "bfs_forward": B * u, "bfs_grid": B * u,
        "bfs_euler": 4 * B * u, "chain_order": 4 * B * u,                 # one call for the 4 trees
        "chain_order_weights": 4 * B * u + B * (3 + 3 * C + 4) * u,       # ... plus the 4 edge-weight builds
        "refine_forward": B * ab["stages"][1]["filter_fwd"],
        "refine_backward_feature": B * (2 * C + 3) * u,
        "refine_backward_weight": B * (2 * C + 2 * C) * u,
        "tree_filter_forward": B * all_stages("filter_fwd") / 4,        # mean over the 4 trees
        "tree_filter_backward": B * all_stages("filter_bwd") / 4,
        "chain_filter_forward": B * all_stages("filter_fwd") / 4,
        "chain_filter_backward": B * all_stages("filter_bwd") / 4,
        "gcrf_fwd_bwd": B * ab["gcrf"],
        "edge_weight_fwd": B * (C + 3) * u, "edge_weight_bwd": B * (2 * C + 4) * u,
        "band_plan": 4 * B * u / 16,
        "channel_softmax_fwd": B * 2 * C * u, "channel_softmax_bwd": B * 3 * C * u,
        "masked_l1_fwd": B * (2 * C + 0.25) * u, "masked_l1_bwd": B * (4 * C + 0.25) * u,
    }


def measure_workload(name, steps, warmup, dev, rank, world, use_graph, flush_buf, attribute=True, sampler_rank=None):
    """Times one workload on all ranks; returns a dict on every rank (attribution only on rank 0)."""
    from fedicra_b200 import _lib
    from fedicra_b200.runner import GraphedLossStack, LossStackStep, PipelinedLossStack
    B, Cin, C, H, W, desc = WORKLOADS[name]
    host = make_inputs(B, Cin, C, H, W, 2022 + rank)     # each rank owns its contiguous shard of the global batch
    pinned = {k: v.pin_memory() for k, v in host.items()}
    data = {k: v.to(dev) for k, v in host.items()}
    eager_step = LossStackStep(C, H, W, TREE_WEIGHT, CRF_WEIGHT, distributed=world > 1, world=world)
    step, graphed = eager_step, False
    if use_graph:  # N > 1: the 2-scalar NCCL all-reduce of the normaliser is captured with the rest
        try:
            step = GraphedLossStack(eager_step, data)
            data = step.static_in
            graphed = True
        except Exception as exc:  # capture not possible: say so and time the eager step
            print(f"[bench] CUDA graph capture failed, timing the eager step: {exc}", file=sys.stderr)
            step = eager_step

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def rank_max(ms):
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    # ---- resident-input timing (value)
    for _ in range(max(warmup, 3)):
        step(data)
    barrier()
    sampler = ClockSampler(dev.index) if (sampler_rank == rank and os.environ.get("BENCH_NO_SAMPLER") != "1") else None
    times = time_steps(lambda: step(data), steps, flush_buf)
    barrier()
    total_ms = rank_max(sum(times))
    if rank == 0:
        print(f"[bench] {name}: per-step ms min {min(times):.4f} median {statistics.median(times):.4f} max {max(times):.4f}", file=sys.stderr)
    launches0 = _lib.lib.fedicra_launch_count()
    # ---- the same step launched eagerly (what the reference's loop, which is not captured, gets)
    n_eager = min(steps, 10)
    eager_step(data)
    launches_per_step = _lib.lib.fedicra_launch_count() - launches0
    barrier()
    eager_ms = rank_max(sum(time_steps(lambda: eager_step(data), n_eager, flush_buf))) / n_eager
    barrier()

    # ---- end to end: host buffers in, loss + gradients out, every copy inside the timed region
    res = {"name": name, "desc": desc, "graphed": graphed, "ms_per_step": total_ms / steps,
           "value": B * world * steps / (total_ms * 1e-3), "eager_ms_per_step": eager_ms,
           "launches_per_step": int(launches_per_step), "shape": (B, Cin, C, H, W)}
    h2d = sum(v.numel() * v.element_size() for v in pinned.values())
    grads_host = None

    def serial_e2e():
        nonlocal grads_host
        if graphed:
            step.load(pinned)
            loss, grads = step()
        else:
            loss, grads = step({k: v.to(dev, non_blocking=True) for k, v in pinned.items()})
        if grads_host is None:
            grads_host = [torch.empty(g.shape, dtype=g.dtype).pin_memory() for g in grads]
        for h, g in zip(grads_host, grads):
            h.copy_(g, non_blocking=True)
        return float(loss)  # D2H read of the loss; also the sync that completes the copies

    for _ in range(3):
        serial_e2e()
    barrier()
    serial_ms = rank_max(sum(time_steps(serial_e2e, min(steps, 10), flush_buf))) / min(steps, 10)
    d2h = 4 + sum(g.numel() * g.element_size() for g in grads_host)
    e2e_ms, pipelined, loss_only_ms = serial_ms, False, None
    if graphed:
        for copy_grads in (True, False):
            try:
                pipe = PipelinedLossStack(eager_step, data, depth=2, copy_gradients=copy_grads)
                for _ in pipe.run([pinned] * 4):
                    pass
                barrier()
                flush_l2(flush_buf)
                torch.cuda.synchronize()
                t0 = time.perf_counter()          # host clock: the pipeline spans three streams
                for loss, grads in pipe.run([pinned] * steps):
                    pass
                torch.cuda.synchronize()
                ms = rank_max((time.perf_counter() - t0) * 1e3) / steps
                if copy_grads:
                    e2e_ms, pipelined = ms, True
                else:
                    loss_only_ms = ms
                barrier()
                pipe.reset()
            except Exception as exc:
                print(f"[bench] pipelined end-to-end run failed, reporting the serial one: {exc}", file=sys.stderr)
    res["e2e"] = {"value": B * world / (e2e_ms * 1e-3), "unit": "images/s", "h2d_bytes_per_step": h2d,
                  "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms, "serial_ms_per_step": serial_ms,
                  "pipelined": pipelined,
                  "loss_only": None if loss_only_ms is None else {
                      "value": B * world / (loss_only_ms * 1e-3), "ms_per_step": loss_only_ms, "d2h_bytes_per_step": 4,
                      "note": "same pipeline, every input from the host every step, only the loss read back (in a "
                              "training loop the gradients stay on the device)"},
                  "how": "fedicra_b200.PipelinedLossStack: two graphed copies of the step, H2D of step k+1 and D2H of step "
                         "k-1 on side streams under the replay of step k; host wall clock over all steps" if pipelined
                         else "serial: H2D, replay, D2H on one stream"}
    res["clocks"] = sampler.stop() if sampler else None
    res["_objects"] = (eager_step, step, data)

    if rank == 0 and attribute:
        local_step = eager_step if world == 1 else LossStackStep(C, H, W, TREE_WEIGHT, CRF_WEIGHT)
        ops_prof = profile_ops(local_step, data, min(steps, 10), flush_buf)
        ab = algorithmic_bytes(Cin, C, H, W)
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (of measured)"
        else:
            peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
        gb = group_bytes_table(B, C, ab)
        traffic_path = os.path.join(ROOT, "profiles", "traffic.json")
        traffic_tab = json.load(open(traffic_path)).get(name, {}) if os.path.exists(traffic_path) else {}
        # the four heavy calls once more, each alone in a graph: their device time without the instrumented pass's host gaps
        iso = {}
        try:
            if os.environ.get("FEDICRA_FILTER", "chain") == "chain" and H * W <= (1 << 18) and C <= 4:
                iso = isolated_call_times(data, C, H, W, flush_buf)
        except Exception as exc:
            print(f"[bench] isolated timing of the heavy calls failed: {exc}", file=sys.stderr)
        for n_, ms in iso.items():
            if n_ in ops_prof:
                ops_prof[n_]["ms_per_call_isolated"] = ms
        per_step = lambda n_: ops_prof[n_]["calls_per_step"] * ops_prof[n_].get("ms_per_call_isolated", ops_prof[n_]["ms_per_call"])
        dominant = max(ops_prof, key=per_step) if ops_prof else None
        if dominant:
            p = ops_prof[dominant]
            ms = p.get("ms_per_call_isolated", p["ms_per_call"])
            achieved = gb[dominant] / (ms * 1e-3) / 1e9
            res["roofline"] = {"kernel": dominant, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                               "frac": achieved / peak, "traffic": traffic_tab.get(dominant), "peak_source": peak_src,
                               "algorithmic_bytes_per_launch": gb[dominant], "ms_per_launch": ms,
                               "timed": "alone, CUDA-graph replay of 10 calls, L2 flushed before each replay" if "ms_per_call_isolated" in p
                                        else "inside the instrumented eager pass (includes host launch gaps)",
                               "share_of_step": per_step(dominant) / (total_ms / steps)}
        res["heavy_calls"] = {n_: {"ms_per_call": ms, "calls_per_step": ops_prof[n_]["calls_per_step"],
                                   "algorithmic_bytes_per_call": gb[n_], "GBps": gb[n_] / (ms * 1e-3) / 1e9,
                                   "frac_of_hbm_peak": gb[n_] / (ms * 1e-3) / 1e9 / peak, "traffic": traffic_tab.get(n_)}
                              for n_, ms in iso.items() if n_ in ops_prof}
        step_bytes = B * (ab["tree"] + ab["gcrf"])
        res["step_roofline"] = {"bound": "hbm", "algorithmic_bytes_per_step": step_bytes,
                                "frac": step_bytes / (total_ms / steps * 1e-3) / 1e9 / peak, "peak": peak,
                                "note": "whole step against HBM; the tree stages are dependency-latency bound"}
        res["kernel_groups"] = ops_prof
    return res


def release(res, world):
    """All ranks together.  A captured graph keeps NCCL kernels of the communicator alive, so it is released first."""
    _, step, _ = res.pop("_objects")
    if world > 1:
        dist.barrier()
    if res["graphed"]:
        step.reset()
    torch.cuda.synchronize()


def run_ours(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path in fedicra_b200)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # 2x the 126 MB L2
    use_graph = not args.no_graph
    main = measure_workload(args.workload, args.steps, args.warmup, dev, rank, world, use_graph, flush_buf,
                            sampler_rank=0)
    B, Cin, C, H, W = main["shape"]
    ref_gpu_blk, cpu = None, None
    if rank == 0 and world == 1 and not args.no_reference_gpu:
        ref_gpu_blk = reference_gpu_block(WORKLOADS[args.workload], main["_objects"][2], 3)
    release(main, world)
    extras = {}
    if not args.no_extra_workloads:
        for name in ("odoc256", "rgb512"):
            if name == args.workload:
                continue
            r = measure_workload(name, max(3, args.steps // 4), 3, dev, rank, world, use_graph, flush_buf)
            release(r, world)
            keep = {"describes": r["desc"], "value": r["value"], "unit": "images/s", "ms_per_step": r["ms_per_step"],
                    "eager_ms_per_step": r["eager_ms_per_step"], "e2e": r["e2e"], "steps": max(3, args.steps // 4)}
            if "roofline" in r:
                keep.update({"roofline": r["roofline"], "step_roofline": r["step_roofline"], "heavy_calls": r.get("heavy_calls"),
                             "kernel_groups": {k: round(v["ms_per_step"], 4) for k, v in r["kernel_groups"].items()}})
            extras[name] = keep
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return

    # ---- CPU baseline (rank 0, N=1 only): the reference's CPU path on this box's host cores, bounded sample
    if world == 1 and not args.no_cpu_baseline:
        cstep, cores, kind, what = cpu_reference_setup()
        sample_b = B if H * W <= 256 * 256 else min(B, 3)
        sample = make_inputs(sample_b, Cin, C, H, W, 2022)
        cstep(sample, C)
        reps = 2
        t0 = time.perf_counter()
        for _ in range(reps):
            cstep(sample, C)
        dt = (time.perf_counter() - t0) / reps
        cpu = {"value": sample_b / dt, "unit": "images/s", "cores": cores, "kind": kind,
               "sample": f"{sample_b} of the {B} images of the batch, {reps} timed passes after 1 warm-up, {cores} threads; {what}"}

    emit(json.dumps({
        "metric": "GatedCRF+TreeEnergy loss fwd+bwd images/s", "value": main["value"], "unit": "images/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": main["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "describes": main["desc"], "batch_per_gpu": B, "global_batch": B * world,
                   "H": H, "W": W, "classes": C, "image_channels": Cin, "parallelism": f"image-sharded x{world}",
                   "l2": "256 MiB device memset between timed steps (not timed); per-step CUDA events on the current stream",
                   "collective": "2-scalar all-reduce of the tree-energy normaliser per step" if world > 1 else "none",
                   "cuda_graph": main["graphed"]},
        "eager": {"value": B * world / (main["eager_ms_per_step"] * 1e-3), "unit": "images/s",
                  "ms_per_step": main["eager_ms_per_step"],
                  "note": "the same step without CUDA-graph capture (host launches included), device-timed"},
        "e2e": main["e2e"],
        "gpu_launches": main["launches_per_step"] * args.steps,
        "clocks": main["clocks"],
        "roofline": main.get("roofline"),
        "step_roofline": main.get("step_roofline"),
        "kernel_groups": main.get("kernel_groups"),
        "heavy_calls": main.get("heavy_calls"),
        "cpu_baseline": cpu,
        "reference_gpu": ref_gpu_blk,
        "extra_workloads": extras,
    }))


# ----------------------------------------------------------------------------------------------- configs[2]: U-Net loop
def run_unet(args):
    """BASELINE.json configs[2]: the reference's own training loop -- MyClient._train, UNet_LC_MultiHead fwd+bwd, pCE +
    tree energy + Gated CRF + loss_lc, AdamW -- unchanged (oracle/_ref/py), eager, on the module-level drop-in.
    `--impl reference` runs the same loop on the reference's own CUDA kernels (rebuilt for sm_100a)."""
    from oracle import ref_gpu, train_harness
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if not ref_gpu.available():
        emit(json.dumps({"impl": args.impl, "unavailable": "oracle/_ref was not built (needs /root/reference at build time)"}))
        return
    backend = "reference" if args.impl == "reference" else "modules"
    iters = max(args.steps, 4)
    res = train_harness.run(backend, "odoc", iters, 3, n_batches=2, batch_size=12, img_size=256,
                            warmup_iters=max(args.warmup, 4))
    line = {"metric": "unet_lc_multihead training iteration (fwd + loss stack + bwd + AdamW) images/s",
            "value": res["images_per_s"], "unit": "images/s", "n_gpus": 1, "steps": iters, "warmup": max(args.warmup, 4),
            "ms_per_step": 1e3 * res["seconds"] / iters, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "odoc256_unet", "describes": "configs[2]: full FedICRA loss stack on ODOC-shape RGB "
                       "12x3x256x256, 3 classes, unet_lc_multihead fwd+bwd, the reference's MyClient._train unchanged, eager, "
                       "host wall clock including the loop's .item() syncs", "backend": backend,
                       "gated_crf_module": res["gated_crf_module"], "tree_filter_module": res["tree_filter_module"]},
            "losses": res["per_iter"][:4]}
    if args.impl == "reference":
        line["impl"] = "reference"
    else:   # the loss stack alone on the same shapes, and the same loop on the reference's kernels in a child process
        from fedicra_b200.runner import LossStackStep
        dev = torch.device("cuda", 0)
        data = {k: v.to(dev) for k, v in make_inputs(12, 3, 3, 256, 256, 2022).items()}
        step = LossStackStep(3, 256, 256, TREE_WEIGHT, CRF_WEIGHT)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        for _ in range(3):
            step(data)
        loss_ms = sum(time_steps(lambda: step(data), 10, flush)) / 10
        line["loss_stack"] = {"eager_ms_per_step": loss_ms, "share_of_iteration": loss_ms / line["ms_per_step"]}
        try:
            r = subprocess.run([sys.executable, "-m", "oracle.train_harness", "--backend", "reference", "--img-class",
                                "odoc", "--iters", str(iters), "--warmup-iters", str(max(args.warmup, 4))], cwd=ROOT,
                               capture_output=True, text=True, timeout=900)
            ref = json.loads(r.stdout.strip().splitlines()[-1])
            line["reference_gpu"] = {"value": ref["images_per_s"], "unit": "images/s",
                                     "ms_per_step": 1e3 * ref["seconds"] / iters, "wheel_load": ref_gpu.wheel_load_status(),
                                     "what": "the same loop on the reference's own modules and CUDA kernels (rebuilt for sm_100a)"}
        except Exception as exc:
            line["reference_gpu"] = {"unavailable": f"{type(exc).__name__}: {exc}"[:300]}
    emit(json.dumps(line))


# ----------------------------------------------------------------------------------------------- configs[3]: federated round
def run_fedround(args):
    """BASELINE.json configs[3]: one FedICRA round per rank = client (torchrun, 5 ranks for the headline): `rep_iters` = 3
    loss-stack steps of configs[1], one pass of the element-wise adaptive aggregation over the decoder (42 tensors,
    633 010 elements when the reference network is available), one FedAvg of the whole flat state over NCCL.  During
    warm-up every rank checks the all-reduced state against oracle.ala_cpu.fedavg on the gathered inputs."""
    from fedicra_b200 import FlatModelState, fused_ala_update
    from fedicra_b200.runner import GraphedLossStack, LossStackStep
    from oracle import ala_cpu, ref_gpu
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    dist.init_process_group("nccl", device_id=dev, rank=rank, world_size=world)
    B, Cin, C, H, W, _ = WORKLOADS["faz256"]
    rep_iters = 3
    batches = [21, 13, 17, 59, 3]                      # len(trainloader) of the five FAZ clients
    n_i = batches[rank % len(batches)]
    torch.manual_seed(100 + rank)
    if ref_gpu.available():
        from oracle import train_harness
        train_harness.install("modules")
        from networks.net_factory import net_factory
        net = net_factory(train_harness.make_args("faz", cid=rank % 5), net_type="unet_lc_multihead", in_chns=1, class_num=2)
        model_name = "UNet_LC_MultiHead (reference networks/unet.py:701-722)"
    else:
        net = torch.nn.Sequential(*[torch.nn.Sequential(torch.nn.Conv2d(64, 64, 3), torch.nn.BatchNorm2d(64)) for _ in range(63)]).to(dev)
        model_name = "synthetic conv/bn stack (oracle/_ref absent)"
    state = FlatModelState(net)
    ignore = ("out_conv", "up4", "up3", "up2", "up1")                               # flower_common.py:507
    params_p = [p for n, p in net.named_parameters() if any(k in n for k in ignore)] or list(net.parameters())[-42:]
    server_p = [torch.randn_like(p) for p in params_p]
    temp_p = [torch.nn.Parameter(p.detach().clone()) for p in params_p]
    for t in temp_p:
        t.grad = torch.randn_like(t) * 1e-2
    fedaa_w = [torch.ones_like(p) for p in params_p]
    data = {k: v.to(dev) for k, v in make_inputs(B, Cin, C, H, W, 2022 + rank).items()}
    step = GraphedLossStack(LossStackStep(C, H, W, TREE_WEIGHT, CRF_WEIGHT), data)

    # ---- correctness of the NCCL FedAvg against the CPU oracle (warm-up, all ranks)
    mine = [a.copy() for a in state.get_weights()]
    gathered = [None] * world
    dist.all_gather_object(gathered, (mine, n_i))
    state.fedavg_(n_i)
    want = ala_cpu.fedavg([g[0] for g in gathered], [g[1] for g in gathered])
    worst = 0.0
    for got, w in zip(state.get_weights(), want):
        if got.dtype.kind == "f":
            worst = max(worst, float(abs(got.astype("float64") - w.astype("float64")).max() / max(abs(w).max(), 1e-12)))
        else:
            assert (got == w.astype(got.dtype)).all(), "integer counters differ from the oracle"
    assert worst < 1e-5, f"NCCL FedAvg differs from the oracle: {worst}"

    def one_round(ev=None):
        for _ in range(rep_iters):
            step(data)
        fused_ala_update(params_p, server_p, temp_p, fedaa_w, 1.0, use_grad=True)
        if ev:
            ev[0].record()
        state.fedavg_(n_i)
        if ev:
            ev[1].record()

    for _ in range(max(args.warmup, 3)):
        one_round()
    dist.barrier()
    torch.cuda.synchronize()
    rounds, ar = [], []
    for _ in range(args.steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        dist.barrier()
        a.record()
        one_round(ev)
        b.record()
        b.synchronize()
        rounds.append(a.elapsed_time(b))
        ar.append(ev[0].elapsed_time(ev[1]))
    t = torch.tensor([sum(rounds), sum(ar)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.barrier()
    step.reset()
    torch.cuda.synchronize()
    dist.destroy_process_group()
    if rank != 0:
        return
    round_ms, ar_ms = float(t[0]) / args.steps, float(t[1]) / args.steps
    nbytes = state.flat.numel() * 4
    emit(json.dumps({
        "metric": "FedICRA round (rep_iters loss steps + adaptive aggregation + FedAvg) images/s",
        "value": world * B * rep_iters / (round_ms * 1e-3), "unit": "images/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round_ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "fedround", "describes": "configs[3]: one client per rank; per round rep_iters=3 loss-stack steps "
                   "(FAZ shape 12x1x256x256), one element-wise adaptive-aggregation pass, one FedAvg over NCCL",
                   "model": model_name, "state_elements": state.numel(), "float_elements": state.flat.numel(),
                   "int64_counters": state.ints.numel(), "ala_tensors": len(params_p),
                   "ala_elements": sum(p.numel() for p in params_p), "client_batches": batches[:world]},
        "collective": {"op": f"ncclAllReduce {state.flat.numel()} fp32 (+ {state.ints.numel()} counters, + 1 count)",
                       "ranks": world, "ms": ar_ms, "bytes": nbytes,
                       "bus_GBps": (2 * (world - 1) / world * nbytes / (ar_ms * 1e-3) / 1e9) if world > 1 else None,
                       "checked": f"all-reduced state == oracle.ala_cpu.fedavg of the gathered client states (max rel err {worst:.2e})"},
    }))


_REAL_STDOUT = None


def emit(line):
    """The one JSON line goes to the real stdout; everything else written to fd 1 (NCCL's version banner, library
    chatter) was redirected to stderr at start-up."""
    data = (line + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(line + "\n")
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="faz256", choices=sorted(WORKLOADS) + ["odoc256_unet", "fedround"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-reference-gpu", action="store_true")
    ap.add_argument("--no-extra-workloads", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="time the eager step instead of a CUDA-graph replay")
    args = ap.parse_args()
    if args.workload == "odoc256_unet":
        run_unet(args)
    elif args.workload == "fedround":
        run_fedround(args)
    elif args.impl == "reference":
        run_reference(args, WORKLOADS[args.workload])
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
