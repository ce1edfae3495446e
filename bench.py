#!/usr/bin/env python
"""bench.py -- GatedCRF + multiscale TreeEnergy loss, forward + backward, images/s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload faz256|odoc256|rgb512]

A step = one pass of the loss hot path over one batch of synthetic input: from (logits, aux1..3, image, label) to
(loss, d/dlogits, d/daux1..3), exactly the calls of the reference training loop
(/root/reference/code/flower_pCE_2D_GateCRFMsacleTreeEnergyLoss_Ours.py:132-151,167) minus the network itself.

One JSON line on stdout (rank 0).  `value` = images/s with the inputs resident in HBM; `e2e` = the same step called
with HOST (pinned) buffers, host->device copies of every input and device->host copies of the loss and all gradients
inside the timed region.  `roofline` describes the dominant kernel group, `cpu_baseline` the CPU oracle timed on this
box's host cores on a bounded sample.  `--impl reference` times the reference's CPU algorithm (oracle port + the
reference's own boruvka.cpp when oracle/_ref was built) on all host threads.
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
class GpuStep:
    """The loss-stack step through the package's public, reference-signature modules."""

    def __init__(self, C, H, W, world, local_batch):
        import fedicra_b200 as fb
        self.C, self.H, self.W = C, H, W
        self.tel = fb.MScaleRecurveTreeEnergyLoss(distributed=world > 1)
        self.crf = fb.ModelLossSemsegGatedCRF()
        self.crf_scale = CRF_WEIGHT / world  # Gated CRF divides by the LOCAL N*H*W; the global batch is world x larger

    def __call__(self, d):
        leaves = [d[k].detach().requires_grad_() for k in ("logits", "aux1", "aux2", "aux3")]
        image = d["image"]
        three = image.repeat(1, 3, 1, 1) if image.shape[1] == 1 else image          # OURS.py:137-141
        roi = d["label"] == self.C                                                    # OURS.py:135
        tel = self.tel(leaves[0], three, leaves[1], leaves[2], leaves[3], roi, TREE_WEIGHT)[0]
        soft = torch.softmax(leaves[0], dim=1)                                        # OURS.py:132
        crf = self.crf(soft, KERNELS_DESC, RADIUS, image, self.H, self.W)["loss"]     # OURS.py:143-150
        loss = tel + self.crf_scale * crf                                             # OURS.py:151 (without pCE)
        loss.backward()                                                               # OURS.py:172
        return loss.detach(), [x.grad for x in leaves]


class GraphedStep:
    """The same step captured once into a CUDA graph and replayed: static input buffers, outputs in the graph's pool.
    Every kernel of the library launches on the capturing stream and none of them synchronises or allocates outside
    torch's allocator, so the step is capture-safe.  Removes the host launch cost (a few dozen library launches + ~20 ATen launches per step)."""

    def __init__(self, step, example):
        self.static_in = {k: v.clone() for k, v in example.items()}
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                step(self.static_in)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
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
             "tree_filter_backward", "band_plan", "chain_order", "chain_filter_forward", "chain_filter_backward", "channel_softmax_fwd", "channel_softmax_bwd", "masked_l1_fwd",
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


def cpu_port_step(d, C, mst_impl):
    """The CPU oracle's loss stack (reference algorithm: unfold-based Gated CRF, sequential Boruvka, restated sweeps)."""
    from oracle import losses_cpu, tree_cpu
    tree_cpu.set_mst_impl(mst_impl)
    xs = [d[k].clone().requires_grad_() for k in ("logits", "aux1", "aux2", "aux3")]
    total, _, _, _ = losses_cpu.loss_stack(xs[0], d["image"], xs[1], xs[2], xs[3], d["label"], C, TREE_WEIGHT, CRF_WEIGHT,
                                           gcrf="unfold")
    total.backward()
    return float(total.detach())


def cpu_setup():
    from oracle import tree_cpu
    tree_cpu.build(ref=True)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    tree_cpu.set_num_threads(cores)
    kind = "reference" if tree_cpu.ref_lib() is not None else "port"
    return cores, kind


def run_reference(args, wl):
    """--impl reference: the reference's CPU implementation of the path on all host threads, bounded sample per step."""
    B, Cin, C, H, W, desc = wl
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores, kind = cpu_setup()
    sample_b = min(B, 4)
    data = make_inputs(sample_b, Cin, C, H, W, 2022)
    mst_impl = "reference" if kind == "reference" else "port"
    for _ in range(args.warmup):
        cpu_port_step(data, C, mst_impl)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_port_step(data, C, mst_impl)
    dt = time.perf_counter() - t0
    value = sample_b * args.steps / dt
    sample = (f"{sample_b} of the {B} images of one batch per step; Gated CRF = unfold formulation of gate_crf_loss.py in "
              f"torch CPU, MST = {'the reference boruvka.cpp (oracle/_ref)' if kind == 'reference' else 'C port of boruvka.cpp'}"
              f", BFS/refine = C restatement, {cores} threads")
    emit(json.dumps({
        "impl": "reference", "metric": "GatedCRF+TreeEnergy loss fwd+bwd images/s", "value": value, "unit": "images/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "describes": desc, "batch_per_step": sample_b, "H": H, "W": W,
                   "classes": C, "image_channels": Cin},
        "cpu_baseline": {"value": value, "unit": "images/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def run_ours(args, wl):
    B, Cin, C, H, W, desc = wl
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path in fedicra_b200)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    from fedicra_b200 import _lib

    host = make_inputs(B, Cin, C, H, W, 2022 + rank)     # each rank owns its contiguous shard of the global batch
    pinned = {k: v.pin_memory() for k, v in host.items()}
    data = {k: v.to(dev) for k, v in host.items()}
    eager_step = GpuStep(C, H, W, world, B)
    step = eager_step
    graphed = False
    if not args.no_graph:  # N > 1: the 2-scalar NCCL all-reduce of the normaliser is captured with the rest
        try:
            step = GraphedStep(eager_step, data)
            data = step.static_in
            graphed = True
        except Exception as exc:  # capture not possible: say so and time the eager step
            print(f"[bench] CUDA graph capture failed, timing the eager step: {exc}", file=sys.stderr)
            step = eager_step
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # 2x the 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident-input timing (value)
    for _ in range(max(args.warmup, 3)):
        step(data)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    launches0 = _lib.lib.fedicra_launch_count()
    times = time_steps(lambda: step(data), args.steps, flush_buf)
    launches = _lib.lib.fedicra_launch_count() - launches0
    if graphed:  # replayed launches are not seen by the host-side counter: count one eager step instead
        launches0 = _lib.lib.fedicra_launch_count()
        eager_step(data)
        launches = (_lib.lib.fedicra_launch_count() - launches0) * args.steps
    barrier()
    total_ms = torch.tensor([sum(times)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms)

    # ---- end-to-end timing: host buffers in, loss + gradients out, copies inside the timed region
    grads_host = None

    def e2e_step():
        nonlocal grads_host
        if graphed:
            step.load(pinned)          # H2D of every input into the graph's static buffers
            loss, grads = step()
        else:
            loss, grads = step({k: v.to(dev, non_blocking=True) for k, v in pinned.items()})
        if grads_host is None:
            grads_host = [torch.empty(g.shape, dtype=g.dtype).pin_memory() for g in grads]
        for h, g in zip(grads_host, grads):
            h.copy_(g, non_blocking=True)
        return float(loss)  # D2H read of the loss; also the sync that completes the copies

    for _ in range(3):
        e2e_step()
    barrier()
    e2e_times = time_steps(e2e_step, args.steps, flush_buf)
    barrier()
    e2e_ms = torch.tensor([sum(e2e_times)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_ms = float(e2e_ms)
    clocks = sampler.stop() if sampler else None
    h2d = sum(v.numel() * v.element_size() for v in pinned.values())
    d2h = 4 + sum(g.numel() * g.element_size() for g in grads_host)

    def teardown():
        """All ranks together.  A captured graph keeps NCCL kernels of the communicator alive, so it is released first."""
        if world > 1:
            dist.barrier()
            if graphed:
                step.graph.reset()
            torch.cuda.synchronize()
            dist.destroy_process_group()

    if rank != 0:
        teardown()  # (waits for rank 0's attribution pass)
        return

    # ---- attribution + roofline of the dominant kernel group (rank 0, separate instrumented pass)
    # (rank 0 alone from here on: the attribution pass uses a step WITHOUT the cross-rank all-reduce)
    local_step = eager_step if world == 1 else GpuStep(C, H, W, 1, B)
    ops_prof = profile_ops(local_step, data, min(args.steps, 10), flush_buf)
    ab = algorithmic_bytes(Cin, C, H, W)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
    all_stages = lambda key: sum(st[key] for st in ab["stages"])
    u = ab["u"]
    group_bytes = {  # algorithmic bytes per CALL of each kernel group (all B images of the batch)
        "mst_grid": lambda: B * (ab["stages"][0]["mst_bfs"] - u),
        "mst_grid_multi": lambda: B * (all_stages("mst_bfs") - 4 * u),          # one launch for the 4 embeddings
        "bfs_forward": lambda: B * u,
        "bfs_grid": lambda: B * u,
        "bfs_euler": lambda: 4 * B * u,                                          # one call for the 4 trees
        "refine_forward": lambda: B * ab["stages"][1]["filter_fwd"],
        "refine_backward_feature": lambda: B * (2 * C + 3) * u,
        "refine_backward_weight": lambda: B * (2 * C + 2 * C) * u,
        "chain_order": lambda: 4 * B * u,                                        # one call for the 4 trees
        "chain_filter_forward": lambda: B * all_stages("filter_fwd") / 4,       # mean over the 4 trees
        "chain_filter_backward": lambda: B * all_stages("filter_bwd") / 4,
        "tree_filter_forward": lambda: B * all_stages("filter_fwd") / 4,        # mean over the 4 trees
        "tree_filter_backward": lambda: B * all_stages("filter_bwd") / 4,
        "gcrf_fwd_bwd": lambda: B * ab["gcrf"],
        "edge_weight_fwd": lambda: B * (C + 3) * u,
        "edge_weight_bwd": lambda: B * (2 * C + 4) * u,
        "band_plan": lambda: 4 * B * u / 16,
        "channel_softmax_fwd": lambda: B * 2 * C * u,
        "channel_softmax_bwd": lambda: B * 3 * C * u,
        "masked_l1_fwd": lambda: B * (2 * C + 0.25) * u,
        "masked_l1_bwd": lambda: B * (4 * C + 0.25) * u,
    }
    # DRAM bytes per launch of the kernel groups, from the committed `ncu --set full` captures (profiles/traffic.json)
    traffic_path = os.path.join(ROOT, "profiles", "traffic.json")
    traffic_tab = json.load(open(traffic_path)).get(args.workload, {}) if os.path.exists(traffic_path) else {}
    dominant = max(ops_prof, key=lambda n: ops_prof[n]["ms_per_step"]) if ops_prof else None
    roofline = None
    if dominant:
        p = ops_prof[dominant]
        nbytes = group_bytes[dominant]()
        achieved = nbytes / (p["ms_per_call"] * 1e-3) / 1e9
        roofline = {"kernel": dominant, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": traffic_tab.get(dominant), "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": nbytes, "ms_per_launch": p["ms_per_call"],
                    "share_of_step": p["ms_per_step"] / (total_ms / args.steps)}
    step_bytes = B * (ab["tree"] + ab["gcrf"])
    step_frac = step_bytes / (total_ms / args.steps * 1e-3) / 1e9 / peak

    # ---- CPU baseline on a bounded sample (rank 0, N=1 only)
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cores, kind = cpu_setup()
        sample_b = min(B, 4)
        sample = make_inputs(sample_b, Cin, C, H, W, 2022)
        impl = "reference" if kind == "reference" else "port"
        cpu_port_step(sample, C, impl)
        reps = 2
        t0 = time.perf_counter()
        for _ in range(reps):
            cpu_port_step(sample, C, impl)
        dt = (time.perf_counter() - t0) / reps
        cpu = {"value": sample_b / dt, "unit": "images/s", "cores": cores, "kind": kind,
               "sample": f"{sample_b} of the {B} images of the batch, {reps} timed passes after 1 warm-up, oracle loss stack "
                         f"(unfold Gated CRF in torch CPU; MST = {'reference boruvka.cpp' if kind == 'reference' else 'C port'}; "
                         f"C restatement of bfs/refine) on {cores} threads"}

    n_images = B * world * args.steps
    emit(json.dumps({
        "metric": "GatedCRF+TreeEnergy loss fwd+bwd images/s", "value": n_images / (total_ms * 1e-3), "unit": "images/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "describes": desc, "batch_per_gpu": B, "global_batch": B * world, "H": H,
                   "W": W, "classes": C, "image_channels": Cin, "parallelism": f"image-sharded x{world}",
                   "l2": "256 MiB device memset between timed steps (not timed); per-step CUDA events on the current stream",
                   "collective": "2-scalar all-reduce of the tree-energy normaliser per step" if world > 1 else "none",
                   "cuda_graph": graphed},
        "e2e": {"value": n_images / (e2e_ms * 1e-3), "unit": "images/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "step_roofline": {"bound": "hbm", "algorithmic_bytes_per_step": step_bytes, "frac": step_frac, "peak": peak,
                          "note": "whole step against HBM; the level-synchronous sweeps are dependency-latency bound"},
        "kernel_groups": ops_prof,
        "cpu_baseline": cpu,
    }))
    teardown()


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
    ap.add_argument("--workload", default="faz256", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="time the eager step instead of a CUDA-graph replay")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
