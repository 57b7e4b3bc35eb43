"""bench.py -- rasterise fwd+bwd throughput (Mpix/s) on the headline workload of BASELINE.json.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one forward + one backward pass of the rasterise path over one batch of synthetic
input.  Prints ONE JSON line (rank 0).  Under torchrun (N > 1) every rank renders its own batch
shard (weak scaling, no data-path collective; the `c5` shared-mesh workload adds the NCCL
allreduce of the vertex gradients).

  value         whole-job Mpix/s, inputs resident in HBM, CUDA events, max over ranks.  The geometry MOVES
                between steps (4 pre-uploaded vertex sets, 0.25 degrees apart); --static gives the
                round-1 protocol, --hint carries the visibility hint from call to call
  extra         c4 and c5 (with / without / overlapped allreduce) device-resident, and the e2e through
                the host-buffer C ABI
  e2e           same metric through the public API with pinned HOST inputs: H2D of the step's
                inputs and D2H of its results inside the timed region
  roofline      dominant kernel: algorithmic bytes per launch / its mean duration (CUDA events
                recorded inside the library around each kernel) against the measured HBM peak
  cpu_baseline  the oracle (scalar C port, oracle/golden_raster.c) on the host cores, bounded sample
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "rasterise fwd+bwd Mpix/s"
UNIT = "Mpix/s"

WORKLOADS = {
    # name: (B per GPU, H, W, C, n_lat, n_lon)  -- sphere recipe of SURVEY.md 8(d)
    "headline": dict(B=64, H=512, W=512, C=3, n_lat=224, n_lon=224,
                     desc="B=64 per GPU, 512x512x3, bumpy UV sphere F=99,904 V=49,954, per-element mesh"),
    "c4": dict(B=256, H=256, W=256, C=3, n_lat=224, n_lon=224,
               desc="B=256 per GPU, 256x256x3, sphere F=99,904"),
    "small": dict(B=4, H=128, W=128, C=3, n_lat=48, n_lon=48, desc="debug size"),
    # config c5: one heightfield mesh shared by all frames; vertex / colour gradients are folded to
    # the shared arrays and (N > 1) summed with one packed NCCL allreduce
    "c5": dict(B=4, H=1024, W=1024, C=3, hill_n=708, shared=True,
               desc="4 frames per GPU, 1024x1024x3, shared hill heightfield F=999,698 V=501,264, grads allreduced"),
    "c5small": dict(B=2, H=256, W=256, C=3, hill_n=96, shared=True, desc="debug size of c5"),
}


def build_scene(wl, B, seed_shift=0):
    import scenes
    if wl.get("shared"):
        return scenes.hill(B, wl["H"], wl["W"], wl["C"], n=wl["hill_n"], seed=seed_shift)
    return scenes.sphere(B, wl["H"], wl["W"], wl["C"], wl["n_lat"], wl["n_lon"],
                         seeds=(1, 2 + seed_shift, 3 + seed_shift, 4 + seed_shift))


def algorithmic_bytes(B, H, W, C, V, F, shared=False):
    """BASELINE.md section 3 / SURVEY.md 8(d): every input read once, every output written once;
    aux buffers and workspace not counted.  Returns per-kernel and total figures.  `shared`:
    faces and colours are broadcast over the batch (multiplicity 1), vertices are per frame."""
    img = 4 * B * H * W * C
    mc = 1 if shared else B
    mesh_v, mesh_c, mesh_f = 16 * B * V, 4 * C * mc * V, 12 * mc * F
    per_kernel = {
        "setup": mesh_v + mesh_f,                       # vertices + faces read once
        "resolve": 2 * img + mesh_c,                    # background read, pixels written, colours read
        "grad_assemble": 3 * img + mesh_f + mesh_v + mesh_c,   # grad_pixels, pixels, grad_background; faces; grad outputs
    }
    fwd = 2 * img + mesh_v + mesh_c + mesh_f
    bwd = 3 * img + mesh_v + mesh_c + mesh_f + mesh_v + mesh_c
    return per_kernel, fwd, bwd


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def count_since(self, t0):
        return sum(1 for t, _ in self.lines if t >= t0)

    def stop(self, t0=0.0, window="timed region"):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, ln in self.lines:
            if t < t0:
                continue
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm), "window": window}


def cpu_baseline_run(wl, target_seconds, threads=None):
    """Times the oracle (fwd+bwd) on a bounded sample of the workload with all host threads."""
    from oracle import golden
    threads = threads or min(golden.max_threads(), os.cpu_count() or 1)
    H, W = wl["H"], wl["W"]
    # calibrate on `threads` elements, then size the sample for ~target_seconds
    nb = threads if not wl.get("shared") else min(threads, wl["B"])
    sc = build_scene(wl, nb)
    g = np.random.RandomState(0).randn(nb, H, W, wl["C"]).astype(np.float32)
    V = sc["vertices"].shape[1]

    def once(sc, g):
        t0 = time.perf_counter()
        px, fid, gb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"], threads=threads)
        golden.backward(sc["faces"], px, g, fid, gb, V, shared_colors=bool(wl.get("shared")), threads=threads)
        return time.perf_counter() - t0

    t = once(sc, g)
    reps = max(1, int(target_seconds / max(t, 1e-3)))
    ts = [t] + [once(sc, g) for _ in range(reps - 1)]
    best = float(np.median(ts))
    return {"value": nb * H * W / best / 1e6, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "%d elements of the workload (%dx%d, F=%d), fwd+bwd, median of %d runs, %d OpenMP threads over batch "
                      "elements; scalar CPU golden rasteriser (oracle/golden_raster.c) -- the reference's OpenGL path is "
                      "unavailable on this box" % (nb, H, W, sc["faces"].shape[1], len(ts), threads),
            "seconds_per_run": best}


def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel, workload):
    """DRAM bytes of `kernel` summed over the launches of ONE step, from the committed ncu --set full
    capture of the headline workload (profiles/roofline_traffic.json, written by
    tools/ncu_summary.py), or None."""
    if workload != "headline":
        return None
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            return json.load(f).get(kernel)
    except Exception:
        return None


def ncu_pipes(kernel, workload):
    """Issue-slot, LSU-data-pipe and DRAM utilisation of `kernel`'s launches in the same committed
    capture: the secondary ceilings SURVEY.md 8(d) asks to be reported beside the HBM roofline."""
    if workload != "headline":
        return None
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            return json.load(f).get("_pipes", {}).get(kernel)
    except Exception:
        return None


def run_reference(args, wl, rank, world):
    """--impl reference: the CPU implementation of the path on the host cores.  The reference has
    no CPU code for this path and cannot be compiled here (needs TensorFlow + EGL/OpenGL), so this
    times the oracle port; rank 0 alone runs."""
    if rank != 0:
        return
    per_step = 8.0 if args.steps * 8.0 < 200 else 200.0 / max(args.steps, 1)
    for _ in range(args.warmup):
        cpu_baseline_run(wl, 0.0)
    vals, base = [], None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        base = cpu_baseline_run(wl, per_step)
        vals.append(base["value"])
    total = time.perf_counter() - t0
    v = float(np.median(vals))
    base["value"] = v
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload + ": " + wl["desc"], "note": "host-CPU run of the oracle port; a step is several runs of a bounded sample (cpu_baseline.sample): "
                                        "ms_per_step is the wall time of all of a step's runs, value comes from the median run"},
        "cpu_baseline": base,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


_ORIG_AFFINITY = None


def bind_to_gpu_numa(index):
    """Pins this process to the CPUs of the NUMA node its GPU hangs off, before any pinned host
    buffer is allocated (first touch puts the pages on that node), the way a production launcher
    binds one rank per GPU.  Returns a short description for the JSON line, or why it was skipped."""
    try:
        import torch
        import pynvml
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(index).uuid)
        h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        bus = bus.lower()
        if len(bus.split(":")[0]) == 8:            # nvml prints an 8-digit domain, sysfs a 4-digit one
            bus = bus[4:]
        base = "/sys/bus/pci/devices/" + bus
        node = open(base + "/numa_node").read().strip()
        cpulist = open(base + "/local_cpulist").read().strip()
        cpus = set()
        for part in cpulist.split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        global _ORIG_AFFINITY
        _ORIG_AFFINITY = os.sched_getaffinity(0)
        cpus &= _ORIG_AFFINITY
        if not cpus:
            return "skipped: no local CPUs in this process's affinity mask"
        os.sched_setaffinity(0, cpus)
        return "rank bound to NUMA node %s (%d CPUs) of GPU %s" % (node, len(cpus), bus)
    except Exception as e:                          # no sysfs / nvml in this container: run unbound
        return "skipped: %s" % (str(e)[:80] or type(e).__name__)


FRAMES = 4            # vertex sets of the timed loops: the geometry moves between steps
FRAME_STEP_DEG = 0.25


def frame_order(n):
    """0 1 2 3 2 1 | 0 1 ...: every step renders the neighbour of the frame before it."""
    return list(range(n)) + list(range(n - 2, 0, -1))


def build_frames(wl, B, rank):
    """Vertex sets 1..FRAMES-1 of the sphere workloads: every element's object turned a further
    FRAME_STEP_DEG degrees per frame (tests/scenes.py:sphere_frames), the way an inverse-rendering loop
    re-renders a scene whose parameters moved a little (tests/optimize_sequence.py:293-352)."""
    import scenes
    seeds = (1, 2 + 10 * rank, 3 + 10 * rank, 4 + 10 * rank)
    return [scenes.sphere_frames(B, wl["H"], wl["W"], wl["C"], wl["n_lat"], wl["n_lon"], seeds=seeds, frame=k,
                                 step_deg=FRAME_STEP_DEG, vertices_only=True) for k in range(1, FRAMES)]


def time_steps(step, steps, warmup, barrier, max_over_ranks):
    """W untimed + K timed calls of step(i) bracketed by barrier + synchronize; returns ms per step (max over ranks)."""
    import torch
    for i in range(warmup):
        step(i)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(steps):
        step(warmup + i)
    ev1.record()
    barrier()
    return max_over_ranks(ev0.elapsed_time(ev1)) / steps


def stage_profile(step, nprof):
    """Per-stage device times per step (CUDA events recorded inside the library around each kernel)."""
    import torch
    from dirt_b200 import _lib
    lib = _lib.load()
    _lib.check(lib.dirt_profile_reset())
    _lib.check(lib.dirt_profile_enable(1))
    for i in range(nprof):
        step(i)
        torch.cuda.synchronize()
        _lib.check(lib.dirt_profile_collect())
    _lib.check(lib.dirt_profile_enable(0))
    times = _lib.stage_times()
    return {k: ms / nprof for k, (ms, n) in times.items()}, {k: n / nprof for k, (ms, n) in times.items()}


def run_ours(args, wl, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from dirt_b200 import rasterise_ops as ops, _lib, sharding
    import dirt_b200

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = bind_to_gpu_numa(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return ms

    B, H, W, C = wl["B"], wl["H"], wl["W"], wl["C"]
    shared = bool(wl.get("shared"))
    sc = build_scene(wl, B, seed_shift=0 if shared else 10 * rank)     # a shared mesh is the same on every rank
    if shared:
        # each rank renders its own frames of the camera path
        full = build_scene(wl, B * world)
        lo, hi = sharding.shard_range(B * world, rank, world)
        sc = {k: (v if k in ("faces", "vertex_colors", "world") else v[lo:hi]) for k, v in full.items()}
        del full
    V, F = sc["vertices"].shape[1], sc["faces"].shape[1]
    host = {k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for k, v in sc.items()}
    host["target"] = torch.from_numpy(np.random.RandomState(7 + rank).uniform(0, 1, (B, H, W, C)).astype(np.float32)).pin_memory()
    d = {k: v.to(dev) for k, v in host.items()}
    grad_pixels = torch.from_numpy(np.random.RandomState(5 + rank).randn(B, H, W, C).astype(np.float32)).to(dev)
    # The geometry MOVES between steps (sphere workloads): FRAMES pre-uploaded vertex sets, visited 0 1 2 3 2 1 0 ...
    # The forward keeps a visibility hint between calls (DirtForwardArgs.visibility_hint): identical inputs every
    # step would be its best case, a sequence of slightly different frames is the case it is made for.
    moving = not shared and not args.static
    host_frames = [host["vertices"]] + ([torch.from_numpy(v).pin_memory() for v in build_frames(wl, B, rank)] if moving else [])
    dev_frames = [h.to(dev) for h in host_frames]
    order = frame_order(len(dev_frames))
    nframes = len(dev_frames)
    hint = ops.visibility_hint(dev, B, F, key="bench") if (F >= ops.HINT_MIN_FACES and args.hint) else None

    def step(i):
        v = dev_frames[order[i % len(order)]]
        px, fid, gb = ops.forward_raw(d["background"], v, d["vertex_colors"], d["faces"], want_aux=True, hint=hint)
        if not shared:
            return ops.backward_raw(d["faces"], px, grad_pixels, fid, gb, V)
        # shared mesh: the backward folds each frame's clip-space vertex gradient to world space
        # with that frame's matrix and sums position + colour gradients over the frames into one
        # packed [V, 8] buffer, which ONE allreduce then sums over the ranks in place
        packed, gbg = ops.backward_shared_raw(d["faces"], px, grad_pixels, fid, gb, V, d["matrices"])
        return sharding.allreduce_packed(packed, C)

    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()                      # nvidia-smi needs ~1 s to deliver its first sample
    warm = max(args.warmup, 3)
    for i in range(warm):
        step(i)
    barrier()
    n0 = _lib.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_timed = time.time()
    ev0.record()
    for i in range(args.steps):
        step(warm + i)
    ev1.record()
    barrier()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    launches = _lib.launch_count() - n0
    # A timed region shorter than a few sampling periods (100 ms each) yields no clock samples:
    # keep the same load running, untimed, until some have arrived (all ranks, to stay in step).
    window = "timed region"
    if ms_total < 1500.0:
        window = "timed region + untimed continuation of the same steps (timed region shorter than the sampling period)"
        for i in range(int(min(5000, 1500.0 / max(ms_total / args.steps, 1e-3)))):    # same count on every rank
            step(i)
        barrier()
    clock_info = clocks.stop(t_timed, window) if rank == 0 else None
    ms_step = ms_total / args.steps
    value = world * B * H * W / (ms_step * 1e-3) / 1e6

    # ---- per-kernel durations, measured live with CUDA events inside the library -------------
    nprof = min(args.steps, 12)
    stages, per_step_launches = stage_profile(step, nprof)
    per_kernel, fwd_bytes, bwd_bytes = algorithmic_bytes(B, H, W, C, V, F, shared)
    dominant = max(per_kernel, key=lambda k: stages.get(k, 0.0))
    peak, peak_src = hbm_peak()
    # A kernel launched several times per step (k_setup: once per depth bucket) is reported per
    # launch: its algorithmic bytes and its measured time are both divided by the launch count.
    nl = max(per_step_launches.get(dominant, 1.0), 1.0)
    achieved = per_kernel[dominant] / (stages[dominant] * 1e-3) / 1e9 if stages[dominant] > 0 else 0.0
    traffic = ncu_traffic("k_" + dominant, args.workload)
    roofline = {"bound": "hbm", "kernel": "k_" + dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic / nl if traffic else None, "peak_source": peak_src,
                "launches_per_step": nl, "algorithmic_bytes_per_launch": per_kernel[dominant] / nl,
                "kernel_ms": stages[dominant] / nl,
                "step_algorithmic_bytes": fwd_bytes + bwd_bytes,
                "step_achieved_GBps": (fwd_bytes + bwd_bytes) / (ms_step * 1e-3) / 1e9,
                "step_frac": (fwd_bytes + bwd_bytes) / (ms_step * 1e-3) / 1e9 / peak,
                "per_kernel_frac": {k: (per_kernel[k] / (stages[k] * 1e-3) / 1e9 / peak if stages.get(k, 0) > 0 else None) for k in per_kernel},
                "binding_pipes_ncu": ncu_pipes("k_" + dominant, args.workload),
                "note": "k_setup is bound by instruction issue, k_grad_assemble / k_resolve by the LSU data pipe "
                        "(profiles/); HBM is the stated roofline, not what limits them"}

    # ---- e2e: public API, pinned host inputs, H2D + D2H inside the timed region -------------
    # Every step uploads ITS inputs (this step's vertices, colours, faces; see below for the background) from
    # pinned host memory and reads back its loss and gradients.  Steps are software-pipelined over two buffer sets
    # and three streams (upload i+1 | compute i | download i-1), the way a training loop feeds a
    # GPU; the target image of the loss stays resident.
    # shared mesh (c5): the caller feeds the mesh once in world space plus one matrix per frame
    # The background is NOT copied: rasterise_batch takes the pinned host tensor as it is and the
    # resolve kernel reads it in place over PCIe, only where no face covers the pixel (the bytes
    # that cross the bus are counted below from the step's own face-id map).
    in_keys = ("world", "matrices", "vertex_colors", "faces") if shared else ("vertices", "vertex_colors", "faces")
    s_in, s_out, s_main = torch.cuda.Stream(dev), torch.cuda.Stream(dev), torch.cuda.current_stream(dev)
    # The step's inputs live in ONE pinned staging block per frame (the caller's upload arena) and cross the bus
    # with ONE cudaMemcpyAsync; the device side is one block per buffer set, the tensors are views into it.
    def carve(block, like):
        views, off = {}, 0
        for k in in_keys:
            n = like[k].numel() * like[k].element_size()
            views[k] = block[off:off + n].view(like[k].dtype).view(like[k].shape)
            off += (n + 255) // 256 * 256
        return views
    arena_bytes = sum((host[k].numel() * host[k].element_size() + 255) // 256 * 256 for k in in_keys)
    host_arena = []
    for fr in range(nframes):
        blk = torch.empty(arena_bytes, dtype=torch.uint8).pin_memory()
        vw = carve(blk, host)
        for k in in_keys:
            vw[k].copy_(host_frames[fr] if k == "vertices" else host[k])
        host_arena.append(blk)
    dev_arena = [torch.empty(arena_bytes, dtype=torch.uint8, device=dev) for _ in range(2)]
    dev_in = [carve(a, host) for a in dev_arena]
    out_gv = [torch.empty((1 if shared else B, V, 4), dtype=torch.float32).pin_memory() for _ in range(2)]
    out_gc = [torch.empty((1 if shared else B, V, C), dtype=torch.float32).pin_memory() for _ in range(2)]
    out_loss = [torch.empty((), dtype=torch.float32).pin_memory() for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    free = [torch.cuda.Event() for _ in range(2)]
    done = [torch.cuda.Event() for _ in range(2)]
    losses = []

    # what one upload moves: the whole block, or (second e2e figure) only the tensors that change from step to step --
    # `in_keys` lists them first and the constant topology (faces) last, so that is a prefix of the block
    upload_bytes = [arena_bytes]
    e2e_bg = [host["background"]]          # pinned host tensor, read in place; the third figure keeps it resident

    def e2e_upload(i):
        slot = i % 2
        with torch.cuda.stream(s_in):
            s_in.wait_event(free[slot])                 # the step that last used this buffer set has finished
            n = upload_bytes[0]
            dev_arena[slot][:n].copy_(host_arena[order[i % len(order)]][:n], non_blocking=True)
            ready[slot].record(s_in)

    def e2e_compute(i):
        slot = i % 2
        s_main.wait_event(ready[slot])
        bg = e2e_bg[0]
        if shared:
            v = dev_in[slot]["world"].detach().requires_grad_(True)
            c = dev_in[slot]["vertex_colors"][0].detach().requires_grad_(True)
            px = dirt_b200.rasterise_batch_shared(bg, v, dev_in[slot]["matrices"], c, dev_in[slot]["faces"][0])
        else:
            v = dev_in[slot]["vertices"].detach().requires_grad_(True)
            c = dev_in[slot]["vertex_colors"].detach().requires_grad_(True)
            px = dirt_b200.rasterise_batch(bg, v, c, dev_in[slot]["faces"])
        loss = 0.5 * ((px - d["target"]) ** 2).sum()
        loss.backward()
        if shared:
            gvo, gco = sharding.allreduce_shared_gradients(v.grad.unsqueeze(0), c.grad.unsqueeze(0))
            gvo, gco = gvo.unsqueeze(0), gco.unsqueeze(0)
        else:
            gvo, gco = v.grad, c.grad
        free[slot].record(s_main)
        if i >= 2:
            done[slot].synchronize()                    # the host buffers of step i-2 have landed
            losses.append(float(out_loss[slot]))        # host read of that step's loss
        with torch.cuda.stream(s_out):
            s_out.wait_event(free[slot])
            for t in (gvo, gco, loss):
                t.record_stream(s_out)
            out_gv[slot].copy_(gvo, non_blocking=True)
            out_gc[slot].copy_(gco, non_blocking=True)
            out_loss[slot].copy_(loss.detach(), non_blocking=True)
            done[slot].record(s_out)

    def e2e_run(n):
        for ev in free:
            ev.record(s_main)
        e2e_upload(0)
        for i in range(n):
            if i + 1 < n:
                e2e_upload(i + 1)
            e2e_compute(i)
        for slot in range(2):
            done[slot].synchronize()
        losses.append(float(out_loss[(n - 1) % 2]))

    e2e_steps = max(2, min(args.steps, 12))

    def e2e_time():
        """ms per step: three timed runs of e2e_steps steps after a warm-up run, the MEDIAN of the three (the loop
        synchronises with the host every step, so one stall of the host -- a page fault, the allocator -- would
        otherwise double a 50 ms measurement); every run is max over ranks."""
        e2e_run(3)
        runs = []
        for _ in range(3):
            barrier()
            ev0.record()
            e2e_run(e2e_steps)
            ev1.record()
            barrier()
            runs.append(max_over_ranks(ev0.elapsed_time(ev1)) / e2e_steps)
        return sorted(runs)[1], runs

    e2e_ms, e2e_runs = e2e_time()
    h2d = sum(host[k].numel() * host[k].element_size() for k in in_keys)
    bg_note = ""
    _, fid_now, _ = ops.forward_raw(d["background"], d["vertices"], d["vertex_colors"], d["faces"], want_aux=True)
    uncovered = int((fid_now < 0).sum().item())
    h2d += uncovered * C * 4
    bg_note = ("; the background stays in pinned host memory and is read in place by the resolve kernel at the %d of %d "
               "pixels no face covers (%d bytes counted; the bus moves whole 32-byte sectors)" % (uncovered, B * H * W, uncovered * C * 4))
    d2h = out_gv[0].numel() * 4 + out_gc[0].numel() * 4 + 4
    e2e = {"value": world * B * H * W / (e2e_ms * 1e-3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms, "steps": e2e_steps, "runs_ms_per_step": e2e_runs,
           "timing": "median of three runs of %d steps" % e2e_steps, "last_loss": losses[-1],
           "path": ("dirt_b200.rasterise_batch_shared" if shared else "dirt_b200.rasterise_batch") +
                   " + autograd backward; each step uploads its inputs (that step's vertex set) from pinned host "
                   "memory (one staging block, one cudaMemcpyAsync per step) and downloads loss + gradients; double-buffered over "
                   "copy/compute/copy streams" + bg_note}

    extra = {}
    # ---- e2e of a loop that keeps the constant topology resident (faces uploaded once, as the reference's caller
    #      holds them in a graph constant, tests/optimize_sequence.py:293-352): vertices and colours still cross every step
    if not shared and not args.no_extras:
        for slot in range(2):
            dev_arena[slot].copy_(host_arena[0])         # faces now live in both buffer sets
        torch.cuda.synchronize()
        upload_bytes[0] = sum((host[k].numel() * host[k].element_size() + 255) // 256 * 256 for k in in_keys if k != "faces")
        ms_r, _ = e2e_time()
        extra["e2e_resident_topology"] = {
            "value": world * B * H * W / (ms_r * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_r, "steps": e2e_steps,
            "h2d_bytes_per_step": h2d - host["faces"].numel() * 4, "d2h_bytes_per_step": d2h,
            "path": "as e2e, but `faces` is uploaded once and stays resident; every step still uploads its vertices and colours"}
        # ... and of a loop in which only the optimised variable crosses the bus: vertices up, loss + gradients down;
        # background, colours and faces resident (the reference's own loop feeds 8 floats per iteration and fetches the
        # image, tests/optimize_sequence.py:158-166: everything else is a graph constant on the device)
        upload_bytes[0] = (host["vertices"].numel() * 4 + 255) // 256 * 256
        e2e_bg[0] = d["background"]
        ms_v, _ = e2e_time()
        extra["e2e_vertices_only"] = {
            "value": world * B * H * W / (ms_v * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_v, "steps": e2e_steps,
            "h2d_bytes_per_step": host["vertices"].numel() * 4, "d2h_bytes_per_step": d2h,
            "path": "as e2e, but background, vertex_colors and faces stay resident; every step uploads its vertices and downloads loss + gradients"}
        upload_bytes[0] = arena_bytes
        e2e_bg[0] = host["background"]
    # ---- the same timed loop with the visibility hint carried from call to call (the other setting of --hint) ----
    if F >= ops.HINT_MIN_FACES and not shared and not args.no_extras:
        other = None if hint is not None else ops.visibility_hint(dev, B, F, key="bench-other")

        def step_other(i):
            v = dev_frames[order[i % len(order)]]
            px, fid, gb = ops.forward_raw(d["background"], v, d["vertex_colors"], d["faces"], want_aux=True, hint=other)
            return ops.backward_raw(d["faces"], px, grad_pixels, fid, gb, V)

        ms_o = time_steps(step_other, min(args.steps, 10), 4, barrier, max_over_ranks)
        st_o, _ = stage_profile(step_other, 6)
        extra["visibility_hint" if other is not None else "no_visibility_hint"] = {
            "ms_per_step": ms_o, "value": world * B * H * W / (ms_o * 1e-3) / 1e6, "unit": UNIT,
            "stages_ms": {k: v for k, v in st_o.items() if v > 0},
            "note": "same loop, same moving geometry; stages: depth_order reads one flag per face instead of three projected corners, "
                    "resolve includes the pass that stamps the visible faces"}
    # ---- e2e through the host-buffer C ABI (what a non-torch caller pays): every tensor crosses the bus ----
    if rank == 0 and not shared and not args.no_extras:
        extra["e2e_host_abi"] = host_abi_e2e(lib, host, host_frames, order, B, H, W, C, V, F)
    del dev_in, dev_arena, host_arena, out_gv, out_gc
    # ---- the other device-resident configs of BASELINE.json, so that a driver-run number exists for them ----
    if args.workload == "headline" and not args.no_extras:
        del d, dev_frames, grad_pixels
        torch.cuda.empty_cache()
        extra["c4"] = extra_sphere(WORKLOADS["c4"], args, rank, world, dev, barrier, max_over_ranks)
        torch.cuda.empty_cache()
        extra["c5"] = extra_c5(WORKLOADS["c5"], args, rank, world, dev, barrier, max_over_ranks)
        torch.cuda.empty_cache()

    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        if _ORIG_AFFINITY:
            os.sched_setaffinity(0, _ORIG_AFFINITY)      # the CPU baseline may use every host core again
        cpu_base = cpu_baseline_run(wl, 12.0)
        one = cpu_baseline_run(wl, 3.0, threads=1)
        cpu_base["single_thread"] = {k: one[k] for k in ("value", "unit", "cores", "sample", "seconds_per_run")}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    print(json.dumps({
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": args.workload + ": " + wl["desc"], "batch_per_gpu": B, "height": H, "width": W, "channels": C,
                   "vertices": V, "faces": F,
                   "geometry": ("moves between steps: %d pre-uploaded vertex sets, each element's object turned a further %.2f deg "
                                "(<= 0.7 px on screen) per frame, visited 0 1 2 3 2 1 0 ...%s" % (nframes, FRAME_STEP_DEG, "; the forward carries its visibility hint from call to call" if hint is not None else "")) if moving
                               else "static (the same inputs every step)" + ("" if hint is None else "; visibility hint carried from call to call"),
                   "visibility_hint": hint is not None,
                   "sharding": ("frames across ranks; one packed all_reduce of [V,4+C] shared-mesh gradients per step" if shared
                                else "batch elements across ranks, no data-path collective"),
                   "l2": "inputs+outputs per step (%.0f MB) exceed the 126 MB L2; no explicit flush" % ((fwd_bytes + bwd_bytes) / 1e6),
                   "host": numa},
        "roofline": roofline, "stages_ms": stages, "cpu_baseline": cpu_base, "e2e": e2e, "gpu_launches": int(launches),
        "clocks": clock_info, "extra": extra}))


def host_abi_e2e(lib, host, host_frames, order, B, H, W, C, V, F, steps=3):
    """fwd+bwd through dirt_rasterise_forward_host / dirt_rasterise_backward_host: plain HOST pointers in and out,
    pixels, aux buffers and all gradients returned to the host -- the cost a caller without device memory pays."""
    import torch
    from dirt_b200 import _lib
    px = torch.empty((B, H, W, C), dtype=torch.float32).pin_memory()
    g = torch.ones((B, H, W, C), dtype=torch.float32).pin_memory()
    gv = torch.empty((B, V, 4), dtype=torch.float32).pin_memory()
    gc = torch.empty((B, V, C), dtype=torch.float32).pin_memory()
    gbg = torch.empty((B, H, W, C), dtype=torch.float32).pin_memory()
    shape = _lib.DirtShape(B, H, W, C, V, F)

    def once(i):
        v = host_frames[order[i % len(order)]]
        # forward without the aux outputs, backward in its stateless form (visibility re-derived on the device from the
        # vertices): face_id and gbuffer -- 20 bytes per pixel -- never cross the bus
        fa = _lib.DirtForwardArgs(shape, host["background"].data_ptr(), H * W * C, v.data_ptr(), V * 4,
                                  host["vertex_colors"].data_ptr(), V * C, host["faces"].data_ptr(), F * 3,
                                  px.data_ptr(), None, None, None, 0, None, 0)
        _lib.check(lib.dirt_rasterise_forward_host(ctypes.byref(fa)))
        ba = _lib.DirtBackwardArgs(shape, host["faces"].data_ptr(), F * 3, px.data_ptr(), g.data_ptr(), None, None,
                                   gv.data_ptr(), V * 4, gc.data_ptr(), V * C, gbg.data_ptr(), v.data_ptr(), V * 4, None, 0, 0)
        _lib.check(lib.dirt_rasterise_backward_host(ctypes.byref(ba)))

    once(0)
    t0 = time.perf_counter()
    for i in range(steps):
        once(1 + i)
    sec = (time.perf_counter() - t0) / steps
    img = B * H * W * C * 4
    mesh = B * (V * 16 + V * C * 4 + F * 12)
    lib.dirt_release_host_staging()
    return {"value": B * H * W / sec / 1e6, "unit": UNIT, "ms_per_step": sec * 1e3, "steps": steps,
            "h2d_bytes_per_step": img + mesh + B * F * 12 + 2 * img + B * V * 16, "d2h_bytes_per_step": img + B * V * (16 + 4 * C) + img,
            "path": "dirt_rasterise_forward_host (pixels only) + dirt_rasterise_backward_host in its stateless form on pinned host "
                    "buffers (synchronous; pixels and every gradient returned to the host, visibility re-derived on the device), "
                    "wall clock, rank 0 only"}


def extra_sphere(wl, args, rank, world, dev, barrier, max_over_ranks):
    """Device-resident fwd+bwd of another sphere config (BASELINE.json config 4), same protocol as the headline."""
    import torch
    from dirt_b200 import rasterise_ops as ops
    B, H, W, C = wl["B"], wl["H"], wl["W"], wl["C"]
    sc = build_scene(wl, B, seed_shift=10 * rank)
    V, F = sc["vertices"].shape[1], sc["faces"].shape[1]
    d = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in sc.items()}
    frames = [d["vertices"]] + [torch.from_numpy(v).to(dev) for v in build_frames(wl, B, rank)]
    order = frame_order(len(frames))
    g = torch.randn((B, H, W, C), device=dev)
    hint = ops.visibility_hint(dev, B, F, key="bench-" + str(H)) if args.hint else None

    def step(i):
        px, fid, gb = ops.forward_raw(d["background"], frames[order[i % len(order)]], d["vertex_colors"], d["faces"], hint=hint)
        return ops.backward_raw(d["faces"], px, g, fid, gb, V)

    steps = min(args.steps, 10)
    ms = time_steps(step, steps, 3, barrier, max_over_ranks)
    stages, _ = stage_profile(step, 6)
    return {"workload": wl["desc"] + "; geometry moves between steps as in the headline", "value": world * B * H * W / (ms * 1e-3) / 1e6,
            "unit": UNIT, "ms_per_step": ms, "steps": steps, "stages_ms": {k: v for k, v in stages.items() if v > 0}}


def extra_c5(wl, args, rank, world, dev, barrier, max_over_ranks):
    """Config c5 (shared hill mesh, frames sharded over the ranks): the step with its packed allreduce issued
    synchronously, without it, and split into two frame groups whose first allreduce overlaps the second group's
    kernels (side stream); the only NCCL traffic on this path's data."""
    import torch
    import torch.distributed as dist
    from dirt_b200 import rasterise_ops as ops, sharding, _lib
    B, H, W, C = wl["B"], wl["H"], wl["W"], wl["C"]
    full = build_scene(wl, B * world)
    lo, hi = sharding.shard_range(B * world, rank, world)
    sc = {k: (v if k in ("faces", "vertex_colors", "world") else v[lo:hi]) for k, v in full.items()}
    del full
    V, F = sc["vertices"].shape[1], sc["faces"].shape[1]
    d = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in sc.items()}
    g = torch.randn((B, H, W, C), device=dev)
    side = torch.cuda.Stream(dev)
    main = torch.cuda.current_stream(dev)

    def fwd_bwd(sl):
        px, fid, gb = ops.forward_raw(d["background"][sl], d["vertices"][sl], d["vertex_colors"], d["faces"])
        return ops.backward_shared_raw(d["faces"], px, g[sl], fid, gb, V, d["matrices"][sl])[0]

    def step_sync(i):
        return sharding.allreduce_packed(fwd_bwd(slice(0, B)), C)

    def step_local(i):
        return fwd_bwd(slice(0, B))

    def step_overlap(i):
        # group 0's gradients travel while group 1 renders; the two sums are added after the second collective
        p0 = fwd_bwd(slice(0, B // 2))
        ev = torch.cuda.Event()
        ev.record(main)
        with torch.cuda.stream(side):
            side.wait_event(ev)
            p0.record_stream(side)
            w0 = dist.all_reduce(p0, async_op=True) if world > 1 else None
        p1 = fwd_bwd(slice(B // 2, B))
        if world > 1:
            dist.all_reduce(p1)
            w0.wait()
        main.wait_stream(side)
        return p0.add_(p1)

    steps = min(args.steps, 10)
    out = {"workload": wl["desc"], "frames_per_gpu": B, "allreduce_bytes": V * lib_stride(C) * 4, "unit": UNIT}
    variants = [("sync_allreduce", step_sync), ("no_allreduce", step_local), ("overlapped_two_groups", step_overlap)]
    # the same sum taken by kernels that work on peer memory directly (symmetric memory over NVLink / NVSwitch)
    sym = None
    if world > 1:
        try:
            sym = sharding.SymmetricPacked(V, C, dev)
        except Exception as e:                       # no symmetric-memory support in this build / on this box
            out["symmetric_memory"] = "unavailable: %s" % (str(e)[:120] or type(e).__name__)
    if sym is not None:
        def sym_step(mode):
            def fn(i):
                px, fid, gb = ops.forward_raw(d["background"], d["vertices"], d["vertex_colors"], d["faces"])
                ops.backward_shared_raw(d["faces"], px, g, fid, gb, V, d["matrices"], out=sym.tensor)
                return sym.allreduce(mode)
            return fn
        ref_sum = step_sync(0)
        ref_sum = torch.cat([ref_sum[0], ref_sum[1]], 1).clone()
        for mode in ("peer", "peer_multicast", "two_shot", "multimem") if sym.has_multicast else ("peer", "two_shot"):
            try:
                got = sym_step(mode)(0)
                torch.cuda.synchronize()
                err = float((torch.cat([got[0], got[1]], 1) - ref_sum).abs().max() / ref_sum.abs().max())
                assert err < 1e-4, "symmetric %s all-reduce differs from NCCL by %g" % (mode, err)
                variants.append(("symm_" + mode, sym_step(mode)))
            except Exception as e:
                out["symm_" + mode] = "failed: %s" % (str(e)[:160] or type(e).__name__)
    for name, fn in variants:
        ms = time_steps(fn, steps, 3, barrier, max_over_ranks)
        out[name] = {"ms_per_step": ms, "value": world * B * H * W / (ms * 1e-3) / 1e6}
    out["allreduce_exposed_ms"] = out["sync_allreduce"]["ms_per_step"] - out["no_allreduce"]["ms_per_step"]
    for name in ("symm_peer", "symm_peer_multicast", "symm_two_shot", "symm_multimem"):
        if isinstance(out.get(name), dict):
            out[name]["exposed_ms"] = out[name]["ms_per_step"] - out["no_allreduce"]["ms_per_step"]
    out["nccl"] = {k: os.environ[k] for k in os.environ if k.startswith("NCCL_")}
    return out


def lib_stride(C):
    from dirt_b200 import _lib
    return int(_lib.load().dirt_shared_grad_stride(C))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="headline", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--static", action="store_true", help="identical inputs every step (round-1 protocol)")
    ap.add_argument("--hint", action="store_true", help="carry the visibility hint between calls in the timed loops (default: "
                    "off -- measured no faster; extra.visibility_hint reports it)")
    ap.add_argument("--no-extras", action="store_true", help="skip extra.c4 / extra.c5 / extra.e2e_host_abi")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl, rank, world)
    else:
        run_ours(args, wl, rank, world, local_rank)


if __name__ == "__main__":
    main()
