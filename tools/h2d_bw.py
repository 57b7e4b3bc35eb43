"""Host<->device bandwidth of the box with 1..N ranks copying at once: pinned H2D and D2H
cudaMemcpyAsync of 64 / 128 / 256 MB, and the rate at which the resolve kernel reads a pinned-host background
IN PLACE (zero-copy over PCIe) -- an empty mesh makes every pixel a background pixel.  One JSON line.
    python tools/h2d_bw.py                                   # one rank
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/h2d_bw.py"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]


def main():
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(v, op):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=op)
        return float(t.item())

    out = {"ranks": world, "copies": {}}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 8
    for mb in (64, 128, 256):
        n = mb << 20
        h = torch.empty(n, dtype=torch.uint8).pin_memory()
        h.fill_(1)
        g = torch.empty(n, dtype=torch.uint8, device=dev)
        for name, (dst, src) in (("h2d", (g, h)), ("d2h", (h, g))):
            for _ in range(2):
                dst.copy_(src, non_blocking=True)
            barrier()
            e0.record()
            reps = 8
            for _ in range(reps):
                dst.copy_(src, non_blocking=True)
            e1.record()
            barrier()
            ms = reduce(e0.elapsed_time(e1), dist.ReduceOp.MAX if world > 1 else None) / reps
            out["copies"]["%s_%dMB" % (name, mb)] = {"GBps_per_rank": n / ms / 1e6, "GBps_total": world * n / ms / 1e6}
        # both directions at once (a training step uploads its inputs while the previous step's results come down)
        h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
        g2 = torch.empty(n, dtype=torch.uint8, device=dev)
        s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        barrier()
        e0.record()
        for _ in range(reps):
            with torch.cuda.stream(s_up):
                g.copy_(h, non_blocking=True)
            with torch.cuda.stream(s_dn):
                h2.copy_(g2, non_blocking=True)
        torch.cuda.current_stream(dev).wait_stream(s_up)
        torch.cuda.current_stream(dev).wait_stream(s_dn)
        e1.record()
        barrier()
        ms = reduce(e0.elapsed_time(e1), dist.ReduceOp.MAX if world > 1 else None) / reps
        out["copies"]["both_%dMB" % mb] = {"GBps_per_rank_each_direction": n / ms / 1e6, "GBps_total_both_directions": 2 * world * n / ms / 1e6}
        del h, g, h2, g2
    # zero-copy: the resolve kernel fetches a pinned background in place
    from dirt_b200 import rasterise_ops as ops
    B, H, W, C = 16, 512, 512, 3
    bg = torch.rand(B, H, W, C).pin_memory()
    v = torch.zeros(B, 3, 4, device=dev)
    c = torch.zeros(B, 3, C, device=dev)
    f = torch.zeros(B, 0, 3, dtype=torch.int32, device=dev)
    for _ in range(2):
        px, _, _ = ops.forward_raw(bg, v, c, f, want_aux=False)
    barrier()
    e0.record()
    reps = 5
    for _ in range(reps):
        px, _, _ = ops.forward_raw(bg, v, c, f, want_aux=False)
    e1.record()
    barrier()
    ms = reduce(e0.elapsed_time(e1), dist.ReduceOp.MAX if world > 1 else None) / reps
    assert torch.equal(px.cpu(), bg)
    nbytes = bg.numel() * 4
    out["zero_copy_background"] = {"bytes": nbytes, "GBps_per_rank": nbytes / ms / 1e6, "GBps_total": world * nbytes / ms / 1e6,
                                   "note": "dirt forward of an empty mesh: k_resolve reads every background pixel (12-byte scalar loads) from pinned host memory"}
    # the same read where the bench does it: a sphere batch, background fetched only at the ~17 % of the pixels no face
    # covers.  The bench counts uncovered * C * 4 bytes; the bus moves whole 32-byte sectors -- NVML's PCIe receive
    # counter (20 ms windows) sampled while the forward repeats says how many.
    if rank == 0:
        try:
            import threading
            import time
            import numpy as np
            import pynvml
            sys.path[:0] = [os.path.join(ROOT, "tests")]
            import scenes
            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(local).uuid)
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            sc = scenes.sphere(16, 512, 512)
            t = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in sc.items() if k != "background"}
            bgp = torch.from_numpy(sc["background"]).pin_memory()
            bgd = bgp.to(dev)
            px, fid, _ = ops.forward_raw(bgd, t["vertices"], t["vertex_colors"], t["faces"])
            uncovered = int((fid < 0).sum())
            res = {}
            for label, bgx in (("resident", bgd), ("pinned_in_place", bgp)):
                samples, stop = [], [False]

                def poll():
                    while not stop[0]:
                        samples.append(pynvml.nvmlDeviceGetPcieThroughput(h, pynvml.NVML_PCIE_UTIL_RX_BYTES))   # KB/s
                th = threading.Thread(target=poll, daemon=True)
                torch.cuda.synchronize()
                th.start()
                t0 = time.perf_counter()
                n = 0
                while time.perf_counter() - t0 < 1.0:
                    for _ in range(20):
                        ops.forward_raw(bgx, t["vertices"], t["vertex_colors"], t["faces"], want_aux=False)
                    torch.cuda.synchronize()
                    n += 20
                sec = (time.perf_counter() - t0) / n
                stop[0] = True
                th.join()
                rx = float(np.median(samples)) * 1e3 if samples else float("nan")       # bytes/s
                res[label] = {"ms_per_forward": sec * 1e3, "pcie_rx_GBps_nvml": rx / 1e9, "pcie_rx_bytes_per_forward": rx * sec}
            out["zero_copy_scattered"] = dict(res, uncovered_pixels=uncovered, counted_bytes=uncovered * 12,
                                              note="B=16 sphere frames of 512x512x3; bytes per forward = NVML receive rate x time per forward; "
                                                   "'resident' is the idle-bus reference (background on the device)")
        except Exception as e:
            out["zero_copy_scattered"] = "unavailable: %s" % (str(e)[:120] or type(e).__name__)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out))


if __name__ == "__main__":
    main()
