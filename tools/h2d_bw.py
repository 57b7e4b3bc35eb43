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
        del h, g
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
    if world > 1:
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out))


if __name__ == "__main__":
    main()
