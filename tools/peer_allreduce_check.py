"""Checks and times dirt_allreduce_peer (csrc/peer_allreduce.cu) against NCCL and torch's symmetric-memory kernels on the
packed shared-mesh gradient message of config c5 (501,264 x 8 floats = 16 MB).  Run under torchrun on one box:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29513 tools/peer_allreduce_check.py
Prints one JSON line (rank 0): max relative deviation from NCCL, identical bits across ranks, microseconds per call."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
from dirt_b200 import sharding  # noqa: E402


def main():
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    V, C = 501264, 3
    sym = sharding.SymmetricPacked(V, C, dev)
    out = {"ranks": world, "bytes": sym.tensor.numel() * 4, "multicast": sym.has_multicast}
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    for it in range(3):                                   # correctness, on fresh random data each time
        src = torch.randn(sym.tensor.shape, device=dev, generator=g)
        ref = src.clone()
        dist.all_reduce(ref)
        for mode in ("peer", "peer_multicast") if sym.has_multicast else ("peer",):
            sym.tensor.copy_(src)
            dist.barrier()
            sym.allreduce(mode)
            torch.cuda.synchronize()
            err = float((sym.tensor - ref).abs().max() / ref.abs().max())
            assert err < 1e-5, "%s all-reduce differs from NCCL by %g" % (mode, err)
            # every rank holds the same bits (one sum per slice, delivered to all)
            mine = sym.tensor.clone()
            other = mine.clone()
            dist.broadcast(other, src=0)
            assert torch.equal(mine, other), "ranks disagree (%s)" % mode
            out["max_rel_dev_vs_nccl"] = max(out.get("max_rel_dev_vs_nccl", 0.0), err)
        assert sym._peer[5].tolist()[1] == 0, "a handshake timed out"
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    modes = ["nccl", "peer", "two_shot"] + (["peer_multicast", "multimem"] if sym.has_multicast else [])
    for mode in modes:
        try:
            for _ in range(5):
                sym.allreduce(mode)
            dist.barrier(); torch.cuda.synchronize()
            e0.record()
            n = 50
            for _ in range(n):
                sym.allreduce(mode)
            e1.record()
            dist.barrier(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / n * 1e3], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            out["us_" + mode] = float(t.item())
        except Exception as e:
            out["us_" + mode] = "failed: %s" % (str(e)[:100])
        sym.tensor.zero_()
    dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out))


if __name__ == "__main__":
    main()
