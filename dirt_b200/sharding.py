"""Multi-GPU host logic for the rasterise path: one process per GPU, batch elements sharded
across ranks, and -- only when the mesh is shared by all elements -- one packed allreduce of the
vertex / colour gradients.

The reference's multi-GPU story is independent per-device ops (tests/multi_gpu_test.py:22-29;
per-CUcontext state csrc/gl_dispatcher.h:27, :101-108): batch elements never interact
(csrc/rasterise_egl.cpp:440-458), so the forward needs no collective at all.  For an inverse-
rendering loop over frames of one mesh (tests/optimize_sequence.py:293-352) every rank holds
gradients w.r.t. the same [V,*] arrays; they are summed with a single torch.distributed
all_reduce over one contiguous [V, 4+C] buffer (NCCL over NVLink on GPUs, gloo in the CPU tests).
Nothing here touches pixels: torch.distributed is plumbing, the kernels are libdirt_b200.so.
"""
import torch
import torch.distributed as dist


def shard_range(batch, rank, world):
    """Contiguous shard [lo, hi) of `batch` elements for `rank` of `world`; sizes differ by at
    most one and earlier ranks take the larger shards."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("rank %r out of range for world size %r" % (rank, world))
    base, extra = divmod(int(batch), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_batch(tensors, rank, world):
    """Slices every tensor of `tensors` (dict or sequence) on dim 0 to this rank's shard.  A leading
    dim of 1 is a broadcast input (shared mesh / shared background) and is passed through."""
    items = tensors.items() if isinstance(tensors, dict) else enumerate(tensors)
    batch = max(int(t.shape[0]) for _, t in items)
    lo, hi = shard_range(batch, rank, world)
    cut = lambda t: t if t.shape[0] == 1 and batch > 1 else t[lo:hi]
    if isinstance(tensors, dict):
        return {k: cut(t) for k, t in tensors.items()}
    return type(tensors)(cut(t) for t in tensors)


def pack_shared_gradients(grad_vertices, grad_vertex_colors):
    """[1|.,V,4] + [1|.,V,C] -> one contiguous [V, 4+C] buffer (summing a leading batch dim)."""
    gv = grad_vertices.reshape(-1, grad_vertices.shape[-2], 4).sum(0)
    gc = grad_vertex_colors.reshape(-1, grad_vertex_colors.shape[-2], grad_vertex_colors.shape[-1]).sum(0)
    return torch.cat([gv, gc], dim=1).contiguous()


def allreduce_shared_gradients(grad_vertices, grad_vertex_colors, group=None, async_op=False):
    """Sums the shared-mesh gradients over all ranks with ONE collective.
    Returns (grad_vertices [V,4], grad_vertex_colors [V,C]) -- or, with async_op=True,
    (work, finish) where finish() waits and returns that pair, so the collective can overlap
    whatever the caller still has to do (e.g. the grad_background copy)."""
    packed = pack_shared_gradients(grad_vertices, grad_vertex_colors)
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1

    def split():
        return packed[:, :4].contiguous(), packed[:, 4:].contiguous()

    if world == 1:
        return (None, split) if async_op else split()
    work = dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
    if async_op:
        def finish():
            work.wait()
            return split()
        return work, finish
    return split()


def allreduce_packed(packed, channels, group=None, async_op=False):
    """All-reduce of the packed shared-mesh gradient buffer [V, stride] that
    `rasterise_ops.backward_shared_raw` fills (in place, ONE collective, no packing copy).
    Returns (grad_world [V,4], grad_vertex_colors [V,C]) as views of the buffer -- or, with
    async_op=True, (work, finish)."""
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1

    def split():
        return packed[:, :4], packed[:, 4:4 + channels]

    if world == 1:
        return (None, split) if async_op else split()
    work = dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
    if async_op:
        def finish():
            work.wait()
            return split()
        return work, finish
    return split()


class SymmetricPacked:
    """The packed shared-mesh gradient buffer [V, stride] allocated in SYMMETRIC memory (every rank's copy mapped into
    every other rank's address space over NVLink; torch.distributed._symmetric_memory), so that the sum over the ranks
    can be taken by a kernel that reads and writes peer memory directly instead of going through NCCL's channels:
        peer       this project's kernel (dirt_allreduce_peer, csrc/peer_allreduce.cu): rank r sums slice r of all
                   buffers in rank order and stores it into every buffer, between two flag handshakes
        peer_multicast   the same kernel on the NVSwitch's multicast mapping: the switch adds (multimem.ld_reduce / .st)
        two_shot   torch's kernel of the same scheme
        multimem   the same with the NVSwitch doing the additions (multimem.ld_reduce / multimem.st, NVLS)
    At the 16 MB of config c5 the collective is latency-bound under NCCL (0.05 ms at 2 GPUs, 0.11 ms at 8 for a
    0.49 ms step).  Allocate once per mesh, pass `.tensor` as `out=` of rasterise_ops.backward_shared_raw, then call
    `allreduce()`."""

    def __init__(self, V, channels, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        from . import _lib
        self.group = group if group is not None else dist.group.WORLD
        self.channels = channels
        stride = int(_lib.load().dirt_shared_grad_stride(channels))
        self.tensor = symm_mem.empty((V, stride), dtype=torch.float32, device=device)
        self.handle = symm_mem.rendezvous(self.tensor, group=self.group)
        self.has_multicast = bool(getattr(self.handle, "multicast_ptr", 0))
        self._peer = None

    def _peer_state(self):
        """Device tables for dirt_allreduce_peer: every rank's buffer and flag pad as this process addresses them.  The
        flags live in a second symmetric allocation of their own (torch's signal pad stays torch's)."""
        if self._peer is None:
            import ctypes
            import torch.distributed._symmetric_memory as symm_mem
            from . import _lib
            dev = self.tensor.device
            world, rank = dist.get_world_size(self.group), dist.get_rank(self.group)
            flags = symm_mem.empty(256, dtype=torch.int32, device=dev)
            fh = symm_mem.rendezvous(flags, group=self.group)
            flags.zero_()
            epoch = torch.zeros(4, dtype=torch.int32, device=dev)     # [0] calls completed, [1] != 0: a peer never arrived
            torch.cuda.synchronize(dev)
            dist.barrier(group=self.group)                 # every pad is zero before anyone raises a flag
            bufs = torch.tensor([int(p) for p in self.handle.buffer_ptrs], dtype=torch.int64, device=dev)
            sigs = torch.tensor([int(p) for p in fh.buffer_ptrs], dtype=torch.int64, device=dev)
            mc = int(getattr(self.handle, "multicast_ptr", 0) or 0)
            args = [_lib.DirtPeerAllreduceArgs(bufs.data_ptr(), sigs.data_ptr(), rank, world, self.tensor.numel(), epoch.data_ptr(), m)
                    for m in (None, mc or None)]                 # [0] peer loads / stores, [1] through the switch's multicast mapping
            self._peer = (args, bufs, sigs, flags, fh, epoch, ctypes)
        return self._peer

    def allreduce(self, mode="two_shot"):
        """Sums `.tensor` over the ranks in place; returns (grad_world [V,4], grad_vertex_colors [V,C]) views."""
        name = self.group.group_name
        if mode == "multimem":
            torch.ops.symm_mem.multimem_all_reduce_(self.tensor, "sum", name)
        elif mode == "two_shot":
            torch.ops.symm_mem.two_shot_all_reduce_(self.tensor, "sum", name)
        elif mode == "nccl":
            dist.all_reduce(self.tensor, op=dist.ReduceOp.SUM, group=self.group)
        elif mode in ("peer", "peer_multicast"):
            # this project's own kernel over the peer mappings, or over the switch's multicast mapping (csrc/peer_allreduce.cu)
            from . import _lib
            st = self._peer_state()
            dev = self.tensor.device
            if mode == "peer_multicast" and not self.has_multicast:
                raise ValueError("no multicast mapping of the buffer on this system")
            with torch.cuda.device(dev):
                _lib.check(_lib.load().dirt_allreduce_peer(st[6].byref(st[0][1 if mode == "peer_multicast" else 0]),
                                                           torch.cuda.current_stream(dev).cuda_stream))
        else:
            raise ValueError("mode must be peer, peer_multicast, two_shot, multimem or nccl")
        return self.tensor[:, :4], self.tensor[:, 4:4 + self.channels]


def rasterise_batch_sharded(background, vertices, vertex_colors, faces, rank=None, world=None):
    """`rasterise_batch` on this rank's shard of the batch (no collective; outputs stay sharded,
    as the reference's per-device graphs do).  Returns (pixels_shard, (lo, hi))."""
    from . import rasterise_ops
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    if world is None:
        world = dist.get_world_size() if dist.is_initialized() else 1
    batch = max(int(t.shape[0]) for t in (background, vertices, vertex_colors, faces))
    lo, hi = shard_range(batch, rank, world)
    bg, v, c, f = shard_batch((background, vertices, vertex_colors, faces), rank, world)
    return rasterise_ops.rasterise_batch(bg, v, c, f), (lo, hi)
