"""CPU, world_size 2, gloo: the N > 1 host logic of the path -- batch sharding and the single
packed allreduce of shared-mesh gradients.  (The kernels need a GPU; what runs here is exactly
the code bench.py and a user run around them under torchrun.)"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dirt_b200 import sharding


def test_shard_ranges_partition_the_batch():
    for batch in (0, 1, 7, 64, 257):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_range(batch, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == batch
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(4, 2, 2)


def test_shard_batch_passes_broadcast_inputs_through():
    t = {"background": torch.zeros(6, 4, 4, 3), "vertices": torch.zeros(6, 5, 4), "faces": torch.zeros(1, 2, 3)}
    s = sharding.shard_batch(t, 1, 4)
    assert s["background"].shape[0] == 2 and s["vertices"].shape[0] == 2 and s["faces"].shape[0] == 1


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, V, C, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.RandomState(100 + rank)
        # each rank holds the gradients of its own frames w.r.t. the shared mesh
        gv = torch.from_numpy(rng.randn(3, V, 4).astype(np.float32))
        gc = torch.from_numpy(rng.randn(1, V, C).astype(np.float32))
        a, b = sharding.allreduce_shared_gradients(gv, gc)
        work, finish = sharding.allreduce_shared_gradients(gv, gc, async_op=True)
        a2, b2 = finish()
        lo, hi = sharding.shard_range(7, rank, world)
        out[rank] = (a.numpy(), b.numpy(), a2.numpy(), b2.numpy(), (lo, hi))
    finally:
        dist.destroy_process_group()


def test_allreduce_of_shared_gradients_world2_gloo():
    world, V, C = 2, 37, 3
    port = _free_port()
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, port, V, C, out), nprocs=world, join=True)
        res = dict(out)
    exp_v = sum(np.random.RandomState(100 + r).randn(3, V, 4).astype(np.float32).sum(0) for r in range(world))
    for r in range(world):
        rng = np.random.RandomState(100 + r)
        rng.randn(3, V, 4)
    exp_c = np.zeros((V, C), np.float32)
    for r in range(world):
        rng = np.random.RandomState(100 + r)
        rng.randn(3, V, 4)
        exp_c += rng.randn(1, V, C).astype(np.float32)[0]
    for r in range(world):
        a, b, a2, b2, span = res[r]
        np.testing.assert_allclose(a, exp_v, rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(b, exp_c, rtol=1e-6, atol=1e-6)
        assert np.array_equal(a, a2) and np.array_equal(b, b2)
    assert res[0][4] == (0, 4) and res[1][4] == (4, 7)


def _worker_packed(rank, world, port, V, C, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        stride = (4 + C + 3) // 4 * 4
        packed = torch.from_numpy(np.random.RandomState(200 + rank).randn(V, stride).astype(np.float32))
        gw, gc = sharding.allreduce_packed(packed.clone(), C)
        work, finish = sharding.allreduce_packed(packed.clone(), C, async_op=True)
        gw2, gc2 = finish()
        out[rank] = (gw.numpy().copy(), gc.numpy().copy(), gw2.numpy().copy(), gc2.numpy().copy())
    finally:
        dist.destroy_process_group()


def test_allreduce_of_the_packed_shared_gradient_buffer_world2_gloo():
    """The buffer `backward_shared_raw` fills ([V, stride]: world-space position gradient, colour
    gradient, padding) is reduced in place with one collective; the results are views of it."""
    world, V, C = 2, 41, 3
    port = _free_port()
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker_packed, args=(world, port, V, C, out), nprocs=world, join=True)
        res = dict(out)
    exp = sum(np.random.RandomState(200 + r).randn(V, 8).astype(np.float32) for r in range(world))
    for r in range(world):
        gw, gc, gw2, gc2 = res[r]
        assert gw.shape == (V, 4) and gc.shape == (V, C)
        np.testing.assert_allclose(gw, exp[:, :4], rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(gc, exp[:, 4:7], rtol=1e-6, atol=1e-6)
        assert np.array_equal(gw, gw2) and np.array_equal(gc, gc2)
