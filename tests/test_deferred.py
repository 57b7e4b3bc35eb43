"""Deferred G-buffer helpers: the fused silhouette dilation and the
single-visibility-pass multi-attribute rasterise, against the oracle / the three-call form of
samples/deferred.py:63-91."""
import numpy as np
import pytest
import torch

import scenes
from oracle import dilate as odil, golden


def _holey(seed, B=2, H=9, W=11, C=3):
    rng = np.random.RandomState(seed)
    x = rng.randn(B, H, W, C).astype(np.float32)
    x[rng.uniform(size=(B, H, W)) < 0.45] = -np.inf         # whole pixels are background, as in a G-buffer
    return x


def test_oracle_dilate_matches_torch_maxpool_form():
    # the reference's own formulation (samples/deferred.py:87-91) with torch ops on the CPU
    x = _holey(0)
    t = torch.from_numpy(x).requires_grad_(True)
    mask = (t == float("-inf")).float()
    pooled = torch.nn.functional.max_pool2d(t.permute(0, 3, 1, 2), 3, 1, 1).permute(0, 2, 3, 1)
    ref = torch.where(mask.bool(), pooled, t)
    assert np.array_equal(odil.dilate(x), ref.detach().numpy())
    g = np.random.RandomState(1).randn(*x.shape).astype(np.float32)
    finite = torch.isfinite(ref)
    (torch.where(finite, ref, torch.zeros_like(ref)) * torch.from_numpy(g)).sum().backward()
    og = odil.dilate_grad(x, np.where(finite.numpy(), g, 0))
    fg = x != -np.inf
    np.testing.assert_allclose(og[fg], t.grad.numpy()[fg], rtol=1e-6, atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(2, 9, 11, 3), (1, 33, 17, 1), (3, 8, 8, 6)])
def test_cuda_dilate_forward_backward(shape):
    import dirt_b200
    x = _holey(3, *shape)
    xt = torch.from_numpy(x).cuda().requires_grad_(True)
    out = dirt_b200.dilate_silhouette(xt)
    assert np.array_equal(out.detach().cpu().numpy(), odil.dilate(x))
    g = np.random.RandomState(4).randn(*shape).astype(np.float32)
    out.backward(torch.from_numpy(g).cuda())
    np.testing.assert_allclose(xt.grad.cpu().numpy(), odil.dilate_grad(x, g), rtol=1e-6, atol=1e-6)
    # unbatched form
    assert np.array_equal(dirt_b200.dilate_silhouette(torch.from_numpy(x[0]).cuda()).cpu().numpy(), odil.dilate(x[:1])[0])


@pytest.mark.gpu
def test_gbuffer_one_pass_equals_three_calls():
    import dirt_b200
    sc = scenes.cube(B=3, H=96, W=96)
    v = torch.from_numpy(sc["vertices"]).cuda()
    f = torch.from_numpy(sc["faces"]).cuda()
    rng = np.random.RandomState(5)
    V = sc["vertices"].shape[1]
    sets = [torch.from_numpy(rng.randn(3, V, 3).astype(np.float32)).cuda() for _ in range(3)]
    bgs = [torch.full((1, 96, 96, 3), float("-inf")).cuda(), torch.zeros(1, 96, 96, 3).cuda(), torch.full((1, 96, 96, 3), float("-inf")).cuda()]
    fused = dirt_b200.rasterise_gbuffer(bgs, v, sets, f)
    for i in range(3):
        single = dirt_b200.rasterise_batch(bgs[i].expand(3, -1, -1, -1).contiguous(), v, sets[i], f)
        assert torch.equal(fused[i], single)
        opx, _, _ = golden.forward(np.broadcast_to(bgs[i].cpu().numpy(), (3, 96, 96, 3)), sc["vertices"], sets[i].cpu().numpy(), sc["faces"])
        fin = np.isfinite(opx)
        np.testing.assert_allclose(fused[i].cpu().numpy()[fin], opx[fin], rtol=1e-5, atol=1e-6)
    dil = dirt_b200.dilate_silhouette(fused[0])
    assert np.array_equal(dil.cpu().numpy(), odil.dilate(fused[0].cpu().numpy()))
