"""geometry / lighting helpers (dirt/lighting.py, dirt/matrices.py).
CPU part: the numpy oracle against hand-derived cases, and the torch composites of
dirt_b200.lighting / dirt_b200.matrices against the oracle (they run on CPU tensors too).
GPU part: the vertex-normal CUDA kernels, forward and backward, against the oracle and against
torch autograd in float64."""
import numpy as np
import pytest
import torch

import scenes
from oracle import lighting as ol
from dirt_b200 import lighting, matrices


def _cube():
    v, f = scenes.cube_mesh()
    return np.asarray(v, np.float64), np.asarray(f, np.int32)


# ---- oracle pinned by hand cases ------------------------------------------------------------

def test_oracle_face_and_vertex_normals_of_one_triangle_and_a_cube():
    tri = np.array([[0., 0., 0.], [1., 0., 0.], [0., 1., 0.]])
    f = np.array([[0, 1, 2]], np.int32)
    assert np.allclose(ol.face_normals(tri, f), [[0., 0., 1.]])
    assert np.allclose(ol.vertex_normals(tri, f), [[0., 0., 1.]] * 3)
    assert np.allclose(ol.face_normals(tri, f[:, ::-1]), [[0., 0., -1.]])        # winding flips the normal
    # two triangles at a right angle sharing the edge v0-v1: normals +z and +y
    v = np.array([[0., 0., 0.], [1., 0., 0.], [0., 1., 0.], [0., 0., 1.]])
    f2 = np.array([[0, 1, 2], [0, 3, 1]], np.int32)
    assert np.allclose(ol.face_normals(v, f2), [[0., 0., 1.], [0., 1., 0.]])
    r = np.sqrt(0.5)
    assert np.allclose(ol.vertex_normals(v, f2), [[0., r, r], [0., r, r], [0., 0., 1.], [0., 1., 0.]])
    assert np.allclose(ol.vertex_normal_sums(v, f2), [[0., 1., 1.], [0., 1., 1.], [0., 0., 1.], [0., 1., 0.]])
    # the pre-split cube of samples/simple.py: every vertex carries its own face's axis-aligned normal
    cv, cf = _cube()
    n = ol.vertex_normals(cv, cf)
    assert np.allclose(np.abs(n).sum(-1), 1.0) and np.allclose(np.abs(n).max(-1), 1.0)
    assert np.allclose(n.reshape(12, 3, 3), ol.face_normals(cv, cf)[:, None, :])


def test_oracle_matrices_known_values():
    R = ol.rodrigues([0., 0., np.pi / 2])
    # K is written [in, out] in dirt/matrices.py:44-48, so a ROW vector turns clockwise about +z
    assert np.allclose(np.array([1., 0., 0., 1.]) @ R, [0., -1., 0., 1.], atol=1e-9)
    # (the scene-fixture helper tests/scenes.py:rodrigues turns the other way; it only poses test meshes)
    assert np.allclose(R[:3, :3], scenes.rodrigues([0., 0., np.pi / 2]).T, atol=1e-9)
    T = ol.translation([1., 2., 3.])
    assert np.allclose(np.array([0., 0., 0., 1.]) @ T, [1., 2., 3., 1.])
    P = ol.perspective_projection(0.1, 10., 0.05, 0.75)
    near_pt = np.array([0.05, 0.0375, -0.1, 1.]) @ P                                  # top-right corner of the near plane
    assert np.allclose(near_pt[:3] / near_pt[3], [1., 1., -1.])
    far_pt = np.array([0., 0., -10., 1.]) @ P
    assert np.allclose(far_pt[2] / far_pt[3], 1.)
    assert np.allclose(ol.compose(T, R), T @ R) and np.allclose(ol.compose(), np.eye(4))
    assert np.allclose(ol.pad_3x3_to_4x4(2 * np.eye(3)), np.diag([2., 2., 2., 1.]))


# ---- torch composites against the oracle (CPU tensors) ----------------------------------------

def test_matrices_match_the_oracle():
    rng = np.random.RandomState(0)
    vecs = rng.randn(5, 3)
    np.testing.assert_allclose(matrices.rodrigues(torch.tensor(vecs)).numpy(), ol.rodrigues(vecs), atol=1e-12)
    np.testing.assert_allclose(matrices.rodrigues(torch.zeros(3, dtype=torch.float64)).numpy(), ol.rodrigues(np.zeros(3)), atol=1e-12)
    np.testing.assert_allclose(matrices.translation(torch.tensor(vecs)).numpy(), ol.translation(vecs))
    np.testing.assert_allclose(matrices.perspective_projection(0.1, 20., 0.05, 0.75).numpy(),
                               ol.perspective_projection(0.1, 20., 0.05, 0.75), rtol=1e-6)
    m3 = rng.randn(2, 3, 3)
    np.testing.assert_allclose(matrices.pad_3x3_to_4x4(torch.tensor(m3)).numpy(), ol.pad_3x3_to_4x4(m3))
    a, b, c = (torch.tensor(rng.randn(4, 4)) for _ in range(3))
    np.testing.assert_allclose(matrices.compose(a, b, c).numpy(), ol.compose(a.numpy(), b.numpy(), c.numpy()))
    assert torch.equal(matrices.compose(), torch.eye(4)) and matrices.compose(a) is a
    # rodrigues is differentiable at exactly zero (the reason for the reference's + 1e-12)
    z = torch.zeros(3, dtype=torch.float64, requires_grad=True)
    matrices.rodrigues(z).sum().backward()
    assert torch.isfinite(z.grad).all()


@pytest.mark.parametrize("double_sided", [True, False])
def test_shading_functions_match_the_oracle(double_sided):
    rng = np.random.RandomState(1)
    B, V, C = 2, 50, 3
    n = rng.randn(B, V, 3)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    p, col = rng.randn(B, V, 3), rng.uniform(0, 1, (B, V, C))
    d = rng.randn(B, 3)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    lc, cam, lp, sh = rng.uniform(0, 1, (B, C)), rng.randn(B, 3) * 3, rng.randn(B, 3) * 3, rng.uniform(1, 8, B)
    t = lambda *xs: [torch.tensor(x) for x in xs]  # noqa: E731
    np.testing.assert_allclose(lighting.diffuse_directional(*t(n, col, d, lc), double_sided=double_sided).numpy(),
                               ol.diffuse_directional(n, col, d, lc, double_sided), rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(lighting.diffuse_point(*t(p, n, col, lp, lc), double_sided=double_sided).numpy(),
                               ol.diffuse_point(p, n, col, lp, lc, double_sided), rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(lighting.specular_directional(*t(p, n, col, d, lc, cam, sh), double_sided=double_sided).numpy(),
                               ol.specular_directional(p, n, col, d, lc, cam, sh, double_sided), rtol=1e-10, atol=1e-14)


def test_shading_known_answers_derived_by_hand():
    """Pins the oracle's shading formulas (and with them the torch composites) by cases worked out on paper from
    dirt/lighting.py:182-343 -- not by another implementation of the same code."""
    n = np.array([[0., 0., 1.]])
    col = np.array([[0.5, 0.25, 1.0]])
    lc = np.array([2., 4., 1.])
    # directional light shining straight down (-z) on an upward normal: cos = n . (-d) = 1; at 60 degrees: 0.5
    np.testing.assert_allclose(ol.diffuse_directional(n, col, np.array([0., 0., -1.]), lc), [[1.0, 1.0, 1.0]], atol=1e-15)
    d60 = np.array([np.sin(np.pi / 3), 0., -np.cos(np.pi / 3)])
    np.testing.assert_allclose(ol.diffuse_directional(n, col, d60, lc), [[0.5, 0.5, 0.5]], atol=1e-15)
    # the same light from below: double sided |cos| = 1, single sided 0
    np.testing.assert_allclose(ol.diffuse_directional(n, col, np.array([0., 0., 1.]), lc, True), [[1.0, 1.0, 1.0]], atol=1e-15)
    np.testing.assert_allclose(ol.diffuse_directional(n, col, np.array([0., 0., 1.]), lc, False), [[0.0, 0.0, 0.0]], atol=1e-15)
    # point light two units above the vertex: the reference's `incident` points FROM the light TO the vertex, so
    # cos = n . (p - l)/|p - l| = -1 -> lit only when double sided (dirt/lighting.py:333-343 as written)
    p = np.array([[1., 2., 3.]])
    lp = np.array([1., 2., 5.])
    np.testing.assert_allclose(ol.diffuse_point(p, n, col, lp, lc, True), [[1.0, 1.0, 1.0]], atol=1e-11)
    np.testing.assert_allclose(ol.diffuse_point(p, n, col, lp, lc, False), [[0.0, 0.0, 0.0]], atol=1e-15)
    # Phong: light straight down, camera straight above: L = +z, R = -L + 2 (n.L) n = +z, T = +z -> cos = 1, any shininess
    cam = np.array([1., 2., 10.])
    np.testing.assert_allclose(ol.specular_directional(p, n, col, np.array([0., 0., -1.]), lc, cam, np.array(7.)), [[1.0, 1.0, 1.0]], atol=1e-10)
    # camera 60 degrees off the mirror direction: cos = 0.5 -> 0.5 ** 3 = 0.125
    cam60 = p[0] + 4. * np.array([np.sin(np.pi / 3), 0., np.cos(np.pi / 3)])
    np.testing.assert_allclose(ol.specular_directional(p, n, col, np.array([0., 0., -1.]), lc, cam60, np.array(3.)),
                               [[0.125, 0.125, 0.125]], atol=1e-10)
    # and the torch composites give the same numbers
    t = lambda *xs: [torch.tensor(x) for x in xs]  # noqa: E731
    np.testing.assert_allclose(lighting.specular_directional(*t(p, n, col, np.array([0., 0., -1.]), lc, cam60, np.array(3.))).numpy(),
                               [[0.125, 0.125, 0.125]], atol=1e-10)
    np.testing.assert_allclose(lighting.diffuse_point(*t(p, n, col, lp, lc), double_sided=False).numpy(), [[0.0, 0.0, 0.0]], atol=1e-15)


def test_split_vertices_by_face_matches_the_oracle():
    v, f = _cube()
    vb = np.stack([v, v * 2.0])
    nv, nf = lighting.split_vertices_by_face(torch.tensor(vb), torch.tensor(f))
    ov, of = ol.split_vertices_by_face(vb, f)
    assert np.array_equal(nv.numpy(), ov) and np.array_equal(nf.numpy(), of) and nf.dtype == torch.int32
    with pytest.raises(ValueError):
        lighting.split_vertices_by_face(torch.tensor(vb), torch.tensor(f[:, :2]))


# ---- the CUDA kernels ---------------------------------------------------------------------------

def _torch_reference(v, f, renormalise):
    p = v[..., :3]
    idx = f.long()
    m = torch.cross(p[:, idx[:, 1]] - p[:, idx[:, 0]], p[:, idx[:, 2]] - p[:, idx[:, 0]], dim=-1)
    n = m / (m.norm(dim=-1, keepdim=True) + 1.e-12)
    s = torch.zeros_like(p)
    for k in range(3):
        s = s.index_add(1, idx[:, k], n)
    return s / (s.norm(dim=-1, keepdim=True) + 1.e-12) if renormalise else s


@pytest.mark.gpu
@pytest.mark.parametrize("D", [3, 4])
def test_vertex_normals_forward_and_backward_on_the_gpu(D):
    mesh_v, mesh_f = scenes.uv_sphere_mesh(24, 32)
    rng = np.random.RandomState(2)
    v = np.stack([mesh_v * s + 0.05 * rng.randn(*mesh_v.shape) for s in (1.0, 1.7, 0.4)]).astype(np.float32)
    if D == 4:
        v = np.concatenate([v, np.ones(v.shape[:2] + (1,), np.float32)], -1)
    f = mesh_f.astype(np.int32)
    vt = torch.tensor(v, device="cuda", requires_grad=True)
    ft = torch.tensor(f, device="cuda")
    out = lighting.vertex_normals(vt, ft)
    np.testing.assert_allclose(out.detach().cpu().numpy(), ol.vertex_normals(v, f), atol=2e-6)
    g = torch.tensor(rng.randn(*out.shape).astype(np.float32), device="cuda")
    out.backward(g)
    v64 = torch.tensor(v, dtype=torch.float64, device="cuda", requires_grad=True)
    _torch_reference(v64, ft, True).backward(g.double())
    exp = v64.grad.cpu().numpy()
    got = vt.grad.cpu().numpy()
    assert got.shape == v.shape
    assert np.abs(got[..., :3] - exp[..., :3]).max() <= 1e-4 * np.abs(exp).max()
    if D == 4:
        assert not got[..., 3].any()
    # leading dims other than one batch axis, and int64 faces
    out2 = lighting.vertex_normals(torch.tensor(v[0], device="cuda"), ft.long())
    assert torch.allclose(out2, out[0].detach(), atol=1e-6)      # float reductions: summation order varies


@pytest.mark.gpu
def test_vertex_normals_pre_split_keeps_the_face_normal():
    v, f = _cube()
    sv, sf = ol.split_vertices_by_face(v[None].astype(np.float32), f)
    vt = torch.tensor(sv, device="cuda", requires_grad=True)
    ft = torch.tensor(sf, device="cuda")
    out = lighting.vertex_normals_pre_split(vt, ft)
    np.testing.assert_allclose(out.detach().cpu().numpy(), ol.vertex_normal_sums(sv, sf), atol=1e-6)
    np.testing.assert_allclose(out.detach().cpu().numpy()[0].reshape(-1, 3, 3)[:, 0], ol.face_normals(v, f), atol=1e-6)
    g = torch.randn_like(out)
    out.backward(g)
    v64 = torch.tensor(sv, dtype=torch.float64, device="cuda", requires_grad=True)
    _torch_reference(v64, ft, False).backward(g.double())
    assert np.abs(vt.grad.cpu().numpy() - v64.grad.cpu().numpy()).max() <= 1e-4 * np.abs(v64.grad.cpu().numpy()).max()


@pytest.mark.gpu
def test_vertex_normals_edge_cases():
    ft = torch.zeros((0, 3), dtype=torch.int32, device="cuda")
    out = lighting.vertex_normals(torch.rand(2, 5, 3, device="cuda"), ft)                 # no faces: all zero
    assert out.shape == (2, 5, 3) and not out.any()
    bad = torch.tensor([[0, 1, 7], [0, 1, 2]], dtype=torch.int32, device="cuda")           # out-of-range face is skipped
    v = torch.tensor([[0., 0., 0.], [1., 0., 0.], [0., 1., 0.]], device="cuda")
    assert torch.allclose(lighting.vertex_normals(v, bad), torch.tensor([[0., 0., 1.]] * 3, device="cuda"))
    with pytest.raises(RuntimeError):
        lighting.vertex_normals(torch.rand(5, 3), torch.zeros((1, 3), dtype=torch.int32))    # CPU tensors: no fallback


@pytest.mark.gpu
@pytest.mark.parametrize("double_sided", [True, False])
@pytest.mark.parametrize("batched_light", [False, True])
@pytest.mark.parametrize("C", [1, 3, 5])
def test_fused_diffuse_directional_matches_the_composite_and_the_oracle(double_sided, batched_light, C):
    """csrc/lighting.cu: one pass forward, one backward, against the reference's formula in torch ops (which
    the CPU test above pins to the numpy oracle) -- values and both gradients."""
    rng = np.random.RandomState(3)
    B, V = 3, 257
    n = rng.randn(B, V, 3).astype(np.float32)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    col = rng.uniform(0, 1, (B, V, C)).astype(np.float32)
    d = rng.randn(*((B, 3) if batched_light else (3,))).astype(np.float32)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    lc = rng.uniform(0.2, 1, (B, C) if batched_light else (C,)).astype(np.float32)
    g = rng.randn(B, V, C).astype(np.float32)
    tn, tc = (torch.tensor(x, device="cuda", requires_grad=True) for x in (n, col))
    td, tl, tg = (torch.tensor(x, device="cuda") for x in (d, lc, g))
    assert lighting._fusable(tn, tc, td, tl)
    out = lighting.diffuse_directional(tn, tc, td, tl, double_sided=double_sided)
    out.backward(tg)
    rn, rc = (torch.tensor(x, device="cuda", requires_grad=True) for x in (n, col))
    dd = td if batched_light else td.expand(B, 3)
    ll = tl if batched_light else tl.expand(B, C)
    cos = torch.matmul(rn, -dd.unsqueeze(-1))
    ref = ll.unsqueeze(-2) * rc * (cos.abs() if double_sided else cos.clamp_min(0.))
    ref.backward(tg)
    torch.testing.assert_close(out, ref, rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(tn.grad, rn.grad, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(tc.grad, rc.grad, rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(out.detach().cpu().numpy(),
                               ol.diffuse_directional(n.astype(np.float64), col.astype(np.float64), np.broadcast_to(d, (B, 3)).astype(np.float64),
                                                      np.broadcast_to(lc, (B, C)).astype(np.float64), double_sided), rtol=1e-5, atol=1e-6)
    # a gradient w.r.t. the light, or CPU tensors, take the torch formula
    assert not lighting._fusable(tn, tc, td.clone().requires_grad_(True), tl)
    assert not lighting._fusable(tn.cpu(), tc.cpu(), td.cpu(), tl.cpu())


def _shade_inputs(rng, B, V, C, batched):
    p = rng.uniform(-1, 1, (B, V, 3)).astype(np.float32)
    n = rng.randn(B, V, 3).astype(np.float32)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    col = rng.uniform(0, 1, (B, V, C)).astype(np.float32)
    lc = rng.uniform(0.2, 1, (B, C) if batched else (C,)).astype(np.float32)
    g = rng.randn(B, V, C).astype(np.float32)
    return p, n, col, lc, g


@pytest.mark.gpu
@pytest.mark.parametrize("double_sided", [True, False])
@pytest.mark.parametrize("batched", [False, True])
@pytest.mark.parametrize("C", [1, 3])
def test_fused_diffuse_point_matches_the_composite_and_the_oracle(double_sided, batched, C):
    """csrc/lighting.cu k_vertex_shade<false, *>: values and the three gradients against the reference's formula in
    torch ops (pinned to the numpy oracle by the CPU test above)."""
    rng = np.random.RandomState(11)
    B, V = 3, 301
    p, n, col, lc, g = _shade_inputs(rng, B, V, C, batched)
    lp = rng.uniform(2, 3, (B, 3) if batched else (3,)).astype(np.float32)
    tp, tn, tc = (torch.tensor(x, device="cuda", requires_grad=True) for x in (p, n, col))
    tlp, tl, tg = (torch.tensor(x, device="cuda") for x in (lp, lc, g))
    assert lighting._fusable_shade(tp, tn, tc, (tlp, tl))
    out = lighting.diffuse_point(tp, tn, tc, tlp, tl, double_sided=double_sided)
    out.backward(tg)
    rp, rn, rc = (torch.tensor(x, device="cuda", requires_grad=True) for x in (p, n, col))
    # the same call with a light that wants a gradient takes the torch formula
    ref = lighting.diffuse_point(rp, rn, rc, tlp.clone().requires_grad_(True), tl, double_sided=double_sided)
    ref.backward(tg)
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(tp.grad, rp.grad, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(tn.grad, rn.grad, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(tc.grad, rc.grad, rtol=1e-5, atol=1e-6)
    want = ol.diffuse_point(p.astype(np.float64), n.astype(np.float64), col.astype(np.float64), np.broadcast_to(lp, (B, 3)).astype(np.float64),
                            np.broadcast_to(lc, (B, C)).astype(np.float64), double_sided)
    np.testing.assert_allclose(out.detach().cpu().numpy(), want, rtol=1e-5, atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("double_sided", [True, False])
@pytest.mark.parametrize("batched", [False, True])
@pytest.mark.parametrize("shininess", [1.0, 4.0, 12.5])
def test_fused_specular_directional_matches_the_composite_and_the_oracle(double_sided, batched, shininess):
    """csrc/lighting.cu k_vertex_shade<true, *>: Phong term, values and the three gradients."""
    rng = np.random.RandomState(12)
    B, V, C = 3, 301, 3
    p, n, col, lc, g = _shade_inputs(rng, B, V, C, batched)
    d = rng.randn(*((B, 3) if batched else (3,))).astype(np.float32)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    cam = rng.uniform(3, 4, (B, 3) if batched else (3,)).astype(np.float32)
    sh = np.full((B,), shininess, np.float32) + (np.arange(B, dtype=np.float32) if batched else 0)
    tp, tn, tc = (torch.tensor(x, device="cuda", requires_grad=True) for x in (p, n, col))
    td, tl, tcam, tg = (torch.tensor(x, device="cuda") for x in (d, lc, cam, g))
    tsh = torch.tensor(sh, device="cuda") if batched else shininess          # a Python number is the usual call
    out = lighting.specular_directional(tp, tn, tc, td, tl, tcam, tsh, double_sided=double_sided)
    assert out.grad_fn.name().startswith("_VertexShade")
    out.backward(tg)
    rp, rn, rc = (torch.tensor(x, device="cuda", requires_grad=True) for x in (p, n, col))
    rsh = torch.tensor(sh if batched else np.float32(shininess), device="cuda")
    ref = lighting.specular_directional(rp, rn, rc, td.clone().requires_grad_(True), tl, tcam, rsh, double_sided=double_sided)
    assert not ref.grad_fn.name().startswith("_VertexShade")
    ref.backward(tg)
    torch.testing.assert_close(out, ref, rtol=2e-5, atol=1e-6)
    torch.testing.assert_close(tp.grad, rp.grad, rtol=2e-4, atol=2e-5)
    torch.testing.assert_close(tn.grad, rn.grad, rtol=2e-4, atol=2e-5)
    torch.testing.assert_close(tc.grad, rc.grad, rtol=2e-5, atol=1e-6)
    want = ol.specular_directional(p.astype(np.float64), n.astype(np.float64), col.astype(np.float64), np.broadcast_to(d, (B, 3)).astype(np.float64),
                                   np.broadcast_to(lc, (B, C)).astype(np.float64), np.broadcast_to(cam, (B, 3)).astype(np.float64),
                                   sh.astype(np.float64), double_sided)
    np.testing.assert_allclose(out.detach().cpu().numpy(), want, rtol=1e-4, atol=1e-6)
