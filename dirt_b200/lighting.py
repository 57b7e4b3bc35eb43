"""Geometry and lighting helpers with the names, argument order and arithmetic of the reference's
dirt/lighting.py, on torch tensors.

`vertex_normals` is the expensive one -- the reference builds a sparse tensor with 9*F*N entries
and reduces it (dirt/lighting.py:74-88) -- and runs as hand-written CUDA kernels behind the C ABI
(`dirt_vertex_normals`, `dirt_vertex_normals_grad`; dirt_b200/csrc/geometry.cu) with autograd wired.
`diffuse_directional`, the shading of the c5 optimisation step, is one fused CUDA pass per direction
(`dirt_diffuse_directional(_grad)`, csrc/lighting.cu: its torch composites took a third of that step,
profiles/r2_c5_step_breakdown.json); `diffuse_point` and `specular_directional` likewise
(`dirt_diffuse_point(_grad)`, `dirt_specular_directional(_grad)`).  Calls the fused kernels do not serve (CPU
tensors, other dtypes, a gradient w.r.t. a light, the camera or the shininess) take the reference's formula in
torch operations, formula by formula as the reference has them, including its `+ 1e-12` terms.
"""
import ctypes

import torch

from . import _lib
from ._lib import InvalidArgumentError


def _prepare_vertices_and_faces(vertices, faces):
    # dirt/lighting.py:12-21
    vertices = torch.as_tensor(vertices)
    faces = torch.as_tensor(faces, device=vertices.device)
    if faces.dtype != torch.int32:
        if faces.dtype != torch.int64:
            raise InvalidArgumentError(-1, "faces must be int32 or int64")
        faces = faces.to(torch.int32)
    if faces.dim() != 2 or faces.shape[1] != 3:
        raise InvalidArgumentError(-1, "faces must be [face count, 3]")
    if vertices.dim() < 2 or vertices.shape[-1] not in (3, 4):
        raise InvalidArgumentError(-1, "vertices must be [*, vertex count, 3] or [*, vertex count, 4]")
    return vertices, faces


class _VertexNormals(torch.autograd.Function):
    @staticmethod
    def forward(ctx, vertices, faces):
        lib = _lib.load()
        if not vertices.is_cuda:
            raise RuntimeError("dirt_b200.lighting.vertex_normals needs CUDA tensors; there is no CPU fallback")
        lead, (V, D) = vertices.shape[:-2], vertices.shape[-2:]
        v = vertices.to(torch.float32).reshape(-1, V, D).contiguous()
        f = faces.contiguous()
        B = v.shape[0]
        sums = torch.empty((B, V, 3), dtype=torch.float32, device=v.device)
        normals = torch.empty((B, V, 3), dtype=torch.float32, device=v.device)
        with torch.cuda.device(v.device):
            _lib.check(lib.dirt_vertex_normals(v.data_ptr(), B, V, D, f.data_ptr(), f.shape[0], sums.data_ptr(),
                                               normals.data_ptr(), torch.cuda.current_stream(v.device).cuda_stream))
        ctx.save_for_backward(v, f, sums)
        ctx.shape = tuple(vertices.shape)
        return normals.reshape(*lead, V, 3)

    @staticmethod
    def backward(ctx, grad_normals):
        lib = _lib.load()
        v, f, sums = ctx.saved_tensors
        B, V, D = v.shape
        g = grad_normals.to(torch.float32).reshape(B, V, 3).contiguous()
        grad_sums = torch.empty_like(sums)
        grad_vertices = torch.empty_like(v)
        with torch.cuda.device(v.device):
            _lib.check(lib.dirt_vertex_normals_grad(v.data_ptr(), B, V, D, f.data_ptr(), f.shape[0], sums.data_ptr(),
                                                    g.data_ptr(), grad_sums.data_ptr(), grad_vertices.data_ptr(),
                                                    torch.cuda.current_stream(v.device).cuda_stream))
        return grad_vertices.reshape(ctx.shape), None


def vertex_normals(vertices, faces, name=None):
    """Vertex normals of a batch of meshes with common topology (dirt/lighting.py:34-93).
    vertices [*, V, 3|4] (w ignored), faces [F, 3]; returns [*, V, 3]: for each vertex the
    normalised sum of the unit normals of the faces that use it."""
    del name
    vertices, faces = _prepare_vertices_and_faces(vertices, faces)
    return _VertexNormals.apply(vertices, faces)


def vertex_normals_pre_split(vertices, faces, name=None, static=False):
    """dirt/lighting.py:101-133: vertex normals of a mesh whose vertices are each used by exactly
    one face (e.g. the output of `split_vertices_by_face`).  NOTE the reference does NOT
    re-normalise here: a vertex gets its face's unit normal, and an unused vertex gets zeros --
    which is what the general kernel's `sums` output holds, so this returns it."""
    del name, static
    vertices, faces = _prepare_vertices_and_faces(vertices, faces)
    return _FaceNormalSums.apply(vertices, faces)


class _FaceNormalSums(torch.autograd.Function):
    """The un-normalised per-vertex sum of unit face normals (first kernel of `vertex_normals`)."""

    @staticmethod
    def forward(ctx, vertices, faces):
        lib = _lib.load()
        if not vertices.is_cuda:
            raise RuntimeError("dirt_b200.lighting.vertex_normals_pre_split needs CUDA tensors; there is no CPU fallback")
        lead, (V, D) = vertices.shape[:-2], vertices.shape[-2:]
        v = vertices.to(torch.float32).reshape(-1, V, D).contiguous()
        f = faces.contiguous()
        B = v.shape[0]
        sums = torch.empty((B, V, 3), dtype=torch.float32, device=v.device)
        scratch = torch.empty((B, V, 3), dtype=torch.float32, device=v.device)
        with torch.cuda.device(v.device):
            _lib.check(lib.dirt_vertex_normals(v.data_ptr(), B, V, D, f.data_ptr(), f.shape[0], sums.data_ptr(),
                                               scratch.data_ptr(), torch.cuda.current_stream(v.device).cuda_stream))
        ctx.save_for_backward(v, f)
        ctx.shape = tuple(vertices.shape)
        return sums.reshape(*lead, V, 3)

    @staticmethod
    def backward(ctx, grad_sums):
        lib = _lib.load()
        v, f = ctx.saved_tensors
        B, V, D = v.shape
        g = grad_sums.to(torch.float32).reshape(B, V, 3).contiguous()
        grad_vertices = torch.empty_like(v)
        with torch.cuda.device(v.device):
            # sums = NULL: skip the final normalisation stage (include/dirt_b200.h)
            _lib.check(lib.dirt_vertex_normals_grad(v.data_ptr(), B, V, D, f.data_ptr(), f.shape[0], None, g.data_ptr(), None,
                                                    grad_vertices.data_ptr(), torch.cuda.current_stream(v.device).cuda_stream))
        return grad_vertices.reshape(ctx.shape), None


def split_vertices_by_face(vertices, faces, name=None):
    """dirt/lighting.py:136-179: an equivalent mesh in which every vertex is used by exactly one
    face.  Returns (new_vertices [*, 3F, D], new_faces [F, 3])."""
    del name
    vertices, faces = _prepare_vertices_and_faces(vertices, faces)
    new_vertices = vertices.index_select(-2, faces.reshape(-1).long())
    new_faces = torch.arange(faces.shape[0] * 3, dtype=torch.int32, device=faces.device).reshape(-1, 3)
    return new_vertices, new_faces


def _cosine_term(cosines, double_sided):
    return cosines.abs() if double_sided else cosines.clamp_min(0.)


class _DiffuseDirectional(torch.autograd.Function):
    """One CUDA pass forward, one backward (csrc/lighting.cu) for the shading formula below."""

    @staticmethod
    def forward(ctx, n, c, d, lc, double_sided):
        lib = _lib.load()
        shape = c.shape
        V, C = shape[-2], shape[-1]
        B = c.numel() // max(V * C, 1)
        n, c = n.contiguous(), c.contiguous()
        d, lc = d.contiguous(), lc.contiguous()
        out = torch.empty_like(c)
        dbs, lbs = (3 if d.dim() > 1 else 0), (C if lc.dim() > 1 else 0)
        with torch.cuda.device(c.device):
            _lib.check(lib.dirt_diffuse_directional(n.data_ptr(), c.data_ptr(), d.data_ptr(), dbs, lc.data_ptr(), lbs,
                                                    int(bool(double_sided)), B, V, C, out.data_ptr(),
                                                    torch.cuda.current_stream(c.device).cuda_stream))
        ctx.save_for_backward(n, c, d, lc)
        ctx.meta = (B, V, C, dbs, lbs, bool(double_sided))
        return out

    @staticmethod
    def backward(ctx, g):
        n, c, d, lc = ctx.saved_tensors
        B, V, C, dbs, lbs, double_sided = ctx.meta
        lib = _lib.load()
        g = g.contiguous()
        gn = torch.empty_like(n) if ctx.needs_input_grad[0] else None
        gc = torch.empty_like(c) if ctx.needs_input_grad[1] else None
        with torch.cuda.device(c.device):
            _lib.check(lib.dirt_diffuse_directional_grad(n.data_ptr(), c.data_ptr(), d.data_ptr(), dbs, lc.data_ptr(), lbs,
                                                         int(double_sided), B, V, C, g.data_ptr(),
                                                         gn.data_ptr() if gn is not None else None,
                                                         gc.data_ptr() if gc is not None else None,
                                                         torch.cuda.current_stream(c.device).cuda_stream))
        return gn, gc, None, None, None


def _fusable(n, c, d, lc):
    """The fused kernels serve the usual call: float32 CUDA tensors, normals / colours [*, V, 3|C] with the same
    leading dims, the light either one for all ([3], [C]) or one per leading index ([*, 3], [*, C]), and no gradient
    wanted for the light itself."""
    if not (n.is_cuda and c.is_cuda and d.is_cuda and lc.is_cuda):
        return False
    if any(t.dtype != torch.float32 for t in (n, c, d, lc)) or d.requires_grad or lc.requires_grad:
        return False
    if n.dim() < 2 or c.dim() != n.dim() or n.shape[:-1] != c.shape[:-1] or n.shape[-1] != 3 or not (1 <= c.shape[-1] <= 16):
        return False
    lead = n.shape[:-2]
    return (d.shape == (3,) or d.shape == lead + (3,)) and (lc.shape == c.shape[-1:] or lc.shape == lead + c.shape[-1:])


def diffuse_directional(vertex_normals, vertex_colors, light_direction, light_color, double_sided=True, name=None):
    """Lambertian reflectance under one directional light (dirt/lighting.py:182-225).
    vertex_normals [*, V, 3], vertex_colors [*, V, C], light_direction [*, 3], light_color [*, C].
    CUDA float32 inputs run as one fused kernel per direction (csrc/lighting.cu); anything else (CPU tensors, other
    dtypes, broadcasting beyond the above, a gradient w.r.t. the light) takes the reference's formula in torch ops."""
    del name
    n, c, d, lc = (torch.as_tensor(x) for x in (vertex_normals, vertex_colors, light_direction, light_color))
    if _fusable(n, c, d, lc):
        return _DiffuseDirectional.apply(n, c, d, lc, double_sided)
    cosines = torch.matmul(n, -d.unsqueeze(-1))                       # [*, V, 1]
    return lc.unsqueeze(-2) * c * _cosine_term(cosines, double_sided)


class _VertexShade(torch.autograd.Function):
    """`diffuse_point` / `specular_directional` as one CUDA pass forward and one backward (csrc/lighting.cu)."""

    @staticmethod
    def _args(specular, p, n, c, lv, lc, cam, sh, double_sided):
        V, C = c.shape[-2], c.shape[-1]
        a = _lib.DirtVertexShadeArgs()
        a.batch, a.vertex_count, a.channels, a.double_sided = c.numel() // max(V * C, 1), V, C, int(bool(double_sided))
        a.positions, a.normals, a.colours = p.data_ptr(), n.data_ptr(), c.data_ptr()
        a.light_vector, a.light_vector_bstride = lv.data_ptr(), (3 if lv.dim() > 1 else 0)
        a.light_color, a.light_color_bstride = lc.data_ptr(), (C if lc.dim() > 1 else 0)
        if specular:
            a.camera_position, a.camera_position_bstride = cam.data_ptr(), (3 if cam.dim() > 1 else 0)
            a.shininess, a.shininess_bstride = sh.data_ptr(), (1 if sh.dim() > 0 else 0)
        return a

    @staticmethod
    def forward(ctx, specular, p, n, c, lv, lc, cam, sh, double_sided):
        lib = _lib.load()
        tensors = [t.contiguous() if t is not None else None for t in (p, n, c, lv, lc, cam, sh)]
        p, n, c, lv, lc, cam, sh = tensors
        out = torch.empty_like(c)
        a = _VertexShade._args(specular, p, n, c, lv, lc, cam, sh, double_sided)
        a.out = out.data_ptr()
        with torch.cuda.device(c.device):
            fn = lib.dirt_specular_directional if specular else lib.dirt_diffuse_point
            _lib.check(fn(ctypes.byref(a), torch.cuda.current_stream(c.device).cuda_stream))
        ctx.save_for_backward(*[t for t in tensors if t is not None])
        ctx.meta = (specular, bool(double_sided))
        return out

    @staticmethod
    def backward(ctx, g):
        specular, double_sided = ctx.meta
        saved = list(ctx.saved_tensors)
        p, n, c, lv, lc = saved[:5]
        cam, sh = (saved[5], saved[6]) if specular else (None, None)
        lib = _lib.load()
        g = g.contiguous()
        gp = torch.empty_like(p) if ctx.needs_input_grad[1] else None
        gn = torch.empty_like(n) if ctx.needs_input_grad[2] else None
        gc = torch.empty_like(c) if ctx.needs_input_grad[3] else None
        a = _VertexShade._args(specular, p, n, c, lv, lc, cam, sh, double_sided)
        a.grad_out = g.data_ptr()
        a.grad_positions = gp.data_ptr() if gp is not None else None
        a.grad_normals = gn.data_ptr() if gn is not None else None
        a.grad_colours = gc.data_ptr() if gc is not None else None
        with torch.cuda.device(c.device):
            fn = lib.dirt_specular_directional_grad if specular else lib.dirt_diffuse_point_grad
            _lib.check(fn(ctypes.byref(a), torch.cuda.current_stream(c.device).cuda_stream))
        return None, gp, gn, gc, None, None, None, None, None


def _fusable_shade(p, n, c, lights, per_batch_scalars=()):
    """As _fusable, for the helpers that also take positions: `lights` are the [3] / [*, 3] / [C] / [*, C] inputs,
    `per_batch_scalars` the [] / [*] ones; none of them may want a gradient."""
    small = tuple(lights) + tuple(per_batch_scalars)
    if not all(t.is_cuda and t.dtype == torch.float32 for t in (p, n, c) + small) or any(t.requires_grad for t in small):
        return False
    if c.dim() < 2 or p.shape != n.shape or p.dim() != c.dim() or p.shape[:-1] != c.shape[:-1] or p.shape[-1] != 3:
        return False
    if not (1 <= c.shape[-1] <= 16):
        return False
    lead = c.shape[:-2]
    for t in lights:
        if t.dim() == 0 or not (t.shape == (t.shape[-1],) or t.shape == lead + (t.shape[-1],)) or t.shape[-1] not in (3, c.shape[-1]):
            return False
    return all(t.dim() == 0 or t.shape == lead for t in per_batch_scalars)


def specular_directional(vertex_positions, vertex_normals, vertex_reflectivities, light_direction, light_color,
                         camera_position, shininess, double_sided=True, name=None):
    """Phong reflectance under one directional light (dirt/lighting.py:228-288), including the
    reference's operator precedence: the 1e-12 is added AFTER the division by the norm.
    CUDA float32 inputs run as one fused kernel per direction (csrc/lighting.cu), everything else in torch ops."""
    del name
    p, n, r, d, lc, cam, sh = (torch.as_tensor(x) for x in (vertex_positions, vertex_normals, vertex_reflectivities,
                                                            light_direction, light_color, camera_position, shininess))
    if r.is_cuda and not sh.is_cuda and not sh.requires_grad:
        sh = sh.to(device=r.device, dtype=torch.float32)          # the usual call passes a Python number
    if _fusable_shade(p, n, r, (d, lc, cam), (sh,)) and d.shape[-1] == 3 and cam.shape[-1] == 3 and lc.shape[-1] == r.shape[-1]:
        return _VertexShade.apply(True, p, n, r, d, lc, cam, sh, double_sided)
    to_light = -d
    reflected = -to_light.unsqueeze(-2) + 2. * torch.matmul(n, to_light.unsqueeze(-1)) * n
    to_camera = cam.unsqueeze(-2) - p
    cosines = ((to_camera / to_camera.norm(dim=-1, keepdim=True) + 1.e-12) * reflected).sum(dim=-1, keepdim=True)
    return lc.unsqueeze(-2) * r * torch.pow(_cosine_term(cosines, double_sided), sh[..., None, None])


def diffuse_point(vertex_positions, vertex_normals, vertex_colors, light_position, light_color, double_sided=True, name=None):
    """Lambertian reflectance under one point light (dirt/lighting.py:291-343).
    CUDA float32 inputs run as one fused kernel per direction (csrc/lighting.cu), everything else in torch ops."""
    del name
    p, n, c, lp, lc = (torch.as_tensor(x) for x in (vertex_positions, vertex_normals, vertex_colors, light_position, light_color))
    if _fusable_shade(p, n, c, (lp, lc)) and lp.shape[-1] == 3 and lc.shape[-1] == c.shape[-1]:
        return _VertexShade.apply(False, p, n, c, lp, lc, None, None, double_sided)
    relative = p - lp.unsqueeze(-2)
    incident = relative / (relative.norm(dim=-1, keepdim=True) + 1.e-12)
    cosines = (n * incident).sum(dim=-1)
    return lc.unsqueeze(-2) * c * _cosine_term(cosines, double_sided).unsqueeze(-1)
