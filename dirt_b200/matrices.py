"""Homogeneous transform matrices for ROW vectors (x' = x @ M), with the names and arithmetic of
the reference's dirt/matrices.py, on torch tensors; differentiable by torch.
Matrices are indexed [*, in, out]."""
import torch


def pad_3x3_to_4x4(matrix, name=None):
    """dirt/matrices.py:119-143: [*, 3, 3] -> [*, 4, 4] with a zero column, a zero row and a 1."""
    del name
    m = torch.as_tensor(matrix)
    top = torch.cat([m, torch.zeros_like(m[..., :, :1])], dim=-1)
    bottom = torch.cat([torch.zeros_like(m[..., :1, :]), torch.ones_like(m[..., :1, :1])], dim=-1)
    return torch.cat([top, bottom], dim=-2)


def rodrigues(vectors, name=None):
    """dirt/matrices.py:15-56: angle-axis vectors [*, 3] -> rotation matrices [*, 4, 4].  As in the
    reference, 1e-12 is added to the vector first (finite derivative at zero)."""
    del name
    v = torch.as_tensor(vectors)
    if not v.is_floating_point():
        v = v.to(torch.float32)
    v = v + 1.e-12
    norms = v.norm(dim=-1, keepdim=True)
    v = v / norms
    norms = norms[..., 0]
    z = torch.zeros_like(v[..., 0])
    # K indexed [*, in, out]
    K = torch.stack([torch.stack([z, -v[..., 2], v[..., 1]], dim=-1),
                     torch.stack([v[..., 2], z, -v[..., 0]], dim=-1),
                     torch.stack([-v[..., 1], v[..., 0], z], dim=-1)], dim=-2)
    c = torch.cos(norms)[..., None, None]
    s = torch.sin(norms)[..., None, None]
    eye = torch.eye(3, dtype=v.dtype, device=v.device)
    return pad_3x3_to_4x4(c * eye + (1 - c) * v[..., :, None] * v[..., None, :] + s * K)


def translation(x, name=None):
    """dirt/matrices.py:59-83: displacements [*, 3] -> [*, 4, 4] (the displacement is the last ROW)."""
    del name
    x = torch.as_tensor(x)
    if not x.is_floating_point():
        x = x.to(torch.float32)
    zeros = torch.zeros_like(x[..., 0])
    ones = torch.ones_like(zeros)
    return torch.stack([torch.stack([ones, zeros, zeros, zeros], dim=-1),
                        torch.stack([zeros, ones, zeros, zeros], dim=-1),
                        torch.stack([zeros, zeros, ones, zeros], dim=-1),
                        torch.stack([x[..., 0], x[..., 1], x[..., 2], ones], dim=-1)], dim=-2)


def perspective_projection(near, far, right, aspect, name=None):
    """dirt/matrices.py:86-116: OpenGL-convention projection (camera looks along -z), 4x4 float32,
    for row vectors (the transpose of the usual column-vector matrix)."""
    del name
    near, far, right, aspect = (torch.as_tensor(x, dtype=torch.float32) for x in (near, far, right, aspect))
    top = right * aspect
    zero, one = torch.zeros_like(near), torch.ones_like(near)
    elements = torch.stack([torch.stack([near / right, zero, zero, zero]),
                            torch.stack([zero, near / top, zero, zero]),
                            torch.stack([zero, zero, -(far + near) / (far - near), -2. * far * near / (far - near)]),
                            torch.stack([zero, zero, -one, zero])])          # [out, in]
    return elements.transpose(0, 1)


def compose(*matrices):
    """dirt/matrices.py:146-170: product of the given matrices; the first is applied first."""
    if len(matrices) == 0:
        return torch.eye(4)
    result = torch.as_tensor(matrices[0])
    for m in matrices[1:]:
        result = torch.matmul(result, torch.as_tensor(m))
    return result
