"""The reference's samples/deferred.py on dirt_b200: a cube rendered into a G-buffer (world
position, albedo, world normal), silhouette dilation, then ambient + diffuse + Phong specular
shading per pixel (samples/deferred.py:15-118).  Here the three attribute sets come out of ONE
visibility pass (dirt_b200.rasterise_gbuffer) and the mask / max-pool / blend of the dilation is one
fused kernel (dirt_b200.dilate_silhouette); both are differentiable.

    python samples/deferred.py [out.png]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dirt_b200 as dirt                      # noqa: E402
from dirt_b200 import lighting, matrices      # noqa: E402
from simple import build_cube                 # noqa: E402

frame_width, frame_height = 640, 480


def unit(v):
    return v / v.norm()


def main(out=None, width=frame_width, height=frame_height):
    dev = torch.device("cuda")
    vertices, faces = build_cube()
    vertices = torch.tensor(vertices, dtype=torch.float32, device=dev)
    vertices, faces = lighting.split_vertices_by_face(vertices, torch.tensor(faces, dtype=torch.int32, device=dev))
    vertices = vertices.clone().requires_grad_(True)
    colours = torch.ones_like(vertices)
    world = torch.cat([vertices, torch.ones_like(vertices[:, -1:])], dim=1) @ matrices.rodrigues(torch.tensor([0., 0.5, 0.], device=dev))
    normals_world = lighting.vertex_normals_pre_split(world, faces)
    view = matrices.compose(matrices.translation(torch.tensor([0., -1.5, -3.5], device=dev)),
                            matrices.rodrigues(torch.tensor([-0.3, 0., 0.], device=dev)))
    projection = matrices.perspective_projection(near=0.1, far=20., right=0.1, aspect=float(height) / width).to(dev)
    clip = world @ view @ projection

    inf = torch.full((1, height, width, 3), float("-inf"), device=dev)
    positions, albedo, normals = dirt.rasterise_gbuffer(
        backgrounds=[inf, torch.zeros(1, height, width, 3, device=dev), inf],
        vertices=clip[None], attribute_sets=[world[None, :, :3], colours[None], normals_world[None]], faces=faces[None])
    # dilate position and normal at the silhouette: the image does not change, the gradients of
    # pixels just outside the silhouette become correct
    positions = dirt.dilate_silhouette(positions)[0]
    normals = dirt.dilate_silhouette(normals)[0]
    albedo = albedo[0]

    ambient = albedo * torch.tensor([0.2, 0.2, 0.2], device=dev)
    light_direction = unit(torch.tensor([1., -0.3, -0.5], device=dev))
    diffuse = lighting.diffuse_directional(normals.reshape(-1, 3), albedo.reshape(-1, 3), light_direction,
                                           light_color=torch.tensor([1., 0., 0.], device=dev), double_sided=False).reshape(height, width, 3)
    camera_position_world = torch.linalg.inv(view)[3, :3]
    specular = lighting.specular_directional(positions.reshape(-1, 3), normals.reshape(-1, 3), albedo.reshape(-1, 3), light_direction,
                                             light_color=torch.tensor([1., 1., 1.], device=dev), camera_position=camera_position_world,
                                             shininess=torch.tensor(6., device=dev), double_sided=False).reshape(height, width, 3)
    covered = (albedo.sum(-1, keepdim=True) > 0)
    pixels = torch.where(covered, diffuse + specular + ambient, torch.zeros_like(ambient))
    pixels.sum().backward()
    if out:
        from _png import save_png
        save_png(out, pixels.detach().cpu().numpy())
    return pixels.detach(), vertices.grad


if __name__ == "__main__":
    px, grad = main(sys.argv[1] if len(sys.argv) > 1 else "deferred.png")
    print("rendered %dx%d, %.1f%% covered; |d sum(pixels) / d vertices| max %.3e" % (
        px.shape[1], px.shape[0], 100 * float((px.sum(-1) > 0).float().mean()), float(grad.abs().max())))
