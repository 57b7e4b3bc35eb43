"""The reference's samples/simple.py on dirt_b200: a Gouraud-lit cube, direct shading, then the
gradient of a loss on the pixels w.r.t. the object-space vertices (what the reference's sample
stops short of).  Same scene, transforms and lighting (samples/simple.py:15-75); torch instead of
TensorFlow, a PNG instead of an OpenCV window.

    python samples/simple.py [out.png]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dirt_b200 as dirt                      # noqa: E402
from dirt_b200 import lighting, matrices      # noqa: E402

frame_width, frame_height = 640, 480


def build_cube():
    vertices = [[x, y, z] for z in [-1, 1] for y in [-1, 1] for x in [-1, 1]]
    quads = [[0, 1, 3, 2], [4, 5, 7, 6], [1, 5, 4, 0], [2, 6, 7, 3], [4, 6, 2, 0], [3, 7, 5, 1]]
    triangles = sum([[[a, b, c], [c, d, a]] for [a, b, c, d] in quads], [])
    return vertices, triangles


def render(cube_vertices_object, cube_faces, width=frame_width, height=frame_height):
    dev = cube_vertices_object.device
    cube_vertex_colors = torch.ones_like(cube_vertices_object)
    homogeneous = torch.cat([cube_vertices_object, torch.ones_like(cube_vertices_object[:, -1:])], dim=1)
    # object -> world: rotate about the vertical axis
    world = homogeneous @ matrices.rodrigues(torch.tensor([0., 0.5, 0.], device=dev))
    normals_world = lighting.vertex_normals_pre_split(world, cube_faces)
    # world -> camera (the camera looks along -z) -> clip
    view = matrices.compose(matrices.translation(torch.tensor([0., -1.5, -3.5], device=dev)),
                            matrices.rodrigues(torch.tensor([-0.3, 0., 0.], device=dev)))
    projection = matrices.perspective_projection(near=0.1, far=20., right=0.1, aspect=float(height) / width).to(dev)
    clip = world @ view @ projection
    lit = lighting.diffuse_directional(normals_world, cube_vertex_colors,
                                       light_direction=torch.tensor([1., 0., 0.], device=dev),
                                       light_color=torch.tensor([1., 1., 1.], device=dev)) * 0.8 + cube_vertex_colors * 0.2
    return dirt.rasterise(vertices=clip, faces=cube_faces, vertex_colors=lit,
                          background=torch.zeros(height, width, 3, device=dev), width=width, height=height, channels=3)


def main(out=None, width=frame_width, height=frame_height):
    dev = torch.device("cuda")
    vertices, faces = build_cube()
    vertices = torch.tensor(vertices, dtype=torch.float32, device=dev)
    # shared vertices are replicated, so normals are per face instead of smoothed
    vertices, faces = lighting.split_vertices_by_face(vertices, torch.tensor(faces, dtype=torch.int32, device=dev))
    vertices = vertices.clone().requires_grad_(True)
    pixels = render(vertices, faces, width, height)
    # a loss the reference's README uses to illustrate gradients: push the cube to the right
    xs = torch.linspace(0, 1, width, device=dev)[None, :, None]
    loss = (pixels * xs).mean()
    loss.backward()
    if out:
        from _png import save_png
        save_png(out, pixels.detach().cpu().numpy())
    return pixels.detach(), vertices.grad


if __name__ == "__main__":
    px, grad = main(sys.argv[1] if len(sys.argv) > 1 else "simple.png")
    print("rendered %dx%d, %.1f%% covered; |d loss / d vertices| max %.3e" % (
        px.shape[1], px.shape[0], 100 * float((px.sum(-1) > 0).float().mean()), float(grad.abs().max())))
