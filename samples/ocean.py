"""The fork's own way of calling the op (tests/square_test.py:39-72, tests/optimize_horizon.py:251-268):
a full-screen quad whose fragments are shaded by a procedural shader driven by `camera_pos`.
Renders every scene op dirt_b200 carries into one contact sheet.

    python samples/ocean.py [out.png]
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dirt_b200 as dirt                      # noqa: E402

canvas_width, canvas_height = 480, 270


def terrain_lookup(height, width):
    """A stand-in for the fork's terrain_lookup.npy (not in its repository): height in [0,1] and
    the normal * 0.5 + 0.5; columns span world z in [5,25], rows (bottom up) world x in [-14,14]."""
    t = np.linspace(1, 0, height)[:, None]
    s = np.linspace(0, 1, width)[None, :]
    h = 0.55 + 0.08 * np.sin(6 * s + 1) * np.cos(5 * t) + 0.05 * np.sin(13 * s * t)
    hw = h * 10.3 - 6.1
    dz = np.gradient(hw, axis=1) / (20.0 / (width - 1))
    dx = np.gradient(hw, axis=0) / (-28.0 / (height - 1))
    n = np.stack([-dx, np.ones_like(hw), -dz], -1)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    return np.concatenate([h[..., None], (n + 1) / 2], -1).astype(np.float32)


def main(out=None, width=canvas_width, height=canvas_height):
    dev = torch.device("cuda")
    square = torch.tensor([[-1, -1, 0, 1], [-1, 1, 0, 1], [1, 1, 0, 1], [1, -1, 0, 1]], dtype=torch.float32, device=dev)
    faces = torch.tensor([[0, 1, 2], [0, 2, 3]], dtype=torch.int32, device=dev)
    colours = torch.ones(4, 3, device=dev)
    noise = torch.randn(height, width, 3, generator=torch.Generator().manual_seed(0)).to(dev) * 0.3
    camera = [0.0, 150.0, 0.0, 0.0, 0.3, 0.0, 1.0, 1.5, 0.5, 1 / 30.0, 30.0, 14.0, 15.0, 0.1, 0.1, 0.1]
    args = dict(vertices=square, faces=faces, vertex_colors=colours, height=height, width=width, channels=3)
    images = {
        "oceanic_horizon": dirt.oceanic_horizon(background=noise, camera_pos=camera[:8], **args),
        "oceanic": dirt.oceanic(background=noise, camera_pos=camera[:8], **args),
        "oceanic_still_cloud": dirt.oceanic_still_cloud(background=noise, camera_pos=camera[:9], **args),
        "oceanic_no_cloud": dirt.oceanic_no_cloud(background=noise, camera_pos=camera[:8], **args),
        "oceanic_simple_proxy": dirt.oceanic_simple_proxy(background=noise, camera_pos=camera[:8], **args),
        "hill": dirt.hill(background=torch.from_numpy(terrain_lookup(height, width)).to(dev),
                          camera_pos=[0.0] * 9 + [0.0, 0.0, 3.0], **args),
    }
    flow = dirt.oceanic_opt_flow(background=noise, camera_pos=camera, **args)
    images["oceanic_opt_flow"] = torch.stack([flow[..., 0] / width, flow[..., 1] / height, torch.zeros_like(flow[..., 0])], -1)
    if out:
        from _png import save_png
        names = list(images)
        rows = [torch.cat([images[n] for n in names[i:i + 4]] + [torch.zeros_like(noise)] * (4 - len(names[i:i + 4])), dim=1)
                for i in range(0, len(names), 4)]
        save_png(out, torch.cat(rows, dim=0).cpu().numpy())
    return images


if __name__ == "__main__":
    imgs = main(sys.argv[1] if len(sys.argv) > 1 else "ocean.png")
    for name, im in imgs.items():
        print("%-22s mean %s" % (name, np.round(im.mean(dim=(0, 1)).cpu().numpy(), 3)))
