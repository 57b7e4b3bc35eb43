"""The fork's horizon fit (tests/optimize_horizon.py:251-268 builds the op, :405-432 is the search):
Powell's method over the three camera angles and the light's z, scoring each candidate by rendering
`oceanic_horizon` and comparing channel 0 (1 = sky, 0 = sea) with a segmentation and channel 1 (the
sun's highlight) with a reflection map.  The fork reads both targets from drone footage that is not
in its repository; here they are rendered from a camera the search does not know, so the sample
also says how close the recovered camera comes.

    python samples/fit_horizon.py [out.png]
"""
import os
import sys

import numpy as np
import scipy.optimize
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dirt_b200 as dirt                      # noqa: E402

canvas_width, canvas_height = 240, 160


def get_dirt_pixels(width=canvas_width, height=canvas_height):
    """tests/optimize_horizon.py:251-268: the full-screen quad over a zero background; returns the
    function of camera_pos the fork gets from session.run on its placeholder."""
    dev = torch.device("cuda")
    square_vertices = torch.tensor([[-1, -1, 0, 1], [-1, 1, 0, 1], [1, 1, 0, 1], [1, -1, 0, 1]], dtype=torch.float32, device=dev)
    faces = torch.tensor([[0, 1, 2], [0, 2, 3]], dtype=torch.int32, device=dev)
    background = torch.zeros(height, width, 3, device=dev)
    colours = torch.ones(4, 3, device=dev)

    def run(camera_pos):
        return dirt.oceanic_horizon(vertices=square_vertices, faces=faces, vertex_colors=colours, background=background,
                                    camera_pos=[float(v) for v in camera_pos], height=height, width=width, channels=3).cpu().numpy()
    return run


def loss_terms(dirt_pixels, seg, refl):
    """tests/optimize_horizon.py:409-420 (loss3 is 0 there)."""
    diff_seg = np.clip(dirt_pixels[:, :, 0], 0.0, 1.0) - seg
    diff_refl = (np.maximum(dirt_pixels[:, :, 1] * 4.0 - 3.0, -1.0) - refl) * (4.5 + 3.5 * np.sign(refl))
    return np.mean(diff_seg ** 2.0), np.mean(diff_refl ** 2.0) / 16


def fit(seg, refl, render, x_init=(0.0, 0.0, 0.0, 1.9), maxiter=None):
    """tests/optimize_horizon.py:396-432: camera at height 200 over the origin, time 0; x = ang1,
    ang2, ang3 and the light's z."""
    feed_dict_arr = np.zeros(8)
    feed_dict_arr[1] = 200.0
    feed_dict_arr[7] = 0.9
    evaluations = [0]

    def opt_func(x):
        feed_dict_arr[3:6] = x[:3]
        feed_dict_arr[7] = x[3]
        evaluations[0] += 1
        loss1, loss2 = loss_terms(render(feed_dict_arr), seg, refl)
        return loss1 + loss2

    options = {"disp": False, "xtol": 1e-3, "ftol": 1e-5}
    if maxiter:
        options["maxiter"] = maxiter
    res = scipy.optimize.minimize(opt_func, np.asarray(x_init, dtype=np.float64), method="Powell", options=options)
    feed_dict_arr[3:6] = res.x[:3]
    feed_dict_arr[7] = res.x[3]
    return res, feed_dict_arr.copy(), evaluations[0]


def main(out=None, truth=(0.12, 0.0, 0.07, 1.2), width=canvas_width, height=canvas_height):
    render = get_dirt_pixels(width, height)
    target_cam = np.zeros(8)
    target_cam[1] = 200.0
    target_cam[3:6] = truth[:3]
    target_cam[7] = truth[3]
    target = render(target_cam)
    seg = np.clip(target[:, :, 0], 0.0, 1.0)
    refl = np.maximum(target[:, :, 1] * 4.0 - 3.0, -1.0)

    start = render(np.array([0, 200.0, 0, 0, 0, 0, 0, 1.9]))
    res, cam, evaluations = fit(seg, refl, render)
    fitted = render(cam)
    report = {
        "loss_start": float(sum(loss_terms(start, seg, refl))),
        "loss_end": float(res.fun),
        "evaluations": evaluations,
        "x": res.x.tolist(),
        "seg_mismatch_start": float(np.mean((start[:, :, 0] > 0.5) != (seg > 0.5))),
        "seg_mismatch_end": float(np.mean((fitted[:, :, 0] > 0.5) != (seg > 0.5))),
    }
    if out:
        from _png import save_png
        comp = np.zeros_like(fitted)                      # the fork's *_seg.png: target green, fit blue
        comp[:, :, 2] = fitted[:, :, 0]
        comp[:, :, 1] = seg
        save_png(out, np.concatenate([np.clip(target, 0, 1), np.clip(start, 0, 1), np.clip(comp, 0, 1)], axis=1))
    return report


if __name__ == "__main__":
    r = main(sys.argv[1] if len(sys.argv) > 1 else "fit_horizon.png")
    for k, v in r.items():
        print("%-20s %s" % (k, v))
