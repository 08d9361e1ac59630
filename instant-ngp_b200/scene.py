"""Synthetic NeRF-synthetic-shape dataset (BASELINE.json): N posed WxH RGBA8 views of the analytic density field,
cameras on an upper-hemisphere Fibonacci lattice.  Views are rendered by the library's
data-synthesis kernel and returned as HOST (pinned) buffers, like images decoded by a loader would be."""
import ctypes as C
import math

import numpy as np
import torch

from ._lib import check, lib

CAMERA_ANGLE_X = 0.6911      # Blender NeRF-synthetic field of view
RADIUS = 4.0 * 0.33          # NeRF radius 4.0 scaled by NERF_SCALE (nerf_loader.h:29)


def focal_px(width, camera_angle_x=CAMERA_ANGLE_X):
    return 0.5 * width / math.tan(0.5 * camera_angle_x)


def camera(view, n_views, radius=RADIUS):
    x = (C.c_float * 12)()
    check(lib().ngp_synth_camera(view, n_views, radius, x))
    return np.array(list(x), np.float32)


def make_views(n_views=100, width=800, height=800, radius=RADIUS, pinned=True):
    """-> list of view dicts for Testbed.set_nerf_dataset (pixels: host uint8 [H,W,4])."""
    f = focal_px(width)
    dev = torch.empty(height, width, 4, dtype=torch.uint8, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    views = []
    for i in range(n_views):
        check(lib().ngp_synth_render_view(s, i, n_views, radius, f, width, height, C.c_void_p(dev.data_ptr())))
        host = torch.empty(height, width, 4, dtype=torch.uint8, pin_memory=pinned)
        host.copy_(dev)
        views.append({"pixels": host, "focal_length": (f, f), "principal_point": (0.5, 0.5), "xform": camera(i, n_views, radius)})
    torch.cuda.synchronize()
    return views


NERF_CONFIG_16x2 = {
    "loss": {"otype": "Huber"},
    "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                  "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}},
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 1.3819},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
    "dir_encoding": {"otype": "Composite", "nested": [{"n_dims_to_encode": 3, "otype": "SphericalHarmonics", "degree": 4}, {"otype": "Identity"}]},
    "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}
