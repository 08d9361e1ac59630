import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_pkg():
    """The package directory is named `instant-ngp_b200` (hyphen): import it through importlib."""
    return importlib.import_module("instant-ngp_b200")


@pytest.fixture(scope="session")
def ngp():
    return load_pkg()


@pytest.fixture(scope="session")
def ops():
    load_pkg()
    return importlib.import_module("instant-ngp_b200.ops")


@pytest.fixture(scope="session")
def ref():
    from oracle import ref_api
    ref_api.lib()
    return ref_api


NERF_CFG_16x2 = {
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16,
                 "per_level_scale": 1.3819},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
    "dir_encoding": {"otype": "Composite", "nested": [{"n_dims_to_encode": 3, "otype": "SphericalHarmonics", "degree": 4}, {"otype": "Identity"}]},
    "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}
# the fork's configs/nerf/base.json network (4 levels x 4 features, per_level_scale default 2.0)
NERF_CFG_FORK = {
    "encoding": {"otype": "HashGrid", "n_levels": 4, "n_features_per_level": 4, "log2_hashmap_size": 19, "base_resolution": 16},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
    "dir_encoding": NERF_CFG_16x2["dir_encoding"],
    "rgb_network": NERF_CFG_16x2["rgb_network"],
}
