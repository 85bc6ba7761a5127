import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
REF_HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden_path(*parts):
    return os.path.join(GOLDEN, *parts)


@pytest.fixture(scope="session")
def flat_scenes():
    from master_b200.scene import load_b2s
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = load_b2s(golden_path("scenes", name + ".b2s"))
        return cache[name]
    return get


@pytest.fixture(scope="session")
def device_scenes(flat_scenes):
    """Scenes uploaded to the GPU through the C ABI (b2pt_scene_create)."""
    from master_b200 import api
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = api.DeviceSceneHandle(flat_scenes(name))
        return cache[name]
    yield get
    for h in cache.values():
        h.close()
