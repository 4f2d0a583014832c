import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def small_scene():
    from gaussian_splat_b200 import scene
    inst, _ = scene.make_instances(20000, 0x5EED0001)
    return inst


@pytest.fixture(scope="session")
def gpu_ctx():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from gaussian_splat_b200 import capi
    ctx = capi.Context(0)
    yield ctx
    ctx.close()
