import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


@pytest.fixture(scope="session")
def sponza_flat():
    """Flattened Sponza stand-in (shared: building it takes a couple of seconds)."""
    from vkrt_b200 import scenes
    from vkrt_b200.raytracer import flatten_model

    model = scenes.sponza_standin()
    flat = flatten_model(model)
    flat["provenance"] = model["provenance"]
    return flat
