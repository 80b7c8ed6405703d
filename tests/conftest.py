import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p_ in (ROOT, os.path.join(ROOT, "tests")):
    if p_ not in sys.path:
        sys.path.insert(0, p_)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def tiny_scene(tmp_path_factory):
    from line3d_b200 import scenes
    p = str(tmp_path_factory.mktemp("scenes") / "tiny.scn")
    scenes.write_scene(scenes.tiny(), p)
    return p


@pytest.fixture(scope="session")
def small_scene(tmp_path_factory):
    """12 views x ~300 segments: large enough for multi-tile paths, seconds on the oracle."""
    from line3d_b200 import scenes
    p = str(tmp_path_factory.mktemp("scenes") / "small.scn")
    scenes.write_scene(scenes.wireframe_building(n_views=12, n_seg=300, seed=11, radius=30.0), p)
    return p
