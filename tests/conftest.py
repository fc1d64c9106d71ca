import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

# k_primary's specular continuation is limited to shallow trees by default (a speed heuristic, csrc/render.cu); the tests lift
# the limit so that the continuation is exercised on every fixture — results are bit-identical either way, which is what
# test_specular_continuation_is_bit_identical_to_queued_paths asserts against the queued form.
os.environ.setdefault("TPT_SPEC_MAX_FACES", "100000000")
# 4-wide compressed nodes are built for scenes of 65 536 triangles and more by default (csrc/capi.cu); the tests build them for
# every scene, so that each traversal test of tpt_trace_rays (brute-force parity on the seven fixtures, the soups) walks the wide
# tree.  test_wide_and_binary_walks_agree (tests/test_gpu_round2.py) runs the binary walk beside it (TPT_NO_WIDE).
os.environ.setdefault("TPT_WIDE_MIN_FACES", "2")

SCENES = ["box", "box1", "box2", "tir", "ball", "light", "square"]
GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `pytest -m gpu`)")


def load_fixture_scene(name: str) -> dict:
    from oracle import gltf_scene
    return gltf_scene.load_scene(GOLDEN / "scenes" / f"{name}.npz")


@pytest.fixture(scope="session")
def scenes():
    return {n: load_fixture_scene(n) for n in SCENES}


@pytest.fixture(scope="session")
def orc():
    from oracle import pyoracle
    pyoracle.lib()
    return pyoracle


@pytest.fixture(scope="session")
def tpt():
    """The product package.  GPU tests must run the CUDA library: no skip if it is missing."""
    import tinypathtracer_b200
    return tinypathtracer_b200


@pytest.fixture(scope="session")
def ctx(tpt):
    c = tpt.Context(0)
    yield c
    c.close()


from tinypathtracer_b200.sky import synthetic_sky  # noqa: E402,F401
