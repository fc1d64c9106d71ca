import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

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


def synthetic_sky(width: int = 1024, height: int = 512) -> np.ndarray:
    """SURVEY §8(d)-2 synthetic sky (replaces the missing kloppenheim_07_puresky.jpg): RGBA8, row 0 = BOTTOM."""
    v = (np.arange(height, dtype=np.float64) + 0.5) / height          # v = 1 - theta/pi ; row 0 bottom
    u = (np.arange(width, dtype=np.float64) + 0.5) / width
    theta = (1.0 - v) * np.pi
    phi = u * 2.0 * np.pi
    dy = np.cos(theta)[:, None] * np.ones((1, width))
    st = np.sin(theta)[:, None]
    dx = st * np.cos(phi)[None, :]
    dz = st * np.sin(phi)[None, :]
    t = np.sqrt(np.maximum(dy, 0.0))
    horizon = np.array([0.9, 0.9, 0.95])
    zenith = np.array([0.25, 0.45, 0.9])
    rgb = horizon[None, None, :] * (1 - t[..., None]) + zenith[None, None, :] * t[..., None]
    rgb = np.where((dy < 0)[..., None], np.array([0.35, 0.33, 0.3])[None, None, :], rgb)
    el, az = np.radians(40.0), np.radians(30.0)
    sun = np.array([np.cos(el) * np.cos(az), np.sin(el), np.cos(el) * np.sin(az)])
    cosang = dx * sun[0] + dy * sun[1] + dz * sun[2]
    rgb = np.where((cosang > np.cos(np.radians(3.0)))[..., None], 1.0, rgb)
    img = np.zeros((height, width, 4), np.uint8)
    img[..., :3] = np.clip(np.floor(rgb * 255.0 + 0.5), 0, 255).astype(np.uint8)
    img[..., 3] = 255
    return img
