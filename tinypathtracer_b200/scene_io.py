"""Flattened-scene files (.npz): the DeviceScene arrays of include/mesh.cuh:80-96 plus the camera, as written by
tests/golden/make_scene_fixtures.py.  Pure I/O — no geometry or shading arithmetic happens here."""
from __future__ import annotations

import numpy as np

from .api import LIGHT_DTYPE, MATERIAL_DTYPE, MTL_INTERVAL_DTYPE


def load_scene_npz(path) -> dict:
    z = np.load(path, allow_pickle=False)
    d = {k: z[k] for k in z.files}
    d["materials"] = np.ascontiguousarray(d["materials"]).view(MATERIAL_DTYPE).reshape(-1)
    d["objects"] = np.ascontiguousarray(d["objects"]).view(MTL_INTERVAL_DTYPE).reshape(-1)
    d["lights"] = np.ascontiguousarray(d["lights"]).view(LIGHT_DTYPE).reshape(-1)
    for k in ("vfov", "aspect", "near"):
        d[k] = np.float32(d[k])
    return d


def scene_upload_bytes(scene: dict) -> int:
    """Bytes tpt_scene_create copies host->device for this scene."""
    keys = ("indices", "vertices", "normals", "materials", "objects", "vert_trans", "normal_trans", "lights")
    return int(sum(np.asarray(scene[k]).nbytes for k in keys))
