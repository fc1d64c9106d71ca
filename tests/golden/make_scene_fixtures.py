"""Regenerates tests/golden/scenes/*.npz: the seven input/*.gltf scenes of the reference flattened to the
DeviceScene arrays (include/mesh.cuh:80-96) by the oracle's restatement of the loader
(oracle/gltf_scene.py <- src/mesh.cu:80-397).  /root/reference does not exist on the GPU box, so the GPU
tests, smoke() and bench.py read these fixtures, never the .gltf files.

    python tests/golden/make_scene_fixtures.py [/root/reference/input]
"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle import gltf_scene  # noqa: E402

SCENES = ["box", "box1", "box2", "tir", "ball", "light", "square"]


def main():
    src = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference/input")
    out = Path(__file__).resolve().parent / "scenes"
    out.mkdir(exist_ok=True)
    for name in SCENES:
        scene = gltf_scene.load_gltf(src / f"{name}.gltf")
        gltf_scene.save_scene(scene, out / f"{name}.npz")
        print(name, "faces", scene["indices"].size // 3, "verts", scene["vertices"].shape[0], "objects", scene["objects"].shape[0],
              "materials", scene["materials"].shape[0], "lights", scene["lights"].shape[0])


if __name__ == "__main__":
    main()
