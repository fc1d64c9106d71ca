"""The C++ host layer (include/tpt/host.hpp, libtpt_host.so): its glTF loader must flatten the reference's scenes to
exactly the DeviceScene arrays of the committed fixtures (which come from the loader restatement that is pinned to the
reference's own loader, tests/test_oracle_pinned_to_reference.py), and its PNG codecs must round-trip.  CPU only: the
loader and the codecs call nothing on the device."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
LIB = ROOT / "tinypathtracer_b200" / "lib" / "libtpt_host.so"
GLTF_DIRS = [Path("/root/reference/input"), ROOT / "oracle" / "_ref" / "input"]


@pytest.fixture(scope="module")
def host():
    if not LIB.exists():
        pytest.fail(f"{LIB} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
    lib = C.CDLL(str(LIB))
    lib.tpt_host_load_gltf.restype = C.c_void_p
    lib.tpt_host_load_gltf.argtypes = [C.c_char_p]
    lib.tpt_host_last_error.restype = C.c_char_p
    lib.tpt_host_scene_free.argtypes = [C.c_void_p]
    lib.tpt_host_scene_sizes.argtypes = [C.c_void_p, C.c_void_p]
    lib.tpt_host_scene_copy.argtypes = [C.c_void_p] + [C.c_void_p] * 10
    lib.tpt_host_write_png.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
    lib.tpt_host_read_image.argtypes = [C.c_char_p, C.c_void_p, C.c_size_t, C.c_void_p]
    return lib


def _load(host, path, tpt_dtypes):
    MATERIAL_DTYPE, MTL_INTERVAL_DTYPE, LIGHT_DTYPE = tpt_dtypes
    h = host.tpt_host_load_gltf(str(path).encode())
    assert h, host.tpt_host_last_error().decode()
    sizes = np.zeros(5, np.uint64)
    host.tpt_host_scene_sizes(h, sizes.ctypes.data)
    ni, nv, nm, no, nl = (int(x) for x in sizes)
    out = dict(indices=np.zeros(ni, np.uint32), vertices=np.zeros((nv, 3), np.float32), normals=np.zeros((nv, 3), np.float32),
               texcoords=np.zeros((nv, 2), np.float32), materials=np.zeros(nm, MATERIAL_DTYPE), objects=np.zeros(no, MTL_INTERVAL_DTYPE),
               vert_trans=np.zeros((no, 16), np.float32), normal_trans=np.zeros((no, 16), np.float32), lights=np.zeros(max(nl, 1), LIGHT_DTYPE),
               cam=np.zeros(19, np.float32))
    host.tpt_host_scene_copy(h, *[out[k].ctypes.data for k in ("indices", "vertices", "normals", "texcoords", "materials", "objects",
                                                              "vert_trans", "normal_trans", "lights", "cam")])
    out["lights"] = out["lights"][:nl]
    host.tpt_host_scene_free(h)
    return out


@pytest.mark.parametrize("name", ["box", "box1", "box2", "tir", "ball", "light", "square"])
def test_cpp_loader_matches_the_fixtures(name, host, scenes):
    from oracle.pyoracle import LIGHT_DTYPE, MATERIAL_DTYPE, MTL_INTERVAL_DTYPE
    src = next((d / f"{name}.gltf" for d in GLTF_DIRS if (d / f"{name}.gltf").exists()), None)
    if src is None:
        pytest.skip("no .gltf inputs on this machine (neither /root/reference/input nor oracle/_ref/input)")
    got = _load(host, src, (MATERIAL_DTYPE, MTL_INTERVAL_DTYPE, LIGHT_DTYPE))
    ref = scenes[name]
    for k in ("indices", "vertices", "normals", "vert_trans"):
        np.testing.assert_array_equal(got[k].reshape(-1), np.asarray(ref[k]).reshape(-1), err_msg=k)
    np.testing.assert_array_equal(got["texcoords"].reshape(-1), np.asarray(ref["texcoords"], np.float32).reshape(-1))
    # the reference inverts with a fully expanded 4x4 cofactor formula; the host layer uses the 3x3 cofactors: same matrix to rounding
    np.testing.assert_allclose(got["normal_trans"], np.asarray(ref["normal_trans"]).reshape(-1, 16), rtol=2e-6, atol=1e-7)
    assert got["materials"].tobytes() == np.asarray(ref["materials"]).tobytes()
    assert got["objects"].tobytes() == np.asarray(ref["objects"]).tobytes()
    assert got["lights"].tobytes() == np.asarray(ref["lights"]).tobytes()
    cam = got["cam"]
    assert cam[0] == np.float32(ref["vfov"]) and cam[1] == np.float32(ref["aspect"]) and cam[2] == np.float32(ref["near"])
    np.testing.assert_array_equal(cam[3:], np.asarray(ref["c2w"], np.float32).reshape(16))


def test_loader_errors_are_exceptions_with_the_reference_messages(host, tmp_path):
    assert not host.tpt_host_load_gltf(str(tmp_path / "missing.gltf").encode())
    assert b"Failed to load .gltf" in host.tpt_host_last_error()
    bad = tmp_path / "bad.gltf"
    bad.write_text('{"nodes": [{"extensions": {"KHR_lights_punctual": {"light": 0}}}], "extensions": {"KHR_lights_punctual": {"lights": [{"type": "area"}]}}}')
    assert not host.tpt_host_load_gltf(str(bad).encode())
    assert b"Unsupported light type" in host.tpt_host_last_error()


def test_png_codecs_round_trip(host, tmp_path):
    rng = np.random.default_rng(3)
    for ch in (3, 4):
        img = rng.integers(0, 256, (37, 53, ch), dtype=np.uint8)
        path = tmp_path / f"rt{ch}.png"
        assert host.tpt_host_write_png(str(path).encode(), img.ctypes.data, 53, 37, ch) == 0
        whc = np.zeros(3, np.int32)
        out = np.zeros_like(img)
        assert host.tpt_host_read_image(str(path).encode(), out.ctypes.data, out.nbytes, whc.ctypes.data) == 0, host.tpt_host_last_error()
        assert tuple(whc) == (53, 37, ch)
        np.testing.assert_array_equal(out, img)
        from PIL import Image                      # an independent decoder agrees, and our decoder reads an independent encoder's file
        np.testing.assert_array_equal(np.asarray(Image.open(path)), img)
        Image.fromarray(img).save(tmp_path / f"pil{ch}.png")
        out2 = np.zeros_like(img)
        assert host.tpt_host_read_image(str(tmp_path / f"pil{ch}.png").encode(), out2.ctypes.data, out2.nbytes, whc.ctypes.data) == 0
        np.testing.assert_array_equal(out2, img)


@pytest.mark.gpu
def test_tpt_render_cli_matches_the_python_binding(tpt, ctx, scenes, tmp_path):
    """src/main.cu's PathTracer(env).render(gltf) through the C++ classes: same seed -> the same framebuffer, byte for byte,
    as PathTracer.doTrace of the ctypes binding on the fixture arrays."""
    import json
    import subprocess
    from PIL import Image
    exe = ROOT / "tinypathtracer_b200" / "lib" / "tpt_render"
    src = next((d / "box.gltf" for d in GLTF_DIRS if (d / "box.gltf").exists()), None)
    if src is None:
        pytest.skip("no .gltf inputs on this machine")
    W, H, spp = 160, 90, 32
    out = tmp_path / "frame.png"
    r = subprocess.run([str(exe), "--scene", str(src), "--w", str(W), "--h", str(H), "--spp", str(spp), "--depth", "8", "--seed", "9", "--out", str(out)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    info = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert info["samples"] == W * H * spp and info["rays"] > info["samples"]
    img = np.asarray(Image.open(out))
    ds = tpt.DeviceScene(ctx, scenes["box"])
    pt = tpt.PathTracer(ctx, W, H)
    fb = np.zeros((H, W, 4), np.uint8)
    pt.doTrace(ds, tpt.Camera.from_scene(scenes["box"]), fb, spp, max_depth=8, seed=9)
    ds.close()
    np.testing.assert_array_equal(img, fb[..., 2::-1])        # PNG is RGB, the framebuffer B,G,R,(A)
    assert img.mean() > 5
    bad = subprocess.run([str(exe), "--scene", str(tmp_path / "nope.gltf")], capture_output=True, text=True)
    assert bad.returncode == 1 and "Failed to load .gltf" in bad.stderr
