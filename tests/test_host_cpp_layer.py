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
    lib.tpt_host_load_scene.restype = C.c_void_p
    lib.tpt_host_load_scene.argtypes = [C.c_char_p, C.c_char_p]
    lib.tpt_host_last_error.restype = C.c_char_p
    lib.tpt_host_scene_free.argtypes = [C.c_void_p]
    lib.tpt_host_scene_sizes.argtypes = [C.c_void_p, C.c_void_p]
    lib.tpt_host_scene_copy.argtypes = [C.c_void_p] + [C.c_void_p] * 10
    lib.tpt_host_write_png.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
    lib.tpt_host_read_image.argtypes = [C.c_char_p, C.c_void_p, C.c_size_t, C.c_void_p]
    return lib


def _load(host, path, tpt_dtypes, kind=None):
    MATERIAL_DTYPE, MTL_INTERVAL_DTYPE, LIGHT_DTYPE = tpt_dtypes
    h = host.tpt_host_load_gltf(str(path).encode()) if kind is None else host.tpt_host_load_scene(str(path).encode(), kind.encode())
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


# ---- "gltf-full": SURVEY §8(f)-1 loader hardening -------------------------------------------------------------------
def _trs(t=(0, 0, 0), q=(0, 0, 0, 1), s=(1, 1, 1)):
    x, y, z, w = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    M = np.eye(4)
    M[:3, :3] = R @ np.diag(s)
    M[:3, 3] = t
    return M


def _hierarchy_gltf(tmp_path, glb):
    """Two meshes: mesh 0 has two primitives (an indexed quad with interleaved, strided POSITION/NORMAL and u16 indices; a
    non-indexed triangle without normals) plus a LINES primitive that must be skipped; mesh 0 is instanced by two nodes
    under a rotated + scaled parent, one of them through a `matrix`; a camera hangs under a translated node; a light
    under the parent."""
    import base64
    import json
    import struct
    quad_p = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]], np.float32)
    quad_n = np.array([[0, 0, 1]] * 4, np.float32)
    inter = np.concatenate([quad_p, quad_n], axis=1).astype(np.float32).tobytes()     # stride 24
    idx = np.array([0, 1, 2, 0, 2, 3], np.uint16).tobytes()
    tri = np.array([[0, 0, 1], [2, 0, 1], [0, 2, 1]], np.float32).tobytes()
    blob = inter + idx + tri
    views = [dict(buffer=0, byteOffset=0, byteLength=len(inter), byteStride=24),
             dict(buffer=0, byteOffset=len(inter), byteLength=len(idx)),
             dict(buffer=0, byteOffset=len(inter) + len(idx), byteLength=len(tri))]
    acc = [dict(bufferView=0, byteOffset=0, componentType=5126, count=4, type="VEC3"),
           dict(bufferView=0, byteOffset=12, componentType=5126, count=4, type="VEC3"),
           dict(bufferView=1, componentType=5123, count=6, type="SCALAR"),
           dict(bufferView=2, componentType=5126, count=3, type="VEC3")]
    parent_q = (0.0, float(np.sin(np.pi / 8)), 0.0, float(np.cos(np.pi / 8)))
    child_m = _trs(t=(0.5, -1, 2), q=(float(np.sin(0.3)), 0, 0, float(np.cos(0.3))), s=(1, 2, 1))
    doc = dict(
        asset=dict(version="2.0"), scene=0, scenes=[dict(nodes=[0, 4])],
        nodes=[dict(name="parent", translation=[1, 2, 3], rotation=list(parent_q), scale=[2, 2, 2], children=[1, 2, 3]),
               dict(name="inst_a", mesh=0, translation=[0, 0, -1]),
               dict(name="inst_b", mesh=0, matrix=[float(v) for v in child_m.T.reshape(-1)]),
               dict(name="lamp", translation=[0, 3, 0], extensions=dict(KHR_lights_punctual=dict(light=0))),
               dict(name="rig", translation=[0, 1, 8], children=[5]),
               dict(name="cam", camera=0, rotation=[0, 0, 0, 1]),
               dict(name="orphan_not_in_scene", mesh=0)],
        meshes=[dict(primitives=[dict(attributes=dict(POSITION=0, NORMAL=1), indices=2, material=1),
                                 dict(attributes=dict(POSITION=3), material=0),
                                 dict(attributes=dict(POSITION=3), mode=1)])],
        materials=[dict(name="same", pbrMetallicRoughness=dict(baseColorFactor=[0.1, 0.2, 0.3, 1], metallicFactor=0)),
                   dict(name="same", pbrMetallicRoughness=dict(baseColorFactor=[0.9, 0.8, 0.7, 1], metallicFactor=0))],
        cameras=[dict(type="perspective", perspective=dict(yfov=0.8, aspectRatio=1.5, znear=0.05))],
        extensions=dict(KHR_lights_punctual=dict(lights=[dict(type="point", intensity=683.0, color=[1, 0.5, 0.25])])),
        accessors=acc, bufferViews=views)
    if glb:
        doc["buffers"] = [dict(byteLength=len(blob))]
        js = json.dumps(doc).encode()
        js += b" " * (-len(js) % 4)
        bn = blob + b"\0" * (-len(blob) % 4)
        body = struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(bn), 0x004E4942) + bn
        path = tmp_path / "h.glb"
        path.write_bytes(b"glTF" + struct.pack("<II", 2, 12 + len(body)) + body)
    else:
        doc["buffers"] = [dict(byteLength=len(blob), uri="data:application/octet-stream;base64," + base64.b64encode(blob).decode())]
        path = tmp_path / "h.gltf"
        path.write_text(json.dumps(doc))
    parent = _trs((1, 2, 3), parent_q, (2, 2, 2))
    worlds = [parent @ _trs(t=(0, 0, -1)), parent @ child_m]
    return path, quad_p, np.frombuffer(tri, np.float32).reshape(3, 3), worlds, parent


@pytest.mark.parametrize("glb", [False, True])
def test_full_loader_walks_the_scene_graph(glb, host, tmp_path):
    from oracle.pyoracle import LIGHT_DTYPE, MATERIAL_DTYPE, MTL_INTERVAL_DTYPE
    path, quad_p, tri_p, worlds, parent = _hierarchy_gltf(tmp_path, glb)
    got = _load(host, path, (MATERIAL_DTYPE, MTL_INTERVAL_DTYPE, LIGHT_DTYPE), "gltf-full")
    # 2 instances x (quad + triangle); the LINES primitive and the node outside the scene are not loaded
    assert got["objects"].shape[0] == 4 and got["indices"].size == 2 * (6 + 3) and got["vertices"].shape[0] == 2 * (4 + 3)
    np.testing.assert_array_equal(got["indices"], [0, 1, 2, 0, 2, 3, 4, 5, 6, 7, 8, 9, 7, 9, 10, 11, 12, 13])
    np.testing.assert_array_equal(got["vertices"], np.concatenate([quad_p, tri_p, quad_p, tri_p]))
    np.testing.assert_array_equal(got["normals"][:4], [[0, 0, 1]] * 4)
    np.testing.assert_allclose(got["normals"][4:7], [[0, 0, 1]] * 3, atol=1e-7)          # computed: the triangle's face normal
    for k, w in enumerate([worlds[0], worlds[0], worlds[1], worlds[1]]):
        np.testing.assert_allclose(got["vert_trans"][k].reshape(4, 4).T, w, rtol=1e-5, atol=1e-6)
        nt = got["normal_trans"][k].reshape(4, 4).T[:3, :3]
        np.testing.assert_allclose(nt, np.linalg.inv(w[:3, :3]).T, rtol=1e-4, atol=1e-6)
    # materials keyed by index (two materials share a name): quad -> material 1, triangle -> material 0
    assert got["materials"].shape[0] == 2
    np.testing.assert_allclose(got["materials"]["base"], [[0.1, 0.2, 0.3], [0.9, 0.8, 0.7]], rtol=1e-6)
    np.testing.assert_array_equal(got["objects"]["mtl"], [1, 0, 1, 0])
    np.testing.assert_array_equal(got["objects"]["begin"], [0, 2, 3, 5])
    # camera under "rig", light under "parent"
    cam = got["cam"]
    assert cam[0] == np.float32(0.8) and cam[1] == np.float32(1.5) and cam[2] == np.float32(0.05)
    np.testing.assert_allclose(cam[3:].reshape(4, 4).T, _trs(t=(0, 1, 8)), atol=1e-6)
    assert got["lights"].shape[0] == 1 and got["lights"]["type"][0] == 0
    np.testing.assert_allclose(got["lights"]["a"][0], (parent @ np.array([0, 3, 0, 1.0]))[:3], rtol=1e-5)
    np.testing.assert_allclose(got["lights"]["intensity"][0], 1.0, rtol=1e-6)


def test_reference_loader_reads_the_same_file_its_own_way(host, tmp_path):
    """type "gltf" keeps the reference's reading of the same file: every node of the flat list (the orphan too), local TRS
    only (no parents, no `matrix`), primitives[0] only, materials keyed by name (first definition wins)."""
    from oracle.pyoracle import LIGHT_DTYPE, MATERIAL_DTYPE, MTL_INTERVAL_DTYPE
    path, quad_p, _, _, _ = _hierarchy_gltf(tmp_path, False)
    got = _load(host, path, (MATERIAL_DTYPE, MTL_INTERVAL_DTYPE, LIGHT_DTYPE), "gltf")
    assert got["objects"].shape[0] == 3 and got["indices"].size == 18 and got["materials"].shape[0] == 1
    np.testing.assert_allclose(got["vert_trans"][0].reshape(4, 4).T, _trs(t=(0, 0, -1)), atol=1e-7)
    np.testing.assert_allclose(got["vert_trans"][1].reshape(4, 4).T, np.eye(4), atol=1e-7)
    np.testing.assert_allclose(got["materials"]["base"][0], [0.9, 0.8, 0.7], rtol=1e-6)
    # ... and it misreads the strided accessor as tightly packed, as the reference does (src/mesh.cu:165-170)
    assert not np.array_equal(got["vertices"][:4], quad_p)


def test_glb_container_in_reference_mode(host, tmp_path, scenes):
    """The reference's scenes repacked as .glb load to the same arrays as their .gltf."""
    import json
    import struct
    from oracle.pyoracle import LIGHT_DTYPE, MATERIAL_DTYPE, MTL_INTERVAL_DTYPE
    src = next((d / "box.gltf" for d in GLTF_DIRS if (d / "box.gltf").exists()), None)
    if src is None:
        pytest.skip("no .gltf inputs on this machine")
    doc = json.loads(src.read_text())
    assert len(doc["buffers"]) == 1
    uri = doc["buffers"][0].pop("uri")
    if uri.startswith("data:"):
        import base64
        blob = base64.b64decode(uri.split(",", 1)[1])
    else:
        blob = (src.parent / uri).read_bytes()
    js = json.dumps(doc).encode()
    js += b" " * (-len(js) % 4)
    bn = blob + b"\0" * (-len(blob) % 4)
    body = struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(bn), 0x004E4942) + bn
    path = tmp_path / "box.glb"
    path.write_bytes(b"glTF" + struct.pack("<II", 2, 12 + len(body)) + body)
    dt = (MATERIAL_DTYPE, MTL_INTERVAL_DTYPE, LIGHT_DTYPE)
    a, b = _load(host, path, dt), _load(host, src, dt)
    for k in a:
        assert a[k].tobytes() == b[k].tobytes(), k


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


def _read_image(host, path):
    whc = np.zeros(3, np.int32)
    assert host.tpt_host_read_image(str(path).encode(), None, 0, whc.ctypes.data) == 0, host.tpt_host_last_error().decode()
    out = np.zeros((whc[1], whc[0], whc[2]), np.uint8)
    assert host.tpt_host_read_image(str(path).encode(), out.ctypes.data, out.nbytes, whc.ctypes.data) == 0
    return out


def test_jpeg_decoder_agrees_with_an_independent_one(host, tmp_path):
    """The reference's environment maps are .jpg (src/main.cu:11, README): baseline JPEG decodes to what libjpeg (PIL)
    gives, up to IDCT rounding (4:4:4, 4:2:2, 4:2:0 with libjpeg's triangle-filter chroma upsampling, grey)."""
    from PIL import Image
    rng = np.random.default_rng(5)
    h, w = 75, 133                                   # not multiples of the MCU size
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.stack([128 + 100 * np.sin(xx / 17.0), 128 + 100 * np.cos(yy / 11.0), (xx * 2 + yy * 3) % 256], axis=-1)
    img = np.clip(img + rng.normal(0, 6, img.shape), 0, 255).astype(np.uint8)
    cases = [("444.jpg", dict(quality=92, subsampling=0), 4, 0.6), ("420.jpg", dict(quality=90, subsampling=2), 4, 0.6),
             ("422.jpg", dict(quality=85, subsampling=1), 4, 0.6), ("opt.jpg", dict(quality=75, subsampling=0, optimize=True), 4, 0.6)]
    for name, kw, max_tol, mean_tol in cases:
        Image.fromarray(img).save(tmp_path / name, **kw)
        ref = np.asarray(Image.open(tmp_path / name).convert("RGB")).astype(int)
        got = _read_image(host, tmp_path / name).astype(int)
        assert got.shape == ref.shape == (h, w, 3)
        d = np.abs(got - ref)
        assert d.max() <= max_tol and d.mean() <= mean_tol, (name, d.max(), d.mean())
    Image.fromarray(img[..., 0]).save(tmp_path / "grey.jpg", quality=90)
    got = _read_image(host, tmp_path / "grey.jpg").astype(int)
    ref = np.asarray(Image.open(tmp_path / "grey.jpg")).astype(int)
    assert got.shape == (h, w, 3) and np.abs(got[..., 0] - ref).max() <= 4 and (got[..., 0] == got[..., 2]).all()
    Image.fromarray(img).save(tmp_path / "prog.jpg", progressive=True)
    whc = np.zeros(3, np.int32)
    assert host.tpt_host_read_image(str(tmp_path / "prog.jpg").encode(), None, 0, whc.ctypes.data) != 0
    assert b"progressive JPEG is not supported" in host.tpt_host_last_error()
    (tmp_path / "trunc.jpg").write_bytes((tmp_path / "444.jpg").read_bytes()[:300])
    assert host.tpt_host_read_image(str(tmp_path / "trunc.jpg").encode(), None, 0, whc.ctypes.data) != 0


def test_jpeg_decoder_refuses_malformed_scan_headers(host, tmp_path):
    """A scan header that names a Huffman table above 3, or a table no DHT / DQT segment defined, or that is shorter than
    its component list, is refused with a message instead of indexing past the decoder's 4-entry table arrays."""
    from PIL import Image
    img = (np.arange(32 * 32 * 3) % 251).astype(np.uint8).reshape(32, 32, 3)
    Image.fromarray(img).save(tmp_path / "ok.jpg", quality=90, subsampling=0)
    data = bytearray((tmp_path / "ok.jpg").read_bytes())
    whc = np.zeros(3, np.int32)

    def refused(blob, needle):
        (tmp_path / "bad.jpg").write_bytes(bytes(blob))
        assert host.tpt_host_read_image(str(tmp_path / "bad.jpg").encode(), None, 0, whc.ctypes.data) != 0
        assert needle in host.tpt_host_last_error(), host.tpt_host_last_error()

    sos = data.index(b"\xff\xda")
    ns = data[sos + 4]
    assert ns == 3
    bad = bytearray(data)
    bad[sos + 6] = 0x7f                      # component 0: Td = 7, Ta = 15
    refused(bad, b"Huffman table above 3")
    bad = bytearray(data)
    bad[sos + 6] = 0x22                      # tables 2/2: in range but never defined by a DHT segment
    refused(bad, b"no DHT segment defined")
    bad = bytearray(data)
    bad[sos + 2:sos + 4] = (4).to_bytes(2, "big")   # segment length too short for 3 components
    refused(bad, b"truncated JPEG scan header")
    # a frame that names quantisation table 3 (never defined)
    sof = data.index(b"\xff\xc0")
    bad = bytearray(data)
    bad[sof + 4 + 8] = 3
    refused(bad, b"no DQT segment defined")


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


@pytest.mark.gpu
def test_progressive_frames_and_checkpoint_resume_equal_one_shot(tmp_path):
    """frames that ADD samples (same seed, continuing sample indices) and a render resumed from a
    checkpoint give the image of one uninterrupted frame with all the samples — the same bytes, since accumulation is exact."""
    import subprocess
    from PIL import Image
    exe = ROOT / "tinypathtracer_b200" / "lib" / "tpt_render"
    src = next((d / "box.gltf" for d in GLTF_DIRS if (d / "box.gltf").exists()), None)
    if src is None:
        pytest.skip("no .gltf inputs on this machine")
    base = [str(exe), "--scene", str(src), "--w", "128", "--h", "72", "--depth", "8", "--seed", "5"]

    def run(*extra):
        r = subprocess.run(base + list(extra), capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        return r

    run("--spp", "32", "--out", str(tmp_path / "one.png"))
    run("--spp", "8", "--frames", "4", "--progressive", "--out", str(tmp_path / "prog.png"))
    run("--spp", "8", "--frames", "2", "--progressive", "--checkpoint", str(tmp_path / "half.ckpt"), "--out", str(tmp_path / "half.png"))
    run("--spp", "16", "--resume", str(tmp_path / "half.ckpt"), "--seed", "777", "--out", str(tmp_path / "resumed.png"))   # seed comes from the checkpoint
    one = np.asarray(Image.open(tmp_path / "one.png"))
    assert one.mean() > 5
    np.testing.assert_array_equal(np.asarray(Image.open(tmp_path / "prog.png")), one)
    np.testing.assert_array_equal(np.asarray(Image.open(tmp_path / "resumed.png")), one)
    assert not np.array_equal(np.asarray(Image.open(tmp_path / "half.png")), one)
    # without --progressive every frame restarts with a new seed, as the reference does
    run("--spp", "8", "--frames", "4", "--out", str(tmp_path / "restart.png"))
    assert not np.array_equal(np.asarray(Image.open(tmp_path / "restart.png")), one)
    bad = subprocess.run(base + ["--w", "64", "--spp", "8", "--resume", str(tmp_path / "half.ckpt")], capture_output=True, text=True)
    assert bad.returncode == 1 and "checkpoint is 128x72" in bad.stderr


# ---- float images: Radiance .hdr and PFM --------------------------------------------------------------
def _read_f32(host, path):
    host.tpt_host_read_image_f32.argtypes = [C.c_char_p, C.c_void_p, C.c_size_t, C.c_void_p]
    wh = np.zeros(2, np.int32)
    assert host.tpt_host_read_image_f32(str(path).encode(), None, 0, wh.ctypes.data) == 0, host.tpt_host_last_error().decode()
    out = np.zeros((wh[1], wh[0], 3), np.float32)
    assert host.tpt_host_read_image_f32(str(path).encode(), out.ctypes.data, out.size, wh.ctypes.data) == 0
    return out


def _rgbe_to_float(rgbe):
    """rgbe.c convention (the one FreeImage's HDR plugin uses): mantissa * 2^(e - 136), e = 0 -> 0."""
    e = rgbe[..., 3].astype(np.int32)
    f = np.where(e > 0, np.ldexp(np.float32(1.0), e - 136), np.float32(0.0)).astype(np.float32)
    return rgbe[..., :3].astype(np.float32) * f[..., None]


def _rle_channel(vals):
    """Radiance new-style run-length coding of one channel of one scanline."""
    out, i, n = bytearray(), 0, len(vals)
    while i < n:
        run = 1
        while i + run < n and run < 127 and vals[i + run] == vals[i]:
            run += 1
        if run >= 4:
            out += bytes([128 + run, vals[i]])
            i += run
        else:
            j = i
            while j < n and j - i < 128:
                r = 1
                while j + r < n and r < 4 and vals[j + r] == vals[j]:
                    r += 1
                if r >= 4:
                    break
                j += 1
            out += bytes([j - i]) + bytes(vals[i:j])
            i = j
    return bytes(out)


def test_radiance_hdr_reader_flat_and_rle(host, tmp_path):
    rng = np.random.default_rng(5)
    W, H = 67, 19
    rgbe = rng.integers(0, 256, (H, W, 4), dtype=np.uint8)
    rgbe[..., 3] = rng.integers(100, 150, (H, W))          # exponents around 1.0
    rgbe[2, 5:40] = (200, 100, 50, 140)                    # long runs, so the RLE path sees both kinds of packet
    rgbe[7, :, 3] = 0                                      # e = 0 -> black
    rgbe[9, 10:30, 0] = 77
    want = _rgbe_to_float(rgbe)
    hdr = b"#?RADIANCE\n# made by the test\nFORMAT=32-bit_rle_rgbe\nEXPOSURE=1.0\n\n-Y %d +X %d\n" % (H, W)
    flat = tmp_path / "flat.hdr"
    flat.write_bytes(hdr + rgbe.tobytes())
    np.testing.assert_array_equal(_read_f32(host, flat), want)
    body = bytearray()
    for y in range(H):
        body += bytes([2, 2, W >> 8, W & 255])
        for c in range(4):
            body += _rle_channel(rgbe[y, :, c].tolist())
    rle = tmp_path / "rle.hdr"
    rle.write_bytes(hdr + bytes(body))
    assert len(body) < rgbe.nbytes + 4 * H + H * W // 16    # it really is run-length coded
    np.testing.assert_array_equal(_read_f32(host, rle), want)
    # errors are exceptions (-> -1 + message through the C hook)
    (tmp_path / "cut.hdr").write_bytes(hdr + bytes(body[:len(body) // 2]))
    wh = np.zeros(2, np.int32)
    assert host.tpt_host_read_image_f32(str(tmp_path / "cut.hdr").encode(), None, 0, wh.ctypes.data) == -1
    assert b"truncated" in host.tpt_host_last_error()
    (tmp_path / "fmt.hdr").write_bytes(hdr.replace(b"rle_rgbe", b"rle_xyze") + rgbe.tobytes())
    assert host.tpt_host_read_image_f32(str(tmp_path / "fmt.hdr").encode(), None, 0, wh.ctypes.data) == -1


def test_pfm_round_trip_and_foreign_files(host, tmp_path):
    host.tpt_host_write_pfm.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int]
    rng = np.random.default_rng(6)
    img = (rng.random((23, 41, 3), dtype=np.float32) * 1000.0).astype(np.float32)
    p = tmp_path / "a.pfm"
    assert host.tpt_host_write_pfm(str(p).encode(), img.ctypes.data, 41, 23) == 0
    raw = p.read_bytes()
    assert raw.startswith(b"PF\n41 23\n-1.0\n")
    body = np.frombuffer(raw[len(b"PF\n41 23\n-1.0\n"):], "<f4").reshape(23, 41, 3)
    np.testing.assert_array_equal(body[::-1], img)          # PFM rows run bottom to top
    np.testing.assert_array_equal(_read_f32(host, p), img)
    # big-endian colour and little-endian grey files from another writer
    (tmp_path / "be.pfm").write_bytes(b"PF\n41 23\n1.0\n" + img[::-1].astype(">f4").tobytes())
    np.testing.assert_array_equal(_read_f32(host, tmp_path / "be.pfm"), img)
    grey = img[..., 0]
    (tmp_path / "g.pfm").write_bytes(b"Pf\n41 23\n-1.0\n" + grey[::-1].astype("<f4").tobytes())
    np.testing.assert_array_equal(_read_f32(host, tmp_path / "g.pfm"), np.repeat(grey[..., None], 3, axis=2))


@pytest.mark.gpu
def test_cli_hdr_environment_and_float_output(tpt, ctx, scenes, tmp_path):
    """tpt_render --env sky.hdr --out-float frame.pfm: the Radiance map reaches the device as float texels (same accumulator as
    the binding given the decoded floats) and the PFM holds the unclamped linear mean radiance."""
    import subprocess
    exe = ROOT / "tinypathtracer_b200" / "lib" / "tpt_render"
    src = next((d / "ball.gltf" for d in GLTF_DIRS if (d / "ball.gltf").exists()), None)
    if src is None:
        pytest.skip("no .gltf inputs on this machine")
    rng = np.random.default_rng(8)
    EW, EH = 64, 32
    rgbe = np.zeros((EH, EW, 4), np.uint8)
    rgbe[..., :3] = rng.integers(64, 256, (EH, EW, 3))
    rgbe[..., 3] = 127                                       # mantissa / 512: a dim sky
    rgbe[4:8, 20:24] = (255, 240, 200, 138)                  # sun, top-down rows: ~1000
    env_file = tmp_path / "sky.hdr"
    env_file.write_bytes(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n" % (EH, EW) + rgbe.tobytes())
    env = _rgbe_to_float(rgbe)[::-1].copy()                  # the device map is bottom-up
    W, H, spp = 128, 72, 16
    out = tmp_path / "frame.pfm"
    r = subprocess.run([str(exe), "--scene", str(src), "--env", str(env_file), "--w", str(W), "--h", str(H), "--spp", str(spp), "--seed", "4",
                        "--out-float", str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    raw = out.read_bytes()
    head = b"PF\n%d %d\n-1.0\n" % (W, H)
    assert raw.startswith(head)
    got = np.frombuffer(raw[len(head):], "<f4").reshape(H, W, 3)       # bottom-up, like the accumulator
    ds = tpt.DeviceScene(ctx, scenes["ball"])
    tpt.BVH(ds).construct()
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=env)
    pt.render_into(ds, tpt.Camera.from_scene(scenes["ball"]), spp=spp, max_depth=8, seed=4)
    want = pt.m_radiance.to_f32() * np.float32(1.0 / spp)
    ds.close()
    np.testing.assert_array_equal(got, want)
    assert got.max() > 1.5                                   # above anything an 8-bit map or framebuffer can hold


@pytest.mark.gpu
def test_cli_single_process_multi_gpu_is_bit_identical(tmp_path):
    """tpt_render --devices a,b,c: one host thread drives one context per entry (here three contexts on GPU 0; on a multi-GPU box
    `--gpus N`), rank g renders the samples = g (mod n), tpt_accum_allreduce sums the buffers with its peer-memory kernel.
    The PNG and the float dump must equal the single-context run's byte for byte."""
    import subprocess
    exe = ROOT / "tinypathtracer_b200" / "lib" / "tpt_render"
    src = next((d / "box2.gltf" for d in GLTF_DIRS if (d / "box2.gltf").exists()), None)
    if src is None:
        pytest.skip("no .gltf inputs on this machine")
    from tinypathtracer_b200.sky import synthetic_sky
    from PIL import Image
    sky = tmp_path / "sky.png"
    Image.fromarray(synthetic_sky(64, 32)[::-1, :, :3].copy()).save(sky)
    outs = []
    for tag, extra in (("one", []), ("three", ["--devices", "0,0,0"])):
        png, pfm = tmp_path / f"{tag}.png", tmp_path / f"{tag}.pfm"
        r = subprocess.run([str(exe), "--scene", str(src), "--env", str(sky), "--w", "96", "--h", "54", "--spp", "24", "--seed", "3", "--frames", "2",
                            "--out", str(png), "--out-float", str(pfm), *extra], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append((png.read_bytes(), pfm.read_bytes()))
    assert outs[0][0] == outs[1][0] and outs[0][1] == outs[1][1]
    bad = subprocess.run([str(exe), "--scene", str(src), "--spp", "25", "--devices", "0,0,0"], capture_output=True, text=True, timeout=300)
    assert bad.returncode == 1 and "multiple of the number of GPUs" in bad.stderr
