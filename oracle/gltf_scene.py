"""Oracle restatement of the reference's scene loader and upload layout.

TEST INFRASTRUCTURE ONLY (see oracle/oracle.cpp header).  Follows
  src/mesh.cu:80-307   Scene::readFromGLTF   (ASCII .gltf, primitives[0], TRS per node, no hierarchy,
                                              KHR_lights_punctual, three KHR material extensions)
  src/mesh.cu:309-397  Scene::copySceneToDevice  (flatten to the DeviceScene of include/mesh.cuh:80-96)
and returns plain numpy arrays with exactly the DeviceScene fields:

  indices      u32[3F]    rebased to the global vertex array (mesh.cu:340-346)
  vertices     f32[V,3]   object space;  normals f32[V,3];  texcoords f32[V,2] (uploaded, never read)
  materials    MATERIAL_DTYPE[M]  in std::map<std::string> (byte-lexicographic) name order (mesh.cu:326-333)
  objects      MTL_INTERVAL_DTYPE[O]  {first face of object, material index}, node order (mesh.cu:357-358)
  vert_trans   f32[O,16]  column-major local-to-world;  normal_trans f32[O,16] (mesh.cu:371-384)
  lights       LIGHT_DTYPE[L] in std::map name order (mesh.cu:389-394)
  vfov, aspect, near, c2w[16]   camera (mesh.cu:143-152; include/camera.h)

One declared deviation: files without a `materials` array (ball.gltf, square.gltf) make the reference
read materials[0] of an empty vector (src/mesh.cu:225,321,358 — undefined behaviour).  Here a single
default-constructed Material (include/material.h:88-103) is injected, as SURVEY §8(d)-2 prescribes.
"""
from __future__ import annotations

import base64
import json
from pathlib import Path

import numpy as np

from . import pyoracle as orc
from .pyoracle import LIGHT_DTYPE, MATERIAL_DTYPE, MTL_INTERVAL_DTYPE

_COMPONENT = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5124: np.int32, 5125: np.uint32,
              5126: np.float32}
WATTS_PER_LUMEN = np.float32(1.0) / np.float32(683.0)   # include/delta_light.h:7


def _buffers(g, base: Path):
    out = []
    for b in g.get("buffers", []):
        uri = b["uri"]
        if uri.startswith("data:"):
            out.append(base64.b64decode(uri.split(",", 1)[1]))
        else:
            out.append((base / uri).read_bytes())
    return out


def _raw(g, bufs, acc_idx, dtype, ncomp):
    """mesh.cu:165-170 — pointer = buffer.data + accessor.byteOffset + bufferView.byteOffset, tightly packed."""
    acc = g["accessors"][acc_idx]
    bv = g["bufferViews"][acc["bufferView"]]
    off = acc.get("byteOffset", 0) + bv.get("byteOffset", 0)
    n = acc["count"] * ncomp
    return np.frombuffer(bufs[bv["buffer"]], dtype=dtype, count=n, offset=off).copy()


def _trs(node):
    """mesh.cu:103-139 readTransform: missing rotation -> Quat() (all zero, which RotateFromQuat maps
    to identity), missing scale -> 1, missing translation -> 0."""
    q = node.get("rotation") or [0.0, 0.0, 0.0, 0.0]
    s = node.get("scale") or [1.0, 1.0, 1.0]
    t = node.get("translation") or [0.0, 0.0, 0.0]
    return orc.trs_to_mat4(t, q, s)


def load_gltf(path) -> dict:
    path = Path(path)
    g = json.loads(path.read_text())
    bufs = _buffers(g, path.parent)
    gl_materials = g.get("materials", [])
    gl_lights = g.get("extensions", {}).get("KHR_lights_punctual", {}).get("lights", [])

    cam = dict(vfov=np.float32(60.0), aspect=np.float32(1.77778), near=np.float32(0.1),   # camera.h:12
               c2w=orc.trs_to_mat4([0, 0, 0], [0, 0, 0, 0], [1, 1, 1]))
    meshes = []          # (indices u32, verts, normals, texcoords, material name, l2w)
    materials = {}       # name -> record (first definition wins, mesh.cu:228)
    lights = {}          # node name -> record

    for node in g.get("nodes", []):
        light_idx = node.get("extensions", {}).get("KHR_lights_punctual", {}).get("light", -1)
        if node.get("camera", -1) > -1:
            ci = g["cameras"][node["camera"]]
            if ci.get("type") == "perspective":
                p = ci["perspective"]
                cam = dict(vfov=np.float32(p["yfov"]), aspect=np.float32(p.get("aspectRatio", 0.0)),
                           near=np.float32(p.get("znear", 0.0)), c2w=_trs(node))
        elif node.get("mesh", -1) > -1:
            prim = g["meshes"][node["mesh"]]["primitives"][0]
            attrs = prim["attributes"]
            pos = _raw(g, bufs, attrs["POSITION"], np.float32, 3).reshape(-1, 3)
            iacc = g["accessors"][prim["indices"]]
            idx = _raw(g, bufs, prim["indices"], _COMPONENT[iacc["componentType"]], 1).astype(np.uint32)
            nrm = _raw(g, bufs, attrs["NORMAL"], np.float32, 3).reshape(-1, 3)
            if "TEXCOORD_0" in attrs:
                tex = _raw(g, bufs, attrs["TEXCOORD_0"], np.float32, 2).reshape(-1, 2)
            else:
                tex = np.zeros((pos.shape[0], 2), np.float32)
            mname = ""
            if gl_materials:
                gm = gl_materials[prim.get("material", 0)]
                mname = gm.get("name", "")
                if mname not in materials:
                    rec = np.zeros((), MATERIAL_DTYPE)
                    rec["base"] = (0.82, 0.67, 0.16)            # Material() defaults, material.h:88-103
                    rec["unused"] = (0, 0.5, 0.5, 0, 0, 0, 0, 0, 1.0)
                    pbr = gm.get("pbrMetallicRoughness", {})
                    rec["unused"][2] = np.float32(pbr.get("roughnessFactor", 1.0))   # roughness (never read)
                    rec["metallic"] = np.float32(pbr.get("metallicFactor", 1.0))     # glTF default 1.0
                    bc = pbr.get("baseColorFactor", [1.0, 1.0, 1.0, 1.0])
                    rec["base"] = (np.float32(bc[0]), np.float32(bc[1]), np.float32(bc[2]))
                    for key, val in gm.get("extensions", {}).items():               # mesh.cu:243-257
                        if key == "KHR_materials_transmission":
                            rec["unused"][1] = np.float32(1.0) - np.float32(val["transmissionFactor"]) / np.float32(5.0)
                        if key == "KHR_materials_emissive_strength":
                            rec["emission"] = np.float32(val["emissiveStrength"])
                        if key == "KHR_materials_ior":
                            rec["eta"] = np.float32(val["ior"])
                    materials[mname] = rec
            meshes.append((idx, pos, nrm, tex, mname, _trs(node)))
        elif light_idx > -1:
            name = node.get("name", "")
            L = gl_lights[light_idx]
            rec = np.zeros((), LIGHT_DTYPE)
            color = L.get("color", [1.0, 1.0, 1.0])
            rec["color"] = tuple(np.float32(c) for c in color)
            inten = np.float32(L.get("intensity", 1.0))
            m = _trs(node)
            if L["type"] == "point":
                rec["type"] = 0
                rec["intensity"] = inten * WATTS_PER_LUMEN
                rec["a"] = orc.mat4_mul_point(m, [0, 0, 0], 1.0)
            elif L["type"] == "directional":
                rec["type"] = 1
                rec["intensity"] = inten
                rec["a"] = orc.mat4_mul_point(m, [0, 0, -1], 0.0)
            elif L["type"] == "spot":
                rec["type"] = 2
                rec["intensity"] = inten * WATTS_PER_LUMEN
                spot = L.get("spot", {})
                inner = np.float32(spot.get("innerConeAngle", 0.0))
                outer = np.float32(spot.get("outerConeAngle", 0.7853981633974483))
                co, ci_ = np.float32(orc.cosf(outer)), np.float32(orc.cosf(inner))
                rec["cos_outer"] = co
                rec["inv_cone_diff"] = np.float32(1.0) / (ci_ - co)
                rec["b"] = orc.mat4_mul_point(m, [0, 0, -1], 0.0)
                rec["a"] = orc.mat4_mul_point(m, [0, 0, 0], 1.0)
            else:
                raise RuntimeError("Unsupported light type")   # mesh.cu:303
            lights[name] = rec

    if not materials:   # declared deviation: default material instead of the reference's OOB read
        rec = np.zeros((), MATERIAL_DTYPE)
        rec["base"] = (0.82, 0.67, 0.16)
        rec["unused"] = (0, 0.5, 0.5, 0, 0, 0, 0, 0, 1.0)
        materials[""] = rec

    def _key(s):   # std::map<std::string> orders by bytes
        return s.encode("utf-8")

    mat_names = sorted(materials, key=_key)
    mat_index = {n: i for i, n in enumerate(mat_names)}
    mats = np.array([materials[n] for n in mat_names], MATERIAL_DTYPE)

    indices, verts, norms, texs, objs, vts, nts = [], [], [], [], [], [], []
    n_idx = n_vtx = 0
    for idx, pos, nrm, tex, mname, l2w in meshes:
        indices.append(idx + np.uint32(n_vtx))
        verts.append(pos)
        norms.append(nrm)
        texs.append(tex)
        objs.append((n_idx // 3, mat_index.get(mname, 0)))
        vts.append(l2w)
        nts.append(orc.normal_matrix(l2w))
        n_idx += idx.size
        n_vtx += pos.shape[0]

    light_names = sorted(lights, key=_key)
    return dict(
        indices=np.ascontiguousarray(np.concatenate(indices) if indices else np.zeros(0, np.uint32), np.uint32),
        vertices=np.ascontiguousarray(np.concatenate(verts), np.float32),
        normals=np.ascontiguousarray(np.concatenate(norms), np.float32),
        texcoords=np.ascontiguousarray(np.concatenate(texs), np.float32),
        materials=mats,
        objects=np.array(objs, MTL_INTERVAL_DTYPE),
        vert_trans=np.ascontiguousarray(np.stack(vts), np.float32),
        normal_trans=np.ascontiguousarray(np.stack(nts), np.float32),
        lights=np.array([lights[n] for n in light_names], LIGHT_DTYPE) if lights else np.zeros(0, LIGHT_DTYPE),
        vfov=cam["vfov"], aspect=cam["aspect"], near=cam["near"], c2w=np.ascontiguousarray(cam["c2w"], np.float32),
        material_names=np.array(mat_names), light_names=np.array(light_names),
    )


_SCENE_KEYS = ["indices", "vertices", "normals", "texcoords", "materials", "objects", "vert_trans", "normal_trans", "lights",
               "vfov", "aspect", "near", "c2w", "material_names", "light_names"]


def save_scene(scene: dict, path) -> None:
    np.savez_compressed(path, **{k: scene[k] for k in _SCENE_KEYS})


def load_scene(path) -> dict:
    z = np.load(path, allow_pickle=False)
    d = {k: z[k] for k in z.files}
    d["materials"] = np.ascontiguousarray(d["materials"]).view(MATERIAL_DTYPE).reshape(-1) if d["materials"].dtype != MATERIAL_DTYPE else d["materials"]
    for k in ("vfov", "aspect", "near"):
        d[k] = np.float32(d[k])
    return d
