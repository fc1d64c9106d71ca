"""Builds the reference's OWN sources from /root/reference into oracle/_ref/ (git-ignored; travels to the GPU box).

Nothing of the reference is copied into the repository.  Outputs:
  oracle/_ref/tpt_ref_headless[_d32]   headless executable: src/{path_tracer,bvh,mesh,env_light,texture}.cu, -O3 sm_100
  oracle/_ref/tpt_ref_headless_G       the same with the shipped `-G` flag (CMakeLists.txt:59), for information
  oracle/_ref/libref_host.so           the reference's CUDA_CALLABLE functions compiled for the host + loader hooks
  oracle/_ref/input/*.gltf             the reference's scenes (ball/square with one default material injected, see below)
  oracle/_ref/sky.png, black.png       environment maps (synthetic sky of SURVEY §8(d)-2; the shipped JPG is a missing blob)
  oracle/_ref/BUILD.json               what was built, with which edits

Edits, all in GENERATED copies under oracle/_ref/patched/ (recorded in BUILD.json):
  src/path_tracer.cu:17   `constexpr int DEPTH_TRACE = 8;`    -> value from -DTPT_REF_DEPTH (default 8)
  src/path_tracer.cu:559  `int nSamplesPerPixel = 64;`         -> tpt_ref_spp()  (run-time --spp)
  window extent (include/vkEngine.h:24) is overridden through the VkExtent2D stand-in's constructor, no source edit.
  input/ball.gltf, input/square.gltf: no `materials` array -> the reference reads materials[0] of an empty vector
  (src/mesh.cu:225,321,358); the copies get one material equal to the default-constructed Material().
"""
from __future__ import annotations

import json
import os
import re
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent
REF = Path("/root/reference")
OUT = HERE / "_ref"
NVCC = "/usr/local/cuda/bin/nvcc"
COMMON = ["-std=c++17", "-rdc=true", "-include", "cfloat", "-DGLFW_INCLUDE_NONE", "-ccbin", "/usr/bin/g++", "-w",
          f"-I{OUT / 'patched'}", f"-I{HERE / 'ref_shims'}", f"-I{REF / 'include'}", f"-I{REF / 'lib'}", f"-I{REF / 'src'}"]


def run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(map(str, cmd)) + "\n" + r.stdout[-3000:] + r.stderr[-3000:])
        raise RuntimeError("reference build step failed")


def patch_sources():
    p = OUT / "patched"
    p.mkdir(parents=True, exist_ok=True)
    src = (REF / "src" / "path_tracer.cu").read_text()
    a = "constexpr int DEPTH_TRACE = 8;"
    b = "int nSamplesPerPixel = 64;"
    assert src.count(a) == 1 and src.count(b) == 1, "reference source changed: patch points not found"
    src = src.replace(a, "#ifndef TPT_REF_DEPTH\n#define TPT_REF_DEPTH 8\n#endif\n\tconstexpr int DEPTH_TRACE = TPT_REF_DEPTH;")
    src = src.replace(b, "int nSamplesPerPixel = tpt_ref_spp();")
    src = "int tpt_ref_spp();\n" + src
    # counting build: one atomicAdd per traverseBVH call in the DEVICE compilation of the generated copy, compiled in
    # only with -DTPT_REF_COUNT_RAYS (tpt_ref_headless_count); the timed builds are untouched
    c = "\tint call_stack[64];\n\tint stackPtr = 1;"
    assert src.count(c) == 1, "reference source changed: traverseBVH patch point not found"
    src = src.replace(c, "#if defined(TPT_REF_COUNT_RAYS) && defined(__CUDA_ARCH__)\n\tatomicAdd(&g_tpt_ref_rays, 1ull);\n#endif\n" + c)
    src = "#ifdef TPT_REF_COUNT_RAYS\n__device__ unsigned long long g_tpt_ref_rays;\n#endif\n" + src
    (p / "path_tracer.cu").write_text(src)


def add_default_material(text: str) -> str:
    g = json.loads(text)
    if g.get("materials"):
        return text
    g["materials"] = [{"name": "", "pbrMetallicRoughness": {"baseColorFactor": [0.82, 0.67, 0.16, 1.0], "metallicFactor": 0.0,
                                                             "roughnessFactor": 0.5}}]
    for m in g["meshes"]:
        for prim in m["primitives"]:
            prim["material"] = 0
    return json.dumps(g)


def copy_inputs():
    d = OUT / "input"
    d.mkdir(parents=True, exist_ok=True)
    for f in sorted((REF / "input").glob("*.gltf")):
        (d / f.name).write_text(add_default_material(f.read_text()))
    sys.path.insert(0, str(ROOT))
    from PIL import Image
    from tinypathtracer_b200.sky import synthetic_sky
    sky = synthetic_sky(4096, 2048)                      # row 0 = bottom -> PNG rows are top-down
    Image.fromarray(sky[::-1, :, :3]).save(OUT / "sky.png")
    import numpy as np
    Image.fromarray(np.zeros((32, 64, 3), np.uint8)).save(OUT / "black.png")


def compile_units(jobs):
    """The translation units of one target, compiled side by side (each nvcc is single-threaded; a target has five)."""
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=min(len(jobs), max(1, (os.cpu_count() or 2)))) as pool:
        list(pool.map(run, jobs))   # re-raises the first failure


def up_to_date(target: Path) -> bool:
    """The target exists and is newer than everything it is made from: the reference's sources, the generated copy's inputs here, this script."""
    if not target.exists():
        return False
    t = target.stat().st_mtime
    deps = list(HERE.glob("ref_*")) + list((REF / "src").glob("*")) + list((REF / "include").rglob("*"))
    return all(d.stat().st_mtime <= t for d in deps if d.is_file())


def build_exe(name: str, opt: list, depth: int, tag: str):
    objs = []
    work = OUT / f"obj_{name}"
    work.mkdir(exist_ok=True)
    units = [(HERE / "ref_headless.cu", "headless"), (HERE / "ref_hooks_bvh.cu", "bvh"), (REF / "src" / "mesh.cu", "mesh"),
             (REF / "src" / "env_light.cu", "env_light"), (REF / "src" / "texture.cu", "texture")]
    jobs = []
    for src, stem in units:
        o = work / f"{stem}.o"
        jobs.append([NVCC, *COMMON, *opt, f"-DTPT_REF_DEPTH={depth}", f'-DTPT_REF_BUILD="{tag}"', "-c", str(src), "-o", str(o)])
        objs.append(str(o))
    compile_units(jobs)
    fo = work / "freeimage_shim.o"
    run(["/usr/bin/g++", "-O2", "-c", f"-I{REF / 'include'}", str(HERE / "ref_freeimage_shim.cpp"), "-o", str(fo)])
    run([NVCC, *opt, "-rdc=true", "-ccbin", "/usr/bin/g++", *objs, str(fo), "-o", str(OUT / name)])


def build_host_lib():
    work = OUT / "obj_host"
    work.mkdir(exist_ok=True)
    opt = ["-O2", "-arch=sm_100", "-Xcompiler", "-fPIC", "-DTPT_REF_NO_MAIN", '-DTPT_REF_BUILD="host"']
    objs, jobs = [], []
    for src, stem in [(HERE / "ref_headless.cu", "headless"), (HERE / "ref_hooks_bvh.cu", "bvh"), (HERE / "ref_hooks_mesh.cu", "mesh"),
                      (REF / "src" / "env_light.cu", "env_light"), (REF / "src" / "texture.cu", "texture")]:
        o = work / f"{stem}.o"
        jobs.append([NVCC, *COMMON, *opt, "-c", str(src), "-o", str(o)])
        objs.append(str(o))
    compile_units(jobs)
    fo = work / "freeimage_shim.o"
    run(["/usr/bin/g++", "-O2", "-fPIC", "-c", f"-I{REF / 'include'}", str(HERE / "ref_freeimage_shim.cpp"), "-o", str(fo)])
    run([NVCC, "-arch=sm_100", "-rdc=true", "-shared", "-Xcompiler", "-fPIC", "-ccbin", "/usr/bin/g++", *objs, str(fo), "-o",
         str(OUT / "libref_host.so")])


def main():
    if not (REF / "src" / "path_tracer.cu").exists():
        print("oracle/ref_build.py: /root/reference is not present; nothing to do")
        return 0
    OUT.mkdir(exist_ok=True)
    patch_sources()
    copy_inputs()
    built = {}
    try:   # a partial rebuild keeps the record of the other targets
        built = json.loads((OUT / "BUILD.json").read_text()).get("built", {})
    except Exception:
        pass
    only = {a for a in sys.argv[1:] if not a.startswith("--")}
    targets = [("libref_host.so", None), ("tpt_ref_headless", (["-O3", "-arch=sm_100"], 8, "-O3 -arch=sm_100")),
               ("tpt_ref_headless_d32", (["-O3", "-arch=sm_100"], 32, "-O3 -arch=sm_100 DEPTH_TRACE=32")),
               ("tpt_ref_headless_G", (["-G", "-arch=sm_100"], 8, "-G -arch=sm_100 (as shipped, CMakeLists.txt:59)")),
               ("tpt_ref_headless_count", (["-O3", "-arch=sm_100", "-DTPT_REF_COUNT_RAYS"], 8, "-O3 -arch=sm_100 + one atomicAdd per traverseBVH call (ray counting; not timed)")),
               ("tpt_ref_headless_count_d32", (["-O3", "-arch=sm_100", "-DTPT_REF_COUNT_RAYS"], 32, "-O3 -arch=sm_100 DEPTH_TRACE=32 + ray counting (not timed)"))]
    for name, spec in targets:
        if only and name not in only:
            continue
        try:
            if "--force" not in sys.argv and built.get(name) == "ok" and up_to_date(OUT / name):
                print(name, "up to date")
                continue
            if spec is None:
                build_host_lib()
            else:
                build_exe(name, *spec)
            built[name] = "ok"
        except Exception as e:   # report, keep going: the oracle port remains the fallback
            built[name] = f"FAILED: {e}"
        print(name, built[name])
    (OUT / "BUILD.json").write_text(json.dumps({"built": built, "edits": ["DEPTH_TRACE -> -DTPT_REF_DEPTH", "nSamplesPerPixel -> --spp",
                                                                           "window extent via VkExtent2D stand-in", "count builds only: one atomicAdd per traverseBVH call (-DTPT_REF_COUNT_RAYS)",
                                                                           "ball.gltf/square.gltf: default material injected"]}, indent=1))
    return 0


if __name__ == "__main__":
    sys.exit(main())
