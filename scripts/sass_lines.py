"""Instructions executed per SOURCE LINE of one kernel: joins `ncu --page source --csv` (per-SASS-instruction
counts) with `nvdisasm -g` line info of the library that was profiled.

  python scripts/sass_lines.py <rep.ncu-rep> <kernel substring> [lib.so] [topN] [launch index]

The library must be the build that was profiled (same SASS); offsets are matched by position in the kernel.
"""
import collections
import csv
import io
import re
import subprocess
import sys
import tempfile
from pathlib import Path

rep, kern = sys.argv[1], sys.argv[2]
lib = sys.argv[3] if len(sys.argv) > 3 else "tinypathtracer_b200/lib/libtpt_b200.so"
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
launch = int(sys.argv[5]) if len(sys.argv) > 5 else 0

# ---- line table from nvdisasm -g
tmp = Path(tempfile.mkdtemp())
subprocess.run(["cuobjdump", "-xelf", "all", str(Path(lib).resolve())], cwd=tmp, capture_output=True)
lines = {}
for cubin in tmp.glob("*.cubin"):
    if "-" in cubin.name.split(".")[0]:
        continue   # the merged image duplicates the per-file ones
    out = subprocess.run(["nvdisasm", "-g", "-c", str(cubin)], capture_output=True, text=True).stdout
    cur_fn, cur_line, tab = None, None, None
    for ln in out.splitlines():
        m = re.match(r"^\.text\.(\S+):", ln)
        if m:
            cur_fn = m.group(1)
            tab = lines.setdefault(cur_fn, {})
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur_line = (Path(m.group(1)).name, int(m.group(2)))
            continue
        m = re.match(r"^\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m and tab is not None:
            tab[int(m.group(1), 16)] = (cur_line, m.group(2).strip())
fn = [f for f in lines if kern in f]
if not fn:
    sys.exit(f"no kernel matching {kern!r} in {lib}")

# ---- per-instruction counts from the report
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], None
for r in csv.reader(io.StringIO(raw)):
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        blocks.append(cur)
        continue
    if cur is not None:
        cur["rows"].append(r)
blocks = [b for b in blocks if kern.split("IL")[0].replace("_ZN3tpt", "").lstrip("0123456789") in b["name"] or kern in b["name"]]
b = blocks[launch]
hdr, data = b["rows"][0], b["rows"][1:]
iA, iS, iE, iT = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
base = int(data[0][iA], 16)


def mismatches(tab):
    bad = 0
    for r in data:
        sass = tab.get(int(r[iA], 16) - base, (None, ""))[1]
        if sass.split()[:1] != r[iS].strip().replace("@", " ").split()[-0:1] and sass[:6] != r[iS].strip()[:6]:
            bad += 1
    return bad


# several template instantiations share the name: take the one whose SASS is the profiled one
tab = min((lines[f] for f in fn), key=mismatches)
per_line = collections.defaultdict(lambda: [0, 0])
tot = totT = 0
mismatch = 0
for r in data:
    off = int(r[iA], 16) - base
    n, t = int(r[iE]), int(r[iT])
    tot += n
    totT += t
    key, sass = tab.get(off, (None, ""))
    if sass.split()[:1] != r[iS].strip().replace("@", " ").split()[-0:1] and sass[:6] != r[iS].strip()[:6]:
        mismatch += 1
    per_line[key][0] += n
    per_line[key][1] += t
print(f"{b['name'][:70]}: {tot} warp-instr, {totT / max(tot, 1):.1f} lanes/instr, {mismatch} SASS mismatches (stale library?)")
src_cache = {}
def src(key):
    if not key:
        return ""
    f, l = key
    if f not in src_cache:
        cands = list(Path("tinypathtracer_b200/csrc").glob(f)) + list(Path("include").glob(f))
        src_cache[f] = cands[0].read_text().splitlines() if cands else []
    s = src_cache[f]
    return s[l - 1].strip()[:90] if 0 < l <= len(s) else ""
for key, (n, t) in sorted(per_line.items(), key=lambda kv: -kv[1][0])[:top]:
    where = f"{key[0]}:{key[1]}" if key else "?"
    print(f"{n:12d} {100 * n / tot:5.1f}% {t / max(n, 1):5.1f}  {where:18s} {src(key)}")
