"""Opcode histogram and hottest SASS instructions from `ncu --page source --csv`:  python scripts/sass_hot.py rep [topN]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + (["--kernel-name", sys.argv[3]] if len(sys.argv) > 3 else []), capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
# find header rows (one block per kernel launch); use the first block
blocks = []; cur = None
for r in rows:
    if r and r[0] == "Kernel Name": cur = {"name": r[1], "rows": []}; blocks.append(cur); continue
    if cur is not None: cur["rows"].append(r)
b = blocks[int(sys.argv[4]) if len(sys.argv) > 4 else 0]
hdr = b["rows"][0]; data = b["rows"][1:]
iS, iE, iT = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
tot = sum(int(r[iE]) for r in data); totT = sum(int(r[iT]) for r in data)
print(b["name"][:80], "warp-instr", tot, "thread-instr", totT, "avg lanes", totT / max(tot, 1))
h = collections.Counter()
for r in data:
    op = r[iS].split()[0] if not r[iS].strip().startswith("@") else r[iS].split()[1]
    h[op.split(".")[0]] += int(r[iE])
for op, n in h.most_common(25): print(f"  {op:12s} {n:12d} {100*n/tot:5.1f}%")
print("--- hottest")
for n, r in sorted(((int(r[iE]), r) for r in data), key=lambda x: -x[0])[:top]:
    print(f"{n:11d} {int(r[iT])/max(n,1):5.1f}  {r[iS].strip()[:90]}")
