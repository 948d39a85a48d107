"""Executed warp instructions of one kernel attributed to source lines.

  python tools/probes/sass_attrib.py <report.ncu-rep> <object.o> <mangled-kernel-substring> [file:lo-hi ...]

Joins the per-instruction 'Instructions Executed' column of an ncu --set full capture (source page) with the inline
line table of the SAME build (nvdisasm -gi on the cubin inside the object), by instruction index.  Prints the totals per
innermost source line and, for each file:lo-hi range given, per line of that range anywhere in the inline chain (i.e. "cost
of everything called from this line")."""
import collections, csv, io, re, subprocess, sys, tempfile, os

rep, obj, kern = sys.argv[1:4]
ranges = []
for r in sys.argv[4:]:
    f, lh = r.split(":"); lo, hi = lh.split("-"); ranges.append((f, int(lo), int(hi)))
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-gi", "-c", cubin], capture_output=True, text=True).stdout.split("\n")
start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and kern in l)
inst, pending, cur = [], [], []
for l in dis[start + 1:]:
    if l.startswith("\t.section") or (l.startswith(".text.") and kern not in l):
        break
    if l.strip().startswith("//## File"):
        pending.append(l)
    elif re.match(r"\s+/\*[0-9a-f]{4,5}\*/", l):
        if pending:
            cur, pending = pending, []
        inst.append((l.strip(), cur))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
ix = {k: i for i, k in enumerate(rows[h])}
R = rows[h + 1:]
assert len(R) == len(inst), (len(R), len(inst), "report and object are different builds")
def frames(a):
    return [(m.group(1).split("/")[-1], int(m.group(2))) for l in a for m in re.finditer(r'File "([^"]+)", line (\d+)', l)]
inner, chain, ops = collections.Counter(), collections.Counter(), collections.Counter()
tot = 0
for (txt, a), r in zip(inst, R):
    ex = int(r[ix["Instructions Executed"]] or 0)
    tot += ex
    fr = frames(a)
    if fr:
        inner[fr[0]] += ex
    for f, lo, hi in ranges:
        hit = [ln for (ff, ln) in fr if ff == f and lo <= ln <= hi]
        if hit:
            chain[(f, hit[-1])] += ex
    m = re.match(r"/\*[0-9a-f]+\*/\s+(@!?U?P\w+\s+)?([A-Z0-9_]+)", txt)
    ops[m.group(2) if m else "?"] += ex
warps = int(R[0][ix["Instructions Executed"]] or 1)
print(f"total {tot} warp instructions, {tot / warps:.1f} per warp ({warps} warps)")
print("-- by opcode:", ", ".join(f"{k} {v / warps:.0f}" for k, v in ops.most_common(24)))
print("-- innermost line")
for k, v in inner.most_common(40):
    print(f"  {k[0]}:{k[1]:<5d} {v / warps:8.1f}")
if ranges:
    print("-- outermost line within the given ranges")
    for k, v in sorted(chain.items(), key=lambda kv: -kv[1])[:40]:
        print(f"  {k[0]}:{k[1]:<5d} {v / warps:8.1f}")
