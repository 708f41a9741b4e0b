"""Attribute ncu's per-SASS-instruction stall samples to CUDA source lines (the CLI's --print-source cuda
view carries no metrics).  usage: python tools/ncu_lines.py <report.ncu-rep> <lib.so> [top]
Needs: ncu, cuobjdump, nvdisasm.  The .so must be the build that was profiled (compiled with -lineinfo)."""
import csv, io, os, re, subprocess, sys, tempfile
rep, so = sys.argv[1], os.path.abspath(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 60
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(k for k, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: k for k, h in enumerate(hdr)}
ins = []
for r in rows[hi + 1:]:                      # first captured launch only
    if r and r[0] in ("Kernel Name", "Address"):
        break
    if len(r) == len(hdr):
        ins.append(r)
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin") and "map_compiler" not in f][0]
sass = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
# sections: name -> list of (opcode text, file, line)
secs, cur, loc = {}, None, ("?", 0)
for ln in sass.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", ln)
    if m: cur = m.group(1); secs[cur] = []; continue
    m = re.match(r'\s*//## File "(.*)", line (\d+)', ln)
    if m: loc = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?);", ln)
    if m and cur: secs[cur].append((m.group(1).strip(), loc))
# ncu lists the kernel's functions back to back; align each section by its opcode sequence
def norm(s): return re.sub(r"\s+", " ", s.split("//")[0]).strip().rstrip(";").split(" ")[0]
nc = [norm(r[col["Source"]]) for r in ins]
lines = [None] * len(ins)
pos = 0
order = sorted(secs.items(), key=lambda kv: -len(kv[1]))
for name, body in order:
    ops = [norm(o) for o, _ in body]
    n = len(ops)
    for start in range(0, len(nc) - n + 1):
        if lines[start] is None and nc[start:start + n] == ops:
            for k in range(n): lines[start + k] = (name, body[k][1])
            break
agg, fagg, tot = {}, {}, 0
S = col["# Samples"]
reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for r, l in zip(ins, lines):
    s = int(r[S]); tot += s
    key = l[1] if l else ("unmapped", 0)
    a = agg.setdefault(key, [0, 0, {}]); a[0] += s; a[1] += int(r[col["Instructions Executed"]])
    for h in reasons:
        v = int(r[col[h]]);
        if v: a[2][h] = a[2].get(h, 0) + v
    fn = l[0] if l else "unmapped"
    fagg[fn] = fagg.get(fn, 0) + s
print("total samples", tot, " instructions", len(ins), " mapped", sum(l is not None for l in lines))
print("by function:")
for f, v in sorted(fagg.items(), key=lambda x: -x[1])[:12]: print(f"  {v / tot:6.1%}  {f[:100]}")
byfile = {}
for (f, l), a in agg.items(): byfile[f] = byfile.get(f, 0) + a[0]
print("by file:", {f: f"{v / tot:.1%}" for f, v in sorted(byfile.items(), key=lambda x: -x[1])})
src_cache = {}
def src(f, l):
    for d in ("nclone_b200/csrc", "include"):
        p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), d, f)
        if os.path.exists(p):
            if p not in src_cache: src_cache[p] = open(p).read().splitlines()
            return src_cache[p][l - 1].strip()[:100] if 0 < l <= len(src_cache[p]) else ""
    return ""
print(f"top {top} source lines by stall samples:")
for (f, l), a in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    rs = sorted(a[2].items(), key=lambda x: -x[1])[:2]
    print(f"  {a[0] / tot:6.2%} inst {a[1]:>9}  {f}:{l:<4} [{', '.join(k[6:] + ' ' + format(v / max(a[0], 1), '.0%') for k, v in rs)}]  {src(f, l)}")
