"""Stall samples per CUDA source line of one kernel: joins the SASS page of an ncu report with the
line table of the built library (nvdisasm --print-line-info), instruction by instruction.
usage: python profiles/ncu_source_lines.py <report.ncu-rep> <kernel regex> <mangled-name substring> [top N] [lib.so]"""
import collections, csv, os, re, subprocess, sys, tempfile
rep, pat, mangled = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
lib = sys.argv[5] if len(sys.argv) > 5 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                          "jaxopt-3d-reconstruction_b200", "libb200ba.so")
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{pat}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; blocks.append(cur); continue
    if r and r[0] == "Address":
        cur["hdr"] = r; continue
    if cur is not None and r:
        cur["rows"].append(r)
sub = os.environ.get('NCU_NAME_SUB')  # pick among template instantiations: substring of the demangled name
b = [x for x in blocks if not sub or sub in x['name']][0]
h = b["hdr"]
isrc, isamp, iex = h.index("Source"), h.index("Warp Stall Sampling (All Samples)"), h.index("Instructions Executed")
# line table
tmp = tempfile.mkdtemp()
subprocess.run(f"cd {tmp} && cuobjdump -xelf all {lib} > /dev/null", shell=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.splitlines()
lines, inside, curline = [], False, ("?", 0)
for ln in dis:
    if ln.startswith("\t.section\t.text."):
        inside = mangled in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        curline = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
        lines.append(curline)
n = min(len(lines), len(b["rows"]))
print(b["name"][:70], f"| sass instrs ncu {len(b['rows'])} / nvdisasm {len(lines)}")
agg = collections.defaultdict(lambda: [0, 0])
tot = 0
for i in range(n):
    r = b["rows"][i]
    s = int(r[isamp] or 0); e = int(r[iex] or 0)
    agg[lines[i]][0] += s; agg[lines[i]][1] += e; tot += s
src_cache = {}
def src(fn, ln):
    if fn not in src_cache:
        path = os.path.join(os.path.dirname(lib), "csrc", fn)
        src_cache[fn] = open(path).read().splitlines() if os.path.exists(path) else []
    L = src_cache[fn]
    return L[ln - 1].strip()[:100] if 0 < ln <= len(L) else ""
print("total samples", tot)
for (fn, ln), (s, e) in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print(f"{100 * s / max(tot, 1):5.1f}% {s:6d}  ex={e:9d}  {fn}:{ln:<4d} {src(fn, ln)}")
