#!/usr/bin/env python
"""Join an ncu SASS source page with nvdisasm line info: instructions / thread-instructions / stall
samples per source line (innermost inlined location) and per source function bucket.
usage: ncu_lines.py report.ncu-rep kernel_mangled_substring [cubin_substring]"""
import csv, io, re, subprocess, sys, glob, os, collections, tempfile
rep, ksub = sys.argv[1], sys.argv[2]
tmp = tempfile.mkdtemp()
lib = os.environ.get("WGPT_LIB") or os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "webgpu_path_tracer_b200", "libwgpt.so")
subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, capture_output=True)
lines = {}
for cb in glob.glob(os.path.join(tmp, "*.cubin")):
    dis = subprocess.run(["nvdisasm", "-g", "-c", cb], capture_output=True, text=True).stdout
    if ksub not in dis: continue
    cur = None; infn = False
    for ln in dis.split("\n"):
        if ln.startswith(".text.") and ln.endswith(":"):
            infn = ksub in ln; continue
        if not infn: continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
        m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*);', ln)
        if m: lines[int(m.group(1), 16)] = (cur, m.group(2).strip())
    if lines: break
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
h = rows[hi]; body = rows[hi + 1:]
base = int(body[0][0], 16)
ci, ti, si = h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("# Samples")
agg = collections.defaultdict(lambda: [0, 0, 0])
tot = [0, 0, 0]
for r in body:
    if len(r) <= ti: continue
    off = int(r[0], 16) - base
    loc = lines.get(off, (None, ""))[0] or ("?", 0)
    v = (int(r[ci] or 0), int(r[ti] or 0), int(r[si] or 0))
    for k in range(3): agg[loc][k] += v[k]; tot[k] += v[k]
print("total warp-inst %d thread-inst %d samples %d" % tuple(tot))
byfile = collections.defaultdict(lambda: [0, 0, 0])
for (f, l), v in agg.items():
    for k in range(3): byfile[f][k] += v[k]
for f, v in sorted(byfile.items(), key=lambda x: -x[1][0]):
    print("  %-22s warp-inst %5.1f%%  thread-inst %5.1f%%  samples %5.1f%%" % (f, 100 * v[0] / tot[0], 100 * v[1] / tot[1], 100 * v[2] / max(1, tot[2])))
print("top lines:")
for (f, l), v in sorted(agg.items(), key=lambda x: -x[1][0])[:45]:
    print("  %-18s:%4d warp-inst %5.2f%% thread-inst %5.2f%% samples %5.2f%%  eff %4.1f" % (f, l, 100 * v[0] / tot[0], 100 * v[1] / tot[1], 100 * v[2] / max(1, tot[2]), v[1] / max(1, v[0])))
