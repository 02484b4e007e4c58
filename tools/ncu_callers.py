#!/usr/bin/env python
"""Caller view of an ncu capture: warp instructions / thread instructions / stall samples per source line,
where a line of the arithmetic helpers (pvec.cuh, the pinned math header, CUDA intrinsic headers) is charged
to the first frame of its inline chain that lies outside them (nvdisasm -gi gives the chain).
usage: ncu_callers.py report.ncu-rep kernel_mangled_substring [n_lines]"""
import csv, io, re, subprocess, sys, glob, os, collections, tempfile
rep, ksub = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 50
HELPERS = ("pvec.cuh", "wgpt_pinned_math.h", "intrinsics.hpp", "cuda_fp16.hpp", "device_atomic_functions.hpp", "math_functions.hpp", "device_functions.hpp")
tmp = tempfile.mkdtemp()
lib = os.environ.get("WGPT_LIB") or os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "webgpu_path_tracer_b200", "libwgpt.so")
subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, capture_output=True)
lines = {}
for cb in glob.glob(os.path.join(tmp, "*.cubin")):
    dis = subprocess.run(["nvdisasm", "-gi", "-c", cb], capture_output=True, text=True).stdout
    if ksub not in dis: continue
    infn = False; chain = []; fresh = True
    for ln in dis.split("\n"):
        if ln.startswith(".text.") and ln.endswith(":"):
            infn = ksub in ln; continue
        if not infn: continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            if fresh: chain = []; fresh = False
            chain.append((os.path.basename(m.group(1)), int(m.group(2))))
            continue
        m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*);', ln)
        if m:
            fresh = True
            loc = next((c for c in chain if not c[0].endswith(HELPERS)), chain[0] if chain else ("?", 0))
            lines[int(m.group(1), 16)] = loc
    if lines: break
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
h = rows[hi]; body = rows[hi + 1:]
base = int(body[0][0], 16)
ci, ti, si = h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("# Samples")
agg = collections.defaultdict(lambda: [0, 0, 0]); tot = [0, 0, 0]
for r in body:
    if len(r) <= ti: continue
    loc = lines.get(int(r[0], 16) - base, ("?", 0))
    v = (int(r[ci] or 0), int(r[ti] or 0), int(r[si] or 0))
    for k in range(3): agg[loc][k] += v[k]; tot[k] += v[k]
print("total warp-inst %d thread-inst %d samples %d" % tuple(tot))
for (f, l), v in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print("  %-18s:%4d warp-inst %5.2f%% thread-inst %5.2f%% samples %5.2f%%  eff %4.1f" % (f, l, 100 * v[0] / tot[0], 100 * v[1] / tot[1], 100 * v[2] / max(1, tot[2]), v[1] / max(1, v[0])))
