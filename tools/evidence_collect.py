#!/usr/bin/env python
"""Turn what tools/evidence.sh left in gpurun_out/ into the tracked files under profiles/:
python tools/evidence_collect.py [r01]"""
import csv, io, json, os, shutil, subprocess, sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
R = sys.argv[1] if len(sys.argv) > 1 else "r01"
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")

UNITS = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2]


traffic = {}
for cfg in ("C2", "C4"):
    traffic[cfg] = {"source": "ncu --set full --clock-control none, launch 20 (-s 20 -c 1) of tools/prof_run.py %s 64 "
                              "(steady state, 8 Mi-slot pool): dram__bytes_read.sum + dram__bytes_write.sum" % cfg}
    for st in ("extend", "shade"):
        rep = os.path.join(G, "%s_%s_%s.ncu-rep" % (st, cfg, R))
        if not os.path.exists(rep):
            continue
        txt = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), rep], capture_output=True, text=True).stdout
        open(os.path.join(P, "%s_%s_%s_summary.txt" % (st, cfg, R)), "w").write(
            "# ncu --set full --clock-control none --import-source on -k regex:k_wf_%s -s 20 -c 1 python tools/prof_run.py %s 64\n" % (st, cfg) + txt)
        h, u, v = raw(rep)
        tot = 0.0
        for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = h.index(k)
            tot += float(v[i].replace(",", "")) * UNITS[u[i]]
        traffic[cfg]["%s_dram_bytes_per_launch" % st] = tot
json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=1)

# launch list of the bench command
src = os.path.join(G, "launches_%s.csv" % R)
if os.path.exists(src):
    shutil.copy(src, os.path.join(P, "launches_%s.csv" % R))
    lines = [l for l in open(src) if not l.startswith("==")]
    rows = list(csv.reader(io.StringIO("".join(lines))))
    h = rows[0]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    tot, cnt = defaultdict(float), defaultdict(int)
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}
    for r in rows[1:]:
        if len(r) <= vi:
            continue
        name = r[ki].split("(")[0][:48]
        tot[name] += float(r[vi].replace(",", "")) * scale.get(r[ui], 1e-3)
        cnt[name] += 1
    all_ms = sum(tot.values())
    with open(os.path.join(P, "launches_%s_summary.txt" % R), "w") as f:
        f.write("kernel launches of `python bench.py --steps 2 --warmup 1 --no-cpu-baseline` under ncu "
                "(gpu__time_duration.sum, --clock-control none; cold cache, serialised)\n")
        f.write("%-50s %6s %12s %10s %7s\n" % ("kernel", "n", "total ms", "mean us", "share"))
        for k in sorted(tot, key=lambda k: -tot[k]):
            f.write("%-50s %6d %12.2f %10.1f %6.1f%%\n" % (k, cnt[k], tot[k], 1e3 * tot[k] / cnt[k], 100 * tot[k] / all_ms))

for name in ("bench_C1_%s.json" % R, "bench_C3_%s.json" % R, "bench_C4_%s.json" % R, "bench_C5_%s.json" % R, "visits_%s.txt" % R, "bench_%s.json" % R, "bench_ref_%s.json" % R, "k4_%s.txt" % R, "stages_%s.txt" % R, "pytest_gpu_%s.log" % R):
    if os.path.exists(os.path.join(G, name)):
        shutil.copy(os.path.join(G, name), os.path.join(P, name))
print(open(os.path.join(P, "traffic.json")).read())
print(open(os.path.join(P, "launches_%s_summary.txt" % R)).read()[:1800])
