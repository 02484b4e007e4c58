#!/usr/bin/env python
"""Turn what tools/evidence.sh left in gpurun_out/ into the tracked files under profiles/:
python tools/evidence_collect.py [r01]"""
import csv, io, json, os, shutil, subprocess, sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
R = sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("--") else "r02"
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")

def wide_counters():
    path = os.path.join(G, "wide_counters.json")
    if os.path.exists(path):
        return json.load(open(path))
    out = {}
    try:
        sys.path.insert(0, ROOT)
        import bench as _b
        for cfg in ("C2", "C4"):
            out[cfg] = _b.work_counters(cfg)
        json.dump(out, open(path, "w"))
    except Exception as ex:      # no GPU here: the file written on the box is used
        print("wide counters unavailable:", ex)
    return out


UNITS = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2]


sys.path.insert(0, ROOT)
import bench  # noqa: E402  (kernel_sha, byte constants)
wide = wide_counters()

COLS = {"ms": "gpu__time_duration.sum", "warp_inst": "smsp__inst_executed.sum",
        "issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "lanes": "smsp__thread_inst_executed_per_inst_executed.ratio",
        "long_sb": "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "l1_hit_pct": "l1tex__t_sector_hit_rate.pct", "l2_hit_pct": "lts__t_sector_hit_rate.pct",
        "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram_pct_of_peak": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"}
TSCALE = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}

# profiles/traffic.json: per (config, stage) the numbers of ONE captured launch (ncu -s 20 -c 1: iteration 20 of a 64-frame
# render, steady state) next to the algorithmic bytes of THAT launch, stamped with the hash of the kernel sources.
traffic = {"kernel_sha": bench.kernel_sha(),
           "source": "ncu --set full --clock-control none -k regex:k_wf_<stage> -s 20 -c 1 python tools/prof_run.py <cfg> 64"}
for cfg in ("C2", "C4"):
    hist_path = os.path.join(G, "prof_hist_%s.json" % cfg)
    hist = json.load(open(hist_path)) if os.path.exists(hist_path) else None
    traffic[cfg] = {}
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
        e = {"dram_bytes": tot}
        for name, col in COLS.items():
            if col in h:
                i = h.index(col)
                x = float(v[i].replace(",", ""))
                e[name] = x * TSCALE.get(u[i], 1.0) if name == "ms" else x
        if hist:
            n = hist["n_extend"][20]
            e["rays_in_launch"] = n
            if st == "extend":
                w = wide.get(cfg)
                if w and n:
                    e["algorithmic_bytes"] = n * w["B_ray_wide"]
                    e["what"] = "%d rays x (80 B x %.2f wide-node visits + 48 B x %.2f triangle tests + 60 B slot / queue I/O)" % (
                        n, w["wide_nodes_per_ray"], w["wide_tris_per_ray"])
                    e["warp_inst_per_unit"] = e.get("warp_inst", 0.0) / n
            else:
                live = hist["pool_slots"]
                e["algorithmic_bytes"] = live * bench.B_SHADE_BOUNCE
                e["what"] = "%d live slots (steady state: the whole pool) x 160 B; finished samples (48 B each) not counted" % live
        traffic[cfg][st] = e
json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=1)
if "--traffic-only" in sys.argv:
    print(json.dumps(traffic, indent=1))
    sys.exit(0)

# launch list of the bench command
src = os.path.join(G, "launches_%s.csv" % R)
if os.path.exists(src):
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
        f.write("kernel launches of `python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras` (C4, 1,024 frames per step) under ncu "
                "(gpu__time_duration.sum, --clock-control none; cold cache, serialised)\n")
        f.write("%-50s %6s %12s %10s %7s\n" % ("kernel", "n", "total ms", "mean us", "share"))
        for k in sorted(tot, key=lambda k: -tot[k]):
            f.write("%-50s %6d %12.2f %10.1f %6.1f%%\n" % (k, cnt[k], tot[k], 1e3 * tot[k] / cnt[k], 100 * tot[k] / all_ms))

for name in ("bench_C2_%s.json" % R, "bench_C1_%s.json" % R, "bench_C3_%s.json" % R, "bench_C4_%s.json" % R, "bench_C5_%s.json" % R, "visits_%s.txt" % R, "bench_%s.json" % R, "bench_ref_%s.json" % R, "k4_%s.txt" % R, "stages_%s.txt" % R, "pytest_gpu_%s.log" % R):
    if os.path.exists(os.path.join(G, name)):
        shutil.copy(os.path.join(G, name), os.path.join(P, name))
print(open(os.path.join(P, "traffic.json")).read())
print(open(os.path.join(P, "launches_%s_summary.txt" % R)).read()[:1800])
