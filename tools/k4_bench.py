#!/usr/bin/env python
"""K4 (accumulate / present) as pure streams against the measured HBM peak: python tools/k4_bench.py [W H]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from webgpu_path_tracer_b200 import PathTracer, MODE_FAST

W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (3840, 2160)
peak = 6555.0
try:
    with open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) as f:
        mp = json.load(f)
    peak = float(mp.get("hbm_gbs_burst", mp.get("hbm_gbs", peak)))
except Exception:
    pass
pt = PathTracer(W, H, mode=MODE_FAST)
for k in ("accumulate", "present_rgba8", "present_f32"):
    best = min(pt.k4_timing(k, 50) for _ in range(3))
    ms, nbytes = best
    gbs = nbytes / ms / 1e6
    print("%-14s %dx%d: %.1f us/launch  %.0f MB  %.0f GB/s  = %.2f of %.0f GB/s" % (k, W, H, ms * 1e3, nbytes / 1e6, gbs, gbs / peak, peak), flush=True)
pt.close()
