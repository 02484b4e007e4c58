#!/usr/bin/env python
"""Developer timing helper (not the contract bench): python tools/quick_bench.py [config] [mode] [frames]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from webgpu_path_tracer_b200 import PathTracer, meshgen, MODE_EXACT, MODE_FAST
import bench_scenes

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
mode = {"exact": MODE_EXACT, "fast": MODE_FAST}[sys.argv[2] if len(sys.argv) > 2 else "exact"]
frames = int(sys.argv[3]) if len(sys.argv) > 3 else 8
scene, cam, W, H, kw = bench_scenes.build(cfg)
pt = PathTracer(W, H, mode=mode, **kw)
pt.write_scene(scene.buffers())
pt.write_uniforms(1, 0, cam.viewMatrix)
pt.render(1, 2); pt.synchronize()
pt.set_counters(True); pt.render(3, 1); c = pt.counters(reset=True); pt.set_counters(False)
pt.render(4, frames)
ms, launches = pt.last_timing()
spp = kw.get("num_samples", 1)
print("%s mode=%s %dx%d frames=%d: %.3f ms/frame, %.1f Msamples/s, %.1f Mrays/s, launches=%d" % (
    cfg, sys.argv[2] if len(sys.argv) > 2 else "exact", W, H, frames, ms / frames, W * H * spp * frames / ms / 1e3,
    W * H * frames * (c["rays"] / max(1, c["samples"])) * spp / ms / 1e3, launches))
print({k: v for k, v in c.items()}, "rays/sample %.3f nodes/ray %.2f tris/ray %.2f" % (
    c["rays"] / max(1, c["samples"]), c["node_visits"] / max(1, c["rays"]), c["tri_tests"] / max(1, c["rays"])))
