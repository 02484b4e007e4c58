#!/usr/bin/env python
"""Wide-BVH visits per ray of the fast path (counters on): python tools/visit_stats.py [configs] [frames]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from webgpu_path_tracer_b200 import PathTracer, MODE_FAST
import bench_scenes

cfgs = (sys.argv[1] if len(sys.argv) > 1 else "C2,C4").split(",")
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 4
for cfg in cfgs:
    scene, cam, W, H, kw = bench_scenes.build(cfg)
    pt = PathTracer(W, H, mode=MODE_FAST, **kw)
    pt.write_scene(scene.buffers())
    pt.write_uniforms(1, 0, cam.viewMatrix)
    pt.set_counters(True)
    pt.counters(reset=True)
    pt.render(1, frames)
    c = pt.counters()
    rays = max(1, c["rays"])
    print("%s: rays/sample %.3f  wide-node visits/ray %.2f  triangle tests/ray %.2f  (%d frames)" % (
        cfg, rays / max(1, c["samples"]), c["node_visits"] / rays, c["tri_tests"] / rays, frames), flush=True)
    pt.close()
