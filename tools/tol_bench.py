#!/usr/bin/env python
"""Fast mode vs the opt-in tolerance mode, device-timed: python tools/tol_bench.py [configs] [frames]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from webgpu_path_tracer_b200 import PathTracer, MODE_FAST, MODE_FAST_TOL
import bench_scenes

cfgs = (sys.argv[1] if len(sys.argv) > 1 else "C1,C2,C3,C4").split(",")
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 64
for cfg in cfgs:
    scene, cam, W, H, kw = bench_scenes.build(cfg)
    for name, mode in (("fast", MODE_FAST), ("tol", MODE_FAST_TOL)):
        pt = PathTracer(W, H, mode=mode, **kw)
        pt.write_scene(scene.buffers())
        pt.write_uniforms(1, 0, cam.viewMatrix)
        pt.render(1, 4)
        pt.set_stage_profiling(True)
        pt.stage_timings()
        best = None
        for rep in range(3):
            pt.render(1, frames)
            ms, _ = pt.last_timing()
            st = pt.stage_timings()
            if best is None or ms < best[0]:
                best = (ms, st)
        ms, st = best
        print("%s %-4s %d frames: %8.1f Msamples/s | extend %.2f ms  shade %.2f ms" % (
            cfg, name, frames, W * H * kw.get("num_samples", 1) * frames / ms / 1e3, st["extend"][0], st["shade"][0]), flush=True)
        pt.close()
