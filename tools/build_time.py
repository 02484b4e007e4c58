#!/usr/bin/env python
"""Device time of the hierarchy build (K5 + K2) per configuration: python tools/build_time.py C2,C4"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from webgpu_path_tracer_b200 import PathTracer, MODE_FAST
import bench_scenes
for cfg in (sys.argv[1] if len(sys.argv) > 1 else "C2,C4").split(","):
    scene, cam, W, H, kw = bench_scenes.build(cfg)
    pt = PathTracer(W, H, mode=MODE_FAST, **kw)
    bufs = scene.buffers()
    pt.write_scene(bufs); pt.write_uniforms(1, 0, cam.viewMatrix); pt.render(1, 1)
    pt.set_stage_profiling(True); pt.stage_timings()
    for rep in range(3):
        pt.write_scene(bufs); pt.render(1, 1)
        st = pt.stage_timings()
        print(cfg, "wide_build %.3f ms (%d launches)" % st["wide_build"], pt.fast_statistics(), flush=True)
