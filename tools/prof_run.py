#!/usr/bin/env python
"""Steady-state render for ncu captures (no counters, one render call): python tools/prof_run.py [config] [frames]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from webgpu_path_tracer_b200 import PathTracer, MODE_FAST
import bench_scenes

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 32
scene, cam, W, H, kw = bench_scenes.build(cfg)
pt = PathTracer(W, H, mode=MODE_FAST, **kw)
pt.write_scene(scene.buffers())
pt.write_uniforms(1, 0, cam.viewMatrix)
pt.render(1, frames)
ms, launches = pt.last_timing()
print("%s %d frames: %.3f ms/frame, %.1f Msamples/s, launches=%d" % (cfg, frames, ms / frames, W * H * kw.get("num_samples", 1) * frames / ms / 1e3, launches))
import json
json.dump({"config": cfg, "frames": frames, "n_extend": [int(x) for x in pt.extend_history()], "pool_slots": pt.fast_statistics()["pool_slots"]},
          open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", "prof_hist_%s.json" % cfg), "w"))
