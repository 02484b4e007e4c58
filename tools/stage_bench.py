#!/usr/bin/env python
"""Per-stage device timings of the fast path: python tools/stage_bench.py [configs, comma separated] [frames]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from webgpu_path_tracer_b200 import PathTracer, MODE_FAST
import bench_scenes

cfgs = (sys.argv[1] if len(sys.argv) > 1 else "C2,C4").split(",")
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 32
for cfg in cfgs:
    scene, cam, W, H, kw = bench_scenes.build(cfg)
    pt = PathTracer(W, H, mode=MODE_FAST, **kw)
    pt.write_scene(scene.buffers())
    pt.write_uniforms(1, 0, cam.viewMatrix)
    pt.render(1, 4)
    pt.set_stage_profiling(True)
    pt.stage_timings()
    best = None
    for rep in range(3):
        pt.render(1, frames)
        ms, launches = pt.last_timing()
        st = pt.stage_timings()
        if best is None or ms < best[0]:
            best = (ms, launches, st)
    ms, launches, st = best
    fs = pt.fast_statistics()
    print("%s %d frames: %.3f ms/frame  %.1f Msamples/s  launches=%d iters=%d fallbacks=%d | %s" % (
        cfg + " levels=%d" % fs["wide_levels"], frames, ms / frames, W * H * kw.get("num_samples", 1) * frames / ms / 1e3, launches, fs["iterations"], fs["fallback_rays"],
        "  ".join("%s %.2f ms (%d)" % (k, v[0], v[1]) for k, v in st.items())), flush=True)
    pt.close()
