#!/usr/bin/env python
"""Overlap potential: two contexts (two streams) rendering half the frames each, concurrently, vs one context."""
import os, sys, time, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from webgpu_path_tracer_b200 import PathTracer, MODE_FAST
import bench_scenes
cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 64
nctx = int(sys.argv[3]) if len(sys.argv) > 3 else 2
scene, cam, W, H, kw = bench_scenes.build(cfg)
pts = []
for i in range(nctx):
    pt = PathTracer(W, H, mode=MODE_FAST, **kw)
    pt.write_scene(scene.buffers()); pt.write_uniforms(1, 0, cam.viewMatrix); pt.render(1, 4); pt.synchronize()
    pts.append(pt)
def run(pt, n):
    pt.render(1, n); pt.synchronize()
for rep in range(3):
    t0 = time.perf_counter(); run(pts[0], frames); t1 = time.perf_counter()
    th = [threading.Thread(target=run, args=(pt, frames // nctx)) for pt in pts]
    t2 = time.perf_counter(); [t.start() for t in th]; [t.join() for t in th]; t3 = time.perf_counter()
    print("%s %d frames: one ctx %.2f ms, %d ctx concurrently %.2f ms" % (cfg, frames, (t1 - t0) * 1e3, nctx, (t3 - t2) * 1e3), flush=True)
