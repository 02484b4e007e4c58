#!/usr/bin/env python
"""Where the per-frame loop's time goes (default scene 900x600, renderer.js:163-215 call pattern): host wall clock of
dispatch / flush / present per frame, iterations a single frame needs (wgpt_extend_history), for several WGPT_BATCH and
WGPT_DISPATCH_ITERS settings (both read by the library at every call).  Developer aid, prints to stdout."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from webgpu_path_tracer_b200 import PathTracer, Scene, Camera, MODE_FAST

s = Scene(); s.init_mesh_data(); s.create_meshes(); s.create_bvh()
cam = Camera(); cam.set_camera([0.5, 0, 2.5], [0.5, 0, 0], [0, 1, 0])
W, H = 900, 600
pt = PathTracer(W, H, mode=MODE_FAST)
pt.write_scene(s.buffers())
f = 1
for _ in range(4):
    pt.write_uniforms(f, 0, cam.viewMatrix); pt.dispatch(); pt.present(rgba8="only"); f += 1
hist = pt.extend_history()
print("rays per iteration of the last single-frame session (first 64):", [int(x) for x in hist if x][:64])
print("iterations with rays:", int((hist > 0).sum()))
for batch, diters in ((8, 4), (4, 4), (16, 4), (32, 4), (8, 1), (8, 16), (8, 64), (16, 64), (64, 64)):
    os.environ["WGPT_BATCH"] = str(batch); os.environ["WGPT_DISPATCH_ITERS"] = str(diters)
    for _ in range(3):
        pt.write_uniforms(f, 0, cam.viewMatrix); pt.dispatch(); pt.present(rgba8="only"); f += 1
    n = 24
    td = ts = tp = 0.0
    pt.synchronize()
    T0 = time.perf_counter()
    for _ in range(n):
        pt.write_uniforms(f, 0, cam.viewMatrix)
        t0 = time.perf_counter(); pt.dispatch(); t1 = time.perf_counter()
        pt.synchronize(); t2 = time.perf_counter()
        pt.present(rgba8="only"); t3 = time.perf_counter()
        td += t1 - t0; ts += t2 - t1; tp += t3 - t2; f += 1
    tot = (time.perf_counter() - T0) / n * 1e3
    print("WGPT_BATCH %2d DISPATCH_ITERS %2d: %.2f ms/frame = dispatch %.2f + flush %.2f + present %.2f" %
          (batch, diters, tot, td / n * 1e3, ts / n * 1e3, tp / n * 1e3), flush=True)
# device time of one frame through wgpt_render (events on the stream)
for _ in range(3):
    pt.render(f, 1); f += 1
pt.render(f, 1)
print("wgpt_render(1 frame): device %.3f ms, launches %d" % tuple(pt.last_timing()[:2]))
pt.render(f, 64)
print("wgpt_render(64 frames): device %.3f ms" % pt.last_timing()[0])
