#!/usr/bin/env python
"""Oracle dumps committed as regression goldens: default scene, 64x64, frame 1
primary hits + radiance, and the 4-frame sum.  They pin the oracle against ITSELF over time (the
reference has no golden vectors); the independent pins are in tests/test_oracle_pins.py."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from webgpu_path_tracer_b200 import Scene, Camera
from oracle.binding import Oracle

s = Scene(); s.init_mesh_data(); s.create_meshes(); s.create_bvh()
cam = Camera(); cam.set_camera([0.5, 0, 2.5], [0.5, 0, 0], [0, 1, 0])
o = Oracle(64, 64); o.write_scene(s.buffers()); o.write_uniforms(1, 0, cam.viewMatrix)
rad, t, i, d = o.probe()
o.render(1, 4)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "oracle_dumps.npz"), c1_type=t, c1_index=i, c1_t=d,
                    c1_radiance=rad, c1_sum4=o.read_framebuffer(),
                    **{"scene_%d" % k: v for k, v in s.buffers().items()}, view=np.asarray(cam.viewMatrix, np.float32))
print("ok")
