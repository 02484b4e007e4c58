#!/usr/bin/env python
"""Generate mesh fixtures from the reference's assets (run in the authoring container only,
where /root/reference exists; the GPU box has no /root/reference).

Outputs (committed):
  webgpu_path_tracer_b200/data/cube.npz     parsed assets/cube.obj  (the default scene's mesh)
  tests/golden/meshes.npz                   parsed cube, icosphere, cone, hole, monkey_968,
                                            monkey_smooth_15744 (vertices/normals, 9 floats per triangle,
                                            exactly what ObjReader.load_model returns, objReader.js:10-69)
"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from webgpu_path_tracer_b200.scene import ObjReader  # noqa: E402

ASSETS = "/root/reference/assets"


def main():
    cube = ObjReader.load_model(os.path.join(ASSETS, "cube.obj"))
    np.savez_compressed(os.path.join(ROOT, "webgpu_path_tracer_b200", "data", "cube.npz"),
                        vertices=cube["vertices"], normals=cube["normals"])
    out = {}
    for name in ("cube", "icosphere", "cone", "hole", "monkey_968", "monkey_smooth_15744"):
        m = ObjReader.load_model(os.path.join(ASSETS, name + ".obj"))
        out[name + "_vertices"] = m["vertices"]
        out[name + "_normals"] = m["normals"]
        print(name, m["vertices"].size // 9, "triangles")
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "meshes.npz"), **out)


if __name__ == "__main__":
    main()
