"""The benchmark configurations of BASELINE.json / SURVEY.md §8(d), built from synthetic stand-in
meshes (the named .obj files are not in the reference mount: /root/reference/.MISSING_LARGE_BLOBS).
Shared by bench.py, tools/ and the full-size GPU tests."""
import math
from webgpu_path_tracer_b200 import Scene, Camera, meshgen
from webgpu_path_tracer_b200 import standins as scenes

_cache = {}


def build(name):
    """-> (scene, camera, W, H, PathTracer kwargs)"""
    if name in _cache:
        return _cache[name]
    if name == "C1":        # default scene, 512x512, HEAD flags (scene.js, index.js:39)
        s = Scene()
        s.init_mesh_data()
        s.create_meshes()
        s.create_bvh()
        cam = Camera()
        cam.set_camera([0.5, 0, 2.5], [0.5, 0, 0], [0, 1, 0])
        out = (s, cam, 512, 512, {})
    elif name == "C2":      # bunny stand-in, diffuse+glossy, 1080p (scene.js:166,198-201)
        m = meshgen.noise_sphere(69451, 2)
        s, cam = scenes.mesh_in_room(m, scenes.DRAGON_MAT, [("scale", 1.2, 1.2, 1.2), ("translate", 0.1, -1.05, 0)])
        out = (s, cam, 1920, 1080, {})
    elif name == "C3":      # dragon stand-in, glass + mixture importance sampling, 1080p (scene.js:216-220)
        m = meshgen.noise_sphere(297972, 3)
        s, cam = scenes.mesh_in_room(m, scenes.GLASS_MAT, [("scale", 1.1, 1.1, 1.1), ("rotate", math.pi / 4, [0, 1, 0]),
                                                           ("translate", 0.65, -0.64, 0)])
        out = (s, cam, 1920, 1080, {"importance_sampling": True})
    elif name == "C4":      # sponza stand-in, camera inside the atrium (scene.js:223-226)
        m = meshgen.atrium()
        s = Scene(default=False)
        s.objs.append([])
        am = s.add_material
        s.add_quad([-0.5, 0.995, -0.4], [1.0, 0, 0], [0, 0, 0.8], am("light", 0, [0, 0, 0], [0, 0, 0], [20, 20, 20], 0, 0, 0))
        s.lights.append(s.quads[0])
        s.objs.append(list(s.quads))
        mesh = s.add_mesh(m, am("stone", 0, [0.76, 0.70, 0.51], [0.76, 0.70, 0.51], [0, 0, 0], 0.05, 0.95, 0))
        mesh.transform.update(mesh.transform.scale(1, 1, 1), mesh.transform.translate(0, 0, 0))
        s.create_meshes()
        s.create_bvh()
        cam = Camera()
        cam.set_camera([-1.8, -0.3, 0.0], [1.0, -0.1, 0.1], [0, 1, 0])
        out = (s, cam, 1920, 1080, {})
    elif name == "C5":      # two meshes at 4K, NUM_SAMPLES=4 stratified (scene.js:204-207, 216-220)
        s = Scene(default=False)
        s.objs.append([])
        scenes.cornell_room(s)
        a = s.add_mesh(meshgen.noise_sphere(100000, 5), s.add_material("m0", *scenes.DRAGON_MAT))
        b = s.add_mesh(meshgen.noise_sphere(100000, 6), s.add_material("m1", *scenes.GLASS_MAT))
        a.transform.update(a.transform.scale(1, 1, 1), a.transform.translate(-0.4, -0.7, 0))
        b.transform.update(b.transform.scale(1.1, 1.1, 1.1), b.transform.rotate(math.pi / 4, [0, 1, 0]),
                           b.transform.translate(0.65, -0.64, 0))
        s.create_meshes()
        s.create_bvh()
        cam = Camera()
        cam.set_camera([0, 0, 2.5], [0, 0, 0], [0, 1, 0])
        out = (s, cam, 3840, 2160, {"num_samples": 4, "stratify": True})
    else:
        raise ValueError(name)
    _cache[name] = out
    return out
