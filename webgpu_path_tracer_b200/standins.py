"""Stand-in scene builders (Cornell room + one mesh, random ray sets) shared by bench_scenes.py, tools/ and the
tests: the benchmark meshes named by BASELINE.json are not in the reference mount."""
import math
import numpy as np
from .scene import Scene, Camera


def cornell_room(scene, light_emission=20.0):
    """The classic 2x2x2 room commented at scene.js:128-132 + the light quad of scene.js:123."""
    am = scene.add_material
    am("red", 0, [0.75, 0.1, 0.1], [0.75, 0.1, 0.1], [0, 0, 0], 0.05, 0.95, 0)
    am("green", 0, [0.05, 0.55, 0.05], [0.05, 0.55, 0.05], [0, 0, 0], 0.05, 0.95, 0)
    am("white", 0, [0.76, 0.70, 0.51], [0.76, 0.70, 0.51], [0, 0, 0], 0.05, 0.95, 0)
    am("glossywhite", 0, [0.76, 0.70, 0.51], [0.76, 0.70, 0.51], [0, 0, 0], 0.3, 0.1, 0)
    am("black", 0, [0.2, 0.2, 0.2], [0.2, 0.2, 0.2], [0, 0, 0], 0.05, 0.95, 0)
    d = scene.material_dict
    e = light_emission
    scene.add_quad([0.15, 0.9999, -0.25], [0.7, 0, 0], [0, 0, 0.5], am("light", 0, [0, 0, 0], [0, 0, 0], [e, e, e], 0, 0, 0))
    scene.add_quad([-1, -1, -1], [2, 0, 0], [0, 2, 0], d["black"])
    scene.add_quad([-1, -1, 1], [0, 0, -2], [0, 2, 0], d["red"])
    scene.add_quad([1, -1, -1], [0, 0, 2], [0, 2, 0], d["green"])
    scene.add_quad([-1, 1, -1], [2, 0, 0], [0, 0, 2], d["white"])
    scene.add_quad([1, -1, -1], [-2, 0, 0], [0, 0, 2], d["glossywhite"])
    scene.lights.append(scene.quads[0])
    scene.objs.append(list(scene.quads))


def mesh_in_room(mesh_data, material, transforms, spheres=()):
    """Room + one mesh.  `material` = add_material args after the name; `transforms` = list of
    ("scale"|"rotate"|"translate", args...) applied in order (lib/transform.js update())."""
    s = Scene(default=False)
    for c, r, mat in spheres:
        s.add_sphere(c, r, s.add_material("sphere", *mat))
    s.objs.append(list(s.spheres))
    cornell_room(s)
    m = s.add_mesh(mesh_data, s.add_material("meshMat", *material))
    mats = []
    for t in transforms:
        mats.append(getattr(m.transform, t[0])(*t[1:]))
    if mats:
        m.transform.update(*mats)
    s.create_meshes()
    s.create_bvh()
    cam = Camera()
    cam.set_camera([0, 0, 2.5], [0, 0, 0], [0, 1, 0])
    return s, cam


DRAGON_MAT = (0, [0.0, 0.37, 0.20], [0.0, 0.95, 0.95], [0, 0, 0], 0.4, 0.3, 2.5)     # scene.js:166
GLASS_MAT = (2, [1, 1, 1], [0, 0, 0], [0, 0, 0], 0, 0, 1.5)
MIRROR_MAT = (1, [0.95, 0.95, 0.95], [0, 0, 0], [0, 0, 0], 0, 0.05, 0)


def random_rays(n, seed, lo=-1.2, hi=1.2):
    """Rays from points in a box towards other points in it, plus axis-parallel ones (0 components)."""
    rng = np.random.default_rng(seed)
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    t = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    d = (t - o).astype(np.float32)
    k = n // 16
    d[:k, 0] = 0.0
    d[k:2 * k, 1] = 0.0
    d[2 * k:3 * k, 2] = 0.0
    d[3 * k:4 * k] /= np.maximum(np.linalg.norm(d[3 * k:4 * k], axis=1, keepdims=True), 1e-6).astype(np.float32)
    seeds = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    return np.concatenate([o, d], axis=1).astype(np.float32), seeds
