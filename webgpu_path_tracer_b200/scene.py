"""Host-side scene description mirroring the reference's JS classes (same names and argument
meaning), producing the flat buffers the hot path consumes.

Reference: /root/reference/lib/scene.js (Scene, add_material, create_spheres/quads/meshes,
create_bvh, get_*), lib/primitives/{sphere,quad,triangle,mesh,objReader}.js, lib/transform.js,
lib/camera.js:25-33.  No JS engine exists on this box, so this restatement is how input buffers
get made; gl-matrix parity is unpinned (see glmatrix.py).
"""
import math
import os
import numpy as np

from . import glmatrix as glm
from .bvh import triangle_boxes, build_flat_bvh

LAMBERTIAN, MIRROR, GLASS, ISOTROPIC = 0, 1, 2, 3


class Transform:
    """lib/transform.js:3-88."""

    def __init__(self):
        self.modelMatrix = glm.mat4_create()
        self.invModelMatrix = glm.mat4_create()

    def getTransform(self):
        return list(self.modelMatrix) + list(self.invModelMatrix)

    def update(self, *transforms):
        if len(transforms) > 0:
            M = glm.mat4_create()
            for t in transforms:
                M = glm.mat4_mul(t, M)
            self.modelMatrix = glm.mat4_mul(M, glm.mat4_create())
            self.invModelMatrix = glm.mat4_invert(self.modelMatrix)

    def translate(self, x, y, z):
        return glm.mat4_from_translation([x, y, z])

    def scale(self, sx, sy, sz):
        return glm.mat4_from_scaling([sx, sy, sz])

    def rotate(self, theta, axis):
        return glm.mat4_from_rotation(theta, glm.f32v(axis))


class Sphere:
    """lib/primitives/sphere.js:25-30 -> 8 f32."""

    def __init__(self, center, r, global_id, local_id, material_id):
        self.type = 0
        self.global_id = global_id
        self.local_id = local_id
        self.data = [center[0], center[1], center[2], r, global_id, local_id, material_id, -1]
        self.transform = Transform()


class Quad:
    """lib/primitives/quad.js:21-36 -> 20 f32."""

    def __init__(self, Q, u, v, global_id, local_id, material_id):
        self.type = 1
        self.global_id = global_id
        self.local_id = local_id
        n = glm.vec3_cross(u, v)
        normal = glm.vec3_normalize(n)
        D = glm.vec3_dot(normal, Q)
        temp = glm.vec3_dot(n, n)
        w = glm.f32v([n[0] / temp, n[1] / temp, n[2] / temp])
        self.data = [Q[0], Q[1], Q[2], -1,
                     u[0], u[1], u[2], local_id,
                     v[0], v[1], v[2], global_id,
                     normal[0], normal[1], normal[2], D,
                     w[0], w[1], w[2], material_id]
        self.transform = Transform()


class ObjReader:
    """lib/primitives/objReader.js:10-69: triangulated `f v/vt/vn` faces only."""

    @staticmethod
    def parse(text):
        verts, norms, vidx, nidx = [], [], [], []
        for raw in text.split("\n"):
            line = raw.strip()
            if line.startswith("#"):
                continue
            if line.startswith("v "):
                verts.append([float(t) for t in line.split(" ")[1:4]])
            elif line.startswith("f "):
                toks = [t for t in line.replace("/", " ").split()][1:]
                vidx.extend(int(t) - 1 for i, t in enumerate(toks) if i % 3 == 0)
                nidx.extend(int(t) - 1 for i, t in enumerate(toks) if i % 3 == 2)
            elif line.startswith("vn "):
                norms.append([float(t) for t in line.split(" ")[1:4]])
        verts = np.asarray(verts, dtype=np.float64).reshape(-1, 3)
        norms = np.asarray(norms, dtype=np.float64).reshape(-1, 3)
        return {"vertices": verts[np.asarray(vidx, dtype=np.int64)].astype(np.float32).reshape(-1),
                "normals": norms[np.asarray(nidx, dtype=np.int64)].astype(np.float32).reshape(-1)}

    @staticmethod
    def load_model(path):
        with open(path, "r") as f:
            return ObjReader.parse(f.read())


class Mesh:
    """lib/primitives/mesh.js:7-63 + lib/primitives/triangle.js:42-52 (24 f32 per triangle)."""

    def __init__(self, data, offset, id, mesh_id, local_id, material_id):
        v = np.asarray(data["vertices"], dtype=np.float32).reshape(-1, 3, 3)
        nrm = np.asarray(data["normals"], dtype=np.float32).reshape(-1, 3, 3)
        n = v.shape[0]
        self.numTriangle = n
        tri = np.full((n, 24), -1.0, dtype=np.float32)
        tri[:, 0:3], tri[:, 4:7], tri[:, 8:11] = v[:, 0], v[:, 1], v[:, 2]
        tri[:, 12:15], tri[:, 16:19], tri[:, 20:23] = nrm[:, 0], nrm[:, 1], nrm[:, 2]
        tri[:, 19] = local_id + np.arange(n)
        tri[:, 23] = mesh_id
        self.triangle_data = tri
        self.object_vertices = v.astype(np.float64)
        self.mesh = [n, offset, id, material_id]
        self.transform = Transform()
        self.box_min = self.box_max = None

    def calc_bbox(self, transform):
        """triangle.js:27-39: world-space f32 vertices -> padded boxes (used for the BVH only)."""
        m = transform.modelMatrix
        A = glm.transform_mat4_many(self.object_vertices[:, 0], m)
        B = glm.transform_mat4_many(self.object_vertices[:, 1], m)
        C = glm.transform_mat4_many(self.object_vertices[:, 2], m)
        self.box_min, self.box_max = triangle_boxes(A, B, C)


class Camera:
    """lib/camera.js:25-33 (the orbit/mouse handling is DOM code and out of scope)."""

    def __init__(self):
        self.viewMatrix = glm.mat4_create()
        self.MOVING = 0
        self.keyPress = 0

    def set_camera(self, eye, center, up):
        self.eye, self.center, self.up = list(eye), list(center), list(up)
        self.viewMatrix = glm.mat4_target_to(eye, center, up)


ASSET_DIR_ENV = "WGPT_ASSET_DIR"


def _packaged_cube():
    """Parsed arrays of the reference's assets/cube.obj (12 triangles), stored by
    tools/make_golden_meshes.py in data/cube.npz because /root/reference does not travel to the
    GPU box.  It is the mesh of the default scene (scene.js:169, 289)."""
    d = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "cube.npz"))
    return {"vertices": d["vertices"], "normals": d["normals"]}


class Scene:
    """lib/scene.js:9-311.  `Scene()` reproduces the hard-coded default scene (19 spheres, 8 quads,
    31 materials, one cube mesh); `Scene(default=False)` starts empty so other scenes can be
    described with the same methods."""

    def __init__(self, default=True, mesh_data=None):
        self.mats = []
        self.material_id = 0
        self.material_dict = {}
        self.global_id = 0
        self.sphere_id = 0
        self.quad_id = 0
        self.triangle_id = 0
        self.mesh_id = 0
        self.triangle_offset = 0
        self.spheres, self.quads, self.meshes, self.objs, self.lights = [], [], [], [], []
        self.mesh_data = dict(mesh_data or {})
        self.bvh_array = np.zeros((0, 12), np.float32)
        self.triangle_array = np.zeros((0, 24), np.float32)
        self.default = default
        if default:
            self.create_spheres()
            self.create_quads()

    # ---- scene.js:261-273 ----
    def add_material(self, name, material_type, color, specularColor, emissionColor, percentSpecular,
                     roughness, eta=float("nan")):
        """`eta` omitted == JS `undefined` -> NaN in the Float32Array (scene.js:119)."""
        self.material_dict[name] = self.material_id
        self.mats.append([color[0], color[1], color[2], -1,
                          specularColor[0], specularColor[1], specularColor[2], -1,
                          emissionColor[0], emissionColor[1], emissionColor[2], percentSpecular,
                          roughness, eta, material_type, -1])
        self.material_id += 1
        return self.material_id - 1

    def add_sphere(self, center, r, material_id):
        s = Sphere(center, r, self.global_id, self.sphere_id, material_id)
        self.global_id += 1
        self.sphere_id += 1
        self.spheres.append(s)
        return s

    def add_quad(self, Q, u, v, material_id):
        q = Quad(Q, u, v, self.global_id, self.quad_id, material_id)
        self.global_id += 1
        self.quad_id += 1
        self.quads.append(q)
        return q

    def add_mesh(self, data, material_id):
        """One `new Mesh(...)` + the triangle_id/offset bookkeeping of scene.js:168-174."""
        m = Mesh(data, self.triangle_offset, self.global_id, self.mesh_id, self.triangle_id, material_id)
        self.global_id += 1
        self.mesh_id += 1
        self.triangle_id += m.numTriangle
        self.triangle_offset += m.numTriangle
        self.meshes.append(m)
        return m

    # ---- scene.js:36-103 ----
    def create_spheres(self):
        self.add_material("default", 0, [1, 0, 0], [0, 0, 0], [0, 0, 0], 0, 0, 0)
        temp = [0.94, 0.70, 0.75]
        green, blue = [0.56, 0.93, 0.56], [0.52, 0.8, 0.92]

        def fog(c, r, col, dens):
            self.add_sphere(c, r, self.add_material("fog", 3, col, [0, 0, 0], [0, 0, 0], 0.00001, dens, 0))

        def glass(c, r, eta):
            self.add_sphere(c, r, self.add_material("gg4t", 2, [1, 1, 1], [0, 0, 0], [0, 0, 0], 0, 0, eta))

        for c, r in (([-0.3, -0.65, 0.3], 0.35), ([-0.3, -0.05, 0.3], 0.25), ([-0.3, 0.3, 0.3], 0.1), ([-0.3, 0.45, 0.3], 0.05)):
            fog(c, r, green, -1 / 4)
            glass(c, r, 1.5)
        fog([0.5, -0.65, -0.2], 0.35, blue, -1 / 7)
        glass([0.5, -0.65, -0.2], 0.35, 1)
        glass([0.5, 0.1, 0.2], 0.2, 1.5)
        for c, r in (([1.3, -0.65, 0.3], 0.35), ([1.3, -0.05, 0.3], 0.25), ([1.3, 0.3, 0.3], 0.1), ([1.3, 0.45, 0.3], 0.05)):
            fog(c, r, temp, -1 / 10)
            glass(c, r, 1)
        self.objs.append(list(self.spheres))

    # ---- scene.js:105-162 ----
    def create_quads(self):
        am = self.add_material
        am("red", 0, [0.75, 0.1, 0.1], [0.75, 0.1, 0.1], [0, 0, 0], 0.05, 0.95, 0)
        am("green", 0, [0.05, 0.55, 0.05], [0.05, 0.55, 0.05], [0, 0, 0], 0.05, 0.95, 0)
        am("blue", 0, [0.05, 0.05, 0.55], [0.05, 0.05, 0.55], [0, 0, 0], 0.05, 0.95, 0)
        am("white", 0, [0.76, 0.70, 0.51], [0.76, 0.70, 0.51], [0, 0, 0], 0.05, 0.95, 0)
        am("glossywhite", 0, [0.76, 0.70, 0.51], [0.76, 0.70, 0.51], [0, 0, 0], 0.3, 0.1, 0)
        am("black", 0, [0.2, 0.2, 0.2], [0.2, 0.2, 0.2], [0, 0, 0], 0.05, 0.95, 0)
        am("glass", 1, [0.95, 0.95, 0.95], [0, 0, 0], [0, 0, 0], 0, 0, 0)
        d = self.material_dict
        self.add_quad([-1, 1, -1], [3, 0, 0], [0, 0, 2], am("tWall", 0, [0, 0, 0], [0, 0, 0], [2, 2, 2], 0, 0))
        self.add_quad([-1, -1, -1], [3, 0, 0], [0, 2, 0], d["black"])
        self.add_quad([-1, -1, 1], [0, 0, -2], [0, 2, 0], d["red"])
        self.add_quad([2, -1, -1], [0, 0, 2], [0, 2, 0], d["green"])
        self.add_quad([-1, 1, -1], [3, 0, 0], [0, 0, 2], d["white"])
        self.add_quad([2, -1, -1], [-3, 0, 0], [0, 0, 2], d["glossywhite"])
        self.add_quad([100, -1, -100], [-200, 0, 0], [0, 0, 200], d["white"])
        self.add_quad([2, -1, 1], [-3, 0, 0], [0, 2, 0], am("fWall", 0, [0.15, 0.15, 0.15], [0, 0, 0], [0, 0, 0], 0, 0, 0))
        self.lights.append(self.quads[0])
        self.objs.append(list(self.quads))

    # ---- scene.js:284-302 ----
    def init_mesh_data(self, asset_dir=None):
        asset_dir = asset_dir or os.environ.get(ASSET_DIR_ENV)
        if "cube" not in self.mesh_data:
            path = os.path.join(asset_dir, "cube.obj") if asset_dir else None
            self.mesh_data["cube"] = ObjReader.load_model(path) if path and os.path.exists(path) else _packaged_cube()

    # ---- scene.js:164-251 ----
    def create_meshes(self):
        if self.default:
            self.add_material("dragonMat", 0, [0.0, 0.37, 0.20], [0.0, 0.95, 0.95], [0, 0, 0], 0.4, 0.3, 2.5)
            m = self.add_mesh(self.mesh_data["cube"],
                              self.add_material("glassBox", 0, [0.95, 0.95, 0.95], [0, 0, 0], [0, 0, 0], 0, 0, 2.5))
            m.transform.update(m.transform.rotate(math.pi / 10, [0, 1, 0]))
        for m in self.meshes:
            m.calc_bbox(m.transform)
        if self.meshes:
            self.triangle_array = np.concatenate([m.triangle_data for m in self.meshes], axis=0)
        self.objs.append(list(self.meshes))

    # ---- scene.js:253-259 + lib/BVH/bvhBuilder.js:6-35 ----
    def create_bvh(self):
        if not self.meshes:
            return
        bmin = np.concatenate([m.box_min for m in self.meshes], axis=0)
        bmax = np.concatenate([m.box_max for m in self.meshes], axis=0)
        self.bvh_array, order = build_flat_bvh(bmin, bmax)
        self.triangle_array = self.triangle_array[order]

    # ---- scene.js:275-282, 304-310 ----
    def get_transforms(self):
        out = []
        for group in self.objs:
            for o in group:
                out.extend(o.transform.getTransform())
        return np.asarray(out, dtype=np.float32)

    def get_bvh(self):
        return np.ascontiguousarray(self.bvh_array, dtype=np.float32).reshape(-1)

    def get_triangles(self):
        return np.ascontiguousarray(self.triangle_array, dtype=np.float32).reshape(-1)

    def get_meshes(self):
        return np.asarray([m.mesh for m in self.meshes], dtype=np.int32).reshape(-1)

    def get_materials(self):
        return np.asarray(self.mats, dtype=np.float32).reshape(-1)

    def get_spheres(self):
        return np.asarray([s.data for s in self.spheres], dtype=np.float32).reshape(-1)

    def get_quads(self):
        return np.asarray([q.data for q in self.quads], dtype=np.float32).reshape(-1)

    def buffers(self):
        """binding number -> flat array, as renderer.js:80-87,109-123 binds them (0 and 3 excluded)."""
        return {1: self.get_spheres(), 2: self.get_quads(), 5: self.get_triangles(), 6: self.get_meshes(),
                7: self.get_transforms(), 8: self.get_materials(), 9: self.get_bvh()}
