"""A SECOND WITNESS for the oracle -- TEST INFRASTRUCTURE.

An independent float32 transliteration of the reference's shaders, written from the WGSL files themselves
(/root/reference/shaders/*.wgsl, cited per function) and NOT from oracle/*.cpp or the CUDA sources: numpy float32
scalars, one operation per WGSL operator, in WGSL's own shape (hitRec is a mutable global that every accepted hit
overwrites, scatterRec / doSpecular / the ONB vectors are globals, the BVH loop is the one of hitRay.wgsl).  What WGSL
leaves loose is taken from SURVEY.md §8(c) "pinned semantics" (left-to-right f32 sums, no contraction, mix = a(1-t)+bt,
normalize = v / length ...); sin / cos / acos / log / pow go through the shared pinned-math header's host hook
(include/wgpt_pinned_math.h via oracle.binding.pinned_math), the one piece both sides must share by construction.
tests/test_second_witness.py compares it with the oracle bit for bit: a misreading of the WGSL shared by the oracle and
the kernels (both written from one reading) would show up here.
"""
import numpy as np

from oracle.binding import pinned_math

f32 = np.float32
u32 = np.uint32
np.seterr(all="ignore")

PI = f32(3.1415926535897932385)
TWO_PI = f32(2 * 3.1415926535897932385)            # `2 * PI` folds in abstract float, then becomes f32
MIN_FLOAT = f32(0.0001)
MAX_FLOAT = f32(999999999.999)
MAX_FLOAT_P1 = f32(999999999.999 + 1)
LAMBERTIAN, MIRROR, GLASS, ISOTROPIC = 0, 1, 2, 3


def pm(fn, x, y=None):
    return f32(pinned_math(fn, np.array([x], np.float32), None if y is None else np.array([y], np.float32))[0])


# ---- vec3 as tuples of float32, WGSL operators spelled out ----
def V(x, y, z):
    return (f32(x), f32(y), f32(z))


def add(a, b): return (a[0] + b[0], a[1] + b[1], a[2] + b[2])
def sub(a, b): return (a[0] - b[0], a[1] - b[1], a[2] - b[2])
def mul(a, b): return (a[0] * b[0], a[1] * b[1], a[2] * b[2])
def scale(a, s): return (a[0] * s, a[1] * s, a[2] * s)
def divs(a, s): return (a[0] / s, a[1] / s, a[2] / s)
def neg(a): return (-a[0], -a[1], -a[2])
def dot(a, b): return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]
def cross(a, b): return (a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0])
def length(a): return np.sqrt(dot(a, a))
def normalize(a): return divs(a, length(a))
def mix(a, b, t): return add(scale(a, f32(1) - t), scale(b, t))
def reflect(e1, e2): return sub(e1, scale(e2, f32(2) * dot(e2, e1)))


def refract(e1, e2, e3):
    d = dot(e2, e1)
    k = f32(1) - e3 * e3 * (f32(1) - d * d)
    if k < 0:
        return V(0, 0, 0)
    return sub(scale(e1, e3), scale(e2, e3 * d + np.sqrt(k)))


def fmin(a, b):
    if np.isnan(a):
        return b
    if np.isnan(b):
        return a
    if a == b:                                    # min(-0, +0) = -0
        return f32(np.array([np.array([a], np.float32).view(np.uint32)[0] | np.array([b], np.float32).view(np.uint32)[0]], np.uint32).view(np.float32)[0])
    return a if a < b else b


def fmax(a, b):
    if np.isnan(a):
        return b
    if np.isnan(b):
        return a
    if a == b:                                    # max(-0, +0) = +0
        return f32(np.array([np.array([a], np.float32).view(np.uint32)[0] & np.array([b], np.float32).view(np.uint32)[0]], np.uint32).view(np.float32)[0])
    return a if a > b else b


def mat_vec4(m, v):
    """mat4x4f * vec4f, m column-major 16 floats: ((c0 x + c1 y) + c2 z) + c3 w per row."""
    return tuple(((m[0 + r] * v[0] + m[4 + r] * v[1]) + m[8 + r] * v[2]) + m[12 + r] * v[3] for r in range(4))


def matT_vec4(m, v):
    """transpose(m) * v"""
    return tuple(((m[4 * r] * v[0] + m[4 * r + 1] * v[1]) + m[4 * r + 2] * v[2]) + m[4 * r + 3] * v[3] for r in range(4))


class HitRecord:
    def __init__(self):
        self.p = V(0, 0, 0)
        self.t = f32(0)
        self.normal = V(0, 0, 0)
        self.front_face = False
        self.material = None           # index into materials (header.wgsl:119-125 copies the struct; an index is the same data)
        self.prim = (-1, -1)           # (type, index) of the accepted primitive: the parity probe, not part of the WGSL


class Witness:
    """One invocation of computeFrameBuffer at a time (main.wgsl:1-28), state in WGSL's private globals."""

    def __init__(self, W, H, buffers, view_matrix, num_samples=1, max_bounces=100, stratify=False, importance_sampling=False,
                 stack_size=20):
        self.W, self.H = f32(W), f32(H)
        b = {k: np.ascontiguousarray(v, np.float32).reshape(-1) for k, v in buffers.items()}
        self.spheres = b[1].reshape(-1, 8)            # center.xyz r | global_id local_id material_id pad   (header.wgsl:68-74)
        self.quads = b[2].reshape(-1, 20)             # Q pad | u local_id | v global_id | normal D | w material_id (:76-86)
        self.tris = b[5].reshape(-1, 24)              # A p B p C p nA p nB local_id nC mesh_id            (:88-98)
        self.meshes = np.ascontiguousarray(buffers[6]).view(np.int32).reshape(-1, 4)   # num_triangles offset global_id material_id
        self.transforms = b[7].reshape(-1, 32)        # modelMatrix | invModelMatrix                         (:63-66)
        self.materials = b[8].reshape(-1, 16)         # color p | specularColor p | emissionColor specularStrength | roughness eta type p
        self.bvh = b[9].reshape(-1, 12)               # min right_offset | max prim_type | prim_id prim_count skip_link axis
        self.view = [f32(x) for x in np.asarray(view_matrix, np.float32).reshape(-1)]
        self.NUM_SAMPLES, self.MAX_BOUNCES, self.STRATIFY = num_samples, max_bounces, stratify
        self.IMPORTANCE_SAMPLING, self.STACK_SIZE = importance_sampling, stack_size
        self.ray_tmin = f32(0.000001)

    # ---------------- common.wgsl ----------------
    def rand2D(self):                                  # common.wgsl:7-12
        self.randState = (self.randState * 747796405 + 2891336453) & 0xffffffff
        s = self.randState
        word = (((s >> ((s >> 28) + 4)) ^ s) * 277803737) & 0xffffffff
        return f32((word >> 22) ^ word) / f32(4294967295)

    def mat(self, i):
        return int(i)

    def hit_sphere(self, i, tmin, tmax, o, d):         # common.wgsl:29-73
        s = self.spheres[i]
        center, r = V(*s[0:3]), f32(s[3])
        oc = sub(o, center)
        a = dot(d, d)
        half_b = dot(d, oc)
        c = dot(oc, oc) - r * r
        disc = half_b * half_b - a * c
        if disc < 0:
            return False
        sqrtd = np.sqrt(disc)
        root = (-half_b - sqrtd) / a
        if root <= tmin or root >= tmax:
            root = (-half_b + sqrtd) / a
            if root <= tmin or root >= tmax:
                return False
        h = self.hitRec
        h.t = root
        h.p = add(o, scale(d, root))
        h.normal = normalize(divs(sub(h.p, center), r))
        h.front_face = bool(dot(d, h.normal) < 0)
        if not h.front_face:
            h.normal = neg(h.normal)
        h.material = self.mat(s[6])
        h.prim = (0, i)
        return True

    def hit_sphere_local(self, i, tmin, tmax, o, d):   # common.wgsl:75-100
        s = self.spheres[i]
        center, r = V(*s[0:3]), f32(s[3])
        oc = sub(o, center)
        a = dot(d, d)
        half_b = dot(d, oc)
        c = dot(oc, oc) - r * r
        disc = half_b * half_b - a * c
        if disc < 0:
            return MAX_FLOAT_P1
        sqrtd = np.sqrt(disc)
        root = (-half_b - sqrtd) / a
        if root <= tmin or root >= tmax:
            root = (-half_b + sqrtd) / a
            if root <= tmin or root >= tmax:
                return MAX_FLOAT_P1
        return root

    def hit_volume(self, i, tmin, tmax, o, d):         # common.wgsl:102-146
        s = self.spheres[i]
        rec1 = self.hit_sphere_local(i, -MAX_FLOAT, MAX_FLOAT, o, d)
        if rec1 == MAX_FLOAT_P1:
            return False
        rec2 = self.hit_sphere_local(i, rec1 + f32(0.0001), MAX_FLOAT, o, d)
        if rec2 == MAX_FLOAT_P1:
            return False
        if rec1 < tmin:
            rec1 = tmin
        if rec2 > tmax:
            rec2 = tmax
        if rec1 >= rec2:
            return False
        if rec1 < 0:
            rec1 = f32(0)
        h = self.hitRec
        h.material = self.mat(s[6])                    # stored BEFORE the accept test (:130)
        ray_length = length(d)
        dist_inside = (rec2 - rec1) * ray_length
        hit_dist = f32(self.materials[h.material][12]) * pm("log", self.rand2D())
        if hit_dist > dist_inside:
            return False
        h.t = rec1 + (hit_dist / ray_length)
        h.p = add(o, scale(d, h.t))
        h.normal = normalize(sub(h.p, V(*s[0:3])))
        h.front_face = True
        h.prim = (0, i)
        return True

    def hit_quad(self, i, tmin, tmax, o, d):           # common.wgsl:148-187
        q = self.quads[i]
        Q, qu, qv, qn, D, qw = V(*q[0:3]), V(*q[4:7]), V(*q[8:11]), V(*q[12:15]), f32(q[15]), V(*q[16:19])
        if dot(d, qn) > 0:
            return False
        denom = dot(qn, d)
        if abs(denom) < f32(1e-8):
            return False
        t = (D - dot(qn, o)) / denom
        if t <= tmin or t >= tmax:
            return False
        inter = add(o, scale(d, t))
        planar = sub(inter, Q)
        alpha = dot(qw, cross(planar, qv))
        beta = dot(qw, cross(qu, planar))
        if alpha < 0 or 1 < alpha or beta < 0 or 1 < beta:
            return False
        h = self.hitRec
        h.t = t
        h.p = inter
        h.normal = normalize(qn)
        h.front_face = bool(dot(d, h.normal) < 0)
        if not h.front_face:
            h.normal = neg(h.normal)
        h.material = self.mat(q[19])
        h.prim = (1, i)
        return True

    def hit_triangle(self, ti, tmin, tmax, o, d):      # common.wgsl:191-242
        tri = self.tris[ti]
        mesh = self.meshes[int(tri[23])]
        tr = self.transforms[mesh[2]]
        inv = [f32(x) for x in tr[16:32]]
        ro = mat_vec4(inv, (o[0], o[1], o[2], f32(1)))[:3]
        rd = mat_vec4(inv, (d[0], d[1], d[2], f32(0)))[:3]
        A, B, C = V(*tri[0:3]), V(*tri[4:7]), V(*tri[8:11])
        AB, AC = sub(B, A), sub(C, A)
        normal = cross(AB, AC)
        det = -dot(rd, normal)
        if abs(det) < tmin:
            return False
        ao = sub(ro, A)
        dao = cross(ao, rd)
        invDet = f32(1) / det
        dst = dot(ao, normal) * invDet
        u = dot(AC, dao) * invDet
        v = -dot(AB, dao) * invDet
        w = f32(1) - u - v
        if dst < tmin or dst > tmax or u < tmin or v < tmin or w < tmin:
            return False
        h = self.hitRec
        h.t = dst
        h.p = add(o, scale(d, dst))
        nA, nB, nC = V(*tri[12:15]), V(*tri[16:19]), V(*tri[20:23])
        n = add(add(scale(nA, w), scale(nB, u)), scale(nC, v))
        h.normal = normalize(matT_vec4(inv, (n[0], n[1], n[2], f32(0)))[:3])
        h.front_face = bool(dot(d, h.normal) < 0)
        if not h.front_face:
            h.normal = neg(h.normal)
        h.material = int(mesh[3])
        h.prim = (2, int(tri[19]))                     # Triangle.local_id
        return True

    def hit_aabb(self, node, tmin, tmax, o, invDir):   # common.wgsl:245-256
        mn, mx = V(*node[0:3]), V(*node[4:7])
        t0 = mul(sub(mn, o), invDir)
        t1 = mul(sub(mx, o), invDir)
        sm = (fmin(t0[0], t1[0]), fmin(t0[1], t1[1]), fmin(t0[2], t1[2]))
        bg = (fmax(t0[0], t1[0]), fmax(t0[1], t1[1]), fmax(t0[2], t1[2]))
        t_min = fmax(tmin, fmax(sm[0], fmax(sm[1], sm[2])))
        t_max = fmin(tmax, fmin(bg[0], fmin(bg[1], bg[2])))
        return bool(t_max > t_min)

    def get_lights(self):                              # common.wgsl:258-269
        self.lights = np.zeros(20, np.float32)
        for i in range(len(self.quads)):
            if self.materials[int(self.quads[i][19])][8] > 0.0:
                self.lights = self.quads[i]
                break

    # ---------------- hitRay.wgsl:1-113 ----------------
    def hitScene(self, o, d):
        closest = MAX_FLOAT
        hit_anything = False
        for i in range(len(self.spheres)):
            medium = self.materials[int(self.spheres[i][6])][14]
            if medium < ISOTROPIC:
                if self.hit_sphere(i, self.ray_tmin, closest, o, d):
                    hit_anything = True
                    closest = self.hitRec.t
            else:
                if self.hit_volume(i, self.ray_tmin, closest, o, d):
                    hit_anything = True
                    closest = self.hitRec.t
        for i in range(len(self.quads)):
            if self.hit_quad(i, self.ray_tmin, closest, o, d):
                hit_anything = True
                closest = self.hitRec.t
        if len(self.bvh) == 0:                         # (arrayLength 0: robust buffer access reads zeros -> prim_type 0, box never hit)
            return hit_anything
        invDir = (f32(1) / d[0], f32(1) / d[1], f32(1) / d[2])
        toVisit, cur = 0, 0
        stack = [0] * max(self.STACK_SIZE, 1)
        while True:
            node = self.bvh[cur]
            if self.hit_aabb(node, self.ray_tmin, closest, o, invDir):
                if int(node[7]) == 2:
                    start, count = int(node[8]), int(node[9])
                    for j in range(count):
                        if self.hit_triangle(start + j, self.ray_tmin, closest, o, d):
                            hit_anything = True
                            closest = self.hitRec.t
                    if toVisit == 0:
                        break
                    toVisit -= 1
                    cur = stack[toVisit]
                else:
                    if d[int(node[11])] < 0:
                        stack[toVisit] = cur + 1
                        toVisit += 1
                        cur = int(node[3])
                    else:
                        stack[toVisit] = int(node[3])
                        toVisit += 1
                        cur += 1
            else:
                if toVisit == 0:
                    break
                toVisit -= 1
                cur = stack[toVisit]
            if toVisit >= self.STACK_SIZE:
                break
        return hit_anything

    # ---------------- importanceSampling.wgsl ----------------
    def reflectance(self, cosine, ref_idx):            # :1-5
        r0 = (f32(1) - ref_idx) / (f32(1) + ref_idx)
        r0 = r0 * r0
        return r0 + (f32(1) - r0) * pm("pow", f32(1) - cosine, f32(5))

    def uniform_random_in_unit_sphere(self):           # :7-16
        phi = self.rand2D() * f32(2.0) * PI
        theta = pm("acos", f32(2.0) * self.rand2D() - f32(1.0))
        x = pm("sin", theta) * pm("cos", phi)
        y = pm("sin", theta) * pm("sin", phi)
        z = pm("cos", theta)
        return normalize((x, y, z))

    def cosine_sampling_wrt_Z(self):                   # :35-45
        r1 = self.rand2D()
        r2 = self.rand2D()
        phi = TWO_PI * r1
        x = pm("cos", phi) * np.sqrt(r2)
        y = pm("sin", phi) * np.sqrt(r2)
        z = np.sqrt(f32(1) - r2)
        return (x, y, z)

    def onb_build_from_w(self, w):                     # :60-67
        self.unit_w = normalize(w)
        a = V(0, 1, 0) if abs(self.unit_w[0]) > f32(0.9) else V(1, 0, 0)
        self.v = normalize(cross(self.unit_w, a))
        self.u = cross(self.unit_w, self.v)

    def onb_get_local(self, a):                        # :69-71
        return add(add(scale(self.u, a[0]), scale(self.v, a[1])), scale(self.unit_w, a[2]))

    def onb_lambertian_scattering_pdf(self, sd):       # :73-76
        cosine_theta = dot(normalize(sd), self.unit_w)
        return fmax(f32(0), cosine_theta / PI)

    def get_random_on_quad(self, q, origin):           # :78-81 (u coefficient drawn first: left-to-right evaluation)
        ru = self.rand2D()
        pu = add(V(*q[0:3]), scale(V(*q[4:7]), ru))
        rv = self.rand2D()
        p = add(pu, scale(V(*q[8:11]), rv))
        return origin, normalize(sub(p, origin))

    def light_pdf(self, o, d, q):                      # :88-125
        qn = V(*q[12:15])
        if dot(d, qn) > 0:
            return MIN_FLOAT
        denom = dot(qn, d)
        if abs(denom) < f32(1e-8):
            return MIN_FLOAT
        t = (f32(q[15]) - dot(qn, o)) / denom
        if t <= f32(0.001) or t >= MAX_FLOAT:
            return MIN_FLOAT
        inter = add(o, scale(d, t))
        planar = sub(inter, V(*q[0:3]))
        alpha = dot(V(*q[16:19]), cross(planar, V(*q[8:11])))
        beta = dot(V(*q[16:19]), cross(V(*q[4:7]), planar))
        if alpha < 0 or 1 < alpha or beta < 0 or 1 < beta:
            return MIN_FLOAT
        hitNormal = qn
        if not (dot(d, qn) < 0):
            hitNormal = neg(hitNormal)
        distance_squared = t * t * length(d) * length(d)
        cosine = abs(dot(d, hitNormal) / length(d))
        L = self.lights
        return distance_squared / (cosine * length(cross(V(*L[4:7]), V(*L[8:11]))))

    # ---------------- scatterRay.wgsl:2-95 ----------------
    def material_scatter(self, ro, rd):
        h = self.hitRec
        m = self.materials[h.material]
        color, spec, emis = V(*m[0:3]), V(*m[4:7]), V(*m[8:11])
        specularStrength, roughness, eta, mtype = f32(m[11]), f32(m[12]), f32(m[13]), m[14]
        so, sd = V(0, 0, 0), V(0, 0, 0)
        self.doSpecular = f32(0)
        if mtype == LAMBERTIAN:
            self.onb_build_from_w(h.normal)
            diffuse_dir = self.cosine_sampling_wrt_Z()
            diffuse_dir = normalize(self.onb_get_local(diffuse_dir))
            self.doSpecular = f32(1.0) if self.rand2D() < specularStrength else f32(0.0)
            specular_dir = reflect(rd, h.normal)
            specular_dir = normalize(mix(specular_dir, diffuse_dir, roughness))
            so, sd = h.p, normalize(mix(diffuse_dir, specular_dir, self.doSpecular))
            self.skip_pdf = False
            if self.doSpecular == 1.0:
                self.skip_pdf = True
                self.skip_pdf_ray = (so, sd)
        elif mtype == MIRROR:
            reflected = reflect(rd, h.normal)
            so, sd = h.p, normalize(add(reflected, scale(self.uniform_random_in_unit_sphere(), roughness)))
            self.skip_pdf = True
            self.skip_pdf_ray = (so, sd)
        elif mtype == GLASS:
            ir = eta
            if h.front_face:
                ir = f32(1.0) / ir
            unit_direction = normalize(rd)
            cos_theta = fmin(dot(neg(unit_direction), h.normal), f32(1.0))
            sin_theta = np.sqrt(f32(1) - cos_theta * cos_theta)
            if ir * sin_theta > f32(1.0) or self.reflectance(cos_theta, ir) > self.rand2D():      # `||` short-circuits
                direction = reflect(unit_direction, h.normal)
            else:
                direction = refract(unit_direction, h.normal, ir)
            # near_zero(): `abs(v) < 0` is never true (common.wgsl:26-28)
            so, sd = h.p, normalize(direction)
            self.skip_pdf = True
            self.skip_pdf_ray = (so, sd)
        elif mtype == ISOTROPIC:
            g = specularStrength
            cos_hg = (f32(1) + g * g - pm("pow", (f32(1) - g * g) / (f32(1) - g + f32(2) * g * self.rand2D()), f32(2))) / (f32(2) * g)
            sin_hg = np.sqrt(f32(1) - cos_hg * cos_hg)
            phi = TWO_PI * self.rand2D()
            hg_dir = (sin_hg * pm("cos", phi), sin_hg * pm("sin", phi), cos_hg)
            self.onb_build_from_w(rd)
            so, sd = h.p, normalize(self.onb_get_local(hg_dir))
            self.skip_pdf = True
            self.skip_pdf_ray = (so, sd)
        return so, sd

    # ---------------- traceRay.wgsl:3-83 ----------------
    def ray_color(self, o, d):
        acc = V(0, 0, 0)
        thr = V(1, 1, 1)
        bg = V(0, 1, 1)
        for i in range(self.MAX_BOUNCES):
            hit = self.hitScene(o, d)
            if self.primary is None:
                self.primary = (self.hitRec.prim[0], self.hitRec.prim[1], self.hitRec.t) if hit else (-1, -1, f32(0))
            if not hit:
                acc = add(acc, mul(bg, thr))
                break
            h = self.hitRec
            m = self.materials[h.material]
            color, spec = V(*m[0:3]), V(*m[4:7])
            emission = V(*m[8:11])
            if not h.front_face:
                emission = V(0, 0, 0)
            if self.IMPORTANCE_SAMPLING:
                s_o, s_d = self.material_scatter(o, d)
                if self.skip_pdf:
                    acc = add(acc, mul(emission, thr))
                    thr = mul(thr, mix(color, spec, self.doSpecular))
                    o, d = self.skip_pdf_ray
                    continue
                l_o, l_d = self.get_random_on_quad(self.lights, h.p)
                sc_o, sc_d = l_o, l_d
                rnd = self.rand2D()
                if rnd > f32(0.2):
                    sc_o, sc_d = s_o, s_d
                lambertian_pdf = self.onb_lambertian_scattering_pdf(sc_d)
                lpdf = self.light_pdf(sc_o, sc_d, self.lights)
                pdf = f32(0.2) * lpdf + f32(0.8) * lambertian_pdf
                if pdf <= f32(0.00001):
                    return mul(emission, thr)
                acc = add(acc, mul(emission, thr))
                thr = mul(thr, divs(scale(mix(color, spec, self.doSpecular), lambertian_pdf), pdf))
                o, d = sc_o, sc_d
            else:
                s_o, s_d = self.material_scatter(o, d)
                acc = add(acc, mul(emission, thr))
                thr = mul(thr, mix(color, spec, self.doSpecular))
                o, d = s_o, s_d
            if i > 2:
                p = fmax(thr[0], fmax(thr[1], thr[2]))
                if self.rand2D() > p:
                    break
                thr = scale(thr, f32(1.0) / p)
        return acc

    # ---------------- shootRay.wgsl ----------------
    def getCameraRay(self, s, t):                      # :54-60
        v4 = mat_vec4(self.view, (s, t, -self.fovFactor, f32(0)))
        l = np.sqrt(((v4[0] * v4[0] + v4[1] * v4[1]) + v4[2] * v4[2]) + v4[3] * v4[3])
        return self.cam_origin, (v4[0] / l, v4[1] / l, v4[2] / l)

    def pathTrace(self):                               # :5-49
        pix = V(0, 0, 0)
        px, py = self.pixelCoords
        if self.STRATIFY:
            sqrt_spp = f32(np.sqrt(np.float64(self.NUM_SAMPLES)))
            recip = f32(1.0) / f32(int(sqrt_spp))
            numSamples = f32(0)
            i = f32(0)
            while i < sqrt_spp:
                j = f32(0)
                while j < sqrt_spp:
                    sx = (self.W / self.H) * (f32(2) * ((px - f32(0.5) + (recip * (i + self.rand2D()))) / self.W) - f32(1))
                    sy = f32(-1) * (f32(2) * ((py - f32(0.5) + (recip * (j + self.rand2D()))) / self.H) - f32(1))
                    o, d = self.getCameraRay(sx, sy)
                    pix = add(pix, self.ray_color(o, d))
                    numSamples = numSamples + f32(1)
                    j = j + f32(1)
                i = i + f32(1)
            pix = divs(pix, numSamples)
        else:
            for _ in range(self.NUM_SAMPLES):
                sx = (self.W / self.H) * (f32(2) * ((px - f32(0.5) + self.rand2D()) / self.W) - f32(1))
                sy = f32(-1) * (f32(2) * ((py - f32(0.5) + self.rand2D()) / self.H) - f32(1))
                o, d = self.getCameraRay(sx, sy)
                pix = add(pix, self.ray_color(o, d))
            pix = divs(pix, f32(self.NUM_SAMPLES))
        return pix

    # ---------------- main.wgsl:1-28 (without the framebuffer write) ----------------
    def invocation(self, pixelIndex, frameNum):
        """-> (pathTracedColor, (type, index, t) of the first closest-hit query)"""
        fi = f32(pixelIndex)
        q = fi / self.W
        self.pixelCoords = (fi - self.W * np.trunc(q), q)
        self.fovFactor = f32(1 / np.tan(60 * (3.1415926535897932385 / 180) / 2))
        self.cam_origin = mat_vec4(self.view, (f32(0), f32(0), f32(0), f32(1)))[:3]
        self.randState = (int(pixelIndex) + int(f32(frameNum)) * 719393) & 0xffffffff
        self.hitRec = HitRecord()
        self.skip_pdf, self.skip_pdf_ray, self.doSpecular = False, (V(0, 0, 0), V(0, 0, 0)), f32(0)
        self.unit_w = self.u = self.v = V(0, 0, 0)
        self.primary = None
        self.get_lights()
        col = self.pathTrace()
        return col, self.primary
