"""Deterministic synthetic stand-in meshes for the assets the reference's benchmark scenes name
but that are not in the mount (/root/reference/.MISSING_LARGE_BLOBS: stanfordBunny, stanfordDragon,
sponzaAtrium, stanfordXYZdragon, buddha, armadillo ...).  SURVEY.md §8(d): a closed genus-0
surface = UV sphere (S segments x R rings -> 2*S*(R-1) triangles) displaced radially by 4-octave
value noise, smooth per-vertex normals; and a box-room-with-columns "atrium".  Output has the
shape ObjReader.load_model returns (objReader.js:66-68): 9 floats per triangle for positions
and for normals, float32.
"""
import numpy as np


def _hash3(ix, iy, iz, seed):
    h = (ix.astype(np.uint64) * np.uint64(73856093)) ^ (iy.astype(np.uint64) * np.uint64(19349663)) ^ \
        (iz.astype(np.uint64) * np.uint64(83492791)) ^ np.uint64(seed * 2654435761 % (1 << 32))
    h = (h ^ (h >> np.uint64(13))) * np.uint64(1274126177)
    h = h ^ (h >> np.uint64(16))
    return (h & np.uint64(0xffffff)).astype(np.float64) / float(0xffffff)


def value_noise(p, seed):
    """Trilinear value noise in [0,1] at points p (N,3)."""
    f = np.floor(p)
    t = p - f
    t = t * t * (3 - 2 * t)
    i = f.astype(np.int64) + 1024
    out = 0.0
    for dx in (0, 1):
        for dy in (0, 1):
            for dz in (0, 1):
                w = (t[:, 0] if dx else 1 - t[:, 0]) * (t[:, 1] if dy else 1 - t[:, 1]) * (t[:, 2] if dz else 1 - t[:, 2])
                out = out + w * _hash3(i[:, 0] + dx, i[:, 1] + dy, i[:, 2] + dz, seed)
    return out


def pick_sphere_resolution(target_tris):
    """S ~ 2R minimising |2*S*(R-1) - target|."""
    best = None
    r0 = max(3, int(round((target_tris / 4.0) ** 0.5)))
    for R in range(max(3, r0 - 40), r0 + 40):
        for S in (2 * R - 3, 2 * R - 2, 2 * R - 1, 2 * R, 2 * R + 1):
            if S < 3:
                continue
            err = abs(2 * S * (R - 1) - target_tris)
            if best is None or err < best[0]:
                best = (err, S, R)
    return best[1], best[2]


def noise_sphere(target_tris, seed, radius=0.4, center=(0.0, 0.45, 0.0), amplitude=0.35, base_freq=2.0):
    S, R = pick_sphere_resolution(target_tris)
    # vertices: north pole, (R-1) rings of S, south pole
    theta = np.pi * np.arange(1, R) / R
    phi = 2 * np.pi * np.arange(S) / S
    tt, pp = np.meshgrid(theta, phi, indexing="ij")
    ring = np.stack([np.sin(tt) * np.cos(pp), np.cos(tt), np.sin(tt) * np.sin(pp)], axis=-1).reshape(-1, 3)
    unit = np.concatenate([[[0.0, 1.0, 0.0]], ring, [[0.0, -1.0, 0.0]]], axis=0)
    disp = np.zeros(unit.shape[0])
    amp, freq = 1.0, base_freq
    for o in range(4):
        disp += amp * (value_noise(unit * freq + 17.0 * o, seed + o) - 0.5)
        amp *= 0.5
        freq *= 2.0
    pos = unit * (radius * (1.0 + amplitude * disp))[:, None] + np.asarray(center)[None, :]
    south = unit.shape[0] - 1

    def vid(r, s):
        return 1 + r * S + (s % S)

    tris = []
    s = np.arange(S)
    tris.append(np.stack([np.zeros(S, np.int64), vid(0, s + 1), vid(0, s)], axis=1))                       # north cap
    for r in range(R - 2):
        a, b, c, d = vid(r, s), vid(r, s + 1), vid(r + 1, s), vid(r + 1, s + 1)
        tris.append(np.stack([a, b, c], axis=1))
        tris.append(np.stack([b, d, c], axis=1))
    tris.append(np.stack([np.full(S, south, np.int64), vid(R - 2, s), vid(R - 2, s + 1)], axis=1))          # south cap
    tris = np.concatenate(tris, axis=0)
    return _finish(pos, tris)


def _finish(pos, tris):
    fn = np.cross(pos[tris[:, 1]] - pos[tris[:, 0]], pos[tris[:, 2]] - pos[tris[:, 0]])
    vn = np.zeros_like(pos)
    for k in range(3):
        np.add.at(vn, tris[:, k], fn)
    ln = np.linalg.norm(vn, axis=1, keepdims=True)
    vn = vn / np.where(ln == 0, 1.0, ln)
    return {"vertices": pos[tris].astype(np.float32).reshape(-1), "normals": vn[tris].astype(np.float32).reshape(-1)}


def _box(lo, hi, n):
    """Tessellated axis-aligned box: each face n x n quads -> 12*n*n triangles; flat normals, outward."""
    lo, hi = np.asarray(lo, float), np.asarray(hi, float)
    V, N = [], []
    g = np.linspace(0.0, 1.0, n + 1)
    for axis in range(3):
        for side in (0, 1):
            u_ax, v_ax = [(1, 2), (2, 0), (0, 1)][axis]
            nrm = np.zeros(3)
            nrm[axis] = 1.0 if side else -1.0
            for i in range(n):
                for j in range(n):
                    def P(a, b):
                        p = np.zeros(3)
                        p[axis] = hi[axis] if side else lo[axis]
                        p[u_ax] = lo[u_ax] + (hi[u_ax] - lo[u_ax]) * a
                        p[v_ax] = lo[v_ax] + (hi[v_ax] - lo[v_ax]) * b
                        return p
                    q = [P(g[i], g[j]), P(g[i + 1], g[j]), P(g[i + 1], g[j + 1]), P(g[i], g[j + 1])]
                    order = [(0, 1, 2), (0, 2, 3)] if side else [(0, 2, 1), (0, 3, 2)]
                    for t in order:
                        V.extend([q[t[0]], q[t[1]], q[t[2]]])
                        N.extend([nrm] * 3)
    return np.asarray(V), np.asarray(N)


def atrium(target_tris=262267, seed=4):
    """Sponza stand-in: an inward-facing box room with two tiers of tessellated square columns and
    lintels (many thin axis-aligned triangles, which exercises AABB.pad, AABB.js:35-51)."""
    parts = []
    cols = [(-1.6 + 0.8 * i, z) for i in range(5) for z in (-0.6, 0.6)]
    n_boxes = 2 * len(cols) + 4
    n = max(1, int(round(((target_tris - 12 * 24 * 24) / (12.0 * n_boxes)) ** 0.5)))
    for tier, (y0, y1) in enumerate(((-1.0, -0.05), (0.05, 0.9))):
        for (x, z) in cols:
            w = 0.09 if tier == 0 else 0.06
            parts.append(_box([x - w, y0, z - w], [x + w, y1, z + w], n))
    for z in (-0.6, 0.6):
        parts.append(_box([-1.75, -0.05, z - 0.12], [1.75, 0.05, z + 0.12], n))
        parts.append(_box([-1.75, 0.9, z - 0.12], [1.75, 0.98, z + 0.12], n))
    V = np.concatenate([p[0] for p in parts], axis=0)
    N = np.concatenate([p[1] for p in parts], axis=0)
    # room shell, inward-facing normals (flip winding and normals of an outward box)
    rv, rn = _box([-2.0, -1.0, -1.2], [2.0, 1.0, 1.2], 24)
    rv = rv.reshape(-1, 3, 3)[:, ::-1, :].reshape(-1, 3)
    V = np.concatenate([V, rv], axis=0)
    N = np.concatenate([N, -rn], axis=0)
    return {"vertices": V.astype(np.float32).reshape(-1), "normals": N.astype(np.float32).reshape(-1)}
