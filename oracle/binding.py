"""ctypes binding of the CPU oracle -- TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package (webgpu_path_tracer_b200) never does.
PARITY UNPINNED: see oracle/oracle_shader.h.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        L.oracle_create.restype = C.c_void_p
        L.oracle_create.argtypes = [C.c_int] * 8
        L.oracle_destroy.argtypes = [C.c_void_p]
        L.oracle_threads.argtypes = [C.c_void_p]
        L.oracle_write_buffer.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        L.oracle_read_framebuffer.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_dispatch_ex.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4
        L.oracle_dispatch.argtypes = [C.c_void_p]
        L.oracle_render.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.oracle_present.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_trace_rays.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 3
        L.oracle_counters.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.oracle_rand_sequence.argtypes = [C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]
        L.oracle_pinned_math.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


COUNTER_NAMES = ("rays", "node_visits", "tri_tests", "sphere_tests", "quad_tests", "rng_draws", "samples", "stack_aborts")


class Oracle:
    """Same call shapes as webgpu_path_tracer_b200.api.PathTracer so parity tests drive both alike."""

    def __init__(self, width, height, num_samples=1, max_bounces=100, stratify=False,
                 importance_sampling=False, stack_size=20, threads=0):
        self.W, self.H = int(width), int(height)
        self.h = lib().oracle_create(self.W, self.H, num_samples, max_bounces, int(stratify),
                                     int(importance_sampling), stack_size, threads)
        self.threads = lib().oracle_threads(self.h)

    def close(self):
        if self.h:
            lib().oracle_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def write_buffer(self, binding, array):
        a = np.ascontiguousarray(array)
        rc = lib().oracle_write_buffer(self.h, binding, _p(a), a.nbytes)
        if rc:
            raise ValueError("oracle_write_buffer(%d) -> %d" % (binding, rc))

    def write_scene(self, buffers):
        for b, a in buffers.items():
            self.write_buffer(b, a)

    def write_uniforms(self, frame_num, reset_buffer, view_matrix):
        u = np.zeros(20, np.float32)
        u[0], u[1], u[2], u[3] = self.W, self.H, frame_num, reset_buffer
        u[4:20] = np.asarray(view_matrix, np.float32)
        self.write_buffer(0, u)

    def dispatch(self):
        lib().oracle_dispatch(self.h)

    def render(self, first_frame, n_frames):
        rc = lib().oracle_render(self.h, first_frame, n_frames)
        if rc:
            raise RuntimeError("oracle_render -> %d" % rc)

    def read_framebuffer(self):
        out = np.empty((self.H * self.W, 4), np.float32)
        lib().oracle_read_framebuffer(self.h, _p(out))
        return out

    def present(self):
        out = np.empty((self.H * self.W, 4), np.float32)
        lib().oracle_present(self.h, _p(out))
        return out

    def probe(self):
        """One dispatch with the current uniforms that does NOT touch the framebuffer; returns
        (radiance (N,3), prim_type (N,), prim_index (N,), prim_t (N,))."""
        n = self.W * self.H
        rad = np.empty((n, 3), np.float32)
        pt = np.empty(n, np.int32)
        pi = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        lib().oracle_dispatch_ex(self.h, 0, _p(rad), _p(pt), _p(pi), _p(t))
        return rad, pt, pi, t

    def trace_rays(self, rays, seeds=None, brute_force=False):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = rays.shape[0]
        seeds = np.ascontiguousarray(seeds, np.uint32) if seeds is not None else np.zeros(n, np.uint32)
        pt = np.empty(n, np.int32)
        pi = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        lib().oracle_trace_rays(self.h, n, _p(rays), _p(seeds), int(brute_force), _p(pt), _p(pi), _p(t))
        return pt, pi, t

    def counters(self, reset=False):
        out = np.zeros(8, np.uint64)
        lib().oracle_counters(self.h, _p(out), int(reset))
        return dict(zip(COUNTER_NAMES, (int(x) for x in out)))


def rand_sequence(seed, n):
    states = np.empty(n, np.uint32)
    values = np.empty(n, np.float32)
    lib().oracle_rand_sequence(seed, n, _p(states), _p(values))
    return states, values


def pinned_math(fn, x, y=None):
    names = {"sin": 0, "cos": 1, "acos": 2, "log": 3, "pow": 4}
    x = np.ascontiguousarray(x, np.float32)
    y = np.ascontiguousarray(y if y is not None else np.zeros_like(x), np.float32)
    out = np.empty_like(x)
    lib().oracle_pinned_math(names[fn], x.size, _p(x), _p(y), _p(out))
    return out
