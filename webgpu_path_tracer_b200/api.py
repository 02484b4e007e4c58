"""ctypes binding of libwgpt.so (include/wgpt.h) -- the only way Python reaches the CUDA path.

There is no CPU fallback here: if the shared library is missing, or no CUDA device exists,
the constructor raises.  Nothing in this package imports oracle/.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libwgpt.so")

MODE_EXACT, MODE_FAST, MODE_FAST_TOL = 0, 1, 2
BINDING_UNIFORMS, BINDING_SPHERES, BINDING_QUADS, BINDING_FRAMEBUFFER = 0, 1, 2, 3
BINDING_TRIANGLES, BINDING_MESHES, BINDING_TRANSFORMS, BINDING_MATERIALS, BINDING_BVH = 5, 6, 7, 8, 9
STRIDES = {0: 80, 1: 32, 2: 80, 3: 16, 5: 96, 6: 16, 7: 128, 8: 64, 9: 48}
COUNTER_NAMES = ("rays", "node_visits", "tri_tests", "sphere_tests", "quad_tests", "rng_draws", "samples", "stack_aborts")

EXPORTS = ("wgpt_default_config", "wgpt_create", "wgpt_destroy", "wgpt_last_error", "wgpt_write_buffer",
           "wgpt_read_buffer", "wgpt_device_ptr", "wgpt_dispatch", "wgpt_render", "wgpt_present", "wgpt_probe",
           "wgpt_trace_rays", "wgpt_counters", "wgpt_set_counters", "wgpt_last_timing", "wgpt_synchronize",
           "wgpt_pinned_math", "wgpt_fast_statistics", "wgpt_set_stage_profiling", "wgpt_stage_timings", "wgpt_k4_timing", "wgpt_present_quant_check",
           "wgpt_timer_mark", "wgpt_timer_elapsed", "wgpt_probe_rng", "wgpt_extend_history", "wgpt_present_rgba8_static", "wgpt_nccl_unique_id", "wgpt_comm_init", "wgpt_reduce_framebuffer", "wgpt_group_create", "wgpt_group_destroy",
           "wgpt_group_last_error", "wgpt_group_size", "wgpt_group_ctx", "wgpt_group_write_buffer", "wgpt_group_render",
           "wgpt_group_reduce")


class WgptError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("width", C.c_int32), ("height", C.c_int32), ("device", C.c_int32),
                ("mode", C.c_int32), ("num_samples", C.c_int32), ("max_bounces", C.c_int32), ("stratify", C.c_int32),
                ("importance_sampling", C.c_int32), ("stack_size", C.c_int32),
                ("tile_rank", C.c_int32), ("tile_world", C.c_int32), ("tile_rows", C.c_int32),
                ("strata_rank", C.c_int32), ("strata_world", C.c_int32)]


class FastStats(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in ("wide_nodes", "wide_tris", "wide_levels", "pool_slots", "ring_frames",
                                          "fallback_rays", "iterations", "trace_fallback_rays")]


class CounterBlock(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in COUNTER_NAMES]


_lib = None


def load_library(path=None):
    """dlopen libwgpt.so and declare the prototypes of every symbol include/wgpt.h exports."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("WGPT_LIB") or LIB_PATH      # WGPT_LIB: developer A/B builds
    if not os.path.exists(p):
        raise WgptError("libwgpt.so not built (%s): run `python -c 'import __graft_entry__ as g; g.build()'`; "
                        "there is no CPU fallback" % p)
    L = C.CDLL(p)
    vp, i32, u32, sz = C.c_void_p, C.c_int, C.c_uint32, C.c_size_t
    L.wgpt_default_config.argtypes = [C.POINTER(Config)]
    L.wgpt_default_config.restype = None
    L.wgpt_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    L.wgpt_destroy.argtypes = [vp]
    L.wgpt_destroy.restype = None
    L.wgpt_last_error.argtypes = [vp]
    L.wgpt_last_error.restype = C.c_char_p
    L.wgpt_write_buffer.argtypes = [vp, i32, vp, sz]
    L.wgpt_read_buffer.argtypes = [vp, i32, vp, sz]
    L.wgpt_device_ptr.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(sz)]
    L.wgpt_dispatch.argtypes = [vp, u32]
    L.wgpt_render.argtypes = [vp, i32, i32]
    L.wgpt_present.argtypes = [vp, vp, vp]
    L.wgpt_probe.argtypes = [vp, vp, vp, vp, vp]
    L.wgpt_trace_rays.argtypes = [vp, i32, vp, vp, vp, vp, vp]
    L.wgpt_counters.argtypes = [vp, C.POINTER(CounterBlock), i32]
    L.wgpt_set_counters.argtypes = [vp, i32]
    L.wgpt_last_timing.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_uint64)]
    L.wgpt_synchronize.argtypes = [vp]
    L.wgpt_pinned_math.argtypes = [vp, i32, i32, vp, vp, vp]
    L.wgpt_k4_timing.argtypes = [vp, i32, i32, vp, vp]
    L.wgpt_present_quant_check.argtypes = [vp, u32, u32, vp, vp]
    L.wgpt_fast_statistics.argtypes = [vp, C.POINTER(FastStats)]
    L.wgpt_set_stage_profiling.argtypes = [vp, i32]
    L.wgpt_stage_timings.argtypes = [vp, C.POINTER(C.c_float * 4), C.POINTER(C.c_uint64 * 4)]
    L.wgpt_present_rgba8_static.argtypes = [vp, vp]
    L.wgpt_extend_history.argtypes = [vp, vp]
    L.wgpt_probe_rng.argtypes = [vp, vp]
    L.wgpt_timer_mark.argtypes = [vp, i32]
    L.wgpt_timer_elapsed.argtypes = [vp, C.POINTER(C.c_float)]
    L.wgpt_nccl_unique_id.argtypes = [vp]
    L.wgpt_comm_init.argtypes = [vp, vp, i32, i32]
    L.wgpt_reduce_framebuffer.argtypes = [vp, i32]
    L.wgpt_group_create.argtypes = [C.POINTER(Config), i32, vp, i32, C.POINTER(vp)]
    L.wgpt_group_destroy.argtypes = [vp]
    L.wgpt_group_destroy.restype = None
    L.wgpt_group_last_error.argtypes = [vp]
    L.wgpt_group_last_error.restype = C.c_char_p
    L.wgpt_group_size.argtypes = [vp]
    L.wgpt_group_ctx.argtypes = [vp, i32]
    L.wgpt_group_ctx.restype = vp
    L.wgpt_group_write_buffer.argtypes = [vp, i32, vp, sz]
    L.wgpt_group_render.argtypes = [vp, i32, i32]
    L.wgpt_group_reduce.argtypes = [vp, i32]
    if path is None:
        _lib = L
    return L


def nccl_unique_id():
    """128 opaque bytes for wgpt_comm_init: generated on ONE rank, shipped to the others by the caller."""
    buf = (C.c_ubyte * 128)()
    rc = load_library().wgpt_nccl_unique_id(buf)
    if rc:
        raise WgptError("wgpt_nccl_unique_id -> %d: %s" % (rc, load_library().wgpt_last_error(None).decode()))
    return bytes(buf)


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class PathTracer:
    """One wgpt_ctx: the bind group + pipelines of renderer.js:105-139 on one CUDA device."""

    def __init__(self, width, height, mode=MODE_EXACT, device=0, num_samples=1, max_bounces=100, stratify=False,
                 importance_sampling=False, stack_size=20, tile_rank=0, tile_world=1, tile_rows=8,
                 strata_rank=0, strata_world=1):
        self.L = load_library()
        cfg = Config()
        self.L.wgpt_default_config(C.byref(cfg))
        cfg.width, cfg.height, cfg.device, cfg.mode = int(width), int(height), int(device), int(mode)
        cfg.num_samples, cfg.max_bounces = int(num_samples), int(max_bounces)
        cfg.stratify, cfg.importance_sampling, cfg.stack_size = int(stratify), int(importance_sampling), int(stack_size)
        cfg.tile_rank, cfg.tile_world, cfg.tile_rows = int(tile_rank), int(tile_world), int(tile_rows)
        cfg.strata_rank, cfg.strata_world = int(strata_rank), int(strata_world)
        self.cfg = cfg
        self.W, self.H = int(width), int(height)
        self.h = C.c_void_p()
        rc = self.L.wgpt_create(C.byref(cfg), C.byref(self.h))
        if rc:
            msg = self.L.wgpt_last_error(None).decode()
            self.h = None
            raise WgptError("wgpt_create -> %d: %s" % (rc, msg))

    def _ck(self, rc, what):
        if rc:
            raise WgptError("%s -> %d: %s" % (what, rc, self.L.wgpt_last_error(self.h).decode()))

    def close(self):
        if getattr(self, "h", None):
            self.L.wgpt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- multi-GPU, one process per GPU (include/wgpt.h, csrc/mgpu.cu) ----
    def comm_init(self, unique_id, rank, world):
        buf = (C.c_ubyte * 128).from_buffer_copy(bytes(unique_id))
        self._ck(self.L.wgpt_comm_init(self.h, buf, int(rank), int(world)), "wgpt_comm_init")

    def reduce_framebuffer(self, root=0):
        """ncclReduce(sum) of binding 3 onto `root` (-1: all ranks), enqueued behind the renders on the context's stream."""
        self._ck(self.L.wgpt_reduce_framebuffer(self.h, int(root)), "wgpt_reduce_framebuffer")

    # ---- buffers ----
    def write_buffer(self, binding, array):
        a = np.ascontiguousarray(array)
        self._ck(self.L.wgpt_write_buffer(self.h, binding, _p(a), a.nbytes), "wgpt_write_buffer(%d)" % binding)

    def write_scene(self, buffers):
        for b, a in buffers.items():
            self.write_buffer(b, a)

    def write_uniforms(self, frame_num, reset_buffer, view_matrix):
        u = np.zeros(20, np.float32)
        u[0], u[1], u[2], u[3] = self.W, self.H, frame_num, reset_buffer
        u[4:20] = np.asarray(view_matrix, np.float32)
        self.write_buffer(0, u)

    def read_framebuffer(self):
        out = np.empty((self.H * self.W, 4), np.float32)
        self._ck(self.L.wgpt_read_buffer(self.h, 3, _p(out), out.nbytes), "wgpt_read_buffer(3)")
        return out

    def framebuffer_device_ptr(self):
        ptr, nbytes = C.c_void_p(), C.c_size_t()
        self._ck(self.L.wgpt_device_ptr(self.h, 3, C.byref(ptr), C.byref(nbytes)), "wgpt_device_ptr")
        return ptr.value, nbytes.value

    # ---- frames ----
    def dispatch(self, workgroups=None):
        if workgroups is None:
            workgroups = (self.W * self.H) // 64 + 1          # renderer.js:187-188 + webgpu-utils.js:130
        self._ck(self.L.wgpt_dispatch(self.h, int(workgroups)), "wgpt_dispatch")

    def render(self, first_frame, n_frames):
        self._ck(self.L.wgpt_render(self.h, int(first_frame), int(n_frames)), "wgpt_render")

    def present(self, rgba8=False):
        """fs output as float RGBA; rgba8=True adds the 8-bit canvas bytes; rgba8="only" returns just the bytes."""
        out = np.empty((self.H * self.W, 4), np.float32) if rgba8 != "only" else None
        out8 = np.empty((self.H * self.W, 4), np.uint8) if rgba8 else None
        self._ck(self.L.wgpt_present(self.h, _p(out), _p(out8)), "wgpt_present")
        if rgba8 == "only":
            return out8
        return (out, out8) if rgba8 else out

    def present_rgba8_static(self):
        """wgpt_present_rgba8_static: the canvas bytes through the allocation-free entry point (static device staging)."""
        out8 = np.empty((self.H * self.W, 4), np.uint8)
        self._ck(self.L.wgpt_present_rgba8_static(self.h, _p(out8)), "wgpt_present_rgba8_static")
        return out8

    def extend_history(self):
        out = np.zeros(64, np.uint32)
        self._ck(self.L.wgpt_extend_history(self.h, _p(out)), "wgpt_extend_history")
        return out

    def synchronize(self):
        self._ck(self.L.wgpt_synchronize(self.h), "wgpt_synchronize")

    def timer_mark(self, which):
        self._ck(self.L.wgpt_timer_mark(self.h, int(which)), "wgpt_timer_mark")

    def timer_elapsed(self):
        ms = C.c_float()
        self._ck(self.L.wgpt_timer_elapsed(self.h, C.byref(ms)), "wgpt_timer_elapsed")
        return ms.value

    def last_timing(self):
        ms, n = C.c_float(), C.c_uint64()
        self._ck(self.L.wgpt_last_timing(self.h, C.byref(ms), C.byref(n)), "wgpt_last_timing")
        return ms.value, n.value

    # ---- parity hooks ----
    def probe(self):
        n = self.W * self.H
        rad, pt = np.empty((n, 3), np.float32), np.empty(n, np.int32)
        pi, t = np.empty(n, np.int32), np.empty(n, np.float32)
        self._ck(self.L.wgpt_probe(self.h, _p(rad), _p(pt), _p(pi), _p(t)), "wgpt_probe")
        return rad, pt, pi, t

    def probe_rng(self, fetch=True):
        """Path signatures (final RNG state per pixel) of the last probe(); probe_rng(fetch=False) switches recording on."""
        out = np.zeros(self.W * self.H, np.uint32) if fetch else None
        self._ck(self.L.wgpt_probe_rng(self.h, _p(out)), "wgpt_probe_rng")
        return out

    def trace_rays(self, rays, seeds=None):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = rays.shape[0]
        seeds = np.ascontiguousarray(seeds, np.uint32) if seeds is not None else None
        pt, pi, t = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n, np.float32)
        self._ck(self.L.wgpt_trace_rays(self.h, n, _p(rays), _p(seeds), _p(pt), _p(pi), _p(t)), "wgpt_trace_rays")
        return pt, pi, t

    def set_counters(self, enabled):
        self._ck(self.L.wgpt_set_counters(self.h, int(enabled)), "wgpt_set_counters")

    def counters(self, reset=False):
        blk = CounterBlock()
        self._ck(self.L.wgpt_counters(self.h, C.byref(blk), int(reset)), "wgpt_counters")
        return {n: int(getattr(blk, n)) for n in COUNTER_NAMES}

    def fast_statistics(self):
        st = FastStats()
        self._ck(self.L.wgpt_fast_statistics(self.h, C.byref(st)), "wgpt_fast_statistics")
        return {n: int(getattr(st, n)) for n, _ in FastStats._fields_}

    def set_stage_profiling(self, enabled):
        self._ck(self.L.wgpt_set_stage_profiling(self.h, int(enabled)), "wgpt_set_stage_profiling")

    def stage_timings(self):
        ms, n = (C.c_float * 4)(), (C.c_uint64 * 4)()
        self._ck(self.L.wgpt_stage_timings(self.h, C.byref(ms), C.byref(n)), "wgpt_stage_timings")
        names = ("extend", "shade", "control_accumulate", "wide_build")
        return {k: (float(ms[i]), int(n[i])) for i, k in enumerate(names)}

    def k4_timing(self, kernel, reps=20):
        """Mean device ms and bytes of one K4 launch: kernel = 'accumulate' | 'present_rgba8' | 'present_f32'."""
        names = {"accumulate": 0, "present_rgba8": 1, "present_f32": 2}
        ms, nbytes = C.c_float(), C.c_uint64()
        self._ck(self.L.wgpt_k4_timing(self.h, names[kernel], int(reps), C.byref(ms), C.byref(nbytes)), "wgpt_k4_timing")
        return float(ms.value), int(nbytes.value)

    def present_quant_check(self, first_bits, count):
        """(mismatching bytes, inputs on the pinned path) of the bytes-only present over a range of float bit patterns."""
        bad, slow = C.c_uint64(), C.c_uint64()
        self._ck(self.L.wgpt_present_quant_check(self.h, int(first_bits), int(count), C.byref(bad), C.byref(slow)), "wgpt_present_quant_check")
        return int(bad.value), int(slow.value)

    def pinned_math(self, fn, x, y=None):
        names = {"sin": 0, "cos": 1, "acos": 2, "log": 3, "pow": 4, "vdiv": 5, "sincos_sin": 6, "sincos_cos": 7}
        x = np.ascontiguousarray(x, np.float32)
        y = np.ascontiguousarray(y, np.float32) if y is not None else None
        out = np.empty_like(x)
        self._ck(self.L.wgpt_pinned_math(self.h, names[fn], x.size, _p(x), _p(y), _p(out)), "wgpt_pinned_math")
        return out


class _GroupMember(PathTracer):
    """A context owned by a Group: same calls as PathTracer, destroyed with the group."""

    def __init__(self, L, handle, W, H):          # no wgpt_create: the group made the context
        self.L, self.h, self.W, self.H = L, C.c_void_p(handle), W, H

    def close(self):
        self.h = None


class Group:
    """wgpt_group: one process, N devices, tile x strata sharding, ncclReduce inside the library."""

    def __init__(self, width, height, n_devices, tile_groups=1, devices=None, mode=MODE_FAST, num_samples=1, max_bounces=100,
                 stratify=False, importance_sampling=False, stack_size=20, tile_rows=8):
        self.L = load_library()
        cfg = Config()
        self.L.wgpt_default_config(C.byref(cfg))
        cfg.width, cfg.height, cfg.mode = int(width), int(height), int(mode)
        cfg.num_samples, cfg.max_bounces = int(num_samples), int(max_bounces)
        cfg.stratify, cfg.importance_sampling, cfg.stack_size = int(stratify), int(importance_sampling), int(stack_size)
        cfg.tile_rows = int(tile_rows)
        self.W, self.H, self.n = int(width), int(height), int(n_devices)
        devs = (C.c_int * self.n)(*devices) if devices is not None else None
        self.h = C.c_void_p()
        rc = self.L.wgpt_group_create(C.byref(cfg), self.n, devs, int(tile_groups), C.byref(self.h))
        if rc:
            self.h = None
            raise WgptError("wgpt_group_create -> %d: %s" % (rc, self.L.wgpt_last_error(None).decode()))

    def _ck(self, rc, what):
        if rc:
            raise WgptError("%s -> %d: %s" % (what, rc, self.L.wgpt_group_last_error(self.h).decode()))

    def member(self, rank):
        return _GroupMember(self.L, self.L.wgpt_group_ctx(self.h, int(rank)), self.W, self.H)

    def write_buffer(self, binding, array):
        a = np.ascontiguousarray(array)
        self._ck(self.L.wgpt_group_write_buffer(self.h, binding, _p(a), a.nbytes), "wgpt_group_write_buffer(%d)" % binding)

    def write_scene(self, buffers):
        for b, a in buffers.items():
            self.write_buffer(b, a)

    def write_uniforms(self, frame_num, reset_buffer, view_matrix):
        u = np.zeros(20, np.float32)
        u[0], u[1], u[2], u[3] = self.W, self.H, frame_num, reset_buffer
        u[4:20] = np.asarray(view_matrix, np.float32)
        self.write_buffer(0, u)

    def render(self, first_frame, n_frames):
        self._ck(self.L.wgpt_group_render(self.h, int(first_frame), int(n_frames)), "wgpt_group_render")

    def reduce(self, root=0):
        self._ck(self.L.wgpt_group_reduce(self.h, int(root)), "wgpt_group_reduce")

    def close(self):
        if getattr(self, "h", None):
            self.L.wgpt_group_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
