#!/usr/bin/env python
"""bench.py -- the contract benchmark.

  python bench.py --gpus N --steps K --warmup W [--impl reference] [--config C4] [--frames F] [--scaling strong|weak]

A "step" is one pass of the hot path over one batch: the progressive render of the configuration's whole job (F frames
= its samples per pixel: 1,024 for C4, the configuration north_star's target is quoted on) at its full resolution,
through libwgpt.so.  metric = Msamples/s (BASELINE.json): W*H*F*NUM_SAMPLES samples per step.  `value` is timed on the
device (CUDA events on the context's OWN stream -- wgpt_timer_mark -- inputs resident in HBM); `e2e` repeats the step
through the host-buffer API (scene + uniforms uploaded from pinned host memory, framebuffer read back) inside the timed
region.

N > 1 (launched under torchrun, one rank per GPU): the SAME job is split over the ranks ("scaling": "strong") by image
tiles x frame strata (rank r renders tile shard r % T of strata shard r // T), and the accumulation
buffers are combined by one ncclReduce(sum) that libwgpt.so issues itself on the context's stream, inside the timed
region.  torch.distributed is plumbing only (NCCL id broadcast, barriers, max over ranks).  --scaling weak keeps the
per-GPU work fixed instead (the job grows to N*F frames).

--impl reference: the reference's WebGPU path cannot run on this box (no graphics stack, no JS engine); its only
runnable form is the CPU oracle (oracle/, a line-by-line restatement of the WGSL), timed here with all host threads on a
bounded sample of the same workload.
"""
import argparse
import glob
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout carries exactly ONE JSON line.  Libraries print there too (NCCL's version banner under NCCL_DEBUG=VERSION / INFO),
# so file descriptor 1 is pointed at stderr for the whole run and the line goes to a saved copy of the real stdout.
sys.stdout.flush()
_REAL_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(line):
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()


import numpy as np  # noqa: E402

CONFIG_FRAMES = {"C1": 16, "C2": 64, "C3": 256, "C4": 1024, "C5": 1024}
CONFIG_DESC = {
    "C1": "C1 default Cornell-box scene (19 spheres, 8 quads, cube) 512x512, 16 spp",
    "C2": "C2 bunny stand-in (69,432-triangle synthetic mesh) diffuse+glossy in the Cornell room, 1920x1080, 64 spp",
    "C3": "C3 dragon stand-in (298,116 triangles) dielectric, mixture importance sampling, 1920x1080, 256 spp",
    "C4": "C4 sponza stand-in (266,112-triangle atrium), camera inside, 1920x1080, 1024 spp",
    "C5": "C5 two 100K-triangle meshes, 3840x2160, NUM_SAMPLES=4 stratified x 1024 frames",
}
B_SHADE_BOUNCE, B_SHADE_SAMPLE = 160.0, 48.0      # 5 x 16-byte slot records read + written; pixel sum + staging
B_SLOT_IO_EXTEND = 4.0 + 16.0 + 16.0 + 4.0 + 16.0 + 4.0   # queue entry, e0, e1, tag read; x written (triangle hit), t lane


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                pw.append(float(r[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def host_threads():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def kernel_sha():
    """Content hash of every kernel source: profiles/traffic.json records the one its ncu captures were taken on, and
    its numbers are only quoted when the kernels are still the same."""
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(ROOT, "webgpu_path_tracer_b200", "csrc", "*")) + glob.glob(os.path.join(ROOT, "include", "*.h"))):
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


def work_counters(cfg_name):
    """Work per sample on this scene, one frame with counters on:
    reference = the reference's own traversal (exact-mode kernel == oracle counter block, tests/test_gpu_fullsize.py):
                the constants behind B_ray / B_sample of SURVEY.md §8(d);
    wide      = what the fast path really does (8-wide node visits, triangle tests): the bytes k_wf_extend itself moves."""
    import bench_scenes
    from webgpu_path_tracer_b200 import PathTracer, MODE_EXACT, MODE_FAST
    scene, cam, W, H, kw = bench_scenes.build(cfg_name)
    out = {}
    for name, mode in (("reference", MODE_EXACT), ("wide", MODE_FAST)):
        pt = PathTracer(W, H, mode=mode, **kw)
        pt.write_scene(scene.buffers())
        pt.write_uniforms(1, 0, cam.viewMatrix)
        pt.set_counters(True)
        pt.render(1, 1)
        out[name] = pt.counters()
        pt.close()
    c, w = out["reference"], out["wide"]
    n_sph, n_quad = len(scene.spheres), len(scene.quads)
    rays = max(1, c["rays"])
    b_ray = (48.0 * c["node_visits"] + 96.0 * c["tri_tests"]) / rays + 32.0 * n_sph + 80.0 * n_quad + 64.0
    rays_per_sample = c["rays"] / max(1, c["samples"])
    wrays = max(1, w["rays"])
    return {"rays_per_sample": rays_per_sample, "nodes_per_ray": c["node_visits"] / rays, "tris_per_ray": c["tri_tests"] / rays,
            "B_ray": b_ray, "B_sample": 32.0 + rays_per_sample * b_ray,
            "wide_nodes_per_ray": w["node_visits"] / wrays, "wide_tris_per_ray": w["tri_tests"] / wrays,
            "B_ray_wide": 80.0 * w["node_visits"] / wrays + 48.0 * w["tri_tests"] / wrays + B_SLOT_IO_EXTEND}


def run_reference(args):
    """CPU arm: the oracle (kind "port") with every host thread, a bounded sample of the workload per step."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    import bench_scenes
    from oracle.binding import Oracle
    scene, cam, W, H, kw = bench_scenes.build(args.config)
    okw = {k: v for k, v in kw.items()}
    # every host thread the box has (torchrun exports OMP_NUM_THREADS=1 to its workers: do not inherit that)
    okw["threads"] = host_threads()
    orc = Oracle(W, H, **okw)
    orc.write_scene(scene.buffers())
    orc.write_uniforms(1, 0, cam.viewMatrix)
    spp = kw.get("num_samples", 1)
    frames_per_step = max(1, args.ref_frames)
    f = 1
    for _ in range(args.warmup):
        orc.render(f, frames_per_step)
        f += frames_per_step
    t0 = time.perf_counter()
    for _ in range(args.steps):
        orc.render(f, frames_per_step)
        f += frames_per_step
    dt = time.perf_counter() - t0
    samples = W * H * spp * frames_per_step * args.steps
    value = samples / dt / 1e6
    sample_desc = "%d frame(s) of %dx%d per step: the rate is EXTRAPOLATED from %d frames to the %d-frame job (frames are i.i.d. work)" % (
        frames_per_step, W, H, frames_per_step, args.frames)
    line = {"metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": CONFIG_DESC[args.config], "frames_per_step": frames_per_step, "sample": sample_desc},
            "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": orc.threads, "kind": "port", "sample": sample_desc},
            "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference WebGPU path unavailable on this box (no graphics stack / JS engine); this is the CPU oracle"}
    emit(line)


def quick_config(cfg_name, frames, device):
    """Short single-GPU run of another BASELINE.json configuration (device-timed, one warm-up render): `other_configs`."""
    import bench_scenes
    from webgpu_path_tracer_b200 import PathTracer, MODE_FAST
    scene, cam, W, H, kw = bench_scenes.build(cfg_name)
    pt = PathTracer(W, H, mode=MODE_FAST, device=device, **kw)
    pt.write_scene(scene.buffers())
    pt.write_uniforms(1, 0, cam.viewMatrix)
    pt.render(1, min(frames, 8))
    pt.set_stage_profiling(True)
    pt.stage_timings()
    best = None
    for _ in range(2):
        pt.write_buffer(3, np.zeros(W * H * 4, np.float32))
        pt.render(1, frames)
        ms, _n = pt.last_timing()
        st = pt.stage_timings()
        if best is None or ms < best[0]:
            best = (ms, st)
    pt.close()
    ms, st = best
    spp = kw.get("num_samples", 1)
    return {"workload": CONFIG_DESC[cfg_name], "frames": frames, "full_job_frames": CONFIG_FRAMES[cfg_name],
            "Msamples_per_s": W * H * spp * frames / ms / 1e3, "ms": ms,
            "stage_ms": {k: v[0] for k, v in st.items()}}


def tol_numbers(cfg_name, frames, device):
    """The OPT-IN tolerance mode (WGPT_MODE_FAST_TOL: shading on fast arithmetic, intersections / camera / RNG pinned) on the
    bench configuration: rate of the same job, and on one full-size frame against the bit-exact mode: primary ids equal,
    fraction of samples outside 1e-4 relative, max relative error on the paths whose signature (draws + hit sequence) is
    equal.  The bit-exact mode stays the headline; tests/test_gpu_tol.py holds the asserts."""
    import bench_scenes
    from webgpu_path_tracer_b200 import PathTracer, MODE_FAST, MODE_FAST_TOL
    scene, cam, W, H, kw = bench_scenes.build(cfg_name)
    out, probes = {}, {}
    for name, mode in (("fast", MODE_FAST), ("tol", MODE_FAST_TOL)):
        pt = PathTracer(W, H, mode=mode, device=device, **kw)
        pt.write_scene(scene.buffers())
        pt.write_uniforms(2, 0, cam.viewMatrix)
        pt.probe_rng(fetch=False)
        rad, ty, ix, t = pt.probe()
        probes[name] = (rad, ty, ix, t, pt.probe_rng())
        if name == "tol":
            pt.write_uniforms(1, 0, cam.viewMatrix)
            pt.render(1, min(frames, 16))
            pt.write_buffer(3, np.zeros(W * H * 4, np.float32))
            pt.render(1, frames)
            ms, _n = pt.last_timing()
            out["Msamples_per_s"] = W * H * kw.get("num_samples", 1) * frames / ms / 1e3
            out["frames"] = frames
        pt.close()
    (rf, tf, xf, df, sf), (rt, tt, xt, dt, st) = probes["fast"], probes["tol"]
    fin = np.isfinite(rf).all(axis=1) & np.isfinite(rt).all(axis=1)
    scale = np.maximum(np.abs(np.where(np.isfinite(rf), rf, 0)).max(axis=1), 1e-3)
    with np.errstate(invalid="ignore"):
        rel = np.abs(rt.astype(np.float64) - rf).max(axis=1) / scale
    same = sf == st
    out.update({"primary_ids_bit_exact": bool(np.array_equal(tf, tt) and np.array_equal(xf, xt) and np.array_equal(df.view(np.uint32), dt.view(np.uint32))),
                "fraction_outside_1e-4": float(np.where(fin, rel > 1e-4, True).mean()),
                "fraction_signature_diverged": float((~same).mean()),
                "max_rel_err_same_signature": float(rel[same & fin].max()) if (same & fin).any() else None,
                "what": "one full-size frame, tolerance mode vs the bit-exact fast mode (= the oracle, bit for bit)"})
    return out


def per_frame_numbers(device):
    """The reference's own call pattern (renderer.js:163-191): per frame one uniform write, one wgpt_dispatch, one
    present.  (a) strict: dispatch + present every frame (the present waits for the frame: fill + drain of the
    wavefront per frame); (b) pipelined: dispatches enqueued back to back (they return at once and join one living
    wavefront), one present at the end -- what a progressive viewer that refreshes every N frames pays per frame."""
    import bench_scenes
    from webgpu_path_tracer_b200 import PathTracer, Scene, Camera, MODE_FAST
    out = {}
    s = Scene()
    s.init_mesh_data()
    s.create_meshes()
    s.create_bvh()
    cam = Camera()
    cam.set_camera([0.5, 0, 2.5], [0.5, 0, 0], [0, 1, 0])
    cases = [("default_scene_900x600", s, cam, 900, 600, {})]
    sc, cm, W2, H2, kw2 = bench_scenes.build("C2")
    cases.append(("C2_1920x1080", sc, cm, W2, H2, kw2))
    for name, scene, c, W, H, kw in cases:
        pt = PathTracer(W, H, mode=MODE_FAST, device=device, **kw)
        pt.write_scene(scene.buffers())
        for f in (1, 2, 3):                                   # warm-up (wide BVH build, pool allocation)
            pt.write_uniforms(f, 0, c.viewMatrix)
            pt.dispatch()
            pt.present(rgba8="only")
        n = 24
        f = 4
        reps, reps_static = [], []
        for rep in range(3):                                   # host wall clock: repeated, best and all reported
            pt.synchronize()
            t0 = time.perf_counter()
            for _ in range(n):
                pt.write_uniforms(f, 0, c.viewMatrix)
                pt.dispatch()
                pt.present(rgba8="only")
                f += 1
            reps.append((time.perf_counter() - t0) / n * 1e3)
        if W * H <= (2 << 20):                                 # the allocation-free present (no cudaMalloc / cudaFree per frame)
            for rep in range(3):
                pt.synchronize()
                t0 = time.perf_counter()
                for _ in range(n):
                    pt.write_uniforms(f, 0, c.viewMatrix)
                    pt.dispatch()
                    pt.present_rgba8_static()
                    f += 1
                reps_static.append((time.perf_counter() - t0) / n * 1e3)
        strict = min(reps)
        n2 = 96
        t0 = time.perf_counter()
        for f in range(f, f + n2):
            pt.write_uniforms(f, 0, c.viewMatrix)
            pt.dispatch()
        t_enq = time.perf_counter() - t0
        pt.present(rgba8="only")
        piped = (time.perf_counter() - t0) / n2 * 1e3
        pt.close()
        out[name] = {"ms_dispatch_plus_present": strict, "fps": 1e3 / strict, "ms_dispatch_plus_present_reps": reps,
                     "ms_dispatch_plus_static_present": min(reps_static) if reps_static else None, "static_present_reps": reps_static,
                     "pipelined_ms_per_frame": piped, "pipelined_fps": 1e3 / piped,
                     "dispatch_call_ms": t_enq / n2 * 1e3, "frames": [n, n2]}
    out["reference_published"] = "benchmarks.txt:14-16: 104 fps / 9.6 ms per frame (default scene, 900x600, the author's GPU)"
    return out


def mgpu_self_check(dist, rank, world, local_rank):
    """Inside the run: a small job rendered sharded (tiles only, then strata only) + reduced, against the same job on
    rank 0's GPU alone.  Tile shards must reproduce the single-GPU buffer bit for bit, strata to f32 rounding."""
    import bench_scenes
    from webgpu_path_tracer_b200 import PathTracer, MODE_FAST, sharding
    scene, cam, _W, _H, kw = bench_scenes.build("C2")
    W, H, frames = 480, 270, 16
    res = {}
    for name, tiles in (("tiles", world), ("strata", 1)):
        tr, tw, sr, sw = sharding.shard_of_rank(rank, world, tiles)
        pt = PathTracer(W, H, mode=MODE_FAST, device=local_rank, tile_rank=tr, tile_world=tw, strata_rank=sr, strata_world=sw, **kw)
        sharding.init_comm(pt, dist, rank, world)
        pt.write_scene(scene.buffers())
        pt.write_uniforms(1, 0, cam.viewMatrix)
        pt.render(1, frames)
        pt.reduce_framebuffer(0)
        res[name] = pt.read_framebuffer()
        pt.close()
    if rank != 0:
        return None
    one = PathTracer(W, H, mode=MODE_FAST, device=local_rank, **kw)
    one.write_scene(scene.buffers())
    one.write_uniforms(1, 0, cam.viewMatrix)
    one.render(1, frames)
    ref = one.read_framebuffer()
    one.close()
    denom = np.maximum(np.abs(ref[:, :3]), 1e-6)
    return {"tile_bit_exact": bool(np.array_equal(ref.view(np.uint32), res["tiles"].view(np.uint32))),
            "strata_max_rel": float(np.nanmax(np.abs(res["strata"][:, :3] - ref[:, :3]) / denom)),
            "job": "C2 scene %dx%d, %d frames, %d ranks" % (W, H, frames, world)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C4", choices=sorted(CONFIG_FRAMES))
    ap.add_argument("--frames", type=int, default=0, help="frames of the job (default: the configuration's spp)")
    ap.add_argument("--mode", default="fast", choices=["fast", "exact"])
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"])
    ap.add_argument("--tiles", type=int, default=0, help="tile groups at N > 1 (default: 1, 1, 2, 4 at 1, 2, 4, 8 GPUs); strata = N / tiles")
    ap.add_argument("--ref-frames", type=int, default=2, help="frames per step of the CPU arm")
    ap.add_argument("--cpu-baseline-frames", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip other_configs / per_frame / mgpu_check")
    args = ap.parse_args()
    if not args.frames:
        args.frames = CONFIG_FRAMES[args.config]
    if args.impl == "reference":
        return run_reference(args)

    rank, local_rank, world = dist_env()
    import torch
    import bench_scenes
    from webgpu_path_tracer_b200 import PathTracer, MODE_EXACT, MODE_FAST, sharding
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    scene, cam, W, H, kw = bench_scenes.build(args.config)
    spp = kw.get("num_samples", 1)
    mode = MODE_FAST if args.mode == "fast" else MODE_EXACT
    tiles = args.tiles or sharding.default_tile_groups(world)
    tr, tw, sr, sw = sharding.shard_of_rank(rank, world, tiles)
    job_frames = args.frames * (world if args.scaling == "weak" else 1)
    pt = PathTracer(W, H, mode=mode, device=local_rank, tile_rank=tr, tile_world=tw, strata_rank=sr, strata_world=sw, **kw)
    if dist:
        sharding.init_comm(pt, dist, rank, world)          # NCCL id over torch.distributed; the reduce is the library's
    bufs = scene.buffers()
    pt.write_scene(bufs)
    pt.write_uniforms(1, 0, cam.viewMatrix)
    samples_per_step = W * H * spp * job_frames
    zeros_fb = np.zeros(W * H * 4, np.float32)
    l2_flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()
        pt.synchronize()

    def device_step():
        """inputs resident; this rank's tiles of its frames, then ONE ncclReduce on the same stream"""
        pt.render(1, job_frames)
        if dist:
            pt.reduce_framebuffer(0)

    # ---------------- device-timed value ----------------
    work = work_counters(args.config) if rank == 0 else None
    pt.set_stage_profiling(True)
    for _ in range(args.warmup):
        pt.write_buffer(3, zeros_fb)
        device_step()
    pt.stage_timings()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    step_ms, launches = [], 0
    t_wall0 = time.perf_counter()
    for _ in range(args.steps):
        pt.write_buffer(3, zeros_fb)
        l2_flush.zero_()                           # 256 MB > L2: nothing of the previous step stays cached
        barrier()
        pt.timer_mark(0)
        device_step()
        pt.timer_mark(1)
        step_ms.append(pt.timer_elapsed())         # CUDA events on the context's (launching) stream: render + reduce
        launches += pt.last_timing()[1] + (1 if dist else 0)
    barrier()
    wall = time.perf_counter() - t_wall0
    clocks = sampler.stop()
    stage = pt.stage_timings()
    t_local = torch.tensor([sum(step_ms)], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(t_local, op=dist.ReduceOp.MAX)
    total_ms = float(t_local.item())
    value = samples_per_step * args.steps / (total_ms * 1e-3) / 1e6
    # where the ranks' time went at N > 1: own render vs waiting in / running the reduce
    render_ms = pt.last_timing()[0]

    # ---------------- end-to-end through the host-buffer API ----------------
    pinned = {b: torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for b, a in bufs.items()}
    pinned_np = {b: t.numpy() for b, t in pinned.items()}
    uni = np.zeros(20, np.float32)
    uni[0], uni[1], uni[2], uni[3] = W, H, 1, 0
    uni[4:] = np.asarray(cam.viewMatrix, np.float32)
    uni_p = torch.from_numpy(uni).pin_memory()
    h2d = sum(a.nbytes for a in bufs.values()) + 80 + zeros_fb.nbytes
    d2h = W * H * 16
    e2e_times = []
    e2e_steps = max(1, min(args.steps, 3))
    for it in range(1 + e2e_steps):
        barrier()
        t0 = time.perf_counter()
        for b, a in pinned_np.items():
            pt.write_buffer(b, a)                  # host -> device, re-derives the hierarchy (K5 + K2) each step
        pt.write_buffer(0, uni_p.numpy())
        pt.write_buffer(3, zeros_fb)
        device_step()
        out = pt.read_framebuffer() if rank == 0 or not dist else None   # device -> host result
        barrier()
        if it >= 1:
            e2e_times.append(time.perf_counter() - t0)
    e2e_t = torch.tensor([sum(e2e_times)], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_value = samples_per_step * e2e_steps / float(e2e_t.item()) / 1e6

    mgpu_check = None
    if dist and not args.no_extras:
        mgpu_check = mgpu_self_check(dist, rank, world, local_rank)
    if rank != 0:
        if dist:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---------------- rooflines of the two kernels that carry the step ----------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    # ncu numbers of ONE steady-state launch per (stage, config) (tools/evidence_collect.py): only quoted when the
    # kernel sources are still the ones they were captured on
    cap_all, cap = {}, {}
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        cap_all = json.load(open(tp))
        if cap_all.get("kernel_sha") == kernel_sha():
            cap = cap_all.get(args.config, {})
    frames_rank = len(sharding.strata_frames(1, job_frames, sr, sw))
    pix_rank = sharding.shard_pixel_count(W, H, tr, tw, 8) if tw > 1 else W * H
    samples_rank = min(pix_rank, W * H) * spp * frames_rank
    rays_rank = samples_rank * work["rays_per_sample"]
    sm_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
    issue_peak = 148 * 4 * sm_hz                     # warp instructions per second: 148 SMs x 4 schedulers x SM clock

    def stage_roofline(kernel, key, alg_bytes_rank, what):
        ms, n = stage[key]
        if not n or ms <= 0:
            return None
        achieved = alg_bytes_rank * args.steps / (ms * 1e-3) / 1e9
        r = {"bound": "hbm", "kernel": kernel, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
             "peak_source": peak_src, "algorithmic_bytes": what,
             "algorithmic_bytes_per_launch": alg_bytes_rank * args.steps / n, "avg_launch_ms": ms / n, "launches": n,
             "share_of_step": ms / sum(step_ms), "traffic": None}
        k = cap.get(key)
        if k and k.get("algorithmic_bytes"):
            # same-launch pair: DRAM bytes ncu measured for the captured launch vs the algorithmic bytes of THAT launch
            r["traffic"] = k["dram_bytes"]
            r["traffic_launch"] = {"dram_bytes": k["dram_bytes"], "algorithmic_bytes": k["algorithmic_bytes"], "ms": k["ms"],
                                   "ratio": k["dram_bytes"] / k["algorithmic_bytes"], "what": k.get("what"),
                                   "physical_dram_GBps": k["dram_bytes"] / (k["ms"] * 1e-3) / 1e9,
                                   "physical_dram_frac_of_peak": k["dram_bytes"] / (k["ms"] * 1e-3) / 1e9 / peak,
                                   "l1_hit_pct": k.get("l1_hit_pct"), "l2_hit_pct": k.get("l2_hit_pct"),
                                   "warps_active_pct": k.get("warps_active_pct")}
            if k.get("warp_inst_per_unit"):
                units = rays_rank
                inst_s = k["warp_inst_per_unit"] * units * args.steps / (ms * 1e-3)
                r["issue"] = {"achieved": inst_s / 1e9, "peak": issue_peak / 1e9, "unit": "G warp-inst/s", "frac": inst_s / issue_peak,
                              "warp_inst_per_ray": k["warp_inst_per_unit"], "sm_mhz": sm_hz / 1e6,
                              "ncu_issue_active_pct": k.get("issue_active_pct"), "ncu_lanes_per_inst": k.get("lanes"),
                              "ncu_long_scoreboard_per_issue": k.get("long_sb")}
        return r

    r_ext = stage_roofline("k_wf_extend", "extend", rays_rank * work["B_ray_wide"],
                           "80 B x wide-node visits + 48 B x triangle tests + 60 B slot / queue I/O per ray (fast-mode counters): %.0f B/ray; "
                           "served by L1 / L2 (the scene lives in L2), so HBM is not what binds this kernel -- see `issue` and DESIGN.md" % work["B_ray_wide"])
    if r_ext:
        ms = stage["extend"][0]
        r_ext["reference_traversal"] = {"B_ray": work["B_ray"], "equivalent_GBps": rays_rank * work["B_ray"] * args.steps / (ms * 1e-3) / 1e9,
                                        "note": "SURVEY.md 8(d) bytes of the REFERENCE's binary traversal per ray over this kernel's time: a speed-up "
                                                "measure against the reference algorithm, not a roofline fraction"}
    r_shd = stage_roofline("k_wf_shade", "shade", rays_rank * B_SHADE_BOUNCE + samples_rank * B_SHADE_SAMPLE,
                           "160 B per path-bounce (5 x 16-byte slot records read + written) + 48 B per finished sample")
    both = [r for r in (r_ext, r_shd) if r]
    roofline = max(both, key=lambda r: r["share_of_step"]) if both else None
    if roofline:
        roofline = dict(roofline)
        roofline.update({"B_ray": work["B_ray"], "B_sample": work["B_sample"], "rays_per_sample": work["rays_per_sample"],
                         "wide_nodes_per_ray": work["wide_nodes_per_ray"], "wide_tris_per_ray": work["wide_tris_per_ray"],
                         "stage_ms": {k: v[0] for k, v in stage.items()},
                         "capture_kernel_sha": cap_all.get("kernel_sha"), "kernel_sha": kernel_sha(),
                         "other_kernel": [r for r in both if r["kernel"] != roofline["kernel"]]})

    # ---------------- CPU baseline (oracle, bounded sample) ----------------
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        from oracle.binding import Oracle
        orc = Oracle(W, H, threads=host_threads(), **kw)
        orc.write_scene(bufs)
        orc.write_uniforms(1, 0, cam.viewMatrix)
        orc.render(1, 1)
        nfr = max(1, args.cpu_baseline_frames)
        t0 = time.perf_counter()
        orc.render(2, nfr)
        dt = time.perf_counter() - t0
        cpu = {"value": W * H * spp * nfr / dt / 1e6, "unit": "Msamples/s", "cores": orc.threads, "kind": "port",
               "sample": "%d frames of %dx%d (%.1f s of CPU work), extrapolated to the %d-frame job of the GPU step" % (nfr, W, H, dt, job_frames)}

    fstats = pt.fast_statistics() if args.mode == "fast" else None
    other, per_frame, tol_mode = None, None, None
    if world == 1 and not args.no_extras and args.mode == "fast":
        del l2_flush
        pt.close()
        other = {}
        for c, fr in (("C1", 256), ("C2", 64), ("C3", 64), ("C5", 16)):
            if c != args.config:
                other[c] = quick_config(c, fr, local_rank)
        per_frame = per_frame_numbers(local_rank)
        tol_mode = tol_numbers(args.config, min(job_frames, 256), local_rank)

    if world > 1:
        par = "%d tile group(s) x %d frame strata, ncclReduce(sum) issued by libwgpt.so on the render stream" % (tw, sw)
    else:
        par = "single GPU"
    line = {"metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": CONFIG_DESC[args.config], "job_frames": job_frames, "frames_rendered_by_rank0": frames_rank,
                       "pixels_rendered_by_rank0": min(pix_rank, W * H), "mode": args.mode, "parallelism": par,
                       "l2": "flushed between timed steps (256 MB device memset); the scene itself (<= 60 MB) is meant to live in L2",
                       "hierarchy": "device-built (K5 PLOC) wide BVH" if os.environ.get("WGPT_BVH", "d")[0] != "r" else "reference tree flattened"},
            "Mrays_per_s": value * work["rays_per_sample"],
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps},
            "gpu_launches": int(launches),
            "roofline": roofline, "cpu_baseline": cpu, "wall_s": wall,
            "other_configs": other, "per_frame": per_frame, "tol_mode": tol_mode, "mgpu_check": mgpu_check,
            "rank0_last_render_ms": render_ms,
            "fast_stats": fstats}
    emit(line)
    if dist:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
