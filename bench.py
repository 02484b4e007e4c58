#!/usr/bin/env python
"""bench.py -- the contract benchmark.

  python bench.py --gpus N --steps K --warmup W [--impl reference] [--config C2] [--frames F]

A "step" is one pass of the hot path over one batch: the progressive render of F frames
(F = the configuration's samples per pixel: 64 for C2) of the configuration's scene at its full
resolution, through libwgpt.so.  metric = Msamples/s (BASELINE.json): W*H*F*NUM_SAMPLES samples
per step per GPU.  `value` is timed on the device (CUDA events on the context's stream, inputs
resident in HBM); `e2e` repeats the step through the host-buffer API (scene + uniforms uploaded
from pinned host memory, framebuffer read back) inside the timed region.

N > 1 (launched under torchrun, one rank per GPU): sample-strata sharding --
rank r renders frames f = r (mod N) of an N*F-frame job on the full image, so per-GPU work is
fixed ("weak"), and the accumulation buffers are combined with one NCCL sum-reduce per step
inside the timed region.

--impl reference: the reference's WebGPU path cannot run on this box (no graphics stack, no JS
engine); its only runnable form is the CPU oracle (oracle/, a line-by-line restatement of the
WGSL), timed here with all host threads on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

CONFIG_FRAMES = {"C1": 16, "C2": 64, "C3": 256, "C4": 1024, "C5": 1024}
CONFIG_DESC = {
    "C1": "C1 default Cornell-box scene (19 spheres, 8 quads, cube) 512x512, 16 spp",
    "C2": "C2 bunny stand-in (69,432-triangle synthetic mesh) diffuse+glossy in the Cornell room, 1920x1080, 64 spp",
    "C3": "C3 dragon stand-in (298,116 triangles) dielectric, mixture importance sampling, 1920x1080, 256 spp",
    "C4": "C4 sponza stand-in (266,112-triangle atrium), camera inside, 1920x1080, 1024 spp",
    "C5": "C5 two 100K-triangle meshes, 3840x2160, NUM_SAMPLES=4 stratified x 1024 frames",
}


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
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def reference_counters(cfg_name):
    """Reference-algorithm work per sample on this scene (binary BVH visit counts), measured with the
    exact-mode kernel's counters on one frame: the constants behind B_ray / B_sample."""
    import bench_scenes
    from webgpu_path_tracer_b200 import PathTracer, MODE_EXACT
    scene, cam, W, H, kw = bench_scenes.build(cfg_name)
    pt = PathTracer(W, H, mode=MODE_EXACT, **kw)
    pt.write_scene(scene.buffers())
    pt.write_uniforms(1, 0, cam.viewMatrix)
    pt.set_counters(True)
    pt.render(1, 1)
    c = pt.counters()
    pt.close()
    n_sph, n_quad = len(scene.spheres), len(scene.quads)
    rays = max(1, c["rays"])
    b_ray = (48.0 * c["node_visits"] + 96.0 * c["tri_tests"]) / rays + 32.0 * n_sph + 80.0 * n_quad + 64.0
    rays_per_sample = c["rays"] / max(1, c["samples"])
    return {"rays_per_sample": rays_per_sample, "nodes_per_ray": c["node_visits"] / rays, "tris_per_ray": c["tri_tests"] / rays,
            "B_ray": b_ray, "B_sample": 32.0 + rays_per_sample * b_ray}


def run_reference(args):
    """CPU arm: the oracle (kind "port") with every host thread, bounded sample per step."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    import bench_scenes
    from oracle.binding import Oracle
    scene, cam, W, H, kw = bench_scenes.build(args.config)
    okw = {k: v for k, v in kw.items()}
    # every host thread the box has (torchrun exports OMP_NUM_THREADS=1 to its workers: do not inherit that)
    okw["threads"] = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
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
    sample_desc = "%d frame(s) of %dx%d per step (the full workload is %d frames; frames are i.i.d. work)" % (
        frames_per_step, W, H, args.frames)
    line = {"metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": CONFIG_DESC[args.config], "frames_per_step": frames_per_step},
            "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": orc.threads, "kind": "port", "sample": sample_desc},
            "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference WebGPU path unavailable on this box (no graphics stack / JS engine); this is the CPU oracle"}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=sorted(CONFIG_FRAMES))
    ap.add_argument("--frames", type=int, default=0, help="frames per step (default: the configuration's spp)")
    ap.add_argument("--mode", default="fast", choices=["fast", "exact"])
    ap.add_argument("--ref-frames", type=int, default=2, help="frames per step of the CPU arm")
    ap.add_argument("--cpu-baseline-frames", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if not args.frames:
        args.frames = CONFIG_FRAMES[args.config]
    if args.impl == "reference":
        return run_reference(args)

    rank, local_rank, world = dist_env()
    import torch
    import bench_scenes
    from webgpu_path_tracer_b200 import PathTracer, MODE_EXACT, MODE_FAST
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
    F = args.frames
    pt = PathTracer(W, H, mode=mode, device=local_rank, strata_rank=rank, strata_world=world, **kw)
    bufs = scene.buffers()
    pt.write_scene(bufs)
    pt.write_uniforms(1, 0, cam.viewMatrix)
    total_frames = F * world                      # frames of the whole job; this rank renders 1/world of them
    samples_per_step = W * H * spp * total_frames

    fb_ptr, fb_bytes = pt.framebuffer_device_ptr()

    class _FB:                                     # zero-copy view of binding 3 for NCCL
        __cuda_array_interface__ = {"shape": (W * H * 4,), "typestr": "<f4", "data": (fb_ptr, False), "version": 3}
    fb_t = torch.as_tensor(_FB(), device="cuda:%d" % local_rank) if world > 1 else None
    zeros_fb = np.zeros(W * H * 4, np.float32)
    l2_flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step():
        """inputs resident; progressive sum restarted by rendering into a cleared buffer"""
        pt.render(1, total_frames)
        if dist:
            pt.synchronize()                       # the reduce runs on torch's stream
            dist.reduce(fb_t, dst=0, op=dist.ReduceOp.SUM)

    # ---------------- device-timed value ----------------
    ref = reference_counters(args.config) if rank == 0 else None
    pt.set_stage_profiling(True)
    for _ in range(args.warmup):
        pt.write_buffer(3, zeros_fb)
        device_step()
    pt.stage_timings()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    step_ms, launches = [], 0
    t_wall0 = time.perf_counter()
    for _ in range(args.steps):
        pt.write_buffer(3, zeros_fb)
        l2_flush.zero_()                           # 256 MB > L2: nothing of the previous step stays cached
        barrier()
        if dist:
            ev0.record()
        device_step()
        ms, n = pt.last_timing()                   # CUDA events on the context's (launching) stream
        if dist:                                   # N > 1: the step includes the NCCL reduce on torch's stream
            ev1.record()
            torch.cuda.synchronize()
            ms = ev0.elapsed_time(ev1)
        step_ms.append(ms)
        launches += n
    barrier()
    wall = time.perf_counter() - t_wall0
    clocks = sampler.stop()
    stage = pt.stage_timings()
    # device time of the K timed steps: sum of per-step CUDA-event times on the context's stream
    t_local = torch.tensor([sum(step_ms)], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(t_local, op=dist.ReduceOp.MAX)
    total_ms = float(t_local.item())
    value = samples_per_step * args.steps / (total_ms * 1e-3) / 1e6

    # ---------------- end-to-end through the host-buffer API ----------------
    pinned = {b: torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for b, a in bufs.items()}
    pinned_np = {b: t.numpy() for b, t in pinned.items()}
    uni = np.zeros(20, np.float32)
    uni[0], uni[1], uni[2], uni[3] = W, H, 1, 0
    uni[4:] = np.asarray(cam.viewMatrix, np.float32)
    uni_p = torch.from_numpy(uni).pin_memory()
    h2d = sum(a.nbytes for a in bufs.values()) + 80
    d2h = W * H * 16
    e2e_times = []
    for it in range(2 + args.steps):
        barrier()
        t0 = time.perf_counter()
        e2e_pt = pt
        for b, a in pinned_np.items():
            e2e_pt.write_buffer(b, a)              # host -> device, re-derives the wide BVH (K2) each step
        e2e_pt.write_buffer(0, uni_p.numpy())
        e2e_pt.write_buffer(3, zeros_fb)
        device_step()
        out = e2e_pt.read_framebuffer() if rank == 0 or not dist else None   # device -> host result
        barrier()
        if it >= 2:
            e2e_times.append(time.perf_counter() - t0)
    e2e_t = torch.tensor([sum(e2e_times)], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_value = samples_per_step * args.steps / float(e2e_t.item()) / 1e6
    h2d += zeros_fb.nbytes

    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return

    # ---------------- roofline of the dominant kernel (extend = K3 traversal) ----------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    # Two kernels carry the step.  Algorithmic bytes:
    #   extend: B_ray per closest-hit query, defined on the REFERENCE's traversal;
    #   shade : B_shade = 160 B per path-bounce (the 5 x 16-byte state records of a slot, read once and
    #           written once) + 48 B per finished sample (pixel sum read, pixel sum / staging-ring write).
    # `roofline` is the kernel with the larger share of the step; the other one is reported beside it.
    traffic_all = {}
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        traffic_all = json.load(open(tp)).get(args.config, {})
    rays_per_step_rank = (W * H * spp * F) * ref["rays_per_sample"]
    samples_per_step_rank = W * H * spp * F

    def stage_roofline(kernel, key, alg_bytes_rank):
        ms, n = stage[key]
        if not n or ms <= 0:
            return None
        achieved = alg_bytes_rank / (ms * 1e-3) / 1e9
        return {"bound": "hbm", "kernel": kernel, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic_all.get(key + "_dram_bytes_per_launch"), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes_rank / n, "avg_launch_ms": ms / n, "launches": n,
                "share_of_step": ms / total_ms}

    B_SHADE_BOUNCE, B_SHADE_SAMPLE = 160.0, 48.0
    r_ext = stage_roofline("k_wf_extend", "extend", rays_per_step_rank * ref["B_ray"] * args.steps)
    r_shd = stage_roofline("k_wf_shade", "shade",
                           (rays_per_step_rank * B_SHADE_BOUNCE + samples_per_step_rank * B_SHADE_SAMPLE) * args.steps)
    both = [r for r in (r_ext, r_shd) if r]
    roofline = max(both, key=lambda r: r["share_of_step"]) if both else None
    if roofline:
        roofline = dict(roofline)
        roofline.update({"B_ray": ref["B_ray"], "B_sample": ref["B_sample"], "rays_per_sample": ref["rays_per_sample"],
                         "B_shade_per_bounce": B_SHADE_BOUNCE, "B_shade_per_sample": B_SHADE_SAMPLE,
                         "stage_ms": {k: v[0] for k, v in stage.items()},
                         "other_kernel": [r for r in both if r["kernel"] != roofline["kernel"]]})

    # ---------------- CPU baseline (oracle, bounded sample) ----------------
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        from oracle.binding import Oracle
        orc = Oracle(W, H, threads=len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1), **kw)
        orc.write_scene(bufs)
        orc.write_uniforms(1, 0, cam.viewMatrix)
        orc.render(1, 1)
        nfr = max(1, args.cpu_baseline_frames)
        t0 = time.perf_counter()
        orc.render(2, nfr)
        dt = time.perf_counter() - t0
        cpu = {"value": W * H * spp * nfr / dt / 1e6, "unit": "Msamples/s", "cores": orc.threads, "kind": "port",
               "sample": "%d frames of %dx%d (%.1f s of CPU work; the GPU step is %d frames)" % (nfr, W, H, dt, F)}

    line = {"metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": CONFIG_DESC[args.config], "frames_per_step_per_gpu": F, "mode": args.mode,
                       "parallelism": "frame-strata x%d + ncclReduce(sum)" % world if world > 1 else "single GPU",
                       "l2": "flushed between timed steps (256 MB device memset); the scene itself (<= 60 MB) is meant to live in L2",
                       "hierarchy": "device-built (K5 PLOC) wide BVH" if os.environ.get("WGPT_BVH", "d")[0] != "r" else "reference tree flattened"},
            "Mrays_per_s": value * ref["rays_per_sample"],
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches),
            "roofline": roofline, "cpu_baseline": cpu, "wall_s": wall,
            "fast_stats": pt.fast_statistics() if args.mode == "fast" else None}
    print(json.dumps(line))
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
