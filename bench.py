#!/usr/bin/env python3
"""bench.py — the reference's headline metric on B200: Mrays/s (and spp/s) of the per-pixel pathTracing() hot
path at 1920x1080 on the 1M-triangle Disney + HDR-importance-sampling + MIS scene (BASELINE.json config 4).

    python bench.py [--gpus N] [--steps K] [--warmup W]              # this repo's CUDA path (libpt.so)
    python bench.py --impl reference [...]                           # the reference's own CPU code for the path

A step = `--spp` frames (samples per pixel) of the whole 1080p image.  Prints ONE JSON line (rank 0).
  value            whole-job Mrays/s with the scene resident in HBM (all hitBVH calls: primary + shadow + bounce)
  e2e              the same through the C ABI from HOST buffers: per step the camera uniforms go host->device and
                   the fp32 framebuffer comes back to pinned host memory, inside the timed region
  roofline         the render kernel: algorithmic bytes (56 B per inner node visited, 36 B per
                   triangle tested, 108 B per shaded hit, 32 B per sample, 12 B per environment texel; the per-ray
                   counts come from the counting build of the same kernel on one frame of the whole image) / launch time
  cpu_baseline     the CPU oracle (OpenMP, all host cores) on a bounded crop of the same workload
N > 1 (torchrun, one rank per GPU): the scene is replicated, a step's frames are dealt round-robin to the ranks
(--split sample: per-rank sums, one launch per rank) or the image is split in diagonal 8x4 tiles (--split tile:
bit-identical to one GPU), and every step ends with an NCCL reduce of the framebuffers to rank 0 — total work
is fixed, so "scaling" is "strong".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (scene kind, size, width, height, hdr resolution)
    "is-1M-1080p": ("is", "1M", 1920, 1080, (2048, 1024)),          # BASELINE.json config 4 = the metric scene
    "disney-1M-1080p": ("disney", "1M", 1920, 1080, (1024, 512)),   # config 3
    "lambert-100k-1080p": ("lambert", "100k", 1920, 1080, (1024, 512)),  # config 2
    "full-10M-4k": ("full", "10M", 3840, 2160, (2048, 1024)),       # config 5 (reference tree from buildBVH, see build_workload)
    "is-1M-4k": ("is", "1M", 3840, 2160, (2048, 1024)),             # the metric scene at 4K (scheduler tuning)
    "is-1M-480p": ("is", "1M", 854, 480, (2048, 1024)),
    "is-100k-1080p": ("is", "100k", 1920, 1080, (2048, 1024)),      # quick variant for development
    "is-10k-480p": ("is", "10k", 854, 480, (512, 256)),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="is-1M-1080p", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=256, help="frames (samples per pixel) per step")
    ap.add_argument("--pipeline", type=int, default=2, help="0 megakernel, 1 wavefront, 2 persistent (production)")
    ap.add_argument("--split", default="sample", choices=["tile", "sample"],
                    help="N > 1: sample = frames dealt round-robin, per-rank sums reduced and scaled (default: every rank keeps the "
                         "whole image, so a tile's consecutive frames stay ~11 resident-warp generations apart, as on one GPU); "
                         "tile = diagonal 8x4 tiles, bit-identical to one GPU")
    ap.add_argument("--mode", type=int, default=0, help="0 fast (production), 1 parity-ref, 2 parity-wide")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-l2-flush", action="store_true")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """SM clock and throttle reasons of one GPU sampled every 5 ms through NVML while the timed region runs
    (nvidia-smi -lms as the fallback when the NVML binding is missing)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, device_index):
        super().__init__(daemon=True)
        self.idx, self.rows, self.proc, self.halt = device_index, [], None, threading.Event()
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # LOCAL_RANK indexes CUDA_VISIBLE_DEVICES when it is set
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[device_index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else device_index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def run(self):
        if self.nvml is not None:
            n = self.nvml
            bits = [getattr(n, "nvmlClocksEventReasonHwSlowdown", 0x8), getattr(n, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                    getattr(n, "nvmlClocksEventReasonSwThermalSlowdown", 0x20), getattr(n, "nvmlClocksEventReasonSwPowerCap", 0x4)]
            while not self.halt.is_set():
                try:
                    sm = float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
                    try:
                        mask = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
                    except Exception:
                        mask = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
                    self.rows.append([str(sm), str(self.max_sm), "0"] + ["Active" if mask & b else "Not Active" for b in bits])
                except Exception:
                    pass
                self.halt.wait(0.005)
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        self.halt.set()
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for n, v in zip(self.NAMES, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except (ValueError, IndexError):
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm),
                "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def build_workload(name):
    from opengl_ray_tracing_b200 import scenes
    kind, size, W, H, hdr_res = WORKLOADS[name]
    t0 = time.time()
    # 10 M triangles: the exact restatement of buildBVHwithSAH needs hours of host time (O(n log^2 n) with the reference's
    # tie order); the reference's other builder, buildBVH (M:787, commented alternative), takes 20 s and is used instead
    sc = scenes.build(kind, size, hdr_res=hdr_res, builder="median" if size == "10M" else "sah")
    return sc, W, H, time.time() - t0


def crop_region(W, H, frac=4):
    """Centre crop used for the bounded CPU samples: (W/frac) x (H/frac) pixels around the image centre."""
    cw, ch = max(8, W // frac), max(8, H // frac)
    x0, y0 = (W - cw) // 2, (H - ch) // 2
    return x0, y0, x0 + cw, y0 + ch


def basic_tracer_baseline():
    """The reference's CPU tracer (BasicRayTracingWithC++/main.cpp compiled by oracle/build_ref.py with g++ -O3 -fopenmp,
    SAMPLE = 200: BASELINE.json config 1), wall-clock on this box's host cores.  None if oracle/_ref was not built."""
    exe = os.path.join(ROOT, "oracle", "_ref", "basic_tracer_200spp")
    if not os.path.exists(exe):
        return None
    import tempfile
    with tempfile.TemporaryDirectory() as tmp:
        t0 = time.perf_counter()
        r = subprocess.run([exe], cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
        dt = time.perf_counter() - t0
        ok = r.returncode == 0 and os.path.exists(os.path.join(tmp, "basic_ray_tracing_with_c++.png"))
    if not ok:
        return None
    return {"value": 256 * 256 * 200 / dt / 1e6, "unit": "Mpaths/s", "seconds": dt, "cores": os.cpu_count(), "kind": "reference",
            "sample": "BasicRayTracingWithC++/main.cpp unmodified but SAMPLE=200, 256x256 Cornell box, 50 OpenMP threads as the "
                      "source requests, wall clock incl. PNG write"}


def algorithmic_bytes(c):
    """bytes the best algorithm on this BVH must read, from the oracle's culled-traversal counts."""
    return 56 * c["cull_inner"] + 36 * c["cull_tris"] + 108 * c["cull_hits"] + 12 * c["texels"]


def cpu_sample(sc, W, H, math=1, frames=1, counters=True):
    """Times the oracle on the centre crop (all host threads); returns (seconds, counters dict, pixels)."""
    from oracle import pyoracle as O
    reg = crop_region(W, H)
    t0 = time.perf_counter()
    _, _, cnt = O.render(sc.variant, sc.max_bounce, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W, H, sc.eye, sc.cam, 0, frames,
                         math=math, counters=counters, region=reg)
    dt = time.perf_counter() - t0
    return dt, cnt, (reg[2] - reg[0]) * (reg[3] - reg[1]), reg


# ---------------------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's own CPU implementation of the path, on the box's host cores, rank 0 only.
    kind = "reference": oracle/_ref/libref_shader_<v>_libm.so — the reference's pass1.fsh compiled as C++ by
    oracle/build_ref.py (prebuilt in the container that has /root/reference); else kind = "port": the oracle."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        # torchrun exports OMP_NUM_THREADS=1 to its workers; this arm is rank 0 alone on ALL host cores (the OpenMP
        # runtime reads the variable when the oracle's library is loaded, below)
        os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    from oracle import pyoracle as O
    sc, W, H, _ = build_workload(args.workload)
    reg = crop_region(W, H)
    npx = (reg[2] - reg[0]) * (reg[3] - reg[1])
    frames = 1
    kind = "reference" if O.ref_shader_available(sc.variant, "libm") else "port"
    # rays in one step of the sample (identical for the reference build and the oracle: same algorithm, libm)
    _, _, cnt = O.render(sc.variant, sc.max_bounce, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W, H, sc.eye, sc.cam, 0, frames,
                         math=0, counters=True, region=reg)

    def step(first):
        if kind == "reference":
            O.ref_shader_render(sc.variant, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W, H, sc.eye, sc.cam, first, frames,
                                math="libm", region=reg)
        else:
            O.render(sc.variant, sc.max_bounce, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W, H, sc.eye, sc.cam, first, frames,
                     math=0, region=reg)

    for w in range(args.warmup):
        step(w)
    t0 = time.perf_counter()
    for k in range(args.steps):
        step(args.warmup + k)
    dt = time.perf_counter() - t0
    ms = dt / args.steps * 1e3
    mrays = cnt["rays"] / (ms * 1e-3) / 1e6
    cores = O.num_threads()
    sample = "centre crop %dx%d of %dx%d, %d frame(s) per step, OpenMP over pixels" % (reg[2] - reg[0], reg[3] - reg[1], W, H, frames)
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "triangles": sc.nTriangles, "nodes": sc.nNodes, "sample": sample},
        "spp_per_s": npx * frames / (ms * 1e-3) / (W * H),
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from opengl_ray_tracing_b200 import renderer as R, dist as D

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — libpt.so has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")      # stdout carries the one JSON line, not NCCL's version banner
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # the host BVH build (the reference's SAH builder, all cores) runs once: rank 0 builds and caches, the others load
    if world > 1 and rank != 0:
        dist.barrier()
    sc, W, H, t_scene = build_workload(args.workload)
    if world > 1 and rank == 0:
        dist.barrier()

    pt = R.PathTracer(W, H, sc.variant, device=local, mode=args.mode, pipeline=args.pipeline)
    t0 = time.perf_counter()
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
    t_upload = time.perf_counter() - t0
    pt.set_camera(sc.eye, sc.cam)
    stream = torch.cuda.ExternalStream(pt.stream(), device=local)
    fb = D.DeviceFramebuffer(pt).tensor(torch.device("cuda", local))
    combined = torch.empty_like(fb) if world > 1 else None
    flush = None if args.no_l2_flush else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    tile, samp = D.TilePlan(W, H, world, rank), D.SamplePlan(world, rank)
    spp = args.spp

    def render_step(first):
        if world == 1:
            pt.render_spp(first, spp)
        elif args.split == "tile":
            D.render_tiles(pt, tile, first, spp)
        else:
            D.render_samples(pt, samp, first, spp)
        if world > 1:
            with torch.cuda.stream(stream):
                D.combine(fb, out=combined, dst=0, scale=(1.0 / (first + spp)) if args.split == "sample" else None)

    def barrier():
        pt.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    frame = 0
    for _ in range(args.warmup):
        render_step(frame); frame += spp
    barrier()
    pt.reset_stats()

    # ---- device-timed region: K steps, CUDA events on the launching stream, L2 flushed between steps ----
    sampler = ClockSampler(local)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    t_wall0 = time.perf_counter()
    for k in range(args.steps):
        if flush is not None:
            with torch.cuda.stream(stream):
                flush.fill_(k & 0xff)                 # 256 MiB > 126 MB L2, outside the per-step events
        ev[k][0].record(stream)
        render_step(frame); frame += spp
        ev[k][1].record(stream)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop()
    st = pt.stats()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    tot = torch.tensor([dev_ms, float(st["rays"]), float(st["samples"]), float(st["launches"]), float(st["render_ms"]),
                        float(clocks["sm_mhz"] or 0.0), float(len(clocks["reasons"]))],
                       dtype=torch.float64, device="cuda")
    if world > 1:
        mx = tot.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tot.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dev_ms, rays, samples, launches, kern_ms = float(mx[0]), float(sm[1]), float(sm[2]), float(sm[3]), float(mx[4])
        allk = [torch.zeros_like(tot) for _ in range(world)]
        dist.all_gather(allk, tot)
        per_rank_kernel_ms = [float(t[4]) / args.steps for t in allk]
        clocks["per_rank_sm_mhz"] = [float(t[5]) for t in allk]          # every rank samples its own GPU
        clocks["ranks_with_reasons"] = int(sum(1 for t in allk if float(t[6]) > 0))
    else:
        dev_ms, rays, samples, launches, kern_ms = [float(x) for x in tot[:5]]
        per_rank_kernel_ms = [kern_ms / args.steps]
    ms_per_step = dev_ms / args.steps
    mrays = rays / (dev_ms * 1e-3) / 1e6
    spp_per_s = samples / (W * H) / (dev_ms * 1e-3)

    # ---- end-to-end through the C ABI with host buffers: camera H2D + render + framebuffer D2H per step ----
    host = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
    eye_h, cam_h = np.array(sc.eye, np.float32), np.array(sc.cam, np.float32)
    pt.reset_stats()
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        pt.set_camera(eye_h, cam_h)
        render_step(frame); frame += spp
        if world > 1:
            if rank == 0:                              # the job's result is the combined image on rank 0
                # copied on torch's OWN stream (ordered after the reduce): a pinned tensor remembers the streams that used
                # it, and the context's stream (an ExternalStream here) is destroyed by pt.close() before torch lets go of it
                cur = torch.cuda.current_stream()
                cur.wait_stream(stream)
                host.copy_(combined, non_blocking=True)
                cur.synchronize()
            else:
                stream.synchronize()
        else:
            pt.readback_into(host.data_ptr(), host.numel() * 4)
    barrier()
    t_e2e = time.perf_counter() - t0
    st2 = pt.stats()
    e2 = torch.tensor([t_e2e, float(st2["rays"])], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = e2.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = e2.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        t_e2e, rays_e2e = float(mx[0]), float(sm[1])
    else:
        rays_e2e = float(e2[1])
    e2e_mrays = rays_e2e / t_e2e / 1e6

    if rank == 0:
        peak, peak_src = load_peaks()
        line = {
            "metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "triangles": sc.nTriangles, "nodes": sc.nNodes, "variant": sc.variant,
                       "max_bounce": sc.max_bounce, "width": W, "height": H, "spp_per_step": spp, "mode": args.mode,
                       "pipeline": {0: "megakernel", 1: "wavefront", 2: "persistent"}[args.pipeline],
                       "split": args.split if world > 1 else "none",
                       "l2": "flushed between steps (256 MiB fill)" if flush is not None else "not flushed",
                       "scene_bytes_hbm": int(sc.tris.nbytes + sc.nodes.nbytes + sc.nTriangles * 96 + st["n_inner_nodes"] * 64
                                              + sc.hdr.shape[0] * sc.hdr.shape[1] * 16 * (4 if sc.hdrCache is not None else 1))},   # hdr4 (+ cache4 + 32 B light table)
            "spp_per_s": spp_per_s, "rays_per_sample": rays / max(1.0, samples),
            "e2e": {"value": e2e_mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 76, "d2h_bytes_per_step": int(host.numel() * 4),
                    "ms_per_step": t_e2e / args.steps * 1e3},
            "gpu_launches": int(launches), "clocks": clocks,
            "scene_build_s": t_scene, "scene_upload_s": t_upload, "wall_s_timed_region": t_wall,
            "kernel_ms_per_step": kern_ms / args.steps, "per_rank_kernel_ms": per_rank_kernel_ms,
        }
        # roofline of the render kernel: algorithmic bytes from the kernel's OWN traversal counts (one frame of the
        # whole image through the counting build of the same kernel, outside the timed region), SURVEY.md §8(d)
        # coefficients: 56 B per inner node visited, 36 B per triangle tested, 108 B per shaded hit, 12 B per
        # environment texel, 32 B per sample (accumulator read + write).
        cpt = R.PathTracer(W, H, sc.variant, device=local, mode=R.MODE_FAST_COUNT, pipeline=R.PERSISTENT)
        cpt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
        cpt.set_camera(sc.eye, sc.cam)
        if world > 1:
            cpt.set_band_interleave(world, rank)
        cpt.render_spp(0, 1)
        cs = cpt.stats()
        cpt.close()
        bytes_per_ray = (56 * cs["inner_visits"] + 36 * cs["tri_tests"] + 108 * cs["shaded_hits"] + 12 * cs["env_texels"]
                         + 32 * cs["samples"]) / max(1, cs["rays"])
        achieved = (rays / world) * bytes_per_ray / (kern_ms * 1e-3) / 1e9          # per GPU, over kernel time
        traffic = issue = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            with open(tpath) as f:
                tj = json.load(f).get(args.workload, {})
            traffic, issue = tj.get("dram_bytes_per_launch"), tj.get("issue")
            if world > 1 or spp != 256:
                traffic = issue = None                 # the committed capture is ONE 256-frame launch of the whole image on one GPU
        line["roofline"] = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                            "traffic": traffic, "peak_source": peak_src, "bytes_per_ray": bytes_per_ray,
                            "inner_per_ray": cs["inner_visits"] / max(1, cs["rays"]), "tris_per_ray": cs["tri_tests"] / max(1, cs["rays"]),
                            "shaded_hits_per_ray": cs["shaded_hits"] / max(1, cs["rays"]), "texels_per_ray": cs["env_texels"] / max(1, cs["rays"]),
                            "kernel": "render_persistent<CfgFast,%d>" % sc.variant, "kernel_ms_per_launch": kern_ms / args.steps,      # one render launch per step and rank (max over ranks)
                            "issue": issue,
                            "note": "the path is SM-issue/latency bound with an L2-resident working set (DESIGN.md); "
                                    "traffic = ncu dram bytes of one launch of the committed capture (profiles/)"}
        if not args.no_cpu_baseline and world == 1:
            from oracle import pyoracle as O
            dt, _, npx, reg = cpu_sample(sc, W, H, math=1, frames=2, counters=False)
            _, cnt, _, _ = cpu_sample(sc, W, H, math=1, frames=1, counters=True)
            cpu_rays = cnt["rays"] * 2          # 2 frames timed; per-frame ray counts differ by < 1 %
            line["cpu_baseline"] = {"value": cpu_rays / dt / 1e6, "unit": "Mrays/s", "cores": O.num_threads(), "kind": "port",
                                    "sample": "oracle (-O2 -fopenmp), centre crop %dx%d of %dx%d, 2 frames"
                                              % (reg[2] - reg[0], reg[3] - reg[1], W, H)}
            basic = basic_tracer_baseline()
            if basic is not None:
                line["cpu_baseline_basic"] = basic
            # BASELINE.json config 1 on the GPU: the same tracer (pt_basic_render), same 256x256 x 200 samples
            from opengl_ray_tracing_b200 import basic as PB
            shapes = PB.default_scene()
            PB.render(shapes, 256, 256, 8)                                     # warm-up
            t0 = time.perf_counter()
            _, ms = PB.render(shapes, 256, 256, 200)
            wall = time.perf_counter() - t0
            line["basic_gpu"] = {"value": 256 * 256 * 200 / wall / 1e6, "unit": "Mpaths/s", "seconds": wall, "device_ms": ms,
                                 "sample": "pt_basic_render, 256x256 Cornell box, 200 samples, wall clock incl. the image copy back"}
        print(json.dumps(line), flush=True)
    # every torch object that touched the context's stream or memory goes before the context does
    del host, fb, combined, flush
    torch.cuda.synchronize()
    pt.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
