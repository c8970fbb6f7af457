#!/usr/bin/env python3
"""bench.py — the reference's headline metric on B200: Mrays/s (and spp/s) of the per-pixel pathTracing() hot path at
1920x1080 on the 1M-triangle Disney + HDR-importance-sampling + MIS scene (BASELINE.json config 4).

    python bench.py [--gpus N] [--steps K] [--warmup W]              # this repo's CUDA path (libpt.so)
    python bench.py --impl reference [...]                           # the reference's own CPU code for the path

A step = `--spp` frames (samples per pixel) of the whole 1080p image.  Prints ONE JSON line (rank 0).
  value            whole-job Mrays/s with the scene resident in HBM (all hitBVH calls: primary + shadow + bounce)
  e2e              the same through the C ABI from HOST buffers: per step the camera uniforms go host->device and
                   the fp32 framebuffer comes back to pinned host memory, inside the timed region
  roofline         the render kernel against the ceiling that binds it — SM issue slots: warp instructions per launch (ncu
                   capture under profiles/, quoted only when the capture's kernel SASS hash equals this build's) / the
                   launch time measured here / (SMs x 4 x max SM clock); the SURVEY.md 8d byte figure rides along as `hbm`
                   (56 B per inner node visited, 36 B per triangle tested, 108 B per shaded hit, 12 B per environment
                   texel, 32 B per sample; per-ray counts from the counting build of the same kernel), and `l2` puts the
                   capture's L1-miss read sectors against the L2 read peak measured live (pt_measure_l2_read_gbs)
  parity           the centre crop the CPU baseline renders, rendered by the GPU for the same frames: ids equal, rRMSE
  cpu_baseline     the CPU oracle (OpenMP, all host cores) on a bounded crop of the same workload
  crop_like_reference   the GPU on exactly the reference arm's sample (same crop, one frame per step): like for like
  e2e_interactive  the reference's display() loop (IS/main.cpp:659-709): one sample per call + the 8-bit display image back
  other_workloads  BASELINE.json configs 2, 3 and 5: three steps each
N > 1 (torchrun, one rank per GPU): the scene is replicated, a step's frames are dealt round-robin to the ranks
(--split sample: per-rank sums, one launch per rank) or the image is split in diagonal 8x4 tiles (--split tile:
bit-identical to one GPU), and every step ends with the fused NVLink combine of the C ABI (pt_combine) onto rank 0 — total
work is fixed, so "scaling" is "strong".  torch.distributed only carries the 128-byte NCCL id and the barriers.
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
OTHER_WORKLOADS = [("lambert-100k-1080p", 64), ("disney-1M-1080p", 64), ("full-10M-4k", 16)]   # (name, frames per step)


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
                    help="N > 1: sample = frames dealt round-robin, per-rank sums combined and scaled (default: every rank keeps the "
                         "whole image, so a tile's consecutive frames stay ~11 resident-warp generations apart, as on one GPU); "
                         "tile = diagonal 8x4 tiles, bit-identical to one GPU")
    ap.add_argument("--combine", default="pipelined", choices=["pipelined", "sync"],
                    help="N > 1 with peer memory: pipelined = PT_COMBINE_PIPELINED (the combine of step k runs on a second stream under "
                         "the rendering of step k + 1; the timed region is then ONE event pair around all K steps, L2 flushes "
                         "included); sync = two NCCL barriers and the combine kernel on the render stream in every step")
    ap.add_argument("--mode", type=int, default=0, help="0 fast (production), 1 parity-ref, 2 parity-wide")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-l2-flush", action="store_true")
    ap.add_argument("--no-other-workloads", action="store_true")
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
            j = json.load(f)
        return float(j["hbm_gbs"]), float(j.get("sm_max_mhz", 1965.0)), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, 1965.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


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


def cpu_sample(sc, W, H, math=1, frames=1, counters=True):
    """Times the oracle on the centre crop (all host threads); returns (seconds, counters dict, pixels, region, image, ids)."""
    from oracle import pyoracle as O
    reg = crop_region(W, H)
    t0 = time.perf_counter()
    img, ids, cnt = O.render(sc.variant, sc.max_bounce, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W, H, sc.eye, sc.cam, 0, frames,
                             math=math, counters=counters, region=reg)
    dt = time.perf_counter() - t0
    return dt, cnt, (reg[2] - reg[0]) * (reg[3] - reg[1]), reg, img, ids


def rrmse(a, b):
    a = np.asarray(a)[..., :3].astype(np.float64)
    b = np.asarray(b)[..., :3].astype(np.float64)
    return float(np.sqrt(((a - b) ** 2).sum() / max(1e-300, (b ** 2).sum())))


# ---------------------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's own CPU implementation of the path, on the box's host cores, rank 0 only.
    kind = "reference": oracle/_ref/libref_shader_<v>_libmO3.so — the reference's pass1.fsh compiled as C++ by
    oracle/build_ref.py (g++ -O3 -fopenmp, the timing build of SURVEY.md 8d; prebuilt in the container that has
    /root/reference); else kind = "port": the oracle."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        # torchrun exports OMP_NUM_THREADS=1 to its workers; this arm is rank 0 alone on ALL host cores (the OpenMP
        # runtime reads the variable when the oracle's library is loaded, below)
        os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    from oracle import pyoracle as O
    # the scene arrays come from libptscene's builders: bit-identical to the reference's own buildBVHwithSAH / calculateHdrCache
    # compiled from /root/reference (tests/test_scene.py), which need minutes single-threaded at this size
    sc, W, H, _ = build_workload(args.workload)
    reg = crop_region(W, H)
    npx = (reg[2] - reg[0]) * (reg[3] - reg[1])
    frames = 1
    flavour = "libmO3" if O.ref_shader_available(sc.variant, "libmO3") else "libm"
    kind = "reference" if O.ref_shader_available(sc.variant, flavour) else "port"
    # rays in one step of the sample (identical for the reference build and the oracle: same algorithm, libm)
    _, _, cnt = O.render(sc.variant, sc.max_bounce, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W, H, sc.eye, sc.cam, 0, frames,
                         math=0, counters=True, region=reg)

    def step(first):
        if kind == "reference":
            O.ref_shader_render(sc.variant, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W, H, sc.eye, sc.cam, first, frames,
                                math=flavour, region=reg)
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
    cores = O.ref_shader_threads(sc.variant, flavour) if kind == "reference" else O.num_threads()
    sample = "centre crop %dx%d of %dx%d, %d frame(s) per step, OpenMP over runs of 16 pixels, %s" % (
        reg[2] - reg[0], reg[3] - reg[1], W, H, frames,
        "pass1.fsh as C++, g++ -O3 -fopenmp, libm" if flavour == "libmO3" else "g++ -O2 -ffp-contract=off")
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "triangles": sc.nTriangles, "nodes": sc.nNodes, "sample": sample,
                   "host_cpus": os.cpu_count(), "omp_threads": cores},
        "spp_per_s": npx * frames / (ms * 1e-3) / (W * H),
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------
def capture_for(workload, world, spp):
    """The committed ncu capture of this workload's render launch (profiles/traffic.json), or None when it describes another
    build of the kernels (SASS hash), another launch shape, or does not exist."""
    from opengl_ray_tracing_b200 import renderer as R
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(path):
        return None, "no profiles/traffic.json"
    with open(path) as f:
        tj = json.load(f).get(workload)
    if not tj:
        return None, "no capture of this workload"
    if world > 1 or spp != tj.get("frames_per_launch", 256):
        return None, "the capture is one %d-frame launch of the whole image on one GPU" % tj.get("frames_per_launch", 256)
    have = R.kernel_sass_hash()
    if tj.get("sass_hash") != have:
        return None, "capture taken from kernels %s, this build is %s" % (tj.get("sass_hash"), have)
    return tj, "profiles/traffic.json (kernel SASS hash %s)" % have


def measure_workload(R, name, frames, steps, device, world=1, rank=0):
    """Three (steps) timed launches of another BASELINE.json config on this rank's GPU (tile split when world > 1)."""
    sc, W, H, t_scene = build_workload(name)
    pt = R.PathTracer(W, H, sc.variant, device=device, mode=R.MODE_FAST, pipeline=R.PERSISTENT)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
    pt.set_camera(sc.eye, sc.cam)
    if world > 1:
        pt.set_tile_interleave(world, rank)
    pt.render_spp(0, frames); pt.synchronize(); pt.reset_stats()
    for k in range(steps):
        pt.render_spp(frames * (k + 1), frames)
    st = pt.stats()
    pt.close()
    return {"rays": st["rays"], "ms": st["render_ms"], "samples": st["samples"], "W": W, "H": H, "triangles": sc.nTriangles,
            "variant": sc.variant, "scene_build_s": t_scene}


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
    peer = None
    if world > 1:
        # the C ABI's own communicator (pt_comm.cu): torch.distributed only hands the 128-byte NCCL id around
        idt = torch.zeros(R.COMM_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(R.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, src=0)
        pt.comm_init(world, rank, bytes(idt.cpu().numpy().tobytes()))
        peer = pt.comm_info()["peer_memory"]
    # 256 MiB > the 126 MB L2, written as 8-byte words (torch's one-byte fill_ runs at 3.7 TB/s, this at 6.8: tools/fill_probe.py)
    flush = None if args.no_l2_flush else torch.empty((256 << 20) // 8, dtype=torch.int64, device="cuda")
    tile, samp = D.TilePlan(W, H, world, rank), D.SamplePlan(world, rank)
    spp = args.spp
    kind = R.COMBINE_DISJOINT if args.split == "tile" else R.COMBINE_SUM
    pipelined = world > 1 and bool(peer) and args.combine == "pipelined"

    def render_step(first, host_ptr=None, overlap=False):
        """One step: this rank's share of frames first .. first+spp-1, then (N > 1) the combine onto rank 0.
        overlap: rank 0's host copy of the combined image rides on a second stream (PT_COMBINE_ASYNC) and lands at the next
        pt_synchronize, overlapping the next step's rendering."""
        if world == 1:
            pt.render_spp(first, spp)
            return
        if args.split == "tile":
            D.render_tiles(pt, tile, first, spp)
        else:
            D.render_samples(pt, samp, first, spp)           # per-rank sums of ALL frames so far (PT_ACC_SUM)
        how = R.COMBINE_PIPELINED if pipelined else (R.COMBINE_ASYNC if overlap else 0)
        pt.combine(kind | how, first + spp, host_ptr if rank == 0 else None, None)

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
    ev_all = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
    barrier()
    t_wall0 = time.perf_counter()
    ev_all[0].record(stream)
    for k in range(args.steps):
        if flush is not None:
            with torch.cuda.stream(stream):
                flush.fill_(k)                 # 256 MiB > 126 MB L2, outside the per-step events
        ev[k][0].record(stream)
        render_step(frame); frame += spp
        ev[k][1].record(stream)
    if pipelined:
        pt.combine_fence()                            # the render stream waits for the last combine, on every rank
    ev_all[1].record(stream)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop()
    st = pt.stats()
    # sync combine / one GPU: the per-step event pairs (the L2 flushes lie between them).  Pipelined combine: the combine of
    # step k runs beside the flush and the render of step k + 1, so only ONE pair around the whole loop — flushes included,
    # closed after the fence — counts everything
    dev_ms = ev_all[0].elapsed_time(ev_all[1]) if pipelined else sum(a.elapsed_time(b) for a, b in ev)
    render_stream_ms = sum(a.elapsed_time(b) for a, b in ev) / args.steps
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

    # ---- end-to-end through the C ABI with host buffers: camera H2D + render (+ combine) + framebuffer D2H per step ----
    host = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
    eye_h, cam_h = np.array(sc.eye, np.float32), np.array(sc.cam, np.float32)

    def e2e_step(first):
        # camera uniforms in, render (+ combine), the step's image out to pinned host memory.  The device-to-host copy of step k
        # is asynchronous (pt_readback_async / PT_COMBINE_ASYNC: a snapshot copied on a second stream) and overlaps the
        # rendering of step k + 1; every copy completes inside the timed region (the closing barrier synchronises).
        pt.set_camera(eye_h, cam_h)
        if world > 1:
            render_step(first, host.data_ptr(), overlap=True)
        else:
            render_step(first)
            pt.readback_async(host.data_ptr(), host.numel() * 4)

    e2e_step(frame); frame += spp                    # one untimed warm-up of this loop (first use of the pinned buffer by the copy engine)
    barrier()
    pt.reset_stats()
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        e2e_step(frame); frame += spp
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

    # ---- parity of what was just timed (outside the timed regions) ----
    parity = {}
    if world > 1:
        # the combined image of one more step from a cleared framebuffer, against rank 0 rendering the same frames alone
        pt.clear()
        render_step(0, host.data_ptr())
        barrier()
        if rank == 0:
            combined = host.numpy().copy()
            solo = R.PathTracer(W, H, sc.variant, device=local, mode=args.mode, pipeline=args.pipeline)
            solo.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
            solo.set_camera(sc.eye, sc.cam)
            solo.render_spp(0, spp)
            alone = solo.readback()
            solo.close()
            parity["multi_gpu"] = {"split": args.split, "rrmse_vs_one_gpu": rrmse(combined, alone),
                                   "bit_identical": bool(combined.tobytes() == alone.tobytes()),
                                   "bar": "bit-identical" if args.split == "tile" else "rRMSE <= 1e-5 (sum order instead of the running mean)",
                                   "ok": bool(combined.tobytes() == alone.tobytes()) if args.split == "tile" else bool(rrmse(combined, alone) <= 1e-5)}
        barrier()

    if rank == 0:
        peak, sm_max_mhz, peak_src = load_peaks()
        line = {
            "metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "triangles": sc.nTriangles, "nodes": sc.nNodes, "variant": sc.variant,
                       "max_bounce": sc.max_bounce, "width": W, "height": H, "spp_per_step": spp, "mode": args.mode,
                       "pipeline": {0: "megakernel", 1: "wavefront", 2: "persistent"}[args.pipeline],
                       "split": args.split if world > 1 else "none",
                       "combine": (("pt_combine | PT_COMBINE_PIPELINED: snapshot + flag on the render stream, one small kernel per GPU on a second stream "
                                    "(in-kernel flag waits over NVLink peer memory) under the next step's render; timed as ONE event pair around "
                                    "all steps incl. the L2 flushes" if pipelined else
                                    "pt_combine: one kernel per GPU over NVLink peer memory between two NCCL barriers") if peer
                                   else "pt_combine: ncclReduce + consume kernel (no peer memory)") if world > 1 else "none",
                       "hdr": "synthHdr(%d, %d, seed 1) — the reference's .hdr files are not on the GPU box" % (sc.hdr.shape[1], sc.hdr.shape[0]),
                       "l2": "flushed between steps (256 MiB fill)" if flush is not None else "not flushed",
                       "scene_bytes_hbm": int(sc.tris.nbytes + sc.nodes.nbytes + sc.nTriangles * 96 + st["n_inner_nodes"] * 64
                                              + sc.hdr.shape[0] * sc.hdr.shape[1] * 16 * (4 if sc.hdrCache is not None else 1))},   # hdr4 (+ cache4 + 32 B light table)
            "spp_per_s": spp_per_s, "rays_per_sample": rays / max(1.0, samples),
            "e2e": {"value": e2e_mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 76, "d2h_bytes_per_step": int(host.numel() * 4),
                    "ms_per_step": t_e2e / args.steps * 1e3,
                    "readback": "every step's fp32 image to pinned host memory, asynchronously: the copy of step k overlaps the render of step k+1"},
            "gpu_launches": int(launches), "clocks": clocks,
            "scene_build_s": t_scene, "scene_upload_s": t_upload, "wall_s_timed_region": t_wall,
            "kernel_ms_per_step": kern_ms / args.steps, "per_rank_kernel_ms": per_rank_kernel_ms,
            "render_stream_ms_per_step_rank0": render_stream_ms,
            "kernel_sass_hash": R.kernel_sass_hash(),
        }
        # ---- roofline of the render kernel ----
        # algorithmic bytes from the kernel's OWN traversal counts (one frame of the whole image through the counting build
        # of the same kernel on ITS tree — conservative against SURVEY.md 8d's "oracle's culled traversal of the reference
        # BVH", which visits more), outside the timed region
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
        kernel_s = kern_ms * 1e-3 / args.steps                                        # one render launch per step and rank (max over ranks)
        rays_per_launch = rays / world / args.steps
        achieved_hbm = rays_per_launch * bytes_per_ray / kernel_s / 1e9
        l2_gbs, hbm_probe_gbs = R.measure_l2_read_gbs(local, 48, 128)
        cap, cap_src = capture_for(args.workload, world, spp)
        issue_peak = 148 * 4 * sm_max_mhz * 1e6                                       # warp instructions per second
        hbm_obj = {"bound": "hbm", "achieved": achieved_hbm, "peak": peak, "unit": "GB/s", "frac": achieved_hbm / peak,
                   "traffic": cap.get("dram_bytes_per_launch") if cap else None, "peak_source": peak_src, "bytes_per_ray": bytes_per_ray,
                   "inner_per_ray": cs["inner_visits"] / max(1, cs["rays"]), "tris_per_ray": cs["tri_tests"] / max(1, cs["rays"]),
                   "shaded_hits_per_ray": cs["shaded_hits"] / max(1, cs["rays"]), "texels_per_ray": cs["env_texels"] / max(1, cs["rays"]),
                   "note": "SURVEY.md 8d byte formula; measured DRAM traffic is well below it (L2-resident working set): not the binding ceiling"}
        if cap:
            wi = float(cap["warp_instructions_per_launch"])
            l2_bytes = float(cap["l2_tex_read_sectors_per_launch"]) * 32.0
            line["roofline"] = {
                "bound": "issue", "achieved": wi / kernel_s / 1e9, "peak": issue_peak / 1e9, "unit": "Gwarp-instr/s",
                "frac": wi / kernel_s / issue_peak, "traffic": cap.get("dram_bytes_per_launch"),
                "lanes": cap.get("threads_per_warp_instruction"), "frac_x_lanes": wi / kernel_s / issue_peak * float(cap.get("threads_per_warp_instruction", 32.0)) / 32.0,
                "warp_instructions_per_launch": wi, "warp_instructions_per_ray": wi / max(1.0, rays_per_launch),
                "peak_source": "148 SMs x 4 schedulers x %.0f MHz (MEASURED_PEAKS.json sm_max_mhz)" % sm_max_mhz,
                "kernel": "render_persistent<CfgFast,%d>" % sc.variant, "kernel_ms_per_launch": kernel_s * 1e3,
                "capture": cap_src,
                "l2": {"achieved": l2_bytes / kernel_s / 1e9, "peak": l2_gbs, "unit": "GB/s", "frac": l2_bytes / kernel_s / 1e9 / l2_gbs,
                       "what": "lts__t_sectors_srcunit_tex_op_read x 32 B (node, triangle, texel and framebuffer fetches that miss L1) / launch time, "
                               "against pt_measure_l2_read_gbs (48 MiB ld.global.nc window, measured in this run)"},
                "hbm": hbm_obj,
                "note": "achieved = warp instructions of the committed ncu capture of THIS kernel build / the launch time measured here"}
        else:
            hbm_obj.update({"kernel": "render_persistent<CfgFast,%d>" % sc.variant, "kernel_ms_per_launch": kernel_s * 1e3,
                            "l2_peak_gbs": l2_gbs, "no_issue_figure": cap_src})
            line["roofline"] = hbm_obj
        # what the traversal itself requests (before any cache): 64 B per inner node visited, 48 B per triangle tested
        node_gbs = rays_per_launch * cs["inner_visits"] / max(1, cs["rays"]) * 64.0 / kernel_s / 1e9
        tri_gbs = rays_per_launch * cs["tri_tests"] / max(1, cs["rays"]) * 48.0 / kernel_s / 1e9
        line["roofline"]["fetches"] = {"node_gbs": node_gbs, "triangle_gbs": tri_gbs, "unit": "GB/s requested by the traversal loop",
                                       "of_l2_read_peak": (node_gbs + tri_gbs) / l2_gbs, "of_hbm_peak": (node_gbs + tri_gbs) / peak,
                                       "note": "an upper bound of what reaches L2: about half of these requests hit L1 (capture)"}
        line["roofline"]["l2_read_peak_gbs"] = l2_gbs
        line["roofline"]["hbm_read_probe_gbs"] = hbm_probe_gbs
        line["parity"] = parity

        if not args.no_cpu_baseline and world == 1:
            from oracle import pyoracle as O
            dt, _, npx, reg, ref_img, ref_ids = cpu_sample(sc, W, H, math=1, frames=2, counters=False)
            _, cnt, _, _, _, _ = cpu_sample(sc, W, H, math=1, frames=1, counters=True)
            cpu_rays = cnt["rays"] * 2          # 2 frames timed; per-frame ray counts differ by < 1 %
            line["cpu_baseline"] = {"value": cpu_rays / dt / 1e6, "unit": "Mrays/s", "cores": O.num_threads(), "kind": "port",
                                    "sample": "oracle (-O2 -fopenmp), centre crop %dx%d of %dx%d, 2 frames"
                                              % (reg[2] - reg[0], reg[3] - reg[1], W, H)}
            # the same crop and frames on the GPU, production mode: does the timed path compute the reference's image?
            x0, y0, x1, y1 = reg
            pt.clear(); pt.set_region(*reg); pt.render_spp(0, 2)
            g_img, g_ids = pt.readback(), pt.readback_primary_ids()
            parity.update({"crop": list(reg), "frames": 2, "mode": args.mode,
                           "ids_equal": bool((g_ids[y0:y1, x0:x1] == ref_ids[y0:y1, x0:x1]).all()),
                           "rrmse": rrmse(g_img[y0:y1, x0:x1], ref_img[y0:y1, x0:x1]), "bar": "ids bit-exact, rRMSE <= 1e-3"})
            parity["ok"] = bool(parity["ids_equal"] and parity["rrmse"] <= 1e-3)
            # like for like with the reference arm: the same crop, ONE frame per step (its sample), and 256-frame steps
            pt.clear(); pt.synchronize(); pt.reset_stats()
            for k in range(64):
                pt.render_spp(k, 1)
            s1 = pt.stats()
            pt.reset_stats()
            for k in range(4):
                pt.render_spp(64 + k * spp, spp)
            s2 = pt.stats()
            line["crop_like_reference"] = {
                "crop": list(reg), "one_frame_steps": {"value": s1["rays"] / s1["render_ms"] / 1e3, "unit": "Mrays/s", "ms_per_step": s1["render_ms"] / 64},
                "%d_frame_steps" % spp: {"value": s2["rays"] / s2["render_ms"] / 1e3, "unit": "Mrays/s", "ms_per_step": s2["render_ms"] / 4},
                "note": "what --impl reference renders per step is the first entry's sample (one frame of this crop)"}
            pt.set_region(0, 0, W, H)
            # the reference's interactive loop: one sample per call, then the display image (pass3) back to the host
            disp_t = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()    # pinned, like the e2e loop's buffer
            disp = disp_t.numpy()
            pt.set_camera(eye_h, cam_h); pt.render_spp(0, 1)                   # untimed first use of the display buffer and the pinned pages
            R._check(R.lib().pt_readback_display(pt._h, disp.ctypes.data, disp.nbytes), "pt_readback_display")
            pt.clear(); pt.synchronize(); pt.reset_stats()
            t0 = time.perf_counter()
            for k in range(64):
                pt.set_camera(eye_h, cam_h)
                pt.render_spp(k, 1)
                R._check(R.lib().pt_readback_display(pt._h, disp.ctypes.data, disp.nbytes), "pt_readback_display")
            dt_i = time.perf_counter() - t0
            s3 = pt.stats()
            # the same loop with the copy of frame k under the rendering of frame k + 1 (pt_readback_display_async, two pinned buffers)
            disp2 = [disp_t, torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()]
            for k in range(2):
                pt.render_spp(k, 1); pt.readback_display_async(disp2[k].data_ptr(), disp2[k].numel())
            pt.clear(); pt.synchronize()
            t0 = time.perf_counter()
            for k in range(64):
                pt.set_camera(eye_h, cam_h)
                pt.render_spp(k, 1)
                pt.readback_display_async(disp2[k & 1].data_ptr(), disp2[k & 1].numel())
            pt.synchronize()
            dt_o = time.perf_counter() - t0
            line["e2e_interactive"] = {"value": s3["rays"] / dt_i / 1e6, "unit": "Mrays/s", "fps": 64 / dt_i, "ms_per_frame": dt_i / 64 * 1e3,
                                       "overlapped": {"fps": 64 / dt_o, "ms_per_frame": dt_o / 64 * 1e3,
                                                      "what": "pt_readback_display_async: the RGBA8 copy of frame k rides under the rendering of frame k + 1"},
                                       "kernel_ms_per_frame": s3["render_ms"] / 64, "h2d_bytes_per_step": 76, "d2h_bytes_per_step": int(disp.nbytes),
                                       "what": "display() of IS/main.cpp:659-709: set uniforms, one sample per pixel, pass3 tone map + RGBA8 readback, 64 frames"}
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
    else:
        line = None

    # ---- the other BASELINE.json configs (2, 3, 5): a driver-side number each ----
    others = {}
    if not args.no_other_workloads:
        del flush
        flush = None
        torch.cuda.empty_cache()
        names = OTHER_WORKLOADS if world == 1 else [w for w in OTHER_WORKLOADS if w[0] == "full-10M-4k"]   # N > 1: config 5, tile-sharded
        for name, frames in names:
            if world > 1 and rank != 0:
                dist.barrier()                                    # rank 0 builds and caches the scene, the others load it
            if world > 1 and rank == 0:
                build_workload(name)
                dist.barrier()
            m = measure_workload(R, name, frames, 3, local, world, rank)
            v = torch.tensor([m["ms"], float(m["rays"]), float(m["samples"])], dtype=torch.float64, device="cuda")
            if world > 1:
                mx = v.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
                sm = v.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
                ms_o, rays_o, samples_o = float(mx[0]), float(sm[1]), float(sm[2])
            else:
                ms_o, rays_o, samples_o = float(v[0]), float(v[1]), float(v[2])
            others[name] = {"value": rays_o / ms_o / 1e3, "unit": "Mrays/s", "ms_per_step": ms_o / 3, "frames_per_step": frames, "steps": 3,
                            "spp_per_s": samples_o / (m["W"] * m["H"]) / (ms_o * 1e-3), "rays_per_sample": rays_o / max(1.0, samples_o),
                            "triangles": m["triangles"], "variant": m["variant"], "width": m["W"], "height": m["H"],
                            "split": "tile (kernel time, max over ranks; no combine)" if world > 1 else "none"}
    if rank == 0:
        line["other_workloads"] = others
        print(json.dumps(line), flush=True)
    # every torch object that touched the context's stream or memory goes before the context does
    del host, flush
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
