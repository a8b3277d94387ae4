#!/usr/bin/env python3
"""Benchmark of the path-tracing hot path on B200 (BASELINE.json metric: Mpaths/s and
time-to-image, beside the reference's CPU renderer).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config 2|3|4]
                    [--shard auto|samples|tiles]

Workloads (BASELINE.json configs; --config, default 2 = the configuration the metric is quoted on):
  2  three 102 400-triangle .obj meshes (metal / ceramic / mirror) + textured floor + lamp under a
     sky, 1920x1080, 64 samples per pixel (octree traversal stress)
  3  glass / frosted glass / liquid primitives, a fog cuboid, ceramic, under a float .exr
     environment, thin lens, 1920x1080, 256 spp (deep-bounce divergence)
  4  one 4.19 M-triangle height field + mirror / glass spheres, thin lens, 3840x2160, 256 spp

One step = one complete image: film clear, all samples traced (sample indices — or interleaved
pixel tiles — sharded over the N ranks), film reduce to rank 0 (NCCL), resolve (mean, gamma,
RGB8).  Strong scaling: the image is fixed, more GPUs finish it sooner.

value  whole-job Mpaths/s with scene and film resident in HBM, per-step CUDA events on the stream
       the kernels are launched on, max over ranks, L2 flushed between steps.
e2e    the same metric through the reference-facing host API (TraceSamples of the C++ mirror:
       camera in, DisplayFrame host buffers out), host<->device copies inside the timed region;
       at N > 1 rank 0 makes that one call on a multi-device context over all N GPUs
       (base::SetTraceDevices -> fs_create_multi), exactly what a user of the reference API does.

--impl reference times the reference's own CPU renderer (oracle/_ref/fs2_ref, compiled
unmodified from the reference sources) on the same scene, all host cores, on a bounded
sample of the workload per step.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "scenes"))

SEED = 20260101
CACHE = os.path.join(REPO, "scenes", "_cache")

# FogMaterial cuboid of config 3 (API-constructed): position, size, rotation axis +
# angle, colour, density — the same 14 numbers go to the host mirror and to the reference harness
FOG3 = (6.0, 6.0, -22.0, 16.0, 12.0, 6.0, 0.0, 1.0, 0.0, 0.25, 0.8, 0.85, 0.9, 0.6)

CONFIGS = {
    2: {"scene": "mesh100k", "width": 1920, "height": 1080, "spp": 64, "fog": None,
        "metric": "Mpaths/s (1080p, 64 spp, ~100k-tri .obj x3 metal/ceramic/mirror)",
        "what": "3 x 102400-tri .obj (metal/ceramic/mirror) + quad + lamp + sky"},
    3: {"scene": "glass3", "width": 1920, "height": 1080, "spp": 256, "fog": FOG3,
        "metric": "Mpaths/s (1080p, 256 spp, glass/liquid/fog/ceramic under an .exr environment, thin lens)",
        "what": "glass, frosted glass, liquid spheres + cuboids, fog cuboid, ceramic, mirror, 1024x512 .exr sky, aperture 1.5"},
    4: {"scene": "grid4m", "width": 3840, "height": 2160, "spp": 256, "fog": None,
        "metric": "Mpaths/s (4K, 256 spp, 4.19 M-tri .obj height field, thin lens)",
        "what": "2048x1024 height field (4 194 304 tris) + mirror and glass spheres + lamp, aperture 1.5"},
}


def scene_path(name, rank=0, world=1):
    """Generates the scene files once (rank 0), everybody else waits for them."""
    import gen_scenes
    marker = os.path.join(CACHE, name + ".scene")
    if rank == 0:
        gen_scenes.generate(CACHE, [name])
    else:
        for _ in range(3000):
            if os.path.exists(marker):
                break
            time.sleep(0.1)
    return marker


class ClockSampler(threading.Thread):
    """nvidia-smi clocks and throttle reasons sampled during the timed region."""
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self.stop_flag = False

    def run(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.FIELDS,
                                      "--format=csv,noheader,nounits"], capture_output=True,
                                     text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.max_mhz = float(out[1])
                for n, v in zip(names, out[2:]):
                    if v.strip().lower() == "active":
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def ref_threads(height, cores):
    """Largest divisor of the image height that does not exceed the core count (the
    reference double-renders boundary rows otherwise)."""
    best = 1
    for t in range(1, min(height, cores) + 1):
        if height % t == 0:
            best = t
    return best


def median(xs):
    s = sorted(xs)
    n = len(s)
    return s[n // 2] if n % 2 else 0.5 * (s[n // 2 - 1] + s[n // 2])


def time_reference(scene, width, height, frames, threads, fog=None):
    """Runs the unmodified reference renderer; returns (Mpaths/s, seconds)."""
    from oracle import binding
    if not binding.have_ref():
        raise RuntimeError("oracle/_ref/fs2_ref is missing (built in the dev container)")
    r = binding.run_ref(scene, width, height, frames=frames, threads=threads, realclock=True,
                        time_only=True, fog=fog)
    return r["mpaths_per_s"], r["seconds"]


def cpu_sample(scene, fog=None, budget_s=12.0):
    """Bounded CPU sample of the workload: 480x270, as many frames (= samples per pixel) as fit
    in about `budget_s` seconds of TraceScene time (scene load and octree build are outside the
    reference's own timer)."""
    cores = host_cores()
    w, h = 480, 270
    threads = ref_threads(h, cores)
    mp1, sec1 = time_reference(scene, w, h, 1, threads, fog)
    frames = int(max(1, min(32, budget_s / max(sec1, 1e-3))))
    mp, sec = time_reference(scene, w, h, frames, threads, fog) if frames > 1 else (mp1, sec1)
    return {"value": mp, "unit": "Mpaths/s", "cores": threads, "kind": "reference",
            "sample": "%dx%d, %d spp of the same scene (%.1f s), %d threads of %d host cores; "
                      "per-path cost does not depend on resolution or spp" %
                      (w, h, frames, sec, threads, cores)}


# --------------------------------------------------------------------------- reference arm
def run_reference(args, cfg, rank, world):
    """The reference's own CPU implementation of the path, timed on the host cores."""
    if rank != 0:
        return
    scene = scene_path(cfg["scene"], 0, 1)
    cores = host_cores()
    w, h = 480, 270
    threads = ref_threads(h, cores)
    for _ in range(args.warmup):
        time_reference(scene, w, h, 1, threads, cfg["fog"])
    paths, secs = 0.0, 0.0
    for _ in range(args.steps):
        mp, sec = time_reference(scene, w, h, 1, threads, cfg["fog"])
        paths += w * h
        secs += sec
    value = paths / secs / 1e6
    sample = ("%dx%d, 1 spp of the same scene per step, %d threads of %d host cores; per-path "
              "cost does not depend on resolution or spp" % (w, h, threads, cores))
    line = {"impl": "reference", "metric": cfg["metric"], "value": value, "unit": "Mpaths/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "%s %dx%d %dspp" % (cfg["scene"], cfg["width"], cfg["height"], cfg["spp"]),
                       "sample": sample},
            "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": threads,
                             "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- roofline
def csrc_hash():
    """sha256 over the kernel sources (same function as scripts/ncu_issue_counters.py): ties the
    committed ncu counters to the build that is being timed."""
    d = os.path.join(REPO, "final-stage-path-tracer-2-0_b200", "csrc")
    h = hashlib.sha256()
    for name in sorted(os.listdir(d)):
        if name.endswith((".cu", ".cuh", ".h", ".inc")):
            h.update(name.encode())
            h.update(open(os.path.join(d, name), "rb").read())
    return h.hexdigest()[:16]


def issue_counters(scene):
    """The newest committed ncu counter file for this scene's dominant kernel
    (profiles/*issue_counters*.json, written by scripts/ncu_issue_counters.py from an
    `ncu --set full` capture); marks it stale when the kernel sources changed since."""
    best = None
    pdir = os.path.join(REPO, "profiles")
    for name in sorted(os.listdir(pdir)) if os.path.isdir(pdir) else []:
        if "issue_counters" not in name or not name.endswith(".json"):
            continue
        try:
            c = json.load(open(os.path.join(pdir, name)))
        except Exception:
            continue
        if c.get("scene", "mesh100k") != scene:
            continue
        c["file"] = "profiles/" + name
        best = c  # names sort by round / step: the last one is the newest
    if best is not None:
        best["stale"] = best.get("csrc_sha16") != csrc_hash()
    return best


def roofline(s, objects, n_pixels, kernel_ms, peaks, counters, sm_mhz, kernel_name):
    """What binds the dominant kernel, and the fractions that do not.

    bound "issue": the traversal kernel is bound by instruction issue at low lane utilisation
    (ncu: DRAM ~1 % and FP32 ~2 % of peak, L2 hit > 90 %).  frac = issue-slot utilisation x
    non-predicated lanes / 32 of the dominant kernel — the share of the SMs' lane-issue capacity
    (148 SMs x 4 schedulers x 32 lanes per cycle) that executes useful thread instructions — from
    the committed `ncu --set full` capture of this build (`counters`).  Next to it, flat:
    hbm_frac_algorithmic (this kernel's own algorithmic bytes, SURVEY 8(d) constants, over the
    measured HBM peak), hbm_frac_measured (ncu DRAM bytes, scaled by ray segments), fp32_frac.
    """
    import numpy as np
    per_type_bytes = {0: 16, 1: 64, 2: 64, 3: 96, 4: 64}
    per_type_flops = {0: 32, 1: 32, 2: 32, 3: 160, 4: 32}
    prim_types = [int(t) for t in objects["type"] if int(t) != 5]
    pb = float(np.mean([per_type_bytes[t] for t in prim_types])) if prim_types else 0.0
    pf = float(np.mean([per_type_flops[t] for t in prim_types])) if prim_types else 0.0
    # the kernel's own algorithmic traffic: octree nodes (32 B), leaf-BVH nodes (128 B), planes
    # actually tested (16 B), barycentric records actually fetched (48 B), shading, film
    nbytes = (32 * s["n_node"] + 128 * s["n_lbnode"] + 16 * s["n_tritest"] + 48 * s["n_bary"] +
              pb * s["n_prim"] + 144 * s["n_shade"] + 16 * n_pixels)
    flops = (54 * s["n_node"] + 60 * s["n_lbnode"] + 13 * s["n_tritest"] + 23 * s["n_bary"] +
             pf * s["n_prim"] + 120 * s["n_shade"])
    # what the reference's loops execute for the same rays (SURVEY.md 8(d) constants)
    ref_bytes = (32 * s["n_node"] + 56 * s["n_tri"] + pb * s["n_prim"] + 144 * s["n_shade"] +
                 16 * n_pixels)
    sec = kernel_ms / 1e3
    hbm_peak = peaks.get("hbm_gbs")
    which = "measured"
    if not hbm_peak:
        hbm_peak, which = 6650.0, "fallback"
    clock = sm_mhz or peaks.get("sm_max_mhz") or 1965.0
    fp32_peak = 148 * 128 * 2 * clock * 1e6 / 1e12
    lane_peak = 148 * 4 * 32 * clock * 1e6 / 1e12  # tera thread-instructions per second
    seg = max(1.0, s["n_segments"])
    out = {"bound": "issue", "unit": "Tlane-inst/s", "peak": lane_peak, "achieved": None, "frac": None,
           "peak_source": "148 SMs x 4 schedulers x 32 lanes x %.0f MHz (SM clock sampled during the run)" % clock,
           "kernel": kernel_name, "kernel_ms": kernel_ms, "traffic": None,
           "hbm_frac_algorithmic": nbytes / sec / 1e9 / hbm_peak,
           "hbm_frac_measured": None,
           "hbm_peak_gbs": hbm_peak, "hbm_peak_source": which,
           "fp32_frac": flops / sec / 1e12 / fp32_peak, "fp32_peak_tflops": fp32_peak,
           "algorithmic_bytes": nbytes, "algorithmic_flops": flops,
           "algorithmic_gbs": nbytes / sec / 1e9, "algorithmic_tflops": flops / sec / 1e12,
           "reference_equivalent_bytes": ref_bytes, "reference_equivalent_gbs": ref_bytes / sec / 1e9,
           "per_segment_n_node": s["n_node"] / seg, "per_segment_n_tri": s["n_tri"] / seg,
           "per_segment_n_bary": s["n_bary"] / seg, "per_segment_n_lbnode": s["n_lbnode"] / seg,
           "per_segment_n_tritest": s["n_tritest"] / seg, "per_segment_n_prim": s["n_prim"] / seg,
           "segments_per_path": s["n_segments"] / max(1.0, s["n_paths"])}
    if counters is not None:
        out["frac"] = counters.get("lane_issue_frac")
        if out["frac"] is not None:
            out["achieved"] = out["frac"] * lane_peak
        for k in ("issue_slot_busy_pct", "active_threads", "nonpred_threads", "achieved_occupancy_pct",
                  "l1_hit_pct", "l2_hit_pct", "registers"):
            out["ncu_" + k] = counters.get(k)
        out["counters_source"] = counters.get("file")
        out["counters_stale"] = counters.get("stale")
        dram = counters.get("dram_bytes_per_segment")
        if dram:
            out["traffic"] = dram * s["n_segments"]
            out["hbm_frac_measured"] = out["traffic"] / sec / 1e9 / hbm_peak
            out["traffic_source"] = ("ncu dram__bytes_read.sum + dram__bytes_write.sum of one k_wf_trace + one "
                                     "k_wf_shade launch per wave (%s), scaled by this run's ray segments"
                                     % counters.get("file"))
    return out


# --------------------------------------------------------------------------- GPU arm
def run_gpu(args, cfg, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    import fs2_b200
    from fs2_b200 import abi, host, multi_gpu

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    W, H, SPP = cfg["width"], cfg["height"], cfg["spp"]

    scene_file = scene_path(cfg["scene"], rank, world)
    # mesh octrees are built on this rank's GPU (csrc/fs2_build.cuh: byte-identical to the host
    # builder, tests/test_gpu_build.py); --host-build keeps the host builder
    host.load_host_library().fs2h_set_gpu_build(-1 if args.host_build else local_rank)
    t0 = time.time()
    hs = host.HostScene(scene_file)
    if cfg["fog"]:
        f = cfg["fog"]
        hs.add_fog_cuboid(f[0:3], f[3:6], f[6:10], f[10:13], f[13])
    desc = hs.desc()
    load_s = time.time() - t0
    cam = hs.camera()
    ctx = fs2_b200.Context(device=local_rank)
    t0 = time.time()
    ctx.upload(desc)
    upload_s = time.time() - t0
    ctx.resize(W, H)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    kernel_name = {"phase": "k_trace_phase<0> (leaf BVH, warp-synchronous phases)",
                   "wavefront": "k_wf_trace<0,%d> (dominant) + k_wf_shade<0> + k_wf_flip: all launches of one fs_trace"
                                % (1 if desc.scene_root >= 0 else 0),
                   "lbvh_lane": "k_trace_mega<0,1> (leaf BVH, free-running lanes)",
                   "coop": "k_trace_coop<0> (linear leaves, warp-cooperative)",
                   "per_thread": "k_trace_mega<0,0> (linear leaves)"}[args.variant]
    variant = {"phase": abi.FS_TRACE_MEGAKERNEL, "wavefront": abi.FS_TRACE_DEFAULT,
               "lbvh_lane": abi.FS_TRACE_LBVH_LANE, "coop": abi.FS_TRACE_COOP,
               "per_thread": abi.FS_TRACE_PER_THREAD}[args.variant]
    tiles = {"auto": None, "samples": False, "tiles": True}[args.shard]
    if tiles and args.variant != "wavefront":
        raise SystemExit("tile sharding needs the default (wavefront) pipeline")

    def step(seed, spp=SPP):
        multi_gpu.render_sharded(ctx, cam, spp, seed, rank, world, flags=variant, tiles=tiles)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn):
        """Device time of fn() on the context's stream (the stream the kernels, the NCCL reduce
        and the resolve are ordered on), bracketed by barriers."""
        flush.zero_()
        barrier()
        with multi_gpu.context_stream(ctx):
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
        barrier()
        return e0.elapsed_time(e1)

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # Untimed ramp before the W warm-up steps: at N = 8 a step is ~35 ms, and three of them do not
    # bring idle GPUs (and NCCL's first reduce) up to steady clocks; keep stepping for about 1.5 s
    # of wall time.  Every rank runs the same count (rank 0's clock decides) so the reduces pair.
    ramp_steps = 0
    t_ramp = time.perf_counter()
    while True:
        step(SEED + 900 + ramp_steps)
        ramp_steps += 1
        barrier()
        go = torch.tensor([1 if time.perf_counter() - t_ramp < 1.5 and ramp_steps < 64 else 0],
                          dtype=torch.int32, device="cuda")
        if world > 1:
            dist.broadcast(go, 0)
        if not int(go.item()):
            break
    for i in range(args.warmup):
        step(SEED + i)
    barrier()
    ctx.reset_stats()
    sampler = ClockSampler(local_rank)
    sampler.start()
    step_ms, kernel_ms = [], []
    for i in range(args.steps):
        step_ms.append(timed(lambda: step(SEED + 1000 + i)))
        kernel_ms.append(ctx.stats()["last_kernel_ms"])
    sampler.stop_flag = True
    sampler.join(timeout=2)
    st = ctx.stats()
    total_ms = max_over_ranks(sum(step_ms))
    segs = torch.tensor([float(st["n_segments"])], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(segs, op=dist.ReduceOp.SUM)
    paths = float(W) * H * SPP * args.steps
    value = paths / total_ms / 1e3

    # BASELINE.json's second figure: time to image at 1080p (config 4: 4K), 256 spp — median of
    # three separately timed frames after one untimed frame (the path pool grows to the job's size)
    if SPP == 256:
        t256 = median([max_over_ranks(t) for t in step_ms])
    else:
        step(SEED + 5000, 256)
        t256 = median([max_over_ranks(timed(lambda: step(SEED + 5001 + k, 256))) for k in range(3)])

    # ---- N-GPU film against the 1-GPU film of the same seed (<= 1 ulp x spp): rank 0
    # re-renders the last timed image alone, all samples, and compares every pixel
    parity = None
    if world > 1:
        pseed = SEED + 7000
        step(pseed)
        barrier()
        if rank == 0:
            film_n = ctx.read_film(aux=True)
            film_n = {k: v.copy() for k, v in film_n.items()}
            multi_gpu.render_sharded(ctx, cam, SPP, pseed, 0, 1, flags=variant)
            film_1 = ctx.read_film(aux=True)
            a, b = film_n["mean"].astype(np.float64), film_1["mean"].astype(np.float64)
            scale = np.maximum(np.maximum(np.abs(a), np.abs(b)), 1e-30)
            rel = np.abs(a - b) / scale / 2.0 ** -23  # in units of one float ulp of the larger value
            parity = {"max_ulp": float(rel.max()), "bound_ulp": float(SPP),
                      "ok": bool(rel.max() <= SPP and (film_n["count"] == film_1["count"]).all() and
                                 (film_n["count"] == SPP).all()),
                      "counts_equal_spp": bool((film_n["count"] == SPP).all()),
                      "aux_identical": bool((film_n["mat_id"] == film_1["mat_id"]).all() and
                                            (film_n["normal"].view(np.uint32) == film_1["normal"].view(np.uint32)).all() and
                                            (film_n["depth"].view(np.uint32) == film_1["depth"].view(np.uint32)).all()),
                      "sharding": "tiles" if (tiles or (tiles is None and SPP < world)) else "samples"}
        barrier()

    # ---- e2e: camera (host POD) in, image in the DisplayFrame's page-locked host buffers out,
    # through the host mirror of the reference API (TraceSamples = n x TraceScene).  N > 1: rank 0
    # makes that one call on a multi-device context over all N GPUs (SetTraceDevices); the other
    # ranks wait.  Median of >= 3 timed calls after one untimed call (scene upload, pool growth).
    e2e_s = []
    n_e2e = max(3, min(args.steps, 5))
    barrier()  # every rank's own GPU work is finished before rank 0 borrows the GPUs
    if rank == 0:
        frame = host.HostFrame(W, H)
        host.set_trace_devices(list(range(world)) if world > 1 else [local_rank],
                               {"auto": 0, "samples": 1, "tiles": 2}[args.shard])
        hs.lib.fs2h_set_seed(SEED + 2000)
        for i in range(1 + n_e2e):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            frame.reset()
            frame.trace(hs, cam, SPP)
            e2e_s.append(time.perf_counter() - t0)
        e2e_counts_ok = bool((frame.count() == SPP).all())
        host.set_trace_devices([local_rank])  # drops the multi-device context
        frame.close()
    if world > 1:
        # The other ranks must wait on the HOST here: an NCCL barrier would park a spinning kernel
        # on every GPU that rank 0's multi-device context is rendering on (two processes' kernels
        # waiting on each other on one GPU).  The rendezvous store is a plain TCP key-value wait.
        import datetime
        store = dist.distributed_c10d._get_default_store()
        if rank == 0:
            store.set("fs2_e2e_done", "1")
        else:
            store.wait(["fs2_e2e_done"], datetime.timedelta(minutes=20))
    barrier()
    n_px = W * H
    # what TraceSamples copies back per call: the picture (float mean, count, display bytes).  The
    # guide planes (normal, depth, material id: 24 more bytes per pixel, private in the reference's
    # DisplayFrame) stay on the device until a getter asks for them.
    d2h = n_px * (12 + 4 + 3)

    if rank == 0:
        e2e_value = float(W) * H * SPP / median(e2e_s[1:]) / 1e6
        # ---- stats pass (untimed): work counters of one launch for the roofline
        sctx = fs2_b200.Context(device=local_rank, stats=True)
        sctx.upload(desc)
        sctx.resize(W, H)
        sctx.trace(cam, 2, seed=SEED, flags=variant)
        sst = sctx.stats()
        scale = float(SPP) / 2.0 / world  # counters grow linearly with samples per pixel; per rank
        for k in ("n_paths", "n_segments", "n_node", "n_tri", "n_bary", "n_prim", "n_shade",
                  "n_lbnode", "n_tritest"):
            sst[k] = sst[k] * scale
        sctx.close()
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        objects = np.frombuffer((__import__("ctypes").c_char * (desc.n_objects * 144))
                                .from_address(desc.objects), dtype=abi.OBJECT_DT)
        kms = sum(kernel_ms) / len(kernel_ms)
        clocks = sampler.summary()
        roof = roofline(sst, objects, n_px, kms, peaks, issue_counters(cfg["scene"]) if args.variant == "wavefront" else None,
                        clocks.get("sm_mhz"), kernel_name)
        try:
            cpu = cpu_sample(scene_file, cfg["fog"]) if (world == 1 and not args.skip_cpu) else None
        except Exception as e:  # the reference binary did not travel
            cpu = {"value": None, "unit": "Mpaths/s", "cores": 0, "kind": "reference",
                   "sample": "unavailable: %s" % e}
        sharding = ("interleaved 16x8-pixel tiles, (tx + ty) mod N" if (tiles or (tiles is None and SPP < world))
                    else "sample indices r, r+N, ... per rank") + "; film reduce to rank 0 (NCCL)"
        line = {"metric": cfg["metric"], "value": value, "unit": "Mpaths/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "%s %dx%d %dspp" % (cfg["scene"], W, H, SPP),
                           "baseline_config": args.config, "scene": cfg["what"],
                           "scene_bytes": hs.nbytes(), "host_load_build_s": round(load_s, 2),
                           "scene_upload_s": round(upload_s, 3),
                           "mesh_octree_build": "host" if args.host_build else "device (k_bd_*)",
                           "sharding": sharding,
                           "l2": "flushed between steps (256 MiB write)", "seed": SEED,
                           "variant": args.variant, "untimed_ramp_steps": ramp_steps,
                           "time_to_image_ms": total_ms / args.steps,
                           "time_to_image_256spp_ms": t256,
                           "time_to_image_256spp_note": "median of 3 timed frames at this resolution",
                           "mrays_per_s": float(segs.item()) / total_ms / 1e3,
                           "step_ms": [round(t, 3) for t in step_ms],
                           "e2e_api": "base::TraceSamples on %d device(s) (SetTraceDevices), median of %d calls"
                                      % (world, n_e2e),
                           "e2e_s": [round(t, 5) for t in e2e_s[1:]], "e2e_counts_ok": e2e_counts_ok},
                "clocks": clocks, "roofline": roof,
                "e2e": {"value": e2e_value, "unit": "Mpaths/s", "h2d_bytes_per_step": 48,
                        "d2h_bytes_per_step": d2h},
                "gpu_launches": int(st["launches"])}
        if args.config == 2:
            line["config"]["time_to_image_1080p_256spp_ms"] = t256
        if parity is not None:
            line["multi_gpu_parity"] = parity
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="fs2")
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS),
                    help="BASELINE.json config (2 = the headline configuration)")
    ap.add_argument("--shard", default="auto", choices=["auto", "samples", "tiles"],
                    help="multi-GPU sharding: sample indices, interleaved tiles, or automatic")
    ap.add_argument("--skip-cpu", action="store_true", help="tuning runs: no cpu_baseline leg")
    ap.add_argument("--host-build", action="store_true", help="build mesh octrees on the host, not the GPU")
    ap.add_argument("--variant", default="wavefront", choices=["phase", "wavefront", "lbvh_lane", "coop", "per_thread"],
                    help="leaf-scan variant of the trace kernel (A/B measurements)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    cfg = CONFIGS[args.config]
    if args.impl == "reference":
        run_reference(args, cfg, rank, world)
    else:
        run_gpu(args, cfg, rank, world, local_rank)


if __name__ == "__main__":
    main()
