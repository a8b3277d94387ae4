#!/usr/bin/env python3
"""Benchmark of the path-tracing hot path on B200 (BASELINE.json metric: Mpaths/s and
time-to-image, beside the reference's CPU renderer).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload = BASELINE.json configs[1]: three 102 400-triangle .obj meshes (metal / ceramic /
mirror) + textured floor + lamp under a sky, 1920x1080, 64 samples per pixel.  One step =
one complete image: film clear, 64 spp traced (sample indices sharded over the N ranks),
film reduce to rank 0 (NCCL), resolve (mean, gamma, RGB8).  Strong scaling: the image is
fixed, more GPUs finish it sooner.

value  whole-job Mpaths/s with scene and film resident in HBM, per-step CUDA events on the
       launching stream, max over ranks, L2 flushed between steps.
e2e    the same metric through the reference-facing host API (TraceSamples of the C++ mirror:
       camera in, DisplayFrame host buffers out), host<->device copies inside the timed region.

--impl reference times the reference's own CPU renderer (oracle/_ref/fs2_ref, compiled
unmodified from the reference sources) on the same scene, all host cores, on a bounded
sample of the workload per step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "scenes"))

WORKLOAD = "mesh100k"
WIDTH, HEIGHT, SPP = 1920, 1080, 64
SEED = 20260101
CACHE = os.path.join(REPO, "scenes", "_cache")
METRIC = "Mpaths/s (1080p, 64 spp, ~100k-tri .obj x3 metal/ceramic/mirror)"


def scene_path(rank=0, world=1):
    """Generates the scene files once (rank 0), everybody else waits for them."""
    import gen_scenes
    marker = os.path.join(CACHE, WORKLOAD + ".scene")
    if rank == 0:
        gen_scenes.generate(CACHE, [WORKLOAD])
    else:
        for _ in range(600):
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


def time_reference(scene, width, height, frames, threads):
    """Runs the unmodified reference renderer; returns (Mpaths/s, seconds)."""
    from oracle import binding
    if not binding.have_ref():
        raise RuntimeError("oracle/_ref/fs2_ref is missing (built in the dev container)")
    r = binding.run_ref(scene, width, height, frames=frames, threads=threads, realclock=True,
                        time_only=True)
    return r["mpaths_per_s"], r["seconds"]


def cpu_sample(scene, budget_s=12.0):
    """Bounded CPU sample of the workload: 480x270 (1/16 of the pixels), as many frames
    (= samples per pixel) as fit in about `budget_s` seconds."""
    cores = host_cores()
    w, h = 480, 270
    threads = ref_threads(h, cores)
    mp1, sec1 = time_reference(scene, w, h, 1, threads)
    frames = int(max(1, min(32, budget_s / max(sec1, 1e-3))))
    mp, sec = time_reference(scene, w, h, frames, threads) if frames > 1 else (mp1, sec1)
    return {"value": mp, "unit": "Mpaths/s", "cores": threads, "kind": "reference",
            "sample": "%dx%d, %d spp of the same scene (%.1f s), %d threads of %d host cores; "
                      "per-path cost does not depend on resolution or spp" %
                      (w, h, frames, sec, threads, cores)}


# --------------------------------------------------------------------------- reference arm
def run_reference(args, rank, world):
    """The reference's own CPU implementation of the path, timed on the host cores."""
    if rank != 0:
        return
    scene = scene_path(0, 1)
    cores = host_cores()
    w, h = 480, 270
    threads = ref_threads(h, cores)
    for _ in range(args.warmup):
        time_reference(scene, w, h, 1, threads)
    paths, secs = 0.0, 0.0
    for _ in range(args.steps):
        mp, sec = time_reference(scene, w, h, 1, threads)
        paths += w * h
        secs += sec
    value = paths / secs / 1e6
    sample = ("%dx%d, 1 spp of the same scene per step, %d threads of %d host cores; per-path "
              "cost does not depend on resolution or spp" % (w, h, threads, cores))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "Mpaths/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "%s %dx%d %dspp" % (WORKLOAD, WIDTH, HEIGHT, SPP),
                       "sample": sample},
            "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": threads,
                             "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


KERNEL_NAME = ["k_trace_phase<0>"]


# --------------------------------------------------------------------------- roofline
# DRAM bytes of one launch of each wavefront kernel on a full 4 194 304-path wave of the bench
# scene: dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` capture each
# (first measured in profiles/r1_h_*: 279.2 + 53.2 MB and 526.6 + 321.3 MB; current numbers from
# r1_l: 295.5 + 60.2 and 526.5 + 322.7 MB).
# Per ray segment that is 287 B, against the ~212 B of path state + hit record the pipeline moves
# by design; the scene itself is served from L2.
PROFILED_DRAM_BYTES_PER_SEGMENT = (295.546368e6 + 60.236032e6 + 526.524672e6 + 322.659072e6) / 4194304.0


def roofline(ctx_stats, objects, n_pixels, kernel_ms, peaks):
    """Algorithmic bytes / flops per SURVEY.md §8(d) from the stats-build counters of one
    launch, against the measured HBM peak and the FP32 pipe peak."""
    import numpy as np
    s = ctx_stats
    per_type_bytes = {0: 16, 1: 64, 2: 64, 3: 96, 4: 64}
    per_type_flops = {0: 32, 1: 32, 2: 32, 3: 160, 4: 32}
    prim_types = [int(t) for t in objects["type"] if int(t) != 5]
    pb = float(np.mean([per_type_bytes[t] for t in prim_types])) if prim_types else 0.0
    pf = float(np.mean([per_type_flops[t] for t in prim_types])) if prim_types else 0.0
    # our kernel's own algorithmic traffic: octree nodes (32 B), leaf-BVH nodes (128 B), planes
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
    sm_mhz = peaks.get("sm_max_mhz") or 1965.0
    fp32_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12
    return {"bound": "hbm", "achieved": nbytes / sec / 1e9, "peak": hbm_peak, "unit": "GB/s",
            "frac": nbytes / sec / 1e9 / hbm_peak, "traffic": None, "peak_source": which,
            "kernel": KERNEL_NAME[0], "kernel_ms": kernel_ms,
            "algorithmic_bytes": nbytes, "algorithmic_flops": flops,
            "reference_equivalent_bytes": ref_bytes,
            "reference_equivalent_gbs": ref_bytes / sec / 1e9,
            "fp32": {"achieved_tflops": flops / sec / 1e12, "peak_tflops": fp32_peak,
                     "frac": flops / sec / 1e12 / fp32_peak},
            "per_segment": {"n_node": s["n_node"] / max(1, s["n_segments"]),
                            "n_tri": s["n_tri"] / max(1, s["n_segments"]),
                            "n_bary": s["n_bary"] / max(1, s["n_segments"]),
                            "n_lbnode": s["n_lbnode"] / max(1, s["n_segments"]),
                            "n_tritest": s["n_tritest"] / max(1, s["n_segments"]),
                            "segments_per_path": s["n_segments"] / max(1, s["n_paths"])}}


# --------------------------------------------------------------------------- GPU arm
def run_gpu(args, rank, world, local_rank):
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

    scene_file = scene_path(rank, world)
    t0 = time.time()
    hs = host.HostScene(scene_file)
    desc = hs.desc()
    load_s = time.time() - t0
    cam = hs.camera()
    ctx = fs2_b200.Context(device=local_rank)
    ctx.upload(desc)
    ctx.resize(WIDTH, HEIGHT)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    film = multi_gpu.film_tensor(ctx)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    first, n_mine, stride = multi_gpu.shard_samples(SPP, rank, world)
    KERNEL_NAME[0] = {"phase": "k_trace_phase<0> (leaf BVH, warp-synchronous phases)",
                      "wavefront": "k_wf_trace<0> + k_wf_shade<0> + k_wf_flip (all launches of one fs_trace)",
                      "lbvh_lane": "k_trace_mega<0,1> (leaf BVH, free-running lanes)",
                      "coop": "k_trace_coop<0> (linear leaves, warp-cooperative)",
                      "per_thread": "k_trace_mega<0,0> (linear leaves)"}[args.variant]
    variant = {"phase": abi.FS_TRACE_MEGAKERNEL, "wavefront": abi.FS_TRACE_DEFAULT, "lbvh_lane": abi.FS_TRACE_LBVH_LANE, "coop": abi.FS_TRACE_COOP,
               "per_thread": abi.FS_TRACE_PER_THREAD}[args.variant]
    flags = abi.FS_TRACE_NO_RESOLVE | abi.FS_TRACE_ASYNC | variant

    def step(seed):
        ctx.reset()
        ctx.trace(cam, n_mine, first_sample=first, stride=stride, seed=seed, flags=flags)
        if world > 1:
            multi_gpu.reduce_film(film, 0)
        if rank == 0:
            ctx.resolve()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

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
        flush.zero_()
        barrier()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        step(SEED + 1000 + i)
        e1.record()
        barrier()
        step_ms.append(e0.elapsed_time(e1))
        kernel_ms.append(ctx.stats()["last_kernel_ms"])
    sampler.stop_flag = True
    sampler.join(timeout=2)
    st = ctx.stats()
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device="cuda")
    segs = torch.tensor([float(st["n_segments"])], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(segs, op=dist.ReduceOp.SUM)
    total_ms = float(total_ms.item())
    # BASELINE.json's second figure: time to image at 1080p, 256 spp (one extra, separately
    # timed frame; not part of `value`)
    f256, n256, s256 = multi_gpu.shard_samples(256, rank, world)
    for timed in (False, True):  # one untimed frame first: the path pool grows to this job's size
        flush.zero_()
        barrier()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        ctx.reset()
        ctx.trace(cam, n256, first_sample=f256, stride=s256, seed=SEED + 5000 + timed, flags=flags)
        if world > 1:
            multi_gpu.reduce_film(film, 0)
        if rank == 0:
            ctx.resolve()
        e1.record()
        barrier()
    t256 = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t256, op=dist.ReduceOp.MAX)
    t256 = float(t256.item())
    paths = float(WIDTH) * HEIGHT * SPP * args.steps
    value = paths / total_ms / 1e3

    # ---- e2e: camera (host POD) in, image in host buffers out on rank 0.
    #  N = 1: through the host mirror of the reference API (TraceSamples = n x TraceScene).
    #  N > 1: every rank calls the C ABI (fs_trace on its sample shard), the films are summed on
    #         the device (NCCL reduce to rank 0), rank 0 resolves and reads the film back through
    #         fs_film_read (mean, count, rgb8, normal, depth, material id: 81 MB D2H).
    frame = host.HostFrame(WIDTH, HEIGHT)
    hs.lib.fs2h_set_device(local_rank)
    e2e_s = []
    host_film = {}
    if world > 1 and rank == 0:
        from fs2_b200 import engine
        host_film = {"mean": np.empty((WIDTH * HEIGHT, 3), np.float32), "count": np.empty(WIDTH * HEIGHT, np.uint32),
                     "rgb8": np.empty((WIDTH * HEIGHT, 3), np.uint8), "normal": np.empty((WIDTH * HEIGHT, 3), np.float32),
                     "depth": np.empty(WIDTH * HEIGHT, np.float32), "mat_id": np.empty(WIDTH * HEIGHT, np.uint64)}
        for a in host_film.values():
            engine.pin_host(a)  # caller-owned, page-locked result buffers
    for i in range(1 + max(1, min(args.steps, 2))):
        barrier()
        t0 = time.perf_counter()
        if world == 1:
            frame.reset()
            frame.trace(hs, cam, n_mine)
        else:
            multi_gpu.render_sharded(ctx, cam, SPP, SEED + 2000 + i, rank, world, flags=variant)
            if rank == 0:
                ctx.read_film(aux=True, out=host_film)
        barrier()
        e2e_s.append(time.perf_counter() - t0)
    e2e_t = torch.tensor([min(e2e_s[1:])], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_value = float(WIDTH) * HEIGHT * SPP / float(e2e_t.item()) / 1e6
    n_px = WIDTH * HEIGHT
    d2h = n_px * (12 + 4 + 3 + 16 + 4)  # mean, count, rgb8, normal+depth, material id

    if rank == 0:
        # ---- stats pass (untimed): work counters of one launch for the roofline
        sctx = fs2_b200.Context(device=local_rank, stats=True)
        sctx.upload(desc)
        sctx.resize(WIDTH, HEIGHT)
        sctx.trace(cam, 2, seed=SEED, flags=variant)
        sst = sctx.stats()
        scale = float(n_mine) / 2.0  # counters grow linearly with samples per pixel
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
        roof = roofline(sst, objects, n_px, kms, peaks)
        if args.variant == "wavefront":
            roof["traffic"] = PROFILED_DRAM_BYTES_PER_SEGMENT * sst["n_segments"]
            roof["traffic_source"] = ("ncu --set full, one k_wf_trace + one k_wf_shade launch on a full "
                                      "wave (profiles/r1_l_*), scaled by this run's ray segments")
        try:
            cpu = cpu_sample(scene_file) if (world == 1 and not args.skip_cpu) else None
        except Exception as e:  # the reference binary did not travel
            cpu = {"value": None, "unit": "Mpaths/s", "cores": 0, "kind": "reference",
                   "sample": "unavailable: %s" % e}
        line = {"metric": METRIC, "value": value, "unit": "Mpaths/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "%s %dx%d %dspp" % (WORKLOAD, WIDTH, HEIGHT, SPP),
                           "scene": "3 x 102400-tri .obj (metal/ceramic/mirror) + quad + lamp + sky",
                           "scene_bytes": hs.nbytes(), "host_load_build_s": round(load_s, 2),
                           "sharding": "sample indices r, r+N, ... per rank; film reduce to rank 0",
                           "l2": "flushed between steps (256 MiB write)", "seed": SEED,
                           "variant": args.variant, "untimed_ramp_steps": ramp_steps,
                           "time_to_image_ms": total_ms / args.steps,
                           "time_to_image_1080p_256spp_ms": t256,
                           "mrays_per_s": float(segs.item()) / total_ms / 1e3},
                "clocks": sampler.summary(), "roofline": roof,
                "e2e": {"value": e2e_value, "unit": "Mpaths/s", "h2d_bytes_per_step": 48,
                        "d2h_bytes_per_step": d2h},
                "gpu_launches": int(st["launches"])}
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
    ap.add_argument("--skip-cpu", action="store_true", help="tuning runs: no cpu_baseline leg")
    ap.add_argument("--variant", default="wavefront", choices=["phase", "wavefront", "lbvh_lane", "coop", "per_thread"],
                    help="leaf-scan variant of the trace kernel (A/B measurements)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
