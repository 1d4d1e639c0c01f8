#!/usr/bin/env python
"""bench.py -- Msamples/s of the path-tracing hot path on N B200s (driver contract in the task brief).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3] [--impl native|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 ... bench.py --gpus N ...

One "step" = one complete render of the workload (every pixel, every sample, reduce across GPUs,
normalise + gamma + pack RGB8).  Ranks share a render by sample ranges (or by interleaved 8-row bands of pixels,
--partition bands); the one collective is an NCCL sum-reduce of the float4 accumulators to rank 0.

  value    : W*H*spp / step time with scene and accumulators resident in HBM (CUDA events, max over ranks)
  e2e      : same metric through the host-buffer C-ABI path: scene upload (H2D) + render + RGB8 image back (D2H)
  roofline : FP32 CUDA-core bound.  achieved = 17 flop x (ray segments x spheres) / trace-kernel time
             (BASELINE.md section 4); peak = FMA-chain microbenchmark measured in this run.
  cpu_baseline / --impl reference : the UNMODIFIED reference's RenderSegmentTask (oracle/_ref, built from
             /root/reference by oracle/Makefile) on all host threads over square tiles, bounded sample.
"""
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

FLOP_PER_TEST = 17  # cd::RaySphereIntersection (Collision.hpp:9-17): 3 sub + 5 + 5 (dots) + mul + sub + mul + sub

# BASELINE.json configs.  tile = square CPU tile edge (the task tracer mis-strides non-square segments).
WORKLOADS = {
    "c1": dict(scene="random", width=640, height=360, spp=16, depth=8, tile=40,
               desc="SceneGenerators random-spheres scene, 640x360, 16 spp, depth 8"),
    "c2": dict(scene="two", width=1920, height=1080, spp=64, depth=4, tile=120,
               desc="single diffuse sphere + ground sphere, 1920x1080, 64 spp, depth 4"),
    "c3": dict(scene="random", width=3840, height=2160, spp=1024, depth=50, tile=240,
               desc="random-spheres scene, 3840x2160, 1024 spp, depth 50"),
    # not bench headlines: parity / stress configurations, timed for DESIGN.md's per-config table
    "c4": dict(scene="metalglass", width=1920, height=1080, spp=256, depth=50, tile=120,
               desc="ground + 500 mirror/glass spheres, 1920x1080, 256 spp, depth 50 (divergence stress)"),
    "c5": dict(scene="lattice", width=1920, height=1080, spp=128, depth=16, tile=120,
               desc="ground + 16383 spheres on a jittered lattice, 1920x1080, 128 spp, depth 16 (staging stress)"),
}
PASS_CAP = 10  # RenderSegmentTask semantics (TaskBasedPathTracer.hpp:81), the tracer BASELINE.json names
SEED = 1


def load_scene(kind):
    import simplepathtracer_b200 as spt

    if kind == "random":
        # GenerateSpheres() of the unmodified reference with its clock seed set to 1, frozen by
        # tests/golden/make_golden.py (the generator is time-seeded, so the dump IS the scene)
        return spt.Scene.load(os.path.join(ROOT, "tests", "golden", "scene_random_seed1.npz"))
    if kind == "two":
        return spt.Scene(
            centers=np.array([[0, -1000.5, 0, 0], [0, 1, 0, 0]], np.float32), radii=np.array([1000.0, 0.5], np.float32),
            materials=np.array([3, 3], np.uint8), colors=np.array([[30, 144, 255, 0], [255, 100, 230, 0]], np.float32),
            fuzz=np.zeros(2, np.float32))
    if kind == "metalglass":
        return spt.generate_metal_glass(4, 500)
    if kind == "lattice":
        return spt.generate_lattice(5, 128)
    raise ValueError(kind)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""

    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        for ts, line in self.rows:
            if ts < t0 - 0.05 or ts > t1 + 0.05:
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------- reference / CPU baseline
def cpu_reference_rate(wl, budget_s, threads=None):
    """Times the reference's own RenderSegmentTask on the host cores over a bounded sample of the workload
    (same scene, same depth; reduced resolution / spp: Msamples/s is spp-invariant).  Returns a dict."""
    from oracle import Oracle, Reference, ref_available

    threads = threads or os.cpu_count() or 1
    scene = load_scene(wl["scene"])
    w, h, tile = 640, 360, 40
    wide = scene.n > 255  # the reference's uint8_t sphere index widened, nothing else touched
    if ref_available(wide=wide):
        R = Reference(wide=wide)
        R.set_camera()
        R.set_scene(scene)
        R.set_config(w, h, 2, wl["depth"])
        _, _, sec = R.render_mt(Reference.TASK, tile, threads, 1)  # calibration
        rate = w * h * 2 / max(sec, 1e-6)
        spp = int(max(2, min(4096, budget_s * rate / (w * h))))
        R.set_config(w, h, spp, wl["depth"])
        _, _, sec = R.render_mt(Reference.TASK, tile, threads, 2)
        return {"value": w * h * spp / sec / 1e6, "unit": "Msamples/s", "cores": threads, "kind": "reference", "width": w, "height": h, "spp": spp,
                "build": "g++ -O2 -msse4.1 -ffp-contract=off over the unmodified headers (oracle/Makefile)",
                "sample": f"{wl['scene']} scene, {w}x{h}, {spp} spp, depth {wl['depth']}, RenderSegmentTask over {tile}x{tile} tiles, "
                          f"{threads} threads, {sec:.1f} s"}
    # the compiled reference did not travel: fall back to the single-threaded C port of the same tracer
    O = Oracle()
    cam = O.camera()
    spp = 2
    t0 = time.time()
    O.render_refseq(scene, cam, w, h, spp, wl["depth"], Oracle.TASK, 1)
    sec = time.time() - t0
    return {"value": w * h * spp / sec / 1e6, "unit": "Msamples/s", "cores": 1, "kind": "port", "width": w, "height": h, "spp": spp,
            "sample": f"{wl['scene']} scene, {w}x{h}, {spp} spp, depth {wl['depth']}, C port of RenderSegmentTask, 1 thread, {sec:.1f} s"}


def run_reference_arm(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rates, sample = [], None
    t_steps = []
    for i in range(args.warmup + args.steps):
        t0 = time.time()
        r = cpu_reference_rate(wl, budget_s=4.0)
        if i >= args.warmup:
            rates.append(r["value"])
            t_steps.append(time.time() - t0)
        sample = r
    value = float(np.mean(rates))
    line = {
        "impl": "reference", "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": float(np.mean(t_steps)) * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        # the native arm's workload is named; width / height / spp are what a step of THIS arm really renders (a bounded
        # sample of it: same scene, same depth, same tracer; Msamples/s is resolution- and spp-invariant to first order)
        "config": {"workload": f"{args.workload}: {wl['desc']}", "spheres": load_scene(wl["scene"]).n, "width": sample["width"],
                   "height": sample["height"], "spp": sample["spp"], "depth": wl["depth"], "pass_cap": PASS_CAP,
                   "native_arm_size": f"{wl['width']}x{wl['height']}x{wl['spp']} spp",
                   "sample": f"bounded sample per step: {sample['sample']}"},
        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": sample["cores"], "kind": sample["kind"], "sample": sample["sample"],
                         "build": sample.get("build")},
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# --------------------------------------------------------------------------- native arm
# FMA-pipe flops the pre-filter really executes per (ray, sphere) test.  Bounce kernel: general form, 7 FMA + 1 ADD.  Camera
# kernel: NONE per sample -- one camera-form test (4 FMA) per sphere, PIXEL and launch decides which few spheres a pixel's
# samples can hit (pixel cone, csrc/spt_filter.cuh); the samples run the exact sequence on that list only.
EXECUTED_FLOP = {"primary": None, "secondary": 15}


def kernel_rooflines(ctr_per_render, n_spheres, primary_ms, secondary_ms, peak):
    """Per-kernel FP32 roofline of split rendering from the device counters of ONE render and the kernels' own durations.
    Camera segments = samples (one per sample, traced by primary_hits_*), secondary segments = the rest (trace_secondary*).
    `frac` is ALGORITHMIC: 17 flop per (segment, sphere) as the reference's test defines it (Collision.hpp:9-17);
    `executed_frac` counts the FMA-pipe flops the pre-filter really executes per test (8 / 15), exact candidates not included."""
    seg_cam = ctr_per_render["samples"]
    seg_sec = ctr_per_render["segments"] - ctr_per_render["samples"]
    out = {}
    for name, segs, ms, key in (("primary_hits_kernel", seg_cam, primary_ms, "primary"), ("trace_secondary", seg_sec, secondary_ms, "secondary")):
        if ms <= 0:
            continue
        tests = segs * n_spheres
        alg = FLOP_PER_TEST * tests / (ms * 1e-3) / 1e12
        out[name] = {"ms_per_render": ms, "segments": segs, "tests": tests, "achieved": alg, "frac": alg / peak if peak else None}
        if EXECUTED_FLOP[key]:
            exe = EXECUTED_FLOP[key] * tests / (ms * 1e-3) / 1e12
            out[name].update({"executed_flop_per_test": EXECUTED_FLOP[key], "executed_frac": exe / peak if peak else None})
        else:
            out[name]["note"] = ("algorithmic rate only: the tests of a camera sample are implied by ONE conservative test per sphere, pixel and "
                                 "launch (pixel-cone candidate lists); per sample the kernel runs the exact sequence on the listed spheres")
        if key == "secondary" and ctr_per_render.get("warp_iterations"):
            # share of the lanes that held a ray in the bounce kernel's closest-hit trips (the rest waited for warp-mates)
            out[name]["lanes_with_a_ray"] = segs / (32.0 * ctr_per_render["warp_iterations"])
    return out


def run_native_arm(args, wl):
    import torch
    import torch.distributed as dist

    import simplepathtracer_b200 as spt

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the native arm has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        # a HOST-side group for the end-to-end leg: while rank 0's process drives all GPUs through one multi-device context, the
        # other ranks must wait without a kernel on their GPU -- an NCCL barrier is a spinning kernel, and two processes on one
        # GPU time-slice (session 8: 1 128 ms per render with the peers parked in an NCCL barrier against 512 ms device-timed)
        host_group = dist.new_group(backend="gloo")

    spt.build_library()
    scene = load_scene(wl["scene"])
    cam = spt.Camera()  # the reference camera (Globals.hpp:22-24 through Renderer.hpp:321)
    W, H, spp, depth = wl["width"], wl["height"], args.spp or wl["spp"], wl["depth"]
    from simplepathtracer_b200 import sharding

    # sample ranges (default): rank r renders samples [s0, s1) of every pixel, so every rank launches the full grid;
    # --partition bands: rank r renders every sample of the 8-row bands b with b % world == r (disjoint pixels, single-GPU
    # bits, but 1/world of the CTAs per launch: slower at 8 GPUs)
    if args.partition == "bands":
        s0, s1 = 0, spp
        bands = sharding.band_params(rank, world) if world > 1 else None
    else:
        s0, s1 = sharding.sample_range(rank, world, spp)
        bands = None

    staging = (args.staging_mib << 20) if args.staging_mib else spt.DEFAULT_STAGING_BYTES
    pt = spt.PathTracer(local)
    pt.set_staging_limit(staging)
    pt.strip_rows, pt.split_chunk = args.strip_rows, args.split_chunk
    # one explicit stream for torch fills, our kernels and the timing events
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    pt.set_stream(stream.cuda_stream)
    pt.set_scene(scene)
    info = pt.device_info()
    accum = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    rgb_dev = torch.zeros((H, W, 3), dtype=torch.uint8, device="cuda")
    rgb_host = torch.empty((H, W, 3), dtype=torch.uint8, pin_memory=True)
    launches = [0]
    variant = spt.WAVEFRONT if args.variant == "wavefront" else spt.MEGAKERNEL

    def step_device(ev=None, flags=0):
        accum.zero_()
        if ev:
            ev[0].record()
        pt.accumulate(cam, W, H, s0, s1, depth, accum, pass_cap=PASS_CAP, seed=SEED, flags=spt.FLAG_COUNTERS | args.flags | flags, variant=variant,
                      bands=bands)
        if ev:
            ev[1].record()
        launches[0] += 1
        sharding.reduce_accumulators(accum, dst=0)  # NCCL sum-reduce over NVLink; a no-op for one rank
        if rank == 0:
            pt.resolve(accum, W, H, rgb_dev)
            launches[0] += 1

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def maxr(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sumr(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- device-resident arm -------------------------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    barrier()
    pt.reset_counters()
    launches[0] = 0
    kernel_events = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    clocks = ClockSampler(local)
    clocks.start()
    time.sleep(0.25)
    t_wall0 = time.time()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for k in range(args.steps):
        step_device(kernel_events[k])
    e1.record()
    barrier()
    t_wall1 = time.time()
    clock_info = clocks.stop(t_wall0, t_wall1)
    ms_total = maxr(e0.elapsed_time(e1))
    ms_step = ms_total / args.steps
    total_samples = W * H * spp
    value = total_samples / (ms_step * 1e-3) / 1e6
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in kernel_events]))
    ctr = pt.counters()
    timing = pt.timing()
    trace_launches = timing["launches"]  # kernels of ONE accumulate call (split rendering: 2 per chunk of samples)
    launches_native = int(sumr(launches[0] + args.steps * (trace_launches - 1)))
    split = trace_launches > 1 and args.variant != "wavefront"
    ctr1 = {k: v / args.steps for k, v in ctr.items()}  # this rank's share of one render

    # ---- roofline (rank 0's kernels) ---------------------------------------------------------------
    peak_scalar, peak_packed = pt.measure_fp32_peak()
    peak = max(peak_scalar, peak_packed)
    tests_per_launch = ctr["tests"] / args.steps
    achieved_step = FLOP_PER_TEST * tests_per_launch / (kernel_ms * 1e-3) / 1e12
    kernels, serial_ms = {}, None
    if split:
        # one more render, untimed for the headline, with the chunks serialised on one stream: the kernels' OWN durations
        # (in the timed region the camera kernel of chunk k+1 runs under the tail of the bounce kernel of chunk k)
        evs = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        step_device(evs, flags=spt.FLAG_SERIAL_CHUNKS)
        torch.cuda.synchronize()
        ts = pt.timing()
        serial_ms = evs[0].elapsed_time(evs[1])
        kernels = kernel_rooflines(ctr1, scene.n, ts["primary_ms"], ts["secondary_ms"], peak)
    traffic = traffic_detail = None
    try:  # measured once per round under ncu --set full and committed; never measured inside a timed run
        t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(f"{args.workload}:{args.variant}")
        if t:
            per = t.get("samples_per_launch_in_capture")
            chunk = (s1 - s0) / max(trace_launches // 2, 1) if split else (s1 - s0)
            scale = (chunk / per) if (per and split) else 1.0
            share = (1.0 / world) if bands else 1.0  # this rank's share of the pixels
            scale *= share
            traffic = (t["dram_bytes_read"] + t["dram_bytes_write"]) * scale  # bytes per launch of the dominant kernel
            traffic_detail = {"kernel": t.get("kernel", "trace_megakernel"), "dram_bytes_per_launch": traffic,
                              # the path's algorithmic HBM bytes are the float4 accumulator, read and written once
                              "algorithmic_bytes_per_launch": W * H * 32 * share,
                              # the cost of the split design, named separately: one 16-byte record per camera sample, written by
                              # the camera kernel and read by the bounce kernel
                              "staging_bytes_per_launch": (W * H * chunk * 16 * share) if split else 0,
                              "source": t["capture"]}
            if "primary_dram_bytes_write" in t and split:
                traffic_detail["primary_hits_dram_bytes_per_launch"] = (t["primary_dram_bytes_read"] + t["primary_dram_bytes_write"]) * scale
    except (OSError, ValueError):
        pass
    dom = kernels.get("trace_secondary")
    roofline = {
        "bound": "fp32", "unit": "TFLOP/s", "peak": peak,
        # the DOMINANT kernel alone (the bounce kernel, ~3/4 of a render) when the render is split; the whole step otherwise
        "achieved": dom["achieved"] if dom else achieved_step, "frac": (dom["frac"] if dom else (achieved_step / peak if peak else None)),
        "kernel": "trace_secondary" if dom else ("trace_" + args.variant),
        "kernels": kernels,
        "step": {"achieved": achieved_step, "frac": achieved_step / peak if peak else None, "kernel_ms": kernel_ms,
                 "kernel_share_of_step": kernel_ms / ms_step,
                 "launches": f"{trace_launches // 2} work item(s) x (primary_hits_kernel + trace_secondary)" if split else f"{trace_launches} launch(es)",
                 "serialised_chunks_ms": serial_ms},
        "traffic": traffic, "traffic_detail": traffic_detail,
        "exact_evals_per_segment": ctr["exact"] / max(ctr["segments"], 1),
        "algorithmic": f"{FLOP_PER_TEST} flop x (ray segments x spheres) (BASELINE.md section 4; Collision.hpp:9-17); this rank, one render: "
                       f"{ctr1['segments']:.4g} segments ({ctr1['samples']:.4g} camera + {ctr1['segments'] - ctr1['samples']:.4g} secondary) x {scene.n} spheres; "
                       f"{ctr['segments'] / max(ctr['samples'], 1):.3f} segments/sample",
        "peak_source": f"measured in this run: dependent-FMA chains on all SMs, scalar {peak_scalar:.1f} / packed f32x2 {peak_packed:.1f} TFLOP/s "
                       f"(MEASURED_PEAKS.json has no FP32 CUDA-core figure; nominal 148 SM x 128 lanes x 2 x 1.965 GHz = 74.4)",
    }

    # ---- end-to-end arm: host buffers through the C ABI, ONE spt_cuda_render call per step ------------------
    # N = 1: the context of the timed region.  N > 1: rank 0 alone calls spt_cuda_render on a multi-device context over the N
    # GPUs of the job (spt_cuda_create_multi: the library splits the samples, combines the accumulators over NVLink with its
    # fused reduce-and-resolve kernel and copies the image back); the other ranks wait at the barrier with their GPUs idle.
    h2d = scene.n * (16 + 4 + 1 + 16 + 4) + 96 + 56
    d2h = W * H * 3
    e2e_steps = max(1, min(args.steps, 3))
    e2e_pt = pt
    if world > 1:
        barrier()
        dist.barrier(group=host_group)
        if rank == 0:
            e2e_pt = spt.PathTracer(list(range(world)))
            e2e_pt.set_staging_limit(staging)
            if args.partition == "bands":
                e2e_pt.set_partition(spt.PARTITION_BANDS)

    def step_e2e():
        e2e_pt.set_scene(scene)  # H2D: the scene globals (Globals.hpp:31-37), to every device of the context
        e2e_pt.render(cam, W, H, spp, depth, pass_cap=PASS_CAP, seed=SEED, rgb8=rgb_host.numpy(), variant=variant, flags=args.flags)  # D2H inside

    e2e_ms = 0.0
    if rank == 0:
        step_e2e()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            step_e2e()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
        e2e_launches = e2e_pt.timing()["launches"]
        if world > 1:
            e2e_pt.close()
    if world > 1:
        dist.barrier(group=host_group)  # host wait: no kernel on the peers' GPUs during rank 0's renders
    barrier()
    e2e = None
    if rank == 0:
        e2e = {"value": total_samples / (e2e_ms * 1e-3) / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d * world),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms, "launches_per_step": e2e_launches,
               "path": "spt_cuda_set_scene + spt_cuda_render (host buffers)" + (f" on one multi-device context over {world} GPUs (spt_cuda_create_multi), rank 0's process" if world > 1 else "")}

    # ---- the stress configurations, short legs after the headline (rank 0, N == 1) ---------------------------
    workloads = None
    if rank == 0 and world == 1 and args.workload == "c3" and not args.no_extra and not args.spp:
        workloads = {}
        for name in ("c4", "c5"):
            workloads[name] = short_leg(pt, spt, torch, cam, WORKLOADS[name], peak, staging)
        # the same two with SPT_FLAG_GRID: the OPT-IN acceleration structure of SURVEY 8(f)-4 (same bits; not the linear loop the
        # roofline is defined on -- its "roofline" fields only say how much faster than the reference's work rate it runs)
        for name in ("c4", "c5"):
            leg = short_leg(pt, spt, torch, cam, WORKLOADS[name], peak, staging, extra_flags=spt.FLAG_GRID)
            leg["note"] = "SPT_FLAG_GRID (opt-in uniform grid, bit-identical results); fractions are algorithmic work rates of the LINEAR loop, not kernel utilisation"
            workloads[name + "_grid"] = leg
        pt.set_scene(scene)

    # ---- CPU baseline (rank 0, N == 1 only) ---------------------------------------------------------
    cpu = cpu4 = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_reference_rate(wl, budget_s=12.0)
        cpu4 = cpu_reference_rate(wl, budget_s=5.0, threads=4)  # the reference's own g_maxThreads (Globals.hpp:11)
        cpu["g_maxThreads_4"] = {"value": cpu4["value"], "cores": 4, "sample": cpu4["sample"]}

    if rank == 0:
        line = {
            "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"{args.workload}: {wl['desc']}" + (f" [spp overridden to {spp}]" if args.spp else ""),
                       "spheres": scene.n, "width": W, "height": H, "spp": spp, "depth": depth, "pass_cap": PASS_CAP,
                       "variant": args.variant, "sharding": (f"interleaved 8-row bands of pixels over {world} rank(s)" if args.partition == "bands" else f"sample ranges over {world} rank(s)")
                                   + " + NCCL sum-reduce of float4 accumulators",
                       "staging": f"first-hit staging limit {staging / 2**30:.0f} GiB per GPU (spt_cuda_set_staging_limit; library default 8 GiB), "
                                  f"{trace_launches // 2 if split else 0} work item(s) (strips of tile rows x chunks of samples) per render",
                       "l2": f"float4 accumulator {W * H * 16 / 1e6:.1f} MB per step vs 126 MB L2; kernel is compute-bound, scene staged in shared memory"},
            "e2e": e2e, "gpu_launches": launches_native, "roofline": roofline, "cpu_baseline": cpu, "workloads": workloads, "clocks": clock_info,
            "device": info["name"], "sm_count": info["sm_count"], "counters": ctr,
        }
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def short_leg(pt, spt, torch, cam, wl, peak, staging, extra_flags=0):
    """One warm-up and one timed render of a stress configuration (BASELINE.json configs 4 / 5) on the context of the headline:
    Msamples/s with the scene resident, the roofline fractions, and what the scene costs to set up (host generation + upload)."""
    t0 = time.perf_counter()
    scene = load_scene(wl["scene"])
    t_gen = time.perf_counter() - t0
    t0 = time.perf_counter()
    pt.set_scene(scene)
    t_up = time.perf_counter() - t0
    W, H, spp, depth = wl["width"], wl["height"], wl["spp"], wl["depth"]
    accum = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    out = {}
    for rep in range(2):
        accum.zero_()
        pt.reset_counters()
        ev[0].record()
        pt.accumulate(cam, W, H, 0, spp, depth, accum, pass_cap=PASS_CAP, seed=SEED,
                      flags=spt.FLAG_COUNTERS | extra_flags | (spt.FLAG_SERIAL_CHUNKS if rep == 0 else 0))
        ev[1].record()
        torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[1])
        ctr = pt.counters()
        if rep == 0:  # warm-up doubles as the serialised pass that gives the kernels' own durations
            t = pt.timing()
            out["kernels"] = kernel_rooflines(ctr, scene.n, t["primary_ms"], t["secondary_ms"], peak)
    alg = FLOP_PER_TEST * ctr["tests"] / (ms * 1e-3) / 1e12
    out.update({"workload": wl["desc"], "spheres": scene.n, "value": W * H * spp / (ms * 1e-3) / 1e6, "unit": "Msamples/s", "ms_per_render": ms,
                "roofline_frac_step": alg / peak if peak else None,
                "roofline_frac": out["kernels"].get("trace_secondary", {}).get("frac"),
                "segments_per_sample": ctr["segments"] / max(ctr["samples"], 1),
                "scene_setup_ms": {"host_generation": t_gen * 1e3, "spt_cuda_set_scene": t_up * 1e3,
                                   "share_of_one_render": (t_gen + t_up) * 1e3 / ms}})
    return out


_REAL_STDOUT = None


def claim_stdout():
    """The contract is ONE JSON line on stdout; libraries (NCCL's version banner, for one) also write to fd 1.
    Point fd 1 at stderr for everybody else and keep the real stdout for the result line."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel (debug only; invalidates the headline)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--variant", default="megakernel", choices=["megakernel", "wavefront"])
    ap.add_argument("--partition", default="samples", choices=["samples", "bands"], help="how N > 1 ranks share a render (sharding.py)")
    ap.add_argument("--flags", type=int, default=0, help="extra SPT_FLAG_* bits for the trace launches (experiments)")
    ap.add_argument("--staging-mib", type=int, default=0, help="first-hit staging limit per GPU in MiB (0 = the library default, 8 GiB)")
    ap.add_argument("--strip-rows", type=int, default=0, help="SptParams.strip_rows (0 = chosen by the library; experiments)")
    ap.add_argument("--split-chunk", type=int, default=0, help="SptParams.split_chunk (0 = chosen by the library; experiments)")
    ap.add_argument("--no-extra", action="store_true", help="skip the short c4 / c5 legs after the headline")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference_arm(args, wl)
    else:
        run_native_arm(args, wl)


if __name__ == "__main__":
    main()
