#!/usr/bin/env python
"""bench.py — throughput of the path-tracing hot path (raygen -> intersect -> shade/NEE -> shadow intersect ->
resolve -> accumulate) on N B200s, one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference]

A "step" renders `frames` frames (samples per pixel) of the workload's scene at its resolution on every GPU
(weak scaling: GPU g of N renders frames g, g+N, ...; the per-step sum images are reduced with one NCCL
all-reduce and normalised).  metric = live ray segments traversed per second (primary + light-sampling rays
whose maxDistance >= 0 at intersect time), the "Mrays/s" of BASELINE.json; Msamples/s (= what the reference's
window title calls Mrays/s, renderer/Renderer.mm:631-637) is reported beside it.

  value  : inputs (noise tiles) already resident in HBM, timed with CUDA events on the launching stream.
  e2e    : the same steps through the public API with HOST buffers — pinned noise tiles copied host->device and
           the accumulated image copied device->host inside the timed region of every step.
  --impl reference : the CPU oracle (scalar C++ restatement of the reference's Metal kernels; the Metal/MPS
           original cannot run on Linux) on all host cores, a bounded number of frames per step.
"""
import argparse
import json
import os
import pathlib
import subprocess
import sys
import tempfile
import threading
import time

ROOT = pathlib.Path(__file__).resolve().parent
sys.path[:0] = [str(ROOT / "metal-renderer_b200")]

import numpy as np  # noqa: E402

WORKLOADS = {   # BASELINE.json configs; SURVEY.md §8(d) "Concrete inputs"
    "spheres-1m-1080p": dict(scene="spheres-1m", seed=3, width=1920, height=1080, frames=128, depth=8, config=3),
    "cornell-water-1080p": dict(scene="cornell-water", seed=2, width=1920, height=1080, frames=64, depth=8, config=2),
    # the reference's own Media/CornellBox-Water-plastic.obj (ABI arrays committed by tests/golden/make_media_fixture.py)
    "media-water-plastic-1080p": dict(scene="media:CornellBox-Water-plastic", seed=2, width=1920, height=1080, frames=64, depth=8, config=2),
    "cornell-256": dict(scene="cornell", seed=1, width=256, height=256, frames=16, depth=4, config=1),
    "spheres-10m-4k": dict(scene="spheres-10m", seed=4, width=3840, height=2160, frames=256, depth=8, config=4),
    "many-lights-1080p": dict(scene="many-lights", seed=5, width=1920, height=1080, frames=512, depth=8, config=5),
}
DEFAULT_WORKLOAD = "spheres-1m-1080p"
BYTES_PER_SEGMENT = 376      # 48 intersect + 224 shade + 48 shadow intersect + 56 resolve
BYTES_PER_SAMPLE = 140       # raygen 16+80, accumulate 12+16+16
BYTES_PER_RAY_TRACE = 48     # one intersect: 32 B of ray read + 16 B Intersection written
NODE_BYTES = 80              # 8-wide quantised node (csrc/bvh8.h)
TRIANGLE_RECORD_BYTES = 48   # leaf-ordered triangle record (csrc/bvh.h)
PARITY_FRAMES = {"spheres-10m-4k": 1}     # frames of the untimed parity step (default 2); the oracle renders them on the host cores


def ncu_summary(kernel, workload):
    """Counters of one steady-state launch of `kernel` ("trace" / "shade") on `workload`, parsed from the committed
    summary profiles/ncu_<kernel>_<workload>.txt (written by scripts/ncu_collect.py from an `ncu --set full` report of
    HEAD's kernels).  None when there is no capture of this workload — never a number from another scene."""
    path = ROOT / "profiles" / f"ncu_{kernel}_{workload}.txt"
    if not path.exists():
        return None
    out = {"source": str(path.relative_to(ROOT))}
    for line in path.read_text().splitlines():
        parts = line.split()
        if line.startswith("#"):
            if "commit" in parts:
                out["commit"] = parts[parts.index("commit") + 1]
            if "rays_in_launch" in parts:
                out["rays_in_launch"] = float(parts[parts.index("rays_in_launch") + 1])
            continue
        if len(parts) >= 2:
            try:
                out[parts[0]] = float(parts[1].replace(",", ""))
            except ValueError:
                pass
    return out


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def __enter__(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "200"], stdout=f, stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None
        return self

    def __exit__(self, *exc):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if not self.path or not os.path.exists(self.path):
            return out
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.path)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


def load_scene(m, wl):
    if wl["scene"].startswith("media:"):
        return m.Scene.from_npz(ROOT / "tests" / "golden" / "media_water.npz", wl["scene"][6:])
    return m.Scene.procedural(wl["scene"], wl["seed"])


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# ------------------------------------------------------------------------------------------------------------------
def run_reference(args, wl):
    """CPU arm: the oracle on the box's host cores, `ref_frames` frames of the same workload per step."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    sys.path.insert(0, str(ROOT / "oracle"))
    import mrt_b200 as m
    import oracle as orc
    scene = load_scene(m, wl)
    o = orc.Oracle(max_path_length=wl["depth"], threads=0)
    o.init_raytracing(scene)
    o.drawable_size_will_change(wl["width"], wl["height"])
    init = m.noise_tile(wl["seed"], 0xFFFFFFFF)
    for s in range(3):
        o.update_noise(s, init)
    f = args.ref_frames
    total = (args.warmup + args.steps) * f
    tiles = m.noise_tiles(wl["seed"], 0, total)
    frame = 0
    for _ in range(args.warmup):
        o.draw(f, tiles[frame:frame + f], first_frame=frame); frame += f
    o.reset_counts()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        o.draw(f, tiles[frame:frame + f], first_frame=frame); frame += f
    dt = time.perf_counter() - t0
    c = o.counts()
    rays = c["primary_rays"] + c["shadow_rays"]
    value = rays / dt / 1e6
    sample = f"{f} frame(s) of {wl['width']}x{wl['height']} per step, {args.steps} steps"
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "msamples_per_s": wl["width"] * wl["height"] * f * args.steps / dt / 1e6,
        "config": workload_config(wl, args, 1, frames=f),
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": o.threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference = scalar C++ restatement of Shaders.metal (oracle/); the Metal/MPS original is unavailable on this box",
    }
    print(json.dumps(line), flush=True)


def workload_config(wl, args, world, frames=None):
    return {"workload": args.workload, "baseline_config": wl["config"], "scene": wl["scene"], "seed": wl["seed"],
            "width": wl["width"], "height": wl["height"], "frames_per_step_per_gpu": frames or wl["frames"],
            "max_path_length": wl["depth"], "parallelism": f"spp-shard x{world}" if world > 1 else "single",
            "l2": "working set (BVH + triangles + shading records) exceeds the 126 MB L2 for the 1M/10M-triangle workloads; "
                  "rays of consecutive steps differ (new noise, new frame indices)"}


def cpu_baseline(wl, m, scene, budget_s=15.0):
    sys.path.insert(0, str(ROOT / "oracle"))
    import oracle as orc
    o = orc.Oracle(max_path_length=wl["depth"], threads=0)
    o.init_raytracing(scene)
    o.drawable_size_will_change(wl["width"], wl["height"])
    init = m.noise_tile(wl["seed"], 0xFFFFFFFF)
    for s in range(3):
        o.update_noise(s, init)
    tiles = m.noise_tiles(wl["seed"], 0, 9)
    t0 = time.perf_counter(); o.draw(1, tiles[:1]); one = time.perf_counter() - t0       # warm-up frame, sizes the sample
    n = int(max(1, min(8, budget_s / max(one, 1e-3))))
    o.reset_counts()
    t0 = time.perf_counter(); o.draw(n, tiles[1:1 + n], first_frame=1); dt = time.perf_counter() - t0
    c = o.counts()
    rays = c["primary_rays"] + c["shadow_rays"]
    return {"value": rays / dt / 1e6, "unit": "Mrays/s", "cores": o.threads, "kind": "port",
            "msamples_per_s": wl["width"] * wl["height"] * n / dt / 1e6,
            "sample": f"{n} frame(s) of {wl['width']}x{wl['height']} ({wl['scene']}, depth {wl['depth']}) after 1 warm-up frame, {dt:.1f} s"}


# ------------------------------------------------------------------------------------------------------------------
def rel_rmse(got, want):
    ok = np.isfinite(got).all(axis=-1) & np.isfinite(want).all(axis=-1)
    a, b = got[ok].astype(np.float64), want[ok].astype(np.float64)
    return float(np.sqrt(np.mean((a - b) ** 2)) / max(np.sqrt(np.mean(b ** 2)), 1e-30)), float(ok.mean())


def parity_step(args, wl, m, D, scene, r, ep, coll, rank, world):
    """Untimed: the gates of BASELINE.md 3 on the benchmarked workload, against the oracle on the host cores (rank 0).
    N = 1: `k` frames at full size — primary-hit triangle ids of every pixel and the accumulated image.
    N > 1: one ShardedRenderer step of k frames per rank; the all-reduced image is held against the oracle's SEQUENTIAL
    render of the same k*N frames on the rows y % stride == 0 (stride bounds the oracle's work), and the primary hits
    of frame 0 on every pixel."""
    import torch
    sys.path.insert(0, str(ROOT / "oracle"))
    W, H, D_, seed = wl["width"], wl["height"], wl["depth"], wl["seed"]
    k = PARITY_FRAMES.get(args.workload, 2)
    init = m.noise_tile(seed, 0xFFFFFFFF)
    r.sync()
    r.clear_image()
    for s in range(3):
        r.update_noise(s, init)
    # primary hits through the staged entry points (raygen + intersect of frame 0)
    r.raygen(0); r.intersect(m.BUF_RAYS)
    hits = r.read_buffer(m.BUF_INTERSECTIONS)["triangleIndex"].copy() if rank == 0 else None
    frames = k * world
    sr = D.ShardedRenderer(ep, rank, world, coll, mode="samples", reduce=args.reduce)   # fresh accumulation state (world 1: running mean)
    plan = D.plan_sample_shard(0, k, rank, world)
    tiles = m.noise_tiles(seed, plan.first_frame, plan.span)
    lead = {f: m.noise_tile(seed, f) for f in D.lead_in_frames(plan.first_frame)}
    if world > 1:
        r.clear_image()
    sr.step(k, tiles, lead_in=lead, sync=True)
    got = ep.read(sr.output()).copy()
    out = None
    if rank == 0:
        import oracle as orc
        t0 = time.perf_counter()
        stride = 1 if world == 1 else max(1, min(16, world))
        o = orc.Oracle(max_path_length=D_, threads=0)
        o.init_raytracing(scene)
        o.drawable_size_will_change(W, H)
        for s in range(3):
            o.update_noise(s, init)
        o.raygen(0); o.intersect(0)
        want_hits = o.read_buffer(2)["triangleIndex"]
        agree = float(np.mean(hits == want_hits))
        if stride > 1:
            o.set_row_shard(0, stride)
        o.draw(frames, m.noise_tiles(seed, 0, frames))
        want = o.save_current_image()
        rr, finite = rel_rmse(got[::stride, :, :3], want[::stride, :, :3])
        same_bits = float(np.mean(got[::stride].view(np.uint32) == want[::stride].view(np.uint32)))
        out = {"hit_id_agree": agree, "rrmse": rr, "frames": frames, "pixels": int(W * H), "rows_compared": int(len(range(0, H, stride))),
               "image_words_bit_equal": same_bits, "finite_pixels": finite, "oracle_threads": o.threads,
               "oracle_s": time.perf_counter() - t0,
               "gates": {"hit_id_agree_min": 0.9999, "rrmse_max": 1e-3},
               "pass": bool(agree >= 0.9999 and rr <= 1e-3),
               "what": ("wavefront pipeline vs oracle, same frames, full image" if world == 1 else
                        f"all-reduced image of one {world}-rank step vs the oracle's sequential render of the same {frames} frames on rows y % {stride} == 0")}
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    del torch
    return out


def run_ours(args, wl):
    import torch
    import mrt_b200 as m
    from mrt_b200 import distributed as D
    rank, local_rank, world = dist_env()
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = local_rank
    torch.cuda.set_device(dev)
    W, H, F, Dp = wl["width"], wl["height"], wl["frames"], wl["depth"]
    t_setup = time.perf_counter()
    scene = load_scene(m, wl)
    t_scene = time.perf_counter()
    r = m.Renderer(device=dev, max_path_length=Dp)
    r.init_raytracing(scene)
    r.drawable_size_will_change(W, H)
    r.sync()
    setup_ms = {"scene_generation": (t_scene - t_setup) * 1e3, "bvh_build_upload_resize": (time.perf_counter() - t_scene) * 1e3}
    init = m.noise_tile(wl["seed"], 0xFFFFFFFF)
    for s in range(3):
        r.update_noise(s, init)
    ep = D.CudaEndpoint(r, dev)
    coll = D.TorchCollective()

    n_tiles = F * world                               # tile j belongs to frame (step start + j); reused every step
    span = (F - 1) * world + 1                        # tiles one rank consumes: frames first .. first+(F-1)*world
    host_tiles = torch.empty((n_tiles, m.NOISE_TILE_FLOATS), dtype=torch.float32, pin_memory=True)
    m.noise_tiles(wl["seed"], 0, n_tiles, out=host_tiles.numpy())
    dev_tiles = host_tiles.to(f"cuda:{dev}")
    host_image = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True)
    torch.cuda.synchronize()
    tile_bytes = m.NOISE_TILE_FLOATS * 4

    def make_stepper(frames_per_rank, reduce=None):
        """One step = ShardedRenderer.step (mrt_b200/distributed.py — the code the sharding tests run): `frames_per_rank`
        frames on every rank, one NCCL all-reduce of the sum image, normalisation.  Tiles are addressed inside the
        resident (device) or pinned (host) tile array: tile j = frame (step start + j) modulo the array."""
        sr = D.ShardedRenderer(ep, rank, world, coll, mode="samples", reduce=reduce or args.reduce)
        n = frames_per_rank * world

        def step(tiles_ptr, readback):
            start = sr.next_frame
            plan = D.plan_sample_shard(start, frames_per_rank, rank, world)
            lead = {f: tiles_ptr + ((f - start) % n) * tile_bytes for f in D.lead_in_frames(plan.first_frame)} if world > 1 else None
            sr.step(frames_per_rank, tiles_ptr + rank * tile_bytes, lead_in=lead)
            if readback:
                ep.read(sr.output(), host_image)      # D2H of the job's image + sync
        return sr, step

    def barrier():
        r.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def reduce_ranks(x, op="max"):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{dev}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
        return float(t.item())

    reduce_note = None
    if world == 1:
        args.reduce = "nccl"                            # nothing to reduce; ShardedRenderer ignores it
    elif args.reduce == "auto":
        bad = 0.0
        try:
            sr_t, step_t = make_stepper(1, reduce="peer")    # raises on every rank together if one cannot map a peer
            step_t(dev_tiles.data_ptr(), False)
            r.peer_status()                                  # synchronises; a wait that gave up is an error
        except Exception as e:                               # noqa: BLE001
            bad, reduce_note = 1.0, str(e)[:200]
        args.reduce = "nccl" if reduce_ranks(bad) > 0 else "peer"
    sr, step = make_stepper(F)
    for _ in range(max(args.warmup, 3)):
        step(dev_tiles.data_ptr(), False)
    barrier()
    r.reset_stats()
    if not args.no_trace_events:
        r.set_profile(m.PROFILE_TRACE_TIME)  # CUDA-event pair around every trace-kernel launch (the dominant kernel) on the ctx stream

    # ---- timed: inputs resident in HBM, CUDA events on the launching stream --------------------------------------
    with ClockSampler(dev) as clocks:
        barrier()
        r.event_record(0)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step(dev_tiles.data_ptr(), False)
        r.event_record(1)
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        dev_ms = r.event_elapsed_ms(0, 1)
    if world > 1 and args.reduce == "peer":
        try:
            r.peer_status()
        except Exception as e:                          # noqa: BLE001 — reported in the line; the parity step below judges the image
            reduce_note = f"a peer wait gave up inside the timed region: {e}"[:200]
    st = r.stats()
    ms = reduce_ranks(dev_ms)
    rays_local = st["primary_rays"] + st["shadow_rays"]
    rays = reduce_ranks(rays_local, "sum")
    samples = float(W) * H * F * world * args.steps
    value = rays / (ms * 1e-3) / 1e6
    launches = st["kernel_launches"]
    trace_ms, trace_launches = st["profile_trace_ms"], st["profile_trace_launches"]      # last step of the timed region
    tail_waves = st["tail_waves"]
    r.set_profile(0)

    # ---- e2e: host buffers, H2D of the noise tiles and D2H of the image inside every step; default schedule (one CUDA
    # graph per batch, no per-launch events) ------------------------------------------------------------------------
    step(host_tiles.data_ptr(), True)        # untimed: first use of the H2D / D2H path and of the wave graph
    barrier()
    r.reset_stats()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(host_tiles.data_ptr(), True)
    barrier()
    e2e_s = reduce_ranks(time.perf_counter() - t0)
    st2 = r.stats()
    e2e_rays = reduce_ranks(st2["primary_rays"] + st2["shadow_rays"], "sum")
    finite = bool(torch.isfinite(host_image[..., :3]).float().mean().item() > 0.999)

    # ---- untimed: one more step through the COUNTING variant of the trace kernel — node tests and triangle tests per ray
    # of this workload, the data-dependent "BVH/scene term" of SURVEY 8(d)'s algorithmic bytes
    barrier()
    r.reset_stats()
    r.set_profile(m.PROFILE_VISITS)
    step(dev_tiles.data_ptr(), False)
    barrier()
    st3 = r.stats()
    r.set_profile(0)
    counted_rays = max(st3["primary_rays"] + st3["shadow_rays"], 1)
    node_tests_per_ray = st3["nodes_visited"] / counted_rays
    tri_tests_per_ray = st3["triangles_tested"] / counted_rays

    # ---- strong scaling (N > 1): the SAME `frames` samples per pixel split over the ranks ---------------------------
    strong = None
    if world > 1 and not args.no_strong:
        fs = max(1, F // world)
        sr_s, step_s = make_stepper(fs)
        sr_s.time_events = (2, 3)                      # CUDA events on the ctx stream around the collective
        for _ in range(3):
            step_s(dev_tiles.data_ptr(), False)
        barrier()
        r.reset_stats()
        r.set_profile(m.PROFILE_TRACE_TIME)
        ar_ms = []
        r.event_record(4)
        for _ in range(args.steps):
            step_s(dev_tiles.data_ptr(), False)
            r.sync()
            ar_ms.append(r.event_elapsed_ms(2, 3))
        r.event_record(5)
        barrier()
        s_ms = reduce_ranks(r.event_elapsed_ms(4, 5))
        sts = r.stats()
        r.set_profile(0)
        s_rays = reduce_ranks(sts["primary_rays"] + sts["shadow_rays"], "sum")
        strong = {"frames_total_per_step": fs * world, "frames_per_gpu": fs, "ms_total": s_ms / args.steps,
                  "mrays_per_s": s_rays / (s_ms * 1e-3) / 1e6, "msamples_per_s": float(W) * H * fs * world * args.steps / (s_ms * 1e-3) / 1e6,
                  "allreduce_ms": reduce_ranks(float(np.median(ar_ms))), "allreduce_bytes": int(W * H * 16),
                  "waves": sts["waves_traced"], "tail_waves": sts["tail_waves"],
                  "bvh_build_ms": reduce_ranks(st["bvh_build_ms"] + st["bvh8_build_ms"]),
                  "note": "same samples per pixel as one GPU's step, split over the ranks (frames_per_gpu = frames / N); ms_total is "
                          "the device time of a step incl. the all-reduce, max over ranks; allreduce_ms is the event-timed collective "
                          "incl. the wait for the slowest rank; tail_waves = waves with fewer extension rays than half the path slots"}

        # the same fixed-work step with the other reduction (csrc/peer.cu: one kernel per GPU over NVLink peer memory, no
        # NCCL on the data path), so that both are on record whichever one --reduce selected
        other = "peer" if args.reduce == "nccl" else "nccl"
        try:
            sr_o, step_o = make_stepper(fs, reduce=other)
            sr_o.time_events = (2, 3)
            for _ in range(3):
                step_o(dev_tiles.data_ptr(), False)
            barrier()
            o_ms = []
            r.event_record(4)
            for _ in range(args.steps):
                step_o(dev_tiles.data_ptr(), False)
                r.sync()
                o_ms.append(r.event_elapsed_ms(2, 3))
            r.event_record(5)
            barrier()
            if other == "peer":
                r.peer_status()
            strong["other_reduce"] = {"reduce": other, "ms_total": reduce_ranks(r.event_elapsed_ms(4, 5)) / args.steps,
                                      "reduce_ms": reduce_ranks(float(np.median(o_ms)))}
        except Exception as e:                          # noqa: BLE001 — e.g. no CUDA IPC between the ranks' processes
            strong["other_reduce"] = {"reduce": other, "unavailable": str(e)[:300]}
        strong["reduce"] = args.reduce

    parity = None if args.no_parity else parity_step(args, wl, m, D, scene, r, ep, coll, rank, world)

    if rank == 0:
        peak, peak_kind = measured_peaks()
        # dominant kernel = k_wf_trace.  SURVEY 8(d): per ray it reads 32 B of ray and writes a 16-B Intersection (48 B for
        # the primary and 48 B for the light-sampling intersect) PLUS the BVH/scene term nodes_visited x sizeof(node) +
        # tris_tested x sizeof(triangle record), with the visit counts of THIS workload from the device counters above;
        # per launch = that x rays traced in the launch.  The stream-only figure (48 B/ray) is reported beside it.
        rays_per_step = rays_local / args.steps
        alg_bytes_step = BYTES_PER_SAMPLE * W * H * F + BYTES_PER_SEGMENT * st["primary_rays"] / args.steps
        trace_launches = max(trace_launches, 1)
        bvh_bytes_per_ray = node_tests_per_ray * NODE_BYTES + tri_tests_per_ray * TRIANGLE_RECORD_BYTES
        rays_per_launch = rays_per_step / trace_launches
        stream_bytes = BYTES_PER_RAY_TRACE * rays_per_launch
        alg_bytes = (BYTES_PER_RAY_TRACE + bvh_bytes_per_ray) * rays_per_launch
        per_launch_ms = trace_ms / trace_launches
        achieved = alg_bytes / max(per_launch_ms * 1e-3, 1e-12) / 1e9
        achieved_stream = stream_bytes / max(per_launch_ms * 1e-3, 1e-12) / 1e9
        # what ncu saw on HEAD's kernels for this workload (profiles/ncu_{trace,shade}_<workload>.txt), or null
        nt, ns = ncu_summary("trace", args.workload), ncu_summary("shade", args.workload)
        traffic = issue = None
        if nt:
            traffic = nt.get("dram__bytes_read.sum", 0.0) + nt.get("dram__bytes_write.sum", 0.0)
            ia = nt.get("smsp__issue_active.avg.pct_of_peak_sustained_active")
            lanes = nt.get("smsp__thread_inst_executed_per_inst_executed.ratio")
            inst = nt.get("smsp__inst_executed.sum")
            issue = {"warp_inst_per_launch": inst, "issue_active": ia / 100.0 if ia is not None else None, "lanes_per_inst": lanes,
                     "thread_issue_frac": (ia / 100.0) * (lanes / 32.0) if ia is not None and lanes else None,
                     "thread_inst_per_ray": inst * lanes / nt["rays_in_launch"] if inst and lanes and nt.get("rays_in_launch") else None,
                     "warp_inst_per_ray": inst / nt["rays_in_launch"] if inst and nt.get("rays_in_launch") else None,
                     "rays_in_profiled_launch": nt.get("rays_in_launch"),
                     "profiled_launch_ms": nt.get("gpu__time_duration.sum", 0.0) / 1e6 if nt.get("gpu__time_duration.sum") else None,
                     "peak_thread_inst_per_s": 148 * 4 * 32 * 1.965e9,
                     "note": "SURVEY 8(d): issue = warp instructions / (cycles x 4 schedulers x 148 SMs) = issue_active; x lanes/32 "
                             "= share of the THREAD-instruction issue peak.  This, not `frac`, is the bound that binds"}
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": dict(workload_config(wl, args, world), **({"reduce": {"peer": "mrt_peer_reduce: one kernel per GPU over NVLink peer memory (csrc/peer.cu)", "nccl": "copy + NCCL all-reduce + scale"}[args.reduce], **({"reduce_note": reduce_note} if reduce_note else {})} if world > 1 else {})),
            "msamples_per_s": samples / (ms * 1e-3) / 1e6,
            "rays_per_sample": rays / samples,
            "wall_ms_per_step": wall_ms / args.steps,
            "e2e": {"value": e2e_rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int((span + (2 if world > 1 else 0)) * tile_bytes),
                    "d2h_bytes_per_step": int(host_image.numel() * 4), "msamples_per_s": samples / e2e_s / 1e6,
                    "image_finite": finite, "gpu_launches_per_step": int(st2["kernel_launches"] / args.steps),
                    "schedule": "one CUDA graph per batch (WHILE node ends the wave loop on the device)"},
            "gpu_launches": int(launches),
            "clocks": clocks.summary(),
            "parity": parity,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic,
                         "source": {"trace": nt and {k: nt.get(k) for k in ("source", "commit")}, "shade": ns and {k: ns.get(k) for k in ("source", "commit")}},
                         "peak_kind": peak_kind, "kernel": "k_wf_trace", "algorithmic_bytes_per_launch": alg_bytes,
                         "algorithmic_bytes_per_ray": {"ray_and_intersection": BYTES_PER_RAY_TRACE, "bvh_term": bvh_bytes_per_ray,
                                                       "node_tests": node_tests_per_ray, "triangle_tests": tri_tests_per_ray,
                                                       "node_bytes": NODE_BYTES, "triangle_record_bytes": TRIANGLE_RECORD_BYTES},
                         "stream_only": {"achieved": achieved_stream, "frac": achieved_stream / peak,
                                         "algorithmic_bytes_per_launch": stream_bytes},
                         "launch_ms": per_launch_ms, "launches_per_step": trace_launches, "rays_per_launch": rays_per_launch,
                         "kernel_share_of_step": trace_ms / max(dev_ms / args.steps, 1e-9),
                         "in_kernel_mrays_per_s": rays_per_launch / max(per_launch_ms * 1e-3, 1e-12) / 1e6,
                         "step_algorithmic_bytes": alg_bytes_step,
                         "step_achieved_gbs": alg_bytes_step / (dev_ms / args.steps * 1e-3) / 1e9,
                         "issue": issue,
                         "shade": ns and {"launch_ms": ns.get("gpu__time_duration.sum", 0.0) / 1e6,
                                          "dram_bytes": ns.get("dram__bytes_read.sum", 0.0) + ns.get("dram__bytes_write.sum", 0.0),
                                          "dram_pct_of_peak": ns.get("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
                                          "lanes_per_inst": ns.get("smsp__thread_inst_executed_per_inst_executed.ratio")},
                         "tail_waves_per_step": tail_waves,
                         "note": "algorithmic bytes per ray = 48 B (32 B ray + 16 B Intersection) + node tests x 80 B + triangle tests "
                                 "x 48 B (SURVEY 8d incl. its BVH/scene term; visit counts of this workload from the device "
                                 "counters, untimed extra step).  The gathers of the BVH term are served mostly by L1/L2 — "
                                 "`traffic` = DRAM bytes of one steady-state launch (ncu --set full, the file named in `source`; null "
                                 "if this workload has no capture) — so `frac` is the contract's algorithmic-bytes rate over the HBM "
                                 "peak, not DRAM utilisation, and it rises when the tree gets worse; `issue` is the bound that binds"},
            "bvh": {"nodes": st["bvh_nodes"], "depth": st["bvh_depth"], "sah_cost": st["bvh_sah_cost"], "build_ms": st["bvh_build_ms"],
                    "wide_nodes": st["bvh8_nodes"], "wide_depth": st["bvh8_depth"], "wide_build_ms": st["bvh8_build_ms"]},
            "setup_ms": setup_ms,
        }
        if strong is not None:
            line["strong"] = strong
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(wl, m, scene)
        print(json.dumps(line), flush=True)
    if world > 1:
        try:
            r.sync()
            r.peer_disconnect()                         # let go of the peers' exchange blocks (CUDA IPC) before anybody frees its own
        except Exception:                               # noqa: BLE001 — nothing was connected
            pass
        dist.barrier()                                  # from here on any rank may free its block (at interpreter exit: torch tensors
        dist.destroy_process_group()                    # allocated on the ctx stream must die before the ctx does, so no r.close() here)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--frames", type=int, default=0, help="override frames per step per GPU")
    ap.add_argument("--ref-frames", type=int, default=1, help="frames per step of the CPU reference arm")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the untimed parity step against the oracle")
    ap.add_argument("--no-strong", action="store_true", help="N > 1: skip the fixed-work (strong scaling) measurement")
    ap.add_argument("--reduce", default="auto", choices=["auto", "nccl", "peer"],
                    help="N > 1: framebuffer reduction = mrt_peer_reduce (one kernel per GPU over NVLink peer memory, csrc/peer.cu) "
                         "or copy + NCCL all-reduce + scale; auto = the peer kernel if every rank can map its peers (CUDA IPC) and "
                         "one trial step succeeds on all of them, else NCCL")
    ap.add_argument("--no-trace-events", action="store_true",
                    help="diagnostic: timed region without the per-launch event pairs (the roofline object is then meaningless)")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    if args.frames:
        wl["frames"] = args.frames
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    # The contract is ONE line on stdout.  Libraries write there too (NCCL prints its version banner on rank 0 when the
    # box sets NCCL_DEBUG), so everything but the result line goes to stderr: fd 1 is pointed at fd 2 for the run and the
    # JSON line is written to the saved descriptor.
    sys.stdout.flush()
    _real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    _print = print

    def print(*a, **k):       # noqa: A001 — the result lines of run_ours / run_reference
        k.pop("flush", None)
        _print(*a, file=_real_stdout, **k)
        _real_stdout.flush()
    main()
