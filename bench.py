#!/usr/bin/env python
"""bench.py -- BASELINE.json's headline metric on its named config.

  metric   Msamples/s (and Mrays/s) of the RTiOW final scene, 1920x1080, 500 spp, max_depth 50
           (BASELINE.json configs[2], "C3"; it fits one GPU, so it is also the N=1 workload)
  step     one full frame: render (all 500 spp, sample-range sharded over the ranks) -> combine on
           rank 0 -> resolve to the 8-bit image, everything resident on the device (`value`);
           `e2e` = the same frame through the public API with HOST buffers: scene upload (H2D) +
           render + resolve + P3 encode + read-back of rgb8 and the P3 text (D2H) every step, frames
           back to back (the copies of a frame overlap the next frame's render; `one_frame_alone_ms`
           is a frame on its own).
  N > 1    one process per GPU (torchrun), strong scaling of the one frame; partial accumulation
           buffers are combined by an NCCL reduce (default) or the fused IPC peer-reduce resolve.

`--impl reference` times the reference's own CPU code (oracle/_ref, the reference compiled in
place; else the C oracle port) on this box's host cores on a bounded sample of the same frame.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WIDTH, SPP, DEPTH = 1920, 500, 50
WORKLOAD = "C3: RTiOW final scene (485 spheres, main.cpp generator), 1920x1080, 500 spp, max_depth 50"
F_RAY = 528.0            # algorithmic flops per ray at N = 485: 48*ceil(log2 N) + 96
B_RAY = 608.0            # algorithmic L2 bytes per ray at N = 485: 64*ceil(log2 N) + 32

# The other BASELINE.json configs, run through the same step / sharding / timing code with --config
# (SURVEY.md 8d "C3 / C4 / C5 inputs").  The default (c3) is the headline config of the bench contract.
#   name: (workload text, width, height, spp, generator half-width g, p_diffuse, p_metal, BVH builder)
CONFIGS = {
    "c3": (WORKLOAD, 1920, 1080, 500, 11, 0.8, 0.95, "sah"),
    "c4": ("C4: 80 % glass/metal final-scene generator (diffuse < 0.2 <= metal < 0.6 <= glass), 1920x1080, 1000 spp, max_depth 50",
           1920, 1080, 1000, 11, 0.2, 0.6, "sah"),
    "c5-10k": ("C5: 10 k random spheres (g = 50), 3840x2160, 64 spp, max_depth 50", 3840, 2160, 64, 50, 0.8, 0.95, "lbvh"),
    "c5-100k": ("C5: 100 k random spheres (g = 159), 3840x2160, 64 spp, max_depth 50", 3840, 2160, 64, 159, 0.8, 0.95, "lbvh"),
    "c5-1m": ("C5: 1 M random spheres (g = 500), 3840x2160, 64 spp, max_depth 50", 3840, 2160, 64, 500, 0.8, 0.95, "lbvh"),
}


def peaks():
    p = {}
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    return p


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [c for c in sm if c > 0]
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_run(spp_sample, threads, steps=1, warmup=0, use_bvh=False):
    """The reference's CPU path on a bounded sample of the C3 frame: full 1920x1080 resolution,
    `spp_sample` of the 500 samples per pixel; linear hittable_list exactly as shipped, or (use_bvh) the
    reference's bvh_node::hit (src/bvh.h:16-24, made to compile and given a builder by the harness)."""
    from oracle import oracle as O
    s, m = O.scene_rtiow()
    cam = O.camera_rtiow(WIDTH, spp_sample, DEPTH)
    kind = "reference" if O.have_ref() and os.path.exists(O.LIB_REF) else "port"
    impl = "ref" if kind == "reference" else "oracle"
    times, rays = [], 0
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        r = O.render(s, m, cam, rng_mode=O.RNG_PHILOX, seed=0, nthreads=threads, impl=impl, use_bvh=use_bvh)
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt); rays = r["rays"]
    dt = sum(times) / len(times)
    samples = WIDTH * r["height"] * spp_sample
    world = ("bvh_node::hit over the harness-built tree (src/bvh.h:16-24; the reference ships no builder)" if use_bvh else
             "linear hittable_list (no BVH, as src/main.cpp:12,70)")
    return {"value": samples / dt / 1e6, "unit": "Msamples/s", "cores": threads, "kind": kind,
            "mrays_per_s": rays / dt / 1e6, "seconds_per_step": dt,
            "sample": f"{WIDTH}x{r['height']} at {spp_sample} of {SPP} spp, depth {DEPTH}, {world}, keyed Philox stream, "
                      f"{threads} host threads over rows"}


def cpu_as_shipped_run():
    """The reference exactly as shipped: src/main.cpp unmodified (400x225, 30 spp, depth 50, global mt19937,
    ONE thread, linear hittable_list), stdout captured by the harness.  ~25-40 s.  None without the reference build."""
    from oracle import oracle as O
    if not (O.have_ref() and os.path.exists(O.LIB_REF)):
        return None
    t0 = time.perf_counter()
    ppm = O.ref_main_ppm()
    dt = time.perf_counter() - t0
    samples = 400 * 225 * 30
    return {"value": samples / dt / 1e6, "unit": "Msamples/s", "cores": 1, "kind": "reference", "seconds": dt,
            "mrays_per_s": 6968725 / dt / 1e6,
            "sample": f"src/main.cpp as shipped: 400x225, 30 spp, depth 50, mt19937, 1 thread, linear list; {len(ppm)} bytes of P3 "
                      "(6 968 725 rays, SURVEY.md 8c); the C3 frame is 384x this many samples"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    spp_sample = max(1, args.ref_spp)
    cb = cpu_reference_run(spp_sample, threads, steps=args.steps, warmup=min(args.warmup, 1))
    line = {"impl": "reference", "metric": "Msamples/s", "value": cb["value"], "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": cb["seconds_per_step"] * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "timed_sample": cb["sample"]},
            "mrays_per_s": cb["mrays_per_s"],
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_c2(args):
    """BASELINE config 2: closest-hit microbenchmark, 2^26 synthetic rays generated in-kernel against the final-scene
    BVH (isolates traversal).  One GPU (rank 0; the path has no multi-GPU form).  `value` = device time of the
    trace kernel alone, rays generated in-kernel, results left on the device; `e2e` = the rt_bench_closest call with the
    per-ray results (prim id + t, 8 B per ray) copied to host arrays.  Untimed: exact primitive-id check of 2^20 of the
    rays against the oracle's linear list (hittable_list::hit, reference src/hittable_list.h:25-39)."""
    import numpy as np
    import torch
    import raytracingproject_b200 as rt
    if int(os.environ.get("RANK", "0")) != 0:
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback")
    n = 1 << 26
    prec = rt.F32 if args.precision == "f32" else rt.F64
    ctx = rt.Context(int(os.environ.get("LOCAL_RANK", "0")))
    s, m = rt.scene_rtiow()
    ctx.upload_scene(s, m)
    cam = rt.camera_rtiow(400, 10, DEPTH)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0")))
    for _ in range(max(3, args.warmup)):
        ctx.bench_closest(cam, n, seed=1, precision=prec)
    sampler.start()
    times = []
    for _ in range(args.steps):
        flush.fill_(1); torch.cuda.synchronize()
        times.append(ctx.bench_closest(cam, n, seed=1, precision=prec)["ms"])         # CUDA events on the ctx stream, inside the call
    clocks = sampler.stop()
    e2e = []
    for _ in range(2):
        t0 = time.perf_counter()
        r = ctx.bench_closest(cam, n, seed=1, precision=prec, want_results=True)
        e2e.append((time.perf_counter() - t0) * 1e3)
    hit_frac = float((r["prim"] >= 0).mean())
    check = None
    if not args.no_check:
        from oracle import oracle as O
        nb = 1 << 20
        d = ctx.bench_closest(cam, nb, seed=1, precision=prec, want_results=True, dump_rays=nb)
        want = O.closest_hit(O.scene_rtiow()[0], d["rays"].view(O.RAY))
        check = {"rays_checked": nb, "oracle": "linear hittable_list::hit on the dumped rays", "prim_id_mismatches": int((d["prim"] != want["prim"]).sum())}
    ms = sum(times) / len(times)
    levels = int(np.ceil(np.log2(len(s))))
    f_ray = 48.0 * levels + 29.0 + 22.0          # SURVEY.md 8d without the scatter / sky term: box tests + one sphere test + hit completion
    pk = peaks()
    sm_mhz = clocks.get("sm_mhz") or pk.get("sm_max_mhz", 1965.0)
    fp32_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12
    achieved = n * f_ray / (ms * 1e-3) / 1e12
    line = {"metric": "Mrays/s", "value": n / ms / 1e3, "unit": "Mrays/s", "n_gpus": 1, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
            "config": {"workload": "C2: closest-hit microbenchmark, 2^26 synthetic rays (even: get_ray of a virtual 8192x8192 image, odd: incoherent) "
                                   "vs the final-scene BVH (485 spheres, host SAH, scene staged in shared memory)",
                       "l2": "256 MB buffer written between timed iterations; the kernel reads no input from HBM (rays from counters)",
                       "hit_fraction": hit_frac},
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak, "traffic": None,
                         "kernel": "bench_closest_kernel", "kernel_ms": ms,
                         "note": f"algorithmic {f_ray:.0f} flop/ray (48 x {levels} levels + 29 sphere test + 22 hit completion) x 2^26 rays; "
                                 f"output 8 B/ray = {8 * n / (ms * 1e-3) / 1e9:.0f} GB/s written"},
            "e2e": {"value": n / (sum(e2e) / len(e2e)) / 1e3, "unit": "Mrays/s", "ms_per_step": sum(e2e) / len(e2e),
                    "h2d_bytes_per_step": 104, "d2h_bytes_per_step": 8 * n, "api": "rt_bench_closest with prim/t copied to host arrays"},
            "gpu_launches": args.steps, "clocks": clocks}
    if check:
        line["id_check"] = check
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="f32", choices=["f32", "f64"])
    ap.add_argument("--combine", default="nccl", choices=["nccl", "ipc"], help="--shard samples: how the partial buffers are combined")
    ap.add_argument("--shard", default="auto", choices=["auto", "tiles", "samples", "jobs"],
                    help="N > 1: tile-interleaved shards (every rank writes its tiles' bytes into rank 0's image over NVLink) or "
                         "sample-range shards (fold + reduce of float partial buffers).  auto: tiles when every rank still owns at "
                         "least twice as many tiles as it runs warps (4K frames), else samples (profiles/r2_tile_sharding_1gpu_diag.txt)")
    ap.add_argument("--variant", default="wavefront", choices=["wavefront", "megakernel", "lockstep"])
    ap.add_argument("--config", default="c3", choices=sorted(CONFIGS) + ["c2"], help="BASELINE.json config (default c3 = the headline one)")
    ap.add_argument("--spp", type=int, default=0, help="override spp (a non-default value is NOT the named config)")
    ap.add_argument("--ref-spp", type=int, default=1, help="spp of the CPU sample per step for --impl reference / cpu_baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cpu-as-shipped", action="store_true", help="skip the ~30 s single-thread run of src/main.cpp as shipped")
    ap.add_argument("--no-check", action="store_true", help="skip the untimed f32-vs-f64 image check")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.config == "c2":
        return run_c2(args)
    args.warmup = max(args.warmup, 3)

    import numpy as np
    import torch
    import torch.distributed as dist
    import raytracingproject_b200 as rt
    from raytracingproject_b200 import sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    prec = rt.F32 if args.precision == "f32" else rt.F64
    variant = {"wavefront": rt.WAVEFRONT, "megakernel": rt.MEGAKERNEL, "lockstep": rt.MEGA_LOCKSTEP}[args.variant]
    kernel_name = {"wavefront": "render_pool_kernel", "megakernel": "render_sched_kernel", "lockstep": "render_mega_kernel"}[args.variant]

    workload, W, H, cfg_spp, g, p_diffuse, p_metal, builder = CONFIGS[args.config]
    headline = args.config == "c3"
    args.spp = args.spp or cfg_spp
    builder_id = rt.BVH_LBVH_GPU if builder == "lbvh" else rt.BVH_SAH_HOST
    ctx = rt.Context(local)
    s, m = rt.scene_rtiow(g, p_diffuse, p_metal)
    # the caller's world lives in pinned host memory (the end-to-end leg uploads it every step)
    pins = []
    def pinned(a):
        t = torch.empty(a.nbytes, dtype=torch.uint8, pin_memory=True)
        v = t.numpy().view(a.dtype).reshape(a.shape)
        v[...] = a
        pins.append(t)
        return v
    s, m = pinned(s), pinned(m)
    ctx.upload_scene(s, m, builder=builder_id)
    cam = rt.camera_rtiow(W, args.spp, DEPTH, scale=g / 11.0)
    n_spheres = len(s)
    levels = max(1, int(np.ceil(np.log2(n_spheres))))
    f_ray, b_ray = 48.0 * levels + 96.0, 64.0 * levels + 32.0          # SURVEY.md 8d, per ray at this N
    stream = torch.cuda.ExternalStream(ctx.stream(), device=torch.device("cuda", local))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")      # > 126 MB L2
    begin, count = sharding.sample_shard(args.spp, rank, world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # warm-up render so buffers exist, then set up the combine
    n_tiles = ((W + 15) // 16) * ((H + 7) // 8)                          # 16 x 8 pixel tiles of the wavefront kernel's 128-path pools
    resident_warps = torch.cuda.get_device_properties(local).multi_processor_count * 24
    # auto: tiles cost the kernel ~3 % at 8100 tiles per rank, 11 % at 2025 (a rank's last layers have fewer jobs than warps:
    # profiles/r2_tile_sharding_1gpu_diag.txt) and save the fold + NCCL reduce (~1 % of a C3 frame at N = 2), so they only pay where
    # the sample shards would be short as well: C5's 64 spp over 8 ranks, not C3's 500 (N = 2: 141.6 ms by tiles, 134.9 by samples)
    want_tiles = args.shard == "tiles" or (args.shard == "auto" and n_tiles // max(1, world) >= 2 * resident_warps and args.spp // max(1, world) < 32)
    tiles = world > 1 and want_tiles and args.variant == "wavefront" and args.shard != "jobs"
    # jobs: one job counter for all GPUs (dynamic sharding over NVLink atomics), buffers combined like sample shards
    # (measured: loses to sample shards by 5 % at N = 2 -- profiles/r2_dynamic_sharding.txt -- so only on request)
    jobs = world > 1 and args.shard == "jobs" and args.variant == "wavefront" and args.precision == "f32"
    if jobs:
        sharding.setup_shared_jobs(ctx, rank, world)
        sharding.render_shared(ctx, cam, prec, world)
    elif tiles:
        ctx.render(cam, precision=prec, seed=0, tile_rank=rank, tile_world=world)
        ctx.synchronize()
        sharding.setup_tile_frame(ctx, rank, world)
    else:
        sharding.render_sharded(ctx, cam, rank, world, prec, variant=variant)
    ctx.synchronize()
    acc = sharding.alias_accum_tensor(ctx, layers=1) if world > 1 and not tiles else None      # layer 0: the rank's partial sum after fold_slices
    if world > 1 and not tiles and args.combine == "ipc":
        handles = sharding.exchange_ipc_handles(ctx.ipc_export())
        if rank == 0:
            ctx.ipc_open_peers([h for r, h in enumerate(handles) if r != 0])

    # end-to-end leg: results land in pinned host buffers owned by the caller (allocated once)
    pin_rgb8 = torch.empty(W * H * 3, dtype=torch.uint8, pin_memory=True).numpy() if rank == 0 else None
    pin_ppm = torch.empty(W * H * 12 + 64, dtype=torch.uint8, pin_memory=True).numpy() if rank == 0 else None

    in_flight = [False]                                              # a split read-back begun and not ended (rank 0)

    def step(e2e, split=False):
        """one frame; returns nothing, everything is enqueued on ctx's stream.  split: the frame's bytes leave through
        rt_readback_begin (copies on the library's copy stream); the previous frame's rt_readback_end comes first"""
        if e2e:
            ctx.upload_scene(s, m, builder=builder_id)               # H2D: the world, every step (BVH build included)
        if jobs:
            sharding.render_shared(ctx, cam, prec, world)            # whatever jobs of the frame this rank gets from the shared counter
        elif tiles:
            # every rank renders all samples of its tiles; ranks != 0 resolve them into rank 0's image over NVLink and signal;
            # rank 0's readback below waits for the signals on its stream.  No host synchronisation inside the step.
            sharding.render_tiles(ctx, cam, rank, world, prec)
        else:
            sharding.render_sharded(ctx, cam, rank, world, prec, variant=variant)
        if world > 1 and not tiles:
            ctx.fold_slices()                                        # one [H][W] layer per rank crosses NVLink, not `slices` of them
            if args.combine == "nccl":
                with torch.cuda.stream(stream):
                    sharding.reduce_partials(acc, dst=0)
            else:
                stream.synchronize(); dist.barrier()                 # peers' partials complete before rank 0 reads them
        if rank == 0:
            if e2e and split:
                n = ctx.readback_end() if in_flight[0] else None     # frame k - 1: its copies ran under this frame's render
                ctx.readback_begin(args.spp, pin_rgb8, pin_ppm)      # waits for this frame, starts its copies, returns
                in_flight[0] = True
                return {"ppm_len": n}
            if e2e:
                return {"ppm_len": ctx.readback_into(args.spp, pin_rgb8, pin_ppm)}
            return ctx.readback(args.spp, want_sum=False, want_rgb8=False)   # resolve on device, nothing copied

    rank_kernel_ms = []

    def timed(e2e, steps, warmup):
        for _ in range(warmup):
            step(e2e)
        times, rays, kernel_ms = [], 0, []
        rank_kernel_ms.clear()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        for _ in range(steps):
            flush.fill_(1)                                           # evict L2 between timed iterations
            barrier()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            with torch.cuda.stream(stream):
                ev0.record(stream)
            out = step(e2e)
            with torch.cuda.stream(stream):
                ev1.record(stream)
            barrier()
            wall = (time.perf_counter() - t0) * 1e3
            dev = ev0.elapsed_time(ev1)
            # device-resident timing uses CUDA events on the launching stream; the end-to-end number
            # includes host work (BVH build, memcpy completion), so it uses the wall clock between barriers
            t = torch.tensor([wall if e2e else dev], dtype=torch.float64, device="cuda")
            st = ctx.stats()
            r = torch.tensor([st["rays"]], dtype=torch.int64, device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dist.all_reduce(r, op=dist.ReduceOp.SUM)
            times.append(t.item()); rays = int(r.item()); kernel_ms.append(st["render_ms"])
            if world > 1:                                            # every rank's render kernel time: the skew the combine has to wait out
                k = torch.zeros(world, dtype=torch.float64, device="cuda"); k[rank] = st["render_ms"]
                dist.all_reduce(k)
                rank_kernel_ms.append([round(x, 3) for x in k.tolist()])
        clocks = sampler.stop() if rank == 0 else None
        return times, rays, kernel_ms, clocks, out

    def timed_pipelined(steps, warmup):
        """the end-to-end frame as a caller rendering frame after frame runs it: K frames back to back between two barriers, the
        device-to-host copies of frame k (rt_readback_begin / _end) under the upload and render of frame k + 1.  Every frame's
        upload, render, combine, resolve, P3 encode and both copies are inside the timed region, the last frame's copies included."""
        def drain():
            n = None
            if rank == 0 and in_flight[0]:
                n = ctx.readback_end(); in_flight[0] = False
            return n
        for _ in range(warmup):
            step(True, split=True)
        drain()
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            with torch.cuda.stream(stream):
                flush.fill_(1)                                       # evict L2 between frames, on the frames' own stream (~0.05 ms)
            step(True, split=True)
        n = drain()
        barrier()
        t = torch.tensor([(time.perf_counter() - t0) * 1e3 / steps], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item(), n

    times, rays, kernel_ms, clocks, _ = timed(False, args.steps, args.warmup)
    rank_kernels = [list(x) for x in rank_kernel_ms]
    e2e_times, _, _, _, e2e_out = timed(True, max(2, min(args.steps, 5)), 1)
    launches = ctx.stats()["kernel_launches"]
    # frames in a row with the read-back split in two (not with tile or job sharding: there the peers write rank 0's image, and
    # the hand-over of frame k + 1 may not start before rank 0 has frame k on the host -- one barrier per frame, nothing to overlap)
    pipelined_ms = None
    if not tiles and not jobs and (world == 1 or args.combine == "nccl") and not os.environ.get("RT_BENCH_NO_PIPELINE"):
        pipelined_ms, n_text = timed_pipelined(max(3, min(2 * args.steps, 10)), 1)   # a frame's kernel time wanders by +- 1 % from step to step: ten frames
        if rank == 0 and e2e_out and n_text != e2e_out["ppm_len"]:
            raise SystemExit(f"bench: the split read-back delivered {n_text} bytes of P3 text, rt_readback {e2e_out['ppm_len']}")

    # untimed, N > 1: the COMBINED frame of the last end-to-end step (folded, reduced over NVLink, resolved on rank 0)
    # against the same frame rendered by rank 0 alone.  The float partial sums re-associate, nothing else may differ.
    multi_check = None
    if world > 1 and not args.no_check:
        combined = pin_rgb8.reshape(H, W, 3).copy() if rank == 0 else None
        if args.combine == "ipc" and rank == 0 and not tiles:
            ctx.ipc_close_peers()
        barrier()
        if rank == 0:
            ctx.render(cam, precision=prec, seed=0, variant=variant)
            solo = ctx.readback(args.spp, want_sum=False)["rgb8"]
            d = np.abs(combined.astype(int) - solo.astype(int)).max(-1)
            mse = float(np.mean((combined.astype(float) - solo.astype(float)) ** 2))
            multi_check = {"combined_vs_single_gpu_max_byte_diff": int(d.max()),
                           "combined_vs_single_gpu_psnr_db": 99.0 if mse == 0 else float(10 * np.log10(255.0 ** 2 / mse)),
                           "frac_pixels_identical": float((d == 0).mean()), "ranks": world,
                           "combine": "tiles over NVLink" if tiles else args.combine, "sharding": "jobs" if jobs else ("tiles" if tiles else "samples")}
        barrier()

    # untimed: the fast mode against parity mode on this very frame (north_star image tolerance)
    check = None
    if rank == 0 and world == 1 and not args.no_check and args.precision == "f32" and headline:
        ctx.render(cam, precision=rt.F32, seed=0, variant=variant); a = ctx.readback(args.spp, want_sum=False)["rgb8"]
        ctx.render(cam, precision=rt.F64, seed=0, variant=rt.MEGAKERNEL); b = ctx.readback(args.spp, want_sum=False)["rgb8"]
        d = np.abs(a.astype(int) - b.astype(int)).max(-1)
        mse = float(np.mean((a.astype(float) - b.astype(float)) ** 2))
        check = {"f32_vs_f64_psnr_db": 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse),
                 "frac_pixels_within_1_of_255": float((d <= 1).mean()), "frac_pixels_identical": float((d == 0).mean())}

    if rank == 0:
        pk = peaks()
        ms = sum(times) / len(times)
        samples = W * H * args.spp
        value = samples / ms / 1e3
        kms = sum(kernel_ms) / len(kernel_ms)                          # rank 0's render kernel, CUDA events
        rank_rays = rays / world
        sm_mhz = (clocks or {}).get("sm_mhz") or pk.get("sm_max_mhz", 1965.0)
        fp32_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12               # TFLOP/s at the clock seen under load
        achieved = rank_rays * f_ray / (kms * 1e-3) / 1e12
        st_scene = ctx.stats()
        e2e_ms = sum(e2e_times) / len(e2e_times)
        if os.environ.get("RT_BENCH_DEBUG"):
            print("device-resident step ms:", [round(t, 3) for t in times], " e2e step ms:", [round(t, 3) for t in e2e_times], file=sys.stderr)
        line = {
            "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": args.precision, "data": "synthetic",
            "config": {"workload": workload if args.spp == cfg_spp else f"NOT the named config: {args.config} at spp={args.spp}",
                       "variant": args.variant, "n_spheres": n_spheres,
                       "bvh": ("host SAH" if builder == "sah" else "GPU Morton LBVH") + f", depth {st_scene['bvh_depth']}, " +
                              ("whole scene staged in shared memory" if st_scene["scene_in_smem"] == 1 else
                               f"nodes and spheres in global memory (L2 / HBM), first {st_scene.get('top_nodes_in_smem', 0)} nodes of the tree also in shared memory"
                               if st_scene["scene_in_smem"] == 2 else "nodes and spheres in global memory (L2 / HBM)"),
                       "seed": 0,
                       "parallelism": (f"tile-interleaved sharding x{world} (pairs of 16x8 tiles), every rank's resolve kernel writes its tiles' "
                                       f"bytes into rank 0's image over NVLink" if tiles else
                                       f"dynamic sharding x{world}: (tile, sample layer) jobs of the frame from one counter in rank 0's memory "
                                       f"(system-scope atomics over NVLink), combine={args.combine}" if jobs else
                                       f"sample-range sharding x{world}, combine={args.combine if world > 1 else 'none'}"),
                       "l2": f"256 MB buffer written between timed iterations (L2 flushed); scene inputs are {(s.nbytes + m.nbytes) / 1e3:.0f} KB"},
            "mrays_per_s": rays / ms / 1e3, "rays_per_sample": rays / samples,
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak,
                         # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of this kernel on this workload, from the
                         # ncu --set full capture profiles/r2_wavefront_c3_500spp_ncu_summary.txt (6.3 MB + 227.8 MB: the 8 sample
                         # layers of the frame written once; fast mode adds samples with RED.ADD, so nothing is read back); the
                         # bounce loop itself reads nothing from HBM (scene staged on chip), this is the accumulator traffic
                         "traffic": 234.1e6 if (headline and args.variant == "wavefront" and world == 1 and args.spp == SPP and args.precision == "f32") else None,
                         "kernel": kernel_name, "kernel_ms": kms,
                         "note": f"algorithmic {f_ray:.0f} flop/ray x {rank_rays:.3e} rays per launch (SURVEY 8d); peak = 148 SMs x 128 "
                                 f"FP32 lanes x 2 x {sm_mhz:.0f} MHz observed under load (not in MEASURED_PEAKS.json); L2-side figure "
                                 f"{rank_rays * b_ray / (kms * 1e-3) / 1e9:.0f} GB/s algorithmic ({b_ray:.0f} B/ray)"},
            # frames in a row (a caller's render loop): the copies of frame k overlap the render of frame k + 1; `one_frame_alone_ms`
            # is the same frame with nothing before or after it (rt_readback, each step between two barriers)
            "e2e": ({"value": samples / pipelined_ms / 1e3, "unit": "Msamples/s", "ms_per_step": pipelined_ms, "one_frame_alone_ms": e2e_ms,
                     "h2d_bytes_per_step": int(s.nbytes + m.nbytes + 104),
                     "d2h_bytes_per_step": int(W * H * 3 + (e2e_out["ppm_len"] if e2e_out else 0)),
                     "api": "rt_upload_scene + rt_render + rt_readback_begin / rt_readback_end (rgb8 + P3 text) per step, frames back to back "
                            "between two barriers: the device-to-host copies of a frame run under the next frame's render, the last frame's "
                            "inside the timed region; scene from pinned host arrays, results into pinned host buffers"}
                    if pipelined_ms else
                    {"value": samples / e2e_ms / 1e3, "unit": "Msamples/s", "ms_per_step": e2e_ms,
                     "h2d_bytes_per_step": int(s.nbytes + m.nbytes + 104),
                     "d2h_bytes_per_step": int(W * H * 3 + (e2e_out["ppm_len"] if e2e_out else 0)),
                     "api": "rt_upload_scene + rt_render + rt_readback(rgb8 + P3 text) per step; scene from pinned host arrays, results into pinned host buffers"}),
            "gpu_launches": int(launches) * args.steps, "clocks": clocks,
            # where a step's time goes (device-resident leg): rank 0's render kernel (CUDA events inside the library) and what
            # follows it until the frame is resolved on rank 0 -- tiles: wait for the peers' bytes + resolve of rank 0's tiles;
            # samples: fold + reduce + resolve; N = 1: resolve alone.  (NVTX ranges of the same names: rt_api.cu)
            "phases_ms": {"render_kernel_rank0": kms, "after_render_until_frame_resolved": ms - kms,
                          "render_kernel_every_rank_last_step": rank_kernels[-1] if rank_kernels else None,
                          "render_kernel_slowest_rank_mean": (sum(max(x) for x in rank_kernels) / len(rank_kernels)) if rank_kernels else None},
        }
        if check:
            line["image_check"] = check
        if multi_check:
            line["image_check"] = multi_check
        if not args.no_cpu_baseline and world == 1 and headline:
            keys = ("value", "unit", "cores", "kind", "sample", "mrays_per_s")
            cb = cpu_reference_run(max(1, args.ref_spp), os.cpu_count() or 1)
            line["cpu_baseline"] = {k: cb[k] for k in keys}
            # how much of the ratio is the BVH, and the reference as it ships
            cbb = cpu_reference_run(4 * max(1, args.ref_spp), os.cpu_count() or 1, use_bvh=True)
            line["cpu_baseline_bvh"] = {k: cbb[k] for k in keys}
            line["cpu_baseline_bvh"]["speedup_over_linear_list"] = cbb["value"] / cb["value"]
            if not args.no_cpu_as_shipped:
                ca = cpu_as_shipped_run()
                if ca:
                    line["cpu_baseline_as_shipped"] = {k: ca[k] for k in keys}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
