#!/usr/bin/env python
"""bench.py — the headline measurement: Msamples/s (and Mrays/s) of the pt+MIS path at 1080p on the
procedural 10 M-triangle displaced mesh + 10 k spheres scene (BASELINE.json configs[3], "C4"),
depth 16, on N B200s of one box.

  python bench.py [--gpus N] [--steps K] [--warmup W]          the CUDA product (libpupil_b200.so)
  python bench.py --impl reference ...                         the reference integrator's own header
                                                               code on the host cores (oracle/_ref)

One STEP = `--spp-per-step` (default 32) consecutive frames (= samples per pixel) of the whole
1920x1080 film through pb_render; the default 8 timed steps are the config's 256 spp.  With N GPUs
rank r renders the frames  s = r (mod N)  (same seeds as 1 GPU) and every step ends with
an NCCL reduce of the fp32 sum buffer to rank 0 inside the timed region ("weak": per-GPU work fixed).

Printed JSON (one line, rank 0): value = whole-job Msamples/s with the scene/BVH resident in HBM and
device output buffers; e2e = the same through the C ABI with HOST (pinned) accumulation buffers —
host->device and device->host copies inside the timed call; roofline = the closest-hit traversal
kernel (k_extend) against the measured HBM copy bandwidth, algorithmic bytes per ray as defined in
SURVEY 8(d) / DESIGN.md; cpu_baseline = the test oracle (reference headers on the host cores) on a
bounded sample of the same workload.  oracle/ is only ever executed here for those two legs.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WIDTH, HEIGHT, MAX_DEPTH, N_SPHERES = 1920, 1080, 16, 10000
METRIC = "Msamples/s at 1080p pt+MIS"
WORKLOAD = "C4: procedural 10M-triangle displaced mesh + 10k spheres, 7-material patchwork, 1920x1080, depth 16, pt+MIS"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """SM clock, power and throttle reasons sampled every 100 ms DURING the timed region, OUT OF PROCESS: a streaming
    `nvidia-smi -lms 100` child started well before the warm-up writes timestamped rows into a pipe that nobody reads
    until the timed region is over (10 rows / s against a 64 KB pipe buffer), so the sampler never competes with the
    enqueuing thread for the GIL (an in-process NVML poller delayed the first enqueue of a timed region by up to 20 ms:
    every hand-over of the GIL can cost a 5 ms switch interval).  Rows are selected by their wall-clock time stamp
    between begin() and end().  In-process NVML (nvidia-ml-py) is only the fallback when nvidia-smi is missing."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int, uuid=None):
        self.rows, self.proc, self.h, self.nv = [], None, None, None
        self.t0 = self.t1 = None
        self.alive = True
        self.wall = False      # rows carry wall-clock stamps (nvidia-smi) instead of perf_counter stamps (NVML thread)
        try:
            sel = ("GPU-" + str(uuid)) if uuid else str(index)
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", sel], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.wall = True
            return
        except Exception:
            self.proc = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode()) if uuid else pynvml.nvmlDeviceGetHandleByIndex(index)
            self._sample()                      # first queries are the slow ones
            self.rows.clear()
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
        except Exception:
            self.h = None

    def _sample(self):
        nv, h = self.nv, self.h
        sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
        flags = [("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4)]
        self.rows.append((time.perf_counter(), float(sm), float(mx), pw, [n for n, b in flags if r & b]))

    def _poll(self):
        while self.alive:
            try:
                self._sample()
            except Exception:
                pass
            time.sleep(0.1)

    def _drain(self):
        """Parse everything the nvidia-smi child has written (called once, after it was terminated)."""
        import datetime
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            out = self.proc.stdout.read()
        except Exception:
            out = ""
        for line in out.splitlines():
            r = [x.strip() for x in line.split(",")]
            try:
                ts = datetime.datetime.strptime(r[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                self.rows.append((ts, float(r[1]), float(r[2]), float(r[3]),
                                  [n for n, v in zip(names, r[4:8]) if v.lower().startswith("active")]))
            except Exception:
                continue

    def _now(self):
        return time.time() if self.wall else time.perf_counter()

    def begin(self):
        self.t0 = self._now()

    def end(self):
        self.t1 = self._now()

    def stop(self) -> dict:
        if self.h is None and not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml / nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.alive = False
        if self.proc:
            self.proc.terminate()
            self._drain()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        return self._summary()

    def _summary(self) -> dict:
        t0 = self.t0 if self.t0 is not None else -1e30
        t1 = self.t1 if self.t1 is not None else 1e30
        rows = [r for r in self.rows if t0 <= r[0] <= t1 + 0.11] or self.rows
        sm, mx, pw = [r[1] for r in rows], [r[2] for r in rows], [r[3] for r in rows]
        reasons = sorted({n for r in rows for n in r[4]})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_min_mhz": min(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": reasons,
                "source": "nvidia-smi -lms 100 (child process)" if self.proc else "nvml (in-process thread)"}


def build_scene(ctx, mesh_tag: str, mesh=None):
    from pupiloptixlab_b200 import scenes
    n = scenes.MESH_N[mesh_tag]
    t0 = time.time()
    mesh = mesh if mesh is not None else scenes.displaced_mesh(n)
    t1 = time.time()
    info = scenes.bench_scene(ctx, n, N_SPHERES, WIDTH, HEIGHT, patchwork=True, mesh=mesh)
    info.update(mesh_s=t1 - t0, commit_s=time.time() - t1)
    return info, mesh


# --------------------------------------------------------------------------------------------------
def cpu_leg(mesh_tag, mesh, seconds_target, steps, warmup, want_line, keep=None):
    """The reference integrator's header code on the host cores (oracle/_ref), bounded sample.  `keep` (a dict) receives
    the oracle context with the scene built, for the roofline's visit-count pass."""
    from oracle import pyoracle
    api = pyoracle.load()
    o = pyoracle.OracleContext(api)
    kind = "port"
    if o.has_reference():
        o.set_backend(1)
        kind = "reference"
    # torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU arm uses every host core it may run on
    avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1 and int(api._max_threads()) < avail:
        api._set_threads(avail)
    cores = int(api._max_threads())
    info, mesh = build_scene(o, mesh_tag, mesh)
    if keep is not None:
        keep["oracle"] = o
    # probe, then the largest film (same camera, same scene) whose 1-spp frame fits the time budget
    t0 = time.time(); o.render(240, 135, MAX_DEPTH, 0, 1, want_aov=False); probe = 240 * 135 / (time.time() - t0)
    film = (240, 135)
    for w, h in ((1920, 1080), (960, 540), (480, 270)):
        if w * h / probe <= seconds_target:
            film = (w, h)
            break
    w, h = film
    rates, rays = [], 0
    for i in range(warmup + steps):
        t0 = time.time()
        o.render(w, h, MAX_DEPTH, i, 1, want_aov=False)
        dt = time.time() - t0
        if i >= warmup:
            st = o.stats()
            rates.append(dt); rays += st["rays_closest"] + st["rays_any"]
    total = sum(rates)
    msps = w * h * steps / total / 1e6
    sample = f"{steps} x 1 spp of the same scene/camera on a {w}x{h} film, depth {MAX_DEPTH} ({w * h * steps} samples, {total:.1f} s)"
    base = {"value": msps, "unit": "Msamples/s", "cores": cores, "kind": kind, "sample": sample,
            "mrays_per_s": rays / total / 1e6}
    return base, total / steps * 1e3, info


def make_config(args, info, N):
    """The `config` object of the JSON line: identical for the product arm and the reference arm (the driver compares them)."""
    S = args.spp_per_step
    return {"workload": WORKLOAD, "mesh": args.mesh, "n_tris": info["n_tris"], "n_spheres": info["n_spheres"],
            "spp_per_step": S, "frames_per_gpu_per_step": S,
            "sharding": f"sample-interleaved x{N} + one ncclReduce(sum, fp32, root 0) per step issued by pb_render (PB_RENDER_REDUCE)" if N > 1 else "none",
            "l2": "inputs larger than L2 (BVH + primitives ~0.7 GB, path state > 1 GB per wave)"}


def reference_main(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    base, ms_per_step, info = cpu_leg(args.mesh, None, 8.0, args.steps, args.warmup, True)
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": "Msamples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": make_config(args, info, max(1, int(os.environ.get("WORLD_SIZE", str(args.gpus))))),   # the product arm's config: same workload;
            # each reference step is a bounded sample of it (cpu_baseline.sample says which)
            "mrays_per_s": base["mrays_per_s"], "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)
    return 0


# --------------------------------------------------------------------------------------------------
def product_main(args):
    import torch
    import torch.distributed as dist
    from pupiloptixlab_b200 import capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product has no CPU path (use --impl reference for the host baseline)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries the one JSON line only
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    N = world
    clocks = None
    if rank == 0:   # started before the scene is built: the child is sampling long before the timed region begins
        try:
            uuid = torch.cuda.get_device_properties(local).uuid
        except Exception:
            uuid = None
        clocks = ClockSampler(local, uuid)
    api = capi.load()
    g = capi.Context(api, local)
    if N > 1:
        # the library's own communicator: rank 0's NCCL unique id travels over the process group, every rank joins;
        # from here on the per-step reduce is the ncclReduce pb_render enqueues itself (PB_RENDER_REDUCE)
        uid = torch.zeros(128, dtype=torch.uint8, device=torch.device("cuda", local))
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(g.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, src=0)
        g.comm_init_rank(bytes(uid.cpu().numpy().tobytes()), N, rank)
    info, mesh = build_scene(g, args.mesh)
    st0 = g.stats()
    if rank == 0:
        log(f"[bench] scene: {info['n_tris']} tris + {info['n_spheres']} spheres, nodes={st0['n_nodes']}, "
            f"bvh levels={st0['bvh_depth']}, build {st0['build_ms']:.0f} ms (host prep + device build), mesh gen {info['mesh_s']:.1f}s")

    n_px = WIDTH * HEIGHT
    S = args.spp_per_step
    dev = torch.device("cuda", local)
    d_sum = torch.zeros((HEIGHT, WIDTH, 4), dtype=torch.float32, device=dev)
    d_total = torch.zeros_like(d_sum)
    stream = torch.cuda.Stream(dev)      # everything (kernels, copies, NCCL) is ordered on this stream
    torch.cuda.set_stream(stream)
    sp = stream.cuda_stream

    def step_params(i, flags):
        return capi.RenderParams(WIDTH, HEIGHT, MAX_DEPTH, i * S * N + rank, S, N, flags, 0)

    REDUCE = capi.RENDER_REDUCE if N > 1 else 0

    def dev_step(i, timing=True, asynchronous=False):
        # N > 1: pb_render sums this rank's frames, enqueues ONE ncclReduce(sum, fp32, root 0) behind its last wave on
        # the same stream and, on rank 0, leaves the reduced film of the step in d_sum
        p = step_params(i, capi.RENDER_CLEAR | REDUCE | (capi.RENDER_STAGE_TIMING if timing else 0) | (capi.RENDER_ASYNC if asynchronous else 0))
        g.render_dev(p, d_sum.data_ptr(), 0, 0, sp)
        if rank == 0:
            d_total.add_(d_sum)
        return None if asynchronous else g.stats()

    def barrier():
        if N > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for i in range(args.warmup):
        dev_step(i)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rays_c = rays_a = launches = 0
    ext_ms = con_ms = shd_ms = 0.0
    ext_l = con_l = 0
    per_step_rays_c = []
    step_ms = []
    if clocks:
        clocks.begin()
    e0.record(stream)
    # the K steps are enqueued back to back on the stream (PB_RENDER_ASYNC; the film of step i is consumed by the
    # reduce / add enqueued right behind it) and synchronised once: the timed region holds device work only,
    # a host hiccup between two steps cannot idle the GPU
    marks, host_ms = [], []
    for i in range(args.steps):
        t_i = time.perf_counter()
        dev_step(args.warmup + i, asynchronous=True)
        host_ms.append(round((time.perf_counter() - t_i) * 1e3, 2))
        ev = torch.cuda.Event(enable_timing=True); ev.record(stream); marks.append(ev)
    e1.record(stream)
    barrier()
    if clocks:
        clocks.end()
    ms = e0.elapsed_time(e1)
    gpu_step_ms = [round((e0 if i == 0 else marks[i - 1]).elapsed_time(marks[i]), 2) for i in range(len(marks))]   # device timeline, step by step
    st = g.stats()            # totals of the K asynchronous renders
    rays_c, rays_a, launches = st["rays_closest"], st["rays_any"], st["kernel_launches"] + (args.steps if rank == 0 else 0)
    ext_ms, con_ms, shd_ms, ext_l, con_l = st["extend_ms"], st["connect_ms"], st["shade_ms"], st["extend_launches"], st["connect_launches"]
    step_ms = [round(st["last_render_ms"] / args.steps, 2)]
    clk = clocks.stop() if clocks else None
    t = torch.tensor([ms, float(rays_c), float(rays_a), float(launches)], dtype=torch.float64, device=dev)
    if N > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms = float(tmax[0]); rays_c_all, rays_a_all, launches_all = float(tsum[1]), float(tsum[2]), float(tsum[3])
    else:
        rays_c_all, rays_a_all, launches_all = float(rays_c), float(rays_a), float(launches)
    samples = float(n_px) * S * args.steps * N
    value = samples / (ms * 1e-3) / 1e6
    mrays = (rays_c_all + rays_a_all) / (ms * 1e-3) / 1e6

    # ---- the same K steps on the shipped launch path (one CUDA graph per wave; the region above launches kernel by
    # kernel because it records events around every traversal launch for the roofline)
    barrier()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dev_step(args.warmup, timing=False)          # captures the graph outside the timed region
    barrier()
    g0.record(stream)
    for i in range(args.steps):
        dev_step(args.warmup + i, timing=False, asynchronous=True)
    g1.record(stream)
    barrier()
    g.stats()
    graph_ms = g0.elapsed_time(g1)
    if N > 1:
        tg = torch.tensor([graph_ms], dtype=torch.float64, device=dev); dist.all_reduce(tg, op=dist.ReduceOp.MAX); graph_ms = float(tg[0])
    graph_value = samples / (graph_ms * 1e-3) / 1e6

    # ---- e2e: the same steps through the C ABI with HOST buffers (pinned), copies inside the call
    h_sum = torch.zeros((HEIGHT, WIDTH, 4), dtype=torch.float32).pin_memory()
    h_ptr = h_sum.data_ptr()

    def e2e_step(i, asynchronous):
        # accumulate into the caller's host buffer: upload of the running sums + download of the new ones (with N > 1 on
        # rank 0 only: the other ranks' frames arrive through the library's ncclReduce, their host buffers are not touched)
        p = step_params(i, capi.RENDER_HOST_BUFFERS | REDUCE | (capi.RENDER_ASYNC if asynchronous else 0))
        g.render_dev(p, h_ptr, 0, 0, sp)
        if not asynchronous:
            torch.cuda.synchronize(dev)

    def e2e_run(asynchronous):
        e2e_step(0, False)
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        f0.record(stream)
        each = []
        for i in range(e2e_steps):
            t_i = time.perf_counter()
            e2e_step(args.warmup + i, asynchronous)
            each.append(round((time.perf_counter() - t_i) * 1e3, 2))
        if asynchronous:
            g.wait()
        f1.record(stream)
        barrier()
        t_ms = max(f0.elapsed_time(f1), (time.perf_counter() - w0) * 1e3 if N == 1 else 0.0)
        if N > 1:
            tt = torch.tensor([t_ms], dtype=torch.float64, device=dev); dist.all_reduce(tt, op=dist.ReduceOp.MAX); t_ms = float(tt[0])
        return t_ms, each

    # e2e: the steps queued through the asynchronous C ABI (PB_RENDER_ASYNC | PB_RENDER_HOST_BUFFERS: every step uploads
    # the running sums from pinned host memory, renders, downloads the new sums; one pb_wait at the end).  Also
    # reported: the same steps as synchronous calls, one at a time ("sync": exposed to host scheduling noise between
    # calls - on shared boxes single steps took up to +50 ms).
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    e2e_ms, e2e_async_host = e2e_run(True)
    sync_ms, e2e_each = e2e_run(False)
    e2e_value = float(n_px) * S * e2e_steps * N / (e2e_ms * 1e-3) / 1e6
    sync_value = float(n_px) * S * e2e_steps * N / (sync_ms * 1e-3) / 1e6
    h2d = n_px * 16 + C.sizeof(capi.RenderParams)
    d2h = n_px * 16

    # ---- N > 1: does the sharded, NCCL-reduced film equal the film one GPU renders for the same frames?  One more step
    # on all ranks (library reduce), then rank 0 alone renders the N * S frames of that step itself and compares.
    parity = None
    if N > 1:
        i_chk = args.warmup
        dev_step(i_chk, timing=False)
        barrier()
        if rank == 0:
            d_one = torch.zeros_like(d_sum)
            g.render_dev(capi.RenderParams(WIDTH, HEIGHT, MAX_DEPTH, i_chk * S * N, S * N, 1, capi.RENDER_CLEAR, 0), d_one.data_ptr(), 0, 0, sp)
            torch.cuda.synchronize(dev)
            a, b = d_sum[..., :3].double(), d_one[..., :3].double()
            # the reference has no NaN / Inf guard: a non-finite pixel must be non-finite on both sides
            fin = torch.isfinite(a).all(-1) & torch.isfinite(b).all(-1)
            nonfinite_equal = bool(torch.equal(torch.isfinite(a), torch.isfinite(b)))
            a = torch.where(fin[..., None], a, torch.zeros_like(a)); b = torch.where(fin[..., None], b, torch.zeros_like(b))
            rel = ((a - b).abs() / (b.abs() + 1e-3)).amax(-1)
            parity = {"frames": S * N, "step": i_chk, "against": "the same frames rendered by rank 0 alone (sample_stride 1)",
                      "counts_equal": bool(torch.equal(d_sum[..., 3], d_one[..., 3])) and float(d_one[0, 0, 3]) == S * N,
                      "non_finite_pixels": int((~fin).sum()), "non_finite_equal": nonfinite_equal,
                      "max_rel": float(rel.max()), "pixels_within_1e-5": float((rel <= 1e-5).double().mean()),
                      "film_mean_ratio": float(a.mean() / b.mean()),
                      "note": "frames are bit-identical on every GPU; the only difference is the order of the float additions (per-rank sums, then the reduce)"}
            del d_one
        barrier()

    # ---- CPU baseline (rank 0, N = 1 only): the oracle on the host cores, bounded sample
    cpu = None
    kept = {}
    if rank == 0 and N == 1 and not args.no_cpu:
        try:
            cpu, _, _ = cpu_leg(args.mesh, mesh, 10.0, 2, 1, False, keep=kept)
        except Exception as e:  # the baseline is a report, never a reason to lose the GPU line
            cpu = {"value": None, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {e}"}

    # ---- roofline of the dominant kernel (k_extend).  Algorithmic bytes per ray = 48 + 80 n_nodes + 48 n_tri
    # + 32 n_sph with the visit counts of the ORACLE's ordered, t-culled traversal of the BVH this context committed
    # (pb_export_bvh -> oracle/orc_cwbvh.h), taken over the rays of a strided sample of the first timed step's first frame
    # (every 8th pixel in x and y, the oracle's own path tracer generates the rays: same seeds, same paths).  The kernel's
    # own visit counters (PB_RENDER_COUNTED: octant order, batched primitive tests, stale stack entries) are reported next
    # to it and are only the fallback when the oracle leg is switched off (--no-cpu).
    roof = None
    extra = {}
    if rank == 0:
        p = step_params(args.warmup, capi.RENDER_CLEAR | capi.RENDER_COUNTED)
        g.render_dev(p, d_sum.data_ptr(), 0, 0, sp)
        sc = g.stats()
        rc, ra = max(1, sc["rays_closest"]), max(1, sc["rays_any"])
        b_closest = 48.0 + (80.0 * sc["node_visits"] + 48.0 * sc["prim_tests_tri"] + 32.0 * sc["prim_tests_sphere"]) / rc
        b_any = 33.0 + (80.0 * sc["node_visits_any"] + 48.0 * sc["prim_tests_tri_any"] + 32.0 * sc["prim_tests_sphere_any"]) / ra
        peak, peak_src = peaks()
        kernel_counters = {"bytes_per_ray": b_closest, "nodes_per_ray": sc["node_visits"] / rc, "tris_per_ray": sc["prim_tests_tri"] / rc,
                           "spheres_per_ray": sc["prim_tests_sphere"] / rc, "bytes_per_ray_any": b_any,
                           "source": "PB_RENDER_COUNTED pass over the first timed step's frames (the kernel's own counters)"}
        bytes_from = kernel_counters["source"] + " — oracle leg off"
        b_min = None
        oc = None
        o = kept.get("oracle")
        if o is not None:
            try:
                t_o = time.time()
                o.set_backend(0)
                bvh = o.import_cwbvh(g)
                step = 8
                o.render_window(WIDTH, HEIGHT, MAX_DEPTH, args.warmup * S * N, 1, 0, 0, (WIDTH + step - 1) // step, (HEIGHT + step - 1) // step,
                                step=step, counted=True)
                so = o.stats()
                orc, ora = max(1, so["rays_closest"]), max(1, so["rays_any"])
                oc = {"nodes_per_ray": so["node_visits"] / orc, "tris_per_ray": so["prim_tests_tri"] / orc, "spheres_per_ray": so["prim_tests_sphere"] / orc,
                      "rays": so["rays_closest"], "nodes_per_ray_any": so["node_visits_any"] / ora, "rays_any": so["rays_any"]}
                b_closest = 48.0 + (80.0 * so["node_visits"] + 48.0 * so["prim_tests_tri"] + 32.0 * so["prim_tests_sphere"]) / orc
                b_any = 33.0 + (80.0 * so["node_visits_any"] + 48.0 * so["prim_tests_tri_any"] + 32.0 * so["prim_tests_sphere_any"]) / ora
                bytes_from = (f"oracle's ordered, t-culled traversal of the exported BVH (oracle/orc_cwbvh.h) over the {so['rays_closest']} closest-hit rays of "
                              f"every {step}th pixel in x and y of frame {args.warmup * S * N}, depth {MAX_DEPTH} ({time.time() - t_o:.1f} s)")
                import math
                levels = max(1, math.ceil(math.log(max(2, bvh["n_leaf_slots"]), 8)))
                per_leaf = bvh["n_prims"] / max(1, bvh["n_leaf_slots"])
                b_min = {"bytes_per_ray": 48.0 + 80.0 * levels + 48.0 * per_leaf, "levels": levels, "prims_per_leaf_slot": per_leaf,
                         "n_leaf_slots": bvh["n_leaf_slots"], "formula": "48 + 80 ceil(log8(leaf slots)) + 48 prims per leaf slot (SURVEY 8d)"}
            except Exception as e:
                bytes_from = kernel_counters["source"] + f" — oracle leg failed: {e}"
        ach = rays_c * b_closest / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else 0.0
        ach_any = rays_a * b_any / (con_ms * 1e-3) / 1e9 if con_ms > 0 else 0.0
        traffic = None
        prof = {}
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                prof = json.load(open(tp))
                traffic = prof.get("k_extend_dram_bytes_per_launch")
            except Exception:
                traffic = None
        roof = {"bound": "hbm", "kernel": "k_extend (closest-hit traversal of the 8-wide BVH)", "achieved": ach, "peak": peak,
                "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                "bytes_per_ray": b_closest, "bytes_per_launch": rays_c * b_closest / max(1, ext_l),
                "avg_launch_ms": ext_ms / max(1, ext_l), "launches": ext_l,
                "nodes_per_ray": (oc or kernel_counters)["nodes_per_ray"], "tris_per_ray": (oc or kernel_counters)["tris_per_ray"],
                "spheres_per_ray": (oc or kernel_counters)["spheres_per_ray"],
                "share_of_step": ext_ms / ms if N == 1 else None,
                "bytes_from": bytes_from, "oracle_counts": oc, "kernel_counters": kernel_counters}
        avg_s = ext_ms * 1e-3 / max(1, ext_l)
        if b_min:      # the implementation-independent floor of SURVEY 8(d)
            roof["b_min"] = b_min
            roof["b_min_frac"] = rays_c * b_min["bytes_per_ray"] / (ext_ms * 1e-3) / 1e9 / peak if ext_ms > 0 else None
        if traffic and avg_s > 0:   # what really crosses the HBM pins (ncu dram__bytes of profiles/, per launch)
            roof["dram_frac"] = traffic / avg_s / 1e9 / peak
            roof["traffic_source"] = prof.get("source")
        wi = prof.get("k_extend_warp_inst_per_ray")
        if wi and ext_ms > 0 and clk and clk.get("sm_mhz"):
            slots = 148 * 4 * clk["sm_mhz"] * 1e6     # warp instructions the SM sub-partitions can issue per second
            roof["issue"] = {"warp_inst_per_ray": wi, "issue_slots_per_s": slots, "frac": wi * rays_c / (ext_ms * 1e-3) / slots,
                             "note": "the binding ceiling: ncu smsp__inst_executed per ray (profiles/) x rays / time / (148 SMs x 4 x SM clock)"}
        extra["roofline_connect"] = {"bound": "hbm", "kernel": "k_connect (any-hit traversal)", "achieved": ach_any, "peak": peak,
                                     "unit": "GB/s", "frac": ach_any / peak, "bytes_per_ray": b_any,
                                     "avg_launch_ms": con_ms / max(1, con_l), "launches": con_l,
                                     "share_of_step": con_ms / ms if N == 1 else None}
        extra["render_ms_each_step"] = step_ms
        extra["gpu_ms_each_step"] = gpu_step_ms
        extra["host_enqueue_ms_each_step"] = host_ms
        extra["shade_share_of_step"] = shd_ms / ms if N == 1 else None
        extra["mrays_per_s_closest_kernel"] = rays_c / (ext_ms * 1e-3) / 1e6 if ext_ms > 0 else None
        extra["mrays_per_s_any_kernel"] = rays_a / (con_ms * 1e-3) / 1e6 if con_ms > 0 else None

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "Msamples/s", "n_gpus": N, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": make_config(args, info, N),
                "mrays_per_s": mrays, "rays_per_sample": (rays_c_all + rays_a_all) / samples,
                "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps, "api": "pb_render(PB_RENDER_HOST_BUFFERS | PB_RENDER_ASYNC) x steps + pb_wait", "host_enqueue_ms_each": e2e_async_host,
                        "sync": {"value": sync_value, "ms_per_step": sync_ms / e2e_steps, "ms_each": e2e_each, "api": "pb_render(PB_RENDER_HOST_BUFFERS), one synchronous call per step"}},
                "launch_path": {"value": "kernel by kernel (PB_RENDER_STAGE_TIMING: CUDA events around every traversal launch feed the roofline)",
                                "graph": {"value": graph_value, "ms_per_step": graph_ms / args.steps,
                                          "note": "the same steps on the shipped default path: one CUDA graph launch per wave"},
                                "e2e": "graph"},
                "parity_check": parity,
                "gpu_launches": int(launches_all), "roofline": roof, "cpu_baseline": cpu, "clocks": clk,
                "bvh_build_ms": st0["build_ms"], "bvh_nodes": st0["n_nodes"]}
        line.update(extra)
        emit(line)
    if N > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


_JSON_FD = None


def claim_stdout():
    """stdout carries the ONE JSON line and nothing else: from here on fd 1 points at stderr, so that libraries which
    printf to it (NCCL's version banner under NCCL_DEBUG=VERSION ignores NCCL_DEBUG_FILE) cannot get in front of it."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="product", choices=["product", "reference"])
    ap.add_argument("--spp-per-step", type=int, default=32)
    ap.add_argument("--e2e-steps", type=int, default=8)
    ap.add_argument("--mesh", default="10M", choices=["1M", "10M", "100M"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--film", default="1080p", choices=["1080p", "4k"], help="4k = BASELINE config C5's 3840x2160 film (not the headline metric)")
    args = ap.parse_args()
    if args.film == "4k":
        global WIDTH, HEIGHT, WORKLOAD, METRIC
        METRIC = "Msamples/s at 4K pt+MIS"
        WIDTH, HEIGHT = 3840, 2160
        WORKLOAD = WORKLOAD.replace("C4:", "C5:").replace("1920x1080", "3840x2160")
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        claim_stdout()
        return reference_main(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus != world and world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under torch.distributed.run, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 1000), os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    claim_stdout()
    return product_main(args)


if __name__ == "__main__":
    sys.exit(main())
