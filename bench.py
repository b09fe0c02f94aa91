#!/usr/bin/env python
"""bench.py — Mrays/s (primary + shadow) and ms/frame at 3840x2160 on the synthetic
10 008 338-triangle terrain (BASELINE.json configs[4]), 1/2/4/8 B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

One step = one full frame: 9 samples per pixel, primary ray + one shadow ray per
directional/point light per hit, shading, SSAA resolve, RGBA8 frame resident on rank 0.
`value` times K frames back to back on the device (scene + hierarchy resident), `e2e`
times the drop-in call render()/rtb_render with host buffers (scene upload + hierarchy
build + frame + download every step).  `--impl reference` times the reference's own CPU
renderer sources (oracle/_ref/libref_cpu.so) on a bounded sample of the same workload.
Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

_JSON_OUT = sys.stdout      # main() points it at the real stdout and sends everything else to stderr

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))

GRID = 2237                 # 2 * 2237^2 = 10 008 338 triangles
WIDTH, HEIGHT = 3840, 2160
ALIASING = 3
SEED = 0xB200
METRIC = "Mrays/s (primary+shadow) at 3840x2160, 3x3 SSAA, 10M-triangle synthetic terrain"
# algorithmic work per ray for T = 10 008 338 (24 node visits; 8 / 4 triangle tests)
OPS_PRIMARY, OPS_SHADOW = 1005, 700
BYTES_PRIMARY, BYTES_SHADOW = 2064, 1760


def peaks():
    p = {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "source": "fallback (B200_PROFILING.md)"}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        p["hbm_gbs"] = float(m["hbm_gbs"])
        p["sm_max_mhz"] = float(m.get("sm_max_mhz", 1965.0))
        p["source"] = "measured (MEASURED_PEAKS.json)"
    except Exception:
        pass
    return p


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


# ------------------------------------------------------------------------------------
# CPU baseline: the reference's cpu/*.c sources (oracle/_ref/libref_cpu.so) on a bounded
# sample of the SAME workload.  Brute force over 10 M triangles per ray, so the sample is
# a few hundred image-plane positions, spread over all host threads.
# ------------------------------------------------------------------------------------
def cpu_reference_run(hs, n_samples, threads, seed):
    """Returns (rays, seconds).  hs: HostScene at OUTPUT resolution (the CPU renderer's own
    convention: camera W H as in the scene file, 2x2 sub-samples at offsets {0, 0.5})."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_py
    L = oracle_py.ref_lib("cpu")
    kind = "reference"
    if L is None:
        raise RuntimeError("oracle/_ref/libref_cpu.so is missing (build it where /root/reference exists)")
    rng = np.random.default_rng(seed)
    # cpu/raytracer.c:50-58: i in (-W/2, W/2], j in (-H/2, H/2], sub-sample offsets 0 / 0.5
    xy = np.stack([rng.integers(-WIDTH // 2 + 1, WIDTH // 2 + 1, n_samples) + 0.5 * rng.integers(0, 2, n_samples),
                   rng.integers(-HEIGHT // 2 + 1, HEIGHT // 2 + 1, n_samples) + 0.5 * rng.integers(0, 2, n_samples)],
                  axis=1).astype(np.float32)
    rgb = np.zeros((n_samples, 3), np.float32)
    rays = np.zeros(2, np.uint64)
    h = L.refcpu_scene_create(C.byref(hs.s))
    try:
        t0 = time.perf_counter()
        L.refcpu_trace_samples(h, xy.ctypes.data, n_samples, rgb.ctypes.data, threads, rays.ctypes.data)
        dt = time.perf_counter() - t0
    finally:
        L.refcpu_scene_destroy(h)
    return int(rays[0] + rays[1]), dt, kind


TRIANGLES = 2 * GRID * GRID
WORKLOAD = ("C5 synthetic terrain %d triangles (grid %d, seed 0x%X), %dx%d x %dx%d SSAA, 1 ambient + 1 directional + 1 point light"
            % (TRIANGLES, GRID, SEED, WIDTH, HEIGHT, ALIASING, ALIASING))


class HostOnlyScene:
    """The C5 scene from librt_host.so (host surface only: no CUDA, none of the repo's tracing code) --
    what the reference arm uses, so that librtb200.so is never mapped into that process."""

    def __init__(self, grid, width, height, seed):
        import pyrtb                      # ctypes struct definitions only; pyrtb.lib() is never called here
        self.L = C.CDLL(os.path.join(ROOT, "raytracing-gpu_b200", "librt_host.so"))
        self.s = pyrtb.SceneStruct()
        self.L.rtb_synth_terrain.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_int, C.c_int]
        if self.L.rtb_synth_terrain(C.byref(self.s), grid, seed, width, height) != 0:
            raise RuntimeError("rtb_synth_terrain failed")
        self.triangle_count = int(self.s.triangle_count)


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    hs = HostOnlyScene(GRID, WIDTH, HEIGHT, SEED)
    per_step = max(threads, args.ref_samples)
    for _ in range(args.warmup):
        cpu_reference_run(hs, max(threads // 4, 1), threads, 1)
    total_rays, total_s = 0, 0.0
    for k in range(args.steps):
        r, dt, kind = cpu_reference_run(hs, per_step, threads, 100 + k)
        total_rays += r
        total_s += dt
    v = total_rays / total_s / 1e6
    sample = ("%d random image-plane samples per step (cpu/raytracer.c sub-sample grid) of the %dx%d frame, "
              "brute force over all %d triangles, sources built -O2" % (per_step, WIDTH, HEIGHT, hs.triangle_count))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total_s / max(args.steps, 1), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample,
                   "semantics": "the reference's CPU renderer (cpu/: 2x2 sub-samples, float colours); a sampled baseline, "
                                "not a parity target"},
        "native_so": "raytracing-gpu_b200/librt_host.so (scene generation only, no CUDA) + oracle/_ref/libref_cpu.so",
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), file=_JSON_OUT, flush=True)


# ------------------------------------------------------------------------------------
# C1-C4 of BASELINE.json (the shipped scenes at their quoted sizes): our frame / render() and the
# reference's own GPU render() (oracle/_ref/libref_gpu.so: its unmodified gpu/*.cu built for sm_100,
# run in a child process) on the same box -- SURVEY 8(d)'s "existing kernel on the same box" bar.
# Context for the headline, not the headline: the reference cannot run C5 (object-level tree).
# ------------------------------------------------------------------------------------
SMALL = [("C1", "triangle", 512, 512), ("C2a", "dir-light-shadows", 1920, 1080), ("C2b", "point-light", 1920, 1080),
         ("C3", "susans_smooth", 1920, 1080), ("C4", "car-on-road", 3840, 2160)]


def small_configs(args):
    import pyrtb
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    out = {}
    try:
        from util import load_golden
    except Exception as ex:
        return {"unavailable": str(ex)}
    ref_so = os.path.join(ROOT, "oracle", "_ref", "libref_gpu.so")
    for cfg, name, w, h in SMALL:
        try:
            hs = load_golden(name)[0].with_resolution(w, h)
            ds = pyrtb.DeviceScene(hs, ALIASING)
            ds.render(count=True)
            st = ds.stats()
            ts = []
            for _ in range(8):
                ds.render()
                ts.append(ds.stats()["frame_ms"])
            ds.close()
            img = np.empty((h, w, 4), np.uint8)
            to = []
            for _ in range(3):
                t0 = time.perf_counter()
                pyrtb.render_oneshot(hs, ALIASING, out=img)
                to.append(1e3 * (time.perf_counter() - t0))
            rays = st["rays_primary"] + st["rays_shadow"]
            row = {"scene": name, "output": "%dx%d" % (w, h), "triangles": hs.triangle_count, "rays": int(rays),
                   "frame_ms": float(np.median(ts[3:])), "mrays_s": rays / float(np.median(ts[3:])) / 1e3,
                   "render_ms": min(to[1:]), "ref_gpu_render_ms": None}
            if os.path.exists(ref_so):
                try:
                    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "reports", "config_table.py"), "--ref-gpu-one",
                                        name, str(w), str(h)], capture_output=True, text=True, timeout=120)
                    ln = [l for l in r.stdout.splitlines() if l.startswith("REFGPU")]
                    if ln:
                        row["ref_gpu_render_ms"] = float(ln[-1].split()[1])
                        row["render_speedup_vs_ref_gpu"] = row["ref_gpu_render_ms"] / row["render_ms"]
                except subprocess.TimeoutExpired:
                    row["ref_gpu_render_ms"] = "timeout (120 s)"
            out[cfg] = row
        except Exception as ex:
            out[cfg] = {"scene": name, "error": str(ex)}
    return out


# ------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------
def run_ours(args):
    import pyrtb
    rank, local_rank, world = dist_env()
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"      # the version banner goes to stdout, where the one JSON line belongs
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if pyrtb.device_count() <= local_rank:
        raise SystemExit("bench.py needs a GPU per rank: librtb200 has no CPU fallback")
    L = pyrtb.lib()

    host_pg = dist.new_group(backend="gloo") if dist is not None else None

    def barrier():
        if dist is not None:
            dist.barrier()

    def host_barrier():
        """the waiting ranks sleep on the host: an NCCL barrier is a kernel that spins on every waiting rank's GPU, and
        rank 0's N-device render() would share the devices with it"""
        if dist is not None:
            import torch
            torch.cuda.synchronize()
            dist.barrier(group=host_pg)

    def max_over_ranks(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    t_gen = time.perf_counter()
    hs = pyrtb.HostScene.terrain(GRID, WIDTH, HEIGHT, SEED)
    t_gen = time.perf_counter() - t_gen
    ds = pyrtb.DeviceScene(hs, ALIASING, device=local_rank)
    st0 = ds.stats()
    frame_bytes = WIDTH * HEIGHT * 4
    pitch = WIDTH * 4

    # rank 0 owns the framebuffer; the other ranks store their tiles straight into it
    # through a CUDA-IPC mapping (NVLink peer stores) — no gather collective.
    # (two frames, for the double-buffered resident loop: frame k + 1 renders while frame k is downloaded)
    def shared_frame():
        d = C.c_void_p()
        if rank == 0:
            pyrtb._check(L.rtb_dev_alloc(local_rank, frame_bytes, C.byref(d)))
            pyrtb._check(L.rtb_dev_memset(d, 0, frame_bytes))
        if world > 1:
            import torch
            handle = torch.zeros(64, dtype=torch.uint8)
            if rank == 0:
                hb = (C.c_ubyte * 64)()
                pyrtb._check(L.rtb_ipc_export(d, hb))
                handle = torch.tensor(list(hb), dtype=torch.uint8)
            handle = handle.cuda()
            dist.broadcast(handle, 0)
            if rank != 0:
                hb = (C.c_ubyte * 64)(*handle.cpu().tolist())
                pyrtb._check(L.rtb_ipc_open(hb, C.byref(d)))
        return d

    d_frame = shared_frame()
    d_frame_b = shared_frame()

    # Sharded projection pass (rtb_scene_exchange_*): every rank projects 1/N of the triangles and stores the
    # items straight into the owners' inboxes through CUDA-IPC mappings of the peers' exchange blocks (NVLink
    # stores inside the kernel, device-side epoch barrier) -- the path's one real exchange step, no collective.
    peer_blocks = []
    if world > 1 and not args.replicated_setup:
        import torch
        blk, blk_bytes = ds.exchange_create(world)
        hb = (C.c_ubyte * 64)()
        pyrtb._check(L.rtb_ipc_export(C.c_void_p(blk), hb))
        mine = torch.tensor(list(hb), dtype=torch.uint8, device="cuda")
        allh = [torch.zeros(64, dtype=torch.uint8, device="cuda") for _ in range(world)]
        dist.all_gather(allh, mine)
        blocks = []
        for q in range(world):
            if q == rank:
                blocks.append(blk)
                continue
            p = C.c_void_p()
            hq = (C.c_ubyte * 64)(*allh[q].cpu().tolist())
            pyrtb._check(L.rtb_ipc_open(hq, C.byref(p)))
            peer_blocks.append(p)
            blocks.append(p.value)
        ds.exchange_attach(rank, world, blocks)
        barrier()

    def frame(count=False, into=None):
        ds.render_device(into or d_frame, pitch, rank=rank, nranks=world, count=count)

    # Camera path: the headline times a fly-through, one new camera per frame (a dolly along the view
    # direction with a little side-slip), so that nothing a frame learns about its camera (the
    # longest-first item order, the projection pass's verdict) is reused on the camera it came from.
    cam0 = pyrtb.Camera()
    C.memmove(C.byref(cam0), C.byref(hs.scaled(ALIASING).camera), C.sizeof(pyrtb.Camera))
    u = np.array([cam0.u.x, cam0.u.y, cam0.u.z]); v = np.array([cam0.v.x, cam0.v.y, cam0.v.z])
    u /= np.linalg.norm(u); v /= np.linalg.norm(v)
    forward = -np.cross(u, v)                       # the scene lies opposite the image plane (gpu/raytracer.cu:181-182)

    def camera_at(k):
        c = pyrtb.Camera()
        C.memmove(C.byref(c), C.byref(cam0), C.sizeof(pyrtb.Camera))
        p = np.array([cam0.position.x, cam0.position.y, cam0.position.z]) + forward * (0.25 * k) + u * (0.1 * k)
        c.position = pyrtb.Vec3(float(p[0]), float(p[1]), float(p[2]))
        return c

    def count_rays(cam):
        """one untimed instrumented frame (device counters); it walks the tree for every ray"""
        ds.set_camera(cam)
        frame(count=True)
        ds.sync()
        st = ds.stats()
        cnt = np.array([st["rays_primary"], st["rays_shadow"], st["rays_bounce"], st["node_visits"], st["tri_tests"]],
                       dtype=np.float64)
        if dist is not None:
            import torch
            t = torch.tensor(cnt, dtype=torch.float64, device="cuda")
            dist.all_reduce(t)
            cnt = t.cpu().numpy()
        return cnt

    path = [camera_at(k) for k in range(args.steps)]
    counts0 = count_rays(cam0)
    counts_path = np.sum([count_rays(c) for c in path], axis=0)
    primary, shadow, bounce, node_visits, tri_tests = [int(x) for x in counts0]
    rays = primary + shadow                                   # per frame, static camera
    rays_path = int(counts_path[0] + counts_path[1])          # over the K frames of the path

    for k in range(args.warmup, 0, -1):                       # warm-up frames on the cameras BEFORE the path
        ds.set_camera(camera_at(-k))
        frame()
    ds.sync()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    # ---- timed region: K frames, one camera each, CUDA events on the launching stream ----
    barrier()
    ds.timer_start()
    for k in range(args.steps):
        ds.set_camera(path[k])
        frame()
    ms_total = ds.timer_stop()
    barrier()
    ms_total = max_over_ranks(ms_total)
    # ---- the same with one camera for all frames (what round 1 reported) ----
    ds.set_camera(cam0)
    for _ in range(args.warmup):
        frame()
    ds.sync()
    barrier()
    ds.timer_start()
    for _ in range(args.steps):
        frame()
    ms_static = ds.timer_stop()
    barrier()
    ms_static = max_over_ranks(ms_static)
    clocks = sampler.stop() if rank == 0 else None
    barrier()       # (stopping the sampler takes rank 0 a moment: the others would wait for its items inside their first frame)
    # per-kernel durations of a frame (events between the launches, same stream), moving camera again
    ds.profile(True)
    kernel_ms, parts = [], []
    for k in range(min(args.steps, 10)):
        ds.set_camera(path[k])
        frame()
        ds.sync()
        kernel_ms.append(ds.stats()["frame_ms"])
        parts.append(ds.kernel_ms())
    ds.profile(False)
    kernel_ms_avg = max_over_ranks(float(np.mean(kernel_ms)))
    parts = np.mean(np.array(parts), axis=0)
    launches_per_frame = int(ds.stats()["kernel_launches"])      # projection pass (3) + frame kernel
    rstate = ds.raster_state()
    per_rank_ms = [float(np.mean(kernel_ms))]
    if dist is not None:
        import torch
        t = torch.zeros(world, dtype=torch.float64, device="cuda")
        t[rank] = per_rank_ms[0]
        dist.all_reduce(t)
        per_rank_ms = [round(float(x), 3) for x in t.cpu().tolist()]
    ds.set_camera(cam0)
    barrier()

    # ---- e2e: the drop-in call with HOST buffers ----
    pinned = pyrtb.PinnedBuffer(frame_bytes)
    host_img = pinned.array((HEIGHT, WIDTH, 4))
    e2e = None
    if world == 1:
        e2e_steps = max(2, min(args.steps, args.e2e_steps))
        for _ in range(2):                                        # warm-up: allocator pools grow over two calls, pinned staging
            pyrtb.render_oneshot(hs, ALIASING, out=host_img)
        per_call = []
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            t1 = time.perf_counter()
            pyrtb.render_oneshot(hs, ALIASING, out=host_img)
            per_call.append(1e3 * (time.perf_counter() - t1))
        mean_s = (time.perf_counter() - t0) / e2e_steps
        oneshot_img = host_img.copy()
        # The call makes ~400 driver calls and ~12 host round trips; on a shared host one call in ten is held up for
        # tens to hundreds of ms somewhere among them (80 calls in a row: 52-53 ms typical, excursions to 450).
        # value = the MEDIAN call; the mean and every call are beside it.
        e2e_s = float(np.median(per_call)) / 1e3
        h2d = hs.triangle_count * 72 + 72 * hs.s.object_count + 28 * hs.s.light_count + 48
        e2e = {"value": rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": frame_bytes, "ms_per_step": 1e3 * e2e_s, "statistic": "median of %d calls" % e2e_steps,
               "ms_per_step_mean": 1e3 * mean_s, "ms_per_call": [round(x, 2) for x in per_call],
               "call": "rtb_render (= render(), gpu/headers/raytracer.h:6): scene upload + hierarchy build + frame + download"}
        # resident-scene variants: camera in, frame out.  (a) rtb_scene_render, one synchronous call per frame;
        # (b) double-buffered (below, the same loop at every N)
        ds.render(out=host_img)
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            ds.render(out=host_img)
        res_sync_s = (time.perf_counter() - t0) / e2e_steps
    else:
        # the SAME call as at N = 1: the one-shot render(), here on N devices of rank 0's process
        # (rtb_render_multi: every device uploads the scene and builds its hierarchy, renders its tiles
        # into device 0's frame over NVLink; rank 0 downloads).  The other ranks wait.
        e2e_steps = max(2, min(args.steps, args.e2e_steps))
        host_barrier()
        e2e_s, mean_s, per_call = 0.0, 0.0, []
        if rank == 0:
            for _ in range(2):                                        # warm-up: contexts, peer access, pools
                pyrtb.render_multi(hs, world, ALIASING, out=host_img)
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                t1 = time.perf_counter()
                pyrtb.render_multi(hs, world, ALIASING, out=host_img)
                per_call.append(1e3 * (time.perf_counter() - t1))
            mean_s = (time.perf_counter() - t0) / e2e_steps
            e2e_s = float(np.median(per_call)) / 1e3          # as at N = 1: the median call, mean and every call beside it
            multi_img = host_img.copy()
        host_barrier()
        h2d = hs.triangle_count * 72 + 72 * hs.s.object_count + 28 * hs.s.light_count + 48
        e2e = {"value": rays / e2e_s / 1e6 if rank == 0 else None, "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d) * world,
               "d2h_bytes_per_step": frame_bytes, "ms_per_step": 1e3 * e2e_s, "statistic": "median of %d calls" % e2e_steps,
               "ms_per_step_mean": 1e3 * mean_s, "ms_per_call": [round(x, 2) for x in per_call],
               "call": "rtb_render_multi (= render() with RT_GPUS = N): scene upload + hierarchy build on every device, "
                       "tiles stored into device 0's frame over NVLink, download"}
        res_sync_s = None
    # resident scene, double-buffered, the same loop at every N: every rank renders its tiles of frame k into one
    # of two frames in rank 0's memory (NVLink peer stores); when all ranks are done rank 0 starts the download of
    # that frame on a copy stream (rtb_dev_download_async) and everybody goes on with frame k + 1 in the other
    # buffer.  Every frame is downloaded (33 MB into pinned memory); the last download is waited for inside the
    # timed region.
    pinned_b = pyrtb.PinnedBuffer(frame_bytes)
    host_b = pinned_b.array((HEIGHT, WIDTH, 4))
    bufs = ((d_frame, host_img), (d_frame_b, host_b))
    res_steps = max(4, min(args.steps, 2 * args.e2e_steps))
    for k in range(2):                              # warm-up: both buffers, the copy stream
        frame(into=bufs[k][0])
        ds.sync()
        barrier()
        if rank == 0:
            pyrtb._check(L.rtb_dev_download_async(bufs[k][1].ctypes.data, bufs[k][0], frame_bytes))
    if rank == 0:
        pyrtb._check(L.rtb_dev_download_wait(d_frame))
    barrier()
    t0 = time.perf_counter()
    for k in range(res_steps):
        ds.set_camera(path[k % len(path)])
        dev_buf, host_buf = bufs[k % 2]
        if rank == 0 and k >= 2:
            pyrtb._check(L.rtb_dev_download_wait(dev_buf))          # (the copy two frames back: long finished)
        frame(into=dev_buf)
        ds.sync()
        barrier()
        if rank == 0:
            pyrtb._check(L.rtb_dev_download_async(host_buf.ctypes.data, dev_buf, frame_bytes))
    if rank == 0:
        pyrtb._check(L.rtb_dev_download_wait(d_frame))
    barrier()
    res_s = max_over_ranks((time.perf_counter() - t0) / res_steps)
    ds.set_camera(cam0)
    e2e["resident"] = {"value": rays_path / len(path) / res_s / 1e6, "ms_per_step": 1e3 * res_s, "h2d_bytes_per_step": 48,
                       "d2h_bytes_per_step": frame_bytes, "steps": res_steps,
                       "call": "rtb_scene_set_camera + rtb_scene_render_device on every rank (tiles stored into rank 0's frame "
                               "over NVLink) + rtb_dev_download_async, double-buffered: frame k is downloaded while frame k + 1 renders"}
    if res_sync_s is not None:
        e2e["resident"]["synchronous_ms_per_step"] = 1e3 * res_sync_s      # rtb_scene_render: frame, then download, one call

    # frame check: the composed frame equals a single-GPU frame of the same scene (camera 0)
    ds.set_camera(cam0)
    frame()
    ds.sync()
    barrier()
    checksum = None
    oneshot_diff = None
    if rank == 0:
        pyrtb._check(L.rtb_dev_download(host_img.ctypes.data, d_frame, frame_bytes))
        composed = host_img.copy()
        if world > 1:
            single = ds.render()
            if not np.array_equal(single, composed):
                raise SystemExit("multi-GPU frame differs from the single-GPU frame")
            # the one-shot call builds the one-frame hierarchy (no cluster rebuild above 2^18 triangles)
            oneshot_diff = int((single != multi_img).any(axis=-1).sum())
            if oneshot_diff > 100:
                raise SystemExit("rtb_render_multi frame differs from the single-GPU frame in %d pixels" % oneshot_diff)
        else:
            # the one-shot call renders without the node array (deferred tree, stray shadow rays by brute force): same pixels?
            oneshot_diff = int((composed != oneshot_img).any(axis=-1).sum())
            if oneshot_diff > 100:
                raise SystemExit("rtb_render frame differs from the resident scene's frame in %d pixels" % oneshot_diff)
        checksum = int(composed.astype(np.uint64).sum())

    if rank == 0:
        pk = peaks()
        sm_count = 148
        issue_peak_tflops = sm_count * 128 * pk["sm_max_mhz"] * 1e6 / 1e12     # un-fused FP32 op/s = issue slots
        step_s = ms_total / 1e3 / args.steps
        static_s = ms_static / 1e3 / args.steps
        kern_s = kernel_ms_avg / 1e3
        alg_ops = primary * OPS_PRIMARY + shadow * OPS_SHADOW + bounce * OPS_PRIMARY
        alg_bytes = primary * BYTES_PRIMARY + shadow * BYTES_SHADOW + bounce * BYTES_PRIMARY
        achieved_tflops = alg_ops / world / kern_s / 1e12        # per GPU: each rank's kernel does 1/world of the rays
        kernels_live = {"clears+k_rast_setup": float(parts[4]), "k_rast_big": float(parts[5]), "k_xchg_sync": float(parts[6]),
                        "k_rast_items": float(parts[1]), "k_frame": float(parts[2]), "frame": float(parts[3])}
        dominant = max(("clears+k_rast_setup", "k_rast_big", "k_rast_items", "k_frame"), key=lambda k: kernels_live[k])
        roofline = {
            "bound": "fp32_issue",
            "kernel": "k_frame<tree>" if launches_per_frame == 1 else "the frame's kernels: k_rast_setup + k_rast_big + k_rast_items + k_frame<tree>",
            "achieved": achieved_tflops, "peak": issue_peak_tflops,
            "unit": "TFLOP/s", "frac": achieved_tflops / issue_peak_tflops, "traffic": None,
            "equivalent_tree_walk_frac": achieved_tflops / issue_peak_tflops,
            "peak_source": "148 SMs x 128 FP32 lanes x sm_max_mhz, %s" % pk["source"],
            "note": "frac is NOMINAL ('equivalent tree-walk throughput'): SURVEY 8(d)'s per-ray op counts describe tree walks "
                    "(24 node visits, 8 / 4 triangle tests), while the timed frame resolves primary rays by the projection "
                    "pass and shadow rays by the light grids; it is algorithmic ops per frame / CUDA-event duration of the "
                    "frame's launches / the un-fused FP32 (= issue) peak.  What the kernels actually achieve is under "
                    "'measured' (ncu: issue-slot utilisation, lanes per instruction, DRAM bytes).  Divergent traversal + scalar "
                    "FP32: neither HBM nor tensor cores bind (SURVEY 8d)",
            "algorithmic_ops_per_launch": alg_ops // world, "kernel_ms": kernel_ms_avg,
            "kernel_ms_per_rank": per_rank_ms,
            "kernels_ms_live": kernels_live, "dominant_kernel": dominant,
            "kernel_share_live": {k: kernels_live[k] / max(kernels_live["frame"], 1e-9) for k in kernels_live if k != "frame"},
            "hbm": None,
            "request_volume_gbs": alg_bytes / world / kern_s / 1e9,   # cache-less: 64 B per node visit, 48 B per triangle test
            "tree_walk_counts_per_ray": {"node_visits": node_visits / max(rays + bounce, 1),
                                         "tri_tests": tri_tests / max(rays + bounce, 1),
                                         "note": "device counters of the UNTIMED counting frame, which walks the tree for every "
                                                 "ray; the timed frame does not (see paths); the ray counts come from that frame too"},
            "paths": {"primary": "projection pass (raster.cuh)" if launches_per_frame > 1 and not rstate["gave_up"] else "tree walk",
                      "shadow": "%d light grid(s) (lightgrid.cuh)" % rstate["light_grids"] if rstate["light_grids"] else "tree walk",
                      "projection_items": rstate["items"], "projection_work_samples": rstate["area"],
                      "light_grid_entries_k": rstate["grid_entries_k"], "launches_per_frame": launches_per_frame},
        }
        # measured counters of the frame's kernels: one ncu --set full capture of tools/perf_frame.py on this
        # workload, summarised by tools/ncu_json.py and committed (a number taken under ncu is never a bench value)
        ncu_file = os.path.join(ROOT, "profiles", "r2_ncu_frame.json")
        if os.path.exists(ncu_file):
            try:
                nj = json.load(open(ncu_file))
                roofline["measured"] = {"source": "profiles/r2_ncu_frame.json (%s)" % nj.get("source", ""), "kernels": nj["kernels"]}
                roofline["traffic"] = int(sum(k["dram_bytes"] for k in nj["kernels"].values()))
                dk = nj["kernels"].get("k_frame")
                if dk:
                    roofline["issue_active_frac_measured"] = dk["issue_active_pct"] / 100.0
            except Exception as ex:
                roofline["measured"] = {"error": str(ex)}
        if roofline["traffic"] and world == 1:
            # the HBM view SURVEY 8(d) asks for beside the issue roofline: DRAM bytes of the ncu capture / launch time
            gbs = roofline["traffic"] / kern_s / 1e9
            roofline["hbm"] = {"achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": gbs / pk["hbm_gbs"],
                               "note": "dram__bytes_read+write of the frame's launches in the ncu capture / live duration of "
                                       "the frame's launches: HBM does not bind these kernels"}
        configs = None
        if world == 1 and not args.no_ref_gpu:
            configs = small_configs(args)
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            n = max(threads, args.ref_samples)
            try:
                r, dt, kind = cpu_reference_run(hs, n, threads, 1234)
                cpu = {"value": r / dt / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind,
                       "sample": "%d random image-plane samples of the same frame through cpu/hit.c collide + "
                                 "cpu/light.c apply_light (brute force over %d triangles), %.1f s, sources built -O2"
                                 % (n, hs.triangle_count, dt)}
            except Exception as ex:   # the checker library is test infrastructure: report, do not fail the bench
                cpu = {"value": None, "unit": "Mrays/s", "cores": threads, "kind": "reference", "sample": "unavailable: %s" % ex}
        out = {
            "metric": METRIC, "value": rays_path / (ms_total / 1e3) / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * step_s, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "static_camera": {"value": rays / static_s / 1e6, "unit": "Mrays/s", "ms_per_step": 1e3 * static_s,
                              "note": "the same K frames with ONE camera: longest-first scheduling then replays the costs of the "
                                      "identical previous frame; the headline moves the camera every frame"},
            "config": {"workload": WORKLOAD,
                       "camera": "fly-through: one new camera per timed frame (0.25 units forward + 0.1 sideways per frame, "
                                 "rtb_scene_set_camera); rays counted per camera in untimed frames",
                       "parallelism": "interleaved 32x32-pixel tiles over %d GPU(s), scene replicated, tiles stored "
                                      "into rank 0's frame over NVLink%s" % (world, "" if world == 1 else (
                                          "; projection pass replicated on every rank" if args.replicated_setup else
                                          "; projection pass sharded by triangles, items stored into the owners' inboxes over NVLink "
                                          "(CUDA IPC), device-side epoch barrier")),
                       "l2": "inputs larger than L2 (triangles 960 MB + nodes %d MB vs 126 MB L2)" % (st0["tree_bytes"] >> 20),
                       "rays_per_frame": {"primary": primary, "shadow": shadow, "bounce": bounce}},
            "e2e": e2e, "gpu_launches": args.steps * launches_per_frame, "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
            "build": {"upload_ms": st0["upload_ms"], "build_ms": st0["build_ms"], "nodes": st0["nodes"],
                      "leaves": st0["leaves"], "max_depth": st0["max_depth"], "build_launches": st0["build_launches"],
                      "scene_generation_s": t_gen},
            "frame_checksum": checksum, "oneshot_vs_resident_pixels_differ": oneshot_diff,
            "configs": configs,
        }
        print(json.dumps(out), file=_JSON_OUT, flush=True)
    barrier()
    pinned.free()
    ds.sync()
    barrier()
    for p in peer_blocks:
        L.rtb_ipc_close(p)
    barrier()
    ds.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--ref-samples", type=int, default=0, help="CPU samples per step (default: one per host thread x4)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--replicated-setup", action="store_true",
                    help="N > 1: every rank projects all triangles (round 1's path) instead of the sharded pass with item exchange")
    ap.add_argument("--no-ref-gpu", action="store_true", help="skip the C1-C4 block (ours and the reference's GPU renderer)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.ref_samples <= 0:
        args.ref_samples = 4 * (os.cpu_count() or 1)
    # stdout carries exactly one line, the JSON: everything libraries print there (NCCL's version
    # banner under torchrun, for one) goes to stderr instead
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
