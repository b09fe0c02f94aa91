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


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    import pyrtb
    threads = os.cpu_count() or 1
    hs = pyrtb.HostScene.terrain(GRID, WIDTH, HEIGHT, SEED)
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
        "config": {"workload": "C5 synthetic terrain %d triangles, %dx%d, 1 ambient + 1 directional + 1 point light"
                   % (hs.triangle_count, WIDTH, HEIGHT), "sample": sample},
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), file=_JSON_OUT, flush=True)


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

    def barrier():
        if dist is not None:
            dist.barrier()

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
    d_frame = C.c_void_p()
    if rank == 0:
        pyrtb._check(L.rtb_dev_alloc(local_rank, frame_bytes, C.byref(d_frame)))
        pyrtb._check(L.rtb_dev_memset(d_frame, 0, frame_bytes))
    if world > 1:
        import torch
        handle = torch.zeros(64, dtype=torch.uint8)
        if rank == 0:
            hb = (C.c_ubyte * 64)()
            pyrtb._check(L.rtb_ipc_export(d_frame, hb))
            handle = torch.tensor(list(hb), dtype=torch.uint8)
        handle = handle.cuda()
        dist.broadcast(handle, 0)
        if rank != 0:
            hb = (C.c_ubyte * 64)(*handle.cpu().tolist())
            pyrtb._check(L.rtb_ipc_open(hb, C.byref(d_frame)))

    def frame(count=False):
        ds.render_device(d_frame, pitch, rank=rank, nranks=world, count=count)

    # ray accounting: one untimed instrumented frame (device counters)
    frame(count=True)
    ds.sync()
    st = ds.stats()
    counts = np.array([st["rays_primary"], st["rays_shadow"], st["rays_bounce"], st["node_visits"], st["tri_tests"]],
                      dtype=np.float64)
    if dist is not None:
        import torch
        t = torch.tensor(counts, dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        counts = t.cpu().numpy()
    primary, shadow, bounce, node_visits, tri_tests = [int(x) for x in counts]
    rays = primary + shadow

    for _ in range(args.warmup):
        frame()
    ds.sync()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    # ---- timed region: K frames, CUDA events on the launching stream ----
    kernel_ms = []
    barrier()
    ds.timer_start()
    for _ in range(args.steps):
        frame()
    ms_total = ds.timer_stop()
    barrier()
    ms_total = max_over_ranks(ms_total)
    clocks = sampler.stop() if rank == 0 else None
    # per-launch duration of the frame kernel (events around each launch, same stream)
    for _ in range(min(args.steps, 10)):
        frame()
        ds.sync()
        kernel_ms.append(ds.stats()["frame_ms"])
    kernel_ms_avg = max_over_ranks(float(np.mean(kernel_ms)))
    launches_per_frame = int(ds.stats()["kernel_launches"])      # projection pass (2) + frame kernel
    rstate = ds.raster_state()
    per_rank_ms = [float(np.mean(kernel_ms))]
    if dist is not None:
        import torch
        t = torch.zeros(world, dtype=torch.float64, device="cuda")
        t[rank] = per_rank_ms[0]
        dist.all_reduce(t)
        per_rank_ms = [round(float(x), 3) for x in t.cpu().tolist()]
    barrier()

    # ---- e2e: the drop-in call with HOST buffers ----
    pinned = pyrtb.PinnedBuffer(frame_bytes)
    host_img = pinned.array((HEIGHT, WIDTH, 4))
    e2e = None
    if world == 1:
        e2e_steps = max(2, min(args.steps, args.e2e_steps))
        pyrtb.render_oneshot(hs, ALIASING, out=host_img)          # warm-up (allocator, pinned staging)
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            pyrtb.render_oneshot(hs, ALIASING, out=host_img)
        e2e_s = (time.perf_counter() - t0) / e2e_steps
        h2d = hs.triangle_count * 72 + 72 * hs.s.object_count + 28 * hs.s.light_count + 48
        e2e = {"value": rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": frame_bytes, "ms_per_step": 1e3 * e2e_s,
               "call": "rtb_render (= render(), gpu/headers/raytracer.h:6): scene upload + hierarchy build + frame + download"}
        # resident-scene variant (rtb_scene_render): camera in, frame out
        ds.render(out=host_img)
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            ds.render(out=host_img)
        res_s = (time.perf_counter() - t0) / e2e_steps
        e2e["resident"] = {"value": rays / res_s / 1e6, "ms_per_step": 1e3 * res_s, "h2d_bytes_per_step": 48,
                           "d2h_bytes_per_step": frame_bytes, "call": "rtb_scene_render"}
    else:
        # all ranks render their tiles into rank 0's device frame; rank 0 downloads it
        e2e_steps = max(2, min(args.steps, args.e2e_steps))
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            frame()
            ds.sync()
            barrier()
            if rank == 0:
                pyrtb._check(L.rtb_dev_download(host_img.ctypes.data, d_frame, frame_bytes))
        barrier()
        e2e_s = max_over_ranks((time.perf_counter() - t0) / e2e_steps)
        e2e = {"value": rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 48,
               "d2h_bytes_per_step": frame_bytes, "ms_per_step": 1e3 * e2e_s,
               "call": "rtb_scene_render_device on every rank (tiles stored into rank 0's frame over NVLink) + download"}

    # frame check: the composed frame equals a single-GPU frame of the same scene
    checksum = None
    if rank == 0:
        pyrtb._check(L.rtb_dev_download(host_img.ctypes.data, d_frame, frame_bytes))
        composed = host_img.copy()
        if world > 1:
            single = ds.render()
            if not np.array_equal(single, composed):
                raise SystemExit("multi-GPU frame differs from the single-GPU frame")
        checksum = int(composed.astype(np.uint64).sum())

    if rank == 0:
        pk = peaks()
        sm_count = 148
        issue_peak_tflops = sm_count * 128 * pk["sm_max_mhz"] * 1e6 / 1e12     # un-fused FP32 op/s = issue slots
        step_s = ms_total / 1e3 / args.steps
        kern_s = kernel_ms_avg / 1e3
        alg_ops = primary * OPS_PRIMARY + shadow * OPS_SHADOW + bounce * OPS_PRIMARY
        alg_bytes = primary * BYTES_PRIMARY + shadow * BYTES_SHADOW + bounce * BYTES_PRIMARY
        achieved_tflops = alg_ops / world / kern_s / 1e12        # per GPU: each rank's kernel does 1/world of the rays
        roofline = {
            "bound": "fp32_issue",
            "kernel": "k_frame<tree>" if launches_per_frame == 1 else "the frame's kernels: k_rast_setup + k_rast_items + k_frame<tree>",
            "achieved": achieved_tflops, "peak": issue_peak_tflops,
            "unit": "TFLOP/s", "frac": achieved_tflops / issue_peak_tflops, "traffic": None,
            "peak_source": "148 SMs x 128 FP32 lanes x sm_max_mhz, %s" % pk["source"],
            "note": "divergent traversal + scalar FP32: the binding ceiling is SM instruction issue, not HBM or "
                    "tensor cores (SURVEY 8d); achieved = algorithmic un-fused FP32 ops per frame (SURVEY 8d per-ray "
                    "figures x rays) / CUDA-event duration of the frame's launches",
            "algorithmic_ops_per_launch": alg_ops // world, "kernel_ms": kernel_ms_avg,
            "kernel_ms_per_rank": per_rank_ms,
            "hbm": None,
            "request_volume_gbs": alg_bytes / world / kern_s / 1e9,   # cache-less: 64 B per node visit, 48 B per triangle test
            "measured_per_ray": {"node_visits": node_visits / max(rays + bounce, 1),
                                 "tri_tests": tri_tests / max(rays + bounce, 1),
                                 "note": "device counters of the untimed counting frame, which walks the tree for every ray"},
            "paths": {"primary": "projection pass (raster.cuh)" if launches_per_frame > 1 and not rstate["gave_up"] else "tree walk",
                      "shadow": "%d light grid(s) (lightgrid.cuh)" % rstate["light_grids"] if rstate["light_grids"] else "tree walk",
                      "projection_items": rstate["items"], "projection_area_samples": rstate["area"],
                      "light_grid_entries_k": rstate["grid_entries_k"], "launches_per_frame": launches_per_frame},
        }
        traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(traffic_file):
            try:
                tj = json.load(open(traffic_file))
                roofline["traffic"] = tj.get("frame_dram_bytes", tj.get("k_frame_dram_bytes_per_launch"))
            except Exception:
                pass
        if roofline["traffic"] and world == 1:
            # the HBM view SURVEY 8(d) asks for beside the issue roofline: DRAM bytes of the ncu capture / launch time
            gbs = roofline["traffic"] / kern_s / 1e9
            roofline["hbm"] = {"achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": gbs / pk["hbm_gbs"],
                               "note": "dram__bytes_read+write of one ncu --set full capture (profiles/traffic.json) / "
                                       "live duration of the frame's launches: HBM does not bind these kernels"}
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
            "metric": METRIC, "value": rays / step_s / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * step_s, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "C5 synthetic terrain %d triangles (grid %d, seed 0x%X), %dx%d x %dx%d SSAA, "
                                   "1 ambient + 1 directional + 1 point light" %
                                   (hs.triangle_count, GRID, SEED, WIDTH, HEIGHT, ALIASING, ALIASING),
                       "parallelism": "interleaved 32x32-pixel tiles over %d GPU(s), scene replicated, tiles stored "
                                      "into rank 0's frame over NVLink" % world,
                       "l2": "inputs larger than L2 (triangles 960 MB + nodes %d MB vs 126 MB L2)" % (st0["tree_bytes"] >> 20),
                       "rays_per_frame": {"primary": primary, "shadow": shadow, "bounce": bounce}},
            "e2e": e2e, "gpu_launches": args.steps * launches_per_frame, "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
            "build": {"upload_ms": st0["upload_ms"], "build_ms": st0["build_ms"], "nodes": st0["nodes"],
                      "leaves": st0["leaves"], "max_depth": st0["max_depth"], "build_launches": st0["build_launches"],
                      "scene_generation_s": t_gen},
            "frame_checksum": checksum,
        }
        print(json.dumps(out), file=_JSON_OUT, flush=True)
    barrier()
    pinned.free()
    ds.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--ref-samples", type=int, default=0, help="CPU samples per step (default: one per host thread x4)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
