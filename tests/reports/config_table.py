"""GPU box: the measurement configs C1..C5 of SURVEY 8(d) on one GPU, with the reference's
own GPU renderer (oracle/_ref/libref_gpu.so, unmodified gpu/*.cu for sm_100) and the
reference's CPU renderer (oracle/_ref/ref_cpu_rt[_O2], cpu/ built with its own flags,
4 hard-coded threads, 2x2 SSAA: a baseline, not a parity target) beside librtb200.
usage: python tests/reports/config_table.py [--cpu C1,C2a] [--frames 20] [--skip-ref-gpu]"""
import argparse
import ctypes as C
import os
import resource
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))  # util.py

CONFIGS = [("C1", "triangle", 512, 512), ("C2a", "dir-light-shadows", 1920, 1080), ("C2b", "point-light", 1920, 1080),
           ("C3", "susans_smooth", 1920, 1080), ("C4", "car-on-road", 3840, 2160), ("C5", None, 3840, 2160)]


def scene_of(name, w, h):
    import pyrtb
    from util import load_golden
    if name is None:
        return pyrtb.HostScene.terrain(2237, w, h)
    hs, _ = load_golden(name)
    return hs.with_resolution(w, h)


def ref_gpu(name, w, h):
    """child process: wall time of the reference's render() (upload + build + 2 kernels + download)"""
    hs = scene_of(name, w, h)
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_gpu.so"))
    L.refgpu_render.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_long, C.c_int, C.c_int]
    s = hs.scaled(3)
    out = np.zeros((h, w, 4), np.uint8)
    devnull, saved = os.open(os.devnull, os.O_WRONLY), os.dup(1)
    t = []
    for _ in range(3):
        os.dup2(devnull, 1)
        t0 = time.perf_counter()
        L.refgpu_render(C.byref(s), out.ctypes.data, 3, w * 4, h, w)
        t.append(time.perf_counter() - t0)
        os.dup2(saved, 1)
    print("REFGPU %.3f %d" % (1e3 * min(t[1:]), int(out.astype(np.uint64).sum())), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpu", default="C1,C2a")
    ap.add_argument("--cpu-scaled", default="C3:8,C4:12",
                    help="cfg:k pairs: reference CPU renderer at 1/k x 1/k of the resolution, time scaled by k*k (cost is linear in pixels)")
    ap.add_argument("--oracle", default="C1,C2a,C2b", help="configs also rendered by the oracle (our CPU restatement, all host threads)")
    ap.add_argument("--frames", type=int, default=20)
    ap.add_argument("--skip-ref-gpu", action="store_true")
    ap.add_argument("--ref-gpu-one", nargs=3, default=None)
    ap.add_argument("--timeout", type=int, default=240)
    a = ap.parse_args()
    if a.ref_gpu_one:
        n, w, h = a.ref_gpu_one
        ref_gpu(None if n == "-" else n, int(w), int(h))
        sys.exit(0)
    import pyrtb
    print("%-4s %-18s %-10s %9s %11s %11s %9s %10s %10s %11s | %13s %9s | %s" % (
        "cfg", "scene", "output", "tris", "primary", "shadow", "bounce", "frame ms", "Mrays/s", "render() ms",
        "ref GPU ms", "speed-up", "ref CPU rt wall (4 threads, 2x2 SSAA)"))
    for cfg, name, w, h in CONFIGS:
        hs = scene_of(name, w, h)
        ds = pyrtb.DeviceScene(hs, 3)
        ds.render(count=True)
        s = ds.stats()
        ts = []
        for _ in range(3 + a.frames):
            img = ds.render()
            ts.append(ds.stats()["frame_ms"])
        frame_ms = float(np.median(ts[3:]))
        ds.close()
        out = np.empty((h, w, 4), np.uint8)
        to = []
        for _ in range(4):
            t0 = time.perf_counter()
            pyrtb.render_oneshot(hs, 3, out=out)
            to.append(1e3 * (time.perf_counter() - t0))
        assert np.array_equal(out, img)
        rays = s["rays_primary"] + s["rays_shadow"]
        ref_ms = None
        if not a.skip_ref_gpu and name is not None:
            try:
                r = subprocess.run([sys.executable, __file__, "--ref-gpu-one", name, str(w), str(h)], capture_output=True,
                                   text=True, timeout=a.timeout)
                ln = [l for l in r.stdout.splitlines() if l.startswith("REFGPU")]
                ref_ms = float(ln[-1].split()[1]) if ln else None
            except subprocess.TimeoutExpired:
                ref_ms = float("inf")
        cpu = ""
        scaled = dict(x.split(":") for x in a.cpu_scaled.split(",") if x)
        if (cfg in a.cpu.split(",") or cfg in scaled) and name is not None:
            k = int(scaled.get(cfg, 1)) if cfg not in a.cpu.split(",") else 1
            with tempfile.TemporaryDirectory() as td:
                sv = os.path.join(td, "s.svati")
                if k > 1:
                    scene_of(name, w // k, h // k).write_svati(sv)
                else:
                    hs.write_svati(sv)

                def unlimited_stack():      # the CPU renderer's frame is a stack VLA (cpu/raytracer.c:91)
                    resource.setrlimit(resource.RLIMIT_STACK, (resource.RLIM_INFINITY, resource.RLIM_INFINITY))
                for exe in ("ref_cpu_rt", "ref_cpu_rt_O2"):
                    t0 = time.perf_counter()
                    try:
                        r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", exe), sv, os.path.join(td, "o.ppm")],
                                           capture_output=True, timeout=600, preexec_fn=unlimited_stack)
                        dt = time.perf_counter() - t0
                        if k > 1:
                            cpu += "%s %.0f s scaled (%.2f s at %dx%d x %d) (rc %d)  " % (exe, dt * k * k, dt, w // k, h // k, k * k, r.returncode)
                        else:
                            cpu += "%s %.2f s (rc %d)  " % (exe, dt, r.returncode)
                    except subprocess.TimeoutExpired:
                        cpu += "%s > 600 s  " % exe
        if cfg in a.oracle.split(",") and name is not None:
            import oracle_py
            t0 = time.perf_counter()
            oracle_py.orc_render(hs, 3, want_ids=False)
            cpu += "| oracle (ours, %d threads, GPU semantics 3x3) %.2f s" % (os.cpu_count() or 1, time.perf_counter() - t0)
        ref_txt = "-" if ref_ms is None else ("> %d s" % a.timeout if ref_ms == float("inf") else "%.2f" % ref_ms)
        sp = "-" if ref_ms in (None, float("inf")) else "%.1fx" % (ref_ms / min(to[1:]))
        print("%-4s %-18s %-10s %9d %11d %11d %9d %10.3f %10.1f %11.2f | %13s %9s | %s" % (
            cfg, name or "terrain 2237", "%dx%d" % (w, h), hs.triangle_count, s["rays_primary"], s["rays_shadow"],
            s["rays_bounce"], frame_ms, rays / frame_ms / 1e3, min(to[1:]), ref_txt, sp, cpu), flush=True)
