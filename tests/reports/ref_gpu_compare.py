"""GPU box: the reference's own GPU renderer (oracle/_ref/libref_gpu.so, unmodified gpu/*.cu
built for sm_100) next to librtb200 on the shipped scenes: wall time of the drop-in call
render() (upload + partitioning build + frame + download, what the reference API offers) and
pixel differences of both against the un-fused semantic definition (tests/golden).
usage: python tests/reports/ref_gpu_compare.py [--scale S] [scene ...]     (one subprocess per scene, with a timeout)"""
import argparse
import ctypes as C
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))  # util.py


def one(name, scale):
    import pyrtb
    from util import load_golden, rgb_report
    hs, z = load_golden(name)
    golden = z["rgba"] if scale == 1 else None
    if scale != 1:
        hs.with_resolution(hs.width * scale, hs.height * scale)
    w, h = hs.width, hs.height
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_gpu.so"))
    L.refgpu_render.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_long, C.c_int, C.c_int]
    s = hs.scaled(3)
    ref = np.zeros((h, w, 4), np.uint8)
    tr = []
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    for _ in range(3):
        os.dup2(devnull, 1)           # the reference prints two lines per call
        t0 = time.perf_counter()
        L.refgpu_render(C.byref(s), ref.ctypes.data, 3, w * 4, h, w)
        tr.append(time.perf_counter() - t0)
        os.dup2(saved, 1)
    ours = np.zeros((h, w, 4), np.uint8)
    to = []
    for _ in range(3):
        t0 = time.perf_counter()
        pyrtb.render_oneshot(hs, 3, out=ours)
        to.append(time.perf_counter() - t0)
    ds = pyrtb.DeviceScene(hs, 3)
    ds.render()
    ds.render()
    frame_ms = ds.stats()["frame_ms"]
    ds.close()
    line = "%-22s %5dx%-5d ref_gpu render() %9.2f ms | rtb_render %8.2f ms (frame kernel %.3f ms) | speed-up %7.1fx" % (
        name, w, h, 1e3 * min(tr[1:]), 1e3 * min(to[1:]), frame_ms, min(tr[1:]) / min(to[1:]))
    d = np.abs(ref[..., :3].astype(int) - ours[..., :3].astype(int)).max(-1)
    line += " | ref_gpu vs ours: %d px differ (%d by >1)" % (int((d > 0).sum()), int((d > 1).sum()))
    if golden is not None:
        dg = np.abs(ref[..., :3].astype(int) - golden[..., :3].astype(int)).max(-1)
        do = np.abs(ours[..., :3].astype(int) - golden[..., :3].astype(int)).max(-1)
        line += " | vs definition: ref_gpu %d px, ours %d px" % (int((dg > 0).sum()), int((do > 0).sum()))
    print(line, flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=int, default=1)
    ap.add_argument("--one", default=None)
    ap.add_argument("--timeout", type=int, default=120)
    ap.add_argument("scenes", nargs="*")
    a = ap.parse_args()
    if a.one:
        one(a.one, a.scale)
        sys.exit(0)
    from util import golden_names
    for name in (a.scenes or golden_names()):
        try:
            r = subprocess.run([sys.executable, __file__, "--one", name, "--scale", str(a.scale)], capture_output=True,
                               text=True, timeout=a.timeout)
            out = [l for l in r.stdout.splitlines() if l.startswith(name)]
            print(out[-1] if out else "%-22s FAILED rc=%d %s" % (name, r.returncode, r.stderr.strip()[-300:]), flush=True)
        except subprocess.TimeoutExpired:
            print("%-22s reference GPU renderer did not finish within %d s" % (name, a.timeout), flush=True)
