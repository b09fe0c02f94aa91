"""GPU box with N devices: wall time of the one-shot render() on 1..N GPUs of one process
(rtb_render / rtb_render_multi, host buffers in, host frame out) on the C5 terrain.
usage: python tools/multi_oneshot.py [--grid 2237]"""
import argparse, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
import pyrtb
ap = argparse.ArgumentParser()
ap.add_argument("--grid", type=int, default=2237)
a = ap.parse_args()
hs = pyrtb.HostScene.terrain(a.grid, 3840, 2160)
n = pyrtb.device_count()
pin = pyrtb.PinnedBuffer(3840 * 2160 * 4)
out = pin.array((2160, 3840, 4))
want = pyrtb.render_oneshot(hs, 3).copy()
for k in [x for x in (1, 2, 4, 8) if x <= n]:
    t = []
    for _ in range(5):
        t0 = time.perf_counter()
        if k == 1:
            pyrtb.render_oneshot(hs, 3, out=out)
        else:
            pyrtb.render_multi(hs, k, 3, out=out)
        t.append(1e3 * (time.perf_counter() - t0))
    print("%d GPU(s): render() wall ms %s  min %.2f  same bytes as 1 GPU: %s" % (
        k, " ".join("%.1f" % x for x in t), min(t[1:]), np.array_equal(out, want)), flush=True)
