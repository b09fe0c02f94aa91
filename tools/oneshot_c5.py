"""GPU box: where the time of the one-shot render() goes on the C5 terrain (wall clock per phase).
usage: python tools/oneshot_c5.py [grid]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import pyrtb
g = int(sys.argv[1]) if len(sys.argv) > 1 else 2237
hs = pyrtb.HostScene.terrain(g, 3840, 2160)
out = np.zeros((hs.height, hs.width, 4), np.uint8)
for _ in range(3):
    pyrtb.render_oneshot(hs, 3, out=out)
t = []
for _ in range(5):
    t0 = time.perf_counter(); pyrtb.render_oneshot(hs, 3, out=out); t.append(time.perf_counter() - t0)
print("oneshot ms:", " ".join("%.2f" % (1e3 * x) for x in t))
for rep in range(3):
    t0 = time.perf_counter(); ds = pyrtb.DeviceScene(hs, 3, flags=8); t_create = time.perf_counter() - t0
    st = ds.stats()
    t0 = time.perf_counter(); ds.render(out=out); t_r1 = time.perf_counter() - t0
    st2 = ds.stats()
    t0 = time.perf_counter(); ds.close(); t_close = time.perf_counter() - t0
    print("create %.2f (upload %.2f build %.2f dev-ms) render %.2f (kernels %.3f) destroy %.2f  sum %.2f" % (
        1e3 * t_create, st["upload_ms"], st["build_ms"], 1e3 * t_r1, st2["frame_ms"], 1e3 * t_close, 1e3 * (t_create + t_r1 + t_close)), flush=True)
