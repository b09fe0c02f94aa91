"""GPU box: where the time of the one-shot call goes on small scenes."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import pyrtb
from util import load_golden
for name in sys.argv[1:] or ["cube", "car", "spheres", "susans_smooth"]:
    hs, z = load_golden(name)
    out = np.zeros((hs.height, hs.width, 4), np.uint8)
    pyrtb.render_oneshot(hs, 3, out=out)
    t = []
    for _ in range(3):
        t0 = time.perf_counter(); pyrtb.render_oneshot(hs, 3, out=out); t.append(time.perf_counter() - t0)
    t0 = time.perf_counter(); ds = pyrtb.DeviceScene(hs, 3); t_create = time.perf_counter() - t0
    st = ds.stats()
    t0 = time.perf_counter(); ds.render(out=out); t_r1 = time.perf_counter() - t0
    t0 = time.perf_counter(); ds.render(out=out); t_r2 = time.perf_counter() - t0
    t0 = time.perf_counter(); ds.close(); t_close = time.perf_counter() - t0
    print("%-16s oneshot %.2f ms | create %.2f (upload %.2f build %.2f dev-ms, depth %d nodes %d) render1 %.2f render2 %.2f (kernel %.3f) destroy %.2f" % (
        name, 1e3 * min(t), 1e3 * t_create, st["upload_ms"], st["build_ms"], st["max_depth"], st["nodes"], 1e3 * t_r1, 1e3 * t_r2, st["frame_ms"], 1e3 * t_close), flush=True)
