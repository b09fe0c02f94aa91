"""Quick iteration probe (GPU box): frame-kernel time on the C5 terrain + frame checksum.
usage: python tools/perf_frame.py [--grid 2237] [--frames 5] [--k 24,8,2] [--ncu]"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
import pyrtb  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--grid", type=int, default=2237)
ap.add_argument("--frames", type=int, default=5)
ap.add_argument("--k", default="")
ap.add_argument("--w", type=int, default=3840)
ap.add_argument("--h", type=int, default=2160)
ap.add_argument("--count", action="store_true")
ap.add_argument("--wave", action="store_true")
ap.add_argument("--nosched", action="store_true")
ap.add_argument("--flags", type=int, default=0, help="rtb_scene_create flags (2: cubic keys only)")
ap.add_argument("--primary", type=int, default=0, help="0 auto, 1 tree walk, 2 projection")
ap.add_argument("--scene", default=None, help="golden scene name instead of the terrain (rendered at --w x --h)")
a = ap.parse_args()
pyrtb.lib().rtb_set_schedule(0 if a.nosched else 1)
if a.scene:
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from util import load_golden
    hs = load_golden(a.scene)[0].with_resolution(a.w, a.h)
else:
    hs = pyrtb.HostScene.terrain(a.grid, a.w, a.h)
ds = pyrtb.DeviceScene(hs, 3, flags=a.flags)
ds.frame_kernel = int(a.wave)
ds.primary = a.primary
ds.profile(True)
st = ds.stats()
print("build %.2f ms upload %.2f ms nodes %d depth %d  builder %d key mode %d (sample-tree cost of modes 1,2,3 vs 0: %s)" %
      (st["build_ms"], st["upload_ms"], st["nodes"], st["max_depth"], st["builder"], st["key_mode"],
       [((st["mode_cost"] >> (16 * k)) & 0xffff) / 1000 for k in range(3)]))
ks = [float(x) for x in a.k.split(",")] if a.k else [None]
for k in ks:
    if k is not None:
        pyrtb.lib().rtb_set_margin_k(k)
    img = ds.render()
    t, km = [], []
    for _ in range(a.frames):
        img = ds.render()
        t.append(ds.stats()["frame_ms"])
        km.append(ds.kernel_ms())
    line = "K=%s frame ms: min %.2f median %.2f  [setup %.2f items %.2f frame %.2f]  checksum %d" % (
        k, min(t), float(np.median(t)), *np.median(np.array(km), axis=0)[:3], int(img.astype(np.uint64).sum()))
    if a.count:
        ds.render(count=True)
        s = ds.stats()
        rays = s["rays_primary"] + s["rays_shadow"] + s["rays_bounce"]
        line += "  rays %d nodes/ray %.2f tris/ray %.2f" % (rays, s["node_visits"] / rays, s["tri_tests"] / rays)
    print(line, flush=True)
    print("raster state", ds.raster_state(), "launches", ds.stats()["kernel_launches"], flush=True)
ds.close()
