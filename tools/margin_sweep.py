"""GPU box: how many primary-ray hit ids / output pixels change with the box-inflation
factor K (reference = a very large K).  usage: python tools/margin_sweep.py [--grid 2237]"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
import pyrtb  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--grid", type=int, default=2237)
ap.add_argument("--w", type=int, default=3840)
ap.add_argument("--h", type=int, default=2160)
ap.add_argument("--ks", default="4000,1000,250,96,48,24,8,2")
a = ap.parse_args()
hs = pyrtb.HostScene.terrain(a.grid, a.w, a.h)
ds = pyrtb.DeviceScene(hs, 3)
ref_tri = ref_img = None
for k in [float(x) for x in a.ks.split(",")]:
    pyrtb.lib().rtb_set_margin_k(k)
    tri, _ = ds.hit_ids(want_obj=False)
    img = ds.render()
    ms = ds.stats()["frame_ms"]
    if ref_tri is None:
        ref_tri, ref_img = tri, img
    print("K=%-6g frame %.1f ms  ids differing from K=max: %d  pixels differing: %d" %
          (k, ms, int((tri != ref_tri).sum()), int((img != ref_img).any(axis=-1).sum())), flush=True)
ds.close()
