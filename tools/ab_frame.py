"""A/B probe (GPU box): frame time of one scene through the library named by RTB200_LIB (default: the tree's).
usage: [RTB200_LIB=tools/variants/x/librtb200.so] python tools/ab_frame.py scene w h [primary] [frames]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pyrtb  # noqa: E402
from util import load_golden  # noqa: E402

name, w, h = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
primary = int(sys.argv[4]) if len(sys.argv) > 4 else 0
frames = int(sys.argv[5]) if len(sys.argv) > 5 else 15
hs = pyrtb.HostScene.terrain(int(name[7:]), w, h) if name.startswith("terrain") else load_golden(name)[0].with_resolution(w, h)
ds = pyrtb.DeviceScene(hs, 3)
ds.primary = primary
t = []
for _ in range(frames):
    img = ds.render()
    t.append(ds.stats()["frame_ms"])
print("%s %s %dx%d primary=%d: frame ms min %.3f median %.3f last %.3f  checksum %d  nodes %d" % (
    os.environ.get("RTB200_LIB", "tree"), name, w, h, primary, min(t), float(np.median(t)), t[-1], int(img.astype(np.uint64).sum()),
    ds.stats()["nodes"]), flush=True)
ds.close()
