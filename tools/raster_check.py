"""GPU box: primary hit ids of the projection pass against the tree walk on the C5 frame, and
both against exhaustive search on the samples where they differ.
usage: python tools/raster_check.py [--grid 2237] [--w 3840] [--h 2160]"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pyrtb  # noqa: E402
from util import primary_rays  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--grid", type=int, default=2237)
ap.add_argument("--w", type=int, default=3840)
ap.add_argument("--h", type=int, default=2160)
a = ap.parse_args()
hs = pyrtb.HostScene.terrain(a.grid, a.w, a.h)
ds = pyrtb.DeviceScene(hs, 3)
ds.primary = 1
tree, _ = ds.hit_ids(pyrtb.TRACER_TREE, want_obj=False)
ds.primary = 2
rast, _ = ds.hit_ids(pyrtb.TRACER_TREE, want_obj=False)
print("raster state", ds.raster_state())
diff = np.argwhere(tree != rast)
print("samples %d, tree != raster: %d" % (tree.size, len(diff)))
rng = np.random.default_rng(3)
extra = np.stack([rng.integers(0, tree.shape[0], 2000), rng.integers(0, tree.shape[1], 2000)], axis=1)
pts = np.concatenate([diff, extra]) if len(diff) else extra
rays = primary_rays(hs, 3, pts[:, ::-1])
bt, _ = ds.trace_rays(rays, pyrtb.TRACER_BRUTE)
t = tree[pts[:, 0], pts[:, 1]]
r = rast[pts[:, 0], pts[:, 1]]
nd = len(diff)
print("on the %d differing samples: tree == brute %d, raster == brute %d" % (nd, int((t[:nd] == bt[:nd]).sum()), int((r[:nd] == bt[:nd]).sum())))
print("on 2000 random samples: tree == brute %d, raster == brute %d" % (int((t[nd:] == bt[nd:]).sum()), int((r[nd:] == bt[nd:]).sum())))
# frames: projection pass + light grids (default) against tree walks for every ray
ds.primary = 0
full = ds.render()
ds.close()
ds2 = pyrtb.DeviceScene(hs, 3, flags=16)      # RTB_BUILD_NO_LIGHT_GRIDS
ds2.primary = 1
ref = ds2.render()
ds2.close()
d = (full != ref).any(axis=-1)
print("frame pixels that differ between (projection + grids) and (tree walk for every ray): %d of %d" % (int(d.sum()), d.size))
