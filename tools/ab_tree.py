"""A/B probe (GPU box): tree statistics and counted visits of one scene through the library named by RTB200_LIB."""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pyrtb  # noqa: E402
from util import load_golden  # noqa: E402

name, w, h = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
hs = load_golden(name)[0].with_resolution(w, h)
ds = pyrtb.DeviceScene(hs, 3)
ds.primary = 1
nodes, perm, root = ds.dump_tree()
ds.render(count=True)
st = ds.stats()
rays = st["rays_primary"] + st["rays_shadow"] + st["rays_bounce"]
b = nodes[:, :12]
fin = np.isfinite(b)
print(os.environ.get("RTB200_LIB", "tree"), "nodes", nodes.shape, "root", root, "md5 nodes", hashlib.md5(nodes.tobytes()).hexdigest()[:10],
      "md5 perm", hashlib.md5(perm.tobytes()).hexdigest()[:10], "box extent sum %.6f" % float(np.abs(b[fin]).sum()),
      "rays", rays, "nodes/ray %.3f tris/ray %.3f" % (st["node_visits"] / rays, st["tri_tests"] / rays), "depth", st["max_depth"], flush=True)
ds.close()
