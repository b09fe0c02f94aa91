"""GPU box: random triangle soups (needles included) — tree vs brute-force primary ids, and
for every mismatch the incidence of the ray on the triangle the tree missed.
usage: python tools/random_scene_probe.py [n_scenes] [tris]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pyrtb
from util import primary_rays, random_scene

n_scenes = int(sys.argv[1]) if len(sys.argv) > 1 else 20
T = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
tot = bad = 0
for seed in range(n_scenes):
    hs = random_scene(seed, T)
    ds = pyrtb.DeviceScene(hs, 3)
    t, _ = ds.hit_ids(pyrtb.TRACER_TREE)
    b, _ = ds.hit_ids(pyrtb.TRACER_BRUTE)
    hit = float((b >= 0).mean())
    yx = np.argwhere(t != b)
    tot += t.size; bad += len(yx)
    line = "seed %2d: %d of %d ids differ (hit fraction %.2f)" % (seed, len(yx), t.size, hit)
    if len(yx):
        V = hs.arrays()["vertex"].astype(np.float64)
        R = primary_rays(hs, 3, yx[:, ::-1]).astype(np.float64)
        cosines = []
        for (y, x), r in zip(yx, R):
            tid = b[y, x] if b[y, x] >= 0 else t[y, x]
            a_, b_, c_ = V[tid]
            nn = np.cross(b_ - a_, c_ - a_); nn /= max(np.linalg.norm(nn), 1e-300)
            cosines.append(abs(nn @ r[3:]) / np.linalg.norm(r[3:]))
        line += "  |cos incidence| of the missed triangles: " + " ".join("%.4f" % c for c in cosines[:12])
        line += "  samples(x,y,tree,brute): " + " ".join("(%d,%d,%d,%d)" % (x, y, t[y, x], b[y, x]) for y, x in yx[:6])
    print(line, flush=True)
    ds.close()
print("total %d of %d (%.2e)" % (bad, tot, bad / max(tot, 1)))
