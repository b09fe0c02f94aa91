"""GPU box: per golden scene, mismatching primary ids (tree and brute vs fixture) and RGB report."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pyrtb  # noqa: E402
from util import golden_names, load_golden, rgb_report  # noqa: E402

names = sys.argv[1:] or golden_names()
for name in names:
    hs, z = load_golden(name)
    ds = pyrtb.DeviceScene(hs, 3)
    tri, obj = ds.hit_ids(pyrtb.TRACER_TREE)
    gt = z["tri"].astype(np.int32)
    bad = np.argwhere(tri != gt)
    img = ds.render()
    ms = ds.stats()["frame_ms"]
    ds.frame_kernel = 1
    img2 = ds.render()
    ms2 = ds.stats()["frame_ms"]
    rep = rgb_report(img, z["rgba"])
    line = "%-24s tree id mismatches %6d  rgb(>1 off %.6f, identical %.5f, max %d)  per-sample %.3f ms wave %.3f ms same=%s" % (
        name, len(bad), rep[0], rep[1], rep[2], ms, ms2, np.array_equal(img, img2))
    if len(bad):
        bt, _ = ds.hit_ids(pyrtb.TRACER_BRUTE)
        line += "  brute mismatches %d" % int((bt != gt).sum())
        for (y, x) in bad[:4]:
            line += "\n      sample (%d,%d): gpu %d golden %d" % (x, y, tri[y, x], gt[y, x])
    print(line, flush=True)
    ds.close()
