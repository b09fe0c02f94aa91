"""GPU box: effect of the sliver-padding policy (A/B builds, tools/build_variant.sh) on
exactness and speed.  Per library: primary-id mismatches against the golden fixtures (native
resolution, all scenes), and for the needle-heavy scenes at 3840x2160: frame ms plus
primary-id mismatches against the brute-force tracer (same arithmetic, no culling).
usage: RTB200_LIB=tools/variants/X.so python tools/pad_study.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pyrtb  # noqa: E402
from util import golden_names, load_golden  # noqa: E402

tot = 0
bad_scenes = []
for name in golden_names():
    hs, z = load_golden(name)
    ds = pyrtb.DeviceScene(hs, 3)
    tri, _ = ds.hit_ids(pyrtb.TRACER_TREE)
    n = int((tri != z["tri"].astype(np.int32)).sum())
    tot += n
    if n:
        bad_scenes.append("%s:%d" % (name, n))
    ds.close()
line = "%-28s golden id mismatches %d %s |" % (os.path.basename(os.environ.get("RTB200_LIB", "default")), tot, ",".join(bad_scenes))
for name in ("car-on-road", "car", "spheres", "secret"):
    hs = load_golden(name)[0].with_resolution(3840, 2160)
    ds = pyrtb.DeviceScene(hs, 3)
    ds.render()
    t = []
    for _ in range(3):
        ds.render()
        t.append(ds.stats()["frame_ms"])
    tri, _ = ds.hit_ids(pyrtb.TRACER_TREE)
    cache = os.path.join(ROOT, "gpurun_out", "brute_%s.npy" % name)
    if os.path.exists(cache):
        bt = np.load(cache)
    else:
        bt, _ = ds.hit_ids(pyrtb.TRACER_BRUTE)
        np.save(cache, bt)
    line += " %s %.2f ms, %d/%d ids off |" % (name, min(t), int((tri != bt).sum()), tri.size)
    ds.close()
print(line, flush=True)
