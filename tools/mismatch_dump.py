"""GPU box: dump the samples where tree and brute-force primary ids differ (scene at WxH)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pyrtb
from util import load_golden
name, w, h = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
hs = load_golden(name)[0].with_resolution(w, h)
ds = pyrtb.DeviceScene(hs, 3)
t, _ = ds.hit_ids(pyrtb.TRACER_TREE)
b, _ = ds.hit_ids(pyrtb.TRACER_BRUTE)
yx = np.argwhere(t != b)
out = np.concatenate([yx[:, ::-1], t[yx[:, 0], yx[:, 1]][:, None], b[yx[:, 0], yx[:, 1]][:, None]], axis=1)
np.save(os.path.join(ROOT, "gpurun_out", "mismatch_%s.npy" % name), out)
print(name, len(out))
