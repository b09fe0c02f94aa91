"""GPU box, one device: rank 0's share of an N-rank C5 frame, four times (for an ncu launch list of what one rank runs
with the replicated projection pass).  usage: ncu ... python tools/one_rank.py N"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
import pyrtb  # noqa: E402

N = int(sys.argv[1])
hs = pyrtb.HostScene.terrain(2237, 3840, 2160)
ds = pyrtb.DeviceScene(hs, 3)
out = np.zeros((2160, 3840, 4), np.uint8)
for _ in range(4):
    ds.render(rank=0, nranks=N, out=out)
ds.close()
