"""GPU box, one device: every rank's share of a multi-rank C5 frame rendered alone (replicated projection
pass), per-kernel times per rank.  Tells imbalance between ranks from what a rank loses by itself.
usage: python tools/rank_balance.py N [frames]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
import pyrtb  # noqa: E402

N = int(sys.argv[1])
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 5
hs = pyrtb.HostScene.terrain(2237, 3840, 2160)
ds = pyrtb.DeviceScene(hs, 3)
ds.profile(True)
out = np.zeros((2160, 3840, 4), np.uint8)
for _ in range(3):
    ds.render(out=out)
full = np.median([(ds.render(out=out), ds.kernel_ms())[1] for _ in range(frames)], axis=0)
print("1 rank: setup %.2f items %.2f frame %.2f total %.2f  [setup kernel %.2f big %.2f sync %.2f]" % tuple(full[:7]))
rows = []
for r in range(N):
    for _ in range(2):
        ds.render(rank=r, nranks=N, out=out)
    km = np.median([(ds.render(rank=r, nranks=N, out=out), ds.kernel_ms())[1] for _ in range(frames)], axis=0)
    rows.append(km)
    print("rank %d of %d: setup %.2f items %.2f frame %.2f total %.2f  [setup kernel %.2f big %.2f]" % (r, N, *km[:6]), flush=True)
rows = np.array(rows)
print("k_frame: sum over ranks %.2f (1 rank %.2f), max %.2f, mean %.2f; k_rast_items: sum %.2f (1 rank %.2f), max %.2f" %
      (rows[:, 2].sum(), full[2], rows[:, 2].max(), rows[:, 2].mean(), rows[:, 1].sum(), full[1], rows[:, 1].max()))
ds.close()
