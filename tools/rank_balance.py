"""GPU box, one device: frame time of each rank's share of the C5 frame for N = 2, 4, 8
(ranks rendered one after the other), to separate rank imbalance from per-GPU tail effects.
usage: python tools/rank_balance.py [--grid 2237]"""
import argparse, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
import pyrtb
ap = argparse.ArgumentParser()
ap.add_argument("--grid", type=int, default=2237)
ap.add_argument("--nosched", action="store_true")
a = ap.parse_args()
pyrtb.lib().rtb_set_schedule(0 if a.nosched else 1)
hs = pyrtb.HostScene.terrain(a.grid, 3840, 2160)
ds = pyrtb.DeviceScene(hs, 3)
for _ in range(4):
    ds.render()
full = ds.stats()["frame_ms"]
print("N=1 %.3f ms" % full)
for n in (2, 4, 8):
    t = []
    for r in range(n):
        for _ in range(4):
            ds.render(rank=r, nranks=n)
        t.append(ds.stats()["frame_ms"])
    print("N=%d per-rank ms: %s  max %.3f  mean %.3f  ideal %.3f  -> speed-up %.2fx (imbalance %.1f%%, overhead vs ideal %.1f%%)" % (
        n, " ".join("%.3f" % x for x in t), max(t), np.mean(t), full / n, full / max(t), 100 * (max(t) / np.mean(t) - 1),
        100 * (np.mean(t) / (full / n) - 1)), flush=True)
ds.close()
