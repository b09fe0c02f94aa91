import sys, time, numpy as np
sys.path.insert(0, "raytracing-gpu_b200")
import pyrtb
hs = pyrtb.HostScene.terrain(2237, 3840, 2160)
pinned = pyrtb.PinnedBuffer(3840 * 2160 * 4)
img = pinned.array((2160, 3840, 4))
for k in range(6):
    t0 = time.perf_counter()
    pyrtb.render_oneshot(hs, 3, out=img)
    print("one-shot %.2f ms" % (1e3 * (time.perf_counter() - t0)), flush=True)
