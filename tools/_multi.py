import sys, time, numpy as np
sys.path.insert(0, "raytracing-gpu_b200")
import pyrtb
n = int(sys.argv[1])
hs = pyrtb.HostScene.terrain(2237, 3840, 2160)
pinned = pyrtb.PinnedBuffer(3840 * 2160 * 4)
img = pinned.array((2160, 3840, 4))
for k in range(5):
    t0 = time.perf_counter()
    pyrtb.render_multi(hs, n, 3, out=img)
    print("render_multi(%d) %.2f ms" % (n, 1e3 * (time.perf_counter() - t0)), file=sys.stderr, flush=True)
