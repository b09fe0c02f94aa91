"""GPU box: the one-shot render() of C5 N times in a row with a resident scene beside it (as in bench.py): per-call wall
clock, to see stalled calls.  RTB_TRACE=1 prints the phases of every call, RTB_TRACE=2 the host time inside render_async.
usage: python tools/oneshot_calls.py [calls]"""
import sys, time, numpy as np
sys.path.insert(0, "raytracing-gpu_b200")
import pyrtb
hs = pyrtb.HostScene.terrain(2237, 3840, 2160)
pinned = pyrtb.PinnedBuffer(3840 * 2160 * 4)
img = pinned.array((2160, 3840, 4))
ds = pyrtb.DeviceScene(hs, 3)
for _ in range(5):
    ds.render(out=img)
ts = []
for k in range(int(sys.argv[1]) if len(sys.argv) > 1 else 30):
    t0 = time.perf_counter()
    pyrtb.render_oneshot(hs, 3, out=img)
    ts.append(1e3 * (time.perf_counter() - t0))
    print("call %d: %.2f ms" % (k, ts[-1]), file=sys.stderr, flush=True)
ts = np.array(ts[2:])
print("one-shot with a resident scene beside it: median %.2f mean %.2f max %.2f ms; calls over 70 ms: %s" % (np.median(ts), ts.mean(), ts.max(), np.round(ts[ts > 70], 1)))
