"""GPU box: node visits / triangle tests per ray by ray type (lights removed one at a time)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
import pyrtb  # noqa: E402

hs = pyrtb.HostScene.terrain(2237, 3840, 2160)
full = hs.s.light_count
prev = None
for nl, label in ((1, "primary only (ambient)"), (2, "+ directional shadow"), (3, "+ point shadow")):
    hs.s.light_count = nl
    ds = pyrtb.DeviceScene(hs, 3, flags=int(sys.argv[1]) if len(sys.argv) > 1 else 0)
    ds.primary = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    ds.render()
    ds.render()
    ms = ds.stats()["frame_ms"]
    ds.render(count=True)
    s = ds.stats()
    rays = s["rays_primary"] + s["rays_shadow"]
    cur = (rays, s["node_visits"], s["tri_tests"], ms)
    if prev is None:
        d = cur
    else:
        d = tuple(c - p for c, p in zip(cur, prev))
    print(ds.raster_state())
    print("%-26s rays %10d  nodes/ray %.1f  tris/ray %.2f  frame %.2f ms (+%.2f ms, %.0f Mrays/s incremental)" %
          (label, d[0], d[1] / d[0], d[2] / d[0], ms, d[3], d[0] / d[3] / 1e3), flush=True)
    prev = cur
    ds.close()
hs.s.light_count = full
