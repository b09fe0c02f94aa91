"""GPU box: device time of upload and hierarchy build for the C5 terrain, 6 scene creations."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
import pyrtb
hs = pyrtb.HostScene.terrain(int(sys.argv[1]) if len(sys.argv) > 1 else 2237, 3840, 2160)
for i in range(6):
    ds = pyrtb.DeviceScene(hs, 3)
    st = ds.stats()
    print("create %d: upload %.2f ms build %.2f ms launches %d nodes %d depth %d" % (i, st["upload_ms"], st["build_ms"], st["build_launches"], st["nodes"], st["max_depth"]), flush=True)
    if i == 5:
        img = ds.render()
        print("checksum", int(img.astype("uint64").sum()))
    ds.close()
