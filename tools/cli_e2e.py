"""GPU box: end-to-end wall time of the CLI, `rt file.svati out.png` (parse + upload
+ build + frame + PNG encode), on the C5 terrain written out as a .svati file and on C4.
usage: python tools/cli_e2e.py [--grid 2237]"""
import argparse, json, os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pyrtb
from util import load_golden
ap = argparse.ArgumentParser()
ap.add_argument("--grid", type=int, default=2237)
a = ap.parse_args()
rt = os.path.join(ROOT, "raytracing-gpu_b200", "rt")
with tempfile.TemporaryDirectory() as td:
    cases = [("C4 car-on-road 3840x2160", load_golden("car-on-road")[0].with_resolution(3840, 2160)),
             ("C5 terrain grid %d 3840x2160" % a.grid, pyrtb.HostScene.terrain(a.grid, 3840, 2160))]
    for name, hs in cases:
        sv, png = os.path.join(td, "s.svati"), os.path.join(td, "o.png")
        t0 = time.perf_counter()
        hs.write_svati(sv)
        tw = time.perf_counter() - t0
        for run in range(2):
            t0 = time.perf_counter()
            r = subprocess.run([rt, sv, png], capture_output=True, text=True, env=dict(os.environ, RT_STATS="1"))
            wall = time.perf_counter() - t0
            st = [l for l in r.stderr.splitlines() if l.startswith("{")]
            print("%-34s run %d: wall %.3f s rc %d  %s  (.svati %.0f MB written in %.1f s, png %.1f MB)" % (
                name, run, wall, r.returncode, st[-1] if st else r.stderr[-200:], os.path.getsize(sv) / 1e6, tw,
                os.path.getsize(png) / 1e6 if os.path.exists(png) else 0), flush=True)
