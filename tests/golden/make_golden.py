"""Generates tests/golden/<scene>.npz for the reference's 20 shipped scenes.

Run in the build container (needs /root/reference):  python tests/golden/make_golden.py [scene ...]

Each fixture holds
  * the scene in memory order (vertex, normal, tri_counts, materials, lights, camera) as the
    reference's own parser reads it (gpu/parser.cpp, parse_obj.cpp via oracle/_ref/libref_parser.so;
    tests/test_parser_pin.py re-checks the committed arrays against it), so the GPU box needs neither
    /root/reference nor the .svati text;
  * rgba: the full-resolution frame (aliasing 3) of the REFERENCE's own device sources
    compiled for the host (oracle/_ref/libref_gpusem.so via oracle/ref_driver.cpp);
  * tri/obj: per-sample primary-ray hit ids from oracle/liboracle.so, accepted only after
    (a) the oracle's frame is byte-identical to the reference's frame and (b) the oracle's
    per-sample hit records (point + normal) are bit-identical to the reference's collide();
  * counts: primary / shadow / bounce ray counts of the frame (SURVEY 8d accounting).
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_py  # noqa: E402
import pyrtb  # noqa: E402

SCENES = os.path.join(oracle_py.REFERENCE, "tests")


def main(argv):
    names = argv or sorted(f[:-6] for f in os.listdir(SCENES) if f.endswith(".svati"))
    for name in names:
        out = os.path.join(HERE, name + ".npz")
        if os.path.exists(out) and not os.environ.get("GOLDEN_FORCE"):
            print(name, "exists, skipped")
            continue
        # scene arrays as the REFERENCE's own loader reads them (oracle/_ref/libref_parser.so), not ours
        from test_parser_pin import ref_parse
        ra = ref_parse(os.path.join(SCENES, name + ".svati"))
        ra["lights"][ra["lights"][:, 0] == pyrtb.AMBIENT, 4:7] = 0      # never set by parse_a_light, never read
        hs = pyrtb.HostScene.from_arrays(ra["vertex"], ra["normal"], ra["tri_counts"], ra["materials"], ra["lights"], ra["camera"])
        t0 = time.time()
        o = oracle_py.orc_render(hs, 3, want_ids=True)
        t1 = time.time()
        ref_rgba, _ = oracle_py.ref_render(hs, 3)
        t2 = time.time()
        if not np.array_equal(ref_rgba, o["rgba"]):
            raise SystemExit("%s: oracle frame differs from the reference frame (%d bytes)" %
                             (name, int((ref_rgba != o["rgba"]).sum())))
        hit_o, tri_o = oracle_py.orc_primary_hits(hs, 3)
        hit_r = oracle_py.ref_primary_hits(hs, 3)
        if not np.array_equal(hit_o.view(np.uint32), hit_r.view(np.uint32)):
            raise SystemExit("%s: oracle hit records differ from the reference's collide()" % name)
        assert np.array_equal(tri_o, o["tri"])
        a = hs.arrays()
        T = hs.triangle_count
        tri = o["tri"].astype(np.int16 if T < 32000 else np.int32)
        obj = o["obj"].astype(np.int8 if len(a["tri_counts"]) < 127 else np.int16)
        np.savez_compressed(out, rgba=ref_rgba, tri=tri, obj=obj, counts=o["counts"], **a)
        print("%-24s %4dx%-4d %5d tris  oracle %.1fs  reference %.1fs  %.0f KB" %
              (name, hs.width, hs.height, T, t1 - t0, t2 - t1, os.path.getsize(out) / 1024), flush=True)


if __name__ == "__main__":
    main(sys.argv[1:])
