"""The C-ABI library loads without a GPU, exports every symbol the headers declare, keeps the
reference's struct layouts, and refuses to compute when no device is present (no fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import pyrtb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for hdr in ("rtb200.h", "rt_host.h"):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names |= set(re.findall(r"\b(rtb_[a-z0-9_]+)\s*\(", text))
    return names - {"rtb_scene_handle"}


def test_library_exports_every_declared_symbol():
    L = pyrtb.lib()
    declared = declared_symbols()
    assert len(declared) >= 35
    missing = [s for s in sorted(declared) if not hasattr(L, s)]
    assert not missing, "declared in include/ but not exported: %s" % missing
    assert set(pyrtb.ABI_SYMBOLS) <= declared | {"rtb_cli_main"}


def test_reference_host_surface_is_exported_with_cxx_linkage():
    # parser / render / scene API keep the reference's C++ signatures (gpu/headers/*.h)
    import subprocess
    out = subprocess.run(["nm", "-DC", pyrtb.LIB_PATH], capture_output=True, text=True).stdout
    for sig in ("parser(char const*)", "parse_object(_IO_FILE*, scene*)", "create_vec(_IO_FILE*)", "empty_scene()",
                "add_object_to_scene(scene*, unsigned int)", "get_vertex(triangles_layout, unsigned int)",
                "get_normal(triangles_layout, unsigned int)", "open_output(char const*, int, int)",
                "print_color(color, _IO_FILE*)", "render(scene const&, char*, int, long, int, int)"):
        assert sig in out, sig


def test_struct_layout_matches_reference_probe():
    # sizes/offsets of the reference headers (LAYOUT_SOA + PARTITIONING_OCTREE)
    assert C.sizeof(pyrtb.Vec3) == 12
    assert C.sizeof(pyrtb.Light) == 28 and pyrtb.Light.v.offset == 16
    assert C.sizeof(pyrtb.Camera) == 48 and pyrtb.Camera.position.offset == 8 and pyrtb.Camera.fov.offset == 44
    O = pyrtb.Object
    assert C.sizeof(O) == 72 and O.triangle_count.offset == 16 and O.ka.offset == 20 and O.kd.offset == 32
    assert O.ks.offset == 44 and O.ns.offset == 56 and O.nr.offset == 64 and O.d.offset == 68
    S = pyrtb.SceneStruct
    assert C.sizeof(S) == 120 and S.objects.offset == 16 and S.lights.offset == 32 and S.camera.offset == 48
    assert S.triangle_count.offset == 96 and S.octree.offset == 112


@pytest.mark.skipif(pyrtb.device_count() > 0, reason="this test describes a box without a GPU")
def test_no_gpu_means_loud_failure_not_fallback():
    hs = pyrtb.HostScene.terrain(8, 16, 16)
    with pytest.raises(pyrtb.RtbError) as e:
        pyrtb.DeviceScene(hs, 3)
    assert e.value.code == pyrtb.RTB_ERR_NOGPU and "no CPU fallback" in str(e.value)
    with pytest.raises(pyrtb.RtbError):
        pyrtb.render_oneshot(hs, 3)
    with pytest.raises(pyrtb.RtbError):
        pyrtb.radix_sort_pairs(np.arange(10, dtype=np.uint32))
    with pytest.raises(pyrtb.RtbError):
        pyrtb.inclusive_scan(np.arange(10, dtype=np.uint64))


def test_product_sources_never_touch_the_oracle():
    pkg = os.path.join(ROOT, "raytracing-gpu_b200")
    for base, _, files in os.walk(pkg):
        if "build" in base.split(os.sep):
            continue
        for f in files:
            if f.endswith((".cu", ".cuh", ".cpp", ".h", ".py", "Makefile")):
                text = open(os.path.join(base, f), errors="replace").read()
                assert "oracle" not in text.lower(), "%s mentions the oracle" % os.path.join(base, f)


def test_ctypes_mirrors_match_the_header_layout(tmp_path):
    """rtb_frame_opts and rtb_stats as gcc lays them out from include/rtb200.h against the ctypes
    structures tests and bench.py pass through the C ABI (a field added on one side only would
    shift everything after it silently)"""
    import shutil
    import subprocess
    gcc = shutil.which("gcc")
    if not gcc:
        pytest.skip("no gcc")
    fo = [n for n, _ in pyrtb.FrameOpts._fields_]
    st = [n for n, _ in pyrtb.Stats._fields_]
    prog = ['#include <stdio.h>', '#include <stddef.h>', '#include "rtb200.h"', 'int main(void) {',
            '  printf("%zu %zu\\n", sizeof(rtb_frame_opts), sizeof(rtb_stats));']
    prog += ['  printf("%%zu\\n", offsetof(rtb_frame_opts, %s));' % n for n in fo]
    prog += ['  printf("%%zu\\n", offsetof(rtb_stats, %s));' % n for n in st]
    prog += ['  return 0; }']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(prog))
    exe = tmp_path / "layout"
    subprocess.run([gcc, "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    assert int(out[0]) == C.sizeof(pyrtb.FrameOpts) and int(out[1]) == C.sizeof(pyrtb.Stats)
    got = [int(x) for x in out[2:]]
    want = [getattr(pyrtb.FrameOpts, n).offset for n in fo] + [getattr(pyrtb.Stats, n).offset for n in st]
    assert got == want
