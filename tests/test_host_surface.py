"""Host surface kept from the reference: .svati parser (gpu/parser.cpp, gpu/parse_obj.cpp;
format in SURVEY Appendix B), scene container (gpu/scene.cu:146-218), PNG/PPM writers."""
import os
import subprocess

import numpy as np
import pytest

import oracle_py
import pyrtb
from util import golden_names, load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SMALL = """camera 64 48 0.0 0.0 -4.0 1.0 0.0 0.0 0.0 -1.0 0.0 90.0
# a comment line at top level
a_light 0.5 0.25 0.125
d_light 1.0 1.0 1.0 0.5 -1.0 1.0
p_light 0.5 0.5 0.5 1 2 3

object 6
Ns 96.078431
Ka 0.1 0.2 0.3
Kd 0.8 0.0 0.0
Ks 0.5 0.5 0.5
Nr 0.25
d 0.9
v 1 0 0
v 2 0 0
v 3 0 0
v 4 0 0
v 5 0 0
v 6 0 0
vn 0 0 1
vn 0 0 2
vn 0 0 3
vn 0 0 4
vn 0 0 5
vn 0 0 6
object 3
Kd 0 1 0
v 0 0 0 v 0 1 0 v 1 0 0
vn 0 0 -1 vn 0 0 -1 vn 0 0 -1
"""


def test_parser_semantics(tmp_path):
    p = tmp_path / "s.svati"
    p.write_text(SMALL)
    hs = pyrtb.HostScene.parse(str(p))
    a = hs.arrays()
    assert hs.width == 64 and hs.height == 48 and hs.triangle_count == 3
    assert a["tri_counts"].tolist() == [2, 1]
    # LIFO reversal (gpu/parse_obj.cpp:22-23,54-61): memory vertex k = file vertex NV-1-k
    assert a["vertex"][:2].reshape(6, 3)[:, 0].tolist() == [6, 5, 4, 3, 2, 1]
    assert a["normal"][:2].reshape(6, 3)[:, 2].tolist() == [6, 5, 4, 3, 2, 1]
    assert a["vertex"][2].tolist() == [[1, 0, 0], [0, 1, 0], [0, 0, 0]]
    m = a["materials"]
    assert np.allclose(m[0], [0.1, 0.2, 0.3, 0.8, 0, 0, 0.5, 0.5, 0.5, 96.078431, 1.0, 0.25, 0.9])
    assert np.allclose(m[1], [0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 1, 0, 1])   # defaults of gpu/scene.cu:199-209
    L = a["lights"]
    assert L[:, 0].tolist() == [pyrtb.AMBIENT, pyrtb.DIRECTIONAL, pyrtb.POINT]
    assert np.allclose(L[0, 1:4], [0.5, 0.25, 0.125]) and np.allclose(L[2, 4:], [1, 2, 3])
    assert np.allclose(a["camera"], [64, 48, 0, 0, -4, 1, 0, 0, 0, -1, 0, 90])


@pytest.mark.parametrize("text,needle", [
    ("camera 1 1 0 0 0 1 0 0 0 1 0 90\nbogus 1 2 3\n", "bogus"),
    ("object 3\nKx 1 2 3\n", "Kx"),
    ("object 3\nv 0 0 0 v 1 0 0\n", "truncated"),
    ("a_light 0.5 zero 0.5\n", "number"),
])
def test_parser_rejects_bad_input(tmp_path, text, needle):
    p = tmp_path / "bad.svati"
    p.write_text(text)
    with pytest.raises(ValueError) as e:
        pyrtb.HostScene.parse(str(p))
    assert needle in str(e.value)


def test_parser_missing_file():
    with pytest.raises(ValueError):
        pyrtb.HostScene.parse("/nonexistent/file.svati")


def test_last_camera_wins_and_empty_scene(tmp_path):
    p = tmp_path / "c.svati"
    p.write_text("camera 10 10 0 0 0 1 0 0 0 1 0 60\ncamera 20 30 1 2 3 1 0 0 0 1 0 70\n")
    hs = pyrtb.HostScene.parse(str(p))
    assert (hs.width, hs.height, hs.triangle_count) == (20, 30, 0) and hs.s.object_count == 0


@pytest.mark.parametrize("name", ["triangle", "cube", "dir-light-shadows"])
def test_write_svati_round_trip(tmp_path, name):
    hs, z = load_golden(name)
    p = str(tmp_path / "rt.svati")
    hs.write_svati(p)
    back = pyrtb.HostScene.parse(p).arrays()
    a = hs.arrays()
    for k in ("vertex", "normal", "tri_counts", "materials", "lights", "camera"):
        assert np.array_equal(np.asarray(a[k]).view(np.uint8), np.asarray(back[k]).view(np.uint8)), k


@pytest.mark.skipif(not oracle_py.have_reference(), reason="needs the reference checkout (build container only)")
@pytest.mark.parametrize("name", golden_names())
def test_parser_reproduces_the_golden_scene_arrays(name):
    hs = pyrtb.HostScene.parse(os.path.join(oracle_py.REFERENCE, "tests", name + ".svati"))
    a, (_, z) = hs.arrays(), load_golden(name)
    for k in ("vertex", "normal", "tri_counts", "materials", "lights", "camera"):
        assert np.array_equal(np.asarray(a[k]), z[k]), k


def test_terrain_generator_is_deterministic_and_sized():
    a = pyrtb.HostScene.terrain(40, 64, 36).arrays()
    b = pyrtb.HostScene.terrain(40, 64, 36).arrays()
    assert a["vertex"].shape == (2 * 40 * 40, 3, 3) and len(a["tri_counts"]) == 16
    assert np.array_equal(a["vertex"], b["vertex"]) and np.array_equal(a["normal"], b["normal"])
    assert a["vertex"][..., 0].min() == -500 and a["vertex"][..., 0].max() == 500
    assert a["vertex"][..., 2].min() == 0 and a["vertex"][..., 2].max() == 1000
    assert np.allclose(np.linalg.norm(a["normal"], axis=-1), 1, atol=1e-5) and (a["normal"][..., 1] > 0).all()
    assert 2 * 2237 * 2237 == 10008338


def test_png_writer_decodes_to_the_same_pixels(tmp_path):
    from PIL import Image
    rng = np.random.default_rng(0)
    for (h, w) in ((1, 1), (37, 53), (540, 960)):
        img = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        stride = w * 4 + 12     # padded rows
        buf = np.zeros((h, stride), np.uint8)
        buf[:, :w * 4] = img.reshape(h, w * 4)
        path = str(tmp_path / ("t%dx%d.png" % (w, h)))
        assert pyrtb.lib().rtb_write_png(buf.ctypes.data, w, h, stride, path.encode()) == 0
        back = np.asarray(Image.open(path).convert("RGBA"))
        assert np.array_equal(back, img)


def test_rt_cli_usage_and_errors(tmp_path):
    rt = os.path.join(ROOT, "raytracing-gpu_b200", "rt")
    r = subprocess.run([rt], capture_output=True, text=True)
    assert r.returncode == 1 and "usage:" in r.stderr and "file.svati output.png" in r.stderr
    r = subprocess.run([rt, "/nonexistent.svati", str(tmp_path / "o.png")], capture_output=True, text=True)
    assert r.returncode == 1
    assert "Using structure of arrays (SOA) layout" in r.stdout      # banner of gpu/rt.cpp:58-64


def _big_scene_text(n_objects=5, tris=4000, seed=3, comment=None, break_at=None):
    """A > 1 MB .svati (the size from which the loader cuts the file at `object` tokens)."""
    rng = np.random.default_rng(seed)
    out = ["camera 64 48 0 0 -4 1 0 0 0 -1 0 90", "a_light 0.5 0.25 0.125", "# heading comment"]
    for o in range(n_objects):
        if comment is not None and o == 2:
            out.append(comment)
        out.append("object %d" % (3 * tris))
        out.append("Kd %.3f 0.5 0.25" % (o / 10))
        if o % 2:
            out.append("Nr 0.5")
        v = rng.standard_normal((3 * tris, 3)).astype(np.float32) * 50
        n = rng.standard_normal((3 * tris, 3)).astype(np.float32)
        for i in range(3 * tris):
            if break_at == (o, i):
                out.append("v 1.0 oops 2.0")
            out.append("v %.9g %+.8e %s" % (v[i, 0], v[i, 1], repr(float(v[i, 2]))))
            out.append("vn %.9g %.9g %.9g" % tuple(n[i]))
        out.append("d_light 1 1 1 0 -1 %d" % o)
    return "\n".join(out) + "\n"


def _parse_with_threads(path, threads):
    old = os.environ.get("RTB_PARSE_THREADS")
    os.environ["RTB_PARSE_THREADS"] = str(threads)
    try:
        return pyrtb.HostScene.parse(path)
    finally:
        if old is None:
            del os.environ["RTB_PARSE_THREADS"]
        else:
            os.environ["RTB_PARSE_THREADS"] = old


def test_parallel_loader_equals_the_one_piece_parse(tmp_path):
    """host/parser.cpp cuts big files at `object` tokens and parses the pieces on all threads;
    arrays, materials, light order and camera must equal the single-threaded walk."""
    p = tmp_path / "big.svati"
    text = _big_scene_text()
    assert len(text) > (1 << 20)
    p.write_text(text)
    one = _parse_with_threads(str(p), 1).arrays()
    many = _parse_with_threads(str(p), 8).arrays()
    assert one["tri_counts"].tolist() == [4000] * 5
    for k in one:
        assert np.array_equal(one[k], many[k]), k
    assert one["lights"][:, 0].tolist() == [pyrtb.AMBIENT] + [pyrtb.DIRECTIONAL] * 5
    assert one["lights"][1:, 6].tolist() == [0, 1, 2, 3, 4]            # file order kept
    # values: correctly rounded decimal -> binary32, whatever spelling the writer used
    rng = np.random.default_rng(3)
    v0 = rng.standard_normal((12000, 3)).astype(np.float32) * 50
    assert np.array_equal(many["vertex"][:4000].reshape(-1, 3), v0[::-1])   # LIFO reversal


def test_parallel_loader_comment_holding_the_word_object(tmp_path):
    """`# ... object ...` at top level: the cut inside the comment must not become an object."""
    p = tmp_path / "c.svati"
    p.write_text(_big_scene_text(comment="# the next object is big"))
    one = _parse_with_threads(str(p), 1).arrays()
    many = _parse_with_threads(str(p), 8).arrays()
    assert many["tri_counts"].tolist() == [4000] * 5
    for k in one:
        assert np.array_equal(one[k], many[k]), k


def test_parallel_loader_reports_the_same_error(tmp_path):
    p = tmp_path / "e.svati"
    p.write_text(_big_scene_text(break_at=(3, 7)))
    msgs = []
    for threads in (1, 8):
        with pytest.raises(ValueError) as e:
            _parse_with_threads(str(p), threads)
        msgs.append(str(e.value))
    assert msgs[0] == msgs[1] and "number" in msgs[0]
