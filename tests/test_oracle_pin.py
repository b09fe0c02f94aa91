"""Pins the CPU oracle (oracle/oracle.cpp).  The reference ships no golden images,
so the pins are
  * tests/golden/*.npz: frames rendered by the REFERENCE's own device sources compiled for
    the host (oracle/_ref/libref_gpusem.so; generating script tests/golden/make_golden.py) —
    the oracle must reproduce them byte for byte, ids included;
  * live, where the reference checkout exists: bit-identical hit records against the
    reference's gpu/hit.cu collide() and its cpu/hit.c collide() on random and primary rays;
  * the colour known-answer values of SURVEY Appendix D."""
import numpy as np
import pytest

import oracle_py
from util import golden_names, load_golden, primary_rays

FAST = ["triangle", "triangle-ambient", "triangle-left-ambient", "cube", "cube-ambient", "dir-light-shadows",
        "point-light"]


@pytest.mark.parametrize("name", [n for n in golden_names() if n in FAST])
def test_oracle_reproduces_reference_frames(name):
    hs, z = load_golden(name)
    o = oracle_py.orc_render(hs, 3)
    assert np.array_equal(o["rgba"], z["rgba"])
    assert np.array_equal(o["tri"], z["tri"].astype(np.int32))
    assert np.array_equal(o["obj"], z["obj"].astype(np.int32))
    assert o["counts"].tolist() == z["counts"].tolist()


@pytest.mark.parametrize("name", [n for n in golden_names() if n not in FAST])
def test_oracle_reproduces_reference_samples_of_heavy_scenes(name):
    """Heavy scenes: 6000 random samples (ids) instead of the whole frame."""
    hs, z = load_golden(name)
    rng = np.random.default_rng(11)
    W, H = hs.width * 3, hs.height * 3
    xy = np.stack([rng.integers(0, W, 6000), rng.integers(0, H, 6000)], axis=1)
    o = oracle_py.orc_render_samples(hs, xy, 3)
    assert np.array_equal(o["tri"], z["tri"][xy[:, 1], xy[:, 0]].astype(np.int32))
    assert np.array_equal(o["obj"], z["obj"][xy[:, 1], xy[:, 0]].astype(np.int32))


def test_fixture_set_is_the_reference_corpus():
    names = golden_names()
    assert len(names) == 20, names
    tris = {n: int(load_golden(n)[1]["tri_counts"].sum()) for n in ("triangle", "cube", "car-on-road", "spheres")}
    assert tris == {"triangle": 1, "cube": 12, "car-on-road": 2586, "spheres": 4812}


@pytest.mark.skipif(not oracle_py.have_reference(), reason="needs the reference checkout (build container only)")
@pytest.mark.parametrize("name", ["cube", "dir-light-shadows", "island_smooth", "susan"])
def test_hit_records_bitwise_vs_reference_gpu_and_cpu_sources(name):
    if name not in golden_names():
        pytest.skip("fixture not generated yet")
    hs, z = load_golden(name)
    rng = np.random.default_rng(5)
    W, H = hs.width * 3, hs.height * 3
    xy = np.stack([rng.integers(0, W, 1500), rng.integers(0, H, 1500)], axis=1)
    rays = primary_rays(hs, 3, xy)
    v = z["vertex"].reshape(-1, 3)
    lo, hi = v.min(0), v.max(0)
    org = rng.uniform(lo - 1, hi + 1, size=(1500, 3)).astype(np.float32)
    rays = np.concatenate([rays, np.concatenate([org, rng.normal(size=(1500, 3)).astype(np.float32)], axis=1)])
    tri, dist, hit = oracle_py.orc_trace_rays(hs, rays)
    ref_gpu = oracle_py.ref_collide(hs, rays)
    ref_cpu = oracle_py.refcpu_collide(hs, rays)
    assert np.array_equal(hit.view(np.uint32), ref_gpu.view(np.uint32))
    assert np.array_equal(hit.view(np.uint32), ref_cpu.view(np.uint32))
    assert (tri >= 0).sum() > 100
    # ids of the primary rays agree with the frame fixture
    assert np.array_equal(tri[:1500], z["tri"][xy[:, 1], xy[:, 0]].astype(np.int32))


def test_colour_known_answers():
    L = oracle_py.oracle()

    def rgb(c):
        return (c & 255, (c >> 8) & 255, (c >> 16) & 255)
    # truncation, not rounding
    assert rgb(L.orc_init_color(0.8, 0.64, 0.65)) == (204, 163, 165)
    assert rgb(L.orc_init_color(0.3, 0.5, 1.5)) == (76, 127, 255)
    assert rgb(L.orc_init_color(-1.0, float("nan"), 0.0)) == (0, 0, 0)
    white = L.orc_init_color(1, 1, 1)
    for c in range(256):
        col = c | (c << 8) | (c << 16) | (255 << 24)
        assert L.orc_color_mul(col, 1.0) == col                   # color_mul(c, 1) is the identity
        assert L.orc_color_mults(white, col) == col               # color_mults(white, c) too
        assert rgb(L.orc_color_add(col, col))[0] == min(255, 2 * c)   # saturating add
    assert rgb(L.orc_color_mul(L.orc_init_color(0.8, 0.8, 0.8), 0.5)) == (102, 102, 102)
