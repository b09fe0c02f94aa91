import glob
import os

import numpy as np

import pyrtb

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    hs = pyrtb.HostScene.from_arrays(z["vertex"], z["normal"], z["tri_counts"], z["materials"], z["lights"], z["camera"])
    return hs, z


def rgb_report(a, b):
    """(fraction of pixels whose RGB differs by more than 1, fraction identical, max abs diff)"""
    d = np.abs(a[..., :3].astype(np.int16) - b[..., :3].astype(np.int16)).max(axis=-1)
    return float((d > 1).mean()), float((d == 0).mean()), int(d.max())


def primary_rays(hs, aliasing, xy):
    """Primary rays of samples xy (n,2) with the arithmetic of gpu/raytracer.cu:178-182,91-110 in numpy float32."""
    f32 = np.float32
    c = hs.s.camera
    W, H = c.width * aliasing, c.height * aliasing

    def norm(v):
        v = np.array(v, f32)
        r = np.sqrt(f32(f32(v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]))
        return (v / r).astype(f32)

    u = norm([c.u.x, c.u.y, c.u.z])
    v = norm([c.v.x, c.v.y, c.v.z])
    w = np.array([u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]], f32)
    L = f32(W / (2 * np.tan(np.float64(c.fov) * np.pi / 360)))
    pos = np.array([c.position.x, c.position.y, c.position.z], f32)
    Cc = (pos + w * L).astype(f32)
    xy = np.asarray(xy)
    ui = (u[None, :] * (xy[:, 0] - W // 2).astype(f32)[:, None]).astype(f32)
    vj = (v[None, :] * (xy[:, 1] - H // 2).astype(f32)[:, None]).astype(f32)
    point = ((Cc[None, :] + ui).astype(f32) + vj).astype(f32)
    d = (pos[None, :] - point).astype(f32)
    ln = np.sqrt(((d[:, 0] * d[:, 0]).astype(f32) + (d[:, 1] * d[:, 1]).astype(f32)).astype(f32) + (d[:, 2] * d[:, 2]).astype(f32)).astype(f32)
    d = (d / ln[:, None]).astype(f32)
    return np.concatenate([point, d], axis=1).astype(f32)


def random_scene(seed, T, w=320, h=200):
    """Random triangle soup: three size classes, 30 % extreme needles, random vertex normals,
    four objects (some mirrors), ambient + directional + point light, camera 9 units out."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-3, 3, (T, 1, 3))
    kind = rng.integers(0, 3, T)
    size = np.where(kind == 0, 0.8, np.where(kind == 1, 0.15, 2.5))[:, None, None]
    v = c + rng.standard_normal((T, 3, 3)) * size
    needle = rng.random(T) < 0.3                        # third vertex almost on the first edge
    t = rng.random((T, 1))
    v[needle, 2] = (v[needle, 0] * (1 - t[needle]) + v[needle, 1] * t[needle]) + rng.standard_normal((needle.sum(), 3)) * 0.003
    n = rng.standard_normal((T, 3, 3))
    mats = [[0.2, 0.2, 0.2, *rng.uniform(0.2, 1, 3), 0.3, 0.3, 0.3, 20, 1, float(rng.choice([0, 0, 0.5, 1.0])), 1] for _ in range(4)]
    counts = [T // 4, T // 4, T // 4, T - 3 * (T // 4)]
    lights = [[pyrtb.AMBIENT, .3, .3, .3, 0, 0, 0], [pyrtb.DIRECTIONAL, 1, 1, 1, *rng.standard_normal(3)], [pyrtb.POINT, 1, 1, 1, *rng.uniform(-6, 6, 3)]]
    pos = rng.standard_normal(3); pos = pos / np.linalg.norm(pos) * 9
    wv = -pos / np.linalg.norm(pos)
    up = np.array([0, 1.0, 0]); u = np.cross(up, wv); u /= np.linalg.norm(u); vv = np.cross(u, wv) * -1
    # the reference's w = u x v must point away from the scene (origin on the image plane behind the camera)
    if np.dot(np.cross(u, vv), wv) > 0: vv = -vv
    cam = [w, h, *pos, *u, *vv, 60]
    return pyrtb.HostScene.from_arrays(v.astype(np.float32), n.astype(np.float32), counts, mats, lights, cam)
