import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
SCENES = os.path.join(GOLDEN, "scenes")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the product library and the oracle in-tree (no-op when up to date)."""
    import __graft_entry__ as g
    g.build()


@pytest.fixture(scope="session")
def rrt():
    import rrt_b200
    return rrt_b200


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import oracle
    return oracle


_scene_cache = {}


def load_scene(name):
    from rrt_b200.loader import load_scene_npz
    if name not in _scene_cache:
        _scene_cache[name] = load_scene_npz(os.path.join(SCENES, name + ".npz"))
    return _scene_cache[name]


@pytest.fixture(scope="session")
def scenes():
    return load_scene


def tiny_scene(tris, kr=0.0, bump=False, tex=None):
    """A few hand-placed triangles with one material. tris: list of 9-float vertex rows."""
    from rrt_b200 import MaterialSpec, SceneData
    from rrt_b200.synthetic import checker_texture
    v = np.asarray(tris, dtype=np.float64).reshape(-1, 9)
    n = len(v)
    e1, e2 = v[:, 3:6] - v[:, 0:3], v[:, 6:9] - v[:, 0:3]
    g = np.cross(e1, e2)
    ln = np.linalg.norm(g, axis=1, keepdims=True)
    g = g / np.where(ln > 0, ln, 1.0)
    vn = np.concatenate([g, g, g], 1)
    vt = np.tile(np.array([0.0, 0.0, 1.0, 0.0, 0.0, 1.0]), (n, 1))
    texs = [checker_texture(64) if tex is None else tex]
    if bump:
        texs.append(checker_texture(64)[::-1].copy())
    mat = MaterialSpec("m", (1, 1, 1), (1, 1, 1), (1, 1, 1), 240.0, kr, 0, 1 if bump else -1)
    return SceneData(v, vt, vn, np.zeros(n, np.int32), [mat], texs, texture_names=["a", "b"][:len(texs)])


def random_scene(n, seed, extent=6.0, size=1.5, kr=0.3):
    rng = np.random.default_rng(seed)
    c = rng.uniform(-extent, extent, size=(n, 1, 3)) + np.array([0.0, 2.0, 4.0])
    p = c + rng.uniform(-size, size, size=(n, 3, 3))
    return tiny_scene(p.reshape(n, 9), kr=kr)


def rgb_planes(fb):
    return ((fb[..., None] >> np.array([16, 8, 0])) & 255).astype(np.int32)


def has_gpu():
    import rrt_b200
    return rrt_b200.abi.lib().rrt_device_count() > 0


def grazing_rays(sd, n, rng, eps_lo=1e-16, eps_hi=1e-8):
    """Adversarial rays for the conservative f32 culling: every ray runs IN the plane of a scene
    triangle or within |eps| rad of it (|eps| log-uniform in [eps_lo, eps_hi]) and is aimed at a point of that triangle,
    so that the reference's determinant `a` (ray.rs:64) sweeps from rounding noise to 1e-8 relative.
    -> origins, directions, index of the triangle each ray was built on."""
    v = np.asarray(sd.v, dtype=np.float64).reshape(-1, 3, 3)
    tri = rng.integers(0, len(v), n)
    a, b, c = v[tri, 0], v[tri, 1], v[tri, 2]
    e1, e2 = b - a, c - a
    nrm = np.cross(e1, e2)
    area2 = np.linalg.norm(nrm, axis=1, keepdims=True)
    ok = area2[:, 0] > 1e-12
    nrm = nrm / np.where(area2 > 0, area2, 1.0)
    w = rng.dirichlet([1, 1, 1], n)
    p = a * w[:, :1] + b * w[:, 1:2] + c * w[:, 2:3]
    th = rng.uniform(0, 2 * np.pi, (n, 1))
    u1 = e1 / np.maximum(np.linalg.norm(e1, axis=1, keepdims=True), 1e-300)
    u2 = np.cross(nrm, u1)
    u = np.cos(th) * u1 + np.sin(th) * u2                      # unit vector in the triangle's plane
    dist = rng.uniform(0.3, 12.0, (n, 1))
    s = 10.0 ** rng.uniform(-1, 1, (n, 1))                     # the reference never normalises camera rays (engine.rs:207-211)
    eps = 10.0 ** rng.uniform(np.log10(eps_lo), np.log10(eps_hi), (n, 1)) * rng.choice([-1.0, 1.0], (n, 1))
    d = u * s + nrm * (eps * s)
    o = p - u * dist + nrm * (-eps * dist)                     # meets the plane near p
    z = rng.random(n) < 0.33                                   # a third of the rays exactly in the plane, up to rounding
    d[z] = (u * s)[z]
    o[z] = (p - u * dist)[z]
    return o[ok], d[ok], tri[ok]


def reference_determinant_rel(sd, tri_ids, d):
    """|a| / (|e1| |e2| |d|) of ray.rs:64 for the given triangles and directions: how far the ray is from the triangle's
    plane, in radians; rounding noise of `a` itself is ~1e-16 on this scale."""
    v = np.asarray(sd.v, dtype=np.float64).reshape(-1, 3, 3)[tri_ids]
    e1, e2 = v[:, 1] - v[:, 0], v[:, 2] - v[:, 0]
    a = np.einsum("ij,ij->i", e1, np.cross(d, e2))
    den = np.linalg.norm(e1, axis=1) * np.linalg.norm(e2, axis=1) * np.linalg.norm(d, axis=1)
    return np.abs(a) / np.maximum(den, 1e-300)
