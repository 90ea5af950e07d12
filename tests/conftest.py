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
