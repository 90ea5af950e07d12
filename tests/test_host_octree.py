"""The product's host octree (csrc/rrt_octree.cpp, flat arrays + iterative insertion) must reproduce
the reference's incremental construction (src/collision/octree.rs:41-241) node for node — the tree
shape decides tie winners and reachability.  Checked against the oracle's Vec-of-Vec restatement."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import random_scene, tiny_scene
from oracle.oracle import OracleScene


def build_product_tree(rrt, sd):
    L = rrt.abi.lib()
    h = C.c_void_p()
    rmin, rmax = (C.c_double * 3)(*sd.root_min), (C.c_double * 3)(*sd.root_max)
    rrt.abi.check(L.rrt_octree_build(sd.v.ctypes.data, sd.n_tris, rmin, rmax, C.byref(h)))
    nodes = []
    for i in range(L.rrt_octree_node_count(h)):
        box = (C.c_double * 6)()
        tc, fc, ll = C.c_uint64(), C.c_uint64(), C.c_uint64()
        rrt.abi.check(L.rrt_octree_node(h, i, box, C.byref(tc), C.byref(fc), C.byref(ll)))
        lst = np.zeros(max(1, ll.value), np.uint32)
        assert L.rrt_octree_node_list(h, i, lst.ctypes.data, ll.value) == ll.value
        nodes.append((list(box), tc.value, fc.value, lst[:ll.value].tolist()))
    L.rrt_octree_destroy(h)
    return nodes


def assert_same_tree(rrt, sd):
    mine = build_product_tree(rrt, sd)
    o = OracleScene(sd)
    assert len(mine) == o.node_count()
    for i, (box, tc, fc, lst) in enumerate(mine):
        ob, otc, ofc, oll = o.node(i)
        assert (box, tc, fc, len(lst)) == (ob, otc, ofc, oll), i
        assert lst == o.node_list(i), i
    return mine


@pytest.mark.parametrize("name", ["model", "model2", "model3"])
def test_shipped_scenes(rrt, scenes, name):
    nodes = assert_same_tree(rrt, scenes(name))
    expect = {"model": (417, 472), "model2": (3265, 1110), "model3": (5593, 801)}[name]
    assert (len(nodes), len(nodes[0][3])) == expect


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_random_soups(rrt, seed):
    assert_same_tree(rrt, random_scene(2000, seed=seed, size=0.4))


def test_known_answers_and_edge_cases(rrt):
    # empty scene: a lone root
    sd = tiny_scene(np.zeros((0, 9)))
    assert build_product_tree(rrt, sd) == [([-20.0] * 3 + [20.0] * 3, 0, 0, [])]
    # dead-test known answer (octree.rs:326-529, live rules): 9 nodes, tri 1 in child 7
    sd = tiny_scene([[5.2, 5.2, 5.2, 5.5, 5.2, 5.2, 5.2, 5.5, 5.5], [0.6, 0.6, 0.6, 0.8, 0.6, 0.6, 0.6, 0.8, 0.8]])
    sd.root_min, sd.root_max = (-10.0,) * 3, (10.0,) * 3
    n = assert_same_tree(rrt, sd)
    assert len(n) == 9 and n[0][3] == [0] and n[7][3] == [1] and n[7][0] == [0.0] * 3 + [10.0] * 3
    # dropped triangle, NaN vertex (fmin/fmax ignore NaN like Rust), triangle exactly on a split plane
    sd = tiny_scene([[1, 1, 1, 2, 1, 1, 1, 2, 1], [30, 30, 30, 31, 30, 30, 30, 31, 30],
                     [np.nan, 1, 1, 3, 1, 1, 3, 2, 1], [0, 0, 0, 0, 1, 0, 0, 0, 1], [4, 4, 4, 5, 4, 4, 4, 5, 4]])
    assert_same_tree(rrt, sd)


def test_depth_limit_is_an_error_not_a_crash(rrt):
    # every coincident duplicate adds one level in the reference (first-in-leaf rule, octree.rs:77-80);
    # the device walk has a fixed short stack, so trees deeper than kMaxOctreeDepth are refused
    t = [0.3] * 9     # degenerate point triangle at a non-dyadic coordinate
    sd = tiny_scene([t] * 100)
    sd.root_min, sd.root_max = (-1e30,) * 3, (1e30,) * 3   # f64 precision bounds the depth at ~58 for the +-20 root
    L = rrt.abi.lib()
    h = C.c_void_p()
    rmin, rmax = (C.c_double * 3)(*sd.root_min), (C.c_double * 3)(*sd.root_max)
    rc = L.rrt_octree_build(sd.v.ctypes.data, sd.n_tris, rmin, rmax, C.byref(h))
    assert rc == -5 and b"deeper" in L.rrt_last_error()


def test_flatten_is_independent_of_thread_count(rrt):
    # the per-list culling hierarchies are built on OpenMP threads and concatenated in node order:
    # the flattened arrays (digest over nodes, records, culling nodes) must not depend on the thread count
    import subprocess
    import sys
    code = (
        "import ctypes as C, sys\n"
        "sys.path.insert(0, %r)\n"
        "import rrt_b200\n"
        "from rrt_b200 import abi\n"
        "from rrt_b200.synthetic import triangle_soup\n"
        "sd = triangle_soup(60000, size=0.5)\n"
        "L = abi.lib(); h = C.c_void_p()\n"
        "lo = (C.c_double * 3)(-20, -20, -20); hi = (C.c_double * 3)(20, 20, 20)\n"
        "abi.check(L.rrt_octree_build(sd.v.ctypes.data, sd.n_tris, lo, hi, C.byref(h)))\n"
        "st = abi.SceneStats(); sec = C.c_double()\n"
        "abi.check(L.rrt_octree_flatten_stats(h, sd.v.ctypes.data, C.byref(st), C.byref(sec)))\n"
        "print(st.flat_digest, st.n_cull_nodes, st.n_tris_stored)\n"
    ) % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for threads in ("1", "5"):
        r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, OMP_NUM_THREADS=threads),
                           capture_output=True, text=True, check=True)
        outs.append(r.stdout.strip())
    assert outs[0] == outs[1] and int(outs[0].split()[2]) == 60000


def _tree_digest(rrt, sd):
    L = rrt.abi.lib()
    h = C.c_void_p()
    rmin, rmax = (C.c_double * 3)(*sd.root_min), (C.c_double * 3)(*sd.root_max)
    rrt.abi.check(L.rrt_octree_build(sd.v.ctypes.data, sd.n_tris, rmin, rmax, C.byref(h)))
    out = (L.rrt_octree_digest(h), L.rrt_octree_node_count(h))
    L.rrt_octree_destroy(h)
    return out


def test_parallel_builder_is_the_sequential_tree(rrt, scenes, monkeypatch):
    # the top-down OpenMP construction must give the tree of sequential insertion: boxes, lists, counts AND the
    # node numbering (digest over every node and list link).  Cases: shipped scenes, soups with many straddlers,
    # coincident triangles (one subdivision per duplicate: a deep chain), triangles outside the root box.
    from rrt_b200.synthetic import triangle_soup, sphere_grid
    cases = [scenes("model"), scenes("model2"), scenes("model3"), triangle_soup(50000, size=0.6),
             triangle_soup(300000, size=0.05), sphere_grid(nx=2, ny=2, nlon=64, nlat=32)]
    dup = triangle_soup(3000, size=0.3)
    dup.v[100:140] = dup.v[100]
    far = triangle_soup(5000, extent=30.0, size=0.5)          # some boxes miss the +-20 root box: dropped
    cases += [dup, far]
    for sd in cases:
        monkeypatch.setenv("RRT_OCTREE_BUILD", "seq")
        a = _tree_digest(rrt, sd)
        monkeypatch.setenv("RRT_OCTREE_BUILD", "par")
        b = _tree_digest(rrt, sd)
        assert a == b and a[1] >= 1


def test_parallel_builder_node_for_node_against_oracle(rrt, monkeypatch):
    monkeypatch.setenv("RRT_OCTREE_BUILD", "par")
    assert_same_tree(rrt, random_scene(2500, seed=11))
