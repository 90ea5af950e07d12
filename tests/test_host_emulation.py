"""The product's traversal headers compiled for the HOST (tests/host_emulation/emu_trace.cpp) fuzzed against the
oracle — no GPU needed.  What runs here is the same source the CUDA kernels inline (csrc/rrt_trace.cuh,
csrc/rrt_device.cuh) over the product's own host octree + flattener (csrc/rrt_octree.cpp): the short-stack walk,
the FMA quotients, the conservative f32 culling hierarchies / list boxes / subtree boxes and the division-free
Moller-Trumbore reject must return, bit for bit, what the reference recursion (the oracle) returns
(DESIGN.md §5).  It is test infrastructure: librrt_b200.so does not contain it and has no CPU fallback
(tests/test_abi.py::test_no_cpu_fallback_without_device)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import grazing_rays, random_scene, reference_determinant_rel, tiny_scene

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
EMU_DIR = os.path.join(HERE, "host_emulation")
EMU_SRC = os.path.join(EMU_DIR, "emu_trace.cpp")
# RRT_EMU_FLAGS="-DFOO=1 ...": the same tests over a compile-time VARIANT of the traversal headers (the flags a
# variant build of the library gets through RRT_EXTRA_FLAGS), in its own library file
EMU_FLAGS = os.environ.get("RRT_EMU_FLAGS", "").split()
EMU_LIB = os.path.join(EMU_DIR, "libemu_trace" + "".join(c if c.isalnum() else "_" for c in "".join(EMU_FLAGS)) + ".so")
CSRC = os.path.join(ROOT, "rust-ray-tracer_b200", "csrc")


def _lib_name(extra):
    return os.path.join(EMU_DIR, "libemu_trace" + "".join(c if c.isalnum() else "_" for c in "".join(extra)) + ".so")


def _build_emu(extra=None):
    """-> path of the emulation library built with the extra -D flags (default: RRT_EMU_FLAGS)."""
    extra = EMU_FLAGS if extra is None else list(extra)
    lib = _lib_name(extra)
    deps = [EMU_SRC, os.path.join(ROOT, "include", "rrt.h")] + [
        os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h", ".cpp"))]
    if os.path.exists(lib) and all(os.path.getmtime(lib) >= os.path.getmtime(d) for d in deps):
        return lib
    flags = ["-O2", "-std=c++17", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-shared", "-w"] + extra
    try:
        if " fma " in open("/proc/cpuinfo").read():
            flags.append("-mfma")          # hardware single-rounding fma(); without it glibc's exact software fma is used
    except OSError:
        pass
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
    subprocess.run(["g++", *flags, "-I" + cuda_inc, "-o", lib, EMU_SRC, os.path.join(CSRC, "rrt_octree.cpp")], check=True)
    return lib


def _load_emu(lib):
    L = C.CDLL(lib)
    vp, dp = C.c_void_p, C.POINTER(C.c_double)
    L.emu_scene_create.argtypes = [vp, C.c_uint64, vp, vp, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.POINTER(vp)]
    L.emu_scene_destroy.argtypes = [vp]
    L.emu_scene_destroy.restype = None
    L.emu_scene_stored.argtypes = [vp]
    L.emu_scene_stored.restype = C.c_uint64
    L.emu_scene_planes_safe.argtypes = [vp]
    L.emu_trace_rays.argtypes = [vp, C.c_uint64, vp, vp, vp, C.c_int, vp, vp]
    L.emu_scene_set_shading.argtypes = [vp, C.c_uint64, vp, vp, vp, C.c_uint32, vp, C.c_uint32, vp]
    L.emu_ray_colours.argtypes = [vp, vp, C.c_uint64, vp, vp, vp]
    L.emu_render_shard.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp]
    L.emu_set_allow.argtypes = [C.c_uint32]
    L.emu_set_allow.restype = None
    L.emu_search_outcomes.argtypes = [vp, C.c_uint64, vp, vp, vp, C.c_int, vp]
    return L


@pytest.fixture(scope="module")
def emu():
    return _load_emu(_build_emu())


# By default the emulation lets traverse() try BOTH searches of the scene-wide hierarchy on every ray they apply to
# (rrt_trace.cuh: scene_search on any-hit rays and on closest-hit rays with max_t = inf — also on rays from
# inside the geometry, which the render kernel leaves to the walk): the whole suite fuzzes their hand-over rules.
ALLOW_ALL, ALLOW_WALK_ONLY = 3, 0


class EmuScene:
    def __init__(self, L, sd, cull_threshold=1, cull_leaf=1, parallel=False, slow_div=False):
        self.L = L
        v = np.ascontiguousarray(sd.v, dtype=np.float64)
        lo, hi = np.asarray(sd.root_min, dtype=np.float64), np.asarray(sd.root_max, dtype=np.float64)
        h = C.c_void_p()
        rc = L.emu_scene_create(v.ctypes.data, v.shape[0], lo.ctypes.data, hi.ctypes.data, cull_threshold, cull_leaf,
                                int(parallel), int(slow_div), C.byref(h))
        assert rc == 0
        self.h = h
        # shading side: the arrays of the same rrt_scene_desc the library would be given
        desc, keep = sd.desc()
        assert L.emu_scene_set_shading(h, sd.n_tris, C.cast(desc.vt, C.c_void_p), C.cast(desc.vn, C.c_void_p),
                                       C.cast(desc.material_id, C.c_void_p), desc.n_materials,
                                       C.cast(desc.materials, C.c_void_p), desc.n_textures,
                                       C.cast(desc.textures, C.c_void_p)) == 0
        del keep

    def ray_colours(self, params, o, d):
        o = np.ascontiguousarray(o, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(d, dtype=np.float64).reshape(-1, 3)
        out = np.zeros((o.shape[0], 3), dtype=np.uint8)
        assert self.L.emu_ray_colours(self.h, C.byref(params), o.shape[0], o.ctypes.data, d.ctypes.data, out.ctypes.data) == 0
        return out

    def trace(self, o, d, max_t=None, any_hit=False):
        o = np.ascontiguousarray(o, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(d, dtype=np.float64).reshape(-1, 3)
        n = o.shape[0]
        mt = None if max_t is None else np.ascontiguousarray(max_t, dtype=np.float64).reshape(n)
        tri = np.full(n, -1, dtype=np.int32)
        tuv = np.zeros((n, 3), dtype=np.float64)
        assert self.L.emu_trace_rays(self.h, n, o.ctypes.data, d.ctypes.data, None if mt is None else mt.ctypes.data,
                                     int(any_hit), tri.ctypes.data, tuv.ctypes.data) == 0
        return tri, tuv

    def render_shard(self, params, tile_rows=1, n_shards=1, shard_index=0, full_frame=None, want_hit_ids=True):
        """-> (fb, hit): the shard's rows compactly (rows, W), or written into `full_frame` (H, W) when given."""
        from rrt_b200.sharding import owned_rows
        W, H = params.width, params.height
        rows = H if n_shards <= 1 else len(owned_rows(H, tile_rows, n_shards, shard_index))
        fb = np.zeros((rows, W), dtype=np.uint32) if full_frame is None else full_frame
        hit = np.full((rows, W, 4), -7, dtype=np.int32) if want_hit_ids else None
        assert self.L.emu_render_shard(self.h, C.byref(params), tile_rows, n_shards, shard_index, int(full_frame is not None),
                                       fb.ctypes.data, None if hit is None else hit.ctypes.data) == 0
        return fb, hit

    def close(self):
        self.L.emu_scene_destroy(self.h)


def fuzz_rays(rng, n, centre, extent):
    """Random rays plus the awkward ones: zero / tiny / huge / non-finite direction components, origins exactly on
    octree planes and far outside the f32-culling eligibility range, finite and negative max_t."""
    o = rng.uniform(-extent, extent, (n, 3)) + centre
    d = rng.normal(size=(n, 3))
    for k in range(3):
        d[k::17, k] = 0.0                                   # one zero component
    d[5::41] = np.eye(3)[rng.integers(0, 3, len(d[5::41]))] * rng.choice([-1.0, 1.0], (len(d[5::41]), 1))  # axis-parallel
    d[7::53] *= 1e-12                                       # below the culling eligibility range
    d[11::59] *= 1e12                                       # above it
    d[13::61, 0] = 1e-300                                   # component whose reciprocal overflows the FMA-quotient range
    o[3::23] = np.round(o[3::23] / 2.5) * 2.5               # origins ON octree planes of the +-20 root box (k * 2.5)
    o[9::67] *= 1e4                                         # |origin| far above 16 x scene scale
    d[19::301, 1] = np.nan
    d[23::307, 2] = np.inf
    o[29::311, 0] = -np.inf
    mt = np.where(rng.random(n) < 0.5, np.inf, rng.uniform(0.05, 3.0 * extent, n))
    mt[31::97] = -1.0
    mt[37::101] = np.nan
    return o, d, mt


def check(emu_scene, oracle_scene, o, d, mt, min_hits):
    tri, tuv = emu_scene.trace(o, d, mt)
    otri, otuv = oracle_scene.trace(o, d, mt)
    bad = np.nonzero(tri != otri)[0]
    assert bad.size == 0, f"{bad.size} hit ids differ, first ray {bad[0]}: {tri[bad[0]]} vs oracle {otri[bad[0]]}"
    hit = tri >= 0
    assert np.array_equal(tuv[hit].view(np.uint64), otuv[hit].view(np.uint64))        # t, u, v bit for bit
    assert hit.sum() >= min_hits
    # shadow-ray mode: Some / None must agree (which triangle is reported is irrelevant to the caller)
    atri, _ = emu_scene.trace(o, d, mt, any_hit=True)
    assert np.array_equal(atri >= 0, otri >= 0)


def test_model2_fuzz(emu, scenes):
    from oracle.oracle import OracleScene
    sd = scenes("model2")
    rng = np.random.default_rng(2024)
    o, d, mt = fuzz_rays(rng, 400_000, np.array([0.0, 3.0, 0.0]), 9.0)
    o[1::29] = np.array([0.0, 2.0, -10.0])                  # the reference camera
    osc = OracleScene(sd)
    for kw in (dict(), dict(slow_div=True), dict(cull_threshold=0), dict(cull_threshold=4, cull_leaf=4), dict(parallel=True)):
        es = EmuScene(emu, sd, **kw)
        check(es, osc, o, d, mt, min_hits=30_000)
        es.close()
    osc.close()


@pytest.mark.parametrize("seed", range(12))
def test_random_soups_fuzz(emu, seed):
    # many small random scenes: different list lengths, depths, degenerate and duplicate triangles, different scales
    from oracle.oracle import OracleScene
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.choice([1, 2, 9, 60, 400, 3000]))
    size = float(rng.choice([0.02, 0.3, 1.5, 6.0]))
    sd = random_scene(n, seed=seed, extent=float(rng.choice([0.5, 6.0, 15.0])), size=size, kr=0.0)
    if n >= 9:
        sd.v[3] = sd.v[2]                                   # duplicate triangle: exact t tie, lower id wins
        sd.v[5, 3:6] = sd.v[5, 0:3]                         # degenerate (zero-length edge)
        sd.v[7] = np.round(sd.v[7])                         # vertices on integer coordinates (on octree planes)
    osc = OracleScene(sd)
    o, d, mt = fuzz_rays(rng, 60_000, np.array([0.0, 2.0, 4.0]), 12.0)
    # half of the rays aimed at triangle interiors so that small scenes are actually hit
    k = rng.integers(0, n, len(o) // 2)
    w = rng.dirichlet([1.0, 1.0, 1.0], len(k))
    target = (sd.v[k].reshape(-1, 3, 3) * w[:, :, None]).sum(1)
    sel = np.arange(0, len(o) - 1, 2)[:len(k)]
    ok = np.isfinite(o[sel]).all(1)
    d[sel[ok]] = (target[ok] - o[sel[ok]]) * rng.uniform(0.2, 3.0, (ok.sum(), 1))
    for kw in (dict(), dict(cull_threshold=0), dict(slow_div=True, cull_threshold=2, cull_leaf=2)):
        es = EmuScene(emu, sd, **kw)
        check(es, osc, o, d, mt, min_hits=1)
        es.close()
    osc.close()


def test_planar_grid_ties_and_axis_rays(emu):
    # a regular grid of coplanar triangles sharing edges and vertices, hit along edges and at vertices by axis-parallel
    # rays: exact t ties between neighbours (the lower id must win) and slab arithmetic full of 0 * inf
    from oracle.oracle import OracleScene
    g = 24
    xs = np.linspace(-6.0, 6.0, g + 1)
    tris = []
    for i in range(g):
        for j in range(g):
            a, b, c, e = (xs[i], xs[j], 5.0), (xs[i + 1], xs[j], 5.0), (xs[i], xs[j + 1], 5.0), (xs[i + 1], xs[j + 1], 5.0)
            tris += [[*a, *b, *c], [*b, *e, *c]]
    sd = tiny_scene(tris)
    osc = OracleScene(sd)
    es = EmuScene(emu, sd)
    rng = np.random.default_rng(5)
    n = 40_000
    gx, gy = rng.integers(0, g + 1, n), rng.integers(0, g + 1, n)
    o = np.stack([xs[gx], xs[gy], np.full(n, -3.0)], 1)
    o[::3, 0] += rng.uniform(0, 0.5, len(o[::3]))           # on an edge line, not at a vertex
    o[1::5] += rng.normal(scale=1e-9, size=(len(o[1::5]), 3))
    d = np.tile(np.array([0.0, 0.0, 1.0]), (n, 1))
    d[2::7] = np.array([0.0, 0.0, 2.5])
    check(es, osc, o, d, np.full(n, np.inf), min_hits=n // 2)
    es.close()
    osc.close()


def camera_rays(rng, n, origin, spread=0.6):
    o = np.tile(np.asarray(origin, dtype=np.float64), (n, 1))
    d = np.stack([rng.uniform(-spread, spread, n), rng.uniform(-spread, spread, n), np.ones(n)], 1)
    return o, d


@pytest.mark.parametrize("name", ["model2", "model", "model3"])
def test_ray_colours_shipped_scenes(emu, rrt, scenes, name):
    # the whole per-ray path of get_ray_colour (textures, bump maps, ambient / point / directional lights with their
    # shadow queries, the point-light `break`, mirror bounces with u8 re-quantisation per level) through the product's
    # RayStateT state machine vs the oracle's recursion: colours must be IDENTICAL
    from oracle.oracle import OracleScene
    sd = scenes(name)
    osc, es = OracleScene(sd), EmuScene(emu, sd)
    rng = np.random.default_rng(77)
    L = rrt.Light
    light_sets = [None, [], [L.Ambient(0.7)], [L.Point(0.9, (0.0, 9.0, -3.0)), L.Ambient(0.3)],
                  [L.Directional(0.8, (1.0, 2.0, -1.0)), L.Point(0.5, (3.0, 4.0, -8.0)), L.Point(0.5, (-3.0, 1.0, 20.0))]]
    total_lit = 0
    for i, lights in enumerate(light_sets):
        kw = {} if lights is None else dict(lights=lights)
        for max_depth in ((5, 0, 1) if name == "model2" and i in (0, 3) else (5,)):
            p = rrt.make_params(320, 180, max_depth=max_depth, **kw)
            o, d = camera_rays(rng, 40_000, (0.0, 2.0, -10.0))
            o2, d2, _ = fuzz_rays(rng, 20_000, np.array([0.0, 3.0, 0.0]), 9.0)
            o, d = np.concatenate([o, o2]), np.concatenate([d, d2])
            got, want = es.ray_colours(p, o, d), osc.ray_colours(p, o, d)
            bad = np.nonzero((got != want).any(1))[0]
            assert bad.size == 0, f"{name} lights#{i} depth {max_depth}: {bad.size} colours differ, first {got[bad[0]]} vs {want[bad[0]]}"
            total_lit += int((want != 255).any(1).sum())
    assert total_lit > 50_000
    es.close()
    osc.close()


def test_ray_colours_random_mirror_scene(emu, rrt):
    from oracle.oracle import OracleScene
    for seed, kr, bump in ((1, 0.3, False), (2, 0.9, True), (3, 0.0, True)):
        sd = random_scene(300, seed=seed, kr=kr)
        if bump:
            sd = tiny_scene(sd.v, kr=kr, bump=True)
        osc, es = OracleScene(sd), EmuScene(emu, sd)
        rng = np.random.default_rng(seed)
        p = rrt.make_params(200, 120, max_depth=5)
        o, d = camera_rays(rng, 60_000, (0.0, 2.0, -10.0), spread=0.9)
        got, want = es.ray_colours(p, o, d), osc.ray_colours(p, o, d)
        assert np.array_equal(got, want)
        assert (want != 255).any(1).sum() > 10_000
        es.close()
        osc.close()


def test_config4_sphere_grid_full_size(emu, rrt):
    # BASELINE configs[3]: the 999 936-triangle sphere grid at full size (parallel octree builder, deep tree, long
    # node-local lists behind culling hierarchies): traversal bits and colours of camera rays and awkward rays
    from oracle.oracle import OracleScene
    from rrt_b200.synthetic import sphere_grid
    sd = sphere_grid()
    osc, es = OracleScene(sd), EmuScene(emu, sd, parallel=True)
    rng = np.random.default_rng(4)
    o, d = camera_rays(rng, 150_000, (0.0, 2.0, -10.0), spread=0.55)
    o2, d2, mt2 = fuzz_rays(rng, 50_000, np.array([0.0, 2.0, 8.0]), 10.0)
    o, d = np.concatenate([o, o2]), np.concatenate([d, d2])
    mt = np.concatenate([np.full(150_000, np.inf), mt2])
    check(es, osc, o, d, mt, min_hits=60_000)
    p = rrt.make_params(3840, 2160)
    assert np.array_equal(es.ray_colours(p, o[:60_000], d[:60_000]), osc.ray_colours(p, o[:60_000], d[:60_000]))
    es.close()
    osc.close()


def test_config5_soup_reduced(emu, rrt):
    # BASELINE configs[4] generator (triangle soup, seed 12345) at 300 000 triangles — the full 10 M scene costs the
    # oracle ~0.1 s per ray; the 10 M case itself is checked on the GPU against an oracle lattice
    from oracle.oracle import OracleScene
    from rrt_b200.synthetic import triangle_soup
    sd = triangle_soup(n=300_000)
    osc, es = OracleScene(sd), EmuScene(emu, sd, parallel=True)
    rng = np.random.default_rng(5)
    o, d = camera_rays(rng, 20_000, (0.0, 2.0, -10.0), spread=0.8)
    o2, d2, mt2 = fuzz_rays(rng, 10_000, np.array([0.0, 0.0, 0.0]), 17.0)
    o, d = np.concatenate([o, o2]), np.concatenate([d, d2])
    mt = np.concatenate([np.full(20_000, np.inf), mt2])
    check(es, osc, o, d, mt, min_hits=10_000)
    p = rrt.make_params(7680, 4320)
    assert np.array_equal(es.ray_colours(p, o[:8_000], d[:8_000]), osc.ray_colours(p, o[:8_000], d[:8_000]))
    es.close()
    osc.close()


@pytest.mark.parametrize("scale", [1e-6, 1e-3, 1e3, 1e7])
def test_scene_scale_invariance_of_the_culling_budget(emu, scale):
    # the f32 culling boxes are padded and the ray eligibility is decided RELATIVE to the scene's coordinate scale
    # (rrt_octree.h kCullPadRel / kCullOriginRel): the same scene and rays scaled by 1e-6 .. 1e7 (root box included)
    # must still match the oracle bit for bit, with and without off-centre placement
    from oracle.oracle import OracleScene
    for shift in (0.0, 0.37):
        sd = random_scene(800, seed=11, extent=6.0, size=1.2, kr=0.0)
        sd.v = np.ascontiguousarray((sd.v + shift) * scale)
        sd.root_min = tuple((-20.0 + shift) * scale for _ in range(3))
        sd.root_max = tuple((20.0 + shift) * scale for _ in range(3))
        osc, es = OracleScene(sd), EmuScene(emu, sd)
        rng = np.random.default_rng(3)
        o, d, mt = fuzz_rays(rng, 80_000, np.array([0.0, 2.0, 4.0]), 12.0)
        k = rng.integers(0, sd.n_tris, len(o) // 2)
        w = rng.dirichlet([1.0, 1.0, 1.0], len(k))
        target = (sd.v[k].reshape(-1, 3, 3) * w[:, :, None]).sum(1) / scale
        sel = np.arange(0, len(o) - 1, 2)[:len(k)]
        ok = np.isfinite(o[sel]).all(1)
        d[sel[ok]] = target[ok] - o[sel[ok]]
        o = (o + shift) * scale
        d = d * rng.choice([1.0, scale, 1.0 / scale], (len(d), 1))
        check(es, osc, o, d, mt * scale, min_hits=5_000)
        es.close()
        osc.close()


@pytest.mark.parametrize("W,H", [(160, 90), (161, 91), (97, 64), (16, 4), (5, 3), (33, 2), (1, 1)])
def test_frames_through_the_kernel_helpers(emu, rrt, scenes, W, H):
    # whole frames the way render_kernel assembles them — queue position -> pixel / supersample (sample_pixel),
    # put_pixel mapping and the rows / columns the reference never writes (primary_ray), the packed per-pixel
    # accumulator and Color::mix (sample_add / mix_pack), all from csrc/rrt_raygen.cuh — against the oracle's
    # draw_scene: frame words and the depth-0 hit id of every supersample, odd and tiny sizes included
    from oracle.oracle import OracleScene
    sd = scenes("model2")
    osc, es = OracleScene(sd), EmuScene(emu, sd)
    origin, rot = rrt.orbit_camera(7)
    for kw in (dict(), dict(origin=origin, camera_rot=rot)):
        p = rrt.make_params(W, H, **kw)
        fb, hit = es.render_shard(p)
        ofb, ohit, _, _ = osc.render(p)
        assert np.array_equal(fb, ofb)
        assert np.array_equal(hit, ohit)
    es.close()
    osc.close()


@pytest.mark.parametrize("W,H,T,G", [(96, 70, 8, 2), (64, 45, 3, 3), (50, 37, 8, 8), (40, 9, 8, 4)])
def test_shards_through_the_kernel_helpers(emu, rrt, scenes, W, H, T, G):
    # interleaved tile-row shards (buffer_row): compact shards are the owned rows of the whole frame, and shards stored
    # straight into one full frame (rrt_render_device_into_frame's mode) reassemble it; ranks without rows included
    from oracle.oracle import OracleScene
    from rrt_b200.sharding import owned_rows
    sd = scenes("model2")
    osc, es = OracleScene(sd), EmuScene(emu, sd)
    p = rrt.make_params(W, H)
    ofb, ohit, _, _ = osc.render(p)
    frame = np.zeros((H, W), dtype=np.uint32)
    for r in range(G):
        rows = owned_rows(H, T, G, r)
        fb, hit = es.render_shard(p, T, G, r)
        assert np.array_equal(fb, ofb[rows]) and np.array_equal(hit, ohit[rows])
        es.render_shard(p, T, G, r, full_frame=frame, want_hit_ids=False)
    assert np.array_equal(frame, ofb)
    es.close()
    osc.close()


def test_walk_only_and_search_outcomes(emu, rrt, scenes):
    # the same answers with the searches switched off (RRT_SEARCH_MODE=0: the pure octree walk), and the searches do
    # answer most of the rays they are meant for: camera rays of the teapot scene, shadow-like rays towards a light
    from oracle.oracle import OracleScene
    sd = scenes("model2")
    rng = np.random.default_rng(77)
    o, d, mt = fuzz_rays(rng, 120_000, np.array([0.0, 3.0, 0.0]), 9.0)
    o[1::29] = np.array([0.0, 2.0, -10.0])
    osc = OracleScene(sd)
    es = EmuScene(emu, sd)
    try:
        emu.emu_set_allow(ALLOW_WALK_ONLY)
        check(es, osc, o, d, mt, min_hits=9_000)
        p = rrt.make_params(96, 54)
        fb, hit = es.render_shard(p)
        ofb, ohit, _, _ = osc.render(p)
        assert np.array_equal(hit, ohit) and np.array_equal(fb, ofb)
    finally:
        emu.emu_set_allow(ALLOW_ALL)
    # camera rays (engine.rs:207-211): direction (x/W, y/H, 1) from (0, 2, -10)
    n = 40_000
    cam_o = np.tile(np.array([0.0, 2.0, -10.0]), (n, 1))
    cam_d = np.stack([rng.uniform(-0.5, 0.5, n), rng.uniform(-0.5, 0.5, n), np.ones(n)], 1)
    out = np.zeros(3, dtype=np.uint64)
    assert emu.emu_search_outcomes(es.h, n, cam_o.ctypes.data, cam_d.ctypes.data, None, 0, out.ctypes.data) == 0
    assert out[2] == 0 and out[0] > 0.8 * n, out          # tried on every ray, answers most of them
    tri, _ = es.trace(cam_o, cam_d)
    otri, _ = osc.trace(cam_o, cam_d)
    assert np.array_equal(tri, otri) and (tri >= 0).mean() > 0.3
    es.close()
    osc.close()


@pytest.mark.parametrize("name", ["model2", "model", "soup"])
def test_grazing_rays_the_documented_exception_of_the_culling(emu, scenes, name):
    # DESIGN.md §5.4, assumption (A): the conservative f32 boxes only ever hide a record the reference accepts when the
    # ray runs IN that triangle's plane to within rounding — then the reference's own t is noise/noise (ray.rs:64-87: `a`
    # and `s.h` are both rounding error) and the "hit" it reports lies anywhere along the ray.  Adversarial rays built
    # in (or within 1e-16 .. 1e-8 rad of) triangle planes:
    #   * with the culling off (RRT_CULL_THRESHOLD=0: infinite boxes, same code) every ray matches the oracle;
    #   * with it on, fewer than 5 in 10 000 of these rays differ, and EVERY one of them is of that class: the
    #     oracle's hit triangle has |a| < 1e-11 |e1||e2||d| (rays 1e-9 rad or more off the plane never differ).
    from oracle.oracle import OracleScene
    if name == "soup":
        from rrt_b200.synthetic import triangle_soup
        sd = triangle_soup(40000, size=0.4)
    else:
        sd = scenes(name)
    rng = np.random.default_rng(2026)
    o, d, built_on = grazing_rays(sd, 150_000, rng)
    osc = OracleScene(sd)
    otri, otuv = osc.trace(o, d)
    assert (otri == built_on).sum() > 5_000                       # the construction does hit the triangles it aims at
    exact = EmuScene(emu, sd, cull_threshold=0)
    t0, tuv0 = exact.trace(o, d)
    assert np.array_equal(t0, otri)
    assert np.array_equal(tuv0[t0 >= 0].view(np.uint64), otuv[otri >= 0].view(np.uint64))
    exact.close()
    es = EmuScene(emu, sd)
    far = reference_determinant_rel(sd, built_on, d) >= 1e-9      # 1e-9 rad or more off the plane it was built on
    for allow in (ALLOW_ALL, ALLOW_WALK_ONLY):
        emu.emu_set_allow(allow)
        try:
            t, _ = es.trace(o, d)
        finally:
            emu.emu_set_allow(ALLOW_ALL)
        bad = np.nonzero(t != otri)[0]
        assert bad.size < 5e-4 * len(o), bad.size
        hidden = bad[otri[bad] >= 0]                              # the oracle's hit was hidden by a box
        assert (reference_determinant_rel(sd, otri[hidden], d[hidden]) < 1e-11).all()
        assert not far[bad].any()
    es.close()
    osc.close()
