"""ctypes binding of oracle/liboracle.so — the CPU restatement of the reference path.
TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's CPU legs."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")


class OracleCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in
                ("n_a", "n_u", "n_v", "n_t", "n_slab", "n_node", "rays_primary", "rays_shadow", "rays_reflect")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


def build(force=False):
    src = os.path.join(_HERE, "rrt_oracle.cpp")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        L.oracle_last_error.restype = C.c_char_p
        L.oracle_octree_node_count.restype = C.c_uint64
        L.oracle_octree_node_count.argtypes = [C.c_void_p]
        L.oracle_octree_node_list.restype = C.c_int64
        L.oracle_octree_node_list.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]
        L.oracle_octree_node.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_double), C.POINTER(C.c_uint64),
                                         C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.oracle_scene_destroy.argtypes = [C.c_void_p]
        L.oracle_scene_destroy.restype = None
        L.oracle_scene_create.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.oracle_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                    C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_double)]
        for f in (L.oracle_trace_rays, L.oracle_trace_bruteforce):
            f.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_ray_colours.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}: {lib().oracle_last_error().decode()}")


class OracleScene:
    """Reference SceneData + Octree built by sequential insertion (utils.rs:192-198, octree.rs:41-108)."""

    def __init__(self, scene_data):
        d, keep = scene_data.desc()
        h = C.c_void_p()
        _check(lib().oracle_scene_create(C.byref(d), C.byref(h)))
        self._h = h
        self.scene_data = scene_data

    def close(self):
        if getattr(self, "_h", None):
            lib().oracle_scene_destroy(self._h)
            self._h = None

    __del__ = close

    def render(self, params, want_hit_ids=True, n_threads=0, row_begin=0, row_end=None, row_step=1, col_step=1,
               counters=False):
        """Scene::draw_scene (engine.rs:186-255). -> (fb (H,W) u32, hit (H,W,4) i32 | None, seconds, counters|None)"""
        W, H = params.width, params.height
        fb = np.zeros((H, W), dtype=np.uint32)
        hit = np.full((H, W, 4), -1, dtype=np.int32) if want_hit_ids else None
        cnt = OracleCounters() if counters else None
        sec = C.c_double(0.0)
        _check(lib().oracle_render(self._h, C.byref(params), fb.ctypes.data, None if hit is None else hit.ctypes.data,
                                   n_threads, row_begin, H if row_end is None else row_end, row_step, col_step,
                                   C.byref(cnt) if cnt is not None else None, C.byref(sec)))
        return fb, hit, sec.value, (cnt.as_dict() if cnt is not None else None)

    def _trace(self, fn, origins, directions, max_t):
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(directions, dtype=np.float64).reshape(-1, 3)
        n = o.shape[0]
        mt = None if max_t is None else np.ascontiguousarray(max_t, dtype=np.float64).reshape(n)
        tri = np.full(n, -1, dtype=np.int32)
        tuv = np.zeros((n, 3), dtype=np.float64)
        _check(fn(self._h, n, o.ctypes.data, d.ctypes.data, None if mt is None else mt.ctypes.data,
                  tri.ctypes.data, tuv.ctypes.data))
        return tri, tuv

    def trace(self, origins, directions, max_t=None):
        """Ray::intersect_with_octant_with_max_t (ray.rs:104-168), recursive."""
        return self._trace(lib().oracle_trace_rays, origins, directions, max_t)

    def trace_bruteforce(self, origins, directions, max_t=None):
        return self._trace(lib().oracle_trace_bruteforce, origins, directions, max_t)

    def ray_colours(self, params, origins, directions):
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(directions, dtype=np.float64).reshape(-1, 3)
        out = np.zeros((o.shape[0], 3), dtype=np.uint8)
        _check(lib().oracle_ray_colours(self._h, C.byref(params), o.shape[0], o.ctypes.data, d.ctypes.data,
                                        out.ctypes.data))
        return out

    # ---- octree inspection
    def node_count(self):
        return int(lib().oracle_octree_node_count(self._h))

    def node(self, i):
        box = (C.c_double * 6)()
        tc, fc, ll = C.c_uint64(), C.c_uint64(), C.c_uint64()
        _check(lib().oracle_octree_node(self._h, i, box, C.byref(tc), C.byref(fc), C.byref(ll)))
        return list(box), int(tc.value), int(fc.value), int(ll.value)

    def node_list(self, i):
        n = int(lib().oracle_octree_node_list(self._h, i, None, 0))
        out = np.zeros(max(n, 1), dtype=np.uint32)
        lib().oracle_octree_node_list(self._h, i, out.ctypes.data, n)
        return out[:n].tolist()
