"""Host-side mirror of the reference's render-call surface, on top of the C ABI (include/rrt.h).

Names follow the reference so callers (and the parity tests) read like src/main.rs:
    SceneData   src/scene/scenedata.rs:6-13      Light      src/scene/entities.rs:5-9
    RayTracer   src/scene/raytracer.rs:22-31     Scene      src/scene/engine.rs:171-255
    Canvas      src/scene/engine.rs:122-167 (headless: a numpy u32 buffer, no minifb window)
All rendering goes through librrt_b200.so (CUDA); nothing here computes pixels on the CPU.
"""
import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

from . import abi


@dataclass
class MaterialSpec:
    """Material  src/scene/material.rs:11-22 (defaults: src/file_management/utils.rs:54-63)."""
    name: str = ""
    ka: Sequence[float] = (0.0, 0.0, 0.0)
    kd: Sequence[float] = (0.0, 0.0, 0.0)
    ks: Sequence[float] = (0.0, 0.0, 0.0)
    ns: float = 240.0
    kr: float = 0.0
    tex_id: int = 0
    bump_id: int = -1


@dataclass
class SceneData:
    """Flat triangle arrays in `f`-line order + materials + decoded RGB8 textures."""
    v: np.ndarray                    # (n, 9) f64
    vt: np.ndarray                   # (n, 6) f64
    vn: np.ndarray                   # (n, 9) f64
    material_id: np.ndarray          # (n,) i32
    materials: List[MaterialSpec]
    textures: List[np.ndarray]       # each (h, w, 3) u8
    root_min: Sequence[float] = (-20.0, -20.0, -20.0)   # utils.rs:145
    root_max: Sequence[float] = (20.0, 20.0, 20.0)
    texture_names: List[str] = field(default_factory=list)

    def __post_init__(self):
        self.v = np.ascontiguousarray(self.v, dtype=np.float64).reshape(-1, 9)
        self.vt = np.ascontiguousarray(self.vt, dtype=np.float64).reshape(-1, 6)
        self.vn = np.ascontiguousarray(self.vn, dtype=np.float64).reshape(-1, 9)
        self.material_id = np.ascontiguousarray(self.material_id, dtype=np.int32).reshape(-1)
        self.textures = [np.ascontiguousarray(t, dtype=np.uint8) for t in self.textures]
        n = self.v.shape[0]
        if not (self.vt.shape[0] == n and self.vn.shape[0] == n and self.material_id.shape[0] == n):
            raise ValueError("triangle arrays disagree on n")

    @property
    def n_tris(self):
        return self.v.shape[0]

    def desc(self):
        """-> (abi.SceneDesc, keepalive list). Pointers stay valid while keepalive lives."""
        d = abi.SceneDesc()
        dp = C.POINTER(C.c_double)
        d.n_tris = self.n_tris
        d.v = self.v.ctypes.data_as(dp)
        d.vt = self.vt.ctypes.data_as(dp)
        d.vn = self.vn.ctypes.data_as(dp)
        d.material_id = self.material_id.ctypes.data_as(C.POINTER(C.c_int32))
        mats = (abi.Material * max(1, len(self.materials)))()
        for i, m in enumerate(self.materials):
            mats[i].ka[:] = list(m.ka)
            mats[i].kd[:] = list(m.kd)
            mats[i].ks[:] = list(m.ks)
            mats[i].ns, mats[i].kr = m.ns, m.kr
            mats[i].tex_id, mats[i].bump_id = m.tex_id, m.bump_id
        texs = (abi.Texture * max(1, len(self.textures)))()
        for i, t in enumerate(self.textures):
            if t.ndim != 3 or t.shape[2] != 3:
                raise ValueError("textures must be (h, w, 3) uint8")
            texs[i].height, texs[i].width = t.shape[0], t.shape[1]
            texs[i].rgb = t.ctypes.data_as(C.POINTER(C.c_uint8))
        d.n_materials, d.materials = len(self.materials), mats
        d.n_textures, d.textures = len(self.textures), texs
        d.root_min[:] = list(self.root_min)
        d.root_max[:] = list(self.root_max)
        return d, [mats, texs, self]


class Light:
    """Light enum  src/scene/entities.rs:5-9."""

    def __init__(self, kind, intensity, vec=(0.0, 0.0, 0.0)):
        self.kind, self.intensity, self.vec = kind, float(intensity), tuple(float(x) for x in vec)

    @staticmethod
    def Ambient(intensity):
        return Light(abi.LIGHT_AMBIENT, intensity)

    @staticmethod
    def Point(intensity, position):
        return Light(abi.LIGHT_POINT, intensity, position)

    @staticmethod
    def Directional(intensity, direction):
        return Light(abi.LIGHT_DIRECTIONAL, intensity, direction)


def default_lights():
    """The four lights main() hard-codes, src/main.rs:32-58."""
    return [Light.Ambient(0.5), Light.Point(0.4, (-7.0, 1.0, -15.0)), Light.Point(0.5, (0.0, 1.0, -41.0)),
            Light.Directional(0.4, (-5.0, 0.0, 20.0))]


DEFAULT_ORIGIN = (0.0, 2.0, -10.0)   # src/main.rs:62-66


def make_params(width=800, height=800, origin=DEFAULT_ORIGIN, lights=None, viewport=(1.0, 1.0, 1.0),
                max_depth=5, surface_offset=1e-4, camera_rot=None):
    """abi.RenderParams with the reference's constants (src/main.rs:14-15,32-66; engine.rs:113-119;
    raytracer.rs:17,20). Pure Python so the oracle harness can build it without the CUDA library."""
    lights = default_lights() if lights is None else lights
    if len(lights) > abi.RRT_MAX_LIGHTS:
        raise ValueError("too many lights")
    p = abi.RenderParams()
    p.width, p.height = int(width), int(height)
    p.origin[:] = [float(x) for x in origin]
    p.viewport[:] = [float(x) for x in viewport]
    p.n_lights, p.max_depth = len(lights), int(max_depth)
    for i, l in enumerate(lights):
        p.lights[i].kind, p.lights[i].intensity = l.kind, l.intensity
        p.lights[i].vec[:] = list(l.vec)
    p.surface_offset = float(surface_offset)
    if camera_rot is None:
        p.use_camera_rot = 0
        p.camera_rot[:] = [1.0, 0, 0, 0, 1.0, 0, 0, 0, 1.0]
    else:
        p.use_camera_rot = 1
        p.camera_rot[:] = [float(x) for x in np.asarray(camera_rot, dtype=np.float64).reshape(9)]
    return p


def orbit_camera(k, n_frames=64, radius=10.0, height=2.0):
    """Config-5 orbit (an EXTENSION — the reference has no rotation): frame k looks at the
    y axis from (r sin th, h, -r cos th); rotation about +y by th. k=0 is the reference camera."""
    th = 2.0 * np.pi * k / n_frames
    origin = (radius * np.sin(th), height, -radius * np.cos(th))
    c, s = np.cos(th), np.sin(th)
    rot = np.array([[c, 0.0, -s], [0.0, 1.0, 0.0], [s, 0.0, c]])
    return origin, (None if k % n_frames == 0 else rot)


class RayTracer:
    """RayTracer { scene_data, lights, origin }  src/scene/raytracer.rs:22-26.
    Owns the device copy of the scene (rrt_scene_create builds the reference octree on the host)."""

    def __init__(self, scene_data: SceneData, lights: Optional[List[Light]] = None, origin=DEFAULT_ORIGIN,
                 device: int = 0):
        self.scene_data, self.origin = scene_data, tuple(origin)
        self.lights = default_lights() if lights is None else lights
        self.device = device
        L = abi.lib()
        d, keep = scene_data.desc()
        h = C.c_void_p()
        abi.check(L.rrt_scene_create(C.byref(d), device, C.byref(h)))
        del keep
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            abi.lib().rrt_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:      # interpreter shutdown
            pass

    @property
    def handle(self):
        return self._h

    def stats(self):
        s = abi.SceneStats()
        abi.check(abi.lib().rrt_scene_get_stats(self._h, C.byref(s)))
        return s.as_dict()

    def params(self, width, height, **kw):
        kw.setdefault("origin", self.origin)
        kw.setdefault("lights", self.lights)
        return make_params(width, height, **kw)

    def get_ray_colours(self, origins, directions, params=None):
        """Batched RayTracer::get_ray_colour (raytracer.rs:29-31) -> (n,3) u8."""
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(directions, dtype=np.float64).reshape(-1, 3)
        out = np.zeros((o.shape[0], 3), dtype=np.uint8)
        p = params or self.params(2, 2)
        abi.check(abi.lib().rrt_ray_colours(self._h, C.byref(p), o.shape[0], o.ctypes.data, d.ctypes.data,
                                            out.ctypes.data))
        return out

    def get_ray_colour(self, origin, direction):
        return tuple(int(x) for x in self.get_ray_colours([origin], [direction])[0])

    def intersect_with_octant_with_max_t(self, origins, directions, max_t=None):
        """Batched Ray::intersect_with_octant_with_max_t(octree, 0, max_t) (ray.rs:104-168).
        -> (tri (n,) i32 with -1 = None, tuv (n,3) f64)."""
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(directions, dtype=np.float64).reshape(-1, 3)
        n = o.shape[0]
        mt = None if max_t is None else np.ascontiguousarray(max_t, dtype=np.float64).reshape(n)
        tri = np.full(n, -1, dtype=np.int32)
        tuv = np.zeros((n, 3), dtype=np.float64)
        abi.check(abi.lib().rrt_trace_rays(self._h, n, o.ctypes.data, d.ctypes.data,
                                           None if mt is None else mt.ctypes.data, tri.ctypes.data, tuv.ctypes.data))
        return tri, tuv


class Canvas:
    """Headless Canvas (engine.rs:122-167): `buffer` is u32 0x00RRGGBB, row-major, row 0 never written."""

    def __init__(self, width, height):
        self.width, self.height = width, height
        self.buffer = np.zeros(width * height, dtype=np.uint32)

    def write_ppm(self, path):
        abi.check(abi.lib().rrt_write_ppm(str(path).encode(), self.buffer.ctypes.data, self.width, self.height))


class Scene:
    """Scene::new(width, height) / Scene::draw_scene(rt)  src/scene/engine.rs:171-255."""

    def __init__(self, width, height):
        self.canvas = Canvas(width, height)
        self.hit_ids = None
        self.device_ms = None

    def reserve(self, rt: RayTracer):
        """Allocate now what the first draw_scene of this canvas size would allocate inside the call (rrt_scene_reserve)."""
        p = rt.params(self.canvas.width, self.canvas.height)
        abi.check(abi.lib().rrt_scene_reserve(rt.handle, C.byref(p), None, 1))

    def draw_scene(self, rt: RayTracer, want_hit_ids=False, **param_overrides):
        W, H = self.canvas.width, self.canvas.height
        p = rt.params(W, H, **param_overrides)
        hit = np.full((H, W, 4), -1, dtype=np.int32) if want_hit_ids else None
        ms = C.c_float(0.0)
        abi.check(abi.lib().rrt_render(rt.handle, C.byref(p), self.canvas.buffer.ctypes.data,
                                       None if hit is None else hit.ctypes.data, C.byref(ms)))
        self.hit_ids, self.device_ms = hit, float(ms.value)
        return self.canvas.buffer.reshape(H, W)

    def draw_frames(self, rt: RayTracer, overrides_per_frame, out=None):
        """A sequence of frames of this resolution (rrt_render_frames): frame i uses rt.params(W, H, **overrides_per_frame[i])
        and lands in out[i] while frame i+1 renders.  `out`: (n, H, W) u32 array (pinned memory makes the copies
        asynchronous); allocated if None.  Returns out; self.device_ms = device time of the whole sequence."""
        W, H = self.canvas.width, self.canvas.height
        n = len(overrides_per_frame)
        if out is None:
            out = np.zeros((n, H, W), dtype=np.uint32)
        if out.dtype != np.uint32 or out.shape != (n, H, W) or not out.flags["C_CONTIGUOUS"]:
            raise ValueError("out must be a C-contiguous (n, H, W) uint32 array")
        params = (abi.RenderParams * max(n, 1))()
        for i, ov in enumerate(overrides_per_frame):
            params[i] = rt.params(W, H, **ov)
        ptrs = (C.c_void_p * max(n, 1))(*[out[i].ctypes.data for i in range(n)])
        ms = C.c_float(0.0)
        abi.check(abi.lib().rrt_render_frames(rt.handle, params, n, ptrs, C.byref(ms)))
        self.device_ms = float(ms.value)
        if n:
            self.canvas.buffer = out[n - 1].reshape(-1)
        return out
