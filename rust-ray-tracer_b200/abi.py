"""ctypes mirror of include/rrt.h and the loader for librrt_b200.so.

The library is the product: if it is missing or cannot be loaded this module raises — there is no
Python / CPU fallback for any rendering call.
"""
import ctypes as C
import os

RRT_MAX_LIGHTS = 16
LIGHT_AMBIENT, LIGHT_POINT, LIGHT_DIRECTIONAL = 0, 1, 2

_HERE = os.path.dirname(os.path.abspath(__file__))
# RRT_LIB_PATH: development override (tools/variant_sweep.py times differently-built copies of the library)
LIB_PATH = os.environ.get("RRT_LIB_PATH") or os.path.join(_HERE, "csrc", "librrt_b200.so")


class RrtError(RuntimeError):
    pass


class Texture(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgb", C.POINTER(C.c_uint8))]


class Material(C.Structure):
    _fields_ = [("ka", C.c_double * 3), ("kd", C.c_double * 3), ("ks", C.c_double * 3),
                ("ns", C.c_double), ("kr", C.c_double), ("tex_id", C.c_int32), ("bump_id", C.c_int32)]


class SceneDesc(C.Structure):
    _fields_ = [("n_tris", C.c_uint64),
                ("v", C.POINTER(C.c_double)), ("vt", C.POINTER(C.c_double)), ("vn", C.POINTER(C.c_double)),
                ("material_id", C.POINTER(C.c_int32)),
                ("n_materials", C.c_uint32), ("materials", C.POINTER(Material)),
                ("n_textures", C.c_uint32), ("textures", C.POINTER(Texture)),
                ("root_min", C.c_double * 3), ("root_max", C.c_double * 3)]


class Light(C.Structure):
    _fields_ = [("kind", C.c_int32), ("_pad", C.c_int32), ("intensity", C.c_double), ("vec", C.c_double * 3)]


class RenderParams(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32),
                ("origin", C.c_double * 3), ("viewport", C.c_double * 3),
                ("n_lights", C.c_int32), ("max_depth", C.c_int32),
                ("lights", Light * RRT_MAX_LIGHTS),
                ("surface_offset", C.c_double),
                ("use_camera_rot", C.c_int32), ("_pad", C.c_int32),
                ("camera_rot", C.c_double * 9)]


class Shard(C.Structure):
    _fields_ = [("tile_rows", C.c_int32), ("n_shards", C.c_int32), ("shard_index", C.c_int32)]


class SceneStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in
                ("n_tris", "n_tris_stored", "n_tris_dropped", "n_nodes", "n_nodes_nonempty", "max_depth",
                 "root_list", "max_list", "n_cull_nodes", "device_bytes", "flat_digest")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in
                ("rays_primary", "rays_shadow", "rays_reflect", "mt_tests", "box_tests", "cull_tests",
                 "nodes_entered")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


RRT_NOALLOC_MAX_PIXELS = 1 << 21   # include/rrt.h

# every symbol include/rrt.h declares (tests check the .so exports all of them)
SYMBOLS = [
    "rrt_abi_version", "rrt_last_error", "rrt_device_count", "rrt_render_params_default",
    "rrt_scene_create", "rrt_scene_destroy", "rrt_scene_get_stats", "rrt_scene_order_hint",
    "rrt_scene_search_mode", "rrt_scene_search_state",
    "rrt_scene_reserve", "rrt_render", "rrt_render_frames", "rrt_render_device", "rrt_render_device_noalloc", "rrt_shard_rows", "rrt_deinterleave_device",
    "rrt_render_device_into_frame", "rrt_frame_alloc", "rrt_frame_free", "rrt_ipc_get_handle", "rrt_ipc_open_handle",
    "rrt_ipc_close_handle", "rrt_copy_frame_to_host",
    "rrt_host_frame_create", "rrt_host_frame_open", "rrt_host_frame_pixels", "rrt_host_frame_geometry",
    "rrt_host_frame_pin", "rrt_host_frame_barrier", "rrt_host_frame_publish_handle", "rrt_host_frame_fetch_handle",
    "rrt_host_frame_close", "rrt_render_to_host_frame", "rrt_copy_shard_to_host_frame", "rrt_synchronize",
    "rrt_trace_rays", "rrt_ray_colours", "rrt_render_counters",
    "rrt_octree_build", "rrt_octree_destroy", "rrt_octree_node_count", "rrt_octree_node", "rrt_octree_node_list",
    "rrt_octree_flatten_stats", "rrt_octree_digest",
    "rrt_obj_load", "rrt_obj_destroy", "rrt_obj_n_tris", "rrt_obj_v", "rrt_obj_vt", "rrt_obj_vn",
    "rrt_obj_material_id", "rrt_obj_n_materials", "rrt_obj_materials", "rrt_obj_material_name",
    "rrt_obj_n_textures", "rrt_obj_texture_name", "rrt_write_ppm", "rrt_microbench_fp64", "rrt_selftest_division",
]

_lib = None


def lib():
    """Load librrt_b200.so (built in-tree by __graft_entry__.build()). Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RrtError(f"{LIB_PATH} not built: run `python __graft_entry__.py build` "
                       "(the CUDA library is required; there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, u64, i32, dp = C.c_void_p, C.c_uint64, C.c_int32, C.POINTER(C.c_double)
    L.rrt_abi_version.restype = C.c_int
    L.rrt_last_error.restype = C.c_char_p
    L.rrt_device_count.restype = C.c_int
    L.rrt_render_params_default.argtypes = [C.POINTER(RenderParams)]
    L.rrt_render_params_default.restype = None
    L.rrt_scene_create.argtypes = [C.POINTER(SceneDesc), C.c_int, C.POINTER(vp)]
    L.rrt_scene_destroy.argtypes = [vp]
    L.rrt_scene_destroy.restype = None
    L.rrt_scene_get_stats.argtypes = [vp, C.POINTER(SceneStats)]
    L.rrt_scene_order_hint.argtypes = [vp, C.c_int]
    L.rrt_scene_search_mode.argtypes = [vp, C.c_int, C.c_int, C.c_int]
    L.rrt_scene_search_state.argtypes = [vp, C.POINTER(C.c_uint32)]
    L.rrt_scene_reserve.argtypes = [vp, C.POINTER(RenderParams), C.POINTER(Shard), C.c_int]
    L.rrt_render.argtypes = [vp, C.POINTER(RenderParams), vp, vp, C.POINTER(C.c_float)]
    L.rrt_render_frames.argtypes = [vp, C.POINTER(RenderParams), C.c_int32, C.POINTER(vp), C.POINTER(C.c_float)]
    L.rrt_render_device.argtypes = [vp, C.POINTER(RenderParams), C.POINTER(Shard), vp, vp, vp]
    L.rrt_render_device_noalloc.argtypes = [vp, C.POINTER(RenderParams), C.POINTER(Shard), vp, vp, vp]
    L.rrt_render_device_into_frame.argtypes = [vp, C.POINTER(RenderParams), C.POINTER(Shard), vp, vp]
    L.rrt_frame_alloc.argtypes = [C.c_int, i32, i32, C.POINTER(vp)]
    L.rrt_frame_free.argtypes = [C.c_int, vp]
    L.rrt_ipc_get_handle.argtypes = [C.c_int, vp, vp]
    L.rrt_ipc_open_handle.argtypes = [C.c_int, vp, C.POINTER(vp)]
    L.rrt_ipc_close_handle.argtypes = [C.c_int, vp]
    L.rrt_copy_frame_to_host.argtypes = [C.c_int, vp, i32, i32, vp, vp]
    L.rrt_host_frame_create.argtypes = [C.c_char_p, i32, i32, i32, C.POINTER(vp)]
    L.rrt_host_frame_open.argtypes = [C.c_char_p, C.c_double, C.POINTER(vp)]
    L.rrt_host_frame_pixels.argtypes = [vp]
    L.rrt_host_frame_pixels.restype = vp
    L.rrt_host_frame_geometry.argtypes = [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]
    L.rrt_host_frame_pin.argtypes = [vp, C.c_int]
    L.rrt_host_frame_barrier.argtypes = [vp, C.c_double]
    L.rrt_host_frame_publish_handle.argtypes = [vp, vp]
    L.rrt_host_frame_fetch_handle.argtypes = [vp, C.c_double, vp]
    L.rrt_host_frame_close.argtypes = [vp]
    L.rrt_host_frame_close.restype = None
    L.rrt_render_to_host_frame.argtypes = [vp, C.POINTER(RenderParams), C.POINTER(Shard), vp, vp]
    L.rrt_synchronize.argtypes = [vp, vp]
    L.rrt_copy_shard_to_host_frame.argtypes = [C.c_int, vp, i32, i32, C.POINTER(Shard), vp, vp]
    L.rrt_shard_rows.argtypes = [i32, C.POINTER(Shard)]
    L.rrt_deinterleave_device.argtypes = [C.c_int, vp, i32, i32, i32, i32, i32, vp, vp]
    L.rrt_trace_rays.argtypes = [vp, u64, vp, vp, vp, vp, vp]
    L.rrt_ray_colours.argtypes = [vp, C.POINTER(RenderParams), u64, vp, vp, vp]
    L.rrt_render_counters.argtypes = [vp, C.POINTER(RenderParams), vp, C.POINTER(Counters)]
    L.rrt_octree_build.argtypes = [vp, u64, dp, dp, C.POINTER(vp)]
    L.rrt_octree_destroy.argtypes = [vp]
    L.rrt_octree_destroy.restype = None
    L.rrt_octree_node_count.argtypes = [vp]
    L.rrt_octree_node_count.restype = u64
    L.rrt_octree_node.argtypes = [vp, u64, dp, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]
    L.rrt_octree_digest.argtypes = [vp]
    L.rrt_octree_digest.restype = u64
    L.rrt_octree_flatten_stats.argtypes = [vp, vp, C.POINTER(SceneStats), dp]
    L.rrt_octree_node_list.argtypes = [vp, u64, vp, u64]
    L.rrt_octree_node_list.restype = C.c_int64
    L.rrt_obj_load.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(vp)]
    L.rrt_obj_destroy.argtypes = [vp]
    L.rrt_obj_destroy.restype = None
    L.rrt_obj_n_tris.argtypes = [vp]
    L.rrt_obj_n_tris.restype = u64
    for name, rt in (("rrt_obj_v", dp), ("rrt_obj_vt", dp), ("rrt_obj_vn", dp),
                     ("rrt_obj_material_id", C.POINTER(C.c_int32)), ("rrt_obj_materials", C.POINTER(Material))):
        getattr(L, name).argtypes = [vp]
        getattr(L, name).restype = rt
    L.rrt_obj_n_materials.argtypes = [vp]
    L.rrt_obj_n_materials.restype = C.c_uint32
    L.rrt_obj_n_textures.argtypes = [vp]
    L.rrt_obj_n_textures.restype = C.c_uint32
    L.rrt_obj_material_name.argtypes = [vp, C.c_uint32]
    L.rrt_obj_material_name.restype = C.c_char_p
    L.rrt_obj_texture_name.argtypes = [vp, C.c_uint32]
    L.rrt_obj_texture_name.restype = C.c_char_p
    L.rrt_write_ppm.argtypes = [C.c_char_p, vp, i32, i32]
    L.rrt_microbench_fp64.argtypes = [C.c_int, dp, dp]
    L.rrt_selftest_division.argtypes = [C.c_int, u64, C.POINTER(u64)]
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise RrtError(f"rrt error {rc}: {lib().rrt_last_error().decode()}")
