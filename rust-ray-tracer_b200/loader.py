"""Scene loading on top of the C++ parser (csrc/rrt_loader.cpp, restating
src/file_management/utils.rs:22-343).  Textures are decoded here with PIL — the reference uses the
Rust `image` crate (utils.rs:345-368); JPEG decoders may differ by an LSB or two, see DESIGN.md."""
import ctypes as C
import io
import os

import numpy as np

from . import abi
from .scene import MaterialSpec, SceneData


def decode_texture(data_or_path):
    """-> (h, w, 3) uint8, the `img.as_bytes().chunks(3)` of utils.rs:353-361 (8-bit RGB only)."""
    from PIL import Image
    src = io.BytesIO(data_or_path) if isinstance(data_or_path, (bytes, bytearray)) else data_or_path
    img = Image.open(src)
    if img.mode != "RGB":
        # the reference reinterprets whatever the decoder returns as packed RGB8; anything else would
        # silently mis-stride there, so refuse it here
        raise ValueError(f"texture mode {img.mode!r}: the reference path assumes 8-bit RGB")
    return np.ascontiguousarray(np.asarray(img, dtype=np.uint8))


def load_obj(path, base_dir=None, texture_loader=None):
    """parse_obj_file_lines (utils.rs:139-213) -> SceneData. `mtllib` and texture names resolve against
    base_dir (default: the .obj's directory; the reference resolves against the cwd)."""
    path = os.fspath(path)
    base_dir = os.path.dirname(os.path.abspath(path)) if base_dir is None else os.fspath(base_dir)
    L = abi.lib()
    h = C.c_void_p()
    abi.check(L.rrt_obj_load(path.encode(), base_dir.encode(), C.byref(h)))
    try:
        n = int(L.rrt_obj_n_tris(h))
        v = np.ctypeslib.as_array(L.rrt_obj_v(h), shape=(n, 9)).copy() if n else np.zeros((0, 9))
        vt = np.ctypeslib.as_array(L.rrt_obj_vt(h), shape=(n, 6)).copy() if n else np.zeros((0, 6))
        vn = np.ctypeslib.as_array(L.rrt_obj_vn(h), shape=(n, 9)).copy() if n else np.zeros((0, 9))
        mid = np.ctypeslib.as_array(L.rrt_obj_material_id(h), shape=(n,)).copy() if n else np.zeros(0, np.int32)
        nm = int(L.rrt_obj_n_materials(h))
        mp = L.rrt_obj_materials(h)
        mats = []
        for i in range(nm):
            m = mp[i]
            mats.append(MaterialSpec(L.rrt_obj_material_name(h, i).decode(), tuple(m.ka), tuple(m.kd), tuple(m.ks),
                                     m.ns, m.kr, m.tex_id, m.bump_id))
        names = [L.rrt_obj_texture_name(h, i).decode() for i in range(int(L.rrt_obj_n_textures(h)))]
    finally:
        L.rrt_obj_destroy(h)
    load = texture_loader or (lambda name: decode_texture(os.path.join(base_dir, name)))
    texs = [load(nm_) for nm_ in names]
    return SceneData(v, vt, vn, mid, mats, texs, texture_names=names)


def save_scene_npz(path, sd, embed_textures=True):
    """Compact fixture: triangle arrays + materials (+ decoded textures unless they live in a shared
    texture store, see save_texture_store)."""
    mats = np.array([[*m.ka, *m.kd, *m.ks, m.ns, m.kr, m.tex_id, m.bump_id] for m in sd.materials], dtype=np.float64)
    d = dict(v=sd.v, vt=sd.vt, vn=sd.vn, material_id=sd.material_id, materials=mats.reshape(len(sd.materials), 13),
             material_names=np.array([m.name for m in sd.materials]),
             texture_names=np.array(sd.texture_names), root_min=np.array(sd.root_min), root_max=np.array(sd.root_max))
    if embed_textures:
        for i, t in enumerate(sd.textures):
            d[f"tex_{i}"] = t
    np.savez_compressed(path, **d)


def save_texture_store(path, files):
    """files: {name: encoded file bytes}. Kept ENCODED (small); decoded again by load_scene_npz."""
    np.savez(path, **{k: np.frombuffer(v, dtype=np.uint8) for k, v in files.items()})


def load_scene_npz(path, texture_store=None):
    z = np.load(path, allow_pickle=False)
    names = [str(x) for x in z["texture_names"]]
    store = None
    texs = []
    for i, nm in enumerate(names):
        if f"tex_{i}" in z:
            texs.append(z[f"tex_{i}"])
        else:
            if store is None:
                sp = texture_store or os.path.join(os.path.dirname(os.path.abspath(path)), "textures.npz")
                store = np.load(sp, allow_pickle=False)
            texs.append(decode_texture(store[nm].tobytes()))
    mats = [MaterialSpec(str(nm), tuple(r[0:3]), tuple(r[3:6]), tuple(r[6:9]), float(r[9]), float(r[10]), int(r[11]), int(r[12]))
            for nm, r in zip(z["material_names"], z["materials"])]
    return SceneData(z["v"], z["vt"], z["vn"], z["material_id"], mats, texs, tuple(z["root_min"]), tuple(z["root_max"]),
                     texture_names=names)
