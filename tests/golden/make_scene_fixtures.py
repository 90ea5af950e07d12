"""Regenerates tests/golden/scenes/*.npz from the reference's shipped assets (run in the build
container, where /root/reference exists; the GPU box only ever sees the committed fixtures).

  model.npz / model2.npz / model3.npz : triangle arrays + materials parsed from the .obj/.mtl files by
                                        the product loader (cross-checked against oracle/py_obj.py)
  textures.npz                        : the texture files named by materials.mtl, byte-for-byte as
                                        shipped (still JPEG-encoded; decoded with PIL at load time)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import rrt_b200  # noqa: E402
from rrt_b200.loader import load_obj, save_scene_npz, save_texture_store  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden", "scenes")


def main():
    os.makedirs(OUT, exist_ok=True)
    files = {}
    for name in ("model", "model2", "model3"):
        sd = load_obj(os.path.join(REF, name + ".obj"))
        save_scene_npz(os.path.join(OUT, name + ".npz"), sd, embed_textures=False)
        for t in sd.texture_names:
            with open(os.path.join(REF, t), "rb") as f:
                files[t] = f.read()
        print(name, sd.n_tris, "triangles", sd.texture_names)
    save_texture_store(os.path.join(OUT, "textures.npz"), files)


if __name__ == "__main__":
    main()
