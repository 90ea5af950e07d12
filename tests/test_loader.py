"""The C++ .obj/.mtl loader (csrc/rrt_loader.cpp) against the pure-Python transcription of
src/file_management/utils.rs (oracle/py_obj.py), including the reference's defaults and panics."""
import os

import numpy as np
import pytest

from oracle import py_obj

MTL = """newmtl a
Ka 1 1 1
Kd 0.5 0.5 0.5
Ks 0.25 0.25 0.25
map_Ka t0.png
bump t1.png

newmtl b
Ns 10
Kr 7
map_Ka t0.png
"""
OBJ = """mtllib m.mtl
v 0 0 5
v 1 0 5
v 0 1 5
v 2 2
vt 0.5 0.25
vt 1 1 9
vn 0 0 -1
usemtl a
f 1/1/1 2/2/1 3/1/1 4/1/1
f 1 2 3
usemtl b
f 1/9/9 2/1 4/2/1
# comment
"""


def _write(tmp_path, obj=OBJ, mtl=MTL):
    (tmp_path / "m.mtl").write_text(mtl)
    p = tmp_path / "s.obj"
    p.write_text(obj)
    return str(p)


def _no_tex(name):
    return np.zeros((2, 2, 3), np.uint8)


def test_subset_semantics(rrt, tmp_path):
    from rrt_b200.loader import load_obj
    sd = load_obj(_write(tmp_path), texture_loader=_no_tex)
    ref = py_obj.parse_obj(_write(tmp_path))
    assert sd.n_tris == 3
    assert np.array_equal(sd.v, np.array(ref["v"])) and np.array_equal(sd.vt, np.array(ref["vt"]))
    assert np.array_equal(sd.vn, np.array(ref["vn"]))
    assert sd.texture_names == ref["texture_names"] == ["t0.png", "t1.png"]
    # quad: only the first three groups; vertex 4 has z defaulted to 0 (utils.rs:231)
    assert sd.v[0].tolist() == [0, 0, 5, 1, 0, 5, 0, 1, 5] and sd.v[2].tolist()[6:] == [2, 2, 0]
    # missing vt/vn and out-of-range vt/vn -> (0,0,0) (utils.rs:285-329)
    assert sd.vt[1].tolist() == [0] * 6 and sd.vn[1].tolist() == [0] * 9
    assert sd.vt[2].tolist() == [0, 0, 0.5, 0.25, 1, 1] and sd.vn[2].tolist() == [0] * 6 + [0, 0, -1]
    a, b = sd.materials[sd.material_id[0]], sd.materials[sd.material_id[2]]
    assert (a.name, a.ns, a.kr, a.tex_id, a.bump_id) == ("a", 240.0, 0.0, 0, 1)      # Ns default utils.rs:60
    assert (b.name, b.ns, b.kr, b.tex_id, b.bump_id) == ("b", 10.0, 1.0, 0, -1)      # Kr clamp utils.rs:131
    assert tuple(b.ka) == (0.0, 0.0, 0.0)                                            # utils.rs:54-59
    for i in range(3):
        m, pm = sd.materials[sd.material_id[i]], ref["materials"][ref["material_id"][i]]
        assert (m.name, tuple(m.ka), tuple(m.kd), tuple(m.ks), m.ns, m.kr, m.tex_id, m.bump_id) == \
               (pm["name"], pm["ka"], pm["kd"], pm["ks"], pm["ns"], pm["kr"], pm["tex"], pm["bump"])


@pytest.mark.parametrize("obj,mtl,msg", [
    (OBJ.replace("usemtl a", "usemtl zz"), MTL, "Material not found"),
    (OBJ.replace("f 1 2 3", "f 1 2 7"), MTL, "No vertex with this index"),
    (OBJ.replace("f 1 2 3", "f 1//1 2 3"), MTL, "Could not parse value"),       # empty vt field panics
    (OBJ.replace("v 2 2", "v 2"), MTL, "Cannot parse vertex"),
    (OBJ.replace("v 2 2", "v 2 x"), MTL, "Could not parse value"),
    (OBJ, MTL.replace("Ka 1 1 1", "Ka 1 1 1.5"), "between 0.0 and 1.0"),
    (OBJ, MTL.replace("map_Ka t0.png\nbump", "bump"), "no map_Ka"),            # utils.rs:61 unwrap
    (OBJ.replace("usemtl a\n", ""), MTL, "before usemtl"),
    (OBJ.replace("m.mtl", "missing.mtl"), MTL, "Could not read file"),
])
def test_reference_panics_become_errors(rrt, tmp_path, obj, mtl, msg):
    from rrt_b200.loader import load_obj
    p = _write(tmp_path, obj, mtl)
    with pytest.raises(rrt.abi.RrtError, match=msg):
        load_obj(p, texture_loader=_no_tex)
    with pytest.raises(py_obj.ParseError):
        py_obj.parse_obj(p)


def test_fixture_scenes_roundtrip(rrt, scenes, tmp_path):
    from rrt_b200.loader import load_scene_npz, save_scene_npz
    sd = scenes("model")
    assert sd.n_tris == 1004 and len(sd.textures) == 6 and sd.textures[0].shape == (1024, 1024, 3)
    save_scene_npz(tmp_path / "x.npz", sd)
    sd2 = load_scene_npz(tmp_path / "x.npz")
    assert np.array_equal(sd.v, sd2.v) and all(np.array_equal(a, b) for a, b in zip(sd.textures, sd2.textures))


@pytest.mark.skipif(not os.path.exists("/root/reference/model2.obj"), reason="reference assets not on this machine")
def test_fixtures_match_reference_assets(rrt, scenes):
    from rrt_b200.loader import load_obj
    for name in ("model", "model2", "model3"):
        a, b = load_obj(f"/root/reference/{name}.obj"), scenes(name)
        assert np.array_equal(a.v, b.v) and np.array_equal(a.vt, b.vt) and np.array_equal(a.vn, b.vn)
        assert np.array_equal(a.material_id, b.material_id) and a.texture_names == b.texture_names
        assert all(np.array_equal(x, y) for x, y in zip(a.textures, b.textures))
