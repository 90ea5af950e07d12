"""Pure-Python restatement of the reference's .obj/.mtl subset parser
(src/file_management/utils.rs:22-343).  TEST INFRASTRUCTURE: the product's loader is the C++
rrt_obj_load; this transcription exists to cross-check it."""
import os


class ParseError(Exception):
    """Stands for a Rust panic (expect/unwrap/assert) in the reference parser."""


def _f64(tok):
    try:
        return float(tok)            # Rust f64::from_str accepts the same decimal / inf / nan spellings
    except ValueError:
        raise ParseError(f"Could not parse value: {tok!r}")   # utils.rs:222


def _vertex(toks):
    """get_vertex utils.rs:228-234: x, y required; z defaults to 0.0."""
    if len(toks) < 2:
        raise ParseError("Cannot parse vertex")
    return (_f64(toks[0]), _f64(toks[1]), _f64(toks[2]) if len(toks) > 2 else 0.0)


def _coeff(toks):
    """get_color_coefficient_from_split_lines utils.rs:370-379."""
    r, g, b = _vertex(toks)
    if not (0.0 <= r <= 1.0 and 0.0 <= g <= 1.0 and 0.0 <= b <= 1.0):
        raise ParseError("All lighting intensity coefficients must be between 0.0 and 1.0")
    return (r, g, b)


def _usize(tok):
    if not tok.isdigit() and not (tok.startswith("+") and tok[1:].isdigit()):
        raise ParseError(f"Could not parse value: {tok!r}")
    return int(tok)


def _attrs(group):
    """get_vertex_attributes utils.rs:236-251: v[/vt[/vn]]; an EMPTY field (as in 1//2) panics."""
    parts = group.split("/")
    idx = _usize(parts[0])
    vt = _usize(parts[1]) if len(parts) > 1 else None
    vn = _usize(parts[2]) if len(parts) > 2 else None
    return idx, vt, vn


def parse_mtl(text, materials, tex_names):
    """parse_mtl_file_lines utils.rs:22-137. materials: dict name -> dict (insertion-ordered, re-insert
    overwrites like HashMap::insert); tex_names: list of texture file names in first-use order."""
    local = {}   # name_texture_map utils.rs:33 — per .mtl file; a new file re-decodes (new Arc)
    cur = None

    def tex_index(name):
        if name not in local:
            tex_names.append(name)           # material_map.textures.push utils.rs:92,106
            local[name] = len(tex_names) - 1
        return local[name]

    def blank():
        return dict(ka=None, kd=None, ks=None, ns=None, tex=None, bump=None, kr=None)
    st = blank()
    for line in text.splitlines() + ["END"]:
        toks = line.split()
        if not toks:
            continue
        k, rest = toks[0], toks[1:]
        if k in ("newmtl", "END"):
            if cur is not None:
                if st["tex"] is None:
                    raise ParseError("material without map_Ka")      # utils.rs:61 unwrap
                materials[cur] = dict(name=cur, ka=st["ka"] or (0.0, 0.0, 0.0), kd=st["kd"] or (0.0, 0.0, 0.0),
                                      ks=st["ks"] or (0.0, 0.0, 0.0), ns=240.0 if st["ns"] is None else st["ns"],
                                      kr=0.0 if st["kr"] is None else st["kr"], tex=st["tex"],
                                      bump=-1 if st["bump"] is None else st["bump"])
                st = blank()
            cur = rest[0] if rest else None
        elif k == "map_Ka":
            if not rest:
                raise ParseError("Expected a texture name")
            st["tex"] = tex_index(rest[0])
        elif k == "bump":
            if not rest:
                raise ParseError("Expected a texture name")
            st["bump"] = tex_index(rest[0])
        elif k == "Ka":
            st["ka"] = _coeff(rest)
        elif k == "Kd":
            st["kd"] = _coeff(rest)
        elif k == "Ks":
            st["ks"] = _coeff(rest)
        elif k == "Ns":
            if not rest:
                raise ParseError("Expected a valid Ns float value")
            st["ns"] = _f64(rest[0])
        elif k == "Kr":
            kr = _f64(rest[0]) if rest else 0.0
            st["kr"] = min(max(kr, 0.0), 1.0) if kr == kr else kr      # f64::clamp keeps NaN


def parse_obj(path, base_dir=None):
    """parse_obj_file_lines utils.rs:139-213. Returns dict(v, vt, vn, material_id (name list index),
    materials (list of dict), texture_names)."""
    base_dir = os.path.dirname(os.path.abspath(path)) if base_dir is None else base_dir
    verts, vts, vns = [], [], []
    materials, tex_names = {}, []
    tri_v, tri_vt, tri_vn, tri_mat = [], [], [], []
    current = None
    zero = (0.0, 0.0, 0.0)
    with open(path, "r") as fh:
        text = fh.read()
    for line in text.splitlines():
        toks = line.split()
        if not toks:
            continue
        k, rest = toks[0], toks[1:]
        if k == "mtllib":
            if not rest:
                raise ParseError("Invalid .mtl file name")
            try:
                with open(os.path.join(base_dir, rest[0]), "r") as fh:
                    mtl = fh.read()
            except OSError:
                raise ParseError("Could not read file")
            parse_mtl(mtl, materials, tex_names)
        elif k == "usemtl":
            if not rest:
                raise ParseError("Invalid material name")
            if rest[0] not in materials:
                raise ParseError("Material not found, is it in your mtl file?")
            current = dict(materials[rest[0]])      # Arc clone: later redefinitions do not affect it
        elif k == "v":
            verts.append(_vertex(rest))
        elif k == "vt":
            vts.append(_vertex(rest))
        elif k == "vn":
            vns.append(_vertex(rest))
        elif k == "f":
            if current is None:
                raise ParseError("face before usemtl")            # utils.rs:193 unwrap
            if len(rest) < 3:
                raise ParseError("No data for vertex")
            a = [_attrs(g) for g in rest[:3]]                     # only the first three groups utils.rs:258-263
            vv, tt, nn = [], [], []
            for idx, ti, ni in a:
                if idx < 1 or idx > len(verts):
                    raise ParseError("No vertex with this index")
                vv.extend(verts[idx - 1])
                if ti is not None and ti < 1:
                    raise ParseError("index underflow")           # usize 0 - 1 (debug panic / release: huge -> default)
                t = vts[ti - 1] if (ti is not None and ti <= len(vts)) else zero
                if ni is not None and ni < 1:
                    raise ParseError("index underflow")
                n = vns[ni - 1] if (ni is not None and ni <= len(vns)) else zero
                tt.extend(t[:2])
                nn.extend(n)
            tri_v.append(vv)
            tri_vt.append(tt)
            tri_vn.append(nn)
            tri_mat.append(current)
    # materials actually referenced keep identity by (name, contents)
    mats, mat_ids = [], []
    for m in tri_mat:
        for i, mm in enumerate(mats):
            if mm == m:
                mat_ids.append(i)
                break
        else:
            mats.append(m)
            mat_ids.append(len(mats) - 1)
    return dict(v=tri_v, vt=tri_vt, vn=tri_vn, material_id=mat_ids, materials=mats, texture_names=tex_names,
                all_materials=list(materials.values()))
