// rrt_loader.cpp — the reference's .obj / .mtl subset (src/file_management/utils.rs:22-343) as host
// C++, plus a PPM writer standing in for the minifb window.  Where the reference panics
// (expect / unwrap / assert) this returns RRT_ERR_PARSE / RRT_ERR_IO with the reference's message.
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/rrt.h"

namespace {

thread_local std::string g_loader_err;

struct ParseFail {
    int code;
    std::string msg;
};

// str::split_whitespace
std::vector<std::string> tokens(const std::string& line) {
    std::vector<std::string> out;
    size_t i = 0, n = line.size();
    while (i < n) {
        while (i < n && std::isspace((unsigned char)line[i])) ++i;
        size_t j = i;
        while (j < n && !std::isspace((unsigned char)line[j])) ++j;
        if (j > i) out.push_back(line.substr(i, j - i));
        i = j;
    }
    return out;
}

// str::lines(): split on \n, strip one trailing \r
std::vector<std::string> lines_of(const std::string& text) {
    std::vector<std::string> out;
    size_t i = 0;
    while (i < text.size()) {
        size_t j = text.find('\n', i);
        if (j == std::string::npos) j = text.size();
        std::string l = text.substr(i, j - i);
        if (!l.empty() && l.back() == '\r') l.pop_back();
        out.push_back(l);
        i = j + 1;
    }
    return out;
}

// f64::from_str (utils.rs:222 "Could not parse value"): decimal / exponent forms, inf, nan; no hex,
// no leading/trailing junk.
double parse_f64(const std::string& s) {
    if (s.empty()) throw ParseFail{RRT_ERR_PARSE, "Could not parse value: ''"};
    for (char c : s)
        if (c == 'x' || c == 'X' || c == 'p' || c == 'P') throw ParseFail{RRT_ERR_PARSE, "Could not parse value: '" + s + "'"};
    char* end = nullptr;
    const double v = std::strtod(s.c_str(), &end);
    if (end != s.c_str() + s.size()) throw ParseFail{RRT_ERR_PARSE, "Could not parse value: '" + s + "'"};
    return v;
}
// usize::from_str: optional '+', digits only
uint64_t parse_usize(const std::string& s) {
    size_t i = (!s.empty() && s[0] == '+') ? 1 : 0;
    if (i >= s.size()) throw ParseFail{RRT_ERR_PARSE, "Could not parse value: '" + s + "'"};
    uint64_t v = 0;
    for (; i < s.size(); ++i) {
        if (s[i] < '0' || s[i] > '9') throw ParseFail{RRT_ERR_PARSE, "Could not parse value: '" + s + "'"};
        v = v * 10 + (uint64_t)(s[i] - '0');
    }
    return v;
}

struct Vec3 {
    double x, y, z;
};
// get_vertex utils.rs:228-234: x and y required, z defaults to 0.0
Vec3 get_vertex(const std::vector<std::string>& t, size_t from) {
    if (t.size() < from + 2) throw ParseFail{RRT_ERR_PARSE, "Cannot parse vertex"};
    Vec3 v{parse_f64(t[from]), parse_f64(t[from + 1]), 0.0};
    if (t.size() > from + 2) v.z = parse_f64(t[from + 2]);
    return v;
}
// get_color_coefficient_from_split_lines utils.rs:370-379
Vec3 get_coeff(const std::vector<std::string>& t) {
    Vec3 v = get_vertex(t, 1);
    if (!(v.x <= 1.0 && v.y <= 1.0 && v.z <= 1.0 && v.x >= 0.0 && v.y >= 0.0 && v.z >= 0.0))
        throw ParseFail{RRT_ERR_PARSE, "All lighting intensity coefficients must be between 0.0 and 1.0"};
    return v;
}

struct MatRec {
    std::string name;
    rrt_material m;
};

}  // namespace

struct rrt_obj {
    std::vector<double> v, vt, vn;
    std::vector<int32_t> material_id;
    std::vector<MatRec> mats;          // one entry per completed `newmtl` block, in file order
    std::vector<rrt_material> mats_c;  // contiguous copy for the C accessor
    std::vector<std::string> tex_names;

    std::map<std::string, int> by_name;  // HashMap<String, Arc<Material>>: latest definition wins (utils.rs:66-68)

    // parse_mtl_file_lines utils.rs:22-137
    void parse_mtl(const std::string& text) {
        std::map<std::string, int> local_tex;  // name_texture_map utils.rs:33 (per call)
        bool have_name = false;
        std::string name;
        struct {
            bool ka = false, kd = false, ks = false, ns = false, kr = false;
            Vec3 a{0, 0, 0}, d{0, 0, 0}, s{0, 0, 0};
            double nsv = 240.0, krv = 0.0;
            int tex = -1, bump = -1;
        } st, blank;
        auto tex_index = [&](const std::string& n) {
            auto it = local_tex.find(n);
            if (it != local_tex.end()) return it->second;
            tex_names.push_back(n);  // material_map.textures.push utils.rs:92,106
            const int id = (int)tex_names.size() - 1;
            local_tex[n] = id;
            return id;
        };
        std::vector<std::string> ls = lines_of(text);
        ls.push_back("END");  // utils.rs:45
        for (const std::string& line : ls) {
            const std::vector<std::string> t = tokens(line);
            if (t.empty()) continue;
            const std::string& k = t[0];
            if (k == "newmtl" || k == "END") {
                if (have_name) {
                    if (st.tex < 0) throw ParseFail{RRT_ERR_PARSE, "material '" + name + "' has no map_Ka (utils.rs:61 unwrap)"};
                    MatRec r;
                    r.name = name;
                    r.m = rrt_material{{st.a.x, st.a.y, st.a.z}, {st.d.x, st.d.y, st.d.z}, {st.s.x, st.s.y, st.s.z},
                                       st.ns ? st.nsv : 240.0, st.kr ? st.krv : 0.0, st.tex, st.bump};
                    mats.push_back(r);
                    by_name[name] = (int)mats.size() - 1;
                    st = blank;
                }
                have_name = t.size() > 1;  // utils.rs:80-81
                if (have_name) name = t[1];
            } else if (k == "map_Ka") {
                if (t.size() < 2) throw ParseFail{RRT_ERR_PARSE, "Expected a texture name"};
                st.tex = tex_index(t[1]);
            } else if (k == "bump") {
                if (t.size() < 2) throw ParseFail{RRT_ERR_PARSE, "Expected a texture name"};
                st.bump = tex_index(t[1]);
            } else if (k == "Ka") {
                st.a = get_coeff(t); st.ka = true;
            } else if (k == "Kd") {
                st.d = get_coeff(t); st.kd = true;
            } else if (k == "Ks") {
                st.s = get_coeff(t); st.ks = true;
            } else if (k == "Ns") {
                if (t.size() < 2) throw ParseFail{RRT_ERR_PARSE, "Expected a valid Ns float value"};
                st.nsv = parse_f64(t[1]); st.ns = true;
            } else if (k == "Kr") {
                double kr = t.size() > 1 ? parse_f64(t[1]) : 0.0;  // utils.rs:130
                if (kr == kr) kr = kr < 0.0 ? 0.0 : (kr > 1.0 ? 1.0 : kr);  // f64::clamp keeps NaN
                st.krv = kr; st.kr = true;
            }
        }
    }

    // parse_obj_file_lines utils.rs:139-213
    void parse_obj(const std::string& text, const std::string& base_dir) {
        std::vector<Vec3> verts, vts, vns;
        int current = -1;
        for (const std::string& line : lines_of(text)) {
            const std::vector<std::string> t = tokens(line);
            if (t.empty()) continue;
            const std::string& k = t[0];
            if (k == "mtllib") {
                if (t.size() < 2) throw ParseFail{RRT_ERR_PARSE, "Invalid .mtl file name"};
                const std::string path = base_dir.empty() ? t[1] : base_dir + "/" + t[1];
                std::ifstream f(path, std::ios::binary);
                if (!f) throw ParseFail{RRT_ERR_IO, "Could not read file: " + path};
                std::stringstream ss;
                ss << f.rdbuf();
                parse_mtl(ss.str());
            } else if (k == "usemtl") {
                if (t.size() < 2) throw ParseFail{RRT_ERR_PARSE, "Invalid material name"};
                auto it = by_name.find(t[1]);
                if (it == by_name.end()) throw ParseFail{RRT_ERR_PARSE, "Material not found, is it in your mtl file?"};
                current = it->second;
            } else if (k == "v") {
                verts.push_back(get_vertex(t, 1));
            } else if (k == "vt") {
                vts.push_back(get_vertex(t, 1));
            } else if (k == "vn") {
                vns.push_back(get_vertex(t, 1));
            } else if (k == "f") {
                if (current < 0) throw ParseFail{RRT_ERR_PARSE, "face before usemtl (utils.rs:193 unwrap)"};
                if (t.size() < 4) throw ParseFail{RRT_ERR_PARSE, "No data for vertex"};
                for (int c = 0; c < 3; ++c) {  // only the first three groups (utils.rs:258-263)
                    // get_vertex_attributes utils.rs:236-251: split("/"); an empty field fails to parse
                    std::vector<std::string> parts;
                    const std::string& g = t[1 + c];
                    size_t i = 0;
                    for (;;) {
                        size_t j = g.find('/', i);
                        if (j == std::string::npos) { parts.push_back(g.substr(i)); break; }
                        parts.push_back(g.substr(i, j - i));
                        i = j + 1;
                    }
                    const uint64_t vi = parse_usize(parts[0]);
                    if (vi < 1 || vi > verts.size()) throw ParseFail{RRT_ERR_PARSE, "No vertex with this index"};
                    const Vec3 p = verts[vi - 1];
                    Vec3 tc{0, 0, 0}, nc{0, 0, 0};  // DEFAULT_VERTICES utils.rs:15-19
                    if (parts.size() > 1) {
                        const uint64_t ti = parse_usize(parts[1]);
                        if (ti < 1) throw ParseFail{RRT_ERR_PARSE, "texture index 0 (usize underflow)"};
                        if (ti <= vts.size()) tc = vts[ti - 1];  // .get(..).unwrap_or(DEFAULT) utils.rs:290-293
                    }
                    if (parts.size() > 2) {
                        const uint64_t ni = parse_usize(parts[2]);
                        if (ni < 1) throw ParseFail{RRT_ERR_PARSE, "normal index 0 (usize underflow)"};
                        if (ni <= vns.size()) nc = vns[ni - 1];
                    }
                    v.insert(v.end(), {p.x, p.y, p.z});
                    vt.insert(vt.end(), {tc.x, tc.y});
                    vn.insert(vn.end(), {nc.x, nc.y, nc.z});
                }
                material_id.push_back(current);
            }
        }
        mats_c.clear();
        for (const MatRec& r : mats) mats_c.push_back(r.m);
    }
};

extern "C" {

// rrt_last_error lives in rrt_api.cu; loader errors are routed through this setter.
void rrt_set_last_error_(const char* msg);

int rrt_obj_load(const char* obj_path, const char* base_dir, rrt_obj** out) {
    if (!obj_path || !out) { rrt_set_last_error_("null argument"); return RRT_ERR_INVALID; }
    *out = nullptr;
    std::ifstream f(obj_path, std::ios::binary);
    if (!f) { rrt_set_last_error_((std::string("Could not read file: ") + obj_path).c_str()); return RRT_ERR_IO; }
    std::stringstream ss;
    ss << f.rdbuf();
    auto o = std::make_unique<rrt_obj>();
    try {
        o->parse_obj(ss.str(), base_dir ? std::string(base_dir) : std::string());
    } catch (const ParseFail& e) {
        rrt_set_last_error_(e.msg.c_str());
        return e.code;
    }
    *out = o.release();
    return RRT_OK;
}
void rrt_obj_destroy(rrt_obj* o) { delete o; }
uint64_t rrt_obj_n_tris(const rrt_obj* o) { return o ? o->material_id.size() : 0; }
const double* rrt_obj_v(const rrt_obj* o) { return o->v.data(); }
const double* rrt_obj_vt(const rrt_obj* o) { return o->vt.data(); }
const double* rrt_obj_vn(const rrt_obj* o) { return o->vn.data(); }
const int32_t* rrt_obj_material_id(const rrt_obj* o) { return o->material_id.data(); }
uint32_t rrt_obj_n_materials(const rrt_obj* o) { return o ? (uint32_t)o->mats.size() : 0; }
const rrt_material* rrt_obj_materials(const rrt_obj* o) { return o->mats_c.data(); }
const char* rrt_obj_material_name(const rrt_obj* o, uint32_t i) { return i < o->mats.size() ? o->mats[i].name.c_str() : ""; }
uint32_t rrt_obj_n_textures(const rrt_obj* o) { return o ? (uint32_t)o->tex_names.size() : 0; }
const char* rrt_obj_texture_name(const rrt_obj* o, uint32_t i) { return i < o->tex_names.size() ? o->tex_names[i].c_str() : ""; }

// Headless stand-in for Canvas::update (engine.rs:162-166): binary PPM of the 0x00RRGGBB buffer.
int rrt_write_ppm(const char* path, const uint32_t* fb, int32_t width, int32_t height) {
    if (!path || !fb || width <= 0 || height <= 0) { rrt_set_last_error_("bad argument"); return RRT_ERR_INVALID; }
    FILE* f = std::fopen(path, "wb");
    if (!f) { rrt_set_last_error_("cannot open output file"); return RRT_ERR_IO; }
    std::fprintf(f, "P6\n%d %d\n255\n", width, height);
    std::vector<unsigned char> row((size_t)width * 3);
    for (int y = 0; y < height; ++y) {
        for (int x = 0; x < width; ++x) {
            const uint32_t c = fb[(size_t)y * width + x];
            row[3 * x] = (c >> 16) & 255; row[3 * x + 1] = (c >> 8) & 255; row[3 * x + 2] = c & 255;
        }
        std::fwrite(row.data(), 1, row.size(), f);
    }
    std::fclose(f);
    return RRT_OK;
}

}  // extern "C"
