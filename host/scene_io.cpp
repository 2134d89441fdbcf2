// scene_io.cpp -- scene JSON, OBJ/MTL, camera and light-list construction for the headless host.
// See scene_io.h for the reference contracts each function mirrors.
#include "scene_io.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <fstream>
#include <array>
#include <cmath>
#include <limits>
#include <map>
#include <set>
#include <sstream>

namespace orbhost {

// ------------------------------------------------------------------------------------------------
// A small JSON reader (objects, arrays, strings, numbers, true/false/null) -- the scene schema has
// seven keys, src/scene.cpp:38-98.
// ------------------------------------------------------------------------------------------------
struct JVal {
    enum Kind { NUL, BOOL, NUM, STR, ARR, OBJ } kind = NUL;
    double num = 0;
    bool b = false;
    std::string str;
    std::vector<JVal> arr;
    std::vector<std::pair<std::string, JVal>> obj;
    const JVal* get(const char* key) const {
        for (auto& kv : obj)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
};

struct JParser {
    const char* p;
    const char* end;
    std::string err;
    void ws() {
        while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p;
    }
    bool fail(const char* m) {
        if (err.empty()) err = m;
        return false;
    }
    bool str(std::string& out) {
        if (p >= end || *p != '"') return fail("expected string");
        ++p;
        out.clear();
        while (p < end && *p != '"') {
            if (*p == '\\' && p + 1 < end) {
                ++p;
                switch (*p) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': {  // keep ASCII, replace the rest
                        if (p + 4 < end) {
                            unsigned v = (unsigned)strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16);
                            out += v < 128 ? (char)v : '?';
                            p += 4;
                        }
                    } break;
                    default: out += *p;
                }
                ++p;
            } else {
                out += *p++;
            }
        }
        if (p >= end) return fail("unterminated string");
        ++p;
        return true;
    }
    bool val(JVal& v, int depth = 0) {
        if (depth > 64) return fail("nesting too deep");
        ws();
        if (p >= end) return fail("unexpected end");
        if (*p == '{') {
            ++p;
            v.kind = JVal::OBJ;
            ws();
            if (p < end && *p == '}') { ++p; return true; }
            for (;;) {
                ws();
                std::string k;
                if (!str(k)) return false;
                ws();
                if (p >= end || *p != ':') return fail("expected ':'");
                ++p;
                JVal c;
                if (!val(c, depth + 1)) return false;
                v.obj.emplace_back(k, std::move(c));
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == '}') { ++p; return true; }
                return fail("expected ',' or '}'");
            }
        }
        if (*p == '[') {
            ++p;
            v.kind = JVal::ARR;
            ws();
            if (p < end && *p == ']') { ++p; return true; }
            for (;;) {
                JVal c;
                if (!val(c, depth + 1)) return false;
                v.arr.push_back(std::move(c));
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == ']') { ++p; return true; }
                return fail("expected ',' or ']'");
            }
        }
        if (*p == '"') {
            v.kind = JVal::STR;
            return str(v.str);
        }
        if (!strncmp(p, "true", 4) && end - p >= 4) { v.kind = JVal::BOOL; v.b = true; p += 4; return true; }
        if (!strncmp(p, "false", 5) && end - p >= 5) { v.kind = JVal::BOOL; v.b = false; p += 5; return true; }
        if (!strncmp(p, "null", 4) && end - p >= 4) { v.kind = JVal::NUL; p += 4; return true; }
        char* e = nullptr;
        v.num = strtod(p, &e);
        if (e == p) return fail("unexpected character");
        v.kind = JVal::NUM;
        p = e;
        return true;
    }
};

static bool read_file(const std::string& path, std::string& out) {
    std::ifstream f(path.c_str(), std::ios::binary);
    if (!f) return false;
    std::stringstream ss;
    ss << f.rdbuf();
    out = ss.str();
    return true;
}

static std::string dir_of(const std::string& path) {
    size_t k = path.rfind('/');
    return k == std::string::npos ? std::string() : path.substr(0, k + 1);
}

static bool vec3_of(const JVal* v, float out[3]) {
    if (!v || v->kind != JVal::ARR || v->arr.size() < 3) return false;
    for (int i = 0; i < 3; i++) {
        if (v->arr[i].kind != JVal::NUM) return false;
        out[i] = (float)v->arr[i].num;
    }
    return true;
}

// ------------------------------------------------------------------------------------------------
// OBJ + MTL  (src/Model.cpp:130-252)
// ------------------------------------------------------------------------------------------------
struct Material {
    std::string name;
    float kd[3] = {0.6f, 0.6f, 0.6f};  // tinyobj's defaults are 0; a missing Kd is unusual either way
    float ke[3] = {0, 0, 0};
    float ns = 1.f;
    std::string map_kd;
    bool has_kd = false;
};

static bool load_mtl(const std::string& path, std::vector<Material>& mats, std::map<std::string, int>& by_name) {
    std::ifstream f(path.c_str());
    if (!f) return false;
    std::string line;
    Material* cur = nullptr;
    while (std::getline(f, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        std::istringstream ss(line);
        std::string tok;
        if (!(ss >> tok) || tok[0] == '#') continue;
        if (tok == "newmtl") {
            Material m;
            ss >> m.name;
            m.kd[0] = m.kd[1] = m.kd[2] = 0.f;
            by_name[m.name] = (int)mats.size();
            mats.push_back(m);
            cur = &mats.back();
        } else if (cur) {
            if (tok == "Kd") { ss >> cur->kd[0] >> cur->kd[1] >> cur->kd[2]; cur->has_kd = true; }
            else if (tok == "Ke") ss >> cur->ke[0] >> cur->ke[1] >> cur->ke[2];
            else if (tok == "Ns") ss >> cur->ns;
            else if (tok == "map_Kd") {
                std::string rest;
                std::getline(ss, rest);
                size_t a = rest.find_first_not_of(" \t");
                cur->map_kd = a == std::string::npos ? std::string() : rest.substr(a);
            }
        }
    }
    return true;
}

bool load_ppm_rgba8(const std::string& path, std::vector<uint8_t>& rgba, int& w, int& h) {
    std::string s;
    if (!read_file(path, s) || s.size() < 11 || s[0] != 'P' || s[1] != '6') return false;
    size_t p = 2;
    int vals[3], n = 0;
    while (n < 3 && p < s.size()) {
        while (p < s.size() && isspace((unsigned char)s[p])) ++p;
        if (p < s.size() && s[p] == '#') {
            while (p < s.size() && s[p] != '\n') ++p;
            continue;
        }
        size_t q = p;
        while (q < s.size() && isdigit((unsigned char)s[q])) ++q;
        if (q == p) return false;
        vals[n++] = atoi(s.substr(p, q - p).c_str());
        p = q;
    }
    if (n < 3 || vals[2] != 255) return false;
    ++p;  // the single whitespace after maxval
    w = vals[0];
    h = vals[1];
    if (w <= 0 || h <= 0 || s.size() < p + (size_t)3 * w * h) return false;
    rgba.resize((size_t)4 * w * h);
    // stored bottom row first, like the y-mirror at Model.cpp:108-116
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            const unsigned char* src = (const unsigned char*)&s[p + 3 * ((size_t)(h - 1 - y) * w + x)];
            uint8_t* dst = &rgba[4 * ((size_t)y * w + x)];
            dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2]; dst[3] = 255;
        }
    return true;
}

struct Corner { int v, t, n; };

static bool parse_corner(const std::string& s, int nv, int nt, int nn, Corner& c) {
    // v, v/t, v//n, v/t/n ; negative = relative to the end
    c.v = c.t = c.n = 0;
    int part = 0;
    size_t i = 0;
    while (part < 3) {
        size_t j = s.find('/', i);
        std::string f = s.substr(i, j == std::string::npos ? std::string::npos : j - i);
        if (!f.empty()) {
            int k = atoi(f.c_str());
            int cnt = part == 0 ? nv : (part == 1 ? nt : nn);
            if (k < 0) k = cnt + k + 1;
            (part == 0 ? c.v : (part == 1 ? c.t : c.n)) = k;
        }
        if (j == std::string::npos) break;
        i = j + 1;
        ++part;
    }
    return c.v >= 1 && c.v <= nv && c.t >= 1 && c.t <= nt && c.n >= 1 && c.n <= nn;
}

// A polygon's triangles as the reference's OBJ reader makes them (the tiny_obj_loader.h it vendors, LoadObj with
// triangulate = true): ear clipping in the coordinate plane picked from the polygon's first non-degenerate corner.
// A convex planar polygon comes out as the fan (0, k, k+1); a concave or non-planar one does not, and the order of the
// triangles and of their corners decides the un-indexed vertex arrays (and through them the primitive ids and the
// light list), so the procedure is followed step by step -- tests/golden/obj_ingest.json pins it on the reference's
// own loader.  `V` = positions read so far (corner indices are 1-based and already absolute).
static void triangulate_polygon(const std::vector<Corner>& face, const std::vector<float>& V,
                                std::vector<std::array<Corner, 3>>& tris) {
    size_t n = face.size();
    auto P = [&](const Corner& c, size_t axis) { return V[3 * (size_t)(c.v - 1) + axis]; };
    size_t ax[2] = {1, 2};
    const float eps = std::numeric_limits<float>::epsilon();
    for (size_t k = 0; k < n; k++) {
        const Corner &a = face[k % n], &b = face[(k + 1) % n], &c = face[(k + 2) % n];
        const float e0x = P(b, 0) - P(a, 0), e0y = P(b, 1) - P(a, 1), e0z = P(b, 2) - P(a, 2);
        const float e1x = P(c, 0) - P(b, 0), e1y = P(c, 1) - P(b, 1), e1z = P(c, 2) - P(b, 2);
        const float cx = std::fabs(e0y * e1z - e0z * e1y), cy = std::fabs(e0z * e1x - e0x * e1z), cz = std::fabs(e0x * e1y - e0y * e1x);
        if (cx > eps || cy > eps || cz > eps) {  // a corner: drop the axis its normal points along most
            if (!(cx > cy && cx > cz)) {
                ax[0] = 0;
                if (cz > cx && cz > cy) ax[1] = 1;
            }
            break;
        }
    }
    float area = 0;
    for (size_t k = 0; k < n; k++) {
        const Corner &a = face[k], &b = face[(k + 1) % n];
        area += (P(a, ax[0]) * P(b, ax[1]) - P(a, ax[1]) * P(b, ax[0])) * 0.5f;
    }
    std::vector<Corner> rest = face;
    size_t guess = 0, tries = n, before = n;
    while (rest.size() > 3 && tries > 0) {
        n = rest.size();
        if (guess >= n) guess -= n;
        if (before != n) { before = n; tries = n; }  // a vertex went: start counting again
        else tries--;
        Corner ind[3];
        float vx[3], vy[3];
        for (size_t k = 0; k < 3; k++) {
            ind[k] = rest[(guess + k) % n];
            vx[k] = P(ind[k], ax[0]);
            vy[k] = P(ind[k], ax[1]);
        }
        const float e0x = vx[1] - vx[0], e0y = vy[1] - vy[0], e1x = vx[2] - vx[1], e1y = vy[2] - vy[1];
        const float cross = e0x * e1y - e0y * e1x;
        if (cross * area < 0.0f) {  // a reflex corner
            guess += 1;
            continue;
        }
        bool overlap = false;  // another vertex inside this candidate ear (crossing-number test)
        for (size_t o = 3; o < n && !overlap; o++) {
            const Corner& q = rest[(guess + o) % n];
            const float tx = P(q, ax[0]), ty = P(q, ax[1]);
            int inside = 0;
            for (int i = 0, j = 2; i < 3; j = i++)
                if (((vy[i] > ty) != (vy[j] > ty)) && (tx < (vx[j] - vx[i]) * (ty - vy[i]) / (vy[j] - vy[i]) + vx[i])) inside = !inside;
            overlap = inside != 0;
        }
        if (overlap) {
            guess += 1;
            continue;
        }
        tris.push_back({ind[0], ind[1], ind[2]});
        rest.erase(rest.begin() + (long)((guess + 1) % n));  // the ear's tip leaves the polygon
    }
    if (rest.size() == 3) tris.push_back({rest[0], rest[1], rest[2]});
}

bool load_obj(const std::string& obj_file, std::vector<Mesh>& out, std::string* err) {
    std::ifstream f(obj_file.c_str());
    if (!f) {
        if (err) *err = "Could not read OBJ model from " + obj_file;
        return false;
    }
    const std::string base = dir_of(obj_file);
    std::vector<float> V, N, T;
    std::vector<Material> mats;
    std::map<std::string, int> mat_by_name;
    struct Face { Corner c[3]; int mat; };
    struct Shape { std::string name; std::vector<Face> faces; };
    std::vector<Shape> shapes(1);
    int cur_mat = -1;
    std::string line;
    int line_no = 0;
    while (std::getline(f, line)) {
        ++line_no;
        if (!line.empty() && line.back() == '\r') line.pop_back();
        std::istringstream ss(line);
        std::string tok;
        if (!(ss >> tok) || tok[0] == '#') continue;
        if (tok == "v") { float x, y, z; ss >> x >> y >> z; V.push_back(x); V.push_back(y); V.push_back(z); }
        else if (tok == "vn") { float x, y, z; ss >> x >> y >> z; N.push_back(x); N.push_back(y); N.push_back(z); }
        else if (tok == "vt") { float u = 0, v = 0; ss >> u >> v; T.push_back(u); T.push_back(v); }
        else if (tok == "mtllib") {
            std::string name;
            ss >> name;
            if (!load_mtl(base + name, mats, mat_by_name)) {
                if (err) *err = "could not read material library " + base + name;
                return false;
            }
        } else if (tok == "usemtl") {
            std::string name;
            ss >> name;
            auto it = mat_by_name.find(name);
            cur_mat = it == mat_by_name.end() ? -1 : it->second;
        } else if (tok == "g" || tok == "o") {  // a new shape starts (tinyobj splits on both)
            if (!shapes.back().faces.empty()) shapes.emplace_back();
            ss >> shapes.back().name;
        } else if (tok == "f") {
            std::vector<Corner> cs;
            std::string s;
            while (ss >> s) {
                Corner c;
                if (!parse_corner(s, (int)V.size() / 3, (int)T.size() / 2, (int)N.size() / 3, c)) {
                    // the reference dereferences normal_index and texcoord_index unconditionally (Model.cpp:193-201)
                    if (err) {
                        char b[256];
                        snprintf(b, sizeof b, "%s:%d: every face corner needs valid v/vt/vn indices", obj_file.c_str(), line_no);
                        *err = b;
                    }
                    return false;
                }
                cs.push_back(c);
            }
            if (cs.size() < 3) continue;
            if (cur_mat < 0) {
                if (err) *err = obj_file + ": face without a material (the reference indexes materials[-1])";
                return false;
            }
            std::vector<std::array<Corner, 3>> tris;
            triangulate_polygon(cs, V, tris);
            for (const auto& t : tris) {
                Face fc;
                fc.c[0] = t[0]; fc.c[1] = t[1]; fc.c[2] = t[2];
                fc.mat = cur_mat;
                shapes.back().faces.push_back(fc);
            }
        }
    }
    if (mats.empty()) {
        if (err) *err = "could not parse materials ...";  // Model.cpp:156-157
        return false;
    }
    std::map<std::string, int> warned;
    for (auto& sh : shapes) {
        std::set<int> ids;
        for (auto& fc : sh.faces) ids.insert(fc.mat);
        for (int id : ids) {  // ascending material id, Model.cpp:172-176
            Mesh m;
            m.name = sh.name;
            const Material& mt = mats[id];
            for (auto& fc : sh.faces) {
                if (fc.mat != id) continue;
                const int b = (int)m.vertex.size() / 3;
                for (int k = 0; k < 3; k++) {
                    const Corner& c = fc.c[k];
                    for (int a = 0; a < 3; a++) m.vertex.push_back(V[3 * (c.v - 1) + a]);
                    for (int a = 0; a < 3; a++) m.normal.push_back(N[3 * (c.n - 1) + a]);
                    for (int a = 0; a < 2; a++) m.texcoord.push_back(T[2 * (c.t - 1) + a]);
                    m.index.push_back(b + k);
                }
            }
            if (m.vertex.empty()) continue;
            for (int a = 0; a < 3; a++) { m.diffuse[a] = mt.kd[a]; m.emit[a] = mt.ke[a]; }
            m.alpha = mt.ns;                     // Model.cpp:210: shininess is used as roughness
            m.material_id = (uint32_t)(id + 1);  // Model.cpp:222
            if (!mt.map_kd.empty()) {
                std::string tp = mt.map_kd;
                for (auto& ch : tp)
                    if (ch == '\\') ch = '/';
                if (!load_image_rgba8(base + tp, m.tex, m.tex_w, m.tex_h) && !warned[tp]++)
                    fprintf(stderr, "Could not load texture from %s (PNG, JPEG, TGA, BMP and binary PPM are read)\n", (base + tp).c_str());
            }
            out.push_back(std::move(m));
        }
    }
    return true;
}

// ------------------------------------------------------------------------------------------------
// scene JSON  (src/scene.cpp:20-101)
// ------------------------------------------------------------------------------------------------
static std::string resolve(const std::string& json_dir, const std::string& p) {
    if (p.empty() || p[0] == '/') return p;
    std::ifstream direct(p.c_str());  // the reference opens the string as given (absolute paths in its scenes)
    if (direct) return p;
    return json_dir + p;
}

bool parse_scene(const std::string& json_file, Scene& scene, std::string* err) {
    auto bad = [&](const std::string& m) {
        if (err) *err = m;
        return false;
    };
    std::string text;
    if (!read_file(json_file, text)) return bad("Could not load scene .json file");
    JParser jp{text.data(), text.data() + text.size(), ""};
    JVal root;
    if (!jp.val(root) || root.kind != JVal::OBJ) return bad("Could not load scene .json file: " + jp.err);
    scene.json_path = json_file;
    const JVal* v;
    if (!(v = root.get("spp")) || v->kind != JVal::NUM) return bad("scene: 'spp' missing");
    scene.spp = (int)v->num;
    if (!(v = root.get("width")) || v->kind != JVal::NUM) return bad("scene: 'width' missing");
    scene.width = (int)v->num;
    if (!(v = root.get("height")) || v->kind != JVal::NUM) return bad("scene: 'height' missing");
    scene.height = (int)v->num;
    // As the reference reads it (src/scene.cpp:44-54; pinned on its parseScene, tests/golden/scene_ingest.json): a
    // missing "renderers" is an empty list, a lone number a list of one; only a non-number element is an error.
    if ((v = root.get("renderers"))) {
        if (v->kind == JVal::NUM) {
            scene.renderers.push_back((int)v->num);
        } else if (v->kind == JVal::ARR) {
            for (auto& r : v->arr) {
                if (r.kind != JVal::NUM) return bad("No renderers defined.");
                scene.renderers.push_back((int)r.num);
            }
        } else if (v->kind != JVal::NUL) {
            return bad("No renderers defined.");
        }
    }
    // (the reference accepts a scene without cameras and dereferences cameras[0] later; refused here with its message)
    if (!(v = root.get("cameras")) || v->kind != JVal::ARR || v->arr.empty()) return bad("No cameras defined.");
    for (auto& c : v->arr) {
        SceneCamera cam;
        const JVal* f = c.get("cos_fovy");
        if (c.kind != JVal::OBJ || !vec3_of(c.get("from"), cam.from) || !vec3_of(c.get("to"), cam.at) ||
            !vec3_of(c.get("up"), cam.up) || !f || f->kind != JVal::NUM)
            return bad("No cameras defined.");
        cam.cos_fovy = (float)f->num;
        scene.cameras.push_back(cam);
    }
    const std::string jd = dir_of(json_file);
    if (!(v = root.get("surface_geometry")) || v->kind != JVal::STR) return bad("No .obj file given/found!");
    std::string e;
    if (!load_obj(resolve(jd, v->str), scene.meshes, &e)) return bad(e);
    scene.n_surface_meshes = scene.meshes.size();
    if (!(v = root.get("area_lights")) || v->kind != JVal::STR) return bad("No .obj file for area lights given/found!");
    std::vector<Mesh> lights;
    if (!load_obj(resolve(jd, v->str), lights, &e)) return bad(e);
    for (auto& l : lights) {  // Scene::syncLights, src/scene.cpp:5-13
        l.is_light = true;
        l.material_id = 0;
        scene.meshes.push_back(std::move(l));
    }
    // Model::bounds of the SURFACE model (Model.cpp:244-246), used by the reference only for the viewer
    for (int a = 0; a < 3; a++) { scene.bounds_lo[a] = INFINITY; scene.bounds_hi[a] = -INFINITY; }
    for (size_t i = 0; i < scene.n_surface_meshes; i++)
        for (size_t k = 0; k < scene.meshes[i].vertex.size(); k++) {
            const float x = scene.meshes[i].vertex[k];
            const int a = (int)(k % 3);
            if (x < scene.bounds_lo[a]) scene.bounds_lo[a] = x;
            if (x > scene.bounds_hi[a]) scene.bounds_hi[a] = x;
        }
    return true;
}

// ------------------------------------------------------------------------------------------------
// light list and camera (include/viewer.hpp:236-265, 635-641) -- fp32, same operation order
// ------------------------------------------------------------------------------------------------
static void sub3(const float* a, const float* b, float* o) { o[0] = a[0] - b[0]; o[1] = a[1] - b[1]; o[2] = a[2] - b[2]; }
static void cross3(const float* a, const float* b, float* o) {
    const float x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
    o[0] = x; o[1] = y; o[2] = z;
}
static float dot3(const float* a, const float* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static void normalize3(float* a) {  // owl: v * (1 / sqrt(dot(v, v)))
    const float s = 1.f / sqrtf(dot3(a, a));
    a[0] *= s; a[1] *= s; a[2] *= s;
}

std::vector<OrbTriLight> build_tri_lights(const Scene& scene) {
    std::vector<OrbTriLight> out;
    for (size_t mi = scene.n_surface_meshes; mi < scene.meshes.size(); mi++) {
        const Mesh& m = scene.meshes[mi];
        for (size_t t = 0; t + 2 < m.index.size(); t += 3) {
            OrbTriLight L;
            const float* v1 = &m.vertex[3 * m.index[t]];
            const float* v2 = &m.vertex[3 * m.index[t + 1]];
            const float* v3 = &m.vertex[3 * m.index[t + 2]];
            const float* n1 = &m.normal[3 * m.index[t]];
            const float* n2 = &m.normal[3 * m.index[t + 1]];
            const float* n3 = &m.normal[3 * m.index[t + 2]];
            for (int a = 0; a < 3; a++) {
                L.v1[a] = v1[a]; L.v2[a] = v2[a]; L.v3[a] = v3[a];
                L.cg[a] = ((v1[a] + v2[a]) + v3[a]) / 3.f;
                L.normal[a] = (n1[a] + n2[a]) + n3[a];
                L.emit[a] = m.emit[a];
            }
            normalize3(L.normal);
            float e1[3], e2[3], c[3];
            sub3(v1, v2, e1);
            sub3(v3, v2, e2);
            cross3(e1, e2, c);
            L.area = 0.5f * sqrtf(dot3(c, c));
            out.push_back(L);
        }
    }
    return out;
}

void camera_vectors(const SceneCamera& cam, int width, int height, float pos[3], float d00[3], float du[3], float dv[3]) {
    float t[3];
    for (int a = 0; a < 3; a++) pos[a] = cam.from[a];
    sub3(cam.at, cam.from, d00);
    normalize3(d00);
    const float aspect = (float)width / (float)height;
    cross3(d00, cam.up, t);
    normalize3(t);
    for (int a = 0; a < 3; a++) du[a] = (cam.cos_fovy * aspect) * t[a];
    cross3(du, d00, t);
    normalize3(t);
    for (int a = 0; a < 3; a++) dv[a] = cam.cos_fovy * t[a];
    for (int a = 0; a < 3; a++) d00[a] -= 0.5f * du[a];
    for (int a = 0; a < 3; a++) d00[a] -= 0.5f * dv[a];
}

// ------------------------------------------------------------------------------------------------
// LTC lookup table: numpy .npy, little-endian float32, C order, shape (3, h, w, 4)
// ------------------------------------------------------------------------------------------------
bool load_ltc_lut_npy(const std::string& path, std::vector<float>& m1, std::vector<float>& m2, std::vector<float>& m3,
                      int& w, int& h, std::string* err) {
    auto bad = [&](const std::string& m) {
        if (err) *err = path + ": " + m;
        return false;
    };
    std::string s;
    if (!read_file(path, s)) return bad("cannot read");
    if (s.size() < 12 || memcmp(s.data(), "\x93NUMPY", 6) != 0) return bad("not a .npy file");
    const int major = (unsigned char)s[6];
    size_t hlen, hoff;
    if (major == 1) { hlen = (unsigned char)s[8] | ((unsigned char)s[9] << 8); hoff = 10; }
    else { hlen = (unsigned char)s[8] | ((unsigned char)s[9] << 8) | ((unsigned char)s[10] << 16) | ((size_t)(unsigned char)s[11] << 24); hoff = 12; }
    if (s.size() < hoff + hlen) return bad("truncated header");
    const std::string hd = s.substr(hoff, hlen);
    if (hd.find("<f4") == std::string::npos || hd.find("'fortran_order': False") == std::string::npos)
        return bad("expected little-endian float32 in C order");
    size_t a = hd.find("'shape': (");
    if (a == std::string::npos) return bad("no shape");
    int dims[4] = {0, 0, 0, 0};
    if (sscanf(hd.c_str() + a + 10, "%d, %d, %d, %d", &dims[0], &dims[1], &dims[2], &dims[3]) != 4 || dims[0] != 3 || dims[3] != 4)
        return bad("expected shape (3, h, w, 4)");
    h = dims[1];
    w = dims[2];
    const size_t n = (size_t)4 * w * h;
    if (s.size() < hoff + hlen + 3 * n * 4) return bad("truncated data");
    const float* d = (const float*)(s.data() + hoff + hlen);
    m1.assign(d, d + n);
    m2.assign(d + n, d + 2 * n);
    m3.assign(d + 2 * n, d + 3 * n);
    return true;
}

}  // namespace orbhost
