// ptb_scene_loader.cpp — see ptb_scene_loader.h.  Compiled without floating-point contraction (build.py) so that
// `v * size + position` rounds twice like the reference's build (-O3 -mavx2, no FMA; object.cpp:107-109).
#include "ptb_scene_loader.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <set>
#include <sstream>

namespace ptb {
namespace {

// ---- text primitives, as tinyobjloader 2.0.0 defines them -----------------------------------------------------------
inline bool is_space(char c) { return c == ' ' || c == '\t'; }
inline bool is_digit(char c) { return c >= '0' && c <= '9'; }
inline bool is_eol(char c) { return c == '\r' || c == '\n' || c == '\0'; }

// Decimal text -> double, the way tiny_obj_loader.h:897-1028 does it (NOT strtod: digits are accumulated in a double,
// fraction digits through a table of 0.1 .. 1e-7 then pow(10, -k), the exponent through ldexp(m * 5^e, e)), so that the
// float it is narrowed to is the reference's bit for bit.  Returns false on malformed text.
bool parse_decimal(const char *s, const char *end, double &out) {
    if (s >= end) return false;
    static const double frac[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
    double m = 0.0;
    bool neg = false, dot_first = false;
    const char *p = s;
    if (*p == '+' || *p == '-') {
        neg = *p == '-';
        ++p;
        dot_first = p != end && *p == '.';
    } else if (*p == '.') {
        dot_first = true;
    } else if (!is_digit(*p)) {
        return false;
    }
    if (!dot_first) {
        int digits = 0;
        for (; p != end && is_digit(*p); ++p, ++digits) m = m * 10 + (*p - '0');
        if (digits == 0) return false;
    }
    int e10 = 0;
    if (p != end && *p == '.') {
        ++p;
        for (int k = 1; p != end && is_digit(*p); ++p, ++k) m += (*p - '0') * (k < 8 ? frac[k] : std::pow(10.0, -k));
    } else if (p != end && (*p == 'e' || *p == 'E')) {
        // exponent directly after the integer part
    } else {
        out = neg ? -m : m;
        return true;
    }
    if (p != end && (*p == 'e' || *p == 'E')) {
        ++p;
        bool eneg = false;
        if (p != end && (*p == '+' || *p == '-')) {
            eneg = *p == '-';
            ++p;
        } else if (p == end || !is_digit(*p)) {
            return false;
        }
        int digits = 0;
        for (; p != end && is_digit(*p); ++p, ++digits) {
            if (e10 > 2147483647 / 10) return false;
            e10 = e10 * 10 + (*p - '0');
        }
        if (digits == 0) return false;
        if (eneg) e10 = -e10;
    }
    const double mag = e10 ? std::ldexp(m * std::pow(5.0, e10), e10) : m;
    out = neg ? -mag : mag;
    return true;
}

struct Cursor { // a position in one line; the accessors advance like tinyobj's parseString / parseReal / parseInt
    const char *p;
    void skip() { p += strspn(p, " \t"); }
    std::string word() { // tiny_obj_loader.h:854-861
        skip();
        const size_t n = strcspn(p, " \t\r");
        std::string s(p, n);
        p += n;
        return s;
    }
    float real_or(double dflt) { // :1030-1038
        skip();
        const char *e = p + strcspn(p, " \t\r");
        double d = dflt;
        parse_decimal(p, e, d);
        p = e;
        return (float)d;
    }
    int integer() { // :863-868
        skip();
        const int i = atoi(p);
        p += strcspn(p, " \t\r");
        return i;
    }
    void on_off() { // :1121-1134, value unused here
        skip();
        p += strcspn(p, " \t\r");
    }
};

bool starts(const char *t, const char *kw) {
    const size_t n = strlen(kw);
    return strncmp(t, kw, n) == 0 && is_space(t[n]);
}

// std::getline that ends a line at \n, \r\n or \r (safeGetline, tiny_obj_loader.h:777-815)
bool next_line(std::istream &is, std::string &line) {
    line.clear();
    if (is.peek() == EOF) return false;
    for (;;) {
        const int c = is.get();
        if (c == EOF) {
            is.clear(std::ios::eofbit);
            return true;
        }
        if (c == '\n') return true;
        if (c == '\r') {
            if (is.peek() == '\n') is.get();
            return true;
        }
        line.push_back((char)c);
    }
}

// ---- MTL ------------------------------------------------------------------------------------------------------------
struct MtlEntry { // the fields of tinyobj::material_t loadObject reads (object.cpp:53-86)
    std::string name, diffuse_tex, bump_tex;
    float diffuse[3] = {0, 0, 0}, emission[3] = {0, 0, 0};
};

// texture statement: options first, the rest of the line is the file name (tiny_obj_loader.h:1274-1352)
std::string texture_name(const char *rest) {
    Cursor c{rest};
    std::string name;
    while (!is_eol(*c.p)) {
        c.skip();
        const char *t = c.p;
        if (starts(t, "-blendu") || starts(t, "-blendv")) { c.p += 8; c.on_off(); }
        else if (starts(t, "-clamp")) { c.p += 7; c.on_off(); }
        else if (starts(t, "-boost")) { c.p += 7; c.real_or(1.0); }
        else if (starts(t, "-bm")) { c.p += 4; c.real_or(1.0); }
        else if (starts(t, "-o") || starts(t, "-s") || starts(t, "-t")) { c.p += 3; c.real_or(0); c.real_or(0); c.real_or(0); }
        else if (starts(t, "-type")) { c.p += 5; c.word(); }
        else if (starts(t, "-texres")) { c.p += 7; c.integer(); }
        else if (starts(t, "-imfchan")) { c.p += 9; c.skip(); if (!is_eol(*c.p)) c.p += strcspn(c.p, " \t\r"); }
        else if (starts(t, "-mm")) { c.p += 4; c.real_or(0); c.real_or(1.0); }
        else if (starts(t, "-colorspace")) { c.p += 12; c.word(); }
        else { // the file name: everything up to the end of the line
            name = std::string(c.p);
            while (!name.empty() && (name.back() == '\r' || name.back() == '\n')) name.pop_back();
            break;
        }
    }
    return name;
}

// LoadMtl, tiny_obj_loader.h:2069-2467: statements other than newmtl / Kd / Ke / map_Kd / map_Bump / bump do not reach
// the reference's Material and are skipped.  The closing entry is pushed even if it has no name (:2459-2462).
void read_mtl(std::istream &is, std::vector<MtlEntry> &materials, std::map<std::string, int> &by_name) {
    MtlEntry cur;
    bool has_kd = false; // never cleared between materials (:2084)
    std::string line;
    while (next_line(is, line)) {
        const size_t last = line.find_last_not_of(" \t");
        line = last == std::string::npos ? std::string() : line.substr(0, last + 1);
        if (line.empty()) continue;
        const char *t = line.c_str();
        t += strspn(t, " \t");
        if (*t == '\0' || *t == '#') continue;
        if (starts(t, "newmtl")) {
            if (!cur.name.empty()) {
                by_name.insert({cur.name, (int)materials.size()});
                materials.push_back(cur);
            }
            cur = MtlEntry();
            Cursor c{t + 7};
            cur.name = c.word();
        } else if (t[0] == 'K' && (t[1] == 'd' || t[1] == 'e') && is_space(t[2])) {
            Cursor c{t + 2};
            float *dst = t[1] == 'd' ? cur.diffuse : cur.emission;
            for (int k = 0; k < 3; ++k) dst[k] = c.real_or(0.0);
            if (t[1] == 'd') has_kd = true;
        } else if (starts(t, "map_Kd")) {
            cur.diffuse_tex = texture_name(t + 7);
            if (!has_kd) cur.diffuse[0] = cur.diffuse[1] = cur.diffuse[2] = (float)0.6; // :2328-2332
        } else if (starts(t, "map_bump") || starts(t, "map_Bump")) {
            cur.bump_tex = texture_name(t + 9);
        } else if (starts(t, "bump")) {
            cur.bump_tex = texture_name(t + 5);
        }
    }
    by_name.insert({cur.name, (int)materials.size()});
    materials.push_back(cur);
}

} // namespace

// ---- OBJ ------------------------------------------------------------------------------------------------------------
namespace {

struct Corner { int v = -1, vt = -1, vn = -1; };
struct Shape { // tinyobj::shape_t, mesh part: three corners per triangle + one material id per triangle
    std::string name;
    std::vector<Corner> corners;
    std::vector<int> material_ids;
};
struct ObjFile {
    std::vector<float> v, vn, vt;
    std::vector<Shape> shapes;
    std::vector<MtlEntry> materials;
};

// OBJ indices are 1-based, negative = relative to the end (fixIndex, tiny_obj_loader.h:819-852)
bool fix_index(int idx, int n, int &out, bool allow_zero) {
    if (idx > 0) { out = idx - 1; return true; }
    if (idx == 0) { out = -1; return allow_zero; }
    out = n + idx;
    return out >= 0;
}

// one "v", "v/vt", "v//vn" or "v/vt/vn" group (parseTriple, :1188-1239)
bool parse_corner(const char *&t, int nv, int nvn, int nvt, Corner &c) {
    c = Corner();
    if (!fix_index(atoi(t), nv, c.v, false)) return false;
    t += strcspn(t, "/ \t\r");
    if (*t != '/') return true;
    ++t;
    if (*t == '/') {
        ++t;
        if (!fix_index(atoi(t), nvn, c.vn, true)) return false;
        t += strcspn(t, "/ \t\r");
        return true;
    }
    if (!fix_index(atoi(t), nvt, c.vt, true)) return false;
    t += strcspn(t, "/ \t\r");
    if (*t != '/') return true;
    ++t;
    if (!fix_index(atoi(t), nvn, c.vn, true)) return false;
    t += strcspn(t, "/ \t\r");
    return true;
}

// exportGroupsToShape (:1481-1960), faces only: triangles as they are, quads cut along the shorter diagonal (ties: 1-3),
// larger polygons are refused — tinyobj ear-clips them, which is not restated here.
bool flush_faces(Shape &shape, const std::vector<std::vector<Corner>> &faces, int material, const std::string &name,
                 const std::vector<float> &v, std::string &log, std::string &err) {
    if (faces.empty()) return false;
    shape.name = name;
    for (const auto &f : faces) {
        if (f.size() < 3) {
            log += "Degenerated face found\n";
            continue;
        }
        if (f.size() == 3) {
            shape.corners.insert(shape.corners.end(), f.begin(), f.end());
            shape.material_ids.push_back(material);
        } else if (f.size() == 4) {
            bool bad = false;
            for (const Corner &c : f) bad |= 3 * (size_t)c.v + 2 >= v.size();
            if (bad) {
                log += "Face with invalid vertex index found.\n";
                continue;
            }
            const float *p0 = &v[3 * (size_t)f[0].v], *p1 = &v[3 * (size_t)f[1].v], *p2 = &v[3 * (size_t)f[2].v], *p3 = &v[3 * (size_t)f[3].v];
            const float ax = p2[0] - p0[0], ay = p2[1] - p0[1], az = p2[2] - p0[2];
            const float bx = p3[0] - p1[0], by = p3[1] - p1[1], bz = p3[2] - p1[2];
            const float d02 = ax * ax + ay * ay + az * az, d13 = bx * bx + by * by + bz * bz;
            const int cut[2][6] = {{0, 1, 2, 0, 2, 3}, {0, 1, 3, 1, 2, 3}};
            for (int k : cut[d02 < d13 ? 0 : 1]) shape.corners.push_back(f[k]);
            shape.material_ids.push_back(material);
            shape.material_ids.push_back(material);
        } else {
            err = "polygon with " + std::to_string(f.size()) + " vertices: triangulate the mesh (only triangles and quads are loaded)";
            return false;
        }
    }
    return true;
}

// LoadObj (tiny_obj_loader.h:2584-3146) for the statements that reach the triangle buffer: v vn vt f usemtl mtllib g o
bool read_obj(const std::string &file, const std::string &mtl_dir, ObjFile &out, std::string &log, std::string &err) {
    std::ifstream is(file, std::ios::binary);
    if (!is) {
        err = "Cannot open file [" + file + "]";
        return false;
    }
    std::map<std::string, int> by_name;
    std::set<std::string> mtl_seen;
    std::vector<std::vector<Corner>> faces;
    Shape shape;
    std::string name, line;
    int material = -1;
    size_t line_no = 0;
    auto flush = [&]() { return flush_faces(shape, faces, material, name, out.v, log, err); };
    while (next_line(is, line)) {
        ++line_no;
        if (line.empty()) continue;
        const char *t = line.c_str();
        t += strspn(t, " \t");
        if (*t == '\0' || *t == '#') continue;
        if (t[0] == 'v' && is_space(t[1])) {
            Cursor c{t + 2};
            for (int k = 0; k < 3; ++k) out.v.push_back(c.real_or(0.0));
        } else if (t[0] == 'v' && t[1] == 'n' && is_space(t[2])) {
            Cursor c{t + 3};
            for (int k = 0; k < 3; ++k) out.vn.push_back(c.real_or(0.0));
        } else if (t[0] == 'v' && t[1] == 't' && is_space(t[2])) {
            Cursor c{t + 3};
            for (int k = 0; k < 2; ++k) out.vt.push_back(c.real_or(0.0));
        } else if (t[0] == 'f' && is_space(t[1])) {
            t += 2;
            t += strspn(t, " \t");
            std::vector<Corner> f;
            while (!is_eol(*t)) {
                Corner c;
                if (!parse_corner(t, (int)(out.v.size() / 3), (int)(out.vn.size() / 3), (int)(out.vt.size() / 2), c)) {
                    err = "Failed to parse `f' line (e.g. a zero value for vertex index or invalid relative vertex index). Line " +
                          std::to_string(line_no);
                    return false;
                }
                f.push_back(c);
                t += strspn(t, " \t\r");
            }
            faces.push_back(std::move(f));
        } else if (strncmp(t, "usemtl", 6) == 0) {
            Cursor c{t + 6};
            const std::string m = c.word();
            const auto it = by_name.find(m);
            const int id = it != by_name.end() ? it->second : -1;
            if (it == by_name.end()) log += "material [ '" + m + "' ] not found in .mtl\n";
            if (id != material) { // faces so far keep the old material; the shape goes on (:2839-2847)
                flush();
                if (!err.empty()) return false;
                faces.clear();
                material = id;
            }
        } else if (starts(t, "mtllib")) {
            std::istringstream names(t + 7);
            std::string fn;
            bool found = false;
            while (!found && names >> fn) {
                if (mtl_seen.count(fn)) {
                    found = true;
                    break;
                }
                std::ifstream ms(mtl_dir + fn, std::ios::binary);
                if (!ms) {
                    log += "Material file [ " + fn + " ] not found in a path : " + mtl_dir + "\n";
                    continue;
                }
                read_mtl(ms, out.materials, by_name);
                mtl_seen.insert(fn);
                found = true;
            }
            if (!found) log += "Failed to load material file(s). Use default material.\n";
        } else if ((t[0] == 'g' || t[0] == 'o') && is_space(t[1])) {
            flush();
            if (!err.empty()) return false;
            if (!shape.corners.empty()) out.shapes.push_back(shape);
            shape = Shape();
            faces.clear();
            if (t[0] == 'o') {
                name = std::string(t + 2); // the rest of the line, spaces included (:2989-2992)
            } else {
                // the statement letter itself is the first "name" the reference collects, the group name starts at the second
                Cursor c{t};
                std::vector<std::string> words;
                while (!is_eol(*c.p)) {
                    words.push_back(c.word());
                    c.p += strspn(c.p, " \t\r");
                }
                name.clear();
                for (size_t i = 1; i < words.size(); ++i) name += (i > 1 ? " " : "") + words[i];
            }
        }
        // everything else (s, l, p, t, vw, ...) does not reach the triangle buffer
    }
    const bool some = flush();
    if (!err.empty()) return false;
    if (some || !shape.corners.empty()) out.shapes.push_back(shape);
    return true;
}

void set_name(char (&dst)[64], const std::string &s) {
    memset(dst, 0, sizeof dst);
    strncpy(dst, s.c_str(), sizeof dst - 1);
}

ptbs_material default_material(const std::string &name) { // pbrProbs / ShaderMix defaults, shader.h:14-35
    ptbs_material m{};
    m.ior1 = 1.0f;
    m.ior2 = 1.5f;
    m.emission = 0.0f;
    m.roughness = 0.5f;
    m.w_lambert = 1.0f;
    set_name(m.name, name.empty() ? "Material" : name); // addMaterial, shader.cpp:31-37
    return m;
}

// Eigen's 3-vector norm: the reduction splits 1 + 2 (vector.cpp:17-19; same order as ptb_sat.h's dot)
float length3(float x, float y, float z) { return std::sqrt(x * x + (y * y + z * z)); }

float surface_area(const float *p) { // calculateTriangleSurfaceArea, triangle.cpp:95-110 (Heron)
    const float l1 = length3(p[6] - p[0], p[7] - p[1], p[8] - p[2]);
    const float l2 = length3(p[3] - p[0], p[4] - p[1], p[5] - p[2]);
    const float l3 = length3(p[6] - p[3], p[7] - p[4], p[8] - p[5]);
    const float s = (l1 + l2 + l3) / 2;
    return sqrtf(s * (s - l1) * (s - l2) * (s - l3));
}

} // namespace

// ---- loadObject (object.cpp:35-174) -----------------------------------------------------------------------------------
bool load_obj(const std::string &file, const float position[3], const float size[3], int default_mat, LoadedScene &sc, std::string &err,
              int flags) {
    const size_t slash = file.find_last_of('/');
    const std::string dir = file.substr(0, slash); // npos: the whole name, as the reference's substr does (:45)
    ObjFile obj;
    if (!read_obj(file, dir + "/", obj, sc.log, err)) return false;

    const int first_mat = (int)sc.materials.size();
    for (const MtlEntry &me : obj.materials) { // :53-86
        ptbs_material m = default_material(me.name);
        float c[3] = {me.diffuse[0], me.diffuse[1], me.diffuse[2]};
        if (me.emission[0] != 0 || me.emission[1] != 0 || me.emission[2] != 0) {
            for (int k = 0; k < 3; ++k) c[k] = me.emission[k];
            const float ma = std::max(c[0], std::max(c[1], c[2]));
            float intensity = 1;
            if (ma >= 1) {
                for (int k = 0; k < 3; ++k) c[k] /= ma;
                intensity = ma;
            }
            m.emission = intensity;
        }
        for (int k = 0; k < 3; ++k) m.albedo[k] = c[k];
        std::array<std::string, 2> tex;
        const std::string *names[2] = {&me.diffuse_tex, &me.bump_tex};
        for (int k = 0; k < 2; ++k) {
            if (names[k]->empty()) continue;
            tex[k] = names[k]->front() == '/' ? *names[k] : dir + "/" + *names[k];
            if (FILE *probe = fopen(tex[k].c_str(), "rb")) {
                fclose(probe);
            } else { // loadTexture prints this and leaves the material untextured (texture.cpp:36-39)
                sc.log += "could not load: " + tex[k] + "\n";
                tex[k].clear();
            }
        }
        if ((!tex[0].empty() || !tex[1].empty()) && !(flags & kSceneIgnoreTextures)) {
            err = "material " + me.name + " names the texture file " + (tex[0].empty() ? tex[1] : tex[0]) +
                  ": image decoding (stbi_load, texture.cpp:34) is not part of this loader; decode it on the caller's side and hand the "
                  "texels to ptb_upload_textures, or load with the ignore-textures flag";
            return false;
        }
        if (flags & kSceneIgnoreTextures) tex = {};
        sc.materials.push_back(m);
        sc.textures.push_back(tex);
        sc.log += "loaded material " + me.name + "\n";
    }

    const size_t nv = obj.v.size() / 3, nn = obj.vn.size() / 3, nt = obj.vt.size() / 2;
    for (const Shape &sh : obj.shapes) { // :89-172
        sc.log += "loaded: " + sh.name + "\n";
        const int start = (int)sc.triangles();
        int shape_mat = sh.material_ids[0];
        shape_mat = shape_mat >= 0 ? shape_mat + first_mat : default_mat;
        std::vector<uint32_t> used;
        std::vector<bool> seen(obj.materials.size(), false);
        float area = 0.0f;
        for (size_t i = 0; i < sh.corners.size(); i += 3) {
            float p[9], n[9], uv[6];
            bool have_normals = true;
            for (int j = 0; j < 3; ++j) {
                const Corner &c = sh.corners[i + j];
                if (c.v < 0 || (size_t)c.v >= nv) {
                    err = "shape " + sh.name + ": vertex index out of bounds";
                    return false;
                }
                for (int k = 0; k < 3; ++k) p[3 * j + k] = obj.v[3 * (size_t)c.v + k] * size[k] + position[k];
                // a corner without vn / vt makes the reference read before the start of its arrays; here: the
                // geometric normal (createTriangle's 4-argument flavour, triangle.cpp:54-65) and uv (0, 0)
                if (c.vn >= 0 && (size_t)c.vn < nn)
                    for (int k = 0; k < 3; ++k) n[3 * j + k] = obj.vn[3 * (size_t)c.vn + k];
                else
                    have_normals = false;
                if (c.vt >= 0 && (size_t)c.vt < nt) {
                    uv[2 * j] = obj.vt[2 * (size_t)c.vt];
                    uv[2 * j + 1] = 1.0f - obj.vt[2 * (size_t)c.vt + 1];
                } else {
                    uv[2 * j] = uv[2 * j + 1] = 0.0f;
                }
            }
            if (!have_normals) {
                const float e1[3] = {p[3] - p[0], p[4] - p[1], p[5] - p[2]}, e2[3] = {p[6] - p[0], p[7] - p[1], p[8] - p[2]};
                float g[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
                const float l = length3(g[0], g[1], g[2]);
                for (int j = 0; j < 3; ++j)
                    for (int k = 0; k < 3; ++k) n[3 * j + k] = l > 0 ? g[k] / l : 0.0f;
            }
            int m = sh.material_ids[i / 3];
            if (m >= 0) {
                if (!seen[m]) {
                    seen[m] = true;
                    used.push_back((uint32_t)(m + first_mat));
                }
                m += first_mat;
            } else {
                m = shape_mat;
            }
            sc.pos9.insert(sc.pos9.end(), p, p + 9);
            sc.nrm9.insert(sc.nrm9.end(), n, n + 9);
            sc.uv6.insert(sc.uv6.end(), uv, uv + 6);
            sc.mat.push_back(m);
            area += surface_area(p); // summed in triangle order (:151-154)
        }
        ptbs_object o{};
        o.start_idx = start;
        o.end_idx = (int)sc.triangles();
        o.surface_area = area;
        o.n_materials = (uint32_t)std::min<size_t>(used.size(), PTBS_MAX_OBJECT_MATERIALS);
        for (uint32_t k = 0; k < o.n_materials; ++k) o.materials[k] = used[k];
        set_name(o.name, sh.name);
        sc.objects.push_back(o);
    }
    return true;
}

// ---- loadScene (SceneFile.cpp:21-110) + initScene's model and material loops (scene.cpp:97-110) -------------------------
bool load_scene_file(const std::string &scene_file, const std::string &base_dir, LoadedScene &sc, std::string &err, int flags) {
    FILE *f = fopen(scene_file.c_str(), "r");
    if (!f) {
        err = "cannot open " + scene_file;
        return false;
    }
    struct Model { std::string file; float pos[3] = {0, 0, 0}, scale[3] = {1, 1, 1}; };
    struct Override { std::string name; ptbs_material m; };
    std::vector<Model> models;
    std::vector<Override> overrides;
    char line[1024], word[1024];
    // The reference's block readers stop at the next line containing '[' and leave it in `line`; the checks below run in
    // the file's usual order Models -> Materials -> Camera on that same buffer, so a header is only seen if it follows in
    // that order (a [Models] block after [Camera] is skipped by the reference as well).
    while (fgets(line, sizeof line, f)) {
        if (sscanf(line, "%1023s", word) != 1) continue;
        if (strstr(line, "[Models")) {
            Model m;
            bool open = false;
            while (fgets(line, sizeof line, f) && !strchr(line, '[')) {
                if (sscanf(line, "file: %1023s", word) == 1) {
                    if (open) models.push_back(m);
                    open = true;
                    m.file = word; // pos / scale carry over from the previous model like in the reference
                } else if (sscanf(line, "pos: %f,%f,%f", &m.pos[0], &m.pos[1], &m.pos[2]) == 3) {
                } else if (sscanf(line, "scale: %f,%f,%f", &m.scale[0], &m.scale[1], &m.scale[2]) == 3) {
                }
            }
            models.push_back(m);
        }
        if (strstr(line, "[Materials")) {
            Override o;
            o.m = default_material("");
            bool open = false;
            while (fgets(line, sizeof line, f) && !strchr(line, '[')) {
                if (sscanf(line, "name: %1023s", word) == 1) {
                    if (open) overrides.push_back(o);
                    open = true;
                    o.name = word;
                } else if (sscanf(line, "color: %f,%f,%f", &o.m.albedo[0], &o.m.albedo[1], &o.m.albedo[2]) == 3) {
                } else if (sscanf(line, "weights: %f,%f,%f", &o.m.w_lambert, &o.m.w_reflection, &o.m.w_refraction) == 3) {
                } else if (sscanf(line, "emission: %f", &o.m.emission) == 1) {
                } else if (sscanf(line, "roughness: %f", &o.m.roughness) == 1) {
                }
            }
            overrides.push_back(o);
        }
        if (strstr(line, "[Camera")) {
            while (fgets(line, sizeof line, f) && !strchr(line, '[')) {
                if (sscanf(line, "pos: %f,%f,%f", &sc.cam_pos[0], &sc.cam_pos[1], &sc.cam_pos[2]) == 3) {
                } else if (sscanf(line, "forward: %f,%f,%f", &sc.cam_forward[0], &sc.cam_forward[1], &sc.cam_forward[2]) == 3) {
                } else if (sscanf(line, "dof: %f", &sc.cam_dof) == 1) {
                } else if (sscanf(line, "focus: %f", &sc.cam_focus) == 1) {
                }
            }
        }
    }
    fclose(f);
    sc.log += "Loaded " + std::to_string(models.size()) + " models, " + std::to_string(overrides.size()) + " materials, and 1 camera.\n";

    for (const Model &m : models) { // every model brings the fallback material "None" with it (scene.cpp:25,100)
        const int fallback = (int)sc.materials.size();
        sc.materials.push_back(default_material("None"));
        sc.textures.push_back({});
        std::string path = m.file;
        if (!base_dir.empty() && !path.empty() && path.front() != '/') path = base_dir + "/" + path;
        if (!load_obj(path, m.pos, m.scale, fallback, sc, err, flags)) return false;
    }
    for (const Override &o : overrides) // weights and pbr replaced as a whole, textures included (scene.cpp:103-110)
        for (size_t i = 0; i < sc.materials.size(); ++i) {
            if (o.name != sc.materials[i].name) continue;
            ptbs_material r = o.m;
            memcpy(r.name, sc.materials[i].name, sizeof r.name);
            sc.materials[i] = r;
            sc.textures[i] = {};
        }
    return true;
}

bool write_ptbs(const std::string &path, const LoadedScene &sc, std::string &err) {
    ptbs_header h{};
    h.magic = PTBS_MAGIC;
    h.version = PTBS_VERSION;
    h.n_tris = (uint32_t)sc.triangles();
    h.n_objects = (uint32_t)sc.objects.size();
    h.n_materials = (uint32_t)sc.materials.size();
    for (int k = 0; k < 3; ++k) {
        h.cam_pos[k] = sc.cam_pos[k];
        h.cam_forward[k] = sc.cam_forward[k];
    }
    h.cam_focus = sc.cam_focus;
    h.cam_dof = sc.cam_dof;
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) {
        err = "cannot write " + path;
        return false;
    }
    bool ok = fwrite(&h, sizeof h, 1, f) == 1;
    auto put = [&](const void *p, size_t bytes) { ok = ok && (bytes == 0 || fwrite(p, 1, bytes, f) == bytes); };
    put(sc.pos9.data(), sc.pos9.size() * 4);
    put(sc.nrm9.data(), sc.nrm9.size() * 4);
    put(sc.uv6.data(), sc.uv6.size() * 4);
    put(sc.mat.data(), sc.mat.size() * 4);
    put(sc.objects.data(), sc.objects.size() * sizeof(ptbs_object));
    put(sc.materials.data(), sc.materials.size() * sizeof(ptbs_material));
    ok = fclose(f) == 0 && ok;
    if (!ok) err = "short write to " + path;
    return ok;
}

} // namespace ptb
