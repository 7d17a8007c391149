// ptb_scene_loader.cpp — see ptb_scene_loader.h.  Compiled without floating-point contraction (build.py) so that
// `v * size + position` rounds twice like the reference's build (-O3 -mavx2, no FMA; object.cpp:107-109).
#include "ptb_scene_loader.h"

#include "ptb_png.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <set>
#include <sstream>
#include <thread>

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

// ---- the file in memory, cut into chunks of whole lines that are scanned and parsed by one thread each ------------------
// v / vn / vt / f statements only depend on how many v / vn / vt came before them (relative indices), so they are counted
// per chunk, prefix-summed, and parsed in parallel into their final places; the statements that carry state from line to
// line (usemtl, mtllib, g, o) are collected with their position among the faces and replayed in file order afterwards.
enum Statement { S_OTHER, S_V, S_VN, S_VT, S_F, S_USEMTL, S_MTLLIB, S_G, S_O };

Statement classify(const char *t) { // the order of tests of LoadObj's if-chain (tiny_obj_loader.h:2640-3000)
    if (t[0] == 'v' && is_space(t[1])) return S_V;
    if (t[0] == 'v' && t[1] == 'n' && is_space(t[2])) return S_VN;
    if (t[0] == 'v' && t[1] == 't' && is_space(t[2])) return S_VT;
    if (t[0] == 'f' && is_space(t[1])) return S_F;
    if (strncmp(t, "usemtl", 6) == 0) return S_USEMTL;
    if (starts(t, "mtllib")) return S_MTLLIB;
    if (t[0] == 'g' && is_space(t[1])) return S_G;
    if (t[0] == 'o' && is_space(t[1])) return S_O;
    return S_OTHER;
}

struct Control {
    Statement kind;
    std::string text;  // the statement's line, leading blanks removed
    uint32_t face;     // faces of its chunk that precede it
    size_t nv_local;   // "v" statements of its chunk that precede it
};
struct Chunk {
    const char *b = nullptr, *e = nullptr;
    size_t nv = 0, nvn = 0, nvt = 0, lines = 0, nfaces = 0, ncorners = 0; // counted by scan()
    size_t v0 = 0, vn0 = 0, vt0 = 0, line0 = 0;                           // prefix sums
    std::vector<Control> ctl;
    std::vector<Corner> corners; // faces of the chunk, CSR
    std::vector<uint32_t> start;
    std::string err;
};

// calls f(text, line_number_in_chunk) for every non-empty, non-comment line; lines end at \n, \r\n or \r (safeGetline)
template <class F> void for_lines(const char *b, const char *e, std::string &line, F f) {
    size_t no = 0;
    while (b < e) {
        const char *p = b;
        while (p < e && *p != '\n' && *p != '\r') ++p;
        line.assign(b, p);
        b = p;
        if (b < e) b += (*b == '\r' && b + 1 < e && b[1] == '\n') ? 2 : 1;
        ++no;
        const char *t = line.c_str();
        t += strspn(t, " \t");
        if (*t == '\0' || *t == '#') continue;
        f(t, no);
    }
}

void scan_chunk(Chunk &c) {
    std::string line;
    for_lines(c.b, c.e, line, [&](const char *t, size_t no) {
        (void)no;
        switch (classify(t)) {
        case S_V: c.nv++; break;
        case S_VN: c.nvn++; break;
        case S_VT: c.nvt++; break;
        case S_F: {
            c.nfaces++;
            const char *p = t + 2;
            p += strspn(p, " \t");
            while (!is_eol(*p)) { // one corner per blank-separated word
                c.ncorners++;
                p += strcspn(p, " \t\r");
                p += strspn(p, " \t\r");
            }
            break;
        }
        case S_USEMTL: case S_MTLLIB: case S_G: case S_O:
            c.ctl.push_back(Control{classify(t), std::string(t), (uint32_t)c.nfaces, c.nv});
            break;
        default: break;
        }
    });
    // lines: for_lines numbers every line, c.lines holds the last statement's; count the rest for the prefix sum
    size_t total = 0;
    for (const char *p = c.b; p < c.e; ++p)
        if (*p == '\n' || (*p == '\r' && !(p + 1 < c.e && p[1] == '\n'))) ++total;
    if (c.e > c.b && c.e[-1] != '\n' && c.e[-1] != '\r') ++total;
    c.lines = total;
}

void parse_chunk(Chunk &c, ObjFile &out) {
    std::string line;
    size_t iv = c.v0, ivn = c.vn0, ivt = c.vt0;
    c.corners.reserve(c.ncorners);
    c.start.reserve(c.nfaces + 1);
    c.start.push_back(0);
    for_lines(c.b, c.e, line, [&](const char *t, size_t no) {
        if (!c.err.empty()) return;
        switch (classify(t)) {
        case S_V: {
            Cursor k{t + 2};
            for (int j = 0; j < 3; ++j) out.v[3 * iv + j] = k.real_or(0.0);
            ++iv;
            break;
        }
        case S_VN: {
            Cursor k{t + 3};
            for (int j = 0; j < 3; ++j) out.vn[3 * ivn + j] = k.real_or(0.0);
            ++ivn;
            break;
        }
        case S_VT: {
            Cursor k{t + 3};
            for (int j = 0; j < 2; ++j) out.vt[2 * ivt + j] = k.real_or(0.0);
            ++ivt;
            break;
        }
        case S_F: {
            const char *p = t + 2;
            p += strspn(p, " \t");
            while (!is_eol(*p)) {
                Corner k;
                if (!parse_corner(p, (int)iv, (int)ivn, (int)ivt, k)) {
                    c.err = "Failed to parse `f' line (e.g. a zero value for vertex index or invalid relative vertex index). Line " +
                            std::to_string(c.line0 + no);
                    return;
                }
                c.corners.push_back(k);
                p += strspn(p, " \t\r");
            }
            c.start.push_back((uint32_t)c.corners.size());
            break;
        }
        default: break;
        }
    });
}

template <class F> void run_parallel(size_t n, F f) {
    if (n <= 1) {
        for (size_t i = 0; i < n; ++i) f(i);
        return;
    }
    std::vector<std::thread> th;
    for (size_t i = 0; i < n; ++i) th.emplace_back([&f, i]() { f(i); });
    for (auto &t : th) t.join();
}

struct FaceRange { const Chunk *c; uint32_t b, e; };

// even-odd point-in-polygon test on a triangle (pnpoly, tiny_obj_loader.h:1438-1450), fp32 like the reference's real_t
bool inside_triangle(const float *x, const float *y, float tx, float ty) {
    bool in = false;
    for (int i = 0, j = 2; i < 3; j = i++)
        if (((y[i] > ty) != (y[j] > ty)) && (tx < (x[j] - x[i]) * (ty - y[i]) / (y[j] - y[i]) + x[i])) in = !in;
    return in;
}

// Polygons of five or more corners: tinyobj's built-in ear clipping (:1740-1958; the reference does not enable the earcut
// variant).  The polygon is projected on the two axes that the first non-degenerate corner's normal is NOT largest along;
// a candidate corner (g, g+1, g+2) is an ear when cross * area >= 0 — with `area` the reference's (x0*y1 - y0*x1)/2 —
// and no other remaining vertex lies inside it; its middle vertex is removed; the search gives up after a full round
// without progress, and a last triangle is emitted when exactly three vertices remain.  Vertices not read yet (index
// >= nv) count as (0, 0) in the tests, as there.
void clip_ears(const Corner *f, size_t n, const std::vector<float> &v, size_t nv, Shape &shape, int material) {
    const size_t vsize = 3 * nv;
    size_t ax[2] = {1, 2};
    for (size_t k = 0; k < n; ++k) {
        const size_t i0 = (size_t)f[k % n].v, i1 = (size_t)f[(k + 1) % n].v, i2 = (size_t)f[(k + 2) % n].v;
        if (3 * i0 + 2 >= vsize || 3 * i1 + 2 >= vsize || 3 * i2 + 2 >= vsize) continue;
        const float e0x = v[3 * i1] - v[3 * i0], e0y = v[3 * i1 + 1] - v[3 * i0 + 1], e0z = v[3 * i1 + 2] - v[3 * i0 + 2];
        const float e1x = v[3 * i2] - v[3 * i1], e1y = v[3 * i2 + 1] - v[3 * i1 + 1], e1z = v[3 * i2 + 2] - v[3 * i1 + 2];
        const float cx = std::fabs(e0y * e1z - e0z * e1y), cy = std::fabs(e0z * e1x - e0x * e1z), cz = std::fabs(e0x * e1y - e0y * e1x);
        const float eps = 1.1920929e-07f; // numeric_limits<float>::epsilon()
        if (cx > eps || cy > eps || cz > eps) {
            if (!(cx > cy && cx > cz)) {
                ax[0] = 0;
                if (cz > cx && cz > cy) ax[1] = 1;
            }
            break;
        }
    }
    std::vector<Corner> rest(f, f + n);
    size_t guess = 0, budget = n, before = n;
    auto emit = [&](const Corner &a, const Corner &b, const Corner &c) {
        shape.corners.push_back(a);
        shape.corners.push_back(b);
        shape.corners.push_back(c);
        shape.material_ids.push_back(material);
    };
    while (rest.size() > 3 && budget > 0) {
        const size_t m = rest.size();
        if (guess >= m) guess -= m;
        if (before != m) {
            before = m;
            budget = m;
        } else {
            --budget;
        }
        Corner ind[3];
        float x[3], y[3];
        for (size_t k = 0; k < 3; ++k) {
            ind[k] = rest[(guess + k) % m];
            const size_t vi = (size_t)ind[k].v;
            const bool out = vi * 3 + ax[0] >= vsize || vi * 3 + ax[1] >= vsize;
            x[k] = out ? 0.0f : v[vi * 3 + ax[0]];
            y[k] = out ? 0.0f : v[vi * 3 + ax[1]];
        }
        const float e0x = x[1] - x[0], e0y = y[1] - y[0], e1x = x[2] - x[1], e1y = y[2] - y[1];
        const float cross = e0x * e1y - e0y * e1x;
        const float area = (x[0] * y[1] - y[0] * x[1]) * 0.5f;
        if (cross * area < 0.0f) { // an internal angle
            ++guess;
            continue;
        }
        bool overlap = false;
        for (size_t other = 3; other < m && !overlap; ++other) {
            const size_t idx = (guess + other) % m;
            const size_t ovi = (size_t)rest[idx].v;
            if (ovi * 3 + ax[0] >= vsize || ovi * 3 + ax[1] >= vsize) continue;
            overlap = inside_triangle(x, y, v[ovi * 3 + ax[0]], v[ovi * 3 + ax[1]]);
        }
        if (overlap) {
            ++guess;
            continue;
        }
        emit(ind[0], ind[1], ind[2]);
        rest.erase(rest.begin() + (long)((guess + 1) % m));
    }
    if (rest.size() == 3) emit(rest[0], rest[1], rest[2]);
}

// exportGroupsToShape (:1481-1960), faces only: triangles as they are, quads cut along the shorter diagonal (ties: 1-3),
// larger polygons ear-clipped (clip_ears).  `nv`: vertices read so far (a quad that names a later one is dropped like the
// reference drops it).
bool flush_faces(Shape &shape, const std::vector<FaceRange> &pending, int material, const std::string &name, const std::vector<float> &v,
                 size_t nv, std::string &log, std::string &err) {
    (void)err; // nothing is refused at this point any more
    size_t total = 0;
    for (const FaceRange &r : pending) total += r.e - r.b;
    if (total == 0) return false;
    shape.name = name;
    for (const FaceRange &r : pending)
        for (uint32_t fi = r.b; fi < r.e; ++fi) {
            const Corner *f = &r.c->corners[r.c->start[fi]];
            const size_t n = r.c->start[fi + 1] - r.c->start[fi];
            if (n < 3) {
                log += "Degenerated face found\n";
            } else if (n == 3) {
                shape.corners.insert(shape.corners.end(), f, f + 3);
                shape.material_ids.push_back(material);
            } else if (n == 4) {
                bool bad = false;
                for (int k = 0; k < 4; ++k) bad |= (size_t)f[k].v >= nv;
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
                clip_ears(f, n, v, nv, shape, material);
            }
        }
    return true;
}

// LoadObj (tiny_obj_loader.h:2584-3146) for the statements that reach the triangle buffer: v vn vt f usemtl mtllib g o
bool read_obj(const std::string &file, const std::string &mtl_dir, ObjFile &out, std::string &log, std::string &err) {
    std::string buf;
    {
        std::ifstream is(file, std::ios::binary);
        if (!is) {
            err = "Cannot open file [" + file + "]";
            return false;
        }
        is.seekg(0, std::ios::end);
        const std::streamoff size = is.tellg();
        is.seekg(0);
        buf.resize(size > 0 ? (size_t)size : 0);
        if (!buf.empty() && !is.read(&buf[0], (std::streamsize)buf.size())) {
            err = "Cannot read file [" + file + "]";
            return false;
        }
    }
    // chunks of at least 1 MB, cut after a '\n' (PTB_LOADER_THREADS=n forces n chunks whatever the size: tests)
    size_t want = std::max<size_t>(1, std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), buf.size() >> 20));
    if (const char *forced = getenv("PTB_LOADER_THREADS"))
        if (atoi(forced) > 0) want = (size_t)std::min(atoi(forced), 256);
    std::vector<Chunk> chunks;
    const char *base = buf.data(), *end = base + buf.size();
    for (size_t i = 0; i < want; ++i) {
        const char *b = chunks.empty() ? base : chunks.back().e;
        const char *e = i + 1 == want ? end : base + buf.size() * (i + 1) / want;
        if (e < b) e = b;
        while (e < end && e > base && e[-1] != '\n') ++e;
        Chunk c;
        c.b = b;
        c.e = e;
        chunks.push_back(std::move(c));
    }
    run_parallel(chunks.size(), [&](size_t i) { scan_chunk(chunks[i]); });
    size_t nv = 0, nvn = 0, nvt = 0, lines = 0;
    for (Chunk &c : chunks) {
        c.v0 = nv; c.vn0 = nvn; c.vt0 = nvt; c.line0 = lines;
        nv += c.nv; nvn += c.nvn; nvt += c.nvt; lines += c.lines;
    }
    out.v.resize(3 * nv);
    out.vn.resize(3 * nvn);
    out.vt.resize(2 * nvt);
    run_parallel(chunks.size(), [&](size_t i) { parse_chunk(chunks[i], out); });
    for (const Chunk &c : chunks)
        if (!c.err.empty()) {
            err = c.err;
            return false;
        }

    // the stateful statements, replayed in file order
    std::map<std::string, int> by_name;
    std::set<std::string> mtl_seen;
    std::vector<FaceRange> pending;
    Shape shape;
    std::string name;
    int material = -1;
    auto flush = [&](size_t nv_now) { return flush_faces(shape, pending, material, name, out.v, nv_now, log, err); };
    for (const Chunk &c : chunks) {
        uint32_t cursor = 0;
        for (const Control &k : c.ctl) {
            if (k.face > cursor) pending.push_back(FaceRange{&c, cursor, k.face});
            cursor = k.face;
            const size_t nv_now = c.v0 + k.nv_local;
            const char *t = k.text.c_str();
            if (k.kind == S_USEMTL) {
                Cursor w{t + 6};
                const std::string m = w.word();
                const auto it = by_name.find(m);
                const int id = it != by_name.end() ? it->second : -1;
                if (it == by_name.end()) log += "material [ '" + m + "' ] not found in .mtl\n";
                if (id != material) { // faces so far keep the old material; the shape goes on (:2839-2847)
                    flush(nv_now);
                    if (!err.empty()) return false;
                    pending.clear();
                    material = id;
                }
            } else if (k.kind == S_MTLLIB) {
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
            } else { // g, o: the shape so far is complete
                flush(nv_now);
                if (!err.empty()) return false;
                if (!shape.corners.empty()) out.shapes.push_back(std::move(shape));
                shape = Shape();
                pending.clear();
                if (k.kind == S_O) {
                    name = std::string(t + 2); // the rest of the line, spaces included (:2989-2992)
                } else {
                    // the statement letter itself is the first "name" the reference collects, the group name starts at the second
                    Cursor w{t};
                    std::vector<std::string> words;
                    while (!is_eol(*w.p)) {
                        words.push_back(w.word());
                        w.p += strspn(w.p, " \t\r");
                    }
                    name.clear();
                    for (size_t i = 1; i < words.size(); ++i) name += (i > 1 ? " " : "") + words[i];
                }
            }
        }
        if (c.nfaces > cursor) pending.push_back(FaceRange{&c, cursor, (uint32_t)c.nfaces});
    }
    const bool some = flush(nv);
    if (!err.empty()) return false;
    if (some || !shape.corners.empty()) out.shapes.push_back(std::move(shape));
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
        if (flags & kSceneIgnoreTextures) tex = {};
        sc.materials.push_back(m);
        sc.textures.push_back(tex);
        sc.log += "loaded material " + me.name + "\n";
    }

    const size_t nv = obj.v.size() / 3, nn = obj.vn.size() / 3, nt = obj.vt.size() / 2;
    for (const Shape &sh : obj.shapes) { // :89-172
        // a group whose faces were all dropped (degenerate, or quads with an invalid index) is still pushed by tinyobjloader
        // with an empty mesh; the reference then reads material_ids[0] of an empty vector (object.cpp:92, undefined
        // behaviour).  Such a shape carries no triangle: skipped here.
        if (sh.material_ids.empty() || sh.corners.empty()) continue;
        sc.log += "loaded: " + sh.name + "\n";
        const int start = (int)sc.triangles();
        int shape_mat = sh.material_ids[0];
        shape_mat = shape_mat >= 0 ? shape_mat + first_mat : default_mat;
        std::vector<uint32_t> used; // the shape's materials in order of first use (:125-131)
        std::vector<bool> seen(obj.materials.size(), false);
        const size_t ntri = sh.corners.size() / 3;
        for (size_t i = 0; i < ntri; ++i) {
            const int m = sh.material_ids[i];
            if (m >= 0 && !seen[m]) {
                seen[m] = true;
                used.push_back((uint32_t)(m + first_mat));
            }
        }
        const size_t base = sc.triangles();
        sc.pos9.resize(9 * (base + ntri));
        sc.nrm9.resize(9 * (base + ntri));
        sc.uv6.resize(6 * (base + ntri));
        sc.mat.resize(base + ntri);
        std::vector<float> areas(ntri);
        const size_t workers = std::max<size_t>(1, std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), ntri >> 14));
        std::vector<int> bad(workers, 0);
        run_parallel(workers, [&](size_t w) {
            for (size_t i = ntri * w / workers; i < ntri * (w + 1) / workers; ++i) {
                float *p = &sc.pos9[9 * (base + i)], *n = &sc.nrm9[9 * (base + i)], *uv = &sc.uv6[6 * (base + i)];
                bool have_normals = true;
                for (int j = 0; j < 3; ++j) {
                    const Corner &c = sh.corners[3 * i + j];
                    if (c.v < 0 || (size_t)c.v >= nv) {
                        bad[w] = 1;
                        return;
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
                const int m = sh.material_ids[i];
                sc.mat[base + i] = m >= 0 ? m + first_mat : shape_mat;
                areas[i] = surface_area(p);
            }
        });
        for (int b : bad)
            if (b) {
                err = "shape " + sh.name + ": vertex index out of bounds";
                return false;
            }
        float area = 0.0f;
        for (size_t i = 0; i < ntri; ++i) area += areas[i]; // summed in triangle order like the reference (:151-154)
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
    return decode_textures(sc, err);
}

bool decode_textures(LoadedScene &sc, std::string &err) {
    sc.texels.clear();
    sc.mat_tex.assign(2 * sc.materials.size(), -1);
    for (size_t i = 0; i < sc.materials.size(); ++i)
        for (int k = 0; k < 2; ++k) {
            const std::string &path = sc.textures[i][k];
            if (path.empty()) continue;
            std::vector<unsigned char> px;
            LoadedScene::Texels t;
            if (!decode_image(path, px, t.width, t.height, err)) {
                err = "material " + std::string(sc.materials[i].name) + ": " + err +
                      " — convert the texture to PNG or TGA, hand texels to ptb_upload_textures yourself, or load with the ignore-textures flag";
                return false;
            }
            t.rgba.resize(px.size());
            for (size_t j = 0; j < px.size(); ++j) t.rgba[j] = px[j] / 255.0f; // texture.cpp:44-47
            sc.mat_tex[2 * i + k] = (int32_t)sc.texels.size();
            sc.texels.push_back(std::move(t));
            sc.log += "loaded texture " + path + "\n";
        }
    return true;
}

bool write_ptbs(const std::string &path, const LoadedScene &sc, std::string &err) {
    ptbs_header h{};
    h.magic = PTBS_MAGIC;
    h.version = sc.texels.empty() ? PTBS_VERSION : PTBS_VERSION_TEXTURES;
    h.reserved[0] = (uint32_t)sc.texels.size();
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
    if (!sc.texels.empty()) {
        put(sc.mat_tex.data(), sc.mat_tex.size() * 4);
        for (const LoadedScene::Texels &t : sc.texels) {
            const ptbs_texture_hdr th{t.width, t.height};
            put(&th, sizeof th);
            put(t.rgba.data(), t.rgba.size() * 4);
        }
    }
    ok = fclose(f) == 0 && ok;
    if (!ok) err = "short write to " + path;
    return ok;
}

} // namespace ptb
