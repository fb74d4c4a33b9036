// OBJ/MTL ingest (host, product code): one pass over the file bytes, no per-token streams.
//
// Two reader conventions are supported because the reference has two readers whose results differ
// (SURVEY.md A.6), and triangle ids / materials must match the one the caller replaces:
//
//   MCPT_LOADER_TINYOBJ    what Quin::Utils::QuinModel gets from tinyobjloader 1.1.1
//       (MCRT/QuinEngine/Utils/Structure.hpp:9-12 -> 3rdparty/include/tiny_obj_loader.h:1713-2050,
//        triangulation :1017-1226, MTL :1237-1621, number grammar :524-637) flattened as
//        ShaderResource.hpp:88-104 / :204-215 do: 0-based indices, shapes in file order, polygons
//        ear-clipped in the projection plane of their first non-degenerate corner, material per face,
//        Tr = 1 - dissolve.
//   MCPT_LOADER_OBJREADER  PW::FileReader::ObjModel (CVMCTracer/CVMCTracer/Framework/ObjReader.cpp:8-259)
//        as PW::Tracer::CreateGeometry consumes it (CUTracer.cu:225-314): 1-based indices with a dummy
//        slot 0 in every table, fan triangulation (v0, v_{k-1}, v_k), `\` line continuation, a `Ks` line
//        forces Ns = 2, and every triangle of a group takes the material of the group's FIRST triangle.
//
// Numbers: tinyobj converts decimals with its own digit-accumulating routine in double precision
// (not strtod), so to reproduce its float tables bit for bit the same procedure is restated in
// decimal_tinyobj(); the ObjReader convention uses operator>>(float), i.e. correctly rounded strtof.
#include "obj_loader.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>

#include "../../include/mcpt.h"

namespace mcpt {
namespace {

bool slurp(const std::string& path, std::string& out) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? static_cast<size_t>(n) : 0);
    const bool ok = out.empty() || std::fread(&out[0], 1, out.size(), f) == out.size();
    std::fclose(f);
    return ok;
}

inline bool blank(char c) { return c == ' ' || c == '\t'; }
inline bool digit(char c) { return static_cast<unsigned>(c - '0') < 10u; }

// Cursor over one line [p, e): e points at the line terminator (or the end of the buffer).
struct Line {
    const char* p;
    const char* e;
    void skip_blank() { while (p < e && blank(*p)) ++p; }
    bool done() const { return p >= e; }
    // next whitespace-delimited token; empty range at end of line
    bool token(const char*& b, const char*& t) {
        while (p < e && (blank(*p) || *p == '\r')) ++p;
        b = p;
        while (p < e && !blank(*p) && *p != '\r') ++p;
        t = p;
        return t > b;
    }
};

// Splits the buffer into lines at '\n' (a preceding '\r' is dropped), like tinyobj's safeGetline (:418-451).
struct LineReader {
    const char* p;
    const char* end;
    explicit LineReader(const std::string& s) : p(s.data()), end(s.data() + s.size()) {}
    bool next(Line& l) {
        if (p >= end) return false;
        const char* nl = static_cast<const char*>(std::memchr(p, '\n', static_cast<size_t>(end - p)));
        const char* e = nl ? nl : end;
        l.p = p;
        l.e = (e > p && e[-1] == '\r') ? e - 1 : e;
        p = nl ? nl + 1 : end;
        return true;
    }
};

// tiny_obj_loader.h:524-637 as a procedure: sign, integer digits accumulated as m = 10 m + d, the k-th
// fractional digit added as d * 10^-k (table for k < 8, pow beyond), optional exponent applied as
// ldexp(m * 5^x, x).  Returns false (value untouched) for anything that does not start with a digit.
bool decimal_tinyobj(const char* s, const char* e, double& value) {
    static const double kNeg10[8] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
    if (s >= e) return false;
    double sign = 1.0;
    if (*s == '+' || *s == '-') { if (*s == '-') sign = -1.0; ++s; }
    else if (!digit(*s)) return false;
    double m = 0.0;
    int nint = 0;
    while (s < e && digit(*s)) { m = m * 10 + (*s - '0'); ++s; ++nint; }
    if (nint == 0) return false;
    int exponent = 0;
    if (s < e && *s == '.') {
        ++s;
        for (int k = 1; s < e && digit(*s); ++s, ++k) m += (*s - '0') * (k < 8 ? kNeg10[k] : std::pow(10.0, -k));
    } else if (s < e && *s != 'e' && *s != 'E') {
        value = sign * m;
        return true;
    }
    if (s < e && (*s == 'e' || *s == 'E')) {
        ++s;
        int esign = 1;
        if (s < e && (*s == '+' || *s == '-')) { if (*s == '-') esign = -1; ++s; }
        else if (!(s < e && digit(*s))) return false;
        int nexp = 0;
        while (s < e && digit(*s)) { exponent = exponent * 10 + (*s - '0'); ++s; ++nexp; }
        if (nexp == 0) return false;
        exponent *= esign;
    }
    value = sign * (exponent ? std::ldexp(m * std::pow(5.0, exponent), exponent) : m);
    return true;
}

// parseReal (:639-647): the next blank-delimited field, `dflt` when it is not a number.
float real_tinyobj(Line& l, double dflt = 0.0) {
    l.skip_blank();
    const char* b = l.p;
    while (l.p < l.e && !blank(*l.p) && *l.p != '\r') ++l.p;
    double v = dflt;
    decimal_tinyobj(b, l.p, v);
    return static_cast<float>(v);
}

// atoi on a bounded range
int int_at(const char* s, const char* e) {
    while (s < e && blank(*s)) ++s;
    bool neg = false;
    if (s < e && (*s == '+' || *s == '-')) { neg = *s == '-'; ++s; }
    long v = 0;
    while (s < e && digit(*s)) { v = v * 10 + (*s - '0'); ++s; }
    return static_cast<int>(neg ? -v : v);
}

struct Corner {
    int v, n;  // 0-based, -1 = absent
};

// fixIndex (:458-479): 1-based -> 0-based, negative = relative to the current count, 0 is an error.
bool fix_index(int idx, int count, int& out) {
    if (idx > 0) { out = idx - 1; return true; }
    if (idx < 0) { out = count + idx; return true; }
    return false;
}

// parseTriple (:774-825): i, i/j, i//k, i/j/k
bool corner_tinyobj(Line& l, int nV, int nN, Corner& c) {
    auto field_end = [&]() { while (l.p < l.e && *l.p != '/' && !blank(*l.p) && *l.p != '\r') ++l.p; };
    c.v = -1; c.n = -1;
    if (!fix_index(int_at(l.p, l.e), nV, c.v)) return false;
    field_end();
    if (l.done() || *l.p != '/') return true;
    ++l.p;
    if (!l.done() && *l.p == '/') {  // i//k
        ++l.p;
        if (!fix_index(int_at(l.p, l.e), nN, c.n)) return false;
        field_end();
        return true;
    }
    int vt;
    if (!fix_index(int_at(l.p, l.e), 1 << 30, vt)) return false;  // texture index: validated, not kept
    field_end();
    if (l.done() || *l.p != '/') return true;
    ++l.p;
    if (!fix_index(int_at(l.p, l.e), nN, c.n)) return false;
    field_end();
    return true;
}

// Crossing-number point-in-triangle test in the projection plane (tiny_obj_loader.h:1002-1015).
bool inside_tri(const float x[3], const float y[3], float tx, float ty) {
    bool in = false;
    for (int i = 0, j = 2; i < 3; j = i++)
        if (((y[i] > ty) != (y[j] > ty)) && (tx < (x[j] - x[i]) * (ty - y[i]) / (y[j] - y[i]) + x[i])) in = !in;
    return in;
}

// Ear clipping of one polygon exactly as tinyobj does it (:1041-1190): projection axes from the first
// corner whose cross product is non-degenerate, signed area in that plane, ears taken at a moving
// guess vertex, at most 10 fruitless sweeps.  Appends corner triples to `out`.
void triangulate_tinyobj(const std::vector<Corner>& poly, const std::vector<float>& V, std::vector<Corner>& out) {
    size_t n = poly.size();
    if (n < 3) return;
    if (n == 3) { out.insert(out.end(), poly.begin(), poly.end()); return; }
    size_t ax0 = 1, ax1 = 2;
    for (size_t k = 0; k < n; ++k) {
        const float* a = &V[3 * static_cast<size_t>(poly[k].v)];
        const float* b = &V[3 * static_cast<size_t>(poly[(k + 1) % n].v)];
        const float* c = &V[3 * static_cast<size_t>(poly[(k + 2) % n].v)];
        const float e0x = b[0] - a[0], e0y = b[1] - a[1], e0z = b[2] - a[2];
        const float e1x = c[0] - b[0], e1y = c[1] - b[1], e1z = c[2] - b[2];
        const float cx = std::fabs(e0y * e1z - e0z * e1y), cy = std::fabs(e0z * e1x - e0x * e1z),
                    cz = std::fabs(e0x * e1y - e0y * e1x);
        const float eps = 1.192092896e-07f;
        if (cx > eps || cy > eps || cz > eps) {
            if (!(cx > cy && cx > cz)) {
                ax0 = 0;
                if (cz > cx && cz > cy) ax1 = 1;
            }
            break;
        }
    }
    float area = 0;
    for (size_t k = 0; k < n; ++k) {
        const float* a = &V[3 * static_cast<size_t>(poly[k].v)];
        const float* b = &V[3 * static_cast<size_t>(poly[(k + 1) % n].v)];
        area += (a[ax0] * b[ax1] - a[ax1] * b[ax0]) * 0.5f;
    }
    std::vector<Corner> rest = poly;
    int rounds = 10;
    size_t guess = 0;
    while (rest.size() > 3 && rounds > 0) {
        n = rest.size();
        if (guess >= n) { --rounds; guess -= n; }
        Corner c[3];
        float x[3], y[3];
        for (size_t k = 0; k < 3; ++k) {
            c[k] = rest[(guess + k) % n];
            x[k] = V[3 * static_cast<size_t>(c[k].v) + ax0];
            y[k] = V[3 * static_cast<size_t>(c[k].v) + ax1];
        }
        const float cross = (x[1] - x[0]) * (y[2] - y[1]) - (y[1] - y[0]) * (x[2] - x[1]);
        if (cross * area < 0.0f) { ++guess; continue; }  // reflex corner
        bool blocked = false;
        for (size_t o = 3; o < n && !blocked; ++o) {
            const size_t vi = static_cast<size_t>(rest[(guess + o) % n].v);
            blocked = inside_tri(x, y, V[3 * vi + ax0], V[3 * vi + ax1]);
        }
        if (blocked) { ++guess; continue; }
        out.push_back(c[0]); out.push_back(c[1]); out.push_back(c[2]);
        rest.erase(rest.begin() + static_cast<long>((guess + 1) % n));
    }
    if (rest.size() == 3) out.insert(out.end(), rest.begin(), rest.end());
}

std::string rest_of_line(const Line& l, size_t skip) {
    const char* b = l.p + skip;
    return b < l.e ? std::string(b, l.e) : std::string();
}

bool starts(const Line& l, const char* word) {
    const size_t n = std::strlen(word);
    return static_cast<size_t>(l.e - l.p) > n && std::memcmp(l.p, word, n) == 0 && blank(l.p[n]);
}

// ---------------------------------------------------------------------------------------------
// tinyobj convention
// ---------------------------------------------------------------------------------------------
struct MtlTiny {
    std::string name;
    float Ka[3] = {0, 0, 0}, Kd[3] = {0, 0, 0}, Ks[3] = {0, 0, 0};
    float shininess = 1.0f, dissolve = 1.0f, ior = 1.0f;  // InitMaterial, tiny_obj_loader.h:963-983
};

// LoadMtl (:1237-1621): the keys the tracer consumes; everything else is skipped.
void load_mtl_tinyobj(const std::string& text, std::vector<MtlTiny>& mats, std::map<std::string, int>& byName) {
    MtlTiny cur;
    bool hasD = false;
    LineReader lr(text);
    Line l;
    auto flush = [&]() {
        byName.insert({cur.name, static_cast<int>(mats.size())});  // first definition of a name wins
        mats.push_back(cur);
    };
    while (lr.next(l)) {
        while (l.e > l.p && blank(l.e[-1])) --l.e;  // trailing blanks are trimmed in MTL files (:1256)
        l.skip_blank();
        if (l.done() || *l.p == '#') continue;
        auto rgb = [&](float* dst) { l.p += 2; dst[0] = real_tinyobj(l); dst[1] = real_tinyobj(l); dst[2] = real_tinyobj(l); };
        if (starts(l, "newmtl")) {
            if (!cur.name.empty()) flush();
            cur = MtlTiny();
            hasD = false;
            cur.name = rest_of_line(l, 7);
        } else if (starts(l, "Ka")) rgb(cur.Ka);
        else if (starts(l, "Kd")) rgb(cur.Kd);
        else if (starts(l, "Ks")) rgb(cur.Ks);
        else if (starts(l, "Ni")) { l.p += 2; cur.ior = real_tinyobj(l); }
        else if (starts(l, "Ns")) { l.p += 2; cur.shininess = real_tinyobj(l); }
        else if (starts(l, "d")) { l.p += 1; cur.dissolve = real_tinyobj(l); hasD = true; }
        else if (starts(l, "Tr")) { l.p += 2; if (!hasD) cur.dissolve = 1.0f - real_tinyobj(l); }  // `d` wins (:1398-1409)
    }
    flush();  // the last material is pushed even when unnamed (:1613-1616)
}

bool load_tinyobj(const std::string& objPath, const std::string& mtlBase, HostScene& out, std::string& err) {
    std::string text;
    if (!slurp(objPath, text)) { err = "Cannot open file [" + objPath + "]"; return false; }
    std::vector<float> V, N;
    size_t numVt = 0;
    std::vector<MtlTiny> mats;
    std::map<std::string, int> matByName;
    int material = -1;
    std::vector<std::vector<Corner>> group;      // faces since the last flush
    std::vector<Corner> shapeCorners;            // triangulated corners of the shape being assembled
    std::vector<int> shapeMat;                   // one per triangle
    std::vector<Corner> allCorners;
    std::vector<int> allMat;
    auto flush_group = [&]() -> bool {           // exportFaceGroupToShape (:1017-1225)
        if (group.empty()) return false;
        for (const auto& face : group) {
            const size_t before = shapeCorners.size();
            triangulate_tinyobj(face, V, shapeCorners);
            shapeMat.insert(shapeMat.end(), (shapeCorners.size() - before) / 3, material);
        }
        group.clear();
        return true;
    };
    auto end_shape = [&](bool keep) {
        if (keep) {
            allCorners.insert(allCorners.end(), shapeCorners.begin(), shapeCorners.end());
            allMat.insert(allMat.end(), shapeMat.begin(), shapeMat.end());
        }
        shapeCorners.clear();
        shapeMat.clear();
    };
    LineReader lr(text);
    Line l;
    size_t lineNo = 0;
    while (lr.next(l)) {
        ++lineNo;
        l.skip_blank();
        if (l.done() || *l.p == '#') continue;
        const char c0 = l.p[0], c1 = l.e - l.p > 1 ? l.p[1] : '\0', c2 = l.e - l.p > 2 ? l.p[2] : '\0';
        if (c0 == 'v' && blank(c1)) {
            l.p += 2;
            for (int i = 0; i < 3; ++i) V.push_back(real_tinyobj(l));
        } else if (c0 == 'v' && c1 == 'n' && blank(c2)) {
            l.p += 3;
            for (int i = 0; i < 3; ++i) N.push_back(real_tinyobj(l));
        } else if (c0 == 'v' && c1 == 't' && blank(c2)) {
            ++numVt;
        } else if (c0 == 'f' && blank(c1)) {
            l.p += 2;
            l.skip_blank();
            std::vector<Corner> face;
            while (!l.done() && *l.p != '\r') {
                Corner c;
                if (!corner_tinyobj(l, static_cast<int>(V.size() / 3), static_cast<int>(N.size() / 3), c)) {
                    err = objPath + ":" + std::to_string(lineNo) + ": bad face index (zero is not allowed)";
                    return false;
                }
                if (c.v < 0 || static_cast<size_t>(c.v) >= V.size() / 3) {
                    err = objPath + ":" + std::to_string(lineNo) + ": vertex index out of range";
                    return false;
                }
                face.push_back(c);
                while (!l.done() && (blank(*l.p) || *l.p == '\r')) ++l.p;
            }
            group.push_back(std::move(face));
        } else if (starts(l, "usemtl")) {
            const auto it = matByName.find(rest_of_line(l, 7));
            const int id = it == matByName.end() ? -1 : it->second;
            if (id != material) { flush_group(); material = id; }  // per-face materials inside one shape (:1842-1850)
        } else if (starts(l, "mtllib")) {
            std::string names = rest_of_line(l, 7);
            size_t pos = 0;
            while (pos <= names.size()) {   // space-separated candidates, first readable file wins (:1860-1891)
                size_t sp = names.find(' ', pos);
                if (sp == std::string::npos) sp = names.size();
                std::string mtlText;
                if (sp > pos && slurp(mtlBase + names.substr(pos, sp - pos), mtlText)) {
                    load_mtl_tinyobj(mtlText, mats, matByName);
                    break;
                }
                pos = sp + 1;
            }
        } else if (c0 == 'g' && blank(c1)) {
            flush_group();
            end_shape(!shapeCorners.empty());   // (:1900-1907)
        } else if (c0 == 'o' && blank(c1)) {
            end_shape(flush_group());           // an `o` keeps the shape only if faces were pending (:1934-1939)
        }
    }
    const bool pending = flush_group();
    end_shape(pending || !shapeCorners.empty());
    (void)numVt;

    if (allCorners.empty()) { err = objPath + ": no faces"; return false; }
    // ---- flatten: ShaderResource.hpp:88-104 (triangles), :204-215 (materials) ----
    out = HostScene();
    out.vertices = std::move(V);
    out.normals = std::move(N);
    for (const auto& m : mats) {
        const float rec[12] = {m.Ka[0], m.Ka[1], m.Ka[2], m.Kd[0], m.Kd[1], m.Kd[2], m.Ks[0], m.Ks[1], m.Ks[2],
                               m.shininess, 1.0f - m.dissolve, m.ior};
        out.materials.insert(out.materials.end(), rec, rec + 12);
    }
    int fallbackMat = -1;
    const size_t nT = allCorners.size() / 3;
    out.triangles.reserve(7 * nT);
    for (size_t t = 0; t < nT; ++t) {
        const Corner* c = &allCorners[3 * t];
        uint32_t n[3];
        if (c[0].n < 0 || c[1].n < 0 || c[2].n < 0) {
            // The shader would read csnormal[-1] (undefined).  Product behaviour: the geometric normal.
            const float *a = &out.vertices[3 * static_cast<size_t>(c[0].v)], *b = &out.vertices[3 * static_cast<size_t>(c[1].v)],
                        *d = &out.vertices[3 * static_cast<size_t>(c[2].v)];
            const float ux = b[0] - a[0], uy = b[1] - a[1], uz = b[2] - a[2], vx = d[0] - a[0], vy = d[1] - a[1], vz = d[2] - a[2];
            float nx = uy * vz - uz * vy, ny = uz * vx - ux * vz, nz = ux * vy - uy * vx;
            const float len = std::sqrt(nx * nx + ny * ny + nz * nz);
            if (len > 0) { nx /= len; ny /= len; nz /= len; } else { nx = 0; ny = 1; nz = 0; }
            n[0] = n[1] = n[2] = static_cast<uint32_t>(out.normals.size() / 3);
            out.normals.push_back(nx); out.normals.push_back(ny); out.normals.push_back(nz);
        } else {
            for (int k = 0; k < 3; ++k) n[k] = static_cast<uint32_t>(c[k].n);
        }
        int m = allMat[t];
        if (m < 0) {  // no usemtl / unknown name: tinyobj's default material (all-zero Kd -> black Lambert)
            if (fallbackMat < 0) {
                fallbackMat = static_cast<int>(out.materials.size() / 12);
                const float rec[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 1.0f, 0.0f, 1.0f};
                out.materials.insert(out.materials.end(), rec, rec + 12);
            }
            m = fallbackMat;
        }
        const uint32_t rec[7] = {static_cast<uint32_t>(c[0].v), static_cast<uint32_t>(c[1].v), static_cast<uint32_t>(c[2].v),
                                 n[0], n[1], n[2], static_cast<uint32_t>(m)};
        out.triangles.insert(out.triangles.end(), rec, rec + 7);
    }
    return true;
}

// ---------------------------------------------------------------------------------------------
// PW::FileReader::ObjModel convention
// ---------------------------------------------------------------------------------------------
struct MtlPw {
    std::string name;
    float Ka[3] = {0, 0, 0}, Kd[3] = {0, 0, 0}, Ks[3] = {0, 0, 0};
    double Ns = 1, Tr = 0, Ni = 1;  // ObjReader.hpp:22
};

// Joins physical lines that end in a backslash (ObjReader.cpp:27-33) and yields whitespace-separated tokens.
struct PwLines {
    LineReader lr;
    std::string joined;
    explicit PwLines(const std::string& s) : lr(s) {}
    bool next(Line& l) {
        Line raw;
        if (!lr.next(raw)) return false;
        if (raw.e > raw.p && raw.e[-1] == '\\') {
            joined.clear();
            while (raw.e > raw.p && raw.e[-1] == '\\') {
                joined.append(raw.p, raw.e - 1);
                if (!lr.next(raw)) { raw.p = raw.e = nullptr; break; }
            }
            if (raw.p) joined.append(raw.p, raw.e);
            l.p = joined.data();
            l.e = joined.data() + joined.size();
        } else {
            l = raw;
        }
        return true;
    }
};

bool tok_is(const char* b, const char* t, const char* w) {
    const size_t n = std::strlen(w);
    return static_cast<size_t>(t - b) == n && std::memcmp(b, w, n) == 0;
}

// operator>>(float): correctly rounded; a missing or malformed field reads as 0.
float float_pw(Line& l) {
    const char *b, *t;
    if (!l.token(b, t)) return 0.0f;
    const std::string s(b, t);
    return std::strtof(s.c_str(), nullptr);
}
double double_pw(Line& l) {
    const char *b, *t;
    if (!l.token(b, t)) return 0.0;
    const std::string s(b, t);
    return std::strtod(s.c_str(), nullptr);
}

// parseFaceVertex (ObjReader.hpp:92-137): v, v/t, v//n, v/t/n; absent fields are 0 (the dummy slot).
bool corner_pw(const char* b, const char* t, int& v, int& n) {
    v = 0; n = 0;
    if (b >= t || !(digit(*b) || *b == '-' || *b == '+')) return false;
    v = int_at(b, t);
    const char* s = static_cast<const char*>(std::memchr(b, '/', static_cast<size_t>(t - b)));
    if (!s) return true;
    ++s;
    if (s < t && *s == '/') {  // v//n
        ++s;
        if (s >= t) return false;
        n = int_at(s, t);
        return true;
    }
    const char* s2 = s < t ? static_cast<const char*>(std::memchr(s, '/', static_cast<size_t>(t - s))) : nullptr;
    if (!s2) return true;      // v/t
    ++s2;
    if (s2 >= t) return false;
    n = int_at(s2, t);
    return true;
}

int find_mtl_pw(const std::vector<MtlPw>& mats, const std::string& name) {
    for (size_t i = 1; i < mats.size(); ++i)
        if (mats[i].name == name) return static_cast<int>(i);
    return 0;  // the unnamed default material in slot 0 (ObjReader.hpp:78-89)
}

bool load_mtl_pw(const std::string& path, std::vector<MtlPw>& mats, std::string& err) {
    std::string text;
    if (!slurp(path, text)) { err = "Can't open file " + path; return false; }
    PwLines lines(text);
    Line l;
    int cur = 0;
    while (lines.next(l)) {
        const char *b, *t;
        if (!l.token(b, t) || *b == '#') continue;
        if (tok_is(b, t, "newmtl")) {
            const char *nb, *nt;
            l.token(nb, nt);
            const std::string name(nb, nt);
            cur = find_mtl_pw(mats, name);
            if (cur == 0) { mats.emplace_back(); mats.back().name = name; cur = static_cast<int>(mats.size()) - 1; }
        } else if (tok_is(b, t, "Ka")) { for (int i = 0; i < 3; ++i) mats[cur].Ka[i] = float_pw(l); }
        else if (tok_is(b, t, "Kd")) { for (int i = 0; i < 3; ++i) mats[cur].Kd[i] = float_pw(l); }
        else if (tok_is(b, t, "Ks")) { for (int i = 0; i < 3; ++i) mats[cur].Ks[i] = float_pw(l); mats[cur].Ns = 2; }  // ObjReader.cpp:232
        else if (tok_is(b, t, "Ns")) mats[cur].Ns = double_pw(l);
        else if (tok_is(b, t, "Tr")) mats[cur].Tr = double_pw(l);
        else if (tok_is(b, t, "Ni")) mats[cur].Ni = double_pw(l);
    }
    return true;
}

bool load_objreader(const std::string& objPath, HostScene& out, std::string& err) {
    std::string text;
    if (!slurp(objPath, text)) { err = "Can't open file " + objPath; return false; }
    std::vector<float> V(3, 0.0f), N(3, 0.0f);      // slot 0 of every table is a dummy (ObjReader.hpp:40-54)
    std::vector<MtlPw> mats(1);
    struct Tri { int v[3], n[3], mat, group; };
    std::vector<Tri> tris;
    std::map<std::string, int> groupId;
    std::vector<int> groupFirstTri;                  // first triangle of each group, -1 = none yet
    auto group_of = [&](const std::string& name) {
        auto it = groupId.find(name);
        if (it != groupId.end()) return it->second;
        groupId[name] = static_cast<int>(groupFirstTri.size());
        groupFirstTri.push_back(-1);
        return static_cast<int>(groupFirstTri.size()) - 1;
    };
    int group = group_of("default"), material = 0;
    const size_t slash = objPath.find_last_of('/');
    const std::string dir = slash == std::string::npos ? std::string(".") : objPath.substr(0, slash);
    PwLines lines(text);
    Line l;
    size_t lineNo = 0;
    while (lines.next(l)) {
        ++lineNo;
        const char *b, *t;
        if (!l.token(b, t) || *b == '#') continue;
        if (tok_is(b, t, "v")) { for (int i = 0; i < 3; ++i) V.push_back(float_pw(l)); }
        else if (tok_is(b, t, "vn")) { for (int i = 0; i < 3; ++i) N.push_back(float_pw(l)); }
        else if (tok_is(b, t, "mtllib")) {
            const char *nb, *nt;
            if (l.token(nb, nt) && !load_mtl_pw(dir + "/" + std::string(nb, nt), mats, err)) return false;
        } else if (tok_is(b, t, "g")) {
            const char *nb, *nt;
            if (l.token(nb, nt)) group = group_of(std::string(nb, nt));
        } else if (tok_is(b, t, "usemtl")) {
            const char *nb, *nt;
            if (l.token(nb, nt)) material = find_mtl_pw(mats, std::string(nb, nt));
        } else if (tok_is(b, t, "f")) {
            int k = 0, v0 = 0, n0 = 0, vp = 0, np = 0;
            const char *cb, *ct;
            Tri tri{};
            while (l.token(cb, ct)) {
                int v, n;
                if (!corner_pw(cb, ct, v, n)) { err = objPath + ":" + std::to_string(lineNo) + ": Invalid OBJ file!"; return false; }
                if (v < 0 || static_cast<size_t>(v) >= V.size() / 3 || n < 0 || static_cast<size_t>(n) >= N.size() / 3) {
                    err = objPath + ":" + std::to_string(lineNo) + ": index out of range (this reader has no relative indices)";
                    return false;
                }
                if (k == 0) { v0 = v; n0 = n; }
                if (k >= 2) {  // fan: (v0, v_{k-1}, v_k)  (ObjReader.cpp:84-104)
                    tri.v[0] = v0; tri.v[1] = vp; tri.v[2] = v; tri.n[0] = n0; tri.n[1] = np; tri.n[2] = n;
                    tri.mat = material; tri.group = group;
                    if (groupFirstTri[static_cast<size_t>(group)] < 0) groupFirstTri[static_cast<size_t>(group)] = static_cast<int>(tris.size());
                    tris.push_back(tri);
                }
                vp = v; np = n;
                ++k;
            }
            if (k < 3) { err = objPath + ":" + std::to_string(lineNo) + ": face with fewer than 3 vertices"; return false; }
        }
    }
    if (tris.empty()) { err = objPath + ": no faces"; return false; }
    out = HostScene();
    out.vertices = std::move(V);
    out.normals = std::move(N);
    for (const auto& m : mats) {   // Geometry::Material (Geometry.h:14-22), doubles narrowed as CUTracer.cu:306-308 does
        const float rec[12] = {m.Ka[0], m.Ka[1], m.Ka[2], m.Kd[0], m.Kd[1], m.Kd[2], m.Ks[0], m.Ks[1], m.Ks[2],
                               static_cast<float>(m.Ns), static_cast<float>(m.Tr), static_cast<float>(m.Ni)};
        out.materials.insert(out.materials.end(), rec, rec + 12);
    }
    out.triangles.reserve(7 * tris.size());
    for (const auto& tr : tris) {
        // CreateGeometry binds ONE material per group: that of the group's first triangle (CUTracer.cu:300-308)
        const int mat = tris[static_cast<size_t>(groupFirstTri[static_cast<size_t>(tr.group)])].mat;
        const uint32_t rec[7] = {static_cast<uint32_t>(tr.v[0]), static_cast<uint32_t>(tr.v[1]), static_cast<uint32_t>(tr.v[2]),
                                 static_cast<uint32_t>(tr.n[0]), static_cast<uint32_t>(tr.n[1]), static_cast<uint32_t>(tr.n[2]),
                                 static_cast<uint32_t>(mat)};
        out.triangles.insert(out.triangles.end(), rec, rec + 7);
    }
    return true;
}

}  // namespace

bool load_obj_scene(const std::string& obj_path, const std::string& mtl_base, int loader, HostScene& out, std::string& err) {
    if (loader == MCPT_LOADER_TINYOBJ) return load_tinyobj(obj_path, mtl_base, out, err);
    if (loader == MCPT_LOADER_OBJREADER) return load_objreader(obj_path, out, err);
    err = "unknown loader kind";
    return false;
}

}  // namespace mcpt
