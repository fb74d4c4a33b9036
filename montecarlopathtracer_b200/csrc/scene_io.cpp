// Baked scene container (*.mcpt): the flat tables mcpt_create_geometry takes, little-endian.
//   char[8]  "MCPTSCN1"
//   uint32   numVertices, numNormals, numTriangles, numMaterials
//   float    vertices[3*V]; float normals[3*N]; uint32 triangles[7*T]; float materials[12*M]
// It exists so that million-triangle scenes load without re-parsing OBJ text
// and so the repo's benchmark scenes travel as data.
#include <cstdio>
#include <cstring>

#include "host_scene.h"

namespace mcpt {

namespace {
const char kMagic[8] = {'M', 'C', 'P', 'T', 'S', 'C', 'N', '1'};

template <class T>
bool read_vec(FILE* f, std::vector<T>& v, size_t n) {
    v.resize(n);
    return n == 0 || fread(v.data(), sizeof(T), n, f) == n;
}
}  // namespace

bool validate_scene(const HostScene& s, std::string& err) {
    if (s.vertices.size() % 3 || s.normals.size() % 3 || s.triangles.size() % 7 || s.materials.size() % 12) {
        err = "scene tables have the wrong stride";
        return false;
    }
    if (s.num_triangles() == 0) { err = "scene has no triangles"; return false; }
    const uint32_t nV = static_cast<uint32_t>(s.num_vertices()), nN = static_cast<uint32_t>(s.num_normals()),
                   nM = static_cast<uint32_t>(s.num_materials());
    // first offending triangle (lowest index), found in parallel for large scenes
    const long long nT = static_cast<long long>(s.num_triangles());
    long long firstBad = nT;
#pragma omp parallel for schedule(static) reduction(min : firstBad) if (nT > 65536)
    for (long long t = 0; t < nT; ++t) {
        const uint32_t* r = &s.triangles[7 * static_cast<size_t>(t)];
        if (r[0] >= nV || r[1] >= nV || r[2] >= nV || r[3] >= nN || r[4] >= nN || r[5] >= nN || r[6] >= nM) firstBad = t < firstBad ? t : firstBad;
    }
    if (firstBad < nT) {
        const size_t t = static_cast<size_t>(firstBad);
        const uint32_t* r = &s.triangles[7 * t];
        if (r[0] >= nV || r[1] >= nV || r[2] >= nV) { err = "triangle " + std::to_string(t) + ": vertex index out of range"; return false; }
        if (r[3] >= nN || r[4] >= nN || r[5] >= nN) { err = "triangle " + std::to_string(t) + ": normal index out of range"; return false; }
        err = "triangle " + std::to_string(t) + ": material index out of range";
        return false;
    }
    return true;
}

bool load_scene_file(const std::string& path, HostScene& out, std::string& err) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) { err = "cannot open " + path; return false; }
    char magic[8];
    uint32_t n[4];
    bool ok = fread(magic, 1, 8, f) == 8 && memcmp(magic, kMagic, 8) == 0 && fread(n, 4, 4, f) == 4;
    if (ok) {
        // The counts come from an untrusted file: they must fit the ABI's int32 counts and add up to exactly the
        // bytes that follow the header, BEFORE anything is sized from them.
        long here = ftell(f);
        ok = here >= 0 && fseek(f, 0, SEEK_END) == 0;
        const long end = ok ? ftell(f) : -1;
        ok = ok && end >= here && fseek(f, here, SEEK_SET) == 0;
        const unsigned long long need = 4ull * (3ull * n[0] + 3ull * n[1] + 7ull * n[2] + 12ull * n[3]);
        if (ok && (n[0] > 0x7FFFFFFFu || n[1] > 0x7FFFFFFFu || n[2] > 0x7FFFFFFFu || n[3] > 0x7FFFFFFFu ||
                   need != static_cast<unsigned long long>(end - here))) {
            fclose(f);
            err = path + ": header counts do not match the file size";
            return false;
        }
    }
    ok = ok && read_vec(f, out.vertices, 3 * static_cast<size_t>(n[0])) && read_vec(f, out.normals, 3 * static_cast<size_t>(n[1])) &&
         read_vec(f, out.triangles, 7 * static_cast<size_t>(n[2])) && read_vec(f, out.materials, 12 * static_cast<size_t>(n[3]));
    fclose(f);
    if (!ok) { err = path + ": not a complete MCPTSCN1 file"; return false; }
    return validate_scene(out, err);
}

bool save_scene_file(const std::string& path, const HostScene& s, std::string& err) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) { err = "cannot create " + path; return false; }
    uint32_t n[4] = {static_cast<uint32_t>(s.num_vertices()), static_cast<uint32_t>(s.num_normals()),
                     static_cast<uint32_t>(s.num_triangles()), static_cast<uint32_t>(s.num_materials())};
    bool ok = fwrite(kMagic, 1, 8, f) == 8 && fwrite(n, 4, 4, f) == 4;
    ok = ok && fwrite(s.vertices.data(), 4, s.vertices.size(), f) == s.vertices.size();
    ok = ok && fwrite(s.normals.data(), 4, s.normals.size(), f) == s.normals.size();
    ok = ok && fwrite(s.triangles.data(), 4, s.triangles.size(), f) == s.triangles.size();
    ok = ok && fwrite(s.materials.data(), 4, s.materials.size(), f) == s.materials.size();
    fclose(f);
    if (!ok) err = "short write to " + path;
    return ok;
}

}  // namespace mcpt
