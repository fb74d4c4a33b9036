// TEST INFRASTRUCTURE -- not product code.  Built only into oracle/_ref/libmcpt_refcv.so.
//
// C bridge around the REFERENCE'S OWN OBJ/MTL reader of the CUDA tracer, compiled from where it lies
// under /root/reference (never copied): CVMCTracer/CVMCTracer/Framework/ObjReader.{hpp,cpp} (+ Math.hpp,
// stdafx.h with the CUDA toolkit headers on the include path).  What is restated here is only the
// host-side table logic of PW::Tracer::CreateGeometry (CUDA/CUTracer.cu:259-311), which lives in a .cu
// file that needs OpenCV and a GPU: triangles in file order without the dummy slot 0, one material per
// OBJ group = the material of the group's first triangle, doubles narrowed to float.
// It also exposes Matrix33f::det (Framework/Math.hpp:157-185), the one explicit determinant order the
// reference gives, so the oracle's triangle test can be pinned against it.
// Used by tests/ to pin libmcpt's MCPT_LOADER_OBJREADER convention.
#include <Framework/ObjReader.hpp>

#include <cstdint>
#include <cstring>

namespace {
struct CvScene {
    PW::FileReader::ObjModel model;
    std::vector<int> triMaterial;  // per triangle (1-based triangle index - 1)
};
}  // namespace

extern "C" {

void* refcv_load_obj(const char* path) {
    auto* s = new CvScene;
    try {
        s->model.readObj(path);
    } catch (...) {
        delete s;
        return nullptr;
    }
    const auto& m = s->model;
    s->triMaterial.assign(m.m_triangles.size() > 0 ? m.m_triangles.size() - 1 : 0, 0);
    for (const auto& g : m.m_groups) {  // CUTracer.cu:288-311
        const auto& ids = g.second.m_triangleIndices;
        if (ids.empty()) continue;
        const int mat = m.m_triangles[static_cast<size_t>(ids[0])].materialIndex;
        for (int id : ids) s->triMaterial[static_cast<size_t>(id) - 1] = mat;
    }
    return s;
}

void refcv_free(void* h) { delete static_cast<CvScene*>(h); }

// out = {vertices, normals, triangles (without the dummy), materials}
void refcv_counts(void* h, int* out) {
    const auto& m = static_cast<CvScene*>(h)->model;
    out[0] = static_cast<int>(m.m_vertices.size());
    out[1] = static_cast<int>(m.m_normals.size());
    out[2] = static_cast<int>(m.m_triangles.size()) - 1;
    out[3] = static_cast<int>(m.m_materials.size());
}

void refcv_copy(void* h, float* verts, float* normals, uint32_t* tris, float* mats) {
    auto* s = static_cast<CvScene*>(h);
    const auto& m = s->model;
    for (size_t i = 0; i < m.m_vertices.size(); ++i) { verts[3 * i] = m.m_vertices[i].x; verts[3 * i + 1] = m.m_vertices[i].y; verts[3 * i + 2] = m.m_vertices[i].z; }
    for (size_t i = 0; i < m.m_normals.size(); ++i) { normals[3 * i] = m.m_normals[i].x; normals[3 * i + 1] = m.m_normals[i].y; normals[3 * i + 2] = m.m_normals[i].z; }
    for (size_t i = 1; i < m.m_triangles.size(); ++i) {
        uint32_t* r = tris + 7 * (i - 1);
        for (int k = 0; k < 3; ++k) { r[k] = m.m_triangles[i].m_vertexIndex[k]; r[3 + k] = m.m_triangles[i].m_normalIndex[k]; }
        r[6] = static_cast<uint32_t>(s->triMaterial[i - 1]);
    }
    for (size_t i = 0; i < m.m_materials.size(); ++i) {
        const auto& a = m.m_materials[i];
        const float rec[12] = {(float)a.Ka.x, (float)a.Ka.y, (float)a.Ka.z, (float)a.Kd.x, (float)a.Kd.y, (float)a.Kd.z,
                               (float)a.Ks.x, (float)a.Ks.y, (float)a.Ks.z, (float)a.Ns, (float)a.Tr, (float)a.Ni};
        std::memcpy(mats + 12 * i, rec, sizeof(rec));
    }
}

// Matrix33f::det of the row-major 3x3 in m (Math.hpp:169-175).
float refcv_det33(const float* m) {
    PW::Math::Matrix33f M(m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], m[8]);
    return M.det();
}

}  // extern "C"
