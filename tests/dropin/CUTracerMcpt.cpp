// The drop-in shim a maintainer of pw1316/MonteCarloPathTracer adds INSTEAD of CUDA/CUTracer.cu: the four functions
// of PW::Tracer (CVMCTracer/CVMCTracer/CUDA/CUTracer.h:9-12), same signatures, implemented over libmcpt.so's C ABI.
// This file is compiled against the reference's REAL CUTracer.h / ObjReader.hpp (never copied: -I into
// /root/reference) by `make -C oracle dropin`, linked with libmcpt.so and run on the GPU box by
// tests/test_dropin.py -- INTEGRATION.md section 1 includes it verbatim.
#include "CUDA/CUTracer.h"   // the reference's own header

#include <cstdint>
#include <vector>

#include "mcpt.h"

namespace PW {
namespace Tracer {

static cudaError_t rc(int e) { return e == MCPT_OK ? cudaSuccess : cudaErrorUnknown; }

cudaError_t Initialize() { return rc(mcpt_initialize(0)); }  // CUTracer.cu:220-223

// CUTracer.cu:225-314: vertex / normal / triangle tables as the reader filled them (1-based indices, dummy slot 0
// kept so that no index moves), ONE material per OBJ group = the material of the group's first triangle (:300-308).
cudaError_t CreateGeometry(const PW::FileReader::ObjModel* m) {
    std::vector<float> v, n, mat;
    std::vector<uint32_t> tri;
    for (const auto& p : m->m_vertices) { v.push_back(static_cast<float>(p.x)); v.push_back(static_cast<float>(p.y)); v.push_back(static_cast<float>(p.z)); }
    for (const auto& p : m->m_normals) { n.push_back(static_cast<float>(p.x)); n.push_back(static_cast<float>(p.y)); n.push_back(static_cast<float>(p.z)); }
    std::vector<int> groupMaterial(m->m_triangles.size(), 0);
    for (const auto& g : m->m_groups) {
        const auto& ids = g.second.m_triangleIndices;
        if (ids.empty()) continue;
        const int first = m->m_triangles[static_cast<size_t>(ids[0])].materialIndex;
        for (int id : ids) groupMaterial[static_cast<size_t>(id)] = first;
    }
    for (size_t i = 1; i < m->m_triangles.size(); ++i) {  // slot 0 is the reader's dummy triangle
        const auto& t = m->m_triangles[i];
        for (int k = 0; k < 3; ++k) tri.push_back(static_cast<uint32_t>(t.m_vertexIndex[k]));
        for (int k = 0; k < 3; ++k) tri.push_back(static_cast<uint32_t>(t.m_normalIndex[k]));
        tri.push_back(static_cast<uint32_t>(groupMaterial[i]));
    }
    for (const auto& a : m->m_materials) {  // Geometry::Material (Framework/Geometry.h:14-22): doubles narrowed to float (:303-308)
        const float r[12] = {(float)a.Ka.x, (float)a.Ka.y, (float)a.Ka.z, (float)a.Kd.x, (float)a.Kd.y, (float)a.Kd.z,
                             (float)a.Ks.x, (float)a.Ks.y, (float)a.Ks.z, (float)a.Ns, (float)a.Tr, (float)a.Ni};
        mat.insert(mat.end(), r, r + 12);
    }
    mcpt_scene_desc d = {v.data(), static_cast<int32_t>(v.size() / 3), n.data(), static_cast<int32_t>(n.size() / 3), tri.data(),
                         static_cast<int32_t>(tri.size() / 7), mat.data(), static_cast<int32_t>(mat.size() / 12), MCPT_BUILD_SAH};
    if (int e = mcpt_initialize(0)) return rc(e);  // the reference's main never calls Initialize() (main.cpp:16-18)
    return rc(mcpt_create_geometry(&d));
}

cudaError_t DestroyGeometry() { return rc(mcpt_destroy_geometry()); }  // CUTracer.cu:316-338

// CUTracer.cu:340-404.  The compile-time knobs of stdafx.h:41-45 become the cfg: CUDA_TWIN mode = +-1 px jitter,
// hfov 60 deg, 7 scattering events + one gather, ILLUM = 10, eye (0,5,17) for scene 1 and (0,5,23) otherwise.
cudaError_t RenderScene(const PWint sceneID, PWVector3f* hostcolor) {
    mcpt_render_cfg cfg;
    mcpt_cuda_twin_render_cfg(&cfg, IMG_WIDTH, IMG_HEIGHT, NUM_KERNELS * NUM_SAMPLES_PER_KERNEL, sceneID);
    return rc(mcpt_render_scene(&cfg, &hostcolor->x, nullptr));  // W*H*3 floats, row-major, top row first, averaged
}

}  // namespace Tracer
}  // namespace PW
