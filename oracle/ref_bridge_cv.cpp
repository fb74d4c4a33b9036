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
//
// Third part: the REFERENCE'S OWN BSDF SAMPLERS, CUDA/Utils.hpp:46-137 (sampleHemi, samplePhong,
// sampleFresnel; the same math as rtx.hlsl:213-302), compiled unmodified for the HOST: without nvcc
// cuda_runtime.h defines __device__ / __inline__ away and curand_kernel.h's XORWOW generator
// (curand_init, curand_uniform) is plain C.  Each call's uniforms are recovered by replaying a copy of
// the curandState, so that the oracle's restatement (shader_restate.cpp) can be fed the very same
// numbers and compared bit for bit (tests/test_oracle_pins.py).
#include <Framework/ObjReader.hpp>
#include <CUDA/Utils.hpp>

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

// ---- CUDA/Utils.hpp:46-137 on the host ----------------------------------------------------------
// n calls on one XORWOW stream (curand_init(seed, 0, 0), as rayTraceKernel seeds it, CUTracer.cu:186-187).
// normals / dirs: 3 floats per call; out: 3 floats per call; uniforms: the 2 (hemi, phong) or 1 (fresnel)
// numbers curand_uniform handed the sampler in that call, in draw order.
void refcv_sample_hemi(unsigned long long seed, int n, const float* normals, float* out, float* uniforms) {
    curandState rng;
    curand_init(seed, 0, 0, &rng);
    for (int i = 0; i < n; ++i) {
        curandState replay = rng;
        uniforms[2 * i] = curand_uniform(&replay);
        uniforms[2 * i + 1] = curand_uniform(&replay);
        const PWVector3f nrm(normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]);
        const PWVector3f o = PW::CUDA::sampleHemi(&rng, nrm);
        out[3 * i] = o.x; out[3 * i + 1] = o.y; out[3 * i + 2] = o.z;
    }
}

void refcv_sample_phong(unsigned long long seed, int n, const float* normals, const float* dirs, const unsigned* Ns,
                        float* out, float* uniforms) {
    curandState rng;
    curand_init(seed, 0, 0, &rng);
    for (int i = 0; i < n; ++i) {
        curandState replay = rng;
        uniforms[2 * i] = curand_uniform(&replay);
        uniforms[2 * i + 1] = curand_uniform(&replay);
        const PWVector3f nrm(normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]);
        const PWVector3f d(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
        const PWVector3f o = PW::CUDA::samplePhong(&rng, nrm, d, Ns[i]);
        out[3 * i] = o.x; out[3 * i + 1] = o.y; out[3 * i + 2] = o.z;
    }
}

void refcv_sample_fresnel(unsigned long long seed, int n, const float* normals, const float* dirs, const float* Tr,
                          const float* Ni, float* out, float* uniforms) {
    curandState rng;
    curand_init(seed, 0, 0, &rng);
    for (int i = 0; i < n; ++i) {
        curandState replay = rng;
        uniforms[i] = curand_uniform(&replay);
        const PWVector3f nrm(normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]);
        const PWVector3f d(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
        const PWVector3f o = PW::CUDA::sampleFresnel(&rng, nrm, d, Tr[i], Ni[i]);
        out[3 * i] = o.x; out[3 * i + 1] = o.y; out[3 * i + 2] = o.z;
    }
}

// CUDA/Utils.hpp:27-34 normalize, :12-15 dot (the twin's path loop uses both, CUTracer.cu:126,152).
void refcv_normalize(float v[3]) {
    PWVector3f a(v[0], v[1], v[2]);
    PW::CUDA::normalize(a);
    v[0] = a.x; v[1] = a.y; v[2] = a.z;
}

}  // extern "C"
