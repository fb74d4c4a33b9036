// TEST INFRASTRUCTURE -- not product code.  Built only into oracle/_ref/libmcpt_ref.so.
//
// Thin C bridge around the REFERENCE'S OWN host code, compiled from where it lies
// under /root/reference (never copied into this repo):
//   * tinyobjloader v1.1.1   MCRT/QuinEngine/3rdparty/include/tiny_obj_loader.h
//   * Quin::Utils::QuinModel MCRT/QuinEngine/Utils/Structure.hpp:7-16
//   * Quin::Utils::KDTree    MCRT/QuinEngine/Utils/KDTree.hpp:58-287
// What is restated here (because the reference wrote it inline in a D3D11 file that
// cannot compile on this box) is ONLY the table/flatten logic of
//   MCRT/QuinEngine/RTX/ShaderResource.hpp:88-104  (triangle table)
//   MCRT/QuinEngine/RTX/ShaderResource.hpp:128-179 (BFS flatten of the pointer tree)
//   MCRT/QuinEngine/RTX/ShaderResource.hpp:204-215 (material table)
// with the 64-slot inline id array replaced by (offset,count) into one id list so a
// leaf deeper than 32 levels cannot overflow.
//
// Used by tests/ to (a) pin oracle/oracle.cpp's restated builder against the real one,
// (b) bake the reference scenes into scenes/*.mcpt fixtures, and by bench.py's
// cpu_baseline / --impl reference legs.  Nothing in the product links against it.
#define TINYOBJLOADER_IMPLEMENTATION
#include <stdafx.h>

#include <Utils/Structure.hpp>
#include <Utils/KDTree.hpp>

#include <chrono>
#include <cstring>
#include <list>

namespace {

struct RefScene {
    tinyobj::attrib_t attr;
    std::vector<tinyobj::shape_t> shapes;
    std::vector<tinyobj::material_t> materials;
    // flattened tree
    std::vector<uint32_t> left, right, splitAxis, leafOffset, leafCount, leafIds;
    std::vector<float> aabb, splitValue;
    uint32_t maxDepth = 0, numLeaves = 0;
    double buildSeconds = 0.0;
};

void flatten(RefScene& s, const Quin::Utils::KDTree& kdtree) {
    // ShaderResource.hpp:128-179: BFS, node k in pop order gets index k, the two
    // children are adjacent, leaves carry ids in std::set (ascending) order.
    using Quin::Utils::KDNode;
    std::list<std::pair<KDNode*, uint32_t>> bfs;
    bfs.push_back({kdtree.m_root, 0u});
    size_t emitted = 0;
    while (!bfs.empty()) {
        auto [node, depth] = bfs.front();
        bfs.pop_front();
        if (depth > s.maxDepth) s.maxDepth = depth;
        for (int i = 0; i < 3; ++i) s.aabb.push_back(node->aabb.min()[i]);
        for (int i = 0; i < 3; ++i) s.aabb.push_back(node->aabb.max()[i]);
        if (node->split.axis != Quin::Utils::KDAxis::AXIS_NONE) {
            s.left.push_back(static_cast<uint32_t>(bfs.size() + emitted + 1));
            s.right.push_back(static_cast<uint32_t>(bfs.size() + emitted + 2));
            s.splitAxis.push_back(static_cast<uint32_t>(node->split.axis));
            s.splitValue.push_back(node->split.value);
            s.leafOffset.push_back(static_cast<uint32_t>(s.leafIds.size()));
            s.leafCount.push_back(0u);
            bfs.push_back({node->left, depth + 1});
            bfs.push_back({node->right, depth + 1});
        } else {
            s.left.push_back(0xFFFFFFFFu);
            s.right.push_back(0xFFFFFFFFu);
            s.splitAxis.push_back(0u);
            s.splitValue.push_back(0.0f);
            s.leafOffset.push_back(static_cast<uint32_t>(s.leafIds.size()));
            s.leafCount.push_back(static_cast<uint32_t>(node->triangleIds.size()));
            for (auto id : node->triangleIds) s.leafIds.push_back(id);
            ++s.numLeaves;
        }
        ++emitted;
    }
}

}  // namespace

extern "C" {

void* ref_scene_load_obj(const char* objPath, const char* mtlBase) {
    auto* s = new RefScene;
    // Structure.hpp:9-12 (QuinModel ctor): same call, same defaults (triangulate = true).
    std::string err;
    bool ok = tinyobj::LoadObj(&s->attr, &s->shapes, &s->materials, &err, objPath, mtlBase);
    if (!ok) { delete s; return nullptr; }
    return s;
}

void* ref_scene_from_arrays(const float* verts, int nV, const uint32_t* triV, int nT) {
    // Geometry-only scene for synthetic meshes: one shape, indices in the given order,
    // which is exactly what KDTree.hpp:60-75 consumes.
    auto* s = new RefScene;
    s->attr.vertices.assign(verts, verts + 3 * static_cast<size_t>(nV));
    tinyobj::shape_t shape;
    shape.mesh.indices.resize(3 * static_cast<size_t>(nT));
    for (size_t i = 0; i < 3 * static_cast<size_t>(nT); ++i) {
        shape.mesh.indices[i].vertex_index = static_cast<int>(triV[i]);
        shape.mesh.indices[i].normal_index = -1;
        shape.mesh.indices[i].texcoord_index = -1;
    }
    shape.mesh.material_ids.assign(static_cast<size_t>(nT), 0);
    s->shapes.push_back(std::move(shape));
    return s;
}

void ref_scene_free(void* h) { delete static_cast<RefScene*>(h); }

// out = {numVertices, numNormals, numTriangles, numMaterials}
void ref_scene_counts(void* h, int* out) {
    auto* s = static_cast<RefScene*>(h);
    size_t nT = 0;
    for (auto& shape : s->shapes) nT += shape.mesh.material_ids.size();
    out[0] = static_cast<int>(s->attr.vertices.size() / 3);
    out[1] = static_cast<int>(s->attr.normals.size() / 3);
    out[2] = static_cast<int>(nT);
    out[3] = static_cast<int>(s->materials.size());
}

// tris: 7 uint32 per triangle {v0,v1,v2,n0,n1,n2,matId}   (ShaderResource.hpp:88-104)
// mats: 12 float per material {Ka3,Kd3,Ks3,Ns,Tr,Ni}       (ShaderResource.hpp:204-215)
void ref_scene_copy(void* h, float* verts, float* normals, uint32_t* tris, float* mats) {
    auto* s = static_cast<RefScene*>(h);
    if (verts) std::memcpy(verts, s->attr.vertices.data(), s->attr.vertices.size() * sizeof(float));
    if (normals) std::memcpy(normals, s->attr.normals.data(), s->attr.normals.size() * sizeof(float));
    if (tris) {
        size_t k = 0;
        for (auto& shape : s->shapes) {
            auto numTries = static_cast<uint32_t>(shape.mesh.material_ids.size());
            for (uint32_t i = 0; i < numTries; ++i) {
                for (int c = 0; c < 3; ++c) tris[7 * k + c] = shape.mesh.indices[3 * i + c].vertex_index;
                for (int c = 0; c < 3; ++c) tris[7 * k + 3 + c] = shape.mesh.indices[3 * i + c].normal_index;
                tris[7 * k + 6] = shape.mesh.material_ids[i];
                ++k;
            }
        }
    }
    if (mats) {
        size_t k = 0;
        for (auto& mat : s->materials) {
            for (int c = 0; c < 3; ++c) mats[12 * k + c] = mat.ambient[c];
            for (int c = 0; c < 3; ++c) mats[12 * k + 3 + c] = mat.diffuse[c];
            for (int c = 0; c < 3; ++c) mats[12 * k + 6 + c] = mat.specular[c];
            mats[12 * k + 9] = mat.shininess;
            mats[12 * k + 10] = 1.0f - mat.dissolve;
            mats[12 * k + 11] = mat.ior;
            ++k;
        }
    }
}

int ref_scene_num_shapes(void* h) { return static_cast<int>(static_cast<RefScene*>(h)->shapes.size()); }
// name copied into buf (NUL-terminated, truncated to cap); returns the shape's triangle count
int ref_scene_shape(void* h, int i, char* buf, int cap) {
    auto* s = static_cast<RefScene*>(h);
    auto& shape = s->shapes[static_cast<size_t>(i)];
    if (buf && cap > 0) {
        std::strncpy(buf, shape.name.c_str(), static_cast<size_t>(cap) - 1);
        buf[cap - 1] = 0;
    }
    return static_cast<int>(shape.mesh.material_ids.size());
}

// Runs the reference KDTree ctor (KDTree.hpp:58-287) and flattens it. Returns node count.
int ref_tree_build(void* h) {
    auto* s = static_cast<RefScene*>(h);
    s->left.clear(); s->right.clear(); s->splitAxis.clear(); s->splitValue.clear();
    s->leafOffset.clear(); s->leafCount.clear(); s->leafIds.clear(); s->aabb.clear();
    s->maxDepth = 0; s->numLeaves = 0;
    auto t0 = std::chrono::steady_clock::now();
    Quin::Utils::KDTree tree(s->attr, s->shapes);
    auto t1 = std::chrono::steady_clock::now();
    s->buildSeconds = std::chrono::duration<double>(t1 - t0).count();
    flatten(*s, tree);
    return static_cast<int>(s->left.size());
}

double ref_tree_build_seconds(void* h) { return static_cast<RefScene*>(h)->buildSeconds; }

// out = {numNodes, numLeaves, numLeafRefs, maxDepth}
void ref_tree_counts(void* h, int* out) {
    auto* s = static_cast<RefScene*>(h);
    out[0] = static_cast<int>(s->left.size());
    out[1] = static_cast<int>(s->numLeaves);
    out[2] = static_cast<int>(s->leafIds.size());
    out[3] = static_cast<int>(s->maxDepth);
}

void ref_tree_copy(void* h, uint32_t* left, uint32_t* right, float* aabb6, uint32_t* splitAxis,
                   float* splitValue, uint32_t* leafOffset, uint32_t* leafCount, uint32_t* leafIds) {
    auto* s = static_cast<RefScene*>(h);
    size_t n = s->left.size();
    std::memcpy(left, s->left.data(), n * 4);
    std::memcpy(right, s->right.data(), n * 4);
    std::memcpy(aabb6, s->aabb.data(), n * 24);
    std::memcpy(splitAxis, s->splitAxis.data(), n * 4);
    std::memcpy(splitValue, s->splitValue.data(), n * 4);
    std::memcpy(leafOffset, s->leafOffset.data(), n * 4);
    std::memcpy(leafCount, s->leafCount.data(), n * 4);
    std::memcpy(leafIds, s->leafIds.data(), s->leafIds.size() * 4);
}

}  // extern "C"
