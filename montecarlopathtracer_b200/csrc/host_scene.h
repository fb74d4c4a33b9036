// Host-side scene tables and the flattened KD-tree, as plain vectors.
// Field meaning follows the reference's buffer contract CSTriangle / CSMaterial / CSKDTree
// (MCRT/QuinEngine/Utils/Structure.hpp:196-223).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace mcpt {

struct HostScene {
    std::vector<float> vertices;     // 3 per vertex
    std::vector<float> normals;      // 3 per normal
    std::vector<uint32_t> triangles; // 7 per triangle: v0 v1 v2 n0 n1 n2 matId
    std::vector<float> materials;    // 12 per material: Ka Kd Ks Ns Tr Ni
    size_t num_vertices() const { return vertices.size() / 3; }
    size_t num_normals() const { return normals.size() / 3; }
    size_t num_triangles() const { return triangles.size() / 7; }
    size_t num_materials() const { return materials.size() / 12; }
    const float* vertex(uint32_t tri, int k) const { return &vertices[3 * static_cast<size_t>(triangles[7 * static_cast<size_t>(tri) + k])]; }
};

// Flattened tree, node k = k-th node in breadth-first order, children adjacent
// (right = left + 1), as the reference uploads it (ShaderResource.hpp:128-179).
struct HostTree {
    std::vector<uint32_t> left;        // 0xFFFFFFFF for a leaf
    std::vector<float> aabb;           // 6 per node: min xyz, max xyz (clipped node box)
    std::vector<uint32_t> split_axis;  // 0 = leaf, 1..3 = x,y,z
    std::vector<float> split_value;
    std::vector<uint32_t> leaf_offset; // into leaf_ids
    std::vector<uint32_t> leaf_count;
    std::vector<uint32_t> leaf_ids;    // ascending inside a leaf
    uint32_t max_depth = 0;
    uint32_t num_leaves = 0;
    double build_seconds = 0.0;
    // A tree built AND packed on the device stays there until something on the host asks for it (context.h: ensure_host_tree):
    // until then only the counts and the root box below are valid and the vectors above are empty.
    bool on_host = true;
    size_t nodes = 0, refs = 0;
    float root_box[6] = {0, 0, 0, 0, 0, 0};
    size_t num_nodes() const { return on_host ? left.size() : nodes; }
    size_t num_refs() const { return on_host ? leaf_ids.size() : refs; }
    float root(int i) const { return on_host ? aabb[static_cast<size_t>(i)] : root_box[i]; }
};

// Tuning and A/B knobs.  The library never reads the caller's environment: the only way to change them is
// mcpt_set_tuning() (values clamped), which the test and tuning scripts call explicitly.  Defaults = shipped
// configuration (profiles/r1_tuning.md, profiles/r2_tuning.md).
struct Tuning {
    int shade_threshold = 18;  // lanes that must have finished rays before a warp leaves TRAVERSE     [1, 32]
    int cam_threshold = 16;    // lanes that must need a camera ray before the warp refills its queue  [1, 32]
    int tri_cap = 4;           // triangle tests per lane and round                                    [1, 64]
    int chunks = 0;            // sample slices per tile, 0 = pick_chunks' choice                      [0, 64]
    int pool = 0;              // 1 = render_kernel_pool (pooled lanes, render_pool.cuh) where it applies; 0 = render_kernel_sm (shipped)
    int min_blocks = 0;        // resident 256-thread blocks per SM: 0 = automatic (4; 8 for scenes that outgrow L2), 4 / 5 / 6 / 8 force it
    int variant = 503;         // MCPT_TUNE builds only: VAR bits of render_kernel_sm
    int leaf_bfs = 0;          // A/B: breadth-first instead of depth-first device order of the leaves
    int sah_resort = 0;        // A/B: the re-sorting SAH builder on every level
    int build_timing = 0;      // print builder phase timings to stderr
    int sm_stats = 0;          // MCPT_TUNE builds, VAR bit 3: print per-warp round statistics
    int device_pack = 1;       // MCPT_BUILD_GPU: render-path tables made on the device as well (0: downloaded tree packed on the host, uploaded)
    int verify_layout = 0;     // run check_ropes on every geometry install (the render kernel's walk relies on consistent tables)
    float sah_ci = 1.0f;       // SAH: cost of a triangle test relative to a node step              [0.05, 100]
    float sah_eb = 1.0f;       // SAH: multiplier when one child is empty                            [0.05, 100]
};
Tuning& tuning();

// kd_build.cpp
// mode 0: node-for-node the tree of KDTree.hpp:58-287; mode 1: own SAH builder.
void build_kd_tree(const HostScene& scene, int mode, HostTree& out);
// kd_build_gpu.cu: level-synchronous SAH build on the bound CUDA device (MCPT_BUILD_GPU), result in the same host layout
// With `pack` the render-path tables (leaf blocks + neighbour links, location nodes, shading records) are made on the device too,
// straight into the caller's buffers (context.h: DeviceTables): nothing of them is computed on the host or uploaded.
struct DeviceTables;
bool build_gpu_tree(const HostScene& scene, HostTree& out, std::string& err, DeviceTables* pack = nullptr);
void release_gpu_builder_scratch();   // the builder's device scratch block is kept between builds; mcpt_shutdown frees it

// scene_io.cpp
bool load_scene_file(const std::string& path, HostScene& out, std::string& err);
bool save_scene_file(const std::string& path, const HostScene& scene, std::string& err);
bool validate_scene(const HostScene& scene, std::string& err);

}  // namespace mcpt
