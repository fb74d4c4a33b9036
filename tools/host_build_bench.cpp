// Host-side timing of geometry ingest (no device needed): SAH build, packing, ropes for an n x n height field.
//   g++ -O2 -std=c++17 -fopenmp -ffp-contract=off -I/usr/local/cuda/include -Imontecarlopathtracer_b200/csrc \
//       tools/host_build_bench.cpp montecarlopathtracer_b200/csrc/{kd_build,kd_build_sah,device_scene,kd_ropes,scene_io}.cpp -o gpurun_out/host_build_bench
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "device_scene.h"
#include "host_scene.h"

using namespace mcpt;

int main(int argc, char** argv) {
    const int n = argc > 1 ? std::atoi(argv[1]) : 708;
    HostScene s;
    if (argc > 1 && std::strstr(argv[1], ".mcpt")) {  // a scene container instead of the height field
        std::string err;
        if (!load_scene_file(argv[1], s, err)) { std::fprintf(stderr, "%s\n", err.c_str()); return 1; }
    } else
    for (int j = 0; j <= n; ++j)
        for (int i = 0; i <= n; ++i) {
            const float x = -5.0f + 10.0f * i / n, z = -5.0f + 10.0f * j / n;
            s.vertices.push_back(x);
            s.vertices.push_back(1.0f + 0.4f * std::sin(3.1f * x) * std::cos(2.7f * z) + 0.05f * std::sin(37.0f * x + 11.0f * z));
            s.vertices.push_back(z);
        }
    s.normals = {0.0f, 1.0f, 0.0f};
    s.materials = {0, 0, 0, 0.7f, 0.7f, 0.7f, 0, 0, 0, 1.0f, 0.0f, 1.0f};
    for (int j = 0; j < n && s.triangles.size() < 7u * 2u * n * n; ++j)
        for (int i = 0; i < n; ++i) {
            const uint32_t a = j * (n + 1) + i, b = a + 1, c = a + n + 1, d = c + 1;
            const uint32_t t0[7] = {a, c, b, 0, 0, 0, 0}, t1[7] = {b, c, d, 0, 0, 0, 0};
            s.triangles.insert(s.triangles.end(), t0, t0 + 7);
            s.triangles.insert(s.triangles.end(), t1, t1 + 7);
        }
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto secs = [](auto a, auto b) { return std::chrono::duration<double>(b - a).count(); };
    HostTree tree;
    auto t0 = now();
    build_kd_tree(s, argc > 2 ? std::atoi(argv[2]) : 1 /* 1 = MCPT_BUILD_SAH, 0 = reference rule */, tree);
    auto t1 = now();
    PackedScene packed;
    pack_scene(s, tree, packed);
    auto t2 = now();
    uint64_t h = 1469598103934665603ull;  // FNV-1a over the flattened tree: must not change when the builder is only made faster
    auto mix = [&](const void* p, size_t bytes) { const unsigned char* c = static_cast<const unsigned char*>(p); for (size_t i = 0; i < bytes; ++i) { h ^= c[i]; h *= 1099511628211ull; } };
    mix(tree.left.data(), tree.left.size() * 4); mix(tree.split_axis.data(), tree.split_axis.size() * 4);
    mix(tree.split_value.data(), tree.split_value.size() * 4); mix(tree.leaf_ids.data(), tree.leaf_ids.size() * 4);
    mix(tree.leaf_count.data(), tree.leaf_count.size() * 4);
    uint64_t hp = 1469598103934665603ull;
    { uint64_t keep = h; h = hp; mix(packed.leaf_blob.data(), packed.leaf_blob.size() * 16); mix(packed.leaf_tris.data(), packed.leaf_tris.size() * 16); mix(packed.rope_nodes.data(), packed.rope_nodes.size() * 8); hp = h; h = keep; }
    size_t hist[10] = {0};
    for (size_t k = 0; k < tree.num_nodes(); ++k)
        if (tree.split_axis[k] == 0) ++hist[std::min<uint32_t>(tree.leaf_count[k], 9u)];
    std::printf("leaf sizes 0..8, 9+:");
    for (int k = 0; k < 10; ++k) std::printf(" %zu", hist[k]);
    std::printf("\n");
    std::printf("tree digest %016llx, packed digest %016llx\n", static_cast<unsigned long long>(h), static_cast<unsigned long long>(hp));
    std::printf("%zu triangles: build %.3f s (%zu nodes, %u leaves, %zu refs, depth %u), pack + ropes %.3f s\n", s.num_triangles(),
                secs(t0, t1), tree.num_nodes(), tree.num_leaves, tree.leaf_ids.size(), tree.max_depth, secs(t1, t2));
    return 0;
}
