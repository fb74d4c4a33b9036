// The call sequence of the reference's own main (CVMCTracer/CVMCTracer/main.cpp:9-29) with the OpenCV lines dropped:
// readObj -> CreateGeometry -> RenderScene, then `hostcolor` is written as raw floats instead of result1.png.
// Linked either with tests/dropin/CUTracerMcpt.cpp + libmcpt.so (the drop-in) or, in oracle/_ref/cutracer_ref, with the
// reference's own CUDA/CUTracer.cu compiled for sm_100a (the statistical witness of MCPT_TRACER_CUDA_TWIN).
//   usage: <binary> scene.obj sceneID out.f32
#include <cstdio>
#include <cstdlib>

#include "Framework/Math.hpp"
#include "Framework/ObjReader.hpp"
#include "CUDA/CUTracer.h"

int main(int argc, char** argv) {
    if (argc < 4) { std::fprintf(stderr, "usage: %s scene.obj sceneID out.f32\n", argv[0]); return 2; }
    PW::FileReader::ObjModel model;
    PWVector3f* hostcolor = new PWVector3f[IMG_HEIGHT * IMG_WIDTH];  // main.cpp:12
    try {
        model.readObj(argv[1]);                                     // main.cpp:16
    } catch (...) { std::fprintf(stderr, "readObj failed\n"); return 1; }
    if (PW::Tracer::CreateGeometry(&model) != cudaSuccess) { std::fprintf(stderr, "CreateGeometry failed\n"); return 1; }   // main.cpp:17
    if (PW::Tracer::RenderScene(std::atoi(argv[2]), hostcolor) != cudaSuccess) { std::fprintf(stderr, "RenderScene failed\n"); return 1; }  // main.cpp:18
    FILE* f = std::fopen(argv[3], "wb");
    if (!f) return 1;
    const unsigned dims[2] = {IMG_WIDTH, IMG_HEIGHT};
    std::fwrite(dims, 4, 2, f);
    std::fwrite(hostcolor, sizeof(PWVector3f), IMG_HEIGHT * IMG_WIDTH, f);
    std::fclose(f);
    std::printf("%u x %u, %u spp\n", IMG_WIDTH, IMG_HEIGHT, NUM_KERNELS * NUM_SAMPLES_PER_KERNEL);
    delete[] hostcolor;
    return 0;
}
