// C ABI, part 3: spp-split rendering over several GPUs behind the drop-in boundary.
//
// The reference is single-GPU (CUTracer.cu:222 is its only device call).  Its only cross-sample coupling is the
// running mean of rtx.hlsl:401-404 -- a sum -- so the samples of a pixel can be split over ranks exactly: one process
// per GPU, rank r renders the r-th contiguous share of the job's sample range (Philox counters carry the ABSOLUTE sample
// index: the union over ranks IS the 1-GPU sample set), and ONE ncclReduce(float4 sums, root 0) over NVLink / NVSwitch
// on the render stream combines the accumulation buffers; the root divides by the sample count on the device and copies
// one float3 image to the host.  16 MiB (1024^2) to 256 MiB (4096^2) per rank against 0.5-60 s of rendering: a plain
// collective on the render stream, nothing to overlap.
//
// NCCL is bound at run time (dlopen "libnccl.so.2"; in a process that already loaded an NCCL of that SONAME, e.g.
// through torch, the loader hands back that copy): a single-GPU caller of libmcpt.so needs no NCCL at all.  Only the
// few entry points used here are declared, with their public prototypes (nccl.h, NCCL 2.x ABI).
#include <dlfcn.h>

#include <cstring>

#include "context.h"
#include "launch.h"

using namespace mcpt;

namespace mcpt {
// abi_render.cpp
int render_core(const mcpt_render_cfg* cfg, cudaStream_t stream, float4* dst, float4* dstSq, const float4** sum, const float4** sumsq);
int finish_counters(cudaStream_t stream);
}  // namespace mcpt

namespace {

// ---- the slice of nccl.h this file needs (types and prototypes as published; NCCL_UNIQUE_ID_BYTES = 128) ----
struct ncclComm;
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;   // ncclSuccess = 0
const int kNcclFloat32 = 7; // ncclFloat32 / ncclFloat
const int kNcclSum = 0;     // ncclSum

struct Nccl {
    void* so = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclComm_t comm = nullptr;
    int rank = 0, nranks = 1, version = 0;
};

Nccl& nccl() {
    static Nccl n;
    return n;
}

int fail_nccl(ncclResult_t r, const char* what) {
    Nccl& n = nccl();
    return fail(MCPT_ERR_NCCL, std::string(what) + ": " + (n.GetErrorString ? n.GetErrorString(r) : "NCCL error") + " (" + std::to_string(r) + ")");
}

int load_nccl() {
    Nccl& n = nccl();
    if (n.so) return MCPT_OK;
    void* so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!so) so = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!so) return fail(MCPT_ERR_NCCL, std::string("cannot load libnccl.so.2: ") + dlerror());
    auto sym = [&](const char* name) { return dlsym(so, name); };
    n.GetVersion = reinterpret_cast<decltype(n.GetVersion)>(sym("ncclGetVersion"));
    n.GetUniqueId = reinterpret_cast<decltype(n.GetUniqueId)>(sym("ncclGetUniqueId"));
    n.CommInitRank = reinterpret_cast<decltype(n.CommInitRank)>(sym("ncclCommInitRank"));
    n.CommDestroy = reinterpret_cast<decltype(n.CommDestroy)>(sym("ncclCommDestroy"));
    n.Reduce = reinterpret_cast<decltype(n.Reduce)>(sym("ncclReduce"));
    n.GetErrorString = reinterpret_cast<decltype(n.GetErrorString)>(sym("ncclGetErrorString"));
    if (!n.GetUniqueId || !n.CommInitRank || !n.CommDestroy || !n.Reduce) {
        dlclose(so);
        return fail(MCPT_ERR_NCCL, "libnccl.so.2 lacks ncclGetUniqueId / ncclCommInitRank / ncclCommDestroy / ncclReduce");
    }
    if (n.GetVersion) n.GetVersion(&n.version);
    n.so = so;
    return MCPT_OK;
}

}  // namespace

extern "C" {

void mcpt_rank_share(int32_t spp, int32_t sample_offset, int rank, int nranks, int32_t* share_offset, int32_t* share_spp) {
    // contiguous shares whose sizes differ by at most one, the first (spp % nranks) ranks take the extra sample
    if (nranks <= 0 || rank < 0 || rank >= nranks || spp < 0) {
        if (share_offset) *share_offset = sample_offset;
        if (share_spp) *share_spp = 0;
        return;
    }
    const int32_t base = spp / nranks, rem = spp % nranks;
    if (share_offset) *share_offset = sample_offset + rank * base + (rank < rem ? rank : rem);
    if (share_spp) *share_spp = base + (rank < rem ? 1 : 0);
}

int mcpt_comm_get_unique_id(void* id128) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!id128) return fail(MCPT_ERR_INVALID, "null id buffer");
    if (const int rc = load_nccl()) return rc;
    ncclUniqueId id;
    const ncclResult_t r = nccl().GetUniqueId(&id);
    if (r != 0) return fail_nccl(r, "ncclGetUniqueId");
    std::memcpy(id128, &id, sizeof(id));
    return MCPT_OK;
    });
}

int mcpt_comm_init(int rank, int nranks, const void* id128) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail(MCPT_ERR_INVALID, "mcpt_comm_init: rank / nranks out of range");
    if (nranks > 1 && !id128) return fail(MCPT_ERR_INVALID, "mcpt_comm_init: null unique id");
    if (const int rc = bind_device()) return rc;   // one process per GPU: the communicator lives on the bound device
    Nccl& n = nccl();
    if (n.comm) return fail(MCPT_ERR_INVALID, "communicator already initialised; call mcpt_comm_destroy first");
    n.rank = rank; n.nranks = nranks;
    if (nranks == 1) return MCPT_OK;               // a 1-rank job needs no NCCL
    if (const int rc = load_nccl()) return rc;
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    const ncclResult_t r = n.CommInitRank(&n.comm, nranks, id, rank);
    if (r != 0) { n.comm = nullptr; n.nranks = 1; n.rank = 0; return fail_nccl(r, "ncclCommInitRank"); }
    return MCPT_OK;
    });
}

int mcpt_comm_destroy(void) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    Nccl& n = nccl();
    if (n.comm) {
        if (c.has_device) cudaSetDevice(c.device);
        const ncclResult_t r = n.CommDestroy(n.comm);
        n.comm = nullptr;
        if (r != 0) return fail_nccl(r, "ncclCommDestroy");
    }
    n.rank = 0; n.nranks = 1;
    return MCPT_OK;
    });
}

int mcpt_comm_info(int* rank, int* nranks, int* nccl_version) {
    const Nccl& n = nccl();
    if (rank) *rank = n.rank;
    if (nranks) *nranks = n.nranks;
    if (nccl_version) *nccl_version = n.version;
    return MCPT_OK;
}

// cfg describes the WHOLE job; this rank renders its share (mcpt_rank_share) and the sums meet on rank 0.
int mcpt_render_accumulate_distributed_device(const mcpt_render_cfg* cfg, void* dev_sum, void* stream) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!cfg || !dev_sum) return fail(MCPT_ERR_INVALID, "null argument");
    Nccl& n = nccl();
    if (n.nranks > 1 && !n.comm) return fail(MCPT_ERR_INVALID, "mcpt_comm_init has not been called");
    if (const int rc = bind_device()) return rc;
    cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : c.stream;
    mcpt_render_cfg mine = *cfg;
    mcpt_rank_share(cfg->spp, cfg->sample_offset, n.rank, n.nranks, &mine.sample_offset, &mine.spp);
    const size_t pixels = static_cast<size_t>(cfg->width) * static_cast<size_t>(cfg->height);
    c.counters = mcpt_counters{};
    if (mine.spp > 0) {
        if (const int rc = render_core(&mine, s, static_cast<float4*>(dev_sum), nullptr, nullptr, nullptr)) return rc;
    }
    if (n.nranks > 1) {
        const ncclResult_t r = n.Reduce(dev_sum, dev_sum, pixels * 4, kNcclFloat32, kNcclSum, 0, n.comm, s);
        if (r != 0) return fail_nccl(r, "ncclReduce");
    }
    if (mine.spp > 0) return finish_counters(s);
    const cudaError_t e = cudaStreamSynchronize(s);
    return e == cudaSuccess ? MCPT_OK : fail_cuda(e, "reduce");
    });
}

int mcpt_render_scene_distributed(const mcpt_render_cfg* cfg, float* host_rgb) {
    return guarded([&]() -> int {
    Context& c = ctx();
    Nccl& n = nccl();
    size_t pixels = 0;
    {
        std::lock_guard<std::mutex> g(c.lock);
        if (!cfg) return fail(MCPT_ERR_INVALID, "null render cfg");
        if (n.rank == 0 && !host_rgb) return fail(MCPT_ERR_INVALID, "rank 0 needs host_rgb");
        if (cfg->width <= 0 || cfg->height <= 0 || cfg->width > 32768 || cfg->height > 32768) return fail(MCPT_ERR_INVALID, "bad image size");
        if (const int rc = bind_device()) return rc;
        pixels = static_cast<size_t>(cfg->width) * static_cast<size_t>(cfg->height);
        cudaError_t e;
        if ((e = c.d_sum.ensure(pixels)) != cudaSuccess) return fail_cuda(e, "alloc sum");
        if ((e = cudaMemsetAsync(c.d_sum.ptr, 0, pixels * sizeof(float4), c.stream)) != cudaSuccess) return fail_cuda(e, "memset");
    }
    if (const int rc = mcpt_render_accumulate_distributed_device(cfg, c.d_sum.ptr, nullptr)) return rc;
    if (n.rank != 0) return MCPT_OK;
    std::lock_guard<std::mutex> g(c.lock);
    cudaError_t e;
    if ((e = c.d_rgb.ensure(3 * pixels)) != cudaSuccess) return fail_cuda(e, "alloc rgb");
    if ((e = launch_finalize(c.d_sum.ptr, pixels, c.d_rgb.ptr, c.stream, &c.counters.launches)) != cudaSuccess) return fail_cuda(e, "finalize");
    if ((e = cudaMemcpyAsync(host_rgb, c.d_rgb.ptr, 3 * pixels * sizeof(float), cudaMemcpyDeviceToHost, c.stream)) != cudaSuccess) return fail_cuda(e, "D2H");
    if ((e = cudaStreamSynchronize(c.stream)) != cudaSuccess) return fail_cuda(e, "distributed render");
    return MCPT_OK;
    });
}

}  // extern "C"
