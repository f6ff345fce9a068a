// One process per GPU: the per-frame sum of the accumulation buffers over NCCL (NVLink 5 / NVSwitch), from the C-ABI.
//
// The reference is single-device (zero collective call sites); north_star adds "the float accumulation
// buffer is summed with one NCCL allreduce over NVLink per frame" (SURVEY.md §8b lists flx_allreduce_pixels, §8e the
// sharding).  Round 1 reached NCCL only through torch.distributed in the Python mirror; this file gives the C++ host the
// same collective without Python: flx_comm_unique_id / flx_comm_init build an ncclComm_t for the context's device, and
// flx_allreduce_pixels / flx_reduce_pixels enqueue ONE collective over float4[w*h] on the context's stream (plus the two
// denoiser AOV sums when that option is on, grouped into the same launch), ordered after the render like any stage.
//
// NCCL is not a link-time dependency: the library is dlopen'ed at the first flx_comm_* call — the copy already loaded in
// the process if there is one (under Python that is the libnccl torch ships, so a process never holds two NCCLs), else
// the system's libnccl.so.2.  Prototypes are declared by hand from NCCL's stable C API (nccl.h 2.x).
#include "flx_internal.h"
#include <dlfcn.h>
#include <mutex>

namespace {

typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId; // NCCL_UNIQUE_ID_BYTES
enum { kNcclSuccess = 0, kNcclSum = 0, kNcclFloat32 = 7 }; // ncclResult_t, ncclRedOp_t, ncclDataType_t values of nccl.h

struct NcclApi
{
    void *lib = nullptr;
    int (*GetUniqueId)(ncclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Reduce)(const void *, void *, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int *) = nullptr;
    std::string error;
};

NcclApi g_nccl;
std::once_flag g_ncclOnce;

void loadNccl()
{
    // FLX_NCCL_LIB names a file explicitly; otherwise the NCCL this process already holds, else the system's.  (A process
    // that will load a library with its OWN newer libnccl.so.2 later — torch — must load that one first: the dynamic
    // loader hands out whichever copy of a soname came first.  fluctus_b200/context.py does so.)
    void *lib = nullptr;
    if (const char *path = getenv("FLX_NCCL_LIB"))
    {
        lib = dlopen(path, RTLD_NOW | RTLD_GLOBAL);
        if (!lib) { g_nccl.error = std::string("FLX_NCCL_LIB: ") + dlerror(); return; }
    }
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) { g_nccl.error = std::string("NCCL not found (dlopen libnccl.so.2; set FLX_NCCL_LIB to its path): ") + dlerror(); return; }
    g_nccl.lib = lib;
    struct { const char *name; void **fn; } syms[] = {
        {"ncclGetUniqueId", (void **)&g_nccl.GetUniqueId}, {"ncclCommInitRank", (void **)&g_nccl.CommInitRank},
        {"ncclCommDestroy", (void **)&g_nccl.CommDestroy}, {"ncclAllReduce", (void **)&g_nccl.AllReduce},
        {"ncclReduce", (void **)&g_nccl.Reduce}, {"ncclGroupStart", (void **)&g_nccl.GroupStart},
        {"ncclGroupEnd", (void **)&g_nccl.GroupEnd}, {"ncclGetErrorString", (void **)&g_nccl.GetErrorString},
        {"ncclGetVersion", (void **)&g_nccl.GetVersion}};
    for (auto &s : syms)
    {
        *s.fn = dlsym(lib, s.name);
        if (!*s.fn) { g_nccl.error = std::string("NCCL symbol missing: ") + s.name; g_nccl.lib = nullptr; return; }
    }
}

int ncclReady(flx_ctx *ctx)
{
    std::call_once(g_ncclOnce, loadNccl);
    if (!g_nccl.lib) return ctx ? flx_fail(ctx, FLX_ERR_STATE, g_nccl.error) : FLX_ERR_STATE;
    return FLX_OK;
}

#define FLX_NCCL(ctx, call)                                                                                     \
    do {                                                                                                        \
        const int r__ = (call);                                                                                 \
        if (r__ != kNcclSuccess)                                                                                \
            return flx_fail(ctx, FLX_ERR_CUDA, std::string(#call) + ": " + g_nccl.GetErrorString(r__));         \
    } while (0)

// the buffers a frame combine moves: the accumulation buffer, and the denoiser AOV sums when they are being accumulated
int frameBuffers(flx_ctx *ctx, float *bufs[3], size_t &count)
{
    if (!ctx->fb.pixels || ctx->fb.numPixels == 0) return flx_fail(ctx, FLX_ERR_STATE, "no frame buffers yet (flx_update_params / flx_resize first)");
    count = (size_t)ctx->fb.numPixels * 4;
    int n = 0;
    bufs[n++] = (float *)ctx->fb.pixels;
    if (ctx->denoiserAOV) { bufs[n++] = (float *)ctx->fb.denoiserAlbedo; bufs[n++] = (float *)ctx->fb.denoiserNormal; }
    return n;
}

} // namespace

extern "C" {

int flx_comm_unique_id(void *id128)
{
    if (!id128) return FLX_ERR_ARG;
    if (ncclReady(nullptr) != FLX_OK) return FLX_ERR_STATE;
    ncclUniqueId id;
    if (g_nccl.GetUniqueId(&id) != kNcclSuccess) return FLX_ERR_CUDA;
    memcpy(id128, &id, sizeof(id));
    return FLX_OK;
}

int flx_comm_init(flx_ctx *ctx, int rank, int world, const void *id128)
{
    if (!ctx) return FLX_ERR_ARG;
    FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!id128 || world < 1 || rank < 0 || rank >= world) return flx_fail(ctx, FLX_ERR_ARG, "flx_comm_init: bad rank / world / id");
    if (ctx->comm) return flx_fail(ctx, FLX_ERR_STATE, "flx_comm_init: this context already has a communicator (flx_comm_destroy first)");
    { const int rc = ncclReady(ctx); if (rc != FLX_OK) return rc; }
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    ncclComm_t comm = nullptr;
    FLX_NCCL(ctx, g_nccl.CommInitRank(&comm, world, id, rank)); // blocks until all `world` ranks have joined
    ctx->comm = comm; ctx->commRank = rank; ctx->commWorld = world;
    return FLX_OK;
}

int flx_comm_destroy(flx_ctx *ctx)
{
    if (!ctx) return FLX_ERR_ARG;
    if (!ctx->comm) return FLX_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    const int r = g_nccl.CommDestroy((ncclComm_t)ctx->comm);
    ctx->comm = nullptr; ctx->commRank = 0; ctx->commWorld = 1;
    return r == kNcclSuccess ? FLX_OK : flx_fail(ctx, FLX_ERR_CUDA, std::string("ncclCommDestroy: ") + g_nccl.GetErrorString(r));
}

int flx_comm_info(flx_ctx *ctx, int *rank, int *world, int *nccl_version)
{
    if (!ctx) return FLX_ERR_ARG;
    if (rank) *rank = ctx->commRank;
    if (world) *world = ctx->comm ? ctx->commWorld : 0;
    if (nccl_version) { *nccl_version = 0; if (g_nccl.lib) g_nccl.GetVersion(nccl_version); }
    return FLX_OK;
}

int flx_allreduce_pixels(flx_ctx *ctx)
{
    if (!ctx) return FLX_ERR_ARG;
    FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!ctx->comm) return flx_fail(ctx, FLX_ERR_STATE, "flx_allreduce_pixels: no communicator (flx_comm_init first)");
    float *bufs[3]; size_t count = 0;
    const int n = frameBuffers(ctx, bufs, count);
    if (n < 0) return n;
    FLX_NCCL(ctx, g_nccl.GroupStart());
    for (int k = 0; k < n; k++)
        FLX_NCCL(ctx, g_nccl.AllReduce(bufs[k], bufs[k], count, kNcclFloat32, kNcclSum, (ncclComm_t)ctx->comm, ctx->stream));
    FLX_NCCL(ctx, g_nccl.GroupEnd());
    return FLX_OK;
}

int flx_reduce_pixels(flx_ctx *ctx, int root)
{
    if (!ctx) return FLX_ERR_ARG;
    FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!ctx->comm) return flx_fail(ctx, FLX_ERR_STATE, "flx_reduce_pixels: no communicator (flx_comm_init first)");
    if (root < 0 || root >= ctx->commWorld) return flx_fail(ctx, FLX_ERR_ARG, "flx_reduce_pixels: root outside the communicator");
    float *bufs[3]; size_t count = 0;
    const int n = frameBuffers(ctx, bufs, count);
    if (n < 0) return n;
    FLX_NCCL(ctx, g_nccl.GroupStart());
    for (int k = 0; k < n; k++)
        FLX_NCCL(ctx, g_nccl.Reduce(bufs[k], bufs[k], count, kNcclFloat32, kNcclSum, root, (ncclComm_t)ctx->comm, ctx->stream));
    FLX_NCCL(ctx, g_nccl.GroupEnd());
    return FLX_OK;
}

} // extern "C"
