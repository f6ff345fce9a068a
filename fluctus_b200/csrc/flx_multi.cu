// Multi-GPU frame combine for contexts that live in ONE process (the C++ host's `--gpus N`).
//
// The reference is a single-device program: there is nothing to mirror.  Sample sharding gives every GPU a full
// scene replica and its own accumulation buffer `float4 pixels[w*h]` = (sum R, sum G, sum B, #samples); a frame is
// the element-wise SUM of those buffers (`.w` adds up to the true sample count, so the tonemap stage is unchanged).
// With one process per GPU (Python / torchrun) that sum is one NCCL all-reduce (fluctus_b200/multi_gpu.py).  With all
// contexts in one address space no library is needed: B200s on NVSwitch map each other's HBM, so the reduction is a
// kernel that READS THE PEERS' BUFFERS DIRECTLY over NVLink — a reduce-scatter whose shards land where the frame is
// wanted.  GPU d owns rows slice d of the frame: it loads that slice from all N buffers (N-1 remote 16-byte loads per
// pixel, coalesced), adds them in rank order 0..N-1 (deterministic, unlike a ring), stores the total into context
// 0's buffer and zeroes the slice in the others.  "Total in context 0, zero elsewhere" keeps progressive rendering
// valid: the sum over all buffers is unchanged, every context can go on accumulating, and a later combine is again
// exact.  Every GPU moves (N-1)/N of a frame over its own NVLink ports at the same time (33 MB frame, 8 GPUs: 29 MB
// per GPU, ~40 us at 900 GB/s) instead of one root pulling 7 frames.
#include "flx_internal.h"
#include <algorithm>

#define FLX_MAX_GPUS 16

namespace {

struct PeerPlanes { float4 *p[FLX_MAX_GPUS]; int n; };

// One plane (accumulation buffer or a denoiser AOV) of one row slice: total -> p[0], zero -> p[1..n-1]
__global__ void __launch_bounds__(256) k_reduce_slice(PeerPlanes f, uint32_t first, uint32_t count)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    {
        const uint32_t k = first + i;
        float4 s = f.p[0][k];
        for (int r = 1; r < f.n; r++)
        {
            const float4 v = f.p[r][k];
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            f.p[r][k] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        }
        f.p[0][k] = s;
    }
}

__global__ void __launch_bounds__(256) k_add_plane(float4 *dst, const float4 *src, uint32_t count)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    {
        float4 s = dst[i];
        const float4 v = src[i];
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        dst[i] = s;
    }
}

#define MULTI_CUDA(ctx, call)                                                                        \
    do {                                                                                             \
        cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess)                                                                      \
            return flx_fail(ctx, FLX_ERR_CUDA, std::string("flx_reduce_frames: " #call ": ") + cudaGetErrorString(e__)); \
    } while (0)

// every device must be able to address every other device's frame
bool enablePeers(flx_ctx *const *ctxs, int n)
{
    for (int a = 0; a < n; a++)
        for (int b = 0; b < n; b++)
        {
            if (ctxs[a]->device == ctxs[b]->device) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, ctxs[a]->device, ctxs[b]->device) != cudaSuccess || !can) return false;
            cudaSetDevice(ctxs[a]->device);
            const cudaError_t e = cudaDeviceEnablePeerAccess(ctxs[b]->device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) { cudaGetLastError(); return false; }
        }
    return true;
}

} // namespace

extern "C" int flx_reduce_frames(flx_ctx *const *ctxs, int n)
{
    if (!ctxs || n < 1 || n > FLX_MAX_GPUS) return flx_fail(nullptr, FLX_ERR_ARG, "flx_reduce_frames: need 1..16 contexts");
    flx_ctx *root = ctxs[0];
    for (int r = 0; r < n; r++)
    {
        if (!ctxs[r]) return flx_fail(nullptr, FLX_ERR_ARG, "flx_reduce_frames: null context");
        if (!ctxs[r]->fbBlock) return flx_fail(ctxs[r], FLX_ERR_STATE, "flx_reduce_frames: no frame buffers (flx_resize)");
        if (ctxs[r]->fb.numPixels != root->fb.numPixels) return flx_fail(ctxs[r], FLX_ERR_ARG, "flx_reduce_frames: frame sizes differ");
        if (ctxs[r]->denoiserAOV != root->denoiserAOV) return flx_fail(ctxs[r], FLX_ERR_ARG, "flx_reduce_frames: denoiser AOV option differs");
        for (int q = 0; q < r; q++)
            if (ctxs[q] == ctxs[r]) return flx_fail(ctxs[r], FLX_ERR_ARG, "flx_reduce_frames: the same context twice");
    }
    if (n == 1) return FLX_OK;
    // the frames must be complete: drain every context's stream
    for (int r = 0; r < n; r++)
    {
        MULTI_CUDA(ctxs[r], cudaSetDevice(ctxs[r]->device));
        MULTI_CUDA(ctxs[r], cudaStreamSynchronize(ctxs[r]->stream));
    }
    const uint32_t np = root->fb.numPixels;
    const int planes = root->denoiserAOV ? 3 : 1;
    auto plane = [](flx_ctx *c, int k) { return k == 0 ? c->fb.pixels : (k == 1 ? c->fb.denoiserAlbedo : c->fb.denoiserNormal); };

    if (enablePeers(ctxs, n))
    {
        // reduce-scatter over peer memory: context d combines slice d on its own GPU and stream
        for (int d = 0; d < n; d++)
        {
            const uint32_t first = (uint32_t)((uint64_t)np * d / n), last = (uint32_t)((uint64_t)np * (d + 1) / n);
            if (last == first) continue;
            MULTI_CUDA(ctxs[d], cudaSetDevice(ctxs[d]->device));
            const int grid = std::min<int>((int)((last - first + 255u) / 256u), ctxs[d]->numSMs * 8);
            for (int k = 0; k < planes; k++)
            {
                PeerPlanes f; f.n = n;
                for (int r = 0; r < n; r++) f.p[r] = plane(ctxs[r], k);
                k_reduce_slice<<<grid, 256, 0, ctxs[d]->stream>>>(f, first, last - first);
                ctxs[d]->launches++;
                MULTI_CUDA(ctxs[d], cudaGetLastError());
            }
        }
        for (int d = 0; d < n; d++)
        {
            MULTI_CUDA(ctxs[d], cudaSetDevice(ctxs[d]->device));
            MULTI_CUDA(ctxs[d], cudaStreamSynchronize(ctxs[d]->stream));
        }
        return FLX_OK;
    }

    // No peer mapping (GPUs without a common NVLink / PCIe domain): stage each frame on the root and add it there.
    MULTI_CUDA(root, cudaSetDevice(root->device));
    float4 *stage = nullptr;
    MULTI_CUDA(root, cudaMalloc(&stage, (size_t)np * sizeof(float4)));
    int rc = FLX_OK;
    for (int r = 1; r < n && rc == FLX_OK; r++)
        for (int k = 0; k < planes && rc == FLX_OK; k++)
        {
            cudaError_t e = cudaMemcpyPeerAsync(stage, root->device, plane(ctxs[r], k), ctxs[r]->device, (size_t)np * sizeof(float4), root->stream);
            if (e == cudaSuccess)
            {
                k_add_plane<<<root->numSMs * 8, 256, 0, root->stream>>>(plane(root, k), stage, np);
                root->launches++;
                e = cudaGetLastError();
            }
            if (e == cudaSuccess) e = cudaStreamSynchronize(root->stream);
            if (e == cudaSuccess) { cudaSetDevice(ctxs[r]->device); e = cudaMemset(plane(ctxs[r], k), 0, (size_t)np * sizeof(float4)); cudaSetDevice(root->device); }
            if (e != cudaSuccess) rc = flx_fail(root, FLX_ERR_CUDA, std::string("flx_reduce_frames (staged): ") + cudaGetErrorString(e));
        }
    cudaFree(stage);
    return rc;
}
