// TEST INFRASTRUCTURE — the buffer set the reference's kernels are bound to: the host-memory
// counterpart of CLContext::deviceBuffers (src/clcontext.hpp:169-214).  Plain C; mirrored by a
// ctypes.Structure in oracle/refcl.py.
#pragma once
#include <stdint.h>

typedef struct refcl_buffers
{
    void *tasks;               // GPUTaskState SoA, 64 x numTasks words (tasksBuffer)
    float *pixels;             // float4 x numPixels accumulation (pixelBuffer)
    float *denoiserAlbedo;     // float4 x numPixels
    float *denoiserNormal;     // float4 x numPixels
    float *pixelsPreview;      // float4 x numPixels (previewBuffer, a GL PBO in the reference)
    float *denoiserAlbedoGL;   // float4 x numPixels
    float *denoiserNormalGL;   // float4 x numPixels
    void *queueCounters;       // QueueCounters (40 B)
    uint32_t *queues[9];       // raygen, extension, shadow, diffuse, glossy, ggxRefl, ggxRefr, delta, emissive
    const void *tris;          // Triangle[] 160 B
    const void *nodes;         // GPUNode[] 48 B
    const uint32_t *indices;
    const void *materials;     // Material[] 96 B
    const uint8_t *texData;
    const void *textures;      // TexDescriptor[] 12 B
    const float *envRGBA;      // image2d_t RGBA32F
    int32_t envW, envH;
    const float *probTable;
    const int32_t *aliasTable;
    const float *pdfTable;
    void *params;              // RenderParams (256 B)
    uint32_t *samplesPerPixel; // numPixels
    uint32_t *currPixelIdx;    // 1 word
    void *stats;               // RenderStats (16 B)
    void *pickResult;          // Hit (64 B)
    uint32_t numTasks, numPixels;
    int32_t threads;           // host threads for the NDRange loops (1 = serial, ascending gid)
} refcl_buffers;
