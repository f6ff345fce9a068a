/* fluctus_abi.h — byte layouts that cross the drop-in boundary.
 *
 * These are the host-side structs the reference uploads verbatim to its device
 * (CLContext::uploadSceneData, reference src/clcontext.cpp:540-629; updateParams :721-725;
 * enqueueGetCounters :686-689).  The CUDA backend accepts exactly these bytes and re-packs them
 * on its side.  Sizes/offsets are the ones measured from the reference headers with the host
 * `float3` being 16 bytes (reference include/math/float3.hpp:31-37) — see SURVEY.md §8(a).
 * Plain C, no dependencies.
 */
#ifndef FLUCTUS_ABI_H
#define FLUCTUS_ABI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float x, y, z, w; } flx_float3; /* 16 B; .w is padding */
typedef struct { float x, y; } flx_float2;

/* reference src/triangle.hpp:7-16 (VertexPNT) */
typedef struct { flx_float3 p, n, t; } flx_vertex; /* 48 B */

/* reference src/triangle.hpp:18-49 (RTTriangle) == src/geom.h:99-112 (Triangle); 160 B */
typedef struct {
    flx_vertex v0, v1, v2;
    int32_t matId;
    int32_t _pad[3];
} flx_triangle;

/* reference src/bvhnode.hpp:50-59 (Node) == src/geom.h:88-97 (GPUNode); 48 B.
 * DFS order, left child = index+1; leaf iff nPrims != 0. */
typedef struct {
    flx_float3 bmin, bmax;
    int32_t parent;
    uint32_t iStartOrRight; /* leaf: first slot in the index list; inner: right child */
    uint8_t nPrims;
    uint8_t _pad[7];
} flx_node;

/* reference src/geom.h:130-143 (Material); 96 B */
typedef struct {
    flx_float3 Kd, Ks, Ke, Kt;
    float Ns; /* GGX alpha after toRoughness (reference src/scene.cpp:13-16) */
    float Ni;
    float d;
    int32_t map_Kd, map_Ks, map_N;
    int32_t type; /* FLX_BXDF_* */
    int32_t _pad;
} flx_material;

/* reference src/geom.h:145-150 (TexDescriptor); 12 B */
typedef struct { uint32_t offset, width, height; } flx_tex_desc;

/* reference src/geom.h:119-127 (AreaLight); 96 B */
typedef struct {
    flx_float3 right, up, N, pos, E;
    flx_float2 size;
    float _pad[2];
} flx_area_light;

/* reference src/geom.h:165-175 (Camera); 80 B */
typedef struct {
    flx_float3 pos, dir, up, right;
    float fov, fovSCALE, apertureSize, focalDist;
} flx_camera;

/* reference src/geom.h:177-181 */
typedef struct { float exposure; uint32_t tmOperator; } flx_pp_params;

/* reference src/geom.h:183-203 (RenderParams); 256 B */
typedef struct {
    flx_area_light areaLight; /*   0 */
    flx_camera camera;        /*  96 */
    flx_pp_params ppParams;   /* 176 */
    uint32_t width;           /* 184 */
    uint32_t height;
    uint32_t n_tris;
    uint32_t useEnvMap;
    uint32_t useAreaLight;
    float envMapStrength;
    uint32_t maxBounces;
    uint32_t sampleImpl;
    uint32_t sampleExpl;
    uint32_t useRoulette;
    uint32_t wfSeparateQueues;
    uint32_t maxSpp;
    float worldRadius;        /* 232 */
    float width1;
    float height1;
    uint32_t _pad[3];
} flx_render_params;

/* reference src/geom.h:152-163 (Hit); 64 B */
typedef struct {
    flx_float3 P, N;
    flx_float2 uvTex;
    float t;
    int32_t i;
    int32_t areaLightHit;
    int32_t matId;
    int32_t _pad[2];
} flx_hit;

/* reference src/geom.h:263-277 (QueueCounters); 40 B, field order is part of the contract
 * (Tracer reads it as a struct, reference src/tracer.cpp:364-367) */
typedef struct {
    uint32_t raygenQueue, extensionQueue, shadowQueue;
    uint32_t diffuseQueue, glossyQueue, ggxReflQueue, ggxRefrQueue, deltaQueue, emissiveQueue;
    uint32_t splattedSamples;
} flx_queue_counters;

/* reference src/geom.h:279-285 (RenderStats); 16 B */
typedef struct { uint32_t primaryRays, extensionRays, shadowRays, samples; } flx_render_stats;

/* reference src/bxdf_types.h:4-12 */
enum {
    FLX_BXDF_DIFFUSE = 1 << 1,
    FLX_BXDF_GLOSSY = 1 << 2,
    FLX_BXDF_GGX_ROUGH_REFLECTION = 1 << 3,
    FLX_BXDF_IDEAL_REFLECTION = 1 << 4,
    FLX_BXDF_GGX_ROUGH_DIELECTRIC = 1 << 5,
    FLX_BXDF_IDEAL_DIELECTRIC = 1 << 6,
    FLX_BXDF_EMISSIVE = 1 << 7,
    FLX_BXDF_MIXED = 1 << 8
};
#define FLX_BXDF_IS_SINGULAR(t) (((t) & (FLX_BXDF_IDEAL_REFLECTION | FLX_BXDF_IDEAL_DIELECTRIC)) != 0)

/* Path-state interchange layout: the reference's GPUTaskState in -DUSE_SOA form
 * (reference src/geom.h:43-67, 222-259): word slot s of task g lives at [s * numTasks + g].
 * 64 four-byte slots.  flx_export_state / flx_import_state speak this layout so that stage
 * parity can be checked field by field whatever the device-side packing is. */
enum {
    FLX_SLOT_ORIG = 0, FLX_SLOT_DIR = 4, FLX_SLOT_SHADOW_ORIG = 8, FLX_SLOT_SHADOW_DIR = 12,
    FLX_SLOT_T = 16, FLX_SLOT_EI = 20, FLX_SLOT_LAST_BSDF = 24, FLX_SLOT_LAST_EMISSION = 28,
    FLX_SLOT_LAST_T = 32, FLX_SLOT_P = 36, FLX_SLOT_N = 40, FLX_SLOT_UV = 44,
    FLX_SLOT_PHASE = 46, FLX_SLOT_LAST_PDF_W = 47, FLX_SLOT_PATH_LEN = 48, FLX_SLOT_SEED = 49,
    FLX_SLOT_LAST_SPECULAR = 50, FLX_SLOT_SHADOW_BLOCKED = 51, FLX_SLOT_BACKFACE = 52,
    FLX_SLOT_PIXEL_INDEX = 53, FLX_SLOT_FIRST_DIFFUSE = 54, FLX_SLOT_LAST_PDF_DIRECT = 55,
    FLX_SLOT_LAST_PDF_IMPLICIT = 56, FLX_SLOT_LAST_COS_TH = 57, FLX_SLOT_LAST_PICK_PROB = 58,
    FLX_SLOT_SHADOW_LEN = 59, FLX_SLOT_HIT_T = 60, FLX_SLOT_HIT_I = 61,
    FLX_SLOT_AREA_LIGHT_HIT = 62, FLX_SLOT_MAT_ID = 63, FLX_NUM_SLOTS = 64
};

/* queue ids for flx_read_queue / flx_write_queue (order of reference src/clcontext.cpp:130-138) */
enum {
    FLX_Q_RAYGEN = 0, FLX_Q_EXTENSION, FLX_Q_SHADOW, FLX_Q_DIFFUSE, FLX_Q_GLOSSY,
    FLX_Q_GGX_REFL, FLX_Q_GGX_REFR, FLX_Q_DELTA, FLX_Q_EMISSIVE, FLX_NUM_QUEUES
};

#ifdef __cplusplus
}
#endif
#endif /* FLUCTUS_ABI_H */
