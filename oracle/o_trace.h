// TEST INFRASTRUCTURE — CPU oracle (see o_math.h header).  Restatement of the reference's
// ray/box, ray/triangle and BVH traversal code: src/intersect.cl, src/bvh.cl.
#pragma once
#include "o_math.h"
#include "../include/fluctus_abi.h"

namespace orc {

struct Ray { f3 orig, dir; };

struct Hit // register form of reference src/geom.h:152-163
{
    f3 P, N;
    f2 uv;
    float t;
    int i, areaLightHit, matId;
};
static inline Hit emptyHit(float tmax) // EMPTY_HIT(tmax), src/geom.h:165
{
    Hit h; h.P = mk3(0.0f); h.N = mk3(0.0f); h.uv.x = h.uv.y = 0.0f;
    h.t = tmax; h.i = -1; h.areaLightHit = 0; h.matId = -1; return h;
}

struct SceneView // borrowed pointers to the host-format arrays
{
    const flx_triangle *tris; const uint32_t *indices; const flx_node *nodes;
    uint32_t n_tris, n_indices, n_nodes;
};

struct TraceCounters // feeds the algorithmic-bytes figure of SURVEY.md §8(d)
{
    uint64_t inner_visits, leaf_visits, tri_tests, hits, max_stack;
};

static inline f3 ld3(const flx_float3 &v) { return mk3(v.x, v.y, v.z); }

bool intersectAABB(const Ray &r, const flx_float3 &bmin, const flx_float3 &bmax, float *tminRet, float *tmaxRet, float tMaxPrev);
bool intersectTriangle(const Ray &r, const flx_triangle &tri, float *t, float *u, float *v);
void intersectLight(Hit *hit, const Ray &r, const flx_render_params *params);

void bvhIntersectStack(const Ray &r, Hit *hit, const SceneView &s, TraceCounters *c);
void bvhIntersectBitstack(const Ray &r, Hit *hit, const SceneView &s, TraceCounters *c);
void bruteIntersect(const Ray &r, Hit *hit, const SceneView &s);
bool bvhOccludedStack(const Ray &r, float maxDist, const SceneView &s, TraceCounters *c);
bool bvhOccludedBitstack(const Ray &r, float maxDist, const SceneView &s, TraceCounters *c);

} // namespace orc
