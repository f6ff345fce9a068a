// TEST INFRASTRUCTURE — CPU oracle.  Restates reference src/intersect.cl and src/bvh.cl.
#include "o_trace.h"

namespace orc {

// reference src/intersect.cl:41-60 (intersectAABB).  dinv is recomputed per box there; the
// value is the same every time so hoisting would not change bits, but we keep the shape.
bool intersectAABB(const Ray &r, const flx_float3 &bmin, const flx_float3 &bmax, float *tminRet, float *tmaxRet, float tMaxPrev)
{
    const f3 dinv = mk3(recip(r.dir.x), recip(r.dir.y), recip(r.dir.z));
    const f3 lo = (ld3(bmin) - r.orig) * dinv;
    f3 hi = (ld3(bmax) - r.orig) * dinv;
    const f3 tminv = fmin3(lo, hi);
    hi = fmax3(lo, hi);

    const float tmin = fmaxf(fmaxf(tminv.x, tminv.y), tminv.z);
    const float tmax = fminf(fminf(hi.x, hi.y), hi.z);

    if (tmax < 0) return false;
    if (tmin > tmax) return false;
    *tminRet = tmin;
    *tmaxRet = tmax;
    return tmin < tMaxPrev;
}

// reference src/intersect.cl:63-93 (Moeller-Trumbore, EPSILON 1e-12f)
static inline bool mollerTrumbore(const Ray &r, f3 p0, f3 p1, f3 p2, float *tret, float *uret, float *vret)
{
    const f3 s1 = p1 - p0;
    const f3 s2 = p2 - p0;
    const f3 pvec = cross(r.dir, s2);
    const float det = dot(s1, pvec);
    if (fabsf(det) < 1e-12f) return false;
    const float iDet = recip(det);

    const f3 tvec = r.orig - p0;
    const float u = dot(tvec, pvec) * iDet;
    if (u < 0.0f || u > 1.0f) return false;

    const f3 qvec = cross(tvec, s1);
    const float v = dot(r.dir, qvec) * iDet;
    if (v < 0.0f || u + v > 1.0f) return false;

    const float t = dot(s2, qvec) * iDet;
    if (t < 0.0f) return false;
    *tret = t; *uret = u; *vret = v;
    return true;
}

bool intersectTriangle(const Ray &r, const flx_triangle &tri, float *t, float *u, float *v)
{
    return mollerTrumbore(r, ld3(tri.v0.p), ld3(tri.v1.p), ld3(tri.v2.p), t, u, v);
}

// reference src/intersect.cl:96-121 (intersectTriangleLocal): also rejects t > *tres
static inline bool triLocal(const Ray &r, f3 p0, f3 p1, f3 p2, float *tres)
{
    float t, u, v;
    if (!mollerTrumbore(r, p0, p1, p2, &t, &u, &v)) return false;
    if (t > *tres) return false;
    *tres = t;
    return true;
}

// reference src/intersect.cl:124-155 (intersectLight)
void intersectLight(Hit *hit, const Ray &r, const flx_render_params *params)
{
    const flx_area_light &al = params->areaLight;
    const f3 N = ld3(al.N), pos = ld3(al.pos), right = ld3(al.right), up = ld3(al.up);
    if (dot(r.dir, N) > 0) return;

    const float sx = al.size.x, sy = al.size.y;
    const f3 tl = pos + sx * right + sy * up;
    const f3 tr = pos - sx * right + sy * up;
    const f3 bl = pos + sx * right - sy * up;
    const f3 br = pos - sx * right - sy * up;

    const bool first = triLocal(r, tl, bl, br, &hit->t);
    const bool second = triLocal(r, tl, br, tr, &hit->t);
    if (first || second)
    {
        hit->areaLightHit = 1;
        hit->P = r.orig + hit->t * r.dir;
        hit->N = N;
        hit->i = 0;
        hit->matId = 0;
    }
}

static inline f3 lerp3(float u, float v, f3 a, f3 b, f3 c) // reference src/utils.cl:25-28
{
    return (1.0f - u - v) * a + u * b + v * c;
}

// Leaf body shared by both traversal variants: reference src/bvh.cl:24-51 == :256-281
static inline void leafClosest(const Ray &r, Hit *hit, const SceneView &s, uint32_t iStart, uint32_t nPrims, TraceCounters *c)
{
    float tmin = FLT_MAX, umin = 0.0f, vmin = 0.0f;
    int imin = -1;
    for (uint32_t i = iStart; i < iStart + nPrims; i++)
    {
        float t, u, v;
        if (c) c->tri_tests++;
        if (intersectTriangle(r, s.tris[s.indices[i]], &t, &u, &v))
        {
            if (t > 0.0f && t < tmin) { imin = (int)i; tmin = t; umin = u; vmin = v; }
        }
    }
    if (imin != -1 && tmin < hit->t)
    {
        const flx_triangle &tri = s.tris[s.indices[imin]];
        hit->i = (int)s.indices[imin];
        hit->matId = tri.matId;
        hit->t = tmin;
        hit->P = r.orig + tmin * r.dir;
        hit->N = normalize(lerp3(umin, vmin, ld3(tri.v0.n), ld3(tri.v1.n), ld3(tri.v2.n)));
        const f3 uv = lerp3(umin, vmin, ld3(tri.v0.t), ld3(tri.v1.t), ld3(tri.v2.t));
        hit->uv.x = uv.x; hit->uv.y = uv.y;
    }
}

static inline bool leafAny(const Ray &r, float maxDist, const SceneView &s, uint32_t iStart, uint32_t nPrims, TraceCounters *c)
{
    for (uint32_t i = iStart; i < iStart + nPrims; i++)
    {
        float t, u, v;
        if (c) c->tri_tests++;
        if (intersectTriangle(r, s.tris[s.indices[i]], &t, &u, &v) && t > 0.0f && t < maxDist) return true;
    }
    return false;
}

// reference src/bvh.cl:234-310 (stack variant — the default, settings.cpp:27)
// Optional per-node visit histogram (analysis aid for DESIGN.md: which share of the inner-node visits falls on
// the top levels of the tree).  Not thread-safe: use with one thread.
uint32_t *g_nodeVisitHistogram = nullptr;

void bvhIntersectStack(const Ray &r, Hit *hit, const SceneView &s, TraceCounters *c)
{
    float lnear, lfar, rnear, rfar;
    uint32_t stack[64];
    int sp = 0;
    stack[0] = 0;
    while (sp >= 0)
    {
        const uint32_t ni = stack[sp--];
        const flx_node &n = s.nodes[ni];
        if (n.nPrims != 0)
        {
            if (c) c->leaf_visits++;
            leafClosest(r, hit, s, n.iStartOrRight, n.nPrims, c);
        }
        else
        {
            if (c) c->inner_visits++;
            if (g_nodeVisitHistogram) g_nodeVisitHistogram[ni]++;
            const uint32_t li = ni + 1, ri = n.iStartOrRight;
            const bool lh = intersectAABB(r, s.nodes[li].bmin, s.nodes[li].bmax, &lnear, &lfar, hit->t);
            const bool rh = intersectAABB(r, s.nodes[ri].bmin, s.nodes[ri].bmax, &rnear, &rfar, hit->t);
            if (lh && rh)
            {
                uint32_t closer = li, farther = ri;
                if (rnear < lnear) { closer = ri; farther = li; }
                stack[++sp] = farther;
                stack[++sp] = closer;
            }
            else if (lh) stack[++sp] = li;
            else if (rh) stack[++sp] = ri;
            if (c && (uint64_t)(sp + 1) > c->max_stack) c->max_stack = (uint64_t)(sp + 1);
        }
    }
    if (c && hit->i >= 0) c->hits++;
}

// reference src/bvh.cl:312-373
bool bvhOccludedStack(const Ray &r, float maxDist, const SceneView &s, TraceCounters *c)
{
    float lnear, lfar, rnear, rfar;
    uint32_t stack[64];
    int sp = 0;
    stack[0] = 0;
    while (sp >= 0)
    {
        const uint32_t ni = stack[sp--];
        const flx_node &n = s.nodes[ni];
        if (n.nPrims != 0)
        {
            if (c) c->leaf_visits++;
            if (leafAny(r, maxDist, s, n.iStartOrRight, n.nPrims, c)) { if (c) c->hits++; return true; }
        }
        else
        {
            if (c) c->inner_visits++;
            const uint32_t li = ni + 1, ri = n.iStartOrRight;
            const bool lh = intersectAABB(r, s.nodes[li].bmin, s.nodes[li].bmax, &lnear, &lfar, maxDist);
            const bool rh = intersectAABB(r, s.nodes[ri].bmin, s.nodes[ri].bmax, &rnear, &rfar, maxDist);
            if (lh && rh)
            {
                uint32_t closer = li, farther = ri;
                if (rnear < lnear) { closer = ri; farther = li; }
                stack[++sp] = farther;
                stack[++sp] = closer;
            }
            else if (lh) stack[++sp] = li;
            else if (rh) stack[++sp] = ri;
        }
    }
    return false;
}

// Descent/backtrack step shared by the two bit-stack routines (reference src/bvh.cl:53-126).
// Returns false when the traversal is finished.
static inline bool bitstackStep(const SceneView &s, bool r1, bool r2, float t1, float t2, bool leaf,
                                int *top, uint32_t *lstack, uint32_t *rstack)
{
    bool trackback = leaf;
    if (!leaf)
    {
        const flx_node &n = s.nodes[*top];
        if (r1 && r2)
        {
            if (t1 <= t2) { *top = *top + 1; *lstack = (*lstack | 1u) << 1; *rstack <<= 1; }
            else { *top = (int)n.iStartOrRight; *rstack = (*rstack | 1u) << 1; *lstack <<= 1; }
        }
        else if (r1) { *top = *top + 1; *lstack <<= 1; *rstack <<= 1; }
        else if (r2) { *top = (int)n.iStartOrRight; *lstack <<= 1; *rstack <<= 1; }
        else trackback = true;
    }
    if (trackback)
    {
        bool f = false;
        while (*lstack != 0 || *rstack != 0)
        {
            const flx_node &n = s.nodes[*top];
            if ((*lstack & 1u) != 0)
            {
                *top = (int)n.iStartOrRight;
                *lstack &= ~1u; *lstack <<= 1; *rstack <<= 1;
                f = true; break;
            }
            else if ((*rstack & 1u) != 0)
            {
                *top = *top + 1;
                *rstack &= ~1u; *lstack <<= 1; *rstack <<= 1;
                f = true; break;
            }
            *top = n.parent;
            *lstack >>= 1; *rstack >>= 1;
        }
        if (!f) return false;
    }
    return true;
}

// reference src/bvh.cl:12-127 (-DUSE_BITSTACK)
void bvhIntersectBitstack(const Ray &r, Hit *hit, const SceneView &s, TraceCounters *c)
{
    int top = 0;
    uint32_t lstack = 0, rstack = 0;
    while (top != -1)
    {
        const flx_node &n = s.nodes[top];
        bool r1 = false, r2 = false; float t1 = 0, t2 = 0, dummy;
        const bool leaf = n.nPrims != 0;
        if (leaf)
        {
            if (c) c->leaf_visits++;
            leafClosest(r, hit, s, n.iStartOrRight, n.nPrims, c);
        }
        else
        {
            if (c) c->inner_visits++;
            r1 = intersectAABB(r, s.nodes[top + 1].bmin, s.nodes[top + 1].bmax, &t1, &dummy, hit->t);
            r2 = intersectAABB(r, s.nodes[n.iStartOrRight].bmin, s.nodes[n.iStartOrRight].bmax, &t2, &dummy, hit->t);
        }
        if (!bitstackStep(s, r1, r2, t1, t2, leaf, &top, &lstack, &rstack)) break;
    }
    if (c && hit->i >= 0) c->hits++;
}

// reference src/bvh.cl:130-230
bool bvhOccludedBitstack(const Ray &r, float maxDist, const SceneView &s, TraceCounters *c)
{
    int top = 0;
    uint32_t lstack = 0, rstack = 0;
    while (top != -1)
    {
        const flx_node &n = s.nodes[top];
        bool r1 = false, r2 = false; float t1 = 0, t2 = 0, dummy;
        const bool leaf = n.nPrims != 0;
        if (leaf)
        {
            if (c) c->leaf_visits++;
            if (leafAny(r, maxDist, s, n.iStartOrRight, n.nPrims, c)) return true;
        }
        else
        {
            if (c) c->inner_visits++;
            r1 = intersectAABB(r, s.nodes[top + 1].bmin, s.nodes[top + 1].bmax, &t1, &dummy, maxDist);
            r2 = intersectAABB(r, s.nodes[n.iStartOrRight].bmin, s.nodes[n.iStartOrRight].bmax, &t2, &dummy, maxDist);
        }
        if (!bitstackStep(s, r1, r2, t1, t2, leaf, &top, &lstack, &rstack)) break;
    }
    return false;
}

// All-triangles loop with the same triangle test and the same "strictly closer wins" rule.
void bruteIntersect(const Ray &r, Hit *hit, const SceneView &s)
{
    for (uint32_t k = 0; k < s.n_tris; k++)
    {
        float t, u, v;
        if (intersectTriangle(r, s.tris[k], &t, &u, &v) && t > 0.0f && t < hit->t)
        {
            hit->t = t; hit->i = (int)k; hit->matId = s.tris[k].matId;
        }
    }
}

} // namespace orc
