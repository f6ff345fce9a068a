// Device-side BVH layout and traversal.
//
// The host hands over the reference's arrays verbatim (Node 48 B, RTTriangle 160 B, index list);
// flx_trace.cu re-packs them once per scene into three 16-byte-vectorised streams:
//
//   inner node  64 B = 4 x float4   {L.min.xyz L.max.x} {L.max.yz R.min.xy} {R.min.z R.max.xyz}
//                                    {leftRef, rightRef, parent, -}
//       both children's boxes live in the parent, so one node visit is one 64-byte fetch instead
//       of the reference's 48 B node + two 48 B child fetches (src/bvh.cl:250, :284-285).
//       ref >= 0: index of an inner node in this array;  ref < 0: ~ref = first leaf slot.
//   leaf slot   48 B = 3 x float4   {v0.xyz, triIndex} {(v1-v0).xyz, lastFlag} {(v2-v0).xyz, matId}
//       first vertex + the two edge vectors of Moeller-Trumbore, stored in index-list order (one slot per entry of the reference's
//       `indices`), so the indices[i] -> tris[...] double indirection of src/bvh.cl:262 is gone;
//       lastFlag != 0 marks the last slot of a leaf, which replaces nPrims.
//   shading rec 64 B = 4 x float4   n0.xyz n1.xyz n2.xyz uv0 uv1 uv2 matId, indexed by triangle id,
//       read once per ray for the final hit only.
//
// Arithmetic and visit order are those of the reference (src/intersect.cl:41-93, src/bvh.cl:234-373):
// near child first, left first on equal entry distance, children tested against the hit distance
// current at the parent (no re-test on pop), leaf triangles in index order, strictly-closer wins.
#pragma once
#include "flx_math.cuh"

struct DevBVH
{
    const float4 *nodes;  // 4 per inner node
    const float4 *slots;  // 3 per leaf slot
    const float4 *shade;  // 4 per triangle
    int rootRef;
    int topN;             // inner nodes [0, topN) are the breadth-first top of the tree (staged in shared memory)
};

// Top of the tree in shared memory: FLX_TOP_NODES x 64 B per CTA, stored as four float4 planes (plane k holds
// float4 k of every node) so that lanes reading different nodes spread over the banks.  440 nodes = 28 160 B per CTA,
// two CTAs of 640 threads per SM (see flx_trace.cu).
#ifndef FLX_TOP_NODES
#define FLX_TOP_NODES 440
#endif
#define FLX_TOP_SMEM_BYTES (FLX_TOP_NODES * 64)

FLX_DEV void stageTopNodes(const DevBVH &bvh, float4 *top)
{
    for (int e = threadIdx.x; e < 4 * bvh.topN; e += blockDim.x)
        top[(e & 3) * FLX_TOP_NODES + (e >> 2)] = __ldg(bvh.nodes + e);
    __syncthreads();
}

#define FLX_STACK_SIZE 64 // reference: uint stack[64] (src/bvh.cl:240)
#define FLX_BITSTACK_MAX_DEPTH 31 // -DUSE_BITSTACK keeps one flag per level in a 32-bit word (src/bvh.cl:14-15)

// One 256-bit load (sm_100: LDG.E.256) = half an inner node.  Rays of a warp sit in different nodes, so every lane of
// a load instruction costs the L1 its own tag look-up whatever the access width: fetching the 64-byte node with two
// 32-byte loads instead of four 16-byte ones halves the L1 wavefronts of a node visit.
#ifndef FLX_SLOT_LD_EARLY
#define FLX_SLOT_LD_EARLY 1
#endif
#ifndef FLX_NODE_LD256
#define FLX_NODE_LD256 0
#endif
FLX_DEV void ldg256(const float4 *p, float4 &a, float4 &b)
{
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
        : "l"(p));
}

// src/intersect.cl:41-60 with dinv hoisted (same value each time => same bits)
FLX_DEV bool slabTest(float bx0, float by0, float bz0, float bx1, float by1, float bz1,
                      f3 o, f3 dinv, float tMaxPrev, float &tnear)
{
    const float lx = (bx0 - o.x) * dinv.x, ly = (by0 - o.y) * dinv.y, lz = (bz0 - o.z) * dinv.z;
    const float hx = (bx1 - o.x) * dinv.x, hy = (by1 - o.y) * dinv.y, hz = (bz1 - o.z) * dinv.z;
    const float tmin = fmaxf(fmaxf(fminf(lx, hx), fminf(ly, hy)), fminf(lz, hz));
    const float tmax = fminf(fminf(fmaxf(lx, hx), fmaxf(ly, hy)), fmaxf(lz, hz));
    tnear = tmin;
    if (tmax < 0.0f) return false;
    if (tmin > tmax) return false;
    return tmin < tMaxPrev;
}

// src/intersect.cl:63-93 (Moeller-Trumbore, EPSILON 1e-12f) with the two edge vectors s1 = v1 - v0 and
// s2 = v2 - v0 supplied by the caller: the leaf slots store them precomputed (one IEEE subtraction per
// component, done once at upload instead of once per test — same operands, same rounding, same bits).
// `never` is a condition the caller knows to be false; see the leaf body of traceLoop for what it is good for.
FLX_DEV bool triTestEdges(f3 o, f3 d, f3 p0, f3 s1, f3 s2, float &t, float &u, float &v, bool never = false)
{
    const f3 pvec = cross(d, s2);
    const float det = dot(s1, pvec);
    if (fabsf(det) < 1e-12f || never) return false;
    const float iDet = recip(det);
    const f3 tvec = o - p0;
    u = dot(tvec, pvec) * iDet;
    if (u < 0.0f || u > 1.0f) return false;
    const f3 qvec = cross(tvec, s1);
    v = dot(d, qvec) * iDet;
    if (v < 0.0f || u + v > 1.0f) return false;
    t = dot(s2, qvec) * iDet;
    if (t < 0.0f) return false;
    return true;
}
FLX_DEV bool triTest(f3 o, f3 d, f3 p0, f3 p1, f3 p2, float &t, float &u, float &v)
{
    return triTestEdges(o, d, p0, p1 - p0, p2 - p0, t, u, v);
}

struct RayHit
{
    float t, u, v;
    int slot; // winning leaf slot, -1 = miss
};

// Visit counters for the algorithmic-bytes figure: inner-node pops, leaf visits,
// triangle tests.  Only the COUNT=true instantiations touch them; the timed kernels do not.
struct TravCount { unsigned int inner, leaf, tri; };

// ---------------------------------------------------------------------------------------------------
// Traversal engine: persistent warps, ray replacement, phase voting.
//
// Closest hit (ANY=false) follows src/bvh.cl:234-310, any hit (ANY=true) src/bvh.cl:312-373 (returns as
// soon as a triangle with 0 < t < tmax is found).  The PER-RAY sequence of box tests, pushes, pops and
// triangle tests is exactly the reference's; what is new is how the work of 32 rays is issued by a warp.
// A lane that owns a ray is in one of three states: INNER (next op: test the two child boxes of node
// `cur`), LEAF (next op: one triangle test at leaf slot ~cur) or DONE (result ready to be written).  In the
// reference — and in a straightforward CUDA port — every loop iteration executes the inner-node body AND
// the leaf body under partial masks, and the warp idles until its slowest ray is finished; ncu showed 10 of
// 32 lanes active per instruction for that design (profiles/r1a_k_ext_full.md).  Here each iteration the warp
// VOTES (ballot + popc) and executes exactly ONE body for the lanes that need it:
//   * retire + refill when >= FLX_VOTE_RETIRE lanes are DONE or idle (or nothing else can run): DONE lanes
//     write their results together and idle lanes take the next rays with ONE atomicAdd per warp;
//   * else the triangle body when >= FLX_VOTE_LEAF lanes wait on a leaf (or no lane is INNER);
//   * else the inner-node body.
// Once a body is chosen the warp stays in it while >= FLX_STAY_* lanes still have work of that kind (one ballot
// per pass instead of the full vote).  Lanes whose state was not chosen wait; they are picked up as soon as
// enough peers join them.  Thresholds were tuned on B200 with tools/tune_vote.sh (kitchen render + cfg5 rays); re-tuned
// for the 2^22-path pool and the fused extension+shadow launch: 12/8/4/8 -> 16/6/3/12 is +7 % on the kitchen, cfg5 unchanged.
// A Job supplies the rays and consumes the results:
//     static constexpr bool BATCH                        // which threshold set the loop uses (see below)
//     bool load(unsigned k, f3 &o, f3 &d, float &tmax)   // false: ray k needs no traversal (already resolved)
//     void store(const RayHit &h, f3 o, f3 d)            // result of the ray loaded last by this lane
#ifndef FLX_VOTE_RETIRE
#define FLX_VOTE_RETIRE 16
#endif
#ifndef FLX_VOTE_LEAF
#define FLX_VOTE_LEAF 6
#endif
#ifndef FLX_STAY_LEAF
#define FLX_STAY_LEAF 3
#endif
#ifndef FLX_STAY_INNER
#define FLX_STAY_INNER 12
#endif
// Ray batches (k_trace_batch / k_any_batch: cheap load / store bodies, incoherent rays of very different lengths) take
// their own set through Job::BATCH; tuned on cfg5 with tools/tune_batch.sh: refilling at 8 free lanes instead of 16
// is +5 % there (2 610 -> 2 740 Mrays/s: the retire body of a batch is two loads and two stores), 4 the same, 24 is -17 %;
// the other three thresholds sit on a plateau (23 settings within 3 %).
#ifndef FLX_VOTE_RETIRE_BATCH
#define FLX_VOTE_RETIRE_BATCH 8
#endif
#ifndef FLX_VOTE_LEAF_BATCH
#define FLX_VOTE_LEAF_BATCH FLX_VOTE_LEAF
#endif
#ifndef FLX_STAY_LEAF_BATCH
#define FLX_STAY_LEAF_BATCH FLX_STAY_LEAF
#endif
#ifndef FLX_STAY_INNER_BATCH
#define FLX_STAY_INNER_BATCH FLX_STAY_INNER
#endif
// any-hit rays (shorter traversals, early exit) may take their own thresholds; default: the same
#ifndef FLX_VOTE_RETIRE_ANY
#define FLX_VOTE_RETIRE_ANY FLX_VOTE_RETIRE
#endif
#ifndef FLX_VOTE_LEAF_ANY
#define FLX_VOTE_LEAF_ANY FLX_VOTE_LEAF
#endif
#ifndef FLX_STAY_LEAF_ANY
#define FLX_STAY_LEAF_ANY FLX_STAY_LEAF
#endif
#ifndef FLX_STAY_INNER_ANY
#define FLX_STAY_INNER_ANY FLX_STAY_INNER
#endif
#define FLX_TRAV_DONE ((int)0x80000000) // not a valid ref: leaf refs are ~slot with slot < 0x7fffffff

// Prefetch hints (experiment switches): FLX_PF_CUR covers the record a lane steps to next (it
// usually waits at least one vote before touching it), FLX_PF_PUSH the record of a postponed far child — the reference
// never re-tests a node on pop (src/bvh.cl:246-309), so for closest-hit rays every pushed node IS visited later and
// the hint is never wasted.  0 = off, 1 = prefetch.global.L2, 2 = prefetch.global.L1.  No registers are held.
#ifndef FLX_PF_CUR
#define FLX_PF_CUR 0
#endif
#ifndef FLX_PF_PUSH
#define FLX_PF_PUSH 0
#endif
template <int LEVEL>
FLX_DEV void prefetchRef(const DevBVH &bvh, int ref)
{
    if (LEVEL == 0) return;
    if (ref == FLX_TRAV_DONE || (ref >= 0 && ref < bvh.topN)) return; // nothing to fetch / staged in shared memory
    const void *p = ref >= 0 ? (const void *)(bvh.nodes + 4 * (size_t)ref) : (const void *)(bvh.slots + 3 * (size_t)(~ref));
    if (LEVEL == 1) asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    else asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
}

// How a ray remembers the far children it postponed — the two halves of src/bvh.cl.
//   StackTrail: `uint stack[64]` of node references (src/bvh.cl:234-373, the reference's default).
//   BitTrail:   -DUSE_BITSTACK (src/bvh.cl:12-230, Settings::clUseBitstack): no stack, two 32-bit trails — bit k of
//               `lst` (`rst`) says "the right (left) sibling at level k is still to be visited" — and parent links to
//               walk back up; depth <= 31.  Same visit order, same box tests (a postponed child is not re-tested), so
//               the same hits bit for bit; what differs is the memory behaviour: no 256-byte local-memory stack per
//               ray, but one 16-byte reference fetch per level climbed on the way back.
// The inner-node record keeps the parent's index in its third reference word for the climb.
// (The state of a trail is scalars only and the stack array lives beside it in traceLoop: an array inside the struct
// keeps the whole struct — the stack pointer included — in local memory.)
struct StackTrail
{
    enum { SLOTS = FLX_STACK_SIZE };
    int sp;
    FLX_DEV void reset() { sp = 0; }
    FLX_DEV void both(int *stack, int, bool, int far) { stack[sp++] = far; } // node, left child first?, postponed child
    FLX_DEV void one(int) {}
    FLX_DEV int back(const int *stack, const DevBVH &, const float4 *) { return sp ? stack[--sp] : FLX_TRAV_DONE; }
};
struct BitTrail
{
    enum { SLOTS = 1 };
    unsigned lst, rst;
    int par; // inner node the current node was entered from
    FLX_DEV void reset() { lst = 0u; rst = 0u; par = -1; }
    FLX_DEV void both(int *, int node, bool leftFirst, int) // src/bvh.cl:61-76
    {
        if (leftFirst) { lst = (lst | 1u) << 1; rst <<= 1; }
        else { rst = (rst | 1u) << 1; lst <<= 1; }
        par = node;
    }
    FLX_DEV void one(int node) { lst <<= 1; rst <<= 1; par = node; } // :77-88
    FLX_DEV int back(const int *, const DevBVH &bvh, const float4 *top) // the trackback loop, :95-124
    {
        if ((lst | rst) == 0u) return FLX_TRAV_DONE;
        int t = par; // first turn of the reference's loop: the flags of the current level are clear, step to the parent
        lst >>= 1; rst >>= 1;
        while ((lst | rst) != 0u)
        {
            const float4 n3 = t < bvh.topN ? top[3 * FLX_TOP_NODES + t] : __ldg(bvh.nodes + 4 * (size_t)t + 3);
            if (lst & 1u) { lst = (lst & ~1u) << 1; rst <<= 1; par = t; return __float_as_int(n3.y); } // right sibling is due
            if (rst & 1u) { rst = (rst & ~1u) << 1; lst <<= 1; par = t; return __float_as_int(n3.x); } // left sibling is due
            t = __float_as_int(n3.z);
            lst >>= 1; rst >>= 1;
        }
        return FLX_TRAV_DONE;
    }
};

template <bool ANY, bool COUNT, class Trail, class Job>
FLX_DEV void traceLoop(const DevBVH &bvh, const float4 *__restrict__ top, Job &job, unsigned int *cursor, unsigned int total, TravCount &cnt)
{
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u, ltMask = (1u << lane) - 1u;
    constexpr bool BATCH = Job::BATCH;
    constexpr int VOTE_RETIRE = BATCH ? FLX_VOTE_RETIRE_BATCH : (ANY ? FLX_VOTE_RETIRE_ANY : FLX_VOTE_RETIRE);
    constexpr int VOTE_LEAF = BATCH ? FLX_VOTE_LEAF_BATCH : (ANY ? FLX_VOTE_LEAF_ANY : FLX_VOTE_LEAF);
    constexpr int STAY_LEAF = BATCH ? FLX_STAY_LEAF_BATCH : (ANY ? FLX_STAY_LEAF_ANY : FLX_STAY_LEAF);
    constexpr int STAY_INNER = BATCH ? FLX_STAY_INNER_BATCH : (ANY ? FLX_STAY_INNER_ANY : FLX_STAY_INNER);
    Trail trail; trail.reset();
    int stack[Trail::SLOTS];
    int cur = FLX_TRAV_DONE;
    bool active = false, poolEmpty = false;
    f3 o = mk3(0.0f), d = mk3(0.0f), dinv = mk3(0.0f);
    RayHit best; best.t = 0.0f; best.u = 0.0f; best.v = 0.0f; best.slot = -1;

    while (true)
    {
        const bool isDone = active && cur == FLX_TRAV_DONE;
        const unsigned mInner = __ballot_sync(FULL, active && cur >= 0);
        const unsigned mLeaf = __ballot_sync(FULL, active && cur < 0 && cur != FLX_TRAV_DONE);
        const unsigned mDone = __ballot_sync(FULL, isDone);
        const unsigned mIdle = poolEmpty ? 0u : __ballot_sync(FULL, !active);
        const unsigned runnable = mInner | mLeaf;
        const int waiting = __popc(mDone | mIdle);

        if (waiting >= VOTE_RETIRE || (runnable == 0u && waiting > 0))
        {
            // ---- retire finished rays, hand the next rays of the pool to the free lanes ----------------
            if (isDone) { job.store(best, o, d); active = false; }
            if (!poolEmpty)
            {
                const unsigned freeMask = mDone | mIdle;
                const unsigned n = (unsigned)__popc(freeMask);
                unsigned base = 0u;
                if (lane == 0u) base = atomicAdd(cursor, n);
                base = __shfl_sync(FULL, base, 0);
                if (!active)
                {
                    const unsigned k = base + (unsigned)__popc(freeMask & ltMask);
                    if (k < total)
                    {
                        float tmax;
                        if (job.load(k, o, d, tmax))
                        {
                            dinv = mk3(recip(d.x), recip(d.y), recip(d.z));
                            best.t = tmax; best.u = 0.0f; best.v = 0.0f; best.slot = -1;
                            cur = bvh.rootRef; trail.reset(); active = true;
                        }
                    }
                }
                if (base + n >= total) poolEmpty = true;
            }
        }
        else if (runnable == 0u) break; // nothing running, nothing to retire, pool empty
        else if (mLeaf != 0u && (__popc(mLeaf) >= VOTE_LEAF || mInner == 0u))
        {
            // ---- triangle tests, one per waiting lane and pass: leaf slots in index order, strictly-closer
            // wins.  The warp stays in this body while enough lanes still hold leaf work (one ballot per pass
            // instead of the full vote).
            unsigned m;
            do
            {
                if (active && cur < 0 && cur != FLX_TRAV_DONE)
                {
                    const int slot = ~cur;
                    if (COUNT) cnt.tri++;
                    const float4 *q = bvh.slots + 3 * (size_t)slot;
                    const float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2);
                    // v0 is not needed before the determinant test, and ptxas sinks its load below that branch: a second,
                    // dependent trip to L1 per triangle.  Making the branch look at the loaded word (a triangle index is
                    // never 0xffffffff) keeps the three loads together; the result of the test is unchanged.
                    // Pays where the traversal waits for memory (ray batches against a tree far larger than the caches: cfg5
                    // +2.6 %), costs where it is bound by issue slots and L1 (the kitchen render: -0.6 %) — batch jobs only.
                    const bool never = BATCH && FLX_SLOT_LD_EARLY && __float_as_uint(q0.w) == 0xffffffffu;
                    float t, u, v;
                    bool found = false;
                    if (triTestEdges(o, d, mk3(q0), mk3(q1), mk3(q2), t, u, v, never) && t > 0.0f && t < best.t)
                    {
                        best.t = t; best.u = u; best.v = v; best.slot = slot;
                        found = true;
                    }
                    if (ANY && found) cur = FLX_TRAV_DONE;
                    else if (__float_as_int(q1.w) != 0) cur = trail.back(stack, bvh, top); // last slot of the leaf
                    else cur = cur - 1;                                                         // ~(slot + 1)
                    prefetchRef<FLX_PF_CUR>(bvh, cur);
                }
                m = __ballot_sync(FULL, active && cur < 0 && cur != FLX_TRAV_DONE);
            } while (__popc(m) >= STAY_LEAF);
        }
        else
        {
            // ---- inner-node steps: both child boxes, near child first --------------------------------------
            unsigned m;
            do
            {
                if (active && cur >= 0)
                {
                    if (COUNT) cnt.inner++;
                    float4 n0, n1, n2, n3;
                    if (cur < bvh.topN)
                    {
                        n0 = top[cur]; n1 = top[FLX_TOP_NODES + cur]; n2 = top[2 * FLX_TOP_NODES + cur]; n3 = top[3 * FLX_TOP_NODES + cur];
                    }
                    else
                    {
                        const float4 *n = bvh.nodes + 4 * (size_t)cur;
#if FLX_NODE_LD256
                        ldg256(n, n0, n1); ldg256(n + 2, n2, n3);
#else
                        n0 = __ldg(n); n1 = __ldg(n + 1); n2 = __ldg(n + 2); n3 = __ldg(n + 3);
#endif
                    }
                    float lnear, rnear;
                    const bool lh = slabTest(n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, o, dinv, best.t, lnear);
                    const bool rh = slabTest(n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, o, dinv, best.t, rnear);
                    const int l = __float_as_int(n3.x), r = __float_as_int(n3.y);
                    if (lh && rh)
                    {
                        const bool swap = rnear < lnear; // left first on equal entry distance (src/bvh.cl:293-300)
                        const int far = swap ? l : r;
                        trail.both(stack, cur, !swap, far);
                        cur = swap ? r : l;
                        if (!ANY) prefetchRef<FLX_PF_PUSH>(bvh, far);
                    }
                    else if (lh) { trail.one(cur); cur = l; }
                    else if (rh) { trail.one(cur); cur = r; }
                    else cur = trail.back(stack, bvh, top);
                    if (COUNT && cur < 0 && cur != FLX_TRAV_DONE) cnt.leaf++;
                    prefetchRef<FLX_PF_CUR>(bvh, cur);
                }
                m = __ballot_sync(FULL, active && cur >= 0);
            } while (__popc(m) >= STAY_INNER);
        }
    }
}

FLX_DEV f3 lerp3(float u, float v, f3 a, f3 b, f3 c) // src/utils.cl:25-28
{
    return (1.0f - u - v) * a + u * b + v * c;
}

// Final-hit attributes (src/bvh.cl:272-280): shading normal and uv from the 64-byte record.
FLX_DEV void fetchHitAttribs(const DevBVH &bvh, const RayHit &h, f3 o, f3 d, int &triIndex, int &matId, f3 &P, f3 &N, float &uvx, float &uvy)
{
    const float4 *q = bvh.slots + 3 * (size_t)h.slot;
    triIndex = __float_as_int(__ldg(q).w);
    const float4 *s = bvh.shade + 4 * (size_t)triIndex;
    const float4 s0 = __ldg(s), s1 = __ldg(s + 1), s2 = __ldg(s + 2), s3 = __ldg(s + 3);
    matId = __float_as_int(s3.w);
    P = o + h.t * d;
    N = normalize(lerp3(h.u, h.v, mk3(s0.x, s0.y, s0.z), mk3(s0.w, s1.x, s1.y), mk3(s1.z, s1.w, s2.x)));
    // only .xy of the interpolated texture coordinate is kept
    const float w0 = 1.0f - h.u - h.v;
    uvx = w0 * s2.y + h.u * s2.w + h.v * s3.y;
    uvy = w0 * s2.z + h.u * s3.x + h.v * s3.z;
}

// src/intersect.cl:96-155 (area-light quad as two local triangles); used by ext and shadow stages
struct AreaLightGeom { f3 tl, tr, bl, br, N; };

FLX_DEV bool triLocal(f3 o, f3 d, f3 p0, f3 p1, f3 p2, float &tres)
{
    float t, u, v;
    if (!triTest(o, d, p0, p1, p2, t, u, v)) return false;
    if (t > tres) return false;
    tres = t;
    return true;
}

FLX_DEV bool intersectLight(const AreaLightGeom &g, f3 o, f3 d, float &t)
{
    if (dot(d, g.N) > 0.0f) return false;
    const bool first = triLocal(o, d, g.tl, g.bl, g.br, t);
    const bool second = triLocal(o, d, g.tl, g.br, g.tr, t);
    return first || second;
}
