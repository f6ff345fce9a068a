// Scene re-packing: the reference's Node[48 B] / RTTriangle[160 B] / index list -> the 16-byte-vectorised
// streams described in flx_bvh.cuh.  Runs once per scene upload (CLContext::uploadSceneData).
#include "flx_internal.h"
#include <cstring>
#include <cstdlib>
#include <utility>

// ---------------------------------------------------------------------------------------------
// Re-pack the reference's Node[48 B] / RTTriangle[160 B] / indices into the streams described in
// flx_bvh.cuh.  Runs once per scene on the host (O(n)), then three bulk uploads.
// ---------------------------------------------------------------------------------------------
int flx_repack_scene(flx_ctx *ctx, const flx_triangle *tris, size_t n_tris, const uint32_t *indices, size_t n_indices,
                     const flx_node *nodes, size_t n_nodes)
{
    if (!tris || !indices || !nodes || n_tris == 0 || n_nodes == 0 || n_indices == 0)
        return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: empty geometry");
    if (n_indices >= 0x7fffffffu || n_nodes >= 0x7fffffffu)
        return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: scene too large for 31-bit references");

    // Compact numbering of inner nodes.  The first FLX_TOP_NODES indices go to the top of the tree in
    // breadth-first order — the traversal kernels stage exactly that prefix in shared memory (on the benchmark
    // scene the first 880 nodes take ~73 % of all inner-node visits, the first 440 most of that, tools/top_level_share.py) — and the rest
    // follow in the input's depth-first order.  Numbering is invisible to the visit order.
    std::vector<int32_t> innerIndex(n_nodes, -1);
    size_t nInner = 0;
    if (nodes[0].nPrims == 0)
    {
        std::vector<uint32_t> bfs;
        bfs.push_back(0);
        for (size_t head = 0; head < bfs.size() && nInner < FLX_TOP_NODES; head++)
        {
            const uint32_t i = bfs[head];
            if (nodes[i].nPrims != 0) continue;
            innerIndex[i] = (int32_t)nInner++;
            const size_t li = (size_t)i + 1, ri = nodes[i].iStartOrRight;
            if (li >= n_nodes || ri >= n_nodes) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: child index out of range");
            bfs.push_back((uint32_t)li);
            bfs.push_back((uint32_t)ri);
        }
    }
    for (size_t i = 0; i < n_nodes; i++)
        if (nodes[i].nPrims == 0 && innerIndex[i] < 0) innerIndex[i] = (int32_t)nInner++;

    // The traversal stack holds one entry per level at most (the far child of every ancestor): the reference
    // uses uint stack[64] unchecked (src/bvh.cl:240); here a deeper tree is refused at upload.
    {
        std::vector<std::pair<uint32_t, uint32_t>> todo(1, std::make_pair(0u, 1u));
        uint32_t maxDepth = 0;
        size_t visited = 0;
        while (!todo.empty())
        {
            const std::pair<uint32_t, uint32_t> it = todo.back(); todo.pop_back();
            if (++visited > n_nodes) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: node array is not a tree");
            if (it.second > maxDepth) maxDepth = it.second;
            const flx_node &n = nodes[it.first];
            if (n.nPrims != 0) continue;
            const size_t li = (size_t)it.first + 1, ri = n.iStartOrRight;
            if (li >= n_nodes || ri >= n_nodes) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: child index out of range");
            todo.push_back(std::make_pair((uint32_t)li, it.second + 1));
            todo.push_back(std::make_pair((uint32_t)ri, it.second + 1));
        }
        if (maxDepth > FLX_STACK_SIZE) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: hierarchy deeper than the 64-entry traversal stack");
        if (ctx->useBitstack && maxDepth > FLX_BITSTACK_MAX_DEPTH)
            return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: hierarchy deeper than the 31 levels the bit-stack traversal can track (flx_set_traversal)");
        ctx->treeDepth = maxDepth;
    }

    auto refOf = [&](size_t child) -> int32_t {
        const flx_node &c = nodes[child];
        return c.nPrims != 0 ? ~(int32_t)c.iStartOrRight : innerIndex[child];
    };

    std::vector<float4> hNodes(4 * (nInner ? nInner : 1));
    std::vector<float4> hSlots(3 * n_indices);
    std::vector<uint8_t> covered(n_indices, 0);

    for (size_t i = 0; i < n_nodes; i++)
    {
        const flx_node &n = nodes[i];
        if (n.nPrims == 0)
        {
            const size_t li = i + 1, ri = n.iStartOrRight;
            if (li >= n_nodes || ri >= n_nodes) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: child index out of range");
            const flx_node &L = nodes[li], &R = nodes[ri];
            float4 *o = &hNodes[4 * (size_t)innerIndex[i]];
            o[0] = make_float4(L.bmin.x, L.bmin.y, L.bmin.z, L.bmax.x);
            o[1] = make_float4(L.bmax.y, L.bmax.z, R.bmin.x, R.bmin.y);
            o[2] = make_float4(R.bmin.z, R.bmax.x, R.bmax.y, R.bmax.z);
            int32_t refs[4] = {refOf(li), refOf(ri), -1, 0}; // [2] = parent (inner index), filled in below
            memcpy(&o[3], refs, 16);
        }
        else
        {
            const size_t s = n.iStartOrRight, e = s + n.nPrims;
            if (e > n_indices) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: leaf range outside index list");
            for (size_t k = s; k < e; k++)
            {
                if (covered[k]) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: overlapping leaf ranges");
                covered[k] = 1;
                const uint32_t ti = indices[k];
                if (ti >= n_tris) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: triangle index out of range");
                const flx_triangle &t = tris[ti];
                const int32_t last = (k + 1 == e) ? 1 : 0;
                float4 *o = &hSlots[3 * k];
                // v0 and the edges s1 = v1 - v0, s2 = v2 - v0 of intersect.cl:66-67: single IEEE binary32
                // subtractions (volatile: no excess precision, no reassociation), identical to the device's FADD
                volatile float e1x = t.v1.p.x - t.v0.p.x, e1y = t.v1.p.y - t.v0.p.y, e1z = t.v1.p.z - t.v0.p.z;
                volatile float e2x = t.v2.p.x - t.v0.p.x, e2y = t.v2.p.y - t.v0.p.y, e2z = t.v2.p.z - t.v0.p.z;
                o[0] = make_float4(t.v0.p.x, t.v0.p.y, t.v0.p.z, 0.f); memcpy(&o[0].w, &ti, 4);
                o[1] = make_float4(e1x, e1y, e1z, 0.f); memcpy(&o[1].w, &last, 4);
                o[2] = make_float4(e2x, e2y, e2z, 0.f); memcpy(&o[2].w, &t.matId, 4);
            }
        }
    }
    // parent links for the bit-stack traversal's climb (src/bvh.cl:121: top = n.parent)
    for (size_t i = 0; i < n_nodes; i++)
    {
        if (nodes[i].nPrims != 0) continue;
        const size_t kids[2] = {i + 1, (size_t)nodes[i].iStartOrRight};
        for (size_t c : kids)
            if (nodes[c].nPrims == 0) memcpy(&hNodes[4 * (size_t)innerIndex[c] + 3].z, &innerIndex[i], 4);
    }
    // uncovered slots (never referenced) still get a terminator so a stray read cannot run away
    for (size_t k = 0; k < n_indices; k++)
        if (!covered[k]) { const int32_t one = 1; hSlots[3 * k] = hSlots[3 * k + 1] = hSlots[3 * k + 2] = make_float4(0, 0, 0, 0); memcpy(&hSlots[3 * k + 1].w, &one, 4); }

    std::vector<float4> hShade(4 * n_tris);
    for (size_t i = 0; i < n_tris; i++)
    {
        const flx_triangle &t = tris[i];
        float4 *o = &hShade[4 * i];
        o[0] = make_float4(t.v0.n.x, t.v0.n.y, t.v0.n.z, t.v1.n.x);
        o[1] = make_float4(t.v1.n.y, t.v1.n.z, t.v2.n.x, t.v2.n.y);
        o[2] = make_float4(t.v2.n.z, t.v0.t.x, t.v0.t.y, t.v1.t.x);
        o[3] = make_float4(t.v1.t.y, t.v2.t.x, t.v2.t.y, 0.f); memcpy(&o[3].w, &t.matId, 4);
    }

    cudaFree(ctx->dNodes); cudaFree(ctx->dSlots); cudaFree(ctx->dShade);
    ctx->dNodes = ctx->dSlots = ctx->dShade = nullptr;
    FLX_CUDA(ctx, cudaMalloc(&ctx->dNodes, hNodes.size() * sizeof(float4)));
    FLX_CUDA(ctx, cudaMalloc(&ctx->dSlots, hSlots.size() * sizeof(float4)));
    FLX_CUDA(ctx, cudaMalloc(&ctx->dShade, hShade.size() * sizeof(float4)));
    FLX_CUDA(ctx, cudaMemcpyAsync(ctx->dNodes, hNodes.data(), hNodes.size() * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    FLX_CUDA(ctx, cudaMemcpyAsync(ctx->dSlots, hSlots.data(), hSlots.size() * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    FLX_CUDA(ctx, cudaMemcpyAsync(ctx->dShade, hShade.data(), hShade.size() * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    FLX_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // host vectors die at return
    ctx->bvh.nodes = ctx->dNodes; ctx->bvh.slots = ctx->dSlots; ctx->bvh.shade = ctx->dShade;
    ctx->bvh.rootRef = refOf(0);
    ctx->bvh.topN = (int)(nInner < FLX_TOP_NODES ? nInner : FLX_TOP_NODES);
    if (const char *lim = getenv("FLX_TOP_NODES_LIMIT")) { const int v = atoi(lim); if (v >= 0 && v < ctx->bvh.topN) ctx->bvh.topN = v; } // tuning aid
    ctx->nInner = nInner; ctx->nSlots = n_indices; ctx->nTris = n_tris;
    return FLX_OK;
}

