// Parallel hierarchy build on the GPU: a linear BVH (Morton order + Karras' radix tree) emitted
// in the REFERENCE's node format, so that it goes through the same upload / repack / traversal path — and through
// the CPU oracle — as a hierarchy from the reference's single-threaded SBVH builder (src/sbvh.cpp, src/bvh.cpp),
// which needs minutes for the 10 M-triangle microbench scene.
//
// Output contract = what BVH hands to CLContext::uploadSceneData (src/clcontext.cpp:540-584): `Node[]`
// (src/bvhnode.hpp:50-59: box, parent, rightChild | iStart, nPrims) in depth-first order with the left child at
// index + 1, and the index list the leaves point into (src/bvh.hpp).
//
// Steps (all kernels are grid-stride, launched on the caller-independent default stream of the chosen device):
//   1. k_scene_bounds      triangle boxes -> scene box (ordered-int atomics)
//   2. k_morton            63-bit Morton code of each triangle's box centre (21 bits per axis)
//   3. radix sort          (code, triangle) pairs — cub::DeviceRadixSort, the one library call of the builder
//   4. k_radix_tree        Karras 2012: internal node i of n-1 finds its leaf range and split from the common prefixes of
//                          neighbouring codes (ties broken by position, so duplicate codes are fine)
//   5. k_refit             bottom-up: boxes, and the number of nodes that stay alive once every subtree holding
//                          <= maxLeaf triangles has been collapsed into ONE leaf (second arrival continues upward)
//   6. k_emit              depth-first index of every live node by walking up to the root (+1 per left step,
//                          +1 + live(left sibling) per right step), node record written in place
// The traversal order, and with it the winner among equal-t hits, depends on the tree: a hierarchy from this builder
// and one from the reference's builder agree on t bit for bit and on the triangle except for exact ties.
#include <cub/device/device_radix_sort.cuh>
#include "flx_internal.h"
#include "flx_bvh.cuh" // FLX_STACK_SIZE

#define FLX_BUILD_THREADS 256

namespace {

struct BuildNode // working form of one node of the binary radix tree; internal i in [0, n-1), leaf j stored at n-1+j
{
    float bmin[3], bmax[3];
    int parent;        // index into the working array, -1 for the root
    int left, right;   // children (working indices); leaves: -1
    int first, last;   // covered range of sorted triangles
    int live;          // live nodes in this subtree after leaf collapsing (1 for a collapsed subtree)
};

__device__ __forceinline__ int orderedInt(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float orderedFloat(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__device__ __forceinline__ void triBox(const flx_triangle &t, float lo[3], float hi[3])
{
    const float *a = &t.v0.p.x, *b = &t.v1.p.x, *c = &t.v2.p.x;
    for (int k = 0; k < 3; k++) { lo[k] = fminf(a[k], fminf(b[k], c[k])); hi[k] = fmaxf(a[k], fmaxf(b[k], c[k])); }
}

__global__ void k_scene_bounds(const flx_triangle *tris, uint32_t n, int *box6)
{
    int lo[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff}, hi[3] = {(int)0x80000000, (int)0x80000000, (int)0x80000000};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        float a[3], b[3];
        triBox(tris[i], a, b);
        for (int k = 0; k < 3; k++) { lo[k] = min(lo[k], orderedInt(a[k])); hi[k] = max(hi[k], orderedInt(b[k])); }
    }
    for (int k = 0; k < 3; k++)
    {
        lo[k] = __reduce_min_sync(0xffffffffu, lo[k]); hi[k] = __reduce_max_sync(0xffffffffu, hi[k]);
        if ((threadIdx.x & 31) == 0) { atomicMin(box6 + k, lo[k]); atomicMax(box6 + 3 + k, hi[k]); }
    }
}

__device__ __forceinline__ unsigned long long spread21(unsigned long long x) // 21 bits -> every third bit
{
    x &= 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void k_morton(const flx_triangle *tris, uint32_t n, const int *box6, unsigned long long *codes, uint32_t *ids)
{
    float lo[3], inv[3];
    for (int k = 0; k < 3; k++)
    {
        lo[k] = orderedFloat(box6[k]);
        const float ext = orderedFloat(box6[3 + k]) - lo[k];
        inv[k] = ext > 0.0f ? 2097152.0f / ext : 0.0f; // 2^21 cells per axis
    }
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        float a[3], b[3];
        triBox(tris[i], a, b);
        unsigned long long code = 0ull;
        for (int k = 0; k < 3; k++)
        {
            const float c = (0.5f * (a[k] + b[k]) - lo[k]) * inv[k];
            const unsigned long long q = (unsigned long long)fminf(fmaxf(c, 0.0f), 2097151.0f);
            code |= spread21(q) << k;
        }
        codes[i] = code;
        ids[i] = i;
    }
}

// length of the common prefix of the keys at sorted positions i and j (position appended to break ties); -1 outside
__device__ __forceinline__ int delta(const unsigned long long *codes, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = codes[i], b = codes[j];
    if (a != b) return __clzll((long long)(a ^ b));
    return 64 + __clz(i ^ j);
}

// Karras, "Maximizing Parallelism in the Construction of BVHs, Octrees, and k-d Trees" (HPG 2012), section 4
__global__ void k_radix_tree(const unsigned long long *codes, int n, BuildNode *nodes)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n - 1; i += gridDim.x * blockDim.x)
    {
        const int d = delta(codes, n, i, i + 1) - delta(codes, n, i, i - 1) >= 0 ? 1 : -1;
        const int dmin = delta(codes, n, i, i - d);
        int lmax = 2;
        while (delta(codes, n, i, i + lmax * d) > dmin) lmax <<= 1;
        int l = 0;
        for (int t = lmax >> 1; t >= 1; t >>= 1)
            if (delta(codes, n, i, i + (l + t) * d) > dmin) l += t;
        const int j = i + l * d;
        const int dnode = delta(codes, n, i, j);
        int s = 0;
        for (int t = (l + 1) >> 1;; t = (t + 1) >> 1)
        {
            if (delta(codes, n, i, i + (s + t) * d) > dnode) s += t;
            if (t == 1) break;
        }
        const int gamma = i + s * d + min(d, 0);
        const int first = min(i, j), last = max(i, j);
        const int left = (first == gamma) ? (n - 1 + gamma) : gamma;
        const int right = (last == gamma + 1) ? (n - 1 + gamma + 1) : (gamma + 1);
        BuildNode &nd = nodes[i];
        nd.left = left; nd.right = right; nd.first = first; nd.last = last;
        nodes[left].parent = i;
        nodes[right].parent = i;
        if (i == 0) nd.parent = -1;
    }
}

__global__ void k_leaf_boxes(const flx_triangle *tris, const uint32_t *ids, int n, BuildNode *nodes)
{
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x)
    {
        BuildNode &nd = nodes[n - 1 + j];
        triBox(tris[ids[j]], nd.bmin, nd.bmax);
        nd.left = nd.right = -1; nd.first = nd.last = j; nd.live = 1;
        if (n == 1) nd.parent = -1;
    }
}

__global__ void k_refit(int n, BuildNode *nodes, int *arrivals, int maxLeaf)
{
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x)
    {
        int cur = nodes[n - 1 + j].parent;
        while (cur >= 0)
        {
            __threadfence();
            if (atomicAdd(arrivals + cur, 1) == 0) break; // the sibling subtree is not done yet
            BuildNode &nd = nodes[cur];
            // the children were finished by other threads (possibly on other SMs): read them past L1
            const volatile BuildNode &a = nodes[nd.left], &b = nodes[nd.right];
            for (int k = 0; k < 3; k++) { nd.bmin[k] = fminf(a.bmin[k], b.bmin[k]); nd.bmax[k] = fmaxf(a.bmax[k], b.bmax[k]); }
            nd.live = (nd.last - nd.first + 1 <= maxLeaf) ? 1 : 1 + a.live + b.live;
            cur = nd.parent;
        }
    }
}

// Depth-first position of every live node and its record in the reference's layout.  A node is live when no proper
// ancestor has been collapsed into a leaf, i.e. when its parent covers more than maxLeaf triangles.
__global__ void k_emit(int n, const BuildNode *nodes, int maxLeaf, flx_node *out, int *maxDepth)
{
    const int total = 2 * n - 1;
    int deepest = 0;
    for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < total; x += gridDim.x * blockDim.x)
    {
        const BuildNode &nd = nodes[x];
        if (nd.parent >= 0 && nodes[nd.parent].last - nodes[nd.parent].first + 1 <= maxLeaf) continue; // swallowed by a collapsed leaf
        int idx = 0, depth = 0, parentIdx = -1;
        for (int c = x, p = nd.parent; p >= 0; c = p, p = nodes[p].parent)
        {
            const int step = (nodes[p].left == c) ? 1 : 1 + nodes[nodes[p].left].live;
            idx += step;
            if (c == x) parentIdx = -step; // remembered relative to this node, fixed up below
            depth++;
        }
        const int count = nd.last - nd.first + 1;
        const bool leaf = count <= maxLeaf;
        flx_node o;
        o.bmin.x = nd.bmin[0]; o.bmin.y = nd.bmin[1]; o.bmin.z = nd.bmin[2]; o.bmin.w = 0.0f;
        o.bmax.x = nd.bmax[0]; o.bmax.y = nd.bmax[1]; o.bmax.z = nd.bmax[2]; o.bmax.w = 0.0f;
        o.parent = nd.parent >= 0 ? idx + parentIdx : -1;
        o.iStartOrRight = leaf ? (uint32_t)nd.first : (uint32_t)(idx + 1 + nodes[nd.left].live);
        o.nPrims = leaf ? (uint8_t)count : 0;
        for (int k = 0; k < 7; k++) o._pad[k] = 0;
        out[idx] = o;
        deepest = max(deepest, depth);
    }
    deepest = __reduce_max_sync(0xffffffffu, deepest);
    if ((threadIdx.x & 31) == 0) atomicMax(maxDepth, deepest);
}

struct DeviceBuffers
{
    void *p[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    ~DeviceBuffers() { for (void *q : p) if (q) cudaFree(q); }
};

} // namespace

#define BUILD_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) \
    return flx_fail(nullptr, FLX_ERR_CUDA, std::string("flx_build_lbvh: " #call ": ") + cudaGetErrorString(e__)); } while (0)

extern "C" int flx_build_lbvh(int device, const flx_triangle *tris, size_t n_tris, uint32_t max_leaf,
                              flx_node *nodes_out, size_t nodes_capacity, size_t *n_nodes_out, uint32_t *indices_out, uint32_t *depth_out)
{
    if (!tris || !nodes_out || !n_nodes_out || !indices_out) return flx_fail(nullptr, FLX_ERR_ARG, "flx_build_lbvh: null argument");
    if (n_tris == 0 || n_tris > 0x3fffffffu) return flx_fail(nullptr, FLX_ERR_ARG, "flx_build_lbvh: triangle count out of range");
    if (max_leaf < 1 || max_leaf > 255) return flx_fail(nullptr, FLX_ERR_ARG, "flx_build_lbvh: max_leaf must be 1..255 (Node::nPrims is 8 bits)");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        return flx_fail(nullptr, FLX_ERR_NO_DEVICE, "flx_build_lbvh: no CUDA device (this builder has no CPU fallback)");
    BUILD_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    BUILD_CUDA(cudaGetDeviceProperties(&prop, device));
    const int n = (int)n_tris, total = 2 * n - 1;
    const int grid = prop.multiProcessorCount * 8;

    DeviceBuffers d;
    flx_triangle *dTris; unsigned long long *dCodes, *dCodesSorted; uint32_t *dIds, *dIdsSorted; BuildNode *dNodes; int *dMisc; flx_node *dOut;
    BUILD_CUDA(cudaMalloc(&d.p[0], n_tris * sizeof(flx_triangle))); dTris = (flx_triangle *)d.p[0];
    BUILD_CUDA(cudaMalloc(&d.p[1], 2 * n_tris * sizeof(unsigned long long))); dCodes = (unsigned long long *)d.p[1]; dCodesSorted = dCodes + n_tris;
    BUILD_CUDA(cudaMalloc(&d.p[2], 2 * n_tris * sizeof(uint32_t))); dIds = (uint32_t *)d.p[2]; dIdsSorted = dIds + n_tris;
    BUILD_CUDA(cudaMalloc(&d.p[3], (size_t)total * sizeof(BuildNode))); dNodes = (BuildNode *)d.p[3];
    BUILD_CUDA(cudaMalloc(&d.p[4], ((size_t)n + 16) * sizeof(int))); dMisc = (int *)d.p[4]; // [0..5] scene box, [6] depth, [16..] arrival flags
    BUILD_CUDA(cudaMemcpy(dTris, tris, n_tris * sizeof(flx_triangle), cudaMemcpyHostToDevice));
    const int init[7] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000, 0};
    BUILD_CUDA(cudaMemcpy(dMisc, init, sizeof(init), cudaMemcpyHostToDevice));
    BUILD_CUDA(cudaMemset(dMisc + 16, 0, (size_t)n * sizeof(int)));

    k_scene_bounds<<<grid, FLX_BUILD_THREADS>>>(dTris, (uint32_t)n, dMisc);
    k_morton<<<grid, FLX_BUILD_THREADS>>>(dTris, (uint32_t)n, dMisc, dCodes, dIds);
    size_t tmpBytes = 0;
    BUILD_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, dCodes, dCodesSorted, dIds, dIdsSorted, n, 0, 63));
    BUILD_CUDA(cudaMalloc(&d.p[5], tmpBytes ? tmpBytes : 16));
    BUILD_CUDA(cub::DeviceRadixSort::SortPairs(d.p[5], tmpBytes, dCodes, dCodesSorted, dIds, dIdsSorted, n, 0, 63));
    k_leaf_boxes<<<grid, FLX_BUILD_THREADS>>>(dTris, dIdsSorted, n, dNodes);
    if (n > 1) k_radix_tree<<<grid, FLX_BUILD_THREADS>>>(dCodesSorted, n, dNodes);
    if (n > 1) k_refit<<<grid, FLX_BUILD_THREADS>>>(n, dNodes, dMisc + 16, (int)max_leaf);
    BUILD_CUDA(cudaGetLastError());

    int liveRoot = 1;
    BUILD_CUDA(cudaMemcpy(&liveRoot, &dNodes[0].live, sizeof(int), cudaMemcpyDeviceToHost)); // node 0 is the root (a leaf when n == 1)
    *n_nodes_out = (size_t)liveRoot;
    if ((size_t)liveRoot > nodes_capacity)
        return flx_fail(nullptr, FLX_ERR_ARG, "flx_build_lbvh: nodes_capacity too small (2 * n_tris - 1 always suffices)");
    BUILD_CUDA(cudaMalloc(&d.p[6], (size_t)liveRoot * sizeof(flx_node))); dOut = (flx_node *)d.p[6];
    k_emit<<<grid, FLX_BUILD_THREADS>>>(n, dNodes, (int)max_leaf, dOut, dMisc + 6);
    BUILD_CUDA(cudaGetLastError());
    BUILD_CUDA(cudaMemcpy(nodes_out, dOut, (size_t)liveRoot * sizeof(flx_node), cudaMemcpyDeviceToHost));
    BUILD_CUDA(cudaMemcpy(indices_out, dIdsSorted, n_tris * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    int depth = 0;
    BUILD_CUDA(cudaMemcpy(&depth, dMisc + 6, sizeof(int), cudaMemcpyDeviceToHost));
    if (depth_out) *depth_out = (uint32_t)depth;
    // k_emit counts edges below the root; the traversal's stack holds one entry per LEVEL.  A radix tree over 63 Morton
    // bits + the index tie-break can reach ~95 levels on inputs with many (near-)coincident centroids — instanced or
    // duplicated geometry — which flx_upload_scene would refuse: say so here, where the caller can still react.
    if (depth + 1 > FLX_STACK_SIZE)
        return flx_fail(nullptr, FLX_ERR_ARG, "flx_build_lbvh: the hierarchy is " + std::to_string(depth + 1) + " levels deep (coincident centroids?) and the "
                        "traversal holds " + std::to_string(FLX_STACK_SIZE) + " (flx_upload_scene would refuse it): build with a larger max_leaf");
    return FLX_OK;
}
