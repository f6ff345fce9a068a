// TEST INFRASTRUCTURE — C entry points of the CPU oracle (ctypes-friendly).  Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
#include "o_wavefront.h"
#include <cstring>
#include <omp.h>

using namespace orc;

namespace orc { extern uint32_t *g_nodeVisitHistogram; } // o_trace.cpp

extern "C" {

void orc_set_node_histogram(uint32_t *hist) { orc::g_nodeVisitHistogram = hist; } // nullptr switches it off
void orc_set_threads(int n) { omp_set_num_threads(n > 0 ? n : 1); }
int orc_max_threads() { return omp_get_max_threads(); }

// Closest-hit for a ray batch.  rays: n x 8 floats (ox,oy,oz,tmax, dx,dy,dz,pad).
// mode: 0 stack traversal (reference default), 1 bit-stack, 2 brute force.
// hits64 (optional): n x flx_hit.  counters (optional): 5 x u64 {inner, leaf, tri, hits, max_stack}.
void orc_trace_closest(const flx_triangle *tris, uint32_t n_tris, const uint32_t *indices, uint32_t n_idx,
                       const flx_node *nodes, uint32_t n_nodes, const float *rays, uint64_t n, int mode,
                       int32_t *hit_i, float *hit_t, flx_hit *hits64, uint64_t *counters)
{
    SceneView s = {tris, indices, nodes, n_tris, n_idx, n_nodes};
    TraceCounters total; memset(&total, 0, sizeof(total));
#pragma omp parallel
    {
        TraceCounters c; memset(&c, 0, sizeof(c));
#pragma omp for schedule(dynamic, 1024)
        for (int64_t k = 0; k < (int64_t)n; k++)
        {
            const float *p = rays + 8 * k;
            Ray r; r.orig = mk3(p[0], p[1], p[2]); r.dir = mk3(p[4], p[5], p[6]);
            Hit h = emptyHit(p[3]);
            if (mode == 0) bvhIntersectStack(r, &h, s, &c);
            else if (mode == 1) bvhIntersectBitstack(r, &h, s, &c);
            else bruteIntersect(r, &h, s);
            if (hit_i) hit_i[k] = h.i;
            if (hit_t) hit_t[k] = h.t;
            if (hits64)
            {
                flx_hit &o = hits64[k]; memset(&o, 0, sizeof(o));
                o.P.x = h.P.x; o.P.y = h.P.y; o.P.z = h.P.z; o.N.x = h.N.x; o.N.y = h.N.y; o.N.z = h.N.z;
                o.uvTex.x = h.uv.x; o.uvTex.y = h.uv.y; o.t = h.t; o.i = h.i; o.areaLightHit = h.areaLightHit; o.matId = h.matId;
            }
        }
#pragma omp critical
        {
            total.inner_visits += c.inner_visits; total.leaf_visits += c.leaf_visits; total.tri_tests += c.tri_tests;
            total.hits += c.hits; if (c.max_stack > total.max_stack) total.max_stack = c.max_stack;
        }
    }
    if (counters)
    {
        counters[0] = total.inner_visits; counters[1] = total.leaf_visits; counters[2] = total.tri_tests;
        counters[3] = total.hits; counters[4] = total.max_stack;
    }
}

// Any-hit with per-ray tmax (rays[3]).  mode 0 stack, 1 bit-stack.
void orc_trace_any(const flx_triangle *tris, uint32_t n_tris, const uint32_t *indices, uint32_t n_idx,
                   const flx_node *nodes, uint32_t n_nodes, const float *rays, uint64_t n, int mode,
                   uint8_t *occluded, uint64_t *counters)
{
    SceneView s = {tris, indices, nodes, n_tris, n_idx, n_nodes};
    TraceCounters total; memset(&total, 0, sizeof(total));
#pragma omp parallel
    {
        TraceCounters c; memset(&c, 0, sizeof(c));
#pragma omp for schedule(dynamic, 1024)
        for (int64_t k = 0; k < (int64_t)n; k++)
        {
            const float *p = rays + 8 * k;
            Ray r; r.orig = mk3(p[0], p[1], p[2]); r.dir = mk3(p[4], p[5], p[6]);
            occluded[k] = (mode == 1 ? bvhOccludedBitstack(r, p[3], s, &c) : bvhOccludedStack(r, p[3], s, &c)) ? 1 : 0;
        }
#pragma omp critical
        { total.inner_visits += c.inner_visits; total.leaf_visits += c.leaf_visits; total.tri_tests += c.tri_tests; total.hits += c.hits; }
    }
    if (counters) { counters[0] = total.inner_visits; counters[1] = total.leaf_visits; counters[2] = total.tri_tests; counters[3] = total.hits; counters[4] = 0; }
}

// ---- wavefront handle -------------------------------------------------------------------------
void *orc_wf_create(uint32_t numTasks, uint32_t width, uint32_t height)
{
    Wavefront *w = new Wavefront();
    memset(&w->params, 0, sizeof(w->params));
    memset(&w->geo, 0, sizeof(w->geo));
    memset(&w->shade, 0, sizeof(w->shade));
    w->params.width = width; w->params.height = height;
    w->allocate(numTasks, width * height);
    return w;
}
void orc_wf_destroy(void *h) { delete (Wavefront *)h; }

// All pointers are borrowed: the caller keeps the arrays alive for the life of the handle.
void orc_wf_set_scene(void *h, const flx_triangle *tris, uint32_t n_tris, const uint32_t *indices, uint32_t n_idx,
                      const flx_node *nodes, uint32_t n_nodes, const flx_material *materials,
                      const flx_tex_desc *textures, const uint8_t *texData)
{
    Wavefront *w = (Wavefront *)h;
    w->geo.tris = tris; w->geo.n_tris = n_tris; w->geo.indices = indices; w->geo.n_indices = n_idx;
    w->geo.nodes = nodes; w->geo.n_nodes = n_nodes;
    w->shade.tris = tris; w->shade.materials = materials; w->shade.textures = textures; w->shade.texData = texData;
}
void orc_wf_set_envmap(void *h, const float *rgba, int width, int height, const float *prob, const int *alias, const float *pdf)
{
    Wavefront *w = (Wavefront *)h;
    w->shade.envRGBA = rgba; w->shade.envW = width; w->shade.envH = height;
    w->shade.probTable = prob; w->shade.aliasTable = alias; w->shade.pdfTable = pdf;
}
void orc_wf_set_params(void *h, const flx_render_params *p) { ((Wavefront *)h)->params = *p; }
void orc_wf_set_options(void *h, int useBitstack, int denoiserAOV, uint32_t seedOffset)
{
    Wavefront *w = (Wavefront *)h; w->useBitstack = useBitstack != 0; w->denoiserAOV = denoiserAOV != 0; w->seedOffset = seedOffset;
}
void orc_wf_set_drop_mixed(void *h, int drop) { ((Wavefront *)h)->dropMixed = drop != 0; }

void orc_wf_reset(void *h) { ((Wavefront *)h)->reset(); }
void orc_wf_raygen(void *h) { ((Wavefront *)h)->raygen(); }
void orc_wf_logic(void *h, int firstIteration) { ((Wavefront *)h)->logic(firstIteration != 0); }
void orc_wf_materials(void *h) { ((Wavefront *)h)->materials(); }
void orc_wf_ext(void *h) { ((Wavefront *)h)->extRays(); }
void orc_wf_shadow(void *h) { ((Wavefront *)h)->shadowRays(); }
void orc_wf_pick(void *h, float ndcX, float ndcY, flx_hit *out)
{
    const Hit hit = ((Wavefront *)h)->pick(ndcX, ndcY);
    memset(out, 0, sizeof(*out));
    out->P.x = hit.P.x; out->P.y = hit.P.y; out->P.z = hit.P.z; out->N.x = hit.N.x; out->N.y = hit.N.y; out->N.z = hit.N.z;
    out->uvTex.x = hit.uv.x; out->uvTex.y = hit.uv.y; out->t = hit.t; out->i = hit.i; out->areaLightHit = hit.areaLightHit; out->matId = hit.matId;
}
void orc_mk_reset(void *h) { ((Wavefront *)h)->mkReset(); }
void orc_mk_raygen(void *h) { ((Wavefront *)h)->mkRaygen(); }
void orc_mk_next_vertex(void *h) { ((Wavefront *)h)->mkNextVertex(); }
void orc_mk_sample_bsdf(void *h) { ((Wavefront *)h)->mkSampleBsdf(); }
void orc_mk_splat(void *h) { ((Wavefront *)h)->mkSplat(); }
void orc_mk_splat_preview(void *h) { ((Wavefront *)h)->mkSplatPreview(); }
void orc_mk_render_single(void *h, uint32_t spp) { ((Wavefront *)h)->mkRenderSingle(spp); }
flx_render_stats *orc_mk_stats(void *h) { return &((Wavefront *)h)->stats; }
void orc_wf_clear_queues(void *h) { ((Wavefront *)h)->clearQueues(); }
void orc_wf_postprocess(void *h, float *outRGBA) { ((Wavefront *)h)->postprocess(outRGBA); }
void orc_wf_update_pixel_index(void *h, uint32_t n) { ((Wavefront *)h)->updatePixelIndex(n); }
void orc_wf_set_pixel_index(void *h, uint32_t v) { ((Wavefront *)h)->currPixelIdx = v; }
uint32_t orc_wf_get_pixel_index(void *h) { return ((Wavefront *)h)->currPixelIdx; }

uint32_t *orc_wf_state(void *h) { return ((Wavefront *)h)->state.data(); }      // 64 x N words
uint32_t *orc_wf_queue(void *h, int q) { return ((Wavefront *)h)->queues[q].data(); }
flx_queue_counters *orc_wf_counters(void *h) { return &((Wavefront *)h)->cnt; }
float *orc_wf_pixels(void *h) { return ((Wavefront *)h)->pixels.data(); }
uint32_t *orc_wf_spp(void *h) { return ((Wavefront *)h)->spp.data(); }
float *orc_wf_denoiser_albedo(void *h) { return ((Wavefront *)h)->denoiserAlbedo.data(); }
float *orc_wf_denoiser_normal(void *h) { return ((Wavefront *)h)->denoiserNormal.data(); }
void orc_wf_trace_counters(void *h, uint64_t *ext5, uint64_t *shadow5)
{
    Wavefront *w = (Wavefront *)h;
    const TraceCounters *c[2] = {&w->extCounters, &w->shadowCounters};
    uint64_t *o[2] = {ext5, shadow5};
    for (int k = 0; k < 2; k++)
        if (o[k]) { o[k][0] = c[k]->inner_visits; o[k][1] = c[k]->leaf_visits; o[k][2] = c[k]->tri_tests; o[k][3] = c[k]->hits; o[k][4] = c[k]->max_stack; }
}

// Fixed-budget wavefront render in the reference's benchmark order
// (Tracer::runBenchmarkFromFile, src/tracer.cpp:633-645 reset+clear, :725-729 priming raygen+ext,
// :737-778 logic -> raygen -> materials -> getCounters -> ext -> shadow -> clear -> updatePixelIndex).
// Stops when the splatted-sample total reaches w*h*maxSpp (maxSpp > 0) or after maxIterations.
// stats4 = {primary, extension, shadow, samples} (64-bit totals, same definitions as tracer.cpp:761-765).
uint64_t orc_wf_render(void *h, uint64_t maxIterations, int doReset, uint64_t *stats4)
{
    Wavefront *w = (Wavefront *)h;
    const uint64_t target = (uint64_t)w->params.width * w->params.height * w->params.maxSpp;
    uint64_t sampleCount = 0, iter = 0;
    uint64_t st[4] = {0, 0, 0, 0};
    if (doReset)
    {
        w->reset(); w->clearQueues(); w->currPixelIdx = 0;
        w->raygen(); w->extRays();
    }
    while (iter < maxIterations && (target == 0 || sampleCount < target))
    {
        w->logic(false);
        w->raygen();
        w->materials();
        const flx_queue_counters cnt = w->cnt;
        w->extRays();
        w->shadowRays();
        w->clearQueues();
        st[0] += cnt.raygenQueue; st[1] += cnt.extensionQueue; st[2] += cnt.shadowQueue;
        st[3] += (iter > 0) ? cnt.raygenQueue : 0;
        sampleCount += cnt.splattedSamples;
        w->updatePixelIndex(cnt.raygenQueue);
        iter++;
    }
    if (stats4) memcpy(stats4, st, sizeof(st));
    return iter;
}

} // extern "C"
