/* fluctus_cuda.h — C-ABI of the B200 (sm_100a) backend for fluctus's wavefront path tracer.
 *
 * One opaque context per GPU replaces the reference's `class CLContext`
 * (reference src/clcontext.hpp:26-215).  Every function below names the CLContext method it
 * stands in for; a C++ wrapper with the reference's method names lives in
 * fluctus_b200/host/cuda_context.hpp and the binding recipe is in INTEGRATION.md.
 *
 * Conventions (mirroring the reference's in-order OpenCL queue, ext/CLT/src/utils.cpp:312):
 *   - every call returns 0 on success or a negative flx_status; flx_last_error() gives the text
 *     (the reference throws std::runtime_error from CLContext::verify, clcontext.cpp:952-957);
 *   - all work is enqueued on ONE CUDA stream per context, in call order; the enqueue calls
 *     return immediately; flx_finish() is the only wait (CLContext::finishQueue);
 *   - host pointers passed in are copied before the call returns (no borrowed host memory);
 *   - a context is not thread-safe; use one per GPU.
 * Plain C: pointers and sizes only.
 */
#ifndef FLUCTUS_CUDA_H
#define FLUCTUS_CUDA_H

#include <stddef.h>
#include <stdint.h>
#include "fluctus_abi.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct flx_ctx flx_ctx;

enum flx_status {
    FLX_OK = 0,
    FLX_ERR_CUDA = -1,       /* a CUDA runtime call failed */
    FLX_ERR_ARG = -2,        /* bad argument */
    FLX_ERR_STATE = -3,      /* call out of order (e.g. stage before scene upload) */
    FLX_ERR_NO_DEVICE = -4   /* no sm_100 device visible: there is no CPU fallback */
};

/* which buffer flx_read_pixels / flx_device_ptr address */
enum flx_buffer {
    FLX_BUF_PIXELS = 0,          /* raw accumulation: float4 (sum R, sum G, sum B, #samples) per pixel */
    FLX_BUF_PREVIEW = 1,         /* tonemapped float4 (what the reference writes into the GL PBO) */
    FLX_BUF_DENOISER_ALBEDO = 2,
    FLX_BUF_DENOISER_NORMAL = 3,
    FLX_BUF_SAMPLES_PER_PIXEL = 4,
    FLX_BUF_DENOISER_ALBEDO_OUT = 5, /* normalised copies written by flx_postprocess (mk_postprocess.cl:49-55) */
    FLX_BUF_DENOISER_NORMAL_OUT = 6
};

/* ---- life cycle: CLContext::CLContext + setup (clcontext.cpp:18-69), NUM_TASKS = wfBufferSize ---- */
int flx_create(flx_ctx **out, int device, uint32_t num_tasks, void *cuda_stream /* NULL = own stream */);
void flx_destroy(flx_ctx *ctx);
const char *flx_last_error(flx_ctx *ctx); /* ctx may be NULL: last creation error */
uint32_t flx_num_tasks(flx_ctx *ctx);     /* CLContext::getNumTasks */
void *flx_stream(flx_ctx *ctx);           /* the context's cudaStream_t (the reference's cmdQueue) */

/* ---- scene: CLContext::uploadSceneData + packTextures (clcontext.cpp:540-629) -------------------
 * Takes the reference's host arrays verbatim (RTTriangle 160 B, Node 48 B, Material 96 B, RGBA8
 * texels back to back with byte offsets in the descriptors) and re-packs them on the device. */
int flx_upload_scene(flx_ctx *ctx, const flx_triangle *tris, size_t n_tris, const uint32_t *indices, size_t n_indices,
                     const flx_node *nodes, size_t n_nodes, const flx_material *materials, size_t n_materials,
                     const flx_tex_desc *textures, size_t n_textures, const uint8_t *texels, size_t texel_bytes);

/* CLContext::createEnvMap (clcontext.cpp:485-529): RGB float triplets + the three alias-method tables */
int flx_upload_envmap(flx_ctx *ctx, const float *rgb, int width, int height, const float *prob_table,
                      const int32_t *alias_table, const float *pdf_table);

/* CLContext::setupPixelStorage (clcontext.cpp:337-402) without the GL PBOs */
int flx_resize(flx_ctx *ctx, uint32_t width, uint32_t height);

/* CLContext::updateParams (clcontext.cpp:721-725); also selects the kernel specialisations the
 * reference obtains by re-JIT with -D flags (kernel_impl.hpp:51-63): no recompilation here. */
int flx_update_params(flx_ctx *ctx, const flx_render_params *params);

/* Options with no RenderParams field: denoiser AOVs (-DUSE_OPTIX_DENOISER accumulation only) and the
 * per-rank RNG offset used when samples are sharded over GPUs (seed = gid + seed_offset). */
int flx_set_options(flx_ctx *ctx, int denoiser_aov, uint32_t seed_offset);

/* Preview / single-spp sharding over GPUs (no reference counterpart): with world > 1 this context
 * generates camera paths only for the image rows y % world == rank, so every rank refines its own interleaved
 * slice of the frame and the slices are combined with one all-gather (or the same all-reduce: foreign rows stay
 * zero).  world = 1 (default) restores the reference's pixel order (wf_raygen.cl:25). */
int flx_set_pixel_sharding(flx_ctx *ctx, uint32_t rank, uint32_t world);

/* CLContext::resetPixelIndex / updatePixelIndex (clcontext.cpp:917-927) */
int flx_reset_pixel_index(flx_ctx *ctx);
int flx_set_pixel_index(flx_ctx *ctx, uint32_t value);
int flx_update_pixel_index(flx_ctx *ctx, uint32_t num_pixels, uint32_t num_new_paths);

/* ---- wavefront stages: CLContext::enqueueWf*Kernel (clcontext.cpp:783-873) ---------------------- */
int flx_wf_reset(flx_ctx *ctx);                     /* enqueueWfResetKernel    -> wf_reset.cl */
int flx_wf_raygen(flx_ctx *ctx);                    /* enqueueWfRaygenKernel   -> wf_raygen.cl */
int flx_wf_logic(flx_ctx *ctx, int first_iteration);/* enqueueWfLogicKernel    -> wf_logic.cl */
int flx_wf_materials(flx_ctx *ctx);                 /* enqueueWfMaterialKernels-> wf_mat_*.cl */
int flx_wf_ext(flx_ctx *ctx);                       /* enqueueWfExtRayKernel   -> wf_extrays.cl */
int flx_wf_shadow(flx_ctx *ctx);                    /* enqueueWfShadowRayKernel-> wf_shadowrays.cl */
int flx_wf_clear_queues(flx_ctx *ctx);              /* enqueueClearWfQueues */
int flx_get_counters(flx_ctx *ctx, flx_queue_counters *out); /* enqueueGetCounters: async, valid after flx_finish */
int flx_postprocess(flx_ctx *ctx);                  /* enqueuePostprocessKernel -> mk_postprocess.cl */
int flx_finish(flx_ctx *ctx);                       /* finishQueue */

/* Headless render entry (new): `iterations` passes of the reference's benchmark inner loop
 * (Tracer::runBenchmarkFromFile, tracer.cpp:737-778: logic -> raygen -> materials -> counters ->
 * ext -> shadow -> clear -> pixel-index advance) replayed from a CUDA graph with the counter
 * bookkeeping done on the device.  do_reset != 0 first runs reset + clear + priming raygen + ext
 * (tracer.cpp:633-645, :725-729).  Asynchronous like the stage calls. */
int flx_render(flx_ctx *ctx, uint32_t iterations, int do_reset);
/* Totals accumulated on the device by flx_render since the last reset: {primary, extension,
 * shadow, samples, splatted, iterations} (definitions of tracer.cpp:761-765).  Synchronises. */
int flx_get_render_totals(flx_ctx *ctx, uint64_t totals[6]);

/* ---- microkernel ("MK") renderer: one task per pixel, exact sample counts, deterministic images ------------
 * CLContext::enqueueResetKernel / RayGenKernel / NextVertexKernel / BsdfSampleKernel / SplatKernel /
 * SplatPreviewKernel (clcontext.cpp:727-768) -> mk_reset.cl, mk_raygen.cl, mk_next_vertex.cl,
 * mk_sample_bsdf.cl, mk_splat.cl, mk_splat_preview.cl.  Needs NUM_TASKS >= width * height to cover the
 * image (tasks beyond min(w*h, NUM_TASKS) idle, as in the reference).  Asynchronous, in order. */
int flx_mk_reset(flx_ctx *ctx);          /* also zeroes the device RenderStats */
int flx_mk_raygen(flx_ctx *ctx);
int flx_mk_next_vertex(flx_ctx *ctx);
int flx_mk_sample_bsdf(flx_ctx *ctx);
int flx_mk_splat(flx_ctx *ctx);
int flx_mk_splat_preview(flx_ctx *ctx);
/* Device RenderStats {primary, extension, shadow rays, samples} the MK kernels maintain
 * (CLContext::fetchStatsAsync, clcontext.cpp:660-664).  Synchronises. */
int flx_mk_get_stats(flx_ctx *ctx, flx_render_stats *out);
/* Tracer::renderSingle (tracer.cpp:108-182): `spp` times raygen -> (nextVertex -> sampleBsdf) x
 * (maxBounces + 1) -> splat, replayed from a CUDA graph; do_reset != 0 runs flx_mk_reset first.
 * Fails when maxBounces == 0 or Russian roulette is on (the reference switches it off, :118-122). */
int flx_mk_render(flx_ctx *ctx, uint32_t spp, int do_reset);

/* ---- results: CLContext::saveImage reads (clcontext.cpp:404-483), pickSingle (:934-949) ---------- */
int flx_read_pixels(flx_ctx *ctx, int which /* flx_buffer */, void *host_out, size_t bytes); /* synchronises */
void *flx_device_ptr(flx_ctx *ctx, int which /* flx_buffer */); /* for the per-frame all-reduce */
int flx_pick(flx_ctx *ctx, float ndc_x, float ndc_y, flx_hit *out);

/* ---- traversal variant: Settings::clUseBitstack / -DUSE_BITSTACK (settings.cpp:27, clcontext.cpp:145-155) -------
 * 0 (default) = the 64-entry stack of bvh.cl:234-373; non-zero = the stackless bit-trail traversal of bvh.cl:12-230
 * (two 32-bit trails + parent links).  Same visit order and box tests, hence the same hits bit for bit; applies to
 * the wavefront ray stages, flx_render and the ray-batch entry points (the microkernel stages, flx_pick and the
 * visit-counting variants keep the stack).  Fails for hierarchies deeper than 31 levels (the reference's variant
 * silently breaks there); flx_upload_scene refuses such a hierarchy while the switch is on. */
int flx_set_traversal(flx_ctx *ctx, int use_bitstack);

/* ---- multi-GPU frame combine inside one process (new; SURVEY.md §8e — the reference is single-device) ---------
 * ctxs[0..n): contexts on different GPUs rendering sample shards of the same frame (same resolution, seeds offset
 * with flx_set_options).  Drains their streams, then sums the accumulation buffers (and the denoiser AOV sums when
 * enabled) element-wise in rank order: afterwards ctxs[0] holds the total — run flx_postprocess / flx_read_pixels
 * on it — and the other buffers are zero, so all contexts may go on accumulating and be combined again.  The sum is
 * a kernel over NVLink peer memory (GPU d combines rows slice d from all N buffers); without peer access the frames
 * are staged through the root.  With one process per GPU use flx_allreduce_pixels below.  samplesPerPixel (the
 * CHECK_SPP budget) stays per context.  Synchronises. */
int flx_reduce_frames(flx_ctx *const *ctxs, int n);

/* ---- multi-GPU frame combine with ONE PROCESS PER GPU: NCCL over NVLink 5 / NVSwitch (new; SURVEY.md §8b/§8e — the
 * reference has no collective call site) -----------------------------------------------------------------------
 * Every rank renders its sample shard of the frame into its own context (same resolution and parameters, seeds offset
 * with flx_set_options) and once per frame the accumulation buffers are summed:
 *     rank 0:     flx_comm_unique_id(id)        -> hand the 128 bytes to the other ranks (file, socket, MPI, env ...)
 *     every rank: flx_comm_init(ctx, rank, world, id)              (collective: returns when all ranks have joined)
 *     per frame:  flx_render(...); flx_allreduce_pixels(ctx)       every rank ends with the total, or
 *                 flx_render(...); flx_reduce_pixels(ctx, 0)       only rank 0 does (the other buffers are unchanged:
 *                                                                   clear them before they go on accumulating)
 *                 rank 0: flx_postprocess / flx_read_pixels
 * The collective is ONE grouped NCCL call over float4[w*h] (+ the two denoiser AOV sums when enabled), enqueued on the
 * context's stream after the render — nothing synchronises; flx_read_pixels / flx_finish do.  samplesPerPixel
 * (CHECK_SPP's per-pixel budget) is per rank and is not combined.  NCCL is not linked: libnccl.so.2 is dlopen'ed at the
 * first call, the copy already loaded in the process if there is one (under Python, torch's), and a missing library is
 * an FLX_ERR_STATE with a message, not a load failure of this library.  flx_destroy destroys the communicator.
 * flx_comm_info: rank, world (0 = no communicator) and the NCCL version in use (0 = not loaded); any may be NULL. */
#define FLX_COMM_ID_BYTES 128
int flx_comm_unique_id(void *id128);
int flx_comm_init(flx_ctx *ctx, int rank, int world, const void *id128);
int flx_comm_destroy(flx_ctx *ctx);
int flx_comm_info(flx_ctx *ctx, int *rank, int *world, int *nccl_version);
int flx_allreduce_pixels(flx_ctx *ctx);
int flx_reduce_pixels(flx_ctx *ctx, int root);

/* ---- parallel hierarchy build (new) --------------------------------------------------------
 * Linear BVH (63-bit Morton order + Karras' radix tree, subtrees of <= max_leaf triangles collapsed into one leaf)
 * built on GPU `device` and returned in the reference's own format — Node[] in depth-first order with the left
 * child at index + 1 (src/bvhnode.hpp:50-59) and the index list its leaves point into — i.e. a drop-in for the
 * arrays BVH / SBVH (src/bvh.cpp, src/sbvh.cpp) hand to CLContext::uploadSceneData.  nodes_capacity >= 2*n_tris - 1
 * always suffices; *n_nodes_out receives the number written; indices_out has n_tris entries; *depth_out (optional)
 * the depth of the tree in edges below the root.  A tree of more than 64 LEVELS — possible with many coincident
 * centroids: Morton bits + index tie-break — is an FLX_ERR_ARG here (flx_upload_scene's traversal stack holds 64
 * entries and would refuse it): rebuild with a larger max_leaf.
 * Stateless and synchronous.  No CPU fallback. */
int flx_build_lbvh(int device, const flx_triangle *tris, size_t n_tris, uint32_t max_leaf,
                   flx_node *nodes_out, size_t nodes_capacity, size_t *n_nodes_out, uint32_t *indices_out, uint32_t *depth_out);

/* ---- ray-cast entry points (parity + microbench; cover kernel_pick.cl as well) -------------------
 * rays: n x 8 floats (ox,oy,oz,tmax, dx,dy,dz,unused).  Host-buffer variants copy in, trace, copy
 * out and synchronise; *_device variants take device pointers and only enqueue. Any output may be NULL. */
int flx_trace_closest(flx_ctx *ctx, const float *rays, size_t n, int32_t *hit_i, float *hit_t, flx_hit *hits);
int flx_trace_any(flx_ctx *ctx, const float *rays, size_t n, uint8_t *occluded);
int flx_trace_closest_device(flx_ctx *ctx, const void *d_rays, size_t n, void *d_hit_i, void *d_hit_t, void *d_hits);
int flx_trace_any_device(flx_ctx *ctx, const void *d_rays, size_t n, void *d_occluded);

/* ---- ray cast without allocation (NO counterpart in the reference) ----
 * flx_trace_closest for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it): same arguments minus the hit records, same results bit for bit, but the device staging
 * of rays and results lives in a file-scope static device array instead of a scratch block that grows with
 * cudaMalloc.  The call path contains no cudaMalloc / cudaFree / allocating library routine.  Limits and rules:
 *   - at most FLX_STATIC_MAX_RAYS rays per call; a larger n is rejected with FLX_ERR_ARG and a message;
 *   - every GPU that loads the library holds the array (10.0 MiB) until the process exits;
 *   - the array is shared by ALL contexts on a device and keeps the last call's contents (the call clears what must
 *     start at zero); calls that overlap on the GPU race on it — serialise them.  The call synchronises the
 *     context's stream before it returns.
 * rays: n x 8 floats as for flx_trace_closest (host memory); hit_i / hit_t: n entries each (host memory), required. */
#define FLX_STATIC_MAX_RAYS 262144u
int flx_trace_closest_static(flx_ctx *ctx, const float *rays, size_t n, int32_t *hit_i, float *hit_t);

/* ---- state inspection for stage-by-stage parity (no reference counterpart; the reference's
 * copyToHost helper, clcontext.cpp:103-115, is the nearest thing) -------------------------------- */
int flx_export_state(flx_ctx *ctx, uint32_t *slots64xN); /* reference SoA layout, see fluctus_abi.h */
int flx_import_state(flx_ctx *ctx, const uint32_t *slots64xN);
int flx_read_queue(flx_ctx *ctx, int queue, uint32_t *out, size_t n);
int flx_write_queue(flx_ctx *ctx, int queue, const uint32_t *in, size_t n);
int flx_set_counters(flx_ctx *ctx, const flx_queue_counters *in);
int flx_write_pixels(flx_ctx *ctx, int which, const void *host_in, size_t bytes);

/* ---- instrumentation --------------------------------------------------------------------------- */
uint64_t flx_launch_count(flx_ctx *ctx);          /* kernels launched by this context so far */
/* CUDA-event timing of the trace kernels on the context's stream (CLContext::checkTracingPerf,
 * clcontext.cpp:691-719).  enable != 0 starts recording; flx_get_kernel_time returns accumulated
 * milliseconds and launch counts per stage id (0 ext, 1 shadow, 2 logic, 3 raygen, 4 materials). */
int flx_enable_timing(flx_ctx *ctx, int enable);
int flx_get_kernel_time(flx_ctx *ctx, int stage, double *ms, uint64_t *launches); /* stage 5 = ray batch */
/* Visit counters of the traversal kernels, for the algorithmic-bytes roofline figure (the oracle
 * counts the same quantities under the reference visit order).  While enabled, instrumented
 * kernel variants run — never time those.  out = 4 groups {ext, shadow, batch closest, batch any}
 * x {inner-node visits, leaf visits, triangle tests, hits, rays}. */
int flx_set_trace_counting(flx_ctx *ctx, int enable);
int flx_get_trace_counters(flx_ctx *ctx, uint64_t out[20]);

#ifdef __cplusplus
}
#endif
#endif /* FLUCTUS_CUDA_H */
