// TEST INFRASTRUCTURE (oracle/): C-ABI shim over the reference's *unmodified* host sources
// (src/scene.cpp, bvh.cpp, sbvh.cpp, envmap.cpp ... compiled in place from $FLUCTUS_REF by
// oracle/ref_host/Makefile into oracle/_ref/libflxref_host.so).  It lets tests and the scene
// cache tool obtain exactly the byte layouts CLContext::uploadSceneData / createEnvMap hand to
// the device (reference src/clcontext.cpp:485-529, 540-629).  Nothing in the product links this file; the
// headless host (fluctus_b200/host) has its own loader on the same glue (ref_glue.hpp).
#include "ref_glue.hpp" // fluctus_b200/host: headless replacements for progressview.cpp / texture.cpp + BVH accessor
#include <cstdio>
#include <memory>

struct RefScene
{
    Scene scene;
    std::unique_ptr<BVH> bvh;
    std::unique_ptr<EnvironmentMap> env;
    ProgressView pv{nullptr};
};

extern "C" {

void *refh_scene_new() { return new RefScene(); }
void refh_scene_free(void *h) { delete (RefScene *)h; }

// Scene::loadModel by extension (reference src/scene.cpp:59-120). scale/translation mimic the
// ModelTransform used by .sc.json scenes (scene.hpp:21-32); pass scale=1, t=0 for a plain load.
int refh_scene_load(void *h, const char *path, float scale, float tx, float ty, float tz, int use_transform)
{
    RefScene *s = (RefScene *)h;
    ModelTransform t; t.scale = scale; t.translation = fr::float3(tx, ty, tz);
    s->scene.loadModel(path, &s->pv, use_transform ? &t : nullptr);
    return (int)s->scene.getTriangles().size();
}

size_t refh_num_tris(void *h) { return ((RefScene *)h)->scene.getTriangles().size(); }
const void *refh_tris(void *h) { return ((RefScene *)h)->scene.getTriangles().data(); }
size_t refh_sizeof_tri() { return sizeof(RTTriangle); }
size_t refh_num_materials(void *h) { return ((RefScene *)h)->scene.getMaterials().size(); }
const void *refh_materials(void *h) { return ((RefScene *)h)->scene.getMaterials().data(); }
size_t refh_sizeof_material() { return sizeof(Material); }
unsigned refh_material_types(void *h) { return ((RefScene *)h)->scene.getMaterialTypes(); }
size_t refh_num_textures(void *h) { return ((RefScene *)h)->scene.getTextures().size(); }
void refh_texture_info(void *h, size_t i, unsigned *w, unsigned *ht, const void **data)
{
    Texture *t = ((RefScene *)h)->scene.getTextures()[i];
    *w = t->getWidth(); *ht = t->getHeight(); *data = t->getData();
}

// new SBVH(tris, SplitMode::SAH, progress) as Tracer::initHierarchy does (src/tracer.cpp:934-952)
int refh_build_sbvh(void *h)
{
    RefScene *s = (RefScene *)h;
    s->bvh.reset(new SBVH(&s->scene.getTriangles(), SplitMode::SAH, &s->pv));
    return (int)CLContext::nodes(s->bvh.get()).size();
}
// plain BVH with the given split mode (src/bvh.cpp), for the consistency tests
int refh_build_bvh(void *h, int mode)
{
    RefScene *s = (RefScene *)h;
    s->bvh.reset(new BVH(&s->scene.getTriangles(), (SplitMode)mode));
    return (int)CLContext::nodes(s->bvh.get()).size();
}
size_t refh_num_nodes(void *h) { return CLContext::nodes(((RefScene *)h)->bvh.get()).size(); }
const void *refh_nodes(void *h) { return CLContext::nodes(((RefScene *)h)->bvh.get()).data(); }
size_t refh_sizeof_node() { return sizeof(Node); }
size_t refh_num_indices(void *h) { return CLContext::indices(((RefScene *)h)->bvh.get()).size(); }
const void *refh_indices(void *h) { return CLContext::indices(((RefScene *)h)->bvh.get()).data(); }
void refh_bounds(void *h, float *mn, float *mx)
{
    AABB_t b = ((RefScene *)h)->bvh->getSceneBounds();
    mn[0] = b.min.x; mn[1] = b.min.y; mn[2] = b.min.z;
    mx[0] = b.max.x; mx[1] = b.max.y; mx[2] = b.max.z;
}

// EnvironmentMap(.hdr) + computeProbabilities (src/envmap.cpp:9-116)
int refh_envmap_load(void *h, const char *path)
{
    RefScene *s = (RefScene *)h;
    s->env.reset(new EnvironmentMap(path));
    return s->env->valid() ? 1 : 0;
}
void refh_envmap_info(void *h, int *w, int *ht, const float **rgb, const float **prob, const int **alias, const float **pdf)
{
    EnvironmentMap *e = ((RefScene *)h)->env.get();
    *w = e->getWidth(); *ht = e->getHeight();
    *rgb = e->getData(); *prob = e->getProbTable(); *alias = e->getAliasTable(); *pdf = e->getPdfTable();
}

} // extern "C"
