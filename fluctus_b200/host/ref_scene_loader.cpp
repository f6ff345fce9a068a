// Scene preparation with the reference's OWN host code (compiled in place from $FLUCTUS_REF, see Makefile):
// Scene::loadModel (OBJ+MTL / PLY / PBRT / PBF / .sc.json, reference src/scene.cpp:59-120), SBVH construction as
// Tracer::initHierarchy does (src/tracer.cpp:934-952, SplitMode::SAH), EnvironmentMap + alias tables
// (src/envmap.cpp:9-116) and the texture packing of CLContext::packTextures (src/clcontext.cpp:588-629).
// Only built when the reference tree is present (FLX_WITH_REFERENCE_HOST); the result is a HostScene that can
// be rendered directly or written to a .flxs file for machines without the reference tree.
#include "ref_glue.hpp"
#define FLX_WITH_REFERENCE_HOST 1
#include "host_scene.hpp"
#include <memory>

static_assert(sizeof(RTTriangle) == sizeof(flx_triangle) && sizeof(Node) == sizeof(flx_node) && sizeof(Material) == sizeof(flx_material),
              "include/fluctus_abi.h must mirror the reference's geom.h / triangle.hpp / bvhnode.hpp layouts");

namespace flx {

HostScene loadSceneWithReference(const std::string &modelPath, const std::string &envMapPath)
{
    ProgressView pv(nullptr);
    Scene scene;
    scene.loadModel(modelPath, &pv);
    std::vector<RTTriangle> &tris = scene.getTriangles();
    if (tris.empty()) throw std::runtime_error("reference loader produced no triangles for " + modelPath);
    std::unique_ptr<BVH> bvh(new SBVH(&tris, SplitMode::SAH, &pv));

    HostScene out;
    out.tris.resize(tris.size());
    memcpy(out.tris.data(), tris.data(), tris.size() * sizeof(RTTriangle)); // after the build: SBVH may reorder / append
    const std::vector<Node> &nodes = CLContext::nodes(bvh.get());
    out.nodes.resize(nodes.size());
    memcpy(out.nodes.data(), nodes.data(), nodes.size() * sizeof(Node));
    out.indices = CLContext::indices(bvh.get());
    const std::vector<Material> &mats = scene.getMaterials();
    out.materials.resize(mats.size());
    memcpy(out.materials.data(), mats.data(), mats.size() * sizeof(Material));

    uint32_t offset = 0; // CLContext::packTextures: RGBA8 texels back to back, byte offsets in the descriptors
    for (Texture *tex : scene.getTextures())
    {
        flx_tex_desc d = {offset, tex->getWidth(), tex->getHeight()};
        out.textures.push_back(d);
        const uint32_t len = tex->getWidth() * tex->getHeight() * 4;
        if (len) out.texels.insert(out.texels.end(), tex->getData(), tex->getData() + len);
        offset += len;
    }

    if (!envMapPath.empty())
    {
        EnvironmentMap env(envMapPath);
        if (!env.valid()) throw std::runtime_error("cannot load environment map " + envMapPath);
        out.envW = env.getWidth(); out.envH = env.getHeight();
        const size_t n = (size_t)out.envW * out.envH;
        out.envRgb.assign(env.getData(), env.getData() + 3 * n);
        out.envProb.assign(env.getProbTable(), env.getProbTable() + n);
        out.envAlias.assign(env.getAliasTable(), env.getAliasTable() + n);
        out.envPdf.assign(env.getPdfTable(), env.getPdfTable() + n);
    }
    return out;
}

} // namespace flx
