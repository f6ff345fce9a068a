// HostScene — the arrays the reference's Scene / BVH / EnvironmentMap objects hand to the device
// (reference src/clcontext.cpp:485-529, 540-629), held in the boundary layout of include/fluctus_abi.h,
// plus a flat binary container (.flxs) so that scenes prepared once — by the reference's own loaders
// and SBVH builder, see ref_scene_loader.cpp — can be rendered on a machine without the reference tree.
//
//   .flxs = "FLXS0001" | u64 nTris nIndices nNodes nMaterials nTextures texelBytes envW envH |
//           tris[160] indices[4] nodes[48] materials[96] textures[12] texels | env rgb[12] prob[4] alias[4] pdf[4]
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/fluctus_abi.h"

namespace flx {

struct HostScene
{
    std::vector<flx_triangle> tris;
    std::vector<uint32_t> indices;
    std::vector<flx_node> nodes;
    std::vector<flx_material> materials;
    std::vector<flx_tex_desc> textures;
    std::vector<uint8_t> texels;
    int envW = 0, envH = 0;
    std::vector<float> envRgb, envProb, envPdf;
    std::vector<int32_t> envAlias;

    bool hasEnvMap() const { return envW > 0 && envH > 0; }
    // params.worldRadius = length(bounds.max - bounds.min) * 0.5f (src/tracer.cpp:79-80); bounds = root box
    float worldRadius() const
    {
        if (nodes.empty()) return 1.0f;
        const float dx = nodes[0].bmax.x - nodes[0].bmin.x, dy = nodes[0].bmax.y - nodes[0].bmin.y, dz = nodes[0].bmax.z - nodes[0].bmin.z;
        return std::sqrt(dx * dx + dy * dy + dz * dz) * 0.5f;
    }
};

namespace detail {
template <class T> inline void writeVec(FILE *f, const std::vector<T> &v)
{
    if (!v.empty() && fwrite(v.data(), sizeof(T), v.size(), f) != v.size()) throw std::runtime_error("flxs: short write");
}
template <class T> inline void readVec(FILE *f, std::vector<T> &v, uint64_t n)
{
    v.resize((size_t)n);
    if (n && fread(v.data(), sizeof(T), (size_t)n, f) != n) throw std::runtime_error("flxs: truncated file");
}
} // namespace detail

inline void saveScene(const std::string &path, const HostScene &s)
{
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("flxs: cannot write " + path);
    const uint64_t hdr[8] = {s.tris.size(), s.indices.size(), s.nodes.size(), s.materials.size(), s.textures.size(),
                             s.texels.size(), (uint64_t)s.envW, (uint64_t)s.envH};
    try
    {
        if (fwrite("FLXS0001", 1, 8, f) != 8 || fwrite(hdr, 8, 8, f) != 8) throw std::runtime_error("flxs: short write");
        detail::writeVec(f, s.tris); detail::writeVec(f, s.indices); detail::writeVec(f, s.nodes); detail::writeVec(f, s.materials);
        detail::writeVec(f, s.textures); detail::writeVec(f, s.texels);
        detail::writeVec(f, s.envRgb); detail::writeVec(f, s.envProb); detail::writeVec(f, s.envAlias); detail::writeVec(f, s.envPdf);
    }
    catch (...) { fclose(f); throw; }
    fclose(f);
}

inline HostScene loadScene(const std::string &path)
{
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("flxs: cannot open " + path);
    HostScene s;
    try
    {
        char magic[8];
        uint64_t hdr[8];
        if (fread(magic, 1, 8, f) != 8 || memcmp(magic, "FLXS0001", 8) != 0) throw std::runtime_error("flxs: bad magic in " + path);
        if (fread(hdr, 8, 8, f) != 8) throw std::runtime_error("flxs: truncated header");
        if (hdr[0] == 0 || hdr[1] == 0 || hdr[2] == 0 || hdr[3] == 0) throw std::runtime_error("flxs: empty scene");
        if (hdr[0] > (1ull << 31) || hdr[1] > (1ull << 31) || hdr[2] > (1ull << 31) || hdr[6] > 65536 || hdr[7] > 65536)
            throw std::runtime_error("flxs: implausible header");
        detail::readVec(f, s.tris, hdr[0]); detail::readVec(f, s.indices, hdr[1]); detail::readVec(f, s.nodes, hdr[2]);
        detail::readVec(f, s.materials, hdr[3]); detail::readVec(f, s.textures, hdr[4]); detail::readVec(f, s.texels, hdr[5]);
        s.envW = (int)hdr[6]; s.envH = (int)hdr[7];
        const uint64_t n = hdr[6] * hdr[7];
        detail::readVec(f, s.envRgb, 3 * n); detail::readVec(f, s.envProb, n); detail::readVec(f, s.envAlias, n); detail::readVec(f, s.envPdf, n);
    }
    catch (...) { fclose(f); throw; }
    fclose(f);
    return s;
}

#ifdef FLX_WITH_REFERENCE_HOST
// Implemented in ref_scene_loader.cpp on top of the reference's own Scene::loadModel, SBVH and
// EnvironmentMap (compiled in place from $FLUCTUS_REF): .obj / .ply / .pbrt / .pbf / .sc.json + .hdr.
HostScene loadSceneWithReference(const std::string &modelPath, const std::string &envMapPath);
#endif

} // namespace flx
