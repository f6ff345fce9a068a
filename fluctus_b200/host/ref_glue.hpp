// Glue for compiling the reference's UNMODIFIED host sources headless (scene loaders, SBVH builder,
// environment-map tables: src/scene.cpp bvh.cpp sbvh.cpp bvhnode.cpp envmap.cpp rgbe/rgbe.cpp utils.cpp +
// ext/pbrt-parser, built in place from $FLUCTUS_REF — never copied).  Two reference TUs cannot be
// used because they need a GUI / DevIL: progressview.cpp (nanogui) and texture.cpp (DevIL); ref_glue.cpp
// supplies no-op / stb_image replacements with the same contract.  Included only by TUs that are
// compiled with the reference's include path (ref_scene_loader.cpp here, oracle/ref_host/ref_host_api.cpp).
#pragma once
#include "scene.hpp"
#include "progressview.hpp"
#include "sbvh.hpp"
#include "envmap.hpp"
#include "texture.hpp"

// BVH keeps its flattened arrays protected but befriends `class CLContext` (reference src/bvh.hpp:15),
// the class that uploads them (src/clcontext.cpp:540-584).  These TUs never see the real CLContext, so
// they define a local one carrying only the accessors the upload needs.
class CLContext
{
public:
    static std::vector<Node> &nodes(BVH *b) { return b->m_nodes; }
    static std::vector<U32> &indices(BVH *b) { return b->m_indices; }
};
