"""Host-side scene containers: the arrays the reference's Scene/BVH/EnvironmentMap objects hold,
kept as numpy records in the boundary layout (include/fluctus_abi.h), with .npz persistence and
an instancing helper that replicates a model + its hierarchy on a grid (the `.sc.json`
multi-model scenes of reference src/scene.cpp:864-897, but re-using the sub-tree instead of
rebuilding it)."""
import numpy as np

from .abi import MAT_DT, NODE_DT, TEX_DT, TRI_DT, BXDF_DIFFUSE, world_radius_of

SCENE_KEYS = ("tris", "indices", "nodes", "materials", "textures", "texels", "bounds")
ENV_KEYS = ("rgb", "prob", "alias", "pdf")


def save_scene(path, scene):
    np.savez_compressed(path, **{k: np.asarray(scene[k]) for k in SCENE_KEYS if k in scene})


def load_scene(path):
    z = np.load(path)
    s = {k: z[k] for k in z.files}
    s["tris"] = np.ascontiguousarray(s["tris"]).view(TRI_DT).reshape(-1) if s["tris"].dtype != TRI_DT else s["tris"]
    s["nodes"] = np.ascontiguousarray(s["nodes"]).view(NODE_DT).reshape(-1) if s["nodes"].dtype != NODE_DT else s["nodes"]
    s["materials"] = np.ascontiguousarray(s["materials"]).view(MAT_DT).reshape(-1) if s["materials"].dtype != MAT_DT else s["materials"]
    if "textures" not in s:
        s["textures"], s["texels"] = np.zeros(0, TEX_DT), np.zeros(0, np.uint8)
    elif s["textures"].dtype != TEX_DT:
        s["textures"] = np.ascontiguousarray(s["textures"]).view(TEX_DT).reshape(-1)
    if "bounds" not in s:
        s["bounds"] = scene_bounds(s)
    return s


def save_envmap(path, env):
    np.savez_compressed(path, **{k: np.asarray(env[k]) for k in ENV_KEYS})


def load_envmap(path):
    z = np.load(path)
    return {k: z[k] for k in ENV_KEYS}


def scene_bounds(scene):
    """BVH::getSceneBounds: the root node's box."""
    n0 = scene["nodes"][0]
    return np.stack([n0["bmin"][:3], n0["bmax"][:3]]).astype(np.float32)


def world_radius(scene):
    return world_radius_of(scene["bounds"])


def default_material():
    """The material Scene::Scene() always installs at index 0 (reference src/scene.cpp:18-32)."""
    m = np.zeros(1, MAT_DT)
    m["Kd"][0, :3] = 0.64
    m["Ni"], m["Ns"] = 1.8, 700.0
    m["map_Kd"] = m["map_Ks"] = m["map_N"] = -1
    m["type"] = BXDF_DIFFUSE
    return m


def replicate_grid(scene, counts, spacing):
    """Instance `scene` counts=(nx,ny,nz) times with translation steps spacing=(sx,sy,sz).

    Triangles are copied with translated positions (float32 add, like ModelTransform::apply with
    scale 1, reference src/scene.hpp:27-31); the model's node array is copied per instance with
    translated boxes and re-based indices, and a median-split top level is laid over the
    instances.  The result obeys the reference's Node[] invariants (DFS order, left child =
    parent + 1, leaf iff nPrims != 0), so every traversal — oracle or CUDA — walks it unchanged."""
    nx, ny, nz = counts
    offs = np.array([(i * spacing[0], j * spacing[1], k * spacing[2]) for i in range(nx) for j in range(ny) for k in range(nz)], np.float32)
    K = len(offs)
    tris, nodes, idx = scene["tris"], scene["nodes"], scene["indices"]
    nt, nn, ni = len(tris), len(nodes), len(idx)

    out_tris = np.tile(tris, K)
    v = out_tris["v"].reshape(K, nt, 3, 3, 4)
    v[:, :, :, 0, :3] += offs[:, None, None, :]  # positions only; normals / uvs unchanged
    out_idx = (idx[None, :].astype(np.uint64) + (np.arange(K, dtype=np.uint64) * nt)[:, None]).astype(np.uint32).reshape(-1)

    # top level over instance boxes: recursive median split along the widest axis of the centres
    root_min, root_max = nodes[0]["bmin"][:3], nodes[0]["bmax"][:3]
    imin, imax = root_min[None, :] + offs, root_max[None, :] + offs
    centre = 0.5 * (imin + imax)
    total_nodes = (K - 1) + K * nn
    out_nodes = np.zeros(total_nodes, NODE_DT)
    inst_base = np.zeros(K, np.int64)    # where each instance's sub-tree starts
    inst_parent = np.full(K, -1, np.int64)

    def build(ids, at, parent):
        if len(ids) == 1:
            inst_base[ids[0]], inst_parent[ids[0]] = at, parent
            return at + nn
        c = centre[ids]
        axis = int(np.argmax(c.max(0) - c.min(0)))
        order = ids[np.argsort(c[:, axis], kind="stable")]
        half = len(order) // 2
        n = out_nodes[at]
        n["bmin"][:3], n["bmax"][:3] = imin[ids].min(0), imax[ids].max(0)
        n["parent"], n["nPrims"] = parent, 0
        right = build(order[:half], at + 1, at)
        out_nodes[at]["iStartOrRight"] = right
        return build(order[half:], right, at)

    end = build(np.arange(K), 0, -1)
    assert end == total_nodes
    is_leaf = nodes["nPrims"] != 0
    for k in range(K):
        b = int(inst_base[k])
        sub = nodes.copy()
        sub["bmin"][:, :3] += offs[k]
        sub["bmax"][:, :3] += offs[k]
        sub["parent"] = np.where(nodes["parent"] < 0, inst_parent[k], nodes["parent"] + b)
        sub["iStartOrRight"] = np.where(is_leaf, nodes["iStartOrRight"].astype(np.int64) + k * ni,
                                        nodes["iStartOrRight"].astype(np.int64) + b).astype(np.uint32)
        out_nodes[b:b + nn] = sub
    out = dict(scene)
    out.update(tris=out_tris, indices=out_idx, nodes=out_nodes)
    out["bounds"] = scene_bounds(out)
    return out


def bench_rays(n, bounds, first=0):
    """Incoherent ray set of SURVEY.md §8(d) cfg5: ray k has seed = k and takes six draws of the
    reference RNG (src/random.cl:7-22): origin uniform in the scene box grown by 25 % per side,
    direction uniform on the sphere, tmax = FLT_MAX.  Returns (n, 8) float32."""
    seed = (np.arange(first, first + n, dtype=np.uint64) & 0xffffffff).astype(np.uint32)

    def draw():
        nonlocal seed
        s = seed
        s = (s ^ np.uint32(61)) ^ (s >> np.uint32(16))
        s = s * np.uint32(9)
        s = s ^ (s >> np.uint32(4))
        s = s * np.uint32(0x27d4eb2d)
        s = s ^ (s >> np.uint32(15))
        seed = s
        return s.astype(np.float32) * np.float32(2.3283064365386963e-10)

    bmin, bmax = np.asarray(bounds[0], np.float32), np.asarray(bounds[1], np.float32)
    ext = bmax - bmin
    r = [draw() for _ in range(6)]
    rays = np.zeros((n, 8), np.float32)
    for a in range(3):
        rays[:, a] = (bmin[a] - np.float32(0.25) * ext[a]) + r[a] * (np.float32(1.5) * ext[a])
    z = np.float32(1.0) - np.float32(2.0) * r[3]
    phi = np.float32(6.2831853071795864) * r[4]
    s = np.sqrt(np.maximum(np.float32(0.0), np.float32(1.0) - z * z)).astype(np.float32)
    rays[:, 3] = np.float32(3.402823466e+38)
    rays[:, 4], rays[:, 5], rays[:, 6] = s * np.cos(phi).astype(np.float32), s * np.sin(phi).astype(np.float32), z
    return rays


# ---- .flxs: the flat container the C++ host reads (fluctus_b200/host/host_scene.hpp) ------------------------
FLXS_MAGIC = b"FLXS0001"


def save_flxs(path, scene, env=None):
    tex = np.ascontiguousarray(scene.get("textures", np.zeros(0, TEX_DT)))
    texels = np.ascontiguousarray(scene.get("texels", np.zeros(0, np.uint8)), np.uint8)
    eh, ew = (env["rgb"].shape[:2] if env is not None else (0, 0))
    hdr = np.array([len(scene["tris"]), len(scene["indices"]), len(scene["nodes"]), len(scene["materials"]), len(tex),
                    texels.size, ew, eh], np.uint64)
    with open(path, "wb") as f:
        f.write(FLXS_MAGIC)
        f.write(hdr.tobytes())
        for a in (scene["tris"], np.asarray(scene["indices"], np.uint32), scene["nodes"], scene["materials"], tex, texels):
            f.write(np.ascontiguousarray(a).tobytes())
        if env is not None:
            f.write(np.ascontiguousarray(env["rgb"], np.float32).tobytes())
            f.write(np.ascontiguousarray(env["prob"], np.float32).tobytes())
            f.write(np.ascontiguousarray(env["alias"], np.int32).tobytes())
            f.write(np.ascontiguousarray(env["pdf"], np.float32).tobytes())


def load_flxs(path):
    """Returns (scene dict, env dict or None)."""
    with open(path, "rb") as f:
        if f.read(8) != FLXS_MAGIC:
            raise ValueError("not a .flxs file: %s" % path)
        nt, ni, nn, nm, ntex, nb, ew, eh = (int(x) for x in np.frombuffer(f.read(64), np.uint64))

        def rd(dtype, n):
            return np.frombuffer(f.read(np.dtype(dtype).itemsize * n), dtype).copy()
        scene = dict(tris=rd(TRI_DT, nt), indices=rd(np.uint32, ni), nodes=rd(NODE_DT, nn), materials=rd(MAT_DT, nm),
                     textures=rd(TEX_DT, ntex), texels=rd(np.uint8, nb))
        scene["bounds"] = scene_bounds(scene)
        env = None
        if ew * eh:
            n = ew * eh
            env = dict(rgb=rd(np.float32, 3 * n).reshape(eh, ew, 3), prob=rd(np.float32, n), alias=rd(np.int32, n), pdf=rd(np.float32, n))
    return scene, env


def build_lbvh(scene, max_leaf=4, device=0):
    """Rebuild the hierarchy of `scene` on the GPU (flx_build_lbvh: Morton order + Karras' radix tree, emitted in the
    reference's Node[] / index-list format).  Returns a copy of the scene with `nodes`, `indices`, `bounds` replaced
    and `lbvh_depth` added; triangles, materials and textures are shared with the input."""
    import ctypes as C
    from ._lib import FluctusError, load
    L = load()
    tris = np.ascontiguousarray(scene["tris"], TRI_DT)
    n = len(tris)
    nodes = np.zeros(2 * n - 1, NODE_DT)
    indices = np.zeros(n, np.uint32)
    n_nodes, depth = C.c_size_t(0), C.c_uint32(0)
    rc = L.flx_build_lbvh(int(device), tris.ctypes.data_as(C.c_void_p), n, int(max_leaf), nodes.ctypes.data_as(C.c_void_p),
                          len(nodes), C.byref(n_nodes), indices.ctypes.data_as(C.c_void_p), C.byref(depth))
    if rc != 0:
        raise FluctusError("%s (flx_status %d)" % (L.flx_last_error(None).decode(), rc))
    out = dict(scene)
    out["nodes"] = nodes[:n_nodes.value].copy()
    out["indices"] = indices
    out["bounds"] = scene_bounds(out)
    out["lbvh_depth"] = int(depth.value)
    return out
