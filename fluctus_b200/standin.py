"""Procedural stand-in scenes for the configs whose geometry is stripped from the reference
snapshot (conference.obj, Country-Kitchen.obj, luxball.obj — reference .MISSING_LARGE_BLOBS).

They are assembled from the committed teapot fixture (reference PLY loader + reference SBVH) so the
same scene can be rebuilt bit-identically on any machine: a closed room with a window wall open
to the environment map, a ceiling area light, and a grid of teapots whose materials follow the mix
of Country-Kitchen.mtl (21 glossy, 8 ideal_reflection, 6 rough_reflection, 4 ideal_dielectric,
57 diffuse/inferred out of 96).  Hierarchies of the parts are joined under a
median-split top level that obeys the reference's Node[] invariants."""
import numpy as np

from . import abi
from .abi import MAT_DT, NODE_DT, TEX_DT, TRI_DT
from .scene import scene_bounds


def _tri(p0, p1, p2, n, mat, uv=((0, 0), (1, 0), (1, 1))):
    t = np.zeros(1, TRI_DT)
    for k, p in enumerate((p0, p1, p2)):
        t["v"][0, k, 0, :3] = p
        t["v"][0, k, 1, :3] = n
        t["v"][0, k, 2, :2] = uv[k]
    t["matId"] = mat
    return t


def quad(a, b, c, d, mat, uv_scale=1.0):
    """Two triangles a-b-c, a-c-d with the geometric normal of a-b-c."""
    a, b, c, d = (np.asarray(x, np.float32) for x in (a, b, c, d))
    n = np.cross(b - a, c - a)
    n = n / np.linalg.norm(n)
    s = uv_scale
    return np.concatenate([_tri(a, b, c, n, mat, ((0, 0), (s, 0), (s, s))), _tri(a, c, d, n, mat, ((0, 0), (s, s), (0, s)))])


def build_median_bvh(tris, leaf_size=4):
    """Small host-side builder (object median split) producing the reference's Node[]/indices layout."""
    P = tris["v"][:, :, 0, :3]
    tmin, tmax = P.min(1), P.max(1)
    cen = 0.5 * (tmin + tmax)
    nodes, indices = [], []

    def rec(ids, parent):
        at = len(nodes)
        n = np.zeros(1, NODE_DT)[0]
        n["bmin"][:3], n["bmax"][:3] = tmin[ids].min(0), tmax[ids].max(0)
        n["parent"] = parent
        nodes.append(n)
        if len(ids) <= leaf_size:
            n["iStartOrRight"], n["nPrims"] = len(indices), len(ids)
            indices.extend(int(i) for i in ids)
            return
        c = cen[ids]
        axis = int(np.argmax(c.max(0) - c.min(0)))
        order = ids[np.argsort(c[:, axis], kind="stable")]
        half = len(order) // 2
        rec(order[:half], at)
        nodes[at]["iStartOrRight"] = len(nodes)
        rec(order[half:], at)

    rec(np.arange(len(tris)), -1)
    return np.array(nodes, NODE_DT), np.array(indices, np.uint32)


def assemble(parts):
    """Join parts = [(tris, nodes, indices, offset xyz, matId override or None)] under one top level."""
    K = len(parts)
    tri_base = np.cumsum([0] + [len(p[0]) for p in parts])
    idx_base = np.cumsum([0] + [len(p[2]) for p in parts])
    bmin = np.array([p[1][0]["bmin"][:3] + np.asarray(p[3], np.float32) for p in parts], np.float32)
    bmax = np.array([p[1][0]["bmax"][:3] + np.asarray(p[3], np.float32) for p in parts], np.float32)
    centre = 0.5 * (bmin + bmax)
    total = (K - 1) + sum(len(p[1]) for p in parts)
    out_nodes = np.zeros(total, NODE_DT)
    base = np.zeros(K, np.int64)
    par = np.full(K, -1, np.int64)

    def build(ids, at, parent):
        if len(ids) == 1:
            base[ids[0]], par[ids[0]] = at, parent
            return at + len(parts[ids[0]][1])
        c = centre[ids]
        axis = int(np.argmax(c.max(0) - c.min(0)))
        order = ids[np.argsort(c[:, axis], kind="stable")]
        half = len(order) // 2
        n = out_nodes[at]
        n["bmin"][:3], n["bmax"][:3] = bmin[ids].min(0), bmax[ids].max(0)
        n["parent"], n["nPrims"] = parent, 0
        right = build(order[:half], at + 1, at)
        out_nodes[at]["iStartOrRight"] = right
        return build(order[half:], right, at)

    assert build(np.arange(K), 0, -1) == total
    out_tris, out_idx = [], []
    for k, (tris, nodes, idx, off, mat) in enumerate(parts):
        off = np.asarray(off, np.float32)
        t = tris.copy()
        t["v"][:, :, 0, :3] += off
        if mat is not None:
            t["matId"] = mat
        out_tris.append(t)
        out_idx.append((idx.astype(np.int64) + tri_base[k]).astype(np.uint32))
        b = int(base[k])
        sub = nodes.copy()
        sub["bmin"][:, :3] += off
        sub["bmax"][:, :3] += off
        leaf = nodes["nPrims"] != 0
        sub["parent"] = np.where(nodes["parent"] < 0, par[k], nodes["parent"] + b)
        sub["iStartOrRight"] = np.where(leaf, nodes["iStartOrRight"].astype(np.int64) + idx_base[k],
                                        nodes["iStartOrRight"].astype(np.int64) + b).astype(np.uint32)
        out_nodes[b:b + len(nodes)] = sub
    return np.concatenate(out_tris), out_nodes, np.concatenate(out_idx)


def _material(type_, Kd=(0.64, 0.64, 0.64), Ks=(0.0, 0.0, 0.0), Ke=(0, 0, 0), Ns=0.1, Ni=1.5, map_Kd=-1):
    m = np.zeros(1, MAT_DT)
    m["Kd"][0, :3], m["Ks"][0, :3], m["Ke"][0, :3] = Kd, Ks, Ke
    m["Ns"], m["Ni"], m["d"] = Ns, Ni, 1.0
    m["map_Kd"], m["map_Ks"], m["map_N"] = map_Kd, -1, -1
    m["type"] = type_
    return m


def checker_texture(size=256, cells=8, a=(205, 190, 160), b=(120, 70, 40)):
    yy, xx = np.mgrid[0:size, 0:size]
    sel = (((xx * cells) // size + (yy * cells) // size) % 2).astype(bool)
    img = np.zeros((size, size, 4), np.uint8)
    img[..., :3] = np.where(sel[..., None], np.array(a, np.uint8), np.array(b, np.uint8))
    img[..., 3] = 255
    return img


def kitchen_standin(teapot, grid=(10, 3, 4), seed=7):
    """Room 16 x 6 x 10 with a window wall, teapots on three shelf levels; ~grid teapots * 3206 tris.
    Returns (scene dict, camera/light keyword overrides for make_render_params)."""
    rng = np.random.default_rng(seed)
    tex = checker_texture()
    textures = np.array([(0, tex.shape[1], tex.shape[0])], TEX_DT)
    texels = tex.reshape(-1)
    mats = [_material(abi.BXDF_DIFFUSE)]                                                   # 0 default (scene.cpp:18-32)
    mats += [_material(abi.BXDF_DIFFUSE, Kd=(0.75, 0.72, 0.68))]                            # 1 walls
    mats += [_material(abi.BXDF_GLOSSY, Kd=(1, 1, 1), Ks=(0.04, 0.04, 0.04), Ns=0.12, Ni=1.5, map_Kd=0)]  # 2 textured floor
    pool = []                                                                              # teapot materials, kitchen mix
    for _ in range(21):
        pool.append(_material(abi.BXDF_GLOSSY, Kd=rng.uniform(0.2, 0.9, 3), Ks=rng.uniform(0.02, 0.3, 3), Ns=rng.uniform(0.05, 0.4), Ni=1.5))
    for _ in range(8):
        pool.append(_material(abi.BXDF_IDEAL_REFLECTION, Ks=rng.uniform(0.7, 0.95, 3)))
    for _ in range(6):
        pool.append(_material(abi.BXDF_GGX_ROUGH_REFLECTION, Ks=rng.uniform(0.6, 0.95, 3), Ns=rng.uniform(0.05, 0.3), Ni=1.0))
    for _ in range(4):
        pool.append(_material(abi.BXDF_IDEAL_DIELECTRIC, Ks=(0.95, 0.97, 0.95), Ni=1.5))
    for _ in range(57):
        pool.append(_material(abi.BXDF_DIFFUSE, Kd=rng.uniform(0.15, 0.85, 3)))
    first_pool = len(mats)
    mats += pool
    materials = np.concatenate(mats)

    X0, X1, Y0, Y1, Z0, Z1 = -8.0, 8.0, 0.0, 6.0, -5.0, 5.0
    room = [quad((X0, Y0, Z1), (X1, Y0, Z1), (X1, Y0, Z0), (X0, Y0, Z0), 2, uv_scale=6.0),   # floor
            quad((X0, Y1, Z0), (X1, Y1, Z0), (X1, Y1, Z1), (X0, Y1, Z1), 1),                # ceiling
            quad((X0, Y0, Z0), (X1, Y0, Z0), (X1, Y1, Z0), (X0, Y1, Z0), 1),                # back wall
            quad((X0, Y0, Z1), (X0, Y0, Z0), (X0, Y1, Z0), (X0, Y1, Z1), 1),                # left wall
            quad((X1, Y0, Z0), (X1, Y0, Z1), (X1, Y1, Z1), (X1, Y1, Z0), 1),                # right wall
            # front (window) wall: a frame around a 10 x 3 opening
            quad((X1, Y0, Z1), (X0, Y0, Z1), (X0, 1.5, Z1), (X1, 1.5, Z1), 1),
            quad((X1, 4.5, Z1), (X0, 4.5, Z1), (X0, Y1, Z1), (X1, Y1, Z1), 1),
            quad((X0 + 3.0, 1.5, Z1), (X0, 1.5, Z1), (X0, 4.5, Z1), (X0 + 3.0, 4.5, Z1), 1),
            quad((X1, 1.5, Z1), (X1 - 3.0, 1.5, Z1), (X1 - 3.0, 4.5, Z1), (X1, 4.5, Z1), 1)]
    nx, ny, nz = grid
    # shelves under each teapot row
    for j in range(ny):
        y = 0.9 + 1.5 * j - 0.47
        room.append(quad((X0 + 1.0, y, 1.1), (X1 - 1.0, y, 1.1), (X1 - 1.0, y, -4.6), (X0 + 1.0, y, -4.6), 1))
    room = np.concatenate(room)
    rn, ri = build_median_bvh(room)
    parts = [(room, rn, ri, (0, 0, 0), None)]
    k = 0
    for i in range(nx):
        for j in range(ny):
            for l in range(nz):
                off = (X0 + 2.0 + (X1 - X0 - 4.0) * i / max(nx - 1, 1), 0.9 + 1.5 * j, -4.0 + 4.5 * l / max(nz - 1, 1))
                parts.append((teapot["tris"], teapot["nodes"], teapot["indices"], off, first_pool + (k * 37) % len(pool)))
                k += 1
    tris, nodes, indices = assemble(parts)
    scene = dict(tris=tris, nodes=nodes, indices=indices, materials=materials, textures=textures, texels=texels)
    scene["bounds"] = scene_bounds(scene)
    view = dict(cam_pos=(0.0, 3.6, 4.8), cam_dir=(0.0, -0.24253563, -0.9701425), cam_up=(0.0, 0.9701425, -0.24253563),
                cam_right=(1.0, 0.0, 0.0), fov=80.0,
                light_pos=(0.0, 5.95, -1.0), light_n=(0.0, -1.0, 0.0), light_right=(1.0, 0.0, 0.0), light_up=(0.0, 0.0, 1.0),
                light_e=(40.0, 38.0, 34.0), light_size=(1.5, 1.0))
    return scene, view


def conference_standin(cat, grid=(7, 1, 3)):
    """BASELINE config 3 stand-in (assets/conference/conference.obj, ~330 k triangles, is stripped from the reference
    snapshot): the reference's egyptcat.obj (16 040 triangles, textured glossy + diffuse + one MIXED material) instanced
    grid = 7 x 1 x 3 times = 336 840 triangles inside a closed room, lit by one area light under the ceiling — an enclosed,
    area-light-only scene like the conference room.  Returns (scene dict, camera / light overrides)."""
    b0, b1 = np.asarray(cat["bounds"][0], np.float32), np.asarray(cat["bounds"][1], np.float32)
    e = b1 - b0
    nx, ny, nz = grid
    sx, sz = float(e[0]) * 1.15, float(e[2]) * 1.3
    mats = np.concatenate([cat["materials"], _material(abi.BXDF_DIFFUSE, Kd=(0.72, 0.7, 0.66))])
    wall = len(mats) - 1
    X0, X1 = float(b0[0]) - 0.6 * sx, float(b0[0]) + (nx - 1) * sx + float(e[0]) + 0.6 * sx
    Z0, Z1 = float(b0[2]) - 0.6 * sz, float(b0[2]) + (nz - 1) * sz + float(e[2]) + 0.9 * sz
    Y0, Y1 = float(b0[1]) - 0.01, float(b0[1]) + 2.6 * float(e[1])
    room = np.concatenate([quad((X0, Y0, Z1), (X1, Y0, Z1), (X1, Y0, Z0), (X0, Y0, Z0), wall),   # floor
                           quad((X0, Y1, Z0), (X1, Y1, Z0), (X1, Y1, Z1), (X0, Y1, Z1), wall),   # ceiling
                           quad((X0, Y0, Z0), (X1, Y0, Z0), (X1, Y1, Z0), (X0, Y1, Z0), wall),   # back
                           quad((X1, Y0, Z1), (X0, Y0, Z1), (X0, Y1, Z1), (X1, Y1, Z1), wall),   # front
                           quad((X0, Y0, Z1), (X0, Y0, Z0), (X0, Y1, Z0), (X0, Y1, Z1), wall),   # left
                           quad((X1, Y0, Z0), (X1, Y0, Z1), (X1, Y1, Z1), (X1, Y1, Z0), wall)])  # right
    rn, ri = build_median_bvh(room)
    parts = [(room, rn, ri, (0, 0, 0), None)]
    for i in range(nx):
        for k in range(nz):
            parts.append((cat["tris"], cat["nodes"], cat["indices"], (i * sx, 0.0, k * sz), None))
    tris, nodes, indices = assemble(parts)
    scene = dict(tris=tris, nodes=nodes, indices=indices, materials=mats,
                 textures=cat.get("textures", np.zeros(0, TEX_DT)), texels=cat.get("texels", np.zeros(0, np.uint8)))
    scene["bounds"] = scene_bounds(scene)
    cx, cz = 0.5 * (X0 + X1), 0.5 * (Z0 + Z1)
    view = dict(cam_pos=(cx, Y0 + 0.62 * (Y1 - Y0), Z1 - 0.25 * sz), cam_dir=(0.0, -0.28734788, -0.9578263),
                cam_up=(0.0, 0.9578263, -0.28734788), cam_right=(1.0, 0.0, 0.0), fov=85.0,
                light_pos=(cx, Y1 - 0.05, cz), light_n=(0.0, -1.0, 0.0), light_right=(1.0, 0.0, 0.0), light_up=(0.0, 0.0, 1.0),
                light_e=(14.0, 13.5, 12.5), light_size=(0.3 * (X1 - X0), 0.3 * (Z1 - Z0)))
    return scene, view
