"""TEST INFRASTRUCTURE — ctypes bindings for the CPU oracle (oracle/liboracle.so) and for the
reference's own host code built in place (oracle/_ref/libflxref_host.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product (fluctus_b200/) never does.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = None

V = C.c_void_p
U32 = C.c_uint32
U64 = C.c_uint64

# numpy views of the boundary structs (include/fluctus_abi.h)
TRI_DT = np.dtype([("v", np.float32, (3, 3, 4)), ("matId", np.int32), ("_pad", np.int32, 3)])  # 160 B
NODE_DT = np.dtype([("bmin", np.float32, 4), ("bmax", np.float32, 4), ("parent", np.int32),
                    ("iStartOrRight", np.uint32), ("nPrims", np.uint8), ("_pad", np.uint8, 7)])  # 48 B
MAT_DT = np.dtype([("Kd", np.float32, 4), ("Ks", np.float32, 4), ("Ke", np.float32, 4), ("Kt", np.float32, 4),
                   ("Ns", np.float32), ("Ni", np.float32), ("d", np.float32), ("map_Kd", np.int32),
                   ("map_Ks", np.int32), ("map_N", np.int32), ("type", np.int32), ("_pad", np.int32)])  # 96 B
TEX_DT = np.dtype([("offset", np.uint32), ("width", np.uint32), ("height", np.uint32)])  # 12 B
HIT_DT = np.dtype([("P", np.float32, 4), ("N", np.float32, 4), ("uv", np.float32, 2), ("t", np.float32),
                   ("i", np.int32), ("areaLightHit", np.int32), ("matId", np.int32), ("_pad", np.int32, 2)])  # 64 B
COUNTERS_DT = np.dtype([(n, np.uint32) for n in ("raygenQueue", "extensionQueue", "shadowQueue", "diffuseQueue",
                                                  "glossyQueue", "ggxReflQueue", "ggxRefrQueue", "deltaQueue",
                                                  "emissiveQueue", "splattedSamples")])  # 40 B
assert TRI_DT.itemsize == 160 and NODE_DT.itemsize == 48 and MAT_DT.itemsize == 96
assert TEX_DT.itemsize == 12 and HIT_DT.itemsize == 64 and COUNTERS_DT.itemsize == 40


def _ptr(a):
    return a.ctypes.data_as(V) if a is not None else None


def lib():
    """Load (building if needed and possible) oracle/liboracle.so."""
    global _LIB
    if _LIB is None:
        # FLX_ORACLE_VARIANT=exactlibm: the build with double-precision transcendental functions (o_math.h, test_exact_libm_gpu.py)
        name = "liboracle_%s.so" % os.environ["FLX_ORACLE_VARIANT"] if os.environ.get("FLX_ORACLE_VARIANT") else "liboracle.so"
        path = os.path.join(_HERE, name)
        if not os.path.exists(path):
            import subprocess
            subprocess.check_call(["make", "-s", "-C", _HERE, name])
        L = C.CDLL(path)
        L.orc_wf_create.restype = V
        for name in ("orc_wf_state", "orc_wf_queue", "orc_wf_counters", "orc_wf_pixels", "orc_wf_spp",
                     "orc_wf_denoiser_albedo", "orc_wf_denoiser_normal", "orc_mk_stats"):
            getattr(L, name).restype = V
        L.orc_wf_render.restype = U64
        L.orc_wf_get_pixel_index.restype = U32
        L.orc_max_threads.restype = C.c_int
        _LIB = L
    return _LIB


def set_threads(n):
    lib().orc_set_threads(C.c_int(int(n)))


def max_threads():
    return lib().orc_max_threads()


def trace_closest(scene, rays, mode=0, want_hits=False, want_counters=False):
    """rays: (n, 8) float32 (ox,oy,oz,tmax,dx,dy,dz,pad).  Returns dict(hit_i, hit_t[, hits][, counters])."""
    L = lib()
    rays = np.ascontiguousarray(rays, np.float32)
    n = rays.shape[0]
    hit_i = np.empty(n, np.int32)
    hit_t = np.empty(n, np.float32)
    hits = np.zeros(n, HIT_DT) if want_hits else None
    cnt = np.zeros(5, np.uint64)
    L.orc_trace_closest(_ptr(scene["tris"]), U32(len(scene["tris"])), _ptr(scene["indices"]), U32(len(scene["indices"])),
                        _ptr(scene["nodes"]), U32(len(scene["nodes"])), _ptr(rays), U64(n), C.c_int(mode),
                        _ptr(hit_i), _ptr(hit_t), _ptr(hits), _ptr(cnt))
    out = {"hit_i": hit_i, "hit_t": hit_t}
    if want_hits:
        out["hits"] = hits
    if want_counters:
        out["counters"] = dict(zip(("inner", "leaf", "tri", "hits", "max_stack"), (int(x) for x in cnt)))
    return out


def trace_any(scene, rays, mode=0, want_counters=False):
    L = lib()
    rays = np.ascontiguousarray(rays, np.float32)
    n = rays.shape[0]
    occ = np.empty(n, np.uint8)
    cnt = np.zeros(5, np.uint64)
    L.orc_trace_any(_ptr(scene["tris"]), U32(len(scene["tris"])), _ptr(scene["indices"]), U32(len(scene["indices"])),
                    _ptr(scene["nodes"]), U32(len(scene["nodes"])), _ptr(rays), U64(n), C.c_int(mode), _ptr(occ), _ptr(cnt))
    if want_counters:
        return occ, dict(zip(("inner", "leaf", "tri", "hits", "max_stack"), (int(x) for x in cnt)))
    return occ


class OracleWavefront:
    """CPU restatement of the wavefront stages over the reference's SoA state layout."""

    def __init__(self, num_tasks, width, height):
        self.L = lib()
        self.N, self.width, self.height = int(num_tasks), int(width), int(height)
        self.h = V(self.L.orc_wf_create(U32(self.N), U32(self.width), U32(self.height)))
        self._keep = {}

    def close(self):
        if self.h:
            self.L.orc_wf_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_scene(self, scene):
        s = {k: np.ascontiguousarray(scene[k]) for k in ("tris", "indices", "nodes", "materials")}
        s["textures"] = np.ascontiguousarray(scene.get("textures", np.zeros(0, TEX_DT)))
        s["texels"] = np.ascontiguousarray(scene.get("texels", np.zeros(0, np.uint8)))
        self._keep["scene"] = s
        self.L.orc_wf_set_scene(self.h, _ptr(s["tris"]), U32(len(s["tris"])), _ptr(s["indices"]), U32(len(s["indices"])),
                                _ptr(s["nodes"]), U32(len(s["nodes"])), _ptr(s["materials"]),
                                _ptr(s["textures"]) if len(s["textures"]) else None,
                                _ptr(s["texels"]) if len(s["texels"]) else None)

    def set_envmap(self, env):
        """env: dict(rgb (h,w,3) f32, prob, alias, pdf) as EnvironmentMap holds them."""
        h, w = env["rgb"].shape[:2]
        rgba = np.ones((h, w, 4), np.float32)
        rgba[..., :3] = env["rgb"]
        e = {"rgba": rgba, "prob": np.ascontiguousarray(env["prob"], np.float32),
             "alias": np.ascontiguousarray(env["alias"], np.int32), "pdf": np.ascontiguousarray(env["pdf"], np.float32)}
        self._keep["env"] = e
        self.L.orc_wf_set_envmap(self.h, _ptr(e["rgba"]), C.c_int(w), C.c_int(h), _ptr(e["prob"]), _ptr(e["alias"]), _ptr(e["pdf"]))

    def set_params(self, params_bytes):
        buf = np.frombuffer(bytes(params_bytes), np.uint8).copy()
        assert buf.size == 256
        self.L.orc_wf_set_params(self.h, _ptr(buf))

    def set_options(self, bitstack=False, denoiser_aov=False, seed_offset=0):
        self.L.orc_wf_set_options(self.h, C.c_int(int(bitstack)), C.c_int(int(denoiser_aov)), U32(seed_offset))

    def set_drop_mixed(self, drop=True):
        """Reference-exact BXDF_MIXED with separate queues: the path is dropped (wf_logic.cl:399-401)."""
        self.L.orc_wf_set_drop_mixed(self.h, C.c_int(int(drop)))

    # stages
    def reset(self): self.L.orc_wf_reset(self.h)
    def raygen(self): self.L.orc_wf_raygen(self.h)
    def logic(self, first_iteration=False): self.L.orc_wf_logic(self.h, C.c_int(int(first_iteration)))
    def materials(self): self.L.orc_wf_materials(self.h)
    def ext(self): self.L.orc_wf_ext(self.h)
    def shadow(self): self.L.orc_wf_shadow(self.h)
    def clear_queues(self): self.L.orc_wf_clear_queues(self.h)
    def update_pixel_index(self, n): self.L.orc_wf_update_pixel_index(self.h, U32(int(n)))
    def set_pixel_index(self, v): self.L.orc_wf_set_pixel_index(self.h, U32(int(v)))
    def get_pixel_index(self): return int(self.L.orc_wf_get_pixel_index(self.h))

    def pick(self, ndc_x, ndc_y):
        """kernel_pick.cl: the Hit under an NDC point (HIT_DT record)."""
        from fluctus_b200.abi import HIT_DT
        out = np.zeros(1, HIT_DT)
        self.L.orc_wf_pick(self.h, C.c_float(ndc_x), C.c_float(ndc_y), _ptr(out))
        return out[0]

    # microkernel renderer (src/mk_*.cl): one task per pixel, phase word in state slot 46
    def mk_reset(self): self.L.orc_mk_reset(self.h)
    def mk_raygen(self): self.L.orc_mk_raygen(self.h)
    def mk_next_vertex(self): self.L.orc_mk_next_vertex(self.h)
    def mk_sample_bsdf(self): self.L.orc_mk_sample_bsdf(self.h)
    def mk_splat(self): self.L.orc_mk_splat(self.h)
    def mk_splat_preview(self): self.L.orc_mk_splat_preview(self.h)
    def mk_render_single(self, spp): self.L.orc_mk_render_single(self.h, U32(int(spp)))
    def mk_stats(self): return self._view(self.L.orc_mk_stats(self.h), (4,), np.uint32)

    def render(self, max_iterations, do_reset=True):
        st = np.zeros(4, np.uint64)
        it = self.L.orc_wf_render(self.h, U64(int(max_iterations)), C.c_int(int(do_reset)), _ptr(st))
        return int(it), dict(zip(("primary", "extension", "shadow", "samples"), (int(x) for x in st)))

    # views (live memory of the oracle)
    def _view(self, ptr, shape, dtype):
        n = int(np.prod(shape))
        buf = (C.c_uint8 * (n * np.dtype(dtype).itemsize)).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype).reshape(shape)

    def state(self): return self._view(self.L.orc_wf_state(self.h), (64, self.N), np.uint32)
    def queue(self, q): return self._view(self.L.orc_wf_queue(self.h, C.c_int(q)), (self.N,), np.uint32)
    def counters(self): return self._view(self.L.orc_wf_counters(self.h), (1,), COUNTERS_DT)[0]
    def pixels(self): return self._view(self.L.orc_wf_pixels(self.h), (self.height * self.width, 4), np.float32)
    def spp(self): return self._view(self.L.orc_wf_spp(self.h), (self.height * self.width,), np.uint32)
    def denoiser_albedo(self): return self._view(self.L.orc_wf_denoiser_albedo(self.h), (self.height * self.width, 4), np.float32)
    def denoiser_normal(self): return self._view(self.L.orc_wf_denoiser_normal(self.h), (self.height * self.width, 4), np.float32)

    def postprocess(self):
        out = np.empty((self.height * self.width, 4), np.float32)
        self.L.orc_wf_postprocess(self.h, _ptr(out))
        return out

    def trace_counters(self):
        e, s = np.zeros(5, np.uint64), np.zeros(5, np.uint64)
        self.L.orc_wf_trace_counters(self.h, _ptr(e), _ptr(s))
        names = ("inner", "leaf", "tri", "hits", "max_stack")
        return dict(zip(names, map(int, e))), dict(zip(names, map(int, s)))


# ---- the reference's own host code (scene loading, SBVH, env-map tables) ---------------------------
def ref_host():
    """oracle/_ref/libflxref_host.so — None when it was never built (no /root/reference)."""
    global _REF
    if _REF is None:
        path = os.path.join(_HERE, "_ref", "libflxref_host.so")
        if not os.path.exists(path):
            return None
        R = C.CDLL(path)
        R.refh_scene_new.restype = V
        for n in ("refh_num_tris", "refh_num_materials", "refh_num_textures", "refh_num_nodes", "refh_num_indices",
                  "refh_sizeof_tri", "refh_sizeof_node", "refh_sizeof_material"):
            getattr(R, n).restype = C.c_size_t
        for n in ("refh_tris", "refh_materials", "refh_nodes", "refh_indices"):
            getattr(R, n).restype = V
        R.refh_material_types.restype = C.c_uint
        for n in ("refh_num_tris", "refh_num_materials", "refh_num_textures", "refh_num_nodes", "refh_num_indices",
                  "refh_tris", "refh_materials", "refh_nodes", "refh_indices", "refh_material_types",
                  "refh_scene_free", "refh_build_sbvh"):
            getattr(R, n).argtypes = [V]
        R.refh_scene_load.argtypes = [V, C.c_char_p, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int]
        R.refh_build_bvh.argtypes = [V, C.c_int]
        R.refh_bounds.argtypes = [V, V, V]
        R.refh_texture_info.argtypes = [V, C.c_size_t, C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.POINTER(V)]
        R.refh_envmap_load.argtypes = [V, C.c_char_p]
        R.refh_envmap_info.argtypes = [V, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(V), C.POINTER(V), C.POINTER(V), C.POINTER(V)]
        _REF = R
    return _REF


def _copy(ptr, count, dtype):
    if count == 0:
        return np.zeros(0, dtype)
    buf = (C.c_uint8 * (count * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).copy()


def ref_load_scene(models, builder="sbvh"):
    """Load model files with the reference's Scene::loadModel and build the hierarchy with the
    reference's SBVH (or BVH split mode int).  models: path or list of (path, scale, (tx,ty,tz))."""
    R = ref_host()
    assert R is not None, "reference host library not built"
    assert R.refh_sizeof_tri() == 160 and R.refh_sizeof_node() == 48 and R.refh_sizeof_material() == 96
    h = V(R.refh_scene_new())
    if isinstance(models, str):
        models = [(models, None, None)]
    for path, scale, t in models:
        use_t = scale is not None
        t = t or (0.0, 0.0, 0.0)
        R.refh_scene_load(h, path.encode(), C.c_float(scale or 1.0), C.c_float(t[0]), C.c_float(t[1]), C.c_float(t[2]), C.c_int(int(use_t)))
    if builder == "sbvh":
        R.refh_build_sbvh(h)
    else:
        R.refh_build_bvh(h, C.c_int(int(builder)))
    scene = {
        "tris": _copy(R.refh_tris(h), R.refh_num_tris(h), TRI_DT),
        "materials": _copy(R.refh_materials(h), R.refh_num_materials(h), MAT_DT),
        "nodes": _copy(R.refh_nodes(h), R.refh_num_nodes(h), NODE_DT),
        "indices": _copy(R.refh_indices(h), R.refh_num_indices(h), np.uint32),
        "material_types": int(R.refh_material_types(h)),
    }
    mn, mx = np.zeros(3, np.float32), np.zeros(3, np.float32)
    R.refh_bounds(h, _ptr(mn), _ptr(mx))
    scene["bounds"] = np.stack([mn, mx])
    descs, texels, off = [], [], 0
    for i in range(R.refh_num_textures(h)):
        w, ht, data = C.c_uint(), C.c_uint(), V()
        R.refh_texture_info(h, C.c_size_t(i), C.byref(w), C.byref(ht), C.byref(data))
        nbytes = w.value * ht.value * 4
        descs.append((off, w.value, ht.value))
        texels.append(_copy(data.value, nbytes, np.uint8) if nbytes else np.zeros(0, np.uint8))
        off += nbytes
    scene["textures"] = np.array(descs, TEX_DT) if descs else np.zeros(0, TEX_DT)
    scene["texels"] = np.concatenate(texels) if texels else np.zeros(0, np.uint8)
    R.refh_scene_free(h)
    return scene


def ref_load_envmap(path):
    R = ref_host()
    assert R is not None, "reference host library not built"
    h = V(R.refh_scene_new())
    ok = R.refh_envmap_load(h, path.encode())
    assert ok, "env map load failed"
    w, ht = C.c_int(), C.c_int()
    rgb, prob, alias, pdf = V(), V(), V(), V()
    R.refh_envmap_info(h, C.byref(w), C.byref(ht), C.byref(rgb), C.byref(prob), C.byref(alias), C.byref(pdf))
    n = w.value * ht.value
    env = {"rgb": _copy(rgb.value, n * 3, np.float32).reshape(ht.value, w.value, 3),
           "prob": _copy(prob.value, n, np.float32), "alias": _copy(alias.value, n, np.int32),
           "pdf": _copy(pdf.value, n, np.float32)}
    R.refh_scene_free(h)
    return env
