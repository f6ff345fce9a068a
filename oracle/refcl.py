"""TEST INFRASTRUCTURE — the reference's own OpenCL kernels, compiled from the sources where they lie
(/root/reference/src/*.cl) as C++ behind oracle/ref_cl/cl_shim.hpp, run on the host.

This is what pins the hand-written oracle (oracle/*.cpp): tests/test_reference_source_cpu.py runs both
on the same inputs and asserts bit-identical results.  It also serves as the `cpu_baseline` /
`--impl reference` arm of bench.py (kind "reference": the reference's kernel code on the host cores).

Build model (mirrors CLContext / clt::Kernel, src/kernel_impl.hpp): one "program" per kernel file,
compiled with the global options of CLContext::setKernelBuildSettings (src/clcontext.cpp:145-155:
-DGPU -DFLT_FLOAT_ATOMICS [-DUSE_BITSTACK] -DUSE_SOA; NVIDIA undefined) plus the kernel's own
getAdditionalBuildOptions.  Objects are cached under oracle/_ref/cl_obj/ keyed by their option string,
so a change of RenderParams "re-JITs" only the kernels whose options changed — when the reference tree
is present.  On a box without /root/reference only prebuilt objects can be used (build() in
__graft_entry__ prebuilds what bench.py and smoke() need).

Only tests/, __graft_entry__ and bench.py's CPU legs may import this module.
"""
import ctypes as C
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("FLUCTUS_REF", "/root/reference")
OUT = os.path.join(_HERE, "_ref")
GEN = os.path.join(OUT, "cl_gen")
OBJ = os.path.join(OUT, "cl_obj")
SRC = os.path.join(_HERE, "ref_cl")

CXX = "/usr/bin/g++"
# pinned float semantics, same as oracle/Makefile: no contraction, no fast-math
CXXFLAGS = ["-O2", "-std=c++17", "-fPIC", "-fopenmp", "-ffp-contract=off", "-fno-fast-math", "-fpermissive", "-w", "-Wno-narrowing"]
GLOBAL_OPTS = ["-DGPU", "-DFLT_FLOAT_ATOMICS", "-DUSE_SOA"]  # src/clcontext.cpp:147-151

BXDF_DEFINES = [  # getBxdfDefines, src/utils.cpp:165-185 (type bit -> define)
    (1 << 1, "BXDF_USE_DIFFUSE"), (1 << 2, "BXDF_USE_GLOSSY"), (1 << 3, "BXDF_USE_GGX_ROUGH_REFLECTION"),
    (1 << 4, "BXDF_USE_IDEAL_REFLECTION"), (1 << 5, "BXDF_USE_GGX_ROUGH_DIELECTRIC"),
    (1 << 6, "BXDF_USE_IDEAL_DIELECTRIC"), (1 << 7, "BXDF_USE_EMISSIVE"),
]

# kernel file -> which option groups it is built with (src/kernel_impl.hpp getAdditionalBuildOptions)
KERNELS = {
    "wf_reset": (), "wf_raygen": (), "wf_extrays": ("bitstack",), "wf_shadowrays": ("bitstack",),
    "wf_logic": ("logic",),
    "wf_mat_diffuse": (), "wf_mat_glossy": (), "wf_mat_ggx_reflection": (), "wf_mat_ggx_refraction": (),
    "wf_mat_delta": (), "wf_mat_emissive": (), "wf_mat_all": ("bxdf",),
    "mk_reset": (), "mk_raygen": (), "mk_next_vertex": ("bitstack", "denoiser"),
    "mk_sample_bsdf": ("bitstack", "denoiser", "bxdf"), "mk_splat": ("spp",), "mk_splat_preview": (),
    "mk_postprocess": ("denoiser",), "kernel_pick": ("bitstack",),
}


def have_reference():
    return os.path.isdir(os.path.join(REF, "src"))


def _generate():
    """Lexical C++-ification of the .cl files into oracle/_ref/cl_gen (see ref_cl/clpp.py)."""
    sys.path.insert(0, SRC)
    try:
        import clpp
    finally:
        sys.path.pop(0)
    clpp.main(os.path.join(REF, "src"), GEN)


class Options:
    """The -D switches derived from RenderParams / Settings / the scene's material mask."""

    def __init__(self, use_area_light=False, use_env_map=False, sample_expl=True, sample_impl=True,
                 separate_queues=False, check_spp=False, denoiser=False, bitstack=False, material_types=0xFE):
        self.use_area_light, self.use_env_map = bool(use_area_light), bool(use_env_map)
        self.sample_expl, self.sample_impl = bool(sample_expl), bool(sample_impl)
        self.separate_queues, self.check_spp = bool(separate_queues), bool(check_spp)
        self.denoiser, self.bitstack = bool(denoiser), bool(bitstack)
        self.material_types = int(material_types)

    @classmethod
    def from_params(cls, p, denoiser=False, bitstack=False, material_types=0xFE):
        """p: numpy record of fluctus_b200.abi.PARAMS_DT (RenderParams)."""
        return cls(int(p["useAreaLight"]), int(p["useEnvMap"]), int(p["sampleExpl"]), int(p["sampleImpl"]),
                   int(p["wfSeparateQueues"]), int(p["maxSpp"]) > 0, denoiser, bitstack, material_types)

    def defines(self, kernel):
        d = []
        for group in KERNELS[kernel]:
            if group == "bitstack" and self.bitstack:
                d.append("USE_BITSTACK")
            elif group == "denoiser" and self.denoiser:
                d.append("USE_OPTIX_DENOISER")
            elif group == "spp" and self.check_spp:
                d.append("CHECK_SPP")
            elif group == "bxdf":
                # MIXED (1 << 8) is shaded as diffuse (bxdf_partial.cl:32-34): the reference's mask has no define for it
                d += [name for bit, name in BXDF_DEFINES if self.material_types & bit]
            elif group == "logic":  # src/kernel_impl.hpp:51-63
                if self.denoiser: d.append("USE_OPTIX_DENOISER")
                if self.use_area_light: d.append("USE_AREA_LIGHT")
                if self.use_env_map: d.append("USE_ENV_MAP")
                if self.sample_expl: d.append("SAMPLE_EXPLICIT")
                if self.sample_impl: d.append("SAMPLE_IMPLICIT")
                if not self.separate_queues: d.append("WF_SINGLE_MAT_QUEUE")
                if self.check_spp: d.append("CHECK_SPP")
        return d


def _obj_path(kernel, defines):
    tag = hashlib.sha1((" ".join(defines)).encode()).hexdigest()[:10] if defines else "base"
    return os.path.join(OBJ, "%s.%s.so" % (kernel, tag))


def _compile(kernel, defines, path):
    cmd = [CXX] + CXXFLAGS + ["-shared", "-I", GEN, "-I", SRC] + GLOBAL_OPTS + ["-D" + d for d in defines] + \
          ["-DREFCL_NS=ns_" + kernel, "-DTU_" + kernel, os.path.join(SRC, "ref_cl_kernels.cpp"), "-o", path + ".tmp"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("refcl: building %s failed:\n%s" % (kernel, r.stderr[-4000:]))
    os.replace(path + ".tmp", path)


_LOADED = {}


def program(kernel, options):
    """The compiled 'program' for one kernel file under the given options (built on demand)."""
    defines = options.defines(kernel)
    path = _obj_path(kernel, defines)
    lib = _LOADED.get(path)
    if lib is None:
        if not os.path.exists(path) or _stale(path):
            if not have_reference():
                raise RuntimeError("refcl: %s with [%s] was not prebuilt and the reference tree is absent"
                                   % (kernel, " ".join(defines)))
            os.makedirs(OBJ, exist_ok=True)
            _generate()
            _compile(kernel, defines, path)
        lib = C.CDLL(path)
        _LOADED[path] = lib
    return lib


def _stale(path):
    """Rebuild when the shim / launchers are newer than the object (only checked where the reference exists)."""
    if not have_reference():
        return False
    t = os.path.getmtime(path)
    return any(os.path.getmtime(os.path.join(SRC, f)) > t for f in os.listdir(SRC))


def prebuild(options_list, kernels=None):
    """Compile (in parallel) every kernel for every option set; used by __graft_entry__.build()."""
    if not have_reference():
        return 0
    os.makedirs(OBJ, exist_ok=True)
    _generate()
    jobs = {}
    for o in options_list:
        for k in (kernels or KERNELS):
            d = o.defines(k)
            p = _obj_path(k, d)
            if not os.path.exists(p) or _stale(p):
                jobs[p] = (k, d)
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        list(ex.map(lambda it: _compile(it[1][0], it[1][1], it[0]), jobs.items()))
    return len(jobs)


# ---- buffers (CLContext::deviceBuffers in host memory) ---------------------------------------------------------
V = C.c_void_p


class Buffers(C.Structure):
    _fields_ = [("tasks", V), ("pixels", V), ("denoiserAlbedo", V), ("denoiserNormal", V), ("pixelsPreview", V),
                ("denoiserAlbedoGL", V), ("denoiserNormalGL", V), ("queueCounters", V), ("queues", V * 9),
                ("tris", V), ("nodes", V), ("indices", V), ("materials", V), ("texData", V), ("textures", V),
                ("envRGBA", V), ("envW", C.c_int32), ("envH", C.c_int32), ("probTable", V), ("aliasTable", V),
                ("pdfTable", V), ("params", V), ("samplesPerPixel", V), ("currPixelIdx", V), ("stats", V),
                ("pickResult", V), ("numTasks", C.c_uint32), ("numPixels", C.c_uint32), ("threads", C.c_int32)]


def _p(a):
    return a.ctypes.data if a is not None and a.size else None


COUNTER_NAMES = ("raygenQueue", "extensionQueue", "shadowQueue", "diffuseQueue", "glossyQueue", "ggxReflQueue",
                 "ggxRefrQueue", "deltaQueue", "emissiveQueue", "splattedSamples")
COUNTERS_DT = np.dtype([(n, np.uint32) for n in COUNTER_NAMES])
MAT_QUEUE_KERNELS = ("wf_mat_diffuse", "wf_mat_glossy", "wf_mat_ggx_reflection", "wf_mat_ggx_refraction",
                     "wf_mat_delta", "wf_mat_emissive")


class RefWavefront:
    """The reference's wavefront + microkernel renderers on the host: CLContext's buffers as numpy arrays,
    its enqueue* methods as calls of the compiled reference kernels.  Same surface as
    oracle.pyoracle.OracleWavefront so that a test can drive both side by side."""

    def __init__(self, num_tasks, width, height):
        self.N, self.width, self.height = int(num_tasks), int(width), int(height)
        npx = self.width * self.height
        self._state = np.zeros((64, self.N), np.uint32)
        self._queues = [np.zeros(self.N, np.uint32) for _ in range(9)]
        self._cnt = np.zeros(1, COUNTERS_DT)
        self._pixels = np.zeros((npx, 4), np.float32)
        self._albedo = np.zeros((npx, 4), np.float32)
        self._normal = np.zeros((npx, 4), np.float32)
        self._preview = np.zeros((npx, 4), np.float32)
        self._albedo_gl = np.zeros((npx, 4), np.float32)
        self._normal_gl = np.zeros((npx, 4), np.float32)
        self._spp = np.zeros(npx, np.uint32)
        self._pix_idx = np.zeros(1, np.uint32)
        self._stats = np.zeros(4, np.uint32)
        self._pick = np.zeros(64, np.uint8)
        self._params = np.zeros(256, np.uint8)
        self._scene = None
        self._env = None
        self.denoiser = False
        self.bitstack = False
        self.material_types = 0xFE
        self.threads = 1
        self.seed_offset = 0
        self._b = Buffers()
        self._opts = None

    # -- configuration ------------------------------------------------------------------------------------------
    def set_scene(self, scene):
        s = {k: np.ascontiguousarray(scene[k]) for k in ("tris", "indices", "nodes", "materials")}
        s["textures"] = np.ascontiguousarray(scene.get("textures", np.zeros(0, np.uint8)))
        s["texels"] = np.ascontiguousarray(scene.get("texels", np.zeros(0, np.uint8)))
        self._scene = s
        if "material_types" in scene:
            self.material_types = int(scene["material_types"])
        else:
            self.material_types = int(np.bitwise_or.reduce(s["materials"]["type"].astype(np.int64))) if len(s["materials"]) else 0xFE
        self._opts = None

    def set_envmap(self, env):
        h, w = env["rgb"].shape[:2]
        rgba = np.ones((h, w, 4), np.float32)
        rgba[..., :3] = env["rgb"]
        self._env = {"rgba": rgba, "prob": np.ascontiguousarray(env["prob"], np.float32),
                     "alias": np.ascontiguousarray(env["alias"], np.int32), "pdf": np.ascontiguousarray(env["pdf"], np.float32)}

    def set_params(self, params_bytes):
        buf = np.frombuffer(bytes(params_bytes), np.uint8)
        assert buf.size == 256
        self._params[:] = buf
        self._opts = None

    def set_options(self, bitstack=False, denoiser_aov=False, seed_offset=0):
        self.bitstack, self.denoiser, self.seed_offset = bool(bitstack), bool(denoiser_aov), int(seed_offset)
        self._opts = None

    def set_threads(self, n):
        self.threads = max(1, int(n))

    def options(self):
        if self._opts is None:
            from fluctus_b200.abi import PARAMS_DT
            p = self._params.view(PARAMS_DT)[0]
            self._opts = Options.from_params(p, self.denoiser, self.bitstack, self.material_types)
        return self._opts

    def _bind(self):
        b, s, e = self._b, self._scene, self._env
        b.tasks, b.pixels, b.denoiserAlbedo, b.denoiserNormal = _p(self._state), _p(self._pixels), _p(self._albedo), _p(self._normal)
        b.pixelsPreview, b.denoiserAlbedoGL, b.denoiserNormalGL = _p(self._preview), _p(self._albedo_gl), _p(self._normal_gl)
        b.queueCounters = _p(self._cnt)
        for i in range(9):
            b.queues[i] = _p(self._queues[i])
        if s is not None:
            b.tris, b.nodes, b.indices, b.materials = _p(s["tris"]), _p(s["nodes"]), _p(s["indices"]), _p(s["materials"])
            b.texData, b.textures = _p(s["texels"]), _p(s["textures"])
        if e is not None:
            b.envRGBA, b.envW, b.envH = _p(e["rgba"]), e["rgba"].shape[1], e["rgba"].shape[0]
            b.probTable, b.aliasTable, b.pdfTable = _p(e["prob"]), _p(e["alias"]), _p(e["pdf"])
        else:
            # CLContext::createEnvMap placeholder, src/clcontext.cpp:531-537: a 1x1 image
            if not hasattr(self, "_dummy_env"):
                self._dummy_env = np.zeros((1, 1, 4), np.float32)
                self._dummy_tab = np.zeros(1, np.float32)
            b.envRGBA, b.envW, b.envH = _p(self._dummy_env), 1, 1
            b.probTable = b.aliasTable = b.pdfTable = _p(self._dummy_tab)
        b.params, b.samplesPerPixel, b.currPixelIdx = _p(self._params), _p(self._spp), _p(self._pix_idx)
        b.stats, b.pickResult = _p(self._stats), _p(self._pick)
        b.numTasks, b.numPixels, b.threads = self.N, self.width * self.height, self.threads
        return C.byref(b)

    def _launch(self, kernel, *extra):
        lib = program(kernel, self.options())
        getattr(lib, "refcl_" + kernel)(self._bind(), *extra)

    # -- wavefront stages (CLContext::enqueueWf*, src/clcontext.cpp:783-873) -----------------------------------
    def reset(self):
        self._launch("wf_reset")
        if self.seed_offset:  # multi-GPU decorrelation — not in the reference, applied after its reset
            self._state[49] += np.uint32(self.seed_offset)

    def raygen(self): self._launch("wf_raygen")
    def logic(self, first_iteration=False): self._launch("wf_logic", C.c_uint32(int(first_iteration)))
    def ext(self): self._launch("wf_extrays")
    def shadow(self): self._launch("wf_shadowrays")

    def materials(self):
        from fluctus_b200.abi import PARAMS_DT
        if int(self._params.view(PARAMS_DT)[0]["wfSeparateQueues"]):  # src/clcontext.cpp:816-831
            for k in MAT_QUEUE_KERNELS:
                self._launch(k)
        else:
            self._launch("wf_mat_all")

    def clear_queues(self): self._cnt[:] = 0  # enqueueClearWfQueues, src/clcontext.cpp:903-909
    def update_pixel_index(self, n):  # src/clcontext.cpp:917-921
        self._pix_idx[0] = np.uint32((int(self._pix_idx[0]) + int(n)) % (self.width * self.height))
    def set_pixel_index(self, v): self._pix_idx[0] = np.uint32(v)
    def get_pixel_index(self): return int(self._pix_idx[0])

    def pick(self, ndc_x, ndc_y):
        from fluctus_b200.abi import HIT_DT
        self._launch("kernel_pick", C.c_float(ndc_x), C.c_float(ndc_y))
        return self._pick.view(HIT_DT)[0].copy()

    # -- microkernel renderer (src/clcontext.cpp:727-768) ---------------------------------------------------------
    def mk_reset(self): self._launch("mk_reset")
    def mk_raygen(self): self._launch("mk_raygen")
    def mk_next_vertex(self): self._launch("mk_next_vertex")
    def mk_sample_bsdf(self): self._launch("mk_sample_bsdf")
    def mk_splat(self): self._launch("mk_splat")
    def mk_splat_preview(self): self._launch("mk_splat_preview")
    def mk_stats(self): return self._stats

    def postprocess(self):
        self._launch("mk_postprocess")
        return self._preview.copy()

    def render(self, max_iterations, do_reset=True):
        """Tracer::runBenchmarkFromFile's loop (src/tracer.cpp:633-645, 725-778), as OracleWavefront.render."""
        from fluctus_b200.abi import PARAMS_DT
        p = self._params.view(PARAMS_DT)[0]
        target = int(p["width"]) * int(p["height"]) * int(p["maxSpp"])
        st = {"primary": 0, "extension": 0, "shadow": 0, "samples": 0}
        sample_count, it = 0, 0
        if do_reset:
            self.reset(); self.clear_queues(); self.set_pixel_index(0)
            self.raygen(); self.ext()
        while it < max_iterations and (target == 0 or sample_count < target):
            self.logic(False); self.raygen(); self.materials()
            cnt = self._cnt[0].copy()
            self.ext(); self.shadow(); self.clear_queues()
            st["primary"] += int(cnt["raygenQueue"]); st["extension"] += int(cnt["extensionQueue"]); st["shadow"] += int(cnt["shadowQueue"])
            st["samples"] += int(cnt["raygenQueue"]) if it > 0 else 0
            sample_count += int(cnt["splattedSamples"])
            self.update_pixel_index(int(cnt["raygenQueue"]))
            it += 1
        return it, st

    # -- views ----------------------------------------------------------------------------------------------------------
    def state(self): return self._state
    def queue(self, q): return self._queues[q]
    def counters(self): return self._cnt[0]
    def pixels(self): return self._pixels
    def spp(self): return self._spp
    def denoiser_albedo(self): return self._albedo
    def denoiser_normal(self): return self._normal
    def close(self): pass


# ---- ray batches through the reference's traversal functions ----------------------------------------------------------
def trace_closest(scene, rays, bitstack=False, want_hits=False, threads=1):
    from oracle.pyoracle import HIT_DT
    lib = program("wf_extrays", Options(bitstack=bitstack))
    rays = np.ascontiguousarray(rays, np.float32)
    n = rays.shape[0]
    hit_i, hit_t = np.empty(n, np.int32), np.empty(n, np.float32)
    hits = np.zeros(n, HIT_DT) if want_hits else None
    tris, nodes, idx = (np.ascontiguousarray(scene[k]) for k in ("tris", "nodes", "indices"))
    lib.refcl_trace_closest(V(_p(tris)), V(_p(nodes)), V(_p(idx)), V(_p(rays)), C.c_uint64(n), V(_p(hit_i)), V(_p(hit_t)),
                            V(_p(hits)) if want_hits else None, C.c_int(threads))
    out = {"hit_i": hit_i, "hit_t": hit_t}
    if want_hits:
        out["hits"] = hits
    return out


def trace_any(scene, rays, bitstack=False, threads=1):
    lib = program("wf_shadowrays", Options(bitstack=bitstack))
    rays = np.ascontiguousarray(rays, np.float32)
    n = rays.shape[0]
    occ = np.empty(n, np.uint8)
    tris, nodes, idx = (np.ascontiguousarray(scene[k]) for k in ("tris", "nodes", "indices"))
    lib.refcl_trace_any(V(_p(tris)), V(_p(nodes)), V(_p(idx)), V(_p(rays)), C.c_uint64(n), V(_p(occ)), C.c_int(threads))
    return occ


def rand_sequence(seed, count):
    lib = program("wf_extrays", Options())
    out = np.empty(count, np.float32)
    s = C.c_uint32(0)
    lib.refcl_rand_sequence(C.c_uint32(seed), C.c_uint32(count), V(_p(out)), C.byref(s))
    return out, s.value


def intersect_light(params_bytes, rays):
    from oracle.pyoracle import HIT_DT
    lib = program("wf_extrays", Options())
    rays = np.ascontiguousarray(rays, np.float32)
    p = np.frombuffer(bytes(params_bytes), np.uint8).copy()
    hits = np.zeros(rays.shape[0], HIT_DT)
    lib.refcl_intersect_light(V(_p(p)), V(_p(rays)), C.c_uint64(rays.shape[0]), V(_p(hits)))
    return hits


# Option sets whose programs must exist on a box without the reference tree: bench.py's workloads (kitchen stand-in:
# both lights, MIS, separate queues; cfg1: env map only, CHECK_SPP; the ray-cast batch) and smoke().
BENCH_OPTIONS = [
    Options(use_area_light=True, use_env_map=True, separate_queues=True),
    Options(use_area_light=True, use_env_map=True, separate_queues=True, check_spp=True),
    Options(use_area_light=False, use_env_map=True),
    Options(use_area_light=False, use_env_map=True, check_spp=True),
    Options(use_area_light=True, use_env_map=False),
]

if __name__ == "__main__":
    print("refcl: %d program(s) built" % prebuild(BENCH_OPTIONS))
