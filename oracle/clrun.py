"""TEST INFRASTRUCTURE — the reference's UNTOUCHED OpenCL kernels on an OpenCL device (the B200 itself).

The GPU box ships NVIDIA's OpenCL platform (libnvidia-opencl.so.1) without an ICD registration
(profiles/r2_opencl_probe*.txt); the CUDA toolkit's loader (libOpenCL.so.1) reaches it through
OCL_ICD_FILENAMES.  This module is a headless stand-in for the reference's CLContext (src/clcontext.cpp):
it builds the kernel files of /root/reference/src — mirrored byte for byte, git-ignored, into
baseline/_ref/src by __graft_entry__.build(), because the reference tree does not exist on the GPU box —
with the reference's own build options (src/clcontext.cpp:145-155, src/kernel_impl.hpp), binds the same
buffers and enqueues the same NDRanges (src/clcontext.cpp:727-873).  No GL, no window.

Uses: (1) parity target — CUDA results against the reference's real kernels on the same GPU
(tests/test_opencl_reference_gpu.py); (2) the same-GPU reference arm of bench.py (`reference_opencl`).
Only tests/, bench.py and tools/ may import this module; the product never does.
"""
import ctypes as C
import os
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
SRC_MIRROR = os.path.join(ROOT, "baseline", "_ref", "src")
VENDOR_LIBS = ("/usr/lib/libnvidia-opencl.so.1", "/usr/lib/x86_64-linux-gnu/libnvidia-opencl.so.1")

V = C.c_void_p
CL_DEVICE_TYPE_ALL = 0xFFFFFFFF
CL_MEM_READ_WRITE, CL_MEM_READ_ONLY, CL_MEM_COPY_HOST_PTR = 1, 4, 32
CL_QUEUE_PROFILING_ENABLE = 2
CL_PROGRAM_BUILD_LOG = 0x1183
CL_RGBA, CL_FLOAT = 0x10B5, 0x10DE
CL_PROFILING_COMMAND_START, CL_PROFILING_COMMAND_END = 0x1282, 0x1283
CL_DEVICE_NAME, CL_PLATFORM_VERSION, CL_DRIVER_VERSION = 0x102B, 0x0901, 0x102D

_RT = None


class CLError(RuntimeError):
    pass


def source_dir():
    ref = os.environ.get("FLUCTUS_REF", "/root/reference")
    if os.path.isdir(os.path.join(ref, "src")):
        return os.path.join(ref, "src")
    return SRC_MIRROR


def mirror_sources():
    """Copy the reference's device sources (inputs of the run-time OpenCL compiler) to baseline/_ref/src,
    which is git-ignored and travels to the GPU box with the snapshot.  Called by __graft_entry__.build()."""
    import shutil
    ref = os.path.join(os.environ.get("FLUCTUS_REF", "/root/reference"), "src")
    if not os.path.isdir(ref):
        return 0
    os.makedirs(SRC_MIRROR, exist_ok=True)
    n = 0
    for name in sorted(os.listdir(ref)):
        if name.endswith(".cl") or name in ("geom.h", "bxdf_types.h"):
            shutil.copyfile(os.path.join(ref, name), os.path.join(SRC_MIRROR, name))
            n += 1
    return n


class Runtime:
    """Platform, device, context and one in-order profiling queue (ext/CLT/src/utils.cpp:312)."""

    def __init__(self):
        vendor = next((p for p in VENDOR_LIBS if os.path.exists(p)), None)
        if vendor is None:
            raise CLError("no vendor OpenCL library on this box")
        os.environ.setdefault("OCL_ICD_FILENAMES", vendor)
        try:
            L = C.CDLL("libOpenCL.so.1")
        except OSError as e:
            raise CLError("libOpenCL.so.1 not loadable: %s" % e)
        self.L = L
        L.clCreateContext.restype = V
        L.clCreateCommandQueue.restype = V
        L.clCreateProgramWithSource.restype = V
        L.clCreateKernel.restype = V
        L.clCreateBuffer.restype = V
        L.clCreateImage2D.restype = V
        L.clCreateContext.argtypes = [V, C.c_uint, V, V, V, V]
        L.clCreateCommandQueue.argtypes = [V, V, C.c_ulong, V]
        L.clCreateProgramWithSource.argtypes = [V, C.c_uint, V, V, V]
        L.clBuildProgram.argtypes = [V, C.c_uint, V, C.c_char_p, V, V]
        L.clGetProgramBuildInfo.argtypes = [V, V, C.c_uint, C.c_size_t, V, V]
        L.clCreateKernel.argtypes = [V, C.c_char_p, V]
        L.clSetKernelArg.argtypes = [V, C.c_uint, C.c_size_t, V]
        L.clCreateBuffer.argtypes = [V, C.c_ulong, C.c_size_t, V, V]
        L.clCreateImage2D.argtypes = [V, C.c_ulong, V, C.c_size_t, C.c_size_t, C.c_size_t, V, V]
        L.clEnqueueNDRangeKernel.argtypes = [V, V, C.c_uint, V, V, V, C.c_uint, V, V]
        L.clEnqueueReadBuffer.argtypes = [V, V, C.c_uint, C.c_size_t, C.c_size_t, V, C.c_uint, V, V]
        L.clEnqueueWriteBuffer.argtypes = [V, V, C.c_uint, C.c_size_t, C.c_size_t, V, C.c_uint, V, V]
        L.clFinish.argtypes = [V]
        L.clReleaseMemObject.argtypes = [V]
        L.clReleaseEvent.argtypes = [V]
        L.clReleaseKernel.argtypes = [V]
        L.clReleaseProgram.argtypes = [V]
        L.clGetEventProfilingInfo.argtypes = [V, C.c_uint, C.c_size_t, V, V]
        L.clGetDeviceIDs.argtypes = [V, C.c_ulong, C.c_uint, V, V]
        L.clGetDeviceInfo.argtypes = [V, C.c_uint, C.c_size_t, V, V]
        L.clGetPlatformInfo.argtypes = [V, C.c_uint, C.c_size_t, V, V]
        n = C.c_uint(0)
        rc = L.clGetPlatformIDs(0, None, C.byref(n))
        if rc != 0 or n.value == 0:
            raise CLError("no OpenCL platform (clGetPlatformIDs rc %d)" % rc)
        plats = (V * n.value)()
        L.clGetPlatformIDs(n.value, plats, None)
        self.platform = V(plats[0])
        nd = C.c_uint(0)
        rc = L.clGetDeviceIDs(self.platform, CL_DEVICE_TYPE_ALL, 0, None, C.byref(nd))
        if rc != 0 or nd.value == 0:
            raise CLError("no OpenCL device (rc %d)" % rc)
        devs = (V * nd.value)()
        L.clGetDeviceIDs(self.platform, CL_DEVICE_TYPE_ALL, nd.value, devs, None)
        self._devs = (V * 1)(devs[int(os.environ.get("FLX_CL_DEVICE", "0"))])
        self.device = V(self._devs[0])
        err = C.c_int(0)
        self.context = V(L.clCreateContext(None, 1, self._devs, None, None, C.byref(err)))
        self._check(err.value, "clCreateContext")
        self.queue = V(L.clCreateCommandQueue(self.context, self.device, CL_QUEUE_PROFILING_ENABLE, C.byref(err)))
        self._check(err.value, "clCreateCommandQueue")
        self.device_name = self._info(L.clGetDeviceInfo, self.device, CL_DEVICE_NAME)
        self.driver = self._info(L.clGetDeviceInfo, self.device, CL_DRIVER_VERSION)
        self.platform_version = self._info(L.clGetPlatformInfo, self.platform, CL_PLATFORM_VERSION)
        self.is_nvidia = "NVIDIA" in self._info(L.clGetPlatformInfo, self.platform, 0x0902)  # platformIsNvidia, clcontext.cpp:151

    @staticmethod
    def _info(fn, obj, code):
        buf = C.create_string_buffer(1024)
        fn(obj, code, 1024, buf, None)
        return buf.value.decode(errors="replace")

    @staticmethod
    def _check(rc, what):
        if rc != 0:
            raise CLError("%s failed with OpenCL error %d" % (what, rc))

    # -- memory --------------------------------------------------------------------------------------------------
    def buffer(self, nbytes, host=None, read_only=False):
        err = C.c_int(0)
        flags = CL_MEM_READ_ONLY if read_only else CL_MEM_READ_WRITE
        ptr = None
        if host is not None:
            host = np.ascontiguousarray(host)
            if host.nbytes < 16:  # keep tiny buffers (counters, cursors) at the 16-byte minimum allocated below
                host = np.concatenate([host.reshape(-1).view(np.uint8), np.zeros(16 - host.nbytes, np.uint8)])
            nbytes = host.nbytes
            if nbytes:
                flags |= CL_MEM_COPY_HOST_PTR
                ptr = host.ctypes.data
        m = self.L.clCreateBuffer(self.context, flags, max(int(nbytes), 16), ptr, C.byref(err))
        self._check(err.value, "clCreateBuffer(%d bytes)" % nbytes)
        return V(m)

    def image2d_rgba32f(self, rgba):
        rgba = np.ascontiguousarray(rgba, np.float32)
        h, w = rgba.shape[:2]
        fmt = (C.c_uint * 2)(CL_RGBA, CL_FLOAT)
        err = C.c_int(0)
        m = self.L.clCreateImage2D(self.context, CL_MEM_READ_ONLY | CL_MEM_COPY_HOST_PTR, fmt, w, h, 0, rgba.ctypes.data, C.byref(err))
        self._check(err.value, "clCreateImage2D")
        return V(m)

    def write(self, mem, arr, offset=0):
        arr = np.ascontiguousarray(arr)
        if arr.nbytes:
            self._check(self.L.clEnqueueWriteBuffer(self.queue, mem, 1, offset, arr.nbytes, arr.ctypes.data, 0, None, None), "clEnqueueWriteBuffer")

    def read(self, mem, out, offset=0):
        assert out.flags["C_CONTIGUOUS"]
        if out.nbytes:
            self._check(self.L.clEnqueueReadBuffer(self.queue, mem, 1, offset, out.nbytes, out.ctypes.data, 0, None, None), "clEnqueueReadBuffer")
        return out

    def release(self, mem):
        if mem:
            self.L.clReleaseMemObject(mem)

    def finish(self):
        self._check(self.L.clFinish(self.queue), "clFinish")


def runtime():
    global _RT
    if _RT is None:
        _RT = Runtime()
    return _RT


def available():
    """(True, description) when an OpenCL device and the kernel sources are both reachable."""
    if not os.path.exists(os.path.join(source_dir(), "wf_extrays.cl")):
        return False, "reference kernel sources not mirrored (baseline/_ref/src)"
    try:
        rt = runtime()
    except CLError as e:
        return False, str(e)
    return True, "%s, %s, driver %s" % (rt.device_name, rt.platform_version, rt.driver)


# ---- programs and kernels (clt::Kernel, ext/CLT/src/Kernel.cpp; options of src/clcontext.cpp:145-155) -----------------
def global_build_options(mode="reference", bitstack=False, nvidia=True):
    """mode "reference": the reference's option string verbatim.  mode "strict": the same program under IEEE
    rules — no -cl-fast-relaxed-math / -cl-denorms-are-zero, correctly rounded divide/sqrt, FP_CONTRACT OFF
    (the pragma is prepended to the program text) — the semantics the bit-exact contract is stated in.
    mode "ieee": strict + native_recip / rsqrt as IEEE divisions."""
    o = ["-DGPU", "-I", source_dir()]
    if mode == "reference":
        o += ["-cl-denorms-are-zero", "-cl-fast-relaxed-math"]
    else:
        o += ["-cl-fp32-correctly-rounded-divide-sqrt"]
        if mode == "ieee":
            # additionally read the native_* / rsqrt built-ins (implementation-defined precision: rcp.approx on NVIDIA) as
            # the IEEE operations the oracle pins them to (o_math.h) — macros on the command line, the source is untouched
            o += ["-Dnative_recip(x)=(1.0f/(x))", "-Drsqrt(x)=(1.0f/sqrt(x))"]
    o += ["-cl-kernel-arg-info", "-DFLT_FLOAT_ATOMICS"]
    if bitstack:
        o += ["-DUSE_BITSTACK"]
    o += ["-DUSE_SOA"]
    if nvidia:
        o += ["-DNVIDIA"]
    return o


class Kernel:
    _cache = {}

    def __init__(self, file_name, kernel_name, defines=(), mode="reference", bitstack=False):
        rt = runtime()
        self.rt, self.name = rt, kernel_name
        opts = " ".join(global_build_options(mode, bitstack, rt.is_nvidia) + ["-D" + d for d in defines])
        key = (file_name, opts)
        hit = Kernel._cache.get(key)
        if hit is None:
            with open(os.path.join(source_dir(), file_name), "rb") as f:
                text = f.read()
            if mode != "reference":
                text = b"#pragma OPENCL FP_CONTRACT OFF\n" + text
            src = C.c_char_p(text)
            srcs = (C.c_char_p * 1)(src)
            lens = (C.c_size_t * 1)(len(text))
            err = C.c_int(0)
            prog = V(rt.L.clCreateProgramWithSource(rt.context, 1, srcs, lens, C.byref(err)))
            rt._check(err.value, "clCreateProgramWithSource(%s)" % file_name)
            t0 = time.time()
            rc = rt.L.clBuildProgram(prog, 1, rt._devs, opts.encode(), None, None)
            if rc != 0:
                n = C.c_size_t(0)
                rt.L.clGetProgramBuildInfo(prog, rt.device, CL_PROGRAM_BUILD_LOG, 0, None, C.byref(n))
                log = C.create_string_buffer(max(n.value, 1))
                rt.L.clGetProgramBuildInfo(prog, rt.device, CL_PROGRAM_BUILD_LOG, n.value, log, None)
                raise CLError("clBuildProgram(%s) failed (%d) with options [%s]:\n%s"
                              % (file_name, rc, opts, log.value.decode(errors="replace")[-6000:]))
            hit = (prog, time.time() - t0)
            Kernel._cache[key] = hit
            self.build_seconds = hit[1]
        else:
            self.build_seconds = 0.0
        self.program = hit[0]
        err = C.c_int(0)
        self.kernel = V(rt.L.clCreateKernel(self.program, kernel_name.encode(), C.byref(err)))  # own cl_kernel: args are per object
        rt._check(err.value, "clCreateKernel(%s)" % kernel_name)
        self.options = opts

    def set_args(self, *args):
        """cl_mem handles (c_void_p), or ctypes scalars."""
        keep = []
        for i, a in enumerate(args):
            if isinstance(a, V):
                box = V(a.value)
                keep.append(box)
                rc = self.rt.L.clSetKernelArg(self.kernel, i, C.sizeof(V), C.byref(box))
            else:
                keep.append(a)
                rc = self.rt.L.clSetKernelArg(self.kernel, i, C.sizeof(a), C.byref(a))
            self.rt._check(rc, "clSetKernelArg(%s, %d)" % (self.name, i))
        self._keep = keep

    def enqueue(self, global_size, timed=False):
        """NullRange local size, as every launch of the reference.  Returns device ms when timed."""
        gs = global_size if isinstance(global_size, (tuple, list)) else (global_size,)
        g = (C.c_size_t * len(gs))(*[int(x) for x in gs])
        ev = V(0)
        rc = self.rt.L.clEnqueueNDRangeKernel(self.rt.queue, self.kernel, len(gs), None, g, None, 0, None, C.byref(ev) if timed else None)
        self.rt._check(rc, "clEnqueueNDRangeKernel(%s, %s)" % (self.name, gs))
        if not timed:
            return None
        self.rt.finish()
        t0, t1 = C.c_ulong(0), C.c_ulong(0)
        self.rt.L.clGetEventProfilingInfo(ev, CL_PROFILING_COMMAND_START, 8, C.byref(t0), None)
        self.rt.L.clGetEventProfilingInfo(ev, CL_PROFILING_COMMAND_END, 8, C.byref(t1), None)
        self.rt.L.clReleaseEvent(ev)
        return (t1.value - t0.value) * 1e-6


# ---- the wavefront renderer on the OpenCL device (CLContext + the loop of Tracer, headless) ------------------------------
U32 = C.c_uint32
QUEUE_NAMES = ("raygen", "extension", "shadow", "diffuse", "glossy", "ggxRefl", "ggxRefr", "delta", "emissive")
MAT_KERNELS = (("wf_mat_diffuse.cl", "wavefrontDiffuse", 3), ("wf_mat_glossy.cl", "wavefrontGlossy", 4),
               ("wf_mat_ggx_reflection.cl", "wavefrontGGXReflection", 5), ("wf_mat_ggx_refraction.cl", "wavefrontGGXRefraction", 6),
               ("wf_mat_delta.cl", "wavefrontDelta", 7), ("wf_mat_emissive.cl", "wavefrontEmissive", 8))


class CLWavefront:
    """Same surface as oracle.refcl.RefWavefront / pyoracle.OracleWavefront, kernels on the OpenCL device."""

    def __init__(self, num_tasks, width, height, mode="reference"):
        from oracle.refcl import COUNTERS_DT
        self.rt = runtime()
        self.mode = mode
        self.N, self.width, self.height = int(num_tasks), int(width), int(height)
        self.npx = self.width * self.height
        rt = self.rt
        self.COUNTERS_DT = COUNTERS_DT
        self.d_tasks = rt.buffer(256 * self.N)
        self.d_queues = [rt.buffer(4 * self.N) for _ in range(9)]
        self.d_counters = rt.buffer(40, np.zeros(10, np.uint32))
        self.d_pixels = rt.buffer(16 * self.npx)
        self.d_albedo = rt.buffer(16 * self.npx)
        self.d_normal = rt.buffer(16 * self.npx)
        self.d_preview = rt.buffer(16 * self.npx)
        self.d_albedo_gl = rt.buffer(16 * self.npx)
        self.d_normal_gl = rt.buffer(16 * self.npx)
        self.d_spp = rt.buffer(4 * self.npx, np.zeros(self.npx, np.uint32))
        self.d_pix_idx = rt.buffer(4, np.zeros(1, np.uint32))
        self.d_params = rt.buffer(256)
        self.d_tris = self.d_nodes = self.d_indices = self.d_materials = self.d_texdata = self.d_textures = None
        # CLContext::createEnvMap placeholder (src/clcontext.cpp:531-537)
        self.d_env = rt.image2d_rgba32f(np.zeros((1, 1, 4), np.float32))
        self.d_prob = rt.buffer(16); self.d_alias = rt.buffer(16); self.d_pdf = rt.buffer(16)
        self._params = np.zeros(256, np.uint8)
        self._pix_idx = 0
        self.denoiser = False
        self.bitstack = False
        self.material_types = 0xFE
        self._kernels = {}
        self.kernel_ms = {}     # accumulated device time per kernel name (when timing is on)
        self.timing = False
        self.build_seconds = 0.0

    # -- configuration (uploadSceneData / createEnvMap / updateParams) -------------------------------------------------------
    def set_scene(self, scene):
        rt = self.rt
        for m in (self.d_tris, self.d_nodes, self.d_indices, self.d_materials, self.d_texdata, self.d_textures):
            rt.release(m)
        self.d_tris = rt.buffer(0, scene["tris"], read_only=True)
        self.d_nodes = rt.buffer(0, scene["nodes"], read_only=True)
        self.d_indices = rt.buffer(0, np.ascontiguousarray(scene["indices"], np.uint32), read_only=True)
        self.d_materials = rt.buffer(0, scene["materials"], read_only=True)
        tex = scene.get("textures", None)
        texels = scene.get("texels", None)
        self.d_textures = rt.buffer(16, tex if tex is not None and len(tex) else None, read_only=True)
        self.d_texdata = rt.buffer(16, texels if texels is not None and len(texels) else None, read_only=True)
        if "material_types" in scene:
            self.material_types = int(scene["material_types"])
        elif len(scene["materials"]):
            self.material_types = int(np.bitwise_or.reduce(scene["materials"]["type"].astype(np.int64)))
        self._kernels.clear()

    def set_envmap(self, env):
        rt = self.rt
        h, w = env["rgb"].shape[:2]
        rgba = np.ones((h, w, 4), np.float32)
        rgba[..., :3] = env["rgb"]
        for m in (self.d_env, self.d_prob, self.d_alias, self.d_pdf):
            rt.release(m)
        self.d_env = rt.image2d_rgba32f(rgba)
        self.d_prob = rt.buffer(0, np.ascontiguousarray(env["prob"], np.float32), read_only=True)
        self.d_alias = rt.buffer(0, np.ascontiguousarray(env["alias"], np.int32), read_only=True)
        self.d_pdf = rt.buffer(0, np.ascontiguousarray(env["pdf"], np.float32), read_only=True)
        self._kernels.clear()

    def set_params(self, params_bytes):
        buf = np.frombuffer(bytes(params_bytes), np.uint8)
        assert buf.size == 256
        self._params = buf.copy()
        self.rt.write(self.d_params, self._params)
        self._kernels.clear()  # recompileKernels: options derive from RenderParams (src/clcontext.cpp:877-900)

    def set_options(self, bitstack=False, denoiser_aov=False, seed_offset=0):
        assert seed_offset == 0
        self.bitstack, self.denoiser = bool(bitstack), bool(denoiser_aov)
        self._kernels.clear()

    def _p(self):
        from fluctus_b200.abi import PARAMS_DT
        return self._params.view(PARAMS_DT)[0]

    def _kernel(self, file_name, kernel_name, group):
        k = self._kernels.get(kernel_name)
        if k is None:
            from oracle.refcl import Options, KERNELS
            o = Options.from_params(self._p(), self.denoiser, self.bitstack, self.material_types)
            defines = [d for d in o.defines(file_name[:-3]) if d != "USE_BITSTACK"]  # bit-stack is a global option
            k = Kernel(file_name, kernel_name, defines, self.mode, self.bitstack)
            self.build_seconds += k.build_seconds
            self._bind(k, kernel_name)
            self._kernels[kernel_name] = k
        return k

    def _bind(self, k, name):
        """Argument lists in the order of the kernels' signatures (src/wf_*.cl, src/mk_postprocess.cl)."""
        q = self.d_queues
        N = U32(self.N)
        if name == "reset":
            k.set_args(self.d_tasks, self.d_pixels, self.d_albedo, self.d_normal, self.d_counters, q[0], self.d_params, self.d_spp, N)
        elif name == "genRays":
            k.set_args(self.d_tasks, self.d_params, self.d_counters, q[0], q[1], self.d_pix_idx, N)
        elif name == "logic":
            k.set_args(self.d_tasks, self.d_pixels, self.d_normal, self.d_albedo, self.d_counters, q[1], q[2], q[0],
                       q[3], q[4], q[5], q[6], q[7], q[8], self.d_tris, self.d_nodes, self.d_indices, self.d_env,
                       self.d_prob, self.d_alias, self.d_pdf, self.d_materials, self.d_texdata, self.d_textures,
                       self.d_params, self.d_spp, N, U32(0))
        elif name == "traceExtension":
            k.set_args(self.d_tasks, self.d_counters, q[1], self.d_tris, self.d_nodes, self.d_indices, self.d_params, N)
        elif name == "traceShadow":
            k.set_args(self.d_tasks, self.d_counters, q[2], self.d_tris, self.d_nodes, self.d_indices, self.d_params, N)
        elif name == "wavefrontAllMaterials":
            k.set_args(self.d_tasks, self.d_counters, q[3], q[1], self.d_materials, self.d_texdata, self.d_textures, self.d_params, N)
        elif name == "process":
            k.set_args(self.d_pixels, self.d_albedo, self.d_normal, self.d_preview, self.d_albedo_gl, self.d_normal_gl, self.d_params, N)
        else:
            for f, kn, qi in MAT_KERNELS:
                if kn == name:
                    k.set_args(self.d_tasks, self.d_counters, q[qi], q[1], self.d_materials, self.d_texdata, self.d_textures, self.d_params, N)
                    return
            raise KeyError(name)

    def _run(self, file_name, kernel_name, global_size):
        k = self._kernel(file_name, kernel_name, None)
        ms = k.enqueue(global_size, timed=self.timing)
        if ms is not None:
            t = self.kernel_ms.setdefault(kernel_name, [0.0, 0])
            t[0] += ms; t[1] += 1

    # -- stages: CLContext::enqueueWf*Kernel (src/clcontext.cpp:783-873), global sizes as there ------------------------------
    def reset(self): self._run("wf_reset.cl", "reset", max(self.N, self.npx))
    def raygen(self): self._run("wf_raygen.cl", "genRays", self.N)
    def ext(self): self._run("wf_extrays.cl", "traceExtension", self.N)
    def shadow(self): self._run("wf_shadowrays.cl", "traceShadow", self.N)

    def logic(self, first_iteration=False):
        k = self._kernel("wf_logic.cl", "logic", None)
        fi = U32(int(first_iteration))
        self.rt._check(self.rt.L.clSetKernelArg(k.kernel, 27, 4, C.byref(fi)), "clSetKernelArg(logic, firstIteration)")
        self._run("wf_logic.cl", "logic", (self.N + 31) & ~31)

    def materials(self):
        if int(self._p()["wfSeparateQueues"]):
            for f, kn, _ in MAT_KERNELS:
                self._run(f, kn, self.N)
        else:
            self._run("wf_mat_all.cl", "wavefrontAllMaterials", self.N)

    def clear_queues(self):  # enqueueClearWfQueues: host write of zeroed counters (src/clcontext.cpp:903-909)
        self.rt.write(self.d_counters, np.zeros(10, np.uint32))

    def update_pixel_index(self, n):  # src/clcontext.cpp:917-921
        self._pix_idx = (self._pix_idx + int(n)) % self.npx
        self.rt.write(self.d_pix_idx, np.array([self._pix_idx], np.uint32))

    def set_pixel_index(self, v):
        self._pix_idx = int(v)
        self.rt.write(self.d_pix_idx, np.array([self._pix_idx], np.uint32))

    def get_pixel_index(self): return self._pix_idx

    def postprocess(self):
        self._run("mk_postprocess.cl", "process", self.npx)
        return self.rt.read(self.d_preview, np.empty((self.npx, 4), np.float32))

    # -- data access ------------------------------------------------------------------------------------------------------------
    def state(self): return self.rt.read(self.d_tasks, np.empty((64, self.N), np.uint32))
    def write_state(self, arr): self.rt.write(self.d_tasks, np.ascontiguousarray(arr, np.uint32))
    def queue(self, qi): return self.rt.read(self.d_queues[qi], np.empty(self.N, np.uint32))
    def write_queue(self, qi, arr): self.rt.write(self.d_queues[qi], np.ascontiguousarray(arr, np.uint32))
    def counters(self): return self.rt.read(self.d_counters, np.zeros(1, self.COUNTERS_DT))[0]
    def write_counters(self, rec): self.rt.write(self.d_counters, np.array([rec], self.COUNTERS_DT))
    def pixels(self): return self.rt.read(self.d_pixels, np.empty((self.npx, 4), np.float32))
    def write_pixels(self, arr): self.rt.write(self.d_pixels, np.ascontiguousarray(arr, np.float32))
    def spp(self): return self.rt.read(self.d_spp, np.empty(self.npx, np.uint32))
    def write_spp(self, arr): self.rt.write(self.d_spp, np.ascontiguousarray(arr, np.uint32))
    def denoiser_albedo(self): return self.rt.read(self.d_albedo, np.empty((self.npx, 4), np.float32))
    def denoiser_normal(self): return self.rt.read(self.d_normal, np.empty((self.npx, 4), np.float32))

    def render(self, max_iterations, do_reset=True, postprocess_every_iteration=False):
        """Tracer::runBenchmarkFromFile's loop (src/tracer.cpp:633-645, 725-778).  One clFinish + counter read per
        iteration, like the reference (src/tracer.cpp:333,336); postprocess per iteration is the viewer's habit."""
        p = self._p()
        target = int(p["width"]) * int(p["height"]) * int(p["maxSpp"])
        st = {"primary": 0, "extension": 0, "shadow": 0, "samples": 0}
        sample_count, it = 0, 0
        if do_reset:
            self.reset(); self.clear_queues(); self.set_pixel_index(0)
            self.raygen(); self.ext()
        while it < max_iterations and (target == 0 or sample_count < target):
            self.logic(False); self.raygen(); self.materials()
            cnt = self.counters()  # enqueueGetCounters + finishQueue
            self.ext(); self.shadow(); self.clear_queues()
            if postprocess_every_iteration:
                self._run("mk_postprocess.cl", "process", self.npx)
            st["primary"] += int(cnt["raygenQueue"]); st["extension"] += int(cnt["extensionQueue"]); st["shadow"] += int(cnt["shadowQueue"])
            st["samples"] += int(cnt["raygenQueue"]) if it > 0 else 0
            sample_count += int(cnt["splattedSamples"])
            self.update_pixel_index(int(cnt["raygenQueue"]))
            it += 1
        self.rt.finish()
        return it, st

    def close(self):
        rt = self.rt
        rt.finish()
        for m in [self.d_tasks, self.d_counters, self.d_pixels, self.d_albedo, self.d_normal, self.d_preview, self.d_albedo_gl,
                  self.d_normal_gl, self.d_spp, self.d_pix_idx, self.d_params, self.d_tris, self.d_nodes, self.d_indices,
                  self.d_materials, self.d_texdata, self.d_textures, self.d_env, self.d_prob, self.d_alias, self.d_pdf] + self.d_queues:
            rt.release(m)
        self.d_tasks = None


# ---- ray batches through the reference's own ray kernels -----------------------------------------------------------------
def _batch_context(scene, n, mode, bitstack):
    from fluctus_b200 import abi
    wf = CLWavefront(n, 1, 1, mode)
    wf.set_scene(scene)
    # sampleImpl && useAreaLight gate the light quad (wf_extrays.cl:28, wf_shadowrays.cl:32): off for plain ray casts
    p = abi.make_render_params(1, 1, n_tris=len(scene["tris"]), use_area_light=False, use_env_map=False)
    wf.set_params(p.tobytes())
    wf.set_options(bitstack=bitstack)
    return wf


def trace_closest(scene, rays, mode="reference", bitstack=False, want_hits=False, chunk=1 << 22):
    """Closest hits of a ray batch by the reference's `traceExtension` kernel (src/wf_extrays.cl:5-36): the rays are
    laid into the orig / dir slots of a GPUTaskState buffer, the extension queue is the identity, and the hit slots
    are read back.  tmax is FLT_MAX there (EMPTY_HIT(FLT_MAX)), as in cfg5.  Returns dict(hit_i, hit_t[, hits], ms)."""
    from fluctus_b200.abi import SLOTS, HIT_DT
    rays = np.ascontiguousarray(rays, np.float32)
    n_all = rays.shape[0]
    hit_i, hit_t = np.empty(n_all, np.int32), np.empty(n_all, np.float32)
    hits = np.zeros(n_all, HIT_DT) if want_hits else None
    n = min(chunk, n_all)
    wf = _batch_context(scene, n, mode, bitstack)
    total_ms = 0.0
    try:
        ident = np.arange(n, dtype=np.uint32)
        wf.write_queue(1, ident)
        for a in range(0, n_all, n):
            m = min(n, n_all - a)
            st = np.zeros((64, n), np.uint32)
            st[0:3, :m] = rays[a:a + m, 0:3].T.view(np.uint32)
            st[4:7, :m] = rays[a:a + m, 4:7].T.view(np.uint32)
            wf.write_state(st)
            cnt = np.zeros(1, wf.COUNTERS_DT); cnt["extensionQueue"] = m
            wf.write_counters(cnt[0])
            k = wf._kernel("wf_extrays.cl", "traceExtension", None)
            total_ms += k.enqueue(n, timed=True)
            out = wf.state()
            hit_i[a:a + m] = out[SLOTS["i"], :m].view(np.int32)
            hit_t[a:a + m] = out[SLOTS["t"], :m].view(np.float32)
            if want_hits:
                h = hits[a:a + m]
                h["P"][:, :3] = out[SLOTS["P"]:SLOTS["P"] + 3, :m].T.view(np.float32)
                h["N"][:, :3] = out[SLOTS["N"]:SLOTS["N"] + 3, :m].T.view(np.float32)
                h["uv"] = out[SLOTS["uvTex"]:SLOTS["uvTex"] + 2, :m].T.view(np.float32)
                h["t"] = hit_t[a:a + m]; h["i"] = hit_i[a:a + m]
                h["areaLightHit"] = out[SLOTS["areaLightHit"], :m].view(np.int32)
                h["matId"] = out[SLOTS["matId"], :m].view(np.int32)
    finally:
        wf.close()
    r = {"hit_i": hit_i, "hit_t": hit_t, "ms": total_ms}
    if want_hits:
        r["hits"] = hits
    return r


def trace_any(scene, rays, mode="reference", bitstack=False, chunk=1 << 22):
    """Any-hit flags by the reference's `traceShadow` kernel (src/wf_shadowrays.cl:6-39); tmax = rays[:, 3]."""
    from fluctus_b200.abi import SLOTS
    rays = np.ascontiguousarray(rays, np.float32)
    n_all = rays.shape[0]
    occ = np.empty(n_all, np.uint8)
    n = min(chunk, n_all)
    wf = _batch_context(scene, n, mode, bitstack)
    total_ms = 0.0
    try:
        wf.write_queue(2, np.arange(n, dtype=np.uint32))
        for a in range(0, n_all, n):
            m = min(n, n_all - a)
            st = np.zeros((64, n), np.uint32)
            so, sd = SLOTS["shadowOrig"], SLOTS["shadowDir"]
            st[so:so + 3, :m] = rays[a:a + m, 0:3].T.view(np.uint32)
            st[sd:sd + 3, :m] = rays[a:a + m, 4:7].T.view(np.uint32)
            st[SLOTS["shadowRayLen"], :m] = rays[a:a + m, 3].view(np.uint32)
            wf.write_state(st)
            cnt = np.zeros(1, wf.COUNTERS_DT); cnt["shadowQueue"] = m
            wf.write_counters(cnt[0])
            k = wf._kernel("wf_shadowrays.cl", "traceShadow", None)
            total_ms += k.enqueue(n, timed=True)
            occ[a:a + m] = wf.state()[SLOTS["shadowRayBlocked"], :m] != 0
    finally:
        wf.close()
    return occ, total_ms
