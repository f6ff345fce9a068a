"""Second probe of the GPU box: reach libnvidia-opencl.so.1 although no /etc/OpenCL/vendors exists.
Routes tried: (1) OCL_ICD_FILENAMES, (2) OCL_ICD_VENDORS pointing into the repo, (3) dlopen of the
vendor library + clIcdGetPlatformIDsKHR + the ICD dispatch table."""
import ctypes as C, os, subprocess, sys

NV = "/usr/lib/libnvidia-opencl.so.1"

def platform_report(lib, get_ids):
    n = C.c_uint(0)
    rc = get_ids(0, None, C.byref(n))
    print("  platform ids rc", rc, "n", n.value)
    if rc != 0 or n.value == 0:
        return None
    ids = (C.c_void_p * n.value)()
    get_ids(n.value, ids, None)
    return ids[0]

if len(sys.argv) > 1 and sys.argv[1] == "loader":
    lib = C.CDLL("libOpenCL.so.1")
    p = platform_report(lib, lib.clGetPlatformIDs)
    if p:
        buf = C.create_string_buffer(256)
        lib.clGetPlatformInfo(C.c_void_p(p), 0x0902, 256, buf, None)
        print("  platform name:", buf.value.decode())
    sys.exit(0)

print("== nm -D", NV)
print(subprocess.run("nm -D --defined-only %s | head -40" % NV, shell=True, capture_output=True, text=True).stdout)
for var, val in (("OCL_ICD_FILENAMES", NV), ("OCL_ICD_VENDORS", os.path.abspath("tools/cl_icd"))):
    print("== loader with %s=%s" % (var, val))
    env = dict(os.environ); env[var] = val
    r = subprocess.run([sys.executable, __file__, "loader"], env=env, capture_output=True, text=True)
    print(r.stdout, r.stderr[-500:])

print("== direct dlopen")
nv = C.CDLL(NV)
nv.clGetExtensionFunctionAddress.restype = C.c_void_p
addr = nv.clGetExtensionFunctionAddress(b"clIcdGetPlatformIDsKHR")
print("  clIcdGetPlatformIDsKHR at", addr)
proto = C.CFUNCTYPE(C.c_int, C.c_uint, C.c_void_p, C.POINTER(C.c_uint))
get_ids = proto(addr)
plat = platform_report(nv, get_ids)
if plat:
    table = C.cast(C.cast(plat, C.POINTER(C.c_void_p))[0], C.POINTER(C.c_void_p))
    clGetPlatformInfo = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint, C.c_size_t, C.c_void_p, C.c_void_p)(table[1])
    clGetDeviceIDs = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_ulong, C.c_uint, C.c_void_p, C.POINTER(C.c_uint))(table[2])
    clGetDeviceInfo = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint, C.c_size_t, C.c_void_p, C.c_void_p)(table[3])
    for what, code in (("name", 0x0902), ("vendor", 0x0903), ("version", 0x0901), ("extensions", 0x0904)):
        buf = C.create_string_buffer(4096)
        rc = clGetPlatformInfo(C.c_void_p(plat), code, 4096, buf, None)
        print("  platform", what, rc, buf.value.decode()[:300])
    nd = C.c_uint(0)
    rc = clGetDeviceIDs(C.c_void_p(plat), 0xFFFFFFFF, 0, None, C.byref(nd))
    print("  devices rc", rc, "n", nd.value)
    if nd.value:
        devs = (C.c_void_p * nd.value)()
        clGetDeviceIDs(C.c_void_p(plat), 0xFFFFFFFF, nd.value, devs, None)
        for what, code in (("name", 0x102B), ("version", 0x102F), ("driver", 0x102D), ("opencl_c", 0x103D)):
            buf = C.create_string_buffer(1024)
            rc = clGetDeviceInfo(C.c_void_p(devs[0]), code, 1024, buf, None)
            print("  device", what, rc, buf.value.decode()[:200])
        cu = C.c_uint(0); clGetDeviceInfo(C.c_void_p(devs[0]), 0x1002, 4, C.byref(cu), None); print("  compute units", cu.value)
        img = C.c_uint(0); clGetDeviceInfo(C.c_void_p(devs[0]), 0x1016, 4, C.byref(img), None); print("  image support", img.value)
