#!/bin/bash
# One-off probe of the GPU box for an OpenCL platform (compute API, not graphics): VERDICT r1 item 1(a).
# Output is committed under profiles/r2_opencl_probe.txt.
out=gpurun_out/opencl_probe.txt
{
echo "== /etc/OpenCL/vendors"; ls -la /etc/OpenCL/vendors 2>&1; cat /etc/OpenCL/vendors/*.icd 2>&1
echo "== ldconfig"; ldconfig -p | grep -i -E 'opencl|pocl' 2>&1
echo "== find libs"; find / -xdev \( -name 'libnvidia-opencl*' -o -name 'libpocl*' -o -name 'libOpenCL*' -o -name '*.icd' -o -name 'libintelocl*' \) 2>/dev/null
echo "== nvidia libs dir"; ls /usr/lib/x86_64-linux-gnu | grep -i -E 'nvidia|cuda' | head -60
echo "== nvidia-smi"; nvidia-smi -L; nproc
echo "== python probe"
python - <<'PY'
import ctypes, glob
cands = ["libOpenCL.so.1", "/usr/local/cuda/targets/x86_64-linux/lib/libOpenCL.so.1"] + glob.glob("/usr/lib/x86_64-linux-gnu/libnvidia-opencl.so*")
for c in cands:
    try:
        lib = ctypes.CDLL(c)
    except OSError as e:
        print(c, "dlopen failed:", e); continue
    try:
        n = ctypes.c_uint(0)
        rc = lib.clGetPlatformIDs(0, None, ctypes.byref(n))
        print(c, "clGetPlatformIDs rc", rc, "platforms", n.value)
        if rc == 0 and n.value:
            ids = (ctypes.c_void_p * n.value)()
            lib.clGetPlatformIDs(n.value, ids, None)
            for p in ids:
                for what, code in (("name", 0x0902), ("vendor", 0x0903), ("version", 0x0901)):
                    buf = ctypes.create_string_buffer(256)
                    lib.clGetPlatformInfo(ctypes.c_void_p(p), code, 256, buf, None)
                    print("   ", what, buf.value.decode())
                nd = ctypes.c_uint(0)
                rc = lib.clGetDeviceIDs(ctypes.c_void_p(p), ctypes.c_ulong(0xFFFFFFFF), 0, None, ctypes.byref(nd))
                print("    devices rc", rc, "n", nd.value)
    except Exception as e:
        print(c, "error", e)
PY
} > $out 2>&1
cat $out
