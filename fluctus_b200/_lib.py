"""ctypes loader for libfluctus_cuda.so (the C-ABI of include/fluctus_cuda.h).

There is deliberately no fallback: if the library is missing or no sm_100 device is visible the
caller gets an exception, never a CPU path."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# FLX_LIB_VARIANT=exactlibm selects the TEST-ONLY build with double-precision transcendental functions (csrc/flx_libm.cuh)
LIB_PATH = os.path.join(_HERE, "libfluctus_cuda_%s.so" % os.environ["FLX_LIB_VARIANT"] if os.environ.get("FLX_LIB_VARIANT") else "libfluctus_cuda.so")
_lib = None

V, U32, U64, I32, SZ, F32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_size_t, C.c_float

# name -> (restype, argtypes); every symbol include/fluctus_cuda.h declares
SIGNATURES = {
    "flx_create": (I32, [C.POINTER(V), I32, U32, V]),
    "flx_destroy": (None, [V]),
    "flx_last_error": (C.c_char_p, [V]),
    "flx_num_tasks": (U32, [V]),
    "flx_stream": (V, [V]),
    "flx_upload_scene": (I32, [V, V, SZ, V, SZ, V, SZ, V, SZ, V, SZ, V, SZ]),
    "flx_upload_envmap": (I32, [V, V, I32, I32, V, V, V]),
    "flx_resize": (I32, [V, U32, U32]),
    "flx_update_params": (I32, [V, V]),
    "flx_set_options": (I32, [V, I32, U32]),
    "flx_set_pixel_sharding": (I32, [V, U32, U32]),
    "flx_set_traversal": (I32, [V, I32]),
    "flx_reset_pixel_index": (I32, [V]),
    "flx_set_pixel_index": (I32, [V, U32]),
    "flx_update_pixel_index": (I32, [V, U32, U32]),
    "flx_wf_reset": (I32, [V]),
    "flx_wf_raygen": (I32, [V]),
    "flx_wf_logic": (I32, [V, I32]),
    "flx_wf_materials": (I32, [V]),
    "flx_wf_ext": (I32, [V]),
    "flx_wf_shadow": (I32, [V]),
    "flx_wf_clear_queues": (I32, [V]),
    "flx_get_counters": (I32, [V, V]),
    "flx_postprocess": (I32, [V]),
    "flx_finish": (I32, [V]),
    "flx_render": (I32, [V, U32, I32]),
    "flx_build_lbvh": (I32, [I32, V, SZ, U32, V, SZ, V, V, V]),
    "flx_reduce_frames": (I32, [V, I32]),
    "flx_comm_unique_id": (I32, [V]),
    "flx_comm_init": (I32, [V, I32, I32, V]),
    "flx_comm_destroy": (I32, [V]),
    "flx_comm_info": (I32, [V, V, V, V]),
    "flx_allreduce_pixels": (I32, [V]),
    "flx_reduce_pixels": (I32, [V, I32]),
    "flx_mk_reset": (I32, [V]),
    "flx_mk_raygen": (I32, [V]),
    "flx_mk_next_vertex": (I32, [V]),
    "flx_mk_sample_bsdf": (I32, [V]),
    "flx_mk_splat": (I32, [V]),
    "flx_mk_splat_preview": (I32, [V]),
    "flx_mk_get_stats": (I32, [V, V]),
    "flx_mk_render": (I32, [V, U32, I32]),
    "flx_get_render_totals": (I32, [V, V]),
    "flx_read_pixels": (I32, [V, I32, V, SZ]),
    "flx_write_pixels": (I32, [V, I32, V, SZ]),
    "flx_device_ptr": (V, [V, I32]),
    "flx_pick": (I32, [V, F32, F32, V]),
    "flx_trace_closest": (I32, [V, V, SZ, V, V, V]),
    "flx_trace_any": (I32, [V, V, SZ, V]),
    "flx_trace_closest_device": (I32, [V, V, SZ, V, V, V]),
    "flx_trace_closest_static": (I32, [V, V, SZ, V, V]),
    "flx_trace_any_device": (I32, [V, V, SZ, V]),
    "flx_export_state": (I32, [V, V]),
    "flx_import_state": (I32, [V, V]),
    "flx_read_queue": (I32, [V, I32, V, SZ]),
    "flx_write_queue": (I32, [V, I32, V, SZ]),
    "flx_set_counters": (I32, [V, V]),
    "flx_launch_count": (U64, [V]),
    "flx_enable_timing": (I32, [V, I32]),
    "flx_get_kernel_time": (I32, [V, I32, V, V]),
    "flx_set_trace_counting": (I32, [V, I32]),
    "flx_get_trace_counters": (I32, [V, V]),
}


class FluctusError(RuntimeError):
    """Raised for any non-zero flx_status (the reference throws std::runtime_error, clcontext.cpp:952-957)."""


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FluctusError(
                "libfluctus_cuda.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C fluctus_b200/csrc`); there is no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError here means header and library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
