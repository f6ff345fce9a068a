"""CPU-only checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/fluctus_cuda.h declares, the boundary structs have the reference's sizes/offsets, and the
product fails loudly (no CPU fallback) when there is no GPU.  No compute calls are made here."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "fluctus_cuda.h")
ABI_HEADER = os.path.join(ROOT, "include", "fluctus_abi.h")


def _declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(flx_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import fluctus_b200 as fx
    lib = fx.load()
    names = _declared_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), "libfluctus_cuda.so does not export %s" % n
    # and the ctypes signature table covers exactly the header
    assert sorted(fx.SIGNATURES) == names


def test_dynamic_symbol_table_has_c_linkage():
    import fluctus_b200 as fx
    out = subprocess.check_output(["nm", "-D", "--defined-only", fx.LIB_PATH], text=True)
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
    for n in _declared_symbols():
        assert n in exported, n  # unmangled => extern "C"


def test_struct_layouts_match_reference_abi(tmp_path):
    """Compile the plain-C header with gcc and compare sizeof/offsetof with SURVEY.md §8(a)."""
    src = tmp_path / "layout.c"
    src.write_text(r'''
#include <stdio.h>
#include <stddef.h>
#include "fluctus_cuda.h"
int main(void) {
  printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof(flx_triangle), sizeof(flx_node), sizeof(flx_material),
         sizeof(flx_tex_desc), sizeof(flx_area_light), sizeof(flx_camera), sizeof(flx_render_params), sizeof(flx_hit),
         sizeof(flx_queue_counters), sizeof(flx_render_stats), sizeof(flx_vertex));
  printf("%zu %zu %zu %zu %zu %zu %zu\n", offsetof(flx_render_params, camera), offsetof(flx_render_params, ppParams),
         offsetof(flx_render_params, width), offsetof(flx_render_params, worldRadius), offsetof(flx_node, parent),
         offsetof(flx_triangle, matId), offsetof(flx_material, type));
  return 0; }
''')
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    l1, l2 = subprocess.check_output([str(exe)], text=True).splitlines()
    assert [int(x) for x in l1.split()] == [160, 48, 96, 12, 96, 80, 256, 64, 40, 16, 48]
    assert [int(x) for x in l2.split()] == [96, 176, 184, 232, 32, 144, 88]


def test_numpy_views_match_header():
    from fluctus_b200 import abi
    assert abi.TRI_DT.itemsize == 160 and abi.NODE_DT.itemsize == 48 and abi.MAT_DT.itemsize == 96
    assert abi.PARAMS_DT.itemsize == 256 and abi.HIT_DT.itemsize == 64 and abi.COUNTERS_DT.itemsize == 40
    assert abi.PARAMS_DT.fields["camera"][1] == 96 and abi.PARAMS_DT.fields["width"][1] == 184
    assert abi.MAT_DT.fields["type"][1] == 88 and abi.NODE_DT.fields["parent"][1] == 32
    # slot table == the enum in fluctus_abi.h
    text = open(ABI_HEADER).read()
    enum = dict((k, int(v)) for k, v in re.findall(r"FLX_SLOT_([A-Z_]+) = (\d+)", text))
    assert enum["SEED"] == abi.SLOTS["seed"] == 49 and enum["MAT_ID"] == abi.SLOTS["matId"] == 63
    assert enum["PIXEL_INDEX"] == abi.SLOTS["pixelIndex"] and enum["HIT_T"] == abi.SLOTS["t"]


def test_no_cpu_fallback_without_device():
    """flx_create must fail with FLX_ERR_NO_DEVICE and a message — never run on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible")
    import fluctus_b200 as fx
    with pytest.raises(fx.FluctusError) as e:
        fx.CudaContext(num_tasks=64)
    assert "no CUDA device" in str(e.value) or "CPU fallback" in str(e.value)
    lib = fx.load()
    h = C.c_void_p()
    assert lib.flx_create(C.byref(h), 0, 64, None) == -4  # FLX_ERR_NO_DEVICE
    assert not h.value
    # NULL-context calls return an argument error instead of crashing
    assert lib.flx_wf_logic(None, 0) == -2 and lib.flx_finish(None) == -2
    assert lib.flx_num_tasks(None) == 0 and lib.flx_launch_count(None) == 0


def test_product_does_not_touch_the_oracle():
    """Nothing under fluctus_b200/ or include/ may import, link or name the oracle."""
    bad = []
    for base in ("fluctus_b200", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", "Makefile")):
                    text = open(os.path.join(dirpath, f), errors="ignore").read()
                    if re.search(r"(from|import)\s+oracle|liboracle|libflxref|pyoracle|#include\s+\"[^\"]*oracle|-[LI]\S*oracle", text):
                        bad.append(os.path.join(dirpath, f))
    assert not bad, bad
    import fluctus_b200 as fx
    out = subprocess.check_output(["ldd", fx.LIB_PATH], text=True)
    assert "oracle" not in out and "flxref" not in out
