"""CPU-side checks of the drop-in boundary: the library loads, exports every symbol declared in
include/spydrpick_b200.h, and refuses to run without a GPU (no CPU fallback)."""
import os
import re

import pytest

import spydrpick_b200.capi as capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    hdr = open(os.path.join(ROOT, "include", "spydrpick_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(spk_[a-z_0-9]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    lib = capi.load()
    names = _declared()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(capi.EXPORTS) == names


def test_struct_layouts_match_header(tmp_path):
    """sizeof/offsetof as the C compiler sees the header vs the ctypes mirrors."""
    import ctypes as C
    import subprocess
    src = tmp_path / "layout.c"
    src.write_text("""
#include <stdio.h>
#include <stddef.h>
#include "spydrpick_b200.h"
int main(void) {
  printf("%zu %zu %zu %zu ", sizeof(spk_scan_params), sizeof(spk_scan_result), sizeof(spk_timings), sizeof(spk_edge));
  printf("%zu %zu %zu ", offsetof(spk_scan_params, band), offsetof(spk_scan_params, keep_on_device), offsetof(spk_scan_result, max_filter_error));
  printf("%zu ", offsetof(spk_timings, gemm_tiles));
  return 0; }
""")
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    exp = [C.sizeof(capi.ScanParams), C.sizeof(capi.ScanResult), C.sizeof(capi.Timings), capi.EDGE_DTYPE.itemsize,
           capi.ScanParams.band.offset, capi.ScanParams.keep_on_device.offset, capi.ScanResult.max_filter_error.offset,
           capi.Timings.gemm_tiles.offset]
    assert got == exp


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(capi.SpkError) as ei:
        capi.Context()
    assert ei.value.code == capi.SPK_E_CUDA
    assert "no CPU fallback" in str(ei.value)


def test_product_path_does_not_import_oracle():
    pkg = os.path.join(ROOT, "spydrpick_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".h", ".cpp")):
                src = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle" not in src.replace("# oracle-free", ""), os.path.join(dp, f)


def test_outlier_thresholds_host_function():
    # host-only entry point: usable without a device
    import ctypes as C
    import numpy as np
    lib = capi.load()
    cm = np.array([0.9, 0.1, 0.5, 0.3, 0.7, 0.2, 0.8, 0.4, 0.6, 1.0])
    a = C.c_double(); b = C.c_double()
    assert lib.spk_outlier_thresholds(cm.ctypes.data_as(C.c_void_p), 10, C.byref(a), C.byref(b)) == 0
    assert a.value == pytest.approx(0.7 + 1.5 * 0.4) and b.value == pytest.approx(0.7 + 3.0 * 0.4)
