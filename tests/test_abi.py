"""The C-ABI library loads and exports every symbol include/b200pt.h declares; without a CUDA
device it fails loudly (no CPU fallback). No compute calls here."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "b200pt.h")).read()
    return sorted(set(re.findall(r"\b(b200pt_[a-z0-9_]+)\s*\(", src)))


def test_header_and_export_list_agree(lib_built):
    assert header_symbols() == sorted(lib_built.EXPORTS)


def test_library_exports_every_declared_symbol(lib_built):
    lib = C.CDLL(lib_built.LIB_PATH)
    for name in header_symbols():
        assert hasattr(lib, name), name


def test_struct_sizes_match_header(lib_built):
    from ipu_path_trace_b200 import _lib
    assert C.sizeof(_lib.Config) == 48
    assert C.sizeof(_lib.NifLayer) == 32
    assert C.sizeof(_lib.Counters) == 64
    assert C.sizeof(_lib.CornellConfig) == 64 and C.sizeof(_lib.CornellStats) == 24
    assert lib_built.TRACE_RECORD.itemsize == 20


def test_no_cpu_fallback(lib_built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(lib_built.B200PTError):
        lib_built.PathTracer(64, 64)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "ipu_path_trace_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", "Makefile")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in txt.lower() or f in ("pt_common.cuh", "pt_device.cuh", "Makefile", "nif_simt.cu", "trace_kernels.cu"), f
                assert "import oracle" not in txt and "from oracle" not in txt and "liboracle" not in txt, f
