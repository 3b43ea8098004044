"""The C-ABI library loads without a GPU and exports every symbol include/radiance_b200.h declares."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "radiance_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rd_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(pkg):
    lib = ctypes.CDLL(pkg.rd.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(lib, s), "missing export: " + s
    assert sorted(pkg.rd.ABI_SYMBOLS) == syms


def test_no_cpu_fallback_without_device(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.rd.RadianceError, match="no CPU path"):
        pkg.rd.Platform(0)


def test_sizes_match_reference_layouts(pkg, ref):
    s = pkg.scenes
    L = ref.lib()
    # radiance/src/core.h:103-158 (RayTraceProperties, Material, MeshInfo, SceneProperties, PhysicalCamera)
    assert (L.ref_sizeof(6), L.ref_sizeof(7), L.ref_sizeof(8), L.ref_sizeof(10), L.ref_sizeof(11)) == (
        s.RTPROP_DTYPE.itemsize, s.MATERIAL_DTYPE.itemsize, s.MESHINFO_DTYPE.itemsize, s.SCENEPROP_DTYPE.itemsize,
        s.CAMERA_DTYPE.itemsize)
