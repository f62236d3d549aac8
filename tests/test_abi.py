"""CPU-side checks of the boundary: the C-ABI library builds, loads and exports every symbol include/henjou_b200.h
declares; ctypes struct layouts match the header; creating a context without a GPU fails loudly (no CPU path)."""
import ctypes as C
import os
import re

import pytest

import henjou_renderer_b200 as hj
from henjou_renderer_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(hj.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return hj.load_library()


def test_exports_match_header(lib):
    hdr = open(os.path.join(ROOT, "include", "henjou_b200.h")).read()
    declared = set(re.findall(r"\b(hj_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(abi.EXPORTS), declared ^ set(abi.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.hj_abi_version() == 3


def test_struct_sizes():
    assert C.sizeof(abi.hj_material) == 25 * 4
    assert C.sizeof(abi.hj_frame_params) == 6 * 4 + 12 * 4 + 2 * 4 + 6 * 4
    assert C.sizeof(abi.hj_stats) == 6 * 8 + 4 * 4 + 8 * 4


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(hj.HenjouError) as e:
        hj.Renderer(0)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "henjou-renderer_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dp, f)).read()
                assert "oracle/" not in src and "oracle_binding" not in src and "hjo_" not in src, f
