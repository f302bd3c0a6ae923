"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/bfq_gpu.h
declares, reports its compile-time mode, and fails loudly (no CPU fallback) without a GPU."""
import ctypes as C
import os
import re

import pytest

from bfqzip_b200 import lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MODES = [(M, B) for M in range(4) for B in range(2)]


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "bfq_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bfq_[a-z_A-Z0-9]+)\s*\(", src)))


def test_header_declares_expected_entry_points():
    syms = declared_symbols()
    for s in ("bfq_build_ebwt", "bfq_smooth_rebuild", "bfq_pipeline", "bfq_pipeline_dev",
              "bfq_ctx_create", "bfq_ctx_destroy", "bfq_last_error"):
        assert s in syms


@pytest.mark.parametrize("mode", MODES, ids=[f"M{m}B{b}" for m, b in MODES])
def test_library_exports_all_symbols(mode):
    M, B = mode
    L = lib.load(M, B)
    for s in declared_symbols():
        assert hasattr(L, s), f"{s} missing from {lib.lib_path(M, B)}"
    assert L.bfq_mode_M() == M and L.bfq_mode_B() == B


def test_params_default_matches_reference_defaults():
    L = lib.load(2, 0)
    p = lib.Params()
    L.bfq_params_default(C.byref(p))
    # bfq_ext.cpp:69-96
    assert (p.K, p.m, p.qthr, p.fthr, p.eos_byte, p.default_q) == (16, 2, 20, 40.0, ord("#"), ord(">"))


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(lib.BfqError) as ei:
        lib.BfqGpu(2, 0, 0)
    assert ei.value.code == -1
    assert "no CPU path" in str(ei.value)


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under bfqzip_b200/ or tools/ may reference it."""
    bad = []
    for base in ("bfqzip_b200", "tools"):
        for dp, _, fns in os.walk(os.path.join(ROOT, base)):
            for fn in fns:
                if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")) or fn == "Makefile":
                    txt = open(os.path.join(dp, fn), errors="ignore").read()
                    if re.search(r"oracle", txt):
                        bad.append(os.path.join(dp, fn))
    assert not bad, bad
