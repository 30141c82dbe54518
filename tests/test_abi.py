"""CPU test: the C-ABI library loads and exports every symbol include/pmcgpu.h declares
(no compute calls without a GPU), and refuses to run without a device."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            txt = open(os.path.join(ROOT, "include", fn)).read()
            txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
            names |= set(re.findall(r"\b(pmcgpu_[a-z0-9_]+)\s*\(", txt))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    import wrf_partmc_b200 as W
    if not os.path.exists(W.library_path()):
        import __graft_entry__
        __graft_entry__.build()
    L = W.load_library()
    decl = _declared()
    assert len(decl) >= 25
    for name in decl:
        assert hasattr(L, name), "libpmcgpu.so does not export %s" % name


def test_no_cpu_fallback():
    """Without a CUDA device the product refuses to run (checked only where no GPU is visible)."""
    import wrf_partmc_b200 as W
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("a GPU is present")
    with pytest.raises(W.PmcGpuError):
        W.PartMCGPU(W.AeroData([1000.0]))


def test_product_does_not_reference_oracle():
    """The product tree must not include, import or link anything under oracle/."""
    pkg = os.path.join(ROOT, "wrf-partmc_b200")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".cu", ".cuh", ".py", ".h", ".F90", "Makefile")):
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                for line in txt.splitlines():
                    if re.search(r"#include.*oracle|import .*oracle|from oracle|pmc_oracle|libpmc_oracle", line):
                        raise AssertionError("%s refers to the oracle: %s" % (fn, line))
