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


# ---------------------------------------------------------------------------------------------------------
# The Fortran side of the boundary (no Fortran compiler exists in this image: these are the checks that can run)
# ---------------------------------------------------------------------------------------------------------
def _gen():
    import importlib.util
    path = os.path.join(ROOT, "wrf-partmc_b200", "fortran", "gen_interfaces.py")
    spec = importlib.util.spec_from_file_location("gen_interfaces", path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_fortran_mirror_is_what_the_header_generates():
    """pmc_gpu_c.F90 is generated from include/pmcgpu.h: the committed file must be current, and it must bind
    every exported function of the header by name"""
    g = _gen()
    structs, funcs = g.parse_header()
    assert open(g.OUT).read() == g.emit(structs, funcs), "run python wrf-partmc_b200/fortran/gen_interfaces.py"
    txt = open(g.OUT).read()
    bound = set(re.findall(r"bind\(c, name='(pmcgpu_\w+)'\)", txt))
    assert bound == set(_declared())


def test_fortran_struct_layouts_match_offsetof(tmp_path):
    """every `type, bind(c)` of pmc_gpu_c.F90 has the members of the C struct of the same name, in the same order,
    at the same offsets (C ABI rules applied to the Fortran member types vs offsetof() compiled from the header)
    and the same total size"""
    import subprocess
    g = _gen()
    ftypes = dict(g.parse_fortran_types(g.OUT))
    structs, _ = g.parse_header()
    assert set(ftypes) == {s + "_t" for s, _ in structs} and len(structs) >= 6
    src = ['#include <stdio.h>', '#include <stddef.h>', '#include "pmcgpu.h"', "int main(void) {"]
    for sname, members in structs:
        src.append('  printf("%s %%zu\\n", sizeof(%s));' % (sname, sname))
        for nm, _ in members:
            src.append('  printf("%s.%s %%zu\\n", offsetof(%s, %s));' % (sname, nm, sname, nm))
    src += ["  return 0;", "}"]
    cfile = tmp_path / "layout.c"
    cfile.write_text("\n".join(src))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(cfile), "-o", str(exe)])
    actual = dict(line.split() for line in subprocess.check_output([str(exe)], text=True).splitlines())
    for sname, _ in structs:
        fm = ftypes[sname + "_t"]
        offs, size = g.c_offsets(fm)
        assert int(actual[sname]) == size, sname
        for nm, off in offs:
            assert int(actual["%s.%s" % (sname, nm)]) == off, (sname, nm)


def test_fortran_dropin_keeps_the_reference_entry_points():
    """pmc_gpu.F90 defines the routines a WRF-PartMC build calls under the reference's own names, with the
    reference's dummy-argument lists (interface/wrf_pmc_driver.F90:46-48, wrf_pmc_trans.F90:37-39,
    wrf_pmc_trans_aero.F90:40-42, coagulation.F90:43-44, condense.F90:147-148, aero_state.F90:3218, :1765)"""
    txt = open(os.path.join(ROOT, "wrf-partmc_b200", "fortran", "pmc_gpu.F90")).read()
    flat = re.sub(r"&\s*\n\s*", " ", txt)
    want = {
        "partmc_timestep": "grid, config_flags, scenario, env_states, aero_data, aero_states, gas_data, gas_states, "
                           "pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke, global_nx, global_ny, global_nz",
        "wrf_pmc_trans_driver": "grid, config_flags, env_states, scenario, aero_data, aero_states, gas_states, pmc_is, "
                                "pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke, global_nx, global_ny, global_nz",
        "trans_aero": "grid, config_flags, aero_data, scenario, env_states, aero_states, pmc_is, pmc_ie, pmc_ks, pmc_ke, "
                      "pmc_js, pmc_je, global_nx, global_ny, global_nz",
        "mc_coag": "coag_kernel_type, env_state, aero_data, aero_state, del_t, tot_n_samp, tot_n_coag",
        "condense_particles": "aero_state, aero_data, env_state_initial, env_state_final, del_t",
        "aero_state_sort": "aero_state, aero_data, bin_grid, all_procs_same",
        "aero_state_rebalance": "aero_state, aero_data, allow_doubling, allow_halving, initial_state_warning",
    }
    for name, args in want.items():
        m = re.search(r"subroutine %s\((.*?)\)" % name, flat)
        assert m, name
        assert [a.strip() for a in m.group(1).split(",")] == [a.strip() for a in args.split(",")], name
    # every C function the shim calls is bound by the generated mirror
    bound = set(re.findall(r"bind\(c, name='(pmcgpu_\w+)'\)",
                           open(os.path.join(ROOT, "wrf-partmc_b200", "fortran", "pmc_gpu_c.F90")).read()))
    called = set(re.findall(r"\b(pmcgpu_[a-z0-9_]+)\s*\(", txt))
    assert called <= bound, called - bound
