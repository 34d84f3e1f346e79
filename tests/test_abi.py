"""The drop-in boundary: libmpas_b200.so loads on a CPU-only machine and exports every symbol that
include/mpas_b200.h declares; the argument-list macros parse to the reference's dummy-argument counts."""
import ctypes
import os
import re

import common

ROOT = common.ROOT


def _declared(header="mpas_b200.h"):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mpas_(?:b200|host|nc)_\w+)\s*\(", src)))


def test_library_loads_and_exports_every_declared_symbol():
    L = common.host.lib()
    names = _declared()
    assert len(names) > 45
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_host_harness_library_has_no_cuda_and_exports_its_header():
    """libmpas_host.so (include/mpas_host.h) is what the CPU oracle and bench.py --impl reference load: it must not pull in
    the CUDA product library."""
    H = common.host.hostlib()
    names = _declared("mpas_host.h")
    assert len(names) > 35
    missing = [n for n in names if not hasattr(H, n)]
    assert not missing, missing
    import subprocess
    needed = subprocess.run(["objdump", "-p", common.host.HOSTLIB_PATH], capture_output=True, text=True).stdout
    assert "libmpas_b200" not in needed and "libcuda" not in needed, needed
    syms = subprocess.run(["nm", "-D", "--defined-only", common.host.HOSTLIB_PATH], capture_output=True, text=True).stdout
    assert "mpas_b200_" not in syms


def test_single_precision_library_exports_the_same_abi_with_4_byte_reals():
    """libmpas_b200_sp.so = the same sources with -DSINGLE_PRECISION (PRECISION=single, mpas_kind_types.F:25-29)."""
    L = ctypes.CDLL(os.path.join(common.PKG, "libmpas_b200_sp.so"), mode=ctypes.RTLD_LOCAL)
    missing = [n for n in _declared() if not hasattr(L, n)]
    assert not missing, missing
    assert L.mpas_b200_real_size() == 4
    assert common.host.lib().mpas_b200_real_size() == 8


def test_no_compute_without_gpu_is_loud():
    if common.has_gpu():
        return
    case, blocks = common.make_case(3, 1, nVertLevels=4, num_scalars=1)
    try:
        common.GpuModel(blocks, common.level_config(3))
    except RuntimeError as e:
        assert "no CPU fallback" in str(e)
    else:
        raise AssertionError("dycore created without a CUDA device")


def test_argument_lists_match_reference_dummy_argument_counts():
    """MPAS_ACOUSTIC_ARGS = the 66 dummies of TI:2447-2455 + nVertLevels, maxEdges (module mpas_atm_dimensions)."""
    src = open(os.path.join(ROOT, "include", "mpas_b200_args.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = src.replace("\\\n", " ")
    macros = dict(re.findall(r"#define\s+(MPAS_\w+_ARGS)\s+(.*)", src))
    rng = macros["MPAS_RANGE_ARGS"]

    def count(name):
        body = macros[name].replace("MPAS_RANGE_ARGS", rng)
        return len([a for a in body.split(",") if a.strip()])

    assert count("MPAS_RANGE_ARGS") == 12
    assert count("MPAS_ACOUSTIC_ARGS") == 66 + 2
    assert count("MPAS_DIVDAMP_ARGS") == 15
    # TI:4481-4499 has 128 dummies; + the 10 host-associated module scratch arrays (TI:35-43) + num_scalars, moist_start/end are already dummies
    assert count("MPAS_DYNTEND_ARGS") >= 128 + 10


def test_fortran_shim_interfaces_are_generated_from_the_headers_and_cover_every_work_routine():
    """shim/mpas_atm_b200_interfaces.inc (the bind(C) interface blocks the Fortran host compiles) must be exactly what
    tools/gen_shim.py makes of include/mpas_b200.h + mpas_b200_args.h, and name all 14 work routines plus the
    halo-exchange entry points, each with as many dummies as the C prototype has parameters."""
    import subprocess
    import sys
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_shim.py"), "--check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    inc = open(os.path.join(ROOT, "shim", "mpas_atm_b200_interfaces.inc")).read()
    hdr = open(os.path.join(ROOT, "include", "mpas_b200.h")).read()
    work = re.findall(r"int\s+(mpas_b200_\w+)\((MPAS_\w+_ARGS)\);", hdr)
    assert len(work) == 15  # the 14 of SURVEY.md 8(b) + atm_compute_output_diagnostics (8f f3)
    src = open(os.path.join(ROOT, "include", "mpas_b200_args.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S).replace("\\\n", " ")
    macros = dict(re.findall(r"#define\s+(MPAS_\w+_ARGS)\s+(.*)", src))
    for name, macro in work:
        m = re.search(r"function %s\((.*?)\)\s*&\s*\n\s*bind\(C, name='%s'\)" % (name, name), inc, flags=re.S)
        assert m, name
        dummies = [a for a in re.sub(r"[&\s]", "", m.group(1)).split(",") if a]
        body = macros[macro].replace("MPAS_RANGE_ARGS", macros["MPAS_RANGE_ARGS"])
        assert len(dummies) == len([a for a in body.split(",") if a.strip()]), name
    for name in ("mpas_b200_halo_register_lists", "mpas_b200_exch_halo_field2d", "mpas_b200_exch_halo_field3d", "mpas_b200_halo_comm_init"):
        assert "bind(C, name='%s')" % name in inc
    shim = open(os.path.join(ROOT, "shim", "mpas_atm_b200_shim.F")).read()
    assert '#include "mpas_atm_b200_interfaces.inc"' in shim
