"""Drop-in flow of the Fortran shim: register HOST arrays, upload, call the work routine with
the host addresses, download -- results equal the oracle called with the same arguments."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


def _arglist(macro):
    src = open(os.path.join(common.ROOT, "include", "mpas_b200_args.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S).replace("\\\n", " ")
    macros = dict(re.findall(r"#define\s+(MPAS_\w+_ARGS)\s+(.*)", src))
    body = macros[macro].replace("MPAS_RANGE_ARGS", macros["MPAS_RANGE_ARGS"])
    out = []
    for a in body.split(","):
        a = a.strip()
        name = re.findall(r"(\w+)\s*$", a)[0]
        out.append((name, "ptr" if "*" in a else ("int" if a.startswith("int") else "real")))
    return out


def _call(lib, fname, macro, values):
    args = []
    for name, kind in _arglist(macro):
        v = values[name]
        if kind == "ptr":
            args.append(C.c_void_p(v.ctypes.data if v is not None else None))
        elif kind == "int":
            args.append(C.c_int(int(v)))
        else:
            args.append(C.c_double(float(v)))
    fn = getattr(lib, fname)
    fn.restype = C.c_int
    rc = fn(*args)
    assert rc == 0, rc


def test_divergence_damping_through_registered_host_arrays():
    case, blocks = common.make_case(4, 1, nVertLevels=26, num_scalars=1)
    b = blocks[0]
    cfg = b.cfg()
    rng = np.random.default_rng(20240611)
    L, nC, nE = 26, int(cfg["nCells"]), int(cfg["nEdges"])
    vals = dict(nVertLevels=L, nCells=nC, nEdges=nE, nCellsSolve=nC, edgeStart=1, edgeEnd=nE,
                theta_m=rng.uniform(250, 500, (nC + 1, L)), ru_p=rng.uniform(-50, 50, (nE + 1, L)),
                rtheta_pp=rng.uniform(-1, 1, (nC + 1, L)), rtheta_pp_old=rng.uniform(-1, 1, (nC + 1, L)),
                cellsOnEdge=b.field("cellsOnEdge").copy(), specZoneMaskEdge=np.zeros(nE + 1), dts=60.0, smdiv=0.1, config_len_disp=480000.0)
    ref = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in vals.items()}
    _call(common.oracle(), "oracle_divergence_damping_3d", "MPAS_DIVDAMP_ARGS", ref)
    Lb = common.host.lib()
    arrays = [v for v in vals.values() if isinstance(v, np.ndarray)]
    for a in arrays:
        assert Lb.mpas_b200_register_field(a.ctypes.data, a.nbytes) == 0
        assert Lb.mpas_b200_upload(a.ctypes.data) == 0
    _call(Lb, "mpas_b200_divergence_damping_3d", "MPAS_DIVDAMP_ARGS", vals)
    assert Lb.mpas_b200_download(vals["ru_p"].ctypes.data) == 0
    assert Lb.mpas_b200_synchronize() == 0
    assert np.array_equal(vals["ru_p"], ref["ru_p"])
    assert not np.array_equal(vals["ru_p"][:-1], 0 * vals["ru_p"][:-1])
    for a in arrays:
        assert Lb.mpas_b200_unregister_field(a.ctypes.data) == 0
