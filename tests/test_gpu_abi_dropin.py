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
        out.append((name, "ptr" if ("*" in a or a.startswith("mpas_halo_fn")) else ("int" if a.startswith("int") else "real")))
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


# ---- the same flow for the five north_star routines, from a mid-integration state --------------------------------------
# Argument name -> pool key where they differ (the wrappers' pool look-ups: TI:2312-2444 acoustic, TI:2798-2906 recover,
# TI:4217-4478 dyn_tend, TI:3509-3609 mono); everything else is the pool field of the same name, a
# dimension / index range of the block, or a namelist value.
RENAME = {
    "acoustic": dict(rho_zz="rho_zz@2", theta_m="theta_m@1", tend_ru="tend_u", tend_rt="tend_theta", tend_rw="tend_w", w="w@2"),
    "recover": dict(w="w@2", rho_zz="rho_zz@2", u="u@2", theta_m="theta_m@2", pressure_b="pressure_base", scalars="scalars@2", zb=None, zb3=None),
    "dyn_tend": dict(rho_zz="rho_zz@2", u="u@2", theta_m="theta_m@2", w="w@2", rb="rho_base", rr="rho_p", pp="pressure_p", pressure_b="pressure_base",
                     theta_m_save="theta_m@1", rr_save="rho_p_save", scalars="scalars@2", u_init=None, v_init=None, qv_init=None, t_init=None,
                     ur_cell=None, vr_cell=None, tend_w_pgf=None, tend_w_buoy=None),
    "mono": dict(uhAvg="ruAvg", scalar_tend="scalars_tend", rho_zz_old="rho_zz@1", rho_zz_new="rho_zz@2", scalars_old="scalars@1", scalars_new="scalars@2",
                 fnm="fzm", fnp="fzp", rdnw="rdzw"),
}
ENTRY = {"acoustic": ("advance_acoustic_step_work", "MPAS_ACOUSTIC_ARGS"), "recover": ("recover_large_step_variables_work", "MPAS_RECOVER_ARGS"),
         "dyn_tend": ("compute_dyn_tend_work", "MPAS_DYNTEND_ARGS"), "mono": ("advance_scalars_mono_work", "MPAS_MONO_ARGS")}


def _values(block, routine, scalars):
    cfg = block.cfg()
    nC, nE, nV = int(cfg["nCells"]), int(cfg["nEdges"]), int(cfg["nVertices"])
    base = dict(nCells=nC, nEdges=nE, nVertices=nV, nVertLevels=int(cfg["nVertLevels"]), maxEdges=int(cfg["maxEdges"]), maxEdges2=int(cfg["maxEdges2"]),
                num_scalars=int(cfg["num_scalars"]), nCellsSolve=int(cfg["nCellsSolve"]), nEdgesSolve=int(cfg["nEdgesSolve"]), vertexDegree=3,
                moist_start=int(cfg.get("moist_start", 1)), moist_end=int(cfg.get("moist_end", 1)),
                cellStart=1, cellEnd=nC, vertexStart=1, vertexEnd=nV, edgeStart=1, edgeEnd=nE, cellSolveStart=1, cellSolveEnd=int(cfg["nCellsSolve"]),
                vertexSolveStart=1, vertexSolveEnd=int(cfg["nVerticesSolve"]), edgeSolveStart=1, edgeSolveEnd=int(cfg["nEdgesSolve"]),
                cf1=cfg["cf1"], cf2=cfg["cf2"], cf3=cfg["cf3"], r_earth=cfg["sphere_radius"], epssm=0.1, coef_3rd_order=0.25, c_s=0.125,
                config_mix_full=1, config_horiz_mixing=1, config_del4u_div_factor=10.0, config_h_mom_eddy_visc2=0.0, config_v_mom_eddy_visc2=0.0,
                config_h_theta_eddy_visc2=0.0, config_v_theta_eddy_visc2=0.0, config_h_theta_eddy_visc4=0.0, config_h_mom_eddy_visc4=0.0,
                config_visc4_2dsmag=0.05, config_mpas_cam_coef=0.0, config_rayleigh_damp_u=0, config_rayleigh_damp_u_timescale_days=5.0,
                config_number_rayleigh_damp_u_levels=6, halo_fn=None, halo_ctx=None)
    base.update(scalars)
    names = set(block.names())
    vals = {}
    for name, kind in _arglist(ENTRY[routine][1]):
        if kind == "ptr":
            key = RENAME[routine].get(name, name)
            if key is None or name in ("halo_fn", "halo_ctx"):
                vals[name] = None
            else:
                assert key in names, (routine, name, key)
                vals[name] = block.field(key)
        else:
            assert name in base, (routine, name)
            vals[name] = base[name]
    return vals


@pytest.fixture(scope="module")
def midstate():
    """one block, integrated two steps and half way into a third by the oracle: realistic contents in every pool array"""
    case, blocks = common.make_case(4, 1, nVertLevels=26, num_scalars=3)
    b = blocks[0]
    s = b.field("scalars@1")
    lat = b.field("latCell")
    s[:-1, :, 1] = (np.sin(2 * lat) ** 2)[:-1, None]
    s[:-1, :, 2] = 1.0
    cfg = common.level_config(4)
    ora = common.OracleModel(blocks, cfg)
    ora.init()
    ora.step(2)
    dts = cfg["config_dt"] / 6
    ora.call("rk_setup"); ora.call("moist"); ora.call("vert_imp_coefs", dts)
    ora.call("dyn_tend", 1, cfg["config_dt"]); ora.call("smlstep")
    ora.call("acoustic", dts, 1); ora.call("divdamp", dts); ora.call("acoustic", dts, 2)
    return b, cfg


@pytest.mark.parametrize("routine,scalars", [
    ("acoustic", dict(dts=100.0, small_step=2)), ("acoustic", dict(dts=100.0, small_step=1)),
    ("recover", dict(dt=200.0, ns=2, rk_step=1)), ("recover", dict(dt=200.0, ns=2, rk_step=3)),
    ("dyn_tend", dict(rk_step=1, dt=600.0)), ("dyn_tend", dict(rk_step=2, dt=600.0)),
    ("mono", dict(dt=600.0, advance_density=1)),
], ids=["acoustic2", "acoustic1", "recover_rk1", "recover_rk3", "dyn_tend_rk1", "dyn_tend_rk2", "mono"])
def test_work_routines_through_registered_host_arrays(midstate, routine, scalars):
    """Layer 1 as the Fortran shim uses it: every array argument is a HOST address (registered + uploaded), the work routine is
    called through its C entry point with the reference's argument list, every array is downloaded again; the CPU oracle gets
    copies of the same tuple.  Bit-equal on every array (recover rk 3: exner / pressure_p within the last place)."""
    b, cfg = midstate
    if "config_len_disp" not in scalars:
        scalars = dict(scalars, config_len_disp=cfg["config_len_disp"])
    vals = _values(b, routine, scalars)
    gpu = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in vals.items()}
    ref = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in vals.items()}
    if routine == "dyn_tend":  # pure outputs: cleared so that "bit-equal to the oracle" cannot be met by leaving them alone
        for k in ("tend_u", "tend_w", "tend_theta", "h_divergence"):
            gpu[k][...] = 0.0
            ref[k][...] = 0.0
            vals[k] = gpu[k].copy()
    # arguments that alias the same pool array must alias in the copies too (none of these four routines has any)
    Lb = common.host.lib()
    arrays = {k: v for k, v in gpu.items() if isinstance(v, np.ndarray)}
    for a in arrays.values():
        assert Lb.mpas_b200_register_field(a.ctypes.data, a.nbytes) == 0
        assert Lb.mpas_b200_upload(a.ctypes.data) == 0
    try:
        _call(common.oracle(), "oracle_" + ENTRY[routine][0], ENTRY[routine][1], ref)
        _call(Lb, "mpas_b200_" + ENTRY[routine][0], ENTRY[routine][1], gpu)
        for a in arrays.values():
            assert Lb.mpas_b200_download(a.ctypes.data) == 0
        assert Lb.mpas_b200_synchronize() == 0
        changed = 0
        loose = ("exner", "pressure_p") if (routine == "recover" and scalars.get("rk_step") == 3) else ()
        # scratch of the monotonic routine that the all-scalars path keeps in library workspace (tests/test_gpu_routines.py)
        skip = ("wdtn", "scalar_old", "scalar_new", "s_max", "s_min", "scale_arr", "flux_arr", "flux_upwind_tmp", "flux_tmp") if routine == "mono" else ()
        for k, a in arrays.items():
            if k in skip:
                continue
            if k in loose:
                assert common.rel_l2(a, ref[k]) < 1e-15, k
            else:
                assert np.array_equal(a, ref[k], equal_nan=True), (routine, k, float(np.nanmax(np.abs(a - ref[k]))))
            changed += int(not np.array_equal(a, vals[k], equal_nan=True))
        assert changed >= 2, "the routine must have written something"
    finally:
        for a in arrays.values():
            Lb.mpas_b200_unregister_field(a.ctypes.data)
