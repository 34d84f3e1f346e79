"""BASELINE.json configs[2]: idealised supercell on a doubly periodic planar mesh of regular hexagons, moist scalars (qv, qc, qr)
transported, physics off.  Planar mesh generator (csrc/host/mesh.cpp make_planar_hex_mesh), the restatement of
init_atm_case_squall_line (src/core_init_atmosphere/mpas_init_atm_cases.F:1313-1895, csrc/host/init_squall.cpp), the
on_a_sphere = .false. branches of the advection / reconstruction set-up, periodic halo topology in the decomposition, and
the code paths the sphere cases never reach: config_horiz_mixing = '2d_fixed', vertical mixing with config_mix_full =
.false. (u_init / t_init, TI:4949-5006, 5360-5406), three moist species in qtot (TI:1903-1905)."""
import numpy as np
import pytest

import common

NX, NY, DC, L, NS = 24, 24, 2000.0, 20, 3
CFG = dict(config_dt=12.0, config_len_disp=DC)


def supercell(nblocks=1, nx=NX, ny=NY, method="metis", **kw):
    case = common.host.HostCase(0, planar=(nx, ny, DC), sfc=True)
    if nblocks > 1:
        case.partition(nblocks, method)
    blocks = [case.build_block(i, nVertLevels=L, init_case=5, num_scalars=NS, full_fields=True, **kw) for i in range(nblocks)]
    return case, blocks


def test_planar_periodic_hex_mesh_is_a_closed_regular_tiling():
    case, blocks = supercell()
    b = blocks[0]
    nC, nE, nV = case.nCells, case.nEdges, case.nVertices
    assert (nC, nE, nV) == (NX * NY, 3 * NX * NY, 2 * NX * NY)  # torus: V - E + F = 0
    assert (b.field("nEdgesOnCell")[:nC] == 6).all() and (b.field("nEdgesOnEdge")[:nE] == 10).all()
    assert int(b.cfg()["on_a_sphere"]) == 0
    h = DC * np.sqrt(3) / 2
    assert np.allclose(b.field("areaCell")[:nC].sum(), NX * DC * NY * h, rtol=1e-13)
    assert np.allclose(b.field("areaTriangle")[:nV].sum(), NX * DC * NY * h, rtol=1e-13)
    assert np.allclose(b.field("dcEdge")[:nE], DC, rtol=1e-13) and np.allclose(b.field("dvEdge")[:nE], DC / np.sqrt(3), rtol=1e-13)
    # every neighbour relation is mutual, also across the periodic boundaries
    coc = b.field("cellsOnCell")[:nC]
    for c in range(nC):
        for n in coc[c]:
            assert c + 1 in coc[n - 1]
    # the normal of an edge points from its first to its second cell (minimum-image displacement)
    coe, ang = b.field("cellsOnEdge")[:nE], b.field("angleEdge")[:nE]
    x, y = b.field("xCell"), b.field("yCell")
    dx = x[coe[:, 1] - 1] - x[coe[:, 0] - 1]
    dy = y[coe[:, 1] - 1] - y[coe[:, 0] - 1]
    dx -= NX * DC * np.round(dx / (NX * DC))
    dy -= NY * h * np.round(dy / (NY * h))
    assert np.allclose(np.cos(ang) * DC, dx, atol=1e-6) and np.allclose(np.sin(ang) * DC, dy, atol=1e-6)
    # TRiSK: antisymmetric un-scaled weights, and the tangential velocity of a UNIFORM flow is reproduced to round-off
    W, eoe, neoe, dc, dv = b.field("weightsOnEdge")[:nE], b.field("edgesOnEdge")[:nE], b.field("nEdgesOnEdge")[:nE], b.field("dcEdge"), b.field("dvEdge")
    w = {(e, int(eoe[e, j]) - 1): W[e, j] * dc[e] / dv[eoe[e, j] - 1] for e in range(nE) for j in range(neoe[e])}
    assert max(abs(v + w[(k[1], k[0])]) for k, v in w.items()) < 1e-14
    for U, V in ((1.0, 0.0), (0.3, -0.8)):
        un = U * np.cos(ang) + V * np.sin(ang)
        vt = (W[:, :10] * un[eoe[:, :10] - 1]).sum(axis=1)  # maxEdges2 = 12 slots, 10 used
        assert np.allclose(vt, -U * np.sin(ang) + V * np.cos(ang), atol=1e-13)


def test_supercell_initial_state_is_the_weisman_klemp_sounding_with_a_warm_bubble():
    case, blocks = supercell()
    b = blocks[0]
    nC = case.nCells
    th, qv = b.field("theta")[:nC], b.field("scalars@1")[:nC, :, 0]
    x, y = b.field("xCell")[:nC], b.field("yCell")[:nC]
    r = np.hypot(x - 0.5 * x.max(), y - 0.5 * y.max())
    far, near = r > 12000.0, r < 3000.0
    assert np.ptp(th[far], axis=0).max() == 0.0  # horizontally uniform outside the bubble (radius 10 km)
    dth = th[near].max(axis=0) - th[far][0]
    assert 2.5 < dth.max() <= 3.0 and dth[L // 2:].max() == 0.0  # +3 K at most, centred at 1.5 km, nothing above 3 km
    assert qv.max() == 0.014 and qv.min() > 0 and (np.diff(qv[far][0][:12]) <= 0).all()  # capped at 14 g/kg, decreasing through the troposphere (12 km)
    assert (np.diff(th[far][0]) > 0).all()  # statically stable
    u, ang = b.field("u@1")[:case.nEdges], b.field("angleEdge")[:case.nEdges]
    prof = b.field("u_init")
    assert prof[0] == 30.0 * 500.0 / 5000.0 - 15.0 and prof[-1] == 15.0  # linear shear to um = 30 at 5 km, minus us = 15 (first level at 500 m)
    assert np.array_equal(u, np.cos(ang)[:, None] * prof[None, :])
    assert not b.field("fEdge").any() and not b.field("zb_cell").any()
    c = b.cfg()
    assert (c["config_horiz_mixing"], c["config_mix_full"], c["moist_end"]) == (2, 0, 3)


def test_supercell_oracle_conserves_mass_keeps_the_far_field_at_rest_and_is_partition_independent():
    case, blocks = supercell()
    m = common.OracleModel(blocks, CFG)
    m.init()
    nC = case.nCells
    area = blocks[0].field("areaCell")[:nC]
    mass0 = (m.field(0, "rho_zz@1")[:nC] * area[:, None]).sum()
    qv0 = (m.field(0, "rho_zz@1")[:nC] * m.field(0, "scalars@1")[:nC, :, 0] * area[:, None]).sum()
    m.step(10)
    w = m.field(0, "w@1")[:nC]
    x, y = blocks[0].field("xCell")[:nC], blocks[0].field("yCell")[:nC]
    r = np.hypot(x - 0.5 * x.max(), y - 0.5 * y.max())
    assert 0.5 < np.abs(w[r < 4000.0]).max() < 5.0   # the bubble rises
    assert np.abs(w[r > 16000.0]).max() < 0.02        # the balanced sounding stays at rest: the hydrostatic iteration converged
    assert abs((m.field(0, "rho_zz@1")[:nC] * area[:, None]).sum() / mass0 - 1) < 1e-13
    assert abs((m.field(0, "rho_zz@1")[:nC] * m.field(0, "scalars@1")[:nC, :, 0] * area[:, None]).sum() / qv0 - 1) < 1e-12
    assert not m.field(0, "scalars@1")[:nC, :, 1:].any()  # qc, qr: no microphysics, nothing to transport yet
    # decomposition test on the periodic topology: 4 blocks == 1 block, bitwise, on every owned element
    case4, blocks4 = supercell(4)
    m4 = common.OracleModel(blocks4, CFG)
    m4.init()
    m4.step(10)
    for ib, b in enumerate(blocks4):
        ids = b.field("indexToCellID")
        nown = int(b.cfg()["nCellsSolve"])
        for n in ("theta_m@1", "rho_zz@1", "w@1", "scalars@1"):
            assert np.array_equal(m4.field(ib, n)[:nown], m.field(0, n)[ids[:nown] - 1]), (ib, n)


STATE = ["u@1", "w@1", "theta_m@1", "rho_zz@1", "scalars@1"]
DIAG = ["ru", "rw", "rho_p", "rtheta_p", "exner", "pressure_p", "rho_edge", "pv_edge", "v", "ke", "divergence", "vorticity", "tend_u", "tend_w", "tend_theta",
        "tend_rho", "ruAvg", "wwAvg", "kdiff", "h_divergence", "uReconstructX", "uReconstructY", "uReconstructZonal", "uReconstructMeridional"]


@pytest.mark.gpu
@pytest.mark.parametrize("nblocks", [1, 3])
def test_supercell_gpu_matches_oracle_after_1_and_10_steps(nblocks):
    common.oracle_set_option("cr_pow", 1)
    try:
        case, blocks = supercell(nblocks)
        gpu = common.GpuModel(blocks, CFG)
        ora = common.OracleModel(blocks, CFG)
        gpu.init(); ora.init()
        for nstep in (1, 9):
            gpu.step(nstep); ora.step(nstep)
            # norms over the owned elements of ALL blocks (what a one-block run measures): a block far from the bubble holds
            # nothing but round-off in its tendencies, a per-block relative error would be noise over noise
            num, den, dmax = {}, {}, {}
            for ib, b in enumerate(blocks):
                for n in STATE + DIAG:
                    g, o = gpu.field(ib, n), ora.field(ib, n)
                    loc = b.info(n)[1]
                    nown = int(b.cfg()[{1: "nCellsSolve", 2: "nEdgesSolve", 3: "nVerticesSolve"}[loc]])
                    assert np.isfinite(g[:nown]).all(), (nstep, ib, n)
                    num[n] = num.get(n, 0.0) + float(((g[:nown] - o[:nown]) ** 2).sum())
                    den[n] = den.get(n, 0.0) + float((o[:nown] ** 2).sum())
                    dmax[n] = max(dmax.get(n, 0.0), float(np.abs(g[:nown] - o[:nown]).max()))
            for n in STATE + DIAG:
                if n in ("vorticity", "pv_edge", "divergence"):
                    # f = 0 and the sheared flow is (nearly) irrotational and non-divergent: these fields are round-off-sized
                    # residuals (1e-15 ... 1e-7 1/s); measured against the velocity gradient scale max|u| / dc instead
                    e = dmax[n] / (15.0 / DC)
                else:
                    e = np.sqrt(num[n] / den[n]) if den[n] > 0 else np.sqrt(num[n])
                # tendencies / vertical velocity: small residuals of the hydrostatic balance of a convectively unstable column --
                # the last-place difference of exner (pow_rcv) grows to a few 1e-12 within 10 steps
                tol = 1e-11 if n in ("tend_u", "tend_w", "tend_theta", "tend_rho", "w@1", "rw", "wwAvg", "h_divergence") else 1e-12
                assert e <= tol, (nstep, n, e)
        # the reference sequence, except that the all-scalars monotonic transport sends the limiter scale factors of all
        # scalars in ONE exchange where the reference's loop over scalars posts one each (TI:4098)
        assert ora.exchange_count() - gpu.exchange_count() == (10 * (NS - 1) if nblocks > 1 else 0)
    finally:
        common.oracle_set_option("cr_pow", 0)
