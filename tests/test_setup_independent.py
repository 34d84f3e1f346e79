"""Independent checks of the set-up code that exists in ONE implementation (csrc/host/*.cpp feeds both the oracle and the
GPU, so "GPU == oracle" says nothing about it).  Two kinds of pins:

  * NumPy restatements written from the reference Fortran (not from the C++): atm_compute_signs
    (src/core_atmosphere/mpas_atm_core.F:987-1074), atm_adv_coef_compression + atm_couple_coef_3rd_order (:1121-1288),
    the inverses of mpas_atm_core.F:339-353 -- compared bit for bit;
  * analytic properties no transcription error survives: TRiSK weights (antisymmetry of the un-scaled weights, Thuburn et
    al. 2009 eq. 33, and the tangential velocity of a solid-body rotation), deriv_two applied to a field with a known
    second derivative along every geodesic, zero deformation of a solid-body rotation through defc_a / defc_b, and the
    Jablonowski-Williamson steady state (init_case = 1): initial tendencies shrink with the mesh spacing.
"""
import numpy as np
import pytest

import common


@pytest.fixture(scope="module")
def blk3():
    case = common.host.HostCase(3, lloyd_iters=10, sfc=True)
    return case, case.build_block(0, nVertLevels=8, init_case=2, num_scalars=1, full_fields=True, keep_zb=True)


def test_compute_signs_kites_and_zb_cell_against_numpy_restatement(blk3):
    case, b = blk3
    nC, nE, nV = case.nCells, case.nEdges, case.nVertices
    eov, voe, eoc, coe = b.field("edgesOnVertex"), b.field("verticesOnEdge"), b.field("edgesOnCell"), b.field("cellsOnEdge")
    voc, cov, nec = b.field("verticesOnCell"), b.field("cellsOnVertex"), b.field("nEdgesOnCell")
    sv = np.zeros((nV, 3))
    for v in range(nV):  # :1025-1037
        for i in range(3):
            e = eov[v, i]
            if e <= nE:
                sv[v, i] = 1.0 if v + 1 == voe[e - 1, 1] else -1.0
    assert np.array_equal(b.field("edgesOnVertex_sign")[:nV], sv)
    zb, zb3 = b.field("zb"), b.field("zb3")  # (nEdges+1, 2, L+1)
    zbc, zb3c = b.field("zb_cell"), b.field("zb3_cell")  # (nCells+1, maxEdges, L+1)
    c3 = b.cfg().get("config_coef_3rd_order", 0.25)
    sc = np.zeros((nC, eoc.shape[1]))
    kite = np.zeros((nC, eoc.shape[1]), dtype=np.int32)
    for c in range(nC):
        for i in range(nec[c]):
            e = eoc[c, i]
            if e <= nE:  # :1039-1055
                first = c + 1 == coe[e - 1, 0]
                sc[c, i] = 1.0 if first else -1.0
                assert np.array_equal(zbc[c, i], zb[e - 1, 0 if first else 1]), (c, i)
                assert np.array_equal(zb3c[c, i], c3 * zb3[e - 1, 0 if first else 1]), (c, i)  # atm_couple_coef_3rd_order :1285
            v = voc[c, i]  # :1057-1072
            kite[c, i] = 1
            if v <= nV:
                kite[c, i] = 1 + cov[v - 1].tolist().index(c + 1)
    got_s, got_k = b.field("edgesOnCell_sign")[:nC], b.field("kiteForCell")[:nC]
    for c in range(nC):
        assert np.array_equal(got_s[c, :nec[c]], sc[c, :nec[c]])
        assert np.array_equal(got_k[c, :nec[c]], kite[c, :nec[c]])


def test_inverse_metrics(blk3):
    case, b = blk3
    for inv, x, n in (("invAreaCell", "areaCell", case.nCells), ("invDvEdge", "dvEdge", case.nEdges), ("invDcEdge", "dcEdge", case.nEdges),
                      ("invAreaTriangle", "areaTriangle", case.nVertices)):
        assert np.array_equal(b.field(inv)[:n], 1.0 / b.field(x)[:n]), inv  # mpas_atm_core.F:339-353


def test_adv_coef_compression_against_numpy_restatement(blk3):
    case, b = blk3
    nC, nE = case.nCells, case.nEdges
    d2 = b.field("deriv_two")  # (nEdges+1, 2, 15)
    coc, coe, nec = b.field("cellsOnCell"), b.field("cellsOnEdge"), b.field("nEdgesOnCell")
    dc, dv = b.field("dcEdge"), b.field("dvEdge")
    c3 = b.cfg().get("config_coef_3rd_order", 0.25)
    nadv, adv, ac, ac3 = b.field("nAdvCellsForEdge"), b.field("advCellsForEdge"), b.field("adv_coefs"), b.field("adv_coefs_3rd")
    for e in range(nE):
        c1, c2 = int(coe[e, 0]), int(coe[e, 1])
        lst = [c1, c2]
        for i in range(nec[c1 - 1]):  # :1168-1173
            if coc[c1 - 1, i] != c2:
                lst.append(int(coc[c1 - 1, i]))
        for i in range(nec[c2 - 1]):  # :1176-1185
            if int(coc[c2 - 1, i]) not in lst:
                lst.append(int(coc[c2 - 1, i]))
        n = len(lst)
        assert nadv[e] == n and adv[e, :n].tolist() == lst, e
        a, a3 = np.zeros(15), np.zeros(15)
        j = lst.index(c1)
        a[j] += d2[e, 0, 0]; a3[j] += d2[e, 0, 0]
        for i in range(nec[c1 - 1]):
            j = lst.index(int(coc[c1 - 1, i]))
            a[j] += d2[e, 0, i + 1]; a3[j] += d2[e, 0, i + 1]
        j = lst.index(c2)
        a[j] += d2[e, 1, 0]; a3[j] -= d2[e, 1, 0]
        for i in range(nec[c2 - 1]):
            j = lst.index(int(coc[c2 - 1, i]))
            a[j] += d2[e, 1, i + 1]; a3[j] -= d2[e, 1, i + 1]
        dc2 = dc[e] * dc[e]  # gfortran expands x**2 to x*x (numpy's float ** 2 goes through libm pow: last-place differences)
        for j in range(n):  # :1231-1234
            a[j] = -dc2 * a[j] / 12.
            a3[j] = -dc2 * a3[j] / 12.
        a[lst.index(c1)] += 0.5
        a[lst.index(c2)] += 0.5
        for j in range(n):
            a[j] = dv[e] * a[j]
            a3[j] = c3 * (dv[e] * a3[j])  # atm_couple_coef_3rd_order :1284
        assert np.array_equal(ac[e], a), e
        assert np.array_equal(ac3[e], a3), e


def _edge_normals(b, n):
    """unit normal of each edge in (east, north) components from angleEdge (the angle of the normal from east)"""
    ang = b.field("angleEdge")[:n]
    return np.cos(ang), np.sin(ang)


def test_trisk_weights_antisymmetry_and_solid_body_tangential_velocity():
    errs = []
    for level in (3, 4, 5):
        case = common.host.HostCase(level, lloyd_iters=10, sfc=True)
        b = case.build_block(0, nVertLevels=4, init_case=2, num_scalars=1, full_fields=False)
        nE = case.nEdges
        W, eoe, neoe = b.field("weightsOnEdge")[:nE], b.field("edgesOnEdge")[:nE], b.field("nEdgesOnEdge")[:nE]
        dc, dv = b.field("dcEdge")[:nE], b.field("dvEdge")[:nE]
        if level == 3:
            # un-scaled weights w(e,e') = W(e,e') dc(e) / dv(e') are antisymmetric (energy conservation of the Coriolis term)
            w = {}
            for e in range(nE):
                for j in range(neoe[e]):
                    w[(e, int(eoe[e, j]) - 1)] = W[e, j] * dc[e] / dv[int(eoe[e, j]) - 1]
            worst = max(abs(v + w[(k[1], k[0])]) for k, v in w.items())
            assert worst < 1e-12, worst
        # solid-body rotation about the polar axis: U = cos(lat) east, V = 0; normal u = U cos(angle), tangential = -U sin(angle)
        lat = b.field("latEdge")[:nE]
        cx, sx = _edge_normals(b, nE)
        un = np.cos(lat) * cx
        vt_exact = -np.cos(lat) * sx  # t = k x n
        vt = np.zeros(nE)
        for j in range(eoe.shape[1]):
            m = j < neoe
            vt[m] += W[m, j] * un[eoe[m, j] - 1]
        errs.append(np.sqrt(np.mean((vt - vt_exact) ** 2)))
    print("tangential reconstruction rms error, levels 3-5:", errs)
    # 0.3 % of U at 960 km and falling (the TRiSK reconstruction is only weakly consistent on a relaxed Voronoi mesh: slow but monotone)
    assert errs[0] < 0.005 and errs[2] < errs[1] < errs[0]


def test_deriv_two_second_derivative_of_a_field_known_along_every_geodesic():
    """f = x (a Cartesian coordinate) on the sphere of radius a: along ANY geodesic through p, d2f/ds2 = -x(p)/a^2.  The
    least-squares polynomial fit of mpas_atm_advection.F:21-394 must reproduce it with an error that shrinks with the mesh."""
    errs = []
    for level in (3, 4, 5):
        case = common.host.HostCase(level, lloyd_iters=10, sfc=True)
        b = case.build_block(0, nVertLevels=4, init_case=2, num_scalars=1, full_fields=False)
        nE, nC = case.nEdges, case.nCells
        x, y, z = b.field("xCell")[:nC], b.field("yCell")[:nC], b.field("zCell")[:nC]
        a = np.sqrt(x[0] ** 2 + y[0] ** 2 + z[0] ** 2)
        d2, coe, coc, nec = b.field("deriv_two")[:nE], b.field("cellsOnEdge")[:nE], b.field("cellsOnCell")[:nC], b.field("nEdgesOnCell")[:nC]
        worst = 0.0
        for f in (x, y, z):
            for side in (0, 1):
                c = coe[:, side] - 1
                val = d2[:, side, 0] * f[c]
                for i in range(coc.shape[1]):
                    m = i < nec[c]
                    val[m] += d2[m, side, i + 1] * f[coc[c[m], i] - 1]
                err = np.abs(val - (-f[c] / a ** 2)) * a  # relative to |f|max / a^2 = 1 / a
                worst = max(worst, float(np.sqrt(np.mean(err ** 2))))
        errs.append(worst)
    print("deriv_two rms error relative to max|f''|, levels 3-5:", errs)
    assert errs[0] < 0.05 and errs[1] < 0.5 * errs[0] and errs[2] < 0.5 * errs[1]


def test_defc_deformation_of_a_solid_body_rotation_vanishes():
    """defc_a / defc_b (mpas_atm_advection.F:744-937) turn (u, v) on the edges of a cell into the two deformation components
    (TI:4677-4693).  A solid-body rotation has none: |D| a / U -> 0 with the mesh spacing, while a pure shear flow
    (zonal wind ~ sin(lat)) keeps an O(1) deformation."""
    res = []
    for level in (3, 4, 5):
        case = common.host.HostCase(level, lloyd_iters=10, sfc=True)
        b = case.build_block(0, nVertLevels=4, init_case=2, num_scalars=1, full_fields=False)
        nE, nC = case.nEdges, case.nCells
        lat = b.field("latEdge")[:nE]
        cx, sx = _edge_normals(b, nE)
        da, db = b.field("defc_a")[:nC], b.field("defc_b")[:nC]
        eoc, nec = b.field("edgesOnCell")[:nC], b.field("nEdgesOnCell")[:nC]
        a = b.cfg()["sphere_radius"]
        out = []
        for U in (np.cos(lat), np.cos(lat) * np.sin(lat)):  # solid body; shear
            un, vt = U * cx, -U * sx
            dd, do = np.zeros(nC), np.zeros(nC)
            for i in range(eoc.shape[1]):
                m = i < nec
                e = eoc[m, i] - 1
                dd[m] += da[m, i] * un[e] - db[m, i] * vt[e]
                do[m] += db[m, i] * un[e] + da[m, i] * vt[e]
            # away from the poles (the rotation of the local east/north frame is singular there)
            ok = np.abs(b.field("latCell")[:nC]) < 1.2
            out.append(float(np.sqrt(np.mean((dd[ok] ** 2 + do[ok] ** 2))) * a))
        res.append(out)
    print("rms deformation * a / U (solid body, shear), levels 3-5:", res)
    assert res[0][0] < 0.1 and res[1][0] < 0.5 * res[0][0] and res[2][0] < 0.5 * res[1][0]
    assert all(r[1] > 0.3 for r in res)  # the shear flow does deform: the operator is not just small


def test_jw_steady_state_initial_tendencies_shrink_with_resolution():
    """init_case = 1 (no perturbation) is an analytic steady state of the continuous equations (mpas_init_atm_cases.F:98):
    what the first dyn_tend evaluation leaves in tend_u is discretisation error only and must fall with the mesh spacing."""
    vals = []
    for level in (3, 4, 5):
        case, blocks = common.make_case(level, 1, nVertLevels=26, init_case=1, num_scalars=1)
        cfg = common.level_config(level)
        m = common.OracleModel(blocks, cfg)
        m.init()
        m.call("rk_setup"); m.call("moist"); m.call("vert_imp_coefs", cfg["config_dt"] / 6)
        m.call("dyn_tend", 1, cfg["config_dt"])
        nE = int(blocks[0].cfg()["nEdgesSolve"])
        tu = m.field(0, "tend_u")[:nE]
        rho_e = m.field(0, "rho_edge")[:nE]
        vals.append(float(np.sqrt(np.mean((tu / rho_e) ** 2))))
    print("rms du/dt of the balanced JW state (m/s^2), levels 3-5:", vals)
    # a 35 m/s jet turning with f ~ 1e-4 has Coriolis and pressure-gradient accelerations of ~3e-3 m/s^2 that must cancel
    assert vals[0] < 1e-3 and vals[1] < 0.7 * vals[0] and vals[2] < 0.7 * vals[1]
