"""mpas_reconstruct (operators/mpas_vector_reconstruction.F:51-305): the radial-basis-function coefficients of the host
library and the per-step reconstruction of the oracle, pinned analytically (the reference ships no test for them):
the interpolant carries a constant basis in the tangent plane, so a constant tangent vector V sampled as u_e = V . n_e on
a cell's edges must come back exactly; and the balanced zonal flow of the Jablonowski-Williamson case must come back
as (u(lat), 0) up to the discretisation error."""
import numpy as np
import pytest

import common


@pytest.fixture(scope="module")
def case():
    case, blocks = common.make_case(4, 1, nVertLevels=8, num_scalars=1, init_case=1)
    return blocks[0]


def test_coefficients_reproduce_a_constant_tangent_vector(case):
    b = case
    nown = int(b.cfg()["nCellsSolve"])
    cr = b.field("coeffs_reconstruct")[:nown]        # (cell, maxEdges, 3)
    env = b.field("edgeNormalVectors")               # (edge, 3)
    ctp = b.field("cellTangentPlane")[:nown]          # (cell, 2, 3)
    eoc = b.field("edgesOnCell")[:nown] - 1
    ne = b.field("nEdgesOnCell")[:nown]
    rng = np.random.default_rng(7)
    worst = 0.0
    for c in range(0, nown, 7):
        a, bb = rng.standard_normal(2)
        V = a * ctp[c, 0] + bb * ctp[c, 1]
        n = int(ne[c])
        ue = env[eoc[c, :n]] @ V
        rec = (cr[c, :n] * ue[:, None]).sum(axis=0)
        worst = max(worst, float(np.abs(rec - V).max() / np.abs(V).max()))
    assert worst < 1e-11, worst
    # unit vectors, orthogonal tangent basis
    assert np.allclose(np.linalg.norm(env[:-1], axis=1), 1.0, atol=1e-14)
    assert np.abs((ctp[:, 0] * ctp[:, 1]).sum(axis=1)).max() < 1e-14
    # cells beyond nCellsSolve keep zero coefficients (includeHalos absent, :112-119)
    assert not b.field("coeffs_reconstruct")[nown:].any()


def test_oracle_reconstructs_the_balanced_zonal_flow(case):
    cfg = common.level_config(4)
    m = common.OracleModel([case], cfg)
    m.init()
    # time level 2 of u is what the step-end call reads (TI:1587); give it the initial state
    m.field(0, "u@2")[...] = m.field(0, "u@1")
    m.call("reconstruct")
    b = case
    nown = int(b.cfg()["nCellsSolve"])
    zon = m.field(0, "uReconstructZonal")[:nown]
    mer = m.field(0, "uReconstructMeridional")[:nown]
    x, y, z = (m.field(0, "uReconstruct" + a)[:nown] for a in "XYZ")
    lat = b.field("latCell")[:nown]
    umax = np.abs(zon).max()
    assert 20.0 < umax < 40.0                      # JW jet, u0 = 35 m/s
    assert np.abs(mer).max() < 0.05 * umax         # zonal flow: meridional part is discretisation error only
    # jet maximum in mid-latitudes of both hemispheres, calm poles
    k = np.unravel_index(np.abs(zon).argmax(), zon.shape)
    assert 0.5 < abs(lat[k[0]]) < 1.1
    assert np.abs(zon[np.abs(lat) > 1.5]).max() < 0.1 * umax
    # the 3-vector is tangent to the sphere
    xc, yc, zc = (b.field(a + "Cell")[:nown, None] for a in "xyz")
    r = np.sqrt(xc ** 2 + yc ** 2 + zc ** 2)
    radial = (x * xc + y * yc + z * zc) / r
    assert np.abs(radial).max() < 1e-9 * umax
    assert np.allclose(x * x + y * y + z * z, zon * zon + mer * mer, rtol=1e-12, atol=1e-12)
