// Restatement of init_atm_case_squall_line (src/core_init_atmosphere/mpas_init_atm_cases.F:1313-1895): the idealised
// squall line (test case 4) and supercell (test case 5) on a doubly periodic planar mesh -- Weisman-Klemp sounding,
// unidirectional shear below zts, a warm bubble of 3 K in the domain centre, hydrostatic balance by fixed-point
// iteration, no terrain, f = 0.  BASELINE.json configs[2].
//
// Decomposition: the reference takes the sounding column from element (k, 1) of the IO task's block and broadcasts it
// (pitop, pibtop, ptop, u_init, t_init, qv_init: mpas_dmpar_bcast_real(s)).  On the flat plane every column is the same
// until the bubble is added, and cell 1 of a periodic_hex mesh sits in the corner, ~40 km from the bubble; here that
// column is carried as a separate 1-D "corner column" (identical arithmetic), so every block of every partition builds
// bit-identical fields without communication.
#include <algorithm>
#include <cmath>
#include <vector>

#include "init_atm.h"

namespace mpas {

void init_case_squall_line(const Mesh& m, const InitConfig& cfg, Pool& p) {
    using K = Constants;
    const int nCells = m.nCells, nEdges = m.nEdges, nVertices = m.nVertices, mE = m.maxEdges;
    const int nz1 = cfg.nVertLevels, nz = nz1 + 1, L = nz1, ns = cfg.num_scalars;
    const int test_case = cfg.init_case;
    const double pii = K::pii, gravity = K::gravity, rgas = K::rgas, cp = K::cp, cv = K::cv;
    const double p0 = 1.e+05, rcp = rgas / cp, rcv = rgas / (cp - rgas);
    const int index_qv = 1;

    // ---- vertical grid (:1503-1553): str = 1, zt = 20 km, pure height coordinate (ah = 1), no terrain
    Field& f_rdzw = p.addr("rdzw", LOC_NONE, {L});
    Field& f_dzu = p.addr("dzu", LOC_NONE, {L});
    Field& f_rdzu = p.addr("rdzu", LOC_NONE, {L});
    Field& f_fzm = p.addr("fzm", LOC_NONE, {L});
    Field& f_fzp = p.addr("fzp", LOC_NONE, {L});
    double *rdzw = f_rdzw.r.data(), *dzu = f_dzu.r.data(), *rdzu = f_rdzu.r.data(), *fzm = f_fzm.r.data(), *fzp = f_fzp.r.data();
    const double str = 1.0, zt = 20000.;
    const double dz = zt / (double)nz1;
    std::vector<double> zc(nz), zw(nz), ah(nz), dzw(nz1), zu(nz1);
    for (int k = 0; k < nz; ++k) {
        zc[k] = zt * std::pow((double)k * dz / zt, str);
        zw[k] = zc[k];
        ah[k] = 1.;
    }
    for (int k = 0; k < nz1; ++k) {
        dzw[k] = zw[k + 1] - zw[k];
        rdzw[k] = 1. / dzw[k];
        zu[k] = .5 * (zw[k] + zw[k + 1]);
    }
    for (int k = 1; k < nz1; ++k) {
        dzu[k] = .5 * (dzw[k] + dzw[k - 1]);
        rdzu[k] = 1. / dzu[k];
        fzp[k] = .5 * dzw[k] / dzu[k];
        fzm[k] = .5 * dzw[k - 1] / dzu[k];
    }
    {
        const double cof1 = (2. * dzu[1] + dzu[2]) / (dzu[1] + dzu[2]) * dzw[0] / dzu[1];
        const double cof2 = dzu[1] / (dzu[1] + dzu[2]) * dzw[0] / dzu[2];
        p.cfg["cf1"] = fzp[1] + cof1;
        p.cfg["cf2"] = fzm[1] - cof1 - cof2;
        p.cfg["cf3"] = cof2;
    }

    Field& f_zgrid = p.addr("zgrid", LOC_CELL, {L + 1, nCells + 1});
    Field& f_zz = p.addr("zz", LOC_CELL, {L, nCells + 1});
    Field& f_zxu = p.addr("zxu", LOC_EDGE, {L, nEdges + 1});
    Field& f_rb = p.addr("rho_base", LOC_CELL, {L, nCells + 1});
    Field& f_tb = p.addr("theta_base", LOC_CELL, {L, nCells + 1});
    Field& f_rho = p.addr("rho", LOC_CELL, {L, nCells + 1});
    Field& f_theta = p.addr("theta", LOC_CELL, {L, nCells + 1});
    Field& f_u = p.addr("u@1", LOC_EDGE, {L, nEdges + 1});
    p.addr("w@1", LOC_CELL, {L + 1, nCells + 1});  // w = rw = 0: no terrain (:1836-1837)
    Field& f_sc = p.addr("scalars@1", LOC_CELL, {ns, L, nCells + 1});
    Field& f_uinit = p.addr("u_init", LOC_NONE, {L});
    p.addr("v_init", LOC_NONE, {L});
    Field& f_qvinit = p.addr("qv_init", LOC_NONE, {L});
    Field& f_tinit = p.addr("t_init", LOC_CELL, {L, nCells + 1});
    double *zgrid = f_zgrid.r.data(), *zz = f_zz.r.data(), *zxu = f_zxu.r.data(), *rb = f_rb.r.data(), *tb = f_tb.r.data();
    double *rho = f_rho.r.data(), *theta = f_theta.r.data(), *u = f_u.r.data(), *scalars = f_sc.r.data();
    auto QV = [&](int k, int c) -> double& { return scalars[((size_t)c * L + k) * ns + (index_qv - 1)]; };

    for (int c = 0; c < nCells; ++c) {
        for (int k = 0; k < nz; ++k) zgrid[(size_t)c * nz + k] = ah[k] * (zc[k] * (1. - 0. / zt) + 0.) + (1. - ah[k]) * zc[k];  // hx = 0
        for (int k = 0; k < nz1; ++k) zz[(size_t)c * L + k] = (zw[k + 1] - zw[k]) / (zgrid[(size_t)c * nz + k + 1] - zgrid[(size_t)c * nz + k]);
    }
    for (int e = 0; e < nEdges; ++e) {
        const int c1 = m.cellsOnEdge[2 * e] - 1, c2 = m.cellsOnEdge[2 * e + 1] - 1;
        if (c1 >= nCells || c2 >= nCells) continue;
        for (int k = 0; k < nz1; ++k)
            zxu[(size_t)e * L + k] = 0.5 * (zgrid[(size_t)c2 * nz + k] - zgrid[(size_t)c1 * nz + k] + zgrid[(size_t)c2 * nz + k + 1] - zgrid[(size_t)c1 * nz + k + 1]) / m.dcEdge[e];
    }

    // ---- convective sounding (:1583-1627); column index nCells = the corner column (see the header)
    const double ztr = 12000., thetar = 343., ttr = 213., thetas = 300.5;
    double um, us, zts;
    if (test_case == 4) { um = 12.; us = 10.; zts = 2500.; }
    else { um = 30.; us = 15.; zts = 5000.; }
    const int NC1 = nCells + 1;  // columns 0..nCells-1 = cells, nCells = corner column
    std::vector<double> t((size_t)NC1 * L), relhum((size_t)NC1 * L), thi((size_t)NC1 * L), tbi((size_t)NC1 * L), cqw((size_t)NC1 * L), cqwb((size_t)NC1 * L);
    std::vector<double> tbw((size_t)NC1 * L), qv((size_t)NC1 * L, 0.0), pb((size_t)NC1 * L), pex((size_t)NC1 * L), rbw((size_t)NC1 * L), rtb((size_t)NC1 * L),
        rr((size_t)NC1 * L), ppb((size_t)NC1 * L), pp((size_t)NC1 * L), rt((size_t)NC1 * L);
    auto zg = [&](int k, int c) { return c < nCells ? zgrid[(size_t)c * nz + k] : zc[k]; };  // flat: the corner column has the same heights
    auto ZZ = [&](int k, int c) { return c < nCells ? zz[(size_t)c * L + k] : 1.0; };
    for (int c = 0; c < NC1; ++c)
        for (int k = 0; k < nz1; ++k) {
            const size_t o = (size_t)c * L + k;
            const double ztemp = .5 * (zg(k, c) + zg(k + 1, c));
            if (ztemp > ztr) {
                t[o] = thetar * std::exp(9.8 * (ztemp - ztr) / (1003. * ttr));
                relhum[o] = 0.25;
            } else {
                t[o] = 300. + 43. * std::pow(ztemp / ztr, 1.25);
                relhum[o] = (1. - 0.75 * std::pow(ztemp / ztr, 1.25));
                if (t[o] < thetas) t[o] = thetas;
            }
            tbw[o] = t[o]; thi[o] = t[o]; tbi[o] = t[o];
            cqw[o] = 1.; cqwb[o] = 1.;
        }
    // ---- velocity (:1633-1651): on a plane, u = cos(angleEdge) * (U(z) - us); u_init = U(z) - us
    for (int k = 0; k < nz1; ++k) {
        const double ztemp = .25 * (zc[k] + zc[k + 1] + zc[k] + zc[k + 1]);
        f_uinit.r[k] = (ztemp < zts ? um * ztemp / zts : um) - us;
    }
    for (int e = 0; e < nEdges; ++e) {
        const int c1 = m.cellsOnEdge[2 * e] - 1, c2 = m.cellsOnEdge[2 * e + 1] - 1;
        if (c1 >= nCells || c2 >= nCells) continue;
        for (int k = 0; k < nz1; ++k) {
            const double ztemp = .25 * (zgrid[(size_t)c1 * nz + k] + zgrid[(size_t)c1 * nz + k + 1] + zgrid[(size_t)c2 * nz + k] + zgrid[(size_t)c2 * nz + k + 1]);
            double uk = ztemp < zts ? um * ztemp / zts : um;
            u[(size_t)e * L + k] = std::cos(m.angleEdge[e]) * (uk - us);
        }
    }

    // ---- reference sounding, 30 fixed-point iterations (:1658-1737)
    std::vector<double> qvb(nz1, 0.0);
    const int C1 = nCells;  // "(k, 1)" of the reference
    double ptopb = 0.0;
    for (int itr = 1; itr <= 30; ++itr) {
        double pitop = 1. - .5 * dzw[0] * gravity * (1. + qv[(size_t)C1 * L]) / (cp * t[(size_t)C1 * L] * ZZ(0, C1));
        double pibtop = 1. - .5 * dzw[0] * gravity * (1. + qvb[0]) / (cp * tbw[(size_t)C1 * L] * ZZ(0, C1));
        for (int k = 1; k < nz1; ++k) {
            pitop = pitop - dzu[k] * gravity / (cp * cqw[(size_t)C1 * L + k] * .5 * (t[(size_t)C1 * L + k] + t[(size_t)C1 * L + k - 1]) * .5 * (ZZ(k, C1) + ZZ(k - 1, C1)));
            pibtop = pibtop - dzu[k] * gravity / (cp * cqwb[(size_t)C1 * L + k] * .5 * (tbw[(size_t)C1 * L + k] + tbw[(size_t)C1 * L + k - 1]) * .5 * (ZZ(k, C1) + ZZ(k - 1, C1)));
        }
        pitop = pitop - .5 * dzw[nz1 - 1] * gravity * (1. + qv[(size_t)C1 * L + nz1 - 1]) / (cp * t[(size_t)C1 * L + nz1 - 1] * ZZ(nz1 - 1, C1));
        pibtop = pibtop - .5 * dzw[nz1 - 1] * gravity * (1. + qvb[nz1 - 1]) / (cp * tbw[(size_t)C1 * L + nz1 - 1] * ZZ(nz1 - 1, C1));
        ptopb = p0 * std::pow(pibtop, 1. / rcp);
#pragma omp parallel for schedule(static)
        for (int c = 0; c < NC1; ++c) {
            const size_t o = (size_t)c * L;
            pb[o + nz1 - 1] = pibtop + .5 * dzw[nz1 - 1] * gravity * (1. + qvb[nz1 - 1]) / (cp * tbw[o + nz1 - 1] * ZZ(nz1 - 1, c));
            pex[o + nz1 - 1] = pitop + .5 * dzw[nz1 - 1] * gravity * (1. + qv[o + nz1 - 1]) / (cp * t[o + nz1 - 1] * ZZ(nz1 - 1, c));
            for (int k = nz1 - 2; k >= 0; --k) {
                pb[o + k] = pb[o + k + 1] + dzu[k + 1] * gravity / (cp * cqwb[o + k + 1] * .5 * (tbw[o + k] + tbw[o + k + 1]) * .5 * (ZZ(k, c) + ZZ(k + 1, c)));
                pex[o + k] = pex[o + k + 1] + dzu[k + 1] * gravity / (cp * cqw[o + k + 1] * .5 * (t[o + k] + t[o + k + 1]) * .5 * (ZZ(k, c) + ZZ(k + 1, c)));
            }
            for (int k = 0; k < nz1; ++k) {
                rbw[o + k] = std::pow(pb[o + k], 1. / rcv) / ((rgas / p0) * tbw[o + k] * ZZ(k, c));
                rtb[o + k] = rbw[o + k] * tbw[o + k];
                rr[o + k] = std::pow(pex[o + k], 1. / rcv) / ((rgas / p0) * t[o + k] * ZZ(k, c)) - rbw[o + k];
                ppb[o + k] = p0 * std::pow(ZZ(k, c) * rgas * rtb[o + k] / p0, cp / cv);
            }
            // water vapour from the humidity profile (:1712-1720)
            for (int k = 0; k < nz1; ++k) {
                const double temp = pex[o + k] * thi[o + k];
                const double pres = p0 * std::pow(pex[o + k], 1. / rcp);
                const double qvs = 380. * std::exp(17.27 * (temp - 273.) / (temp - 36.)) / pres;
                qv[o + k] = std::min(0.014, relhum[o + k] * qvs);
            }
        }
        for (int k = 0; k < nz1; ++k) qvb[k] = 0.;  // dry reference state (:1729)
#pragma omp parallel for schedule(static)
        for (int c = 0; c < NC1; ++c) {
            const size_t o = (size_t)c * L;
            for (int k = 0; k < nz1; ++k) {
                t[o + k] = thi[o + k] * (1. + 1.61 * qv[o + k]);
                tbw[o + k] = tbi[o + k] * (1. + 1.61 * qvb[k]);
            }
            for (int k = 1; k < nz1; ++k) {
                cqw[o + k] = 1. / (1. + .5 * (qv[o + k] + qv[o + k - 1]));
                cqwb[o + k] = 1. / (1. + .5 * (qvb[k] + qvb[k - 1]));
            }
        }
    }

    // ---- potential temperature perturbation: the warm bubble (:1748-1785)
    const double delt = 3., radx = 10000., radz = 1500., zcent = 1500.;
    const double xmid = m.planar_xmax * 0.5, ymid = test_case == 4 ? 0.0 : m.planar_ymax * 0.5;  // global maxima (mpas_dmpar_max_real)
    for (int c = 0; c < nCells; ++c) {
        const double xloc = m.xCell[c] - xmid;
        const double yloc = test_case == 4 ? 0. : m.yCell[c] - ymid;
        for (int k = 0; k < nz1; ++k) {
            const size_t o = (size_t)c * L + k;
            const double ztemp = .5 * (zg(k + 1, c) + zg(k, c));
            const double rad = std::sqrt((xloc / radx) * (xloc / radx) + (yloc / radx) * (yloc / radx) + ((ztemp - zcent) / radz) * ((ztemp - zcent) / radz));
            if (rad < 1) {
                const double cc = std::cos(.5 * pii * rad);
                thi[o] = thi[o] + delt * (cc * cc);
            }
            t[o] = thi[o] * (1. + 1.61 * qv[o]);
        }
    }
    for (int k = 0; k < nz1; ++k) t[(size_t)C1 * L + k] = thi[(size_t)C1 * L + k] * (1. + 1.61 * qv[(size_t)C1 * L + k]);

    // ---- re-balance, 30 iterations (:1787-1825)
    for (int itr = 1; itr <= 30; ++itr) {
        double pitop = 1. - .5 * dzw[0] * gravity * (1. + qv[(size_t)C1 * L]) / (cp * t[(size_t)C1 * L] * ZZ(0, C1));
        for (int k = 1; k < nz1; ++k)
            pitop = pitop - dzu[k] * gravity / (cp * cqw[(size_t)C1 * L + k] * .5 * (t[(size_t)C1 * L + k] + t[(size_t)C1 * L + k - 1]) * .5 * (ZZ(k, C1) + ZZ(k - 1, C1)));
        pitop = pitop - .5 * dzw[nz1 - 1] * gravity * (1. + qv[(size_t)C1 * L + nz1 - 1]) / (cp * t[(size_t)C1 * L + nz1 - 1] * ZZ(nz1 - 1, C1));
        const double ptop = p0 * std::pow(pitop, 1. / rcp);
#pragma omp parallel for schedule(static)
        for (int c = 0; c < NC1; ++c) {
            const size_t o = (size_t)c * L;
            pp[o + nz1 - 1] = ptop - ptopb + .5 * dzw[nz1 - 1] * gravity * (rr[o + nz1 - 1] + (rr[o + nz1 - 1] + rbw[o + nz1 - 1]) * qv[o + nz1 - 1]);
            for (int k = nz1 - 2; k >= 0; --k)
                pp[o + k] = pp[o + k + 1] + dzu[k + 1] * gravity * (fzm[k + 1] * (rbw[o + k + 1] * (qv[o + k + 1] - qvb[k + 1]) + rr[o + k + 1] * (1. + qv[o + k + 1])) +
                                                                   fzp[k + 1] * (rbw[o + k] * (qv[o + k] - qvb[k]) + rr[o + k] * (1. + qv[o + k])));
            for (int k = 0; k < nz1; ++k) {
                rt[o + k] = (pp[o + k] / (rgas * ZZ(k, c)) - rtb[o + k] * (pex[o + k] - pb[o + k])) / pex[o + k];
                pex[o + k] = std::pow(ZZ(k, c) * (rgas / p0) * (rtb[o + k] + rt[o + k]), rcv);
                rr[o + k] = (rt[o + k] - rbw[o + k] * (t[o + k] - tbw[o + k])) / t[o + k];
            }
        }
    }

    // ---- results into the pool (:1828-1891)
    for (int k = 0; k < nz1; ++k) f_qvinit.r[k] = qv[(size_t)C1 * L + k];
    for (int c = 0; c < nCells; ++c)
        for (int k = 0; k < nz1; ++k) {
            const size_t o = (size_t)c * L + k;
            f_tinit.r[o] = t[(size_t)C1 * L + k];
            QV(k, c) = qv[o];
            rb[o] = rbw[o];
            tb[o] = tbw[o];
            const double rho_zz = rbw[o] + rr[o];
            rho[o] = rho_zz * zz[o];
            theta[o] = t[o] / (1.0 + 1.61 * qv[o]);
        }
    // f = 0 on the plane (:1845-1851); no terrain: zb = zb3 = 0 (:1839-1840)
    p.addr("fEdge", LOC_EDGE, {nEdges + 1});
    p.addr("fVertex", LOC_VERTEX, {nVertices + 1});
    p.addr("zb_cell", LOC_CELL, {L + 1, mE, nCells + 1});
    p.addr("zb3_cell", LOC_CELL, {L + 1, mE, nCells + 1});
    if (cfg.keep_zb) {
        p.addr("zb", LOC_EDGE, {L + 1, 2, nEdges + 1});
        p.addr("zb3", LOC_EDGE, {L + 1, 2, nEdges + 1});
    }
    p.cfg["nVertLevels"] = L;
    p.cfg["num_scalars"] = ns;
    // qv, qc, qr are the moist species (Registry.xml:1433-1453 with the Kessler package): all of them enter qtot (TI:1903-1905)
    p.cfg["moist_start"] = 1;
    p.cfg["moist_end"] = std::min(ns, 3);
    p.cfg["index_qv"] = index_qv;
    // the namelist of the reference's idealised supercell case: fixed eddy viscosities, mixing of the perturbation from the
    // initial sounding (config_mix_full = false uses u_init, v_init, t_init: TI:4949-5006, 5360-5406), no upper damping layer
    p.cfg["config_horiz_mixing"] = 2;  // '2d_fixed'
    p.cfg["config_h_mom_eddy_visc2"] = 500.;
    p.cfg["config_h_theta_eddy_visc2"] = 500.;
    p.cfg["config_v_mom_eddy_visc2"] = 500.;
    p.cfg["config_v_theta_eddy_visc2"] = 500.;
    p.cfg["config_h_mom_eddy_visc4"] = 0.;
    p.cfg["config_h_theta_eddy_visc4"] = 0.;
    p.cfg["config_mix_full"] = 0;
    p.cfg["config_zd"] = 12000.;
    p.cfg["config_xnutr"] = 0.;
}

}  // namespace mpas
