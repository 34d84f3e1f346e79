// Restatement of init_atm_case_jw (src/core_init_atmosphere/mpas_init_atm_cases.F:367-1160)
// with init_atm_calc_flux_zonal (:1163-1211) and init_atm_recompute_geostrophic_wind
// (:1215-1310): vertical grid, terrain-following metrics, balanced Jablonowski-Williamson
// state.  Dry (moisture=.false., :454).  Deviation from the reference, documented in
// the scatter loop that accumulates rw over edges (:1098-1119) is evaluated in
// gather form (per cell, edgesOnCell slot order) so the result does not depend on edge
// numbering or decomposition.
#include <cmath>
#include <vector>

#include "init_atm.h"

namespace mpas {
namespace {
inline double p2(double x) { return x * x; }
inline double p3(double x) { return x * x * x; }
inline double p5(double x) { return x * x * x * x * x; }
inline double p6(double x) { double y = x * x * x; return y * y; }
inline double sgn1(double x) { return std::signbit(x) ? -1.0 : 1.0; }  // Fortran sign(1.0,x)

double sphere_distance(double lat1, double lon1, double lat2, double lon2, double radius) {
    // src/core_init_atmosphere/mpas_init_atm_cases.F sphere_distance (haversine form)
    double arg1 = std::sqrt(p2(std::sin(0.5 * (lat2 - lat1))) + std::cos(lat1) * std::cos(lat2) * p2(std::sin(0.5 * (lon2 - lon1))));
    return 2. * radius * std::asin(arg1);
}

// :1163-1211
void calc_flux_zonal(const std::vector<double>& u_2d, const std::vector<double>& lat_2d, double* flux_zonal, double lat1_in,
                     double lat2_in, double dvEdge, double a, double u0, int nz1, int nlat) {
    double lat1 = std::fabs(lat1_in), lat2 = std::fabs(lat2_in);
    if (lat2 <= lat1) { lat1 = std::fabs(lat2_in); lat2 = std::fabs(lat1_in); }
    for (int k = 0; k < nz1; ++k) flux_zonal[k] = 0.;
    double dlat = 0.;
    for (int i = 0; i < nlat - 1; ++i) {
        if ((lat1 <= lat_2d[i + 1]) && (lat2 >= lat_2d[i])) {
            dlat = lat_2d[i + 1] - lat_2d[i];
            double da = (std::max(lat1, lat_2d[i]) - lat_2d[i]) / dlat;
            double db = (std::min(lat2, lat_2d[i + 1]) - lat_2d[i]) / dlat;
            double w1 = (db - da) - 0.5 * p2(db - da);
            double w2 = 0.5 * p2(db - da);
            for (int k = 0; k < nz1; ++k) flux_zonal[k] = flux_zonal[k] + w1 * u_2d[(size_t)i * nz1 + k] + w2 * u_2d[(size_t)(i + 1) * nz1 + k];
        }
    }
    // note: like the reference, dlat here is the value left over from the loop (:1208)
    if (dlat == 0.) dlat = lat_2d[1] - lat_2d[0];
    for (int k = 0; k < nz1; ++k) flux_zonal[k] = sgn1(lat2_in - lat1_in) * flux_zonal[k] * dlat * a / dvEdge / u0;
}
}  // namespace

void init_case_jw(const Mesh& m, const InitConfig& cfg, Pool& p) {
    using K = Constants;
    const int nCells = m.nCells, nEdges = m.nEdges, nVertices = m.nVertices, mE = m.maxEdges;
    const int nz1 = cfg.nVertLevels, nz = nz1 + 1, L = nz1;
    const double u0 = 35.0, t0b = 250., t0 = 288., delta_t = 4.8e+05, dtdz = 0.005, eta_t = 0.2;
    const double u_perturbation = 1., pert_radius = 0.1, latitude_pert = 40., longitude_pert = 20.;
    const double k_x = 9.;
    const double pii = K::pii, gravity = K::gravity, rgas = K::rgas, cp = K::cp;
    const double sphere_radius = m.sphere_radius, r_earth = m.sphere_radius, omega_e = K::omega, p0 = 1.0e+05;
    const double znut = eta_t;
    const double etavs = (1.0 - 0.252) * pii / 2.;
    const int nlat = 721;

    // ---- vertical grid (:617-669)
    std::vector<double> sh(nz), zw(nz), ah(nz), dzw(nz1), zu(nz1);
    Field& f_rdzw = p.addr("rdzw", LOC_NONE, {L});
    Field& f_dzu = p.addr("dzu", LOC_NONE, {L});
    Field& f_rdzu = p.addr("rdzu", LOC_NONE, {L});
    Field& f_fzm = p.addr("fzm", LOC_NONE, {L});
    Field& f_fzp = p.addr("fzp", LOC_NONE, {L});
    double *rdzw = f_rdzw.r.data(), *dzu = f_dzu.r.data(), *rdzu = f_rdzu.r.data(), *fzm = f_fzm.r.data(), *fzp = f_fzp.r.data();
    const double str = 1.5, zt = 45000.;
    const double dz = zt / (double)nz1;
    for (int k = 0; k < nz; ++k) {
        sh[k] = std::pow((double)k * dz / zt, str);
        zw[k] = (double)k * dz;
        ah[k] = 1. - p6(std::cos(.5 * pii * k * dz / zt));
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
    double cof1 = (2. * dzu[1] + dzu[2]) / (dzu[1] + dzu[2]) * dzw[0] / dzu[1];
    double cof2 = dzu[1] / (dzu[1] + dzu[2]) * dzw[0] / dzu[2];
    double cf1 = fzp[1] + cof1, cf2 = fzm[1] - cof1 - cof2, cf3 = cof2;
    p.cfg["cf1"] = cf1; p.cfg["cf2"] = cf2; p.cfg["cf3"] = cf3;

    auto hx_of = [&](double phi) {  // :606-611
        return u0 / gravity * std::pow(std::cos(etavs), 1.5) *
               ((-2. * p6(std::sin(phi)) * (p2(std::cos(phi)) + 1. / 3.) + 10. / 63.) * (u0)*std::pow(std::cos(etavs), 1.5) +
                (1.6 * p3(std::cos(phi)) * (p2(std::sin(phi)) + 2. / 3.) - pii / 4.) * r_earth * omega_e);
    };
    auto temperature = [&](double teta, double eta, double etav, double phi, double qv) {  // :875-881
        return teta + .75 * eta * pii * u0 / rgas * std::sin(etav) * std::sqrt(std::cos(etav)) *
                          ((-2. * p6(std::sin(phi)) * (p2(std::cos(phi)) + 1. / 3.) + 10. / 63.) * 2. * u0 * std::pow(std::cos(etav), 1.5) +
                           (1.6 * p3(std::cos(phi)) * (p2(std::sin(phi)) + 2. / 3.) - pii / 4.) * r_earth * omega_e) /
                          (1. + 0.61 * qv);
    };
    // balanced column: given zgrid column & zz column, returns pp, rr, tt (and base state)
    auto balance_column = [&](const double* zg, const double* zzc, double phi, double* ppb, double* pb, double* rb, double* tb,
                              double* pp, double* rr, double* tt) {
        std::vector<double> eta(nz1), etav(nz1), teta(nz1), ppi(nz1);
        for (int k = 0; k < nz1; ++k) {
            double ztemp = .5 * (zg[k + 1] + zg[k]);
            ppb[k] = p0 * std::exp(-gravity * ztemp / (rgas * t0b));
            pb[k] = std::pow(ppb[k] / p0, rgas / cp);
            rb[k] = ppb[k] / (rgas * t0b * zzc[k]);
            tb[k] = t0b / pb[k];
            pp[k] = 0.;
            rr[k] = 0.;
        }
        for (int itr = 0; itr < 10; ++itr) {
            for (int k = 0; k < nz1; ++k) {
                eta[k] = (ppb[k] + pp[k]) / p0;
                etav[k] = (eta[k] - .252) * pii / 2.;
                if (eta[k] >= znut) teta[k] = t0 * std::pow(eta[k], rgas * dtdz / gravity);
                else teta[k] = t0 * std::pow(eta[k], rgas * dtdz / gravity) + delta_t * p5(znut - eta[k]);
            }
            for (int k = 0; k < nz1; ++k) {
                double t1d = temperature(teta[k], eta[k], etav[k], phi, 0.0);
                tt[k] = t1d * (1. + 1.61 * 0.0);
            }
            for (int itrp = 0; itrp < 25; ++itrp) {
                for (int k = 0; k < nz1; ++k) rr[k] = (pp[k] / (rgas * zzc[k]) - rb[k] * (tt[k] - t0b)) / tt[k];
                ppi[0] = p0 - .5 * dzw[0] * gravity * (1.25 * (rr[0] + rb[0]) * (1. + 0.) - .25 * (rr[1] + rb[1]) * (1. + 0.));
                ppi[0] = ppi[0] - ppb[0];
                for (int k = 0; k < nz1 - 1; ++k)
                    ppi[k + 1] = ppi[k] - dzu[k + 1] * gravity * ((rr[k] + (rr[k] + rb[k]) * 0.) * fzp[k + 1] + (rr[k + 1] + (rr[k + 1] + rb[k + 1]) * 0.) * fzm[k + 1]);
                for (int k = 0; k < nz1; ++k) pp[k] = .2 * ppi[k] + .8 * pp[k];
            }
        }
    };

    // ---- 2-D (lat,z) table for the balanced zonal wind (:720-832)
    std::vector<double> lat_2d(nlat), u_2d((size_t)nlat * nz1), pp_2d((size_t)nlat * nz1), rho_2d((size_t)nlat * nz1),
        zz_2d((size_t)nlat * nz1), zgrid_2d((size_t)nlat * nz), zx_2d((size_t)(nlat - 1) * nz1);
    const double dlat = 0.5 * pii / (double)(nlat - 1);
#pragma omp parallel for schedule(dynamic, 8)
    for (int i = 0; i < nlat; ++i) {
        lat_2d[i] = (double)i * dlat;
        double phi = lat_2d[i];
        double hx_1d = hx_of(phi);
        double* zg = &zgrid_2d[(size_t)i * nz];
        for (int k = 0; k < nz; ++k) zg[k] = (1. - ah[k]) * (sh[k] * (zt - hx_1d) + hx_1d) + ah[k] * sh[k] * zt;
        double* zzc = &zz_2d[(size_t)i * nz1];
        for (int k = 0; k < nz1; ++k) zzc[k] = (zw[k + 1] - zw[k]) / (zg[k + 1] - zg[k]);
        std::vector<double> ppb(nz1), pb(nz1), rb(nz1), tb(nz1), rr(nz1), tt(nz1);
        balance_column(zg, zzc, phi, ppb.data(), pb.data(), rb.data(), tb.data(), &pp_2d[(size_t)i * nz1], rr.data(), tt.data());
        for (int k = 0; k < nz1; ++k) {
            rho_2d[(size_t)i * nz1 + k] = rr[k] + rb[k];
            double etavs_2d = ((ppb[k] + pp_2d[(size_t)i * nz1 + k]) / p0 - 0.252) * pii / 2.;
            u_2d[(size_t)i * nz1 + k] = u0 * p2(std::sin(2. * lat_2d[i])) * std::pow(std::cos(etavs_2d), 1.5);
        }
    }
    for (int i = 0; i < nlat - 1; ++i)
        for (int k = 0; k < nz1; ++k) zx_2d[(size_t)i * nz1 + k] = (zgrid_2d[(size_t)(i + 1) * nz + k] - zgrid_2d[(size_t)i * nz + k]) / (dlat * r_earth);
    {  // init_atm_recompute_geostrophic_wind (:1215-1310), qv = 0
        std::vector<double> pgrad((size_t)(nlat - 1) * nz1), ru((size_t)(nlat - 1) * nz1), u((size_t)(nlat - 1) * nz1), dpzx(nz1 + 1);
        double rdx = 1. / (dlat * r_earth);
        auto PP = [&](int k, int i) { return pp_2d[(size_t)i * nz1 + k]; };
        auto ZZ = [&](int k, int i) { return zz_2d[(size_t)i * nz1 + k]; };
        auto ZX = [&](int k, int i) { return zx_2d[(size_t)i * nz1 + k]; };
        for (int i = 0; i < nlat - 1; ++i) {
            for (int k = 0; k < nz1; ++k) pgrad[(size_t)i * nz1 + k] = rdx * (PP(k, i + 1) / ZZ(k, i + 1) - PP(k, i) / ZZ(k, i));
            for (auto& d : dpzx) d = 0.;
            dpzx[0] = .5 * ZX(0, i) * (cf1 * (PP(0, i + 1) + PP(0, i)) + cf2 * (PP(1, i + 1) + PP(1, i)) + cf3 * (PP(2, i + 1) + PP(2, i)));
            for (int k = 1; k < nz1; ++k) dpzx[k] = .5 * ZX(k, i) * (fzm[k] * (PP(k, i + 1) + PP(k, i)) + fzp[k] * (PP(k - 1, i + 1) + PP(k - 1, i)));
            for (int k = 0; k < nz1; ++k) pgrad[(size_t)i * nz1 + k] = pgrad[(size_t)i * nz1 + k] - rdzw[k] * (dpzx[k + 1] - dpzx[k]);
        }
        for (int i = 0; i < nlat - 1; ++i)
            for (int k = 0; k < nz1; ++k) {
                u[(size_t)i * nz1 + k] = .5 * (u_2d[(size_t)i * nz1 + k] + u_2d[(size_t)(i + 1) * nz1 + k]);
                ru[(size_t)i * nz1 + k] = u[(size_t)i * nz1 + k] * (rho_2d[(size_t)i * nz1 + k] + rho_2d[(size_t)(i + 1) * nz1 + k]) * .5;
            }
        for (int itr = 0; itr < 50; ++itr)
            for (int i = 0; i < nlat - 1; ++i) {
                double phi = (lat_2d[i] + lat_2d[i + 1]) / 2.;
                double f = 2. * omega_e * std::sin(phi);
                for (int k = 0; k < nz1; ++k) {
                    size_t q = (size_t)i * nz1 + k;
                    if (f == 0.) ru[q] = 0.;
                    else ru[q] = -(1. / (1. + 0.) * pgrad[q] + std::tan(phi) / r_earth * u[q] * ru[q]) / f;
                    u[q] = ru[q] * 2. / (rho_2d[q] + rho_2d[q + nz1]);
                }
            }
        for (int i = 1; i < nlat - 1; ++i)
            for (int k = 0; k < nz1; ++k) u_2d[(size_t)i * nz1 + k] = (ru[(size_t)(i - 1) * nz1 + k] + ru[(size_t)i * nz1 + k]) * .5;
        for (int k = 0; k < nz1; ++k) u_2d[k] = (3. * u_2d[(size_t)1 * nz1 + k] - u_2d[(size_t)2 * nz1 + k]) * .5;
        for (int k = 0; k < nz1; ++k)
            u_2d[(size_t)(nlat - 1) * nz1 + k] = (3. * u_2d[(size_t)(nlat - 2) * nz1 + k] - u_2d[(size_t)(nlat - 3) * nz1 + k]) * .5;
    }

    // ---- 3-D metrics (:680-696)
    Field& f_zgrid = p.addr("zgrid", LOC_CELL, {L + 1, nCells + 1});
    Field& f_zz = p.addr("zz", LOC_CELL, {L, nCells + 1});
    Field& f_zxu = p.addr("zxu", LOC_EDGE, {L, nEdges + 1});
    double *zgrid = f_zgrid.r.data(), *zz = f_zz.r.data(), *zxu = f_zxu.r.data();
    Field& f_rb = p.addr("rho_base", LOC_CELL, {L, nCells + 1});
    Field& f_tb = p.addr("theta_base", LOC_CELL, {L, nCells + 1});
    Field& f_rho = p.addr("rho", LOC_CELL, {L, nCells + 1});
    Field& f_theta = p.addr("theta", LOC_CELL, {L, nCells + 1});
    Field& f_u = p.addr("u@1", LOC_EDGE, {L, nEdges + 1});
    Field& f_w = p.addr("w@1", LOC_CELL, {L + 1, nCells + 1});
    Field& f_sc = p.addr("scalars@1", LOC_CELL, {cfg.num_scalars, L, nCells + 1});
    (void)f_sc;
    double *rb = f_rb.r.data(), *tb = f_tb.r.data(), *rho = f_rho.r.data(), *theta = f_theta.r.data();
    double *u = f_u.r.data(), *w = f_w.r.data();
    std::vector<double> rho_zz((size_t)L * (nCells + 1), 0.0), pp((size_t)L * (nCells + 1), 0.0), rr((size_t)L * (nCells + 1), 0.0);
    std::vector<double> ru((size_t)L * (nEdges + 1), 0.0);

#pragma omp parallel for schedule(static)
    for (int c = 0; c < nCells; ++c) {
        double hx = hx_of(m.latCell[c]);
        double* zg = &zgrid[(size_t)c * nz];
        for (int k = 0; k < nz; ++k) zg[k] = (1. - ah[k]) * (sh[k] * (zt - hx) + hx) + ah[k] * sh[k] * zt;
        for (int k = 0; k < nz1; ++k) zz[(size_t)c * L + k] = (zw[k + 1] - zw[k]) / (zg[k + 1] - zg[k]);
    }
#pragma omp parallel for schedule(static)
    for (int e = 0; e < nEdges; ++e) {
        int c1 = m.cellsOnEdge[2 * e] - 1, c2 = m.cellsOnEdge[2 * e + 1] - 1;
        for (int k = 0; k < nz1; ++k)
            zxu[(size_t)e * L + k] = 0.5 * (zgrid[(size_t)c2 * nz + k] - zgrid[(size_t)c1 * nz + k] + zgrid[(size_t)c2 * nz + k + 1] - zgrid[(size_t)c1 * nz + k + 1]) / m.dcEdge[e];
    }
    // ---- balanced state per cell (:841-958)
#pragma omp parallel for schedule(dynamic, 64)
    for (int c = 0; c < nCells; ++c) {
        std::vector<double> ppb(nz1), pb(nz1), tt(nz1);
        size_t o = (size_t)c * L;
        balance_column(&zgrid[(size_t)c * nz], &zz[o], m.latCell[c], ppb.data(), pb.data(), &rb[o], &tb[o], &pp[o], &rr[o], tt.data());
        for (int k = 0; k < nz1; ++k) {
            double pk = std::pow((ppb[k] + pp[o + k]) / p0, rgas / cp);
            double t = tt[k] / pk;  // theta_m (qv = 0)
            rho_zz[o + k] = rb[o + k] + rr[o + k];
            rho[o + k] = rho_zz[o + k] * zz[o + k];     // :1155
            theta[o + k] = t / (1.0 + 1.61 * 0.0);      // :1156
        }
    }
    // ---- u on edges (:971-1029)
    Field& f_fE = p.addr("fEdge", LOC_EDGE, {nEdges + 1});
    Field& f_fV = p.addr("fVertex", LOC_VERTEX, {nVertices + 1});
    const double lat_pert = latitude_pert * pii / 180., lon_pert = longitude_pert * pii / 180.;
#pragma omp parallel for schedule(dynamic, 256)
    for (int e = 0; e < nEdges; ++e) {
        int v1 = m.verticesOnEdge[2 * e] - 1, v2 = m.verticesOnEdge[2 * e + 1] - 1;
        int c1 = m.cellsOnEdge[2 * e] - 1, c2 = m.cellsOnEdge[2 * e + 1] - 1;
        f_fE.r[e] = 2.0 * omega_e * std::sin(m.latEdge[e]);
        if (v1 >= nVertices || v2 >= nVertices) continue;  // incomplete outer-halo edge
        double lat1 = m.latVertex[v1], lat2 = m.latVertex[v2];
        double u_pert = 0.0;
        if (cfg.init_case == 2) {
            double r_pert = sphere_distance(m.latEdge[e], m.lonEdge[e], lat_pert, lon_pert, 1.0) / pert_radius;
            u_pert = u_perturbation * std::exp(-p2(r_pert)) * (lat2 - lat1) * sphere_radius / m.dvEdge[e];
        } else if (cfg.init_case == 3) {
            u_pert = u_perturbation * std::cos(k_x * (m.lonEdge[e] - lon_pert)) * (0.5 * (lat2 - lat1) - 0.125 * (std::sin(4. * lat2) - std::sin(4. * lat1))) * sphere_radius / m.dvEdge[e];
        }
        std::vector<double> flux_zonal(nz1);
        calc_flux_zonal(u_2d, lat_2d, flux_zonal.data(), lat1, lat2, m.dvEdge[e], sphere_radius, u0, nz1, nlat);
        for (int k = 0; k < nz1; ++k) {
            double fluxk = u0 * flux_zonal[k] / (0.5 * (rb[(size_t)c1 * L + k] + rb[(size_t)c2 * L + k] + rr[(size_t)c1 * L + k] + rr[(size_t)c2 * L + k]));
            u[(size_t)e * L + k] = fluxk + u_pert;
            ru[(size_t)e * L + k] = 0.5 * (rho_zz[(size_t)c1 * L + k] + rho_zz[(size_t)c2 * L + k]) * u[(size_t)e * L + k];
        }
    }
    for (int v = 0; v < nVertices; ++v) f_fV.r[v] = 2.0 * omega_e * std::sin(m.latVertex[v]);

    // ---- zb, zb3 (:1045-1093) stored cell-wise as atm_compute_signs does (mpas_atm_core.F:1037-1053)
    Field& f_zbc = p.addr("zb_cell", LOC_CELL, {L + 1, mE, nCells + 1});
    Field& f_zb3c = p.addr("zb3_cell", LOC_CELL, {L + 1, mE, nCells + 1});
    double *zb_cell = f_zbc.r.data(), *zb3_cell = f_zb3c.r.data();
    double *zb = nullptr, *zb3 = nullptr;
    if (cfg.keep_zb) {
        zb = p.addr("zb", LOC_EDGE, {L + 1, 2, nEdges + 1}).r.data();
        zb3 = p.addr("zb3", LOC_EDGE, {L + 1, 2, nEdges + 1}).r.data();
    }
    const double* deriv_two = p.R("deriv_two");
#pragma omp parallel for schedule(dynamic, 256)
    for (int e = 0; e < nEdges; ++e) {
        int cell1 = m.cellsOnEdge[2 * e], cell2 = m.cellsOnEdge[2 * e + 1];
        if (cell1 > nCells || cell2 > nCells) continue;
        int slot1 = -1, slot2 = -1;
        for (int i = 0; i < m.nEdgesOnCell[cell1 - 1]; ++i)
            if (m.edgesOnCell[(size_t)(cell1 - 1) * mE + i] == e + 1) slot1 = i;
        for (int i = 0; i < m.nEdgesOnCell[cell2 - 1]; ++i)
            if (m.edgesOnCell[(size_t)(cell2 - 1) * mE + i] == e + 1) slot2 = i;
        for (int k = 0; k < nz1; ++k) {
            double z_edge, z_edge3 = 0.;
            double zg1 = zgrid[(size_t)(cell1 - 1) * nz + k], zg2 = zgrid[(size_t)(cell2 - 1) * nz + k];
            if (cfg.theta_adv_order == 2) {
                z_edge = (zg1 + zg2) / 2.;
            } else {
                double d1 = deriv_two[(size_t)e * 30 + 0] * zg1;
                double d2 = deriv_two[(size_t)e * 30 + 15] * zg2;
                for (int i = 1; i <= m.nEdgesOnCell[cell1 - 1]; ++i) {
                    int cc = m.cellsOnCell[(size_t)(cell1 - 1) * mE + (i - 1)];
                    if (cc > 0) d1 = d1 + deriv_two[(size_t)e * 30 + i] * zgrid[(size_t)(cc - 1) * nz + k];
                }
                for (int i = 1; i <= m.nEdgesOnCell[cell2 - 1]; ++i) {
                    int cc = m.cellsOnCell[(size_t)(cell2 - 1) * mE + (i - 1)];
                    if (cc > 0) d2 = d2 + deriv_two[(size_t)e * 30 + 15 + i] * zgrid[(size_t)(cc - 1) * nz + k];
                }
                z_edge = 0.5 * (zg1 + zg2) - p2(m.dcEdge[e]) * (d1 + d2) / 12.;
                if (cfg.theta_adv_order == 3) z_edge3 = -p2(m.dcEdge[e]) * (d1 - d2) / 12.;
            }
            double b1 = (z_edge - zg1) * m.dvEdge[e] / m.areaCell[cell1 - 1];
            double b2 = (z_edge - zg2) * m.dvEdge[e] / m.areaCell[cell2 - 1];
            double b31 = z_edge3 * m.dvEdge[e] / m.areaCell[cell1 - 1];
            double b32 = z_edge3 * m.dvEdge[e] / m.areaCell[cell2 - 1];
            if (zb) {
                zb[((size_t)e * 2 + 0) * nz + k] = b1; zb[((size_t)e * 2 + 1) * nz + k] = b2;
                zb3[((size_t)e * 2 + 0) * nz + k] = b31; zb3[((size_t)e * 2 + 1) * nz + k] = b32;
            }
            if (slot1 >= 0) { zb_cell[((size_t)(cell1 - 1) * mE + slot1) * nz + k] = b1; zb3_cell[((size_t)(cell1 - 1) * mE + slot1) * nz + k] = b31; }
            if (slot2 >= 0) { zb_cell[((size_t)(cell2 - 1) * mE + slot2) * nz + k] = b2; zb3_cell[((size_t)(cell2 - 1) * mE + slot2) * nz + k] = b32; }
        }
    }
    // ---- rw (terrain part) and w (:1096-1126), gather form
#pragma omp parallel for schedule(static)
    for (int c = 0; c < nCells; ++c) {
        std::vector<double> rw(nz, 0.0);
        for (int i = 0; i < m.nEdgesOnCell[c]; ++i) {
            int e = m.edgesOnCell[(size_t)c * mE + i] - 1;
            if (e >= nEdges) continue;
            bool is1 = (m.cellsOnEdge[2 * e] - 1 == c);
            for (int k = 1; k < nz1; ++k) {
                double flux = (fzm[k] * ru[(size_t)e * L + k] + fzp[k] * ru[(size_t)e * L + k - 1]);
                double zzt = (fzm[k] * zz[(size_t)c * L + k] + fzp[k] * zz[(size_t)c * L + k - 1]);
                double b = zb_cell[((size_t)c * mE + i) * nz + k], b3 = zb3_cell[((size_t)c * mE + i) * nz + k];
                double s = sgn1(ru[(size_t)e * L + k]);
                if (is1) {
                    rw[k] = rw[k] - zzt * b * flux;
                    if (cfg.theta_adv_order == 3) rw[k] = rw[k] + s * cfg.coef_3rd_order * zzt * b3 * flux;
                } else {
                    rw[k] = rw[k] + zzt * b * flux;
                    if (cfg.theta_adv_order == 3) rw[k] = rw[k] - s * cfg.coef_3rd_order * zzt * b3 * flux;
                }
            }
        }
        for (int k = 1; k < nz1; ++k) w[(size_t)c * nz + k] = rw[k] / (fzp[k] * rho_zz[(size_t)c * L + k - 1] + fzm[k] * rho_zz[(size_t)c * L + k]);
    }
    p.cfg["nVertLevels"] = L;
    p.cfg["num_scalars"] = cfg.num_scalars;
}

}  // namespace mpas
