// atm_compute_dyn_tend_work (TI:4481-5416), TI = src/core_atmosphere/dynamics/mpas_atm_time_integration.F.
//
// The reference routine is ~25 loop nests with OpenMP barriers between the ones that consume a
// neighbour's result.  Grouped by those dependences it becomes 4 kernels on every RK stage plus 3
// on rk_step 1 (forward-Euler mixing terms), each one thread per (k, element) with k fastest:
//   dt_cell_pre_kernel   all cells   [rk1: Smagorinsky kdiff TI:4677-4720] h_divergence TI:4729-4748
//                                    [rk1: tend_rho, dpdz TI:4755-4766]
//   dt_edge_mix_kernel   all edges   rk1: pressure gradient into tend_u_euler TI:4781-4788, del2 u +
//                                    delsq_u TI:4856-4883
//   dt_vertex_mix_kernel all verts   rk1: delsq_vorticity TI:4893-4904
//   dt_cell_mix_kernel   all cells   rk1: delsq_divergence TI:4908-4920, del2 w + delsq_w TI:5097-5128,
//                                    del2 theta + delsq_theta TI:5276-5300
//   dt_edge_flux_kernel  edges of    high-order horizontal fluxes of theta_m and w, once per edge (library workspace)
//                        owned cells
//   dt_edge_kernel       owned edges vertical advection of u TI:4792-4807, nonlinear Coriolis TI:4811-4820,
//                                    KE gradient TI:4823-4838, rk1: del4 u TI:4924-4944, vertical mixing
//                                    TI:4949-5006, Rayleigh damping TI:5015-5028, final sum TI:5030-5036
//   dt_cell_kernel       owned cells w: flux divergence TI:5046-5074, rk1 del4 TI:5134-5154, vertical advection,
//                                    pgf + buoyancy, final sum TI:5167-5223; theta_m: flux divergence TI:5231-5269,
//                                    rk1 del4 TI:5304-5325, vertical advection, diabatic, vertical mixing, final
//                                    sum TI:5331-5414
// The last three are warp-per-column kernels (lanes along k, per-column stencil metadata staged once in shared
// memory, vertical flux columns in shared memory); the others one thread per (k, element), k fastest.
// Every sum over edgesOnCell / edgesOnEdge / advCellsForEdge keeps the reference's index order and
// is a gather (no atomics), so results are independent of the launch geometry.
#include "common.cuh"

using namespace mpasb200;

namespace {

struct DtP {
    KMagic km;
    int L, mE, mE2, vD, rk_step, nCellsSolve;
    int cellStart, nCellsLoop, cellSolveStart, nCellsSolveLoop, edgeStart, nEdgesLoop, edgeSolveStart, edgeSolveEnd, nEdgesSolveLoop;
    int vertexStart, nVerticesLoop;
    int mixing, mix_full, rayleigh_damp_u, n_rayleigh_levels, pf;
    real c_s, len_disp, invDt, cam_coef, h_theta_eddy_visc2_fixed, h_mom_eddy_visc4, h_theta_eddy_visc4, del4u_div_factor;
    real v_mom_eddy_visc2, v_theta_eddy_visc2, coef_3rd_order, rayleigh_coef_inverse, prandtl_inv;
    const real *fEdge, *dvEdge, *dcEdge, *invDcEdge, *invDvEdge, *invAreaCell, *invAreaTriangle, *meshScalingDel2, *meshScalingDel4;
    const real *weightsOnEdge, *zgrid, *rho_edge, *rho_zz, *ru, *u, *v, *divergence, *vorticity, *ke, *pv_edge, *theta_m, *rw;
    const real *rt_diabatic_tend, *w, *cqw, *rb, *pp, *zz, *zxu, *cqu, *edgesOnCell_sign, *edgesOnVertex_sign, *rw_save, *ru_save;
    const real *theta_m_save, *rr_save, *angleEdge, *u_init, *v_init, *adv_coefs, *adv_coefs_3rd, *rdzu, *rdzw, *fzm, *fzp, *t_init;
    const real *defc_a, *defc_b, *qtot, *tend_ru_physics, *tend_rtheta_physics, *tend_rho_physics;
    const int *cellsOnEdge, *verticesOnEdge, *edgesOnCell, *edgesOnEdge, *edgesOnVertex, *nEdgesOnCell, *nEdgesOnEdge, *advCellsForEdge,
        *nAdvCellsForEdge;
    real *tend_u, *tend_rho, *tend_theta, *tend_w, *h_divergence, *kdiff, *tend_u_euler, *tend_w_euler, *tend_theta_euler;
    real *tend_rtheta_adv, *rthdynten, *delsq_theta, *delsq_w, *delsq_divergence, *delsq_u, *delsq_vorticity, *dpdz;
};

// Load-first flat kernels: the index loads of all stencil slots, then the data loads (in batches that fit the register
// budget), then the sums in the reference's order.  Slots beyond nEdgesOnCell repeat slot 1 and are left out of the sums.
template <int NE, bool SMAG, int MINB = (SMAG ? 3 : 4)>
__global__ void __launch_bounds__(256, MINB) dt_cell_pre_kernel(DtP p) {
    const int L = p.L, L1 = L + 1, mE = p.mE;
    FLAT_KC(p.nCellsLoop, L);
    const int iCell = p.cellStart + _c0;
    const size_t ic = IX2(k, iCell, L);
    const int ne = p.nEdgesOnCell[iCell - 1];
    const int* eoc = p.edgesOnCell + (size_t)(iCell - 1) * mE;
    const real* sg = p.edgesOnCell_sign + (size_t)(iCell - 1) * mE;
    int e[NE];
#pragma unroll
    for (int i = 0; i < NE; ++i) e[i] = eoc[i < ne ? i : 0];
    real hdiv = 0.0, d_diag = 0.0, d_off_diag = 0.0;
    {
        real dv[NE], rue[NE], sv[NE];
#pragma unroll
        for (int i = 0; i < NE; ++i) {
            dv[i] = p.dvEdge[e[i] - 1];
            rue[i] = p.ru[IX2(k, e[i], L)];
            sv[i] = sg[i < ne ? i : 0];
        }
#pragma unroll
        for (int i = 0; i < NE; ++i) {
            const real edge_sign = sv[i] * dv[i];
            if (i < ne) hdiv = hdiv + edge_sign * rue[i];
        }
    }
    if (SMAG) {  // TI:4677-4693
        const real* da = p.defc_a + (size_t)(iCell - 1) * mE;
        const real* db = p.defc_b + (size_t)(iCell - 1) * mE;
        real a[NE], b[NE], ue[NE], ve[NE];
#pragma unroll
        for (int i = 0; i < NE; ++i) {
            const int j = i < ne ? i : 0;
            a[i] = da[j];
            b[i] = db[j];
            ue[i] = p.u[IX2(k, e[i], L)];
            ve[i] = p.v[IX2(k, e[i], L)];
        }
#pragma unroll
        for (int i = 0; i < NE; ++i) {
            if (i < ne) {
                d_diag = d_diag + a[i] * ue[i] - b[i] * ve[i];
                d_off_diag = d_off_diag + b[i] * ue[i] + a[i] * ve[i];
            }
        }
    }
    hdiv = hdiv * p.invAreaCell[iCell - 1];
    p.h_divergence[ic] = hdiv;
    if (p.rk_step == 1) {
        if (p.mixing == 1 || p.mixing == 2) {
            real kd;
            if (p.mixing == 1) {
                const real cl = (p.c_s * p.len_disp);
                kd = fmin((cl * cl) * sqrt(d_diag * d_diag + d_off_diag * d_off_diag), ((real)0.01 * (p.len_disp * p.len_disp)) * p.invDt);
            } else {
                kd = p.h_theta_eddy_visc2_fixed;
            }
            if (p.cam_coef > (real)0.0) {  // TI:4708-4720
                if (k == L - 2) kd = fmax(kd, (real)2.0833 * p.len_disp * p.cam_coef);
                if (k == L - 1) kd = fmax(kd, (real)2.0 * (real)2.0833 * p.len_disp * p.cam_coef);
                if (k == L) kd = fmax(kd, (real)4.0 * (real)2.0833 * p.len_disp * p.cam_coef);
            }
            p.kdiff[ic] = kd;
        } else if (p.cam_coef > (real)0.0) {
            real kd = p.kdiff[ic];
            if (k == L - 2) kd = fmax(kd, (real)2.0833 * p.len_disp * p.cam_coef);
            if (k == L - 1) kd = fmax(kd, (real)2.0 * (real)2.0833 * p.len_disp * p.cam_coef);
            if (k == L) kd = fmax(kd, (real)4.0 * (real)2.0833 * p.len_disp * p.cam_coef);
            p.kdiff[ic] = kd;
        }
        p.tend_rho[ic] = -hdiv - p.rdzw[k - 1] * (p.rw[IX2(k + 1, iCell, L1)] - p.rw[IX2(k, iCell, L1)]) +
                         (p.tend_rho_physics ? p.tend_rho_physics[ic] : (real)0.0);
        const real qt = p.qtot[ic];
        p.dpdz[ic] = -MPAS_GRAVITY * (p.rb[ic] * (qt) + p.rr_save[ic] * ((real)1. + qt));
    }
}

template <int MINB>
__global__ void __launch_bounds__(256, MINB) dt_edge_mix_kernel(DtP p) {  // rk_step == 1 only
    const int L = p.L;
    FLAT_KC(p.nEdgesLoop, L);
    const int iEdge = p.edgeStart + _c0;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    const int vertex1 = p.verticesOnEdge[IX2(1, iEdge, 2)], vertex2 = p.verticesOnEdge[IX2(2, iEdge, 2)];
    const size_t ie = IX2(k, iEdge, L), i1 = IX2(k, cell1, L), i2 = IX2(k, cell2, L);
    const real idc = p.invDcEdge[iEdge - 1];
    real te;
    if (iEdge >= p.edgeSolveStart && iEdge <= p.edgeSolveEnd) {
        te = -p.cqu[ie] * ((p.pp[i2] - p.pp[i1]) * idc / ((real).5 * (p.zz[i2] + p.zz[i1])) - (real)0.5 * p.zxu[ie] * (p.dpdz[i1] + p.dpdz[i2]));
    } else {
        te = p.tend_u_euler[ie];
    }
    const real r_dc = idc;
    const real r_dv = fmin(p.invDvEdge[iEdge - 1], 4 * idc);
    const real u_diffusion = (p.divergence[i2] - p.divergence[i1]) * r_dc - (p.vorticity[IX2(k, vertex2, L)] - p.vorticity[IX2(k, vertex1, L)]) * r_dv;
    p.delsq_u[ie] = (real)0.0 + u_diffusion;
    const real kdiffu = (real)0.5 * (p.kdiff[i1] + p.kdiff[i2]);
    te = te + p.rho_edge[ie] * kdiffu * u_diffusion * p.meshScalingDel2[iEdge - 1];
    p.tend_u_euler[ie] = te;
}

__global__ void __launch_bounds__(256) dt_vertex_mix_kernel(DtP p) {  // rk_step == 1, h_mom_eddy_visc4 > 0
    const int L = p.L, vD = p.vD;
    FLAT_KC(p.nVerticesLoop, L);
    const int iVertex = p.vertexStart + _c0;
    real s = 0.0;
    const real iat = p.invAreaTriangle[iVertex - 1];
    for (int i = 1; i <= vD; ++i) {
        const int iEdge = p.edgesOnVertex[IX2(i, iVertex, vD)];
        const real edge_sign = iat * p.dcEdge[iEdge - 1] * p.edgesOnVertex_sign[IX2(i, iVertex, vD)];
        s = s + edge_sign * p.delsq_u[IX2(k, iEdge, L)];
    }
    p.delsq_vorticity[IX2(k, iVertex, L)] = s;
}

template <int MINB>
__global__ void __launch_bounds__(256, MINB) dt_cell_mix_kernel(DtP p) {  // rk_step == 1
    const int L = p.L, L1 = L + 1, mE = p.mE;
    FLAT_KC(p.nCellsLoop, L1);
    const int iCell = p.cellStart + _c0;
    const int ne = p.nEdgesOnCell[iCell - 1];
    const real r_areaCell = p.invAreaCell[iCell - 1];
    if (k == L1) {
        p.tend_w_euler[IX2(k, iCell, L1)] = 0.0;
        return;
    }
    real dsq_div = 0.0, dsq_w = 0.0, twe = 0.0, dsq_th = 0.0, tte = 0.0;
    for (int i = 1; i <= ne; ++i) {
        const int iEdge = p.edgesOnCell[IX2(i, iCell, mE)];
        const real sgn = p.edgesOnCell_sign[IX2(i, iCell, mE)];
        const real dv = p.dvEdge[iEdge - 1], idc = p.invDcEdge[iEdge - 1];
        const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
        const size_t ie = IX2(k, iEdge, L), i1 = IX2(k, cell1, L), i2 = IX2(k, cell2, L);
        const real ms2 = p.meshScalingDel2[iEdge - 1];
        const real rho_e = p.rho_edge[ie];
        const real kd1 = p.kdiff[i1], kd2 = p.kdiff[i2];
        if (p.h_mom_eddy_visc4 > (real)0.0) {
            const real edge_sign = r_areaCell * dv * sgn;
            dsq_div = dsq_div + edge_sign * p.delsq_u[ie];
        }
        if (k >= 2) {
            const real edge_sign = (real)0.5 * r_areaCell * sgn * dv * idc;
            real w_turb_flux = edge_sign * (rho_e + p.rho_edge[IX2(k - 1, iEdge, L)]) * (p.w[IX2(k, cell2, L1)] - p.w[IX2(k, cell1, L1)]);
            dsq_w = dsq_w + w_turb_flux;
            w_turb_flux = w_turb_flux * ms2 * (real)0.25 * (kd1 + kd2 + p.kdiff[IX2(k - 1, cell1, L)] + p.kdiff[IX2(k - 1, cell2, L)]);
            twe = twe + w_turb_flux;
        }
        {
            const real edge_sign = r_areaCell * sgn * dv * idc;
            const real pr_scale = p.prandtl_inv * ms2;
            real theta_turb_flux = edge_sign * (p.theta_m[i2] - p.theta_m[i1]) * rho_e;
            dsq_th = dsq_th + theta_turb_flux;
            theta_turb_flux = theta_turb_flux * (real)0.5 * (kd1 + kd2) * pr_scale;
            tte = tte + theta_turb_flux;
        }
    }
    const size_t ic = IX2(k, iCell, L);
    if (p.h_mom_eddy_visc4 > (real)0.0) p.delsq_divergence[ic] = dsq_div;
    p.delsq_w[ic] = dsq_w;
    p.tend_w_euler[IX2(k, iCell, L1)] = twe;
    p.delsq_theta[ic] = dsq_th;
    p.tend_theta_euler[ic] = tte;
}

// dt_cell_mix_kernel as one cell per warp, two levels per lane: the edge metadata (offsets of the edge and of its two cells,
// the sign * metric factors in the reference's multiplication order, meshScalingDel2) is staged once per cell instead of
// once per level; cells with more than 6 edges and meshes with maxEdges > 8 take the flat kernel's path per slot count.
// Padding slots carry zero factors (sum + 0 * x == sum: the sums start at +0 and are never -0).
template <int MINB>
__global__ void __launch_bounds__(WPB * 32, MINB) dt_cell_mix_warp_kernel(DtP p) {  // rk_step == 1
    __shared__ unsigned s_eo[WPB][8], s_c1[WPB][8], s_c2[WPB][8], s_w1[WPB][8], s_w2[WPB][8];
    __shared__ real s_fd[WPB][8], s_fw[WPB][8], s_ft[WPB][8], s_ms2[WPB][8];
    const int L = p.L, L1 = L + 1, mE = p.mE;
    WARP_ENTITY(p.cellStart, p.nCellsLoop);
    const int iCell = ent;
    const int ne = p.nEdgesOnCell[iCell - 1];
    const real r_areaCell = p.invAreaCell[iCell - 1];
    const int npad = ne <= 6 ? 6 : 8;
    if (lane < 8) {
        const bool in = lane < ne;
        const int iEdge = p.edgesOnCell[IX2(in ? lane + 1 : 1, iCell, mE)];
        const real sgn = p.edgesOnCell_sign[IX2(in ? lane + 1 : 1, iCell, mE)];
        const real dv = p.dvEdge[iEdge - 1], idc = p.invDcEdge[iEdge - 1];
        const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
        s_eo[warp][lane] = (unsigned)(iEdge - 1) * (unsigned)L;
        s_c1[warp][lane] = (unsigned)(cell1 - 1) * (unsigned)L;
        s_c2[warp][lane] = (unsigned)(cell2 - 1) * (unsigned)L;
        s_w1[warp][lane] = (unsigned)(cell1 - 1) * (unsigned)L1;
        s_w2[warp][lane] = (unsigned)(cell2 - 1) * (unsigned)L1;
        s_fd[warp][lane] = in ? r_areaCell * dv * sgn : (real)0.0;
        s_fw[warp][lane] = in ? (real)0.5 * r_areaCell * sgn * dv * idc : (real)0.0;
        s_ft[warp][lane] = in ? r_areaCell * sgn * dv * idc : (real)0.0;
        s_ms2[warp][lane] = p.meshScalingDel2[iEdge - 1];
    }
    __syncwarp();
    const bool visc4 = p.h_mom_eddy_visc4 > (real)0.0;
    const size_t coff = (size_t)(iCell - 1) * L, woff = (size_t)(iCell - 1) * L1;
    if (lane == 0) p.tend_w_euler[woff + L] = 0.0;  // k = nVertLevels + 1
    for (int ka = lane; ka < L; ka += 64) {  // 0-based levels ka and ka + 32
        const bool vb = ka + 32 < L;
        const int bo = vb ? 32 : 0;
        const int ma = ka > 0 ? 1 : 0;  // distance to the level below (level 1 has none: the w terms are skipped there)
        real dsq_div[2] = {0, 0}, dsq_w[2] = {0, 0}, twe[2] = {0, 0}, dsq_th[2] = {0, 0}, tte[2] = {0, 0};
#pragma unroll 1
        for (int t = 0; t < npad; ++t) {
            const unsigned eo = s_eo[warp][t], c1 = s_c1[warp][t], c2 = s_c2[warp][t];
            const real* __restrict__ re = elem_ptr(p.rho_edge + ka, eo);
            const real* __restrict__ k1 = elem_ptr(p.kdiff + ka, c1);
            const real* __restrict__ k2 = elem_ptr(p.kdiff + ka, c2);
            const real* __restrict__ t1 = elem_ptr(p.theta_m + ka, c1);
            const real* __restrict__ t2 = elem_ptr(p.theta_m + ka, c2);
            const real* __restrict__ w1 = elem_ptr(p.w + ka, s_w1[warp][t]);
            const real* __restrict__ w2 = elem_ptr(p.w + ka, s_w2[warp][t]);
            real rho_e[2], rho_m[2], kd1[2], kd2[2], kd1m[2], kd2m[2], th1[2], th2[2], ww1[2], ww2[2], du[2] = {0, 0};
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int o = h ? bo : 0, m = h ? (bo ? 1 : ma) : ma;
                rho_e[h] = re[o]; rho_m[h] = re[o - m];
                kd1[h] = k1[o]; kd2[h] = k2[o]; kd1m[h] = k1[o - m]; kd2m[h] = k2[o - m];
                th1[h] = t1[o]; th2[h] = t2[o];
                ww1[h] = w1[o]; ww2[h] = w2[o];
            }
            if (visc4) {
                const real* __restrict__ dq = elem_ptr(p.delsq_u + ka, eo);
                du[0] = dq[0];
                du[1] = dq[bo];
            }
            const real fd = s_fd[warp][t], fw = s_fw[warp][t], ft = s_ft[warp][t], ms2 = s_ms2[warp][t];
            const real pr_scale = p.prandtl_inv * ms2;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (visc4) dsq_div[h] = dsq_div[h] + fd * du[h];
                if (h == 1 || ka > 0) {  // k >= 2
                    real w_turb_flux = fw * (rho_e[h] + rho_m[h]) * (ww2[h] - ww1[h]);
                    dsq_w[h] = dsq_w[h] + w_turb_flux;
                    w_turb_flux = w_turb_flux * ms2 * (real)0.25 * (kd1[h] + kd2[h] + kd1m[h] + kd2m[h]);
                    twe[h] = twe[h] + w_turb_flux;
                }
                real theta_turb_flux = ft * (th2[h] - th1[h]) * rho_e[h];
                dsq_th[h] = dsq_th[h] + theta_turb_flux;
                theta_turb_flux = theta_turb_flux * (real)0.5 * (kd1[h] + kd2[h]) * pr_scale;
                tte[h] = tte[h] + theta_turb_flux;
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !vb) break;
            const int k0 = ka + 32 * h;
            if (visc4) p.delsq_divergence[coff + k0] = dsq_div[h];
            p.delsq_w[coff + k0] = dsq_w[h];
            p.tend_w_euler[woff + k0] = twe[h];
            p.delsq_theta[coff + k0] = dsq_th[h];
            p.tend_theta_euler[coff + k0] = tte[h];
        }
    }
}

// ---- warp-per-column kernels: one mesh element per warp, lanes along k.  The k-independent stencil metadata
// (neighbour indices, weights, signs) is fetched ONCE per warp into shared memory instead of once per level. ----

// High-order horizontal fluxes of theta_m and w through every edge of an owned cell, evaluated once per edge
// (the reference evaluates each edge's stencil from both adjacent cells, TI:5046-5074, 5231-5252):
//   ft (k,e) = ru(k,e)        * sum_j (adv_coefs(j,e) +- adv_coefs_3rd(j,e)) * theta_m(k, advCell_j)     k = 1..L
//   fw (k,e) = ru_edge_w(k,e) * sum_j (adv_coefs(j,e) +- adv_coefs_3rd(j,e)) * w(k, advCell_j)           k = 2..L
//   ft2(k,e) = dvEdge * (ru_save - ru) * 0.5 * (theta_m_save(c2) + theta_m_save(c1))   (rk_step > 1, TI:5256-5269)
// +- is Fortran sign(1, flux); a + (+-1)*b == a +- b bit for bit, and multiplying by edgesOnCell_sign = +-1 later is
// exact, so the cell kernels reproduce the reference's sums exactly.
// One edge per warp; each lane owns levels (lane, lane + 32) of a 64-level chunk so that one address computation
// serves two gathers.  The stencil is staged once per warp as element offsets plus the two coefficient tables
// (c + c3, c - c3), padded with zero coefficients to a multiple of FUNR (flux + 0*x == flux exactly, flux is never -0);
// each lane selects its table by the sign of its own ru once per level.
#define FUNR 5
template <bool RK1, int MINB>
__global__ void __launch_bounds__(WPB * 32, MINB) dt_edge_flux_kernel(DtP p, real* __restrict__ ft, real* __restrict__ fw, real* __restrict__ ft2) {
    __shared__ unsigned s_ot[WPB][20], s_ow[WPB][20];
    __shared__ real s_cp[WPB][20], s_cm[WPB][20];
    const int L = p.L, L1 = L + 1;
    WARP_ENTITY(p.edgeStart, p.nEdgesLoop);
    const int iEdge = ent;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    if (!(cell1 <= p.nCellsSolve || cell2 <= p.nCellsSolve)) return;
    const int n = p.nAdvCellsForEdge[iEdge - 1];
    const int npad = (n + FUNR - 1) / FUNR * FUNR;
    if (lane < npad) {
        const bool in = lane < n;
        const int c = in ? p.advCellsForEdge[IX2(lane + 1, iEdge, 15)] : cell1;
        const real a = in ? p.adv_coefs[IX2(lane + 1, iEdge, 15)] : (real)0.0, a3 = in ? p.adv_coefs_3rd[IX2(lane + 1, iEdge, 15)] : (real)0.0;
        s_ot[warp][lane] = (unsigned)(c - 1) * (unsigned)L;
        s_ow[warp][lane] = (unsigned)(c - 1) * (unsigned)L1;
        s_cp[warp][lane] = a + a3;
        s_cm[warp][lane] = a - a3;
    }
    __syncwarp();
    const real dv = p.dvEdge[iEdge - 1];
    const unsigned eoff = (unsigned)(iEdge - 1) * (unsigned)L;
    const unsigned c1off = (unsigned)(cell1 - 1) * (unsigned)L, c2off = (unsigned)(cell2 - 1) * (unsigned)L;
    for (int ka = lane; ka < L; ka += 64) {  // 0-based levels ka and ka + 32
        const bool vb = ka + 32 < L;
        const int bo = vb ? 32 : 0;  // lanes without a second level recompute the first one
        const real* __restrict__ rup = elem_ptr(p.ru + ka, eoff);
        const real rue_a = rup[0], rue_b = rup[bo];
        const int mo = ka > 0 ? -1 : 0;  // level below (k = 1 has no w flux; keeps the load in bounds)
        const real rum_a = rup[mo], rum_b = rup[bo - 1 + (vb ? 0 : 1 + mo)];
        real rs_a = 0, rs_b = 0, t1_a = 0, t1_b = 0, t2_a = 0, t2_b = 0;
        if (!RK1) {  // operands of ft2, loaded ahead of the two gathers
            const real* __restrict__ rs = elem_ptr(p.ru_save + ka, eoff);
            const real* __restrict__ t1 = elem_ptr(p.theta_m_save + ka, c1off);
            const real* __restrict__ t2 = elem_ptr(p.theta_m_save + ka, c2off);
            rs_a = rs[0]; rs_b = rs[bo];
            t1_a = t1[0]; t1_b = t1[bo];
            t2_a = t2[0]; t2_b = t2[bo];
        }
        {
            const real* tab_a = signbit(rue_a) ? s_cm[warp] : s_cp[warp];
            const real* tab_b = signbit(rue_b) ? s_cm[warp] : s_cp[warp];
            const real* __restrict__ th = p.theta_m + ka;
            real f_a = 0.0, f_b = 0.0;
            for (int j = 0; j < npad; j += FUNR) {
                real xa[FUNR], xb[FUNR];
#pragma unroll
                for (int t = 0; t < FUNR; ++t) {
                    const real* __restrict__ q = elem_ptr(th, s_ot[warp][j + t]);
                    xa[t] = q[0];
                    xb[t] = q[bo];
                }
#pragma unroll
                for (int t = 0; t < FUNR; ++t) {
                    f_a = f_a + tab_a[j + t] * xa[t];
                    f_b = f_b + tab_b[j + t] * xb[t];
                }
            }
            real* o = ft + (size_t)eoff + ka;
            o[0] = rue_a * f_a;
            if (vb) o[32] = rue_b * f_b;
        }
        {
            const real ruw_a = p.fzm[ka] * rue_a + p.fzp[ka] * rum_a;
            const real ruw_b = p.fzm[ka + bo] * rue_b + p.fzp[ka + bo] * rum_b;
            const real* tab_a = signbit(ruw_a) ? s_cm[warp] : s_cp[warp];
            const real* tab_b = signbit(ruw_b) ? s_cm[warp] : s_cp[warp];
            const real* __restrict__ wp = p.w + ka;
            real f_a = 0.0, f_b = 0.0;
            for (int j = 0; j < npad; j += FUNR) {
                real xa[FUNR], xb[FUNR];
#pragma unroll
                for (int t = 0; t < FUNR; ++t) {
                    const real* __restrict__ q = elem_ptr(wp, s_ow[warp][j + t]);
                    xa[t] = q[0];
                    xb[t] = q[bo];
                }
#pragma unroll
                for (int t = 0; t < FUNR; ++t) {
                    f_a = f_a + tab_a[j + t] * xa[t];
                    f_b = f_b + tab_b[j + t] * xb[t];
                }
            }
            real* o = fw + (size_t)eoff + ka;
            if (ka > 0) o[0] = ruw_a * f_a;
            if (vb) o[32] = ruw_b * f_b;
        }
        if (!RK1) {
            real* o = ft2 + (size_t)eoff + ka;
            o[0] = dv * (rs_a - rue_a) * (real)0.5 * (t2_a + t1_a);
            if (vb) o[32] = dv * (rs_b - rue_b) * (real)0.5 * (t2_b + t1_b);
        }
    }
}

// The same fluxes with the stencil columns staged in shared memory (tiles.cu): one block per tile of consecutive edges.
// Every thread below nc issues the bulk (TMA) copies of ONE cell's theta_m and w columns (16-byte window around the
// column, as everywhere) and posts their byte count on the block's mbarrier; the block waits once, then each warp takes
// every WPB-th edge of the tile and evaluates exactly the sums of dt_edge_flux_kernel with shared-memory reads in place
// of the 2 x 10 gathered global loads per level pair.  About 1.1 columns per edge and array cross the L1/L2 boundary
// instead of 10, coalesced and without a register or scoreboard.  Identical arithmetic, identical results.
// dynamic shared memory: 2 * ncmax * tma_slot(L + 1) reals.
#define FT_NCMAX 112
#define FT_TEMAX 64
#define FT1_NCMAX 88  // one-field form: 88 columns + the per-edge tables = 53 KB per block, 4 blocks per SM
template <bool RK1>
__global__ void __launch_bounds__(WPB * 32, 2) dt_edge_flux_tile_kernel(DtP p, EdgeTilePlan tp, real* __restrict__ ft, real* __restrict__ fw, real* __restrict__ ft2) {
    extern __shared__ __align__(16) real s_dyn[];
    __shared__ __align__(8) unsigned long long s_bar;
    __shared__ unsigned short s_base_t[FT_NCMAX], s_base_w[FT_NCMAX];  // element offset of level 1 of each staged column
    __shared__ unsigned short s_ot[WPB][20], s_ow[WPB][20];
    __shared__ real s_cp[WPB][20], s_cm[WPB][20];
    const int L = p.L, L1 = L + 1, LS = tma_slot(L1);
    const int tile = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int e0 = tp.first[tile], e1 = tp.first[tile + 1], nc = tp.nc[tile];
    const unsigned mbar = (unsigned)__cvta_generic_to_shared(&s_bar);
    if (tid == 0) mbar_init(mbar, (unsigned)nc);
    __syncthreads();
    if (tid < nc) {
        const int c = tp.cells[(size_t)tile * tp.ncmax + tid];
        const real* gt = p.theta_m + (size_t)(c - 1) * L;
        const real* gw = p.w + (size_t)(c - 1) * L1;
        const unsigned s0 = (unsigned)__cvta_generic_to_shared(s_dyn), cs = (unsigned)(LS * sizeof(real));
        mbar_arrive_expect_tx(mbar, tma_col_bytes(gt, L) + tma_col_bytes(gw, L1));
        unsigned dummy = 0;
        tma_issue_col(s0 + (unsigned)tid * cs, gt, L, mbar, dummy);
        tma_issue_col(s0 + (unsigned)(tp.ncmax + tid) * cs, gw, L1, mbar, dummy);
        s_base_t[tid] = (unsigned short)(tid * LS + tma_shift(gt));
        s_base_w[tid] = (unsigned short)((tp.ncmax + tid) * LS + tma_shift(gw));
    }
    __syncthreads();
    mbar_wait(mbar, 0);
    for (int er = e0 + warp; er < e1; er += WPB) {
        const int iEdge = p.edgeStart + er;
        const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
        if (!(cell1 <= p.nCellsSolve || cell2 <= p.nCellsSolve)) continue;
        const int n = p.nAdvCellsForEdge[iEdge - 1];
        const int npad = (n + FUNR - 1) / FUNR * FUNR;
        __syncwarp();
        if (lane < npad) {
            const bool in = lane < n;
            const int li = in ? (int)tp.lidx[(size_t)er * 16 + lane] : 0;
            const real a = in ? p.adv_coefs[IX2(lane + 1, iEdge, 15)] : (real)0.0, a3 = in ? p.adv_coefs_3rd[IX2(lane + 1, iEdge, 15)] : (real)0.0;
            s_ot[warp][lane] = s_base_t[li];
            s_ow[warp][lane] = s_base_w[li];
            s_cp[warp][lane] = a + a3;
            s_cm[warp][lane] = a - a3;
        }
        __syncwarp();
        const real dv = p.dvEdge[iEdge - 1];
        const unsigned eoff = (unsigned)(iEdge - 1) * (unsigned)L;
        const unsigned c1off = (unsigned)(cell1 - 1) * (unsigned)L, c2off = (unsigned)(cell2 - 1) * (unsigned)L;
        for (int ka = lane; ka < L; ka += 64) {  // 0-based levels ka and ka + 32
            const bool vb = ka + 32 < L;
            const int bo = vb ? 32 : 0;
            const real* __restrict__ rup = elem_ptr(p.ru + ka, eoff);
            const real rue_a = rup[0], rue_b = rup[bo];
            const int mo = ka > 0 ? -1 : 0;
            const real rum_a = rup[mo], rum_b = rup[bo - 1 + (vb ? 0 : 1 + mo)];
            real rs_a = 0, rs_b = 0, t1_a = 0, t1_b = 0, t2_a = 0, t2_b = 0;
            if (!RK1) {
                const real* __restrict__ rs = elem_ptr(p.ru_save + ka, eoff);
                const real* __restrict__ t1 = elem_ptr(p.theta_m_save + ka, c1off);
                const real* __restrict__ t2 = elem_ptr(p.theta_m_save + ka, c2off);
                rs_a = rs[0]; rs_b = rs[bo];
                t1_a = t1[0]; t1_b = t1[bo];
                t2_a = t2[0]; t2_b = t2[bo];
            }
            {
                const real* tab_a = signbit(rue_a) ? s_cm[warp] : s_cp[warp];
                const real* tab_b = signbit(rue_b) ? s_cm[warp] : s_cp[warp];
                const real* th = s_dyn + ka;
                real f_a = 0.0, f_b = 0.0;
                for (int j = 0; j < npad; j += FUNR) {
                    real xa[FUNR], xb[FUNR];
#pragma unroll
                    for (int t = 0; t < FUNR; ++t) {
                        const real* q = th + s_ot[warp][j + t];
                        xa[t] = q[0];
                        xb[t] = q[bo];
                    }
#pragma unroll
                    for (int t = 0; t < FUNR; ++t) {
                        f_a = f_a + tab_a[j + t] * xa[t];
                        f_b = f_b + tab_b[j + t] * xb[t];
                    }
                }
                real* o = ft + (size_t)eoff + ka;
                o[0] = rue_a * f_a;
                if (vb) o[32] = rue_b * f_b;
            }
            {
                const real ruw_a = p.fzm[ka] * rue_a + p.fzp[ka] * rum_a;
                const real ruw_b = p.fzm[ka + bo] * rue_b + p.fzp[ka + bo] * rum_b;
                const real* tab_a = signbit(ruw_a) ? s_cm[warp] : s_cp[warp];
                const real* tab_b = signbit(ruw_b) ? s_cm[warp] : s_cp[warp];
                const real* wp = s_dyn + ka;
                real f_a = 0.0, f_b = 0.0;
                for (int j = 0; j < npad; j += FUNR) {
                    real xa[FUNR], xb[FUNR];
#pragma unroll
                    for (int t = 0; t < FUNR; ++t) {
                        const real* q = wp + s_ow[warp][j + t];
                        xa[t] = q[0];
                        xb[t] = q[bo];
                    }
#pragma unroll
                    for (int t = 0; t < FUNR; ++t) {
                        f_a = f_a + tab_a[j + t] * xa[t];
                        f_b = f_b + tab_b[j + t] * xb[t];
                    }
                }
                real* o = fw + (size_t)eoff + ka;
                if (ka > 0) o[0] = ruw_a * f_a;
                if (vb) o[32] = ruw_b * f_b;
            }
            if (!RK1) {
                real* o = ft2 + (size_t)eoff + ka;
                o[0] = dv * (rs_a - rue_a) * (real)0.5 * (t2_a + t1_a);
                if (vb) o[32] = dv * (rs_b - rue_b) * (real)0.5 * (t2_b + t1_b);
            }
        }
    }
}

// Second form of the staged kernel: ONE field per launch (FIELD 0: theta_m -> ft [, ft2]; FIELD 1: w -> fw), so a block holds
// one set of columns (half the shared memory: 4 blocks per SM), and everything an edge needs besides its own operand columns
// -- the shared-memory offsets of its stencil columns and the two coefficient tables -- is put into shared memory by the whole
// block BEFORE the wait for the bulk copies, so the latency of those index / coefficient loads hides behind the copy.  In
// the edge loop a warp then has only its operand loads (ru [, ru_save, theta_m_save of the two cells]) in front of the sums,
// and it requests those of its next edge before it works on the current one.
template <int FIELD, bool RK1>
__global__ void __launch_bounds__(WPB * 32, 4) dt_edge_flux_tile1_kernel(DtP p, EdgeTilePlan tp, real* __restrict__ ft, real* __restrict__ fw, real* __restrict__ ft2) {
    extern __shared__ __align__(16) real s_dyn[];
    __shared__ __align__(8) unsigned long long s_bar;
    __shared__ unsigned short s_base[FT1_NCMAX];
    __shared__ unsigned short s_o[FT_TEMAX][10];
    __shared__ real s_cp[FT_TEMAX][10], s_cm[FT_TEMAX][10];
    __shared__ int s_c1[FT_TEMAX], s_c2[FT_TEMAX];  // cells of the edge; s_c1 = 0: edge not needed (no owned cell) or not a 10-cell stencil
    const int L = p.L, L1 = L + 1, NL = FIELD ? L1 : L, LS = tma_slot(NL);
    const int tile = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int e0 = tp.first[tile], ne = tp.first[tile + 1] - e0, nc = tp.nc[tile];
    const unsigned mbar = (unsigned)__cvta_generic_to_shared(&s_bar);
    if (tid == 0) mbar_init(mbar, (unsigned)nc);
    __syncthreads();
    if (tid < nc) {
        const int c = tp.cells[(size_t)tile * tp.ncmax + tid];
        const real* g = FIELD ? p.w + (size_t)(c - 1) * L1 : p.theta_m + (size_t)(c - 1) * L;
        mbar_arrive_expect_tx(mbar, tma_col_bytes(g, NL));
        unsigned dummy = 0;
        tma_issue_col((unsigned)__cvta_generic_to_shared(s_dyn) + (unsigned)tid * (unsigned)(LS * sizeof(real)), g, NL, mbar, dummy);
        s_base[tid] = (unsigned short)(tid * LS + tma_shift(g));
    }
    if (tid < ne) {
        const int iEdge = p.edgeStart + e0 + tid;
        const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
        const bool need = (cell1 <= p.nCellsSolve || cell2 <= p.nCellsSolve);
        s_c1[tid] = need ? cell1 : 0;
        s_c2[tid] = cell2;
    }
    __syncthreads();
    for (int t = tid; t < ne * 10; t += WPB * 32) {
        const int er = t / 10, j = t - er * 10;
        const int iEdge = p.edgeStart + e0 + er;
        const int n = p.nAdvCellsForEdge[iEdge - 1];
        const bool in = j < n;
        const real a = in ? p.adv_coefs[IX2(j + 1, iEdge, 15)] : (real)0.0, a3 = in ? p.adv_coefs_3rd[IX2(j + 1, iEdge, 15)] : (real)0.0;
        s_o[er][j] = s_base[in ? (int)tp.lidx[(size_t)(e0 + er) * 16 + j] : 0];
        s_cp[er][j] = a + a3;
        s_cm[er][j] = a - a3;
        if (j == 0 && n > 10) s_c1[er] = -s_c1[er];  // longer stencil than the tables hold: the warp takes the direct path below
    }
    __syncthreads();
    mbar_wait(mbar, 0);
    // operands of the first edge of this warp
    real rue_a = 0, rue_b = 0, rum_a = 0, rum_b = 0, rs_a = 0, rs_b = 0, t1_a = 0, t1_b = 0, t2_a = 0, t2_b = 0;
    const int ka = lane;  // levels ka and ka + 32 (chunks beyond 64 levels: outer loop below)
    auto load_ops = [&](int er, int kb) {
        if (er >= ne) return;
        const int c1 = s_c1[er];
        if (c1 <= 0) return;
        const int iEdge = p.edgeStart + e0 + er;
        const unsigned eoff = (unsigned)(iEdge - 1) * (unsigned)L;
        const bool vb = kb + 32 < L;
        const int bo = vb ? 32 : 0;
        if (kb < L) {
            const real* __restrict__ rup = elem_ptr(p.ru + kb, eoff);
            rue_a = rup[0]; rue_b = rup[bo];
            if (FIELD) {
                const int mo = kb > 0 ? -1 : 0;
                rum_a = rup[mo]; rum_b = rup[bo - 1 + (vb ? 0 : 1 + mo)];
            } else if (!RK1) {
                const real* __restrict__ rs = elem_ptr(p.ru_save + kb, eoff);
                const real* __restrict__ t1 = elem_ptr(p.theta_m_save + kb, (unsigned)(c1 - 1) * (unsigned)L);
                const real* __restrict__ t2 = elem_ptr(p.theta_m_save + kb, (unsigned)(s_c2[er] - 1) * (unsigned)L);
                rs_a = rs[0]; rs_b = rs[bo];
                t1_a = t1[0]; t1_b = t1[bo];
                t2_a = t2[0]; t2_b = t2[bo];
            }
        }
    };
    for (int k0 = 0; k0 < L; k0 += 64) {
        const int kb = k0 + ka;
        load_ops(warp, kb);
        for (int er = warp; er < ne; er += WPB) {
            const real c_rue_a = rue_a, c_rue_b = rue_b, c_rum_a = rum_a, c_rum_b = rum_b;
            const real c_rs_a = rs_a, c_rs_b = rs_b, c_t1_a = t1_a, c_t1_b = t1_b, c_t2_a = t2_a, c_t2_b = t2_b;
            load_ops(er + WPB, kb);  // next edge of this warp: in flight while this one is summed
            const int c1 = s_c1[er];
            if (c1 == 0 || kb >= L) continue;
            const int iEdge = p.edgeStart + e0 + er;
            const unsigned eoff = (unsigned)(iEdge - 1) * (unsigned)L;
            const bool vb = kb + 32 < L;
            const int bo = vb ? 32 : 0;
            if (c1 < 0) continue;  // (handled by the direct kernel launched for such meshes; never on Voronoi meshes with <= 10-cell stencils)
            real ua, ub;
            if (FIELD) {
                ua = p.fzm[kb] * c_rue_a + p.fzp[kb] * c_rum_a;
                ub = p.fzm[kb + bo] * c_rue_b + p.fzp[kb + bo] * c_rum_b;
            } else {
                ua = c_rue_a;
                ub = c_rue_b;
            }
            const real* tab_a = signbit(ua) ? s_cm[er] : s_cp[er];
            const real* tab_b = signbit(ub) ? s_cm[er] : s_cp[er];
            const real* col = s_dyn + kb;
            real f_a = 0.0, f_b = 0.0;
#pragma unroll
            for (int j = 0; j < 10; j += FUNR) {
                real xa[FUNR], xb[FUNR];
#pragma unroll
                for (int t = 0; t < FUNR; ++t) {
                    const real* q = col + s_o[er][j + t];
                    xa[t] = q[0];
                    xb[t] = q[bo];
                }
#pragma unroll
                for (int t = 0; t < FUNR; ++t) {
                    f_a = f_a + tab_a[j + t] * xa[t];
                    f_b = f_b + tab_b[j + t] * xb[t];
                }
            }
            if (FIELD) {
                real* o = fw + (size_t)eoff + kb;
                if (kb > 0) o[0] = ua * f_a;
                if (vb) o[32] = ub * f_b;
            } else {
                real* o = ft + (size_t)eoff + kb;
                o[0] = ua * f_a;
                if (vb) o[32] = ub * f_b;
                if (!RK1) {
                    const real dv = p.dvEdge[iEdge - 1];
                    real* o2 = ft2 + (size_t)eoff + kb;
                    o2[0] = dv * (c_rs_a - c_rue_a) * (real)0.5 * (c_t2_a + c_t1_a);
                    if (vb) o2[32] = dv * (c_rs_b - c_rue_b) * (real)0.5 * (c_t2_b + c_t1_b);
                }
            }
        }
    }
}

// vertical mixing of u (TI:4949-5006), off by default: kept out of line so that it costs the common path no registers
__device__ __noinline__ real u_vmix_term(const DtP& p, int k, int iEdge, int cell1, int cell2, real rho_e) {
    const int L = p.L, L1 = L + 1;
    real um[3];
    for (int d = -1; d <= 1; ++d) {
        const int kk = k + d;
        real x = p.u[IX2(kk, iEdge, L)];
        if (!p.mix_full) x = x - p.u_init[kk - 1] * cos(p.angleEdge[iEdge - 1]) - p.v_init[kk - 1] * sin(p.angleEdge[iEdge - 1]);
        um[d + 1] = x;
    }
    const real z1 = (real)0.5 * (p.zgrid[IX2(k - 1, cell1, L1)] + p.zgrid[IX2(k - 1, cell2, L1)]);
    const real z2 = (real)0.5 * (p.zgrid[IX2(k, cell1, L1)] + p.zgrid[IX2(k, cell2, L1)]);
    const real z3 = (real)0.5 * (p.zgrid[IX2(k + 1, cell1, L1)] + p.zgrid[IX2(k + 1, cell2, L1)]);
    const real z4 = (real)0.5 * (p.zgrid[IX2(k + 2, cell1, L1)] + p.zgrid[IX2(k + 2, cell2, L1)]);
    const real zm = (real)0.5 * (z1 + z2), z0 = (real)0.5 * (z2 + z3), zp = (real)0.5 * (z3 + z4);
    return rho_e * p.v_mom_eddy_visc2 * ((um[2] - um[1]) / (zp - z0) - (um[1] - um[0]) / (z0 - zm)) / ((real)0.5 * (zp - zm));
}

// Shared-memory column slots of one warp of dt_edge_kernel (tma_slot(L + 1) reals each), filled by tma_stage_col.
// (rk 1: the delsq_vorticity columns reuse the rw slots in a second copy phase once the vertical flux is done, which keeps
// the block at 13 slots = 4 blocks per SM)
enum { EC_RW1, EC_RW2, EC_U, EC_PV, EC_RHO, EC_TE, EC_KE1, EC_KE2, EC_HD1, EC_HD2, EC_WZ, EC_NCOL_RK23, EC_DD1 = EC_NCOL_RK23, EC_DD2, EC_NCOL_RK1, EC_DV1 = EC_RW1, EC_DV2 = EC_RW2 };

// One owned edge per warp.  Every column the edge reads exactly once per level (its own u, pv_edge, rho_edge,
// tend_u_euler; rw, ke, h_divergence [, delsq_divergence] of its two cells [, delsq_vorticity of its two vertices]) is
// staged into shared memory by one bulk (TMA) copy per column, all issued back to back by lane 0, so the warp waits for
// memory once and the L1 data path carries only the gathers; the vertical stencils of u then read shared memory.
// The edgesOnEdge gather (TI:4811-4820) stays a direct gather: each lane owns levels (lane, lane + 32) of a 64-level
// chunk so one address serves two loads, the neighbour list holds element offsets and is padded with zero-weight self
// references to a multiple of UNR (q + 0*x == q exactly, q is never -0), and the unrolled body issues 4*UNR
// independent loads before the first dependent add.
// dynamic shared memory: WPB * ncol * tma_slot(L + 1) reals.
template <int UNR, bool RK1, bool VMIX, int MINB = 4>
__global__ void __launch_bounds__(WPB * 32, VMIX ? 1 : MINB) dt_edge_kernel(DtP p) {
    extern __shared__ __align__(16) real s_dyn[];
    __shared__ __align__(8) unsigned long long s_bar[WPB];
    __shared__ unsigned s_off[WPB][20];
    __shared__ real s_wt[WPB][20];
    const int L = p.L, L1 = L + 1, mE2 = p.mE2, LS = tma_slot(L1);
    WARP_ENTITY(p.edgeSolveStart, p.nEdgesSolveLoop);
    const int iEdge = ent;
    // rk 2, 3: the vertical flux of a level overwrites the rw column of cell 1 it was computed from (same lane, same level),
    // 10 slots per warp; rk 1 keeps its own slot (the rw slots are reused by the delsq_vorticity copies)
    real* sc = s_dyn + (size_t)warp * (RK1 ? EC_NCOL_RK1 : EC_NCOL_RK23 - 1) * LS;
    const unsigned mbar = (unsigned)__cvta_generic_to_shared(&s_bar[warp]);
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    const bool del4 = (RK1 && p.h_mom_eddy_visc4 > (real)0.0);
    int vertex1 = 0, vertex2 = 0;
    if (del4) {
        vertex1 = p.verticesOnEdge[IX2(1, iEdge, 2)];
        vertex2 = p.verticesOnEdge[IX2(2, iEdge, 2)];
    }
    const int neoe = p.nEdgesOnEdge[iEdge - 1];
    const unsigned eoff = (unsigned)(iEdge - 1) * (unsigned)L;
    const size_t c1off = (size_t)(cell1 - 1) * L, c2off = (size_t)(cell2 - 1) * L;
    const unsigned s0 = (unsigned)__cvta_generic_to_shared(sc), cs = (unsigned)(LS * sizeof(real));
    const real* g_u = p.u + eoff;
    const real* g_rw1 = p.rw + (c1off + (cell1 - 1));
    const real* g_rw2 = p.rw + (c2off + (cell2 - 1));
    const real* g_pv = p.pv_edge + eoff;
    const real* g_rho = p.rho_edge + eoff;
    const real* g_te = p.tend_u_euler + eoff;
    const real* g_ke1 = p.ke + c1off;
    const real* g_ke2 = p.ke + c2off;
    const real* g_hd1 = p.h_divergence + c1off;
    const real* g_hd2 = p.h_divergence + c2off;
    const real *g_dd1 = nullptr, *g_dd2 = nullptr, *g_dv1 = nullptr, *g_dv2 = nullptr;
    if (del4) {
        g_dd1 = p.delsq_divergence + c1off;
        g_dd2 = p.delsq_divergence + c2off;
        g_dv1 = p.delsq_vorticity + (size_t)(vertex1 - 1) * L;
        g_dv2 = p.delsq_vorticity + (size_t)(vertex2 - 1) * L;
    }
    if (elect_one()) {  // one bulk copy per column, all in flight together
        mbar_init(mbar, 1);
        const unsigned s0 = (unsigned)__cvta_generic_to_shared(sc), cs = (unsigned)(LS * sizeof(real));
        unsigned total = 0;
        tma_issue_col(s0 + EC_U * cs, g_u, L, mbar, total);
        tma_issue_col(s0 + EC_RW1 * cs, g_rw1, L1, mbar, total);
        tma_issue_col(s0 + EC_RW2 * cs, g_rw2, L1, mbar, total);
        tma_issue_col(s0 + EC_PV * cs, g_pv, L, mbar, total);
        tma_issue_col(s0 + EC_RHO * cs, g_rho, L, mbar, total);
        tma_issue_col(s0 + EC_TE * cs, g_te, L, mbar, total);
        tma_issue_col(s0 + EC_KE1 * cs, g_ke1, L, mbar, total);
        tma_issue_col(s0 + EC_KE2 * cs, g_ke2, L, mbar, total);
        tma_issue_col(s0 + EC_HD1 * cs, g_hd1, L, mbar, total);
        tma_issue_col(s0 + EC_HD2 * cs, g_hd2, L, mbar, total);
        if (del4) {
            tma_issue_col(s0 + EC_DD1 * cs, g_dd1, L, mbar, total);
            tma_issue_col(s0 + EC_DD2 * cs, g_dd2, L, mbar, total);
        }
        mbar_arrive_expect_tx(mbar, total);
    }
    // x_<col>: index of level 1 of the column inside sc[]
    const int x_u = EC_U * LS + tma_shift(g_u), x_rw1 = EC_RW1 * LS + tma_shift(g_rw1), x_rw2 = EC_RW2 * LS + tma_shift(g_rw2);
    const int x_pv = EC_PV * LS + tma_shift(g_pv), x_rho = EC_RHO * LS + tma_shift(g_rho), x_te = EC_TE * LS + tma_shift(g_te);
    const int x_ke1 = EC_KE1 * LS + tma_shift(g_ke1), x_ke2 = EC_KE2 * LS + tma_shift(g_ke2);
    const int x_hd1 = EC_HD1 * LS + tma_shift(g_hd1), x_hd2 = EC_HD2 * LS + tma_shift(g_hd2);
    const int x_dd1 = EC_DD1 * LS + tma_shift(g_dd1), x_dd2 = EC_DD2 * LS + tma_shift(g_dd2);
    const int x_dv1 = EC_DV1 * LS + tma_shift(g_dv1), x_dv2 = EC_DV2 * LS + tma_shift(g_dv2);
    const int npad = (neoe + UNR - 1) / UNR * UNR;
    if (lane < npad) {
        const bool real_nb = lane < neoe;
        s_off[warp][lane] = real_nb ? (unsigned)(p.edgesOnEdge[IX2(lane + 1, iEdge, mE2)] - 1) * (unsigned)L : eoff;
        s_wt[warp][lane] = real_nb ? p.weightsOnEdge[IX2(lane + 1, iEdge, mE2)] : (real)0.0;
    }
    const real idc = p.invDcEdge[iEdge - 1];
    real r_dc = 0, r_dv = 0;
    if (del4) {  // TI:4924-4944
        const real u_mix_scale = p.meshScalingDel4[iEdge - 1] * p.h_mom_eddy_visc4;
        r_dc = u_mix_scale * p.del4u_div_factor * idc;
        r_dv = u_mix_scale * fmin(p.invDvEdge[iEdge - 1], 4 * idc);
    }
    __syncwarp();
    mbar_wait(mbar, 0);
    const real* su = sc + x_u - 1;  // 1-based views: su[k] = u(k, iEdge)
    const real* srw1 = sc + x_rw1 - 1;
    const real* srw2 = sc + x_rw2 - 1;
    real* wz = RK1 ? sc + EC_WZ * LS - 1 : sc + x_rw1 - 1;
    for (int kk = lane + 1; kk <= L1; kk += 32) {  // vertical flux of u at the interfaces, TI:4792-4804
        // branch-free: every lane evaluates both the 3rd-order flux (stencil indices clamped into the column) and the
        // 2nd-order boundary form, then selects; a divergent if/else would make the warp execute both anyway
        const int km2 = max(kk - 2, 1), km1 = max(kk - 1, 1), kc = min(kk, L), kp1 = min(kk + 1, L);
        const real rwe = (real)0.5 * (srw1[kk] + srw2[kk]);
        const real u_m2 = su[km2], u_m1 = su[km1], u_c = su[kc], u_p1 = su[kp1];
        const real f3 = flux3(u_m2, u_m1, u_c, u_p1, rwe, (real)1.0);
        const real f2 = rwe * (p.fzm[kc - 1] * u_c + p.fzp[kc - 1] * u_m1);
        real f = (kk == 2 || kk == L) ? f2 : f3;
        if (kk == 1 || kk == L1) f = 0.;
        wz[kk] = f;
    }
    __syncwarp();
    if (del4 && elect_one()) {  // second copy phase into the rw slots (their last read is above)
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        const unsigned s0 = (unsigned)__cvta_generic_to_shared(sc), cs = (unsigned)(LS * sizeof(real));
        unsigned total = 0;
        tma_issue_col(s0 + EC_DV1 * cs, g_dv1, L, mbar, total);
        tma_issue_col(s0 + EC_DV2 * cs, g_dv2, L, mbar, total);
        mbar_arrive_expect_tx(mbar, total);
    }
    bool dv_ready = !del4;
    for (int ka = lane; ka < L; ka += 64) {  // 0-based levels ka and ka + 32
        const bool vb = ka + 32 < L;
        const int bo = vb ? 32 : 0;  // lanes without a second level recompute the first one (no predicated loads)
        const real* __restrict__ pvp = p.pv_edge + ka;
        const real* __restrict__ up = p.u + ka;
        const real pv_a = sc[x_pv + ka], pv_b = sc[x_pv + ka + bo];
        // q = sum_j w_j u_j (0.5 (pv + pv_j)): the factor 0.5 is a power of two, so it commutes with every rounding
        // of the sum and is applied once at the end (identical bits unless an intermediate underflows).
        real q_a = 0.0, q_b = 0.0;
        for (int j = 0; j < npad; j += UNR) {  // TI:4811-4820
            real pa[UNR], ua[UNR], pb[UNR], ub[UNR], wt[UNR];
#pragma unroll
            for (int t = 0; t < UNR; ++t) {
                const unsigned o = s_off[warp][j + t];
                const real* __restrict__ pj = elem_ptr(pvp, o);
                const real* __restrict__ uj = elem_ptr(up, o);
                wt[t] = s_wt[warp][j + t];
                pa[t] = pj[0];
                ua[t] = uj[0];
                pb[t] = pj[bo];
                ub[t] = uj[bo];
            }
#pragma unroll
            for (int t = 0; t < UNR; ++t) {
                q_a = q_a + wt[t] * ua[t] * (pv_a + pa[t]);
                q_b = q_b + wt[t] * ub[t] * (pv_b + pb[t]);
            }
        }
        q_a = (real)0.5 * q_a;
        q_b = (real)0.5 * q_b;
        if (!dv_ready) {
            mbar_wait(mbar, 1);
            dv_ready = true;
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !vb) break;
            const int k0 = ka + 32 * h, k = k0 + 1;  // k0: 0-based level
            const real q = h ? q_b : q_a;
            const size_t ie = (size_t)eoff + k0;
            real tu = -p.rdzw[k0] * (wz[k + 1] - wz[k]);
            const real ue = sc[x_u + k0], rho_e = sc[x_rho + k0];
            tu = tu + rho_e * (q - (sc[x_ke2 + k0] - sc[x_ke1 + k0]) * idc) - ue * (real)0.5 * (sc[x_hd1 + k0] + sc[x_hd2 + k0]);
            real te = sc[x_te + k0];
            if (RK1) {
                bool changed = false;
                if (del4) {
                    const real u_diffusion = rho_e * ((sc[x_dd2 + k0] - sc[x_dd1 + k0]) * r_dc - (sc[x_dv2 + k0] - sc[x_dv1 + k0]) * r_dv);
                    te = te - u_diffusion;
                    changed = true;
                }
                if (VMIX && k >= 2 && k <= L - 1) {  // TI:4949-5006
                    te = te + u_vmix_term(p, k, iEdge, cell1, cell2, rho_e);
                    changed = true;
                }
                if (changed) p.tend_u_euler[ie] = te;
            }
            if (p.rayleigh_damp_u && k >= L - p.n_rayleigh_levels + 1) {  // TI:5015-5028
                const real coef = (real)(k - (L - p.n_rayleigh_levels)) * p.rayleigh_coef_inverse;
                tu = tu - rho_e * ue * coef;
            }
            p.tend_u[ie] = tu + te + (p.tend_ru_physics ? p.tend_ru_physics[ie] : (real)0.0);
        }
    }
}

// acc -= sum_t sg[t] * col_t(level), two levels per lane, all 2*NE loads issued before the first dependent subtract
// (the sum keeps the list order; one memory round trip per call instead of one per edge)
template <int NE>
__device__ __forceinline__ void gather_sub(const real* __restrict__ base_ka, const unsigned* eo, const real* sg, int bo, real& acc_a, real& acc_b) {
    real xa[NE], xb[NE];
#pragma unroll
    for (int t = 0; t < NE; ++t) {
        const real* __restrict__ q = elem_ptr(base_ka, eo[t]);
        xa[t] = q[0];
        xb[t] = q[bo];
    }
#pragma unroll
    for (int t = 0; t < NE; ++t) {
        const real s = sg[t];
        acc_a = acc_a - s * xa[t];
        acc_b = acc_b - s * xb[t];
    }
}
// acc -= sum_t coef[t] * (col2_t(level) - col1_t(level))
template <int NE>
__device__ __forceinline__ void gather_sub_diff(const real* __restrict__ base_ka, const unsigned* o2, const unsigned* o1, const real* coef, int bo, real& acc_a,
                                                real& acc_b) {
    real ya[NE], yb[NE], za[NE], zb[NE];
#pragma unroll
    for (int t = 0; t < NE; ++t) {
        const real* __restrict__ q2 = elem_ptr(base_ka, o2[t]);
        const real* __restrict__ q1 = elem_ptr(base_ka, o1[t]);
        ya[t] = q2[0];
        yb[t] = q2[bo];
        za[t] = q1[0];
        zb[t] = q1[bo];
    }
#pragma unroll
    for (int t = 0; t < NE; ++t) {
        const real e = coef[t];
        acc_a = acc_a - e * (ya[t] - za[t]);
        acc_b = acc_b - e * (yb[t] - zb[t]);
    }
}

// Shared-memory column slots of one warp of dt_cell_kernel (tma_slot(L + 1) reals each)
enum { CC_W, CC_RW, CC_TH, CC_RWS, CC_THS, CC_TWE, CC_TTE, CC_RHO, CC_WDW, CC_WDT, CC_NCOL_RK23, CC_CQW = CC_NCOL_RK23, CC_PP, CC_DPDZ, CC_NCOL_RK1 };

// w and theta_m tendencies of one owned cell per warp (w: flux divergence TI:5046-5074, rk1 del4 TI:5134-5154, vertical
// advection, pgf + buoyancy, final sum TI:5167-5223; theta_m: TI:5231-5269, 5304-5325, 5331-5414).
// The cell's own columns (w, rw, theta_m, rw_save, theta_m_save, tend_w_euler, tend_theta_euler, rho_zz [, cqw, pp, dpdz])
// are staged into shared memory by one bulk (TMA) copy each, issued by lane 0 before the stencil metadata is even
// read, so every vertical stencil and every single-use operand costs a shared-memory read and the warp waits for
// memory once.  The edge gathers (fw, ft, ft2, delsq_*) are direct loads: each lane owns levels (lane, lane + 32), the
// edge list is staged as element offsets with signs (and the rk1 del4 coefficients), padded with zero entries to 6 or 8
// (x - 0*y == x exactly, the sums are never -0), and all loads of a sum are issued before its first subtract.
// dynamic shared memory: WPB * ncol * tma_slot(L + 1) reals.
template <bool RK1, int MINB>
__global__ void __launch_bounds__(WPB * 32, MINB) dt_cell_kernel(DtP p, const real* __restrict__ ft, const real* __restrict__ fw, const real* __restrict__ ft2) {
    extern __shared__ __align__(16) real s_dyn[];
    __shared__ __align__(8) unsigned long long s_bar[WPB];
    __shared__ unsigned s_eo[WPB][8], s_c1[WPB][8], s_c2[WPB][8];
    __shared__ real s_sign[WPB][8], s_esw[WPB][8], s_est[WPB][8];
    const int L = p.L, L1 = L + 1, mE = p.mE, LS = tma_slot(L1);
    WARP_ENTITY(p.cellSolveStart, p.nCellsSolveLoop);
    const int iCell = ent;
    real* sc = s_dyn + (size_t)warp * (RK1 ? CC_NCOL_RK1 : CC_NCOL_RK23) * LS;
    const unsigned mbar = (unsigned)__cvta_generic_to_shared(&s_bar[warp]);
    const unsigned coff = (unsigned)(iCell - 1) * (unsigned)L, woff = (unsigned)(iCell - 1) * (unsigned)L1;
    const real *g_w = p.w + woff, *g_rw = p.rw + woff, *g_th = p.theta_m + coff, *g_rws = p.rw_save + woff, *g_ths = p.theta_m_save + coff;
    const real *g_twe = p.tend_w_euler + woff, *g_tte = p.tend_theta_euler + coff, *g_rho = p.rho_zz + coff;
    const real *g_cqw = p.cqw + coff, *g_pp = p.pp + coff, *g_dpdz = p.dpdz + coff;
    if (elect_one()) {
        mbar_init(mbar, 1);
        const unsigned s0 = (unsigned)__cvta_generic_to_shared(sc), cs = (unsigned)(LS * sizeof(real));
        unsigned total = 0;
        tma_issue_col(s0 + CC_W * cs, g_w, L1, mbar, total);
        tma_issue_col(s0 + CC_RW * cs, g_rw, L1, mbar, total);
        tma_issue_col(s0 + CC_TH * cs, g_th, L, mbar, total);
        tma_issue_col(s0 + CC_RWS * cs, g_rws, L1, mbar, total);
        tma_issue_col(s0 + CC_THS * cs, g_ths, L, mbar, total);
        tma_issue_col(s0 + CC_TWE * cs, g_twe, L1, mbar, total);
        tma_issue_col(s0 + CC_TTE * cs, g_tte, L, mbar, total);
        tma_issue_col(s0 + CC_RHO * cs, g_rho, L, mbar, total);
        if (RK1) {
            tma_issue_col(s0 + CC_CQW * cs, g_cqw, L, mbar, total);
            tma_issue_col(s0 + CC_PP * cs, g_pp, L, mbar, total);
            tma_issue_col(s0 + CC_DPDZ * cs, g_dpdz, L, mbar, total);
        }
        mbar_arrive_expect_tx(mbar, total);
    }
    // 1-based views of the staged columns: s_w[k] = w(k, iCell), ...
    const real* s_w = sc + CC_W * LS + tma_shift(g_w) - 1;
    const real* s_rw = sc + CC_RW * LS + tma_shift(g_rw) - 1;
    const real* s_th = sc + CC_TH * LS + tma_shift(g_th) - 1;
    const real* s_rws = sc + CC_RWS * LS + tma_shift(g_rws) - 1;
    const real* s_ths = sc + CC_THS * LS + tma_shift(g_ths) - 1;
    const real* s_twe = sc + CC_TWE * LS + tma_shift(g_twe) - 1;
    const real* s_tte = sc + CC_TTE * LS + tma_shift(g_tte) - 1;
    const real* s_rho = sc + CC_RHO * LS + tma_shift(g_rho) - 1;
    const real* s_cqw = sc + CC_CQW * LS + tma_shift(g_cqw) - 1;
    const real* s_pp = sc + CC_PP * LS + tma_shift(g_pp) - 1;
    const real* s_dpdz = sc + CC_DPDZ * LS + tma_shift(g_dpdz) - 1;
    real* wdw = sc + CC_WDW * LS - 1;
    real* wdt = sc + CC_WDT * LS - 1;

    const int ne = p.nEdgesOnCell[iCell - 1];
    const int npad = ne <= 6 ? 6 : 8;  // padded edge count (zero signs / coefficients beyond ne)
    const real invArea = p.invAreaCell[iCell - 1];
    const bool del4w = (RK1 && p.h_mom_eddy_visc4 > (real)0.0);
    const bool del4t = (RK1 && p.h_theta_eddy_visc4 > (real)0.0);
    if (lane < npad) {
        const bool in = lane < ne;
        const int iEdge = p.edgesOnCell[IX2(in ? lane + 1 : 1, iCell, mE)];
        const real sgn = in ? p.edgesOnCell_sign[IX2(lane + 1, iCell, mE)] : (real)0.0;
        s_eo[warp][lane] = (unsigned)(iEdge - 1) * (unsigned)L;
        s_sign[warp][lane] = sgn;
        if (del4w || del4t) {
            s_c1[warp][lane] = (unsigned)(p.cellsOnEdge[IX2(1, iEdge, 2)] - 1) * (unsigned)L;
            s_c2[warp][lane] = (unsigned)(p.cellsOnEdge[IX2(2, iEdge, 2)] - 1) * (unsigned)L;
            const real ms4 = p.meshScalingDel4[iEdge - 1], dv = p.dvEdge[iEdge - 1], idc = p.invDcEdge[iEdge - 1];
            s_esw[warp][lane] = ms4 * (p.h_mom_eddy_visc4 * invArea) * dv * sgn * idc;                    // TI:5141-5146
            s_est[warp][lane] = ms4 * (p.h_theta_eddy_visc4 * p.prandtl_inv * invArea) * dv * sgn * idc;  // TI:5309-5315
        }
    }
    __syncwarp();
    mbar_wait(mbar, 0);
    // ---- vertical fluxes at the interfaces (TI:5167-5185, 5331-5353), branch-free: stencil levels are clamped into the
    // column, both the 3rd-order and the 2nd-order (boundary) forms are evaluated and one is selected
    for (int kk = lane + 1; kk <= L1; kk += 32) {
        const int k = min(max(kk, 2), L);  // evaluation level (results for kk = 1, L+1 are discarded)
        const int km2 = max(k - 2, 1), kp1 = min(k + 1, L);
        const real rwk = s_rw[k], rwkm = s_rw[k - 1];
        const real wk = s_w[k], wkm = s_w[k - 1];
        const bool bnd = (k == 2 || k == L), ends = (kk == 1 || kk == L1);
        {
            const real f2 = (real)0.25 * (rwk + rwkm) * (wk + wkm);
            const real f3 = flux3(s_w[km2], wkm, wk, s_w[kp1], (real)0.5 * (rwk + rwkm), (real)1.0);
            wdw[kk] = ends ? (real)0.0 : (bnd ? f2 : f3);
        }
        {
            const real fm = p.fzm[k - 1], fp = p.fzp[k - 1];
            const real th_k = s_th[k], th_km = s_th[k - 1];
            const real rws = s_rws[k];
            const real c2 = fm * th_k + fp * th_km;
            const real f3 = flux3(s_th[km2], th_km, th_k, s_th[kp1], rwk, p.coef_3rd_order);
            const real corr = (rws - rwk) * (fm * s_ths[k] + fp * s_ths[k - 1]);
            const real inner = ((k == 2) ? rwk * c2 : f3) + corr;
            wdt[kk] = ends ? (real)0.0 : ((k == L) ? rws * c2 : inner);
        }
    }
    __syncwarp();
    // ---- w, TI:5046-5074, 5134-5154, 5167-5223: levels k = ka + 1 and ka + 33 of (1..L+1)
    for (int ka = lane; ka < L1; ka += 64) {
        const bool vb = ka + 32 < L1;
        const int bo = vb ? 32 : 0;
        real tw_a = 0.0, tw_b = 0.0;
        real twe_a = s_twe[ka + 1], twe_b = s_twe[ka + 1 + bo];
        {
            const real* __restrict__ fwp = fw + ka;  // fw(k, e): element (e-1)*L + k-1
            gather_sub<6>(fwp, s_eo[warp], s_sign[warp], bo, tw_a, tw_b);
            if (npad > 6) gather_sub<2>(fwp, s_eo[warp] + 6, s_sign[warp] + 6, bo, tw_a, tw_b);
            if (del4w) {
                const real* __restrict__ dsw = p.delsq_w + ka;
                gather_sub_diff<6>(dsw, s_c2[warp], s_c1[warp], s_esw[warp], bo, twe_a, twe_b);
                if (npad > 6) gather_sub_diff<2>(dsw, s_c2[warp] + 6, s_c1[warp] + 6, s_esw[warp] + 6, bo, twe_a, twe_b);
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !vb) break;
            const int k = ka + 1 + 32 * h;
            const size_t iw = (size_t)woff + (k - 1);
            if (k == 1 || k == L1) {
                p.tend_w[iw] = 0.0;
                continue;
            }
            real tw = h ? tw_b : tw_a, twe = h ? twe_b : twe_a;
            tw = tw * invArea - p.rdzu[k - 1] * (wdw[k + 1] - wdw[k]);
            if (RK1) {
                twe = twe - s_cqw[k] * (p.rdzu[k - 1] * (s_pp[k] - s_pp[k - 1]) - (p.fzm[k - 1] * s_dpdz[k] + p.fzp[k - 1] * s_dpdz[k - 1]));
                if (p.v_mom_eddy_visc2 > (real)0.0) {
                    const real wk = s_w[k];
                    twe = twe + p.v_mom_eddy_visc2 * (real)0.5 * (s_rho[k] + s_rho[k - 1]) *
                                    ((s_w[k + 1] - wk) * p.rdzw[k - 1] - (wk - s_w[k - 1]) * p.rdzw[k - 2]) * p.rdzu[k - 1];
                }
                p.tend_w_euler[iw] = twe;
            }
            p.tend_w[iw] = tw + twe;
        }
    }
    // ---- theta_m, TI:5231-5269, 5304-5325, 5331-5414
    for (int ka = lane; ka < L; ka += 64) {
        const bool vb = ka + 32 < L;
        const int bo = vb ? 32 : 0;
        real tt_a = 0.0, tt_b = 0.0;
        real tte_a = s_tte[ka + 1], tte_b = s_tte[ka + 1 + bo];
        {
            const real* __restrict__ ftp = ft + ka;
            gather_sub<6>(ftp, s_eo[warp], s_sign[warp], bo, tt_a, tt_b);
            if (npad > 6) gather_sub<2>(ftp, s_eo[warp] + 6, s_sign[warp] + 6, bo, tt_a, tt_b);
            if (!RK1) {
                const real* __restrict__ f2p = ft2 + ka;
                gather_sub<6>(f2p, s_eo[warp], s_sign[warp], bo, tt_a, tt_b);
                if (npad > 6) gather_sub<2>(f2p, s_eo[warp] + 6, s_sign[warp] + 6, bo, tt_a, tt_b);
            }
            if (del4t) {
                const real* __restrict__ dst = p.delsq_theta + ka;
                gather_sub_diff<6>(dst, s_c2[warp], s_c1[warp], s_est[warp], bo, tte_a, tte_b);
                if (npad > 6) gather_sub_diff<2>(dst, s_c2[warp] + 6, s_c1[warp] + 6, s_est[warp] + 6, bo, tte_a, tte_b);
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !vb) break;
            const int k = ka + 1 + 32 * h;
            const size_t ic = (size_t)coff + (k - 1);
            real tt = h ? tt_b : tt_a, tte = h ? tte_b : tte_a;
            tt = tt * invArea - p.rdzw[k - 1] * (wdt[k + 1] - wdt[k]);
            const real rho = s_rho[k];
            if (p.tend_rtheta_adv) p.tend_rtheta_adv[ic] = tt;
            if (p.rthdynten) p.rthdynten[ic] = tt / rho;
            tt = tt + rho * (p.rt_diabatic_tend ? p.rt_diabatic_tend[ic] : (real)0.0);
            if (RK1 && p.v_theta_eddy_visc2 > (real)0.0 && k >= 2 && k <= L - 1) {  // TI:5360-5406
                const real z1 = p.zgrid[IX2(k - 1, iCell, L1)], z2 = p.zgrid[IX2(k, iCell, L1)], z3 = p.zgrid[IX2(k + 1, iCell, L1)], z4 = p.zgrid[IX2(k + 2, iCell, L1)];
                const real zm = (real)0.5 * (z1 + z2), z0 = (real)0.5 * (z2 + z3), zp = (real)0.5 * (z3 + z4);
                real a = s_th[k + 1], b = s_th[k], c = s_th[k - 1];
                if (p.mix_full) {
                    tte = tte + p.v_theta_eddy_visc2 * p.prandtl_inv * rho * ((a - b) / (zp - z0) - (b - c) / (z0 - zm)) / ((real)0.5 * (zp - zm));
                } else {
                    const real ta = p.t_init[ic + 1], tb = p.t_init[ic], tc = p.t_init[ic - 1];
                    tte = tte + p.v_theta_eddy_visc2 * p.prandtl_inv * rho * (((a - ta) - (b - tb)) / (zp - z0) - ((b - tb) - (c - tc)) / (z0 - zm)) / ((real)0.5 * (zp - zm));
                }
            }
            if (RK1 && (del4t || p.v_theta_eddy_visc2 > (real)0.0)) p.tend_theta_euler[ic] = tte;
            p.tend_theta[ic] = tt + tte + (p.tend_rtheta_physics ? p.tend_rtheta_physics[ic] : (real)0.0);
        }
    }
}

}  // namespace

extern "C" int mpas_b200_compute_dyn_tend_work(MPAS_DYNTEND_ARGS) {
    (void)nCells; (void)nVertices; (void)nEdgesSolve; (void)num_scalars; (void)moist_start; (void)moist_end;
    (void)fEdge; (void)rr; (void)pressure_b; (void)exner; (void)scalars; (void)deriv_two; (void)cellsOnCell; (void)latCell; (void)latEdge;
    (void)qv_init; (void)r_earth; (void)ur_cell; (void)vr_cell; (void)tend_w_pgf; (void)tend_w_buoy; (void)vertexSolveStart; (void)vertexSolveEnd;
    if (nVertLevels < 4) return fail_msg("mpas_b200_compute_dyn_tend_work: nVertLevels must be >= 4");
    if (vertexDegree != 3) return fail_msg("mpas_b200_compute_dyn_tend_work: vertexDegree must be 3");
    DtP p;
    p.L = nVertLevels; p.km.set(nVertLevels, nVertLevels + 1); p.mE = maxEdges; p.mE2 = maxEdges2; p.vD = vertexDegree; p.rk_step = rk_step; p.nCellsSolve = nCellsSolve;
    p.cellStart = cellStart; p.nCellsLoop = cellEnd - cellStart + 1; p.cellSolveStart = cellSolveStart; p.nCellsSolveLoop = cellSolveEnd - cellSolveStart + 1;
    p.edgeStart = edgeStart; p.nEdgesLoop = edgeEnd - edgeStart + 1; p.edgeSolveStart = edgeSolveStart; p.edgeSolveEnd = edgeSolveEnd;
    p.nEdgesSolveLoop = edgeSolveEnd - edgeSolveStart + 1; p.vertexStart = vertexStart; p.nVerticesLoop = vertexEnd - vertexStart + 1;
    static const int pf_mode = tune("dt_pf", 0);
    p.pf = pf_mode;
    p.mixing = config_horiz_mixing; p.mix_full = config_mix_full; p.rayleigh_damp_u = config_rayleigh_damp_u;
    p.n_rayleigh_levels = config_number_rayleigh_damp_u_levels;
    p.c_s = c_s; p.len_disp = config_len_disp; p.invDt = (real)1.0 / dt; p.cam_coef = config_mpas_cam_coef;
    p.h_theta_eddy_visc2_fixed = config_h_theta_eddy_visc2; p.del4u_div_factor = config_del4u_div_factor;
    p.h_mom_eddy_visc4 = 0.0; p.h_theta_eddy_visc4 = 0.0;
    if (rk_step == 1) {  // TI:4697-4704: locals set only on the first RK stage
        if (config_horiz_mixing == 1) {
            p.h_mom_eddy_visc4 = config_visc4_2dsmag * (config_len_disp * config_len_disp * config_len_disp);
            p.h_theta_eddy_visc4 = p.h_mom_eddy_visc4;
        } else if (config_horiz_mixing == 2) {
            p.h_mom_eddy_visc4 = config_h_mom_eddy_visc4;
            p.h_theta_eddy_visc4 = config_h_theta_eddy_visc4;
        }
    }
    (void)config_h_mom_eddy_visc2;
    p.v_mom_eddy_visc2 = config_v_mom_eddy_visc2; p.v_theta_eddy_visc2 = config_v_theta_eddy_visc2; p.coef_3rd_order = coef_3rd_order;
    p.rayleigh_coef_inverse = config_rayleigh_damp_u
                                  ? (real)1.0 / ((real)config_number_rayleigh_damp_u_levels * (config_rayleigh_damp_u_timescale_days * (real)86400.0))
                                  : (real)0.0;
    p.prandtl_inv = (real)1.0 / MPAS_PRANDTL;
    p.fEdge = X(fEdge); p.dvEdge = X(dvEdge); p.dcEdge = X(dcEdge); p.invDcEdge = X(invDcEdge); p.invDvEdge = X(invDvEdge); p.invAreaCell = X(invAreaCell);
    p.invAreaTriangle = X(invAreaTriangle); p.meshScalingDel2 = X(meshScalingDel2); p.meshScalingDel4 = X(meshScalingDel4);
    p.weightsOnEdge = X(weightsOnEdge); p.zgrid = X(zgrid); p.rho_edge = X(rho_edge); p.rho_zz = X(rho_zz); p.ru = X(ru); p.u = X(u); p.v = X(v);
    p.divergence = X(divergence); p.vorticity = X(vorticity); p.ke = X(ke); p.pv_edge = X(pv_edge); p.theta_m = X(theta_m); p.rw = X(rw);
    p.rt_diabatic_tend = X(rt_diabatic_tend); p.w = X(w); p.cqw = X(cqw); p.rb = X(rb); p.pp = X(pp); p.zz = X(zz); p.zxu = X(zxu); p.cqu = X(cqu);
    p.edgesOnCell_sign = X(edgesOnCell_sign); p.edgesOnVertex_sign = X(edgesOnVertex_sign); p.rw_save = X(rw_save); p.ru_save = X(ru_save);
    p.theta_m_save = X(theta_m_save); p.rr_save = X(rr_save); p.angleEdge = X(angleEdge); p.u_init = X(u_init); p.v_init = X(v_init);
    p.adv_coefs = X(adv_coefs); p.adv_coefs_3rd = X(adv_coefs_3rd); p.rdzu = X(rdzu); p.rdzw = X(rdzw); p.fzm = X(fzm); p.fzp = X(fzp); p.t_init = X(t_init);
    p.defc_a = X(defc_a); p.defc_b = X(defc_b); p.qtot = X(qtot); p.tend_ru_physics = X(tend_ru_physics); p.tend_rtheta_physics = X(tend_rtheta_physics);
    p.tend_rho_physics = X(tend_rho_physics);
    p.cellsOnEdge = X(cellsOnEdge); p.verticesOnEdge = X(verticesOnEdge); p.edgesOnCell = X(edgesOnCell); p.edgesOnEdge = X(edgesOnEdge);
    p.edgesOnVertex = X(edgesOnVertex); p.nEdgesOnCell = X(nEdgesOnCell); p.nEdgesOnEdge = X(nEdgesOnEdge); p.advCellsForEdge = X(advCellsForEdge);
    p.nAdvCellsForEdge = X(nAdvCellsForEdge);
    p.tend_u = X(tend_u); p.tend_rho = X(tend_rho); p.tend_theta = X(tend_theta); p.tend_w = X(tend_w); p.h_divergence = X(h_divergence); p.kdiff = X(kdiff);
    p.tend_u_euler = X(tend_u_euler); p.tend_w_euler = X(tend_w_euler); p.tend_theta_euler = X(tend_theta_euler); p.tend_rtheta_adv = X(tend_rtheta_adv);
    p.rthdynten = X(rthdynten); p.delsq_theta = X(delsq_theta); p.delsq_w = X(delsq_w); p.delsq_divergence = X(delsq_divergence); p.delsq_u = X(delsq_u);
    p.delsq_vorticity = X(delsq_vorticity); p.dpdz = X(dpdz);
    if (!p.mix_full && (p.v_mom_eddy_visc2 > 0 || p.v_theta_eddy_visc2 > 0) && (!u_init || !v_init || !angleEdge || !t_init))
        return fail_msg("mpas_b200_compute_dyn_tend_work: config_mix_full=false needs u_init, v_init, angleEdge, t_init");
    const int L = p.L;
    if (maxEdges > 8) return fail_msg("mpas_b200_compute_dyn_tend_work: maxEdges > 8 or maxEdges2 > 16 not supported");
    {
        const bool smag = (rk_step == 1 && config_horiz_mixing == 1);
        const long long nt = (long long)p.nCellsLoop * L;
        if (maxEdges <= 6) {
            static const int pre_minb = tune("dt_pre_minb", 4);
            if (smag && pre_minb == 4) KLAUNCH((dt_cell_pre_kernel<6, true, 4>), nt, 256, p);
            else if (smag) KLAUNCH((dt_cell_pre_kernel<6, true>), nt, 256, p);
            else KLAUNCH((dt_cell_pre_kernel<6, false>), nt, 256, p);
        } else {
            if (smag) KLAUNCH((dt_cell_pre_kernel<8, true>), nt, 256, p);
            else KLAUNCH((dt_cell_pre_kernel<8, false>), nt, 256, p);
        }
    }
    KCHECK("dt_cell_pre_kernel");
    if (rk_step == 1) {
        static const int emix_minb = tune("dt_emix_minb", 6), cmix_minb = tune("dt_cmix_minb", 0);
        if (emix_minb == 6) KLAUNCH(dt_edge_mix_kernel<6>, (long long)p.nEdgesLoop * L, 256, p);
        else KLAUNCH(dt_edge_mix_kernel<0>, (long long)p.nEdgesLoop * L, 256, p);
        KCHECK("dt_edge_mix_kernel");
        if (p.h_mom_eddy_visc4 > (real)0.0) {
            KLAUNCH(dt_vertex_mix_kernel, (long long)p.nVerticesLoop * L, 256, p);
            KCHECK("dt_vertex_mix_kernel");
        }
        static const int cmix_warp = tune("dt_cmix_warp", 3);  // blocks per SM of the warp form, 0 = flat kernel
        if (cmix_warp == 3 && maxEdges <= 8) WLAUNCH(dt_cell_mix_warp_kernel<3>, p.nCellsLoop, 0, p);
        else if (cmix_warp == 4 && maxEdges <= 8) WLAUNCH(dt_cell_mix_warp_kernel<4>, p.nCellsLoop, 0, p);
        else if (cmix_warp == 5 && maxEdges <= 8) WLAUNCH(dt_cell_mix_warp_kernel<5>, p.nCellsLoop, 0, p);
        else if (cmix_warp && maxEdges <= 8) WLAUNCH(dt_cell_mix_warp_kernel<6>, p.nCellsLoop, 0, p);
        else if (cmix_minb == 5) KLAUNCH(dt_cell_mix_kernel<5>, (long long)p.nCellsLoop * (L + 1), 256, p);
        else if (cmix_minb == 6) KLAUNCH(dt_cell_mix_kernel<6>, (long long)p.nCellsLoop * (L + 1), 256, p);
        else KLAUNCH(dt_cell_mix_kernel<0>, (long long)p.nCellsLoop * (L + 1), 256, p);
        KCHECK("dt_cell_mix_kernel");
    }
    if (maxEdges > 8 || maxEdges2 > 16) return fail_msg("mpas_b200_compute_dyn_tend_work: maxEdges > 8 or maxEdges2 > 16 not supported");
    const size_t nEL = (size_t)(nEdges + 1) * L;
    // per-edge high-order fluxes of theta and w: per-block workspace keyed by the block's tend_u (runtime.cu)
    const size_t nELp = (nEL + 1) & ~(size_t)1;
    real* ft = keyed_workspace(p.tend_u, 3 * nELp, 1);
    if (!ft) return fail_msg("mpas_b200_compute_dyn_tend_work: cannot allocate the edge-flux workspace (the first call must be outside stream capture)");
    real* fw = ft + nELp;
    real* ft2 = fw + nELp;
    static const int flux_minb = tune("dt_flux_minb", 6), flux_tile = tune("dt_flux_tile", 2);
    // shared-memory staged stencil (tiles.cu) when the plan exists and two column sets of ncmax cells fit; else direct gathers
    const EdgeTilePlan* tp = nullptr;
    const size_t tile_shm = (size_t)2 * FT_NCMAX * tma_slot(L + 1) * sizeof(real);
    if (flux_tile && 2 * FT_NCMAX * tma_slot(L + 1) < 65536 && p.nEdgesLoop > 0 &&
        ensure_dyn_smem(rk_step == 1 ? (const void*)dt_edge_flux_tile_kernel<true> : (const void*)dt_edge_flux_tile_kernel<false>, tile_shm))
        tp = edge_tile_plan(p.advCellsForEdge, p.nAdvCellsForEdge, p.edgeStart, p.nEdgesLoop, nCells, flux_tile == 2 ? FT1_NCMAX : FT_NCMAX, FT_TEMAX);
    if (tp && flux_tile == 2) {
        const size_t shm1 = (size_t)FT1_NCMAX * tma_slot(L + 1) * sizeof(real);
        if (!ensure_dyn_smem((const void*)dt_edge_flux_tile1_kernel<0, true>, shm1) || !ensure_dyn_smem((const void*)dt_edge_flux_tile1_kernel<0, false>, shm1) ||
            !ensure_dyn_smem((const void*)dt_edge_flux_tile1_kernel<1, true>, shm1) || tp->max_stencil > 10)
            tp = nullptr;
    }
    if (tp && flux_tile == 2) {
        const size_t shm1 = (size_t)FT1_NCMAX * tma_slot(L + 1) * sizeof(real);
        ktime_begin(rk_step == 1 ? "dt_edge_flux_tile1_kernel<0,true>" : "dt_edge_flux_tile1_kernel<0,false>");
        if (rk_step == 1) dt_edge_flux_tile1_kernel<0, true><<<tp->nTiles, WPB * 32, shm1, stream()>>>(p, *tp, ft, fw, ft2);
        else dt_edge_flux_tile1_kernel<0, false><<<tp->nTiles, WPB * 32, shm1, stream()>>>(p, *tp, ft, fw, ft2);
        ktime_end();
        ktime_begin("dt_edge_flux_tile1_kernel<1>");
        dt_edge_flux_tile1_kernel<1, true><<<tp->nTiles, WPB * 32, shm1, stream()>>>(p, *tp, ft, fw, ft2);
        ktime_end();
        count_launch(2);
    } else if (tp) {
        ktime_begin(rk_step == 1 ? "dt_edge_flux_tile_kernel<true>" : "dt_edge_flux_tile_kernel<false>");
        if (rk_step == 1) dt_edge_flux_tile_kernel<true><<<tp->nTiles, WPB * 32, tile_shm, stream()>>>(p, *tp, ft, fw, ft2);
        else dt_edge_flux_tile_kernel<false><<<tp->nTiles, WPB * 32, tile_shm, stream()>>>(p, *tp, ft, fw, ft2);
        ktime_end();
        count_launch();
    } else if (rk_step == 1) WLAUNCH((dt_edge_flux_kernel<true, 0>), p.nEdgesLoop, 0, p, ft, fw, ft2);
    else if (flux_minb == 5) WLAUNCH((dt_edge_flux_kernel<false, 5>), p.nEdgesLoop, 0, p, ft, fw, ft2);
    else if (flux_minb == 6) WLAUNCH((dt_edge_flux_kernel<false, 6>), p.nEdgesLoop, 0, p, ft, fw, ft2);
    else WLAUNCH((dt_edge_flux_kernel<false, 0>), p.nEdgesLoop, 0, p, ft, fw, ft2);
    KCHECK("dt_edge_flux_kernel");
    {
        const size_t shm = (size_t)WPB * (rk_step == 1 ? EC_NCOL_RK1 : EC_NCOL_RK23 - 1) * tma_slot(L + 1) * sizeof(real);
        static const int de_minb = tune("dt_edge_minb", 5);
        if (rk_step != 1 && de_minb == 5) WLAUNCH((dt_edge_kernel<5, false, false, 5>), p.nEdgesSolveLoop, shm, p);
        else if (rk_step != 1 && de_minb == 3) WLAUNCH((dt_edge_kernel<5, false, false, 3>), p.nEdgesSolveLoop, shm, p);
        else if (rk_step != 1) WLAUNCH((dt_edge_kernel<5, false, false>), p.nEdgesSolveLoop, shm, p);
        else if (p.v_mom_eddy_visc2 > (real)0.0) WLAUNCH((dt_edge_kernel<5, true, true>), p.nEdgesSolveLoop, shm, p);
        else WLAUNCH((dt_edge_kernel<5, true, false>), p.nEdgesSolveLoop, shm, p);
        KCHECK("dt_edge_kernel");
    }
    {
        static const int minb1 = tune("dt_cell_minb_rk1", 4), minb23 = tune("dt_cell_minb", 5);  // measured: 438 / 464 us (rk 1), 258 / 246 us (rk 2, 3) at 4 / 5
        const int minb = rk_step == 1 ? minb1 : minb23;
        const size_t shm = (size_t)WPB * (rk_step == 1 ? CC_NCOL_RK1 : CC_NCOL_RK23) * tma_slot(L + 1) * sizeof(real);
        if (rk_step == 1) {
            if (minb == 3) WLAUNCH((dt_cell_kernel<true, 3>), p.nCellsSolveLoop, shm, p, ft, fw, ft2);
            else if (minb == 5) WLAUNCH((dt_cell_kernel<true, 5>), p.nCellsSolveLoop, shm, p, ft, fw, ft2);
            else if (minb == 6) WLAUNCH((dt_cell_kernel<true, 6>), p.nCellsSolveLoop, shm, p, ft, fw, ft2);
            else WLAUNCH((dt_cell_kernel<true, 4>), p.nCellsSolveLoop, shm, p, ft, fw, ft2);
        } else {
            if (minb == 3) WLAUNCH((dt_cell_kernel<false, 3>), p.nCellsSolveLoop, shm, p, ft, fw, ft2);
            else if (minb == 5) WLAUNCH((dt_cell_kernel<false, 5>), p.nCellsSolveLoop, shm, p, ft, fw, ft2);
            else if (minb == 6) WLAUNCH((dt_cell_kernel<false, 6>), p.nCellsSolveLoop, shm, p, ft, fw, ft2);
            else WLAUNCH((dt_cell_kernel<false, 4>), p.nCellsSolveLoop, shm, p, ft, fw, ft2);
        }
    }
    KCHECK("dt_cell_kernel");
    return 0;
}
