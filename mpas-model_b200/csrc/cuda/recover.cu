// atm_recover_large_step_variables_work (TI:2909-3099) and atm_set_smlstep_pert_variables_work
// (TI:2224-2309).  TI = src/core_atmosphere/dynamics/mpas_atm_time_integration.F.
//
// The reference routine has three phases separated by OpenMP barriers (TI:3046, 3061); they are the
// three kernels here (a cell needs rho_zz of its edge neighbours, an edge the ru of ... etc.):
//   recover_cell_kernel   (k in 1..L+1, all cells)  rho_p, rho_zz, wwAvg, rw, first part of w, rtheta_p,
//                         theta_m, and on rk_step 3 exner (pow) and pressure_p      TI:2998-3039
//   recover_edge_kernel   (k, all edges)            ruAvg, ru, u                       TI:3048-3059
//   recover_w_kernel      (k, all cells)            terrain contribution to w through zb_cell/zb3_cell,
//                                                   division by the interface density  TI:3063-3097
// All are one thread per (k, column) with k fastest (coalesced 8-byte lanes along the column).
#include "common.cuh"

using namespace mpasb200;

namespace {

struct RecCellP {
    KMagic km;
    int L, cellStart, nCellsLoop, rk_step, nCells;
    real dt, invNs;
    const real *rw_save, *rw_p, *rtheta_pp, *rtheta_p_save, *rt_diabatic_tend, *rho_p_save, *rho_pp, *rho_base;
    const real *exner_base, *rtheta_base, *zz, *fzm, *fzp;
    real *wwAvg, *w, *rw, *rtheta_p, *rho_p, *rho_zz, *exner, *pressure_p, *theta_m;
};

__global__ void __launch_bounds__(256, 8) recover_cell_kernel(RecCellP p) {
    const int L = p.L, L1 = L + 1;
    FLAT_KC(p.nCellsLoop, L1);
    const int iCell = p.cellStart + _c0;
    const real rcv = MPAS_RGAS / (MPAS_CP - MPAS_RGAS);
    const real p0 = 1.0e+05;
    const size_t iw = IX2(k, iCell, L1);
    if (k <= L) {
        const size_t ic = IX2(k, iCell, L);
        const real rho_p = p.rho_p_save[ic] + p.rho_pp[ic];
        const real rho_zz = rho_p + p.rho_base[ic];
        p.rho_p[ic] = rho_p;
        p.rho_zz[ic] = rho_zz;
        real rtheta_p;
        if (p.rk_step == 3) {
            rtheta_p = p.rtheta_p_save[ic] + p.rtheta_pp[ic] - p.dt * rho_zz * (p.rt_diabatic_tend ? p.rt_diabatic_tend[ic] : (real)0.0);
        } else {
            rtheta_p = p.rtheta_p_save[ic] + p.rtheta_pp[ic];
        }
        p.rtheta_p[ic] = rtheta_p;
        const real rtb = p.rtheta_base[ic];
        p.theta_m[ic] = (rtheta_p + rtb) / rho_zz;
        if (p.rk_step == 3) {
            const real zzk = p.zz[ic];
            const real ex = pow_rcv(zzk * (MPAS_RGAS / p0) * (rtheta_p + rtb), rcv);
            p.exner[ic] = ex;
            p.pressure_p[ic] = zzk * MPAS_RGAS * (ex * rtheta_p + rtb * (ex - p.exner_base[ic]));
        }
        if (k >= 2) {
            p.wwAvg[iw] = p.rw_save[iw] + (p.wwAvg[iw] * p.invNs);
            const real rw = p.rw_save[iw] + p.rw_p[iw];
            p.rw[iw] = rw;
            p.w[iw] = rw / (p.fzm[k - 1] * p.zz[ic] + p.fzp[k - 1] * p.zz[IX2(k - 1, iCell, L)]);
        } else {
            p.w[iw] = 0.0;
        }
    } else {
        p.w[iw] = 0.0;
    }
}

__global__ void set_garbage_column_kernel(real* a, int L, int col, real v) {  // TI:2989-2991
    const int k = blockIdx.x * blockDim.x + threadIdx.x + 1;
    if (k <= L) a[IX2(k, col, L)] = v;
}

struct RecEdgeP {
    KMagic km;
    int L, edgeStart, nEdgesLoop;
    real invNs;
    const int* cellsOnEdge;
    const real *ru_save, *ru_p, *rho_zz;
    real *ruAvg, *ru, *u;
};

__global__ void __launch_bounds__(256) recover_edge_kernel(RecEdgeP p) {
    const int L = p.L;
    FLAT_KC(p.nEdgesLoop, L);
    const int iEdge = p.edgeStart + _c0;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    const size_t ie = IX2(k, iEdge, L);
    const real rs = p.ru_save[ie];
    p.ruAvg[ie] = rs + (p.ruAvg[ie] * p.invNs);
    const real ru = rs + p.ru_p[ie];
    p.ru[ie] = ru;
    p.u[ie] = (real)2. * ru / (p.rho_zz[IX2(k, cell1, L)] + p.rho_zz[IX2(k, cell2, L)]);
}

struct RecWP {
    KMagic km;
    int L, mE, cellStart, nCellsLoop;
    real cf1, cf2, cf3;
    const int *bdyMaskCell, *nEdgesOnCell, *edgesOnCell;
    const real *edgesOnCell_sign, *zb_cell, *zb3_cell, *ru, *rho_zz, *fzm, *fzp;
    real* w;
};

// Flat (k, cell) mapping; the edge loop is unrolled to NE slots with the loads of ALL edges issued before the first
// dependent add (index loads, then ru / zb / zb3 loads, then the sum in the reference's order), so a thread waits for
// memory twice instead of twice per edge.  Slots beyond nEdgesOnCell repeat edge 1 and are skipped in the sum.
template <int NE>
__global__ void __launch_bounds__(256, NE <= 6 ? 4 : 2) recover_w_kernel(RecWP p) {
    const int L = p.L, L1 = L + 1, mE = p.mE;
    FLAT_KC(p.nCellsLoop, L);
    const int iCell = p.cellStart + _c0;
    if (p.bdyMaskCell[iCell - 1] > MPAS_NRELAXZONE) return;
    const size_t iw = IX2(k, iCell, L1);
    const int ne = p.nEdgesOnCell[iCell - 1];
    const int* eoc = p.edgesOnCell + (size_t)(iCell - 1) * mE;
    const real* sgn = p.edgesOnCell_sign + (size_t)(iCell - 1) * mE;
    const real* zb = p.zb_cell + (size_t)(iCell - 1) * mE * L1 + (k - 1);
    const real* zb3 = p.zb3_cell + (size_t)(iCell - 1) * mE * L1 + (k - 1);
    int e[NE];
#pragma unroll
    for (int i = 0; i < NE; ++i) e[i] = eoc[i < ne ? i : 0];
    real a[NE], a3[NE], r0[NE], r1[NE];
    const int km = k >= 2 ? k - 1 : 2;  // second level of the flux (k = 1 uses levels 1, 2, 3)
#pragma unroll
    for (int i = 0; i < NE; ++i) {
        const int j = i < ne ? i : 0;
        const real* ru = p.ru + (size_t)(e[i] - 1) * L;
        a[i] = zb[(size_t)j * L1];
        a3[i] = zb3[(size_t)j * L1];
        r0[i] = ru[k - 1];
        r1[i] = ru[km - 1];
    }
    real wk = p.w[iw];
    // k = 1: flux = cf1 ru(1) + cf2 ru(2) + cf3 ru(3) (TI:3069-3082); k >= 2: flux = fzm(k) ru(k) + fzp(k) ru(k-1) (TI:3085-3095)
    const bool first = (k == 1);
    const real c0 = first ? p.cf1 : p.fzm[k - 1], c1 = first ? p.cf2 : p.fzp[k - 1];
    const real* rho = p.rho_zz + (size_t)(iCell - 1) * L;
    real den = c0 * rho[k - 1] + c1 * rho[km - 1];
    if (first) den = den + p.cf3 * rho[2];
#pragma unroll
    for (int i = 0; i < NE; ++i) {
        real flux = c0 * r0[i] + c1 * r1[i];
        if (first) flux = flux + p.cf3 * p.ru[(size_t)(e[i] - 1) * L + 2];
        if (i < ne) wk = wk + sgn[i] * (a[i] + fsign((real)1.0, flux) * a3[i]) * flux;
    }
    p.w[iw] = wk / den;
}

struct SmlP {
    int L, mE, cellStart, nCellsLoop;
    const int *bdyMaskCell, *nEdgesOnCell, *edgesOnCell;
    const real *edgesOnCell_sign, *zb_cell, *zb3_cell, *u_tend, *zz, *fzm, *fzp;
    real* w_tend;
};

// TI:2287-2307: omega tendency from the w tendency and the terrain flux of the u tendency
template <int NE>
__global__ void __launch_bounds__(256) smlstep_kernel(SmlP p) {
    const int L = p.L, L1 = L + 1, mE = p.mE;
    FLAT_KC_DIV(p.nCellsLoop, L);
    if (k < 2) return;
    const int iCell = p.cellStart + _c0;
    if (p.bdyMaskCell[iCell - 1] > MPAS_NRELAXZONE) return;
    const size_t iw = IX2(k, iCell, L1);
    const int ne = p.nEdgesOnCell[iCell - 1];
    const int* eoc = p.edgesOnCell + (size_t)(iCell - 1) * mE;
    const real* sgn = p.edgesOnCell_sign + (size_t)(iCell - 1) * mE;
    const real* zb = p.zb_cell + (size_t)(iCell - 1) * mE * L1 + (k - 1);
    const real* zb3 = p.zb3_cell + (size_t)(iCell - 1) * mE * L1 + (k - 1);
    int e[NE];
#pragma unroll
    for (int i = 0; i < NE; ++i) e[i] = eoc[i < ne ? i : 0];
    real s[NE], a[NE], a3[NE], u0[NE], u1[NE];
#pragma unroll
    for (int i = 0; i < NE; ++i) {  // every load of the edge sum is issued before its first use
        const int j = i < ne ? i : 0;
        const real* ut = p.u_tend + (size_t)(e[i] - 1) * L + (k - 1);
        s[i] = sgn[j];
        a[i] = zb[(size_t)j * L1];
        a3[i] = zb3[(size_t)j * L1];
        u0[i] = ut[0];
        u1[i] = ut[-1];
    }
    real wt = p.w_tend[iw];
    const real fm = p.fzm[k - 1], fp = p.fzp[k - 1];
    const real zf = fm * p.zz[IX2(k, iCell, L)] + fp * p.zz[IX2(k - 1, iCell, L)];
#pragma unroll
    for (int i = 0; i < NE; ++i) {
        const real flux = s[i] * (fm * u0[i] + fp * u1[i]);
        if (i < ne) wt = wt - (a[i] + fsign((real)1.0, u0[i]) * a3[i]) * flux;
    }
    p.w_tend[iw] = zf * wt;
}

}  // namespace

extern "C" int mpas_b200_recover_large_step_variables_work(MPAS_RECOVER_ARGS) {
    (void)num_scalars; (void)nEdges; (void)nCellsSolve; (void)nEdgesSolve; (void)pressure_b; (void)scalars; (void)zb; (void)zb3;
    (void)vertexStart; (void)vertexEnd; (void)cellSolveStart; (void)cellSolveEnd; (void)vertexSolveStart; (void)vertexSolveEnd;
    (void)edgeSolveStart; (void)edgeSolveEnd;
    const int L = nVertLevels;
    real* d_rho_zz = X(rho_zz);
    set_garbage_column_kernel<<<cdiv(L, 128), 128, 0, stream()>>>(d_rho_zz, L, nCells + 1, (real)1.0);
    count_launch();
    const real invNs = 1 / (real)ns;
    {
        RecCellP p;
        p.L = L; p.km.set(L, L + 1); p.cellStart = cellStart; p.nCellsLoop = cellEnd - cellStart + 1; p.rk_step = rk_step; p.nCells = nCells;
        p.dt = dt; p.invNs = invNs;
        p.rw_save = X(rw_save); p.rw_p = X(rw_p); p.rtheta_pp = X(rtheta_pp); p.rtheta_p_save = X(rtheta_p_save);
        p.rt_diabatic_tend = X(rt_diabatic_tend); p.rho_p_save = X(rho_p_save); p.rho_pp = X(rho_pp); p.rho_base = X(rho_base);
        p.exner_base = X(exner_base); p.rtheta_base = X(rtheta_base); p.zz = X(zz); p.fzm = X(fzm); p.fzp = X(fzp);
        p.wwAvg = X(wwAvg); p.w = X(w); p.rw = X(rw); p.rtheta_p = X(rtheta_p); p.rho_p = X(rho_p); p.rho_zz = d_rho_zz;
        p.exner = X(exner); p.pressure_p = X(pressure_p); p.theta_m = X(theta_m);
        KLAUNCH(recover_cell_kernel, (long long)p.nCellsLoop * (L + 1), 256, p);
        KCHECK("recover_cell_kernel");
    }
    {
        RecEdgeP p;
        p.L = L; p.km.set(L, L + 1); p.edgeStart = edgeStart; p.nEdgesLoop = edgeEnd - edgeStart + 1; p.invNs = invNs;
        p.cellsOnEdge = X(cellsOnEdge); p.ru_save = X(ru_save); p.ru_p = X(ru_p); p.rho_zz = d_rho_zz;
        p.ruAvg = X(ruAvg); p.ru = X(ru); p.u = X(u);
        KLAUNCH(recover_edge_kernel, (long long)p.nEdgesLoop * L, 256, p);
        KCHECK("recover_edge_kernel");
    }
    {
        RecWP p;
        p.L = L; p.km.set(L, L + 1); p.mE = maxEdges; p.cellStart = cellStart; p.nCellsLoop = cellEnd - cellStart + 1;
        p.cf1 = cf1; p.cf2 = cf2; p.cf3 = cf3;
        p.bdyMaskCell = X(bdyMaskCell); p.nEdgesOnCell = X(nEdgesOnCell); p.edgesOnCell = X(edgesOnCell);
        p.edgesOnCell_sign = X(edgesOnCell_sign); p.zb_cell = X(zb_cell); p.zb3_cell = X(zb3_cell); p.ru = X(ru); p.rho_zz = d_rho_zz;
        p.fzm = X(fzm); p.fzp = X(fzp); p.w = X(w);
        if (p.mE <= 6) KLAUNCH(recover_w_kernel<6>, (long long)p.nCellsLoop * L, 256, p);
        else if (p.mE <= 8) KLAUNCH(recover_w_kernel<8>, (long long)p.nCellsLoop * L, 256, p);
        else KLAUNCH(recover_w_kernel<12>, (long long)p.nCellsLoop * L, 256, p);
        KCHECK("recover_w_kernel");
    }
    return 0;
}

extern "C" int mpas_b200_set_smlstep_pert_variables_work(MPAS_SMLSTEP_ARGS) {
    (void)nCells; (void)nEdges; (void)nCellsSolve; (void)cellsOnEdge; (void)cellStart; (void)cellEnd; (void)edgeStart; (void)edgeEnd;
    (void)edgeSolveStart; (void)edgeSolveEnd;
    SmlP p;
    p.L = nVertLevels; p.mE = maxEdges; p.cellStart = cellSolveStart; p.nCellsLoop = cellSolveEnd - cellSolveStart + 1;
    p.bdyMaskCell = X(bdyMaskCell); p.nEdgesOnCell = X(nEdgesOnCell); p.edgesOnCell = X(edgesOnCell);
    p.edgesOnCell_sign = X(edgesOnCell_sign); p.zb_cell = X(zb_cell); p.zb3_cell = X(zb3_cell); p.u_tend = X(u_tend); p.zz = X(zz);
    p.fzm = X(fzm); p.fzp = X(fzp); p.w_tend = X(w_tend);
    if (p.mE <= 6) KLAUNCH(smlstep_kernel<6>, (long long)p.nCellsLoop * p.L, 256, p);
    else if (p.mE <= 8) KLAUNCH(smlstep_kernel<8>, (long long)p.nCellsLoop * p.L, 256, p);
    else KLAUNCH(smlstep_kernel<12>, (long long)p.nCellsLoop * p.L, 256, p);
    KCHECK("smlstep_kernel");
    return 0;
}
