// atm_advance_acoustic_step_work (TI:2447-2723) and atm_divergence_damping_3d (TI:2726-2795)
// TI = src/core_atmosphere/dynamics/mpas_atm_time_integration.F of the reference.
//
// Kernels
//   acoustic_edge_kernel   one thread per (k, edge), k fastest: forward step of ru_p + ruAvg accumulation
//                          (TI:2552-2599).  HBM-bound, 9 edge/cell words per element.
//   acoustic_cell_kernel   one warp per column, lanes along k: flux divergence rs/ts over edgesOnCell
//                          (TI:2629-2639), explicit parts (TI:2646-2670), the tridiagonal sweep
//                          (TI:2675-2682) run by lane 0 out of shared memory in the reference's order
//                          (bit-exact), Rayleigh damping, wwAvg, rho_pp / rtheta_pp update
//                          (TI:2687-2708).  Also performs the rtheta_pp_old <- rtheta_pp copy of
//                          TI:2603-2611 for its own column (halo columns only do that copy).
//   divdamp_kernel         one thread per (k, edge) (TI:2768-2793).
#include "common.cuh"

using namespace mpasb200;

namespace {

struct AcEdgeP {
    int L, nCellsSolve, edgeStart, nEdgesLoop, small_step;
    real dts, c2;
    const int* cellsOnEdge;
    const real *rtheta_pp, *rho_pp, *zz, *exner, *cqu, *zxu, *tend_ru, *invDcEdge, *specZoneMaskEdge;
    real *ru_p, *ruAvg;
};

__global__ void __launch_bounds__(256) acoustic_edge_kernel(AcEdgeP p) {
    FLAT_KC(p.nEdgesLoop, p.L);
    const int L = p.L;
    const int iEdge = p.edgeStart + _c0;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    if (!(cell1 <= p.nCellsSolve || cell2 <= p.nCellsSolve)) return;
    const size_t ie = IX2(k, iEdge, L);
    if (p.small_step != 1) {
        const size_t i1 = IX2(k, cell1, L), i2 = IX2(k, cell2, L);
        real pgrad = ((p.rtheta_pp[i2] - p.rtheta_pp[i1]) * p.invDcEdge[iEdge - 1]) / ((real).5 * (p.zz[i2] + p.zz[i1]));
        pgrad = p.cqu[ie] * (real)0.5 * p.c2 * (p.exner[i1] + p.exner[i2]) * pgrad;
        pgrad = pgrad + (real)0.5 * p.zxu[ie] * MPAS_GRAVITY * (p.rho_pp[i1] + p.rho_pp[i2]);
        const real r = p.ru_p[ie] + p.dts * (p.tend_ru[ie] - ((real)1.0 - p.specZoneMaskEdge[iEdge - 1]) * pgrad);
        p.ru_p[ie] = r;
        p.ruAvg[ie] = p.ruAvg[ie] + r;
    } else {
        const real r = p.dts * p.tend_ru[ie];
        p.ru_p[ie] = r;
        p.ruAvg[ie] = r;
    }
}

struct AcCellP {
    int L, mE, cellSolveStart, cellSolveEnd, cellStart, cellEnd, small_step;
    real dts, epssm, resm;
    const int *nEdgesOnCell, *cellsOnEdge, *edgesOnCell;
    const real *edgesOnCell_sign, *dvEdge, *invAreaCell, *specZoneMaskCell;
    const real *rho_zz, *theta_m, *zz, *cofwt, *coftz, *a_tri, *alpha_tri, *gamma_tri, *dss;
    const real *tend_rho, *tend_rt, *tend_rw, *cofwr, *cofwz, *w, *rw, *rw_save;
    const real *fzm, *fzp, *rdzw, *cofrz, *ru_p;
    real *rw_p, *rtheta_pp, *rtheta_pp_old, *wwAvg, *rho_pp;
};

// shared memory per warp: ts[L+1] rs[L+1] rwp[L+2] a[L+1] al[L+1] ga[L+1]  (1-based slots)
__global__ void __launch_bounds__(256) acoustic_cell_kernel(AcCellP p) {
    extern __shared__ real smem[];
    const int L = p.L, L1 = L + 1, mE = p.mE;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    real* ts = smem + (size_t)warp * (6 * (L + 2));
    real* rs = ts + (L + 2);
    real* rwp = rs + (L + 2);
    real* sa = rwp + (L + 2);
    real* sal = sa + (L + 2);
    real* sga = sal + (L + 2);
    const real dts = p.dts, epssm = p.epssm, resm = p.resm;
    const bool first = p.small_step == 1;

    for (int iCell = p.cellStart + blockIdx.x * wpb + warp; iCell <= p.cellEnd; iCell += gridDim.x * wpb) {
        // TI:2603-2611 (all cells)
        for (int k = lane + 1; k <= L; k += 32) {
            const size_t ic = IX2(k, iCell, L);
            p.rtheta_pp_old[ic] = first ? (real)0.0 : p.rtheta_pp[ic];
        }
        if (iCell > p.cellSolveEnd || iCell < p.cellSolveStart) continue;  // halo column: copy only

        if (p.specZoneMaskCell[iCell - 1] == (real)0.0) {
            const real invArea = p.invAreaCell[iCell - 1];
            // stage the old rw_p column and the Thomas coefficients
            for (int k = lane + 1; k <= L1; k += 32) rwp[k] = first ? (real)0.0 : p.rw_p[IX2(k, iCell, L1)];
            for (int k = lane + 1; k <= L; k += 32) {
                const size_t ic = IX2(k, iCell, L);
                sa[k] = p.a_tri[ic];
                sal[k] = p.alpha_tri[ic];
                sga[k] = p.gamma_tri[ic];
            }
            __syncwarp();
            const int ne = p.nEdgesOnCell[iCell - 1];
            for (int k = lane + 1; k <= L; k += 32) {
                real tsk = 0.0, rsk = 0.0;
                for (int i = 1; i <= ne; ++i) {
                    const int iEdge = p.edgesOnCell[IX2(i, iCell, mE)];
                    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
                    const real flux = p.edgesOnCell_sign[IX2(i, iCell, mE)] * dts * p.dvEdge[iEdge - 1] * p.ru_p[IX2(k, iEdge, L)] * invArea;
                    rsk = rsk - flux;
                    tsk = tsk - flux * (real)0.5 * (p.theta_m[IX2(k, cell2, L)] + p.theta_m[IX2(k, cell1, L)]);
                }
                const size_t ic = IX2(k, iCell, L);
                const real rho_pp_k = first ? (real)0.0 : p.rho_pp[ic];
                const real rth_pp_k = first ? (real)0.0 : p.rtheta_pp[ic];
                rsk = rho_pp_k + dts * p.tend_rho[ic] + rsk - p.cofrz[k - 1] * resm * (rwp[k + 1] - rwp[k]);
                tsk = rth_pp_k + dts * p.tend_rt[ic] + tsk -
                      resm * p.rdzw[k - 1] * (p.coftz[IX2(k + 1, iCell, L1)] * rwp[k + 1] - p.coftz[IX2(k, iCell, L1)] * rwp[k]);
                rs[k] = rsk;
                ts[k] = tsk;
            }
            __syncwarp();
            // explicit part of rw_p, TI:2659-2670 (reads only rwp[k] of its own level)
            for (int k = lane + 1; k <= L; k += 32) {
                if (k < 2) continue;
                const size_t ic = IX2(k, iCell, L), icm = IX2(k - 1, iCell, L);
                const real rtpp_k = first ? (real)0.0 : p.rtheta_pp[ic], rtpp_km = first ? (real)0.0 : p.rtheta_pp[icm];
                const real rpp_k = first ? (real)0.0 : p.rho_pp[ic], rpp_km = first ? (real)0.0 : p.rho_pp[icm];
                const real zzk = p.zz[ic], zzkm = p.zz[icm];
                rwp[k] = rwp[k] + dts * p.tend_rw[IX2(k, iCell, L1)] -
                         p.cofwz[ic] * ((zzk * ts[k] - zzkm * ts[k - 1]) + resm * (zzk * rtpp_k - zzkm * rtpp_km)) -
                         p.cofwr[ic] * ((rs[k] + rs[k - 1]) + resm * (rpp_k + rpp_km)) + p.cofwt[ic] * (ts[k] + resm * rtpp_k) +
                         p.cofwt[icm] * (ts[k - 1] + resm * rtpp_km);
            }
            // first half of the wwAvg accumulation uses the OLD rw_p (TI:2655-2657): keep it in a register
            // by recomputing from global below; here we only need ordering, so do it before rwp is overwritten.
            // (old rw_p is still in global memory: rwp[] in shared memory is the working copy)
            __syncwarp();
            if (lane == 0) {  // TI:2675-2682, reference order
                real prev = rwp[1];
                for (int k = 2; k <= L; ++k) {
                    prev = (rwp[k] - sa[k] * prev) * sal[k];
                    rwp[k] = prev;
                }
                real next = rwp[L1];
                for (int k = L; k >= 1; --k) {
                    next = rwp[k] - sga[k] * next;
                    rwp[k] = next;
                }
            }
            __syncwarp();
            for (int k = lane + 1; k <= L; k += 32) {
                if (k < 2) continue;
                const size_t ic = IX2(k, iCell, L), icm = IX2(k - 1, iCell, L), iw = IX2(k, iCell, L1);
                const real rw_old = first ? (real)0.0 : p.rw_p[iw];
                real ww = first ? (real)0.0 : p.wwAvg[iw];
                ww = ww + (real)0.5 * ((real)1.0 - epssm) * rw_old;
                const real d = p.rw_save[iw] - p.rw[iw];
                const real dk = p.dss[ic];
                const real r = (rwp[k] + d -
                                dts * dk * (p.fzm[k - 1] * p.zz[ic] + p.fzp[k - 1] * p.zz[icm]) *
                                    (p.fzm[k - 1] * p.rho_zz[ic] + p.fzp[k - 1] * p.rho_zz[icm]) * p.w[iw]) /
                                   ((real)1.0 + dts * dk) -
                               d;
                rwp[k] = r;
                ww = ww + (real)0.5 * ((real)1.0 + epssm) * r;
                p.wwAvg[iw] = ww;
            }
            __syncwarp();
            for (int k = lane + 1; k <= L; k += 32) {
                const size_t ic = IX2(k, iCell, L);
                p.rw_p[IX2(k, iCell, L1)] = rwp[k];
                p.rho_pp[ic] = rs[k] - p.cofrz[k - 1] * (rwp[k + 1] - rwp[k]);
                p.rtheta_pp[ic] = ts[k] - p.rdzw[k - 1] * (p.coftz[IX2(k + 1, iCell, L1)] * rwp[k + 1] - p.coftz[IX2(k, iCell, L1)] * rwp[k]);
            }
            if (first) {  // TI:2617-2622: the levels the loops above never write
                if (lane == 0) {
                    p.wwAvg[IX2(1, iCell, L1)] = 0.0;
                    p.wwAvg[IX2(L1, iCell, L1)] = 0.0;
                    p.rw_p[IX2(L1, iCell, L1)] = 0.0;
                }
            }
            __syncwarp();
        } else {  // specified zone of a regional run, TI:2710-2719
            for (int k = lane + 1; k <= L; k += 32) {
                const size_t ic = IX2(k, iCell, L), iw = IX2(k, iCell, L1);
                const real rho0 = first ? (real)0.0 : p.rho_pp[ic], rt0 = first ? (real)0.0 : p.rtheta_pp[ic];
                const real rw0 = first ? (real)0.0 : p.rw_p[iw], ww0 = first ? (real)0.0 : p.wwAvg[iw];
                p.rho_pp[ic] = rho0 + dts * p.tend_rho[ic];
                p.rtheta_pp[ic] = rt0 + dts * p.tend_rt[ic];
                const real r = rw0 + dts * p.tend_rw[iw];
                p.rw_p[iw] = r;
                p.wwAvg[iw] = ww0 + (real)0.5 * ((real)1.0 + epssm) * r;
            }
            if (first && lane == 0) {
                p.wwAvg[IX2(L1, iCell, L1)] = 0.0;
                p.rw_p[IX2(L1, iCell, L1)] = 0.0;
            }
        }
    }
}

struct DivDampP {
    int L, nCellsSolve, edgeStart, nEdgesLoop;
    real coef_divdamp;
    const int* cellsOnEdge;
    const real *theta_m, *rtheta_pp, *rtheta_pp_old, *specZoneMaskEdge;
    real* ru_p;
};

__global__ void __launch_bounds__(256) divdamp_kernel(DivDampP p) {
    FLAT_KC(p.nEdgesLoop, p.L);
    const int L = p.L;
    const int iEdge = p.edgeStart + _c0;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    if (!(cell1 <= p.nCellsSolve || cell2 <= p.nCellsSolve)) return;
    const size_t i1 = IX2(k, cell1, L), i2 = IX2(k, cell2, L), ie = IX2(k, iEdge, L);
    const real divCell1 = -(p.rtheta_pp[i1] - p.rtheta_pp_old[i1]);
    const real divCell2 = -(p.rtheta_pp[i2] - p.rtheta_pp_old[i2]);
    p.ru_p[ie] = p.ru_p[ie] + p.coef_divdamp * (divCell2 - divCell1) * ((real)1.0 - p.specZoneMaskEdge[iEdge - 1]) / (p.theta_m[i1] + p.theta_m[i2]);
}

}  // namespace

extern "C" int mpas_b200_advance_acoustic_step_work(MPAS_ACOUSTIC_ARGS) {
    (void)nCells; (void)nEdges; (void)vertexStart; (void)vertexEnd; (void)vertexSolveStart; (void)vertexSolveEnd;
    (void)edgeSolveStart; (void)edgeSolveEnd; (void)zgrid; (void)ru; (void)ru_save; (void)dcEdge; (void)cf1; (void)cf2; (void)cf3;
    const real rcv = MPAS_RGAS / (MPAS_CP - MPAS_RGAS);
    const real c2 = MPAS_CP * rcv;
    const real resm = ((real)1.0 - epssm) / ((real)1.0 + epssm);
    const int L = nVertLevels;
    {
        AcEdgeP p;
        p.L = L; p.nCellsSolve = nCellsSolve; p.edgeStart = edgeStart; p.nEdgesLoop = edgeEnd - edgeStart + 1; p.small_step = small_step;
        p.dts = dts; p.c2 = c2;
        p.cellsOnEdge = X(cellsOnEdge);
        p.rtheta_pp = X(rtheta_pp); p.rho_pp = X(rho_pp); p.zz = X(zz); p.exner = X(exner); p.cqu = X(cqu); p.zxu = X(zxu);
        p.tend_ru = X(tend_ru); p.invDcEdge = X(invDcEdge); p.specZoneMaskEdge = X(specZoneMaskEdge);
        p.ru_p = X(ru_p); p.ruAvg = X(ruAvg);
        KLAUNCH(acoustic_edge_kernel, (long long)p.nEdgesLoop * L, 256, p);
        KCHECK("acoustic_edge_kernel");
    }
    {
        AcCellP p;
        p.L = L; p.mE = maxEdges; p.cellSolveStart = cellSolveStart; p.cellSolveEnd = cellSolveEnd; p.cellStart = cellStart; p.cellEnd = cellEnd; p.small_step = small_step;
        p.dts = dts; p.epssm = epssm; p.resm = resm;
        p.nEdgesOnCell = X(nEdgesOnCell); p.cellsOnEdge = X(cellsOnEdge); p.edgesOnCell = X(edgesOnCell);
        p.edgesOnCell_sign = X(edgesOnCell_sign); p.dvEdge = X(dvEdge); p.invAreaCell = X(invAreaCell); p.specZoneMaskCell = X(specZoneMaskCell);
        p.rho_zz = X(rho_zz); p.theta_m = X(theta_m); p.zz = X(zz); p.cofwt = X(cofwt); p.coftz = X(coftz);
        p.a_tri = X(a_tri); p.alpha_tri = X(alpha_tri); p.gamma_tri = X(gamma_tri); p.dss = X(dss);
        p.tend_rho = X(tend_rho); p.tend_rt = X(tend_rt); p.tend_rw = X(tend_rw); p.cofwr = X(cofwr); p.cofwz = X(cofwz);
        p.w = X(w); p.rw = X(rw); p.rw_save = X(rw_save); p.fzm = X(fzm); p.fzp = X(fzp); p.rdzw = X(rdzw); p.cofrz = X(cofrz);
        p.ru_p = X(ru_p);
        p.rw_p = X(rw_p); p.rtheta_pp = X(rtheta_pp); p.rtheta_pp_old = X(rtheta_pp_old); p.wwAvg = X(wwAvg); p.rho_pp = X(rho_pp);
        const int ncols = cellEnd - cellStart + 1;
        if (ncols > 0) {
            const int wpb = 8;
            const size_t shm = (size_t)wpb * 6 * (L + 2) * sizeof(real);
            static bool attr_set = false;
            if (!attr_set) {
                cudaFuncSetAttribute(acoustic_cell_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
                attr_set = true;
            }
            if (shm > 200 * 1024) return fail_msg("acoustic_cell_kernel: nVertLevels too large for the shared-memory staging");
            int nblocks = cdiv(ncols, wpb);
            ktime_begin("acoustic_cell_kernel");
            acoustic_cell_kernel<<<nblocks, wpb * 32, shm, stream()>>>(p);
            ktime_end();
            count_launch();
            KCHECK("acoustic_cell_kernel");
        }
    }
    return 0;
}

extern "C" int mpas_b200_divergence_damping_3d(MPAS_DIVDAMP_ARGS) {
    (void)nCells; (void)nEdges;
    DivDampP p;
    const real rdts = (real)1.0 / dts;
    p.coef_divdamp = (real)2.0 * smdiv * config_len_disp * rdts;
    p.L = nVertLevels; p.nCellsSolve = nCellsSolve; p.edgeStart = edgeStart; p.nEdgesLoop = edgeEnd - edgeStart + 1;
    p.cellsOnEdge = X(cellsOnEdge); p.theta_m = X(theta_m); p.rtheta_pp = X(rtheta_pp); p.rtheta_pp_old = X(rtheta_pp_old);
    p.specZoneMaskEdge = X(specZoneMaskEdge); p.ru_p = X(ru_p);
    KLAUNCH(divdamp_kernel, (long long)p.nEdgesLoop * p.L, 256, p);
    KCHECK("divdamp_kernel");
    return 0;
}
