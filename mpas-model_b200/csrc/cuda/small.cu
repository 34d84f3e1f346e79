// The short routines atm_srk3 calls between the big ones (SURVEY.md section 8a "adjacent"):
//   atm_rk_integration_setup            TI:1799-1859   state copies (device-to-device)
//   atm_compute_moist_coefficients      TI:1862-1933   qtot, cqw, cqu
//   atm_compute_vert_imp_coefs_work     TI:2012-2131   implicit-solve coefficients + Thomas pre-factorisation
//   atm_rk_dynamics_substep_finish      TI:5993-6081   level copies and split-step averages
//   atm_init_coupled_diagnostics        TI:5825-5990   start-up coupling of the prognostic state
// TI = src/core_atmosphere/dynamics/mpas_atm_time_integration.F.
#include "common.cuh"

using namespace mpasb200;

namespace {

int d2d(void* dst, const void* src, size_t bytes) {
    if (bytes == 0) return 0;
    cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, stream());
    if (e != cudaSuccess) return fail("cudaMemcpyAsync(D2D)", e);
    count_launch();
    return 0;
}
#define D2D_COLS(dst, src, ld, cStart, cEnd)                                                                       \
    do {                                                                                                           \
        int _rc = d2d(X(dst) + IX2(1, cStart, ld), X(src) + IX2(1, cStart, ld), (size_t)((cEnd) - (cStart) + 1) * (ld) * sizeof(real)); \
        if (_rc) return _rc;                                                                                       \
    } while (0)

struct MoistP {
    int L, ns, ms, me, cellStart, nCellsLoop, edgeStart, nEdgesLoop, nCellsSolve;
    const int* cellsOnEdge;
    const real* scalars;
    real *cqw, *cqu, *qtot;
};

__global__ void __launch_bounds__(256) moist_cell_kernel(MoistP p) {
    const int L = p.L, ns = p.ns;
    FLAT_KC(p.nCellsLoop, L);
    const int iCell = p.cellStart + _c0;
    real q = 0.0;
    for (int iq = p.ms; iq <= p.me; ++iq) q = q + p.scalars[IX3(iq, k, iCell, ns, L)];
    p.qtot[IX2(k, iCell, L)] = q;
    if (k >= 2) {
        real qm = 0.0;
        for (int iq = p.ms; iq <= p.me; ++iq) qm = qm + p.scalars[IX3(iq, k - 1, iCell, ns, L)];
        const real qtotal = (real)0.5 * (q + qm);
        p.cqw[IX2(k, iCell, L)] = (real)1.0 / ((real)1.0 + qtotal);
    }
}

__global__ void __launch_bounds__(256) moist_edge_kernel(MoistP p) {
    const int L = p.L, ns = p.ns;
    FLAT_KC(p.nEdgesLoop, L);
    const int iEdge = p.edgeStart + _c0;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    if (!(cell1 <= p.nCellsSolve || cell2 <= p.nCellsSolve)) return;
    real qtotal = 0.0;
    for (int iq = p.ms; iq <= p.me; ++iq) qtotal = qtotal + (real)0.5 * (p.scalars[IX3(iq, k, cell1, ns, L)] + p.scalars[IX3(iq, k, cell2, ns, L)]);
    p.cqu[IX2(k, iEdge, L)] = (real)1.0 / ((real)1.0 + qtotal);
}

struct VicP {
    int L, cellSolveStart, nCellsLoop;
    real dtseps, rcv, c2;
    const real *zz, *cqw, *p_, *t, *rb, *rtb, *pb, *rt, *rdzw, *fzm, *fzp, *rdzu, *qtot;
    real *cofwr, *cofwz, *coftz, *cofwt, *a_tri, *alpha_tri, *gamma_tri, *cofrz;
};

__global__ void cofrz_kernel(VicP p) {  // TI:2070-2072
    const int k = blockIdx.x * blockDim.x + threadIdx.x + 1;
    if (k <= p.L) p.cofrz[k - 1] = p.dtseps * p.rdzw[k - 1];
}

// one warp per column; lanes along k for the coefficient arrays, lane 0 runs the recurrence
// alpha(k) = 1/(b(k) - a(k)*gamma(k-1)), gamma(k) = c(k)*alpha(k) (TI:2124-2127) in the reference's order.
// shared memory per warp: b_tri[L+2] c_tri[L+2] a[L+2] al[L+2] ga[L+2]
__global__ void __launch_bounds__(256) vert_imp_kernel(VicP p) {
    extern __shared__ real smem[];
    const int L = p.L, L1 = L + 1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    real* b_tri = smem + (size_t)warp * 5 * (L + 2);
    real* c_tri = b_tri + (L + 2);
    real* sa = c_tri + (L + 2);
    real* sal = sa + (L + 2);
    real* sga = sal + (L + 2);
    const real dtseps = p.dtseps, rcv = p.rcv, c2 = p.c2;
    for (int iCell = p.cellSolveStart + blockIdx.x * wpb + warp; iCell < p.cellSolveStart + p.nCellsLoop; iCell += gridDim.x * wpb) {
        // pass 1: cofwr, cofwz, coftz, cofwt
        for (int k = lane + 1; k <= L1; k += 32) {
            if (k == 1 || k == L1) {
                p.coftz[IX2(k, iCell, L1)] = 0.0;
            }
            if (k >= 2 && k <= L) {
                const size_t ic = IX2(k, iCell, L), icm = IX2(k - 1, iCell, L);
                const real fm = p.fzm[k - 1], fp = p.fzp[k - 1];
                const real zzk = p.zz[ic], zzkm = p.zz[icm];
                p.cofwr[ic] = (real).5 * dtseps * MPAS_GRAVITY * (fm * zzk + fp * zzkm);
                p.cofwz[ic] = dtseps * c2 * (fm * zzk + fp * zzkm) * p.rdzu[k - 1] * p.cqw[ic] * (fm * p.p_[ic] + fp * p.p_[icm]);
                p.coftz[IX2(k, iCell, L1)] = dtseps * (fm * p.t[ic] + fp * p.t[icm]);
            }
            if (k <= L) {
                const size_t ic = IX2(k, iCell, L);
                const real qtotal = p.qtot[ic];
                p.cofwt[ic] = (real).5 * dtseps * rcv * p.zz[ic] * MPAS_GRAVITY * p.rb[ic] / ((real)1. + qtotal) * p.p_[ic] / ((p.rtb[ic] + p.rt[ic]) * p.pb[ic]);
            }
        }
        __syncwarp();  // the pass below re-reads this warp's own global writes (same thread block: visible after the barrier)
        __threadfence_block();
        for (int k = lane + 1; k <= L; k += 32) {
            if (k == 1) {
                sa[1] = 0.;
                b_tri[1] = 1.;
                c_tri[1] = 0.;
            } else {
                const size_t ic = IX2(k, iCell, L), icm = IX2(k - 1, iCell, L);
                const real cofwz = p.cofwz[ic], cofwr = p.cofwr[ic], cofwt = p.cofwt[ic], cofwtm = p.cofwt[icm];
                const real coftz_k = p.coftz[IX2(k, iCell, L1)], coftz_km = p.coftz[IX2(k - 1, iCell, L1)], coftz_kp = p.coftz[IX2(k + 1, iCell, L1)];
                const real zzk = p.zz[ic], zzkm = p.zz[icm];
                const real rdzw_k = p.rdzw[k - 1], rdzw_km = p.rdzw[k - 2];
                const real cofrz_k = dtseps * rdzw_k, cofrz_km = dtseps * rdzw_km;
                sa[k] = -cofwz * coftz_km * rdzw_km * zzkm + cofwr * cofrz_km - cofwtm * coftz_km * rdzw_km;
                b_tri[k] = (real)1. + cofwz * (coftz_k * rdzw_k * zzk + coftz_k * rdzw_km * zzkm) - coftz_k * (cofwt * rdzw_k - cofwtm * rdzw_km) +
                           cofwr * (cofrz_k - cofrz_km);
                c_tri[k] = -cofwz * coftz_kp * rdzw_k * zzk - cofwr * cofrz_k + cofwt * coftz_kp * rdzw_k;
            }
        }
        __syncwarp();
        if (lane == 0) {
            sga[1] = 0.;
            sal[1] = 0.;
            real g = 0.;
            for (int k = 2; k <= L; ++k) {
                const real al = (real)1. / (b_tri[k] - sa[k] * g);
                g = c_tri[k] * al;
                sal[k] = al;
                sga[k] = g;
            }
        }
        __syncwarp();
        for (int k = lane + 1; k <= L; k += 32) {
            const size_t ic = IX2(k, iCell, L);
            p.a_tri[ic] = sa[k];
            p.alpha_tri[ic] = sal[k];
            p.gamma_tri[ic] = sga[k];
        }
        __syncwarp();
    }
}

struct AxP {
    long long n;
    real s;
    const real *a, *b;
    real* out;
};
// out = a + b  (mode 0)   or   out = a * s  (mode 1), over a contiguous column range
template <int MODE>
__global__ void __launch_bounds__(256) ax_kernel(AxP p) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= p.n) return;
    if (MODE == 0) p.out[t] = p.a[t] + p.b[t];
    else p.out[t] = p.a[t] * p.s;
}

struct IcdP {
    int L, mE, ns, index_qv, cellStart, nCellsLoop, edgeStart, nEdgesLoop;
    const int *cellsOnEdge, *nEdgesOnCell, *edgesOnCell;
    const real *edgesOnCell_sign, *theta, *rho, *rho_base, *theta_base, *zz, *scalars, *u, *w, *fzm, *fzp, *zb_cell, *zb3_cell;
    real *theta_m, *rho_zz, *rho_p, *rtheta_base, *rtheta_p, *ru, *rw, *pressure_p, *pressure_base, *exner, *exner_base;
};

__global__ void __launch_bounds__(256) icd_cell1_kernel(IcdP p) {  // TI:5909-5917, 5954-5990
    const int L = p.L;
    FLAT_KC(p.nCellsLoop, L);
    const int iCell = p.cellStart + _c0;
    const size_t ic = IX2(k, iCell, L);
    const real rcv = MPAS_RGAS / (MPAS_CP - MPAS_RGAS);
    const real p0 = 1.e5;
    const real theta_m = p.theta[ic] * ((real)1. + MPAS_RVORD * p.scalars[IX3(p.index_qv, k, iCell, p.ns, L)]);
    const real zzk = p.zz[ic];
    const real rho_zz = p.rho[ic] / zzk;
    p.theta_m[ic] = theta_m;
    p.rho_zz[ic] = rho_zz;
    const real rb = p.rho_base[ic], tb = p.theta_base[ic];
    const real rho_p = rho_zz - rb;
    const real rtb = tb * rb;
    const real rtp = theta_m * rho_p + rb * (theta_m - tb);
    p.rho_p[ic] = rho_p;
    p.rtheta_base[ic] = rtb;
    p.rtheta_p[ic] = rtp;
    const real ex = pow_rcv(zzk * (MPAS_RGAS / p0) * (rtp + rtb), rcv);
    const real exb = pow_rcv(zzk * (MPAS_RGAS / p0) * (rtb), rcv);
    p.exner[ic] = ex;
    p.exner_base[ic] = exb;
    p.pressure_p[ic] = zzk * MPAS_RGAS * (ex * rtp + rtb * (ex - exb));
    p.pressure_base[ic] = zzk * MPAS_RGAS * exb * rtb;
}

__global__ void __launch_bounds__(256) icd_edge_kernel(IcdP p) {  // TI:5919-5927
    const int L = p.L;
    FLAT_KC(p.nEdgesLoop, L);
    const int iEdge = p.edgeStart + _c0;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    const size_t ie = IX2(k, iEdge, L);
    p.ru[ie] = (real)0.5 * p.u[ie] * (p.rho_zz[IX2(k, cell1, L)] + p.rho_zz[IX2(k, cell2, L)]);
}

__global__ void __launch_bounds__(256) icd_cell2_kernel(IcdP p) {  // TI:5931-5952
    const int L = p.L, L1 = L + 1, mE = p.mE;
    FLAT_KC(p.nCellsLoop, L1);
    const int iCell = p.cellStart + _c0;
    const size_t iw = IX2(k, iCell, L1);
    if (k == 1 || k == L1) {
        p.rw[iw] = 0.0;
        return;
    }
    const real fm = p.fzm[k - 1], fp = p.fzp[k - 1];
    const size_t ic = IX2(k, iCell, L), icm = IX2(k - 1, iCell, L);
    const real zint = (fp * p.zz[icm] + fm * p.zz[ic]);
    real rw = p.w[iw] * (fp * p.rho_zz[icm] + fm * p.rho_zz[ic]) * zint;
    const int ne = p.nEdgesOnCell[iCell - 1];
    for (int i = 1; i <= ne; ++i) {
        const int iEdge = p.edgesOnCell[IX2(i, iCell, mE)];
        const real flux = (fm * p.ru[IX2(k, iEdge, L)] + fp * p.ru[IX2(k - 1, iEdge, L)]);
        rw = rw - p.edgesOnCell_sign[IX2(i, iCell, mE)] * (p.zb_cell[IX3(k, i, iCell, L1, mE)] + fsign((real)1.0, flux) * p.zb3_cell[IX3(k, i, iCell, L1, mE)]) * flux * zint;
    }
    p.rw[iw] = rw;
}

}  // namespace

extern "C" int mpas_b200_rk_integration_setup(MPAS_RKSETUP_ARGS) {
    (void)nCells; (void)nEdges; (void)vertexStart; (void)vertexEnd; (void)cellSolveStart; (void)cellSolveEnd; (void)vertexSolveStart; (void)vertexSolveEnd;
    (void)edgeSolveStart; (void)edgeSolveEnd;
    const int L = nVertLevels, L1 = L + 1;
    D2D_COLS(ru_save, ru, L, edgeStart, edgeEnd);
    D2D_COLS(u_2, u_1, L, edgeStart, edgeEnd);
    D2D_COLS(rw_save, rw, L1, cellStart, cellEnd);
    D2D_COLS(w_2, w_1, L1, cellStart, cellEnd);
    D2D_COLS(rtheta_p_save, rtheta_p, L, cellStart, cellEnd);
    D2D_COLS(rho_p_save, rho_p, L, cellStart, cellEnd);
    D2D_COLS(theta_m_2, theta_m_1, L, cellStart, cellEnd);
    D2D_COLS(rho_zz_2, rho_zz_1, L, cellStart, cellEnd);
    D2D_COLS(rho_zz_old_split, rho_zz_1, L, cellStart, cellEnd);
    D2D_COLS(scalars_2, scalars_1, L * num_scalars, cellStart, cellEnd);
    return 0;
}

extern "C" int mpas_b200_compute_moist_coefficients(MPAS_MOIST_ARGS) {
    (void)nCells; (void)nEdges; (void)vertexStart; (void)vertexEnd; (void)cellSolveStart; (void)cellSolveEnd; (void)vertexSolveStart; (void)vertexSolveEnd;
    (void)edgeSolveStart; (void)edgeSolveEnd;
    MoistP p;
    p.L = nVertLevels; p.ns = num_scalars; p.ms = moist_start; p.me = moist_end; p.cellStart = cellStart; p.nCellsLoop = cellEnd - cellStart + 1;
    p.edgeStart = edgeStart; p.nEdgesLoop = edgeEnd - edgeStart + 1; p.nCellsSolve = nCellsSolve;
    p.cellsOnEdge = X(cellsOnEdge); p.scalars = X(scalars); p.cqw = X(cqw); p.cqu = X(cqu); p.qtot = X(qtot);
    KLAUNCH(moist_cell_kernel, (long long)p.nCellsLoop * p.L, 256, p);
    KCHECK("moist_cell_kernel");
    KLAUNCH(moist_edge_kernel, (long long)p.nEdgesLoop * p.L, 256, p);
    KCHECK("moist_edge_kernel");
    return 0;
}

extern "C" int mpas_b200_compute_vert_imp_coefs_work(MPAS_VERTIMP_ARGS) {
    (void)num_scalars; (void)nCells; (void)moist_start; (void)moist_end; (void)scalars; (void)cellStart; (void)cellEnd; (void)edgeStart; (void)edgeEnd;
    (void)edgeSolveStart; (void)edgeSolveEnd;
    VicP v;
    v.L = nVertLevels; v.cellSolveStart = cellSolveStart; v.nCellsLoop = cellSolveEnd - cellSolveStart + 1;
    v.dtseps = (real).5 * dts * ((real)1. + epssm);
    v.rcv = MPAS_RGAS / (MPAS_CP - MPAS_RGAS);
    v.c2 = MPAS_CP * v.rcv;
    v.zz = X(zz); v.cqw = X(cqw); v.p_ = X(p); v.t = X(t); v.rb = X(rb); v.rtb = X(rtb); v.pb = X(pb); v.rt = X(rt);
    v.rdzw = X(rdzw); v.fzm = X(fzm); v.fzp = X(fzp); v.rdzu = X(rdzu); v.qtot = X(qtot);
    v.cofwr = X(cofwr); v.cofwz = X(cofwz); v.coftz = X(coftz); v.cofwt = X(cofwt); v.a_tri = X(a_tri); v.alpha_tri = X(alpha_tri);
    v.gamma_tri = X(gamma_tri); v.cofrz = X(cofrz);
    cofrz_kernel<<<cdiv(v.L, 128), 128, 0, stream()>>>(v);
    count_launch();
    KCHECK("cofrz_kernel");
    if (v.nCellsLoop > 0) {
        const int wpb = 8;
        const size_t shm = (size_t)wpb * 5 * (v.L + 2) * sizeof(real);
        static bool attr_set = false;
        if (!attr_set) {
            cudaFuncSetAttribute(vert_imp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            attr_set = true;
        }
        if (shm > 200 * 1024) return fail_msg("vert_imp_kernel: nVertLevels too large for the shared-memory staging");
        ktime_begin("vert_imp_kernel");
        vert_imp_kernel<<<cdiv(v.nCellsLoop, wpb), wpb * 32, shm, stream()>>>(v);
        ktime_end();
        count_launch();
        KCHECK("vert_imp_kernel");
    }
    return 0;
}

extern "C" int mpas_b200_rk_dynamics_substep_finish(MPAS_SUBSTEPFINISH_ARGS) {
    (void)nCells; (void)nEdges; (void)vertexStart; (void)vertexEnd; (void)cellSolveStart; (void)cellSolveEnd; (void)vertexSolveStart; (void)vertexSolveEnd;
    (void)edgeSolveStart; (void)edgeSolveEnd;
    const int L = nVertLevels, L1 = L + 1;
    const real inv_dynamics_split = (real)1.0 / (real)dynamics_split;
    const long long nE = (long long)(edgeEnd - edgeStart + 1) * L, nC1 = (long long)(cellEnd - cellStart + 1) * L1;
    if (dynamics_substep < dynamics_split) {
        D2D_COLS(ru_save, ru, L, edgeStart, edgeEnd);
        D2D_COLS(u_1, u_2, L, edgeStart, edgeEnd);
        D2D_COLS(rw_save, rw, L1, cellStart, cellEnd);
        D2D_COLS(w_1, w_2, L1, cellStart, cellEnd);
        D2D_COLS(rtheta_p_save, rtheta_p, L, cellStart, cellEnd);
        D2D_COLS(rho_p_save, rho_p, L, cellStart, cellEnd);
        D2D_COLS(theta_m_1, theta_m_2, L, cellStart, cellEnd);
        D2D_COLS(rho_zz_1, rho_zz_2, L, cellStart, cellEnd);
    }
    if (dynamics_substep == 1) {
        D2D_COLS(ruAvg_split, ruAvg, L, edgeStart, edgeEnd);
        D2D_COLS(wwAvg_split, wwAvg, L1, cellStart, cellEnd);
    } else {
        AxP a;
        a.s = 0;
        a.n = nE; a.a = X(ruAvg) + IX2(1, edgeStart, L); a.b = X(ruAvg_split) + IX2(1, edgeStart, L); a.out = X(ruAvg_split) + IX2(1, edgeStart, L);
        KLAUNCH(ax_kernel<0>, a.n, 256, a);
        a.n = nC1; a.a = X(wwAvg) + IX2(1, cellStart, L1); a.b = X(wwAvg_split) + IX2(1, cellStart, L1); a.out = X(wwAvg_split) + IX2(1, cellStart, L1);
        KLAUNCH(ax_kernel<0>, a.n, 256, a);
        KCHECK("substep_finish accumulate");
    }
    if (dynamics_substep == dynamics_split) {
        AxP a;
        a.s = inv_dynamics_split; a.b = nullptr;
        a.n = nE; a.a = X(ruAvg_split) + IX2(1, edgeStart, L); a.out = X(ruAvg) + IX2(1, edgeStart, L);
        KLAUNCH(ax_kernel<1>, a.n, 256, a);
        a.n = nC1; a.a = X(wwAvg_split) + IX2(1, cellStart, L1); a.out = X(wwAvg) + IX2(1, cellStart, L1);
        KLAUNCH(ax_kernel<1>, a.n, 256, a);
        KCHECK("substep_finish average");
        D2D_COLS(rho_zz_1, rho_zz_old_split, L, cellStart, cellEnd);
    }
    return 0;
}

extern "C" int mpas_b200_init_coupled_diagnostics(MPAS_INITCOUPLED_ARGS) {
    (void)nCells; (void)nEdges; (void)vertexStart; (void)vertexEnd; (void)cellSolveStart; (void)cellSolveEnd; (void)vertexSolveStart; (void)vertexSolveEnd;
    (void)edgeSolveStart; (void)edgeSolveEnd;
    IcdP p;
    p.L = nVertLevels; p.mE = maxEdges; p.ns = num_scalars; p.index_qv = index_qv; p.cellStart = cellStart; p.nCellsLoop = cellEnd - cellStart + 1;
    p.edgeStart = edgeStart; p.nEdgesLoop = edgeEnd - edgeStart + 1;
    p.cellsOnEdge = X(cellsOnEdge); p.nEdgesOnCell = X(nEdgesOnCell); p.edgesOnCell = X(edgesOnCell); p.edgesOnCell_sign = X(edgesOnCell_sign);
    p.theta = X(theta); p.rho = X(rho); p.rho_base = X(rho_base); p.theta_base = X(theta_base); p.zz = X(zz); p.scalars = X(scalars); p.u = X(u); p.w = X(w);
    p.fzm = X(fzm); p.fzp = X(fzp); p.zb_cell = X(zb_cell); p.zb3_cell = X(zb3_cell);
    p.theta_m = X(theta_m); p.rho_zz = X(rho_zz); p.rho_p = X(rho_p); p.rtheta_base = X(rtheta_base); p.rtheta_p = X(rtheta_p); p.ru = X(ru); p.rw = X(rw);
    p.pressure_p = X(pressure_p); p.pressure_base = X(pressure_base); p.exner = X(exner); p.exner_base = X(exner_base);
    KLAUNCH(icd_cell1_kernel, (long long)p.nCellsLoop * p.L, 256, p);
    KCHECK("icd_cell1_kernel");
    KLAUNCH(icd_edge_kernel, (long long)p.nEdgesLoop * p.L, 256, p);
    KCHECK("icd_edge_kernel");
    KLAUNCH(icd_cell2_kernel, (long long)p.nCellsLoop * (p.L + 1), 256, p);
    KCHECK("icd_cell2_kernel");
    return 0;
}
