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
    KMagic km;
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
    KMagic km;
    int L, cellSolveStart, nCellsLoop;
    real dtseps, rcv, c2;
    const real *zz, *cqw, *p_, *t, *rb, *rtb, *pb, *rt, *rdzw, *fzm, *fzp, *rdzu, *qtot;
    real *cofwr, *cofwz, *coftz, *cofwt, *a_tri, *alpha_tri, *gamma_tri, *cofrz;
};

__global__ void cofrz_kernel(VicP p) {  // TI:2070-2072
    const int k = blockIdx.x * blockDim.x + threadIdx.x + 1;
    if (k <= p.L) p.cofrz[k - 1] = p.dtseps * p.rdzw[k - 1];
}

// One warp per column, lanes along k (two levels per lane and pass so that the 9 input columns of a 64-level chunk arrive
// in one round trip); the coefficients stay in shared memory between the passes.  The recurrence
// alpha(k) = 1/(b(k) - a(k)*gamma(k-1)), gamma(k) = c(k)*alpha(k) (TI:2124-2127) is serial in k, so one LANE runs it for one
// column: the block's WPB columns are swept together by the first WPB lanes of warp 0 between two block barriers, in the
// reference's order.  shared memory per warp: VI_NCOL columns of L + 2 reals.
enum { VI_ZZ, VI_P, VI_T, VI_CQW, VI_WT, VI_WR, VI_TZ, VI_C, VI_NCOL, VI_WZ = VI_CQW, VI_A = VI_P, VI_B = VI_T };
__global__ void __launch_bounds__(WPB * 32) vert_imp_kernel(VicP p) {
    extern __shared__ real smem[];
    const int L = p.L, L1 = L + 1, LS = L + 2;
    const int lane = threadIdx.x & 31, warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    real* sw = smem + (size_t)warp * VI_NCOL * LS;
    real *s_zz = sw + VI_ZZ * LS, *s_p = sw + VI_P * LS, *s_t = sw + VI_T * LS, *s_cqw = sw + VI_CQW * LS, *s_wt = sw + VI_WT * LS;
    real *s_wr = sw + VI_WR * LS, *s_tz = sw + VI_TZ * LS, *s_c = sw + VI_C * LS;
    const real dtseps = p.dtseps, rcv = p.rcv, c2 = p.c2;
    const int iCell = p.cellSolveStart + blockIdx.x * WPB + warp;
    const bool exists = iCell < p.cellSolveStart + p.nCellsLoop;
    if (exists) {
        const size_t c0 = IX2(1, iCell, L);
        // pass 1: inputs of levels ka, ka + 32; cofwt (TI:2100-2107)
        for (int k0 = 0; k0 < L; k0 += 64) {
            const int ka = k0 + lane + 1, kb = ka + 32;
            const bool va = ka <= L, vb = kb <= L;
            const size_t ia = c0 + ((va ? ka : L) - 1), ib = c0 + ((vb ? kb : L) - 1);
            const real zz_a = p.zz[ia], zz_b = p.zz[ib], p_a = p.p_[ia], p_b = p.p_[ib], t_a = p.t[ia], t_b = p.t[ib];
            const real cq_a = p.cqw[ia], cq_b = p.cqw[ib], q_a = p.qtot[ia], q_b = p.qtot[ib], rb_a = p.rb[ia], rb_b = p.rb[ib];
            const real rtb_a = p.rtb[ia], rtb_b = p.rtb[ib], rt_a = p.rt[ia], rt_b = p.rt[ib], pb_a = p.pb[ia], pb_b = p.pb[ib];
            if (va) {
                const real wt = (real).5 * dtseps * rcv * zz_a * MPAS_GRAVITY * rb_a / ((real)1. + q_a) * p_a / ((rtb_a + rt_a) * pb_a);
                p.cofwt[ia] = wt;
                s_wt[ka] = wt; s_zz[ka] = zz_a; s_p[ka] = p_a; s_t[ka] = t_a; s_cqw[ka] = cq_a;
            }
            if (vb) {
                const real wt = (real).5 * dtseps * rcv * zz_b * MPAS_GRAVITY * rb_b / ((real)1. + q_b) * p_b / ((rtb_b + rt_b) * pb_b);
                p.cofwt[ib] = wt;
                s_wt[kb] = wt; s_zz[kb] = zz_b; s_p[kb] = p_b; s_t[kb] = t_b; s_cqw[kb] = cq_b;
            }
        }
        __syncwarp();
        // pass 2: cofwr, cofwz, coftz (TI:2083-2098)
        for (int k = lane + 1; k <= L1; k += 32) {
            if (k == 1 || k == L1) {
                p.coftz[IX2(k, iCell, L1)] = 0.0;
                s_tz[k] = 0.0;
            } else {
                const size_t ic = c0 + (k - 1);
                const real fm = p.fzm[k - 1], fp = p.fzp[k - 1];
                const real zzk = s_zz[k], zzkm = s_zz[k - 1];
                const real wr = (real).5 * dtseps * MPAS_GRAVITY * (fm * zzk + fp * zzkm);
                const real wz = dtseps * c2 * (fm * zzk + fp * zzkm) * p.rdzu[k - 1] * s_cqw[k] * (fm * s_p[k] + fp * s_p[k - 1]);
                const real tz = dtseps * (fm * s_t[k] + fp * s_t[k - 1]);
                p.cofwr[ic] = wr;
                p.cofwz[ic] = wz;
                p.coftz[IX2(k, iCell, L1)] = tz;
                s_wr[k] = wr;
                s_tz[k] = tz;
                s_cqw[k] = wz;  // VI_WZ: only this lane reads s_cqw[k]
            }
        }
        __syncwarp();
        // pass 3: a, b, c of the tridiagonal system (TI:2109-2122); a and b replace the p and t columns (no longer read)
        for (int k = lane + 1; k <= L; k += 32) {
            real a, b, c;
            if (k == 1) {
                a = 0.;
                b = 1.;
                c = 0.;
            } else {
                const real cofwz = s_cqw[k], cofwr = s_wr[k], cofwt = s_wt[k], cofwtm = s_wt[k - 1];
                const real coftz_k = s_tz[k], coftz_km = s_tz[k - 1], coftz_kp = s_tz[k + 1];
                const real zzk = s_zz[k], zzkm = s_zz[k - 1];
                const real rdzw_k = p.rdzw[k - 1], rdzw_km = p.rdzw[k - 2];
                const real cofrz_k = dtseps * rdzw_k, cofrz_km = dtseps * rdzw_km;
                a = -cofwz * coftz_km * rdzw_km * zzkm + cofwr * cofrz_km - cofwtm * coftz_km * rdzw_km;
                b = (real)1. + cofwz * (coftz_k * rdzw_k * zzk + coftz_k * rdzw_km * zzkm) - coftz_k * (cofwt * rdzw_k - cofwtm * rdzw_km) +
                    cofwr * (cofrz_k - cofrz_km);
                c = -cofwz * coftz_kp * rdzw_k * zzk - cofwr * cofrz_k + cofwt * coftz_kp * rdzw_k;
            }
            s_p[k] = a;  // VI_A
            s_t[k] = b;  // VI_B -> alpha after the sweep
            s_c[k] = c;  // -> gamma after the sweep
        }
    }
    __syncthreads();
    if (warp == 0 && lane < WPB && p.cellSolveStart + blockIdx.x * WPB + lane < p.cellSolveStart + p.nCellsLoop) {
        real* __restrict__ xs = smem + (size_t)lane * VI_NCOL * LS;
        const real* __restrict__ xa = xs + VI_A * LS;
        real* __restrict__ xb = xs + VI_B * LS;
        real* __restrict__ xc = xs + VI_C * LS;
        xb[1] = 0.;
        xc[1] = 0.;
        real g = 0.;
        int k = 2;
        for (; k + 3 <= L; k += 4) {  // operands of four steps are read ahead of the dependent chain
            const real a0 = xa[k], a1 = xa[k + 1], a2 = xa[k + 2], a3 = xa[k + 3];
            const real b0 = xb[k], b1 = xb[k + 1], b2 = xb[k + 2], b3 = xb[k + 3];
            const real c0 = xc[k], c1 = xc[k + 1], c2_ = xc[k + 2], c3 = xc[k + 3];
            const real al0 = (real)1. / (b0 - a0 * g);
            const real g0 = c0 * al0;
            const real al1 = (real)1. / (b1 - a1 * g0);
            const real g1 = c1 * al1;
            const real al2 = (real)1. / (b2 - a2 * g1);
            const real g2 = c2_ * al2;
            const real al3 = (real)1. / (b3 - a3 * g2);
            g = c3 * al3;
            xb[k] = al0; xb[k + 1] = al1; xb[k + 2] = al2; xb[k + 3] = al3;
            xc[k] = g0; xc[k + 1] = g1; xc[k + 2] = g2; xc[k + 3] = g;
        }
        for (; k <= L; ++k) {
            const real al = (real)1. / (xb[k] - xa[k] * g);
            g = xc[k] * al;
            xb[k] = al;
            xc[k] = g;
        }
    }
    __syncthreads();
    if (exists) {
        const size_t c0 = IX2(1, iCell, L);
        for (int k = lane + 1; k <= L; k += 32) {
            p.a_tri[c0 + (k - 1)] = s_p[k];
            p.alpha_tri[c0 + (k - 1)] = s_t[k];
            p.gamma_tri[c0 + (k - 1)] = s_c[k];
        }
    }
}

struct AxP {
    KMagic km;
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
    KMagic km;
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
    p.L = nVertLevels; p.km.set(nVertLevels, nVertLevels + 1); p.ns = num_scalars; p.ms = moist_start; p.me = moist_end; p.cellStart = cellStart; p.nCellsLoop = cellEnd - cellStart + 1;
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
        const int wpb = WPB;
        const size_t shm = (size_t)wpb * VI_NCOL * (v.L + 2) * sizeof(real);
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

// ---- mpas_reconstruct_2d (operators/mpas_vector_reconstruction.F:195-305)
namespace {
struct RecP {
    KMagic km;
    int L, mE, nCellsLoop, on_a_sphere;
    const real *u, *coeffs, *latCell, *lonCell;
    const int *nEdgesOnCell, *edgesOnCell;
    real *ux, *uy, *uz, *uzonal, *umerid;
};
// one thread per (k, cell); the NE edge indices, then all coefficients and velocities are loaded before the ordered sums
// (:254-269).  The rotation (:275-288) evaluates cos/sin of the cell's latitude and longitude on the device (<= 2 ulp from
// the host's libm, the only non-bit-exact part of this routine).
template <int NE>
__global__ void __launch_bounds__(256) reconstruct_kernel(RecP p) {
    const int L = p.L, mE = p.mE;
    FLAT_KC(p.nCellsLoop, L);
    const int iCell = 1 + _c0;
    const int ne = p.nEdgesOnCell[iCell - 1];
    const int* eoc = p.edgesOnCell + (size_t)(iCell - 1) * mE;
    const real* cf = p.coeffs + (size_t)(iCell - 1) * mE * 3;
    int e[NE];
#pragma unroll
    for (int i = 0; i < NE; ++i) e[i] = eoc[i < ne ? i : 0];
    real c1[NE], c2[NE], c3[NE], ue[NE];
#pragma unroll
    for (int i = 0; i < NE; ++i) {
        const int j = i < ne ? i : 0;
        c1[i] = cf[3 * j];
        c2[i] = cf[3 * j + 1];
        c3[i] = cf[3 * j + 2];
        ue[i] = p.u[IX2(k, e[i], L)];
    }
    real x = 0.0, y = 0.0, z = 0.0;
#pragma unroll
    for (int i = 0; i < NE; ++i)
        if (i < ne) {
            x = x + c1[i] * ue[i];
            y = y + c2[i] * ue[i];
            z = z + c3[i] * ue[i];
        }
    for (int i = NE + 1; i <= ne; ++i) {
        const real uu = p.u[IX2(k, eoc[i - 1], L)];
        x = x + cf[3 * (i - 1)] * uu;
        y = y + cf[3 * (i - 1) + 1] * uu;
        z = z + cf[3 * (i - 1) + 2] * uu;
    }
    const size_t ic = IX2(k, iCell, L);
    p.ux[ic] = x;
    p.uy[ic] = y;
    p.uz[ic] = z;
    if (p.on_a_sphere) {
        real slat, clat, slon, clon;
        sincos(p.latCell[iCell - 1], &slat, &clat);
        sincos(p.lonCell[iCell - 1], &slon, &clon);
        p.uzonal[ic] = -x * slon + y * clon;
        p.umerid[ic] = -(x * clon + y * slon) * slat + z * clat;
    } else {
        p.uzonal[ic] = x;
        p.umerid[ic] = y;
    }
}
}  // namespace

extern "C" int mpas_b200_reconstruct_2d(MPAS_RECONSTRUCT_ARGS) {
    (void)nCells; (void)nEdges;
    RecP p;
    p.L = nVertLevels; p.km.set(nVertLevels, nVertLevels + 1); p.mE = maxEdges; p.nCellsLoop = nCellsLoop; p.on_a_sphere = on_a_sphere;
    p.u = X(u); p.coeffs = X(coeffs_reconstruct); p.latCell = X(latCell); p.lonCell = X(lonCell);
    p.nEdgesOnCell = X(nEdgesOnCell); p.edgesOnCell = X(edgesOnCell);
    p.ux = X(uReconstructX); p.uy = X(uReconstructY); p.uz = X(uReconstructZ); p.uzonal = X(uReconstructZonal); p.umerid = X(uReconstructMeridional);
    if (p.nCellsLoop <= 0) return 0;
    if (maxEdges <= 6) KLAUNCH((reconstruct_kernel<6>), (long long)p.nCellsLoop * p.L, 256, p);
    else KLAUNCH((reconstruct_kernel<8>), (long long)p.nCellsLoop * p.L, 256, p);
    KCHECK("reconstruct_kernel");
    return 0;
}

extern "C" int mpas_b200_init_coupled_diagnostics(MPAS_INITCOUPLED_ARGS) {
    (void)nCells; (void)nEdges; (void)vertexStart; (void)vertexEnd; (void)cellSolveStart; (void)cellSolveEnd; (void)vertexSolveStart; (void)vertexSolveEnd;
    (void)edgeSolveStart; (void)edgeSolveEnd;
    IcdP p;
    p.L = nVertLevels; p.km.set(nVertLevels, nVertLevels + 1); p.mE = maxEdges; p.ns = num_scalars; p.index_qv = index_qv; p.cellStart = cellStart; p.nCellsLoop = cellEnd - cellStart + 1;
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

// ---- atm_compute_output_diagnostics (src/core_atmosphere/mpas_atm_core.F:753-802): theta, rho, pressure for the history stream
namespace {
struct OutDiagP {
    KMagic km;
    int L, ns, iqv, nCells;
    const real *theta_m, *rho_zz, *scalars, *zz, *pressure_base, *pressure_p;
    real *theta, *rho, *pressure;
};
__global__ void __launch_bounds__(256) output_diag_kernel(OutDiagP p) {
    const int L = p.L;
    FLAT_KC(p.nCells, L);
    const size_t ic = IX2(k, _c0 + 1, L);
    p.theta[ic] = p.theta_m[ic] / ((real)1. + MPAS_RVORD * p.scalars[ic * (size_t)p.ns + (size_t)(p.iqv - 1)]);
    p.rho[ic] = p.rho_zz[ic] * p.zz[ic];
    p.pressure[ic] = p.pressure_base[ic] + p.pressure_p[ic];
}
}  // namespace

extern "C" int mpas_b200_compute_output_diagnostics(MPAS_OUTPUTDIAG_ARGS) {
    OutDiagP p;
    p.L = nVertLevels; p.km.set(nVertLevels); p.ns = num_scalars; p.iqv = index_qv; p.nCells = nCells;
    p.theta_m = X(theta_m); p.rho_zz = X(rho_zz); p.scalars = X(scalars); p.zz = X(zz); p.pressure_base = X(pressure_base); p.pressure_p = X(pressure_p);
    p.theta = X(theta); p.rho = X(rho); p.pressure = X(pressure);
    KLAUNCH(output_diag_kernel, (long long)nCells * nVertLevels, 256, p);
    KCHECK("output_diag_kernel");
    return 0;
}
