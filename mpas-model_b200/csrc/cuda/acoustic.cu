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
#include <map>
#include <mutex>

#include "common.cuh"

using namespace mpasb200;

namespace {

struct AcEdgeP {
    KMagic km;
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
    const real* tend_ru_fold;  // FOLD: tend_ru in place of ru_p
    real *rw_p, *rtheta_pp, *rtheta_pp_old, *wwAvg, *rho_pp;
};

// Shared-memory column slots of one warp of acoustic_cell_kernel (tma_slot(L + 1) reals each).  The first AC_NTMA_* are
// filled by bulk (TMA) copies, the rest is scratch.
enum { AC_ZZ, AC_COFTZ, AC_COFWT, AC_RHOZZ, AC_ATRI, AC_ALTRI, AC_GATRI, AC_NTMA_FIRST, AC_RTPP = AC_NTMA_FIRST, AC_RHOPP, AC_RWPOLD, AC_NTMA, AC_TS = AC_NTMA, AC_RS, AC_RWP, AC_NCOL };

// rs/ts -= flux over NE edges for the two levels of a lane; all 6*NE loads are issued before the first dependent subtract
// SCALE: the gathered column holds tend_ru and ru_p = scale * tend_ru is formed on the fly (first substep, folded edge pass)
template <int NE, bool SCALE>
__device__ __forceinline__ void ac_flux_sub(const real* __restrict__ rup_ka, const real* __restrict__ thp_ka, const unsigned* eo, const unsigned* c1o,
                                            const unsigned* c2o, const real* coef, real invArea, int bo, real scale, real& rs_a, real& rs_b, real& ts_a,
                                            real& ts_b) {
    real ua[NE], ub[NE], t1a[NE], t1b[NE], t2a[NE], t2b[NE];
#pragma unroll
    for (int t = 0; t < NE; ++t) {
        const real* __restrict__ q = elem_ptr(rup_ka, eo[t]);
        const real* __restrict__ q1 = elem_ptr(thp_ka, c1o[t]);
        const real* __restrict__ q2 = elem_ptr(thp_ka, c2o[t]);
        ua[t] = SCALE ? scale * q[0] : q[0];
        ub[t] = SCALE ? scale * q[bo] : q[bo];
        t1a[t] = q1[0];
        t1b[t] = q1[bo];
        t2a[t] = q2[0];
        t2b[t] = q2[bo];
    }
#pragma unroll
    for (int t = 0; t < NE; ++t) {
        const real c = coef[t];
        const real fa = c * ua[t] * invArea, fb = c * ub[t] * invArea;
        rs_a = rs_a - fa;
        rs_b = rs_b - fb;
        ts_a = ts_a - fa * (real)0.5 * (t2a[t] + t1a[t]);
        ts_b = ts_b - fb * (real)0.5 * (t2b[t] + t1b[t]);
    }
}

// One column per warp, lanes along k.  Columns with a vertical stencil or several uses (zz, coftz, cofwt, rho_zz, the
// tridiagonal coefficients, and after the first substep rtheta_pp, rho_pp, rw_p) are staged by one bulk (TMA) copy each,
// issued before the stencil metadata is read; the flux divergence over edgesOnCell is a direct gather with two levels
// per lane, the edge list staged as element offsets + the coefficient (sign*dts)*dvEdge (zero beyond nEdgesOnCell).
// FIRST = small_step 1: rtheta_pp, rho_pp, rw_p, wwAvg start from zero and are not read (TI:2603-2622).
// dynamic shared memory: WPB * AC_NCOL * tma_slot(L + 1) reals.
// FOLD (first substep of the device-resident driver only): the edge pass ru_p = dts * tend_ru is not run; the fluxes read
// tend_ru and scale it, the divergence damping that follows writes ru_p and ruAvg (divdamp_kernel<true>)
template <bool FIRST, bool FOLD = false, int MINB = 4>
__global__ void __launch_bounds__(WPB * 32, MINB) acoustic_cell_kernel(AcCellP p) {
    extern __shared__ __align__(16) real s_dyn[];
    __shared__ __align__(8) unsigned long long s_bar[WPB];
    __shared__ unsigned s_eo[WPB][8], s_c1[WPB][8], s_c2[WPB][8];
    __shared__ real s_coef[WPB][8];
    const int L = p.L, L1 = L + 1, mE = p.mE, LS = tma_slot(L1);
    const int lane = threadIdx.x & 31, warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform for the compiler
    __shared__ int s_tv[WPB][5];  // per warp: active flag and the s_dyn indices of its rwp, a, alpha, gamma columns (level 1)
    const int iCell0 = p.cellStart + blockIdx.x * WPB + warp;
    const bool exists = iCell0 <= p.cellEnd;
    const int iCell = exists ? iCell0 : p.cellEnd;
    const real dts = p.dts, epssm = p.epssm, resm = p.resm;
    const unsigned coff = (unsigned)(iCell - 1) * (unsigned)L, woff = (unsigned)(iCell - 1) * (unsigned)L1;
    const bool solve = exists && !(iCell > p.cellSolveEnd || iCell < p.cellSolveStart);
    const bool spec = solve && p.specZoneMaskCell[iCell - 1] != (real)0.0;
    if (!solve || spec) {  // halo column: TI:2603-2611 copy only; specified zone of a regional run: TI:2710-2719
        if (lane == 0) s_tv[warp][0] = 0;
        if (exists)
        for (int k = lane + 1; k <= L; k += 32) {
            const size_t ic = (size_t)coff + (k - 1), iw = (size_t)woff + (k - 1);
            const real rt0 = FIRST ? (real)0.0 : p.rtheta_pp[ic];
            p.rtheta_pp_old[ic] = rt0;
            if (spec) {
                const real rho0 = FIRST ? (real)0.0 : p.rho_pp[ic];
                const real rw0 = FIRST ? (real)0.0 : p.rw_p[iw], ww0 = FIRST ? (real)0.0 : p.wwAvg[iw];
                p.rho_pp[ic] = rho0 + dts * p.tend_rho[ic];
                p.rtheta_pp[ic] = rt0 + dts * p.tend_rt[ic];
                const real r = rw0 + dts * p.tend_rw[iw];
                p.rw_p[iw] = r;
                p.wwAvg[iw] = ww0 + (real)0.5 * ((real)1.0 + epssm) * r;
            }
        }
        if (spec && FIRST && lane == 0) {
            p.wwAvg[IX2(L1, iCell, L1)] = 0.0;
            p.rw_p[IX2(L1, iCell, L1)] = 0.0;
        }
        __syncthreads();  // the two block barriers around the tridiagonal sweep of the other warps
        __syncthreads();
        return;
    }
    // the first substep stages no rtheta_pp / rho_pp / rw_p columns: its scratch columns move down, 10 slots per warp
    constexpr int C_TS = FIRST ? (int)AC_NTMA_FIRST : (int)AC_TS, C_RS = C_TS + 1, C_RWP = C_TS + 2, NCOL = C_TS + 3;
    real* sc = s_dyn + (size_t)warp * NCOL * LS;
    const unsigned mbar = (unsigned)__cvta_generic_to_shared(&s_bar[warp]);
    const real *g_zz = p.zz + coff, *g_coftz = p.coftz + woff, *g_cofwt = p.cofwt + coff, *g_rhozz = p.rho_zz + coff;
    const real *g_a = p.a_tri + coff, *g_al = p.alpha_tri + coff, *g_ga = p.gamma_tri + coff;
    const real *g_rtpp = p.rtheta_pp + coff, *g_rhopp = p.rho_pp + coff, *g_rwp = p.rw_p + woff;
    if (elect_one()) {
        mbar_init(mbar, 1);
        const unsigned s0 = (unsigned)__cvta_generic_to_shared(sc), cs = (unsigned)(LS * sizeof(real));
        unsigned total = 0;
        tma_issue_col(s0 + AC_ZZ * cs, g_zz, L, mbar, total);
        tma_issue_col(s0 + AC_COFTZ * cs, g_coftz, L1, mbar, total);
        tma_issue_col(s0 + AC_COFWT * cs, g_cofwt, L, mbar, total);
        tma_issue_col(s0 + AC_RHOZZ * cs, g_rhozz, L, mbar, total);
        tma_issue_col(s0 + AC_ATRI * cs, g_a, L, mbar, total);
        tma_issue_col(s0 + AC_ALTRI * cs, g_al, L, mbar, total);
        tma_issue_col(s0 + AC_GATRI * cs, g_ga, L, mbar, total);
        if (!FIRST) {
            tma_issue_col(s0 + AC_RTPP * cs, g_rtpp, L, mbar, total);
            tma_issue_col(s0 + AC_RHOPP * cs, g_rhopp, L, mbar, total);
            tma_issue_col(s0 + AC_RWPOLD * cs, g_rwp, L1, mbar, total);
        }
        mbar_arrive_expect_tx(mbar, total);
    }
    // 1-based views
    const real* s_zz = sc + AC_ZZ * LS + tma_shift(g_zz) - 1;
    const real* s_coftz = sc + AC_COFTZ * LS + tma_shift(g_coftz) - 1;
    const real* s_cofwt = sc + AC_COFWT * LS + tma_shift(g_cofwt) - 1;
    const real* s_rhozz = sc + AC_RHOZZ * LS + tma_shift(g_rhozz) - 1;
    const real* sa = sc + AC_ATRI * LS + tma_shift(g_a) - 1;
    const real* sal = sc + AC_ALTRI * LS + tma_shift(g_al) - 1;
    const real* sga = sc + AC_GATRI * LS + tma_shift(g_ga) - 1;
    const real* s_rtpp = sc + AC_RTPP * LS + tma_shift(g_rtpp) - 1;
    const real* s_rhopp = sc + AC_RHOPP * LS + tma_shift(g_rhopp) - 1;
    const real* s_rwold = sc + AC_RWPOLD * LS + tma_shift(g_rwp) - 1;
    real* ts = sc + C_TS * LS - 1;
    real* rs = sc + C_RS * LS - 1;
    real* rwp = sc + C_RWP * LS - 1;

    const int ne = p.nEdgesOnCell[iCell - 1];
    const int npad = ne <= 6 ? 6 : 8;
    const real invArea = p.invAreaCell[iCell - 1];
    if (lane < npad) {
        const bool in = lane < ne;
        const int iEdge = p.edgesOnCell[IX2(in ? lane + 1 : 1, iCell, mE)];
        s_eo[warp][lane] = (unsigned)(iEdge - 1) * (unsigned)L;
        s_c1[warp][lane] = (unsigned)(p.cellsOnEdge[IX2(1, iEdge, 2)] - 1) * (unsigned)L;
        s_c2[warp][lane] = (unsigned)(p.cellsOnEdge[IX2(2, iEdge, 2)] - 1) * (unsigned)L;
        s_coef[warp][lane] = in ? p.edgesOnCell_sign[IX2(lane + 1, iCell, mE)] * dts * p.dvEdge[iEdge - 1] : (real)0.0;
    }
    __syncwarp();
    // ---- flux divergence, TI:2629-2639 (direct gather; the first chunk overlaps the bulk copies), then TI:2646-2653.
    // Each lane owns levels k0 + lane + 1 and k0 + lane + 33 of a 64-level chunk.
    bool staged = false;
    for (int k0 = 0; k0 < L; k0 += 64) {
        const int ka = k0 + lane;
        const bool va = ka < L, vb = ka + 32 < L;
        const int kl = va ? ka : L - 1;  // lanes past the column recompute its last level (loads stay in bounds)
        const int bo = vb ? 32 : 0;
        real rs_a = 0.0, rs_b = 0.0, ts_a = 0.0, ts_b = 0.0;
        {
            const real* __restrict__ rup = (FOLD ? p.tend_ru_fold : p.ru_p) + kl;
            const real* __restrict__ thp = p.theta_m + kl;
            ac_flux_sub<3, FOLD>(rup, thp, s_eo[warp], s_c1[warp], s_c2[warp], s_coef[warp], invArea, bo, dts, rs_a, rs_b, ts_a, ts_b);
            ac_flux_sub<3, FOLD>(rup, thp, s_eo[warp] + 3, s_c1[warp] + 3, s_c2[warp] + 3, s_coef[warp] + 3, invArea, bo, dts, rs_a, rs_b, ts_a, ts_b);
            if (npad > 6) ac_flux_sub<2, FOLD>(rup, thp, s_eo[warp] + 6, s_c1[warp] + 6, s_c2[warp] + 6, s_coef[warp] + 6, invArea, bo, dts, rs_a, rs_b, ts_a, ts_b);
        }
        // single-use operands of this chunk, loaded now so that they share the round trip of the gather
        const real trho_a = p.tend_rho[coff + kl], trho_b = p.tend_rho[coff + kl + bo];
        const real trt_a = p.tend_rt[coff + kl], trt_b = p.tend_rt[coff + kl + bo];
        if (!staged) {
            mbar_wait(mbar, 0);
            // TI:2603-2611 and the working copy of the old rw_p column
            for (int k = lane + 1; k <= L1; k += 32) rwp[k] = FIRST ? (real)0.0 : s_rwold[k];
            for (int k = lane + 1; k <= L; k += 32) p.rtheta_pp_old[(size_t)coff + (k - 1)] = FIRST ? (real)0.0 : s_rtpp[k];
            __syncwarp();
            staged = true;
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 0 ? !va : !vb) continue;
            const int k = ka + 1 + 32 * h;
            real rsk = h ? rs_b : rs_a, tsk = h ? ts_b : ts_a;
            const real rho_pp_k = FIRST ? (real)0.0 : s_rhopp[k];
            const real rth_pp_k = FIRST ? (real)0.0 : s_rtpp[k];
            rsk = rho_pp_k + dts * (h ? trho_b : trho_a) + rsk - p.cofrz[k - 1] * resm * (rwp[k + 1] - rwp[k]);
            tsk = rth_pp_k + dts * (h ? trt_b : trt_a) + tsk - resm * p.rdzw[k - 1] * (s_coftz[k + 1] * rwp[k + 1] - s_coftz[k] * rwp[k]);
            rs[k] = rsk;
            ts[k] = tsk;
        }
    }
    __syncwarp();
    // explicit part of rw_p, TI:2659-2670 (reads only rwp[k] of its own level); the single-use operands of a lane's two
    // levels share one round trip
    for (int k0 = 0; k0 < L; k0 += 64) {
        const int ka = k0 + lane + 1, kb = ka + 32;
        const bool va = ka >= 2 && ka <= L, vb = kb <= L;
        const int kla = va ? ka : 2, klb = vb ? kb : 2;
        const real trw_a = p.tend_rw[(size_t)woff + (kla - 1)], trw_b = p.tend_rw[(size_t)woff + (klb - 1)];
        const real cwz_a = p.cofwz[(size_t)coff + (kla - 1)], cwz_b = p.cofwz[(size_t)coff + (klb - 1)];
        const real cwr_a = p.cofwr[(size_t)coff + (kla - 1)], cwr_b = p.cofwr[(size_t)coff + (klb - 1)];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 0 ? !va : !vb) continue;
            const int k = h ? kb : ka;
            const real rtpp_k = FIRST ? (real)0.0 : s_rtpp[k], rtpp_km = FIRST ? (real)0.0 : s_rtpp[k - 1];
            const real rpp_k = FIRST ? (real)0.0 : s_rhopp[k], rpp_km = FIRST ? (real)0.0 : s_rhopp[k - 1];
            const real zzk = s_zz[k], zzkm = s_zz[k - 1];
            rwp[k] = rwp[k] + dts * (h ? trw_b : trw_a) -
                     (h ? cwz_b : cwz_a) * ((zzk * ts[k] - zzkm * ts[k - 1]) + resm * (zzk * rtpp_k - zzkm * rtpp_km)) -
                     (h ? cwr_b : cwr_a) * ((rs[k] + rs[k - 1]) + resm * (rpp_k + rpp_km)) + s_cofwt[k] * (ts[k] + resm * rtpp_k) +
                     s_cofwt[k - 1] * (ts[k - 1] + resm * rtpp_km);
        }
    }
    // operands of the Rayleigh-damping pass of the first 64 levels, requested before the sweep so that they arrive during it
    const int pka = lane + 1, pkb = lane + 33;
    const bool pva = pka >= 2 && pka <= L, pvb = pkb <= L;
    const int pla = pva ? pka : 2, plb = pvb ? pkb : 2;
    const real ww_a = FIRST ? (real)0.0 : p.wwAvg[(size_t)woff + (pla - 1)], ww_b = FIRST ? (real)0.0 : p.wwAvg[(size_t)woff + (plb - 1)];
    const real d_a = p.rw_save[(size_t)woff + (pla - 1)] - p.rw[(size_t)woff + (pla - 1)];
    const real d_b = p.rw_save[(size_t)woff + (plb - 1)] - p.rw[(size_t)woff + (plb - 1)];
    const real dk_a = p.dss[(size_t)coff + (pla - 1)], dk_b = p.dss[(size_t)coff + (plb - 1)];
    const real w_a = p.w[(size_t)woff + (pla - 1)], w_b = p.w[(size_t)woff + (plb - 1)];
    // ---- tridiagonal sweep, TI:2675-2682, in the reference's order.  The recurrence is serial in k, so one LANE sweeps one
    // column: the block's WPB columns are swept together by the first WPB lanes of warp 0 (1/WPB of the instructions of
    // a sweep per warp), between two block barriers.
    if (lane == 0) {
        s_tv[warp][0] = 1;
        s_tv[warp][1] = (int)(rwp - s_dyn);
        s_tv[warp][2] = (int)(sa - s_dyn);
        s_tv[warp][3] = (int)(sal - s_dyn);
        s_tv[warp][4] = (int)(sga - s_dyn);
    }
    __syncthreads();
    if (warp == 0 && lane < WPB && s_tv[lane][0]) {
        real* __restrict__ x = s_dyn + s_tv[lane][1];
        const real* __restrict__ ca = s_dyn + s_tv[lane][2];
        const real* __restrict__ cal = s_dyn + s_tv[lane][3];
        const real* __restrict__ cga = s_dyn + s_tv[lane][4];
        // operands of four steps are read ahead of the dependent chain (mul, sub, mul per step)
        real prev = x[1];
        int k = 2;
        for (; k + 3 <= L; k += 4) {
            const real b0 = x[k], b1 = x[k + 1], b2 = x[k + 2], b3 = x[k + 3];
            const real a0 = ca[k], a1 = ca[k + 1], a2 = ca[k + 2], a3 = ca[k + 3];
            const real l0 = cal[k], l1 = cal[k + 1], l2 = cal[k + 2], l3 = cal[k + 3];
            const real y0 = (b0 - a0 * prev) * l0;
            const real y1 = (b1 - a1 * y0) * l1;
            const real y2 = (b2 - a2 * y1) * l2;
            const real y3 = (b3 - a3 * y2) * l3;
            x[k] = y0;
            x[k + 1] = y1;
            x[k + 2] = y2;
            x[k + 3] = y3;
            prev = y3;
        }
        for (; k <= L; ++k) {
            prev = (x[k] - ca[k] * prev) * cal[k];
            x[k] = prev;
        }
        real next = x[L1];
        k = L;
        for (; k - 3 >= 1; k -= 4) {
            const real b0 = x[k], b1 = x[k - 1], b2 = x[k - 2], b3 = x[k - 3];
            const real g0 = cga[k], g1 = cga[k - 1], g2 = cga[k - 2], g3 = cga[k - 3];
            const real y0 = b0 - g0 * next;
            const real y1 = b1 - g1 * y0;
            const real y2 = b2 - g2 * y1;
            const real y3 = b3 - g3 * y2;
            x[k] = y0;
            x[k - 1] = y1;
            x[k - 2] = y2;
            x[k - 3] = y3;
            next = y3;
        }
        for (; k >= 1; --k) {
            next = x[k] - cga[k] * next;
            x[k] = next;
        }
    }
    __syncthreads();
    for (int k = lane + 1; k <= L; k += 32) {  // TI:2655-2657, 2687-2699
        if (k < 2) continue;
        const size_t ic = (size_t)coff + (k - 1), iw = (size_t)woff + (k - 1);
        const bool pre = k <= 64;  // preloaded above
        const bool hb = k > 32;
        const real rw_old = FIRST ? (real)0.0 : s_rwold[k];
        real ww = pre ? (hb ? ww_b : ww_a) : (FIRST ? (real)0.0 : p.wwAvg[iw]);
        ww = ww + (real)0.5 * ((real)1.0 - epssm) * rw_old;
        const real d = pre ? (hb ? d_b : d_a) : p.rw_save[iw] - p.rw[iw];
        const real dk = pre ? (hb ? dk_b : dk_a) : p.dss[ic];
        const real wk = pre ? (hb ? w_b : w_a) : p.w[iw];
        const real r = (rwp[k] + d -
                        dts * dk * (p.fzm[k - 1] * s_zz[k] + p.fzp[k - 1] * s_zz[k - 1]) *
                            (p.fzm[k - 1] * s_rhozz[k] + p.fzp[k - 1] * s_rhozz[k - 1]) * wk) /
                           ((real)1.0 + dts * dk) -
                       d;
        rwp[k] = r;
        ww = ww + (real)0.5 * ((real)1.0 + epssm) * r;
        p.wwAvg[iw] = ww;
    }
    __syncwarp();
    for (int k = lane + 1; k <= L; k += 32) {  // TI:2701-2708
        const size_t ic = (size_t)coff + (k - 1);
        p.rw_p[(size_t)woff + (k - 1)] = rwp[k];
        p.rho_pp[ic] = rs[k] - p.cofrz[k - 1] * (rwp[k + 1] - rwp[k]);
        p.rtheta_pp[ic] = ts[k] - p.rdzw[k - 1] * (s_coftz[k + 1] * rwp[k + 1] - s_coftz[k] * rwp[k]);
    }
    if (FIRST && lane == 0) {  // TI:2617-2622: the levels the loops above never write
        p.wwAvg[IX2(1, iCell, L1)] = 0.0;
        p.wwAvg[IX2(L1, iCell, L1)] = 0.0;
        p.rw_p[IX2(L1, iCell, L1)] = 0.0;
    }
}

struct DivDampP {
    KMagic km;
    int L, nCellsSolve, edgeStart, nEdgesLoop;
    real coef_divdamp;
    const int* cellsOnEdge;
    const real *theta_m, *rtheta_pp, *rtheta_pp_old, *specZoneMaskEdge;
    real* ru_p;
    const real* tend_ru;  // FOLD: ru_p = dts * tend_ru has not been stored yet
    real* ruAvg;
    real dts;
    // NEXT: the edge pass of the substep that follows (acoustic_edge_kernel, small_step > 1)
    real c2;
    const real *rho_pp, *zz, *exner, *cqu, *zxu, *invDcEdge;
};

// FOLD: this call also does the first substep's edge pass (TI:2592-2597: ru_p = dts * tend_ru, ruAvg = ru_p) that the acoustic
// step left out -- the same products, one read of tend_ru instead of a write and a read of ru_p
// NEXT: a substep follows, and this call also runs ITS edge pass (TI:2552-2590 with the ru_p just damped): ru_p is read and written
// once instead of twice, rtheta_pp of the two cells is gathered once, 3 launches per step less.  The driver issues the rho_pp halo
// exchange of that substep before this call.
template <bool FOLD, bool NEXT>
__global__ void __launch_bounds__(256) divdamp_kernel(DivDampP p) {
    FLAT_KC(p.nEdgesLoop, p.L);
    const int L = p.L;
    const int iEdge = p.edgeStart + _c0;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    if (!(cell1 <= p.nCellsSolve || cell2 <= p.nCellsSolve)) return;
    const size_t i1 = IX2(k, cell1, L), i2 = IX2(k, cell2, L), ie = IX2(k, iEdge, L);
    const real divCell1 = -(p.rtheta_pp[i1] - p.rtheta_pp_old[i1]);
    const real divCell2 = -(p.rtheta_pp[i2] - p.rtheta_pp_old[i2]);
    real r, avg = 0;
    if (FOLD) {
        r = p.dts * p.tend_ru[ie];
        avg = r;
    } else {
        r = p.ru_p[ie];
    }
    r = r + p.coef_divdamp * (divCell2 - divCell1) * ((real)1.0 - p.specZoneMaskEdge[iEdge - 1]) / (p.theta_m[i1] + p.theta_m[i2]);
    if (NEXT) {
        real pgrad = ((p.rtheta_pp[i2] - p.rtheta_pp[i1]) * p.invDcEdge[iEdge - 1]) / ((real).5 * (p.zz[i2] + p.zz[i1]));
        pgrad = p.cqu[ie] * (real)0.5 * p.c2 * (p.exner[i1] + p.exner[i2]) * pgrad;
        pgrad = pgrad + (real)0.5 * p.zxu[ie] * MPAS_GRAVITY * (p.rho_pp[i1] + p.rho_pp[i2]);
        const real r2 = r + p.dts * (p.tend_ru[ie] - ((real)1.0 - p.specZoneMaskEdge[iEdge - 1]) * pgrad);
        if (!FOLD) avg = p.ruAvg[ie];
        p.ru_p[ie] = r2;
        p.ruAvg[ie] = avg + r2;
    } else {
        p.ru_p[ie] = r;
        if (FOLD) p.ruAvg[ie] = avg;
    }
}

// Folded first substep: which ru_p arrays wait for their edge pass (one entry per block between the acoustic step and the
// divergence damping of the same substep)
struct FoldEntry {
    AcEdgeP ep;   // the edge pass's parameters (same arrays for every substep of an RK stage)
    bool first;   // the first substep's edge pass is still to be done
    bool next;    // the next substep's edge pass is to be done as well
};
std::mutex g_fold_mu;
std::map<const void*, FoldEntry> g_fold;
std::map<const void*, int> g_edge_done;  // ru_p arrays whose next edge pass has been run by the damping
thread_local int tl_fold_flags = 0;

}  // namespace

namespace mpasb200 {
// device-resident driver (srk3.cu): the next acoustic step on this thread may leave its first-substep edge pass to the
// divergence damping that follows it
void acoustic_fold_next_call(int flags) { tl_fold_flags = flags; }
}  // namespace mpasb200

extern "C" int mpas_b200_advance_acoustic_step_work(MPAS_ACOUSTIC_ARGS) {
    (void)nCells; (void)nEdges; (void)vertexStart; (void)vertexEnd; (void)vertexSolveStart; (void)vertexSolveEnd;
    (void)edgeSolveStart; (void)edgeSolveEnd; (void)zgrid; (void)ru; (void)ru_save; (void)dcEdge; (void)cf1; (void)cf2; (void)cf3;
    const real rcv = MPAS_RGAS / (MPAS_CP - MPAS_RGAS);
    const real c2 = MPAS_CP * rcv;
    const real resm = ((real)1.0 - epssm) / ((real)1.0 + epssm);
    const int L = nVertLevels;
    bool fold = false;
    {
        AcEdgeP p;
        p.L = L; p.km.set(L, L + 1); p.nCellsSolve = nCellsSolve; p.edgeStart = edgeStart; p.nEdgesLoop = edgeEnd - edgeStart + 1; p.small_step = small_step;
        p.dts = dts; p.c2 = c2;
        p.cellsOnEdge = X(cellsOnEdge);
        p.rtheta_pp = X(rtheta_pp); p.rho_pp = X(rho_pp); p.zz = X(zz); p.exner = X(exner); p.cqu = X(cqu); p.zxu = X(zxu);
        p.tend_ru = X(tend_ru); p.invDcEdge = X(invDcEdge); p.specZoneMaskEdge = X(specZoneMaskEdge);
        p.ru_p = X(ru_p); p.ruAvg = X(ruAvg);
        const int flags = tl_fold_flags;
        tl_fold_flags = 0;
        fold = (flags & ACOUSTIC_FOLD_FIRST) && small_step == 1;
        const bool next = (flags & ACOUSTIC_FOLD_NEXT) != 0;
        bool done = false;
        {
            std::lock_guard<std::mutex> lk(g_fold_mu);
            if (g_fold.count(p.ru_p)) return fail_msg("mpas_b200_advance_acoustic_step_work: folded edge pass of the previous call was never consumed");
            if ((flags & ACOUSTIC_FOLD_DONE) && small_step > 1) {
                if (!g_edge_done.erase(p.ru_p)) return fail_msg("mpas_b200_advance_acoustic_step_work: the previous divergence damping did not run this substep's edge pass");
                done = true;
            }
            if (fold || next) g_fold[p.ru_p] = FoldEntry{p, fold, next};
        }
        if (!fold && !done) {
            KLAUNCH(acoustic_edge_kernel, (long long)p.nEdgesLoop * L, 256, p);
            KCHECK("acoustic_edge_kernel");
        }
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
        const size_t shm = (size_t)WPB * (small_step == 1 ? AC_NCOL - 3 : AC_NCOL) * tma_slot(L + 1) * sizeof(real);
        p.tend_ru_fold = X(tend_ru);
        static const int first_minb = tune("ac_first_minb", 5);
        if (fold && first_minb == 5) WLAUNCH((acoustic_cell_kernel<true, true, 5>), ncols, shm, p);
        else if (fold) WLAUNCH((acoustic_cell_kernel<true, true>), ncols, shm, p);
        else if (small_step == 1 && first_minb == 5) WLAUNCH((acoustic_cell_kernel<true, false, 5>), ncols, shm, p);
        else if (small_step == 1) WLAUNCH(acoustic_cell_kernel<true>, ncols, shm, p);
        else WLAUNCH(acoustic_cell_kernel<false>, ncols, shm, p);
        KCHECK("acoustic_cell_kernel");
    }
    return 0;
}

extern "C" int mpas_b200_divergence_damping_3d(MPAS_DIVDAMP_ARGS) {
    (void)nCells; (void)nEdges;
    DivDampP p;
    p.km.set(nVertLevels, nVertLevels + 1);
    const real rdts = (real)1.0 / dts;
    p.coef_divdamp = (real)2.0 * smdiv * config_len_disp * rdts;
    p.L = nVertLevels; p.nCellsSolve = nCellsSolve; p.edgeStart = edgeStart; p.nEdgesLoop = edgeEnd - edgeStart + 1;
    p.cellsOnEdge = X(cellsOnEdge); p.theta_m = X(theta_m); p.rtheta_pp = X(rtheta_pp); p.rtheta_pp_old = X(rtheta_pp_old);
    p.specZoneMaskEdge = X(specZoneMaskEdge); p.ru_p = X(ru_p);
    p.tend_ru = nullptr; p.ruAvg = nullptr; p.dts = 0; p.c2 = 0;
    p.rho_pp = p.zz = p.exner = p.cqu = p.zxu = p.invDcEdge = nullptr;
    bool fold = false, next = false;
    {
        std::lock_guard<std::mutex> lk(g_fold_mu);
        auto it = g_fold.find(p.ru_p);
        if (it != g_fold.end()) {
            const FoldEntry& f = it->second;
            if (f.ep.edgeStart != p.edgeStart || f.ep.nEdgesLoop != p.nEdgesLoop || f.ep.nCellsSolve != p.nCellsSolve)
                return fail_msg("mpas_b200_divergence_damping_3d: edge range differs from the folded acoustic step's");
            p.tend_ru = f.ep.tend_ru; p.ruAvg = f.ep.ruAvg; p.dts = f.ep.dts; p.c2 = f.ep.c2;
            p.rho_pp = f.ep.rho_pp; p.zz = f.ep.zz; p.exner = f.ep.exner; p.cqu = f.ep.cqu; p.zxu = f.ep.zxu; p.invDcEdge = f.ep.invDcEdge;
            fold = f.first;
            next = f.next;
            if (next) g_edge_done[p.ru_p] = 1;
            g_fold.erase(it);
        }
    }
    if (fold && next) KLAUNCH((divdamp_kernel<true, true>), (long long)p.nEdgesLoop * p.L, 256, p);
    else if (fold) KLAUNCH((divdamp_kernel<true, false>), (long long)p.nEdgesLoop * p.L, 256, p);
    else if (next) KLAUNCH((divdamp_kernel<false, true>), (long long)p.nEdgesLoop * p.L, 256, p);
    else KLAUNCH((divdamp_kernel<false, false>), (long long)p.nEdgesLoop * p.L, 256, p);
    KCHECK("divdamp_kernel");
    return 0;
}
