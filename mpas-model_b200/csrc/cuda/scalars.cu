// Scalar transport: atm_advance_scalars_work (TI:3238-3506) and the flux-corrected monotonic
// atm_advance_scalars_mono_work (TI:3612-4214).  TI = src/core_atmosphere/dynamics/mpas_atm_time_integration.F.
//
// scalars arrays are (num_scalars, nVertLevels, nCells+1) with the scalar index fastest (TI:3685).
//
// advance_scalars:   as_edge_kernel (q,k,edge)   high-order horizontal flux for every scalar   TI:3330-3387
//                    as_cell_kernel (q,k,cell)   flux divergence + vertical flux3 + update     TI:3400-3500
// mono, once:        mono_pre_kernel             scalars_old += dt*tend/rho, tend = 0          TI:3737-3752
//                    mono_rho_int_kernel         rho_zz_int                                    TI:3772-3792
//       per scalar:  mono_copy_kernel            2-D working copies                            TI:3800-3812
//                    mono_bounds_kernel          vertical hi-order flux wdtn, s_min/s_max      TI:3843-3908
//                    mono_edge_flux_kernel       flux_arr (hi-order), flux_upwind_tmp, flux_tmp TI:3916-3961, 4007-4022
//                    mono_limiter_kernel         upwind update + scale factors                 TI:3969-4000, 4024-4076
//                    [halo exchange of scale_arr, TI:4098]
//                    mono_rescale_edge_kernel    limited horizontal fluxes                     TI:4102-4138
//                    mono_update_kernel          limited vertical fluxes, final update, store  TI:4145-4210
// One thread per (k, element) everywhere; all neighbour sums are gathers in the reference's order.
#include "common.cuh"

using namespace mpasb200;

namespace {

const bool config_apply_lbcs = false;  // regional LBCs are out of scope

// x / ns for the small flattened (scalar, level) indices of this file without the ~20-instruction run-time division:
// magic = floor(2^32 / ns) + 1 gives the exact quotient for x * ns < 2^32 (ns = 1 is passed through).
struct FastDiv {
    unsigned d, magic;
    __host__ void set(int ns) { d = (unsigned)ns; magic = ns > 1 ? (unsigned)(0x100000000ULL / (unsigned)ns) + 1u : 0u; }
    __device__ __forceinline__ int operator()(int x) const { return d == 1 ? x : (int)__umulhi((unsigned)x, magic); }
};

struct AsP {
    FastDiv nsdiv;
    int L, mE, ns, edgeStart, nEdgesLoop, cellSolveStart, nCellsSolveLoop;
    real dt, coef_3rd_order, weight_time_old, weight_time_new;
    const real *scalar_old, *rho_zz_old, *rho_zz_new, *uhAvg, *wwAvg, *edgesOnCell_sign, *invAreaCell, *fnm, *fnp, *rdnw, *adv_coefs, *adv_coefs_3rd;
    const int *edgesOnCell, *nEdgesOnCell, *bdyMaskCell, *bdyMaskEdge, *nAdvCellsForEdge, *advCellsForEdge;
    const real* scalar_tend_save;
    real *scalar_new, *horiz_flux_arr;
};

// High-order horizontal flux of every scalar through one edge per warp (TI:3330-3387).  Lanes run over the flattened
// (scalar, level) index of the edge (num_scalars is the fastest dimension of the scalar arrays), two elements per
// lane; the stencil is staged once per warp as element offsets plus the two coefficient tables (c + c3, c - c3),
// padded with zero coefficients to a multiple of 5; each element selects its table by the sign of uhAvg at its level.
// The reference assigns the first of exactly 10 terms and otherwise adds the terms to zero (TI:3339-3375): kept.
__global__ void __launch_bounds__(WPB * 32) as_edge_kernel(AsP p) {
    __shared__ unsigned s_o[WPB][20];
    __shared__ real s_cp[WPB][20], s_cm[WPB][20];
    const int L = p.L, ns = p.ns, nel = L * ns;
    WARP_ENTITY(p.edgeStart, p.nEdgesLoop);
    const int iEdge = ent;
    if (config_apply_lbcs && !(p.bdyMaskEdge[iEdge - 1] < MPAS_NRELAXZONE - 1)) return;
    const int n = p.nAdvCellsForEdge[iEdge - 1];
    const int npad = n <= 5 ? 5 : ((n + 4) / 5 * 5);
    if (lane < npad) {
        const bool in = lane < n;
        const int c = in ? p.advCellsForEdge[IX2(lane + 1, iEdge, 15)] : p.advCellsForEdge[IX2(1, iEdge, 15)];
        const real a = in ? p.adv_coefs[IX2(lane + 1, iEdge, 15)] : (real)0.0, a3 = in ? p.adv_coefs_3rd[IX2(lane + 1, iEdge, 15)] : (real)0.0;
        s_o[warp][lane] = (unsigned)(max(c, 1) - 1) * (unsigned)nel;
        s_cp[warp][lane] = a + a3;
        s_cm[warp][lane] = a - a3;
    }
    __syncwarp();
    const bool assign_first = (n == 10);
    const real* __restrict__ uh = p.uhAvg + (size_t)(iEdge - 1) * L;
    real* __restrict__ out = p.horiz_flux_arr + (size_t)(iEdge - 1) * nel;
    for (int ea = lane; ea < nel; ea += 64) {
        const bool vb = ea + 32 < nel;
        const int bo = vb ? 32 : 0;
        const real* tab_a = signbit(uh[p.nsdiv(ea)]) ? s_cm[warp] : s_cp[warp];
        const real* tab_b = signbit(uh[p.nsdiv(ea + bo)]) ? s_cm[warp] : s_cp[warp];
        const real* __restrict__ sp = p.scalar_new + ea;
        real s_a = 0.0, s_b = 0.0;
        for (int j = 0; j < npad; j += 5) {
            real xa[5], xb[5];
#pragma unroll
            for (int t = 0; t < 5; ++t) {
                const real* __restrict__ q = elem_ptr(sp, s_o[warp][j + t]);
                xa[t] = q[0];
                xb[t] = q[bo];
            }
#pragma unroll
            for (int t = 0; t < 5; ++t) {
                const real ta = tab_a[j + t] * xa[t], tb = tab_b[j + t] * xb[t];
                if (t == 0 && j == 0 && assign_first) {
                    s_a = ta;
                    s_b = tb;
                } else {
                    s_a = s_a + ta;
                    s_b = s_b + tb;
                }
            }
        }
        out[ea] = s_a;
        if (vb) out[ea + 32] = s_b;
    }
}

enum { ASC_SCAL, ASC_WW, ASC_NTMA };

// Flux divergence, vertical 3rd-order flux and update of every scalar of one owned cell per warp (TI:3400-3500).
// The cell's scalar_new block (num_scalars x nVertLevels, contiguous) and its wwAvg column are staged by bulk (TMA)
// copies, so the update can be stored in place: every read of the pre-update column comes from shared memory (the
// reference fills wdtn for the whole column before it updates, TI:3436-3500).  The vertical fluxes are evaluated once
// per (scalar, interface) into shared memory; the edge sum is a direct gather, two elements per lane, the edge list
// staged as offsets and signs padded to 6 or 8 with zero signs (x - 0*y == x).
// dynamic shared memory per warp: tma_slot(ns*L) + tma_slot(L+1) + ns*(L+1) reals.
template <int MINB>
__global__ void __launch_bounds__(WPB * 32, MINB) as_cell_kernel(AsP p) {
    extern __shared__ __align__(16) real s_dyn[];
    __shared__ __align__(8) unsigned long long s_bar[WPB];
    __shared__ unsigned s_eL[WPB][8], s_eF[WPB][8];
    __shared__ real s_sign[WPB][8];
    const int L = p.L, L1 = L + 1, mE = p.mE, ns = p.ns, nel = L * ns;
    const int slotS = tma_slot(nel), slotW = tma_slot(L1), per_warp = slotS + slotW + ((ns * L1 + 3) & ~3);  // multiple of 16 bytes in either precision
    WARP_ENTITY(p.cellSolveStart, p.nCellsSolveLoop);
    const int iCell = ent;
    if (p.bdyMaskCell[iCell - 1] > MPAS_NRELAXZONE) return;
    real* sc = s_dyn + (size_t)warp * per_warp;
    const unsigned mbar = (unsigned)__cvta_generic_to_shared(&s_bar[warp]);
    const real* g_scal = p.scalar_new + (size_t)(iCell - 1) * nel;
    const real* g_ww = p.wwAvg + (size_t)(iCell - 1) * L1;
    if (elect_one()) {
        mbar_init(mbar, 1);
        const unsigned s0 = (unsigned)__cvta_generic_to_shared(sc);
        unsigned total = 0;
        tma_issue_col(s0, g_scal, nel, mbar, total);
        tma_issue_col(s0 + (unsigned)(slotS * sizeof(real)), g_ww, L1, mbar, total);
        mbar_arrive_expect_tx(mbar, total);
    }
    const real* s_scal = sc + tma_shift(g_scal);          // s_scal[(kk-1)*ns + q-1]
    const real* s_ww = sc + slotS + tma_shift(g_ww) - 1;  // s_ww[kk]
    real* wd = sc + slotS + slotW;                        // wd[(kk-1)*ns + q-1], kk = 1..L+1
    const int ne = p.nEdgesOnCell[iCell - 1];
    const int npad = ne <= 6 ? 6 : 8;
    if (lane < npad) {
        const bool in = lane < ne;
        const int iEdge = p.edgesOnCell[IX2(in ? lane + 1 : 1, iCell, mE)];
        s_eL[warp][lane] = (unsigned)(iEdge - 1) * (unsigned)L;
        s_eF[warp][lane] = (unsigned)(iEdge - 1) * (unsigned)nel;
        s_sign[warp][lane] = in ? p.edgesOnCell_sign[IX2(lane + 1, iCell, mE)] : (real)0.0;
    }
    const real invArea = p.invAreaCell[iCell - 1];
    __syncwarp();
    mbar_wait(mbar, 0);
    // ---- vertical flux at every (scalar, interface), TI:3436-3470, branch-free
    for (int t = lane; t < ns * L1; t += 32) {
        const int kk = p.nsdiv(t) + 1, q0 = t - (kk - 1) * ns;  // interface kk, scalar q0 (0-based)
        const int k = min(max(kk, 2), L);
        const int km2 = max(k - 2, 1), kp1 = min(k + 1, L);
        const real ww = s_ww[k];
        const real x_m2 = s_scal[(km2 - 1) * ns + q0], x_m1 = s_scal[(k - 2) * ns + q0], x_c = s_scal[(k - 1) * ns + q0], x_p1 = s_scal[(kp1 - 1) * ns + q0];
        const real f3 = flux3(x_m2, x_m1, x_c, x_p1, ww, p.coef_3rd_order);
        const real f2 = ww * (p.fnm[k - 1] * x_c + p.fnp[k - 1] * x_m1);
        wd[t] = (kk == 1 || kk == L1) ? (real)0.0 : ((k == 2 || k == L) ? f2 : f3);
    }
    __syncwarp();
    // ---- flux divergence and update, TI:3400-3432, 3474-3500
    const real* __restrict__ rho_old_col = p.rho_zz_old + (size_t)(iCell - 1) * L;
    const real* __restrict__ rho_new_col = p.rho_zz_new + (size_t)(iCell - 1) * L;
    const real* __restrict__ tend_col = p.scalar_tend_save + (size_t)(iCell - 1) * nel;
    const real* __restrict__ old_col = p.scalar_old + (size_t)(iCell - 1) * nel;
    real* __restrict__ new_col = p.scalar_new + (size_t)(iCell - 1) * nel;
    for (int ea = lane; ea < nel; ea += 64) {
        const bool vb = ea + 32 < nel;
        const int bo = vb ? 32 : 0;
        const int ka = p.nsdiv(ea), kb = p.nsdiv(ea + bo);  // 0-based levels
        real col_a = 0.0, col_b = 0.0;
        {
            const real* __restrict__ hf = p.horiz_flux_arr + ea;
            const real* __restrict__ uha = p.uhAvg + ka;
            const real* __restrict__ uhb = p.uhAvg + kb;
            real fa[8], fb[8], ua[8], ub[8];
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const real* __restrict__ q = elem_ptr(hf, s_eF[warp][i]);
                fa[i] = q[0];
                fb[i] = q[bo];
                ua[i] = *elem_ptr(uha, s_eL[warp][i]);
                ub[i] = *elem_ptr(uhb, s_eL[warp][i]);
            }
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const real s = s_sign[warp][i];
                col_a = col_a - s * ua[i] * fa[i];
                col_b = col_b - s * ub[i] * fb[i];
            }
            if (npad > 6) {
#pragma unroll
                for (int i = 6; i < 8; ++i) {
                    const real* __restrict__ q = elem_ptr(hf, s_eF[warp][i]);
                    const real s = s_sign[warp][i];
                    col_a = col_a - s * *elem_ptr(uha, s_eL[warp][i]) * q[0];
                    col_b = col_b - s * *elem_ptr(uhb, s_eL[warp][i]) * q[bo];
                }
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !vb) break;
            const int e = ea + 32 * h, k0 = h ? kb : ka;
            real col = h ? col_b : col_a;
            col = col * invArea + tend_col[e];
            const real flo = wd[e], fhi = wd[e + ns];
            const real rho_old = rho_old_col[k0];
            const real rho_zz_new_inv = (real)1.0 / (p.weight_time_old * rho_old + p.weight_time_new * rho_new_col[k0]);
            new_col[e] = (old_col[e] * rho_old + p.dt * (col - p.rdnw[k0] * (fhi - flo))) * rho_zz_new_inv;
        }
    }
}

// ---------------------------------------------------------------------------------------------
struct MonoP {
    KMagic km;
    int L, mE, ns, iScalar, nCells;
    int cellStart, nCellsLoop, cellSolveStart, nCellsSolveLoop, edgeStart, nEdgesLoop, nCellsSolve;
    int advance_density;
    real dt, coef_3rd_order;
    const real *uhAvg, *wwAvg, *rho_zz_old, *rho_lim, *invAreaCell, *dvEdge, *edgesOnCell_sign, *fnm, *fnp, *rdnw, *adv_coefs, *adv_coefs_3rd;
    const int *cellsOnEdge, *cellsOnCell, *edgesOnCell, *nEdgesOnCell, *nAdvCellsForEdge, *advCellsForEdge, *bdyMaskCell, *bdyMaskEdge;
    real *scalar_tend, *scalars_old, *scalars_new, *scalar_old, *scalar_new, *s_max, *s_min, *wdtn, *scale_arr, *flux_arr, *flux_upwind_tmp, *flux_tmp,
        *rho_zz_int;
};

__global__ void __launch_bounds__(256) mono_pre_kernel(MonoP p) {
    const int L = p.L, ns = p.ns;
    FLAT_KC(p.nCellsSolveLoop, L * ns);
    const int iCell = p.cellSolveStart + _c0;
    const int kk = (k - 1) / ns + 1, q = (k - 1) % ns + 1;
    const size_t is = IX3(q, kk, iCell, ns, L);
    p.scalars_old[is] = p.scalars_old[is] + p.dt * p.scalar_tend[is] / p.rho_zz_old[IX2(kk, iCell, L)];
    p.scalar_tend[is] = 0.0;
}

__global__ void __launch_bounds__(256) mono_rho_int_kernel(MonoP p) {
    const int L = p.L, L1 = L + 1, mE = p.mE;
    FLAT_KC(p.nCellsLoop, L);
    const int iCell = p.cellStart + _c0;
    const size_t ic = IX2(k, iCell, L);
    if (iCell < p.cellSolveStart || iCell >= p.cellSolveStart + p.nCellsSolveLoop) {
        p.rho_zz_int[ic] = 0.0;
        return;
    }
    real r = 0.0;
    const int ne = p.nEdgesOnCell[iCell - 1];
    const real invArea = p.invAreaCell[iCell - 1];
    for (int i = 1; i <= ne; ++i) {
        const int iEdge = p.edgesOnCell[IX2(i, iCell, mE)];
        r = r - p.edgesOnCell_sign[IX2(i, iCell, mE)] * p.uhAvg[IX2(k, iEdge, L)] * p.dvEdge[iEdge - 1] * invArea;
    }
    p.rho_zz_int[ic] = p.rho_zz_old[ic] + p.dt * (r - p.rdnw[k - 1] * (p.wwAvg[IX2(k + 1, iCell, L1)] - p.wwAvg[IX2(k, iCell, L1)]));
}

__global__ void __launch_bounds__(256) mono_copy_kernel(MonoP p) {
    const int L = p.L, ns = p.ns;
    FLAT_KC(p.nCellsLoop + 1, L);  // + the garbage column, zeroed (TI:3808-3812)
    const int iCell = p.cellStart + _c0;
    const size_t ic = IX2(k, iCell, L);
    if (iCell == p.nCells + 1) {
        p.scalar_old[ic] = 0.;
        p.scalar_new[ic] = 0.;
        return;
    }
    p.scalar_old[ic] = p.scalars_old[IX3(p.iScalar, k, iCell, ns, L)];
    p.scalar_new[ic] = p.scalars_new[IX3(p.iScalar, k, iCell, ns, L)];
}

__global__ void __launch_bounds__(256) mono_bounds_kernel(MonoP p) {
    const int L = p.L, L1 = L + 1, mE = p.mE;
    FLAT_KC(p.nCellsSolveLoop, L1);
    const int iCell = p.cellSolveStart + _c0;
    const size_t iw = IX2(k, iCell, L1);
    // vertical high-order flux at interface k
    real f;
    if (k == 1 || k == L1) f = 0.0;
    else if (k == 2 || k == L) f = p.wwAvg[iw] * (p.fnm[k - 1] * p.scalar_new[IX2(k, iCell, L)] + p.fnp[k - 1] * p.scalar_new[IX2(k - 1, iCell, L)]);
    else
        f = flux3(p.scalar_new[IX2(k - 2, iCell, L)], p.scalar_new[IX2(k - 1, iCell, L)], p.scalar_new[IX2(k, iCell, L)], p.scalar_new[IX2(k + 1, iCell, L)],
                  p.wwAvg[iw], p.coef_3rd_order);
    p.wdtn[iw] = f;
    if (k > L) return;
    const size_t ic = IX2(k, iCell, L);
    real smax, smin;
    const real s0 = p.scalar_old[ic];
    if (k == 1) {
        const real s1 = p.scalar_old[IX2(2, iCell, L)];
        smax = fmax(s0, s1);
        smin = fmin(s0, s1);
    } else if (k == L) {
        const real sm = p.scalar_old[IX2(k - 1, iCell, L)];
        smax = fmax(s0, sm);
        smin = fmin(s0, sm);
    } else {
        const real sm = p.scalar_old[IX2(k - 1, iCell, L)], sp = p.scalar_old[IX2(k + 1, iCell, L)];
        smax = fmax(fmax(sm, s0), sp);
        smin = fmin(fmin(sm, s0), sp);
    }
    const int ne = p.nEdgesOnCell[iCell - 1];
    for (int i = 1; i <= ne; ++i) {
        const real sc = p.scalar_old[IX2(k, p.cellsOnCell[IX2(i, iCell, mE)], L)];
        smax = fmax(smax, sc);
        smin = fmin(smin, sc);
    }
    p.s_max[ic] = smax;
    p.s_min[ic] = smin;
}

// High-order and upwind horizontal fluxes of the current scalar through one edge per warp (TI:3916-3961, 4007-4022).
// Two levels per lane; for the usual 10-cell stencil the offsets and the two pre-combined coefficient tables are staged
// once per warp and every level selects its table by uhAvg > 0 (TI:3925-3945); other stencil sizes take the reference's
// general form (TI:3947-3957) in a plain loop.
__global__ void __launch_bounds__(WPB * 32) mono_edge_flux_kernel(MonoP p) {
    __shared__ unsigned s_o[WPB][10];
    __shared__ real s_cp[WPB][10], s_cm[WPB][10];
    const int L = p.L;
    WARP_ENTITY(p.edgeStart, p.nEdgesLoop);
    const int iEdge = ent;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    const bool owned = (cell1 <= p.nCellsSolve || cell2 <= p.nCellsSolve);
    const int n = owned ? p.nAdvCellsForEdge[iEdge - 1] : 0;
    if (n == 10 && lane < 10) {
        const real c = p.adv_coefs[IX2(lane + 1, iEdge, 15)], c3 = p.adv_coefs_3rd[IX2(lane + 1, iEdge, 15)];
        s_o[warp][lane] = (unsigned)(p.advCellsForEdge[IX2(lane + 1, iEdge, 15)] - 1) * (unsigned)L;
        s_cp[warp][lane] = c + c3;
        s_cm[warp][lane] = c - c3;
    }
    __syncwarp();
    const real dvdt = p.dvEdge[iEdge - 1] * p.dt;
    const int bm = p.bdyMaskEdge[iEdge - 1];
    const bool zero_corr = (config_apply_lbcs && (bm == MPAS_NRELAXZONE)) || (bm == MPAS_NRELAXZONE - 1);
    const unsigned eoff = (unsigned)(iEdge - 1) * (unsigned)L, c1off = (unsigned)(cell1 - 1) * (unsigned)L, c2off = (unsigned)(cell2 - 1) * (unsigned)L;
    for (int ka = lane; ka < L; ka += 64) {
        const bool vb = ka + 32 < L;
        const int bo = vb ? 32 : 0;
        const real* __restrict__ uhp = elem_ptr(p.uhAvg + ka, eoff);
        const real uh_a = uhp[0], uh_b = uhp[bo];
        const real* __restrict__ o1 = elem_ptr(p.scalar_old + ka, c1off);
        const real* __restrict__ o2 = elem_ptr(p.scalar_old + ka, c2off);
        const real so1_a = o1[0], so1_b = o1[bo], so2_a = o2[0], so2_b = o2[bo];
        real fa_a = 0.0, fa_b = 0.0;
        if (n == 10) {
            const real* tab_a = (uh_a > 0) ? s_cp[warp] : s_cm[warp];
            const real* tab_b = (uh_b > 0) ? s_cp[warp] : s_cm[warp];
            const real* __restrict__ sp = p.scalar_new + ka;
            real s_a = 0.0, s_b = 0.0;
#pragma unroll
            for (int j = 0; j < 10; j += 5) {
                real xa[5], xb[5];
#pragma unroll
                for (int t = 0; t < 5; ++t) {
                    const real* __restrict__ q = elem_ptr(sp, s_o[warp][j + t]);
                    xa[t] = q[0];
                    xb[t] = q[bo];
                }
#pragma unroll
                for (int t = 0; t < 5; ++t) {
                    const real ta = tab_a[j + t] * xa[t], tb = tab_b[j + t] * xb[t];
                    if (j == 0 && t == 0) {  // the first term is assigned (TI:3928), not added to zero
                        s_a = ta;
                        s_b = tb;
                    } else {
                        s_a = s_a + ta;
                        s_b = s_b + tb;
                    }
                }
            }
            fa_a = uh_a * s_a;
            fa_b = uh_b * s_b;
        } else if (n > 0) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int k = ka + 1 + (h ? bo : 0);
                const real uh = h ? uh_b : uh_a;
                const real s3 = fsign((real)1.0, uh);
                real fa = 0.0;
                for (int i = 1; i <= n; ++i) {
                    const real scalar_weight = uh * (p.adv_coefs[IX2(i, iEdge, 15)] + s3 * p.adv_coefs_3rd[IX2(i, iEdge, 15)]);
                    fa = fa + scalar_weight * p.scalar_new[IX2(k, p.advCellsForEdge[IX2(i, iEdge, 15)], L)];
                }
                if (h) fa_b = fa;
                else fa_a = fa;
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !vb) break;
            const real uh = h ? uh_b : uh_a;
            real fa = h ? fa_b : fa_a;
            // TI:4007-4022 (all edges)
            const real fup = dvdt * (fmax((real)0.0, uh) * (h ? so1_b : so1_a) + fmin((real)0.0, uh) * (h ? so2_b : so2_a));
            real ft = p.dt * fa - fup;
            if (zero_corr) {
                ft = 0.;
                fa = fup;
            }
            const size_t ie = (size_t)eoff + ka + 32 * h;
            p.flux_arr[ie] = fa;
            p.flux_upwind_tmp[ie] = fup;
            p.flux_tmp[ie] = ft;
        }
    }
}

__device__ __forceinline__ real mono_upwind_at(const MonoP& p, int kk, int iCell) {  // flux_upwind_arr(kk), kk in 2..L
    const int L = p.L, L1 = L + 1;
    const real ww = p.wwAvg[IX2(kk, iCell, L1)];
    return p.dt * (fmax((real)0.0, ww) * p.scalar_old[IX2(kk - 1, iCell, L)] + fmin((real)0.0, ww) * p.scalar_old[IX2(kk, iCell, L)]);
}

// vertical upwind update (TI:3969-4000) + horizontal accumulation + scale factors (TI:4024-4076).
// The reference rewrites wdtn(k) = dt*wdtn(k) - flux_upwind(k) in place; thread k needs the rewritten
// wdtn(k) and wdtn(k+1): both are recomputed in registers from the high-order values still in memory,
// and the in-place rewrite itself is done afterwards by mono_wdtn_store_kernel.
__global__ void __launch_bounds__(256) mono_limiter_kernel(MonoP p) {
    const int L = p.L, L1 = L + 1, mE = p.mE;
    const real eps = 1.e-20;
    {
        FLAT_KC(p.nCellsSolveLoop, L);
        const int iCell = p.cellSolveStart + _c0;
        const size_t ic = IX2(k, iCell, L);
        const size_t iw = IX2(k, iCell, L1);
        const real so = p.scalar_old[ic];
        real sn = so * p.rho_zz_old[ic];
        const real fu_k = (k >= 2) ? mono_upwind_at(p, k, iCell) : (real)0.0;
        const real fu_kp = (k <= L - 1) ? mono_upwind_at(p, k + 1, iCell) : (real)0.0;
        const real rd = p.rdnw[k - 1];
        if (k <= L - 1) sn = sn - fu_kp * rd;
        if (k >= 2) sn = sn + fu_k * rd;
        const real wd_k = (k >= 2) ? (p.dt * p.wdtn[iw] - fu_k) : p.wdtn[iw];
        const real wd_kp = (k <= L - 1) ? (p.dt * p.wdtn[IX2(k + 1, iCell, L1)] - fu_kp) : p.wdtn[IX2(k + 1, iCell, L1)];
        real sc_in = -rd * (fmin((real)0.0, wd_kp) - fmax((real)0.0, wd_k));
        real sc_out = -rd * (fmax((real)0.0, wd_kp) - fmin((real)0.0, wd_k));
        const int ne = p.nEdgesOnCell[iCell - 1];
        const real invArea = p.invAreaCell[iCell - 1];
        for (int i = 1; i <= ne; ++i) {
            const int iEdge = p.edgesOnCell[IX2(i, iCell, mE)];
            const real sgn = p.edgesOnCell_sign[IX2(i, iCell, mE)];
            const size_t ie = IX2(k, iEdge, L);
            sn = sn - sgn * p.flux_upwind_tmp[ie] * invArea;
            const real sf = sgn * p.flux_tmp[ie];
            sc_out = sc_out - fmax((real)0.0, sf) * invArea;
            sc_in = sc_in - fmin((real)0.0, sf) * invArea;
        }
        p.scalar_new[ic] = sn;
        const real rl = p.rho_lim[ic];
        real scale_factor = (p.s_max[ic] * rl - sn) / (sc_in + eps);
        p.scale_arr[IX3(k, 1, iCell, L, 2)] = fmin((real)1.0, fmax((real)0.0, scale_factor));
        scale_factor = (p.s_min[ic] * rl - sn) / (sc_out - eps);
        p.scale_arr[IX3(k, 2, iCell, L, 2)] = fmin((real)1.0, fmax((real)0.0, scale_factor));
    }
}

// second half of TI:3990-3993: wdtn(k) = dt*wdtn(k) - flux_upwind_arr(k), k = 2..L, run after mono_limiter_kernel
__global__ void __launch_bounds__(256) mono_wdtn_store_kernel(MonoP p) {
    const int L = p.L, L1 = L + 1;
    FLAT_KC(p.nCellsSolveLoop, L);
    if (k < 2) return;
    const int iCell = p.cellSolveStart + _c0;
    const size_t iw = IX2(k, iCell, L1);
    p.wdtn[iw] = p.dt * p.wdtn[iw] - mono_upwind_at(p, k, iCell);
}

__global__ void __launch_bounds__(256) mono_rescale_edge_kernel(MonoP p) {
    const int L = p.L;
    FLAT_KC(p.nEdgesLoop, L);
    const int iEdge = p.edgeStart + _c0;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    if (!(cell1 <= p.nCellsSolve || cell2 <= p.nCellsSolve)) return;
    const size_t ie = IX2(k, iEdge, L);
    const real uh = p.uhAvg[ie];
    const real flux_upwind = p.dvEdge[iEdge - 1] * p.dt * (fmax((real)0.0, uh) * p.scalar_old[IX2(k, cell1, L)] + fmin((real)0.0, uh) * p.scalar_old[IX2(k, cell2, L)]);
    real flux = p.dt * p.flux_arr[ie] - flux_upwind;
    const int bm = p.bdyMaskEdge[iEdge - 1];
    if ((config_apply_lbcs && (bm == MPAS_NRELAXZONE)) || (bm == MPAS_NRELAXZONE - 1)) flux = 0.;
    flux = fmax((real)0.0, flux) * fmin(p.scale_arr[IX3(k, 2, cell1, L, 2)], p.scale_arr[IX3(k, 1, cell2, L, 2)]) +
           fmin((real)0.0, flux) * fmin(p.scale_arr[IX3(k, 1, cell1, L, 2)], p.scale_arr[IX3(k, 2, cell2, L, 2)]);
    p.flux_arr[ie] = flux;
}

__device__ __forceinline__ real mono_wdtn_limited(const MonoP& p, int kk, int iCell) {  // TI:4145-4153
    const int L = p.L, L1 = L + 1;
    const real flux = p.wdtn[IX2(kk, iCell, L1)];
    if (kk < 2 || kk > L) return flux;
    return fmax((real)0.0, flux) * fmin(p.scale_arr[IX3(kk - 1, 2, iCell, L, 2)], p.scale_arr[IX3(kk, 1, iCell, L, 2)]) +
           fmin((real)0.0, flux) * fmin(p.scale_arr[IX3(kk, 2, iCell, L, 2)], p.scale_arr[IX3(kk - 1, 1, iCell, L, 2)]);
}

__global__ void __launch_bounds__(256) mono_update_kernel(MonoP p) {
    const int L = p.L, mE = p.mE, ns = p.ns;
    FLAT_KC(p.nCellsLoop, L);
    const int iCell = p.cellStart + _c0;
    const size_t ic = IX2(k, iCell, L);
    real sn = p.scalar_new[ic];
    if (iCell >= p.cellSolveStart && iCell < p.cellSolveStart + p.nCellsSolveLoop) {
        const real wlo = mono_wdtn_limited(p, k, iCell), whi = mono_wdtn_limited(p, k + 1, iCell);
        const int ne = p.nEdgesOnCell[iCell - 1];
        const real invArea = p.invAreaCell[iCell - 1];
        for (int i = 1; i <= ne; ++i) {
            const int iEdge = p.edgesOnCell[IX2(i, iCell, mE)];
            sn = sn - p.edgesOnCell_sign[IX2(i, iCell, mE)] * p.flux_arr[IX2(k, iEdge, L)] * invArea;
        }
        sn = (sn + (-p.rdnw[k - 1] * (whi - wlo))) / p.rho_lim[ic];
        p.scalar_new[ic] = sn;
    }
    if (p.bdyMaskCell[iCell - 1] <= MPAS_NSPECZONE) p.scalars_new[IX3(p.iScalar, k, iCell, ns, L)] = fmax((real)0.0, sn);
}

// ---------------------------------------------------------------------------------------------
// All scalars in one pass (the default path of mpas_b200_advance_scalars_mono_work).  The reference loops over the
// scalars (TI:3798-4212) with 2-D working copies; nothing in that loop couples two scalars, so the same arithmetic is
// done here on the 3-D (num_scalars, nVertLevels, .) layout with the flattened (scalar, level) index e = (k-1)*ns + q-1
// along the lanes: no working copies, one read of the stencil metadata for all scalars, 4 launches and ONE scale_arr
// exchange instead of 7 and 1 per scalar.  Scratch lives in the library's per-block workspace:
//   w_fa, w_fup (ns*L, nEdges+1)   high-order / upwind horizontal flux        (flux_arr, flux_upwind_tmp; flux_tmp is recomputed)
//   w_wdtn      (ns*(L+1), nCells+1)  dt*wdtn - flux_upwind_arr               (wdtn after TI:3990-3993)
//   w_snew      (ns*L, nCells+1)   upwind-updated scalar_new                  (TI:3969-4039)
//   w_scale     (2*ns*L, nCells+1) [scale_in | scale_out] per cell            (scale_arr)
// s_max / s_min stay in registers.  Every expression and every summation order is the reference's.
struct MonoNP {
    KMagic km;
    FastDiv nsdiv;
    int L, mE, ns, nCells;
    int cellStart, nCellsLoop, cellSolveStart, nCellsSolveLoop, edgeStart, nEdgesLoop, nCellsSolve;
    real dt, coef_3rd_order;
    const real *uhAvg, *wwAvg, *rho_zz_old, *rho_lim, *invAreaCell, *dvEdge, *edgesOnCell_sign, *fnm, *fnp, *rdnw, *adv_coefs, *adv_coefs_3rd;
    const int *cellsOnEdge, *cellsOnCell, *edgesOnCell, *nEdgesOnCell, *nAdvCellsForEdge, *advCellsForEdge, *bdyMaskCell, *bdyMaskEdge;
    const real* scalars_old;
    real *scalars_new, *w_fa, *w_fup, *w_wdtn, *w_snew, *w_scale;
};

__device__ __forceinline__ bool mono_zero_corr(int bm) { return (config_apply_lbcs && (bm == MPAS_NRELAXZONE)) || (bm == MPAS_NRELAXZONE - 1); }

// TI:3916-3961 + 4007-4022 for every scalar of one edge per warp: w_fa = high-order flux (0 off the owned edges),
// w_fup = upwind flux (all edges).  Two (scalar, level) elements per lane, 10-cell stencil staged as in mono_edge_flux_kernel.
template <int MINB>
__global__ void __launch_bounds__(WPB * 32, MINB) monoN_edge_flux_kernel(MonoNP p) {
    __shared__ unsigned s_o[WPB][10];
    __shared__ real s_cp[WPB][10], s_cm[WPB][10];
    const int L = p.L, ns = p.ns, nel = L * ns;
    WARP_ENTITY(p.edgeStart, p.nEdgesLoop);
    const int iEdge = ent;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    const bool owned = (cell1 <= p.nCellsSolve || cell2 <= p.nCellsSolve);
    const int n = owned ? p.nAdvCellsForEdge[iEdge - 1] : 0;
    if (n == 10 && lane < 10) {
        const real c = p.adv_coefs[IX2(lane + 1, iEdge, 15)], c3 = p.adv_coefs_3rd[IX2(lane + 1, iEdge, 15)];
        s_o[warp][lane] = (unsigned)(p.advCellsForEdge[IX2(lane + 1, iEdge, 15)] - 1) * (unsigned)nel;
        s_cp[warp][lane] = c + c3;
        s_cm[warp][lane] = c - c3;
    }
    __syncwarp();
    const real dvdt = p.dvEdge[iEdge - 1] * p.dt;
    const bool zero_corr = mono_zero_corr(p.bdyMaskEdge[iEdge - 1]);
    const real* __restrict__ uh = p.uhAvg + (size_t)(iEdge - 1) * L;
    const real* __restrict__ o1 = p.scalars_old + (size_t)(cell1 - 1) * nel;
    const real* __restrict__ o2 = p.scalars_old + (size_t)(cell2 - 1) * nel;
    real* __restrict__ out_fa = p.w_fa + (size_t)(iEdge - 1) * nel;
    real* __restrict__ out_fup = p.w_fup + (size_t)(iEdge - 1) * nel;
    for (int ea = lane; ea < nel; ea += 64) {
        const bool vb = ea + 32 < nel;
        const int bo = vb ? 32 : 0;
        const real uh_a = uh[p.nsdiv(ea)], uh_b = uh[p.nsdiv(ea + bo)];
        const real so1_a = o1[ea], so1_b = o1[ea + bo], so2_a = o2[ea], so2_b = o2[ea + bo];
        real fa_a = 0.0, fa_b = 0.0;
        if (n == 10) {
            const real* tab_a = (uh_a > 0) ? s_cp[warp] : s_cm[warp];
            const real* tab_b = (uh_b > 0) ? s_cp[warp] : s_cm[warp];
            const real* __restrict__ sp = p.scalars_new + ea;
            real s_a = 0.0, s_b = 0.0;
#pragma unroll
            for (int j = 0; j < 10; j += 5) {
                real xa[5], xb[5];
#pragma unroll
                for (int t = 0; t < 5; ++t) {
                    const real* __restrict__ q = elem_ptr(sp, s_o[warp][j + t]);
                    xa[t] = q[0];
                    xb[t] = q[bo];
                }
#pragma unroll
                for (int t = 0; t < 5; ++t) {
                    const real ta = tab_a[j + t] * xa[t], tb = tab_b[j + t] * xb[t];
                    if (j == 0 && t == 0) {  // the first term is assigned (TI:3928), not added to zero
                        s_a = ta;
                        s_b = tb;
                    } else {
                        s_a = s_a + ta;
                        s_b = s_b + tb;
                    }
                }
            }
            fa_a = uh_a * s_a;
            fa_b = uh_b * s_b;
        } else if (n > 0) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int e = ea + (h ? bo : 0);
                const real uhv = h ? uh_b : uh_a;
                const real s3 = fsign((real)1.0, uhv);
                real fa = 0.0;
                for (int i = 1; i <= n; ++i) {
                    const real scalar_weight = uhv * (p.adv_coefs[IX2(i, iEdge, 15)] + s3 * p.adv_coefs_3rd[IX2(i, iEdge, 15)]);
                    fa = fa + scalar_weight * p.scalars_new[(size_t)(p.advCellsForEdge[IX2(i, iEdge, 15)] - 1) * nel + e];
                }
                if (h) fa_b = fa;
                else fa_a = fa;
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !vb) break;
            const real uhv = h ? uh_b : uh_a;
            real fa = h ? fa_b : fa_a;
            const real fup = dvdt * (fmax((real)0.0, uhv) * (h ? so1_b : so1_a) + fmin((real)0.0, uhv) * (h ? so2_b : so2_a));
            if (zero_corr) fa = fup;  // TI:4017-4021
            out_fa[ea + 32 * h] = fa;
            out_fup[ea + 32 * h] = fup;
        }
    }
}

// One horizontal batch of the limiter kernel: CNT edges starting at table slot I0, both elements of the lane.
// All loads first, then the reference's three accumulations per edge (TI:4024-4039) with flux_tmp = dt*fa - fup
// recomputed (TI:4012); s_sT carries a zero where the reference zeroes flux_tmp (TI:4017-4021).
#define MONO_EDGE_BATCH(I0, CNT)                                                        \
    {                                                                                   \
        real xa[CNT], xb[CNT], ya[CNT], yb[CNT];                                        \
        _Pragma("unroll") for (int t = 0; t < CNT; ++t) {                               \
            const real* __restrict__ qf = elem_ptr(fa_base, s_eF[warp][(I0) + t]);      \
            const real* __restrict__ qu = elem_ptr(fup_base, s_eF[warp][(I0) + t]);     \
            xa[t] = qf[0];                                                              \
            xb[t] = qf[bo];                                                             \
            ya[t] = qu[0];                                                              \
            yb[t] = qu[bo];                                                             \
        }                                                                               \
        _Pragma("unroll") for (int t = 0; t < CNT; ++t) {                               \
            const real sg = s_sign[warp][(I0) + t], sT = s_sT[warp][(I0) + t];          \
            sn_a = sn_a - sg * ya[t] * invArea;                                         \
            sn_b = sn_b - sg * yb[t] * invArea;                                         \
            const real sfa = sT * (p.dt * xa[t] - ya[t]), sfb = sT * (p.dt * xb[t] - yb[t]); \
            so_a = so_a - fmax((real)0.0, sfa) * invArea;                               \
            so_b = so_b - fmax((real)0.0, sfb) * invArea;                               \
            si_a = si_a - fmin((real)0.0, sfa) * invArea;                               \
            si_b = si_b - fmin((real)0.0, sfb) * invArea;                               \
        }                                                                               \
    }
#define MONO_NBR_BATCH(I0, CNT)                                                         \
    {                                                                                   \
        real xa[CNT], xb[CNT];                                                          \
        _Pragma("unroll") for (int t = 0; t < CNT; ++t) {                               \
            const real* __restrict__ q = elem_ptr(old_base, s_cF[warp][(I0) + t]);      \
            xa[t] = q[0];                                                               \
            xb[t] = q[bo];                                                              \
        }                                                                               \
        _Pragma("unroll") for (int t = 0; t < CNT; ++t) {                               \
            mx_a = fmax(mx_a, xa[t]);                                                   \
            mn_a = fmin(mn_a, xa[t]);                                                   \
            mx_b = fmax(mx_b, xb[t]);                                                   \
            mn_b = fmin(mn_b, xb[t]);                                                   \
        }                                                                               \
    }

// TI:3843-3908 (vertical high-order flux, bounds), 3969-4000 (vertical upwind update), 4024-4076 (horizontal
// accumulation, scale factors) for every scalar of one owned cell per warp.  The cell's scalars_new / scalars_old blocks
// and its wwAvg column are staged by bulk (TMA) copies; the rewritten vertical flux dt*wdtn - flux_upwind_arr goes to
// shared memory (and to w_wdtn for the final update); the edge sums and the cellsOnCell bounds are direct gathers with the
// edge / neighbour lists staged as offsets, padded to 6 or 8 with zero signs (x - 0*y == x) and the cell itself
// (max/min are idempotent).
// dynamic shared memory per warp: 2*tma_slot(ns*L) + tma_slot(L+1) + ns*(L+1) reals.
template <int MINB>
__global__ void __launch_bounds__(WPB * 32, MINB) monoN_cell_kernel(MonoNP p) {
    extern __shared__ __align__(16) real s_dyn[];
    __shared__ __align__(8) unsigned long long s_bar[WPB];
    __shared__ unsigned s_eF[WPB][8], s_cF[WPB][8];
    __shared__ real s_sign[WPB][8], s_sT[WPB][8];
    const int L = p.L, L1 = L + 1, mE = p.mE, ns = p.ns, nel = L * ns, nel1 = ns * L1;
    const int slotS = tma_slot(nel), slotW = tma_slot(L1), per_warp = 2 * slotS + slotW + ((nel1 + 3) & ~3);
    const real eps = 1.e-20;
    WARP_ENTITY(p.cellSolveStart, p.nCellsSolveLoop);
    const int iCell = ent;
    real* sc = s_dyn + (size_t)warp * per_warp;
    const unsigned mbar = (unsigned)__cvta_generic_to_shared(&s_bar[warp]);
    const real* g_new = p.scalars_new + (size_t)(iCell - 1) * nel;
    const real* g_old = p.scalars_old + (size_t)(iCell - 1) * nel;
    const real* g_ww = p.wwAvg + (size_t)(iCell - 1) * L1;
    if (elect_one()) {
        mbar_init(mbar, 1);
        const unsigned s0 = (unsigned)__cvta_generic_to_shared(sc);
        unsigned total = 0;
        tma_issue_col(s0, g_new, nel, mbar, total);
        tma_issue_col(s0 + (unsigned)(slotS * sizeof(real)), g_old, nel, mbar, total);
        tma_issue_col(s0 + (unsigned)(2 * slotS * sizeof(real)), g_ww, L1, mbar, total);
        mbar_arrive_expect_tx(mbar, total);
    }
    const real* s_new = sc + tma_shift(g_new);                  // s_new[(k-1)*ns + q-1]
    const real* s_old = sc + slotS + tma_shift(g_old);          // s_old[(k-1)*ns + q-1]
    const real* s_ww = sc + 2 * slotS + tma_shift(g_ww) - 1;    // s_ww[kk], kk = 1..L+1
    real* wd = sc + 2 * slotS + slotW;                          // wd[(kk-1)*ns + q-1], kk = 1..L+1
    const int ne = p.nEdgesOnCell[iCell - 1];
    const int npad = ne <= 6 ? 6 : 8;
    if (lane < npad) {
        const bool in = lane < ne;
        const int iEdge = p.edgesOnCell[IX2(in ? lane + 1 : 1, iCell, mE)];
        const int cc = in ? p.cellsOnCell[IX2(lane + 1, iCell, mE)] : iCell;
        s_eF[warp][lane] = (unsigned)(iEdge - 1) * (unsigned)nel;
        s_cF[warp][lane] = (unsigned)(cc - 1) * (unsigned)nel;
        const real sg = in ? p.edgesOnCell_sign[IX2(lane + 1, iCell, mE)] : (real)0.0;
        s_sign[warp][lane] = sg;
        s_sT[warp][lane] = mono_zero_corr(p.bdyMaskEdge[iEdge - 1]) ? (real)0.0 : sg;
    }
    const real invArea = p.invAreaCell[iCell - 1];
    __syncwarp();
    mbar_wait(mbar, 0);
    // ---- vertical high-order flux at every (scalar, interface) (TI:3843-3876, branch-free) and its rewrite
    //      wdtn(k) = dt*wdtn(k) - flux_upwind_arr(k), k = 2..L (TI:3977-3993); interfaces 1 and L+1 stay 0
    real* __restrict__ g_wd = p.w_wdtn + (size_t)(iCell - 1) * nel1;
    for (int t = lane; t < nel1; t += 32) {
        const int kk = p.nsdiv(t) + 1, q0 = t - (kk - 1) * ns;
        const int k = min(max(kk, 2), L);
        const int km2 = max(k - 2, 1), kp1 = min(k + 1, L);
        const real ww = s_ww[k];
        const real x_m2 = s_new[(km2 - 1) * ns + q0], x_m1 = s_new[(k - 2) * ns + q0], x_c = s_new[(k - 1) * ns + q0], x_p1 = s_new[(kp1 - 1) * ns + q0];
        const real f3 = flux3(x_m2, x_m1, x_c, x_p1, ww, p.coef_3rd_order);
        const real f2 = ww * (p.fnm[k - 1] * x_c + p.fnp[k - 1] * x_m1);
        const real f = (k == 2 || k == L) ? f2 : f3;
        const real fu = p.dt * (fmax((real)0.0, ww) * s_old[(k - 2) * ns + q0] + fmin((real)0.0, ww) * s_old[(k - 1) * ns + q0]);
        const real w = (kk == 1 || kk == L1) ? (real)0.0 : (p.dt * f - fu);
        wd[t] = w;
        g_wd[t] = w;
    }
    __syncwarp();
    const real* __restrict__ rho_old_col = p.rho_zz_old + (size_t)(iCell - 1) * L;
    const real* __restrict__ rho_lim_col = p.rho_lim + (size_t)(iCell - 1) * L;
    real* __restrict__ snew_col = p.w_snew + (size_t)(iCell - 1) * nel;
    real* __restrict__ scale_col = p.w_scale + (size_t)(iCell - 1) * 2 * nel;
    for (int ea = lane; ea < nel; ea += 64) {
        const bool vb = ea + 32 < nel;
        const int bo = vb ? 32 : 0;
        const int ka = p.nsdiv(ea), kb = p.nsdiv(ea + bo);  // 0-based levels
        real sn_a, sn_b, si_a, si_b, so_a, so_b, mx_a, mx_b, mn_a, mn_b;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int e = ea + (h ? bo : 0), k0 = h ? kb : ka, q0 = e - k0 * ns;
            const real so = s_old[e];
            const real so_m = s_old[max(e - ns, q0)], so_p = s_old[min(e + ns, nel - ns + q0)];
            const real ww_k = s_ww[k0 + 1], ww_kp = s_ww[k0 + 2];
            const real rd = p.rdnw[k0];
            real sn = so * rho_old_col[k0];
            const real fu_k = p.dt * (fmax((real)0.0, ww_k) * so_m + fmin((real)0.0, ww_k) * so);     // flux_upwind_arr(k),   k >= 2
            const real fu_kp = p.dt * (fmax((real)0.0, ww_kp) * so + fmin((real)0.0, ww_kp) * so_p);  // flux_upwind_arr(k+1), k <= L-1
            if (k0 <= L - 2) sn = sn - fu_kp * rd;
            if (k0 >= 1) sn = sn + fu_k * rd;
            const real wd_k = wd[e], wd_kp = wd[e + ns];
            const real si = -rd * (fmin((real)0.0, wd_kp) - fmax((real)0.0, wd_k));
            const real sout = -rd * (fmax((real)0.0, wd_kp) - fmin((real)0.0, wd_k));
            real mx, mn;
            if (k0 == 0) {
                mx = fmax(so, so_p);
                mn = fmin(so, so_p);
            } else if (k0 == L - 1) {
                mx = fmax(so, so_m);
                mn = fmin(so, so_m);
            } else {
                mx = fmax(fmax(so_m, so), so_p);
                mn = fmin(fmin(so_m, so), so_p);
            }
            if (h) { sn_b = sn; si_b = si; so_b = sout; mx_b = mx; mn_b = mn; }
            else { sn_a = sn; si_a = si; so_a = sout; mx_a = mx; mn_a = mn; }
        }
        const real* __restrict__ fa_base = p.w_fa + ea;
        const real* __restrict__ fup_base = p.w_fup + ea;
        const real* __restrict__ old_base = p.scalars_old + ea;
        MONO_EDGE_BATCH(0, 3)
        MONO_EDGE_BATCH(3, 3)
        if (npad > 6) MONO_EDGE_BATCH(6, 2)
        MONO_NBR_BATCH(0, 6)
        if (npad > 6) MONO_NBR_BATCH(6, 2)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !vb) break;
            const int e = ea + 32 * h, k0 = h ? kb : ka;
            const real sn = h ? sn_b : sn_a, rl = rho_lim_col[k0];
            snew_col[e] = sn;
            real scale_factor = ((h ? mx_b : mx_a) * rl - sn) / ((h ? si_b : si_a) + eps);
            scale_col[e] = fmin((real)1.0, fmax((real)0.0, scale_factor));
            scale_factor = ((h ? mn_b : mn_a) * rl - sn) / ((h ? so_b : so_a) - eps);
            scale_col[nel + e] = fmin((real)1.0, fmax((real)0.0, scale_factor));
        }
    }
}
#undef MONO_EDGE_BATCH
#undef MONO_NBR_BATCH

// TI:4102-4138 for every scalar: limited perturbation flux on the edges of owned cells, one thread per (scalar, level, edge).
// The upwind flux is the value stored by monoN_edge_flux_kernel (the reference recomputes the same expression).
__global__ void __launch_bounds__(256) monoN_rescale_edge_kernel(MonoNP p) {
    const int nel = p.L * p.ns;
    FLAT_KC(p.nEdgesLoop, nel);
    const int iEdge = p.edgeStart + _c0, e = k - 1;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    if (!(cell1 <= p.nCellsSolve || cell2 <= p.nCellsSolve)) return;
    const size_t ie = (size_t)(iEdge - 1) * nel + e;
    real flux = p.dt * p.w_fa[ie] - p.w_fup[ie];
    if (mono_zero_corr(p.bdyMaskEdge[iEdge - 1])) flux = 0.;
    const real* __restrict__ s1 = p.w_scale + (size_t)(cell1 - 1) * 2 * nel + e;  // [0] scale_in, [nel] scale_out
    const real* __restrict__ s2 = p.w_scale + (size_t)(cell2 - 1) * 2 * nel + e;
    flux = fmax((real)0.0, flux) * fmin(s1[nel], s2[0]) + fmin((real)0.0, flux) * fmin(s1[0], s2[nel]);
    p.w_fa[ie] = flux;
}

// TI:4145-4210 for every scalar: limited vertical fluxes, final update of the owned cells, store (clamped at zero) for
// all cells -- cells outside the owned range keep their predictor value, as the reference's scalar_new copy does.
__global__ void __launch_bounds__(256, 5) monoN_update_kernel(MonoNP p) {
    const int L = p.L, mE = p.mE, ns = p.ns, nel = L * ns;
    FLAT_KC(p.nCellsLoop, nel);
    const int iCell = p.cellStart + _c0, e = k - 1, k0 = p.nsdiv(e);
    const size_t is = (size_t)(iCell - 1) * nel + e;
    real sn;
    if (iCell >= p.cellSolveStart && iCell < p.cellSolveStart + p.nCellsSolveLoop) {
        sn = p.w_snew[is];
        const real* __restrict__ sc = p.w_scale + (size_t)(iCell - 1) * 2 * nel;  // [e'] scale_in, [nel + e'] scale_out
        const real* __restrict__ wdc = p.w_wdtn + (size_t)(iCell - 1) * (size_t)(ns * (L + 1));
        real wlo = wdc[e], whi = wdc[e + ns];
        const real in_k = sc[e], out_k = sc[nel + e];
        if (k0 >= 1) wlo = fmax((real)0.0, wlo) * fmin(sc[nel + e - ns], in_k) + fmin((real)0.0, wlo) * fmin(out_k, sc[e - ns]);
        if (k0 <= L - 2) whi = fmax((real)0.0, whi) * fmin(out_k, sc[e + ns]) + fmin((real)0.0, whi) * fmin(sc[nel + e + ns], in_k);
        const int ne = p.nEdgesOnCell[iCell - 1];
        const real invArea = p.invAreaCell[iCell - 1];
        // load-first: edge indices and signs of all (<= 8) slots, then the fluxes, then the sum in the reference's order;
        // slots beyond nEdgesOnCell repeat slot 1 and are left out of the sum
        const int* eoc = p.edgesOnCell + (size_t)(iCell - 1) * mE;
        const real* sgc = p.edgesOnCell_sign + (size_t)(iCell - 1) * mE;
        int ed[8];
        real sg[8], fl[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) ed[i] = eoc[i < ne ? i : 0];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (i < 6 || ne > 6) {
                sg[i] = sgc[i < ne ? i : 0];
                fl[i] = p.w_fa[(size_t)(ed[i] - 1) * nel + e];
            }
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
            if (i < ne) sn = sn - sg[i] * fl[i] * invArea;
        sn = (sn + (-p.rdnw[k0] * (whi - wlo))) / p.rho_lim[IX2(k0 + 1, iCell, L)];
    } else {
        sn = p.scalars_new[is];
    }
    if (p.bdyMaskCell[iCell - 1] <= MPAS_NSPECZONE) p.scalars_new[is] = fmax((real)0.0, sn);
}

}  // namespace

extern "C" int mpas_b200_advance_scalars_work(MPAS_ADVSCALARS_ARGS) {
    (void)nCells; (void)nEdges; (void)nCellsSolve; (void)kdiff; (void)deriv_two; (void)dvEdge; (void)cellsOnEdge; (void)meshScalingDel2; (void)meshScalingDel4;
    (void)rho_edge; (void)qv_init; (void)zgrid; (void)cellStart; (void)cellEnd; (void)vertexStart; (void)vertexEnd; (void)vertexSolveStart; (void)vertexSolveEnd;
    (void)edgeSolveStart; (void)edgeSolveEnd;
    AsP p;
    p.L = nVertLevels; p.mE = maxEdges; p.ns = num_scalars; p.nsdiv.set(num_scalars); p.edgeStart = edgeStart; p.nEdgesLoop = edgeEnd - edgeStart + 1;
    p.cellSolveStart = cellSolveStart; p.nCellsSolveLoop = cellSolveEnd - cellSolveStart + 1;
    real weight_time_new = 1.;
    if (advance_density) {  // TI:3296-3304
        if ((rk_step == 1) && config_time_integration_order == 3) weight_time_new = (real)1. / 3;
        if ((rk_step == 1) && config_time_integration_order == 2) weight_time_new = (real)1. / 2;
        if (rk_step == 2) weight_time_new = (real)1. / 2;
        if (rk_step == 3) weight_time_new = 1.;
    }
    p.weight_time_new = weight_time_new; p.weight_time_old = (real)1. - weight_time_new;
    p.dt = dt; p.coef_3rd_order = coef_3rd_order;
    p.scalar_old = X(scalar_old); p.rho_zz_old = X(rho_zz_old); p.rho_zz_new = X(rho_zz_new); p.uhAvg = X(uhAvg); p.wwAvg = X(wwAvg);
    p.edgesOnCell_sign = X(edgesOnCell_sign); p.invAreaCell = X(invAreaCell); p.fnm = X(fnm); p.fnp = X(fnp); p.rdnw = X(rdnw);
    p.adv_coefs = X(adv_coefs); p.adv_coefs_3rd = X(adv_coefs_3rd); p.edgesOnCell = X(edgesOnCell); p.nEdgesOnCell = X(nEdgesOnCell);
    p.bdyMaskCell = X(bdyMaskCell); p.bdyMaskEdge = X(bdyMaskEdge); p.nAdvCellsForEdge = X(nAdvCellsForEdge); p.advCellsForEdge = X(advCellsForEdge);
    p.scalar_tend_save = X(scalar_tend_save); p.scalar_new = X(scalar_new); p.horiz_flux_arr = X(horiz_flux_arr);
    WLAUNCH(as_edge_kernel, p.nEdgesLoop, 0, p);
    KCHECK("as_edge_kernel");
    if (p.nCellsSolveLoop > 0) {
        if (p.mE > 8) return fail_msg("mpas_b200_advance_scalars_work: maxEdges > 8 not supported");
        const size_t shm = (size_t)WPB * (tma_slot(p.L * p.ns) + tma_slot(p.L + 1) + ((p.ns * (p.L + 1) + 3) & ~3)) * sizeof(real);
        static const int ascell_minb = tune("as_cell_minb", 4);
        if (ascell_minb == 5) WLAUNCH(as_cell_kernel<5>, p.nCellsSolveLoop, shm, p);
        else if (ascell_minb == 6) WLAUNCH(as_cell_kernel<6>, p.nCellsSolveLoop, shm, p);
        else WLAUNCH(as_cell_kernel<4>, p.nCellsSolveLoop, shm, p);
        KCHECK("as_cell_kernel");
    }
    return 0;
}

extern "C" int mpas_b200_advance_scalars_mono_work(MPAS_MONO_ARGS) {
    (void)nEdges; (void)vertexStart; (void)vertexEnd; (void)vertexSolveStart; (void)vertexSolveEnd; (void)edgeSolveStart; (void)edgeSolveEnd;
    MonoP p;
    p.L = nVertLevels; p.km.set(nVertLevels, nVertLevels + 1); p.mE = maxEdges; p.ns = num_scalars; p.iScalar = 0; p.nCells = nCells;
    p.cellStart = cellStart; p.nCellsLoop = cellEnd - cellStart + 1; p.cellSolveStart = cellSolveStart; p.nCellsSolveLoop = cellSolveEnd - cellSolveStart + 1;
    p.edgeStart = edgeStart; p.nEdgesLoop = edgeEnd - edgeStart + 1; p.nCellsSolve = nCellsSolve; p.advance_density = advance_density;
    p.dt = dt; p.coef_3rd_order = coef_3rd_order;
    p.uhAvg = X(uhAvg); p.wwAvg = X(wwAvg); p.rho_zz_old = X(rho_zz_old); p.invAreaCell = X(invAreaCell); p.dvEdge = X(dvEdge);
    p.edgesOnCell_sign = X(edgesOnCell_sign); p.fnm = X(fnm); p.fnp = X(fnp); p.rdnw = X(rdnw); p.adv_coefs = X(adv_coefs); p.adv_coefs_3rd = X(adv_coefs_3rd);
    p.cellsOnEdge = X(cellsOnEdge); p.cellsOnCell = X(cellsOnCell); p.edgesOnCell = X(edgesOnCell); p.nEdgesOnCell = X(nEdgesOnCell);
    p.nAdvCellsForEdge = X(nAdvCellsForEdge); p.advCellsForEdge = X(advCellsForEdge); p.bdyMaskCell = X(bdyMaskCell); p.bdyMaskEdge = X(bdyMaskEdge);
    p.scalar_tend = X(scalar_tend); p.scalars_old = X(scalars_old); p.scalars_new = X(scalars_new); p.scalar_old = X(scalar_old); p.scalar_new = X(scalar_new);
    p.s_max = X(s_max); p.s_min = X(s_min); p.wdtn = X(wdtn); p.scale_arr = X(scale_arr); p.flux_arr = X(flux_arr); p.flux_upwind_tmp = X(flux_upwind_tmp);
    p.flux_tmp = X(flux_tmp); p.rho_zz_int = X(rho_zz_int);
    p.rho_lim = advance_density ? p.rho_zz_int : X(rho_zz_new);
    if (advance_density && !rho_zz_int) return fail_msg("mpas_b200_advance_scalars_mono_work: advance_density needs rho_zz_int (TI:3768)");
    const int L = p.L;
    KLAUNCH(mono_pre_kernel, (long long)p.nCellsSolveLoop * L * p.ns, 256, p);
    KCHECK("mono_pre_kernel");
    if (halo_fn) {
        int rc = halo_fn(halo_ctx, 0, scalars_old);  // TI:3757
        if (rc) return rc;
    }
    if (advance_density) {
        KLAUNCH(mono_rho_int_kernel, (long long)p.nCellsLoop * L, 256, p);
        KCHECK("mono_rho_int_kernel");
    }
    // ---- all scalars at once (default): needs the per-block workspace and the 6/8-slot edge tables
    if (tune("mono_fused", 1) && maxEdges <= 8 && p.nCellsLoop > 0) {
        const size_t nel = (size_t)L * p.ns, nel1 = (size_t)(L + 1) * p.ns;
        const MonoWsLayout lay(L, p.ns, nCells, nEdges);
        const size_t n_edge = lay.n_edge, n_wd = lay.n_wd, n_cell = lay.n_cell;
        const size_t shm = (size_t)WPB * (2 * tma_slot((int)nel) + tma_slot(L + 1) + (((int)nel1 + 3) & ~3)) * sizeof(real);
        static const int cell_minb = tune("monoN_cell_minb", 4), edge_minb = tune("monoN_edge_minb", 5);
        const void* cell_fn = cell_minb == 2 ? (const void*)monoN_cell_kernel<2> : cell_minb == 3 ? (const void*)monoN_cell_kernel<3> : (const void*)monoN_cell_kernel<4>;
        const bool fits = ensure_dyn_smem(cell_fn, shm);
        real* ws = fits ? keyed_workspace(X(scale_arr), 2 * n_edge + n_wd + 3 * n_cell) : nullptr;
        // no silent fallback: ranks / blocks taking different paths would post different numbers of exchanges
        if (fits && !ws)
            return fail_msg("mpas_b200_advance_scalars_mono_work: cannot allocate the all-scalars workspace (" +
                            std::to_string((2 * n_edge + n_wd + 3 * n_cell) * sizeof(real) >> 20) +
                            " MiB; first call must be outside stream capture); MPAS_B200_TUNE_mono_fused=0 selects the per-scalar path");
        if (ws) {
            MonoNP q;
            q.L = p.L; q.km.set(p.L * p.ns); q.mE = p.mE; q.ns = p.ns; q.nsdiv.set(p.ns); q.nCells = nCells; q.cellStart = p.cellStart; q.nCellsLoop = p.nCellsLoop;
            q.cellSolveStart = p.cellSolveStart; q.nCellsSolveLoop = p.nCellsSolveLoop; q.edgeStart = p.edgeStart; q.nEdgesLoop = p.nEdgesLoop;
            q.nCellsSolve = p.nCellsSolve; q.dt = p.dt; q.coef_3rd_order = p.coef_3rd_order;
            q.uhAvg = p.uhAvg; q.wwAvg = p.wwAvg; q.rho_zz_old = p.rho_zz_old; q.rho_lim = p.rho_lim; q.invAreaCell = p.invAreaCell; q.dvEdge = p.dvEdge;
            q.edgesOnCell_sign = p.edgesOnCell_sign; q.fnm = p.fnm; q.fnp = p.fnp; q.rdnw = p.rdnw; q.adv_coefs = p.adv_coefs; q.adv_coefs_3rd = p.adv_coefs_3rd;
            q.cellsOnEdge = p.cellsOnEdge; q.cellsOnCell = p.cellsOnCell; q.edgesOnCell = p.edgesOnCell; q.nEdgesOnCell = p.nEdgesOnCell;
            q.nAdvCellsForEdge = p.nAdvCellsForEdge; q.advCellsForEdge = p.advCellsForEdge; q.bdyMaskCell = p.bdyMaskCell; q.bdyMaskEdge = p.bdyMaskEdge;
            q.scalars_old = p.scalars_old; q.scalars_new = p.scalars_new;
            q.w_fa = ws; q.w_fup = ws + n_edge; q.w_wdtn = ws + 2 * n_edge; q.w_snew = q.w_wdtn + n_wd; q.w_scale = q.w_snew + n_cell;
            if (edge_minb == 5) WLAUNCH(monoN_edge_flux_kernel<5>, q.nEdgesLoop, 0, q);
            else if (edge_minb == 6) WLAUNCH(monoN_edge_flux_kernel<6>, q.nEdgesLoop, 0, q);
            else WLAUNCH(monoN_edge_flux_kernel<4>, q.nEdgesLoop, 0, q);
            KCHECK("monoN_edge_flux_kernel");
            if (cell_minb == 2) WLAUNCH(monoN_cell_kernel<2>, q.nCellsSolveLoop, shm, q);
            else if (cell_minb == 3) WLAUNCH(monoN_cell_kernel<3>, q.nCellsSolveLoop, shm, q);
            else WLAUNCH(monoN_cell_kernel<4>, q.nCellsSolveLoop, shm, q);
            KCHECK("monoN_cell_kernel");
            if (halo_fn) {
                int rc = halo_fn(halo_ctx, 2, q.w_scale);  // TI:4098, all scalars in one message
                if (rc) return rc;
            }
            KLAUNCH(monoN_rescale_edge_kernel, (long long)q.nEdgesLoop * (long long)nel, 256, q);
            KCHECK("monoN_rescale_edge_kernel");
            KLAUNCH(monoN_update_kernel, (long long)q.nCellsLoop * (long long)nel, 256, q);
            KCHECK("monoN_update_kernel");
            return 0;
        }
    }
    // ---- one scalar at a time with the reference's 2-D scratch arrays (TI:3798-4212)
    for (int iScalar = 1; iScalar <= num_scalars; ++iScalar) {
        p.iScalar = iScalar;
        KLAUNCH(mono_copy_kernel, (long long)(p.nCellsLoop + 1) * L, 256, p);
        KCHECK("mono_copy_kernel");
        KLAUNCH(mono_bounds_kernel, (long long)p.nCellsSolveLoop * (L + 1), 256, p);
        KCHECK("mono_bounds_kernel");
        WLAUNCH(mono_edge_flux_kernel, p.nEdgesLoop, 0, p);
        KCHECK("mono_edge_flux_kernel");
        KLAUNCH(mono_limiter_kernel, (long long)p.nCellsSolveLoop * L, 256, p);
        KCHECK("mono_limiter_kernel");
        KLAUNCH(mono_wdtn_store_kernel, (long long)p.nCellsSolveLoop * L, 256, p);
        KCHECK("mono_wdtn_store_kernel");
        if (halo_fn) {
            int rc = halo_fn(halo_ctx, 1, scale_arr);  // TI:4098
            if (rc) return rc;
        }
        KLAUNCH(mono_rescale_edge_kernel, (long long)p.nEdgesLoop * L, 256, p);
        KCHECK("mono_rescale_edge_kernel");
        KLAUNCH(mono_update_kernel, (long long)p.nCellsLoop * L, 256, p);
        KCHECK("mono_update_kernel");
    }
    return 0;
}
