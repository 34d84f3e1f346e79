// atm_compute_solve_diagnostics_work (TI:5513-5822), TI = src/core_atmosphere/dynamics/mpas_atm_time_integration.F.
//
// The reference sweeps edges, vertices and cells eleven times with barriers in between.  Data
// dependences allow three kernels.  sd_vertex is one thread per (k, element), k fastest; sd_cell and sd_edge run one
// element per warp with two levels per lane (sd_cell_warp_kernel, sd_edge_warp_kernel below: 258 -> 170 us, 596 -> 437 us
// with the tangential velocity, 273 -> 223 us without, 60 km mesh); the flat forms stay as the fallback for maxEdges > 8
// and behind MPAS_B200_TUNE_sd_warp=0:
//   sd_vertex_kernel  vorticity (TI:5606-5620), ke_vertex (TI:5683-5690, with ke_edge = dc*dv*u^2 of
//                     TI:5598 recomputed in registers -- the same expression, so the same bits),
//                     pv_vertex (TI:5745-5749)
//   sd_cell_kernel    divergence (TI:5627-5640), ke incl. the Hollingsworth blend with ke_vertex
//                     (TI:5650-5712), pv_cell (TI:5779-5790)
//   sd_edge_kernel    h_edge (TI:5589-5591), ke_edge (kept for callers that read the scratch array),
//                     v (TI:5722-5733, only when reconstruct_v), pv_edge with APVM upwinding
//                     (TI:5756-5765, 5800-5816), gradPVt / gradPVn
#include "common.cuh"

using namespace mpasb200;

namespace {

struct SdP {
    KMagic km;
    int L, mE, mE2, vD;
    int cellStart, nCellsLoop, vertexStart, nVerticesLoop, edgeStart, nEdgesLoop;
    int reconstruct_v;
    real dt, apvm;
    const real *fVertex, *invAreaTriangle, *invAreaCell, *dvEdge, *dcEdge, *invDvEdge, *invDcEdge, *weightsOnEdge, *kiteAreasOnVertex;
    const real *h, *u, *edgesOnVertex_sign, *edgesOnCell_sign;
    const int *cellsOnEdge, *verticesOnEdge, *edgesOnCell, *edgesOnEdge, *edgesOnVertex, *kiteForCell, *verticesOnCell, *nEdgesOnCell, *nEdgesOnEdge;
    real *h_edge, *v, *vorticity, *ke, *pv_edge, *pv_vertex, *pv_cell, *gradPVn, *gradPVt, *divergence, *ke_vertex, *ke_edge;
};

// sd_vertex and the fixed part of sd_edge are written load-first: the index loads of every stencil slot, then every data load, then the sums
// in the reference's order.  A thread therefore waits for memory twice instead of twice per stencil entry (the edge /
// vertex loops have run-time trip counts the compiler will not batch).  Slots beyond the run-time count repeat slot 1
// and are left out of the sums.
__global__ void __launch_bounds__(256) sd_vertex_kernel(SdP p) {
    const int L = p.L;
    FLAT_KC(p.nVerticesLoop, L);
    const int iVertex = p.vertexStart + _c0;
    const int* eov = p.edgesOnVertex + (size_t)(iVertex - 1) * 3;
    const real* sg = p.edgesOnVertex_sign + (size_t)(iVertex - 1) * 3;
    int e[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) e[i] = eov[i];
    real dc[3], dv[3], ue[3], sv[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        dc[i] = p.dcEdge[e[i] - 1];
        dv[i] = p.dvEdge[e[i] - 1];
        sv[i] = sg[i];
        ue[i] = p.u[IX2(k, e[i], L)];
    }
    const real iat = p.invAreaTriangle[iVertex - 1];
    const real fv = p.fVertex[iVertex - 1];
    real vort = 0.0;
    real kev[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const real s = sv[i] * dc[i];
        vort = vort + s * ue[i];
        const real efac = dc[i] * dv[i];
        kev[i] = efac * (ue[i] * ue[i]);
    }
    vort = vort * iat;
    const size_t iv = IX2(k, iVertex, L);
    p.vorticity[iv] = vort;
    const real r = (real)0.25 * iat;
    p.ke_vertex[iv] = (kev[0] + kev[1] + kev[2]) * r;
    p.pv_vertex[iv] = (fv + vort);
}

__global__ void __launch_bounds__(256) sd_cell_kernel(SdP p) {
    const int L = p.L, mE = p.mE;
    FLAT_KC(p.nCellsLoop, L);
    const int iCell = p.cellStart + _c0;
    const int ne = p.nEdgesOnCell[iCell - 1];
    const real r = p.invAreaCell[iCell - 1];
    real div = 0.0, ke = 0.0;
    for (int i = 1; i <= ne; ++i) {
        const int iEdge = p.edgesOnCell[IX2(i, iCell, mE)];
        const real dv = p.dvEdge[iEdge - 1];
        const real s = p.edgesOnCell_sign[IX2(i, iCell, mE)] * dv;
        const real ue = p.u[IX2(k, iEdge, L)];
        div = div + s * ue;
        const real efac = p.dcEdge[iEdge - 1] * dv;
        ke = ke + (real)0.25 * (efac * (ue * ue));
    }
    div = div * r;
    ke = ke * r;
    const real ke_fact = (real)1.0 - (real).375;
    ke = ke_fact * ke;
    real pvc = 0.0;
    for (int i = 1; i <= ne; ++i) {
        const int iVertex = p.verticesOnCell[IX2(i, iCell, mE)];
        const int j = p.kiteForCell[IX2(i, iCell, mE)];
        const real kite = p.kiteAreasOnVertex[IX2(j, iVertex, 3)];
        const size_t iv = IX2(k, iVertex, L);
        ke = ke + ((real)1. - ke_fact) * kite * p.ke_vertex[iv] * r;
        pvc = pvc + kite * p.pv_vertex[iv] * r;
    }
    const size_t ic = IX2(k, iCell, L);
    p.divergence[ic] = div;
    p.ke[ic] = ke;
    if (p.apvm > (real)0.0) p.pv_cell[ic] = pvc;
}

template <bool RECON>
__global__ void __launch_bounds__(256) sd_edge_kernel(SdP p) {
    const int L = p.L, mE2 = p.mE2;
    FLAT_KC(p.nEdgesLoop, L);
    const int iEdge = p.edgeStart + _c0;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    const int v1 = p.verticesOnEdge[IX2(1, iEdge, 2)], v2 = p.verticesOnEdge[IX2(2, iEdge, 2)];
    const size_t ie = IX2(k, iEdge, L);
    const real h1 = p.h[IX2(k, cell1, L)], h2 = p.h[IX2(k, cell2, L)];
    const real ue = p.u[ie];
    const real pv1 = p.pv_vertex[IX2(k, v1, L)], pv2 = p.pv_vertex[IX2(k, v2, L)];
    const bool apvm = p.apvm > (real)0.0;
    real pc1 = 0, pc2 = 0, idv = 0, idc = 0;
    if (apvm) {
        pc1 = p.pv_cell[IX2(k, cell1, L)];
        pc2 = p.pv_cell[IX2(k, cell2, L)];
        idv = p.invDvEdge[iEdge - 1];
        idc = p.invDcEdge[iEdge - 1];
    }
    p.h_edge[ie] = (real)0.5 * (h1 + h2);
    if (p.ke_edge) {
        const real efac = p.dcEdge[iEdge - 1] * p.dvEdge[iEdge - 1];
        p.ke_edge[ie] = efac * (ue * ue);
    }
    real ve;
    if (RECON) {
        // TI:5641-5650.  The first RN edges: all indices, then all weights and velocities are loaded before the ordered sum
        // (two round trips instead of two per edge); entries beyond nEdgesOnEdge repeat the first index and are not added.
        ve = 0.0;
        const int n = p.nEdgesOnEdge[iEdge - 1];
        constexpr int RN = 10;
        const int* eoep = p.edgesOnEdge + (size_t)(iEdge - 1) * mE2;
        const real* wp = p.weightsOnEdge + (size_t)(iEdge - 1) * mE2;
        const int nn = mE2 < RN ? mE2 : RN;
        int eo[RN];
#pragma unroll
        for (int i = 0; i < RN; ++i) eo[i] = eoep[i < n && i < nn ? i : 0];
        real wv[RN], uv[RN];
#pragma unroll
        for (int i = 0; i < RN; ++i) {
            wv[i] = wp[i < nn ? i : 0];
            uv[i] = p.u[IX2(k, eo[i], L)];
        }
#pragma unroll
        for (int i = 0; i < RN; ++i)
            if (i < n && i < nn) ve = ve + wv[i] * uv[i];
        for (int i = RN + 1; i <= n; ++i) {
            const int eoe = p.edgesOnEdge[IX2(i, iEdge, mE2)];
            ve = ve + p.weightsOnEdge[IX2(i, iEdge, mE2)] * p.u[IX2(k, eoe, L)];
        }
        p.v[ie] = ve;
    } else {
        ve = apvm ? p.v[ie] : (real)0.0;
    }
    real pve = (real)0.5 * (pv1 + pv2);
    if (apvm) {
        const real r = p.apvm * p.dt;
        const real r1 = (real)1.0 * idv, r2 = (real)1.0 * idc;
        const real gt = (pv2 - pv1) * r1;
        const real gn = (pc2 - pc1) * r2;
        if (p.gradPVt) p.gradPVt[ie] = gt;
        if (p.gradPVn) p.gradPVn[ie] = gn;
        pve = pve - r * (ve * gt + ue * gn);
    }
    p.pv_edge[ie] = pve;
}

// ---- one mesh element per warp, lanes along k (two levels per lane) --------------------------------------------------
// The flat kernels above load an element's stencil metadata (indices, signs, weights, metric factors) once per THREAD,
// 55 times per element; here one lane per stencil slot loads it once per element into shared memory as element offsets
// and ready-made factors, and the gathers of a sum are all issued before the first dependent add.  The sums keep the
// reference's order.  Lists are padded with zero-factor self references (x + 0*y == x exactly: the sums start at +0 and
// are never -0).

// sd_vertex: the three edges' metadata are warp-uniform loads, once per vertex
template <int MINB>
__global__ void __launch_bounds__(WPB * 32, MINB) sd_vertex_warp_kernel(SdP p) {
    const int L = p.L;
    WARP_ENTITY(p.vertexStart, p.nVerticesLoop);
    const int iVertex = ent;
    const int* eov = p.edgesOnVertex + (size_t)(iVertex - 1) * 3;
    const real* sg = p.edgesOnVertex_sign + (size_t)(iVertex - 1) * 3;
    unsigned eo[3];
    real sdc[3], efac[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const int e = eov[i];
        const real dc = p.dcEdge[e - 1], dv = p.dvEdge[e - 1];
        eo[i] = (unsigned)(e - 1) * (unsigned)L;
        sdc[i] = sg[i] * dc;
        efac[i] = dc * dv;
    }
    const real iat = p.invAreaTriangle[iVertex - 1];
    const real fv = p.fVertex[iVertex - 1];
    const real r = (real)0.25 * iat;
    const size_t voff = (size_t)(iVertex - 1) * L;
    for (int ka = lane; ka < L; ka += 64) {
        const bool vb = ka + 32 < L;
        const int bo = vb ? 32 : 0;
        real ua[3], ub[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const real* __restrict__ q = elem_ptr(p.u + ka, eo[i]);
            ua[i] = q[0];
            ub[i] = q[bo];
        }
        real va = 0.0, vbb = 0.0, ka_[3], kb_[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            va = va + sdc[i] * ua[i];
            vbb = vbb + sdc[i] * ub[i];
            ka_[i] = efac[i] * (ua[i] * ua[i]);
            kb_[i] = efac[i] * (ub[i] * ub[i]);
        }
        va = va * iat;
        vbb = vbb * iat;
        p.vorticity[voff + ka] = va;
        p.ke_vertex[voff + ka] = (ka_[0] + ka_[1] + ka_[2]) * r;
        p.pv_vertex[voff + ka] = (fv + va);
        if (vb) {
            p.vorticity[voff + ka + 32] = vbb;
            p.ke_vertex[voff + ka + 32] = (kb_[0] + kb_[1] + kb_[2]) * r;
            p.pv_vertex[voff + ka + 32] = (fv + vbb);
        }
    }
}

// sd_cell: divergence, ke (with the ke_vertex blend), pv_cell.  NE = padded edge count (6 or 8).
template <int NE>
__device__ __forceinline__ void sd_cell_levels(const SdP& p, int L, int lane, const unsigned* eo, const real* sdv, const real* efac, const unsigned* vo,
                                               const real* kite, real r, size_t coff) {
    const real ke_fact = (real)1.0 - (real).375;
    for (int ka = lane; ka < L; ka += 64) {
        const bool vb = ka + 32 < L;
        const int bo = vb ? 32 : 0;
        const real* __restrict__ up = p.u + ka;
        const real* __restrict__ kvp = p.ke_vertex + ka;
        const real* __restrict__ pvp = p.pv_vertex + ka;
        real ua[NE], ub[NE];
#pragma unroll
        for (int t = 0; t < NE; ++t) {
            const real* __restrict__ q = elem_ptr(up, eo[t]);
            ua[t] = q[0];
            ub[t] = q[bo];
        }
        real div_a = 0.0, div_b = 0.0, ke_a = 0.0, ke_b = 0.0;
#pragma unroll
        for (int t = 0; t < NE; ++t) {  // TI:5627-5640, 5650-5660
            div_a = div_a + sdv[t] * ua[t];
            div_b = div_b + sdv[t] * ub[t];
            ke_a = ke_a + (real)0.25 * (efac[t] * (ua[t] * ua[t]));
            ke_b = ke_b + (real)0.25 * (efac[t] * (ub[t] * ub[t]));
        }
        div_a = div_a * r; div_b = div_b * r;
        ke_a = ke_fact * (ke_a * r); ke_b = ke_fact * (ke_b * r);
        real ka_[NE], kb_[NE], pa[NE], pb[NE];
#pragma unroll
        for (int t = 0; t < NE; ++t) {
            const real* __restrict__ q = elem_ptr(kvp, vo[t]);
            const real* __restrict__ q2 = elem_ptr(pvp, vo[t]);
            ka_[t] = q[0];
            kb_[t] = q[bo];
            pa[t] = q2[0];
            pb[t] = q2[bo];
        }
        real pvc_a = 0.0, pvc_b = 0.0;
#pragma unroll
        for (int t = 0; t < NE; ++t) {  // TI:5683-5712, 5779-5790
            const real c = ((real)1. - ke_fact) * kite[t];
            ke_a = ke_a + c * ka_[t] * r;
            ke_b = ke_b + c * kb_[t] * r;
            pvc_a = pvc_a + kite[t] * pa[t] * r;
            pvc_b = pvc_b + kite[t] * pb[t] * r;
        }
        const size_t ic = coff + ka;
        p.divergence[ic] = div_a;
        p.ke[ic] = ke_a;
        if (p.apvm > (real)0.0) p.pv_cell[ic] = pvc_a;
        if (vb) {
            p.divergence[ic + 32] = div_b;
            p.ke[ic + 32] = ke_b;
            if (p.apvm > (real)0.0) p.pv_cell[ic + 32] = pvc_b;
        }
    }
}

// cells with more than 6 edges (rare): the same sums, one stencil slot at a time
__device__ __forceinline__ void sd_cell_levels_any(const SdP& p, int L, int lane, int ne, const unsigned* eo, const real* sdv, const real* efac, const unsigned* vo,
                                                const real* kite, real r, size_t coff) {
    const real ke_fact = (real)1.0 - (real).375;
    for (int k0 = lane; k0 < L; k0 += 32) {
        real div = 0.0, ke = 0.0;
#pragma unroll 1
        for (int t = 0; t < ne; ++t) {
            const real ue = p.u[(size_t)eo[t] + k0];
            div = div + sdv[t] * ue;
            ke = ke + (real)0.25 * (efac[t] * (ue * ue));
        }
        div = div * r;
        ke = ke_fact * (ke * r);
        real pvc = 0.0;
#pragma unroll 1
        for (int t = 0; t < ne; ++t) {
            ke = ke + ((real)1. - ke_fact) * kite[t] * p.ke_vertex[(size_t)vo[t] + k0] * r;
            pvc = pvc + kite[t] * p.pv_vertex[(size_t)vo[t] + k0] * r;
        }
        const size_t ic = coff + k0;
        p.divergence[ic] = div;
        p.ke[ic] = ke;
        if (p.apvm > (real)0.0) p.pv_cell[ic] = pvc;
    }
}

// MINB = 6 spills ~200 bytes per thread and is still the fastest (170 us against 192 at 4 and 247 at 3 blocks per SM, 60 km mesh):
// the gathers want the warps more than the registers.  One level per lane instead of two was 1.8x slower.
template <int MINB>
__global__ void __launch_bounds__(WPB * 32, MINB) sd_cell_warp_kernel(SdP p) {
    __shared__ unsigned s_eo[WPB][8], s_vo[WPB][8];
    __shared__ real s_sdv[WPB][8], s_efac[WPB][8], s_kite[WPB][8];
    const int L = p.L, mE = p.mE;
    WARP_ENTITY(p.cellStart, p.nCellsLoop);
    const int iCell = ent;
    const int ne = p.nEdgesOnCell[iCell - 1];
    const real r = p.invAreaCell[iCell - 1];
    const unsigned coff = (unsigned)(iCell - 1) * (unsigned)L;
    if (lane < 8) {
        const bool in = lane < ne;
        const int iEdge = in ? p.edgesOnCell[IX2(lane + 1, iCell, mE)] : 1;
        const real dv = p.dvEdge[iEdge - 1];
        s_eo[warp][lane] = (unsigned)(iEdge - 1) * (unsigned)L;
        s_sdv[warp][lane] = in ? p.edgesOnCell_sign[IX2(lane + 1, iCell, mE)] * dv : (real)0.0;
        s_efac[warp][lane] = in ? p.dcEdge[iEdge - 1] * dv : (real)0.0;
        const int iVertex = in ? p.verticesOnCell[IX2(lane + 1, iCell, mE)] : 1;
        const int j = in ? p.kiteForCell[IX2(lane + 1, iCell, mE)] : 1;
        s_vo[warp][lane] = (unsigned)(iVertex - 1) * (unsigned)L;
        s_kite[warp][lane] = in ? p.kiteAreasOnVertex[IX2(j, iVertex, 3)] : (real)0.0;
    }
    __syncwarp();
    if (ne <= 6) sd_cell_levels<6>(p, L, lane, s_eo[warp], s_sdv[warp], s_efac[warp], s_vo[warp], s_kite[warp], r, coff);
    else sd_cell_levels_any(p, L, lane, ne, s_eo[warp], s_sdv[warp], s_efac[warp], s_vo[warp], s_kite[warp], r, coff);
}

// sd_edge with the tangential velocity (TI:5722-5733): the edgesOnEdge gather of dt_edge_kernel (dyn_tend.cu), UNR slots
// at a time, and the edge's own operands as coalesced loads.
template <int UNR, int MINB, bool RECON>
__global__ void __launch_bounds__(WPB * 32, MINB) sd_edge_warp_kernel(SdP p) {
    __shared__ unsigned s_off[WPB][20];
    __shared__ real s_wt[WPB][20];
    const int L = p.L, mE2 = p.mE2;
    WARP_ENTITY(p.edgeStart, p.nEdgesLoop);
    const int iEdge = ent;
    const int cell1 = p.cellsOnEdge[IX2(1, iEdge, 2)], cell2 = p.cellsOnEdge[IX2(2, iEdge, 2)];
    const int v1 = p.verticesOnEdge[IX2(1, iEdge, 2)], v2 = p.verticesOnEdge[IX2(2, iEdge, 2)];
    const int n = RECON ? p.nEdgesOnEdge[iEdge - 1] : 0;
    const unsigned eoff = (unsigned)(iEdge - 1) * (unsigned)L;
    const int npad = (n + UNR - 1) / UNR * UNR;
    if (RECON && lane < npad) {
        const bool in = lane < n;
        s_off[warp][lane] = in ? (unsigned)(p.edgesOnEdge[IX2(lane + 1, iEdge, mE2)] - 1) * (unsigned)L : eoff;
        s_wt[warp][lane] = in ? p.weightsOnEdge[IX2(lane + 1, iEdge, mE2)] : (real)0.0;
    }
    const bool apvm = p.apvm > (real)0.0;
    real idv = 0, idc = 0;
    if (apvm) {
        idv = p.invDvEdge[iEdge - 1];
        idc = p.invDcEdge[iEdge - 1];
    }
    const real efac = p.dcEdge[iEdge - 1] * p.dvEdge[iEdge - 1];
    const real* __restrict__ hc1 = p.h + (size_t)(cell1 - 1) * L;
    const real* __restrict__ hc2 = p.h + (size_t)(cell2 - 1) * L;
    const real* __restrict__ pvv1 = p.pv_vertex + (size_t)(v1 - 1) * L;
    const real* __restrict__ pvv2 = p.pv_vertex + (size_t)(v2 - 1) * L;
    const real* __restrict__ pcc1 = p.pv_cell + (size_t)(cell1 - 1) * L;
    const real* __restrict__ pcc2 = p.pv_cell + (size_t)(cell2 - 1) * L;
    __syncwarp();
    for (int ka = lane; ka < L; ka += 64) {
        const bool vb = ka + 32 < L;
        const int bo = vb ? 32 : 0;
        const real* __restrict__ up = p.u + ka;
        // the edge's own operands first: they are in flight during the gather
        real h1[2], h2[2], ue[2], pv1[2], pv2[2], pc1[2] = {0, 0}, pc2[2] = {0, 0};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int k0 = ka + (h ? bo : 0);
            h1[h] = hc1[k0]; h2[h] = hc2[k0];
            ue[h] = p.u[(size_t)eoff + k0];
            pv1[h] = pvv1[k0]; pv2[h] = pvv2[k0];
            if (apvm) { pc1[h] = pcc1[k0]; pc2[h] = pcc2[k0]; }
        }
        real ve_a = 0.0, ve_b = 0.0;
        if (!RECON && apvm) {
            ve_a = p.v[(size_t)eoff + ka];
            ve_b = p.v[(size_t)eoff + ka + bo];
        }
        for (int j = 0; RECON && j < npad; j += UNR) {
            real ua[UNR], ub[UNR], wt[UNR];
#pragma unroll
            for (int t = 0; t < UNR; ++t) {
                const real* __restrict__ q = elem_ptr(up, s_off[warp][j + t]);
                wt[t] = s_wt[warp][j + t];
                ua[t] = q[0];
                ub[t] = q[bo];
            }
#pragma unroll
            for (int t = 0; t < UNR; ++t) {
                ve_a = ve_a + wt[t] * ua[t];
                ve_b = ve_b + wt[t] * ub[t];
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !vb) break;
            const size_t ie = (size_t)eoff + ka + 32 * h;
            const real ve = h ? ve_b : ve_a;
            p.h_edge[ie] = (real)0.5 * (h1[h] + h2[h]);
            if (p.ke_edge) p.ke_edge[ie] = efac * (ue[h] * ue[h]);
            if (RECON) p.v[ie] = ve;
            real pve = (real)0.5 * (pv1[h] + pv2[h]);
            if (apvm) {
                const real rr = p.apvm * p.dt;
                const real r1 = (real)1.0 * idv, r2 = (real)1.0 * idc;
                const real gt = (pv2[h] - pv1[h]) * r1;
                const real gn = (pc2[h] - pc1[h]) * r2;
                if (p.gradPVt) p.gradPVt[ie] = gt;
                if (p.gradPVn) p.gradPVn[ie] = gn;
                pve = pve - rr * (ve * gt + ue[h] * gn);
            }
            p.pv_edge[ie] = pve;
        }
    }
}

}  // namespace

extern "C" int mpas_b200_compute_solve_diagnostics_work(MPAS_SOLVEDIAG_ARGS) {
    (void)nCells; (void)nEdges; (void)nVertices; (void)cellsOnVertex;
    SdP p;
    p.L = nVertLevels; p.km.set(nVertLevels, nVertLevels + 1); p.mE = maxEdges; p.mE2 = maxEdges2; p.vD = vertexDegree;
    if (vertexDegree != 3) return fail_msg("mpas_b200_compute_solve_diagnostics_work: vertexDegree must be 3 (TI:5683-5690 hard-codes three edges)");
    p.cellStart = cellStart; p.nCellsLoop = cellEnd - cellStart + 1;
    p.vertexStart = vertexStart; p.nVerticesLoop = vertexEnd - vertexStart + 1;
    p.edgeStart = edgeStart; p.nEdgesLoop = edgeEnd - edgeStart + 1;
    p.reconstruct_v = (rk_step >= 0 && rk_step != 3) ? 0 : 1;  // TI:5716-5720
    p.dt = dt; p.apvm = config_apvm_upwinding;
    p.fVertex = X(fVertex); p.invAreaTriangle = X(invAreaTriangle); p.invAreaCell = X(invAreaCell); p.dvEdge = X(dvEdge); p.dcEdge = X(dcEdge);
    p.invDvEdge = X(invDvEdge); p.invDcEdge = X(invDcEdge); p.weightsOnEdge = X(weightsOnEdge); p.kiteAreasOnVertex = X(kiteAreasOnVertex);
    p.h = X(h); p.u = X(u); p.edgesOnVertex_sign = X(edgesOnVertex_sign); p.edgesOnCell_sign = X(edgesOnCell_sign);
    p.cellsOnEdge = X(cellsOnEdge); p.verticesOnEdge = X(verticesOnEdge); p.edgesOnCell = X(edgesOnCell); p.edgesOnEdge = X(edgesOnEdge);
    p.edgesOnVertex = X(edgesOnVertex); p.kiteForCell = X(kiteForCell); p.verticesOnCell = X(verticesOnCell); p.nEdgesOnCell = X(nEdgesOnCell);
    p.nEdgesOnEdge = X(nEdgesOnEdge);
    p.h_edge = X(h_edge); p.v = X(v); p.vorticity = X(vorticity); p.ke = X(ke); p.pv_edge = X(pv_edge); p.pv_vertex = X(pv_vertex); p.pv_cell = X(pv_cell);
    p.gradPVn = X(gradPVn); p.gradPVt = X(gradPVt); p.divergence = X(divergence); p.ke_vertex = X(ke_vertex); p.ke_edge = X(ke_edge);
    static const int vw = tune("sd_vertex_warp", 8);  // blocks per SM of the warp form, 0 = flat kernel
    if (vw == 4) WLAUNCH(sd_vertex_warp_kernel<4>, p.nVerticesLoop, 0, p);
    else if (vw == 8) WLAUNCH(sd_vertex_warp_kernel<8>, p.nVerticesLoop, 0, p);
    else if (vw) WLAUNCH(sd_vertex_warp_kernel<6>, p.nVerticesLoop, 0, p);
    else KLAUNCH(sd_vertex_kernel, (long long)p.nVerticesLoop * p.L, 256, p);
    KCHECK("sd_vertex_kernel");
    static const int warp_form = tune("sd_warp", 7);  // bit 0: sd_cell, bit 1: sd_edge with reconstruct_v, bit 2: sd_edge without (0 = the flat kernels)
    static const int cell_minb = tune("sd_cell_minb", 6), edge_minb = tune("sd_edge_minb", 6);
    const bool cw = (warp_form & 1) && maxEdges <= 8;
    if (cw && cell_minb == 4) WLAUNCH(sd_cell_warp_kernel<4>, p.nCellsLoop, 0, p);
    else if (cw && cell_minb == 8) WLAUNCH(sd_cell_warp_kernel<8>, p.nCellsLoop, 0, p);
    else if (cw) WLAUNCH(sd_cell_warp_kernel<6>, p.nCellsLoop, 0, p);
    else KLAUNCH(sd_cell_kernel, (long long)p.nCellsLoop * p.L, 256, p);
    KCHECK("sd_cell_kernel");
    if (p.reconstruct_v && (warp_form & 2) && maxEdges2 <= 20) WLAUNCH((sd_edge_warp_kernel<5, 4, true>), p.nEdgesLoop, 0, p);
    else if (!p.reconstruct_v && (warp_form & 4) && edge_minb == 6) WLAUNCH((sd_edge_warp_kernel<5, 6, false>), p.nEdgesLoop, 0, p);
    else if (!p.reconstruct_v && (warp_form & 4) && edge_minb == 8) WLAUNCH((sd_edge_warp_kernel<5, 8, false>), p.nEdgesLoop, 0, p);
    else if (!p.reconstruct_v && (warp_form & 4)) WLAUNCH((sd_edge_warp_kernel<5, 4, false>), p.nEdgesLoop, 0, p);
    else if (p.reconstruct_v) KLAUNCH(sd_edge_kernel<true>, (long long)p.nEdgesLoop * p.L, 256, p);
    else KLAUNCH(sd_edge_kernel<false>, (long long)p.nEdgesLoop * p.L, 256, p);
    KCHECK("sd_edge_kernel");
    return 0;
}
