// atm_compute_solve_diagnostics_work (TI:5513-5822), TI = src/core_atmosphere/dynamics/mpas_atm_time_integration.F.
//
// The reference sweeps edges, vertices and cells eleven times with barriers in between.  Data
// dependences allow three kernels, all one thread per (k, element), k fastest:
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
    KLAUNCH(sd_vertex_kernel, (long long)p.nVerticesLoop * p.L, 256, p);
    KCHECK("sd_vertex_kernel");
    KLAUNCH(sd_cell_kernel, (long long)p.nCellsLoop * p.L, 256, p);
    KCHECK("sd_cell_kernel");
    if (p.reconstruct_v) KLAUNCH(sd_edge_kernel<true>, (long long)p.nEdgesLoop * p.L, 256, p);
    else KLAUNCH(sd_edge_kernel<false>, (long long)p.nEdgesLoop * p.L, 256, p);
    KCHECK("sd_edge_kernel");
    return 0;
}
