// Coefficients of the cell-centre velocity reconstruction, restating
//   mpas_initialize_vectors        src/operators/mpas_vector_operations.F:653-769
//   mpas_init_reconstruct          src/operators/mpas_vector_reconstruction.F:51-176
//   mpas_rbf_interp_func_3D_plane_vec_const_dir_comp_coeffs   src/operators/mpas_rbf_interpolation.F:1079-1145
//   mpas_set_up_vector_dirichlet_rbf_matrix_and_rhs :1527-1559, evaluate_rbf :1369-1376, mpas_legs :1670-1698, elgs :1782-1846
// (called from atm_mpas_init_block, mpas_atm_core.F:409-410).  Non-periodic meshes only (the sphere).
#include <cmath>
#include <vector>

#include "init_atm.h"

namespace mpas {

namespace {

inline void unit3(double* x) {  // mpas_unit_vec_in_r3, mpas_vector_operations.F:95-101
    const double mag = std::sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
    x[0] = x[0] / mag; x[1] = x[1] / mag; x[2] = x[2] / mag;
}

// Gaussian elimination with scaled partial pivoting; A is n x n, column-major as in the Fortran (A(i,j) = a[j*n+i]), 0-based
void elgs(std::vector<double>& a, int n, std::vector<int>& indx) {
    std::vector<double> c(n);
    for (int i = 0; i < n; ++i) indx[i] = i;
    for (int i = 0; i < n; ++i) {
        double c1 = 0.0;
        for (int j = 0; j < n; ++j) c1 = std::max(c1, std::fabs(a[(size_t)j * n + i]));
        c[i] = c1;
    }
    int k = 0;
    for (int j = 0; j < n - 1; ++j) {
        double pi1 = 0.0;
        for (int i = j; i < n; ++i) {
            const double pi = std::fabs(a[(size_t)j * n + indx[i]]) / c[indx[i]];
            if (pi > pi1) { pi1 = pi; k = i; }
        }
        const int itmp = indx[j];
        indx[j] = indx[k];
        indx[k] = itmp;
        for (int i = j + 1; i < n; ++i) {
            const double pj = a[(size_t)j * n + indx[i]] / a[(size_t)j * n + indx[j]];
            a[(size_t)j * n + indx[i]] = pj;
            for (int kk = j + 1; kk < n; ++kk) a[(size_t)kk * n + indx[i]] = a[(size_t)kk * n + indx[i]] - pj * a[(size_t)kk * n + indx[j]];
        }
    }
}

void legs(std::vector<double>& a, int n, std::vector<double>& b, std::vector<double>& x, std::vector<int>& indx) {
    elgs(a, n, indx);
    for (int i = 0; i < n - 1; ++i)
        for (int j = i + 1; j < n; ++j) b[indx[j]] = b[indx[j]] - a[(size_t)i * n + indx[j]] * b[indx[i]];
    x[n - 1] = b[indx[n - 1]] / a[(size_t)(n - 1) * n + indx[n - 1]];
    for (int i = n - 2; i >= 0; --i) {
        x[i] = b[indx[i]];
        for (int j = i + 1; j < n; ++j) x[i] = x[i] - a[(size_t)j * n + indx[i]] * x[j];
        x[i] = x[i] / a[(size_t)i * n + indx[i]];
    }
}

inline double evaluate_rbf(double rSquared) { return 1 / std::sqrt(1 + rSquared); }  // inverse multiquadric

}  // namespace

void init_reconstruct(const Mesh& m, Pool& p) {
    const int nCells = m.nCells, nEdges = m.nEdges, mE = m.maxEdges;
    Field& lvu = p.addr("localVerticalUnitVectors", LOC_CELL, {3, nCells + 1});
    Field& env = p.addr("edgeNormalVectors", LOC_EDGE, {3, nEdges + 1});
    Field& ctp = p.addr("cellTangentPlane", LOC_CELL, {3, 2, nCells + 1});
    Field& cr = p.addr("coeffs_reconstruct", LOC_CELL, {3, mE, nCells + 1});
    // ---- mpas_initialize_vectors
    for (int c = 0; c < nCells; ++c) {
        double* v = &lvu.r[3 * (size_t)c];
        if (m.on_a_sphere) {
            v[0] = m.xCell[c]; v[1] = m.yCell[c]; v[2] = m.zCell[c];
            unit3(v);
        } else {
            v[0] = 0.; v[1] = 0.; v[2] = 1.;
        }
    }
    for (int e = 0; e < nEdges; ++e) {
        const int cell1 = m.cellsOnEdge[2 * (size_t)e], cell2 = m.cellsOnEdge[2 * (size_t)e + 1];
        double* n = &env.r[3 * (size_t)e];
        if (cell1 == nCells + 1) {  // boundary edge: from the edge to cell 2
            n[0] = m.xCell[cell2 - 1] - m.xEdge[e]; n[1] = m.yCell[cell2 - 1] - m.yEdge[e]; n[2] = m.zCell[cell2 - 1] - m.zEdge[e];
        } else if (cell2 == nCells + 1) {  // from cell 1 to the edge
            n[0] = m.xEdge[e] - m.xCell[cell1 - 1]; n[1] = m.yEdge[e] - m.yCell[cell1 - 1]; n[2] = m.zEdge[e] - m.zCell[cell1 - 1];
        } else {  // from cell 1 to cell 2
            n[0] = m.xCell[cell2 - 1] - m.xCell[cell1 - 1]; n[1] = m.yCell[cell2 - 1] - m.yCell[cell1 - 1]; n[2] = m.zCell[cell2 - 1] - m.zCell[cell1 - 1];
        }
        unit3(n);
    }
    for (int c = 0; c < nCells; ++c) {
        const int e = m.edgesOnCell[(size_t)c * mE] - 1;  // xHat points toward the first edge
        const double* rHat = &lvu.r[3 * (size_t)c];
        const double* nv = &env.r[3 * (size_t)e];
        const double normalDotRHat = nv[0] * rHat[0] + nv[1] * rHat[1] + nv[2] * rHat[2];
        double xHat[3] = {nv[0] - normalDotRHat * rHat[0], nv[1] - normalDotRHat * rHat[1], nv[2] - normalDotRHat * rHat[2]};
        unit3(xHat);
        double yHat[3] = {rHat[1] * xHat[2] - rHat[2] * xHat[1], rHat[2] * xHat[0] - rHat[0] * xHat[2], rHat[0] * xHat[1] - rHat[1] * xHat[0]};
        unit3(yHat);
        for (int d = 0; d < 3; ++d) {
            ctp.r[(size_t)c * 6 + d] = xHat[d];
            ctp.r[(size_t)c * 6 + 3 + d] = yHat[d];
        }
    }
    // ---- mpas_init_reconstruct over the owned cells (includeHalos absent)
#pragma omp parallel for schedule(static)
    for (int c = 0; c < m.nCellsSolve; ++c) {
        const int n = m.nEdgesOnCell[c];
        const double cc[3] = {m.xCell[c], m.yCell[c], m.zCell[c]};
        std::vector<double> loc(3 * (size_t)n), nrm(3 * (size_t)n);
        for (int i = 0; i < n; ++i) {
            const int e = m.edgesOnCell[(size_t)c * mE + i] - 1;
            loc[3 * i] = m.xEdge[e]; loc[3 * i + 1] = m.yEdge[e]; loc[3 * i + 2] = m.zEdge[e];
            for (int d = 0; d < 3; ++d) nrm[3 * i + d] = env.r[3 * (size_t)e + d];
        }
        double alpha = 0.0;
        for (int i = 0; i < n; ++i) {
            const double dx = cc[0] - loc[3 * i], dy = cc[1] - loc[3 * i + 1], dz = cc[2] - loc[3 * i + 2];
            alpha = alpha + std::sqrt(dx * dx + dy * dy + dz * dz);
        }
        alpha = alpha / n;
        const double* t1 = &ctp.r[(size_t)c * 6];
        const double* t2 = t1 + 3;
        // planar source points, unit vectors and destination point
        std::vector<double> ps(2 * (size_t)n), pu(2 * (size_t)n);
        for (int i = 0; i < n; ++i) {
            ps[2 * i] = loc[3 * i] * t1[0] + loc[3 * i + 1] * t1[1] + loc[3 * i + 2] * t1[2];
            ps[2 * i + 1] = loc[3 * i] * t2[0] + loc[3 * i + 1] * t2[1] + loc[3 * i + 2] * t2[2];
            pu[2 * i] = nrm[3 * i] * t1[0] + nrm[3 * i + 1] * t1[1] + nrm[3 * i + 2] * t1[2];
            pu[2 * i + 1] = nrm[3 * i] * t2[0] + nrm[3 * i + 1] * t2[1] + nrm[3 * i + 2] * t2[2];
        }
        const double pd[2] = {cc[0] * t1[0] + cc[1] * t1[1] + cc[2] * t1[2], cc[0] * t2[0] + cc[1] * t2[1] + cc[2] * t2[2]};
        const int N = n + 2;  // room for the constant vector in the plane
        std::vector<double> A((size_t)N * N, 0.0), A2, rhs1(N, 0.0), rhs2(N, 0.0), x1(N, 0.0), x2(N, 0.0);
        std::vector<int> piv(N);
        for (int j = 0; j < n; ++j)
            for (int i = j; i < n; ++i) {
                const double dx = ps[2 * i] - ps[2 * j], dy = ps[2 * i + 1] - ps[2 * j + 1];
                const double rSquared = (dx * dx + dy * dy) / (alpha * alpha);
                const double dot = pu[2 * i] * pu[2 * j] + pu[2 * i + 1] * pu[2 * j + 1];
                const double v = evaluate_rbf(rSquared) * dot;
                A[(size_t)j * N + i] = v;
                A[(size_t)i * N + j] = v;
            }
        for (int j = 0; j < n; ++j) {
            const double dx = pd[0] - ps[2 * j], dy = pd[1] - ps[2 * j + 1];
            const double rb = evaluate_rbf((dx * dx + dy * dy) / (alpha * alpha));
            rhs1[j] = rb * pu[2 * j];
            rhs2[j] = rb * pu[2 * j + 1];
        }
        for (int i = 0; i < n; ++i)
            for (int d = 0; d < 2; ++d) {
                A[(size_t)(n + d) * N + i] = pu[2 * i + d];
                A[(size_t)i * N + (n + d)] = pu[2 * i + d];
            }
        rhs1[n] = 1.0;      // the unit vector in the first plane direction
        rhs2[n + 1] = 1.0;  // ... and the second
        A2 = A;
        legs(A, N, rhs1, x1, piv);
        legs(A2, N, rhs2, x2, piv);
        for (int i = 0; i < n; ++i)
            for (int d = 0; d < 3; ++d) cr.r[((size_t)c * mE + i) * 3 + d] = t1[d] * x1[i] + t2[d] * x2[i];
    }
}

}  // namespace mpas
