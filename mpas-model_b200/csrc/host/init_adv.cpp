// Restatement of src/core_init_atmosphere/mpas_atm_advection.F: second-derivative stencil
// coefficients deriv_two (atm_initialize_advection_rk :21-394) and Smagorinsky deformation
// weights defc_a/defc_b (atm_initialize_deformation_weights :744-937), with the helpers
// sphere_angle :403, plane_angle :456, arc_length :506, arc_bisect :539, poly_fit_2 :567,
// migs :633, elgs :678.  Per-cell computations: independent of numbering and decomposition.
#include <algorithm>
#include <cmath>
#include <vector>

#include "init_atm.h"

namespace mpas {
namespace {

double arc_length(double ax, double ay, double az, double bx, double by, double bz) {  // :506-529
    double cx = bx - ax, cy = by - ay, cz = bz - az;
    double r = std::sqrt(ax * ax + ay * ay + az * az);
    double c = std::sqrt(cx * cx + cy * cy + cz * cz);
    return r * 2.0 * std::asin(c / (2.0 * r));
}

double sphere_angle(double ax, double ay, double az, double bx, double by, double bz, double cx, double cy,
                    double cz) {  // :403-447
    double a = arc_length(bx, by, bz, cx, cy, cz);
    double b = arc_length(ax, ay, az, cx, cy, cz);
    double c = arc_length(ax, ay, az, bx, by, bz);
    double ABx = bx - ax, ABy = by - ay, ABz = bz - az;
    double ACx = cx - ax, ACy = cy - ay, ACz = cz - az;
    double Dx = (ABy * ACz) - (ABz * ACy);
    double Dy = -((ABx * ACz) - (ABz * ACx));
    double Dz = (ABx * ACy) - (ABy * ACx);
    double s = 0.5 * (a + b + c);
    double sin_angle = std::sqrt(std::min(1.0, std::max(0.0, (std::sin(s - b) * std::sin(s - c)) / (std::sin(b) * std::sin(c)))));
    if ((Dx * ax + Dy * ay + Dz * az) >= 0.0) return 2.0 * std::asin(std::max(std::min(sin_angle, 1.0), -1.0));
    return -2.0 * std::asin(std::max(std::min(sin_angle, 1.0), -1.0));
}

double plane_angle(double ax, double ay, double az, double bx, double by, double bz, double cx, double cy, double cz,
                   double u, double v, double w) {  // :456-496
    double ABx = bx - ax, ABy = by - ay, ABz = bz - az;
    double mAB = std::sqrt(ABx * ABx + ABy * ABy + ABz * ABz);
    double ACx = cx - ax, ACy = cy - ay, ACz = cz - az;
    double mAC = std::sqrt(ACx * ACx + ACy * ACy + ACz * ACz);
    double Dx = (ABy * ACz) - (ABz * ACy);
    double Dy = -((ABx * ACz) - (ABz * ACx));
    double Dz = (ABx * ACy) - (ABy * ACx);
    double cos_angle = (ABx * ACx + ABy * ACy + ABz * ACz) / (mAB * mAC);
    if ((Dx * u + Dy * v + Dz * w) >= 0.0) return std::acos(std::max(std::min(cos_angle, 1.0), -1.0));
    return -std::acos(std::max(std::min(cos_angle, 1.0), -1.0));
}

void arc_bisect(double ax, double ay, double az, double bx, double by, double bz, double& cx, double& cy,
                double& cz) {  // :539-564
    double r = std::sqrt(ax * ax + ay * ay + az * az);
    cx = 0.5 * (ax + bx); cy = 0.5 * (ay + by); cz = 0.5 * (az + bz);
    double d = std::sqrt(cx * cx + cy * cy + cz * cz);
    cx = r * cx / d; cy = r * cy / d; cz = r * cz / d;
}

// ELGS :678-741 (scaled partial pivoting), A is n x n row-major a[i*n+j], 0-based
void elgs(std::vector<double>& A, int n, std::vector<int>& indx) {
    std::vector<double> C(n);
    for (int i = 0; i < n; ++i) indx[i] = i;
    for (int i = 0; i < n; ++i) {
        double c1 = 0.0;
        for (int j = 0; j < n; ++j) c1 = std::max(c1, std::fabs(A[i * n + j]));
        C[i] = c1;
    }
    int k = 0;
    for (int j = 0; j < n - 1; ++j) {
        double pi1 = 0.0;
        for (int i = j; i < n; ++i) {
            double pi = std::fabs(A[indx[i] * n + j]) / C[indx[i]];
            if (pi > pi1) { pi1 = pi; k = i; }
        }
        std::swap(indx[j], indx[k]);
        for (int i = j + 1; i < n; ++i) {
            double pj = A[indx[i] * n + j] / A[indx[j] * n + j];
            A[indx[i] * n + j] = pj;
            for (int kk = j + 1; kk < n; ++kk) A[indx[i] * n + kk] = A[indx[i] * n + kk] - pj * A[indx[j] * n + kk];
        }
    }
}

// MIGS :633-675: X = inverse(A)
void migs(std::vector<double>& A, int n, std::vector<double>& X) {
    std::vector<double> B((size_t)n * n, 0.0);
    std::vector<int> indx(n);
    for (int i = 0; i < n; ++i) B[i * n + i] = 1.0;
    elgs(A, n, indx);
    for (int i = 0; i < n - 1; ++i)
        for (int j = i + 1; j < n; ++j)
            for (int k = 0; k < n; ++k) B[indx[j] * n + k] = B[indx[j] * n + k] - A[indx[j] * n + i] * B[indx[i] * n + k];
    for (int i = 0; i < n; ++i) {
        X[(n - 1) * n + i] = B[indx[n - 1] * n + i] / A[indx[n - 1] * n + (n - 1)];
        for (int j = n - 2; j >= 0; --j) {
            X[j * n + i] = B[indx[j] * n + i];
            for (int k = j + 1; k < n; ++k) X[j * n + i] = X[j * n + i] - A[indx[j] * n + k] * X[k * n + i];
            X[j * n + i] = X[j * n + i] / A[indx[j] * n + j];
        }
    }
}

// poly_fit_2 :567-614 with unit weights: b(n x m) = (A^T A)^-1 A^T ; a is m x n
void poly_fit_2(const double amatrix[25][25], double bmatrix[25][25], int m, int n) {
    std::vector<double> atha((size_t)n * n, 0.0), inv((size_t)n * n, 0.0);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            double s = 0.0;
            for (int k = 0; k < m; ++k) s += amatrix[k][i] * amatrix[k][j];
            atha[i * n + j] = s;
        }
    migs(atha, n, inv);
    for (int i = 0; i < 25; ++i)
        for (int j = 0; j < 25; ++j) bmatrix[i][j] = 0.0;
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < m; ++j) {
            double s = 0.0;
            for (int k = 0; k < n; ++k) s += inv[i * n + k] * amatrix[j][k];
            bmatrix[i][j] = s;
        }
}

}  // namespace

void init_advection_and_deformation(const Mesh& m, Pool& p) {
    const int nC = m.nCells, nE = m.nEdges, mE = m.maxEdges;
    const double R = m.sphere_radius;
    const double pii = 2. * std::asin(1.0);
    Field& d2 = p.addr("deriv_two", LOC_EDGE, {15, 2, nE + 1});
    Field& fa = p.addr("defc_a", LOC_CELL, {mE, nC + 1});
    Field& fb = p.addr("defc_b", LOC_CELL, {mE, nC + 1});
    double* deriv_two = d2.r.data();
    double* defc_a = fa.r.data();
    double* defc_b = fb.r.data();
    auto COC = [&](int i, int c) { return m.cellsOnCell[(size_t)(c - 1) * mE + (i - 1)]; };
    auto EOC = [&](int i, int c) { return m.edgesOnCell[(size_t)(c - 1) * mE + (i - 1)]; };
    auto VOC = [&](int i, int c) { return m.verticesOnCell[(size_t)(c - 1) * mE + (i - 1)]; };

#pragma omp parallel for schedule(dynamic, 256)
    for (int iCell = 1; iCell <= nC; ++iCell) {
        int cell_list[26];
        double xc[26], yc[26], zc[26], thetav[26], thetat[26], dl_sphere[26], xp[26], yp[26];
        double amatrix[25][25], bmatrix[25][25];
        int nEoC = m.nEdgesOnCell[iCell - 1];
        cell_list[1] = iCell;
        for (int i = 2; i <= nEoC + 1; ++i) cell_list[i] = COC(i - 1, iCell);
        int n = nEoC + 1;
        bool do_the_cell = true;
        for (int i = 1; i <= n; ++i)
            if (cell_list[i] > nC) do_the_cell = false;
        if (!do_the_cell) continue;

        // ---------------- deriv_two (:129-390)
        double theta_abs = 0.0;
        double angle_2d[26];
        if (m.on_a_sphere) {
            for (int i = 1; i <= n; ++i) {
                int c = cell_list[i] - 1;
                xc[i] = m.xCell[c] / R; yc[i] = m.yCell[c] / R; zc[i] = m.zCell[c] / R;
            }
            if (zc[1] == 1.0) theta_abs = pii / 2.;
            else theta_abs = pii / 2. - sphere_angle(xc[1], yc[1], zc[1], xc[2], yc[2], zc[2], 0.0, 0.0, 1.0);
            for (int i = 1; i <= n - 1; ++i) {
                int ip2 = i + 2;
                if (ip2 > n) ip2 = 2;
                thetav[i] = sphere_angle(xc[1], yc[1], zc[1], xc[i + 1], yc[i + 1], zc[i + 1], xc[ip2], yc[ip2], zc[ip2]);
                dl_sphere[i] = R * arc_length(xc[1], yc[1], zc[1], xc[i + 1], yc[i + 1], zc[i + 1]);
            }
            thetat[1] = theta_abs;
            for (int i = 2; i <= n - 1; ++i) thetat[i] = thetat[i - 1] + thetav[i - 1];
            for (int i = 1; i <= n - 1; ++i) {
                xp[i] = std::cos(thetat[i]) * dl_sphere[i];
                yp[i] = std::sin(thetat[i]) * dl_sphere[i];
            }
        } else {  // on an x-y plane (:184-200): neighbour positions from angleEdge / dcEdge, safe across periodic boundaries
            bool edges_ok = true;
            for (int i = 1; i <= n - 1; ++i) {
                const int iEdge = EOC(i, iCell);
                if (iEdge > nE) { edges_ok = false; break; }
                angle_2d[i] = m.angleEdge[iEdge - 1];
                if (iCell != m.cellsOnEdge[2 * (iEdge - 1)]) angle_2d[i] = angle_2d[i] - pii;
                xp[i] = m.dcEdge[iEdge - 1] * std::cos(angle_2d[i]);
                yp[i] = m.dcEdge[iEdge - 1] * std::sin(angle_2d[i]);
            }
            if (!edges_ok) continue;
        }
        int ma = n - 1;
        for (int i = 0; i < 25; ++i)
            for (int j = 0; j < 25; ++j) amatrix[i][j] = 0.0;
        const int na = 6;
        ma = ma + 1;
        amatrix[0][0] = 1.;
        for (int i = 2; i <= ma; ++i) {
            amatrix[i - 1][0] = 1.;
            amatrix[i - 1][1] = xp[i - 1];
            amatrix[i - 1][2] = yp[i - 1];
            amatrix[i - 1][3] = xp[i - 1] * xp[i - 1];
            amatrix[i - 1][4] = xp[i - 1] * yp[i - 1];
            amatrix[i - 1][5] = yp[i - 1] * yp[i - 1];
        }
        poly_fit_2(amatrix, bmatrix, ma, na);
        for (int i = 1; i <= nEoC; ++i) {
            int iEdge = EOC(i, iCell);
            if (iEdge > nE) continue;
            double thetae;
            if (m.on_a_sphere) {
                int v1 = m.verticesOnEdge[2 * (iEdge - 1)] - 1, v2 = m.verticesOnEdge[2 * (iEdge - 1) + 1] - 1;
                double xec, yec, zec;
                arc_bisect(m.xVertex[v1] / R, m.yVertex[v1] / R, m.zVertex[v1] / R, m.xVertex[v2] / R, m.yVertex[v2] / R,
                           m.zVertex[v2] / R, xec, yec, zec);
                thetae = sphere_angle(xc[1], yc[1], zc[1], xc[i + 1], yc[i + 1], zc[i + 1], xec, yec, zec);
                thetae = thetae + thetat[i];
            } else {
                thetae = angle_2d[i];  // :358-362
            }
            double cos2t = std::cos(thetae), sin2t = std::sin(thetae);
            double costsint = cos2t * sin2t;
            cos2t = cos2t * cos2t;
            sin2t = sin2t * sin2t;
            int slot = (iCell == m.cellsOnEdge[2 * (iEdge - 1)]) ? 0 : 1;
            for (int j = 1; j <= n; ++j)
                deriv_two[(size_t)(iEdge - 1) * 30 + slot * 15 + (j - 1)] =
                    2. * cos2t * bmatrix[3][j - 1] + 2. * costsint * bmatrix[4][j - 1] + 2. * sin2t * bmatrix[5][j - 1];
        }

        // ---------------- deformation weights (:821-933)
        bool verts_ok = true;
        for (int i = 2; i <= n; ++i)
            if (VOC(i - 1, iCell) > m.nVertices) verts_ok = false;
        if (!verts_ok) continue;
        if (m.on_a_sphere) {
            xc[1] = m.xCell[iCell - 1] / R; yc[1] = m.yCell[iCell - 1] / R; zc[1] = m.zCell[iCell - 1] / R;
            for (int i = 2; i <= n; ++i) {
                int iv = VOC(i - 1, iCell);
                xc[i] = m.xVertex[iv - 1] / R; yc[i] = m.yVertex[iv - 1] / R; zc[i] = m.zVertex[iv - 1] / R;
            }
            if (zc[1] == 1.0) theta_abs = pii / 2.;
            else theta_abs = pii / 2. - sphere_angle(xc[1], yc[1], zc[1], xc[2], yc[2], zc[2], 0.0, 0.0, 1.0);
            for (int i = 1; i <= n - 1; ++i) {
                int ip2 = i + 2;
                if (ip2 > n) ip2 = 2;
                thetav[i] = sphere_angle(xc[1], yc[1], zc[1], xc[i + 1], yc[i + 1], zc[i + 1], xc[ip2], yc[ip2], zc[ip2]);
                dl_sphere[i] = R * arc_length(xc[1], yc[1], zc[1], xc[i + 1], yc[i + 1], zc[i + 1]);
            }
            thetat[1] = 0.;
            for (int i = 2; i <= n - 1; ++i) thetat[i] = thetat[i - 1] + thetav[i - 1];
            for (int i = 1; i <= n - 1; ++i) {
                xp[i] = std::cos(thetat[i]) * dl_sphere[i];
                yp[i] = std::sin(thetat[i]) * dl_sphere[i];
            }
        } else {
            // on an x-y plane, literally as the reference has it (:889-901): xp(1) is the CELL CENTRE and xp(2..n) the vertices in raw
            // mesh coordinates, while the loops below use xp(1..n-1) as the polygon.  The reference's own planar cases run with
            // config_horiz_mixing = '2d_fixed' (the supercell here too), which never reads defc_a / defc_b.
            theta_abs = 0.0;
            xp[1] = m.xCell[iCell - 1];
            yp[1] = m.yCell[iCell - 1];
            for (int i = 2; i <= n; ++i) {
                int iv = VOC(i - 1, iCell);
                xp[i] = m.xVertex[iv - 1];
                yp[i] = m.yVertex[iv - 1];
            }
        }
        thetat[1] = theta_abs;
        for (int i = 2; i <= n - 1; ++i) {
            int ip1 = i + 1;
            if (ip1 == n) ip1 = 1;
            thetat[i] = plane_angle(0., 0., 0., xp[i] - xp[i - 1], yp[i] - yp[i - 1], 0., xp[ip1] - xp[i], yp[ip1] - yp[i],
                                    0., 0., 0., 1.);
            thetat[i] = thetat[i] + thetat[i - 1];
        }
        double area_cell = 0.;
        for (int i = 1; i <= n - 1; ++i) {
            int ip1 = i + 1;
            if (ip1 == n) ip1 = 1;
            area_cell = area_cell + 0.25 * (xp[i] + xp[ip1]) * (yp[ip1] - yp[i]) - 0.25 * (yp[i] + yp[ip1]) * (xp[ip1] - xp[i]);
        }
        for (int i = 1; i <= n - 1; ++i) {
            int ip1 = i + 1;
            if (ip1 == n) ip1 = 1;
            double dl = std::sqrt((xp[ip1] - xp[i]) * (xp[ip1] - xp[i]) + (yp[ip1] - yp[i]) * (yp[ip1] - yp[i]));
            double sint2 = std::sin(thetat[i]) * std::sin(thetat[i]);
            double cost2 = std::cos(thetat[i]) * std::cos(thetat[i]);
            double sint_cost = std::sin(thetat[i]) * std::cos(thetat[i]);
            double da = dl * (cost2 - sint2) / area_cell;
            double db = dl * 2. * sint_cost / area_cell;
            int e = EOC(i, iCell);
            if (e <= nE && m.cellsOnEdge[2 * (e - 1)] != iCell) { da = -da; db = -db; }
            defc_a[(size_t)(iCell - 1) * mE + (i - 1)] = da;
            defc_b[(size_t)(iCell - 1) * mE + (i - 1)] = db;
        }
    }
}

}  // namespace mpas
