// Core-init derived arrays, restating src/core_atmosphere/mpas_atm_core.F:
// atm_compute_signs :987-1074, inverse geometry :339-353, atm_adv_coef_compression :1121-1266,
// atm_couple_coef_3rd_order :1269-1288, atm_compute_mesh_scaling :927-984,
// atm_compute_damping_coefs :1077-1118; plus pool plumbing (mesh -> Registry-named fields).
#include <cmath>

#include "init_atm.h"

namespace mpas {

void scale_mesh_to_sphere(Mesh& m, double R) {  // mpas_init_atm_cases.F:497-510
    auto S = [&](std::vector<double>& v, double f) { for (auto& x : v) x = x * f; };
    S(m.xCell, R); S(m.yCell, R); S(m.zCell, R);
    S(m.xVertex, R); S(m.yVertex, R); S(m.zVertex, R);
    S(m.xEdge, R); S(m.yEdge, R); S(m.zEdge, R);
    S(m.dvEdge, R); S(m.dcEdge, R);
    S(m.areaCell, R * R); S(m.areaTriangle, R * R); S(m.kiteAreasOnVertex, R * R);
    m.sphere_radius = m.sphere_radius * R;
}

static void putr(Pool& p, const char* name, FieldLoc loc, std::vector<int64_t> dims, const std::vector<double>& v) {
    Field& f = p.addr(name, loc, dims);
    if ((size_t)f.count() != v.size()) throw std::runtime_error(std::string("mesh_to_pool size mismatch: ") + name);
    f.r = v;
}
static void puti(Pool& p, const char* name, FieldLoc loc, std::vector<int64_t> dims, const std::vector<int>& v) {
    Field& f = p.addi(name, loc, dims);
    if ((size_t)f.count() != v.size()) throw std::runtime_error(std::string("mesh_to_pool size mismatch: ") + name);
    f.i.assign(v.begin(), v.end());
}

void mesh_to_pool(const Mesh& m, Pool& p) {
    const int nC = m.nCells, nE = m.nEdges, nV = m.nVertices, mE = m.maxEdges, mE2 = m.maxEdges2;
    p.cfg["nCells"] = nC; p.cfg["nEdges"] = nE; p.cfg["nVertices"] = nV;
    p.cfg["nCellsSolve"] = m.nCellsSolve; p.cfg["nEdgesSolve"] = m.nEdgesSolve; p.cfg["nVerticesSolve"] = m.nVerticesSolve;
    p.cfg["maxEdges"] = mE; p.cfg["maxEdges2"] = mE2; p.cfg["vertexDegree"] = 3;
    p.cfg["sphere_radius"] = m.sphere_radius; p.cfg["on_a_sphere"] = m.on_a_sphere ? 1 : 0;
    puti(p, "indexToCellID", LOC_CELL, {nC + 1}, m.indexToCellID);
    puti(p, "indexToEdgeID", LOC_EDGE, {nE + 1}, m.indexToEdgeID);
    puti(p, "indexToVertexID", LOC_VERTEX, {nV + 1}, m.indexToVertexID);
    puti(p, "nEdgesOnCell", LOC_CELL, {nC + 1}, m.nEdgesOnCell);
    puti(p, "cellsOnCell", LOC_CELL, {mE, nC + 1}, m.cellsOnCell);
    puti(p, "edgesOnCell", LOC_CELL, {mE, nC + 1}, m.edgesOnCell);
    puti(p, "verticesOnCell", LOC_CELL, {mE, nC + 1}, m.verticesOnCell);
    puti(p, "cellsOnEdge", LOC_EDGE, {2, nE + 1}, m.cellsOnEdge);
    puti(p, "verticesOnEdge", LOC_EDGE, {2, nE + 1}, m.verticesOnEdge);
    puti(p, "nEdgesOnEdge", LOC_EDGE, {nE + 1}, m.nEdgesOnEdge);
    puti(p, "edgesOnEdge", LOC_EDGE, {mE2, nE + 1}, m.edgesOnEdge);
    puti(p, "cellsOnVertex", LOC_VERTEX, {3, nV + 1}, m.cellsOnVertex);
    puti(p, "edgesOnVertex", LOC_VERTEX, {3, nV + 1}, m.edgesOnVertex);
    putr(p, "latCell", LOC_CELL, {nC + 1}, m.latCell); putr(p, "lonCell", LOC_CELL, {nC + 1}, m.lonCell);
    putr(p, "latEdge", LOC_EDGE, {nE + 1}, m.latEdge); putr(p, "lonEdge", LOC_EDGE, {nE + 1}, m.lonEdge);
    putr(p, "latVertex", LOC_VERTEX, {nV + 1}, m.latVertex); putr(p, "lonVertex", LOC_VERTEX, {nV + 1}, m.lonVertex);
    putr(p, "xCell", LOC_CELL, {nC + 1}, m.xCell); putr(p, "yCell", LOC_CELL, {nC + 1}, m.yCell); putr(p, "zCell", LOC_CELL, {nC + 1}, m.zCell);
    putr(p, "dcEdge", LOC_EDGE, {nE + 1}, m.dcEdge); putr(p, "dvEdge", LOC_EDGE, {nE + 1}, m.dvEdge);
    putr(p, "angleEdge", LOC_EDGE, {nE + 1}, m.angleEdge);
    putr(p, "areaCell", LOC_CELL, {nC + 1}, m.areaCell); putr(p, "areaTriangle", LOC_VERTEX, {nV + 1}, m.areaTriangle);
    putr(p, "meshDensity", LOC_CELL, {nC + 1}, m.meshDensity);
    putr(p, "kiteAreasOnVertex", LOC_VERTEX, {3, nV + 1}, m.kiteAreasOnVertex);
    putr(p, "weightsOnEdge", LOC_EDGE, {mE2, nE + 1}, m.weightsOnEdge);
}

void core_init_derived(const Mesh& m, const InitConfig& cfg, Pool& p) {
    const int nCells = m.nCells, nEdges = m.nEdges, nVertices = m.nVertices, mE = m.maxEdges, L = cfg.nVertLevels;
    // ---- atm_compute_signs (:1023-1072)
    Field& eovs = p.addr("edgesOnVertex_sign", LOC_VERTEX, {3, nVertices + 1});
    Field& eocs = p.addr("edgesOnCell_sign", LOC_CELL, {mE, nCells + 1});
    Field& kfc = p.addi("kiteForCell", LOC_CELL, {mE, nCells + 1});
    for (int v = 1; v <= nVertices; ++v)
        for (int i = 1; i <= 3; ++i) {
            int e = m.edgesOnVertex[3 * (v - 1) + i - 1];
            double s = 0.0;
            if (e <= nEdges) s = (v == m.verticesOnEdge[2 * (e - 1) + 1]) ? 1.0 : -1.0;
            eovs.r[3 * (v - 1) + i - 1] = s;
        }
    for (int c = 1; c <= nCells; ++c)
        for (int i = 1; i <= m.nEdgesOnCell[c - 1]; ++i) {
            int e = m.edgesOnCell[(size_t)(c - 1) * mE + i - 1];
            double s = 0.0;
            if (e <= nEdges) s = (c == m.cellsOnEdge[2 * (e - 1)]) ? 1.0 : -1.0;
            eocs.r[(size_t)(c - 1) * mE + i - 1] = s;
            int iv = m.verticesOnCell[(size_t)(c - 1) * mE + i - 1];
            int kite = 1;
            if (iv <= nVertices)
                for (int j = 1; j <= 3; ++j)
                    if (c == m.cellsOnVertex[3 * (iv - 1) + j - 1]) { kite = j; break; }
            kfc.i[(size_t)(c - 1) * mE + i - 1] = kite;
        }
    // ---- inverses (:339-353)
    Field& iac = p.addr("invAreaCell", LOC_CELL, {nCells + 1});
    Field& idv = p.addr("invDvEdge", LOC_EDGE, {nEdges + 1});
    Field& idc = p.addr("invDcEdge", LOC_EDGE, {nEdges + 1});
    Field& iat = p.addr("invAreaTriangle", LOC_VERTEX, {nVertices + 1});
    for (int c = 0; c < nCells; ++c) iac.r[c] = 1.0 / m.areaCell[c];
    for (int e = 0; e < nEdges; ++e) { idv.r[e] = 1.0 / m.dvEdge[e]; idc.r[e] = 1.0 / m.dcEdge[e]; }
    for (int v = 0; v < nVertices; ++v) iat.r[v] = 1.0 / m.areaTriangle[v];

    // ---- atm_adv_coef_compression (:1154-1264)
    Field& fn = p.addi("nAdvCellsForEdge", LOC_EDGE, {nEdges + 1});
    Field& fa = p.addi("advCellsForEdge", LOC_EDGE, {15, nEdges + 1});
    Field& fc = p.addr("adv_coefs", LOC_EDGE, {15, nEdges + 1});
    Field& fc3 = p.addr("adv_coefs_3rd", LOC_EDGE, {15, nEdges + 1});
    const double* deriv_two = p.R("deriv_two");
    auto D2 = [&](int j, int s, int e) { return deriv_two[(size_t)(e - 1) * 30 + (s - 1) * 15 + (j - 1)]; };
    auto COC = [&](int i, int c) { return m.cellsOnCell[(size_t)(c - 1) * mE + (i - 1)]; };
#pragma omp parallel for schedule(static)
    for (int iEdge = 1; iEdge <= nEdges; ++iEdge) {
        int* adv = &fa.i[(size_t)(iEdge - 1) * 15];
        double* ac = &fc.r[(size_t)(iEdge - 1) * 15];
        double* ac3 = &fc3.r[(size_t)(iEdge - 1) * 15];
        for (int j = 0; j < 15; ++j) adv[j] = nCells + 1;  // unused slots -> garbage cell
        int cell1 = m.cellsOnEdge[2 * (iEdge - 1)], cell2 = m.cellsOnEdge[2 * (iEdge - 1) + 1];
        if (!(cell1 <= nCells || cell2 <= nCells)) continue;
        int cell_list[21];
        cell_list[1] = cell1; cell_list[2] = cell2;
        int n = 2;
        int n1 = m.nEdgesOnCell[cell1 - 1], n2 = m.nEdgesOnCell[cell2 - 1];
        for (int i = 1; i <= n1; ++i)
            if (COC(i, cell1) != cell2) { n = n + 1; cell_list[n] = COC(i, cell1); }
        for (int ic = 1; ic <= n2; ++ic) {
            bool addcell = true;
            for (int i = 1; i <= n; ++i)
                if (cell_list[i] == COC(ic, cell2)) addcell = false;
            if (addcell) { n = n + 1; cell_list[n] = COC(ic, cell2); }
        }
        if (n > 15) n = 15;
        fn.i[iEdge - 1] = n;
        for (int i = 1; i <= n; ++i) adv[i - 1] = cell_list[i];
        for (int j = 0; j < 15; ++j) { ac[j] = 0.; ac3[j] = 0.; }
        auto find = [&](int c) { int j_in = 0; for (int j = 1; j <= n; ++j) if (cell_list[j] == c) j_in = j; return j_in; };
        int j_in = find(cell1);
        ac[j_in - 1] = ac[j_in - 1] + D2(1, 1, iEdge);
        ac3[j_in - 1] = ac3[j_in - 1] + D2(1, 1, iEdge);
        for (int ic = 1; ic <= n1; ++ic) {
            j_in = find(COC(ic, cell1));
            ac[j_in - 1] = ac[j_in - 1] + D2(ic + 1, 1, iEdge);
            ac3[j_in - 1] = ac3[j_in - 1] + D2(ic + 1, 1, iEdge);
        }
        j_in = find(cell2);
        ac[j_in - 1] = ac[j_in - 1] + D2(1, 2, iEdge);
        ac3[j_in - 1] = ac3[j_in - 1] - D2(1, 2, iEdge);
        for (int ic = 1; ic <= n2; ++ic) {
            j_in = find(COC(ic, cell2));
            ac[j_in - 1] = ac[j_in - 1] + D2(ic + 1, 2, iEdge);
            ac3[j_in - 1] = ac3[j_in - 1] - D2(ic + 1, 2, iEdge);
        }
        double dc = m.dcEdge[iEdge - 1], dv = m.dvEdge[iEdge - 1];
        for (int j = 0; j < n; ++j) {
            ac[j] = -(dc * dc) * ac[j] / 12.;
            ac3[j] = -(dc * dc) * ac3[j] / 12.;
        }
        j_in = find(cell1); ac[j_in - 1] = ac[j_in - 1] + 0.5;
        j_in = find(cell2); ac[j_in - 1] = ac[j_in - 1] + 0.5;
        for (int j = 0; j < n; ++j) { ac[j] = dv * ac[j]; ac3[j] = dv * ac3[j]; }
        // atm_couple_coef_3rd_order (:1285)
        for (int j = 0; j < 15; ++j) ac3[j] = cfg.coef_3rd_order * ac3[j];
    }
    // atm_couple_coef_3rd_order (:1286) for zb3_cell
    if (p.has("zb3_cell")) {
        Field& z3 = p.get("zb3_cell");
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < (int64_t)z3.r.size(); ++i) z3.r[i] = cfg.coef_3rd_order * z3.r[i];
    }
    // ---- atm_compute_mesh_scaling (:956-965)
    Field& ms2 = p.addr("meshScalingDel2", LOC_EDGE, {nEdges + 1});
    Field& ms4 = p.addr("meshScalingDel4", LOC_EDGE, {nEdges + 1});
    for (int e = 0; e <= nEdges; ++e) { ms2.r[e] = 1.0; ms4.r[e] = 1.0; }
    if (cfg.h_ScaleWithMesh)
        for (int e = 0; e < nEdges; ++e) {
            int c1 = m.cellsOnEdge[2 * e] - 1, c2 = m.cellsOnEdge[2 * e + 1] - 1;
            ms2.r[e] = 1.0 / std::pow((m.meshDensity[c1] + m.meshDensity[c2]) / 2.0, 0.25);
            ms4.r[e] = 1.0 / std::pow((m.meshDensity[c1] + m.meshDensity[c2]) / 2.0, 0.75);
        }
    // ---- atm_compute_damping_coefs (:1105-1116)
    Field& dss = p.addr("dss", LOC_CELL, {L, nCells + 1});
    const double* zgrid = p.R("zgrid");
    const double pii = std::acos(-1.0);
#pragma omp parallel for schedule(static)
    for (int c = 0; c < nCells; ++c) {
        double zt = zgrid[(size_t)c * (L + 1) + L];
        for (int k = 0; k < L; ++k) {
            double z = 0.5 * (zgrid[(size_t)c * (L + 1) + k] + zgrid[(size_t)c * (L + 1) + k + 1]);
            if (z > cfg.config_zd) {
                double s = std::sin(0.5 * pii * (z - cfg.config_zd) / (zt - cfg.config_zd));
                double d = cfg.config_xnutr * s * s;
                dss.r[(size_t)c * L + k] = d / std::pow(m.meshDensity[c], 0.25 * 1.0);
            }
        }
    }
    // regional masks: global meshes have none (mpas_atm_setup_bdy_masks with config_apply_lbcs=false)
    p.addi("bdyMaskCell", LOC_CELL, {nCells + 1});
    p.addi("bdyMaskEdge", LOC_EDGE, {nEdges + 1});
    p.addr("specZoneMaskCell", LOC_CELL, {nCells + 1});
    p.addr("specZoneMaskEdge", LOC_EDGE, {nEdges + 1});
}

// State/diag/tend fields of the dycore (Registry.xml:1412 state, :1506 diag, :1709 tend).
// full=false allocates only what a device-resident run still needs on the host.
// full = 2 ("lean"): the same fields, but without host storage (Field::zero) -- they are all zero at this point, the last
// step of building a block, and the device mirrors are created zero-filled from the metadata; a 15 km block then needs
// 3 GB of host memory instead of 20.
void allocate_dycore_fields(const Mesh& m, const InitConfig& cfg, Pool& p, int full) {
    const int nC = m.nCells + 1, nE = m.nEdges + 1, nV = m.nVertices + 1, L = cfg.nVertLevels, ns = cfg.num_scalars;
    auto add = [&](const char* n, FieldLoc loc, std::vector<int64_t> dims) {
        if (p.has(n)) return;
        if (full == 2) p.addz(n, loc, dims);
        else p.addr(n, loc, dims);
    };
    auto C = [&](const char* n) { add(n, LOC_CELL, {L, nC}); };
    auto C1 = [&](const char* n) { add(n, LOC_CELL, {L + 1, nC}); };
    auto E = [&](const char* n) { add(n, LOC_EDGE, {L, nE}); };
    auto V = [&](const char* n) { add(n, LOC_VERTEX, {L, nV}); };
    if (!p.has("scalars@1")) p.addr("scalars@1", LOC_CELL, {ns, L, nC});
    if (!full) return;
    for (const char* n : {"u@2", "ru", "ru_save", "ru_p", "ruAvg", "ruAvg_split", "rho_edge", "pv_edge", "v", "cqu", "gradPVn", "gradPVt",
                          "tend_u", "tend_u_euler", "tend_ru_physics", "delsq_u", "ke_edge", "flux_arr", "flux_upwind_tmp", "flux_tmp"}) E(n);
    for (const char* n : {"w@2", "rw", "rw_save", "rw_p", "wwAvg", "wwAvg_split", "coftz", "tend_w", "tend_w_euler", "wdtn"}) C1(n);
    for (const char* n : {"theta_m@1", "theta_m@2", "rho_zz@1", "rho_zz@2", "rtheta_p", "rtheta_p_save", "rtheta_pp", "rtheta_pp_old",
                          "rtheta_base", "rho_p", "rho_p_save", "rho_pp", "rho_zz_old_split", "exner", "exner_base", "pressure_p",
                          "pressure_base", "cqw", "cofwr", "cofwz", "cofwt", "a_tri", "alpha_tri", "gamma_tri", "h_divergence", "kdiff",
                          "divergence", "ke", "pv_cell", "qtot", "dpdz", "delsq_theta", "delsq_w", "delsq_divergence", "tend_rtheta_adv",
                          "rthdynten", "tend_rho_physics", "tend_rtheta_physics", "tend_theta", "tend_theta_euler", "tend_rho",
                          "rt_diabatic_tend", "scalar_old", "scalar_new", "s_max", "s_min", "rho_zz_int", "uReconstructX", "uReconstructY",
                          "uReconstructZ", "uReconstructZonal", "uReconstructMeridional", "pressure"}) C(n);
    for (const char* n : {"vorticity", "pv_vertex", "ke_vertex", "delsq_vorticity"}) V(n);
    add("scalars@2", LOC_CELL, {ns, L, nC});
    add("scalars_tend", LOC_CELL, {ns, L, nC});
    add("scale_arr", LOC_CELL, {L, 2, nC});
    add("horiz_flux_arr", LOC_EDGE, {ns, L, nE});
    add("cofrz", LOC_NONE, {L});
}

}  // namespace mpas
