// CPU ORACLE -- TEST INFRASTRUCTURE ONLY (see dycore_oracle.cpp header; parity unpinned by the
// reference's own tests).  Driver restating the call sequence of atm_srk3
// (TI = src/core_atmosphere/dynamics/mpas_atm_time_integration.F:142-1796), the start-up sequence of
// atm_mpas_init_block / atm_core_init (src/core_atmosphere/mpas_atm_core.F:145-186, 311-402) and the
// time-level shift (mpas_atm_core.F:671), for P "virtual ranks" (blocks) in one process.  Halo
// exchanges follow mpas_dmpar_exch_halo_field (src/framework/mpas_dmpar.F:5386-5552, 5565-5748):
// pack buffer((destList-1)*dim1+j) = array(j,srcList), unpack array(j,destList) = buffer((srcList-1)*dim1+j),
// with a memcpy between blocks in place of MPI (the reference's own no-MPI path, :5481-5504).
#include <omp.h>

#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <list>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "../include/mpas_b200_args.h"

typedef mpas_real real;

extern "C" {
int oracle_compute_output_diagnostics(MPAS_OUTPUTDIAG_ARGS);
int oracle_rk_integration_setup(MPAS_RKSETUP_ARGS);
int oracle_compute_moist_coefficients(MPAS_MOIST_ARGS);
int oracle_compute_vert_imp_coefs_work(MPAS_VERTIMP_ARGS);
int oracle_set_smlstep_pert_variables_work(MPAS_SMLSTEP_ARGS);
int oracle_advance_acoustic_step_work(MPAS_ACOUSTIC_ARGS);
int oracle_divergence_damping_3d(MPAS_DIVDAMP_ARGS);
int oracle_recover_large_step_variables_work(MPAS_RECOVER_ARGS);
int oracle_compute_solve_diagnostics_work(MPAS_SOLVEDIAG_ARGS);
int oracle_init_coupled_diagnostics(MPAS_INITCOUPLED_ARGS);
int oracle_reconstruct_2d(MPAS_RECONSTRUCT_ARGS);
int oracle_rk_dynamics_substep_finish(MPAS_SUBSTEPFINISH_ARGS);
int oracle_compute_dyn_tend_work(MPAS_DYNTEND_ARGS);
int oracle_advance_scalars_work(MPAS_ADVSCALARS_ARGS);
int oracle_advance_scalars_mono_work(MPAS_MONO_ARGS);
}

namespace {

struct OField {
    void* ptr = nullptr;
    int inner = 1;  // product of all dims but the last
    int loc = 0;    // 1 cell, 2 edge, 3 vertex
};
struct XList {
    int endPoint, layer;
    std::vector<int> src, dest;
};
struct OBlock {
    std::map<std::string, OField> f;
    std::map<std::string, double> cfg;
    std::vector<XList> send[4], recv[4];  // by loc
    real* R(const std::string& n) {
        auto it = f.find(n);
        if (it == f.end()) throw std::runtime_error("oracle: missing field " + n);
        return (real*)it->second.ptr;
    }
    real* Ropt(const std::string& n) {
        auto it = f.find(n);
        return it == f.end() ? nullptr : (real*)it->second.ptr;
    }
    int* I(const std::string& n) {
        auto it = f.find(n);
        if (it == f.end()) throw std::runtime_error("oracle: missing field " + n);
        return (int*)it->second.ptr;
    }
    double c(const std::string& n) {
        auto it = cfg.find(n);
        if (it == cfg.end()) throw std::runtime_error("oracle: missing config " + n);
        return it->second;
    }
    double c(const std::string& n, double dflt) {
        auto it = cfg.find(n);
        return it == cfg.end() ? dflt : it->second;
    }
    void swap_levels(const std::string& n) { std::swap(f[n + "@1"].ptr, f[n + "@2"].ptr); }
};

struct Barrier {
    std::mutex m;
    std::condition_variable cv;
    int count = 0, gen = 0, n = 1;
    void wait() {
        std::unique_lock<std::mutex> lk(m);
        int g = gen;
        if (++count == n) { count = 0; ++gen; cv.notify_all(); }
        else cv.wait(lk, [&] { return g != gen; });
    }
};

struct OModel {
    std::vector<OBlock> blocks;
    std::list<std::vector<real>> owned;  // single-precision build: float copies of the double host pools
    Barrier bar;
    std::string err;
    long nExchanges = 0;
    // wall-clock per phase under the reference's own timer names (mpas_timer_start calls at TI:341...1369, src/framework/mpas_timer.F:88)
    std::map<std::string, std::pair<double, long>> timers;
    std::vector<std::string> timer_order;

    // exchange one field over the given halo layers (empty = all)
    void exchange_ptr(const std::vector<real*>& arrays, int loc, int inner, const std::vector<int>& layers) {
        ++nExchanges;
        const int P = (int)blocks.size();
        if (P == 1) return;
        auto want = [&](int l) {
            if (layers.empty()) return true;
            for (int q : layers)
                if (q == l) return true;
            return false;
        };
        std::vector<real> buffer;
        for (int b = 0; b < P; ++b)
            for (const XList& s : blocks[b].send[loc]) {
                if (!want(s.layer)) continue;
                OBlock& dstB = blocks[s.endPoint];
                const XList* r = nullptr;
                for (const XList& q : dstB.recv[loc])
                    if (q.endPoint == b && q.layer == s.layer) r = &q;
                if (!r) throw std::runtime_error("oracle: unmatched exchange list");
                buffer.assign((size_t)s.src.size() * inner, 0.0);
                for (size_t i = 0; i < s.src.size(); ++i)
                    std::memcpy(&buffer[(size_t)(s.dest[i] - 1) * inner], arrays[b] + (size_t)(s.src[i] - 1) * inner, sizeof(real) * inner);
                for (size_t i = 0; i < r->src.size(); ++i)
                    std::memcpy(arrays[s.endPoint] + (size_t)(r->dest[i] - 1) * inner, &buffer[(size_t)(r->src[i] - 1) * inner], sizeof(real) * inner);
            }
    }
    void exchange(const std::string& name, const std::vector<int>& layers = {}) {
        std::vector<real*> arrays;
        for (auto& b : blocks) arrays.push_back(b.R(name));
        const OField& f0 = blocks[0].f[name];
        exchange_ptr(arrays, f0.loc, f0.inner, layers);
    }
};

struct Ranges {
    int cS, cE, vS, vE, eS, eE, csS, csE, vsS, vsE, esS, esE;
};
Ranges ranges(OBlock& b) {
    Ranges r;
    r.cS = 1; r.cE = (int)b.c("nCells"); r.vS = 1; r.vE = (int)b.c("nVertices"); r.eS = 1; r.eE = (int)b.c("nEdges");
    r.csS = 1; r.csE = (int)b.c("nCellsSolve"); r.vsS = 1; r.vsE = (int)b.c("nVerticesSolve"); r.esS = 1; r.esE = (int)b.c("nEdgesSolve");
    return r;
}
#define RANGES r.cS, r.cE, r.vS, r.vE, r.eS, r.eE, r.csS, r.csE, r.vsS, r.vsE, r.esS, r.esE

struct Dims {
    int L, mE, mE2, ns, nC, nE, nV, nCS, nES, ms, me;
};
Dims dims(OBlock& b) {
    Dims d;
    d.L = (int)b.c("nVertLevels"); d.mE = (int)b.c("maxEdges"); d.mE2 = (int)b.c("maxEdges2"); d.ns = (int)b.c("num_scalars");
    d.nC = (int)b.c("nCells"); d.nE = (int)b.c("nEdges"); d.nV = (int)b.c("nVertices"); d.nCS = (int)b.c("nCellsSolve"); d.nES = (int)b.c("nEdgesSolve");
    d.ms = (int)b.c("moist_start", 1); d.me = (int)b.c("moist_end", 1);
    return d;
}

void call_solve_diagnostics(OBlock& b, int lev, real dt, int rk_step) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    std::string l = lev == 1 ? "@1" : "@2";
    oracle_compute_solve_diagnostics_work(d.L, d.mE, d.mE2, d.nC, d.nE, d.nV, 3, dt, b.c("config_apvm_upwinding", 0.5), b.R("fVertex"), b.R("fEdge"),
                                          b.R("invAreaTriangle"), b.R("invAreaCell"), b.R("dvEdge"), b.R("dcEdge"), b.R("invDvEdge"), b.R("invDcEdge"),
                                          b.R("weightsOnEdge"), b.R("kiteAreasOnVertex"), b.R("rho_edge"), b.R("rho_zz" + l), b.R("u" + l), b.R("v"),
                                          b.R("vorticity"), b.R("ke"), b.R("pv_edge"), b.R("pv_vertex"), b.R("pv_cell"), b.Ropt("gradPVn"), b.Ropt("gradPVt"),
                                          b.R("divergence"), b.I("cellsOnEdge"), b.I("cellsOnVertex"), b.I("verticesOnEdge"), b.I("edgesOnCell"),
                                          b.I("edgesOnEdge"), b.I("edgesOnVertex"), b.I("kiteForCell"), b.I("verticesOnCell"), b.R("edgesOnVertex_sign"),
                                          b.R("edgesOnCell_sign"), b.I("nEdgesOnCell"), b.I("nEdgesOnEdge"), r.cS, r.cE, r.vS, r.vE, r.eS, r.eE, rk_step,
                                          b.R("ke_vertex"), b.R("ke_edge"));
}

void call_vert_imp_coefs(OBlock& b, real dts) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    oracle_compute_vert_imp_coefs_work(d.L, d.ns, d.nC, d.ms, d.me, dts, b.c("config_epssm", 0.1), b.R("zz"), b.R("cqw"), b.R("exner"), b.R("theta_m@2"),
                                       b.R("rho_base"), b.R("rtheta_base"), b.R("exner_base"), b.R("rtheta_p"), b.R("cofwr"), b.R("cofwz"), b.R("coftz"),
                                       b.R("cofwt"), b.R("a_tri"), b.R("alpha_tri"), b.R("gamma_tri"), b.R("cofrz"), b.R("rdzw"), b.R("fzm"), b.R("fzp"),
                                       b.R("rdzu"), b.R("scalars@2"), r.cS, r.cE, r.eS, r.eE, r.csS, r.csE, r.esS, r.esE, b.R("qtot"));
}

void call_dyn_tend(OBlock& b, int rk_step, real dt) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    oracle_compute_dyn_tend_work(
        d.nC, d.nE, d.nV, d.L, d.nCS, d.nES, 3, d.mE, d.mE2, d.ns, d.ms, d.me, b.R("fEdge"), b.R("dvEdge"), b.R("dcEdge"), b.R("invDcEdge"), b.R("invDvEdge"),
        b.R("invAreaCell"), b.R("invAreaTriangle"), b.R("meshScalingDel2"), b.R("meshScalingDel4"), b.R("weightsOnEdge"), b.R("zgrid"), b.R("rho_edge"),
        b.R("rho_zz@2"), b.R("ru"), b.R("u@2"), b.R("v"), b.R("tend_u"), b.R("divergence"), b.R("vorticity"), b.R("ke"), b.R("pv_edge"), b.R("theta_m@2"),
        b.R("rw"), b.R("tend_rho"), b.Ropt("rt_diabatic_tend"), b.R("tend_theta"), b.R("tend_w"), b.R("w@2"), b.R("cqw"), b.R("rho_base"), b.R("rho_p"),
        b.R("pressure_p"), b.Ropt("pressure_base"), b.R("zz"), b.R("zxu"), b.R("cqu"), b.R("h_divergence"), b.R("kdiff"), b.R("edgesOnCell_sign"),
        b.R("edgesOnVertex_sign"), b.R("rw_save"), b.R("ru_save"), b.R("theta_m@1"), b.R("exner"), b.R("rho_p_save"), b.R("scalars@2"), b.R("tend_u_euler"),
        b.R("tend_w_euler"), b.R("tend_theta_euler"), b.Ropt("deriv_two"), b.I("cellsOnEdge"), b.I("verticesOnEdge"), b.I("edgesOnCell"), b.I("edgesOnEdge"),
        b.I("cellsOnCell"), b.I("edgesOnVertex"), b.I("nEdgesOnCell"), b.I("nEdgesOnEdge"), b.Ropt("latCell"), b.Ropt("latEdge"), b.Ropt("angleEdge"), b.Ropt("u_init"),
        b.Ropt("v_init"), b.I("advCellsForEdge"), b.I("nAdvCellsForEdge"), b.R("adv_coefs"), b.R("adv_coefs_3rd"), b.R("rdzu"), b.R("rdzw"), b.R("fzm"), b.R("fzp"),
        b.Ropt("qv_init"), b.Ropt("t_init"), b.c("cf1"), b.c("cf2"), b.c("cf3"), b.c("sphere_radius"), nullptr, nullptr, b.R("defc_a"), b.R("defc_b"), nullptr, nullptr,
        b.c("config_coef_3rd_order", 0.25), b.c("config_smagorinsky_coef", 0.125), (int)b.c("config_mix_full", 1), (int)b.c("config_horiz_mixing", 1),
        b.c("config_del4u_div_factor", 10.0), b.c("config_h_mom_eddy_visc2", 0.0), b.c("config_v_mom_eddy_visc2", 0.0), b.c("config_h_theta_eddy_visc2", 0.0),
        b.c("config_v_theta_eddy_visc2", 0.0), b.c("config_h_theta_eddy_visc4", 0.0), b.c("config_h_mom_eddy_visc4", 0.0), b.c("config_visc4_2dsmag", 0.05),
        b.c("config_len_disp"), rk_step, dt, b.c("config_mpas_cam_coef", 0.0), (int)b.c("config_rayleigh_damp_u", 0),
        b.c("config_rayleigh_damp_u_timescale_days", 5.0), (int)b.c("config_number_rayleigh_damp_u_levels", 6), b.Ropt("tend_rtheta_adv"), b.Ropt("rthdynten"),
        RANGES, b.R("qtot"), b.Ropt("tend_ru_physics"), b.Ropt("tend_rtheta_physics"), b.Ropt("tend_rho_physics"), b.R("delsq_theta"), b.R("delsq_w"),
        b.R("delsq_divergence"), b.R("delsq_u"), b.R("delsq_vorticity"), b.R("dpdz"));
}

void call_acoustic(OBlock& b, real dts, int small_step) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    oracle_advance_acoustic_step_work(d.L, d.mE, d.nC, d.nE, d.nCS, RANGES, b.R("rho_zz@2"), b.R("theta_m@1"), b.R("ru_p"), b.R("rw_p"), b.R("rtheta_pp"),
                                      b.R("rtheta_pp_old"), b.R("zz"), b.R("exner"), b.R("cqu"), b.R("ruAvg"), b.R("wwAvg"), b.R("rho_pp"), b.R("cofwt"),
                                      b.R("coftz"), b.R("zxu"), b.R("a_tri"), b.R("alpha_tri"), b.R("gamma_tri"), b.R("dss"), b.R("tend_u"), b.R("tend_rho"),
                                      b.R("tend_theta"), b.R("tend_w"), b.R("zgrid"), b.R("cofwr"), b.R("cofwz"), b.R("w@2"), b.R("ru"), b.R("ru_save"),
                                      b.R("rw"), b.R("rw_save"), b.R("fzm"), b.R("fzp"), b.R("rdzw"), b.R("dcEdge"), b.R("invDcEdge"), b.R("invAreaCell"),
                                      b.R("cofrz"), b.R("dvEdge"), b.I("nEdgesOnCell"), b.I("cellsOnEdge"), b.I("edgesOnCell"), b.R("edgesOnCell_sign"), dts,
                                      small_step, b.c("config_epssm", 0.1), b.c("cf1"), b.c("cf2"), b.c("cf3"), b.R("specZoneMaskEdge"), b.R("specZoneMaskCell"));
}

void call_recover(OBlock& b, real dt, int ns_, int rk_step) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    oracle_recover_large_step_variables_work(d.L, d.mE, d.ns, d.nC, d.nE, d.nCS, d.nES, dt, ns_, rk_step, b.R("wwAvg"), b.R("rw_save"), b.R("w@2"), b.R("rw"),
                                             b.R("rw_p"), b.R("rtheta_p"), b.R("rtheta_pp"), b.R("rtheta_p_save"), b.Ropt("rt_diabatic_tend"), b.R("rho_p"),
                                             b.R("rho_p_save"), b.R("rho_pp"), b.R("rho_zz@2"), b.R("rho_base"), b.R("ruAvg"), b.R("ru_save"), b.R("ru_p"),
                                             b.R("u@2"), b.R("ru"), b.R("exner"), b.R("exner_base"), b.R("rtheta_base"), b.R("pressure_p"), b.R("zz"),
                                             b.R("theta_m@2"), b.Ropt("pressure_base"), b.R("scalars@2"), b.R("fzm"), b.R("fzp"), nullptr, nullptr,
                                             b.R("zb_cell"), b.R("zb3_cell"), b.R("edgesOnCell_sign"), b.I("cellsOnEdge"), b.I("edgesOnCell"),
                                             b.I("nEdgesOnCell"), b.c("cf1"), b.c("cf2"), b.c("cf3"), b.I("bdyMaskCell"), RANGES);
}

void call_advance_scalars(OBlock& b, real dt, int rk_step, int order, int advance_density) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    oracle_advance_scalars_work(d.ns, d.nC, d.L, d.mE, dt, RANGES, b.c("config_coef_3rd_order", 0.25), b.R("scalars@1"), b.R("scalars@2"), b.R("rho_zz@1"),
                                b.R("rho_zz@2"), nullptr, b.R("ruAvg"), b.R("wwAvg"), nullptr, b.R("dvEdge"), b.I("cellsOnEdge"), b.I("edgesOnCell"),
                                b.I("nEdgesOnCell"), b.R("edgesOnCell_sign"), b.R("invAreaCell"), b.R("scalars_tend"), b.R("fzm"), b.R("fzp"), b.R("rdzw"),
                                nullptr, nullptr, b.I("bdyMaskCell"), b.I("bdyMaskEdge"), b.I("nAdvCellsForEdge"), b.I("advCellsForEdge"), b.R("adv_coefs"),
                                b.R("adv_coefs_3rd"), nullptr, nullptr, nullptr, d.nCS, d.nE, b.R("horiz_flux_arr"), rk_step, order, advance_density);
}

struct MonoCtx {
    OModel* m;
    int iblock;
};
int mono_halo(void* ctx, int which, real* /*array*/) {
    MonoCtx* c = (MonoCtx*)ctx;
    OModel* m = c->m;
    m->bar.wait();
    if (c->iblock == 0) {
        try {
            if (which == 0) m->exchange("scalars@1");
            else m->exchange("scale_arr", {1});
        } catch (const std::exception& e) { m->err = e.what(); }
    }
    m->bar.wait();
    return 0;
}

void call_mono(OModel& m, int ib, real dt, int advance_density) {
    OBlock& b = m.blocks[ib];
    Dims d = dims(b);
    Ranges r = ranges(b);
    MonoCtx ctx{&m, ib};
    oracle_advance_scalars_mono_work(m.blocks.size() > 1 ? mono_halo : nullptr, &ctx, d.nC, d.nE, d.L, d.mE, dt, RANGES, b.c("config_coef_3rd_order", 0.25),
                                     d.nCS, d.ns, b.R("ruAvg"), b.R("wwAvg"), b.R("scalars_tend"), b.R("rho_zz@1"), b.R("rho_zz@2"), b.R("scalars@1"),
                                     b.R("scalars@2"), b.R("invAreaCell"), b.R("dvEdge"), b.I("cellsOnEdge"), b.I("cellsOnCell"), b.I("edgesOnCell"),
                                     b.R("edgesOnCell_sign"), b.I("nEdgesOnCell"), b.R("fzm"), b.R("fzp"), b.R("rdzw"), b.I("nAdvCellsForEdge"),
                                     b.I("advCellsForEdge"), b.R("adv_coefs"), b.R("adv_coefs_3rd"), b.R("scalar_old"), b.R("scalar_new"), b.R("s_max"),
                                     b.R("s_min"), b.R("wdtn"), b.R("scale_arr"), b.R("flux_arr"), b.R("flux_upwind_tmp"), b.R("flux_tmp"),
                                     b.I("bdyMaskCell"), b.I("bdyMaskEdge"), advance_density, b.R("rho_zz_int"));
}

void run_mono_all(OModel& m, real dt, int advance_density) {
    const int P = (int)m.blocks.size();
    if (P == 1) { call_mono(m, 0, dt, advance_density); return; }
    m.bar.n = P;
    std::vector<std::thread> th;
    for (int ib = 0; ib < P; ++ib) th.emplace_back([&m, ib, dt, advance_density] { call_mono(m, ib, dt, advance_density); });
    for (auto& t : th) t.join();
    if (!m.err.empty()) throw std::runtime_error(m.err);
}

}  // namespace

namespace {

// atm_mpas_init_block + atm_core_init exchanges (mpas_atm_core.F:145, 163, 180-186, 390-402)
void model_init(OModel& m) {
    // The reference reads the initial state INCLUDING halo elements from the global init file ('input' stream,
    // mpas_atm_core.F:95).  The harness initialises each block on its own, so fields built with horizontal
    // stencils are wrong in the outer halo; fill them from their owners first (not counted as a step exchange).
    {
        long keep = m.nExchanges;
        for (const char* n : {"w@1", "zb_cell", "zb3_cell", "zxu", "defc_a", "defc_b"})
            if (m.blocks[0].f.count(n)) m.exchange(n);
        m.nExchanges = keep;
    }
    m.exchange("u@1");  // :145
    for (auto& b : m.blocks) {
        Dims d = dims(b);
        Ranges r = ranges(b);
        if (b.c("config_do_restart", 0) == 0)  // mpas_atm_core.F:389: a restart file already holds the coupled diagnostics
        oracle_init_coupled_diagnostics(d.L, d.mE, d.ns, (int)b.c("index_qv", 1), d.nC, d.nE, RANGES, b.I("cellsOnEdge"), b.I("nEdgesOnCell"), b.I("edgesOnCell"),
                                        b.R("edgesOnCell_sign"), b.R("theta_m@1"), b.R("theta"), b.R("rho_zz@1"), b.R("rho"), b.R("rho_p"), b.R("rho_base"),
                                        b.R("rtheta_base"), b.R("theta_base"), b.R("rtheta_p"), b.R("zz"), b.R("scalars@1"), b.R("ru"), b.R("rw"), b.R("u@1"),
                                        b.R("w@1"), b.R("pressure_p"), b.R("pressure_base"), b.R("exner"), b.R("exner_base"), b.R("fzm"), b.R("fzp"),
                                        b.R("zb_cell"), b.R("zb3_cell"));
        call_solve_diagnostics(b, 1, (real)b.c("config_dt"), -1);
    }
    m.exchange("pv_edge");
    m.exchange("ru");
    m.exchange("rw");
}

struct OPhase {  // mpas_timer_start / mpas_timer_stop of the reference, same names
    OModel& m;
    const char* name;
    std::chrono::steady_clock::time_point t0;
    OPhase(OModel& m_, const char* n) : m(m_), name(n), t0(std::chrono::steady_clock::now()) {}
    ~OPhase() {
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        auto it = m.timers.find(name);
        if (it == m.timers.end()) {
            m.timer_order.push_back(name);
            it = m.timers.emplace(name, std::make_pair(0.0, 0L)).first;
        }
        it->second.first += ms;
        it->second.second += 1;
    }
};

// atm_srk3, TI:142-1796 (config_apply_lbcs = false, physics off, IAU off)
// mpas_reconstruct at the end of atm_srk3 (TI:1594): u time level 2, owned cells
void call_reconstruct(OBlock& b) {
    if (!b.f.count("coeffs_reconstruct") || !b.f.count("uReconstructX")) return;
    Dims d = dims(b);
    oracle_reconstruct_2d(d.L, d.mE, d.nC, d.nE, d.nCS, (int)b.c("on_a_sphere", 1), b.R("u@2"), b.R("coeffs_reconstruct"), b.I("nEdgesOnCell"), b.I("edgesOnCell"),
                          b.R("latCell"), b.R("lonCell"), b.R("uReconstructX"), b.R("uReconstructY"), b.R("uReconstructZ"), b.R("uReconstructZonal"),
                          b.R("uReconstructMeridional"));
}

void srk3_step(OModel& m) {
    OBlock& b0 = m.blocks[0];
    const real dt = (real)b0.c("config_dt");
    const int order = (int)b0.c("config_time_integration_order", 2);
    const int number_of_sub_steps = (int)b0.c("config_number_of_sub_steps", 2);
    const bool split = b0.c("config_split_dynamics_transport", 1) != 0;
    const bool scalar_advection = b0.c("config_scalar_advection", 1) != 0;
    const bool monotonic = b0.c("config_monotonic", 1) != 0;
    const bool positive_definite = b0.c("config_positive_definite", 0) != 0;
    int dynamics_split = (int)b0.c("config_dynamics_split_steps", 3);
    real dt_dynamics;
    if (split) dt_dynamics = dt / (real)dynamics_split;
    else { dynamics_split = 1; dt_dynamics = dt; }
    real rk_timestep[4], rk_sub_timestep[4];
    int number_sub_steps[4];
    if (order == 3) {
        rk_timestep[1] = dt_dynamics / 3.; rk_timestep[2] = dt_dynamics / 2.; rk_timestep[3] = dt_dynamics;
        rk_sub_timestep[1] = dt_dynamics / 3.; rk_sub_timestep[2] = dt_dynamics / (real)number_of_sub_steps; rk_sub_timestep[3] = dt_dynamics / (real)number_of_sub_steps;
        number_sub_steps[1] = 1; number_sub_steps[2] = std::max(1, number_of_sub_steps / 2); number_sub_steps[3] = number_of_sub_steps;
    } else {
        rk_timestep[1] = dt_dynamics / 2.; rk_timestep[2] = dt_dynamics / 2.; rk_timestep[3] = dt_dynamics;
        rk_sub_timestep[1] = rk_sub_timestep[2] = rk_sub_timestep[3] = dt_dynamics / (real)number_of_sub_steps;
        number_sub_steps[1] = std::max(1, number_of_sub_steps / 2); number_sub_steps[2] = std::max(1, number_of_sub_steps / 2); number_sub_steps[3] = number_of_sub_steps;
    }
    m.exchange("theta_m@1");  // TI:329-338
    m.exchange("scalars@1");
    m.exchange("pressure_p");
    m.exchange("rtheta_p");
    for (auto& b : m.blocks) {  // TI:369
        OPhase ph(m, "atm_rk_integration_setup");
        Dims d = dims(b);
        Ranges r = ranges(b);
        oracle_rk_integration_setup(d.L, d.ns, d.nC, d.nE, RANGES, b.R("ru"), b.R("ru_save"), b.R("rw"), b.R("rw_save"), b.R("rtheta_p"), b.R("rtheta_p_save"),
                                    b.R("rho_p"), b.R("rho_p_save"), b.R("rho_zz_old_split"), b.R("u@1"), b.R("u@2"), b.R("w@1"), b.R("w@2"), b.R("theta_m@1"),
                                    b.R("theta_m@2"), b.R("rho_zz@1"), b.R("rho_zz@2"), b.R("scalars@1"), b.R("scalars@2"));
    }
    for (auto& b : m.blocks) {  // TI:410
        OPhase ph(m, "atm_compute_moist_coefficients");
        Dims d = dims(b);
        Ranges r = ranges(b);
        oracle_compute_moist_coefficients(d.L, d.ns, d.nC, d.nE, d.nCS, d.ms, d.me, RANGES, b.I("cellsOnEdge"), b.R("scalars@2"), b.R("cqw"), b.R("cqu"), b.R("qtot"));
    }
    for (int dynamics_substep = 1; dynamics_substep <= dynamics_split; ++dynamics_substep) {
        {
            OPhase ph(m, "atm_compute_vert_imp_coefs");
            for (auto& b : m.blocks) call_vert_imp_coefs(b, rk_sub_timestep[1]);  // TI:501
        }
        m.exchange("exner");                                                  // TI:513
        for (int rk_step = 1; rk_step <= 3; ++rk_step) {
            if (order == 3 && rk_step == 2) {
                OPhase ph(m, "atm_compute_vert_imp_coefs");
                for (auto& b : m.blocks) call_vert_imp_coefs(b, rk_sub_timestep[rk_step]);  // TI:550
            }
            {
                OPhase ph(m, "atm_compute_dyn_tend");
                for (auto& b : m.blocks) call_dyn_tend(b, rk_step, dt);  // TI:609
            }
            m.exchange("tend_u", {1});                                                       // TI:642
            for (auto& b : m.blocks) {                                                       // TI:669
                OPhase ph(m, "small_step_prep");
                Dims d = dims(b);
                Ranges r = ranges(b);
                oracle_set_smlstep_pert_variables_work(d.L, d.mE, d.nC, d.nE, d.nCS, b.I("nEdgesOnCell"), b.I("cellsOnEdge"), b.I("edgesOnCell"), b.R("fzm"),
                                                       b.R("fzp"), b.R("zb_cell"), b.R("zb3_cell"), b.R("zz"), b.R("tend_w"), b.R("tend_u"), b.I("bdyMaskCell"),
                                                       b.R("edgesOnCell_sign"), r.cS, r.cE, r.eS, r.eE, r.csS, r.csE, r.esS, r.esE);
            }
            for (int small_step = 1; small_step <= number_sub_steps[rk_step]; ++small_step) {
                m.exchange("rho_pp", {1});  // TI:792
                {
                    OPhase ph(m, "atm_advance_acoustic_step");
                    for (auto& b : m.blocks) call_acoustic(b, rk_sub_timestep[rk_step], small_step);  // TI:824
                }
                m.exchange("rtheta_pp", {1});  // TI:845
                for (auto& b : m.blocks) {     // TI:862
                    OPhase ph(m, "atm_divergence_damping_3d");
                    Dims d = dims(b);
                    oracle_divergence_damping_3d(d.L, d.nC, d.nE, d.nCS, 1, d.nE, b.R("theta_m@1"), b.R("ru_p"), b.R("rtheta_pp"), b.R("rtheta_pp_old"),
                                                 b.I("cellsOnEdge"), b.R("specZoneMaskEdge"), rk_sub_timestep[rk_step], b.c("config_smdiv", 0.1), b.c("config_len_disp"));
                }
            }
            m.exchange("rw_p");  // TI:876-887
            m.exchange("ru_p");
            m.exchange("rho_pp");
            m.exchange("rtheta_pp", {2});
            {
                OPhase ph(m, "atm_recover_large_step_variables");
                for (auto& b : m.blocks) call_recover(b, rk_timestep[rk_step], number_sub_steps[rk_step], rk_step);  // TI:917
            }
            m.exchange("u@2");                                                                                   // TI:988
            if (scalar_advection && !split) {                                                                    // TI:993-1099
                if (rk_step < 3 || (!monotonic && !positive_definite)) {
                    OPhase ph(m, "atm_advance_scalars");
                    for (auto& b : m.blocks) call_advance_scalars(b, rk_timestep[rk_step], rk_step, order, 0);
                } else {
                    OPhase ph(m, "atm_advance_scalars_mono");
                    run_mono_all(m, rk_timestep[rk_step], 0);
                }
            }
            {
                OPhase ph(m, "atm_compute_solve_diagnostics");
                for (auto& b : m.blocks) call_solve_diagnostics(b, 2, dt, rk_step);  // TI:1216
            }
            m.exchange("w@2");                                                   // TI:1234-1249
            m.exchange("pv_edge");
            m.exchange("rho_edge");
            if (scalar_advection && !split) m.exchange("scalars@2");
        }
        if (dynamics_substep < dynamics_split) {  // TI:1282-1288
            m.exchange("theta_m@2");
            m.exchange("pressure_p");
            m.exchange("rtheta_p");
        }
        for (auto& b : m.blocks) {  // TI:1329
            OPhase ph(m, "atm_rk_dynamics_substep_finish");
            Dims d = dims(b);
            Ranges r = ranges(b);
            oracle_rk_dynamics_substep_finish(d.L, d.nC, d.nE, dynamics_substep, dynamics_split, RANGES, b.R("ru"), b.R("ru_save"), b.R("rw"), b.R("rw_save"),
                                              b.R("rtheta_p"), b.R("rtheta_p_save"), b.R("rho_p"), b.R("rho_p_save"), b.R("rho_zz_old_split"), b.R("ruAvg"),
                                              b.R("ruAvg_split"), b.R("wwAvg"), b.R("wwAvg_split"), b.R("u@1"), b.R("u@2"), b.R("w@1"), b.R("w@2"),
                                              b.R("theta_m@1"), b.R("theta_m@2"), b.R("rho_zz@1"), b.R("rho_zz@2"));
        }
    }
    if (scalar_advection && split) {  // TI:1355-1576
        real rk_t[4] = {0, dt / (real)3., dt / (real)2., dt};
        if (order == 2) rk_t[1] = dt / 2.;
        for (int rk_step = 1; rk_step <= 3; ++rk_step) {
            if (rk_step < 3 || (!monotonic && !positive_definite)) {
                OPhase ph(m, "atm_advance_scalars");
                for (auto& b : m.blocks) call_advance_scalars(b, rk_t[rk_step], rk_step, order, 1);
            } else {
                OPhase ph(m, "atm_advance_scalars_mono");
                run_mono_all(m, rk_t[rk_step], 1);
            }
            if (rk_step < 3) m.exchange("scalars@2");
        }
    }
    // DO_PHYSICS (default build, src/core_atmosphere/Makefile:3): negative mixing ratios reset, TI:1646-1647
    for (auto& b : m.blocks) {
        Dims d = dims(b);
        real* s2 = b.R("scalars@2");
        size_t n = (size_t)d.ns * d.L * (d.nC + 1);
        for (size_t i = 0; i < n; ++i)
            if (s2[i] < 0.0) s2[i] = 0.0;
    }
    {
        OPhase ph(m, "mpas_reconstruct");
        for (auto& b : m.blocks) call_reconstruct(b);  // TI:1594
    }
    // mpas_pool_shift_time_levels(state), mpas_atm_core.F:671
    for (auto& b : m.blocks)
        for (const char* n : {"u", "w", "theta_m", "rho_zz", "scalars"}) b.swap_levels(n);
}

}  // namespace

extern "C" {

static std::string g_oracle_err;
const char* oracle_last_error() { return g_oracle_err.c_str(); }

void* oracle_model_create(int nBlocks) {
    OModel* m = new OModel();
    m->blocks.resize(nBlocks);
    return m;
}
void oracle_model_destroy(void* h) { delete (OModel*)h; }
int oracle_model_set_field(void* h, int ib, const char* name, void* ptr, int inner, int loc) {
    OModel* m = (OModel*)h;
    OField f;
    f.ptr = ptr; f.inner = inner; f.loc = loc;
    m->blocks[ib].f[name] = f;
    return 0;
}
// sizeof(mpas_real) of this build (liboracle_sp.so = -DSINGLE_PRECISION, the reference's PRECISION=single)
int oracle_real_size() { return (int)sizeof(real); }
// A REAL field of the (double) host pool.  The double build works in place on `ptr`; the single-precision build keeps
// its own float copy (owned by the model, handed out by oracle_model_get_field).
int oracle_model_set_field_f64(void* h, int ib, const char* name, double* ptr, long long count, int inner, int loc) {
    OModel* m = (OModel*)h;
    OField f;
    f.inner = inner; f.loc = loc;
    if (sizeof(real) == sizeof(double)) {
        f.ptr = ptr;
    } else {
        m->owned.emplace_back((size_t)count);
        std::vector<real>& v = m->owned.back();
        for (long long i = 0; i < count; ++i) v[(size_t)i] = (real)ptr[i];
        f.ptr = v.data();
    }
    m->blocks[ib].f[name] = f;
    return 0;
}
void* oracle_model_get_field(void* h, int ib, const char* name) {
    OModel* m = (OModel*)h;
    auto it = m->blocks[ib].f.find(name);
    return it == m->blocks[ib].f.end() ? nullptr : it->second.ptr;
}
int oracle_model_set_cfg(void* h, int ib, const char* name, double v) {
    ((OModel*)h)->blocks[ib].cfg[name] = v;
    return 0;
}
int oracle_model_add_exchange(void* h, int ib, int loc, int which, int endPoint, int layer, int n, const int* src, const int* dest) {
    OModel* m = (OModel*)h;
    XList x;
    x.endPoint = endPoint; x.layer = layer;
    x.src.assign(src, src + n);
    x.dest.assign(dest, dest + n);
    (which == 0 ? m->blocks[ib].send[loc] : m->blocks[ib].recv[loc]).push_back(x);
    return 0;
}
int oracle_model_exchange(void* h, const char* name, int layer /*0 = all*/) {
    try {
        OModel* m = (OModel*)h;
        if (layer == 0) m->exchange(name);
        else m->exchange(name, {layer});
        return 0;
    } catch (const std::exception& e) { g_oracle_err = e.what(); return 1; }
}
int oracle_model_init(void* h) {
    try { model_init(*(OModel*)h); return 0; } catch (const std::exception& e) { g_oracle_err = e.what(); return 1; }
}
int oracle_model_step(void* h, int nsteps) {
    try {
        for (int i = 0; i < nsteps; ++i) srk3_step(*(OModel*)h);
        return 0;
    } catch (const std::exception& e) { g_oracle_err = e.what(); return 1; }
}
// run one work routine on the current fields of every block (routine-level parity tests); parameters as in
// mpas-model_b200/csrc/cuda/srk3.cu dycore_call
int oracle_model_call(void* h, const char* routine, const double* q, int nq) {
    try {
        OModel& m = *(OModel*)h;
        std::string rt = routine;
        auto need = [&](int n) { if (nq < n) throw std::runtime_error("oracle_model_call: too few parameters for " + rt); };
        if (rt == "mono") { need(2); run_mono_all(m, (real)q[0], (int)q[1]); return 0; }
        for (auto& b : m.blocks) {
            Dims d = dims(b);
            Ranges r = ranges(b);
            if (rt == "acoustic") { need(2); call_acoustic(b, (real)q[0], (int)q[1]); }
            else if (rt == "divdamp") {
                need(1);
                oracle_divergence_damping_3d(d.L, d.nC, d.nE, d.nCS, 1, d.nE, b.R("theta_m@1"), b.R("ru_p"), b.R("rtheta_pp"), b.R("rtheta_pp_old"),
                                             b.I("cellsOnEdge"), b.R("specZoneMaskEdge"), (real)q[0], b.c("config_smdiv", 0.1), b.c("config_len_disp"));
            }
            else if (rt == "recover") { need(3); call_recover(b, (real)q[0], (int)q[1], (int)q[2]); }
            else if (rt == "dyn_tend") { need(2); call_dyn_tend(b, (int)q[0], (real)q[1]); }
            else if (rt == "advance_scalars") { need(4); call_advance_scalars(b, (real)q[0], (int)q[1], (int)q[2], (int)q[3]); }
            else if (rt == "solve_diagnostics") { need(3); call_solve_diagnostics(b, (int)q[0], (real)q[1], (int)q[2]); }
            else if (rt == "vert_imp_coefs") { need(1); call_vert_imp_coefs(b, (real)q[0]); }
            else if (rt == "smlstep") {
                oracle_set_smlstep_pert_variables_work(d.L, d.mE, d.nC, d.nE, d.nCS, b.I("nEdgesOnCell"), b.I("cellsOnEdge"), b.I("edgesOnCell"), b.R("fzm"),
                                                       b.R("fzp"), b.R("zb_cell"), b.R("zb3_cell"), b.R("zz"), b.R("tend_w"), b.R("tend_u"), b.I("bdyMaskCell"),
                                                       b.R("edgesOnCell_sign"), r.cS, r.cE, r.eS, r.eE, r.csS, r.csE, r.esS, r.esE);
            }
            else if (rt == "rk_setup") {
                oracle_rk_integration_setup(d.L, d.ns, d.nC, d.nE, RANGES, b.R("ru"), b.R("ru_save"), b.R("rw"), b.R("rw_save"), b.R("rtheta_p"), b.R("rtheta_p_save"),
                                            b.R("rho_p"), b.R("rho_p_save"), b.R("rho_zz_old_split"), b.R("u@1"), b.R("u@2"), b.R("w@1"), b.R("w@2"), b.R("theta_m@1"),
                                            b.R("theta_m@2"), b.R("rho_zz@1"), b.R("rho_zz@2"), b.R("scalars@1"), b.R("scalars@2"));
            }
            else if (rt == "moist") {
                oracle_compute_moist_coefficients(d.L, d.ns, d.nC, d.nE, d.nCS, d.ms, d.me, RANGES, b.I("cellsOnEdge"), b.R("scalars@2"), b.R("cqw"), b.R("cqu"), b.R("qtot"));
            }
            else if (rt == "substep_finish") {
                need(2);
                oracle_rk_dynamics_substep_finish(d.L, d.nC, d.nE, (int)q[0], (int)q[1], RANGES, b.R("ru"), b.R("ru_save"), b.R("rw"), b.R("rw_save"),
                                                  b.R("rtheta_p"), b.R("rtheta_p_save"), b.R("rho_p"), b.R("rho_p_save"), b.R("rho_zz_old_split"), b.R("ruAvg"),
                                                  b.R("ruAvg_split"), b.R("wwAvg"), b.R("wwAvg_split"), b.R("u@1"), b.R("u@2"), b.R("w@1"), b.R("w@2"),
                                                  b.R("theta_m@1"), b.R("theta_m@2"), b.R("rho_zz@1"), b.R("rho_zz@2"));
            }
            else if (rt == "reconstruct") call_reconstruct(b);
            else if (rt == "output_diagnostics") {  // time level 1 after a step, as atm_core_run calls it (mpas_atm_core.F:717)
                oracle_compute_output_diagnostics(d.L, d.ns, (int)b.c("index_qv", 1), d.nC, b.R("theta_m@1"), b.R("rho_zz@1"), b.R("scalars@1"), b.R("zz"),
                                                  b.R("pressure_base"), b.R("pressure_p"), b.R("theta"), b.R("rho"), b.R("pressure"));
            }
            else if (rt == "init_coupled") {
                oracle_init_coupled_diagnostics(d.L, d.mE, d.ns, (int)b.c("index_qv", 1), d.nC, d.nE, RANGES, b.I("cellsOnEdge"), b.I("nEdgesOnCell"), b.I("edgesOnCell"),
                                                b.R("edgesOnCell_sign"), b.R("theta_m@1"), b.R("theta"), b.R("rho_zz@1"), b.R("rho"), b.R("rho_p"), b.R("rho_base"),
                                                b.R("rtheta_base"), b.R("theta_base"), b.R("rtheta_p"), b.R("zz"), b.R("scalars@1"), b.R("ru"), b.R("rw"), b.R("u@1"),
                                                b.R("w@1"), b.R("pressure_p"), b.R("pressure_base"), b.R("exner"), b.R("exner_base"), b.R("fzm"), b.R("fzp"),
                                                b.R("zb_cell"), b.R("zb3_cell"));
            }
            else throw std::runtime_error("oracle_model_call: unknown routine " + rt);
        }
        return 0;
    } catch (const std::exception& e) { g_oracle_err = e.what(); return 1; }
}
long oracle_model_exchange_count(void* h) { return ((OModel*)h)->nExchanges; }

// per-phase wall-clock of the steps taken so far, reference timer names
int oracle_model_timer_count(void* h) { return (int)((OModel*)h)->timer_order.size(); }
int oracle_model_timer_get(void* h, int idx, char* name, int cap, double* ms, long* calls) {
    OModel* m = (OModel*)h;
    if (idx < 0 || idx >= (int)m->timer_order.size()) return 1;
    const std::string& n = m->timer_order[idx];
    std::strncpy(name, n.c_str(), cap - 1);
    name[cap - 1] = 0;
    *ms = m->timers[n].first;
    *calls = m->timers[n].second;
    return 0;
}
void oracle_model_timer_reset(void* h) {
    for (auto& kv : ((OModel*)h)->timers) kv.second = {0.0, 0L};
}
// OpenMP threads of every later call (torch.distributed.run exports OMP_NUM_THREADS=1 to its workers)
int oracle_set_num_threads(int n) {
    if (n < 1) n = omp_get_num_procs();
    omp_set_num_threads(n);
    return n;
}
int oracle_get_max_threads() { return omp_get_max_threads(); }

}  // extern "C"
