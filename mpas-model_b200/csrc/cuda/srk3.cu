// The atm_srk3 call sequence (TI = src/core_atmosphere/dynamics/mpas_atm_time_integration.F:142-1796),
// start-up (src/core_atmosphere/mpas_atm_core.F:145-186, 311-402) and the time-level shift
// (mpas_atm_core.F:671) over device-resident fields.  Every compute call goes through the same
// extern "C" mpas_b200_*_work entry points a Fortran host would bind (include/mpas_b200.h), passing
// device pointers; pool keys follow SURVEY.md Appendix D.  config_apply_lbcs = false, physics off.
#include <stdexcept>
#include <thread>

#include "../../../include/mpas_b200.h"
#include "dycore.h"

namespace mpasb200 {

namespace {

struct Ranges {
    int cS, cE, vS, vE, eS, eE, csS, csE, vsS, vsE, esS, esE;
};
Ranges ranges(const DBlock& b) {
    Ranges r;
    r.cS = 1; r.cE = (int)b.c("nCells"); r.vS = 1; r.vE = (int)b.c("nVertices"); r.eS = 1; r.eE = (int)b.c("nEdges");
    r.csS = 1; r.csE = (int)b.c("nCellsSolve"); r.vsS = 1; r.vsE = (int)b.c("nVerticesSolve"); r.esS = 1; r.esE = (int)b.c("nEdgesSolve");
    return r;
}
#define RANGES r.cS, r.cE, r.vS, r.vE, r.eS, r.eE, r.csS, r.csE, r.vsS, r.vsE, r.esS, r.esE

struct Dims {
    int L, mE, mE2, ns, nC, nE, nV, nCS, nES, ms, me;
};
Dims dims(const DBlock& b) {
    Dims d;
    d.L = (int)b.c("nVertLevels"); d.mE = (int)b.c("maxEdges"); d.mE2 = (int)b.c("maxEdges2"); d.ns = (int)b.c("num_scalars");
    d.nC = (int)b.c("nCells"); d.nE = (int)b.c("nEdges"); d.nV = (int)b.c("nVertices"); d.nCS = (int)b.c("nCellsSolve"); d.nES = (int)b.c("nEdgesSolve");
    d.ms = (int)b.c("moist_start", 1); d.me = (int)b.c("moist_end", 1);
    return d;
}

// ---- phase timers (names follow the reference's mpas_timer_start calls) ------------
cudaEvent_t get_event(Dycore& m) {
    if (!m.event_pool.empty()) {
        cudaEvent_t e = m.event_pool.back();
        m.event_pool.pop_back();
        return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}
struct Phase {
    Dycore& m;
    PhaseTimer* t = nullptr;
    cudaEvent_t e0 = nullptr;
    Phase(Dycore& m_, const char* name) : m(m_) {
        if (!m.timing) return;
        auto it = m.timers.find(name);
        if (it == m.timers.end()) {
            m.timer_order.push_back(name);
            it = m.timers.emplace(name, PhaseTimer()).first;
        }
        t = &it->second;
        e0 = get_event(m);
        cudaEventRecord(e0, stream());
    }
    ~Phase() {
        if (!t) return;
        cudaEvent_t e1 = get_event(m);
        cudaEventRecord(e1, stream());
        t->pending.push_back({e0, e1});
        t->calls++;
    }
};
void flush_timers(Dycore& m) {
    if (!m.timing) return;
    cudaStreamSynchronize(stream());
    for (auto& kv : m.timers) {
        for (auto& pr : kv.second.pending) {
            float ms = 0;
            cudaEventElapsedTime(&ms, pr.first, pr.second);
            kv.second.ms += ms;
            m.event_pool.push_back(pr.first);
            m.event_pool.push_back(pr.second);
        }
        kv.second.pending.clear();
    }
}

#define CK(call)             \
    do {                     \
        int _rc = (call);    \
        if (_rc) return _rc; \
    } while (0)

int call_solve_diagnostics(DBlock& b, int lev, real dt, int rk_step) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    std::string l = lev == 1 ? "@1" : "@2";
    return mpas_b200_compute_solve_diagnostics_work(
        d.L, d.mE, d.mE2, d.nC, d.nE, d.nV, 3, dt, (real)b.c("config_apvm_upwinding", 0.5), b.R("fVertex"), b.R("fEdge"), b.R("invAreaTriangle"),
        b.R("invAreaCell"), b.R("dvEdge"), b.R("dcEdge"), b.R("invDvEdge"), b.R("invDcEdge"), b.R("weightsOnEdge"), b.R("kiteAreasOnVertex"), b.R("rho_edge"),
        b.R("rho_zz" + l), b.R("u" + l), b.R("v"), b.R("vorticity"), b.R("ke"), b.R("pv_edge"), b.R("pv_vertex"), b.R("pv_cell"), b.Ropt("gradPVn"),
        b.Ropt("gradPVt"), b.R("divergence"), b.I("cellsOnEdge"), b.I("cellsOnVertex"), b.I("verticesOnEdge"), b.I("edgesOnCell"), b.I("edgesOnEdge"),
        b.I("edgesOnVertex"), b.I("kiteForCell"), b.I("verticesOnCell"), b.R("edgesOnVertex_sign"), b.R("edgesOnCell_sign"), b.I("nEdgesOnCell"),
        b.I("nEdgesOnEdge"), r.cS, r.cE, r.vS, r.vE, r.eS, r.eE, rk_step, b.R("ke_vertex"), b.Ropt("ke_edge"));
}

int call_vert_imp_coefs(DBlock& b, real dts) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    return mpas_b200_compute_vert_imp_coefs_work(d.L, d.ns, d.nC, d.ms, d.me, dts, (real)b.c("config_epssm", 0.1), b.R("zz"), b.R("cqw"), b.R("exner"),
                                                 b.R("theta_m@2"), b.R("rho_base"), b.R("rtheta_base"), b.R("exner_base"), b.R("rtheta_p"), b.R("cofwr"),
                                                 b.R("cofwz"), b.R("coftz"), b.R("cofwt"), b.R("a_tri"), b.R("alpha_tri"), b.R("gamma_tri"), b.R("cofrz"),
                                                 b.R("rdzw"), b.R("fzm"), b.R("fzp"), b.R("rdzu"), b.R("scalars@2"), r.cS, r.cE, r.eS, r.eE, r.csS, r.csE, r.esS,
                                                 r.esE, b.R("qtot"));
}

int call_dyn_tend(DBlock& b, int rk_step, real dt) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    return mpas_b200_compute_dyn_tend_work(
        d.nC, d.nE, d.nV, d.L, d.nCS, d.nES, 3, d.mE, d.mE2, d.ns, d.ms, d.me, b.R("fEdge"), b.R("dvEdge"), b.R("dcEdge"), b.R("invDcEdge"), b.R("invDvEdge"),
        b.R("invAreaCell"), b.R("invAreaTriangle"), b.R("meshScalingDel2"), b.R("meshScalingDel4"), b.R("weightsOnEdge"), b.R("zgrid"), b.R("rho_edge"),
        b.R("rho_zz@2"), b.R("ru"), b.R("u@2"), b.R("v"), b.R("tend_u"), b.R("divergence"), b.R("vorticity"), b.R("ke"), b.R("pv_edge"), b.R("theta_m@2"),
        b.R("rw"), b.R("tend_rho"), b.Ropt("rt_diabatic_tend"), b.R("tend_theta"), b.R("tend_w"), b.R("w@2"), b.R("cqw"), b.R("rho_base"), b.R("rho_p"),
        b.R("pressure_p"), b.Ropt("pressure_base"), b.R("zz"), b.R("zxu"), b.R("cqu"), b.R("h_divergence"), b.R("kdiff"), b.R("edgesOnCell_sign"),
        b.R("edgesOnVertex_sign"), b.R("rw_save"), b.R("ru_save"), b.R("theta_m@1"), b.R("exner"), b.R("rho_p_save"), b.R("scalars@2"), b.R("tend_u_euler"),
        b.R("tend_w_euler"), b.R("tend_theta_euler"), b.Ropt("deriv_two"), b.I("cellsOnEdge"), b.I("verticesOnEdge"), b.I("edgesOnCell"), b.I("edgesOnEdge"),
        b.I("cellsOnCell"), b.I("edgesOnVertex"), b.I("nEdgesOnCell"), b.I("nEdgesOnEdge"), b.Ropt("latCell"), b.Ropt("latEdge"), b.Ropt("angleEdge"), b.Ropt("u_init"),
        b.Ropt("v_init"), b.I("advCellsForEdge"), b.I("nAdvCellsForEdge"), b.R("adv_coefs"), b.R("adv_coefs_3rd"), b.R("rdzu"), b.R("rdzw"), b.R("fzm"), b.R("fzp"),
        b.Ropt("qv_init"), b.Ropt("t_init"), (real)b.c("cf1"), (real)b.c("cf2"), (real)b.c("cf3"), (real)b.c("sphere_radius"), nullptr, nullptr, b.R("defc_a"), b.R("defc_b"),
        nullptr, nullptr, (real)b.c("config_coef_3rd_order", 0.25), (real)b.c("config_smagorinsky_coef", 0.125), (int)b.c("config_mix_full", 1),
        (int)b.c("config_horiz_mixing", 1), (real)b.c("config_del4u_div_factor", 10.0), (real)b.c("config_h_mom_eddy_visc2", 0.0),
        (real)b.c("config_v_mom_eddy_visc2", 0.0), (real)b.c("config_h_theta_eddy_visc2", 0.0), (real)b.c("config_v_theta_eddy_visc2", 0.0),
        (real)b.c("config_h_theta_eddy_visc4", 0.0), (real)b.c("config_h_mom_eddy_visc4", 0.0), (real)b.c("config_visc4_2dsmag", 0.05),
        (real)b.c("config_len_disp"), rk_step, dt, (real)b.c("config_mpas_cam_coef", 0.0), (int)b.c("config_rayleigh_damp_u", 0),
        (real)b.c("config_rayleigh_damp_u_timescale_days", 5.0), (int)b.c("config_number_rayleigh_damp_u_levels", 6), b.Ropt("tend_rtheta_adv"),
        b.Ropt("rthdynten"), RANGES, b.R("qtot"), b.Ropt("tend_ru_physics"), b.Ropt("tend_rtheta_physics"), b.Ropt("tend_rho_physics"), b.R("delsq_theta"),
        b.R("delsq_w"), b.R("delsq_divergence"), b.R("delsq_u"), b.R("delsq_vorticity"), b.R("dpdz"));
}

int call_smlstep(DBlock& b) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    return mpas_b200_set_smlstep_pert_variables_work(d.L, d.mE, d.nC, d.nE, d.nCS, b.I("nEdgesOnCell"), b.I("cellsOnEdge"), b.I("edgesOnCell"), b.R("fzm"),
                                                     b.R("fzp"), b.R("zb_cell"), b.R("zb3_cell"), b.R("zz"), b.R("tend_w"), b.R("tend_u"), b.I("bdyMaskCell"),
                                                     b.R("edgesOnCell_sign"), r.cS, r.cE, r.eS, r.eE, r.csS, r.csE, r.esS, r.esE);
}

int call_acoustic(DBlock& b, real dts, int small_step) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    return mpas_b200_advance_acoustic_step_work(
        d.L, d.mE, d.nC, d.nE, d.nCS, RANGES, b.R("rho_zz@2"), b.R("theta_m@1"), b.R("ru_p"), b.R("rw_p"), b.R("rtheta_pp"), b.R("rtheta_pp_old"), b.R("zz"),
        b.R("exner"), b.R("cqu"), b.R("ruAvg"), b.R("wwAvg"), b.R("rho_pp"), b.R("cofwt"), b.R("coftz"), b.R("zxu"), b.R("a_tri"), b.R("alpha_tri"),
        b.R("gamma_tri"), b.R("dss"), b.R("tend_u"), b.R("tend_rho"), b.R("tend_theta"), b.R("tend_w"), b.R("zgrid"), b.R("cofwr"), b.R("cofwz"), b.R("w@2"),
        b.R("ru"), b.R("ru_save"), b.R("rw"), b.R("rw_save"), b.R("fzm"), b.R("fzp"), b.R("rdzw"), b.R("dcEdge"), b.R("invDcEdge"), b.R("invAreaCell"),
        b.R("cofrz"), b.R("dvEdge"), b.I("nEdgesOnCell"), b.I("cellsOnEdge"), b.I("edgesOnCell"), b.R("edgesOnCell_sign"), dts, small_step,
        (real)b.c("config_epssm", 0.1), (real)b.c("cf1"), (real)b.c("cf2"), (real)b.c("cf3"), b.R("specZoneMaskEdge"), b.R("specZoneMaskCell"));
}

int call_divdamp(DBlock& b, real dts) {
    Dims d = dims(b);
    return mpas_b200_divergence_damping_3d(d.L, d.nC, d.nE, d.nCS, 1, d.nE, b.R("theta_m@1"), b.R("ru_p"), b.R("rtheta_pp"), b.R("rtheta_pp_old"),
                                           b.I("cellsOnEdge"), b.R("specZoneMaskEdge"), dts, (real)b.c("config_smdiv", 0.1), (real)b.c("config_len_disp"));
}

int call_recover(DBlock& b, real dt, int ns_, int rk_step) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    return mpas_b200_recover_large_step_variables_work(
        d.L, d.mE, d.ns, d.nC, d.nE, d.nCS, d.nES, dt, ns_, rk_step, b.R("wwAvg"), b.R("rw_save"), b.R("w@2"), b.R("rw"), b.R("rw_p"), b.R("rtheta_p"),
        b.R("rtheta_pp"), b.R("rtheta_p_save"), b.Ropt("rt_diabatic_tend"), b.R("rho_p"), b.R("rho_p_save"), b.R("rho_pp"), b.R("rho_zz@2"), b.R("rho_base"),
        b.R("ruAvg"), b.R("ru_save"), b.R("ru_p"), b.R("u@2"), b.R("ru"), b.R("exner"), b.R("exner_base"), b.R("rtheta_base"), b.R("pressure_p"), b.R("zz"),
        b.R("theta_m@2"), b.Ropt("pressure_base"), b.R("scalars@2"), b.R("fzm"), b.R("fzp"), nullptr, nullptr, b.R("zb_cell"), b.R("zb3_cell"),
        b.R("edgesOnCell_sign"), b.I("cellsOnEdge"), b.I("edgesOnCell"), b.I("nEdgesOnCell"), (real)b.c("cf1"), (real)b.c("cf2"), (real)b.c("cf3"),
        b.I("bdyMaskCell"), RANGES);
}

int call_advance_scalars(DBlock& b, real dt, int rk_step, int order, int advance_density) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    return mpas_b200_advance_scalars_work(d.ns, d.nC, d.L, d.mE, dt, RANGES, (real)b.c("config_coef_3rd_order", 0.25), b.R("scalars@1"), b.R("scalars@2"),
                                          b.R("rho_zz@1"), b.R("rho_zz@2"), nullptr, b.R("ruAvg"), b.R("wwAvg"), nullptr, b.R("dvEdge"), b.I("cellsOnEdge"),
                                          b.I("edgesOnCell"), b.I("nEdgesOnCell"), b.R("edgesOnCell_sign"), b.R("invAreaCell"), b.R("scalars_tend"), b.R("fzm"),
                                          b.R("fzp"), b.R("rdzw"), nullptr, nullptr, b.I("bdyMaskCell"), b.I("bdyMaskEdge"), b.I("nAdvCellsForEdge"),
                                          b.I("advCellsForEdge"), b.R("adv_coefs"), b.R("adv_coefs_3rd"), nullptr, nullptr, nullptr, d.nCS, d.nE,
                                          b.R("horiz_flux_arr"), rk_step, order, advance_density);
}

int call_rk_setup(DBlock& b) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    return mpas_b200_rk_integration_setup(d.L, d.ns, d.nC, d.nE, RANGES, b.R("ru"), b.R("ru_save"), b.R("rw"), b.R("rw_save"), b.R("rtheta_p"),
                                          b.R("rtheta_p_save"), b.R("rho_p"), b.R("rho_p_save"), b.R("rho_zz_old_split"), b.R("u@1"), b.R("u@2"), b.R("w@1"),
                                          b.R("w@2"), b.R("theta_m@1"), b.R("theta_m@2"), b.R("rho_zz@1"), b.R("rho_zz@2"), b.R("scalars@1"), b.R("scalars@2"));
}

int call_moist(DBlock& b) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    return mpas_b200_compute_moist_coefficients(d.L, d.ns, d.nC, d.nE, d.nCS, d.ms, d.me, RANGES, b.I("cellsOnEdge"), b.R("scalars@2"), b.R("cqw"), b.R("cqu"),
                                                b.R("qtot"));
}

int call_substep_finish(DBlock& b, int dynamics_substep, int dynamics_split) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    return mpas_b200_rk_dynamics_substep_finish(d.L, d.nC, d.nE, dynamics_substep, dynamics_split, RANGES, b.R("ru"), b.R("ru_save"), b.R("rw"), b.R("rw_save"),
                                                b.R("rtheta_p"), b.R("rtheta_p_save"), b.R("rho_p"), b.R("rho_p_save"), b.R("rho_zz_old_split"), b.R("ruAvg"),
                                                b.R("ruAvg_split"), b.R("wwAvg"), b.R("wwAvg_split"), b.R("u@1"), b.R("u@2"), b.R("w@1"), b.R("w@2"),
                                                b.R("theta_m@1"), b.R("theta_m@2"), b.R("rho_zz@1"), b.R("rho_zz@2"));
}

int call_init_coupled(DBlock& b) {
    Dims d = dims(b);
    Ranges r = ranges(b);
    return mpas_b200_init_coupled_diagnostics(d.L, d.mE, d.ns, (int)b.c("index_qv", 1), d.nC, d.nE, RANGES, b.I("cellsOnEdge"), b.I("nEdgesOnCell"),
                                              b.I("edgesOnCell"), b.R("edgesOnCell_sign"), b.R("theta_m@1"), b.R("theta"), b.R("rho_zz@1"), b.R("rho"),
                                              b.R("rho_p"), b.R("rho_base"), b.R("rtheta_base"), b.R("theta_base"), b.R("rtheta_p"), b.R("zz"), b.R("scalars@1"),
                                              b.R("ru"), b.R("rw"), b.R("u@1"), b.R("w@1"), b.R("pressure_p"), b.R("pressure_base"), b.R("exner"),
                                              b.R("exner_base"), b.R("fzm"), b.R("fzp"), b.R("zb_cell"), b.R("zb3_cell"));
}

// mpas_reconstruct at the end of atm_srk3 (TI:1594): u time level 2, owned cells
int call_reconstruct(DBlock& b) {
    if (!b.f.count("coeffs_reconstruct") || !b.f.count("uReconstructX")) return 0;
    Dims d = dims(b);
    return mpas_b200_reconstruct_2d(d.L, d.mE, d.nC, d.nE, d.nCS, (int)b.c("on_a_sphere", 1), b.R("u@2"), b.R("coeffs_reconstruct"), b.I("nEdgesOnCell"),
                                    b.I("edgesOnCell"), b.R("latCell"), b.R("lonCell"), b.R("uReconstructX"), b.R("uReconstructY"), b.R("uReconstructZ"),
                                    b.R("uReconstructZonal"), b.R("uReconstructMeridional"));
}

// ---- mono: the two embedded exchanges (TI:3757, 4098) need every block of the process at the same point
struct MonoCtx {
    Dycore* m;
    int iblock;
};
int mono_halo(void* ctx, int which, real* array) {
    MonoCtx* c = (MonoCtx*)ctx;
    Dycore* m = c->m;
    const bool threaded = m->blocks.size() > 1;
    if (which == 2) {
        // all-scalars scale factors (2*num_scalars*nVertLevels, nCells+1) in the library's per-block workspace: a view,
        // registered by every block's thread before the barrier
        DBlock& b = m->blocks[c->iblock];
        Dims d = dims(b);
        DField df;
        df.dev = array; df.kind = 0; df.loc = 1; df.inner = 2 * d.ns * d.L; df.owned = false;
        df.dims = {2LL * d.ns * d.L, (long long)d.nC + 1};
        df.bytes = (size_t)df.inner * (size_t)(d.nC + 1) * sizeof(real);
        b.f["mono_scale_all"] = df;
    }
    if (threaded && !m->bar.wait()) return 1;  // another block's thread failed: do not wait for it
    int rc = 0;
    if (c->iblock == 0) {
        try {
            if (which == 0) rc = dycore_exchange(*m, {{"scalars@1", 0}});
            else if (which == 2) rc = dycore_exchange(*m, {{"mono_scale_all", 1}});
            else rc = dycore_exchange(*m, {{"scale_arr", 1}});
        } catch (const std::exception& e) {
            m->err = e.what();
            rc = 1;
        }
    }
    if (rc) m->bar.abort();
    if (threaded && !m->bar.wait()) return 1;
    return rc;
}

int call_mono(Dycore& m, int ib, real dt, int advance_density) {
    DBlock& b = m.blocks[ib];
    Dims d = dims(b);
    Ranges r = ranges(b);
    MonoCtx ctx{&m, ib};
    const bool need_halo = m.blocks.size() > 1 || m.nranks > 1;
    return mpas_b200_advance_scalars_mono_work(need_halo ? mono_halo : nullptr, &ctx, d.nC, d.nE, d.L, d.mE, dt, RANGES, (real)b.c("config_coef_3rd_order", 0.25),
                                               d.nCS, d.ns, b.R("ruAvg"), b.R("wwAvg"), b.R("scalars_tend"), b.R("rho_zz@1"), b.R("rho_zz@2"), b.R("scalars@1"),
                                               b.R("scalars@2"), b.R("invAreaCell"), b.R("dvEdge"), b.I("cellsOnEdge"), b.I("cellsOnCell"), b.I("edgesOnCell"),
                                               b.R("edgesOnCell_sign"), b.I("nEdgesOnCell"), b.R("fzm"), b.R("fzp"), b.R("rdzw"), b.I("nAdvCellsForEdge"),
                                               b.I("advCellsForEdge"), b.R("adv_coefs"), b.R("adv_coefs_3rd"), b.R("scalar_old"), b.R("scalar_new"), b.R("s_max"),
                                               b.R("s_min"), b.R("wdtn"), b.R("scale_arr"), b.R("flux_arr"), b.R("flux_upwind_tmp"), b.R("flux_tmp"),
                                               b.I("bdyMaskCell"), b.I("bdyMaskEdge"), advance_density, b.R("rho_zz_int"));
}

int run_mono_all(Dycore& m, real dt, int advance_density) {
    const int P = (int)m.blocks.size();
    if (P == 1) return call_mono(m, 0, dt, advance_density);
    m.bar.reset(P);
    m.err.clear();
    std::vector<std::thread> th;
    std::vector<int> rcs(P, 0);
    for (int ib = 0; ib < P; ++ib)
        th.emplace_back([&m, ib, dt, advance_density, &rcs] {
            try {
                rcs[ib] = call_mono(m, ib, dt, advance_density);
            } catch (const std::exception& e) {
                std::lock_guard<std::mutex> lk(m.err_mu);
                m.err = e.what();
                rcs[ib] = 1;
            }
            if (rcs[ib]) m.bar.abort();  // a thread that fails before a barrier releases the others
        });
    for (auto& t : th) t.join();
    if (!m.err.empty()) return fail_msg(m.err);
    for (int rc : rcs)
        if (rc) return rc;
    return 0;
}

__global__ void clamp_nonneg_kernel(real* a, long long n) {  // TI:1646-1647
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n && a[t] < (real)0.0) a[t] = 0.0;
}

// exchange helper honouring the "aggregate" option: adjacent reference exchanges either go out as one
// grouped message per neighbour or one by one in the reference's order
int xchg(Dycore& m, const std::vector<XField>& fields) {
    Phase ph(m, "halo_exchange");
    if (m.aggregate) return dycore_exchange(m, fields);
    for (const XField& f : fields) CK(dycore_exchange(m, {f}));
    return 0;
}

}  // namespace

// atm_mpas_init_block + atm_core_init exchanges (mpas_atm_core.F:145, 163, 180-186, 390-402)
int dycore_init(Dycore& m) {
    // The reference reads the initial state INCLUDING halo elements from the global init file ('input' stream,
    // mpas_atm_core.F:95).  The harness initialises each block on its own, so fields built with horizontal
    // stencils are wrong in the outer halo; fill them from their owners first (not counted as a step exchange).
    {
        const long keep = m.nExchanges;
        for (const char* n : {"w@1", "zb_cell", "zb3_cell", "zxu", "defc_a", "defc_b"})
            if (m.blocks[0].f.count(n)) CK(dycore_exchange(m, {{n, 0}}));
        m.nExchanges = keep;
    }
    CK(dycore_exchange(m, {{"u@1", 0}}));  // :145
    for (auto& b : m.blocks) {
        if (b.c("config_do_restart", 0) == 0) CK(call_init_coupled(b));  // mpas_atm_core.F:389
        CK(call_solve_diagnostics(b, 1, (real)b.c("config_dt"), -1));
    }
    CK(dycore_exchange(m, {{"pv_edge", 0}, {"ru", 0}, {"rw", 0}}));  // :180-186
    cudaError_t e = cudaStreamSynchronize(stream());
    if (e != cudaSuccess) return fail("dycore_init", e);
    return 0;
}

// atm_srk3, TI:142-1796
static int srk3_body(Dycore& m) {
    DBlock& b0 = m.blocks[0];
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
    if (order == 3) {  // TI:298-310
        rk_timestep[1] = dt_dynamics / (real)3.; rk_timestep[2] = dt_dynamics / (real)2.; rk_timestep[3] = dt_dynamics;
        rk_sub_timestep[1] = dt_dynamics / (real)3.; rk_sub_timestep[2] = dt_dynamics / (real)number_of_sub_steps; rk_sub_timestep[3] = dt_dynamics / (real)number_of_sub_steps;
        number_sub_steps[1] = 1; number_sub_steps[2] = std::max(1, number_of_sub_steps / 2); number_sub_steps[3] = number_of_sub_steps;
    } else {  // TI:312-326
        rk_timestep[1] = dt_dynamics / (real)2.; rk_timestep[2] = dt_dynamics / (real)2.; rk_timestep[3] = dt_dynamics;
        rk_sub_timestep[1] = rk_sub_timestep[2] = rk_sub_timestep[3] = dt_dynamics / (real)number_of_sub_steps;
        number_sub_steps[1] = std::max(1, number_of_sub_steps / 2); number_sub_steps[2] = std::max(1, number_of_sub_steps / 2); number_sub_steps[3] = number_of_sub_steps;
    }
    CK(xchg(m, {{"theta_m@1", 0}, {"scalars@1", 0}, {"pressure_p", 0}, {"rtheta_p", 0}}));  // TI:329-338
    {
        Phase ph(m, "atm_rk_integration_setup");
        for (auto& b : m.blocks) CK(call_rk_setup(b));  // TI:369
    }
    {
        Phase ph(m, "atm_compute_moist_coefficients");
        for (auto& b : m.blocks) CK(call_moist(b));  // TI:410
    }
    for (int dynamics_substep = 1; dynamics_substep <= dynamics_split; ++dynamics_substep) {
        {
            Phase ph(m, "atm_compute_vert_imp_coefs");
            for (auto& b : m.blocks) CK(call_vert_imp_coefs(b, rk_sub_timestep[1]));  // TI:501
        }
        CK(xchg(m, {{"exner", 0}}));  // TI:513
        for (int rk_step = 1; rk_step <= 3; ++rk_step) {
            if (order == 3 && rk_step == 2) {
                Phase ph(m, "atm_compute_vert_imp_coefs");
                for (auto& b : m.blocks) CK(call_vert_imp_coefs(b, rk_sub_timestep[rk_step]));  // TI:550
            }
            {
                Phase ph(m, "atm_compute_dyn_tend");
                for (auto& b : m.blocks) {
                    if (rk_step == 1) CK(dycore_zero_garbage(b, 1));  // TI:593-605
                    CK(call_dyn_tend(b, rk_step, dt));  // TI:609
                }
            }
            CK(xchg(m, {{"tend_u", 1}}));  // TI:642
            {
                Phase ph(m, "small_step_prep");
                for (auto& b : m.blocks) CK(call_smlstep(b));  // TI:669
            }
            // The edge pass of an acoustic substep (TI:2552-2599) is folded into the divergence damping of the substep before it (the
            // first one: into its own), DESIGN.md section 3 rule 12: the damping kernel then needs the rho_pp halo of the NEXT
            // substep, so that exchange (TI:792) moves up from the top of the next substep to just before the damping.
            static const int fold_tune = tune("ac_fold_first", 1);
            const bool fold = fold_tune && m.fold_first;
            for (int small_step = 1; small_step <= number_sub_steps[rk_step]; ++small_step) {
                const bool next = fold && small_step < number_sub_steps[rk_step], done = fold && small_step > 1;
                if (!done) CK(xchg(m, {{"rho_pp", 1}}));  // TI:792
                {
                    Phase ph(m, "atm_advance_acoustic_step");
                    for (auto& b : m.blocks) {
                        if (fold) acoustic_fold_next_call((small_step == 1 ? ACOUSTIC_FOLD_FIRST : 0) | (next ? ACOUSTIC_FOLD_NEXT : 0) | (done ? ACOUSTIC_FOLD_DONE : 0));
                        CK(call_acoustic(b, rk_sub_timestep[rk_step], small_step));  // TI:824
                    }
                }
                CK(xchg(m, {{"rtheta_pp", 1}}));  // TI:845
                if (next) CK(xchg(m, {{"rho_pp", 1}}));  // TI:792 of the next substep
                {
                    Phase ph(m, "atm_divergence_damping_3d");
                    for (auto& b : m.blocks) CK(call_divdamp(b, rk_sub_timestep[rk_step]));  // TI:862
                }
            }
            CK(xchg(m, {{"rw_p", 0}, {"ru_p", 0}, {"rho_pp", 0}, {"rtheta_pp", 2}}));  // TI:876-887
            {
                Phase ph(m, "atm_recover_large_step_variables");
                for (auto& b : m.blocks) CK(call_recover(b, rk_timestep[rk_step], number_sub_steps[rk_step], rk_step));  // TI:917
            }
            CK(xchg(m, {{"u@2", 0}}));             // TI:988
            if (scalar_advection && !split) {  // TI:993-1099
                if (rk_step < 3 || (!monotonic && !positive_definite)) {
                    Phase ph(m, "atm_advance_scalars");
                    for (auto& b : m.blocks) {
                        CK(dycore_zero_garbage(b, 3));  // TI:1044-1045
                        CK(call_advance_scalars(b, rk_timestep[rk_step], rk_step, order, 0));
                    }
                } else {
                    Phase ph(m, "atm_advance_scalars_mono");
                    for (auto& b : m.blocks) CK(dycore_zero_garbage(b, 2));  // TI:1029-1050
                    CK(run_mono_all(m, rk_timestep[rk_step], 0));
                }
            }
            {
                Phase ph(m, "atm_compute_solve_diagnostics");
                for (auto& b : m.blocks) CK(call_solve_diagnostics(b, 2, dt, rk_step));  // TI:1216
            }
            if (scalar_advection && !split) CK(xchg(m, {{"w@2", 0}, {"pv_edge", 0}, {"rho_edge", 0}, {"scalars@2", 0}}));  // TI:1234-1249
            else CK(xchg(m, {{"w@2", 0}, {"pv_edge", 0}, {"rho_edge", 0}}));
        }
        if (dynamics_substep < dynamics_split) CK(xchg(m, {{"theta_m@2", 0}, {"pressure_p", 0}, {"rtheta_p", 0}}));  // TI:1282-1288
        {
            Phase ph(m, "atm_rk_dynamics_substep_finish");
            for (auto& b : m.blocks) CK(call_substep_finish(b, dynamics_substep, dynamics_split));  // TI:1329
        }
    }
    if (scalar_advection && split) {  // TI:1355-1576
        real rk_t[4] = {0, dt / (real)3., dt / (real)2., dt};
        if (order == 2) rk_t[1] = dt / (real)2.;
        for (int rk_step = 1; rk_step <= 3; ++rk_step) {
            if (rk_step < 3 || (!monotonic && !positive_definite)) {
                Phase ph(m, "atm_advance_scalars");
                for (auto& b : m.blocks) {
                    CK(dycore_zero_garbage(b, 3));  // TI:1417-1418
                    CK(call_advance_scalars(b, rk_t[rk_step], rk_step, order, 1));
                }
            } else {
                Phase ph(m, "atm_advance_scalars_mono");
                for (auto& b : m.blocks) CK(dycore_zero_garbage(b, 2));  // TI:1402-1423
                CK(run_mono_all(m, rk_t[rk_step], 1));
            }
            if (rk_step < 3) CK(xchg(m, {{"scalars@2", 0}}));  // TI:1569-1572
        }
    }
    {
        Phase ph(m, "mpas_reconstruct");
        for (auto& b : m.blocks) CK(call_reconstruct(b));  // TI:1594
    }
    // DO_PHYSICS (default build, src/core_atmosphere/Makefile:3): negative mixing ratios reset, TI:1646-1647
    for (auto& b : m.blocks) {
        Dims d = dims(b);
        const long long n = (long long)d.ns * d.L * (d.nC + 1);
        clamp_nonneg_kernel<<<cdiv(n, 256), 256, 0, stream()>>>(b.R("scalars@2"), n);
        count_launch();
    }
    return 0;
}

int dycore_step(Dycore& m) {
    int rc = 0;
    // The launch sequence of a step depends only on which buffer currently is time level 1: two graphs
    // (one per parity) replay the whole step with a single launch each.  NCCL send/recv are captured too.
    // The first step of each parity runs eagerly: it sizes the library workspace and builds the halo plans
    // (both allocate, which a capture forbids).
    const bool graphable = m.use_graph && !m.timing && m.blocks.size() == 1 && m.steps_done >= 2;
    if (graphable) {
        auto capture = [&](int par) -> int {
            cudaGraph_t g;
            const long long before = launches();
            const long xbefore = m.nExchanges;
            cudaError_t e = cudaStreamBeginCapture(stream(), cudaStreamCaptureModeThreadLocal);
            if (e != cudaSuccess) return fail("cudaStreamBeginCapture", e);
            // an exception out of the captured body (a missing field or config) must not leave the stream in capture mode:
            // every later call on it would fail with a capture error that hides the real message
            int r = 0;
            std::string what;
            try {
                r = srk3_body(m);
            } catch (const std::exception& ex) {
                what = ex.what();
                r = 1;
            }
            e = cudaStreamEndCapture(stream(), &g);
            if (!what.empty()) {
                if (e == cudaSuccess && g) cudaGraphDestroy(g);
                cudaGetLastError();
                count_launch(-(int)(launches() - before));
                m.nExchanges = xbefore;
                return fail_msg(what);
            }
            m.graph_nodes[par] = launches() - before;  // kernels / copies recorded, not yet executed
            m.graph_exchanges[par] = m.nExchanges - xbefore;
            count_launch(-(int)m.graph_nodes[par]);
            m.nExchanges = xbefore;
            if (r) {
                if (e == cudaSuccess && g) cudaGraphDestroy(g);
                return r;
            }
            if (e != cudaSuccess) return fail("cudaStreamEndCapture", e);
            e = cudaGraphInstantiate(&m.graph_exec[par], g, 0);
            cudaGraphDestroy(g);
            if (e != cudaSuccess) return fail("cudaGraphInstantiate", e);
            return 0;
        };
        auto swap_all = [&]() {
            for (auto& b : m.blocks)
                for (const char* n : {"u", "w", "theta_m", "rho_zz", "scalars"}) b.swap_levels(n);
            m.parity ^= 1;
        };
        if (!m.graph_exec[m.parity]) {
            // both graphs are built on the first graphable step (the third): capture records without executing, so the
            // other parity is captured with the time levels swapped and swapped back
            if ((rc = capture(m.parity))) return rc;
            if (!m.graph_exec[m.parity ^ 1]) {
                swap_all();
                rc = capture(m.parity);
                swap_all();
                if (rc) return rc;
            }
        }
        cudaGraphExec_t& ge = m.graph_exec[m.parity];
        cudaError_t e = cudaGraphLaunch(ge, stream());
        if (e != cudaSuccess) return fail("cudaGraphLaunch", e);
        count_launch((int)m.graph_nodes[m.parity]);
        m.nExchanges += m.graph_exchanges[m.parity];
    } else {
        rc = srk3_body(m);
        if (rc) return rc;
    }
    // mpas_pool_shift_time_levels(state), mpas_atm_core.F:671
    for (auto& b : m.blocks)
        for (const char* n : {"u", "w", "theta_m", "rho_zz", "scalars"}) b.swap_levels(n);
    m.parity ^= 1;
    m.steps_done++;
    flush_timers(m);
    // both time-level parities have built their exchange plans: switch the exchanges from NCCL send/recv to
    // peer-to-peer stores when every rank can map every other rank's window (collective, before the graphs are captured)
    if (m.steps_done == 2 && m.nranks > 1) return dycore_p2p_setup(m);
    return 0;
}

// Run one work routine on the current fields of every local block (routine-level parity tests).
// params: acoustic (dts, small_step) | divdamp (dts) | recover (dt, ns, rk_step) | dyn_tend (rk_step, dt) |
// mono (dt, advance_density) | advance_scalars (dt, rk_step, order, advance_density) |
// solve_diagnostics (time_lev, dt, rk_step) | vert_imp_coefs (dts) | substep_finish (substep, split)
int dycore_call(Dycore& m, const std::string& rt, const double* q, int nq) {
    auto need = [&](int n) { if (nq < n) throw std::runtime_error("dycore_call: too few parameters for " + rt); };
    if (rt == "mono") {
        need(2);
        for (auto& b : m.blocks) CK(dycore_zero_garbage(b, 2));
        return run_mono_all(m, (real)q[0], (int)q[1]);
    }
    for (auto& b : m.blocks) {
        if (rt == "acoustic") { need(2); CK(call_acoustic(b, (real)q[0], (int)q[1])); }
        else if (rt == "divdamp") { need(1); CK(call_divdamp(b, (real)q[0])); }
        else if (rt == "recover") { need(3); CK(call_recover(b, (real)q[0], (int)q[1], (int)q[2])); }
        else if (rt == "dyn_tend") { need(2); if ((int)q[0] == 1) CK(dycore_zero_garbage(b, 1)); CK(call_dyn_tend(b, (int)q[0], (real)q[1])); }
        else if (rt == "advance_scalars") { need(4); CK(dycore_zero_garbage(b, 3)); CK(call_advance_scalars(b, (real)q[0], (int)q[1], (int)q[2], (int)q[3])); }
        else if (rt == "solve_diagnostics") { need(3); CK(call_solve_diagnostics(b, (int)q[0], (real)q[1], (int)q[2])); }
        else if (rt == "vert_imp_coefs") { need(1); CK(call_vert_imp_coefs(b, (real)q[0])); }
        else if (rt == "smlstep") CK(call_smlstep(b));
        else if (rt == "rk_setup") CK(call_rk_setup(b));
        else if (rt == "moist") CK(call_moist(b));
        else if (rt == "substep_finish") { need(2); CK(call_substep_finish(b, (int)q[0], (int)q[1])); }
        else if (rt == "init_coupled") CK(call_init_coupled(b));
        else if (rt == "reconstruct") CK(call_reconstruct(b));
        else if (rt == "output_diagnostics") {  // time level 1 after a step, as atm_core_run calls it (mpas_atm_core.F:717)
            Dims d = dims(b);
            CK(mpas_b200_compute_output_diagnostics(d.L, d.ns, (int)b.c("index_qv", 1), d.nC, b.R("theta_m@1"), b.R("rho_zz@1"), b.R("scalars@1"), b.R("zz"),
                                                    b.R("pressure_base"), b.R("pressure_p"), b.R("theta"), b.R("rho"), b.R("pressure")));
        }
        else return fail_msg("dycore_call: unknown routine " + rt);
    }
    return 0;
}

// summarize_timestep analogue (TI:6675-7018): global max |field| over owned elements
namespace {
__global__ void max_abs_kernel(const real* a, long long n, real* out) {
    __shared__ real sm[256];
    real v = 0.0;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) v = fmax(v, fabs(a[t]));
    sm[threadIdx.x] = v;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) sm[threadIdx.x] = fmax(sm[threadIdx.x], sm[threadIdx.x + s]);
        __syncthreads();
    }
    if (threadIdx.x == 0) out[blockIdx.x] = sm[0];
}
}  // namespace

double dycore_global_max_abs(Dycore& m, const std::string& name) {
    const int NB = 256;
    if (!m.d_reduce) cudaMalloc(&m.d_reduce, (NB + 2) * sizeof(real));
    double result = 0.0;
    for (auto& b : m.blocks) {
        const DField& f = b.f.at(name);
        long long nOwned = f.loc == 1 ? (long long)b.c("nCellsSolve") : f.loc == 2 ? (long long)b.c("nEdgesSolve") : (long long)b.c("nVerticesSolve");
        const long long n = nOwned * f.inner;
        max_abs_kernel<<<NB, 256, 0, stream()>>>((const real*)f.dev, n, m.d_reduce);
        max_abs_kernel<<<1, 256, 0, stream()>>>(m.d_reduce, NB, m.d_reduce + NB);
        count_launch(2);
        if (m.nranks > 1 && m.have_comm) m.nccl.allreduce_max(m.d_reduce + NB, m.d_reduce + NB, 1, stream());
        real v = 0;
        cudaMemcpyAsync(&v, m.d_reduce + NB, sizeof(real), cudaMemcpyDeviceToHost, stream());
        cudaStreamSynchronize(stream());
        result = std::max(result, (double)v);
    }
    return result;
}

// global min and max over the owned elements of a field (every stride-th element from offset: one scalar of the
// (num_scalars, nVertLevels, nCells) array), the reductions summarize_timestep prints (TI:6758-6790 w, :6815-6850 u, :6985-7010 scalars)
namespace {
__global__ void minmax_kernel(const real* a, long long n, long long offset, long long stride, real* out, int nout) {
    __shared__ real smin[256], smax[256];
    real lo = INFINITY, hi = -INFINITY;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const real v = a[offset + t * stride];
        lo = fmin(lo, v);
        hi = fmax(hi, v);
    }
    smin[threadIdx.x] = lo;
    smax[threadIdx.x] = hi;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
            smin[threadIdx.x] = fmin(smin[threadIdx.x], smin[threadIdx.x + s]);
            smax[threadIdx.x] = fmax(smax[threadIdx.x], smax[threadIdx.x + s]);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {  // out[2b] = max, out[2b + 1] = -min: both reduce with max
        out[2 * blockIdx.x] = smax[0];
        out[2 * blockIdx.x + 1] = -smin[0];
    }
    (void)nout;
}
__global__ void minmax_final_kernel(const real* part, int nb, real* out) {
    __shared__ real sa[256], sb[256];
    real a = -INFINITY, b = -INFINITY;
    for (int i = threadIdx.x; i < nb; i += blockDim.x) {
        a = fmax(a, part[2 * i]);
        b = fmax(b, part[2 * i + 1]);
    }
    sa[threadIdx.x] = a;
    sb[threadIdx.x] = b;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
            sa[threadIdx.x] = fmax(sa[threadIdx.x], sa[threadIdx.x + s]);
            sb[threadIdx.x] = fmax(sb[threadIdx.x], sb[threadIdx.x + s]);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        out[0] = sa[0];
        out[1] = sb[0];
    }
}
}  // namespace

int dycore_global_minmax(Dycore& m, const std::string& name, long long offset, long long stride, double* mn, double* mx) {
    const int NB = 256;
    if (stride < 1 || offset < 0 || offset >= stride) return fail_msg("mpas_b200_dycore_global_minmax: need 0 <= offset < stride");
    if (!m.d_minmax) {
        cudaError_t e = cudaMalloc(&m.d_minmax, (2 * NB + 2) * sizeof(real));
        if (e != cudaSuccess) return fail("cudaMalloc minmax", e);
    }
    double lo = INFINITY, hi = -INFINITY;
    for (auto& b : m.blocks) {
        auto it = b.f.find(name);
        if (it == b.f.end()) return fail_msg("mpas_b200_dycore_global_minmax: no field " + name);
        const DField& f = it->second;
        if (f.kind != 0) return fail_msg("mpas_b200_dycore_global_minmax: " + name + " is not a real field");
        const long long nOwned = f.loc == 1 ? (long long)b.c("nCellsSolve") : f.loc == 2 ? (long long)b.c("nEdgesSolve") : (long long)b.c("nVerticesSolve");
        if ((nOwned * f.inner) % stride) return fail_msg("mpas_b200_dycore_global_minmax: stride does not divide the field");
        const long long n = nOwned * f.inner / stride;
        minmax_kernel<<<NB, 256, 0, stream()>>>((const real*)f.dev, n, offset, stride, m.d_minmax, NB);
        minmax_final_kernel<<<1, 256, 0, stream()>>>(m.d_minmax, NB, m.d_minmax + 2 * NB);
        count_launch(2);
        if (m.nranks > 1 && m.have_comm) m.nccl.allreduce_max(m.d_minmax + 2 * NB, m.d_minmax + 2 * NB, 2, stream());
        real v[2] = {0, 0};
        cudaMemcpyAsync(v, m.d_minmax + 2 * NB, 2 * sizeof(real), cudaMemcpyDeviceToHost, stream());
        cudaError_t e = cudaStreamSynchronize(stream());
        if (e != cudaSuccess) return fail("mpas_b200_dycore_global_minmax", e);
        hi = std::max(hi, (double)v[0]);
        lo = std::min(lo, -(double)v[1]);
    }
    *mn = lo;
    *mx = hi;
    return dycore_check_p2p(m);
}

// Order-independent checksum of the owned elements of a field: the sum (mod 2^64) of their bit patterns.  Owned-element
// arithmetic does not depend on the partition, so the sum over all ranks is the same number for 1, 2, 4, 8 GPUs -- the
// decomposition test of the reference (COMPASS decomp_test: P ranks == 1 rank, bitwise) as one printable integer.
namespace {
__global__ void checksum_kernel(const real* a, long long n, unsigned long long* out) {
    unsigned long long v = 0;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        real x = a[t];
        if (x == (real)0.0) x = (real)0.0;  // -0 and +0 are the same value
        if (sizeof(real) == 8) v += (unsigned long long)__double_as_longlong((double)x);
        else v += (unsigned long long)__float_as_uint((float)x);
    }
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, v);
}
}  // namespace

int dycore_checksum(Dycore& m, const std::string& name, unsigned long long* out) {
    unsigned long long* d = nullptr;
    cudaError_t e = cudaMalloc(&d, sizeof(unsigned long long));
    if (e != cudaSuccess) return fail("cudaMalloc checksum", e);
    cudaMemsetAsync(d, 0, sizeof(unsigned long long), stream());
    for (auto& b : m.blocks) {
        auto it = b.f.find(name);
        if (it == b.f.end() || it->second.kind != 0 || it->second.loc < 1 || it->second.loc > 3) {
            cudaFree(d);
            return fail_msg("mpas_b200_dycore_checksum: no decomposed real field " + name);
        }
        const DField& f = it->second;
        const long long nOwned = f.loc == 1 ? (long long)b.c("nCellsSolve") : f.loc == 2 ? (long long)b.c("nEdgesSolve") : (long long)b.c("nVerticesSolve");
        checksum_kernel<<<592, 256, 0, stream()>>>((const real*)f.dev, nOwned * f.inner, d);
        count_launch();
    }
    cudaMemcpyAsync(out, d, sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream());
    e = cudaStreamSynchronize(stream());
    cudaFree(d);
    if (e != cudaSuccess) return fail("mpas_b200_dycore_checksum", e);
    return dycore_check_p2p(m);
}

}  // namespace mpasb200

extern "C" int mpas_b200_dycore_checksum(void* h, const char* name, unsigned long long* out) {
    try {
        return mpasb200::dycore_checksum(*(mpasb200::Dycore*)h, name, out);
    } catch (const std::exception& e) { return mpasb200::fail_msg(e.what()); }
}

extern "C" int mpas_b200_dycore_global_minmax(void* h, const char* name, long long offset, long long stride, double* mn, double* mx) {
    try {
        return mpasb200::dycore_global_minmax(*(mpasb200::Dycore*)h, name, offset, stride, mn, mx);
    } catch (const std::exception& e) { return mpasb200::fail_msg(e.what()); }
}

extern "C" double mpas_b200_dycore_global_max_abs(void* h, const char* name) {
    try {
        return mpasb200::dycore_global_max_abs(*(mpasb200::Dycore*)h, name);
    } catch (const std::exception& e) {
        mpasb200::fail_msg(e.what());
        return -1.0;
    }
}
