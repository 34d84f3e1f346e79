/* Argument lists of the dycore work routines, shared by the product C ABI (mpas_b200_*) and
 * the CPU oracle (oracle_*) so both are called with the identical tuple.
 *
 * Each list follows the dummy-argument order of the reference subroutine it replaces
 * (file:line cited per macro; TI = src/core_atmosphere/dynamics/mpas_atm_time_integration.F),
 * preceded by the dimensions the reference takes from module mpas_atm_dimensions
 * (src/core_atmosphere/mpas_atm_dimensions.F:18-27) and followed by the module-level
 * scratch arrays the reference reaches by host association (TI:35-43, 70-71).
 * Arrays: contiguous, Fortran order (k fastest), extent includes the +1 garbage column.
 * Integers are 32-bit, index arrays 1-based, index ranges inclusive 1-based.
 */
#ifndef MPAS_B200_ARGS_H
#define MPAS_B200_ARGS_H

#ifdef SINGLE_PRECISION
typedef float mpas_real;
#else
typedef double mpas_real;
#endif

#define MPAS_RANGE_ARGS                                                                            \
    int cellStart, int cellEnd, int vertexStart, int vertexEnd, int edgeStart, int edgeEnd,        \
        int cellSolveStart, int cellSolveEnd, int vertexSolveStart, int vertexSolveEnd,            \
        int edgeSolveStart, int edgeSolveEnd

/* atm_advance_acoustic_step_work, TI:2447-2455 */
#define MPAS_ACOUSTIC_ARGS                                                                         \
    int nVertLevels, int maxEdges, int nCells, int nEdges, int nCellsSolve, MPAS_RANGE_ARGS,       \
        const mpas_real *rho_zz, const mpas_real *theta_m, mpas_real *ru_p, mpas_real *rw_p,       \
        mpas_real *rtheta_pp, mpas_real *rtheta_pp_old, const mpas_real *zz, const mpas_real *exner, \
        const mpas_real *cqu, mpas_real *ruAvg, mpas_real *wwAvg, mpas_real *rho_pp,               \
        const mpas_real *cofwt, const mpas_real *coftz, const mpas_real *zxu,                      \
        const mpas_real *a_tri, const mpas_real *alpha_tri, const mpas_real *gamma_tri,            \
        const mpas_real *dss, const mpas_real *tend_ru, const mpas_real *tend_rho,                 \
        const mpas_real *tend_rt, const mpas_real *tend_rw, const mpas_real *zgrid,                \
        const mpas_real *cofwr, const mpas_real *cofwz, const mpas_real *w, const mpas_real *ru,   \
        const mpas_real *ru_save, const mpas_real *rw, const mpas_real *rw_save,                   \
        const mpas_real *fzm, const mpas_real *fzp, const mpas_real *rdzw,                         \
        const mpas_real *dcEdge, const mpas_real *invDcEdge, const mpas_real *invAreaCell,         \
        const mpas_real *cofrz, const mpas_real *dvEdge, const int *nEdgesOnCell,                  \
        const int *cellsOnEdge, const int *edgesOnCell, const mpas_real *edgesOnCell_sign,         \
        mpas_real dts, int small_step, mpas_real epssm, mpas_real cf1, mpas_real cf2,              \
        mpas_real cf3, const mpas_real *specZoneMaskEdge, const mpas_real *specZoneMaskCell

/* atm_divergence_damping_3d, TI:2726 (pool arguments resolved to arrays as at TI:2751-2763) */
#define MPAS_DIVDAMP_ARGS                                                                          \
    int nVertLevels, int nCells, int nEdges, int nCellsSolve, int edgeStart, int edgeEnd,          \
        const mpas_real *theta_m, mpas_real *ru_p, const mpas_real *rtheta_pp,                     \
        const mpas_real *rtheta_pp_old, const int *cellsOnEdge,                                    \
        const mpas_real *specZoneMaskEdge, mpas_real dts, mpas_real smdiv, mpas_real config_len_disp

/* atm_recover_large_step_variables_work, TI:2909-2917 (zb, zb3 edge forms are unused by the
 * routine and may be NULL) */
#define MPAS_RECOVER_ARGS                                                                          \
    int nVertLevels, int maxEdges, int num_scalars, int nCells, int nEdges, int nCellsSolve,       \
        int nEdgesSolve, mpas_real dt, int ns, int rk_step, mpas_real *wwAvg,                      \
        const mpas_real *rw_save, mpas_real *w, mpas_real *rw, const mpas_real *rw_p,              \
        mpas_real *rtheta_p, const mpas_real *rtheta_pp, const mpas_real *rtheta_p_save,           \
        const mpas_real *rt_diabatic_tend, mpas_real *rho_p, const mpas_real *rho_p_save,          \
        const mpas_real *rho_pp, mpas_real *rho_zz, const mpas_real *rho_base, mpas_real *ruAvg,   \
        const mpas_real *ru_save, const mpas_real *ru_p, mpas_real *u, mpas_real *ru,              \
        mpas_real *exner, const mpas_real *exner_base, const mpas_real *rtheta_base,               \
        mpas_real *pressure_p, const mpas_real *zz, mpas_real *theta_m,                            \
        const mpas_real *pressure_b, const mpas_real *scalars, const mpas_real *fzm,               \
        const mpas_real *fzp, const mpas_real *zb, const mpas_real *zb3,                           \
        const mpas_real *zb_cell, const mpas_real *zb3_cell, const mpas_real *edgesOnCell_sign,    \
        const int *cellsOnEdge, const int *edgesOnCell, const int *nEdgesOnCell, mpas_real cf1,    \
        mpas_real cf2, mpas_real cf3, const int *bdyMaskCell, MPAS_RANGE_ARGS

/* atm_compute_vert_imp_coefs_work, TI:2012-2016 (+ module scratch qtot) */
#define MPAS_VERTIMP_ARGS                                                                          \
    int nVertLevels, int num_scalars, int nCells, int moist_start, int moist_end, mpas_real dts,   \
        mpas_real epssm, const mpas_real *zz, const mpas_real *cqw, const mpas_real *p,            \
        const mpas_real *t, const mpas_real *rb, const mpas_real *rtb, const mpas_real *pb,        \
        const mpas_real *rt, mpas_real *cofwr, mpas_real *cofwz, mpas_real *coftz,                 \
        mpas_real *cofwt, mpas_real *a_tri, mpas_real *alpha_tri, mpas_real *gamma_tri,            \
        mpas_real *cofrz, const mpas_real *rdzw, const mpas_real *fzm, const mpas_real *fzp,       \
        const mpas_real *rdzu, const mpas_real *scalars, int cellStart, int cellEnd,               \
        int edgeStart, int edgeEnd, int cellSolveStart, int cellSolveEnd, int edgeSolveStart,      \
        int edgeSolveEnd, const mpas_real *qtot

/* atm_set_smlstep_pert_variables_work, TI:2224-2231 (only the arrays the routine touches) */
#define MPAS_SMLSTEP_ARGS                                                                          \
    int nVertLevels, int maxEdges, int nCells, int nEdges, int nCellsSolve,                        \
        const int *nEdgesOnCell, const int *cellsOnEdge, const int *edgesOnCell,                   \
        const mpas_real *fzm, const mpas_real *fzp, const mpas_real *zb_cell,                      \
        const mpas_real *zb3_cell, const mpas_real *zz, mpas_real *w_tend,                         \
        const mpas_real *u_tend, const int *bdyMaskCell, const mpas_real *edgesOnCell_sign,        \
        int cellStart, int cellEnd, int edgeStart, int edgeEnd, int cellSolveStart,                \
        int cellSolveEnd, int edgeSolveStart, int edgeSolveEnd

/* atm_compute_solve_diagnostics_work, TI:5513-5520 (+ module scratch ke_vertex, ke_edge);
 * rk_step < 0 means "not present" (reconstruct v) */
#define MPAS_SOLVEDIAG_ARGS                                                                        \
    int nVertLevels, int maxEdges, int maxEdges2, int nCells, int nEdges, int nVertices,           \
        int vertexDegree, mpas_real dt, mpas_real config_apvm_upwinding, const mpas_real *fVertex, \
        const mpas_real *fEdge, const mpas_real *invAreaTriangle, const mpas_real *invAreaCell,    \
        const mpas_real *dvEdge, const mpas_real *dcEdge, const mpas_real *invDvEdge,              \
        const mpas_real *invDcEdge, const mpas_real *weightsOnEdge,                                \
        const mpas_real *kiteAreasOnVertex, mpas_real *h_edge, const mpas_real *h,                 \
        const mpas_real *u, mpas_real *v, mpas_real *vorticity, mpas_real *ke, mpas_real *pv_edge, \
        mpas_real *pv_vertex, mpas_real *pv_cell, mpas_real *gradPVn, mpas_real *gradPVt,          \
        mpas_real *divergence, const int *cellsOnEdge, const int *cellsOnVertex,                   \
        const int *verticesOnEdge, const int *edgesOnCell, const int *edgesOnEdge,                 \
        const int *edgesOnVertex, const int *kiteForCell, const int *verticesOnCell,               \
        const mpas_real *edgesOnVertex_sign, const mpas_real *edgesOnCell_sign,                    \
        const int *nEdgesOnCell, const int *nEdgesOnEdge, int cellStart, int cellEnd,              \
        int vertexStart, int vertexEnd, int edgeStart, int edgeEnd, int rk_step,                   \
        mpas_real *ke_vertex, mpas_real *ke_edge

/* atm_compute_dyn_tend_work, TI:4481-4499 (+ module scratch TI:35-43: qtot, tend_ru_physics,
 * tend_rtheta_physics, tend_rho_physics, delsq_theta, delsq_w, delsq_divergence, delsq_u,
 * delsq_vorticity, dpdz).  config_horiz_mixing: 1 = '2d_smagorinsky', 2 = '2d_fixed', 0 = other.
 * Arrays only used under the never-defined CURVATURE macro (latCell, latEdge, ur_cell, vr_cell)
 * and the unused rr, pressure_b, exner, scalars, deriv_two, cellsOnCell, qv_init, tend_w_pgf,
 * tend_w_buoy may be NULL.  The *_physics arrays may be NULL (treated as zero: physics off). */
#define MPAS_DYNTEND_ARGS                                                                          \
    int nCells, int nEdges, int nVertices, int nVertLevels, int nCellsSolve, int nEdgesSolve,      \
        int vertexDegree, int maxEdges, int maxEdges2, int num_scalars, int moist_start,           \
        int moist_end, const mpas_real *fEdge, const mpas_real *dvEdge, const mpas_real *dcEdge,   \
        const mpas_real *invDcEdge, const mpas_real *invDvEdge, const mpas_real *invAreaCell,      \
        const mpas_real *invAreaTriangle, const mpas_real *meshScalingDel2,                        \
        const mpas_real *meshScalingDel4, const mpas_real *weightsOnEdge, const mpas_real *zgrid,  \
        const mpas_real *rho_edge, const mpas_real *rho_zz, const mpas_real *ru,                   \
        const mpas_real *u, const mpas_real *v, mpas_real *tend_u, const mpas_real *divergence,    \
        const mpas_real *vorticity, const mpas_real *ke, const mpas_real *pv_edge,                 \
        const mpas_real *theta_m, const mpas_real *rw, mpas_real *tend_rho,                        \
        const mpas_real *rt_diabatic_tend, mpas_real *tend_theta, mpas_real *tend_w,               \
        const mpas_real *w, const mpas_real *cqw, const mpas_real *rb, const mpas_real *rr,        \
        const mpas_real *pp, const mpas_real *pressure_b, const mpas_real *zz,                     \
        const mpas_real *zxu, const mpas_real *cqu, mpas_real *h_divergence, mpas_real *kdiff,     \
        const mpas_real *edgesOnCell_sign, const mpas_real *edgesOnVertex_sign,                    \
        const mpas_real *rw_save, const mpas_real *ru_save, const mpas_real *theta_m_save,         \
        const mpas_real *exner, const mpas_real *rr_save, const mpas_real *scalars,                \
        mpas_real *tend_u_euler, mpas_real *tend_w_euler, mpas_real *tend_theta_euler,             \
        const mpas_real *deriv_two, const int *cellsOnEdge, const int *verticesOnEdge,             \
        const int *edgesOnCell, const int *edgesOnEdge, const int *cellsOnCell,                    \
        const int *edgesOnVertex, const int *nEdgesOnCell, const int *nEdgesOnEdge,                \
        const mpas_real *latCell, const mpas_real *latEdge, const mpas_real *angleEdge,            \
        const mpas_real *u_init, const mpas_real *v_init, const int *advCellsForEdge,              \
        const int *nAdvCellsForEdge, const mpas_real *adv_coefs, const mpas_real *adv_coefs_3rd,   \
        const mpas_real *rdzu, const mpas_real *rdzw, const mpas_real *fzm, const mpas_real *fzp,  \
        const mpas_real *qv_init, const mpas_real *t_init, mpas_real cf1, mpas_real cf2,           \
        mpas_real cf3, mpas_real r_earth, const mpas_real *ur_cell, const mpas_real *vr_cell,      \
        const mpas_real *defc_a, const mpas_real *defc_b, mpas_real *tend_w_pgf,                   \
        mpas_real *tend_w_buoy, mpas_real coef_3rd_order, mpas_real c_s, int config_mix_full,      \
        int config_horiz_mixing, mpas_real config_del4u_div_factor,                                \
        mpas_real config_h_mom_eddy_visc2, mpas_real config_v_mom_eddy_visc2,                      \
        mpas_real config_h_theta_eddy_visc2, mpas_real config_v_theta_eddy_visc2,                  \
        mpas_real config_h_theta_eddy_visc4, mpas_real config_h_mom_eddy_visc4,                    \
        mpas_real config_visc4_2dsmag, mpas_real config_len_disp, int rk_step, mpas_real dt,       \
        mpas_real config_mpas_cam_coef, int config_rayleigh_damp_u,                                \
        mpas_real config_rayleigh_damp_u_timescale_days, int config_number_rayleigh_damp_u_levels, \
        mpas_real *tend_rtheta_adv, mpas_real *rthdynten, MPAS_RANGE_ARGS, const mpas_real *qtot,  \
        const mpas_real *tend_ru_physics, const mpas_real *tend_rtheta_physics,                    \
        const mpas_real *tend_rho_physics, mpas_real *delsq_theta, mpas_real *delsq_w,             \
        mpas_real *delsq_divergence, mpas_real *delsq_u, mpas_real *delsq_vorticity,               \
        mpas_real *dpdz

/* atm_advance_scalars_work, TI:3238-3248 (optional scalar_tend / rho_zz_int are unused by the
 * routine body and dropped; kdiff, deriv_two, meshScaling*, rho_edge, qv_init, zgrid unused -> may be NULL) */
#define MPAS_ADVSCALARS_ARGS                                                                       \
    int num_scalars, int nCells, int nVertLevels, int maxEdges, mpas_real dt, MPAS_RANGE_ARGS,     \
        mpas_real coef_3rd_order, const mpas_real *scalar_old, mpas_real *scalar_new,              \
        const mpas_real *rho_zz_old, const mpas_real *rho_zz_new, const mpas_real *kdiff,          \
        const mpas_real *uhAvg, const mpas_real *wwAvg, const mpas_real *deriv_two,                \
        const mpas_real *dvEdge, const int *cellsOnEdge, const int *edgesOnCell,                   \
        const int *nEdgesOnCell, const mpas_real *edgesOnCell_sign, const mpas_real *invAreaCell,  \
        mpas_real *scalar_tend_save, const mpas_real *fnm, const mpas_real *fnp,                   \
        const mpas_real *rdnw, const mpas_real *meshScalingDel2, const mpas_real *meshScalingDel4, \
        const int *bdyMaskCell, const int *bdyMaskEdge, const int *nAdvCellsForEdge,               \
        const int *advCellsForEdge, const mpas_real *adv_coefs, const mpas_real *adv_coefs_3rd,    \
        const mpas_real *rho_edge, const mpas_real *qv_init, const mpas_real *zgrid,               \
        int nCellsSolve, int nEdges, mpas_real *horiz_flux_arr, int rk_step,                       \
        int config_time_integration_order, int advance_density

/* Halo-exchange hook for the two exchanges embedded in atm_advance_scalars_mono_work
 * (TI:3757 scalars_old, all halo layers; TI:4098 scale_arr, halo layer 1).  which: 0 = scalars_old
 * (num_scalars, nVertLevels, nCells+1), 1 = scale_arr (nVertLevels, 2, nCells+1) of the per-scalar path,
 * 2 = the scale factors of ALL scalars at once, a library workspace array (2*num_scalars*nVertLevels, nCells+1)
 * passed as `array` (default all-scalars path, halo layer 1).  NULL = single block. */
typedef int (*mpas_halo_fn)(void *ctx, int which, mpas_real *array);

/* atm_advance_scalars_mono_work, TI:3612-3621 (block/state replaced by the exchange hook) */
#define MPAS_MONO_ARGS                                                                             \
    mpas_halo_fn halo_fn, void *halo_ctx, int nCells, int nEdges, int nVertLevels, int maxEdges,   \
        mpas_real dt, MPAS_RANGE_ARGS, mpas_real coef_3rd_order, int nCellsSolve, int num_scalars, \
        const mpas_real *uhAvg, const mpas_real *wwAvg, mpas_real *scalar_tend,                    \
        const mpas_real *rho_zz_old, const mpas_real *rho_zz_new, mpas_real *scalars_old,          \
        mpas_real *scalars_new, const mpas_real *invAreaCell, const mpas_real *dvEdge,             \
        const int *cellsOnEdge, const int *cellsOnCell, const int *edgesOnCell,                    \
        const mpas_real *edgesOnCell_sign, const int *nEdgesOnCell, const mpas_real *fnm,          \
        const mpas_real *fnp, const mpas_real *rdnw, const int *nAdvCellsForEdge,                  \
        const int *advCellsForEdge, const mpas_real *adv_coefs, const mpas_real *adv_coefs_3rd,    \
        mpas_real *scalar_old, mpas_real *scalar_new, mpas_real *s_max, mpas_real *s_min,          \
        mpas_real *wdtn, mpas_real *scale_arr, mpas_real *flux_arr, mpas_real *flux_upwind_tmp,    \
        mpas_real *flux_tmp, const int *bdyMaskCell, const int *bdyMaskEdge, int advance_density,  \
        mpas_real *rho_zz_int

/* atm_rk_integration_setup, TI:1799-1859 (pool arrays resolved as at TI:1826-1845) */
#define MPAS_RKSETUP_ARGS                                                                          \
    int nVertLevels, int num_scalars, int nCells, int nEdges, MPAS_RANGE_ARGS,                     \
        const mpas_real *ru, mpas_real *ru_save, const mpas_real *rw, mpas_real *rw_save,          \
        const mpas_real *rtheta_p, mpas_real *rtheta_p_save, const mpas_real *rho_p,               \
        mpas_real *rho_p_save, mpas_real *rho_zz_old_split, const mpas_real *u_1, mpas_real *u_2,  \
        const mpas_real *w_1, mpas_real *w_2, const mpas_real *theta_m_1, mpas_real *theta_m_2,    \
        const mpas_real *rho_zz_1, mpas_real *rho_zz_2, const mpas_real *scalars_1,                \
        mpas_real *scalars_2

/* atm_compute_moist_coefficients, TI:1862-1933 (+ module scratch qtot) */
#define MPAS_MOIST_ARGS                                                                            \
    int nVertLevels, int num_scalars, int nCells, int nEdges, int nCellsSolve, int moist_start,    \
        int moist_end, MPAS_RANGE_ARGS, const int *cellsOnEdge, const mpas_real *scalars,          \
        mpas_real *cqw, mpas_real *cqu, mpas_real *qtot

/* atm_compute_output_diagnostics, src/core_atmosphere/mpas_atm_core.F:753-802 (pool arguments resolved to arrays) */
#define MPAS_OUTPUTDIAG_ARGS                                                                       \
    int nVertLevels, int num_scalars, int index_qv, int nCells, const mpas_real *theta_m,          \
        const mpas_real *rho_zz, const mpas_real *scalars, const mpas_real *zz,                    \
        const mpas_real *pressure_base, const mpas_real *pressure_p, mpas_real *theta,             \
        mpas_real *rho, mpas_real *pressure

/* atm_rk_dynamics_substep_finish, TI:5993-6081 */
#define MPAS_SUBSTEPFINISH_ARGS                                                                    \
    int nVertLevels, int nCells, int nEdges, int dynamics_substep, int dynamics_split,             \
        MPAS_RANGE_ARGS, const mpas_real *ru, mpas_real *ru_save, const mpas_real *rw,             \
        mpas_real *rw_save, const mpas_real *rtheta_p, mpas_real *rtheta_p_save,                   \
        const mpas_real *rho_p, mpas_real *rho_p_save, const mpas_real *rho_zz_old_split,          \
        mpas_real *ruAvg, mpas_real *ruAvg_split, mpas_real *wwAvg, mpas_real *wwAvg_split,        \
        mpas_real *u_1, const mpas_real *u_2, mpas_real *w_1, const mpas_real *w_2,                \
        mpas_real *theta_m_1, const mpas_real *theta_m_2, mpas_real *rho_zz_1,                     \
        const mpas_real *rho_zz_2

/* atm_init_coupled_diagnostics, TI:5825-5990 (pool arrays resolved as at TI:5874-5903) */
#define MPAS_INITCOUPLED_ARGS                                                                      \
    int nVertLevels, int maxEdges, int num_scalars, int index_qv, int nCells, int nEdges,          \
        MPAS_RANGE_ARGS, const int *cellsOnEdge, const int *nEdgesOnCell, const int *edgesOnCell,  \
        const mpas_real *edgesOnCell_sign, mpas_real *theta_m, const mpas_real *theta,             \
        mpas_real *rho_zz, const mpas_real *rho, mpas_real *rho_p, const mpas_real *rho_base,      \
        mpas_real *rtheta_base, const mpas_real *theta_base, mpas_real *rtheta_p,                  \
        const mpas_real *zz, const mpas_real *scalars, mpas_real *ru, mpas_real *rw,               \
        const mpas_real *u, const mpas_real *w, mpas_real *pressure_p, mpas_real *pressure_base,   \
        mpas_real *exner, mpas_real *exner_base, const mpas_real *fzm, const mpas_real *fzp,       \
        const mpas_real *zb_cell, const mpas_real *zb3_cell

/* mpas_reconstruct_2d, operators/mpas_vector_reconstruction.F:195-305 (called once per dt at TI:1594 with u time level 2;
 * pool arrays resolved as at :229-243).  nCellsLoop = nCellsSolve, or nCells when includeHalos is given.  coeffs_reconstruct
 * (3, maxEdges, nCells+1) comes from mpas_init_reconstruct (:51-176). */
#define MPAS_RECONSTRUCT_ARGS                                                                      \
    int nVertLevels, int maxEdges, int nCells, int nEdges, int nCellsLoop, int on_a_sphere,        \
        const mpas_real *u, const mpas_real *coeffs_reconstruct, const int *nEdgesOnCell,          \
        const int *edgesOnCell, const mpas_real *latCell, const mpas_real *lonCell,                \
        mpas_real *uReconstructX, mpas_real *uReconstructY, mpas_real *uReconstructZ,              \
        mpas_real *uReconstructZonal, mpas_real *uReconstructMeridional

#endif
