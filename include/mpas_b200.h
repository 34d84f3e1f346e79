/* C ABI of libmpas_b200.so -- the B200-native drop-in for the MPAS-Atmosphere SRK3 dycore step.
 *
 * Layer 1 (drop-in work routines): one entry point per reference *_work routine, same dummy-argument
 * order (argument lists in mpas_b200_args.h, each citing TI = src/core_atmosphere/dynamics/
 * mpas_atm_time_integration.F file:line).  Arrays may be DEVICE pointers, or HOST addresses inside a
 * range registered with mpas_b200_register_field (the Fortran host keeps passing the pool arrays it
 * always passed; the library maps them to device mirrors).  All calls are asynchronous on the
 * library stream; they return 0 or a non-zero code with a message in mpas_b200_last_error()
 * (the reference's error convention is mpas_log_write(..., MPAS_LOG_CRIT), TI:123-124).
 *
 * Layer 2 (device-resident dycore, the atm_srk3 call sequence TI:142-1796 for the harness, the
 * tests and bench.py): mpas_b200_dycore_*.  Halo exchanges (mpas_dmpar_exch_halo_field,
 * src/framework/mpas_dmpar.F:5386-5748) run as pack kernel -> NCCL grouped send/recv (or
 * device copies between blocks of one process) -> unpack kernel.
 *
 * Layer 0 (host harness: mesh, decomposition, initial conditions) lives in its own library without CUDA:
 * libmpas_host.so, include/mpas_host.h.
 */
#ifndef MPAS_B200_H
#define MPAS_B200_H
#include <stddef.h>

#include "mpas_b200_args.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- lifecycle / registry (no reference analogue) ---- */
const char* mpas_b200_last_error(void);
int mpas_b200_set_device(int device);
int mpas_b200_device_count(void);
int mpas_b200_register_field(const void* host, size_t bytes);
int mpas_b200_unregister_field(const void* host);
int mpas_b200_upload(const void* host);
int mpas_b200_download(void* host);
void* mpas_b200_device_ptr(const void* host);
int mpas_b200_synchronize(void);
long long mpas_b200_launch_count(void);
/* sizeof(mpas_real) of this build: 8 (default) or 4 (PRECISION=single, libmpas_b200_sp.so built with -DSINGLE_PRECISION;
 * the reference's RKIND switch, src/framework/mpas_kind_types.F:25-29) */
int mpas_b200_real_size(void);
int mpas_b200_timer_start(void);    /* CUDA-event stopwatch on the library stream */
double mpas_b200_timer_stop(void);  /* elapsed ms since timer_start, after synchronising */
int mpas_b200_ktime_enable(int on);                /* per-kernel CUDA-event timing of eager launches */
int mpas_b200_ktime_report(char* buf, int cap);    /* lines "kernel ms calls"; returns needed size if cap is too small */
int mpas_b200_finalize(void);

/* ---- Layer 1: work routines ---- */
int mpas_b200_advance_acoustic_step_work(MPAS_ACOUSTIC_ARGS);              /* TI:2447 */
int mpas_b200_divergence_damping_3d(MPAS_DIVDAMP_ARGS);                    /* TI:2726 */
int mpas_b200_recover_large_step_variables_work(MPAS_RECOVER_ARGS);        /* TI:2909 */
int mpas_b200_compute_dyn_tend_work(MPAS_DYNTEND_ARGS);                    /* TI:4481 */
int mpas_b200_advance_scalars_mono_work(MPAS_MONO_ARGS);                   /* TI:3612 */
int mpas_b200_advance_scalars_work(MPAS_ADVSCALARS_ARGS);                  /* TI:3238 */
int mpas_b200_compute_solve_diagnostics_work(MPAS_SOLVEDIAG_ARGS);         /* TI:5513 */
int mpas_b200_compute_vert_imp_coefs_work(MPAS_VERTIMP_ARGS);              /* TI:2012 */
int mpas_b200_set_smlstep_pert_variables_work(MPAS_SMLSTEP_ARGS);          /* TI:2224 */
int mpas_b200_rk_integration_setup(MPAS_RKSETUP_ARGS);                     /* TI:1799 */
int mpas_b200_compute_moist_coefficients(MPAS_MOIST_ARGS);                 /* TI:1862 */
int mpas_b200_rk_dynamics_substep_finish(MPAS_SUBSTEPFINISH_ARGS);         /* TI:5993 */
int mpas_b200_init_coupled_diagnostics(MPAS_INITCOUPLED_ARGS);             /* TI:5825 */
int mpas_b200_compute_output_diagnostics(MPAS_OUTPUTDIAG_ARGS);           /* mpas_atm_core.F:753 */
int mpas_b200_reconstruct_2d(MPAS_RECONSTRUCT_ARGS);                       /* mpas_vector_reconstruction.F:195, called TI:1594 */

/* ---- Layer 1: halo exchange for a host that keeps its own blocks and exchange lists (the bodies of
 * mpas_dmpar_exch_halo_field2d_real / 3d_real, src/framework/mpas_dmpar.F:5386-5552, 5565-5748).
 * kind: 1 cell, 2 edge, 3 vertex (the decomposed, LAST dimension); which: 0 send lists, 1 recv lists;
 * layer_mask: bit (l-1) selects halo layer l, 0 = all layers (the optional haloLayersIn of the reference).
 * Lists are the host's own, as mpas_dmpar_get_exch_list built them (mpas_dmpar.F:1483-2038): per list endPointID (rank),
 * halo layer, nList and the 1-based srcList / destList (all lists of a side concatenated).  endPoint == own rank is the
 * reference's copyList.  Arrays: registered host addresses or device pointers, extent includes the +1 garbage column. ---- */
int mpas_b200_halo_comm_init(int rank, int nranks, const void* id128);  /* id from mpas_b200_nccl_unique_id, broadcast by the host */
int mpas_b200_halo_register_lists(int kind, int which, int nLists, const int* endPoint, const int* haloLayer, const int* nList, const int* srcList,
                                  const int* destList);
int mpas_b200_exch_halo_field1d(mpas_real* array, int kind, int layer_mask);                    /* mpas_dmpar.F:5253 */
int mpas_b200_exch_halo_field2d(mpas_real* array, int dim1, int kind, int layer_mask);          /* mpas_dmpar.F:5386 */
int mpas_b200_exch_halo_field3d(mpas_real* array, int dim1, int dim2, int kind, int layer_mask); /* mpas_dmpar.F:5565 */
long mpas_b200_halo_exchange_count(void);
/* the halo_fn of mpas_b200_advance_scalars_mono_work for such a host (TI:3757, 4098): ctx -> int[2] = {nVertLevels, num_scalars} */
int mpas_b200_mono_halo_default(void* ctx, int which, mpas_real* array);
int mpas_b200_halo_finalize(void);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it).  One same-process halo copy (the copyList leg of the
 * exchange above) that allocates nothing: array(dim1, nElems+1); pack  buffer(:, sendDest(i)) = array(:, sendSrc(i)),
 * then unpack  array(:, recvDest(i)) = buffer(:, recvSrc(i)),  i = 1..n, all four lists 1-based and device-resident (device
 * pointers or registered host addresses); sendDest / recvSrc are buffer slots in 1..n.  The message buffer is a fixed
 * device array inside the library: n * dim1 <= MPAS_B200_NO_ALLOC_MAX_REALS or the call fails with a clear error.  All
 * calls share that buffer: calls that overlap on the GPU race (calls on the library stream are ordered and safe). */
#define MPAS_B200_NO_ALLOC_MAX_REALS (1u << 20)
int mpas_b200_halo_copy_no_alloc(mpas_real* array, int dim1, int n, const int* sendSrc, const int* sendDest, const int* recvSrc,
                                 const int* recvDest);

/* ---- Layer 2: device-resident dycore ---- */
void* mpas_b200_dycore_create(int nLocalBlocks, int rank, int nranks);
void mpas_b200_dycore_destroy(void* h);
/* upload every pool field, config value and exchange list of a host-harness block (mpas_host_case_block) */
int mpas_b200_dycore_attach_block(void* h, int iLocal, void* hostBlock);
int mpas_b200_dycore_set_cfg(void* h, int iLocal, const char* name, double value);
double mpas_b200_dycore_get_cfg(void* h, int iLocal, const char* name);
/* NCCL: rank 0 fills id (128 bytes) with mpas_b200_nccl_unique_id, the host broadcasts it, every rank calls comm_init */
int mpas_b200_nccl_unique_id(void* id128);
int mpas_b200_dycore_comm_init(void* h, const void* id128);
int mpas_b200_dycore_init(void* h);             /* atm_mpas_init_block + exchanges, mpas_atm_core.F:145-186, 311-402 */
int mpas_b200_dycore_step(void* h, int nsteps); /* atm_srk3 + time-level shift, TI:142, mpas_atm_core.F:671 */
int mpas_b200_dycore_get_field(void* h, int iLocal, const char* name, void* host_out, size_t bytes);
int mpas_b200_dycore_set_field(void* h, int iLocal, const char* name, const void* host_in, size_t bytes);
void* mpas_b200_dycore_field_device_ptr(void* h, int iLocal, const char* name);
int mpas_b200_dycore_exchange(void* h, const char* name, int layer /* 0 = all */);
/* run ONE work routine on the model's current fields (tests): name = "acoustic", "divdamp", "recover",
 * "dyn_tend", "mono", "advance_scalars", "solve_diagnostics", "vert_imp_coefs", "smlstep", "rk_setup",
 * "moist", "substep_finish"; params as documented in dycore.cu */
int mpas_b200_dycore_call(void* h, const char* routine, const double* params, int nparams);
long mpas_b200_dycore_exchange_count(void* h);
/* how halo messages travel between the ranks: 0 one rank (no messages), 1 NCCL grouped send/recv (the replacement of the
 * MPI_Irecv/Isend pairs of mpas_dmpar.F:5430-5452, 5466-5476), 2 peer-to-peer stores into the neighbours' receive windows
 * (CUDA IPC over NVLink, switched on after the second step when every rank can map every other); -2 a peer-to-peer wait
 * timed out since the last call.  MPAS_B200_TUNE_p2p=0 keeps NCCL. */
int mpas_b200_dycore_halo_transport(void* h);
int mpas_b200_dycore_set_option(void* h, const char* name, int value); /* "timing", "graph", "aggregate"; "diet" (before attach_block); "fold_first" (before the first step) */
/* memory diet (default on; option "diet" = 0 before attach_block gives every pool array its own mirror): arrays only physics or
 * output use (gradPVn/t, ke_edge, tend_*_physics, rthdynten, tend_rtheta_adv, rt_diabatic_tend: optional arguments of the work
 * routines) get no mirror, and phase-local scratch (the delsq_* / dpdz arrays of TI:592-605, the mono scratch of TI:1029-1050,
 * horiz_flux_arr) plus the library workspace of the same phases are views of ONE arena per block, as the reference allocates and
 * frees them around each phase.  1 = `name` is such a view (valid only inside its phase), 0 = own storage, -1 = no mirror */
int mpas_b200_dycore_field_is_scratch_view(void* h, int iLocal, const char* name);
long long mpas_b200_dycore_device_bytes(void* h); /* bytes of device memory held by the blocks' mirrors and arenas */
int mpas_b200_dycore_timer_count(void* h);
int mpas_b200_dycore_timer_get(void* h, int idx, char* name, int cap, double* ms, long* calls);
int mpas_b200_dycore_timer_reset(void* h);
double mpas_b200_dycore_global_max_abs(void* h, const char* name); /* summarize_timestep analogue, TI:6675 */
/* global min / max over owned elements of every stride-th element from offset (one scalar of scalars(:,k,iCell)): the
 * reductions summarize_timestep prints for w, u and the scalars, TI:6758-7010 (mpas_dmpar_min_real / max_real) */
int mpas_b200_dycore_global_minmax(void* h, const char* name, long long offset, long long stride, double* min_out, double* max_out);
/* sum (mod 2^64) of the bit patterns of this process's OWNED elements of a decomposed real field; added over the ranks it is
 * the same number for every partition of the mesh (owned-element arithmetic is partition independent) */
int mpas_b200_dycore_checksum(void* h, const char* name, unsigned long long* out);

#ifdef __cplusplus
}
#endif
#endif
