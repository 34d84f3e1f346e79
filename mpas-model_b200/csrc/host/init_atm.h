// Harness inputs the reference defines in core_init_atmosphere and core_atmosphere init
// (SURVEY.md Appendix A): vertical grid, Jablonowski-Williamson state, deriv_two, defc_a/b,
// advection coefficient compression, signs, inverses, mesh scaling, damping coefficients.
#pragma once
#include "mesh.h"
#include "pool.h"

namespace mpas {

struct Constants {
    static constexpr double pii = 3.141592653589793;  // src/framework/mpas_constants.F:25
    static constexpr double a = 6371229.0;            // :26
    static constexpr double omega = 7.29212e-5;       // :27
    static constexpr double gravity = 9.80616;        // :28
    static constexpr double rgas = 287.0;             // :29
    static constexpr double rv = 461.6;               // :30
    static constexpr double rvord = rv / rgas;        // :31
    static constexpr double cp = 7. * rgas / 2.;      // :33
    static constexpr double cv = cp - rgas;           // :34
    static constexpr double prandtl = 1.0;            // :36
};

struct InitConfig {
    int nVertLevels = 26;
    int init_case = 2;          // 1 steady JW, 2 JW + perturbation, 3 normal-mode; 4 squall line, 5 supercell (planar mesh)
    int num_scalars = 1;        // qv + passive tracers
    int theta_adv_order = 3;
    double coef_3rd_order = 0.25;
    double config_zd = 22000.0, config_xnutr = 0.2;  // Registry.xml nhyd_model defaults
    bool h_ScaleWithMesh = true;
    bool keep_zb = false;       // also store edge-based zb, zb3 (small meshes / tests)
};

// multiply all lengths/areas of a unit-sphere mesh by radius (mpas_init_atm_cases.F:497-510)
void scale_mesh_to_sphere(Mesh& m, double radius);

// copy the mesh (connectivity + metrics) into the pool under Registry names
void mesh_to_pool(const Mesh& m, Pool& p);

// atm_initialize_advection_rk (mpas_atm_advection.F:21-394) + atm_initialize_deformation_weights (:744-937)
void init_advection_and_deformation(const Mesh& m, Pool& p);

// init_atm_case_jw (mpas_init_atm_cases.F:367-1160) on all local cells/edges of m
void init_case_jw(const Mesh& m, const InitConfig& cfg, Pool& p);

// init_atm_case_squall_line (mpas_init_atm_cases.F:1313-1895): test case 4 squall line, 5 supercell, planar periodic mesh
void init_case_squall_line(const Mesh& m, const InitConfig& cfg, Pool& p);

// core-init derived arrays (mpas_atm_core.F:311-360, 456-463, 927-1288)
void core_init_derived(const Mesh& m, const InitConfig& cfg, Pool& p);

// mpas_initialize_vectors + mpas_init_reconstruct (mpas_vector_operations.F:653, mpas_vector_reconstruction.F:51):
// edgeNormalVectors, cellTangentPlane, localVerticalUnitVectors, coeffs_reconstruct
void init_reconstruct(const Mesh& m, Pool& p);

// allocate the remaining state/diag/tend fields the dycore uses (zero filled)
void allocate_dycore_fields(const Mesh& m, const InitConfig& cfg, Pool& p, int full);  // 0 none, 1 with host storage, 2 lean (Field::zero)

}  // namespace mpas
