/* Host harness of the B200 dycore library (libmpas_host.so): icosahedral Voronoi mesh generator, METIS / file partition,
 * block decomposition with the reference's local ordering and exchange lists, Jablonowski-Williamson initial state,
 * classic-netCDF files.  No CUDA in here: the CPU oracle / the --impl reference arm of bench.py load this library alone;
 * libmpas_b200.so (include/mpas_b200.h) takes the blocks it builds (mpas_b200_dycore_attach_block).
 * Reference counterparts: src/framework/mpas_block_decomp.F:51-506, mpas_block_creator.F:52-938, mpas_dmpar.F:1483-2038,
 * src/core_init_atmosphere/mpas_init_atm_cases.F:367-1310, mpas_atm_advection.F, src/core_atmosphere/mpas_atm_core.F:927-1288. */
#ifndef MPAS_HOST_H
#define MPAS_HOST_H
#ifdef __cplusplus
extern "C" {
#endif

const char* mpas_host_last_error(void);
int mpas_host_set_num_threads(int n); /* OpenMP threads of the harness; n < 1 = all processors; returns the count set */

/* ---- Layer 0: host harness (mesh, partition, JW initial state) ---- */
void* mpas_host_case_create(int level, int lloyd_iters, int sfc);
/* doubly periodic planar mesh of nx x ny regular hexagons, spacing dc (m); build blocks with init_case 4 (squall line) / 5 (supercell),
 * src/core_init_atmosphere/mpas_init_atm_cases.F:1313-1895 */
void* mpas_host_case_create_planar(int nx, int ny, double dc, int sfc);
void mpas_host_case_destroy(void* h);
int mpas_host_case_dims(void* h, int* nCells, int* nEdges, int* nVertices, int* maxEdges, int* maxEdges2);
int mpas_host_case_permutation(void* h, int kind, int* out);
int mpas_host_case_partition(void* h, int nBlocks, int method, const char* prefix);
int mpas_host_case_get_partition(void* h, int* out);
int mpas_host_case_build_block(void* h, int iBlock, int nVertLevels, int init_case, int num_scalars, int full_fields, int keep_zb);
void* mpas_host_case_block(void* h, int iBlock);
int mpas_host_block_nfields(void* b);
int mpas_host_block_field_info(void* b, int idx, char* name, int name_cap, int* kind, int* loc, int* ndims, long long* dims);
void* mpas_host_block_field_ptr(void* b, const char* name);
/* lean blocks (mpas_host_case_build_block full_fields = 2): the work arrays that are all zero when a block has been built
 * get no host storage until mpas_host_block_field_ptr is asked for them; 1 = no storage yet, 0 = has storage, -1 = no such field */
int mpas_host_block_field_is_lean(void* block, const char* name);
int mpas_host_block_ncfg(void* b);
int mpas_host_block_cfg_info(void* b, int idx, char* name, int name_cap, double* value);
int mpas_host_block_nexch(void* b, int kind, int which);
int mpas_host_block_exch_info(void* b, int kind, int which, int idx, int* endPoint, int* haloLayer, int* nlist);
int mpas_host_block_exch_lists(void* b, int kind, int which, int idx, int* srcList, int* destList);

/* ---- on-disk files: MPAS grid / restart files in the netCDF classic formats (CDF-1/2/5), the formats
 * of the reference's streams (Registry.xml:397-1125, mpas_bootstrapping.F:79-423); no netCDF library is used ---- */
int mpas_host_case_write_grid(void* h, const char* path, int version);
void* mpas_host_case_create_from_grid(const char* path, int sfc);
int mpas_host_case_create_restart(void* h, int iBlock, const char* path, int version, const char* names_csv, const char* xtime);
int mpas_host_block_write_restart(void* b, const char* path, const char* names_csv);
int mpas_host_block_read_restart(void* b, const char* path);
void* mpas_nc_open(const char* path);
void* mpas_nc_create(const char* path, int version);
int mpas_nc_close(void* h);
int mpas_nc_inq(void* h, int* version, int* ndims, int* nvars, int* ngatts, long long* numrecs);
int mpas_nc_dim_info(void* h, int idx, char* name, int name_cap, long long* len);
int mpas_nc_var_info(void* h, int idx, char* name, int name_cap, int* type, int* ndims, int* dimids, long long* count);
int mpas_nc_gatt_text(void* h, const char* name, char* out, int cap);
int mpas_nc_get_double(void* h, const char* name, double* out);
int mpas_nc_get_int(void* h, const char* name, int* out);
int mpas_nc_def_dim(void* h, const char* name, long long len);
int mpas_nc_def_var(void* h, const char* name, int type, int ndims, const int* dimids);
int mpas_nc_put_gatt_text(void* h, const char* name, const char* value);
int mpas_nc_enddef(void* h, long long numrecs);
int mpas_nc_put_double(void* h, int varid, long long rec, long long start, long long count, const double* data);
int mpas_nc_put_int(void* h, int varid, long long rec, long long start, long long count, const int* data);

#ifdef __cplusplus
}
#endif
#endif
