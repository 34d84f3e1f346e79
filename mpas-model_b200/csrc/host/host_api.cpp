// C ABI of the host-side harness (mesh generation, decomposition, initial conditions).
// Declared in include/mpas_b200.h.  Handles are opaque pointers; errors return non-zero /
// NULL and leave a message retrievable through mpas_b200_last_error().
#include <omp.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>

#include "decomp.h"
#include "init_atm.h"
#include "mesh.h"
#include "mpas_io.h"
#include "nc_classic.h"
#include "pool.h"

using namespace mpas;

namespace mpas {
std::string g_last_error;
}

struct HostCase {
    Mesh global;                 // global mesh (SFC order if requested), unit sphere scaled to radius a
    MeshPermutation perm;        // SFC permutation (new -> generator order)
    bool has_perm = false;
    InitConfig cfg;
    std::vector<int> part;       // partition (0-based block id per global cell)
    int nBlocks = 0;
    std::vector<std::unique_ptr<Block>> blocks;
};

#define API_TRY try {
#define API_CATCH(ret)                      \
    }                                       \
    catch (const std::exception& e) {       \
        mpas::g_last_error = e.what();      \
        return ret;                         \
    }

extern "C" {

// OpenMP threads of the harness (mesh generation, initial state); n < 1 = all processors.  torch.distributed.run exports
// OMP_NUM_THREADS=1 to its workers, which would make the set-up of a large mesh single-threaded.
int mpas_host_set_num_threads(int n) {
    if (n < 1) n = omp_get_num_procs();
    omp_set_num_threads(n);
    return n;
}
const char* mpas_host_last_error() { return mpas::g_last_error.c_str(); }

// Create a global icosahedral mesh (unit sphere -> radius a), optionally SFC-reordered.
void* mpas_host_case_create(int level, int lloyd_iters, int sfc) {
    API_TRY
    auto* hc = new HostCase();
    Mesh gen;
    make_icosahedral_mesh(level, gen, lloyd_iters);
    if (sfc) {
        sfc_reorder(gen, hc->global, hc->perm);
        hc->has_perm = true;
    } else {
        hc->global = gen;
    }
    scale_mesh_to_sphere(hc->global, Constants::a);
    return hc;
    API_CATCH(nullptr)
}

// Doubly periodic planar mesh of nx x ny regular hexagons with spacing dc (m): the idealised squall-line / supercell cases
// (BASELINE.json configs[2]); build blocks with init_case 4 (squall line) or 5 (supercell).
void* mpas_host_case_create_planar(int nx, int ny, double dc, int sfc) {
    API_TRY
    auto* hc = new HostCase();
    Mesh gen;
    make_planar_hex_mesh(nx, ny, dc, gen);
    if (sfc) {
        sfc_reorder(gen, hc->global, hc->perm);
        hc->has_perm = true;
    } else {
        hc->global = gen;
    }
    return hc;
    API_CATCH(nullptr)
}

void mpas_host_case_destroy(void* h) { delete (HostCase*)h; }

int mpas_host_case_dims(void* h, int* nCells, int* nEdges, int* nVertices, int* maxEdges, int* maxEdges2) {
    auto* hc = (HostCase*)h;
    *nCells = hc->global.nCells; *nEdges = hc->global.nEdges; *nVertices = hc->global.nVertices;
    *maxEdges = hc->global.maxEdges; *maxEdges2 = hc->global.maxEdges2;
    return 0;
}

// SFC permutation: kind 0 cells, 1 edges, 2 vertices; out[new] = old (0-based)
int mpas_host_case_permutation(void* h, int kind, int* out) {
    auto* hc = (HostCase*)h;
    if (!hc->has_perm) { mpas::g_last_error = "case has no SFC permutation"; return 1; }
    const std::vector<int>& v = kind == 0 ? hc->perm.cellNewToOld : kind == 1 ? hc->perm.edgeNewToOld : hc->perm.vertexNewToOld;
    std::memcpy(out, v.data(), v.size() * sizeof(int));
    return 0;
}

// Partition into nBlocks: method 0 = METIS k-way (libmetis), 1 = contiguous SFC chunks,
// 2 = read "<prefix><nBlocks>" partition file (reference format).  Writes graph.info and the
// partition file when prefix is non-empty.
int mpas_host_case_partition(void* h, int nBlocks, int method, const char* prefix) {
    API_TRY
    auto* hc = (HostCase*)h;
    std::string pre = prefix ? prefix : "";
    hc->nBlocks = nBlocks;
    if (nBlocks == 1) {
        hc->part.assign(hc->global.nCells, 0);
    } else if (method == 2) {
        if (!read_partition(pre + std::to_string(nBlocks), hc->global.nCells, hc->part))
            throw std::runtime_error("partition file missing or wrong number of lines");
    } else {
        partition_cells(hc->global, nBlocks, method, hc->part);
    }
    if (!pre.empty() && method != 2) {
        write_graph_info(hc->global, pre.substr(0, pre.size() - 6));  // "graph.info.part." -> "graph.info"
        write_partition(hc->part, pre + std::to_string(nBlocks));
    }
    return 0;
    API_CATCH(1)
}

int mpas_host_case_get_partition(void* h, int* out) {
    auto* hc = (HostCase*)h;
    std::memcpy(out, hc->part.data(), hc->part.size() * sizeof(int));
    return 0;
}

// Build block `iBlock` (owned + 2 cell halo layers / 3 edge+vertex layers, reference ordering and
// exchange lists) and initialise its pool with the JW case.  Other blocks need not exist in this
// process (one process per GPU builds only its own block).
int mpas_host_case_build_block(void* h, int iBlock, int nVertLevels, int init_case, int num_scalars, int full_fields, int keep_zb) {
    API_TRY
    auto* hc = (HostCase*)h;
    if (hc->nBlocks == 0) { hc->nBlocks = 1; hc->part.assign(hc->global.nCells, 0); }
    if ((int)hc->blocks.size() < hc->nBlocks) hc->blocks.resize(hc->nBlocks);
    hc->cfg.nVertLevels = nVertLevels;
    hc->cfg.init_case = init_case;
    hc->cfg.num_scalars = num_scalars;
    hc->cfg.keep_zb = keep_zb != 0;
    auto blk = std::make_unique<Block>();
    const bool timing = getenv("MPAS_HOST_TIMING") != nullptr;  // per-phase wall times of the block set-up on stderr
    auto t0 = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "mpas_host_case_build_block: %-32s %.2f s\n", what, std::chrono::duration<double>(t1 - t0).count());
        t0 = t1;
    };
    build_block(hc->global, hc->part, hc->nBlocks, iBlock, 2, *blk);
    lap("build_block (decomposition)");
    Pool& p = blk->pool;
    mesh_to_pool(blk->mesh, p);
    lap("mesh_to_pool");
    init_advection_and_deformation(blk->mesh, p);
    lap("init_advection_and_deformation");
    if (init_case == 4 || init_case == 5) {
        if (blk->mesh.on_a_sphere) throw std::runtime_error("init_case 4 / 5 (squall line / supercell) need a planar mesh (mpas_host_case_create_planar)");
        hc->cfg.config_zd = 12000.0;  // no upper damping layer in the idealised convective cases (mpas_init_atm_cases.F:1483-1484)
        hc->cfg.config_xnutr = 0.0;
        init_case_squall_line(blk->mesh, hc->cfg, p);
    } else {
        if (!blk->mesh.on_a_sphere) throw std::runtime_error("the Jablonowski-Williamson cases need a spherical mesh");
        init_case_jw(blk->mesh, hc->cfg, p);
    }
    lap("init_case");
    core_init_derived(blk->mesh, hc->cfg, p);
    lap("core_init_derived");
    init_reconstruct(blk->mesh, p);
    lap("init_reconstruct");
    allocate_dycore_fields(blk->mesh, hc->cfg, p, full_fields);
    p.cfg["config_coef_3rd_order"] = hc->cfg.coef_3rd_order;
    hc->blocks[iBlock] = std::move(blk);
    return 0;
    API_CATCH(1)
}

void* mpas_host_case_block(void* h, int iBlock) {
    auto* hc = (HostCase*)h;
    if (iBlock < 0 || iBlock >= (int)hc->blocks.size() || !hc->blocks[iBlock]) return nullptr;
    return hc->blocks[iBlock].get();
}

// ---- pool access -----------------------------------------------------------------------
int mpas_host_block_nfields(void* b) { return (int)((Block*)b)->pool.f.size(); }

int mpas_host_block_field_info(void* b, int idx, char* name, int name_cap, int* kind, int* loc, int* ndims, long long* dims) {
    Pool& p = ((Block*)b)->pool;
    auto it = p.f.begin();
    std::advance(it, idx);
    std::strncpy(name, it->first.c_str(), name_cap - 1);
    name[name_cap - 1] = 0;
    *kind = it->second.kind; *loc = it->second.loc; *ndims = (int)it->second.dims.size();
    for (size_t k = 0; k < it->second.dims.size(); ++k) dims[k] = it->second.dims[k];
    return 0;
}

void* mpas_host_block_field_ptr(void* b, const char* name) {
    Pool& p = ((Block*)b)->pool;
    auto it = p.f.find(name);
    if (it == p.f.end()) { mpas::g_last_error = std::string("no field ") + name; return nullptr; }
    it->second.materialize();  // a lean block's zero field gets its host storage when somebody looks at it
    return it->second.kind == FK_REAL ? (void*)it->second.r.data() : (void*)it->second.i.data();
}

// 1: the field is an all-zero real field of a lean block that has no host storage (yet); 0 otherwise; -1 unknown name
int mpas_host_block_field_is_lean(void* b, const char* name) {
    Pool& p = ((Block*)b)->pool;
    auto it = p.f.find(name);
    if (it == p.f.end()) return -1;
    return (it->second.kind == FK_REAL && it->second.zero && it->second.r.empty()) ? 1 : 0;
}

int mpas_host_block_ncfg(void* b) { return (int)((Block*)b)->pool.cfg.size(); }
int mpas_host_block_cfg_info(void* b, int idx, char* name, int name_cap, double* value) {
    Pool& p = ((Block*)b)->pool;
    auto it = p.cfg.begin();
    std::advance(it, idx);
    std::strncpy(name, it->first.c_str(), name_cap - 1);
    name[name_cap - 1] = 0;
    *value = it->second;
    return 0;
}

// ---- exchange lists (reference form, src/framework/mpas_dmpar.F:1483-2038) -----------------
// kind: 0 cells, 1 edges, 2 vertices; which: 0 send, 1 recv.  Returns the number of
// (neighbour, layer) lists; each list is queried by index.
int mpas_host_block_nexch(void* b, int kind, int which) {
    Block* blk = (Block*)b;
    const ExchangeSet& s = kind == 0 ? blk->cellExch : kind == 1 ? blk->edgeExch : blk->vertexExch;
    return (int)(which == 0 ? s.send.size() : s.recv.size());
}
int mpas_host_block_exch_info(void* b, int kind, int which, int idx, int* endPoint, int* haloLayer, int* nlist) {
    Block* blk = (Block*)b;
    const ExchangeSet& s = kind == 0 ? blk->cellExch : kind == 1 ? blk->edgeExch : blk->vertexExch;
    const ExchangeList& l = which == 0 ? s.send[idx] : s.recv[idx];
    *endPoint = l.endPointID; *haloLayer = l.haloLayer; *nlist = (int)l.srcList.size();
    return 0;
}
int mpas_host_block_exch_lists(void* b, int kind, int which, int idx, int* srcList, int* destList) {
    Block* blk = (Block*)b;
    const ExchangeSet& s = kind == 0 ? blk->cellExch : kind == 1 ? blk->edgeExch : blk->vertexExch;
    const ExchangeList& l = which == 0 ? s.send[idx] : s.recv[idx];
    std::memcpy(srcList, l.srcList.data(), l.srcList.size() * sizeof(int));
    std::memcpy(destList, l.destList.data(), l.destList.size() * sizeof(int));
    return 0;
}


// ---- on-disk files: grid.nc, restart, generic classic-netCDF access -------------------------

static std::vector<std::string> split_csv(const char* s) {
    std::vector<std::string> out;
    std::string cur;
    for (const char* p = s; *p; ++p) {
        if (*p == ',') { if (!cur.empty()) out.push_back(cur); cur.clear(); }
        else cur.push_back(*p);
    }
    if (!cur.empty()) out.push_back(cur);
    return out;
}

// the global mesh of the case as an MPAS grid file (version 1, 2 or 5 = CDF-1 / CDF-2 / CDF-5)
int mpas_host_case_write_grid(void* h, const char* path, int version) {
    API_TRY
    write_grid_nc(((HostCase*)h)->global, path, version);
    return 0;
    API_CATCH(1)
}

// a case from an MPAS grid file instead of the mesh generator; a unit-sphere file is scaled to the Earth radius as
// init_atm_case_jw does (mpas_init_atm_cases.F:497-510)
void* mpas_host_case_create_from_grid(const char* path, int sfc) {
    API_TRY
    auto* hc = new HostCase();
    Mesh in;
    read_grid_nc(path, in);
    if (sfc) {
        sfc_reorder(in, hc->global, hc->perm);
        hc->has_perm = true;
    } else {
        hc->global = in;
    }
    if (hc->global.sphere_radius == 1.0) scale_mesh_to_sphere(hc->global, Constants::a);
    return hc;
    API_CATCH(nullptr)
}

// restart file: header from the global dimensions and the fields of block iBlock (names: comma-separated pool names)
int mpas_host_case_create_restart(void* h, int iBlock, const char* path, int version, const char* names_csv, const char* xtime) {
    API_TRY
    auto* hc = (HostCase*)h;
    if (iBlock < 0 || iBlock >= (int)hc->blocks.size() || !hc->blocks[iBlock]) throw std::runtime_error("create_restart: block not built");
    create_restart_nc(path, version, hc->global, *hc->blocks[iBlock], split_csv(names_csv), xtime ? xtime : "");
    return 0;
    API_CATCH(1)
}
int mpas_host_block_write_restart(void* b, const char* path, const char* names_csv) {
    API_TRY
    write_restart_block(path, *(Block*)b, split_csv(names_csv));
    return 0;
    API_CATCH(1)
}
int mpas_host_block_read_restart(void* b, const char* path) {  // number of pool fields filled, -1 on error
    API_TRY
    return read_restart_block(path, *(Block*)b);
    API_CATCH(-1)
}

void* mpas_nc_open(const char* path) {
    API_TRY
    auto* f = new nc::File();
    try { f->open(path); } catch (...) { delete f; throw; }
    return f;
    API_CATCH(nullptr)
}
void* mpas_nc_create(const char* path, int version) {
    API_TRY
    auto* f = new nc::File();
    try { f->create(path, version); } catch (...) { delete f; throw; }
    return f;
    API_CATCH(nullptr)
}
int mpas_nc_close(void* h) {
    API_TRY
    auto* f = (nc::File*)h;
    try { f->close(); } catch (...) { delete f; throw; }
    delete f;
    return 0;
    API_CATCH(1)
}
int mpas_nc_inq(void* h, int* version, int* ndims, int* nvars, int* ngatts, long long* numrecs) {
    auto* f = (nc::File*)h;
    *version = f->version; *ndims = (int)f->dims.size(); *nvars = (int)f->vars.size(); *ngatts = (int)f->gatts.size(); *numrecs = f->numrecs;
    return 0;
}
int mpas_nc_dim_info(void* h, int idx, char* name, int name_cap, long long* len) {
    auto* f = (nc::File*)h;
    if (idx < 0 || idx >= (int)f->dims.size()) return 1;
    std::strncpy(name, f->dims[idx].name.c_str(), name_cap - 1);
    name[name_cap - 1] = 0;
    *len = f->dims[idx].len;
    return 0;
}
int mpas_nc_var_info(void* h, int idx, char* name, int name_cap, int* type, int* ndims, int* dimids, long long* count) {
    auto* f = (nc::File*)h;
    if (idx < 0 || idx >= (int)f->vars.size()) return 1;
    const nc::Var& v = f->vars[idx];
    std::strncpy(name, v.name.c_str(), name_cap - 1);
    name[name_cap - 1] = 0;
    *type = v.type; *ndims = (int)v.dimids.size();
    for (size_t i = 0; i < v.dimids.size() && i < 8; ++i) dimids[i] = v.dimids[i];
    *count = f->var_count(idx);
    return 0;
}
int mpas_nc_gatt_text(void* h, const char* name, char* out, int cap) {
    auto* f = (nc::File*)h;
    const nc::Att* a = f->find_gatt(name);
    if (!a) return 1;
    std::string t = a->type == nc::NC_CHAR ? a->text() : std::to_string(a->number());
    std::strncpy(out, t.c_str(), cap - 1);
    out[cap - 1] = 0;
    return 0;
}
int mpas_nc_get_double(void* h, const char* name, double* out) {
    API_TRY
    auto* f = (nc::File*)h;
    const int id = f->find_var(name);
    if (id < 0) throw std::runtime_error(std::string("no variable ") + name);
    f->get_double(id, out);
    return 0;
    API_CATCH(1)
}
int mpas_nc_get_int(void* h, const char* name, int* out) {
    API_TRY
    auto* f = (nc::File*)h;
    const int id = f->find_var(name);
    if (id < 0) throw std::runtime_error(std::string("no variable ") + name);
    f->get_int(id, out);
    return 0;
    API_CATCH(1)
}
int mpas_nc_def_dim(void* h, const char* name, long long len) {
    API_TRY
    return ((nc::File*)h)->def_dim(name, len);
    API_CATCH(-1)
}
int mpas_nc_def_var(void* h, const char* name, int type, int ndims, const int* dimids) {
    API_TRY
    return ((nc::File*)h)->def_var(name, type, std::vector<int>(dimids, dimids + ndims));
    API_CATCH(-1)
}
int mpas_nc_put_gatt_text(void* h, const char* name, const char* value) {
    API_TRY
    ((nc::File*)h)->put_gatt_text(name, value);
    return 0;
    API_CATCH(1)
}
int mpas_nc_enddef(void* h, long long numrecs) {
    API_TRY
    ((nc::File*)h)->enddef(numrecs);
    return 0;
    API_CATCH(1)
}
int mpas_nc_put_double(void* h, int varid, long long rec, long long start, long long count, const double* data) {
    API_TRY
    ((nc::File*)h)->put_double(varid, rec, start, count, data);
    return 0;
    API_CATCH(1)
}
int mpas_nc_put_int(void* h, int varid, long long rec, long long start, long long count, const int* data) {
    API_TRY
    ((nc::File*)h)->put_int(varid, rec, start, count, data);
    return 0;
    API_CATCH(1)
}
}  // extern "C"
