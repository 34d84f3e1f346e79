// MPAS on-disk files on top of nc_classic: the horizontal mesh (`grid.nc`: the 'input' stream's mesh variables,
// src/core_atmosphere/Registry.xml:397-560, read by src/framework/mpas_bootstrapping.F:79-423) and the restart state
// (the 'restart' stream, Registry.xml restart var lists; the reference writes it through PIO with one decomposition per
// dimension pair, src/framework/mpas_io.F).  Global files are in GLOBAL element order with C-order dimensions
// (Time, nCells, nVertLevels) = Fortran (k fastest) columns; a block writes / reads the columns of its own elements at
// `indexTo{Cell,Edge,Vertex}ID` — what PIO's decomposed write does — so several processes can fill one file.
#include <cmath>
#include <cstring>
#include <stdexcept>

#include "decomp.h"
#include "init_atm.h"
#include "mesh.h"
#include "mpas_io.h"
#include "nc_classic.h"

namespace mpas {

namespace {
void put_all_double(nc::File& f, int id, const std::vector<double>& v, int64_t n) { f.put_double(id, 0, 0, n, v.data()); }
void put_all_int(nc::File& f, int id, const std::vector<int>& v, int64_t n) { f.put_int(id, 0, 0, n, v.data()); }
}  // namespace

// ---------------------------------------------------------------------------------------------- grid.nc
void write_grid_nc(const Mesh& m, const std::string& path, int version) {
    nc::File f;
    f.create(path, version);
    const int dC = f.def_dim("nCells", m.nCells), dE = f.def_dim("nEdges", m.nEdges), dV = f.def_dim("nVertices", m.nVertices);
    const int dM = f.def_dim("maxEdges", m.maxEdges), dM2 = f.def_dim("maxEdges2", m.maxEdges2), d2 = f.def_dim("TWO", 2);
    const int d3 = f.def_dim("vertexDegree", 3);
    f.def_dim("Time", 0);
    f.put_gatt_text("on_a_sphere", m.on_a_sphere ? "YES" : "NO");
    f.put_gatt_double("sphere_radius", m.sphere_radius);
    f.put_gatt_text("is_periodic", "NO");
    f.put_gatt_text("Conventions", "MPAS");
    f.put_gatt_text("source", "libmpas_b200 host harness");
    struct RV { const char* name; const std::vector<double>* v; std::vector<int> dims; int64_t n; };
    struct IV { const char* name; const std::vector<int>* v; std::vector<int> dims; int64_t n; };
    const int64_t nC = m.nCells, nE = m.nEdges, nV = m.nVertices;
    std::vector<RV> rv = {
        {"latCell", &m.latCell, {dC}, nC}, {"lonCell", &m.lonCell, {dC}, nC}, {"xCell", &m.xCell, {dC}, nC}, {"yCell", &m.yCell, {dC}, nC},
        {"zCell", &m.zCell, {dC}, nC}, {"latEdge", &m.latEdge, {dE}, nE}, {"lonEdge", &m.lonEdge, {dE}, nE}, {"xEdge", &m.xEdge, {dE}, nE},
        {"yEdge", &m.yEdge, {dE}, nE}, {"zEdge", &m.zEdge, {dE}, nE}, {"latVertex", &m.latVertex, {dV}, nV}, {"lonVertex", &m.lonVertex, {dV}, nV},
        {"xVertex", &m.xVertex, {dV}, nV}, {"yVertex", &m.yVertex, {dV}, nV}, {"zVertex", &m.zVertex, {dV}, nV},
        {"dvEdge", &m.dvEdge, {dE}, nE}, {"dcEdge", &m.dcEdge, {dE}, nE}, {"angleEdge", &m.angleEdge, {dE}, nE}, {"areaCell", &m.areaCell, {dC}, nC},
        {"areaTriangle", &m.areaTriangle, {dV}, nV}, {"meshDensity", &m.meshDensity, {dC}, nC},
        {"kiteAreasOnVertex", &m.kiteAreasOnVertex, {dV, d3}, nV * 3}, {"weightsOnEdge", &m.weightsOnEdge, {dE, dM2}, nE * m.maxEdges2}};
    std::vector<IV> iv = {
        {"indexToCellID", &m.indexToCellID, {dC}, nC}, {"indexToEdgeID", &m.indexToEdgeID, {dE}, nE}, {"indexToVertexID", &m.indexToVertexID, {dV}, nV},
        {"nEdgesOnCell", &m.nEdgesOnCell, {dC}, nC}, {"nEdgesOnEdge", &m.nEdgesOnEdge, {dE}, nE},
        {"cellsOnCell", &m.cellsOnCell, {dC, dM}, nC * m.maxEdges}, {"edgesOnCell", &m.edgesOnCell, {dC, dM}, nC * m.maxEdges},
        {"verticesOnCell", &m.verticesOnCell, {dC, dM}, nC * m.maxEdges}, {"cellsOnEdge", &m.cellsOnEdge, {dE, d2}, nE * 2},
        {"verticesOnEdge", &m.verticesOnEdge, {dE, d2}, nE * 2}, {"edgesOnEdge", &m.edgesOnEdge, {dE, dM2}, nE * m.maxEdges2},
        {"cellsOnVertex", &m.cellsOnVertex, {dV, d3}, nV * 3}, {"edgesOnVertex", &m.edgesOnVertex, {dV, d3}, nV * 3}};
    std::vector<int> rid, iid;
    for (auto& v : rv) rid.push_back(f.def_var(v.name, nc::NC_DOUBLE, v.dims));
    for (auto& v : iv) iid.push_back(f.def_var(v.name, nc::NC_INT, v.dims));
    f.enddef(0);
    for (size_t i = 0; i < rv.size(); ++i) put_all_double(f, rid[i], *rv[i].v, rv[i].n);
    for (size_t i = 0; i < iv.size(); ++i) put_all_int(f, iid[i], *iv[i].v, iv[i].n);
    f.close();
}

void read_grid_nc(const std::string& path, Mesh& m) {
    nc::File f;
    f.open(path);
    auto dim = [&](const char* n) -> int64_t {
        const int id = f.find_dim(n);
        if (id < 0) throw std::runtime_error(std::string("grid file has no dimension ") + n);
        if (f.dims[id].len > 2147483647LL) throw std::runtime_error(std::string("grid file: dimension ") + n + " exceeds the 32-bit index range of the model");
        return f.dims[id].len;
    };
    const int nC = (int)dim("nCells"), nE = (int)dim("nEdges"), nV = (int)dim("nVertices");
    m.maxEdges = (int)dim("maxEdges");
    m.maxEdges2 = (int)dim("maxEdges2");
    if (dim("vertexDegree") != 3) throw std::runtime_error("grid file: vertexDegree must be 3");
    m.alloc(nC, nE, nV);
    m.nCellsSolve = nC; m.nEdgesSolve = nE; m.nVerticesSolve = nV;
    m.nCellsHalo = {nC}; m.nEdgesHalo = {nE}; m.nVerticesHalo = {nV};
    const nc::Att* a = f.find_gatt("on_a_sphere");
    m.on_a_sphere = !a || a->text().substr(0, 3) == "YES";
    a = f.find_gatt("sphere_radius");
    m.sphere_radius = a ? a->number() : 1.0;
    auto R = [&](const char* n, std::vector<double>& v, int64_t cnt, bool required = true) {
        const int id = f.find_var(n);
        if (id < 0) {
            if (required) throw std::runtime_error(std::string("grid file has no variable ") + n);
            return false;
        }
        if (f.var_count(id) != cnt) throw std::runtime_error(std::string("grid file: unexpected size of ") + n);
        f.get_double(id, v.data());
        return true;
    };
    // connectivity: a non-positive entry (unused slot in files written by the MPAS mesh tools) becomes the garbage element
    auto I = [&](const char* n, std::vector<int>& v, int64_t cnt, int garbage, bool required = true) {
        const int id = f.find_var(n);
        if (id < 0) {
            if (required) throw std::runtime_error(std::string("grid file has no variable ") + n);
            return false;
        }
        if (f.var_count(id) != cnt) throw std::runtime_error(std::string("grid file: unexpected size of ") + n);
        f.get_int(id, v.data());
        if (garbage > 0)
            for (int64_t i = 0; i < cnt; ++i)
                if (v[i] <= 0) v[i] = garbage;
        return true;
    };
    R("latCell", m.latCell, nC); R("lonCell", m.lonCell, nC); R("xCell", m.xCell, nC); R("yCell", m.yCell, nC); R("zCell", m.zCell, nC);
    R("latEdge", m.latEdge, nE); R("lonEdge", m.lonEdge, nE); R("xEdge", m.xEdge, nE); R("yEdge", m.yEdge, nE); R("zEdge", m.zEdge, nE);
    R("latVertex", m.latVertex, nV); R("lonVertex", m.lonVertex, nV); R("xVertex", m.xVertex, nV); R("yVertex", m.yVertex, nV); R("zVertex", m.zVertex, nV);
    R("dvEdge", m.dvEdge, nE); R("dcEdge", m.dcEdge, nE); R("angleEdge", m.angleEdge, nE); R("areaCell", m.areaCell, nC);
    R("areaTriangle", m.areaTriangle, nV); R("kiteAreasOnVertex", m.kiteAreasOnVertex, (int64_t)nV * 3);
    R("weightsOnEdge", m.weightsOnEdge, (int64_t)nE * m.maxEdges2);
    if (!R("meshDensity", m.meshDensity, nC, false)) m.meshDensity.assign((size_t)nC + 1, 1.0);
    if (!I("indexToCellID", m.indexToCellID, nC, 0, false)) for (int i = 0; i < nC; ++i) m.indexToCellID[i] = i + 1;
    if (!I("indexToEdgeID", m.indexToEdgeID, nE, 0, false)) for (int i = 0; i < nE; ++i) m.indexToEdgeID[i] = i + 1;
    if (!I("indexToVertexID", m.indexToVertexID, nV, 0, false)) for (int i = 0; i < nV; ++i) m.indexToVertexID[i] = i + 1;
    I("nEdgesOnCell", m.nEdgesOnCell, nC, 0); I("nEdgesOnEdge", m.nEdgesOnEdge, nE, 0);
    I("cellsOnCell", m.cellsOnCell, (int64_t)nC * m.maxEdges, nC + 1); I("edgesOnCell", m.edgesOnCell, (int64_t)nC * m.maxEdges, nE + 1);
    I("verticesOnCell", m.verticesOnCell, (int64_t)nC * m.maxEdges, nV + 1); I("cellsOnEdge", m.cellsOnEdge, (int64_t)nE * 2, nC + 1);
    I("verticesOnEdge", m.verticesOnEdge, (int64_t)nE * 2, nV + 1); I("edgesOnEdge", m.edgesOnEdge, (int64_t)nE * m.maxEdges2, nE + 1);
    I("cellsOnVertex", m.cellsOnVertex, (int64_t)nV * 3, nC + 1); I("edgesOnVertex", m.edgesOnVertex, (int64_t)nV * 3, nE + 1);
    // garbage element, as the mesh generator leaves it (mesh.cpp)
    m.nEdgesOnCell[nC] = 0;
    m.cellsOnEdge[2 * (size_t)nE] = nC + 1; m.cellsOnEdge[2 * (size_t)nE + 1] = nC + 1;
    m.verticesOnEdge[2 * (size_t)nE] = nV + 1; m.verticesOnEdge[2 * (size_t)nE + 1] = nV + 1;
    for (int s = 0; s < 3; ++s) {
        m.cellsOnVertex[3 * (size_t)nV + s] = nC + 1;
        m.edgesOnVertex[3 * (size_t)nV + s] = nE + 1;
    }
    f.close();
}

// ---------------------------------------------------------------------------------------------- restart
// Variable naming: pool name without the "@1" time-level suffix; scalars(num_scalars, k, iCell) is written as one variable
// per constituent (the reference's var_array convention): qv for index_qv, tracer<N> for the others.
namespace {
std::string restart_var_name(const std::string& pool_name) {
    const size_t at = pool_name.find('@');
    return at == std::string::npos ? pool_name : pool_name.substr(0, at);
}
std::string scalar_name(int iq, int index_qv) { return iq == index_qv ? std::string("qv") : "tracer" + std::to_string(iq); }
const std::vector<int>& ids_of(const Mesh& m, FieldLoc loc) {
    return loc == LOC_CELL ? m.indexToCellID : loc == LOC_EDGE ? m.indexToEdgeID : m.indexToVertexID;
}
int local_count(const Mesh& m, FieldLoc loc) { return loc == LOC_CELL ? m.nCells : loc == LOC_EDGE ? m.nEdges : m.nVertices; }
int owned_count(const Mesh& m, FieldLoc loc) { return loc == LOC_CELL ? m.nCellsSolve : loc == LOC_EDGE ? m.nEdgesSolve : m.nVerticesSolve; }
}  // namespace

// Header of a restart file for the global mesh dimensions; `names` are pool names of real fields of `proto` (any block).
void create_restart_nc(const std::string& path, int version, const Mesh& global, const Block& proto, const std::vector<std::string>& names,
                       const std::string& xtime) {
    nc::File f;
    f.create(path, version);
    const Pool& p = proto.pool;
    const int L = (int)p.cfg.at("nVertLevels");
    const int dT = f.def_dim("Time", 0);
    const int dC = f.def_dim("nCells", global.nCells), dE = f.def_dim("nEdges", global.nEdges), dV = f.def_dim("nVertices", global.nVertices);
    const int dL = f.def_dim("nVertLevels", L), dL1 = f.def_dim("nVertLevelsP1", L + 1), dS = f.def_dim("StrLen", 64);
    f.put_gatt_text("Conventions", "MPAS");
    f.put_gatt_text("core_name", "atmosphere");
    f.put_gatt_text("on_a_sphere", global.on_a_sphere ? "YES" : "NO");
    f.put_gatt_double("sphere_radius", global.sphere_radius);
    f.put_gatt_double("config_dt", p.cfg.count("config_dt") ? p.cfg.at("config_dt") : 0.0);
    f.def_var("xtime", nc::NC_CHAR, {dT, dS});
    const int index_qv = p.cfg.count("index_qv") ? (int)p.cfg.at("index_qv") : 1;
    for (const std::string& n : names) {
        const Field& fl = p.f.at(n);
        if (fl.kind != FK_REAL || fl.loc == LOC_NONE) throw std::runtime_error("restart: " + n + " is not a real field on cells, edges or vertices");
        const int dH = fl.loc == LOC_CELL ? dC : fl.loc == LOC_EDGE ? dE : dV;
        if (fl.dims.size() == 3) {  // (num_scalars, nVertLevels, n)
            if (fl.dims[1] != L) throw std::runtime_error("restart: unexpected shape of " + n);
            for (int iq = 1; iq <= (int)fl.dims[0]; ++iq) f.def_var(scalar_name(iq, index_qv), nc::NC_DOUBLE, {dT, dH, dL});
        } else if (fl.dims.size() == 2 && (fl.dims[0] == L || fl.dims[0] == L + 1)) {
            f.def_var(restart_var_name(n), nc::NC_DOUBLE, {dT, dH, fl.dims[0] == L ? dL : dL1});
        } else if (fl.dims.size() == 1) {
            f.def_var(restart_var_name(n), nc::NC_DOUBLE, {dT, dH});
        } else {
            throw std::runtime_error("restart: unsupported shape of " + n);
        }
    }
    f.enddef(1);
    std::vector<int> t(64, ' ');
    for (size_t i = 0; i < xtime.size() && i < 64; ++i) t[i] = (unsigned char)xtime[i];
    f.put_int(f.find_var("xtime"), 0, 0, 64, t.data());
    f.close();
}

// The block's OWNED columns of every field into the (already created) file.
void write_restart_block(const std::string& path, const Block& b, const std::vector<std::string>& names) {
    nc::File f;
    f.open_for_update(path);
    const Pool& p = b.pool;
    const int index_qv = p.cfg.count("index_qv") ? (int)p.cfg.at("index_qv") : 1;
    std::vector<double> col;
    for (const std::string& n : names) {
        const_cast<Field&>(p.f.at(n)).materialize();  // lean block: a zero field gets storage before it is read (contents unchanged)
        const Field& fl = p.f.at(n);
        const std::vector<int>& ids = ids_of(b.mesh, fl.loc);
        const int nown = owned_count(b.mesh, fl.loc);
        if (fl.dims.size() == 3) {
            const int ns = (int)fl.dims[0], L = (int)fl.dims[1];
            col.resize(L);
            for (int iq = 1; iq <= ns; ++iq) {
                const int id = f.find_var(scalar_name(iq, index_qv));
                if (id < 0) throw std::runtime_error("restart file has no variable " + scalar_name(iq, index_qv));
                for (int i = 0; i < nown; ++i) {
                    for (int k = 0; k < L; ++k) col[k] = fl.r[((size_t)i * L + k) * ns + (iq - 1)];
                    f.put_double(id, 0, (int64_t)(ids[i] - 1) * L, L, col.data());
                }
            }
        } else {
            const int64_t inner = fl.inner();
            const int id = f.find_var(restart_var_name(n));
            if (id < 0) throw std::runtime_error("restart file has no variable " + restart_var_name(n));
            // runs of consecutive global IDs go out as one write
            int i = 0;
            while (i < nown) {
                int j = i + 1;
                while (j < nown && ids[j] == ids[j - 1] + 1) ++j;
                f.put_double(id, 0, (int64_t)(ids[i] - 1) * inner, (int64_t)(j - i) * inner, &fl.r[(size_t)i * inner]);
                i = j;
            }
        }
    }
    f.close();
}

// Every variable of the file that names a real pool field (or a scalar constituent) into the block: owned AND halo columns
// (the reference reads the halo from the file as well, mpas_atm_core.F:95).  Returns the number of pool fields filled.
int read_restart_block(const std::string& path, Block& b) {
    nc::File f;
    f.open(path);
    if (f.numrecs < 1) throw std::runtime_error("restart file has no record: " + path);
    Pool& p = b.pool;
    const int index_qv = p.cfg.count("index_qv") ? (int)p.cfg.at("index_qv") : 1;
    // the file must be of THIS mesh and level count: variables are matched by name only, and a file whose sizes merely
    // divide evenly would otherwise be read from the wrong offsets
    {
        auto flen = [&](const char* n) -> int64_t { const int id = f.find_dim(n); return id < 0 ? -1 : f.dims[id].len; };
        auto maxid = [&](FieldLoc loc) { int mx = 0; for (int g : ids_of(b.mesh, loc)) mx = std::max(mx, g); return (int64_t)mx; };
        const struct { const char* dim; FieldLoc loc; } chk[3] = {{"nCells", LOC_CELL}, {"nEdges", LOC_EDGE}, {"nVertices", LOC_VERTEX}};
        for (auto& c : chk) {
            const int64_t n = flen(c.dim);
            if (n >= 0 && n < maxid(c.loc))
                throw std::runtime_error(std::string("restart file ") + path + ": dimension " + c.dim + " = " + std::to_string(n) +
                                         " is smaller than a global ID of this block (" + std::to_string(maxid(c.loc)) + "): file of another mesh");
        }
        const int64_t nl = flen("nVertLevels");
        if (nl >= 0 && p.cfg.count("nVertLevels") && nl != (int64_t)p.cfg.at("nVertLevels"))
            throw std::runtime_error("restart file " + path + ": nVertLevels = " + std::to_string(nl) + " but the block has " +
                                     std::to_string((int64_t)p.cfg.at("nVertLevels")));
    }
    int nfilled = 0;
    std::vector<double> col;
    for (auto& kv : p.f) {
        Field& fl = kv.second;
        if (fl.kind != FK_REAL || fl.loc == LOC_NONE) continue;
        const std::vector<int>& ids = ids_of(b.mesh, fl.loc);
        const int nloc = local_count(b.mesh, fl.loc);
        const size_t at = kv.first.find('@');
        if (at != std::string::npos && kv.first.substr(at) != "@1") continue;  // time level 1 is what a restart holds
        if (fl.dims.size() == 3) {
            const int ns = (int)fl.dims[0], L = (int)fl.dims[1];
            if (restart_var_name(kv.first) != "scalars") continue;
            fl.materialize();
            col.resize(L);
            bool any = false;
            for (int iq = 1; iq <= ns; ++iq) {
                const int id = f.find_var(scalar_name(iq, index_qv));
                if (id < 0) continue;
                any = true;
                for (int i = 0; i < nloc; ++i) {
                    f.get_double_range(id, f.numrecs - 1, (int64_t)(ids[i] - 1) * L, L, col.data());
                    for (int k = 0; k < L; ++k) fl.r[((size_t)i * L + k) * ns + (iq - 1)] = col[k];
                }
            }
            if (any) ++nfilled;
        } else {
            const int id = f.find_var(restart_var_name(kv.first));
            if (id < 0) continue;
            fl.materialize();
            const nc::Var& v = f.vars[id];
            if (!v.record) continue;
            const int64_t inner = fl.inner();
            if (v.count_per_record % inner) throw std::runtime_error("restart: shape mismatch for " + kv.first);
            {
                int mx = 0;
                for (int i = 0; i < nloc; ++i) mx = std::max(mx, ids[i]);
                if ((int64_t)mx * inner > v.count_per_record)
                    throw std::runtime_error("restart: variable " + kv.first + " of " + path + " is too small for this mesh");
            }
            int i = 0;
            while (i < nloc) {
                int j = i + 1;
                while (j < nloc && ids[j] == ids[j - 1] + 1) ++j;
                f.get_double_range(id, f.numrecs - 1, (int64_t)(ids[i] - 1) * inner, (int64_t)(j - i) * inner, &fl.r[(size_t)i * inner]);
                i = j;
            }
            ++nfilled;
        }
    }
    f.close();
    return nfilled;
}

}  // namespace mpas
