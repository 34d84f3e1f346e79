// Host-side field container ("pool") for one mesh block.
//
// Mirrors the role of the reference's mpas_pool_type (src/framework/mpas_pool_routines.F:4282
// mpas_pool_get_array): every array the dycore touches is looked up by its Registry name
// (src/core_atmosphere/Registry.xml:1127 mesh, :1412 state, :1506 diag, :1709 tend).
// Arrays keep the reference layout: Fortran column-major, fastest index first,
// one extra trailing "garbage" element (nCells+1, nEdges+1, nVertices+1) where the
// reference has it (src/framework/mpas_block_creator.F:1445-1508).  Index arrays are
// 1-based exactly as in the reference.
#pragma once
#include <cstdint>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace mpas {

typedef double real_t;  // RKIND, double build (Makefile:221 -fdefault-real-8)

enum FieldKind { FK_REAL = 0, FK_INT = 1 };
// which mesh entity the LAST dimension runs over (for halo exchange / decomposition)
enum FieldLoc { LOC_NONE = 0, LOC_CELL = 1, LOC_EDGE = 2, LOC_VERTEX = 3 };

struct Field {
    std::string name;
    FieldKind kind = FK_REAL;
    FieldLoc loc = LOC_NONE;
    std::vector<int64_t> dims;  // Fortran order, fastest first; last dim includes the +1 element
    std::vector<real_t> r;
    std::vector<int32_t> i;
    // zero = true: a real field whose every element is 0 and whose host storage has not been allocated (the scratch /
    // tendency / work arrays of a "lean" block: 83 % of a block's words).  The device mirror is created zero-filled from
    // the metadata alone; materialize() gives it host storage when somebody wants to look at it.
    bool zero = false;
    void materialize() {
        if (zero && r.empty()) r.assign((size_t)count(), 0.0);
        zero = false;
    }
    int64_t count() const {
        int64_t n = 1;
        for (auto d : dims) n *= d;
        return n;
    }
    int64_t inner() const {  // product of all dims except the last
        int64_t n = 1;
        for (size_t k = 0; k + 1 < dims.size(); ++k) n *= dims[k];
        return n;
    }
};

struct Pool {
    std::map<std::string, Field> f;
    std::map<std::string, double> cfg;  // namelist scalars + dimensions (as doubles)

    Field& addr(const std::string& name, FieldLoc loc, std::vector<int64_t> dims) {
        Field& fl = f[name];
        fl.name = name;
        fl.kind = FK_REAL;
        fl.loc = loc;
        fl.dims = dims;
        fl.r.assign((size_t)fl.count(), 0.0);
        return fl;
    }
    Field& addz(const std::string& name, FieldLoc loc, std::vector<int64_t> dims) {  // all-zero real field without host storage
        Field& fl = f[name];
        fl.name = name;
        fl.kind = FK_REAL;
        fl.loc = loc;
        fl.dims = dims;
        fl.r.clear();
        fl.r.shrink_to_fit();
        fl.zero = true;
        return fl;
    }
    Field& addi(const std::string& name, FieldLoc loc, std::vector<int64_t> dims) {
        Field& fl = f[name];
        fl.name = name;
        fl.kind = FK_INT;
        fl.loc = loc;
        fl.dims = dims;
        fl.i.assign((size_t)fl.count(), 0);
        return fl;
    }
    bool has(const std::string& name) const { return f.count(name) != 0; }
    Field& get(const std::string& name) {
        auto it = f.find(name);
        if (it == f.end()) throw std::runtime_error("pool: no field '" + name + "'");
        return it->second;
    }
    real_t* R(const std::string& name) {
        Field& fl = get(name);
        fl.materialize();
        return fl.r.data();
    }
    int32_t* I(const std::string& name) { return get(name).i.data(); }
    double c(const std::string& name) const {
        auto it = cfg.find(name);
        if (it == cfg.end()) throw std::runtime_error("pool: no config '" + name + "'");
        return it->second;
    }
    int ci(const std::string& name) const { return (int)c(name); }
};

}  // namespace mpas
