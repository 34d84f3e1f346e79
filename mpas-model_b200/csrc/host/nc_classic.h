// Minimal reader / writer of the netCDF CLASSIC file formats (CDF-1, CDF-2 "64-bit offset", CDF-5 "64-bit data") — the
// formats MPAS streams are written in through PIO/pnetcdf (io_type "pnetcdf" = CDF-2, "pnetcdf,cdf5" = CDF-5,
// src/framework/mpas_io.F; stream definitions src/core_atmosphere/Registry.xml:397-1125).  No netCDF library exists in this
// image, so the on-disk layout is restated from the published classic-format specification: big-endian header
// (magic, numrecs, dim_list, gatt_list, var_list) followed by the fixed-size variables and then the records of the
// variables whose first dimension is the unlimited one.
#pragma once
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

namespace mpas {
namespace nc {

enum Type { NC_BYTE = 1, NC_CHAR = 2, NC_SHORT = 3, NC_INT = 4, NC_FLOAT = 5, NC_DOUBLE = 6, NC_UBYTE = 7, NC_USHORT = 8, NC_UINT = 9, NC_INT64 = 10, NC_UINT64 = 11 };
int type_size(int t);

struct Dim {
    std::string name;
    int64_t len = 0;  // 0 = the unlimited (record) dimension
};
struct Att {
    std::string name;
    int type = NC_CHAR;
    int64_t nelems = 0;
    std::vector<unsigned char> data;  // host byte order
    std::string text() const { return std::string(data.begin(), data.end()); }
    double number(int64_t i = 0) const;
};
struct Var {
    std::string name;
    std::vector<int> dimids;
    std::vector<Att> atts;
    int type = NC_DOUBLE;
    int64_t vsize = 0, begin = 0;
    bool record = false;
    int64_t count_per_record = 0;  // elements in one record (or in the whole variable when !record)
};

struct File {
    int version = 2;
    int64_t numrecs = 0;
    std::vector<Dim> dims;
    std::vector<Att> gatts;
    std::vector<Var> vars;
    int64_t recsize = 0;
    std::FILE* fp = nullptr;
    bool writing = false, defined = false;
    std::string path;

    ~File();
    // ---- reading
    void open(const std::string& path);  // throws std::runtime_error
    int find_var(const std::string& name) const;
    int find_dim(const std::string& name) const;
    const Att* find_gatt(const std::string& name) const;
    int64_t var_count(int varid) const;  // all records
    // whole variable (all records), converted to the requested host type
    void get_double(int varid, double* out) const;
    void get_int(int varid, int* out) const;
    // ---- writing: define everything, enddef(), then put data
    void create(const std::string& path, int version);
    int def_dim(const std::string& name, int64_t len);
    int def_var(const std::string& name, int type, const std::vector<int>& dimids);
    void put_gatt_text(const std::string& name, const std::string& value);
    void put_gatt_double(const std::string& name, double v);
    void put_att_text(int varid, const std::string& name, const std::string& value);
    void enddef(int64_t numrecs);
    void open_for_update(const std::string& path);  // existing file, header already written (parallel writers)
    // `count` elements starting at element `start` of record `rec` (rec ignored for fixed variables)
    void put_double(int varid, int64_t rec, int64_t start, int64_t count, const double* data);
    void put_int(int varid, int64_t rec, int64_t start, int64_t count, const int* data);
    void get_double_range(int varid, int64_t rec, int64_t start, int64_t count, double* out) const;
    void close();
};

}  // namespace nc
}  // namespace mpas
