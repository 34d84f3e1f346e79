// netCDF classic (CDF-1/2/5) reader and writer; see nc_classic.h.  Layout restated from the classic-format specification:
//   header  = magic numrecs dim_list gatt_list var_list          (all integers big-endian)
//   list    = ABSENT (tag 0, count 0) | tag count [entries]      tags: NC_DIMENSION 0x0A, NC_VARIABLE 0x0B, NC_ATTRIBUTE 0x0C
//   name    = length chars, padded with zeros to a multiple of 4
//   var     = name ndims [dimid] vatt_list nc_type vsize begin   vsize = bytes of one record (or of the variable), padded to 4
//   counts, dimension lengths, dimids and vsize are 4 bytes in CDF-1/2 and 8 bytes in CDF-5; begin is 4 bytes in CDF-1,
//   8 bytes otherwise.  Fixed-size variables follow the header in definition order; then numrecs records, each holding one
//   slab of every record variable (recsize = sum of their vsize; a single record variable is not padded).
#define _FILE_OFFSET_BITS 64
#include "nc_classic.h"

#include <cstring>
#include <stdexcept>

namespace mpas {
namespace nc {

namespace {
const uint32_t TAG_DIM = 0x0A, TAG_VAR = 0x0B, TAG_ATT = 0x0C;

[[noreturn]] void die(const std::string& m) { throw std::runtime_error("netcdf: " + m); }

inline void swap_copy(unsigned char* dst, const unsigned char* src, int64_t n, int sz) {  // host <-> big-endian
    if (sz == 1) { std::memcpy(dst, src, (size_t)n); return; }
    for (int64_t i = 0; i < n; ++i)
        for (int b = 0; b < sz; ++b) dst[i * sz + b] = src[i * sz + (sz - 1 - b)];
}

struct Buf {  // big-endian serialiser
    std::vector<unsigned char> d;
    void u32(uint32_t v) { for (int b = 3; b >= 0; --b) d.push_back((unsigned char)(v >> (8 * b))); }
    void u64(uint64_t v) { for (int b = 7; b >= 0; --b) d.push_back((unsigned char)(v >> (8 * b))); }
    void size(int version, int64_t v) {
        if (version == 5) u64((uint64_t)v);
        else {
            if (v < 0 || v > 0xFFFFFFFFLL) die("value does not fit the 32-bit fields of CDF-1/2; use CDF-5");
            u32((uint32_t)v);
        }
    }
    void pad4() { while (d.size() % 4) d.push_back(0); }
    void name(int version, const std::string& s) {
        size(version, (int64_t)s.size());
        d.insert(d.end(), s.begin(), s.end());
        pad4();
    }
};

struct Reader {
    std::FILE* fp;
    int version = 1;
    uint32_t u32() {
        unsigned char b[4];
        if (std::fread(b, 1, 4, fp) != 4) die("truncated header");
        return ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3];
    }
    uint64_t u64() {
        const uint64_t hi = u32();
        return (hi << 32) | u32();
    }
    int64_t size() { return version == 5 ? (int64_t)u64() : (int64_t)u32(); }
    std::string name() {
        const int64_t n = size();
        if (n < 0 || n > (1 << 20)) die("implausible name length");
        std::string s((size_t)n, '\0');
        if (n && std::fread(&s[0], 1, (size_t)n, fp) != (size_t)n) die("truncated name");
        const int64_t pad = (4 - n % 4) % 4;
        if (pad) std::fseek(fp, (long)pad, SEEK_CUR);
        return s;
    }
    void atts(std::vector<Att>& out) {
        const uint32_t tag = u32();
        const int64_t n = size();
        if (tag == 0 && n == 0) return;
        if (tag != TAG_ATT) die("attribute list expected");
        for (int64_t i = 0; i < n; ++i) {
            Att a;
            a.name = name();
            a.type = (int)u32();
            a.nelems = size();
            const int sz = type_size(a.type);
            const int64_t bytes = a.nelems * sz;
            std::vector<unsigned char> raw((size_t)bytes);
            if (bytes && std::fread(raw.data(), 1, (size_t)bytes, fp) != (size_t)bytes) die("truncated attribute");
            const int64_t pad = (4 - bytes % 4) % 4;
            if (pad) std::fseek(fp, (long)pad, SEEK_CUR);
            a.data.resize((size_t)bytes);
            swap_copy(a.data.data(), raw.data(), a.nelems, sz);
            out.push_back(a);
        }
    }
};

void write_atts(Buf& b, int version, const std::vector<Att>& atts) {
    if (atts.empty()) { b.u32(0); b.size(version, 0); return; }
    b.u32(TAG_ATT);
    b.size(version, (int64_t)atts.size());
    for (const Att& a : atts) {
        b.name(version, a.name);
        b.u32((uint32_t)a.type);
        b.size(version, a.nelems);
        const int sz = type_size(a.type);
        std::vector<unsigned char> be(a.data.size());
        swap_copy(be.data(), a.data.data(), a.nelems, sz);
        b.d.insert(b.d.end(), be.begin(), be.end());
        b.pad4();
    }
}

void serialise_header(const File& f, Buf& b) {
    b.d.push_back('C'); b.d.push_back('D'); b.d.push_back('F'); b.d.push_back((unsigned char)f.version);
    b.size(f.version, f.numrecs);
    if (f.dims.empty()) { b.u32(0); b.size(f.version, 0); }
    else {
        b.u32(TAG_DIM);
        b.size(f.version, (int64_t)f.dims.size());
        for (const Dim& d : f.dims) { b.name(f.version, d.name); b.size(f.version, d.len); }
    }
    write_atts(b, f.version, f.gatts);
    if (f.vars.empty()) { b.u32(0); b.size(f.version, 0); }
    else {
        b.u32(TAG_VAR);
        b.size(f.version, (int64_t)f.vars.size());
        for (const Var& v : f.vars) {
            b.name(f.version, v.name);
            b.size(f.version, (int64_t)v.dimids.size());
            for (int id : v.dimids) b.size(f.version, id);
            write_atts(b, f.version, v.atts);
            b.u32((uint32_t)v.type);
            b.size(f.version, v.vsize);
            if (f.version == 1) b.size(1, v.begin);
            else b.u64((uint64_t)v.begin);
        }
    }
}

void finish_layout(File& f) {  // record flags, element counts, recsize (after the header is known)
    int nrec = 0;
    for (Var& v : f.vars) {
        v.record = !v.dimids.empty() && f.dims.at(v.dimids[0]).len == 0;
        int64_t n = 1;
        for (size_t i = v.record ? 1 : 0; i < v.dimids.size(); ++i) n *= f.dims.at(v.dimids[i]).len;
        v.count_per_record = n;
        if (v.record) ++nrec;
    }
    f.recsize = 0;
    for (Var& v : f.vars)
        if (v.record) f.recsize += (nrec == 1) ? v.count_per_record * type_size(v.type) : v.vsize;
}

int64_t var_offset(const File& f, const Var& v, int64_t rec, int64_t start) {
    return v.begin + (v.record ? rec * f.recsize : 0) + start * type_size(v.type);
}
}  // namespace

int type_size(int t) {
    switch (t) {
        case NC_BYTE: case NC_CHAR: case NC_UBYTE: return 1;
        case NC_SHORT: case NC_USHORT: return 2;
        case NC_INT: case NC_FLOAT: case NC_UINT: return 4;
        case NC_DOUBLE: case NC_INT64: case NC_UINT64: return 8;
    }
    die("unknown nc_type " + std::to_string(t));
}

double Att::number(int64_t i) const {
    const unsigned char* p = data.data() + i * type_size(type);
    switch (type) {
        case NC_DOUBLE: { double v; std::memcpy(&v, p, 8); return v; }
        case NC_FLOAT: { float v; std::memcpy(&v, p, 4); return v; }
        case NC_INT: { int32_t v; std::memcpy(&v, p, 4); return v; }
        case NC_SHORT: { int16_t v; std::memcpy(&v, p, 2); return v; }
        case NC_INT64: { int64_t v; std::memcpy(&v, p, 8); return (double)v; }
        default: return (double)*p;
    }
}

File::~File() {
    if (fp) std::fclose(fp);
}

void File::close() {
    if (fp) {
        if (std::fclose(fp) != 0) { fp = nullptr; die("close failed: " + path); }
        fp = nullptr;
    }
}

// ------------------------------------------------------------------ reading
void File::open(const std::string& p) {
    path = p;
    fp = std::fopen(p.c_str(), "rb");
    if (!fp) die("cannot open " + p);
    unsigned char magic[4];
    if (std::fread(magic, 1, 4, fp) != 4 || magic[0] != 'C' || magic[1] != 'D' || magic[2] != 'F') die("not a netCDF classic file: " + p);
    version = magic[3];
    if (version != 1 && version != 2 && version != 5) die("unsupported netCDF version byte " + std::to_string(version) + " (HDF5-based netCDF-4 is not classic)");
    Reader r{fp, version};
    numrecs = r.size();
    if (version != 5 && numrecs == 0xFFFFFFFFLL) numrecs = -1;  // STREAMING: derive from the file size below
    {
        const uint32_t tag = r.u32();
        const int64_t n = r.size();
        if (!(tag == 0 && n == 0)) {
            if (tag != TAG_DIM) die("dimension list expected");
            for (int64_t i = 0; i < n; ++i) {
                Dim d;
                d.name = r.name();
                d.len = r.size();
                dims.push_back(d);
            }
        }
    }
    r.atts(gatts);
    {
        const uint32_t tag = r.u32();
        const int64_t n = r.size();
        if (!(tag == 0 && n == 0)) {
            if (tag != TAG_VAR) die("variable list expected");
            for (int64_t i = 0; i < n; ++i) {
                Var v;
                v.name = r.name();
                const int64_t nd = r.size();
                for (int64_t j = 0; j < nd; ++j) v.dimids.push_back((int)r.size());
                r.atts(v.atts);
                v.type = (int)r.u32();
                v.vsize = r.size();
                v.begin = version == 1 ? (int64_t)r.u32() : (int64_t)r.u64();
                vars.push_back(v);
            }
        }
    }
    finish_layout(*this);
    if (numrecs < 0) {
        fseeko(fp, 0, SEEK_END);
        const int64_t end = ftello(fp);
        int64_t first = end;
        for (const Var& v : vars)
            if (v.record) first = std::min(first, v.begin);
        numrecs = recsize > 0 ? (end - first) / recsize : 0;
    }
    defined = true;
}

int File::find_var(const std::string& name) const {
    for (size_t i = 0; i < vars.size(); ++i)
        if (vars[i].name == name) return (int)i;
    return -1;
}
int File::find_dim(const std::string& name) const {
    for (size_t i = 0; i < dims.size(); ++i)
        if (dims[i].name == name) return (int)i;
    return -1;
}
const Att* File::find_gatt(const std::string& name) const {
    for (const Att& a : gatts)
        if (a.name == name) return &a;
    return nullptr;
}
int64_t File::var_count(int varid) const {
    const Var& v = vars.at(varid);
    return v.record ? v.count_per_record * numrecs : v.count_per_record;
}

namespace {
template <class T>
void read_converted(const File& f, const Var& v, int64_t rec, int64_t start, int64_t count, T* out) {
    const int sz = type_size(v.type);
    const int64_t CH = 1 << 16;
    std::vector<unsigned char> raw((size_t)(CH * sz)), host((size_t)(CH * sz));
    // a range outside the variable would silently read the neighbouring variable / record (write_converted checks the same)
    if (start < 0 || count < 0 || start + count > v.count_per_record) die("read outside variable " + v.name + " in " + f.path);
    if (rec < 0 || (v.record && rec >= f.numrecs) || (!v.record && rec != 0)) die("no such record of variable " + v.name + " in " + f.path);
    if (fseeko(f.fp, var_offset(f, v, rec, start), SEEK_SET) != 0) die("seek failed in " + f.path);
    for (int64_t done = 0; done < count; done += CH) {
        const int64_t n = std::min(CH, count - done);
        if (std::fread(raw.data(), (size_t)sz, (size_t)n, f.fp) != (size_t)n) die("short read of variable " + v.name);
        swap_copy(host.data(), raw.data(), n, sz);
        for (int64_t i = 0; i < n; ++i) {
            const unsigned char* p = host.data() + i * sz;
            switch (v.type) {
                case NC_DOUBLE: { double x; std::memcpy(&x, p, 8); out[done + i] = (T)x; break; }
                case NC_FLOAT: { float x; std::memcpy(&x, p, 4); out[done + i] = (T)x; break; }
                case NC_INT: { int32_t x; std::memcpy(&x, p, 4); out[done + i] = (T)x; break; }
                case NC_UINT: { uint32_t x; std::memcpy(&x, p, 4); out[done + i] = (T)x; break; }
                case NC_SHORT: { int16_t x; std::memcpy(&x, p, 2); out[done + i] = (T)x; break; }
                case NC_USHORT: { uint16_t x; std::memcpy(&x, p, 2); out[done + i] = (T)x; break; }
                case NC_INT64: { int64_t x; std::memcpy(&x, p, 8); out[done + i] = (T)x; break; }
                case NC_UINT64: { uint64_t x; std::memcpy(&x, p, 8); out[done + i] = (T)x; break; }
                case NC_BYTE: out[done + i] = (T)(signed char)*p; break;
                default: out[done + i] = (T)*p; break;
            }
        }
    }
}
template <class T>
void read_all(const File& f, int varid, T* out) {
    const Var& v = f.vars.at(varid);
    if (!v.record) { read_converted(f, v, 0, 0, v.count_per_record, out); return; }
    for (int64_t r = 0; r < f.numrecs; ++r) read_converted(f, v, r, 0, v.count_per_record, out + r * v.count_per_record);
}
}  // namespace

void File::get_double(int varid, double* out) const { read_all(*this, varid, out); }
void File::get_int(int varid, int* out) const { read_all(*this, varid, out); }
void File::get_double_range(int varid, int64_t rec, int64_t start, int64_t count, double* out) const {
    read_converted(*this, vars.at(varid), rec, start, count, out);
}

// ------------------------------------------------------------------ writing
void File::create(const std::string& p, int ver) {
    if (ver != 1 && ver != 2 && ver != 5) die("create: version must be 1, 2 or 5");
    path = p;
    version = ver;
    writing = true;
    fp = std::fopen(p.c_str(), "wb");
    if (!fp) die("cannot create " + p);
}
int File::def_dim(const std::string& name, int64_t len) {
    if (defined) die("def_dim after enddef");
    if (len == 0 && [&] { for (auto& d : dims) if (d.len == 0) return true; return false; }()) die("only one unlimited dimension");
    Dim d;
    d.name = name;
    d.len = len;
    dims.push_back(d);
    return (int)dims.size() - 1;
}
int File::def_var(const std::string& name, int type, const std::vector<int>& dimids) {
    if (defined) die("def_var after enddef");
    for (size_t i = 0; i < dimids.size(); ++i)
        if (dimids[i] < 0 || dimids[i] >= (int)dims.size() || (i > 0 && dims[dimids[i]].len == 0)) die("bad dimension list for " + name);
    Var v;
    v.name = name;
    v.type = type;
    v.dimids = dimids;
    vars.push_back(v);
    return (int)vars.size() - 1;
}
static Att text_att(const std::string& name, const std::string& value) {
    Att a;
    a.name = name;
    a.type = NC_CHAR;
    a.nelems = (int64_t)value.size();
    a.data.assign(value.begin(), value.end());
    return a;
}
void File::put_gatt_text(const std::string& name, const std::string& value) { gatts.push_back(text_att(name, value)); }
void File::put_gatt_double(const std::string& name, double v) {
    Att a;
    a.name = name;
    a.type = NC_DOUBLE;
    a.nelems = 1;
    a.data.resize(8);
    std::memcpy(a.data.data(), &v, 8);
    gatts.push_back(a);
}
void File::put_att_text(int varid, const std::string& name, const std::string& value) { vars.at(varid).atts.push_back(text_att(name, value)); }

void File::enddef(int64_t nrecs) {
    if (!writing || defined) die("enddef: not in define mode");
    numrecs = nrecs;
    finish_layout(*this);
    for (Var& v : vars) {
        int64_t bytes = v.count_per_record * type_size(v.type);
        bytes = (bytes + 3) / 4 * 4;
        if (version != 5 && bytes > 0xFFFFFFFCLL) die("variable " + v.name + " needs CDF-5");
        v.vsize = bytes;
    }
    finish_layout(*this);
    Buf probe;
    serialise_header(*this, probe);  // begins are still 0: the size does not depend on their value
    int64_t off = (int64_t)probe.d.size();
    for (Var& v : vars)
        if (!v.record) { v.begin = off; off += v.vsize; }
    for (Var& v : vars)
        if (v.record) { v.begin = off; off += (recsize == v.count_per_record * type_size(v.type)) ? recsize : v.vsize; }
    if (version == 1 && off + numrecs * recsize > 0x7FFFFFFFLL) die("file too large for CDF-1");
    Buf b;
    serialise_header(*this, b);
    if (std::fwrite(b.d.data(), 1, b.d.size(), fp) != b.d.size()) die("header write failed");
    // extend the file to its full size so that independent writers can fill disjoint ranges
    int64_t end = 0;
    for (const Var& v : vars)
        if (!v.record) end = std::max(end, v.begin + v.vsize);
    int64_t first_rec = end;
    for (const Var& v : vars)
        if (v.record) { first_rec = v.begin; break; }
    if (recsize > 0) end = std::max(end, first_rec + numrecs * recsize);
    if (end > (int64_t)b.d.size()) {
        if (fseeko(fp, end - 1, SEEK_SET) != 0) die("seek failed");
        const unsigned char z = 0;
        if (std::fwrite(&z, 1, 1, fp) != 1) die("extend failed");
    }
    std::fflush(fp);
    defined = true;
}

void File::open_for_update(const std::string& p) {
    open(p);
    std::fclose(fp);
    fp = std::fopen(p.c_str(), "r+b");
    if (!fp) die("cannot reopen for update: " + p);
    writing = true;
}

namespace {
template <class T>
void write_converted(File& f, const Var& v, int64_t rec, int64_t start, int64_t count, const T* data) {
    if (!f.writing || !f.defined) die("put before enddef");
    if (start < 0 || start + count > v.count_per_record) die("put outside variable " + v.name);
    if (v.record && (rec < 0 || rec >= f.numrecs)) die("record index outside numrecs for " + v.name);
    const int sz = type_size(v.type);
    const int64_t CH = 1 << 16;
    std::vector<unsigned char> host((size_t)(CH * sz)), be((size_t)(CH * sz));
    if (fseeko(f.fp, var_offset(f, v, rec, start), SEEK_SET) != 0) die("seek failed in " + f.path);
    for (int64_t done = 0; done < count; done += CH) {
        const int64_t n = std::min(CH, count - done);
        for (int64_t i = 0; i < n; ++i) {
            unsigned char* p = host.data() + i * sz;
            const T x = data[done + i];
            switch (v.type) {
                case NC_DOUBLE: { double y = (double)x; std::memcpy(p, &y, 8); break; }
                case NC_FLOAT: { float y = (float)x; std::memcpy(p, &y, 4); break; }
                case NC_INT: { int32_t y = (int32_t)x; std::memcpy(p, &y, 4); break; }
                case NC_INT64: { int64_t y = (int64_t)x; std::memcpy(p, &y, 8); break; }
                case NC_SHORT: { int16_t y = (int16_t)x; std::memcpy(p, &y, 2); break; }
                default: *p = (unsigned char)x; break;
            }
        }
        swap_copy(be.data(), host.data(), n, sz);
        if (std::fwrite(be.data(), (size_t)sz, (size_t)n, f.fp) != (size_t)n) die("short write of variable " + v.name);
    }
}
}  // namespace

void File::put_double(int varid, int64_t rec, int64_t start, int64_t count, const double* data) { write_converted(*this, vars.at(varid), rec, start, count, data); }
void File::put_int(int varid, int64_t rec, int64_t start, int64_t count, const int* data) { write_converted(*this, vars.at(varid), rec, start, count, data); }

}  // namespace nc
}  // namespace mpas
