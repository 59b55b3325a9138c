// netcdf_classic.hpp — self-contained reader/writer for the netCDF *classic* family
// (CDF-1 classic, CDF-2 64-bit offset, CDF-5 64-bit data), header-only, no libnetcdf / libhdf5.
//
// It gives the command-line driver (cli/test_rte_rrtmgp.cpp) the handful of calls the reference's
// driver makes on its Netcdf_file wrapper (include_test/Netcdf_interface.h:256-300):
//   get_dimension_size, variable_exists, get_variable_dimensions, get_variable<T>(name[, dims]),
//   add_dimension, add_variable<T>(name, dim names) + insert(values, start).
// netCDF-4 files are HDF5 containers: they are opened through hdf5_lite.hpp (read-only, no libhdf5) and
// presented through the same calls, so the files the reference's python recipes write and rrtmgp-data
// ships are read as they are.  Output is always written as netCDF classic (CDF-2), which libnetcdf and
// netCDF4-python open transparently.
//
// On-disk layout (big-endian): magic "CDF" + version {1,2,5}; numrecs; dim_list; gatt_list;
// var_list {name, dimids, vatt_list, nc_type, vsize, begin}; then the fixed-size variables and
// the interleaved records.  Sizes/counts are 32-bit in CDF-1/2 and 64-bit in CDF-5; `begin` is
// 32-bit in CDF-1 only.
#pragma once
#include <cstdint>
#include <cstring>
#include <fstream>
#include <map>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>
#include <memory>
#include <utility>
#include "hdf5_lite.hpp"

namespace rrtmgp_b200 { namespace nc {

enum Type { T_BYTE = 1, T_CHAR = 2, T_SHORT = 3, T_INT = 4, T_FLOAT = 5, T_DOUBLE = 6,
            T_UBYTE = 7, T_USHORT = 8, T_UINT = 9, T_INT64 = 10, T_UINT64 = 11 };

inline size_t type_size(const int t)
{
    switch (t)
    {
        case T_BYTE: case T_CHAR: case T_UBYTE: return 1;
        case T_SHORT: case T_USHORT: return 2;
        case T_INT: case T_FLOAT: case T_UINT: return 4;
        case T_DOUBLE: case T_INT64: case T_UINT64: return 8;
    }
    throw std::runtime_error("netcdf: unknown nc_type " + std::to_string(t));
}

template<typename T> constexpr int type_of()
{
    if (std::is_same<T, float>::value) return T_FLOAT;
    if (std::is_same<T, double>::value) return T_DOUBLE;
    if (std::is_same<T, int>::value) return T_INT;
    if (std::is_same<T, char>::value) return T_CHAR;
    if (std::is_same<T, signed char>::value) return T_BYTE;
    return 0;
}

// big-endian <-> host, element by element
template<typename U> inline U load_be(const unsigned char* p)
{
    typename std::conditional<sizeof(U) == 1, uint8_t, typename std::conditional<sizeof(U) == 2, uint16_t,
             typename std::conditional<sizeof(U) == 4, uint32_t, uint64_t>::type>::type>::type v = 0;
    for (size_t i = 0; i < sizeof(U); ++i) v = (decltype(v))((v << 8) | p[i]);
    U out; std::memcpy(&out, &v, sizeof(U)); return out;
}
template<typename U> inline void store_be(unsigned char* p, const U x)
{
    typename std::conditional<sizeof(U) == 1, uint8_t, typename std::conditional<sizeof(U) == 2, uint16_t,
             typename std::conditional<sizeof(U) == 4, uint32_t, uint64_t>::type>::type>::type v;
    std::memcpy(&v, &x, sizeof(U));
    for (size_t i = 0; i < sizeof(U); ++i) p[i] = (unsigned char)(v >> (8 * (sizeof(U) - 1 - i)));
}

template<typename T> inline T element_as(const unsigned char* p, const int t)
{
    switch (t)
    {
        case T_BYTE:   return (T)load_be<int8_t>(p);
        case T_CHAR:   return (T)load_be<uint8_t>(p);
        case T_UBYTE:  return (T)load_be<uint8_t>(p);
        case T_SHORT:  return (T)load_be<int16_t>(p);
        case T_USHORT: return (T)load_be<uint16_t>(p);
        case T_INT:    return (T)load_be<int32_t>(p);
        case T_UINT:   return (T)load_be<uint32_t>(p);
        case T_FLOAT:  return (T)load_be<float>(p);
        case T_DOUBLE: return (T)load_be<double>(p);
        case T_INT64:  return (T)load_be<int64_t>(p);
        case T_UINT64: return (T)load_be<uint64_t>(p);
    }
    throw std::runtime_error("netcdf: unknown nc_type");
}

struct Dim { std::string name; uint64_t len = 0; };                 // len 0 in the file = the record dimension
struct Var
{
    std::string name; std::vector<int> dimids; int type = 0;
    uint64_t vsize = 0, begin = 0; bool record = false;
    std::vector<unsigned char> bytes;                                // create mode: big-endian payload
    uint64_t count = 0;                                              // create mode: number of elements
    bool hdf5 = false;                                               // read through Netcdf_file::h5file
};

inline uint64_t pad4(const uint64_t n) { return (n + 3) & ~uint64_t(3); }

}  // namespace nc

enum class Netcdf_mode { Create, Read };

class Netcdf_file;

// handle returned by add_variable<T>: insert() stores the values (whole variable, start = 0...) in the file image
template<typename T>
class Netcdf_variable
{
    public:
        Netcdf_variable(Netcdf_file& f, const int id) : file(f), var_id(id) {}
        void insert(const std::vector<T>& values, const std::vector<int>& start);
        void insert(const T value);
    private:
        Netcdf_file& file; int var_id;
};

class Netcdf_file
{
    public:
        Netcdf_file(const std::string& name, const Netcdf_mode mode_in) : file_name(name), mode(mode_in)
        {
            if (mode == Netcdf_mode::Read) read_header();
        }
        ~Netcdf_file() { if (mode == Netcdf_mode::Create && !written) { try { sync(); } catch (...) {} } }
        Netcdf_file(const Netcdf_file&) = delete;

        // ---------------- read side
        int get_dimension_size(const std::string& name) const
        {
            for (const auto& d : dims) if (d.name == name) return (int)(d.len ? d.len : numrecs);
            throw std::runtime_error("netcdf: dimension \"" + name + "\" not in " + file_name);
        }
        bool dimension_exists(const std::string& name) const
        { for (const auto& d : dims) if (d.name == name) return true; return false; }
        bool variable_exists(const std::string& name) const { return find(name) >= 0; }
        std::vector<std::string> variable_names() const { std::vector<std::string> n; for (const auto& v : vars) n.push_back(v.name); return n; }
        bool variable_is_char(const std::string& name) const { return var(name).type == nc::T_CHAR; }
        // (dimension name, size) in file order, slowest first
        std::vector<std::pair<std::string, int>> get_variable_dimension_list(const std::string& name) const
        {
            std::vector<std::pair<std::string, int>> l;
            for (const int id : var(name).dimids) l.push_back({dims[id].name, (int)dim_len(id)});
            return l;
        }
        std::map<std::string, int> get_variable_dimensions(const std::string& name) const
        {
            std::map<std::string, int> m;
            for (const int id : var(name).dimids) m[dims[id].name] = (int)(dims[id].len ? dims[id].len : numrecs);
            return m;
        }
        // scalar variable
        template<typename T> T get_variable(const std::string& name) const
        {
            const std::vector<T> v = read<T>(var(name));
            if (v.size() != 1) throw std::runtime_error("netcdf: variable \"" + name + "\" is not a scalar");
            return v[0];
        }
        // array variable; `shape` must be the variable's shape in file order (slowest first)
        template<typename T> std::vector<T> get_variable(const std::string& name, const std::vector<int>& shape) const
        {
            const nc::Var& v = var(name);
            if (shape.size() != v.dimids.size())
                throw std::runtime_error("netcdf: variable \"" + name + "\" has " + std::to_string(v.dimids.size()) + " dimensions, " +
                                         std::to_string(shape.size()) + " requested");
            for (size_t i = 0; i < shape.size(); ++i)
                if ((uint64_t)shape[i] != dim_len(v.dimids[i]))
                    throw std::runtime_error("netcdf: variable \"" + name + "\": dimension " + dims[v.dimids[i]].name + " is " +
                                             std::to_string(dim_len(v.dimids[i])) + ", requested " + std::to_string(shape[i]));
            return read<T>(v);
        }

        // ---------------- create side
        void add_dimension(const std::string& name, const int size)
        {
            require_create();
            if (dimension_exists(name)) throw std::runtime_error("netcdf: dimension \"" + name + "\" defined twice");
            nc::Dim d; d.name = name; d.len = (uint64_t)size; dims.push_back(d);
        }
        template<typename T> Netcdf_variable<T> add_variable(const std::string& name, const std::vector<std::string>& dim_names = {})
        {
            require_create();
            static_assert(nc::type_of<T>() != 0, "netcdf: unsupported variable type");
            if (variable_exists(name)) throw std::runtime_error("netcdf: variable \"" + name + "\" defined twice");
            nc::Var v; v.name = name; v.type = nc::type_of<T>(); v.count = 1;
            for (const auto& dn : dim_names)
            {
                int id = -1;
                for (size_t i = 0; i < dims.size(); ++i) if (dims[i].name == dn) id = (int)i;
                if (id < 0) throw std::runtime_error("netcdf: dimension \"" + dn + "\" of variable \"" + name + "\" not defined");
                v.dimids.push_back(id); v.count *= dims[id].len;
            }
            vars.push_back(std::move(v));
            return Netcdf_variable<T>(*this, (int)vars.size() - 1);
        }
        template<typename T> void store(const int id, const T* values, const size_t n, const std::vector<int>& start)
        {
            nc::Var& v = vars[id];
            for (const int s : start) if (s != 0) throw std::runtime_error("netcdf: partial insert into \"" + v.name + "\" is not supported");
            if (n != v.count) throw std::runtime_error("netcdf: insert into \"" + v.name + "\": " + std::to_string(n) + " values for " +
                                                       std::to_string(v.count) + " elements");
            v.bytes.resize(n * sizeof(T));
            for (size_t i = 0; i < n; ++i) nc::store_be<T>(v.bytes.data() + i * sizeof(T), values[i]);
        }
        void sync();     // write the file (CDF-2); also called by the destructor

    private:
        void require_create() const { if (mode != Netcdf_mode::Create) throw std::runtime_error("netcdf: " + file_name + " is open read-only"); }
        int find(const std::string& name) const
        { for (size_t i = 0; i < vars.size(); ++i) if (vars[i].name == name) return (int)i; return -1; }
        const nc::Var& var(const std::string& name) const
        {
            const int i = find(name);
            if (i < 0) throw std::runtime_error("netcdf: variable \"" + name + "\" not in " + file_name);
            return vars[i];
        }
        uint64_t dim_len(const int id) const { return dims[id].len ? dims[id].len : numrecs; }
        void read_header();
        void read_hdf5_header();
        std::shared_ptr<h5::File> h5file;
        template<typename T> std::vector<T> read(const nc::Var& v) const;

        std::string file_name; Netcdf_mode mode; bool written = false;
        int version = 2; uint64_t numrecs = 0, recsize = 0;
        std::vector<nc::Dim> dims; std::vector<nc::Var> vars;
};

// ---------------------------------------------------------------- implementation
namespace nc {
// sequential big-endian reader over the header
struct Cursor
{
    std::ifstream& in; const int version; const std::string& file;
    void raw(void* dst, const size_t n)
    {
        in.read(reinterpret_cast<char*>(dst), (std::streamsize)n);
        if (!in) throw std::runtime_error("netcdf: " + file + " ends inside the header");
    }
    uint32_t u32() { unsigned char b[4]; raw(b, 4); return load_be<uint32_t>(b); }
    uint64_t u64() { unsigned char b[8]; raw(b, 8); return load_be<uint64_t>(b); }
    uint64_t size() { return version == 5 ? u64() : (uint64_t)u32(); }       // NON_NEG: 32-bit in CDF-1/2, 64-bit in CDF-5
    std::string name()
    {
        const uint64_t n = size();
        std::string s(pad4(n), '\0');
        if (n) raw(&s[0], s.size());
        s.resize(n);
        return s;
    }
    void skip_attributes()
    {
        const uint32_t tag = u32(); const uint64_t n = size();
        if (tag == 0 && n == 0) return;
        if (tag != 0x0C) throw std::runtime_error("netcdf: " + file + ": attribute list expected");
        for (uint64_t i = 0; i < n; ++i)
        {
            name();
            const int t = (int)u32(); const uint64_t ne = size();
            in.seekg((std::streamoff)pad4(ne * type_size(t)), std::ios::cur);
        }
    }
};
}  // namespace nc

inline void Netcdf_file::read_header()
{
    std::ifstream in(file_name, std::ios::binary);
    if (!in) throw std::runtime_error("netcdf: cannot open " + file_name);
    unsigned char magic[4] = {0, 0, 0, 0};
    in.read(reinterpret_cast<char*>(magic), 4);
    if (in && magic[0] == 0x89 && magic[1] == 'H' && magic[2] == 'D' && magic[3] == 'F') { in.close(); read_hdf5_header(); return; }
    if (!in || magic[0] != 'C' || magic[1] != 'D' || magic[2] != 'F' || (magic[3] != 1 && magic[3] != 2 && magic[3] != 5))
        throw std::runtime_error("netcdf: " + file_name + " is not a netCDF classic (CDF-1/2/5) file");
    version = magic[3];
    nc::Cursor c{in, version, file_name};
    numrecs = c.size();
    if (version != 5 && numrecs == 0xFFFFFFFFu) numrecs = 0;                    // STREAMING marker
    {   // dim_list
        const uint32_t tag = c.u32(); const uint64_t n = c.size();
        if (!(tag == 0 && n == 0))
        {
            if (tag != 0x0A) throw std::runtime_error("netcdf: " + file_name + ": dimension list expected");
            for (uint64_t i = 0; i < n; ++i) { nc::Dim d; d.name = c.name(); d.len = c.size(); dims.push_back(d); }
        }
    }
    c.skip_attributes();
    {   // var_list
        const uint32_t tag = c.u32(); const uint64_t n = c.size();
        if (!(tag == 0 && n == 0))
        {
            if (tag != 0x0B) throw std::runtime_error("netcdf: " + file_name + ": variable list expected");
            for (uint64_t i = 0; i < n; ++i)
            {
                nc::Var v; v.name = c.name();
                const uint64_t nd = c.size();
                for (uint64_t k = 0; k < nd; ++k)
                {
                    const uint64_t id = c.size();
                    if (id >= dims.size()) throw std::runtime_error("netcdf: " + file_name + ": bad dimension id in \"" + v.name + "\"");
                    v.dimids.push_back((int)id);
                }
                c.skip_attributes();
                v.type = (int)c.u32(); v.vsize = c.size();
                v.begin = version == 1 ? (uint64_t)c.u32() : c.u64();
                v.record = !v.dimids.empty() && dims[v.dimids[0]].len == 0;
                vars.push_back(std::move(v));
            }
        }
    }
    // record stride: sum of the per-record sizes; a single record variable is stored unpadded
    int nrec = 0; uint64_t one = 0;
    for (const auto& v : vars) if (v.record)
    {
        uint64_t per = nc::type_size(v.type);
        for (size_t k = 1; k < v.dimids.size(); ++k) per *= dims[v.dimids[k]].len;
        recsize += nc::pad4(per); one = per; ++nrec;
    }
    if (nrec == 1) recsize = one;
}

// netCDF-4: dimensions are the HDF5 dimension scales, variables the other datasets; a variable whose element type this
// reader does not decode (variable-length strings) is left out, a variable whose dimension names are stored out of reach
// (dense attribute storage) gets anonymous dimensions
inline void Netcdf_file::read_hdf5_header()
{
    h5file = std::make_shared<h5::File>(file_name);
    const std::vector<std::string> names = h5file->names();
    for (const std::string& n : names)
    {
        const h5::Object& o = h5file->get(n);
        if (o.is_dataset && o.is_scale && o.shape.size() == 1) { nc::Dim d; d.name = n; d.len = o.shape[0]; dims.push_back(d); }
    }
    const auto dim_id = [this](const std::string& n, const uint64_t len)
    {
        for (size_t i = 0; i < dims.size(); ++i) if (dims[i].name == n) return (int)i;
        nc::Dim d; d.name = n; d.len = len; dims.push_back(d);
        return (int)dims.size() - 1;
    };
    for (const std::string& n : names)
    {
        const h5::Object& o = h5file->get(n);
        if (!o.is_dataset || o.pure_dimension || !o.has_type) continue;
        nc::Var v; v.name = n; v.hdf5 = true;
        const h5::Type& t = o.type;
        if (t.cls == 1) v.type = t.size == 4 ? nc::T_FLOAT : nc::T_DOUBLE;
        else if (t.cls == 0) v.type = t.size == 1 ? (t.is_signed ? nc::T_BYTE : nc::T_UBYTE) : t.size == 2 ? (t.is_signed ? nc::T_SHORT : nc::T_USHORT)
                                    : t.size == 4 ? (t.is_signed ? nc::T_INT : nc::T_UINT) : (t.is_signed ? nc::T_INT64 : nc::T_UINT64);
        else if (t.cls == 3 && t.size == 1) v.type = nc::T_CHAR;
        else continue;
        std::vector<std::string> dn;
        try { dn = h5file->dims(n); }
        catch (const std::runtime_error&) { dn.clear(); for (size_t i = 0; i < o.shape.size(); ++i) dn.push_back(n + "_dim" + std::to_string(i)); }
        if (dn.size() != o.shape.size()) throw std::runtime_error("netcdf: " + file_name + ": \"" + n + "\" has " + std::to_string(o.shape.size()) +
                                                                  " dimensions but names " + std::to_string(dn.size()));
        v.count = 1;
        for (size_t i = 0; i < dn.size(); ++i) { v.dimids.push_back(dim_id(dn[i], o.shape[i])); v.count *= o.shape[i]; }
        for (size_t i = 0; i < dn.size(); ++i)
            if (dims[v.dimids[i]].len != o.shape[i])      // an unlimited dimension: the variable's own extent counts
                v.dimids[i] = dim_id(n + "_dim" + std::to_string(i), o.shape[i]);
        vars.push_back(std::move(v));
    }
}

template<typename T> inline std::vector<T> Netcdf_file::read(const nc::Var& v) const
{
    if (mode != Netcdf_mode::Read) throw std::runtime_error("netcdf: " + file_name + " is open for writing");
    if (v.hdf5) return h5file->read<T>(v.name);
    const size_t ts = nc::type_size(v.type);
    uint64_t per = 1;                                   // elements per record (record var) or in total
    for (size_t k = v.record ? 1 : 0; k < v.dimids.size(); ++k) per *= dims[v.dimids[k]].len;
    const uint64_t nrec = v.record ? numrecs : 1;
    std::ifstream in(file_name, std::ios::binary);
    if (!in) throw std::runtime_error("netcdf: cannot open " + file_name);
    std::vector<T> out((size_t)(per * nrec));
    std::vector<unsigned char> buf;
    const uint64_t chunk = uint64_t(1) << 22;          // elements per read
    for (uint64_t r = 0; r < nrec; ++r)
    {
        in.seekg((std::streamoff)(v.begin + r * recsize));
        for (uint64_t e0 = 0; e0 < per; e0 += chunk)
        {
            const uint64_t ne = std::min<uint64_t>(chunk, per - e0);
            buf.resize((size_t)(ne * ts));
            in.read(reinterpret_cast<char*>(buf.data()), (std::streamsize)buf.size());
            if (!in) throw std::runtime_error("netcdf: " + file_name + " ends inside variable \"" + v.name + "\"");
            T* o = out.data() + r * per + e0;
            for (uint64_t i = 0; i < ne; ++i) o[i] = nc::element_as<T>(buf.data() + i * ts, v.type);
        }
    }
    return out;
}

inline void Netcdf_file::sync()
{
    require_create();
    // CDF-2 (64-bit offset): 32-bit counts, 64-bit `begin`; a variable larger than 4 GiB stores vsize = 2^32-1
    std::vector<unsigned char> h;
    auto u32 = [&h](const uint32_t x) { unsigned char b[4]; nc::store_be<uint32_t>(b, x); h.insert(h.end(), b, b + 4); };
    auto u64 = [&h](const uint64_t x) { unsigned char b[8]; nc::store_be<uint64_t>(b, x); h.insert(h.end(), b, b + 8); };
    auto str = [&h, &u32](const std::string& s) { u32((uint32_t)s.size()); h.insert(h.end(), s.begin(), s.end()); h.resize(nc::pad4(h.size()), 0); };
    std::vector<size_t> begin_at;                       // header offsets of the `begin` fields
    auto header = [&](const std::vector<uint64_t>& begins)
    {
        h.clear(); begin_at.clear();
        h.insert(h.end(), {'C', 'D', 'F', 2});
        u32(0);                                          // numrecs: no record variables are written
        if (dims.empty()) { u32(0); u32(0); }
        else { u32(0x0A); u32((uint32_t)dims.size()); for (const auto& d : dims) { str(d.name); u32((uint32_t)d.len); } }
        u32(0); u32(0);                                  // no global attributes
        if (vars.empty()) { u32(0); u32(0); }
        else
        {
            u32(0x0B); u32((uint32_t)vars.size());
            for (size_t i = 0; i < vars.size(); ++i)
            {
                const nc::Var& v = vars[i];
                str(v.name); u32((uint32_t)v.dimids.size());
                for (const int id : v.dimids) u32((uint32_t)id);
                u32(0); u32(0);                          // no attributes
                u32((uint32_t)v.type);
                const uint64_t vs = nc::pad4(v.count * nc::type_size(v.type));
                u32(vs > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)vs);
                begin_at.push_back(h.size()); u64(begins.empty() ? 0 : begins[i]);
            }
        }
    };
    header({});
    std::vector<uint64_t> begins(vars.size());
    uint64_t off = h.size();
    for (size_t i = 0; i < vars.size(); ++i) { begins[i] = off; off += nc::pad4(vars[i].count * nc::type_size(vars[i].type)); }
    header(begins);
    std::ofstream out(file_name, std::ios::binary | std::ios::trunc);
    if (!out) throw std::runtime_error("netcdf: cannot create " + file_name);
    out.write(reinterpret_cast<const char*>(h.data()), (std::streamsize)h.size());
    const char zeros[4] = {0, 0, 0, 0};
    for (const auto& v : vars)
    {
        const uint64_t nbytes = v.count * nc::type_size(v.type);
        if (v.bytes.size() == nbytes) out.write(reinterpret_cast<const char*>(v.bytes.data()), (std::streamsize)nbytes);
        else { const std::vector<char> z((size_t)nbytes, 0); out.write(z.data(), (std::streamsize)nbytes); }   // defined, never inserted
        out.write(zeros, (std::streamsize)(nc::pad4(nbytes) - nbytes));
    }
    if (!out) throw std::runtime_error("netcdf: write to " + file_name + " failed");
    written = true;
}

template<typename T> inline void Netcdf_variable<T>::insert(const std::vector<T>& values, const std::vector<int>& start)
{ file.template store<T>(var_id, values.data(), values.size(), start); }
template<typename T> inline void Netcdf_variable<T>::insert(const T value)
{ file.template store<T>(var_id, &value, 1, {}); }

}  // namespace rrtmgp_b200
