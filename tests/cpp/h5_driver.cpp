// h5_driver — prints, for every variable of an HDF5/netCDF-4 file (or of a netCDF classic file, through the
// same Netcdf_file class the command-line driver uses), its dimension names, shape, and a checksum of its
// values as doubles; tests/test_h5lite.py compares the lines with the python reader's.
//   h5_driver <file>
#include <cstdio>
#include <cmath>
#include "rrtmgp_b200/netcdf_classic.hpp"

int main(int argc, char** argv)
{
    if (argc != 2) { std::fprintf(stderr, "usage: h5_driver <file>\n"); return 2; }
    try
    {
        rrtmgp_b200::Netcdf_file f(argv[1], rrtmgp_b200::Netcdf_mode::Read);
        for (const std::string& v : f.variable_names())
        {
            std::printf("%s |", v.c_str());
            std::vector<int> shape;
            for (const auto& d : f.get_variable_dimension_list(v)) { std::printf(" %s=%d", d.first.c_str(), d.second); shape.push_back(d.second); }
            if (f.variable_is_char(v)) { std::printf(" | char\n"); continue; }
            const std::vector<double> a = shape.empty() ? std::vector<double>{f.get_variable<double>(v)} : f.get_variable<double>(v, shape);
            double s = 0, w = 0;
            for (size_t i = 0; i < a.size(); ++i) { s += a[i]; w += a[i] * (double)((i % 97) + 1); }
            std::printf(" | n=%zu sum=%.17g wsum=%.17g\n", a.size(), s, w);
        }
    }
    catch (const std::exception& e) { std::fprintf(stderr, "error: %s\n", e.what()); return 1; }
    return 0;
}
