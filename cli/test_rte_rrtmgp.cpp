// test_rte_rrtmgp_b200 — command-line driver with the flags and the file contract of the reference's
// `test_rte_rrtmgp` (src_test/test_rte_rrtmgp.cu:171-818, CPU twin src_test/test_rte_rrtmgp.cpp):
//   reads   rte_rrtmgp_input.nc, coefficients_{lw,sw}.nc, cloud_coefficients_{lw,sw}.nc, aerosol_optics.nc   (cwd)
//   writes  rte_rrtmgp_output.nc  — same variable names, dimension names and dimension orders
//   flags   --[no-]shortwave --[no-]longwave --[no-]fluxes --[no-]cloud-optics --[no-]output-optical
//           --[no-]output-bnd-fluxes --[no-]delta-cloud --[no-]timings --[no-]aerosol-optics --[no-]delta-aerosol
// Files are netCDF classic (CDF-1/2/5); see include/rrtmgp_b200/netcdf_classic.hpp.  All radiation work
// goes through include/rrtmgp_b200/classes.hpp -> librrtmgp_b200.so; nothing is computed on the host.
// Build: g++ -std=c++17 [-DRTE_USE_SP] -Iinclude cli/test_rte_rrtmgp.cpp -lrrtmgp_b200 -lcudart
#include <cstdio>
#include <iomanip>
#include <iostream>
#include <sstream>

#include "rrtmgp_b200/classes.hpp"
#include "rrtmgp_b200/coefficients.hpp"

using rrtmgp_b200::Netcdf_file;
using rrtmgp_b200::Netcdf_mode;

namespace {

void say(const std::string& s) { std::cout << s << std::endl; }

struct Option { const char* name; bool value; const char* help; };

// ---- option table: same names and defaults as src_test/test_rte_rrtmgp.cu:178-189
std::vector<Option> default_options()
{
    return {
        {"shortwave", true, "Enable computation of shortwave radiation."},
        {"longwave", true, "Enable computation of longwave radiation."},
        {"fluxes", true, "Enable computation of fluxes."},
        {"cloud-optics", false, "Enable cloud optics."},
        {"aerosol-optics", false, "Enable aerosol optics."},
        {"output-optical", false, "Enable output of optical properties."},
        {"output-bnd-fluxes", false, "Enable output of band fluxes."},
        {"timings", false, "Repeat computation 10x for run times."},
        {"delta-cloud", true, "delta-scaling of cloud optical properties"},
        {"delta-aerosol", false, "delta-scaling of aerosol optical properties"}};
}

// returns true when only the help text was asked for
bool parse(std::vector<Option>& opts, int argc, char** argv)
{
    for (int i = 1; i < argc; ++i)
    {
        std::string a(argv[i]);
        a.erase(0, a.find_first_not_of(" \t")); a.erase(a.find_last_not_of(" \t") + 1);
        if (a == "-h" || a == "--help")
        {
            say("Possible usage:");
            for (const auto& o : opts) { std::ostringstream ss; ss << std::left << std::setw(30) << ("--" + std::string(o.name)) << o.help; say(ss.str()); }
            return true;
        }
        if (a.rfind("--", 0) != 0) throw std::runtime_error(a + " is an illegal command line option.");
        std::string key = a.substr(2);
        bool enable = true;
        if (key.rfind("no-", 0) == 0) { enable = false; key.erase(0, 3); }
        bool found = false;
        for (auto& o : opts) if (key == o.name) { o.value = enable; found = true; }
        if (!found) throw std::runtime_error(key + " is an illegal command line option.");
    }
    return false;
}

bool opt(const std::vector<Option>& opts, const std::string& name)
{
    for (const auto& o : opts) if (name == o.name) return o.value;
    throw std::runtime_error("unknown option " + name);
}

// vmr_<gas>: scalar, (lay) or (lay, y, x) — src_test/test_rte_rrtmgp.cpp:31-68
void read_and_set_vmr(const std::string& gas, const int nx, const int ny, const int nlay, const Netcdf_file& in, Gas_concs_gpu& concs)
{
    const std::string var = "vmr_" + gas;
    if (!in.variable_exists(var)) { say("WARNING: Gas \"" + gas + "\" not available in input file."); return; }
    const std::map<std::string, int> dims = in.get_variable_dimensions(var);
    const auto bad = [&gas]() { return std::runtime_error("Illegal dimensions of gas \"" + gas + "\" in input"); };
    if (dims.empty())
        concs.set_vmr(gas, in.get_variable<Float>(var));
    else if (dims.size() == 1)
    {
        if (!dims.count("lay") || dims.at("lay") != nlay) throw bad();
        concs.set_vmr(gas, in.get_variable<Float>(var, {nlay}), 1, nlay);
    }
    else if (dims.size() == 3)
    {
        if (!dims.count("lay") || !dims.count("x") || !dims.count("y") || dims.at("lay") != nlay || dims.at("x") != nx || dims.at("y") != ny) throw bad();
        concs.set_vmr(gas, in.get_variable<Float>(var, {nlay, ny, nx}), nx * ny, nlay);
    }
    else throw bad();
}

// aermr01..aermr11: (lay) or (lay, y, x) — src_test/test_rte_rrtmgp.cpp:69-100
void read_and_set_aer(const std::string& name, const int nx, const int ny, const int nlay, const Netcdf_file& in, Aerosol_concs_gpu& concs)
{
    if (!in.variable_exists(name)) throw std::runtime_error("Aerosol type \"" + name + "\" not available in input file.");
    const std::map<std::string, int> dims = in.get_variable_dimensions(name);
    const auto bad = [&name]() { return std::runtime_error("Illegal dimensions of \"" + name + "\" in input"); };
    if (dims.size() == 1)
    {
        if (!dims.count("lay") || dims.at("lay") != nlay) throw std::runtime_error("Illegal dimensions of gas \"" + name + "\" in input");
        concs.set_vmr(name, in.get_variable<Float>(name, {nlay}), 1, nlay);
    }
    else if (dims.size() == 3)
    {
        if (!dims.count("lay") || !dims.count("x") || !dims.count("y") || dims.at("lay") != nlay || dims.at("x") != nx || dims.at("y") != ny) throw bad();
        concs.set_vmr(name, in.get_variable<Float>(name, {nlay, ny, nx}), nx * ny, nlay);
    }
    else throw bad();
}

template<int N> Array_gpu<Float,N> upload(const Netcdf_file& in, const std::string& var, const std::vector<int>& file_shape, const std::array<int,N>& dims)
{ return Array_gpu<Float,N>(in.get_variable<Float>(var, file_shape), dims); }

template<int N> void write(Netcdf_file& out, const std::string& var, const std::vector<std::string>& dim_names, const Array_gpu<Float,N>& a)
{
    auto v = out.add_variable<Float>(var, dim_names);
    v.insert(a.to_host(), std::vector<int>(dim_names.size(), 0));
}

// run once; with --timings ten more times between CUDA events (the reference's run_solver lambdas, :418-470)
template<typename F> void run_timed(const bool timings, const std::string& what, F&& solve)
{
    cudaEvent_t t0, t1;
    rrtmgp_b200::cuda_check(cudaEventCreate(&t0)); rrtmgp_b200::cuda_check(cudaEventCreate(&t1));
    const int n = timings ? 11 : 1;
    for (int i = 0; i < n; ++i)
    {
        rrtmgp_b200::cuda_check(cudaDeviceSynchronize());
        rrtmgp_b200::cuda_check(cudaEventRecord(t0, rrtmgp_b200::stream()));
        solve();
        rrtmgp_b200::cuda_check(cudaEventRecord(t1, rrtmgp_b200::stream()));
        rrtmgp_b200::cuda_check(cudaEventSynchronize(t1));
        float ms = 0.f;
        rrtmgp_b200::cuda_check(cudaEventElapsedTime(&ms, t0, t1));
        say("Duration " + what + " solver: " + std::to_string(ms) + " (ms)");
    }
    cudaEventDestroy(t0); cudaEventDestroy(t1);
}

}  // namespace

static void solve_radiation(int argc, char** argv)
{
    say("###### Starting RTE+RRTMGP solver (B200) ######");
    std::vector<Option> opts = default_options();
    if (parse(opts, argc, argv)) return;
    const bool sw_on = opt(opts, "shortwave"), lw_on = opt(opts, "longwave"), fluxes = opt(opts, "fluxes");
    const bool cloud = opt(opts, "cloud-optics"), optical = opt(opts, "output-optical"), bnd = opt(opts, "output-bnd-fluxes");
    const bool timings = opt(opts, "timings"), delta_cloud = opt(opts, "delta-cloud");
    say("Solver settings:");
    for (const auto& o : opts) { std::ostringstream ss; ss << std::left << std::setw(20) << o.name << " = " << std::boolalpha << o.value; say(ss.str()); }
    const bool aerosol = opt(opts, "aerosol-optics"), delta_aerosol = opt(opts, "delta-aerosol");

    // ---- atmosphere: (lay|lev, y, x) in the file = (col, lay|lev) column-major in memory
    say("Reading atmospheric input data from NetCDF.");
    Netcdf_file in("rte_rrtmgp_input.nc", Netcdf_mode::Read);
    const int nx = in.get_dimension_size("x"), ny = in.get_dimension_size("y");
    const int ncol = nx * ny, nlay = in.get_dimension_size("lay"), nlev = in.get_dimension_size("lev");

    const std::vector<Float> p_lay_h = in.get_variable<Float>("p_lay", {nlay, ny, nx});
    const std::vector<Float> p_lev_h = in.get_variable<Float>("p_lev", {nlev, ny, nx});
    Array_gpu<Float,2> p_lay(p_lay_h, {ncol, nlay}), p_lev(p_lev_h, {ncol, nlev});
    Array_gpu<Float,2> t_lay = upload<2>(in, "t_lay", {nlay, ny, nx}, {ncol, nlay});
    Array_gpu<Float,2> t_lev = upload<2>(in, "t_lev", {nlev, ny, nx}, {ncol, nlev});
    Array_gpu<Float,2> col_dry;
    if (in.variable_exists("col_dry")) col_dry = upload<2>(in, "col_dry", {nlay, ny, nx}, {ncol, nlay});

    Gas_concs_gpu gas_concs;
    for (const char* g : {"h2o", "co2", "o3", "n2o", "co", "ch4", "o2", "n2", "ccl4", "cfc11", "cfc12", "cfc22",
                          "hfc143a", "hfc125", "hfc23", "hfc32", "hfc134a", "cf4", "no2"})
        read_and_set_vmr(g, nx, ny, nlay, in, gas_concs);

    Array_gpu<Float,2> lwp, iwp, rel, dei;
    if (cloud)
    {
        lwp = upload<2>(in, "lwp", {nlay, ny, nx}, {ncol, nlay}); iwp = upload<2>(in, "iwp", {nlay, ny, nx}, {ncol, nlay});
        rel = upload<2>(in, "rel", {nlay, ny, nx}, {ncol, nlay}); dei = upload<2>(in, "dei", {nlay, ny, nx}, {ncol, nlay});
    }

    Array_gpu<Float,2> rh;
    Aerosol_concs_gpu aerosol_concs;
    if (aerosol)      // src_test/test_rte_rrtmgp.cpp:274-293
    {
        rh = upload<2>(in, "rh", {nlay, ny, nx}, {ncol, nlay});
        for (int i = 1; i <= 11; ++i)
            read_and_set_aer(std::string("aermr") + (i < 10 ? "0" : "") + std::to_string(i), nx, ny, nlay, in, aerosol_concs);
    }

    // ---- output file
    say("Preparing NetCDF output file.");
    Netcdf_file out("rte_rrtmgp_output.nc", Netcdf_mode::Create);
    out.add_dimension("x", nx); out.add_dimension("y", ny); out.add_dimension("lay", nlay); out.add_dimension("lev", nlev);
    out.add_dimension("pair", 2);
    out.add_variable<Float>("p_lay", {"lay", "y", "x"}).insert(p_lay_h, {0, 0, 0});
    out.add_variable<Float>("p_lev", {"lev", "y", "x"}).insert(p_lev_h, {0, 0, 0});

    if (lw_on)
    {
        say("Initializing the longwave solver.");
        rrtmgp_b200::Kdist_tables kd("coefficients_lw.nc");
        std::unique_ptr<rrtmgp_b200::Cloud_tables> ct;
        if (cloud) ct = std::make_unique<rrtmgp_b200::Cloud_tables>("cloud_coefficients_lw.nc");
        Radiation_solver_longwave rad_lw(gas_concs, kd.file(), ct ? &ct->file() : nullptr);
        const int nbnd = rad_lw.get_n_bnd_gpu(), ngpt = rad_lw.get_n_gpt_gpu();

        Array_gpu<Float,2> emis_sfc = upload<2>(in, "emis_sfc", {ny, nx, nbnd}, {nbnd, ncol});
        Array_gpu<Float,1> t_sfc = upload<1>(in, "t_sfc", {ny, nx}, {ncol});

        Array_gpu<Float,3> lw_tau, lay_source, lev_source; Array_gpu<Float,2> sfc_source;
        Array_gpu<Float,2> lw_flux_up, lw_flux_dn, lw_flux_net;
        Array_gpu<Float,3> lw_bnd_flux_up, lw_bnd_flux_dn, lw_bnd_flux_net;
        if (optical)
        { lw_tau.set_dims({ncol, nlay, ngpt}); lay_source.set_dims({ncol, nlay, ngpt}); lev_source.set_dims({ncol, nlev, ngpt}); sfc_source.set_dims({ncol, ngpt}); }
        if (fluxes)
        {
            lw_flux_up.set_dims({ncol, nlev}); lw_flux_dn.set_dims({ncol, nlev}); lw_flux_net.set_dims({ncol, nlev});
            if (bnd) { lw_bnd_flux_up.set_dims({ncol, nlev, nbnd}); lw_bnd_flux_dn.set_dims({ncol, nlev, nbnd}); lw_bnd_flux_net.set_dims({ncol, nlev, nbnd}); }
        }

        say("Solving the longwave radiation.");
        run_timed(timings, "longwave", [&]()
        {
            rad_lw.solve_gpu(fluxes, cloud, optical, bnd, gas_concs, p_lay, p_lev, t_lay, t_lev, col_dry, t_sfc, emis_sfc,
                             lwp, iwp, rel, dei, lw_tau, lay_source, lev_source, sfc_source,
                             lw_flux_up, lw_flux_dn, lw_flux_net, lw_bnd_flux_up, lw_bnd_flux_dn, lw_bnd_flux_net);
        });

        say("Storing the longwave output.");
        out.add_dimension("gpt_lw", ngpt); out.add_dimension("band_lw", nbnd);
        out.add_variable<Float>("lw_band_lims_wvn", {"band_lw", "pair"}).insert(rad_lw.get_band_lims_wavenumber_gpu().v(), {0, 0});
        if (optical)
        {
            out.add_variable<int>("lw_band_lims_gpt", {"band_lw", "pair"}).insert(rad_lw.get_band_lims_gpoint_gpu().v(), {0, 0});
            write<3>(out, "lw_tau", {"gpt_lw", "lay", "y", "x"}, lw_tau);
            write<3>(out, "lay_source", {"gpt_lw", "lay", "y", "x"}, lay_source);
            write<3>(out, "lev_source", {"gpt_lw", "lev", "y", "x"}, lev_source);
            write<2>(out, "sfc_source", {"gpt_lw", "y", "x"}, sfc_source);
        }
        if (fluxes)
        {
            write<2>(out, "lw_flux_up", {"lev", "y", "x"}, lw_flux_up);
            write<2>(out, "lw_flux_dn", {"lev", "y", "x"}, lw_flux_dn);
            write<2>(out, "lw_flux_net", {"lev", "y", "x"}, lw_flux_net);
            if (bnd)
            {
                write<3>(out, "lw_bnd_flux_up", {"band_lw", "lev", "y", "x"}, lw_bnd_flux_up);
                write<3>(out, "lw_bnd_flux_dn", {"band_lw", "lev", "y", "x"}, lw_bnd_flux_dn);
                write<3>(out, "lw_bnd_flux_net", {"band_lw", "lev", "y", "x"}, lw_bnd_flux_net);
            }
        }
    }

    if (sw_on)
    {
        say("Initializing the shortwave solver.");
        rrtmgp_b200::Kdist_tables kd("coefficients_sw.nc");
        std::unique_ptr<rrtmgp_b200::Cloud_tables> ct;
        if (cloud) ct = std::make_unique<rrtmgp_b200::Cloud_tables>("cloud_coefficients_sw.nc");
        std::unique_ptr<rrtmgp_b200::Aerosol_tables> at;
        if (aerosol) at = std::make_unique<rrtmgp_b200::Aerosol_tables>("aerosol_optics.nc");
        Radiation_solver_shortwave rad_sw(gas_concs, kd.file(), ct ? &ct->file() : nullptr, at ? &at->file() : nullptr);
        const int nbnd = rad_sw.get_n_bnd_gpu(), ngpt = rad_sw.get_n_gpt_gpu();

        Array_gpu<Float,1> mu0 = upload<1>(in, "mu0", {ny, nx}, {ncol});
        Array_gpu<Float,2> sfc_alb_dir = upload<2>(in, "sfc_alb_dir", {ny, nx, nbnd}, {nbnd, ncol});
        Array_gpu<Float,2> sfc_alb_dif = upload<2>(in, "sfc_alb_dif", {ny, nx, nbnd}, {nbnd, ncol});

        // tsi (y, x) / tsi_scaling (scalar) / neither: src_test/test_rte_rrtmgp.cu:561-579
        std::vector<Float> tsi_scaling_h(ncol, Float(1.));
        if (in.variable_exists("tsi"))
        {
            const std::vector<Float> tsi = in.get_variable<Float>("tsi", {ny, nx});
            const Float tsi_ref = rad_sw.get_tsi_gpu();
            for (int i = 0; i < ncol; ++i) tsi_scaling_h[i] = tsi[i] / tsi_ref;
        }
        else if (in.variable_exists("tsi_scaling"))
            tsi_scaling_h.assign(ncol, in.get_variable<Float>("tsi_scaling"));
        Array_gpu<Float,1> tsi_scaling(tsi_scaling_h, {ncol});

        Array_gpu<Float,3> sw_tau, ssa, g; Array_gpu<Float,2> toa_source;
        Array_gpu<Float,2> sw_flux_up, sw_flux_dn, sw_flux_dn_dir, sw_flux_net;
        Array_gpu<Float,3> sw_bnd_flux_up, sw_bnd_flux_dn, sw_bnd_flux_dn_dir, sw_bnd_flux_net;
        if (optical)
        { sw_tau.set_dims({ncol, nlay, ngpt}); ssa.set_dims({ncol, nlay, ngpt}); g.set_dims({ncol, nlay, ngpt}); toa_source.set_dims({ncol, ngpt}); }
        if (fluxes)
        {
            sw_flux_up.set_dims({ncol, nlev}); sw_flux_dn.set_dims({ncol, nlev}); sw_flux_dn_dir.set_dims({ncol, nlev}); sw_flux_net.set_dims({ncol, nlev});
            if (bnd)
            {
                sw_bnd_flux_up.set_dims({ncol, nlev, nbnd}); sw_bnd_flux_dn.set_dims({ncol, nlev, nbnd});
                sw_bnd_flux_dn_dir.set_dims({ncol, nlev, nbnd}); sw_bnd_flux_net.set_dims({ncol, nlev, nbnd});
            }
        }

        say("Solving the shortwave radiation.");
        run_timed(timings, "shortwave", [&]()
        {
            rad_sw.solve_gpu(fluxes, cloud, aerosol, optical, bnd, delta_cloud, delta_aerosol, gas_concs, p_lay, p_lev, t_lay, t_lev, col_dry,
                             sfc_alb_dir, sfc_alb_dif, tsi_scaling, mu0, lwp, iwp, rel, dei, rh, aerosol_concs,
                             sw_tau, ssa, g, toa_source, sw_flux_up, sw_flux_dn, sw_flux_dn_dir, sw_flux_net,
                             sw_bnd_flux_up, sw_bnd_flux_dn, sw_bnd_flux_dn_dir, sw_bnd_flux_net);
        });

        say("Storing the shortwave output.");
        out.add_dimension("gpt_sw", ngpt); out.add_dimension("band_sw", nbnd);
        out.add_variable<Float>("sw_band_lims_wvn", {"band_sw", "pair"}).insert(rad_sw.get_band_lims_wavenumber_gpu().v(), {0, 0});
        if (optical)
        {
            out.add_variable<int>("sw_band_lims_gpt", {"band_sw", "pair"}).insert(rad_sw.get_band_lims_gpoint_gpu().v(), {0, 0});
            write<3>(out, "sw_tau", {"gpt_sw", "lay", "y", "x"}, sw_tau);
            write<3>(out, "ssa", {"gpt_sw", "lay", "y", "x"}, ssa);
            write<3>(out, "g", {"gpt_sw", "lay", "y", "x"}, g);
            write<2>(out, "toa_source", {"gpt_sw", "y", "x"}, toa_source);
        }
        if (fluxes)
        {
            write<2>(out, "sw_flux_up", {"lev", "y", "x"}, sw_flux_up);
            write<2>(out, "sw_flux_dn", {"lev", "y", "x"}, sw_flux_dn);
            write<2>(out, "sw_flux_dn_dir", {"lev", "y", "x"}, sw_flux_dn_dir);
            write<2>(out, "sw_flux_net", {"lev", "y", "x"}, sw_flux_net);
            if (bnd)
            {
                write<3>(out, "sw_bnd_flux_up", {"band_sw", "lev", "y", "x"}, sw_bnd_flux_up);
                write<3>(out, "sw_bnd_flux_dn", {"band_sw", "lev", "y", "x"}, sw_bnd_flux_dn);
                write<3>(out, "sw_bnd_flux_dn_dir", {"band_sw", "lev", "y", "x"}, sw_bnd_flux_dn_dir);
                write<3>(out, "sw_bnd_flux_net", {"band_sw", "lev", "y", "x"}, sw_bnd_flux_net);
            }
        }
    }
    out.sync();
    say("###### Finished RTE+RRTMGP solver ######");
}

int main(int argc, char** argv)
{
    try { solve_radiation(argc, argv); }
    catch (const std::exception& e) { say("EXCEPTION: " + std::string(e.what())); return 1; }
    catch (...) { say("UNHANDLED EXCEPTION!"); return 1; }
    return 0;
}
