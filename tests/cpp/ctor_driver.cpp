// Test driver for the set-up half of the drop-in boundary (include/rrtmgp_b200/{array,classes}.hpp):
//   * host Array<T,N> / Array_gpu<T,N> semantics of the reference (include/Array.h:94-297, 355-650);
//   * the reference's positional constructors: Gas_optics_rrtmgp_gpu (32 / 35 arguments,
//     include/Gas_optics_rrtmgp.h:293-364), Cloud_optics_gpu (include/Cloud_optics.h:84-89), Aerosol_optics_gpu
//     (include/Aerosol_optics.h:53-56), fed the way load_and_init_* feeds them (src_test/Radiation_solver.cu:93-402):
//     every file variable as an Array whose dims are the file's reversed.
// Test infrastructure only; not part of the shipped library.
#include <cstring>
#include <sstream>
#include "rrtmgp_b200/classes.hpp"

namespace
{
    std::string g_msg;
    int fail(const std::string& m) { g_msg = m; return 1; }
#define REQUIRE(...) do { if (!(__VA_ARGS__)) { std::ostringstream o_; o_ << "line " << __LINE__ << ": " #__VA_ARGS__; return fail(o_.str()); } } while (0)

    std::vector<std::string> strings(const char* block, int n, int len)
    {
        std::vector<std::string> out;
        for (int i = 0; i < n; ++i)
        {
            std::string s(block + (size_t)i * len, len);
            const size_t e = s.find_last_not_of(" \0", std::string::npos, 2);
            out.push_back(e == std::string::npos ? "" : s.substr(0, e + 1));
        }
        return out;
    }
    template<int N> Array<Float,N> farr(const double* p, const std::array<int,N>& dims)
    {
        size_t n = 1; for (int d : dims) n *= (size_t)d;
        std::vector<Float> v(n);
        for (size_t i = 0; i < n; ++i) v[i] = (Float)p[i];
        return Array<Float,N>(std::move(v), dims);
    }
    template<int N> Array<int,N> iarr(const int* p, const std::array<int,N>& dims)
    {
        size_t n = 1; for (int d : dims) n *= (size_t)d;
        return Array<int,N>(std::vector<int>(p, p + n), dims);
    }
    Array<Bool,1> barr(const int* p, int n)
    {
        std::vector<Bool> v(n);
        for (int i = 0; i < n; ++i) v[i] = (Bool)(p[i] != 0);
        return Array<Bool,1>(std::move(v), {n});
    }
    Array<std::string,1> sarr(const char* block, int n, int len) { return Array<std::string,1>(strings(block, n, len), {n}); }

    // the argument list of load_and_init_gas_optics (src_test/Radiation_solver.cu:125-318), variables from the struct
    Gas_optics_rrtmgp_gpu positional_kdist(const Gas_concs_gpu& gas_concs, const rrtmgp_b200_kdist_file* f)
    {
        const int L = f->string_len;
        Array<std::string,1> gas_names = sarr(f->gas_names, f->n_absorbers, L);
        Array<int,3> key_species = iarr<3>(f->key_species, {2, f->n_layers, f->n_bnds});
        Array<Float,2> band_lims = farr<2>(f->bnd_limits_wavenumber, {2, f->n_bnds});
        Array<int,2> band2gpt = iarr<2>(f->bnd_limits_gpt, {2, f->n_bnds});
        Array<Float,1> press_ref = farr<1>(f->press_ref, {f->n_press});
        Array<Float,1> temp_ref = farr<1>(f->temp_ref, {f->n_temps});
        const Float temp_ref_p = (Float)f->absorption_coefficient_ref_P, temp_ref_t = (Float)f->absorption_coefficient_ref_T;
        const Float press_ref_trop = (Float)f->press_ref_trop;
        Array<Float,3> kminor_lower = farr<3>(f->kminor_lower, {f->n_contrib_lower, f->n_mixingfracs, f->n_temps});
        Array<Float,3> kminor_upper = farr<3>(f->kminor_upper, {f->n_contrib_upper, f->n_mixingfracs, f->n_temps});
        Array<std::string,1> gas_minor = sarr(f->gas_minor, f->n_minorabsorbers, L);
        Array<std::string,1> identifier_minor = sarr(f->identifier_minor, f->n_minorabsorbers, L);
        Array<std::string,1> minor_gases_lower = sarr(f->minor_gases_lower, f->n_minor_lower, L);
        Array<std::string,1> minor_gases_upper = sarr(f->minor_gases_upper, f->n_minor_upper, L);
        Array<int,2> minor_limits_gpt_lower = iarr<2>(f->minor_limits_gpt_lower, {f->n_pairs, f->n_minor_lower});
        Array<int,2> minor_limits_gpt_upper = iarr<2>(f->minor_limits_gpt_upper, {f->n_pairs, f->n_minor_upper});
        Array<Bool,1> minor_scales_with_density_lower = barr(f->minor_scales_with_density_lower, f->n_minor_lower);
        Array<Bool,1> minor_scales_with_density_upper = barr(f->minor_scales_with_density_upper, f->n_minor_upper);
        Array<Bool,1> scale_by_complement_lower = barr(f->scale_by_complement_lower, f->n_minor_lower);
        Array<Bool,1> scale_by_complement_upper = barr(f->scale_by_complement_upper, f->n_minor_upper);
        Array<std::string,1> scaling_gas_lower = sarr(f->scaling_gas_lower, f->n_minor_lower, L);
        Array<std::string,1> scaling_gas_upper = sarr(f->scaling_gas_upper, f->n_minor_upper, L);
        Array<int,1> kminor_start_lower = iarr<1>(f->kminor_start_lower, {f->n_minor_lower});
        Array<int,1> kminor_start_upper = iarr<1>(f->kminor_start_upper, {f->n_minor_upper});
        Array<Float,3> vmr_ref = farr<3>(f->vmr_ref, {f->n_layers, f->n_extabsorbers, f->n_temps});
        Array<Float,4> kmajor = farr<4>(f->kmajor, {f->n_gpts, f->n_mixingfracs, f->n_press + 1, f->n_temps});
        Array<Float,3> rayl_lower, rayl_upper;
        if (f->rayl_lower)
        {
            rayl_lower = farr<3>(f->rayl_lower, {f->n_gpts, f->n_mixingfracs, f->n_temps});
            rayl_upper = farr<3>(f->rayl_upper, {f->n_gpts, f->n_mixingfracs, f->n_temps});
        }
        if (f->totplnk)
        {
            Array<Float,2> totplnk = farr<2>(f->totplnk, {f->n_planck_temps, f->n_bnds});
            Array<Float,4> planck_frac = farr<4>(f->plank_fraction, {f->n_gpts, f->n_mixingfracs, f->n_press + 1, f->n_temps});
            return Gas_optics_rrtmgp_gpu(
                    gas_concs, gas_names, key_species, band2gpt, band_lims, press_ref, press_ref_trop, temp_ref, temp_ref_p, temp_ref_t,
                    vmr_ref, kmajor, kminor_lower, kminor_upper, gas_minor, identifier_minor, minor_gases_lower, minor_gases_upper,
                    minor_limits_gpt_lower, minor_limits_gpt_upper, minor_scales_with_density_lower, minor_scales_with_density_upper,
                    scaling_gas_lower, scaling_gas_upper, scale_by_complement_lower, scale_by_complement_upper,
                    kminor_start_lower, kminor_start_upper, totplnk, planck_frac, rayl_lower, rayl_upper);
        }
        Array<Float,1> solar_src_quiet = farr<1>(f->solar_source_quiet, {f->n_gpts});
        Array<Float,1> solar_src_facular = farr<1>(f->solar_source_facular, {f->n_gpts});
        Array<Float,1> solar_src_sunspot = farr<1>(f->solar_source_sunspot, {f->n_gpts});
        return Gas_optics_rrtmgp_gpu(
                gas_concs, gas_names, key_species, band2gpt, band_lims, press_ref, press_ref_trop, temp_ref, temp_ref_p, temp_ref_t,
                vmr_ref, kmajor, kminor_lower, kminor_upper, gas_minor, identifier_minor, minor_gases_lower, minor_gases_upper,
                minor_limits_gpt_lower, minor_limits_gpt_upper, minor_scales_with_density_lower, minor_scales_with_density_upper,
                scaling_gas_lower, scaling_gas_upper, scale_by_complement_lower, scale_by_complement_upper,
                kminor_start_lower, kminor_start_upper, solar_src_quiet, solar_src_facular, solar_src_sunspot,
                (Float)f->tsi_default, (Float)f->mg_default, (Float)f->sb_default, rayl_lower, rayl_upper);
    }
    // load_and_init_cloud_optics (src_test/Radiation_solver.cu:321-364)
    Cloud_optics_gpu positional_cloud(const rrtmgp_b200_cloud_file* c)
    {
        Array<Float,2> band_lims_wvn = farr<2>(c->bnd_limits_wavenumber, {2, c->n_band});
        Array<Float,2> lut_extliq = farr<2>(c->lut_extliq, {c->n_size_liq, c->n_band});
        Array<Float,2> lut_ssaliq = farr<2>(c->lut_ssaliq, {c->n_size_liq, c->n_band});
        Array<Float,2> lut_asyliq = farr<2>(c->lut_asyliq, {c->n_size_liq, c->n_band});
        Array<Float,3> lut_extice = farr<3>(c->lut_extice, {c->n_size_ice, c->n_band, c->n_rghice});
        Array<Float,3> lut_ssaice = farr<3>(c->lut_ssaice, {c->n_size_ice, c->n_band, c->n_rghice});
        Array<Float,3> lut_asyice = farr<3>(c->lut_asyice, {c->n_size_ice, c->n_band, c->n_rghice});
        return Cloud_optics_gpu(band_lims_wvn, (Float)c->radliq_lwr, (Float)c->radliq_upr, (Float)c->radliq_fac,
                                (Float)c->diamice_lwr, (Float)c->diamice_upr, (Float)c->diamice_fac,
                                lut_extliq, lut_ssaliq, lut_asyliq, lut_extice, lut_ssaice, lut_asyice);
    }
    // load_and_init_aerosol_optics (src_test/Radiation_solver.cu:366-402)
    Aerosol_optics_gpu positional_aerosol(const rrtmgp_b200_aerosol_file* a)
    {
        Array<Float,2> band_lims_wvn({2, a->n_band});
        Array<Float,2> mext_phobic = farr<2>(a->mass_ext_sw_hydrophobic, {a->n_band, a->n_phobic});
        Array<Float,2> ssa_phobic = farr<2>(a->ssa_sw_hydrophobic, {a->n_band, a->n_phobic});
        Array<Float,2> g_phobic = farr<2>(a->asymmetry_sw_hydrophobic, {a->n_band, a->n_phobic});
        Array<Float,3> mext_philic = farr<3>(a->mass_ext_sw_hydrophilic, {a->n_band, a->n_hum, a->n_philic});
        Array<Float,3> ssa_philic = farr<3>(a->ssa_sw_hydrophilic, {a->n_band, a->n_hum, a->n_philic});
        Array<Float,3> g_philic = farr<3>(a->asymmetry_sw_hydrophilic, {a->n_band, a->n_hum, a->n_philic});
        Array<Float,1> rh_upper = farr<1>(a->relative_humidity2, {a->n_hum});
        return Aerosol_optics_gpu(band_lims_wvn, rh_upper, mext_phobic, ssa_phobic, g_phobic, mext_philic, ssa_philic, g_philic);
    }
    // available gases without touching the device: (1,1) views on no memory (only the names are read at construction)
    Gas_concs_gpu names_only(int n, const char* names, int len)
    {
        Gas_concs_gpu g;
        for (auto& s : strings(names, n, len)) g.set_vmr(s, Array_gpu<Float,2>(nullptr, {1, 1}));
        return g;
    }
    template<int N> Array_gpu<Float,N> up(const double* h, const std::array<int,N>& dims) { return Array_gpu<Float,N>(farr<N>(h, dims)); }
    template<int N> void down(const Array_gpu<Float,N>& a, double* h)
    {
        if (!h) return;
        const Array<Float,N> v(a);
        for (int i = 0; i < v.size(); ++i) h[i] = (double)v.v()[i];
    }
}

extern "C" const char* ctor_last_message() { return g_msg.c_str(); }
extern "C" int ctor_float_bytes() { return (int)sizeof(Float); }

// host Array<T,N>: the reference's semantics, no device needed
extern "C" int ctor_array_host_selftest()
{
    try
    {
        Array<int,3> a({4, 3, 2});
        REQUIRE(a.size() == 24 && a.dim(1) == 4 && a.dim(2) == 3 && a.dim(3) == 2 && !a.is_empty());
        REQUIRE(a.min() == 0 && a.max() == 0);                                   // zeros
        for (int k = 1; k <= 2; ++k) for (int j = 1; j <= 3; ++j) for (int i = 1; i <= 4; ++i) a({i, j, k}) = 100 * k + 10 * j + i;
        REQUIRE(a.v()[0] == 111 && a.v()[1] == 112 && a.v()[4] == 121 && a.v()[12] == 211);      // column-major, 1-based
        REQUIRE(a.ptr() == a.v().data());
        const Array<int,3> s = a.subset({{ {2, 3}, {1, 3}, {2, 2} }});
        REQUIRE(s.dim(1) == 2 && s.dim(2) == 3 && s.dim(3) == 1 && s({1, 1, 1}) == 212 && s({2, 3, 1}) == 233);
        Array<int,2> b(std::vector<int>{7, 8, 9}, {1, 3});                       // extent-1 dimension is broadcast by subset
        const Array<int,2> bs = b.subset({{ {5, 8}, {2, 3} }});
        REQUIRE(bs.dim(1) == 4 && bs.dim(2) == 2 && bs({1, 1}) == 8 && bs({4, 1}) == 8 && bs({3, 2}) == 9);
        Array<double,3> g({2, 2, 3});                                            // col_gas-style zero-based last index
        g.set_offsets({0, 0, -1});
        g({1, 1, 0}) = 5.; g({2, 2, 2}) = 6.;
        REQUIRE(g.v()[0] == 5. && g.v()[11] == 6.);
        Array<int,3> m(std::move(a));
        REQUIRE(a.size() == 0 && a.is_empty() && m.size() == 24 && m({4, 3, 2}) == 234);
        Array<int,1> e;
        REQUIRE(e.is_empty());
        e.set_dims({5});
        REQUIRE(e.size() == 5);
        bool threw = false;
        try { e.set_dims({6}); } catch (const std::runtime_error&) { threw = true; }
        REQUIRE(threw);
        e.expand_dims({8});
        REQUIRE(e.size() == 8 && e.dim(1) == 8);
        threw = false;
        try { e.expand_dims({3}); } catch (const std::runtime_error&) { threw = true; }
        REQUIRE(threw);
        e.fill(3);
        REQUIRE(e.min() == 3 && e.max() == 3);
        Array<std::string,1> names(std::vector<std::string>{"h2o", "co2"}, {2});
        REQUIRE(names({2}) == "co2");
        Array<Float,1> r(std::vector<Float>{Float(0.1), Float(0.5), Float(1.5)}, {3});
        REQUIRE(any_vals_outside(r, Float(0.), Float(1.)) && !any_vals_less_than(r, Float(0.)));
        // Gas_concs: scalar / profile / field, range check, column subset
        Gas_concs gc;
        gc.set_vmr("co2", Float(4e-4));
        gc.set_vmr("o3", Array<Float,1>(std::vector<Float>{Float(1e-6), Float(2e-6)}, {2}));
        Array<Float,2> h2o({3, 2});
        for (int j = 1; j <= 2; ++j) for (int i = 1; i <= 3; ++i) h2o({i, j}) = Float(0.001) * Float(10 * j + i);
        gc.set_vmr("h2o", h2o);
        REQUIRE(gc.exists("co2") && !gc.exists("ch4") && gc.get_vmr("o3").dim(1) == 1 && gc.get_vmr("o3").dim(2) == 2);
        const Gas_concs sub(gc, 2, 2);
        REQUIRE(sub.get_vmr("h2o").dim(1) == 2 && sub.get_vmr("h2o")({1, 2}) == h2o({2, 2}) && sub.get_vmr("co2").size() == 1);
        threw = false;
        try { gc.set_vmr("bad", Float(1.5)); } catch (const std::range_error&) { threw = true; }
        REQUIRE(threw);
        return 0;
    }
    catch (const std::exception& ex) { return fail(std::string("exception: ") + ex.what()); }
}

// Array_gpu<T,N>: conversions, views, subset kernel against the host subset, fill, element access (needs a GPU)
extern "C" int ctor_array_gpu_selftest()
{
    try
    {
        Array<Float,3> h({7, 5, 3});
        for (int i = 0; i < h.size(); ++i) h.v()[i] = Float(i) * Float(0.25);
        const Array_gpu<Float,3> d(h);
        REQUIRE(d.size() == 105 && d.dim(2) == 5);
        const Array<Float,3> back(d);
        REQUIRE(back.v() == h.v());
        REQUIRE(d({7, 5, 3}) == h({7, 5, 3}) && d({2, 1, 1}) == h({2, 1, 1}));        // blocking element read
        const std::array<std::pair<int,int>,3> win = {{ {3, 6}, {2, 5}, {1, 2} }};
        REQUIRE(Array<Float,3>(d.subset(win)).v() == h.subset(win).v());
        Array<Float,2> row(std::vector<Float>{Float(1), Float(2), Float(3), Float(4)}, {1, 4});      // broadcast dimension
        const std::array<std::pair<int,int>,2> bw = {{ {10, 14}, {2, 3} }};
        REQUIRE(Array<Float,2>(Array_gpu<Float,2>(row).subset(bw)).v() == row.subset(bw).v());
        Array<signed char,2> flags({6, 4});
        for (int i = 0; i < flags.size(); ++i) flags.v()[i] = (signed char)(i % 3);
        const std::array<std::pair<int,int>,2> fw = {{ {2, 5}, {2, 4} }};
        REQUIRE((Array<signed char,2>(Array_gpu<signed char,2>(flags).subset(fw)).v() == flags.subset(fw).v()));
        Array_gpu<Float,2> f({9, 4});
        f.fill(Float(2.5));
        const Array<Float,2> fh(f);
        REQUIRE(fh.min() == Float(2.5) && fh.max() == Float(2.5));
        f.insert({3, 2}, Float(-1.));
        REQUIRE(f({3, 2}) == Float(-1.) && f({4, 2}) == Float(2.5));
        Array_gpu<Float,2> view(f.ptr(), {9, 4});                                        // a view shares memory and never frees it
        { Array_gpu<Float,2> copy_of_view(view); REQUIRE(copy_of_view.ptr() == f.ptr()); }
        REQUIRE(f({3, 2}) == Float(-1.));
        Array_gpu<Float,2> deep(f);
        REQUIRE(deep.ptr() != f.ptr() && deep({3, 2}) == Float(-1.));
        Array_gpu<Float,2> moved(std::move(deep));
        REQUIRE(deep.size() == 0 && deep.ptr() == nullptr && moved({3, 2}) == Float(-1.));
        bool threw = false;
        try { Array_gpu<Float,2> small({2, 2}); small = f; } catch (const std::runtime_error&) { threw = true; }
        REQUIRE(threw);                                                                  // "Initialised arrays can not be resized"
        Array_gpu<Float,2> lazy;
        lazy = f;
        REQUIRE(lazy.dim(1) == 9 && lazy({3, 2}) == Float(-1.));
        // Gas_concs_gpu: upload + column subset (what the reference does per column block, src_test/Radiation_solver.cu:476)
        Gas_concs gc;
        gc.set_vmr("co2", Float(4e-4));
        Array<Float,2> q({6, 3});
        for (int i = 0; i < q.size(); ++i) q.v()[i] = Float(1e-3) * Float(i);
        gc.set_vmr("h2o", q);
        const Gas_concs_gpu gg(gc);
        const Gas_concs_gpu gs(gg, 3, 2);
        REQUIRE(gs.get_vmr("h2o").dim(1) == 2 && gs.get_vmr("h2o").dim(2) == 3 && gs.get_vmr("co2").size() == 1);
        REQUIRE((Array<Float,2>(gs.get_vmr("h2o")).v() == q.subset({{ {3, 4}, {1, 3} }}).v()));
        return 0;
    }
    catch (const std::exception& ex) { return fail(std::string("exception: ") + ex.what()); }
}

// positional constructor vs struct constructor: every prepared table bit-identical.  host_only != 0 runs without a GPU.
extern "C" int ctor_tables_equal(const rrtmgp_b200_kdist_file* kd, int n_avail, const char* avail, int len, int host_only)
{
    try
    {
        rrtmgp_b200::create_flags() = host_only ? RRTMGP_B200_HOST_ONLY : 0;
        const Gas_concs_gpu gases = names_only(n_avail, avail, len);
        const Gas_optics_rrtmgp_gpu a(gases, *kd);
        const Gas_optics_rrtmgp_gpu b = positional_kdist(gases, kd);
        rrtmgp_b200::create_flags() = 0;
        unsigned long long ha = 0, hb = 1;
        rrtmgp_b200::check(rrtmgp_b200_gas_optics_table_hash(a.handle(), &ha));
        rrtmgp_b200::check(rrtmgp_b200_gas_optics_table_hash(b.handle(), &hb));
        REQUIRE(ha == hb);
        REQUIRE(a.get_ngpt() == b.get_ngpt() && a.get_nband() == b.get_nband() && a.get_ngpt() == kd->n_gpts);
        REQUIRE(a.get_band_lims_gpoint().v() == b.get_band_lims_gpoint().v());
        REQUIRE(a.get_band_lims_wavenumber().v() == b.get_band_lims_wavenumber().v());
        REQUIRE(a.get_gpoint_bands().v() == b.get_gpoint_bands().v());
        REQUIRE(a.get_gas_names() == b.get_gas_names());
        REQUIRE(a.get_press_ref_min() == b.get_press_ref_min() && a.get_temp_max() == b.get_temp_max() && a.get_tsi() == b.get_tsi());
        REQUIRE(b.get_band_lims_gpoint()({2, b.get_nband()}) == b.get_ngpt() && b.get_gpoint_bands()({1}) == 1);
        return 0;
    }
    catch (const std::exception& ex) { rrtmgp_b200::create_flags() = 0; return fail(std::string("exception: ") + ex.what()); }
}

// Fluxes of one all-sky call through solvers whose optics objects came from the positional constructors; the caller
// compares them with the struct-constructed solvers' (tests/test_classes_cpp.py): must be bit-identical.
extern "C" int ctor_solve(const rrtmgp_b200_kdist_file* kd_lw, const rrtmgp_b200_kdist_file* kd_sw,
        const rrtmgp_b200_cloud_file* cloud_lw, const rrtmgp_b200_cloud_file* cloud_sw, const rrtmgp_b200_aerosol_file* aerosol,
        int positional, int ncol, int nlay, int ngas, const char* names, const double* const* vmr, const int* vnc, const int* vnl,
        const double* play, const double* plev, const double* tlay, const double* tlev, const double* tsfc, const double* emis,
        const double* alb_dir, const double* alb_dif, const double* mu0,
        const double* lwp, const double* iwp, const double* rel, const double* dei, const double* rh, const double* const* aermr,
        double* lw_up, double* lw_dn, double* sw_up, double* sw_dn, double* sw_dir)
{
    try
    {
        Gas_concs gc_host;                                      // host concentrations -> device, like the reference driver
        const auto nm = strings(names, ngas, 32);
        for (int i = 0; i < ngas; ++i) gc_host.set_vmr(nm[i], farr<2>(vmr[i], {vnc[i], vnl[i]}));
        const Gas_concs_gpu gas_concs(gc_host);
        const int nlev = nlay + 1;
        std::unique_ptr<Radiation_solver_longwave> rad_lw;
        std::unique_ptr<Radiation_solver_shortwave> rad_sw;
        if (positional)
        {
            rad_lw = std::make_unique<Radiation_solver_longwave>(positional_kdist(gas_concs, kd_lw), positional_cloud(cloud_lw));
            rad_sw = std::make_unique<Radiation_solver_shortwave>(positional_kdist(gas_concs, kd_sw), positional_cloud(cloud_sw), positional_aerosol(aerosol));
        }
        else
        {
            rad_lw = std::make_unique<Radiation_solver_longwave>(gas_concs, *kd_lw, cloud_lw);
            rad_sw = std::make_unique<Radiation_solver_shortwave>(gas_concs, *kd_sw, cloud_sw, aerosol);
        }
        const int nbnd_lw = rad_lw->get_n_bnd_gpu(), nbnd_sw = rad_sw->get_n_bnd_gpu();
        auto p_lay = up<2>(play, {ncol, nlay}); auto p_lev = up<2>(plev, {ncol, nlev});
        auto t_lay = up<2>(tlay, {ncol, nlay}); auto t_lev = up<2>(tlev, {ncol, nlev});
        auto t_sfc = up<1>(tsfc, {ncol}); auto emis_sfc = up<2>(emis, {nbnd_lw, ncol});
        auto a_dir = up<2>(alb_dir, {nbnd_sw, ncol}); auto a_dif = up<2>(alb_dif, {nbnd_sw, ncol}); auto a_mu0 = up<1>(mu0, {ncol});
        auto a_lwp = up<2>(lwp, {ncol, nlay}); auto a_iwp = up<2>(iwp, {ncol, nlay}); auto a_rel = up<2>(rel, {ncol, nlay}); auto a_dei = up<2>(dei, {ncol, nlay});
        auto a_rh = up<2>(rh, {ncol, nlay});
        Aerosol_concs_gpu aer;
        for (int i = 0; i < 11; ++i)
            aer.set_vmr(std::string("aermr") + (i < 9 ? "0" : "") + std::to_string(i + 1), up<2>(aermr[i], {ncol, nlay}));
        Array_gpu<Float,2> col_dry({ncol, nlay});
        Gas_optics_rrtmgp_gpu::get_col_dry(col_dry, gas_concs.get_vmr("h2o"), p_lev);
        Array_gpu<Float,3> o3a, o3b, o3c; Array_gpu<Float,2> o2a;
        Array_gpu<Float,2> f_up({ncol, nlev}), f_dn({ncol, nlev}), f_dir({ncol, nlev}), f_net({ncol, nlev});
        Array_gpu<Float,3> b0, b1, b2, b3;
        rad_lw->solve_gpu(true, true, false, false, gas_concs, p_lay, p_lev, t_lay, t_lev, col_dry, t_sfc, emis_sfc,
                          a_lwp, a_iwp, a_rel, a_dei, o3a, o3b, o3c, o2a, f_up, f_dn, f_net, b0, b1, b2);
        down(f_up, lw_up); down(f_dn, lw_dn);
        rad_sw->solve_gpu(true, true, true, false, false, true, false, gas_concs, p_lay, p_lev, t_lay, t_lev, col_dry, a_dir, a_dif,
                          Array_gpu<Float,1>(), a_mu0, a_lwp, a_iwp, a_rel, a_dei, a_rh, aer, o3a, o3b, o3c, o2a,
                          f_up, f_dn, f_dir, f_net, b0, b1, b2, b3);
        down(f_up, sw_up); down(f_dn, sw_dn); down(f_dir, sw_dir);
        // the same two calls prepared and launched together (rrtmgp_b200::solve_gpu_pair): bit-identical fluxes
        {
            Array_gpu<Float,2> l_up({ncol, nlev}), l_dn({ncol, nlev}), l_net({ncol, nlev});
            Array_gpu<Float,2> s_up({ncol, nlev}), s_dn({ncol, nlev}), s_dir({ncol, nlev}), s_net({ncol, nlev});
            const rrtmgp_b200::Fused_call clw = rad_lw->prepare_gpu(true, false, gas_concs, p_lay, p_lev, t_lay, t_lev, col_dry, t_sfc, emis_sfc,
                    a_lwp, a_iwp, a_rel, a_dei, l_up, l_dn, l_net, b0, b1, b2);
            const rrtmgp_b200::Fused_call csw = rad_sw->prepare_gpu(true, true, false, true, false, gas_concs, p_lay, p_lev, t_lay, t_lev, col_dry,
                    a_dir, a_dif, Array_gpu<Float,1>(), a_mu0, a_lwp, a_iwp, a_rel, a_dei, a_rh, aer, s_up, s_dn, s_dir, s_net, b0, b1, b2, b3);
            rrtmgp_b200::solve_gpu_pair(clw, csw);
            const size_t n = (size_t)ncol * nlev;
            std::vector<double> h(n);
            const std::pair<const Array_gpu<Float,2>*, const double*> chk[5] = {{&l_up, lw_up}, {&l_dn, lw_dn}, {&s_up, sw_up}, {&s_dn, sw_dn}, {&s_dir, sw_dir}};
            for (int k = 0; k < 5; ++k)
            {
                down(*chk[k].first, h.data());
                for (size_t i = 0; i < n; ++i)
                    if (h[i] != chk[k].second[i]) return fail("solve_gpu_pair differs from the two solve_gpu calls (field " + std::to_string(k) + ")");
            }
        }
        return 0;
    }
    catch (const std::exception& ex) { return fail(std::string("exception: ") + ex.what()); }
}
