// classes.hpp — header-only C++ mirror of the reference's L3 `_gpu` class API over the C ABI
// (include/rrtmgp_b200.h).  Same class and method names, argument order and error behaviour as
// the reference (paths relative to the reference root):
//   Array_gpu<T,N>            include/Array.h:355-650          (1-based dims, owning or view)
//   Gas_concs_gpu             include/Gas_concs.h:73-91
//   Optical_props[_1scl|_2str]_gpu, add_to   include/Optical_props.h:165-293
//   Source_func_lw_gpu        include/Source_functions.h:70-93
//   Gas_optics_rrtmgp_gpu     include/Gas_optics_rrtmgp.h:293-408
//   Cloud_optics_gpu          include/Cloud_optics.h
//   Aerosol_optics_gpu, Aerosol_concs_gpu   include/Aerosol_optics.h:20,48-84
//   Rte_lw_gpu / Rte_sw_gpu   include/Rte_lw.h:62-77, include/Rte_sw.h:62-80
//   Fluxes_broadband_gpu / Fluxes_byband_gpu   include/Fluxes.h:127-210
// Precision follows the reference switch: -DRTE_USE_SP -> Float = float (include/types.h:19-25).
// All work is enqueued on `rrtmgp_b200::stream()` (default: the legacy default stream, like the
// reference); nothing here computes on the host.  Link: -lrrtmgp_b200 -lcudart.
#pragma once
#include <cuda_runtime.h>
#include <array>
#include <deque>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include "../rrtmgp_b200.h"

#ifdef RTE_USE_SP
using Float = float;
#else
using Float = double;
#endif
using Bool = signed char;

#include "array.hpp"       // Array<T,N>, Array_gpu<T,N>, rrtmgp_b200::stream() / check()

namespace rrtmgp_b200 {
constexpr int dtype = sizeof(Float) == 4 ? RRTMGP_B200_F32 : RRTMGP_B200_F64;
// OR-ed into dtype when a k-distribution object is constructed.  RRTMGP_B200_HOST_ONLY builds the host tables only,
// so constructor logic can be checked on a machine without a GPU (compute calls on such an object fail).
inline int& create_flags() { static int f = 0; return f; }
}  // namespace rrtmgp_b200

// include/Gas_concs.h:40-70, src/Gas_concs.cpp:28-117 (host volume mixing ratios: scalar (1,1), profile (1,nlay), field (ncol,nlay))
class Gas_concs
{
    public:
        Gas_concs() = default;
        Gas_concs(const Gas_concs& gas_concs_ref, const int start, const int size)
        {
            for (auto& g : gas_concs_ref.gas_concs_map)
                gas_concs_map[g.first] = g.second.dim(1) == 1 ? g.second
                                       : g.second.subset({{ {start, start + size - 1}, {1, g.second.dim(2)} }});
        }
        void set_vmr(const std::string& name, const Float data)
        {
            if (data < Float(0.) || data > Float(1.)) throw std::range_error("Gas concentration " + name + " is out of range");
            gas_concs_map[name] = Array<Float,2>(std::vector<Float>{data}, {1, 1});
        }
        void set_vmr(const std::string& name, const Array<Float,1>& data)
        {
            if (any_vals_outside(data, Float(0.), Float(1.))) throw std::range_error("Gas concentration " + name + " is out of range");
            gas_concs_map[name] = Array<Float,2>(data.v(), {1, data.dim(1)});
        }
        void set_vmr(const std::string& name, const Array<Float,2>& data)
        {
            if (any_vals_outside(data, Float(0.), Float(1.))) throw std::range_error("Gas concentration " + name + " is out of range");
            gas_concs_map[name] = data;
        }
        const Array<Float,2>& get_vmr(const std::string& name) const { return gas_concs_map.at(name); }
        Bool exists(const std::string& name) const { return gas_concs_map.count(name) != 0; }
    private:
        std::map<std::string, Array<Float,2>> gas_concs_map;
        friend class Gas_concs_gpu;
};

// include/Gas_concs.h:73-91, src_cuda/Gas_concs.cu:29-80
class Gas_concs_gpu
{
    public:
        Gas_concs_gpu() = default;
        Gas_concs_gpu(const Gas_concs& gas_concs_ref)
        {
            for (auto& g : gas_concs_ref.gas_concs_map) gas_concs_map.emplace(g.first, Array_gpu<Float,2>(g.second));
        }
        // columns [start, start+size-1] (1-based) of every field; scalars and profiles are shared
        Gas_concs_gpu(const Gas_concs_gpu& gas_concs_ref, const int start, const int size)
        {
            for (auto& g : gas_concs_ref.gas_concs_map)
            {
                if (g.second.dim(1) == 1) gas_concs_map.emplace(g.first, g.second);
                else gas_concs_map.emplace(g.first, g.second.subset({{ {start, start + size - 1}, {1, g.second.dim(2)} }}));
            }
        }
        void set_vmr(const std::string& name, const Array<Float,2>& data)
        {
            if (any_vals_outside(data, Float(0.), Float(1.))) throw std::range_error("Gas concentration " + name + " is out of range");
            gas_concs_map.erase(name);
            gas_concs_map.emplace(name, Array_gpu<Float,2>(data));
        }
        void set_vmr(const std::string& name, const Array_gpu<Float,2>& data) { gas_concs_map.erase(name); gas_concs_map.emplace(name, data); }   // caller validated
        // extensions (not in the reference): scalar and flat-vector setters
        void set_vmr(const std::string& name, const Float data)
        {
            if (data < Float(0.) || data > Float(1.)) throw std::range_error("Gas concentration " + name + " is out of range");
            set_vmr(name, Array_gpu<Float,2>(std::vector<Float>{data}, {1, 1}));
        }
        void set_vmr(const std::string& name, const std::vector<Float>& data, const int ncol, const int nlay)
        {
            for (const Float v : data)
                if (v < Float(0.) || v > Float(1.)) throw std::range_error("Gas concentration " + name + " is out of range");
            set_vmr(name, Array_gpu<Float,2>(data, {ncol, nlay}));
        }
        const Array_gpu<Float,2>& get_vmr(const std::string& name) const
        {
            auto it = gas_concs_map.find(name);
            if (it == gas_concs_map.end()) throw std::runtime_error("Gas_concs: gas " + name + " not set");
            return it->second;
        }
        Bool exists(const std::string& name) const { return gas_concs_map.count(name) != 0; }
        std::vector<std::string> names() const { std::vector<std::string> n; for (auto& kv : gas_concs_map) n.push_back(kv.first); return n; }
    private:
        std::map<std::string, Array_gpu<Float,2>> gas_concs_map;
};

// band <-> g-point bookkeeping (include/Optical_props.h:178-210, src/Optical_props.cpp:31-72); limits are (2, nbnd), 1-based
class Optical_props_gpu
{
    public:
        Optical_props_gpu() = default;
        Optical_props_gpu(const Array<Float,2>& band_lims_wvn, const Array<int,2>& band_lims_gpt)
            : band2gpt(band_lims_gpt), band_lims_wvn(band_lims_wvn) { init(); }
        // one g-point per band
        Optical_props_gpu(const Array<Float,2>& band_lims_wvn) : band2gpt({2, band_lims_wvn.dim(2)}), band_lims_wvn(band_lims_wvn)
        {
            for (int b = 1; b <= band_lims_wvn.dim(2); ++b) { band2gpt({1, b}) = b; band2gpt({2, b}) = b; }
            init();
        }
        virtual ~Optical_props_gpu() = default;
        Optical_props_gpu(const Optical_props_gpu& o) : band2gpt(o.band2gpt), gpt2band(o.gpt2band), band_lims_wvn(o.band_lims_wvn) {}
        Optical_props_gpu& operator=(const Optical_props_gpu& o)
        {
            band2gpt = o.band2gpt; gpt2band = o.gpt2band; band_lims_wvn = o.band_lims_wvn;
            band2gpt_gpu.reset(); gpt2band_gpu.reset();
            return *this;
        }
        int get_nband() const { return band2gpt.dim(2); }
        int get_ngpt() const { return band2gpt.is_empty() ? 0 : band2gpt.max(); }
        Array<int,1> get_gpoint_bands() const { return gpt2band; }
        const Array<int,2>& get_band_lims_gpoint() const { return band2gpt; }
        const Array<Float,2>& get_band_lims_wavenumber() const { return band_lims_wvn; }
        // device copies, uploaded on first use
        const Array_gpu<int,1>& get_gpoint_bands_gpu() const
        {
            if (!gpt2band_gpu) gpt2band_gpu = std::make_shared<Array_gpu<int,1>>(gpt2band);
            return *gpt2band_gpu;
        }
        const Array_gpu<int,2>& get_band_lims_gpoint_gpu() const
        {
            if (!band2gpt_gpu) band2gpt_gpu = std::make_shared<Array_gpu<int,2>>(band2gpt);
            return *band2gpt_gpu;
        }
    protected:
        void init()
        {
            const int ngpt = get_ngpt();
            gpt2band = Array<int,1>({ngpt});
            for (int b = 1; b <= get_nband(); ++b)
                for (int g = band2gpt({1, b}); g <= band2gpt({2, b}); ++g) gpt2band({g}) = b;
            band2gpt_gpu.reset(); gpt2band_gpu.reset();
        }
        void set_spectral(const std::vector<double>& wvn, const std::vector<int>& lims)       // from the C ABI's flat tables
        {
            const int nb = (int)lims.size() / 2;
            band2gpt = Array<int,2>(lims, {2, nb});
            band_lims_wvn = Array<Float,2>(std::vector<Float>(wvn.begin(), wvn.end()), {2, nb});
            init();
        }
        Array<int,2> band2gpt;
        Array<int,1> gpt2band;
        Array<Float,2> band_lims_wvn;
        mutable std::shared_ptr<Array_gpu<int,2>> band2gpt_gpu;
        mutable std::shared_ptr<Array_gpu<int,1>> gpt2band_gpu;
};

class Optical_props_arry_gpu : public Optical_props_gpu
{
    public:
        Optical_props_arry_gpu(const Optical_props_gpu& o) : Optical_props_gpu(o) {}
        virtual Array_gpu<Float,3>& get_tau() = 0;
        virtual Array_gpu<Float,3>& get_ssa() = 0;
        virtual Array_gpu<Float,3>& get_g() = 0;
        virtual const Array_gpu<Float,3>& get_tau() const = 0;
        virtual const Array_gpu<Float,3>& get_ssa() const = 0;
        virtual const Array_gpu<Float,3>& get_g() const = 0;
        virtual void delta_scale(const Array_gpu<Float,3>& forward_frac = Array_gpu<Float,3>()) = 0;
        virtual int get_ncol() const = 0;
        virtual int get_nlay() const = 0;
};

class Optical_props_1scl_gpu : public Optical_props_arry_gpu
{
    public:
        Optical_props_1scl_gpu(const int ncol, const int nlay, const Optical_props_gpu& o)
            : Optical_props_arry_gpu(o), tau({ncol, nlay, o.get_ngpt()}) {}
        Array_gpu<Float,3>& get_tau() override { return tau; }
        const Array_gpu<Float,3>& get_tau() const override { return tau; }
        Array_gpu<Float,3>& get_ssa() override { throw std::runtime_error("ssa is not available in this class"); }
        Array_gpu<Float,3>& get_g() override { throw std::runtime_error("g is not available in this class"); }
        const Array_gpu<Float,3>& get_ssa() const override { throw std::runtime_error("ssa is not available in this class"); }
        const Array_gpu<Float,3>& get_g() const override { throw std::runtime_error("g is not available in this class"); }
        void delta_scale(const Array_gpu<Float,3>& = Array_gpu<Float,3>()) override {}
        int get_ncol() const override { return tau.dim(1); }
        int get_nlay() const override { return tau.dim(2); }
    private:
        Array_gpu<Float,3> tau;
};

class Optical_props_2str_gpu : public Optical_props_arry_gpu
{
    public:
        Optical_props_2str_gpu(const int ncol, const int nlay, const Optical_props_gpu& o)
            : Optical_props_arry_gpu(o), tau({ncol, nlay, o.get_ngpt()}), ssa({ncol, nlay, o.get_ngpt()}), g({ncol, nlay, o.get_ngpt()}) {}
        Array_gpu<Float,3>& get_tau() override { return tau; }
        Array_gpu<Float,3>& get_ssa() override { return ssa; }
        Array_gpu<Float,3>& get_g() override { return g; }
        const Array_gpu<Float,3>& get_tau() const override { return tau; }
        const Array_gpu<Float,3>& get_ssa() const override { return ssa; }
        const Array_gpu<Float,3>& get_g() const override { return g; }
        void delta_scale(const Array_gpu<Float,3>& = Array_gpu<Float,3>()) override     // src/Optical_props.cpp:207-216
        {
            rrtmgp_b200::check(rrtmgp_b200_delta_scale_2str(rrtmgp_b200::dtype, tau.dim(1), tau.dim(2), tau.dim(3),
                                                            tau.ptr(), ssa.ptr(), g.ptr(), rrtmgp_b200::stream()));
        }
        int get_ncol() const override { return tau.dim(1); }
        int get_nlay() const override { return tau.dim(2); }
    private:
        Array_gpu<Float,3> tau, ssa, g;
};

// Optical_props increment (src/Optical_props.cpp:219-268)
inline void add_to(Optical_props_1scl_gpu& op_inout, const Optical_props_1scl_gpu& op_in)
{
    const int ncol = op_inout.get_ncol(), nlay = op_inout.get_nlay(), ngpt = op_inout.get_ngpt();
    if (ngpt == op_in.get_ngpt())
        rrtmgp_b200::check(rrtmgp_b200_increment_1scl(rrtmgp_b200::dtype, ncol, nlay, ngpt, op_inout.get_tau().ptr(),
                                                      op_in.get_tau().ptr(), rrtmgp_b200::stream()));
    else
    {
        if (op_in.get_ngpt() != op_inout.get_nband())
            throw std::runtime_error("Cannot add optical properties with incompatible band - gpoint combination");
        rrtmgp_b200::check(rrtmgp_b200_increment_1scl_bybnd(rrtmgp_b200::dtype, ncol, nlay, ngpt, op_inout.get_tau().ptr(),
                op_in.get_tau().ptr(), op_inout.get_nband(), op_inout.get_band_lims_gpoint().ptr(), rrtmgp_b200::stream()));
    }
}
inline void add_to(Optical_props_2str_gpu& op_inout, const Optical_props_2str_gpu& op_in)
{
    const int ncol = op_inout.get_ncol(), nlay = op_inout.get_nlay(), ngpt = op_inout.get_ngpt();
    if (ngpt == op_in.get_ngpt())
        rrtmgp_b200::check(rrtmgp_b200_increment_2str(rrtmgp_b200::dtype, ncol, nlay, ngpt, op_inout.get_tau().ptr(),
                op_inout.get_ssa().ptr(), op_inout.get_g().ptr(), op_in.get_tau().ptr(), op_in.get_ssa().ptr(),
                op_in.get_g().ptr(), rrtmgp_b200::stream()));
    else
    {
        if (op_in.get_ngpt() != op_inout.get_nband())
            throw std::runtime_error("Cannot add optical properties with incompatible band - gpoint combination");
        rrtmgp_b200::check(rrtmgp_b200_increment_2str_bybnd(rrtmgp_b200::dtype, ncol, nlay, ngpt, op_inout.get_tau().ptr(),
                op_inout.get_ssa().ptr(), op_inout.get_g().ptr(), op_in.get_tau().ptr(), op_in.get_ssa().ptr(),
                op_in.get_g().ptr(), op_inout.get_nband(), op_inout.get_band_lims_gpoint().ptr(), rrtmgp_b200::stream()));
    }
}

class Source_func_lw_gpu : public Optical_props_gpu
{
    public:
        Source_func_lw_gpu(const int n_col, const int n_lay, const Optical_props_gpu& o)
            : Optical_props_gpu(o), sfc_source({n_col, o.get_ngpt()}), sfc_source_jac({n_col, o.get_ngpt()}),
              lay_source({n_col, n_lay, o.get_ngpt()}), lev_source({n_col, n_lay + 1, o.get_ngpt()}) {}
        Array_gpu<Float,2>& get_sfc_source() { return sfc_source; }
        Array_gpu<Float,2>& get_sfc_source_jac() { return sfc_source_jac; }
        Array_gpu<Float,3>& get_lay_source() { return lay_source; }
        Array_gpu<Float,3>& get_lev_source() { return lev_source; }
        const Array_gpu<Float,2>& get_sfc_source() const { return sfc_source; }
        const Array_gpu<Float,2>& get_sfc_source_jac() const { return sfc_source_jac; }
        const Array_gpu<Float,3>& get_lay_source() const { return lay_source; }
        const Array_gpu<Float,3>& get_lev_source() const { return lev_source; }
    private:
        Array_gpu<Float,2> sfc_source, sfc_source_jac;
        Array_gpu<Float,3> lay_source, lev_source;
};

namespace rrtmgp_b200 {
// Gas_concs -> the C struct, in the order of the k-distribution object's gas list
struct Concs_view
{
    std::vector<const void*> p; std::vector<int> nc, nl; rrtmgp_b200_gas_concs c{};
    Concs_view(const Gas_concs_gpu& g, const std::vector<std::string>& names)
    {
        for (auto& n : names) { const auto& a = g.get_vmr(n); p.push_back(a.ptr()); nc.push_back(a.dim(1)); nl.push_back(a.dim(2)); }
        c.ngas = (int)names.size(); c.d_vmr = p.data(); c.vmr_ncol = nc.data(); c.vmr_nlay = nl.data();
    }
};
inline std::string pad_names(const std::vector<std::string>& names, int len)
{
    std::string s;
    for (auto n : names) { n.resize(len, ' '); s += n; }
    return s;
}
}  // namespace rrtmgp_b200

namespace rrtmgp_b200 {
// Helpers of the positional constructors: the reference hands its constructors Array objects whose memory order IS
// the coefficient file's (column-major dims reversed, src_test/Radiation_solver.cu:125-243), so packing them into the
// C ABI's file-order structs is a pointer hand-over (plus a widening copy when Float is float).
struct Table_pack
{
    std::deque<std::vector<double>> dbl; std::deque<std::vector<int>> ints; std::deque<std::string> strs;   // deque: element addresses are stable
    template<int N> const double* d(const Array<Float,N>& a)
    {
        if (a.size() == 0) return nullptr;
        dbl.emplace_back(a.v().begin(), a.v().end());
        return dbl.back().data();
    }
    const int* b(const Array<Bool,1>& a)
    {
        ints.emplace_back(a.v().begin(), a.v().end());
        return ints.back().data();
    }
    const char* s(const Array<std::string,1>& a, int len)
    {
        strs.push_back(pad_names(a.v(), len));
        return strs.back().c_str();
    }
};
constexpr int kNameLen = 32;
}  // namespace rrtmgp_b200

class Gas_optics_rrtmgp_gpu : public Optical_props_gpu
{
    public:
        // Longwave constructor, the reference's 32 positional arguments (include/Gas_optics_rrtmgp.h:293-325)
        Gas_optics_rrtmgp_gpu(
                const Gas_concs_gpu& available_gases,
                const Array<std::string,1>& gas_names,
                const Array<int,3>& key_species,
                const Array<int,2>& band2gpt,
                const Array<Float,2>& band_lims_wavenum,
                const Array<Float,1>& press_ref,
                const Float press_ref_trop,
                const Array<Float,1>& temp_ref,
                const Float temp_ref_p,
                const Float temp_ref_t,
                const Array<Float,3>& vmr_ref,
                const Array<Float,4>& kmajor,
                const Array<Float,3>& kminor_lower,
                const Array<Float,3>& kminor_upper,
                const Array<std::string,1>& gas_minor,
                const Array<std::string,1>& identifier_minor,
                const Array<std::string,1>& minor_gases_lower,
                const Array<std::string,1>& minor_gases_upper,
                const Array<int,2>& minor_limits_gpt_lower,
                const Array<int,2>& minor_limits_gpt_upper,
                const Array<Bool,1>& minor_scales_with_density_lower,
                const Array<Bool,1>& minor_scales_with_density_upper,
                const Array<std::string,1>& scaling_gas_lower,
                const Array<std::string,1>& scaling_gas_upper,
                const Array<Bool,1>& scale_by_complement_lower,
                const Array<Bool,1>& scale_by_complement_upper,
                const Array<int,1>& kminor_start_lower,
                const Array<int,1>& kminor_start_upper,
                const Array<Float,2>& totplnk,
                const Array<Float,4>& planck_frac,
                const Array<Float,3>& rayl_lower,
                const Array<Float,3>& rayl_upper)
        {
            rrtmgp_b200::Table_pack t;
            rrtmgp_b200_kdist_file f = pack_common(t, gas_names, key_species, band2gpt, band_lims_wavenum, press_ref, press_ref_trop,
                    temp_ref, temp_ref_p, temp_ref_t, vmr_ref, kmajor, kminor_lower, kminor_upper, gas_minor, identifier_minor,
                    minor_gases_lower, minor_gases_upper, minor_limits_gpt_lower, minor_limits_gpt_upper,
                    minor_scales_with_density_lower, minor_scales_with_density_upper, scaling_gas_lower, scaling_gas_upper,
                    scale_by_complement_lower, scale_by_complement_upper, kminor_start_lower, kminor_start_upper, rayl_lower, rayl_upper);
            f.n_planck_temps = totplnk.dim(1);
            f.totplnk = t.d(totplnk); f.plank_fraction = t.d(planck_frac);
            create(available_gases, f);
        }
        // Shortwave constructor, 35 positional arguments (include/Gas_optics_rrtmgp.h:328-364)
        Gas_optics_rrtmgp_gpu(
                const Gas_concs_gpu& available_gases,
                const Array<std::string,1>& gas_names,
                const Array<int,3>& key_species,
                const Array<int,2>& band2gpt,
                const Array<Float,2>& band_lims_wavenum,
                const Array<Float,1>& press_ref,
                const Float press_ref_trop,
                const Array<Float,1>& temp_ref,
                const Float temp_ref_p,
                const Float temp_ref_t,
                const Array<Float,3>& vmr_ref,
                const Array<Float,4>& kmajor,
                const Array<Float,3>& kminor_lower,
                const Array<Float,3>& kminor_upper,
                const Array<std::string,1>& gas_minor,
                const Array<std::string,1>& identifier_minor,
                const Array<std::string,1>& minor_gases_lower,
                const Array<std::string,1>& minor_gases_upper,
                const Array<int,2>& minor_limits_gpt_lower,
                const Array<int,2>& minor_limits_gpt_upper,
                const Array<Bool,1>& minor_scales_with_density_lower,
                const Array<Bool,1>& minor_scales_with_density_upper,
                const Array<std::string,1>& scaling_gas_lower,
                const Array<std::string,1>& scaling_gas_upper,
                const Array<Bool,1>& scale_by_complement_lower,
                const Array<Bool,1>& scale_by_complement_upper,
                const Array<int,1>& kminor_start_lower,
                const Array<int,1>& kminor_start_upper,
                const Array<Float,1>& solar_src_quiet,
                const Array<Float,1>& solar_src_facular,
                const Array<Float,1>& solar_src_sunspot,
                const Float tsi_default,
                const Float mg_default,
                const Float sb_default,
                const Array<Float,3>& rayl_lower,
                const Array<Float,3>& rayl_upper)
        {
            rrtmgp_b200::Table_pack t;
            rrtmgp_b200_kdist_file f = pack_common(t, gas_names, key_species, band2gpt, band_lims_wavenum, press_ref, press_ref_trop,
                    temp_ref, temp_ref_p, temp_ref_t, vmr_ref, kmajor, kminor_lower, kminor_upper, gas_minor, identifier_minor,
                    minor_gases_lower, minor_gases_upper, minor_limits_gpt_lower, minor_limits_gpt_upper,
                    minor_scales_with_density_lower, minor_scales_with_density_upper, scaling_gas_lower, scaling_gas_upper,
                    scale_by_complement_lower, scale_by_complement_upper, kminor_start_lower, kminor_start_upper, rayl_lower, rayl_upper);
            f.solar_source_quiet = t.d(solar_src_quiet); f.solar_source_facular = t.d(solar_src_facular);
            f.solar_source_sunspot = t.d(solar_src_sunspot);
            f.tsi_default = tsi_default; f.mg_default = mg_default; f.sb_default = sb_default;
            create(available_gases, f);
        }
        // The same variables as one struct in file order (include/rrtmgp_b200_tables.h): what this repository's own
        // loaders (include/rrtmgp_b200/coefficients.hpp) produce.
        Gas_optics_rrtmgp_gpu(const Gas_concs_gpu& available_gases, const rrtmgp_b200_kdist_file& kdist) { create(available_gases, kdist); }
        ~Gas_optics_rrtmgp_gpu() { if (h) rrtmgp_b200_gas_optics_destroy(h); }
        Gas_optics_rrtmgp_gpu(const Gas_optics_rrtmgp_gpu&) = delete;
        Gas_optics_rrtmgp_gpu& operator=(const Gas_optics_rrtmgp_gpu&) = delete;
        // returned by value by load_and_init_gas_optics (src_test/Radiation_solver.cu:243, :428)
        Gas_optics_rrtmgp_gpu(Gas_optics_rrtmgp_gpu&& o) noexcept
            : Optical_props_gpu(o), h(o.h), info(o.info), gas_names(std::move(o.gas_names)) { o.h = nullptr; }

        static void get_col_dry(Array_gpu<Float,2>& col_dry, const Array_gpu<Float,2>& vmr_h2o, const Array_gpu<Float,2>& plev)
        {
            rrtmgp_b200::check(rrtmgp_b200_get_col_dry(rrtmgp_b200::dtype, vmr_h2o.dim(1), vmr_h2o.dim(2), vmr_h2o.ptr(),
                                                       plev.ptr(), col_dry.ptr(), rrtmgp_b200::stream()));
        }
        bool source_is_internal() const { return info.is_longwave != 0; }
        bool source_is_external() const { return info.is_longwave == 0; }
        Float get_press_ref_min() const { return (Float)info.press_ref_min; }
        Float get_press_ref_max() const { return (Float)info.press_ref_max; }
        Float get_temp_min() const { return (Float)info.temp_ref_min; }
        Float get_temp_max() const { return (Float)info.temp_ref_max; }
        Float get_tsi() const { return (Float)info.tsi; }
        rrtmgp_b200_gas_optics* handle() const { return h; }
        const std::vector<std::string>& get_gas_names() const { return gas_names; }

        // Longwave (include/Gas_optics_rrtmgp.h:389-398).  Range checks as the CPU path throws them.
        void gas_optics(const Array_gpu<Float,2>& play, const Array_gpu<Float,2>& plev, const Array_gpu<Float,2>& tlay,
                        const Array_gpu<Float,1>& tsfc, const Gas_concs_gpu& gas_desc,
                        std::unique_ptr<Optical_props_arry_gpu>& optical_props, Source_func_lw_gpu& sources,
                        const Array_gpu<Float,2>& col_dry, const Array_gpu<Float,2>& tlev)
        {
            validate(play, "play", true); validate(plev, "plev", true); validate(tlay, "tlay", false);
            validate(tlev, "tlev", false); validate(tsfc, "tsfc", false);
            rrtmgp_b200::Concs_view cv(gas_desc, gas_names);
            rrtmgp_b200::check(rrtmgp_b200_gas_optics_lw(h, play.dim(1), play.dim(2), play.ptr(), plev.ptr(), tlay.ptr(),
                    tsfc.ptr(), tlev.ptr(), col_dry.size() ? col_dry.ptr() : nullptr, &cv.c, optical_props->get_tau().ptr(),
                    sources.get_lay_source().ptr(), sources.get_lev_source().ptr(), sources.get_sfc_source().ptr(),
                    sources.get_sfc_source_jac().ptr(), rrtmgp_b200::stream()));
        }
        // Shortwave (include/Gas_optics_rrtmgp.h:401-408)
        void gas_optics(const Array_gpu<Float,2>& play, const Array_gpu<Float,2>& plev, const Array_gpu<Float,2>& tlay,
                        const Gas_concs_gpu& gas_desc, std::unique_ptr<Optical_props_arry_gpu>& optical_props,
                        Array_gpu<Float,2>& toa_src, const Array_gpu<Float,2>& col_dry)
        {
            validate(play, "play", true); validate(plev, "plev", true); validate(tlay, "tlay", false);
            rrtmgp_b200::Concs_view cv(gas_desc, gas_names);
            const bool two = dynamic_cast<Optical_props_2str_gpu*>(optical_props.get()) != nullptr;
            rrtmgp_b200::check(rrtmgp_b200_gas_optics_sw(h, play.dim(1), play.dim(2), play.ptr(), plev.ptr(), tlay.ptr(),
                    col_dry.size() ? col_dry.ptr() : nullptr, &cv.c, optical_props->get_tau().ptr(),
                    two ? optical_props->get_ssa().ptr() : nullptr, two ? optical_props->get_g().ptr() : nullptr,
                    toa_src.ptr(), rrtmgp_b200::stream()));
        }
    private:
        template<int N> void validate(const Array_gpu<Float,N>& a, const char* name, bool pressure) const
        {
            const double lo = pressure ? info.press_ref_min : info.temp_ref_min, hi = pressure ? info.press_ref_max : info.temp_ref_max;
            rrtmgp_b200::check(rrtmgp_b200_check_range(rrtmgp_b200::dtype, a.size(), a.ptr(), lo, hi, name, rrtmgp_b200::stream()));
        }
        void create(const Gas_concs_gpu& available_gases, const rrtmgp_b200_kdist_file& kdist)
        {
            const auto names = available_gases.names();
            const std::string blk = rrtmgp_b200::pad_names(names, rrtmgp_b200::kNameLen);
            rrtmgp_b200::check(rrtmgp_b200_gas_optics_create(&kdist, (int)names.size(), blk.c_str(), rrtmgp_b200::kNameLen,
                                                             rrtmgp_b200::dtype | rrtmgp_b200::create_flags(), &h));
            rrtmgp_b200::check(rrtmgp_b200_gas_optics_get_info(h, &info));
            std::vector<int> lims(2 * (size_t)info.nbnd), g2b((size_t)info.ngpt);
            std::vector<double> wvn(2 * (size_t)info.nbnd), sol((size_t)info.ngpt);
            std::vector<char> gn((size_t)info.ngas * rrtmgp_b200::kNameLen);
            rrtmgp_b200::check(rrtmgp_b200_gas_optics_get_tables(h, lims.data(), wvn.data(), g2b.data(), nullptr,
                                                                 nullptr, gn.data(), rrtmgp_b200::kNameLen, sol.data()));
            set_spectral(wvn, lims);
            for (int i = 0; i < info.ngas; ++i)
            {
                std::string s(gn.data() + (size_t)i * rrtmgp_b200::kNameLen, rrtmgp_b200::kNameLen);
                gas_names.push_back(s.substr(0, s.find_last_not_of(' ') + 1));
            }
        }
        // the 28 arguments the two constructors share -> the file struct (dimension names: rrtmgp_b200_tables.h)
        static rrtmgp_b200_kdist_file pack_common(rrtmgp_b200::Table_pack& t,
                const Array<std::string,1>& gas_names, const Array<int,3>& key_species, const Array<int,2>& band2gpt,
                const Array<Float,2>& band_lims_wavenum, const Array<Float,1>& press_ref, const Float press_ref_trop,
                const Array<Float,1>& temp_ref, const Float temp_ref_p, const Float temp_ref_t, const Array<Float,3>& vmr_ref,
                const Array<Float,4>& kmajor, const Array<Float,3>& kminor_lower, const Array<Float,3>& kminor_upper,
                const Array<std::string,1>& gas_minor, const Array<std::string,1>& identifier_minor,
                const Array<std::string,1>& minor_gases_lower, const Array<std::string,1>& minor_gases_upper,
                const Array<int,2>& minor_limits_gpt_lower, const Array<int,2>& minor_limits_gpt_upper,
                const Array<Bool,1>& dens_lower, const Array<Bool,1>& dens_upper,
                const Array<std::string,1>& scaling_gas_lower, const Array<std::string,1>& scaling_gas_upper,
                const Array<Bool,1>& comp_lower, const Array<Bool,1>& comp_upper,
                const Array<int,1>& kminor_start_lower, const Array<int,1>& kminor_start_upper,
                const Array<Float,3>& rayl_lower, const Array<Float,3>& rayl_upper)
        {
            const int L = rrtmgp_b200::kNameLen;
            rrtmgp_b200_kdist_file f{};
            f.n_temps = temp_ref.dim(1); f.n_press = press_ref.dim(1); f.n_absorbers = gas_names.dim(1);
            f.n_minorabsorbers = gas_minor.dim(1); f.n_extabsorbers = vmr_ref.dim(2); f.n_mixingfracs = kmajor.dim(2);
            f.n_layers = vmr_ref.dim(1); f.n_bnds = band2gpt.dim(2); f.n_gpts = kmajor.dim(1); f.n_pairs = 2;
            f.n_minor_lower = minor_limits_gpt_lower.dim(2); f.n_minor_upper = minor_limits_gpt_upper.dim(2);
            f.n_contrib_lower = kminor_lower.dim(1); f.n_contrib_upper = kminor_upper.dim(1);
            f.n_planck_temps = 0; f.string_len = L;
            if (kmajor.dim(3) != f.n_press + 1 || kmajor.dim(4) != f.n_temps || key_species.dim(1) != 2 || key_species.dim(3) != f.n_bnds)
                throw std::runtime_error("Gas_optics_rrtmgp_gpu: inconsistent table dimensions");
            f.gas_names = t.s(gas_names, L); f.key_species = key_species.ptr();
            f.bnd_limits_wavenumber = t.d(band_lims_wavenum); f.bnd_limits_gpt = band2gpt.ptr();
            f.press_ref = t.d(press_ref); f.temp_ref = t.d(temp_ref);
            f.absorption_coefficient_ref_P = temp_ref_p; f.absorption_coefficient_ref_T = temp_ref_t; f.press_ref_trop = press_ref_trop;
            f.vmr_ref = t.d(vmr_ref); f.kmajor = t.d(kmajor); f.kminor_lower = t.d(kminor_lower); f.kminor_upper = t.d(kminor_upper);
            f.gas_minor = t.s(gas_minor, L); f.identifier_minor = t.s(identifier_minor, L);
            f.minor_gases_lower = t.s(minor_gases_lower, L); f.minor_gases_upper = t.s(minor_gases_upper, L);
            f.minor_limits_gpt_lower = minor_limits_gpt_lower.ptr(); f.minor_limits_gpt_upper = minor_limits_gpt_upper.ptr();
            f.minor_scales_with_density_lower = t.b(dens_lower); f.minor_scales_with_density_upper = t.b(dens_upper);
            f.scale_by_complement_lower = t.b(comp_lower); f.scale_by_complement_upper = t.b(comp_upper);
            f.scaling_gas_lower = t.s(scaling_gas_lower, L); f.scaling_gas_upper = t.s(scaling_gas_upper, L);
            f.kminor_start_lower = kminor_start_lower.ptr(); f.kminor_start_upper = kminor_start_upper.ptr();
            f.rayl_lower = t.d(rayl_lower); f.rayl_upper = t.d(rayl_upper);
            return f;
        }
        rrtmgp_b200_gas_optics* h = nullptr;
        rrtmgp_b200_gas_optics_info info{};
        std::vector<std::string> gas_names;
};

class Cloud_optics_gpu : public Optical_props_gpu
{
    public:
        // the reference's positional constructor (include/Cloud_optics.h:84-89; arrays as load_and_init_cloud_optics
        // builds them, src_test/Radiation_solver.cu:327-364: band_lims_wvn (2, nband), lut_*liq (nsize_liq, nband),
        // lut_*ice (nsize_ice, nband, nrghice))
        Cloud_optics_gpu(
                const Array<Float,2>& band_lims_wvn,
                const Float radliq_lwr, const Float radliq_upr, const Float radliq_fac,
                const Float diamice_lwr, const Float diamice_upr, const Float diamice_fac,
                const Array<Float,2>& lut_extliq, const Array<Float,2>& lut_ssaliq, const Array<Float,2>& lut_asyliq,
                const Array<Float,3>& lut_extice, const Array<Float,3>& lut_ssaice, const Array<Float,3>& lut_asyice)
            : Optical_props_gpu(band_lims_wvn)
        {
            rrtmgp_b200::Table_pack t;
            rrtmgp_b200_cloud_file f{};
            f.n_band = band_lims_wvn.dim(2); f.n_rghice = lut_extice.dim(3); f.n_size_liq = lut_extliq.dim(1); f.n_size_ice = lut_extice.dim(1);
            if (lut_extliq.dim(2) != f.n_band || lut_extice.dim(2) != f.n_band) throw std::runtime_error("Cloud_optics_gpu: inconsistent table dimensions");
            f.bnd_limits_wavenumber = t.d(band_lims_wvn);
            f.radliq_lwr = radliq_lwr; f.radliq_upr = radliq_upr; f.radliq_fac = radliq_fac;
            f.diamice_lwr = diamice_lwr; f.diamice_upr = diamice_upr; f.diamice_fac = diamice_fac;
            f.lut_extliq = t.d(lut_extliq); f.lut_ssaliq = t.d(lut_ssaliq); f.lut_asyliq = t.d(lut_asyliq);
            f.lut_extice = t.d(lut_extice); f.lut_ssaice = t.d(lut_ssaice); f.lut_asyice = t.d(lut_asyice);
            rrtmgp_b200::check(rrtmgp_b200_cloud_optics_create(&f, rrtmgp_b200::dtype, &h));
        }
        explicit Cloud_optics_gpu(const rrtmgp_b200_cloud_file& lut)
            : Optical_props_gpu(Array<Float,2>(std::vector<Float>(lut.bnd_limits_wavenumber, lut.bnd_limits_wavenumber + 2 * lut.n_band), {2, lut.n_band}))
        { rrtmgp_b200::check(rrtmgp_b200_cloud_optics_create(&lut, rrtmgp_b200::dtype, &h)); }
        ~Cloud_optics_gpu() { if (h) rrtmgp_b200_cloud_optics_destroy(h); }
        Cloud_optics_gpu(const Cloud_optics_gpu&) = delete;
        Cloud_optics_gpu(Cloud_optics_gpu&& o) noexcept : Optical_props_gpu(o), h(o.h) { o.h = nullptr; }
        void cloud_optics(const Array_gpu<Float,2>& clwp, const Array_gpu<Float,2>& ciwp, const Array_gpu<Float,2>& reliq,
                          const Array_gpu<Float,2>& deice, Optical_props_2str_gpu& optical_props)
        {
            rrtmgp_b200::check(rrtmgp_b200_cloud_optics_2str(h, clwp.dim(1), clwp.dim(2), clwp.ptr(), ciwp.ptr(), reliq.ptr(),
                    deice.ptr(), optical_props.get_tau().ptr(), optical_props.get_ssa().ptr(), optical_props.get_g().ptr(),
                    rrtmgp_b200::stream()));
        }
        void cloud_optics(const Array_gpu<Float,2>& clwp, const Array_gpu<Float,2>& ciwp, const Array_gpu<Float,2>& reliq,
                          const Array_gpu<Float,2>& deice, Optical_props_1scl_gpu& optical_props)
        {
            rrtmgp_b200::check(rrtmgp_b200_cloud_optics_1scl(h, clwp.dim(1), clwp.dim(2), clwp.ptr(), ciwp.ptr(), reliq.ptr(),
                    deice.ptr(), optical_props.get_tau().ptr(), rrtmgp_b200::stream()));
        }
        rrtmgp_b200_cloud_optics* handle() const { return h; }
    private:
        rrtmgp_b200_cloud_optics* h = nullptr;
};

using Aerosol_concs_gpu = Gas_concs_gpu;       // include/Aerosol_optics.h:20

namespace rrtmgp_b200 {
// Aerosol_concs -> the C struct (aermr01..aermr11)
inline rrtmgp_b200_aerosol_concs aerosol_view(const Aerosol_concs_gpu& a)
{
    rrtmgp_b200_aerosol_concs c{};
    for (int i = 0; i < 11; ++i)
    {
        const std::string name = std::string("aermr") + (i < 9 ? "0" : "") + std::to_string(i + 1);
        const auto& m = a.get_vmr(name);
        c.d_mmr[i] = m.ptr(); c.mmr_ncol[i] = m.dim(1); c.mmr_nlay[i] = m.dim(2);
    }
    return c;
}
}  // namespace rrtmgp_b200

// include/Aerosol_optics.h:48-84.  Like the reference's loader (src_test/Radiation_solver.cu:366-402) the band
// limits of the struct constructor stay zero.
class Aerosol_optics_gpu : public Optical_props_gpu
{
    public:
        // the reference's positional constructor (include/Aerosol_optics.h:53-56; arrays as load_and_init_aerosol_optics
        // builds them, src_test/Radiation_solver.cu:366-402: m*_phobic (band, hydrophobic), m*_philic (band, humidity, hydrophilic))
        Aerosol_optics_gpu(
                const Array<Float,2>& band_lims_wvn, const Array<Float,1>& rh_upper,
                const Array<Float,2>& mext_phobic, const Array<Float,2>& ssa_phobic, const Array<Float,2>& g_phobic,
                const Array<Float,3>& mext_philic, const Array<Float,3>& ssa_philic, const Array<Float,3>& g_philic)
            : Optical_props_gpu(band_lims_wvn)
        {
            rrtmgp_b200::Table_pack t;
            rrtmgp_b200_aerosol_file f{};
            f.n_band = band_lims_wvn.dim(2); f.n_hum = rh_upper.dim(1); f.n_philic = mext_philic.dim(3); f.n_phobic = mext_phobic.dim(2);
            if (mext_phobic.dim(1) != f.n_band || mext_philic.dim(1) != f.n_band || mext_philic.dim(2) != f.n_hum)
                throw std::runtime_error("Aerosol_optics_gpu: inconsistent table dimensions");
            f.mass_ext_sw_hydrophobic = t.d(mext_phobic); f.ssa_sw_hydrophobic = t.d(ssa_phobic); f.asymmetry_sw_hydrophobic = t.d(g_phobic);
            f.mass_ext_sw_hydrophilic = t.d(mext_philic); f.ssa_sw_hydrophilic = t.d(ssa_philic); f.asymmetry_sw_hydrophilic = t.d(g_philic);
            f.relative_humidity2 = t.d(rh_upper);
            rrtmgp_b200::check(rrtmgp_b200_aerosol_optics_create(&f, rrtmgp_b200::dtype, &h));
        }
        explicit Aerosol_optics_gpu(const rrtmgp_b200_aerosol_file& lut)
            : Optical_props_gpu(Array<Float,2>({2, lut.n_band}))
        { rrtmgp_b200::check(rrtmgp_b200_aerosol_optics_create(&lut, rrtmgp_b200::dtype, &h)); }
        ~Aerosol_optics_gpu() { if (h) rrtmgp_b200_aerosol_optics_destroy(h); }
        Aerosol_optics_gpu(const Aerosol_optics_gpu&) = delete;
        Aerosol_optics_gpu(Aerosol_optics_gpu&& o) noexcept : Optical_props_gpu(o), h(o.h) { o.h = nullptr; }
        void aerosol_optics(const Aerosol_concs_gpu& aerosol_concs, const Array_gpu<Float,2>& rh, const Array_gpu<Float,2>& plev,
                            Optical_props_2str_gpu& optical_props) const
        {
            const rrtmgp_b200_aerosol_concs ac = rrtmgp_b200::aerosol_view(aerosol_concs);
            rrtmgp_b200::check(rrtmgp_b200_aerosol_optics_2str(h, rh.dim(1), rh.dim(2), &ac, rh.ptr(), plev.ptr(),
                    optical_props.get_tau().ptr(), optical_props.get_ssa().ptr(), optical_props.get_g().ptr(), rrtmgp_b200::stream()));
        }
        rrtmgp_b200_aerosol_optics* handle() const { return h; }
    private:
        rrtmgp_b200_aerosol_optics* h = nullptr;
};

class Rte_lw_gpu
{
    public:
        static void expand_and_transpose(const std::unique_ptr<Optical_props_arry_gpu>& ops, const Array_gpu<Float,2>& arr_in,
                                         Array_gpu<Float,2>& arr_out)
        {
            rrtmgp_b200::check(rrtmgp_b200_expand_and_transpose(rrtmgp_b200::dtype, arr_in.dim(2), arr_in.dim(1), ops->get_ngpt(),
                    ops->get_band_lims_gpoint().ptr(), arr_in.ptr(), arr_out.ptr(), rrtmgp_b200::stream()));
        }
        // include/Rte_lw.h:62-77; broadband iff gpt_flux_up.dim(3) == 1 (src/Rte_lw.cpp:170)
        static void rte_lw(const std::unique_ptr<Optical_props_arry_gpu>& optical_props, const Bool top_at_1,
                           const Source_func_lw_gpu& sources, const Array_gpu<Float,2>& sfc_emis,
                           const Array_gpu<Float,2>& inc_flux, Array_gpu<Float,3>& gpt_flux_up, Array_gpu<Float,3>& gpt_flux_dn,
                           const int n_gauss_angles, const Bool do_rescaling = false)
        {
            const int ncol = optical_props->get_ncol(), nlay = optical_props->get_nlay(), ngpt = optical_props->get_ngpt();
            Array_gpu<Float,2> emis_gpt({ncol, ngpt});
            expand_and_transpose(optical_props, sfc_emis, emis_gpt);
            if (do_rescaling)   // false in the reference (src/Rte_lw.cpp:178); the kernel's re-scaled transport, 2-stream properties
            {
                rrtmgp_b200::check(rrtmgp_b200_lw_solver_rescaled(rrtmgp_b200::dtype, ncol, nlay, ngpt, top_at_1, n_gauss_angles,
                        optical_props->get_tau().ptr(), optical_props->get_ssa().ptr(), optical_props->get_g().ptr(),
                        sources.get_lay_source().ptr(), sources.get_lev_source().ptr(),
                        emis_gpt.ptr(), sources.get_sfc_source().ptr(), inc_flux.size() ? inc_flux.ptr() : nullptr,
                        gpt_flux_up.ptr(), gpt_flux_dn.ptr(), gpt_flux_up.dim(3) == 1, nullptr, nullptr, rrtmgp_b200::stream()));
                return;
            }
            rrtmgp_b200::check(rrtmgp_b200_lw_solver_noscat(rrtmgp_b200::dtype, ncol, nlay, ngpt, top_at_1, n_gauss_angles,
                    optical_props->get_tau().ptr(), sources.get_lay_source().ptr(), sources.get_lev_source().ptr(),
                    emis_gpt.ptr(), sources.get_sfc_source().ptr(), inc_flux.size() ? inc_flux.ptr() : nullptr,
                    gpt_flux_up.ptr(), gpt_flux_dn.ptr(), gpt_flux_up.dim(3) == 1, nullptr, nullptr, rrtmgp_b200::stream()));
        }
};

class Rte_sw_gpu
{
    public:
        // include/Rte_sw.h:62-80; has_dif_bc iff inc_flux_dif.size() > 0
        static void rte_sw(const std::unique_ptr<Optical_props_arry_gpu>& optical_props, const Bool top_at_1,
                           const Array_gpu<Float,1>& mu0, const Array_gpu<Float,2>& inc_flux_dir,
                           const Array_gpu<Float,2>& sfc_alb_dir, const Array_gpu<Float,2>& sfc_alb_dif,
                           const Array_gpu<Float,2>& inc_flux_dif, Array_gpu<Float,3>& gpt_flux_up,
                           Array_gpu<Float,3>& gpt_flux_dn, Array_gpu<Float,3>& gpt_flux_dir)
        {
            const int ncol = optical_props->get_ncol(), nlay = optical_props->get_nlay(), ngpt = optical_props->get_ngpt();
            Array_gpu<Float,2> a_dir({ncol, ngpt}), a_dif({ncol, ngpt});
            Rte_lw_gpu::expand_and_transpose(optical_props, sfc_alb_dir, a_dir);
            Rte_lw_gpu::expand_and_transpose(optical_props, sfc_alb_dif, a_dif);
            rrtmgp_b200::check(rrtmgp_b200_sw_solver_2stream(rrtmgp_b200::dtype, ncol, nlay, ngpt, top_at_1,
                    optical_props->get_tau().ptr(), optical_props->get_ssa().ptr(), optical_props->get_g().ptr(), mu0.ptr(),
                    a_dir.ptr(), a_dif.ptr(), inc_flux_dir.ptr(), inc_flux_dif.size() ? inc_flux_dif.ptr() : nullptr,
                    gpt_flux_up.ptr(), gpt_flux_dn.ptr(), gpt_flux_dir.ptr(), gpt_flux_up.dim(3) == 1, rrtmgp_b200::stream()));
        }
        // mu0 (ncol, nlay): the kernel's own shape (mo_rte_solver_kernels.F90:517), a height-dependent zenith cosine
        static void rte_sw(const std::unique_ptr<Optical_props_arry_gpu>& optical_props, const Bool top_at_1,
                           const Array_gpu<Float,2>& mu0_lay, const Array_gpu<Float,2>& inc_flux_dir,
                           const Array_gpu<Float,2>& sfc_alb_dir, const Array_gpu<Float,2>& sfc_alb_dif,
                           const Array_gpu<Float,2>& inc_flux_dif, Array_gpu<Float,3>& gpt_flux_up,
                           Array_gpu<Float,3>& gpt_flux_dn, Array_gpu<Float,3>& gpt_flux_dir)
        {
            const int ncol = optical_props->get_ncol(), nlay = optical_props->get_nlay(), ngpt = optical_props->get_ngpt();
            if (mu0_lay.dim(1) != ncol || mu0_lay.dim(2) != nlay) throw std::runtime_error("rte_sw: mu0 must be (ncol, nlay)");
            Array_gpu<Float,2> a_dir({ncol, ngpt}), a_dif({ncol, ngpt});
            Rte_lw_gpu::expand_and_transpose(optical_props, sfc_alb_dir, a_dir);
            Rte_lw_gpu::expand_and_transpose(optical_props, sfc_alb_dif, a_dif);
            rrtmgp_b200::check(rrtmgp_b200_sw_solver_2stream_mu0lay(rrtmgp_b200::dtype, ncol, nlay, ngpt, top_at_1,
                    optical_props->get_tau().ptr(), optical_props->get_ssa().ptr(), optical_props->get_g().ptr(), mu0_lay.ptr(),
                    a_dir.ptr(), a_dif.ptr(), inc_flux_dir.ptr(), inc_flux_dif.size() ? inc_flux_dif.ptr() : nullptr,
                    gpt_flux_up.ptr(), gpt_flux_dn.ptr(), gpt_flux_dir.ptr(), gpt_flux_up.dim(3) == 1, rrtmgp_b200::stream()));
        }
};

class Fluxes_broadband_gpu
{
    public:
        Fluxes_broadband_gpu(const int ncol, const int nlev)
            : flux_up({ncol, nlev}), flux_dn({ncol, nlev}), flux_dn_dir({ncol, nlev}), flux_net({ncol, nlev}) {}
        virtual ~Fluxes_broadband_gpu() = default;
        virtual void reduce(const Array_gpu<Float,3>& gpt_flux_up, const Array_gpu<Float,3>& gpt_flux_dn,
                            const std::unique_ptr<Optical_props_arry_gpu>& spectral_disc, const Bool top_at_1)
        {
            sum(gpt_flux_up, flux_up); sum(gpt_flux_dn, flux_dn); net_flux();
        }
        virtual void reduce(const Array_gpu<Float,3>& gpt_flux_up, const Array_gpu<Float,3>& gpt_flux_dn,
                            const Array_gpu<Float,3>& gpt_flux_dn_dir,
                            const std::unique_ptr<Optical_props_arry_gpu>& spectral_disc, const Bool top_at_1)
        {
            reduce(gpt_flux_up, gpt_flux_dn, spectral_disc, top_at_1);
            sum(gpt_flux_dn_dir, flux_dn_dir);
        }
        void net_flux()
        {
            rrtmgp_b200::check(rrtmgp_b200_net_flux(rrtmgp_b200::dtype, flux_up.size(), flux_dn.ptr(), flux_up.ptr(),
                                                    flux_net.ptr(), rrtmgp_b200::stream()));
        }
        Array_gpu<Float,2>& get_flux_up() { return flux_up; }
        Array_gpu<Float,2>& get_flux_dn() { return flux_dn; }
        Array_gpu<Float,2>& get_flux_dn_dir() { return flux_dn_dir; }
        Array_gpu<Float,2>& get_flux_net() { return flux_net; }
        // include/Fluxes.h:169-172
        virtual Array_gpu<Float,3>& get_bnd_flux_up() { throw std::runtime_error("Band fluxes are not available"); }
        virtual Array_gpu<Float,3>& get_bnd_flux_dn() { throw std::runtime_error("Band fluxes are not available"); }
        virtual Array_gpu<Float,3>& get_bnd_flux_dn_dir() { throw std::runtime_error("Band fluxes are not available"); }
        virtual Array_gpu<Float,3>& get_bnd_flux_net() { throw std::runtime_error("Band fluxes are not available"); }
    protected:
        static void sum(const Array_gpu<Float,3>& g, Array_gpu<Float,2>& f)
        {
            rrtmgp_b200::check(rrtmgp_b200_sum_broadband(rrtmgp_b200::dtype, g.dim(1), g.dim(2), g.dim(3), g.ptr(), f.ptr(),
                                                         rrtmgp_b200::stream()));
        }
        Array_gpu<Float,2> flux_up, flux_dn, flux_dn_dir, flux_net;
};

class Fluxes_byband_gpu : public Fluxes_broadband_gpu
{
    public:
        Fluxes_byband_gpu(const int ncol, const int nlev, const int nbnd)
            : Fluxes_broadband_gpu(ncol, nlev), bnd_flux_up({ncol, nlev, nbnd}), bnd_flux_dn({ncol, nlev, nbnd}),
              bnd_flux_dn_dir({ncol, nlev, nbnd}), bnd_flux_net({ncol, nlev, nbnd}) {}
        void reduce(const Array_gpu<Float,3>& gpt_flux_up, const Array_gpu<Float,3>& gpt_flux_dn,
                    const std::unique_ptr<Optical_props_arry_gpu>& spectral_disc, const Bool top_at_1) override
        {
            Fluxes_broadband_gpu::reduce(gpt_flux_up, gpt_flux_dn, spectral_disc, top_at_1);
            sumb(gpt_flux_up, bnd_flux_up, spectral_disc); sumb(gpt_flux_dn, bnd_flux_dn, spectral_disc);
            rrtmgp_b200::check(rrtmgp_b200_net_flux(rrtmgp_b200::dtype, bnd_flux_up.size(), bnd_flux_dn.ptr(), bnd_flux_up.ptr(),
                                                    bnd_flux_net.ptr(), rrtmgp_b200::stream()));
        }
        void reduce(const Array_gpu<Float,3>& gpt_flux_up, const Array_gpu<Float,3>& gpt_flux_dn,
                    const Array_gpu<Float,3>& gpt_flux_dn_dir,
                    const std::unique_ptr<Optical_props_arry_gpu>& spectral_disc, const Bool top_at_1) override
        {
            reduce(gpt_flux_up, gpt_flux_dn, spectral_disc, top_at_1);
            sum(gpt_flux_dn_dir, flux_dn_dir); sumb(gpt_flux_dn_dir, bnd_flux_dn_dir, spectral_disc);
        }
        Array_gpu<Float,3>& get_bnd_flux_up() override { return bnd_flux_up; }
        Array_gpu<Float,3>& get_bnd_flux_dn() override { return bnd_flux_dn; }
        Array_gpu<Float,3>& get_bnd_flux_dn_dir() override { return bnd_flux_dn_dir; }
        Array_gpu<Float,3>& get_bnd_flux_net() override { return bnd_flux_net; }
    private:
        static void sumb(const Array_gpu<Float,3>& g, Array_gpu<Float,3>& f, const std::unique_ptr<Optical_props_arry_gpu>& d)
        {
            rrtmgp_b200::check(rrtmgp_b200_sum_byband(rrtmgp_b200::dtype, g.dim(1), g.dim(2), g.dim(3), d->get_nband(),
                    d->get_band_lims_gpoint().ptr(), g.ptr(), f.ptr(), rrtmgp_b200::stream()));
        }
        Array_gpu<Float,3> bnd_flux_up, bnd_flux_dn, bnd_flux_dn_dir, bnd_flux_net;
};

// One prepared fused call: the C-ABI arguments of a solver's solve_gpu(switch_fluxes = true, switch_output_optical = false),
// not launched yet.  Radiation_solver_{longwave,shortwave}::prepare_gpu fill it; solve_gpu launches its own straight away;
// rrtmgp_b200::solve_gpu_pair launches a longwave and a shortwave one of the same columns together (no counterpart in the
// reference, whose driver calls the two solvers one after the other, src_test/test_rte_rrtmgp.cu).  The arrays it points to
// must stay alive until the launch.
namespace rrtmgp_b200 {
struct Fused_call
{
    rrtmgp_b200_gas_optics* handle = nullptr;
    rrtmgp_b200_columns c{};
    rrtmgp_b200_fluxes f{};
    std::unique_ptr<Concs_view> concs;       // owns the per-gas pointer lists c.concs points into
};
// Longwave and shortwave of the same columns in one library call: on large domains band i of both runs in one launch with
// longwave and shortwave CTAs co-resident on every SM; small calls run side by side on two streams.  Fluxes are bit-identical
// to lw.solve_gpu(...) followed by sw.solve_gpu(...).
inline void solve_gpu_pair(const Fused_call& lw, const Fused_call& sw)
{
    check(rrtmgp_b200_lwsw_fluxes(lw.handle, &lw.c, &lw.f, sw.handle, &sw.c, &sw.f, stream()));
}
}  // namespace rrtmgp_b200

// ---- Radiation_solver_{longwave,shortwave}::solve_gpu (include_test/Radiation_solver.h:73-88,176-199;
// src_test/Radiation_solver.cu:419-657,690-950).  Same switches and argument order (aerosols are out of
// scope).  With switch_output_optical=false the whole chain runs in the fused band
// kernels (one C-ABI call); with it true the stage-by-stage class path runs so tau/sources are returned.
class Radiation_solver_longwave
{
    public:
        Radiation_solver_longwave(const Gas_concs_gpu& gas_concs, const rrtmgp_b200_kdist_file& kdist,
                                  const rrtmgp_b200_cloud_file* cloud = nullptr)
            : kdist_gpu(std::make_unique<Gas_optics_rrtmgp_gpu>(gas_concs, kdist))
        { if (cloud) cloud_optics_gpu = std::make_unique<Cloud_optics_gpu>(*cloud); }
        // from objects a caller built itself with the reference's positional constructors (what the reference's own
        // constructor does with load_and_init_gas_optics / load_and_init_cloud_optics, src_test/Radiation_solver.cu:404-416)
        explicit Radiation_solver_longwave(Gas_optics_rrtmgp_gpu&& kdist)
            : kdist_gpu(std::make_unique<Gas_optics_rrtmgp_gpu>(std::move(kdist))) {}
        Radiation_solver_longwave(Gas_optics_rrtmgp_gpu&& kdist, Cloud_optics_gpu&& cloud)
            : kdist_gpu(std::make_unique<Gas_optics_rrtmgp_gpu>(std::move(kdist))),
              cloud_optics_gpu(std::make_unique<Cloud_optics_gpu>(std::move(cloud))) {}

        int get_n_gpt_gpu() const { return kdist_gpu->get_ngpt(); }
        int get_n_bnd_gpu() const { return kdist_gpu->get_nband(); }
        const Array<int,2>& get_band_lims_gpoint_gpu() const { return kdist_gpu->get_band_lims_gpoint(); }
        const Array<Float,2>& get_band_lims_wavenumber_gpu() const { return kdist_gpu->get_band_lims_wavenumber(); }
        Gas_optics_rrtmgp_gpu& get_kdist_gpu() { return *kdist_gpu; }
        void set_n_gauss_angles(const int n) { n_gauss_angles = n; }

        // the fused call of solve_gpu(true, switch_cloud_optics, false, switch_output_bnd_fluxes, ...), prepared but not launched
        rrtmgp_b200::Fused_call prepare_gpu(const bool switch_cloud_optics, const bool switch_output_bnd_fluxes, const Gas_concs_gpu& gas_concs,
                       const Array_gpu<Float,2>& p_lay, const Array_gpu<Float,2>& p_lev,
                       const Array_gpu<Float,2>& t_lay, const Array_gpu<Float,2>& t_lev,
                       const Array_gpu<Float,2>& col_dry, const Array_gpu<Float,1>& t_sfc, const Array_gpu<Float,2>& emis_sfc,
                       const Array_gpu<Float,2>& lwp, const Array_gpu<Float,2>& iwp,
                       const Array_gpu<Float,2>& rel, const Array_gpu<Float,2>& dei,
                       Array_gpu<Float,2>& lw_flux_up, Array_gpu<Float,2>& lw_flux_dn, Array_gpu<Float,2>& lw_flux_net,
                       Array_gpu<Float,3>& lw_bnd_flux_up, Array_gpu<Float,3>& lw_bnd_flux_dn, Array_gpu<Float,3>& lw_bnd_flux_net) const
        {
            if (switch_cloud_optics && !cloud_optics_gpu) throw std::runtime_error("cloud optics requested without cloud tables");
            rrtmgp_b200::Fused_call fc;
            fc.handle = kdist_gpu->handle();
            fc.concs = std::make_unique<rrtmgp_b200::Concs_view>(gas_concs, kdist_gpu->get_gas_names());
            rrtmgp_b200_columns& c = fc.c;
            c.ncol = p_lay.dim(1); c.nlay = p_lay.dim(2); c.top_at_1 = -1;
            c.d_play = p_lay.ptr(); c.d_plev = p_lev.ptr(); c.d_tlay = t_lay.ptr(); c.d_tlev = t_lev.ptr();
            c.d_tsfc = t_sfc.ptr(); c.d_col_dry = col_dry.size() ? col_dry.ptr() : nullptr; c.concs = fc.concs->c;
            c.d_sfc_emis = emis_sfc.ptr(); c.n_gauss_angles = n_gauss_angles;
            if (switch_cloud_optics)
            { c.cloud = cloud_optics_gpu->handle(); c.d_lwp = lwp.ptr(); c.d_iwp = iwp.ptr(); c.d_rel = rel.ptr(); c.d_dei = dei.ptr(); }
            rrtmgp_b200_fluxes& f = fc.f;
            f.d_flux_up = lw_flux_up.ptr(); f.d_flux_dn = lw_flux_dn.ptr(); f.d_flux_net = lw_flux_net.ptr();
            if (switch_output_bnd_fluxes)
            { f.d_bnd_flux_up = lw_bnd_flux_up.ptr(); f.d_bnd_flux_dn = lw_bnd_flux_dn.ptr(); f.d_bnd_flux_net = lw_bnd_flux_net.ptr(); }
            return fc;
        }

        void solve_gpu(const bool switch_fluxes, const bool switch_cloud_optics, const bool switch_output_optical,
                       const bool switch_output_bnd_fluxes, const Gas_concs_gpu& gas_concs,
                       const Array_gpu<Float,2>& p_lay, const Array_gpu<Float,2>& p_lev,
                       const Array_gpu<Float,2>& t_lay, const Array_gpu<Float,2>& t_lev,
                       const Array_gpu<Float,2>& col_dry, const Array_gpu<Float,1>& t_sfc, const Array_gpu<Float,2>& emis_sfc,
                       const Array_gpu<Float,2>& lwp, const Array_gpu<Float,2>& iwp,
                       const Array_gpu<Float,2>& rel, const Array_gpu<Float,2>& dei,
                       Array_gpu<Float,3>& tau, Array_gpu<Float,3>& lay_source, Array_gpu<Float,3>& lev_source,
                       Array_gpu<Float,2>& sfc_source,
                       Array_gpu<Float,2>& lw_flux_up, Array_gpu<Float,2>& lw_flux_dn, Array_gpu<Float,2>& lw_flux_net,
                       Array_gpu<Float,3>& lw_bnd_flux_up, Array_gpu<Float,3>& lw_bnd_flux_dn, Array_gpu<Float,3>& lw_bnd_flux_net)
        {
            const int n_col = p_lay.dim(1), n_lay = p_lay.dim(2), n_lev = p_lev.dim(2);
            if (switch_cloud_optics && !cloud_optics_gpu) throw std::runtime_error("cloud optics requested without cloud tables");
            if (!switch_output_optical)
            {
                if (!switch_fluxes) return;
                const rrtmgp_b200::Fused_call fc = prepare_gpu(switch_cloud_optics, switch_output_bnd_fluxes, gas_concs, p_lay, p_lev, t_lay, t_lev,
                        col_dry, t_sfc, emis_sfc, lwp, iwp, rel, dei, lw_flux_up, lw_flux_dn, lw_flux_net, lw_bnd_flux_up, lw_bnd_flux_dn, lw_bnd_flux_net);
                rrtmgp_b200::check(rrtmgp_b200_lw_fluxes(fc.handle, &fc.c, &fc.f, rrtmgp_b200::stream()));
                return;
            }
            // staged path: gas_optics -> [cloud_optics -> add_to] -> rte_lw -> reduce
            std::unique_ptr<Optical_props_arry_gpu> optical_props = std::make_unique<Optical_props_1scl_gpu>(n_col, n_lay, *kdist_gpu);
            Source_func_lw_gpu sources(n_col, n_lay, *kdist_gpu);
            kdist_gpu->gas_optics(p_lay, p_lev, t_lay, t_sfc, gas_concs, optical_props, sources, col_dry, t_lev);
            if (switch_cloud_optics)
            {
                Optical_props_1scl_gpu clouds(n_col, n_lay, *cloud_optics_gpu);
                cloud_optics_gpu->cloud_optics(lwp, iwp, rel, dei, clouds);
                add_to(dynamic_cast<Optical_props_1scl_gpu&>(*optical_props), clouds);
            }
            tau = optical_props->get_tau(); lay_source = sources.get_lay_source();
            lev_source = sources.get_lev_source(); sfc_source = sources.get_sfc_source();
            if (!switch_fluxes) return;
            const std::vector<Float> pl = top_probe(p_lay);
            const Bool top_at_1 = pl[0] < pl[1];
            Array_gpu<Float,3> gpt_flux_up({n_col, n_lev, get_n_gpt_gpu()}), gpt_flux_dn({n_col, n_lev, get_n_gpt_gpu()});
            Rte_lw_gpu::rte_lw(optical_props, top_at_1, sources, emis_sfc, Array_gpu<Float,2>(), gpt_flux_up, gpt_flux_dn, n_gauss_angles);
            std::unique_ptr<Fluxes_broadband_gpu> fl;
            if (switch_output_bnd_fluxes) fl = std::make_unique<Fluxes_byband_gpu>(n_col, n_lev, get_n_bnd_gpu());
            else fl = std::make_unique<Fluxes_broadband_gpu>(n_col, n_lev);
            fl->reduce(gpt_flux_up, gpt_flux_dn, optical_props, top_at_1);
            lw_flux_up = fl->get_flux_up(); lw_flux_dn = fl->get_flux_dn(); lw_flux_net = fl->get_flux_net();
            if (switch_output_bnd_fluxes)
            {
                auto& fb = dynamic_cast<Fluxes_byband_gpu&>(*fl);
                lw_bnd_flux_up = fb.get_bnd_flux_up(); lw_bnd_flux_dn = fb.get_bnd_flux_dn(); lw_bnd_flux_net = fb.get_bnd_flux_net();
            }
        }
        // p_lay(1,1) and p_lay(1,n_lay), as src_test/Radiation_solver.cu:442 reads them
        static std::vector<Float> top_probe(const Array_gpu<Float,2>& p_lay)
        {
            std::vector<Float> v(2);
            rrtmgp_b200::cuda_check(cudaStreamSynchronize(rrtmgp_b200::stream()));
            rrtmgp_b200::cuda_check(cudaMemcpy(&v[0], p_lay.ptr(), sizeof(Float), cudaMemcpyDeviceToHost));
            rrtmgp_b200::cuda_check(cudaMemcpy(&v[1], p_lay.ptr() + (size_t)(p_lay.dim(2) - 1) * p_lay.dim(1), sizeof(Float), cudaMemcpyDeviceToHost));
            return v;
        }
    private:
        std::unique_ptr<Gas_optics_rrtmgp_gpu> kdist_gpu;
        std::unique_ptr<Cloud_optics_gpu> cloud_optics_gpu;
        int n_gauss_angles = 1;
};

class Radiation_solver_shortwave
{
    public:
        Radiation_solver_shortwave(const Gas_concs_gpu& gas_concs, const rrtmgp_b200_kdist_file& kdist,
                                   const rrtmgp_b200_cloud_file* cloud = nullptr, const rrtmgp_b200_aerosol_file* aerosol = nullptr)
            : kdist_gpu(std::make_unique<Gas_optics_rrtmgp_gpu>(gas_concs, kdist))
        {
            if (cloud) cloud_optics_gpu = std::make_unique<Cloud_optics_gpu>(*cloud);
            if (aerosol) aerosol_optics_gpu = std::make_unique<Aerosol_optics_gpu>(*aerosol);
        }
        explicit Radiation_solver_shortwave(Gas_optics_rrtmgp_gpu&& kdist)
            : kdist_gpu(std::make_unique<Gas_optics_rrtmgp_gpu>(std::move(kdist))) {}
        Radiation_solver_shortwave(Gas_optics_rrtmgp_gpu&& kdist, Cloud_optics_gpu&& cloud)
            : kdist_gpu(std::make_unique<Gas_optics_rrtmgp_gpu>(std::move(kdist))),
              cloud_optics_gpu(std::make_unique<Cloud_optics_gpu>(std::move(cloud))) {}
        Radiation_solver_shortwave(Gas_optics_rrtmgp_gpu&& kdist, Cloud_optics_gpu&& cloud, Aerosol_optics_gpu&& aerosol)
            : kdist_gpu(std::make_unique<Gas_optics_rrtmgp_gpu>(std::move(kdist))),
              cloud_optics_gpu(std::make_unique<Cloud_optics_gpu>(std::move(cloud))),
              aerosol_optics_gpu(std::make_unique<Aerosol_optics_gpu>(std::move(aerosol))) {}

        int get_n_gpt_gpu() const { return kdist_gpu->get_ngpt(); }
        int get_n_bnd_gpu() const { return kdist_gpu->get_nband(); }
        Float get_tsi_gpu() const { return kdist_gpu->get_tsi(); }
        const Array<int,2>& get_band_lims_gpoint_gpu() const { return kdist_gpu->get_band_lims_gpoint(); }
        const Array<Float,2>& get_band_lims_wavenumber_gpu() const { return kdist_gpu->get_band_lims_wavenumber(); }
        Gas_optics_rrtmgp_gpu& get_kdist_gpu() { return *kdist_gpu; }

        // the fused call of solve_gpu(true, ..., switch_output_optical = false, ...), prepared but not launched
        rrtmgp_b200::Fused_call prepare_gpu(const bool switch_cloud_optics, const bool switch_aerosol_optics, const bool switch_output_bnd_fluxes,
                       const bool switch_delta_cloud, const bool switch_delta_aerosol, const Gas_concs_gpu& gas_concs,
                       const Array_gpu<Float,2>& p_lay, const Array_gpu<Float,2>& p_lev,
                       const Array_gpu<Float,2>& t_lay, const Array_gpu<Float,2>& t_lev, const Array_gpu<Float,2>& col_dry,
                       const Array_gpu<Float,2>& sfc_alb_dir, const Array_gpu<Float,2>& sfc_alb_dif,
                       const Array_gpu<Float,1>& tsi_scaling, const Array_gpu<Float,1>& mu0,
                       const Array_gpu<Float,2>& lwp, const Array_gpu<Float,2>& iwp,
                       const Array_gpu<Float,2>& rel, const Array_gpu<Float,2>& dei,
                       const Array_gpu<Float,2>& rh, const Aerosol_concs_gpu& aerosol_concs,
                       Array_gpu<Float,2>& sw_flux_up, Array_gpu<Float,2>& sw_flux_dn,
                       Array_gpu<Float,2>& sw_flux_dn_dir, Array_gpu<Float,2>& sw_flux_net,
                       Array_gpu<Float,3>& sw_bnd_flux_up, Array_gpu<Float,3>& sw_bnd_flux_dn,
                       Array_gpu<Float,3>& sw_bnd_flux_dn_dir, Array_gpu<Float,3>& sw_bnd_flux_net) const
        {
            (void)t_lev;
            if (switch_cloud_optics && !cloud_optics_gpu) throw std::runtime_error("cloud optics requested without cloud tables");
            if (switch_aerosol_optics && !aerosol_optics_gpu) throw std::runtime_error("aerosol optics requested without aerosol tables");
            rrtmgp_b200::Fused_call fc;
            fc.handle = kdist_gpu->handle();
            fc.concs = std::make_unique<rrtmgp_b200::Concs_view>(gas_concs, kdist_gpu->get_gas_names());
            rrtmgp_b200_columns& c = fc.c;
            c.ncol = p_lay.dim(1); c.nlay = p_lay.dim(2); c.top_at_1 = -1;
            c.d_play = p_lay.ptr(); c.d_plev = p_lev.ptr(); c.d_tlay = t_lay.ptr();
            c.d_col_dry = col_dry.size() ? col_dry.ptr() : nullptr; c.concs = fc.concs->c;
            c.d_sfc_alb_dir = sfc_alb_dir.ptr(); c.d_sfc_alb_dif = sfc_alb_dif.ptr(); c.d_mu0 = mu0.ptr();
            c.d_tsi_scaling = tsi_scaling.size() ? tsi_scaling.ptr() : nullptr;
            if (switch_cloud_optics)
            {
                c.cloud = cloud_optics_gpu->handle(); c.delta_cloud = switch_delta_cloud;
                c.d_lwp = lwp.ptr(); c.d_iwp = iwp.ptr(); c.d_rel = rel.ptr(); c.d_dei = dei.ptr();
            }
            if (switch_aerosol_optics)
            {
                c.aerosol = aerosol_optics_gpu->handle(); c.delta_aerosol = switch_delta_aerosol;
                c.d_rh = rh.ptr(); c.aerosol_concs = rrtmgp_b200::aerosol_view(aerosol_concs);
            }
            rrtmgp_b200_fluxes& f = fc.f;
            f.d_flux_up = sw_flux_up.ptr(); f.d_flux_dn = sw_flux_dn.ptr(); f.d_flux_net = sw_flux_net.ptr();
            f.d_flux_dn_dir = sw_flux_dn_dir.ptr();
            if (switch_output_bnd_fluxes)
            {
                f.d_bnd_flux_up = sw_bnd_flux_up.ptr(); f.d_bnd_flux_dn = sw_bnd_flux_dn.ptr();
                f.d_bnd_flux_net = sw_bnd_flux_net.ptr(); f.d_bnd_flux_dn_dir = sw_bnd_flux_dn_dir.ptr();
            }
            return fc;
        }

        // the aerosol-free call of earlier callers
        void solve_gpu(const bool switch_fluxes, const bool switch_cloud_optics, const bool switch_output_optical,
                       const bool switch_output_bnd_fluxes, const bool switch_delta_cloud, const Gas_concs_gpu& gas_concs,
                       const Array_gpu<Float,2>& p_lay, const Array_gpu<Float,2>& p_lev,
                       const Array_gpu<Float,2>& t_lay, const Array_gpu<Float,2>& t_lev, const Array_gpu<Float,2>& col_dry,
                       const Array_gpu<Float,2>& sfc_alb_dir, const Array_gpu<Float,2>& sfc_alb_dif,
                       const Array_gpu<Float,1>& tsi_scaling, const Array_gpu<Float,1>& mu0,
                       const Array_gpu<Float,2>& lwp, const Array_gpu<Float,2>& iwp,
                       const Array_gpu<Float,2>& rel, const Array_gpu<Float,2>& dei,
                       Array_gpu<Float,3>& tau, Array_gpu<Float,3>& ssa, Array_gpu<Float,3>& g, Array_gpu<Float,2>& toa_src,
                       Array_gpu<Float,2>& sw_flux_up, Array_gpu<Float,2>& sw_flux_dn,
                       Array_gpu<Float,2>& sw_flux_dn_dir, Array_gpu<Float,2>& sw_flux_net,
                       Array_gpu<Float,3>& sw_bnd_flux_up, Array_gpu<Float,3>& sw_bnd_flux_dn,
                       Array_gpu<Float,3>& sw_bnd_flux_dn_dir, Array_gpu<Float,3>& sw_bnd_flux_net)
        {
            solve_gpu(switch_fluxes, switch_cloud_optics, false, switch_output_optical, switch_output_bnd_fluxes, switch_delta_cloud, false,
                      gas_concs, p_lay, p_lev, t_lay, t_lev, col_dry, sfc_alb_dir, sfc_alb_dif, tsi_scaling, mu0, lwp, iwp, rel, dei,
                      Array_gpu<Float,2>(), Aerosol_concs_gpu(), tau, ssa, g, toa_src, sw_flux_up, sw_flux_dn, sw_flux_dn_dir, sw_flux_net,
                      sw_bnd_flux_up, sw_bnd_flux_dn, sw_bnd_flux_dn_dir, sw_bnd_flux_net);
        }
        // the reference's argument list (src_test/Radiation_solver.cu:684-711)
        void solve_gpu(const bool switch_fluxes, const bool switch_cloud_optics, const bool switch_aerosol_optics,
                       const bool switch_output_optical, const bool switch_output_bnd_fluxes,
                       const bool switch_delta_cloud, const bool switch_delta_aerosol, const Gas_concs_gpu& gas_concs,
                       const Array_gpu<Float,2>& p_lay, const Array_gpu<Float,2>& p_lev,
                       const Array_gpu<Float,2>& t_lay, const Array_gpu<Float,2>& t_lev, const Array_gpu<Float,2>& col_dry,
                       const Array_gpu<Float,2>& sfc_alb_dir, const Array_gpu<Float,2>& sfc_alb_dif,
                       const Array_gpu<Float,1>& tsi_scaling, const Array_gpu<Float,1>& mu0,
                       const Array_gpu<Float,2>& lwp, const Array_gpu<Float,2>& iwp,
                       const Array_gpu<Float,2>& rel, const Array_gpu<Float,2>& dei,
                       const Array_gpu<Float,2>& rh, const Aerosol_concs_gpu& aerosol_concs,
                       Array_gpu<Float,3>& tau, Array_gpu<Float,3>& ssa, Array_gpu<Float,3>& g, Array_gpu<Float,2>& toa_src,
                       Array_gpu<Float,2>& sw_flux_up, Array_gpu<Float,2>& sw_flux_dn,
                       Array_gpu<Float,2>& sw_flux_dn_dir, Array_gpu<Float,2>& sw_flux_net,
                       Array_gpu<Float,3>& sw_bnd_flux_up, Array_gpu<Float,3>& sw_bnd_flux_dn,
                       Array_gpu<Float,3>& sw_bnd_flux_dn_dir, Array_gpu<Float,3>& sw_bnd_flux_net)
        {
            const int n_col = p_lay.dim(1), n_lay = p_lay.dim(2), n_lev = p_lev.dim(2);
            (void)t_lev;
            if (switch_cloud_optics && !cloud_optics_gpu) throw std::runtime_error("cloud optics requested without cloud tables");
            if (switch_aerosol_optics && !aerosol_optics_gpu) throw std::runtime_error("aerosol optics requested without aerosol tables");
            if (!switch_output_optical)
            {
                if (!switch_fluxes) return;
                const rrtmgp_b200::Fused_call fc = prepare_gpu(switch_cloud_optics, switch_aerosol_optics, switch_output_bnd_fluxes, switch_delta_cloud,
                        switch_delta_aerosol, gas_concs, p_lay, p_lev, t_lay, t_lev, col_dry, sfc_alb_dir, sfc_alb_dif, tsi_scaling, mu0, lwp, iwp, rel, dei,
                        rh, aerosol_concs, sw_flux_up, sw_flux_dn, sw_flux_dn_dir, sw_flux_net, sw_bnd_flux_up, sw_bnd_flux_dn, sw_bnd_flux_dn_dir, sw_bnd_flux_net);
                rrtmgp_b200::check(rrtmgp_b200_sw_fluxes(fc.handle, &fc.c, &fc.f, rrtmgp_b200::stream()));
                return;
            }
            std::unique_ptr<Optical_props_arry_gpu> optical_props = std::make_unique<Optical_props_2str_gpu>(n_col, n_lay, *kdist_gpu);
            Array_gpu<Float,2> toa({n_col, get_n_gpt_gpu()});
            kdist_gpu->gas_optics(p_lay, p_lev, t_lay, gas_concs, optical_props, toa, col_dry);
            if (tsi_scaling.size())      // scaling_to_subset, src_test/Radiation_solver.cu:42-70,772
                rrtmgp_b200::check(rrtmgp_b200_scale_columns(rrtmgp_b200::dtype, n_col, get_n_gpt_gpu(), toa.ptr(), tsi_scaling.ptr(),
                                                             rrtmgp_b200::stream()));
            if (switch_cloud_optics)
            {
                Optical_props_2str_gpu clouds(n_col, n_lay, *cloud_optics_gpu);
                cloud_optics_gpu->cloud_optics(lwp, iwp, rel, dei, clouds);
                if (switch_delta_cloud) clouds.delta_scale();
                add_to(dynamic_cast<Optical_props_2str_gpu&>(*optical_props), clouds);
            }
            if (switch_aerosol_optics)      // src_test/Radiation_solver.cu:794-810
            {
                Optical_props_2str_gpu aerosols(n_col, n_lay, *aerosol_optics_gpu);
                aerosol_optics_gpu->aerosol_optics(aerosol_concs, rh, p_lev, aerosols);
                if (switch_delta_aerosol) aerosols.delta_scale();
                add_to(dynamic_cast<Optical_props_2str_gpu&>(*optical_props), aerosols);
            }
            tau = optical_props->get_tau(); ssa = optical_props->get_ssa(); g = optical_props->get_g(); toa_src = toa;
            if (!switch_fluxes) return;
            const std::vector<Float> pl = Radiation_solver_longwave::top_probe(p_lay);
            const Bool top_at_1 = pl[0] < pl[1];
            const int ngpt = get_n_gpt_gpu();
            Array_gpu<Float,3> gpt_flux_up({n_col, n_lev, ngpt}), gpt_flux_dn({n_col, n_lev, ngpt}), gpt_flux_dn_dir({n_col, n_lev, ngpt});
            Rte_sw_gpu::rte_sw(optical_props, top_at_1, mu0, toa, sfc_alb_dir, sfc_alb_dif, Array_gpu<Float,2>(),
                               gpt_flux_up, gpt_flux_dn, gpt_flux_dn_dir);
            std::unique_ptr<Fluxes_broadband_gpu> fl;
            if (switch_output_bnd_fluxes) fl = std::make_unique<Fluxes_byband_gpu>(n_col, n_lev, get_n_bnd_gpu());
            else fl = std::make_unique<Fluxes_broadband_gpu>(n_col, n_lev);
            fl->reduce(gpt_flux_up, gpt_flux_dn, gpt_flux_dn_dir, optical_props, top_at_1);
            sw_flux_up = fl->get_flux_up(); sw_flux_dn = fl->get_flux_dn(); sw_flux_dn_dir = fl->get_flux_dn_dir(); sw_flux_net = fl->get_flux_net();
            if (switch_output_bnd_fluxes)
            {
                auto& fb = dynamic_cast<Fluxes_byband_gpu&>(*fl);
                sw_bnd_flux_up = fb.get_bnd_flux_up(); sw_bnd_flux_dn = fb.get_bnd_flux_dn();
                sw_bnd_flux_dn_dir = fb.get_bnd_flux_dn_dir(); sw_bnd_flux_net = fb.get_bnd_flux_net();
            }
        }
    private:
        std::unique_ptr<Gas_optics_rrtmgp_gpu> kdist_gpu;
        std::unique_ptr<Cloud_optics_gpu> cloud_optics_gpu;
        std::unique_ptr<Aerosol_optics_gpu> aerosol_optics_gpu;
};
