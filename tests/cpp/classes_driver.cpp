// Test driver for include/rrtmgp_b200/classes.hpp: runs the reference's solve_gpu call sequence
// (src_test/test_rte_rrtmgp.cu:560-700 longwave, :730-900 shortwave) through the C++ class mirror,
// on host arrays handed in from pytest.  Built twice: Float=double and -DRTE_USE_SP.
// Test infrastructure only; not part of the shipped library.
#include <cstring>
#include "rrtmgp_b200/classes.hpp"

namespace
{
    template<int N> Array_gpu<Float,N> up(const double* h, const std::array<int,N>& dims)
    {
        size_t n = 1; for (int d : dims) n *= (size_t)d;
        std::vector<Float> v(n);
        for (size_t i = 0; i < n; ++i) v[i] = (Float)h[i];
        return Array_gpu<Float,N>(v, dims);
    }
    template<int N> void down(const Array_gpu<Float,N>& a, double* h)
    {
        if (!h) return;
        const std::vector<Float> v = a.to_host();
        for (size_t i = 0; i < v.size(); ++i) h[i] = (double)v[i];
    }
    Gas_concs_gpu make_concs(int ngas, const char* names, const double* const* vmr, const int* vnc, const int* vnl)
    {
        Gas_concs_gpu g;
        for (int i = 0; i < ngas; ++i)
        {
            std::string s(names + 32 * i, 32);
            s = s.substr(0, s.find_last_not_of(' ') + 1);
            std::vector<Float> v((size_t)vnc[i] * vnl[i]);
            for (size_t k = 0; k < v.size(); ++k) v[k] = (Float)vmr[i][k];
            if (v.size() == 1) g.set_vmr(s, v[0]); else g.set_vmr(s, v, vnc[i], vnl[i]);
        }
        return g;
    }
    int fail(const std::exception& e, char* err, int errlen, int code)
    {
        if (err && errlen > 0) { std::strncpy(err, e.what(), errlen - 1); err[errlen - 1] = 0; }
        return code;
    }
}

extern "C" int classes_float_bytes() { return (int)sizeof(Float); }

// mode: 0 fused (switch_output_optical=false), 1 staged (optical properties returned too)
extern "C" int classes_run_lw(const rrtmgp_b200_kdist_file* kd, const rrtmgp_b200_cloud_file* cloud, int mode, int bnd, int n_ang,
        int ncol, int nlay, int ngas, const char* names, const double* const* vmr, const int* vnc, const int* vnl,
        const double* play, const double* plev, const double* tlay, const double* tlev, const double* tsfc, const double* emis,
        const double* lwp, const double* iwp, const double* rel, const double* dei,
        double* tau, double* lay_src, double* lev_src, double* sfc_src,
        double* flux_up, double* flux_dn, double* flux_net, double* bnd_up, double* bnd_dn, double* bnd_net, char* err, int errlen)
{
    try
    {
        const Gas_concs_gpu gas_concs = make_concs(ngas, names, vmr, vnc, vnl);
        Radiation_solver_longwave rad_lw(gas_concs, *kd, cloud);
        rad_lw.set_n_gauss_angles(n_ang);
        const int nlev = nlay + 1, ngpt = rad_lw.get_n_gpt_gpu(), nbnd = rad_lw.get_n_bnd_gpu();
        auto p_lay = up<2>(play, {ncol, nlay}); auto p_lev = up<2>(plev, {ncol, nlev});
        auto t_lay = up<2>(tlay, {ncol, nlay}); auto t_lev = up<2>(tlev, {ncol, nlev});
        auto t_sfc = up<1>(tsfc, {ncol}); auto emis_sfc = up<2>(emis, {nbnd, ncol});
        Array_gpu<Float,2> col_dry, a_lwp, a_iwp, a_rel, a_dei;
        if (cloud) { a_lwp = up<2>(lwp, {ncol, nlay}); a_iwp = up<2>(iwp, {ncol, nlay}); a_rel = up<2>(rel, {ncol, nlay}); a_dei = up<2>(dei, {ncol, nlay}); }
        Array_gpu<Float,3> a_tau, a_lay, a_lev; Array_gpu<Float,2> a_sfc;
        Array_gpu<Float,2> f_up({ncol, nlev}), f_dn({ncol, nlev}), f_net({ncol, nlev});
        Array_gpu<Float,3> b_up({ncol, nlev, bnd ? nbnd : 0}), b_dn({ncol, nlev, bnd ? nbnd : 0}), b_net({ncol, nlev, bnd ? nbnd : 0});
        rad_lw.solve_gpu(true, cloud != nullptr, mode == 1, bnd != 0, gas_concs, p_lay, p_lev, t_lay, t_lev, col_dry, t_sfc, emis_sfc,
                         a_lwp, a_iwp, a_rel, a_dei, a_tau, a_lay, a_lev, a_sfc, f_up, f_dn, f_net, b_up, b_dn, b_net);
        if (mode == 1) { down(a_tau, tau); down(a_lay, lay_src); down(a_lev, lev_src); down(a_sfc, sfc_src); }
        down(f_up, flux_up); down(f_dn, flux_dn); down(f_net, flux_net);
        if (bnd) { down(b_up, bnd_up); down(b_dn, bnd_dn); down(b_net, bnd_net); }
        (void)ngpt;
        return 0;
    }
    catch (const std::range_error& e) { return fail(e, err, errlen, 2); }
    catch (const std::exception& e) { return fail(e, err, errlen, 1); }
}

extern "C" int classes_run_sw(const rrtmgp_b200_kdist_file* kd, const rrtmgp_b200_cloud_file* cloud, int mode, int bnd, int delta_cloud,
        int ncol, int nlay, int ngas, const char* names, const double* const* vmr, const int* vnc, const int* vnl,
        const double* play, const double* plev, const double* tlay, const double* alb_dir, const double* alb_dif,
        const double* tsi_scaling, const double* mu0,
        const double* lwp, const double* iwp, const double* rel, const double* dei,
        double* tau, double* ssa, double* g, double* toa_src,
        double* flux_up, double* flux_dn, double* flux_dn_dir, double* flux_net,
        double* bnd_up, double* bnd_dn, double* bnd_dn_dir, double* bnd_net, char* err, int errlen)
{
    try
    {
        const Gas_concs_gpu gas_concs = make_concs(ngas, names, vmr, vnc, vnl);
        Radiation_solver_shortwave rad_sw(gas_concs, *kd, cloud);
        const int nlev = nlay + 1, nbnd = rad_sw.get_n_bnd_gpu();
        auto p_lay = up<2>(play, {ncol, nlay}); auto p_lev = up<2>(plev, {ncol, nlev}); auto t_lay = up<2>(tlay, {ncol, nlay});
        Array_gpu<Float,2> t_lev, col_dry, a_lwp, a_iwp, a_rel, a_dei;
        auto a_dir = up<2>(alb_dir, {nbnd, ncol}); auto a_dif = up<2>(alb_dif, {nbnd, ncol});
        auto a_tsi = up<1>(tsi_scaling, {ncol}); auto a_mu0 = up<1>(mu0, {ncol});
        if (cloud) { a_lwp = up<2>(lwp, {ncol, nlay}); a_iwp = up<2>(iwp, {ncol, nlay}); a_rel = up<2>(rel, {ncol, nlay}); a_dei = up<2>(dei, {ncol, nlay}); }
        Array_gpu<Float,3> a_tau, a_ssa, a_g; Array_gpu<Float,2> a_toa;
        Array_gpu<Float,2> f_up({ncol, nlev}), f_dn({ncol, nlev}), f_dir({ncol, nlev}), f_net({ncol, nlev});
        const int nb = bnd ? nbnd : 0;
        Array_gpu<Float,3> b_up({ncol, nlev, nb}), b_dn({ncol, nlev, nb}), b_dir({ncol, nlev, nb}), b_net({ncol, nlev, nb});
        rad_sw.solve_gpu(true, cloud != nullptr, mode == 1, bnd != 0, delta_cloud != 0, gas_concs, p_lay, p_lev, t_lay, t_lev, col_dry,
                         a_dir, a_dif, a_tsi, a_mu0, a_lwp, a_iwp, a_rel, a_dei, a_tau, a_ssa, a_g, a_toa,
                         f_up, f_dn, f_dir, f_net, b_up, b_dn, b_dir, b_net);
        if (mode == 1) { down(a_tau, tau); down(a_ssa, ssa); down(a_g, g); down(a_toa, toa_src); }
        down(f_up, flux_up); down(f_dn, flux_dn); down(f_dir, flux_dn_dir); down(f_net, flux_net);
        if (bnd) { down(b_up, bnd_up); down(b_dn, bnd_dn); down(b_dir, bnd_dn_dir); down(b_net, bnd_net); }
        return 0;
    }
    catch (const std::range_error& e) { return fail(e, err, errlen, 2); }
    catch (const std::exception& e) { return fail(e, err, errlen, 1); }
}
