// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE ONLY (never imported by the product path).
//
// Thin extern "C" driver around the UNMODIFIED reference sources (compiled from where they lie under
// /root/reference/cpp by oracle/Makefile into oracle/_ref/libkgref.so). It exposes the reference's own
// chi / get_R / omega / w_integral / transfer lengths / rdf / matrices / transport properties so that
//   (1) the C restatement in oracle/kg_oracle.c can be pinned against the real thing, and
//   (2) golden fixtures under tests/golden/ can be generated (oracle/make_golden.py).
// EoS-derived inputs enter through a table-backed GenericEoS (cpp/eos_interface.h:57-119) because
// thermopack's Fortran core cannot be built in this image.
#include "MieKinGas.h"
#include "QuantumMie.h"
#include "LJSpline.h"
#include "LJTS.h"
#include "multiparam.h"
#include "Integration/Integration.h"
#include <cstring>
#include <set>
#include <mutex>
#ifdef KG_WITH_GPU_SHIM
#include "gpu_saft_eos.h"
#include "gpu_miekingas.h"      // integration/gpu_miekingas.h: the reference-side shim over libkgcuda.so (integration/Makefile)
#endif

namespace {

struct TableEoS {
    int VAPPH{2};
    int nc{0};
    std::vector<double> dmudn_tab;   // nc*nc, J/mol^2, for n = x (1 mol total), V = Vm
    std::vector<double> cp_id;       // nc
    double p{0.}, vm{0.};
    std::vector<double> dvdn_tab;    // nc
    std::vector<std::vector<double>> dmudn(double, double, const std::vector<double>) const {
        std::vector<std::vector<double>> out(nc, std::vector<double>(nc));
        for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) out[i][j] = dmudn_tab[i * nc + j];
        return out;
    }
    double Cp_ideal(double, int c) const { return cp_id.at(c - 1); }
    double Cp_real(double, double, const std::vector<double>) const { throw std::runtime_error("TableEoS: Cp_real"); }
    double pressure_tv(double, double, const std::vector<double>) const { return p; }
    double specific_volume(double, double, const std::vector<double>, int) const { return vm; }
    std::vector<double> dvdn(double, double, const std::vector<double>, int) const { return dvdn_tab; }
};

struct Handle {
    std::unique_ptr<Spherical> kin;
    MieKinGas* mie = nullptr;          // set when kin is a MieKinGas
    QuantumMie* qmie = nullptr;        // set when kin is a QuantumMie
    std::function<void(double)> set_T; // T hook of Feynman-Hibbs corrected models (FH_Corrected::set_internals)
    std::function<void()> clear;       // cache.clear() of the non-Mie models
    std::function<void(double*)> scales;
    int nc;
    std::string err;
};

vector2d to2d(const double* a, int n) {
    vector2d v(n, vector1d(n));
    for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) v[i][j] = a[i * n + j];
    return v;
}

template<class F> int guarded(Handle* h, F&& f) {
    try { f(); return 0; }
    catch (const std::exception& e) { h->err = e.what(); return -1; }
}

// Expose protected/private-free members we need through a derived type
struct MieAccess : public MieKinGas {
    using MieKinGas::MieKinGas;
    using KineticGas::cache;
    void clear_cache() { cache.clear(); }
};

// set_internals / cache are protected: open them up per model type
template<class B> struct Open : public B {
    using B::B;
    using B::set_internals;
    void clear_cache() { this->cache.clear(); }
    void scales(double* o) { o[0] = this->sigma[0][0]; o[1] = this->eps[0][0]; o[2] = this->m[0]; o[3] = pow(HBAR, 2) / (24 * this->red_mass[0][0]); }
};
template<class M, class... A> Handle* make_open(int nc, A&&... args) {
    Handle* h = new Handle;
    h->nc = nc;
    auto* k = new Open<M>(std::forward<A>(args)...);
    h->kin.reset(k);
    h->set_T = [k](double T) { k->set_internals(0., T, vector1d{1.}); };
    h->clear = [k]() { k->clear_cache(); };
    h->scales = [k](double* o) { k->scales(o); };
    return h;
}

} // namespace

extern "C" {

// kind: 1 LJSpline, 2 LJTS, 3 ModTangToennis / FH_ModTangToennies, 4 HFD_B2 / FH_HFD_B2, 5 Patowski / PatowskiFH
// (cpp/LJSpline.h, cpp/LJTS.h, cpp/multiparam.h:36-131). comps names the fluid file; fh_order 0 = uncorrected class with
// set_quantum_active(false), so that omega is the classical Spherical::omega (cpp/Quantum.cpp:1261-1268).
void* kgref_generic_create(int kind, const char* comps, int fh_order, char* errbuf, int errcap) {
    try {
        Handle* h = nullptr;
        if (kind == 1) h = make_open<LJSpline>(2, false, true);
        else if (kind == 2) {
            const double s = 3.42e-10, e = 1.71200476e-21;
            h = make_open<LJTS>(2, vector1d{6.64e-26, 6.64e-26}, vector2d{{s, s}, {s, s}}, vector2d{{e, e}, {e, e}}, true, true);
        }
        else if (kind == 3 && fh_order == 0) { h = make_open<ModTangToennis>(2, std::string(comps), std::string("default")); static_cast<Quantum*>(h->kin.get())->set_quantum_active(false); }
        else if (kind == 3) h = make_open<FH_ModTangToennies>(2, std::string(comps), (size_t) fh_order, std::string("default"));
        else if (kind == 4 && fh_order == 0) { h = make_open<HFD_B2>(2, std::string(comps)); static_cast<Quantum*>(h->kin.get())->set_quantum_active(false); }
        else if (kind == 4) h = make_open<FH_HFD_B2>(2, std::string(comps), (size_t) fh_order);
        else if (kind == 5 && fh_order == 0) { h = make_open<Patowski>(2, std::string(comps)); static_cast<Quantum*>(h->kin.get())->set_quantum_active(false); }
        else if (kind == 5) h = make_open<PatowskiFH>(2, std::string(comps), (size_t) fh_order);
        else throw std::runtime_error("unknown kind");
        return h;
    } catch (const std::exception& e) { if (errbuf) { strncpy(errbuf, e.what(), errcap - 1); errbuf[errcap - 1] = 0; } return nullptr; }
}
// sigma, eps, mass of component 0 and the FH D factor hbar^2 / (24 mu_00) (cpp/extentions.h:308-314)
int kgref_generic_scales(void* hv, double* out4) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ if (!h->scales) throw std::runtime_error("not a generic model"); h->scales(out4); });
}
// ---- quantum-mechanical collision integrals (cpp/Quantum.cpp): phase shifts, cross sections; omega follows the active switch ----
static Quantum* as_quantum(void* hv) {
    Quantum* q = dynamic_cast<Quantum*>(static_cast<Handle*>(hv)->kin.get());
    if (!q) throw std::runtime_error("not a Quantum model");
    return q;
}
int kgref_set_quantum_active(void* hv, int active) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ as_quantum(hv)->set_quantum_active(active != 0); });
}
int kgref_phase_shift(void* hv, int i, int j, int l, double E_red, int which, double* out) {      // which: 0 phase_shift, 1 Numerov, 2 JKWB
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{
        Quantum* q = as_quantum(hv);
        double r_lev = 0;
        *out = (which == 0) ? q->phase_shift(i, j, l, E_red) : (which == 1) ? q->quantum_phase_shift(i, j, l, E_red, r_lev) : q->JKWB_phase_shift(i, j, l, E_red);
    });
}
int kgref_quantum_cross_section(void* hv, int i, int j, int n, double E_red, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ *out = as_quantum(hv)->cross_section(i, j, n, E_red); });
}
int kgref_symmetry_weights(void* hv, int i, int j, double* out2) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ auto w = as_quantum(hv)->get_symmetry_weights(i, j); out2[0] = w[0]; out2[1] = w[1]; });
}
int kgref_r_classical_forbidden(void* hv, int i, int j, int l, double E_red, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ *out = as_quantum(hv)->r_classical_forbidden(i, j, l, E_red); });
}
double kgref_potential_rr(void* hv, int i, int j, double r) { return static_cast<Handle*>(hv)->kin->potential_dblderivative_rr(i, j, r); }

void* kgref_mie_create(int nc, const double* m, const double* sigma, const double* eps,
                       const double* la, const double* lr, int is_idealgas, int is_singlecomp) {
    Handle* h = new Handle;
    h->nc = nc;
    MieAccess* mk = new MieAccess(vector1d(m, m + nc), to2d(sigma, nc), to2d(eps, nc), to2d(la, nc), to2d(lr, nc),
                                  is_idealgas != 0, is_singlecomp != 0);
    h->kin.reset(mk);
    h->mie = mk;
    return h;
}
// QuantumMie through its vector constructor (cpp/QuantumMie.cpp:5-49)
void* kgref_qmie_create(int nc, const double* m, const double* sigma, const double* eps, const double* la, const double* lr,
                        int fh_order, int is_idealgas, int is_singlecomp) {
    Handle* h = new Handle;
    h->nc = nc;
    QuantumMie* q = new QuantumMie(vector1d(m, m + nc), to2d(sigma, nc), to2d(eps, nc), to2d(la, nc), to2d(lr, nc),
                                   std::vector<int>(nc, fh_order), is_idealgas != 0, is_singlecomp != 0);
    h->kin.reset(q);
    h->qmie = q;
    return h;
}
// T-dependent potentials: ExtSutherland::set_internals -> set_effective_params (cpp/ExtendedSutherland.h:140, .cpp:168-180)
int kgref_set_internals(void* hv, double rho, double T) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ if (h->qmie) h->qmie->set_internals(rho, T, vector1d(h->nc, 1.0 / h->nc)); else if (h->set_T) h->set_T(T); });
}
// effective coefficients C_eff[k][i][j] after set_internals (what the device path receives as coeff_T)
int kgref_c_eff(void* hv, int i, int j, double* out, int* nterms) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{
        if (!h->qmie) throw std::runtime_error("not an ExtSutherland model");
        *nterms = (int) h->qmie->nterms;
        for (size_t k = 0; k < h->qmie->nterms; k++) out[k] = static_cast<double>(h->qmie->C_eff[k][i][j]);
    });
}
void kgref_destroy(void* hv) { delete static_cast<Handle*>(hv); }
const char* kgref_last_error(void* hv) { return static_cast<Handle*>(hv)->err.c_str(); }
void kgref_clear_cache(void* hv) { Handle* h = static_cast<Handle*>(hv); if (h->clear) h->clear(); else if (h->mie) static_cast<MieAccess*>(h->mie)->clear_cache(); }
void kgref_set_omega_correlation_active(void* hv, int active) { Handle* h = static_cast<Handle*>(hv); if (h->mie) h->mie->set_omega_correlation_active(active != 0); }
void kgref_set_tl_model(void* hv, int id) { static_cast<Handle*>(hv)->kin->set_transfer_length_model(id); }

void kgref_set_eos_tables(void* hv, const double* dmudn, const double* cp_id, double p, double vm, const double* dvdn) {
    Handle* h = static_cast<Handle*>(hv);
    TableEoS e;
    e.nc = h->nc;   // note: for single-component models the EoS sees 1 component; caller passes matching tables
    if (dmudn) e.dmudn_tab.assign(dmudn, dmudn + h->nc * h->nc);
    if (cp_id) e.cp_id.assign(cp_id, cp_id + h->nc);
    if (dvdn) e.dvdn_tab.assign(dvdn, dvdn + h->nc);
    e.p = p; e.vm = vm;
    h->kin->set_eos(GenericEoS(std::move(e)));
}
void kgref_set_eos_tables_nc(void* hv, int nc_eos, const double* dmudn, const double* cp_id, double p, double vm, const double* dvdn) {
    Handle* h = static_cast<Handle*>(hv);
    TableEoS e;
    e.nc = nc_eos;
    if (dmudn) e.dmudn_tab.assign(dmudn, dmudn + nc_eos * nc_eos);
    if (cp_id) e.cp_id.assign(cp_id, cp_id + nc_eos);
    if (dvdn) e.dvdn_tab.assign(dvdn, dvdn + nc_eos);
    e.p = p; e.vm = vm;
    h->kin->set_eos(GenericEoS(std::move(e)));
}

// ---- collision dynamics (cpp/Spherical.cpp) ----
double kgref_potential(void* hv, int i, int j, double r) { return static_cast<Handle*>(hv)->kin->potential(i, j, r); }
double kgref_potential_r(void* hv, int i, int j, double r) { return static_cast<Handle*>(hv)->kin->potential_derivative_r(i, j, r); }
double kgref_chi(void* hv, int i, int j, double T, double g, double b) { return static_cast<Handle*>(hv)->kin->chi(i, j, T, g, b); }
double kgref_theta(void* hv, int i, int j, double T, double g, double b) { return static_cast<Handle*>(hv)->kin->theta(i, j, T, g, b); }
double kgref_get_R(void* hv, int i, int j, double T, double g, double b) { return static_cast<Handle*>(hv)->kin->get_R(i, j, T, g, b); }
int kgref_chi_batch(void* hv, int i, int j, double T, int n, const double* g, const double* b, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ for (int k = 0; k < n; k++) out[k] = h->kin->chi(i, j, T, g[k], b[k]); });
}
int kgref_get_R_batch(void* hv, int i, int j, double T, int n, const double* g, const double* b, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ for (int k = 0; k < n; k++) out[k] = h->kin->get_R(i, j, T, g[k], b[k]); });
}
double kgref_w_integrand(void* hv, int i, int j, double T, double g, double b, int l, int r) {
    return static_cast<Handle*>(hv)->kin->w_integrand(i, j, T, g, b, l, r);
}
// classical cross sections (cpp/Spherical.cpp:104-127) and the classical second virial coefficient (:140-160)
int kgref_cross_section(void* hv, int i, int j, int l, double E, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ *out = h->kin->cross_section(i, j, l, E); });
}
int kgref_second_virial(void* hv, int i, int j, double T, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ *out = h->kin->second_virial(i, j, T); });
}
// omega through the normal (cached, correlation-aware) path: MieKinGas::omega -> Spherical::omega
int kgref_omega(void* hv, int i, int j, int l, int r, double T, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ *out = h->kin->Spherical::omega(i, j, l, r, T); });
}
int kgref_omega_model(void* hv, int i, int j, int l, int r, double T, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ *out = h->kin->omega(i, j, l, r, T); });
}
// dimensionless W via the tester entry (no cache, explicit lattice)
int kgref_w_integral_tester(void* hv, int i, int j, double T, int l, int r,
                            double end_g, double end_b, double dg, double db, int rlg, int rlb, double lim, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{
        IntegrationParam p({1e-7, 1e-7}, {end_g, end_b}, dg, db, rlg, rlb, lim);
        *out = h->kin->w_integral_tester(i, j, T, l, r, p);
    });
}
// Same integral, but through integrate2d with a recording integrand: returns the lattice points
// (first-visit coordinates) the reference's adaptive mesh evaluates. pts: [max_pts][3] = (g, b, value)
int kgref_w_integral_trace(void* hv, int i, int j, double T, int l, int r,
                           double end_g, double end_b, double dg, double db, int rlg, int rlb, double lim,
                           double* out, double* pts, int max_pts, int* npts) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{
        int n = 0;
        auto f = [&](double g, double b) {
            double v = h->kin->w_integrand(i, j, T, g, b, l, r);
            if (n < max_pts) { pts[3 * n] = g; pts[3 * n + 1] = b; pts[3 * n + 2] = v; }
            n++;
            return v;
        };
        *out = integrate2d(Point{1e-7, 1e-7}, Point{end_g, end_b}, dg, db, rlg, rlb, lim, f);
        *npts = n;
    });
}
// integrate2d on a caller-supplied table f[Nx][Ny] (pins the mesh replay independently of chi)
double kgref_integrate2d_table(const double* tab, int nx, int ny, double ox, double oy, double ex, double ey,
                               double dx, double dy, int rlx, int rly, double lim, int* nevals) {
    int n = 0;
    auto f = [&](double x, double y) {
        int ix = (int) ((x - ox) / dx + 0.5), iy = (int) ((y - oy) / dy + 0.5);
        n++;
        if (ix < 0 || iy < 0 || ix >= nx || iy >= ny) return 0.0;
        return tab[ix * ny + iy];
    };
    double v = integrate2d(Point{ox, oy}, Point{ex, ey}, dx, dy, rlx, rly, lim, f);
    if (nevals) *nevals = n;
    return v;
}
double kgref_tanh_sinh_test(int which, double a, double h_, double tol) {
    // which=0: f(u)=1/sqrt(1-u*u) ; which=1: f(u)=exp(a*u)
    if (which == 0) return tanh_sinh([&](double u){ return 1.0 / sqrt(1.0 - u * u); }, h_, tol);
    return tanh_sinh([&](double u){ return exp(a * u); }, h_, tol);
}

// ---- state-point level (cpp/KineticGas.cpp, KineticGas_properties.cpp, transfer_lengths.cpp, MieKinGas.cpp) ----
static void flat(const vector2d& v, double* out) { size_t k = 0; for (auto& r : v) for (double x : r) out[k++] = x; }
int kgref_mtl(void* hv, double rho, double T, const double* x, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ flat(h->kin->get_mtl(rho, T, vector1d(x, x + h->nc)), out); });
}
int kgref_etl(void* hv, double rho, double T, const double* x, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ flat(h->kin->get_etl(rho, T, vector1d(x, x + h->nc)), out); });
}
int kgref_rdf(void* hv, double rho, double T, const double* x, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ flat(h->kin->get_rdf(rho, T, vector1d(x, x + h->nc)), out); });
}
int kgref_bh_diameters(void* hv, double T, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ if (!h->mie) throw std::runtime_error("not a MieKinGas model"); flat(h->mie->get_BH_diameters(T), out); });
}
int kgref_K_factors(void* hv, double rho, double T, const double* x, double* K, double* Kp) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{
        vector1d xv(x, x + h->nc);
        vector1d a = h->kin->get_K_factors(rho, T, xv), b = h->kin->get_K_prime_factors(rho, T, xv);
        for (int i = 0; i < h->nc; i++) { K[i] = a[i]; Kp[i] = b[i]; }
    });
}
// which: 0 conductivity, 1 diffusion, 2 viscosity. mat is (dim x dim) row-major, vec dim. returns dim via *dim
int kgref_matrix_vector(void* hv, int which, double rho, double T, const double* x, int N, double* mat, double* vec, int* dim) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{
        vector1d xv(x, x + h->nc);
        vector2d M; vector1d v;
        if (which == 0) { M = h->kin->get_conductivity_matrix(rho, T, xv, N); v = h->kin->get_conductivity_vector(rho, T, xv, N); }
        else if (which == 1) { M = h->kin->get_diffusion_matrix(rho, T, xv, N); v = h->kin->get_diffusion_vector(rho, T, xv, N); }
        else { M = h->kin->get_viscosity_matrix(rho, T, xv, N); v = h->kin->get_viscosity_vector(rho, T, xv, N); }
        *dim = (int) v.size();
        if (mat) flat(M, mat);
        if (vec) for (size_t k = 0; k < v.size(); k++) vec[k] = v[k];
    });
}
int kgref_viscosity(void* hv, double T, double Vm, const double* x, int nx, int N, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ *out = h->kin->viscosity(T, Vm, vector1d(x, x + nx), N); });
}
int kgref_thermal_conductivity(void* hv, double T, double Vm, const double* x, int nx, int N, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{ *out = h->kin->thermal_conductivity(T, Vm, vector1d(x, x + nx), N); });
}
int kgref_thermal_conductivity_contribs(void* hv, double T, double Vm, const double* x, int nx, int N, double* out3) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{
        auto m = h->kin->thermal_conductivity_contributions(T, Vm, vector1d(x, x + nx), N, "tdi");
        out3[0] = m["t"]; out3[1] = m["d"]; out3[2] = m["i"];
    });
}
int kgref_interdiffusion(void* hv, double T, double Vm, const double* x, int nx, int N, int frame, int dep, int solvent,
                         int do_compress, double* out, int* rows, int* cols) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{
        Eigen::MatrixXd D = h->kin->interdiffusion(T, Vm, vector1d(x, x + nx), N, frame, dep, solvent, do_compress != 0);
        *rows = (int) D.rows(); *cols = (int) D.cols();
        for (int i = 0; i < D.rows(); i++) for (int j = 0; j < D.cols(); j++) out[i * D.cols() + j] = D(i, j);
    });
}
int kgref_thermal_diffusion_coeff(void* hv, double T, double Vm, const double* x, int nx, int N, int frame, int dep, int solvent,
                                  double* out, int* n) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{
        Eigen::VectorXd D = h->kin->thermal_diffusion_coeff(T, Vm, vector1d(x, x + nx), N, frame, dep, solvent);
        *n = (int) D.size();
        for (int i = 0; i < D.size(); i++) out[i] = D(i);
    });
}

int kgref_thermal_diffusion_ratio(void* hv, double T, double Vm, const double* x, int nx, int N, double* out, int* n) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{
        Eigen::VectorXd k = h->kin->thermal_diffusion_ratio(T, Vm, vector1d(x, x + nx), N);
        *n = (int) k.size();
        for (int i = 0; i < k.size(); i++) out[i] = k(i);
    });
}

int kgref_soret_coefficient(void* hv, double T, double Vm, const double* x, int nx, int N, int dep, double* out, int* n) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{
        Eigen::VectorXd k = h->kin->soret_coefficient(T, Vm, vector1d(x, x + nx), N, dep);
        *n = (int) k.size();
        for (int i = 0; i < k.size(); i++) out[i] = k(i);
    });
}

} // extern "C"

#ifdef KG_WITH_GPU_SHIM
// integration/_build/libkgshim.so only: the same driver over a GpuMieKinGas (the reference's MieKinGas with precompute_* and the
// transfer-length entry routed to the device). Every kgref_* property call works on the returned handle unchanged.
extern "C" {
// installs the device SAFT-VR Mie EoS (integration/gpu_saft_eos.h) into any handle through the reference's own set_eos
// (cpp/eos_interface.h:57-135): pair matrices [nc_eos][nc_eos], Cp correlations as thermopack's database holds them
int kgshim_set_gpu_eos(void* hv, int nc_eos, const double* m, const double* sigma, const double* eps_div_k, const double* la, const double* lr,
                       const int32_t* cp_type, const double* cp_par, const double* mw, char* errbuf, int errcap) {
    try {
        Handle* h = static_cast<Handle*>(hv);
        const size_t n2 = (size_t) nc_eos * nc_eos;
        GpuSaftEoS e(nc_eos, vector1d(m, m + nc_eos), vector1d(sigma, sigma + n2), vector1d(eps_div_k, eps_div_k + n2), vector1d(la, la + n2),
                     vector1d(lr, lr + n2), std::vector<int32_t>(cp_type, cp_type + nc_eos), vector1d(cp_par, cp_par + (size_t) nc_eos * KG_EOS_CP_PARAMS),
                     vector1d(mw, mw + nc_eos));
        h->kin->set_eos(GenericEoS(std::move(e)));
        return 0;
    } catch (const std::exception& e) { if (errbuf) { strncpy(errbuf, e.what(), errcap - 1); errbuf[errcap - 1] = 0; } return -1; }
}
void* kgshim_mie_create(int nc, const double* m, const double* sigma, const double* eps, const double* la, const double* lr,
                        int is_idealgas, int is_singlecomp, int device, char* errbuf, int errcap) {
    try {
        Handle* h = new Handle;
        h->nc = nc;
        GpuMieKinGas* mk = new GpuMieKinGas(vector1d(m, m + nc), to2d(sigma, nc), to2d(eps, nc), to2d(la, nc), to2d(lr, nc),
                                            is_idealgas != 0, is_singlecomp != 0, device);
        h->kin.reset(mk);
        h->mie = mk;
        h->clear = [mk]() { mk->clear_cache(); };
        return h;
    } catch (const std::exception& e) { if (errbuf) { strncpy(errbuf, e.what(), errcap - 1); errbuf[errcap - 1] = 0; } return nullptr; }
}
// out[0] CPU theta-integrand evaluations, out[1] device feeds, out[2] Omega values stored into the reference cache
int kgshim_counters(void* hv, long* out3, int reset) {
    Handle* h = static_cast<Handle*>(hv);
    GpuMieKinGas* g = dynamic_cast<GpuMieKinGas*>(h->kin.get());
    if (!g) return -1;
    out3[0] = g->cpu_theta_integrand_calls(); out3[1] = g->device_feeds(); out3[2] = g->omega_values_stored();
    if (reset) g->reset_counters();
    return 0;
}
int kgshim_omega_table(void* hv, int nT, const double* T, int nlr, const int* l, const int* r, double* out) {
    Handle* h = static_cast<Handle*>(hv);
    return guarded(h, [&]{
        GpuMieKinGas* g = dynamic_cast<GpuMieKinGas*>(h->kin.get());
        if (!g) throw std::runtime_error("not a GpuMieKinGas");
        const std::vector<double> om = g->omega_table(std::vector<double>(T, T + nT), std::vector<int>(l, l + nlr), std::vector<int>(r, r + nlr));
        std::copy(om.begin(), om.end(), out);
    });
}
}
#endif
