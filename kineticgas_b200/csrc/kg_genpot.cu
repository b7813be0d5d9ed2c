// kg_genpot.cu -- host side of the generic potentials: builds the piecewise P(r) exp(Q(r)) description of each supported
// potential from its parameter block (the numbers the reference's C++ class holds) and folds the Feynman-Hibbs correction
// into the prefactors with exact Laurent-polynomial algebra, once per temperature. No device code in this file.
//
// Reference constructors mirrored: LJSpline (cpp/LJSpline.h:15-25,45-82), LJTS (cpp/LJTS.h:15-43),
// ModTangToennis (cpp/multiparam.cpp:5-30), HFD_B2 (cpp/multiparam.cpp:91-131), Patowski (cpp/multiparam.cpp:184-226),
// FH_Corrected<T>::potential / potential_derivative_r (cpp/extentions.h:228-255).
#include "kg_internal.cuh"
#include <cmath>
#include <cstring>
#include <limits>

namespace {

// Laurent polynomial sum_k c[k] r^(kmin+k)
struct Laurent {
    int kmin = 0;
    std::vector<double> c;
    Laurent() {}
    Laurent(int kmin_, std::vector<double> c_) : kmin(kmin_), c(std::move(c_)) {}
    bool empty() const { return c.empty(); }
    int kmax() const { return kmin + (int) c.size() - 1; }
    double coef(int k) const { return (k < kmin || k > kmax()) ? 0.0 : c[k - kmin]; }
    void trim() {
        while (!c.empty() && c.back() == 0.0) c.pop_back();
        size_t lead = 0;
        while (lead < c.size() && c[lead] == 0.0) lead++;
        if (lead) { c.erase(c.begin(), c.begin() + lead); kmin += (int) lead; }
        if (c.empty()) kmin = 0;
    }
};

Laurent add(const Laurent& a, const Laurent& b, double sb = 1.0) {
    if (a.empty() && b.empty()) return Laurent();
    const int lo = a.empty() ? b.kmin : b.empty() ? a.kmin : std::min(a.kmin, b.kmin);
    const int hi = a.empty() ? b.kmax() : b.empty() ? a.kmax() : std::max(a.kmax(), b.kmax());
    Laurent r(lo, std::vector<double>(hi - lo + 1, 0.0));
    for (int k = lo; k <= hi; k++) r.c[k - lo] = a.coef(k) + sb * b.coef(k);
    r.trim();
    return r;
}
Laurent mul(const Laurent& a, const Laurent& b) {
    if (a.empty() || b.empty()) return Laurent();
    Laurent r(a.kmin + b.kmin, std::vector<double>(a.c.size() + b.c.size() - 1, 0.0));
    for (size_t i = 0; i < a.c.size(); i++) for (size_t j = 0; j < b.c.size(); j++) r.c[i + j] += a.c[i] * b.c[j];
    r.trim();
    return r;
}
Laurent scale(const Laurent& a, double s) { Laurent r = a; for (double& x : r.c) x *= s; r.trim(); return r; }
Laurent shift(const Laurent& a, int dk) { Laurent r = a; if (!r.empty()) r.kmin += dk; return r; }   // times r^dk
Laurent deriv(const Laurent& a) {
    if (a.empty()) return Laurent();
    Laurent r(a.kmin - 1, std::vector<double>(a.c.size(), 0.0));
    for (size_t i = 0; i < a.c.size(); i++) r.c[i] = (double) (a.kmin + (int) i) * a.c[i];
    r.trim();
    return r;
}
// Polynomial(k_min, k_max, coeff, k_step) of cpp/Factorial.cpp:447
Laurent stepped(int kmin, int kmax, const std::vector<double>& coeff, int step = 1) {
    Laurent r(kmin, std::vector<double>(kmax - kmin + 1, 0.0));
    size_t ci = 0;
    for (int k = kmin; k <= kmax; k += step) r.c[k - kmin] = coeff[ci++];
    return r;    // not trimmed: callers keep zero coefficients harmlessly
}

// spur[n]: constant the reference's n-th derivative of this term carries on top of the true one (see HFD-B2 below)
struct Term { Laurent P, Q; std::vector<double> spur; };
// no_fh: the reference's potential_dn of this piece lacks the length-unit factor, so its Feynman-Hibbs terms vanish
struct Piece { double r_hi; std::vector<Term> terms; bool no_fh = false; };
struct Potential { double len_scale, e_scale; int zero_on_negative = 0; std::vector<Piece> pieces; };

const double KB = 1.38064852e-23;
double factorial(int k) { double f = 1; for (int i = 2; i <= k; i++) f *= i; return f; }

bool build_base(const kg_potential_desc_ex& d, Potential& pot, std::string& err) {
    const double inf = std::numeric_limits<double>::infinity();
    const double* q = d.params;
    auto need = [&](int n) { if (d.nparams != n) { err = "kg_model_create_ex: potential kind " + std::to_string(d.kind) + " expects " + std::to_string(n) + " parameters"; return false; } return true; };
    switch (d.kind) {
    case KG_POT_LJ_SPLINE: {
        // internal unit sigma, energy eps. a, b, rs, rc as in the constructor (cpp/LJSpline.h:20-23)
        if (!need(0)) return false;
        pot.len_scale = 1.0; pot.e_scale = d.eps; pot.zero_on_negative = 1;
        const double a = -24192. / 3211., b = -387072. / 61009.;
        const double rs = pow(26. / 7., 1. / 6.), rc = 67. * rs / 48.;
        Piece lj{rs, {Term{Laurent(-12, {4., 0, 0, 0, 0, 0, -4.}), Laurent()}}};
        // a y^2 + b y^3, y = (r - rc)/rs
        const Laurent y(0, {-rc / rs, 1.0 / rs});
        const Laurent y2 = mul(y, y);
        Piece sp{rc, {Term{add(scale(y2, a), scale(mul(y2, y), b)), Laurent()}}};
        pot.pieces = {lj, sp};
        return true;
    }
    case KG_POT_LJTS: {
        // r <= rc = 2.5 sigma: LJ - LJ(rc) (cpp/LJTS.h:15,27,32-38)
        if (!need(0)) return false;
        pot.len_scale = 1.0; pot.e_scale = d.eps;
        const double rc = 2.5;
        const double uc = 4 * (pow(1.0 / rc, 12) - pow(1.0 / rc, 6));
        std::vector<double> c(13, 0.0); c[0] = 4.; c[6] = -4.; c[12] = -uc;
        pot.pieces = {Piece{std::nextafter(rc, inf), {Term{Laurent(-12, c), Laurent()}}}};
        return true;
    }
    case KG_POT_TANG_TOENNIES: {
        // params: A, b, A_tilde, a_tilde, a1, a2, am1, am2, L_unit, short_range_lim, C6..C16 (energies in K)
        if (!need(16)) return false;
        const double A = q[0], b = q[1], At = q[2], at = q[3], a1 = q[4], a2 = q[5], am1 = q[6], am2 = q[7], L = q[8], lim = q[9];
        const double* C = q + 10;
        pot.len_scale = d.sigma / L; pot.e_scale = KB;
        Piece shortr{lim, {Term{Laurent(-1, {At}), Laurent(1, {-at})}}};
        Piece main{inf, {}};
        main.terms.push_back(Term{Laurent(0, {A}), Laurent(-2, {am2, am1, 0., a1, a2})});
        main.terms.push_back(Term{stepped(-16, -6, {-C[5], -C[4], -C[3], -C[2], -C[1], -C[0]}, 2), Laurent()});
        std::vector<double> coeff(17, 0.);     // coeff[m] multiplies r^-m exp(-b r)  (cpp/multiparam.cpp:20-26)
        for (int n = 3; n <= 8; n++) for (int k = 0; k <= 2 * n; k++) coeff[2 * n - k] += C[n - 3] * pow(b, k) / factorial(k);
        std::vector<double> cr(coeff.rbegin(), coeff.rend());
        main.terms.push_back(Term{Laurent(-16, cr), Laurent(1, {-b})});
        pot.pieces = {shortr, main};
        return true;
    }
    case KG_POT_HFD_B2: {
        // params: A, alpha, c6, c8, c10, beta_star, D, rm [nm]; internal unit x = r / rm, energy eps
        if (!need(8)) return false;
        const double A = q[0], alpha = q[1], c6 = q[2], c8 = q[3], c10 = q[4], bs = q[5], D = q[6], rm = q[7];
        pot.len_scale = d.sigma * 1e9 / rm; pot.e_scale = d.eps;
        // Reference quirk kept for parity: PolyExp::get_Gk (cpp/Factorial.cpp:560-579) leaves the loop over a partition with
        // `break` when a part exceeds the degree of a non-negative-power exponent polynomial, so such partitions
        // contribute their multiplicity instead of 0. For exp(-alpha x + beta* x^2) the n-th derivative (n >= 3) gains
        // A S_n e^Q with S_n = sum of multiplicities of the partitions of n having a part > 2: S_3 = 1, S_4 = 5, S_5 = 26.
        const Term rep{Laurent(0, {A}), Laurent(1, {-alpha, bs}), {0., 0., 0., A * 1., A * 5., A * 26.}};
        const Laurent disp = stepped(-10, -6, {-c10, -c8, -c6}, 2);
        Piece damped{D, {rep, Term{disp, Laurent(-2, {-D * D, 2 * D, -1.})}}};
        Piece undamped{inf, {rep, Term{disp, Laurent()}}};
        pot.pieces = {damped, undamped};
        return true;
    }
    case KG_POT_PATKOWSKI: {
        // params: Rc, Ac, Bc, Cex1, Cex2, Csp1..4, delta, C6, C8, C10; internal unit Angstrom, energies in K
        if (!need(13)) return false;
        const double Rc = q[0], Ac = q[1], Bc = q[2], Cex1 = q[3], Cex2 = q[4], delta = q[9];
        const double Cn[3] = {q[10], q[11], q[12]};
        pot.len_scale = d.sigma * 1e10; pot.e_scale = KB;
        // Patowski::potential_dn below Rc returns Ac Bc^n exp(..) kB without the 1e10^n of the main branch
        // (cpp/multiparam.cpp:268-270): in SI units those derivatives are ~1e-20 of the true ones, i.e. the reference applies
        // no Feynman-Hibbs correction inside the core extrapolation. Kept.
        Piece core{Rc, {Term{Laurent(0, {Ac}), Laurent(0, {-Bc * Rc, Bc})}}, true};
        Piece main{inf, {}};
        main.terms.push_back(Term{Laurent(0, {q[5], q[6], q[7], q[8]}), Laurent(0, {Cex1, Cex2})});
        main.terms.push_back(Term{stepped(-10, -6, {Cn[2], Cn[1], Cn[0]}, 2), Laurent()});
        Laurent damp;                       // the three damping sums share exp(-delta r): one term
        for (int n = 3; n <= 5; n++) {
            std::vector<double> coeff;
            for (int k = 0; k <= 2 * n; k++) coeff.push_back(-Cn[n - 3] * pow(delta, k) / factorial(k));
            damp = add(damp, Laurent(-2 * n, coeff));
        }
        main.terms.push_back(Term{damp, Laurent(1, {-delta})});
        pot.pieces = {core, main};
        return true;
    }
    default:
        err = "kg_model_create_ex: unknown potential kind " + std::to_string(d.kind);
        return false;
    }
}

// V_{k+1} = V_k' + V_k Q'  (prefactor of the (k+1)-th derivative of P exp(Q))
Laurent next_deriv(const Laurent& V, const Laurent& dQ) { return add(deriv(V), mul(V, dQ)); }

bool store(const Laurent& a, int* kmin, int* n, double* c, int cap, std::string& err) {
    if ((int) a.c.size() > cap) { err = "generic potential: prefactor needs " + std::to_string(a.c.size()) + " coefficients"; return false; }
    *kmin = a.kmin; *n = (int) a.c.size();
    for (int i = 0; i < cap; i++) c[i] = i < *n ? a.c[i] : 0.0;
    return true;
}

} // namespace

// Effective potential at temperature T: base potential for fh_order == 0, else the Feynman-Hibbs corrected one.
// Db = D / (kT) [m^2] enters in internal length units.
int kg_genpot_build(const kg_potential_desc_ex& d, double T, kg::GenPot& out, std::string& err) {
    Potential pot;
    if (!build_base(d, pot, err)) return -1;
    if (d.fh_order < 0 || d.fh_order > 2) { err = "kg_model_create_ex: fh_order must be 0, 1 or 2"; return -1; }
    if (pot.pieces.size() > (size_t) kg::GP_MAXP) { err = "generic potential: too many pieces"; return -1; }
    memset(&out, 0, sizeof(out));
    out.len_scale = pot.len_scale; out.e_scale = pot.e_scale; out.zero_on_negative = pot.zero_on_negative;
    out.npieces = (int) pot.pieces.size();
    const double L = d.sigma / pot.len_scale;                       // internal length unit in m
    const double Db = d.fh_order > 0 ? d.fh_D / (KB * T) / (L * L) : 0.0;
    for (int p = 0; p < out.npieces; p++) {
        const Piece& pc = pot.pieces[p];
        if (pc.terms.size() > (size_t) kg::GP_MAXT) { err = "generic potential: too many terms"; return -1; }
        out.piece[p].r_hi = pc.r_hi; out.piece[p].nterms = (int) pc.terms.size();
        for (size_t t = 0; t < pc.terms.size(); t++) {
            Laurent Q = pc.terms[t].Q; Q.trim();
            const Laurent dQ = deriv(Q);
            const int F = pc.no_fh ? 0 : d.fh_order;
            std::vector<Laurent> V(2 * F + 2);                      // V[n] e^Q = n-th derivative of the term
            V[0] = pc.terms[t].P; V[0].trim();
            for (size_t k = 1; k < V.size(); k++) V[k] = next_deriv(V[k - 1], dQ);
            const Laurent Vtrue1 = V[1];
            for (size_t k = 1; k < V.size() && k < pc.terms[t].spur.size(); k++) V[k] = add(V[k], Laurent(0, {pc.terms[t].spur[k]}));
            Laurent Pe = V[0], De = Vtrue1;                         // order 0 and 1 of the base class are exact (autodiff / closed forms)
            double Dn = 1;
            for (int n = 1; n <= F; n++) {
                Dn *= Db / n;
                Pe = add(Pe, add(V[2 * n], shift(V[2 * n - 1], -1), 2.0 * n), Dn);                            // cpp/extentions.h:231-236
                const Laurent inner = add(shift(V[2 * n], -1), shift(V[2 * n - 1], -2), -1.0);
                De = add(De, add(V[2 * n + 1], inner, 2.0 * n), Dn);                                          // cpp/extentions.h:240-255
            }
            kg::GenTerm& g = out.piece[p].t[t];
            if (!store(Pe, &g.pmin, &g.pn, g.pc, kg::GP_MAXC, err) || !store(De, &g.dmin, &g.dn, g.dc, kg::GP_MAXC, err) ||
                !store(Q, &g.qmin, &g.qn, g.qc, kg::GP_MAXQ, err)) return -1;
        }
    }
    return 0;
}

// Host-only inspection of the effective potential the device will evaluate (tests pin the Laurent algebra against the
// reference's potential / potential_derivative_r without a GPU). out receives the GenPot as raw doubles.
extern "C" int kg_genpot_describe(const kg_potential_desc_ex* d, double T, double* out, int cap_doubles) {
    if (!d || !out || cap_doubles < kg::GP_DOUBLES) return -kg::GP_DOUBLES;
    kg::GenPot gp; std::string err;
    if (kg_genpot_build(*d, T, gp, err)) return -1;
    memcpy(out, &gp, sizeof(gp));
    return kg::GP_DOUBLES;
}
