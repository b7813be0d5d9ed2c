// kg_saft.cuh -- SAFT-VR Mie residual Helmholtz energy (Lafitte et al., J. Chem. Phys. 139, 154504 (2013)) as one generic function
// F(T; V, n) = A_res / (R T), written once for any scalar type: evaluated in hyper-dual numbers it gives the exact second derivatives
// the RET property code asks its EoS for:  d mu_i / d n_j  (cpp/eos_interface.h:121-135, chemical_potential_tv),
// p(T, V, n) and dV/dn_i at constant T, p (pressure_tv, specific_volume; CoV frame, cpp/KineticGas_properties.cpp:459-500).
//
// Reference replaced: thermopack's saftvrmie model, which pykingas.MieKinGas constructs as its EoS (pykingas/MieKinGas.py:55-65) -- a
// third-party dependency of the reference (vendored: cpp/external/thermopack, release 2.2.3; src/saftvrmie_{hardsphere,dispersion,chain}.f90)
// with the model options of its default "LAFITTE" set (src/saftvrmie_options.f90:129-144): BMCSL hard-sphere mixture of the Barker-Henderson
// diameters d_ii, d_ij = (d_ii + d_jj)/2, a1 / a2 / a3 with the mixture packing fractions zeta_x, zeta_x*, K_HS of Carnahan-Starling, chain term
// from g_ii^Mie(sigma_ii) to second order, no quantum corrections, no association. This file restates the PUBLISHED model; nothing is
// translated from the Fortran (which carries hand-written derivatives; here derivatives come from the number type).
// Parity is pinned term by term against the thermopack wheel (hard-sphere, dispersion and total Helmholtz energy, their n-Hessians, mu, p): tests/test_eos_host.py, tests/test_gpu_eos.py.
//
// The header compiles for the device (kg_eos.cu) and, with KG_HD empty, for the host (tests/host/eos_host.cpp: the same code checked
// on CPU against the fixtures in the `-m "not gpu"` suite).
#pragma once
#include <math.h>

#ifndef KG_HD
#ifdef __CUDACC__
#define KG_HD __host__ __device__ __forceinline__
#else
#define KG_HD inline
#endif
#endif

namespace kgsaft {

constexpr int MAXC = 6;                       // = KG_MAX_COMPS
constexpr double KB = 1.380649e-23;           // thermopack_constants.f90:7-10 (CODATA 2018)
constexpr double NA = 6.02214076e23;
constexpr double RGAS = NA * KB;
constexpr double PI = 3.14159265358979323846;

// ---- hyper-dual number: value, two first-order parts, one mixed second-order part. f(x + a e1 + b e2 + ab e1e2) carries
// f, f' a, f' b, f' ab + f'' a b: seeding variables p and q with e1 and e2 returns d2f/dp dq exactly (no step size).
struct HD {
    double v, a, b, ab;
    KG_HD HD() : v(0), a(0), b(0), ab(0) {}
    KG_HD HD(double v_) : v(v_), a(0), b(0), ab(0) {}
    KG_HD HD(double v_, double a_, double b_, double ab_) : v(v_), a(a_), b(b_), ab(ab_) {}
};
KG_HD HD chain(const HD& x, double f, double f1, double f2) { return HD(f, f1 * x.a, f1 * x.b, f1 * x.ab + f2 * x.a * x.b); }
KG_HD HD operator+(const HD& x, const HD& y) { return HD(x.v + y.v, x.a + y.a, x.b + y.b, x.ab + y.ab); }
KG_HD HD operator-(const HD& x, const HD& y) { return HD(x.v - y.v, x.a - y.a, x.b - y.b, x.ab - y.ab); }
KG_HD HD operator-(const HD& x) { return HD(-x.v, -x.a, -x.b, -x.ab); }
KG_HD HD operator*(const HD& x, const HD& y) {
    return HD(x.v * y.v, x.a * y.v + x.v * y.a, x.b * y.v + x.v * y.b, x.ab * y.v + x.a * y.b + x.b * y.a + x.v * y.ab);
}
KG_HD HD inv(const HD& x) { const double r = 1.0 / x.v; return chain(x, r, -r * r, 2 * r * r * r); }
KG_HD HD operator/(const HD& x, const HD& y) { return x * inv(y); }
KG_HD HD operator+(const HD& x, double c) { return HD(x.v + c, x.a, x.b, x.ab); }
KG_HD HD operator+(double c, const HD& x) { return x + c; }
KG_HD HD operator-(const HD& x, double c) { return HD(x.v - c, x.a, x.b, x.ab); }
KG_HD HD operator-(double c, const HD& x) { return HD(c - x.v, -x.a, -x.b, -x.ab); }
KG_HD HD operator*(const HD& x, double c) { return HD(x.v * c, x.a * c, x.b * c, x.ab * c); }
KG_HD HD operator*(double c, const HD& x) { return x * c; }
KG_HD HD operator/(const HD& x, double c) { return x * (1.0 / c); }
KG_HD HD operator/(double c, const HD& x) { return inv(x) * c; }
KG_HD HD& operator+=(HD& x, const HD& y) { x = x + y; return x; }
KG_HD HD exp(const HD& x) { const double e = ::exp(x.v); return chain(x, e, e, e); }
KG_HD HD log(const HD& x) { const double r = 1.0 / x.v; return chain(x, ::log(x.v), r, -r * r); }
KG_HD HD tanh(const HD& x) { const double t = ::tanh(x.v), s = 1 - t * t; return chain(x, t, s, -2 * t * s); }
KG_HD double val(const HD& x) { return x.v; }
KG_HD double val(double x) { return x; }
// plain doubles use the global math functions through these names as well
KG_HD double inv(double x) { return 1.0 / x; }
using ::exp; using ::log; using ::tanh;

// ---- model ------------------------------------------------------------------------------------------------------------
struct PairC {                  // temperature-independent constants of one pair (i, j)
    double sigma, epsk, la, lr; // m, K (epsilon / k_B), exponents
    double C, alpha;            // Mie prefactor, van der Waals alpha = C (1/(la-3) - 1/(lr-3))
    double f[6];                // f_1 .. f_6 (alpha) of the chi and a3 correlations
    double cz[5][4];            // zeta_eff coefficients c_1..c_4 for the exponents la, lr, 2 la, la + lr, 2 lr
    double lam[5];              // those exponents
};
struct Model {
    int nc;
    double m[MAXC];             // segments per molecule
    PairC p[MAXC][MAXC];
};

// published constants (Lafitte 2013, Table II and eq. for zeta_eff); thermopack holds the same numbers in
// src/saftvrmie_containers.f90:670-677 and src/saftvrmie_dispersion.f90:383-385
KG_HD double phi_tab(int n, int i) {            // n = 0..6 coefficient index, i = 0..5 function index (f_1..f_6)
    const double t[6][7] = {{7.5365557, -37.60463, 71.745953, -46.83552, -2.467982, -0.50272, 8.0956883},
                            {-359.44, 1825.6, -3168.0, 1884.2, -0.82376, -3.1935, 3.709},
                            {1550.9, -5070.1, 6534.6, -3288.7, -2.7171, 2.0883, 0.0},
                            {-1.19932, 9.063632, -17.9482, 11.34027, 20.52142, -56.6377, 40.53683},
                            {-1911.28, 21390.175, -51320.7, 37064.54, 1103.742, -3264.61, 2556.181},
                            {9236.9, -129430.0, 357230.0, -315530.0, 1390.2, -4518.2, 4241.6}};
    return t[i][n];
}
KG_HD void init_pair(PairC& q) {
    const double la = q.la, lr = q.lr;
    q.C = (lr / (lr - la)) * pow(lr / la, la / (lr - la));
    q.alpha = q.C * (1.0 / (la - 3.0) - 1.0 / (lr - 3.0));
    const double a = q.alpha, a2 = a * a, a3 = a2 * a;
    for (int i = 0; i < 6; i++)
        q.f[i] = (phi_tab(0, i) + phi_tab(1, i) * a + phi_tab(2, i) * a2 + phi_tab(3, i) * a3) /
                 (1.0 + phi_tab(4, i) * a + phi_tab(5, i) * a2 + phi_tab(6, i) * a3);
    const double lc[4][4] = {{0.81096, 1.7888, -37.578, 92.284}, {1.0205, -19.341, 151.26, -463.50},
                             {-1.9057, 22.845, -228.14, 973.92}, {1.0885, -6.1962, 106.98, -677.64}};
    q.lam[0] = la; q.lam[1] = lr; q.lam[2] = 2 * la; q.lam[3] = la + lr; q.lam[4] = 2 * lr;
    for (int k = 0; k < 5; k++) {
        const double il = 1.0 / q.lam[k];
        for (int i = 0; i < 4; i++) q.cz[k][i] = lc[i][0] + il * (lc[i][1] + il * (lc[i][2] + il * lc[i][3]));
    }
}

// Barker-Henderson diameter d = int_0^sigma (1 - exp(-u(r)/kT)) dr of a Mie potential. Below r0, where exp(-u/kT) falls under the
// machine precision, the integrand is 1 (thermopack does the same split, src/saftvrmie_hardsphere.f90:655-666); [r0, sigma] by
// 48-point Gauss-Legendre in three panels (converged to 1e-14 relative: the integrand is analytic there).
KG_HD double bh_diameter(const PairC& q, double T) {
    const double be = q.C * q.epsk / T;
    auto bu = [&](double x) { return be * (pow(x, -q.lr) - pow(x, -q.la)); };      // u / kT at r = x sigma
    double x0 = 0.7;                                                              // bu(x0) = -ln(eps_mach)
    const double target = 36.04365338911715;                                      // -ln(2^-52)
    for (int it = 0; it < 60; it++) {
        const double fv = bu(x0) - target;
        const double dv = be * (-q.lr * pow(x0, -q.lr - 1) + q.la * pow(x0, -q.la - 1));
        const double xn = x0 - fv / dv;
        const double xc = xn <= 0.0 ? 0.5 * x0 : (xn >= 1.0 ? 0.5 * (x0 + 1.0) : xn);
        if (fabs(xc - x0) < 1e-15 * x0) { x0 = xc; break; }
        x0 = xc;
    }
    // 16-point Gauss-Legendre nodes / weights on [-1, 1] (positive half)
    const double gx[8] = {0.0950125098376374401853193, 0.2816035507792589132304605, 0.4580167776572273863424194, 0.6178762444026437484466718,
                          0.7554044083550030338951012, 0.8656312023878317438804679, 0.9445750230732325760779884, 0.9894009349916499325961542};
    const double gw[8] = {0.1894506104550684962853967, 0.1826034150449235888667637, 0.1691565193950025381893121, 0.1495959888165767320815017,
                          0.1246289712555338720524763, 0.0951585116824927848099251, 0.0622535239386478928628438, 0.0271524594117540948517806};
    double sum = 0.0;
    const int NP = 3;
    for (int pnl = 0; pnl < NP; pnl++) {
        const double lo = x0 + (1.0 - x0) * pnl / NP, hi = x0 + (1.0 - x0) * (pnl + 1) / NP, h = 0.5 * (hi - lo), c = 0.5 * (hi + lo);
        for (int k = 0; k < 8; k++) {
            sum += gw[k] * h * ((1.0 - ::exp(-bu(c - h * gx[k]))) + (1.0 - ::exp(-bu(c + h * gx[k]))));
        }
    }
    return q.sigma * (x0 + sum);
}

// temperature-dependent quantities of a state point (plain doubles: nothing here is differentiated)
struct TPre {
    double T;
    double d[MAXC];                 // d_ii
    double d3[MAXC][MAXC];          // d_ij^3, d_ij = (d_ii + d_jj) / 2
    double x0p[MAXC][MAXC][5];      // (sigma_ij / d_ij)^lam_k
    double I[MAXC][MAXC][5], J[MAXC][MAXC][5];
    double x0[MAXC][MAXC];
};
KG_HD void init_T(const Model& M, double T, TPre& P) {
    P.T = T;
    for (int i = 0; i < M.nc; i++) P.d[i] = bh_diameter(M.p[i][i], T);
    for (int i = 0; i < M.nc; i++) for (int j = 0; j < M.nc; j++) {
        const PairC& q = M.p[i][j];
        const double dij = 0.5 * (P.d[i] + P.d[j]);
        P.d3[i][j] = dij * dij * dij;
        const double x0 = q.sigma / dij;
        P.x0[i][j] = x0;
        for (int k = 0; k < 5; k++) {
            const double l = q.lam[k];
            P.x0p[i][j][k] = pow(x0, l);
            P.I[i][j][k] = -(pow(x0, 3.0 - l) - 1.0) / (l - 3.0);
            P.J[i][j][k] = -(pow(x0, 4.0 - l) * (l - 3.0) - pow(x0, 3.0 - l) * (l - 4.0) - 1.0) / ((l - 3.0) * (l - 4.0));
        }
    }
}

// ---- the Helmholtz function ---------------------------------------------------------------------------------------------
// (a1s + B)(lam_k) of pair (i, j) [K m^3-free: energy in K] at segment density rho_s and packing fraction zeta_x, and rho_s times
// its derivative with respect to rho_s at constant composition (zeta_x is proportional to rho_s):
//   aB  = 2 pi rho_s d^3 eps [ -(1 - ze/2) / ((lam-3)(1-ze)^3) + (1 - zx/2)/(1-zx)^3 I - 9 zx (1+zx) / (2 (1-zx)^3) J ]      (Lafitte eqs. A16, A12)
//   rho_s d(aB)/d(rho_s) = aB + 2 pi rho_s d^3 eps zx d[...]/d(zx)
template<class S>
KG_HD void aB_pair(const PairC& q, const TPre& P, int i, int j, int k, const S& rho_s, const S& zx, S& aB, S* rho_daB) {
    const double* c = q.cz[k];
    const double lam = q.lam[k], I = P.I[i][j][k], J = P.J[i][j][k];
    const S ze = zx * (c[0] + zx * (c[1] + zx * (c[2] + zx * c[3])));
    const S ome = 1.0 - ze, omx = 1.0 - zx;
    const S ie3 = inv(ome * ome * ome), ix3 = inv(omx * omx * omx);
    const S t1 = (1.0 - 0.5 * ze) * ie3 * (-1.0 / (lam - 3.0));
    const S t2 = (1.0 - 0.5 * zx) * ix3 * I;
    const S t3 = zx * (1.0 + zx) * ix3 * (-4.5 * J);
    const S pre = rho_s * (2.0 * PI * P.d3[i][j] * q.epsk);
    aB = pre * (t1 + t2 + t3);
    if (rho_daB) {
        const S dze = c[0] + zx * (2.0 * c[1] + zx * (3.0 * c[2] + zx * 4.0 * c[3]));
        // d/dze [(1 - ze/2)/(1-ze)^3] = (5/2 - ze) / (1-ze)^4 ; the same form in zx ; d/dzx [zx(1+zx)/(1-zx)^3] = (1 + 4 zx + zx^2)/(1-zx)^4
        const S d1 = (2.5 - ze) * ie3 * inv(ome) * dze * (-1.0 / (lam - 3.0));
        const S d2 = (2.5 - zx) * ix3 * inv(omx) * I;
        const S d3 = (1.0 + zx * (4.0 + zx)) * ix3 * inv(omx) * (-4.5 * J);
        *rho_daB = aB + pre * zx * (d1 + d2 + d3);
    }
}

// F = A_res / (R T) [mol] of n[] mol in volume V [m^3]; the switches select terms (term-by-term pinning against thermopack).
enum { TERM_HS = 1, TERM_A1 = 2, TERM_A2 = 4, TERM_A3 = 8, TERM_CHAIN = 16, TERM_ALL = 31 };
template<class S>
KG_HD S helmholtz(const Model& M, const TPre& P, const S& V, const S* n, int terms = TERM_ALL) {
    const int nc = M.nc;
    const double T = P.T;
    S ns = S(0.0);
    for (int i = 0; i < nc; i++) ns += n[i] * M.m[i];
    const S ins = inv(ns);
    S xs[MAXC];
    for (int i = 0; i < nc; i++) xs[i] = n[i] * M.m[i] * ins;
    const S rho_s = ns * inv(V) * NA;
    // packing fractions
    S z0 = S(0.0), z1 = S(0.0), z2 = S(0.0), z3 = S(0.0), sx = S(0.0), sxs = S(0.0);
    for (int i = 0; i < nc; i++) {
        const double d = P.d[i];
        z0 += xs[i]; z1 += xs[i] * d; z2 += xs[i] * (d * d); z3 += xs[i] * (d * d * d);
        for (int j = 0; j < nc; j++) {
            const S xx = xs[i] * xs[j];
            sx += xx * P.d3[i][j];
            const double sg = M.p[i][j].sigma;
            sxs += xx * (sg * sg * sg);
        }
    }
    const S pr = rho_s * (PI / 6.0);
    z0 = z0 * pr; z1 = z1 * pr; z2 = z2 * pr; z3 = z3 * pr;
    const S zx = sx * pr, zxs = sxs * pr;
    S a_seg = S(0.0);                                   // per segment, in units of kT
    if (terms & TERM_HS) {                              // Boublik-Mansoori-Carnahan-Starling-Leland (Lafitte eq. 10)
        const S om = 1.0 - z3, z23 = z2 * z2 * z2;
        a_seg += (6.0 / PI) * inv(rho_s) * ((z23 * inv(z3 * z3) - z0) * log(om) + 3.0 * z1 * z2 * inv(om) + z23 * inv(z3 * om * om));
    }
    const S omx = 1.0 - zx;
    const S omx2 = omx * omx;
    const S KHS = omx2 * omx2 * inv(1.0 + zx * (4.0 + zx * (4.0 + zx * (-4.0 + zx))));       // Lafitte eq. A21
    const S zxs2 = zxs * zxs, zxs5 = zxs2 * zxs2 * zxs, zxs8 = zxs5 * zxs2 * zxs;
    if (terms & (TERM_A1 | TERM_A2 | TERM_A3)) {
        S a1 = S(0.0), a2 = S(0.0), a3 = S(0.0);
        for (int i = 0; i < nc; i++) for (int j = i; j < nc; j++) {
            const PairC& q = M.p[i][j];
            const S w = xs[i] * xs[j] * (i == j ? 1.0 : 2.0);
            S aB[5];
            for (int k = 0; k < 5; k++) aB_pair<S>(q, P, i, j, k, rho_s, zx, aB[k], (S*) nullptr);
            const double* xp = P.x0p[i][j];
            if (terms & TERM_A1) a1 += w * (q.C * (xp[0] * aB[0] - xp[1] * aB[1]));
            if (terms & TERM_A2) {
                const S chi = q.f[0] * zxs + q.f[1] * zxs5 + q.f[2] * zxs8;
                a2 += w * (0.5 * q.epsk * q.C * q.C) * KHS * (1.0 + chi) * (xp[2] * aB[2] - 2.0 * xp[3] * aB[3] + xp[4] * aB[4]);
            }
            if (terms & TERM_A3) a3 += w * (-(q.epsk * q.epsk * q.epsk) * q.f[3]) * zxs * exp(q.f[4] * zxs + q.f[5] * zxs2);
        }
        const double b = 1.0 / T;
        a_seg += a1 * b + a2 * (b * b) + a3 * (b * b * b);
    }
    S F = ns * a_seg;
    if (terms & TERM_CHAIN) {
        // -sum_i n_i (m_i - 1) ln g_ii^Mie(sigma_ii), g^Mie = g_dHS exp(beta eps g1 / g_dHS + (beta eps)^2 g2 / g_dHS)   (Lafitte eqs. 46-65)
        const S ix3 = inv(omx2 * omx);
        const S k0 = -log(omx) + zx * (42.0 + zx * (-39.0 + zx * (9.0 - 2.0 * zx))) * ix3 * (1.0 / 6.0);
        const S k1 = zx * (-12.0 + zx * (6.0 + zx * zx)) * ix3 * 0.5;
        const S k2 = zx * zx * inv(omx2) * (-0.375);
        const S k3 = zx * (3.0 + zx * (3.0 - zx * zx)) * ix3 * (1.0 / 6.0);
        // dK_HS/dzx times zx (rho_s d/d rho_s of K_HS)
        const S den = 1.0 + zx * (4.0 + zx * (4.0 + zx * (-4.0 + zx)));
        const S dK = (-4.0 * omx2 * omx * den - omx2 * omx2 * (4.0 + zx * (8.0 + zx * (-12.0 + 4.0 * zx)))) * inv(den * den);
        for (int i = 0; i < nc; i++) {
            if (M.m[i] == 1.0) continue;
            const PairC& q = M.p[i][i];
            const double x0 = P.x0[i][i];
            const S gd = exp(k0 + x0 * (k1 + x0 * (k2 + x0 * k3)));
            S aB[5], rd[5];
            for (int k = 0; k < 5; k++) aB_pair<S>(q, P, i, i, k, rho_s, zx, aB[k], &rd[k]);
            const double* xp = P.x0p[i][i];
            const S irho = inv(rho_s);
            const double pre1 = 1.0 / (2.0 * PI * q.epsk * P.d3[i][i]);
            // g1 = [3 d a1/d rho_s - C la x0^la aB(la)/rho_s + C lr x0^lr aB(lr)/rho_s] / (2 pi eps d^3)
            const S g1 = pre1 * irho * (3.0 * q.C * (xp[0] * rd[0] - xp[1] * rd[1]) - q.C * q.la * xp[0] * aB[0] + q.C * q.lr * xp[1] * aB[1]);
            // g2_MCA = [3 d(a2/(1+chi))/d rho_s - eps K C^2 (lr x0^2lr aB(2lr) - (la+lr) x0^(la+lr) aB(la+lr) + la x0^2la aB(2la)) / rho_s] / (2 pi eps^2 d^3)
            const S Q = xp[2] * aB[2] - 2.0 * xp[3] * aB[3] + xp[4] * aB[4];
            const S rdQ = xp[2] * rd[2] - 2.0 * xp[3] * rd[3] + xp[4] * rd[4];
            const double ec2 = q.epsk * q.C * q.C;
            const S rd_a2 = (0.5 * ec2) * (dK * zx * Q + KHS * rdQ);
            const S g2m = (pre1 / q.epsk) * irho * (3.0 * rd_a2 - ec2 * KHS * (q.lr * xp[4] * aB[4] - (q.la + q.lr) * xp[3] * aB[3] + q.la * xp[2] * aB[2]));
            const double be = q.epsk / T, th = ::exp(be) - 1.0;
            const S gam = (10.0 * (1.0 - ::tanh(10.0 * (0.57 - q.alpha))) * th) * zxs * exp(-6.7 * zxs - 8.0 * zxs2);      // phi_7 = (10, 10, 0.57, -6.7, -8)
            const S g2 = (1.0 + gam) * g2m;
            const S lng = log(gd) + (be * g1 + (be * be) * g2) * inv(gd);
            F = F - n[i] * (M.m[i] - 1.0) * lng;
        }
    }
    return F;
}

}   // namespace kgsaft
