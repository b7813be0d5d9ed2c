// kg_transport.cu -- Revised-Enskog state-point assembly on B200 (sm_100a), behind the C-ABI of include/kgcuda.h.
//
// Reference path replaced (thermotools/KineticGas @ 75848b1):
//   bracket weights A, A', B', B'' and sums H_ij, H_i, L_ij, L_i      cpp/KineticGas.cpp:520-705
//   MieKinGas::omega correlation branch (Fokin, Popov, Kalashnikov)   cpp/MieKinGas.cpp:81-148
//   SAFT-VR Mie RDF at contact, Barker-Henderson diameters            cpp/MieKinGas.cpp:157-276,314-786
//   K-factors, Lambda / B matrices and right-hand sides               cpp/KineticGas.cpp:207-457
//   partial-pivot LU solves and the property formulas                 cpp/KineticGas_properties.cpp:28-500
//
// Layout (not a translation). Everything that depends on temperature only is computed once per distinct T into
// device tables (kg_tcache): the Omega table [nT][npairs][nlr], transfer lengths, BH diameters and the bracket
// integrals [nT][nc][nc][N][N]. The combinatorial weights are T-independent: the host evaluates them once per
// (model, N) in extended precision and ships dense weight tables. A state point then costs one RDF evaluation per
// pair (rdf_kernel, one thread per (state point, density slot, pair)) and one thread of assembly + LU + property
// formulas (state_kernel). The (N nc^2)^2 diffusion system of the reference is block diagonal with nc copies of the
// conductivity matrix (cpp/KineticGas.cpp:334-388), so Lambda is factorised once and solved for nc + 1 right-hand sides.
#include "kg_internal.cuh"
#include <cmath>
#include <cfloat>
#include <cstring>
#include <map>
#include <algorithm>
#include <chrono>

using namespace kg;

namespace {

constexpr int MAXC = KG_MAX_COMPS;
constexpr int MAXO = KG_MAX_ORDER;
constexpr int MAXLR = 40;                 // (l,r) combinations at order <= 4: viscosity needs l <= N+1, r <= 2N+2-l
constexpr double GAS_CONSTANT = 8.31446261815324;     // cpp/global_params.h:5
constexpr double AVOGADRO = GAS_CONSTANT / 1.38064852e-23;

// ---------------------------------------------------------------------------------------------------------------
// host: combinatorial weights of the square-bracket integrals (Thompson, Tipton, Loyalka; cpp/KineticGas.cpp:520-629).
// Evaluated in long double from the closed forms; the reference cancels factorials symbolically before dividing,
// which is an overflow guard, not a different value.
// ---------------------------------------------------------------------------------------------------------------
typedef long double ld;
ld fac(int n) { ld r = 1; for (int k = 2; k <= n; k++) r *= k; return r; }
ld ipw(ld b, int e) { return powl(b, (ld) e); }
ld poch(int z, int n) { return fac(z + n - 1) / fac(z - 1); }
int kdelta(int a, int b) { return a == b; }

double weight_A(int p, int q, int r, int l) {
    ld v = 0;
    const int max_i = std::min(std::min(p, q), std::min(r, p + q + 1 - r));
    for (int i = l - 1; i <= max_i; i++) {
        const ld num = ipw(8, i) * fac(p + q - 2 * i) * ipw(-1, l + r + i) * fac(r + 1) * fac(2 * (p + q + 2 - i)) * ipw(4, r);
        const ld den = fac(p - i) * fac(q - i) * fac(l) * fac(i + 1 - l) * fac(r - i) * fac(p + q + 1 - i - r) * fac(2 * r + 2)
                     * fac(p + q + 2 - i) * ipw(4, p + q + 1);
        v += (num / den) * ((i + 1 - l) * (p + q + 1 - i - r) - l * (r - i));
    }
    return (double) v;
}

double weight_A_prime(int p, int q, int r, int l, double M1d, double M2d) {
    const ld M1 = M1d, M2 = M2d;
    const ld F = (M1 * M1 + M2 * M2) / (2 * M1 * M2), G = (M1 - M2) / M2;
    const int max_i = std::min(p, std::min(q, std::min(r, p + q + 1 - r)));
    ld v = 0;
    for (int i = l - 1; i <= max_i; i++) {
        const int max_w = std::min(p, std::min(q, p + q + 1 - r)) - i, max_k = std::min(l, i);
        for (int k = l - 1; k <= max_k; k++) for (int w = 0; w <= max_w; w++) {
            const ld p1 = ipw(8, i) * fac(p + q - 2 * i - w) * ipw(-1, r + i) * fac(r + 1) * fac(2 * (p + q + 2 - i - w)) * ipw(2, 2 * r)
                        * ipw(F, i - k) * ipw(G, w) * (ipw(2, 2 * w - 1) * ipw(M1, i) * ipw(M2, p + q - i - w) * 2)
                        * (M1 * (p + q + 1 - i - r - w) * kdelta(k, l) - M2 * (r - i) * kdelta(k, l - 1));
            const ld p2 = fac(p - i - w) * fac(q - i - w) * fac(r - i) * fac(p + q + 1 - i - r - w) * fac(2 * r + 2) * fac(p + q + 2 - i - w)
                        * ipw(4, p + q + 1) * fac(k) * fac(i - k) * fac(w);
            v += p1 / p2;
        }
    }
    return (double) v;
}

double weight_B_prime(int p, int q, int r, int l, double M1d, double M2d) {
    if (r == p + q + 3) return 0.0;
    const ld M1 = M1d, M2 = M2d;
    const ld G = (M1 - M2) / M2, F = (M1 * M1 + M2 * M2) / (2 * M1 * M2);
    const int i_max = std::min(p, std::min(q, std::min(r, p + q + 2 - r)));
    ld val = 0;
    for (int i = std::max(l - 2, 0); i <= i_max; i++) {        // the i = -1 term carries the factor (1 - delta(i,-1)) = 0
        ld inner = 0;
        const int w_max = std::min(p, std::min(q, p + q + 2 - r)) - i;
        for (int w = 0; w <= w_max; w++) {
            const ld num = poch(p + 1 - i - w, w) * poch(q + 1 - i - w, w) * poch(p + q + 3 - i - r - w, w) * ipw(2, 2 * w - 2) * ipw(G, w)
                         * poch(p + q + 4 - i - w, w) * ipw(M1 * M2, i) * ipw(M2, p + q - 2 * i - w) * ipw(F, i + 2 - l) * M1 * M1 * 4;
            const ld den = fac(w) * poch(p + q + 1 - 2 * i - w, w) * poch(2 * (p + q + 3 - i) - 2 * w + 1, w)
                         * poch(2 * (p + q + 3 - i) - w + 1, w) * fac(l) * fac(i + 2 - l);
            const ld m21 = M2 / M1;
            inner += (num / den) * ((ld) 1.5 * m21 * m21 * l * (l - 1) * (r - i) * (r - i - 1)
                                    - (2 / F) * m21 * l * (i + 2 - l) * (r - i) * (p + q + 2 - i - r - w)
                                    + (1 / (F * F)) * (i + 1 - l) * (i + 2 - l)
                                        * ((p + q + 1 - i - r - w) * (p + q + 2 - i - r - w) - (ld) 0.5 * m21 * m21 * (r - i) * (r - i - 1)));
        }
        const ld num = ipw(2, 2 * r) * ipw(8, i) * fac(p + q - 2 * i) * ipw(-1, r + i) * fac(r + 1) * fac(2 * (p + q + 3 - i));
        const ld den = ipw(4, p + q + 2) * fac(p - i) * fac(q - i) * fac(r - i) * fac(p + q + 2 - i - r) * fac(2 * r + 2) * fac(p + q + 3 - i);
        val += (num / den) * inner;
    }
    return (double) val;
}

double weight_B_dblprime(int p, int q, int r, int l, double M1d, double M2d) {
    if (r == p + q + 3) return 0.0;
    const ld M1 = M1d, M2 = M2d;
    const int i_max = std::min(p, std::min(q, std::min(r, p + q + 2 - r)));
    ld val = 0;
    for (int i = std::max(l - 2, 0); i <= i_max; i++) {
        const ld num = ipw(2, 2 * r) * ipw(8, i) * fac(p + q - 2 * i) * ipw(-1, r + i) * fac(r + 1) * fac(2 * (p + q + 3 - i)) * ipw(-1, l);
        const ld den = ipw(4, p + q + 2) * fac(p - i) * fac(q - i) * fac(r - i) * fac(p + q + 2 - i - r) * fac(2 * r + 2) * fac(p + q + 3 - i)
                     * fac(l) * fac(i + 2 - l);
        const ld pre = num / den;
        val += pre * (i + 1 - l) * (i + 2 - l) * ((p + q + 1 - i - r) * (p + q + 2 - i - r) - (ld) 0.5 * (r - i) * (r - i - 1));
        val += pre * (ld) 1.5 * (l - 1) * l * (r - i) * (r - i - 1);
        val += pre * (-2) * l * (i + 2 - l) * (r - i) * (p + q + 2 - i - r);
    }
    val *= ipw(M2, p + 1) * ipw(M1, q + 1);
    return (double) val;
}

} // namespace

// ---------------------------------------------------------------------------------------------------------------
// device constants: regressed SAFT-VR Mie / collision-integral correlation coefficients (published data, same values
// the reference carries in cpp/MieKinGas.h:199-276)
// ---------------------------------------------------------------------------------------------------------------
namespace {

// 20-point Gauss-Legendre rule on [-1,1] (BH diameter integral, cpp/MieKinGas.cpp:314-332)
__constant__ double c_gl_x[20] = {-0.9931285991850949, -0.9639719272779139, -0.912234428251326, -0.8391169718222187, -0.7463319064601508,
    -0.636053680726515, -0.5108670019508271, -0.37370608871541955, -0.22778585114164507, -0.07652652113349737, 0.07652652113349737,
    0.22778585114164507, 0.37370608871541955, 0.5108670019508271, 0.636053680726515, 0.7463319064601508, 0.8391169718222187,
    0.912234428251326, 0.9639719272779139, 0.9931285991850949};
__constant__ double c_gl_w[20] = {0.017614007139152742, 0.040601429800386134, 0.06267204833410799, 0.08327674157670514, 0.1019301198172403,
    0.11819453196151845, 0.13168863844917675, 0.14209610931838218, 0.149172986472604, 0.15275338713072611, 0.15275338713072611,
    0.149172986472604, 0.14209610931838218, 0.13168863844917675, 0.11819453196151845, 0.1019301198172403, 0.08327674157670514,
    0.06267204833410799, 0.040601429800386134, 0.017614007139152742};
// effective packing fraction coefficients c_k(lambda) = sum_j Cm[k][j] lambda^-j (cpp/MieKinGas.h:243-249; T. Maltby 2024, no LRC)
__constant__ double c_Cm[4][4] = {{0.6382, 0.9540, -31.9965, 83.7914}, {2.8596, -18.8315, 150.6853, -468.9992},
                                  {-7.0331, 21.9390, -229.9483, 970.6560}, {5.5441, -4.0284, 106.4364, -679.1889}};
// f_k(alpha) rational fits, Lafitte et al. J. Chem. Phys. 139, 154504 (2013) Table II (cpp/MieKinGas.h:265-275)
__constant__ double c_phi[7][3] = {{7.5365557, -359.44, 1550.9}, {-37.60463, 1825.6, -5070.1}, {71.745953, -3168.0, 6534.6},
                                   {-46.83552, 1884.2, -3288.7}, {-2.467982, -0.82376, -2.7171}, {-0.50272, -3.1935, 2.0883},
                                   {8.0956883, 3.7090, 0.0}};
// ln Omega* correlation, Fokin, Popov, Kalashnikov, High Temp. 37 (1999) (cpp/MieKinGas.h:199-217)
__constant__ double c_ocf[2][6][4] = {
    {{0., -0.145269e1, 0.294682e2, 0.242508e1}, {0.107782e-1, 0.587725, -0.180714e3, .595694e2}, {0.546646e-1, -0.651465e1, 0.374457e3, -0.137807e3},
     {0.485352, 0.245523e2, -0.336782e3, 0.814187e2}, {-0.385355, -0.206868e2, 0.132246e3, 0.}, {0.847232e-1, 0.521812e1, -0.181140e2, -0.747215e1}},
    {{0., 0.113086e1, 0.234799e2, 0.310127e1}, {0., 0.551559e1, -0.137023e3, 0.185848e2}, {0.325909e-1, -0.292925e2, 0.243761e3, 0.},
     {0.697682, 0.590792e2, -0.143670e3, -0.123518e3}, {-0.564238, -0.430549e2, 0., 0.137282e3}, {0.126508, 0.104273e2, 0.150601e2, -0.408911e2}}};

// Mie pair block as the assembly sees it (SI)
struct MiePair { double sigma, eps, la, lr, C, alpha; };

struct TModel {                 // per-model constants of the state-point stage
    int nc;                     // internal component count (2 for single-component models)
    int npairs_dev;             // pairs of the kg_model (1 for single-component models)
    int is_singlecomp, is_idealgas, N, nlr;
    double m[MAXC];             // particle masses
    double sqmm[MAXC][MAXC];    // sqrt(m_i m_j)                 (host libm, once per model: the state kernel used to take ~150 pow / sqrt per point)
    double sqred[MAXC][MAXC];   // sqrt(m_i m_j / (m_i + m_j))
    int pair_of[MAXC][MAXC];    // (i,j) -> pair index in the kg_model's tables
    int lr_l[MAXLR], lr_r[MAXLR];
};

__device__ __forceinline__ double ipowd(double x, int n) { double r = 1.0; for (int k = 0; k < n; k++) r *= x; return r; }

// ---- MieKinGas::omega correlation branch (cpp/MieKinGas.cpp:95-148) and the hard-sphere integral (cpp/MieKinGas.h:52-60) ----
__device__ double corr_am(int l, int n, double lr) {
    return c_ocf[l - 1][n - 1][0] + (c_ocf[l - 1][n - 1][1] / lr) + pow(lr, -2.) * (c_ocf[l - 1][n - 1][2] + c_ocf[l - 1][n - 1][3] * log(lr));
}
__device__ double corr_diag(int l, double lr, double Ts) {      // omega_correlation with l == r
    double lnw = -(2. / lr) * log(Ts);
    for (int n = 1; n < 7; n++) lnw += corr_am(l, n, lr) * pow(Ts, (1. - n) * 0.5);
    if (l == 2) lnw += log(1. - (2. / (3. * lr)));
    return exp(lnw);
}
// omega_correlation(l, r) = prod_{rr=l}^{r-1} omega_recursive_factor(l, rr) * omega_correlation(l, l), r <= 3.
// Integer quirk kept: the reference writes (1 / (r + 2)) with int operands, which is 0, so the a_m sums of the
// recursion factors vanish and only the leading terms survive (cpp/MieKinGas.cpp:129,141).
__device__ double omega_correlation(int l, int r, double lr, double Ts) {
    const double diag = corr_diag(l, lr, Ts);
    double v = diag, f = 0.0;
    for (int rr = l; rr < r; rr++) {
        if (rr == l) f = 1. + ((-(2 / (lr * Ts))) * diag) / (rr + 2);
        else f = 1 + (0.0 / ((rr + 2) * f)) + (f - 1) * (rr + 1) / (rr + 2);
        v *= f;
    }
    return v;
}

__device__ double ljs_omega_correlation(int l, int r, double Ts);      // LJ/spline table, defined with the LJ/s RDF below

// Omega of the correlation branch: omega_correlation * omega_hs. One thread per (t, pair, lr); mask says where it applies.
__global__ void omega_correlation_kernel(int nT, int npairs, int nlr, const double* __restrict__ T, const MiePair* __restrict__ mp,
                                         const int* __restrict__ pair_i, const int* __restrict__ pair_j, const double* __restrict__ mass,
                                         const int* __restrict__ lr_l, const int* __restrict__ lr_r, const unsigned char* __restrict__ is_corr,
                                         int ljs, double* __restrict__ omega) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nT * npairs * nlr) return;
    if (!is_corr[e]) return;
    const int k = e % nlr, p = (e / nlr) % npairs, t = e / (nlr * npairs);
    const int l = lr_l[k], r = lr_r[k], i = pair_i[p], j = pair_j[p];
    const MiePair q = mp[p];
    const double Ts = T[t] * BOLTZMANN / q.eps;
    // omega_hs, cpp/MieKinGas.h:52-60
    double w = PI_REF * pow(q.sigma, 2) * 0.5 * (r + 1);
    for (int ri = r; ri > 1; ri--) w *= ri;
    if (l % 2 == 0) w *= (1. - (1.0 / (l + 1.)));
    double hs;
    if (i == j) hs = sqrt((BOLTZMANN * T[t]) / (PI_REF * mass[i])) * w;
    else hs = sqrt(BOLTZMANN * T[t] * (mass[i] + mass[j]) / (2. * PI_REF * mass[i] * mass[j])) * w;
    omega[e] = (ljs ? ljs_omega_correlation(l, r, Ts) : omega_correlation(l, r, q.lr, Ts)) * hs;
}

// Barker-Henderson diameters of the pure components, d_ii(T) (cpp/MieKinGas.cpp:314-332). One thread per (t, i).
__global__ void bh_kernel(int nT, int nc, const double* __restrict__ T, const MiePair* __restrict__ diag, double* __restrict__ dbh) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nT * nc) return;
    const int i = e % nc, t = e / nc;
    const MiePair q = diag[i];
    const double beta = 1. / (BOLTZMANN * T[t]);
    double d = 2.;
    for (int n = 0; n < 20; n++) {
        const double r = q.sigma * (c_gl_x[n] / 4. + 3. / 4.);
        const double u = q.C * q.eps * (pow(q.sigma / r, q.lr) - pow(q.sigma / r, q.la));
        d += c_gl_w[n] * (1. - exp(-beta * u));
    }
    dbh[e] = d * (q.sigma / 4.);
}

// Bracket integrals from the Omega table and the host-built weights (cpp/KineticGas.cpp:644-705):
//   H_ij = 8 M2^(p+1/2) M1^(q+1/2) sum A Omega ; H_i = 8 sum A' Omega ; L_ij = 16/3 sum B'' Omega ; L_i = 16/3 sum B' Omega
// One thread per (t, i, j, p, q). Weights: wA [N][N][nlr]; wAp, wBp, wBpp [nc][nc][N][N][nlr]; hpre [nc][nc][N][N].
// Output brk [nT][4][nc][nc][N][N].
__global__ void bracket_kernel(int nT, TModel tm, const double* __restrict__ omega, const double* __restrict__ wA, const double* __restrict__ wAp,
                               const double* __restrict__ wBp, const double* __restrict__ wBpp, const double* __restrict__ hpre,
                               double* __restrict__ brk) {
    const int N = tm.N, nc = tm.nc, nlr = tm.nlr;
    const int per_t = nc * nc * N * N;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nT * per_t) return;
    const int t = e / per_t, rem = e % per_t;
    const int q = rem % N, p = (rem / N) % N, j = (rem / (N * N)) % nc, i = rem / (N * N * nc);
    const double* om = omega + ((size_t) t * tm.npairs_dev + tm.pair_of[i][j]) * nlr;
    const double* a = wA + (size_t) (p * N + q) * nlr;
    const size_t wo = ((size_t) ((i * nc + j) * N + p) * N + q) * nlr;
    double hij = 0, hi = 0, lij = 0, li = 0;
    for (int k = 0; k < nlr; k++) {
        const double o = om[k];
        if (a[k] != 0) hij += a[k] * o;
        if (wAp[wo + k] != 0) hi += wAp[wo + k] * o;
        if (wBpp[wo + k] != 0) lij += wBpp[wo + k] * o;
        if (wBp[wo + k] != 0) li += wBp[wo + k] * o;
    }
    double* out = brk + (size_t) t * 4 * per_t;
    out[rem] = hij * hpre[rem];
    out[per_t + rem] = hi * 8;
    out[2 * per_t + rem] = lij * (16.0 / 3.0);
    out[3 * per_t + rem] = li * (16.0 / 3.0);
}

// ---- SAFT-VR Mie RDF at contact, second order (cpp/MieKinGas.cpp:157-276, 345-786) ----
struct RdfCommon {      // what every pair of one (state point, density) shares
    double rho, zx, dzx, zxs, KHS, dKHS, k0, k1, k2, k3;
};

__device__ __forceinline__ double p2(double x) { return x * x; }
__device__ __forceinline__ double p3(double x) { return x * x * x; }
__device__ __forceinline__ double p4(double x) { const double y = x * x; return y * y; }

// Everything of MieKinGas::saft_rdf (cpp/MieKinGas.cpp:157-276, 314-786) that depends on the temperature and the pair only, computed once
// per (T, i, j) by rdf_T_kernel: the per-state-point kernel used to spend ~280 pow() calls per pair on these.
struct RdfT {
    double d3, sig3, x0, eps, la, lr, C;
    double lam[5];             // la, lr, 2 la, 2 lr, la + lr
    double ck[5][4];           // zeta_eff coefficients c_k(lam) = sum_j c_Cm[k][j] lam^-j
    double I[5], J[5];         // I(lam), J(lam) of the B term
    double xp[5];              // x0^lam
    double f[3];               // f_i(alpha) of the chi correction
    double gam0;               // 10 (1 - tanh(10 (0.57 - alpha))) (exp(eps / kT) - 1)
    double be;                 // eps / kT
};

__global__ void rdf_T_kernel(int nT, int nc, const double* __restrict__ T, const double* __restrict__ dbh, const MiePair* __restrict__ mp,
                             RdfT* __restrict__ out) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nT * nc * nc) return;
    const int j = e % nc, i = (e / nc) % nc, t = e / (nc * nc);
    const MiePair q = mp[i * nc + j];
    const double* d = dbh + (size_t) t * nc;
    const double dij = (i == j) ? d[i] : (d[i] + d[j]) / 2.0;
    RdfT r;
    r.d3 = pow(dij, 3); r.sig3 = pow(q.sigma, 3); r.x0 = q.sigma / dij; r.eps = q.eps; r.la = q.la; r.lr = q.lr; r.C = q.C;
    r.lam[0] = q.la; r.lam[1] = q.lr; r.lam[2] = 2 * q.la; r.lam[3] = 2 * q.lr; r.lam[4] = q.la + q.lr;
    for (int k = 0; k < 5; k++) {
        const double lam = r.lam[k];
        for (int a = 0; a < 4; a++) { double c = 0; for (int b = 0; b < 4; b++) c += c_Cm[a][b] * pow(lam, (double) -b); r.ck[k][a] = c; }
        r.I[k] = -(pow(r.x0, 3. - lam) - 1.) / (lam - 3.);
        r.J[k] = -(pow(r.x0, 4. - lam) * (lam - 3.) - pow(r.x0, 3. - lam) * (lam - 4.) - 1.) / ((lam - 3.) * (lam - 4.));
        r.xp[k] = pow(r.x0, lam);
    }
    for (int a = 0; a < 3; a++) {
        double num = 0.0, den = 1.0;
        for (int n = 0; n <= 3; n++) num += c_phi[n][a] * pow(q.alpha, (double) n);
        for (int n = 4; n <= 6; n++) den += c_phi[n][a] * pow(q.alpha, (double) (n - 3));
        r.f[a] = num / den;
    }
    const double theta = exp(q.eps / (BOLTZMANN * T[t])) - 1.;
    r.gam0 = 10. * (1 - tanh(10. * (0.57 - q.alpha))) * theta;
    r.be = (1. / (BOLTZMANN * T[t])) * q.eps;
    out[e] = r;
}

__device__ void rdf_common(int nc, double rho, const double* x, const RdfT* __restrict__ rt /*[nc*nc] of this T*/, RdfCommon& c) {
    double s = 0, ss = 0;
    for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) {
        s += x[i] * x[j] * rt[i * nc + j].d3;
        ss += x[i] * x[j] * rt[i * nc + j].sig3;
    }
    c.rho = rho;
    c.zx = s * (PI_REF * rho / 6.0);           // zeta_x_func
    c.dzx = s * (PI_REF / 6.0);                // dzetax_drho_func
    c.zxs = ss * (PI_REF * rho / 6.0);         // zeta_x with sigma (rdf_chi_func, gamma_corr)
    const double z = c.zx, z2 = z * z, z3 = z2 * z, z4 = z2 * z2, omz = 1 - z, omz2 = omz * omz, omz3 = omz2 * omz;
    c.KHS = (omz2 * omz2) / (1 + 4 * z + 4 * z2 - 4 * z3 + z4);
    // sign of the z^4 term differs from K_HS in the reference (cpp/MieKinGas.h:176-179); kept
    c.dKHS = -4 * c.dzx * omz3 * (2 + 5 * z - z2 - 2 * z3) / p2(1 + 4 * z + 4 * z2 - 4 * z3 - z4);
    c.k0 = -log(1 - z) + (42. * z - 39. * z2 + 9. * z3 - 2 * z4) / (6 * omz3);
    c.k1 = (z4 + 6. * z2 - 12. * z) / (2. * omz3);
    c.k2 = -3. * z2 / (8. * omz2);
    c.k3 = (-z4 + 3. * z2 + 3. * z) / (6. * omz3);
}

// a1s + B and d(a1s + B)/drho for exponent k (lam[k]) of one pair
__device__ __forceinline__ void a1s_B(const RdfCommon& c, const RdfT& r, int k, double& aB, double& daB) {
    const double lam = r.lam[k], eps = r.eps, d3 = r.d3, zx = c.zx, zx2 = zx * zx, zx3 = zx2 * zx;
    const double ze = r.ck[k][0] * zx + r.ck[k][1] * zx2 + r.ck[k][2] * zx3 + r.ck[k][3] * (zx2 * zx2);
    const double dze = (r.ck[k][0] + r.ck[k][1] * 2 * zx + r.ck[k][2] * 3 * zx2 + r.ck[k][3] * 4 * zx3) * c.dzx;
    const double omze = 1. - ze, omze3 = p3(omze), omze4 = omze3 * omze;
    const double a1s = (-2. * c.rho * PI_REF * eps * d3 / (lam - 3)) * (1. - ze / 2.) / omze3;
    const double da1s = -2. * (PI_REF * eps * d3 / (lam - 3)) * (((1. - ze / 2.) / omze3) + c.rho * dze * (2.5 - ze) / omze4);
    const double I = r.I[k], J = r.J[k];
    const double omz = 1. - zx, omz3 = p3(omz), omz4 = omz3 * omz;
    const double B = 2. * PI_REF * c.rho * d3 * eps * (I * (1. - zx / 2.) / omz3 - J * 9. * zx * (1. + zx) / (2. * omz3));
    const double dB = B / c.rho + 2. * PI_REF * c.rho * eps * d3
                      * (I * (-0.5 * (1 - zx) + 3. * (1. - 0.5 * zx)) / omz4
                         - J * (9. * (1. + 2. * zx) * (1. - zx) + 27. * zx * (1. + zx)) / (2. * omz4)) * c.dzx;
    aB = a1s + B;
    daB = da1s + dB;
}

__device__ double rdf_pair(const RdfCommon& c, const RdfT& r) {
    const double x0 = r.x0, d3 = r.d3, eps = r.eps, la = r.la, lr = r.lr, C = r.C, rho = c.rho;
    const double gHS = exp(c.k0 + c.k1 * x0 + c.k2 * (x0 * x0) + c.k3 * (x0 * x0 * x0));
    double aB_la, daB_la, aB_lr, daB_lr, aB_2la, daB_2la, aB_2lr, daB_2lr, aB_lalr, daB_lalr;
    a1s_B(c, r, 0, aB_la, daB_la);
    a1s_B(c, r, 1, aB_lr, daB_lr);
    a1s_B(c, r, 2, aB_2la, daB_2la);
    a1s_B(c, r, 3, aB_2lr, daB_2lr);
    a1s_B(c, r, 4, aB_lalr, daB_lalr);
    const double x_la = r.xp[0], x_lr = r.xp[1], x_2la = r.xp[2], x_2lr = r.xp[3], x_lalr = r.xp[4];
    // g1 (rdf_g1_func)
    const double da1 = C * (x_la * daB_la - x_lr * daB_lr);
    const double g1 = (1.0 / (2. * PI_REF * eps * d3)) * (3. * da1 - C * la * x_la * aB_la / rho + C * lr * x_lr * aB_lr / rho);
    // g2 (rdf_g2_func with g2_correction = true)
    const double zs = c.zxs, zs2 = zs * zs, zs4 = zs2 * zs2;
    const double chi = r.f[0] * zs + r.f[1] * (zs4 * zs) + r.f[2] * (zs4 * zs4);
    const double C2 = C * C;
    const double a2 = .5 * c.KHS * (1 + chi) * eps * C2 * (x_2la * aB_2la - 2 * x_lalr * aB_lalr + x_2lr * aB_2lr);
    const double da2_div_chi = c.dKHS * (a2 / (c.KHS * (1 + chi)))
                             + .5 * c.KHS * eps * C2 * (x_2la * daB_2la - 2 * x_lalr * daB_lalr + x_2lr * daB_2lr);
    const double gamma_c = r.gam0 * zs * exp(-6.7 * zs + -8. * zs2);
    const double g2 = (1 + gamma_c) * (1. / (2 * PI_REF * (eps * eps) * d3))
                    * (3 * da2_div_chi + eps * c.KHS * C2 * (-lr * x_2lr * aB_2lr / rho + (la + lr) * x_lalr * aB_lalr / rho - la * x_2la * aB_2la / rho));
    const double be = r.be;
    return gHS * exp(be * (g1 / gHS) + (be * be) * (g2 / gHS));
}

// ---- LJ/spline: collision-integral correlation (cpp/LJSpline.cpp:123-165, table cpp/LJSpline.h:138-142) and the RDF at
// contact g = g0 + beta* g1 + beta*^2 g2 (cpp/LJSpline.h:113-125, cpp/LJSpline.cpp:16-118, constants cpp/LJSpline.h:167-205) ----
__constant__ double c_ljs_ocf[2][6] = {{0.12381066, -0.19473855, -0.12605004, 1.42026419, -1.14753069, 0.25563792},
                                       {0.2525206, -0.12502955, -1.02633768, 3.31321146, -2.58470784, 0.62745361}};
__constant__ double c_ljs_q[4][11] = {
    {-0.53572874, -0.37806724, -0.23016988, 0.34630713, 0.07308556, 5.27289035, -6.43165849, -1.63392187, 16.74650716, -27.36299942, -0.43778915},      // q_I1
    {1.60718622, -7.09047972e-02, -9.43081376e-01, -5.09423352e+00, 1.36949006e+00, -5.81631691e+01, -1.39795273e+00, 3.66229139e+00, -2.77904034e+02,
     5.65474349e+01, 1.01557784e+01},                                                                                                                  // q_J1
    {0.3592528, 0.27758414, 0.25126487, -0.22023453, -0.11005555, -4.46920946, 6.117315, 0.95067716, -13.88450783, 24.30452001, -0.63636868},          // q_I2
    {-0.5388792, 1.16814143e-02, 8.83906862e-02, 2.36081648e+00, -5.31094654e-01, 2.62590123e+01, -2.13956077e+00, -1.20815977e+00, 1.18753081e+02,
     -2.84345337e+01, -2.53564086e+00}};                                                                                                               // q_J2
__constant__ double c_ljs_phi[4] = {13.2795715, 3.84921999, -2.16074213, -1.40190623};

__device__ double ljs_recursive_factor(int l, int r, double Ts) {
    double d = -(1. / 6.);
    for (int n = 2; n < 7; n++) d += c_ljs_ocf[l - 1][n - 1] * pow(Ts, -(n - 1.) / 2.) * (1. - n) / 2.;
    return 1. + d / (r + 2);
}
__device__ double ljs_omega_correlation(int l, int r, double Ts) {
    double lnw = -(1. / 6.) * log(Ts);
    for (int n = 1; n < 7; n++) lnw += c_ljs_ocf[l - 1][n - 1] * pow(Ts, (1. - n) * 0.5);
    double v = exp(lnw);
    for (int rr = l; rr < r; rr++) v = ljs_recursive_factor(l, rr, Ts) * v;
    return v;
}
__device__ double ljs_fit(const double* q, double rs, double d) {          // I1, J1, I2, J2
    return q[0] + q[1] * rs + q[2] * pow(rs, 2) + q[3] * pow(rs, 3) + q[4] * pow(rs, 4) +
           rs * (q[5] * pow(rs, 2) + q[6] * rs + q[7]) * (d - 1) + rs * (q[8] * pow(rs, 2) + q[9] * rs + q[10]) * pow((d - 1), 2);
}
__device__ double ljs_fit_diff(const double* q, double rs, double d) {     // I1_diff, I2_diff
    return q[1] + 2 * q[2] * rs + 3 * q[3] * pow(rs, 2) + 4 * q[4] * pow(rs, 3) +
           (3 * q[5] * pow(rs, 2) + 2 * q[6] * rs + q[7]) * (d - 1) + (3 * q[8] * pow(rs, 2) + 2 * q[9] * rs + q[10]) * pow((d - 1), 2);
}
__device__ double ljs_rdf(double rho, double T, double sigma, double eps, double dbh) {
    const double eta = PI_REF * rho * pow(dbh, 3) / 6., x0 = sigma / dbh, d = dbh / sigma;
    const double k0 = -log(1. - eta) + (42. * eta - 39. * pow(eta, 2) + 9 * pow(eta, 3) - 2. * pow(eta, 4)) / (6. * pow((1. - eta), 3));
    const double k1 = (pow(eta, 4) + 6. * pow(eta, 2) - 12. * eta) / (2. * pow((1 - eta), 3));
    const double k2 = -3. * pow(eta, 2) / (8. * pow((1 - eta), 2));
    const double k3 = (-pow(eta, 4) + 3. * pow(eta, 2) + 3. * eta) / (6. * pow((1 - eta), 3));
    const double g0 = exp(k0 + k1 * x0 + k2 * pow(x0, 2) + k3 * pow(x0, 3));
    const double rs = rho * pow(sigma, 3);
    const double g1 = (3 * ljs_fit_diff(c_ljs_q[0], rs, d) * rs + 3 * ljs_fit(c_ljs_q[0], rs, d) + ljs_fit(c_ljs_q[1], rs, d)) / pow(d, 3);
    const double i2 = ljs_fit(c_ljs_q[2], rs, d), di2 = ljs_fit_diff(c_ljs_q[2], rs, d), j2 = ljs_fit(c_ljs_q[3], rs, d);
    const double khs = pow((1 - eta), 4) / (1. + 4. * eta + 4. * pow(eta, 2) - 4. * pow(eta, 3) + pow(eta, 4));
    const double tmp = pow(eta, 4) - 4. * pow(eta, 3) + 4. * pow(eta, 2) + 4. * eta + 1.;
    const double dkhs = (4. * (pow(eta, 2) - 5. * eta - 2.) * pow(1. - eta, 3) / pow(tmp, 2)) * (PI_REF * pow(d, 3)) / 6;
    const double g2_mca = (-3 / (2 * pow(d, 3))) * (i2 * khs + rs * i2 * dkhs + rs * khs * di2 + 2 * khs * j2 / 3);
    const double Tr = T * BOLTZMANN / eps, theta = exp(1 / Tr) - 1;
    const double gamma = c_ljs_phi[0] * eta * pow(x0, 3) * theta * exp(c_ljs_phi[1] * eta * pow(x0, 3) + c_ljs_phi[2] * pow(eta, 2) * pow(x0, 6)) * exp(c_ljs_phi[3] / Tr);
    const double g2 = (1 + gamma) * g2_mca;
    const double bs = eps / (BOLTZMANN * T);
    return g0 + bs * g1 + pow(bs, 2) * g2;
}
// One thread per (state point, slot): the LJ/spline model is single-component, every pair shares g.
__global__ void ljs_rdf_kernel(int n, int nslots, int nc, const int* __restrict__ tidx, const double* __restrict__ T_of, const double* __restrict__ Vm,
                               const double* __restrict__ dbh, MiePair q, double* __restrict__ rdf) {
    const long long e = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long) n * nslots) return;
    const int slot = (int) (e % nslots), sp = (int) (e / nslots), t = tidx[sp];
    if (t < 0) return;                              // temperature not in the cache: state_kernel reports KG_STATUS_T_NOT_CACHED
    const double rho = (slot == 0) ? AVOGADRO / Vm[sp] : AVOGADRO / 1e12;
    const double g = ljs_rdf(rho, T_of[t], q.sigma, q.eps, dbh[(size_t) t * nc]);
    for (int k = 0; k < nc * nc; k++) rdf[((size_t) sp * nslots + slot) * nc * nc + k] = g;
}

// ---- extended Sutherland / Feynman-Hibbs Mie (QuantumMie): effective parameters per temperature and the RDF at contact of
// ExtSutherland::saft_rdf (cpp/ExtendedSutherland.cpp:168-247, 292-590). u = eps sum_k C_eff,k(T) (sigma/r)^lam_k; QuantumMie has no
// density exponents (cpp/QuantumMie.cpp:73-109), so sigma_eff, eps_eff, alpha_vdw and the Barker-Henderson diameters depend on T only
// and the reference's forward-mode d/drho (autodiff duals, :388-394, 483-489) acts on the explicit rho of a_1s, B, zeta_x, K_HS only.
struct EsPair { double sigma, eps; int nterms, pad; double lam[KG_MAX_TERMS]; };
struct EsParams { double C[KG_MAX_TERMS]; double sigma_eff, eps_eff, alpha, dbh; };     // per (T, internal pair i*nc+j)

__device__ void es_pot(const EsPair& q, const double* C, double r, double& u, double& ur, double& urr) {
    u = ur = urr = 0.0;
    for (int k = 0; k < q.nterms; k++) {
        const double t = C[k] * pow(q.sigma / r, q.lam[k]);
        u += t; ur -= q.lam[k] * t / r; urr += q.lam[k] * (q.lam[k] + 1.) * t / (r * r);
    }
    u *= q.eps; ur *= q.eps; urr *= q.eps;
}

// One thread per (t, i <= j): Newton for sigma_eff and r_min from sigma (set_sigma_eff / set_epsilon_eff), alpha_vdw, and for
// i == j the 20-point Gauss-Legendre Barker-Henderson diameter. Cross diameters are filled by es_dbh_mix_kernel.
__global__ void es_params_kernel(int nT, int nc, const double* __restrict__ T, const EsPair* __restrict__ pairs, const double* __restrict__ coeff,
                                 const int* __restrict__ pair_of, int npairs_dev, EsParams* __restrict__ out) {
    const int npu = nc * (nc + 1) / 2;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nT * npu) return;
    const int t = e / npu;
    int i = 0, rest = e % npu;
    while (rest >= nc - i) { rest -= nc - i; i++; }
    const int j = i + rest;
    const int pd = pair_of[i * nc + j];
    const EsPair q = pairs[pd];
    EsParams P;
    for (int k = 0; k < KG_MAX_TERMS; k++) P.C[k] = (k < q.nterms) ? coeff[((size_t) t * npairs_dev + pd) * KG_MAX_TERMS + k] : 0.0;
    double u, ur, urr, s = q.sigma;
    int guard = 0;
    do { es_pot(q, P.C, s, u, ur, urr); s -= u / ur; } while (fabs(u) / q.eps > 1e-6 && guard++ < 200);
    P.sigma_eff = s;
    double rm = q.sigma;
    guard = 0;
    do { es_pot(q, P.C, rm, u, ur, urr); rm -= ur / urr; } while (fabs(ur) * rm / q.eps > 1e-10 && guard++ < 200);
    es_pot(q, P.C, rm, u, ur, urr);
    P.eps_eff = -u;
    double a = 0.;
    for (int k = 0; k < q.nterms; k++) a -= P.C[k] * q.eps * pow(q.sigma / P.sigma_eff, q.lam[k]) / (q.lam[k] - 3.);
    P.alpha = a / P.eps_eff;
    P.dbh = 0.0;
    if (i == j) {
        const double beta = 1. / (BOLTZMANN * T[t]);
        double d = 2.;
        for (int n = 0; n < 20; n++) {
            es_pot(q, P.C, P.sigma_eff * (c_gl_x[n] / 4. + 3. / 4.), u, ur, urr);
            d += c_gl_w[n] * (1. - exp(-beta * u));
        }
        P.dbh = d * (P.sigma_eff / 4.);
    }
    out[(size_t) t * nc * nc + i * nc + j] = P;
    if (i != j) out[(size_t) t * nc * nc + j * nc + i] = P;
}
__global__ void es_dbh_mix_kernel(int nT, int nc, EsParams* __restrict__ out) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nT * nc * nc) return;
    const int t = e / (nc * nc), i = (e / nc) % nc, j = e % nc;
    if (i == j) return;
    EsParams* base = out + (size_t) t * nc * nc;
    base[i * nc + j].dbh = (base[i * nc + i].dbh + base[j * nc + j].dbh) / 2.0;
}

// first-order dual number in rho
struct D1 { double v, d; };
__device__ __forceinline__ D1 operator+(D1 a, D1 b) { return {a.v + b.v, a.d + b.d}; }
__device__ __forceinline__ D1 operator-(D1 a, D1 b) { return {a.v - b.v, a.d - b.d}; }
__device__ __forceinline__ D1 operator*(D1 a, D1 b) { return {a.v * b.v, a.d * b.v + a.v * b.d}; }
__device__ __forceinline__ D1 operator/(D1 a, D1 b) { const double q = a.v / b.v; return {q, (a.d - q * b.d) / b.v}; }
__device__ __forceinline__ D1 operator*(double a, D1 b) { return {a * b.v, a * b.d}; }
__device__ __forceinline__ D1 operator+(double a, D1 b) { return {a + b.v, b.d}; }
__device__ __forceinline__ D1 operator-(double a, D1 b) { return {a - b.v, -b.d}; }
__device__ __forceinline__ D1 dpowi(D1 a, int n) { const double p = pow(a.v, (double) (n - 1)); return {p * a.v, n * p * a.d}; }

struct EsState { D1 rho, zeta; double zeta_hs; };

__device__ D1 es_zeta_eff(D1 zeta_x, double lam) {
    D1 z = {0.0, 0.0};
    for (int i = 0; i < 4; i++) {
        double c = 0.0;
        for (int j = 0; j < 4; j++) c += c_Cm[i][j] * pow(lam, (double) -j);
        z = z + c * dpowi(zeta_x, i + 1);
    }
    return z;
}
// a_1s + B for exponent lam (a_1s_func, B_func, I_func, J_func; cpp/ExtendedSutherland.cpp:399-441)
__device__ D1 es_a1s_plus_B(const EsState& S, double eps, double d, double xeff, double lam) {
    const D1 ze = es_zeta_eff(S.zeta, lam);
    const D1 one_m = 1.0 - ze;
    const D1 a1s = ((-2. * PI_REF * eps * pow(d, 3) / (lam - 3)) * S.rho) * (1. - 0.5 * ze) / dpowi(one_m, 3);
    const double I = -(pow(xeff, 3. - lam) - 1.) / (lam - 3.);
    const double J = -(pow(xeff, 4. - lam) * (lam - 3.) - pow(xeff, 3. - lam) * (lam - 4.) - 1.) / ((lam - 3.) * (lam - 4.));
    const D1 om = 1.0 - S.zeta;
    const D1 om3 = dpowi(om, 3);
    const D1 B = ((2. * PI_REF * pow(d, 3) * eps) * S.rho) * (I * ((1. - 0.5 * S.zeta) / om3) - (J * 9.) * (S.zeta * (1. + S.zeta)) / (2. * om3));
    return a1s + B;
}
__device__ __forceinline__ D1 es_KHS(D1 z) {
    return dpowi(1.0 - z, 4) / (1.0 + 4.0 * z + 4.0 * dpowi(z, 2) - 4.0 * dpowi(z, 3) + dpowi(z, 4));
}

// One thread per (state point, slot, i <= j): ExtSutherland::saft_rdf order 2 with the g2 correction (model_rdf default).
__global__ void es_rdf_kernel(int n, int nslots, TModel tm, const int* __restrict__ tidx, const double* __restrict__ T_of, const double* __restrict__ Vm,
                              const double* __restrict__ x, const EsPair* __restrict__ pairs_int, const EsParams* __restrict__ prm, double* __restrict__ rdf) {
    const int nc = tm.nc, npu = nc * (nc + 1) / 2;
    const long long e = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long) n * nslots * npu) return;
    const int pu = (int) (e % npu), slot = (int) ((e / npu) % nslots), sp = (int) (e / ((long long) npu * nslots));
    int i = 0, rest = pu;
    while (rest >= nc - i) { rest -= nc - i; i++; }
    const int j = i + rest;
    const int t = tidx[sp];
    if (t < 0) return;                              // temperature not in the cache: state_kernel reports KG_STATUS_T_NOT_CACHED
    const double T = T_of[t];
    const double rho = (slot == 0) ? AVOGADRO / Vm[sp] : AVOGADRO / 1e12;
    const EsParams* P = prm + (size_t) t * nc * nc;
    double zx = 0.0, zhs = 0.0;
    for (int a = 0; a < nc; a++) for (int b = 0; b < nc; b++) {
        const double xa = tm.is_singlecomp ? 0.5 : x[(size_t) sp * nc + a], xb = tm.is_singlecomp ? 0.5 : x[(size_t) sp * nc + b];
        zx += xa * xb * pow(P[a * nc + b].dbh, 3);
        zhs += xa * xb * pow(P[a * nc + b].sigma_eff, 3);
    }
    EsState S;
    S.rho = {rho, 1.0};
    S.zeta = {zx * (PI_REF * rho / 6.0), zx * (PI_REF / 6.0)};
    S.zeta_hs = zhs * (PI_REF * rho / 6.0);
    const EsPair q = pairs_int[i * nc + j];
    const EsParams p = P[i * nc + j];
    const double d = p.dbh, xeff = p.sigma_eff / d, x0 = q.sigma / d, z = S.zeta.v;
    // g0, rdf_g0_func :346-362
    const double k0 = -log(1 - z) + (42. * z - 39. * pow(z, 2) + 9. * pow(z, 3) - 2 * pow(z, 4)) / (6 * pow(1 - z, 3));
    const double k1 = (pow(z, 4) + 6. * pow(z, 2) - 12. * z) / (2. * pow(1 - z, 3));
    const double k2 = -3. * pow(z, 2) / (8. * pow(1 - z, 2));
    const double k3 = (-pow(z, 4.) + 3. * pow(z, 2) + 3. * z) / (6. * pow(1 - z, 3));
    const double g0 = exp(k0 + k1 * xeff + k2 * pow(xeff, 2) + k3 * pow(xeff, 3));
    // g1, rdf_g1_func :376-397 ; a1 = - sum_k C_k x0^lam_k (a1s_k + B_k)
    double g1 = 0.0;
    D1 a1 = {0.0, 0.0};
    for (int k = 0; k < q.nterms; k++) {
        const D1 ab = es_a1s_plus_B(S, q.eps, d, xeff, q.lam[k]);
        const double w = p.C[k] * pow(x0, q.lam[k]);
        g1 += q.lam[k] * w * ab.v / rho;
        a1 = a1 - w * ab;
    }
    g1 += 3. * a1.d;
    g1 *= (1.0 / (2. * PI_REF * q.eps * pow(d, 3)));
    // g2, rdf_g2_func :443-478 ; a2/chi = K_HS eps / 2 sum_kl C_k C_l x0^lam_kl (a1s_kl + B_kl)
    const D1 KHS = es_KHS(S.zeta);
    double g2 = 0.0;
    D1 a2 = {0.0, 0.0};
    for (int k = 0; k < q.nterms; k++) for (int l = 0; l < q.nterms; l++) {
        const double lkl = q.lam[k] + q.lam[l];
        const D1 ab = es_a1s_plus_B(S, q.eps, d, xeff, lkl);
        const double w = p.C[k] * p.C[l] * pow(x0, lkl);
        g2 -= KHS.v * q.eps * w * q.lam[k] * ab.v / rho;
        a2 = a2 + w * ab;
    }
    a2 = (0.5 * q.eps) * (KHS * a2);
    g2 += 3. * a2.d;
    const double theta = exp(p.eps_eff / (BOLTZMANN * T)) - 1.;
    const double gamma = 10. * (1 - tanh(10. * (0.57 - p.alpha))) * S.zeta_hs * theta * exp(-6.7 * S.zeta_hs + -8. * pow(S.zeta_hs, 2));
    g2 *= (1 + gamma) * (1. / (2. * PI_REF * pow(q.eps, 2) * pow(d, 3)));
    const double be = q.eps / (BOLTZMANN * T);
    const double g = g0 * exp(be * (g1 / g0) + pow(be, 2) * (g2 / g0));
    double* out = rdf + ((size_t) sp * nslots + slot) * nc * nc;
    out[i * nc + j] = g;
    out[j * nc + i] = g;
}

// One thread per (state point, slot, i <= j). slot 0: the state point's density; slot 1: N_A / 1e12 (the dilute viscosity of the
// internal-conductivity term, cpp/KineticGas_properties.cpp:133). rdf [n][nslots][nc][nc].
__global__ void rdf_kernel(int n, int nslots, TModel tm, const int* __restrict__ tidx, const double* __restrict__ Vm,
                           const double* __restrict__ x, const RdfT* __restrict__ rdfT, double* __restrict__ rdf) {
    const int nc = tm.nc, npu = nc * (nc + 1) / 2;
    const long long e = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long) n * nslots * npu) return;
    const int pu = (int) (e % npu), slot = (int) ((e / npu) % nslots), sp = (int) (e / ((long long) npu * nslots));
    int i = 0, rest = pu;
    while (rest >= nc - i) { rest -= nc - i; i++; }
    const int j = i + rest;
    double xs[MAXC];
    for (int k = 0; k < nc; k++) xs[k] = tm.is_singlecomp ? 0.5 : x[(size_t) sp * nc + k];
    const int t = tidx[sp];
    if (t < 0) return;                              // temperature not in the cache: state_kernel reports KG_STATUS_T_NOT_CACHED
    const double rho = (slot == 0) ? AVOGADRO / Vm[sp] : AVOGADRO / 1e12;
    const RdfT* rt = rdfT + (size_t) t * nc * nc;
    RdfCommon c;
    rdf_common(nc, rho, xs, rt, c);
    const double g = rdf_pair(c, rt[i * nc + j]);
    double* out = rdf + ((size_t) sp * nslots + slot) * nc * nc;
    out[i * nc + j] = g;
    out[j * nc + i] = g;
}

// ---- per-state-point assembly, LU and property formulas ----
struct StateArgs {
    int n, nslots, nT;
    kg_transport_opts opt;
    const int* tidx; const double* T_of; const double* Vm; const double* x;
    const double* rdf;          // [n][nslots][nc][nc]
    const double *mtl, *etl;    // [nT][npairs_dev] (null for ideal gas)
    const double* tlpow;        // [nT][npairs_dev][4]: etl^3, mtl^3, mtl^4, etl^4 (host libm pow, as the reference forms them)
    const double* brk;          // [nT][4][nc][nc][N][N]
    const double *dmudn, *cp_ideal, *pressure, *dvdn;
    double *viscosity, *conductivity, *contribs, *D, *DT, *kT, *soret, *K, *Kp, *cond_matrix, *cond_vector, *visc_matrix, *visc_vector;
    int* status;
};

// Unblocked partial-pivot LU, the algorithm Eigen's partialPivLu runs for sizes <= 16 (row-max pivot, first maximum wins,
// column scaled by division, rank-1 update with separately rounded multiply and subtract). A is n x n row-major, stride ld.
// ES = element stride: 1 for the small matrices in thread-local memory; the block size for the Sonine system, which lives in shared
// memory interleaved over the threads of the block (element e of thread t at [e * ES + t]: conflict-free, and an order of magnitude
// closer than the L2-backed local memory a 10 x 10 system per thread otherwise streams through).
template<int ES>
__device__ int lu_factor(double* A, int n, int ld, int* piv, double zero_pivot_rel = 0.0) {
#define AT(i, j) A[((i) * ld + (j)) * ES]
    int singular = 0;
    double scale = 0.0;
    if (zero_pivot_rel > 0.0) for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) scale = fmax(scale, fabs(AT(i, j)));
    for (int k = 0; k < n; k++) {
        int pr = k; double best = fabs(AT(k, k));
        for (int i = k + 1; i < n; i++) { const double v = fabs(AT(i, k)); if (v > best) { best = v; pr = i; } }
        piv[k] = pr;
        if (best != 0.0) {
            if (pr != k) for (int j = 0; j < n; j++) { const double t = AT(k, j); AT(k, j) = AT(pr, j); AT(pr, j) = t; }
            const double d = AT(k, k);
            for (int i = k + 1; i < n; i++) AT(i, k) /= d;
        }
        else if (zero_pivot_rel > 0.0) AT(k, k) = zero_pivot_rel * scale;
        else singular = 1;
        for (int i = k + 1; i < n; i++) {
            const double lik = AT(i, k);
            for (int j = k + 1; j < n; j++) AT(i, j) = __dsub_rn(AT(i, j), __dmul_rn(lik, AT(k, j)));
        }
    }
    return singular;
}
template<int ES>
__device__ void lu_solve(const double* A, int n, int ld, const int* piv, double* b) {
    for (int k = 0; k < n; k++) if (piv[k] != k) { const double t = b[k]; b[k] = b[piv[k]]; b[piv[k]] = t; }
    for (int i = 1; i < n; i++) { double s = b[i]; for (int j = 0; j < i; j++) s = __dsub_rn(s, __dmul_rn(AT(i, j), b[j])); b[i] = s; }
    for (int i = n - 1; i >= 0; i--) { double s = b[i]; for (int j = i + 1; j < n; j++) s = __dsub_rn(s, __dmul_rn(AT(i, j), b[j])); b[i] = s / AT(i, i); }
}

struct SP {                     // one state point's scalars, all lanes keep them in registers / local memory
    int nc, N, t;
    double T, rho, Vm;
    double x[MAXC], wt[MAXC];
};

__device__ __forceinline__ double brk_at(const StateArgs& a, const TModel& tm, int t, int which, int i, int j, int p, int q) {
    const int per_t = tm.nc * tm.nc * tm.N * tm.N;
    return a.brk[((size_t) t * 4 + which) * per_t + ((i * tm.nc + j) * tm.N + p) * tm.N + q];
}

// Lambda (cpp/KineticGas.cpp:271-318) into A (n x n, n = N nc)
template<int ES>
__device__ void build_cond_matrix(const StateArgs& a, const TModel& tm, const SP& s, const double* g, double* A, int ld) {
    const int nc = s.nc, N = s.N, n = N * nc;
    for (int r = 0; r < n; r++) for (int c = 0; c < n; c++) AT(r, c) = 0.0;
    for (int i = 0; i < nc; i++) AT(0, i) = s.wt[i];
    const double c75 = 75. * (BOLTZMANN * BOLTZMANN) * s.T;
    for (int p = 0; p < N; p++) for (int i = 0; i < nc; i++) {
        if (i == 0 && p == 0) continue;                     // replaced by the weight-fraction row
        const int row = (i == 0) ? p : N + (nc - 1) * p + (i - 1);
        for (int q = 0; q < N; q++) for (int j = 0; j < nc; j++) {
            double v = s.x[i] * s.x[j] * g[i * nc + j] * brk_at(a, tm, s.t, 0, i, j, p, q);
            if (i == j) for (int l = 0; l < nc; l++) v += s.x[i] * s.x[l] * g[i * nc + l] * brk_at(a, tm, s.t, 1, i, l, p, q);
            v *= 8. * tm.sqmm[i][j] / c75;
            AT(row, nc * q + j) = v;
        }
    }
}

// B (cpp/KineticGas.cpp:418-439)
template<int ES>
__device__ void build_visc_matrix(const StateArgs& a, const TModel& tm, const SP& s, const double* g, double* A, int ld) {
    const int nc = s.nc, N = s.N;
    const double c5 = 2. / (5. * BOLTZMANN * s.T);
    for (int p = 0; p < N; p++) for (int i = 0; i < nc; i++) for (int q = 0; q < N; q++) for (int j = 0; j < nc; j++) {
        double v = s.x[i] * s.x[j] * g[i * nc + j] * brk_at(a, tm, s.t, 2, i, j, p, q);
        if (i == j) for (int k = 0; k < nc; k++) v += s.x[i] * s.x[k] * g[i * nc + k] * brk_at(a, tm, s.t, 3, i, k, p, q);
        v *= c5;
        AT(p * nc + i, q * nc + j) = v;
    }
}

// K and K' (cpp/KineticGas.cpp:207-239); tl3 = etl^3 or mtl^3 by pair
__device__ void k_factors(const StateArgs& a, const TModel& tm, const SP& s, double rho, const double* g, double* K, double* Kp) {
    const int nc = s.nc;
    for (int i = 0; i < nc; i++) {
        if (tm.is_idealgas) { K[i] = 1.; Kp[i] = 1.; continue; }
        double k = 0, kp = 0;
        for (int j = 0; j < nc; j++) {
            const int pr = tm.pair_of[i][j];
            const double Mij = tm.m[i] / (tm.m[i] + tm.m[j]), Mji = tm.m[j] / (tm.m[j] + tm.m[i]);
            const double* tp = a.tlpow + ((size_t) s.t * tm.npairs_dev + pr) * 4;
            k += s.x[j] * tp[0] * Mij * Mji * g[i * nc + j];
            kp += s.x[j] * Mji * tp[1] * g[i * nc + j];
        }
        K[i] = k * (24. * PI_REF * rho / 15) + 1;
        Kp[i] = kp * (8. * PI_REF * rho / 15) + 1.;
    }
}

// KineticGas::viscosity at (T, rho) with the given rdf (cpp/KineticGas_properties.cpp:28-58). A, piv, b are scratch.
template<int ES>
__device__ double viscosity_at(const StateArgs& a, const TModel& tm, const SP& s, double rho, const double* g, const double* Kp,
                               double* A, int ld, int* piv, double* b, int& st, double* mat_out, double* vec_out) {
    const int nc = s.nc, N = s.N, n = N * nc;
    build_visc_matrix<ES>(a, tm, s, g, A, ld);
    for (int i = 0; i < n; i++) b[i] = 0.0;
    for (int i = 0; i < nc; i++) b[i] = 2. * s.x[i] * Kp[i] / (BOLTZMANN * s.T);
    if (mat_out) for (int r = 0; r < n; r++) for (int c = 0; c < n; c++) mat_out[r * n + c] = AT(r, c);
    if (vec_out) for (int i = 0; i < n; i++) vec_out[i] = b[i];
    if (lu_factor<ES>(A, n, ld, piv)) st |= KG_STATUS_SINGULAR;
    lu_solve<ES>(A, n, ld, piv, b);
    double eta_p = 0.0;
    for (int i = 0; i < nc; i++) eta_p += Kp[i] * s.x[i] * b[i];
    eta_p *= BOLTZMANN * s.T / 2.0;
    if (tm.is_idealgas) return eta_p;
    double eta_pp = 0.0;
    for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) {
        const double mtl4 = a.tlpow[((size_t) s.t * tm.npairs_dev + tm.pair_of[i][j]) * 4 + 2];
        eta_pp += tm.sqred[i][j] * s.x[i] * s.x[j] * mtl4 * g[i * nc + j];
    }
    eta_pp *= 4.0 * (rho * rho) * sqrt(2.0 * PI_REF * BOLTZMANN * s.T) / 15.0;
    return eta_p + eta_pp;
}

__host__ __device__ constexpr int state_tpb(int maxn) { return maxn <= 12 ? 64 : 32; }      // threads per block: the block's Sonine systems fill its shared memory
template<int MAXN>
__global__ void __launch_bounds__(state_tpb(MAXN)) state_kernel(TModel tm, StateArgs a) {
    extern __shared__ double s_sys[];                   // [MAXN * MAXN][block] : the N nc x N nc system of every thread, interleaved
    constexpr int ES = state_tpb(MAXN);
    const int sp = blockIdx.x * blockDim.x + threadIdx.x;
    if (sp >= a.n) return;
    const int nc = tm.nc, N = tm.N, n = N * nc, ld = MAXN;
    const unsigned props = a.opt.properties;
    int st = 0;
    SP s; s.nc = nc; s.N = N; s.t = a.tidx[sp];
    const int nD = a.opt.do_compress ? (nc - 1) : nc;
    if (s.t < 0) {          // temperature not in the cache: poison the outputs, flag, leave
        const double qnan = nan("");
        if (a.viscosity) a.viscosity[sp] = qnan;
        if (a.conductivity) a.conductivity[sp] = qnan;
        if (a.contribs) for (int k = 0; k < 3; k++) a.contribs[(size_t) sp * 3 + k] = qnan;
        if (a.D) for (int k = 0; k < nD * nD; k++) a.D[(size_t) sp * nD * nD + k] = qnan;
        if (a.DT) for (int k = 0; k < nc; k++) a.DT[(size_t) sp * nc + k] = qnan;
        if (a.kT) for (int k = 0; k < nc; k++) a.kT[(size_t) sp * nc + k] = qnan;
        if (a.soret) for (int k = 0; k < nc; k++) a.soret[(size_t) sp * nc + k] = qnan;
        if (a.K) for (int k = 0; k < nc; k++) a.K[(size_t) sp * nc + k] = qnan;
        if (a.Kp) for (int k = 0; k < nc; k++) a.Kp[(size_t) sp * nc + k] = qnan;
        if (a.cond_matrix) for (int k = 0; k < n * n; k++) a.cond_matrix[(size_t) sp * n * n + k] = qnan;
        if (a.cond_vector) for (int k = 0; k < n; k++) a.cond_vector[(size_t) sp * n + k] = qnan;
        if (a.visc_matrix) for (int k = 0; k < n * n; k++) a.visc_matrix[(size_t) sp * n * n + k] = qnan;
        if (a.visc_vector) for (int k = 0; k < n; k++) a.visc_vector[(size_t) sp * n + k] = qnan;
        if (a.status) a.status[sp] = KG_STATUS_T_NOT_CACHED;
        return;
    }
    s.T = a.T_of[s.t]; s.Vm = a.Vm[sp]; s.rho = AVOGADRO / s.Vm;
    double wsum = 0.0, xsum = 0.0;
    for (int i = 0; i < nc; i++) { s.x[i] = tm.is_singlecomp ? 0.5 : a.x[(size_t) sp * nc + i]; wsum += s.x[i] * tm.m[i]; xsum += s.x[i]; }
    if (fabs(xsum - 1.0) > 1e-12) st |= KG_STATUS_COMPOSITION;      // warning bit: the check of pykingas/py_KineticGas.py:119-141 (FLTEPS = 1e-12)
    for (int i = 0; i < nc; i++) s.wt[i] = s.x[i] * tm.m[i] / wsum;
    double g[MAXC * MAXC], K[MAXC], Kp[MAXC];
    for (int k = 0; k < nc * nc; k++) g[k] = tm.is_idealgas ? 1.0 : a.rdf[((size_t) sp * a.nslots) * nc * nc + k];
    k_factors(a, tm, s, s.rho, g, K, Kp);
    if (a.K) for (int i = 0; i < nc; i++) a.K[(size_t) sp * nc + i] = K[i];
    if (a.Kp) for (int i = 0; i < nc; i++) a.Kp[(size_t) sp * nc + i] = Kp[i];

    double* A = s_sys + threadIdx.x;
    double b[MAXN], th[MAXN], dsol[MAXC * MAXN];
    int piv[MAXN];
    const bool single = tm.is_singlecomp != 0;
    if (props & (KG_PROP_CONDUCTIVITY | KG_PROP_DIFFUSION | KG_PROP_THERMAL_DIFFUSION)) {
        build_cond_matrix<ES>(a, tm, s, g, A, ld);
        if (a.cond_matrix) for (int r = 0; r < n; r++) for (int c = 0; c < n; c++) a.cond_matrix[((size_t) sp * n + r) * n + c] = AT(r, c);
        if (lu_factor<ES>(A, n, ld, piv)) st |= KG_STATUS_SINGULAR;
        const bool need_th = (props & (KG_PROP_CONDUCTIVITY | KG_PROP_THERMAL_DIFFUSION)) != 0;
        const bool need_d = (props & (KG_PROP_DIFFUSION | KG_PROP_THERMAL_DIFFUSION)) || ((props & KG_PROP_CONDUCTIVITY) && !single && N > 1);
        if (need_th || a.cond_vector) {
            // get_conductivity_vector, cpp/KineticGas.cpp:320-330 (needs N >= 2)
            for (int i = 0; i < n; i++) th[i] = 0.0;
            if (N > 1) {
                th[1] = s.x[0] * K[0] * 4. / (5. * BOLTZMANN);
                for (int i = 1; i < nc; i++) th[N + (nc - 1) + (i - 1)] = s.x[i] * K[i] * 4. / (5. * BOLTZMANN);
            }
            if (a.cond_vector) for (int i = 0; i < n; i++) a.cond_vector[(size_t) sp * n + i] = th[i];
            lu_solve<ES>(A, n, ld, piv, th);
        }
        if (need_d) {
            // get_diffusion_vector, cpp/KineticGas.cpp:390-403: block k of the block-diagonal system, d[q][i][k] = dsol[k][nc q + i]
            for (int k = 0; k < nc; k++) {
                double* dk = dsol + k * MAXN;
                for (int i = 0; i < n; i++) dk[i] = 0.0;
                for (int i = 1; i < nc; i++) {
                    double v = -8. * s.wt[i] / (25. * BOLTZMANN);
                    if (i == k) v += 8. / (25. * BOLTZMANN);
                    dk[N + (i - 1)] = v;
                }
                lu_solve<ES>(A, n, ld, piv, dk);
            }
        }
        // E_ij and ksi (cpp/KineticGas_properties.cpp:459-500), frame matrices (:348-396), D in the centre-of-mass frame (:332-346)
        double E[MAXC * MAXC], ksi[MAXC], psi[MAXC * MAXC], Dcom[MAXC * MAXC], Dd[MAXC * MAXC];
        int dep = a.opt.dependent_idx;
        if (dep < 0 && a.opt.frame_of_reference == KG_FOR_SOLVENT) dep = a.opt.solvent_idx;
        while (dep < 0) dep += nc;
        const bool want_E = a.dmudn != nullptr && (props & (KG_PROP_DIFFUSION | KG_PROP_THERMAL_DIFFUSION));
        if (want_E) {
            if (single) {
                const double dp = s.Vm * a.dmudn[sp], RT = GAS_CONSTANT * s.T, rm = 1 / s.Vm;
                E[0] = (dp + RT * s.x[1] / (rm * s.x[0])) / AVOGADRO;
                E[1] = E[2] = (dp - RT / rm) / AVOGADRO;
                E[3] = (dp + RT * s.x[0] / (rm * s.x[1])) / AVOGADRO;
            }
            else for (int k = 0; k < nc * nc; k++) E[k] = s.Vm * a.dmudn[(size_t) sp * nc * nc + k] / AVOGADRO;
            for (int i = 0; i < nc; i++) { const double ni = s.x[i] / s.Vm; for (int j = 0; j < nc; j++) E[i * nc + j] = (ni / (BOLTZMANN * s.T)) * E[i * nc + j]; }
            for (int i = 0; i < nc; i++) { ksi[i] = 0.; for (int j = 0; j < nc; j++) ksi[i] += E[j * nc + i]; }
            for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) {
                double v = (i == j) ? 1.0 : 0.0;
                switch (a.opt.frame_of_reference) {
                    case KG_FOR_CON: case KG_FOR_ZARATE: case KG_FOR_ZARATE_X: case KG_FOR_ZARATE_W:
                        v += s.x[i] * ((s.wt[j] / s.x[j]) - 1); break;
                    case KG_FOR_SOLVENT: v += s.x[i] * ((s.wt[j] / s.x[j]) - ((j == a.opt.solvent_idx ? 1 : 0) / s.x[a.opt.solvent_idx])); break;
                    case KG_FOR_COV: v += s.x[i] * ((s.wt[j] / s.x[j]) - (a.dvdn[(size_t) sp * (single ? 1 : nc) + (single ? 0 : j)] / s.Vm)); break;
                    default: break;
                }
                psi[i * nc + j] = v;
            }
            for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) {
                double v = 0.0;
                for (int k = 0; k < nc; k++) v += dsol[k * MAXN + i] * E[k * nc + j];       // d[0][i][k]
                Dcom[i * nc + j] = v * (s.x[i] / (2 * s.rho));
            }
            for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) {
                double v = 0.0;
                for (int k = 0; k < nc; k++) v += psi[i * nc + k] * Dcom[k * nc + j];
                Dd[i * nc + j] = v;
            }
        }
        // Ortiz de Zarate frames (cpp/KineticGas_properties.cpp:179-193, 262-277, 325-330, 397-418): similarity transforms of
        // the compressed molar-frame matrix D^(x) with X = diag(x) - x x^T (mole fractions) and W (weight fractions)
        const int fr = a.opt.frame_of_reference;
        const bool zar = fr >= KG_FOR_ZARATE;
        const int nm = nc - 1;
        double Dz[MAXC * MAXC], Xm[MAXC * MAXC], Xlu[MAXC * MAXC];
        int idm[MAXC], pz[MAXC];
        if (want_E && zar) {
            int q = 0;
            for (int i = 0; i < nc; i++) if (i != dep) idm[q++] = i;
            double Dx[MAXC * MAXC];
            for (int i = 0; i < nm; i++) for (int j = 0; j < nm; j++)
                Dx[i * MAXC + j] = Dd[idm[i] * nc + idm[j]] - (ksi[idm[j]] / ksi[dep]) * Dd[idm[i] * nc + dep];
            for (int i = 0; i < nm; i++) for (int j = 0; j < nm; j++) {
                Xm[i * MAXC + j] = ((i == j) ? s.x[idm[i]] : 0.0) - s.x[idm[i]] * s.x[idm[j]];
                Xlu[i * MAXC + j] = Xm[i * MAXC + j];
            }
            if (lu_factor<1>(Xlu, nm, MAXC, pz)) st |= KG_STATUS_SINGULAR;
            if (fr == KG_FOR_ZARATE_X) { for (int k = 0; k < nm; k++) for (int j = 0; j < nm; j++) Dz[k * MAXC + j] = Dx[k * MAXC + j]; }
            else {
                // D = X^-1 D^(x) X, column by column
                for (int cidx = 0; cidx < nm; cidx++) {
                    double col[MAXC];
                    for (int i = 0; i < nm; i++) { double v = 0.; for (int k = 0; k < nm; k++) v += Dx[i * MAXC + k] * Xm[k * MAXC + cidx]; col[i] = v; }
                    lu_solve<1>(Xlu, nm, MAXC, pz, col);
                    for (int i = 0; i < nm; i++) Dz[i * MAXC + cidx] = col[i];
                }
                if (fr == KG_FOR_ZARATE_W) {
                    // D^(w) = W D W^-1: rows m_r solve W^T m_r = (W D)_r^T
                    double Wm[MAXC * MAXC], WT[MAXC * MAXC], WD[MAXC * MAXC]; int pw[MAXC];
                    for (int i = 0; i < nm; i++) for (int j = 0; j < nm; j++) {
                        Wm[i * MAXC + j] = ((i == j) ? s.wt[idm[i]] : 0.0) - s.wt[idm[i]] * s.wt[idm[j]];
                        WT[j * MAXC + i] = Wm[i * MAXC + j];
                    }
                    for (int i = 0; i < nm; i++) for (int j = 0; j < nm; j++) { double v = 0.; for (int k = 0; k < nm; k++) v += Wm[i * MAXC + k] * Dz[k * MAXC + j]; WD[i * MAXC + j] = v; }
                    if (lu_factor<1>(WT, nm, MAXC, pw)) st |= KG_STATUS_SINGULAR;
                    for (int r = 0; r < nm; r++) {
                        double row[MAXC];
                        for (int j = 0; j < nm; j++) row[j] = WD[r * MAXC + j];
                        lu_solve<1>(WT, nm, MAXC, pw, row);
                        for (int j = 0; j < nm; j++) Dz[r * MAXC + j] = row[j];
                    }
                }
            }
            if ((props & KG_PROP_DIFFUSION) && a.D) for (int i = 0; i < nm; i++) for (int j = 0; j < nm; j++) a.D[((size_t) sp * nD + i) * nD + j] = Dz[i * MAXC + j];
        }
        if (want_E && !zar && (props & KG_PROP_DIFFUSION) && a.D) {
            int oi = 0;
            for (int i = 0; i < nc; i++) {
                if (a.opt.do_compress && i == dep) continue;
                int oj = 0;
                for (int j = 0; j < nc; j++) {
                    if (a.opt.do_compress && j == dep) continue;
                    a.D[((size_t) sp * nD + oi) * nD + oj] = Dd[i * nc + j] - (ksi[j] / ksi[dep]) * Dd[i * nc + dep];
                    oj++;
                }
                oi++;
            }
        }
        if (want_E && (props & KG_PROP_THERMAL_DIFFUSION) && (a.DT || a.kT || a.soret)) {
            // cpp/KineticGas_properties.cpp:208-258
            double bm[MAXC * MAXC], DT[MAXC], DT2[MAXC];
            for (int j = 0; j < nc; j++) for (int k = 0; k < nc; k++) {
                double v = (j == k) ? 1.0 : 0.0;
                if (!tm.is_idealgas) {
                    const double etl3 = a.tlpow[((size_t) s.t * tm.npairs_dev + tm.pair_of[j][k]) * 4];
                    v += (4. * PI_REF / 3) * s.rho * s.x[k] * etl3 * (tm.m[j] / (tm.m[j] + tm.m[k])) * g[j * nc + k];
                }
                bm[j * nc + k] = v;
            }
            for (int i = 0; i < nc; i++) {
                double outer = 0.;
                for (int j = 0; j < nc; j++) {
                    double inner = 0.;
                    for (int k = 0; k < nc; k++) inner += bm[j * nc + k];
                    outer += dsol[j * MAXN + i] * s.x[j] * inner;
                }
                DT[i] = (s.x[i] / 2.) * (th[i] - outer);
            }
            for (int i = 0; i < nc; i++) { double v = 0.; for (int k = 0; k < nc; k++) v += psi[i * nc + k] * DT[k]; DT2[i] = v; }
            double tmp = 0.;
            for (int k = 0; k < nc; k++) for (int mm = 0; mm < nc; mm++) tmp += s.rho * s.x[mm] * bm[mm * nc + k];
            if (a.DT && !zar) for (int i = 0; i < nc; i++) a.DT[(size_t) sp * nc + i] = (DT2[i] + (Dd[i * nc + dep] / ksi[dep]) * tmp) / AVOGADRO;
            if ((a.DT || a.soret) && zar) {
                // D_T of the Zarate relation (:262-277): (-c X) D_T = (D_T,indep^(CoN) + c D^(x) x_indep) / T, c = 1 / Vm (nc - 1 entries)
                const double cm = 1.0 / s.Vm;
                double rhs[MAXC], XL[MAXC * MAXC]; int px[MAXC];
                for (int i = 0; i < nm; i++) {
                    double v = (DT2[idm[i]] + (Dd[idm[i] * nc + dep] / ksi[dep]) * tmp) / AVOGADRO;
                    for (int j = 0; j < nm; j++) v += cm * (Dd[idm[i] * nc + idm[j]] - (ksi[idm[j]] / ksi[dep]) * Dd[idm[i] * nc + dep]) * s.x[idm[j]];
                    rhs[i] = v / s.T;
                    for (int j = 0; j < nm; j++) XL[i * MAXC + j] = -cm * Xm[i * MAXC + j];
                }
                if (lu_factor<1>(XL, nm, MAXC, px)) st |= KG_STATUS_SINGULAR;
                lu_solve<1>(XL, nm, MAXC, px, rhs);
                if (a.DT) for (int i = 0; i < nm; i++) a.DT[(size_t) sp * nc + i] = rhs[i];
                if (a.soret) {
                    // KineticGas::soret_coefficient (:325-330): S_T = D^-1 D_T, both in the Zarate frame
                    double DL[MAXC * MAXC]; int pd[MAXC];
                    for (int i = 0; i < nm; i++) for (int j = 0; j < nm; j++) DL[i * MAXC + j] = Dz[i * MAXC + j];
                    if (lu_factor<1>(DL, nm, MAXC, pd)) st |= KG_STATUS_SINGULAR;
                    lu_solve<1>(DL, nm, MAXC, pd, rhs);
                    for (int i = 0; i < nm; i++) a.soret[(size_t) sp * nc + i] = rhs[i];
                }
            }
            if (a.kT) {
                // KineticGas::thermal_diffusion_ratio (cpp/KineticGas_properties.cpp:282-312): always from the centre-of-mass
                // frame quantities with the last component dependent, whatever frame the D / D_T outputs use.
                const int last = nc - 1;
                double Am[MAXC * MAXC], rhs[MAXC]; int pv[MAXC];
                for (int i = 0; i < last; i++) {
                    for (int j = 0; j < nc; j++) {
                        const double Dij = Dcom[i * nc + j] - (ksi[j] / ksi[last]) * Dcom[i * nc + last];
                        Am[i * MAXC + j] = -Dij * s.x[j] * (1 / s.Vm);
                    }
                    Am[last * MAXC + i] = s.x[i] * ksi[i];
                    rhs[i] = (DT[i] + (Dcom[i * nc + last] / ksi[last]) * tmp) / AVOGADRO;
                }
                Am[last * MAXC + last] = s.x[last] * ksi[last];
                if (tm.is_idealgas) rhs[last] = 1;
                else {
                    double v = 0.;
                    for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) v += s.x[i] * bm[i * nc + j];     // bm = delta_ij + (4 pi / 3) rho x_j etl^3 M_ij g_ij
                    rhs[last] = v;
                }
                if (lu_factor<1>(Am, nc, MAXC, pv)) st |= KG_STATUS_SINGULAR;
                lu_solve<1>(Am, nc, MAXC, pv, rhs);
                for (int i = 0; i < nc; i++) a.kT[(size_t) sp * nc + i] = rhs[i];
            }
        }
        if ((props & KG_PROP_CONDUCTIVITY) && a.cp_ideal) {
            // thermal_conductivity_contributions "tdi", cpp/KineticGas_properties.cpp:74-141
            double dth[MAXC];
            for (int i = 0; i < nc; i++) dth[i] = 0.0;
            if (!single && N > 1) {
                // compute_dth_vector (:432-447): (d[0] / d000) dth = l / d000
                double dm[MAXC * MAXC]; int pv[MAXC];
                const double d000 = dsol[0];
                for (int i = 0; i < nc; i++) { for (int j = 0; j < nc; j++) dm[i * MAXC + j] = dsol[j * MAXN + i] / d000; dth[i] = th[i] / d000; }
                // d[0] has rank nc - 1 (the diffusive expansion coefficients of the nc driving forces are linearly dependent), so
                // the reference's partialPivLu ends on a pivot that is pure rounding noise and lambda' does not depend on it
                // (d[1] annihilates the null vector). When that noise happens to be exactly 0.0 the reference would return NaN;
                // here the pivot is replaced by one ulp of the matrix scale, which is what it is in every other case.
                lu_factor<1>(dm, nc, MAXC, pv, DBL_EPSILON);
                lu_solve<1>(dm, nc, MAXC, pv, dth);
            }
            double lam_t = 0.;
            for (int i = 0; i < nc; i++) {
                double tmp = 0.;
                if (!single && N > 1) for (int k = 0; k < nc; k++) tmp += dsol[k * MAXN + nc + i] * dth[k];      // d[1][i][k]
                lam_t += s.x[i] * K[i] * (th[nc + i] - tmp);
            }
            lam_t *= (5. * BOLTZMANN / 4.);
            double lam_d = 0.;
            if (!tm.is_idealgas) {
                for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) {
                    const double etl4 = a.tlpow[((size_t) s.t * tm.npairs_dev + tm.pair_of[i][j]) * 4 + 3];
                    lam_d += (s.rho * s.rho) * sqrt(2 * PI_REF * tm.m[i] * tm.m[j] * BOLTZMANN * s.T / (tm.m[i] + tm.m[j]))
                             * (s.x[i] * s.x[j]) / (tm.m[i] + tm.m[j]) * etl4 * g[i * nc + j];
                }
                lam_d *= (4. * BOLTZMANN / 3.);
            }
            double lam_i = 0., cpf = 0.;
            if (single) cpf = (a.cp_ideal[sp] - 5. * GAS_CONSTANT / 2.) / (tm.m[0] * AVOGADRO * 1e3);
            else for (int i = 0; i < nc; i++) cpf += s.x[i] * (a.cp_ideal[(size_t) sp * nc + i] - 5. * GAS_CONSTANT / 2.) / (tm.m[i] * AVOGADRO * 1e3);
            if (1.32e3 * cpf > 1e-8) {
                // viscosity(T, Vm = 1e12, x, N): rdf slot 1
                double g0[MAXC * MAXC], K0[MAXC], Kp0[MAXC];
                const double rho0 = AVOGADRO / 1e12;
                for (int k = 0; k < nc * nc; k++) g0[k] = tm.is_idealgas ? 1.0 : a.rdf[((size_t) sp * a.nslots + 1) * nc * nc + k];
                k_factors(a, tm, s, rho0, g0, K0, Kp0);
                const double eta0 = viscosity_at<ES>(a, tm, s, rho0, g0, Kp0, A, ld, piv, b, st, nullptr, nullptr);
                lam_i = 1.32e3 * eta0 * cpf;
            }
            if (a.contribs) { a.contribs[(size_t) sp * 3] = lam_t; a.contribs[(size_t) sp * 3 + 1] = lam_d; a.contribs[(size_t) sp * 3 + 2] = lam_i; }
            if (a.conductivity) a.conductivity[sp] = lam_t + lam_d + lam_i;
        }
    }
    if (props & KG_PROP_VISCOSITY) {
        const double eta = viscosity_at<ES>(a, tm, s, s.rho, g, Kp, A, ld, piv, b, st,
                                        a.visc_matrix ? a.visc_matrix + (size_t) sp * n * n : nullptr, a.visc_vector ? a.visc_vector + (size_t) sp * n : nullptr);
        if (a.viscosity) a.viscosity[sp] = eta;
    }
    if (a.status) a.status[sp] = st;
}

#undef AT
} // namespace

// ================================================================================================================
// host side
// ================================================================================================================
struct kg_tcache {
    kg_model* m = nullptr;
    int device = 0;
    kg_transport_opts opt{};
    TModel tm{};
    int nT = 0;
    bool mie = false, ljs = false, es = false;      // es: extended Sutherland RDF (power-law sums other than plain Mie)
    std::vector<double> T;
    std::map<double, int> index_of;
    std::vector<unsigned char> is_corr;              // [nT][npairs][nlr]
    double *d_T = nullptr, *d_omega = nullptr, *d_mtl = nullptr, *d_etl = nullptr, *d_dbh = nullptr, *d_brk = nullptr, *d_tlpow = nullptr;
    MiePair* d_mp_int = nullptr;                     // [nc][nc] internal pairs
    void* d_rdfT = nullptr;                          // RdfT [nT][nc][nc] (Mie RDF: T- and pair-only part)
    EsPair* d_es_int = nullptr;                      // [nc][nc] internal pairs (es)
    EsParams* d_es_prm = nullptr;                    // [nT][nc][nc] effective parameters (es)
    MiePair mp00{};                                  // pair (0,0) on the host (LJ/spline RDF kernel argument)
    std::vector<double> h_mtl, h_etl;
    double ms_omega = 0, ms_tl = 0;                  // stage timings of the build (host clock around synchronised stages)
};
void kg_register_child(void* child, kg_model* m, int kind);      // kg_omega.cu

namespace {

template<class T> cudaError_t upload(kg_model* m, T** dptr, const std::vector<T>& v) {
    cudaError_t e = kg_dmalloc(m, dptr, std::max<size_t>(1, v.size()) * sizeof(T));
    if (e != cudaSuccess) return e;
    return cudaMemcpy(*dptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
}

bool desc_is_mie(const kg_potential_desc& d) {
    return d.kind == KG_POT_POWER_SUM && d.nterms == 2 && d.C[0] > 0 && d.C[1] == -d.C[0] && d.lam[0] > d.lam[1];
}

} // namespace

static void tcache_free(kg_tcache* c) {
    cudaSetDevice(c->device);
    void* ptrs[] = {c->d_T, c->d_omega, c->d_mtl, c->d_etl, c->d_dbh, c->d_brk, c->d_tlpow, c->d_mp_int, c->d_es_int, c->d_es_prm, c->d_rdfT};
    for (void* p : ptrs) c->m->dfree(p);
    delete c;
}
extern "C" void kg_tcache_destroy(kg_tcache* c) {
    KG_TRACE();
    if (!c || !KgRegistry::get().remove(c)) return;       // unknown handle: already destroyed together with its model
    tcache_free(c);
}

extern "C" int kg_tcache_lr(const kg_tcache* c, int32_t* l, int32_t* r) {
    KG_TRACE();
    if (!c) return -1;
    for (int k = 0; k < c->tm.nlr; k++) { if (l) l[k] = c->tm.lr_l[k]; if (r) r[k] = c->tm.lr_r[k]; }
    return c->tm.nlr;
}

extern "C" int kg_tcache_fetch(kg_tcache* c, double* omega, double* mtl, double* etl, uint8_t* is_correlation) {
    KG_TRACE();
    if (!c) return -1;
    kg_model* m = c->m;
    auto set_error = [&](const std::string& s) { m->err = s; };
    CUDA_OK(cudaSetDevice(m->device));
    const size_t no = (size_t) c->nT * m->npairs * c->tm.nlr;
    if (omega) CUDA_OK(cudaMemcpy(omega, c->d_omega, no * sizeof(double), cudaMemcpyDeviceToHost));
    const size_t nt = (size_t) c->nT * m->npairs;
    if (mtl) for (size_t k = 0; k < nt; k++) mtl[k] = c->h_mtl.empty() ? 0.0 : c->h_mtl[k];
    if (etl) for (size_t k = 0; k < nt; k++) etl[k] = c->h_etl.empty() ? 0.0 : c->h_etl[k];
    if (is_correlation) memcpy(is_correlation, c->is_corr.data(), no);
    return 0;
}

extern "C" int kg_tcache_create(kg_model* m, const kg_transport_opts* opts, int nT, const double* T, const double* coeff_T, kg_tcache** out) {
    KG_TRACE();
    if (!m) return -1;
    auto set_error = [&](const std::string& s) { m->err = s; };
    if (!opts || nT < 1 || !T || !out) { set_error("kg_tcache_create: invalid argument"); return -1; }
    if (m->coeff_T_required && !coeff_T) { set_error("this model's coefficients depend on temperature: coeff_T is required (kg_model_set_temperature_dependent)"); return -1; }
    const int N = opts->N;
    if (N < 1 || N > MAXO) { set_error("kg_tcache_create: Enskog order N out of range"); return -1; }
    const int nc = m->is_singlecomp ? 2 : m->ncomps;
    if (nc > MAXC) { set_error("kg_tcache_create: too many components"); return -1; }
    if (m->is_singlecomp && m->ncomps != 1) { set_error("kg_tcache_create: single-component models are created with ncomps = 1"); return -1; }
    bool mie = true;
    for (auto& d : m->descs) mie = mie && desc_is_mie(d);
    bool ljs = m->generic;      // LJSpline: own correlation (cpp/LJSpline.cpp:123-134) and own RDF model (cpp/LJSpline.h:89-125)
    for (auto& g : m->gdescs) ljs = ljs && g.kind == KG_POT_LJ_SPLINE;
    // power-law sums that are not MieKinGas (ExtSutherland / QuantumMie): SAFT-VR Mie style RDF of cpp/ExtendedSutherland.cpp:315-333
    const bool es = !m->generic && (!mie || coeff_T);
    if (!opts->is_idealgas && m->generic && !ljs) {
        set_error("RDF model is not implemented. Model is only available for ideal gas!");     // cpp/KineticGas.h:259
        return -1;
    }
    if (!opts->is_idealgas && es) for (auto& d : m->descs) for (int k = 0; k < d.nterms; k++)
        if (d.lam[k] <= 4.0) { set_error("kg_tcache_create: the extended Sutherland RDF needs exponents > 4"); return -1; }
    CUDA_OK(cudaSetDevice(m->device));
    kg_tcache* c = new kg_tcache;
    c->m = m; c->device = m->device; c->opt = *opts; c->nT = nT; c->mie = mie; c->ljs = ljs; c->es = es;
    c->T.assign(T, T + nT);
    for (int t = 0; t < nT; t++) c->index_of.emplace(T[t], t);
    TModel& tm = c->tm;
    tm.nc = nc; tm.npairs_dev = m->npairs; tm.is_singlecomp = m->is_singlecomp; tm.is_idealgas = opts->is_idealgas; tm.N = N;
    for (int i = 0; i < nc; i++) tm.m[i] = m->mass[m->is_singlecomp ? 0 : i];
    for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) {
        tm.sqmm[i][j] = sqrt(tm.m[i] * tm.m[j]);
        tm.sqred[i][j] = sqrt(tm.m[i] * tm.m[j] / (tm.m[i] + tm.m[j]));
    }
    std::vector<int> pair_i, pair_j;
    for (int i = 0; i < m->ncomps; i++) for (int j = i; j < m->ncomps; j++) { pair_i.push_back(i); pair_j.push_back(j); }
    for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) {
        if (m->is_singlecomp) { tm.pair_of[i][j] = 0; continue; }
        const int a = std::min(i, j), b = std::max(i, j);
        tm.pair_of[i][j] = a * m->ncomps - a * (a - 1) / 2 + (b - a);
    }
    // (l,r) sets: conductivity / diffusion l = 1..N, r = l..2N-l (cpp/KineticGas_mthr.cpp:62-73); viscosity l = 1..N+1,
    // r = l..2N+2-l where some B' or B'' is non-zero (:102-125)
    const unsigned props = opts->properties;
    const bool want_cond = props & (KG_PROP_CONDUCTIVITY | KG_PROP_DIFFUSION | KG_PROP_THERMAL_DIFFUSION);
    const bool want_visc = props & (KG_PROP_VISCOSITY | KG_PROP_CONDUCTIVITY);
    auto Mof = [&](int i, int j) { return tm.m[i] / (tm.m[i] + tm.m[j]); };
    tm.nlr = 0;
    auto lr_index = [&](int l, int r) { for (int k = 0; k < tm.nlr; k++) if (tm.lr_l[k] == l && tm.lr_r[k] == r) return k; return -1; };
    for (int l = 1; l <= N + 1; l++) for (int r = l; r <= 2 * N + 2 - l; r++) {
        bool use = want_cond && l <= N && r <= 2 * N - l;
        if (!use && want_visc) {
            for (int i = 0; i < nc && !use; i++) for (int j = 0; j < nc && !use; j++) for (int p = 0; p < N && !use; p++) for (int q = 0; q < N && !use; q++)
                use = weight_B_dblprime(p, q, r, l, Mof(i, j), Mof(j, i)) != 0 || weight_B_prime(p, q, r, l, Mof(i, j), Mof(j, i)) != 0;
        }
        if (use) {
            if (tm.nlr >= MAXLR) { set_error("kg_tcache_create: (l,r) table overflow"); delete c; return -1; }
            tm.lr_l[tm.nlr] = l; tm.lr_r[tm.nlr] = r; tm.nlr++;
        }
    }
    const int nlr = tm.nlr, Np = m->npairs;
    // weight tables
    std::vector<double> wA((size_t) N * N * nlr, 0.0), wAp((size_t) nc * nc * N * N * nlr, 0.0), wBp(wAp.size(), 0.0), wBpp(wAp.size(), 0.0),
                        hpre((size_t) nc * nc * N * N, 0.0);
    std::vector<unsigned char> pair_needs((size_t) Np * nlr, 0);
    for (int p = 0; p < N; p++) for (int q = 0; q < N; q++) {
        if (want_cond) for (int l = 1; l <= std::min(p, q) + 1; l++) for (int r = l; r <= p + q + 2 - l; r++) {
            const int k = lr_index(l, r);
            wA[(size_t) (p * N + q) * nlr + k] = weight_A(p, q, r, l);
        }
        for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) {
            const double M1 = Mof(i, j), M2 = Mof(j, i);
            const size_t wo = ((size_t) ((i * nc + j) * N + p) * N + q) * nlr;
            hpre[(size_t) ((i * nc + j) * N + p) * N + q] = 8 * pow(M2, p + 0.5) * pow(M1, q + 0.5);
            const int pr = tm.pair_of[i][j];
            if (want_cond) for (int l = 1; l <= std::min(p, q) + 1; l++) for (int r = l; r <= p + q + 2 - l; r++) {
                const int k = lr_index(l, r);
                wAp[wo + k] = weight_A_prime(p, q, r, l, M1, M2);
                pair_needs[(size_t) pr * nlr + k] = 1;       // the reference evaluates omega for every (l,r) of the H sums
            }
            if (want_visc) for (int l = 1; l <= std::min(p, q) + 2; l++) for (int r = l; r <= p + q + 4 - l; r++) {
                const int k = lr_index(l, r);
                if (k < 0) continue;                          // all weights zero for every pair
                wBpp[wo + k] = weight_B_dblprime(p, q, r, l, M1, M2);
                wBp[wo + k] = weight_B_prime(p, q, r, l, M1, M2);
                if (wBpp[wo + k] != 0 || wBp[wo + k] != 0) pair_needs[(size_t) pr * nlr + k] = 1;
            }
        }
    }
    // where MieKinGas::omega takes the correlation branch (cpp/MieKinGas.cpp:81-93)
    c->is_corr.assign((size_t) nT * Np * nlr, 0);
    std::vector<unsigned char> need((size_t) nT * Np * nlr, 0);
    bool any_numeric = false, any_corr = false;
    for (int t = 0; t < nT; t++) for (int p = 0; p < Np; p++) for (int k = 0; k < nlr; k++) {
        if (!pair_needs[(size_t) p * nlr + k]) continue;
        const size_t o = ((size_t) t * Np + p) * nlr + k;
        const auto& d = m->descs[p];
        const int l = tm.lr_l[k], r = tm.lr_r[k];
        bool corr = opts->omega_correlation_active && mie && !coeff_T && (l <= 2) && (r >= l) && (r <= 3) && (fabs(d.lam[1] - 6.) < 1e-10);
        if (ljs) corr = opts->omega_correlation_active && (l <= 2) && (r >= l) && (r <= 3);
        if (corr) corr = (T[t] * 1.38064852e-23 / d.eps) > 0.4;
        c->is_corr[o] = corr; need[o] = !corr;
        any_numeric |= !corr; any_corr |= corr;
    }
    int rc = 0;
#define TC_CUDA(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { set_error(std::string(#expr) + ": " + cudaGetErrorString(e_)); tcache_free(c); return -2; } } while (0)
    TC_CUDA(upload(m, &c->d_T, c->T));
    TC_CUDA(kg_dmalloc(m, &c->d_omega, need.size() * sizeof(double)));
    cudaStream_t s = m->stream;
    TC_CUDA(cudaMemsetAsync(c->d_omega, 0, need.size() * sizeof(double), s));
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms_since = [&](std::chrono::steady_clock::time_point t0) { return std::chrono::duration<double, std::milli>(now() - t0).count(); };
    auto t_stage = now();
    m->timings[1] = m->timings[2] = m->timings[3] = m->timings[4] = 0;
    if (m->quantum_active) {
        // Quantum::omega (cpp/Quantum.cpp:1266-1268): every collision integral from partial-wave phase shifts
        std::vector<int32_t> ll(tm.lr_l, tm.lr_l + nlr), rr(tm.lr_r, tm.lr_r + nlr);
        std::vector<double> hom((size_t) nT * Np * nlr, 0.0), tmp((size_t) nT * nlr);
        std::vector<int32_t> qst(nT);
        for (int p = 0; p < Np; p++) {
            rc = kg_quantum_omega(m, p, m->q_wsym[p], m->q_wanti[p], nT, T, nlr, ll.data(), rr.data(), tmp.data(), qst.data());
            if (rc) { tcache_free(c); return rc; }
            for (int t = 0; t < nT; t++) for (int k = 0; k < nlr; k++) hom[((size_t) t * Np + p) * nlr + k] = tmp[(size_t) t * nlr + k];
        }
        TC_CUDA(cudaMemcpyAsync(c->d_omega, hom.data(), hom.size() * sizeof(double), cudaMemcpyHostToDevice, s));
        TC_CUDA(cudaStreamSynchronize(s));
        std::fill(c->is_corr.begin(), c->is_corr.end(), 0);
        any_numeric = any_corr = false;
    }
    if (any_numeric) {
        kg_omega_plan* plan = nullptr;
        std::vector<int32_t> ll(tm.lr_l, tm.lr_l + nlr), rr(tm.lr_r, tm.lr_r + nlr);
        rc = kg_omega_plan_create_masked(m, nT, T, nlr, ll.data(), rr.data(), coeff_T, need.data(), &plan);
        if (!rc) rc = kg_omega_plan_run(plan);
        if (!rc) rc = kg_omega_plan_export_dev(plan, c->d_omega, nullptr);
        std::vector<int32_t> pst((size_t) nT * Np, 0);
        if (!rc) rc = kg_omega_plan_fetch(plan, nullptr, nullptr, pst.data());
        if (!rc) { double cnt[16]; if (!kg_omega_plan_counters(plan, cnt)) m->timings[4] += cnt[7]; }
        if (plan) kg_omega_plan_destroy(plan);
        if (rc) { tcache_free(c); return rc; }
        int bad = 0; for (int v : pst) bad |= v;
        if (bad & KG_STATUS_GETR_MAXITER) { set_error("get_R reached max iter!"); tcache_free(c); return -4; }
    }
    m->timings[1] = ms_since(t_stage); t_stage = now();
    std::vector<MiePair> mp_dev(Np), mp_int((size_t) nc * nc), mp_diag(nc);
    if (mie || ljs) {
        for (int p = 0; p < Np; p++) {
            const auto& d = m->descs[p];
            MiePair q; q.sigma = d.sigma; q.eps = d.eps; q.lr = d.lam[0]; q.la = d.lam[1]; q.C = d.C[0];
            if (ljs) { q.lr = 12.; q.la = 6.; q.C = 4.; }     // inside [sigma/2, sigma] (the Barker-Henderson integral) LJ/s is plain LJ
            q.alpha = q.C * ((1.0 / (q.la - 3.0)) - (1.0 / (q.lr - 3.0)));          // cpp/MieKinGas.cpp:52
            mp_dev[p] = q;
        }
        for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) mp_int[i * nc + j] = mp_dev[tm.pair_of[i][j]];
        for (int i = 0; i < nc; i++) mp_diag[i] = mp_int[i * nc + i];
    }
    TC_CUDA(upload(m, &c->d_mp_int, mp_int));
    c->mp00 = mp_int[0];
    if (any_corr) {
        MiePair* d_mp = nullptr; int *d_pi = nullptr, *d_pj = nullptr, *d_l = nullptr, *d_r = nullptr; double* d_mass = nullptr; unsigned char* d_ic = nullptr;
        std::vector<int> ll(tm.lr_l, tm.lr_l + nlr), rr(tm.lr_r, tm.lr_r + nlr);
        TC_CUDA(upload(m, &d_mp, mp_dev)); TC_CUDA(upload(m, &d_pi, pair_i)); TC_CUDA(upload(m, &d_pj, pair_j)); TC_CUDA(upload(m, &d_l, ll)); TC_CUDA(upload(m, &d_r, rr));
        TC_CUDA(upload(m, &d_mass, m->mass)); TC_CUDA(upload(m, &d_ic, c->is_corr));
        const int tot = nT * Np * nlr;
        omega_correlation_kernel<<<(tot + 127) / 128, 128, 0, s>>>(nT, Np, nlr, c->d_T, d_mp, d_pi, d_pj, d_mass, d_l, d_r, d_ic, ljs ? 1 : 0, c->d_omega);
        TC_CUDA(cudaGetLastError());
        TC_CUDA(cudaStreamSynchronize(s));
        for (void* q : {(void*) d_mp, (void*) d_pi, (void*) d_pj, (void*) d_l, (void*) d_r, (void*) d_mass, (void*) d_ic}) m->dfree(q);
    }
    if (!opts->is_idealgas) {
        c->h_mtl.resize((size_t) nT * Np); c->h_etl.resize((size_t) nT * Np);
        std::vector<int32_t> tst((size_t) nT * Np, 0);
        rc = kg_transfer_lengths_batch(m, nT, T, coeff_T, c->h_mtl.data(), c->h_etl.data(), tst.data());
        if (rc) { tcache_free(c); return rc; }
        int bad = 0; for (int v : tst) bad |= v;
        if (bad & KG_STATUS_BMID) { set_error("Unable to converge bmid!"); tcache_free(c); return -4; }
        if (bad & KG_STATUS_GETR_MAXITER) { set_error("get_R reached max iter!"); tcache_free(c); return -4; }
        m->timings[2] = ms_since(t_stage); m->timings[4] += m->tl_launches;
        TC_CUDA(upload(m, &c->d_mtl, c->h_mtl)); TC_CUDA(upload(m, &c->d_etl, c->h_etl));
        {
            std::vector<double> tp((size_t) nT * Np * 4);
            for (size_t k = 0; k < (size_t) nT * Np; k++) {
                tp[4 * k] = pow(c->h_etl[k], 3); tp[4 * k + 1] = pow(c->h_mtl[k], 3.); tp[4 * k + 2] = pow(c->h_mtl[k], 4); tp[4 * k + 3] = pow(c->h_etl[k], 4);
            }
            TC_CUDA(upload(m, &c->d_tlpow, tp));
        }
        if (es) {
            std::vector<EsPair> ep_dev(Np), ep_int((size_t) nc * nc);
            for (int p = 0; p < Np; p++) {
                const auto& d = m->descs[p];
                EsPair q{}; q.sigma = d.sigma; q.eps = d.eps; q.nterms = d.nterms;
                for (int k = 0; k < d.nterms; k++) q.lam[k] = d.lam[k];
                ep_dev[p] = q;
            }
            std::vector<int> pof((size_t) nc * nc);
            for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) { pof[i * nc + j] = tm.pair_of[i][j]; ep_int[i * nc + j] = ep_dev[tm.pair_of[i][j]]; }
            std::vector<double> cf((size_t) nT * Np * KG_MAX_TERMS);
            for (int t = 0; t < nT; t++) for (int p = 0; p < Np; p++) for (int k = 0; k < KG_MAX_TERMS; k++)
                cf[((size_t) t * Np + p) * KG_MAX_TERMS + k] = coeff_T ? coeff_T[((size_t) t * Np + p) * KG_MAX_TERMS + k] : m->descs[p].C[k];
            EsPair* d_ep = nullptr; double* d_cf = nullptr; int* d_pof = nullptr;
            TC_CUDA(upload(m, &d_ep, ep_dev)); TC_CUDA(upload(m, &d_cf, cf)); TC_CUDA(upload(m, &d_pof, pof));
            TC_CUDA(upload(m, &c->d_es_int, ep_int));
            TC_CUDA(kg_dmalloc(m, &c->d_es_prm, (size_t) nT * nc * nc * sizeof(EsParams)));
            const int npu = nc * (nc + 1) / 2;
            es_params_kernel<<<(nT * npu + 63) / 64, 64, 0, s>>>(nT, nc, c->d_T, d_ep, d_cf, d_pof, Np, c->d_es_prm);
            es_dbh_mix_kernel<<<(nT * nc * nc + 63) / 64, 64, 0, s>>>(nT, nc, c->d_es_prm);
            TC_CUDA(cudaGetLastError());
            TC_CUDA(cudaStreamSynchronize(s));
            for (void* q : {(void*) d_ep, (void*) d_cf, (void*) d_pof}) m->dfree(q);
        } else {
            MiePair* d_diag = nullptr;
            TC_CUDA(upload(m, &d_diag, mp_diag));
            TC_CUDA(kg_dmalloc(m, &c->d_dbh, (size_t) nT * nc * sizeof(double)));
            bh_kernel<<<(nT * nc + 127) / 128, 128, 0, s>>>(nT, nc, c->d_T, d_diag, c->d_dbh);
            TC_CUDA(cudaGetLastError());
            if (!ljs) {
                TC_CUDA(kg_dmalloc(m, (RdfT**) &c->d_rdfT, (size_t) nT * nc * nc * sizeof(RdfT)));
                rdf_T_kernel<<<(nT * nc * nc + 63) / 64, 64, 0, s>>>(nT, nc, c->d_T, c->d_dbh, c->d_mp_int, (RdfT*) c->d_rdfT);
                TC_CUDA(cudaGetLastError());
            }
            TC_CUDA(cudaStreamSynchronize(s));
            m->dfree(d_diag);
        }
    }
    {
        double *d_wA = nullptr, *d_wAp = nullptr, *d_wBp = nullptr, *d_wBpp = nullptr, *d_hpre = nullptr;
        TC_CUDA(upload(m, &d_wA, wA)); TC_CUDA(upload(m, &d_wAp, wAp)); TC_CUDA(upload(m, &d_wBp, wBp)); TC_CUDA(upload(m, &d_wBpp, wBpp)); TC_CUDA(upload(m, &d_hpre, hpre));
        const int per_t = nc * nc * N * N;
        TC_CUDA(kg_dmalloc(m, &c->d_brk, (size_t) nT * 4 * per_t * sizeof(double)));
        bracket_kernel<<<(nT * per_t + 127) / 128, 128, 0, s>>>(nT, tm, c->d_omega, d_wA, d_wAp, d_wBp, d_wBpp, d_hpre, c->d_brk);
        TC_CUDA(cudaGetLastError());
        TC_CUDA(cudaStreamSynchronize(s));
        for (void* q : {(void*) d_wA, (void*) d_wAp, (void*) d_wBp, (void*) d_wBpp, (void*) d_hpre}) m->dfree(q);
    }
#undef TC_CUDA
    m->timings[3] = ms_since(t_stage) - m->timings[2];
    m->timings[4] += (any_corr ? 1 : 0) + (opts->is_idealgas ? 0 : 1) + 1;
    kg_register_child(c, m, 1);
    *out = c;
    return 0;
}

namespace {

struct DevBuf {      // scoped device allocations of one batch call
    kg_model* m;
    std::vector<void*> ptrs;
    explicit DevBuf(kg_model* model) : m(model) {}
    ~DevBuf() { for (void* p : ptrs) m->dfree(p); }
    template<class T> T* alloc(size_t n) { T* p = nullptr; if (kg_dmalloc(m, &p, std::max<size_t>(1, n) * sizeof(T)) != cudaSuccess) return nullptr; ptrs.push_back(p); return p; }
    template<class T> T* put(const T* h, size_t n, cudaStream_t s) {
        T* p = alloc<T>(n);
        if (p && h && cudaMemcpyAsync(p, h, n * sizeof(T), cudaMemcpyHostToDevice, s) != cudaSuccess) return nullptr;
        return p;
    }
    // source in host OR device memory (unified addressing decides): the EoS inputs may come straight from kg_eos_tv_batch_dev
    template<class T> T* put_any(const T* src, size_t n, cudaStream_t s) {
        T* p = alloc<T>(n);
        if (p && src && cudaMemcpyAsync(p, src, n * sizeof(T), cudaMemcpyDefault, s) != cudaSuccess) return nullptr;
        return p;
    }
};

int launch_state(kg_tcache* c, StateArgs& a, cudaStream_t s, const double* d_x_for_rdf, double* d_rdf) {
    const TModel& tm = c->tm;
    const int n = a.n;
    if (!tm.is_idealgas) {
        const long long tot = (long long) n * a.nslots * (tm.nc * (tm.nc + 1) / 2);
        if (c->ljs) ljs_rdf_kernel<<<(unsigned) (((long long) n * a.nslots + 127) / 128), 128, 0, s>>>(n, a.nslots, tm.nc, a.tidx, a.T_of, a.Vm, c->d_dbh, c->mp00, d_rdf);
        else if (c->es) es_rdf_kernel<<<(unsigned) ((tot + 63) / 64), 64, 0, s>>>(n, a.nslots, tm, a.tidx, a.T_of, a.Vm, d_x_for_rdf, c->d_es_int, c->d_es_prm, d_rdf);
        else rdf_kernel<<<(unsigned) ((tot + 127) / 128), 128, 0, s>>>(n, a.nslots, tm, a.tidx, a.Vm, d_x_for_rdf, (const RdfT*) c->d_rdfT, d_rdf);
    }
    const int nn = tm.N * tm.nc;
    // the Sonine system of a state point lives in shared memory (interleaved over the block's threads), sized by the tightest
    // leading dimension that holds N nc
#define LAUNCH_STATE(MX) do { constexpr int TPB = state_tpb(MX); const size_t smem = (size_t) MX * MX * TPB * sizeof(double); \
        cudaFuncSetAttribute(state_kernel<MX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem); \
        state_kernel<MX><<<(n + TPB - 1) / TPB, TPB, smem, s>>>(tm, a); } while (0)
    if (nn <= 6) LAUNCH_STATE(6);
    else if (nn <= 8) LAUNCH_STATE(8);
    else if (nn <= 10) LAUNCH_STATE(10);
    else if (nn <= 12) LAUNCH_STATE(12);
    else if (nn <= 16) LAUNCH_STATE(16);
    else LAUNCH_STATE(24);
#undef LAUNCH_STATE
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { c->m->err = std::string("state kernels: ") + cudaGetErrorString(e); return -2; }
    return 0;
}

} // namespace

extern "C" int kg_transport_batch(kg_model* m, kg_tcache* cache, int n, const double* T, const double* Vm, const double* x,
                                  const kg_transport_opts* opts, const kg_eos_inputs* eos, const kg_transport_out* out, int32_t* status) {
    KG_TRACE();
    if (!m) return -1;
    const auto t_enter = std::chrono::steady_clock::now();
    auto set_error = [&](const std::string& s) { m->err = s; };
    if (n < 0 || !T || !Vm || !opts || !out) { set_error("kg_transport_batch: invalid argument"); return -1; }
    if (n == 0) return 0;
    if (!m->is_singlecomp && !x) { set_error("kg_transport_batch: mole fractions required"); return -1; }
    const unsigned props = opts->properties;
    const int nc = m->is_singlecomp ? 2 : m->ncomps, nce = m->is_singlecomp ? 1 : m->ncomps;
    const bool want_D = ((props & KG_PROP_DIFFUSION) && out->D) || ((props & KG_PROP_THERMAL_DIFFUSION) && (out->DT || out->thermal_diffusion_ratio || out->soret_zarate));
    const bool want_lambda = (props & KG_PROP_CONDUCTIVITY) && (out->conductivity || out->conductivity_contribs);
    if (want_D && (!eos || !eos->dmudn)) { set_error("kg_transport_batch: EoS is not set (dmudn required)"); return -1; }
    if (want_lambda && (!eos || !eos->cp_ideal)) { set_error("kg_transport_batch: EoS is not set (cp_ideal required)"); return -1; }
    if ((props & KG_PROP_CONDUCTIVITY) && opts->N < 2) { set_error("kg_transport_batch: thermal conductivity needs N >= 2"); return -1; }
    if (want_D && opts->frame_of_reference == KG_FOR_COV && (!eos || !eos->dvdn)) {
        set_error("kg_transport_batch: EoS is not set (dvdn required for the CoV frame)"); return -1; }
    if (opts->frame_of_reference < KG_FOR_COM || opts->frame_of_reference > KG_FOR_ZARATE_W) { set_error("Unknown (not-implemented) Frame of Reference."); return -1; }
    const bool zarate = opts->frame_of_reference >= KG_FOR_ZARATE;
    if (zarate && want_D && !opts->is_idealgas) {
        set_error((props & KG_PROP_THERMAL_DIFFUSION) && !(props & KG_PROP_DIFFUSION) ? "Zarate relation for thermal diffusion only implemented for ideal gas!"
                                                                                   : "Zarate relation for diffusion only implemented for ideal gas!");     // cpp/KineticGas_properties.cpp:180,263
        return -1;
    }
    if ((props & KG_PROP_THERMAL_DIFFUSION) && (out->DT || out->soret_zarate) && opts->frame_of_reference > KG_FOR_ZARATE) {
        set_error("Unknown (not-implemented) Frame of Reference."); return -1; }       // CoM_to_FoR_matr has no zarate_x / zarate_w (:348-396)
    if (out->soret_zarate && opts->frame_of_reference != KG_FOR_ZARATE) { set_error("kg_transport_batch: soret_zarate needs frame = zarate"); return -1; }
    if (opts->frame_of_reference == KG_FOR_SOLVENT && (opts->solvent_idx < 0 || opts->solvent_idx >= nc)) { set_error("kg_transport_batch: bad solvent index"); return -1; }
    if (opts->dependent_idx >= nc || opts->dependent_idx < -nc) { set_error("kg_transport_batch: bad dependent index"); return -1; }
    CUDA_OK(cudaSetDevice(m->device));
    kg_tcache* c = cache;
    bool own = false;
    if (!c) {
        std::vector<double> Tu(T, T + n);
        std::sort(Tu.begin(), Tu.end());
        Tu.erase(std::unique(Tu.begin(), Tu.end()), Tu.end());
        const int rc = kg_tcache_create(m, opts, (int) Tu.size(), Tu.data(), nullptr, &c);
        if (rc) return rc;
        own = true;
    }
    else if (c->m != m || c->opt.N != opts->N || c->opt.is_idealgas != opts->is_idealgas || (props & ~c->opt.properties)
             || c->opt.omega_correlation_active != opts->omega_correlation_active) {
        set_error("kg_transport_batch: cache was built for other options"); return -1;
    }
    struct Guard { kg_tcache* c; bool own; ~Guard() { if (own) kg_tcache_destroy(c); } } guard{c, own};
    cudaStream_t s = m->stream;
    std::vector<int> tidx(n);
    {   // state points usually come in runs of equal temperature: one map look-up per run
        double lastT = std::nan(""); int lastI = -1;
        for (int k = 0; k < n; k++) {
            if (T[k] != lastT) { auto f = c->index_of.find(T[k]); lastI = (f == c->index_of.end()) ? -1 : f->second; lastT = T[k]; }
            tidx[k] = lastI;
        }
    }
    DevBuf db(m);
    StateArgs a; memset(&a, 0, sizeof(a));
    a.n = n; a.nT = c->nT; a.opt = *opts; a.opt.properties = props;
    if (zarate) a.opt.do_compress = 1;            // the Zarate matrices are (nc - 1) x (nc - 1) by construction
    a.nslots = (props & KG_PROP_CONDUCTIVITY) ? 2 : 1;
    const int N = opts->N, nn = N * nc, nD = a.opt.do_compress ? nc - 1 : nc;
#define NEED(p) do { if (!(p)) { set_error("kg_transport_batch: device allocation or copy failed"); return -2; } } while (0)
    const bool chunked = n >= 262144;             // very large batches: the copies go chunk by chunk on the auxiliary streams (below)
    a.tidx = db.put(chunked ? nullptr : tidx.data(), n, s); NEED(a.tidx);
    a.T_of = c->d_T;
    a.Vm = db.put(chunked ? nullptr : Vm, n, s); NEED(a.Vm);
    a.x = m->is_singlecomp ? nullptr : db.put(chunked ? nullptr : x, (size_t) n * nc, s); if (!m->is_singlecomp) NEED(a.x);
    a.mtl = c->d_mtl; a.etl = c->d_etl; a.tlpow = c->d_tlpow; a.brk = c->d_brk;
    double* d_rdf = nullptr;
    if (!opts->is_idealgas) { d_rdf = db.alloc<double>((size_t) n * a.nslots * nc * nc); NEED(d_rdf); }
    a.rdf = d_rdf;
    if (eos && eos->dmudn && (props & (KG_PROP_DIFFUSION | KG_PROP_THERMAL_DIFFUSION))) { a.dmudn = db.put_any(chunked ? nullptr : eos->dmudn, (size_t) n * nce * nce, s); NEED(a.dmudn); }
    if (eos && eos->cp_ideal && (props & KG_PROP_CONDUCTIVITY)) { a.cp_ideal = db.put_any(chunked ? nullptr : eos->cp_ideal, (size_t) n * nce, s); NEED(a.cp_ideal); }
    if (eos && eos->dvdn && opts->frame_of_reference == KG_FOR_COV) { a.dvdn = db.put_any(chunked ? nullptr : eos->dvdn, (size_t) n * nce, s); NEED(a.dvdn); }
    struct OutMap { double* host; double** dev; size_t count; };
    OutMap maps[] = {{out->viscosity, &a.viscosity, (size_t) n}, {out->conductivity, &a.conductivity, (size_t) n},
                     {out->conductivity_contribs, &a.contribs, (size_t) n * 3}, {out->D, &a.D, (size_t) n * nD * nD}, {out->DT, &a.DT, (size_t) n * nc}, {out->thermal_diffusion_ratio, &a.kT, (size_t) n * nc}, {out->soret_zarate, &a.soret, (size_t) n * nc},
                     {out->K, &a.K, (size_t) n * nc}, {out->K_prime, &a.Kp, (size_t) n * nc}, {out->cond_matrix, &a.cond_matrix, (size_t) n * nn * nn},
                     {out->cond_vector, &a.cond_vector, (size_t) n * nn}, {out->visc_matrix, &a.visc_matrix, (size_t) n * nn * nn},
                     {out->visc_vector, &a.visc_vector, (size_t) n * nn}};
    for (auto& mp : maps) if (mp.host) { *mp.dev = db.alloc<double>(mp.count); NEED(*mp.dev); CUDA_OK(cudaMemsetAsync(*mp.dev, 0, mp.count * sizeof(double), s)); }
    a.status = db.alloc<int>(n); NEED(a.status);
#undef NEED
    // Chunked pipeline over two auxiliary streams: H2D of chunk k+1, the rdf + state kernels of chunk k and the D2H of chunk k-1
    // overlap (truly so when the caller's buffers are pinned, kg_host_alloc; pageable buffers are staged by the driver and only
    // the kernels overlap). Inputs were enqueued above on the model stream in one piece when the batch is small.
    std::vector<int> st(n);
    const int nchunk = chunked ? 4 : 1;
    float kms = 0;
    if (nchunk == 1) {
        cudaEvent_t ev0, ev1;
        CUDA_OK(cudaEventCreate(&ev0)); CUDA_OK(cudaEventCreate(&ev1));
        CUDA_OK(cudaEventRecord(ev0, s));
        if (launch_state(c, a, s, a.x, d_rdf)) { cudaEventDestroy(ev0); cudaEventDestroy(ev1); set_error("kg_transport_batch: kernel launch failed (" + m->err + ")"); return -2; }
        CUDA_OK(cudaEventRecord(ev1, s));
        for (auto& mp : maps) if (mp.host) CUDA_OK(cudaMemcpyAsync(mp.host, *mp.dev, mp.count * sizeof(double), cudaMemcpyDeviceToHost, s));
        if (out->rdf && !opts->is_idealgas)
            CUDA_OK(cudaMemcpy2DAsync(out->rdf, (size_t) nc * nc * sizeof(double), d_rdf, (size_t) a.nslots * nc * nc * sizeof(double),
                                      (size_t) nc * nc * sizeof(double), n, cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaMemcpyAsync(st.data(), a.status, n * sizeof(int), cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));
        cudaEventElapsedTime(&kms, ev0, ev1); cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    } else {
        if (!m->aux_stream[0]) for (int k = 0; k < 2; k++) {
            CUDA_OK(cudaStreamCreateWithFlags(&m->aux_stream[k], cudaStreamNonBlocking));
            CUDA_OK(cudaEventCreateWithFlags(&m->aux_done[k], cudaEventDisableTiming));
        }
        cudaEvent_t fork, kev[8];
        CUDA_OK(cudaEventCreateWithFlags(&fork, cudaEventDisableTiming));
        for (auto& e : kev) CUDA_OK(cudaEventCreate(&e));
        CUDA_OK(cudaEventRecord(fork, s));                      // allocations, memsets and the small tables are on the model stream
        for (int k = 0; k < 2; k++) CUDA_OK(cudaStreamWaitEvent(m->aux_stream[k], fork, 0));
        const int per = ((n + nchunk - 1) / nchunk + 127) & ~127;
        int rc_launch = 0;
        for (int k = 0; k < nchunk && !rc_launch; k++) {
            const int lo = k * per, cnt = std::min(per, n - lo);
            if (cnt <= 0) break;
            cudaStream_t cs = m->aux_stream[k & 1];
            StateArgs ac = a;
            ac.n = cnt;
            ac.tidx = a.tidx + lo; ac.Vm = a.Vm + lo; ac.status = a.status + lo;
            auto h2d = [&](const double* host, const double* dev, size_t per_pt) {
                return cudaMemcpyAsync(const_cast<double*>(dev) + (size_t) lo * per_pt, host + (size_t) lo * per_pt, (size_t) cnt * per_pt * sizeof(double), cudaMemcpyDefault, cs);
            };
            CUDA_OK(cudaMemcpyAsync(const_cast<int*>(ac.tidx), tidx.data() + lo, cnt * sizeof(int), cudaMemcpyHostToDevice, cs));
            CUDA_OK(h2d(Vm, a.Vm, 1));
            if (a.x) { CUDA_OK(h2d(x, a.x, nc)); ac.x = a.x + (size_t) lo * nc; }
            if (a.dmudn) { CUDA_OK(h2d(eos->dmudn, a.dmudn, (size_t) nce * nce)); ac.dmudn = a.dmudn + (size_t) lo * nce * nce; }
            if (a.cp_ideal) { CUDA_OK(h2d(eos->cp_ideal, a.cp_ideal, nce)); ac.cp_ideal = a.cp_ideal + (size_t) lo * nce; }
            if (a.dvdn) { CUDA_OK(h2d(eos->dvdn, a.dvdn, nce)); ac.dvdn = a.dvdn + (size_t) lo * nce; }
            double* rdf_c = d_rdf ? d_rdf + (size_t) lo * a.nslots * nc * nc : nullptr;
            ac.rdf = rdf_c;
            double** outs[] = {&ac.viscosity, &ac.conductivity, &ac.contribs, &ac.D, &ac.DT, &ac.kT, &ac.soret, &ac.K, &ac.Kp, &ac.cond_matrix, &ac.cond_vector,
                               &ac.visc_matrix, &ac.visc_vector};
            const size_t per_pt[] = {1, 1, 3, (size_t) nD * nD, (size_t) nc, (size_t) nc, (size_t) nc, (size_t) nc, (size_t) nc, (size_t) nn * nn, (size_t) nn,
                                     (size_t) nn * nn, (size_t) nn};
            for (int q = 0; q < 13; q++) if (*outs[q]) *outs[q] += (size_t) lo * per_pt[q];
            CUDA_OK(cudaEventRecord(kev[2 * k], cs));
            rc_launch = launch_state(c, ac, cs, ac.x, rdf_c);
            CUDA_OK(cudaEventRecord(kev[2 * k + 1], cs));
            int q = 0;
            for (auto& mp : maps) { if (mp.host) CUDA_OK(cudaMemcpyAsync(mp.host + (size_t) lo * per_pt[q], *outs[q], (size_t) cnt * per_pt[q] * sizeof(double), cudaMemcpyDeviceToHost, cs)); q++; }
            if (out->rdf && !opts->is_idealgas)
                CUDA_OK(cudaMemcpy2DAsync(out->rdf + (size_t) lo * nc * nc, (size_t) nc * nc * sizeof(double), rdf_c, (size_t) a.nslots * nc * nc * sizeof(double),
                                          (size_t) nc * nc * sizeof(double), cnt, cudaMemcpyDeviceToHost, cs));
            CUDA_OK(cudaMemcpyAsync(st.data() + lo, ac.status, cnt * sizeof(int), cudaMemcpyDeviceToHost, cs));
        }
        for (int k = 0; k < 2; k++) { CUDA_OK(cudaEventRecord(m->aux_done[k], m->aux_stream[k])); CUDA_OK(cudaStreamWaitEvent(s, m->aux_done[k], 0)); }
        CUDA_OK(cudaStreamSynchronize(s));
        for (int k = 0; k < nchunk; k++) { float t = 0; if (cudaEventElapsedTime(&t, kev[2 * k], kev[2 * k + 1]) == cudaSuccess) kms += t; }
        cudaGetLastError();
        cudaEventDestroy(fork);
        for (auto& e : kev) cudaEventDestroy(e);
        if (rc_launch) { set_error("kg_transport_batch: kernel launch failed (" + m->err + ")"); return -2; }
    }
    if (out->rdf && opts->is_idealgas) for (size_t k = 0; k < (size_t) n * nc * nc; k++) out->rdf[k] = 1.0;
    m->timings[0] = kms; m->timings[5] = (opts->is_idealgas ? 1 : 2) * nchunk;
    m->timings[6] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_enter).count();     // whole call on the host clock
    if (status) for (int k = 0; k < n; k++) status[k] = st[k];
    return 0;
}

extern "C" int kg_last_timings(const kg_model* m, double* out8) {
    KG_TRACE();
    if (!m || !out8) return -1;
    for (int k = 0; k < 8; k++) out8[k] = m->timings[k];
    return 0;
}
