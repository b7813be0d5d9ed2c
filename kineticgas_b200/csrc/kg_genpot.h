// kg_genpot.h -- piecewise "polynomial x exponential" pair potentials (host + device layout).
//
// Every non-power-law potential of the reference is, on each radial piece, a sum of terms P_t(r) exp(Q_t(r)) with Laurent
// polynomials P_t, Q_t in the potential's own length unit:
//   LJ/spline   cpp/LJSpline.h:45-82       LJ for r < rs, cubic spline for rs <= r < rc, 0 beyond
//   LJTS        cpp/LJTS.h:28-60           LJ - u(rc) for r <= rc, 0 beyond
//   mod. Tang-Toennies  cpp/multiparam.cpp:5-89    (short-range A~/r exp(-a~ r) below short_range_lim)
//   HFD-B2      cpp/multiparam.cpp:91-182  (damping switched off for x > D)
//   Patkowski   cpp/multiparam.cpp:184-279 (exponential extrapolation below Rc)
// The Feynman-Hibbs correction of order F (cpp/extentions.h:228-255),
//   u_FH = u + sum_{n=1..F} (D/kT)^n / n! [ u^(2n) + 2n u^(2n-1) / r ],
// keeps that form: d/dr (P e^Q) = (P' + P Q') e^Q, so the host folds it into an effective prefactor per temperature
// (kg_genpot.cu) and the device only ever evaluates  U = sum_t P_t e^{Q_t}  and  dU/dr = sum_t D_t e^{Q_t}.
#pragma once

namespace kg {

constexpr int GP_MAXC = 40;   // Laurent coefficients per prefactor (Tang-Toennies with FH order 2 needs 37)
constexpr int GP_MAXQ = 6;
constexpr int GP_MAXT = 4;    // terms per piece
constexpr int GP_MAXP = 3;    // radial pieces

struct GenTerm {
    int pmin, pn;             // P(r)  = sum_{k<pn} pc[k] r^(pmin+k)
    int dmin, dn;             // D(r)  = P' + P Q' = sum_{k<dn} dc[k] r^(dmin+k)
    int qmin, qn;             // Q(r)  = sum_{k<qn} qc[k] r^(qmin+k); qn = 0: no exponential
    int pad0, pad1;
    double pc[GP_MAXC];
    double dc[GP_MAXC];
    double qc[GP_MAXQ];
};

struct GenPiece {
    double r_hi;              // piece is valid for r < r_hi (internal length unit); the last piece may have r_hi = +inf
    int nterms, pad;
    GenTerm t[GP_MAXT];
};

struct GenPot {
    double len_scale;         // r_internal = (r / sigma) * len_scale
    double e_scale;           // internal energy unit in J
    int npieces;
    int zero_on_negative;     // LJSpline::theta_integrand override: radicand < 0 -> integrand 0 (cpp/LJSpline.h:37-42)
    GenPiece piece[GP_MAXP];  // beyond the last piece the potential is exactly 0
};

constexpr int GP_DOUBLES = (int) (sizeof(GenPot) / sizeof(double));
static_assert(sizeof(GenPot) % sizeof(double) == 0, "GenPot must be a whole number of doubles");

} // namespace kg
