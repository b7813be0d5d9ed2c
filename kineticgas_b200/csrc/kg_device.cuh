// kg_device.cuh -- device-side collision dynamics for sm_100a (FP64, register resident).
//
// What the reference computes (cpp/Spherical.cpp:259-465, cpp/Integration/Integration.cpp:417-510):
//   R     = outermost root of  u(r)/(kT g^2) + (b/r)^2 - 1                       (get_R: shrink x0.95, bisect, Newton)
//   I     = sum_k w_k (R/u_k^2) [ r^4/b^2 (1 - u(r)/(kT g^2)) - r^2 ]^(-1/2), r = R/u_k   (tanh-sinh, h = 1e-3, truncated)
//   theta = I(R,g,b) - I(R_lim,g,100 sigma) + PI/2 ;  chi = PI - 2 theta
// How it is laid out here: one lane per deflection angle. The tanh-sinh abscissae are problem independent, so the
// host tabulates, per pair, rows  { w_k/(2 u_k), 1/u_k^2, u_k^lam_0, .., u_k^lam_{NT-1} }  (same libm calls as the
// reference). With s = sigma/r = u (sigma/R), every power-law term factorises, (sigma/r)^lam = u^lam (sigma/R)^lam, so
//   term_k = (w_k/u_k) * rsqrt( X_k ),   X_k = (R/b)^2 (1/u_k^2) (1 - sum_t A_t u_k^lam_t) - 1,
//   A_t = eps C_t (sigma/R)^lam_t / (kT g^2)   (per lane, once)
// which is algebraically the reference integrand; no transcendental is left in the node loop. The sequential stop
// rule (add, then test the next term against the running sum, break on NaN/Inf) is kept node for node.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "kg_genpot.h"

namespace kg {

// cpp/global_params.h:4-7 -- PI is truncated in the reference and must be truncated here (parity)
__device__ constexpr double BOLTZMANN = 1.38064852e-23;
__device__ constexpr double PI_REF = 3.14159265359;

constexpr int KMAX = 3328;          // tanh-sinh rows tabulated (u_k == 1.0 in double from k ~ 3170)
constexpr int MAX_TERMS = 6;
constexpr double TS_TOL = 1e-5;     // tanh_sinh default tol, cpp/Integration/Integration.h

struct PairParams {                 // staged in shared memory per CTA
    double sigma, eps;
    int nterms, ntpad;              // ntpad in {2,4,6}: row width = 2 + ntpad doubles
    double C[MAX_TERMS];
    double lam[MAX_TERMS];
    const GenPot* gen;              // device pointer: non-power-law potential of this pair (ntpad == 1), else null
};

// Per-lane collision context in reduced units (lengths / sigma)
template<int NT>
struct Lane {
    double Bt[NT];     // eps C_t / (kT g^2)
    double lam[NT];
    double b;          // impact parameter / sigma
    int ipath;         // every exponent is a small integer: powers by repeated squaring instead of exp(lam ln s)
};

// NT == 1 is the tag of the generic (piecewise polynomial x exponential) potentials: node rows { w/2u, 1/u^2, 1/u }
template<>
struct Lane<1> {
    const GenPot* gp;
    double invE;       // 1 / (kT g^2)
    double b;
};

__device__ __forceinline__ double ipow_pos(double x, int n) {      // x^n, n >= 0, by squaring
    double r = 1.0;
    while (n) { if (n & 1) r *= x; x *= x; n >>= 1; }
    return r;
}

// sum_{k<n} c[k] r^(kmin+k) ; s = 1/r. Horner runs in the variable that is <= 1 so no partial sum overflows.
__device__ __forceinline__ double laurent(const double* __restrict__ c, int kmin, int n, double r, double s) {
    if (n <= 0) return 0.0;
    double acc; int e;
    if (r >= 1.0) {
        acc = c[0];
        for (int k = 1; k < n; k++) acc = fma(acc, s, c[k]);
        e = kmin + n - 1;
    } else {
        acc = c[n - 1];
        for (int k = n - 2; k >= 0; k--) acc = fma(acc, r, c[k]);
        e = kmin;
    }
    return acc * (e >= 0 ? ipow_pos(r, e) : ipow_pos(s, -e));
}

// u(r) and du/d(r/sigma) in J of a generic potential at r = ri sigma
static __device__ __noinline__ void gen_eval(const GenPot* __restrict__ gp, double ri, int want_d, double& U, double& dU) {
    const double r = ri * gp->len_scale;
    int p = 0;
    while (p < gp->npieces && !(r < gp->piece[p].r_hi)) p++;
    U = 0.0; dU = 0.0;
    if (p == gp->npieces) return;
    const GenPiece& pc = gp->piece[p];
    const double s = 1.0 / r;
    double u = 0.0, du = 0.0;
    for (int t = 0; t < pc.nterms; t++) {
        const GenTerm& tm = pc.t[t];
        const double E = tm.qn > 0 ? exp(laurent(tm.qc, tm.qmin, tm.qn, r, s)) : 1.0;
        u = fma(laurent(tm.pc, tm.pmin, tm.pn, r, s), E, u);
        if (want_d) du = fma(laurent(tm.dc, tm.dmin, tm.dn, r, s), E, du);
    }
    U = u * gp->e_scale;
    dU = du * (gp->e_scale * gp->len_scale);
}

// ---- the same evaluation for the serial marches of the quantum engine (kg_quantum.cuh), where a lane calls it millions of times --------
// x^(2^i) once per call; a power is then a product of the selected squares in the order ipow_pos multiplies them (bit-identical results)
struct PowLadder { double x[6]; };
__device__ __forceinline__ PowLadder make_ladder(double x) {
    PowLadder L;
    L.x[0] = x;
#pragma unroll
    for (int i = 1; i < 6; i++) L.x[i] = L.x[i - 1] * L.x[i - 1];
    return L;
}
__device__ __forceinline__ double ladder_pow(const PowLadder& L, int n) {      // 0 <= n < 64
    // branches, not selects: the exponent belongs to the term, so a warp takes them together (selects cost two FSEL per bit and
    // were 20 % of the cross-section kernel's instructions)
    double r = (n & 1) ? L.x[0] : 1.0;
#pragma unroll
    for (int i = 1; i < 6; i++) if ((n >> i) & 1) r *= L.x[i];
    return r;
}
__device__ __forceinline__ double laurent_scale(int e, double r, double s, const PowLadder& Lr, const PowLadder& Ls) {
    const int a = e >= 0 ? e : -e;
    if (a >= 64) return e >= 0 ? ipow_pos(r, a) : ipow_pos(s, a);
    if (e >= 0) return ladder_pow(Lr, a);
    return ladder_pow(Ls, a);
}

// gen_eval with the prefactor and derivative polynomials of a term run as two interleaved Horner chains, coefficients fetched four
// ahead, powers from the ladders: same operations in the same order per chain, so U and dU are bit-identical to gen_eval's.
// `gp` may point to a copy of the potential in shared memory.
static __device__ __noinline__ void gen_eval_march(const GenPot* __restrict__ gp, double ri, double& U, double& dU) {
    const double r = ri * gp->len_scale;
    int p = 0;
    while (p < gp->npieces && !(r < gp->piece[p].r_hi)) p++;
    U = 0.0; dU = 0.0;
    if (p == gp->npieces) return;
    const GenPiece& pc = gp->piece[p];
    const double s = 1.0 / r;
    const bool up = r >= 1.0;                   // Horner in s from the lowest power, else in r from the highest
    const double x = up ? s : r;
    const PowLadder Lr = make_ladder(r), Ls = make_ladder(s);
    double u = 0.0, du = 0.0;
    for (int t = 0; t < pc.nterms; t++) {
        const GenTerm& tm = pc.t[t];
        double E = 1.0;
        if (tm.qn > 0) {
            const int n = tm.qn;
            double acc = up ? tm.qc[0] : tm.qc[n - 1];
            for (int k = 1; k < n; k++) acc = fma(acc, x, tm.qc[up ? k : n - 1 - k]);
            E = exp(acc * laurent_scale(up ? tm.qmin + n - 1 : tm.qmin, r, s, Lr, Ls));
        }
        const int pn = tm.pn, dn = tm.dn;
        double ap = 0.0, ad = 0.0;
        if (pn > 0) ap = up ? tm.pc[0] : tm.pc[pn - 1];
        if (dn > 0) ad = up ? tm.dc[0] : tm.dc[dn - 1];
        const int nmax = pn > dn ? pn : dn;
        const int pb = up ? 0 : pn - 1, db = up ? 0 : dn - 1, sg = up ? 1 : -1;
        int k = 1;
        for (; k + 3 < nmax; k += 4) {
            double cp[4], cd[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                cp[q] = (k + q < pn) ? tm.pc[pb + sg * (k + q)] : 0.0;
                cd[q] = (k + q < dn) ? tm.dc[db + sg * (k + q)] : 0.0;
            }
#pragma unroll
            for (int q = 0; q < 4; q++) {
                if (k + q < pn) ap = fma(ap, x, cp[q]);
                if (k + q < dn) ad = fma(ad, x, cd[q]);
            }
        }
        for (; k < nmax; k++) {
            if (k < pn) ap = fma(ap, x, tm.pc[pb + sg * k]);
            if (k < dn) ad = fma(ad, x, tm.dc[db + sg * k]);
        }
        if (pn > 0) u = fma(ap * laurent_scale(up ? tm.pmin + pn - 1 : tm.pmin, r, s, Lr, Ls), E, u);
        if (dn > 0) du = fma(ad * laurent_scale(up ? tm.dmin + dn - 1 : tm.dmin, r, s, Lr, Ls), E, du);
    }
    U = u * gp->e_scale;
    dU = du * (gp->e_scale * gp->len_scale);
}

// Bt_t (sigma/r)^lam_t at r = sigma / inv: binary powers for integer exponents (the pair-uniform common case, LJ and most
// Mie fluids), else s^lam = exp(lam ln s) with one shared log
template<int NT>
__device__ __forceinline__ void lane_powers(const Lane<NT>& c, double inv, double (&p)[NT]) {
    if (c.ipath) {
#pragma unroll
        for (int t = 0; t < NT; t++) p[t] = c.Bt[t] * ipow_pos(inv, (int) c.lam[t]);
    } else {
        const double ls = log(inv);
#pragma unroll
        for (int t = 0; t < NT; t++) p[t] = c.Bt[t] * exp(c.lam[t] * ls);
    }
}

// u(r)/(kT g^2) and sigma u'(r)/(kT g^2) at r = ri sigma
template<int NT>
__device__ __forceinline__ void pot_eval(const Lane<NT>& c, double ri, double& U, double& dU) {
    const double inv = 1.0 / ri;
    double p[NT];
    lane_powers<NT>(c, inv, p);
    double u = 0.0, du = 0.0;
#pragma unroll
    for (int t = 0; t < NT; t++) {
        u += p[t];
        du -= c.lam[t] * p[t];
    }
    U = u;
    dU = du * inv;
}
template<>
__device__ __forceinline__ void pot_eval<1>(const Lane<1>& c, double ri, double& U, double& dU) {
    gen_eval(c.gp, ri, 1, U, dU);
    U *= c.invE; dU *= c.invE;
}

template<int NT>
__device__ __forceinline__ double rootfun(const Lane<NT>& c, double ri) {
    // cpp/Spherical.cpp:368
    double U, dU;
    pot_eval<NT>(c, ri, U, dU);
    const double q = c.b / ri;
    return U + q * q - 1.0;
}

template<int NT>
__device__ __forceinline__ void rootfun_d(const Lane<NT>& c, double ri, double& f, double& df) {
    // cpp/Spherical.cpp:368-369
    double U, dU;
    pot_eval<NT>(c, ri, U, dU);
    const double q = c.b / ri;
    f = U + q * q - 1.0;
    df = dU - (2.0 / c.b) * (q * q * q);
}

// cpp/Integration/Integration.cpp:456-471 (ftol = -1: only the step criterion is live)
template<int NT>
__device__ __noinline__ double newton_usafe(const Lane<NT>& c, double x0, double dtol, int& ierr) {
    int niter = 0;
    double prev_x, f, df;
    do {
        prev_x = x0;
        rootfun_d<NT>(c, x0, f, df);
        x0 -= f / df;
        if (niter++ > 50) { ierr = 1; return x0; }
    } while (fabs(x0 - prev_x) > dtol);
    if (isnan(x0)) { ierr = 2; return x0; }
    if (isinf(x0)) { ierr = 3; return x0; }
    return x0;
}

// cpp/Integration/Integration.cpp:494-510 (ftol = -1)
template<int NT>
__device__ __noinline__ void bracket_root_usafe(const Lane<NT>& c, double& x0, double& x1, double& f0, double& f1, double dtol) {
    while (fabs(x0 - x1) > dtol) {
        const double xm = 0.5 * (x0 + x1);
        const double fm = rootfun<NT>(c, xm);
        if (f0 * fm > 0) { x0 = xm; f0 = fm; }
        else { x1 = xm; f1 = fm; }
    }
    if (fabs(f0) > fabs(f1)) {
        double t = x0; x0 = x1; x1 = t;
        t = f0; f0 = f1; f1 = t;
    }
}

// cpp/Spherical.cpp:437-453 : head-on turning point, u(R0) = kT g^2. Reduced units; Bt already carries 1/(kT g^2),
// so f = U - 1 and the convergence test |f / (dfdr sigma)| is unchanged by the common scale.
template<int NT>
__device__ __noinline__ double get_R0(const Lane<NT>& c) {
    double r = 0.9, U, dU;
    pot_eval<NT>(c, r, U, dU);
    double f = U - 1.0;
    int guard = 0;
    while (f < 0 && guard++ < 2000) {
        r *= 0.8;
        pot_eval<NT>(c, r, U, dU);
        f = U - 1.0;
    }
    guard = 0;
    while (fabs(f / dU) > 1e-6 && guard++ < 10000) {
        r -= f / dU;
        pot_eval<NT>(c, r, U, dU);
        f = U - 1.0;
    }
    return r;
}

// The shrink loop of get_R (cpp/Spherical.cpp:381-384): r0 *= 0.95 until f(r0) >= 0. Each step scales every power-law term
// by the constant 0.95^-lam_t, so the loop multiplies instead of re-evaluating log / exp (up to ~30 steps for b ~ 5 sigma;
// r0 itself takes the reference's exact sequence of values, f0 differs in the last bits only and is used for its sign).
template<int NT>
__device__ __forceinline__ void shrink_until_nonneg(const Lane<NT>& c, double& r0, double& f0) {
    if (!(f0 < 0)) return;
    double p[NT], F[NT];
    lane_powers<NT>(c, 1.0 / r0, p);
#pragma unroll
    for (int t = 0; t < NT; t++) F[t] = c.ipath ? ipow_pos(1.0 / 0.95, (int) c.lam[t]) : exp(c.lam[t] * 0.05129329438755058);    // 0.95^-lam
    int guard = 0;
    while (f0 < 0 && guard++ < 4000) {
        r0 *= 0.95;
        double u = 0.0;
#pragma unroll
        for (int t = 0; t < NT; t++) { p[t] *= F[t]; u += p[t]; }
        const double q = c.b / r0;
        f0 = u + q * q - 1.0;
    }
}
template<>
__device__ __forceinline__ void shrink_until_nonneg<1>(const Lane<1>& c, double& r0, double& f0) {
    int guard = 0;
    while (f0 < 0 && guard++ < 4000) { r0 *= 0.95; f0 = rootfun<1>(c, r0); }
}

// cpp/Spherical.cpp:358-435 ; returns R / sigma. g enters through Bt; g < 1e-6 or b < 1e-5 -> get_R0.
template<int NT>
__device__ double get_R(const Lane<NT>& c, double g, int& status) {
    if ((c.b < 1e-5) || (g < 1e-6)) return get_R0<NT>(c);
    double r0, r1, bracket_dtol;
    const double newton_dtol = 1e-6;
    if (c.b > 1) { r0 = r1 = c.b; bracket_dtol = 0.1; }
    else { r0 = r1 = 1; bracket_dtol = 0.05; }
    double f0 = rootfun<NT>(c, r0);
    double f1 = f0;
    shrink_until_nonneg<NT>(c, r0, f0);
    if (r1 - r0 > bracket_dtol) bracket_root_usafe<NT>(c, r0, r1, f0, f1, bracket_dtol);
    if (fabs(f0) < 1e-8) return r0;

    int ierr = 0;
    double R = newton_usafe<NT>(c, r0, newton_dtol, ierr);
    if (ierr != 0) { ierr = 0; R = newton_usafe<NT>(c, r1, newton_dtol, ierr); }
    if (ierr != 0) { ierr = 0; R = newton_usafe<NT>(c, 0.5 * (r0 + r1), newton_dtol, ierr); }
    int niter = 0;
    while (ierr != 0) {
        const double prev_r0 = r0;
        while ((prev_r0 == r0) && (bracket_dtol > newton_dtol)) {
            if (fabs(r1 - r0) < bracket_dtol) bracket_dtol /= 2;
            bracket_root_usafe<NT>(c, r0, r1, f0, f1, bracket_dtol);
        }
        if (bracket_dtol < newton_dtol) { R = r0; break; }
        ierr = 0;
        R = newton_usafe<NT>(c, r0, newton_dtol, ierr);
        if (niter++ > 50) { status |= 1; return nan(""); }   // "get_R reached max iter!"
    }
    return R;
}

// ---- get_R for integer exponents, written for the set-up kernel (every lane of the grid is in here) -------------------------
// Same algorithm, same sequence of iterates as get_R above (cpp/Spherical.cpp:358-435); what changes is the arithmetic around it:
// the lane context stays in registers (nothing is called out of line on the common path), the powers of one evaluation share a
// squaring chain, 1/r and f/f' use a MUFU reciprocal seed with two Newton steps (1-2 ulp instead of correctly rounded, no range
// check), and the shrink loop carries (b/r0)^2 along instead of dividing. Iterates therefore differ from the reference's in the last
// bits only; the step criterion sees the same values except on a knife edge, where R moves by the convergence noise (< 1e-12).
// Everything past the first Newton attempt (its failure modes) goes to the general code.
__device__ __forceinline__ double rcp_fast(double x) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x, y, 1.0);
    y = fma(y, e, y);
    e = fma(-x, y, 1.0);
    return fma(y, e, y);
}

template<int NT>
struct LaneIP {
    double Bt[NT];
    int n[NT];
    int nbits;         // bits of the largest exponent
    bool lj;           // exponents (12, 6)
    double b, two_over_b;
};

template<int NT>
__device__ __forceinline__ void ip_powers(const LaneIP<NT>& c, double inv, double (&p)[NT]) {
    if constexpr (NT == 2) {
        if (c.lj) {                                            // 12-6 (Lennard-Jones and the default Mie): five multiplications, no exponent loop
            const double x2 = inv * inv, x6 = (x2 * x2) * x2;
            p[1] = c.Bt[1] * x6;
            p[0] = c.Bt[0] * (x6 * x6);
            return;
        }
    }
#pragma unroll
    for (int t = 0; t < NT; t++) p[t] = c.Bt[t];
    double x = inv;
    for (int bit = 0; bit < c.nbits; bit++) {
#pragma unroll
        for (int t = 0; t < NT; t++) if ((c.n[t] >> bit) & 1) p[t] *= x;
        x *= x;
    }
}
template<int NT>
__device__ __forceinline__ double ip_rootfun(const LaneIP<NT>& c, double ri) {
    const double inv = rcp_fast(ri);
    double p[NT];
    ip_powers<NT>(c, inv, p);
    double u = 0.0;
#pragma unroll
    for (int t = 0; t < NT; t++) u += p[t];
    const double q = c.b * inv;
    return fma(q, q, u) - 1.0;
}
template<int NT>
__device__ __forceinline__ void ip_rootfun_d(const LaneIP<NT>& c, double ri, double& f, double& df) {
    const double inv = rcp_fast(ri);
    double p[NT];
    ip_powers<NT>(c, inv, p);
    double u = 0.0, du = 0.0;
#pragma unroll
    for (int t = 0; t < NT; t++) { u += p[t]; du = fma(-(double) c.n[t], p[t], du); }
    const double q = c.b * inv, q2 = q * q;
    f = (u + q2) - 1.0;
    df = fma(du, inv, -c.two_over_b * (q2 * q));
}
template<int NT>
__device__ __forceinline__ double ip_newton(const LaneIP<NT>& c, double x0, double dtol, int& ierr) {
    int niter = 0;
    double prev_x, f, df;
    do {
        prev_x = x0;
        ip_rootfun_d<NT>(c, x0, f, df);
        x0 = fma(-f, rcp_fast(df), x0);
        if (niter++ > 50) { ierr = 1; return x0; }
    } while (fabs(x0 - prev_x) > dtol);
    if (isnan(x0)) { ierr = 2; return x0; }
    if (isinf(x0)) { ierr = 3; return x0; }
    return x0;
}

// the rare remainder of get_R: the first Newton attempt failed
template<int NT>
__device__ __noinline__ double get_R_after_failure(const Lane<NT>& c, double r0, double r1, double f0, double f1, double bracket_dtol, int& status) {
    const double newton_dtol = 1e-6;
    int ierr = 0;
    double R = newton_usafe<NT>(c, r1, newton_dtol, ierr);
    if (ierr != 0) { ierr = 0; R = newton_usafe<NT>(c, 0.5 * (r0 + r1), newton_dtol, ierr); }
    int niter = 0;
    while (ierr != 0) {
        const double prev_r0 = r0;
        while ((prev_r0 == r0) && (bracket_dtol > newton_dtol)) {
            if (fabs(r1 - r0) < bracket_dtol) bracket_dtol /= 2;
            bracket_root_usafe<NT>(c, r0, r1, f0, f1, bracket_dtol);
        }
        if (bracket_dtol < newton_dtol) { R = r0; break; }
        ierr = 0;
        R = newton_usafe<NT>(c, r0, newton_dtol, ierr);
        if (niter++ > 50) { status |= 1; return nan(""); }
    }
    return R;
}

template<int NT>
__device__ __forceinline__ double get_R_ip(const Lane<NT>& lane, double g, int& status) {
    if ((lane.b < 1e-5) || (g < 1e-6)) return get_R0<NT>(lane);
    LaneIP<NT> c;
    int nmax = 1;
#pragma unroll
    for (int t = 0; t < NT; t++) { c.Bt[t] = lane.Bt[t]; c.n[t] = (int) lane.lam[t]; nmax = max(nmax, c.n[t]); }
    c.nbits = 32 - __clz(nmax);
    c.lj = (NT == 2) && (c.n[0] == 12) && (c.n[NT - 1] == 6);
    c.b = lane.b; c.two_over_b = 2.0 * rcp_fast(lane.b);
    double r0, r1, bracket_dtol;
    if (c.b > 1) { r0 = r1 = c.b; bracket_dtol = 0.1; }
    else { r0 = r1 = 1; bracket_dtol = 0.05; }
    double f0 = ip_rootfun<NT>(c, r0);
    double f1 = f0;
    if (f0 < 0) {                                              // r0 *= 0.95 until f(r0) >= 0 (cpp/Spherical.cpp:381-384)
        double p[NT], F[NT];
        const double inv0 = rcp_fast(r0);
        ip_powers<NT>(c, inv0, p);
        {   // F[t] = 0.95^-n[t] on the same chain
            LaneIP<NT> one = c;
#pragma unroll
            for (int t = 0; t < NT; t++) one.Bt[t] = 1.0;
            ip_powers<NT>(one, 1.0 / 0.95, F);
        }
        double q2 = (c.b * inv0) * (c.b * inv0);
        const double k2 = (1.0 / 0.95) * (1.0 / 0.95);
        int guard = 0;
        while (f0 < 0 && guard++ < 4000) {
            r0 *= 0.95;
            q2 *= k2;
            double u = 0.0;
#pragma unroll
            for (int t = 0; t < NT; t++) { p[t] *= F[t]; u += p[t]; }
            f0 = (u + q2) - 1.0;
        }
    }
    if (r1 - r0 > bracket_dtol) {                              // cpp/Integration/Integration.cpp:494-510
        while (fabs(r0 - r1) > bracket_dtol) {
            const double xm = 0.5 * (r0 + r1);
            const double fm = ip_rootfun<NT>(c, xm);
            if (f0 * fm > 0) { r0 = xm; f0 = fm; }
            else { r1 = xm; f1 = fm; }
        }
        if (fabs(f0) > fabs(f1)) {
            double t = r0; r0 = r1; r1 = t;
            t = f0; f0 = f1; f1 = t;
        }
    }
    if (fabs(f0) < 1e-8) return r0;
    int ierr = 0;
    const double R = ip_newton<NT>(c, r0, 1e-6, ierr);
    if (ierr == 0) return R;
    return get_R_after_failure<NT>(lane, r0, r1, f0, f1, bracket_dtol, status);
}

// MUFU.RSQ64H seed (max relative error ~2^-22); negative -> NaN, +0 -> +Inf, which is exactly what the stop rule needs
__device__ __forceinline__ double rsqrt_seed(double x) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}

#ifndef KG_RSQRT_ORDER
#define KG_RSQRT_ORDER 3
#endif

// Where the tanh-sinh node rows of the current pair live: the first KS rows in shared memory (s = byte address in the shared
// window, sp = the same as a generic pointer for the generic-potential path), all KMAX rows in global memory (g). Row j holds
// node k = j + 1. The hot loops read shared rows with explicit 128-bit ld.shared (two per row at NT = 2): handed a plain pointer
// the compiler cannot prove the address space once a row may come from either table and falls back to 64-bit generic loads with
// 64-bit address arithmetic -- 2.5x the load instructions of a loop whose FP64 work is 7 instructions per node.
template<int NT>
struct NodeRows {
    unsigned s;
    const double* sp;
    const double* g;
    int KS;
    static constexpr int W = 2 + NT;
    __device__ __forceinline__ void load_s(int k, double (&r)[2 + NT]) const {          // node k <= KS
        const unsigned a = s + (unsigned) (k - 1) * (unsigned) (W * 8);
#pragma unroll
        for (int q = 0; q < W / 2; q++)
            asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(r[2 * q]), "=d"(r[2 * q + 1]) : "r"(a + 16u * q));
    }
    __device__ __forceinline__ void load_at(unsigned a, double (&r)[2 + NT]) const {    // shared byte address of the row
#pragma unroll
        for (int q = 0; q < W / 2; q++)
            asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(r[2 * q]), "=d"(r[2 * q + 1]) : "r"(a + 16u * q));
    }
    __device__ __forceinline__ void load_g(int k, double (&r)[2 + NT]) const {
        const double2* __restrict__ p = reinterpret_cast<const double2*>(g + (size_t) (k - 1) * W);
#pragma unroll
        for (int q = 0; q < W / 2; q++) { const double2 v = __ldg(p + q); r[2 * q] = v.x; r[2 * q + 1] = v.y; }
    }
    __device__ __forceinline__ void load(int k, double (&r)[2 + NT]) const { if (k <= KS) load_s(k, r); else load_g(k, r); }
    __device__ __forceinline__ unsigned addr(int k) const { return s + (unsigned) (k - 1) * (unsigned) (W * 8); }
};

// One tanh-sinh term  w_k f(u_k) = (w_k/u_k) X_k^(-1/2)  from a table row { (w/u)/2, 1/u^2, u^lam_0 .. }.
// Ap[t] = c2 A_t, so X = (1/u^2) (c2 - sum_t Ap_t u^lam_t) - 1 costs NT+1 DFMA. The reciprocal square root is a MUFU
// seed y0 plus one Householder step on e = 1 - X y0^2, fused with the weight: t = 2 z (1 + e/2 + 3e^2/8), z = (w/2u) y0
// (relative error < 1e-19 for order 3, < 3e-14 for order 2). X <= 0 or NaN gives t = NaN, which fails the stop test.
// This exact form is used for the first node and wherever the stop rule may trigger.
template<int NT>
__device__ __forceinline__ double ts_term(const double (&row)[2 + NT], const double (&Ap)[NT], double c2) {
    double P = c2;
#pragma unroll
    for (int t = 0; t < NT; t++) P = fma(-Ap[t], row[2 + t], P);
    const double X = fma(row[1], P, -1.0);
    const double y0 = rsqrt_seed(X);
    const double e = fma(-X, y0 * y0, 1.0);
    const double z = (row[0] + row[0]) * y0;
#if KG_RSQRT_ORDER == 3
    return fma(z * e, fma(e, 0.375, 0.5), z);
#else
    return fma(z * e, 0.5, z);
#endif
}

// Fast form of the same term for the bulk of the sum:
//   X  = (1/u^2) (c2 - sum_t Ap_t u^lam_t) - 1                                                     NT+1 DFMA
//   q  = MUFU.RSQ64H seed of X (relative error delta < 2^-21)
//   c  = 3 - (X q) q = 2 (3/2 - X q^2 / 2)   (one Newton step: rsqrt(X) = q c / 2, error 1.5 delta^2 < 4e-13)   DMUL + DFMA
//   z  = (w/2u) q ,  I += z c                                                                      DMUL + DFMA
// i.e. NT+5 FP64-pipe instructions per node and no FP64 compare: the stop rule is checked conservatively for a block
// of nodes at once on the high words (positive doubles order like integers): with c = 2 (1 +- 2^-20) and I_i <= I_n
// (all terms positive) it is enough that every z_i exceeds (tol/2) I_n by two high-word ulps and that I_n is finite
// and < 1e100. Anything else -- the end of the sum, a non-positive radicand, NaN/Inf -- falls through to the exact
// sequential replay, which is the reference's rule term by term.
template<int NT>
__device__ __forceinline__ void ts_fast(const double (&row)[2 + NT], const double (&Ap)[NT], double c2, double& z, double& c) {
    double P = c2;
#pragma unroll
    for (int t = 0; t < NT; t++) P = fma(-Ap[t], row[2 + t], P);
    const double X = fma(row[1], P, -1.0);
    const double q = rsqrt_seed(X);
    const double a = X * q;
    c = fma(-a, q, 3.0);
    z = row[0] * q;
}

constexpr int HI_1E100 = 0x54B249AD;      // high word of 1e100

// 1: every theta integral is summed node by node (validation switch, kg_set_exact_quadrature); 0: Euler-Maclaurin path below
static __constant__ int c_ts_exact = 0;
// acceptance threshold of the first (stride 16) tier on |i16 - i32| / |i16| (tuning knob: environment KG_EM_ACCEPT16 at model creation)
static __constant__ double c_em_accept16 = 1e-7;

// Exact replay of four consecutive nodes k+1 .. k+4 with the reference's sequential stop rule: a term is added only if all
// earlier ones were and |w f / I| > tol (Integration.cpp:431); NaN/Inf of the new integrand value (radicand <= 0) ends the sum
// (Integration.cpp:430). I is the running sum, k = index of the last node examined (what the reference's loop variable
// holds at exit). Returns true when the sum has ended.
template<int NT>
__device__ __noinline__ bool ts_exact4(const NodeRows<NT>& R, const double (&Ap)[NT], double cc, double& I, int& k) {
    double r[2 + NT];
    R.load(k + 1, r); const double t0 = ts_term<NT>(r, Ap, cc);
    R.load(k + 2, r); const double t1 = ts_term<NT>(r, Ap, cc);
    R.load(k + 3, r); const double t2 = ts_term<NT>(r, Ap, cc);
    R.load(k + 4, r); const double t3 = ts_term<NT>(r, Ap, cc);
    const double S1 = I + t0, S2 = S1 + t1, S3 = S2 + t2, S4 = S3 + t3;
    const bool ok0 = fabs(t0) > fabs(TS_TOL * I);
    const bool ok1 = ok0 && (fabs(t1) > fabs(TS_TOL * S1));
    const bool ok2 = ok1 && (fabs(t2) > fabs(TS_TOL * S2));
    const bool ok3 = ok2 && (fabs(t3) > fabs(TS_TOL * S3));
    if (ok3) { I = S4; k += 4; return false; }
    I = ok2 ? S3 : ok1 ? S2 : ok0 ? S1 : I;
    k += 1 + (int) ok0 + (int) ok1 + (int) ok2;
    return true;
}

#ifndef KG_TS_BLOCK
#define KG_TS_BLOCK 8
#endif

// cpp/Spherical.cpp:321-329 + cpp/Integration/Integration.cpp:417-437, on tabulated nodes.
// A[t] = eps C_t (sigma/R)^lam_t / (kT g^2), c2 = (R/b)^2. nodes += k at exit.
// The reference's loop continued from node k (running sum I): blocks of KG_TS_BLOCK fast nodes with one conservative
// high-word stop check per block; a block that may contain the stop (or anything irregular) is replayed exactly, four nodes at
// a time. On return k = index of the last node examined.
template<int NT>
__device__ __forceinline__ void ts_tail(const NodeRows<NT>& R, const double (&Ap)[NT], double c2, double& I, int& k, int& status) {
    for (;;) {
#ifndef KG_TS_NOFAST
        while (k + KG_TS_BLOCK <= R.KS) {
            const unsigned a0 = R.addr(k + 1);
            double In = I;
            int zmin = 0x7fffffff;
#pragma unroll
            for (int j = 0; j < KG_TS_BLOCK; j++) {
                double r[2 + NT], z, c;
                R.load_at(a0 + (unsigned) (j * NodeRows<NT>::W * 8), r);
                ts_fast<NT>(r, Ap, c2, z, c);
                In = fma(z, c, In);
                zmin = min(zmin, __double2hiint(z));
            }
            const int lim = __double2hiint((0.5 * TS_TOL) * In);
            if (!((zmin > lim + 2) && (lim < HI_1E100) && (lim >= 0))) break;
            I = In; k += KG_TS_BLOCK;
        }
#endif
        if (k + 4 > KMAX) { status |= 2; break; }
        if (ts_exact4<NT>(R, Ap, c2, I, k)) break;
    }
}

template<int NT>
__device__ __noinline__ double theta_integral(const NodeRows<NT>& R, const double (&A)[NT], double c2, unsigned& nodes, int& status) {
    double Ap[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) Ap[t] = c2 * A[t];
    double r[2 + NT];
    R.load_s(1, r);
    double I = ts_term<NT>(r, Ap, c2);              // k = 1 is added unconditionally (Integration.cpp:425)
    int k = 1;
    ts_tail<NT>(R, Ap, c2, I, k, status);
    nodes += (unsigned) k;
    return I;
}

// ------------------------------------------------------------------------------------------------------------
// Strided evaluation of the same truncated sum. The reference's step h = 1e-3 over-resolves the integrand by two orders
// of magnitude: in the node index k the terms g(k) = w_k f(u_k), k >= 2, are samples of a function that changes by about
// 1 % per node (node 1 carries the first-weight quirk and is added on its own). The Euler-Maclaurin formula relates the
// unit-stride sum to the stride-s trapezoid sum of the same samples,
//   sum_{k=a}^{b} g(k) = T_s + (g_a + g_b)/2 + (1-s^2)/12 D1 - (1-s^4)/720 D3 + (1-s^6)/30240 D5 + ...,  Dn = g^(n)(b) - g^(n)(a),
// with the end derivatives taken from 7-point forward differences of fine nodes. So: march with stride S in blocks of four
// coarse nodes until a block may contain the reference's stop (terms fall below 1.02 tol x running sum), close the coarse
// range at an even coarse node b, then continue the reference's loop from b + 1 (ts_tail) to find the stop index exactly.
// Every other coarse node gives the stride-2S sum for free; the two end-corrected integrals must agree to `accept`, otherwise
// (near-orbiting trajectories whose integrand has structure on the scale of a few nodes, a non-positive radicand, the
// early stop of the g = 1e-7 column, ...) the lane reports `redo` and the entry goes to the next tier.
//   S = 16, accept 1e-7 (first tier):  the stride-32 estimate is far worse than the stride-16 one (256x where the end corrections
//                                      dominate, quadratically where an interior feature does), so the difference over-estimates
//                                      the error of the stride-16 sum by orders of magnitude: measured on configs[1] against the
//                                      node-by-node sums, Omega agrees to 3e-11 and the reference node totals are identical
//                                      (tools/tune_accept.py, profiles/r2_tune_accept.md); ~97 % of the integrals are accepted
//   S = 8,  accept 1e-10 (second tier): error < 5e-13
// (random LJ 12-6 trajectories: tools/em_stride_probe.py; whole lattices: tests/test_gpu_chi_omega.py, tests/test_gpu_regimes.py).
// Third tier: node by node.
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void fd7(const double (&f)[7], double& d1, double& d3, double& d5) {
    d1 = -2.45 * f[0] + 6. * f[1] - 7.5 * f[2] + (20. / 3.) * f[3] - 3.75 * f[4] + 1.2 * f[5] - (1. / 6.) * f[6];
    d3 = -6.125 * f[0] + 29. * f[1] - 57.625 * f[2] + 62. * f[3] - 38.375 * f[4] + 13. * f[5] - 1.875 * f[6];
    d5 = -3.5 * f[0] + 20. * f[1] - 47.5 * f[2] + 60. * f[3] - 42.5 * f[4] + 16. * f[5] - 2.5 * f[6];
}

template<int NT, int S>
__device__ double theta_integral_em(const NodeRows<NT>& R, const double (&A)[NT], double c2, unsigned& nodes, unsigned& exec_nodes,
                                    int& status, bool& redo, bool& redo_early) {
    constexpr unsigned RB = NodeRows<NT>::W * 8;              // bytes per row
    const double ACCEPT = (S >= 16) ? c_em_accept16 : 1e-10;
    constexpr double S2 = (double) S * S, S4 = S2 * S2, S6 = S4 * S2;
    const int KS = R.KS;
    redo_early = false;
    double Ap[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) Ap[t] = c2 * A[t];
    redo = true;
    auto fast = [&](int k) {                                  // node k (1-based); NaN when the radicand is not positive
        double r[2 + NT], z, c;
        R.load(k, r);
        ts_fast<NT>(r, Ap, c2, z, c);
        return z * c;
    };
    double row[2 + NT];
    R.load_s(1, row);
    const double t1 = ts_term<NT>(row, Ap, c2);
    double f[7];
#pragma unroll
    for (int i = 0; i < 7; i++) { R.load_s(2 + i, row); double z, c; ts_fast<NT>(row, Ap, c2, z, c); f[i] = z * c; }
    double d1a, d3a, d5a;
    fd7(f, d1a, d3a, d5a);
    const double ga = f[0];
    // coarse sums in units of g: Sg = sum over coarse nodes j >= 1, Eg = the even ones; running sum I = base + S Sg
    const double base = t1 + (0.5 * S) * ga;
    double Sg = 0.0, Eg = 0.0, gb;
    int j0 = 1, bj;
    // Blocks whose four coarse rows are staged in shared memory (practically all: sums end near node 2000 of the 2560 staged)
    // step a shared-memory address; the remainder selects the table per row.
    const int j_smem = (KS - 2 - 3 * S) / S;                  // largest j0 with 2 + S (j0 + 3) <= KS
    unsigned ra = R.addr(2 + S);                              // row of coarse node j0 = 1 (node 2 + S)
    double z0, c0, z1, c1, z2, c2_, z3, c3, In;
    bool closing = false;
    while (j0 <= j_smem) {                                    // hot loop: 4 coarse nodes per trip
        double r0[2 + NT], r1[2 + NT], r2[2 + NT], r3[2 + NT];
        R.load_at(ra, r0); R.load_at(ra + S * RB, r1); R.load_at(ra + 2 * S * RB, r2); R.load_at(ra + 3 * S * RB, r3);
        ts_fast<NT>(r0, Ap, c2, z0, c0);
        ts_fast<NT>(r1, Ap, c2, z1, c1);
        ts_fast<NT>(r2, Ap, c2, z2, c2_);
        ts_fast<NT>(r3, Ap, c2, z3, c3);
        const double Sa = fma(z1, c1, fma(z0, c0, Sg));
        const double Sn = fma(z3, c3, fma(z2, c2_, Sa));
        In = fma((double) S, Sn, base);
        // every term of the block against 1.02 tol x the sum after the block, on the high words (z = g / 2 to 2^-20); a NaN sum
        // (radicand <= 0 somewhere) has a high word that fails one of the two integer tests
        const int lim = __double2hiint((0.5 * 1.02 * TS_TOL) * In);
        const int zmin = min(min(__double2hiint(z0), __double2hiint(z1)), min(__double2hiint(z2), __double2hiint(z3)));
        if (!((zmin > lim + 2) && (lim >= 0) && (lim < HI_1E100))) { closing = true; break; }
        Eg = fma(z3, c3, fma(z1, c1, Eg));
        Sg = Sn; j0 += 4; ra += 4 * S * RB;
    }
    if (!closing) {                                           // ran off the staged rows (sums that long are rare): table chosen per row
        for (;;) {
            const int kb = 2 + S * j0;
            if (kb + 3 * S + 8 > KMAX) { exec_nodes += (unsigned) (j0 + 8); return 0.0; }
            double r0[2 + NT];
            R.load(kb, r0); ts_fast<NT>(r0, Ap, c2, z0, c0);
            R.load(kb + S, r0); ts_fast<NT>(r0, Ap, c2, z1, c1);
            R.load(kb + 2 * S, r0); ts_fast<NT>(r0, Ap, c2, z2, c2_);
            R.load(kb + 3 * S, r0); ts_fast<NT>(r0, Ap, c2, z3, c3);
            const double Sa = fma(z1, c1, fma(z0, c0, Sg));
            const double Sn = fma(z3, c3, fma(z2, c2_, Sa));
            In = fma((double) S, Sn, base);
            const int lim = __double2hiint((0.5 * 1.02 * TS_TOL) * In);
            const int zmin = min(min(__double2hiint(z0), __double2hiint(z1)), min(__double2hiint(z2), __double2hiint(z3)));
            if (!((zmin > lim + 2) && (lim >= 0) && (lim < HI_1E100))) break;
            Eg = fma(z3, c3, fma(z1, c1, Eg));
            Sg = Sn; j0 += 4;
        }
    }
    if (!(In < 1e300)) { exec_nodes += (unsigned) (j0 + 11); return 0.0; }      // NaN (radicand <= 0 somewhere) or overflow
    {
        // the block that may contain the stop: close the coarse range at an even coarse node
        const double g0 = z0 * c0, g1 = z1 * c1, dl = (1.02 * TS_TOL) * In;
        if (g0 > dl && g1 > dl) { Sg += g0 + g1; Eg += g1; bj = j0 + 1; gb = g1; }
        else {
            bj = j0 - 1;
            if (j0 == 1) gb = ga;
            else { double r0[2 + NT], z, c; R.load(2 + S * bj, r0); ts_fast<NT>(r0, Ap, c2, z, c); gb = z * c; }     // g of the last accepted coarse node
        }
    }
    const double E = ga + Eg, O = Sg - Eg;
    if (bj * S < 64) { redo_early = true; exec_nodes += (unsigned) (j0 + 11); return 0.0; }      // the sum ends within the first ~70 nodes (g = 1e-7 column): cheap node by node
    const int b = 2 + S * bj;
    f[0] = gb;
#pragma unroll 1
    for (int i = 1; i < 7; i++) f[i] = fast(b + i);
    double D1, D3, D5;
    fd7(f, D1, D3, D5);
    D1 -= d1a; D3 -= d3a; D5 -= d5a;
    const double ends = 0.5 * (ga + gb);
    // T_s - integral = s^2/12 D1 - s^4/720 D3 + s^6/30240 D5
    const double is = S * (E + O - ends) - ((S2 / 12.) * D1 - (S4 / 720.) * D3 + (S6 / 30240.) * D5);
    const double i2s = 2 * S * (E - ends) - ((4. * S2 / 12.) * D1 - (16. * S4 / 720.) * D3 + (64. * S6 / 30240.) * D5);
    if (!(fabs(is - i2s) <= ACCEPT * fabs(is))) { exec_nodes += (unsigned) (bj + 17); return 0.0; }
    double I = t1 + (is + ends + ((1. / 12.) * D1 - (1. / 720.) * D3 + (1. / 30240.) * D5));
    int k = b;
    ts_tail<NT>(R, Ap, c2, I, k, status);                     // the reference's loop from node b + 1
    redo = false;
    nodes += (unsigned) k;
    exec_nodes += (unsigned) (bj + 13 + (k - b) + 4);
    return I;
}

// Node-by-node sum of ONE theta integral by a whole warp (third tier: the few integrals both strided sums reject, so that no
// lane ever walks ~2000 nodes alone while 31 others wait). Lane l owns nodes base + 1 + l + 32 i. Far from the stop the lanes add
// fast terms to private partial sums for WARP_RUN iterations (32 WARP_RUN nodes) and one warp reduction then checks, conservatively,
// that no node of the run could have triggered the reference's stop rule (every term above 1.02 tol x the sum after the run,
// everything finite); a run that fails the check is replayed from its start in the exact form with a warp scan, which gives the
// running sums the reference's rule needs term by term, and a ballot finds the first node that fails it.
// All lanes must call with the same (Ap, c2). Returns I; k_out = index of the last node examined.
constexpr int WARP_RUN = 8;
template<int NT>
__device__ double theta_integral_warp(const NodeRows<NT>& R, const double (&Ap)[NT], double c2, int& k_out, int& status) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    double row[2 + NT];
    R.load_s(1, row);
    double I = ts_term<NT>(row, Ap, c2);                      // node 1, added unconditionally (Integration.cpp:425)
    int base = 1;                                             // nodes <= base are summed
    // fast runs
    while (base + 32 * WARP_RUN <= R.KS) {
        double P = 0.0;
        int zmin = 0x7fffffff;
        unsigned a = R.addr(base + 1 + lane);
#pragma unroll
        for (int i = 0; i < WARP_RUN; i++) {
            double z, c;
            R.load_at(a, row);
            ts_fast<NT>(row, Ap, c2, z, c);
            P = fma(z, c, P);
            zmin = min(zmin, __double2hiint(z));
            a += 32u * NodeRows<NT>::W * 8u;
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) { P += __shfl_xor_sync(full, P, off); zmin = min(zmin, __shfl_xor_sync(full, zmin, off)); }
        const double In = I + P;
        const int lim = __double2hiint((0.5 * 1.02 * TS_TOL) * In);
        if (!((zmin > lim + 2) && (lim < HI_1E100) && (lim >= 0))) break;      // NaN sums give lim < 0 or a failed compare
        I = In; base += 32 * WARP_RUN;
    }
    // exact replay from `base`
    for (;; base += 32) {
        const int k = base + 1 + lane;
        double t = 0.0;
        bool ok = false;
        if (k <= KMAX) {
            R.load(k, row);
            t = ts_term<NT>(row, Ap, c2);
            ok = !(isnan(t) || isinf(t));
        }
        double P = ok ? t : 0.0;                              // inclusive prefix sum over the lanes
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) { const double v = __shfl_up_sync(full, P, off); if (lane >= off) P += v; }
        ok = ok && (fabs(t) > fabs(TS_TOL * (I + (P - t))));   // against the sum before this node (Integration.cpp:431)
        const unsigned pass = __ballot_sync(full, ok);
        if (pass == full) { I += __shfl_sync(full, P, 31); continue; }
        const int ff = __ffs(~pass) - 1;                      // first failing lane
        if (ff > 0) I += __shfl_sync(full, P, ff - 1);
        k_out = base + 1 + ff;
        if (k_out > KMAX) { status |= 2; k_out = KMAX; }
        return I;
    }
}

// Lane set-up shared by every caller: Bt = eps C_t / (kT g^2) (coeff overrides C[] for T-dependent potentials)
template<int NT>
__device__ __forceinline__ void make_lane(Lane<NT>& c, const PairParams& pp, const double* coeff, double T, double g, double b) {
    const double E = BOLTZMANN * T * (g * g);
#pragma unroll
    for (int t = 0; t < NT; t++) {
        const double Ct = (t < pp.nterms) ? (coeff ? coeff[t] : pp.C[t]) : 0.0;
        c.Bt[t] = pp.eps * Ct / E;
        c.lam[t] = (t < pp.nterms) ? pp.lam[t] : 1.0;
    }
    c.b = b;
    bool ip = true;
#pragma unroll
    for (int t = 0; t < NT; t++) ip = ip && (c.lam[t] == (double) (int) c.lam[t]) && (c.lam[t] >= 1.0) && (c.lam[t] <= 64.0);
    c.ipath = ip;
}
template<>
__device__ __forceinline__ void make_lane<1>(Lane<1>& c, const PairParams& pp, const double* coeff, double T, double g, double b) {
    c.gp = coeff ? reinterpret_cast<const GenPot*>(coeff) : pp.gen;     // per-T effective potential (Feynman-Hibbs) or the base one
    c.invE = 1.0 / (BOLTZMANN * T * (g * g));
    c.b = b;
}

// get_R for a caller that has no lane yet (the chi short-cuts b < 1e-3 still need R for the transfer lengths)
template<int NT>
__device__ double lane_get_R(const PairParams& pp, const double* coeff, double T, double g, double b, int& status) {
    Lane<NT> c;
    make_lane<NT>(c, pp, coeff, T, g, b);
    return get_R<NT>(c, g, status);
}

// Generic potentials: the reference's loop node for node (cpp/Integration/Integration.cpp:417-437) with
// f(u) = (R/u^2) [ r^4/b^2 (1 - u(r)/(kT g^2)) - r^2 ]^(-1/2) = (1/u) X^(-1/2), X = (R/b)^2 (1/u^2) (1 - U) - 1, r = R/u.
static __device__ double theta_integral_gen(const NodeRows<1>& NR, const Lane<1>& c, double R, unsigned& nodes, int& status) {
    const double* __restrict__ srows = NR.sp;
    const double* __restrict__ grows = NR.g;
    const int KS = NR.KS;
    const double q = R / c.b, c2 = q * q;
    const int zero_neg = c.gp->zero_on_negative;
    auto term = [&](int k, bool& bad) {
        const double* __restrict__ row = ((k <= KS) ? srows : grows) + (size_t) (k - 1) * 3;
        double U, dU;
        gen_eval(c.gp, R * row[2], 0, U, dU);
        const double X = fma(row[1], c2 * (1.0 - U * c.invE), -1.0);
        if (zero_neg && X < 0) { bad = false; return 0.0; }          // cpp/LJSpline.h:40
        const double f = (row[0] + row[0]) / sqrt(X);
        bad = isnan(f) || isinf(f);
        return f;
    };
    bool bad;
    double t = term(1, bad);
    double I = 0.0;
    int k = 1;
    for (;;) {
        I += t;                                           // Integration.cpp:425 (k = 1 unconditionally), :431
        k += 1;
        if (k > KMAX) { status |= 2; break; }
        t = term(k, bad);
        if (bad) break;                                   // Integration.cpp:430
        if (!(fabs(t) > fabs(TS_TOL * I))) break;         // Integration.cpp:431
    }
    nodes += (unsigned) k;
    return I;
}

// Full single-lane evaluation of the theta integral for (g, b): R, A_t, c2, then the sum. Returns I; R_out in sigma.
template<int NT>
__device__ double lane_theta_integral(const PairParams& pp, const double* coeff, const NodeRows<NT>& NR, double T, double g, double b,
                                      double& R_out, unsigned& nodes, int& status) {
    Lane<NT> c;
    make_lane<NT>(c, pp, coeff, T, g, b);
    const double R = get_R<NT>(c, g, status);
    R_out = R;
    double A[NT];
    lane_powers<NT>(c, 1.0 / R, A);                                        // Bt (sigma/R)^lam, as pot_eval forms it
    const double q = R / b;
    if (!c_ts_exact) {
        bool redo, early; unsigned ex = 0;
        double I = theta_integral_em<NT, 16>(NR, A, q * q, nodes, ex, status, redo, early);
        if (!redo) return I;
        if (!early) {
            I = theta_integral_em<NT, 8>(NR, A, q * q, nodes, ex, status, redo, early);
            if (!redo) return I;
        }
    }
    return theta_integral<NT>(NR, A, q * q, nodes, status);
}

// For the batched Omega kernel of the generic potentials (declaration; the power-law kernels use lane_setup + lane_sum_defer)
template<int NT>
__device__ double lane_theta_integral_defer(const PairParams& pp, const double* coeff, const NodeRows<NT>& NR, double T, double g, double b,
                                            unsigned& nodes, unsigned& exec_nodes, int& status, bool& redo, double (&Ap_out)[NT], double& c2_out);
// The same work in two phases, for the batched Omega kernels: lane_setup does the root finding (its own kernel: every lane
// is in get_R, small code), lane_sum_defer the strided sum from the stored A_t = Bt (sigma/R)^lam_t and c2 = (R/b)^2.
template<int NT>
__device__ void lane_setup(const PairParams& pp, const double* coeff, double T, double g, double b, double (&A)[NT], double& c2, int& status) {
    Lane<NT> c;
    make_lane<NT>(c, pp, coeff, T, g, b);
    const double R = c.ipath ? get_R_ip<NT>(c, g, status) : get_R<NT>(c, g, status);
    lane_powers<NT>(c, 1.0 / R, A);
    const double q = R / b;
    c2 = q * q;
}
template<int NT, int S>
__device__ double lane_sum_defer(const NodeRows<NT>& NR, const double (&A)[NT], double c2,
                                 unsigned& nodes, unsigned& exec_nodes, int& status, bool& redo, double (&Ap_out)[NT]) {
#pragma unroll
    for (int t = 0; t < NT; t++) Ap_out[t] = c2 * A[t];
    unsigned nd = 0;
    int st = 0;
    double I;
    if (c_ts_exact) { redo = false; I = theta_integral<NT>(NR, A, c2, nd, st); exec_nodes += nd; }
    else {
        bool early;
        I = theta_integral_em<NT, S>(NR, A, c2, nd, exec_nodes, st, redo, early);
        // whole columns end early together (neighbouring lanes share g), so summing them here costs no divergence
        if (redo && early) { redo = false; nd = 0; I = theta_integral<NT>(NR, A, c2, nd, st); exec_nodes += nd; }
    }
    if (!redo) { nodes += nd; status |= st; }
    return I;
}
template<>
__device__ inline double lane_theta_integral_defer<1>(const PairParams& pp, const double* coeff, const NodeRows<1>& NR, double T, double g, double b,
                                                      unsigned& nodes, unsigned& exec_nodes, int& status, bool& redo, double (&Ap_out)[1], double& c2_out);
template<>
__device__ inline double lane_theta_integral<1>(const PairParams& pp, const double* coeff, const NodeRows<1>& NR, double T, double g, double b,
                                                double& R_out, unsigned& nodes, int& status) {
    Lane<1> c;
    make_lane<1>(c, pp, coeff, T, g, b);
    const double R = get_R<1>(c, g, status);
    R_out = R;
    return theta_integral_gen(NR, c, R, nodes, status);
}
template<>
__device__ inline double lane_theta_integral_defer<1>(const PairParams& pp, const double* coeff, const NodeRows<1>& NR, double T, double g, double b,
                                                      unsigned& nodes, unsigned& exec_nodes, int& status, bool& redo, double (&Ap_out)[1], double& c2_out) {
    Ap_out[0] = 0.0; c2_out = 0.0;
    redo = false;                  // generic potentials are always summed node by node (piecewise potentials are not smooth in k)
    double R; unsigned nd = 0;
    const double I = lane_theta_integral<1>(pp, coeff, NR, T, g, b, R, nd, status);
    nodes += nd; exec_nodes += nd;
    return I;
}

} // namespace kg
