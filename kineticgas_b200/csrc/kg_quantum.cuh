// kg_quantum.cuh -- quantum-mechanical collision integrals on the GPU. Included by kg_omega.cu.
//
// Reference replaced: the second Omega engine of the reference, cpp/Quantum.cpp -- partial-wave phase shifts from the radial Schroedinger
// equation (Quantum::quantum_phase_shift, :418-520: Numerov outwards from inside the classically forbidden region until the particles
// are free, local phase shift from a 5-point derivative and spherical Bessel functions), the JKWB (semiclassical) phase shift above the
// energy / angular-momentum limits (Quantum::JKWB_phase_shift, :251-318; limits cpp/Quantum.h:153-154), the transport cross sections
// Q^(n)(E) as sums over l with Bose / Fermi / Boltzmann weights (cross_section_kernel, cross_section, :1176-1236) and their thermal
// average Omega^(n,s)(T) (quantum_omega, :1242-1259, simpson_inf over E/kT). This is what ModTangToennis / HFD_B2 / Patowski models
// return by default (pykingas/multiparam.py: quantum_active=True).
//
// Units: lengths in sigma, energies in eps; the only model constant of the wave equation is kappa^2 = 2 mu eps sigma^2 / hbar^2.
// Parallelism: one lane per partial wave -- a phase shift is a few thousand dependent Numerov steps -- and one warp per energy: the lanes
// take the l values of the parity class 32 at a time, the reference's sequential stop rule of the l-sum is applied to the chunk by lane 0.
#pragma once
#include <array>
#include <map>

namespace {

constexpr double Q_HBAR = 6.62607015e-34 / (2.0 * 3.14159265359);      // cpp/global_params.h: PLANCK / (2 PI)
constexpr double Q_JKWB_E = 120.0;                                      // cpp/Quantum.h:153-154
constexpr int Q_JKWB_L = 1000;

// u / eps and d(u / eps) / dx at r = x sigma
template<int NT>
__device__ __forceinline__ void qpot(const PairParams& pp, double x, double& U, double& dU) {
    if constexpr (NT == 1) {
        gen_eval_march(pp.gen, x, U, dU);
        const double ie = 1.0 / pp.eps;
        U *= ie; dU *= ie;
    } else {
        const double ix = 1.0 / x;
        double u = 0.0, du = 0.0;
#pragma unroll
        for (int t = 0; t < NT; t++) {
            if (t < pp.nterms) {
                const double lam = pp.lam[t];
                const double p = pp.C[t] * ((lam == (double) (int) lam && lam <= 64.0) ? ipow_pos(ix, (int) lam) : pow(ix, lam));
                u += p; du -= lam * p;
            }
        }
        U = u; dU = du * ix;
    }
}

// j_l, j_(l-1), y_l, y_(l-1) at X by upward recurrence (used at the end of the Numerov range, where X = k r exceeds l; for l = 0 the
// "l - 1" slots hold j_1, y_1)
__device__ void sph_jy(int l, double X, double& jl, double& jm, double& yl, double& ym) {
    double s, c;
    sincos(X, &s, &c);
    const double iX = 1.0 / X;
    double j0 = s * iX, j1 = (s * iX - c) * iX, y0 = -c * iX, y1 = (-c * iX - s) * iX;
    if (l == 0) { jl = j0; jm = j1; yl = y0; ym = y1; return; }
    for (int n = 1; n < l; n++) {
        const double f = (2 * n + 1) * iX;
        const double j2 = f * j1 - j0, y2 = f * y1 - y0;
        j0 = j1; j1 = j2; y0 = y1; y1 = y2;
    }
    jl = j1; jm = j0; yl = y1; ym = y0;
}

// innermost classically allowed radius (x = r / sigma) for angular momentum l at reduced energy E: cpp/Quantum.cpp:537-551
template<int NT>
__device__ double q_r_forbidden(const PairParams& pp, double kappa2, int l, double E, int& st) {
    const double L2 = (double) l * (l + 1) / kappa2;
    auto f = [&](double x, double& df) { double U, dU; qpot<NT>(pp, x, U, dU); df = dU - 2.0 * L2 / (x * x * x); return U + L2 / (x * x) - E; };
    double x = 0.3, df;
    int guard = 0;
    while (f(x, df) > 0 && guard++ < 2000) x *= 1.1;
    while (f(x, df) < 0 && guard++ < 4000) x *= 0.9;
    for (int it = 0;; it++) {                        // Integration.cpp:456-471 with ftol = 1e-10, dtol = -1
        const double fv = f(x, df);
        x -= fv / df;
        if (it > 50) { st |= KG_STATUS_GETR_MAXITER; break; }
        if (!(fabs(fv) > 1e-10)) break;
    }
    return x;
}

// cpp/Quantum.cpp:418-520. Returns the phase shift in (-pi/2, pi/2).
template<int NT>
__device__ double q_numerov_phase(const PairParams& pp, double kappa2, int l, double E, int& st) {
    const double ks = sqrt(kappa2 * E);              // k sigma
    const double h = 0.025 * fmin(1.0 / ks, 1.0);
    const double s2 = h * h / 12.0;
    const double LL = (double) l * (l + 1);
    auto gfun = [&](double x) { double U, dU; qpot<NT>(pp, x, U, dU); return LL / (x * x) + kappa2 * (U - E); };
    auto is_free = [&](double x) { double U, dU; qpot<NT>(pp, x, U, dU); return (fabs(U / E) < 1e-7) && (fabs(dU / (ks * E)) < 1e-7); };
    const double xcf = q_r_forbidden<NT>(pp, kappa2, l, E, st);
    if (is_free(xcf)) return 0.0;
    const double x0 = fmax(xcf - 3.0 * (2.0 * 3.14159265359 / ks), 0.5 * xcf);
    double r0 = 0, r1 = 0, r2 = 0, r3 = x0, r4 = x0 + h;
    double g0 = 0, g1 = 0, g2 = 0, g3 = gfun(r3), g4 = gfun(r4);
    double p0 = 0, p1 = 0, p2 = 0, p3 = 0, p4 = h;
    double Ul = 0, dUl = 0;                          // potential at the newest node: one evaluation serves the recurrence and the free test
    auto step = [&]() {
        r0 = r1; r1 = r2; r2 = r3; r3 = r4; g0 = g1; g1 = g2; g2 = g3; g3 = g4; p0 = p1; p1 = p2; p2 = p3; p3 = p4;
        r4 = r3 + h;
        qpot<NT>(pp, r4, Ul, dUl);
        g4 = LL / (r4 * r4) + kappa2 * (Ul - E);
        p4 = (p3 * (2.0 + 10.0 * s2 * g3) - p2 * (1.0 - s2 * g2)) / (1.0 - s2 * g4);
    };
    step(); step();
    int guard = 0;
    do { step(); } while (!((fabs(Ul / E) < 1e-7) && (fabs(dUl / (ks * E)) < 1e-7)) && guard++ < 4000000);
    if (guard >= 4000000) st |= KG_STATUS_TS_EXHAUSTED;
    const double A1 = 0.5 * (p3 - p1), A2 = 0.5 * (p4 - p0);
    const double B1 = s2 * (g3 * p3 - g1 * p1), B2 = s2 * (g4 * p4 - g0 * p0);
    const double dpsi = (16.0 / (21.0 * h)) * (-A1 + (37.0 * A2 / 32.0) - (37.0 * B1 / 5.0) - (17.0 * B2 / 40.0));
    const double gamma = dpsi / p2 - 1.0 / r2;
    const double X = ks * r2;
    double jl, jm, yl, ym;
    sph_jy(l, X, jl, jm, yl, ym);
    const double djl = (l == 0) ? -jm : jm - (l + 1) * jl / X;
    const double dyl = (l == 0) ? -ym : ym - (l + 1) * yl / X;
    return atan((ks * djl - gamma * jl) / (ks * dyl - gamma * yl));
}

// serial versions of the reference's quadrature helpers for one lane (cpp/Integration/Integration.cpp:332-373)
template<class F>
__device__ double q_simpson(F&& f, double x0, double xN, int N) {      // even N
    const double dx = (xN - x0) / N;
    double val = f(x0) + f(xN);
    for (int i = 1; i <= N / 2 - 1; i++) val += 4. * f(x0 + (2 * i - 1) * dx) + 2. * f(x0 + (2 * i) * dx);
    val += 4. * f(xN - dx);
    return val * (dx / 3.);
}
template<class F>
__device__ double q_simpson_inf(F&& f, double x0, double init_end, int& st, double tol = 1e-8) {
    double I = q_simpson(f, x0, init_end, 10);
    double dx = (init_end - x0) / 10.;
    double part_tol = tol * 1e6;
    int guard = 0;
    do {
        if (isnan(I)) { st |= KG_STATUS_NAN; break; }
        x0 = init_end;
        init_end += 10 * dx;
        const double Ip = q_simpson(f, x0, init_end, 10);
        I += Ip;
        const double conv = fabs(Ip / I);
        if (conv < tol) break;
        if (conv < part_tol) { dx *= 2; part_tol *= 0.1; }
    } while (part_tol > tol && guard++ < 100000);
    return I;
}
template<class F>
__device__ double q_bracket_positive(F&& f, double x0, double x1, int& st, double tol = 1e-10) {      // Integration.cpp:473-492
    double f0 = f(x0), f1 = f(x1);
    if (f0 * f1 > 0) { st |= KG_STATUS_BRACKET; return x1; }
    int guard = 0;
    while (((fabs(f0) > tol) || (fabs(f1) > tol)) && guard++ < 2000) {
        const double xm = 0.5 * (x0 + x1), fm = f(xm);
        if (f0 * fm > 0) { x0 = xm; f0 = fm; } else { x1 = xm; f1 = fm; }
    }
    if (f0 >= 0) return x0;
    if (f1 < 0) st |= KG_STATUS_BRACKET;
    return x1;
}

// semiclassical (JKWB) phase shift, cpp/Quantum.cpp:251-318; s0 = root of the potential / sigma (Spherical::get_potential_root)
template<int NT>
__device__ double q_jkwb_phase(const PairParams& pp, double kappa2, double s0, int l, double E, int& st) {
    if (E < 1e-10) return 0.0;
    const double ks = sqrt(kappa2 * E);
    const double T = 50.0;
    const double g = sqrt(E * pp.eps / (BOLTZMANN * T));
    const double b = (l + 0.5) / ks;
    double R = lane_get_R<NT>(pp, nullptr, T, g, b, st);
    auto Ured = [&](double r) { double U, dU; qpot<NT>(pp, r, U, dU); return U / E; };
    auto R_fun = [&](double r) { const double q = b / r; return 1.0 - q * q - Ured(r); };
    if (b > s0) {
        if (R_fun(R) < 0) R = q_bracket_positive(R_fun, R, b, st);
    } else if (R_fun(R) < 0) {
        double R2 = R;
        int guard = 0;
        while (R_fun(R2) < 0 && guard++ < 100000) R2 += 1e-2;
        R = q_bracket_positive(R_fun, R, R2, st);
    }
    auto integrand1 = [&](double r) { const double q = b / r; return sqrt(1.0 - q * q - Ured(r)) - sqrt(1.0 - q * q); };
    double I1, I2;
    if (b <= R) {
        double R2 = 1.5 * R;
        const double I0 = integrand1(R);
        while (integrand1(R2) < 0.5 * I0 && (R2 / R - 1) > 1e-6) R2 = 0.5 * (R + R2);
        I1 = q_simpson_inf(integrand1, R, R2, st);
        I2 = (b == R) ? 0.0 : -q_simpson([&](double r) { const double q = b / r; return sqrt(1.0 - q * q); }, b, R, 10);
    } else {
        double b2 = 1.5 * b;
        const double I0 = integrand1(b);
        while (integrand1(b2) < 0.5 * I0 && (b2 / b - 1) > 1e-3) b2 = 0.5 * (b + b2);
        I1 = q_simpson_inf(integrand1, b, b2, st);
        I2 = q_simpson([&](double r) { const double q = b / r; return sqrt(1.0 - q * q - Ured(r)); }, R, b, 10);
    }
    return ks * (I1 + I2);
}

// Spherical::get_potential_root / sigma (cpp/Spherical.cpp:236-240): Newton on u / eps from x = 1
template<int NT>
__device__ double q_potential_root(const PairParams& pp, int& st) {
    double x = 1.0;
    for (int it = 0;; it++) {
        double U, dU;
        qpot<NT>(pp, x, U, dU);
        x -= U / dU;
        if (it > 50) { st |= KG_STATUS_GETR_MAXITER; break; }
        if (!(fabs(U) > 1e-10)) break;
    }
    return x;
}

// Quantum::phase_shift (cpp/Quantum.cpp:522-528); which: 0 the reference's switch, 1 Numerov, 2 JKWB
template<int NT>
__device__ double q_phase(const PairParams& pp, double kappa2, double s0, int l, double E, int which, int& st, double jE = Q_JKWB_E, int jL = Q_JKWB_L) {
    const bool jk = (which == 2) || (which == 0 && (l >= jL || E > jE));
    return jk ? q_jkwb_phase<NT>(pp, kappa2, s0, l, E, st) : q_numerov_phase<NT>(pp, kappa2, l, E, st);
}

// A CTA's copy of the generic potential in shared memory (8.7 KB): the Numerov / JKWB marches read its coefficients at every step
template<int NT>
__device__ __forceinline__ PairParams q_stage_potential(const PairParams& pp, double* s_gp) {
    PairParams pq = pp;
    if constexpr (NT == 1) {
        const double* src = reinterpret_cast<const double*>(pp.gen);
        for (int i = threadIdx.x; i < GP_DOUBLES; i += blockDim.x) s_gp[i] = src[i];
        __syncthreads();
        pq.gen = reinterpret_cast<const GenPot*>(s_gp);
    }
    return pq;
}

template<int NT>
__global__ void __launch_bounds__(128) q_phase_kernel(PairParams pp_in, double kappa2, double jE, int jL, int n, const int* __restrict__ l, const double* __restrict__ E,
                                                      int which, double* __restrict__ out, int* __restrict__ status) {
    __shared__ double s_gp[NT == 1 ? GP_DOUBLES : 1];
    const PairParams pp = q_stage_potential<NT>(pp_in, s_gp);
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    int st = 0;
    const double s0 = q_potential_root<NT>(pp, st);
    out[e] = q_phase<NT>(pp, kappa2, s0, l[e], E[e], which, st, jE, jL);
    if (status) status[e] = st;
}

// cross_section_A (cpp/Quantum.cpp:1156-1174), k = 0, 1, 2 (the only values cross_section_kernel asks for)
__device__ double q_xsec_A(int n, int l, int k) {
    if (k == 0) return 1.0;
    if (k == 1) return (double) (n * (2 * l * l + 2 * (n * l - l) - n + 1)) / (double) (2 * (2 * l - 1) * (2 * (l + n) - 1));
    const double L = l;
    return ((double) (n * (n - 1)) / 8.) * (4. * L * L * L * L + (n - 3) * (8. * L * L * L + (3 - 8 * l) * (n - 2)) + 4 * L * L * ((double) n * n - 8 * n + 13))
           / (double) ((2 * l - 3) * (2 * l - 1) * (2 * l + 2 * n - 5) * (2 * l + 2 * n - 3));
}
// cross_section_kernel (cpp/Quantum.cpp:1176-1196) from cached phase shifts d[l]
__device__ double q_xsec_term(int n, int l, const double* __restrict__ d) {
    auto s2 = [&](int a, int b) { const double v = sin(d[a] - d[b]); return v * v; };
    switch (n) {
        case 0: { const double v = sin(d[l]); return (2 * l + 1) * v * v; }
        case 1: return (l + 1) * s2(l, l + 1);
        case 2: return (double) ((l + 1) * (l + 2)) * s2(l, l + 2) / (2 * l + 3);
        default: {
            double q = 0.0;
            for (int k = 0; (k < 3) && (2 * k + 1 <= n); k++) {
                double B = (2 * l + 1) * s2(l, l + n - 2 * k);
                // the reference's inner sum runs in unsigned integer arithmetic, (l + p) / (2 (l + p) - 1): 1 for l + p = 1, else 0
                double Bt = 0.0;
                for (int p = 1; p <= n - 2 * k; p++) Bt += (double) ((unsigned) (l + p) / (unsigned) (2 * (l + p) - 1));
                B *= Bt;
                q += q_xsec_A(n, l, k) * B;
            }
            return q;
        }
    }
}

// Quantum::cross_section (cpp/Quantum.cpp:1200-1236): one warp per energy. The lanes compute the phase shifts of 32 consecutive l at a time
// into the warp's cache; lanes 0 .. 2 nn - 1 each own one (n, parity) sum and advance it with the reference's stop rule as far as the cache allows.
constexpr int Q_MAXN = 8;
#ifdef KG_Q_MIN_CTAS               // build-time experiment: 6 (80 registers) and 8 (64) CTAs per SM were measured, see DESIGN 4c. The default
#define KG_Q_BOUNDS __launch_bounds__(128, KG_Q_MIN_CTAS)          // leaves the allocation to ptxas (128 registers, 4 CTAs): stating
#else                                                              // "4" explicitly gave a different, slower schedule (329 against 358 Omega/s)
#define KG_Q_BOUNDS __launch_bounds__(128)
#endif
template<int NT>
__global__ void KG_Q_BOUNDS q_xsec_kernel(PairParams pp_in, double kappa2, double jE, int jL, double w_sym, double w_anti, int nE, const double* __restrict__ E_in,
                                                     int nn, const int* __restrict__ nlist, double* __restrict__ dcache, int Lcap,
                                                     double* __restrict__ out, int* __restrict__ status) {
    __shared__ double s_gp[NT == 1 ? GP_DOUBLES : 1];
    const PairParams pp = q_stage_potential<NT>(pp_in, s_gp);
    const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (wid >= nE) return;
    const unsigned full = 0xffffffffu;
    int st = 0;
    const double E = fmax(E_in[wid], 5e-4);
    const double s0 = q_potential_root<NT>(pp, st);
    double* d = dcache + (size_t) wid * Lcap;
    int nmax = 0;
    for (int k = 0; k < nn; k++) nmax = max(nmax, nlist[k]);
    // my sum: lane = 2 k + parity
    const int myk = lane >> 1, par = lane & 1;
    const bool mine = (lane < 2 * nn) && ((par == 0) ? (w_sym > 0.0) : (w_anti > 0.0));
    const int myn = mine ? nlist[myk] : 0;
    double Q = 0.0;
    int lnext = par;
    bool done = !mine;
    int Lhave = 0;
    while (__any_sync(full, !done)) {
        if (Lhave + 32 > Lcap) { st |= KG_STATUS_TS_EXHAUSTED; break; }
        d[Lhave + lane] = q_phase<NT>(pp, kappa2, s0, Lhave + lane, E, 0, st, jE, jL);
        Lhave += 32;
        __syncwarp();
        while (!done && lnext + myn < Lhave) {
            const double q = q_xsec_term(myn, lnext, d);
            Q += q;
            lnext += 2;
            if (!(fabs(q) > 1e-6 * Q)) done = true;
        }
        __syncwarp();
    }
    const double pref = 4.0 * 3.14159265359 / (2.0 * kappa2 * E) * (pp.sigma * pp.sigma);
    const double contrib = mine ? ((par == 0) ? w_sym : w_anti) * Q : 0.0;
    const double other = __shfl_xor_sync(full, contrib, 1);
    if (lane < 2 * nn && par == 0) out[(size_t) wid * nn + myk] = (contrib + other) * pref;
    for (int off = 16; off > 0; off >>= 1) st |= __shfl_xor_sync(full, st, off);
    if (lane == 0 && status) status[wid] = st;
}

}   // namespace

// Partial-wave phase shifts of pair `pair` at reduced energies E[n] = E / eps and angular momenta l[n] (Quantum::phase_shift,
// quantum_phase_shift, JKWB_phase_shift: cpp/Quantum.cpp:251-318,418-528). which: 0 as the reference switches, 1 Numerov, 2 JKWB.
extern "C" int kg_quantum_phase_shifts(kg_model* m, int pair, int n, const int32_t* l, const double* E_red, int which, double* out, int32_t* status) {
    KG_TRACE();
    if (!m) return -1;
    auto set_error = [&](const std::string& s) { m->err = s; };
    if (pair < 0 || pair >= m->npairs || n < 0 || !l || !E_red || !out || which < 0 || which > 2) { set_error("kg_quantum_phase_shifts: invalid argument"); return -1; }
    if (m->coeff_T_required || (m->generic && m->gen_T_dependent)) { set_error("kg_quantum_phase_shifts: temperature-dependent (Feynman-Hibbs) potentials have no quantum engine"); return -1; }
    if (n == 0) return 0;
    for (int i = 0; i < n; i++) if (l[i] < 0 || !(E_red[i] > 0)) { set_error("kg_quantum_phase_shifts: l >= 0 and E > 0 required"); return -1; }
    CUDA_OK(cudaSetDevice(m->device));
    int ci = 0, cj = 0;
    { int p = 0; for (int i = 0; i < m->ncomps; i++) for (int j = i; j < m->ncomps; j++, p++) if (p == pair) { ci = i; cj = j; } }
    const double mu = m->mass[ci] * m->mass[cj] / (m->mass[ci] + m->mass[cj]);
    const PairParams& pp = m->pairs[pair];
    const double kappa2 = 2.0 * mu * pp.eps * pp.sigma * pp.sigma / (Q_HBAR * Q_HBAR);
    int *dl = nullptr, *dst = nullptr; double *dE = nullptr, *dout = nullptr;
    CUDA_OK(kg_dmalloc(m, &dl, n * sizeof(int))); CUDA_OK(kg_dmalloc(m, &dst, n * sizeof(int)));
    CUDA_OK(kg_dmalloc(m, &dE, n * sizeof(double))); CUDA_OK(kg_dmalloc(m, &dout, n * sizeof(double)));
    cudaStream_t s = m->stream;
    CUDA_OK(cudaMemcpyAsync(dl, l, n * sizeof(int), cudaMemcpyHostToDevice, s));
    CUDA_OK(cudaMemcpyAsync(dE, E_red, n * sizeof(double), cudaMemcpyHostToDevice, s));
    const int grid = (n + 127) / 128;
    switch (m->ntpad) {
        case 1: q_phase_kernel<1><<<grid, 128, 0, s>>>(pp, kappa2, m->q_jkwb_E, m->q_jkwb_l, n, dl, dE, which, dout, dst); break;
        case 2: q_phase_kernel<2><<<grid, 128, 0, s>>>(pp, kappa2, m->q_jkwb_E, m->q_jkwb_l, n, dl, dE, which, dout, dst); break;
        case 4: q_phase_kernel<4><<<grid, 128, 0, s>>>(pp, kappa2, m->q_jkwb_E, m->q_jkwb_l, n, dl, dE, which, dout, dst); break;
        default: q_phase_kernel<6><<<grid, 128, 0, s>>>(pp, kappa2, m->q_jkwb_E, m->q_jkwb_l, n, dl, dE, which, dout, dst); break;
    }
    CUDA_OK(cudaGetLastError());
    CUDA_OK(cudaMemcpyAsync(out, dout, n * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (status) CUDA_OK(cudaMemcpyAsync(status, dst, n * sizeof(int), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    m->dfree(dl); m->dfree(dst); m->dfree(dE); m->dfree(dout);
    return 0;
}

static double q_reduced_mass(const kg_model* m, int pair) {
    int ci = 0, cj = 0, p = 0;
    for (int i = 0; i < m->ncomps; i++) for (int j = i; j < m->ncomps; j++, p++) if (p == pair) { ci = i; cj = j; }
    return m->mass[ci] * m->mass[cj] / (m->mass[ci] + m->mass[cj]);
}

// device pass shared by the two entry points below: Q^(n)(E) for nE energies x nn orders -> host array q[nE][nn]
static int q_cross_sections_dev(kg_model* m, int pair, double w_sym, double w_anti, int nn, const int32_t* nlist, int nE, const double* E_red,
                                double* q, int32_t* status) {
    auto set_error = [&](const std::string& s) { m->err = s; };
    const PairParams& pp = m->pairs[pair];
    const double kappa2 = 2.0 * q_reduced_mass(m, pair) * pp.eps * pp.sigma * pp.sigma / (Q_HBAR * Q_HBAR);
    const int Lcap = 8192;
    int *dn = nullptr, *dst = nullptr; double *dE = nullptr, *dout = nullptr, *dcache = nullptr;
    CUDA_OK(kg_dmalloc(m, &dn, Q_MAXN * sizeof(int))); CUDA_OK(kg_dmalloc(m, &dst, nE * sizeof(int)));
    CUDA_OK(kg_dmalloc(m, &dE, nE * sizeof(double))); CUDA_OK(kg_dmalloc(m, &dout, (size_t) nE * nn * sizeof(double)));
    CUDA_OK(kg_dmalloc(m, &dcache, (size_t) nE * Lcap * sizeof(double)));
    cudaStream_t s = m->stream;
    // Longest first: a warp's serial work grows from ~1e3 to ~1e5 potential evaluations with the energy (Numerov: steps x chunks of 32 partial
    // waves; JKWB: quadratures x chunks), and CTAs are dispatched in index order, so the energies are launched in descending estimated cost
    // (the launch no longer ends on a few heavy warps that started last) and the results are put back in the caller's order.
    std::vector<int> order(nE);
    std::vector<double> cost(nE), E_sorted(nE);
    for (int e = 0; e < nE; e++) {
        const double E = std::max(E_red[e], 5e-4), ks = std::sqrt(kappa2 * E);
        const double chunks = std::ceil((4.0 * ks + 8.0) / 32.0);
        const double steps = (E > m->q_jkwb_E) ? 600.0 : std::max(2.0, std::pow(1e7 / E, 1.0 / 6.0)) * std::max(ks, 1.0) / 0.025;
        cost[e] = steps * chunks;
        order[e] = e;
    }
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cost[a] > cost[b]; });
    for (int e = 0; e < nE; e++) E_sorted[e] = E_red[order[e]];
    std::vector<double> q_sorted((size_t) nE * nn);
    std::vector<int32_t> st_sorted(nE);
    CUDA_OK(cudaMemcpyAsync(dn, nlist, nn * sizeof(int), cudaMemcpyHostToDevice, s));
    CUDA_OK(cudaMemcpyAsync(dE, E_sorted.data(), nE * sizeof(double), cudaMemcpyHostToDevice, s));
    const int grid = (nE + 3) / 4;                       // 4 warps per CTA, one warp per energy
    switch (m->ntpad) {
        case 1: q_xsec_kernel<1><<<grid, 128, 0, s>>>(pp, kappa2, m->q_jkwb_E, m->q_jkwb_l, w_sym, w_anti, nE, dE, nn, dn, dcache, Lcap, dout, dst); break;
        case 2: q_xsec_kernel<2><<<grid, 128, 0, s>>>(pp, kappa2, m->q_jkwb_E, m->q_jkwb_l, w_sym, w_anti, nE, dE, nn, dn, dcache, Lcap, dout, dst); break;
        case 4: q_xsec_kernel<4><<<grid, 128, 0, s>>>(pp, kappa2, m->q_jkwb_E, m->q_jkwb_l, w_sym, w_anti, nE, dE, nn, dn, dcache, Lcap, dout, dst); break;
        default: q_xsec_kernel<6><<<grid, 128, 0, s>>>(pp, kappa2, m->q_jkwb_E, m->q_jkwb_l, w_sym, w_anti, nE, dE, nn, dn, dcache, Lcap, dout, dst); break;
    }
    CUDA_OK(cudaGetLastError());
    CUDA_OK(cudaMemcpyAsync(q_sorted.data(), dout, (size_t) nE * nn * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaMemcpyAsync(st_sorted.data(), dst, nE * sizeof(int), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    m->dfree(dn); m->dfree(dst); m->dfree(dE); m->dfree(dout); m->dfree(dcache);
    for (int e = 0; e < nE; e++) {
        for (int k = 0; k < nn; k++) q[(size_t) order[e] * nn + k] = q_sorted[(size_t) e * nn + k];
        if (status) status[order[e]] = st_sorted[e];
    }
    return 0;
}

static int q_check(kg_model* m, int pair, double w_sym, double w_anti, const char* who) {
    if (pair < 0 || pair >= m->npairs || !(w_sym >= 0) || !(w_anti >= 0) || !(w_sym + w_anti > 0)) { m->err = std::string(who) + ": invalid argument"; return -1; }
    if (m->coeff_T_required || (m->generic && m->gen_T_dependent)) { m->err = std::string(who) + ": temperature-dependent (Feynman-Hibbs) potentials have no quantum engine"; return -1; }
    return 0;
}

// Quantum::cross_section (cpp/Quantum.cpp:1200-1236): Q^(n)(E) [m^2] for nE reduced energies x nn orders n[k] (0 <= n <= 6), out[nE][nn].
// w_sym / w_anti: weights of the even-l and odd-l sums (Quantum::get_symmetry_weights, :208-223: 0.5 / 0.5 for unlike particles,
// (1, 0) for spin-0 bosons, spin-dependent otherwise).
extern "C" int kg_quantum_cross_sections(kg_model* m, int pair, double w_sym, double w_anti, int nn, const int32_t* n, int nE, const double* E_red,
                                         double* out, int32_t* status) {
    KG_TRACE();
    if (!m) return -1;
    if (q_check(m, pair, w_sym, w_anti, "kg_quantum_cross_sections")) return -1;
    auto set_error = [&](const std::string& s) { m->err = s; };
    if (nn < 1 || nn > Q_MAXN || !n || nE < 0 || !E_red || !out) { set_error("kg_quantum_cross_sections: invalid argument"); return -1; }
    for (int k = 0; k < nn; k++) if (n[k] < 0 || n[k] > 6) { set_error("kg_quantum_cross_sections: order n out of range (0..6)"); return -1; }
    for (int k = 0; k < nE; k++) if (!(E_red[k] >= 0)) { set_error("kg_quantum_cross_sections: energies must be non-negative"); return -1; }
    if (nE == 0) return 0;
    CUDA_OK(cudaSetDevice(m->device));
    return q_cross_sections_dev(m, pair, w_sym, w_anti, nn, n, nE, E_red, out, status);
}

// Quantum::quantum_omega (cpp/Quantum.cpp:1242-1259): Omega^(n,s)(T) = sqrt(2 kT / (pi mu)) int_0^inf e^-x x^(s+1) Q^(n)(x kT / eps) dx by the
// reference's simpson_inf(kernel, 0, (s+1)/3) -- blocks of ten Simpson intervals marched outwards, the interval doubled as the blocks' share
// falls, until a block adds less than 1e-8 (cpp/Integration/Integration.cpp:346-373). The march is driven from the host: every round evaluates
// the next block of ALL unfinished (T, n, s) integrals in one device pass (one warp per distinct energy).
extern "C" int kg_quantum_omega(kg_model* m, int pair, double w_sym, double w_anti, int nT, const double* T, int nlr, const int32_t* l, const int32_t* r,
                                double* omega, int32_t* status) {
    KG_TRACE();
    if (!m) return -1;
    if (q_check(m, pair, w_sym, w_anti, "kg_quantum_omega")) return -1;
    auto set_error = [&](const std::string& s) { m->err = s; };
    if (nT < 1 || !T || nlr < 1 || !l || !r || !omega) { set_error("kg_quantum_omega: invalid argument"); return -1; }
    std::vector<int32_t> nlist;
    for (int k = 0; k < nlr; k++) {
        if (l[k] < 0 || l[k] > 6 || r[k] < 0 || r[k] > 24) { set_error("kg_quantum_omega: (l, r) out of range"); return -1; }
        if (std::find(nlist.begin(), nlist.end(), l[k]) == nlist.end()) nlist.push_back(l[k]);
    }
    if ((int) nlist.size() > Q_MAXN) { set_error("kg_quantum_omega: too many distinct orders"); return -1; }
    for (int t = 0; t < nT; t++) if (!(T[t] > 0)) { set_error("kg_quantum_omega: temperatures must be positive"); return -1; }
    CUDA_OK(cudaSetDevice(m->device));
    const PairParams& pp = m->pairs[pair];
    const double mu = q_reduced_mass(m, pair);
    const int nn = (int) nlist.size();
    struct Seq { int t, k, ni; double x0, xe, dx, part_tol, I; bool done, first; };
    std::vector<Seq> seqs;
    for (int t = 0; t < nT; t++) for (int k = 0; k < nlr; k++) {
        const int ni = (int) (std::find(nlist.begin(), nlist.end(), l[k]) - nlist.begin());
        const double end = (r[k] + 1.) / 3.;
        seqs.push_back(Seq{t, k, ni, 0.0, end, end / 10., 1e-8 * 1e6, 0.0, false, true});
    }
    if (status) for (int t = 0; t < nT; t++) status[t] = 0;
    const double tol = 1e-8;
    constexpr int AHEAD_MAX = 8;
    int AHEAD = 3;                 // blocks evaluated per round and integral: the pending one and more on the assumption that the interval
                                   // is not doubled (it is doubled a handful of times per integral; those speculative blocks are thrown away).
                                   // A round cannot be shorter than the slowest single energy (25-40 ms: one warp's serial march), so fewer,
                                   // larger rounds win until the discarded blocks outweigh them (KG_Q_AHEAD: tuning hook)
    if (const char* ev = getenv("KG_Q_AHEAD")) AHEAD = std::min(AHEAD_MAX, std::max(1, atoi(ev)));
    auto node_x = [](double x0, double xe, int i) {                  // simpson(func, x0, xe, 10)'s node expressions
        const double dxb = (xe - x0) / 10.;
        return (i == 0) ? x0 : (i == 10) ? xe : (i == 9) ? xe - dxb : x0 + i * dxb;
    };
    for (int round = 0; round < 4000; round++) {
        std::vector<double> E; std::map<double, int> index;
        std::vector<std::array<int, 11 * AHEAD_MAX>> node_of(seqs.size());
        bool any = false;
        for (size_t q = 0; q < seqs.size(); q++) {
            const Seq& sq = seqs[q];
            if (sq.done) continue;
            any = true;
            double x0 = sq.x0, xe = sq.xe;
            for (int bk = 0; bk < AHEAD; bk++) {
                for (int i = 0; i <= 10; i++) {
                    const double Er = node_x(x0, xe, i) * BOLTZMANN * T[sq.t] / pp.eps;
                    auto f = index.find(Er);
                    if (f == index.end()) { f = index.emplace(Er, (int) E.size()).first; E.push_back(Er); }
                    node_of[q][11 * bk + i] = f->second;
                }
                x0 = xe; xe += 10 * sq.dx;
            }
        }
        if (!any) break;
        std::vector<double> Q((size_t) E.size() * nn); std::vector<int32_t> st(E.size());
        const int rc = q_cross_sections_dev(m, pair, w_sym, w_anti, nn, nlist.data(), (int) E.size(), E.data(), Q.data(), st.data());
        if (rc) return rc;
        for (size_t q = 0; q < seqs.size(); q++) {
            Seq& sq = seqs[q];
            for (int bk = 0; bk < AHEAD && !sq.done; bk++) {
                auto kern = [&](int i) {
                    const double x = node_x(sq.x0, sq.xe, i);
                    const int e = node_of[q][11 * bk + i];
                    if (status) status[sq.t] |= st[e];
                    return exp(-x) * pow(x, r[sq.k] + 1) * Q[(size_t) e * nn + sq.ni];
                };
                double val = kern(0) + kern(10);
                for (int i = 1; i <= 4; i++) val += 4. * kern(2 * i - 1) + 2. * kern(2 * i);
                val += 4. * kern(9);
                const double Ip = val * ((sq.xe - sq.x0) / 10. / 3.);
                sq.I += Ip;
                bool doubled = false;
                if (sq.first) sq.first = false;
                else {
                    const double conv = fabs(Ip / sq.I);
                    if (conv < tol) { sq.done = true; break; }
                    if (conv < sq.part_tol) { sq.dx *= 2; sq.part_tol *= 0.1; doubled = true; }
                    if (!(sq.part_tol > tol)) { sq.done = true; break; }
                }
                if (std::isnan(sq.I)) { if (status) status[sq.t] |= KG_STATUS_NAN; sq.done = true; break; }
                sq.x0 = sq.xe;
                sq.xe += 10 * sq.dx;
                if (doubled) break;                                  // the blocks evaluated ahead assumed the old interval
            }
        }
    }
    for (const Seq& sq : seqs) omega[(size_t) sq.t * nlr + sq.k] = sq.I * sqrt(2. * BOLTZMANN * T[sq.t] / (3.14159265359 * mu));
    return 0;
}

// Quantum::set_quantum_active (cpp/Quantum.cpp:1520-1528) for the device tables: while active, kg_tcache_create takes every Omega^(l,r)_ij(T)
// from kg_quantum_omega with the given symmetry weights per pair (w_sym[npairs], w_anti[npairs]) instead of the classical quadrature.
extern "C" int kg_model_set_quantum(kg_model* m, int active, const double* w_sym, const double* w_anti) {
    KG_TRACE();
    if (!m) return -1;
    if (active && (!w_sym || !w_anti)) { m->err = "kg_model_set_quantum: symmetry weights required"; return -1; }
    if (active && (m->coeff_T_required || (m->generic && m->gen_T_dependent))) { m->err = "kg_model_set_quantum: temperature-dependent (Feynman-Hibbs) potentials have no quantum engine"; return -1; }
    m->quantum_active = active != 0;
    if (active) { m->q_wsym.assign(w_sym, w_sym + m->npairs); m->q_wanti.assign(w_anti, w_anti + m->npairs); }
    return 0;
}

// Quantum::set_JKWB_limits (cpp/Quantum.h:128-130): the semiclassical phase shift replaces the Numerov solution for E / eps > E_limit or l >= l_limit
extern "C" int kg_model_set_jkwb_limits(kg_model* m, double E_limit, int l_limit) {
    KG_TRACE();
    if (!m) return -1;
    if (!(E_limit > 0) || l_limit < 1) { m->err = "kg_model_set_jkwb_limits: invalid argument"; return -1; }
    m->q_jkwb_E = E_limit; m->q_jkwb_l = l_limit;
    return 0;
}
