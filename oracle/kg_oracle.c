/* oracle/kg_oracle.c -- TEST INFRASTRUCTURE ONLY (see kg_oracle.h for the rules and the parity status).
 *
 * Scalar restatement, statement for statement, of the reference's collision-integral path. Every function
 * cites the reference file:line it follows (paths relative to /root/reference). Arithmetic is kept in the
 * reference's operation order (same pow() calls, same truncated PI) so that, compiled with the same compiler
 * and libm, results agree with oracle/_ref to rounding.
 */
#include "kg_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* cpp/global_params.h:4-7 */
static const double BOLTZMANN = 1.38064852e-23;
static const double PI = 3.14159265359;
#define PLANE_FLTEPS 1e-12 /* cpp/Integration/Integration.cpp:17 */

static long g_nodes_theta = 0, g_nodes_lim = 0, g_nchi = 0;
static int g_in_lim = 0;
long kgo_last_nodes(void) { return g_nodes_theta + g_nodes_lim; }

/* ------------------------------------------------------------------------------------------------ potentials */
void kgo_mie_pair(double sigma, double eps, double la, double lr, kgo_pair* out) {
    /* cpp/MieKinGas.cpp:46-56 (set_C_alpha) */
    memset(out, 0, sizeof(*out));
    out->sigma = sigma; out->eps = eps; out->nterms = 2;
    double C = (lr / (lr - la)) * pow(lr / la, (la / (lr - la)));
    out->C[0] = C;  out->lam[0] = lr;
    out->C[1] = -C; out->lam[1] = la;
}

double kgo_potential(const kgo_pair* p, double r) {
    if (p->nterms == 2 && p->C[1] == -p->C[0]) {
        /* cpp/MieKinGas.h:30-32 */
        return p->C[0] * p->eps * (pow(p->sigma / r, p->lam[0]) - pow(p->sigma / r, p->lam[1]));
    }
    /* cpp/ExtendedSutherland.cpp:113-119 (C_k already the effective, T-dependent coefficients) */
    double u = 0.0;
    for (int k = 0; k < p->nterms; k++) u += p->C[k] * pow(p->sigma / r, p->lam[k]);
    return p->eps * u;
}

double kgo_potential_r(const kgo_pair* p, double r) {
    if (p->nterms == 2 && p->C[1] == -p->C[0]) {
        /* cpp/MieKinGas.h:34-37 */
        double la = p->lam[1], lr = p->lam[0];
        return p->C[0] * p->eps * ((la * pow(p->sigma / r, la) / r) - (lr * pow(p->sigma / r, lr) / r));
    }
    double u = 0.0;
    for (int k = 0; k < p->nterms; k++) u -= p->C[k] * p->lam[k] * pow(p->sigma / r, p->lam[k]) / r;
    return p->eps * u;
}

/* ------------------------------------------------------------------------------------------------ lattice */
void kgo_lattice_for_T(double T, kgo_lattice* L) {
    /* cpp/Spherical.cpp:60-84 */
    L->end_g = 8; L->end_b = 5; L->dg = 0.5; L->db = 0.03125; L->rlg = 4; L->rlb = 16; L->lim = 1e-5;
    if (T < 70) { L->dg /= 4.; L->end_g = 6; L->end_b = 3.5; }
    if (T < 50) { L->rlb /= 2; }
    if (T < 35) { L->dg /= 2; L->rlg *= 2; L->db /= 2; L->end_g = 3.5; L->end_b = 3.5; }
    if (T < 20) { L->dg /= 2; L->rlg *= 2; L->db /= 2; L->rlb *= 2; }
}

/* ------------------------------------------------------------------------------------------------ tanh-sinh */
typedef double (*f1d)(double u, void* ctx);

static double tanh_sinh(f1d func, void* ctx, double h, double tol, long* nodes) {
    /* cpp/Integration/Integration.cpp:417-437, including the (k*h*1e-3) first-weight quirk at :422 */
    const int n_max_intervals = 10000;
    /* the reference calls tanh_sinh with a run-time h: keep the compiler from folding the k = 1 node with MPFR
       (correctly rounded) instead of the libm calls the reference makes at run time (1-ulp differences otherwise) */
    volatile double h_runtime = h;
    h = h_runtime;
    double I = 0.0;
    int k = 1;
    double u = tanh(PI * sinh(k * h) / 2.);
    double w = (PI / 2.) * h * cosh(k * h) / pow(cosh(PI * sinh(k * h * 1e-3) / 2.), 2);
    double f = func(u, ctx);
    do {
        I += w * f;
        k += 1;
        u = tanh(PI * sinh(k * h) / 2.);
        w = (PI / 2) * h * cosh(k * h) / pow(cosh(PI * sinh(k * h) / 2.), 2);
        f = func(u, ctx);
        if (isnan(f) || isnan(w) || isinf(f) || isinf(w)) break;
    } while ((fabs((f * w) / I) > tol) && (k < n_max_intervals));
    if (nodes) *nodes += k;
    return I;
}

static double ts_test0(double u, void* c) { (void) c; return 1.0 / sqrt(1.0 - u * u); }
static double ts_test1(double u, void* c) { return exp(*(double*) c * u); }
double kgo_tanh_sinh_test(int which, double a, double h, double tol) {
    return which == 0 ? tanh_sinh(ts_test0, 0, h, tol, 0) : tanh_sinh(ts_test1, &a, h, tol, 0);
}

/* ------------------------------------------------------------------------------------------------ get_R */
typedef struct { const kgo_pair* p; double T, g, b; } rootctx; /* b reduced by sigma */

static double rootfun(const rootctx* c, double ri) {
    /* cpp/Spherical.cpp:368 */
    const kgo_pair* p = c->p;
    return (kgo_potential(p, ri * p->sigma) / (BOLTZMANN * c->T * pow(c->g, 2))) + pow(c->b / ri, 2) - 1;
}
static double d_rootfun(const rootctx* c, double ri) {
    /* cpp/Spherical.cpp:369 */
    const kgo_pair* p = c->p;
    return (p->sigma * kgo_potential_r(p, ri * p->sigma) / (BOLTZMANN * c->T * pow(c->g, 2))) - (2 / c->b) * pow(c->b / ri, 3);
}

static double newton_usafe(const rootctx* c, double x0, double ftol, double dtol, int* ierr) {
    /* cpp/Integration/Integration.cpp:456-471 */
    double f_val;
    int niter = 0;
    int max_iter = 50;
    double prev_x;
    do {
        prev_x = x0;
        f_val = rootfun(c, x0);
        x0 -= f_val / d_rootfun(c, x0);
        if (niter++ > max_iter) { *ierr = 1; return x0; }
    } while (fabs(f_val) > ftol && fabs(x0 - prev_x) > dtol);
    if (isnan(x0)) { *ierr = 2; return x0; }
    if (isinf(x0)) { *ierr = 3; return x0; }
    return x0;
}

static void bracket_root_usafe(const rootctx* c, double* x0, double* x1, double* f0, double* f1, double dtol, double ftol) {
    /* cpp/Integration/Integration.cpp:494-510 */
    while ((fabs(*x0 - *x1) > dtol) && (fabs(*f0) > ftol) && (fabs(*f1) > ftol)) {
        double x_mid = 0.5 * (*x0 + *x1);
        double f_mid = rootfun(c, x_mid);
        if ((*f0 * f_mid > 0)) { *x0 = x_mid; *f0 = f_mid; }
        else { *x1 = x_mid; *f1 = f_mid; }
    }
    if (fabs(*f0) > fabs(*f1)) {
        double t = *x0; *x0 = *x1; *x1 = t;
        t = *f0; *f0 = *f1; *f1 = t;
    }
}

static double get_R0(const kgo_pair* p, double T, double g) {
    /* cpp/Spherical.cpp:437-453 */
    double tol = 1e-6;
    double Ek = BOLTZMANN * T * pow(g, 2);
    double r = 0.9 * p->sigma;
    double f = kgo_potential(p, r) - Ek;
    while (f < 0) {
        r *= 0.8;
        f = kgo_potential(p, r) - Ek;
    }
    double dfdr = kgo_potential_r(p, r);
    while (fabs(f / (dfdr * p->sigma)) > tol) {
        r -= f / dfdr;
        f = kgo_potential(p, r) - Ek;
        dfdr = kgo_potential_r(p, r);
    }
    return r;
}

double kgo_get_R(const kgo_pair* p, double T, double g, double b, int* status) {
    /* cpp/Spherical.cpp:358-435 */
    b /= p->sigma;
    if ((b < 1e-5) || (g < 1e-6)) return get_R0(p, T, g);
    rootctx c = {p, T, g, b};
    double r0, r1;
    double bracket_dtol;
    double newton_dtol = 1e-6;
    double bracket_ftol = -1;
    double newton_ftol = -1;
    if (b > 1) { r0 = r1 = b; bracket_dtol = 0.1; }
    else { r0 = r1 = 1; bracket_dtol = 0.05; }
    double f0 = rootfun(&c, r0);
    double f1 = f0;
    while (f0 < 0) { r0 *= 0.95; f0 = rootfun(&c, r0); }
    if (r1 - r0 > bracket_dtol) bracket_root_usafe(&c, &r0, &r1, &f0, &f1, bracket_dtol, bracket_ftol);
    if (fabs(f0) < 1e-8) return r0 * p->sigma;

    int ierr = 0;
    double R = newton_usafe(&c, r0, newton_ftol, newton_dtol, &ierr);
    if (ierr != 0) { ierr = 0; R = newton_usafe(&c, r1, newton_ftol, newton_dtol, &ierr); }
    if (ierr != 0) { ierr = 0; R = newton_usafe(&c, 0.5 * (r0 + r1), newton_ftol, newton_dtol, &ierr); }
    int niter = 0;
    while (ierr != 0) {
        double prev_r0 = r0;
        while ((prev_r0 == r0) && (bracket_dtol > newton_dtol)) {
            if (fabs(r1 - r0) < bracket_dtol) bracket_dtol /= 2;
            bracket_root_usafe(&c, &r0, &r1, &f0, &f1, bracket_dtol, bracket_ftol);
        }
        if (bracket_dtol < newton_dtol) { R = r0; break; }
        ierr = 0;
        R = newton_usafe(&c, r0, newton_ftol, newton_dtol, &ierr);
        if (niter++ > 50) { if (status) *status = 1; /* "get_R reached max iter!" */ return NAN; }
    }
    return R * p->sigma;
}

/* ------------------------------------------------------------------------------------------------ theta, chi */
typedef struct { const kgo_pair* p; double T, R, g, b; } thetactx;

static double theta_integrand(const kgo_pair* p, double T, double r, double g, double b) {
    /* cpp/Spherical.cpp:327-329 */
    return pow((pow(r, 4) / pow(b, 2)) * (1.0 - kgo_potential(p, r) / (BOLTZMANN * T * pow(g, 2))) - pow(r, 2), -0.5);
}
static double theta_integrand_u(double u, void* cv) {
    /* cpp/Spherical.cpp:323 */
    const thetactx* c = (const thetactx*) cv;
    return (c->R / pow(u, 2)) * theta_integrand(c->p, c->T, c->R / u, c->g, c->b);
}
static double theta_integral(const kgo_pair* p, double T, double R, double g, double b) {
    /* cpp/Spherical.cpp:321-325 ; tanh_sinh default tol = 1e-5 (cpp/Integration/Integration.h) */
    thetactx c = {p, T, R, g, b};
    return tanh_sinh(theta_integrand_u, &c, 1.0e-3, 1e-5, g_in_lim ? &g_nodes_lim : &g_nodes_theta);
}
static double theta_lim(const kgo_pair* p, double T, double g, int* status) {
    /* cpp/Spherical.cpp:315-319 */
    double b = 100 * p->sigma;
    double R = kgo_get_R(p, T, g, b, status);
    g_in_lim = 1;
    double v = theta_integral(p, T, R, g, b);
    g_in_lim = 0;
    return v;
}
double kgo_theta(const kgo_pair* p, double T, double g, double b, int* status) {
    /* cpp/Spherical.cpp:259-270 */
    if (b / p->sigma > 100) return PI / 2;
    if (b / p->sigma < 1e-3) return 0;
    double R = kgo_get_R(p, T, g, b, status);
    return theta_integral(p, T, R, g, b) - theta_lim(p, T, g, status) + PI / 2;
}
double kgo_chi(const kgo_pair* p, double T, double g, double b, int* status) {
    /* cpp/Spherical.cpp:455-465 */
    if (b / p->sigma > 100) return 0;
    double t = kgo_theta(p, T, g, b, status);
    g_nchi++;
    return PI - 2.0 * t;
}
double kgo_w_integrand(const kgo_pair* p, double T, double g, double b, int l, int r, int* status) {
    /* cpp/Spherical.cpp:96-102 */
    if ((b > 2.5) && (g > 1.5)) return 0;
    const double chi_val = kgo_chi(p, T, g, b * p->sigma, status);
    return 2 * exp(-pow(g, 2)) * pow(g, 2.0 * r + 3.0) * (1 - pow(cos(chi_val), l)) * b;
}

/* ------------------------------------------------------------------------------------------------ integrate2d */
typedef struct { double x, y, z; } pt;
typedef double (*f2d)(double x, double y, void* ctx);
typedef struct {
    f2d func; void* ctx;
    double* memo; unsigned char* have; int nx, ny;   /* dense stand-in for the std::map memo (Integration.cpp:116-141) */
    double dx, dy, lim;
    int nevals;
} mesh;

static pt pt_add(pt a, pt b) { pt r = {a.x + b.x, a.y + b.y, a.z + b.z}; return r; }

static double eval_function(mesh* m, pt p, int Nx, int Ny) {
    /* cpp/Integration/Integration.cpp:116-141 : first visitor's coordinates define the stored value */
    if (Nx < 0 || Ny < 0 || Nx >= m->nx || Ny >= m->ny) abort();
    int idx = Nx * m->ny + Ny;
    if (!m->have[idx]) {
        m->memo[idx] = m->func(p.x, p.y, m->ctx);
        m->have[idx] = 1;
        m->nevals++;
    }
    return m->memo[idx];
}

static double integrate_plane(pt p1, pt p2, pt p3) {
    /* cpp/Integration/Integration.cpp:19-32 (get_plane, get_line) and :43-112 */
    if ((p1.x == p2.x) && (p1.x == p3.x)) return 0;
    double pa, pb, pc;
    {
        const double a = -p1.y * p3.z - p2.y * p1.z + p2.y * p3.z + p1.y * p2.z + p3.y * p1.z - p3.y * p2.z;
        const double b = p1.x * p3.z + p2.x * p1.z - p2.x * p3.z - p1.x * p2.z - p3.x * p1.z + p3.x * p2.z;
        const double c = -p1.x * p3.y - p2.x * p1.y + p2.x * p3.y + p1.x * p2.y + p3.x * p1.y - p3.x * p2.y;
        const double d = -(a * p1.x + b * p1.y + c * p1.z);
        pa = -a / c; pb = -b / c; pc = -d / c;
    }
    pt t;
#define SWAP(u, v) { t = u; u = v; v = t; }
    if ((p1.x <= p2.x) && (p1.x <= p3.x)) {
        if (p2.x > p3.x) SWAP(p2, p3)
    }
    else if ((p1.x >= p2.x) && (p1.x >= p3.x)) {
        SWAP(p1, p3)
        if (p1.x > p2.x) SWAP(p1, p2)
    }
    else {
        if (p2.x < p3.x) SWAP(p1, p2)
        else { SWAP(p1, p3) SWAP(p2, p3) }
    }
#undef SWAP
    /* get_line(p, q): A = (q.y - p.y) / (q.x - p.x); B = q.y - A * q.x */
    double u1a = (p3.y - p1.y) / (p3.x - p1.x), u1b = p3.y - u1a * p3.x;   /* l_upper_1 = line(p1, p3) */
    double l1a = (p2.y - p1.y) / (p2.x - p1.x), l1b = p2.y - l1a * p2.x;   /* l_lower_1 = line(p1, p2) */
    double u2a = u1a, u2b = u1b;                                           /* l_upper_2 = l_upper_1 */
    double l2a = (p3.y - p2.y) / (p3.x - p2.x), l2b = p3.y - l2a * p3.x;   /* l_lower_2 = line(p2, p3) */
    if (p3.y < p2.y) {
        double s;
        s = l1a; l1a = u1a; u1a = s; s = l1b; l1b = u1b; u1b = s;
        s = l2a; l2a = u2a; u2a = s; s = l2b; l2b = u2b; u2b = s;
    }
    double integral = 0;
    if (fabs(p1.x - p2.x) > PLANE_FLTEPS) {
        const double A31_star = u1a - l1a;
        const double B31_star = u1b - l1b;
        const double A31_dblstar = pow(u1a, 2) - pow(l1a, 2);
        const double B31_dblstar = 2 * (u1a * u1b - l1a * l1b);
        const double C31_dblstar = pow(u1b, 2) - pow(l1b, 2);
        const double A12_tilde = pa * A31_star + pb * A31_dblstar / 2;
        const double B12_tilde = pa * B31_star + pc * A31_star + pb * B31_dblstar / 2;
        const double C12_tilde = pc * B31_star + pb * C31_dblstar / 2;
        integral += (A12_tilde / 3) * (pow(p2.x, 3) - pow(p1.x, 3)) + (B12_tilde / 2) * (pow(p2.x, 2) - pow(p1.x, 2)) + C12_tilde * (p2.x - p1.x);
    }
    if (fabs(p2.x - p3.x) > PLANE_FLTEPS) {
        const double A23_star = u2a - l2a;
        const double B23_star = u2b - l2b;
        const double A23_dblstar = pow(u2a, 2) - pow(l2a, 2);
        const double B23_dblstar = 2 * (u2a * u2b - l2a * l2b);
        const double C23_dblstar = pow(u2b, 2) - pow(l2b, 2);
        const double A23_tilde = pa * A23_star + pb * A23_dblstar / 2;
        const double B23_tilde = pa * B23_star + pc * A23_star + pb * B23_dblstar / 2;
        const double C23_tilde = pc * B23_star + pb * C23_dblstar / 2;
        integral += (A23_tilde / 3) * (pow(p3.x, 3) - pow(p2.x, 3)) + (B23_tilde / 2) * (pow(p3.x, 2) - pow(p2.x, 2)) + C23_tilde * (p3.x - p2.x);
    }
    return integral;
}

static double integrate_adaptive(mesh* m, pt origin, int Nx_origin, int Ny_origin, int Nx_end, int Ny_end,
                                 int* Nxsteps, int Nysteps, double lim);

static void integration_step(mesh* m, pt* p1, pt* p2, pt* p3, int* Nx, int* Ny, double* integral,
                             int* Nxsteps_p, int Nysteps, double lim) {
    /* cpp/Integration/Integration.cpp:145-236 */
    const double dx = m->dx, dy = m->dy;
    const int Nxsteps = *Nxsteps_p;
    pt ystep = {0, dy * Nysteps, 0};
    pt xstep = {dx * Nxsteps, -dy * Nysteps, 0};
    double hx = fabs(dx * Nxsteps);
    double hy = fabs(dy * Nysteps);
    const int ax = abs(Nxsteps), ay = abs(Nysteps);

    pt q;
    double f = eval_function(m, *p3, *Nx, *Ny);
    q = *p3; q.x += hx;                     double f1x = eval_function(m, q, *Nx + ax, *Ny);
    q = *p3; q.x += 2 * hx;                 double f2x = eval_function(m, q, *Nx + 2 * ax, *Ny);
    q = *p3; q.y += hy;                     double f1y = eval_function(m, q, *Nx, *Ny + ay);
    q = *p3; q.y += 2 * hy;                 double f2y = eval_function(m, q, *Nx, *Ny + 2 * ay);
    q = *p3; q.x += hx; q.y += hy;          double f1x1y = eval_function(m, q, *Nx + ax, *Ny + ay);
    q = *p3; q.x += 2 * hx; q.y += 2 * hy;  double f2x2y = eval_function(m, q, *Nx + 2 * ax, *Ny + 2 * ay);
    double fxx = (f - 2 * f1x + f2x) / pow(dx * Nxsteps, 2);
    double fyy = (f - 2 * f1y + f2y) / pow(dy * Nysteps, 2);
    double fxy = (f2x2y - 2 * f1x1y + f - pow(hx, 2) * fxx - pow(hy, 2) * fyy) / (2 * hx * hy);
    double dblder = fabs(fxx) + fabs(fyy) + fabs(fxy);

    if ((dblder * pow(hx + hy, 3) > lim) && ((ax > 1) || (ay > 1))) {
        pt sub_origin = *p3;
        int sub_Nxsteps = Nxsteps;
        int sub_Nysteps = Nysteps;
        int sub_Nx_end = *Nx + Nxsteps;
        int sub_Ny_end = *Ny + Nysteps;
        if (fabs(fxx) * pow(hx, 3) > lim && (fabs(fyy) * pow(hy, 3) < lim) && (fabs(fxy) < lim) && (ax > 1)) {
            sub_Nxsteps = sub_Nxsteps / 2;
        }
        else if (fabs(fyy) * pow(hy, 3) > lim && (fabs(fxx) * pow(hx, 3) < lim) && (fabs(fxy) < lim) && (ay > 1)) {
            sub_Nysteps = sub_Nysteps / 2;
        }
        else {
            if (ax > 1) sub_Nxsteps = sub_Nxsteps / 2;
            if (ay > 1) sub_Nysteps = sub_Nysteps / 2;
        }
        *integral += integrate_adaptive(m, sub_origin, *Nx, *Ny, sub_Nx_end, sub_Ny_end, &sub_Nxsteps, sub_Nysteps, lim);
        *p1 = *p2;
        *p2 = *p3;
        *p3 = pt_add(*p2, ystep);
        *Ny += Nysteps;
        p3->z = eval_function(m, *p3, *Nx, *Ny);
        *integral += integrate_plane(*p1, *p2, *p3);
        /* p1 = p2 = p3 alias one point that is then moved by xstep (Integration.cpp:211-216) */
        *p3 = pt_add(*p3, xstep);
        *Nx += Nxsteps;
        *Ny -= Nysteps;
        p3->z = eval_function(m, *p3, *Nx, *Ny);
        *p1 = *p3;
        *p2 = *p3;
    }
    else {
        *p1 = *p2;
        *p2 = *p3;
        *p3 = pt_add(*p2, ystep);
        *Ny += Nysteps;
        p3->z = eval_function(m, *p3, *Nx, *Ny);
        *integral += integrate_plane(*p1, *p2, *p3);
        *p1 = *p2;
        *p2 = *p3;
        *p3 = pt_add(*p2, xstep);
        *Nx += Nxsteps;
        *Ny -= Nysteps;
        p3->z = eval_function(m, *p3, *Nx, *Ny);
        *integral += integrate_plane(*p1, *p2, *p3);
    }
}

static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

static double integrate_adaptive(mesh* m, pt origin, int Nx_origin, int Ny_origin, int Nx_end, int Ny_end,
                                 int* Nxsteps, int Nysteps, double lim) {
    /* cpp/Integration/Integration.cpp:240-280 */
    double integral = 0;
    pt p1 = origin, p2 = origin, p3 = origin;
    int Nx = Nx_origin, Ny = Ny_origin;
    p3.z = eval_function(m, p3, Nx, Ny);
    p1 = p2 = p3; /* the three shared_ptr alias one Point (Integration.cpp:253-260) */
    integration_step(m, &p1, &p2, &p3, &Nx, &Ny, &integral, Nxsteps, Nysteps, lim);
    while (Ny < Ny_end) {
        while (imin(Nx_origin, Nx_end) < Nx && Nx < imax(Nx_origin, Nx_end)) {
            integration_step(m, &p1, &p2, &p3, &Nx, &Ny, &integral, Nxsteps, Nysteps, lim);
        }
        pt ystep = {0, m->dy * Nysteps, 0};
        p1 = p2;
        p2 = p3;
        p3 = pt_add(p2, ystep);
        Ny += Nysteps;
        p3.z = eval_function(m, p3, Nx, Ny);
        integral += integrate_plane(p1, p2, p3);
        if (Ny < Ny_end) {
            *Nxsteps *= -1;
            integration_step(m, &p1, &p2, &p3, &Nx, &Ny, &integral, Nxsteps, Nysteps, lim);
        }
    }
    return integral;
}

static double integrate2d(f2d func, void* ctx, double ox, double oy, double ex, double ey, double dx, double dy,
                          int rlx, int rly, double lim, int* nevals, unsigned char** have_out, int* nx_out, int* ny_out) {
    /* cpp/Integration/Integration.cpp:283-307 */
    double delta_x = ex - ox;
    double delta_y = ey - oy;
    int Nx_end = (int) (delta_x / dx + 0.5);
    int Ny_end = (int) (delta_y / dy + 0.5);
    int Nxsteps = rlx;
    int Nysteps = rly;
    mesh m;
    m.func = func; m.ctx = ctx; m.dx = dx; m.dy = dy; m.lim = lim; m.nevals = 0;
    m.nx = Nx_end + 2 * rlx + 1; m.ny = Ny_end + 2 * rly + 1;
    m.memo = (double*) calloc((size_t) m.nx * m.ny, sizeof(double));
    m.have = (unsigned char*) calloc((size_t) m.nx * m.ny, 1);
    pt origin = {ox, oy, 0};
    double val = integrate_adaptive(&m, origin, 0, 0, Nx_end, Ny_end, &Nxsteps, Nysteps, lim);
    if (nevals) *nevals = m.nevals;
    free(m.memo);
    if (have_out) { *have_out = m.have; *nx_out = m.nx; *ny_out = m.ny; } else free(m.have);
    return val;
}

typedef struct { const double* tab; int nx, ny; double ox, oy, dx, dy; } tabctx;
static double tab_func(double x, double y, void* cv) {
    const tabctx* c = (const tabctx*) cv;
    int ix = (int) ((x - c->ox) / c->dx + 0.5), iy = (int) ((y - c->oy) / c->dy + 0.5);
    if (ix < 0 || iy < 0 || ix >= c->nx || iy >= c->ny) return 0.0;
    return c->tab[ix * c->ny + iy];
}
double kgo_integrate2d_table(const double* tab, int nx, int ny, double ox, double oy, double ex, double ey,
                             double dx, double dy, int rlx, int rly, double lim, int* nevals) {
    tabctx c = {tab, nx, ny, ox, oy, dx, dy};
    return integrate2d(tab_func, &c, ox, oy, ex, ey, dx, dy, rlx, rly, lim, nevals, 0, 0, 0);
}

typedef struct { const kgo_pair* p; double T; int l, r; int* status; } wctx;
static double w_func(double g, double b, void* cv) {
    const wctx* c = (const wctx*) cv;
    return kgo_w_integrand(c->p, c->T, g, b, c->l, c->r, c->status);
}
double kgo_w_integral(const kgo_pair* p, double T, int l, int r, const kgo_lattice* lat, int* status,
                      long* n_chi, long* nodes_theta, long* nodes_lim) {
    /* cpp/Spherical.cpp:52-94 (lattice chosen by the caller: kgo_lattice_for_T) */
    wctx c = {p, T, l, r, status};
    g_nodes_theta = g_nodes_lim = g_nchi = 0;
    double I = integrate2d(w_func, &c, 1e-7, 1e-7, lat->end_g, lat->end_b, lat->dg, lat->db, lat->rlg, lat->rlb, lat->lim, 0, 0, 0, 0);
    if (n_chi) *n_chi = g_nchi;
    if (nodes_theta) *nodes_theta = g_nodes_theta;
    if (nodes_lim) *nodes_lim = g_nodes_lim;
    return I;
}

double kgo_omega_prefactor(double sigma, double T, double m_i, double m_j, int same) {
    /* cpp/Spherical.cpp:46-47 ; M_ij = m_i/(m_i+m_j), m0 = m_i+m_j (cpp/KineticGas.cpp:90-99) */
    if (same) return pow(sigma, 2) * sqrt((PI * BOLTZMANN * T) / m_i);
    double m0 = m_i + m_j, Mij = m_i / m0, Mji = m_j / m0;
    return 0.5 * pow(sigma, 2) * sqrt(2 * PI * BOLTZMANN * T / (m0 * Mij * Mji));
}

/* ================================================================================================ transfer lengths
 * Exchange-weighted closest approach (EWCA), cpp/transfer_lengths.cpp:151-266 */
static double dimless_relative_vdf(double g) {
    /* cpp/transfer_lengths.cpp:161-163 */
    return sqrt(2. / PI) * pow(g, 2) * exp(-0.5 * pow(g, 2));
}
static double momentum_transfer(const kgo_pair* p, double T, double g, double b, int* st) {
    /* cpp/transfer_lengths.cpp:207-212 */
    double chi_val = kgo_chi(p, T, g, b * p->sigma, st);
    return g * sqrt(2 * (1 - cos(chi_val))) * fabs(sin(chi_val / 2.));
}
static double energy_transfer(const kgo_pair* p, double T, double g, double b, int* st) {
    /* cpp/transfer_lengths.cpp:214-219 */
    double chi_val = kgo_chi(p, T, g, b * p->sigma, st);
    return g * fabs(cos(chi_val) - sin(chi_val) - 1);
}
static double ewca_weight(const kgo_pair* p, double T, double g, double b, int property, int* st) {
    /* cpp/transfer_lengths.cpp:196-205 */
    if (property == 0) return momentum_transfer(p, T, g, b, st) * dimless_relative_vdf(g) * g * b;
    return energy_transfer(p, T, g, b, st) * dimless_relative_vdf(g) * g * b;
}
double kgo_get_b_max_g(const kgo_pair* p, double g, double T, double bmid, int* st) {
    /* cpp/transfer_lengths.cpp:221-232 */
    double b, db, chi_val;
    if (g == 0) return 10.;
    b = 2 * bmid;
    db = 0.5;
    chi_val = kgo_chi(p, T, g, b * p->sigma, st);
    while (fabs(chi_val + 1e-5) > 1e-3) {
        b += db;
        chi_val = kgo_chi(p, T, g, b * p->sigma, st);
    }
    return b;
}
double kgo_get_bmid(const kgo_pair* p, double g, double T, int* st) {
    /* cpp/transfer_lengths.cpp:234-266 */
    if (g == 0) return 5.;
    double b = 0.9;
    double db = 0.1;
    double chi_val = kgo_chi(p, T, g, b * p->sigma, st);
    double next_chi_val;
    double dchidb = 10;
    double step_tol = 1e-6;
    double chi_tol = 1e-6;
    while ((fabs(chi_val) > chi_tol) && (fabs(db) > step_tol)) {
        if ((chi_val < 0 && db > 0) || (chi_val > 0 && db < 0)) db *= -0.9;
        next_chi_val = kgo_chi(p, T, g, (b + db) * p->sigma, st);
        while (fabs(next_chi_val) >= fabs(chi_val)) {
            db *= 0.1;
            next_chi_val = kgo_chi(p, T, g, (b + db) * p->sigma, st);
            if (fabs(db) < step_tol) break;
        }
        if (fabs(db) > step_tol) {
            b += db;
            dchidb = (chi_val - next_chi_val) / db;
            chi_val = next_chi_val;
            db = -chi_val / dchidb;
        }
        if ((b > 10) || (b < 0)) { if (st) *st |= 8; /* "Unable to converge bmid!" */ return NAN; }
    }
    return b;
}
typedef struct { const kgo_pair* p; double T, g; int property; int* st; } ewctx;
static void ws_funcs(const ewctx* c, double b, double* f, double* w) {
    /* the func / wt lambdas of ewca_inner, cpp/transfer_lengths.cpp:186-187 */
    *w = ewca_weight(c->p, c->T, c->g, b, c->property, c->st);
    *f = kgo_get_R(c->p, c->T, c->g, b * c->p->sigma, c->st);
}
static void weighted_simpson_fw(const ewctx* c, double x0, double xN, int N_intervals, double* F_out, double* W_out) {
    /* cpp/Integration/Integration.cpp:375-394 (note (N/2)-1 with integer division: an odd N drops a node) */
    double dx = (xN - x0) / N_intervals;
    double F = 0, W = 0, w1 = 0, w2 = 0, f1, f2;
    ws_funcs(c, x0, &f1, &w1); ws_funcs(c, xN, &f2, &w2);
    W += w1 + w2;
    F += f1 * w1 + f2 * w2;
    for (size_t i = 1; i <= (size_t) ((N_intervals / 2) - 1); i++) {
        ws_funcs(c, x0 + (2 * i - 1) * dx, &f1, &w1);
        ws_funcs(c, x0 + (2 * i) * dx, &f2, &w2);
        W += 4. * w1 + 2. * w2;
        F += 4. * f1 * w1 + 2. * f2 * w2;
    }
    ws_funcs(c, xN - dx, &f1, &w1);
    W += 4. * w1;
    F += 4. * f1 * w1;
    W *= dx / 3.;
    F *= dx / 3.;
    *F_out = F; *W_out = W;
}
static void ewca_inner(const kgo_pair* p, double T, double g, int property, int* st, double* F, double* W) {
    /* cpp/transfer_lengths.cpp:181-194 */
    const double bmid = kgo_get_bmid(p, g, T, st);
    const double bmax = kgo_get_b_max_g(p, g, T, bmid, st);
    ewctx c = {p, T, g, property, st};
    double F1, W1, F2, W2;
    weighted_simpson_fw(&c, 0., bmid, 15, &F1, &W1);
    weighted_simpson_fw(&c, bmid, bmax, 20, &F2, &W2);
    *F = F1 + F2; *W = W1 + W2;
}
static void weighted_simpson_pair(const kgo_pair* p, double T, int property, int* st, double x0, double xN, int N_intervals, double* F_out, double* W_out) {
    /* cpp/Integration/Integration.cpp:396-415 */
    double dx = (xN - x0) / N_intervals;
    double F = 0, W = 0, a1, b1, a2, b2;
    ewca_inner(p, T, x0, property, st, &a1, &b1); ewca_inner(p, T, xN, property, st, &a2, &b2);
    F += a1 + a2;
    W += b1 + b2;
    for (size_t i = 1; i <= (size_t) ((N_intervals / 2) - 1); i++) {
        ewca_inner(p, T, x0 + (2 * i - 1) * dx, property, st, &a1, &b1);
        ewca_inner(p, T, x0 + (2 * i) * dx, property, st, &a2, &b2);
        F += 4. * a1 + 2. * a2;
        W += 4. * b1 + 2. * b2;
    }
    ewca_inner(p, T, xN - dx, property, st, &a1, &b1);
    F += 4. * a1;
    W += 4. * b1;
    F *= dx / 3;
    W *= dx / 3;
    *F_out = F; *W_out = W;
}
double kgo_tl_ewca(const kgo_pair* p, double T, int property, int* st) {
    /* cpp/transfer_lengths.cpp:165-179 ; property 0 = MTL, 1 = ETL ; metres */
    double g0 = 0., g1 = 3.5, gmax = 5.;
    double F1, W1, F2, W2;
    weighted_simpson_pair(p, T, property, st, g0, g1, 15, &F1, &W1);
    weighted_simpson_pair(p, T, property, st, g1, gmax, 10, &F2, &W2);
    double F = F1 + F2;
    double W = W1 + W2;
    return F / W;
}
