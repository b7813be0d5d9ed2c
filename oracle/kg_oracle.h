/* oracle/kg_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C, scalar, single-threaded restatement of the reference's collision-integral hot path
 * (thermotools/KineticGas @ 75848b1). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it; the product path (kineticgas_b200/) never does.
 *
 * Parity status: PINNED. Every function here is checked in tests/test_oracle_vs_reference.py against
 * oracle/_ref/libkgref.so (the unmodified reference C++ compiled by oracle/Makefile) when that library is
 * present, and against tests/golden/ fixtures generated from it (oracle/make_golden.py) otherwise.
 */
#ifndef KG_ORACLE_H
#define KG_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

#define KGO_MAX_TERMS 6

/* Pair potential u(r) = eps * sum_k C_k (sigma/r)^lam_k  (Mie: 2 terms, C_0 = +C, C_1 = -C;
 * reference: cpp/MieKinGas.h:30-42, cpp/ExtendedSutherland.cpp:113-119) */
typedef struct {
    double sigma, eps;
    int nterms;
    double C[KGO_MAX_TERMS];
    double lam[KGO_MAX_TERMS];
} kgo_pair;

typedef struct {
    double end_g, end_b, dg, db;
    int rlg, rlb;
    double lim;
} kgo_lattice;

void   kgo_mie_pair(double sigma, double eps, double la, double lr, kgo_pair* out);
void   kgo_lattice_for_T(double T, kgo_lattice* out);                    /* cpp/Spherical.cpp:60-84 */
double kgo_potential(const kgo_pair* p, double r);
double kgo_potential_r(const kgo_pair* p, double r);
double kgo_tanh_sinh_test(int which, double a, double h, double tol);
double kgo_get_R(const kgo_pair* p, double T, double g, double b, int* status);     /* b in metres */
double kgo_theta(const kgo_pair* p, double T, double g, double b, int* status);
double kgo_chi(const kgo_pair* p, double T, double g, double b, int* status);
double kgo_w_integrand(const kgo_pair* p, double T, double g, double b_red, int l, int r, int* status);
/* Adaptive 2-D quadrature on a caller-supplied table (pins the mesh replay on its own). */
double kgo_integrate2d_table(const double* tab, int nx, int ny, double ox, double oy, double ex, double ey,
                             double dx, double dy, int rlx, int rly, double lim, int* nevals);
/* Dimensionless collision integral W(l,r) on an explicit lattice. touched (optional, [nxmax*nymax] bytes) receives
 * the lattice nodes evaluated; nodes_* the tanh-sinh node counts (theta and theta_lim separately). */
double kgo_w_integral(const kgo_pair* p, double T, int l, int r, const kgo_lattice* lat, int* status,
                      long* n_chi, long* nodes_theta, long* nodes_lim);
double kgo_omega_prefactor(double sigma, double T, double m_i, double m_j, int same);   /* cpp/Spherical.cpp:46-47 */
long   kgo_last_nodes(void);
/* Exchange-weighted closest approach transfer lengths (cpp/transfer_lengths.cpp:165-266); property 0 = MTL, 1 = ETL; metres.
 * status bit 8 = "Unable to converge bmid!" */
double kgo_get_bmid(const kgo_pair* p, double g, double T, int* status);
double kgo_get_b_max_g(const kgo_pair* p, double g, double T, double bmid, int* status);
double kgo_tl_ewca(const kgo_pair* p, double T, int property, int* status);

#ifdef __cplusplus
}
#endif
#endif
