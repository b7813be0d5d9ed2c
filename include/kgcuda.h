/* include/kgcuda.h -- C-ABI of libkgcuda.so, the B200 (sm_100a) implementation of KineticGas's collision-integral
 * hot path. Plain C: opaque handles, plain pointers and sizes, int return codes (0 = ok, <0 = error, text via
 * kg_last_error). No exceptions cross this boundary. Host pointers unless the name ends in _dev.
 *
 * Each entry point names the reference interface it replaces (paths relative to thermotools/KineticGas @ 75848b1).
 * INTEGRATION.md shows the binding a reference maintainer adds inside cpp/KineticGas_mthr.cpp / cpp/bindings.cpp.
 */
#ifndef KGCUDA_H
#define KGCUDA_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define KG_MAX_TERMS 6

/* Per-pair potential block, filled by the C++ subclass that owns sigma/eps/lambda/C
 * (cpp/MieKinGas.h:16,26-42; cpp/ExtendedSutherland.h:124-128, ExtendedSutherland.cpp:113-119):
 *     u(r) = eps * sum_k C[k] * (sigma / r)^lam[k]
 * Mie: nterms = 2, C = {+C, -C}, lam = {lambda_r, lambda_a}. */
typedef enum { KG_POT_POWER_SUM = 0 } kg_potential_kind;
typedef struct {
    int32_t kind;        /* kg_potential_kind */
    int32_t nterms;      /* <= KG_MAX_TERMS */
    double sigma;        /* m */
    double eps;          /* J */
    double C[KG_MAX_TERMS];
    double lam[KG_MAX_TERMS];
} kg_potential_desc;

/* Per-work-item status bits (mirror the reference's throw sites) */
#define KG_STATUS_OK            0
#define KG_STATUS_GETR_MAXITER  1   /* "get_R reached max iter!"  cpp/Spherical.cpp:432 */
#define KG_STATUS_TS_EXHAUSTED  2   /* tanh-sinh ran past the node table (reference would warn at 10000, Integration.cpp:432) */
#define KG_STATUS_NAN           4   /* non-finite collision integral */
#define KG_STATUS_BMID          8   /* "Unable to converge bmid!"  cpp/transfer_lengths.cpp:259-262 */

typedef struct kg_model kg_model;
typedef struct kg_omega_plan kg_omega_plan;

/* One kg_model per KineticGas object (cpp/KineticGas.h:70-74 constructors). pairs: npairs = nc(nc+1)/2 blocks ordered
 * (0,0),(0,1),..,(0,nc-1),(1,1),.. ; for is_singlecomp models pass ncomps = 1 (cpp/Spherical.cpp:38-40 forces i=j=0).
 * mass: particle masses [kg], length ncomps. device: CUDA ordinal. */
int kg_model_create(const kg_potential_desc* pairs, int ncomps, const double* mass, int is_singlecomp, int device, kg_model** out);
void kg_model_destroy(kg_model* m);
const char* kg_last_error(const kg_model* m);   /* m may be NULL: error of the last failed create */
int kg_model_npairs(const kg_model* m);
/* Run all work of this model on a caller-owned CUDA stream (cudaStream_t passed as void*), e.g. the host framework's
 * current stream, so that the caller's events bracket the kernels. */
int kg_model_set_stream(kg_model* m, void* cuda_stream);

/* Replaces Spherical::chi / Spherical::get_R point calls (cpp/Spherical.cpp:455-465, 358-435; bound at
 * cpp/bindings.cpp:198-199). g dimensionless, b in units of sigma_ij. chi_out / R_out (R in units of sigma) may be NULL.
 * coeff (optional, nterms doubles) overrides C[] for this call (T-dependent effective coefficients). */
int kg_chi_points(kg_model* m, int pair, double T, int n, const double* g, const double* b_red,
                  const double* coeff, double* chi_out, double* R_out, int32_t* status);

/* Replaces the omega fan-out KineticGas::precompute_omega / precompute_conductivity / precompute_viscosity
 * (cpp/KineticGas_mthr.cpp:40-46,58-143) plus Spherical::omega -> w_integral (cpp/Spherical.cpp:37-94):
 * all Omega^(l,r)_ij(T) for nT temperatures x all pairs x nlr (l,r) combinations, by numerical quadrature on the
 * reference's T-dependent lattice. omega is [nT][npairs][nlr] in SI units (what cache.omega.store expects,
 * cpp/KineticGasCache.h:25-28); w (optional) the dimensionless W integrals, same shape.
 * coeff_T (optional) [nT][npairs][KG_MAX_TERMS]: per-temperature effective coefficients (Feynman-Hibbs / extended
 * Sutherland, cpp/ExtendedSutherland.cpp:182-200). status [nT][npairs]. */
int kg_omega_batch(kg_model* m, int nT, const double* T, int nlr, const int32_t* l, const int32_t* r,
                   const double* coeff_T, double* omega, double* w, int32_t* status);

/* Replaces Spherical::get_transfer_length for the default EWCA model (cpp/transfer_lengths.cpp:38-80,165-266; reached from
 * KineticGas::get_mtl / get_etl, cpp/KineticGas.h:78-85): momentum and energy transfer lengths [m] for nT temperatures x all
 * pairs, [nT][npairs]. One pass serves both (the reference recomputes every deflection angle per property). */
int kg_transfer_lengths_batch(kg_model* m, int nT, const double* T, const double* coeff_T, double* mtl, double* etl, int32_t* status);

/* Plan API: same computation with everything resident in HBM between calls (what a sweep driver / bench uses). */
int kg_omega_plan_create(kg_model* m, int nT, const double* T, int nlr, const int32_t* l, const int32_t* r,
                         const double* coeff_T, kg_omega_plan** out);
int kg_omega_plan_run(kg_omega_plan* p);                  /* asynchronous on the model's stream */
int kg_omega_plan_sync(kg_omega_plan* p);
int kg_omega_plan_fetch(kg_omega_plan* p, double* omega, double* w, int32_t* status);  /* D2H, synchronises */
const double* kg_omega_plan_omega_dev(const kg_omega_plan* p);   /* device pointer, [nT][npairs][nlr] */
int kg_omega_plan_export_dev(kg_omega_plan* p, double* omega_dev, double* w_dev);   /* D2D into caller buffers, async on the model stream */
void kg_omega_plan_destroy(kg_omega_plan* p);

/* Counters of the last run of a plan (algorithmic work and kernel time, for roofline accounting):
 *  out[0] lattice nodes whose chi was evaluated (touched by the adaptive mesh)   out[1] tanh-sinh nodes of those theta integrals
 *  out[2] columns whose theta_lim was evaluated                                   out[3] tanh-sinh nodes of those
 *  out[4] milliseconds in the chi kernel (CUDA events on the model stream)        out[5] ms in the theta_lim kernel
 *  out[6] ms in the mesh kernels          out[7] number of kernel launches        out[8] refinement passes
 *  out[9] H2D bytes of plan creation      out[10] D2H bytes of the last fetch     out[11] bytes of lattice + mask pools in HBM */
int kg_omega_plan_counters(kg_omega_plan* p, double* out12);

/* FP64 FMA-pipe peak probe: runs a register-resident DFMA chain kernel and returns the measured TFLOP/s
 * (2 flop per DFMA). Used as the roofline denominator (MEASURED_PEAKS.json has no FP64 figure). */
int kg_measure_fp64_peak(int device, double* tflops, double* sm_mhz_est);

#ifdef __cplusplus
}
#endif
#endif
