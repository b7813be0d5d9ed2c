/* include/kgcuda.h -- C-ABI of libkgcuda.so, the B200 (sm_100a) implementation of KineticGas's collision-integral
 * hot path. Plain C: opaque handles, plain pointers and sizes, int return codes (0 = ok, <0 = error, text via
 * kg_last_error). No exceptions cross this boundary. Host pointers unless the name ends in _dev.
 *
 * Each entry point names the reference interface it replaces (paths relative to thermotools/KineticGas @ 75848b1).
 * INTEGRATION.md shows the binding a reference maintainer adds inside cpp/KineticGas_mthr.cpp / cpp/bindings.cpp.
 */
#ifndef KGCUDA_H
#define KGCUDA_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define KG_MAX_TERMS 6

/* Per-pair potential block, filled by the C++ subclass that owns sigma/eps/lambda/C
 * (cpp/MieKinGas.h:16,26-42; cpp/ExtendedSutherland.h:124-128, ExtendedSutherland.cpp:113-119):
 *     u(r) = eps * sum_k C[k] * (sigma / r)^lam[k]
 * Mie: nterms = 2, C = {+C, -C}, lam = {lambda_r, lambda_a}. */
typedef enum {
    KG_POT_POWER_SUM = 0,      /* Mie, Feynman-Hibbs Mie, extended Sutherland: kg_potential_desc below */
    KG_POT_LJ_SPLINE = 1,      /* cpp/LJSpline.h:45-82        no parameters beyond sigma, eps */
    KG_POT_LJTS = 2,           /* cpp/LJTS.h:28-60            truncated (2.5 sigma) and shifted LJ */
    KG_POT_TANG_TOENNIES = 3,  /* cpp/multiparam.cpp:5-89     params {A, b, A_tilde, a_tilde, a1, a2, am1, am2, L_unit, short_range_lim, C6, C8, C10, C12, C14, C16} (TangToennisParam, energies in K) */
    KG_POT_HFD_B2 = 4,         /* cpp/multiparam.cpp:91-182   params {A, alpha, c6, c8, c10, beta_star, D, rm[nm]} (HFD_B2_Param) */
    KG_POT_PATKOWSKI = 5       /* cpp/multiparam.cpp:184-279  params {Rc, Ac, Bc, Cex1, Cex2, Csp1, Csp2, Csp3, Csp4, delta, C6, C8, C10} (PatowskiParam) */
} kg_potential_kind;
typedef struct {
    int32_t kind;        /* kg_potential_kind */
    int32_t nterms;      /* <= KG_MAX_TERMS */
    double sigma;        /* m */
    double eps;          /* J */
    double C[KG_MAX_TERMS];
    double lam[KG_MAX_TERMS];
} kg_potential_desc;

/* Per-pair block of the non-power-law potentials (kinds 1-5). sigma/eps are Spherical::sigma / eps of the pair (the
 * length and energy scales of the reduced (g, b) lattice). fh_order > 0 adds the Feynman-Hibbs correction of
 * FH_Corrected<T> (cpp/extentions.h:228-273) with fh_D = hbar^2 / (24 mu_ij) [J m^2] (set_D_factors, :308-314); the
 * T-dependent effective potential is rebuilt inside the library for every temperature of a batch. */
#define KG_MAX_POT_PARAMS 24
typedef struct {
    int32_t kind;        /* kg_potential_kind, 1..5 */
    int32_t fh_order;    /* 0, 1 or 2 */
    double sigma;        /* m */
    double eps;          /* J */
    double fh_D;         /* J m^2; ignored when fh_order == 0 */
    int32_t nparams;
    int32_t reserved;
    double params[KG_MAX_POT_PARAMS];
} kg_potential_desc_ex;

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
/* Same for models built on kg_potential_desc_ex blocks: LJSpline, LJTS, ModTangToennis, HFD_B2, Patowski and their
 * FH_Corrected variants (cpp/multiparam.h:36-131; Quantum::omega forwards to Spherical::omega for them,
 * cpp/Quantum.cpp:1261-1268). Every entry point below accepts such a model; coeff / coeff_T arguments are ignored. */
int kg_model_create_ex(const kg_potential_desc_ex* pairs, int ncomps, const double* mass, int is_singlecomp, int device, kg_model** out);
/* Host-only: the piecewise polynomial-times-exponential form of one potential block at temperature T as the device will
 * evaluate it (layout: struct GenPot of kineticgas_b200/csrc/kg_genpot.h, as raw doubles). Returns the number of doubles
 * written, or minus the required capacity. Used by the CPU tests to pin the Feynman-Hibbs folding. */
int kg_genpot_describe(const kg_potential_desc_ex* d, double T, double* out, int cap_doubles);
void kg_model_destroy(kg_model* m);
const char* kg_last_error(const kg_model* m);   /* m may be NULL: error of the last failed create */
/* "libkgcuda sm_100a src-sha256:<16 hex>": hash of kineticgas_b200/csrc/* + include/kgcuda.h the binary was built from
 * (__graft_entry__.source_sha() recomputes it from the tree; bench.py prints both so a stale binary is visible). */
const char* kg_build_info(void);
/* Page-locked host buffers (cudaHostAlloc). Arrays passed to kg_transport_batch / kg_omega_plan_fetch from such memory are copied by
 * asynchronous DMA, overlapped with the kernels chunk by chunk; pageable memory works too (staged by the driver). NULL on failure. */
void* kg_host_alloc(size_t bytes);
void kg_host_free(void* p);
int kg_model_npairs(const kg_model* m);
/* Declares that the coefficients C[k] of a power-sum model depend on temperature (Feynman-Hibbs Mie: C_eff(T), cpp/ExtendedSutherland.cpp:
 * 182-200). Afterwards every call that integrates -- kg_omega_batch, kg_omega_plan_create*, kg_transfer_lengths_batch, kg_tcache_create,
 * kg_chi_points and kg_transport_batch with cache == NULL -- fails with an error unless it is given coeff_T, instead of silently using the
 * base coefficients. */
int kg_model_set_temperature_dependent(kg_model* m, int flag);
/* Run all work of this model on a caller-owned CUDA stream (cudaStream_t passed as void*), e.g. the host framework's
 * current stream, so that the caller's events bracket the kernels. */
int kg_model_set_stream(kg_model* m, void* cuda_stream);

/* Replaces Spherical::chi / Spherical::get_R point calls (cpp/Spherical.cpp:455-465, 358-435; bound at
 * cpp/bindings.cpp:198-199). g dimensionless, b in units of sigma_ij. chi_out / R_out (R in units of sigma) may be NULL.
 * coeff (optional, nterms doubles) overrides C[] for this call (T-dependent effective coefficients). */
int kg_chi_points(kg_model* m, int pair, double T, int n, const double* g, const double* b_red,
                  const double* coeff, double* chi_out, double* R_out, int32_t* status);

/* kg_chi_points_noalloc: kg_chi_points without any device allocation, for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it). NO COUNTERPART IN THE REFERENCE.
 * Same arguments, results and status bits as kg_chi_points. The device staging buffers are one fixed file-scope array in the library
 * (36 bytes per point, KG_CHI_NOALLOC_MAX_POINTS points = 2.25 MiB, held by every GPU that loads the library until exit); n beyond the
 * maximum is rejected (-1, kg_last_error says so), as are temperature-dependent generic (Feynman-Hibbs tabulated) potentials. All calls
 * on one device share the array: calls that overlap on the GPU race on it -- serialise them. Calls cudaMalloc / cudaFree never.
 * kg_chi_points_noalloc_max() returns the limit. */
#define KG_CHI_NOALLOC_MAX_POINTS 65536
int kg_chi_points_noalloc(kg_model* m, int pair, double T, int n, const double* g, const double* b_red,
                          const double* coeff, double* chi_out, double* R_out, int32_t* status);
int kg_chi_points_noalloc_max(void);

/* Replaces Spherical::potential / potential_derivative_r (cpp/bindings.cpp:187-189) for n radii r [m] at temperature T
 * (T matters for Feynman-Hibbs corrected models only): u [J], du/dr [J/m]. Either output may be NULL. */
int kg_potential_points(kg_model* m, int pair, double T, int n, const double* r, double* u_out, double* du_out);

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

/* ---- classical cross sections and second virial coefficient on the same collision dynamics ----
 * kg_cross_sections: Spherical::cross_section (cpp/Spherical.cpp:104-127; bound at cpp/bindings.cpp:253-255 for the Quantum classes as
 * classical_cross_section): Q^(l)(E) [m^2] for E[nE] in units of the pair's eps and l[nl] (1 <= nl <= 8); out [nE][nl]. Bisection for the
 * root of chi, the two Simpson segments and the outward Simpson march are the reference's, block for block. status [nE] (may be NULL).
 * kg_second_virial: Spherical::second_virial (cpp/Spherical.cpp:140-160) [m^3/mol] at T[nT]; coeff_T [nT][KG_MAX_TERMS] of this pair for
 * T-dependent coefficients. Power-sum models only. */
#define KG_STATUS_BRACKET 64        /* "Bracket solver: Initial values have same sign!" (cpp/Integration/Integration.h:224-226) */
int kg_cross_sections(kg_model* m, int pair, int nE, const double* E_red, int nl, const int32_t* l, const double* coeff,
                      double* out, int32_t* status);
int kg_second_virial(kg_model* m, int pair, int nT, const double* T, const double* coeff_T, double* out);
/* Plan API: same computation with everything resident in HBM between calls (what a sweep driver / bench uses). */
int kg_omega_plan_create(kg_model* m, int nT, const double* T, int nlr, const int32_t* l, const int32_t* r,
                         const double* coeff_T, kg_omega_plan** out);
/* need (optional) [nT][npairs][nlr] bytes: 0 = this Omega is not wanted (e.g. it comes from MieKinGas::omega's correlation
 * branch, cpp/MieKinGas.cpp:81-93); its output slot is 0 and pairs/temperatures with no wanted entry cost nothing. */
int kg_omega_plan_create_masked(kg_model* m, int nT, const double* T, int nlr, const int32_t* l, const int32_t* r,
                                const double* coeff_T, const uint8_t* need, kg_omega_plan** out);
int kg_omega_plan_run(kg_omega_plan* p);                  /* asynchronous on the model's stream */
int kg_omega_plan_sync(kg_omega_plan* p);
int kg_omega_plan_fetch(kg_omega_plan* p, double* omega, double* w, int32_t* status);  /* D2H, synchronises */
const double* kg_omega_plan_omega_dev(const kg_omega_plan* p);   /* device pointer, [nT][npairs][nlr] */
int kg_omega_plan_export_dev(kg_omega_plan* p, double* omega_dev, double* w_dev);   /* D2D into caller buffers, async on the model stream */
void kg_omega_plan_destroy(kg_omega_plan* p);

/* Counters of the last run of a plan (algorithmic work and kernel time, for roofline accounting):
 *  out[0] lattice nodes whose chi was evaluated (touched by the adaptive mesh)   out[1] tanh-sinh nodes of those theta integrals
 *  out[2] columns whose theta_lim was evaluated                                   out[3] tanh-sinh nodes of those
 *  out[4] milliseconds in the theta kernels (CUDA events on the model stream)      out[5] tanh-sinh node evaluations actually executed
 *  out[6] ms in the mesh kernels          out[7] number of kernel launches        out[8] refinement passes
 *  out[9] H2D bytes of plan creation      out[10] D2H bytes of the last fetch     out[11] bytes of lattice + mask pools in HBM
 *  out[12] largest cell list of one refinement level   out[13] cell-list capacity   out[14] its hard bound   out[15] times it was grown */
int kg_omega_plan_counters(kg_omega_plan* p, double* out16);
/* Quadrature accounting of the last run (8 doubles): out[0] tanh-sinh nodes of the reference's sums (the algorithmic count of
 * SURVEY 8d), out[1] node evaluations executed (the strided Euler-Maclaurin sums touch ~1/12 of them; rejected attempts are
 * counted), out[2] theta integrals that both strided tiers rejected and that were summed node by node, out[3] theta integrals in
 * total, out[4] theta integrals whose stride-16 sum was rejected (second tier, stride 8), out[5] mesh cells processed over all
 * refinement levels, out[6..7] reserved (0). */
int kg_omega_plan_quadrature_stats(kg_omega_plan* p, double* out8);
/* Validation switch, process wide: 1 = every theta integral is summed node by node exactly like tanh_sinh
 * (cpp/Integration/Integration.cpp:417-437); 0 (default) = strided sum with end corrections (same result to ~4e-13, same
 * stop index). The environment variable KG_TS_EXACT=1 selects the exact mode at model creation. */
int kg_set_exact_quadrature(int exact);


/* ================================================================================================================
 * State-point assembly (Revised Enskog Theory): everything KineticGas does per (T, Vm, x) after the collision
 * integrals exist -- RDF at contact, K-factors, square-bracket sums, the Sonine-expansion linear systems and the
 * property formulas -- batched over state points on the GPU.
 * ================================================================================================================ */
#define KG_MAX_COMPS 6          /* internal component count (a single-component model is a 2-component mixture inside,
                                   cpp/KineticGas.cpp:139-144) */
#define KG_MAX_ORDER 4          /* Enskog approximation order N */

/* FrameOfReference, same values as cpp/utils.h:37-45 */
#define KG_FOR_COM 0
#define KG_FOR_CON 1
#define KG_FOR_COV 2
#define KG_FOR_SOLVENT 3
#define KG_FOR_ZARATE 4     /* Ortiz de Zarate frames (ideal gas only): D is always (nc-1) x (nc-1), D_T has nc-1 entries */
#define KG_FOR_ZARATE_X 5
#define KG_FOR_ZARATE_W 6

#define KG_PROP_VISCOSITY          1u   /* KineticGas::viscosity                 cpp/KineticGas_properties.cpp:28-58   */
#define KG_PROP_CONDUCTIVITY       2u   /* KineticGas::thermal_conductivity      cpp/KineticGas_properties.cpp:69-141  */
#define KG_PROP_DIFFUSION          4u   /* KineticGas::interdiffusion            cpp/KineticGas_properties.cpp:167-206 */
#define KG_PROP_THERMAL_DIFFUSION  8u   /* KineticGas::thermal_diffusion_coeff   cpp/KineticGas_properties.cpp:208-280 */

#define KG_STATUS_SINGULAR      16  /* zero pivot in a Sonine-expansion system */
#define KG_STATUS_T_NOT_CACHED  32  /* state-point temperature absent from the kg_tcache it was run against */
#define KG_STATUS_COMPOSITION   128 /* WARNING only, results are computed: "Mole fractions do not sum to unity" (pykingas/py_KineticGas.py:119-141,
                                       |sum x - 1| > 1e-12), reported per state point so that callers need no host pass over x */

typedef struct {
    int32_t N;                        /* Enskog approximation order (>= 1; conductivity needs >= 2) */
    int32_t is_idealgas;              /* KineticGas::is_idealgas: rdf = 1, K = K' = 1, no density terms */
    int32_t omega_correlation_active; /* MieKinGas::set_omega_correlation_active (cpp/MieKinGas.cpp:150-155) */
    int32_t frame_of_reference;       /* KG_FOR_* */
    int32_t dependent_idx;            /* python-style negative index allowed (-1 = last) */
    int32_t solvent_idx;
    int32_t do_compress;              /* drop the dependent row/column of D (compress_diffusion_matrix) */
    uint32_t properties;              /* KG_PROP_* bit mask */
} kg_transport_opts;

/* EoS-derived inputs (thermopack on the host in the reference, cpp/eos_interface.h:57-135); nc_eos = 1 for a
 * single-component model else ncomps. Any pointer may be NULL when the requested properties do not need it. The arrays may live in host
 * or in device memory (unified addressing; e.g. the outputs of kg_eos_tv_batch_dev). */
typedef struct {
    const double* dmudn;      /* [n][nc_eos][nc_eos]  (dmu_i/dn_j)_{T,V} at V = Vm, n = x (1 mol)  [J/mol^2]  -> diffusion */
    const double* cp_ideal;   /* [n][nc_eos]          ideal-gas Cp of each pure component [J/mol K]         -> conductivity */
    const double* pressure;   /* [n]                  (unused unless frame = CoV) */
    const double* dvdn;       /* [n][nc_eos]          partial molar volumes                                 -> frame = CoV */
} kg_eos_inputs;

/* Outputs; NULL = not wanted. nc = internal component count (2 for single-component models). */
typedef struct {
    double* viscosity;              /* [n]            Pa s */
    double* conductivity;           /* [n]            W/m K */
    double* conductivity_contribs;  /* [n][3]         translational, density, internal (thermal_conductivity_contributions "tdi") */
    double* D;                      /* [n][nc][nc] or [n][nc-1][nc-1] when do_compress; m^2/s */
    double* DT;                     /* [n][nc]        thermal diffusion coefficients, mol/(m s) */
    double* rdf;                    /* [n][nc][nc]    KineticGas::get_rdf */
    double* K;                      /* [n][nc]        get_K_factors */
    double* K_prime;                /* [n][nc]        get_K_prime_factors */
    double* cond_matrix;            /* [n][N nc][N nc] get_conductivity_matrix (cpp/KineticGas.cpp:271-318) */
    double* cond_vector;            /* [n][N nc]       get_conductivity_vector */
    double* visc_matrix;            /* [n][N nc][N nc] get_viscosity_matrix    (cpp/KineticGas.cpp:418-439) */
    double* visc_vector;            /* [n][N nc]       get_viscosity_vector */
    double* thermal_diffusion_ratio; /* [n][nc]       k_T, KineticGas::thermal_diffusion_ratio (cpp/KineticGas_properties.cpp:282-312); needs
                                                      KG_PROP_THERMAL_DIFFUSION. thermal_diffusion_factor is alpha_ij = k_T,i - k_T,j (:314-323) */
    double* soret_zarate;            /* [n][nc]       first nc-1 entries: KineticGas::soret_coefficient (:325-330), S_T = D^-1 D_T in the Zarate
                                                      frame; needs frame = KG_FOR_ZARATE and KG_PROP_DIFFUSION | KG_PROP_THERMAL_DIFFUSION */
} kg_transport_out;

/* Device-resident per-temperature tables = what the reference keeps in KineticGasCache (cpp/KineticGasCache.h): every
 * Omega^(l,r)_ij the requested properties need at order N (numerical quadrature, or the Fokin-Popov-Kalashnikov correlation
 * where MieKinGas::omega would use it), MTL/ETL, Barker-Henderson diameters and the bracket integrals H_ij, H_i, L_ij, L_i
 * (cpp/KineticGas.cpp:644-705). Replaces precompute_conductivity / precompute_viscosity (cpp/KineticGas_mthr.cpp:58-143). */
typedef struct kg_tcache kg_tcache;
/* coeff_T (optional) [nT][npairs][KG_MAX_TERMS]: per-temperature effective coefficients of T-dependent potentials.
 * Power-law sums other than plain Mie (or any model given coeff_T) use the extended Sutherland RDF of cpp/ExtendedSutherland.cpp:315-333
 * when opts->is_idealgas == 0 (all exponents must exceed 4). */
int kg_tcache_create(kg_model* m, const kg_transport_opts* opts, int nT, const double* T, const double* coeff_T, kg_tcache** out);
void kg_tcache_destroy(kg_tcache* c);
/* Shape of the Omega table: nlr entries (l[k], r[k]); pass NULL arrays to query nlr only. */
int kg_tcache_lr(const kg_tcache* c, int32_t* l, int32_t* r);
/* Host copies, to feed the reference's cache.omega.store / cache.mtl / cache.etl (cpp/KineticGasCache.h:25-28):
 * omega [nT][npairs][nlr] (0 where not needed), mtl/etl [nT][npairs] (NULL ok; zeros for ideal-gas models), is_correlation
 * [nT][npairs][nlr] (1 where the value came from the correlation branch). */
int kg_tcache_fetch(kg_tcache* c, double* omega, double* mtl, double* etl, uint8_t* is_correlation);

/* All requested properties for n state points. T, Vm [n]; x [n][ncomps] (ignored for single-component models, which use
 * {0.5,0.5} like pykingas/py_KineticGas.py:119-141). cache may be NULL: a temporary one is built for the distinct T.
 * status [n] (NULL ok). */
int kg_transport_batch(kg_model* m, kg_tcache* cache, int n, const double* T, const double* Vm, const double* x,
                       const kg_transport_opts* opts, const kg_eos_inputs* eos, const kg_transport_out* out, int32_t* status);

/* Stage timings of the last kg_tcache_create / kg_transport_batch on this model (bench accounting):
 *  out[0] ms in rdf_kernel + state_kernel of the last batch (CUDA events on the model stream)
 *  out[1] ms of the tcache Omega stage (numerical quadrature, host clock around the synchronised stage)
 *  out[2] ms of the tcache transfer-length stage      out[3] ms of correlation + BH diameters + bracket sums
 *  out[4] kernel launches of the last tcache build     out[5] kernel launches of the last batch */
int kg_last_timings(const kg_model* m, double* out8);
/* ---- device-resident Omega(T) interpolation tables with persistence ----
 * Replaces the per-object 0.1 K-bucket memo of cpp/KineticGasCache.h:92-164 for sweeps: built once from exact device quadrature at nodes
 * uniform in ln T (nodes_per_decade), segmented at the lattice-regime temperatures 20 / 35 / 50 / 70 K (cpp/Spherical.cpp:60-84) so that no
 * interpolation stencil crosses one, evaluated by a kernel (4-point Lagrange in ln T on ln Omega), saved to / loaded from a file that
 * carries a fingerprint of the model. ACCURACY CONTRACT: not the bit-parity of the batched path -- the interpolant follows the exact path to
 * the level of the reference's own mesh noise (its adaptive lattice changes refinement pattern with T: median ~1e-6, 99 % < 1e-3, worst
 * ~5e-3 relative at 64 nodes per decade; tests/test_gpu_table.py measures it). T outside
 * the table gives NaN. Models with temperature-dependent potentials are refused. */
typedef struct kg_omega_table kg_omega_table;
int kg_omega_table_create(kg_model* m, double Tmin, double Tmax, int nodes_per_decade, int nlr, const int32_t* l, const int32_t* r,
                          kg_omega_table** out);
int kg_omega_table_eval(kg_omega_table* t, int nT, const double* T, double* omega /* [nT][npairs][nlr], host */);
int kg_omega_table_eval_dev(kg_omega_table* t, int nT, const double* T_dev, double* omega_dev);   /* device pointers, async on the model stream */
int kg_omega_table_save(const kg_omega_table* t, const char* path);
int kg_omega_table_load(kg_model* m, const char* path, kg_omega_table** out);   /* fails if the fingerprint is another model's */
int kg_omega_table_lr(const kg_omega_table* t, int32_t* l, int32_t* r);         /* returns nlr */
/* out8: Tmin, Tmax, nodes, segments, npairs, nlr, bytes in HBM, 0 */
int kg_omega_table_info(const kg_omega_table* t, double* out8);
void kg_omega_table_destroy(kg_omega_table* t);


/* FP64 FMA-pipe peak probe: runs a register-resident DFMA chain kernel and returns the measured TFLOP/s
 * (2 flop per DFMA). Used as the roofline denominator (MEASURED_PEAKS.json has no FP64 figure). */
int kg_measure_fp64_peak(int device, double* tflops, double* sm_mhz_est);

/* ---------------------------------------------------------------------------------------------------------------------------
 * Quantum-mechanical collision integrals (cpp/Quantum.cpp). Reduced energies E_red = E / eps_ij.
 * kg_quantum_phase_shifts: Quantum::phase_shift / quantum_phase_shift / JKWB_phase_shift (cpp/Quantum.cpp:251-318,418-528) for n pairs
 *   (l[k], E_red[k]); which = 0 as the reference switches (JKWB for l >= 1000 or E_red > 120), 1 Numerov, 2 JKWB. */
int kg_quantum_phase_shifts(kg_model* m, int pair, int n, const int32_t* l, const double* E_red, int which, double* out, int32_t* status);
/* kg_quantum_cross_sections: Quantum::cross_section (cpp/Quantum.cpp:1176-1236): Q^(n)(E) [m^2], out[nE][nn], orders n[k] in 0..6.
 *   w_sym / w_anti weight the even-l / odd-l partial-wave sums (Quantum::get_symmetry_weights, :208-223: 0.5 / 0.5 for unlike particles,
 *   1 / 0 for spin-0 bosons such as He-4, (s+1)/(2s+1) and s/(2s+1) swapped for fermions). status[nE] may be NULL.
 * kg_quantum_omega: Quantum::quantum_omega (:1242-1259): Omega^(l,r)(T) [m^3/s ... the reference's units] for nT temperatures x nlr (l, r),
 *   omega[nT][nlr]; status[nT] may be NULL. Not available for temperature-dependent (Feynman-Hibbs) potentials. */
int kg_quantum_cross_sections(kg_model* m, int pair, double w_sym, double w_anti, int nn, const int32_t* n, int nE, const double* E_red,
                              double* out, int32_t* status);
int kg_quantum_omega(kg_model* m, int pair, double w_sym, double w_anti, int nT, const double* T, int nlr, const int32_t* l, const int32_t* r,
                     double* omega, int32_t* status);
/* kg_model_set_quantum: Quantum::set_quantum_active (cpp/Quantum.cpp:1520-1528) for the device tables: while active, kg_tcache_create (and
 * with it kg_transport_batch) takes every Omega^(l,r)_ij(T) from kg_quantum_omega with the weights w_sym[npairs], w_anti[npairs]. */
int kg_model_set_quantum(kg_model* m, int active, const double* w_sym, const double* w_anti);
/* Quantum::set_JKWB_limits (cpp/Quantum.h:128-130; defaults 120, 1000): JKWB phase shifts for E_red > E_limit or l >= l_limit. */
int kg_model_set_jkwb_limits(kg_model* m, double E_limit, int l_limit);

/* ---------------------------------------------------------------------------------------------------------------------------
 * EoS-derived inputs on the device. Replaces, for batches of state points, the calls the reference makes on its
 * thermopack EoS object one point at a time (cpp/eos_interface.h:121-181; pykingas/MieKinGas.py:55-65 constructs `saftvrmie`):
 *   chemical_potential_tv(T, V, n, dmudn)  ->  dmudn     [n][nc][nc]  J/mol^2  (total: ideal + residual), at n = x (one mole)
 *   idealenthalpysingle(T, j, dhdt)        ->  cp_ideal  [n][nc]      J/(mol K)
 *   pressure_tv(T, V, n)                   ->  pressure  [n]          Pa
 *   specific_volume(T, p, n, dvdn)         ->  dvdn      [n][nc]      m^3/mol  (dV/dn_i at constant T, p; CoV frame)
 *   specific_volume(T, p, x, VAPPH)        ->  kg_eos_specific_volume_batch: vapour-like root Vm [m^3/mol], status 1 = not converged
 * Model: SAFT-VR Mie (Lafitte et al. 2013) with thermopack's default options; m[nc] segments, pair matrices [nc][nc] sigma [m],
 * eps_div_k [K], la, lr (the caller applies its combining rules / k_ij). cp_type[nc], cp_par[nc][KG_EOS_CP_PARAMS], mw[nc] (g/mol):
 * thermopack's ideal-gas Cp correlation of each component (compdata.f90:14-26; types 10 and 11 are not available) -- may be NULL
 * when cp_ideal is never requested. Single-component models take nc = 1 (x may be NULL). Any output pointer may be NULL. Host
 * buffers in and out; the handle owns a stream and grow-only device buffers. */
#define KG_EOS_CP_PARAMS 21
typedef struct kg_eos kg_eos;
int kg_eos_saftvrmie_create(int device, int nc, const double* m, const double* sigma, const double* eps_div_k, const double* la,
                            const double* lr, const int32_t* cp_type, const double* cp_par, const double* mw, kg_eos** out);
void kg_eos_destroy(kg_eos* e);
const char* kg_eos_last_error(const kg_eos* e);
int kg_eos_tv_batch(kg_eos* e, int n, const double* T, const double* Vm, const double* x,
                    double* dmudn, double* cp_ideal, double* pressure, double* dvdn);
/* The same with the results LEFT ON THE DEVICE: dev[0..3] = device addresses of dmudn, cp_ideal, pressure, dvdn (NULL where `want` -- bit 0
 * dmudn, 1 cp_ideal, 2 pressure, 3 dvdn -- does not select them), inside the handle's buffers and valid until its next call. They can be
 * passed to kg_transport_batch as kg_eos_inputs: its EoS arrays may live in host or device memory (unified addressing). */
int kg_eos_tv_batch_dev(kg_eos* e, int n, const double* T, const double* Vm, const double* x, unsigned want, const double** dev);
int kg_eos_specific_volume_batch(kg_eos* e, int n, const double* T, const double* p, const double* x, double* Vm, int32_t* status);
double kg_eos_last_kernel_ms(const kg_eos* e);     /* summed kernel time of the last batch call (CUDA events) */

#ifdef __cplusplus
}
#endif
#endif
