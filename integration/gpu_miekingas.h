// integration/gpu_miekingas.h -- the reference-side half of the drop-in boundary, as real code.
//
// `GpuMieKinGas` derives from the reference's UNMODIFIED `MieKinGas` (cpp/MieKinGas.h) and overrides exactly the virtuals the
// reference declares "to facilitate multithreading" (cpp/KineticGas.h:277-280) plus the transfer-length entry:
//
//   precompute_conductivity / precompute_viscosity   replace the std::thread fan-out of cpp/KineticGas_mthr.cpp:58-143 by ONE
//                                                    device batch (kg_tcache_create) and store the device-computed collision
//                                                    integrals under the reference's own keys:
//                                                    cache.omega.store(get_omega_point(i,j,l,r,T), value)
//                                                    (cpp/KineticGasCache.h:25-28, cpp/KineticGas.h:244-246)
//   precompute_omega                                 same, for an explicit (i,j,l,r) list (kg_omega_plan_create_masked)
//   model_mtl / model_etl                            replace Spherical::get_transfer_length (EWCA, cpp/transfer_lengths.cpp:38-80)
//                                                    by kg_transfer_lengths_batch; stored in cache.mtl / cache.etl
//
// Everything else -- bracket integrals, RDF, K-factors, matrix assembly, Eigen solves, property formulas, the pybind11 surface
// of cpp/bindings.cpp and pykingas above it -- runs as the stock reference code: after the feed every `omega()`,
// `get_mtl()`, `get_etl()` is a cache hit, so no CPU quadrature is ever started.
//
// Cache-key semantics are the reference's: values are stored only where the key is ABSENT (ThreadSafeMap::store_if_absent
// behaviour of compute_if_absent), so two temperatures in one 0.1 K bucket share the first one's collision integrals exactly
// as in the stock build.
//
// Build: compiled against the reference headers where they lie and linked with libkgcuda.so (integration/Makefile).
// No file under the reference tree is modified.
#pragma once
#include "MieKinGas.h"
#include <atomic>

struct kg_model;

class GpuMieKinGas : public MieKinGas {
public:
    GpuMieKinGas(vector1d mole_weights, vector2d sigma, vector2d eps, vector2d la, vector2d lr, bool is_idealgas, bool is_singlecomp,
                 int device = 0);
    ~GpuMieKinGas() override;
    GpuMieKinGas(const GpuMieKinGas&) = delete;
    GpuMieKinGas& operator=(const GpuMieKinGas&) = delete;

    // cpp/KineticGas.h:277-280
    void precompute_conductivity(int N, double T, double rho, bool precompute_etl = true) override;
    void precompute_viscosity(int N, double T, double rho) override;
    void precompute_omega(const std::vector<int>& i_vec, const std::vector<int>& j_vec, const std::vector<int>& l_vec,
                          const std::vector<int>& r_vec, double T) override;
    // cpp/transfer_lengths.cpp:24-36
    vector2d model_mtl(double rho, double T, const vector1d& x) override;
    vector2d model_etl(double rho, double T, const vector1d& x) override;

    // Additive batched entry points (bound as `omega_table` / `transport_batch` in integration/bindings_addon.cpp)
    // omega_table: [nT][npairs][nlr] by quadrature on the device; also feeds the cache.
    std::vector<double> omega_table(const std::vector<double>& T, const std::vector<int>& l, const std::vector<int>& r);
    // transport_batch: n state points in one device pass (kg_tcache_create on the distinct temperatures + kg_transport_batch).
    // x is [n][Ncomps] (ignored for single-component models), dmudn [n][nce][nce] and cp_ideal [n][nce] are the EoS-derived inputs
    // (nce = 1 for single-component models); either may be empty, the properties that need it are then skipped.
    struct BatchResult { std::vector<double> viscosity, conductivity, D, DT; int nD = 0; };
    BatchResult transport_batch(const std::vector<double>& T, const std::vector<double>& Vm, const std::vector<double>& x, int N,
                                const std::vector<double>& dmudn, const std::vector<double>& cp_ideal);

    // Test hooks: CPU theta-integrand evaluations (every CPU chi / w_integral / transfer-length quadrature goes through it),
    // device feeds, and values stored into the reference cache.
    long cpu_theta_integrand_calls() const { return n_cpu_theta.load(); }
    long device_feeds() const { return n_feeds.load(); }
    long omega_values_stored() const { return n_stored.load(); }
    void reset_counters() { n_cpu_theta = 0; n_feeds = 0; n_stored = 0; }
    void clear_cache() { cache.clear(); }

private:
    // cpp/Spherical.cpp:327-329, restated only to count the calls (private virtual in Spherical: overridable, not callable)
    double theta_integrand(int i, int j, double T, double r, double g, double b) const override;
    void feed(int N, double T, unsigned props);
    void feed_transfer_lengths(double rho, double T, const vector1d& x);
    [[noreturn]] void fail(const char* what) const;

    kg_model* gpu = nullptr;
    mutable std::atomic<long> n_cpu_theta{0};
    std::atomic<long> n_feeds{0}, n_stored{0};
};
