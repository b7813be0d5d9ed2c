// kg_internal.cuh -- definitions shared by the translation units of libkgcuda.so (not part of the C-ABI).
#pragma once
#include "kg_device.cuh"
#include "../../include/kgcuda.h"
#include <string>
#include <mutex>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <map>
#include <unordered_map>
#include <algorithm>

#define CUDA_OK(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { \
    set_error(std::string(#expr) + ": " + cudaGetErrorString(e_)); return -2; } } while (0)

// Debug aid (KG_TRACE_ERRORS=1): report the entry point that leaves a CUDA error behind on exit
struct KgTrace {
    const char* name;
    explicit KgTrace(const char* n) : name(n) {
        static const bool on = getenv("KG_TRACE_ERRORS") != nullptr;
        if (!on) { name = nullptr; return; }
        const cudaError_t e = cudaPeekAtLastError();
        if (e != cudaSuccess) fprintf(stderr, "[kg trace] error present on ENTRY of %s: %s\n", n, cudaGetErrorString(e));
    }
    ~KgTrace() {
        if (!name) return;
        const cudaError_t e = cudaPeekAtLastError();
        if (e != cudaSuccess) fprintf(stderr, "[kg trace] error present on EXIT of %s: %s\n", name, cudaGetErrorString(e));
    }
};
#define KG_TRACE() KgTrace kg_trace_(__func__)

struct kg_model {
    int device = 0, ncomps = 0, npairs = 0, is_singlecomp = 0, ntpad = 2;
    std::vector<double> mass;
    std::vector<kg_potential_desc> descs;
    std::vector<kg::PairParams> pairs;
    // generic (non-power-law) models: ntpad == 1, node rows { w/2u, 1/u^2, 1/u }
    bool generic = false, gen_T_dependent = false;
    bool coeff_T_required = false;      // power-sum model whose coefficients depend on T (kg_model_set_temperature_dependent)
    std::vector<kg_potential_desc_ex> gdescs;
    kg::GenPot* d_gpots = nullptr;      // base potentials, one per pair
    int coeff_stride = KG_MAX_TERMS;    // doubles per (T, pair) block of a coefficient pool (GP_DOUBLES for generic models)
    std::vector<double*> d_rows;        // per pair: [KMAX][2+ntpad]
    double** d_rows_by_pair = nullptr;
    kg::PairParams* d_pairs = nullptr;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    cudaStream_t aux_stream[2] = {nullptr, nullptr};   // chunk pipeline of kg_transport_batch (created on first use)
    cudaEvent_t aux_done[2] = {nullptr, nullptr};
    std::string err;
    int sm_count = 148;
    // stage timings of the last kg_tcache_create / kg_transport_batch on this model (kg_last_timings)
    double timings[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int tl_launches = 0;       // kernel launches of the last kg_transfer_lengths_batch
    void set_error(const std::string& s) { err = s; }

    // Device-memory pool of this model. A batch call allocates a few dozen buffers whose sizes repeat from call to call;
    // cudaMalloc/cudaFree are driver calls that serialise across the processes of a multi-GPU job, so freed blocks are
    // kept and handed out again (smallest cached block within 25 % of the request). Everything a model does runs on its
    // stream, and dfree() waits for that stream before the block can be reused by a host-synchronous copy.
    std::multimap<size_t, void*> pool_free;
    std::unordered_map<void*, size_t> pool_live;
    size_t pool_cached = 0;
    bool quantum_active = false;        // kg_model_set_quantum: device tables take Omega from the partial-wave engine (kg_quantum_omega)
    std::vector<double> q_wsym, q_wanti;   // per pair: weights of the even-l / odd-l sums
    double q_jkwb_E = 120.0; int q_jkwb_l = 1000;   // Quantum::set_JKWB_limits (defaults cpp/Quantum.h:153-154)
    bool tl_warp_search = false;        // KG_TL_WARP_SEARCH=1: one warp per b_mid / b_max search (comparison runs; measured slower)
    int* h_pinned = nullptr;            // page-locked host memory: 16 ints of the model + 32 slots of 8 ints handed to plans (run flags)
    unsigned pinned_used = 0;           // slot bitmap (cudaMallocHost / cudaFreeHost per plan cost ~1 ms each and synchronise the device)
    int* pinned_slot_acquire(int& slot) {
        for (slot = 0; slot < 32; slot++) if (!(pinned_used >> slot & 1u)) { pinned_used |= 1u << slot; return h_pinned + 16 + 8 * slot; }
        slot = -1;
        return nullptr;
    }
    void pinned_slot_release(int slot) { if (slot >= 0) pinned_used &= ~(1u << slot); }
    cudaError_t dmalloc_bytes(void** out, size_t bytes) {
        bytes = (std::max<size_t>(bytes, 1) + 511) & ~(size_t) 511;
        auto f = pool_free.lower_bound(bytes);
        if (f != pool_free.end() && f->first <= bytes + bytes / 4 + 65536) {
            *out = f->second; pool_live[f->second] = f->first; pool_cached -= f->first; pool_free.erase(f);
            return cudaSuccess;
        }
        cudaError_t e = cudaMalloc(out, bytes);
        if (e != cudaSuccess) { cudaGetLastError(); pool_release(); e = cudaMalloc(out, bytes); }
        if (e == cudaSuccess) pool_live[*out] = bytes;
        return e;
    }
    void dfree(void* p) {
        if (!p) return;
        auto f = pool_live.find(p);
        if (f == pool_live.end()) { cudaFree(p); return; }
        if (stream) cudaStreamSynchronize(stream);
        const size_t bytes = f->second;
        pool_live.erase(f);
        if (pool_cached + bytes > ((size_t) 48 << 30)) { cudaFree(p); return; }
        pool_free.emplace(bytes, p); pool_cached += bytes;
    }
    void pool_release() {
        for (auto& kv : pool_free) cudaFree(kv.second);
        pool_free.clear(); pool_cached = 0;
    }
};

// kg_genpot.cu: effective (Feynman-Hibbs folded) generic potential of one pair at temperature T
int kg_genpot_build(const kg_potential_desc_ex& d, double T, kg::GenPot& out, std::string& err);
// Per-(T, pair) coefficient pool of a batch: the caller's coeff_T for power-sum models, the folded potentials for
// T-dependent generic models (built into buf), NULL otherwise. Layout [nT][npairs][m->coeff_stride].
int kg_effective_coeff(kg_model* m, int nT, const double* T, const double* coeff_T_user, std::vector<double>& buf, const double** out);

// Live plans / caches, process wide. A model destroys its remaining children, and destroying a child whose model is already
// gone (garbage collectors finalise in any order) is a no-op instead of a use-after-free.
struct KgRegistry {
    std::mutex mu;
    std::map<void*, kg_model*> owner;       // child handle -> model
    static KgRegistry& get() { static KgRegistry r; return r; }
    void add(void* child, kg_model* m) { std::lock_guard<std::mutex> g(mu); owner[child] = m; }
    bool remove(void* child) { std::lock_guard<std::mutex> g(mu); return owner.erase(child) > 0; }     // false: unknown / already destroyed
    std::vector<void*> children_of(kg_model* m) {
        std::lock_guard<std::mutex> g(mu);
        std::vector<void*> v;
        for (auto& kv : owner) if (kv.second == m) v.push_back(kv.first);
        return v;
    }
};
void kg_destroy_child(void* child);          // kg_omega.cu: dispatches to kg_omega_plan_destroy / kg_tcache_destroy

template<class T> inline cudaError_t kg_dmalloc(kg_model* m, T** out, size_t bytes) { return m->dmalloc_bytes((void**) out, bytes); }

