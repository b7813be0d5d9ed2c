// kg_eos.cu -- EoS-derived inputs of the RET property code on the GPU: SAFT-VR Mie d mu_i/d n_j, p(T,V,n), dV/dn_i at
// constant (T,p), Vm(T,p,x) and the ideal-gas heat capacities, for batches of state points, behind the C-ABI of include/kgcuda.h.
//
// What it replaces: the reference asks a thermopack object for these numbers one state point at a time through Python
// (cpp/eos_interface.h:121-181: chemical_potential_tv -> dmudn, idealenthalpysingle -> Cp_id, specific_volume(T,p), pressure_tv;
// callers cpp/KineticGas_properties.cpp:459-500 and pykingas/py_KineticGas.py:642-715). After the transport kernels those host calls
// dominate a sweep. The model is kg_saft.cuh (the published SAFT-VR Mie of Lafitte et al. 2013 with thermopack's default
// options); derivatives come from hyper-dual evaluation of ONE generic Helmholtz function, so every output is exact to rounding.
//
// Kernels
//   eos_tpre_kernel : one thread per run of equal temperatures: Barker-Henderson diameters and the x0-power tables of that T
//   eos_hess_kernel : one thread per (state point, entry of the symmetric Hessian of F = A_res/RT in y = (n_1..n_nc, V)); each evaluates
//                     F once in hyper-dual numbers seeded on its two variables. The model constants sit in shared memory.
//   eos_out_kernel  : one thread per state point: dmudn, p, dV/dn from the Hessian; Cp_ideal per component
//   eos_vol_kernel  : one thread per state point: Newton iteration in ln V on p(T,V,x) = p from the ideal-gas volume (vapour-like root)
#include "kg_saft.cuh"
#include "../../include/kgcuda.h"
#include <cuda_runtime.h>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

using namespace kgsaft;

struct kg_eos {
    int device = 0, nc = 0;
    Model M;
    int cp_type[MAXC];
    double cp_par[MAXC][KG_EOS_CP_PARAMS];
    double mw[MAXC];                 // g/mol (the mass-based Cp correlations)
    Model* d_M = nullptr;
    int* d_cp_type = nullptr; double* d_cp_par = nullptr; double* d_mw = nullptr;
    cudaStream_t stream = nullptr;
    std::string err;
    double last_ms = 0;              // kernels of the last batch call (CUDA events)
};
static thread_local std::string g_eos_create_error;

#define EOS_OK(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { \
    set_error(std::string(#expr) + ": " + cudaGetErrorString(e_)); return -2; } } while (0)

namespace {

// one CTA per temperature run: the nc diameters side by side, then the nc x nc x 5 table entries side by side (kgsaft::init_T does the same
// serially; one thread per run took 0.5 ms for a 5-component model -- a fifth of the whole batch)
__global__ void __launch_bounds__(128) eos_tpre_kernel(const Model* __restrict__ M, int nruns, const double* __restrict__ Trun, TPre* __restrict__ pre) {
    const int r = blockIdx.x;
    if (r >= nruns) return;
    const int nc = M->nc, t = threadIdx.x;
    TPre& P = pre[r];
    const double T = Trun[r];
    if (t == 0) P.T = T;
    if (t < nc) P.d[t] = bh_diameter(M->p[t][t], T);
    __syncthreads();
    for (int idx = t; idx < nc * nc * 5; idx += blockDim.x) {
        const int k = idx % 5, j = (idx / 5) % nc, i = idx / (5 * nc);
        const PairC& q = M->p[i][j];
        const double dij = 0.5 * (P.d[i] + P.d[j]), x0 = q.sigma / dij, l = q.lam[k];
        if (k == 0) { P.d3[i][j] = dij * dij * dij; P.x0[i][j] = x0; }
        P.x0p[i][j][k] = pow(x0, l);
        P.I[i][j][k] = -(pow(x0, 3.0 - l) - 1.0) / (l - 3.0);
        P.J[i][j][k] = -(pow(x0, 4.0 - l) * (l - 3.0) - pow(x0, 3.0 - l) * (l - 4.0) - 1.0) / ((l - 3.0) * (l - 4.0));
    }
}

// entry e of the upper triangle of an nv x nv symmetric matrix, row-major: (a, b), a <= b
__device__ __forceinline__ void tri_index(int e, int nv, int& a, int& b) {
    a = 0;
    while (e >= nv - a) { e -= nv - a; a++; }
    b = a + e;
}

__global__ void __launch_bounds__(128) eos_hess_kernel(const Model* __restrict__ Mg, const TPre* __restrict__ pre, const int* __restrict__ run_of,
                                                       int n, int nv_used, const double* __restrict__ Vm, const double* __restrict__ x,
                                                       double* __restrict__ hess /*[n][nent]*/, double* __restrict__ gradV /*[n]*/) {
    __shared__ Model M;
    {
        const int words = sizeof(Model) / sizeof(double);
        const double* src = reinterpret_cast<const double*>(Mg);
        double* dst = reinterpret_cast<double*>(&M);
        for (int i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const int nc = M.nc, nent = nv_used * (nv_used + 1) / 2;
    const long long t = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long long) n * nent) return;
    const int sp = (int) (t / nent), e = (int) (t % nent);
    int a, b;
    tri_index(e, nv_used, a, b);
    HD y[MAXC + 1];
    for (int k = 0; k < nc; k++) y[k] = HD(nc == 1 ? 1.0 : x[(size_t) sp * nc + k]);
    y[nc] = HD(Vm[sp]);
    y[a].a = 1.0; y[b].b = 1.0;
    const HD f = helmholtz<HD>(M, pre[run_of[sp]], y[nc], y, TERM_ALL);
    hess[(size_t) sp * nent + e] = f.ab;
    if (gradV && a == nc && b == nc) gradV[sp] = f.a;
}

__device__ double cp_ideal(int type, const double* __restrict__ cp, double mw, double T) {
    // thermopack's correlation types (cpp/external/thermopack/src/compdata.f90:14-26, evaluated as in src/ideal.f90:203-309), J/(mol K)
    switch (type) {
        case 1: return (cp[0] + cp[1] * T + cp[2] * T * T + cp[3] * T * T * T) * 4.1868;
        case 2: return (cp[1] + T * (2.0 * cp[2] + T * (3.0 * cp[3] + T * (4.0 * cp[4] + T * 5.0 * cp[5])))) * mw;
        case 3: return (cp[0] + cp[1] * T * 1.8 + cp[2] * (T * 1.8) * (T * 1.8)) * 4.1868 * mw;
        case 4: return cp[0] + cp[1] * T + cp[2] * T * T + cp[3] * T * T * T;
        case 5: return (cp[0] + cp[1] * T + cp[2] * T * T + cp[3] * T * T * T + cp[4] / (T * T)) * mw;
        case 6: return (cp[0] + cp[1] * T + cp[2] * T * T + cp[3] * T * T * T + cp[4] * T * T * T * T) * mw;
        case 7: {
            const double u = cp[2] / T, w = cp[4] / T;
            const double s = u / sinh(u), c = w / cosh(w);
            return 1.0e-3 * (cp[0] + cp[1] * s * s + cp[3] * c * c);
        }
        case 8: return (cp[0] + cp[1] * T + cp[2] * T * T + cp[3] * T * T * T + cp[4] * T * T * T * T) * RGAS;
        case 9: return cp[0] + cp[1] * T + cp[2] / (T + cp[3]);
        case 12: { const double Ts = T * 1.0e-3; return cp[0] + cp[1] * Ts + cp[2] * Ts * Ts + cp[3] * Ts * Ts * Ts + cp[4] / (Ts * Ts); }
        case 13: {
            double v = cp[0];
            for (int j = 1; j <= (KG_EOS_CP_PARAMS - 1) / 2; j++) {
                if (fabs(cp[2 * j - 1]) < 1.0e-12) break;
                const double xx = cp[2 * j] / T;
                if (xx > 350.0) continue;
                const double ex = ::exp(xx);
                v += cp[2 * j - 1] * xx * xx * ex / ((ex - 1.0) * (ex - 1.0));
            }
            return v * RGAS;
        }
        default: return nan("");
    }
}

__global__ void eos_out_kernel(int nc, int n, int nv_used, const double* __restrict__ T, const double* __restrict__ Vm, const double* __restrict__ x,
                               const double* __restrict__ hess, const double* __restrict__ gradV,
                               const int* __restrict__ cp_type, const double* __restrict__ cp_par, const double* __restrict__ mw,
                               double* __restrict__ dmudn, double* __restrict__ cp_out, double* __restrict__ p_out, double* __restrict__ dvdn) {
    const int sp = blockIdx.x * blockDim.x + threadIdx.x;
    if (sp >= n) return;
    const double RT = RGAS * T[sp], V = Vm[sp];
    const int nent = nv_used * (nv_used + 1) / 2;
    auto H = [&](int a, int b) {                     // a <= b
        int off = 0;
        for (int r = 0; r < a; r++) off += nv_used - r;
        return hess[(size_t) sp * nent + off + (b - a)];
    };
    double ntot = 0.0;
    for (int i = 0; i < nc; i++) ntot += (nc == 1) ? 1.0 : x[(size_t) sp * nc + i];
    if (dmudn) for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) {
        const double xi = (nc == 1) ? 1.0 : x[(size_t) sp * nc + i];
        dmudn[((size_t) sp * nc + i) * nc + j] = RT * (H(i < j ? i : j, i < j ? j : i) + (i == j ? 1.0 / xi : 0.0));
    }
    if (p_out) p_out[sp] = ntot * RT / V - RT * gradV[sp];
    if (dvdn) {
        const double dpdv = -ntot * RT / (V * V) - RT * H(nc, nc);
        for (int i = 0; i < nc; i++) dvdn[(size_t) sp * nc + i] = -(RT / V - RT * H(i, nc)) / dpdv;
    }
    if (cp_out) for (int i = 0; i < nc; i++) cp_out[(size_t) sp * nc + i] = cp_ideal(cp_type[i], cp_par + (size_t) i * KG_EOS_CP_PARAMS, mw[i], T[sp]);
}

__global__ void __launch_bounds__(128) eos_vol_kernel(const Model* __restrict__ Mg, const TPre* __restrict__ pre, const int* __restrict__ run_of, int n,
                                                      const double* __restrict__ T, const double* __restrict__ p, const double* __restrict__ x,
                                                      double* __restrict__ Vm, int* __restrict__ status) {
    __shared__ Model M;
    {
        const int words = sizeof(Model) / sizeof(double);
        const double* src = reinterpret_cast<const double*>(Mg);
        double* dst = reinterpret_cast<double*>(&M);
        for (int i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const int sp = blockIdx.x * blockDim.x + threadIdx.x;
    if (sp >= n) return;
    const int nc = M.nc;
    const TPre& P = pre[run_of[sp]];
    HD y[MAXC + 1];
    double ntot = 0.0;
    for (int k = 0; k < nc; k++) { const double xk = (nc == 1) ? 1.0 : x[(size_t) sp * nc + k]; y[k] = HD(xk); ntot += xk; }
    const double RT = RGAS * T[sp], pt = p[sp];
    double V = ntot * RT / pt;                      // ideal gas: the vapour-like root is approached from above
    int st = 1;
    for (int it = 0; it < 60; it++) {
        y[nc] = HD(V, 1.0, 1.0, 0.0);
        const HD f = helmholtz<HD>(M, P, y[nc], y, TERM_ALL);
        const double pv = ntot * RT / V - RT * f.a, dpdv = -ntot * RT / (V * V) - RT * f.ab;
        // Newton in ln V; where the isotherm is not falling (inside a van der Waals loop) halve the volume instead
        double dlnV = (dpdv < 0.0) ? -(pv - pt) / (dpdv * V) : -0.7;
        dlnV = fmax(-0.7, fmin(0.7, dlnV));
        V *= ::exp(dlnV);
        if (fabs(dlnV) < 1e-13) { st = 0; break; }
    }
    Vm[sp] = V;
    if (status) status[sp] = st;
}

}   // namespace

// ------------------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------------------
struct EosScratch {                 // grow-only device buffers of one handle (a sweep calls with the same sizes again and again)
    void* p = nullptr; size_t cap = 0;
    cudaError_t need(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        const cudaError_t e = cudaMalloc(&p, bytes + bytes / 4);
        if (e == cudaSuccess) cap = bytes + bytes / 4;
        return e;
    }
};
struct kg_eos_bufs { EosScratch T, V, x, p, run_of, Trun, pre, hess, gradV, dmudn, cp, pout, dvdn, st; cudaEvent_t ev[2] = {nullptr, nullptr}; };
static std::vector<std::pair<kg_eos*, kg_eos_bufs*>>& eos_registry() { static std::vector<std::pair<kg_eos*, kg_eos_bufs*>> r; return r; }
static kg_eos_bufs* bufs_of(kg_eos* e) { for (auto& kv : eos_registry()) if (kv.first == e) return kv.second; return nullptr; }

extern "C" const char* kg_eos_last_error(const kg_eos* e) { return e ? e->err.c_str() : g_eos_create_error.c_str(); }

extern "C" int kg_eos_saftvrmie_create(int device, int nc, const double* m, const double* sigma, const double* eps_div_k, const double* la,
                                       const double* lr, const int32_t* cp_type, const double* cp_par, const double* mw, kg_eos** out) {
    auto set_error = [&](const std::string& s) { g_eos_create_error = s; };
    if (nc < 1 || nc > MAXC || !m || !sigma || !eps_div_k || !la || !lr || !out) { set_error("kg_eos_saftvrmie_create: invalid argument"); return -1; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { set_error("kg_eos_saftvrmie_create: no CUDA device (the GPU path has no CPU fallback)"); return -1; }
    if (device < 0 || device >= ndev) { set_error("kg_eos_saftvrmie_create: bad device ordinal"); return -1; }
    for (int i = 0; i < nc * nc; i++)
        if (!(sigma[i] > 0) || !(eps_div_k[i] > 0) || !(la[i] > 4.0) || !(lr[i] > la[i])) { set_error("kg_eos_saftvrmie_create: Mie parameters out of range (need lr > la > 4)"); return -1; }
    for (int i = 0; i < nc; i++) if (!(m[i] >= 1.0)) { set_error("kg_eos_saftvrmie_create: segment numbers must be >= 1"); return -1; }
    EOS_OK(cudaSetDevice(device));
    kg_eos* e = new kg_eos;
    e->device = device; e->nc = nc;
    memset(&e->M, 0, sizeof(Model));
    e->M.nc = nc;
    for (int i = 0; i < nc; i++) e->M.m[i] = m[i];
    for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) {
        PairC& q = e->M.p[i][j];
        q.sigma = sigma[i * nc + j]; q.epsk = eps_div_k[i * nc + j]; q.la = la[i * nc + j]; q.lr = lr[i * nc + j];
        init_pair(q);
    }
    for (int i = 0; i < MAXC; i++) {
        e->cp_type[i] = (cp_type && i < nc) ? cp_type[i] : 0;
        e->mw[i] = (mw && i < nc) ? mw[i] : 0.0;
        for (int k = 0; k < KG_EOS_CP_PARAMS; k++) e->cp_par[i][k] = (cp_par && i < nc) ? cp_par[i * KG_EOS_CP_PARAMS + k] : 0.0;
    }
    auto fail = [&](cudaError_t ce, const char* what) { set_error(std::string(what) + ": " + cudaGetErrorString(ce)); kg_eos_destroy(e); return -2; };
    kg_eos_bufs* b = new kg_eos_bufs;
    eos_registry().push_back({e, b});
    cudaError_t ce;
    if ((ce = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail(ce, "cudaStreamCreate");
    if ((ce = cudaMalloc(&e->d_M, sizeof(Model))) != cudaSuccess) return fail(ce, "cudaMalloc");
    if ((ce = cudaMalloc(&e->d_cp_type, MAXC * sizeof(int))) != cudaSuccess) return fail(ce, "cudaMalloc");
    if ((ce = cudaMalloc(&e->d_cp_par, sizeof(e->cp_par))) != cudaSuccess) return fail(ce, "cudaMalloc");
    if ((ce = cudaMalloc(&e->d_mw, sizeof(e->mw))) != cudaSuccess) return fail(ce, "cudaMalloc");
    if ((ce = cudaMemcpy(e->d_M, &e->M, sizeof(Model), cudaMemcpyHostToDevice)) != cudaSuccess) return fail(ce, "H2D");
    if ((ce = cudaMemcpy(e->d_cp_type, e->cp_type, MAXC * sizeof(int), cudaMemcpyHostToDevice)) != cudaSuccess) return fail(ce, "H2D");
    if ((ce = cudaMemcpy(e->d_cp_par, e->cp_par, sizeof(e->cp_par), cudaMemcpyHostToDevice)) != cudaSuccess) return fail(ce, "H2D");
    if ((ce = cudaMemcpy(e->d_mw, e->mw, sizeof(e->mw), cudaMemcpyHostToDevice)) != cudaSuccess) return fail(ce, "H2D");
    for (auto& ev : b->ev) if ((ce = cudaEventCreate(&ev)) != cudaSuccess) return fail(ce, "cudaEventCreate");
    *out = e;
    return 0;
}

extern "C" void kg_eos_destroy(kg_eos* e) {
    if (!e) return;
    auto& reg = eos_registry();
    kg_eos_bufs* b = nullptr;
    for (size_t k = 0; k < reg.size(); k++) if (reg[k].first == e) { b = reg[k].second; reg.erase(reg.begin() + k); break; }
    if (!b) return;                                        // unknown handle: already destroyed
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    for (EosScratch* s : {&b->T, &b->V, &b->x, &b->p, &b->run_of, &b->Trun, &b->pre, &b->hess, &b->gradV, &b->dmudn, &b->cp, &b->pout, &b->dvdn, &b->st})
        if (s->p) cudaFree(s->p);
    for (auto& ev : b->ev) if (ev) cudaEventDestroy(ev);
    delete b;
    for (void* q : {(void*) e->d_M, (void*) e->d_cp_type, (void*) e->d_cp_par, (void*) e->d_mw}) if (q) cudaFree(q);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

// runs of equal temperature: run_of[sp], Trun[]
static void find_runs(int n, const double* T, std::vector<int>& run_of, std::vector<double>& Trun) {
    run_of.resize(n); Trun.clear();
    for (int i = 0; i < n; i++) {
        if (i == 0 || T[i] != T[i - 1]) Trun.push_back(T[i]);
        run_of[i] = (int) Trun.size() - 1;
    }
}

// shared front half of the two batch calls: inputs to the device, temperature runs, per-run tables
static int eos_stage(kg_eos* e, kg_eos_bufs* b, int n, const double* T, const double* second, const double* x) {
    auto set_error = [&](const std::string& s) { e->err = s; };
    const int nc = e->nc;
    cudaStream_t s = e->stream;
    std::vector<int> run_of; std::vector<double> Trun;
    find_runs(n, T, run_of, Trun);
    const int nruns = (int) Trun.size();
    if ((size_t) nruns * sizeof(TPre) > ((size_t) 2 << 30)) { set_error("kg_eos: more than 4e5 distinct temperature runs in one call; split the batch"); return -1; }
    EOS_OK(b->T.need(n * sizeof(double))); EOS_OK(b->V.need(n * sizeof(double)));
    EOS_OK(b->x.need((size_t) n * nc * sizeof(double)));
    EOS_OK(b->run_of.need(n * sizeof(int))); EOS_OK(b->Trun.need(nruns * sizeof(double))); EOS_OK(b->pre.need((size_t) nruns * sizeof(TPre)));
    EOS_OK(cudaMemcpyAsync(b->T.p, T, n * sizeof(double), cudaMemcpyHostToDevice, s));
    EOS_OK(cudaMemcpyAsync(b->V.p, second, n * sizeof(double), cudaMemcpyHostToDevice, s));
    if (nc > 1) EOS_OK(cudaMemcpyAsync(b->x.p, x, (size_t) n * nc * sizeof(double), cudaMemcpyHostToDevice, s));
    EOS_OK(cudaMemcpyAsync(b->run_of.p, run_of.data(), n * sizeof(int), cudaMemcpyHostToDevice, s));
    EOS_OK(cudaMemcpyAsync(b->Trun.p, Trun.data(), nruns * sizeof(double), cudaMemcpyHostToDevice, s));
    EOS_OK(cudaStreamSynchronize(s));                        // run_of / Trun are stack-lifetime host vectors
    EOS_OK(cudaEventRecord(b->ev[0], s));
    eos_tpre_kernel<<<nruns, 128, 0, s>>>(e->d_M, nruns, (const double*) b->Trun.p, (TPre*) b->pre.p);
    EOS_OK(cudaGetLastError());
    return 0;
}

static int eos_tv(kg_eos* e, int n, const double* T, const double* Vm, const double* x, double* dmudn, double* cp_ideal_out, double* pressure,
                  double* dvdn, const double** dev_out);
extern "C" int kg_eos_tv_batch(kg_eos* e, int n, const double* T, const double* Vm, const double* x,
                               double* dmudn, double* cp_ideal_out, double* pressure, double* dvdn) {
    return eos_tv(e, n, T, Vm, x, dmudn, cp_ideal_out, pressure, dvdn, nullptr);
}
// Results left on the device: dev[0..3] receive the addresses of dmudn, cp_ideal, pressure, dvdn inside the handle's buffers (NULL for
// what `want` does not select: bit 0 dmudn, 1 cp_ideal, 2 pressure, 3 dvdn), valid until the next call on this handle. The stream is
// synchronised before return, so the arrays can be handed to kg_transport_batch as kg_eos_inputs right away.
extern "C" int kg_eos_tv_batch_dev(kg_eos* e, int n, const double* T, const double* Vm, const double* x, unsigned want, const double** dev) {
    if (!e) return -1;
    if (!dev) { e->err = "kg_eos_tv_batch_dev: invalid argument"; return -1; }
    double dummy = 0;       // non-NULL selects an output; nothing is written through these in device-output mode
    return eos_tv(e, n, T, Vm, x, (want & 1u) ? &dummy : nullptr, (want & 2u) ? &dummy : nullptr, (want & 4u) ? &dummy : nullptr,
                  (want & 8u) ? &dummy : nullptr, dev);
}
static int eos_tv(kg_eos* e, int n, const double* T, const double* Vm, const double* x, double* dmudn, double* cp_ideal_out, double* pressure,
                  double* dvdn, const double** dev_out) {
    if (!e) return -1;
    auto set_error = [&](const std::string& s) { e->err = s; };
    kg_eos_bufs* b = bufs_of(e);
    const int nc = e->nc;
    if (!b || n < 0 || !T || !Vm || (nc > 1 && !x)) { set_error("kg_eos_tv_batch: invalid argument"); return -1; }
    if (n == 0) return 0;
    if (cp_ideal_out) for (int i = 0; i < nc; i++) {
        const int t = e->cp_type[i];
        if (!(t >= 1 && t <= 13) || t == 10 || t == 11) { set_error("kg_eos_tv_batch: ideal-gas Cp correlation type " + std::to_string(t) + " of component " + std::to_string(i) + " is not available on the device"); return -1; }
    }
    EOS_OK(cudaSetDevice(e->device));
    cudaStream_t s = e->stream;
    { const int rc = eos_stage(e, b, n, T, Vm, x); if (rc) return rc; }
    const bool need_hess = dmudn || pressure || dvdn;
    const int nv_used = (pressure || dvdn) ? nc + 1 : nc, nent = nv_used * (nv_used + 1) / 2;
    if (need_hess) {
        EOS_OK(b->hess.need((size_t) n * nent * sizeof(double))); EOS_OK(b->gradV.need(n * sizeof(double)));
        const long long nthreads = (long long) n * nent;
        eos_hess_kernel<<<(unsigned) ((nthreads + 127) / 128), 128, 0, s>>>(e->d_M, (const TPre*) b->pre.p, (const int*) b->run_of.p, n, nv_used,
            (const double*) b->V.p, (const double*) b->x.p, (double*) b->hess.p, (nv_used == nc + 1) ? (double*) b->gradV.p : nullptr);
        EOS_OK(cudaGetLastError());
    }
    if (dmudn) EOS_OK(b->dmudn.need((size_t) n * nc * nc * sizeof(double)));
    if (cp_ideal_out) EOS_OK(b->cp.need((size_t) n * nc * sizeof(double)));
    if (pressure) EOS_OK(b->pout.need(n * sizeof(double)));
    if (dvdn) EOS_OK(b->dvdn.need((size_t) n * nc * sizeof(double)));
    eos_out_kernel<<<(n + 127) / 128, 128, 0, s>>>(nc, n, nv_used, (const double*) b->T.p, (const double*) b->V.p, (const double*) b->x.p,
        (const double*) b->hess.p, (const double*) b->gradV.p, e->d_cp_type, e->d_cp_par, e->d_mw,
        dmudn ? (double*) b->dmudn.p : nullptr, cp_ideal_out ? (double*) b->cp.p : nullptr, pressure ? (double*) b->pout.p : nullptr,
        dvdn ? (double*) b->dvdn.p : nullptr);
    EOS_OK(cudaGetLastError());
    EOS_OK(cudaEventRecord(b->ev[1], s));
    if (dev_out) {
        dev_out[0] = dmudn ? (const double*) b->dmudn.p : nullptr; dev_out[1] = cp_ideal_out ? (const double*) b->cp.p : nullptr;
        dev_out[2] = pressure ? (const double*) b->pout.p : nullptr; dev_out[3] = dvdn ? (const double*) b->dvdn.p : nullptr;
        EOS_OK(cudaStreamSynchronize(s));
        float msd = 0; if (cudaEventElapsedTime(&msd, b->ev[0], b->ev[1]) == cudaSuccess) e->last_ms = msd;
        return 0;
    }
    if (dmudn) EOS_OK(cudaMemcpyAsync(dmudn, b->dmudn.p, (size_t) n * nc * nc * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (cp_ideal_out) EOS_OK(cudaMemcpyAsync(cp_ideal_out, b->cp.p, (size_t) n * nc * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (pressure) EOS_OK(cudaMemcpyAsync(pressure, b->pout.p, n * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (dvdn) EOS_OK(cudaMemcpyAsync(dvdn, b->dvdn.p, (size_t) n * nc * sizeof(double), cudaMemcpyDeviceToHost, s));
    EOS_OK(cudaStreamSynchronize(s));
    float ms = 0; if (cudaEventElapsedTime(&ms, b->ev[0], b->ev[1]) == cudaSuccess) e->last_ms = ms;
    return 0;
}

extern "C" int kg_eos_specific_volume_batch(kg_eos* e, int n, const double* T, const double* p, const double* x, double* Vm, int32_t* status) {
    if (!e) return -1;
    auto set_error = [&](const std::string& s) { e->err = s; };
    kg_eos_bufs* b = bufs_of(e);
    const int nc = e->nc;
    if (!b || n < 0 || !T || !p || !Vm || (nc > 1 && !x)) { set_error("kg_eos_specific_volume_batch: invalid argument"); return -1; }
    if (n == 0) return 0;
    for (int i = 0; i < n; i++) if (!(p[i] > 0) || !(T[i] > 0)) { set_error("kg_eos_specific_volume_batch: T and p must be positive"); return -1; }
    EOS_OK(cudaSetDevice(e->device));
    cudaStream_t s = e->stream;
    { const int rc = eos_stage(e, b, n, T, p, x); if (rc) return rc; }
    EOS_OK(b->pout.need(n * sizeof(double))); EOS_OK(b->st.need(n * sizeof(int)));
    eos_vol_kernel<<<(n + 127) / 128, 128, 0, s>>>(e->d_M, (const TPre*) b->pre.p, (const int*) b->run_of.p, n, (const double*) b->T.p,
        (const double*) b->V.p, (const double*) b->x.p, (double*) b->pout.p, (int*) b->st.p);
    EOS_OK(cudaGetLastError());
    EOS_OK(cudaEventRecord(b->ev[1], s));
    EOS_OK(cudaMemcpyAsync(Vm, b->pout.p, n * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (status) EOS_OK(cudaMemcpyAsync(status, b->st.p, n * sizeof(int), cudaMemcpyDeviceToHost, s));
    EOS_OK(cudaStreamSynchronize(s));
    float ms = 0; if (cudaEventElapsedTime(&ms, b->ev[0], b->ev[1]) == cudaSuccess) e->last_ms = ms;
    return 0;
}

extern "C" double kg_eos_last_kernel_ms(const kg_eos* e) { return e ? e->last_ms : 0.0; }
