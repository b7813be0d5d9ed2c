// integration/gpu_miekingas.cpp -- see gpu_miekingas.h. The reference-side binding of libkgcuda.so (include/kgcuda.h).
#include "gpu_miekingas.h"
#include "kgcuda.h"
#include <stdexcept>
#include <algorithm>

namespace {
int pair_count(int nc) { return nc * (nc + 1) / 2; }
}

void GpuMieKinGas::fail(const char* what) const {
    // the C-ABI carries the reference's own error texts ("get_R reached max iter!", "Unable to converge bmid!", ...)
    const char* e = kg_last_error(gpu);
    throw std::runtime_error(std::string(e && *e ? e : what));
}

GpuMieKinGas::GpuMieKinGas(vector1d mole_weights, vector2d sigma_, vector2d eps_, vector2d la_, vector2d lr_, bool is_idealgas_,
                           bool is_singlecomp_, int device)
    : MieKinGas(mole_weights, sigma_, eps_, la_, lr_, is_idealgas_, is_singlecomp_) {
    // one kg_model per KineticGas object; a single-component model is ONE device component (cpp/Spherical.cpp:38-40 forces i=j=0)
    const int nc = is_singlecomp ? 1 : (int) Ncomps;
    std::vector<kg_potential_desc> d;
    for (int i = 0; i < nc; i++) for (int j = i; j < nc; j++) {
        kg_potential_desc p{};
        p.kind = KG_POT_POWER_SUM; p.nterms = 2;
        p.sigma = sigma[i][j]; p.eps = eps[i][j];
        p.C[0] = C[i][j]; p.lam[0] = lr[i][j];          // u = C eps [(sigma/r)^lr - (sigma/r)^la], cpp/MieKinGas.h:26-28
        p.C[1] = -C[i][j]; p.lam[1] = la[i][j];
        d.push_back(p);
    }
    if (kg_model_create(d.data(), nc, m.data(), is_singlecomp ? 1 : 0, device, &gpu))
        throw std::runtime_error(std::string("GpuMieKinGas: ") + kg_last_error(nullptr));
}

GpuMieKinGas::~GpuMieKinGas() { kg_model_destroy(gpu); }

double GpuMieKinGas::theta_integrand(int i, int j, double T, double r, double g, double b) const {
    n_cpu_theta.fetch_add(1, std::memory_order_relaxed);
    return pow((pow(r, 4) / pow(b, 2)) * (1.0 - potential(i, j, r) / (BOLTZMANN * T * pow(g, 2))) - pow(r, 2), -0.5);
}

// One device batch for everything the order-N property needs at temperature T: kg_tcache_create picks the (l,r) set exactly as
// cpp/KineticGas_mthr.cpp:62-73 (conductivity / diffusion) and :102-125 (viscosity) do, integrates where MieKinGas::omega would
// integrate (correlation branch excluded, cpp/MieKinGas.cpp:81-93) and, for non-ideal models, evaluates MTL and ETL.
void GpuMieKinGas::feed(int N, double T, unsigned props) {
    kg_transport_opts o{};
    o.N = N; o.is_idealgas = is_idealgas ? 1 : 0; o.omega_correlation_active = omega_correlation_active ? 1 : 0;
    o.frame_of_reference = KG_FOR_CON; o.dependent_idx = -1; o.solvent_idx = -1; o.do_compress = 1; o.properties = props;
    const int nc = is_singlecomp ? 1 : (int) Ncomps, np = pair_count(nc);
    const StatePoint tl_point = get_transfer_length_point(0., T, vector1d(Ncomps, 0.));
    const bool tl_on_device = !is_idealgas && transfer_length_model_id == TransferLengthModel::EWCA;
    if (!tl_on_device) o.is_idealgas = 1;                  // other transfer-length models stay on the reference's CPU code
    kg_tcache* c = nullptr;
    if (kg_tcache_create(gpu, &o, 1, &T, nullptr, &c)) fail("kg_tcache_create failed");
    n_feeds++;
    const int nlr = kg_tcache_lr(c, nullptr, nullptr);
    std::vector<int32_t> l(nlr), r(nlr);
    kg_tcache_lr(c, l.data(), r.data());
    std::vector<double> om((size_t) np * nlr), mtl(np), etl(np);
    std::vector<uint8_t> corr((size_t) np * nlr);
    const int rc = kg_tcache_fetch(c, om.data(), mtl.data(), etl.data(), corr.data());
    kg_tcache_destroy(c);
    if (rc) fail("kg_tcache_fetch failed");
    int p = 0;
    for (int i = 0; i < nc; i++) for (int j = i; j < nc; j++, p++) for (int k = 0; k < nlr; k++) {
        const double v = om[(size_t) p * nlr + k];
        if (corr[(size_t) p * nlr + k] || v == 0.0) continue;      // correlation values: MieKinGas::omega recomputes them, never cached
        const OmegaPoint point = get_omega_point(i, j, l[k], r[k], T);
        if (!cache.omega.has(point)) { cache.omega.store(point, v); n_stored++; }   // first temperature of a 0.1 K bucket wins, as in the stock build
    }
    if (tl_on_device) {
        vector2d M(Ncomps, vector1d(Ncomps, 0.)), E(Ncomps, vector1d(Ncomps, 0.));
        if (is_singlecomp) { M.assign(Ncomps, vector1d(Ncomps, mtl[0])); E.assign(Ncomps, vector1d(Ncomps, etl[0])); }
        else {
            p = 0;
            for (int i = 0; i < nc; i++) for (int j = i; j < nc; j++, p++) { M[i][j] = M[j][i] = mtl[p]; E[i][j] = E[j][i] = etl[p]; }
        }
        if (!cache.mtl.has(tl_point)) cache.mtl.store(tl_point, M);
        if (!cache.etl.has(tl_point)) cache.etl.store(tl_point, E);
    }
}

void GpuMieKinGas::precompute_conductivity(int N, double T, double rho, bool precompute_etl) {
    feed(N, T, KG_PROP_DIFFUSION | KG_PROP_THERMAL_DIFFUSION);
}

void GpuMieKinGas::precompute_viscosity(int N, double T, double rho) {
    feed(N, T, KG_PROP_VISCOSITY);
}

void GpuMieKinGas::precompute_omega(const std::vector<int>& i_vec, const std::vector<int>& j_vec, const std::vector<int>& l_vec,
                                    const std::vector<int>& r_vec, double T) {
    // explicit work list: the distinct (l,r) become the columns of one masked plan, (pair, l, r) not requested are skipped
    const int nc = is_singlecomp ? 1 : (int) Ncomps, np = pair_count(nc);
    std::vector<int32_t> l, r;
    auto column = [&](int ll, int rr) {
        for (size_t k = 0; k < l.size(); k++) if (l[k] == ll && r[k] == rr) return (int) k;
        l.push_back(ll); r.push_back(rr);
        return (int) l.size() - 1;
    };
    auto pair_index = [&](int i, int j) { if (is_singlecomp) return 0; if (i > j) std::swap(i, j); return i * nc - i * (i - 1) / 2 + (j - i); };
    std::vector<std::array<int, 4>> todo;
    for (size_t k = 0; k < i_vec.size(); k++) {
        const int i = is_singlecomp ? 0 : i_vec[k], j = is_singlecomp ? 0 : j_vec[k];
        const bool corr = omega_correlation_active && (l_vec[k] <= 2) && (r_vec[k] >= l_vec[k]) && (r_vec[k] <= 3) && (abs(la[i][j] - 6.) < 1e-10)
                          && (T * BOLTZMANN / eps[i][j] > 0.4);                   // cpp/MieKinGas.cpp:82-88
        if (corr || cache.omega.has(get_omega_point(i, j, l_vec[k], r_vec[k], T))) continue;
        todo.push_back({i, j, pair_index(i, j), column(l_vec[k], r_vec[k])});
    }
    if (todo.empty()) return;
    const int nlr = (int) l.size();
    std::vector<uint8_t> need((size_t) np * nlr, 0);
    for (auto& t : todo) need[(size_t) t[2] * nlr + t[3]] = 1;
    std::vector<double> om((size_t) np * nlr);
    std::vector<int32_t> st(np);
    kg_omega_plan* plan = nullptr;
    if (kg_omega_plan_create_masked(gpu, 1, &T, nlr, l.data(), r.data(), nullptr, need.data(), &plan)) fail("kg_omega_plan_create failed");
    int rc = kg_omega_plan_run(plan);
    if (!rc) rc = kg_omega_plan_fetch(plan, om.data(), nullptr, st.data());
    kg_omega_plan_destroy(plan);
    if (rc) fail("kg_omega_plan_run failed");
    n_feeds++;
    for (int s : st) if (s & KG_STATUS_GETR_MAXITER) throw std::runtime_error("get_R reached max iter!");     // cpp/Spherical.cpp:432
    for (auto& t : todo) {
        const OmegaPoint point = get_omega_point(t[0], t[1], l[t[3]], r[t[3]], T);
        if (!cache.omega.has(point)) { cache.omega.store(point, om[(size_t) t[2] * nlr + t[3]]); n_stored++; }
    }
}

void GpuMieKinGas::feed_transfer_lengths(double rho, double T, const vector1d& x) {
    const int nc = is_singlecomp ? 1 : (int) Ncomps, np = pair_count(nc);
    std::vector<double> mtl(np), etl(np);
    std::vector<int32_t> st(np);
    if (kg_transfer_lengths_batch(gpu, 1, &T, nullptr, mtl.data(), etl.data(), st.data())) fail("kg_transfer_lengths_batch failed");
    n_feeds++;
    for (int s : st) {
        if (s & KG_STATUS_BMID) throw std::runtime_error("Unable to converge bmid!");           // cpp/transfer_lengths.cpp:259-262
        if (s & KG_STATUS_GETR_MAXITER) throw std::runtime_error("get_R reached max iter!");
    }
    vector2d M(Ncomps, vector1d(Ncomps, 0.)), E(Ncomps, vector1d(Ncomps, 0.));
    if (is_singlecomp) { M.assign(Ncomps, vector1d(Ncomps, mtl[0])); E.assign(Ncomps, vector1d(Ncomps, etl[0])); }
    else {
        int p = 0;
        for (int i = 0; i < nc; i++) for (int j = i; j < nc; j++, p++) { M[i][j] = M[j][i] = mtl[p]; E[i][j] = E[j][i] = etl[p]; }
    }
    const StatePoint point = get_transfer_length_point(rho, T, x);
    if (!cache.mtl.has(point)) cache.mtl.store(point, M);
    if (!cache.etl.has(point)) cache.etl.store(point, E);
}

vector2d GpuMieKinGas::model_mtl(double rho, double T, const vector1d& x) {
    if (transfer_length_model_id != TransferLengthModel::EWCA) return Spherical::model_mtl(rho, T, x);
    const StatePoint point = get_transfer_length_point(rho, T, x);
    if (!cache.mtl.has(point)) feed_transfer_lengths(rho, T, x);
    return *cache.mtl.get(point);
}

vector2d GpuMieKinGas::model_etl(double rho, double T, const vector1d& x) {
    if (transfer_length_model_id != TransferLengthModel::EWCA) return Spherical::model_etl(rho, T, x);
    const StatePoint point = get_transfer_length_point(rho, T, x);
    if (!cache.etl.has(point)) feed_transfer_lengths(rho, T, x);
    return *cache.etl.get(point);
}

std::vector<double> GpuMieKinGas::omega_table(const std::vector<double>& T, const std::vector<int>& l, const std::vector<int>& r) {
    const int nc = is_singlecomp ? 1 : (int) Ncomps, np = pair_count(nc), nlr = (int) l.size(), nT = (int) T.size();
    std::vector<int32_t> ll(l.begin(), l.end()), rr(r.begin(), r.end());
    std::vector<double> om((size_t) nT * np * nlr);
    std::vector<int32_t> st((size_t) nT * np);
    if (kg_omega_batch(gpu, nT, T.data(), nlr, ll.data(), rr.data(), nullptr, om.data(), nullptr, st.data())) fail("kg_omega_batch failed");
    n_feeds++;
    for (int s : st) if (s & KG_STATUS_GETR_MAXITER) throw std::runtime_error("get_R reached max iter!");
    for (int t = 0; t < nT; t++) {
        int p = 0;
        for (int i = 0; i < nc; i++) for (int j = i; j < nc; j++, p++) for (int k = 0; k < nlr; k++) {
            const OmegaPoint point = get_omega_point(i, j, l[k], r[k], T[t]);
            if (!cache.omega.has(point)) { cache.omega.store(point, om[((size_t) t * np + p) * nlr + k]); n_stored++; }
        }
    }
    return om;
}

GpuMieKinGas::BatchResult GpuMieKinGas::transport_batch(const std::vector<double>& T, const std::vector<double>& Vm, const std::vector<double>& x,
                                                        int N, const std::vector<double>& dmudn, const std::vector<double>& cp_ideal) {
    const int n = (int) T.size(), nc = (int) Ncomps;
    if ((int) Vm.size() != n || (!is_singlecomp && (int) x.size() != n * nc)) throw std::invalid_argument("transport_batch: array sizes do not match");
    kg_transport_opts o{};
    o.N = N; o.is_idealgas = is_idealgas ? 1 : 0; o.omega_correlation_active = omega_correlation_active ? 1 : 0;
    o.frame_of_reference = KG_FOR_CON; o.dependent_idx = -1; o.solvent_idx = -1; o.do_compress = 1;
    o.properties = KG_PROP_VISCOSITY | (cp_ideal.empty() ? 0u : KG_PROP_CONDUCTIVITY) | (dmudn.empty() ? 0u : (KG_PROP_DIFFUSION | KG_PROP_THERMAL_DIFFUSION));
    std::vector<double> Tu(T);
    std::sort(Tu.begin(), Tu.end());
    Tu.erase(std::unique(Tu.begin(), Tu.end()), Tu.end());
    kg_tcache* c = nullptr;
    if (kg_tcache_create(gpu, &o, (int) Tu.size(), Tu.data(), nullptr, &c)) fail("kg_tcache_create failed");
    BatchResult res;
    res.nD = nc - 1;
    res.viscosity.resize(n);
    kg_transport_out out{};
    out.viscosity = res.viscosity.data();
    kg_eos_inputs eos{};
    if (!cp_ideal.empty()) { res.conductivity.resize(n); out.conductivity = res.conductivity.data(); eos.cp_ideal = cp_ideal.data(); }
    if (!dmudn.empty()) {
        res.D.resize((size_t) n * res.nD * res.nD); res.DT.resize((size_t) n * nc);
        out.D = res.D.data(); out.DT = res.DT.data(); eos.dmudn = dmudn.data();
    }
    std::vector<int32_t> st(n);
    const int rc = kg_transport_batch(gpu, c, n, T.data(), Vm.data(), is_singlecomp ? nullptr : x.data(), &o, &eos, &out, st.data());
    kg_tcache_destroy(c);
    if (rc) fail("kg_transport_batch failed");
    n_feeds++;
    for (int s_ : st) if (s_ & KG_STATUS_SINGULAR) throw std::runtime_error("singular Sonine-expansion system");
    return res;
}
