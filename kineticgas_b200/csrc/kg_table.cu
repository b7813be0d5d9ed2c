// kg_table.cu -- device-resident Omega(T) interpolation tables with persistence.
//
// What it replaces: the reference memoises collision integrals in a hash map keyed on 0.1 K buckets (cpp/KineticGasCache.h:92-164), per
// object and per process; a sweep over new temperatures integrates everything again. Here a table is built ONCE from exact device
// quadrature (kg_omega_plan) at nodes spaced uniformly in ln T, kept in HBM, evaluated by a kernel (4-point Lagrange interpolation of
// ln Omega in ln T, one thread per (T, pair, (l,r))), and can be written to / read from a file that carries a fingerprint of the
// model it belongs to.
//
// Accuracy contract (different from the bit-parity of the batched path, and stated as such): the reference's Omega(T) is not a smooth
// function -- the integration lattice switches at 20 / 35 / 50 / 70 K (cpp/Spherical.cpp:60-84) and the adaptive mesh changes shape
// with T -- so the table is SEGMENTED at those temperatures (no interpolation stencil crosses one) and inside a segment the
// interpolant reproduces the exact path to the level of the mesh noise of the reference itself (measured: tests/test_gpu_table.py).
#include "kg_internal.cuh"
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>

struct kg_omega_table {
    kg_model* m = nullptr;
    int npairs = 0, nlr = 0, nseg = 0, nnodes = 0;
    std::vector<int32_t> l, r;
    std::vector<double> seg_lo, seg_hi;       // [nseg] temperature range of each segment (hi exclusive except for the last)
    std::vector<int> seg_first, seg_n;        // [nseg] first node, node count
    std::vector<double> T;                    // [nnodes]
    std::vector<double> lnom;                 // [nnodes][npairs][nlr]  ln Omega
    double *d_lnom = nullptr, *d_seg = nullptr;   // d_seg: [nseg][4] = ln lo, ln hi, first node, node count
    uint64_t fingerprint = 0;
};

namespace {

uint64_t fnv1a(const void* p, size_t n, uint64_t h = 1469598103934665603ull) {
    const unsigned char* b = (const unsigned char*) p;
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}
uint64_t model_fingerprint(const kg_model* m) {
    uint64_t h = fnv1a(&m->ncomps, sizeof(int));
    h = fnv1a(&m->is_singlecomp, sizeof(int), h);
    h = fnv1a(m->mass.data(), m->mass.size() * sizeof(double), h);
    for (auto& d : m->descs) { h = fnv1a(&d.kind, sizeof(d.kind), h); h = fnv1a(&d.nterms, sizeof(d.nterms), h); h = fnv1a(&d.sigma, sizeof(double), h);
                               h = fnv1a(&d.eps, sizeof(double), h); h = fnv1a(d.C, sizeof(d.C), h); h = fnv1a(d.lam, sizeof(d.lam), h); }
    for (auto& g : m->gdescs) h = fnv1a(&g, sizeof(g), h);
    return h;
}

// one thread per (T, pair, lr)
__global__ void table_eval_kernel(int nT, int npl /* npairs * nlr */, int nseg, const double* __restrict__ seg, const double* __restrict__ lnom,
                                  const double* __restrict__ T, double* __restrict__ out) {
    const long long e = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long) nT * npl) return;
    const int t = (int) (e / npl), k = (int) (e % npl);
    const double x = log(T[t]);
    int s = nseg - 1;
    for (int q = 0; q < nseg; q++) if (x <= seg[4 * q + 1]) { s = q; break; }
    const double x0 = seg[4 * s], x1 = seg[4 * s + 1];
    const int first = (int) seg[4 * s + 2], n = (int) seg[4 * s + 3];
    if (x < x0 - 1e-12 || (s == nseg - 1 && x > x1 + 1e-12)) { out[e] = nan(""); return; }      // outside the table
    const double h = (x1 - x0) / (n - 1), u = (x - x0) / h;
    int i = (int) floor(u) - 1;
    i = max(0, min(i, n - 4));
    const double p = u - i;                                  // position in units of h relative to node i, stencil i .. i+3
    const double y0 = lnom[(size_t) (first + i) * npl + k], y1 = lnom[(size_t) (first + i + 1) * npl + k];
    const double y2 = lnom[(size_t) (first + i + 2) * npl + k], y3 = lnom[(size_t) (first + i + 3) * npl + k];
    const double w0 = -(p - 1) * (p - 2) * (p - 3) / 6., w1 = p * (p - 2) * (p - 3) / 2., w2 = -p * (p - 1) * (p - 3) / 2., w3 = p * (p - 1) * (p - 2) / 6.;
    out[e] = exp(w0 * y0 + w1 * y1 + w2 * y2 + w3 * y3);
}

int upload_table(kg_omega_table* t) {
    kg_model* m = t->m;
    if (cudaSetDevice(m->device) != cudaSuccess) return -2;
    std::vector<double> seg((size_t) t->nseg * 4);
    for (int s = 0; s < t->nseg; s++) {
        seg[4 * s] = log(t->T[t->seg_first[s]]); seg[4 * s + 1] = log(t->T[t->seg_first[s] + t->seg_n[s] - 1]);
        seg[4 * s + 2] = t->seg_first[s]; seg[4 * s + 3] = t->seg_n[s];
    }
    if (kg_dmalloc(m, &t->d_lnom, t->lnom.size() * sizeof(double)) != cudaSuccess) return -2;
    if (kg_dmalloc(m, &t->d_seg, seg.size() * sizeof(double)) != cudaSuccess) return -2;
    if (cudaMemcpy(t->d_lnom, t->lnom.data(), t->lnom.size() * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess) return -2;
    if (cudaMemcpy(t->d_seg, seg.data(), seg.size() * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess) return -2;
    return 0;
}

} // namespace

void kg_register_child(void* child, kg_model* m, int kind);      // kg_omega.cu (kind 2 = table)

extern "C" void kg_omega_table_destroy(kg_omega_table* t) {
    if (!t) return;
    if (t->d_lnom || t->d_seg) {
        if (!KgRegistry::get().remove(t)) return;          // unknown handle: already destroyed together with its model
        cudaSetDevice(t->m->device); t->m->dfree(t->d_lnom); t->m->dfree(t->d_seg);
    }
    delete t;
}

extern "C" int kg_omega_table_create(kg_model* m, double Tmin, double Tmax, int nodes_per_decade, int nlr, const int32_t* l, const int32_t* r,
                                     kg_omega_table** out) {
    KG_TRACE();
    if (!m) return -1;
    auto set_error = [&](const std::string& s) { m->err = s; };
    if (!(Tmin > 0) || !(Tmax > Tmin) || nodes_per_decade < 4 || nlr < 1 || !l || !r || !out) { set_error("kg_omega_table_create: invalid argument"); return -1; }
    if (m->coeff_T_required || m->gen_T_dependent) { set_error("kg_omega_table_create: models with temperature-dependent potentials are not tabulated"); return -1; }
    kg_omega_table* t = new kg_omega_table;
    t->m = m; t->npairs = m->npairs; t->nlr = nlr; t->l.assign(l, l + nlr); t->r.assign(r, r + nlr);
    t->fingerprint = model_fingerprint(m);
    // segments: [Tmin, Tmax] cut at the lattice-regime temperatures of cpp/Spherical.cpp:60-84
    const double edges[4] = {20.0, 35.0, 50.0, 70.0};
    std::vector<double> cuts = {Tmin};
    for (double e : edges) if (e > Tmin && e < Tmax) cuts.push_back(e);
    cuts.push_back(Tmax);
    for (size_t s = 0; s + 1 < cuts.size(); s++) {
        const double lo = cuts[s];
        const double hi = (s + 2 < cuts.size()) ? std::nextafter(cuts[s + 1], 0.0) : cuts[s + 1];      // last node of a segment stays in its regime
        const int n = std::max(5, (int) std::ceil(nodes_per_decade * std::log10(hi / lo)) + 1);
        t->seg_lo.push_back(lo); t->seg_hi.push_back(hi); t->seg_first.push_back((int) t->T.size()); t->seg_n.push_back(n);
        for (int i = 0; i < n; i++) t->T.push_back(i == 0 ? lo : i == n - 1 ? hi : std::exp(std::log(lo) + (std::log(hi) - std::log(lo)) * i / (n - 1)));
    }
    t->nseg = (int) t->seg_lo.size(); t->nnodes = (int) t->T.size();
    std::vector<double> om((size_t) t->nnodes * t->npairs * nlr);
    std::vector<int32_t> st((size_t) t->nnodes * t->npairs);
    const int rc = kg_omega_batch(m, t->nnodes, t->T.data(), nlr, l, r, nullptr, om.data(), nullptr, st.data());
    if (rc) { delete t; return rc; }
    for (int32_t s : st) if (s & (KG_STATUS_GETR_MAXITER | KG_STATUS_NAN)) { set_error("kg_omega_table_create: a node failed to integrate"); delete t; return -4; }
    t->lnom.resize(om.size());
    for (size_t k = 0; k < om.size(); k++) t->lnom[k] = std::log(om[k]);
    kg_register_child(t, m, 2);
    if (upload_table(t)) { set_error("kg_omega_table_create: device upload failed"); kg_omega_table_destroy(t); return -2; }
    *out = t;
    return 0;
}

extern "C" int kg_omega_table_eval_dev(kg_omega_table* t, int nT, const double* T_dev, double* omega_dev) {
    KG_TRACE();
    if (!t || nT < 0 || !T_dev || !omega_dev) return -1;
    if (nT == 0) return 0;
    kg_model* m = t->m;
    auto set_error = [&](const std::string& s) { m->err = s; };
    CUDA_OK(cudaSetDevice(m->device));
    const int npl = t->npairs * t->nlr;
    const long long tot = (long long) nT * npl;
    table_eval_kernel<<<(unsigned) ((tot + 255) / 256), 256, 0, m->stream>>>(nT, npl, t->nseg, t->d_seg, t->d_lnom, T_dev, omega_dev);
    CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int kg_omega_table_eval(kg_omega_table* t, int nT, const double* T, double* omega) {
    KG_TRACE();
    if (!t || nT < 0 || !T || !omega) return -1;
    if (nT == 0) return 0;
    kg_model* m = t->m;
    auto set_error = [&](const std::string& s) { m->err = s; };
    CUDA_OK(cudaSetDevice(m->device));
    const size_t no = (size_t) nT * t->npairs * t->nlr;
    double *dT = nullptr, *dO = nullptr;
    CUDA_OK(kg_dmalloc(m, &dT, nT * sizeof(double))); CUDA_OK(kg_dmalloc(m, &dO, no * sizeof(double)));
    CUDA_OK(cudaMemcpyAsync(dT, T, nT * sizeof(double), cudaMemcpyHostToDevice, m->stream));
    int rc = kg_omega_table_eval_dev(t, nT, dT, dO);
    if (!rc) {
        CUDA_OK(cudaMemcpyAsync(omega, dO, no * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
        CUDA_OK(cudaStreamSynchronize(m->stream));
    }
    m->dfree(dT); m->dfree(dO);
    return rc;
}

extern "C" int kg_omega_table_info(const kg_omega_table* t, double* out8) {
    if (!t || !out8) return -1;
    out8[0] = t->seg_lo.front(); out8[1] = t->seg_hi.back(); out8[2] = t->nnodes; out8[3] = t->nseg; out8[4] = t->npairs; out8[5] = t->nlr;
    out8[6] = (double) (t->lnom.size() * sizeof(double)); out8[7] = 0;
    return 0;
}

// file: "KGOT1\0\0\0" | fingerprint u64 | npairs, nlr, nseg, nnodes (int32) | l[nlr], r[nlr] (int32) | seg_first[nseg], seg_n[nseg] (int32)
//       | seg_lo[nseg], seg_hi[nseg], T[nnodes], lnom[nnodes * npairs * nlr] (double)
extern "C" int kg_omega_table_save(const kg_omega_table* t, const char* path) {
    if (!t || !path) return -1;
    FILE* f = fopen(path, "wb");
    if (!f) { t->m->err = std::string("kg_omega_table_save: cannot open ") + path; return -1; }
    const char magic[8] = {'K', 'G', 'O', 'T', '1', 0, 0, 0};
    const int32_t hdr[4] = {t->npairs, t->nlr, t->nseg, t->nnodes};
    bool ok = fwrite(magic, 1, 8, f) == 8 && fwrite(&t->fingerprint, 8, 1, f) == 1 && fwrite(hdr, 4, 4, f) == 4;
    ok = ok && fwrite(t->l.data(), 4, t->nlr, f) == (size_t) t->nlr && fwrite(t->r.data(), 4, t->nlr, f) == (size_t) t->nlr;
    ok = ok && fwrite(t->seg_first.data(), 4, t->nseg, f) == (size_t) t->nseg && fwrite(t->seg_n.data(), 4, t->nseg, f) == (size_t) t->nseg;
    ok = ok && fwrite(t->seg_lo.data(), 8, t->nseg, f) == (size_t) t->nseg && fwrite(t->seg_hi.data(), 8, t->nseg, f) == (size_t) t->nseg;
    ok = ok && fwrite(t->T.data(), 8, t->nnodes, f) == (size_t) t->nnodes && fwrite(t->lnom.data(), 8, t->lnom.size(), f) == t->lnom.size();
    fclose(f);
    if (!ok) { t->m->err = "kg_omega_table_save: write failed"; return -1; }
    return 0;
}

extern "C" int kg_omega_table_load(kg_model* m, const char* path, kg_omega_table** out) {
    KG_TRACE();
    if (!m || !path || !out) return -1;
    auto set_error = [&](const std::string& s) { m->err = s; };
    FILE* f = fopen(path, "rb");
    if (!f) { set_error(std::string("kg_omega_table_load: cannot open ") + path); return -1; }
    kg_omega_table* t = new kg_omega_table;
    t->m = m;
    char magic[8]; int32_t hdr[4];
    bool ok = fread(magic, 1, 8, f) == 8 && memcmp(magic, "KGOT1", 5) == 0 && fread(&t->fingerprint, 8, 1, f) == 1 && fread(hdr, 4, 4, f) == 4;
    if (ok) {
        t->npairs = hdr[0]; t->nlr = hdr[1]; t->nseg = hdr[2]; t->nnodes = hdr[3];
        ok = t->npairs > 0 && t->nlr > 0 && t->nlr <= 64 && t->nseg > 0 && t->nseg <= 8 && t->nnodes >= 5 && t->nnodes < (1 << 24);
    }
    if (ok) {
        t->l.resize(t->nlr); t->r.resize(t->nlr); t->seg_first.resize(t->nseg); t->seg_n.resize(t->nseg); t->seg_lo.resize(t->nseg); t->seg_hi.resize(t->nseg);
        t->T.resize(t->nnodes); t->lnom.resize((size_t) t->nnodes * t->npairs * t->nlr);
        ok = fread(t->l.data(), 4, t->nlr, f) == (size_t) t->nlr && fread(t->r.data(), 4, t->nlr, f) == (size_t) t->nlr;
        ok = ok && fread(t->seg_first.data(), 4, t->nseg, f) == (size_t) t->nseg && fread(t->seg_n.data(), 4, t->nseg, f) == (size_t) t->nseg;
        ok = ok && fread(t->seg_lo.data(), 8, t->nseg, f) == (size_t) t->nseg && fread(t->seg_hi.data(), 8, t->nseg, f) == (size_t) t->nseg;
        ok = ok && fread(t->T.data(), 8, t->nnodes, f) == (size_t) t->nnodes && fread(t->lnom.data(), 8, t->lnom.size(), f) == t->lnom.size();
    }
    fclose(f);
    if (!ok) { set_error("kg_omega_table_load: not a table file, or truncated"); delete t; return -1; }
    if (t->fingerprint != model_fingerprint(m) || t->npairs != m->npairs) {
        set_error("kg_omega_table_load: the table was built for a different model (potential parameters or masses differ)"); delete t; return -1; }
    kg_register_child(t, m, 2);
    if (upload_table(t)) { set_error("kg_omega_table_load: device upload failed"); kg_omega_table_destroy(t); return -2; }
    *out = t;
    return 0;
}

extern "C" int kg_omega_table_lr(const kg_omega_table* t, int32_t* l, int32_t* r) {
    if (!t) return -1;
    for (int k = 0; k < t->nlr; k++) { if (l) l[k] = t->l[k]; if (r) r[k] = t->r[k]; }
    return t->nlr;
}
