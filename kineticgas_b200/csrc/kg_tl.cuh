// kg_tl.cuh -- exchange-weighted closest-approach (EWCA) transfer lengths on the GPU. Included by kg_omega.cu.
//
// Reference replaced: Spherical::get_transfer_length / tl_ewca / ewca_inner / ewca_weight / get_bmid / get_b_max_g
// (cpp/transfer_lengths.cpp:38-80,165-266) and weighted_simpson (cpp/Integration/Integration.cpp:375-415).
//   TL_ij(T) = sum_g c_g sum_b c_b R(g,b) w(g,b) / sum_g c_g sum_b c_b w(g,b),  w = transfer(chi) vdf(g) g b
// on the reference's fixed Simpson nodes: g in [0,3.5] (15 intervals; the odd count drops node 13 -- kept) and
// [3.5,5] (10); b in [0,b_mid] (15) and [b_mid,b_max] (20), with b_mid (chi = 0) and b_max (|chi + 1e-5| <= 1e-3)
// found by the reference's own sequential searches. The reference evaluates all of this twice (MTL, ETL); here one pass
// serves both, because only the weight differs.
//   tl_search_kernel : one lane per (pair,T,g-node): theta_lim(g) once, then the b_mid / b_max searches
//   tl_node_kernel   : one lane per (pair,T,g-node,b-node): R, chi, both weights
//   tl_reduce_kernel : one thread per (pair,T): the two nested Simpson sums in the reference's summation order
#pragma once

namespace {

constexpr int TL_NG = 26;      // g nodes in evaluation order (15 + 11)
constexpr int TL_NB = 36;      // b nodes per g in evaluation order (15 + 21)

struct TlItem { int pair; int coeff_off; double T; };

// chi for arbitrary (g, b/sigma) with the column's theta_lim integral already known (cpp/Spherical.cpp:259-270,455-465)
template<int NT>
__device__ double chi_with_lim(const PairParams& pp, const double* coeff, const NodeRows<NT>& NR,
                               double T, double g, double b, double I_lim, double& R, unsigned& nodes, int& st) {
    if (b > 100) return 0.0;
    if (b < 1e-3) return PI_REF;
    const double I = lane_theta_integral<NT>(pp, coeff, NR, T, g, b, R, nodes, st);
    const double th = __dadd_rn(__dsub_rn(I, I_lim), PI_REF / 2);
    return __dsub_rn(PI_REF, __dmul_rn(2.0, th));
}

template<int NT>
__global__ void __launch_bounds__(256) tl_search_kernel(const PairParams* __restrict__ pairs, const double* const* __restrict__ rows_by_pair,
                                                        const TlItem* __restrict__ items, int pair0, int nitems, const double* __restrict__ coeff_pool,
                                                        const double* __restrict__ gnodes, double* __restrict__ ilim,
                                                        double* __restrict__ bmid_out, double* __restrict__ bmax_out, int* __restrict__ status) {
    extern __shared__ double srows[];
    __shared__ PairParams pp;
    constexpr int W = 2 + NT;
    const int pair = pair0 + blockIdx.y, item0 = pair * nitems;       // grid.y walks the pairs of this launch, nitems = nT
    const double* __restrict__ rows = rows_by_pair[pair];
    for (int i = threadIdx.x; i < KS_ROWS * W; i += blockDim.x) srows[i] = rows[i];
    if (threadIdx.x == 0) pp = pairs[pair];
    __syncthreads();
    const NodeRows<NT> NR{(unsigned) __cvta_generic_to_shared(srows), srows, rows, KS_ROWS};
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nitems * TL_NG) return;
    const int it = item0 + e / TL_NG, q = e % TL_NG;
    const TlItem item = items[it];
    const double* coeff = (item.coeff_off >= 0) ? coeff_pool + item.coeff_off : nullptr;
    const double g = gnodes[q], T = item.T;
    const long long o = (long long) it * TL_NG + q;
    if (g == 0) { ilim[o] = 0.0; bmid_out[o] = 5.0; bmax_out[o] = 10.0; return; }   // transfer_lengths.cpp:223,235
    int st = 0; unsigned nodes = 0; double R;
    const double I_lim = lane_theta_integral<NT>(pp, coeff, NR, T, g, 100.0, R, nodes, st);
    auto chi = [&](double b) { return chi_with_lim<NT>(pp, coeff, NR, T, g, b, I_lim, R, nodes, st); };
    // get_bmid, cpp/transfer_lengths.cpp:234-266
    double b = 0.9, db = 0.1;
    double chi_val = chi(b), next_chi_val, dchidb;
    const double step_tol = 1e-6, chi_tol = 1e-6;
    int guard = 0;
    while ((fabs(chi_val) > chi_tol) && (fabs(db) > step_tol) && guard++ < 10000) {
        if ((chi_val < 0 && db > 0) || (chi_val > 0 && db < 0)) db *= -0.9;
        next_chi_val = chi(b + db);
        while (fabs(next_chi_val) >= fabs(chi_val)) {
            db *= 0.1;
            next_chi_val = chi(b + db);
            if (fabs(db) < step_tol) break;
        }
        if (fabs(db) > step_tol) {
            b += db;
            dchidb = (chi_val - next_chi_val) / db;
            chi_val = next_chi_val;
            db = -chi_val / dchidb;
        }
        if ((b > 10) || (b < 0)) { st |= 8; break; }       // "Unable to converge bmid!"
    }
    const double bmid = b;
    // get_b_max_g, cpp/transfer_lengths.cpp:221-232
    b = 2 * bmid;
    chi_val = chi(b);
    guard = 0;
    while (fabs(chi_val + 1e-5) > 1e-3 && guard++ < 400) { b += 0.5; chi_val = chi(b); }
    if (guard >= 400) st |= 8;
    ilim[o] = I_lim; bmid_out[o] = bmid; bmax_out[o] = b;
    if (st) atomicOr(&status[it], st);
}

// The same searches with one WARP per (pair, T, g-node), for the power-law potentials (opt-in: KG_TL_WARP_SEARCH=1). The searches are
// sequential in b and a transfer-length call has only nT x 26 of them per pair, so one lane per search leaves the GPU mostly empty;
// here all 32 lanes follow the same search and share each theta integral node by node (theta_integral_warp). MEASURED SLOWER than the
// lane version on B200 (configs[2]: TL stage 1.26 ms against 0.90 ms; configs[4]: 1.80 against 1.16 ms): the shared sum visits every
// node (the lane version's strided sum visits 9 %), and 63 nodes per lane plus the reductions of each run of 256 is no shorter a chain
// than ~190 strided nodes in one lane. Kept for comparison runs; the lane kernel is the product path.
template<int NT>
__global__ void __launch_bounds__(256) tl_search_warp_kernel(const PairParams* __restrict__ pairs, const double* const* __restrict__ rows_by_pair,
                                                             const TlItem* __restrict__ items, int pair0, int nitems, const double* __restrict__ coeff_pool,
                                                             const double* __restrict__ gnodes, double* __restrict__ ilim,
                                                             double* __restrict__ bmid_out, double* __restrict__ bmax_out, int* __restrict__ status) {
    if constexpr (NT != 1) {
    extern __shared__ double srows[];
    __shared__ PairParams pp;
    constexpr int W = 2 + NT;
    const int pair = pair0 + blockIdx.y, item0 = pair * nitems;
    const double* __restrict__ rows = rows_by_pair[pair];
    for (int i = threadIdx.x; i < KS_ROWS * W; i += blockDim.x) srows[i] = rows[i];
    if (threadIdx.x == 0) pp = pairs[pair];
    __syncthreads();
    const NodeRows<NT> NR{(unsigned) __cvta_generic_to_shared(srows), srows, rows, KS_ROWS};
    const int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;      // one search per warp
    if (e >= nitems * TL_NG) return;
    const int it = item0 + e / TL_NG, q = e % TL_NG;
    const TlItem item = items[it];
    const double* coeff = (item.coeff_off >= 0) ? coeff_pool + item.coeff_off : nullptr;
    const double g = gnodes[q], T = item.T;
    const long long o = (long long) it * TL_NG + q;
    if (g == 0) { if (lane == 0) { ilim[o] = 0.0; bmid_out[o] = 5.0; bmax_out[o] = 10.0; } return; }   // transfer_lengths.cpp:223,235
    int st = 0;
    auto theta = [&](double b) {                               // theta integral at (g, b): root finding in every lane (same result), sum shared
        double A[NT], c2, Ap[NT];
        lane_setup<NT>(pp, coeff, T, g, b, A, c2, st);
#pragma unroll
        for (int t = 0; t < NT; t++) Ap[t] = c2 * A[t];
        int kk = 0;
        return theta_integral_warp<NT>(NR, Ap, c2, kk, st);
    };
    const double I_lim = theta(100.0);
    auto chi = [&](double b) {                                 // cpp/Spherical.cpp:259-270,455-465
        if (b > 100) return 0.0;
        if (b < 1e-3) return PI_REF;
        const double th = __dadd_rn(__dsub_rn(theta(b), I_lim), PI_REF / 2);
        return __dsub_rn(PI_REF, __dmul_rn(2.0, th));
    };
    // get_bmid, cpp/transfer_lengths.cpp:234-266
    double b = 0.9, db = 0.1;
    double chi_val = chi(b), next_chi_val, dchidb;
    const double step_tol = 1e-6, chi_tol = 1e-6;
    int guard = 0;
    while ((fabs(chi_val) > chi_tol) && (fabs(db) > step_tol) && guard++ < 10000) {
        if ((chi_val < 0 && db > 0) || (chi_val > 0 && db < 0)) db *= -0.9;
        next_chi_val = chi(b + db);
        while (fabs(next_chi_val) >= fabs(chi_val)) {
            db *= 0.1;
            next_chi_val = chi(b + db);
            if (fabs(db) < step_tol) break;
        }
        if (fabs(db) > step_tol) {
            b += db;
            dchidb = (chi_val - next_chi_val) / db;
            chi_val = next_chi_val;
            db = -chi_val / dchidb;
        }
        if ((b > 10) || (b < 0)) { st |= 8; break; }       // "Unable to converge bmid!"
    }
    const double bmid = b;
    // get_b_max_g, cpp/transfer_lengths.cpp:221-232
    b = 2 * bmid;
    chi_val = chi(b);
    guard = 0;
    while (fabs(chi_val + 1e-5) > 1e-3 && guard++ < 400) { b += 0.5; chi_val = chi(b); }
    if (guard >= 400) st |= 8;
    if (lane == 0) {
        ilim[o] = I_lim; bmid_out[o] = bmid; bmax_out[o] = b;
        if (st) atomicOr(&status[it], st);
    }
    }
}

// b node k (evaluation order of weighted_simpson, Integration.cpp:375-394) of the two b intervals
__device__ __forceinline__ double tl_bnode(int k, double bmid, double bmax, double& coef) {
    double x0, xN; int N, kk;
    if (k < 15) { x0 = 0.0; xN = bmid; N = 15; kk = k; } else { x0 = bmid; xN = bmax; N = 20; kk = k - 15; }
    const double dx = (xN - x0) / N;
    const int npair = N / 2 - 1;
    coef = 1.0;
    if (kk == 0) return x0;
    if (kk == 1) return xN;
    if (kk < 2 + 2 * npair) {
        const int i = (kk - 2) / 2 + 1;
        if (((kk - 2) & 1) == 0) { coef = 4.0; return x0 + (2 * i - 1) * dx; }
        coef = 2.0; return x0 + (2 * i) * dx;
    }
    coef = 4.0;
    return xN - dx;
}

template<int NT>
__global__ void __launch_bounds__(256) tl_node_kernel(const PairParams* __restrict__ pairs, const double* const* __restrict__ rows_by_pair,
                                                      const TlItem* __restrict__ items, int pair0, int nitems, const double* __restrict__ coeff_pool,
                                                      const double* __restrict__ gnodes, const double* __restrict__ ilim,
                                                      const double* __restrict__ bmid_in, const double* __restrict__ bmax_in,
                                                      double4* __restrict__ fw, int* __restrict__ status) {
    extern __shared__ double srows[];
    __shared__ PairParams pp;
    constexpr int W = 2 + NT;
    const int pair = pair0 + blockIdx.y, item0 = pair * nitems;
    const double* __restrict__ rows = rows_by_pair[pair];
    for (int i = threadIdx.x; i < KS_ROWS * W; i += blockDim.x) srows[i] = rows[i];
    if (threadIdx.x == 0) pp = pairs[pair];
    __syncthreads();
    const NodeRows<NT> NR{(unsigned) __cvta_generic_to_shared(srows), srows, rows, KS_ROWS};
    const long long e = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long) nitems * TL_NG * TL_NB) return;
    const int k = (int) (e % TL_NB), q = (int) ((e / TL_NB) % TL_NG), it = item0 + (int) (e / (TL_NB * TL_NG));
    const TlItem item = items[it];
    const double* coeff = (item.coeff_off >= 0) ? coeff_pool + item.coeff_off : nullptr;
    const double g = gnodes[q], T = item.T;
    const long long o = (long long) it * TL_NG + q;
    double4 out = make_double4(0.0, 0.0, 0.0, 0.0);      // (R w_mtl, w_mtl, R w_etl, w_etl), Simpson coefficient not applied
    if (g != 0) {
        double coef;
        const double b = tl_bnode(k, bmid_in[o], bmax_in[o], coef);
        int st = 0; unsigned nodes = 0; double R = 0.0;
        const double chi = chi_with_lim<NT>(pp, coeff, NR, T, g, b, ilim[o], R, nodes, st);
        if (b < 1e-3 || b > 100) {        // chi short-cuts do not compute R; the integrand's func(b) = get_R still needs it
            R = lane_get_R<NT>(pp, coeff, T, g, b, st);
        }
        // cpp/transfer_lengths.cpp:161-163,196-219
        const double vdf = sqrt(2. / PI_REF) * (g * g) * exp(-0.5 * (g * g));
        const double mt = g * sqrt(2 * (1 - cos(chi))) * fabs(sin(chi / 2.));
        const double et = g * fabs(cos(chi) - sin(chi) - 1);
        const double wm = mt * vdf * g * b, we = et * vdf * g * b;
        const double Rm = R * pp.sigma;
        out = make_double4(Rm * wm, wm, Rm * we, we);
        if (st) atomicOr(&status[it], st);
    }
    fw[(long long) blockIdx.y * nitems * TL_NG * TL_NB + e] = out;      // fw holds the pairs of one launch
}

// nested Simpson sums in the reference's order (Integration.cpp:375-415, transfer_lengths.cpp:165-194)
// One warp per (pair, T): lane q owns g node q (its two b sums, 36 nodes read in the reference's order), lane 0 then adds the 26 results
// in the reference's order. (One thread per item walked 936 double4 reads alone: 0.33 ms for the 1500 items of configs[4].)
__global__ void __launch_bounds__(128) tl_reduce_kernel(int item0, int nitems, const double* __restrict__ bmid_in, const double* __restrict__ bmax_in,
                                 const double4* __restrict__ fw, double* __restrict__ mtl, double* __restrict__ etl) {
    const int li = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, q = threadIdx.x & 31;     // item within this launch (pairs of the launch are contiguous)
    if (li >= nitems) return;
    const int it = item0 + li;
    double F[2] = {0, 0}, Wt[2] = {0, 0};
    if (q < TL_NG) {
        const long long o = (long long) it * TL_NG + q;
        const double bmid = bmid_in[o], bmax = bmax_in[o];
        for (int iv = 0; iv < 2; iv++) {
            const int k0 = iv ? 15 : 0, nk = iv ? 21 : 15, N = iv ? 20 : 15;
            const double dx = iv ? (bmax - bmid) / N : (bmid - 0.0) / N;
            double f[2] = {0, 0}, w[2] = {0, 0};
            const double4* p = fw + ((long long) li * TL_NG + q) * TL_NB + k0;
            // x0 and xN first, then the (4,2) pairs, then the last 4 -- same association as the reference
            w[0] += p[0].y + p[1].y; f[0] += p[0].x + p[1].x;
            w[1] += p[0].w + p[1].w; f[1] += p[0].z + p[1].z;
            for (int k = 2; k + 1 < nk - 1; k += 2) {
                w[0] += 4. * p[k].y + 2. * p[k + 1].y; f[0] += 4. * p[k].x + 2. * p[k + 1].x;
                w[1] += 4. * p[k].w + 2. * p[k + 1].w; f[1] += 4. * p[k].z + 2. * p[k + 1].z;
            }
            w[0] += 4. * p[nk - 1].y; f[0] += 4. * p[nk - 1].x;
            w[1] += 4. * p[nk - 1].w; f[1] += 4. * p[nk - 1].z;
            for (int pr = 0; pr < 2; pr++) { w[pr] *= dx / 3.; f[pr] *= dx / 3.; }
            if (iv == 0) { F[0] = f[0]; F[1] = f[1]; Wt[0] = w[0]; Wt[1] = w[1]; }
            else { F[0] = F[0] + f[0]; F[1] = F[1] + f[1]; Wt[0] = Wt[0] + w[0]; Wt[1] = Wt[1] + w[1]; }
        }
    }
    double Fg[2][2] = {{0, 0}, {0, 0}}, Wg[2][2] = {{0, 0}, {0, 0}};   // [g interval][property]
    for (int qs = 0; qs < TL_NG; qs++) {
        // outer Simpson over g: q order is x0, xN, pairs.., last; coefficients 1,1,(4,2).., 4
        const int gi = qs < 15 ? 0 : 1, qq = qs < 15 ? qs : qs - 15, nq = qs < 15 ? 15 : 11;
        const double cg = (qq < 2) ? 1.0 : (qq == nq - 1) ? 4.0 : (((qq - 2) & 1) == 0 ? 4.0 : 2.0);
        for (int pr = 0; pr < 2; pr++) {
            const double Fq = __shfl_sync(0xffffffffu, F[pr], qs), Wq = __shfl_sync(0xffffffffu, Wt[pr], qs);
            Fg[gi][pr] += cg * Fq; Wg[gi][pr] += cg * Wq;
        }
    }
    if (q != 0) return;
    const double dxg[2] = {(3.5 - 0.) / 15, (5. - 3.5) / 10};
    double Fm = 0, Wm = 0, Fe = 0, We = 0;
    Fm = Fg[0][0] * (dxg[0] / 3) + Fg[1][0] * (dxg[1] / 3); Wm = Wg[0][0] * (dxg[0] / 3) + Wg[1][0] * (dxg[1] / 3);
    Fe = Fg[0][1] * (dxg[0] / 3) + Fg[1][1] * (dxg[1] / 3); We = Wg[0][1] * (dxg[0] / 3) + Wg[1][1] * (dxg[1] / 3);
    mtl[it] = Fm / Wm;
    etl[it] = Fe / We;
}

// g nodes in the evaluation order of weighted_simpson(FW, x0, xN, N) (Integration.cpp:396-415), same expressions
void tl_gnodes(double* g) {
    int n = 0;
    const double bounds[3] = {0., 3.5, 5.};
    const int Ns[2] = {15, 10};
    for (int iv = 0; iv < 2; iv++) {
        const double x0 = bounds[iv], xN = bounds[iv + 1];
        const int N = Ns[iv];
        const double dx = (xN - x0) / N;
        g[n++] = x0; g[n++] = xN;
        for (size_t i = 1; i <= (size_t) ((N / 2) - 1); i++) { g[n++] = x0 + (2 * i - 1) * dx; g[n++] = x0 + (2 * i) * dx; }
        g[n++] = xN - dx;
    }
}

} // namespace

extern "C" int kg_transfer_lengths_batch(kg_model* m, int nT, const double* T, const double* coeff_T, double* mtl, double* etl, int32_t* status) {
    KG_TRACE();
    if (!m) return -1;
    auto set_error = [&](const std::string& s) { m->err = s; };
    if (nT < 1 || !T || !mtl || !etl) { set_error("kg_transfer_lengths_batch: invalid argument"); return -1; }
    if (m->coeff_T_required && !coeff_T) { set_error(KG_NEED_COEFF); return -1; }
    CUDA_OK(cudaSetDevice(m->device));
    const int Np = m->npairs, nitems = nT * Np;
    std::vector<double> coeff_buf;
    if (kg_effective_coeff(m, nT, T, coeff_T, coeff_buf, &coeff_T)) return -1;
    std::vector<TlItem> items(nitems);
    for (int p = 0; p < Np; p++) for (int t = 0; t < nT; t++) {
        TlItem& it = items[p * nT + t];
        it.pair = p; it.T = T[t]; it.coeff_off = coeff_T ? (t * Np + p) * m->coeff_stride : -1;
    }
    double gn[TL_NG]; tl_gnodes(gn);
    TlItem* d_items = nullptr; double *d_g = nullptr, *d_ilim = nullptr, *d_bmid = nullptr, *d_bmax = nullptr, *d_mtl = nullptr, *d_etl = nullptr, *d_coeff = nullptr;
    double4* d_fw = nullptr; int* d_status = nullptr;
    cudaStream_t s = m->stream;
    CUDA_OK(kg_dmalloc(m, &d_items, nitems * sizeof(TlItem)));
    CUDA_OK(kg_dmalloc(m, &d_g, TL_NG * sizeof(double)));
    CUDA_OK(kg_dmalloc(m, &d_ilim, (size_t) nitems * TL_NG * sizeof(double)));
    CUDA_OK(kg_dmalloc(m, &d_bmid, (size_t) nitems * TL_NG * sizeof(double)));
    CUDA_OK(kg_dmalloc(m, &d_bmax, (size_t) nitems * TL_NG * sizeof(double)));
    // all pairs in one launch (grid.y) unless the node buffer would exceed 1 GiB
    const size_t fw_per_pair = (size_t) nT * TL_NG * TL_NB * sizeof(double4);
    const int pairs_per_launch = (int) std::max<size_t>(1, std::min<size_t>((size_t) Np, ((size_t) 1 << 30) / fw_per_pair));
    CUDA_OK(kg_dmalloc(m, &d_fw, fw_per_pair * pairs_per_launch));
    m->tl_launches = 0;
    CUDA_OK(kg_dmalloc(m, &d_mtl, nitems * sizeof(double)));
    CUDA_OK(kg_dmalloc(m, &d_etl, nitems * sizeof(double)));
    CUDA_OK(kg_dmalloc(m, &d_status, nitems * sizeof(int)));
    CUDA_OK(cudaMemsetAsync(d_status, 0, nitems * sizeof(int), s));
    CUDA_OK(cudaMemcpyAsync(d_items, items.data(), nitems * sizeof(TlItem), cudaMemcpyHostToDevice, s));
    CUDA_OK(cudaMemcpyAsync(d_g, gn, sizeof(gn), cudaMemcpyHostToDevice, s));
    if (coeff_T) {
        CUDA_OK(kg_dmalloc(m, &d_coeff, (size_t) nT * Np * m->coeff_stride * sizeof(double)));
        CUDA_OK(cudaMemcpyAsync(d_coeff, coeff_T, (size_t) nT * Np * m->coeff_stride * sizeof(double), cudaMemcpyHostToDevice, s));
    }
    const size_t smem = (size_t) KS_ROWS * (2 + m->ntpad) * sizeof(double);
    for (int p0 = 0; p0 < Np; p0 += pairs_per_launch) {
        const int np = std::min(pairs_per_launch, Np - p0);
        const dim3 g1((nT * TL_NG + 255) / 256, np);
        const dim3 g1w((nT * TL_NG + 7) / 8, np);          // one warp per search
        const long long n2 = (long long) nT * TL_NG * TL_NB;
        const dim3 g2((unsigned) ((n2 + 255) / 256), np);
#define TL_LAUNCH(NT) do { \
        cudaFuncSetAttribute(tl_search_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem); \
        cudaFuncSetAttribute(tl_node_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem); \
        if (NT == 1 || !m->tl_warp_search) tl_search_kernel<NT><<<g1, 256, smem, s>>>(m->d_pairs, m->d_rows_by_pair, d_items, p0, nT, d_coeff, d_g, d_ilim, d_bmid, d_bmax, d_status); \
        else { cudaFuncSetAttribute(tl_search_warp_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem); \
               tl_search_warp_kernel<NT><<<g1w, 256, smem, s>>>(m->d_pairs, m->d_rows_by_pair, d_items, p0, nT, d_coeff, d_g, d_ilim, d_bmid, d_bmax, d_status); } \
        tl_node_kernel<NT><<<g2, 256, smem, s>>>(m->d_pairs, m->d_rows_by_pair, d_items, p0, nT, d_coeff, d_g, d_ilim, d_bmid, d_bmax, d_fw, d_status); } while (0)
        switch (m->ntpad) { case 1: TL_LAUNCH(1); break; case 2: TL_LAUNCH(2); break; case 4: TL_LAUNCH(4); break; default: TL_LAUNCH(6); break; }
#undef TL_LAUNCH
        tl_reduce_kernel<<<(np * nT + 3) / 4, 128, 0, s>>>(p0 * nT, np * nT, d_bmid, d_bmax, d_fw, d_mtl, d_etl);
        CUDA_OK(cudaGetLastError());
        m->tl_launches += 3;
    }
    std::vector<double> hm(nitems), he(nitems); std::vector<int> hs(nitems);
    CUDA_OK(cudaMemcpyAsync(hm.data(), d_mtl, nitems * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaMemcpyAsync(he.data(), d_etl, nitems * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaMemcpyAsync(hs.data(), d_status, nitems * sizeof(int), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    for (int p = 0; p < Np; p++) for (int t = 0; t < nT; t++) {
        mtl[t * Np + p] = hm[p * nT + t]; etl[t * Np + p] = he[p * nT + t];
        if (status) status[t * Np + p] = hs[p * nT + t];
    }
    for (void* q : {(void*) d_items, (void*) d_g, (void*) d_ilim, (void*) d_bmid, (void*) d_bmax, (void*) d_fw, (void*) d_mtl, (void*) d_etl, (void*) d_status, (void*) d_coeff}) m->dfree(q);
    return 0;
}
