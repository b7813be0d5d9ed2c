// kg_xsec.cuh -- classical collision cross sections and the classical second virial coefficient on the chi kernels
// (included at the end of kg_omega.cu).
//
// Reference: Spherical::cross_section (cpp/Spherical.cpp:104-127)
//     Q^(l)(E) = sigma^2 [ S(1e-7, b0; 20) + S(b0, b0 + 0.5; 50) + S_inf(b0 + 0.5, b0 + 1) ],  integrand (1 - cos^l chi(g, b)) b,
//     g = sqrt(E eps / (k 300 K)) at the dummy temperature 300 K, b0 = root of chi(b) by bisection on [0.9, 5] down to 1e-5
//     (bracket_root, cpp/Integration/Integration.h:221-229, .cpp:494-510), S = composite Simpson (.cpp:332-344), S_inf = Simpson
//     marching outwards in blocks of ten intervals whose width doubles as the blocks stop mattering (.cpp:346-373);
// and Spherical::second_virial (cpp/Spherical.cpp:140-160).
//
// One warp per energy. The bisection is a chain of 19 deflection angles, every lane walks it (uniform, no divergence); the Simpson
// nodes of the two fixed segments are spread over the lanes, their chi values parked in shared memory, and lane k sums them for
// l[k] in the reference's order; the outward march runs per l (its block sequence depends on the integrand), eleven lanes per block.
#pragma once

namespace {

constexpr int XS_MAXL = 8;
constexpr double AVOGADRO_REF = 8.31446261815324 / 1.38064852e-23;        // cpp/global_params.h:5-6
constexpr int XS_WARPS = 4;

template<int NT>
__global__ void __launch_bounds__(32 * XS_WARPS) xsec_kernel(PairParams pp, const double* __restrict__ rows, const double* __restrict__ coeff,
                                                             int nE, const double* __restrict__ E_red, int nl, const int* __restrict__ lval,
                                                             double* __restrict__ out, int* __restrict__ status) {
    extern __shared__ double srows[];
    __shared__ double s_chi[XS_WARPS][64];
    constexpr int W = 2 + NT;
    for (int i = threadIdx.x; i < KS_ROWS * W; i += blockDim.x) srows[i] = rows[i];
    __syncthreads();
    const NodeRows<NT> NR{(unsigned) __cvta_generic_to_shared(srows), srows, rows, KS_ROWS};
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int e = blockIdx.x * XS_WARPS + wid;
    if (e >= nE) return;
    const double T = 300.0;                                                 // dummy temperature, cpp/Spherical.cpp:116
    const double g = sqrt(E_red[e] * pp.eps / (BOLTZMANN * T));
    int st = 0; unsigned nodes = 0; double R;
    const double I_lim = lane_theta_integral<NT>(pp, coeff, NR, T, g, 100.0, R, nodes, st);
    auto chi = [&](double b) { return chi_with_lim<NT>(pp, coeff, NR, T, g, b, I_lim, R, nodes, st); };
    // b0: bracket_root(chi, 0.9, 5, dtol = 1e-5, ftol = -1)
    double x0 = 0.9, x1 = 5.0, f0 = chi(x0), f1 = chi(x1);
    if (f0 * f1 > 0) st |= 64;                                              // "Bracket solver: Initial values have same sign!"
    while (fabs(x0 - x1) > 1e-5) {
        const double xm = 0.5 * (x0 + x1), fm = chi(xm);
        if (f0 * fm > 0) { x0 = xm; f0 = fm; } else { x1 = xm; f1 = fm; }
    }
    if (fabs(f0) > fabs(f1)) { const double t = x0; x0 = x1; x1 = t; }
    const double b0 = x0;
    // integrand (1 - cos^l chi) b for this lane's l (lanes >= nl idle in the sums)
    const int myl = (lane < nl) ? lval[lane] : 1;
    auto integrand = [&](double chi_v, double b) {
        const double c = cos(chi_v);
        double p = 1.0;
        for (int k = 0; k < myl; k++) p *= c;
        return (1. - p) * b;
    };
    // composite Simpson over [a, bN] with N (even) intervals: node i at the reference's expressions, chi of all nodes by the lanes
    auto simpson_seg = [&](double a, double bN, int N) {
        const double dx = (bN - a) / N;
        auto node = [&](int i) {                                            // i = 0 .. N in evaluation-independent indexing
            if (i == 0) return a;
            if (i == N) return bN;
            if (i == N - 1) return bN - dx;
            return a + i * dx;                                              // a + (2k-1) dx and a + (2k) dx
        };
        for (int i0 = 0; i0 <= N; i0 += 32) {
            const int i = i0 + lane;
            if (i <= N) s_chi[wid][i] = chi(node(i));
        }
        __syncwarp();
        double val = integrand(s_chi[wid][0], a) + integrand(s_chi[wid][N], bN);
        for (int k = 1; k <= N / 2 - 1; k++)
            val += 4. * integrand(s_chi[wid][2 * k - 1], a + (2 * k - 1) * dx) + 2. * integrand(s_chi[wid][2 * k], a + (2 * k) * dx);
        val += 4. * integrand(s_chi[wid][N - 1], bN - dx);
        __syncwarp();
        return val * (dx / 3.);
    };
    double I = simpson_seg(1e-7, b0, 20);
    I += simpson_seg(b0, b0 + 0.5, 50);
    // simpson_inf(integrand, b0 + 0.5, b0 + 1, tol = 1e-8, I): per l, its own block sequence
    const double tol = 1e-8;
    double result = 0.0;
    for (int k = 0; k < nl; k++) {
        const int lk = lval[k];
        auto f_l = [&](double chi_v, double b) { const double c = cos(chi_v); double p = 1.0; for (int q = 0; q < lk; q++) p *= c; return (1. - p) * b; };
        auto block = [&](double a, double bN) {                            // simpson(func, a, bN, 10), all lanes get the value
            const double dx = (bN - a) / 10;
            double fv = 0.0;
            if (lane <= 10) {
                const double b = (lane == 0) ? a : (lane == 10) ? bN : (lane == 9) ? bN - dx : a + lane * dx;
                fv = f_l(chi(b), b);
            }
            double v[11];
#pragma unroll
            for (int i = 0; i < 11; i++) v[i] = __shfl_sync(full, fv, i);
            double val = v[0] + v[10];
            for (int i = 1; i <= 4; i++) val += 4. * v[2 * i - 1] + 2. * v[2 * i];
            val += 4. * v[9];
            return val * (dx / 3.);
        };
        double Il = __shfl_sync(full, I, k);
        double xa = b0 + 0.5, xb = b0 + 1.0;
        Il += block(xa, xb);
        double dx = (xb - xa) / 10.;
        double part_tol = tol * 1e6;
        int guard = 0;
        do {
            if (isnan(Il)) { st |= KG_STATUS_NAN; break; }                  // "Encountered NAN in simpson_inf"
            xa = xb;
            xb += 10 * dx;
            const double Ip = block(xa, xb);
            Il += Ip;
            const double conv = fabs(Ip / Il);
            if (conv < tol) break;
            if (conv < part_tol) { dx *= 2; part_tol *= 0.1; }
        } while (part_tol > tol && guard++ < 10000);
        if (lane == k) result = Il;
    }
    if (lane < nl) out[(size_t) e * nl + lane] = (pp.sigma * pp.sigma) * result;
    if (lane == 0 && status) status[e] = st;
}

// Spherical::second_virial (cpp/Spherical.cpp:140-160) for potentials whose get_sigma_eff is sigma (MieKinGas and the generic models);
// one thread per temperature
template<int NT>
__global__ void second_virial_kernel(PairParams pp, const double* __restrict__ coeff_T, int coeff_stride, int nT, const double* __restrict__ T,
                                     double* __restrict__ out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nT) return;
    Lane<NT> c;
    make_lane<NT>(c, pp, coeff_T ? coeff_T + (size_t) t * coeff_stride : nullptr, 1.0 / BOLTZMANN, 1.0, 1.0);     // kT g^2 = 1: pot_eval returns joules
    const double Tk = T[t];
    auto u_div_k = [&](double r) { double U, dU; pot_eval<NT>(c, r, U, dU); return U / BOLTZMANN; };        // r in units of sigma
    auto f = [&](double r) { return (r * r) * (exp(-u_div_k(r) / Tk) - 1); };
    auto simpson = [&](double a, double b, int N) {
        const double dx = (b - a) / N;
        double val = f(a) + f(b);
        for (int i = 1; i <= N / 2 - 1; i++) val += 4. * f(a + (2 * i - 1) * dx) + 2. * f(a + (2 * i) * dx);
        val += 4. * f(b - dx);
        return val * (dx / 3.);
    };
    double r0 = 0.9;
    int guard = 0;
    while (exp(-u_div_k(r0) / Tk) > 1e-8 && guard++ < 100) r0 -= 0.01;
    const double r1 = 2;
    double I = -((r0 * r0 * r0) / 3) + simpson(r0, r1, 100);
    {   // simpson_inf(integrand, r1, 3): tol = 1e-8, I = 0
        const double tol = 1e-8;
        double xa = r1, xb = 3, Ii = simpson(xa, xb, 10), dx = (xb - xa) / 10., part_tol = tol * 1e6;
        guard = 0;
        do {
            xa = xb; xb += 10 * dx;
            const double Ip = simpson(xa, xb, 10);
            Ii += Ip;
            const double conv = fabs(Ip / Ii);
            if (conv < tol) break;
            if (conv < part_tol) { dx *= 2; part_tol *= 0.1; }
        } while (part_tol > tol && guard++ < 10000);
        I += Ii;
    }
    out[t] = -0.5 * 4 * PI_REF * (pp.sigma * pp.sigma * pp.sigma) * AVOGADRO_REF * I;
}

} // namespace

extern "C" int kg_cross_sections(kg_model* m, int pair, int nE, const double* E_red, int nl, const int32_t* l, const double* coeff,
                                 double* out, int32_t* status) {
    KG_TRACE();
    if (!m) return -1;
    auto set_error = [&](const std::string& s) { m->err = s; };
    if (pair < 0 || pair >= m->npairs || nE < 0 || !E_red || nl < 1 || nl > XS_MAXL || !l || !out) { set_error("kg_cross_sections: invalid argument"); return -1; }
    if (m->generic) { set_error("kg_cross_sections: power-law potentials only"); return -1; }
    if (m->coeff_T_required && !coeff) { set_error(KG_NEED_COEFF); return -1; }
    for (int k = 0; k < nl; k++) if (l[k] < 1 || l[k] > 64) { set_error("kg_cross_sections: l out of range"); return -1; }
    if (nE == 0) return 0;
    CUDA_OK(cudaSetDevice(m->device));
    double *dE = nullptr, *dout = nullptr, *dcoef = nullptr; int *dl = nullptr, *dst = nullptr;
    CUDA_OK(kg_dmalloc(m, &dE, nE * sizeof(double))); CUDA_OK(kg_dmalloc(m, &dout, (size_t) nE * nl * sizeof(double)));
    CUDA_OK(kg_dmalloc(m, &dl, nl * sizeof(int))); CUDA_OK(kg_dmalloc(m, &dst, nE * sizeof(int)));
    cudaStream_t s = m->stream;
    CUDA_OK(cudaMemcpyAsync(dE, E_red, nE * sizeof(double), cudaMemcpyHostToDevice, s));
    CUDA_OK(cudaMemcpyAsync(dl, l, nl * sizeof(int), cudaMemcpyHostToDevice, s));
    if (coeff) {
        CUDA_OK(kg_dmalloc(m, &dcoef, KG_MAX_TERMS * sizeof(double)));
        CUDA_OK(cudaMemcpyAsync(dcoef, coeff, m->descs[pair].nterms * sizeof(double), cudaMemcpyHostToDevice, s));
    }
    const size_t smem = (size_t) KS_ROWS * (2 + m->ntpad) * sizeof(double);
    const int grid = (nE + XS_WARPS - 1) / XS_WARPS;
#define XS_LAUNCH(NT) do { cudaFuncSetAttribute(xsec_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem); \
        xsec_kernel<NT><<<grid, 32 * XS_WARPS, smem, s>>>(m->pairs[pair], m->d_rows[pair], dcoef, nE, dE, nl, dl, dout, dst); } while (0)
    switch (m->ntpad) { case 2: XS_LAUNCH(2); break; case 4: XS_LAUNCH(4); break; default: XS_LAUNCH(6); break; }
#undef XS_LAUNCH
    CUDA_OK(cudaGetLastError());
    CUDA_OK(cudaMemcpyAsync(out, dout, (size_t) nE * nl * sizeof(double), cudaMemcpyDeviceToHost, s));
    std::vector<int> st(nE);
    CUDA_OK(cudaMemcpyAsync(st.data(), dst, nE * sizeof(int), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    if (status) for (int k = 0; k < nE; k++) status[k] = st[k];
    for (void* q : {(void*) dE, (void*) dout, (void*) dl, (void*) dst, (void*) dcoef}) m->dfree(q);
    for (int k = 0; k < nE; k++) if (st[k] & 64) { set_error("Bracket solver: Initial values have same sign!"); return -4; }
    return 0;
}

extern "C" int kg_second_virial(kg_model* m, int pair, int nT, const double* T, const double* coeff_T, double* out) {
    KG_TRACE();
    if (!m) return -1;
    auto set_error = [&](const std::string& s) { m->err = s; };
    if (pair < 0 || pair >= m->npairs || nT < 0 || !T || !out) { set_error("kg_second_virial: invalid argument"); return -1; }
    if (m->generic) { set_error("kg_second_virial: power-law potentials only"); return -1; }
    if (m->coeff_T_required || coeff_T) { set_error("kg_second_virial: temperature-dependent coefficients are not supported (the reference scales r by sigma_eff(T))"); return -1; }
    if (nT == 0) return 0;
    CUDA_OK(cudaSetDevice(m->device));
    double *dT = nullptr, *dout = nullptr, *dcoef = nullptr;
    cudaStream_t s = m->stream;
    CUDA_OK(kg_dmalloc(m, &dT, nT * sizeof(double))); CUDA_OK(kg_dmalloc(m, &dout, nT * sizeof(double)));
    CUDA_OK(cudaMemcpyAsync(dT, T, nT * sizeof(double), cudaMemcpyHostToDevice, s));
    if (coeff_T) {      // [nT][KG_MAX_TERMS] of this pair
        CUDA_OK(kg_dmalloc(m, &dcoef, (size_t) nT * KG_MAX_TERMS * sizeof(double)));
        CUDA_OK(cudaMemcpyAsync(dcoef, coeff_T, (size_t) nT * KG_MAX_TERMS * sizeof(double), cudaMemcpyHostToDevice, s));
    }
    const int grid = (nT + 63) / 64;
    switch (m->ntpad) {
        case 2: second_virial_kernel<2><<<grid, 64, 0, s>>>(m->pairs[pair], dcoef, KG_MAX_TERMS, nT, dT, dout); break;
        case 4: second_virial_kernel<4><<<grid, 64, 0, s>>>(m->pairs[pair], dcoef, KG_MAX_TERMS, nT, dT, dout); break;
        default: second_virial_kernel<6><<<grid, 64, 0, s>>>(m->pairs[pair], dcoef, KG_MAX_TERMS, nT, dT, dout); break;
    }
    CUDA_OK(cudaGetLastError());
    CUDA_OK(cudaMemcpyAsync(out, dout, nT * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    for (void* q : {(void*) dT, (void*) dout, (void*) dcoef}) m->dfree(q);
    return 0;
}
