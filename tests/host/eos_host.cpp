// Host build of the product's SAFT-VR Mie header (kineticgas_b200/csrc/kg_saft.cuh), for the CPU test suite: the same generic code the
// device kernels instantiate, evaluated in hyper-dual numbers on the host and compared with fixtures produced by the thermopack wheel
// (tests/test_eos_host.py). Test infrastructure only -- the product path is kg_eos.cu on the GPU.
#define KG_HD inline
#include "../../kineticgas_b200/csrc/kg_saft.cuh"
#include <cstring>
using namespace kgsaft;

static void build_model(int nc, const double* m, const double* sigma, const double* epsk, const double* la, const double* lr, Model& M) {
    memset(&M, 0, sizeof(M));
    M.nc = nc;
    for (int i = 0; i < nc; i++) M.m[i] = m[i];
    for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) {
        PairC& q = M.p[i][j];
        q.sigma = sigma[i * nc + j]; q.epsk = epsk[i * nc + j]; q.la = la[i * nc + j]; q.lr = lr[i * nc + j];
        init_pair(q);
    }
}

extern "C" {
// d_ii [m] at T
void kgsaft_host_diameters(int nc, const double* m, const double* sigma, const double* epsk, const double* la, const double* lr, double T, double* d) {
    Model M; build_model(nc, m, sigma, epsk, la, lr, M);
    for (int i = 0; i < nc; i++) d[i] = bh_diameter(M.p[i][i], T);
}
// F = A_res/(RT) of the selected terms with all first and second derivatives in y = (n_1..n_nc, V): grad[nc+1], hess[(nc+1)^2]
void kgsaft_host_helmholtz(int nc, const double* m, const double* sigma, const double* epsk, const double* la, const double* lr,
                           double T, double V, const double* n, int terms, double* F, double* grad, double* hess) {
    Model M; build_model(nc, m, sigma, epsk, la, lr, M);
    TPre P; init_T(M, T, P);
    const int nv = nc + 1;
    for (int a = 0; a < nv; a++) for (int b = a; b < nv; b++) {
        HD y[MAXC + 1];
        for (int k = 0; k < nc; k++) y[k] = HD(n[k]);
        y[nc] = HD(V);
        y[a].a = 1.0; y[b].b = 1.0;
        const HD f = helmholtz<HD>(M, P, y[nc], y, terms);
        *F = f.v; grad[a] = f.a; grad[b] = f.b;
        hess[a * nv + b] = hess[b * nv + a] = f.ab;
    }
}
}
