// kg_omega.cu -- batched collision integrals Omega^(l,r)_ij(T) on B200 (sm_100a), behind the C-ABI of include/kgcuda.h.
//
// Reference path replaced (thermotools/KineticGas @ 75848b1):
//   KineticGas::precompute_omega fan-out              cpp/KineticGas_mthr.cpp:40-46,58-143
//   Spherical::omega -> w_integral -> integrate2d     cpp/Spherical.cpp:37-102, cpp/Integration/Integration.cpp:145-307
//   chi / theta / get_R / tanh_sinh                   cpp/Spherical.cpp:259-465, cpp/Integration/Integration.cpp:417-510
//
// Design (not a translation): the reference evaluates the integrand lazily, point by point, inside a recursive
// triangulation, recomputing the deflection angle for every (l,r). Here one *item* = (pair, T) owns a lattice of
// cos(chi) in HBM shared by all (l,r); the adaptive mesh is a decision tree over that lattice and is walked
// level-synchronously:
//     mesh_level    : one thread per cell of the current refinement level (breadth-first over the tree): the reference's
//                     exact refinement predicate decides leaf (integrate two triangles) or split (emit the sub-domain
//                     sweep's cells to the next level and mark the lattice nodes they read)
//     count/scan/fill: compaction of the newly marked nodes of all items into one work list (sorted by pair, item,
//                     column, row -> neighbouring lanes integrate neighbouring trajectories)
//     theta_setup   : one thread per work entry: distance of closest approach (get_R), stores A_t and (R/b)^2 (power-law potentials)
//     theta_kernel  : one lane per theta integral (kg_device.cuh), persistent CTAs, 32-entry tiles per warp, tanh-sinh rows of
//                     the pair staged in shared memory; strided sum with Euler-Maclaurin end corrections, exact replay of the stop
//     finalize      : cos(chi) from I - I_lim, publish nodes as known
// repeated until no mesh thread meets an unknown node (depth of the refinement tree + 1 passes). Only nodes the
// reference's mesh touches are ever integrated.
#include "kg_device.cuh"
#include "kg_internal.cuh"
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <vector>
#include <algorithm>

using namespace kg;

static thread_local std::string g_create_error;

namespace {

constexpr int THETA_THREADS = 512;
constexpr int KS_ROWS = 2560;          // tanh-sinh rows staged in shared memory (rest read through L1/L2)
constexpr unsigned LIM_ROW = 0xFFFFu;  // work-list entry that is a theta_lim evaluation of its column

struct Regime {            // the T-dependent lattice of Spherical::w_integral (cpp/Spherical.cpp:60-84)
    int nx, ny;            // allocated lattice incl. the +2-step over-reach of the refinement probes
    int Nx_end, Ny_end, rlg, rlb, nwords;
    int ldg, ldb;          // dg = 2^ldg, db = 2^ldb (all lattice steps are powers of two: exact scaling)
    int nx_cut, ny_cut;    // first Nx with g > 1.5, first Ny with b > 2.5 (integrand cut-off, cpp/Spherical.cpp:99)
    double dg, db, lim;
};

struct Item {              // one (pair, T)
    int pair, regime;
    double T;
    double pref;           // Omega = pref * W (cpp/Spherical.cpp:46-47)
    long long tab_off;     // into the cos(chi) lattice pool   [nx][ny]
    long long mask_off;    // into the want/known bit pools     [nx][nwords]
    long long col_off;     // into the per-column pools (I_lim, have_lim, task index) [nx]
    int coeff_off;         // into coeff pool (doubles) or -1
    int pad;
    long long cell0_off;   // (unused, kept for alignment)
};

struct WorkEntry { int item; unsigned short nx, ny; };

__device__ __forceinline__ double lattice_coord(int n, double d) { return 1e-7 + n * d; }   // cpp/Spherical.cpp:60

// ------------------------------------------------------------------------------------------------------------
// mesh kernels: the reference's recursive triangulation as a breadth-first walk over cell lists
// ------------------------------------------------------------------------------------------------------------
// A cell is one integration_step of the reference (cpp/Integration/Integration.cpp:145-236): entry corner (Nx,Ny), signed
// x step, y step. The refinement trees of the (l,r) of one item share most of their geometry, so a cell carries the
// SET of (l,r) that reach it: the lattice is read once per cell and the predicate is evaluated per member.
// lo: [0,22) item, [22,30) Nx, [30,40) Ny, [40,43) log2|sx|, 43 sx<0, [44,47) log2 sy.   hi: (l,r) membership bits.
struct Cell { unsigned long long lo, hi; };
__host__ __device__ __forceinline__ unsigned long long pack_cell(unsigned item, int Nx, int Ny, int sx, int sy) {
    const int ax = sx < 0 ? -sx : sx;
    unsigned lx = 0, ly = 0;
    while ((1 << lx) < ax) lx++;
    while ((1 << ly) < sy) ly++;
    return (unsigned long long) item | ((unsigned long long) Nx << 22) | ((unsigned long long) Ny << 30) | ((unsigned long long) lx << 40)
         | ((unsigned long long) (sx < 0) << 43) | ((unsigned long long) ly << 44);
}

struct MeshView {
    const double* tab;
    unsigned* want;
    int ny, nwords, nx_cut, ny_cut, ldg, ldb;
    double db, lim, cospi;
};

__device__ __forceinline__ double pow2(int e) { return __longlong_as_double((long long) (1023 + e) << 52); }

__device__ __forceinline__ bool node_is_free(const MeshView& v, int Nx, int Ny) {
    // no chi needed: integrand cut-off b > 2.5 && g > 1.5 (cpp/Spherical.cpp:99) or b/sigma < 1e-3 -> chi = PI (:267)
    return ((Ny >= v.ny_cut) && (Nx >= v.nx_cut)) || Ny == 0;
}

__device__ __forceinline__ void mesh_mark(const MeshView& v, int Nx, int Ny) {
    if (node_is_free(v, Nx, Ny)) return;
    const int w = Nx * v.nwords + (Ny >> 5);
    const unsigned bit = 1u << (Ny & 31);
    if (!(v.want[w] & bit)) atomicOr(&v.want[w], bit);
}

// every lattice node a cell reads: the 7 probes of the refinement test (Integration.cpp:159-165) and, for a sweep in -x,
// the two exit corners of its triangles
__device__ __forceinline__ void mesh_mark_cell(const MeshView& v, int Nx, int Ny, int sx, int sy) {
    const int ax = abs(sx);
    mesh_mark(v, Nx, Ny); mesh_mark(v, Nx + ax, Ny); mesh_mark(v, Nx + 2 * ax, Ny);
    mesh_mark(v, Nx, Ny + sy); mesh_mark(v, Nx, Ny + 2 * sy);
    mesh_mark(v, Nx + ax, Ny + sy); mesh_mark(v, Nx + 2 * ax, Ny + 2 * sy);
    if (sx < 0) { mesh_mark(v, Nx + sx, Ny); mesh_mark(v, Nx + sx, Ny + sy); }
}

// cos(chi) at lattice node (Nx,Ny) (known by construction). Cut-off nodes (integrand identically 0, cpp/Spherical.cpp:99) return
// exactly 1.0: 1 - 1^l = 0 for every l, so the integrand formula below gives +0.0 there without a special case.
__device__ __forceinline__ double mesh_cos(const MeshView& v, int Nx, int Ny) {
    if ((Ny >= v.ny_cut) && (Nx >= v.nx_cut)) return 1.0;
    if (Ny == 0) return v.cospi;
    const double t = v.tab[(long long) Nx * v.ny + Ny];
    return (t != t) ? 1.0 : t;                       // a NaN table entry (failed root finding, flagged in status) contributes 0 as before
}

// integrand value, cpp/Spherical.cpp:96-102: gf = 2 exp(-g^2) g^(2r+3), cl = cos(chi)^l, b = impact parameter / sigma
__device__ __forceinline__ double mesh_f(double gf, double cl, double b) {
    return __dmul_rn(__dmul_rn(gf, __dsub_rn(1.0, cl)), b);
}

__device__ __forceinline__ MeshView make_view(const Item& item, const Regime& rg, const double* tab_pool, unsigned* want_pool, double cospi) {
    MeshView v;
    v.tab = tab_pool ? tab_pool + item.tab_off : nullptr;
    v.want = want_pool + item.mask_off;
    v.ny = rg.ny; v.nwords = rg.nwords;
    v.nx_cut = rg.nx_cut; v.ny_cut = rg.ny_cut; v.ldg = rg.ldg; v.ldb = rg.ldb;
    v.db = rg.db; v.lim = rg.lim; v.cospi = cospi;
    return v;
}

// tasks[col_off + nx] = (item, nx) for every lattice column of every active item (col_off counts columns in item order)
__global__ void build_tasks_kernel(const Item* __restrict__ items, const Regime* __restrict__ regimes, const int* __restrict__ active, int nactive,
                                   int2* __restrict__ tasks) {
    const int a = blockIdx.x;
    if (a >= nactive) return;
    const int it = active[a];
    const Item item = items[it];
    const int nx = regimes[item.regime].nx;
    for (int k = threadIdx.x; k < nx; k += blockDim.x) tasks[item.col_off + k] = make_int2(it, k);
}

// level 0: the coarse cells of every active item in the reference's sweep order (Integration.cpp:240-280: row k runs +x
// for even k, -x for odd k), owned by all requested (l,r) of the item, and the lattice nodes they read. grid.x = active items.
__global__ void __launch_bounds__(128) mesh_init_kernel(const Item* __restrict__ items, const Regime* __restrict__ regimes,
                                                        unsigned* __restrict__ want_pool, const int* __restrict__ active,
                                                        const unsigned long long* __restrict__ active_mask,
                                                        const long long* __restrict__ active_cell_off, Cell* __restrict__ cells) {
    const int it = active[blockIdx.x];
    const Item item = items[it];
    const Regime rg = regimes[item.regime];
    MeshView v = make_view(item, rg, nullptr, want_pool, 0.0);
    const int ncx = rg.Nx_end / rg.rlg, ncy = rg.Ny_end / rg.rlb, nc = ncx * ncy;
    const long long c0 = active_cell_off[blockIdx.x];
    const unsigned long long mask = active_mask[blockIdx.x];
    for (int cell = threadIdx.x; cell < nc; cell += blockDim.x) {
        const int row = cell / ncx, col = cell % ncx;
        const int d = (row & 1) ? -1 : 1;
        const int ex = ((row & 1) ? rg.Nx_end : 0) + col * d * rg.rlg, ey = row * rg.rlb;
        cells[c0 + cell] = Cell{pack_cell((unsigned) it, ex, ey, d * rg.rlg, rg.rlb), mask};
        mesh_mark_cell(v, ex, ey, d * rg.rlg, rg.rlb);
    }
}

// children of one split class: the sub-domain sweep of integrate_adaptive (Integration.cpp:172-202) over the parent cell,
// origin (Nx,Ny), far corner (Nx+sx, Ny+sy); row k runs in direction sgn for even k, -sgn for odd k
__device__ __forceinline__ int mesh_emit(const MeshView& v, Cell* __restrict__ out, int o, unsigned item, unsigned long long members,
                                         int Nx, int Ny, int sx, int sy, int sub_sx, int sub_sy) {
    const int ax = abs(sx), asx = abs(sub_sx), ncols = ax / asx, nrows = sy / sub_sy, sgn = (sx > 0) ? 1 : -1;
    for (int k = 0; k < nrows; k++) {
        const int d = (k & 1) ? -sgn : sgn;
        const int xs = (k & 1) ? Nx + sx : Nx;
        for (int j = 0; j < ncols; j++) {
            const int cxn = xs + j * d * asx, cyn = Ny + k * sub_sy;
            out[o++] = Cell{pack_cell(item, cxn, cyn, d * asx, sub_sy), members};
            mesh_mark_cell(v, cxn, cyn, d * asx, sub_sy);
        }
    }
    return o;
}

// One level of the refinement trees: every member (l,r) of a cell either integrates it (leaf) or splits it; members that
// split the same way share the child cells, which go to cells_out with their nodes marked.
__global__ void __launch_bounds__(256, 4) mesh_level_kernel(const Item* __restrict__ items, const Regime* __restrict__ regimes,
                                                         const double* __restrict__ tab_pool, unsigned* __restrict__ want_pool,
                                                         const double* __restrict__ gfac_pool, const int* __restrict__ lr_l,
                                                         const int* __restrict__ lr_r_index, int nr_distinct, int gfac_stride, double cospi,
                                                         int nlr, const Cell* __restrict__ cells_in, int n_in, Cell* __restrict__ cells_out,
                                                         int* __restrict__ n_out, int cap_out, double* __restrict__ W, int* __restrict__ flags) {
    // cos(chi) of the nine nodes a cell reads, [node][thread]: the member loop keeps their l-th powers in registers and
    // multiplies them up from here as l grows along the (l,r) list (l-major lists never restart)
    __shared__ double s_c[9][256];
    const unsigned full = 0xffffffffu;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = e < n_in;
    Cell c = Cell{0ull, 0ull};
    if (active) c = cells_in[e];
    const unsigned item_id = (unsigned) (c.lo & 0x3FFFFFu);
    const int Nx = (int) ((c.lo >> 22) & 0xFF), Ny = (int) ((c.lo >> 30) & 0x3FF);
    const int lx = (int) ((c.lo >> 40) & 7), ly = (int) ((c.lo >> 44) & 7);
    const int ax = 1 << lx, sy = 1 << ly;
    const int sx = ((c.lo >> 43) & 1) ? -ax : ax;
    MeshView v = {};
    double c0 = 0, c1x = 0, c2x = 0, c1y = 0, c2y = 0, c11 = 0, c22 = 0, cXb = 0, cXt = 0, b0 = 0, b1 = 0, b2 = 0;
    int regime = 0;
    if (active) {
        const Item item = items[item_id];
        const Regime rg = regimes[item.regime];
        regime = item.regime;
        v = make_view(item, rg, tab_pool, want_pool, cospi);
        c0 = mesh_cos(v, Nx, Ny); c1x = mesh_cos(v, Nx + ax, Ny); c1y = mesh_cos(v, Nx, Ny + sy); c11 = mesh_cos(v, Nx + ax, Ny + sy);
        // a cell at the finest step in both directions cannot split: it is integrated whatever the refinement test says, so the
        // three far probes of the test (and the test) are skipped for it -- most cells of the last level are of this kind
        if (ax > 1 || sy > 1) { c2x = mesh_cos(v, Nx + 2 * ax, Ny); c2y = mesh_cos(v, Nx, Ny + 2 * sy); c22 = mesh_cos(v, Nx + 2 * ax, Ny + 2 * sy); }
        if (sx < 0) { cXb = mesh_cos(v, Nx + sx, Ny); cXt = mesh_cos(v, Nx + sx, Ny + sy); }
        b0 = lattice_coord(Ny, v.db); b1 = lattice_coord(Ny + sy, v.db); b2 = lattice_coord(Ny + 2 * sy, v.db);
    }
    {
        const int tx = threadIdx.x;
        s_c[0][tx] = c0; s_c[1][tx] = c1x; s_c[2][tx] = c2x; s_c[3][tx] = c1y; s_c[4][tx] = c2y; s_c[5][tx] = c11; s_c[6][tx] = c22;
        s_c[7][tx] = cXb; s_c[8][tx] = cXt;
    }
    int lcur = 1;                                    // c0 .. cXt hold cos^lcur from here on
    // hx, hy are powers of two, so the reference's divisions are exact scalings and are done as multiplications
    const int ehx = v.ldg + lx, ehy = v.ldb + ly;
    const double hx = pow2(ehx), hy = pow2(ehy), hx2 = pow2(2 * ehx), hy2 = pow2(2 * ehy);
    const double ihx2 = pow2(-2 * ehx), ihy2 = pow2(-2 * ehy), ihxy = pow2(-(1 + ehx + ehy));
    const double hs = __dadd_rn(hx, hy);
    const double hs3 = __dmul_rn(__dmul_rn(hs, hs), hs);
    const double hx3 = __dmul_rn(hx2, hx), hy3 = __dmul_rn(hy2, hy);
    const bool can_x = ax > 1, can_y = sy > 1;
    const unsigned item0 = __shfl_sync(full, item_id, 0);
    const bool uniform = __all_sync(full, !active || item_id == item0);
    unsigned long long mx = 0, my = 0, mb = 0;
    const unsigned um_lo = __reduce_or_sync(full, (unsigned) c.hi), um_hi = __reduce_or_sync(full, (unsigned) (c.hi >> 32));
    unsigned long long um = ((unsigned long long) um_hi << 32) | um_lo;
    while (um) {
        const int lr = __ffsll((long long) um) - 1;
        um &= um - 1;
        double contrib = 0.0;
        const int l = lr_l[lr];
        if (l != lcur) {                                 // warp-uniform: same successive products as pow by repeated multiplication
            const int tx = threadIdx.x;
            if (l < lcur) {
                c0 = s_c[0][tx]; c1x = s_c[1][tx]; c2x = s_c[2][tx]; c1y = s_c[3][tx]; c2y = s_c[4][tx]; c11 = s_c[5][tx]; c22 = s_c[6][tx];
                cXb = s_c[7][tx]; cXt = s_c[8][tx];
                lcur = 1;
            }
            for (; lcur < l; lcur++) {
                c0 *= s_c[0][tx]; c1x *= s_c[1][tx]; c2x *= s_c[2][tx]; c1y *= s_c[3][tx]; c2y *= s_c[4][tx]; c11 *= s_c[5][tx]; c22 *= s_c[6][tx];
                cXb *= s_c[7][tx]; cXt *= s_c[8][tx];
            }
        }
        if (active && ((c.hi >> lr) & 1ull)) {
            const double* gf = gfac_pool + ((long long) regime * nr_distinct + lr_r_index[lr]) * gfac_stride;
            const double g0 = gf[Nx], g1 = gf[Nx + ax];
            const double f = mesh_f(g0, c0, b0), f1x = mesh_f(g1, c1x, b0), f1y = mesh_f(g0, c1y, b1), f1x1y = mesh_f(g1, c11, b1);
            bool split = false;
            if (can_x || can_y) {
            const double g2 = gf[Nx + 2 * ax];
            const double f2x = mesh_f(g2, c2x, b0), f2y = mesh_f(g0, c2y, b2), f2x2y = mesh_f(g2, c22, b2);
            // second-difference estimator in the reference's operation order, no FMA contraction (decision predicate)
            const double fxx = __dmul_rn(__dadd_rn(__dsub_rn(f, __dmul_rn(2.0, f1x)), f2x), ihx2);
            const double fyy = __dmul_rn(__dadd_rn(__dsub_rn(f, __dmul_rn(2.0, f1y)), f2y), ihy2);
            double t = __dadd_rn(__dsub_rn(f2x2y, __dmul_rn(2.0, f1x1y)), f);
            t = __dsub_rn(t, __dmul_rn(hx2, fxx));
            t = __dsub_rn(t, __dmul_rn(hy2, fyy));
            const double fxy = __dmul_rn(t, ihxy);
            const double dblder = __dadd_rn(__dadd_rn(fabs(fxx), fabs(fyy)), fabs(fxy));
            if (__dmul_rn(dblder, hs3) > v.lim) {
                split = true;
                const double cx = __dmul_rn(fabs(fxx), hx3);
                const double cy = __dmul_rn(fabs(fyy), hy3);
                const unsigned long long bit = 1ull << lr;
                if (cx > v.lim && (cy < v.lim) && (fabs(fxy) < v.lim) && can_x) mx |= bit;
                else if (cy > v.lim && (cx < v.lim) && (fabs(fxy) < v.lim) && can_y) my |= bit;
                else if (can_x && can_y) mb |= bit;
                else if (can_x) mx |= bit;
                else my |= bit;
            }
            }
            if (!split) {
                // leaf: the two triangles the sweep adds for this cell, (Eb,Et,Xb) and (Et,Xb,Xt) (Integration.cpp:218-234);
                // a linear interpolant integrates to area * mean of its corner values
                double fXb = f1x, fXt = f1x1y;
                if (sx < 0) { const double gm = gf[Nx + sx]; fXb = mesh_f(gm, cXb, b0); fXt = mesh_f(gm, cXt, b1); }
                contrib = (hx * hy / 6.0) * (f + 2.0 * f1y + 2.0 * fXb + fXt);
            }
        }
        // leaf sums: one atomic per warp when the warp belongs to one item, else one per lane
        if (uniform) {
            double sum = contrib;
            for (int off = 16; off > 0; off >>= 1) sum += __shfl_down_sync(full, sum, off);
            if ((threadIdx.x & 31) == 0 && sum != 0.0) atomicAdd(&W[(size_t) item0 * nlr + lr], sum);
        }
        else if (contrib != 0.0) atomicAdd(&W[(size_t) item_id * nlr + lr], contrib);
    }
    // children: warp-aggregated allocation in cells_out
    const int nchild = (mx ? 2 : 0) + (my ? 2 : 0) + (mb ? 4 : 0);
    int pre = nchild;
    for (int off = 1; off < 32; off <<= 1) { const int nb = __shfl_up_sync(full, pre, off); if ((threadIdx.x & 31) >= off) pre += nb; }
    const int warp_total = __shfl_sync(full, pre, 31);
    int base = 0;
    if ((threadIdx.x & 31) == 31 && warp_total > 0) base = atomicAdd(n_out, warp_total);
    base = __shfl_sync(full, base, 31);
    if (nchild > 0) {
        int o = base + pre - nchild;
        if (o + nchild > cap_out) { flags[1] = 1; return; }
        if (mx) o = mesh_emit(v, cells_out, o, item_id, mx, Nx, Ny, sx, sy, sx / 2, sy);
        if (my) o = mesh_emit(v, cells_out, o, item_id, my, Nx, Ny, sx, sy, sx, sy / 2);
        if (mb) o = mesh_emit(v, cells_out, o, item_id, mb, Nx, Ny, sx, sy, sx / 2, sy / 2);
    }
}

// ------------------------------------------------------------------------------------------------------------
// work-list compaction: count -> scan -> fill.  A task is one lattice column (item, Nx).
// ------------------------------------------------------------------------------------------------------------
__global__ void count_kernel(const Item* __restrict__ items, const Regime* __restrict__ regimes, const int2* __restrict__ tasks, int ntasks,
                             const unsigned* __restrict__ want_pool, const unsigned* __restrict__ known_pool,
                             const unsigned char* __restrict__ have_lim, int* __restrict__ counts) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntasks) return;
    const Item item = items[tasks[t].x];
    const Regime rg = regimes[item.regime];
    const long long w0 = item.mask_off + (long long) tasks[t].y * rg.nwords;
    int c = 0;
    for (int w = 0; w < rg.nwords; w++) c += __popc(want_pool[w0 + w] & ~known_pool[w0 + w]);
    if (c > 0 && !have_lim[item.col_off + tasks[t].y]) c += 1;
    counts[t] = c;
}

// exclusive scan of counts[0..n) into offsets[0..n], offsets[n] = total: 4096-element chunks per CTA (coalesced), a single CTA over
// the chunk totals, then the chunk bases are added
constexpr int SCAN_CHUNK = 4096;
__global__ void __launch_bounds__(1024) scan_chunk_kernel(const int* __restrict__ counts, int* __restrict__ offsets, int n, int* __restrict__ chunk_sums) {
    __shared__ int warp_sums[32];
    const int base = blockIdx.x * SCAN_CHUNK + threadIdx.x * 4;
    int v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = (base + k < n) ? counts[base + k] : 0;
    const int mine = v[0] + v[1] + v[2] + v[3];
    int pre = mine;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int off = 1; off < 32; off <<= 1) { const int u = __shfl_up_sync(0xffffffffu, pre, off); if (lane >= off) pre += u; }
    if (lane == 31) warp_sums[wid] = pre;
    __syncthreads();
    if (wid == 0) {
        int ws = warp_sums[lane];
        for (int off = 1; off < 32; off <<= 1) { const int u = __shfl_up_sync(0xffffffffu, ws, off); if (lane >= off) ws += u; }
        warp_sums[lane] = ws;
    }
    __syncthreads();
    int run = pre - mine + (wid ? warp_sums[wid - 1] : 0);
#pragma unroll
    for (int k = 0; k < 4; k++) { if (base + k < n) offsets[base + k] = run; run += v[k]; }
    if (threadIdx.x == 1023) chunk_sums[blockIdx.x] = warp_sums[31];
}
__global__ void __launch_bounds__(1024) scan_sums_kernel(int* __restrict__ chunk_sums, int nchunks, int* __restrict__ offsets, int n) {
    // nchunks <= 1024 (n < 4 M columns); exclusive scan in place, total to offsets[n]
    __shared__ int sh[1024];
    const int mine = (threadIdx.x < nchunks) ? chunk_sums[threadIdx.x] : 0;
    sh[threadIdx.x] = mine;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
        const int u = (threadIdx.x >= off) ? sh[threadIdx.x - off] : 0;
        __syncthreads();
        sh[threadIdx.x] += u;
        __syncthreads();
    }
    if (threadIdx.x < nchunks) chunk_sums[threadIdx.x] = sh[threadIdx.x] - mine;
    if (threadIdx.x == 1023) offsets[n] = sh[1023];
}
__global__ void __launch_bounds__(1024) scan_add_kernel(int* __restrict__ offsets, int n, const int* __restrict__ chunk_sums) {
    const int base = blockIdx.x * SCAN_CHUNK + threadIdx.x * 4, add = chunk_sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < 4; k++) if (base + k < n) offsets[base + k] += add;
}

// one warp per column task, one lane per 32-node mask word (a column has at most 16 words)
__global__ void __launch_bounds__(256) fill_kernel(const Item* __restrict__ items, const Regime* __restrict__ regimes, const int2* __restrict__ tasks, int ntasks,
                            const unsigned* __restrict__ want_pool, const unsigned* __restrict__ known_pool,
                            const unsigned char* __restrict__ have_lim, const int* __restrict__ offsets, WorkEntry* __restrict__ list) {
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (t >= ntasks) return;
    int o = offsets[t];
    if (offsets[t + 1] == o) return;
    const int it = tasks[t].x, nx = tasks[t].y;
    const Item item = items[it];
    const Regime rg = regimes[item.regime];
    const long long w0 = item.mask_off + (long long) nx * rg.nwords;
    if (!have_lim[item.col_off + nx]) { if (lane == 0) list[o] = WorkEntry{it, (unsigned short) nx, (unsigned short) LIM_ROW}; o++; }
    for (int wb = 0; wb < rg.nwords; wb += 32) {
        const int w = wb + lane;
        unsigned m = (w < rg.nwords) ? (want_pool[w0 + w] & ~known_pool[w0 + w]) : 0u;
        const int cnt = __popc(m);
        int pre = cnt;
        for (int off = 1; off < 32; off <<= 1) { const int v = __shfl_up_sync(0xffffffffu, pre, off); if (lane >= off) pre += v; }
        int q = o + pre - cnt;
        while (m) {
            const int bpos = __ffs(m) - 1;
            m &= m - 1;
            list[q++] = WorkEntry{it, (unsigned short) nx, (unsigned short) (w * 32 + bpos)};
        }
        o += __shfl_sync(0xffffffffu, pre, 31);
    }
}

// ------------------------------------------------------------------------------------------------------------
// theta set-up kernel (power-law potentials): one thread per work entry finds the distance of closest approach and stores
// A_t = eps C_t (sigma/R)^lam_t / (kT g^2) and c2 = (R/b)^2, structure-of-arrays with stride `stride`. Splitting the root
// finding from the sums keeps each kernel's code small (the fused kernel stalled on instruction fetch) and every warp in one phase.
// ------------------------------------------------------------------------------------------------------------
template<int NT>
__global__ void __launch_bounds__(256) theta_setup_kernel(const Item* __restrict__ items, const Regime* __restrict__ regimes, const PairParams* __restrict__ pairs,
                                                          const double* __restrict__ coeff_pool, const WorkEntry* __restrict__ list, int nlist,
                                                          double* __restrict__ prm, long long stride, int* __restrict__ status_pool) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nlist) return;
    if constexpr (NT != 1) {
        const WorkEntry we = list[e];
        const Item item = items[we.item];
        const Regime rg = regimes[item.regime];
        const double g = lattice_coord(we.nx, rg.dg);
        const double b = (we.ny == LIM_ROW) ? 100.0 : lattice_coord(we.ny, rg.db);
        const double* coeff = (item.coeff_off >= 0) ? coeff_pool + item.coeff_off : nullptr;
        double A[NT], c2;
        int st = 0;
        lane_setup<NT>(pairs[item.pair], coeff, item.T, g, b, A, c2, st);
#pragma unroll
        for (int t = 0; t < NT; t++) prm[(long long) t * stride + e] = A[t];
        prm[(long long) NT * stride + e] = c2;
        if (st) atomicOr(&status_pool[we.item], st);
    }
}

// ------------------------------------------------------------------------------------------------------------
// theta kernel: one lane per theta integral. Persistent CTAs; the work list is sorted by pair, so a CTA walks the pair
// segments in order, stages that pair's tanh-sinh rows in shared memory once, and its warps pull 32-entry tiles from
// the segment's counter independently (no CTA barrier between tiles).
// ------------------------------------------------------------------------------------------------------------
template<int NT>
__global__ void __launch_bounds__(THETA_THREADS, 2)
theta_kernel(const Item* __restrict__ items, const Regime* __restrict__ regimes, const PairParams* __restrict__ pairs,
             const double* const* __restrict__ rows_by_pair, const double* __restrict__ coeff_pool,
             const WorkEntry* __restrict__ list, const int* __restrict__ offsets, const int* __restrict__ pair_task_start, int npairs,
             int* __restrict__ tile_counters, double* __restrict__ tab_pool, double* __restrict__ lim_pool, int* __restrict__ status_pool,
             unsigned long long* __restrict__ counters, const double* __restrict__ prm, long long prm_stride) {
    extern __shared__ double srows[];
    __shared__ PairParams spair;
    constexpr int W = 2 + NT;
    const int lane = threadIdx.x & 31;
    unsigned long long my_nodes = 0, my_evals = 0, my_lim_nodes = 0, my_lim_evals = 0, my_exec = 0, my_redo = 0;
    for (int p = 0; p < npairs; p++) {
        const int seg0 = offsets[pair_task_start[p]], seg1 = offsets[pair_task_start[p + 1]];
        if (seg1 == seg0) continue;
        __syncthreads();                                   // previous pair's rows no longer in use
        const double* grows = rows_by_pair[p];
        for (int i = threadIdx.x; i < KS_ROWS * W; i += THETA_THREADS) srows[i] = grows[i];
        if (threadIdx.x == 0) spair = pairs[p];
        __syncthreads();
        const int ntiles = (seg1 - seg0 + 31) / 32;
        for (;;) {
            int tile = 0;
            if (lane == 0) tile = atomicAdd(&tile_counters[p], 1);
            tile = __shfl_sync(0xffffffffu, tile, 0);
            if (tile >= ntiles) break;
            const int e = seg0 + tile * 32 + lane;
            bool redo = false;
            double Ap[NT], cc = 0.0, I = 0.0;
#pragma unroll
            for (int t = 0; t < NT; t++) Ap[t] = 0.0;
            int st = 0; unsigned nodes = 0;
            WorkEntry we = WorkEntry{0, 0, 0};
            if (e < seg1) {
                we = list[e];
                const Item item = items[we.item];
                const Regime rg = regimes[item.regime];
                const double g = lattice_coord(we.nx, rg.dg);
                const double b = (we.ny == LIM_ROW) ? 100.0 : lattice_coord(we.ny, rg.db);   // theta_lim: b = 100 sigma (Spherical.cpp:316)
                const double* coeff = (item.coeff_off >= 0) ? coeff_pool + item.coeff_off : nullptr;
                unsigned ex = 0;
                if constexpr (NT == 1) I = lane_theta_integral_defer<NT>(spair, coeff, srows, grows, KS_ROWS, item.T, g, b, nodes, ex, st, redo, Ap, cc);
                else {
                    double A[NT];
#pragma unroll
                    for (int t = 0; t < NT; t++) A[t] = prm[(long long) t * prm_stride + e];
                    cc = prm[(long long) NT * prm_stride + e];
                    I = lane_sum_defer<NT>(srows, grows, KS_ROWS, A, cc, nodes, ex, st, redo, Ap);
                }
                my_exec += ex;
            }
            // integrals the strided sum rejected: the warp sums them node by node together, one after the other
            if constexpr (NT != 1) {
                unsigned pending = __ballot_sync(0xffffffffu, redo);
                while (pending) {
                    const int src = __ffs(pending) - 1;
                    pending &= pending - 1;
                    double Aw[NT];
#pragma unroll
                    for (int t = 0; t < NT; t++) Aw[t] = __shfl_sync(0xffffffffu, Ap[t], src);
                    const double cw = __shfl_sync(0xffffffffu, cc, src);
                    int kk = 0, stw = 0;
                    const double Iw = theta_integral_warp<NT>(srows, grows, KS_ROWS, Aw, cw, kk, stw);
                    if (lane == src) { I = Iw; nodes = (unsigned) kk; st |= stw; my_exec += (unsigned) (((kk + 31) / 32) * 32); my_redo++; }
                }
            }
            if (e < seg1) {
                const Item item = items[we.item];
                const Regime rg = regimes[item.regime];
                if (we.ny == LIM_ROW) { lim_pool[item.col_off + we.nx] = I; my_lim_nodes += nodes; my_lim_evals++; }
                else { tab_pool[item.tab_off + (long long) we.nx * rg.ny + we.ny] = I; my_nodes += nodes; my_evals++; }
                if (st) atomicOr(&status_pool[we.item], st);
            }
            __syncwarp();
        }
    }
    // counters: [0] chi evals [1] chi nodes [2] lim evals [3] lim nodes [4] node evaluations actually executed [5] redo entries
    for (int off = 16; off > 0; off >>= 1) {
        my_exec += __shfl_down_sync(0xffffffffu, my_exec, off);
        my_redo += __shfl_down_sync(0xffffffffu, my_redo, off);
        my_nodes += __shfl_down_sync(0xffffffffu, my_nodes, off);
        my_evals += __shfl_down_sync(0xffffffffu, my_evals, off);
        my_lim_nodes += __shfl_down_sync(0xffffffffu, my_lim_nodes, off);
        my_lim_evals += __shfl_down_sync(0xffffffffu, my_lim_evals, off);
    }
    if (lane == 0) {
        if (my_evals) { atomicAdd(&counters[0], my_evals); atomicAdd(&counters[1], my_nodes); }
        if (my_lim_evals) { atomicAdd(&counters[2], my_lim_evals); atomicAdd(&counters[3], my_lim_nodes); }
        if (my_exec) atomicAdd(&counters[4], my_exec);
        if (my_redo) atomicAdd(&counters[5], my_redo);
    }
}

// finalize: chi = PI - 2 (I - I_lim + PI/2) (cpp/Spherical.cpp:269,464), store cos(chi), publish the node as known.
__global__ void finalize_kernel(const Item* __restrict__ items, const Regime* __restrict__ regimes, const WorkEntry* __restrict__ list,
                                const int* __restrict__ nlist_ptr, double* __restrict__ tab_pool, const double* __restrict__ lim_pool,
                                unsigned* __restrict__ known_pool, unsigned char* __restrict__ have_lim) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= *nlist_ptr) return;
    const WorkEntry we = list[e];
    const Item item = items[we.item];
    const Regime rg = regimes[item.regime];
    if (we.ny == LIM_ROW) { have_lim[item.col_off + we.nx] = 1; return; }
    const long long o = item.tab_off + (long long) we.nx * rg.ny + we.ny;
    const double I = tab_pool[o];
    const double th = __dadd_rn(__dsub_rn(I, lim_pool[item.col_off + we.nx]), PI_REF / 2);
    const double chi = __dsub_rn(PI_REF, __dmul_rn(2.0, th));
    tab_pool[o] = cos(chi);
    atomicOr(&known_pool[item.mask_off + (long long) we.nx * rg.nwords + (we.ny >> 5)], 1u << (we.ny & 31));
}

// ------------------------------------------------------------------------------------------------------------
// point kernel (kg_chi_points): one lane per (g,b), theta_lim evaluated by the same lane
// ------------------------------------------------------------------------------------------------------------
template<int NT>
__global__ void __launch_bounds__(256) chi_points_kernel(PairParams pp, const double* __restrict__ rows, const double* __restrict__ coeff,
                                                         double T, int n, const double* __restrict__ g, const double* __restrict__ b,
                                                         double* __restrict__ chi_out, double* __restrict__ R_out, int* __restrict__ status) {
    extern __shared__ double srows[];
    constexpr int W = 2 + NT;
    for (int i = threadIdx.x; i < KS_ROWS * W; i += blockDim.x) srows[i] = rows[i];
    __syncthreads();
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    const double gi = g[e], bi = b[e];
    int st = 0; unsigned nodes = 0;
    double R = nan(""), chi;
    if (bi > 100) chi = 0.0;                                  // cpp/Spherical.cpp:462
    else if (bi < 1e-3) chi = PI_REF;                         // theta = 0 (cpp/Spherical.cpp:267)
    else {
        double Rl;
        const double I = lane_theta_integral<NT>(pp, coeff, srows, rows, KS_ROWS, T, gi, bi, R, nodes, st);
        const double Il = lane_theta_integral<NT>(pp, coeff, srows, rows, KS_ROWS, T, gi, 100.0, Rl, nodes, st);
        const double th = __dadd_rn(__dsub_rn(I, Il), PI_REF / 2);
        chi = __dsub_rn(PI_REF, __dmul_rn(2.0, th));
    }
    if (bi <= 100 && bi < 1e-3 && R_out) {
        // get_R is still defined there (used by the transfer-length integrals)
        R = lane_get_R<NT>(pp, coeff, T, gi, bi, st);
    }
    if (chi_out) chi_out[e] = chi;
    if (R_out) R_out[e] = R;
    if (status) status[e] = st;
}

// FP64 FMA-pipe probe: 8 independent DFMA chains per thread
__global__ void __launch_bounds__(256) dfma_probe_kernel(double* out, int iters, double a, double b) {
    double x0 = threadIdx.x * 1e-3, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

} // namespace

// ================================================================================================================
// host side
// ================================================================================================================

namespace {

void regime_for_T(double T, Regime& R) {
    // cpp/Spherical.cpp:60-84
    double end_g = 8, end_b = 5, dg = 0.5, db = 0.03125;
    int rlg = 4, rlb = 16;
    if (T < 70) { dg /= 4.; end_g = 6; end_b = 3.5; }
    if (T < 50) { rlb /= 2; }
    if (T < 35) { dg /= 2; rlg *= 2; db /= 2; end_g = 3.5; end_b = 3.5; }
    if (T < 20) { dg /= 2; rlg *= 2; db /= 2; rlb *= 2; }
    R.dg = dg; R.db = db; R.rlg = rlg; R.rlb = rlb; R.lim = 1e-5;
    // cpp/Integration/Integration.cpp:290-293
    R.Nx_end = (int) ((end_g - 1e-7) / dg + 0.5);
    R.Ny_end = (int) ((end_b - 1e-7) / db + 0.5);
    R.nx = R.Nx_end + 2 * rlg + 1;
    R.ny = R.Ny_end + 2 * rlb + 1;
    R.nwords = (R.ny + 31) / 32;
    R.ldg = (int) std::lround(std::log2(dg)); R.ldb = (int) std::lround(std::log2(db));
    R.nx_cut = 0; while (1e-7 + R.nx_cut * dg <= 1.5) R.nx_cut++;
    R.ny_cut = 0; while (1e-7 + R.ny_cut * db <= 2.5) R.ny_cut++;
}
int regime_index(double T) { return T < 20 ? 4 : T < 35 ? 3 : T < 50 ? 2 : T < 70 ? 1 : 0; }
const double REGIME_T[5] = {300.0, 60.0, 40.0, 30.0, 10.0};

// tanh-sinh abscissae/weights with the reference's libm calls (cpp/Integration/Integration.cpp:421-428)
void build_rows(const kg_potential_desc& d, int ntpad, std::vector<double>& rows) {
    const double PI = 3.14159265359, h = 1.0e-3;
    const int W = 2 + ntpad;
    rows.assign((size_t) KMAX * W, 0.0);
    for (int k = 1; k <= KMAX; k++) {
        const double u = tanh(PI * sinh(k * h) / 2.);
        double w;
        if (k == 1) w = (PI / 2.) * h * cosh(k * h) / pow(cosh(PI * sinh(k * h * 1e-3) / 2.), 2);   // the 1e-3 quirk, :422
        else w = (PI / 2) * h * cosh(k * h) / pow(cosh(PI * sinh(k * h) / 2.), 2);
        double* r = &rows[(size_t) (k - 1) * W];
        r[0] = 0.5 * (w / u);       // halved: the fast node form carries the factor 2 in its Newton step (kg_device.cuh)
        r[1] = 1.0 / (u * u);
        if (ntpad == 1) r[2] = 1.0 / u;      // generic potentials evaluate u(r) at r = R / u
        else for (int t = 0; t < ntpad; t++) r[2 + t] = (t < d.nterms) ? pow(u, d.lam[t]) : 0.0;
    }
}

template<int NT> cudaError_t set_theta_smem() {
    const int bytes = KS_ROWS * (2 + NT) * (int) sizeof(double);
    cudaError_t e = cudaFuncSetAttribute(theta_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(chi_points_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

} // namespace

extern "C" const char* kg_last_error(const kg_model* m) { return m ? m->err.c_str() : g_create_error.c_str(); }
extern "C" int kg_model_npairs(const kg_model* m) { return m ? m->npairs : -1; }

namespace {
// device resources shared by both constructors: stream, node-row tables, pair blocks
int finish_model(kg_model* m, int device, kg_model** out) {
    auto cu = [&](cudaError_t e, const char* what) { if (e != cudaSuccess) { g_create_error = std::string(what) + ": " + cudaGetErrorString(e); return false; } return true; };
    if (!cu(cudaSetDevice(device), "cudaSetDevice")) { delete m; return -2; }
    cudaDeviceProp prop;
    if (!cu(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties")) { delete m; return -2; }
    m->sm_count = prop.multiProcessorCount;
    if (!cu(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking), "cudaStreamCreate")) { delete m; return -2; }
    if (const char* ev = getenv("KG_TS_EXACT")) { const int v = atoi(ev) ? 1 : 0; cudaMemcpyToSymbol(c_ts_exact, &v, sizeof(int)); }
    bool ok = cu(set_theta_smem<1>(), "smem attr") && cu(set_theta_smem<2>(), "smem attr") && cu(set_theta_smem<4>(), "smem attr") && cu(set_theta_smem<6>(), "smem attr");
    if (ok && m->generic) {
        std::vector<GenPot> base(m->npairs);
        for (int p = 0; ok && p < m->npairs; p++)
            if (kg_genpot_build(m->gdescs[p], 300.0, base[p], g_create_error)) ok = false;     // T only matters when fh_order > 0 (then per-T pools are used)
        ok = ok && cu(cudaMalloc(&m->d_gpots, m->npairs * sizeof(GenPot)), "cudaMalloc") &&
             cu(cudaMemcpy(m->d_gpots, base.data(), m->npairs * sizeof(GenPot), cudaMemcpyHostToDevice), "H2D");
    }
    std::vector<double> rows;
    for (int p = 0; ok && p < m->npairs; p++) {
        const auto& d = m->descs[p];
        PairParams pp; memset(&pp, 0, sizeof(pp));
        pp.sigma = d.sigma; pp.eps = d.eps; pp.nterms = d.nterms; pp.ntpad = m->ntpad;
        for (int t = 0; t < d.nterms; t++) { pp.C[t] = d.C[t]; pp.lam[t] = d.lam[t]; }
        if (m->generic) pp.gen = m->d_gpots + p;
        m->pairs.push_back(pp);
        build_rows(d, m->ntpad, rows);
        double* dptr = nullptr;
        ok = ok && cu(cudaMalloc(&dptr, rows.size() * sizeof(double)), "cudaMalloc rows");
        if (ok) { m->d_rows.push_back(dptr); ok = cu(cudaMemcpy(dptr, rows.data(), rows.size() * sizeof(double), cudaMemcpyHostToDevice), "H2D rows"); }
    }
    ok = ok && cu(cudaMalloc(&m->d_rows_by_pair, m->npairs * sizeof(double*)), "cudaMalloc") &&
         cu(cudaMemcpy(m->d_rows_by_pair, m->d_rows.data(), m->npairs * sizeof(double*), cudaMemcpyHostToDevice), "H2D") &&
         cu(cudaMalloc(&m->d_pairs, m->npairs * sizeof(PairParams)), "cudaMalloc") &&
         cu(cudaMemcpy(m->d_pairs, m->pairs.data(), m->npairs * sizeof(PairParams), cudaMemcpyHostToDevice), "H2D");
    ok = ok && cu(cudaMallocHost(&m->h_pinned, 16 * sizeof(int)), "cudaMallocHost");
    if (!ok) { kg_model_destroy(m); return -2; }
    *out = m;
    return 0;
}
int check_device(int device) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { g_create_error = "kg_model_create: no CUDA device (the GPU path has no CPU fallback)"; return -1; }
    if (device < 0 || device >= ndev) { g_create_error = "kg_model_create: bad device ordinal"; return -1; }
    return 0;
}
} // namespace

extern "C" int kg_model_create(const kg_potential_desc* pairs, int ncomps, const double* mass, int is_singlecomp, int device, kg_model** out) {
    KG_TRACE();
    auto fail = [&](const std::string& s) { g_create_error = s; return -1; };
    if (!pairs || !mass || !out || ncomps < 1) return fail("kg_model_create: invalid argument");
    if (check_device(device)) return -1;
    kg_model* m = new kg_model;
    m->device = device; m->ncomps = ncomps; m->npairs = ncomps * (ncomps + 1) / 2; m->is_singlecomp = is_singlecomp;
    m->mass.assign(mass, mass + ncomps);
    m->descs.assign(pairs, pairs + m->npairs);
    int ntmax = 0;
    for (auto& d : m->descs) {
        if (d.kind != KG_POT_POWER_SUM || d.nterms < 1 || d.nterms > KG_MAX_TERMS) { delete m; return fail("kg_model_create: unsupported potential block"); }
        ntmax = std::max(ntmax, d.nterms);
    }
    m->ntpad = ntmax <= 2 ? 2 : ntmax <= 4 ? 4 : 6;
    return finish_model(m, device, out);
}

extern "C" int kg_model_create_ex(const kg_potential_desc_ex* pairs, int ncomps, const double* mass, int is_singlecomp, int device, kg_model** out) {
    KG_TRACE();
    auto fail = [&](const std::string& s) { g_create_error = s; return -1; };
    if (!pairs || !mass || !out || ncomps < 1) return fail("kg_model_create_ex: invalid argument");
    if (check_device(device)) return -1;
    kg_model* m = new kg_model;
    m->device = device; m->ncomps = ncomps; m->npairs = ncomps * (ncomps + 1) / 2; m->is_singlecomp = is_singlecomp;
    m->mass.assign(mass, mass + ncomps);
    m->generic = true; m->ntpad = 1; m->coeff_stride = GP_DOUBLES;
    m->gdescs.assign(pairs, pairs + m->npairs);
    for (auto& g : m->gdescs) {
        if (g.kind < KG_POT_LJ_SPLINE || g.kind > KG_POT_PATKOWSKI || !(g.sigma > 0) || !(g.eps > 0) || g.nparams < 0 || g.nparams > KG_MAX_POT_PARAMS) { delete m; return fail("kg_model_create_ex: unsupported potential block"); }
        if (g.fh_order > 0) { m->gen_T_dependent = true; if (!(g.fh_D > 0)) { delete m; return fail("kg_model_create_ex: fh_order > 0 needs fh_D > 0"); } }
        GenPot probe;
        if (kg_genpot_build(g, 300.0, probe, g_create_error)) { delete m; return -1; }
        kg_potential_desc d; memset(&d, 0, sizeof(d));       // sigma / eps for the Omega prefactors and the TL scaling
        d.kind = g.kind; d.sigma = g.sigma; d.eps = g.eps;
        m->descs.push_back(d);
    }
    return finish_model(m, device, out);
}

namespace { std::map<void*, int> g_child_kind; std::mutex g_child_mu; }     // 0 = plan, 1 = tcache
void kg_register_child(void* child, kg_model* m, int kind) {
    { std::lock_guard<std::mutex> g(g_child_mu); g_child_kind[child] = kind; }
    KgRegistry::get().add(child, m);
}
extern "C" void kg_tcache_destroy(kg_tcache* c);
void kg_destroy_child(void* child) {
    int kind = -1;
    { std::lock_guard<std::mutex> g(g_child_mu); auto f = g_child_kind.find(child); if (f != g_child_kind.end()) { kind = f->second; g_child_kind.erase(f); } }
    if (kind == 0) kg_omega_plan_destroy((kg_omega_plan*) child);
    else if (kind == 1) kg_tcache_destroy((kg_tcache*) child);
}

extern "C" void kg_model_destroy(kg_model* m) {
    KG_TRACE();
    if (!m) return;
    cudaSetDevice(m->device);
    for (void* child : KgRegistry::get().children_of(m)) kg_destroy_child(child);
    for (double* p : m->d_rows) cudaFree(p);
    if (m->d_rows_by_pair) cudaFree(m->d_rows_by_pair);
    if (m->d_pairs) cudaFree(m->d_pairs);
    if (m->d_gpots) cudaFree(m->d_gpots);
    if (m->stream) cudaStreamSynchronize(m->stream);
    m->pool_release();
    for (auto& kv : m->pool_live) cudaFree(kv.first);
    if (m->h_pinned) cudaFreeHost(m->h_pinned);
    if (m->stream && m->own_stream) cudaStreamDestroy(m->stream);
    delete m;
}

extern "C" int kg_model_set_stream(kg_model* m, void* stream) {
    KG_TRACE();
    if (!m) return -1;
    cudaSetDevice(m->device);
    if (m->stream && m->own_stream) { cudaStreamSynchronize(m->stream); cudaStreamDestroy(m->stream); }
    m->stream = (cudaStream_t) stream;
    m->own_stream = false;
    return 0;
}

extern "C" int kg_chi_points(kg_model* m, int pair, double T, int n, const double* g, const double* b_red,
                             const double* coeff, double* chi_out, double* R_out, int32_t* status) {
    KG_TRACE();
    if (!m) return -1;
    auto set_error = [&](const std::string& s) { m->err = s; };
    if (pair < 0 || pair >= m->npairs || n < 0 || !g || !b_red) { set_error("kg_chi_points: invalid argument"); return -1; }
    if (n == 0) return 0;
    CUDA_OK(cudaSetDevice(m->device));
    double *dg = nullptr, *db = nullptr, *dchi = nullptr, *dR = nullptr, *dcoef = nullptr; int* dst = nullptr;
    CUDA_OK(kg_dmalloc(m, &dg, n * sizeof(double))); CUDA_OK(kg_dmalloc(m, &db, n * sizeof(double)));
    CUDA_OK(kg_dmalloc(m, &dchi, n * sizeof(double))); CUDA_OK(kg_dmalloc(m, &dR, n * sizeof(double)));
    CUDA_OK(kg_dmalloc(m, &dst, n * sizeof(int)));
    CUDA_OK(cudaMemcpyAsync(dg, g, n * sizeof(double), cudaMemcpyHostToDevice, m->stream));
    CUDA_OK(cudaMemcpyAsync(db, b_red, n * sizeof(double), cudaMemcpyHostToDevice, m->stream));
    GenPot gp_T;
    if (m->generic) {
        if (m->gen_T_dependent) {          // Feynman-Hibbs corrected: the effective potential of this temperature
            if (kg_genpot_build(m->gdescs[pair], T, gp_T, m->err)) return -1;
            CUDA_OK(kg_dmalloc(m, &dcoef, sizeof(GenPot)));
            CUDA_OK(cudaMemcpyAsync(dcoef, &gp_T, sizeof(GenPot), cudaMemcpyHostToDevice, m->stream));
        }
    } else if (coeff) {
        CUDA_OK(kg_dmalloc(m, &dcoef, KG_MAX_TERMS * sizeof(double)));
        CUDA_OK(cudaMemcpyAsync(dcoef, coeff, m->descs[pair].nterms * sizeof(double), cudaMemcpyHostToDevice, m->stream));
    }
    const int W = 2 + m->ntpad;
    const size_t smem = (size_t) KS_ROWS * W * sizeof(double);
    const int grid = (n + 255) / 256;
    switch (m->ntpad) {
        case 1: chi_points_kernel<1><<<grid, 256, smem, m->stream>>>(m->pairs[pair], m->d_rows[pair], dcoef, T, n, dg, db, dchi, dR, dst); break;
        case 2: chi_points_kernel<2><<<grid, 256, smem, m->stream>>>(m->pairs[pair], m->d_rows[pair], dcoef, T, n, dg, db, dchi, dR, dst); break;
        case 4: chi_points_kernel<4><<<grid, 256, smem, m->stream>>>(m->pairs[pair], m->d_rows[pair], dcoef, T, n, dg, db, dchi, dR, dst); break;
        default: chi_points_kernel<6><<<grid, 256, smem, m->stream>>>(m->pairs[pair], m->d_rows[pair], dcoef, T, n, dg, db, dchi, dR, dst); break;
    }
    CUDA_OK(cudaGetLastError());
    if (chi_out) CUDA_OK(cudaMemcpyAsync(chi_out, dchi, n * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    if (R_out) CUDA_OK(cudaMemcpyAsync(R_out, dR, n * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    if (status) CUDA_OK(cudaMemcpyAsync(status, dst, n * sizeof(int), cudaMemcpyDeviceToHost, m->stream));
    CUDA_OK(cudaStreamSynchronize(m->stream));
    m->dfree(dg); m->dfree(db); m->dfree(dchi); m->dfree(dR); m->dfree(dst); m->dfree(dcoef);
    return 0;
}

int kg_effective_coeff(kg_model* m, int nT, const double* T, const double* coeff_T_user, std::vector<double>& buf, const double** out) {
    if (!m->generic) { *out = coeff_T_user; return 0; }
    *out = nullptr;
    if (!m->gen_T_dependent) return 0;
    if ((long long) nT * m->npairs * GP_DOUBLES >= (1LL << 31)) { m->err = "generic potential: batch too large, split the T grid"; return -1; }
    buf.resize((size_t) nT * m->npairs * GP_DOUBLES);
    for (int t = 0; t < nT; t++) for (int p = 0; p < m->npairs; p++) {
        GenPot gp;
        if (kg_genpot_build(m->gdescs[p], T[t], gp, m->err)) return -1;
        memcpy(&buf[((size_t) t * m->npairs + p) * GP_DOUBLES], &gp, sizeof(GenPot));
    }
    *out = buf.data();
    return 0;
}

namespace {
template<int NT>
__global__ void potential_points_kernel(PairParams pp, const double* __restrict__ coeff, int n, const double* __restrict__ r,
                                        double* __restrict__ u_out, double* __restrict__ du_out) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    Lane<NT> c;
    make_lane<NT>(c, pp, coeff, 1.0 / BOLTZMANN, 1.0, 1.0);      // kT g^2 = 1: pot_eval returns joules
    double U, dU;
    pot_eval<NT>(c, r[e] / pp.sigma, U, dU);
    if (u_out) u_out[e] = U;
    if (du_out) du_out[e] = dU / pp.sigma;
}
}

extern "C" int kg_potential_points(kg_model* m, int pair, double T, int n, const double* r, double* u_out, double* du_out) {
    KG_TRACE();
    if (!m) return -1;
    auto set_error = [&](const std::string& s) { m->err = s; };
    if (pair < 0 || pair >= m->npairs || n < 0 || !r) { set_error("kg_potential_points: invalid argument"); return -1; }
    if (n == 0) return 0;
    CUDA_OK(cudaSetDevice(m->device));
    double *dr = nullptr, *du = nullptr, *ddu = nullptr, *dcoef = nullptr;
    CUDA_OK(kg_dmalloc(m, &dr, n * sizeof(double))); CUDA_OK(kg_dmalloc(m, &du, n * sizeof(double))); CUDA_OK(kg_dmalloc(m, &ddu, n * sizeof(double)));
    CUDA_OK(cudaMemcpyAsync(dr, r, n * sizeof(double), cudaMemcpyHostToDevice, m->stream));
    GenPot gp_T;
    if (m->generic && m->gen_T_dependent) {
        if (kg_genpot_build(m->gdescs[pair], T, gp_T, m->err)) return -1;
        CUDA_OK(kg_dmalloc(m, &dcoef, sizeof(GenPot)));
        CUDA_OK(cudaMemcpyAsync(dcoef, &gp_T, sizeof(GenPot), cudaMemcpyHostToDevice, m->stream));
    }
    const int grid = (n + 127) / 128;
    switch (m->ntpad) {
        case 1: potential_points_kernel<1><<<grid, 128, 0, m->stream>>>(m->pairs[pair], dcoef, n, dr, du, ddu); break;
        case 2: potential_points_kernel<2><<<grid, 128, 0, m->stream>>>(m->pairs[pair], dcoef, n, dr, du, ddu); break;
        case 4: potential_points_kernel<4><<<grid, 128, 0, m->stream>>>(m->pairs[pair], dcoef, n, dr, du, ddu); break;
        default: potential_points_kernel<6><<<grid, 128, 0, m->stream>>>(m->pairs[pair], dcoef, n, dr, du, ddu); break;
    }
    CUDA_OK(cudaGetLastError());
    if (u_out) CUDA_OK(cudaMemcpyAsync(u_out, du, n * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    if (du_out) CUDA_OK(cudaMemcpyAsync(du_out, ddu, n * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    CUDA_OK(cudaStreamSynchronize(m->stream));
    m->dfree(dr); m->dfree(du); m->dfree(ddu); m->dfree(dcoef);
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------- plan
struct kg_omega_plan {
    kg_model* m = nullptr;
    int device = 0;
    int nT = 0, nlr = 0, nitems = 0, ntasks = 0, nr_distinct = 0, gfac_stride = 0;
    long long max_list = 0;
    std::vector<Item> items;            // sorted by pair
    std::vector<int> item_out_index;    // item -> iT * npairs + pair
    Item* d_items = nullptr; Regime* d_regimes = nullptr;
    double* d_tab = nullptr; unsigned *d_want = nullptr, *d_known = nullptr;
    double* d_lim = nullptr; unsigned char* d_have_lim = nullptr;
    int2* d_tasks = nullptr; int *d_counts = nullptr, *d_offsets = nullptr;
    WorkEntry* d_list = nullptr;
    double* d_prm = nullptr;          // [ntpad + 1][max_list] A_t and c2 of every work entry (theta_setup_kernel -> theta_kernel)
    double* d_gfac = nullptr; int *d_lr_l = nullptr, *d_lr_ri = nullptr;
    double* d_coeff = nullptr;
    double *d_W = nullptr, *d_omega = nullptr;          // [item][lr] in plan order
    Cell* d_cells[2] = {nullptr, nullptr}; int* d_ncells = nullptr; long long cell_cap = 0, ncells0 = 0, cell_cap_limit = 0;
    double *d_W_out = nullptr, *d_omega_out = nullptr;  // [nT][npairs][nlr]
    int *d_status = nullptr, *d_flags = nullptr, *d_tile_counter = nullptr;
    int* d_perm = nullptr;                              // item -> output slot
    int* d_chunk_sums = nullptr;                        // scan scratch
    int* d_active = nullptr; unsigned long long* d_active_mask = nullptr; long long* d_active_cell_off = nullptr; int nactive = 0;   // items with work, their (l,r) sets
    int max_cells_seen = 0, regrows = 0;
    int* d_pair_task_start = nullptr;                   // [npairs+1] first column task of each pair
    unsigned long long* d_counters = nullptr;
    long long tab_doubles = 0, mask_words = 0, ncols_total = 0;
    double cospi = -1.0;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    double ms_theta = 0, ms_mesh = 0, ms_other = 0; int launches = 0, passes = 0;
    long long h2d_bytes = 0, d2h_bytes = 0;
    int* h_pinned = nullptr;   // [0] flag, [1] nlist
};

namespace {
__global__ void permute_kernel(const double* __restrict__ W, const Item* __restrict__ items, const int* __restrict__ perm, int nitems, int nlr,
                               double* __restrict__ W_out, double* __restrict__ om_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nitems * nlr) return;
    const int it = i / nlr, lr = i % nlr;
    const long long o = (long long) perm[it] * nlr + lr;
    W_out[o] = W[i];
    om_out[o] = items[it].pref * W[i];     // cpp/Spherical.cpp:46-47
}
}

static void plan_free(kg_omega_plan* p);
extern "C" int kg_omega_plan_create(kg_model* m, int nT, const double* T, int nlr, const int32_t* l, const int32_t* r,
                                    const double* coeff_T, kg_omega_plan** out) {
    return kg_omega_plan_create_masked(m, nT, T, nlr, l, r, coeff_T, nullptr, out);
}

extern "C" int kg_omega_plan_create_masked(kg_model* m, int nT, const double* T, int nlr, const int32_t* l, const int32_t* r,
                                           const double* coeff_T, const uint8_t* need, kg_omega_plan** out) {
    KG_TRACE();
    if (!m) return -1;
    auto set_error = [&](const std::string& s) { m->err = s; };
    if (nT < 1 || !T || nlr < 1 || !l || !r || !out) { set_error("kg_omega_plan_create: invalid argument"); return -1; }
    std::vector<double> coeff_buf;
    if (kg_effective_coeff(m, nT, T, coeff_T, coeff_buf, &coeff_T)) return -1;
    for (int i = 0; i < nlr; i++) if (l[i] < 1 || l[i] > 12 || r[i] < 0 || r[i] > 24) { set_error("kg_omega_plan_create: (l,r) out of range"); return -1; }
    for (int i = 0; i < nT; i++) if (!(T[i] > 0) || !std::isfinite(T[i])) { set_error("kg_omega_plan_create: temperatures must be positive and finite"); return -1; }
    CUDA_OK(cudaSetDevice(m->device));
    kg_omega_plan* p = new kg_omega_plan;
    p->m = m; p->device = m->device; p->nT = nT; p->nlr = nlr;
    const int Np = m->npairs;
    p->nitems = nT * Np;
    Regime regs[5];
    for (int k = 0; k < 5; k++) regime_for_T(REGIME_T[k], regs[k]);
    // items sorted by pair (theta tiles keep one pair's node rows in shared memory), then by T index
    p->items.reserve(p->nitems);
    std::vector<int> perm; perm.reserve(p->nitems);
    std::vector<int> active; std::vector<unsigned long long> active_mask; std::vector<long long> active_cell_off;   // items with work, their (l,r) sets, level-0 cell offsets
    std::vector<char> item_active;
    const double PI = 3.14159265359, KB = 1.38064852e-23;
    int pi = 0;
    for (int i = 0; i < m->ncomps; i++) for (int j = i; j < m->ncomps; j++, pi++) {
        for (int t = 0; t < nT; t++) {
            Item it; memset(&it, 0, sizeof(it));
            it.pair = pi; it.T = T[t]; it.regime = regime_index(T[t]);
            const double sig = m->descs[pi].sigma;
            if (i == j) it.pref = pow(sig, 2) * sqrt((PI * KB * T[t]) / m->mass[i]);            // cpp/Spherical.cpp:46
            else {
                const double m0 = m->mass[i] + m->mass[j], Mij = m->mass[i] / m0, Mji = m->mass[j] / m0;
                it.pref = 0.5 * pow(sig, 2) * sqrt(2 * PI * KB * T[t] / (m0 * Mij * Mji));       // cpp/Spherical.cpp:47
            }
            const Regime& rg = regs[it.regime];
            const int item_index = (int) p->items.size();
            int nact = 0;
            unsigned long long members = 0;
            for (int k = 0; k < nlr; k++) {
                if (need && !need[((size_t) t * Np + pi) * nlr + k]) continue;
                members |= 1ull << k;
                nact++;
            }
            item_active.push_back(nact > 0);
            if (nact > 0) {      // items with nothing to integrate own no lattice, no columns, no cells
                active.push_back(item_index); active_mask.push_back(members); active_cell_off.push_back(p->ncells0);
                p->ncells0 += (long long) (rg.Nx_end / rg.rlg) * (rg.Ny_end / rg.rlb);
                it.tab_off = p->tab_doubles; p->tab_doubles += (long long) rg.nx * rg.ny;
                it.mask_off = p->mask_words; p->mask_words += (long long) rg.nx * rg.nwords;
                it.col_off = p->ncols_total; p->ncols_total += rg.nx;
                // cells of one level: the members of an item mostly share their cells, so one tiling of the domain at the
                // finest step is the working estimate; the hard bound is one tiling per member (kg_omega_plan_run grows the
                // lists on overflow)
                p->cell_cap += (long long) rg.Nx_end * rg.Ny_end * 5 / 4 + 64;
                p->cell_cap_limit += (long long) nact * rg.Nx_end * rg.Ny_end + 64;
            }
            it.coeff_off = coeff_T ? (t * Np + pi) * m->coeff_stride : -1;
            p->items.push_back(it);
            perm.push_back(t * Np + pi);
        }
    }
    // tasks (one per lattice column of an active item, in item order = pair-major) are written on the device by
    // build_tasks_kernel; the first task of pair q is the col_off of its first active item
    p->nactive = (int) active.size();
    p->ntasks = (int) p->ncols_total;
    std::vector<int> pair_task_start(Np + 1, p->ntasks);
    for (int a = (int) active.size() - 1; a >= 0; a--) { const Item& it = p->items[active[a]]; pair_task_start[it.pair] = (int) it.col_off; }
    for (int q = Np - 1; q >= 0; q--) if (pair_task_start[q] > pair_task_start[q + 1]) pair_task_start[q] = pair_task_start[q + 1];
    // g-factor tables 2 exp(-g^2) g^(2r+3) per (regime, distinct r): same libm calls, same order as cpp/Spherical.cpp:101
    std::vector<int> rvals, lr_ri(nlr);
    for (int i = 0; i < nlr; i++) {
        auto f = std::find(rvals.begin(), rvals.end(), r[i]);
        if (f == rvals.end()) { lr_ri[i] = (int) rvals.size(); rvals.push_back(r[i]); } else lr_ri[i] = (int) (f - rvals.begin());
    }
    p->nr_distinct = (int) rvals.size();
    p->gfac_stride = 160;
    std::vector<double> gfac((size_t) 5 * p->nr_distinct * p->gfac_stride, 0.0);
    for (int k = 0; k < 5; k++) for (int ri = 0; ri < p->nr_distinct; ri++) for (int nx = 0; nx < regs[k].nx; nx++) {
        const double g = 1e-7 + nx * regs[k].dg;
        gfac[((size_t) k * p->nr_distinct + ri) * p->gfac_stride + nx] = 2 * exp(-pow(g, 2)) * pow(g, 2.0 * rvals[ri] + 3.0);
    }
    p->cospi = cos(PI - 2.0 * 0.0);
    // The worst case list (every node of every lattice) is never reached: nodes in the cut-off region (b>2.5, g>1.5) are
    // never requested. Cap by the count of requestable nodes.
    long long cap = 0, cap_of_regime[5];
    for (int k = 0; k < 5; k++) {
        const Regime& rg = regs[k];
        cap_of_regime[k] = 0;
        for (int nx = 0; nx < rg.nx; nx++) {
            const double g = 1e-7 + nx * rg.dg;
            int nyreq = rg.ny;
            if (g > 1.5) nyreq = std::min(rg.ny, (int) (2.5 / rg.db) + 2);
            cap_of_regime[k] += nyreq + 1;
        }
    }
    for (int it = 0; it < p->nitems; it++) if (item_active[it]) cap += cap_of_regime[p->items[it].regime];
    p->max_list = cap;

#define PLAN_CUDA(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { set_error(std::string(#expr) + ": " + cudaGetErrorString(e_)); plan_free(p); return -2; } } while (0)
    PLAN_CUDA(kg_dmalloc(m, &p->d_items, p->nitems * sizeof(Item)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_regimes, 5 * sizeof(Regime)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_tab, p->tab_doubles * sizeof(double)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_want, p->mask_words * sizeof(unsigned)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_known, p->mask_words * sizeof(unsigned)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_lim, p->ncols_total * sizeof(double)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_have_lim, p->ncols_total));
    PLAN_CUDA(kg_dmalloc(m, &p->d_tasks, p->ntasks * sizeof(int2)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_counts, (p->ntasks + 1) * sizeof(int)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_offsets, (p->ntasks + 1) * sizeof(int)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_list, p->max_list * sizeof(WorkEntry)));
    if (m->ntpad != 1) PLAN_CUDA(kg_dmalloc(m, &p->d_prm, (size_t) p->max_list * (m->ntpad + 1) * sizeof(double)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_gfac, gfac.size() * sizeof(double)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_lr_l, nlr * sizeof(int)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_lr_ri, nlr * sizeof(int)));
    // small batches take the hard bound outright (never grows); large ones start from the estimate, at least 32 M cells
    p->cell_cap = std::min(p->cell_cap_limit, std::max(p->cell_cap, (long long) 32 << 20));
    p->cell_cap = std::max(p->cell_cap, p->ncells0);
    if (p->nitems >= (1 << 22) || nlr > 64 || p->cell_cap >= (1LL << 31)) { set_error("kg_omega_plan_create: batch too large, split the T grid"); plan_free(p); return -1; }
    PLAN_CUDA(kg_dmalloc(m, &p->d_cells[0], p->cell_cap * sizeof(Cell)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_cells[1], p->cell_cap * sizeof(Cell)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_ncells, 2 * sizeof(int)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_W, (size_t) p->nitems * nlr * sizeof(double)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_omega, (size_t) p->nitems * nlr * sizeof(double)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_W_out, (size_t) p->nitems * nlr * sizeof(double)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_omega_out, (size_t) p->nitems * nlr * sizeof(double)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_status, p->nitems * sizeof(int)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_flags, 4 * sizeof(int)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_tile_counter, (Np + 1) * sizeof(int)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_pair_task_start, (Np + 1) * sizeof(int)));
    PLAN_CUDA(cudaMemcpy(p->d_pair_task_start, pair_task_start.data(), (Np + 1) * sizeof(int), cudaMemcpyHostToDevice));
    PLAN_CUDA(kg_dmalloc(m, &p->d_perm, p->nitems * sizeof(int)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_active, (active.size() + 1) * sizeof(int)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_active_mask, (active.size() + 1) * sizeof(unsigned long long)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_active_cell_off, (active.size() + 1) * sizeof(long long)));
    PLAN_CUDA(cudaMemcpy(p->d_active, active.data(), active.size() * sizeof(int), cudaMemcpyHostToDevice));
    PLAN_CUDA(cudaMemcpy(p->d_active_mask, active_mask.data(), active.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice));
    PLAN_CUDA(cudaMemcpy(p->d_active_cell_off, active_cell_off.data(), active.size() * sizeof(long long), cudaMemcpyHostToDevice));
    PLAN_CUDA(kg_dmalloc(m, &p->d_counters, 8 * sizeof(unsigned long long)));
    PLAN_CUDA(kg_dmalloc(m, &p->d_chunk_sums, 1024 * sizeof(int)));
    if (p->ntasks > 1024 * SCAN_CHUNK) { set_error("kg_omega_plan_create: batch too large, split the T grid"); plan_free(p); return -1; }
    p->h_pinned = m->h_pinned;
    if (coeff_T) {
        PLAN_CUDA(kg_dmalloc(m, &p->d_coeff, (size_t) nT * Np * m->coeff_stride * sizeof(double)));
        PLAN_CUDA(cudaMemcpy(p->d_coeff, coeff_T, (size_t) nT * Np * m->coeff_stride * sizeof(double), cudaMemcpyHostToDevice));
    }
    PLAN_CUDA(cudaMemcpy(p->d_items, p->items.data(), p->nitems * sizeof(Item), cudaMemcpyHostToDevice));
    PLAN_CUDA(cudaMemcpy(p->d_regimes, regs, sizeof(regs), cudaMemcpyHostToDevice));
    if (p->nactive > 0) {
        build_tasks_kernel<<<p->nactive, 64, 0, m->stream>>>(p->d_items, p->d_regimes, p->d_active, p->nactive, p->d_tasks);
        PLAN_CUDA(cudaGetLastError());
    }
    PLAN_CUDA(cudaMemcpy(p->d_gfac, gfac.data(), gfac.size() * sizeof(double), cudaMemcpyHostToDevice));
    PLAN_CUDA(cudaMemcpy(p->d_lr_l, l, nlr * sizeof(int), cudaMemcpyHostToDevice));
    PLAN_CUDA(cudaMemcpy(p->d_lr_ri, lr_ri.data(), nlr * sizeof(int), cudaMemcpyHostToDevice));
    PLAN_CUDA(cudaMemcpy(p->d_perm, perm.data(), p->nitems * sizeof(int), cudaMemcpyHostToDevice));
    for (int k = 0; k < 4; k++) PLAN_CUDA(cudaEventCreate(&p->ev[k]));
    p->h2d_bytes = (long long) p->nitems * sizeof(Item) + sizeof(regs) + gfac.size() * sizeof(double)
                 + 2LL * nlr * sizeof(int) + (long long) p->nitems * sizeof(int) + (coeff_T ? (long long) nT * Np * m->coeff_stride * sizeof(double) : 0);
#undef PLAN_CUDA
    kg_register_child(p, m, 0);
    *out = p;
    return 0;
}

static void plan_free(kg_omega_plan* p);
extern "C" void kg_omega_plan_destroy(kg_omega_plan* p) {
    KG_TRACE();
    if (!p || !KgRegistry::get().remove(p)) return;       // unknown handle: already destroyed together with its model
    plan_free(p);
}
static void plan_free(kg_omega_plan* p) {
    cudaSetDevice(p->device);
    void* ptrs[] = {p->d_items, p->d_regimes, p->d_tab, p->d_want, p->d_known, p->d_lim, p->d_have_lim, p->d_tasks, p->d_counts,
                    p->d_offsets, p->d_list, p->d_prm, p->d_gfac, p->d_lr_l, p->d_lr_ri, p->d_coeff, p->d_W, p->d_omega, p->d_W_out,
                    p->d_omega_out, p->d_status, p->d_flags, p->d_tile_counter, p->d_perm, p->d_counters, p->d_cells[0], p->d_cells[1], p->d_ncells, p->d_pair_task_start,
                    p->d_active, p->d_active_mask, p->d_active_cell_off, p->d_chunk_sums};
    for (void* q : ptrs) p->m->dfree(q);
    for (int k = 0; k < 4; k++) if (p->ev[k]) cudaEventDestroy(p->ev[k]);
    delete p;
}

extern "C" int kg_omega_plan_run(kg_omega_plan* p) {
    KG_TRACE();
    if (!p) return -1;
    kg_model* m = p->m;
    auto set_error = [&](const std::string& s) { m->err = s; };
    CUDA_OK(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
#define LAUNCHED(name) do { cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess) { set_error(std::string("kg_omega_plan_run: launch of " name ": ") + cudaGetErrorString(e_)); return -2; } } while (0)
    LAUNCHED("(error left by an earlier call)");
    CUDA_OK(cudaMemsetAsync(p->d_want, 0, p->mask_words * sizeof(unsigned), s));
    CUDA_OK(cudaMemsetAsync(p->d_known, 0, p->mask_words * sizeof(unsigned), s));
    CUDA_OK(cudaMemsetAsync(p->d_have_lim, 0, p->ncols_total, s));
    CUDA_OK(cudaMemsetAsync(p->d_status, 0, p->nitems * sizeof(int), s));
    CUDA_OK(cudaMemsetAsync(p->d_counters, 0, 8 * sizeof(unsigned long long), s));
    p->ms_theta = p->ms_mesh = p->ms_other = 0; p->launches = 0; p->passes = 0;
    CUDA_OK(cudaMemsetAsync(p->d_W, 0, (size_t) p->nitems * p->nlr * sizeof(double), s));
    CUDA_OK(cudaMemsetAsync(p->d_flags, 0, 4 * sizeof(int), s));
    const int W = 2 + m->ntpad;
    const size_t smem = (size_t) KS_ROWS * W * sizeof(double);
    const int tgrid = std::max(1, (p->ntasks + 255) / 256);
    CUDA_OK(cudaEventRecord(p->ev[0], s));
    if (p->nactive > 0)
        mesh_init_kernel<<<p->nactive, 128, 0, s>>>(p->d_items, p->d_regimes, p->d_want, p->d_active, p->d_active_mask, p->d_active_cell_off, p->d_cells[0]);
    LAUNCHED("mesh_init_kernel");
    CUDA_OK(cudaEventRecord(p->ev[1], s));
    p->launches += 1;
    long long n_in = p->ncells0;
    int cur = 0;
    float ms = 0;
    for (int pass = 0; pass < 32 && n_in > 0; pass++) {
        p->passes = pass + 1;
        // 1. theta integrals of every newly marked lattice node
        count_kernel<<<tgrid, 256, 0, s>>>(p->d_items, p->d_regimes, p->d_tasks, p->ntasks, p->d_want, p->d_known, p->d_have_lim, p->d_counts);
        LAUNCHED("count_kernel");
        {
            const int nchunks = (p->ntasks + SCAN_CHUNK - 1) / SCAN_CHUNK;
            scan_chunk_kernel<<<nchunks, 1024, 0, s>>>(p->d_counts, p->d_offsets, p->ntasks, p->d_chunk_sums);
            scan_sums_kernel<<<1, 1024, 0, s>>>(p->d_chunk_sums, nchunks, p->d_offsets, p->ntasks);
            scan_add_kernel<<<nchunks, 1024, 0, s>>>(p->d_offsets, p->ntasks, p->d_chunk_sums);
        }
        LAUNCHED("scan kernels");
        CUDA_OK(cudaMemcpyAsync(&p->h_pinned[1], p->d_offsets + p->ntasks, sizeof(int), cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));
        cudaEventElapsedTime(&ms, p->ev[0], p->ev[1]); p->ms_mesh += ms;
        p->launches += 4;
        const int nlist = p->h_pinned[1];
        if (nlist > p->max_list) { set_error("kg_omega_plan_run: work list overflow"); return -3; }
        if (nlist > 0) {
            fill_kernel<<<(p->ntasks + 7) / 8, 256, 0, s>>>(p->d_items, p->d_regimes, p->d_tasks, p->ntasks, p->d_want, p->d_known, p->d_have_lim, p->d_offsets, p->d_list);
            LAUNCHED("fill_kernel");
            CUDA_OK(cudaMemsetAsync(p->d_tile_counter, 0, (m->npairs + 1) * sizeof(int), s));
            const int ntiles = (nlist + THETA_THREADS - 1) / THETA_THREADS;
            const int grid = std::min(ntiles, 2 * m->sm_count);
            CUDA_OK(cudaEventRecord(p->ev[2], s));
#define LAUNCH_THETA(NT) if (NT != 1) theta_setup_kernel<NT><<<(nlist + 255) / 256, 256, 0, s>>>(p->d_items, p->d_regimes, m->d_pairs, p->d_coeff, p->d_list, nlist, \
                p->d_prm, p->max_list, p->d_status); \
            theta_kernel<NT><<<grid, THETA_THREADS, smem, s>>>(p->d_items, p->d_regimes, m->d_pairs, m->d_rows_by_pair, p->d_coeff, \
                p->d_list, p->d_offsets, p->d_pair_task_start, m->npairs, p->d_tile_counter, p->d_tab, p->d_lim, p->d_status, p->d_counters, \
                p->d_prm, p->max_list); \
            LAUNCHED("theta kernels")
            switch (m->ntpad) { case 1: LAUNCH_THETA(1); break; case 2: LAUNCH_THETA(2); break; case 4: LAUNCH_THETA(4); break; default: LAUNCH_THETA(6); break; }
#undef LAUNCH_THETA
            CUDA_OK(cudaEventRecord(p->ev[3], s));
            finalize_kernel<<<(nlist + 255) / 256, 256, 0, s>>>(p->d_items, p->d_regimes, p->d_list, p->d_offsets + p->ntasks, p->d_tab, p->d_lim,
                                                               p->d_known, p->d_have_lim);
            p->launches += (m->ntpad != 1) ? 4 : 3;
        }
        // 2. evaluate this level's cells, emit the next level
        CUDA_OK(cudaMemsetAsync(p->d_ncells, 0, sizeof(int), s));
        CUDA_OK(cudaEventRecord(p->ev[0], s));
        mesh_level_kernel<<<(unsigned) ((n_in + 255) / 256), 256, 0, s>>>(p->d_items, p->d_regimes, p->d_tab, p->d_want, p->d_gfac, p->d_lr_l, p->d_lr_ri,
            p->nr_distinct, p->gfac_stride, p->cospi, p->nlr, p->d_cells[cur], (int) n_in, p->d_cells[cur ^ 1], p->d_ncells, (int) p->cell_cap,
            p->d_W, p->d_flags);
        CUDA_OK(cudaEventRecord(p->ev[1], s));
        CUDA_OK(cudaMemcpyAsync(&p->h_pinned[2], p->d_ncells, sizeof(int), cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaMemcpyAsync(&p->h_pinned[0], p->d_flags, 2 * sizeof(int), cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaGetLastError());
        CUDA_OK(cudaStreamSynchronize(s));
        if (nlist > 0) { cudaEventElapsedTime(&ms, p->ev[2], p->ev[3]); p->ms_theta += ms; }
        p->launches += 1;
        if (p->h_pinned[1]) {
            // the working estimate of the cell lists was too small: grow them and start over (results are rebuilt from scratch)
            if (p->cell_cap >= p->cell_cap_limit) { set_error("kg_omega_plan_run: cell list overflow"); return -3; }
            const long long grown = std::min(p->cell_cap_limit, 2 * p->cell_cap);
            if (grown >= (1LL << 31)) { set_error("kg_omega_plan_run: batch too large, split the T grid"); return -3; }
            m->dfree(p->d_cells[0]); m->dfree(p->d_cells[1]); p->d_cells[0] = p->d_cells[1] = nullptr;
            CUDA_OK(kg_dmalloc(m, &p->d_cells[0], grown * sizeof(Cell)));
            CUDA_OK(kg_dmalloc(m, &p->d_cells[1], grown * sizeof(Cell)));
            p->cell_cap = grown; p->regrows++;
            return kg_omega_plan_run(p);
        }
        p->max_cells_seen = std::max(p->max_cells_seen, p->h_pinned[2]);
        n_in = p->h_pinned[2];
        cur ^= 1;
    }
    cudaEventElapsedTime(&ms, p->ev[0], p->ev[1]); p->ms_mesh += ms;
    if (n_in > 0) { set_error("kg_omega_plan_run: refinement did not terminate"); return -3; }
    permute_kernel<<<(p->nitems * p->nlr + 255) / 256, 256, 0, s>>>(p->d_W, p->d_items, p->d_perm, p->nitems, p->nlr, p->d_W_out, p->d_omega_out);
    p->launches += 1;
    CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int kg_omega_plan_sync(kg_omega_plan* p) {
    KG_TRACE();
    if (!p) return -1;
    auto set_error = [&](const std::string& s) { p->m->err = s; };
    CUDA_OK(cudaStreamSynchronize(p->m->stream));
    return 0;
}

extern "C" int kg_omega_plan_fetch(kg_omega_plan* p, double* omega, double* w, int32_t* status) {
    KG_TRACE();
    if (!p) return -1;
    kg_model* m = p->m;
    auto set_error = [&](const std::string& s) { m->err = s; };
    const size_t n = (size_t) p->nitems * p->nlr;
    if (omega) CUDA_OK(cudaMemcpyAsync(omega, p->d_omega_out, n * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    if (w) CUDA_OK(cudaMemcpyAsync(w, p->d_W_out, n * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    std::vector<int> st;
    if (status) { st.resize(p->nitems); CUDA_OK(cudaMemcpyAsync(st.data(), p->d_status, p->nitems * sizeof(int), cudaMemcpyDeviceToHost, m->stream)); }
    CUDA_OK(cudaStreamSynchronize(m->stream));
    p->d2h_bytes = (omega ? n * sizeof(double) : 0) + (w ? n * sizeof(double) : 0) + (status ? p->nitems * sizeof(int) : 0);
    if (status) {
        // plan order is pair-major; caller order is [nT][npairs]
        const int Np = m->npairs;
        for (int pi = 0; pi < Np; pi++) for (int t = 0; t < p->nT; t++) status[t * Np + pi] = st[pi * p->nT + t];
        if (omega) for (int i = 0; i < p->nitems; i++) for (int k = 0; k < p->nlr; k++) if (!std::isfinite(omega[(size_t) i * p->nlr + k])) status[i] |= KG_STATUS_NAN;
    }
    return 0;
}

extern "C" const double* kg_omega_plan_omega_dev(const kg_omega_plan* p) { return p ? p->d_omega_out : nullptr; }

extern "C" int kg_omega_plan_export_dev(kg_omega_plan* p, double* omega_dev, double* w_dev) {
    KG_TRACE();
    if (!p) return -1;
    auto set_error = [&](const std::string& s) { p->m->err = s; };
    const size_t n = (size_t) p->nitems * p->nlr;
    if (omega_dev) CUDA_OK(cudaMemcpyAsync(omega_dev, p->d_omega_out, n * sizeof(double), cudaMemcpyDeviceToDevice, p->m->stream));
    if (w_dev) CUDA_OK(cudaMemcpyAsync(w_dev, p->d_W_out, n * sizeof(double), cudaMemcpyDeviceToDevice, p->m->stream));
    return 0;
}

extern "C" int kg_omega_plan_counters(kg_omega_plan* p, double* out16) {
    KG_TRACE();
    if (!p || !out16) return -1;
    auto set_error = [&](const std::string& s) { p->m->err = s; };
    unsigned long long c[8];
    CUDA_OK(cudaMemcpy(c, p->d_counters, sizeof(c), cudaMemcpyDeviceToHost));
    out16[0] = (double) c[0]; out16[1] = (double) c[1]; out16[2] = (double) c[2]; out16[3] = (double) c[3];
    out16[4] = p->ms_theta; out16[5] = 0.0; out16[6] = p->ms_mesh; out16[7] = p->launches; out16[8] = p->passes;
    out16[9] = (double) p->h2d_bytes; out16[10] = (double) p->d2h_bytes;
    out16[11] = (double) (p->tab_doubles * 8 + p->mask_words * 8);
    out16[5] = (double) c[4]; out16[12] = (double) p->max_cells_seen; out16[13] = (double) p->cell_cap; out16[14] = (double) p->cell_cap_limit; out16[15] = (double) p->regrows;
    return 0;
}

extern "C" int kg_omega_plan_quadrature_stats(kg_omega_plan* p, double* out4) {
    KG_TRACE();
    if (!p || !out4) return -1;
    auto set_error = [&](const std::string& s) { p->m->err = s; };
    unsigned long long c[8];
    CUDA_OK(cudaMemcpy(c, p->d_counters, sizeof(c), cudaMemcpyDeviceToHost));
    out4[0] = (double) (c[1] + c[3]); out4[1] = (double) c[4]; out4[2] = (double) c[5]; out4[3] = (double) (c[0] + c[2]);
    return 0;
}

extern "C" int kg_set_exact_quadrature(int exact) {
    KG_TRACE();
    const int v = exact ? 1 : 0;
    return cudaMemcpyToSymbol(c_ts_exact, &v, sizeof(int)) == cudaSuccess ? 0 : -2;
}

extern "C" int kg_omega_batch(kg_model* m, int nT, const double* T, int nlr, const int32_t* l, const int32_t* r,
                              const double* coeff_T, double* omega, double* w, int32_t* status) {
    KG_TRACE();
    kg_omega_plan* p = nullptr;
    int rc = kg_omega_plan_create(m, nT, T, nlr, l, r, coeff_T, &p);
    if (rc) return rc;
    rc = kg_omega_plan_run(p);
    if (!rc) rc = kg_omega_plan_fetch(p, omega, w, status);
    kg_omega_plan_destroy(p);
    return rc;
}

extern "C" int kg_measure_fp64_peak(int device, double* tflops, double* sm_mhz_est) {
    KG_TRACE();
    if (!tflops) return -1;
    if (cudaSetDevice(device) != cudaSuccess) return -2;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -2;
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
    double* d = nullptr;
    if (cudaMalloc(&d, (size_t) blocks * threads * sizeof(double)) != cudaSuccess) return -2;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    double best = 0;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(a);
        dfma_probe_kernel<<<blocks, threads>>>(d, iters, 0.999999, 1e-9);
        cudaEventRecord(b);
        if (cudaEventSynchronize(b) != cudaSuccess) { cudaFree(d); return -2; }
        float ms = 0; cudaEventElapsedTime(&ms, a, b);
        const double flops = 2.0 * 64.0 * iters * (double) blocks * threads;
        best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(d);
    *tflops = best;
    if (sm_mhz_est) *sm_mhz_est = best * 1e12 / (2.0 * 64.0 * prop.multiProcessorCount) / 1e6;
    return 0;
}

#include "kg_tl.cuh"
