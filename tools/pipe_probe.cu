// tools/pipe_probe.cu -- micro-benchmark: do DFMA and MUFU.RSQ64H issue independently on sm_100a?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/pipe_probe tools/pipe_probe.cu
// Each variant runs the same number of DFMA per thread; variants add K MUFU.RSQ64H per 8 DFMA (independent chains).
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ double rsq64h(double x) { double y; asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x)); return y; }

template<int MUFU_PER_8>
__global__ void __launch_bounds__(512, 2) probe(double* out, int iters, double a, double b) {
    double x0 = threadIdx.x * 1e-3 + 1, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    double m0 = x0 + 0.5, m1 = x0 + 1.5;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
            if (MUFU_PER_8 >= 1) m0 = rsq64h(m0);
            if (MUFU_PER_8 >= 2) m1 = rsq64h(m1);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7 + m0 + m1;
}

// DFMA with three distinct per-thread register operands (what the node loop issues), optionally fed from shared memory
template<int LDS>
__global__ void __launch_bounds__(512, 2) probe3(double* out, int iters, const double* __restrict__ in) {
    __shared__ double sh[2048];
    for (int i = threadIdx.x; i < 2048; i += 512) sh[i] = in[i];
    __syncthreads();
    double x0 = threadIdx.x * 1e-3 + 1, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    double y0 = in[threadIdx.x], y1 = in[threadIdx.x + 1], y2 = in[threadIdx.x + 2], y3 = in[threadIdx.x + 3];
    double z0 = in[threadIdx.x + 4], z1 = in[threadIdx.x + 5], z2 = in[threadIdx.x + 6], z3 = in[threadIdx.x + 7];
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (LDS) {
                const double4 r0 = *reinterpret_cast<const double4*>(&sh[((i * 4 + j) * 8) & 2047]);
                const double4 r1 = *reinterpret_cast<const double4*>(&sh[((i * 4 + j) * 8 + 4) & 2047]);
                x0 = fma(x0, r0.x, z0); x1 = fma(x1, r0.y, z1); x2 = fma(x2, r0.z, z2); x3 = fma(x3, r0.w, z3);
                x4 = fma(x4, r1.x, z0); x5 = fma(x5, r1.y, z1); x6 = fma(x6, r1.z, z2); x7 = fma(x7, r1.w, z3);
            } else {
                x0 = fma(x0, y0, z0); x1 = fma(x1, y1, z1); x2 = fma(x2, y2, z2); x3 = fma(x3, y3, z3);
                x4 = fma(x4, y0, z1); x5 = fma(x5, y1, z2); x6 = fma(x6, y2, z3); x7 = fma(x7, y3, z0);
            }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}
template<int LDS> float run3(double* d, const double* in, int blocks, int iters) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(a);
        probe3<LDS><<<blocks, 512>>>(d, iters, in);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    return best;
}

template<int M> float run(double* d, int blocks, int iters) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(a);
        probe<M><<<blocks, 512>>>(d, iters, 0.999999, 1e-9);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int blocks = p.multiProcessorCount * 2, iters = 8192;
    double* d; cudaMalloc(&d, (size_t) blocks * 512 * sizeof(double));
    const double dfma = 32.0 * iters * blocks * 512;
    const float t0 = run<0>(d, blocks, iters), t1 = run<1>(d, blocks, iters), t2 = run<2>(d, blocks, iters);
    printf("{\"sms\": %d, \"dfma_only_ms\": %.4f, \"dfma_tflops\": %.2f, \"dfma_plus_1mufu_per8_ms\": %.4f, \"dfma_plus_2mufu_per8_ms\": %.4f, "
           "\"ratio_1\": %.4f, \"ratio_2\": %.4f}\n", p.multiProcessorCount, t0, 2 * dfma / (t0 * 1e-3) / 1e12, t1, t2, t1 / t0, t2 / t0);
    double* in; cudaMalloc(&in, 4096 * sizeof(double));
    { double h[4096]; for (int i = 0; i < 4096; i++) h[i] = 0.999999 + 1e-9 * i; cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice); }
    const float t3 = run3<0>(d, in, blocks, iters), t4 = run3<1>(d, in, blocks, iters);
    printf("{\"dfma_3reg_ms\": %.4f, \"ratio_3reg\": %.4f, \"dfma_lds_operand_ms\": %.4f, \"ratio_lds\": %.4f}\n", t3, t3 / t0, t4, t4 / t0);
    // expectation if the pipes are independent: ratio 1.0 until the XU pipe (8 cycles / warp instruction vs 2 for DFMA) saturates:
    // 1 MUFU per 8 DFMA = 8 vs 16 cycles -> 1.0 ; 2 per 8 = 16 vs 16 -> ~1.0. If they share an issue port: 1.5 and 2.0.
    return 0;
}
