// Is the FP64 tensor path (mma.sync m8n8k4 f64, "DMMA") on B200 a pipe of its own next to the FP64 vector
// pipe (DFMA)?  Three kernels with the same loop structure: DFMA only, DMMA only, both interleaved.
// If time(both) ~ max(time(dfma), time(dmma)) the pipes are separate and a dense-gate pass that is FP64-bound
// could split its arithmetic between them; if ~ sum, they share the unit.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o dmma_probe dmma_probe.cu && ./dmma_probe
#include <cstdio>
#include <cuda_runtime.h>

#define CHAINS 8
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double a, double b) {
    double x[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) x[i] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) x[i] = fma(x[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

#define MCHAINS 4
__global__ void __launch_bounds__(256) k_dmma(double* out, int iters, double a, double b) {
    double c[MCHAINS][2];
#pragma unroll
    for (int i = 0; i < MCHAINS; ++i) { c[i][0] = threadIdx.x * 1e-3 + i; c[i][1] = i; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < MCHAINS; ++i) dmma(c[i][0], c[i][1], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < MCHAINS; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// per iteration: CHAINS DFMA (CHAINS flops*2 per thread) and MCHAINS DMMA (8 FMA per thread each)
__global__ void __launch_bounds__(256) k_both(double* out, int iters, double a, double b) {
    double x[CHAINS], c[MCHAINS][2];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) x[i] = threadIdx.x * 1e-3 + i;
#pragma unroll
    for (int i = 0; i < MCHAINS; ++i) { c[i][0] = threadIdx.x * 1e-3 + i; c[i][1] = i; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < MCHAINS; ++i) {
            dmma(c[i][0], c[i][1], a, b);
            x[2 * i] = fma(x[2 * i], a, b);
            x[2 * i + 1] = fma(x[2 * i + 1], a, b);
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += x[i];
#pragma unroll
    for (int i = 0; i < MCHAINS; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static float timed(F launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int blocks = p.multiProcessorCount * 8, threads = 256, iters = 20000;
    double* out; cudaMalloc(&out, sizeof(double) * blocks * threads);
    const double a = 0.999999, b = 1e-9;
    float t_f = timed([&] { k_dfma<<<blocks, threads>>>(out, iters, a, b); });
    float t_m = timed([&] { k_dmma<<<blocks, threads>>>(out, iters, a, b); });
    float t_b = timed([&] { k_both<<<blocks, threads>>>(out, iters, a, b); });
    const double n_thr = (double)blocks * threads;
    const double fl_f = n_thr * iters * CHAINS * 2.0;          // flops (FMA = 2)
    const double fl_m = n_thr * iters * MCHAINS * 8.0 * 2.0;   // m8n8k4 = 256 FMA per warp = 8 per thread
    printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
    printf("DFMA only : %8.3f ms  %7.2f TFLOP/s (vector)\n", t_f, fl_f / t_f / 1e9);
    printf("DMMA only : %8.3f ms  %7.2f TFLOP/s (tensor, m8n8k4)\n", t_m, fl_m / t_m / 1e9);
    printf("both      : %8.3f ms  %7.2f TFLOP/s total; sum of the two alone %8.3f ms, max %8.3f ms\n", t_b,
           (fl_f + fl_m) / t_b / 1e9, t_f + t_m, t_f > t_m ? t_f : t_m);
    printf("verdict   : both/sum = %.2f, both/max = %.2f  (1.00 vs sum = shared unit; 1.00 vs max = separate pipes)\n",
           t_b / (t_f + t_m), t_b / (t_f > t_m ? t_f : t_m));
    if (cudaGetLastError() != cudaSuccess) { printf("CUDA error\n"); return 1; }
    return 0;
}
