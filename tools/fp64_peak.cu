// FP64 peak calibration for B200 (sm_100a): DFMA (CUDA-core) vs DMMA (mma.sync f64) issue rates,
// plus the auxiliary pipes the SampleZ path leans on (fp64 exp, int->fp64 conversion).
// MEASURED_PEAKS.json carries no FP64 figure, so bench.py reads the number this prints.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/fp64_peak tools/fp64_peak.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

template <int NACC>
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double a, double b) {
    double acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = threadIdx.x + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++) acc[i] = fma(acc[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) k_dmma884(double* out, int iters, double a, double b) {
    double c[NACC][2];
#pragma unroll
    for (int i = 0; i < NACC; i++) { c[i][0] = threadIdx.x; c[i][1] = i; }
    double fa = a + threadIdx.x * 1e-9, fb = b - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(fa), "d"(fb));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) k_dmma1688(double* out, int iters, double a, double b) {
    double c[NACC][4];
#pragma unroll
    for (int i = 0; i < NACC; i++) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
    double a0 = a + threadIdx.x * 1e-9, a1 = a * 0.5, a2 = a * 0.25, a3 = a * 0.125;
    double b0 = b - threadIdx.x * 1e-9, b1 = b * 0.5;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                         : "d"(a0), "d"(a1), "d"(a2), "d"(a3), "d"(b0), "d"(b1));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// half the warps of each CTA issue DFMA, the other half DMMA: do they share one pipe?
__global__ void __launch_bounds__(256) k_mixed(double* out, int iters, double a, double b) {
    int warp = threadIdx.x >> 5;
    double s = 0;
    if (warp & 1) {
        double acc[8];
#pragma unroll
        for (int i = 0; i < 8; i++) acc[i] = threadIdx.x + i;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int i = 0; i < 8; i++) acc[i] = fma(acc[i], a, b);
        }
#pragma unroll
        for (int i = 0; i < 8; i++) s += acc[i];
    } else {
        double c[4][2];
#pragma unroll
        for (int i = 0; i < 4; i++) { c[i][0] = threadIdx.x; c[i][1] = i; }
        double fa = a + threadIdx.x * 1e-9, fb = b;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int i = 0; i < 4; i++)
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(fa), "d"(fb));
        }
#pragma unroll
        for (int i = 0; i < 4; i++) s += c[i][0] + c[i][1];
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) k_exp(double* out, int iters, double a) {
    double x[4];
#pragma unroll
    for (int i = 0; i < 4; i++) x[i] = -1e-3 * (threadIdx.x + i);
    double s = 0;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 4; i++) { s += exp(x[i]); x[i] -= a; }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) k_i2d(double* out, int iters, int a) {
    int x[4];
#pragma unroll
    for (int i = 0; i < 4; i++) x[i] = threadIdx.x + i;
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    for (int it = 0; it < iters; it++) {
        s0 += (double)x[0]; s1 += (double)x[1]; s2 += (double)x[2]; s3 += (double)x[3];
#pragma unroll
        for (int i = 0; i < 4; i++) x[i] += a;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s0 + s1 + s2 + s3;
}

template <typename F>
static double time_ms(F launch, int reps) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; i++) launch();
    CK(cudaDeviceSynchronize());
    double best = 1e30;
    for (int r = 0; r < reps; r++) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

int main(int argc, char** argv) {
    int dev = 0; CK(cudaSetDevice(dev));
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, dev));
    int sms = p.multiProcessorCount;
    int clk_khz = 0; CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev));
    double* out; CK(cudaMalloc(&out, sizeof(double) * sms * 8 * 256));
    const int iters = 20000;
    const int grid = sms * 8;
    const double thr = (double)grid * 256;
    auto gf = [&](double flops_per_thread_iter, double ms) { return thr * iters * flops_per_thread_iter / (ms * 1e-3) / 1e12; };

    double ms;
    ms = time_ms([&] { k_dfma<8><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); }, 5);
    double dfma8 = gf(8 * 2, ms);
    ms = time_ms([&] { k_dfma<16><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); }, 5);
    double dfma16 = gf(16 * 2, ms);
    // one m8n8k4 = 8*8*4*2 = 512 flop per warp = 16 flop per thread
    ms = time_ms([&] { k_dmma884<4><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); }, 5);
    double dmma4 = gf(4 * 16, ms);
    ms = time_ms([&] { k_dmma884<8><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); }, 5);
    double dmma8 = gf(8 * 16, ms);
    // one m16n8k8 = 16*8*8*2 = 2048 flop per warp = 64 flop per thread
    ms = time_ms([&] { k_dmma1688<4><<<grid, 256>>>(out, iters / 4, 1.0000001, 1e-9); }, 5);
    double dmma1688 = thr * (iters / 4) * 4 * 64 / (ms * 1e-3) / 1e12;
    ms = time_ms([&] { k_mixed<<<grid, 256>>>(out, iters, 1.0000001, 1e-9); }, 5);
    // per CTA: 4 warps x 8 DFMA x 2 flop x 32 thr + 4 warps x 4 DMMA x 512 flop
    double mixed = (double)grid * iters * (4 * 8 * 2 * 32 + 4 * 4 * 512) / (ms * 1e-3) / 1e12;
    ms = time_ms([&] { k_exp<<<grid, 256>>>(out, iters / 10, 1e-4); }, 5);
    double gexp = thr * (iters / 10) * 4 / (ms * 1e-3) / 1e9;
    ms = time_ms([&] { k_i2d<<<grid, 256>>>(out, iters, 3); }, 5);
    double gi2d = thr * iters * 4 / (ms * 1e-3) / 1e9;

    double best = dfma16 > dfma8 ? dfma16 : dfma8;
    double bestm = dmma8 > dmma4 ? dmma8 : dmma4;
    if (dmma1688 > bestm) bestm = dmma1688;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_max_mhz\": %.0f, "
           "\"dfma_tflops\": %.3f, \"dfma8_tflops\": %.3f, \"dmma884_x4_tflops\": %.3f, \"dmma884_x8_tflops\": %.3f, "
           "\"dmma1688_tflops\": %.3f, \"mixed_dfma_dmma_tflops\": %.3f, \"exp_f64_gops\": %.1f, \"i2d_gops\": %.1f, "
           "\"fp64_peak_tflops\": %.3f, \"nominal_dfma_tflops_at_max_clock\": %.3f}\n",
           p.name, sms, clk_khz / 1e3, dfma16, dfma8, dmma4, dmma8, dmma1688, mixed, gexp, gi2d,
           best > bestm ? best : bestm, sms * 128.0 * clk_khz * 1e3 / 1e12);
    return 0;
}
