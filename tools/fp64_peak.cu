// FP64 peak calibration for B200 (sm_100a): DFMA (CUDA-core) vs DMMA (mma.sync f64) issue rates,
// plus the auxiliary pipes the SampleZ path leans on (fp64 exp, int->fp64 conversion).
// MEASURED_PEAKS.json carries no FP64 figure, so bench.py reads the number this prints.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/fp64_peak tools/fp64_peak.cu
#include <cstdio>
#include <string>
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

// --occupancy: how many DMMA warps per SM sub-partition does it take to fill the pipe?  Register-resident operands
// (k_occ<false>) and the chain kernel's k8-step shape -- 8 + 8 LDS.64 fragment loads, then 8 m16n8k8 on 8 independent
// accumulator tiles -- with operands in shared memory (k_occ<true>); one CTA per SM.
template <bool LDS>
__global__ void __launch_bounds__(1024, 1) k_occ(double* out, int iters, double a, double b) {
    extern __shared__ double sm[];
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    for (int i = threadIdx.x; i < 64 * 20 + 16 * 132; i += blockDim.x) sm[i] = a + i * 1e-9;
    __syncthreads();
    const double* A = sm;            // [64][20]
    const double* B = sm + 64 * 20;  // [16][132]
    double c[8][4];
#pragma unroll
    for (int i = 0; i < 8; i++) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
    double af[2][4] = {{a, a * 0.5, a * 0.25, a * 0.125}, {b, b * 0.5, b * 0.25, b * 0.125}};
    double bf[4][2] = {{b, a}, {b * 0.5, a}, {b * 0.25, a}, {b * 0.125, a}};
    for (int it = 0; it < iters; it++) {
        if (LDS) {
            asm volatile("" ::: "memory");   // reload every iteration, but let ptxas schedule the loads as it likes
            const int kk = (it & 1) * 8;
#pragma unroll
            for (int mt = 0; mt < 2; mt++) {
                const double* ap = &A[(mt * 16 + g) * 20 + kk + t];
                af[mt][0] = ap[0]; af[mt][1] = ap[8 * 20]; af[mt][2] = ap[4]; af[mt][3] = ap[8 * 20 + 4];
            }
#pragma unroll
            for (int nt = 0; nt < 4; nt++) {
                const double* bp = &B[(kk + t) * 132 + nt * 8 + g];
                bf[nt][0] = bp[0]; bf[nt][1] = bp[4 * 132];
            }
        }
#pragma unroll
        for (int mt = 0; mt < 2; mt++)
#pragma unroll
            for (int nt = 0; nt < 4; nt++)
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+d"(c[mt * 4 + nt][0]), "+d"(c[mt * 4 + nt][1]), "+d"(c[mt * 4 + nt][2]), "+d"(c[mt * 4 + nt][3])
                             : "d"(af[mt][0]), "d"(af[mt][1]), "d"(af[mt][2]), "d"(af[mt][3]), "d"(bf[nt][0]), "d"(bf[nt][1]));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
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
    if (argc > 1 && std::string(argv[1]) == "--occupancy") {
        const int smem = (64 * 20 + 16 * 132) * 8;
        printf("{\"dmma_tflops_by_warps_per_subpartition\": {");
        const int ws[] = {1, 2, 3, 4, 8};
        for (int lds = 0; lds < 2; lds++) {
            printf("%s\"%s\": {", lds ? ", " : "", lds ? "operands_from_shared_memory" : "register_operands");
            for (int i = 0; i < 5; i++) {
                const int w = ws[i], threads = 128 * w, it2 = 40000;
                ms = time_ms([&] {
                    if (lds) k_occ<true><<<sms, threads, smem>>>(out, it2, 1.0000001, 1e-9);
                    else k_occ<false><<<sms, threads, smem>>>(out, it2, 1.0000001, 1e-9);
                }, 5);
                printf("%s\"%d\": %.3f", i ? ", " : "", w, (double)sms * threads / 32 * it2 * 8 * 2048.0 / (ms * 1e-3) / 1e12);
            }
            printf("}");
        }
        printf("}}\n");
        return 0;
    }
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
