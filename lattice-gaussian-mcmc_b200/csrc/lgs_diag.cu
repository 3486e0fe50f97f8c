// Device-side chain diagnostics on kept states that never leave the GPU: per-chain mean / variance /
// autocorrelation / integrated autocorrelation time of tracked coordinates, and the per-coordinate sums the
// Gelman-Rubin statistic needs.  Reference arithmetic:
//   compute_autocorrelation          src/diagnostics/convergence_diag.py:75-112, mcmc_diag.py:12-33
//       acf[k] = sum_{t < N-k} (x_t - mean)(x_{t+k} - mean) / sum_t (x_t - mean)^2      (biased, normalised by lag 0;
//       the reference gets it from an FFT / np.correlate, the same numbers up to rounding)
//   integrated_autocorrelation_time  convergence_diag.py:115-145, mcmc_diag.py:36-57
//       tau = 1 + sum_{k=1..K} 2 acf[k],  K = first k with k >= c * (sum so far), else max_lag
//   gelman_rubin_statistic           convergence_diag.py:176-213   (chain means, ddof=1 variances)
#include "lgs_internal.h"

#include <mutex>

namespace lgs {

namespace {

constexpr int kDiagThreads = 128;

// V[chain][t][n] (the layout lgs_imhk_run writes) -> Y[chain][dim][t], coalesced writes
__global__ void __launch_bounds__(256)
diag_gather_kernel(const double* __restrict__ V, int64_t chains, int64_t keep, int n, const int32_t* __restrict__ dims,
                   int n_dims, double* __restrict__ Y) {
    const int64_t total = chains * (int64_t)n_dims * keep;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = i % keep;
        const int64_t cd = i / keep;
        const int d = (int)(cd % n_dims);
        const int64_t c = cd / n_dims;
        Y[i] = V[((size_t)c * keep + t) * n + dims[d]];
    }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// one CTA per series.  out[series][kDiagFields] = {mean, var (ddof 1), tau_int, window, acf[1]}
__global__ void __launch_bounds__(kDiagThreads)
diag_series_kernel(const double* __restrict__ Y, int64_t keep, double cwin, int max_lag, double* __restrict__ out,
                   double* __restrict__ acf_out, int acf_len) {
    __shared__ double red[kDiagThreads / 32];
    __shared__ double lagv[kDiagThreads / 32];
    __shared__ double s_mean, s_c0;
    __shared__ int s_stop;
    const double* __restrict__ x = Y + (size_t)blockIdx.x * keep;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int NW = kDiagThreads / 32;
    // mean
    double s = 0.0;
    for (int64_t t = tid; t < keep; t += kDiagThreads) s += x[t];
    s = warp_sum(s);
    if (lane == 0) red[warp] = s;
    __syncthreads();
    if (tid == 0) {
        double a = 0.0;
        for (int w = 0; w < NW; w++) a += red[w];
        s_mean = a / (double)keep;
    }
    __syncthreads();
    const double mean = s_mean;
    // lag 0
    s = 0.0;
    for (int64_t t = tid; t < keep; t += kDiagThreads) {
        const double d = x[t] - mean;
        s = fma(d, d, s);
    }
    s = warp_sum(s);
    __syncthreads();
    if (lane == 0) red[warp] = s;
    __syncthreads();
    if (tid == 0) {
        double a = 0.0;
        for (int w = 0; w < NW; w++) a += red[w];
        s_c0 = a;
        s_stop = 0;
    }
    __syncthreads();
    const double c0 = s_c0;
    double* acf_row = acf_out ? acf_out + (size_t)blockIdx.x * acf_len : nullptr;
    if (acf_row && tid == 0) acf_row[0] = 1.0;   // c0 / c0 (nan for a constant series, like the reference)
    // lags 1 .. max_lag, one warp per lag, NW lags per round; thread 0 applies Sokal's window in lag order
    double tau = 0.0, acf1 = 0.0;
    int window = max_lag;
    for (int k0 = 1; k0 <= max_lag; k0 += NW) {
        const int k = k0 + warp;
        double a = 0.0;
        if (k <= max_lag) {
            for (int64_t t = lane; t + k < keep; t += 32) a = fma(x[t] - mean, x[t + k] - mean, a);
            a = warp_sum(a);
        }
        if (lane == 0) lagv[warp] = a / c0;
        __syncthreads();
        if (tid == 0) {
            for (int w = 0; w < NW && k0 + w <= max_lag; w++) {
                const int kk = k0 + w;
                if (acf_row && kk < acf_len) acf_row[kk] = lagv[w];
                if (s_stop) continue;
                if (kk == 1) acf1 = lagv[w];
                tau += 2.0 * lagv[w];
                if ((double)kk >= cwin * tau) {
                    window = kk;
                    s_stop = 1;
                }
            }
        }
        __syncthreads();
        // the window has closed: later lags are only needed when the caller asked for the ACF itself
        if (s_stop && (!acf_row || k0 + NW >= acf_len)) break;
    }
    if (tid == 0) {
        double* o = out + (size_t)blockIdx.x * kDiagFields;
        o[0] = mean;
        o[1] = c0 / (double)(keep - 1);
        o[2] = 1.0 + tau;
        o[3] = (double)window;
        o[4] = acf1;
    }
}

// sums[d][3] += {sum_c mean, sum_c mean^2, sum_c var} over the chains of this rank; one CTA per coordinate
__global__ void __launch_bounds__(256)
diag_gr_sums_kernel(const double* __restrict__ stats, int64_t chains, int n_dims, double* __restrict__ sums) {
    __shared__ double red[3][8];
    const int d = blockIdx.x;
    double a = 0.0, b = 0.0, c = 0.0;
    for (int64_t ch = threadIdx.x; ch < chains; ch += blockDim.x) {
        const double* o = stats + ((size_t)ch * n_dims + d) * kDiagFields;
        a += o[0];
        b = fma(o[0], o[0], b);
        c += o[1];
    }
    a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { red[0][warp] = a; red[1][warp] = b; red[2][warp] = c; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double sa = 0.0, sb = 0.0, sc = 0.0;
        for (int w = 0; w < 8; w++) { sa += red[0][w]; sb += red[1][w]; sc += red[2][w]; }
        sums[d * 3 + 0] += sa;
        sums[d * 3 + 1] += sb;
        sums[d * 3 + 2] += sc;
    }
}

}  // namespace

cudaError_t launch_chain_diagnostics(const double* V, int64_t chains, int64_t keep, int n, const int32_t* dims_dev,
                                     int n_dims, double cwin, int max_lag, double* scratch, double* stats, double* acf,
                                     int acf_len, cudaStream_t st) {
    const int64_t total = chains * (int64_t)n_dims * keep;
    int64_t blocks = (total + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    diag_gather_kernel<<<(unsigned)blocks, 256, 0, st>>>(V, chains, keep, n, dims_dev, n_dims, scratch);
    diag_series_kernel<<<(unsigned)(chains * n_dims), kDiagThreads, 0, st>>>(scratch, keep, cwin, max_lag, stats, acf,
                                                                             acf_len);
    return cudaGetLastError();
}

cudaError_t launch_gelman_rubin_sums(const double* stats, int64_t chains, int n_dims, double* sums, cudaStream_t st) {
    diag_gr_sums_kernel<<<(unsigned)n_dims, 256, 0, st>>>(stats, chains, n_dims, sums);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// lgs_chain_diagnostics_static: the same two kernels for callers that must not allocate device memory.
//
// launch_chain_diagnostics above needs a work array Y[chain][dim][t] (the tracked series, gathered so that the series
// kernel reads them contiguously) and a device copy of dims[]; lgs_chain_diagnostics obtains both with cudaMalloc and
// grows the work array when a larger request arrives.  Some users of this library cannot afford that in their
// steady-state loop: a service whose framework has already handed the whole device to its own memory pool (a further
// cudaMalloc fails or fragments the pool's budget), a sampler/diagnostics loop with a latency bound (cudaMalloc and
// cudaFree synchronise the whole device), and code that runs in a phase where the CUDA runtime forbids allocation.
// For them the work array below is part of the library image: it exists from the moment a GPU loads this module
// until the process exits, on EVERY GPU that loads it (kDiagStaticDoubles * 8 bytes = 32 MiB each), and is never
// resized.  dims[] travels by value in the kernel parameters, so nothing else is needed.
//
// g_diag_static is a device variable.  In host code its name only stands for the symbol; it is not a device address
// and must not be handed to a kernel.  diag_static_base() asks the runtime for the real address, once per device,
// with that device current, and the launcher passes that pointer on.
// ---------------------------------------------------------------------------------------------------------------
namespace {

constexpr int64_t kDiagStaticDoubles = LGS_DIAG_STATIC_MAX_DOUBLES;
static __device__ __align__(256) double g_diag_static[kDiagStaticDoubles];

constexpr int kDiagStaticMaxDevices = 64;
std::mutex g_diag_static_mu;
double* g_diag_static_ptr[kDiagStaticMaxDevices];   // device address of g_diag_static on device i; nullptr = not asked yet

struct DiagDims {
    int32_t d[256];   // 1 KB of kernel parameters
};

// as diag_gather_kernel, with the coordinate list in the parameters instead of a device array
__global__ void __launch_bounds__(256)
diag_gather_byvalue_kernel(const double* __restrict__ V, int64_t chains, int64_t keep, int n, const DiagDims dims,
                           int n_dims, double* __restrict__ Y) {
    const int64_t total = chains * (int64_t)n_dims * keep;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = i % keep;
        const int64_t cd = i / keep;
        const int d = (int)(cd % n_dims);
        const int64_t c = cd / n_dims;
        Y[i] = V[((size_t)c * keep + t) * n + dims.d[d]];
    }
}

// The caller has made `device` current.  First call per device: cudaGetSymbolAddress (which also makes the runtime
// load this module on that device, if it has not yet).  Later calls: a table look-up.
cudaError_t diag_static_base(int device, double** base) {
    if (device < 0 || device >= kDiagStaticMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_diag_static_mu);
    if (!g_diag_static_ptr[device]) {
        int current = -1;
        cudaError_t e = cudaGetDevice(&current);
        if (e != cudaSuccess) return e;
        if (current != device) return cudaErrorInvalidDevice;
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_diag_static);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_diag_static_ptr[device] = static_cast<double*>(p);
    }
    *base = g_diag_static_ptr[device];
    return cudaSuccess;
}

}  // namespace

int64_t diag_static_capacity() { return kDiagStaticDoubles; }

// No cudaMalloc / cudaFree / cudaMemcpy here or below.  Nothing in the work array has to start at zero: the gather
// kernel overwrites every element [0, chains * n_dims * keep) the series kernel will read, so whatever the previous
// call left behind is never seen.  All calls on a device share the one array: two calls whose kernels can overlap on
// the GPU (different streams without an ordering between them, or two host threads) race on it.
cudaError_t launch_chain_diagnostics_static(int device, const double* V, int64_t chains, int64_t keep, int n,
                                            const int32_t* dims_host, int n_dims, double cwin, int max_lag,
                                            double* stats, double* acf, int acf_len, cudaStream_t st) {
    double* work = nullptr;
    cudaError_t e = diag_static_base(device, &work);
    if (e != cudaSuccess) return e;
    DiagDims dims;
    for (int i = 0; i < 256; i++) dims.d[i] = i < n_dims ? dims_host[i] : 0;
    const int64_t total = chains * (int64_t)n_dims * keep;
    int64_t blocks = (total + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    diag_gather_byvalue_kernel<<<(unsigned)blocks, 256, 0, st>>>(V, chains, keep, n, dims, n_dims, work);
    diag_series_kernel<<<(unsigned)(chains * n_dims), kDiagThreads, 0, st>>>(work, keep, cwin, max_lag, stats, acf,
                                                                             acf_len);
    return cudaGetLastError();
}

}  // namespace lgs
