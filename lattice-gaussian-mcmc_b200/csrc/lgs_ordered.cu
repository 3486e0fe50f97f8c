// Ordered path: one thread per sample / chain, every floating-point operation in the order the reference
// performs it (klein.py:181-220, 222-271; imhk.py:102-177).  Compiled with -fmad=false so that
// `sum += R[i,j] * x[j]` keeps its two roundings.  This is the bit-faithful path replay mode runs on; the blocked
// DMMA path (lgs_blocked.cu) is the throughput path.
#include "lgs_internal.h"

namespace lgs {

namespace {

constexpr int kThreads = 128;

struct Draws {
    RngSpec rng;
    Philox gen;
    __device__ Draws(const RngSpec& r) : rng(r), gen(r.seed) {}
    // uniform number k of unit `unit` in stream `tag`; replay streams are laid out by the caller
    __device__ __forceinline__ double get(int64_t unit, int64_t replay_pos, uint32_t k, uint32_t tag) const {
        if (rng.U) return rng.U[unit * rng.stride + replay_pos];
        return philox_uniform(gen, rng.first_id + (uint64_t)unit, k, tag);
    }
};

// One pass of sample_single (klein.py:181-220) for unit `s`, writing the coefficient vector as doubles into
// column `col` of Xc ([n][stride]); `unit` is the sample / chain index that selects the uniform stream.  Accumulates the Klein log-density (klein.py:246-269 with log_Z == 0) into *lp and
// sum_i log rho_{sigma_i,m_i}(Z) into *lz.  Returns the number of uniforms consumed.
__device__ int klein_draw(const DeviceCtx& ctx, double* __restrict__ Xc, int64_t stride, int64_t col, int64_t unit,
                          const Draws& dr, int64_t replay_base, uint32_t tag, int samplez, double* lp_out,
                          double* lz_out, int32_t* fallback) {
    const int n = ctx.n;
    double lp = 0.0, lz = 0.0;
    int used = 0;
    for (int i = n - 1; i >= 0; i--) {
        const double* __restrict__ Ri = ctx.R + (size_t)i * n;
        double csum = 0.0;
        for (int j = i + 1; j < n; j++) csum = __dadd_rn(csum, __dmul_rn(Ri[j], Xc[(size_t)j * stride + col]));
        double mean = __ddiv_rn(__dsub_rn(ctx.t[i], csum), ctx.rdiag[i]);
        double sg = ctx.sig[i];
        long long x;
        if (ctx.mode[i]) {
            x = __double2ll_rn(mean);
        } else {
            double u = dr.get(unit, replay_base + used, (uint32_t)ctx.uidx[i], tag);
            if (samplez == LGS_SAMPLEZ_TABLE) {
                bool ok = samplez_table(mean, sg, ctx.precision, ctx.max_range, u, &x);
                if (ok) used++; else if (fallback) atomicAdd(fallback, 1);
                if (lz_out) {
                    TableWindow tw = table_window(mean, sg, ctx.precision, ctx.max_range);
                    LogpGen lpg{(double)tw.lower, mean, sg};
                    lz += table_lse(lpg, tw.len);
                }
            } else {
                FastZ fz{ctx.fz_inv2s2[i], ctx.fz_q[i], ctx.fz_sq[i], ctx.fz_znorm[i], ctx.fz_eps1[i],
                         ctx.fz_kind[i], ctx.fz_dmax[i]};
                double zs;
                x = samplez_fast(mean, fz, u, &zs);
                used++;
                lz += log(zs);
            }
        }
        double xd = (double)x;
        Xc[(size_t)i * stride + col] = xd;
        double z = __ddiv_rn(__dsub_rn(xd, mean), sg);
        lp = __dadd_rn(lp, __dmul_rn(-0.5, __dmul_rn(z, z)));
        lp = __dsub_rn(lp, ctx.lognorm[i]);
    }
    *lp_out = lp;
    if (lz_out) *lz_out = lz;
    return used;
}

// lattice_point = basis @ x (klein.py:218) into column s of Vc, and ||v - c||^2
__device__ double lattice_point(const DeviceCtx& ctx, const double* __restrict__ Xc, double* __restrict__ Vc,
                                int64_t stride, int64_t s) {
    const int n = ctx.n;
    double d2 = 0.0;
    for (int r = 0; r < n; r++) {
        const double* __restrict__ Br = ctx.B + (size_t)r * n;
        double acc = 0.0;
        for (int j = 0; j < n; j++) acc = __dadd_rn(acc, __dmul_rn(Br[j], Xc[(size_t)j * stride + s]));
        if (Vc) Vc[(size_t)r * stride + s] = acc;
        double d = __dsub_rn(acc, ctx.center[r]);
        d2 = __dadd_rn(d2, __dmul_rn(d, d));
    }
    return d2;
}

__global__ void __launch_bounds__(kThreads)
klein_ordered_kernel(DeviceCtx ctx, double* __restrict__ Xw, int64_t stride, int64_t s0, int64_t S, RngSpec rng,
                     int samplez, double* __restrict__ V, int64_t* __restrict__ X, double* __restrict__ logq,
                     int32_t* fallback) {
    int64_t l = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // column in the workspace
    int64_t s = s0 + l;                                            // sample index of this call
    if (s >= S) return;
    Draws dr(rng);
    double lp;
    const int n = ctx.n;
    double* Xc = Xw;
    klein_draw(ctx, Xc, stride, l, s, dr, 0, 0u, samplez, &lp, nullptr, fallback);
    if (logq) logq[s] = lp;
    if (X) {
        for (int i = 0; i < n; i++) X[s * n + i] = (int64_t)Xc[(size_t)i * stride + l];
    }
    if (V) {
        for (int r = 0; r < n; r++) {
            const double* __restrict__ Br = ctx.B + (size_t)r * n;
            double acc = 0.0;
            for (int j = 0; j < n; j++) acc = __dadd_rn(acc, __dmul_rn(Br[j], Xc[(size_t)j * stride + l]));
            V[s * n + r] = acc;
        }
    }
}

// compute_log_density (klein.py:246-269) for given coefficient vectors, one thread per point
__global__ void __launch_bounds__(kThreads)
klein_logdensity_kernel(DeviceCtx ctx, int64_t S, const int64_t* __restrict__ X, double* __restrict__ logq) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    const int n = ctx.n;
    const int64_t* __restrict__ xs = X + s * n;
    double lp = 0.0;
    for (int i = n - 1; i >= 0; i--) {
        const double* __restrict__ Ri = ctx.R + (size_t)i * n;
        double csum = 0.0;
        for (int j = i + 1; j < n; j++) csum = __dadd_rn(csum, __dmul_rn(Ri[j], (double)xs[j]));
        double mean = __ddiv_rn(__dsub_rn(ctx.t[i], csum), ctx.rdiag[i]);
        double z = __ddiv_rn(__dsub_rn((double)xs[i], mean), ctx.sig[i]);
        lp = __dadd_rn(lp, __dmul_rn(-0.5, __dmul_rn(z, z)));
        lp = __dsub_rn(lp, ctx.lognorm[i]);
    }
    logq[s] = lp;
}

// imhk.py:126-177 for one chain per thread; burn-in and thinning inside the kernel.
__global__ void __launch_bounds__(kThreads)
imhk_ordered_kernel(DeviceCtx ctx, ChainSet cs, int64_t burn_in, int64_t keep, int64_t thin, RngSpec rng,
                    int samplez, int weight_mode, double* __restrict__ V, int64_t* __restrict__ X,
                    uint8_t* __restrict__ accept, double* __restrict__ logw_out, int32_t* fallback) {
    int64_t ch = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (ch >= cs.chains) return;
    const int n = ctx.n;
    const int64_t C = cs.chains;
    const int64_t steps = burn_in + keep * thin;
    Draws dr(rng);
    int cur = cs.cur[ch];
    double lw_cur = cs.logw[ch];
    int64_t stepno = cs.step_no[ch];
    int64_t n_acc = cs.counts[ch * 2], n_tot = cs.counts[ch * 2 + 1];
    int64_t upos = 0;
    const size_t bufsz = (size_t)n * C;
    const bool exact = weight_mode == LGS_WEIGHT_EXACT;

    if (!cs.init[ch]) {
        // _initialize_chain: one Klein sample and its importance weight
        double lp, lz;
        double* Xc = cs.Xbuf + cur * bufsz;
        upos += klein_draw(ctx, Xc, C, ch, ch, dr, upos, 0u, samplez, &lp, exact ? &lz : nullptr, fallback);
        double d2 = lattice_point(ctx, Xc, cs.Vbuf + cur * bufsz, C, ch);
        lw_cur = exact ? lz : __dsub_rn(__ddiv_rn(-d2, ctx.two_sigma2), lp);
        cs.init[ch] = 1;
    }
    if (logw_out) logw_out[ch * (steps + 1)] = lw_cur;

    for (int64_t st = 0; st < steps; st++) {
        stepno++;
        int nxt = cur ^ 1;
        double lp, lz;
        double* Xc = cs.Xbuf + nxt * bufsz;
        upos += klein_draw(ctx, Xc, C, ch, ch, dr, upos, (uint32_t)stepno, samplez, &lp, exact ? &lz : nullptr, fallback);
        double d2 = lattice_point(ctx, Xc, cs.Vbuf + nxt * bufsz, C, ch);
        double lw_prop = exact ? lz : __dsub_rn(__ddiv_rn(-d2, ctx.two_sigma2), lp);
        // imhk.py:118-122 only clamps the *returned weight*, the log weight is used as is
        double ratio = (lw_cur == -INFINITY) ? 1.0 : fmin(1.0, exp(__dsub_rn(lw_prop, lw_cur)));
        double ua = dr.get(ch, upos, kAcceptDraw, (uint32_t)stepno);
        upos++;
        bool acc = ua < ratio;
        n_tot++;
        if (acc) {
            cur = nxt;
            lw_cur = lw_prop;
            n_acc++;
        }
        if (accept) accept[ch * steps + st] = acc ? 1 : 0;
        if (logw_out) logw_out[ch * (steps + 1) + st + 1] = lw_prop;
        if (st >= burn_in && (st - burn_in + 1) % thin == 0) {
            int64_t kidx = (st - burn_in + 1) / thin - 1;
            const double* Xs = cs.Xbuf + cur * bufsz;
            const double* Vs = cs.Vbuf + cur * bufsz;
            size_t o = ((size_t)ch * keep + kidx) * n;
            if (X) for (int i = 0; i < n; i++) X[o + i] = (int64_t)Xs[(size_t)i * C + ch];
            if (V) for (int i = 0; i < n; i++) V[o + i] = Vs[(size_t)i * C + ch];
        }
    }
    cs.cur[ch] = cur;
    cs.logw[ch] = lw_cur;
    cs.step_no[ch] = stepno;
    cs.counts[ch * 2] = n_acc;
    cs.counts[ch * 2 + 1] = n_tot;
}

__global__ void imhk_get_state_kernel(DeviceCtx ctx, ChainSet cs, int64_t* __restrict__ counts,
                                      int64_t* __restrict__ X, double* __restrict__ V, double* __restrict__ logw) {
    int64_t ch = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (ch >= cs.chains) return;
    const int n = ctx.n;
    const size_t bufsz = (size_t)n * cs.chains;
    int cur = cs.cur[ch];
    if (counts) {
        counts[ch * 2] = cs.counts[ch * 2];
        counts[ch * 2 + 1] = cs.counts[ch * 2 + 1];
    }
    if (logw) logw[ch] = cs.logw[ch];
    if (X) for (int i = 0; i < n; i++) X[ch * n + i] = (int64_t)cs.Xbuf[cur * bufsz + (size_t)i * cs.chains + ch];
    if (V) for (int i = 0; i < n; i++) V[ch * n + i] = cs.Vbuf[cur * bufsz + (size_t)i * cs.chains + ch];
}

// lgs_imhk_set_state for the ordered kernels' state: coefficients into buffer 0, its lattice point recomputed
// (klein.py:218), log-weight / counters / step index as given
__global__ void imhk_set_state_kernel(DeviceCtx ctx, ChainSet cs, const int64_t* __restrict__ X,
                                      const double* __restrict__ logw, const int64_t* __restrict__ counts,
                                      int64_t steps_done) {
    int64_t ch = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (ch >= cs.chains) return;
    const int n = ctx.n;
    for (int i = 0; i < n; i++) cs.Xbuf[(size_t)i * cs.chains + ch] = (double)X[ch * n + i];
    lattice_point(ctx, cs.Xbuf, cs.Vbuf, cs.chains, ch);
    cs.cur[ch] = 0;
    cs.init[ch] = 1;
    cs.logw[ch] = logw[ch];
    cs.counts[ch * 2] = counts ? counts[ch * 2] : 0;
    cs.counts[ch * 2 + 1] = counts ? counts[ch * 2 + 1] : 0;
    cs.step_no[ch] = steps_done;
}

__global__ void reduce_counts_kernel(const int64_t* __restrict__ counts, int64_t chains, unsigned long long* out2) {
    unsigned long long a = 0, t = 0;
    for (int64_t ch = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; ch < chains; ch += (int64_t)gridDim.x * blockDim.x) {
        a += (unsigned long long)counts[ch * 2];
        t += (unsigned long long)counts[ch * 2 + 1];
    }
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_down_sync(0xffffffffu, a, o);
        t += __shfl_down_sync(0xffffffffu, t, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(out2, a);
        atomicAdd(out2 + 1, t);
    }
}

// compute_tvd's per-coordinate histograms (convergence_diag.py:15-72) on device-resident samples.  The tracked
// coordinates travel BY VALUE as a kernel parameter (up to 256 per launch): no per-handle device buffer, so calls on
// different streams cannot race on it.  Two variants: block-private shared-memory bins with one global atomic per bin
// and block (n_dims * nbins * 4 <= 160 KB), or global atomics straight into the int64 bins (any size; the bins of a
// 1024-dim marginal histogram over +-1500 are 25 MB and live in L2).
struct HistDims {
    int32_t d[kHistChunk];
};

__global__ void histogram_kernel(const double* __restrict__ V, int64_t S, int n, const __grid_constant__ HistDims dims,
                                 int n_dims, double lo, double inv_width, int nbins, unsigned long long* hist) {
    extern __shared__ unsigned int sbins[];
    const int total = n_dims * nbins;
    for (int i = threadIdx.x; i < total; i += blockDim.x) sbins[i] = 0;
    __syncthreads();
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < S * n_dims; e += (int64_t)gridDim.x * blockDim.x) {
        int64_t s = e / n_dims;
        int d = (int)(e - s * n_dims);
        double v = V[s * n + dims.d[d]];
        double b = floor((v - lo) * inv_width);
        int bi = b < 0.0 ? 0 : (b >= (double)nbins ? nbins - 1 : (int)b);
        atomicAdd(&sbins[d * nbins + bi], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < total; i += blockDim.x)
        if (sbins[i]) atomicAdd(&hist[i], (unsigned long long)sbins[i]);
}

__global__ void histogram_global_kernel(const double* __restrict__ V, int64_t S, int n,
                                        const __grid_constant__ HistDims dims, int n_dims, double lo, double inv_width,
                                        int nbins, unsigned long long* hist) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < S * n_dims; e += (int64_t)gridDim.x * blockDim.x) {
        int64_t s = e / n_dims;
        int d = (int)(e - s * n_dims);
        double v = V[s * n + dims.d[d]];
        double b = floor((v - lo) * inv_width);
        int bi = b < 0.0 ? 0 : (b >= (double)nbins ? nbins - 1 : (int)b);
        atomicAdd(&hist[(size_t)d * nbins + bi], 1ull);
    }
}

}  // namespace

cudaError_t launch_klein_ordered(const DeviceCtx& ctx, Scratch& sc, int64_t S, const RngSpec& rng, int samplez,
                                 double* V, int64_t* X, double* logq, int32_t* fallback, cudaStream_t st) {
    if (S <= 0) return cudaSuccess;
    const int64_t chunk = S < 32768 ? S : 32768;
    if (sc.cap < chunk) {
        if (sc.Xw) cudaFree(sc.Xw);
        sc.Xw = nullptr;
        sc.cap = 0;
        cudaError_t e = cudaMalloc(&sc.Xw, sizeof(double) * (size_t)ctx.n * chunk);
        if (e != cudaSuccess) return e;
        sc.cap = chunk;
    }
    for (int64_t s0 = 0; s0 < S; s0 += chunk) {
        int64_t cnt = (S - s0) < chunk ? (S - s0) : chunk;
        int blocks = (int)((cnt + kThreads - 1) / kThreads);
        klein_ordered_kernel<<<blocks, kThreads, 0, st>>>(ctx, sc.Xw, sc.cap, s0, S, rng, samplez, V, X, logq, fallback);
    }
    return cudaGetLastError();
}

cudaError_t launch_klein_logdensity(const DeviceCtx& ctx, int64_t S, const int64_t* X, double* logq, cudaStream_t st) {
    if (S <= 0) return cudaSuccess;
    int blocks = (int)((S + kThreads - 1) / kThreads);
    klein_logdensity_kernel<<<blocks, kThreads, 0, st>>>(ctx, S, X, logq);
    return cudaGetLastError();
}

cudaError_t launch_imhk_ordered(const DeviceCtx& ctx, ChainSet& cs, int64_t burn_in, int64_t keep, int64_t thin,
                                const RngSpec& rng, int samplez, int weight_mode, double* V, int64_t* X,
                                uint8_t* accept, double* logw_out, int32_t* fallback, cudaStream_t st) {
    if (cs.chains <= 0) return cudaSuccess;
    int blocks = (int)((cs.chains + kThreads - 1) / kThreads);
    imhk_ordered_kernel<<<blocks, kThreads, 0, st>>>(ctx, cs, burn_in, keep, thin, rng, samplez, weight_mode, V, X,
                                                     accept, logw_out, fallback);
    return cudaGetLastError();
}

cudaError_t launch_imhk_get_state(const DeviceCtx& ctx, ChainSet& cs, int64_t* counts, int64_t* X, double* V,
                                  double* logw, cudaStream_t st) {
    if (cs.chains <= 0) return cudaSuccess;
    int blocks = (int)((cs.chains + 127) / 128);
    imhk_get_state_kernel<<<blocks, 128, 0, st>>>(ctx, cs, counts, X, V, logw);
    return cudaGetLastError();
}

cudaError_t launch_imhk_set_state(const DeviceCtx& ctx, ChainSet& cs, const int64_t* X, const double* logw,
                                  const int64_t* counts, int64_t steps_done, cudaStream_t st) {
    if (cs.chains <= 0) return cudaSuccess;
    int blocks = (int)((cs.chains + 127) / 128);
    imhk_set_state_kernel<<<blocks, 128, 0, st>>>(ctx, cs, X, logw, counts, steps_done);
    return cudaGetLastError();
}

cudaError_t launch_reduce_counts(const ChainSet& cs, int64_t* out2, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(out2, 0, 2 * sizeof(int64_t), st);
    if (e != cudaSuccess) return e;
    if (cs.chains <= 0) return cudaSuccess;
    int blocks = (int)((cs.chains + 255) / 256);
    if (blocks > 1024) blocks = 1024;
    reduce_counts_kernel<<<blocks, 256, 0, st>>>(cs.counts, cs.chains, (unsigned long long*)out2);
    return cudaGetLastError();
}

cudaError_t launch_histogram(const DeviceCtx& ctx, const double* V, int64_t S, const int32_t* dims_host, int n_dims,
                             double lo, double width, int nbins, int64_t* hist, cudaStream_t st) {
    if (S <= 0 || n_dims <= 0) return cudaSuccess;
    for (int d0 = 0; d0 < n_dims; d0 += kHistChunk) {
        const int nd = n_dims - d0 < kHistChunk ? n_dims - d0 : kHistChunk;
        HistDims hd{};
        for (int i = 0; i < nd; i++) hd.d[i] = dims_host[d0 + i];
        unsigned long long* out = (unsigned long long*)hist + (size_t)d0 * nbins;
        const size_t smem = sizeof(unsigned int) * (size_t)nd * nbins;
        const int64_t work = S * nd;
        int blocks = (int)((work + 255) / 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        if (smem <= 160 * 1024) {
            if (smem > 48 * 1024) {
                cudaError_t e = cudaFuncSetAttribute(histogram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                if (e != cudaSuccess) return e;
            }
            histogram_kernel<<<blocks, 256, smem, st>>>(V, S, ctx.n, hd, nd, lo, 1.0 / width, nbins, out);
        } else {
            histogram_global_kernel<<<blocks * 2, 256, 0, st>>>(V, S, ctx.n, hd, nd, lo, 1.0 / width, nbins, out);
        }
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace lgs
