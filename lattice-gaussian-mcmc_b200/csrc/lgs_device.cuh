// Device-side building blocks shared by the ordered and the blocked kernels:
// Philox4x32-10 counter RNG, the reference-order table SampleZ, and the fast SampleZ.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace lgs {

// ---------------------------------------------------------------------------------------------------------
// Per-sampler constants living in device memory (built once by lgs_create()).
// ---------------------------------------------------------------------------------------------------------
struct DeviceCtx {
    int n;            // lattice dimension
    int npad;         // n rounded up to the panel width of the blocked path
    int precision;    // window factor (klein.py:26)
    int max_range;    // 1000 (klein.py:122)
    int ups;          // uniforms per sample in replay mode
    double sigma;
    double inv_2sigma2;     // 1 / (2 sigma^2)   (base.py:144 divides by 2*sigma**2; kept as a divisor below)
    double two_sigma2;      // 2 * sigma**2
    const double* R;        // [n*n] row-major
    const double* B;        // [n*n] row-major, columns = basis vectors
    const double* t;        // [n]
    const double* center;   // [n]
    const double* rdiag;    // [n] R_ii
    const double* sig;      // [n] conditional sigma_i = sigma/|R_ii|, clamped as klein.py:205-208
    const double* lognorm;  // [n] 0.5*log(2*pi) + log(sigma_i)   (klein.py:262), computed on the host in numpy
    const uint8_t* mode;    // [n] 0 = sample, 1 = round (sigma_i < 1e-10, klein.py:201-204)
    const int* uidx;        // [n] index of coordinate i's uniform inside one sample's replay stream
    // fast SampleZ constants per coordinate
    const double* fz_inv2s2;  // 1/(2 sigma_i^2)
    const double* fz_q;       // exp(-1/sigma_i^2)
    const double* fz_sq;      // exp(-1/(2 sigma_i^2))
    const double* fz_znorm;   // sigma_i*sqrt(2 pi)
    const double* fz_eps1;    // 2*exp(-2 pi^2 sigma_i^2)   (first theta-series correction)
    const int* fz_kind;       // 0 analytic Z, 1 analytic Z + one cosine term, 2 summed Z (small sigma_i or truncated window),
                              // 3 wide truncated window (sigma_i >= 50): closed-form masses (window_mass_em)
    const int* fz_dmax;       // walk bound on each side of round(mean)
};

// ---------------------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011).  key = 64-bit seed, counter = (id_lo, id_hi, c2, c3).
// ---------------------------------------------------------------------------------------------------------
struct Philox {
    uint32_t k0, k1;
    __device__ __forceinline__ Philox(uint64_t seed) : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)) {}

    __device__ __forceinline__ uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const {
        uint32_t a = k0, b = k1;
#pragma unroll
        for (int r = 0; r < 10; r++) {
            uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            uint32_t n0 = hi1 ^ c1 ^ a, n1 = lo1, n2 = hi0 ^ c3 ^ b, n3 = lo0;
            c0 = n0; c1 = n1; c2 = n2; c3 = n3;
            a += 0x9E3779B9u; b += 0xBB67AE85u;
        }
        return make_uint4(c0, c1, c2, c3);
    }
};

// 53-bit uniform in [0,1) from two 32-bit words, the same bit recipe numpy's random_sample uses
__device__ __forceinline__ double u53(uint32_t a, uint32_t b) {
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) * (1.0 / 9007199254740992.0);
}

// Uniform for draw number k (k = 0 for coordinate n-1, ... as the reference consumes them) of unit `id`
// in stream `tag` (0 = Klein sample / IMHK initial state, s>=1 = proposal of IMHK step s).
__device__ __forceinline__ double philox_uniform(const Philox& g, uint64_t id, uint32_t k, uint32_t tag) {
    uint4 w = g((uint32_t)id, (uint32_t)(id >> 32), k >> 1, tag);
    return (k & 1) ? u53(w.z, w.w) : u53(w.x, w.y);
}
constexpr uint32_t kAcceptDraw = 0xFFFFFFFEu;  // draw index of the accept/reject uniform of a step

// ---------------------------------------------------------------------------------------------------------
// numpy pairwise summation over a generated sequence f(lo) .. f(lo+n-1)  (np.sum order; see the oracle).
// ---------------------------------------------------------------------------------------------------------
template <class F>
__device__ __noinline__ double pw_leaf(const F& f, int lo, int n) {
    if (n < 8) {
        double res = 0.0;
        for (int i = 0; i < n; i++) res = __dadd_rn(res, f(lo + i));
        return res;
    }
    double r0 = f(lo), r1 = f(lo + 1), r2 = f(lo + 2), r3 = f(lo + 3);
    double r4 = f(lo + 4), r5 = f(lo + 5), r6 = f(lo + 6), r7 = f(lo + 7);
    int i;
    for (i = 8; i < n - (n % 8); i += 8) {
        r0 = __dadd_rn(r0, f(lo + i));     r1 = __dadd_rn(r1, f(lo + i + 1));
        r2 = __dadd_rn(r2, f(lo + i + 2)); r3 = __dadd_rn(r3, f(lo + i + 3));
        r4 = __dadd_rn(r4, f(lo + i + 4)); r5 = __dadd_rn(r5, f(lo + i + 5));
        r6 = __dadd_rn(r6, f(lo + i + 6)); r7 = __dadd_rn(r7, f(lo + i + 7));
    }
    double res = __dadd_rn(__dadd_rn(__dadd_rn(r0, r1), __dadd_rn(r2, r3)),
                           __dadd_rn(__dadd_rn(r4, r5), __dadd_rn(r6, r7)));
    for (; i < n; i++) res = __dadd_rn(res, f(lo + i));
    return res;
}

template <int D, class F>
__device__ __forceinline__ double pw_sum(const F& f, int lo, int n) {
    if constexpr (D == 0) {
        return pw_leaf(f, lo, n);
    } else {
        if (n <= 128) return pw_leaf(f, lo, n);
        int n2 = n / 2;
        n2 -= n2 % 8;
        double a = pw_sum<D - 1>(f, lo, n2);
        double b = pw_sum<D - 1>(f, lo + n2, n - n2);
        return __dadd_rn(a, b);
    }
}
constexpr int kPwDepth = 4;  // enough for tables of up to 2048 entries (reference cap: 1001)

// ---------------------------------------------------------------------------------------------------------
// Reference-order table SampleZ  (klein.py:101-179 + numpy choice + scipy logsumexp; mirrors
// oracle/lgs_oracle.c:table_build/samplez_table but recomputes entries instead of storing them).
// Returns true when a uniform was consumed.
// ---------------------------------------------------------------------------------------------------------
struct TableWindow {
    long long lower;
    int len;
};

__device__ __forceinline__ TableWindow table_window(double mean, double sigma, int precision, int max_range) {
    double rf = (sigma < 0.1) ? (double)(precision > 3 ? precision : 3) : (double)precision;
    double w = __dmul_rn(rf, sigma);
    long long lower = (long long)floor(__dsub_rn(mean, w));
    long long upper = (long long)ceil(__dadd_rn(mean, w));
    if (upper - lower > max_range) {
        long long c = __double2ll_rn(mean);  // round half to even, as np.round
        lower = c - max_range / 2;
        upper = c + max_range / 2;
    }
    TableWindow tw;
    tw.lower = lower;
    tw.len = (int)(upper - lower + 1);
    return tw;
}

struct LogpGen {  // unnormalised log-probability of support point lower + k
    double lower, mean, sigma;
    __device__ __forceinline__ double operator()(int k) const {
        double z = __ddiv_rn(__dsub_rn(__dadd_rn(lower, (double)k), mean), sigma);
        return __dmul_rn(-0.5, __dmul_rn(z, z));
    }
};
struct ShiftExpGen {  // exp(a_k - a_max) with the maxima removed (scipy _logsumexp)
    LogpGen lp;
    double a_max;
    __device__ __forceinline__ double operator()(int k) const {
        double a = lp(k);
        return (a == a_max) ? 0.0 : exp(__dsub_rn(a, a_max));
    }
};
struct PlainExpGen {
    LogpGen lp;
    __device__ __forceinline__ double operator()(int k) const { return exp(lp(k)); }
};
struct ProbGen {  // exp(logp_k - lse)
    LogpGen lp;
    double lse;
    __device__ __forceinline__ double operator()(int k) const { return exp(__dsub_rn(lp(k), lse)); }
};

// log-sum-exp of the window, scipy order.  Also usable on its own for the exact importance weight.
__device__ __forceinline__ double table_lse(const LogpGen& lp, int len) {
    double a_max = -INFINITY;
    for (int k = 0; k < len; k++) a_max = fmax(a_max, lp(k));
    double m = 0.0;
    for (int k = 0; k < len; k++) m += (lp(k) == a_max) ? 1.0 : 0.0;
    ShiftExpGen se{lp, a_max};
    double s = pw_sum<kPwDepth>(se, 0, len);
    if (s != 0.0) s = __ddiv_rn(s, m);
    double lse = __dadd_rn(__dadd_rn(log1p(s), log(m)), a_max);
    if (!isfinite(lse)) {
        PlainExpGen pe{lp};
        lse = log(pw_sum<kPwDepth>(pe, 0, len));
    }
    return lse;
}

__device__ __forceinline__ bool samplez_table(double mean, double sigma, int precision, int max_range, double u,
                                              long long* out) {
    TableWindow tw = table_window(mean, sigma, precision, max_range);
    LogpGen lp{(double)tw.lower, mean, sigma};
    double lse = table_lse(lp, tw.len);
    ProbGen pg{lp, lse};
    double tot = pw_sum<kPwDepth>(pg, 0, tw.len);
    // cdf = cumsum(p / tot); cdf /= cdf[-1]
    double run = 0.0;
    bool bad = false;
    for (int k = 0; k < tw.len; k++) {
        double p = __ddiv_rn(pg(k), tot);
        bad |= !(p >= 0.0);
        run = (k == 0) ? p : __dadd_rn(run, p);
    }
    if (bad) {
        *out = __double2ll_rn(mean);
        return false;
    }
    double last = run;
    // searchsorted(cdf, u, side='right') == number of entries <= u (cdf is non-decreasing)
    int idx = 0;
    run = 0.0;
    for (int k = 0; k < tw.len; k++) {
        double p = __ddiv_rn(pg(k), tot);
        run = (k == 0) ? p : __dadd_rn(run, p);
        if (__ddiv_rn(run, last) <= u) idx = k + 1; else break;
    }
    if (idx >= tw.len) idx = tw.len - 1;
    *out = tw.lower + idx;
    return true;
}

// ---------------------------------------------------------------------------------------------------------
// Fast SampleZ: same distribution D_{Z, sigma_i, mean} restricted to the reference window, inverse transform
// over the ordering c0, c0+1, c0-1, c0+2, ... (c0 = round(mean)), weights by the two-term recurrence
//   rho(c0+d+1)/rho(c0+d) = r_d,  r_{d+1} = r_d * exp(-1/sigma_i^2),
// normaliser from the theta-series identity  sum_k rho(k) = sigma sqrt(2 pi) (1 + 2 e^{-2 pi^2 sigma^2} cos(2 pi m) + ...)
// when sigma_i is large enough for the omitted terms to be < 1e-17, otherwise summed.
// Returns the draw and, through *zsum, the window mass sum_k rho_{sigma_i,mean}(k) (used by LGS_WEIGHT_EXACT).
// ---------------------------------------------------------------------------------------------------------
struct FastZ {
    double inv2s2, q, sq, znorm, eps1;
    int kind, dmax;
};

// sum_{j=-d}^{d} exp(-a (j + delta)^2) in closed form, for the wide truncated windows (kind 3: sigma_i >= 50, where the
// reference cuts the table to max_range + 1 entries, klein.py:122-126).  Euler-Maclaurin: the integral (two erf), the
// end-point mean and the B2 / B4 / B6 derivative terms; g^(k) = (-1)^k a^(k/2) H_k(sqrt(a) s) g falls off as a^(k/2),
// so at a <= 2e-4 the first omitted term is < 1e-16 of the sum.  gu = g(d), gd = g(-d) are passed in (the caller
// needs them as walk weights anyway).  d = 0 gives g(0) exactly.
__device__ __forceinline__ double window_mass_em(double a, double sqrt_a, double delta, double d, double gu, double gd) {
    const double su = d + delta, sd = d - delta;          // |s| at the two ends; every odd derivative is odd in s
    const double integ = __ddiv_rn(0.88622692545275801365, sqrt_a) * (erf(sqrt_a * su) + erf(sqrt_a * sd));
    const double a2 = a * a, a3 = a2 * a;
    const double su2 = su * su, sd2 = sd * sd;
    const double t1 = -2.0 * a * (su * gu + sd * gd);                                          // 1st derivative, d vs -d
    const double t3 = (12.0 * a2 - 8.0 * a3 * su2) * su * gu + (12.0 * a2 - 8.0 * a3 * sd2) * sd * gd;   // 3rd
    const double t5 = -(((32.0 * a3 * a2 * su2 - 160.0 * a2 * a2) * su2 + 120.0 * a3) * su * gu +
                        ((32.0 * a3 * a2 * sd2 - 160.0 * a2 * a2) * sd2 + 120.0 * a3) * sd * gd);        // 5th
    return integ + 0.5 * (gu + gd) + t1 * (1.0 / 12.0) - t3 * (1.0 / 720.0) + t5 * (1.0 / 30240.0);
}

// WIDE = false compiles the closed-form branch out: the chain kernel instantiates it for lattices without kind-3
// coordinates, whose register allocation (and speed) must not pay for code they never run.
template <bool WIDE = true>
__device__ __forceinline__ long long samplez_fast(double mean, const FastZ& c, double u, double* zsum) {
    double c0 = rint(mean);
    double delta = c0 - mean;                     // in [-0.5, 0.5]
    double w0 = exp(-delta * delta * c.inv2s2);   // rho(c0)
    double ru, rd;                                // rho(c0+1)/rho(c0), rho(c0-1)/rho(c0)
    if (c.inv2s2 > 300.0) {
        // sigma_i < 0.041 (e.g. sigma / q on the q-vectors of an NTRU public basis): e1 below would overflow or
        // underflow and ru, rd become inf * 0.  The two ratios directly -- both exponents are <= 0 for |delta| <= 1/2.
        ru = exp(-c.inv2s2 * (1.0 + 2.0 * delta));
        rd = exp(-c.inv2s2 * (1.0 - 2.0 * delta));
    } else {
        const double e1 = exp(-2.0 * delta * c.inv2s2);     // ru = e1 * sq ; rd = sq / e1
        ru = e1 * c.sq;
        rd = c.sq / e1;
    }
    double Z;
    if (c.kind == 0) {
        Z = c.znorm;
    } else if (c.kind == 1) {
        Z = c.znorm * (1.0 + c.eps1 * (double)cospif(2.0f * (float)delta));
    } else if (WIDE && c.kind == 3) {
        const double D = (double)c.dmax;
        Z = window_mass_em(c.inv2s2, sqrt(c.inv2s2), delta, D, exp(-c.inv2s2 * (D + delta) * (D + delta)),
                           exp(-c.inv2s2 * (D - delta) * (D - delta)));
    } else {
        // summed normaliser over the (possibly truncated) window
        double su = w0, wu = w0, wd = w0, a = ru, b = rd;
        for (int d = 1; d <= c.dmax; d++) {
            wu *= a; a *= c.q;
            wd *= b; b *= c.q;
            su += wu + wd;
        }
        Z = su;
    }
    *zsum = Z;
    double target = u * Z;
    double cum = w0;
    long long x = (long long)c0;
    if (target < cum) return x;
    double wu = w0, wd = w0;
    int d = 1;
    if (WIDE && c.kind == 3) {
        // jump: an FP32 estimate of the exit distance from the symmetric integral, the exact mass inside distance
        // ds - 1 in closed form, then the ordinary walk from ds (a few steps).  The estimate only chooses where the walk
        // starts: if the mass there already exceeds the target (or the estimate is useless) the walk starts at 1.
        const double sa = sqrt(c.inv2s2);
        const float y = (float)(target * sa * 0.56418958354775628695);
        if (y < 0.99999f) {
            int ds = (int)floorf(erfinvf(y) / (float)sa - 0.5f) - 1;
            ds = ds > c.dmax ? c.dmax : ds;
            if (ds >= 3) {
                const double e = (double)(ds - 1);
                const double su = e + delta, sd = e - delta;
                const double gu = exp(-c.inv2s2 * su * su), gd = exp(-c.inv2s2 * sd * sd);
                const double base = window_mass_em(c.inv2s2, sa, delta, e, gu, gd);
                if (base <= target) {
                    cum = base, wu = gu, wd = gd;
                    ru = exp(-c.inv2s2 * (2.0 * su + 1.0));
                    rd = exp(-c.inv2s2 * (2.0 * sd + 1.0));
                    d = ds;
                }
            }
        }
    }
    for (; d <= c.dmax; d++) {
        wu *= ru; ru *= c.q;
        cum += wu;
        if (target < cum) return x + d;
        wd *= rd; rd *= c.q;
        cum += wd;
        if (target < cum) return x - d;
    }
    // u*Z beyond the accumulated mass (rounding, ~1e-16): the last point visited
    return x - c.dmax;
}

// ---------------------------------------------------------------------------------------------------------
// Wide truncated windows from per-coordinate TABLES (chain kernel: the FP64 draw behind samplez_wide_screen).
// The same map u -> x as samplez_fast<true> on a kind-3 coordinate, without a transcendental call:
//   M(d; delta) = sum_{|j| <= d} e^{-a (j + delta)^2}
//               = e^{-a delta^2} [1 + 2 sum_k (2 a delta)^{2k} / (2k)!  S_k(d)],   S_k(d) = sum_{j=1..d} j^{2k} e^{-a j^2},
// (the odd powers of the cosh series cancel over the symmetric window); with a <= 2e-4, |delta| <= 1/2, d <= 500 the
// argument 2 a delta d is below 0.1 and six terms leave < 3e-17 of the sum.  The walk weights are
// e^{-a (d +- delta)^2} = G(d) e^{-a delta^2} e^{-+ 2 a d delta}, G(d) = e^{-a d^2} tabulated, the last factor a short
// series.  Table layout per coordinate (chain_init): tab[k * (D + 1) + d] = S_k(d) for k = 0 .. 5, tab[6 (D + 1) + d] = G(d).
// Why: the closed form (two erf and six exp in FP64, ~800 instructions) ran on a whole warp whenever ONE of its 32
// lanes failed the FP32 screen -- 27 % of the warp-draws on a flat window, 22 % of BASELINE config C3's time.
// ---------------------------------------------------------------------------------------------------------
constexpr int kWideTabK = 6;   // S_0 .. S_5
__device__ __forceinline__ double exp_small(double x) {   // e^x, |x| <= 0.11: degree-10 Taylor, error < 1e-17
    double p = 1.0 / 3628800.0;
    p = fma(p, x, 1.0 / 362880.0);
    p = fma(p, x, 1.0 / 40320.0);
    p = fma(p, x, 1.0 / 5040.0);
    p = fma(p, x, 1.0 / 720.0);
    p = fma(p, x, 1.0 / 120.0);
    p = fma(p, x, 1.0 / 24.0);
    p = fma(p, x, 1.0 / 6.0);
    p = fma(p, x, 0.5);
    p = fma(p, x, 1.0);
    return fma(p, x, 1.0);
}
// M(d; delta) / e^{-a delta^2}; y2 = (2 a delta)^2
__device__ __forceinline__ double wide_mass_tab(const double* __restrict__ tab, int ld, int d, double y2) {
    double acc = __ldg(tab + 5 * ld + d) * (1.0 / 3628800.0);
    acc = fma(acc, y2, __ldg(tab + 4 * ld + d) * (1.0 / 40320.0));
    acc = fma(acc, y2, __ldg(tab + 3 * ld + d) * (1.0 / 720.0));
    acc = fma(acc, y2, __ldg(tab + 2 * ld + d) * (1.0 / 24.0));
    acc = fma(acc, y2, __ldg(tab + 1 * ld + d) * 0.5);
    acc = fma(acc, y2, __ldg(tab + d));
    return fma(2.0, acc, 1.0);
}
// zc[k] = 2 S_k(D) / (2k)!, k = 0 .. 4: the normaliser needs no table row (the k = 5 term is < 3e-17 of it)
__device__ __forceinline__ long long samplez_wide_table(double mean, double a, float sqrt_a, int dmax, double u,
                                                        const double* __restrict__ tab, const double* zc, double* zsum) {
    const int ld = dmax + 1;
    const double c0 = rint(mean);
    const double delta = c0 - mean;                   // in [-0.5, 0.5]
    const double w0 = exp_small(-a * delta * delta);  // rho(c0) = e^{-a delta^2}
    const double y = 2.0 * a * delta, y2 = y * y;
    const double Z = w0 * fma(fma(fma(fma(fma(zc[4], y2, zc[3]), y2, zc[2]), y2, zc[1]), y2, zc[0]), 1.0, 1.0);
    *zsum = Z;
    const double target = u * Z;
    const long long x = (long long)c0;
    if (target < w0) return x;
    // the exit distance: FP32 estimate from the symmetric integral, exact mass inside distance ds - 1 from the tables,
    // then the outward walk (the estimate only chooses where the walk starts)
    int d = 1;
    double cum = w0;
    const double* __restrict__ G = tab + kWideTabK * ld;
    double g0 = 0.0, g1 = 0.0, g2 = 0.0;
    int dg = 0;   // g0..g2 = G(dg .. dg + 2) when dg > 0
    {
        const float yy = (float)target * sqrt_a * 0.56418958354775628695f;
        if (yy < 0.99999f) {
            int ds = (int)floorf(erfinvf(yy) / sqrt_a - 0.5f) - 1;
            ds = ds > dmax ? dmax : ds;
            if (ds >= 3) {
                // the first weights of the walk are requested together with the row of masses
                if (ds + 2 <= dmax) { g0 = __ldg(G + ds); g1 = __ldg(G + ds + 1); g2 = __ldg(G + ds + 2); dg = ds; }
                const double base = w0 * wide_mass_tab(tab, ld, ds - 1, y2);
                if (base <= target) cum = base, d = ds;
            }
        }
    }
    for (; d <= dmax; d++) {
        const double gt = (dg > 0 && d == dg) ? g0 : (dg > 0 && d == dg + 1) ? g1 : (dg > 0 && d == dg + 2) ? g2 : __ldg(G + d);
        const double gd = gt * w0;
        const double yd = y * (double)d;              // 2 a d delta, |.| <= 0.1
        cum += gd * exp_small(-yd);                   // e^{-a (d + delta)^2}
        if (target < cum) return x + d;
        cum += gd * exp_small(yd);                    // e^{-a (d - delta)^2}
        if (target < cum) return x - d;
    }
    return x - dmax;   // u Z beyond the accumulated mass (rounding, ~1e-16): the last point visited
}

// ---------------------------------------------------------------------------------------------------------
// Fast SampleZ, smooth regime (kind 0 / 1: sigma_i >= 1, full window): the same inverse transform as samplez_fast
// (same ordering c0, c0+1, c0-1, ..., same one uniform), arranged for few instructions and short dependency chains:
//   * weights in units of rho(c0), so only ratios are needed: with a = 1/(2 sigma_i^2), delta = c0 - mean,
//       rho(c0 +- (d+1)) / rho(c0 +- d) = e^{-a} q^d E^{+-1},   E = e^{-2 a delta},  q = e^{-2a};
//   * |2 a delta| <= a <= 1/2 here, so E and 1/E come from the even and odd parts of one Taylor polynomial
//     (E = C + S, 1/E = C - S; truncation < 1e-16), with no range reduction, no division and no second exp;
//   * the target u Z / rho(c0) = u sigma sqrt(2 pi) (1 + eps cos 2 pi delta) e^{a delta^2}, the last factor a
//     degree-10 polynomial on [0, 1/8];
//   * one branch per step of the outward walk (both sides are added, the side is decided after the exit).
// log Z (for LGS_WEIGHT_EXACT) is lognz + eps cos(2 pi delta) to < 1e-16: no log call.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float cos2pi_half(float dl) {  // cos(2 pi dl), |dl| <= 1/2, abs. error ~1e-7
    const float s2 = dl * dl;
    float p = -2.1915353e-05f;
    p = fmaf(p, s2, 4.6630281e-04f);
    p = fmaf(p, s2, -7.3704309e-03f);
    p = fmaf(p, s2, 8.2145887e-02f);
    p = fmaf(p, s2, -0.59926453f);
    p = fmaf(p, s2, 2.5501640f);
    p = fmaf(p, s2, -5.1677128f);
    p = fmaf(p, s2, 3.14159265f);
    const float sn = p * dl;  // sin(pi dl)
    return fmaf(-2.0f * sn, sn, 1.0f);
}

// Taylor coefficients in the constant bank: FP64 instructions take them as operands directly (an FP64 immediate
// would cost two moves per use)
static __constant__ double kEvenC[8] = {1.0, 0.5, 0.041666666666666664, 0.001388888888888889, 2.48015873015873e-05,
                                        2.755731922398589e-07, 2.08767569878681e-09, 1.1470745597729725e-11};
static __constant__ double kOddC[7] = {1.0, 0.16666666666666666, 0.008333333333333333, 0.0001984126984126984,
                                       2.7557319223985893e-06, 2.505210838544172e-08, 1.6059043836821613e-10};
static __constant__ double kExpC[11] = {1.0, 1.0, 0.5, 0.16666666666666666, 0.041666666666666664,
                                        0.008333333333333333, 0.001388888888888889, 0.0001984126984126984,
                                        2.48015873015873e-05, 2.7557319223985893e-06, 2.755731922398589e-07};

template <bool WANT_LOGZ>
__device__ __forceinline__ double samplez_smooth(double mean, double a, double q, double sq, double znorm, double eps1,
                                                 double lognz, int kind, int dmax, double u, double* logz) {
    const double c0 = rint(mean);
    const double delta = c0 - mean;  // in [-1/2, 1/2]
    const double ad = a * delta;
    const double x = -2.0 * ad;      // E = e^x, |x| <= 1/2
    const double y = x * x;
    double C = kEvenC[7];
#pragma unroll
    for (int k = 6; k >= 0; k--) C = fma(C, y, kEvenC[k]);
    double Sx = kOddC[6];
#pragma unroll
    for (int k = 5; k >= 0; k--) Sx = fma(Sx, y, kOddC[k]);
    Sx *= x;
    const double g = ad * delta;     // a delta^2 in [0, 1/8]
    double G = kExpC[10];
#pragma unroll
    for (int k = 9; k >= 0; k--) G = fma(G, g, kExpC[k]);
    double t = u * znorm;
    if (WANT_LOGZ) *logz = lognz;
    if (kind == 1) {
        const double corr = eps1 * (double)cos2pi_half((float)delta);
        t = fma(t, corr, t);
        if (WANT_LOGZ) *logz = lognz + corr;
    }
    t *= G;
    double ru = sq * (C + Sx), rd = sq * (C - Sx);
    double vu = 1.0, vd = 1.0, cum = 1.0, cu = 2.0;
    int d = 0;
    if (!(t < 1.0)) {
        // outward walk; leaving at d == dmax without t < cum means u Z lies beyond the accumulated mass (rounding,
        // ~1e-16) and takes the last point visited, c0 - dmax, like samplez_fast
        do {
            d++;
            vu *= ru;
            vd *= rd;
            cu = cum + vu;
            cum = cu + vd;
            ru *= q;
            rd *= q;
        } while (!(t < cum) && d < dmax);
    }
    const double sd = (double)d;
    return (t < cu) ? c0 + sd : c0 - sd;   // d == 0: both are c0
}

// ---------------------------------------------------------------------------------------------------------
// Single-precision screen of samplez_smooth (chain kernel).  The same inverse transform evaluated in FP32 (24 bits of
// the uniform, ex2-based exponentials, FP32 walk); the draw is accepted only when the target is clear of the three
// thresholds that decide it (the cumulative sums before / inside / after the step the walk stopped at) by more than a
// bound on everything FP32 can have got wrong: tol = tol_r * cum + tol_a, tol_a covering the 29 dropped bits of the
// uniform, tol_r the rounding of the exponentials and of the walk.  Otherwise the caller recomputes the coordinate
// with samplez_smooth in FP64 (a few draws in 10^5), so the map u -> x is the FP64 one bit for bit.  Why: in the chain
// kernel the panel warps share the FP64 pipe with back-to-back DMMA, every FP64 instruction of a panel warp waits for
// a DMMA slot, and SampleZ was 3/4 of them.
// ---------------------------------------------------------------------------------------------------------
struct SmoothF32 {
    float a, sq, q, znorm, eps1, tol_r, tol_a, pad;
};

// dl = (float)(c0 - mean); uhi = the high word of the coordinate's uniform.  Returns true when certain; *off is the
// signed offset of the draw from c0 and *corr the theta-series correction eps cos(2 pi dl) (0 for kind 0).
__device__ __forceinline__ bool samplez_smooth_screen(float dl, const SmoothF32& k, int kind, int dmax, uint32_t uhi,
                                                      float* off, float* corr_out) {
    const float ad = k.a * dl;
    const float x = -2.0f * ad;
    const float E = __expf(x), Ei = __expf(-x);
    const float G = __expf(ad * dl);
    float t = (float)(uhi >> 8) * 5.9604644775390625e-8f * k.znorm;
    float corr = 0.0f;
    if (kind == 1) {
        corr = k.eps1 * cos2pi_half(dl);
        t = fmaf(t, corr, t);
    }
    *corr_out = corr;
    t *= G;
    float ru = k.sq * E, rd = k.sq * Ei;
    float vu = 1.0f, vd = 1.0f, cum = 1.0f, cu = 2.0f, cprev = 0.0f;
    int d = 0;
    if (!(t < 1.0f)) {
        do {
            d++;
            cprev = cum;
            vu *= ru;
            vd *= rd;
            cu = cum + vu;
            cum = cu + vd;
            ru *= k.q;
            rd *= k.q;
        } while (!(t < cum) && d < dmax);
    }
    const float tol = fmaf(k.tol_r, cum, k.tol_a);
    const float sd = (float)d;
    *off = (t < cu) ? sd : -sd;
    return (t < cum) && fabsf(t - cprev) > tol && fabsf(t - cu) > tol && fabsf(t - cum) > tol;
}

// ---------------------------------------------------------------------------------------------------------
// Single-precision screen of the wide truncated windows (kind 3; samplez_fast<WIDE>): sigma_i >= 50, the reference's
// table cut to max_range + 1 entries (klein.py:122-126).  The draw is the index d >= 1 with M(d - 1) <= u Z < M(d),
// M(d) = sum_{|j| <= d} exp(-a (j + delta)^2), and the side of c0 it falls on.  In FP32: Z = M(D) and M(d0 - 1) from the
// closed form (integral + end-point mean + the B2 term; the B4 term is < 1e-9 of the sum at a <= 2e-4), d0 from erfinv,
// then at most four outward steps with directly evaluated weights.  Error budget against the FP64 path, relative to Z:
// erff 2 ulp and three roundings per closed form (3.6e-7, twice), the 29 dropped bits of u (6e-8), <= 8 additions of
// the walk (5e-7) -- under 1.4e-6; a draw is accepted only when u Z clears every threshold it is compared with by
// tol = 4e-6 T + 1.2e-7 Z, otherwise the caller runs the FP64 draw (about 0.4 % of the draws on a flat window), so the map
// u -> x stays the FP64 one.  Constants (SmoothF32): a, sq = sqrt(a), q = sqrt(pi) / (2 sqrt(a)), znorm = sqrt(a / pi),
// eps1 = 1 / sqrt(a), tol_r = 4e-6.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool samplez_wide_screen(float dl, const SmoothF32& k, int dmax, uint32_t uhi, float* off) {
    const float a = k.a, s = k.sq, c1 = k.q;
    const float a6 = a * (1.0f / 6.0f);
    const float D = (float)dmax;
    float su = D + dl, sd = D - dl;
    float gu = __expf(-a * su * su), gd = __expf(-a * sd * sd);
    const float Z = fmaf(c1, erff(s * su) + erff(s * sd), fmaf(0.5f, gu + gd, -a6 * fmaf(su, gu, sd * gd)));
    const float T = (float)(uhi >> 8) * 5.9604644775390625e-8f * Z;
    // every FP32 error above is relative to the masses compared (T and the cumulative mass next to it), except the 29
    // dropped bits of u, which are 6e-8 of Z whatever T is
    const float tol = fmaf(k.tol_r, T, 1.2e-7f * Z);
    const float y = T * k.znorm;
    if (!(y < 0.9999f)) return false;
    int d0 = (int)floorf(fmaf(erfinvf(y), k.eps1, -0.5f)) - 1;
    d0 = d0 < 1 ? 1 : d0;
    const float e = (float)(d0 - 1);
    su = e + dl, sd = e - dl;
    gu = __expf(-a * su * su), gd = __expf(-a * sd * sd);
    float cum = fmaf(c1, erff(s * su) + erff(s * sd), fmaf(0.5f, gu + gd, -a6 * fmaf(su, gu, sd * gd)));
    if (!(T > cum + tol)) return false;   // at or below M(d0 - 1), or too close to call
#pragma unroll 1
    for (int i = 0; i < 4; i++) {
        const float d = (float)(d0 + i);
        if (d > D) return false;
        su = d + dl, sd = d - dl;
        cum += __expf(-a * su * su);
        if (T < cum) {
            *off = d;
            return T < cum - tol;
        }
        if (!(T > cum + tol)) return false;
        cum += __expf(-a * sd * sd);
        if (T < cum) {
            *off = -d;
            return T < cum - tol;
        }
        if (!(T > cum + tol)) return false;
    }
    return false;
}

// delta = rint(mean) - mean as a float WITHOUT FP64 arithmetic: the mantissa of `mean` is shifted to a fixed-point
// number with 32 fractional bits with integer instructions and the fraction is taken from there (absolute error
// < 2^-32 plus one FP32 rounding, inside the screen's tolerance).  Returns false for a fraction of exactly one half
// at that resolution (round-half-even is decided by the FP64 path) and for |mean| >= 2^31.
__device__ __forceinline__ bool rint_delta_f32(double mean, float* dl) {
    const uint32_t hi = (uint32_t)__double2hiint(mean), lo = (uint32_t)__double2loint(mean);
    const int e = (int)((hi >> 20) & 0x7ffu) - 1023;
    if (e > 30) return false;
    if (e < -30) {  // |mean| < 2^-30: rint(mean) = 0
        *dl = -(float)mean;
        return true;
    }
    const uint64_t M = ((uint64_t)((hi & 0xfffffu) | 0x100000u) << 32) | lo;  // 1.m * 2^52
    const int sh = e - 20;                                                    // a * 2^32 = M * 2^(e - 20)
    const uint64_t A = sh >= 0 ? (M << sh) : (M >> (-sh));
    const uint32_t F = (uint32_t)A;
    if (F == 0x80000000u) return false;
    // fraction below one half: rint = floor, delta_a = -F / 2^32; above: rint = floor + 1, delta_a = (2^32 - F) / 2^32
    const float d = (F < 0x80000000u) ? -(float)F : (float)(0u - F);
    const float ds = d * 2.3283064365386963e-10f;
    *dl = (hi >> 31) ? -ds : ds;
    return true;
}

}  // namespace lgs
