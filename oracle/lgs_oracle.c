/*
 * lgs_oracle.c -- CPU restatement of the reference Klein / IMHK sampling path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under lattice-gaussian-mcmc_b200/ links, loads or calls this file;
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks this restatement against golden vectors
 * produced by importing the reference itself (tests/golden/make_golden.py, run in the build container
 * where /root/reference is mounted): integer coefficient vectors bit-exact, lattice points exact,
 * log-densities / log-weights within 1e-12 relative.
 *
 * What it follows (paths relative to the reference tree):
 *   src/samplers/klein.py:101-141   _compute_1d_probabilities   -> table_build()
 *   src/samplers/klein.py:143-179   _sample_1d_discrete_gaussian -> samplez_table()
 *   src/samplers/klein.py:181-220   sample_single               -> klein_one()
 *   src/samplers/klein.py:222-271   compute_log_density         -> klein_logq_one()
 *   src/samplers/base.py:133-144    log_gaussian_weight         -> log_target()
 *   src/samplers/imhk.py:102-177    _compute_importance_weight, _initialize_chain, step
 *   src/samplers/imhk.py:196-229    sample(num_samples, thin)
 * plus the third-party arithmetic the reference calls and that is not under /root/reference:
 *   numpy 2.3.5  legacy RandomState: mt19937 seeding, random_sample(), choice(p=...) =
 *                cumsum -> /cdf[-1] -> searchsorted(side='right'); np.sum = pairwise summation
 *                (blocks of 128, 8 accumulators); np.round = round-half-even
 *   scipy 1.18.1 special.logsumexp: max element(s) split out, s = sum(exp(a - a_max)) / m,
 *                log1p(s) + log(m) + a_max
 * Differences that cannot be removed: numpy's AVX512 exp/log1p differ from glibc's in the last bit for
 * a few percent of arguments; that moves CDF entries by <= 1 ulp and flips a drawn index with
 * probability ~1e-15 per draw.  The harness counts mismatches instead of assuming zero.
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define LGS_MAX_RANGE 1000 /* klein.py:122 */

/* ------------------------------------------------------------------ numpy legacy MT19937 ---------- */

typedef struct {
    uint32_t key[624];
    int pos;
} mt_state;

static void mt_seed(mt_state *st, uint32_t seed) {
    for (int i = 0; i < 624; i++) {
        st->key[i] = seed;
        seed = 1812433253u * (seed ^ (seed >> 30)) + (uint32_t)i + 1u;
    }
    st->pos = 624;
}

static void mt_refill(mt_state *st) {
    uint32_t *k = st->key;
    for (int i = 0; i < 624; i++) {
        uint32_t y = (k[i] & 0x80000000u) | (k[(i + 1) % 624] & 0x7fffffffu);
        k[i] = k[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    st->pos = 0;
}

static uint32_t mt_next32(mt_state *st) {
    if (st->pos >= 624) mt_refill(st);
    uint32_t y = st->key[st->pos++];
    y ^= y >> 11;
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= y >> 18;
    return y;
}

static double mt_next_double(mt_state *st) {
    uint32_t a = mt_next32(st) >> 5, b = mt_next32(st) >> 6;
    return (a * 67108864.0 + b) / 9007199254740992.0;
}

/* np.random.RandomState(seed).random_sample(count) for a 32-bit integer seed */
void lgs_oracle_uniform_stream(uint32_t seed, int64_t count, double *out) {
    mt_state st;
    mt_seed(&st, seed);
    for (int64_t i = 0; i < count; i++) out[i] = mt_next_double(&st);
}

/* ------------------------------------------------------------------ numpy reductions -------------- */

/* numpy's pairwise summation of a contiguous double vector (the order np.sum uses) */
static double np_pairwise_sum(const double *a, int64_t n) {
    if (n < 8) {
        double res = 0.0;
        for (int64_t i = 0; i < n; i++) res += a[i];
        return res;
    } else if (n <= 128) {
        double r[8];
        int64_t i;
        for (i = 0; i < 8; i++) r[i] = a[i];
        for (i = 8; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; j++) r[j] += a[i + j];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; i++) res += a[i];
        return res;
    } else {
        int64_t n2 = n / 2;
        n2 -= n2 % 8;
        return np_pairwise_sum(a, n2) + np_pairwise_sum(a + n2, n - n2);
    }
}

double lgs_oracle_np_sum(const double *a, int64_t n) { return np_pairwise_sum(a, n); }

/* ------------------------------------------------------------------ the 1-D table sampler --------- */

typedef struct {
    int64_t lower;  /* support[0] */
    int len;        /* number of support points, <= LGS_MAX_RANGE + 1 */
    double logp[LGS_MAX_RANGE + 2];
    double work[LGS_MAX_RANGE + 2];
} table_t;

/* klein.py:101-141 with use_log_space=True, and scipy's logsumexp */
static void table_build(table_t *tb, double mean, double sigma, int precision) {
    double range_factor = (sigma < 0.1) ? (double)(precision > 3 ? precision : 3) : (double)precision;
    int64_t lower = (int64_t)floor(mean - range_factor * sigma);
    int64_t upper = (int64_t)ceil(mean + range_factor * sigma);
    if (upper - lower > LGS_MAX_RANGE) {
        int64_t c = (int64_t)nearbyint(mean); /* np.round: half to even */
        lower = c - LGS_MAX_RANGE / 2;
        upper = c + LGS_MAX_RANGE / 2;
    }
    int len = (int)(upper - lower + 1);
    tb->lower = lower;
    tb->len = len;
    double a_max = -INFINITY;
    for (int k = 0; k < len; k++) {
        double z = ((double)(lower + k) - mean) / sigma;
        double lp = -0.5 * (z * z);
        tb->logp[k] = lp;
        if (lp > a_max) a_max = lp;
    }
    /* scipy logsumexp: elements equal to the max are removed from the sum and counted in m */
    double m = 0.0;
    for (int k = 0; k < len; k++) {
        if (tb->logp[k] == a_max) {
            m += 1.0;
            tb->work[k] = 0.0; /* exp(-inf) */
        } else {
            tb->work[k] = exp(tb->logp[k] - a_max);
        }
    }
    double s = np_pairwise_sum(tb->work, len);
    if (s != 0.0) s = s / m;
    double lse = log1p(s) + log(m) + a_max;
    if (!isfinite(lse)) {
        /* scipy falls back to log(sum(exp(a))) when the shifted result is not finite */
        for (int k = 0; k < len; k++) tb->work[k] = exp(tb->logp[k]);
        lse = log(np_pairwise_sum(tb->work, len));
    }
    for (int k = 0; k < len; k++) tb->logp[k] = tb->logp[k] - lse;
}

/* logsumexp of an already-normalised table: the `log_Z` of klein.py:265-269 */
static double table_log_z(table_t *tb) {
    int len = tb->len;
    double a_max = -INFINITY;
    for (int k = 0; k < len; k++)
        if (tb->logp[k] > a_max) a_max = tb->logp[k];
    double m = 0.0;
    for (int k = 0; k < len; k++) {
        if (tb->logp[k] == a_max) {
            m += 1.0;
            tb->work[k] = 0.0;
        } else {
            tb->work[k] = exp(tb->logp[k] - a_max);
        }
    }
    double s = np_pairwise_sum(tb->work, len);
    if (s != 0.0) s = s / m;
    double out = log1p(s) + log(m) + a_max;
    if (!isfinite(out)) {
        for (int k = 0; k < len; k++) tb->work[k] = exp(tb->logp[k]);
        out = log(np_pairwise_sum(tb->work, len));
    }
    return out;
}

/*
 * klein.py:143-179 + numpy RandomState.choice(support, p=probs).
 * Returns 1 and stores the draw when a uniform was consumed, 0 when the reference's
 * "Sampling fallback" (ValueError -> round(mean), no uniform consumed) would trigger.
 */
static int samplez_table(table_t *tb, double mean, double sigma, int precision, double u, int64_t *out) {
    table_build(tb, mean, sigma, precision);
    int len = tb->len;
    double *p = tb->work;
    for (int k = 0; k < len; k++) p[k] = exp(tb->logp[k]);
    double tot = np_pairwise_sum(p, len);
    int bad = 0;
    for (int k = 0; k < len; k++) {
        p[k] = p[k] / tot;
        if (!(p[k] >= 0.0)) bad = 1; /* NaN or negative: choice() raises ValueError */
    }
    if (bad) {
        *out = (int64_t)nearbyint(mean);
        return 0;
    }
    /* cdf = p.cumsum(); cdf /= cdf[-1]; idx = searchsorted(cdf, u, side='right') */
    double run = p[0];
    for (int k = 1; k < len; k++) {
        run = run + p[k];
        p[k] = run;
    }
    double last = p[len - 1];
    int lo = 0, hi = len; /* first index with cdf[idx] > u */
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (p[mid] / last <= u) lo = mid + 1; else hi = mid;
    }
    if (lo >= len) lo = len - 1; /* unreachable for u < 1 */
    *out = tb->lower + lo;
    return 1;
}

/* standalone entry used by the unit tests of the 1-D sampler */
int lgs_oracle_samplez(double mean, double sigma, int precision, double u, int64_t *out) {
    table_t *tb = (table_t *)malloc(sizeof(table_t));
    int used = samplez_table(tb, mean, sigma, precision, u, out);
    free(tb);
    return used;
}

/* table export for tests: writes support lower bound, length and normalised log-probabilities */
int lgs_oracle_table(double mean, double sigma, int precision, int64_t *lower, double *logp_out) {
    table_t *tb = (table_t *)malloc(sizeof(table_t));
    table_build(tb, mean, sigma, precision);
    *lower = tb->lower;
    int len = tb->len;
    memcpy(logp_out, tb->logp, sizeof(double) * len);
    free(tb);
    return len;
}

/* ------------------------------------------------------------------ Klein ------------------------- */

typedef struct {
    int n;
    const double *R;  /* row-major n x n, upper triangular, positive diagonal (klein.py:56-79) */
    const double *B;  /* row-major n x n, columns are basis vectors (klein.py:59,218) */
    const double *t;  /* Q^T center */
    const double *c;  /* center */
    double sigma;
    int precision;
} ctx_t;

/* number of uniforms one sample_single() consumes: coordinates with sigma_i >= 1e-10 (klein.py:201) */
static int uniforms_per_sample(const ctx_t *cx) {
    int cnt = 0;
    for (int i = 0; i < cx->n; i++)
        if (!(cx->sigma / fabs(cx->R[(size_t)i * cx->n + i]) < 1e-10)) cnt++;
    return cnt;
}

int lgs_oracle_uniforms_per_sample(int n, const double *R, double sigma) {
    ctx_t cx = {n, R, 0, 0, 0, sigma, 10};
    return uniforms_per_sample(&cx);
}

/* klein.py:181-220.  `u` points at this sample's uniforms; returns how many were consumed. */
static int klein_one(const ctx_t *cx, table_t *tb, const double *u, int64_t *x, double *v) {
    int n = cx->n, used = 0;
    for (int i = n - 1; i >= 0; i--) {
        const double *Ri = cx->R + (size_t)i * n;
        double csum = 0.0;
        for (int j = i + 1; j < n; j++) csum += Ri[j] * (double)x[j];
        double rii = Ri[i];
        double mean = (cx->t[i] - csum) / rii;
        double sg = cx->sigma / fabs(rii);
        if (sg < 1e-10) {
            x[i] = (int64_t)nearbyint(mean);
        } else {
            if (sg > 1e10) sg = fmin(sg, 1e6);
            used += samplez_table(tb, mean, sg, cx->precision, u[used], &x[i]);
        }
    }
    if (v) {
        /* lattice_point = basis @ x */
        for (int r = 0; r < n; r++) {
            const double *Br = cx->B + (size_t)r * n;
            double acc = 0.0;
            for (int j = 0; j < n; j++) acc += Br[j] * (double)x[j];
            v[r] = acc;
        }
    }
    return used;
}

/*
 * klein.py:222-271 for a point already known by its integer coefficients (the reference recovers them
 * with an O(n^3) solve + round, klein.py:230-231).  `exact_logz` = 1 evaluates the reference's log_Z term
 * (a logsumexp of an already normalised table, ~1e-16) instead of taking it as 0.
 */
static double klein_logq_one(const ctx_t *cx, table_t *tb, const int64_t *x, int exact_logz) {
    int n = cx->n;
    const double log_2pi = log(2.0 * M_PI);
    double lp = 0.0;
    for (int i = n - 1; i >= 0; i--) {
        const double *Ri = cx->R + (size_t)i * n;
        double csum = 0.0;
        for (int j = i + 1; j < n; j++) csum += Ri[j] * (double)x[j];
        double rii = Ri[i];
        double mean = (cx->t[i] - csum) / rii;
        double sg = cx->sigma / fabs(rii);
        double z = ((double)x[i] - mean) / sg;
        lp += -0.5 * (z * z);
        lp -= 0.5 * log_2pi + log(sg);
        if (exact_logz) {
            table_build(tb, mean, sg, cx->precision);
            lp -= table_log_z(tb);
        }
    }
    return lp;
}

/* base.py:133-144 */
static double log_target(const ctx_t *cx, const double *v) {
    double d2 = 0.0;
    for (int i = 0; i < cx->n; i++) {
        double d = v[i] - cx->c[i];
        d2 += d * d;
    }
    return -d2 / (2.0 * (cx->sigma * cx->sigma));
}

/*
 * Conditional-law check of sampler OUTPUT (no uniforms involved; checker for native-RNG samples).
 * For every sample x (integer coefficients, as returned by the sampler under test) and every coordinate i the
 * reference's conditional mean m_i is recomputed in the reference's order (klein.py:189-195) and the reference's own
 * table for (m_i, sigma_i) is built (klein.py:101-141, 160-170: probs = exp(log_p); probs /= probs.sum()).  Under the
 * reference's law the offset x_i - round(m_i) given m_i is distributed by that table, whatever the other coordinates
 * did.  Per coordinate i, per bin f of frac(m_i) = m_i - floor(m_i) (FB equal bins) and per clipped offset
 * k in [-K, K] (|offset| >= K pooled into the end bins) this accumulates
 *     obs[i][f][k + K]  += 1
 *     expc[i][f][k + K] += table probability of the offsets that clip to k
 * so that the caller can form sum (obs - expc)^2 / expc per coordinate (a sum of independent, non-identically
 * distributed multinomial draws: Var(obs) <= expc, i.e. the chi-square reference distribution is conservative).
 * Coordinates with sigma_i < 1e-10 (deterministic, klein.py:201-204) are skipped.  `width` (int32[n] or NULL = all 1)
 * coarsens the offsets of coordinate i into bins of width[i] integers (offset -> nearest multiple) before the clip, for
 * the wide truncated windows (sigma_i >= 50: offsets spread over +-500).
 */
static int64_t bin_offset(int64_t off, int64_t w, int K) {
    int64_t k = off >= 0 ? (off + w / 2) / w : -((-off + w / 2) / w);
    if (k > K) k = K;
    if (k < -K) k = -K;
    return k;
}

int lgs_oracle_conditional_counts(int n, const double *R, const double *B, const double *t, const double *center,
                                  double sigma, int precision, int64_t S, const int64_t *X, int K, int FB,
                                  const int32_t *width, int64_t *obs, double *expc, int threads) {
    ctx_t cx = {n, R, B, t, center, sigma, precision};
    const int W = 2 * K + 1;
    const size_t cells = (size_t)n * FB * W;
    memset(obs, 0, sizeof(int64_t) * cells);
    for (size_t c = 0; c < cells; c++) expc[c] = 0.0;
    int err = 0;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel
#endif
    {
        table_t *tb = (table_t *)malloc(sizeof(table_t));
        int64_t *lo = (int64_t *)calloc(cells, sizeof(int64_t));
        double *le = (double *)calloc(cells, sizeof(double));
        if (!tb || !lo || !le) {
#ifdef _OPENMP
#pragma omp atomic write
#endif
            err = 1;
        } else {
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 16)
#endif
            for (int64_t s = 0; s < S; s++) {
                const int64_t *x = X + (size_t)s * n;
                for (int i = n - 1; i >= 0; i--) {
                    const double *Ri = cx.R + (size_t)i * n;
                    double csum = 0.0;
                    for (int j = i + 1; j < n; j++) csum += Ri[j] * (double)x[j];
                    double rii = Ri[i];
                    double mean = (cx.t[i] - csum) / rii;
                    double sg = cx.sigma / fabs(rii);
                    if (sg < 1e-10) continue;
                    if (sg > 1e10) sg = fmin(sg, 1e6);
                    table_build(tb, mean, sg, cx.precision);
                    int len = tb->len;
                    double *p = tb->work;
                    for (int k = 0; k < len; k++) p[k] = exp(tb->logp[k]);
                    double tot = np_pairwise_sum(p, len);
                    int64_t c0 = (int64_t)nearbyint(mean);
                    double fr = mean - floor(mean);
                    int f = (int)(fr * FB);
                    if (f >= FB) f = FB - 1;
                    if (f < 0) f = 0;
                    size_t base = ((size_t)i * FB + f) * W;
                    const int64_t w = width && width[i] > 1 ? width[i] : 1;
                    lo[base + (size_t)(bin_offset(x[i] - c0, w, K) + K)] += 1;
                    for (int k = 0; k < len; k++)
                        le[base + (size_t)(bin_offset(tb->lower + k - c0, w, K) + K)] += p[k] / tot;
                }
            }
#ifdef _OPENMP
#pragma omp critical
#endif
            {
                for (size_t c = 0; c < cells; c++) {
                    obs[c] += lo[c];
                    expc[c] += le[c];
                }
            }
        }
        free(tb);
        free(lo);
        free(le);
    }
    return err ? -1 : 0;
}

/*
 * Replay of KleinSampler.sample(S): sample s consumes U[s*ups .. (s+1)*ups) with ups = uniforms per sample.
 * X int64[S*n], V double[S*n] or NULL, logq double[S] or NULL (Klein log-density, klein.py:222-271).
 * Returns 0, or -1 if a draw fell into the no-uniform fallback (stream layout then undefined).
 */
int lgs_oracle_klein_replay(int n, const double *R, const double *B, const double *t, const double *center,
                            double sigma, int precision, int64_t S, const double *U, int64_t *X, double *V,
                            double *logq, int exact_logz, int threads) {
    ctx_t cx = {n, R, B, t, center, sigma, precision};
    int ups = uniforms_per_sample(&cx);
    int rc = 0;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel
#endif
    {
        table_t *tb = (table_t *)malloc(sizeof(table_t));
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
        for (int64_t s = 0; s < S; s++) {
            int used = klein_one(&cx, tb, U + s * ups, X + s * n, V ? V + s * n : 0);
            if (used != ups) rc = -1;
            if (logq) logq[s] = klein_logq_one(&cx, tb, X + s * n, exact_logz);
        }
        free(tb);
    }
    return rc;
}

/*
 * KleinSampler.sample(S) after np.random.seed(seed) (single stream, serial -- the reference's behaviour),
 * or, with chains > 1, the reference's multi-process pattern: worker k reseeds with seed + 1000*k and draws
 * S/chains samples (experiments/dimension_scaling.py:877-884).  Used as the timed CPU baseline.
 */
int lgs_oracle_klein_seeded(int n, const double *R, const double *B, const double *t, const double *center,
                            double sigma, int precision, int64_t S, uint32_t seed, int chains, int64_t *X,
                            double *V, int threads) {
    ctx_t cx = {n, R, B, t, center, sigma, precision};
    if (chains < 1) chains = 1;
    int64_t per = (S + chains - 1) / chains;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel
#endif
    {
        table_t *tb = (table_t *)malloc(sizeof(table_t));
        double *u = (double *)malloc(sizeof(double) * n);
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
        for (int k = 0; k < chains; k++) {
            mt_state st;
            mt_seed(&st, seed + 1000u * (uint32_t)k);
            int64_t lo = k * per, hi = lo + per < S ? lo + per : S;
            for (int64_t s = lo; s < hi; s++) {
                /* uniforms are drawn lazily by the reference; count first, then draw exactly that many */
                int ups = uniforms_per_sample(&cx);
                for (int i = 0; i < ups; i++) u[i] = mt_next_double(&st);
                klein_one(&cx, tb, u, X + s * n, V ? V + s * n : 0);
            }
        }
        free(u);
        free(tb);
    }
    return 0;
}

/* ------------------------------------------------------------------ IMHK -------------------------- */

/*
 * Replay of `chains` independent IMHKSampler objects, each doing _initialize_chain() then `steps` step()s
 * (imhk.py:126-177).  Per chain the uniform stream is [ups init] + steps * [ups proposal, 1 accept]
 * (SURVEY.md appendix A.4).  Outputs, all optional except X_state:
 *   X_keep  int64[chains * n_keep * n]   coefficient vector of the chain state after every `thin`-th step
 *   V_keep  double[chains * n_keep * n]  the same states as lattice points
 *   accept  uint8[chains * steps]        accept bit of every step
 *   logw    double[chains * (steps+1)]   log importance weight of the initial state and of every proposal
 * n_keep = steps / thin.  This is IMHKSampler.sample(n_keep, thin) on a fresh sampler.
 */
int lgs_oracle_imhk_replay(int n, const double *R, const double *B, const double *t, const double *center,
                           double sigma, int precision, int64_t chains, int64_t steps, int64_t thin,
                           const double *U, int64_t *X_keep, double *V_keep, uint8_t *accept, double *logw,
                           int64_t *accepted_total, int exact_logz, int threads) {
    ctx_t cx = {n, R, B, t, center, sigma, precision};
    int ups = uniforms_per_sample(&cx);
    int64_t per_chain = ups + steps * (int64_t)(ups + 1);
    int64_t n_keep = thin > 0 ? steps / thin : 0;
    int rc = 0;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel
#endif
    {
        table_t *tb = (table_t *)malloc(sizeof(table_t));
        int64_t *xc = (int64_t *)malloc(sizeof(int64_t) * n), *xp = (int64_t *)malloc(sizeof(int64_t) * n);
        double *vc = (double *)malloc(sizeof(double) * n), *vp = (double *)malloc(sizeof(double) * n);
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
        for (int64_t ch = 0; ch < chains; ch++) {
            const double *u = U + ch * per_chain;
            int64_t pos = 0, acc_count = 0;
            /* _initialize_chain */
            int used = klein_one(&cx, tb, u + pos, xc, vc);
            if (used != ups) rc = -1;
            pos += used;
            double lw_cur = log_target(&cx, vc) - klein_logq_one(&cx, tb, xc, exact_logz);
            if (logw) logw[ch * (steps + 1)] = lw_cur;
            for (int64_t st = 0; st < steps; st++) {
                used = klein_one(&cx, tb, u + pos, xp, vp);
                if (used != ups) rc = -1;
                pos += used;
                double lw_prop = log_target(&cx, vp) - klein_logq_one(&cx, tb, xp, exact_logz);
                double ratio;
                if (lw_cur == -INFINITY) {
                    ratio = 1.0;
                } else {
                    ratio = fmin(1.0, exp(lw_prop - lw_cur));
                }
                double ua = u[pos++];
                int acc = ua < ratio;
                if (acc) {
                    memcpy(xc, xp, sizeof(int64_t) * n);
                    memcpy(vc, vp, sizeof(double) * n);
                    lw_cur = lw_prop;
                    acc_count++;
                }
                if (accept) accept[ch * steps + st] = (uint8_t)acc;
                if (logw) logw[ch * (steps + 1) + st + 1] = lw_prop;
                if (thin > 0 && (st + 1) % thin == 0) {
                    int64_t kidx = (st + 1) / thin - 1;
                    if (X_keep) memcpy(X_keep + (ch * n_keep + kidx) * n, xc, sizeof(int64_t) * n);
                    if (V_keep) memcpy(V_keep + (ch * n_keep + kidx) * n, vc, sizeof(double) * n);
                }
            }
            if (accepted_total) accepted_total[ch] = acc_count;
        }
        free(tb); free(xc); free(xp); free(vc); free(vp);
    }
    return rc;
}

/* IMHK driven by the oracle's own MT19937 (chain k seeded seed + 1000*k): the timed CPU baseline. */
int lgs_oracle_imhk_seeded(int n, const double *R, const double *B, const double *t, const double *center,
                           double sigma, int precision, int64_t chains, int64_t steps, uint32_t seed,
                           int64_t *X_last, int64_t *accepted_total, int exact_logz, int threads) {
    ctx_t cx = {n, R, B, t, center, sigma, precision};
    int ups = uniforms_per_sample(&cx);
    int64_t per_chain = ups + steps * (int64_t)(ups + 1);
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel
#endif
    {
        double *u = (double *)malloc(sizeof(double) * per_chain);
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
        for (int64_t ch = 0; ch < chains; ch++) {
            mt_state st;
            mt_seed(&st, seed + 1000u * (uint32_t)ch);
            for (int64_t i = 0; i < per_chain; i++) u[i] = mt_next_double(&st);
            /* one chain, kept state = last state */
            lgs_oracle_imhk_replay(n, R, B, t, center, sigma, precision, 1, steps, steps > 0 ? steps : 1, u,
                                   X_last ? X_last + ch * n : 0, 0, 0, 0,
                                   accepted_total ? accepted_total + ch : 0, exact_logz, -1);
        }
        free(u);
    }
    return 0;
}
