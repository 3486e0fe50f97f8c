// lgs_reduce.cpp -- host-side lattice basis reduction behind the C ABI (include/lgs_b200.h, "basis reduction").
//
// the step BEFORE the sampling path.  The reference reduces through Sage's Matrix.LLL() /
// Matrix.BKZ() (fplll; src/lattices/reduction.py:103,275, klein_scaling_analysis.py:95-108), which is neither vendored
// nor installable here.  This file is an own implementation of the published algorithms, host C++ only:
//   * LLL in the L2 form of Nguyen-Stehle: the basis stays in exact integers, every inner product that feeds the
//     Gram-Schmidt data is computed EXACTLY (double FMA when the operands provably fit 53 bits, __int128 otherwise),
//     only the Cholesky recurrence r_kj = <b_k,b_j> - sum_i mu_ji r_ki runs in floating point (double, x87 long
//     double or double-double, tried in that order; a tier that fails hands its exact basis to the next), and size
//     reduction is iterated until the freshly recomputed mu_kj are all <= eta.
//   * BKZ-beta of Schnorr-Euchner: full (or linearly pruned) zig-zag enumeration of every projected block, insertion
//     of a shorter vector by a unimodular (Euclid) transformation of the block, LLL from the insertion point.
// Rows are basis vectors (the reference's Sage convention for reduction); the sampler-side column convention is
// applied by the Python caller exactly as the reference's scripts do.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/lgs_b200.h"

namespace {

using i64 = int64_t;
using i128 = __int128;

constexpr double kTwo53 = 9007199254740992.0;
constexpr int64_t kMaxEntry = (int64_t)1 << 56;  // basis entries beyond this are refused (LGS_E_RANGE)

// double-double (about 106 bits): the third precision tier, and the arithmetic of the final reducedness check
struct dd {
    double hi, lo;
    dd() : hi(0), lo(0) {}
    dd(double h) : hi(h), lo(0) {}
    dd(double h, double l) : hi(h), lo(l) {}
};
inline dd quick_two_sum(double a, double b) { const double s = a + b; return dd(s, b - (s - a)); }
inline dd two_sum(double a, double b) {
    const double s = a + b, bb = s - a;
    return dd(s, (a - (s - bb)) + (b - bb));
}
inline dd operator+(const dd& a, const dd& b) {
    dd s = two_sum(a.hi, b.hi);
    const dd t = two_sum(a.lo, b.lo);
    s.lo += t.hi;
    s = quick_two_sum(s.hi, s.lo);
    s.lo += t.lo;
    return quick_two_sum(s.hi, s.lo);
}
inline dd operator-(const dd& a) { return dd(-a.hi, -a.lo); }
inline dd operator-(const dd& a, const dd& b) { return a + (-b); }
inline dd operator*(const dd& a, const dd& b) {
    const double p = a.hi * b.hi;
    double e = std::fma(a.hi, b.hi, -p);
    e += a.hi * b.lo + a.lo * b.hi;
    return quick_two_sum(p, e);
}
inline dd operator/(const dd& a, const dd& b) {
    const double q1 = a.hi / b.hi;
    dd rem = a - b * dd(q1);
    const double q2 = rem.hi / b.hi;
    rem = rem - b * dd(q2);
    const double q3 = rem.hi / b.hi;
    return quick_two_sum(q1, q2) + dd(q3);
}
inline dd& operator+=(dd& a, const dd& b) { return a = a + b; }
inline dd& operator-=(dd& a, const dd& b) { return a = a - b; }
inline bool operator<=(const dd& a, const dd& b) { return a.hi < b.hi || (a.hi == b.hi && a.lo <= b.lo); }
inline bool operator<(const dd& a, const dd& b) { return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo); }
inline bool operator>(const dd& a, const dd& b) { return b < a; }

// the few scalar helpers the reducer needs, for each float type
inline double to_double(double x) { return x; }
inline double to_double(long double x) { return (double)x; }
inline double to_double(const dd& x) { return x.hi + x.lo; }
inline double round_ft(double x) { return std::nearbyint(x); }
inline long double round_ft(long double x) { return std::nearbyint(x); }
inline dd round_ft(const dd& x) {
    const double rh = std::nearbyint(x.hi);
    if (rh != x.hi) return dd(rh);                     // the fraction lives in hi
    return quick_two_sum(rh, std::nearbyint(x.lo));    // hi is an integer: round the tail
}
inline int64_t to_i64(double x) { return (int64_t)x; }
inline int64_t to_i64(long double x) { return (int64_t)x; }
inline int64_t to_i64(const dd& x) { return (int64_t)x.hi + (int64_t)x.lo; }
template <class FT> inline FT from_i128(__int128 s) { return (FT)(long double)s; }
template <> inline dd from_i128<dd>(__int128 s) {
    const double h = (double)s;
    return dd(h, (double)(s - (__int128)h));
}

template <class FT>
struct Reducer {
    int n, m;
    std::vector<i64> b;      // n x m, rows = basis vectors (exact)
    std::vector<double> bd;  // the same entries as doubles (valid per row while rowmax < 2^53)
    std::vector<i64> rowmax; // max |entry| per row
    std::vector<i64> u;      // n x n unimodular transform (rows follow the basis rows); empty when not tracked
    std::vector<FT> r, mu;   // n x n, lower triangles: r[k][j] = <b_k, b*_j>, mu[k][j] = r[k][j] / r[j][j]
    std::vector<FT> g, tmp;  // scratch: exact Gram row, working copy of mu_k
    std::vector<i64> X;
    int valid = 0;           // rows [0, valid) carry up-to-date r / mu
    int status = LGS_OK;
    i64 swaps = 0, passes = 0, insertions = 0, enum_nodes = 0;
    double delta, eta;

    Reducer(const i64* rows, int n_, int m_, const i64* u0, double delta_, double eta_)
        : n(n_), m(m_), b(rows, rows + (size_t)n_ * m_), bd((size_t)n_ * m_), rowmax(n_), r((size_t)n_ * n_),
          mu((size_t)n_ * n_), g(n_), tmp(n_), X(n_), delta(delta_), eta(eta_) {
        if (u0) u.assign(u0, u0 + (size_t)n * n);
        for (int i = 0; i < n; i++) refresh_row(i);
    }

    void refresh_row(int i) {
        i64 mx = 0;
        const i64* p = &b[(size_t)i * m];
        double* q = &bd[(size_t)i * m];
        for (int c = 0; c < m; c++) {
            const i64 a = p[c] < 0 ? -p[c] : p[c];
            if (a < 0 || a >= kMaxEntry) { status = LGS_E_RANGE; return; }
            mx = std::max(mx, a);
            q[c] = (double)p[c];
        }
        rowmax[i] = mx;
    }

    // exact <b_i, b_j>, rounded once to FT
    FT dot(int i, int j) const {
        const long double bound = (long double)rowmax[i] * (long double)rowmax[j] * (long double)m;
        if (bound < (long double)kTwo53) {  // every product and every partial sum is an exact double
            const double* p = &bd[(size_t)i * m];
            const double* q = &bd[(size_t)j * m];
            double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
            int c = 0;
            for (; c + 4 <= m; c += 4) {
                s0 += p[c] * q[c];
                s1 += p[c + 1] * q[c + 1];
                s2 += p[c + 2] * q[c + 2];
                s3 += p[c + 3] * q[c + 3];
            }
            for (; c < m; c++) s0 += p[c] * q[c];
            return FT((s0 + s1) + (s2 + s3));
        }
        const i64* p = &b[(size_t)i * m];
        const i64* q = &b[(size_t)j * m];
        i128 s = 0;  // |entries| < 2^56 (kMaxEntry, enforced on input and by row_axpy): m < 2^15 products cannot overflow
        for (int c = 0; c < m; c++) s += (i128)p[c] * q[c];
        return from_i128<FT>(s);
    }

    // Gram-Schmidt data of row k from rows < k (which must be valid)
    void gso_row(int k) {
        for (int j = 0; j <= k; j++) g[j] = dot(k, j);
        FT* rk = &r[(size_t)k * n];
        FT* mk = &mu[(size_t)k * n];
        for (int j = 0; j < k; j++) {
            const FT* mj = &mu[(size_t)j * n];
            FT a0 = 0, a1 = 0, a2 = 0, a3 = 0;
            int i = 0;
            for (; i + 4 <= j; i += 4) {
                a0 += mj[i] * rk[i];
                a1 += mj[i + 1] * rk[i + 1];
                a2 += mj[i + 2] * rk[i + 2];
                a3 += mj[i + 3] * rk[i + 3];
            }
            for (; i < j; i++) a0 += mj[i] * rk[i];
            rk[j] = g[j] - ((a0 + a1) + (a2 + a3));
            mk[j] = rk[j] / r[(size_t)j * n + j];
        }
        FT a = 0;
        for (int i = 0; i < k; i++) a += mk[i] * rk[i];
        rk[k] = g[k] - a;
    }

    // b_k -= x * b_j (and the same on U), with overflow detection
    void row_axpy(int k, int j, i64 x) {
        const long double worst = (long double)rowmax[k] + fabsl((long double)x) * (long double)rowmax[j];
        if (worst >= (long double)kMaxEntry) { status = LGS_E_RANGE; return; }
        i64* p = &b[(size_t)k * m];
        const i64* q = &b[(size_t)j * m];
        for (int c = 0; c < m; c++) p[c] -= x * q[c];
        if (!u.empty()) {
            i64* pu = &u[(size_t)k * n];
            const i64* qu = &u[(size_t)j * n];
            for (int c = 0; c < n; c++) {
                i64 t;
                if (__builtin_mul_overflow(x, qu[c], &t) || __builtin_sub_overflow(pu[c], t, &pu[c])) {
                    status = LGS_E_RANGE;
                    return;
                }
            }
        }
    }

    void swap_rows(int a, int c) {
        std::swap_ranges(&b[(size_t)a * m], &b[(size_t)a * m] + m, &b[(size_t)c * m]);
        std::swap_ranges(&bd[(size_t)a * m], &bd[(size_t)a * m] + m, &bd[(size_t)c * m]);
        std::swap(rowmax[a], rowmax[c]);
        if (!u.empty()) std::swap_ranges(&u[(size_t)a * n], &u[(size_t)a * n] + n, &u[(size_t)c * n]);
    }

    // move row `from` up to position `to` (to <= from), shifting the rows between down by one
    void rotate_up(int to, int from) {
        for (int i = from; i > to; i--) swap_rows(i, i - 1);
    }

    // iterated Babai size reduction of row k against rows < k; leaves r / mu of row k fresh
    void size_reduce(int k) {
        for (int iter = 0;; iter++) {
            if (iter > 200) { status = LGS_E_PRECISION; return; }
            gso_row(k);
            passes++;
            const FT* mk = &mu[(size_t)k * n];
            for (int j = 0; j < k; j++) tmp[j] = mk[j];
            bool any = false;
            for (int j = k - 1; j >= 0; j--) {
                X[j] = 0;
                if (std::fabs(to_double(tmp[j])) > eta) {
                    const FT xr = round_ft(tmp[j]);
                    if (std::fabs(to_double(xr)) > 4.0e18) { status = LGS_E_RANGE; return; }
                    X[j] = to_i64(xr);
                    const FT* mj = &mu[(size_t)j * n];
                    for (int i = 0; i < j; i++) tmp[i] -= xr * mj[i];
                    any = true;
                }
            }
            if (!any) return;
            for (int j = 0; j < k; j++)
                if (X[j]) {
                    row_axpy(k, j, X[j]);
                    if (status) return;
                }
            refresh_row(k);
            if (status) return;
        }
    }

    // LLL on rows [0, kend), assuming rows [0, min(kstart, valid)) are already reduced among themselves
    void lll(int kstart, int kend) {
        if (n == 0) return;
        if (valid == 0) {
            gso_row(0);
            valid = 1;
        }
        int k = std::max(1, std::min(kstart, valid));
        while (k < kend && status == LGS_OK) {
            size_reduce(k);
            if (status) return;
            const FT rkk = r[(size_t)k * n + k];
            const FT rp = r[(size_t)(k - 1) * n + (k - 1)];
            if (!(rp > FT(0.0)) || !std::isfinite(to_double(rkk))) { status = LGS_E_INVALID; return; }  // dependent rows
            const FT mkk = mu[(size_t)k * n + (k - 1)];
            const FT s = rkk + mkk * mkk * rp;  // squared length of b_k projected orthogonally to rows < k-1
            if (FT(delta) * rp <= s) {
                valid = k + 1;
                k++;
            } else {
                swap_rows(k, k - 1);
                swaps++;
                valid = k - 1;
                if (k - 1 == 0) {
                    gso_row(0);
                    valid = 1;
                }
                k = std::max(k - 1, 1);
            }
        }
        if (status == LGS_OK && kend > 0 && !(r[(size_t)(kend - 1) * n + (kend - 1)] > FT(0.0))) status = LGS_E_INVALID;
    }

    // Schnorr-Euchner enumeration of the projected block [j, k]: the coefficient vector of a lattice vector whose
    // projection orthogonal to rows < j is shorter than sqrt(bound); returns false if none.  prune = 1: linear
    // pruning of the partial lengths (heuristic, used for large blocks only).
    bool enumerate(int j, int k, double bound, int prune, std::vector<i64>& best) {
        const int d = k - j + 1;
        std::vector<double> c(d), y(d + 1, 0.0), ct(d + 2, 0.0), lim(d);
        std::vector<i64> ut(d + 1, 0), v(d + 1, 0), Dl(d + 1, 0), dl(d + 1, 1);
        std::vector<double> mub((size_t)d * d, 0.0);
        for (int a = 0; a < d; a++) {
            c[a] = to_double(r[(size_t)(j + a) * n + (j + a)]);
            for (int e = 0; e < a; e++) mub[(size_t)a * d + e] = to_double(mu[(size_t)(j + a) * n + (j + e)]);
        }
        double cbar = bound;
        auto set_limits = [&]() {
            for (int t = 0; t < d; t++) {
                // level t has d - t coordinates fixed
                lim[t] = prune ? cbar * std::min(1.0, 1.05 * (double)(d - t) / (double)d + 0.05) : cbar;
            }
            lim[0] = cbar;
        };
        set_limits();
        bool found = false;
        ut[0] = 1;
        int s = 0, t = 0;
        while (t < d) {
            enum_nodes++;
            const double w = y[t] + (double)ut[t];
            ct[t] = ct[t + 1] + w * w * c[t];
            if (ct[t] < lim[t]) {
                if (t > 0) {
                    t--;
                    double acc = 0;
                    for (int i = t + 1; i <= s; i++) acc += (double)ut[i] * mub[(size_t)i * d + t];
                    y[t] = acc;
                    const double rr = std::nearbyint(-acc);
                    ut[t] = v[t] = (i64)rr;
                    Dl[t] = 0;
                    dl[t] = (rr > -acc) ? -1 : 1;
                } else {
                    cbar = ct[0];
                    set_limits();
                    best.assign(ut.begin(), ut.begin() + d);
                    found = true;
                }
            } else {
                t++;
                if (t >= d) break;
                s = std::max(s, t);
                if (t < s) Dl[t] = -Dl[t];
                if (Dl[t] * dl[t] >= 0) Dl[t] += dl[t];
                ut[t] = v[t] + Dl[t];
            }
        }
        return found;
    }

    // make w = sum_a coef[a] b_{j+a} a member of the basis by unimodular row operations inside the block, then move it
    // to position j.  coef must be primitive (gcd 1), which holds for any shortest vector of the projected block.
    bool insert_combination(int j, std::vector<i64> coef) {
        const int d = (int)coef.size();
        for (;;) {
            int a = -1, c2 = -1;  // a: smallest nonzero |coef|, c2: any other nonzero
            for (int i = 0; i < d; i++)
                if (coef[i] && (a < 0 || std::llabs(coef[i]) < std::llabs(coef[a]))) a = i;
            if (a < 0) return false;
            for (int i = 0; i < d; i++)
                if (i != a && coef[i]) { c2 = i; break; }
            if (c2 < 0) {
                if (std::llabs(coef[a]) != 1) return false;  // not primitive: refuse (nothing was lost, ops were unimodular)
                rotate_up(j, j + a);
                return true;
            }
            // w = coef[c2] b_c2 + coef[a] b_a + ... ; coef[c2] = q coef[a] + rem  =>  b_a += q b_c2, coef[c2] = rem
            const i64 q = coef[c2] / coef[a];
            row_axpy(j + a, j + c2, -q);
            if (status) return false;
            refresh_row(j + a);
            coef[c2] -= q * coef[a];
        }
    }

    void bkz(int beta, int max_tours, int prune) {
        lll(0, n);
        if (status || beta < 2 || n < 2) return;
        beta = std::min(beta, n);
        int z = 0, j = -1;
        i64 blocks = 0;
        const i64 max_blocks = max_tours > 0 ? (i64)max_tours * (n - 1) : INT64_MAX;
        std::vector<i64> coef;
        while (z < n - 1 && blocks < max_blocks && status == LGS_OK) {
            j = (j + 1) % (n - 1);
            const int k = std::min(j + beta - 1, n - 1);
            blocks++;
            const double rjj = to_double(r[(size_t)j * n + j]);
            const bool found = enumerate(j, k, 0.99 * rjj, prune, coef);
            bool trivial = found;
            if (found)
                for (int a = 1; a < (int)coef.size(); a++) trivial = trivial && coef[a] == 0;
            if (!found || trivial) {
                z++;
                continue;
            }
            if (!insert_combination(j, coef)) {
                if (status) return;
                z++;
                valid = std::min(valid, j);
                lll(j, n);
                continue;
            }
            insertions++;
            z = 0;
            valid = std::min(valid, j);
            lll(j, n);
        }
    }
};

// independent check in double-double on the final basis: (delta, eta)-reduced up to a small slack
bool verify_reduced(const i64* rows, int n, int m, double delta, double eta) {
    Reducer<dd> chk(rows, n, m, nullptr, delta, eta);
    if (chk.status) return false;
    for (int k = 0; k < n; k++) chk.gso_row(k);
    for (int k = 0; k < n; k++) {
        const double rkk = to_double(chk.r[(size_t)k * n + k]);
        if (!(rkk > 0)) return false;
        for (int j = 0; j < k; j++)
            if (std::fabs(to_double(chk.mu[(size_t)k * n + j])) > eta + 1e-6) return false;
        if (k) {
            const double mkk = to_double(chk.mu[(size_t)k * n + k - 1]);
            const double rp = to_double(chk.r[(size_t)(k - 1) * n + k - 1]);
            if (rkk + mkk * mkk * rp < (delta - 1e-6) * rp) return false;
        }
    }
    return true;
}

struct Job {
    int n, m, beta, max_tours, prune;
    double delta, eta;
    std::vector<i64> b, u;  // current basis / transform: every tier continues from where the previous one stopped
    i64 stats[5] = {0, 0, 0, 0, 0};
};

// one precision tier; the row operations are exact, so whatever it leaves in job.b is a basis of the same lattice
template <class FT>
int attempt(Job& job, int tag) {
    Reducer<FT> red(job.b.data(), job.n, job.m, job.u.empty() ? nullptr : job.u.data(), job.delta, job.eta);
    if (red.status) return red.status;
    if (job.beta >= 2) red.bkz(job.beta, job.max_tours, job.prune);
    else red.lll(0, job.n);
    job.stats[0] += red.swaps;
    job.stats[1] += red.passes;
    job.stats[2] += red.insertions;
    job.stats[3] += red.enum_nodes;
    job.stats[4] = tag;
    if (red.status == LGS_OK || red.status == LGS_E_PRECISION) {
        job.b = red.b;
        job.u = red.u;
    }
    if (red.status) return red.status;
    return verify_reduced(job.b.data(), job.n, job.m, job.delta, job.eta) ? LGS_OK : LGS_E_PRECISION;
}

int reduce_entry(i64* rows, int n, int m, double delta, double eta, int beta, int max_tours, int prune, int precision,
                 i64* U, i64* stats) {
    if (!rows || n <= 0 || m < n || m >= (1 << 15) || !(delta > 0.25) || !(delta <= 1.0) || !(eta >= 0.5) ||
        !(eta < std::sqrt(delta)))
        return LGS_E_INVALID;
    if (precision != 0 && precision != 53 && precision != 64 && precision != 106) return LGS_E_INVALID;
    Job job;
    job.n = n, job.m = m, job.beta = beta, job.max_tours = max_tours, job.prune = prune;
    job.delta = delta, job.eta = eta;
    job.b.assign(rows, rows + (size_t)n * m);
    if (U) {
        job.u.assign((size_t)n * n, 0);
        for (int i = 0; i < n; i++) job.u[(size_t)i * n + i] = 1;
    }
    int rc = LGS_E_PRECISION;
    const int first = precision ? precision : (n <= 96 ? 53 : 64);
    if (first == 53) rc = attempt<double>(job, 53);
    if (rc == LGS_E_PRECISION && (first == 64 || (precision == 0 && first < 64))) rc = attempt<long double>(job, 64);
    if (rc == LGS_E_PRECISION && (first == 106 || precision == 0)) rc = attempt<dd>(job, 106);
    if (rc) return rc;
    std::memcpy(rows, job.b.data(), sizeof(i64) * (size_t)n * m);
    if (U) std::memcpy(U, job.u.data(), sizeof(i64) * (size_t)n * n);
    if (stats) std::memcpy(stats, job.stats, sizeof(job.stats));
    return LGS_OK;
}

}  // namespace

// this file is compiled with -mfma (std::fma must be the single-rounding hardware instruction for the exact products
// and the double-double arithmetic): refuse to run, instead of dying with SIGILL, on a host without it
static bool host_has_fma() {
    static const bool ok = __builtin_cpu_supports("fma");
    return ok;
}

extern "C" {

int lgs_lll_reduce(int64_t* rows, int32_t n, int32_t m, double delta, double eta, int32_t precision, int64_t* U,
                   int64_t* stats) {
    if (!host_has_fma()) return LGS_E_PRECISION;
    try {
        return reduce_entry(rows, n, m, delta, eta, 0, 0, 0, precision, U, stats);
    } catch (...) {  // std::bad_alloc from the work arrays: nothing throws across the ABI
        return LGS_E_NOMEM;
    }
}

int lgs_bkz_reduce(int64_t* rows, int32_t n, int32_t m, int32_t block_size, double delta, int32_t max_tours,
                   int32_t prune, int32_t precision, int64_t* U, int64_t* stats) {
    if (block_size < 2) return LGS_E_INVALID;
    if (!host_has_fma()) return LGS_E_PRECISION;
    if (prune < 0) prune = block_size > 32 ? 1 : 0;
    try {
        return reduce_entry(rows, n, m, delta, 0.501, block_size, max_tours, prune, precision, U, stats);
    } catch (...) {
        return LGS_E_NOMEM;
    }
}

int lgs_gso_norms(const int64_t* rows, int32_t n, int32_t m, double* norms) {
    if (!rows || !norms || n <= 0 || m < n) return LGS_E_INVALID;
    if (!host_has_fma()) return LGS_E_PRECISION;
    try {
        Reducer<dd> red(rows, n, m, nullptr, 0.99, 0.501);
        if (red.status) return red.status;
        for (int k = 0; k < n; k++) {
            red.gso_row(k);
            const double d = to_double(red.r[(size_t)k * n + k]);
            if (!(d > 0)) return LGS_E_INVALID;
            norms[k] = std::sqrt(d);
        }
        return LGS_OK;
    } catch (...) {
        return LGS_E_NOMEM;
    }
}

}  // extern "C"
