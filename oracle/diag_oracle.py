"""CPU restatement of the reference's chain diagnostics -- TEST INFRASTRUCTURE (only tests/ may import it).

Follows src/diagnostics/convergence_diag.py:75-145 (compute_autocorrelation, integrated_autocorrelation_time),
:176-213 (gelman_rubin_statistic) and src/diagnostics/mcmc_diag.py:59-78 (effective_sample_size, 'autocorr').
Pinned: tests/golden/diag_ar1.npz holds outputs of the reference's own functions (minted by
tests/golden/make_golden_diag.py, which imports /root/reference/src/diagnostics); tests/test_oracle_golden.py requires
this file to reproduce them.
"""
import numpy as np


def compute_autocorrelation(x, max_lag=None):
    """convergence_diag.py:75-112: biased autocovariance via FFT, normalised by lag 0, lags 0..max_lag (len // 4)."""
    x = np.asarray(x, dtype=np.float64).flatten()
    if max_lag is None:
        max_lag = len(x) // 4
    x = x - np.mean(x)
    n = len(x)
    padded = 2 ** int(np.ceil(np.log2(2 * n - 1)))
    f = np.fft.fft(x, n=padded)
    acf = np.fft.ifft(f * np.conj(f)).real[:n]
    acf = acf / acf[0]
    return acf[:max_lag + 1]


def integrated_autocorrelation_time(x, c=5.0, max_lag=None):
    """convergence_diag.py:115-145: tau = 1 + sum 2 acf[k], stopped at the first k >= c * (sum so far)."""
    acf = compute_autocorrelation(x, max_lag)
    tau, window = 0.0, len(acf) - 1
    for k in range(1, len(acf)):
        tau += 2 * acf[k]
        if k >= c * tau:
            window = k
            break
    return 1 + tau, window


def effective_sample_size(x):
    """mcmc_diag.py:59-78 ('autocorr') and :98-105 (multivariate: minimum over coordinates)."""
    x = np.asarray(x, dtype=np.float64)
    if x.ndim == 1:
        return len(x) / integrated_autocorrelation_time(x)[0]
    return min(effective_sample_size(x[:, i]) for i in range(x.shape[1]))


def gelman_rubin_statistic(chains):
    """convergence_diag.py:176-213."""
    m = len(chains)
    n = min(len(c) for c in chains)
    chains = [np.asarray(c[:n], dtype=np.float64) for c in chains]
    means = [np.mean(c) for c in chains]
    overall = np.mean(means)
    B = n / (m - 1) * sum((mu - overall) ** 2 for mu in means)
    W = np.mean([np.var(c, ddof=1) for c in chains])
    return float(np.sqrt((((n - 1) / n) * W + B / n) / W))


def compute_tvd(samples1, samples2, bins=None):
    """convergence_diag.py:15-72.  1-D, bins=None: exact-value counting over the union of values,
    0.5 * sum |n1(v)/N1 - n2(v)/N2|; 1-D with bins: np.histogram of both on the shared [min, max] range, each
    normalised by its own total; multivariate: MEAN of the per-coordinate TVDs.  Pinned: tests/golden/tvd_d1.npz
    (minted by tests/golden/make_golden_tvd.py from the reference's function)."""
    a, b = np.asarray(samples1), np.asarray(samples2)
    if a.ndim == 1 and b.ndim == 1:
        if bins is None:
            values, inv = np.unique(np.concatenate([a, b]), return_inverse=True)
            c1 = np.bincount(inv[:len(a)], minlength=len(values)) / len(a)
            c2 = np.bincount(inv[len(a):], minlength=len(values)) / len(b)
            tvd = 0.0
            for p1, p2 in zip(c1, c2):      # the reference accumulates value by value, in sorted order
                tvd += abs(p1 - p2)
            return 0.5 * tvd
        lo, hi = min(a.min(), b.min()), max(a.max(), b.max())
        h1, _ = np.histogram(a, bins=bins, range=(lo, hi))
        h2, _ = np.histogram(b, bins=bins, range=(lo, hi))
        return 0.5 * np.abs(h1 / h1.sum() - h2 / h2.sum()).sum()
    return float(np.mean([compute_tvd(a[:, i], b[:, i], bins) for i in range(a.shape[1])]))


def compute_acceptance_rate(accepted):
    """mcmc_diag.py:107-117."""
    return np.mean(accepted)
