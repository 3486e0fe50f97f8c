/*
 * lgs_b200.h -- C ABI of the B200 lattice Gaussian sampler (liblgs_b200.so).
 *
 * The reference (NickQrumpton/lattice-gaussian-mcmc) is pure Python and has no FFI layer: its boundary for this
 * path is the class surface exported by src/samplers/__init__.py:3-7.  Each entry point below names the
 * reference method it replaces; INTEGRATION.md shows the ctypes stub a reference maintainer would add.
 *
 * Conventions
 *   - every function returns 0 on success or a negative LGS_E_* code; nothing throws across the ABI;
 *     lgs_last_error() gives the text of the last failure on a handle (or of the last lgs_create()).
 *   - one handle per CUDA device; calls on one handle are serialised by the caller; handles are independent.
 *     A handle owns scratch buffers (coefficient workspaces, diagnostics scratch) that its calls write in stream
 *     order: issue the calls of one handle on ONE stream at a time, or order the streams with events yourself.
 *   - `*_dev` pointers are device memory on the handle's device (e.g. a torch tensor's data_ptr());
 *     `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).  Calls are asynchronous
 *     with respect to the host unless the name ends in `_host`.
 *   - `*_host` entry points take ordinary host buffers and perform the host<->device copies themselves
 *     (this is what the reference-facing Python classes and bench.py's `e2e` figure go through).
 *   - matrices are row-major.  The basis has COLUMNS = basis vectors, as the reference samplers use it
 *     (src/samplers/klein.py:59 `np.linalg.qr(self.lattice.basis)`, :218 `self.lattice.basis @ x`).
 *   - there is no CPU fallback: without a CUDA device lgs_create() fails with LGS_E_CUDA.
 */
#ifndef LGS_B200_H
#define LGS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define LGS_API __attribute__((visibility("default")))
#else
#define LGS_API
#endif

#define LGS_ABI_VERSION 1

enum {
    LGS_OK = 0,
    LGS_E_INVALID = -1, /* bad argument (n <= 0, sigma <= 0, NULL pointer, ...) */
    LGS_E_CUDA = -2,    /* CUDA runtime error; text in lgs_last_error() */
    LGS_E_NOMEM = -3,   /* device (or, in the host-side reducer, host) allocation failed */
    LGS_E_STATE = -4,   /* call not valid in the current chain state */
    LGS_E_FALLBACK = -5, /* replay: a draw hit the reference's no-uniform fallback (klein.py:176-179) */
    LGS_E_RANGE = -6,    /* basis reduction: an integer left the exact range (|entry| >= 2^56 or a transform overflow) */
    LGS_E_PRECISION = -7 /* basis reduction: the floating-point Gram-Schmidt data was not accurate enough */
};

/* 1-D sampler used for each coordinate */
enum {
    LGS_SAMPLEZ_TABLE = 0, /* the reference's normalised table + inverse CDF (klein.py:101-179), same arithmetic order */
    LGS_SAMPLEZ_FAST = 1   /* same distribution, recurrence-built weights walked outward from round(mean); wide truncated
                            * windows (sigma_i >= 50: the reference's max_range + 1 entry tables, klein.py:122-126) use
                            * closed-form window masses (LGS_TRUNC_CLOSED_FORM=0 at lgs_create: the summed form, A/B) */
};

/* importance weight used by the IMHK accept step */
enum {
    LGS_WEIGHT_REFERENCE = 0, /* imhk.py:102-124 as written: log rho(v) - compute_log_density(v) (a constant +- 1e-14) */
    LGS_WEIGHT_EXACT = 1      /* Wang-Ling: w(x) = prod_i rho_{sigma_i, m_i(x)}(Z), the table normalisers */
};

/* which kernels carry the contraction work */
enum {
    LGS_PATH_AUTO = 0,
    LGS_PATH_ORDERED = 1, /* one thread per sample, reference summation order (bit-faithful conditional means) */
    LGS_PATH_BLOCKED = 2  /* panel-blocked back-substitution + lattice-point contraction on FP64 DMMA */
};

typedef struct lgs_handle lgs_handle;

/* Host-side description of one sampler.  Replaces the state RefinedKleinSampler.__init__ /
 * _precompute_qr_stable build (src/samplers/klein.py:26-79): the QR factor is computed on the host by the
 * caller (numpy/LAPACK, as the reference does) and uploaded once. */
typedef struct {
    int32_t n;            /* lattice dimension */
    const double *R;      /* [n*n] upper-triangular factor, rows sign-fixed to a positive diagonal (klein.py:68-72) */
    const double *B;      /* [n*n] basis, columns are basis vectors */
    const double *t;      /* [n]   Q^T center (klein.py:75) */
    const double *center; /* [n]   center c */
    double sigma;         /* > 0 (base.py:74) */
    int32_t precision;    /* window half-width in conditional sigmas (klein.py:26, default 10) */
    int32_t max_range;    /* window cap, reference constant 1000 (klein.py:122); 0 = 1000 */
    int32_t device;       /* CUDA device ordinal */
    int32_t path;         /* LGS_PATH_* */
} lgs_desc;

LGS_API int lgs_abi_version(void);

/* RefinedKleinSampler.__init__ (klein.py:26-54) / IMHKSampler.__init__ (imhk.py:32-66): upload R, B, t, c. */
LGS_API int lgs_create(const lgs_desc *desc, lgs_handle **out);
LGS_API void lgs_destroy(lgs_handle *h);
LGS_API const char *lgs_last_error(const lgs_handle *h); /* h may be NULL: last lgs_create() failure */

/* number of uniforms one Klein sample consumes in replay mode: coordinates with sigma_i >= 1e-10 (klein.py:201) */
LGS_API int lgs_uniforms_per_sample(const lgs_handle *h);

/*
 * Samples (or chains) per call that fill the device evenly: the blocked path works on tiles of 128 samples and a
 * full batch gives every SM the same whole number of tiles (148 SMs x 9 tiles x 128 = 170 496 on a B200).  Any S is
 * accepted by the sampling calls; multiples of this figure run at the quoted throughput.  No reference counterpart.
 */
LGS_API int64_t lgs_preferred_batch(const lgs_handle *h);

/*
 * KleinSampler.sample(S) (klein.py:324-337 -> sample_single 181-220), native RNG.
 * Sample s uses Philox4x32-10 with key = seed and counter = (first_sample_id + s, draw pair, 0): the output does
 * not depend on how [0, S) is split over calls, streams or GPUs.
 *   V_dev    [S*n] lattice points B x, or NULL
 *   X_dev    [S*n] integer coefficient vectors x, or NULL
 *   logq_dev [S]   Klein log-density of each sample as compute_log_density returns it (klein.py:222-271), or NULL
 */
LGS_API int lgs_klein_sample(lgs_handle *h, int64_t S, uint64_t seed, uint64_t first_sample_id, int samplez,
                     double *V_dev, int64_t *X_dev, double *logq_dev, void *stream);

/*
 * The same with the reference's recorded uniform stream: sample s consumes
 * U_dev[s*ups .. (s+1)*ups), ups = lgs_uniforms_per_sample(), in coordinate order i = n-1 .. 0 -- i.e.
 * np.random.RandomState(seed).random_sample(S*ups) reproduces np.random.seed(seed); sampler.sample(S).
 * Always uses LGS_SAMPLEZ_TABLE.  fallback_count_dev (int32[1], may be NULL) counts draws that took the
 * reference's no-uniform fallback.  Runs on the kernel lgs_replay_path() names.
 */
LGS_API int lgs_klein_replay(lgs_handle *h, int64_t S, const double *U_dev, double *V_dev, int64_t *X_dev,
                     double *logq_dev, int32_t *fallback_count_dev, void *stream);

/*
 * Which kernel the replay entry points of this handle run on: LGS_PATH_ORDERED (one thread per sample, the reference's
 * summation order: klein_ordered_kernel / imhk_ordered_kernel) or LGS_PATH_BLOCKED (the throughput kernel itself --
 * klein_chain_kernel's REPLAY instantiation: the conditional means of klein.py:189-195 come from the FP64 DMMA
 * contraction, the draw is the reference's table SampleZ on the recorded uniform).  A handle created with
 * lgs_desc.path = LGS_PATH_BLOCKED replays on the chain kernel whenever that kernel serves the lattice.
 */
LGS_API int lgs_replay_path(const lgs_handle *h);

/* Host-buffer forms (pageable or pinned host memory; copies inside; synchronous).
 *
 * Transport of lattice points to the host.  The caller always receives the reference's float64 [S, n] array
 * (klein.py:322).  For an integer basis every lattice point is an integer vector, so when only V_host is requested the
 * bytes that cross PCIe are int16 (a quarter of fp64) and are widened into V_host by a pool of host threads:
 *   LGS_TRANSPORT_INT16_ZEROCOPY  chain kernel, n % 8 == 0: the kernel's lattice-point epilogue writes int16 straight into
 *                                 mapped pinned host memory and publishes a completion word per 128-sample tile; the
 *                                 host widens finished tiles while the kernel is still running (one launch per batch);
 *   LGS_TRANSPORT_INT16_COPY      any other kernel path: sample + pack kernel + copy in chunks, widened two chunks behind;
 *   LGS_TRANSPORT_FP64            real-valued basis, or X_host / logq_host requested: plain fp64 copies.
 * lgs_imhk_run_host ships kept states of an integer basis the same way (pack kernel + chunked copies + widening) once
 * they exceed 4 Mi values.
 * A batch in which a value does not fit int16 is delivered through the fp64 route instead (same samples).  The array
 * handed back is bit-identical in all three (tests/test_gpu_hostpack.py).  LGS_HOST_THREADS sets the pool size
 * (default: usable CPUs / LOCAL_WORLD_SIZE; with fewer than four threads the fp64 route is used); LGS_HOST_PACK=0 / LGS_HOST_ZEROCOPY=0 select the plainer routes (development A/B).
 * fp64 / int64 results bound for PAGEABLE host memory (>= 8 MB) are staged through a pinned ring and moved on by the
 * same pool (LGS_HOST_BOUNCE=0: the driver's pageable copy); page-locked destinations are copied directly. */
enum { LGS_TRANSPORT_FP64 = 0, LGS_TRANSPORT_INT16_COPY = 1, LGS_TRANSPORT_INT16_ZEROCOPY = 2 };
LGS_API int lgs_host_transport(const lgs_handle *h); /* the route lgs_klein_sample_host(V_host only) takes on this handle */
/* the host half of the int16 transport on its own (no handle, no CUDA): dst[i] = (double)src[i], non-temporal stores,
 * split over `threads` pool threads (0 or 1: the calling thread) */
LGS_API int lgs_host_widen_i16(const int16_t *src, double *dst, int64_t count, int32_t threads);
LGS_API int lgs_klein_sample_host(lgs_handle *h, int64_t S, uint64_t seed, uint64_t first_sample_id, int samplez,
                          double *V_host, int64_t *X_host, double *logq_host);
LGS_API int lgs_klein_replay_host(lgs_handle *h, int64_t S, const double *U_host, double *V_host, int64_t *X_host,
                          double *logq_host, int32_t *fallback_count_host);

/*
 * RefinedKleinSampler.compute_log_density (klein.py:222-271) for S points given by their integer coefficient
 * vectors X[S*n] (the reference recovers them from the point with an O(n^3) solve + round, klein.py:230-231; the
 * host wrapper does the same before calling this).  logq[S].
 */
LGS_API int lgs_klein_logdensity(lgs_handle *h, int64_t S, const int64_t *X_dev, double *logq_dev, void *stream);
LGS_API int lgs_klein_logdensity_host(lgs_handle *h, int64_t S, const int64_t *X_host, double *logq_host);

/*
 * IMHK chains.  A handle owns one set of `chains` independent Markov chains (IMHKSampler objects,
 * imhk.py:19-66).  lgs_imhk_reset() discards their state; the next run initialises each chain lazily with one
 * Klein sample exactly as _initialize_chain does (imhk.py:126-139).
 * Native RNG: chain k uses Philox key = seed, counter = (first_chain_id + k, draw pair, step index).
 */
LGS_API int lgs_imhk_reset(lgs_handle *h, int64_t chains, uint64_t seed, uint64_t first_chain_id);

/*
 * IMHKSampler.sample(keep, thin) (imhk.py:196-229) preceded by `burn_in` discarded steps
 * (the README / src/core/base_sampler.py:45-76 form sample(n_samples, burn_in, thin)):
 * every chain performs burn_in + keep*thin step()s (imhk.py:141-177) and its state is stored after each
 * thin-th step of the last keep*thin.
 *   V_dev      [chains*keep*n] kept states as lattice points (chain-major), or NULL
 *   X_dev      [chains*keep*n] kept states as coefficient vectors, or NULL
 * Burn-in and thinning happen inside the kernel; only kept states are written.
 */
LGS_API int lgs_imhk_run(lgs_handle *h, int64_t burn_in, int64_t keep, int64_t thin, int samplez, int weight_mode,
                 double *V_dev, int64_t *X_dev, void *stream);

/*
 * Replay form: chain k consumes U_dev[k*stride ...): [ups for the initial state, only if the chain is not yet
 * initialised] then per step [ups for the proposal, 1 for the accept test].
 *   accept_dev [chains*steps] accept bit per step, or NULL;  steps = burn_in + keep*thin
 *   logw_dev   [chains*(steps+1)] log importance weight of the state at entry and of every proposal, or NULL
 */
LGS_API int lgs_imhk_replay(lgs_handle *h, int64_t burn_in, int64_t keep, int64_t thin, int weight_mode,
                    const double *U_dev, int64_t stride, double *V_dev, int64_t *X_dev, uint8_t *accept_dev,
                    double *logw_dev, int32_t *fallback_count_dev, void *stream);

/*
 * Chain statistics and state (IMHKSampler.accepted_proposals / total_proposals / current_state /
 * current_coeffs / current_log_weight, imhk.py:54-61).  Device buffers, any may be NULL.
 *   counts_dev [chains*2] (accepted, total) per chain
 *   X_dev [chains*n], V_dev [chains*n], logw_dev [chains]
 */
LGS_API int lgs_imhk_get_state(lgs_handle *h, int64_t *counts_dev, int64_t *X_dev, double *V_dev, double *logw_dev,
                       void *stream);
/*
 * Checkpoint / resume.  The reference's chain state is three plain attributes a caller can assign
 * (IMHKSampler.current_state / current_coeffs / current_log_weight, imhk.py:59-61) plus two counters (imhk.py:54-55).
 * lgs_imhk_set_state installs, after lgs_imhk_reset(chains, seed, first_chain_id), for every chain: its integer
 * coefficient vector X_dev [chains*n] (the lattice point is B x), its log importance weight logw_dev [chains], its
 * counters counts_dev [chains*2] (NULL: zeros), and `steps_done`, the number of Metropolis-Hastings steps the chain had
 * performed when the state was taken (lgs_imhk_steps_done) -- the index of its Philox stream, so a resumed native-mode
 * run continues the uninterrupted run draw for draw.  `samplez` names the kernels that will continue the chains
 * (as in lgs_imhk_run: LGS_SAMPLEZ_FAST on a non-ordered handle = the chain kernel, else the ordered kernels).
 */
LGS_API int lgs_imhk_set_state(lgs_handle *h, int samplez, const int64_t *X_dev, const double *logw_dev,
                       const int64_t *counts_dev, int64_t steps_done, void *stream);
LGS_API int lgs_imhk_set_state_host(lgs_handle *h, int samplez, const int64_t *X_host, const double *logw_host,
                            const int64_t *counts_host, int64_t steps_done);
LGS_API int64_t lgs_imhk_steps_done(const lgs_handle *h);
/* sum over chains of (accepted, total), copied to the host: the two numbers acceptance_rate is made of */
LGS_API int lgs_imhk_counts_host(lgs_handle *h, int64_t *accepted, int64_t *total);

/* Host-buffer forms of the IMHK entry points (copies inside; synchronous).  Any output may be NULL. */
LGS_API int lgs_imhk_run_host(lgs_handle *h, int64_t burn_in, int64_t keep, int64_t thin, int samplez, int weight_mode,
                      double *V_host, int64_t *X_host);
LGS_API int lgs_imhk_replay_host(lgs_handle *h, int64_t burn_in, int64_t keep, int64_t thin, int weight_mode,
                         const double *U_host, int64_t stride, double *V_host, int64_t *X_host, uint8_t *accept_host,
                         double *logw_host, int32_t *fallback_count_host);
LGS_API int lgs_imhk_get_state_host(lgs_handle *h, int64_t *counts_host, int64_t *X_host, double *V_host,
                            double *logw_host);

/*
 * Marginal histograms for src/diagnostics/convergence_diag.py:15-72 (compute_tvd) on device-resident samples:
 * for each tracked coordinate d in dims[], counts of floor((V[s, d] - lo) / width) over [0, nbins), values
 * outside are dropped into the first / last bin.  hist_dev is int64[n_dims*nbins] and is ACCUMULATED into, so
 * per-rank results can be all-reduced (sum) afterwards.  Any n_dims and nbins: coordinates are processed 256 per launch,
 * with block-private shared-memory bins while 256 * nbins * 4 bytes fit 160 KB and global (L2) atomics beyond that
 * (e.g. all 1024 marginals of the NTRU lattice over +-1500).  Touches no per-handle scratch.
 */
LGS_API int lgs_histogram(lgs_handle *h, const double *V_dev, int64_t S, const int32_t *dims_host, int32_t n_dims,
                  double lo, double width, int32_t nbins, int64_t *hist_dev, void *stream);

/*
 * The same counting FUSED into the sampling kernel ("fuse the per-rank accumulation into the sampling-kernel
 * epilogue so the allreduce is the only extra launch"): while a histogram is attached, every NEW lattice point the chain
 * kernel writes to a V (lgs_klein_sample; the kept states of lgs_imhk_run / lgs_imhk_replay and of their _host forms)
 * is also counted, for the tracked coordinates only, by the lattice-point epilogue that holds it in registers -- same bins and
 * clamping as lgs_histogram, ACCUMULATED into hist_dev (int64[n_dims * nbins], caller-owned, must stay valid while
 * attached).  hist_dev = NULL detaches.  Handles whose lattice the chain kernel does not serve (path = LGS_PATH_ORDERED,
 * clamped sigma_i) return LGS_E_STATE: call lgs_histogram on V instead.  Not counted: the chains' current states read
 * by lgs_imhk_get_state[_host] (they are not new samples; polling must not change the histogram) and the batches of
 * lgs_klein_sample_host / lgs_klein_replay_host, on either host transport.  The call synchronises the device (it replaces a small per-handle table).
 */
LGS_API int lgs_attach_histogram(lgs_handle *h, const int32_t *dims_host, int32_t n_dims, double lo, double width,
                         int32_t nbins, int64_t *hist_dev);

/*
 * Chain diagnostics on kept states that stay on the device (V_dev[chains][keep][n], the layout lgs_imhk_run writes).
 * For every chain and every tracked coordinate dims_host[d] the time series x_t = V[chain][t][dim] gives
 *   stats_dev[chain][d][LGS_DIAG_FIELDS] = { mean, variance (ddof = 1), tau_int, window, acf[1] }
 * with the reference's arithmetic: compute_autocorrelation (src/diagnostics/convergence_diag.py:75-112,
 * mcmc_diag.py:12-33: biased autocovariance normalised by lag 0) and integrated_autocorrelation_time
 * (convergence_diag.py:115-145, mcmc_diag.py:36-57: tau = 1 + 2 sum acf[k] up to the first k >= c_window * sum, or
 * max_lag).  max_lag <= 0 selects len // 4 (convergence_diag.py:86).  effective_sample_size (mcmc_diag.py:59-104) is
 * keep / tau_int.  acf_dev (may be NULL) receives acf[0 .. acf_len) per series.
 * lgs_gelman_rubin_sums adds, per coordinate, { sum of chain means, sum of squared chain means, sum of chain
 * variances } over the chains into sums_dev[n_dims][3] -- the only quantities gelman_rubin_statistic
 * (convergence_diag.py:176-213) needs, so that ranks can all-reduce them (a few doubles) and form R-hat on the host.
 */
#define LGS_DIAG_FIELDS 5
LGS_API int lgs_chain_diagnostics(lgs_handle *h, const double *V_dev, int64_t chains, int64_t keep,
                                  const int32_t *dims_host, int32_t n_dims, double c_window, int32_t max_lag,
                                  double *stats_dev, double *acf_dev, int32_t acf_len, void *stream);
LGS_API int lgs_gelman_rubin_sums(lgs_handle *h, const double *stats_dev, int64_t chains, int32_t n_dims,
                                  double *sums_dev, void *stream);

/*
 * lgs_chain_diagnostics without any device allocation.  (No counterpart in the reference, which runs on numpy and
 * never manages device memory; this exists for the hosts this library gets embedded in.)
 *
 * Who needs it: lgs_chain_diagnostics gathers the tracked series into a work array that it cudaMalloc's on first use
 * and re-allocates when a larger request comes.  That is unusable for (a) a process whose framework has reserved
 * the GPU's memory for its own pool up front, where a stray cudaMalloc fails or eats into another tenant's budget,
 * (b) a monitoring loop that runs beside live sampling and must not stall it -- cudaMalloc / cudaFree synchronise
 * the whole device, (c) any phase of the host program in which allocation is simply off limits.  Such callers
 * bring V_dev, stats_dev and acf_dev out of their own memory and call this entry point, which obtains nothing: the
 * work array is a fixed array inside the library image, and dims_host is passed to the kernel by value.
 *
 * Arguments, arithmetic and results are those of lgs_chain_diagnostics (same kernels on the same gathered series).
 * Differences a caller has to know:
 *   - Limit: chains * n_dims * keep <= LGS_DIAG_STATIC_MAX_DOUBLES (4 Mi values = 32 MiB).  Larger requests return
 *     LGS_E_INVALID with the sizes in lgs_last_error(); split the chains over several calls (per-chain results are
 *     independent) or use lgs_chain_diagnostics.
 *   - Cost: every GPU on which this library's device code gets loaded holds those 32 MiB until the process exits,
 *     whether or not this function is ever called.
 *   - First call on a device: looks up the array's device address (cudaGetSymbolAddress, once per device, under a
 *     lock); the CUDA runtime may load the library's device code at that moment.  Make this first call during
 *     set-up, before any stream capture begins; later calls only launch the two kernels.
 *   - One array per device, shared by every handle and every call, never cleared (nothing in it needs to start at
 *     zero: each call overwrites what it reads).  Calls that can overlap on the GPU -- different streams with no
 *     ordering between them, or concurrent host threads -- race on it; order them on one stream or with events.
 *   - The result buffers are read by whoever waits on `stream`, as with lgs_chain_diagnostics.
 */
#define LGS_DIAG_STATIC_MAX_DOUBLES 4194304
LGS_API int lgs_chain_diagnostics_static(lgs_handle *h, const double *V_dev, int64_t chains, int64_t keep,
                                         const int32_t *dims_host, int32_t n_dims, double c_window, int32_t max_lag,
                                         double *stats_dev, double *acf_dev, int32_t acf_len, void *stream);

/*
 * Basis reduction on the host (no handle, no CUDA): the step before the sampling path.  Replaces the reference's
 * LatticeReduction.lll_reduce / bkz_reduce (src/lattices/reduction.py:68-133, 238-300), which call Sage's
 * Matrix.LLL(delta) / Matrix.BKZ(block_size) (fplll; also klein_scaling_analysis.py:95-108).  ROWS are basis vectors
 * (Sage's convention there).  `rows` [n*m] int64 row-major is reduced in place; `U` [n*n] (or NULL) receives the
 * unimodular transform with reduced = U * input (reduction.py:104 `B.solve_left(B_reduced)`); `stats` [5] (or NULL)
 * receives {swaps, size-reduction passes, BKZ insertions, enumeration nodes, float precision used (53 | 64 | 106)}.
 * precision: 0 = automatic (double for n <= 96, else x87 extended, then double-double; a tier whose Gram-Schmidt data
 *   proves too inaccurate hands its exact basis to the next), or one fixed tier 53 / 64 / 106.
 *   lgs_lll_reduce   L2-style LLL, (delta, eta)-reduced output, e.g. (0.99, 0.501) = Sage's LLL(delta=0.99).
 *   lgs_bkz_reduce   Schnorr-Euchner BKZ with block size `block_size`; max_tours <= 0: run until no block improves;
 *                    prune: 0 full enumeration, 1 linear pruning, < 0 automatic (pruned above block size 32).
 *   lgs_gso_norms    ||b*_i|| of the rows (reduction.py:386-391 GSO(B).update_gso()).
 * Errors: LGS_E_INVALID (NULL, m < n, linearly dependent rows, bad delta / eta), LGS_E_RANGE, LGS_E_PRECISION.
 */
LGS_API int lgs_lll_reduce(int64_t *rows, int32_t n, int32_t m, double delta, double eta, int32_t precision,
                           int64_t *U, int64_t *stats);
LGS_API int lgs_bkz_reduce(int64_t *rows, int32_t n, int32_t m, int32_t block_size, double delta, int32_t max_tours,
                           int32_t prune, int32_t precision, int64_t *U, int64_t *stats);
LGS_API int lgs_gso_norms(const int64_t *rows, int32_t n, int32_t m, double *norms);

/*
 * Measurement hooks (no reference counterpart; used by bench.py for the roofline figures).
 * While enabled, every kernel launch of the blocked path is bracketed by CUDA events on its own stream;
 * lgs_profile_read() synchronises the device, returns the summed device time [ms] and the launch count per kernel
 * class, and clears the records.  The launch count doubles as the "how many of our kernels ran" figure.
 */
enum {
    LGS_PROF_PANEL = 0,   /* klein_panel_kernel: 32 coordinates of every sample, SampleZ + in-block triangle */
    LGS_PROF_CENTERS = 1, /* contract_kernel, centers of a block pair from all finished coordinates (DMMA) */
    LGS_PROF_ACCUM = 2,   /* contract_kernel, 32x32 in-pair update (DMMA) */
    LGS_PROF_POINT = 3,   /* contract_kernel, lattice points V = X B^T (DMMA) */
    LGS_PROF_OTHER = 4,   /* accept / commit / export / log-density finish */
    LGS_PROF_CHAIN = 5,   /* klein_chain_kernel: the persistent warp-specialised kernel (centers + panels + lattice points) */
    LGS_PROF_CLASSES = 6
};
LGS_API int lgs_profile_enable(lgs_handle *h, int on);
LGS_API int lgs_profile_read(lgs_handle *h, double *ms_by_class, int64_t *launches_by_class);

#ifdef __cplusplus
}
#endif
#endif /* LGS_B200_H */
