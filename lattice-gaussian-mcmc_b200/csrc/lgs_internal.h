// Internal declarations shared by the translation units of liblgs_b200.so.
#pragma once
#include <cstdint>
#include <string>
#include <cuda_runtime.h>

#include <vector>

#include "../../include/lgs_b200.h"
#include "lgs_device.cuh"

namespace lgs {

// How the uniforms of a launch are obtained.
struct RngSpec {
    const double* U;     // replay stream (device) or nullptr for Philox
    int64_t stride;      // replay: doubles per unit (sample or chain)
    uint64_t seed;       // Philox key
    uint64_t first_id;   // global id of unit 0 of this launch
};

// IMHK chain set owned by a handle (device memory).
struct ChainSet {
    int64_t chains = 0;
    uint64_t seed = 0, first_id = 0;
    // ordered path: double-buffered states, column layout [2][n][chains]
    double* Xbuf = nullptr;   // coefficient vectors as doubles
    double* Vbuf = nullptr;   // lattice points
    double* logw = nullptr;   // [chains] current log importance weight
    int32_t* cur = nullptr;   // [chains] which buffer holds the current state
    int32_t* init = nullptr;  // [chains] 0 until _initialize_chain ran
    int64_t* counts = nullptr;    // [chains*2] accepted, total
    int64_t* step_no = nullptr;   // [chains] steps performed so far (Philox stream index)
    int64_t* upos = nullptr;      // [chains] unused (replay positions are per call)
    // which implementation owns the state: 0 = none yet, 1 = ordered kernels, 2 = blocked path
    int state_path = 0;
    // blocked path: all chains step together; current coefficient vectors coordinate-major [npad][cap]
    double* Xcur = nullptr;
    uint8_t* flags = nullptr;     // [cap] where the current state of the chain lives: 0 = Xcur, 1 / 2 = proposal buffer 0 / 1
    int64_t bl_cap = 0;
    int64_t bl_steps = 0;         // steps performed so far by every chain (Philox stream index)
    bool bl_init = false;
    int64_t steps_total = 0;      // host mirror, both paths: Metropolis-Hastings steps every chain has performed
    bool bl_loc_dirty = false;    // chain kernel: some current states may live in a proposal buffer (flags != 0)
    // step-parallel IMHK (lgs_chain.cu): scan outputs and the staging of kept states
    int* sp_src = nullptr;        // [kMaxKeepPerLaunch + 1][cap]
    size_t sp_src_cap = 0;
    double* sp_keep = nullptr;    // [kept slots][cap / 128][npad][132]
    size_t sp_keep_cap = 0;
    // kept states inside a step-by-step launch (lgs_chain.cu, ChainArgs::keep_thin): tiled snapshots of the chain set
    double* Xkeep = nullptr;
    size_t xkeep_cap = 0;
};

// ---- blocked (DMMA) path -------------------------------------------------------------------------------
constexpr int kPB = 32;  // coordinates per panel-kernel block

// Constants of one 32-coordinate block, passed by value as kernel parameters so that the in-block triangle of R
// and the SampleZ constants are read from the constant bank (no shared-memory or global traffic in the serial loop).
struct CoordConst {  // 64 bytes of per-coordinate SampleZ / log-density constants
    double inv_r;    // 1 / R_ii
    double r2;       // R_ii^2                      (||Bx - c||^2 = sum r_ii^2 (x_i - m_i)^2)
    double a;        // 1 / (2 sigma_i^2)
    double q;        // exp(-1/sigma_i^2)
    double sq;       // exp(-1/(2 sigma_i^2))
    double znorm;    // sigma_i sqrt(2 pi)
    double eps1;     // 2 exp(-2 pi^2 sigma_i^2)
    double lognz;    // log(sigma_i sqrt(2 pi))
};
struct PanelParams {
    double nrt[kPB * kPB];  // row i: [kSub shifted triangle entries | rectangle entries for the rows below the sub-block], see klein_panel_kernel
    CoordConst cc[kPB];
    int4 ci[kPB];           // x = kind, y = dmax, z = draw index in the reference's consumption order,
                            // w = 1 for a round-to-mean coordinate (sigma_i < 1e-10, klein.py:201-204) or padding
};

// ---- chain kernel (lgs_chain.cu): constants of one 8-coordinate sub-block, staged into shared memory by a bulk copy
constexpr int kChainRowLd = 132;  // doubles per coordinate row of a 128-sample tile (128 + 4 padding)
constexpr int kRectLd = 12;       // row stride of SubPack::rect (conflict-free LDS.64 A-fragment loads)
struct SubPack {
    CoordConst cc[8];
    int4 ci[8];           // as PanelParams::ci; x carries bit 2 when the coordinate is not eligible for the FP32 screen
    SmoothF32 cf[8];      // FP32 screen constants
    double win[8][8];     // row j: [0..6] rotating-window triangle entries of coordinate base + j
    double rect[64][kRectLd];  // rect[r][j] = -R[lo + r][base + j] / R[lo + r][lo + r] for the rows r < base below the
                               // sub-block (A operand of the panel warps' DMMA update), zero elsewhere; copied up to
                               // the 16-row tile that contains row base - 1
};
static_assert(sizeof(SubPack) == 7552 && sizeof(SubPack) % 16 == 0, "SubPack is copied as raw bytes");

// one batch worth of device workspace for the blocked path
struct BlockedWork {
    double* X = nullptr;    // [npad][cap] coordinate-major centers / coefficients
    double* lp = nullptr;   // [cap] sum (x_i - m_i)^2 / (2 sigma_i^2)
    double* d2 = nullptr;   // [cap] sum r_ii^2 (x_i - m_i)^2
    double* lz = nullptr;   // [cap] sum log rho_{sigma_i,m_i}(Z)
    int64_t cap = 0;        // samples per batch, multiple of 128
};

// optional per-launch timing of the blocked path (lgs_profile_enable / lgs_profile_read)
struct ProfRec {
    int cls;
    cudaEvent_t e0, e1;
};
struct Prof {
    bool on = false;
    std::vector<ProfRec> recs;
    std::vector<cudaEvent_t> pool;  // recycled events
};

// brackets one launch with events when profiling is on (class ids: LGS_PROF_* in include/lgs_b200.h)
struct ProfScope {
    Prof& p;
    cudaStream_t st;
    cudaEvent_t e1 = nullptr;
    ProfScope(Prof& prof, int cls, cudaStream_t s) : p(prof), st(s) {
        if (!p.on) return;
        cudaEvent_t ev[2];
        for (auto& e : ev) {
            if (!p.pool.empty()) { e = p.pool.back(); p.pool.pop_back(); }
            else cudaEventCreate(&e);
        }
        cudaEventRecord(ev[0], st);
        e1 = ev[1];
        p.recs.push_back(ProfRec{cls, ev[0], ev[1]});
    }
    ~ProfScope() { if (e1) cudaEventRecord(e1, st); }
};

// chain kernel (lgs_chain.cu): packed operands and tiled workspaces
struct ChainCtx {
    bool ready = false;
    int npad = 0;                 // n padded to a multiple of 64
    double* Rpk = nullptr;        // [npad/64][npad/16][64][20] shared-memory images of the 64 x 16 slivers of R
    double* Bpk = nullptr;        // same for the basis
    double* tp = nullptr;         // [npad] t_i / R_ii
    bool t_zero = true;           // t = Q^T c = 0
    // histogram attached to the lattice-point epilogue (lgs_attach_histogram); hist == nullptr: none
    int16_t* hist_map = nullptr;  // [npad] tracked coordinate -> its row of bins, -1 otherwise
    unsigned long long* hist = nullptr;
    double hist_lo = 0.0, hist_inv_width = 0.0;
    int hist_nbins = 0;
    // set around launches that write lattice points which are not NEW samples (state polls) or that serve a host
    // route (lgs_klein_sample_host counts on neither of its transports): the epilogue then leaves the histogram alone
    bool hist_pause = false;
    double* sig = nullptr;        // [npad] sigma_i (the replay instantiations' table sampler takes sigma_i itself)
    int precision = 10, max_range = 1000;
    SubPack* subs = nullptr;      // [npad/8]
    double* wtab = nullptr;       // tables of the wide truncated windows (lgs_device.cuh samplez_wide_table)
    int64_t wtab_ld = 0;          // doubles per kind-3 coordinate: 7 (dmax + 1)
    BlockedWork work[3];          // tiled: X = [cap/128][npad][132]; [0] Klein, [1]/[2] IMHK proposals
    int64_t batch = 0;            // samples per full batch: sm_count x 128 x 9
    bool wide = false;            // some coordinate has a wide truncated window (kind 3): the WIDE kernel instantiations
    void* pt_buf = nullptr;       // global lattice-point queue of a launch: [slots handed out, next item, list ...]
    size_t pt_cap = 0;
};

struct BlockedCtx {
    bool ready = false;
    Prof prof;
    ChainCtx chain;
    int use_chain = 1;                 // LGS_CHAIN=0 selects the multi-kernel v1 path (development A/B)
    int n = 0, npad = 0, mpad = 0;     // npad: coordinates padded to 32; mpad: rows of B padded to 64
    double* Rp = nullptr;              // [npad][npad] R, identity on the padding
    double* Bp = nullptr;              // [mpad][npad] basis, zero padding
    double* tp = nullptr;              // [npad] Q^T c, zero padding
    double sum_lognorm = 0.0;          // sum_i 0.5 log(2 pi) + log sigma_i   (klein.py:262)
    double two_sigma2 = 0.0;
    std::vector<PanelParams> panels;   // npad / 32 blocks, index = i0 / 32
    BlockedWork work[3];               // [0] Klein batches, [1] and [2] IMHK proposals of even / odd steps
    int sm_count = 0;
    int64_t wave = 0;                  // samples in one full wave of 64 x 128 contraction tiles (2 CTAs per SM)
};

// per-coordinate constants computed once on the host by lgs_create()
struct HostConsts {
    std::vector<double> rdiag, sig, lognorm, inv2s2, q, sq, znorm, eps1;
    std::vector<uint8_t> mode;
    std::vector<int> uidx, kind, dmax;
};

struct Scratch {
    double* Xw = nullptr;  // [n][S_cap] coefficient workspace of the ordered Klein kernel
    int64_t cap = 0;
};

}  // namespace lgs

#include "lgs_internal_host.h"

struct lgs_handle {
    int device = 0;
    int path = LGS_PATH_AUTO;
    int n = 0;
    lgs::DeviceCtx ctx{};       // host copy; device pointers inside
    void* arena = nullptr;      // one allocation holding every per-sampler constant
    lgs::Scratch scratch;
    lgs::ChainSet chains;
    lgs::BlockedCtx blocked;
    // staging for the *_host entry points
    void* stage_dev = nullptr;
    size_t stage_cap = 0;
    int32_t* hist_dims = nullptr;  // device copy of the tracked coordinates of the last lgs_histogram call
    double* diag_scratch = nullptr;  // series-major copy of the tracked coordinates (lgs_chain_diagnostics)
    size_t diag_cap = 0;
    cudaStream_t host_compute = nullptr, host_copy = nullptr;  // compute / transfer overlap of the *_host calls
    cudaEvent_t ev_done[2] = {}, ev_free[2] = {};
    // packed transport of lattice points to host buffers (integer bases only; lgs_hostpack.cpp)
    bool pack_ok = false;             // every basis entry is an integer => every lattice point is
    void* pack_dev = nullptr;         // device ring: kPackDev slots of {fp64 chunk, int16 chunk + flag}
    size_t pack_dev_cap = 0;
    void* pack_host = nullptr;        // pinned host ring: kPackHost slots of {int16 chunk + flag}
    size_t pack_host_cap = 0;
    cudaEvent_t pk_done[3] = {}, pk_copied[3] = {};
    lgs::HostPool* pool = nullptr;
    // zero-copy variant (chain kernel, n % 8 == 0): the kernel writes int16 lattice points into mapped pinned memory
    void* zc_v16[2] = {};             // mapped pinned: [batch][n] int16, double-buffered over batches
    unsigned* zc_done[2] = {};        // mapped pinned: per-tile completion epochs
    unsigned* zc_cnt = nullptr;       // device: [2][tiles] arrival counters
    int* zc_ovf = nullptr;            // device: [2]
    size_t zc_cap = 0;                // samples per buffer
    unsigned zc_epoch = 0;
    cudaEvent_t zc_ev[2] = {};
    // streamed variant: int16 in device memory, groups of tiles moved by the copy engine behind cuStreamWaitValue32
    void* zs_dev[2] = {};             // device: [batch][n] int16
    unsigned* zs_gcnt = nullptr;      // device: [2][groups]
    cudaEvent_t zs_reset[2] = {}, zs_copied[2] = {};
    // pinned bounce ring for fp64 / int64 results that go to pageable host memory (staged_d2h in lgs_capi.cu)
    void* bounce[2] = {};
    cudaEvent_t bounce_ev[2] = {};
    std::string err;
};

namespace lgs {

// lgs_ordered.cu  (compiled with -fmad=false: reference rounding order)
cudaError_t launch_klein_ordered(const DeviceCtx& ctx, Scratch& sc, int64_t S, const RngSpec& rng, int samplez,
                                 double* V, int64_t* X, double* logq, int32_t* fallback, cudaStream_t st);
cudaError_t launch_klein_logdensity(const DeviceCtx& ctx, int64_t S, const int64_t* X, double* logq, cudaStream_t st);
cudaError_t launch_imhk_ordered(const DeviceCtx& ctx, ChainSet& cs, int64_t burn_in, int64_t keep, int64_t thin,
                                const RngSpec& rng, int samplez, int weight_mode, double* V, int64_t* X,
                                uint8_t* accept, double* logw_out, int32_t* fallback, cudaStream_t st);
cudaError_t launch_imhk_get_state(const DeviceCtx& ctx, ChainSet& cs, int64_t* counts, int64_t* X, double* V,
                                  double* logw, cudaStream_t st);
cudaError_t launch_imhk_set_state(const DeviceCtx& ctx, ChainSet& cs, const int64_t* X, const double* logw,
                                  const int64_t* counts, int64_t steps_done, cudaStream_t st);
constexpr int kHistChunk = 256;  // tracked coordinates per histogram launch (passed by value as a kernel parameter)
cudaError_t launch_histogram(const DeviceCtx& ctx, const double* V, int64_t S, const int32_t* dims_host, int n_dims,
                             double lo, double width, int nbins, int64_t* hist, cudaStream_t st);
cudaError_t launch_reduce_counts(const ChainSet& cs, int64_t* out2, cudaStream_t st);

// lgs_blocked.cu  (DMMA contractions + panel kernel)
cudaError_t blocked_init(BlockedCtx& bc, const lgs_desc& d, const HostConsts& hc);
void blocked_free(BlockedCtx& bc);
cudaError_t blocked_profile_read(BlockedCtx& bc, double* ms, int64_t* launches, int nclass);
// packed zero-copy output of one chain-kernel batch (lgs_chain.cu, ChainArgs::V16)
struct ChainZeroCopy {
    int16_t* v16;         // mapped pinned host memory, device-side address: [S][n] int16
    unsigned* tile_cnt;   // device: [tiles]
    unsigned* tile_done;  // mapped pinned host memory, device-side address: [tiles]
    int* ovf;             // device
    unsigned epoch;
    unsigned* group_cnt;  // streamed variant (v16 in device memory): device [groups]; nullptr = publish tile_done instead
    int group_tiles;
};
cudaError_t launch_klein_chain_zc(BlockedCtx& bc, int64_t S, uint64_t seed, uint64_t first_id, const ChainZeroCopy& zc,
                                  cudaStream_t st);
cudaError_t launch_klein_blocked(BlockedCtx& bc, int64_t S, uint64_t seed, uint64_t first_id, double* V, int64_t* X,
                                 double* logq, cudaStream_t st);
cudaError_t launch_imhk_blocked(BlockedCtx& bc, ChainSet& cs, int64_t burn_in, int64_t keep, int64_t thin,
                                int weight_mode, double* V, int64_t* X, cudaStream_t st);
cudaError_t launch_imhk_blocked_get_state(BlockedCtx& bc, ChainSet& cs, int64_t* counts, int64_t* X, double* V,
                                          double* logw, cudaStream_t st);

// lgs_diag.cu  (device-side chain diagnostics)
constexpr int kDiagFields = LGS_DIAG_FIELDS;
cudaError_t launch_chain_diagnostics(const double* V, int64_t chains, int64_t keep, int n, const int32_t* dims_dev,
                                     int n_dims, double cwin, int max_lag, double* scratch, double* stats, double* acf,
                                     int acf_len, cudaStream_t st);
cudaError_t launch_gelman_rubin_sums(const double* stats, int64_t chains, int n_dims, double* sums, cudaStream_t st);
// the variant that allocates nothing: work array in a static __device__ array of diag_static_capacity() doubles,
// dims by value; `device` must be current (lgs_chain_diagnostics_static)
int64_t diag_static_capacity();
cudaError_t launch_chain_diagnostics_static(int device, const double* V, int64_t chains, int64_t keep, int n,
                                            const int32_t* dims_host, int n_dims, double cwin, int max_lag,
                                            double* stats, double* acf, int acf_len, cudaStream_t st);

// lgs_chain.cu  (persistent warp-specialised kernel)
// Replay through the chain kernel: the recorded uniform stream (device), `stride` doubles per sample / chain, and the
// optional per-step outputs of lgs_imhk_replay.  nullptr = native Philox.
struct ChainReplay {
    const double* U;
    int64_t stride;
    int32_t* fallback;   // or nullptr
    uint8_t* accept;     // IMHK: [chains][steps] or nullptr
    double* logw;        // IMHK: [chains][steps + 1] or nullptr
};
cudaError_t chain_init(BlockedCtx& bc, const lgs_desc& d, const HostConsts& hc);
void chain_free(BlockedCtx& bc);
cudaError_t launch_klein_chain(BlockedCtx& bc, int64_t S, uint64_t seed, uint64_t first_id, double* V, int64_t* X,
                               double* logq, cudaStream_t st, const ChainReplay* rp = nullptr);
cudaError_t launch_imhk_chain(BlockedCtx& bc, ChainSet& cs, int64_t burn_in, int64_t keep, int64_t thin, int weight_mode,
                              double* V, int64_t* X, cudaStream_t st, const ChainReplay* rp = nullptr);
cudaError_t launch_imhk_chain_get_state(BlockedCtx& bc, ChainSet& cs, int64_t* counts, int64_t* X, double* V,
                                        double* logw, cudaStream_t st);
cudaError_t chain_attach_histogram(BlockedCtx& bc, const int32_t* dims, int n_dims, double lo, double width, int nbins,
                                   int64_t* hist);
cudaError_t launch_imhk_chain_set_state(BlockedCtx& bc, ChainSet& cs, const int64_t* X, const double* logw,
                                        const int64_t* counts, int64_t steps_done, cudaStream_t st);

}  // namespace lgs
