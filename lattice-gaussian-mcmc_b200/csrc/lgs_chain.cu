// Chain kernel: the whole Klein sampler (klein.py:181-220) for a batch of samples in ONE persistent,
// warp-specialised kernel, one CTA per SM.
//
// A CTA owns sample tiles of 128 samples from their first coordinate to their lattice point.  Roles:
//   * 8 DMMA warps (two warpgroups): consume a ring of shared-memory stages, each a 64 x 16 sliver of R (or of the
//     basis) and the matching 16 x 128 sliver of the tile's coefficient matrix, with mma.sync m16n8k8 f64; the
//     epilogue writes the centers of a 64-coordinate block pair (t - R[rows, done] X[done, tile]) or 64 rows of
//     the tile's lattice points;
//   * 1 producer warp: fills the ring with bulk asynchronous copies (cp.async.bulk = the TMA engine, completion on
//     an mbarrier) and is the scheduler: it hands the DMMA warps the next ready work item -- centers of a tile whose
//     panel warps have finished the pair above (priority), else 64 rows of lattice points of a finished tile;
//   * 3 panel warpgroups, each rotating through one to three tiles ("slots"): one thread per sample walks the 64
//     coordinates of a pair serially (conditional mean -> Philox -> SampleZ -> log-density terms -> in-pair triangle
//     update), hands the tile back to the scheduler and turns to its next slot while the pair's successor is contracted.
// While a tile is in its serial panel phase the DMMA warps work for the other tiles (and on the lattice points of
// finished ones, which every CTA draws from one global queue), so the FP64 tensor pipe stays busy; registers are moved
// from the producer / panel warpgroups to the DMMA warpgroups with setmaxnreg.
//
// HBM layout of the coefficient workspace ("tiled"): Xt[tile][coordinate][132], 128 samples + 4 doubles of padding
// per row, so that a 16-coordinate sliver of a tile is one contiguous 16 896-byte bulk copy whose shared-memory
// image already has the conflict-free row stride of the B fragments.  R and the basis are pre-packed once per
// sampler into the shared-memory image of their 64 x 16 slivers (row stride 20 doubles).
#include "lgs_internal.h"

#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace lgs {

namespace {

// ---------------------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* b, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok)
        : "r"(smem_u32(b)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ bool mbar_test_wait(uint64_t* b, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok)
        : "r"(smem_u32(b)), "r"(parity)
        : "memory");
    return ok != 0;
}
// the same on a precomputed shared-memory address
__device__ __forceinline__ void mbar_wait_u32(uint32_t addr, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
            : "=r"(ok)
            : "r"(addr), "r"(parity)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ void mbar_arrive_u32(uint32_t addr) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(addr) : "memory");
}
__device__ __forceinline__ double lds_f64(uint32_t addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void lds_v4(uint32_t addr, int& a, int& b, int& c, int& d) {
    asm volatile("ld.shared.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr));
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    while (!mbar_try_wait(b, parity)) {
    }
}
// bulk asynchronous copy global -> shared (TMA engine), completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(b))
                 : "memory");
}
// orders this thread's generic-proxy writes before later async-proxy (bulk copy) reads of the same memory
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }
__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void dmma_16x8x8(double (&d)[4], const double (&a)[4], const double (&b)[2]) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
        : "+d"(d[0]), "+d"(d[1]), "+d"(d[2]), "+d"(d[3])
        : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}

// ---------------------------------------------------------------------------------------------------------
// geometry
// ---------------------------------------------------------------------------------------------------------
constexpr int kTile = 128;            // samples per tile
constexpr int kRowLd = kChainRowLd;   // 132: doubles per coordinate row of a tile in HBM and in shared memory
constexpr int kPair = 64;             // coordinates per work item (a pair of 32-blocks of the v1 path)
#ifndef LGS_CHAIN_BK
#define LGS_CHAIN_BK 16
#endif
constexpr int kBK = LGS_CHAIN_BK;     // contraction depth of one stage
constexpr int kLdA = kBK + 4;         // 20: row stride of an A sliver (conflict-free LDS.64 fragment loads)
constexpr int kSubW = 8;              // coordinates per sub-block of the panel (width of the register window)
#ifndef LGS_CHAIN_TPW
#define LGS_CHAIN_TPW 3   // most tiles a panel warpgroup rotates through (ChainArgs::tpw chooses per launch)
#endif
#ifndef LGS_CHAIN_STAGES
#define LGS_CHAIN_STAGES 5
#endif
constexpr int kStages = LGS_CHAIN_STAGES;
constexpr int kWgs = 3;                  // panel warpgroups
// Tiles in flight per CTA ("slots"): every panel warpgroup alternates between kTpw tiles, pair by pair -- while the
// centers of one tile's next pair are being contracted it draws a pair of the other, and the DMMA warps always have
// the look-ahead items of 2 x 3 tiles to choose from.  Slot = warpgroup * tpw + k.  (The replay instantiations keep
// one tile per warpgroup: their stream position lives in a register across pairs.)
constexpr int kTpw = LGS_CHAIN_TPW;
constexpr int kMaxSlots = kWgs * kTpw;
constexpr int kDmmaThreads = 256, kProdThreads = 128, kPanelThreads = kWgs * 128;
constexpr int kThreads = kDmmaThreads + kProdThreads + kPanelThreads;  // 768
// thread layout: panel warpgroups first (lowest warp ids), then the producer warpgroup, then the DMMA warpgroups
#ifndef LGS_CHAIN_PANEL_FIRST
#define LGS_CHAIN_PANEL_FIRST 1
#endif
constexpr int kPanelBase = LGS_CHAIN_PANEL_FIRST ? 0 : kDmmaThreads + kProdThreads;
constexpr int kProdBase = LGS_CHAIN_PANEL_FIRST ? kPanelThreads : kDmmaThreads;
constexpr int kDmmaBase = LGS_CHAIN_PANEL_FIRST ? kPanelThreads + kProdThreads : 0;
// registers: 768 x 80 at launch, rebalanced between the roles with setmaxnreg
constexpr int kRegsDmma = 120, kRegsProd = 24, kRegsPanel = 72;
constexpr uint32_t kABytes = kPair * kLdA * 8;   // 10 240
constexpr uint32_t kBBytes = kBK * kRowLd * 8;   // 16 896

// F_NODATA: a stage without operand bytes (the top pair has nothing above it: its centers are t itself);
// F_NOSTORE: no epilogue stores, only the cready signal (wakes a panel warpgroup that has no tile left)
// F_PART_OUT / F_PART_IN: look-ahead.  The centers of pair p are contracted in two items: the columns of the pairs
// above p - 1 as soon as those are drawn (while the tile's panel warps are still busy with pair p - 1; the partial
// result t - acc is parked in the pair's own rows), and the 64 columns of pair p - 1 when its panel is done -- so a
// finished panel waits for a 4-stage item, not for a contraction over every finished coordinate.
enum : int { F_FIRST = 1, F_LAST = 2, F_POINT = 4, F_NOSTORE = 8, F_EXIT = 16, F_NODATA = 32, F_PART_OUT = 64,
             F_PART_IN = 128 };

struct StageHdr {
    int flags;
    int slot;      // centers items: whose tile this is (cready[slot] is signalled by the epilogue)
    int r0;        // first row of the 64-row block, bits 0..29; bit 30: which workspace (Xbuf[0 / 1]) the centers go to
    int tile;
};

// panel_role's bookkeeping, kept in shared memory so that it costs the coordinate loop no registers
struct PanelSlot {
    int tile;        // tile being processed; -1: the slot waits for its next tile; -2: no tile left
    int step;        // Metropolis-Hastings step of the launch (Klein: 0)
    int pair;        // next pair of the step
    uint32_t crp;    // parity of the slot's centers-ready barrier
};
struct PanelIssue {
    uint32_t q_iss;  // constants loads issued so far (load number q goes to buffer q & 1)
    int cur;         // next sub-block of the current job to request
    int end;         // one past the current job's last sub-block
    int nstart;      // first sub-block of the job that follows, or -1 while that depends on a tile assignment
    int nj;          // sub-blocks of the following job already requested
};

struct ChainSmem {
    double A[kStages][kPair][kLdA];
    double B[kStages][kBK][kRowLd];
    SubPack sub[kWgs][2];              // double-buffered constants of the sub-block being drawn (per warpgroup)
    alignas(16) double xs[kWgs][kSubW][kRowLd];    // the sub-block's draws (B operand of the rows-below update), then
                                                   // the next window on its way back to one-thread-per-sample
    StageHdr hdr[kStages];
    uint64_t full[kStages], empty[kStages];
    uint64_t cready[kMaxSlots], pdone[kMaxSlots], cfull[kWgs][2];
    int slot_tile[kMaxSlots];
    PanelSlot pst[kWgs][4][kTpw];         // per panel warp: where each of its warpgroup's slots stands (panel_role)
    PanelIssue iss[kWgs];              // per warpgroup: the constants loads in flight (thread 0 of the warpgroup only)
    // packed lattice-point output (ChainArgs::V16): per DMMA warp, its 32 samples x 32 rows as int16, one 80-byte
    // row per sample so that a lane reads its sample's 64 bytes as four conflict-free 16-byte words
    alignas(16) int16_t pk16[kDmmaThreads / 32][32][40];
};
static_assert(sizeof(ChainSmem) <= 232448, "ChainSmem exceeds the 227 KB of shared memory a CTA can opt into");

struct ChainArgs {
    const double* Rpk;      // packed slivers of R: [npad/64][npad/16][64][20]
    const double* Bpk;      // packed slivers of the basis, same shape
    const double* tvec;     // [npad] t_i / R_ii
    int t_zero;             // t = 0 (center c = 0)
    const SubPack* subs;    // [npad/8] sub-block constants, index = first coordinate / 8
    double* Xbuf[2];        // tiled coefficient workspace(s): Klein uses [0]; an IMHK step with Philox tag g draws its
                            // proposal into Xbuf[g & 1]
    double* lp;             // [cap] per-sample accumulators (outputs)
    double* d2;
    double* lz;
    double* V;              // lattice points [S][ldv] or nullptr
    int64_t ldv;
    // Packed zero-copy output (lgs_klein_sample_host on an integer basis, n % 8 == 0): lattice points as int16
    // [S][n_valid] written straight into MAPPED PINNED HOST memory by the lattice-point epilogue, 64 contiguous bytes
    // per sample and warp; tile_cnt[tile] (device) counts the DMMA warps that have stored the tile's last rows, and the warp that
    // completes a tile publishes tile_done[tile] = epoch (mapped host memory) after a system-scope fence, so host
    // threads widen finished tiles into the caller's float64 buffer while the kernel is still running.
    int16_t* V16;
    unsigned* tile_cnt;
    unsigned* tile_done;
    int* ovf;               // set if a value does not survive the trip through int16 (the host then re-runs in fp64)
    unsigned tile_target, epoch;
    // Streamed variant of the same (preferred: stores into host memory throttle the DMMA warps): V16 is DEVICE memory
    // and a finished tile is counted in group_cnt[tile / group_tiles] (device); the copy stream sits on
    // cuStreamWaitValue32(group_cnt[g] >= tiles of the group), then copies the group's int16 block to pinned host memory
    // with the copy engine and writes the group's completion word for the host pool -- all while the kernel runs.
    unsigned* group_cnt;
    int group_tiles;
    int64_t S;
    int npad;               // multiple of 64
    int n_valid;            // n
    int ntiles;
    int want_lz;
    uint64_t seed, first_id;
    uint32_t tag;
    double inv_two_sigma2;  // 1 / (2 sigma^2)
    unsigned long long* stats;  // development counters (LGS_CHAIN_STATS=1), 12 per CTA, or nullptr
    // IMHK (imhk.py:126-177): nsteps Metropolis-Hastings steps per chain in this launch, tags tag .. tag + nsteps - 1
    int nsteps;             // Klein: 1
    int first_is_init;      // the first step is _initialize_chain (tag 0): no accept test, no count
    int exact;              // LGS_WEIGHT_EXACT
    int flush;              // after the last step every chain's current state is moved into Xcur (and, with V, its
                            // lattice point is contracted from there)
    double sum_lognorm;
    double* Xcur;           // tiled current states
    uint8_t* loc;           // [cap] where a chain's current state lives: 0 = Xcur, 1 / 2 = Xbuf[0] / Xbuf[1]
    double* lw_cur;         // [cap] current log importance weight
    long long* counts;      // [cap][2] accepted, total
    // Step-parallel IMHK (launch_imhk_chain, few chains): the proposals of an independent Metropolis-Hastings chain do
    // not depend on its state (imhk.py:150 draws y from the Klein sampler alone), so the proposals of MANY steps of
    // the same chains are drawn as one Klein-mode batch -- tile t holds chains [(t % tiles_per_step) * 128, +128) at step
    // t / tiles_per_step, Philox key (chain id, tag + step) exactly as the step-by-step launch -- and the accept /
    // reject recursion runs afterwards over the per-proposal log-weights (imhk_scan_kernel).  0 = off.
    // Global lattice-point queue (gq != 0).  The lattice points of a finished tile are npairs work items (64 rows
    // each) that stream the tile's whole coefficient block; served per CTA as filler they are spread over the launch and
    // every one of them re-reads the block from DRAM (the blocks in flight, 3 x 148 x 1 MB, do not fit the L2).  With the
    // queue a finished tile is published in pt_list and EVERY CTA claims items from it in order, so at any time the
    // SMs work on the row blocks of the same ~10 tiles: their blocks stay in L2 and are read from DRAM once.  It also
    // balances the tail of a launch.  pt_list[i] = tile + 1 once published (zeroed per launch); pt_ctr[0] = list slots
    // handed out, pt_ctr[1] = next ticket (item = list position * npairs + row block), drawn by fetch-and-add.
    int gq;
    int lookahead;          // centers of a pair in two items (see F_PART_OUT): 0 never, 1 always, 2 when the DMMA warps
                            // would otherwise idle
    int stagger;            // slots started 1 / nslots of a step apart (producer_role), in per cent of that spacing; 0: together
    int tpw;                // tiles per panel warpgroup: 1 .. kTpw
    int* pt_list;
    unsigned* pt_ctr;
    // Kept states inside a launch (IMHK, thin <= 64): the steps keep_first, keep_first + keep_thin, ... (nkeep of them) end
    // with a snapshot of every chain's current state into Xkeep[m] (tiled, keep_stride doubles per snapshot), and the
    // lattice points of snapshot m -- V + m * v_slot_stride -- are contracted from there as filler items WHILE the chains
    // go on to the next steps: what makes a Klein batch fast (two thirds of its DMMA work can run at any time) then holds
    // for IMHK too.  keep_thin = 0: off.
    int keep_thin, keep_first, nkeep;
    double* Xkeep;
    int64_t keep_stride;
    int tiles_per_step;
    int skip0;              // all-padding sub-blocks at the top of the first pair (see panel_role)
    int64_t v_slot_stride;  // lattice points of a step-parallel points launch: slot (= tile / tiles_per_step) stride in V
    // Replay (the REPLAY instantiations; lgs_klein_replay / lgs_imhk_replay on LGS_PATH_BLOCKED): the panel threads take
    // their uniforms from the reference's recorded stream instead of Philox -- unit s reads U[s * u_stride + position],
    // one uniform per drawn coordinate in the order i = n-1 .. 0 (klein.py:181-220), then one for the accept test of a
    // Metropolis-Hastings step (imhk.py:167) -- and draw with the reference's table arithmetic (samplez_table:
    // klein.py:101-179 + numpy choice), so the integers are the reference's except where the DMMA summation order of
    // the conditional mean moves a CDF comparison (counted by the tests, ~1e-15 per draw).
    const double* U;
    int64_t u_stride;
    long long* upos;        // IMHK: [cap] position of every chain inside its stream, carried over the launches of one
                            // call; Klein: nullptr (every sample starts at 0)
    const double* sig;      // [npad] sigma_i = sigma / |R_ii|
    // histogram fused into the lattice-point epilogue (lgs_attach_histogram): bins of the tracked coordinates
    const int16_t* hist_map;    // [npad] row of bins of a tracked coordinate, -1 otherwise; nullptr: no histogram
    unsigned long long* hist;   // [tracked][hist_nbins], accumulated with global atomics
    double hist_lo, hist_inv_width;
    int hist_nbins;
    const double* wtab;     // tables of the wide truncated windows (samplez_wide_table), wtab_ld doubles per kind-3 coordinate
    int64_t wtab_ld;
    int precision, max_range;
    int32_t* fallback;      // counts draws that took the reference's no-uniform fallback (klein.py:176-179), or nullptr
    uint8_t* accept_out;    // IMHK: [S][call_steps] accept bit of every step of the call, or nullptr
    double* logw_out;       // IMHK: [S][call_steps + 1] log weight at entry and of every proposal, or nullptr
    int64_t call_steps;     // Metropolis-Hastings steps of the whole call (row length of accept_out)
    int64_t step_off;       // steps of the call performed by earlier launches
};

// Development cycle counters (read with LGS_CHAIN_STATS=1 in the environment) exist only in a
// -DLGS_CHAIN_STATS_BUILD=1 build (tools/build_variant.sh): even predicated-off clock reads cost the panel loop
// 0.5 % (Klein) to 2.7 % (IMHK).  LGS_CHAIN_STATS_DETAIL=1 adds the tile / constants / sub-block barrier waits.
#ifndef LGS_CHAIN_STATS_BUILD
#define LGS_CHAIN_STATS_BUILD 0
#endif
constexpr bool kStatsBuild = LGS_CHAIN_STATS_BUILD != 0;
#ifndef LGS_CHAIN_STATS_DETAIL
#define LGS_CHAIN_STATS_DETAIL 0
#endif
#ifndef LGS_ZC_NOFENCE
#define LGS_ZC_NOFENCE 0   // development: time the zero-copy epilogue without its system-scope fences (NOT correct)
#endif
// An integer-valued double with |v| < 2^15 as an int, by integer arithmetic on its bit pattern -- the DMMA warps own the
// FP64 pipe, and F2I / I2F / DSETP would each take one of its slots.  `bad` is raised for anything else (a fraction,
// |v| >= 2^15, inf, nan), in which case the host re-runs the batch through the fp64 route.
__device__ __forceinline__ int pack_i16(double v, bool& bad) {
    const int a = __double2hiint(v);
    const unsigned b = (unsigned)__double2loint(v);
    const int e = ((a >> 20) & 0x7ff) - 1023;
    const unsigned m = ((unsigned)a & 0xfffffu) | 0x100000u;   // leading 21 bits of the significand
    const bool in = (unsigned)e <= 14u;
    const unsigned sh = in ? (unsigned)(20 - e) : 20u;
    const unsigned mag = in ? (m >> sh) : 0u;
    const bool exact = in ? ((m & ((1u << sh) - 1u)) == 0u) : (((unsigned)a << 1) == 0u);   // else branch: +-0
    bad |= !exact || b != 0u;
    return a < 0 ? -(int)mag : (int)mag;
}

// ---------------------------------------------------------------------------------------------------------
// DMMA warps: consume stages until the producer sends F_EXIT
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void dmma_role(ChainSmem& sm, const ChainArgs& P) {
    const int tid = threadIdx.x - kDmmaBase;
    const int warp = tid >> 5, lane = tid & 31;
    const int wm = warp >> 2, wn = warp & 3;
    const int g = lane >> 2, t = lane & 3;
    double acc[2][4][4];
    // Shared-memory addresses as 32-bit offsets, computed once: the stage loop's bookkeeping (barrier wait, header read,
    // fragment addresses) sits between the last DMMA of a stage and the first of the next on every warp, and derived
    // from the generic `sm` it cost a special-register read, a window-base computation and a re-derivation of the lane's
    // fragment offsets from threadIdx per stage (ncu, profiles/r02e_stall_summary_by_role.txt: 15 % of the DMMA warps'
    // samples, a third of it the genuine wait for a full stage).
    // (made opaque to the compiler, which otherwise re-derives them from %tid inside the loop rather than hold three
    // registers)
    uint32_t sbase = smem_u32(&sm);
    uint32_t a_thr = smem_u32(&sm.A[0][wm * 32 + g][t]);   // + stage * kABytes + (16 mt + 8 h) rows + (kk + 4 c) doubles
    uint32_t b_thr = smem_u32(&sm.B[0][t][wn * 32 + g]);   // + stage * kBBytes + (kk + 4 c) rows + 8 nt doubles
    asm volatile("" : "+r"(sbase), "+r"(a_thr), "+r"(b_thr));
    const uint32_t full0 = sbase + (uint32_t)offsetof(ChainSmem, full), empty0 = sbase + (uint32_t)offsetof(ChainSmem, empty),
                   hdr0 = sbase + (uint32_t)offsetof(ChainSmem, hdr);
    uint32_t stage = 0, par = 0;
    uint32_t it = 0;
    const bool rec = kStatsBuild && P.stats && tid == 0;
    long long t_wait = 0;
    const long long t_begin = rec ? clock64() : 0;
    for (;; it++, stage = (stage + 1 == (uint32_t)kStages) ? 0u : stage + 1, par ^= (stage == 0u) ? 1u : 0u) {
        if (rec) {
            const long long t0 = clock64();
            mbar_wait_u32(full0 + stage * 8u, par);
            t_wait += clock64() - t0;
        } else {
            mbar_wait_u32(full0 + stage * 8u, par);
        }
        StageHdr h;
        lds_v4(hdr0 + stage * (uint32_t)sizeof(StageHdr), h.flags, h.slot, h.r0, h.tile);
        if (h.flags & F_EXIT) break;
        if (h.flags & F_FIRST) {
#pragma unroll
            for (int a = 0; a < 2; a++)
#pragma unroll
                for (int b = 0; b < 4; b++)
#pragma unroll
                    for (int c = 0; c < 4; c++) acc[a][b][c] = 0.0;
        }
        if (!(h.flags & F_NODATA)) {
            const uint32_t as = a_thr + stage * kABytes, bs = b_thr + stage * kBBytes;
#pragma unroll
            for (int kk = 0; kk < kBK; kk += 8) {
                double af[2][4], bf[4][2];
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
                    af[mt][0] = lds_f64(as + (uint32_t)(((mt * 16) * kLdA + kk) * 8));
                    af[mt][1] = lds_f64(as + (uint32_t)(((mt * 16 + 8) * kLdA + kk) * 8));
                    af[mt][2] = lds_f64(as + (uint32_t)(((mt * 16) * kLdA + kk + 4) * 8));
                    af[mt][3] = lds_f64(as + (uint32_t)(((mt * 16 + 8) * kLdA + kk + 4) * 8));
                }
#pragma unroll
                for (int nt = 0; nt < 4; nt++) {
                    bf[nt][0] = lds_f64(bs + (uint32_t)((kk * kRowLd + nt * 8) * 8));
                    bf[nt][1] = lds_f64(bs + (uint32_t)(((kk + 4) * kRowLd + nt * 8) * 8));
                }
#pragma unroll
                for (int mt = 0; mt < 2; mt++)
#pragma unroll
                    for (int nt = 0; nt < 4; nt++) dmma_16x8x8(acc[mt][nt], af[mt], bf[nt]);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_u32(empty0 + stage * 8u);
        if (!(h.flags & F_LAST)) continue;
        if (!(h.flags & F_NOSTORE)) {
            // fragment (mt, nt): rows g / g+8 of the 16-row tile, columns 2t, 2t+1 of the 8-column tile
#define LGS_ACC(mt, hh, nt, c) acc[mt][nt][(hh) * 2 + (c)]
            if ((h.flags & F_POINT) && P.V16) {
                const int64_t s0 = (int64_t)(h.tile & 0xffffff) * kTile;
                int16_t (*stg)[40] = sm.pk16[warp];
                bool bad = false;
#pragma unroll
                for (int mt = 0; mt < 2; mt++)
#pragma unroll
                    for (int hh = 0; hh < 2; hh++)
#pragma unroll
                        for (int nt = 0; nt < 4; nt++)
#pragma unroll
                            for (int c = 0; c < 2; c++) {
                                stg[nt * 8 + t * 2 + c][mt * 16 + hh * 8 + g] = (short)pack_i16(LGS_ACC(mt, hh, nt, c), bad);
                            }
                __syncwarp();
                // store j: lane L carries 16 bytes (8 rows) of sample 8 j + L / 4 -- four neighbouring lanes write the
                // 64 contiguous bytes of one sample, i.e. whole 32-byte sectors go out over PCIe
                const int row0 = h.r0 + wm * 32 + (lane & 3) * 8;
                if (row0 + 8 <= P.n_valid) {
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const int sl = j * 8 + (lane >> 2);
                        const int64_t col = s0 + wn * 32 + sl;
                        if (col < P.S)
                            *reinterpret_cast<uint4*>(P.V16 + (size_t)col * P.n_valid + row0) =
                                *reinterpret_cast<const uint4*>(&stg[sl][(lane & 3) * 8]);
                    }
                }
                if (bad) atomicOr(P.ovf, 1);
                __syncwarp();   // the staging rows are rewritten by the warp's next lattice-point item
                if (P.gq || h.r0 + kPair == P.npad) {
                    // Per-CTA order (gq == 0): the tile's last 64 rows (its items arrive in row order): everything this
                    // warp has stored for the tile becomes visible system-wide, then the warp is counted; the eighth
                    // warp publishes.  Global queue (V16 in device memory): the row blocks of a tile finish on
                    // different SMs in any order, so every (row block, warp) is counted after a device-scope fence and
                    // the last of the 8 npairs arrivals publishes behind a system-scope fence.
#if !LGS_ZC_NOFENCE
                    if (P.gq) __threadfence(); else __threadfence_system();
#endif
                    __syncwarp();
                    if (lane == 0) {
                        const unsigned prev = atomicAdd(&P.tile_cnt[h.tile], 1u);
                        if (prev + 1 == P.tile_target) {
#if !LGS_ZC_NOFENCE
                            __threadfence_system();
#endif
                            if (P.group_cnt) atomicAdd(&P.group_cnt[h.tile / P.group_tiles], 1u);
                            else *reinterpret_cast<volatile unsigned*>(&P.tile_done[h.tile]) = P.epoch;
                        }
                    }
                }
            } else if (h.flags & F_POINT) {
                // step-parallel points launch: tile -> (kept slot, chain tile)
                // ... or tile | kept-state number << 24 (ChainArgs::keep_thin)
                const int tps = P.tiles_per_step;
                const int ptile = h.tile & 0xffffff, pm = h.tile >> 24;
                const int64_t s0 = (int64_t)(tps ? ptile % tps : ptile) * kTile;
                double* __restrict__ Vb = P.V + (tps ? (size_t)(ptile / tps) : (size_t)pm) * P.v_slot_stride;
#pragma unroll
                for (int mt = 0; mt < 2; mt++)
#pragma unroll
                    for (int hh = 0; hh < 2; hh++) {
                        const int row = h.r0 + wm * 32 + mt * 16 + g + hh * 8;
                        if (row >= P.n_valid) continue;
#pragma unroll
                        for (int nt = 0; nt < 4; nt++) {
                            const int64_t col = s0 + wn * 32 + nt * 8 + t * 2;
                            if (col < P.S) Vb[(size_t)col * P.ldv + row] = LGS_ACC(mt, hh, nt, 0);
                            if (col + 1 < P.S) Vb[(size_t)(col + 1) * P.ldv + row] = LGS_ACC(mt, hh, nt, 1);
                        }
                    }
                if (P.hist_map) {
                    // marginal histograms of the tracked coordinates (src/diagnostics/convergence_diag.py:15-72), counted
                    // while the lattice points are in registers: same bins and clamping as histogram_kernel
#pragma unroll
                    for (int mt = 0; mt < 2; mt++)
#pragma unroll
                        for (int hh = 0; hh < 2; hh++) {
                            const int row = h.r0 + wm * 32 + mt * 16 + g + hh * 8;
                            if (row >= P.n_valid) continue;
                            const int m = P.hist_map[row];
                            if (m < 0) continue;
                            unsigned long long* hb = P.hist + (size_t)m * P.hist_nbins;
#pragma unroll
                            for (int nt = 0; nt < 4; nt++)
#pragma unroll
                                for (int c = 0; c < 2; c++) {
                                    const int64_t col = s0 + wn * 32 + nt * 8 + t * 2 + c;
                                    if (col >= P.S) continue;
                                    const double b = floor((LGS_ACC(mt, hh, nt, c) - P.hist_lo) * P.hist_inv_width);
                                    const int bi = b < 0.0 ? 0 : (b >= (double)P.hist_nbins ? P.hist_nbins - 1 : (int)b);
                                    atomicAdd(hb + bi, 1ull);
                                }
                        }
                }
            } else {
                double* xt = P.Xbuf[(h.r0 >> 30) & 1] + (size_t)h.tile * P.npad * kRowLd;
                const int r0 = h.r0 & 0x3fffffff;
#pragma unroll
                for (int mt = 0; mt < 2; mt++)
#pragma unroll
                    for (int hh = 0; hh < 2; hh++) {
                        const int row = r0 + wm * 32 + mt * 16 + g + hh * 8;
                        // centers are kept divided by their R_ii: the rows of the packed R and t already are (chain_init),
                        // so the epilogue has nothing to load for the usual center c = 0 (t = 0)
                        const double tv = P.t_zero ? 0.0 : P.tvec[row];
#pragma unroll
                        for (int nt = 0; nt < 4; nt++) {
                            const int col = wn * 32 + nt * 8 + t * 2;
                            double2* dst = reinterpret_cast<double2*>(xt + (size_t)row * kRowLd + col);
                            double2 o = make_double2(tv, tv);
                            if (h.flags & F_PART_IN) o = *dst;  // written by these same threads (F_PART_OUT item)
                            o.x -= LGS_ACC(mt, hh, nt, 0);
                            o.y -= LGS_ACC(mt, hh, nt, 1);
                            *dst = o;
                        }
                    }
            }
        }
        if (!(h.flags & (F_POINT | F_PART_OUT))) {
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.cready[h.slot]);
        }
    }
    if (rec) {
        P.stats[blockIdx.x * 12 + 0] = (unsigned long long)(clock64() - t_begin);
        P.stats[blockIdx.x * 12 + 1] = (unsigned long long)t_wait;
        P.stats[blockIdx.x * 12 + 2] = it;
    }
}

// ---------------------------------------------------------------------------------------------------------
// producer / scheduler: one thread
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void producer_role(ChainSmem& sm, const ChainArgs& P, const int nslots) {
    const int G = gridDim.x;
    const int npairs = P.npad / kPair;
    const int nkt = P.npad / kBK;
    uint32_t it = 0;
    long long t_empty = 0, t_idle = 0;
    const bool rec = kStatsBuild && P.stats != nullptr;
    auto push = [&](int flags, int slot, int r0, int tile, const double* asrc, const double* bsrc) {
        const int stage = it % kStages;
        if (rec) {
            const long long t0 = clock64();
            mbar_wait(&sm.empty[stage], ((it / kStages) & 1) ^ 1);
            t_empty += clock64() - t0;
        } else {
            mbar_wait(&sm.empty[stage], ((it / kStages) & 1) ^ 1);
        }
        sm.hdr[stage] = StageHdr{flags, slot, r0, tile};
        if (flags & (F_NODATA | F_EXIT)) {
            mbar_arrive(&sm.full[stage]);
        } else {
            mbar_arrive_expect_tx(&sm.full[stage], kABytes + kBBytes);
            bulk_g2s(&sm.A[stage][0][0], asrc, kABytes, &sm.full[stage]);
            bulk_g2s(&sm.B[stage][0][0], bsrc, kBBytes, &sm.full[stage]);
        }
        it++;
    };
    // centers of pair p (0 = top) of a tile, rows [lo, lo + 64):
    //   part A (p >= 2): the columns of pairs 0 .. p - 2, issued ahead (look-ahead), parks t - acc in the rows;
    //   part B (p >= 1): the 64 columns of pair p - 1, finishes the centers and wakes the panel warpgroup.
    auto issue_range = [&](int flags, int slot, int lo, int tile, int xsel, int c0, int c1) {
        const int rt = lo / kPair;
        const int kt0 = c0 / kBK, kt1 = c1 / kBK;
        const double* a = P.Rpk + ((size_t)rt * nkt + kt0) * (kPair * kLdA);
        const double* b = P.Xbuf[xsel] + (size_t)tile * P.npad * kRowLd + (size_t)kt0 * kBK * kRowLd;
        for (int kt = kt0; kt < kt1; kt++, a += kPair * kLdA, b += kBK * kRowLd)
            push(flags | (kt == kt0 ? F_FIRST : 0) | (kt == kt1 - 1 ? F_LAST : 0), slot, lo | (xsel << 30), tile, a, b);
    };
    auto issue_A = [&](int slot, int p, int tile, int xsel) {
        const int lo = P.npad - kPair * (p + 1);
        issue_range(F_PART_OUT, slot, lo, tile, xsel, lo + 2 * kPair, P.npad);
    };
    auto issue_B = [&](int slot, int p, int tile, int xsel) {
        const int lo = P.npad - kPair * (p + 1);
        if (p == 0) {
            push(F_FIRST | F_LAST | F_NODATA, slot, lo | (xsel << 30), tile, nullptr, nullptr);
            return;
        }
        issue_range(p >= 2 ? F_PART_IN : 0, slot, lo, tile, xsel, lo + kPair, lo + 2 * kPair);
    };
    // 64 rows of the lattice points of a finished tile (Klein: from its workspace; IMHK: from the current states)
    // (`code` = tile | kept-state number << 24, see ChainArgs::keep_thin; the epilogue decodes it from the header)
    auto issue_point = [&](int code, int rt) {
        const int tile = code & 0xffffff, m = code >> 24;
        const double* a = P.Bpk + (size_t)rt * nkt * (kPair * kLdA);
        const double* src = P.keep_thin ? P.Xkeep + (size_t)m * P.keep_stride : (P.flush ? P.Xcur : P.Xbuf[0]);
        const double* b = src + (size_t)tile * P.npad * kRowLd;
        const int nkv = (P.n_valid + kBK - 1) / kBK;   // k-tiles beyond the last coordinate multiply zeros
        for (int kt = 0; kt < nkv; kt++, a += kPair * kLdA, b += kBK * kRowLd)
            push(F_POINT | (kt == 0 ? F_FIRST : 0) | (kt == nkv - 1 ? F_LAST : 0), 0, rt * kPair, code, a, b);
    };

    // global lattice-point queue (ChainArgs::gq)
    const unsigned gq_total = (unsigned)P.ntiles * (unsigned)npairs * (unsigned)(P.keep_thin ? P.nkeep : 1);
    auto gq_publish = [&](int tile) {
        // the tile's coefficients were stored by the panel threads (generic proxy, fence.proxy.async, pdone observed
        // above); the release below orders them before the list entry for whichever CTA claims the tile's items
        const unsigned slot = atomicAdd(&P.pt_ctr[0], 1u);
        __threadfence();
        st_release_gpu(&P.pt_list[slot], tile + 1);
    };
    // Tickets: an item is claimed with one fetch-and-add (it cannot fail, however many CTAs claim at once -- a
    // compare-and-swap loop here retried for most of its attempts: 148 schedulers claim ~2 items per microsecond in
    // total and a claim takes longer than that).  The ticket may be ahead of the published tiles; it is then kept
    // (gq_ticket) and re-examined on later passes of the scheduling loop, which goes on serving this CTA's own
    // panels in between -- so a pending ticket never holds up the tile it waits for.
    unsigned gq_ticket = 0xffffffffu;   // none
    bool gq_drained = !(P.gq && (P.V || P.V16));   // this CTA has drawn a ticket past the last item
    auto gq_claim = [&](int& tile, int& rt) -> bool {
        if (gq_drained) return false;
        if (gq_ticket == 0xffffffffu) {
            gq_ticket = atomicAdd(&P.pt_ctr[1], 1u);
            if (gq_ticket >= gq_total) {
                gq_drained = true;
                return false;
            }
        }
        const int t = ld_acquire_gpu(&P.pt_list[gq_ticket / (unsigned)npairs]);
        if (t == 0) return false;       // the ticket's tile is not finished yet
        tile = t - 1;
        rt = (int)(gq_ticket % (unsigned)npairs);
        gq_ticket = 0xffffffffu;
        fence_proxy_async();            // the claimed tile's rows are read next by this thread's bulk copies
        return true;
    };
    int next_tile = blockIdx.x;
    int s_pair[kMaxSlots] = {}, s_tile[kMaxSlots] = {};   // s_pair: pairs whose centers have been completed (part B issued)
    int s_ahead[kMaxSlots] = {};          // next pair whose part A has not been issued yet
    int s_step[kMaxSlots] = {};           // Metropolis-Hastings step of the launch the slot's tile is in (Klein: 0)
    uint32_t pd_par[kMaxSlots];
    bool s_active[kMaxSlots];
    int n_active = 0;
    auto xsel_of = [&](int step) { return (int)((P.tag + (uint32_t)step) & 1u); };
    auto start_step = [&](int s) {
        s_pair[s] = 1;
        s_ahead[s] = 2;
        issue_B(s, 0, s_tile[s], xsel_of(s_step[s]));
    };
    auto assign = [&](int s) {
        if (next_tile < P.ntiles) {
            s_tile[s] = next_tile;
            next_tile += G;
            s_step[s] = 0;
            sm.slot_tile[s] = s_tile[s];
            if (P.nsteps > 0) {
                start_step(s);
            } else {  // no step, only the flush (+ lattice points) of the current states: wake the panel warpgroup
                s_pair[s] = npairs;
                s_ahead[s] = npairs;   // nothing to look ahead to
                push(F_FIRST | F_LAST | F_NODATA | F_NOSTORE, s, 0, s_tile[s], nullptr, nullptr);
            }
        } else {
            s_active[s] = false;
            n_active--;
            sm.slot_tile[s] = -1;
            push(F_FIRST | F_LAST | F_NODATA | F_NOSTORE, s, 0, -1, nullptr, nullptr);
        }
    };
    // Staggered start.  The centers of pair p cost 4 p stages while a panel takes the same time for every pair, so
    // tiles that walk their pairs in phase starve the DMMA warps at the top of a step and queue behind them at the
    // bottom.  The slots are therefore started 1 / nslots of a step apart -- warpgroup w's first slot once slot 0 has
    // drawn w npairs / nslots pairs, a warpgroup's slot k once its first has drawn k npairs / tpw -- and the rotation
    // inside a warpgroup and the equal length of all tiles keep them there for the rest of the launch.  A warpgroup's
    // slot that is not started yet says so at each of its turns (slot_tile = -3) and the warpgroup goes on to the next.
    const int tpw = nslots / kWgs;
    const bool stagger = P.stagger && npairs >= 12 && P.nsteps > 0 && P.ntiles >= nslots * G;
    int s_jobs[kMaxSlots] = {};           // pairs the slot's panel has drawn since the launch began
    bool s_defer[kMaxSlots] = {};
    auto not_yet = [&](int s) {
        sm.slot_tile[s] = -3;
        push(F_FIRST | F_LAST | F_NODATA | F_NOSTORE, s, 0, -1, nullptr, nullptr);
    };
    for (int s = 0; s < nslots; s++) {
        pd_par[s] = 0;
        s_active[s] = true;
        n_active++;
        if (stagger && s > 0) s_defer[s] = true;
        else assign(s);
    }
    // finished tiles whose lattice points are still to be contracted (filler work for the DMMA warps)
    constexpr int kQ = 32;
    int pq[kQ];
    int pq_head = 0, pq_cnt = 0, pq_rt = 0;
    int rr = 0, ra = 0;
    for (;;) {
        bool progressed = false;
        if (pq_cnt < kQ) {
            for (int i = 0; i < nslots; i++) {
                const int s = (rr + i) % nslots;
                if (!s_active[s] || s_defer[s] || !mbar_test_wait(&sm.pdone[s], pd_par[s])) continue;
                pd_par[s] ^= 1;
                if (stagger) {
                    s_jobs[s]++;
                    if (s == 0)
                        for (int w = 1; w < kWgs; w++)
                            if (s_defer[w * tpw] && s_jobs[0] >= w * npairs * P.stagger / (100 * nslots)) {
                                s_defer[w * tpw] = false;
                                assign(w * tpw);
                            }
                    // the warpgroup turns to its following slots next: those not started yet are started now (the
                    // warpgroup's first slot has drawn their share of a step, or there is no tile left to wait for)
                    // or send the warpgroup on (not_yet), up to the first one that has a tile
                    const int w0 = s / tpw * tpw;
                    for (int j = 1; j < tpw; j++) {
                        const int o = w0 + (s - w0 + j) % tpw;
                        if (!s_active[o]) continue;
                        if (!s_defer[o]) break;
                        if (s_jobs[w0] >= (o - w0) * npairs * P.stagger / (100 * tpw) || next_tile >= P.ntiles) {
                            s_defer[o] = false;
                            assign(o);
                            break;
                        }
                        not_yet(o);
                    }
                }
                if (s_pair[s] == npairs) {
                    if (P.keep_thin && P.V && s_step[s] >= P.keep_first && (s_step[s] - P.keep_first) % P.keep_thin == 0) {
                        // a kept step: the panel has just stored the tile's snapshot; its lattice points are filler now
                        const int code = s_tile[s] | ((s_step[s] - P.keep_first) / P.keep_thin) << 24;
                        if (P.gq) gq_publish(code);
                        else { pq[(pq_head + pq_cnt) % kQ] = code; pq_cnt++; }
                    }
                    if (s_step[s] + 1 < P.nsteps) {  // next Metropolis-Hastings step of the same chains
                        s_step[s]++;
                        start_step(s);
                    } else {
                        if ((P.V || P.V16) && !P.keep_thin) {
                            if (P.gq) gq_publish(s_tile[s]);
                            else { pq[(pq_head + pq_cnt) % kQ] = s_tile[s]; pq_cnt++; }
                        }
                        assign(s);
                    }
                } else {
                    const int pr = s_pair[s];  // its panel may start once part B is through
                    const int xs_ = xsel_of(s_step[s]);
                    if (P.lookahead == 1 && pr >= 2 && s_ahead[s] <= pr) {   // eager look-ahead: every pair in two items
                        issue_A(s, pr, s_tile[s], xs_);
                        s_ahead[s] = pr + 1;
                    }
                    if (pr >= 2 && s_ahead[s] > pr) {
                        issue_B(s, pr, s_tile[s], xs_);   // part A is parked in the pair's rows
                    } else {
                        // one item over every column above the pair: no parked partial sums, one epilogue
                        const int lo = P.npad - kPair * (pr + 1);
                        issue_range(0, s, lo, s_tile[s], xs_, lo + kPair, P.npad);
                        s_ahead[s] = pr + 1;
                    }
                    s_pair[s]++;
                }
                rr = s + 1;
                progressed = true;
                break;
            }
        }
        if (progressed) continue;
        // Look-ahead: part A of pair s_pair (its columns, the pairs up to s_pair - 2, are all drawn) while the slot's
        // panel warps work on pair s_pair - 1.  It costs a second epilogue and the re-read of the parked sums by the
        // DMMA warps (8.7 % of their time when every pair is split, ncu).  With two tiles per panel warpgroup a
        // finished pair no longer waits for its centers -- the warpgroup draws its other tile meanwhile -- and the
        // split costs more than it hides: default 0 (one item per pair).  1 = every pair split (the round-1 kernel),
        // 2 = split only when the ring has run dry and there are no lattice points to contract either; measured at
        // n = 1024 / 512: IMHK 73.4 / 73.9 / 73.4 ms per 10 steps, Klein without points 4.61 / 4.95 / 4.98 ms at n = 512.
        auto look_ahead = [&]() -> bool {
            for (int i = 0; i < nslots; i++) {
                const int s = (ra + i) % nslots;
                if (!s_active[s] || s_defer[s] || s_ahead[s] > s_pair[s] || s_ahead[s] >= npairs) continue;
                issue_A(s, s_ahead[s], s_tile[s], xsel_of(s_step[s]));
                s_ahead[s]++;
                ra = s + 1;
                return true;
            }
            return false;
        };
        if (P.lookahead == 1 && look_ahead()) continue;
        if (pq_cnt > 0) {
            issue_point(pq[pq_head], pq_rt);
            if (++pq_rt == npairs) {
                pq_rt = 0;
                pq_head = (pq_head + 1) % kQ;
                pq_cnt--;
            }
            continue;
        }
        const bool use_gq = P.gq && (P.V || P.V16);
        if (use_gq) {
            int gt, grt;
            if (gq_claim(gt, grt)) {
                issue_point(gt, grt);
                continue;
            }
        }
        if (P.lookahead == 2) {
            // every stage but the last one pushed has been consumed
            const uint32_t j = it - 2u;
            const bool ring_low = it < 2u || mbar_test_wait(&sm.empty[j % kStages], (j / kStages) & 1);
            if (ring_low && look_ahead()) continue;
        }
        // leave when this CTA has no tile left and (global queue) the queue has no item left for it
        if (n_active == 0 && (!use_gq || gq_drained)) break;
        if (rec) {
            const long long t0 = clock64();
            __nanosleep(200);
            t_idle += clock64() - t0;
        } else {
            __nanosleep(200);
        }
    }
    push(F_EXIT, 0, 0, -1, nullptr, nullptr);
    if (rec) {
        P.stats[blockIdx.x * 12 + 3] = (unsigned long long)t_empty;
        P.stats[blockIdx.x * 12 + 4] = (unsigned long long)t_idle;
    }
}

// ---------------------------------------------------------------------------------------------------------
// panel warpgroup: one thread per sample of the slot's tile
// ---------------------------------------------------------------------------------------------------------
template <bool WIDE>
__device__ __noinline__ double chain_samplez_fast_call(double mean, double inv2s2, double q, double sq, double znorm,
                                                       double eps1, int kind, int dmax, double u, double* zsum) {
    FastZ fz{inv2s2, q, sq, znorm, eps1, kind, dmax};
    return (double)samplez_fast<WIDE>(mean, fz, u, zsum);
}

// (the coordinate's CoordConst carries the normaliser's coefficients in q, sq, znorm, eps1, lognz: chain_init)
__device__ __noinline__ double chain_samplez_wide_table_call(double mean, const CoordConst* cc, float sqrt_a, int dmax,
                                                             double u, const double* tab, double* zsum) {
    const double zc[5] = {cc->q, cc->sq, cc->znorm, cc->eps1, cc->lognz};
    return (double)samplez_wide_table(mean, cc->a, sqrt_a, dmax, u, tab, zc, zsum);
}

// Replay draw: the reference's table SampleZ (klein.py:101-179) on the recorded uniform; *lse (exact weights) is the
// table's log-sum-exp, the log of the window mass (imhk: Wang-Ling weight).  Returns false when the reference's
// no-uniform fallback was taken (klein.py:176-179): the uniform is then not consumed.
template <bool WANT_LZ>
__device__ __noinline__ bool chain_samplez_table_call(double mean, double sig, int precision, int max_range, double u,
                                                      double* x, double* lse) {
    long long xi;
    const bool ok = samplez_table(mean, sig, precision, max_range, u, &xi);
    *x = (double)xi;
    if (WANT_LZ) {
        const TableWindow tw = table_window(mean, sig, precision, max_range);
        const LogpGen lpg{(double)tw.lower, mean, sig};
        *lse = table_lse(lpg, tw.len);
    }
    return ok;
}

// WIDE: the lattice has wide truncated windows (kind 3, closed-form masses); a separate instantiation, because the
// larger callee costs the panel loop spills (+1 % on the NTRU workload, which has none)
// REPLAY: uniforms from the recorded stream + the reference's table SampleZ (see ChainArgs::U)
template <bool WANT_LZ, bool WANT_D2, bool IMHK, bool WIDE, bool REPLAY>
__device__ __forceinline__ void panel_role(ChainSmem& sm, const ChainArgs& P, const int wg) {
    // The warpgroup's work is a sequence of JOBS: one job = one 64-coordinate pair of one tile.  With TPW = 2 the
    // warpgroup owns two slots (tiles) and takes their jobs in strict alternation, so the 4-stage item that finishes
    // the centers of a tile's next pair (and the look-ahead item before it) runs while the warpgroup draws a pair of
    // the other tile.  Nothing per-thread survives a job in registers except with TPW = 1: the partial log-density
    // sums are parked in their output arrays, the chain state (log weight, location, counters) is read and written
    // back around the accept test, the window is reloaded from the workspace at every pair as before.
    const int TPW = REPLAY ? 1 : P.tpw;   // 1 or 2, the same for the whole launch
    const int sl = (threadIdx.x - kPanelBase) & 127;
    const int lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int wbase = sl & ~31;  // first sample of this warp inside the tile
    const int npairs = P.npad / kPair;
    const int nsub = P.npad / kSubW;
    // Sub-blocks of the top pair that consist of padding only (n not a multiple of 64: coordinates n .. npad - 1 are
    // centred at 0 with x = 0) are not walked at all: their rows keep the zeros the workspace was cleared to / the top
    // pair's centers item writes (t = 0 there).  n = 16 walks 2 sub-blocks instead of 8, n = 10 two (16 coordinates).
    // (ChainArgs::skip0, set by chain_launch: a kernel parameter costs the panel loop no register)
    const Philox gen(P.seed);
    double* __restrict__ xs = &sm.xs[wg][0][sl];        // this thread's column of the sub-block's draws
    double* __restrict__ xsw = &sm.xs[wg][0][wbase];    // the warp's 32 columns (DMMA fragments)
    const bool rec = kStatsBuild && P.stats && wg == 0 && sl == 0;
    constexpr bool kDetail = LGS_CHAIN_STATS_DETAIL != 0;
    const bool rec2 = kDetail && rec;
    long long t_cwait = 0, t_cfull = 0, t_bar = 0;
    const long long t_begin = rec ? clock64() : 0;
    // bytes of the constants of the sub-block whose rows-below count is `base` (the rect rows come last)
    auto sub_bytes = [](int base) -> uint32_t {
        return (uint32_t)(offsetof(SubPack, rect) + ((base + 15) / 16 * 16) * kRectLd * sizeof(double));
    };
    // Constants of the sub-blocks, double-buffered per warpgroup: load number q goes to buffer q & 1.  q_cons counts the
    // loads consumed and released (every thread); the issuing side is thread 0 of the warpgroup alone, its state in
    // sm.iss[wg].  Loads are issued in processing order, at most two ahead, across job boundaries when the next job is
    // certain (PanelIssue::nstart): its first sub-blocks are then in flight before the job begins.
    uint32_t q_cons = 0;
    PanelIssue& iss = sm.iss[wg];
    // constants of sub-block i of a step (i = 0 is the top one): SubPack nsub - 1 - i, rows below = 8 (7 - i % 8)
    auto topup = [&]() {   // thread 0 of the warpgroup
        uint32_t qi = iss.q_iss;
        int cur = iss.cur, nj = iss.nj;
        const int end = iss.end, nstart = iss.nstart;
        while (qi - q_cons < 2u) {
            int i;
            if (cur < end) i = cur++;
            else if (nstart >= 0 && nstart + nj < (nstart / (kPair / kSubW) + 1) * (kPair / kSubW)) i = nstart + nj++;
            else break;
            const uint32_t nb = sub_bytes(kSubW * (7 - (i & 7)));
            mbar_arrive_expect_tx(&sm.cfull[wg][qi & 1], nb);
            bulk_g2s(&sm.sub[wg][qi & 1], &P.subs[nsub - 1 - i], nb, &sm.cfull[wg][qi & 1]);
            qi++;
        }
        iss.q_iss = qi;
        iss.cur = cur;
        iss.nj = nj;
    };
    if (sl == 0) iss = PanelIssue{0u, 0, 0, -1, 0};
    PanelSlot* pst = sm.pst[wg][sl >> 5];
    if (lane == 0) {
        for (int j = 0; j < kTpw; j++) pst[j] = PanelSlot{j < TPW ? -1 : -2, 0, 0, 0u};
    }
    __syncwarp();
    // TPW = 1: these stay in registers from pair to pair
    double d2a = 0.0, lza = 0.0;
    long long upos = 0;   // replay: position inside this unit's recorded stream
    int k = 0;
    for (;;) {
        // the next slot of the rotation that still has tiles to come
        PanelSlot me = pst[k];
        {
            int tries = 1;
            while (me.tile == -2 && tries < TPW) {
                k = k + 1 == TPW ? 0 : k + 1;
                me = pst[k];
                tries++;
            }
            if (me.tile == -2) break;
        }
        const int slot = wg * TPW + k;
        {
            const long long t0 = rec ? clock64() : 0;
            mbar_wait(&sm.cready[slot], me.crp);
            if (rec) t_cwait += clock64() - t0;
        }
        me.crp ^= 1;
        const bool tile_start = me.tile < 0;
        if (tile_start) {
            me.tile = sm.slot_tile[slot];
            me.step = 0;
            me.pair = 0;
            if (me.tile == -3) me.tile = -1;       // not started yet (staggered start): ask again at the next turn
            else if (me.tile < 0) me.tile = -2;
        }
        __syncwarp();
        if (lane == 0) pst[k] = me;   // re-read at the end of the job: nothing of it stays in registers meanwhile
        __syncwarp();
        if (me.tile < 0) {
            if (me.tile == -1) k = k + 1 == TPW ? 0 : k + 1;
            continue;
        }
        const int tile = me.tile;
        const int step = me.step;
        const int p = me.pair;
        const int64_t s = (int64_t)tile * kTile + sl;
        // step-parallel proposals: the tile's chains and its step (see ChainArgs::tiles_per_step)
        const uint64_t id = P.first_id + (uint64_t)((!IMHK && P.tiles_per_step) ? (int64_t)(tile % P.tiles_per_step) * kTile + sl : s);
        const double* __restrict__ urow = REPLAY ? P.U + (size_t)(s < P.S ? s : 0) * P.u_stride : nullptr;
        // column copy of one chain's coefficients between tiled matrices (predicated per lane, coalesced per row)
        auto copy_column = [&](const double* __restrict__ src, double* __restrict__ dst) {
            const size_t o = (size_t)tile * P.npad * kRowLd + sl;
#pragma unroll 8
            for (int i = 0; i < P.npad; i++) dst[o + (size_t)i * kRowLd] = src[o + (size_t)i * kRowLd];
        };
        if (REPLAY && tile_start) upos = 0;
        if (IMHK && tile_start) {
            if (REPLAY && s < P.S) {
                if (P.upos) upos = P.upos[s];
                // the state a call starts from (launches of initialised chains): entry 0 of the log-weight row
                if (P.logw_out && P.step_off == 0 && !P.first_is_init) P.logw_out[s * (P.call_steps + 1)] = P.lw_cur[s];
            }
            if (P.nsteps == 0) {  // flush only: every current state into Xcur
                if (s < P.S) {
                    const int loc = P.loc[s];
                    if (loc) copy_column(P.Xbuf[loc - 1], P.Xcur);
                    P.loc[s] = 0;
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(&sm.pdone[slot]);
                    pst[k].tile = -1;
                }
                __syncwarp();
                k = k + 1 == TPW ? 0 : k + 1;
                continue;
            }
        }
        if (sl == 0) {
            // first sub-block of the job that follows this one, or -1 when that depends on a tile assignment not made yet
            int nstart = -1;
            int o = k;   // the slot the rotation reaches next
            for (int j = 1; j < TPW; j++) {
                const int c = k + j < TPW ? k + j : k + j - TPW;
                if (pst[c].tile != -2) { o = c; break; }
            }
            if (o != k) {
                const PanelSlot ot = pst[o];
                if (ot.tile >= 0) nstart = ot.pair == 0 ? P.skip0 : ot.pair * (kPair / kSubW);
            } else if (p + 1 < npairs) {
                nstart = (p + 1) * (kPair / kSubW);
            } else if (step + 1 < P.nsteps) {
                nstart = P.skip0;
            }
            // the job starts at sub-block skip0 / 8 p; iss.nj of its sub-blocks were requested during the previous job
            iss.cur = (p == 0 ? P.skip0 : p * (kPair / kSubW)) + iss.nj;
            iss.end = (p + 1) * (kPair / kSubW);
            iss.nstart = nstart;
            iss.nj = 0;
            topup();
        }
        const uint32_t tagv = P.tag + (uint32_t)step + ((!IMHK && P.tiles_per_step) ? (uint32_t)(tile / P.tiles_per_step) : 0u);
        double* __restrict__ xtile = P.Xbuf[tagv & 1] + (size_t)tile * P.npad * kRowLd;
        double* __restrict__ xt = xtile + sl;
        if (p == 0) {
            d2a = 0.0;
            lza = 0.0;
        } else if (TPW > 1) {  // partial sums of the pairs above, parked in the output arrays
            if (WANT_D2) d2a = P.d2[s];
            if (WANT_LZ) lza = P.lz[s];
        }
        uint32_t wz = 0, ww = 0;
        bool wfresh = true;
        {
            const int lo = P.npad - kPair * (p + 1);
            const int base_top = kPair - kSubW - (p == 0 ? P.skip0 * kSubW : 0);   // first sub-block with a coordinate < n
            double c[kSubW];
            {
                const double* xl = xt + (size_t)(lo + base_top) * kRowLd;
#pragma unroll
                for (int kk = 0; kk < kSubW; kk++) c[kk] = __ldcg(xl + (size_t)kk * kRowLd);
            }
#pragma unroll 1
            for (int base = base_top; base >= 0; base -= kSubW) {
                const int buf = q_cons & 1;
                // done with this buffer: once all four warps of the warpgroup say so it is refilled with the constants
                // of the sub-block after next
                auto release_sub = [&]() {
                    {
                        const long long t0 = rec2 ? clock64() : 0;
                        named_bar_sync(1 + wg, 128);
                        if (rec2) t_bar += clock64() - t0;
                    }
                    q_cons++;
                    if (sl == 0) topup();
                };
                {
                    const long long t0 = rec2 ? clock64() : 0;
                    mbar_wait(&sm.cfull[wg][buf], (q_cons >> 1) & 1);
                    if (rec2) t_cfull += clock64() - t0;
                }
                const SubPack& sp = sm.sub[wg][buf];
                double* __restrict__ xp = xt + (size_t)(lo + base + kSubW - 1) * kRowLd;
#pragma unroll 1
                for (int j = kSubW - 1; j >= 0; j--, xp -= kRowLd) {
                    const CoordConst& cc = sp.cc[j];
                    const int4 ci = sp.ci[j];
                    const double mean = c[kSubW - 1];  // 1/r_ii is folded into the centers and the triangle
                    double x;
                    // rounded once for all three uses (ptxas if-converts the deterministic-coordinate branch below and used
                    // to issue a second, predicated-off FRND.F64 on every coordinate)
                    const double rmean = rint(mean);
                    if (ci.w) {
                        x = rmean;
                    } else if constexpr (REPLAY) {
                        const double u = s < P.S ? __ldg(urow + upos) : 0.5;
                        double lse = 0.0;
                        const bool ok = chain_samplez_table_call<WANT_LZ>(mean, __ldg(P.sig + lo + base + j), P.precision,
                                                                          P.max_range, u, &x, &lse);
                        if (ok) upos++;
                        else if (P.fallback && s < P.S) atomicAdd(P.fallback, 1);
                        if (WANT_LZ) lza += lse;
                    } else {
                        // one Philox call serves the draws 2m and 2m + 1: computed at the even one (or at the odd one
                        // when a job starts there), its second half kept for the odd one
                        const bool odd = ci.z & 1;
                        uint32_t uhi = wz, ulo = ww;
                        if (!odd || wfresh) {
                            const uint4 w = gen((uint32_t)id, (uint32_t)(id >> 32), (uint32_t)(ci.z >> 1), tagv);
                            wfresh = false;
                            wz = w.z;
                            ww = w.w;
                            uhi = odd ? w.z : w.x;
                            ulo = odd ? w.w : w.y;
                        }
                        const int kind = ci.x & 3;
                        bool done = false;
                        const float dlf = (float)(rmean - mean);
                        if (ci.x <= 1) {  // smooth regime, eligible for the FP32 screen
                            // no FP64 instruction between the window update above and x below: a panel warp's first
                            // FP64 instruction after a stretch of other work waits for a whole DMMA burst
                            float off, corr;
                            done = samplez_smooth_screen(dlf, sp.cf[j], kind, ci.y, uhi, &off, &corr);
                            x = rmean + (double)off;
                            if (WANT_LZ && done) lza += cc.lognz + (double)corr;
                        } else if (WIDE && !WANT_LZ && ci.x == 3) {
                            // wide truncated window: FP32 screen of the closed-form draw (the exact-weight
                            // instantiations need log Z in FP64 and stay on the FP64 path)
                            float off;
                            done = samplez_wide_screen(dlf, sp.cf[j], ci.y, uhi, &off);
                            x = rmean + (double)off;
                        }
                        if (!done) {
                            const double u = u53(uhi, ulo);
                            if (kind <= 1) {
                                double logz;
                                x = samplez_smooth<WANT_LZ>(mean, cc.a, cc.q, cc.sq, cc.znorm, cc.eps1, cc.lognz, kind,
                                                            ci.y, u, &logz);
                                if (WANT_LZ) lza += logz;
                            } else {
                                double zs;
                                const int wt = __float_as_int(sp.cf[j].pad);   // wide truncated window: its tables + 1
                                if (WIDE && kind == 3 && P.wtab && wt > 0)
                                    x = chain_samplez_wide_table_call(mean, &cc, sp.cf[j].sq, ci.y, u,
                                                                      P.wtab + (size_t)(wt - 1) * P.wtab_ld, &zs);
                                else
                                    x = chain_samplez_fast_call<WIDE>(mean, cc.a, cc.q, cc.sq, cc.znorm, cc.eps1, kind, ci.y,
                                                                      u, &zs);
                                if (WANT_LZ) lza += log(zs);
                            }
                        }
                    }
                    *xp = x;
                    xs[j * kRowLd] = x;
                    if (WANT_D2) {
                        const double dl = x - mean;
                        d2a = fma(dl * dl, cc.r2, d2a);
                    }
                    const double* __restrict__ T = sp.win[j];
#pragma unroll
                    for (int jj = kSubW - 2; jj >= 0; jj--) c[jj + 1] = fma(T[jj], x, c[jj]);
                    c[0] = 0.0;
                }
                if (base == 0) {
                    release_sub();
                    continue;
                }
                // Contribution of the sub-block to the rows below it (rows [0, base) of the pair) as warp-level DMMA:
                // C[16 rows, 8 samples] += rect[16 rows, 8 coordinates] * x[8 coordinates, 8 samples], one instruction
                // instead of 128 scalar FMAs per lane group -- every FP64 instruction of a panel warp queues behind
                // the DMMA warps' stream, so their count is what the panel's speed depends on.  The 8 rows right
                // below the sub-block are the next window: they go to the (now free) xs rows instead of HBM.
                __syncwarp();
                double bf[4][2];
#pragma unroll
                for (int nt = 0; nt < 4; nt++) {
                    bf[nt][0] = xsw[t * kRowLd + nt * 8 + g];
                    bf[nt][1] = xsw[(t + 4) * kRowLd + nt * 8 + g];
                }
                __syncwarp();
                const int wrow = base - kSubW;  // first row of the next window
#pragma unroll 1
                for (int r0 = 0; r0 < base; r0 += 16) {
                    double af[4];
                    af[0] = sp.rect[r0 + g][t];
                    af[1] = sp.rect[r0 + g + 8][t];
                    af[2] = sp.rect[r0 + g][t + 4];
                    af[3] = sp.rect[r0 + g + 8][t + 4];
                    const bool hi_ok = r0 + 8 < base;
                    double* __restrict__ r_lo = xtile + (size_t)(lo + r0 + g) * kRowLd + wbase + 2 * t;
                    double* __restrict__ r_hi = r_lo + 8 * kRowLd;
                    double d[4][4];
#pragma unroll
                    for (int nt = 0; nt < 4; nt++) {
                        const double2 a = __ldcg(reinterpret_cast<const double2*>(r_lo + nt * 8));
                        d[nt][0] = a.x;
                        d[nt][1] = a.y;
                        double2 b = make_double2(0.0, 0.0);
                        if (hi_ok) b = __ldcg(reinterpret_cast<const double2*>(r_hi + nt * 8));
                        d[nt][2] = b.x;
                        d[nt][3] = b.y;
                    }
#pragma unroll
                    for (int nt = 0; nt < 4; nt++) dmma_16x8x8(d[nt], af, bf[nt]);
                    const bool lo_win = r0 == wrow, hi_win = r0 + 8 == wrow;
#pragma unroll
                    for (int nt = 0; nt < 4; nt++) {
                        if (lo_win)
                            *reinterpret_cast<double2*>(&xsw[g * kRowLd + nt * 8 + 2 * t]) = make_double2(d[nt][0], d[nt][1]);
                        else
                            *reinterpret_cast<double2*>(r_lo + nt * 8) = make_double2(d[nt][0], d[nt][1]);
                        if (hi_win)
                            *reinterpret_cast<double2*>(&xsw[g * kRowLd + nt * 8 + 2 * t]) = make_double2(d[nt][2], d[nt][3]);
                        else if (hi_ok)
                            *reinterpret_cast<double2*>(r_hi + nt * 8) = make_double2(d[nt][2], d[nt][3]);
                    }
                }
                release_sub();
#pragma unroll
                for (int kk = 0; kk < kSubW; kk++) c[kk] = xs[kk * kRowLd];
            }
        }
        const PanelSlot fin = pst[k];
        const bool last_pair = fin.pair == npairs - 1;
        if (IMHK && last_pair) {
            const int step = fin.step;
            // imhk.py:102-124 + 156-174.  Reference weights: log rho(v) - log q(v) is the constant sum_lognorm
            // here (both are the same sum r_ii^2 (x_i - m_i)^2 / (2 sigma^2)): always accepted, as the reference.
            // The chain's state (log weight, location byte, counters) lives in its arrays between steps.
            const int buf = 1 + (int)(tagv & 1);
            const double lw_prop = P.exact ? lza : P.sum_lognorm;
            const bool live = s < P.S;
            double lw = live ? P.lw_cur[s] : 0.0;
            int loc = live ? (int)P.loc[s] : 0;
            if (P.first_is_init && step == 0) {
                lw = lw_prop;
                loc = buf;
                if (REPLAY && P.logw_out && live) P.logw_out[s * (P.call_steps + 1)] = lw;
            } else {
                bool acc = true;
                if (REPLAY) {  // imhk.py:156-174 on the recorded uniform (always drawn, imhk.py:167)
                    const double ua = live ? __ldg(urow + upos) : 0.0;
                    upos++;
                    const double ratio = (lw == -INFINITY) ? 1.0 : fmin(1.0, exp(lw_prop - lw));
                    acc = ua < ratio;
                    const int64_t kk = P.step_off + step - (P.first_is_init ? 1 : 0);
                    if (live) {
                        if (P.accept_out) P.accept_out[s * P.call_steps + kk] = acc ? 1 : 0;
                        if (P.logw_out) P.logw_out[s * (P.call_steps + 1) + kk + 1] = lw_prop;
                    }
                } else if (P.exact) {
                    const double ratio = (lw == -INFINITY) ? 1.0 : fmin(1.0, exp(lw_prop - lw));
                    acc = philox_uniform(gen, id, kAcceptDraw, tagv) < ratio;
                }
                if (live) {
                    if (acc) P.counts[2 * s] += 1;
                    P.counts[2 * s + 1] += 1;
                }
                if (acc) {
                    lw = lw_prop;
                    loc = buf;
                }
            }
            // a kept step inside the launch (ChainArgs::keep_thin): the chain's state as it stands after this step goes
            // into snapshot m; the scheduler hands its lattice points out as filler from there
            if (P.keep_thin && P.V && step >= P.keep_first && (step - P.keep_first) % P.keep_thin == 0) {
                const int m = (step - P.keep_first) / P.keep_thin;
                copy_column(loc ? P.Xbuf[loc - 1] : P.Xcur, P.Xkeep + (size_t)m * P.keep_stride);
            }
            // The accepted proposal is not copied: loc records the buffer it lives in.  A state still living in the
            // buffer the NEXT step draws into moves to Xcur now (before that step's first centers are written);
            // at a flush every state moves there.
            const int nbuf = 1 + (int)((tagv + 1) & 1);
            if (loc && (loc == nbuf || (P.flush && step == P.nsteps - 1))) {
                copy_column(P.Xbuf[loc - 1], P.Xcur);
                loc = 0;
            }
            if (live) {
                P.lw_cur[s] = lw;
                P.loc[s] = (uint8_t)loc;
                if (REPLAY && P.upos && step == P.nsteps - 1) P.upos[s] = upos;
            }
        }
        if (!IMHK) {
            if (last_pair) {
                // sum (x_i - m_i)^2 / (2 sigma_i^2) = sum r_ii^2 (x_i - m_i)^2 / (2 sigma^2)
                if (WANT_D2) {
                    P.lp[s] = d2a * P.inv_two_sigma2;
                    P.d2[s] = d2a;
                }
                if (WANT_LZ) P.lz[s] = lza;
            } else if (TPW > 1) {
                if (WANT_D2) P.d2[s] = d2a;
                if (WANT_LZ) P.lz[s] = lza;
            }
        } else if (TPW > 1 && WANT_LZ && !last_pair) {
            P.lz[s] = lza;
        }
        // the pair's coefficients are read next by bulk copies (async proxy)
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.pdone[slot]);
        // the slot's next job
        if (lane == 0) {
            PanelSlot nx = fin;
            if (++nx.pair == npairs) {
                nx.pair = 0;
                if (++nx.step >= P.nsteps) nx.tile = -1;   // the tile is finished: the slot waits for its next assignment
            }
            pst[k] = nx;
        }
        __syncwarp();
        k = k + 1 == TPW ? 0 : k + 1;
    }
    if (rec) {
        P.stats[blockIdx.x * 12 + 5] = (unsigned long long)(clock64() - t_begin);
        P.stats[blockIdx.x * 12 + 6] = (unsigned long long)t_cwait;
        P.stats[blockIdx.x * 12 + 7] = 0ull;
        P.stats[blockIdx.x * 12 + 8] = (unsigned long long)t_cfull;
        P.stats[blockIdx.x * 12 + 9] = (unsigned long long)t_bar;
    }
}

template <bool WANT_LZ, bool WANT_D2, bool IMHK, bool WIDE, bool REPLAY = false>
__global__ void __launch_bounds__(kThreads, 1) klein_chain_kernel(const __grid_constant__ ChainArgs P) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    ChainSmem& sm = *reinterpret_cast<ChainSmem*>(smem_raw);
    if (threadIdx.x == 0) {
        for (int i = 0; i < kStages; i++) {
            mbar_init(&sm.full[i], 1);
            mbar_init(&sm.empty[i], kDmmaThreads / 32);
        }
        for (int s = 0; s < kMaxSlots; s++) {
            mbar_init(&sm.cready[s], kDmmaThreads / 32);
            mbar_init(&sm.pdone[s], 4);
        }
        for (int s = 0; s < kWgs; s++) {
            mbar_init(&sm.cfull[s][0], 1);
            mbar_init(&sm.cfull[s][1], 1);
        }
        fence_mbar_init();
    }
    __syncthreads();
    const int tid = threadIdx.x;
    if (tid >= kDmmaBase && tid < kDmmaBase + kDmmaThreads) {
        reg_inc<kRegsDmma>();
        dmma_role(sm, P);
    } else if (tid >= kProdBase && tid < kProdBase + kProdThreads) {
        reg_dec<kRegsProd>();
        if (tid == kProdBase) producer_role(sm, P, kWgs * (REPLAY ? 1 : P.tpw));
    } else {
        reg_dec<kRegsPanel>();
        panel_role<WANT_LZ, WANT_D2, IMHK, WIDE, REPLAY>(sm, P, (tid - kPanelBase) >> 7);
    }
}

// tiled workspace -> int64 [S][n] row-major, through a 32x32 shared-memory transpose
__global__ void __launch_bounds__(256)
export_coeffs_tiled_kernel(const double* __restrict__ Xt, int npad, int n, int64_t S, int64_t* __restrict__ out,
                           int64_t ldo) {
    __shared__ double tile[32][33];
    const int64_t s0 = (int64_t)blockIdx.x * 32;
    const int i0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {
        const int i = i0 + r;
        const int64_t s = s0 + tx;
        tile[r][tx] = (i < n && s < S) ? Xt[((size_t)(s >> 7) * npad + i) * kRowLd + (s & 127)] : 0.0;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int64_t s = s0 + r;
        const int i = i0 + tx;
        if (s < S && i < n) out[(size_t)s * ldo + i] = (int64_t)tile[tx][r];
    }
}

// int64 [S][n] row-major -> tiled workspace (the inverse of export_coeffs_tiled_kernel); padding coordinates untouched
__global__ void __launch_bounds__(256)
import_coeffs_tiled_kernel(const int64_t* __restrict__ in, int64_t ldi, int npad, int n, int64_t S, double* __restrict__ Xt) {
    __shared__ double tile[32][33];
    const int64_t s0 = (int64_t)blockIdx.x * 32;
    const int i0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {
        const int64_t s = s0 + r;
        const int i = i0 + tx;
        tile[tx][r] = (s < S && i < n) ? (double)in[(size_t)s * ldi + i] : 0.0;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int i = i0 + r;
        const int64_t s = s0 + tx;
        if (i < n && s < S) Xt[((size_t)(s >> 7) * npad + i) * kRowLd + (s & 127)] = tile[r][tx];
    }
}

// ---------------------------------------------------------------------------------------------------------
// step-parallel IMHK (ChainArgs::tiles_per_step): the accept / reject recursion over the proposals of one launch
// ---------------------------------------------------------------------------------------------------------
struct ScanArgs {
    int64_t C, cpad;            // chains, chains padded to whole tiles
    int nsteps;                 // proposals per chain in this launch (the first one may be _initialize_chain)
    int first_is_init, exact;
    uint64_t seed, first_id;
    uint32_t tag;               // Philox tag of proposal 0
    double sum_lognorm;
    const double* lz;           // [nsteps][cpad] log-weights of the proposals (exact weights)
    double* lw_cur;             // [C]
    long long* counts;          // [C][2]
    int keep_first, thin, nkeep;   // kept states: after proposals keep_first, keep_first + thin, ...  (nkeep of them)
    int* keep_src;              // [nkeep][cpad] which proposal holds the state kept in slot m (-1: the state at entry)
    int* final_src;             // [cpad] the state after the last proposal
};

// imhk.py:141-177 for one chain per thread, given every proposal's log-weight: the same expressions, in the same
// order, as the step-by-step launches (panel_role, IMHK) -- chain states are bit-identical between the two.
__global__ void __launch_bounds__(128) imhk_scan_kernel(const ScanArgs A) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= A.C) return;
    const Philox gen(A.seed);
    const uint64_t id = A.first_id + (uint64_t)c;
    double lw = A.lw_cur[c];
    int src = -1, acc_n = 0, tot_n = 0;
    int next_keep = A.keep_first, slot = 0;
    for (int j = 0; j < A.nsteps; j++) {
        const double lw_prop = A.exact ? A.lz[(size_t)j * A.cpad + c] : A.sum_lognorm;
        if (A.first_is_init && j == 0) {
            lw = lw_prop;
            src = 0;
        } else {
            bool acc = true;
            if (A.exact) {
                const double ratio = (lw == -INFINITY) ? 1.0 : fmin(1.0, exp(lw_prop - lw));
                acc = philox_uniform(gen, id, kAcceptDraw, A.tag + (uint32_t)j) < ratio;
            }
            tot_n++;
            if (acc) {
                acc_n++;
                lw = lw_prop;
                src = j;
            }
        }
        if (slot < A.nkeep && j == next_keep) {
            A.keep_src[(size_t)slot * A.cpad + c] = src;
            slot++;
            next_keep += A.thin;
        }
    }
    A.final_src[c] = src;
    A.lw_cur[c] = lw;
    if (tot_n) {
        A.counts[2 * c] += acc_n;
        A.counts[2 * c + 1] += tot_n;
    }
}

// Column copies between tiled matrices: slot m of `dst` (tile = m * tps + chain tile) <- the proposal src[m][c] of
// `prop` (tile = src * tps + chain tile), or the chain's column of `cur` when src < 0 (skipped when skip_neg: the
// destination already is `cur`).  One block = the 128 chains of a tile x a strip of coordinates.
__global__ void __launch_bounds__(128)
imhk_gather_kernel(const double* __restrict__ prop, const double* __restrict__ cur, double* __restrict__ dst,
                   const int* __restrict__ src, int64_t C, int64_t cpad, int tps, int npad, int skip_neg) {
    const int ct = blockIdx.x, m = blockIdx.z;
    const int64_t c = (int64_t)ct * kTile + threadIdx.x;
    if (c >= C) return;
    const int sj = src[(size_t)m * cpad + c];
    if (sj < 0 && skip_neg) return;
    const double* __restrict__ from = (sj < 0 ? cur + (size_t)ct * npad * kRowLd
                                              : prop + ((size_t)sj * tps + ct) * npad * kRowLd) + threadIdx.x;
    double* __restrict__ to = dst + ((size_t)m * tps + ct) * npad * kRowLd + threadIdx.x;
    const int rows = (npad + gridDim.y - 1) / gridDim.y;
    const int r0 = blockIdx.y * rows, r1 = r0 + rows < npad ? r0 + rows : npad;
#pragma unroll 4
    for (int i = r0; i < r1; i++) to[(size_t)i * kRowLd] = from[(size_t)i * kRowLd];
}

__global__ void chain_finish_logq_kernel(const double* __restrict__ lp, double sum_lognorm, int64_t S,
                                         double* __restrict__ logq) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s < S) logq[s] = -lp[s] - sum_lognorm;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
void chain_free(BlockedCtx& bc) {
    ChainCtx& cc = bc.chain;
    cudaFree(cc.Rpk);
    cudaFree(cc.Bpk);
    cudaFree(cc.tp);
    cudaFree(cc.sig);
    cudaFree(cc.subs);
    cudaFree(cc.wtab);
    cudaFree(cc.hist_map);
    cudaFree(cc.pt_buf);
    for (auto& w : cc.work) {
        cudaFree(w.X); cudaFree(w.lp); cudaFree(w.d2); cudaFree(w.lz);
        w = BlockedWork{};
    }
    cc = ChainCtx{};
}

// shared-memory images of the 64 x 16 slivers of an [rows][cols] row-major matrix M (ldm = its row length):
// out[rt][kt][r][20], zero outside the matrix
static void pack_slivers(const double* M, int rows, int cols, int ldm, int npad, std::vector<double>& out,
                         bool identity_pad) {
    const int nrt = npad / kPair, nkt = npad / kBK;
    out.assign((size_t)nrt * nkt * kPair * kLdA, 0.0);
    for (int rt = 0; rt < nrt; rt++)
        for (int kt = 0; kt < nkt; kt++) {
            double* t = &out[((size_t)rt * nkt + kt) * kPair * kLdA];
            for (int r = 0; r < kPair; r++)
                for (int k = 0; k < kBK; k++) {
                    const int i = rt * kPair + r, j = kt * kBK + k;
                    double v = 0.0;
                    if (i < rows && j < cols) v = M[(size_t)i * ldm + j];
                    else if (identity_pad && i == j) v = 1.0;
                    t[r * kLdA + k] = v;
                }
        }
}

cudaError_t chain_init(BlockedCtx& bc, const lgs_desc& d, const HostConsts& hc) {
    ChainCtx& cc = bc.chain;
    const int n = d.n;
    const int npad = (n + kPair - 1) / kPair * kPair;
    cc.npad = npad;
    std::vector<double> invr(npad, 1.0);
    for (int i = 0; i < n; i++) {
        invr[i] = 1.0 / hc.rdiag[i];
        // the chain kernel derives sum (x-m)^2/(2 sigma_i^2) from sum r_ii^2 (x-m)^2: needs sigma_i = sigma/|r_ii| unclamped
        if (hc.sig[i] != d.sigma / std::fabs(hc.rdiag[i])) return cudaSuccess;  // not ready: the v1 path serves
    }
    // the centers are kept divided by their R_ii (as the window triangle and the rows-below rectangles are): row i of
    // the packed R and t_i carry the factor, so the centers epilogue is t' - acc with nothing else to load or multiply
    std::vector<double> Rup((size_t)n * n, 0.0);
    for (int i = 0; i < n; i++)
        for (int j = i; j < n; j++) Rup[(size_t)i * n + j] = d.R[(size_t)i * n + j] * invr[i];
    std::vector<double> Rpk, Bpk, tp(npad, 0.0);
    pack_slivers(Rup.data(), n, n, n, npad, Rpk, true);
    pack_slivers(d.B, n, n, n, npad, Bpk, false);
    cc.t_zero = true;
    for (int i = 0; i < n; i++) {
        tp[i] = d.t[i] * invr[i];
        cc.t_zero = cc.t_zero && d.t[i] == 0.0;
    }
    // kind-3 coordinates share one window half-width (max_range / 2); tables only for those
    int wide_dmax = 0;
    for (int i = 0; i < n; i++)
        if (hc.kind[i] == 3) wide_dmax = hc.dmax[i];
    std::vector<int> wide_rows;
    auto int_as_float_host = [](int v) { float f; memcpy(&f, &v, sizeof(f)); return f; };
    std::vector<SubPack> subs(npad / kSubW);
    for (int gs = 0; gs < npad / kSubW; gs++) {
        SubPack& sp = subs[gs];
        memset(&sp, 0, sizeof(sp));
        const int g0 = gs * kSubW;
        const int lo = g0 / kPair * kPair;
        const int base = g0 - lo;
        for (int j = 0; j < kSubW; j++) {
            const int gi = g0 + j;
            if (gi < n) {
                for (int jj = 0; jj + 1 < kSubW; jj++) {
                    const int k = j - (kSubW - 1) + jj;  // sub-block-relative coordinate held by window slot jj
                    if (k >= 0) sp.win[j][jj] = -d.R[(size_t)(g0 + k) * n + gi] * invr[g0 + k];
                }
                for (int r = 0; r < base; r++) sp.rect[r][j] = -d.R[(size_t)(lo + r) * n + gi] * invr[lo + r];
                CoordConst& c = sp.cc[j];
                c.inv_r = invr[gi];
                c.r2 = hc.rdiag[gi] * hc.rdiag[gi];
                c.a = hc.inv2s2[gi];
                c.q = hc.q[gi];
                c.sq = hc.sq[gi];
                c.znorm = hc.znorm[gi];
                c.eps1 = hc.eps1[gi];
                c.lognz = std::log(hc.znorm[gi]);
                // FP32 screen (samplez_smooth_screen): eligible while its error bound keeps the FP64 fallback rare
                const double sg = hc.sig[gi];
                const bool f32ok = hc.kind[gi] <= 1 && sg <= 16.0;
                SmoothF32& f = sp.cf[j];
                f.a = (float)c.a; f.sq = (float)c.sq; f.q = (float)c.q; f.znorm = (float)c.znorm; f.eps1 = (float)c.eps1;
                f.tol_r = (float)(4e-6 + 4e-7 * sg * sg);
                f.tol_a = (float)(1.6e-7 * c.znorm);
                bool wide_f32 = false;
                // tables of samplez_wide_table -- only where its series are exact to FP64: exp_small / the cosh series
                // take arguments up to 2 a |delta| d <= a dmax (|delta| <= 1/2) and hold their < 1e-17 error for
                // |x| <= 0.11.  With the default window (precision 10, max_range 1000) a dmax <= 0.1 always; a handle
                // made with a larger precision (adaptive_precision_sample passes 20) can exceed it, and such a
                // coordinate keeps the closed-form FP64 draw (pad = 0).
                if (hc.kind[gi] == 3 && hc.dmax[gi] == wide_dmax && hc.inv2s2[gi] * (double)wide_dmax <= 0.1) {
                    wide_rows.push_back(gi);
                    f.pad = int_as_float_host((int)wide_rows.size());   // table number + 1; 0 = none
                }
                if (hc.kind[gi] == 3) {  // constants of samplez_wide_screen
                    static const bool wide_screen = [] { const char* v = getenv("LGS_WIDE_SCREEN"); return !(v && v[0] == '0'); }();
                    const double sa = std::sqrt(c.a);
                    f.a = (float)c.a; f.sq = (float)sa; f.q = (float)(0.88622692545275801365 / sa);
                    f.znorm = (float)(sa * 0.56418958354775628695); f.eps1 = (float)(1.0 / sa);
                    static const float wide_tol = [] { const char* v = getenv("LGS_WIDE_TOL"); return v ? (float)atof(v) : 4e-6f; }();
                    f.tol_r = wide_tol; f.tol_a = 0.0f;   // LGS_WIDE_TOL: development (what the FP64 fallbacks cost)
                    wide_f32 = wide_screen;
                }
                sp.ci[j] = make_int4(hc.kind[gi] | ((f32ok || wide_f32) ? 0 : 4), hc.dmax[gi], hc.uidx[gi], hc.mode[gi]);
            } else {  // padding coordinate: center 0, R_ii = 1, rounded => x = 0
                sp.cc[j].inv_r = 1.0;
                sp.ci[j] = make_int4(0, 0, 0, 1);
            }
        }
    }
    cudaError_t e;
    auto up = [&](void** dst, const void* src, size_t bytes) -> cudaError_t {
        cudaError_t er = cudaMalloc(dst, bytes);
        if (er != cudaSuccess) return er;
        return cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice);
    };
    if ((e = up((void**)&cc.Rpk, Rpk.data(), sizeof(double) * Rpk.size())) != cudaSuccess) return e;
    if ((e = up((void**)&cc.Bpk, Bpk.data(), sizeof(double) * Bpk.size())) != cudaSuccess) return e;
    if ((e = up((void**)&cc.tp, tp.data(), sizeof(double) * tp.size())) != cudaSuccess) return e;
    {
        std::vector<double> sigp(npad, 1.0);
        for (int i = 0; i < n; i++) sigp[i] = hc.sig[i];
        if ((e = up((void**)&cc.sig, sigp.data(), sizeof(double) * sigp.size())) != cudaSuccess) return e;
    }
    cc.wtab = nullptr;
    cc.wtab_ld = 0;
    {
        static const bool use_tab = [] { const char* v = getenv("LGS_WIDE_TABLE"); return !(v && v[0] == '0'); }();
        if (use_tab && !wide_rows.empty()) {
            const int ld = wide_dmax + 1;
            const size_t per = (size_t)(kWideTabK + 1) * ld;
            std::vector<double> tab(per * wide_rows.size());
            for (size_t w = 0; w < wide_rows.size(); w++) {
                const long double a = (long double)hc.inv2s2[wide_rows[w]];
                double* t = &tab[w * per];
                long double S[kWideTabK] = {};
                for (int k = 0; k < kWideTabK; k++) t[(size_t)k * ld] = 0.0;
                t[(size_t)kWideTabK * ld] = 1.0;
                for (int dd = 1; dd <= wide_dmax; dd++) {
                    const long double j2 = (long double)dd * dd, g = expl(-a * j2);
                    long double pw = g;
                    for (int k = 0; k < kWideTabK; k++) {
                        S[k] += pw;
                        t[(size_t)k * ld + dd] = (double)S[k];
                        pw *= j2;
                    }
                    t[(size_t)kWideTabK * ld + dd] = (double)g;
                }
                // normaliser coefficients 2 S_k(D) / (2k)! into the CoordConst fields a wide window does not use
                CoordConst& c = subs[wide_rows[w] / kSubW].cc[wide_rows[w] % kSubW];
                const long double fact[5] = {1.0L, 2.0L, 24.0L, 720.0L, 40320.0L};
                double* dst[5] = {&c.q, &c.sq, &c.znorm, &c.eps1, &c.lognz};
                for (int k = 0; k < 5; k++) *dst[k] = (double)(2.0L * S[k] / fact[k]);
            }
            if ((e = up((void**)&cc.wtab, tab.data(), sizeof(double) * tab.size())) != cudaSuccess) return e;
            cc.wtab_ld = (int64_t)per;
        }
    }
    cc.precision = d.precision > 0 ? d.precision : 10;
    cc.max_range = d.max_range > 0 ? d.max_range : 1000;
    if ((e = up((void**)&cc.subs, subs.data(), sizeof(SubPack) * subs.size())) != cudaSuccess) return e;
#define LGS_CHAIN_ATTR(LZ, D2, MH)                                                                                 \
    if ((e = cudaFuncSetAttribute(klein_chain_kernel<LZ, D2, MH, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                  (int)sizeof(ChainSmem))) != cudaSuccess)                                        \
        return e;                                                                                                  \
    if ((e = cudaFuncSetAttribute(klein_chain_kernel<LZ, D2, MH, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                  (int)sizeof(ChainSmem))) != cudaSuccess)                                        \
        return e;
    LGS_CHAIN_ATTR(false, false, false)
    LGS_CHAIN_ATTR(false, true, false)
    LGS_CHAIN_ATTR(true, false, false)
    LGS_CHAIN_ATTR(false, false, true)
    LGS_CHAIN_ATTR(true, false, true)
#undef LGS_CHAIN_ATTR
#define LGS_CHAIN_ATTR_REPLAY(LZ, D2, MH)                                                                            \
    if ((e = cudaFuncSetAttribute(klein_chain_kernel<LZ, D2, MH, false, true>,                                       \
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ChainSmem))) != cudaSuccess) \
        return e;
    LGS_CHAIN_ATTR_REPLAY(false, true, false)
    LGS_CHAIN_ATTR_REPLAY(false, false, true)
    LGS_CHAIN_ATTR_REPLAY(true, false, true)
#undef LGS_CHAIN_ATTR_REPLAY
    {
        // a full batch gives every SM the same whole number of rounds of kWgs tiles
        // ... and more rounds per launch for small dimensions, where a tile is short and the ramps at both ends of a
        // launch weigh more (+3 % at n = 512 to +7 % at n = 256 / 64); the workspace stays ~1.4 GB either way
        const char* env = getenv("LGS_CHAIN_ROUNDS");
        int by_n = 3 * 1024 / npad;
        by_n = by_n < 3 ? 3 : by_n > 24 ? 24 : by_n;
        const int rounds = env ? atoi(env) : by_n;
        cc.batch = (int64_t)(bc.sm_count > 0 ? bc.sm_count : 148) * kTile * kWgs * (rounds > 0 ? rounds : 3);
    }
    cc.wide = false;
    for (int i = 0; i < bc.n; i++) cc.wide = cc.wide || hc.kind[i] == 3;
    cc.ready = true;
    return cudaSuccess;
}

static cudaError_t chain_ensure_work(ChainCtx& cc, BlockedWork& w, int64_t cap) {
    if (w.cap >= cap) return cudaSuccess;
    cudaFree(w.X); cudaFree(w.lp); cudaFree(w.d2); cudaFree(w.lz);
    w = BlockedWork{};
    cudaError_t e;
    const size_t xbytes = sizeof(double) * (size_t)(cap / kTile) * cc.npad * kRowLd;
    if ((e = cudaMalloc(&w.X, xbytes)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&w.lp, sizeof(double) * cap)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&w.d2, sizeof(double) * cap)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&w.lz, sizeof(double) * cap)) != cudaSuccess) return e;
    // the padding columns of every row are read by the bulk copies: keep them finite
    if ((e = cudaMemset(w.X, 0, xbytes)) != cudaSuccess) return e;
    w.cap = cap;
    return cudaSuccess;
}

// one launch of the chain kernel
static cudaError_t chain_launch(BlockedCtx& bc, ChainArgs& a, bool imhk, bool want_lz, bool want_d2, cudaStream_t st) {
    ChainCtx& cc = bc.chain;
    a.Rpk = cc.Rpk; a.Bpk = cc.Bpk; a.tvec = cc.tp; a.t_zero = cc.t_zero ? 1 : 0; a.subs = cc.subs;
    a.wtab = cc.wtab; a.wtab_ld = cc.wtab_ld;
    a.hist_map = (cc.hist && a.V && !cc.hist_pause) ? cc.hist_map : nullptr;
    a.hist = cc.hist; a.hist_lo = cc.hist_lo; a.hist_inv_width = cc.hist_inv_width; a.hist_nbins = cc.hist_nbins;
    a.inv_two_sigma2 = 1.0 / bc.two_sigma2;
    a.sum_lognorm = bc.sum_lognorm;
    a.npad = cc.npad; a.n_valid = bc.n;
    a.skip0 = (cc.npad - (bc.n + kSubW - 1) / kSubW * kSubW) / kSubW;
    if (!(a.tiles_per_step > 0 && a.ntiles > 0)) a.ntiles = (int)((a.S + kTile - 1) / kTile);
    a.want_lz = want_lz ? 1 : 0;
    {
        const int la = [] { const char* v = getenv("LGS_CHAIN_LOOKAHEAD"); return v ? atoi(v) : 0; }();   // read per launch: the tests switch it
        a.lookahead = la;
        // staggered start: IMHK launches and proposal-only batches (+3.9 % / +1.5 %); a Klein batch with lattice points
        // has filler for the top of its steps and only pays the later start (-1 %)
        const int sg = [] { const char* v = getenv("LGS_CHAIN_STAGGER"); return v ? atoi(v) : -1; }();
        a.stagger = sg >= 0 ? sg : ((a.nsteps > 1 || !(a.V || a.V16)) ? 100 : 0);   // per cent of the even spacing
        if (a.U) a.stagger = 0;   // replay runs one tile per warpgroup and is not a throughput path: plain start
        // Two or three tiles per panel warpgroup where nothing else hides the wait for a pair's centers: IMHK launches
        // (nine steps out of ten have no lattice points to contract: 81.0 / 74.3 / 73.3 ms per 10 steps at n = 1024 with
        // one / two / three; three is slower than two at n <= 512), proposal-only batches (7.74 -> 6.83 ms), and tiles so
        // short that the hand-over to the next tile shows (n <= 32: +8 %).  A Klein batch
        // with lattice points has that filler and is 1-3 % faster with one tile per warpgroup; its finished tiles also
        // reach the host transport sooner (mapped int16 route: 19.6 against 22.3 ms per call).
        const int tp = [] { const char* v = getenv("LGS_CHAIN_TPW"); return v ? atoi(v) : 0; }();
        const bool no_filler = a.nsteps > 1 || (a.nsteps == 1 && !(a.V || a.V16));
        a.tpw = (tp >= 1 && tp <= kTpw) ? tp : (no_filler ? (cc.npad >= 12 * kPair ? 3 : 2) : (bc.n <= 32 ? 2 : 1));
        if (a.tpw > kTpw) a.tpw = kTpw;
    }
    const int sms = bc.sm_count > 0 ? bc.sm_count : 148;
    const int grid = a.ntiles < sms ? a.ntiles : sms;
    {
        // global lattice-point queue (ChainArgs::gq): every launch that produces lattice points, except the mapped
        // zero-copy transport, whose per-tile completion word relies on a tile's row blocks finishing in order on one SM.
        // LGS_CHAIN_GQ=0: per-CTA filler order (development A/B)
        const bool gq_on = [] { const char* v = getenv("LGS_CHAIN_GQ"); return !(v && v[0] == '0'); }();
        a.gq = 0;
        // Small dimensions keep the per-CTA order: a tile's lattice points are then one or two short items and the
        // ticket costs more than the re-read (n = 64: -2 %; n = 512: +3.6 %; n = 1024: +0.3 %, DRAM reads per launch cut)
        if (gq_on && (a.V || (a.V16 && a.group_cnt)) && grid > 1 && cc.npad >= 4 * kPair) {
            // one list entry per (tile, kept state of the launch)
            const size_t need = sizeof(int) * ((size_t)a.ntiles * (size_t)(a.keep_thin ? a.nkeep : 1) + 2);
            if (cc.pt_cap < need) {
                cudaFree(cc.pt_buf);
                cc.pt_buf = nullptr, cc.pt_cap = 0;
                const size_t cap = need < 65536 ? 65536 : need * 2;
                cudaError_t e = cudaMalloc(&cc.pt_buf, cap);
                if (e != cudaSuccess) return e;
                cc.pt_cap = cap;
            }
            cudaError_t e = cudaMemsetAsync(cc.pt_buf, 0, need, st);
            if (e != cudaSuccess) return e;
            a.gq = 1;
            a.pt_ctr = reinterpret_cast<unsigned*>(cc.pt_buf);
            a.pt_list = reinterpret_cast<int*>(cc.pt_buf) + 2;
            if (a.V16) a.tile_target = (unsigned)(kDmmaThreads / 32) * (unsigned)(cc.npad / kPair);
        }
    }
    static const bool want_stats = kStatsBuild && getenv("LGS_CHAIN_STATS") != nullptr;
    static unsigned long long* stats_dev = nullptr;
    if (want_stats && !stats_dev) cudaMalloc(&stats_dev, sizeof(unsigned long long) * 12 * 1024);
    a.stats = want_stats ? stats_dev : nullptr;
    {
        ProfScope ps(bc.prof, LGS_PROF_CHAIN, st);
#define LGS_CHAIN_GO(LZ, D2, MH)                                                                  \
    do {                                                                                          \
        if (cc.wide) klein_chain_kernel<LZ, D2, MH, true><<<grid, kThreads, sizeof(ChainSmem), st>>>(a);   \
        else klein_chain_kernel<LZ, D2, MH, false><<<grid, kThreads, sizeof(ChainSmem), st>>>(a);          \
    } while (0)
        if (a.U) {  // replay instantiations (the table sampler covers every kind of window: no WIDE split)
            a.sig = cc.sig; a.precision = cc.precision; a.max_range = cc.max_range;
            if (imhk && want_lz) klein_chain_kernel<true, false, true, false, true><<<grid, kThreads, sizeof(ChainSmem), st>>>(a);
            else if (imhk) klein_chain_kernel<false, false, true, false, true><<<grid, kThreads, sizeof(ChainSmem), st>>>(a);
            else klein_chain_kernel<false, true, false, false, true><<<grid, kThreads, sizeof(ChainSmem), st>>>(a);
        } else if (imhk && want_lz) LGS_CHAIN_GO(true, false, true);
        else if (imhk) LGS_CHAIN_GO(false, false, true);
        else if (want_lz) LGS_CHAIN_GO(true, false, false);   // step-parallel IMHK proposals with exact weights
        else if (want_d2) LGS_CHAIN_GO(false, true, false);
        else LGS_CHAIN_GO(false, false, false);
#undef LGS_CHAIN_GO
    }
    if (want_stats) {  // development only: synchronous read-back, mean over CTAs
        std::vector<unsigned long long> hs((size_t)grid * 12);
        cudaStreamSynchronize(st);
        cudaMemcpy(hs.data(), stats_dev, sizeof(unsigned long long) * hs.size(), cudaMemcpyDeviceToHost);
        double m[12] = {};
        for (int b = 0; b < grid; b++)
            for (int k = 0; k < 12; k++) m[k] += (double)hs[(size_t)b * 12 + k] / grid;
        fprintf(stderr,
                "[chain stats] grid %d tiles %d V=%d steps=%d | dmma total %.0f kcyc, wait-full %.0f (%.1f%%), stages %.0f | "
                "producer wait-empty %.0f idle %.0f | panel0 total %.0f wait-centers %.0f (%.1f%%) wait-tile %.0f "
                "wait-constants %.0f sub-block barrier %.0f\n",
                grid, a.ntiles, a.V ? 1 : 0, a.nsteps, m[0] / 1e3, m[1] / 1e3, 100.0 * m[1] / (m[0] + 1), m[2], m[3] / 1e3,
                m[4] / 1e3, m[5] / 1e3, m[6] / 1e3, 100.0 * m[6] / (m[5] + 1), m[7] / 1e3, m[8] / 1e3, m[9] / 1e3);
    }
    return cudaGetLastError();
}

// one batch of S <= w.cap samples: coefficients into w.X (tiled), accumulators into w.lp / d2, lattice points into V
// (may be nullptr)
static cudaError_t chain_batch(BlockedCtx& bc, BlockedWork& w, int64_t S, uint64_t seed, uint64_t first_id,
                               bool want_d2, double* V, int64_t ldv, const ChainReplay* rp, int64_t s0, cudaStream_t st) {
    ChainArgs a{};
    a.Xbuf[0] = a.Xbuf[1] = w.X;
    a.lp = w.lp; a.d2 = w.d2; a.lz = w.lz;
    a.V = V; a.ldv = ldv; a.S = S;
    a.seed = seed; a.first_id = first_id; a.tag = 0u;
    a.nsteps = 1;
    if (rp) {
        a.U = rp->U + (size_t)s0 * rp->stride;
        a.u_stride = rp->stride;
        a.fallback = rp->fallback;
    }
    return chain_launch(bc, a, false, false, want_d2 || rp != nullptr, st);
}

// one batch (S <= chain.batch) with the packed zero-copy output described at ChainArgs::V16
cudaError_t launch_klein_chain_zc(BlockedCtx& bc, int64_t S, uint64_t seed, uint64_t first_id, const ChainZeroCopy& zc,
                                  cudaStream_t st) {
    ChainCtx& cc = bc.chain;
    if (S <= 0 || S > cc.batch || bc.n % 8) return cudaErrorInvalidValue;
    cudaError_t e;
    BlockedWork& w = cc.work[0];
    const int64_t cap = S < cc.batch ? (S + kTile - 1) / kTile * kTile : cc.batch;
    if ((e = chain_ensure_work(cc, w, cap)) != cudaSuccess) return e;
    const int ntiles = (int)((S + kTile - 1) / kTile);
    if ((e = cudaMemsetAsync(zc.tile_cnt, 0, sizeof(unsigned) * ntiles, st)) != cudaSuccess) return e;
    ChainArgs a{};
    a.Xbuf[0] = a.Xbuf[1] = w.X;
    a.lp = w.lp; a.d2 = w.d2; a.lz = w.lz;
    a.V = nullptr; a.ldv = 0; a.S = S;
    a.V16 = zc.v16; a.tile_cnt = zc.tile_cnt; a.tile_done = zc.tile_done; a.ovf = zc.ovf; a.epoch = zc.epoch;
    a.tile_target = kDmmaThreads / 32;
    a.group_cnt = zc.group_cnt; a.group_tiles = zc.group_tiles;
    a.seed = seed; a.first_id = first_id; a.tag = 0u;
    a.nsteps = 1;
    return chain_launch(bc, a, false, false, false, st);
}

cudaError_t launch_klein_chain(BlockedCtx& bc, int64_t S, uint64_t seed, uint64_t first_id, double* V, int64_t* X,
                               double* logq, cudaStream_t st, const ChainReplay* rp) {
    if (S <= 0) return cudaSuccess;
    ChainCtx& cc = bc.chain;
    cudaError_t e;
    const int n = bc.n;
    const int64_t cap = S < cc.batch ? (S + kTile - 1) / kTile * kTile : cc.batch;
    BlockedWork& w = cc.work[0];
    if ((e = chain_ensure_work(cc, w, cap)) != cudaSuccess) return e;
    for (int64_t s0 = 0; s0 < S; s0 += w.cap) {
        const int64_t cnt = (S - s0) < w.cap ? (S - s0) : w.cap;
        if ((e = chain_batch(bc, w, cnt, seed, first_id + (uint64_t)s0, logq != nullptr, V ? V + (size_t)s0 * n : nullptr, n,
                             rp, s0, st)) != cudaSuccess) return e;
        if (X) {
            ProfScope ps(bc.prof, LGS_PROF_OTHER, st);
            dim3 grid((unsigned)((cnt + 31) / 32), (unsigned)((n + 31) / 32));
            export_coeffs_tiled_kernel<<<grid, 256, 0, st>>>(w.X, cc.npad, n, cnt, X + (size_t)s0 * n, n);
        }
        if (logq) {
            ProfScope ps2(bc.prof, LGS_PROF_OTHER, st);
            chain_finish_logq_kernel<<<(unsigned)((cnt + 255) / 256), 256, 0, st>>>(w.lp, bc.sum_lognorm, cnt, logq + s0);
        }
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// ---------------------------------------------------------------------------------------------------------
// IMHK on the chain kernel (imhk.py:126-229): all chains of the handle, `nsteps` Metropolis-Hastings steps per launch.
// A launch ends at a kept step when kept states are wanted: its last step moves every current state into Xcur and
// the lattice points are contracted from there inside the same launch.
// ---------------------------------------------------------------------------------------------------------
static cudaError_t chain_ensure_chains(BlockedCtx& bc, ChainSet& cs, bool with_proposal_buffers = true) {
    ChainCtx& cc = bc.chain;
    const int64_t cap = (cs.chains + kTile - 1) / kTile * kTile;
    cudaError_t e;
    if (!(cs.bl_cap >= cap && cs.Xcur)) {
        cudaFree(cs.Xcur);
        cudaFree(cs.flags);
        cs.Xcur = nullptr;
        cs.flags = nullptr;
        cs.bl_cap = 0;
        const size_t xbytes = sizeof(double) * (size_t)(cap / kTile) * cc.npad * kRowLd;
        if ((e = cudaMalloc(&cs.Xcur, xbytes)) != cudaSuccess) return e;
        if ((e = cudaMalloc(&cs.flags, cap)) != cudaSuccess) return e;
        if ((e = cudaMemset(cs.Xcur, 0, xbytes)) != cudaSuccess) return e;
        if ((e = cudaMemset(cs.flags, 0, cap)) != cudaSuccess) return e;
        cs.bl_cap = cap;
    }
    if (!with_proposal_buffers) return cudaSuccess;
    for (int b = 1; b <= 2; b++) {
        BlockedWork& w = cc.work[b];
        if (w.cap != cs.bl_cap) {
            cudaFree(w.X); cudaFree(w.lp); cudaFree(w.d2); cudaFree(w.lz);
            w = BlockedWork{};
            if ((e = chain_ensure_work(cc, w, cs.bl_cap)) != cudaSuccess) return e;
        }
    }
    return cudaSuccess;
}

static void chain_imhk_args(BlockedCtx& bc, ChainSet& cs, ChainArgs& a, int weight_mode) {
    ChainCtx& cc = bc.chain;
    a.Xbuf[0] = cc.work[1].X;   // proposals of even tags (buffer id 1)
    a.Xbuf[1] = cc.work[2].X;   // proposals of odd tags (buffer id 2)
    a.lp = cc.work[1].lp; a.d2 = cc.work[1].d2;
    a.lz = cc.work[1].lz;       // exact weights: the proposal's partial log-weight is parked here between pairs
    a.Xcur = cs.Xcur;
    a.loc = cs.flags;
    a.lw_cur = cs.logw;
    a.counts = reinterpret_cast<long long*>(cs.counts);
    a.S = cs.chains;
    a.seed = cs.seed;
    a.first_id = cs.first_id;
    a.exact = weight_mode == LGS_WEIGHT_EXACT ? 1 : 0;
}

// every chain's current state into Xcur (states accepted by step-by-step launches may still live in a proposal buffer)
static cudaError_t chain_imhk_flush_states(BlockedCtx& bc, ChainSet& cs, cudaStream_t st) {
    if (!cs.bl_loc_dirty) return cudaSuccess;
    ChainArgs a{};
    chain_imhk_args(bc, cs, a, LGS_WEIGHT_REFERENCE);
    a.nsteps = 0;
    a.flush = 1;
    cudaError_t e = chain_launch(bc, a, true, false, false, st);
    if (e == cudaSuccess) cs.bl_loc_dirty = false;
    return e;
}

// Step-parallel IMHK: see ChainArgs::tiles_per_step.  Used when the chain set is small against one batch of the chain
// kernel (BASELINE config 2: 4096 chains at n = 64 are 32 tiles -- a step-by-step launch occupies 32 of 148 SMs and every
// chain waits for its own serial SampleZ walk; here one launch draws the proposals of up to batch / chains steps at the
// kernel's full batched rate).  Proposals, accept draws and weights are keyed and computed exactly as in the
// step-by-step launches: the chains are bit-identical either way (tests/test_gpu_chain.py).
constexpr int kMaxKeepPerLaunch = 32;
constexpr int64_t kMaxKeepInLaunch = 4;   // snapshots per step-by-step launch (ChainArgs::keep_thin): 4 x the chain set in HBM

static cudaError_t launch_imhk_chain_steppar(BlockedCtx& bc, ChainSet& cs, int64_t burn_in, int64_t keep, int64_t thin,
                                             int weight_mode, double* V, int64_t* X, cudaStream_t st) {
    ChainCtx& cc = bc.chain;
    cudaError_t e;
    if ((e = chain_ensure_chains(bc, cs, false)) != cudaSuccess) return e;
    if ((e = chain_imhk_flush_states(bc, cs, st)) != cudaSuccess) return e;
    const bool exact = weight_mode == LGS_WEIGHT_EXACT;
    const int64_t C = cs.chains, cpad = cs.bl_cap;
    const int tps = (int)(cpad / kTile);
    const int n = bc.n;
    int64_t max_steps = cc.batch / cpad;
    if (max_steps > 4096) max_steps = 4096;
    BlockedWork& w = cc.work[0];
    const bool want_states = (V || X) && keep > 0;
    if (!cs.bl_init && (e = cudaMemsetAsync(cs.flags, 0, (size_t)cs.bl_cap, st)) != cudaSuccess) return e;
    // scan outputs + kept-state staging
    const size_t need_src = sizeof(int) * (size_t)(kMaxKeepPerLaunch + 1) * cpad;
    if (cs.sp_src_cap < need_src) {
        cudaFree(cs.sp_src);
        cs.sp_src = nullptr, cs.sp_src_cap = 0;
        if ((e = cudaMalloc(&cs.sp_src, need_src)) != cudaSuccess) return e;
        cs.sp_src_cap = need_src;
    }
    const size_t slot_doubles = (size_t)tps * cc.npad * kRowLd;
    if (want_states) {
        const int64_t kmax = keep < kMaxKeepPerLaunch ? keep : kMaxKeepPerLaunch;
        const size_t need = sizeof(double) * slot_doubles * (size_t)kmax;
        if (cs.sp_keep_cap < need) {
            cudaFree(cs.sp_keep);
            cs.sp_keep = nullptr, cs.sp_keep_cap = 0;
            if ((e = cudaMalloc(&cs.sp_keep, need)) != cudaSuccess) return e;
            if ((e = cudaMemsetAsync(cs.sp_keep, 0, need, st)) != cudaSuccess) return e;   // padding columns stay finite
            cs.sp_keep_cap = need;
        }
    }
    int64_t remaining = burn_in + keep * thin, done = 0;
    while (remaining > 0 || !cs.bl_init) {
        const int init = cs.bl_init ? 0 : 1;
        int64_t now = remaining + init < max_steps ? remaining + init : max_steps;
        // kept states inside this launch: call-steps g in (done, done + now - init], kept when g > burn_in and
        // (g - burn_in) % thin == 0; local proposal index j = g - done - 1 + init
        int64_t g0 = -1, nkeep = 0;
        if (want_states) {
            int64_t g = done + 1 > burn_in + thin ? done + 1 : burn_in + thin;
            const int64_t r = (g - burn_in) % thin;
            if (r) g += thin - r;
            if (g <= done + now - init) {
                g0 = g;
                nkeep = (done + now - init - g) / thin + 1;
                if (nkeep > kMaxKeepPerLaunch) {
                    nkeep = kMaxKeepPerLaunch;
                    now = g0 + (nkeep - 1) * thin - done + init;   // end the launch at its last kept state
                }
            }
        }
        if ((e = chain_ensure_work(cc, w, now * cpad)) != cudaSuccess) return e;
        {   // proposals of `now` steps of every chain: one Klein-mode batch
            ChainArgs a{};
            a.Xbuf[0] = a.Xbuf[1] = w.X;
            a.lp = w.lp; a.d2 = w.d2; a.lz = w.lz;
            a.S = now * cpad;
            a.seed = cs.seed; a.first_id = cs.first_id;
            a.tag = init ? 0u : (uint32_t)(cs.bl_steps + 1);
            a.nsteps = 1;
            a.tiles_per_step = tps;
            a.ntiles = (int)(now * tps);
            if ((e = chain_launch(bc, a, false, exact, false, st)) != cudaSuccess) return e;
        }
        int* keep_src = cs.sp_src;
        int* final_src = cs.sp_src + (size_t)kMaxKeepPerLaunch * cpad;
        {
            ProfScope ps(bc.prof, LGS_PROF_OTHER, st);
            ScanArgs sa{};
            sa.C = C; sa.cpad = cpad; sa.nsteps = (int)now; sa.first_is_init = init; sa.exact = exact ? 1 : 0;
            sa.seed = cs.seed; sa.first_id = cs.first_id; sa.tag = init ? 0u : (uint32_t)(cs.bl_steps + 1);
            sa.sum_lognorm = bc.sum_lognorm; sa.lz = w.lz; sa.lw_cur = cs.logw;
            sa.counts = reinterpret_cast<long long*>(cs.counts);
            sa.keep_first = nkeep ? (int)(g0 - done - 1 + init) : -1; sa.thin = (int)thin; sa.nkeep = (int)nkeep;
            sa.keep_src = keep_src; sa.final_src = final_src;
            imhk_scan_kernel<<<(unsigned)((C + 127) / 128), 128, 0, st>>>(sa);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
            const int ystrips = cc.npad >= 256 ? 4 : 1;
            if (nkeep) {
                dim3 grid((unsigned)tps, (unsigned)ystrips, (unsigned)nkeep);
                imhk_gather_kernel<<<grid, 128, 0, st>>>(w.X, cs.Xcur, cs.sp_keep, keep_src, C, cpad, tps, cc.npad, 0);
                if ((e = cudaGetLastError()) != cudaSuccess) return e;
            }
        }
        if (nkeep) {
            const int64_t kidx0 = (g0 - burn_in) / thin - 1;
            if (V) {   // lattice points of the kept states: a points-only launch over nkeep x chains columns
                ChainArgs a{};
                chain_imhk_args(bc, cs, a, LGS_WEIGHT_REFERENCE);
                a.Xcur = cs.sp_keep;
                a.nsteps = 0;
                a.flush = 1;
                a.V = V + (size_t)kidx0 * n;
                a.ldv = keep * (int64_t)n;
                a.v_slot_stride = n;
                a.tiles_per_step = tps;
                a.ntiles = (int)(nkeep * tps);
                if ((e = chain_launch(bc, a, true, false, false, st)) != cudaSuccess) return e;
            }
            if (X) {
                ProfScope ps(bc.prof, LGS_PROF_OTHER, st);
                dim3 grid((unsigned)((C + 31) / 32), (unsigned)((n + 31) / 32));
                for (int64_t m = 0; m < nkeep; m++) {
                    export_coeffs_tiled_kernel<<<grid, 256, 0, st>>>(cs.sp_keep + (size_t)m * slot_doubles, cc.npad, n, C,
                                                                     X + (size_t)(kidx0 + m) * n, keep * (int64_t)n);
                }
                if ((e = cudaGetLastError()) != cudaSuccess) return e;
            }
        }
        {   // the state after the launch's last proposal
            ProfScope ps(bc.prof, LGS_PROF_OTHER, st);
            const int ystrips = cc.npad >= 256 ? 4 : 1;
            dim3 grid((unsigned)tps, (unsigned)ystrips, 1u);
            imhk_gather_kernel<<<grid, 128, 0, st>>>(w.X, cs.Xcur, cs.Xcur, final_src, C, cpad, tps, cc.npad, 1);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
        }
        const int64_t stepped = now - init;
        cs.bl_steps += stepped;
        cs.bl_init = true;
        remaining -= stepped;
        done += stepped;
    }
    return cudaSuccess;
}

cudaError_t launch_imhk_chain(BlockedCtx& bc, ChainSet& cs, int64_t burn_in, int64_t keep, int64_t thin, int weight_mode,
                              double* V, int64_t* X, cudaStream_t st, const ChainReplay* rp) {
    cudaError_t e;
    {
        // few chains: draw the proposals of many steps per launch (LGS_IMHK_STEPPAR=0 / 1 forces a mode: development A/B)
        const char* fv = getenv("LGS_IMHK_STEPPAR");   // read per call: the tests switch it between calls
        const int force = fv ? atoi(fv) : -1;
        const int64_t cpad = (cs.chains + kTile - 1) / kTile * kTile;
        const bool few = 4 * cpad <= bc.chain.batch;
        if (!rp && (force == 1 || (force != 0 && few)) && 2 * cpad <= bc.chain.batch)
            return launch_imhk_chain_steppar(bc, cs, burn_in, keep, thin, weight_mode, V, X, st);
    }
    if ((e = chain_ensure_chains(bc, cs)) != cudaSuccess) return e;
    cs.bl_loc_dirty = true;
    int64_t call_done = 0;   // replay: steps of this call performed by earlier launches
    if (rp) {  // stream positions are per call: chain k reads U[k * stride ...) from its start
        if (!cs.upos) return cudaErrorInvalidValue;
        if ((e = cudaMemsetAsync(cs.upos, 0, sizeof(int64_t) * (size_t)cs.chains, st)) != cudaSuccess) return e;
    }
    const bool exact = weight_mode == LGS_WEIGHT_EXACT;
    const int64_t C = cs.chains;
    const int n = bc.n;
    constexpr int64_t kMaxSteps = 64;  // steps per launch when nothing is kept (bounds one launch to well under a second)
    // run `cnt` steps (the first may be the initial state); flush != 0: kept state number kidx ends the launch
    auto segment = [&](int64_t cnt, bool flush, int64_t kidx) -> cudaError_t {
        while (cnt > 0 || flush || !cs.bl_init) {
            const bool init = !cs.bl_init;
            int64_t now = cnt + (init ? 1 : 0);
            bool last = true;
            if (now > kMaxSteps) { now = kMaxSteps; last = false; }
            ChainArgs a{};
            chain_imhk_args(bc, cs, a, weight_mode);
            a.first_is_init = init ? 1 : 0;
            a.tag = init ? 0u : (uint32_t)(cs.bl_steps + 1);
            a.nsteps = (int)now;
            a.flush = (flush && last) ? 1 : 0;
            if (a.flush && V) { a.V = V + (size_t)kidx * n; a.ldv = keep * (int64_t)n; }
            if (rp) {
                a.U = rp->U; a.u_stride = rp->stride; a.fallback = rp->fallback;
                a.upos = reinterpret_cast<long long*>(cs.upos);
                a.accept_out = rp->accept; a.logw_out = rp->logw;
                a.call_steps = burn_in + keep * thin;
                a.step_off = call_done;
            }
            cudaError_t er = chain_launch(bc, a, true, exact, false, st);
            if (er != cudaSuccess) return er;
            const int64_t done = now - (init ? 1 : 0);
            call_done += done;
            cs.bl_steps += done;
            cs.bl_init = true;
            cnt -= done;
            if (a.flush && X) {
                ProfScope ps(bc.prof, LGS_PROF_OTHER, st);
                dim3 grid((unsigned)((C + 31) / 32), (unsigned)((n + 31) / 32));
                export_coeffs_tiled_kernel<<<grid, 256, 0, st>>>(cs.Xcur, bc.chain.npad, n, C, X + (size_t)kidx * n,
                                                                 keep * (int64_t)n);
                if ((er = cudaGetLastError()) != cudaSuccess) return er;
            }
            if (last) break;
        }
        return cudaSuccess;
    };
    if (!cs.bl_init) {
        if ((e = cudaMemsetAsync(cs.flags, 0, (size_t)cs.bl_cap, st)) != cudaSuccess) return e;
    }
    const bool want_states = (V || X) && keep > 0;
    if (!want_states) {
        const int64_t steps = burn_in + keep * thin;
        if (steps > 0 || !cs.bl_init) return segment(steps, false, 0);
        return cudaSuccess;
    }
    if (burn_in > 0 || !cs.bl_init) {
        if ((e = segment(burn_in, false, 0)) != cudaSuccess) return e;
    }
    const bool keep_in_launch = [] { const char* v = getenv("LGS_IMHK_KEEP_IN_LAUNCH"); return !(v && v[0] == '0'); }();   // per call
    if (!keep_in_launch || thin > kMaxSteps || !V) {
        // one launch per kept state: its last step moves every state into Xcur and the points are contracted from there
        for (int64_t k = 0; k < keep; k++)
            if ((e = segment(thin, true, k)) != cudaSuccess) return e;
        return cudaSuccess;
    }
    // Several kept states per launch (ChainArgs::keep_thin): a launch of g * thin steps snapshots the states after every
    // thin-th step and contracts the snapshots' lattice points as filler while the chains step on.
    const int64_t gmax = kMaxKeepInLaunch < kMaxSteps / thin ? kMaxKeepInLaunch : kMaxSteps / thin;
    const size_t snap = (size_t)(cs.bl_cap / kTile) * bc.chain.npad * kRowLd;   // doubles per snapshot
    {
        const int64_t g = keep < gmax ? keep : gmax;
        if (cs.xkeep_cap < sizeof(double) * snap * (size_t)g) {
            cudaFree(cs.Xkeep);
            cs.Xkeep = nullptr, cs.xkeep_cap = 0;
            if ((e = cudaMalloc(&cs.Xkeep, sizeof(double) * snap * (size_t)g)) != cudaSuccess) return e;
            if ((e = cudaMemsetAsync(cs.Xkeep, 0, sizeof(double) * snap * (size_t)g, st)) != cudaSuccess) return e;   // padding columns stay finite
            cs.xkeep_cap = sizeof(double) * snap * (size_t)g;
        }
    }
    for (int64_t k0 = 0; k0 < keep; k0 += gmax) {
        const int64_t g = keep - k0 < gmax ? keep - k0 : gmax;
        ChainArgs a{};
        chain_imhk_args(bc, cs, a, weight_mode);
        a.tag = (uint32_t)(cs.bl_steps + 1);
        a.nsteps = (int)(g * thin);
        a.keep_thin = (int)thin; a.keep_first = (int)thin - 1; a.nkeep = (int)g;
        a.Xkeep = cs.Xkeep; a.keep_stride = (int64_t)snap;
        a.V = V + (size_t)k0 * n; a.ldv = keep * (int64_t)n; a.v_slot_stride = n;
        if (rp) {
            a.U = rp->U; a.u_stride = rp->stride; a.fallback = rp->fallback;
            a.upos = reinterpret_cast<long long*>(cs.upos);
            a.accept_out = rp->accept; a.logw_out = rp->logw;
            a.call_steps = burn_in + keep * thin;
            a.step_off = call_done;
        }
        if ((e = chain_launch(bc, a, true, exact, false, st)) != cudaSuccess) return e;
        call_done += g * thin;
        cs.bl_steps += g * thin;
        if (X) {
            ProfScope ps(bc.prof, LGS_PROF_OTHER, st);
            dim3 grid((unsigned)((C + 31) / 32), (unsigned)((n + 31) / 32));
            for (int64_t m = 0; m < g; m++)
                export_coeffs_tiled_kernel<<<grid, 256, 0, st>>>(cs.Xkeep + (size_t)m * snap, bc.chain.npad, n, C,
                                                                 X + (size_t)(k0 + m) * n, keep * (int64_t)n);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
        }
    }
    return cudaSuccess;
}

// lgs_attach_histogram: the coordinate -> bins table of the lattice-point epilogue (synchronous; hist == nullptr detaches)
cudaError_t chain_attach_histogram(BlockedCtx& bc, const int32_t* dims, int n_dims, double lo, double width, int nbins,
                                   int64_t* hist) {
    ChainCtx& cc = bc.chain;
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return e;
    cc.hist = nullptr;
    if (!hist) return cudaSuccess;
    std::vector<int16_t> map((size_t)cc.npad, (int16_t)-1);
    for (int i = 0; i < n_dims; i++) map[dims[i]] = (int16_t)i;   // a coordinate listed twice counts into its last row
    if (!cc.hist_map && (e = cudaMalloc(&cc.hist_map, sizeof(int16_t) * cc.npad)) != cudaSuccess) return e;
    if ((e = cudaMemcpy(cc.hist_map, map.data(), sizeof(int16_t) * cc.npad, cudaMemcpyHostToDevice)) != cudaSuccess) return e;
    cc.hist = reinterpret_cast<unsigned long long*>(hist);
    cc.hist_lo = lo;
    cc.hist_inv_width = 1.0 / width;
    cc.hist_nbins = nbins;
    return cudaSuccess;
}

cudaError_t launch_imhk_chain_get_state(BlockedCtx& bc, ChainSet& cs, int64_t* counts, int64_t* X, double* V,
                                        double* logw, cudaStream_t st) {
    const int64_t C = cs.chains;
    const int n = bc.n;
    cudaError_t e;
    if (counts && (e = cudaMemcpyAsync(counts, cs.counts, sizeof(int64_t) * 2 * C, cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
    if (logw && (e = cudaMemcpyAsync(logw, cs.logw, sizeof(double) * C, cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
    if (X || V) {
        if ((e = chain_ensure_chains(bc, cs)) != cudaSuccess) return e;
        ChainArgs a{};
        chain_imhk_args(bc, cs, a, LGS_WEIGHT_REFERENCE);
        a.nsteps = 0;
        a.flush = 1;
        a.V = V;
        a.ldv = n;
        // the chains' current states are not new samples: an attached histogram must not see them again
        bc.chain.hist_pause = true;
        e = chain_launch(bc, a, true, false, false, st);
        bc.chain.hist_pause = false;
        if (e != cudaSuccess) return e;
        cs.bl_loc_dirty = false;
        if (X) {
            dim3 grid((unsigned)((C + 31) / 32), (unsigned)((n + 31) / 32));
            export_coeffs_tiled_kernel<<<grid, 256, 0, st>>>(cs.Xcur, bc.chain.npad, n, C, X, n);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
        }
    }
    return cudaSuccess;
}

// lgs_imhk_set_state on the chain kernel's state: coefficients into the tiled current-state matrix, every chain's
// location = Xcur, log-weights and counters as given, `steps_done` = the Philox stream index the run continues from
cudaError_t launch_imhk_chain_set_state(BlockedCtx& bc, ChainSet& cs, const int64_t* X, const double* logw,
                                        const int64_t* counts, int64_t steps_done, cudaStream_t st) {
    cudaError_t e;
    if ((e = chain_ensure_chains(bc, cs)) != cudaSuccess) return e;
    const int64_t C = cs.chains;
    const int n = bc.n;
    dim3 grid((unsigned)((C + 31) / 32), (unsigned)((n + 31) / 32));
    import_coeffs_tiled_kernel<<<grid, 256, 0, st>>>(X, n, bc.chain.npad, n, C, cs.Xcur);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(cs.flags, 0, (size_t)cs.bl_cap, st)) != cudaSuccess) return e;
    if ((e = cudaMemcpyAsync(cs.logw, logw, sizeof(double) * C, cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
    if (counts) e = cudaMemcpyAsync(cs.counts, counts, sizeof(int64_t) * 2 * C, cudaMemcpyDeviceToDevice, st);
    else e = cudaMemsetAsync(cs.counts, 0, sizeof(int64_t) * 2 * C, st);
    if (e != cudaSuccess) return e;
    cs.bl_steps = steps_done;
    cs.bl_init = true;
    cs.bl_loc_dirty = false;
    return cudaSuccess;
}

}  // namespace lgs
