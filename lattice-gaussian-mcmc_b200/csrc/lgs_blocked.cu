// Blocked path: the throughput implementation of Klein's sampler (klein.py:181-220) for a batch of samples.
//
// Coordinates are cut into blocks of kPB = 32.  For each block, from the last coordinate to the first:
//   1. contraction  C[rows of the block, s] = t - R[rows, done] * X[done, s]   (all samples s at once)
//      -- a dense fp64 GEMM on the DMMA pipe (mma.sync m16n8k8 f64), operands staged with cp.async;
//   2. panel kernel: one thread per sample walks the 32 coordinates of the block serially (the in-block triangle of
//      R sits in the constant bank as kernel parameters), drawing x_i ~ D_{Z, sigma_i, c_i} with the fast SampleZ and
//      Philox, and accumulates the Klein log-density terms.
// Two 32-blocks share one 64-row contraction over the already finished coordinates (halves the re-reads of X);
// the second block of a pair gets the first block's contribution from a small 32x32 contraction.
// Finally the lattice points V = X B^T are one more dense GEMM whose epilogue writes the reference's row-major
// [S, n] layout (klein.py:218,322).
//
// Data layout (HBM): X is coordinate-major, X[i * ld + s], one fp64 per (coordinate, sample); rows of the block being
// processed hold the running centers C until the panel kernel overwrites them with the integers it draws.
#include "lgs_internal.h"

#include <cmath>
#include <cstdlib>
#include <vector>

namespace lgs {

namespace {

// ---------------------------------------------------------------------------------------------------------
// panel kernel
// ---------------------------------------------------------------------------------------------------------
constexpr int kPanelThreads = 128;

__device__ __noinline__ double samplez_fast_call(double mean, double inv2s2, double q, double sq, double znorm,
                                                 double eps1, int kind, int dmax, double u, double* zsum) {
    FastZ fz{inv2s2, q, sq, znorm, eps1, kind, dmax};
    return (double)samplez_fast(mean, fz, u, zsum);
}

// One thread per sample.  The 32 coordinates of the block are walked in sub-blocks of kSub (upper first).  Only the
// kSub running centers of the sub-block live in registers, in a ROTATING window: slot kSub-1 always holds the
// coordinate being drawn, and the triangle update writes every slot one place up (c[j+1] = fma(T[j], x, c[j])), so the
// coordinate loop is a rolled loop with a compact body (the fully unrolled form was 10k instructions per kernel and
// stalled on instruction fetch) and the register budget allows 4-6 CTAs per SM -- the serial SampleZ chain is
// latency-bound, warps in flight are what hides it.  Constants come from the kernel-parameter constant bank through
// uniform loads: P.nrt row gi = [kSub shifted triangle entries (zero where the slot is empty) | rectangle entries for
// the rows below the sub-block].  After a sub-block is drawn its contribution to the rows below goes through the
// thread's own column of X (coalesced, L1/L2-resident).
#ifndef LGS_PANEL_SUB
#define LGS_PANEL_SUB 8
#endif
#ifndef LGS_PANEL_MINB
#define LGS_PANEL_MINB 8
#endif
constexpr int kSub = LGS_PANEL_SUB;
static_assert(kPB % kSub == 0, "sub-block width must divide the panel width");

template <bool WANT_LZ>
__global__ void __launch_bounds__(kPanelThreads, LGS_PANEL_MINB)
klein_panel_kernel(const __grid_constant__ PanelParams P, double* __restrict__ X, int64_t ld, int i0, int64_t S,
                   uint64_t seed, uint64_t first_id, uint32_t tag, const int64_t* __restrict__ tag_dev, int first,
                   double* __restrict__ lp, double* __restrict__ d2, double* __restrict__ lz) {
    const int64_t s = (int64_t)blockIdx.x * kPanelThreads + threadIdx.x;
    if (s >= S) return;
    double* __restrict__ Xs = X + (size_t)i0 * ld + s;
    double lpa = 0.0, d2a = 0.0, lza = 0.0;
    if (!first) {
        lpa = lp[s];
        d2a = d2[s];
        if (WANT_LZ) lza = lz[s];
    }
    if (tag_dev) tag = (uint32_t)(*tag_dev);
    const Philox gen(seed);
    const uint64_t id = first_id + (uint64_t)s;
    uint4 w = make_uint4(0, 0, 0, 0);
    int wpair = -1;
#pragma unroll 1
    for (int base = kPB - kSub; base >= 0; base -= kSub) {
        double c[kSub];
        {
            const double* __restrict__ xl = Xs + (size_t)base * ld;
#pragma unroll
            for (int k = 0; k < kSub; k++, xl += ld) c[k] = *xl;
        }
        double* __restrict__ xp = Xs + (size_t)(base + kSub - 1) * ld;  // row of the coordinate being drawn
#pragma unroll 1
        for (int gi = base + kSub - 1; gi >= base; gi--, xp -= ld) {
            const CoordConst& cc = P.cc[gi];
            const int4 ci = P.ci[gi];
            const double mean = c[kSub - 1] * cc.inv_r;
            double x;
            if (ci.w) {
                x = rint(mean);
            } else {
                if ((ci.z >> 1) != wpair) {
                    wpair = ci.z >> 1;
                    w = gen((uint32_t)id, (uint32_t)(id >> 32), (uint32_t)wpair, tag);
                }
                const bool odd = ci.z & 1;
                const double u = u53(odd ? w.z : w.x, odd ? w.w : w.y);
                if (ci.x <= 1) {
                    double logz;
                    x = samplez_smooth<WANT_LZ>(mean, cc.a, cc.q, cc.sq, cc.znorm, cc.eps1, cc.lognz, ci.x, ci.y, u,
                                                &logz);
                    if (WANT_LZ) lza += logz;
                } else {
                    double zs;
                    x = samplez_fast_call(mean, cc.a, cc.q, cc.sq, cc.znorm, cc.eps1, ci.x, ci.y, u, &zs);
                    if (WANT_LZ) lza += log(zs);
                }
            }
            *xp = x;
            const double dl = x - mean;
            const double dl2 = dl * dl;
            lpa = fma(dl2, cc.a, lpa);
            d2a = fma(dl2, cc.r2, d2a);
            const double* __restrict__ T = &P.nrt[gi * kPB];
#pragma unroll
            for (int j = kSub - 2; j >= 0; j--) c[j + 1] = fma(T[j], x, c[j]);
            c[0] = 0.0;
        }
        // contribution of this sub-block to the rows below it: rows [0, base) of the block, kSub at a time
#pragma unroll 1
        for (int r0 = 0; r0 < base; r0 += kSub) {
            double a[kSub];
            double* __restrict__ xr = Xs + (size_t)r0 * ld;
#pragma unroll
            for (int r = 0; r < kSub; r++) a[r] = xr[(size_t)r * ld];
            const double* __restrict__ xj = Xs + (size_t)(base + kSub - 1) * ld;
            const double* __restrict__ N = &P.nrt[(base + kSub - 1) * kPB + kSub + r0];
#pragma unroll 1
            for (int j = 0; j < kSub; j++, xj -= ld, N -= kPB) {
                const double xv = *xj;
#pragma unroll
                for (int r = 0; r < kSub; r++) a[r] = fma(N[r], xv, a[r]);
            }
#pragma unroll
            for (int r = 0; r < kSub; r++) xr[(size_t)r * ld] = a[r];
        }
    }
    lp[s] = lpa;
    d2[s] = d2a;
    if (WANT_LZ) lz[s] = lza;
}

// ---------------------------------------------------------------------------------------------------------
// DMMA contraction kernel:   D[m, s] = sum_{k in [k0, k1)} A[r0 + m, k] * X[k, s]
//   MODE_PANEL : X[r0 + m, s]  = tvec[r0 + m] - D      (start the centers of a block)
//   MODE_ACCUM : X[r0 + m, s] -= D                      (add a finished block's contribution)
//   MODE_POINT : V[s * ldv + r0 + m] = D                (lattice points, row-major per sample)
// CTA tile BM x 128 samples, warps 32 x 32, k-tile 16, kStages-deep cp.async pipeline.
// ---------------------------------------------------------------------------------------------------------
constexpr int MODE_PANEL = 0, MODE_ACCUM = 1, MODE_POINT = 2;
#ifndef LGS_GEMM_STAGES
#define LGS_GEMM_STAGES 4
#endif
constexpr int kBN = 128, kBK = 16, kStages = LGS_GEMM_STAGES;
constexpr int kLdA = kBK + 4;   // 20 doubles: conflict-free LDS.64 for the A fragments (row stride = 4 mod 16)
constexpr int kLdB = kBN + 4;   // 132 doubles: same for the B fragments

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ void dmma_16x8x8(double (&d)[4], const double (&a)[4], const double (&b)[2]) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
        : "+d"(d[0]), "+d"(d[1]), "+d"(d[2]), "+d"(d[3])
        : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}

template <int BM>
struct GemmSmem {
    double A[kStages][BM][kLdA];
    double B[kStages][kBK][kLdB];
};

template <int BM, int MODE>
__global__ void __launch_bounds__(BM * 4)
contract_kernel(const double* __restrict__ A, int64_t lda, double* __restrict__ X, int64_t ldx, int r0_base, int k0,
                int k1, const double* __restrict__ tvec, int64_t S, double* __restrict__ V, int64_t ldv, int n_valid) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GemmSmem<BM>& sm = *reinterpret_cast<GemmSmem<BM>*>(smem_raw);
    constexpr int NT = BM * 4;
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int wm = warp >> 2, wn = warp & 3;
    const int g = lane >> 2, t = lane & 3;
    const int r0 = r0_base + blockIdx.x * BM;            // blockIdx.x: row tile (fastest, shares the X tile in L2)
    const int64_t s0 = (int64_t)blockIdx.y * kBN;         // blockIdx.y: sample tile
    const int nk = (k1 - k0) / kBK;

    double acc[2][4][4];
#pragma unroll
    for (int a = 0; a < 2; a++)
#pragma unroll
        for (int b = 0; b < 4; b++)
#pragma unroll
            for (int c = 0; c < 4; c++) acc[a][b][c] = 0.0;

    auto load_tile = [&](int kt, int stage) {
        const int kb = k0 + kt * kBK;
        // A tile: BM rows x 8 chunks of 16 bytes
#pragma unroll
        for (int c = tid; c < BM * 8; c += NT) {
            int row = c >> 3, ch = c & 7;
            cp_async16(&sm.A[stage][row][ch * 2], A + (size_t)(r0 + row) * lda + kb + ch * 2);
        }
        // B tile: 16 rows x 64 chunks
#pragma unroll
        for (int c = tid; c < kBK * 64; c += NT) {
            int row = c >> 6, ch = c & 63;
            cp_async16(&sm.B[stage][row][ch * 2], X + (size_t)(kb + row) * ldx + s0 + ch * 2);
        }
    };

#pragma unroll
    for (int st = 0; st < kStages - 1; st++) {
        if (st < nk) load_tile(st, st);
        cp_async_commit();
    }
    for (int kt = 0; kt < nk; kt++) {
        cp_async_wait<kStages - 2>();
        __syncthreads();
        {
            int nx = kt + kStages - 1;
            if (nx < nk) load_tile(nx, nx % kStages);
            cp_async_commit();
        }
        const int stage = kt % kStages;
#pragma unroll
        for (int kk = 0; kk < kBK; kk += 8) {
            double af[2][4], bf[4][2];
#pragma unroll
            for (int mt = 0; mt < 2; mt++) {
                const double* ap = &sm.A[stage][wm * 32 + mt * 16 + g][kk + t];
                af[mt][0] = ap[0];
                af[mt][1] = ap[8 * kLdA];
                af[mt][2] = ap[4];
                af[mt][3] = ap[8 * kLdA + 4];
            }
#pragma unroll
            for (int nt = 0; nt < 4; nt++) {
                const double* bp = &sm.B[stage][kk + t][wn * 32 + nt * 8 + g];
                bf[nt][0] = bp[0];
                bf[nt][1] = bp[4 * kLdB];
            }
#pragma unroll
            for (int mt = 0; mt < 2; mt++)
#pragma unroll
                for (int nt = 0; nt < 4; nt++) dmma_16x8x8(acc[mt][nt], af[mt], bf[nt]);
        }
    }
    cp_async_wait<0>();

    // epilogue: fragment (mt, nt): rows g / g+8 of the 16-row tile, columns 2t, 2t+1 of the 8-column tile
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int row = r0 + wm * 32 + mt * 16 + g + h * 8;
#pragma unroll
            for (int nt = 0; nt < 4; nt++) {
                const int64_t col = s0 + wn * 32 + nt * 8 + t * 2;
                const double v0 = acc[mt][nt][h * 2], v1 = acc[mt][nt][h * 2 + 1];
                if (MODE == MODE_POINT) {
                    if (row < n_valid) {
                        if (col < S) V[(size_t)col * ldv + row] = v0;
                        if (col + 1 < S) V[(size_t)(col + 1) * ldv + row] = v1;
                    }
                } else {
                    double2* dst = reinterpret_cast<double2*>(X + (size_t)row * ldx + col);
                    double2 o;
                    if (MODE == MODE_PANEL) {
                        const double tv = tvec[row];
                        o.x = tv - v0;
                        o.y = tv - v1;
                    } else {
                        o = *dst;
                        o.x -= v0;
                        o.y -= v1;
                    }
                    *dst = o;
                }
            }
        }
    }
}

// X (coordinate-major doubles) -> int64 [S, n] row-major, through a 32x32 shared-memory transpose
__global__ void __launch_bounds__(256)
export_coeffs_kernel(const double* __restrict__ X, int64_t ld, int n, int64_t S, int64_t* __restrict__ out) {
    __shared__ double tile[32][33];
    const int64_t s0 = (int64_t)blockIdx.x * 32;
    const int i0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {
        int i = i0 + r;
        int64_t s = s0 + tx;
        tile[r][tx] = (i < n && s < S) ? X[(size_t)i * ld + s] : 0.0;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        int64_t s = s0 + r;
        int i = i0 + tx;
        if (s < S && i < n) out[(size_t)s * n + i] = (int64_t)tile[tx][r];
    }
}

__global__ void finish_logq_kernel(const double* __restrict__ lp, double sum_lognorm, int64_t S,
                                   double* __restrict__ logq) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s < S) logq[s] = -lp[s] - sum_lognorm;
}

template <int BM, int MODE>
cudaError_t launch_contract(Prof& prof, const double* A, int64_t lda, double* X, int64_t ldx, int r0, int rows, int k0,
                            int k1, const double* tvec, int64_t S, double* V, int64_t ldv, int n_valid,
                            cudaStream_t st) {
    const size_t smem = sizeof(GemmSmem<BM>);
    ProfScope ps(prof, MODE == MODE_POINT ? LGS_PROF_POINT : (MODE == MODE_ACCUM ? LGS_PROF_ACCUM : LGS_PROF_CENTERS), st);
    dim3 grid((unsigned)(rows / BM), (unsigned)((S + kBN - 1) / kBN));
    contract_kernel<BM, MODE><<<grid, BM * 4, smem, st>>>(A, lda, X, ldx, r0, k0, k1, tvec, S, V, ldv, n_valid);
    return cudaGetLastError();
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
cudaError_t blocked_profile_read(BlockedCtx& bc, double* ms, int64_t* launches, int nclass) {
    for (int i = 0; i < nclass; i++) { ms[i] = 0.0; launches[i] = 0; }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return e;
    for (auto& r : bc.prof.recs) {
        float t = 0.f;
        if ((e = cudaEventElapsedTime(&t, r.e0, r.e1)) != cudaSuccess) return e;
        if (r.cls >= 0 && r.cls < nclass) { ms[r.cls] += t; launches[r.cls] += 1; }
        bc.prof.pool.push_back(r.e0);
        bc.prof.pool.push_back(r.e1);
    }
    bc.prof.recs.clear();
    return cudaSuccess;
}

void blocked_free(BlockedCtx& bc) {
    chain_free(bc);
    for (auto& r : bc.prof.recs) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
    for (auto& e : bc.prof.pool) cudaEventDestroy(e);
    bc.prof.recs.clear();
    bc.prof.pool.clear();
    cudaFree(bc.Rp);
    cudaFree(bc.Bp);
    cudaFree(bc.tp);
    auto free_work = [](BlockedWork& w) {
        cudaFree(w.X);
        cudaFree(w.lp);
        cudaFree(w.d2);
        cudaFree(w.lz);
        w = BlockedWork{};
    };
    for (auto& w : bc.work) free_work(w);
    bc.Rp = bc.Bp = bc.tp = nullptr;
    bc.ready = false;
}

cudaError_t blocked_init(BlockedCtx& bc, const lgs_desc& d, const HostConsts& hc) {
    const int n = d.n;
    const int npad = (n + kPB - 1) / kPB * kPB;
    const int mpad = (n + 63) / 64 * 64;
    bc.n = n;
    bc.npad = npad;
    bc.mpad = mpad;
    bc.two_sigma2 = 2.0 * d.sigma * d.sigma;
    std::vector<double> Rp((size_t)npad * npad, 0.0), Bp((size_t)mpad * npad, 0.0), tp(npad, 0.0);
    for (int i = 0; i < n; i++) {
        for (int j = i; j < n; j++) Rp[(size_t)i * npad + j] = d.R[(size_t)i * n + j];
        for (int j = 0; j < n; j++) Bp[(size_t)i * npad + j] = d.B[(size_t)i * n + j];
        tp[i] = d.t[i];
    }
    for (int i = n; i < npad; i++) Rp[(size_t)i * npad + i] = 1.0;
    bc.sum_lognorm = 0.0;
    for (int i = 0; i < n; i++) bc.sum_lognorm += hc.lognorm[i];

    bc.panels.assign(npad / kPB, PanelParams{});
    for (int b = 0; b < npad / kPB; b++) {
        PanelParams& P = bc.panels[b];
        const int i0 = b * kPB;
        for (int i = 0; i < kPB; i++) {
            const int gi = i0 + i;
            if (gi < n) {
                // row i of the packed constants (see klein_panel_kernel): slot j of the rotating window holds
                // coordinate i - (kSub - 1) + j of the block while coordinate i is drawn
                const int base = i / kSub * kSub;
                for (int j = 0; j + 1 < kSub; j++) {
                    const int k = i - (kSub - 1) + j;
                    if (k >= base) P.nrt[i * kPB + j] = -d.R[(size_t)(i0 + k) * n + gi];
                }
                for (int r = 0; r < base; r++) P.nrt[i * kPB + kSub + r] = -d.R[(size_t)(i0 + r) * n + gi];
                CoordConst& cc = P.cc[i];
                cc.inv_r = 1.0 / hc.rdiag[gi];
                cc.r2 = hc.rdiag[gi] * hc.rdiag[gi];
                cc.a = hc.inv2s2[gi];
                cc.q = hc.q[gi];
                cc.sq = hc.sq[gi];
                cc.znorm = hc.znorm[gi];
                cc.eps1 = hc.eps1[gi];
                cc.lognz = std::log(hc.znorm[gi]);
                P.ci[i] = make_int4(hc.kind[gi], hc.dmax[gi], hc.uidx[gi], hc.mode[gi]);
            } else {  // padding coordinate: center 0, R_ii = 1, rounded => x = 0, contributes nothing
                P.cc[i].inv_r = 1.0;
                P.ci[i] = make_int4(0, 0, 0, 1);
            }
        }
    }
    cudaError_t e;
    if ((e = cudaMalloc(&bc.Rp, sizeof(double) * Rp.size())) != cudaSuccess) return e;
    if ((e = cudaMalloc(&bc.Bp, sizeof(double) * Bp.size())) != cudaSuccess) return e;
    if ((e = cudaMalloc(&bc.tp, sizeof(double) * tp.size())) != cudaSuccess) return e;
    if ((e = cudaMemcpy(bc.Rp, Rp.data(), sizeof(double) * Rp.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return e;
    if ((e = cudaMemcpy(bc.Bp, Bp.data(), sizeof(double) * Bp.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return e;
    if ((e = cudaMemcpy(bc.tp, tp.data(), sizeof(double) * tp.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return e;
#define LGS_SMEM_ATTR(BM, MODE)                                                                                  \
    if ((e = cudaFuncSetAttribute(contract_kernel<BM, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,        \
                                  (int)sizeof(GemmSmem<BM>))) != cudaSuccess)                                   \
        return e;
    LGS_SMEM_ATTR(64, MODE_PANEL)
    LGS_SMEM_ATTR(32, MODE_PANEL)
    LGS_SMEM_ATTR(32, MODE_ACCUM)
    LGS_SMEM_ATTR(64, MODE_POINT)
#undef LGS_SMEM_ATTR
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&bc.sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (bc.sm_count <= 0) bc.sm_count = 148;
        bc.wave = (int64_t)bc.sm_count * 2 * kBN;  // 37 888 samples on a 148-SM B200
    }
    {
        const char* env = getenv("LGS_CHAIN");
        bc.use_chain = env ? atoi(env) : 1;
        if ((e = chain_init(bc, d, hc)) != cudaSuccess) return e;
    }
    bc.ready = true;
    return cudaSuccess;
}

static cudaError_t ensure_work(BlockedCtx& bc, BlockedWork& w, int64_t cap) {
    if (w.cap >= cap) return cudaSuccess;
    cudaFree(w.X); cudaFree(w.lp); cudaFree(w.d2); cudaFree(w.lz);
    w = BlockedWork{};
    cudaError_t e;
    if ((e = cudaMalloc(&w.X, sizeof(double) * (size_t)bc.npad * cap)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&w.lp, sizeof(double) * cap)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&w.d2, sizeof(double) * cap)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&w.lz, sizeof(double) * cap)) != cudaSuccess) return e;
    // columns past S of a partial tile feed the contractions (their results are never stored to V): keep them finite
    if ((e = cudaMemset(w.X, 0, sizeof(double) * (size_t)bc.npad * cap)) != cudaSuccess) return e;
    w.cap = cap;
    return cudaSuccess;
}

// One batch of S <= w.cap Klein proposals into w.X (coefficients), w.lp, w.d2, w.lz.
static cudaError_t klein_batch(BlockedCtx& bc, BlockedWork& w, int64_t S, uint64_t seed, uint64_t first_id,
                               uint32_t tag, const int64_t* tag_dev, bool want_lz, cudaStream_t st) {
    const int nb = bc.npad / kPB;
    const int64_t ld = w.cap;
    const int pblocks = (int)((S + kPanelThreads - 1) / kPanelThreads);
    cudaError_t e;
    auto panel = [&](int b, bool first) {
        ProfScope ps(bc.prof, LGS_PROF_PANEL, st);
        if (want_lz)
            klein_panel_kernel<true><<<pblocks, kPanelThreads, 0, st>>>(bc.panels[b], w.X, ld, b * kPB, S, seed, first_id,
                                                                        tag, tag_dev, first ? 1 : 0, w.lp, w.d2, w.lz);
        else
            klein_panel_kernel<false><<<pblocks, kPanelThreads, 0, st>>>(bc.panels[b], w.X, ld, b * kPB, S, seed,
                                                                         first_id, tag, tag_dev, first ? 1 : 0, w.lp,
                                                                         w.d2, w.lz);
        return cudaGetLastError();
    };
    int b = nb - 1;
    bool first = true;
    if (nb & 1) {  // odd number of 32-blocks: the top one stands alone
        if ((e = launch_contract<32, MODE_PANEL>(bc.prof, bc.Rp, bc.npad, w.X, ld, b * kPB, kPB, bc.npad, bc.npad, bc.tp, S,
                                                 nullptr, 0, 0, st)) != cudaSuccess) return e;
        if ((e = panel(b, first)) != cudaSuccess) return e;
        first = false;
        b--;
    }
    for (; b >= 1; b -= 2) {
        const int lo = (b - 1) * kPB;  // first coordinate of the pair
        // centers of both blocks from everything already drawn (coordinates >= lo + 64)
        if ((e = launch_contract<64, MODE_PANEL>(bc.prof, bc.Rp, bc.npad, w.X, ld, lo, 64, lo + 64, bc.npad, bc.tp, S, nullptr,
                                                 0, 0, st)) != cudaSuccess) return e;
        if ((e = panel(b, first)) != cudaSuccess) return e;
        first = false;
        // the upper block's contribution to the lower block
        if ((e = launch_contract<32, MODE_ACCUM>(bc.prof, bc.Rp, bc.npad, w.X, ld, lo, kPB, lo + kPB, lo + 64, bc.tp, S,
                                                 nullptr, 0, 0, st)) != cudaSuccess) return e;
        if ((e = panel(b - 1, false)) != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// outputs of one finished batch held in w: lattice points, int64 coefficients, log-densities
static cudaError_t klein_outputs(BlockedCtx& bc, BlockedWork& w, int64_t cnt, double* V, int64_t* X, double* logq,
                                 cudaStream_t st) {
    cudaError_t e;
    if (V) {
        if ((e = launch_contract<64, MODE_POINT>(bc.prof, bc.Bp, bc.npad, w.X, w.cap, 0, bc.mpad, 0, bc.npad, nullptr, cnt,
                                                 V, bc.n, bc.n, st)) != cudaSuccess) return e;
    }
    if (X) {
        ProfScope ps(bc.prof, LGS_PROF_OTHER, st);
        dim3 grid((unsigned)((cnt + 31) / 32), (unsigned)((bc.n + 31) / 32));
        export_coeffs_kernel<<<grid, 256, 0, st>>>(w.X, w.cap, bc.n, cnt, X);
    }
    if (logq) {
        ProfScope ps2(bc.prof, LGS_PROF_OTHER, st);
        finish_logq_kernel<<<(unsigned)((cnt + 255) / 256), 256, 0, st>>>(w.lp, bc.sum_lognorm, cnt, logq);
    }
    return cudaGetLastError();
}

cudaError_t launch_klein_blocked(BlockedCtx& bc, int64_t S, uint64_t seed, uint64_t first_id, double* V, int64_t* X,
                                 double* logq, cudaStream_t st) {
    if (S <= 0) return cudaSuccess;
    if (bc.use_chain && bc.chain.ready) return launch_klein_chain(bc, S, seed, first_id, V, X, logq, st);
    cudaError_t e;
    const int n = bc.n;
    // batches of up to four full waves of contraction tiles; a last partial wave of <= sm_count tiles runs one CTA
    // per SM at twice the speed, so any S costs what its flops cost to within half a wave
    const int64_t kMaxBatch = bc.wave * 4;
    int64_t cap = S < kMaxBatch ? (S + 127) / 128 * 128 : kMaxBatch;
    BlockedWork& w = bc.work[0];
    if ((e = ensure_work(bc, w, cap)) != cudaSuccess) return e;
    for (int64_t s0 = 0; s0 < S; s0 += w.cap) {
        const int64_t cnt = (S - s0) < w.cap ? (S - s0) : w.cap;
        if ((e = klein_batch(bc, w, cnt, seed, first_id + (uint64_t)s0, 0u, nullptr, false, st)) != cudaSuccess) return e;
        if ((e = klein_outputs(bc, w, cnt, V ? V + (size_t)s0 * n : nullptr, X ? X + (size_t)s0 * n : nullptr,
                               logq ? logq + s0 : nullptr, st)) != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// ---------------------------------------------------------------------------------------------------------
// IMHK on the blocked path (imhk.py:126-177): all chains of the handle step together.
// One step = one blocked Klein batch (the proposals) + the accept kernel + a predicated copy of accepted
// proposals into the current-state matrix; kept states go through the lattice-point contraction straight into
// the caller's [chain][keep][n] array.  Burn-in and thinning never leave the device.
// ---------------------------------------------------------------------------------------------------------
namespace {

// imhk.py:102-124 + 156-174 for one chain per thread.  init != 0: _initialize_chain (no accept test, no count).
// The accepted proposal is NOT copied anywhere: loc[ch] records that the chain's current state now lives in
// proposal buffer `buf` (1 or 2).  A location pointing at this buffer from two steps ago was saved into Xcur by
// imhk_materialize_kernel before the buffer was overwritten, so it is reset here.
__global__ void __launch_bounds__(256)
imhk_accept_kernel(int64_t C, const double* __restrict__ lp, const double* __restrict__ d2,
                   const double* __restrict__ lz, double sum_lognorm, double two_sigma2, int exact, int init,
                   uint64_t seed, uint64_t first_id, uint32_t tag, double* __restrict__ lw_cur,
                   int64_t* __restrict__ counts, uint8_t* __restrict__ loc, int buf) {
    const int64_t ch = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (ch >= C) return;
    // log pi(x) - log q(x): base.py:133-144 minus klein.py:246-269, with ||Bx - c||^2 = sum r_ii^2 (x_i - m_i)^2
    const double logq = -lp[ch] - sum_lognorm;
    const double lw_prop = exact ? lz[ch] : (-d2[ch] / two_sigma2) - logq;
    if (init) {
        lw_cur[ch] = lw_prop;
        loc[ch] = (uint8_t)buf;
        return;
    }
    const double lw = lw_cur[ch];
    const double ratio = (lw == -INFINITY) ? 1.0 : fmin(1.0, exp(lw_prop - lw));
    const Philox gen(seed);
    const double u = philox_uniform(gen, first_id + (uint64_t)ch, kAcceptDraw, tag);
    const bool acc = u < ratio;
    counts[ch * 2 + 1] += 1;
    if (acc) {
        counts[ch * 2] += 1;
        lw_cur[ch] = lw_prop;
        loc[ch] = (uint8_t)buf;
    } else if (loc[ch] == (uint8_t)buf) {
        loc[ch] = 0;
    }
}

// current_state <- the proposal buffer that holds it (imhk.py:169-171, deferred), for the chains whose location is
// in `which` (bit 0: buffer 1, bit 1: buffer 2); coalesced along chains.  Locations are reset by the caller
// (imhk_accept_kernel for the buffer about to be reused, a memset after a full flush).
__global__ void __launch_bounds__(256)
imhk_materialize_kernel(int64_t C, int npad, int64_t ld, const uint8_t* __restrict__ loc, int which,
                        const double* __restrict__ P1, const double* __restrict__ P2, double* __restrict__ Xcur) {
    const int64_t ch = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (ch >= C) return;
    const int l0 = loc[ch], l1 = (ch + 1 < C) ? loc[ch + 1] : 0;
    const bool a0 = l0 && ((which >> (l0 - 1)) & 1), a1 = l1 && ((which >> (l1 - 1)) & 1);
    if (!a0 && !a1) return;
    const double* __restrict__ s0 = (l0 == 1) ? P1 : P2;
    const double* __restrict__ s1 = (l1 == 1) ? P1 : P2;
    for (int i = blockIdx.y; i < npad; i += gridDim.y) {
        const size_t o = (size_t)i * ld + ch;
        if (a0 && a1 && s0 == s1) {
            *reinterpret_cast<double2*>(Xcur + o) = *reinterpret_cast<const double2*>(s0 + o);
        } else {
            if (a0) Xcur[o] = s0[o];
            if (a1) Xcur[o + 1] = s1[o + 1];
        }
    }
}

// X (coordinate-major doubles) -> int64 out[s * ldo + i]
__global__ void __launch_bounds__(256)
export_coeffs_strided_kernel(const double* __restrict__ X, int64_t ld, int n, int64_t S, int64_t* __restrict__ out,
                             int64_t ldo) {
    __shared__ double tile[32][33];
    const int64_t s0 = (int64_t)blockIdx.x * 32;
    const int i0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {
        int i = i0 + r;
        int64_t s = s0 + tx;
        tile[r][tx] = (i < n && s < S) ? X[(size_t)i * ld + s] : 0.0;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        int64_t s = s0 + r;
        int i = i0 + tx;
        if (s < S && i < n) out[(size_t)s * ldo + i] = (int64_t)tile[tx][r];
    }
}

}  // namespace

static cudaError_t ensure_blocked_chains(BlockedCtx& bc, ChainSet& cs) {
    const int64_t cap = (cs.chains + 127) / 128 * 128;
    if (cs.bl_cap >= cap && cs.Xcur) return cudaSuccess;
    cudaFree(cs.Xcur);
    cudaFree(cs.flags);
    cs.Xcur = nullptr;
    cs.flags = nullptr;
    cs.bl_cap = 0;
    cudaError_t e;
    if ((e = cudaMalloc(&cs.Xcur, sizeof(double) * (size_t)bc.npad * cap)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&cs.flags, cap)) != cudaSuccess) return e;
    if ((e = cudaMemset(cs.Xcur, 0, sizeof(double) * (size_t)bc.npad * cap)) != cudaSuccess) return e;
    cs.bl_cap = cap;
    return cudaSuccess;
}

// make Xcur hold every chain's current state (needed before it is read as a matrix: kept states, get_state)
static cudaError_t imhk_flush(BlockedCtx& bc, ChainSet& cs, cudaStream_t st) {
    const int64_t C = cs.chains;
    if (!bc.work[1].X || !bc.work[2].X) return cudaSuccess;
    ProfScope ps(bc.prof, LGS_PROF_OTHER, st);
    dim3 cgrid((unsigned)((C / 2 + 255) / 256 + 1), (unsigned)(bc.npad < 64 ? bc.npad : 64));
    imhk_materialize_kernel<<<cgrid, 256, 0, st>>>(C, bc.npad, cs.bl_cap, cs.flags, 3, bc.work[1].X, bc.work[2].X,
                                                   cs.Xcur);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    return cudaMemsetAsync(cs.flags, 0, (size_t)cs.bl_cap, st);
}

cudaError_t launch_imhk_blocked(BlockedCtx& bc, ChainSet& cs, int64_t burn_in, int64_t keep, int64_t thin,
                                int weight_mode, double* V, int64_t* X, cudaStream_t st) {
    if (bc.use_chain && bc.chain.ready) return launch_imhk_chain(bc, cs, burn_in, keep, thin, weight_mode, V, X, st);
    const int64_t C = cs.chains;
    cudaError_t e;
    if ((e = ensure_blocked_chains(bc, cs)) != cudaSuccess) return e;
    for (int b = 1; b <= 2; b++) {  // the two proposal buffers share the state matrix's leading dimension
        BlockedWork& w = bc.work[b];
        if (w.cap != cs.bl_cap) {
            cudaFree(w.X); cudaFree(w.lp); cudaFree(w.d2); cudaFree(w.lz);
            w = BlockedWork{};
            if ((e = ensure_work(bc, w, cs.bl_cap)) != cudaSuccess) return e;
        }
    }
    const int64_t ld = cs.bl_cap;
    const bool exact = weight_mode == LGS_WEIGHT_EXACT;
    const unsigned ablocks = (unsigned)((C + 255) / 256);
    dim3 cgrid((unsigned)((C / 2 + 255) / 256 + 1), (unsigned)(bc.npad < 64 ? bc.npad : 64));
    auto one_step = [&](bool init) -> cudaError_t {
        const uint32_t tag = init ? 0u : (uint32_t)(++cs.bl_steps);
        const int buf = 1 + (int)(cs.bl_steps & 1);  // proposal buffer of this step
        BlockedWork& w = bc.work[buf];
        {   // states still living in this buffer (accepted two steps ago, rejected since) move to Xcur first
            ProfScope ps(bc.prof, LGS_PROF_OTHER, st);
            imhk_materialize_kernel<<<cgrid, 256, 0, st>>>(C, bc.npad, ld, cs.flags, 1 << (buf - 1), bc.work[1].X,
                                                           bc.work[2].X, cs.Xcur);
        }
        cudaError_t er = klein_batch(bc, w, C, cs.seed, cs.first_id, tag, nullptr, exact, st);
        if (er != cudaSuccess) return er;
        ProfScope ps(bc.prof, LGS_PROF_OTHER, st);
        imhk_accept_kernel<<<ablocks, 256, 0, st>>>(C, w.lp, w.d2, w.lz, bc.sum_lognorm, bc.two_sigma2, exact ? 1 : 0,
                                                    init ? 1 : 0, cs.seed, cs.first_id, tag, cs.logw, cs.counts,
                                                    cs.flags, buf);
        return cudaGetLastError();
    };
    if (!cs.bl_init) {
        if ((e = cudaMemsetAsync(cs.flags, 0, (size_t)cs.bl_cap, st)) != cudaSuccess) return e;
        if ((e = one_step(true)) != cudaSuccess) return e;
        cs.bl_init = true;
    }
    const int64_t steps = burn_in + keep * thin;
    for (int64_t stp = 0; stp < steps; stp++) {
        if ((e = one_step(false)) != cudaSuccess) return e;
        if (stp >= burn_in && (stp - burn_in + 1) % thin == 0 && (V || X)) {
            const int64_t kidx = (stp - burn_in + 1) / thin - 1;
            if ((e = imhk_flush(bc, cs, st)) != cudaSuccess) return e;
            if (V) {
                if ((e = launch_contract<64, MODE_POINT>(bc.prof, bc.Bp, bc.npad, cs.Xcur, ld, 0, bc.mpad, 0, bc.npad, nullptr,
                                                         C, V + (size_t)kidx * bc.n, keep * (int64_t)bc.n, bc.n,
                                                         st)) != cudaSuccess) return e;
            }
            if (X) {
                dim3 grid((unsigned)((C + 31) / 32), (unsigned)((bc.n + 31) / 32));
                export_coeffs_strided_kernel<<<grid, 256, 0, st>>>(cs.Xcur, ld, bc.n, C, X + (size_t)kidx * bc.n,
                                                                   keep * (int64_t)bc.n);
                if ((e = cudaGetLastError()) != cudaSuccess) return e;
            }
        }
    }
    return cudaSuccess;
}

cudaError_t launch_imhk_blocked_get_state(BlockedCtx& bc, ChainSet& cs, int64_t* counts, int64_t* X, double* V,
                                          double* logw, cudaStream_t st) {
    if (bc.use_chain && bc.chain.ready) return launch_imhk_chain_get_state(bc, cs, counts, X, V, logw, st);
    const int64_t C = cs.chains;
    cudaError_t e;
    if (counts && (e = cudaMemcpyAsync(counts, cs.counts, sizeof(int64_t) * 2 * C, cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
    if (logw && (e = cudaMemcpyAsync(logw, cs.logw, sizeof(double) * C, cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
    if ((X || V) && (e = imhk_flush(bc, cs, st)) != cudaSuccess) return e;
    if (X) {
        dim3 grid((unsigned)((C + 31) / 32), (unsigned)((bc.n + 31) / 32));
        export_coeffs_strided_kernel<<<grid, 256, 0, st>>>(cs.Xcur, cs.bl_cap, bc.n, C, X, bc.n);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    if (V) {
        if ((e = launch_contract<64, MODE_POINT>(bc.prof, bc.Bp, bc.npad, cs.Xcur, cs.bl_cap, 0, bc.mpad, 0, bc.npad, nullptr, C,
                                                 V, bc.n, bc.n, st)) != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace lgs
