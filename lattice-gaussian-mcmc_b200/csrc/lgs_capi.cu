// C ABI of liblgs_b200.so (include/lgs_b200.h): handle management, per-sampler constants, dispatch.
#include <sched.h>

#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "lgs_internal.h"

using namespace lgs;

namespace {

thread_local std::string g_create_err;

int fail(lgs_handle* h, int code, const std::string& msg) {
    if (h) h->err = msg; else g_create_err = msg;
    return code;
}

int cuda_fail(lgs_handle* h, cudaError_t e, const char* where) {
    std::string msg = std::string(where) + ": " + cudaGetErrorString(e);
    int code = (e == cudaErrorMemoryAllocation) ? LGS_E_NOMEM : LGS_E_CUDA;
    return fail(h, code, msg);
}

#define LGS_CUDA(h, call)                                        \
    do {                                                         \
        cudaError_t e__ = (call);                                \
        if (e__ != cudaSuccess) return cuda_fail(h, e__, #call); \
    } while (0)

// NVTX range around a C-ABI call (tracing): visible to nsys / ncu --nvtx, free when no tool is
// attached (the header-only NVTX v3 stub resolves to no-ops)
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};
#define LGS_TRACE() NvtxRange nvtx_range__(__func__)

struct ArenaBuilder {
    std::vector<uint8_t> bytes;
    size_t add(const void* src, size_t nbytes) {
        size_t off = (bytes.size() + 255) & ~size_t(255);
        bytes.resize(off + nbytes);
        if (src) memcpy(bytes.data() + off, src, nbytes);
        return off;
    }
};

int ensure_stage(lgs_handle* h, size_t nbytes) {
    if (h->stage_cap >= nbytes) return LGS_OK;
    if (h->stage_dev) cudaFree(h->stage_dev);
    h->stage_dev = nullptr;
    h->stage_cap = 0;
    LGS_CUDA(h, cudaMalloc(&h->stage_dev, nbytes));
    h->stage_cap = nbytes;
    return LGS_OK;
}

void free_chains(ChainSet& cs) {
    cudaFree(cs.Xbuf); cudaFree(cs.Vbuf); cudaFree(cs.logw); cudaFree(cs.cur); cudaFree(cs.init);
    cudaFree(cs.counts); cudaFree(cs.step_no); cudaFree(cs.upos); cudaFree(cs.Xcur); cudaFree(cs.flags);
    cudaFree(cs.sp_src); cudaFree(cs.sp_keep); cudaFree(cs.Xkeep);
    cs = ChainSet{};
}

inline size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

// fp64 lattice points -> int16 for the trip over PCIe; *flag is set if any value does not survive the round trip
// (non-integer or outside int16).  Pure streaming: 8 doubles in, one 16-byte store out per thread and iteration.
__global__ void __launch_bounds__(256) pack_points_i16(const double* __restrict__ V, int16_t* __restrict__ P,
                                                       size_t count, int32_t* __restrict__ flag) {
    const size_t stride = (size_t)gridDim.x * blockDim.x * 8;
    bool bad = false;
    for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 8; i + 8 <= count; i += stride) {
        const double2* src = reinterpret_cast<const double2*>(V + i);
        const double2 a = src[0], b = src[1], c = src[2], d = src[3];
        const double v[8] = {a.x, a.y, b.x, b.y, c.x, c.y, d.x, d.y};
        unsigned w[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int lo = __double2int_rn(v[2 * k]), hi = __double2int_rn(v[2 * k + 1]);
            bad |= (double)lo != v[2 * k] || (double)hi != v[2 * k + 1] || lo < -32768 || lo > 32767 || hi < -32768 ||
                   hi > 32767;
            w[k] = ((unsigned)lo & 0xffffu) | ((unsigned)hi << 16);
        }
        *reinterpret_cast<uint4*>(P + i) = make_uint4(w[0], w[1], w[2], w[3]);
    }
    const size_t tail = count & ~(size_t)7;
    if (blockIdx.x == 0 && tail + threadIdx.x < count) {
        const double x = V[tail + threadIdx.x];
        const int r = __double2int_rn(x);
        bad |= (double)r != x || r < -32768 || r > 32767;
        P[tail + threadIdx.x] = (int16_t)r;
    }
    if (bad) atomicOr(flag, 1);
}

constexpr int kPackDev = 3;   // device slots: the GPU runs up to two chunks ahead of the host-side widening
constexpr int kPackHost = 3;  // pinned slots

}  // namespace

extern "C" {

int lgs_abi_version(void) { return LGS_ABI_VERSION; }

const char* lgs_last_error(const lgs_handle* h) { return h ? h->err.c_str() : g_create_err.c_str(); }

static int create_impl(const lgs_desc* d, lgs_handle** out);

int lgs_create(const lgs_desc* d, lgs_handle** out) {
    try {
        return create_impl(d, out);
    } catch (const std::bad_alloc&) {  // host-side setup arrays: nothing throws across the ABI
        return fail(nullptr, LGS_E_NOMEM, "lgs_create: host allocation failed");
    } catch (...) {
        return fail(nullptr, LGS_E_INVALID, "lgs_create: unexpected failure in host setup");
    }
}

static int create_impl(const lgs_desc* d, lgs_handle** out) {
    NvtxRange nvtx_range__("lgs_create");
    if (!d || !out) return fail(nullptr, LGS_E_INVALID, "lgs_create: NULL argument");
    *out = nullptr;
    if (d->n <= 0) return fail(nullptr, LGS_E_INVALID, "lgs_create: n must be positive");
    if (!(d->sigma > 0.0)) return fail(nullptr, LGS_E_INVALID, "lgs_create: Standard deviation must be positive");
    if (!d->R || !d->B || !d->t || !d->center) return fail(nullptr, LGS_E_INVALID, "lgs_create: NULL matrix");
    int max_range = d->max_range > 0 ? d->max_range : 1000;
    if (max_range > 2000) return fail(nullptr, LGS_E_INVALID, "lgs_create: max_range > 2000 unsupported");
    int precision = d->precision > 0 ? d->precision : 10;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, LGS_E_CUDA, std::string("lgs_create: no CUDA device (there is no CPU fallback): ") +
                                             cudaGetErrorString(e));
    if (d->device < 0 || d->device >= ndev) return fail(nullptr, LGS_E_INVALID, "lgs_create: bad device ordinal");
    e = cudaSetDevice(d->device);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaSetDevice");

    const int n = d->n;
    const double sigma = d->sigma;
    HostConsts hc;
    hc.rdiag.resize(n); hc.sig.resize(n); hc.lognorm.resize(n); hc.inv2s2.resize(n); hc.q.resize(n); hc.sq.resize(n);
    hc.znorm.resize(n); hc.eps1.resize(n); hc.mode.resize(n); hc.uidx.resize(n); hc.kind.resize(n); hc.dmax.resize(n);
    std::vector<double>&rdiag = hc.rdiag, &sig = hc.sig, &lognorm = hc.lognorm, &inv2s2 = hc.inv2s2, &q = hc.q,
                       &sq = hc.sq, &znorm = hc.znorm, &eps1 = hc.eps1;
    std::vector<uint8_t>& mode = hc.mode;
    std::vector<int>&uidx = hc.uidx, &kind = hc.kind, &dmax = hc.dmax;
    const double log_2pi = std::log(2.0 * M_PI);
    for (int i = 0; i < n; i++) {
        double rii = d->R[(size_t)i * n + i];
        if (rii == 0.0 || !std::isfinite(rii)) return fail(nullptr, LGS_E_INVALID, "lgs_create: singular R");
        rdiag[i] = rii;
        double s = sigma / std::fabs(rii);
        lognorm[i] = 0.5 * log_2pi + std::log(s);  // klein.py:262, unclamped sigma_i
        mode[i] = s < 1e-10 ? 1 : 0;               // klein.py:201
        if (s > 1e10) s = std::fmin(s, 1e6);       // klein.py:205-208
        sig[i] = s;
        inv2s2[i] = 1.0 / (2.0 * s * s);
        q[i] = std::exp(-1.0 / (s * s));
        sq[i] = std::exp(-inv2s2[i]);
        znorm[i] = s * std::sqrt(2.0 * M_PI);
        eps1[i] = 2.0 * std::exp(-2.0 * M_PI * M_PI * s * s);
        double rf = (s < 0.1) ? (double)(precision > 3 ? precision : 3) : (double)precision;
        double half = rf * s;
        bool truncated = 2.0 * half >= (double)(max_range - 1);
        if (truncated) {
            // wide truncated windows take closed-form masses instead of an O(window) walk (LGS_TRUNC_CLOSED_FORM=0:
            // the summed form, development A/B)
            const char* tc = getenv("LGS_TRUNC_CLOSED_FORM");
            kind[i] = (!(tc && tc[0] == '0') && s >= 50.0 && max_range >= 16) ? 3 : 2;
            dmax[i] = max_range / 2;
        } else {
            dmax[i] = (int)std::ceil(half) + 1;
            if (precision < 9 || s < 1.0) kind[i] = 2;
            else if (s < 1.45) kind[i] = 1;
            else kind[i] = 0;
        }
    }
    int ups = 0;
    for (int i = n - 1; i >= 0; i--) {
        uidx[i] = ups;
        if (!mode[i]) ups++;
    }

    ArenaBuilder ab;
    size_t oR = ab.add(d->R, sizeof(double) * (size_t)n * n);
    size_t oB = ab.add(d->B, sizeof(double) * (size_t)n * n);
    size_t ot = ab.add(d->t, sizeof(double) * n);
    size_t oc = ab.add(d->center, sizeof(double) * n);
    size_t ord = ab.add(rdiag.data(), sizeof(double) * n);
    size_t osg = ab.add(sig.data(), sizeof(double) * n);
    size_t oln = ab.add(lognorm.data(), sizeof(double) * n);
    size_t omo = ab.add(mode.data(), n);
    size_t oui = ab.add(uidx.data(), sizeof(int) * n);
    size_t o1 = ab.add(inv2s2.data(), sizeof(double) * n);
    size_t o2 = ab.add(q.data(), sizeof(double) * n);
    size_t o3 = ab.add(sq.data(), sizeof(double) * n);
    size_t o4 = ab.add(znorm.data(), sizeof(double) * n);
    size_t o5 = ab.add(eps1.data(), sizeof(double) * n);
    size_t o6 = ab.add(kind.data(), sizeof(int) * n);
    size_t o7 = ab.add(dmax.data(), sizeof(int) * n);

    lgs_handle* h = new lgs_handle();
    h->device = d->device;
    h->n = n;
    h->path = d->path;
    e = cudaMalloc(&h->arena, ab.bytes.size());
    if (e != cudaSuccess) { delete h; return cuda_fail(nullptr, e, "cudaMalloc(arena)"); }
    e = cudaMemcpy(h->arena, ab.bytes.data(), ab.bytes.size(), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { cudaFree(h->arena); delete h; return cuda_fail(nullptr, e, "cudaMemcpy(arena)"); }
    uint8_t* base = (uint8_t*)h->arena;
    DeviceCtx& c = h->ctx;
    c.n = n;
    c.npad = (n + 31) / 32 * 32;
    c.precision = precision;
    c.max_range = max_range;
    c.ups = ups;
    c.sigma = sigma;
    c.two_sigma2 = 2.0 * (sigma * sigma);
    c.inv_2sigma2 = 1.0 / c.two_sigma2;
    c.R = (const double*)(base + oR);
    c.B = (const double*)(base + oB);
    c.t = (const double*)(base + ot);
    c.center = (const double*)(base + oc);
    c.rdiag = (const double*)(base + ord);
    c.sig = (const double*)(base + osg);
    c.lognorm = (const double*)(base + oln);
    c.mode = (const uint8_t*)(base + omo);
    c.uidx = (const int*)(base + oui);
    c.fz_inv2s2 = (const double*)(base + o1);
    c.fz_q = (const double*)(base + o2);
    c.fz_sq = (const double*)(base + o3);
    c.fz_znorm = (const double*)(base + o4);
    c.fz_eps1 = (const double*)(base + o5);
    c.fz_kind = (const int*)(base + o6);
    c.fz_dmax = (const int*)(base + o7);
    if (d->path != LGS_PATH_ORDERED) {
        e = blocked_init(h->blocked, *d, hc);
        if (e != cudaSuccess) {
            blocked_free(h->blocked);
            cudaFree(h->arena);
            delete h;
            return cuda_fail(nullptr, e, "blocked_init");
        }
    }
    h->pack_ok = true;
    for (size_t i = 0; i < (size_t)n * n && h->pack_ok; i++)
        h->pack_ok = d->B[i] == std::nearbyint(d->B[i]) && std::fabs(d->B[i]) < 9007199254740992.0;
    if (const char* e = getenv("LGS_HOST_PACK")) h->pack_ok = h->pack_ok && atoi(e) != 0;
    h->pack_ok = h->pack_ok && host_simd_ok();   // the widening loops are AVX2: without it, the fp64 route
    *out = h;
    return LGS_OK;
}

void lgs_destroy(lgs_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    free_chains(h->chains);
    blocked_free(h->blocked);
    cudaFree(h->scratch.Xw);
    cudaFree(h->stage_dev);
    cudaFree(h->hist_dims);
    cudaFree(h->diag_scratch);
    cudaFree(h->arena);
    cudaFree(h->pack_dev);
    if (h->pack_host) cudaFreeHost(h->pack_host);
    for (int i = 0; i < 3; i++) {
        if (h->pk_done[i]) cudaEventDestroy(h->pk_done[i]);
        if (h->pk_copied[i]) cudaEventDestroy(h->pk_copied[i]);
    }
    hostpool_destroy(h->pool);
    for (int i = 0; i < 2; i++) {
        if (h->bounce[i]) cudaFreeHost(h->bounce[i]);
        if (h->bounce_ev[i]) cudaEventDestroy(h->bounce_ev[i]);
    }
    for (int i = 0; i < 2; i++) {
        if (h->zc_v16[i]) cudaFreeHost(h->zc_v16[i]);
        if (h->zc_done[i]) cudaFreeHost(h->zc_done[i]);
        if (h->zc_ev[i]) cudaEventDestroy(h->zc_ev[i]);
    }
    for (int i = 0; i < 2; i++) {
        cudaFree(h->zs_dev[i]);
        if (h->zs_reset[i]) cudaEventDestroy(h->zs_reset[i]);
        if (h->zs_copied[i]) cudaEventDestroy(h->zs_copied[i]);
    }
    cudaFree(h->zs_gcnt);
    cudaFree(h->zc_cnt);
    cudaFree(h->zc_ovf);
    if (h->host_compute) cudaStreamDestroy(h->host_compute);
    if (h->host_copy) cudaStreamDestroy(h->host_copy);
    for (int i = 0; i < 2; i++) {
        if (h->ev_done[i]) cudaEventDestroy(h->ev_done[i]);
        if (h->ev_free[i]) cudaEventDestroy(h->ev_free[i]);
    }
    delete h;
}

int lgs_uniforms_per_sample(const lgs_handle* h) { return h ? h->ctx.ups : LGS_E_INVALID; }

int64_t lgs_preferred_batch(const lgs_handle* h) {
    if (!h) return LGS_E_INVALID;
    if (h->blocked.ready && h->blocked.use_chain && h->blocked.chain.ready) return h->blocked.chain.batch;
    if (h->blocked.ready) return h->blocked.wave * 4;
    return 65536;
}

int lgs_klein_sample(lgs_handle* h, int64_t S, uint64_t seed, uint64_t first_sample_id, int samplez, double* V_dev,
                     int64_t* X_dev, double* logq_dev, void* stream) {
    LGS_TRACE();
    if (!h) return LGS_E_INVALID;
    if (S < 0) return fail(h, LGS_E_INVALID, "lgs_klein_sample: S < 0");
    if (samplez != LGS_SAMPLEZ_TABLE && samplez != LGS_SAMPLEZ_FAST)
        return fail(h, LGS_E_INVALID, "lgs_klein_sample: unknown samplez");
    LGS_CUDA(h, cudaSetDevice(h->device));
    if (samplez == LGS_SAMPLEZ_FAST && h->path != LGS_PATH_ORDERED && h->blocked.ready) {
        LGS_CUDA(h, launch_klein_blocked(h->blocked, S, seed, first_sample_id, V_dev, X_dev, logq_dev,
                                         (cudaStream_t)stream));
        return LGS_OK;
    }
    if (h->path == LGS_PATH_BLOCKED)
        return fail(h, LGS_E_INVALID, "lgs_klein_sample: the blocked path needs LGS_SAMPLEZ_FAST");
    RngSpec rng{nullptr, 0, seed, first_sample_id};
    LGS_CUDA(h, launch_klein_ordered(h->ctx, h->scratch, S, rng, samplez, V_dev, X_dev, logq_dev, nullptr,
                                     (cudaStream_t)stream));
    return LGS_OK;
}

// replay calls of a handle created with LGS_PATH_BLOCKED run on the chain kernel (when it serves this lattice)
static bool replay_on_chain(const lgs_handle* h) {
    return h->path == LGS_PATH_BLOCKED && h->blocked.ready && h->blocked.use_chain && h->blocked.chain.ready;
}

int lgs_replay_path(const lgs_handle* h) {
    if (!h) return LGS_E_INVALID;
    return replay_on_chain(h) ? LGS_PATH_BLOCKED : LGS_PATH_ORDERED;
}

// one replay launch sequence; the fallback counter is accumulated into, not reset
static int klein_replay_impl(lgs_handle* h, int64_t S, const double* U_dev, double* V_dev, int64_t* X_dev,
                             double* logq_dev, int32_t* fallback_count_dev, cudaStream_t stream) {
    if (replay_on_chain(h)) {
        // LGS_PATH_BLOCKED: the recorded stream goes through the throughput kernel itself (klein_chain_kernel, REPLAY
        // instantiation: DMMA-contracted conditional means + the reference's table SampleZ)
        const ChainReplay rp{U_dev, h->ctx.ups, fallback_count_dev, nullptr, nullptr};
        LGS_CUDA(h, launch_klein_chain(h->blocked, S, 0, 0, V_dev, X_dev, logq_dev, stream, &rp));
        return LGS_OK;
    }
    RngSpec rng{U_dev, h->ctx.ups, 0, 0};
    LGS_CUDA(h, launch_klein_ordered(h->ctx, h->scratch, S, rng, LGS_SAMPLEZ_TABLE, V_dev, X_dev, logq_dev,
                                     fallback_count_dev, stream));
    return LGS_OK;
}

int lgs_klein_replay(lgs_handle* h, int64_t S, const double* U_dev, double* V_dev, int64_t* X_dev, double* logq_dev,
                     int32_t* fallback_count_dev, void* stream) {
    LGS_TRACE();
    if (!h) return LGS_E_INVALID;
    if (S < 0 || (S > 0 && !U_dev)) return fail(h, LGS_E_INVALID, "lgs_klein_replay: bad arguments");
    LGS_CUDA(h, cudaSetDevice(h->device));
    if (fallback_count_dev) LGS_CUDA(h, cudaMemsetAsync(fallback_count_dev, 0, sizeof(int32_t), (cudaStream_t)stream));
    return klein_replay_impl(h, S, U_dev, V_dev, X_dev, logq_dev, fallback_count_dev, (cudaStream_t)stream);
}

// Host-buffer Klein call for an integer basis, lattice points only: chunk k is sampled and packed to int16 on the
// compute stream, crosses PCIe on the copy stream (a quarter of the fp64 bytes) into a pinned ring, and is widened
// into the caller's float64 buffer by the host pool while the GPU is already two chunks ahead.  A chunk whose pack
// kernel raised its flag (a value outside int16) is fetched again as fp64 from its device slot, which is only
// reused after the flag has been read.  Output is bit-identical to the plain fp64 copy.
static int klein_host_packed(lgs_handle* h, int64_t S, uint64_t seed, uint64_t first_id, int samplez, double* V_host,
                             int64_t chunk) {
    const size_t n = h->n;
    const size_t bV = align256(sizeof(double) * chunk * n);
    const size_t bP = align256(sizeof(int16_t) * chunk * n) + 256;  // + flag
    if (h->pack_dev_cap < kPackDev * (bV + bP)) {
        cudaFree(h->pack_dev);
        h->pack_dev = nullptr, h->pack_dev_cap = 0;
        LGS_CUDA(h, cudaMalloc(&h->pack_dev, kPackDev * (bV + bP)));
        h->pack_dev_cap = kPackDev * (bV + bP);
    }
    if (h->pack_host_cap < kPackHost * bP) {
        if (h->pack_host) cudaFreeHost(h->pack_host);
        h->pack_host = nullptr, h->pack_host_cap = 0;
        LGS_CUDA(h, cudaHostAlloc(&h->pack_host, kPackHost * bP, cudaHostAllocDefault));
        h->pack_host_cap = kPackHost * bP;
    }
    if (!h->pool) h->pool = hostpool_create(0);
    if (!h->pk_done[0])
        for (int i = 0; i < 3; i++) {
            LGS_CUDA(h, cudaEventCreateWithFlags(&h->pk_done[i], cudaEventDisableTiming));
            LGS_CUDA(h, cudaEventCreateWithFlags(&h->pk_copied[i], cudaEventDisableTiming));
        }
    cudaStream_t cs = h->host_compute, ps = h->host_copy;
    const int64_t K = (S + chunk - 1) / chunk;
    const int lag = 2;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
    for (int64_t k = 0; k < K + lag; k++) {
        if (k < K) {
            const int64_t s0 = k * chunk, cnt = (S - s0) < chunk ? (S - s0) : chunk;
            const int b = (int)(k % kPackDev);
            uint8_t* slot = (uint8_t*)h->pack_dev + (size_t)b * (bV + bP);
            double* dV = (double*)slot;
            int16_t* dP = (int16_t*)(slot + bV);
            int32_t* dflag = (int32_t*)(slot + bV + bP - 256);
            // slot b was last used by chunk k-3: its packed copy completed before the host widened it (iteration k-1)
            int rc = lgs_klein_sample(h, cnt, seed, first_id + (uint64_t)s0, samplez, dV, nullptr, nullptr, cs);
            if (rc) return rc;
            LGS_CUDA(h, cudaMemsetAsync(dflag, 0, sizeof(int32_t), cs));
            pack_points_i16<<<sms * 8, 256, 0, cs>>>(dV, dP, (size_t)cnt * n, dflag);
            LGS_CUDA(h, cudaGetLastError());
            LGS_CUDA(h, cudaEventRecord(h->pk_done[b], cs));
            LGS_CUDA(h, cudaStreamWaitEvent(ps, h->pk_done[b], 0));
            uint8_t* hp = (uint8_t*)h->pack_host + (size_t)(k % kPackHost) * bP;
            const size_t payload = align256(sizeof(int16_t) * cnt * n);
            LGS_CUDA(h, cudaMemcpyAsync(hp, dP, sizeof(int16_t) * cnt * n, cudaMemcpyDeviceToHost, ps));
            LGS_CUDA(h, cudaMemcpyAsync(hp + bP - 256, dflag, sizeof(int32_t), cudaMemcpyDeviceToHost, ps));
            (void)payload;
            LGS_CUDA(h, cudaEventRecord(h->pk_copied[k % kPackHost], ps));
        }
        const int64_t j = k - lag;
        if (j >= 0) {
            const int64_t s0 = j * chunk, cnt = (S - s0) < chunk ? (S - s0) : chunk;
            LGS_CUDA(h, cudaEventSynchronize(h->pk_copied[j % kPackHost]));
            const uint8_t* hp = (const uint8_t*)h->pack_host + (size_t)(j % kPackHost) * bP;
            if (*(const int32_t*)(hp + bP - 256)) {  // out of int16 range: take the fp64 chunk itself
                const double* dV = (const double*)((uint8_t*)h->pack_dev + (size_t)(j % kPackDev) * (bV + bP));
                LGS_CUDA(h, cudaMemcpyAsync(V_host + (size_t)s0 * n, dV, sizeof(double) * cnt * n, cudaMemcpyDeviceToHost, ps));
                LGS_CUDA(h, cudaStreamSynchronize(ps));
            } else {
                hostpool_expand_i16(h->pool, (const int16_t*)hp, V_host + (size_t)s0 * n, (size_t)cnt * n);
            }
        }
    }
    LGS_CUDA(h, cudaStreamSynchronize(cs));
    return LGS_OK;
}

static int klein_host_common(lgs_handle* h, int64_t S, const double* U_host, uint64_t seed, uint64_t first_id,
                             int samplez, double* V_host, int64_t* X_host, double* logq_host, int32_t* fb_host,
                             bool allow_pack);

static int ensure_host_streams(lgs_handle* h) {
    if (h->host_compute) return LGS_OK;
    LGS_CUDA(h, cudaStreamCreateWithFlags(&h->host_compute, cudaStreamNonBlocking));
    LGS_CUDA(h, cudaStreamCreateWithFlags(&h->host_copy, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) {
        LGS_CUDA(h, cudaEventCreateWithFlags(&h->ev_done[i], cudaEventDisableTiming));
        LGS_CUDA(h, cudaEventCreateWithFlags(&h->ev_free[i], cudaEventDisableTiming));
    }
    return LGS_OK;
}

// Device -> host copy on stream ps for results whose destination may be PAGEABLE memory (a fresh numpy array): the
// driver's own pageable path is a single-threaded bounce; here the bytes land in a two-deep pinned ring at full PCIe
// rate and the host pool moves each piece on with non-temporal stores (faulting the destination's pages in parallel)
// while the next piece is in flight.  Pinned / registered destinations and small copies are a plain cudaMemcpyAsync.
// Returns with the copy either complete (staged) or enqueued on ps (direct): callers synchronise ps before returning.
static const size_t kBouncePiece = [] {
    const char* e = std::getenv("LGS_BOUNCE_MB");   // development override of the piece size
    const int mb = e ? std::atoi(e) : 0;
    return (size_t)(mb > 0 && mb <= 1024 ? mb : 32) << 20;
}();
static constexpr size_t kBounceMin = (size_t)8 << 20;

static bool host_memory_is_pinned(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

static bool bounce_enabled() {
    static const bool on = [] {
        const char* e = std::getenv("LGS_HOST_BOUNCE");
        return !(e && e[0] == '0') && hostpool_default_threads() >= 2 && host_simd_ok();
    }();
    return on;
}

static int staged_d2h(lgs_handle* h, void* dst, const void* src_dev, size_t bytes, cudaStream_t ps) {
    if (!bytes) return LGS_OK;
    if (bytes < kBounceMin || !bounce_enabled() || host_memory_is_pinned(dst)) {
        LGS_CUDA(h, cudaMemcpyAsync(dst, src_dev, bytes, cudaMemcpyDeviceToHost, ps));
        return LGS_OK;
    }
    for (int i = 0; i < 2; i++) {
        if (!h->bounce[i]) LGS_CUDA(h, cudaHostAlloc(&h->bounce[i], kBouncePiece, cudaHostAllocDefault));
        if (!h->bounce_ev[i]) LGS_CUDA(h, cudaEventCreateWithFlags(&h->bounce_ev[i], cudaEventDisableTiming));
    }
    if (!h->pool) h->pool = hostpool_create(0);
    const size_t pieces = (bytes + kBouncePiece - 1) / kBouncePiece;
    for (size_t i = 0; i <= pieces; i++) {
        if (i < pieces) {
            const size_t off = i * kBouncePiece, len = bytes - off < kBouncePiece ? bytes - off : kBouncePiece;
            LGS_CUDA(h, cudaMemcpyAsync(h->bounce[i & 1], (const uint8_t*)src_dev + off, len, cudaMemcpyDeviceToHost, ps));
            LGS_CUDA(h, cudaEventRecord(h->bounce_ev[i & 1], ps));
        }
        if (i >= 1) {  // piece i-1 has landed: move it on while piece i crosses PCIe
            const size_t j = i - 1, off = j * kBouncePiece;
            const size_t len = bytes - off < kBouncePiece ? bytes - off : kBouncePiece;
            LGS_CUDA(h, cudaEventSynchronize(h->bounce_ev[j & 1]));
            hostpool_copy(h->pool, h->bounce[j & 1], (uint8_t*)dst + off, len);
        }
    }
    return LGS_OK;
}

// Lattice points that already sit in device memory as fp64 [rows][n] (complete: the producing work has been
// synchronised) -> the caller's float64 host array, over the int16 transport: pack kernel and copy in chunks on two
// streams, widened by the host pool two chunks behind; a chunk with a value outside int16 is copied as fp64.
static int ship_points_packed(lgs_handle* h, const double* dV, int64_t rows, double* V_host) {
    const size_t n = h->n;
    int rc = ensure_host_streams(h);
    if (rc) return rc;
    int64_t chunk = (int64_t)((size_t)48 << 20) / (int64_t)n;   // ~48 M values = 96 MB of int16 per chunk
    if (chunk < 1024) chunk = 1024;
    if (chunk > rows) chunk = rows;
    const size_t bP = align256(sizeof(int16_t) * chunk * n) + 256;  // + flag
    if (h->pack_dev_cap < kPackDev * bP) {
        cudaFree(h->pack_dev);
        h->pack_dev = nullptr, h->pack_dev_cap = 0;
        LGS_CUDA(h, cudaMalloc(&h->pack_dev, kPackDev * bP));
        h->pack_dev_cap = kPackDev * bP;
    }
    if (h->pack_host_cap < kPackHost * bP) {
        if (h->pack_host) cudaFreeHost(h->pack_host);
        h->pack_host = nullptr, h->pack_host_cap = 0;
        LGS_CUDA(h, cudaHostAlloc(&h->pack_host, kPackHost * bP, cudaHostAllocDefault));
        h->pack_host_cap = kPackHost * bP;
    }
    if (!h->pool) h->pool = hostpool_create(0);
    if (!h->pk_done[0])
        for (int i = 0; i < 3; i++) {
            LGS_CUDA(h, cudaEventCreateWithFlags(&h->pk_done[i], cudaEventDisableTiming));
            LGS_CUDA(h, cudaEventCreateWithFlags(&h->pk_copied[i], cudaEventDisableTiming));
        }
    cudaStream_t cs = h->host_compute, ps = h->host_copy;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
    const int64_t K = (rows + chunk - 1) / chunk;
    const int lag = 2;
    for (int64_t k = 0; k < K + lag; k++) {
        if (k < K) {
            const int64_t r0 = k * chunk, cnt = (rows - r0) < chunk ? (rows - r0) : chunk;
            const int b = (int)(k % kPackDev);
            uint8_t* slot = (uint8_t*)h->pack_dev + (size_t)b * bP;   // last used by chunk k-3, widened at iteration k-1
            int16_t* dP = (int16_t*)slot;
            int32_t* dflag = (int32_t*)(slot + bP - 256);
            LGS_CUDA(h, cudaMemsetAsync(dflag, 0, sizeof(int32_t), cs));
            pack_points_i16<<<sms * 8, 256, 0, cs>>>(dV + (size_t)r0 * n, dP, (size_t)cnt * n, dflag);
            LGS_CUDA(h, cudaGetLastError());
            LGS_CUDA(h, cudaEventRecord(h->pk_done[b], cs));
            LGS_CUDA(h, cudaStreamWaitEvent(ps, h->pk_done[b], 0));
            uint8_t* hp = (uint8_t*)h->pack_host + (size_t)(k % kPackHost) * bP;
            LGS_CUDA(h, cudaMemcpyAsync(hp, dP, sizeof(int16_t) * cnt * n, cudaMemcpyDeviceToHost, ps));
            LGS_CUDA(h, cudaMemcpyAsync(hp + bP - 256, dflag, sizeof(int32_t), cudaMemcpyDeviceToHost, ps));
            LGS_CUDA(h, cudaEventRecord(h->pk_copied[k % kPackHost], ps));
        }
        const int64_t j = k - lag;
        if (j >= 0) {
            const int64_t r0 = j * chunk, cnt = (rows - r0) < chunk ? (rows - r0) : chunk;
            LGS_CUDA(h, cudaEventSynchronize(h->pk_copied[j % kPackHost]));
            const uint8_t* hp = (const uint8_t*)h->pack_host + (size_t)(j % kPackHost) * bP;
            if (*(const int32_t*)(hp + bP - 256)) {
                LGS_CUDA(h, cudaMemcpyAsync(V_host + (size_t)r0 * n, dV + (size_t)r0 * n, sizeof(double) * cnt * n,
                                            cudaMemcpyDeviceToHost, ps));
                LGS_CUDA(h, cudaStreamSynchronize(ps));
            } else {
                hostpool_expand_i16(h->pool, (const int16_t*)hp, V_host + (size_t)r0 * n, (size_t)cnt * n);
            }
        }
    }
    LGS_CUDA(h, cudaStreamSynchronize(cs));
    LGS_CUDA(h, cudaStreamSynchronize(ps));
    return LGS_OK;
}

// Host-buffer Klein call on the chain kernel for an integer basis with n % 8 == 0, lattice points only.  ONE launch per
// full batch: the kernel's lattice-point epilogue writes int16 straight into mapped pinned host memory and publishes a
// per-tile completion word; the host pool picks finished tiles up in order and widens them into the caller's float64
// buffer while the kernel is still running, so what remains after the kernel ends is the last tile of every SM.  The
// next batch is launched before the current one is drained.  A batch in which a value left int16 is re-run through
// the plain fp64 route (same Philox keys, same samples).
// cuStreamWaitValue32 / cuStreamWriteValue32 through the runtime's driver entry point query (no link-time libcuda)
typedef int (*StreamValue32Fn)(cudaStream_t, unsigned long long, unsigned, unsigned);
struct StreamMemOps {
    StreamValue32Fn wait = nullptr, write = nullptr;
    bool ok = false;
};
static const StreamMemOps& stream_mem_ops() {
    static StreamMemOps ops = [] {
        StreamMemOps o;
        void *w = nullptr, *x = nullptr;
        cudaDriverEntryPointQueryResult q1, q2;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &w, cudaEnableDefault, &q1) == cudaSuccess &&
            cudaGetDriverEntryPoint("cuStreamWriteValue32", &x, cudaEnableDefault, &q2) == cudaSuccess &&
            q1 == cudaDriverEntryPointSuccess && q2 == cudaDriverEntryPointSuccess && w && x) {
            o.wait = (StreamValue32Fn)w;
            o.write = (StreamValue32Fn)x;
            o.ok = true;
        }
        return o;
    }();
    return ops;
}

static int klein_host_zerocopy(lgs_handle* h, int64_t S, uint64_t seed, uint64_t first_id, double* V_host) {
    const size_t n = h->n;
    const int64_t batch = h->blocked.chain.batch;
    const int64_t cap = S < batch ? (S + 127) / 128 * 128 : batch;
    const int tiles_cap = (int)(cap / 128);
    // Two hand-over mechanisms (both measured, tools/quick_e2e.py):
    //   mapped    the kernel stores int16 into mapped pinned memory, one completion word per tile -- the finest-grained
    //             hand-over and the shortest tail, but stores into host memory throttle the DMMA warps (~13 %);
    //   streamed  the kernel keeps its int16 output in HBM; the copy engine ships finished groups of tiles behind
    //             cuStreamWaitValue32 -- the kernel runs at full speed and batch k's transfer overlaps batch k+1.
    // A call of one batch is bound by first-tile latency + host widening, where mapped wins (20.3 vs 21.3 ms at
    // n = 1024); from two batches on streamed wins (18.4 vs 20.1 ms per batch).  LGS_HOST_ZEROCOPY=2 / 3 force one.
    const char* zc_env = getenv("LGS_HOST_ZEROCOPY");
    const int zc_mode = zc_env ? atoi(zc_env) : 1;
    const bool streamed = stream_mem_ops().ok && zc_mode != 2 && (zc_mode == 3 || S > cap);
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
    int G = sms >= 4 ? sms / 4 : 1;   // tiles per group: a quarter of one round of tiles over the SMs
    if (const char* ge = getenv("LGS_ZC_GROUP")) G = atoi(ge) > 0 ? atoi(ge) : G;   // development
    if (h->zc_cap < (size_t)cap) {
        for (int i = 0; i < 2; i++) {
            if (h->zc_v16[i]) cudaFreeHost(h->zc_v16[i]);
            if (h->zc_done[i]) cudaFreeHost(h->zc_done[i]);
            cudaFree(h->zs_dev[i]);
            h->zc_v16[i] = nullptr, h->zc_done[i] = nullptr, h->zs_dev[i] = nullptr;
        }
        cudaFree(h->zc_cnt);
        cudaFree(h->zs_gcnt);
        h->zc_cnt = nullptr, h->zs_gcnt = nullptr, h->zc_cap = 0;
        for (int i = 0; i < 2; i++) {
            LGS_CUDA(h, cudaHostAlloc(&h->zc_v16[i], sizeof(int16_t) * cap * n, cudaHostAllocMapped));
            LGS_CUDA(h, cudaHostAlloc((void**)&h->zc_done[i], sizeof(unsigned) * tiles_cap, cudaHostAllocMapped));
            memset(h->zc_done[i], 0, sizeof(unsigned) * tiles_cap);
            LGS_CUDA(h, cudaMalloc(&h->zs_dev[i], sizeof(int16_t) * cap * n));
        }
        LGS_CUDA(h, cudaMalloc((void**)&h->zc_cnt, sizeof(unsigned) * 2 * tiles_cap));
        LGS_CUDA(h, cudaMalloc((void**)&h->zs_gcnt, sizeof(unsigned) * 2 * tiles_cap));
        if (!h->zc_ovf) LGS_CUDA(h, cudaMalloc((void**)&h->zc_ovf, sizeof(int) * 2));
        h->zc_cap = (size_t)cap;
        h->zc_epoch = 0;
    }
    if (!h->zc_ev[0])
        for (int i = 0; i < 2; i++) {
            LGS_CUDA(h, cudaEventCreateWithFlags(&h->zc_ev[i], cudaEventDisableTiming));
            LGS_CUDA(h, cudaEventCreateWithFlags(&h->zs_reset[i], cudaEventDisableTiming));
            LGS_CUDA(h, cudaEventCreateWithFlags(&h->zs_copied[i], cudaEventDisableTiming));
        }
    if (!h->pool) h->pool = hostpool_create(0);
    cudaStream_t cs = h->host_compute, ps = h->host_copy;
    const int64_t K = (S + cap - 1) / cap;
    const int tiles_stride = (int)(h->zc_cap / 128);
    unsigned epochs[2] = {0, 0};
    auto launch = [&](int64_t k) -> int {
        const int b = (int)(k & 1);
        const int64_t s0 = k * cap, cnt = (S - s0) < cap ? (S - s0) : cap;
        const int ntiles = (int)((cnt + 127) / 128);
        ChainZeroCopy zc{};
        void* dv = nullptr;
        void* dd = nullptr;
        LGS_CUDA(h, cudaHostGetDevicePointer(&dv, h->zc_v16[b], 0));
        LGS_CUDA(h, cudaHostGetDevicePointer(&dd, h->zc_done[b], 0));
        zc.v16 = streamed ? (int16_t*)h->zs_dev[b] : (int16_t*)dv;
        zc.tile_done = (unsigned*)dd;
        zc.tile_cnt = h->zc_cnt + (size_t)b * tiles_stride;
        zc.ovf = h->zc_ovf + b;
        zc.epoch = epochs[b] = ++h->zc_epoch;
        zc.group_cnt = streamed ? h->zs_gcnt + (size_t)b * tiles_stride : nullptr;
        zc.group_tiles = G;
        LGS_CUDA(h, cudaMemsetAsync(zc.ovf, 0, sizeof(int), cs));
        if (streamed) {
            LGS_CUDA(h, cudaMemsetAsync(zc.group_cnt, 0, sizeof(unsigned) * ((ntiles + G - 1) / G), cs));
            LGS_CUDA(h, cudaEventRecord(h->zs_reset[b], cs));
        }
        LGS_CUDA(h, launch_klein_chain_zc(h->blocked, cnt, seed, first_id + (uint64_t)s0, zc, cs));
        LGS_CUDA(h, cudaEventRecord(h->zc_ev[b], cs));
        if (streamed) {
            const StreamMemOps& ops = stream_mem_ops();
            LGS_CUDA(h, cudaStreamWaitEvent(ps, h->zs_reset[b], 0));   // the counters have been cleared
            for (int g0 = 0, g = 0; g0 < ntiles; g0 += G, g++) {
                const int gt = ntiles - g0 < G ? ntiles - g0 : G;
                const int64_t r0 = (int64_t)g0 * 128, r1 = (int64_t)(g0 + gt) * 128 < cnt ? (int64_t)(g0 + gt) * 128 : cnt;
                if (ops.wait(ps, (unsigned long long)(uintptr_t)(zc.group_cnt + g), (unsigned)gt, 0x1 /* GEQ */))
                    return fail(h, LGS_E_CUDA, "cuStreamWaitValue32 failed");
                LGS_CUDA(h, cudaMemcpyAsync((int16_t*)h->zc_v16[b] + (size_t)r0 * n, (const int16_t*)h->zs_dev[b] + (size_t)r0 * n,
                                            sizeof(int16_t) * (size_t)(r1 - r0) * n, cudaMemcpyDeviceToHost, ps));
                if (ops.write(ps, (unsigned long long)(uintptr_t)((unsigned*)dd + g), zc.epoch, 0))
                    return fail(h, LGS_E_CUDA, "cuStreamWriteValue32 failed");
            }
            LGS_CUDA(h, cudaEventRecord(h->zs_copied[b], ps));
        }
        return LGS_OK;
    };
    int rc = launch(0);
    if (rc) return rc;
    for (int64_t k = 0; k < K; k++) {
        if (k + 1 < K && (rc = launch(k + 1))) return rc;  // its buffers were drained in iteration k - 1
        const int b = (int)(k & 1);
        const int64_t s0 = k * cap, cnt = (S - s0) < cap ? (S - s0) : cap;
        const int ntiles = (int)((cnt + 127) / 128);
        const int16_t* src = (const int16_t*)h->zc_v16[b];
        volatile unsigned* done = h->zc_done[b];
        const unsigned epoch = epochs[b];
        cudaEvent_t ev = streamed ? h->zs_copied[b] : h->zc_ev[b];
        const int fdiv = streamed ? G : 1;   // one completion word per group / per tile
        std::atomic<int> next{0};
        std::atomic<int> failed{0};
        const int device = h->device;
        hostpool_run(h->pool, [&, src, done, epoch, ev, ntiles, cnt, s0, device, fdiv](int, int) {
            cudaSetDevice(device);
            for (;;) {
                const int t = next.fetch_add(1);
                if (t >= ntiles || failed.load()) return;
                const int f = t / fdiv;
                unsigned spins = 0;
                while (done[f] != epoch) {
                    __builtin_ia32_pause();
                    if ((++spins & 0xffu) == 0) sched_yield();   // be polite when the host is oversubscribed
                    if ((spins & 0xfffu) == 0 && cudaEventQuery(ev) != cudaErrorNotReady) {
                        if (done[f] == epoch) break;  // everything has been issued: the word is out, or the run failed
                        failed.store(1);
                        return;
                    }
                }
                std::atomic_thread_fence(std::memory_order_acquire);
                const int64_t r0 = (int64_t)t * 128, r1 = r0 + 128 < cnt ? r0 + 128 : cnt;
                expand_i16_range(src + (size_t)r0 * n, V_host + (size_t)(s0 + r0) * n, (size_t)(r1 - r0) * n);
            }
        });
        LGS_CUDA(h, cudaEventSynchronize(h->zc_ev[b]));
        if (streamed) LGS_CUDA(h, cudaEventSynchronize(h->zs_copied[b]));
        if (failed.load()) return fail(h, LGS_E_CUDA, "lgs_klein_sample_host: the chain kernel ended without publishing every tile");
        int ovf = 0;
        LGS_CUDA(h, cudaMemcpy(&ovf, h->zc_ovf + b, sizeof(int), cudaMemcpyDeviceToHost));
        if (ovf) {
            // A value outside int16: this batch again, as fp64.  host_compute / host_copy are cudaStreamNonBlocking, so the
            // legacy-stream cudaMemcpy above has NOT drained them: batch k + 1's kernel and its gated copies may still be
            // in flight.  The re-run is ordered after them because it is enqueued on the same two streams (and shares
            // chain.work[0] with batch k + 1 in stream order); its output buffers are this batch's alone.
            rc = klein_host_common(h, cnt, nullptr, seed, first_id + (uint64_t)s0, LGS_SAMPLEZ_FAST, V_host + (size_t)s0 * n,
                                   nullptr, nullptr, nullptr, false);
            if (rc) return rc;
        }
    }
    LGS_CUDA(h, cudaStreamSynchronize(cs));
    LGS_CUDA(h, cudaStreamSynchronize(ps));
    return LGS_OK;
}

// Host-buffer Klein call.  The batch is cut into chunks; chunk k is computed on the compute stream into one of two
// staging slots while chunk k-1 travels to the host on the copy stream, so the PCIe transfer (the larger cost:
// 8n bytes per sample) hides the sampling.  Works with pageable host memory too (the copy then blocks the host, not
// the device).
static int klein_host_common_impl(lgs_handle* h, int64_t S, const double* U_host, uint64_t seed, uint64_t first_id,
                                  int samplez, double* V_host, int64_t* X_host, double* logq_host, int32_t* fb_host,
                                  bool allow_pack);

// Every early error return below can leave work in flight on the handle's two streams that reads or writes
// handle-owned buffers: drain them before the error reaches the caller, so that a later call or lgs_destroy cannot
// free or reuse a buffer under a running kernel.
static int klein_host_common(lgs_handle* h, int64_t S, const double* U_host, uint64_t seed, uint64_t first_id,
                             int samplez, double* V_host, int64_t* X_host, double* logq_host, int32_t* fb_host,
                             bool allow_pack) {
    const int rc = klein_host_common_impl(h, S, U_host, seed, first_id, samplez, V_host, X_host, logq_host, fb_host,
                                          allow_pack);
    if (rc != LGS_OK && h && h->host_compute) {
        const std::string keep = h->err;
        cudaStreamSynchronize(h->host_compute);
        cudaStreamSynchronize(h->host_copy);
        cudaGetLastError();
        h->err = keep;
    }
    return rc;
}

static int klein_host_common_impl(lgs_handle* h, int64_t S, const double* U_host, uint64_t seed, uint64_t first_id,
                                  int samplez, double* V_host, int64_t* X_host, double* logq_host, int32_t* fb_host,
                                  bool allow_pack) {
    if (!h) return LGS_E_INVALID;
    if (S < 0) return fail(h, LGS_E_INVALID, "S < 0");
    if (S == 0) return LGS_OK;
    LGS_CUDA(h, cudaSetDevice(h->device));
    const size_t n = h->n;
    const int ups = h->ctx.ups;
    {
        const int rcs = ensure_host_streams(h);
        if (rcs) return rcs;
    }
    // chunk = two full waves of contraction tiles on the blocked path (75 776 samples on a B200), else 64 Ki
    int64_t chunk = h->blocked.ready && h->blocked.wave > 0 ? 2 * h->blocked.wave : (int64_t)65536;
    if (h->blocked.ready && h->blocked.use_chain && h->blocked.chain.ready)
        chunk = h->blocked.chain.batch / 3;  // three tiles per SM (one per slot): short chunks overlap the PCIe copy best
    if (chunk > S) chunk = S;
    // widening needs host threads: below four per rank the plain fp64 copy (copy engine only) is the faster delivery
    if (allow_pack && h->pack_ok && V_host && !X_host && !logq_host && !U_host && hostpool_default_threads() >= 4) {
        const char* zc_env = getenv("LGS_HOST_ZEROCOPY");  // development A/B: 0 = chunked packed copies instead
        const bool zc_off = zc_env && atoi(zc_env) == 0;
        const bool chain = h->blocked.ready && h->blocked.use_chain && h->blocked.chain.ready &&
                           samplez == LGS_SAMPLEZ_FAST && h->path != LGS_PATH_ORDERED;
        if (chain && n % 8 == 0 && !zc_off) return klein_host_zerocopy(h, S, seed, first_id, V_host);
        return klein_host_packed(h, S, seed, first_id, samplez, V_host, chunk);
    }
    const size_t bV = V_host ? align256(sizeof(double) * chunk * n) : 0;
    const size_t bX = X_host ? align256(sizeof(int64_t) * chunk * n) : 0;
    const size_t bq = logq_host ? align256(sizeof(double) * chunk) : 0;
    const size_t bU = U_host ? align256(sizeof(double) * chunk * ups) : 0;
    const size_t slot = bV + bX + bq + bU;
    int rc = ensure_stage(h, 2 * slot + 256);
    if (rc) return rc;
    int32_t* dfb = (int32_t*)((uint8_t*)h->stage_dev + 2 * slot);
    cudaStream_t cs = h->host_compute, ps = h->host_copy;
    if (U_host && fb_host) LGS_CUDA(h, cudaMemsetAsync(dfb, 0, sizeof(int32_t), cs));
    // chunk k is enqueued on the compute stream first, then chunk k-1 leaves for the host: the staged copy of a
    // pageable destination blocks this thread, and the device works on chunk k meanwhile
    const int64_t K = (S + chunk - 1) / chunk;
    auto slot_ptrs = [&](int b, double*& dV, int64_t*& dX, double*& dq, double*& dU) {
        uint8_t* p = (uint8_t*)h->stage_dev + b * slot;
        dV = V_host ? (double*)p : nullptr; p += bV;
        dX = X_host ? (int64_t*)p : nullptr; p += bX;
        dq = logq_host ? (double*)p : nullptr; p += bq;
        dU = U_host ? (double*)p : nullptr;
    };
    for (int64_t k = 0; k <= K; k++) {
        double *dV, *dq, *dU;
        int64_t* dX;
        if (k < K) {
            const int64_t s0 = k * chunk, cnt = (S - s0) < chunk ? (S - s0) : chunk;
            const int b = (int)(k & 1);
            slot_ptrs(b, dV, dX, dq, dU);
            if (k >= 2) LGS_CUDA(h, cudaStreamWaitEvent(cs, h->ev_free[b], 0));  // slot b has left for the host
            if (U_host) {
                LGS_CUDA(h, cudaMemcpyAsync(dU, U_host + (size_t)s0 * ups, sizeof(double) * cnt * ups,
                                            cudaMemcpyHostToDevice, cs));
                // the fallback counter accumulates over chunks (no per-chunk reset)
                // host routes of Klein batches do not feed an attached histogram, whichever transport carries them
                h->blocked.chain.hist_pause = true;
                rc = klein_replay_impl(h, cnt, dU, dV, dX, dq, fb_host ? dfb : nullptr, cs);
                h->blocked.chain.hist_pause = false;
                if (rc) return rc;
            } else {
                h->blocked.chain.hist_pause = true;
                rc = lgs_klein_sample(h, cnt, seed, first_id + (uint64_t)s0, samplez, dV, dX, dq, cs);
                h->blocked.chain.hist_pause = false;
                if (rc) return rc;
            }
            LGS_CUDA(h, cudaEventRecord(h->ev_done[b], cs));
        }
        if (k >= 1) {
            const int64_t j = k - 1, s0 = j * chunk, cnt = (S - s0) < chunk ? (S - s0) : chunk;
            const int b = (int)(j & 1);
            slot_ptrs(b, dV, dX, dq, dU);
            LGS_CUDA(h, cudaStreamWaitEvent(ps, h->ev_done[b], 0));
            if (V_host && (rc = staged_d2h(h, V_host + (size_t)s0 * n, dV, sizeof(double) * cnt * n, ps))) return rc;
            if (X_host && (rc = staged_d2h(h, X_host + (size_t)s0 * n, dX, sizeof(int64_t) * cnt * n, ps))) return rc;
            if (logq_host) LGS_CUDA(h, cudaMemcpyAsync(logq_host + s0, dq, sizeof(double) * cnt, cudaMemcpyDeviceToHost, ps));
            LGS_CUDA(h, cudaEventRecord(h->ev_free[b], ps));
        }
    }
    LGS_CUDA(h, cudaStreamSynchronize(ps));
    if (fb_host && U_host) {
        LGS_CUDA(h, cudaMemcpyAsync(fb_host, dfb, sizeof(int32_t), cudaMemcpyDeviceToHost, cs));
    }
    LGS_CUDA(h, cudaStreamSynchronize(cs));
    return LGS_OK;
}

int lgs_host_widen_i16(const int16_t* src, double* dst, int64_t count, int32_t threads) {
    if (count < 0 || (count > 0 && (!src || !dst)) || threads < 0) return LGS_E_INVALID;
    if (!host_simd_ok()) {  // no AVX2 on this host: plain loop
        for (int64_t i = 0; i < count; i++) dst[i] = (double)src[i];
        return LGS_OK;
    }
    if (threads <= 1) {
        expand_i16_range(src, dst, (size_t)count);
        return LGS_OK;
    }
    HostPool* pool = hostpool_create(threads);
    hostpool_expand_i16(pool, src, dst, (size_t)count);
    hostpool_destroy(pool);
    return LGS_OK;
}

int lgs_host_transport(const lgs_handle* h) {
    if (!h) return LGS_E_INVALID;
    if (!h->pack_ok || hostpool_default_threads() < 4) return LGS_TRANSPORT_FP64;
    const char* zc_env = getenv("LGS_HOST_ZEROCOPY");
    const bool chain = h->blocked.ready && h->blocked.use_chain && h->blocked.chain.ready && h->path != LGS_PATH_ORDERED;
    return chain && h->n % 8 == 0 && !(zc_env && atoi(zc_env) == 0) ? LGS_TRANSPORT_INT16_ZEROCOPY : LGS_TRANSPORT_INT16_COPY;
}

int lgs_klein_sample_host(lgs_handle* h, int64_t S, uint64_t seed, uint64_t first_sample_id, int samplez,
                          double* V_host, int64_t* X_host, double* logq_host) {
    LGS_TRACE();
    return klein_host_common(h, S, nullptr, seed, first_sample_id, samplez, V_host, X_host, logq_host, nullptr, true);
}

int lgs_klein_replay_host(lgs_handle* h, int64_t S, const double* U_host, double* V_host, int64_t* X_host,
                          double* logq_host, int32_t* fallback_count_host) {
    LGS_TRACE();
    if (S > 0 && !U_host) return fail(h, LGS_E_INVALID, "lgs_klein_replay_host: NULL uniform stream");
    return klein_host_common(h, S, U_host, 0, 0, LGS_SAMPLEZ_TABLE, V_host, X_host, logq_host, fallback_count_host, true);
}

int lgs_klein_logdensity(lgs_handle* h, int64_t S, const int64_t* X_dev, double* logq_dev, void* stream) {
    LGS_TRACE();
    if (!h) return LGS_E_INVALID;
    if (S < 0 || (S > 0 && (!X_dev || !logq_dev))) return fail(h, LGS_E_INVALID, "lgs_klein_logdensity: bad arguments");
    LGS_CUDA(h, cudaSetDevice(h->device));
    LGS_CUDA(h, launch_klein_logdensity(h->ctx, S, X_dev, logq_dev, (cudaStream_t)stream));
    return LGS_OK;
}

int lgs_klein_logdensity_host(lgs_handle* h, int64_t S, const int64_t* X_host, double* logq_host) {
    if (!h) return LGS_E_INVALID;
    if (S < 0 || (S > 0 && (!X_host || !logq_host))) return fail(h, LGS_E_INVALID, "lgs_klein_logdensity_host: bad arguments");
    if (S == 0) return LGS_OK;
    LGS_CUDA(h, cudaSetDevice(h->device));
    size_t bX = align256(sizeof(int64_t) * S * h->n);
    int rc = ensure_stage(h, bX + sizeof(double) * S);
    if (rc) return rc;
    int64_t* dX = (int64_t*)h->stage_dev;
    double* dq = (double*)((uint8_t*)h->stage_dev + bX);
    LGS_CUDA(h, cudaMemcpy(dX, X_host, sizeof(int64_t) * S * h->n, cudaMemcpyHostToDevice));
    rc = lgs_klein_logdensity(h, S, dX, dq, nullptr);
    if (rc) return rc;
    LGS_CUDA(h, cudaMemcpy(logq_host, dq, sizeof(double) * S, cudaMemcpyDeviceToHost));
    return LGS_OK;
}

int lgs_imhk_reset(lgs_handle* h, int64_t chains, uint64_t seed, uint64_t first_chain_id) {
    LGS_TRACE();
    if (!h) return LGS_E_INVALID;
    if (chains <= 0) return fail(h, LGS_E_INVALID, "lgs_imhk_reset: chains must be positive");
    LGS_CUDA(h, cudaSetDevice(h->device));
    ChainSet& cs = h->chains;
    if (cs.chains != chains) {
        free_chains(cs);
        LGS_CUDA(h, cudaMalloc(&cs.logw, sizeof(double) * chains));
        LGS_CUDA(h, cudaMalloc(&cs.cur, sizeof(int32_t) * chains));
        LGS_CUDA(h, cudaMalloc(&cs.init, sizeof(int32_t) * chains));
        LGS_CUDA(h, cudaMalloc(&cs.counts, sizeof(int64_t) * 2 * chains));
        LGS_CUDA(h, cudaMalloc(&cs.step_no, sizeof(int64_t) * chains));
        LGS_CUDA(h, cudaMalloc(&cs.upos, sizeof(int64_t) * chains));
        cs.chains = chains;
    }
    // IMHK chains draw from their own Philox key: with the plain seed, chain c would start exactly at sample c of a
    // Klein sampler (and of its own proposal_sampler) created with the same seed, correlating the two arms of the
    // reference's IMHK-vs-Klein convergence studies at step 0
    cs.seed = seed ^ 0x494D484B5F4C4753ull;
    cs.first_id = first_chain_id;
    cs.state_path = 0;
    cs.bl_init = false;
    cs.bl_steps = 0;
    cs.steps_total = 0;
    cs.bl_loc_dirty = false;
    LGS_CUDA(h, cudaMemset(cs.logw, 0, sizeof(double) * chains));
    LGS_CUDA(h, cudaMemset(cs.cur, 0, sizeof(int32_t) * chains));
    LGS_CUDA(h, cudaMemset(cs.init, 0, sizeof(int32_t) * chains));
    LGS_CUDA(h, cudaMemset(cs.counts, 0, sizeof(int64_t) * 2 * chains));
    LGS_CUDA(h, cudaMemset(cs.step_no, 0, sizeof(int64_t) * chains));
    LGS_CUDA(h, cudaMemset(cs.upos, 0, sizeof(int64_t) * chains));
    return LGS_OK;
}

// the ordered kernels keep two state buffers per chain (current / proposal), allocated on first use
static int ensure_ordered_chains(lgs_handle* h) {
    ChainSet& cs = h->chains;
    if (cs.state_path == 2) return fail(h, LGS_E_STATE, "imhk: this chain set is owned by the blocked path; call lgs_imhk_reset()");
    if (!cs.Xbuf) {
        size_t buf = sizeof(double) * 2 * (size_t)h->n * cs.chains;
        LGS_CUDA(h, cudaMalloc(&cs.Xbuf, buf));
        LGS_CUDA(h, cudaMalloc(&cs.Vbuf, buf));
        LGS_CUDA(h, cudaMemset(cs.Xbuf, 0, buf));
        LGS_CUDA(h, cudaMemset(cs.Vbuf, 0, buf));
    }
    cs.state_path = 1;
    return LGS_OK;
}

// The step index of a chain is the Philox tag of its proposal / accept draws (tag 0 is the initial draw) and that tag
// is 32 bits wide: beyond 2^32 - 2 steps it would wrap onto streams already used.
constexpr int64_t kMaxImhkSteps = 0xFFFFFFFELL;

// Steps every chain has performed.  On the blocked / chain path the launch loops advance cs.bl_steps launch by launch
// (it IS the Philox position), so it is the one source there: a call that failed part-way leaves steps_total behind.
static int64_t imhk_steps_done_of(const lgs_handle* h) {
    return h->chains.state_path == 2 ? h->chains.bl_steps : h->chains.steps_total;
}

static int imhk_check(lgs_handle* h, int64_t burn_in, int64_t keep, int64_t thin) {
    if (!h) return LGS_E_INVALID;
    if (h->chains.chains <= 0) return fail(h, LGS_E_STATE, "imhk: call lgs_imhk_reset() first");
    if (burn_in < 0 || keep < 0 || thin < 1) return fail(h, LGS_E_INVALID, "imhk: need burn_in >= 0, keep >= 0, thin >= 1");
    const int64_t room = kMaxImhkSteps - imhk_steps_done_of(h);
    if (burn_in > room || (keep > 0 && thin > (room - burn_in) / keep))
        return fail(h, LGS_E_INVALID, "imhk: a chain cannot go beyond 2^32 - 2 steps (32-bit Philox step tag); "
                                      "lgs_imhk_reset() with another seed continues the run");
    return LGS_OK;
}

int lgs_imhk_run(lgs_handle* h, int64_t burn_in, int64_t keep, int64_t thin, int samplez, int weight_mode,
                 double* V_dev, int64_t* X_dev, void* stream) {
    LGS_TRACE();
    int rc = imhk_check(h, burn_in, keep, thin);
    if (rc) return rc;
    LGS_CUDA(h, cudaSetDevice(h->device));
    if (samplez != LGS_SAMPLEZ_TABLE && samplez != LGS_SAMPLEZ_FAST) return fail(h, LGS_E_INVALID, "imhk: unknown samplez");
    if (weight_mode != LGS_WEIGHT_REFERENCE && weight_mode != LGS_WEIGHT_EXACT) return fail(h, LGS_E_INVALID, "imhk: unknown weight mode");
    ChainSet& cs = h->chains;
    const bool want_blocked = samplez == LGS_SAMPLEZ_FAST && h->path != LGS_PATH_ORDERED && h->blocked.ready;
    if (cs.state_path == 2 || (cs.state_path == 0 && want_blocked)) {
        if (!want_blocked) return fail(h, LGS_E_STATE, "imhk: this chain set is owned by the blocked path (fast SampleZ); call lgs_imhk_reset()");
        cs.state_path = 2;
        LGS_CUDA(h, launch_imhk_blocked(h->blocked, cs, burn_in, keep, thin, weight_mode, V_dev, X_dev,
                                        (cudaStream_t)stream));
        cs.steps_total += burn_in + keep * thin;
        return LGS_OK;
    }
    rc = ensure_ordered_chains(h);
    if (rc) return rc;
    RngSpec rng{nullptr, 0, h->chains.seed, h->chains.first_id};
    LGS_CUDA(h, launch_imhk_ordered(h->ctx, h->chains, burn_in, keep, thin, rng, samplez, weight_mode, V_dev, X_dev,
                                    nullptr, nullptr, nullptr, (cudaStream_t)stream));
    cs.steps_total += burn_in + keep * thin;
    return LGS_OK;
}

int lgs_imhk_replay(lgs_handle* h, int64_t burn_in, int64_t keep, int64_t thin, int weight_mode, const double* U_dev,
                    int64_t stride, double* V_dev, int64_t* X_dev, uint8_t* accept_dev, double* logw_dev,
                    int32_t* fallback_count_dev, void* stream) {
    LGS_TRACE();
    int rc = imhk_check(h, burn_in, keep, thin);
    if (rc) return rc;
    if (!U_dev) return fail(h, LGS_E_INVALID, "lgs_imhk_replay: NULL uniform stream");
    LGS_CUDA(h, cudaSetDevice(h->device));
    if (fallback_count_dev) LGS_CUDA(h, cudaMemsetAsync(fallback_count_dev, 0, sizeof(int32_t), (cudaStream_t)stream));
    if (replay_on_chain(h) && h->chains.state_path != 1) {
        h->chains.state_path = 2;
        const ChainReplay rp{U_dev, stride, fallback_count_dev, accept_dev, logw_dev};
        LGS_CUDA(h, launch_imhk_chain(h->blocked, h->chains, burn_in, keep, thin, weight_mode, V_dev, X_dev,
                                      (cudaStream_t)stream, &rp));
        h->chains.steps_total += burn_in + keep * thin;
        return LGS_OK;
    }
    rc = ensure_ordered_chains(h);
    if (rc) return rc;
    RngSpec rng{U_dev, stride, 0, 0};
    LGS_CUDA(h, launch_imhk_ordered(h->ctx, h->chains, burn_in, keep, thin, rng, LGS_SAMPLEZ_TABLE, weight_mode, V_dev,
                                    X_dev, accept_dev, logw_dev, fallback_count_dev, (cudaStream_t)stream));
    h->chains.steps_total += burn_in + keep * thin;
    return LGS_OK;
}

int lgs_imhk_get_state(lgs_handle* h, int64_t* counts_dev, int64_t* X_dev, double* V_dev, double* logw_dev,
                       void* stream) {
    LGS_TRACE();
    if (!h) return LGS_E_INVALID;
    if (h->chains.chains <= 0) return fail(h, LGS_E_STATE, "imhk: call lgs_imhk_reset() first");
    LGS_CUDA(h, cudaSetDevice(h->device));
    if (h->chains.state_path == 0) return fail(h, LGS_E_STATE, "imhk: the chains have no state yet");
    if (h->chains.state_path == 2) {
        LGS_CUDA(h, launch_imhk_blocked_get_state(h->blocked, h->chains, counts_dev, X_dev, V_dev, logw_dev,
                                                  (cudaStream_t)stream));
        return LGS_OK;
    }
    LGS_CUDA(h, launch_imhk_get_state(h->ctx, h->chains, counts_dev, X_dev, V_dev, logw_dev, (cudaStream_t)stream));
    return LGS_OK;
}

int64_t lgs_imhk_steps_done(const lgs_handle* h) { return h ? imhk_steps_done_of(h) : (int64_t)LGS_E_INVALID; }

int lgs_imhk_set_state(lgs_handle* h, int samplez, const int64_t* X_dev, const double* logw_dev,
                       const int64_t* counts_dev, int64_t steps_done, void* stream) {
    LGS_TRACE();
    if (!h) return LGS_E_INVALID;
    if (h->chains.chains <= 0) return fail(h, LGS_E_STATE, "imhk: call lgs_imhk_reset() first");
    if (!X_dev || !logw_dev || steps_done < 0) return fail(h, LGS_E_INVALID, "lgs_imhk_set_state: bad arguments");
    if (steps_done > kMaxImhkSteps)
        return fail(h, LGS_E_INVALID, "lgs_imhk_set_state: steps_done beyond 2^32 - 2 (32-bit Philox step tag)");
    if (samplez != LGS_SAMPLEZ_TABLE && samplez != LGS_SAMPLEZ_FAST) return fail(h, LGS_E_INVALID, "imhk: unknown samplez");
    LGS_CUDA(h, cudaSetDevice(h->device));
    ChainSet& cs = h->chains;
    const bool want_blocked = samplez == LGS_SAMPLEZ_FAST && h->path != LGS_PATH_ORDERED && h->blocked.ready;
    if (want_blocked) {
        if (cs.state_path == 1) return fail(h, LGS_E_STATE, "imhk: this chain set is owned by the ordered kernels; call lgs_imhk_reset()");
        if (!(h->blocked.use_chain && h->blocked.chain.ready))
            return fail(h, LGS_E_STATE, "lgs_imhk_set_state: not offered by the LGS_CHAIN=0 development path");
        LGS_CUDA(h, launch_imhk_chain_set_state(h->blocked, cs, X_dev, logw_dev, counts_dev, steps_done, (cudaStream_t)stream));
        cs.state_path = 2;
    } else {
        int rc = ensure_ordered_chains(h);
        if (rc) return rc;
        LGS_CUDA(h, launch_imhk_set_state(h->ctx, cs, X_dev, logw_dev, counts_dev, steps_done, (cudaStream_t)stream));
    }
    cs.steps_total = steps_done;
    return LGS_OK;
}

int lgs_imhk_set_state_host(lgs_handle* h, int samplez, const int64_t* X_host, const double* logw_host,
                            const int64_t* counts_host, int64_t steps_done) {
    if (!h) return LGS_E_INVALID;
    if (h->chains.chains <= 0) return fail(h, LGS_E_STATE, "imhk: call lgs_imhk_reset() first");
    if (!X_host || !logw_host) return fail(h, LGS_E_INVALID, "lgs_imhk_set_state_host: NULL argument");
    LGS_CUDA(h, cudaSetDevice(h->device));
    const size_t n = h->n;
    const int64_t C = h->chains.chains;
    const size_t bX = align256(sizeof(int64_t) * C * n), bw = align256(sizeof(double) * C), bc = align256(sizeof(int64_t) * 2 * C);
    int rc = ensure_stage(h, bX + bw + bc);
    if (rc) return rc;
    uint8_t* p = (uint8_t*)h->stage_dev;
    int64_t* dX = (int64_t*)p; p += bX;
    double* dw = (double*)p; p += bw;
    int64_t* dc = (int64_t*)p;
    LGS_CUDA(h, cudaMemcpy(dX, X_host, sizeof(int64_t) * C * n, cudaMemcpyHostToDevice));
    LGS_CUDA(h, cudaMemcpy(dw, logw_host, sizeof(double) * C, cudaMemcpyHostToDevice));
    if (counts_host) LGS_CUDA(h, cudaMemcpy(dc, counts_host, sizeof(int64_t) * 2 * C, cudaMemcpyHostToDevice));
    rc = lgs_imhk_set_state(h, samplez, dX, dw, counts_host ? dc : nullptr, steps_done, nullptr);
    if (rc) return rc;
    LGS_CUDA(h, cudaDeviceSynchronize());
    return LGS_OK;
}

static int imhk_host_common(lgs_handle* h, int64_t burn_in, int64_t keep, int64_t thin, int samplez, int weight_mode,
                            const double* U_host, int64_t stride, double* V_host, int64_t* X_host,
                            uint8_t* accept_host, double* logw_host, int32_t* fb_host) {
    int rc = imhk_check(h, burn_in, keep, thin);
    if (rc) return rc;
    LGS_CUDA(h, cudaSetDevice(h->device));
    const size_t n = h->n;
    const int64_t C = h->chains.chains;
    const int64_t steps = burn_in + keep * thin;
    size_t bV = V_host ? align256(sizeof(double) * C * keep * n) : 0;
    size_t bX = X_host ? align256(sizeof(int64_t) * C * keep * n) : 0;
    size_t bA = accept_host ? align256((size_t)C * steps) : 0;
    size_t bW = logw_host ? align256(sizeof(double) * C * (steps + 1)) : 0;
    size_t bU = U_host ? align256(sizeof(double) * C * stride) : 0;
    rc = ensure_stage(h, bV + bX + bA + bW + bU + 256);
    if (rc) return rc;
    uint8_t* p = (uint8_t*)h->stage_dev;
    double* dV = V_host ? (double*)p : nullptr; p += bV;
    int64_t* dX = X_host ? (int64_t*)p : nullptr; p += bX;
    uint8_t* dA = accept_host ? p : nullptr; p += bA;
    double* dW = logw_host ? (double*)p : nullptr; p += bW;
    double* dU = U_host ? (double*)p : nullptr; p += bU;
    int32_t* dfb = (int32_t*)p;
    if (U_host) {
        LGS_CUDA(h, cudaMemcpy(dU, U_host, sizeof(double) * C * stride, cudaMemcpyHostToDevice));
        rc = lgs_imhk_replay(h, burn_in, keep, thin, weight_mode, dU, stride, dV, dX, dA, dW, dfb, nullptr);
    } else {
        rc = lgs_imhk_run(h, burn_in, keep, thin, samplez, weight_mode, dV, dX, nullptr);
    }
    if (rc) return rc;
    if (V_host && keep) {
        // kept states of an integer basis travel as int16 (same transport as lgs_klein_sample_host), once they are
        // large enough for the pipeline to matter
        if (h->pack_ok && hostpool_default_threads() >= 4 && (size_t)C * keep * n >= ((size_t)1 << 22)) {
            LGS_CUDA(h, cudaDeviceSynchronize());
            rc = ship_points_packed(h, dV, C * keep, V_host);
            if (rc) return rc;
        } else {
            LGS_CUDA(h, cudaDeviceSynchronize());
            if ((rc = ensure_host_streams(h))) return rc;
            if ((rc = staged_d2h(h, V_host, dV, sizeof(double) * C * keep * n, h->host_copy))) return rc;
            LGS_CUDA(h, cudaStreamSynchronize(h->host_copy));
        }
    }
    if (X_host && keep) {
        LGS_CUDA(h, cudaDeviceSynchronize());
        if ((rc = ensure_host_streams(h))) return rc;
        if ((rc = staged_d2h(h, X_host, dX, sizeof(int64_t) * C * keep * n, h->host_copy))) return rc;
        LGS_CUDA(h, cudaStreamSynchronize(h->host_copy));
    }
    if (accept_host && steps) LGS_CUDA(h, cudaMemcpy(accept_host, dA, (size_t)C * steps, cudaMemcpyDeviceToHost));
    if (logw_host) LGS_CUDA(h, cudaMemcpy(logw_host, dW, sizeof(double) * C * (steps + 1), cudaMemcpyDeviceToHost));
    if (fb_host && U_host) LGS_CUDA(h, cudaMemcpy(fb_host, dfb, sizeof(int32_t), cudaMemcpyDeviceToHost));
    LGS_CUDA(h, cudaDeviceSynchronize());
    return LGS_OK;
}

int lgs_imhk_run_host(lgs_handle* h, int64_t burn_in, int64_t keep, int64_t thin, int samplez, int weight_mode,
                      double* V_host, int64_t* X_host) {
    LGS_TRACE();
    return imhk_host_common(h, burn_in, keep, thin, samplez, weight_mode, nullptr, 0, V_host, X_host, nullptr, nullptr,
                            nullptr);
}

int lgs_imhk_replay_host(lgs_handle* h, int64_t burn_in, int64_t keep, int64_t thin, int weight_mode,
                         const double* U_host, int64_t stride, double* V_host, int64_t* X_host, uint8_t* accept_host,
                         double* logw_host, int32_t* fallback_count_host) {
    LGS_TRACE();
    if (!U_host) return h ? fail(h, LGS_E_INVALID, "lgs_imhk_replay_host: NULL uniform stream") : LGS_E_INVALID;
    return imhk_host_common(h, burn_in, keep, thin, LGS_SAMPLEZ_TABLE, weight_mode, U_host, stride, V_host, X_host,
                            accept_host, logw_host, fallback_count_host);
}

int lgs_imhk_get_state_host(lgs_handle* h, int64_t* counts_host, int64_t* X_host, double* V_host, double* logw_host) {
    LGS_TRACE();
    if (!h) return LGS_E_INVALID;
    if (h->chains.chains <= 0) return fail(h, LGS_E_STATE, "imhk: call lgs_imhk_reset() first");
    LGS_CUDA(h, cudaSetDevice(h->device));
    const size_t n = h->n;
    const int64_t C = h->chains.chains;
    size_t bc = align256(sizeof(int64_t) * 2 * C), bX = align256(sizeof(int64_t) * C * n);
    size_t bV = align256(sizeof(double) * C * n), bw = align256(sizeof(double) * C);
    int rc = ensure_stage(h, bc + bX + bV + bw);
    if (rc) return rc;
    uint8_t* p = (uint8_t*)h->stage_dev;
    int64_t* dc = (int64_t*)p; p += bc;
    int64_t* dX = (int64_t*)p; p += bX;
    double* dV = (double*)p; p += bV;
    double* dw = (double*)p;
    // only what the caller asked for: a counts-only poll (IMHKSampler.step) launches no kernel and copies 16 bytes per chain
    rc = lgs_imhk_get_state(h, counts_host ? dc : nullptr, X_host ? dX : nullptr, V_host ? dV : nullptr,
                            logw_host ? dw : nullptr, nullptr);
    if (rc) return rc;
    if (counts_host) LGS_CUDA(h, cudaMemcpy(counts_host, dc, sizeof(int64_t) * 2 * C, cudaMemcpyDeviceToHost));
    if (X_host) LGS_CUDA(h, cudaMemcpy(X_host, dX, sizeof(int64_t) * C * n, cudaMemcpyDeviceToHost));
    if (V_host) LGS_CUDA(h, cudaMemcpy(V_host, dV, sizeof(double) * C * n, cudaMemcpyDeviceToHost));
    if (logw_host) LGS_CUDA(h, cudaMemcpy(logw_host, dw, sizeof(double) * C, cudaMemcpyDeviceToHost));
    return LGS_OK;
}

int lgs_imhk_counts_host(lgs_handle* h, int64_t* accepted, int64_t* total) {
    LGS_TRACE();
    if (!h || !accepted || !total) return LGS_E_INVALID;
    if (h->chains.chains <= 0) return fail(h, LGS_E_STATE, "imhk: call lgs_imhk_reset() first");
    LGS_CUDA(h, cudaSetDevice(h->device));
    int rc = ensure_stage(h, 256);
    if (rc) return rc;
    LGS_CUDA(h, launch_reduce_counts(h->chains, (int64_t*)h->stage_dev, nullptr));
    int64_t out[2];
    LGS_CUDA(h, cudaMemcpy(out, h->stage_dev, sizeof(out), cudaMemcpyDeviceToHost));
    *accepted = out[0];
    *total = out[1];
    return LGS_OK;
}

int lgs_profile_enable(lgs_handle* h, int on) {
    if (!h) return LGS_E_INVALID;
    h->blocked.prof.on = on != 0;
    return LGS_OK;
}

int lgs_profile_read(lgs_handle* h, double* ms_by_class, int64_t* launches_by_class) {
    if (!h || !ms_by_class || !launches_by_class) return LGS_E_INVALID;
    LGS_CUDA(h, cudaSetDevice(h->device));
    LGS_CUDA(h, blocked_profile_read(h->blocked, ms_by_class, launches_by_class, LGS_PROF_CLASSES));
    return LGS_OK;
}

int lgs_histogram(lgs_handle* h, const double* V_dev, int64_t S, const int32_t* dims_host, int32_t n_dims, double lo,
                  double width, int32_t nbins, int64_t* hist_dev, void* stream) {
    LGS_TRACE();
    if (!h) return LGS_E_INVALID;
    if (!V_dev || !dims_host || !hist_dev || n_dims <= 0 || nbins <= 0 || !(width > 0.0))
        return fail(h, LGS_E_INVALID, "lgs_histogram: bad arguments");
    for (int i = 0; i < n_dims; i++)
        if (dims_host[i] < 0 || dims_host[i] >= h->n) return fail(h, LGS_E_INVALID, "lgs_histogram: dim out of range");
    LGS_CUDA(h, cudaSetDevice(h->device));
    // the tracked coordinates go to the kernels by value (256 per launch): nothing per-handle is written here
    LGS_CUDA(h, launch_histogram(h->ctx, V_dev, S, dims_host, n_dims, lo, width, nbins, hist_dev, (cudaStream_t)stream));
    return LGS_OK;
}

int lgs_attach_histogram(lgs_handle* h, const int32_t* dims_host, int32_t n_dims, double lo, double width, int32_t nbins,
                         int64_t* hist_dev) {
    LGS_TRACE();
    if (!h) return LGS_E_INVALID;
    LGS_CUDA(h, cudaSetDevice(h->device));
    if (!h->blocked.ready || !h->blocked.use_chain || !h->blocked.chain.ready || h->path == LGS_PATH_ORDERED)
        return fail(h, LGS_E_STATE, "lgs_attach_histogram: this handle is not served by the chain kernel (use lgs_histogram)");
    if (hist_dev) {
        if (!dims_host || n_dims <= 0 || n_dims > 32767 || nbins <= 0 || !(width > 0.0))
            return fail(h, LGS_E_INVALID, "lgs_attach_histogram: bad arguments");
        for (int i = 0; i < n_dims; i++)
            if (dims_host[i] < 0 || dims_host[i] >= h->n)
                return fail(h, LGS_E_INVALID, "lgs_attach_histogram: dim out of range");
    }
    LGS_CUDA(h, chain_attach_histogram(h->blocked, dims_host, n_dims, lo, width, nbins, hist_dev));
    return LGS_OK;
}

int lgs_chain_diagnostics(lgs_handle* h, const double* V_dev, int64_t chains, int64_t keep, const int32_t* dims_host,
                          int32_t n_dims, double c_window, int32_t max_lag, double* stats_dev, double* acf_dev,
                          int32_t acf_len, void* stream) {
    LGS_TRACE();
    if (!h) return LGS_E_INVALID;
    if (!V_dev || !dims_host || !stats_dev || chains <= 0 || keep < 2 || n_dims <= 0 || n_dims > 256 || !(c_window > 0.0))
        return fail(h, LGS_E_INVALID, "lgs_chain_diagnostics: bad arguments (need keep >= 2, 1 <= n_dims <= 256)");
    if (acf_dev && acf_len <= 0) return fail(h, LGS_E_INVALID, "lgs_chain_diagnostics: acf_len must be positive");
    for (int i = 0; i < n_dims; i++)
        if (dims_host[i] < 0 || dims_host[i] >= h->n) return fail(h, LGS_E_INVALID, "lgs_chain_diagnostics: dim out of range");
    LGS_CUDA(h, cudaSetDevice(h->device));
    if (max_lag <= 0) max_lag = (int32_t)(keep / 4);                 // convergence_diag.py:86
    if (max_lag > keep - 1) max_lag = (int32_t)(keep - 1);
    if (acf_dev && acf_len > max_lag + 1) return fail(h, LGS_E_INVALID, "lgs_chain_diagnostics: acf_len > max_lag + 1");
    if (!h->hist_dims) LGS_CUDA(h, cudaMalloc(&h->hist_dims, sizeof(int32_t) * 256));
    const size_t need = sizeof(double) * (size_t)chains * n_dims * keep;
    if (h->diag_cap < need) {
        cudaFree(h->diag_scratch);
        h->diag_scratch = nullptr;
        h->diag_cap = 0;
        LGS_CUDA(h, cudaMalloc(&h->diag_scratch, need));
        h->diag_cap = need;
    }
    LGS_CUDA(h, cudaMemcpyAsync(h->hist_dims, dims_host, sizeof(int32_t) * n_dims, cudaMemcpyHostToDevice,
                                (cudaStream_t)stream));
    LGS_CUDA(h, launch_chain_diagnostics(V_dev, chains, keep, h->n, h->hist_dims, n_dims, c_window, max_lag,
                                         h->diag_scratch, stats_dev, acf_dev, acf_dev ? acf_len : 0,
                                         (cudaStream_t)stream));
    return LGS_OK;
}

// lgs_chain_diagnostics for callers that cannot allocate device memory (see the block comment in lgs_diag.cu): same
// checks, same kernels on the same gathered series, the work array in the library's static device array.  This
// function and everything it calls is free of cudaMalloc / cudaFree / cudaMemcpy; it does not touch the handle's
// own diag_scratch / hist_dims buffers either, so it may be mixed freely with lgs_chain_diagnostics.
int lgs_chain_diagnostics_static(lgs_handle* h, const double* V_dev, int64_t chains, int64_t keep,
                                 const int32_t* dims_host, int32_t n_dims, double c_window, int32_t max_lag,
                                 double* stats_dev, double* acf_dev, int32_t acf_len, void* stream) {
    LGS_TRACE();
    if (!h) return LGS_E_INVALID;
    if (!V_dev || !dims_host || !stats_dev || chains <= 0 || keep < 2 || n_dims <= 0 || n_dims > 256 || !(c_window > 0.0))
        return fail(h, LGS_E_INVALID, "lgs_chain_diagnostics_static: bad arguments (need keep >= 2, 1 <= n_dims <= 256)");
    if (acf_dev && acf_len <= 0) return fail(h, LGS_E_INVALID, "lgs_chain_diagnostics_static: acf_len must be positive");
    for (int i = 0; i < n_dims; i++)
        if (dims_host[i] < 0 || dims_host[i] >= h->n)
            return fail(h, LGS_E_INVALID, "lgs_chain_diagnostics_static: dim out of range");
    // chains * n_dims * keep doubles of work space, compared without forming a product that could overflow
    const int64_t cap = diag_static_capacity();
    const int64_t per_chain = (int64_t)n_dims * keep;            // n_dims <= 256: overflows only beyond any capacity
    if (keep > cap || per_chain > cap || chains > cap / per_chain)
        return fail(h, LGS_E_INVALID,
                    "lgs_chain_diagnostics_static: chains * n_dims * keep = " + std::to_string(chains) + " * " +
                        std::to_string(n_dims) + " * " + std::to_string(keep) +
                        " series values exceed the static work array (LGS_DIAG_STATIC_MAX_DOUBLES = " +
                        std::to_string(cap) + "); split the chains over several calls or use lgs_chain_diagnostics");
    LGS_CUDA(h, cudaSetDevice(h->device));
    if (max_lag <= 0) max_lag = (int32_t)(keep / 4);                 // convergence_diag.py:86
    if (max_lag > keep - 1) max_lag = (int32_t)(keep - 1);
    if (acf_dev && acf_len > max_lag + 1)
        return fail(h, LGS_E_INVALID, "lgs_chain_diagnostics_static: acf_len > max_lag + 1");
    LGS_CUDA(h, launch_chain_diagnostics_static(h->device, V_dev, chains, keep, h->n, dims_host, n_dims, c_window,
                                                max_lag, stats_dev, acf_dev, acf_dev ? acf_len : 0,
                                                (cudaStream_t)stream));
    return LGS_OK;
}

int lgs_gelman_rubin_sums(lgs_handle* h, const double* stats_dev, int64_t chains, int32_t n_dims, double* sums_dev,
                          void* stream) {
    LGS_TRACE();
    if (!h) return LGS_E_INVALID;
    if (!stats_dev || !sums_dev || chains <= 0 || n_dims <= 0) return fail(h, LGS_E_INVALID, "lgs_gelman_rubin_sums: bad arguments");
    LGS_CUDA(h, cudaSetDevice(h->device));
    LGS_CUDA(h, launch_gelman_rubin_sums(stats_dev, chains, n_dims, sums_dev, (cudaStream_t)stream));
    return LGS_OK;
}

}  // extern "C"
