// lgs_hostpack.cpp -- host side of the packed device-to-host transport used by lgs_klein_sample_host.
//
// For an integer basis every lattice point has integer coordinates, so the fp64 [S, n] result (klein.py:322) crosses
// PCIe as int16 (a quarter of the bytes; the device-side pack kernel proves that every value round-trips, otherwise the
// chunk is re-sent as fp64) and is widened back into the caller's float64 buffer here, by a small pool of host threads
// with non-temporal stores.  This is transport only: the values delivered are bit-identical to the fp64 copy.
#include <immintrin.h>
#include <sched.h>

#include <condition_variable>
#include <cstdint>
#include <cstdlib>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#include "lgs_internal_host.h"

namespace lgs {

struct HostPool {
    std::vector<std::thread> workers;
    std::mutex mtx;
    std::condition_variable cv_go, cv_done;
    std::function<void(int, int)> job;
    uint64_t generation = 0;
    int pending = 0;
    bool stop = false;

    explicit HostPool(int threads) {
        try {
            for (int t = 0; t < threads; t++) workers.emplace_back([this, t, threads] { loop(t, threads); });
        } catch (...) {  // thread limit reached: release the ones that did start, let the caller fall back
            shutdown();
            throw;
        }
    }
    ~HostPool() { shutdown(); }
    void shutdown() {
        {
            std::lock_guard<std::mutex> lk(mtx);
            stop = true;
        }
        cv_go.notify_all();
        for (auto& w : workers) w.join();
        workers.clear();
    }
    void loop(int tid, int threads) {
        uint64_t seen = 0;
        for (;;) {
            std::function<void(int, int)> f;
            {
                std::unique_lock<std::mutex> lk(mtx);
                cv_go.wait(lk, [&] { return stop || generation != seen; });
                if (stop) return;
                seen = generation;
                f = job;
            }
            f(tid, threads);
            {
                std::lock_guard<std::mutex> lk(mtx);
                if (--pending == 0) cv_done.notify_one();
            }
        }
    }
    void run(std::function<void(int, int)> f) {
        std::unique_lock<std::mutex> lk(mtx);
        job = std::move(f);
        pending = (int)workers.size();
        generation++;
        cv_go.notify_all();
        cv_done.wait(lk, [&] { return pending == 0; });
    }
};

bool host_simd_ok() {
    static const bool ok = __builtin_cpu_supports("avx2");
    return ok;
}

// LGS_HOST_THREADS, else the CPUs this process may run on divided by the ranks sharing the host (LOCAL_WORLD_SIZE of
// torchrun): every rank of a multi-GPU job brings its own pool, and the pool threads spin on completion words
int hostpool_default_threads() {
    if (const char* e = std::getenv("LGS_HOST_THREADS")) {
        const int t = std::atoi(e);
        if (t > 0) return t > 64 ? 64 : t;
    }
    int cpus = 0;
    cpu_set_t set;
    if (sched_getaffinity(0, sizeof(set), &set) == 0) cpus = CPU_COUNT(&set);
    if (cpus <= 0) cpus = (int)std::thread::hardware_concurrency();
    if (cpus <= 0) cpus = 4;
    int ranks = 1;
    if (const char* e = std::getenv("LOCAL_WORLD_SIZE")) ranks = std::atoi(e) > 0 ? std::atoi(e) : 1;
    int t = cpus / ranks;
    // Ranks that share a host compete for its memory system, not for its cores: a widening thread spends most of its
    // time on DRAM reads of freshly DMA-written int16 and on store buffers, and half again as many threads as cores
    // moved the 8-rank delivery from 12.9 to 14.2 M samples/s (tools/e2e_ranks.py; 3 / 4 / 6 threads per rank on 32 CPUs)
    if (ranks >= 4) t += t / 2;
    if (t < 1) t = 1;
    return t > 64 ? 64 : t;
}

HostPool* hostpool_create(int threads) {
    if (threads <= 0) threads = hostpool_default_threads();
    try {
        return new HostPool(threads);
    } catch (...) {
        return nullptr;  // every user below runs the job on the calling thread when there is no pool
    }
}

void hostpool_destroy(HostPool* p) { delete p; }

int hostpool_threads(const HostPool* p) { return p ? (int)p->workers.size() : 0; }

// dst[i] = (double)src[i] for i in [0, count)
void expand_i16_range(const int16_t* src, double* dst, size_t count) {
    size_t i = 0;
    while (i < count && ((uintptr_t)(dst + i) & 31)) {  // reach a 32-byte boundary of the destination
        dst[i] = (double)src[i];
        i++;
    }
    for (; i + 8 <= count; i += 8) {
        if ((i & 31) == 0) _mm_prefetch((const char*)(src + i + 512), _MM_HINT_NTA);   // the source is read once
        const __m128i v = _mm_loadu_si128((const __m128i*)(src + i));
        const __m256i w = _mm256_cvtepi16_epi32(v);
        const __m256d lo = _mm256_cvtepi32_pd(_mm256_castsi256_si128(w));
        const __m256d hi = _mm256_cvtepi32_pd(_mm256_extracti128_si256(w, 1));
        _mm256_stream_pd(dst + i, lo);
        _mm256_stream_pd(dst + i + 4, hi);
    }
    for (; i < count; i++) dst[i] = (double)src[i];
    _mm_sfence();
}

void hostpool_run(HostPool* p, const std::function<void(int, int)>& job) {
    if (p)
        p->run(job);
    else
        job(0, 1);
}

void hostpool_expand_i16(HostPool* p, const int16_t* src, double* dst, size_t count) {
    if (!p || count < (size_t)1 << 16) {
        expand_i16_range(src, dst, count);
        return;
    }
    p->run([=](int tid, int threads) {
        const size_t per = ((count + threads - 1) / threads + 63) & ~(size_t)63;
        const size_t lo = per * (size_t)tid;
        if (lo >= count) return;
        const size_t hi = lo + per < count ? lo + per : count;
        expand_i16_range(src + lo, dst + lo, hi - lo);
    });
}

// dst[0, bytes) = src[0, bytes) on the calling thread; non-temporal stores once the destination is 32-byte aligned
static void copy_range_nt(const uint8_t* src, uint8_t* dst, size_t bytes) {
    size_t i = 0;
    while (i < bytes && ((uintptr_t)(dst + i) & 31)) {
        dst[i] = src[i];
        i++;
    }
    for (; i + 64 <= bytes; i += 64) {
        const __m256i a = _mm256_loadu_si256((const __m256i*)(src + i));
        const __m256i b = _mm256_loadu_si256((const __m256i*)(src + i + 32));
        _mm256_stream_si256((__m256i*)(dst + i), a);
        _mm256_stream_si256((__m256i*)(dst + i + 32), b);
    }
    for (; i < bytes; i++) dst[i] = src[i];
    _mm_sfence();
}

void hostpool_copy(HostPool* p, const void* src, void* dst, size_t bytes) {
    if (!p || bytes < (size_t)1 << 18) {
        copy_range_nt((const uint8_t*)src, (uint8_t*)dst, bytes);
        return;
    }
    p->run([=](int tid, int threads) {
        const size_t per = ((bytes + threads - 1) / threads + 4095) & ~(size_t)4095;
        const size_t lo = per * (size_t)tid;
        if (lo >= bytes) return;
        const size_t hi = lo + per < bytes ? lo + per : bytes;
        copy_range_nt((const uint8_t*)src + lo, (uint8_t*)dst + lo, hi - lo);
    });
}

}  // namespace lgs
