// Host-only helpers shared by lgs_capi.cu and lgs_hostpack.cpp (no CUDA types here).
#pragma once
#include <cstddef>
#include <cstdint>
#include <functional>

namespace lgs {

bool host_simd_ok();                     // the CPU has AVX2 (the widening / non-temporal copy loops use it)
struct HostPool;
int hostpool_default_threads();          // LGS_HOST_THREADS, else usable CPUs / LOCAL_WORLD_SIZE
HostPool* hostpool_create(int threads);  // threads <= 0: hostpool_default_threads()
void hostpool_destroy(HostPool* p);
int hostpool_threads(const HostPool* p);
// dst[i] = (double)src[i], split over the pool, non-temporal stores
void hostpool_expand_i16(HostPool* p, const int16_t* src, double* dst, size_t count);

// dst[0, bytes) = src[0, bytes), split over the pool, non-temporal stores (pinned staging -> pageable destination)
void hostpool_copy(HostPool* p, const void* src, void* dst, size_t bytes);

// run job(thread id, thread count) on every pool thread and wait for all of them
void hostpool_run(HostPool* p, const std::function<void(int, int)>& job);
// dst[i] = (double)src[i] on the calling thread, non-temporal stores
void expand_i16_range(const int16_t* src, double* dst, size_t count);

}  // namespace lgs
