// Internal declarations shared by the translation units of liblgs_b200.so.
#pragma once
#include <cstdint>
#include <string>
#include <cuda_runtime.h>

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
};

struct Scratch {
    double* Xw = nullptr;  // [n][S_cap] coefficient workspace of the ordered Klein kernel
    int64_t cap = 0;
};

}  // namespace lgs

struct lgs_handle {
    int device = 0;
    int path = LGS_PATH_AUTO;
    int n = 0;
    lgs::DeviceCtx ctx{};       // host copy; device pointers inside
    void* arena = nullptr;      // one allocation holding every per-sampler constant
    lgs::Scratch scratch;
    lgs::ChainSet chains;
    // staging for the *_host entry points
    void* stage_dev = nullptr;
    size_t stage_cap = 0;
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
cudaError_t launch_histogram(const DeviceCtx& ctx, const double* V, int64_t S, const int32_t* dims_dev, int n_dims,
                             double lo, double width, int nbins, int64_t* hist, cudaStream_t st);
cudaError_t launch_reduce_counts(const ChainSet& cs, int64_t* out2, cudaStream_t st);

}  // namespace lgs
