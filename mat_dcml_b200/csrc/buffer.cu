// Minibatch gather of the rollout buffer for sm_100a: ONE launch copies row idx[i] of every buffer array into
// row i of the minibatch arrays (SharedReplayBuffer.feed_forward_generator_transformer,
// mat_src/mat/utils/shared_buffer.py:265-314, which fancy-indexes each of its ten arrays separately).  A "row" is
// one (t, env) sample carrying all n_agent agents (the reference's `[rows, cols]` grid with cols = arange,
// shared_buffer.py:15-19,262); the share_obs row is one (W+2) vector per sample in the buffer and is replicated
// n_agent times on output, which is the reference's (M * n_agent, W + 2) share_obs_batch.  HBM-bound copy:
// one CTA per output row, 128-bit accesses where the row width and base allow.
#include "../../include/matas_b200.h"
#include "common.cuh"

namespace {

constexpr int MAX_ARR = 12;
struct GatherJobs {
    int n;
    const float* src[MAX_ARR];
    float* dst[MAX_ARR];
    int width[MAX_ARR];      // floats per source row
    int rep[MAX_ARR];        // the source row is written `rep` times back to back (share_obs: n_agent)
};

__global__ void __launch_bounds__(128) gather_rows_kernel(GatherJobs jobs, const long* __restrict__ idx, long n_rows) {
    const long row = blockIdx.x;
    if (row >= n_rows) return;
    const long src_row = idx[row];
    for (int a = 0; a < jobs.n; ++a) {
        const int w = jobs.width[a], rep = jobs.rep[a];
        const float* s = jobs.src[a] + src_row * w;
        float* d = jobs.dst[a] + row * (long)w * rep;
        const bool vec = (w & 3) == 0 && ((reinterpret_cast<uintptr_t>(s) | reinterpret_cast<uintptr_t>(d)) & 15) == 0;
        if (vec) {
            const int w4 = w >> 2;
            for (int i = threadIdx.x; i < w4 * rep; i += blockDim.x) {
                const int r = i / w4, c = i - r * w4;
                reinterpret_cast<float4*>(d + (long)r * w)[c] = __ldg(reinterpret_cast<const float4*>(s) + c);
            }
        } else {
            for (int i = threadIdx.x; i < w * rep; i += blockDim.x) {
                const int r = i / w, c = i - r * w;
                d[(long)r * w + c] = __ldg(s + c);
            }
        }
    }
}

}  // namespace

extern "C" int matas_gather_rows(int n_arrays, const float* const* src, float* const* dst, const int* width, const int* repeat,
                                 const long* idx, long n_rows, void* stream) {
    ARG_CHECK(n_arrays > 0 && n_arrays <= MAX_ARR, "n_arrays %d not in [1,%d]", n_arrays, MAX_ARR);
    ARG_CHECK(src && dst && width && repeat && idx && n_rows > 0, "bad argument (n_rows=%ld)", n_rows);
    GatherJobs j;
    j.n = n_arrays;
    for (int a = 0; a < n_arrays; ++a) {
        ARG_CHECK(src[a] && dst[a] && width[a] > 0 && repeat[a] > 0, "array %d: null pointer or empty row", a);
        j.src[a] = src[a]; j.dst[a] = dst[a]; j.width[a] = width[a]; j.rep[a] = repeat[a];
    }
    ARG_CHECK(n_rows < (1L << 31), "n_rows %ld too large", n_rows);
    gather_rows_kernel<<<(unsigned)n_rows, 128, 0, (cudaStream_t)stream>>>(j, idx, n_rows);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
