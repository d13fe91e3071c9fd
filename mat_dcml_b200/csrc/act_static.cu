// matas_policy_act_autoregressive_static: the autoregressive act without any allocation.
//
// Who needs this.  Some hosts of this library may not allocate device memory once they are up and
// running: a decision server that shares the GPU with another framework whose caching allocator
// owns the remaining memory (a raw cudaMalloc there can fail or fragment its pool), or a
// latency-critical decision loop that cannot afford the device-wide synchronisation that
// cudaMalloc / cudaFree imply while other streams are busy.  The regular entry points grow their
// scratch (cross-attention queries and the decode K/V cache) with cudaMalloc the first time a larger
// batch arrives; this variant instead keeps that scratch in a fixed array that is part of the
// library image, so after start-up (matas_policy_create + matas_policy_set_weights) a caller can
// take decisions without the library ever touching the allocator.
//
// Cost and limits.  The array below is MATAS_STATIC_MAX_TOKENS * 640 floats = 5 MiB of device
// memory on EVERY GPU that loads the library, held until the process exits, whether or not this
// entry point is used.  A call may cover at most MATAS_STATIC_MAX_TOKENS = 2048 (env, agent) pairs
// (B * n_agent), e.g. 20 envs at the 101-agent DCML default or 97 envs at 21 agents; larger calls
// are rejected with an error, they are never split or silently routed to the allocating path.
//
// Sharing.  All calls on a device use the same array and find the previous call's contents in it.
// Nothing has to start at zero: the kernels write every query row and every K/V row of a position
// before any read of it, so no clearing is done or needed.  Two calls that overlap on the GPU
// (different streams, or different handles on one device) would overwrite each other's cache and
// produce garbage: serialise them on one stream.
//
// Set-up.  In host code `g_static_scratch` is only a host-side stand-in for the device array, not
// an address a kernel can use, so the first call on each device (with that device current) asks the
// runtime for the real device address with cudaGetSymbolAddress and remembers it.  Make that first
// call before starting any stream capture.
#include <stdlib.h>

#include <mutex>

#include "../../include/matas_b200.h"
#include "policy_fwd.h"

#define MATAS_STATIC_MAX_TOKENS 2048
// per token: up to 10 rows of 64 floats -- 4 folded cross-attention query rows + 4 cached attention-input
// rows are in use (2 spare)
#define MATAS_STATIC_FLOATS ((size_t)MATAS_STATIC_MAX_TOKENS * 64 * (2 + 8))

static __device__ __align__(256) float g_static_scratch[MATAS_STATIC_FLOATS];

namespace {
constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
float* g_dev_ptr[MAX_DEVICES] = {};

// device address of g_static_scratch on the CURRENT device, resolved once per device
int static_scratch(int device, float** out) {
    ARG_CHECK(device >= 0 && device < MAX_DEVICES, "device index %d out of range", device);
    std::lock_guard<std::mutex> lock(g_mu);
    if (g_dev_ptr[device] == nullptr) {
        float* ptr = nullptr;
        CUDA_TRY(cudaGetSymbolAddress((void**)&ptr, g_static_scratch));
        ARG_CHECK(ptr != nullptr, "cudaGetSymbolAddress returned a null address");
        g_dev_ptr[device] = ptr;
    }
    *out = g_dev_ptr[device];
    return MATAS_OK;
}
}  // namespace

extern "C" int matas_policy_act_autoregressive_static(matas_policy* p, const float* obs_rep, const float* ava, long B,
                                                      int deterministic, int stride, uint64_t seed, uint32_t step,
                                                      long env_offset, const float* u, const float* z, float* actions,
                                                      float* logp, void* stream) {
    PolicyView v;
    int rc = matas_policy_view(p, &v);
    if (rc) return rc;
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    ARG_CHECK(dev == v.device, "policy handle belongs to device %d but device %d is current", v.device, dev);
    ARG_CHECK(v.weights_set, "matas_policy_set_weights has not been called");
    ARG_CHECK(obs_rep && actions && logp && B > 0, "bad argument (B=%ld)", B);
    const int L = v.n_agent;
    if (B * L > MATAS_STATIC_MAX_TOKENS) {
        matas_set_error("matas_policy_act_autoregressive_static: %ld envs x %d agents = %ld tokens exceeds the static "
                        "scratch limit of %d tokens; use matas_policy_act_autoregressive (which allocates) or send "
                        "smaller batches", B, L, B * L, MATAS_STATIC_MAX_TOKENS);
        return MATAS_ERR_LIMIT;
    }
    float* scratch = nullptr;
    if ((rc = static_scratch(dev, &scratch))) return rc;
    const long n_tok = B * L;
    PolicyBufs b = {};
    b.w = v.packed;
    b.x1tab = v.x1tab;                        // block-0 table of the stochastic decode (allocated at set_weights)
    b.rep = const_cast<float*>(obs_rep);      // read-only here: q2_from_rep copies only when rep_in != rep
    for (int k = 0; k < 2; ++k)
        for (int h = 0; h < 2; ++h) b.Q2p[k][h] = scratch + (size_t)(k * 2 + h) * n_tok * 64;
    b.kv = scratch + (size_t)MATAS_STATIC_MAX_TOKENS * 64 * 4;   // [B env slots][4][L][64] <= MAX_TOKENS*256 floats
    if ((rc = launch_q2_from_rep(b, obs_rep, B, L, (cudaStream_t)stream, 1))) return rc;
    DecodeArgs a;
    a.B = B; a.L = L; a.mode = v.mode; a.deterministic = deterministic ? 1 : 0; a.stride = stride;
    a.n_steps = policy_decode_steps(L, a.mode, a.deterministic, stride);
    a.seed = seed; a.step = step; a.env_offset = env_offset;
    a.ava = ava; a.u = u; a.z = z; a.actions = actions; a.logp = logp;
    int grid, rows; long slots;
    if ((rc = policy_decode_geometry(L, B, v.n_sm, &grid, &slots, &rows))) return rc;
    // The kernel addresses the cache as [grid * rows env slots][4][L][64]; tiles are padded to a multiple of four envs,
    // so the slot count can exceed B: the padded geometry must fit the part of the array behind the query rows.
    ARG_CHECK((size_t)slots * 4 * L * 64 <= MATAS_STATIC_FLOATS - (size_t)MATAS_STATIC_MAX_TOKENS * 64 * 4,
              "static scratch too small for %ld padded env slots x %d agents", slots, L);
    a.rows = rows;
    return launch_decode_tc(b, a, grid, (cudaStream_t)stream);
}
