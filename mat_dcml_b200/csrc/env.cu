// Vectorised DCML simulator for sm_100a: one thread per (env, worker), whole envs per CTA.
//
// Restates Env.step / Env.reset (DCML_BID_FIRST_MA_ENV_SingleProcess.py:57-144,157-274),
// Worker.process / bid (DCML_Worker_TIMESLOT_MultiProcess.py:46-135), Master.reset /
// get_workload (DCML_Master.py:35-56) and the ShareDummyVecEnv double reset
// (mat_src/mat/envs/env_wrappers.py:730-736).  All scalar time arithmetic is float64 in the
// reference's evaluation order (this file is compiled with -fmad=false), because the simulator
// branches on floor/ceil/compare of those values.  Randomness: named
// Philox draws (common.cuh / DESIGN.md).  HBM traffic per env-step is the 160 B/worker workload
// table read + written, the Pr / flag state, and the observation rows.
#include "../../include/matas_b200.h"
#include <stdlib.h>

#include "common.cuh"

namespace {

constexpr int PERIOD = 20;
constexpr double DATA_RATE = 157286400.0;   // 150 * 2^20
constexpr double FREQUENCY = 2e9;
constexpr double PR_MAX = 0.95;
constexpr double DONE_P = 0.8;
constexpr int R_MIN = 1024, R_MAX = 1048576, C_MIN = 32, C_MAX = 1024;
constexpr int R_HI = 1153434, C_HI = 1126;   // round(1.1 * max), DCML_Master.py:47-48
constexpr double NOISE = 0x1.4f8b588e368f1p-17;   // python 10**(-50/10), Shannon.py:9
constexpr double B_TOTAL = 100e9;                 // DCML_Master.py:29
constexpr int RETRY_CAP = 4096, SLOT_CAP = 4096;
constexpr int LWS = 21;                      // padded per-thread stride of the shared workload copy

enum { ENV_INIT = 0, ENV_STEP = 1, ENV_OBSERVE = 2, ENV_RESET = 3 };   // RESET: a new context, workload tables kept

struct EnvState {
    long B;
    int W, tpe;            // threads per env (W rounded up to a warp multiple)
    uint64_t seed;
    long env_offset;
    int double_reset;
    int fixed, standalone; // Env(fixed=True) strategy; step(standalone=True) branch on every step
    int shannon;           // Shannon link rates instead of the fixed NON_SHANNON_DATA_RATE
    double* rate_down;     // [B][W] download rate used by Worker.process for both transfers (Shannon mode)
    double* rate_up;       // [B][W] upload rate (observed only)
    const double* base;    // [W][20]
    double* lw;            // [B][W][20] current jittered workload == next step's local workload
    double* R;             // [B]
    double* C;             // [B]
    double* Pr;            // [B][W]
    int* arrive;           // [B]
    int* kdis;             // [B]
    uint8_t* disabled;     // [B][W]
};

struct EnvIO {
    const float* actions;  // [B][L]
    float* obs;            // [B][L][7]
    float* sob;            // [B][W+2]
    float* ava;            // [B][L][2]
    float* reward;         // [B]
    uint8_t* done;         // [B]
    float* delay;          // [B]
    float* payment;        // [B]
    int binary;            // reset(binary=True): share_obs rows start with the 32 + 32 bits of R and C (then width + 62 wide)
};

// The n-th geometric-retry uniform of a worker (stream RETRY: block n / 4, lane n % 4), consumed as the comparison
// u < Pr.  u = m 2^-24 with m the draw's top 24 bits, so u < Pr  <=>  m < ceil(Pr 2^24): an integer compare, no float64
// conversion per draw.  Two Philox blocks are kept: the one being consumed and the next, which `ensure()` generates at
// points where the warp is converged (loop entry, every fourth trip) -- the lanes of a warp sit at different draw
// indices, and a refill inside the per-draw path would make the whole warp pay a Philox call on almost every trip.
__device__ __noinline__ uint4 philox_nl(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint64_t seed) {
    return philox4x32(c0, c1, c2, c3, seed);
}
// Shannon.upload / download (Shannon.py:14-21) for one worker; kept out of line: pow / log2 in float64 are long
// instruction sequences that the default (non-Shannon) configuration never executes
__device__ __noinline__ void shannon_rates(double u_master, double u_worker, double u_dist, int W, double& down, double& up) {
    const double pm = 50.0 + (60.0 - 50.0) * u_master;
    const double pw = 10.0 + (20.0 - 10.0) * u_worker;
    const double dist = 10.0 + (100.0 - 10.0) * u_dist;
    const double d4 = pow(dist, -4.0);
    const double bw = B_TOTAL / (double)W;
    down = bw * log2(1.0 + (pm * d4) / NOISE);
    up = bw * log2(1.0 + (pw * d4) / NOISE);
}
struct RetryStream {
    uint32_t env, step, tag, thr;
    uint64_t seed;
    uint4 b0, b1;
    int k, blk;
    bool have1;
    __device__ void init(uint32_t env_, uint32_t step_, int w, uint64_t seed_, double Pr) {
        env = env_; step = step_; tag = rng_tag(S_RETRY, (uint32_t)w); seed = seed_; k = 0; blk = 0; have1 = false;
        thr = (uint32_t)ceil(Pr * 16777216.0);
        b0 = philox_nl(env, step, tag, 0u, seed);
    }
    __device__ void ensure() {
        if (!have1) { b1 = philox_nl(env, step, tag, (uint32_t)(blk + 1), seed); have1 = true; }
    }
    __device__ void advance() { ensure(); b0 = b1; ++blk; have1 = false; }
    // Consume draws up to and including the first one that is >= Pr (`while U < Pr: n++`, WORKER:54-56,100-102) and return
    // the number of draws that were below: a whole Philox block is compared at once and the first failure found with a
    // find-first-set, so a geometric retry costs one short trip per FOUR draws instead of a dependent chain per draw.
    // At most `room` retries are counted (the RETRY_CAP guard of the per-draw loop, which stops after the draw that
    // reaches the cap).
    __device__ int run(int room) {
        int c = 0;
        for (;;) {
            const uint32_t lt = (uint32_t)((b0.x >> 8) < thr) | ((uint32_t)((b0.y >> 8) < thr) << 1) |
                                ((uint32_t)((b0.z >> 8) < thr) << 2) | ((uint32_t)((b0.w >> 8) < thr) << 3);
            const int lane = k & 3;
            const uint32_t fail = (~lt >> lane) & (0xFu >> lane);       // bit i: draw lane + i is >= Pr
            if (fail) {
                const int f = __ffs(fail) - 1;
                if (c + f < room) {
                    c += f; k += f + 1;
                    if (lane + f == 3) advance();
                    return c;
                }
            }
            const int take = fail ? room - c : min(4 - lane, room - c); // successes taken from this block
            if (c + take >= room) {                                     // the cap: the draw that reached it was the last one consumed
                k += take;
                if (((k - 1) & 3) == 3 && take > 0) advance();
                return room;
            }
            c += take; k += take;
            advance();
        }
    }
};

struct ProcOut { double delay, price0; int tp0, nslots; };

// Worker.process, DCML_Worker_TIMESLOT_MultiProcess.py:46-112; lw = this worker's 20 local slots
__device__ ProcOut process(const double* lw, double r, double c, int arrive, RetryStream& rs, double DATA_RATE) {
    ProcOut o;
    double cost = ceil((9.0 * r - 3.0) * c) / FREQUENCY;                    // :49-50
    rs.ensure();
    int n = 1 + rs.run(RETRY_CAP - 1);                                      // :54-56
    const double t_down = ((ceil((r + 1.0) * c) * 1.0 * 4.0) / DATA_RATE + 0.001) * (double)n;   // :60
    double price = floor(t_down) * 0.1;                                     // :65
    const int slot = (int)floor(t_down + (double)arrive);                   // :66
    int tp = slot % PERIOD;                                                 // :67-69
    int finish = slot;
    double avail = 0.0;
    const double frac = t_down - floor(t_down);                             // t_down % 1 (exact: t_down > 0)
    if (frac > lw[tp]) cost += frac - lw[tp];                               // :85-86
    o.price0 = price; o.tp0 = tp;
    int nslots = 0;
    double t_up = 0.0;
    while (avail < cost && nslots < SLOT_CAP) {                             // :87-106
        const double fr = 1.0 - lw[tp];
        price += fr * 1.0;
        avail += fr;
        ++nslots; ++finish; ++tp;
        if (tp > PERIOD - 1) tp -= PERIOD;
        rs.ensure();
        if (n < RETRY_CAP) n += rs.run(RETRY_CAP - n);
        else rs.run(0);
        t_up = ((ceil(r) * 1.0 * 4.0) / DATA_RATE + 0.001) * (double)n + 0.02;
    }
    o.nslots = nslots;
    o.delay = (double)(finish - arrive) - (avail - cost) + t_up;            // :108
    return o;
}

// cumulative price after m compute slots (the reference's prices[m-1]), same addition order
__device__ double price_after(const double* lw, const ProcOut& o, int m) {
    double price = o.price0;
    int tp = o.tp0;
    for (int s = 0; s < m; ++s) {
        price += (1.0 - lw[tp]) * 1.0;
        if (++tp > PERIOD - 1) tp -= PERIOD;
    }
    return price;
}

__global__ void env_kernel(EnvState st, EnvIO io, int what, uint32_t t, long env_first, long env_end) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int W = st.W, L = W + 1, tpe = st.tpe;
    const int epb = blockDim.x / tpe;
    const int le = threadIdx.x / tpe;            // local env
    const int w = threadIdx.x - le * tpe;        // worker (>= W: idle lane)
    const long env = env_first + (long)blockIdx.x * epb + le;
    const bool env_ok = env < env_end;
    const bool active = env_ok && w < W;
    const uint32_t genv = (uint32_t)(st.env_offset + env);

    // shared carve-up (per block)
    double* s_lw = reinterpret_cast<double*>(smem_raw);                 // [threads][LWS] old workload
    double* s_d = s_lw + (size_t)blockDim.x * LWS;                      // [threads] delay, later price / Pr
    double* s_f = s_d + blockDim.x;                                     // [threads][3] observed slots
    double* s_rate = s_f + (size_t)blockDim.x * 3;                      // [2][threads] Shannon down / up rates
    double* s_sc = s_rate + (size_t)blockDim.x * 2;                     // [epb][8] scalars
    uint32_t* s_key = reinterpret_cast<uint32_t*>(s_sc + (size_t)epb * 8);   // [threads]
    float* s_sel = reinterpret_cast<float*>(s_key + blockDim.x);        // [threads]
    int* s_i = reinterpret_cast<int*>(s_sel + blockDim.x);              // [epb][8] int scalars
    uint8_t* s_dis = reinterpret_cast<uint8_t*>(s_i + (size_t)epb * 8); // [threads]

    // reductions over the threads of one env (tpe / 32 whole warps): warp shuffles, then the env's warp partials through
    // shared memory; every thread of the CTA takes part (two CTA barriers)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwe = tpe >> 5, w0e = le * nwe;
    double* s_red = reinterpret_cast<double*>((reinterpret_cast<uintptr_t>(s_dis + blockDim.x) + 15) & ~(uintptr_t)15);   // [warps][4]
    auto env_sum4 = [&](double (&v)[4]) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
            for (int o = 16; o; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
        if (lane == 0) { s_red[warp * 4] = v[0]; s_red[warp * 4 + 1] = v[1]; s_red[warp * 4 + 2] = v[2]; s_red[warp * 4 + 3] = v[3]; }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i] = 0.0;
        for (int j = 0; j < nwe; ++j) { v[0] += s_red[(w0e + j) * 4]; v[1] += s_red[(w0e + j) * 4 + 1]; v[2] += s_red[(w0e + j) * 4 + 2]; v[3] += s_red[(w0e + j) * 4 + 3]; }
        __syncthreads();
    };
    // count of `pred` over the env, and the count over the env's threads BEFORE this one
    auto env_count = [&](bool pred, int& before) -> int {
        const unsigned m = __ballot_sync(0xffffffffu, pred);
        int* s_cnt = reinterpret_cast<int*>(s_red);
        if (lane == 0) s_cnt[warp] = __popc(m);
        __syncthreads();
        int tot = 0; before = __popc(m & ((1u << lane) - 1u));
        for (int j = 0; j < nwe; ++j) { const int c = s_cnt[w0e + j]; tot += c; if (w0e + j < warp) before += c; }
        __syncthreads();
        return tot;
    };
    double* my_lw = s_lw + (size_t)threadIdx.x * LWS;
    double* sc = s_sc + le * 8;     // 0 final delay, 1 payment, 2 R, 3 C
    int* si = s_i + le * 8;         // 0 N, 1 K, 2 n0 flag, 3 arrive, 4 kdis, 5 done
    double* e_d = s_d + le * tpe;
    float* e_sel = s_sel + le * tpe;
    uint32_t* e_key = s_key + le * tpe;
    uint8_t* e_dis = s_dis + le * tpe;
    double* e_rd = s_rate + le * tpe;
    double* e_ru = s_rate + blockDim.x + le * tpe;
    // workload table of the env, slot-pair major: double2 (slots 2i, 2i+1) of worker w at [i][w], so that the lanes of a
    // warp read / write consecutive 16-byte words (coalesced 128-bit accesses) instead of 160-byte-strided rows
    double2* glw2 = reinterpret_cast<double2*>(st.lw) + env * (long)(PERIOD / 2) * W + w;
    auto lw_at = [&](int slot) -> double {
        const double2 v = glw2[(long)(slot >> 1) * W];
        return (slot & 1) ? v.y : v.x;
    };

    if (what == ENV_INIT) {
        // Worker.__init__ :39  all_workload = clip(base * U(0.8,1.2), 0, 1), step-0 jitter stream
        if (active) {
#pragma unroll 1
            for (int blk = 0; blk < 5; ++blk) {
                const uint4 x = philox_nl(genv, 0u, rng_tag(S_JITTER, (uint32_t)w), (uint32_t)blk, st.seed);
                double nv[4];
                for (int i = 0; i < 4; ++i) {
                    const double v = st.base[w * PERIOD + blk * 4 + i] * (0.8 + (1.2 - 0.8) * u01d(lane_of(x, i)));
                    nv[i] = fmin(fmax(v, 0.0), 1.0);
                }
                glw2[(long)(blk * 2) * W] = make_double2(nv[0], nv[1]);
                glw2[(long)(blk * 2 + 1) * W] = make_double2(nv[2], nv[3]);
            }
        }
    }

    if (what == ENV_STEP) {
        // ---- Env.step :65-74, 96-105 ---------------------------------------------------------
        if (active) {
            e_sel[w] = st.fixed ? (st.disabled[env * W + w] ? 0.f : 1.f) : io.actions[env * L + w];   // :58-59 vs :73
            for (int i = 0; i < PERIOD; i += 2) {
                const double2 v = glw2[(long)(i >> 1) * W];
                my_lw[i] = v.x; my_lw[i + 1] = v.y;
            }
        }
        // N = int(sum(action[:-1])) (:71).  The reference adds fp32 values in order; the policy's selections are exactly 0 / 1,
        // for which a count is that sum -- anything else takes the in-order sum by one thread
        const float selv = active ? e_sel[w] : 0.f;
        const bool odd = selv != 0.f && selv != 1.f;
        int before_;
        int N = env_count(selv == 1.f, before_);
        if (__syncthreads_or(odd)) {
            if (env_ok && w == 0) {
                float sum = 0.f;
                for (int i = 0; i < W; ++i) sum = __fadd_rn(sum, e_sel[i]);
                si[0] = (int)sum;
            }
            __syncthreads();
            if (env_ok) N = si[0];
        }
        const double Rv = env_ok ? st.R[env] : 1.0, Cv = env_ok ? st.C[env] : 1.0;
        const int arr = env_ok ? st.arrive[env] : 0;
        int K;
        bool n0 = false;
        {
            const float ratio = env_ok ? io.actions[env * L + W] : 0.f;
            K = st.fixed ? (int)floor((double)N * 0.7)                     // :61
                         : (int)ceilf(__fmul_rn((float)N, ratio));         // :72 (fp32 product of fp32 actions)
            n0 = env_ok && (N == 0 || st.standalone);                      // :81
            N = min(max(N, 1), W);
            K = min(max(K, 1), N);
            if (n0) K = 1;                                                 // :82 update_strategy(K=1,N=1)
        }
        if (env_ok && w == 0) sc[0] = 0.0;
        ProcOut po;
        po.delay = 0.0; po.price0 = 0.0; po.tp0 = 0; po.nslots = 0;
        const bool runs = active && (!n0 || w == 0);                       // :83 vs :106-111
        if (runs) {
            RetryStream rs;
            rs.init(genv, t, w, st.seed, st.Pr[env * W + w]);
            const double r = ceil(Rv / (double)K);                         // Master.get_workload, DCML_Master.py:39-40
            po = process(my_lw, r, Cv, arr, rs, st.shannon ? st.rate_down[env * W + w] : DATA_RATE);
            e_d[w] = po.delay;
            // Worker.process :111 redraw the workload (step-t jitter stream)
#pragma unroll 1
            for (int blk = 0; blk < 5; ++blk) {
                const uint4 x = philox_nl(genv, t, rng_tag(S_JITTER, (uint32_t)w), (uint32_t)blk, st.seed);
                double nv[4];
                for (int i = 0; i < 4; ++i) {
                    const double v = st.base[w * PERIOD + blk * 4 + i] * (0.8 + (1.2 - 0.8) * u01d(lane_of(x, i)));
                    nv[i] = fmin(fmax(v, 0.0), 1.0);
                }
                glw2[(long)(blk * 2) * W] = make_double2(nv[0], nv[1]);
                glw2[(long)(blk * 2 + 1) * W] = make_double2(nv[2], nv[3]);
            }
        }
        __syncthreads();
        // ---- k-of-n completion order statistic :128-129 -------------------------------------
        // final_delay = K-th smallest delay among the selected workers: every thread ranks its own delay (unselected
        // workers carry +inf) -- against its own warp by register shuffles, against the env's other warps through the
        // delays in shared memory -- and the thread of rank K-1 publishes it.  Ties break by worker index (a stable sort).
        {
            const bool selected = active && !n0 && e_sel[w] != 0.f;
            const double d = selected ? po.delay : INFINITY;
            if (active) e_d[w] = d;                         // other warps read the padded key
            int rank = 0;
#pragma unroll 8
            for (int j = 0; j < 32; ++j) {
                const double dj = __shfl_sync(0xffffffffu, d, j);
                rank += (dj < d) || (dj == d && j < lane);
            }
            __syncthreads();
            const int wi = w >> 5;
            for (int ow = 0; ow < nwe; ++ow) {
                if (ow == wi) continue;
                const int lim = min(32, W - ow * 32);
                for (int j = 0; j < lim; ++j) {
                    const double dj = e_d[ow * 32 + j];
                    rank += (dj < d) || (dj == d && ow < wi);
                }
            }
            if (env_ok && n0 && w == 0) sc[0] = po.delay;
            if (selected && rank == K - 1) sc[0] = d;
        }
        __syncthreads();
        // ---- payment :130-137 ----------------------------------------------------------------
        double red[4] = {0.0, 0.0, 0.0, 0.0};
        if (runs) {
            const int e = (int)ceil(sc[0]);
            const int m = n0 ? po.nslots : (po.nslots >= e ? e : po.nslots);
            const double price = price_after(my_lw, po, m);
            red[0] = n0 ? price : (double)e_sel[w] * price;
        }
        env_sum4(red);                                                     // the reference adds in worker order: a tree here, fp32 out
        if (env_ok && w == 0) {
            const double pay = red[0];
            const double fd = sc[0];
            double rew = -(fd * 99.0) - (pay * 1.0);                       // DCML_ENV_Functions.py:15-17
            if (n0) rew = 1.5 * rew;                                       // :88
            io.reward[env] = (float)rew;
            io.delay[env] = (float)fd;
            io.payment[env] = (float)pay;
            const uint4 x = philox_nl(genv, t, rng_tag(S_RESET, 0), 1u, st.seed);
            const int dn = u01d(x.y) < DONE_P;                             // :142
            si[5] = dn;
            io.done[env] = (uint8_t)dn;
        }
        __syncthreads();
    }

    // ---- Env.reset :157-199 (twice when done under ShareDummyVecEnv) ---------------------------
    if (what != ENV_OBSERVE) {
#pragma unroll 1
        for (int r = 0; r < 2; ++r) {
            const bool apply = env_ok && (r == 0 || (what == ENV_STEP && st.double_reset && si[5]));
            if (apply && w == 0) {
                const uint4 x = philox_nl(genv, t, rng_tag(S_RESET, 0), (uint32_t)(16 * r), st.seed);
                const int kmax = (int)floor(0.8 * (double)W);
                si[4] = 1 + (int)bounded(x.x, (uint32_t)kmax);                              // :158
                si[3] = (int)bounded(x.y, PERIOD);                                           // :160
                sc[2] = (double)(R_MIN + (int)bounded(x.z, (uint32_t)(R_HI - R_MIN + 1)));   // DCML_Master.py:47
                sc[3] = (double)(C_MIN + (int)bounded(x.w, (uint32_t)(C_HI - C_MIN + 1)));   // DCML_Master.py:48
            }
            if (apply && w < W) {
                const uint4 x = philox_nl(genv, t, rng_tag(S_WORKER, (uint32_t)w), (uint32_t)r, st.seed);
                e_d[w] = 0.0 + (PR_MAX - 0.0) * u01d(x.x);                                   // :197
                e_key[w] = x.y;
                if (st.shannon) {
                    // update_workers_transmission(True): DCML_Basic_Env.py:18-27, DCML_Master.py:41-45, Shannon.py:14-21
                    const uint4 xm = philox_nl(genv, t, rng_tag(S_RESET, 0), (uint32_t)(16 * r + 2), st.seed);
                    shannon_rates(u01d(xm.x), u01d(x.z), u01d(x.w), W, e_rd[w], e_ru[w]);
                }
            }
            __syncthreads();
            if (apply && w < W) {
                // :199 choice(W, k, replace=False): the k smallest (key, index) pairs are disabled
                const uint32_t key = e_key[w];
                int rank = 0;
                for (int j = 0; j < W; ++j) rank += (e_key[j] < key) || (e_key[j] == key && j < w);
                e_dis[w] = rank < si[4];
            }
            __syncthreads();
        }
        if (active) {
            st.Pr[env * W + w] = e_d[w];
            st.disabled[env * W + w] = e_dis[w];
            if (st.shannon) { st.rate_down[env * W + w] = e_rd[w]; st.rate_up[env * W + w] = e_ru[w]; }
        }
        if (env_ok && w == 0) {
            st.R[env] = sc[2]; st.C[env] = sc[3]; st.arrive[env] = si[3]; st.kdis[env] = si[4];
        }
    } else {
        if (active) {
            e_d[w] = st.Pr[env * W + w]; e_dis[w] = st.disabled[env * W + w];
            if (st.shannon) { e_rd[w] = st.rate_down[env * W + w]; e_ru[w] = st.rate_up[env * W + w]; }
        }
        if (env_ok && w == 0) { sc[2] = st.R[env]; sc[3] = st.C[env]; si[3] = st.arrive[env]; si[4] = st.kdis[env]; }
        __syncthreads();
    }

    // ---- observation :180-271 -------------------------------------------------------------------
    const double Rn = env_ok ? (sc[2] - (double)R_MIN) / (double)(R_MAX - R_MIN) : 0.0;   // DCML_Master.py:54
    const double Cn = env_ok ? (sc[3] - (double)C_MIN) / (double)(C_MAX - C_MIN) : 0.0;
    // reset(binary=True) :164-172: share_obs starts with R and C as 32-bit strings (most significant bit first) and the
    // observation rows carry its first two entries -- the two top bits of R -- where Rn, Cn would be (:203 shared_obs[:2])
    const uint32_t Rb = env_ok ? (uint32_t)sc[2] : 0u, Cb = env_ok ? (uint32_t)sc[3] : 0u;
    const float o0 = io.binary ? (float)(Rb >> 31) : (float)Rn, o1 = io.binary ? (float)((Rb >> 30) & 1u) : (float)Cn;
    const long sw = (st.shannon ? 2 * W + 2 : W + 2) + (io.binary ? 62 : 0);    // floats per share_obs row
    const int s0 = io.binary ? 64 : 2;                                          // first per-worker column
    double* my_f = s_f + (size_t)threadIdx.x * 3;
    int en_before = 0;
    env_count(active && !e_dis[w], en_before);                             // enabled workers with a smaller index (:226 i - disabled)
    if (active) {
        const int a = si[3];
        const bool dis = e_dis[w];
        const int n_en = W - si[4];
        float* o = io.obs + (env * L + w) * 7;
        o[0] = o0; o[1] = o1;
        if (dis) {
            // :207-213: ones, then the previous row's 7th feature (0 for the first row)
            const double f7 = en_before > 0 ? (double)(en_before - 1) / (double)n_en : 0.0;
            o[2] = 1.f; o[3] = 1.f; o[4] = 1.f; o[5] = 1.f; o[6] = (float)f7;
            my_f[0] = 0.0; my_f[1] = 0.0; my_f[2] = 0.0;
        } else {
            for (int s = 0; s < 3; ++s) {                                  // :222
                const double v = lw_at((a + s) % PERIOD);
                my_f[s] = v;
                o[2 + s] = (float)v;
            }
            o[5] = (float)e_d[w];
            o[6] = (float)((double)en_before / (double)n_en);              // :226
        }
        io.ava[(env * L + w) * 2] = 1.f;                                   // Worker.get_availability :144-148
        io.ava[(env * L + w) * 2 + 1] = dis ? 0.f : 1.f;
        if (st.shannon) {                                                  // :248-250
            io.sob[env * sw + s0 + w] = (float)(e_ru[w] / 1e7);
            io.sob[env * sw + s0 + W + w] = (float)(e_rd[w] / 1e7);
        } else {
            io.sob[env * sw + s0 + w] = (float)e_d[w];                     // :252-253
        }
        if (io.binary && w < 64) {
            const int bit = 31 - (w & 31);
            io.sob[env * sw + w] = (float)(((w < 32 ? Rb : Cb) >> bit) & 1u);
        }
    }
    double mean[4] = {0.0, 0.0, 0.0, 0.0};                                 // :235-239 sums over the enabled workers
    if (active && !e_dis[w]) { mean[0] = my_f[0]; mean[1] = my_f[1]; mean[2] = my_f[2]; mean[3] = e_d[w]; }
    env_sum4(mean);
    if (env_ok && w == 0) {
        const double m0 = mean[0], m1 = mean[1], m2 = mean[2], sp = mean[3];
        const double n_en = (double)(W - si[4]);
        float* o = io.obs + (env * L + W) * 7;                             // :235-239
        o[0] = o0; o[1] = o1;
        o[2] = (float)(m0 / n_en); o[3] = (float)(m1 / n_en); o[4] = (float)(m2 / n_en);
        o[5] = (float)(sp / n_en); o[6] = 1.1f;
        io.ava[(env * L + W) * 2] = 1.f; io.ava[(env * L + W) * 2 + 1] = 1.f;
        if (!io.binary) { io.sob[env * sw] = (float)Rn; io.sob[env * sw + 1] = (float)Cn; }
        else for (int b = W; b < 64; ++b) io.sob[env * sw + b] = (float)(((b < 32 ? Rb : Cb) >> (31 - (b & 31))) & 1u);   // W < 64
    }
}

__global__ void set_context_kernel(EnvState st, const double* R, const double* C, const double* pr, const int32_t* nd) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < st.B) {
        if (R) st.R[i] = R[i];
        if (C) st.C[i] = C[i];
        if (nd) st.kdis[i] = nd[i];
    }
    if (pr && i < st.B * st.W) st.Pr[i] = pr[i];
}

// preset disable counts: re-rank the stored keys is not possible (keys are not kept), so the
// disabled set for a new count is re-drawn from the WORKER stream of the env's current step.
__global__ void redraw_disabled_kernel(EnvState st, uint32_t t) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= st.B * st.W) return;
    const long env = i / st.W;
    const int w = (int)(i - env * st.W);
    const uint32_t genv = (uint32_t)(st.env_offset + env);
    const uint32_t key = philox_nl(genv, t, rng_tag(S_WORKER, (uint32_t)w), 0u, st.seed).y;
    int rank = 0;
    for (int j = 0; j < st.W; ++j) {
        const uint32_t kj = philox_nl(genv, t, rng_tag(S_WORKER, (uint32_t)j), 0u, st.seed).y;
        rank += (kj < key) || (kj == key && j < w);
    }
    st.disabled[i] = rank < st.kdis[env];
}

}  // namespace

struct dcml_env {
    dcml_env_cfg cfg;
    int device;
    EnvState st;
    double* d_base = nullptr;
    void* d_state = nullptr;
    uint32_t t = 0;
    int block, epb;
    size_t smem;
    bool reset_done = false;
};

static int env_launch(dcml_env* e, const EnvIO& io, int what, cudaStream_t stream, long env0 = 0, long n_env = -1) {
    if (n_env < 0) n_env = e->cfg.n_envs;
    const unsigned grid = (unsigned)((n_env + e->epb - 1) / e->epb);
    env_kernel<<<grid, e->block, e->smem, stream>>>(e->st, io, what, e->t, env0, env0 + n_env);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}

static int env_check(dcml_env* e) {
    ARG_CHECK(e != nullptr, "null env handle");
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    ARG_CHECK(dev == e->device, "env handle belongs to device %d but device %d is current", e->device, dev);
    return MATAS_OK;
}

extern "C" int dcml_env_create(const dcml_env_cfg* cfg, const double* base_host, dcml_env** out) {
    ARG_CHECK(cfg && base_host && out, "null argument");
    ARG_CHECK(cfg->n_envs > 0, "n_envs %ld <= 0", cfg->n_envs);
    ARG_CHECK(cfg->n_workers >= 2 && cfg->n_workers <= 1024, "n_workers %d not in [2,1024]", cfg->n_workers);
    dcml_env* e = new dcml_env();
    e->cfg = *cfg;
    const int W = cfg->n_workers;
    const long B = cfg->n_envs;
    if (cudaGetDevice(&e->device) != cudaSuccess) {
        matas_set_error("no usable CUDA device");
        delete e;
        return MATAS_ERR_CUDA;
    }
    const int tpe = (W + 31) / 32 * 32;
    // envs per CTA: the CTA barriers of a step wait for the slowest worker of every env in the block (the slot loops of
    // Worker.process have data-dependent trip counts), so small blocks lose less to that imbalance (MATAS_ENV_THREADS tunes)
    static int target = -1;
    if (target < 0) { const char* ev = getenv("MATAS_ENV_THREADS"); target = ev ? atoi(ev) : 64; if (target < 32) target = 32; }
    e->epb = tpe >= target ? 1 : target / tpe;
    e->block = e->epb * tpe;
    e->smem = (size_t)e->block * (LWS + 1 + 3 + 2) * 8 + (size_t)e->epb * 8 * 8 + (size_t)e->block * (4 + 4) +
              (size_t)e->epb * 8 * 4 + e->block + 16 + (size_t)(e->block / 32) * 32 + 32;
    size_t off = 0;
    auto carve = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t o_lw = carve((size_t)B * W * PERIOD * 8), o_R = carve(B * 8), o_C = carve(B * 8), o_Pr = carve((size_t)B * W * 8),
                 o_arr = carve(B * 4), o_k = carve(B * 4), o_dis = carve((size_t)B * W),
                 o_rd = carve((size_t)B * W * 8), o_ru = carve((size_t)B * W * 8);
    auto fail = [&](const char* what, cudaError_t err) {
        matas_set_error("%s: %s", what, cudaGetErrorString(err));
        dcml_env_destroy(e);
        return MATAS_ERR_CUDA;
    };
    cudaError_t err;
    if ((err = cudaMalloc(&e->d_state, off)) != cudaSuccess) return fail("cudaMalloc(env state)", err);
    if ((err = cudaMemset(e->d_state, 0, off)) != cudaSuccess) return fail("cudaMemset(env state)", err);
    if ((err = cudaMalloc(&e->d_base, (size_t)W * PERIOD * 8)) != cudaSuccess) return fail("cudaMalloc(base)", err);
    if ((err = cudaMemcpy(e->d_base, base_host, (size_t)W * PERIOD * 8, cudaMemcpyHostToDevice)) != cudaSuccess)
        return fail("cudaMemcpy(base)", err);
    if ((err = cudaFuncSetAttribute(env_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)e->smem)) != cudaSuccess)
        return fail("cudaFuncSetAttribute(env_kernel)", err);
    unsigned char* s = static_cast<unsigned char*>(e->d_state);
    EnvState& st = e->st;
    st.B = B; st.W = W; st.tpe = tpe; st.seed = cfg->seed; st.env_offset = cfg->env_offset; st.double_reset = cfg->double_reset;
    st.base = e->d_base;
    st.lw = reinterpret_cast<double*>(s + o_lw); st.R = reinterpret_cast<double*>(s + o_R); st.C = reinterpret_cast<double*>(s + o_C);
    st.Pr = reinterpret_cast<double*>(s + o_Pr); st.arrive = reinterpret_cast<int*>(s + o_arr); st.kdis = reinterpret_cast<int*>(s + o_k);
    st.disabled = s + o_dis;
    st.shannon = cfg->shannon ? 1 : 0;
    st.fixed = cfg->fixed ? 1 : 0; st.standalone = cfg->standalone ? 1 : 0;
    st.rate_down = reinterpret_cast<double*>(s + o_rd); st.rate_up = reinterpret_cast<double*>(s + o_ru);
    *out = e;
    return MATAS_OK;
}

extern "C" int dcml_env_destroy(dcml_env* e) {
    if (!e) return MATAS_OK;
    cudaFree(e->d_state); cudaFree(e->d_base);
    delete e;
    return MATAS_OK;
}

static int env_reset(dcml_env* e, float* obs, float* share_obs, float* ava, int binary, void* stream) {
    int rc = env_check(e);
    if (rc) return rc;
    ARG_CHECK(obs && share_obs && ava, "null output");
    EnvIO io = {};
    io.obs = obs; io.sob = share_obs; io.ava = ava; io.binary = binary;
    // The first reset builds the workers (Worker.__init__: step-0 jitter stream).  A later reset is the reference's
    // env.reset() on a live env: a fresh context from the next step's named draws, the carried workload tables untouched --
    // it never rewinds the step counter, so no context is ever replayed.
    const int what = e->reset_done ? ENV_RESET : ENV_INIT;
    if (e->reset_done) e->t += 1; else e->t = 0;
    e->reset_done = true;
    return env_launch(e, io, what, (cudaStream_t)stream);
}

extern "C" int dcml_env_reset(dcml_env* e, float* obs, float* share_obs, float* ava, void* stream) {
    return env_reset(e, obs, share_obs, ava, 0, stream);
}

extern "C" int dcml_env_reset_binary(dcml_env* e, float* obs, float* share_obs_bits, float* ava, void* stream) {
    return env_reset(e, obs, share_obs_bits, ava, 1, stream);
}

extern "C" int dcml_env_step(dcml_env* e, const float* actions, float* obs, float* share_obs, float* ava, float* reward,
                             uint8_t* done, float* delay, float* payment, void* stream) {
    int rc = env_check(e);
    if (rc) return rc;
    ARG_CHECK(e->reset_done, "dcml_env_step before dcml_env_reset");
    ARG_CHECK(actions && obs && share_obs && ava && reward && done && delay && payment, "null argument");
    EnvIO io = {};
    io.actions = actions; io.obs = obs; io.sob = share_obs; io.ava = ava; io.reward = reward; io.done = done;
    io.delay = delay; io.payment = payment;
    e->t += 1;
    return env_launch(e, io, ENV_STEP, (cudaStream_t)stream);
}

extern "C" int dcml_env_step_slice(dcml_env* e, const float* actions, long env0, long n_env, int first, float* obs, float* share_obs,
                                   float* ava, float* reward, uint8_t* done, float* delay, float* payment, void* stream) {
    int rc = env_check(e);
    if (rc) return rc;
    ARG_CHECK(e->reset_done, "dcml_env_step_slice before dcml_env_reset");
    ARG_CHECK(actions && obs && share_obs && ava && reward && done && delay && payment, "null argument");
    ARG_CHECK(env0 >= 0 && n_env > 0 && env0 + n_env <= e->cfg.n_envs, "bad slice (env0=%ld n_env=%ld of %ld)", env0, n_env, e->cfg.n_envs);
    EnvIO io = {};
    io.actions = actions; io.obs = obs; io.sob = share_obs; io.ava = ava; io.reward = reward; io.done = done;
    io.delay = delay; io.payment = payment;
    if (first) e->t += 1;
    return env_launch(e, io, ENV_STEP, (cudaStream_t)stream, env0, n_env);
}

extern "C" int dcml_env_set_context(dcml_env* e, const double* R, const double* C, const double* worker_pr,
                                    const int32_t* n_disabled, float* obs, float* share_obs, float* ava, void* stream) {
    int rc = env_check(e);
    if (rc) return rc;
    ARG_CHECK(e->reset_done, "dcml_env_set_context before dcml_env_reset");
    ARG_CHECK(obs && share_obs && ava, "null output");
    const long n = e->cfg.n_envs * e->cfg.n_workers;
    set_context_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(e->st, R, C, worker_pr, n_disabled);
    if (n_disabled) redraw_disabled_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(e->st, e->t);
    CUDA_TRY(cudaGetLastError());
    EnvIO io = {};
    io.obs = obs; io.sob = share_obs; io.ava = ava;
    return env_launch(e, io, ENV_OBSERVE, (cudaStream_t)stream);
}
