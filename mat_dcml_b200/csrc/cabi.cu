// extern "C" entry points of libmatas_b200.so (declared in include/matas_b200.h): policy handle,
// weight repacking, scratch management.  The DCML simulator and GAE entry points live in env.cu
// and gae.cu.
#include <cuda_fp16.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/matas_b200.h"
#include "layout.h"
#include "policy_bwd.h"
#include "policy_fwd.h"

static thread_local char g_err[512] = "";

void matas_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* matas_last_error(void) { return g_err; }
extern "C" int matas_version(void) { return 100; }

// one repack instruction: copy a vector (cols == 0) or transpose W[rows][cols] -> dst[c*64 + r]
struct PackEntry { int src, dst, rows, cols; };

__global__ void pack_kernel(const float* __restrict__ flat, float* __restrict__ packed, const PackEntry* __restrict__ tab) {
    const PackEntry e = tab[blockIdx.x];
    if (e.cols == 0) {
        for (int i = threadIdx.x; i < e.rows; i += blockDim.x) packed[e.dst + i] = flat[e.src + i];
    } else {
        const int n = e.rows * e.cols;
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const int r = i / e.cols, c = i - r * e.cols;
            packed[e.dst + c * 64 + r] = flat[e.src + i];
        }
    }
}

// one folded 64x64 map of the absorbed decode attention (layout.h): built from two Linear layers
struct FoldEntry {
    int kind;                // 0: query map  Wt[j][i] = scale * sum_d Wa[32h+d][i] Wb[32h+d][j], bias_i = scale * sum_d Wa[32h+d][i] bb[32h+d]
                             //    (Wa = key, Wb = query);  1: output map Wt[i][o] = sum_d Wa[o][32h+d] Wb[32h+d][i]  (Wa = proj, Wb = value)
    int h;                   // head
    int wa, wb, bb;          // flat-parameter offsets; kind 1: bb = value bias
    int bp;                  // kind 1, h == 1: proj bias (the constant Wp bv + bp goes into this blob), else -1
    int lnw, lnb;            // kind 1, h == 1: LayerNorm applied after the attention block, else -1
    int dst;                 // packed offset of the blob
};

__global__ void fold_kernel(const float* __restrict__ flat, float* __restrict__ packed, const FoldEntry* __restrict__ tab) {
    const FoldEntry e = tab[blockIdx.x];
    const float scale = 0.17677669529663687f * 1.4426950408889634f;     // log2(e)/sqrt(32): the decode softmax uses exp2
    float* blob = packed + e.dst;
    for (int idx = threadIdx.x; idx < 4096; idx += blockDim.x) {
        const int k = idx >> 6, n = idx & 63;        // Wt[k][n]
        float acc = 0.f;
        if (e.kind == 0) {
            for (int d = 0; d < 32; ++d) acc += flat[e.wa + (32 * e.h + d) * 64 + n] * flat[e.wb + (32 * e.h + d) * 64 + k];
            acc *= scale;
        } else {
            for (int d = 0; d < 32; ++d) acc += flat[e.wa + n * 64 + 32 * e.h + d] * flat[e.wb + (32 * e.h + d) * 64 + k];
        }
        blob[idx] = acc;
    }
    for (int n = threadIdx.x; n < 64; n += blockDim.x) {
        float bias = 0.f;
        if (e.kind == 0) {
            for (int d = 0; d < 32; ++d) bias += flat[e.wa + (32 * e.h + d) * 64 + n] * flat[e.bb + 32 * e.h + d];
            bias *= scale;
        } else if (e.bp >= 0) {
            for (int d = 0; d < 64; ++d) bias += flat[e.wa + n * 64 + d] * flat[e.bb + d];      // Wp bv over both heads
            bias += flat[e.bp + n];
        }
        blob[lay::B_BIAS + n] = bias;
        blob[lay::B_LNW + n] = e.lnw >= 0 ? flat[e.lnw + n] : 1.f;
        blob[lay::B_LNB + n] = e.lnb >= 0 ? flat[e.lnb + n] : 0.f;
    }
}

// tensor-core copy of one forward / folded blob (layout.h: hb()): fp16 hi + scaled-lo parts of W[n][k],
// K-major with the 128-byte swizzle tcgen05 shared-memory descriptors expect, plus the fp32 tail.
__global__ void half_kernel(float* __restrict__ packed, const int* __restrict__ offs) {
    const int off = offs[blockIdx.x];
    const float* src = packed + off;                                  // Wt[k][n]
    unsigned char* dst = reinterpret_cast<unsigned char*>(packed + lay::hb(off));
    for (int idx = threadIdx.x; idx < 4096; idx += blockDim.x) {
        const int n = idx >> 6, k = idx & 63;
        const float w = src[k * 64 + n];
        const __half hi = __float2half_rn(w);
        const __half lo = __float2half_rn((w - __half2float(hi)) * 2048.0f);
        const int byte = n * 128 + ((((k >> 3) ^ (n & 7)) << 4) | ((k & 7) * 2));
        *reinterpret_cast<__half*>(dst + byte) = hi;
        *reinterpret_cast<__half*>(dst + lay::HB_LO_BYTES + byte) = lo;
    }
    for (int i = threadIdx.x; i < 192; i += blockDim.x) packed[lay::hb(off) + lay::B_BIAS + i] = src[lay::B_BIAS + i];
}

struct matas_policy {
    matas_model_cfg cfg;
    int device;
    int n_sm;
    long n_params;
    std::vector<PackEntry> table;
    PackEntry* d_table = nullptr;
    std::vector<FoldEntry> folds;
    FoldEntry* d_folds = nullptr;
    int n_half = 0;
    int* d_half = nullptr;       // float offsets of the blobs that get a tensor-core copy
    float* d_packed = nullptr;
    bool weights_set = false;
    // scratch
    long tok_cap = 0;
    float* d_tok = nullptr;      // 21 token buffers [cap][64]
    size_t kv_cap = 0;
    float* d_kv = nullptr;
    float* d_x1tab = nullptr;    // [L (L + 1)][64] block-0 cross-attention inputs of the stochastic decode (decode_tc.cu)
    float* d_x1spec = nullptr;   // speculative rows of the deterministic schedule for `spec_stride` (decode_tc.cu)
    long spec_cap = 0;
    int spec_stride = -1;        // -1: the rows must be recomputed (weights changed)
    PolicyBufs bufs = {};
    // training: saved activations + backward work buffers, gradient offsets
    GradMap gm;
    TrainBufs tb;
    long train_cap = 0;
    float* d_train = nullptr;
    long train_tokens = 0;        // > 0: the saved activations belong to a training forward of that many tokens
    // optional per-launch timing of the decode kernel (CUDA events on the launching stream)
    static constexpr int NEV = 64;
    bool timing = false;
    cudaEvent_t ev0[NEV], ev1[NEV];
    bool ev_made = false, ev_pending[NEV];
    int ev_next = 0;
    double t_ms = 0.0;
    long t_n = 0;
};

static int harvest(matas_policy* p, int i) {
    if (p->ev_pending[i]) {
        float ms = 0.f;
        CUDA_TRY(cudaEventSynchronize(p->ev1[i]));
        CUDA_TRY(cudaEventElapsedTime(&ms, p->ev0[i], p->ev1[i]));
        p->t_ms += ms; p->t_n += 1; p->ev_pending[i] = false;
    }
    return MATAS_OK;
}

namespace {

// walks the parameters in the reference's state_dict order (ma_transformer.py:119-203)
struct LayoutBuilder {
    std::vector<PackEntry>& tab;
    GradMap& gm;
    int src = 0;
    LayoutBuilder(std::vector<PackEntry>& t, GradMap& g) : tab(t), gm(g) {
        for (int i = 0; i < lay::N_BLOBS; ++i) gm.w[i] = gm.b[i] = gm.lnw[i] = gm.lnb[i] = -1;
    }
    static int idx(int blob) { return (blob - lay::ENC_BLOBS) / lay::BLOB; }
    void vec(int n, int dst) { if (dst >= 0) tab.push_back({src, dst, n, 0}); src += n; }
    void mat(int rows, int cols, int dst) { if (dst >= 0) tab.push_back({src, dst, rows, cols}); src += rows * cols; }
    void linear(int out, int in, int dst_wt, int dst_b, bool bias = true) {
        mat(out, in, dst_wt);
        if (bias) vec(out, dst_b);
    }
    void blob_linear(int blob) {
        tab.push_back({src, lay::bw(blob), 4096, 0});      // untransposed copy for the backward pass
        gm.w[idx(blob)] = src; gm.b[idx(blob)] = src + 4096;
        linear(64, 64, blob, blob + lay::B_BIAS);
    }
    void blob_ln(int blob) {
        gm.lnw[idx(blob)] = src; gm.lnb[idx(blob)] = src + 64;
        vec(64, blob + lay::B_LNW); vec(64, blob + lay::B_LNB);
    }
};

long build_layout(const matas_model_cfg& c, std::vector<PackEntry>& tab, GradMap& gm) {
    LayoutBuilder b(tab, gm);
    using namespace lay;
    // encoder.state_encoder (unused when encode_state=False, ma_transformer.py:144-149)
    b.vec(c.state_dim, -1); b.vec(c.state_dim, -1); b.linear(64, c.state_dim, -1, -1);
    gm.enc_ln7_w = b.src; gm.enc_ln7_b = b.src + c.obs_dim;
    b.vec(c.obs_dim, ENC_EMB + ENC_EMB_LNW); b.vec(c.obs_dim, ENC_EMB + ENC_EMB_LNB);
    gm.enc_wobs = b.src; gm.enc_bobs = b.src + 64 * c.obs_dim;
    b.linear(64, c.obs_dim, ENC_EMB + ENC_EMB_WT, ENC_EMB + ENC_EMB_B);
    gm.enc_ln_w = b.src; gm.enc_ln_b = b.src + 64;
    b.vec(64, ENC_EMB + ENC_EMB_LN2W); b.vec(64, ENC_EMB + ENC_EMB_LN2B);
    for (int k = 0; k < 2; ++k) {
        b.blob_ln(enc_blob(k, EP));
        b.blob_ln(enc_blob(k, EM2));
        b.blob_linear(enc_blob(k, EK)); b.blob_linear(enc_blob(k, EQ)); b.blob_linear(enc_blob(k, EV)); b.blob_linear(enc_blob(k, EP));
        b.blob_linear(enc_blob(k, EM0)); b.blob_linear(enc_blob(k, EM2));
    }
    b.blob_linear(ENC_H0); b.blob_ln(ENC_H0);
    gm.enc_h3_w = b.src; gm.enc_h3_b = b.src + 64;
    b.vec(64, ENC_H3); b.vec(1, ENC_H3 + 64);
    gm.log_std = -1;
    if (c.mode == MATAS_MODE_SEMI) { gm.log_std = b.src; b.vec(c.act_dim, DEC_H3 + DEC_H3_LOGSTD); }
    gm.dec_wa = b.src;
    b.mat(64, c.act_dim + 1, DEC_EMB + DEC_EMB_WT);
    // decoder.obs_encoder (unused unless dec_actor, ma_transformer.py:198-199)
    b.vec(c.obs_dim, -1); b.vec(c.obs_dim, -1); b.linear(64, c.obs_dim, -1, -1);
    gm.dec_ln_w = b.src; gm.dec_ln_b = b.src + 64;
    b.vec(64, DEC_EMB + DEC_EMB_LNW); b.vec(64, DEC_EMB + DEC_EMB_LNB);
    for (int k = 0; k < 2; ++k) {
        b.blob_ln(dec_blob(k, DP1)); b.blob_ln(dec_blob(k, DP2)); b.blob_ln(dec_blob(k, DM2));
        b.blob_linear(dec_blob(k, DK1)); b.blob_linear(dec_blob(k, DQ1)); b.blob_linear(dec_blob(k, DV1)); b.blob_linear(dec_blob(k, DP1));
        b.blob_linear(dec_blob(k, DK2)); b.blob_linear(dec_q2(k)); b.blob_linear(dec_blob(k, DV2)); b.blob_linear(dec_blob(k, DP2));
        b.blob_linear(dec_blob(k, DM0)); b.blob_linear(dec_blob(k, DM2));
    }
    b.blob_linear(DEC_H0); b.blob_ln(DEC_H0);
    gm.dec_h3_w = b.src; gm.dec_h3_b = b.src + 128;
    b.vec(2 * 64, DEC_H3); b.vec(2, DEC_H3 + DEC_H3_B);
    return b.src;
}

int check_handle(const matas_policy* p) {
    ARG_CHECK(p != nullptr, "null policy handle");
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    ARG_CHECK(dev == p->device, "policy handle belongs to device %d but device %d is current", p->device, dev);
    ARG_CHECK(p->weights_set, "matas_policy_set_weights has not been called");
    return MATAS_OK;
}

}  // namespace

extern "C" int matas_policy_create(const matas_model_cfg* cfg, matas_policy** out) {
    ARG_CHECK(cfg && out, "null argument");
    if (cfg->n_embd != 64 || cfg->n_head != 2 || cfg->n_block != 2 || cfg->act_dim != 2) {
        matas_set_error("unsupported model: kernels are built for n_embd=64 n_head=2 n_block=2 act_dim=2 "
                        "(got %d %d %d %d)", cfg->n_embd, cfg->n_head, cfg->n_block, cfg->act_dim);
        return MATAS_ERR_UNSUPPORTED;
    }
    ARG_CHECK(cfg->obs_dim >= 1 && cfg->obs_dim <= 8, "obs_dim %d not in [1,8]", cfg->obs_dim);
    ARG_CHECK(cfg->n_agent >= 2, "n_agent %d < 2", cfg->n_agent);
    ARG_CHECK(cfg->state_dim >= 1, "state_dim %d < 1", cfg->state_dim);
    ARG_CHECK(cfg->mode == MATAS_MODE_SEMI || cfg->mode == MATAS_MODE_DISCRETE, "bad mode %d", cfg->mode);
    int rc = policy_check_agents(cfg->n_agent);
    if (rc) return rc;
    matas_policy* p = new matas_policy();
    p->cfg = *cfg;
    memset(&p->bufs, 0, sizeof(p->bufs));
    if (cudaGetDevice(&p->device) != cudaSuccess ||
        cudaDeviceGetAttribute(&p->n_sm, cudaDevAttrMultiProcessorCount, p->device) != cudaSuccess) {
        matas_set_error("no usable CUDA device: %s", cudaGetErrorString(cudaGetLastError()));
        delete p;
        return MATAS_ERR_CUDA;
    }
    memset(&p->tb, 0, sizeof(p->tb));
    p->n_params = build_layout(p->cfg, p->table, p->gm);
    {
        using namespace lay;
        auto ix = [](int blob) { return (blob - ENC_BLOBS) / BLOB; };
        const GradMap& g = p->gm;
        for (int k = 0; k < 2; ++k) {
            const int q1 = ix(dec_blob(k, DQ1)), k1 = ix(dec_blob(k, DK1)), v1 = ix(dec_blob(k, DV1)), p1 = ix(dec_blob(k, DP1));
            const int q2 = ix(dec_q2(k)), k2 = ix(dec_blob(k, DK2)), v2 = ix(dec_blob(k, DV2)), p2 = ix(dec_blob(k, DP2));
            for (int h = 0; h < 2; ++h) {
                p->folds.push_back({0, h, g.w[k1], g.w[q1], g.b[q1], -1, -1, -1, fold_blob(k, FA0 + h)});
                p->folds.push_back({1, h, g.w[p1], g.w[v1], g.b[v1], h ? g.b[p1] : -1, h ? g.lnw[p1] : -1, h ? g.lnb[p1] : -1, fold_blob(k, FB0 + h)});
                p->folds.push_back({1, h, g.w[p2], g.w[v2], g.b[v2], h ? g.b[p2] : -1, h ? g.lnw[p2] : -1, h ? g.lnb[p2] : -1, fold_blob(k, FC0 + h)});
                p->folds.push_back({0, h, g.w[k2], g.w[q2], g.b[q2], -1, -1, -1, fold_blob(k, FQ0 + h)});
            }
        }
    }
    auto fail = [&](cudaError_t e, const char* what) {
        matas_set_error("%s: %s", what, cudaGetErrorString(e));
        matas_policy_destroy(p);
        return MATAS_ERR_CUDA;
    };
    cudaError_t e;
    if ((e = cudaMalloc(&p->d_table, p->table.size() * sizeof(PackEntry))) != cudaSuccess) return fail(e, "cudaMalloc(table)");
    if ((e = cudaMemcpy(p->d_table, p->table.data(), p->table.size() * sizeof(PackEntry), cudaMemcpyHostToDevice)) != cudaSuccess)
        return fail(e, "cudaMemcpy(table)");
    if ((e = cudaMalloc(&p->d_folds, p->folds.size() * sizeof(FoldEntry))) != cudaSuccess) return fail(e, "cudaMalloc(folds)");
    if ((e = cudaMemcpy(p->d_folds, p->folds.data(), p->folds.size() * sizeof(FoldEntry), cudaMemcpyHostToDevice)) != cudaSuccess)
        return fail(e, "cudaMemcpy(folds)");
    {
        std::vector<int> offs;
        for (int o = lay::ENC_BLOBS; o < lay::TOTAL; o += lay::BLOB) offs.push_back(o);
        for (int o = lay::FOLD_BASE; o < lay::TOTAL_ALL; o += lay::BLOB) offs.push_back(o);
        p->n_half = (int)offs.size();
        if ((e = cudaMalloc(&p->d_half, offs.size() * sizeof(int))) != cudaSuccess) return fail(e, "cudaMalloc(half table)");
        if ((e = cudaMemcpy(p->d_half, offs.data(), offs.size() * sizeof(int), cudaMemcpyHostToDevice)) != cudaSuccess)
            return fail(e, "cudaMemcpy(half table)");
    }
    if ((e = cudaMalloc(&p->d_packed, lay::TOTAL_HB * sizeof(float))) != cudaSuccess) return fail(e, "cudaMalloc(packed)");
    if ((e = cudaMemset(p->d_packed, 0, lay::TOTAL_HB * sizeof(float))) != cudaSuccess) return fail(e, "cudaMemset(packed)");
    p->bufs.w = p->d_packed;
    *out = p;
    return MATAS_OK;
}

extern "C" int matas_policy_destroy(matas_policy* p) {
    if (!p) return MATAS_OK;
    cudaFree(p->d_table); cudaFree(p->d_folds); cudaFree(p->d_half); cudaFree(p->d_packed); cudaFree(p->d_tok); cudaFree(p->d_kv); cudaFree(p->d_x1tab); cudaFree(p->d_x1spec); cudaFree(p->d_train);
    if (p->ev_made) for (int i = 0; i < matas_policy::NEV; ++i) { cudaEventDestroy(p->ev0[i]); cudaEventDestroy(p->ev1[i]); }
    delete p;
    return MATAS_OK;
}

extern "C" long matas_policy_param_count(const matas_policy* p) { return p ? p->n_params : -1; }

// read-only view of a handle for entry points that live in other translation units
int matas_policy_view(const matas_policy* p, PolicyView* v) {
    ARG_CHECK(p && v, "null policy handle");
    v->n_agent = p->cfg.n_agent; v->mode = p->cfg.mode; v->device = p->device; v->n_sm = p->n_sm;
    v->weights_set = p->weights_set ? 1 : 0; v->packed = p->d_packed; v->x1tab = p->d_x1tab;
    return MATAS_OK;
}

extern "C" int matas_policy_set_weights(matas_policy* p, const float* flat, void* stream) {
    ARG_CHECK(p && flat, "null argument");
    pack_kernel<<<(unsigned)p->table.size(), 256, 0, (cudaStream_t)stream>>>(flat, p->d_packed, p->d_table);
    fold_kernel<<<(unsigned)p->folds.size(), 256, 0, (cudaStream_t)stream>>>(flat, p->d_packed, p->d_folds);
    half_kernel<<<(unsigned)p->n_half, 256, 0, (cudaStream_t)stream>>>(p->d_packed, p->d_half);
    CUDA_TRY(cudaGetLastError());
    {   // block-0 cross-attention inputs by (position, ones so far, previous action): decode_tc.cu
        const int L = p->cfg.n_agent;
        if (!p->d_x1tab) CUDA_TRY(cudaMalloc(&p->d_x1tab, (size_t)L * (L + 1) * 64 * sizeof(float)));
        int rc_ = launch_x1_table(p->d_packed, p->d_x1tab, L, (cudaStream_t)stream);
        if (rc_) return rc_;
        p->bufs.x1tab = p->d_x1tab;
        p->spec_stride = -1;
    }
    p->weights_set = true;
    return MATAS_OK;
}

static int alloc_tokens(matas_policy* p, long n_tok) {
    p->train_tokens = 0;            // any other forward overwrites K/V/rep that a pending backward would need
    if (n_tok <= p->tok_cap) return MATAS_OK;
    if (p->d_tok) CUDA_TRY(cudaFree(p->d_tok));
    p->d_tok = nullptr; p->tok_cap = 0;
    const long cap = n_tok + 64;
    constexpr int NBUF = 22;
    CUDA_TRY(cudaMalloc(&p->d_tok, (size_t)cap * 64 * sizeof(float) * NBUF));
    float* f = p->d_tok;
    auto take = [&]() { float* r = f; f += cap * 64; return r; };
    PolicyBufs& b = p->bufs;
    b.X = take(); b.Q = take(); b.rep = take(); b.Y = take(); b.Q2[0] = take(); b.Q2[1] = take();
    for (int k = 0; k < 2; ++k) for (int h = 0; h < 2; ++h) b.Q2p[k][h] = take();
    for (int k = 0; k < 2; ++k) { b.K[k] = take(); b.V[k] = take(); b.K1[k] = take(); b.V1[k] = take(); b.K2[k] = take(); b.V2[k] = take(); }
    p->tok_cap = cap;
    return MATAS_OK;
}

extern "C" int matas_policy_get_values(matas_policy* p, const float* obs, long B, float* values, void* stream) {
    int rc = check_handle(p);
    if (rc) return rc;
    ARG_CHECK(obs && values && B > 0, "bad argument (B=%ld)", B);
    if ((rc = alloc_tokens(p, B * p->cfg.n_agent))) return rc;
    return launch_encoder(p->bufs, obs, values, B, p->cfg.n_agent, p->cfg.obs_dim, p->n_sm, (cudaStream_t)stream);
}

static int run_decode(matas_policy* p, const float* ava, long B, int deterministic, int stride, uint64_t seed,
                      uint32_t step, long env_offset, const float* u, const float* z, float* actions, float* logp,
                      cudaStream_t stream) {
    const int L = p->cfg.n_agent;
    int grid, rows; long slots;
    { int rc_ = policy_decode_geometry(L, B, p->n_sm, &grid, &slots, &rows); if (rc_) return rc_; }
    const size_t need = (size_t)slots * 4 * L * 64 * sizeof(float);
    if (need > p->kv_cap) {
        if (p->d_kv) CUDA_TRY(cudaFree(p->d_kv));
        p->d_kv = nullptr; p->kv_cap = 0;
        CUDA_TRY(cudaMalloc(&p->d_kv, need));
        p->kv_cap = need;
    }
    p->bufs.kv = p->d_kv;
    p->bufs.x1spec = nullptr;
    if (deterministic && p->bufs.x1tab && p->cfg.mode == MATAS_MODE_SEMI) {
        // speculative block-0 rows of this stride's blocks (recomputed when the stride or the weights changed)
        const int sd = stride < 1 ? 1 : stride;
        const long rows = decode_tc_spec_rows(L, sd);
        if (rows > p->spec_cap) {
            if (p->d_x1spec) CUDA_TRY(cudaFree(p->d_x1spec));
            p->d_x1spec = nullptr; p->spec_cap = 0; p->spec_stride = -1;
            CUDA_TRY(cudaMalloc(&p->d_x1spec, (size_t)rows * 64 * sizeof(float)));
            p->spec_cap = rows;
        }
        if (rows > 0 && p->spec_stride != sd) {
            int rc_ = launch_x1_spec_table(p->d_packed, p->d_x1spec, L, sd, stream);
            if (rc_) return rc_;
            p->spec_stride = sd;
        }
        p->bufs.x1spec = p->d_x1spec ? p->d_x1spec : p->d_x1tab;      // no speculative rows (stride 1): any valid pointer
    }
    DecodeArgs a;
    a.B = B; a.L = L; a.mode = p->cfg.mode; a.deterministic = deterministic ? 1 : 0; a.stride = stride;
    a.n_steps = policy_decode_steps(L, a.mode, a.deterministic, stride);
    a.seed = seed; a.step = step; a.env_offset = env_offset;
    a.rows = rows;
    a.ava = ava; a.u = u; a.z = z; a.actions = actions; a.logp = logp;
    if (!p->timing) return launch_decode_tc(p->bufs, a, grid, stream);
    const int i = p->ev_next;
    p->ev_next = (i + 1) % matas_policy::NEV;
    int rc = harvest(p, i);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(p->ev0[i], stream));
    rc = launch_decode_tc(p->bufs, a, grid, stream);
    CUDA_TRY(cudaEventRecord(p->ev1[i], stream));
    p->ev_pending[i] = true;
    return rc;
}

extern "C" int matas_policy_timing(matas_policy* p, int enable) {
    ARG_CHECK(p != nullptr, "null policy handle");
    if (enable && !p->ev_made) {
        for (int i = 0; i < matas_policy::NEV; ++i) {
            CUDA_TRY(cudaEventCreate(&p->ev0[i]));
            CUDA_TRY(cudaEventCreate(&p->ev1[i]));
            p->ev_pending[i] = false;
        }
        p->ev_made = true;
    }
    if (p->ev_made) for (int i = 0; i < matas_policy::NEV; ++i) { int rc = harvest(p, i); if (rc) return rc; }
    p->timing = enable != 0;
    p->t_ms = 0.0; p->t_n = 0;
    return MATAS_OK;
}

extern "C" int matas_policy_decode_time(matas_policy* p, double* ms_total, long* launches) {
    ARG_CHECK(p && ms_total && launches, "null argument");
    if (p->ev_made) for (int i = 0; i < matas_policy::NEV; ++i) { int rc = harvest(p, i); if (rc) return rc; }
    *ms_total = p->t_ms; *launches = p->t_n;
    return MATAS_OK;
}

// the device copy of the named-draw generator (common.cuh), exposed so that tests can pin it to the Random123 known answers
__global__ void philox_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, long n) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t* c = in + i * 6;
    const uint4 x = philox4x32(c[0], c[1], c[2], c[3], (uint64_t)c[4] | ((uint64_t)c[5] << 32));
    out[i * 4] = x.x; out[i * 4 + 1] = x.y; out[i * 4 + 2] = x.z; out[i * 4 + 3] = x.w;
}

extern "C" int matas_philox4x32_10(const uint32_t* counter_key, uint32_t* out, long n, void* stream) {
    ARG_CHECK(counter_key && out && n > 0, "bad argument (n=%ld)", n);
    philox_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(counter_key, out, n);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}

// dependent-FFMA probe: 8 independent chains per thread, 2 FLOP per FFMA
__global__ void fp32_peak_kernel(float* out, int iters) {
    float a[8];
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-3f + i;
    const float b = 1.0000001f, c = 1e-7f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], b, c);
    }
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += a[i];
    if (s == 123.456f) out[0] = s;
}

extern "C" int matas_fp32_peak(int iters, double* tflops) {
    ARG_CHECK(iters > 0 && tflops, "bad argument");
    int dev = 0, n_sm = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    CUDA_TRY(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
    float* d = nullptr;
    CUDA_TRY(cudaMalloc(&d, 4));
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0)); CUDA_TRY(cudaEventCreate(&e1));
    const int grid = n_sm * 8, block = 256;
    fp32_peak_kernel<<<grid, block>>>(d, iters / 10 + 1);
    double best = 0.0;
    for (int rep = 0; rep < 3; ++rep) {
        CUDA_TRY(cudaEventRecord(e0));
        fp32_peak_kernel<<<grid, block>>>(d, iters);
        CUDA_TRY(cudaEventRecord(e1));
        CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0.f;
        CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
        const double fl = 2.0 * 64.0 * (double)iters * (double)grid * block;
        if (fl / (ms * 1e-3) / 1e12 > best) best = fl / (ms * 1e-3) / 1e12;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    *tflops = best;
    return MATAS_OK;
}

extern "C" int matas_policy_get_actions(matas_policy* p, const float* obs, const float* ava, long B, int deterministic,
                                        int stride, uint64_t seed, uint32_t step, long env_offset, const float* u,
                                        const float* z, float* values, float* actions, float* logp, void* stream) {
    int rc = check_handle(p);
    if (rc) return rc;
    ARG_CHECK(obs && values && actions && logp && B > 0, "bad argument (B=%ld)", B);
    if ((rc = alloc_tokens(p, B * p->cfg.n_agent))) return rc;
    if ((rc = launch_encoder(p->bufs, obs, values, B, p->cfg.n_agent, p->cfg.obs_dim, p->n_sm, (cudaStream_t)stream, nullptr, 1))) return rc;
    return run_decode(p, ava, B, deterministic, stride, seed, step, env_offset, u, z, actions, logp, (cudaStream_t)stream);
}

// PolicyBufs whose token buffers start `tok0` tokens in (one env-aligned slice of a larger batch)
static PolicyBufs shifted(const PolicyBufs& b, long tok0) {
    PolicyBufs s = b;
    const long o = tok0 * 64;
    s.X += o; s.Q += o; s.rep += o; s.Y += o;
    for (int k = 0; k < 2; ++k) {
        s.K[k] += o; s.V[k] += o; s.K1[k] += o; s.V1[k] += o; s.K2[k] += o; s.V2[k] += o; s.Q2[k] += o;
        for (int h = 0; h < 2; ++h) s.Q2p[k][h] += o;
    }
    return s;
}

/* get_actions in two parts, so that a caller whose observations arrive from the host in slices can run the encoder of
 * one slice while the next slice is still on the PCIe bus: encoder of envs [env0, env0 + n_env) of a batch of B_total
 * (obs, values point at the slice's first row), then -- once every slice is in -- the decode of the whole batch. */
extern "C" int matas_policy_encode_slice(matas_policy* p, const float* obs, long env0, long n_env, long B_total, float* values,
                                         void* stream) {
    int rc = check_handle(p);
    if (rc) return rc;
    ARG_CHECK(obs && values && n_env > 0 && env0 >= 0 && env0 + n_env <= B_total, "bad slice (env0=%ld n_env=%ld B=%ld)", env0, n_env, B_total);
    const int L = p->cfg.n_agent;
    if (B_total * L > p->tok_cap) {
        ARG_CHECK(env0 == 0, "the first slice must start at env 0 (it sizes the token buffers)");
        if ((rc = alloc_tokens(p, B_total * L))) return rc;
    } else {
        p->train_tokens = 0;
    }
    const PolicyBufs b = shifted(p->bufs, env0 * L);
    return launch_encoder(b, obs, values, n_env, L, p->cfg.obs_dim, p->n_sm, (cudaStream_t)stream, nullptr, 1);
}

extern "C" int matas_policy_decode(matas_policy* p, const float* ava, long B, int deterministic, int stride, uint64_t seed,
                                   uint32_t step, long env_offset, const float* u, const float* z, float* actions, float* logp,
                                   void* stream) {
    int rc = check_handle(p);
    if (rc) return rc;
    ARG_CHECK(actions && logp && B > 0, "bad argument (B=%ld)", B);
    ARG_CHECK(B * p->cfg.n_agent <= p->tok_cap, "matas_policy_decode: the token buffers hold no encoder output for %ld envs", B);
    return run_decode(p, ava, B, deterministic, stride, seed, step, env_offset, u, z, actions, logp, (cudaStream_t)stream);
}

extern "C" int matas_policy_act_autoregressive(matas_policy* p, const float* obs_rep, const float* ava, long B,
                                               int deterministic, int stride, uint64_t seed, uint32_t step,
                                               long env_offset, const float* u, const float* z, float* actions,
                                               float* logp, void* stream) {
    int rc = check_handle(p);
    if (rc) return rc;
    ARG_CHECK(obs_rep && actions && logp && B > 0, "bad argument (B=%ld)", B);
    if ((rc = alloc_tokens(p, B * p->cfg.n_agent))) return rc;
    if ((rc = launch_q2_from_rep(p->bufs, obs_rep, B, p->cfg.n_agent, (cudaStream_t)stream, 1))) return rc;
    return run_decode(p, ava, B, deterministic, stride, seed, step, env_offset, u, z, actions, logp, (cudaStream_t)stream);
}

extern "C" int matas_policy_act_parallel(matas_policy* p, const float* obs_rep, const float* ava, const float* actions,
                                         long M, float* logp, float* entropy, void* stream) {
    int rc = check_handle(p);
    if (rc) return rc;
    ARG_CHECK(obs_rep && actions && logp && entropy && M > 0, "bad argument (M=%ld)", M);
    const int L = p->cfg.n_agent;
    if ((rc = alloc_tokens(p, M * L))) return rc;
    if ((rc = launch_q2_from_rep(p->bufs, obs_rep, M, L, (cudaStream_t)stream, 0))) return rc;
    return launch_teacher_forced(p->bufs, actions, ava, logp, entropy, M, L, p->cfg.mode, p->n_sm, (cudaStream_t)stream);
}

static int alloc_train(matas_policy* p, long n_tok) {
    if (n_tok > p->train_cap) {
        if (p->d_train) CUDA_TRY(cudaFree(p->d_train));
        p->d_train = nullptr; p->train_cap = 0;
        const long cap = n_tok + 64;
        constexpr int WIDE = 45, NARROW = 27;
        CUDA_TRY(cudaMalloc(&p->d_train, (size_t)cap * (WIDE * 64 + NARROW + 5) * sizeof(float)));
        float* f = p->d_train;
        auto wide = [&]() { float* r = f; f += cap * 64; return r; };
        TrainBufs& t = p->tb;
        t.e_h0 = wide(); t.e_xh0 = wide(); t.e_x0 = wide(); t.d_x0 = wide(); t.e_hh = wide(); t.e_xhh = wide();
        t.d_hh = wide(); t.d_xhh = wide(); t.gX = wide(); t.gRep = wide(); t.gYa = wide(); t.gYb = wide();
        for (int k = 0; k < 9; ++k) t.gS[k] = wide();
        for (int k = 0; k < 2; ++k) {
            t.e_q[k] = wide(); t.e_y[k] = wide(); t.e_xh1[k] = wide(); t.e_m0[k] = wide(); t.e_xh2[k] = wide();
            t.d_q1[k] = wide(); t.d_y1[k] = wide(); t.d_xh1[k] = wide(); t.d_y2[k] = wide(); t.d_xh2[k] = wide();
            t.d_m0[k] = wide(); t.d_xh3[k] = wide();
        }
        auto narrow = [&](int w) { float* r = f; f += ((cap * w + 3) & ~3L); return r; };
        t.e_rs0 = narrow(1); t.e_rsh = narrow(1); t.d_rsh = narrow(1); t.gDa = narrow(2); t.gDb = narrow(2);
        for (int k = 0; k < 2; ++k) {
            t.e_rs1[k] = narrow(1); t.e_rs2[k] = narrow(1); t.e_lse[k] = narrow(2);
            t.d_rs1[k] = narrow(1); t.d_rs2[k] = narrow(1); t.d_rs3[k] = narrow(1); t.d_lse1[k] = narrow(2); t.d_lse2[k] = narrow(2);
        }
        p->train_cap = cap;
    }
    p->tb.on = 1;
    return MATAS_OK;
}

/* MultiAgentTransformer.forward in training mode: as matas_policy_evaluate, and additionally keeps the
 * activations the backward needs (until the next call on this handle that runs a forward). */
extern "C" int matas_policy_evaluate_train(matas_policy* p, const float* obs, const float* ava, const float* actions, long M,
                                           float* values, float* logp, float* entropy, void* stream) {
    int rc = check_handle(p);
    if (rc) return rc;
    ARG_CHECK(obs && actions && values && logp && entropy && M > 0, "bad argument (M=%ld)", M);
    const int L = p->cfg.n_agent;
    if (L > 256) { matas_set_error("training supports n_agent <= 256 (got %d)", L); return MATAS_ERR_LIMIT; }
    if ((rc = alloc_tokens(p, M * L))) return rc;
    if ((rc = alloc_train(p, M * L))) return rc;
    if ((rc = launch_encoder(p->bufs, obs, values, M, L, p->cfg.obs_dim, p->n_sm, (cudaStream_t)stream, &p->tb))) return rc;
    if ((rc = launch_teacher_forced(p->bufs, actions, ava, logp, entropy, M, L, p->cfg.mode, p->n_sm, (cudaStream_t)stream, &p->tb))) return rc;
    p->train_tokens = M * L;
    return MATAS_OK;
}

/* loss.backward() through MultiAgentTransformer.forward (mat_trainer.py:146-147): given the
 * gradients of the loss w.r.t. values / log-probs / entropies [M*L], ACCUMULATES the gradient w.r.t.
 * every parameter into flat_grad (same layout as the flat parameter vector). */
extern "C" int matas_policy_evaluate_backward(matas_policy* p, const float* obs, const float* ava, const float* actions,
                                              long M, const float* d_values, const float* d_logp, const float* d_entropy,
                                              float* flat_grad, void* stream) {
    int rc = check_handle(p);
    if (rc) return rc;
    ARG_CHECK(obs && actions && d_values && d_logp && d_entropy && flat_grad && M > 0, "bad argument (M=%ld)", M);
    const int L = p->cfg.n_agent;
    ARG_CHECK(p->train_tokens == M * L, "matas_policy_evaluate_backward must directly follow matas_policy_evaluate_train "
              "with the same batch (saved %ld tokens, asked %ld)", p->train_tokens, M * L);
    return launch_evaluate_backward(p->bufs, p->tb, p->gm, flat_grad, obs, actions, ava, d_values, d_logp, d_entropy, M, L,
                                    p->cfg.obs_dim, p->cfg.mode, p->n_sm, (cudaStream_t)stream);
}

/* encoder only, exposing obs_rep (Encoder.forward, ma_transformer.py:141-154) */
extern "C" int matas_policy_encode(matas_policy* p, const float* obs, long B, float* values, float* obs_rep, void* stream) {
    int rc = check_handle(p);
    if (rc) return rc;
    ARG_CHECK(obs && values && obs_rep && B > 0, "bad argument (B=%ld)", B);
    const long n_tok = B * p->cfg.n_agent;
    if ((rc = alloc_tokens(p, n_tok))) return rc;
    if ((rc = launch_encoder(p->bufs, obs, values, B, p->cfg.n_agent, p->cfg.obs_dim, p->n_sm, (cudaStream_t)stream))) return rc;
    CUDA_TRY(cudaMemcpyAsync(obs_rep, p->bufs.rep, (size_t)n_tok * 64 * sizeof(float), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return MATAS_OK;
}

extern "C" int matas_policy_evaluate(matas_policy* p, const float* obs, const float* ava, const float* actions, long M,
                                     float* values, float* logp, float* entropy, void* stream) {
    int rc = check_handle(p);
    if (rc) return rc;
    ARG_CHECK(obs && actions && values && logp && entropy && M > 0, "bad argument (M=%ld)", M);
    const int L = p->cfg.n_agent;
    if ((rc = alloc_tokens(p, M * L))) return rc;
    if ((rc = launch_encoder(p->bufs, obs, values, M, L, p->cfg.obs_dim, p->n_sm, (cudaStream_t)stream))) return rc;
    return launch_teacher_forced(p->bufs, actions, ava, logp, entropy, M, L, p->cfg.mode, p->n_sm, (cudaStream_t)stream);
}
