// Host side of tma.cuh: tensor maps of the token-major activation arrays, encoded through the driver entry point
// (cuTensorMapEncodeTiled has no runtime-API twin; the library does not link libcuda).
#include "tma.cuh"

#include "common.cuh"

namespace tma {

namespace {
typedef CUresult(CUDAAPI* EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                       const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                       CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiled encode_fn() {
    static EncodeTiled fn = nullptr;                 // benign race: every thread resolves the same pointer
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiled>(p);
    }
    return fn;
}
// the encode is pure host arithmetic, but a rollout step builds the same maps every time (15 per encoder pass, times the
// four env slices of a host-array call): keep the last ones
constexpr int N_CACHED = 128;
struct Cached { const float* base; long n_rows; CUtensorMap map; };
thread_local Cached t_cache[N_CACHED];
thread_local int t_next = 0;
}  // namespace

int rows_map(CUtensorMap* out, const float* base, long n_rows) {
    if (!base || n_rows <= 0) {                      // an output this launch does not produce: never dereferenced by the kernel
        memset(out, 0, sizeof(*out));
        return MATAS_OK;
    }
    for (int i = 0; i < N_CACHED; ++i)
        if (t_cache[i].base == base && t_cache[i].n_rows == n_rows) { *out = t_cache[i].map; return MATAS_OK; }
    EncodeTiled fn = encode_fn();
    if (!fn) {
        matas_set_error("cuTensorMapEncodeTiled is not available from this CUDA driver");
        return MATAS_ERR_CUDA;
    }
    const cuuint64_t dims[2] = {64, (cuuint64_t)n_rows};
    const cuuint64_t strides[1] = {64 * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)BOX_COLS, (cuuint32_t)BOX_ROWS};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult rc = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) {
        matas_set_error("cuTensorMapEncodeTiled failed (%d) for %ld rows at %p", (int)rc, n_rows, (const void*)base);
        return MATAS_ERR_CUDA;
    }
    Cached& c = t_cache[t_next];
    t_next = (t_next + 1) % N_CACHED;
    c.base = base; c.n_rows = n_rows; c.map = *out;
    return MATAS_OK;
}

}  // namespace tma
