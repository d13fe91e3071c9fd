// Tensor-map TMA for the token-major activation arrays ([n_tok][64] fp32) of the token-parallel kernels (sm_100a).
//
// A 128-token tile of such an array is moved as two boxes of [128 rows][32 floats] (128 bytes per row, the widest inner
// extent the 128-byte swizzle allows) between global memory and a DENSE shared tile, by one thread:
//     cp.async.bulk.tensor.2d  (SASS UTMALDG / UTMASTG)
// Shared layout of a box: row r at r * 128 bytes, its 16-byte chunk c at chunk position c ^ (r % 8) (SWIZZLE_128B), so that
// the thread-per-row epilogues (32 consecutive rows per warp instruction) read and write it without bank conflicts.
// Rows past the end of the array are clipped on stores and zero-filled on loads by the tensor map itself (its row extent
// is the launch's n_tok), which is what the ragged last tile needs.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace tma {

constexpr int BOX_ROWS = 128, BOX_COLS = 32;
constexpr int BOX_FLOATS = BOX_ROWS * BOX_COLS;        // 4096 floats = 16 KB; a tile is two boxes (column halves)
constexpr int MAX_MAPS = 6;
struct RowMaps { CUtensorMap m[MAX_MAPS]; };           // passed to kernels as one `const __grid_constant__` parameter

// host: tensor map of rows [0, n_rows) x 64 fp32 at `base` (cudaMalloc'd, 16-byte aligned); returns a MATAS_* code
int rows_map(CUtensorMap* out, const float* base, long n_rows);

#ifdef __CUDACC__
__device__ __forceinline__ void store_box(const CUtensorMap* map, uint32_t smem_box, int col, int row) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(reinterpret_cast<uint64_t>(map)),
                 "r"(col), "r"(row), "r"(smem_box)
                 : "memory");
}
__device__ __forceinline__ void store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// the issuing thread: every committed store has finished READING shared memory (the tile may be rewritten)
__device__ __forceinline__ void store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void load_box(uint32_t smem_box, const CUtensorMap* map, int col, int row, uint32_t mbar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_box),
                 "l"(reinterpret_cast<uint64_t>(map)), "r"(col), "r"(row), "r"(mbar)
                 : "memory");
}
// 32 consecutive columns [32 half, 32 half + 32) of row `row` of the dense swizzled tile sD
__device__ __forceinline__ void row_store32(float* sD, int row, int half, const float (&v)[32]) {
    float* p = sD + half * BOX_FLOATS + row * BOX_COLS;
#pragma unroll
    for (int j = 0; j < 8; ++j) *reinterpret_cast<float4*>(p + ((j ^ (row & 7)) << 2)) = make_float4(v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
}
__device__ __forceinline__ void row_load32(const float* sD, int row, int half, float (&v)[32]) {
    const float* p = sD + half * BOX_FLOATS + row * BOX_COLS;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float4 x = *reinterpret_cast<const float4*>(p + ((j ^ (row & 7)) << 2));
        v[j * 4] = x.x; v[j * 4 + 1] = x.y; v[j * 4 + 2] = x.z; v[j * 4 + 3] = x.w;
    }
}
#endif

}  // namespace tma
