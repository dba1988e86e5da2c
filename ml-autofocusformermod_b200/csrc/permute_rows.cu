// permute_rows.cu — token reorder of a stage: dst[b][i][:] = src[b][perm[b][i]][:]  (gather) and its adjoint
// dsrc[b][perm[b][i]][:] = ddst[b][i][:] (scatter).  Reference: `feat = feat[b, reorder]` after the space-filling
// clustering (models/aff_transformer.py:466,472; mixres_neighbour.py:503-504,514-515).  pos_ranking is a permutation
// of the tokens, so the adjoint is a plain scatter: no atomics, no zero fill, run-to-run identical.  The stock path is
// an index gather forward and a zero-fill + scatter_add backward.  Rows are moved as 16-byte chunks, consecutive
// lanes on consecutive chunks of a row.
#include "common.cuh"
#include <algorithm>

namespace affb200 {

template <bool SCATTER>
__global__ void __launch_bounds__(256)
permute_rows_kernel(const uint4 *__restrict__ src, const int64_t *__restrict__ perm, uint4 *__restrict__ dst, int N,
                    int chunks, int64_t total) {
    for (int64_t t = int64_t(blockIdx.x) * 256 + threadIdx.x; t < total; t += int64_t(gridDim.x) * 256) {
        const int64_t row = t / chunks;          // b*N + i
        const int ch = int(t - row * chunks);
        const int64_t b = row / N;
        const int64_t other = b * N + __ldg(perm + row);
        if (SCATTER) dst[other * chunks + ch] = __ldg(src + t);
        else dst[t] = __ldg(src + other * chunks + ch);
    }
}

}  // namespace affb200

using namespace affb200;

extern "C" int affb200_permute_rows(const void *src, const int64_t *perm, void *dst, int64_t B, int64_t N,
                                    int64_t row_bytes, int scatter, void *stream) {
    if (B < 0 || N < 0 || row_bytes < 0 || N >= (int64_t(1) << 31) || row_bytes % 16 != 0 ||
        row_bytes / 16 >= (int64_t(1) << 31)) {
        set_error("permute_rows: rows must be a multiple of 16 bytes, N < 2^31");
        return AFFB200_E_ARG;
    }
    const int64_t total = B * N * (row_bytes / 16);
    if (total == 0) return AFFB200_OK;
    if (!src || !perm || !dst || !aligned16(src) || !aligned16(dst)) {
        set_error("permute_rows: null or misaligned pointer");
        return AFFB200_E_ARG;
    }
    const unsigned grid = unsigned(std::min<int64_t>((total + 255) / 256, int64_t(kNumSMs) * 32));
    auto st = (cudaStream_t)stream;
    if (scatter)
        permute_rows_kernel<true><<<grid, 256, 0, st>>>((const uint4 *)src, perm, (uint4 *)dst, int(N),
                                                        int(row_bytes / 16), total);
    else
        permute_rows_kernel<false><<<grid, 256, 0, st>>>((const uint4 *)src, perm, (uint4 *)dst, int(N),
                                                         int(row_bytes / 16), total);
    return check_launch("permute_rows");
}
