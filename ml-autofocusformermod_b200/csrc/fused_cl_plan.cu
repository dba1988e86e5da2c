// fused_cl_plan.cu — plans of the block-structured fused attention (see fused_cl.cuh).
//
// affb200_cl_qplan_build: verifies that an (nbhd_idx, cluster_mask, pe_idx, pos) set has the structure the reference's
// index build produces (models/aff_transformer.py:476-487, models/point_utils.py:186-206):
//   nbhd_idx[b,i,8q+r] = blk[b,i,q]*8 + r for rows < N, padded rows (>= N) masked out,
//   pe_idx[b,i,j] = (y_u - y_i + w)*tw + (x_u - x_i + w) for integer positions,
// and emits the compact form: block ids per token, per tile of 64 tokens the ascending union of referenced blocks,
// per token a 64-bit membership word over that union, and packed 16-bit positions.  Anything else (random indices of
// the reference's unit tests, clamped offsets, fractional positions, a token listing a block twice) sets the tile's
// BAD flag and is counted in stats: the caller then keeps the generic row-union path.
//
// affb200_cl_kplan_build: the transposed view for the dK/dV kernel: per tile of 64 key rows (8 blocks) the ascending
// list of query tokens that reference any of them, each with an 8-bit "which of the 8 blocks" word.
#include "fused_cl.cuh"

namespace affb200 {

constexpr int kPlanThreads = 128;

__global__ void __launch_bounds__(kPlanThreads)
cl_qplan_kernel(const int64_t *__restrict__ nbhd, const int64_t *__restrict__ mask, const int64_t *__restrict__ pe_idx,
                const float *__restrict__ pos, int N, int M, int w, int tw, uint16_t *__restrict__ blk_out,
                int32_t *__restrict__ hdr, uint16_t *__restrict__ ucl, unsigned long long *__restrict__ qmask,
                uint32_t *__restrict__ pos16, int32_t *__restrict__ stats) {
    unsigned char *smem = affb200_dsm;
    const int nblk = (N + 7) >> 3, W = (nblk + 31) >> 5, NB = M >> 3;
    uint32_t *bitmap = reinterpret_cast<uint32_t *>(smem);          // [W]
    uint32_t *pre = bitmap + W;                                      // [W] exclusive popcount prefix
    unsigned long long *qm = reinterpret_cast<unsigned long long *>(pre + W);  // [64] (2W words: 8-byte aligned)
    uint16_t *blk_s = reinterpret_cast<uint16_t *>(qm + kClT);       // [64 * NB]
    __shared__ int bad_s, warp_tot[kPlanThreads / 32];
    const int tid = threadIdx.x, tile = blockIdx.x, b = blockIdx.y, n_tiles = gridDim.x;
    const int t0 = tile * kClT, ti = min(kClT, N - t0);
    const int64_t bN = int64_t(b) * N;
    for (int i = tid; i < W; i += kPlanThreads) bitmap[i] = 0u;
    if (tid < kClT) qm[tid] = 0ull;
    if (tid == 0) bad_s = 0;
    __syncthreads();
    int bad = 0;
    // packed positions of the tile's own tokens (and the sentinel of the padded tail)
    const int64_t bP = int64_t(b) * n_tiles * kClT;   // qmask / pos16 rows are padded to whole tiles
    for (int t = tid; t < kClT; t += kPlanThreads) {
        uint32_t pk = kClPosInvalid;
        if (t < ti) {
            const float x = pos[(bN + t0 + t) * 2], y = pos[(bN + t0 + t) * 2 + 1];
            if (!(x >= 0.f && x <= 32767.f && y >= 0.f && y <= 32767.f) || x != floorf(x) || y != floorf(y)) bad = 1;
            pk = (uint32_t(int(y)) << 16) | uint32_t(int(x));
        }
        pos16[bP + t0 + t] = pk;
    }
    // structure check + block ids
    for (int it = tid; it < ti * NB; it += kPlanThreads) {
        const int t = it / NB, q = it - t * NB;
        const int64_t so = (bN + t0 + t) * M + q * 8;
        const int64_t base = nbhd[so];
        int blk = 0;
        if (base < 0 || base >= N || (base & 7) != 0) {
            bad = 1;
        } else {
            blk = int(base >> 3);
            const float xi = pos[(bN + t0 + t) * 2], yi = pos[(bN + t0 + t) * 2 + 1];
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int64_t row = base + r;
                if (row < N) {
                    if (nbhd[so + r] != row || (mask && mask[so + r] == 0)) bad = 1;
                    const float xu = pos[(bN + row) * 2], yu = pos[(bN + row) * 2 + 1];
                    const int64_t want = int64_t(int(yu) - int(yi) + w) * tw + (int(xu) - int(xi) + w);
                    if (pe_idx[so + r] != want) bad = 1;
                } else if (!mask || mask[so + r] != 0) {
                    bad = 1;
                }
            }
            atomicOr(&bitmap[blk >> 5], 1u << (blk & 31));
        }
        blk_s[it] = uint16_t(blk);
        blk_out[(bN + t0 + t) * NB + q] = uint16_t(blk);
    }
    if (bad) atomicOr(&bad_s, 1);
    __syncthreads();
    // exclusive prefix of the per-word popcounts
    const int wpt = (W + kPlanThreads - 1) / kPlanThreads;
    int local = 0;
    for (int k = 0; k < wpt; ++k) {
        const int wi = tid * wpt + k;
        if (wi < W) local += __popc(bitmap[wi]);
    }
    int incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if ((tid & 31) >= o) incl += v;
    }
    if ((tid & 31) == 31) warp_tot[tid >> 5] = incl;
    __syncthreads();
    int woff = 0, U = 0;
#pragma unroll
    for (int k = 0; k < kPlanThreads / 32; ++k) {
        if (k < (tid >> 5)) woff += warp_tot[k];
        U += warp_tot[k];
    }
    int run = woff + incl - local;
    for (int k = 0; k < wpt; ++k) {
        const int wi = tid * wpt + k;
        if (wi < W) {
            pre[wi] = uint32_t(run);
            uint32_t bits = bitmap[wi];
            uint16_t *ul = ucl + (int64_t(b) * n_tiles + tile) * kClUMax;
            while (bits) {
                const int bit = __ffs(bits) - 1;
                bits &= bits - 1;
                if (run < kClUMax) ul[run] = uint16_t(wi * 32 + bit);
                ++run;
            }
        }
    }
    __syncthreads();
    if (U <= kClUMax) {
        for (int it = tid; it < ti * NB; it += kPlanThreads) {
            const int t = it / NB, blk = blk_s[it];
            const int loc = int(pre[blk >> 5]) + __popc(bitmap[blk >> 5] & ((1u << (blk & 31)) - 1u));
            const unsigned long long bitv = 1ull << loc;
            const unsigned long long old = atomicOr(&qm[t], bitv);
            if (old & bitv) atomicOr(&bad_s, 1);   // a token lists a block twice
        }
    }
    __syncthreads();
    for (int t = tid; t < kClT; t += kPlanThreads) qmask[bP + t0 + t] = t < ti ? qm[t] : 0ull;
    if (tid == 0) {
        const bool over = U > kClUMax;
        const bool isbad = bad_s != 0;
        hdr[int64_t(b) * n_tiles + tile] = U | (over ? AFFB200_PLAN_OVERFLOW : 0) | (isbad ? AFFB200_CL_BAD : 0);
        if (over) atomicAdd(&stats[0], 1);
        if (isbad) atomicAdd(&stats[1], 1);
        atomicMax(&stats[4], U);
    }
}

// one CTA per (key tile, image): scan all query tokens in order, keep those referencing blocks [8*kt, 8*kt+8)
__global__ void __launch_bounds__(kPlanThreads)
cl_kplan_kernel(const uint16_t *__restrict__ blk, int N, int NB, int kumax, int32_t *__restrict__ hdr,
                uint32_t *__restrict__ utok, int32_t *__restrict__ stats) {
    __shared__ int warp_tot[kPlanThreads / 32];
    __shared__ int base_s;
    const int tid = threadIdx.x, kt = blockIdx.x, b = blockIdx.y, n_tiles = gridDim.x;
    const int lane = tid & 31, warp = tid >> 5;
    uint32_t *out = utok + (int64_t(b) * n_tiles + kt) * kumax;
    if (tid == 0) base_s = 0;
    __syncthreads();
    for (int i0 = 0; i0 < N; i0 += kPlanThreads) {
        const int i = i0 + tid;
        uint32_t bits = 0u;
        if (i < N) {
            const uint16_t *bp = blk + (int64_t(b) * N + i) * NB;
            for (int q = 0; q < NB; ++q) {
                const int v = bp[q];
                if ((v >> 3) == kt) bits |= 1u << (v & 7);
            }
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, bits != 0u);
        if (lane == 0) warp_tot[warp] = __popc(bal);
        __syncthreads();
        int off = base_s;
        for (int k = 0; k < warp; ++k) off += warp_tot[k];
        off += __popc(bal & ((1u << lane) - 1u));
        if (bits != 0u && off < kumax) out[off] = (uint32_t(i) << 8) | bits;
        __syncthreads();
        if (tid == 0) {
            int t = 0;
            for (int k = 0; k < kPlanThreads / 32; ++k) t += warp_tot[k];
            base_s += t;
        }
        __syncthreads();
    }
    if (tid == 0) {
        const int U = base_s;
        hdr[int64_t(b) * n_tiles + kt] = U | (U > kumax ? AFFB200_PLAN_OVERFLOW : 0);
        if (U > kumax) atomicAdd(&stats[2], 1);
        atomicMax(&stats[5], U);
    }
}

}  // namespace affb200

using namespace affb200;

extern "C" int64_t affb200_cl_tiles(int64_t n) { return n <= 0 ? 0 : (n + kClT - 1) / kClT; }

extern "C" int affb200_cl_qplan_build(const int64_t *nbhd_idx, const int64_t *mask, const int64_t *pe_idx,
                                      const float *pos, int64_t B, int64_t N, int64_t M, int64_t rel_pos_width,
                                      int64_t table_width, uint16_t *blk, int32_t *hdr, uint16_t *ucl, uint64_t *qmask,
                                      uint32_t *pos16, int32_t *stats, void *stream) {
    if (B <= 0 || B > 65535 || N <= 0 || N > 32767 * 8 || M <= 0 || M % 8 != 0 || M > 8 * kClUMax
        || rel_pos_width < 0 || table_width != 2 * rel_pos_width + 1) {
        set_error("cl_qplan_build: unsupported dimensions (B=%lld N=%lld M=%lld w=%lld tw=%lld)", (long long)B,
                  (long long)N, (long long)M, (long long)rel_pos_width, (long long)table_width);
        return AFFB200_E_ARG;
    }
    if (!nbhd_idx || !pe_idx || !pos || !blk || !hdr || !ucl || !qmask || !pos16 || !stats) {
        set_error("cl_qplan_build: null pointer");
        return AFFB200_E_ARG;
    }
    const int nblk = int((N + 7) / 8), W = (nblk + 31) / 32, NB = int(M / 8);
    const size_t smem = size_t(2 * W) * 4 + size_t(kClT) * 8 + size_t(kClT) * NB * 2;
    const dim3 grid{unsigned((N + kClT - 1) / kClT), unsigned(B), 1u};
    cl_qplan_kernel<<<grid, kPlanThreads, smem, (cudaStream_t)stream>>>(
        nbhd_idx, mask, pe_idx, pos, int(N), int(M), int(rel_pos_width), int(table_width), blk, hdr, ucl,
        reinterpret_cast<unsigned long long *>(qmask), pos16, stats);
    return check_launch("cl_qplan_build");
}

extern "C" int affb200_cl_kplan_build(const uint16_t *blk, int64_t B, int64_t N, int64_t M, int64_t kumax,
                                      int32_t *hdr, uint32_t *utok, int32_t *stats, void *stream) {
    if (B <= 0 || B > 65535 || N <= 0 || N >= (1 << 24) || M <= 0 || M % 8 != 0 || kumax <= 0 || kumax % 16 != 0) {
        set_error("cl_kplan_build: unsupported dimensions");
        return AFFB200_E_ARG;
    }
    if (!blk || !hdr || !utok || !stats) { set_error("cl_kplan_build: null pointer"); return AFFB200_E_ARG; }
    const dim3 grid{unsigned((N + kClT - 1) / kClT), unsigned(B), 1u};
    cl_kplan_kernel<<<grid, kPlanThreads, 0, (cudaStream_t)stream>>>(blk, int(N), int(M / 8), int(kumax), hdr, utok,
                                                                   stats);
    return check_launch("cl_kplan_build");
}
