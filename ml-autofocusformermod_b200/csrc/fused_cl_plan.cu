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

// Ascending list of the set bits of bitmap[W] (block ids) -> ul[0..63], and pre[w] = number of set bits below word w.
// Returns the number of set bits to every thread.  Ends on a barrier.
__device__ __forceinline__ int cl_union_list(const uint32_t *bitmap, uint32_t *pre, int W, int *warp_tot, uint16_t *ul) {
    const int tid = threadIdx.x;
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
            while (bits) {
                const int bit = __ffs(bits) - 1;
                bits &= bits - 1;
                if (run < kClUMax) ul[run] = uint16_t(wi * 32 + bit);
                ++run;
            }
        }
    }
    __syncthreads();
    return U;
}

// membership words of tokens [ta, tb): bit = position of each of the token's blocks in the union list
__device__ __forceinline__ void cl_membership(const uint32_t *bitmap, const uint32_t *pre, const uint16_t *blk_s,
                                              unsigned long long *qm, int *bad_s, int ta, int tb, int NB) {
    for (int it = ta * NB + threadIdx.x; it < tb * NB; it += kPlanThreads) {
        const int t = it / NB, blk = blk_s[it];
        if (blk == kClAbsent) continue;
        const int loc = int(pre[blk >> 5]) + __popc(bitmap[blk >> 5] & ((1u << (blk & 31)) - 1u));
        const unsigned long long bitv = 1ull << loc;
        const unsigned long long old = atomicOr(&qm[t], bitv);
        if (old & bitv) atomicOr(bad_s, 1);   // a token lists a block twice
    }
}

__global__ void __launch_bounds__(kPlanThreads)
cl_qplan_kernel(const int64_t *__restrict__ nbhd, const int64_t *__restrict__ mask, const int64_t *__restrict__ pe_idx,
                const float *__restrict__ pos, int N, int M, int w, int tw, uint16_t *__restrict__ blk_out,
                int32_t *__restrict__ hdr, uint16_t *__restrict__ ucl, int32_t *__restrict__ hdr2,
                uint16_t *__restrict__ ucl2, unsigned long long *__restrict__ qmask, uint32_t *__restrict__ pos16,
                int32_t *__restrict__ stats) {
    unsigned char *smem = affb200_dsm;
    const int nblk = (N + 7) >> 3, W = (nblk + 31) >> 5, NB = M >> 3;
    uint32_t *bitmap = reinterpret_cast<uint32_t *>(smem);          // [W]
    uint32_t *pre = bitmap + W;                                      // [W] exclusive popcount prefix
    unsigned long long *qm = reinterpret_cast<unsigned long long *>(pre + W);  // [64] (2W words: 8-byte aligned)
    uint16_t *blk_s = reinterpret_cast<uint16_t *>(qm + kClT);       // [64 * NB]
    __shared__ int bad_s, off_s, warp_tot[kPlanThreads / 32];
    const int tid = threadIdx.x, tile = blockIdx.x, b = blockIdx.y, n_tiles = gridDim.x;
    const int t0 = tile * kClT, ti = min(kClT, N - t0);
    const int64_t bN = int64_t(b) * N;
    for (int i = tid; i < W; i += kPlanThreads) bitmap[i] = 0u;
    if (tid < kClT) qm[tid] = 0ull;
    if (tid == 0) { bad_s = 0; off_s = 0; }
    __syncthreads();
    int bad = 0, max_off = 0;
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
        if (mask && mask[so] == 0) {
            // a block that lies entirely in the padding of the last cluster (clusters of 16, 24 ... rows): all 8 slots
            // masked; it joins no union and its slots sort behind the token's real blocks
            for (int r = 1; r < 8; ++r)
                if (mask[so + r] != 0) bad = 1;
            blk = kClAbsent;
        } else if (base < 0 || base >= N || (base & 7) != 0) {
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
                    max_off = max(max_off, max(abs(int(yu) - int(yi)), abs(int(xu) - int(xi))));
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
    if (max_off > 0) atomicMax(&off_s, max_off);
    __syncthreads();
    uint16_t *ul = ucl + (int64_t(b) * n_tiles + tile) * kClUMax;
    int U = cl_union_list(bitmap, pre, W, warp_tot, ul), U2 = 0;
    bool split = false;
    if (U > kClUMax && hdr2 != nullptr && ti > kClT / 2) {
        // more than 64 blocks for the 64 tokens (mixed-resolution token sets): one union per half of 32 tokens
        split = true;
        for (int half = 0; half < 2; ++half) {
            const int ta = half * (kClT / 2), tb = min(ti, ta + kClT / 2);
            __syncthreads();
            for (int i = tid; i < W; i += kPlanThreads) bitmap[i] = 0u;
            __syncthreads();
            for (int it = ta * NB + tid; it < tb * NB; it += kPlanThreads) {
                const int blk = blk_s[it];
                if (blk != kClAbsent) atomicOr(&bitmap[blk >> 5], 1u << (blk & 31));
            }
            __syncthreads();
            const int Uh = cl_union_list(bitmap, pre, W, warp_tot,
                                         half == 0 ? ul : ucl2 + (int64_t(b) * n_tiles + tile) * kClUMax);
            if (Uh <= kClUMax) cl_membership(bitmap, pre, blk_s, qm, &bad_s, ta, tb, NB);
            if (half == 0) U = Uh; else U2 = Uh;
        }
    } else if (U <= kClUMax) {
        cl_membership(bitmap, pre, blk_s, qm, &bad_s, 0, ti, NB);
    }
    __syncthreads();
    for (int t = tid; t < kClT; t += kPlanThreads) qmask[bP + t0 + t] = t < ti ? qm[t] : 0ull;
    if (tid == 0) {
        const bool over = U > kClUMax || U2 > kClUMax;
        const bool isbad = bad_s != 0;
        hdr[int64_t(b) * n_tiles + tile] = U | (over ? AFFB200_PLAN_OVERFLOW : 0) | (isbad ? AFFB200_CL_BAD : 0) |
                                           (split ? AFFB200_CL_SPLIT : 0);
        if (hdr2) hdr2[int64_t(b) * n_tiles + tile] = U2;
        if (over) atomicAdd(&stats[0], 1);
        if (isbad) atomicAdd(&stats[1], 1);
        if (split) atomicAdd(&stats[3], 1);
        atomicMax(&stats[4], max(U, U2));
        atomicMax(&stats[6], off_s);     // largest |dx| or |dy| of a (token, key) pair: bounds the table rows in use
    }
}

// One CTA per (tile of 8 key blocks, image): every query token's block list is scanned once, in token order, and the
// token is appended to the list of each of the tile's blocks it references (ordered compaction by ballots), so every
// list is ascending — the summation order of the dK / dV kernel.
constexpr int kKpThreads = 256;
__global__ void __launch_bounds__(kKpThreads)
cl_kplan_kernel(const uint16_t *__restrict__ blk, int N, int NB, int kbcap, int32_t *__restrict__ kb_cnt,
                uint32_t *__restrict__ kb_tok, int32_t *__restrict__ stats) {
    __shared__ int warp_tot[kKpThreads / 32][8];
    __shared__ int base_s[8];
    const int tid = threadIdx.x, kt = blockIdx.x, b = blockIdx.y;
    const int lane = tid & 31, warp = tid >> 5, nblk = (N + 7) >> 3;
    if (tid < 8) base_s[tid] = 0;
    __syncthreads();
    for (int i0 = 0; i0 < N; i0 += kKpThreads) {
        const int i = i0 + tid;
        uint32_t bits = 0u;
        if (i < N) {
            const uint16_t *bp = blk + (int64_t(b) * N + i) * NB;
            for (int q = 0; q < NB; ++q) {
                const int v = bp[q];
                if (v != kClAbsent && (v >> 3) == kt) bits |= 1u << (v & 7);
            }
        }
        uint32_t bal[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            bal[r] = __ballot_sync(0xffffffffu, (bits >> r) & 1u);
            if (lane == r) warp_tot[warp][r] = __popc(bal[r]);
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if ((bits >> r) & 1u) {
                int off = base_s[r];
                for (int k = 0; k < warp; ++k) off += warp_tot[k][r];
                off += __popc(bal[r] & ((1u << lane) - 1u));
                const int kb = kt * 8 + r;
                if (off < kbcap && kb < nblk) kb_tok[(int64_t(b) * nblk + kb) * kbcap + off] = uint32_t(i);
            }
        }
        __syncthreads();
        if (tid < 8) {
            int t = 0;
            for (int k = 0; k < kKpThreads / 32; ++k) t += warp_tot[k][tid];
            base_s[tid] += t;
        }
        __syncthreads();
    }
    if (tid < 8 && kt * 8 + tid < nblk) {
        const int U = base_s[tid];
        kb_cnt[int64_t(b) * nblk + kt * 8 + tid] = U;
        if (U > kbcap) atomicAdd(&stats[2], 1);
        atomicMax(&stats[5], U);
    }
}

// packed bias indices in ascending-block slot order (the order fused_cl_bwd writes dS in)
__global__ void __launch_bounds__(256)
cl_pack_pe_sorted_kernel(const int64_t *__restrict__ pe_idx, const int64_t *__restrict__ mask, const uint16_t *__restrict__ blk,
                         int64_t total, int NB, uint16_t *__restrict__ out, int32_t *__restrict__ bad) {
    const int64_t t = int64_t(blockIdx.x) * 256 + threadIdx.x;   // (token, group of 8 slots)
    if (t >= total) return;
    const int64_t tokn = t / NB;
    const int q = int(t - tokn * NB);
    const uint16_t *bp = blk + tokn * NB;
    const int mine = bp[q];
    int rank = 0;
    for (int k = 0; k < NB; ++k) rank += (bp[k] < mine || (bp[k] == mine && k < q)) ? 1 : 0;
    const int64_t src = t * 8, dst = (tokn * NB + rank) * 8;
    uint16_t v[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int64_t pe = __ldg(pe_idx + src + r);
        const bool masked = mask && __ldg(mask + src + r) == 0;
        if (pe < 0 || pe >= 32768) { if (!masked) atomicAdd(bad, 1); v[r] = masked ? 0x8000u : 0u; }
        else v[r] = uint16_t(pe) | (masked ? 0x8000u : 0u);
    }
    *reinterpret_cast<uint4 *>(out + dst) = *reinterpret_cast<const uint4 *>(v);
}

}  // namespace affb200

using namespace affb200;

extern "C" int64_t affb200_cl_tiles(int64_t n) { return n <= 0 ? 0 : (n + kClT - 1) / kClT; }

extern "C" int affb200_cl_qplan_build_split(const int64_t *nbhd_idx, const int64_t *mask, const int64_t *pe_idx,
                                            const float *pos, int64_t B, int64_t N, int64_t M, int64_t rel_pos_width,
                                            int64_t table_width, uint16_t *blk, int32_t *hdr, uint16_t *ucl,
                                            int32_t *hdr2, uint16_t *ucl2, uint64_t *qmask, uint32_t *pos16,
                                            int32_t *stats, void *stream) {
    if (B <= 0 || B > 65535 || N <= 0 || N > 32767 * 8 || M <= 0 || M % 8 != 0 || M > 8 * kClUMax
        || rel_pos_width < 0 || table_width != 2 * rel_pos_width + 1) {
        set_error("cl_qplan_build: unsupported dimensions (B=%lld N=%lld M=%lld w=%lld tw=%lld)", (long long)B,
                  (long long)N, (long long)M, (long long)rel_pos_width, (long long)table_width);
        return AFFB200_E_ARG;
    }
    if (!nbhd_idx || !pe_idx || !pos || !blk || !hdr || !ucl || !qmask || !pos16 || !stats || (hdr2 != nullptr) != (ucl2 != nullptr)) {
        set_error("cl_qplan_build: null pointer");
        return AFFB200_E_ARG;
    }
    const int nblk = int((N + 7) / 8), W = (nblk + 31) / 32, NB = int(M / 8);
    const size_t smem = size_t(2 * W) * 4 + size_t(kClT) * 8 + size_t(kClT) * NB * 2;
    const dim3 grid{unsigned((N + kClT - 1) / kClT), unsigned(B), 1u};
    cl_qplan_kernel<<<grid, kPlanThreads, smem, (cudaStream_t)stream>>>(
        nbhd_idx, mask, pe_idx, pos, int(N), int(M), int(rel_pos_width), int(table_width), blk, hdr, ucl, hdr2, ucl2,
        reinterpret_cast<unsigned long long *>(qmask), pos16, stats);
    return check_launch("cl_qplan_build");
}

extern "C" int affb200_cl_qplan_build(const int64_t *nbhd_idx, const int64_t *mask, const int64_t *pe_idx,
                                      const float *pos, int64_t B, int64_t N, int64_t M, int64_t rel_pos_width,
                                      int64_t table_width, uint16_t *blk, int32_t *hdr, uint16_t *ucl, uint64_t *qmask,
                                      uint32_t *pos16, int32_t *stats, void *stream) {
    return affb200_cl_qplan_build_split(nbhd_idx, mask, pe_idx, pos, B, N, M, rel_pos_width, table_width, blk, hdr, ucl,
                                        nullptr, nullptr, qmask, pos16, stats, stream);
}

extern "C" int affb200_cl_kplan_build(const uint16_t *blk, int64_t B, int64_t N, int64_t M, int64_t kbcap,
                                      int32_t *kb_cnt, uint32_t *kb_tok, int32_t *stats, void *stream) {
    if (B <= 0 || B > 65535 || N <= 0 || N >= (1 << 24) || M <= 0 || M % 8 != 0 || kbcap <= 0) {
        set_error("cl_kplan_build: unsupported dimensions");
        return AFFB200_E_ARG;
    }
    if (!blk || !kb_cnt || !kb_tok || !stats) { set_error("cl_kplan_build: null pointer"); return AFFB200_E_ARG; }
    const dim3 grid{unsigned((N + kClT - 1) / kClT), unsigned(B), 1u};
    cl_kplan_kernel<<<grid, kKpThreads, 0, (cudaStream_t)stream>>>(blk, int(N), int(M / 8), int(kbcap), kb_cnt, kb_tok, stats);
    return check_launch("cl_kplan_build");
}

extern "C" int affb200_cl_pack_pe_sorted(const int64_t *pe_idx, const int64_t *mask, const uint16_t *blk, int64_t B,
                                         int64_t N, int64_t M, uint16_t *out, int32_t *bad, void *stream) {
    if (B <= 0 || N <= 0 || M <= 0 || M % 8 != 0 || !pe_idx || !blk || !out || !bad) {
        set_error("cl_pack_pe_sorted: bad arguments");
        return AFFB200_E_ARG;
    }
    const int64_t total = B * N * (M / 8);
    cl_pack_pe_sorted_kernel<<<unsigned((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(pe_idx, mask, blk, total, int(M / 8),
                                                                                              out, bad);
    return check_launch("cl_pack_pe_sorted");
}
