// tileplan.cu — per-nbhd_idx "tile plans": the index pre-pass that lets the dense-tile kernels (clusten_dense.cu)
// treat a neighbourhood gather as a small dense contraction.
//
// Tokens are stored in space-filling-curve order (reference models/point_utils.py:158-159), so a tile of T
// consecutive query tokens references only a small UNION of key rows (the tokens of the ~6 nearest clusters of each
// token overlap heavily: reference models/aff_transformer.py:476-480).  The plan stores, per query tile,
//     q_rows[u]      the sorted distinct key rows the tile references (u < U <= 256)
//     q_loc[t][j]    uint8 position of nbhd_idx[b,t,j] inside q_rows
// and, for the scatter halves of the backward passes, per tile of T consecutive KEY rows,
//     k_utok[u]      the sorted distinct query tokens that reference any row of the tile (u < Uq <= 256)
//     k_rloc[u][j]   uint8 local key row (nbhd_idx[b,utok[u],j] - r0) or 0xFF when outside the tile.
// nbhd_idx is constant across the 2..18 blocks of a stage (aff_transformer.py:489-495), so one build is reused by
// every forward and backward CLUSTEN call of the stage; the compact forms replace the 8-byte indices on every call.
// Nothing here assumes the cluster structure — any index tensor with locality qualifies; tiles that do not
// (U > 256) are counted in `stats` and the host then routes the whole tensor to the generic kernels.
#include "common.cuh"

namespace affb200 {

constexpr int kPlanThreads = 256;
constexpr int kUMax = AFFB200_PLAN_UMAX;

__device__ __forceinline__ int block_exclusive_scan_256(int v, int *warp_buf, int &total) {
    // 256 threads; returns exclusive prefix of v, total = block sum.  warp_buf: 8 ints of shared memory.
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int s = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, s, d);
        if (lane >= d) s += t;
    }
    if (lane == 31) warp_buf[wid] = s;
    __syncthreads();
    int woff = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kPlanThreads / 32; ++w) {
        const int x = warp_buf[w];
        if (w < wid) woff += x;
        tot += x;
    }
    __syncthreads();
    total = tot;
    return woff + s - v;
}

// ---------------------------------------------------------------------------------------------------------
// query-tile plan.  grid (n_tiles, B).  Shared: bitmap[nw] | prefix[nw] | loc_s[T*M] bytes | small.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kPlanThreads)
qplan_kernel(const int64_t *__restrict__ idx, int Nq, int Nk, int M, int T, int32_t *__restrict__ hdr,
             int32_t *__restrict__ rows, uint8_t *__restrict__ loc, int32_t *__restrict__ stats) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nw = (Nk + 31) >> 5;
    uint32_t *bm = reinterpret_cast<uint32_t *>(smem_raw);
    int *pre = reinterpret_cast<int *>(bm + nw);
    uint8_t *loc_s = reinterpret_cast<uint8_t *>(pre + nw);
    __shared__ int warp_buf[8];
    __shared__ int dup_flag;

    const int b = blockIdx.y, tile = blockIdx.x, n_tiles = gridDim.x;
    const int t0 = tile * T;
    const int ti = min(T, Nq - t0);
    const int tid = threadIdx.x;
    const int n_ent = ti * M;
    const int64_t *ig = idx + (int64_t(b) * Nq + t0) * M;

    for (int w = tid; w < nw; w += kPlanThreads) bm[w] = 0u;
    if (tid == 0) dup_flag = 0;
    __syncthreads();
    for (int p = tid; p < n_ent; p += kPlanThreads) {
        const int r = int(__ldg(ig + p));
        atomicOr(&bm[r >> 5], 1u << (r & 31));
    }
    __syncthreads();
    int carry = 0;
    for (int base = 0; base < nw; base += kPlanThreads) {
        const int w = base + tid;
        const int v = w < nw ? __popc(bm[w]) : 0;
        int total;
        const int ex = block_exclusive_scan_256(v, warp_buf, total);
        if (w < nw) pre[w] = carry + ex;
        carry += total;
    }
    __syncthreads();
    const int U = carry;
    int32_t *h = hdr + int64_t(b) * n_tiles + tile;
    if (U > kUMax) {
        if (tid == 0) {
            *h = U | AFFB200_PLAN_OVERFLOW;
            atomicAdd(stats + 0, 1);
        }
        return;
    }
    int32_t *rg = rows + (int64_t(b) * n_tiles + tile) * kUMax;
    for (int w = tid; w < nw; w += kPlanThreads) {
        uint32_t bits = bm[w];
        int o = pre[w];
        while (bits) {
            const int bit = __ffs(bits) - 1;
            bits &= bits - 1;
            rg[o++] = (w << 5) + bit;
        }
    }
    // pad the row list to a multiple of 16 with the last valid row (the dense kernels stage whole MMA k-steps)
    const int Upad = (U + 15) & ~15;
    if (U > 0 && tid < Upad - U) rg[U + tid] = -1;
    for (int p = tid; p < n_ent; p += kPlanThreads) {
        const int r = int(__ldg(ig + p));
        const uint32_t word = bm[r >> 5];
        loc_s[p] = uint8_t(pre[r >> 5] + __popc(word & ((1u << (r & 31)) - 1u)));
    }
    __syncthreads();
    // duplicate neighbours inside one token's list (e.g. the padded slots of the last cluster, which all point at
    // row 0: point_utils.py:204) — the wsum-type kernels then build their operand row sequentially
    if (tid < ti) {
        uint32_t seen[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        bool dup = false;
        const uint8_t *lr = loc_s + tid * M;
        for (int j = 0; j < M; ++j) {
            const int v = lr[j];
            const uint32_t bit = 1u << (v & 31);
#pragma unroll
            for (int q = 0; q < 8; ++q)
                if (q == (v >> 5)) {
                    dup |= (seen[q] & bit) != 0;
                    seen[q] |= bit;
                }
        }
        if (dup) dup_flag = 1;
    }
    __syncthreads();
    uint8_t *lg = loc + (int64_t(b) * Nq + t0) * M;
    for (int p = tid; p < n_ent; p += kPlanThreads) lg[p] = loc_s[p];
    if (tid == 0) {
        *h = U | (dup_flag ? AFFB200_PLAN_DUP : 0);
        if (dup_flag) atomicAdd(stats + 1, 1);
        atomicMax(stats + 4, U);
    }
}

// ---------------------------------------------------------------------------------------------------------
// key-tile plan, pass A: every token emits itself once to each key tile its neighbours fall in.
// grid (ceil(Nq/TOK), B); TOK tokens per block; thread per (token, slot) decides "first slot of this token that
// hits this key tile".
// ---------------------------------------------------------------------------------------------------------
constexpr int kEmitTok = 32;

__global__ void __launch_bounds__(kPlanThreads)
kplan_emit_kernel(const int64_t *__restrict__ idx, int Nq, int M, FastDiv divT, FastDiv divM, int nk_tiles,
                  int32_t *__restrict__ cnt, int32_t *__restrict__ utok) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int *kt_s = reinterpret_cast<int *>(smem_raw);  // [kEmitTok][MS]
    const int MS = M | 1;
    const int b = blockIdx.y;
    const int t0 = blockIdx.x * kEmitTok;
    const int ti = min(kEmitTok, Nq - t0);
    const int n_ent = ti * M;
    const int64_t *ig = idx + (int64_t(b) * Nq + t0) * M;
    for (int p = threadIdx.x; p < n_ent; p += kPlanThreads) {
        const int tl = int(divM.div(uint32_t(p)));
        kt_s[tl * MS + (p - tl * M)] = int(divT.div(uint32_t(__ldg(ig + p))));
    }
    __syncthreads();
    for (int p = threadIdx.x; p < n_ent; p += kPlanThreads) {
        const int tl = int(divM.div(uint32_t(p)));
        const int j = p - tl * M;
        const int *row = kt_s + tl * MS;
        const int kt = row[j];
        bool first = true;
        for (int q = 0; q < j; ++q) first &= (row[q] != kt);
        if (first) {
            const int64_t tile = int64_t(b) * nk_tiles + kt;
            const int slot = atomicAdd(cnt + tile, 1);
            if (slot < kUMax) utok[tile * kUMax + slot] = t0 + tl;
        }
    }
}

// pass B: grid (nk_tiles, B).  Sort the token list, then write rloc and the header.
__global__ void __launch_bounds__(kPlanThreads)
kplan_finish_kernel(const int64_t *__restrict__ idx, int Nq, int Nk, int M, int T, const int32_t *__restrict__ cnt,
                    int32_t *__restrict__ hdr, int32_t *__restrict__ utok, uint8_t *__restrict__ rloc,
                    int32_t *__restrict__ stats) {
    __shared__ int tok_s[kUMax];
    __shared__ int dup_flag;
    const int b = blockIdx.y, tile = blockIdx.x, nk_tiles = gridDim.x;
    const int64_t tix = int64_t(b) * nk_tiles + tile;
    const int n = cnt[tix];
    const int tid = threadIdx.x;
    if (n > kUMax) {
        if (tid == 0) {
            hdr[tix] = n | AFFB200_PLAN_OVERFLOW;
            atomicAdd(stats + 2, 1);
        }
        return;
    }
    static_assert(kUMax == kPlanThreads, "one list element per thread");
    int32_t *ug = utok + tix * kUMax;
    tok_s[tid] = tid < n ? ug[tid] : 0x7fffffff;
    if (tid == 0) dup_flag = 0;
    __syncthreads();
    // bitonic sort of 256 ints (ascending); entries are distinct
    for (int k = 2; k <= kUMax; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int u = tid ^ j;
            if (u > tid) {
                const int a = tok_s[tid], c = tok_s[u];
                const bool up = (tid & k) == 0;
                if ((a > c) == up) { tok_s[tid] = c; tok_s[u] = a; }
            }
            __syncthreads();
        }
    }
    const int npad = (n + 15) & ~15;
    if (tid < npad) ug[tid] = tid < n ? tok_s[tid] : -1;
    const int r0 = tile * T;
    const int r1 = min(r0 + T, Nk);
    uint8_t *rg = rloc + tix * int64_t(kUMax) * M;
    const int n_ent = n * M;
    for (int p = tid; p < n_ent; p += kPlanThreads) {
        const int u = p / M;
        const int j = p - u * M;
        const int r = int(__ldg(idx + (int64_t(b) * Nq + tok_s[u]) * M + j));
        rg[p] = (r >= r0 && r < r1) ? uint8_t(r - r0) : uint8_t(0xFF);
    }
    __syncthreads();
    // a token that lists the same key row twice (both inside this tile)
    if (tid < n) {
        unsigned long long seen0 = 0ull, seen1 = 0ull;
        bool dup = false;
        const int64_t *ir = idx + (int64_t(b) * Nq + tok_s[tid]) * M;
        for (int j = 0; j < M; ++j) {
            const int r = int(__ldg(ir + j));
            if (r >= r0 && r < r1) {
                const int v = r - r0;
                const unsigned long long bit = 1ull << (v & 63);
                if (v < 64) { dup |= (seen0 & bit) != 0; seen0 |= bit; }
                else        { dup |= (seen1 & bit) != 0; seen1 |= bit; }
            }
        }
        if (dup) dup_flag = 1;
    }
    __syncthreads();
    if (tid == 0) {
        hdr[tix] = n | (dup_flag ? AFFB200_PLAN_DUP : 0);
        if (dup_flag) atomicAdd(stats + 3, 1);
        atomicMax(stats + 5, n);
    }
}

}  // namespace affb200

using namespace affb200;

static bool plan_dims_ok(const char *what, int64_t B, int64_t Nq, int64_t Nk, int64_t M, int64_t T) {
    if (B < 0 || Nq < 0 || Nk < 0 || M < 0 || (T != 8 && T != 16 && T != 32 && T != 64) || B > 65535 ||
        Nq * M >= (int64_t(1) << 31) || Nk > AFFB200_PLAN_MAX_NK || M > AFFB200_PLAN_MAX_M) {
        set_error("%s: unsupported dimensions (T in {8,16,32,64}, Nk <= %d, M <= %d, B <= 65535)", what,
                  AFFB200_PLAN_MAX_NK, AFFB200_PLAN_MAX_M);
        return false;
    }
    return true;
}

extern "C" int64_t affb200_plan_tiles(int64_t n, int64_t T) { return T > 0 ? (n + T - 1) / T : 0; }

extern "C" int affb200_qplan_build(const int64_t *nbhd_idx, int64_t B, int64_t Nq, int64_t Nk, int64_t M, int64_t T,
                                   int32_t *hdr, int32_t *rows, uint8_t *loc, int32_t *stats, void *stream) {
    if (!plan_dims_ok("qplan_build", B, Nq, Nk, M, T)) return AFFB200_E_ARG;
    if (B * Nq * M == 0) return AFFB200_OK;
    if (!nbhd_idx || !hdr || !rows || !loc || !stats) {
        set_error("qplan_build: null pointer");
        return AFFB200_E_ARG;
    }
    const int nw = int((Nk + 31) >> 5);
    const size_t smem = size_t(nw) * 8 + size_t(T) * M;
    auto kern = qplan_kernel;
    if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    dim3 grid(unsigned(affb200_plan_tiles(Nq, T)), unsigned(B));
    kern<<<grid, kPlanThreads, smem, (cudaStream_t)stream>>>(nbhd_idx, int(Nq), int(Nk), int(M), int(T), hdr, rows,
                                                              loc, stats);
    return check_launch("qplan_build");
}

extern "C" int affb200_kplan_build(const int64_t *nbhd_idx, int64_t B, int64_t Nq, int64_t Nk, int64_t M, int64_t T,
                                   int32_t *hdr, int32_t *utok, uint8_t *rloc, int32_t *cnt_work, int32_t *stats,
                                   void *stream) {
    if (!plan_dims_ok("kplan_build", B, Nq, Nk, M, T)) return AFFB200_E_ARG;
    const int64_t nk_tiles = affb200_plan_tiles(Nk, T);
    if (B * nk_tiles == 0) return AFFB200_OK;
    if (!hdr || !utok || !rloc || !cnt_work || !stats || (!nbhd_idx && Nq * M > 0)) {
        set_error("kplan_build: null pointer");
        return AFFB200_E_ARG;
    }
    auto st = (cudaStream_t)stream;
    if (cudaMemsetAsync(cnt_work, 0, size_t(B * nk_tiles) * 4, st) != cudaSuccess) return check_launch("kplan memset");
    if (Nq * M > 0) {
        const size_t smem = size_t(kEmitTok) * (int(M) | 1) * 4;
        dim3 grid(unsigned((Nq + kEmitTok - 1) / kEmitTok), unsigned(B));
        kplan_emit_kernel<<<grid, kPlanThreads, smem, st>>>(nbhd_idx, int(Nq), int(M), FastDiv(uint32_t(T)),
                                                            FastDiv(uint32_t(M)), int(nk_tiles), cnt_work, utok);
    }
    dim3 grid2((unsigned)nk_tiles, (unsigned)B);
    kplan_finish_kernel<<<grid2, kPlanThreads, 0, st>>>(nbhd_idx, int(Nq), int(Nk), int(M), int(T), cnt_work, hdr,
                                                        utok, rloc, stats);
    return check_launch("kplan_build", Nq * M > 0 ? 2 : 1);
}
