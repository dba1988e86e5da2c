// clusten_dense.cu — dense-tile CLUSTEN kernels for f16/bf16 with head dim 16 or 32 (sm_100a).
//
// With a tile plan (tileplan.cu) a tile of T consecutive tokens and the UNION of the U <= 256 rows it touches turn
// the gather ops into small dense contractions on the tensor cores, the only place the path is a real contraction
// (per-cluster (tokens x head_dim)·(head_dim x neighbourhood) tiles):
//   gather-dot   S[T x U]  = X[T x c] · Y_U[U x c]^T ; out[t, j] = S[t, loc[t, j]]          QK fwd, AV bwd (d_attn)
//   gather-wsum  out[T x c] = P[T x U] · Y_U[U x c],  P[t, loc[t, j]] = w[t, j]              AV fwd, QK bwd (d_q)
//   scatter-wsum dY[T x c]  = P'[T x Uq] · X_Uq[Uq x c], P'[rloc[u, j], u] = w[utok[u], j]   QK bwd (d_k), AV bwd (d_v)
// Every union row is fetched from L2/HBM once per tile and head with 16-byte cp.async into padded shared memory
// (conflict-free ldmatrix), instead of once per (token, slot) as a row gather would (~6x fewer bytes through L1),
// the 8-byte indices are replaced by the plan's 1-byte positions, outputs leave as 16-byte streaming stores, and
// all sums are fp32 MMA accumulators combined in a fixed order (run-to-run deterministic, no atomics).
#include "common.cuh"

namespace affb200 {

constexpr int kDT = 256;              // threads per CTA (8 warps)
constexpr int kSS = kUMaxDense + 8;   // element stride of the [T][U] operand matrices: (2*kSS/16) is odd

// ---- PTX wrappers ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void cp_async16(void *dst, const void *src, bool valid) {
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(sz) : "memory");
}
// 16-bit store to a 32-bit shared-window address: a store through a generic pointer derived from the dynamic shared
// array makes ptxas rebuild the shared address per (predicated) store from SR_CgaCtaId
__device__ __forceinline__ void st_shared_u16(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"((unsigned short)v) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void *p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(p)));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const void *p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(p)));
}

template <typename T> __device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1);
template <> __device__ __forceinline__ void mma16816<__nv_bfloat16>(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <> __device__ __forceinline__ void mma16816<__half>(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <typename T> __device__ __forceinline__ uint32_t pack2(float lo, float hi);
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t *>(&v);
}
template <> __device__ __forceinline__ uint32_t pack2<__half>(float lo, float hi) {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t *>(&v);
}

// stage `n_rows` rows of C elements into padded shared memory rows; row_of(i) gives the global row or -1 (zero-fill)
template <typename T, int C, typename RowOf>
__device__ __forceinline__ void stage_rows(T *dst, const T *__restrict__ src_base, int n_rows, RowOf row_of,
                                           int64_t row_stride = C) {
    constexpr int CS = C + 8;
    constexpr int CH = C * 2 / 16;
    for (int q = threadIdx.x; q < n_rows * CH; q += kDT) {
        const int i = q / CH, ch = q - i * CH;
        const int r = row_of(i);
        cp_async16(dst + i * CS + ch * 8, src_base + int64_t(r < 0 ? 0 : r) * row_stride + ch * 8, r >= 0);
    }
}

// The same staging with the row offsets held in registers for the whole tile and the copies of one batch issued back
// to back (ptxas separates an LDGSTS from earlier shared-memory traffic with filler instructions, so per-chunk index
// loads between the copies cost more than the copies): thread owns chunk `ch` of rows r0, r0+RPP, ...
template <typename T, int C>
struct RowStager {
    static constexpr int CS = C + 8;
    static constexpr int CH = C / 8;
    static constexpr int RPP = kDT / CH;
    static constexpr int KMAX = kUMaxDense / RPP;
    int off[KMAX];   // element offset row*stride + ch*8, < 0: no copy (the row stays as the tile prologue left it)
    uint32_t dst0;

    __device__ __forceinline__ void init(const int *rows_s, int n_rows, int stride) {
        const int r0 = threadIdx.x / CH, ch = threadIdx.x % CH;
        dst0 = uint32_t(r0 * CS + ch * 8) * 2u;
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
            const int row = r0 + k * RPP;
            const int r = row < n_rows ? rows_s[row] : -1;
            off[k] = r >= 0 ? r * stride + ch * 8 : -1;
        }
    }
    __device__ __forceinline__ void issue(uint32_t buf_addr, const T *slab) const {
#pragma unroll
        for (int k = 0; k < KMAX; ++k)
            if (off[k] >= 0)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(buf_addr + dst0 + uint32_t(k * RPP * CS * 2)),
                             "l"(slab + off[k]) : "memory");
    }
};
// rows [r_lo, r_hi) of a padded staging buffer <- 0
template <typename T, int C>
__device__ __forceinline__ void zero_rows(T *buf, int r_lo, int r_hi) {
    constexpr int CS = C + 8, CH = C / 8;
    for (int q = threadIdx.x; q < (r_hi - r_lo) * CH; q += kDT)
        *reinterpret_cast<uint4 *>(buf + (r_lo + q / CH) * CS + (q % CH) * 8) = make_uint4(0u, 0u, 0u, 0u);
}

struct TileSmem {
    int *rows;        // [256]
    uint8_t *loc;     // [T*M] (padded to 16 bytes)
    void *xs;         // [T][C+8]          (gather-dot only)
    void *ys;         // [256][C+8]
    void *ps;         // [T][kSS]
    float *os;        // [8*16*(16+... )]  fp32 partial outputs (wsum kernels)
};

// strides (in elements) of a [B][H][N][C]-indexed tensor whose memory layout may differ (e.g. a [B][N][H][C] buffer)
struct StridedRows {
    int64_t sB, sH, sN;
};

template <int C> __host__ __device__ constexpr int os_floats() { return 8 * 16 * 16 * (C + 8) / C; }

// =========================================================================================================
// gather-dot
// =========================================================================================================
// FAST (decided by the launcher: M % 8 == 0, M <= 64, 16-byte aligned operands): the tile's 1-byte positions live in
// registers as groups of 8 slots (at most 2 groups per thread) for all heads instead of in shared memory.
// A tile that a speculated plan could not describe (see plan.py, static mode) is skipped; its outputs are written as
// zeros so that the rest of the step computes on defined values until the host's deferred check redoes the step.
template <typename T>
__device__ __forceinline__ void zero_span(T *p, int64_t n) {
    for (int64_t e = threadIdx.x; e < n; e += blockDim.x) p[e] = Elem<T>::from_acc(0.f);
}

template <typename T, int C, bool FAST>
__global__ void __launch_bounds__(kDT)
gather_dot_dense_kernel(const T *__restrict__ x, const T *__restrict__ y, const int32_t *__restrict__ q_hdr,
                        const int32_t *__restrict__ q_rows, const uint8_t *__restrict__ q_loc, T *__restrict__ out,
                        int H, int Nq, int Nk, int M, int TT, int TS) {
    // TS tokens per tile (the plan's tile size), TT >= TS rows of capacity (a multiple of the 16-row MMA tile)
    constexpr int CS = C + 8;
    constexpr int CH = C / 8;
    unsigned char *const smem_raw = affb200_dsm;
    int *rows_s = reinterpret_cast<int *>(smem_raw);
    T *ys = reinterpret_cast<T *>(smem_raw + 1024);
    T *ss = ys + kUMaxDense * CS;
    T *xs = ss + TT * kSS;
    uint8_t *loc_s = reinterpret_cast<uint8_t *>(xs + TT * CS);

    const int b = blockIdx.y, tile = blockIdx.x, n_tiles = gridDim.x;
    const int t0 = tile * TS;
    const int ti = min(TS, Nq - t0);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int hdr0 = q_hdr[int64_t(b) * n_tiles + tile];
    if (hdr0 & AFFB200_PLAN_OVERFLOW) {  // only reachable with a speculated plan; the caller's deferred check reports it
        for (int h = 0; h < H; ++h) zero_span<T>(out + ((int64_t(b) * H + h) * Nq + t0) * M, int64_t(ti) * M);
        return;
    }
    const int U = hdr0 & AFFB200_PLAN_COUNT_MASK;
    const int Upad = (U + 15) & ~15;
    const int n_ent = ti * M;

    const int32_t *rg = q_rows + (int64_t(b) * n_tiles + tile) * kUMaxDense;
    for (int u = tid; u < Upad; u += kDT) rows_s[u] = u < U ? rg[u] : -1;
    const uint8_t *lg = q_loc + (int64_t(b) * Nq + t0) * M;
    uint2 lc[2];
    int srow[2];
    if constexpr (FAST) {
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int p = (tid + k * kDT) * 8;
            lc[k] = make_uint2(0u, 0u);
            srow[k] = -1;
            if (p < n_ent) {
                lc[k] = __ldg(reinterpret_cast<const uint2 *>(lg + p));
                srow[k] = (p / M) * kSS;
            }
        }
    } else {
        if ((reinterpret_cast<uintptr_t>(lg) & 15) == 0) {
            for (int p = tid; p < (n_ent + 15) / 16; p += kDT)
                reinterpret_cast<uint4 *>(loc_s)[p] = __ldg(reinterpret_cast<const uint4 *>(lg) + p);  // tail over-read stays inside the [B][Nq][M] array + 16-byte slack kept by the host
        } else {
            for (int p = tid; p < n_ent; p += kDT) loc_s[p] = lg[p];
        }
    }
    zero_rows<T, C>(ys, U, Upad);
    zero_rows<T, C>(xs, ti, TT);
    __syncthreads();
    RowStager<T, C> st;
    st.init(rows_s, U, C);
    const uint32_t ys_a = smem_u32(ys), xs_a = smem_u32(xs);
    const int xrow = tid / CH, xch = tid % CH;
    const bool xok = tid < TT * CH && xrow < ti;
    auto issue = [&](int h) {
        const int64_t z = int64_t(b) * H + h;
        if (xok)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(xs_a + uint32_t(xrow * CS + xch * 8) * 2u),
                         "l"(x + (z * Nq + t0 + xrow) * C + xch * 8) : "memory");
        st.issue(ys_a, y + z * Nk * C);
        cp_async_commit();
    };
    issue(0);

    const int MT = TT >> 4;
    const int mt = warp % MT;
    const int nstep = 8 / MT;  // warps sharing one m-tile
    const int g = lane >> 2, q4 = lane & 3;

    for (int h = 0; h < H; ++h) {
        const int64_t z = int64_t(b) * H + h;
        cp_async_wait_all();
        __syncthreads();

        // A fragments of this warp's m-tile
        uint32_t a[C / 16][4];
#pragma unroll
        for (int ks = 0; ks < C / 16; ++ks)
            ldsm_x4(a[ks], xs + (mt * 16 + (lane & 15)) * CS + ks * 16 + (lane >> 4) * 8);
        if constexpr (C == 32) {
            for (int nt = warp / MT; nt < Upad / 8; nt += nstep) {
                uint32_t bf[4];  // (k 0-7, 8-15, 16-23, 24-31) of keys nt*8 .. nt*8+7
                ldsm_x4(bf, ys + (nt * 8 + (lane & 7)) * CS + (lane >> 3) * 8);
                float d[4] = {0.f, 0.f, 0.f, 0.f};
                mma16816<T>(d, a[0], bf[0], bf[1]);
                mma16816<T>(d, a[1], bf[2], bf[3]);
                T *sp = ss + (mt * 16 + g) * kSS + nt * 8 + q4 * 2;
                *reinterpret_cast<uint32_t *>(sp) = pack2<T>(d[0], d[1]);
                *reinterpret_cast<uint32_t *>(sp + 8 * kSS) = pack2<T>(d[2], d[3]);
            }
        } else {
            for (int np = warp / MT; np < Upad / 16; np += nstep) {
                uint32_t bf[4];  // keys np*16..+7: k 0-7, 8-15 ; keys np*16+8..+15: k 0-7, 8-15
                ldsm_x4(bf, ys + (np * 16 + (lane & 7) + (lane >> 4) * 8) * CS + ((lane >> 3) & 1) * 8);
                float d0[4] = {0.f, 0.f, 0.f, 0.f}, d1[4] = {0.f, 0.f, 0.f, 0.f};
                mma16816<T>(d0, a[0], bf[0], bf[1]);
                mma16816<T>(d1, a[0], bf[2], bf[3]);
                T *sp = ss + (mt * 16 + g) * kSS + np * 16 + q4 * 2;
                *reinterpret_cast<uint32_t *>(sp) = pack2<T>(d0[0], d0[1]);
                *reinterpret_cast<uint32_t *>(sp + 8 * kSS) = pack2<T>(d0[2], d0[3]);
                *reinterpret_cast<uint32_t *>(sp + 8) = pack2<T>(d1[0], d1[1]);
                *reinterpret_cast<uint32_t *>(sp + 8 * kSS + 8) = pack2<T>(d1[2], d1[3]);
            }
        }
        __syncthreads();
        if (h + 1 < H) issue(h + 1);  // the staging buffers are free: the next head loads while this one is written out

        T *og = out + (z * Nq + t0) * M;
        if constexpr (FAST) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                if (srow[k] < 0) continue;
                const uint16_t *sr = reinterpret_cast<const uint16_t *>(ss + srow[k]);
                uint32_t w4[4];
#pragma unroll
                for (int e = 0; e < 8; e += 2) {
                    const uint32_t l0 = ((e < 4 ? lc[k].x : lc[k].y) >> (8 * (e & 3))) & 0xffu;
                    const uint32_t l1 = ((e < 4 ? lc[k].x : lc[k].y) >> (8 * ((e + 1) & 3))) & 0xffu;
                    w4[e >> 1] = uint32_t(sr[l0]) | (uint32_t(sr[l1]) << 16);
                }
                st_cs_u4(og + (tid + k * kDT) * 8, make_uint4(w4[0], w4[1], w4[2], w4[3]));
            }
        } else {
            for (int p = tid; p < n_ent; p += kDT) og[p] = ss[(p / M) * kSS + loc_s[p]];
        }
    }
}

// =========================================================================================================
// shared tail of the two wsum kernels: out[T x C] = A[T x Kpad] (smem, stride kSS) · Bm[Kpad x C] (smem, stride C+8)
// K is split over the warps that share an output item; partials meet in `os` and are summed in slice order.
// =========================================================================================================
template <typename T, int C>
__device__ __forceinline__ void wsum_mma_store(const T *ps, const T *bs, float *os, T *__restrict__ og, int TT,
                                                int Kpad, int n_valid_rows, int64_t out_stride = C, float alpha = 1.f) {
    constexpr int CS = C + 8;
    constexpr int OS = C + 8;  // fp32 row stride of the partial-output buffer
    constexpr int NP = C / 16;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int MT = TT >> 4;
    const int items = MT * NP;
    const int nslices = 8 / items;
    const int item = warp % items, slice = warp / items;
    const int mt = item % MT, np = item / MT;
    const int g = lane >> 2, q4 = lane & 3;

    float d0[4] = {0.f, 0.f, 0.f, 0.f}, d1[4] = {0.f, 0.f, 0.f, 0.f};
    for (int ks = slice; ks < Kpad / 16; ks += nslices) {
        uint32_t a[4], bf[4];
        ldsm_x4(a, ps + (mt * 16 + (lane & 15)) * kSS + ks * 16 + (lane >> 4) * 8);
        // B stored [k][n]: matrices (k 0-7, n 0-7), (k 8-15, n 0-7), (k 0-7, n 8-15), (k 8-15, n 8-15), transposed on load
        ldsm_x4_t(bf, bs + (ks * 16 + (lane & 15)) * CS + np * 16 + (lane >> 4) * 8);
        mma16816<T>(d0, a, bf[0], bf[1]);
        mma16816<T>(d1, a, bf[2], bf[3]);
    }
    float *op = os + (slice * TT + mt * 16 + g) * OS + np * 16 + q4 * 2;
    *reinterpret_cast<float2 *>(op) = make_float2(d0[0], d0[1]);
    *reinterpret_cast<float2 *>(op + 8 * OS) = make_float2(d0[2], d0[3]);
    *reinterpret_cast<float2 *>(op + 8) = make_float2(d1[0], d1[1]);
    *reinterpret_cast<float2 *>(op + 8 * OS + 8) = make_float2(d1[2], d1[3]);
    __syncthreads();
    // 8 consecutive channels per thread -> one 16-byte store; the tile's rows are contiguous in global memory
    for (int e = tid; e < n_valid_rows * (C / 8); e += kDT) {
        const int row = e / (C / 8), c0 = (e - row * (C / 8)) * 8;
        float acc[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = 0.f;
        for (int s = 0; s < nslices; ++s) {
            const float4 v0 = *reinterpret_cast<const float4 *>(os + (s * TT + row) * OS + c0);
            const float4 v1 = *reinterpret_cast<const float4 *>(os + (s * TT + row) * OS + c0 + 4);
            acc[0] += v0.x; acc[1] += v0.y; acc[2] += v0.z; acc[3] += v0.w;
            acc[4] += v1.x; acc[5] += v1.y; acc[6] += v1.z; acc[7] += v1.w;
        }
        uint4 v;
        v.x = pack2<T>(acc[0] * alpha, acc[1] * alpha); v.y = pack2<T>(acc[2] * alpha, acc[3] * alpha);
        v.z = pack2<T>(acc[4] * alpha, acc[5] * alpha); v.w = pack2<T>(acc[6] * alpha, acc[7] * alpha);
        *reinterpret_cast<uint4 *>(og + int64_t(row) * out_stride + c0) = v;
    }
}

template <typename T>
__device__ __forceinline__ void zero_operand(T *ps, int n_rows, int Kpad) {
    const int chunks = Kpad / 8;  // 16-byte chunks per row
    for (int q = threadIdx.x; q < n_rows * chunks; q += kDT) {
        const int r = q / chunks, ch = q - r * chunks;
        *reinterpret_cast<uint4 *>(ps + r * kSS + ch * 8) = make_uint4(0u, 0u, 0u, 0u);
    }
}

// =========================================================================================================
// gather-wsum
// =========================================================================================================
template <typename T, int C, bool FAST>
__global__ void __launch_bounds__(kDT)
gather_wsum_dense_kernel(const T *__restrict__ w, const T *__restrict__ y, const int32_t *__restrict__ q_hdr,
                         const int32_t *__restrict__ q_rows, const uint8_t *__restrict__ q_loc, T *__restrict__ out,
                         int H, int Nq, int Nk, int M, int TT, int TS) {
    constexpr int CS = C + 8;
    unsigned char *const smem_raw = affb200_dsm;
    int *rows_s = reinterpret_cast<int *>(smem_raw);
    T *ys = reinterpret_cast<T *>(smem_raw + 1024);
    T *ps = ys + kUMaxDense * CS;
    float *os = reinterpret_cast<float *>(ps + TT * kSS);
    uint8_t *loc_s = reinterpret_cast<uint8_t *>(os + os_floats<C>());

    const int b = blockIdx.y, tile = blockIdx.x, n_tiles = gridDim.x;
    const int t0 = tile * TS;
    const int ti = min(TS, Nq - t0);
    const int tid = threadIdx.x;
    const int hdr = q_hdr[int64_t(b) * n_tiles + tile];
    if (hdr & AFFB200_PLAN_OVERFLOW) {  // speculated plan only (see gather_dot_dense_kernel)
        for (int h = 0; h < H; ++h) zero_span<T>(out + ((int64_t(b) * H + h) * Nq + t0) * C, int64_t(ti) * C);
        return;
    }
    const int U = hdr & AFFB200_PLAN_COUNT_MASK;
    const bool dup = (hdr & AFFB200_PLAN_DUP) != 0;
    const int Upad = (U + 15) & ~15;
    const int n_ent = ti * M;

    const int32_t *rg = q_rows + (int64_t(b) * n_tiles + tile) * kUMaxDense;
    for (int u = tid; u < Upad; u += kDT) rows_s[u] = u < U ? rg[u] : -1;
    const uint8_t *lg = q_loc + (int64_t(b) * Nq + t0) * M;
    const bool regs = FAST && !dup;  // slot groups in registers, operand zeroed once (every head writes the same entries)
    uint2 lc[2];
    int prow[2];
    if (regs) {
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int p = (tid + k * kDT) * 8;
            lc[k] = make_uint2(0u, 0u);
            prow[k] = -1;
            if (p < n_ent) {
                lc[k] = __ldg(reinterpret_cast<const uint2 *>(lg + p));
                prow[k] = (p / M) * kSS;
            }
        }
        zero_operand<T>(ps, TT, Upad);
    } else {
        if ((reinterpret_cast<uintptr_t>(lg) & 15) == 0) {
            for (int p = tid; p < (n_ent + 15) / 16; p += kDT)
                reinterpret_cast<uint4 *>(loc_s)[p] = __ldg(reinterpret_cast<const uint4 *>(lg) + p);
        } else {
            for (int p = tid; p < n_ent; p += kDT) loc_s[p] = lg[p];
        }
    }
    zero_rows<T, C>(ys, U, Upad);
    __syncthreads();
    RowStager<T, C> st;
    st.init(rows_s, U, C);
    const uint32_t ys_a = smem_u32(ys), ps_a = dsm_addr() + 1024u + uint32_t(kUMaxDense * CS * 2);

    for (int h = 0; h < H; ++h) {
        const int64_t z = int64_t(b) * H + h;
        const T *wg = w + (z * Nq + t0) * M;
        uint4 wv[2];
        if (regs) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                wv[k] = make_uint4(0u, 0u, 0u, 0u);
                if (prow[k] >= 0) wv[k] = __ldg(reinterpret_cast<const uint4 *>(wg) + tid + k * kDT);
            }
        }
        st.issue(ys_a, y + z * Nk * C);
        cp_async_commit();
        if (regs) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                if (prow[k] < 0) continue;
                const uint32_t pr = ps_a + uint32_t(prow[k]) * 2u;
                const uint32_t wr[4] = {wv[k].x, wv[k].y, wv[k].z, wv[k].w};
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const uint32_t l = ((e < 4 ? lc[k].x : lc[k].y) >> (8 * (e & 3))) & 0xffu;
                    st_shared_u16(pr + l * 2u, wr[e >> 1] >> (16 * (e & 1)));
                }
            }
        } else {
            zero_operand<T>(ps, TT, Upad);
            __syncthreads();
            if (!dup) {
                if (M % 8 == 0 && (reinterpret_cast<uintptr_t>(wg) & 15) == 0) {
                    for (int p8 = tid; p8 < n_ent / 8; p8 += kDT) {
                        const int p = p8 * 8;
                        const uint4 wv8 = __ldg(reinterpret_cast<const uint4 *>(wg + p));
                        const uint2 l2 = *reinterpret_cast<const uint2 *>(loc_s + p);
                        const uint16_t *e = reinterpret_cast<const uint16_t *>(&wv8);
                        const uint8_t *l8 = reinterpret_cast<const uint8_t *>(&l2);
                        uint16_t *pr = reinterpret_cast<uint16_t *>(ps + (p / M) * kSS);
#pragma unroll
                        for (int k = 0; k < 8; ++k) pr[l8[k]] = e[k];
                    }
                } else {
                    for (int p = tid; p < n_ent; p += kDT) ps[(p / M) * kSS + loc_s[p]] = wg[p];
                }
            } else if (tid < ti) {
                // a token lists a row more than once: accumulate that token's operand row in slot order
                T *pr = ps + tid * kSS;
                for (int j = 0; j < M; ++j) {
                    const int l = loc_s[tid * M + j];
                    pr[l] = Elem<T>::from_acc(Elem<T>::to_acc(pr[l]) + Elem<T>::to_acc(wg[tid * M + j]));
                }
            }
        }
        cp_async_wait_all();
        __syncthreads();
        wsum_mma_store<T, C>(ps, ys, os, out + (z * Nq + t0) * C, TT, Upad, ti);
        __syncthreads();
    }
}

// =========================================================================================================
// scatter-wsum over key tiles: dY[r0 + r, :] = sum_{u, j : rloc[u][j] == r} w[utok[u], j] * x[utok[u], :]
// =========================================================================================================
template <typename T, int C, int GPT>
__global__ void __launch_bounds__(kDT, GPT == 8 ? 2 : 3)
scatter_wsum_dense_kernel(const T *__restrict__ w, const T *__restrict__ x, const int32_t *__restrict__ k_hdr,
                          const int32_t *__restrict__ k_utok, const uint8_t *__restrict__ k_rloc,
                          T *__restrict__ dy, int H, int Nq, int Nk, int M, int TT, StridedRows xs_, StridedRows ys_,
                          float alpha) {
    constexpr int CS = C + 8;
    unsigned char *const smem_raw = affb200_dsm;
    int *tok_s = reinterpret_cast<int *>(smem_raw);
    T *xs = reinterpret_cast<T *>(smem_raw + 1024);
    T *ps = xs + kUMaxDense * CS;
    float *os = reinterpret_cast<float *>(ps + TT * kSS);

    const int b = blockIdx.y, tile = blockIdx.x, n_tiles = gridDim.x;
    const int HS = (H + int(gridDim.z) - 1) / int(gridDim.z);   // heads [h_lo, h_hi) of this CTA
    const int h_lo = int(blockIdx.z) * HS, h_hi = min(H, h_lo + HS);
    const int r0 = tile * TT;
    const int ri = min(TT, Nk - r0);
    const int tid = threadIdx.x;
    const int64_t tix = int64_t(b) * n_tiles + tile;
    const int hdr = k_hdr[tix];
    if (hdr & AFFB200_PLAN_OVERFLOW) {  // speculated plan only (see gather_dot_dense_kernel)
        for (int h = h_lo; h < h_hi; ++h)
            for (int e = tid; e < ri * C; e += kDT)
                dy[int64_t(b) * ys_.sB + int64_t(h) * ys_.sH + int64_t(r0 + e / C) * ys_.sN + e % C] = Elem<T>::from_acc(0.f);
        return;
    }
    const int Uq = hdr & AFFB200_PLAN_COUNT_MASK;
    const bool dup = (hdr & AFFB200_PLAN_DUP) != 0;
    const int Upad = (Uq + 15) & ~15;
    const int n_ent = Uq * M;

    const int32_t *ug = k_utok + tix * kUMaxDense;
    for (int u = tid; u < Upad; u += kDT) tok_s[u] = u < Uq ? ug[u] : -1;
    __syncthreads();
    const uint8_t *rl = k_rloc + tix * int64_t(kUMaxDense) * M;
    const T *xb = x + int64_t(b) * xs_.sB;
    T *yb = dy + int64_t(b) * ys_.sB + int64_t(r0) * ys_.sN;
    RowStager<T, C> st;
    st.init(tok_s, Uq, int(xs_.sN));
    zero_rows<T, C>(xs, Uq, Upad);
    const uint32_t xs_a = smem_u32(xs);

    // Fast path (GPT > 0: M <= 8*GPT, M % 8 == 0, Nq*M < 2^24, checked by the launcher): each thread keeps its GPT
    // groups of 8 slots — the 8 positions, and source offset | operand column << 24 — in registers for all heads.
    // rloc does not depend on the head, so neither does the set of operand entries that get written: the operand is
    // zeroed once per tile and every head overwrites the same entries.
    if constexpr (GPT > 0) {
        if (!dup) {
            uint2 lc[GPT];
            uint32_t src[GPT];
#pragma unroll
            for (int g = 0; g < GPT; ++g) {
                const int p = (tid + g * kDT) * 8;
                lc[g] = make_uint2(0xffffffffu, 0xffffffffu);
                src[g] = 0u;
                if (p < n_ent) {
                    const int u = p / M;
                    lc[g] = __ldg(reinterpret_cast<const uint2 *>(rl + p));
                    src[g] = uint32_t(tok_s[u] * M + (p - u * M)) | (uint32_t(u) << 24);
                }
            }
            zero_operand<T>(ps, TT, Upad);
            // shared address of ps formed from the array symbol itself (a link-time constant plus an offset)
            const uint32_t ps_a = dsm_addr() + 1024u + uint32_t(kUMaxDense * CS * 2);
            for (int h = h_lo; h < h_hi; ++h) {
                const T *wz = w + (int64_t(b) * H + h) * Nq * M;
                uint4 wv[GPT];
#pragma unroll
                for (int g = 0; g < GPT; ++g) {
                    wv[g] = make_uint4(0u, 0u, 0u, 0u);
                    if ((lc[g].x & lc[g].y) != 0xffffffffu)
                        wv[g] = __ldg(reinterpret_cast<const uint4 *>(wz + (src[g] & 0xffffffu)));
                }
                st.issue(xs_a, xb + int64_t(h) * xs_.sH);
                cp_async_commit();
                if (h == h_lo) __syncthreads();  // the zero fill above
                uint32_t hz;  // an opaque zero per head: keeps ptxas from hoisting the 48 store addresses out of the
                              // head loop into registers it does not have (248 bytes of spills otherwise)
                asm volatile("mov.u32 %0, 0;" : "=r"(hz));
#pragma unroll
                for (int g = 0; g < GPT; ++g) {
                    const uint32_t pcol = ps_a + hz + (src[g] >> 24) * 2u;
                    const uint32_t wr[4] = {wv[g].x, wv[g].y, wv[g].z, wv[g].w};
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const uint32_t l = ((k < 4 ? lc[g].x : lc[g].y) >> (8 * (k & 3))) & 0xffu;
                        if (l != 0xffu) st_shared_u16(pcol + l * uint32_t(kSS * 2), wr[k >> 1] >> (16 * (k & 1)));
                    }
                }
                cp_async_wait_all();
                __syncthreads();
                wsum_mma_store<T, C>(ps, xs, os, yb + int64_t(h) * ys_.sH, TT, Upad, ri, ys_.sN, alpha);
                __syncthreads();
            }
            return;
        }
    }

    for (int h = h_lo; h < h_hi; ++h) {
        const int64_t z = int64_t(b) * H + h;
        st.issue(xs_a, xb + int64_t(h) * xs_.sH);
        cp_async_commit();
        zero_operand<T>(ps, TT, Upad);
        __syncthreads();
        const T *wz = w + z * Nq * M;
        if (!dup) {
            if (M % 8 == 0 && (reinterpret_cast<uintptr_t>(wz) & 15) == 0 && (reinterpret_cast<uintptr_t>(rl) & 7) == 0) {
                for (int p8 = tid; p8 < n_ent / 8; p8 += kDT) {  // wide neighbourhoods (M > 64): groups do not fit registers
                    const int p = p8 * 8;
                    const int u = p / M, j = p - u * M;
                    const uint2 lc = __ldg(reinterpret_cast<const uint2 *>(rl + p));
                    if ((lc.x & lc.y) == 0xffffffffu) continue;
                    const uint4 wv = __ldg(reinterpret_cast<const uint4 *>(wz + int64_t(tok_s[u]) * M + j));
                    const uint16_t *e = reinterpret_cast<const uint16_t *>(&wv);
                    const uint8_t *l8 = reinterpret_cast<const uint8_t *>(&lc);
                    uint16_t *pcol = reinterpret_cast<uint16_t *>(ps) + u;
#pragma unroll
                    for (int k = 0; k < 8; ++k)
                        if (l8[k] != 0xFF) pcol[l8[k] * kSS] = e[k];
                }
            } else {
                for (int p = tid; p < n_ent; p += kDT) {
                    const int r = rl[p];
                    if (r != 0xFF) {
                        const int u = p / M;
                        ps[r * kSS + u] = wz[int64_t(tok_s[u]) * M + (p - u * M)];
                    }
                }
            }
        } else if (tid < Uq) {
            const T *wr = wz + int64_t(tok_s[tid]) * M;
            for (int j = 0; j < M; ++j) {
                const int r = rl[tid * M + j];
                if (r != 0xFF) {
                    T *pe = ps + r * kSS + tid;
                    *pe = Elem<T>::from_acc(Elem<T>::to_acc(*pe) + Elem<T>::to_acc(wr[j]));
                }
            }
        }
        cp_async_wait_all();
        __syncthreads();
        wsum_mma_store<T, C>(ps, xs, os, yb + int64_t(h) * ys_.sH, TT, Upad, ri, ys_.sN, alpha);
        __syncthreads();
    }
}

// =========================================================================================================
// launchers
// =========================================================================================================
static size_t smem_gather_dot(int C, int T, int M) {
    return 1024 + size_t(kUMaxDense) * (C + 8) * 2 + size_t(T) * kSS * 2 + size_t(T) * (C + 8) * 2 +
           ((size_t(T) * M + 15) & ~size_t(15));
}
static size_t smem_gather_wsum(int C, int T, int M) {
    return 1024 + size_t(kUMaxDense) * (C + 8) * 2 + size_t(T) * kSS * 2 + size_t(8 * 16 * 16 * (C + 8) / C) * 4 +
           ((size_t(T) * M + 15) & ~size_t(15));
}
static size_t smem_scatter_wsum(int C, int T) {
    return 1024 + size_t(kUMaxDense) * (C + 8) * 2 + size_t(T) * kSS * 2 + size_t(8 * 16 * 16 * (C + 8) / C) * 4;
}

template <typename K> static int set_smem(K kern, size_t smem) {
    if (smem > 48 * 1024 &&
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)) != cudaSuccess)
        return check_launch("cudaFuncSetAttribute");
    return AFFB200_OK;
}

template <typename T>
int launch_gather_dot_dense(const T *x, const T *y, const affb200_nbhd_plan *pl, T *out, int64_t B, int64_t H,
                            int64_t Nq, int64_t Nk, int64_t M, int64_t c, cudaStream_t st) {
    const int TS = pl->tq, TT = (TS + 15) & ~15;
    dim3 grid(unsigned((Nq + TS - 1) / TS), unsigned(B));
    const size_t smem = smem_gather_dot(int(c), TT, int(M));
    const bool fast = (M % 8 == 0) && M <= 64 && int64_t(TT) * M <= 4096 && aligned16(out) &&
                      (reinterpret_cast<uintptr_t>(pl->q_loc) & 7) == 0;
#define AFF_L(C_, F_)                                                                                           \
    do {                                                                                                        \
        auto kern = gather_dot_dense_kernel<T, C_, F_>;                                                         \
        if (int e = set_smem(kern, smem)) return e;                                                             \
        kern<<<grid, kDT, smem, st>>>(x, y, pl->q_hdr, pl->q_rows, pl->q_loc, out, int(H), int(Nq), int(Nk),    \
                                      int(M), TT, TS);                                                          \
    } while (0)
    if (c == 32) { if (fast) AFF_L(32, true); else AFF_L(32, false); }
    else { if (fast) AFF_L(16, true); else AFF_L(16, false); }
#undef AFF_L
    return check_launch("gather_dot_dense");
}

template <typename T>
int launch_gather_wsum_dense(const T *w, const T *y, const affb200_nbhd_plan *pl, T *out, int64_t B, int64_t H,
                             int64_t Nq, int64_t Nk, int64_t M, int64_t c, cudaStream_t st) {
    const int TS = pl->tq, TT = (TS + 15) & ~15;
    dim3 grid(unsigned((Nq + TS - 1) / TS), unsigned(B));
    const size_t smem = smem_gather_wsum(int(c), TT, int(M));
    const bool fast = (M % 8 == 0) && M <= 64 && int64_t(TT) * M <= 4096 && aligned16(w) &&
                      (reinterpret_cast<uintptr_t>(pl->q_loc) & 7) == 0;
#define AFF_L(C_, F_)                                                                                           \
    do {                                                                                                        \
        auto kern = gather_wsum_dense_kernel<T, C_, F_>;                                                        \
        if (int e = set_smem(kern, smem)) return e;                                                             \
        kern<<<grid, kDT, smem, st>>>(w, y, pl->q_hdr, pl->q_rows, pl->q_loc, out, int(H), int(Nq), int(Nk),    \
                                      int(M), TT, TS);                                                          \
    } while (0)
    if (c == 32) { if (fast) AFF_L(32, true); else AFF_L(32, false); }
    else { if (fast) AFF_L(16, true); else AFF_L(16, false); }
#undef AFF_L
    return check_launch("gather_wsum_dense");
}

template <typename T>
int launch_scatter_wsum_dense(const T *w, const T *x, const affb200_nbhd_plan *pl, T *dy, int64_t B, int64_t H,
                              int64_t Nq, int64_t Nk, int64_t M, int64_t c, cudaStream_t st) {
    const int64_t xs[3] = {H * Nq * c, Nq * c, c}, ys[3] = {H * Nk * c, Nk * c, c};
    return launch_scatter_wsum_dense_strided<T>(w, x, pl, dy, B, H, Nq, Nk, M, c, xs, ys, 1.f, st);
}

template <typename T>
int launch_scatter_wsum_dense_strided(const T *w, const T *x, const affb200_nbhd_plan *pl, T *dy, int64_t B, int64_t H,
                                      int64_t Nq, int64_t Nk, int64_t M, int64_t c, const int64_t *x_str,
                                      const int64_t *y_str, float alpha, cudaStream_t st) {
    const StridedRows xs_{x_str[0], x_str[1], x_str[2]}, ys_{y_str[0], y_str[1], y_str[2]};
    if (xs_.sN <= 0 || Nq * xs_.sN >= (int64_t(1) << 31) || Nq * M >= (int64_t(1) << 31)) {
        set_error("scatter_wsum_dense: source rows are addressed with 32-bit element offsets (Nq * row stride < 2^31)");
        return AFFB200_E_ARG;
    }
    const int TT = pl->tk;
    dim3 grid(unsigned((Nk + TT - 1) / TT), unsigned(B));
    // few key tiles: split the heads over gridDim.z (power-of-two divisor of H) up to about four waves
    while (int64_t(grid.z) * 2 <= H && H % (int64_t(grid.z) * 2) == 0 &&
           int64_t(grid.x) * grid.y * grid.z < int64_t(4) * kNumSMs * 3)
        grid.z *= 2;
    const size_t smem = smem_scatter_wsum(int(c), TT);
    // register-resident slot groups when they fit (see the kernel): 6 groups per thread cover M <= 48, 8 cover M <= 64
    const bool regs_ok = (M % 8 == 0) && Nq * M < (int64_t(1) << 24) && aligned16(w) &&
                         (reinterpret_cast<uintptr_t>(pl->k_rloc) & 7) == 0;
    const int gpt = !regs_ok ? 0 : (M <= 48 ? 6 : (M <= 64 ? 8 : 0));
#define AFF_L(C_, G_)                                                                                           \
    do {                                                                                                        \
        auto kern = scatter_wsum_dense_kernel<T, C_, G_>;                                                       \
        if (int e = set_smem(kern, smem)) return e;                                                             \
        kern<<<grid, kDT, smem, st>>>(w, x, pl->k_hdr, pl->k_utok, pl->k_rloc, dy, int(H), int(Nq), int(Nk),    \
                                      int(M), TT, xs_, ys_, alpha);                                             \
    } while (0)
#define AFF_LG(C_)                                                          \
    do {                                                                    \
        if (gpt == 6) AFF_L(C_, 6); else if (gpt == 8) AFF_L(C_, 8); else AFF_L(C_, 0); \
    } while (0)
    if (c == 32) AFF_LG(32); else AFF_LG(16);
#undef AFF_LG
#undef AFF_L
    return check_launch("scatter_wsum_dense");
}

#define AFF_INST(T)                                                                                                 \
    template int launch_gather_dot_dense<T>(const T *, const T *, const affb200_nbhd_plan *, T *, int64_t, int64_t,   \
                                            int64_t, int64_t, int64_t, int64_t, cudaStream_t);                        \
    template int launch_gather_wsum_dense<T>(const T *, const T *, const affb200_nbhd_plan *, T *, int64_t, int64_t,  \
                                             int64_t, int64_t, int64_t, int64_t, cudaStream_t);                       \
    template int launch_scatter_wsum_dense<T>(const T *, const T *, const affb200_nbhd_plan *, T *, int64_t, int64_t, \
                                              int64_t, int64_t, int64_t, int64_t, cudaStream_t);                      \
    template int launch_scatter_wsum_dense_strided<T>(const T *, const T *, const affb200_nbhd_plan *, T *, int64_t,  \
                                                      int64_t, int64_t, int64_t, int64_t, int64_t, const int64_t *,   \
                                                      const int64_t *, float, cudaStream_t);
AFF_INST(__half)
AFF_INST(__nv_bfloat16)
#undef AFF_INST

}  // namespace affb200
