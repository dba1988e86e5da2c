// fused_attn.cu — fused neighbourhood attention of ClusterAttention (reference models/aff_transformer.py:103-156):
//
//   S[i,j]   = scale * q_i · k_{nbhd[i,j]} + table[pe_idx[i,j], h] + (mask[i,j] ? 0 : -100)      j < M
//   S[i,M]   = scale * q_i · blank_k[h]                                                           (blank token)
//   P        = softmax over the M+1 logits
//   out_i    = sum_j P[i,j] v_{nbhd[i,j]} + P[i,M] blank_v[h]
//
// One kernel replaces QK -> bias gather/permute/add -> mask add -> blank logit -> cat -> softmax -> split -> AV ->
// blank_v add (8+ launches and 5-6 round trips of the [B,H,N,M] tensor).  It reads q/k/v straight out of the Linear
// outputs ([B,N,H*c] and [B,N,H,2,c]: no permute copies, no q*scale pass) and writes out as [B,N,H*c].
//
// Per CTA: a tile of T consecutive tokens (query-tile plan of tileplan.cu).  The union of the U <= 255 key rows the
// tile references is staged once per head in shared memory (K_U and V_U, 16-byte cp.async, padded rows), with the
// blank token appended as row U, so the blank logit and the blank_v term fall out of the same two tensor-core
// contractions:  S = Q·K_U^T (fp32, kept dense in shared memory), per-row softmax over the gathered entries,
// P scattered back into a dense bf16 tile that aliases S row by row, then O = P·V_U.
// Backward (fused_attn_bwd_kernel): recomputes S and P from the saved log-sum-exp, forms dP = dO·V_U^T,
// dS = P∘(dP − δ), dQ = scale·dS·K_U, and emits dS and P (bf16, [B,H,N,M]) for the key-tile scatter kernels
// (clusten_dense.cu) that produce dK and dV without atomics, and for the table-gradient kernel.
#include "common.cuh"
#include <type_traits>

namespace affb200 {

constexpr int kFT = 256;  // threads per CTA

__device__ __forceinline__ uint32_t f_smem_u32(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void f_cp_async16(void *dst, const void *src, bool valid) {
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(f_smem_u32(dst)), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void f_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void f_cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void f_ldsm_x4(uint32_t (&r)[4], const void *p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(f_smem_u32(p)));
}
__device__ __forceinline__ void f_ldsm_x4_t(uint32_t (&r)[4], const void *p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(f_smem_u32(p)));
}
template <typename T> __device__ __forceinline__ void f_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1);
template <> __device__ __forceinline__ void f_mma<__nv_bfloat16>(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <> __device__ __forceinline__ void f_mma<__half>(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <typename T> __device__ __forceinline__ uint32_t f_pack2(float lo, float hi);
template <> __device__ __forceinline__ uint32_t f_pack2<__nv_bfloat16>(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t *>(&v);
}
template <> __device__ __forceinline__ uint32_t f_pack2<__half>(float lo, float hi) {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t *>(&v);
}


// strided [B][N][H][c] view of a Linear output: element (b, i, h, 0) at base + b*sB + i*sN + h*sH
template <typename T> struct View {
    const T *p;
    int64_t sB, sN, sH;
    __device__ __forceinline__ const T *row(int64_t b, int64_t i, int h) const { return p + b * sB + i * sN + h * sH; }
};
template <typename T> struct ViewW {
    T *p;
    int64_t sB, sN, sH;
    __device__ __forceinline__ T *row(int64_t b, int64_t i, int h) const { return p + b * sB + i * sN + h * sH; }
};

// compact per-(token, slot) bias index: pe | (masked << 15) — built once per stage from pe_idx / cluster_mask
__global__ void __launch_bounds__(256)
pe_pack_kernel(const int64_t *__restrict__ pe_idx, const int64_t *__restrict__ mask, int64_t total,
               uint16_t *__restrict__ out, int32_t *__restrict__ bad) {
    const int64_t t = int64_t(blockIdx.x) * 256 + threadIdx.x;
    if (t >= total) return;
    const int64_t p = __ldg(pe_idx + t);
    if (p < 0 || p >= 32768) { atomicAdd(bad, 1); out[t] = 0; return; }
    out[t] = uint16_t(p) | ((mask && __ldg(mask + t) == 0) ? 0x8000u : 0u);
}

// ---- shared-memory carve-up (all offsets multiples of 16 bytes) --------------------------------------------
// rows   union rows of the tile as element offsets into the k/v slab of one (batch, head)
// ks/vs  K_U / V_U of the current head (padded rows of C+8); the backward keeps two K buffers so that the next
//        head's keys load while dQ = dS·K_U still reads the current ones
// xs/dos/outs  the tile's own Q / dO / O rows
// sp     S (fp32) and, aliased row by row, P or dS (bf16);  dp  dP (fp32);  os  split-K partials of the output
//        contraction (the backward lays it over dp, which is dead by then);  bl  blank-gradient partials
struct FusedSmem {
    int UC;       // row capacity of the K/V staging buffers (multiple of 32, >= U+1)
    int RS;       // byte stride of an S/P row: UC*4 + 16  (== 16 mod 128: conflict-free ldmatrix on the bf16 view;
                  // the 4 spare floats hold the row's sentinel for unused slots)
    size_t rows, ks, ks2, vs, xs, dos, outs, sp, dp, os, bl, total;
};
__host__ __device__ inline FusedSmem fused_smem(int UC, int TT, int C, bool bwd) {
    FusedSmem s;
    s.UC = UC;
    s.RS = UC * 4 + 16;
    const size_t kv = size_t(UC) * (C + 8) * 2, tile = size_t(TT) * (C + 8) * 2, srow = size_t(TT) * s.RS;
    const size_t osz = size_t(8 * 16 * 16 * (C + 8) / C) * 4;
    size_t o = 0;
    s.rows = o; o += size_t(UC) * 4;
    s.ks = o;   o += kv;
    s.ks2 = o;  o += bwd ? kv : 0;
    s.vs = o;   o += kv;
    s.xs = o;   o += tile;
    s.dos = o;  o += bwd ? tile : 0;
    s.outs = o; o += bwd ? tile : 0;
    s.sp = o;   o += srow;
    if (bwd) {
        s.dp = o; s.os = o; o += srow > osz ? srow : osz;
    } else {
        s.dp = o; s.os = o; o += osz;
    }
    s.bl = o;   o += bwd ? size_t(8) * 2 * C * 4 : 0;
    s.total = o;
    return s;
}

// ---- staging ---------------------------------------------------------------------------------------------------
// 16-byte cp.async into padded rows of stride C+8.  ptxas separates an LDGSTS from earlier shared-memory traffic
// with three filler instructions, so the copies of one batch are issued back to back from offsets that already sit
// in registers: each thread owns chunk `ch` of rows r0, r0+RPP, r0+2·RPP, ... and keeps their element offsets for
// the whole tile (K and V of every head are gathered with the same offsets).
template <typename T, int C>
struct GatherStager {
    static constexpr int CS = C + 8;
    static constexpr int CH = C / 8;         // 16-byte chunks per row
    static constexpr int RPP = kFT / CH;     // rows per pass of the CTA
    static constexpr int KMAX = AFFB200_PLAN_UMAX / RPP;
    int off[KMAX];   // element offset of this thread's chunk (row offset + ch*8); < 0: nothing to copy
    uint32_t dst0;   // byte offset of the chunk inside a staging buffer, row r0

    __device__ __forceinline__ void init(const int *rows_s, int U) {
        const int r0 = threadIdx.x / CH, ch = threadIdx.x % CH;
        dst0 = uint32_t(r0 * CS + ch * 8) * 2u;
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
            const int row = r0 + k * RPP;
            off[k] = row < U ? rows_s[row] + ch * 8 : -1;
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

// the tile's own rows: row i from base + i*stride, i < n_valid (TT*CH <= 256 chunks: at most one per thread)
template <typename T, int C, int TT>
__device__ __forceinline__ void f_stage_tile(uint32_t buf_addr, const T *base, int stride, int n_valid) {
    constexpr int CS = C + 8, CH = C / 8;
    static_assert(TT * CH <= kFT, "one chunk per thread");
    const int row = threadIdx.x / CH, ch = threadIdx.x % CH;
    if (threadIdx.x < TT * CH && row < n_valid)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(buf_addr + uint32_t(row * CS + ch * 8) * 2u),
                     "l"(base + row * stride + ch * 8) : "memory");
}
template <int n> __device__ __forceinline__ void f_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(n) : "memory"); }

// rows [r_lo, r_hi) of a padded staging buffer <- 0 (the C payload elements of each row)
template <typename T, int C>
__device__ __forceinline__ void f_zero_rows(T *buf, int r_lo, int r_hi) {
    constexpr int CS = C + 8, CH = C / 8;
    for (int q = threadIdx.x; q < (r_hi - r_lo) * CH; q += kFT)
        *reinterpret_cast<uint4 *>(buf + (r_lo + q / CH) * CS + (q % CH) * 8) = make_uint4(0u, 0u, 0u, 0u);
}

// S[TT x Upad] (fp32, smem row stride RS bytes) = alpha * X[TT x C] · Y[Upad x C]^T
template <typename T, int C>
__device__ __forceinline__ void f_mma_xyT(const T *xs, const T *ys, unsigned char *sp, int RS, int TT, int Upad,
                                          float alpha) {
    constexpr int CS = C + 8;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int MT = TT >> 4;
    const int mt = warp % MT;
    const int nstep = 8 / MT;
    const int g = lane >> 2, q4 = lane & 3;
    uint32_t a[C / 16][4];
#pragma unroll
    for (int ks = 0; ks < C / 16; ++ks)
        f_ldsm_x4(a[ks], xs + (mt * 16 + (lane & 15)) * CS + ks * 16 + (lane >> 4) * 8);
    if constexpr (C == 32) {
        for (int nt = warp / MT; nt < Upad / 8; nt += nstep) {
            uint32_t bf[4];
            f_ldsm_x4(bf, ys + (nt * 8 + (lane & 7)) * CS + (lane >> 3) * 8);
            float d[4] = {0.f, 0.f, 0.f, 0.f};
            f_mma<T>(d, a[0], bf[0], bf[1]);
            f_mma<T>(d, a[1], bf[2], bf[3]);
            float *r0 = reinterpret_cast<float *>(sp + size_t(mt * 16 + g) * RS) + nt * 8 + q4 * 2;
            float *r1 = reinterpret_cast<float *>(sp + size_t(mt * 16 + g + 8) * RS) + nt * 8 + q4 * 2;
            *reinterpret_cast<float2 *>(r0) = make_float2(d[0] * alpha, d[1] * alpha);
            *reinterpret_cast<float2 *>(r1) = make_float2(d[2] * alpha, d[3] * alpha);
        }
    } else {
        for (int np = warp / MT; np < Upad / 16; np += nstep) {
            uint32_t bf[4];
            f_ldsm_x4(bf, ys + (np * 16 + (lane & 7) + (lane >> 4) * 8) * CS + ((lane >> 3) & 1) * 8);
            float d0[4] = {0.f, 0.f, 0.f, 0.f}, d1[4] = {0.f, 0.f, 0.f, 0.f};
            f_mma<T>(d0, a[0], bf[0], bf[1]);
            f_mma<T>(d1, a[0], bf[2], bf[3]);
            float *r0 = reinterpret_cast<float *>(sp + size_t(mt * 16 + g) * RS) + np * 16 + q4 * 2;
            float *r1 = reinterpret_cast<float *>(sp + size_t(mt * 16 + g + 8) * RS) + np * 16 + q4 * 2;
            *reinterpret_cast<float2 *>(r0) = make_float2(d0[0] * alpha, d0[1] * alpha);
            *reinterpret_cast<float2 *>(r1) = make_float2(d0[2] * alpha, d0[3] * alpha);
            *reinterpret_cast<float2 *>(r0 + 8) = make_float2(d1[0] * alpha, d1[1] * alpha);
            *reinterpret_cast<float2 *>(r1 + 8) = make_float2(d1[2] * alpha, d1[3] * alpha);
        }
    }
}

// out[TT x C] = alpha * A[TT x Kpad] (bf16/f16 view of the S/P buffer, row stride RS bytes) · Bm[Kpad x C]
// K is split over the warps that share an output item; partials meet in `os` and are summed in slice order.
template <typename T, int C>
__device__ __forceinline__ void f_mma_pv_store(const unsigned char *sp, int RS, const T *bs, float *os, int TT,
                                               int Kpad, int n_valid, float alpha, T *out_base, int64_t out_stride) {
    constexpr int CS = C + 8;
    constexpr int OS = C + 8;
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
        f_ldsm_x4(a, sp + size_t(mt * 16 + (lane & 15)) * RS + (ks * 16 + (lane >> 4) * 8) * 2);
        f_ldsm_x4_t(bf, bs + (ks * 16 + (lane & 15)) * CS + np * 16 + (lane >> 4) * 8);
        f_mma<T>(d0, a, bf[0], bf[1]);
        f_mma<T>(d1, a, bf[2], bf[3]);
    }
    float *op = os + (slice * TT + mt * 16 + g) * OS + np * 16 + q4 * 2;
    *reinterpret_cast<float2 *>(op) = make_float2(d0[0], d0[1]);
    *reinterpret_cast<float2 *>(op + 8 * OS) = make_float2(d0[2], d0[3]);
    *reinterpret_cast<float2 *>(op + 8) = make_float2(d1[0], d1[1]);
    *reinterpret_cast<float2 *>(op + 8 * OS + 8) = make_float2(d1[2], d1[3]);
    __syncthreads();
    for (int e = tid; e < n_valid * (C / 8); e += kFT) {
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
        v.x = f_pack2<T>(acc[0] * alpha, acc[1] * alpha); v.y = f_pack2<T>(acc[2] * alpha, acc[3] * alpha);
        v.z = f_pack2<T>(acc[4] * alpha, acc[5] * alpha); v.w = f_pack2<T>(acc[6] * alpha, acc[7] * alpha);
        *reinterpret_cast<uint4 *>(out_base + int64_t(row) * out_stride + c0) = v;
    }
}

// Row phase mapping: TPR = 256 / TT threads share one token row (a contiguous lane group).  Thread `sub` owns the
// PER = NE-1 consecutive neighbour slots j = sub*PER + e and, for sub == 0, the blank logit as slot NE-1; all TT rows
// of the tile are processed in one pass.  What a slot needs is packed once per tile into a 32-bit descriptor held
// in registers for every head:  bits 0-8 column of S (the slot's position in the union, or UC = the row's sentinel
// column for unused slots), bits 9-23 bias-table row, bit 24 masked slot (-100), bit 31 unused slot.
constexpr float kLog2e = 1.4426950408889634f;
constexpr uint32_t kSlotUnused = 0x80000000u;

struct TileHdr {
    int U, Upad, ti, t0;
    bool dup, bad;  // bad: overflowed tile or more rows than the launch sized its buffers for (speculated plans only)
};

// common prologue: tile header and union rows (as element offsets row*kv_stride, -1 above U) -> shared memory
__device__ __forceinline__ TileHdr f_load_tile(const int32_t *q_hdr, const int32_t *q_rows, int N, int TT, int UC,
                                               int kv_stride, int *rows_s) {
    TileHdr t;
    const int b = blockIdx.y, tile = blockIdx.x, n_tiles = gridDim.x, tid = threadIdx.x;
    t.t0 = tile * TT;
    t.ti = min(TT, N - t.t0);
    const int hdr = q_hdr[int64_t(b) * n_tiles + tile];
    t.U = hdr & AFFB200_PLAN_COUNT_MASK;
    t.bad = (hdr & AFFB200_PLAN_OVERFLOW) != 0 || t.U + 1 > UC;
    t.dup = false;
    t.Upad = 0;
    if (t.bad) return t;
    t.dup = (hdr & AFFB200_PLAN_DUP) != 0;
    t.Upad = (t.U + 1 + 15) & ~15;
    const int32_t *rg = q_rows + (int64_t(b) * n_tiles + tile) * AFFB200_PLAN_UMAX;
    for (int u = tid; u < UC; u += kFT) rows_s[u] = u < t.U ? rg[u] * kv_stride : -1;
    return t;
}

// slot descriptors of this thread, straight from the plan's 1-byte positions and the packed bias indices
template <int PER>
__device__ __forceinline__ void f_load_desc(const uint8_t *__restrict__ loc_r, const uint16_t *__restrict__ pe_r, int M,
                                            int sub, bool live, int UC, uint32_t (&desc)[PER]) {
    const int j0 = sub * PER;
    auto mk = [](uint32_t loc, uint32_t pe) { return loc | ((pe & 0x7fffu) << 9) | ((pe >> 15) << 24); };
    if ((PER % 2 == 0) && (M % 2 == 0)) {  // slots come in aligned pairs that are valid together
#pragma unroll
        for (int e = 0; e < PER; e += 2) {
            desc[e] = desc[e + 1] = kSlotUnused | uint32_t(UC);
            if (live && j0 + e < M) {
                const uint32_t l2 = __ldg(reinterpret_cast<const uint16_t *>(loc_r + j0 + e));
                const uint32_t p2 = __ldg(reinterpret_cast<const uint32_t *>(pe_r + j0 + e));
                desc[e] = mk(l2 & 0xffu, p2 & 0xffffu);
                desc[e + 1] = mk(l2 >> 8, p2 >> 16);
            }
        }
    } else {
#pragma unroll
        for (int e = 0; e < PER; ++e) {
            desc[e] = kSlotUnused | uint32_t(UC);
            if (live && j0 + e < M) desc[e] = mk(__ldg(loc_r + j0 + e), __ldg(pe_r + j0 + e));
        }
    }
}

__device__ __forceinline__ float f_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// n contiguous elements (n*sizeof(T) in {2,4,8,16} bytes, naturally aligned) -> fp32
template <typename T, int n> __device__ __forceinline__ void f_load_n(const T *p, float (&o)[n]) {
    static_assert(sizeof(T) == 2 && (n == 1 || n == 2 || n == 4 || n == 8), "2-byte elements");
    if constexpr (n == 1) {
        o[0] = Elem<T>::to_acc(*p);
    } else if constexpr (n == 2) {
        const uint32_t r = *reinterpret_cast<const uint32_t *>(p);
        const T *e = reinterpret_cast<const T *>(&r);
        o[0] = Elem<T>::to_acc(e[0]); o[1] = Elem<T>::to_acc(e[1]);
    } else if constexpr (n == 4) {
        const uint2 r = *reinterpret_cast<const uint2 *>(p);
        const T *e = reinterpret_cast<const T *>(&r);
#pragma unroll
        for (int k = 0; k < 4; ++k) o[k] = Elem<T>::to_acc(e[k]);
    } else {
        const uint4 r = *reinterpret_cast<const uint4 *>(p);
        const T *e = reinterpret_cast<const T *>(&r);
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = Elem<T>::to_acc(e[k]);
    }
}

template <int TPR> __device__ __forceinline__ float f_group_max(float v) {
#pragma unroll
    for (int o = TPR / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
template <int TPR> __device__ __forceinline__ float f_group_sum(float v) {
#pragma unroll
    for (int o = TPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// the M+1 logits of one token, gathered by its TPR threads: S[row][col] + bias (+ -100 at masked slots); the blank
// logit S[row][U] is slot NE-1 of the group's first thread.  Unused slots read the row's -inf sentinel.
template <int NE>
__device__ __forceinline__ void f_row_logits(const float *srow, const uint32_t (&desc)[NE - 1],
                                             const float *__restrict__ table_h, int H, int U, bool has_blank,
                                             float (&pv)[NE]) {
#pragma unroll
    for (int e = 0; e < NE - 1; ++e) {
        const uint32_t d = desc[e];
        pv[e] = srow[d & 511u] + __ldg(table_h + ((d >> 9) & 0x7fffu) * H) + ((d & (1u << 24)) ? -100.f : 0.f);
    }
    pv[NE - 1] = has_blank ? srow[U] : -INFINITY;
}

// softmax over the row's group; leaves the probabilities in pv[] and returns the log-sum-exp
template <int TPR, int NE>
__device__ __forceinline__ float f_row_softmax(bool live, float (&pv)[NE]) {
    float mx = -INFINITY;
#pragma unroll
    for (int e = 0; e < NE; ++e) mx = fmaxf(mx, pv[e]);
    mx = f_group_max<TPR>(mx);
    if (!live) mx = 0.f;
    float sum = 0.f;
#pragma unroll
    for (int e = 0; e < NE; ++e) {
        pv[e] = f_ex2((pv[e] - mx) * kLog2e);  // ex2(-inf) = 0 for the unused slots
        sum += pv[e];
    }
    sum = f_group_sum<TPR>(sum);
    const float inv = live ? __fdividef(1.f, sum) : 0.f;
#pragma unroll
    for (int e = 0; e < NE; ++e) pv[e] *= inv;
    return mx + __logf(sum);
}

// write a token's values into its dense operand row (bf16/f16 view at the start of its S row), zero elsewhere.
// Tiles in which a token lists a row twice (the padded slots of the last cluster all point at row 0) accumulate in
// fp32 with shared-memory atomics over the row's fp32 footprint and convert in place afterwards.
template <typename T, int TPR, int NE>
__device__ __forceinline__ void f_scatter_row(T *prow, const uint32_t (&desc)[NE - 1], int Upad, int U, bool dup,
                                              int sub, bool has_blank, const float (&pv)[NE]) {
    if (!dup) {
        for (int w = sub; w < Upad / 8; w += TPR) reinterpret_cast<uint4 *>(prow)[w] = make_uint4(0u, 0u, 0u, 0u);
        __syncwarp();
#pragma unroll
        for (int e = 0; e < NE - 1; ++e)
            if (int(desc[e]) >= 0) prow[desc[e] & 511u] = Elem<T>::from_acc(pv[e]);
        if (has_blank) prow[U] = Elem<T>::from_acc(pv[NE - 1]);
    } else {
        float *frow = reinterpret_cast<float *>(prow);
        for (int w = sub; w < Upad; w += TPR) frow[w] = 0.f;
        __syncwarp();
#pragma unroll
        for (int e = 0; e < NE - 1; ++e)
            if (int(desc[e]) >= 0) atomicAdd(frow + (desc[e] & 511u), pv[e]);
        if (has_blank) atomicAdd(frow + U, pv[NE - 1]);
        __syncwarp();
        for (int base = 0; base < Upad; base += TPR) {  // in-place fp32 -> T, front to back
            const float vv = frow[base + sub];
            __syncwarp();
            prow[base + sub] = Elem<T>::from_acc(vv);
            __syncwarp();
        }
    }
}

// blank token of head h -> row U of a K or V staging buffer
template <typename T, int C>
__device__ __forceinline__ void f_put_blank(T *buf, int U, const float *__restrict__ blank, int h) {
    if (threadIdx.x < C) buf[U * (C + 8) + threadIdx.x] = Elem<T>::from_acc(__ldg(blank + h * C + threadIdx.x));
}

// Forward.  Loads run one head ahead of the math: Q and K of head h+1 are requested as soon as S = Q·K^T of head h
// has left the staging buffers, V of head h+1 as soon as O = P·V of head h has; every wait is "all but the newest
// group" (cp.async groups retire in order: ... QK_h, V_h, QK_h+1, V_h+1 ...).
template <typename T, int C, int TT, int NE>
__global__ void __launch_bounds__(kFT, NE <= 7 ? 3 : 2)
fused_attn_fwd_kernel(View<T> q, View<T> k, View<T> v, ViewW<T> out, const float *__restrict__ table,
                      const float *__restrict__ blank_k, const float *__restrict__ blank_v,
                      const int32_t *__restrict__ q_hdr, const int32_t *__restrict__ q_rows,
                      const uint8_t *__restrict__ q_loc, const uint16_t *__restrict__ q_pe, float *__restrict__ lse,
                      int H, int N, int M, float scale, FusedSmem L) {
    constexpr int TPR = kFT / TT;
    extern __shared__ __align__(128) unsigned char smem[];
    int *rows_s = reinterpret_cast<int *>(smem + L.rows);
    T *ks = reinterpret_cast<T *>(smem + L.ks);
    T *vs = reinterpret_cast<T *>(smem + L.vs);
    T *xs = reinterpret_cast<T *>(smem + L.xs);
    unsigned char *sp = smem + L.sp;
    float *os = reinterpret_cast<float *>(smem + L.os);
    const uint32_t ks_a = f_smem_u32(ks), vs_a = f_smem_u32(vs), xs_a = f_smem_u32(xs);

    const int b = blockIdx.y, tid = threadIdx.x;
    const TileHdr t = f_load_tile(q_hdr, q_rows, N, TT, L.UC, int(k.sN), rows_s);
    if (t.bad) return;  // uniform over the CTA; the host's deferred plan check reports the failed speculation
    const int i = tid / TPR, sub = tid % TPR;  // this thread's token row and its position in the row's group
    const bool live = i < t.ti, has_blank = live && sub == 0;
    uint32_t desc[NE - 1];
    {
        const int64_t r = (int64_t(b) * N + t.t0 + (live ? i : 0)) * M;
        f_load_desc<NE - 1>(q_loc + r, q_pe + r, M, sub, live, L.UC, desc);
    }
    T *prow = reinterpret_cast<T *>(sp + i * L.RS);
    const float *srow = reinterpret_cast<const float *>(prow);
    if (sub == 0) reinterpret_cast<float *>(prow)[L.UC] = -INFINITY;  // what unused slots read
    f_zero_rows<T, C>(ks, t.U + 1, t.Upad);
    f_zero_rows<T, C>(vs, t.U + 1, t.Upad);
    f_zero_rows<T, C>(xs, t.ti, TT);
    __syncthreads();

    GatherStager<T, C> st;
    st.init(rows_s, t.U);
    // heads [h_lo, h_hi) of this CTA: small problems split the heads over gridDim.z to fill the machine
    const int HS = (H + int(gridDim.z) - 1) / int(gridDim.z);
    const int h_lo = int(blockIdx.z) * HS, h_hi = min(H, h_lo + HS);
    const T *qb = q.row(b, t.t0, h_lo), *kb = k.row(b, 0, h_lo), *vb = v.row(b, 0, h_lo);
    f_stage_tile<T, C, TT>(xs_a, qb, int(q.sN), t.ti);
    st.issue(ks_a, kb);
    f_cp_async_commit();
    st.issue(vs_a, vb);
    f_cp_async_commit();
    f_put_blank<T, C>(ks, t.U, blank_k, h_lo);
    f_put_blank<T, C>(vs, t.U, blank_v, h_lo);

    for (int h = h_lo; h < h_hi; ++h) {
        f_cp_async_wait<1>();  // Q_h, K_h
        __syncthreads();
        f_mma_xyT<T, C>(xs, ks, sp, L.RS, TT, t.Upad, scale);
        __syncthreads();
        if (h + 1 < h_hi) {
            f_stage_tile<T, C, TT>(xs_a, qb + (h + 1 - h_lo) * q.sH, int(q.sN), t.ti);
            st.issue(ks_a, kb + (h + 1 - h_lo) * k.sH);
            f_put_blank<T, C>(ks, t.U, blank_k, h + 1);
        }
        f_cp_async_commit();
        {
            float pv[NE];
            f_row_logits<NE>(srow, desc, table + h, H, t.U, has_blank, pv);
            const float l = f_row_softmax<TPR, NE>(live, pv);
            if (has_blank) lse[(int64_t(b) * H + h) * N + t.t0 + i] = l;
            __syncwarp();
            f_scatter_row<T, TPR, NE>(prow, desc, t.Upad, t.U, t.dup, sub, has_blank, pv);
        }
        f_cp_async_wait<1>();  // V_h
        __syncthreads();
        f_mma_pv_store<T, C>(sp, L.RS, vs, os, TT, t.Upad, t.ti, 1.f, out.row(b, t.t0, h), out.sN);
        __syncthreads();
        if (h + 1 < h_hi) {
            st.issue(vs_a, vb + (h + 1 - h_lo) * v.sH);
            f_put_blank<T, C>(vs, t.U, blank_v, h + 1);
        }
        f_cp_async_commit();
    }
    f_cp_async_wait<0>();
}

// ---------------------------------------------------------------------------------------------------------
// Backward, query-tile half.  Emits dQ, and dS / P in [B][H][N][M] for the key-tile scatter kernels and the table
// gradient; per-tile partial sums of the blank-token gradients go to blank_part [B][tiles][H][2][C] (fp32).
// Per head: S = Q·K_U^T and dP = dO·V_U^T on the tensor cores, then per row P = exp(S + bias − lse),
// dS = P∘(dP − δ) with δ = dO·O, then dQ = scale·dS·K_U.  V and K of head h+1 (second K buffer) are requested right
// after the two contractions of head h, Q/dO/O of head h+1 right after its row phase.
// ---------------------------------------------------------------------------------------------------------
template <typename T, int C, int TT, int NE>
__global__ void __launch_bounds__(kFT, 2)
fused_attn_bwd_kernel(View<T> q, View<T> k, View<T> v, View<T> out, View<T> dout, ViewW<T> dq,
                      const float *__restrict__ table, const float *__restrict__ blank_k,
                      const float *__restrict__ blank_v, const int32_t *__restrict__ q_hdr,
                      const int32_t *__restrict__ q_rows, const uint8_t *__restrict__ q_loc,
                      const uint16_t *__restrict__ q_pe, const float *__restrict__ lse, T *__restrict__ ds_out,
                      T *__restrict__ p_out, float *__restrict__ blank_part, int H, int N, int M, float scale,
                      FusedSmem L) {
    constexpr int CS = C + 8;
    constexpr int TPR = kFT / TT;
    constexpr int PER = NE - 1;
    constexpr int CPT = (C + TPR - 1) / TPR;  // channels per thread in the per-row channel reductions
    extern __shared__ __align__(128) unsigned char smem[];
    int *rows_s = reinterpret_cast<int *>(smem + L.rows);
    T *ks0 = reinterpret_cast<T *>(smem + L.ks);
    T *ks1 = reinterpret_cast<T *>(smem + L.ks2);
    T *vs = reinterpret_cast<T *>(smem + L.vs);
    T *xs = reinterpret_cast<T *>(smem + L.xs);
    T *dos = reinterpret_cast<T *>(smem + L.dos);
    T *outs = reinterpret_cast<T *>(smem + L.outs);
    unsigned char *sp = smem + L.sp;
    unsigned char *dp = smem + L.dp;
    float *os = reinterpret_cast<float *>(smem + L.os);  // over dp: dP is dead when the dQ partials are written
    float *bl = reinterpret_cast<float *>(smem + L.bl);  // [8 warps][2][C]
    const uint32_t vs_a = f_smem_u32(vs), xs_a = f_smem_u32(xs), dos_a = f_smem_u32(dos), outs_a = f_smem_u32(outs);

    const int b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const TileHdr t = f_load_tile(q_hdr, q_rows, N, TT, L.UC, int(k.sN), rows_s);
    if (t.bad) return;
    const int i = tid / TPR, sub = tid % TPR;
    const bool live = i < t.ti, has_blank = live && sub == 0;
    const int64_t tok = t.t0 + (live ? i : 0);
    uint32_t desc[PER];
    {
        const int64_t r = (int64_t(b) * N + tok) * M;
        f_load_desc<PER>(q_loc + r, q_pe + r, M, sub, live, L.UC, desc);
    }
    T *prow = reinterpret_cast<T *>(sp + i * L.RS);
    const float *srow = reinterpret_cast<const float *>(prow);
    const float *dprow = reinterpret_cast<const float *>(dp + i * L.RS);
    if (sub == 0) reinterpret_cast<float *>(prow)[L.UC] = -INFINITY;  // unused slots: P = 0 ...
    // ... and dP = 0 (set again per head below, the dQ partials overwrite it)
    f_zero_rows<T, C>(ks0, t.U + 1, t.Upad);
    f_zero_rows<T, C>(ks1, t.U + 1, t.Upad);
    f_zero_rows<T, C>(vs, t.U + 1, t.Upad);
    f_zero_rows<T, C>(xs, t.ti, TT);
    f_zero_rows<T, C>(dos, t.ti, TT);
    f_zero_rows<T, C>(outs, t.ti, TT);
    __syncthreads();

    GatherStager<T, C> st;
    st.init(rows_s, t.U);
    const int HS = (H + int(gridDim.z) - 1) / int(gridDim.z);
    const int h_lo = int(blockIdx.z) * HS, h_hi = min(H, h_lo + HS);
    const T *qb = q.row(b, t.t0, h_lo), *dob = dout.row(b, t.t0, h_lo), *ob = out.row(b, t.t0, h_lo);
    const T *kb = k.row(b, 0, h_lo), *vb = v.row(b, 0, h_lo);
    f_stage_tile<T, C, TT>(xs_a, qb, int(q.sN), t.ti);
    f_stage_tile<T, C, TT>(dos_a, dob, int(dout.sN), t.ti);
    f_stage_tile<T, C, TT>(outs_a, ob, int(out.sN), t.ti);
    st.issue(f_smem_u32(ks0), kb);
    st.issue(vs_a, vb);
    f_cp_async_commit();
    f_put_blank<T, C>(ks0, t.U, blank_k, h_lo);
    f_put_blank<T, C>(vs, t.U, blank_v, h_lo);
    const bool pair_ok = (PER % 2 == 0) && (M % 2 == 0);

    for (int h = h_lo; h < h_hi; ++h) {
        T *kcur = ((h - h_lo) & 1) ? ks1 : ks0, *knext = ((h - h_lo) & 1) ? ks0 : ks1;
        f_cp_async_wait<0>();
        __syncthreads();
        f_mma_xyT<T, C>(xs, kcur, sp, L.RS, TT, t.Upad, scale);  // S
        f_mma_xyT<T, C>(dos, vs, dp, L.RS, TT, t.Upad, 1.f);     // dP = dO · V_U^T
        if (sub == 0) const_cast<float *>(dprow)[L.UC] = 0.f;
        __syncthreads();
        if (h + 1 < h_hi) {
            st.issue(f_smem_u32(knext), kb + (h + 1 - h_lo) * k.sH);
            st.issue(vs_a, vb + (h + 1 - h_lo) * v.sH);
            f_put_blank<T, C>(knext, t.U, blank_k, h + 1);
            f_put_blank<T, C>(vs, t.U, blank_v, h + 1);
        }
        f_cp_async_commit();

        float acc_bk[CPT], acc_bv[CPT];
        {
            // delta = dO_i · O_i over this thread's channels, then over the row's group
            float dl = 0.f, qv[CPT], dov[CPT], ov[CPT];
#pragma unroll
            for (int cc = 0; cc < CPT; ++cc) qv[cc] = dov[cc] = ov[cc] = 0.f;
            if (sub * CPT < C) {  // rows past the end of the tile are zero in all three buffers
                f_load_n<T, CPT>(dos + i * CS + sub * CPT, dov);
                f_load_n<T, CPT>(xs + i * CS + sub * CPT, qv);
                f_load_n<T, CPT>(outs + i * CS + sub * CPT, ov);
            }
#pragma unroll
            for (int cc = 0; cc < CPT; ++cc) dl += dov[cc] * ov[cc];
            dl = f_group_sum<TPR>(dl);
            // P from the saved log-sum-exp: no max / sum reductions
            const float nl = live ? -__ldg(lse + (int64_t(b) * H + h) * N + tok) * kLog2e : 0.f;
            float pv[NE];
            f_row_logits<NE>(srow, desc, table + h, H, t.U, has_blank, pv);
            T *pg = p_out + ((int64_t(b) * H + h) * N + tok) * M + sub * PER;
            T *dg = ds_out + ((int64_t(b) * H + h) * N + tok) * M + sub * PER;
            if (pair_ok) {
#pragma unroll
                for (int e = 0; e + 1 < PER; e += 2) {
                    const float p0 = f_ex2(fmaf(pv[e], kLog2e, nl)), p1 = f_ex2(fmaf(pv[e + 1], kLog2e, nl));
                    const float d0 = p0 * (dprow[desc[e] & 511u] - dl), d1 = p1 * (dprow[desc[e + 1] & 511u] - dl);
                    if (int(desc[e]) >= 0) {
                        *reinterpret_cast<uint32_t *>(pg + e) = f_pack2<T>(p0, p1);
                        *reinterpret_cast<uint32_t *>(dg + e) = f_pack2<T>(d0, d1);
                    }
                    pv[e] = d0; pv[e + 1] = d1;
                }
            } else {
#pragma unroll
                for (int e = 0; e < PER; ++e) {
                    const float p0 = f_ex2(fmaf(pv[e], kLog2e, nl));
                    const float d0 = p0 * (dprow[desc[e] & 511u] - dl);
                    if (int(desc[e]) >= 0) {
                        pg[e] = Elem<T>::from_acc(p0);
                        dg[e] = Elem<T>::from_acc(d0);
                    }
                    pv[e] = d0;
                }
            }
            float pb = f_ex2(fmaf(pv[NE - 1], kLog2e, nl));  // 0 unless this thread holds the blank slot
            float dsb = has_blank ? pb * (dprow[t.U] - dl) : 0.f;
            pv[NE - 1] = dsb;
            // blank-token gradients: broadcast the blank slot's (p, ds) within the row's group
            const int src = lane & ~(TPR - 1);
            pb = __shfl_sync(0xffffffffu, pb, src);
            dsb = __shfl_sync(0xffffffffu, dsb, src);
#pragma unroll
            for (int cc = 0; cc < CPT; ++cc) {
                acc_bk[cc] = scale * dsb * qv[cc];
                acc_bv[cc] = pb * dov[cc];
            }
            __syncwarp();
            f_scatter_row<T, TPR, NE>(prow, desc, t.Upad, t.U, t.dup, sub, has_blank, pv);
        }
        // sum the blank partials over the rows of this warp (lanes with equal `sub`), then over the warps
#pragma unroll
        for (int cc = 0; cc < CPT; ++cc) {
#pragma unroll
            for (int o = TPR; o < 32; o <<= 1) {
                acc_bk[cc] += __shfl_xor_sync(0xffffffffu, acc_bk[cc], o);
                acc_bv[cc] += __shfl_xor_sync(0xffffffffu, acc_bv[cc], o);
            }
            const int c = sub * CPT + cc;
            if (lane < TPR && c < C) {
                bl[(warp * 2 + 0) * C + c] = acc_bk[cc];
                bl[(warp * 2 + 1) * C + c] = acc_bv[cc];
            }
        }
        __syncthreads();
        if (h + 1 < h_hi) {
            f_stage_tile<T, C, TT>(xs_a, qb + (h + 1 - h_lo) * q.sH, int(q.sN), t.ti);
            f_stage_tile<T, C, TT>(dos_a, dob + (h + 1 - h_lo) * dout.sH, int(dout.sN), t.ti);
            f_stage_tile<T, C, TT>(outs_a, ob + (h + 1 - h_lo) * out.sH, int(out.sN), t.ti);
        }
        f_cp_async_commit();
        if (tid < 2 * C) {
            float sacc = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) sacc += bl[w * 2 * C + tid];
            blank_part[((int64_t(b) * gridDim.x + blockIdx.x) * H + h) * 2 * C + tid] = sacc;
        }
        f_mma_pv_store<T, C>(sp, L.RS, kcur, os, TT, t.Upad, t.ti, scale, dq.row(b, t.t0, h), dq.sN);  // dQ = scale·dS·K_U
        __syncthreads();
    }
    f_cp_async_wait<0>();
}

}  // namespace affb200

// =========================================================================================================
// C ABI
// =========================================================================================================
using namespace affb200;

template <typename T> static View<T> mk_view(const affb200_view *v) { return View<T>{(const T *)v->p, v->sB, v->sN, v->sH}; }
template <typename T> static ViewW<T> mk_vieww(const affb200_view *v) { return ViewW<T>{(T *)v->p, v->sB, v->sN, v->sH}; }


// heads per CTA: problems with few tiles split the heads over gridDim.z (a power-of-two divisor of H) until the grid
// holds about four waves of the CTAs that fit the machine
static unsigned head_split(int64_t n_cta, int64_t H, int ctas_per_sm) {
    unsigned z = 1;
    while (int64_t(z) * 2 <= H && H % (int64_t(z) * 2) == 0 && n_cta * z < int64_t(4) * kNumSMs * ctas_per_sm) z *= 2;
    return z;
}

static bool kv_ok(const affb200_view *k, const affb200_view *v, int64_t N) {
    // the kernels index gathered rows with 32-bit element offsets row * sN shared by k and v
    return k->sN == v->sN && k->sN > 0 && N * k->sN < (int64_t(1) << 31);
}
static bool view_ok(const affb200_view *v, int64_t c) {
    // rows must be 16-byte aligned for cp.async / 128-bit stores
    return v && v->p && (reinterpret_cast<uintptr_t>(v->p) & 15) == 0 && (v->sB * 2) % 16 == 0 && (v->sN * 2) % 16 == 0 &&
           (v->sH * 2) % 16 == 0 && (c == 16 || c == 32);
}

static int fused_common_check(const char *what, const affb200_nbhd_plan *plan, const uint16_t *q_pe, int64_t B, int64_t H,
                              int64_t N, int64_t M, int64_t c, int umax, int dtype) {
    if (B <= 0 || H <= 0 || N <= 0 || M <= 0 || B > 65535 || M > AFFB200_PLAN_MAX_M || H > 65535) {
        set_error("%s: unsupported dimensions", what);
        return AFFB200_E_ARG;
    }
    if (dtype != AFFB200_F16 && dtype != AFFB200_BF16) { set_error("%s: f16 / bf16 only", what); return AFFB200_E_ARG; }
    if (c != 16 && c != 32) { set_error("%s: head dim must be 16 or 32", what); return AFFB200_E_ARG; }
    if (!plan || plan->tq <= 0 || !plan->q_hdr || !plan->q_rows || !plan->q_loc || !q_pe) {
        set_error("%s: a query-tile plan and packed bias indices are required", what);
        return AFFB200_E_ARG;
    }
    if (N > AFFB200_PLAN_MAX_NK) { set_error("%s: N > %d", what, AFFB200_PLAN_MAX_NK); return AFFB200_E_ARG; }
    if (umax < 0 || umax > AFFB200_PLAN_UMAX - 1) {
        set_error("%s: a tile references %d distinct rows; the fused kernel holds at most %d plus the blank token", what,
                  umax, AFFB200_PLAN_UMAX - 1);
        return AFFB200_E_ARG;
    }
    return AFFB200_OK;
}

extern "C" int affb200_fused_attn_pack_pe(const int64_t *pe_idx, const int64_t *mask, int64_t total, uint16_t *out,
                                          int32_t *bad, void *stream) {
    if (total < 0) { set_error("pack_pe: negative size"); return AFFB200_E_ARG; }
    if (total == 0) return AFFB200_OK;
    if (!pe_idx || !out || !bad) { set_error("pack_pe: null pointer"); return AFFB200_E_ARG; }
    pe_pack_kernel<<<unsigned((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(pe_idx, mask, total, out, bad);
    return check_launch("pack_pe");
}

extern "C" int affb200_fused_attn_fwd(const affb200_view *q, const affb200_view *k, const affb200_view *v,
                                      const affb200_view *out, const float *table, const float *blank_k,
                                      const float *blank_v, const affb200_nbhd_plan *plan, const uint16_t *q_pe,
                                      int32_t umax, float *lse, int64_t B, int64_t H, int64_t N, int64_t M, int64_t c,
                                      float scale, int dtype, void *stream) {
    if (int e = fused_common_check("fused_attn_fwd", plan, q_pe, B, H, N, M, c, umax, dtype)) return e;
    if (!view_ok(q, c) || !view_ok(k, c) || !view_ok(v, c) || !view_ok(out, c) || !table || !blank_k || !blank_v || !lse) {
        set_error("fused_attn_fwd: null or misaligned operand (rows must be 16-byte aligned)");
        return AFFB200_E_ARG;
    }
    if (!kv_ok(k, v, N) || N * q->sN >= (int64_t(1) << 31)) { set_error("fused_attn_fwd: k/v token strides must match and fit int32"); return AFFB200_E_ARG; }
    const int TT = plan->tq;
    const int UC = (umax + 1 + 31) & ~31;
    const FusedSmem L = fused_smem(UC, TT, int(c), false);
    if (L.total > 227 * 1024) { set_error("fused_attn_fwd: tile needs %zu bytes of shared memory", L.total); return AFFB200_E_ARG; }
    dim3 grid(unsigned((N + TT - 1) / TT), unsigned(B));
    grid.z = head_split(int64_t(grid.x) * grid.y, H, 3);
    auto st = (cudaStream_t)stream;
    return dispatch_dtype(dtype, [&](auto tag) -> int {
        using T0 = decltype(tag);
        if constexpr (std::is_same<T0, __half>::value || std::is_same<T0, __nv_bfloat16>::value) {
            using T = T0;
#define AFF_F(C_, TT_, NE_)                                                                                                 \
    do {                                                                                                           \
        auto kern = fused_attn_fwd_kernel<T, C_, TT_, NE_>;                                                        \
        if (L.total > 48 * 1024 &&                                                                                 \
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(L.total)) != cudaSuccess)  \
            return check_launch("cudaFuncSetAttribute", 0);                                                        \
        kern<<<grid, kFT, L.total, st>>>(mk_view<T>(q), mk_view<T>(k), mk_view<T>(v), mk_vieww<T>(out), table,     \
                                         blank_k, blank_v, plan->q_hdr, plan->q_rows, plan->q_loc, q_pe, lse,      \
                                         int(H), int(N), int(M), scale, L);                                        \
    } while (0)
            // (tile size, slots per thread): 256/TT threads per row, each NE-1 neighbour slots: (NE-1)*256/TT >= M
            const int need = int(M);
#define AFF_FC(C_)                                                                                   \
    do {                                                                                             \
        if (TT == 16 && need <= 12 * 16) AFF_F(C_, 16, 13);                                          \
        else if (TT == 32 && need <= 6 * 8) AFF_F(C_, 32, 7);                                        \
        else if (TT == 32 && need <= 12 * 8) AFF_F(C_, 32, 13);                                      \
        else if (TT == 64 && need <= 16 * 4) AFF_F(C_, 64, 17);                                      \
        else { set_error("fused_attn: tile size %d does not support M = %d", TT, int(M)); return AFFB200_E_ARG; } \
    } while (0)
            if (c == 32) AFF_FC(32); else AFF_FC(16);
#undef AFF_FC
#undef AFF_F
            return check_launch("fused_attn_fwd");
        } else {
            return AFFB200_E_ARG;
        }
    });
}

extern "C" int affb200_fused_attn_bwd(const affb200_view *q, const affb200_view *k, const affb200_view *v,
                                      const affb200_view *out, const affb200_view *dout, const affb200_view *dq,
                                      const affb200_view *dk, const affb200_view *dv, const float *table,
                                      const float *blank_k, const float *blank_v, const affb200_nbhd_plan *plan,
                                      const uint16_t *q_pe, int32_t umax, const float *lse, void *ds_work, void *p_work,
                                      float *blank_part, float *d_table, int64_t P, int64_t B, int64_t H, int64_t N,
                                      int64_t M, int64_t c, float scale, int dtype, void *stream) {
    if (int e = fused_common_check("fused_attn_bwd", plan, q_pe, B, H, N, M, c, umax, dtype)) return e;
    if (!view_ok(q, c) || !view_ok(k, c) || !view_ok(v, c) || !view_ok(out, c) || !view_ok(dout, c) || !view_ok(dq, c) ||
        !view_ok(dk, c) || !view_ok(dv, c) || !table || !blank_k || !blank_v || !lse || !ds_work || !p_work ||
        !blank_part || !d_table) {
        set_error("fused_attn_bwd: null or misaligned operand");
        return AFFB200_E_ARG;
    }
    if (!kv_ok(k, v, N) || N * q->sN >= (int64_t(1) << 31) || N * dout->sN >= (int64_t(1) << 31)) { set_error("fused_attn_bwd: k/v token strides must match and fit int32"); return AFFB200_E_ARG; }
    if (plan->tk <= 0 || !plan->k_hdr || !plan->k_utok || !plan->k_rloc) {
        set_error("fused_attn_bwd: a key-tile plan is required");
        return AFFB200_E_ARG;
    }
    if (P <= 0 || P > 32768) { set_error("fused_attn_bwd: bias table must have 1..32768 rows"); return AFFB200_E_ARG; }
    const int TT = plan->tq;
    const int UC = (umax + 1 + 31) & ~31;
    const FusedSmem L = fused_smem(UC, TT, int(c), true);
    if (L.total > 227 * 1024) { set_error("fused_attn_bwd: tile needs %zu bytes of shared memory", L.total); return AFFB200_E_ARG; }
    dim3 grid(unsigned((N + TT - 1) / TT), unsigned(B));
    grid.z = head_split(int64_t(grid.x) * grid.y, H, 2);
    auto st = (cudaStream_t)stream;
    return dispatch_dtype(dtype, [&](auto tag) -> int {
        using T0 = decltype(tag);
        if constexpr (std::is_same<T0, __half>::value || std::is_same<T0, __nv_bfloat16>::value) {
            using T = T0;
#define AFF_B(C_, TT_, NE_)                                                                                                 \
    do {                                                                                                           \
        auto kern = fused_attn_bwd_kernel<T, C_, TT_, NE_>;                                                        \
        if (L.total > 48 * 1024 &&                                                                                 \
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(L.total)) != cudaSuccess)  \
            return check_launch("cudaFuncSetAttribute", 0);                                                        \
        kern<<<grid, kFT, L.total, st>>>(mk_view<T>(q), mk_view<T>(k), mk_view<T>(v), mk_view<T>(out),             \
                                         mk_view<T>(dout), mk_vieww<T>(dq), table, blank_k, blank_v, plan->q_hdr,  \
                                         plan->q_rows, plan->q_loc, q_pe, lse, (T *)ds_work, (T *)p_work,          \
                                         blank_part, int(H), int(N), int(M), scale, L);                            \
    } while (0)
            // (tile size, slots per thread): 256/TT threads per row, each NE-1 neighbour slots: (NE-1)*256/TT >= M
            const int need = int(M);
#define AFF_BC(C_)                                                                                   \
    do {                                                                                             \
        if (TT == 16 && need <= 12 * 16) AFF_B(C_, 16, 13);                                          \
        else if (TT == 32 && need <= 6 * 8) AFF_B(C_, 32, 7);                                        \
        else if (TT == 32 && need <= 12 * 8) AFF_B(C_, 32, 13);                                      \
        else if (TT == 64 && need <= 16 * 4) AFF_B(C_, 64, 17);                                      \
        else { set_error("fused_attn: tile size %d does not support M = %d", TT, int(M)); return AFFB200_E_ARG; } \
    } while (0)
            if (c == 32) AFF_BC(32); else AFF_BC(16);
#undef AFF_BC
#undef AFF_B
            if (int e = check_launch("fused_attn_bwd")) return e;
            // dK = scale * scatter(dS, Q) ; dV = scatter(P, dO): key-tile ownership, no atomics
            const int64_t qs[3] = {q->sB, q->sH, q->sN}, dos_[3] = {dout->sB, dout->sH, dout->sN};
            const int64_t dks[3] = {dk->sB, dk->sH, dk->sN}, dvs[3] = {dv->sB, dv->sH, dv->sN};
            if (int e = launch_scatter_wsum_dense_strided<T>((const T *)ds_work, (const T *)q->p, plan, (T *)dk->p, B, H, N,
                                                             N, M, c, qs, dks, scale, st))
                return e;
            if (int e = launch_scatter_wsum_dense_strided<T>((const T *)p_work, (const T *)dout->p, plan, (T *)dv->p, B, H,
                                                             N, N, M, c, dos_, dvs, 1.f, st))
                return e;
            return launch_table_grad_u16(ds_work, q_pe, B, H, N * M, P, d_table, dtype, st);
        } else {
            return AFFB200_E_ARG;
        }
    });
}
