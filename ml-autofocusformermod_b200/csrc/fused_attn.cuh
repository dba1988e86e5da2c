// fused_attn.cuh — device helpers shared by the fused neighbourhood-attention kernels (fused_attn.cu: one CTA per
// query tile; fused_attn_img.cu: one CTA per (image, head) with the image's keys resident in shared memory).
#pragma once
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
    t.t0 = tile * TT;   // TT here = tokens per tile of the plan (<= the kernel's row capacity)
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

// The same in two steps, so that the global loads of tile t+1 can be issued while tile t is still being computed:
// raw words now (PER % 2 == 0, M % 2 == 0: aligned pairs), descriptors when the tile starts.
template <int PER> struct DescRaw {
    uint32_t loc[PER / 2];  // two 1-byte columns
    uint32_t pe[PER / 2];   // two packed bias indices
};
template <int PER>
__device__ __forceinline__ void f_load_desc_raw(const uint8_t *__restrict__ loc_r, const uint16_t *__restrict__ pe_r, int M,
                                                int sub, bool live, DescRaw<PER> &r) {
    const int j0 = sub * PER;
#pragma unroll
    for (int e = 0; e < PER; e += 2) {
        r.loc[e / 2] = 0xffffffffu;   // marks an unused pair
        r.pe[e / 2] = 0u;
        if (live && j0 + e < M) {
            r.loc[e / 2] = __ldg(reinterpret_cast<const uint16_t *>(loc_r + j0 + e));
            r.pe[e / 2] = __ldg(reinterpret_cast<const uint32_t *>(pe_r + j0 + e));
        }
    }
}
template <int PER>
__device__ __forceinline__ void f_make_desc(const DescRaw<PER> &r, int UC, uint32_t (&desc)[PER]) {
    auto mk = [](uint32_t loc, uint32_t pe) { return loc | ((pe & 0x7fffu) << 9) | ((pe >> 15) << 24); };
#pragma unroll
    for (int e = 0; e < PER; e += 2) {
        if (r.loc[e / 2] == 0xffffffffu) {
            desc[e] = desc[e + 1] = kSlotUnused | uint32_t(UC);
        } else {
            desc[e] = mk(r.loc[e / 2] & 0xffu, r.pe[e / 2] & 0xffffu);
            desc[e + 1] = mk(r.loc[e / 2] >> 8, r.pe[e / 2] >> 16);
        }
    }
}
// bias-table entries of the slots, requested ahead of the contraction that produces S
template <int PER>
__device__ __forceinline__ void f_load_bias(const uint32_t (&desc)[PER], const float *__restrict__ table_h, int H,
                                            float (&bias)[PER]) {
#pragma unroll
    for (int e = 0; e < PER; ++e) bias[e] = __ldg(table_h + ((desc[e] >> 9) & 0x7fffu) * H);
}
template <int NE>
__device__ __forceinline__ void f_row_logits_b(const float *srow, const uint32_t (&desc)[NE - 1],
                                               const float (&bias)[NE - 1], int U, bool has_blank, float (&pv)[NE]) {
#pragma unroll
    for (int e = 0; e < NE - 1; ++e) {
        const uint32_t d = desc[e];
        pv[e] = srow[d & 511u] + bias[e] + ((d & (1u << 24)) ? -100.f : 0.f);
    }
    pv[NE - 1] = has_blank ? srow[U] : -INFINITY;
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
// SKIPM: masked slots (bit 24) are left out — their probability is exp(-100)-small, zero in the operand's precision —
// so that the padded slots, which all repeat row 0, do not force the sequential path.
template <typename T, int TPR, int NE, bool SKIPM = false>
__device__ __forceinline__ void f_scatter_row(T *prow, const uint32_t (&desc)[NE - 1], int Upad, int U, bool dup,
                                              int sub, bool has_blank, const float (&pv)[NE]) {
    constexpr uint32_t kSkip = SKIPM ? (kSlotUnused | (1u << 24)) : kSlotUnused;
    if (!dup) {
        for (int w = sub; w < Upad / 8; w += TPR) reinterpret_cast<uint4 *>(prow)[w] = make_uint4(0u, 0u, 0u, 0u);
        __syncwarp();
#pragma unroll
        for (int e = 0; e < NE - 1; ++e)
            if ((desc[e] & kSkip) == 0u) prow[desc[e] & 511u] = Elem<T>::from_acc(pv[e]);
        if (has_blank) prow[U] = Elem<T>::from_acc(pv[NE - 1]);
    } else {
        float *frow = reinterpret_cast<float *>(prow);
        for (int w = sub; w < Upad; w += TPR) frow[w] = 0.f;
        __syncwarp();
#pragma unroll
        for (int e = 0; e < NE - 1; ++e)
            if ((desc[e] & kSkip) == 0u) atomicAdd(frow + (desc[e] & 511u), pv[e]);
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

}  // namespace affb200
