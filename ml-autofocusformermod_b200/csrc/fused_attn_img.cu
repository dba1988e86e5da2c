// fused_attn_img.cu — fused neighbourhood attention for SMALL token sets (N + 1 <= 256 rows: the later stages of
// the backbone, e.g. 125 tokens per image at stage 3 of AFF-Tiny-1/5, where 18 of the 25 blocks live).
//
// When an image has that few tokens, the union of rows a query tile references is nearly the whole image, so the
// per-tile kernels of fused_attn.cu stage the same K/V rows once per tile, and the backward pays two key-tile
// scatter launches plus a round trip of P and dS through HBM to transpose the contraction.  Here one CTA owns one
// (image, head): K and V of the image are loaded ONCE (contiguous rows, no gather list), the blank token is row N,
// and the CTA walks the image's query tiles with the next tile's Q/dO/O prefetched.  Column j of S is key row j
// itself, so the slot -> column map is the neighbour index narrowed to a byte.  In the backward the CTA sees every
// query of the image, therefore
//     dK = scale · dS^T · Q        dV = P^T · dO
// accumulate in tensor-core registers across the tiles (transposed ldmatrix of the dense dS / P tiles) and are written
// once — no scatter kernels, no P in HBM, and row N of the two accumulators IS the blank-token gradient.
// dS still goes to [B][H][N][M] for the bias-table gradient kernel.
#include "fused_attn.cuh"

namespace affb200 {

struct ImgSmem {
    int UC, RS;
    size_t ks, vs, xs, dos, outs, sp, dp, os, total;
};
__host__ __device__ inline ImgSmem img_smem(int UC, int TT, int C, bool bwd) {
    ImgSmem s;
    s.UC = UC;
    s.RS = UC * 4 + 16;
    const size_t kv = size_t(UC) * (C + 8) * 2, tile = size_t(TT) * (C + 8) * 2, srow = size_t(TT) * s.RS;
    const size_t osz = size_t(8 * 16 * 16 * (C + 8) / C) * 4;
    size_t o = 0;
    s.ks = o;   o += kv;
    s.vs = o;   o += kv;
    s.xs = o;   o += 2 * tile;                 // double-buffered over the tiles
    s.dos = o;  o += bwd ? 2 * tile : 0;
    s.outs = o; o += bwd ? 2 * tile : 0;
    s.sp = o;   o += srow;
    s.dp = o;   o += bwd ? srow : 0;
    s.os = o;   o += osz;
    s.total = o;
    return s;
}

// rows [0, n) of a [rows][C] slab (row stride `stride` elements) -> padded staging rows
template <typename T, int C>
__device__ __forceinline__ void img_stage_rows(uint32_t buf_addr, const T *base, int stride, int n) {
    constexpr int CS = C + 8, CH = C / 8;
    // lane -> chunk so that the 8 lanes of a 128-bit shared-memory phase write the SAME chunk of 8 consecutive rows:
    // with the 16-byte row padding those 8 stores fall into distinct bank groups (rows of one phase in the natural
    // order, 2 rows x 4 chunks, collide: row r chunk 0 and row r+1 chunk 3 share banks)
    const int n8 = n & ~7;
    for (int e = threadIdx.x; e < n * CH; e += kFT) {
        int row, ch;
        if (e < n8 * CH) {
            const int blk = e / (8 * CH), in = e - blk * (8 * CH);
            row = blk * 8 + (in & 7);
            ch = in >> 3;
        } else {
            row = e / CH;
            ch = e % CH;
        }
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(buf_addr + uint32_t(row * CS + ch * 8) * 2u),
                     "l"(base + row * stride + ch * 8) : "memory");
    }
}

// dupf[b][tile] != 0: some token of the 32-row tile lists a key row twice among its UNMASKED slots (the masked,
// padded slots all repeat row 0 and are skipped by the scatter).  One warp per token row, a 256-bit map per warp.
__global__ void __launch_bounds__(256)
img_dup_flags_kernel(const uint8_t *__restrict__ col, const uint16_t *__restrict__ q_pe, int N, int M,
                     uint8_t *__restrict__ dupf) {
    __shared__ uint32_t bm[8][8];
    __shared__ int any;
    const int b = blockIdx.y, t0 = blockIdx.x * 32, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) any = 0;
    __syncthreads();
    bool dup = false;
    for (int i = t0 + warp; i < min(N, t0 + 32); i += 8) {
        if (lane < 8) bm[warp][lane] = 0u;
        __syncwarp();
        const int64_t r = (int64_t(b) * N + i) * M;
        for (int j = lane; j < M; j += 32) {
            if (__ldg(q_pe + r + j) & 0x8000u) continue;
            const uint32_t c = __ldg(col + r + j), bit = 1u << (c & 31u);
            if (atomicOr(&bm[warp][c >> 5], bit) & bit) dup = true;
        }
        __syncwarp();
    }
    if (dup) any = 1;
    __syncthreads();
    if (threadIdx.x == 0) dupf[int64_t(b) * gridDim.x + blockIdx.x] = uint8_t(any);
}

__global__ void __launch_bounds__(256)
img_cols_kernel(const int64_t *__restrict__ idx, int64_t total, uint8_t *__restrict__ col) {
    const int64_t t = int64_t(blockIdx.x) * 256 + threadIdx.x;
    if (t < total) col[t] = uint8_t(__ldg(idx + t));
}

// acc[w] += A^T · Bm for the 16-row slices mt = warp + 8w of the M dimension:
//   A  [TT x Mpad] dense bf16/f16 tile at `ap` (row stride RS bytes)  -> read transposed
//   Bm [TT x C] staging rows
template <typename T, int C, int TT, int MW>
__device__ __forceinline__ void img_mma_atb(const unsigned char *ap, int RS, const T *bs, int Mpad,
                                            float (&acc)[MW][C / 8][4]) {
    constexpr int CS = C + 8;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int j = lane >> 3, r = lane & 7;
#pragma unroll
    for (int w = 0; w < MW; ++w) {
        const int mt = warp + w * 8;
        if (mt * 16 < Mpad) {
#pragma unroll
            for (int ks = 0; ks < TT / 16; ++ks) {
                uint32_t a[4];
                f_ldsm_x4_t(a, ap + size_t(ks * 16 + r + (j >> 1) * 8) * RS + (mt * 16 + (j & 1) * 8) * 2);
#pragma unroll
                for (int np = 0; np < C / 16; ++np) {
                    uint32_t bf[4];
                    f_ldsm_x4_t(bf, bs + (ks * 16 + (lane & 15)) * CS + np * 16 + (lane >> 4) * 8);
                    f_mma<T>(acc[w][np * 2], a, bf[0], bf[1]);
                    f_mma<T>(acc[w][np * 2 + 1], a, bf[2], bf[3]);
                }
            }
        }
    }
}

template <typename T, int C, int TT, int NE>
__global__ void __launch_bounds__(kFT, NE <= 7 ? 3 : 2)
fused_attn_img_fwd_kernel(View<T> q, View<T> k, View<T> v, ViewW<T> out, const float *__restrict__ table,
                          const float *__restrict__ blank_k, const float *__restrict__ blank_v,
                          const uint8_t *__restrict__ dupf, const uint8_t *__restrict__ col,
                          const uint16_t *__restrict__ q_pe, float *__restrict__ lse, int H, int N, int M, float scale,
                          ImgSmem L) {
    constexpr int TPR = kFT / TT;
    constexpr int CS = C + 8;
    extern __shared__ __align__(128) unsigned char smem[];
    T *ks = reinterpret_cast<T *>(smem + L.ks);
    T *vs = reinterpret_cast<T *>(smem + L.vs);
    T *xs = reinterpret_cast<T *>(smem + L.xs);
    unsigned char *sp = smem + L.sp;
    float *os = reinterpret_cast<float *>(smem + L.os);

    const int h = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const int Upad = (N + 1 + 15) & ~15;
    const int n_tiles = (N + TT - 1) / TT;
    const int i = tid / TPR, sub = tid % TPR;
    img_stage_rows<T, C>(f_smem_u32(ks), k.row(b, 0, h), int(k.sN), N);
    img_stage_rows<T, C>(f_smem_u32(vs), v.row(b, 0, h), int(v.sN), N);
    f_stage_tile<T, C, TT>(f_smem_u32(xs), q.row(b, 0, h), int(q.sN), min(TT, N));
    f_cp_async_commit();
    f_zero_rows<T, C>(ks, N + 1, Upad);
    f_zero_rows<T, C>(vs, N + 1, Upad);
    f_put_blank<T, C>(ks, N, blank_k, h);
    f_put_blank<T, C>(vs, N, blank_v, h);
    // rows past the end of the last tile, in the buffer that tile will use: never-written shared memory must not
    // reach the tensor cores (0 x NaN).  With three or more tiles the buffer holds an earlier tile's (finite) rows.
    if (n_tiles <= 2 && n_tiles * TT > N) {
        T *xl = xs + ((n_tiles - 1) & 1) * TT * CS;
        f_zero_rows<T, C>(xl, N - (n_tiles - 1) * TT, TT);
    }
    T *prow = reinterpret_cast<T *>(sp + i * L.RS);
    const float *srow = reinterpret_cast<const float *>(prow);
    if (sub == 0) reinterpret_cast<float *>(prow)[L.UC] = -INFINITY;  // what unused slots read

    // slot words of tile t+1 are requested during tile t (M even and NE-1 even: checked by the launcher)
    DescRaw<NE - 1> raw;
    {
        const bool live0 = i < min(TT, N);
        const int64_t r = (int64_t(b) * N + (live0 ? i : 0)) * M;
        f_load_desc_raw<NE - 1>(col + r, q_pe + r, M, sub, live0, raw);
    }
    for (int t = 0; t < n_tiles; ++t) {
        const int t0 = t * TT, ti = min(TT, N - t0);
        const bool live = i < ti, has_blank = live && sub == 0;
        const bool dup = dupf[int64_t(b) * n_tiles + t] != 0;
        uint32_t desc[NE - 1];
        float bias[NE - 1];
        f_make_desc<NE - 1>(raw, L.UC, desc);
        f_load_bias<NE - 1>(desc, table + h, H, bias);
        if (t + 1 < n_tiles) {
            const bool live1 = i < min(TT, N - t0 - TT);
            const int64_t r = (int64_t(b) * N + t0 + TT + (live1 ? i : 0)) * M;
            f_load_desc_raw<NE - 1>(col + r, q_pe + r, M, sub, live1, raw);
        }
        const T *xc = xs + (t & 1) * TT * CS;
        f_cp_async_wait<0>();
        __syncthreads();
        if (t + 1 < n_tiles)
            f_stage_tile<T, C, TT>(f_smem_u32(xs + ((t + 1) & 1) * TT * CS), q.row(b, t0 + TT, h), int(q.sN),
                                   min(TT, N - t0 - TT));
        f_cp_async_commit();
        f_mma_xyT<T, C>(xc, ks, sp, L.RS, TT, Upad, scale);
        __syncthreads();
        {
            float pv[NE];
            f_row_logits_b<NE>(srow, desc, bias, N, has_blank, pv);
            const float l = f_row_softmax<TPR, NE>(live, pv);
            if (has_blank) lse[(int64_t(b) * H + h) * N + t0 + i] = l;
            __syncwarp();
            f_scatter_row<T, TPR, NE, true>(prow, desc, Upad, N, dup, sub, has_blank, pv);
        }
        __syncthreads();
        f_mma_pv_store<T, C>(sp, L.RS, vs, os, TT, Upad, ti, 1.f, out.row(b, t0, h), out.sN);
        __syncthreads();
    }
    f_cp_async_wait<0>();
}

// Backward.  Per tile: S = Q·K^T and dP = dO·V^T, the row phase (P from the saved log-sum-exp, dS = P∘(dP − δ)),
// dS and P scattered into the dense bf16 tiles that alias the S and dP rows, dQ = scale·dS·K, then the two transposed
// contractions into the dK / dV accumulators.  blank_part is [B][H][2][C] (fp32): row N of the accumulators.
template <typename T, int C, int TT, int NE, int MW>
__global__ void __launch_bounds__(kFT, 2)
fused_attn_img_bwd_kernel(View<T> q, View<T> k, View<T> v, View<T> out, View<T> dout, ViewW<T> dq, ViewW<T> dk,
                          ViewW<T> dv, const float *__restrict__ table, const float *__restrict__ blank_k,
                          const float *__restrict__ blank_v, const uint8_t *__restrict__ dupf,
                          const uint8_t *__restrict__ col, const uint16_t *__restrict__ q_pe,
                          const float *__restrict__ lse, T *__restrict__ ds_out, float *__restrict__ blank_part, int H,
                          int N, int M, float scale, ImgSmem L) {
    constexpr int CS = C + 8;
    constexpr int TPR = kFT / TT;
    constexpr int PER = NE - 1;
    constexpr int CPT = (C + TPR - 1) / TPR;
    extern __shared__ __align__(128) unsigned char smem[];
    T *ks = reinterpret_cast<T *>(smem + L.ks);
    T *vs = reinterpret_cast<T *>(smem + L.vs);
    T *xs = reinterpret_cast<T *>(smem + L.xs);
    T *dos = reinterpret_cast<T *>(smem + L.dos);
    T *outs = reinterpret_cast<T *>(smem + L.outs);
    unsigned char *sp = smem + L.sp;
    unsigned char *dp = smem + L.dp;
    float *os = reinterpret_cast<float *>(smem + L.os);

    const int h = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int Upad = (N + 1 + 15) & ~15;
    const int n_tiles = (N + TT - 1) / TT;
    const int i = tid / TPR, sub = tid % TPR;
    img_stage_rows<T, C>(f_smem_u32(ks), k.row(b, 0, h), int(k.sN), N);
    img_stage_rows<T, C>(f_smem_u32(vs), v.row(b, 0, h), int(v.sN), N);
    {
        const int t1 = min(TT, N);
        f_stage_tile<T, C, TT>(f_smem_u32(xs), q.row(b, 0, h), int(q.sN), t1);
        f_stage_tile<T, C, TT>(f_smem_u32(dos), dout.row(b, 0, h), int(dout.sN), t1);
        f_stage_tile<T, C, TT>(f_smem_u32(outs), out.row(b, 0, h), int(out.sN), t1);
    }
    f_cp_async_commit();
    f_zero_rows<T, C>(ks, N + 1, Upad);
    f_zero_rows<T, C>(vs, N + 1, Upad);
    f_put_blank<T, C>(ks, N, blank_k, h);
    f_put_blank<T, C>(vs, N, blank_v, h);
    if (n_tiles <= 2 && n_tiles * TT > N) {
        const int o = ((n_tiles - 1) & 1) * TT * CS, lo = N - (n_tiles - 1) * TT;
        f_zero_rows<T, C>(xs + o, lo, TT);
        f_zero_rows<T, C>(dos + o, lo, TT);
        f_zero_rows<T, C>(outs + o, lo, TT);
    }
    T *prow = reinterpret_cast<T *>(sp + i * L.RS);
    T *pprow = reinterpret_cast<T *>(dp + i * L.RS);
    const float *srow = reinterpret_cast<const float *>(prow);
    const float *dprow = reinterpret_cast<const float *>(pprow);
    if (sub == 0) {
        reinterpret_cast<float *>(prow)[L.UC] = -INFINITY;  // unused slots: P = 0
        reinterpret_cast<float *>(pprow)[L.UC] = 0.f;       // ... and dP = 0
    }
    float acc_k[MW][C / 8][4], acc_v[MW][C / 8][4];
#pragma unroll
    for (int w = 0; w < MW; ++w)
#pragma unroll
        for (int n = 0; n < C / 8; ++n)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc_k[w][n][e] = acc_v[w][n][e] = 0.f;
    const bool pair_ok = (PER % 2 == 0) && (M % 2 == 0);

    DescRaw<PER> raw;
    float lse_next;
    {
        const bool live0 = i < min(TT, N);
        const int64_t tok0 = live0 ? i : 0;
        const int64_t r = (int64_t(b) * N + tok0) * M;
        f_load_desc_raw<PER>(col + r, q_pe + r, M, sub, live0, raw);
        lse_next = live0 ? __ldg(lse + (int64_t(b) * H + h) * N + tok0) : 0.f;
    }
    for (int t = 0; t < n_tiles; ++t) {
        const int t0 = t * TT, ti = min(TT, N - t0);
        const bool live = i < ti, has_blank = live && sub == 0;
        const bool dup = dupf[int64_t(b) * n_tiles + t] != 0;
        const int64_t tok = t0 + (live ? i : 0);
        uint32_t desc[PER];
        float bias[PER];
        f_make_desc<PER>(raw, L.UC, desc);
        f_load_bias<PER>(desc, table + h, H, bias);
        const float lse_cur = lse_next;
        if (t + 1 < n_tiles) {
            const bool live1 = i < min(TT, N - t0 - TT);
            const int64_t tok1 = t0 + TT + (live1 ? i : 0);
            const int64_t r = (int64_t(b) * N + tok1) * M;
            f_load_desc_raw<PER>(col + r, q_pe + r, M, sub, live1, raw);
            lse_next = live1 ? __ldg(lse + (int64_t(b) * H + h) * N + tok1) : 0.f;
        }
        const int cur = (t & 1) * TT * CS, nxt = ((t + 1) & 1) * TT * CS;
        f_cp_async_wait<0>();
        __syncthreads();
        if (t + 1 < n_tiles) {
            const int t1 = min(TT, N - t0 - TT);
            f_stage_tile<T, C, TT>(f_smem_u32(xs + nxt), q.row(b, t0 + TT, h), int(q.sN), t1);
            f_stage_tile<T, C, TT>(f_smem_u32(dos + nxt), dout.row(b, t0 + TT, h), int(dout.sN), t1);
            f_stage_tile<T, C, TT>(f_smem_u32(outs + nxt), out.row(b, t0 + TT, h), int(out.sN), t1);
        }
        f_cp_async_commit();
        f_mma_xyT<T, C>(xs + cur, ks, sp, L.RS, TT, Upad, scale);  // S
        f_mma_xyT<T, C>(dos + cur, vs, dp, L.RS, TT, Upad, 1.f);   // dP = dO · V^T
        __syncthreads();
        {
            float dl = 0.f, dov[CPT], ov[CPT];
#pragma unroll
            for (int cc = 0; cc < CPT; ++cc) dov[cc] = ov[cc] = 0.f;
            if (sub * CPT < C) {
                f_load_n<T, CPT>(dos + cur + i * CS + sub * CPT, dov);
                f_load_n<T, CPT>(outs + cur + i * CS + sub * CPT, ov);
            }
#pragma unroll
            for (int cc = 0; cc < CPT; ++cc) dl += dov[cc] * ov[cc];
            dl = f_group_sum<TPR>(dl);
            const float nl = -lse_cur * kLog2e;   // 0 for rows past the end of the tile
            float pv[NE], dsv[NE];
            f_row_logits_b<NE>(srow, desc, bias, N, has_blank, pv);
            T *dg = ds_out + ((int64_t(b) * H + h) * N + tok) * M + sub * PER;
#pragma unroll
            for (int e = 0; e < PER; ++e) {
                pv[e] = f_ex2(fmaf(pv[e], kLog2e, nl));
                dsv[e] = pv[e] * (dprow[desc[e] & 511u] - dl);
            }
            if (pair_ok) {
#pragma unroll
                for (int e = 0; e + 1 < PER; e += 2)
                    if (int(desc[e]) >= 0) *reinterpret_cast<uint32_t *>(dg + e) = f_pack2<T>(dsv[e], dsv[e + 1]);
            } else {
#pragma unroll
                for (int e = 0; e < PER; ++e)
                    if (int(desc[e]) >= 0) dg[e] = Elem<T>::from_acc(dsv[e]);
            }
            pv[NE - 1] = f_ex2(fmaf(pv[NE - 1], kLog2e, nl));  // 0 unless this thread holds the blank slot
            dsv[NE - 1] = has_blank ? pv[NE - 1] * (dprow[N] - dl) : 0.f;
            __syncwarp();
            f_scatter_row<T, TPR, NE, true>(prow, desc, Upad, N, dup, sub, has_blank, dsv);
            f_scatter_row<T, TPR, NE, true>(pprow, desc, Upad, N, dup, sub, has_blank, pv);
        }
        __syncthreads();
        f_mma_pv_store<T, C>(sp, L.RS, ks, os, TT, Upad, ti, scale, dq.row(b, t0, h), dq.sN);  // dQ = scale·dS·K
        img_mma_atb<T, C, TT, MW>(sp, L.RS, xs + cur, Upad, acc_k);                             // dK += dS^T·Q
        img_mma_atb<T, C, TT, MW>(dp, L.RS, dos + cur, Upad, acc_v);                            // dV += P^T·dO
        __syncthreads();
    }
    f_cp_async_wait<0>();
    // rows < N -> dK / dV, row N -> the blank-token gradients
    const int g = lane >> 2, q4 = lane & 3;
    float *bp = blank_part + (int64_t(b) * H + h) * 2 * C;
#pragma unroll
    for (int w = 0; w < MW; ++w) {
        const int mt = warp + w * 8;
        if (mt * 16 >= Upad) continue;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int u = mt * 16 + g + half * 8;
#pragma unroll
            for (int n = 0; n < C / 8; ++n) {
                const int c = n * 8 + q4 * 2;
                const float k0 = acc_k[w][n][half * 2] * scale, k1 = acc_k[w][n][half * 2 + 1] * scale;
                const float v0 = acc_v[w][n][half * 2], v1 = acc_v[w][n][half * 2 + 1];
                if (u < N) {
                    *reinterpret_cast<uint32_t *>(dk.row(b, u, h) + c) = f_pack2<T>(k0, k1);
                    *reinterpret_cast<uint32_t *>(dv.row(b, u, h) + c) = f_pack2<T>(v0, v1);
                } else if (u == N) {
                    bp[c] = k0; bp[c + 1] = k1;
                    bp[C + c] = v0; bp[C + c + 1] = v1;
                }
            }
        }
    }
}

}  // namespace affb200

// =========================================================================================================
// C ABI
// =========================================================================================================
using namespace affb200;

template <typename T> static View<T> img_view(const affb200_view *v) { return View<T>{(const T *)v->p, v->sB, v->sN, v->sH}; }
template <typename T> static ViewW<T> img_vieww(const affb200_view *v) { return ViewW<T>{(T *)v->p, v->sB, v->sN, v->sH}; }

constexpr int kImgTT = 32;

static bool img_view_ok(const affb200_view *v, int64_t N) {
    return v && v->p && (reinterpret_cast<uintptr_t>(v->p) & 15) == 0 && (v->sB * 2) % 16 == 0 && (v->sN * 2) % 16 == 0 &&
           (v->sH * 2) % 16 == 0 && v->sN > 0 && N * v->sN < (int64_t(1) << 31);
}

static int img_common_check(const char *what, const uint8_t *dupf, const void *col, const void *q_pe, int64_t B,
                            int64_t H, int64_t N, int64_t M, int64_t c, int dtype) {
    if (B <= 0 || H <= 0 || N <= 0 || M <= 0 || B > 65535 || H > 65535 || N > AFFB200_FUSED_IMG_MAX_N || M > 96 || M % 2) {
        set_error("%s: needs 1 <= N <= %d tokens per image and an even M <= 96 (got N=%lld M=%lld)", what,
                  AFFB200_FUSED_IMG_MAX_N, (long long)N, (long long)M);
        return AFFB200_E_ARG;
    }
    if (dtype != AFFB200_F16 && dtype != AFFB200_BF16) { set_error("%s: f16 / bf16 only", what); return AFFB200_E_ARG; }
    if (c != 16 && c != 32) { set_error("%s: head dim must be 16 or 32", what); return AFFB200_E_ARG; }
    if (!dupf || !col || !q_pe) {
        set_error("%s: duplicate flags, byte neighbour columns and packed bias indices are required", what);
        return AFFB200_E_ARG;
    }
    return AFFB200_OK;
}

extern "C" int affb200_fused_attn_img_prepare(const int64_t *nbhd_idx, const uint16_t *q_pe, int64_t B, int64_t N,
                                              int64_t M, uint8_t *col, uint8_t *dupf, void *stream) {
    if (B <= 0 || N <= 0 || M <= 0 || B > 65535 || N > AFFB200_FUSED_IMG_MAX_N || !nbhd_idx || !q_pe || !col || !dupf) {
        set_error("fused_attn_img_prepare: bad argument (needs N <= %d)", AFFB200_FUSED_IMG_MAX_N);
        return AFFB200_E_ARG;
    }
    auto st = (cudaStream_t)stream;
    const int64_t total = B * N * M;
    img_cols_kernel<<<unsigned((total + 255) / 256), 256, 0, st>>>(nbhd_idx, total, col);
    const dim3 grid{unsigned((N + 31) / 32), unsigned(B)};
    img_dup_flags_kernel<<<grid, 256, 0, st>>>(col, q_pe, int(N), int(M), dupf);
    return check_launch("fused_attn_img_prepare", 2);
}

extern "C" int affb200_fused_attn_img_fwd(const affb200_view *q, const affb200_view *k, const affb200_view *v,
                                          const affb200_view *out, const float *table, const float *blank_k,
                                          const float *blank_v, const uint8_t *dupf, const uint8_t *col,
                                          const uint16_t *q_pe, float *lse, int64_t B, int64_t H, int64_t N, int64_t M,
                                          int64_t c, float scale, int dtype, void *stream) {
    if (int e = img_common_check("fused_attn_img_fwd", dupf, col, q_pe, B, H, N, M, c, dtype)) return e;
    if (!img_view_ok(q, N) || !img_view_ok(k, N) || !img_view_ok(v, N) || !img_view_ok(out, N) || !table || !blank_k ||
        !blank_v || !lse) {
        set_error("fused_attn_img_fwd: null or misaligned operand (rows must be 16-byte aligned)");
        return AFFB200_E_ARG;
    }
    const int UC = int(N + 1 + 31) & ~31;
    const ImgSmem L = img_smem(UC, kImgTT, int(c), false);
    const dim3 grid{unsigned(H), unsigned(B)};
    auto st = (cudaStream_t)stream;
    return dispatch_dtype(dtype, [&](auto tag) -> int {
        using T0 = decltype(tag);
        if constexpr (std::is_same<T0, __half>::value || std::is_same<T0, __nv_bfloat16>::value) {
            using T = T0;
#define AFF_IF(C_, NE_)                                                                                            \
    do {                                                                                                           \
        auto kern = fused_attn_img_fwd_kernel<T, C_, kImgTT, NE_>;                                                 \
        if (L.total > 48 * 1024 &&                                                                                 \
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(L.total)) != cudaSuccess)  \
            return check_launch("cudaFuncSetAttribute", 0);                                                        \
        kern<<<grid, kFT, L.total, st>>>(img_view<T>(q), img_view<T>(k), img_view<T>(v), img_vieww<T>(out), table, \
                                         blank_k, blank_v, dupf, col, q_pe, lse, int(H), int(N), int(M), scale, L);  \
    } while (0)
            if (c == 32) { if (M <= 48) AFF_IF(32, 7); else AFF_IF(32, 13); }
            else { if (M <= 48) AFF_IF(16, 7); else AFF_IF(16, 13); }
#undef AFF_IF
            return check_launch("fused_attn_img_fwd");
        } else {
            return AFFB200_E_ARG;
        }
    });
}

extern "C" int affb200_fused_attn_img_bwd(const affb200_view *q, const affb200_view *k, const affb200_view *v,
                                          const affb200_view *out, const affb200_view *dout, const affb200_view *dq,
                                          const affb200_view *dk, const affb200_view *dv, const float *table,
                                          const float *blank_k, const float *blank_v, const uint8_t *dupf,
                                          const uint8_t *col, const uint16_t *q_pe, const float *lse, void *ds_work,
                                          float *blank_part, float *d_table, int64_t P, int64_t B, int64_t H, int64_t N,
                                          int64_t M, int64_t c, float scale, int dtype, void *stream) {
    if (int e = img_common_check("fused_attn_img_bwd", dupf, col, q_pe, B, H, N, M, c, dtype)) return e;
    if (!img_view_ok(q, N) || !img_view_ok(k, N) || !img_view_ok(v, N) || !img_view_ok(out, N) || !img_view_ok(dout, N) ||
        !img_view_ok(dq, N) || !img_view_ok(dk, N) || !img_view_ok(dv, N) || !table || !blank_k || !blank_v || !lse ||
        !ds_work || !blank_part) {
        set_error("fused_attn_img_bwd: null or misaligned operand");
        return AFFB200_E_ARG;
    }
    if (P <= 0 || P > 32768) { set_error("fused_attn_img_bwd: bias table must have 1..32768 rows"); return AFFB200_E_ARG; }
    const int UC = int(N + 1 + 31) & ~31;
    const ImgSmem L = img_smem(UC, kImgTT, int(c), true);
    const bool wide = ((N + 1 + 15) & ~15) > 128;  // more than 8 sixteen-row slices of dK/dV: two per warp
    const dim3 grid{unsigned(H), unsigned(B)};
    auto st = (cudaStream_t)stream;
    return dispatch_dtype(dtype, [&](auto tag) -> int {
        using T0 = decltype(tag);
        if constexpr (std::is_same<T0, __half>::value || std::is_same<T0, __nv_bfloat16>::value) {
            using T = T0;
#define AFF_IB(C_, NE_, MW_)                                                                                       \
    do {                                                                                                           \
        auto kern = fused_attn_img_bwd_kernel<T, C_, kImgTT, NE_, MW_>;                                            \
        if (L.total > 48 * 1024 &&                                                                                 \
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(L.total)) != cudaSuccess)  \
            return check_launch("cudaFuncSetAttribute", 0);                                                        \
        kern<<<grid, kFT, L.total, st>>>(img_view<T>(q), img_view<T>(k), img_view<T>(v), img_view<T>(out),         \
                                         img_view<T>(dout), img_vieww<T>(dq), img_vieww<T>(dk), img_vieww<T>(dv),  \
                                         table, blank_k, blank_v, dupf, col, q_pe, lse,                            \
                                         (T *)ds_work, blank_part, int(H), int(N), int(M), scale, L);              \
    } while (0)
#define AFF_IBC(C_)                                                                      \
    do {                                                                                 \
        if (M <= 48) { if (wide) AFF_IB(C_, 7, 2); else AFF_IB(C_, 7, 1); }              \
        else { if (wide) AFF_IB(C_, 13, 2); else AFF_IB(C_, 13, 1); }                    \
    } while (0)
            if (c == 32) AFF_IBC(32); else AFF_IBC(16);
#undef AFF_IBC
#undef AFF_IB
            if (int e = check_launch("fused_attn_img_bwd")) return e;
            if (!d_table) return AFFB200_OK;   // the caller takes the bias gradient from ds_work (affb200_pe_feat_grad)
            return launch_table_grad_u16(ds_work, q_pe, B, H, N * M, P, d_table, dtype, st);
        } else {
            return AFFB200_E_ARG;
        }
    });
}
