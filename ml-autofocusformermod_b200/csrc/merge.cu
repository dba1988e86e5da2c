// merge.cu — the merge half of ClusterMerging as three kernels (reference models/aff_transformer.py:343-361):
//
//   W[b,i,j,:]   = table[pe_idx[b,i,j], :] * lp[b, nbhd[b,i,j]] * mask[b,i,j]          (PointConv weights, 4 per slot)
//   out[b,i,q,c] = sum_j W[b,i,j,q] * feat[b, nbhd[b,i,j], c]                          (CLUSTENWF)
//
// The reference materialises W ([B,N',M,4], via a table gather, a gather of lp, two multiplies) and its gradient, and
// runs the WF operator on it.  Here the weights are formed in registers from their factors, forward and backward:
//   forward      one warp per kept token: the M weight quadruples go to shared memory, then the M neighbour rows are
//                streamed once, each lane holding C/32 channels of the 4 output rows;
//   backward/w   same mapping with d_out in registers: dW[j,:] = feat[nbhd_j,:]·d_out[q,:] by a 6-shuffle transposing
//                reduction, turned at once into the table-gradient operand dW*lp*mask ([B,N',M,4], consumed by the
//                shared-memory histogram kernel of attn_glue.cu) and the lp gradient (one fp32 atomic per slot);
//   backward/f   d_feat[b,r,:] = lp[b,r] * sum over the slots that reference r (transposed index) of
//                mask * sum_q table[pe][q] * d_out[b,i,q,:] — the common factor lp[b,r] leaves the sum, W is never read.
// fp32 arithmetic throughout (under autocast the reference runs this block in fp32 as well: weight_net ends in a
// LayerNorm); feat / lp are read in their storage type.
#include "common.cuh"
#include <algorithm>

namespace affb200 {

constexpr int kMT = 256;   // 8 warps = 8 tokens per CTA
constexpr int kMW = kMT / 32;

template <typename T, int CPL> __device__ __forceinline__ void mg_load(const T *p, float (&o)[CPL]) {
    // CPL contiguous elements per lane: 1, 2, 4, 8 or 16
    if constexpr (sizeof(T) == 4) {
        if constexpr (CPL >= 4) {
#pragma unroll
            for (int k = 0; k < CPL / 4; ++k) {
                const float4 v = __ldg(reinterpret_cast<const float4 *>(p) + k);
                o[4 * k] = v.x; o[4 * k + 1] = v.y; o[4 * k + 2] = v.z; o[4 * k + 3] = v.w;
            }
        } else if constexpr (CPL == 2) {
            const float2 v = __ldg(reinterpret_cast<const float2 *>(p));
            o[0] = v.x; o[1] = v.y;
        } else {
            o[0] = __ldg(reinterpret_cast<const float *>(p));
        }
    } else {
        if constexpr (CPL >= 8) {
#pragma unroll
            for (int k = 0; k < CPL / 8; ++k) {
                const uint4 r = __ldg(reinterpret_cast<const uint4 *>(p) + k);
                const T *e = reinterpret_cast<const T *>(&r);
#pragma unroll
                for (int q = 0; q < 8; ++q) o[8 * k + q] = Elem<T>::to_acc(e[q]);
            }
        } else if constexpr (CPL == 4) {
            const uint2 r = __ldg(reinterpret_cast<const uint2 *>(p));
            const T *e = reinterpret_cast<const T *>(&r);
#pragma unroll
            for (int q = 0; q < 4; ++q) o[q] = Elem<T>::to_acc(e[q]);
        } else if constexpr (CPL == 2) {
            const uint32_t r = __ldg(reinterpret_cast<const uint32_t *>(p));
            const T *e = reinterpret_cast<const T *>(&r);
            o[0] = Elem<T>::to_acc(e[0]); o[1] = Elem<T>::to_acc(e[1]);
        } else {
            o[0] = Elem<T>::to_acc(__ldg(p));
        }
    }
}
template <typename T, int CPL> __device__ __forceinline__ void mg_store(T *p, const float (&v)[CPL]) {
#pragma unroll
    for (int q = 0; q < CPL; ++q) p[q] = Elem<T>::from_acc(v[q]);  // contiguous per lane; the compiler merges the stores
}
template <typename TL> __device__ __forceinline__ float mg_lp(const TL *lp, int64_t i) {
    return lp ? Elem<TL>::to_acc(__ldg(lp + i)) : 1.f;
}

// weights of one token's slots -> shared memory (w_s [M][4], row_s [M] = neighbour row or -1 for a zero weight)
template <typename TL>
__device__ __forceinline__ void mg_weights(const float *__restrict__ table, const int64_t *__restrict__ pe,
                                           const int64_t *__restrict__ nb, const int64_t *__restrict__ mask,
                                           const TL *__restrict__ lp_b, int M, int lane, float4 *w_s, int *row_s) {
    for (int j = lane; j < M; j += 32) {
        const int r = int(__ldg(nb + j));
        const float4 t = __ldg(reinterpret_cast<const float4 *>(table) + __ldg(pe + j));
        // the reference multiplies lp by the mask in lp's dtype (exact), then the fp32 table row by that value
        const float s = mg_lp<TL>(lp_b, r) * ((mask && __ldg(mask + j) == 0) ? 0.f : 1.f);
        w_s[j] = make_float4(t.x * s, t.y * s, t.z * s, t.w * s);
        row_s[j] = r;
    }
}

template <typename TF, typename TL, int CPL>
__global__ void __launch_bounds__(kMT)
merge_fwd_kernel(const float *__restrict__ table, const int64_t *__restrict__ pe_idx, const int64_t *__restrict__ nbhd,
                 const int64_t *__restrict__ mask, const TL *__restrict__ lp, const TF *__restrict__ feat,
                 float *__restrict__ out, int64_t n_tok, int Nq, int Nk, int M) {
    constexpr int C = CPL * 32;
    extern __shared__ __align__(16) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float4 *w_s = reinterpret_cast<float4 *>(smem) + warp * M;
    int *row_s = reinterpret_cast<int *>(smem + size_t(kMW) * M * 16) + warp * M;
    const int64_t tok = int64_t(blockIdx.x) * kMW + warp;
    if (tok >= n_tok) return;
    const int64_t b = tok / Nq;
    const int64_t e0 = tok * M;
    mg_weights<TL>(table, pe_idx + e0, nbhd + e0, mask ? mask + e0 : nullptr, lp ? lp + b * Nk : nullptr, M, lane, w_s, row_s);
    __syncwarp();
    const TF *fb = feat + b * Nk * C + lane * CPL;
    float acc[4][CPL];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int c = 0; c < CPL; ++c) acc[q][c] = 0.f;
    int j = 0;
    for (; j + 4 <= M; j += 4) {  // four independent row loads in flight
        float f[4][CPL];
#pragma unroll
        for (int u = 0; u < 4; ++u) mg_load<TF, CPL>(fb + int64_t(row_s[j + u]) * C, f[u]);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float4 w = w_s[j + u];
#pragma unroll
            for (int c = 0; c < CPL; ++c) {
                acc[0][c] += w.x * f[u][c]; acc[1][c] += w.y * f[u][c];
                acc[2][c] += w.z * f[u][c]; acc[3][c] += w.w * f[u][c];
            }
        }
    }
    for (; j < M; ++j) {
        float f[CPL];
        mg_load<TF, CPL>(fb + int64_t(row_s[j]) * C, f);
        const float4 w = w_s[j];
#pragma unroll
        for (int c = 0; c < CPL; ++c) {
            acc[0][c] += w.x * f[c]; acc[1][c] += w.y * f[c]; acc[2][c] += w.z * f[c]; acc[3][c] += w.w * f[c];
        }
    }
    float *ob = out + tok * 4 * C + lane * CPL;
#pragma unroll
    for (int q = 0; q < 4; ++q) mg_store<float, CPL>(ob + q * C, acc[q]);
}

// d_wt[b,i,j,:] = (feat[nbhd]·d_out[q]) * lp * mask   and   d_lp[b, nbhd] += mask * sum_q (feat·d_out[q]) * table[pe][q]
template <typename TF, typename TL, int CPL>
__global__ void __launch_bounds__(kMT)
merge_bwd_w_kernel(const float *__restrict__ d_out, const float *__restrict__ table, const int64_t *__restrict__ pe_idx,
                   const int64_t *__restrict__ nbhd, const int64_t *__restrict__ mask, const TL *__restrict__ lp,
                   const TF *__restrict__ feat, float *__restrict__ d_wt, float *__restrict__ d_lp, int64_t n_tok,
                   int Nq, int Nk, int M) {
    constexpr int C = CPL * 32;
    extern __shared__ __align__(16) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float4 *dw_s = reinterpret_cast<float4 *>(smem) + warp * M;
    const int64_t tok = int64_t(blockIdx.x) * kMW + warp;
    if (tok >= n_tok) return;
    const int64_t b = tok / Nq;
    const int64_t e0 = tok * M;
    float g[4][CPL];
    const float *gb = d_out + tok * 4 * C + lane * CPL;
#pragma unroll
    for (int q = 0; q < 4; ++q) mg_load<float, CPL>(gb + q * C, g[q]);
    const TF *fb = feat + b * Nk * C + lane * CPL;
    const int64_t *nb = nbhd + e0;
    for (int j0 = 0; j0 < M; j0 += 2) {  // two slots per round: two row loads in flight, 8 partial sums to reduce
        const bool two = j0 + 1 < M;
        float f0[CPL], f1[CPL];
        mg_load<TF, CPL>(fb + __ldg(nb + j0) * C, f0);
        mg_load<TF, CPL>(fb + __ldg(nb + (two ? j0 + 1 : j0)) * C, f1);
        float v[8];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float a0 = 0.f, a1 = 0.f;
#pragma unroll
            for (int c = 0; c < CPL; ++c) { a0 += g[q][c] * f0[c]; a1 += g[q][c] * f1[c]; }
            v[q] = a0; v[4 + q] = a1;
        }
        // transposing butterfly: 8 -> 4 -> 2 -> 1 values per lane, then two plain steps: 4+2+1+2 = 9 shuffles for 8 sums
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float send = (lane & 16) ? v[k] : v[k + 4], keep = (lane & 16) ? v[k + 4] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const float send = (lane & 8) ? v[k] : v[k + 2], keep = (lane & 8) ? v[k + 2] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
        {
            const float send = (lane & 4) ? v[0] : v[1], keep = (lane & 4) ? v[1] : v[0];
            v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
        v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
        v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
        // lane holds sum index s = (bit4 ? 4 : 0) + (bit3 ? 2 : 0) + (bit2 ? 1 : 0): slot s>>2, component s&3
        if ((lane & 3) == 0) {
            const int s = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
            const int jj = j0 + (s >> 2);
            if (jj < M) reinterpret_cast<float *>(dw_s + jj)[s & 3] = v[0];
        }
    }
    __syncwarp();
    const TL *lp_b = lp ? lp + b * Nk : nullptr;
    for (int j = lane; j < M; j += 32) {
        const int r = int(__ldg(nb + j));
        const float4 dw = dw_s[j];
        const float4 t = __ldg(reinterpret_cast<const float4 *>(table) + __ldg(pe_idx + e0 + j));
        const float mk = (mask && __ldg(mask + e0 + j) == 0) ? 0.f : 1.f;
        const float s = mg_lp<TL>(lp_b, r) * mk;
        reinterpret_cast<float4 *>(d_wt)[e0 + j] = make_float4(dw.x * s, dw.y * s, dw.z * s, dw.w * s);
        if (d_lp && mk != 0.f) atomicAdd(d_lp + b * Nk + r, dw.x * t.x + dw.y * t.y + dw.z * t.z + dw.w * t.w);
    }
}

// d_feat[b,r,:] = lp[b,r] * sum_{e -> r} mask_e * sum_q table[pe_e][q] * d_out[b, i_e, q, :]   (one warp per source row)
template <typename TF, typename TL, int CPL>
__global__ void __launch_bounds__(kMT)
merge_bwd_f_kernel(const float *__restrict__ d_out, const float *__restrict__ table, const int64_t *__restrict__ pe_idx,
                   const int64_t *__restrict__ mask, const TL *__restrict__ lp, const int32_t *__restrict__ off,
                   const int32_t *__restrict__ ent, TF *__restrict__ d_feat, int64_t n_rows, int Nq, int Nk, int M,
                   FastDiv divM) {
    constexpr int C = CPL * 32;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row = int64_t(blockIdx.x) * kMW + warp;
    if (row >= n_rows) return;
    const int64_t b = row / Nk;
    const int r = int(row - b * Nk);
    const int32_t *o = off + b * (Nk + 1) + r;
    const int beg = o[0], end = o[1];
    const int32_t *eb = ent + b * int64_t(Nq) * M;
    const int64_t base = b * int64_t(Nq) * M;
    const float *gb = d_out + b * int64_t(Nq) * 4 * C + lane * CPL;
    float acc[CPL];
#pragma unroll
    for (int c = 0; c < CPL; ++c) acc[c] = 0.f;
    for (int k0 = beg; k0 < end; k0 += 32) {
        // lanes fetch the slot descriptors of up to 32 referencing entries, then the warp walks them
        const int k = k0 + lane;
        int e = 0;
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < end) {
            e = __ldg(eb + k);
            if (!(mask && __ldg(mask + base + e) == 0)) t = __ldg(reinterpret_cast<const float4 *>(table) + __ldg(pe_idx + base + e));
        }
        const int n = min(32, end - k0);
        for (int u = 0; u < n; ++u) {
            const int eu = __shfl_sync(0xffffffffu, e, u);
            const float tx = __shfl_sync(0xffffffffu, t.x, u), ty = __shfl_sync(0xffffffffu, t.y, u);
            const float tz = __shfl_sync(0xffffffffu, t.z, u), tw = __shfl_sync(0xffffffffu, t.w, u);
            const float *gi = gb + int64_t(divM.div(uint32_t(eu))) * 4 * C;
            float g0[CPL], g1[CPL], g2[CPL], g3[CPL];
            mg_load<float, CPL>(gi, g0);
            mg_load<float, CPL>(gi + C, g1);
            mg_load<float, CPL>(gi + 2 * C, g2);
            mg_load<float, CPL>(gi + 3 * C, g3);
#pragma unroll
            for (int c = 0; c < CPL; ++c) acc[c] += tx * g0[c] + ty * g1[c] + tz * g2[c] + tw * g3[c];
        }
    }
    const float s = lp ? Elem<TL>::to_acc(__ldg(lp + row)) : 1.f;
#pragma unroll
    for (int c = 0; c < CPL; ++c) acc[c] *= s;
    mg_store<TF, CPL>(d_feat + row * C + lane * CPL, acc);
}

// The same gradient with a warp per ALIGNED BLOCK OF 8 source rows.  Neighbourhoods are whole clusters of 8 consecutive
// rows (reference models/point_utils.py:203-206, aff_transformer.py:476-480), so the 8 rows of a block are referenced by
// the same kept tokens, in slots j .. j+7 of each: the token's four d_out rows (the bulk of the traffic: 4*C fp32) are
// loaded once and serve all 8 destination rows (the per-row kernel re-reads them 8 times: 7.7 GB through L2 at the first
// merge of AFF-Tiny-1/5).  The structure is checked on the block's own CSR lists first (equal lengths, entry e + u for row
// u); a block without it takes the per-row walk, so any index tensor gives the same sums in the same order.
template <typename TF, typename TL, int CPL>
__global__ void __launch_bounds__(kMT)
merge_bwd_f8_kernel(const float *__restrict__ d_out, const float *__restrict__ table, const int64_t *__restrict__ pe_idx,
                    const int64_t *__restrict__ mask, const TL *__restrict__ lp, const int32_t *__restrict__ off,
                    const int32_t *__restrict__ ent, TF *__restrict__ d_feat, int64_t n_blocks, int Nq, int Nk, int M,
                    FastDiv divM) {
    constexpr int C = CPL * 32;
    __shared__ float4 ts_all[kMW][4][8];
    __shared__ int ti_all[kMW][4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t blk = int64_t(blockIdx.x) * kMW + warp;
    if (blk >= n_blocks) return;
    const int bpi = (Nk + 7) >> 3;
    const int64_t b = blk / bpi;
    const int r0 = int(blk - b * bpi) * 8;
    const int nrow = min(8, Nk - r0);
    const int32_t *o = off + b * (Nk + 1) + r0;
    const int32_t *eb = ent + b * int64_t(Nq) * M;
    const int64_t base = b * int64_t(Nq) * M;
    const float *gb = d_out + b * int64_t(Nq) * 4 * C + lane * CPL;
    const int kk = lane >> 3, u = lane & 7;
    const int beg_u = u < nrow ? o[u] : 0, cnt_u = u < nrow ? o[u + 1] - beg_u : 0;
    const int beg0 = __shfl_sync(0xffffffffu, beg_u, 0), cnt = __shfl_sync(0xffffffffu, cnt_u, 0);
    // structure check
    bool bad = u < nrow && cnt_u != cnt;
    for (int k = kk; k < cnt && !bad; k += 4) {
        const int e0 = __ldg(eb + beg0 + k);
        if (u < nrow && u > 0 && __ldg(eb + beg_u + k) != e0 + u) bad = true;
    }
    if (__any_sync(0xffffffffu, bad)) {     // per-row walk (identical to merge_bwd_f_kernel)
        for (int uu = 0; uu < nrow; ++uu) {
            const int beg = o[uu], end = o[uu + 1];
            float acc[CPL];
#pragma unroll
            for (int c = 0; c < CPL; ++c) acc[c] = 0.f;
            for (int k0 = beg; k0 < end; k0 += 32) {
                const int k = k0 + lane;
                int e = 0;
                float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
                if (k < end) {
                    e = __ldg(eb + k);
                    if (!(mask && __ldg(mask + base + e) == 0)) t = __ldg(reinterpret_cast<const float4 *>(table) + __ldg(pe_idx + base + e));
                }
                const int n = min(32, end - k0);
                for (int v = 0; v < n; ++v) {
                    const int eu = __shfl_sync(0xffffffffu, e, v);
                    const float tx = __shfl_sync(0xffffffffu, t.x, v), ty = __shfl_sync(0xffffffffu, t.y, v);
                    const float tz = __shfl_sync(0xffffffffu, t.z, v), tw = __shfl_sync(0xffffffffu, t.w, v);
                    const float *gi = gb + int64_t(divM.div(uint32_t(eu))) * 4 * C;
                    float g0[CPL], g1[CPL], g2[CPL], g3[CPL];
                    mg_load<float, CPL>(gi, g0);
                    mg_load<float, CPL>(gi + C, g1);
                    mg_load<float, CPL>(gi + 2 * C, g2);
                    mg_load<float, CPL>(gi + 3 * C, g3);
#pragma unroll
                    for (int c = 0; c < CPL; ++c) acc[c] += tx * g0[c] + ty * g1[c] + tz * g2[c] + tw * g3[c];
                }
            }
            const int64_t row = b * Nk + r0 + uu;
            const float sc = lp ? Elem<TL>::to_acc(__ldg(lp + row)) : 1.f;
#pragma unroll
            for (int c = 0; c < CPL; ++c) acc[c] *= sc;
            mg_store<TF, CPL>(d_feat + row * C + lane * CPL, acc);
        }
        return;
    }
    float acc[8][CPL];
#pragma unroll
    for (int uu = 0; uu < 8; ++uu)
#pragma unroll
        for (int c = 0; c < CPL; ++c) acc[uu][c] = 0.f;
    float4(*ts)[8] = ts_all[warp];
    int *ti = ti_all[warp];
    for (int k0 = 0; k0 < cnt; k0 += 4) {
        // lane (kk, u): the weight quadruple of (referencing token k0 + kk, destination row u)
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k0 + kk < cnt) {
            const int e = __ldg(eb + beg0 + k0 + kk) + u;
            if (u < nrow && !(mask && __ldg(mask + base + e) == 0)) t = __ldg(reinterpret_cast<const float4 *>(table) + __ldg(pe_idx + base + e));
            if (u == 0) ti[kk] = int(divM.div(uint32_t(e)));
        }
        ts[kk][u] = t;
        __syncwarp();
        const int n = min(4, cnt - k0);
        for (int v = 0; v < n; ++v) {
            const float *gi = gb + int64_t(ti[v]) * 4 * C;
            float g0[CPL], g1[CPL], g2[CPL], g3[CPL];
            mg_load<float, CPL>(gi, g0);
            mg_load<float, CPL>(gi + C, g1);
            mg_load<float, CPL>(gi + 2 * C, g2);
            mg_load<float, CPL>(gi + 3 * C, g3);
#pragma unroll
            for (int uu = 0; uu < 8; ++uu) {
                const float4 tv = ts[v][uu];
#pragma unroll
                for (int c = 0; c < CPL; ++c) acc[uu][c] += tv.x * g0[c] + tv.y * g1[c] + tv.z * g2[c] + tv.w * g3[c];
            }
        }
        __syncwarp();
    }
#pragma unroll
    for (int uu = 0; uu < 8; ++uu) {
        if (uu < nrow) {
            const int64_t row = b * Nk + r0 + uu;
            const float sc = lp ? Elem<TL>::to_acc(__ldg(lp + row)) : 1.f;
            float outv[CPL];
#pragma unroll
            for (int c = 0; c < CPL; ++c) outv[c] = acc[uu][c] * sc;
            mg_store<TF, CPL>(d_feat + row * C + lane * CPL, outv);
        }
    }
}

template <typename F> static int mg_dispatch(int dtype, F &&f) {
    switch (dtype) {
        case AFFB200_F32: return f(float{});
        case AFFB200_F16: return f(__half{});
        case AFFB200_BF16: return f(__nv_bfloat16{});
        default: set_error("merge: unsupported dtype enum %d", dtype); return AFFB200_E_ARG;
    }
}
template <typename F> static int mg_cpl(int64_t C, F &&f) {
    switch (C) {
        case 32: return f(std::integral_constant<int, 1>{});
        case 64: return f(std::integral_constant<int, 2>{});
        case 128: return f(std::integral_constant<int, 4>{});
        case 256: return f(std::integral_constant<int, 8>{});
        case 512: return f(std::integral_constant<int, 16>{});
        default: set_error("merge: channel count must be 32, 64, 128, 256 or 512"); return AFFB200_E_ARG;
    }
}
static int mg_check(const char *what, int64_t B, int64_t Nq, int64_t Nk, int64_t M) {
    if (B < 0 || Nq < 0 || Nk <= 0 || M <= 0 || M > 1024 || Nq * M >= (int64_t(1) << 31) || Nk >= (int64_t(1) << 31)) {
        set_error("%s: unsupported dimensions", what);
        return AFFB200_E_ARG;
    }
    return AFFB200_OK;
}

}  // namespace affb200

using namespace affb200;

extern "C" int affb200_merge_fwd(const float *table, const int64_t *pe_idx, const int64_t *nbhd, const int64_t *mask,
                                 const void *lp, const void *feat, float *out, int64_t B, int64_t Nq, int64_t Nk,
                                 int64_t M, int64_t C, int lp_dtype, int feat_dtype, void *stream) {
    if (int e = mg_check("merge_fwd", B, Nq, Nk, M)) return e;
    if (B * Nq == 0) return AFFB200_OK;
    if (!table || !pe_idx || !nbhd || !feat || !out || !aligned16(table) || !aligned16(feat) || !aligned16(out)) {
        set_error("merge_fwd: null or misaligned pointer");
        return AFFB200_E_ARG;
    }
    const int64_t n_tok = B * Nq;
    const size_t smem = size_t(kMW) * M * 20;
    auto st = (cudaStream_t)stream;
    return mg_cpl(C, [&](auto cpl) {
        constexpr int CPL = decltype(cpl)::value;
        return mg_dispatch(feat_dtype, [&](auto tf) {
            using TF = decltype(tf);
            return mg_dispatch(lp ? lp_dtype : AFFB200_F32, [&](auto tl) {
                using TL = decltype(tl);
                auto kern = merge_fwd_kernel<TF, TL, CPL>;
                if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
                kern<<<unsigned((n_tok + kMW - 1) / kMW), kMT, smem, st>>>(table, pe_idx, nbhd, mask, (const TL *)lp,
                                                                           (const TF *)feat, out, n_tok, int(Nq), int(Nk), int(M));
                return check_launch("merge_fwd");
            });
        });
    });
}

extern "C" int affb200_merge_bwd(const float *d_out, const float *table, const int64_t *pe_idx, const int64_t *nbhd,
                                 const int64_t *mask, const void *lp, const void *feat, const int32_t *csr_off,
                                 const int32_t *csr_ent, float *d_wt, float *d_lp, void *d_feat, int64_t B, int64_t Nq,
                                 int64_t Nk, int64_t M, int64_t C, int lp_dtype, int feat_dtype, void *stream) {
    if (int e = mg_check("merge_bwd", B, Nq, Nk, M)) return e;
    auto st = (cudaStream_t)stream;
    if (d_lp && cudaMemsetAsync(d_lp, 0, size_t(B * Nk) * 4, st) != cudaSuccess) return check_launch("merge_bwd memset", 0);
    if (B * Nq == 0) {
        if (d_feat && B * Nk > 0) {
            const size_t eb = feat_dtype == AFFB200_F32 ? 4 : 2;
            if (cudaMemsetAsync(d_feat, 0, size_t(B * Nk * C) * eb, st) != cudaSuccess) return check_launch("merge_bwd memset", 0);
        }
        return AFFB200_OK;
    }
    if (!d_out || !table || !pe_idx || !nbhd || !feat || !d_wt || !aligned16(d_out) || !aligned16(table) ||
        !aligned16(feat) || !aligned16(d_wt) || (d_feat && (!csr_off || !csr_ent || !aligned16(d_feat)))) {
        set_error("merge_bwd: null or misaligned pointer");
        return AFFB200_E_ARG;
    }
    const int64_t n_tok = B * Nq, n_rows = B * Nk;
    const size_t smem = size_t(kMW) * M * 16;
    return mg_cpl(C, [&](auto cpl) {
        constexpr int CPL = decltype(cpl)::value;
        return mg_dispatch(feat_dtype, [&](auto tf) {
            using TF = decltype(tf);
            return mg_dispatch(lp ? lp_dtype : AFFB200_F32, [&](auto tl) {
                using TL = decltype(tl);
                auto kw = merge_bwd_w_kernel<TF, TL, CPL>;
                if (smem > 48 * 1024) cudaFuncSetAttribute(kw, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
                kw<<<unsigned((n_tok + kMW - 1) / kMW), kMT, smem, st>>>(d_out, table, pe_idx, nbhd, mask, (const TL *)lp,
                                                                         (const TF *)feat, d_wt, d_lp, n_tok, int(Nq), int(Nk), int(M));
                if (int e = check_launch("merge_bwd_w")) return e;
                if (!d_feat) return int(AFFB200_OK);
                if constexpr (CPL <= 8) {      // a warp per block of 8 source rows: d_out rows loaded once per referencing token
                    const int64_t n_blocks = B * ((Nk + 7) / 8);
                    merge_bwd_f8_kernel<TF, TL, CPL><<<unsigned((n_blocks + kMW - 1) / kMW), kMT, 0, st>>>(
                        d_out, table, pe_idx, mask, (const TL *)lp, csr_off, csr_ent, (TF *)d_feat, n_blocks, int(Nq), int(Nk),
                        int(M), FastDiv{uint32_t(M)});
                } else {
                    merge_bwd_f_kernel<TF, TL, CPL><<<unsigned((n_rows + kMW - 1) / kMW), kMT, 0, st>>>(
                        d_out, table, pe_idx, mask, (const TL *)lp, csr_off, csr_ent, (TF *)d_feat, n_rows, int(Nq), int(Nk), int(M),
                        FastDiv{uint32_t(M)});
                }
                return check_launch("merge_bwd_f");
            });
        });
    });
}
