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
#include "fused_attn.cuh"

namespace affb200 {

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

// Forward.  Loads run one head ahead of the math: Q and K of head h+1 are requested as soon as S = Q·K^T of head h
// has left the staging buffers, V of head h+1 as soon as O = P·V of head h has; every wait is "all but the newest
// group" (cp.async groups retire in order: ... QK_h, V_h, QK_h+1, V_h+1 ...).
template <typename T, int C, int TT, int NE>
__global__ void __launch_bounds__(kFT, NE <= 7 ? 3 : 2)
fused_attn_fwd_kernel(View<T> q, View<T> k, View<T> v, ViewW<T> out, const float *__restrict__ table,
                      const float *__restrict__ blank_k, const float *__restrict__ blank_v,
                      const int32_t *__restrict__ q_hdr, const int32_t *__restrict__ q_rows,
                      const uint8_t *__restrict__ q_loc, const uint16_t *__restrict__ q_pe, float *__restrict__ lse,
                      int H, int N, int M, float scale, FusedSmem L, int TS) {
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
    const TileHdr t = f_load_tile(q_hdr, q_rows, N, TS, L.UC, int(k.sN), rows_s);
    if (t.bad) {  // uniform over the CTA; the host's deferred plan check reports the failed speculation and redoes
                  // the step; until then the tile's outputs are zeros, not uninitialised memory
        const int HS0 = (H + int(gridDim.z) - 1) / int(gridDim.z);
        for (int h = int(blockIdx.z) * HS0; h < min(H, (int(blockIdx.z) + 1) * HS0); ++h) {
            for (int e = tid; e < t.ti * C; e += kFT) out.row(b, t.t0 + e / C, h)[e % C] = Elem<T>::from_acc(0.f);
            for (int e = tid; e < t.ti; e += kFT) lse[(int64_t(b) * H + h) * N + t.t0 + e] = 0.f;
        }
        return;
    }
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
                      FusedSmem L, int TS) {
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
    const TileHdr t = f_load_tile(q_hdr, q_rows, N, TS, L.UC, int(k.sN), rows_s);
    if (t.bad) {  // failed speculation (see the forward kernel): zeros for everything this tile would have written
        const int HS0 = (H + int(gridDim.z) - 1) / int(gridDim.z);
        for (int h = int(blockIdx.z) * HS0; h < min(H, (int(blockIdx.z) + 1) * HS0); ++h) {
            for (int e = tid; e < t.ti * C; e += kFT) dq.row(b, t.t0 + e / C, h)[e % C] = Elem<T>::from_acc(0.f);
            const int64_t o = ((int64_t(b) * H + h) * N + t.t0) * M;
            for (int e = tid; e < t.ti * M; e += kFT) { ds_out[o + e] = Elem<T>::from_acc(0.f); p_out[o + e] = Elem<T>::from_acc(0.f); }
            if (tid < 2 * C) blank_part[((int64_t(b) * gridDim.x + blockIdx.x) * H + h) * 2 * C + tid] = 0.f;
        }
        return;
    }
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
    const int TS = plan->tq, TT = (TS + 15) & ~15;   // tokens per tile; row capacity of the kernel
    const int UC = (umax + 1 + 31) & ~31;
    const FusedSmem L = fused_smem(UC, TT, int(c), false);
    if (L.total > 227 * 1024) { set_error("fused_attn_fwd: tile needs %zu bytes of shared memory", L.total); return AFFB200_E_ARG; }
    dim3 grid(unsigned((N + TS - 1) / TS), unsigned(B));
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
                                         int(H), int(N), int(M), scale, L, TS);                                        \
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
        !blank_part) {
        set_error("fused_attn_bwd: null or misaligned operand");
        return AFFB200_E_ARG;
    }
    if (!kv_ok(k, v, N) || N * q->sN >= (int64_t(1) << 31) || N * dout->sN >= (int64_t(1) << 31)) { set_error("fused_attn_bwd: k/v token strides must match and fit int32"); return AFFB200_E_ARG; }
    if (plan->tk <= 0 || !plan->k_hdr || !plan->k_utok || !plan->k_rloc) {
        set_error("fused_attn_bwd: a key-tile plan is required");
        return AFFB200_E_ARG;
    }
    if (P <= 0 || P > 32768) { set_error("fused_attn_bwd: bias table must have 1..32768 rows"); return AFFB200_E_ARG; }
    const int TS = plan->tq, TT = (TS + 15) & ~15;   // tokens per tile; row capacity of the kernel
    const int UC = (umax + 1 + 31) & ~31;
    const FusedSmem L = fused_smem(UC, TT, int(c), true);
    if (L.total > 227 * 1024) { set_error("fused_attn_bwd: tile needs %zu bytes of shared memory", L.total); return AFFB200_E_ARG; }
    dim3 grid(unsigned((N + TS - 1) / TS), unsigned(B));
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
                                         blank_part, int(H), int(N), int(M), scale, L, TS);                            \
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
            if (!d_table) return AFFB200_OK;   // the caller takes the bias gradient from ds_work (affb200_pe_feat_grad)
            return launch_table_grad_u16(ds_work, q_pe, B, H, N * M, P, d_table, dtype, st);
        } else {
            return AFFB200_E_ARG;
        }
    });
}
