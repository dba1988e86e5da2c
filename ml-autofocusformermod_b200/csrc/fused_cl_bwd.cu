// fused_cl_bwd.cu — backward of the block-structured fused neighbourhood attention (forward: fused_cl_fwd.cu; operator:
// reference models/aff_transformer.py:103-156 under autograd).
//
//   P = exp(S - lse),  dP = dO·V^T,  delta_i = dO_i·O_i,  dS = P∘(dP - delta),
//   dQ = scale·dS·K,   dK = scale·dS^T·Q,   dV = P^T·dO,   d blank_k = scale·sum_i dS_blank[i]·q_i,  d blank_v = sum_i P_blank[i]·dO_i
//
// Two kernels, neither of which materialises P in HBM:
//   fused_cl_bwd_dq_kernel    query-major, the forward's structure (tile of 64 tokens, union blocks staged by TMA, a warp
//                             = 8 tokens x 2 heads): recomputes S and P block by block in registers, forms dP against the
//                             staged V rows, and accumulates dQ += dS·K in registers.  It also emits delta, the blank-token
//                             gradients (per-CTA partial sums) and dS in "ascending block" slot order — the one [B,H,N,M]
//                             tensor left, read by the bias-table gradient (affb200_pe_feat_grad / table histogram).
//   fused_cl_bwd_dkv_kernel   key-major: a warp owns one 8-row block of K|V (both heads of a group) and walks the list of
//                             query tokens that reference it, 16 at a time (their Q and dO rows gathered with one bulk copy
//                             each, double-buffered per warp): S^T = K·Q^T and dP^T = V·dO^T in accumulators, then
//                             dV += P^T·dO and dK += dS^T·Q stay in registers for the whole list.  No atomics, no scatter:
//                             every dK/dV row is written once, in a fixed summation order.
#include "fused_cl.cuh"

namespace affb200 {

int launch_cl_bwd_dkv(const affb200_view *q, const affb200_view *k, const affb200_view *dout, const affb200_view *dk,
                      const float *table, const affb200_cl_plan *plan, const float *lse, const float *delta, int64_t B,
                      int64_t H, int64_t N, int64_t c, int64_t rel_pos_width, int64_t table_width, float scale, int dtype,
                      cudaStream_t st);

// ---------------------------------------------------------------------------------------------------------------------
// dQ kernel
// ---------------------------------------------------------------------------------------------------------------------
template <int C> struct ClGeomB {
    using G = ClGeom<C>;
    static constexpr int OFF_Q = G::OFF_Q;                 // Q tile, later the dQ staging rows
    static constexpr int OFF_DO = OFF_Q + kClT * 128;      // dO tile
    static constexpr int OFF_KV = OFF_DO + kClT * 128;
    __host__ __device__ static constexpr int total(int ucap) { return OFF_KV + (ucap + 1) * G::BLK; }
};

struct ClBwdQParams {
    ClView o, dq;
    const float *table, *blank_k, *blank_v;
    const int32_t *hdr;
    const uint16_t *ucl;
    const unsigned long long *qmask;
    const uint32_t *pos16;
    const float *lse;
    float *delta;          // [B][H][N]
    void *ds;              // [B][H][N][M] of T, slot order = ascending block id
    float *blank_part;     // [grid][8 warps][H][2][C]: running sums of d blank_k / d blank_v per warp
    int32_t *guard;
    int B, H, N, M, n_tiles, ucap, w, tw;
    float scale, scale_log2;
};

template <typename T, int C>
__global__ void __launch_bounds__(kClThreads, 2)
fused_cl_bwd_dq_kernel(const __grid_constant__ CUtensorMap tm_kv, const __grid_constant__ CUtensorMap tm_q,
                       const __grid_constant__ CUtensorMap tm_do, const ClBwdQParams p) {
    using G = ClGeom<C>;
    using GB = ClGeomB<C>;
    constexpr int NT = C / 8, KS = C / 16;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, q4 = lane & 3, r8 = lane & 7;
    const uint32_t sm = dsm_addr();
    const uint32_t bar = sm, s_act = sm + G::OFF_ACT + warp * 64, s_win = sm + G::OFF_WIN, s_pu = sm + G::OFF_PU;
    const uint32_t s_q = sm + GB::OFF_Q, s_do = sm + GB::OFF_DO, s_kv = sm + GB::OFF_KV;
    unsigned char *gen = affb200_dsm;
    const int ucap = p.ucap;
    float *bp = p.blank_part + (int64_t(blockIdx.x) * 8 + warp) * p.H * 2 * C;
    const uint32_t items = uint32_t(p.H / 2) * p.B * p.n_tiles;
    const T *op = static_cast<const T *>(p.o.p);
    T *dqp = static_cast<T *>(const_cast<void *>(p.dq.p));
    T *dsp = static_cast<T *>(p.ds);
    const int64_t npad = int64_t(p.n_tiles) * kClT;

    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    for (int e = tid; e < G::BLK / 16; e += kClThreads)
        *reinterpret_cast<uint4 *>(gen + GB::OFF_KV + ucap * G::BLK + e * 16) = make_uint4(0u, 0u, 0u, 0u);
    for (int e = tid; e < 8 * p.H * 2 * C; e += kClThreads) p.blank_part[int64_t(blockIdx.x) * 8 * p.H * 2 * C + e] = 0.f;
    __syncthreads();

    uint32_t phase = 0;
    int cur_hg = -1;
    uint32_t item = blockIdx.x;
    int hdr_n = 0;
    uint32_t ucl_n = 0;
    if (item < items) {
        const uint32_t bt = item % (uint32_t(p.B) * p.n_tiles);
        hdr_n = __ldg(p.hdr + bt);
        if (tid < kClUMax) ucl_n = __ldg(p.ucl + int64_t(bt) * kClUMax + tid);
    }
    while (item < items) {
        const int hdr = hdr_n;
        const int blk_mine = int(ucl_n);
        const int tile = int(item % uint32_t(p.n_tiles));
        const uint32_t ib = item / uint32_t(p.n_tiles);
        const int b = int(ib % uint32_t(p.B)), hg = int(ib / uint32_t(p.B));
        const int h0 = hg * 2;
        const int t0 = tile * kClT, ti = min(kClT, p.N - t0);
        const uint32_t next = item + gridDim.x;
        if (next < items) {
            const uint32_t bt = next % (uint32_t(p.B) * p.n_tiles);
            hdr_n = __ldg(p.hdr + bt);
            if (tid < kClUMax) ucl_n = __ldg(p.ucl + int64_t(bt) * kClUMax + tid);
        }
        item = next;
        const int U = hdr & AFFB200_PLAN_COUNT_MASK;
        // (split tiles are a forward-only feature: a plan that carries them is not handed to the backward)
        const bool skip = (hdr & (AFFB200_PLAN_OVERFLOW | AFFB200_CL_BAD | AFFB200_CL_SPLIT)) != 0 || U > ucap;
        T *dq_tile = dqp + int64_t(b) * p.dq.sB + int64_t(t0) * p.dq.sN + h0 * p.dq.sH;
        const int tok = warp * 8 + g;
        const bool live = tok < ti;
        if (skip) {
            for (int e = tid; e < ti * (G::QB / 16); e += kClThreads) {
                const int row = e / (G::QB / 16), ch = e % (G::QB / 16);
                *reinterpret_cast<uint4 *>(dq_tile + int64_t(row) * p.dq.sN + ch * 8) = make_uint4(0u, 0u, 0u, 0u);
            }
            for (int e = tid; e < ti * 2; e += kClThreads)
                p.delta[(int64_t(b) * p.H + h0 + e / ti) * p.N + t0 + e % ti] = 0.f;
            for (int e = tid; e < ti * 2 * (p.M / 8); e += kClThreads) {   // dS rows of the tile <- 0
                const int hh = e / (ti * (p.M / 8)), rem = e % (ti * (p.M / 8));
                *reinterpret_cast<uint4 *>(dsp + ((int64_t(b) * p.H + h0 + hh) * p.N + t0) * p.M + int64_t(rem) * 8) = make_uint4(0u, 0u, 0u, 0u);
            }
            if (tid == 0 && hg == 0) atomicAdd(p.guard, 1);
            continue;
        }
        if (hg != cur_hg) {
            cur_hg = hg;
            cl_fill_window<2>(reinterpret_cast<float *>(gen + G::OFF_WIN), p.table, p.H, h0, p.w, p.tw);
            for (int e = tid; e < 4 * C; e += kClThreads) {
                const int hh = e / (2 * C), rr = e % (2 * C);
                const float v = rr < C ? __ldg(p.blank_k + (h0 + hh) * C + rr) : __ldg(p.blank_v + (h0 + hh) * C + rr - C);
                reinterpret_cast<T *>(gen + GB::OFF_KV + ucap * G::BLK + (C == 32 ? hh * 1024 : hh * 4 * C))[rr] = Elem<T>::from_acc(v);
            }
            __syncthreads();
        }
        if (tid == 0) mbar_expect_tx(bar, uint32_t(U) * (G::BLK + 32u) + 2u * kClT * G::QB + 512u + 256u);
        if (tid < U) {
#pragma unroll
            for (int t = 0; t < G::TPB; ++t)
                tma_load_3d(s_kv + (tid * G::TPB + t) * 1024, &tm_kv, h0 * 2 * C + t * 64, blk_mine * 8, b, bar);
            bulk_g2s(s_pu + tid * 32, p.pos16 + int64_t(b) * npad + blk_mine * 8, 32u, bar);
        } else if (tid == 64) {
            tma_load_3d(s_q, &tm_q, h0 * C, t0, b, bar);
        } else if (tid == 65) {
            bulk_g2s(sm + G::OFF_MK, p.qmask + int64_t(b) * npad + t0, 512u, bar);
        } else if (tid == 66) {
            bulk_g2s(sm + G::OFF_PI, p.pos16 + int64_t(b) * npad + t0, 256u, bar);
        } else if (tid == 67) {
            tma_load_3d(s_do, &tm_do, h0 * C, t0, b, bar);
        }
        // ---- per-(token, head) scalars from global while the copies fly -----------------------------------------
        float lse2[2], dlt[2];
        uint4 orow[2];   // this lane's eighth (C = 32) / quarter-pair (C = 16) of the token's O rows
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            lse2[hh] = live ? __ldg(p.lse + (int64_t(b) * p.H + h0 + hh) * p.N + t0 + tok) * kLog2e : 0.f;
            orow[hh] = make_uint4(0u, 0u, 0u, 0u);
            if (live) {
                const T *src = op + int64_t(b) * p.o.sB + int64_t(t0 + tok) * p.o.sN + (h0 + hh) * p.o.sH + q4 * (C / 4);
                if constexpr (C == 32) orow[hh] = __ldg(reinterpret_cast<const uint4 *>(src));
                else { const uint2 v = __ldg(reinterpret_cast<const uint2 *>(src)); orow[hh].x = v.x; orow[hh].y = v.y; }
            }
        }
        mbar_wait(bar, phase);
        phase ^= 1u;

        unsigned long long mk;
        {
            const uint2 mw = lds_u64(sm + G::OFF_MK + tok * 8);
            mk = (static_cast<unsigned long long>(mw.y) << 32) | mw.x;
        }
        const uint32_t pib = lds_u32(sm + G::OFF_PI + tok * 4) - kClPosBias;
        int n_act;
        {
            const uint32_t alo = __reduce_or_sync(0xffffffffu, uint32_t(mk)), ahi = __reduce_or_sync(0xffffffffu, uint32_t(mk >> 32));
            const uint32_t lt = (1u << lane) - 1u;
            const int clo = __popc(alo);
            if ((alo >> lane) & 1u) gen[G::OFF_ACT + warp * 64 + __popc(alo & lt)] = uint8_t(lane);
            if ((ahi >> lane) & 1u) gen[G::OFF_ACT + warp * 64 + clo + __popc(ahi & lt)] = uint8_t(32 + lane);
            n_act = clo + __popc(ahi);
        }
        // delta = dO·O per (token, head): this lane's C/4 elements, then the quad
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const uint32_t a = s_do + cl_tile_off<C>(tok, hh * (C / 8) + (C == 32 ? q4 : q4 >> 1)) + (C == 32 ? 0 : (q4 & 1) * 8);
            uint4 dv = make_uint4(0u, 0u, 0u, 0u);
            if constexpr (C == 32) {
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(dv.x), "=r"(dv.y), "=r"(dv.z), "=r"(dv.w) : "r"(a));
            } else {
                const uint2 v = lds_u64(a);
                dv.x = v.x; dv.y = v.y;
            }
            const T *x = reinterpret_cast<const T *>(&dv), *y = reinterpret_cast<const T *>(&orow[hh]);
            float acc = 0.f;
#pragma unroll
            for (int e = 0; e < C / 4; ++e) acc = fmaf(Elem<T>::to_acc(x[e]), Elem<T>::to_acc(y[e]), acc);
            acc += __shfl_xor_sync(0xffffffffu, acc, 1);
            acc += __shfl_xor_sync(0xffffffffu, acc, 2);
            dlt[hh] = acc;
            if (q4 == 0 && live) p.delta[(int64_t(b) * p.H + h0 + hh) * p.N + t0 + tok] = acc;
        }
        uint32_t qa[KS][4], da[KS][4];
        cl_load_afrag<C>(qa, s_q, warp, lane);
        cl_load_afrag<C>(da, s_do, warp, lane);
        float dq0[NT][4], dq1[NT][4];
        {
            // ---- blank token: P_b, dS_b; dQ starts at dS_b·blank_k; running sums of d blank_k / d blank_v -----------
            const uint32_t bb = s_kv + ucap * G::BLK;
            float s0[4], s1[4], d0[4], d1[4];
            cl_scores<T, C, false>(s0, s1, qa, bb, lane);
            cl_scores<T, C, true>(d0, d1, da, bb, lane);
            float pb[2], dsb[2];
            pb[0] = f_ex2(__shfl_sync(0xffffffffu, s0[0], lane & ~3) * p.scale_log2 - lse2[0]);
            pb[1] = f_ex2(__shfl_sync(0xffffffffu, s1[2], lane & ~3) * p.scale_log2 - lse2[1]);
            dsb[0] = pb[0] * (__shfl_sync(0xffffffffu, d0[0], lane & ~3) - dlt[0]);
            dsb[1] = pb[1] * (__shfl_sync(0xffffffffu, d1[2], lane & ~3) - dlt[1]);
            if (!live) { pb[0] = pb[1] = dsb[0] = dsb[1] = 0.f; }
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const uint32_t a0 = nt * 16 + q4 * 4;   // blank_k as staged (rounded to T): row 0 of the blank block, K chunks
                const uint32_t k0 = lds_u32(bb + a0), k1 = lds_u32(bb + (C == 32 ? 1024 : 4 * C) + a0);
                const T *e0 = reinterpret_cast<const T *>(&k0), *e1 = reinterpret_cast<const T *>(&k1);
                dq0[nt][0] = dsb[0] * Elem<T>::to_acc(e0[0]); dq0[nt][1] = dsb[0] * Elem<T>::to_acc(e0[1]); dq0[nt][2] = dq0[nt][3] = 0.f;
                dq1[nt][2] = dsb[1] * Elem<T>::to_acc(e1[0]); dq1[nt][3] = dsb[1] * Elem<T>::to_acc(e1[1]); dq1[nt][0] = dq1[nt][1] = 0.f;
            }
            // column sums over the warp's 8 tokens with one m16n8k16 per (tensor, head, n-tile): A row 0 = the head-0
            // weights of tokens 0..7 (k), row 8 = the head-1 weights; B = the tokens' dO (or Q) rows through ldmatrix.trans
            uint32_t aw[2][4];   // [0]: P_b (for d blank_v), [1]: dS_b (for d blank_k)
#pragma unroll
            for (int kind = 0; kind < 2; ++kind) {
                const float w0 = kind == 0 ? pb[0] : dsb[0], w1 = kind == 0 ? pb[1] : dsb[1];
                const float a00 = __shfl_sync(0xffffffffu, w0, q4 * 8), a01 = __shfl_sync(0xffffffffu, w0, q4 * 8 + 4);
                const float a10 = __shfl_sync(0xffffffffu, w1, q4 * 8), a11 = __shfl_sync(0xffffffffu, w1, q4 * 8 + 4);
                aw[kind][0] = g == 0 ? f_pack2<T>(a00, a01) : 0u;
                aw[kind][1] = g == 0 ? f_pack2<T>(a10, a11) : 0u;
                aw[kind][2] = aw[kind][3] = 0u;
            }
#pragma unroll
            for (int kind = 0; kind < 2; ++kind) {
                const uint32_t tile = kind == 0 ? s_do : s_q;
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    uint32_t bf[4];
                    const int row = warp * 8 + r8;
                    if constexpr (C == 32) cl_ldsm_x4_t(bf, tile + cl_tile_off<C>(row, hh * 4 + (lane >> 3)));
                    else cl_ldsm_x4_t(bf, tile + cl_tile_off<C>(row, hh * 2 + ((lane >> 3) & 1)));
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) {
                        float acc[4] = {0.f, 0.f, 0.f, 0.f};
                        f_mma<T>(acc, aw[kind], bf[nt], 0u);
                        if (g == 0) {   // kind 0 -> d blank_v (index 1), kind 1 -> d blank_k (index 0)
                            float *dst = bp + ((h0 + hh) * 2 + (1 - kind)) * C + nt * 8 + q4 * 2;
                            const float sc = kind == 0 ? 1.f : p.scale;   // d blank_k carries the logit scale
                            dst[0] += acc[2 * hh] * sc;
                            dst[1] += acc[2 * hh + 1] * sc;
                        }
                    }
                }
            }
        }
        __syncwarp();

        T *ds_tok = dsp + ((int64_t(b) * p.H + h0) * p.N + t0 + tok) * p.M;
        const int64_t ds_head = int64_t(p.N) * p.M;
        {
            // slots behind the token's real blocks (fully padded groups of the last cluster): dS = 0
            const int first = __popcll(mk) * 8;
            if (live)
                for (int e = first + q4 * 2; e < p.M; e += 8) {
                    *reinterpret_cast<uint32_t *>(ds_tok + e) = 0u;
                    *reinterpret_cast<uint32_t *>(ds_tok + ds_head + e) = 0u;
                }
        }
        for (int k = 0; k < n_act; k += 2) {
            const uint32_t two = lds_u16(s_act + k);
            const bool has1 = k + 1 < n_act;
            int pj[2];
            pj[0] = int(two & 0xffu);
            pj[1] = has1 ? int(two >> 8) : pj[0];
            bool far[2][2], anyfar = false, mem[2];
            float2 bias[2][2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const uint2 pu = lds_u64(s_pu + (pj[j] * 8 + q4 * 2) * 4);
                mem[j] = ((mk >> pj[j]) & 1ull) != 0ull && (j == 0 || has1);
                const uint32_t i0 = cl_win_index(pu.x, pib, mem[j], far[j][0]), i1 = cl_win_index(pu.y, pib, mem[j], far[j][1]);
                bias[j][0] = lds_f2(s_win + i0 * 8u);
                bias[j][1] = lds_f2(s_win + i1 * 8u);
                anyfar = anyfar || far[j][0] || far[j][1];
            }
            if (__any_sync(0xffffffffu, anyfar)) {
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const uint2 pu = lds_u64(s_pu + (pj[j] * 8 + q4 * 2) * 4);
#pragma unroll
                    for (int cc = 0; cc < 2; ++cc)
                        if (far[j][cc]) {
                            bias[j][cc].x = cl_bias_far(p.table, p.H, h0, p.w, p.tw, cc ? pu.y : pu.x, pib);
                            bias[j][cc].y = cl_bias_far(p.table, p.H, h0 + 1, p.w, p.tw, cc ? pu.y : pu.x, pib);
                        }
                }
            }
            float s0[2][4], s1[2][4], d0[2][4], d1[2][4];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                cl_scores<T, C, false>(s0[j], s1[j], qa, s_kv + pj[j] * G::BLK, lane);
                cl_scores<T, C, true>(d0[j], d1[j], da, s_kv + pj[j] * G::BLK, lane);
            }
            uint32_t pa[4];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                float ds0[2], ds1[2];
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const float p0 = f_ex2(fmaf(s0[j][cc], p.scale_log2, bias[j][cc].x) - lse2[0]);
                    const float p1 = f_ex2(fmaf(s1[j][2 + cc], p.scale_log2, bias[j][cc].y) - lse2[1]);
                    ds0[cc] = p0 * (d0[j][cc] - dlt[0]);
                    ds1[cc] = p1 * (d1[j][2 + cc] - dlt[1]);
                }
                pa[2 * j] = f_pack2<T>(ds0[0], ds0[1]);
                pa[2 * j + 1] = f_pack2<T>(ds1[0], ds1[1]);
                if (mem[j] && live) {   // dS in ascending-block slot order, for the bias-table gradient
                    const unsigned long long below = mk & ((1ull << pj[j]) - 1ull);
                    const int slot = __popcll(below) * 8 + q4 * 2;
                    *reinterpret_cast<uint32_t *>(ds_tok + slot) = pa[2 * j];
                    *reinterpret_cast<uint32_t *>(ds_tok + ds_head + slot) = pa[2 * j + 1];
                }
            }
            cl_accum_rows<T, C, false>(dq0, dq1, pa, s_kv + pj[(lane >> 3) & 1] * G::BLK, lane);
        }

        // ---- dQ = scale * acc through the warp's Q rows ---------------------------------------------------------
#pragma unroll
        for (int hh = 0; hh < 2; ++hh)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const float x0 = hh == 0 ? dq0[nt][0] : dq1[nt][2], x1 = hh == 0 ? dq0[nt][1] : dq1[nt][3];
                *reinterpret_cast<uint32_t *>(gen + GB::OFF_Q + cl_tile_off<C>(tok, hh * (C / 8) + nt) + q4 * 4) =
                    f_pack2<T>(x0 * p.scale, x1 * p.scale);
            }
        __syncwarp();
#pragma unroll
        for (int e = lane; e < 8 * (G::QB / 16); e += 32) {
            const int row = warp * 8 + e / (G::QB / 16), ch = e % (G::QB / 16);
            if (row < ti)
                *reinterpret_cast<uint4 *>(dq_tile + int64_t(row) * p.dq.sN + ch * 8) =
                    *reinterpret_cast<const uint4 *>(gen + GB::OFF_Q + cl_tile_off<C>(row, ch));
        }
        fence_proxy_async();
        __syncthreads();
    }
}

template <typename T, int C>
static int launch_cl_bwd_dq(const CUtensorMap &tm_kv, const CUtensorMap &tm_q, const CUtensorMap &tm_do,
                            const ClBwdQParams &p, int grid, cudaStream_t st) {
    const int smem = ClGeomB<C>::total(p.ucap);
    if (smem > 227 * 1024) { set_error("fused_cl_bwd: union of %d blocks does not fit shared memory", p.ucap); return AFFB200_E_ARG; }
    auto kern = fused_cl_bwd_dq_kernel<T, C>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("fused_cl_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return AFFB200_E_CUDA; }
    kern<<<grid, kClThreads, smem, st>>>(tm_kv, tm_q, tm_do, p);
    return check_launch("fused_cl_bwd_dq");
}


// ---------------------------------------------------------------------------------------------------------------------
// dK / dV kernel
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kKvThreads = 128;   // 4 independent warps per CTA, three CTAs per SM

template <int C> struct ClGeomK {
    using G = ClGeom<C>;
    static constexpr int QS = G::QB + 16;              // padded row of a gathered Q / dO row (conflict-free ldmatrix)
    static constexpr int OFF_WIN = 0;
    static constexpr int OFF_WARP = 9216;              // 9 x 1024: per-warp regions start 1024-aligned
    static constexpr int W_KV = 0;                     // the warp's K|V block (TMA, swizzled); later dK|dV staging
    static constexpr int W_ST = G::BLK;                // 2 stages x (16 Q rows + 16 dO rows)
    static constexpr int W_SC = W_ST + 2 * 32 * QS;    // 2 stages x (16 float4 scalars + 16 packed positions)
    static constexpr int W_PU = W_SC + 2 * 320;        // packed positions of the block's 8 rows
    static constexpr int W_BAR = W_PU + 32;            // kv barrier, 2 stage barriers
    static constexpr int W_BYTES = (W_BAR + 32 + 1023) / 1024 * 1024;
    static constexpr int TOTAL = OFF_WARP + 4 * W_BYTES;
};

struct ClBwdKParams {
    ClView q, dout, dkv;
    const float *table;
    const int32_t *kb_cnt;
    const uint32_t *kb_tok;
    const uint32_t *pos16;
    const float *lse, *delta;
    int kbcap, B, H, N, w, tw;
    int64_t npad;
    float scale, scale_log2;
};

template <typename T, int C>
__global__ void __launch_bounds__(kKvThreads, 3)
fused_cl_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tm_kv, const ClBwdKParams p) {
    using G = ClGeom<C>;
    using GK = ClGeomK<C>;
    constexpr int NT = C / 8, KS = C / 16, QS = GK::QS;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, q4 = lane & 3, r8 = lane & 7;
    const uint32_t sm = dsm_addr();
    unsigned char *gen = affb200_dsm;
    const uint32_t s_win = sm + GK::OFF_WIN;
    const uint32_t wbase = sm + GK::OFF_WARP + warp * GK::W_BYTES;
    unsigned char *wgen = gen + GK::OFF_WARP + warp * GK::W_BYTES;
    const uint32_t s_kv = wbase + GK::W_KV, s_st = wbase + GK::W_ST, s_sc = wbase + GK::W_SC, s_pu = wbase + GK::W_PU;
    const uint32_t kvbar = wbase + GK::W_BAR, fullbar = kvbar + 8;
    const int nblk = (p.N + 7) >> 3;
    const T *qp = static_cast<const T *>(p.q.p), *dop = static_cast<const T *>(p.dout.p);
    T *dkvp = static_cast<T *>(const_cast<void *>(p.dkv.p));

    if (lane == 0) {
        mbar_init(kvbar, 1);
        mbar_init(fullbar, 1);
        mbar_init(fullbar + 8, 1);
        mbar_fence_init();
    }
    // gathered rows of a short last batch keep whatever the stage held before: start from zeros (0 x finite = 0)
    for (int e = lane; e < 2 * 32 * QS / 16; e += 32) *reinterpret_cast<uint4 *>(wgen + GK::W_ST + e * 16) = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async();
    __syncwarp();
    uint32_t kvphase = 0, stphase[2] = {0u, 0u};
    const uint32_t items = uint32_t(p.B) * nblk, gw = blockIdx.x * 4 + warp, nw = gridDim.x * 4;

    for (int hg = 0; hg < p.H / 2; ++hg) {
        const int h0 = hg * 2;
        __syncthreads();
        cl_fill_window<2>(reinterpret_cast<float *>(gen + GK::OFF_WIN), p.table, p.H, h0, p.w, p.tw);
        __syncthreads();
        for (uint32_t item = gw; item < items; item += nw) {
            const int b = int(item / uint32_t(nblk)), blk = int(item % uint32_t(nblk));
            const int n = min(__ldg(p.kb_cnt + item), p.kbcap);
            const uint32_t *list = p.kb_tok + int64_t(item) * p.kbcap;
            const T *q_img = qp + int64_t(b) * p.q.sB + h0 * p.q.sH, *do_img = dop + int64_t(b) * p.dout.sB + h0 * p.dout.sH;
            const float *lse_img = p.lse + (int64_t(b) * p.H + h0) * p.N, *dl_img = p.delta + (int64_t(b) * p.H + h0) * p.N;
            if (lane == 0) {
                mbar_expect_tx(kvbar, G::BLK + 32u);
#pragma unroll
                for (int t = 0; t < G::TPB; ++t) tma_load_3d(s_kv + t * 1024, &tm_kv, h0 * 2 * C + t * 64, blk * 8, b, kvbar);
                bulk_g2s(s_pu, p.pos16 + int64_t(b) * p.npad + blk * 8, 32u, kvbar);
            }
            const int nb = (n + 15) >> 4;
            // batch kb: gathers of 16 query rows (Q, dO) into stage kb & 1; the per-query scalars travel in registers
            float4 sc_r = make_float4(0.f, 0.f, 0.f, 0.f);
            uint32_t pos_r = 0u;
            auto issue_batch = [&](int kb) {
                const int st = kb & 1, idx = kb * 16 + lane;
                const bool valid = lane < 16 && idx < n;
                const int nvalid = min(16, n - kb * 16);
                if (lane == 0) mbar_expect_tx(fullbar + st * 8, uint32_t(nvalid) * 2u * G::QB);
                __syncwarp();
                sc_r = make_float4(INFINITY, INFINITY, 0.f, 0.f);   // padding queries: P = 2^(x - inf) = 0
                pos_r = 0u;
                if (valid) {
                    const int tok = int(__ldg(list + idx));
                    bulk_g2s(s_st + st * 32 * QS + lane * QS, q_img + int64_t(tok) * p.q.sN, G::QB, fullbar + st * 8);
                    bulk_g2s(s_st + st * 32 * QS + (16 + lane) * QS, do_img + int64_t(tok) * p.dout.sN, G::QB, fullbar + st * 8);
                    sc_r.x = __ldg(lse_img + tok) * kLog2e;
                    sc_r.y = __ldg(lse_img + p.N + tok) * kLog2e;
                    sc_r.z = __ldg(dl_img + tok);
                    sc_r.w = __ldg(dl_img + p.N + tok);
                    pos_r = __ldg(p.pos16 + int64_t(b) * p.npad + tok) - kClPosBias;
                }
            };
            auto store_scalars = [&](int kb) {
                const int st = kb & 1;
                if (lane < 16) {
                    *reinterpret_cast<float4 *>(wgen + GK::W_SC + st * 320 + lane * 16) = sc_r;
                    *reinterpret_cast<uint32_t *>(wgen + GK::W_SC + st * 320 + 256 + lane * 4) = pos_r;
                }
            };
            if (nb > 0) { issue_batch(0); store_scalars(0); }
            mbar_wait(kvbar, kvphase);
            kvphase ^= 1u;
            // A fragments of the block: rows 0-7 = its rows under head 0, rows 8-15 = under head 1
            uint32_t ka[KS][4], va[KS][4];
            {
                const int i = lane >> 3;
#pragma unroll
                for (int ks = 0; ks < KS; ++ks) {
                    if constexpr (C == 32) {
                        cl_ldsm_x4(ka[ks], s_kv + (i & 1) * 1024 + cl_sw(r8, 2 * ks + (i >> 1)));
                        cl_ldsm_x4(va[ks], s_kv + (i & 1) * 1024 + cl_sw(r8, 4 + 2 * ks + (i >> 1)));
                    } else {
                        cl_ldsm_x4(ka[ks], s_kv + cl_sw(r8, (i & 1) * 4 + (i >> 1)));
                        cl_ldsm_x4(va[ks], s_kv + cl_sw(r8, (i & 1) * 4 + 2 + (i >> 1)));
                    }
                }
            }
            const uint32_t pu = lds_u32(s_pu + g * 4);   // this thread's key row (both heads)
            float dk0[NT][4], dk1[NT][4], dv0[NT][4], dv1[NT][4];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int e = 0; e < 4; ++e) dk0[nt][e] = dk1[nt][e] = dv0[nt][e] = dv1[nt][e] = 0.f;
            __syncwarp();

            for (int kb = 0; kb < nb; ++kb) {
                const int st = kb & 1;
                if (kb + 1 < nb) issue_batch(kb + 1);   // the other stage was consumed in the previous iteration
                mbar_wait(fullbar + st * 8, stphase[st]);
                stphase[st] ^= 1u;
                __syncwarp();   // the scalars of this stage were stored by lanes 0-15 at the end of the previous iteration
                const uint32_t qst = s_st + st * 32 * QS, dst = qst + 16 * QS, scs = s_sc + st * 320;
                float s0[2][4], s1[2][4], d0[2][4], d1[2][4];
#pragma unroll
                for (int j = 0; j < 2; ++j) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) s0[j][e] = s1[j][e] = d0[j][e] = d1[j][e] = 0.f;
                    uint32_t f[4];
                    if constexpr (C == 32) {
                        const uint32_t ro = (j * 8 + r8) * QS + (lane >> 3) * 16;
                        cl_ldsm_x4(f, qst + ro);
                        f_mma<T>(s0[j], ka[0], f[0], f[1]); f_mma<T>(s0[j], ka[1], f[2], f[3]);
                        cl_ldsm_x4(f, qst + ro + 64);
                        f_mma<T>(s1[j], ka[0], f[0], f[1]); f_mma<T>(s1[j], ka[1], f[2], f[3]);
                        cl_ldsm_x4(f, dst + ro);
                        f_mma<T>(d0[j], va[0], f[0], f[1]); f_mma<T>(d0[j], va[1], f[2], f[3]);
                        cl_ldsm_x4(f, dst + ro + 64);
                        f_mma<T>(d1[j], va[0], f[0], f[1]); f_mma<T>(d1[j], va[1], f[2], f[3]);
                    } else {
                        const uint32_t ro = (j * 8 + r8) * QS + (lane >> 4) * 32 + ((lane >> 3) & 1) * 16;
                        cl_ldsm_x4(f, qst + ro);
                        f_mma<T>(s0[j], ka[0], f[0], f[1]); f_mma<T>(s1[j], ka[0], f[2], f[3]);
                        cl_ldsm_x4(f, dst + ro);
                        f_mma<T>(d0[j], va[0], f[0], f[1]); f_mma<T>(d1[j], va[0], f[2], f[3]);
                    }
                }
                uint32_t pp[4], pd[4];
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    float pr0[2], pr1[2], ds0[2], ds1[2];
                    const uint2 pq = lds_u64(scs + 256 + (j * 8 + q4 * 2) * 4);
#pragma unroll
                    for (int cc = 0; cc < 2; ++cc) {
                        const int t = j * 8 + q4 * 2 + cc;
                        float4 sc;
                        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(sc.x), "=f"(sc.y), "=f"(sc.z), "=f"(sc.w) : "r"(scs + t * 16));
                        const uint32_t pib = cc ? pq.y : pq.x;
                        const bool mem = kb * 16 + t < n;
                        bool far;
                        const uint32_t wi = cl_win_index(pu, pib, mem, far);
                        float2 bias = lds_f2(s_win + wi * 8u);
                        if (far) {
                            bias.x = cl_bias_far(p.table, p.H, h0, p.w, p.tw, pu, pib);
                            bias.y = cl_bias_far(p.table, p.H, h0 + 1, p.w, p.tw, pu, pib);
                        }
                        pr0[cc] = f_ex2(fmaf(s0[j][cc], p.scale_log2, bias.x) - sc.x);
                        pr1[cc] = f_ex2(fmaf(s1[j][2 + cc], p.scale_log2, bias.y) - sc.y);
                        ds0[cc] = pr0[cc] * (d0[j][cc] - sc.z);
                        ds1[cc] = pr1[cc] * (d1[j][2 + cc] - sc.w);
                    }
                    pp[2 * j] = f_pack2<T>(pr0[0], pr0[1]); pp[2 * j + 1] = f_pack2<T>(pr1[0], pr1[1]);
                    pd[2 * j] = f_pack2<T>(ds0[0], ds0[1]); pd[2 * j + 1] = f_pack2<T>(ds1[0], ds1[1]);
                }
                // dV += P^T·dO, dK += dS^T·Q: B = the 16 gathered rows through ldmatrix.trans
                {
                    const uint32_t ro = (((lane >> 3) & 1) * 8 + r8) * QS + (lane >> 4) * 16;
                    uint32_t f[4];
#pragma unroll
                    for (int np = 0; np < C / 16; ++np) {
                        cl_ldsm_x4_t(f, dst + ro + np * 32);
                        f_mma<T>(dv0[2 * np], pp, f[0], f[1]); f_mma<T>(dv0[2 * np + 1], pp, f[2], f[3]);
                        cl_ldsm_x4_t(f, dst + ro + C * 2 + np * 32);
                        f_mma<T>(dv1[2 * np], pp, f[0], f[1]); f_mma<T>(dv1[2 * np + 1], pp, f[2], f[3]);
                        cl_ldsm_x4_t(f, qst + ro + np * 32);
                        f_mma<T>(dk0[2 * np], pd, f[0], f[1]); f_mma<T>(dk0[2 * np + 1], pd, f[2], f[3]);
                        cl_ldsm_x4_t(f, qst + ro + C * 2 + np * 32);
                        f_mma<T>(dk1[2 * np], pd, f[0], f[1]); f_mma<T>(dk1[2 * np + 1], pd, f[2], f[3]);
                    }
                }
                __syncwarp();                 // every lane is done with this stage's rows and scalars
                fence_proxy_async();
                if (kb + 1 < nb) store_scalars(kb + 1);
            }

            // ---- dK (scaled) and dV of the block's 8 rows, staged in the block's own tile layout ---------------------
            __syncwarp();
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const int kc = nt, vc = (C == 32 ? 4 : 2) + nt;
                const uint32_t t1 = C == 32 ? 1024u : 0u;
                const int hc = C == 32 ? 0 : 4;   // head 1's chunks follow head 0's inside the single tile at head dim 16
                *reinterpret_cast<uint32_t *>(wgen + GK::W_KV + cl_sw(g, kc) + q4 * 4) = f_pack2<T>(dk0[nt][0] * p.scale, dk0[nt][1] * p.scale);
                *reinterpret_cast<uint32_t *>(wgen + GK::W_KV + cl_sw(g, vc) + q4 * 4) = f_pack2<T>(dv0[nt][0], dv0[nt][1]);
                *reinterpret_cast<uint32_t *>(wgen + GK::W_KV + t1 + cl_sw(g, hc + kc) + q4 * 4) = f_pack2<T>(dk1[nt][2] * p.scale, dk1[nt][3] * p.scale);
                *reinterpret_cast<uint32_t *>(wgen + GK::W_KV + t1 + cl_sw(g, hc + vc) + q4 * 4) = f_pack2<T>(dv1[nt][2], dv1[nt][3]);
            }
            __syncwarp();
            T *dst_img = dkvp + int64_t(b) * p.dkv.sB + h0 * p.dkv.sH;
#pragma unroll
            for (int e = lane; e < 8 * (G::KVB / 16); e += 32) {
                const int row = e / (G::KVB / 16), ch = e % (G::KVB / 16);
                if (blk * 8 + row < p.N)
                    *reinterpret_cast<uint4 *>(dst_img + int64_t(blk * 8 + row) * p.dkv.sN + ch * 8) =
                        *reinterpret_cast<const uint4 *>(wgen + GK::W_KV + (ch >> 3) * 1024 + cl_sw(row, ch & 7));
            }
            __syncwarp();
            fence_proxy_async();   // the staging writes precede the next item's TMA into the same tile
        }
    }
}

template <typename T, int C>
static int launch_cl_bwd_dkv_t(const CUtensorMap &tm_kv, const ClBwdKParams &p, cudaStream_t st) {
    const int smem = ClGeomK<C>::TOTAL;
    auto kern = fused_cl_bwd_dkv_kernel<T, C>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("fused_cl_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return AFFB200_E_CUDA; }
    const int64_t items = int64_t(p.B) * ((p.N + 7) / 8);
    const int64_t ctas = (items + 3) / 4;
    const int grid = int(ctas < 3 * kNumSMs ? ctas : 3 * kNumSMs);
    kern<<<grid, kKvThreads, smem, st>>>(tm_kv, p);
    return check_launch("fused_cl_bwd_dkv");
}

int launch_cl_bwd_dkv(const affb200_view *q, const affb200_view *k, const affb200_view *dout, const affb200_view *dk,
                      const float *table, const affb200_cl_plan *plan, const float *lse, const float *delta, int64_t B,
                      int64_t H, int64_t N, int64_t c, int64_t rel_pos_width, int64_t table_width, float scale, int dtype,
                      cudaStream_t st) {
    ClBwdKParams p;
    p.q = ClView{q->p, q->sB, q->sN, q->sH};
    p.dout = ClView{dout->p, dout->sB, dout->sN, dout->sH};
    p.dkv = ClView{dk->p, dk->sB, dk->sN, dk->sH};
    p.table = table; p.kb_cnt = plan->kb_cnt; p.kb_tok = plan->kb_tok; p.kbcap = plan->kbcap; p.pos16 = plan->pos16;
    p.lse = lse; p.delta = delta;
    p.B = int(B); p.H = int(H); p.N = int(N); p.w = int(rel_pos_width); p.tw = int(table_width);
    p.npad = ((N + kClT - 1) / kClT) * kClT;
    p.scale = scale; p.scale_log2 = scale * kLog2e;
    CUtensorMap tm_kv;
    int rc = cl_make_tensor_map(&tm_kv, k->p, dtype, 2 * H * c, N, B, k->sN, k->sB, 64, 8, false);
    if (rc) return rc;
    if (dtype == AFFB200_BF16) return c == 32 ? launch_cl_bwd_dkv_t<__nv_bfloat16, 32>(tm_kv, p, st) : launch_cl_bwd_dkv_t<__nv_bfloat16, 16>(tm_kv, p, st);
    return c == 32 ? launch_cl_bwd_dkv_t<__half, 32>(tm_kv, p, st) : launch_cl_bwd_dkv_t<__half, 16>(tm_kv, p, st);
}

}  // namespace affb200

using namespace affb200;

extern "C" int64_t affb200_fused_cl_bwd_grid(int64_t B, int64_t H, int64_t N) {
    const int64_t items = (H / 2) * B * ((N + kClT - 1) / kClT);
    return items < 2 * kNumSMs ? items : 2 * kNumSMs;
}

extern "C" int affb200_fused_cl_bwd(const affb200_view *q, const affb200_view *k, const affb200_view *v,
                                    const affb200_view *out, const affb200_view *dout, const affb200_view *dq,
                                    const affb200_view *dk, const affb200_view *dv, const float *table,
                                    const float *blank_k, const float *blank_v, const affb200_cl_plan *plan,
                                    const float *lse, float *delta, void *ds_work, float *blank_part, int32_t *guard,
                                    int64_t B, int64_t H, int64_t N, int64_t M, int64_t c, int64_t rel_pos_width,
                                    int64_t table_width, float scale, int dtype, void *stream) {
    if (!q || !k || !v || !out || !dout || !dq || !dk || !dv || !plan || !table || !blank_k || !blank_v || !lse || !delta ||
        !ds_work || !blank_part || !guard) {
        set_error("fused_cl_bwd: null pointer");
        return AFFB200_E_ARG;
    }
    if (B <= 0 || N <= 0 || H <= 0 || H % 2 != 0 || H * c > 512 || (c != 16 && c != 32) || M % 8 != 0 ||
        (dtype != AFFB200_F16 && dtype != AFFB200_BF16)) {
        set_error("fused_cl_bwd: needs an even head count, head dim 16 or 32, H*c <= 512 and f16/bf16");
        return AFFB200_E_ARG;
    }
    const int64_t es = 2;
    auto kv_ok = [&](const affb200_view *a, const affb200_view *b2) {
        return reinterpret_cast<uintptr_t>(b2->p) == reinterpret_cast<uintptr_t>(a->p) + uintptr_t(c * es) && a->sH == 2 * c &&
               b2->sH == 2 * c && a->sN == b2->sN && a->sB == b2->sB && aligned16(a->p) && (a->sN * es) % 16 == 0 && (a->sB * es) % 16 == 0;
    };
    auto row_ok = [&](const affb200_view *a) {
        return a->sH == c && aligned16(a->p) && (a->sN * es) % 16 == 0 && (a->sB * es) % 16 == 0;
    };
    if (!kv_ok(k, v) || !kv_ok(dk, dv) || !row_ok(q) || !row_ok(out) || !row_ok(dout) || !row_ok(dq)) {
        set_error("fused_cl_bwd: k/v (dk/dv) must be the [H][2][c] halves of one row, q/out/dout/dq rows [H][c], 16-byte aligned");
        return AFFB200_E_ARG;
    }
    if (plan->ucap < 1 || plan->ucap > AFFB200_CL_UMAX || !plan->q_hdr || !plan->q_ucl || !plan->q_mask || !plan->pos16 ||
        !plan->kb_cnt || !plan->kb_tok || plan->kbcap < 1) {
        set_error("fused_cl_bwd: incomplete plan");
        return AFFB200_E_ARG;
    }
    auto st = (cudaStream_t)stream;
    ClBwdQParams p;
    p.o = ClView{out->p, out->sB, out->sN, out->sH};
    p.dq = ClView{dq->p, dq->sB, dq->sN, dq->sH};
    p.table = table; p.blank_k = blank_k; p.blank_v = blank_v;
    p.hdr = plan->q_hdr; p.ucl = plan->q_ucl; p.qmask = reinterpret_cast<const unsigned long long *>(plan->q_mask);
    p.pos16 = plan->pos16; p.lse = lse; p.delta = delta; p.ds = ds_work; p.blank_part = blank_part; p.guard = guard;
    p.B = int(B); p.H = int(H); p.N = int(N); p.M = int(M); p.n_tiles = int((N + kClT - 1) / kClT); p.ucap = plan->ucap;
    p.w = int(rel_pos_width); p.tw = int(table_width);
    p.scale = scale; p.scale_log2 = scale * kLog2e;
    CUtensorMap tm_kv, tm_q, tm_do;
    int rc = cl_make_tensor_map(&tm_kv, k->p, dtype, 2 * H * c, N, B, k->sN, k->sB, 64, 8, false);
    if (rc) return rc;
    rc = cl_make_tensor_map(&tm_q, q->p, dtype, H * c, N, B, q->sN, q->sB, int(2 * c), kClT, c == 16);
    if (rc) return rc;
    rc = cl_make_tensor_map(&tm_do, dout->p, dtype, H * c, N, B, dout->sN, dout->sB, int(2 * c), kClT, c == 16);
    if (rc) return rc;
    const int grid = int(affb200_fused_cl_bwd_grid(B, H, N));
    if (dtype == AFFB200_BF16)
        rc = c == 32 ? launch_cl_bwd_dq<__nv_bfloat16, 32>(tm_kv, tm_q, tm_do, p, grid, st) : launch_cl_bwd_dq<__nv_bfloat16, 16>(tm_kv, tm_q, tm_do, p, grid, st);
    else
        rc = c == 32 ? launch_cl_bwd_dq<__half, 32>(tm_kv, tm_q, tm_do, p, grid, st) : launch_cl_bwd_dq<__half, 16>(tm_kv, tm_q, tm_do, p, grid, st);
    if (rc) return rc;
    return affb200::launch_cl_bwd_dkv(q, k, dout, dk, table, plan, lse, delta, B, H, N, c, rel_pos_width, table_width, scale,
                                      dtype, st);
}
