// fused_cl_fwd.cu — forward of the block-structured fused neighbourhood attention (see fused_cl.cuh for the idea;
// same operator as fused_attn.cu: reference models/aff_transformer.py:103-156).
//
// A work item is (head group of 2 heads, image, tile of 64 tokens).  Persistent CTAs of 8 warps (two per SM) walk the
// items; per item
//   1. the tile's union blocks (<= 64 x 8 rows), its own 64 Q rows and the packed positions of the union rows are staged
//      with bulk-async copies — one per row, K and V of both heads at once (256 B for head dim 32) — all completing on
//      one mbarrier; rows are padded by 16 B so that ldmatrix is conflict-free;
//   2. each warp owns 8 tokens and both heads (rows 0-7 of its m16 tiles = head 0, rows 8-15 = head 1; the A halves of
//      the other head are zero registers): it ORs the tokens' membership words and visits only those union blocks, two
//      at a time: S = Q·K^T, x = S*scale*log2e + bias (shared-memory window of the table indexed by pos_u - pos_i, one
//      lookup for both heads; non-members read a -inf entry), p = 2^(x - m) with a lazily updated running maximum,
//      then the two blocks' probabilities are the A fragment of O += P·V (V through ldmatrix.trans) and of a ones
//      column that accumulates the row sums — nothing goes back to shared memory;
//   3. the blank token (models/aff_transformer.py:139-150) is a pseudo-block behind the union: its logit initialises the
//      running maximum, its value row initialises O;
//   4. O / l is staged in the warp's (dead) Q rows and leaves as 16-byte coalesced stores; lse = m + ln(l) is kept for
//      the backward.
#include "fused_cl.cuh"

namespace affb200 {

struct ClFwdParams {
    ClView out;
    const float *table, *blank_k, *blank_v;
    const int32_t *hdr;
    const uint16_t *ucl;
    const unsigned long long *qmask;
    const uint32_t *pos16;
    const int32_t *hdr2;   // second-half union size of split tiles (AFFB200_CL_SPLIT), or nullptr
    const uint16_t *ucl2;  // second-half union list
    float *lse;
    int32_t *guard;        // incremented when a tile had to be skipped (speculated plans)
    int B, H, N, n_tiles, ucap, w, tw;
    float scale_log2;
};

template <typename T, int C>
__global__ void __launch_bounds__(kClThreads, 2)
fused_cl_fwd_kernel(const __grid_constant__ CUtensorMap tm_kv, const __grid_constant__ CUtensorMap tm_q,
                    const ClFwdParams p) {
    using G = ClGeom<C>;
    constexpr int NT = C / 8;     // output n-tiles per head
    constexpr int KS = C / 16;    // k-steps of the S contraction per head
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, q4 = lane & 3, r8 = lane & 7;
    const uint32_t sm = dsm_addr();
    const uint32_t bar = sm, s_act = sm + G::OFF_ACT + warp * 64, s_win = sm + G::OFF_WIN, s_pu = sm + G::OFF_PU;
    const uint32_t s_q = sm + G::OFF_Q, s_kv = sm + G::OFF_KV;
    unsigned char *gen = affb200_dsm;
    const int ucap = p.ucap;
    const uint32_t items = uint32_t(p.H / 2) * p.B * p.n_tiles;
    T *outp = static_cast<T *>(const_cast<void *>(p.out.p));
    const int64_t npad = int64_t(p.n_tiles) * kClT;
    const uint32_t onesb = g == 0 ? cl_ones2<T>() : 0u;   // B fragment of a column of ones (column 0)

    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    // rows 1..7 of the blank pseudo-block stay zero for the whole kernel
    for (int e = tid; e < G::BLK / 16; e += kClThreads)
        *reinterpret_cast<uint4 *>(gen + G::OFF_KV + ucap * G::BLK + e * 16) = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();

    uint32_t phase = 0;
    int cur_hg = -1;
    uint32_t item = blockIdx.x;
    int hdr_n = 0;
    uint32_t ucl_n = 0;
    if (item < items) {
        const uint32_t bt = item % (uint32_t(p.B) * p.n_tiles);   // (image, tile) index: hdr / ucl are shared by the head groups
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
        if (next < items) {   // next item's header and union list travel while this item is computed
            const uint32_t bt = next % (uint32_t(p.B) * p.n_tiles);
            hdr_n = __ldg(p.hdr + bt);
            if (tid < kClUMax) ucl_n = __ldg(p.ucl + int64_t(bt) * kClUMax + tid);
        }
        item = next;
        const int U = hdr & AFFB200_PLAN_COUNT_MASK;
        // a split tile (its 64 tokens reference more than 64 blocks) carries one union per half of 32 tokens: two
        // passes over the tile, warps 0-3 compute in the first, warps 4-7 in the second
        const bool split = (hdr & AFFB200_CL_SPLIT) != 0;
        const int64_t bt_cur = int64_t(b) * p.n_tiles + tile;
        const int U2 = split && p.hdr2 ? (__ldg(p.hdr2 + bt_cur) & AFFB200_PLAN_COUNT_MASK) : 0;
        const bool skip = (hdr & (AFFB200_PLAN_OVERFLOW | AFFB200_CL_BAD)) != 0 || U > ucap || U2 > ucap || (split && !p.hdr2);
        T *out_tile = outp + int64_t(b) * p.out.sB + int64_t(t0) * p.out.sN + h0 * p.out.sH;
        if (skip) {   // speculated plan that did not hold: defined output, guard tripped
            for (int e = tid; e < ti * (G::QB / 16); e += kClThreads) {
                const int row = e / (G::QB / 16), ch = e % (G::QB / 16);
                *reinterpret_cast<uint4 *>(out_tile + int64_t(row) * p.out.sN + ch * 8) = make_uint4(0u, 0u, 0u, 0u);
            }
            for (int e = tid; e < ti * 2; e += kClThreads)
                p.lse[(int64_t(b) * p.H + h0 + e / ti) * p.N + t0 + e % ti] = 0.f;
            if (tid == 0 && hg == 0) atomicAdd(p.guard, 1);
            continue;
        }
        if (hg != cur_hg) {   // head group changed: table window and blank row of this group (all warps are idle here)
            cur_hg = hg;
            cl_fill_window<2>(reinterpret_cast<float *>(gen + G::OFF_WIN), p.table, p.H, h0, p.w, p.tw);
            for (int e = tid; e < 4 * C; e += kClThreads) {
                const int hh = e / (2 * C), rr = e % (2 * C);
                const float v = rr < C ? __ldg(p.blank_k + (h0 + hh) * C + rr) : __ldg(p.blank_v + (h0 + hh) * C + rr - C);
                // row 0 of a swizzled tile is stored linearly; head dim 32 has one tile per head
                reinterpret_cast<T *>(gen + G::OFF_KV + ucap * G::BLK + (C == 32 ? hh * 1024 : hh * 4 * C))[rr] = Elem<T>::from_acc(v);
            }
            __syncthreads();
        }
        for (int pass = 0; pass < (split ? 2 : 1); ++pass) {
        int Up = U, blk_p = blk_mine;
        if (pass == 1) {
            Up = U2;
            blk_p = tid < kClUMax ? int(__ldg(p.ucl2 + bt_cur * kClUMax + tid)) : 0;
        }
        const bool wact = !split || (warp >> 2) == pass;
        // ---- stage: TMA boxes of 8 rows x 128 bytes, the Q tile, and the tile's plan rows, on one mbarrier ----------
        if (tid == 0) mbar_expect_tx(bar, uint32_t(Up) * (G::BLK + 32u) + uint32_t(kClT) * G::QB + 512u + 256u);
        if (tid < Up) {
#pragma unroll
            for (int t = 0; t < G::TPB; ++t)
                tma_load_3d(s_kv + (tid * G::TPB + t) * 1024, &tm_kv, h0 * 2 * C + t * 64, blk_p * 8, b, bar);
            bulk_g2s(s_pu + tid * 32, p.pos16 + int64_t(b) * npad + blk_p * 8, 32u, bar);
        } else if (tid == 64) {
            tma_load_3d(s_q, &tm_q, h0 * C, t0, b, bar);
        } else if (tid == 65) {
            bulk_g2s(sm + G::OFF_MK, p.qmask + int64_t(b) * npad + t0, 512u, bar);
        } else if (tid == 66) {
            bulk_g2s(sm + G::OFF_PI, p.pos16 + int64_t(b) * npad + t0, 256u, bar);
        }
        mbar_wait(bar, phase);
        phase ^= 1u;

        if (wact) {
        // ---- per-token state and the warp's list of referenced blocks -------------------------------------------
        const int tok = warp * 8 + g;
        const bool live = tok < ti;
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
        // ---- Q fragments: registers 0, 2 = head 0 (rows 0-7), 1, 3 = head 1 (rows 8-15); blank token --------------
        uint32_t qa[KS][4];
        cl_load_afrag<C>(qa, s_q, warp, lane);
        float oh0[NT][4], oh1[NT][4], ol[4], m[2];
        {
            float s0[4], s1[4];
            cl_scores<T, C>(s0, s1, qa, s_kv + ucap * G::BLK, lane);
            m[0] = __shfl_sync(0xffffffffu, s0[0], lane & ~3) * p.scale_log2;
            m[1] = __shfl_sync(0xffffffffu, s1[2], lane & ~3) * p.scale_log2;
            ol[0] = ol[2] = q4 == 0 ? 1.f : 0.f;
            ol[1] = ol[3] = 0.f;
            // blank_v as staged (rounded to T): row 0 of the blank block, V chunks
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const uint32_t a0 = C * 2 + nt * 16 + q4 * 4;
                const uint32_t v0 = lds_u32(s_kv + ucap * G::BLK + a0);
                const uint32_t v1 = lds_u32(s_kv + ucap * G::BLK + (C == 32 ? 1024 : 4 * C) + a0);
                const T *e0 = reinterpret_cast<const T *>(&v0), *e1 = reinterpret_cast<const T *>(&v1);
                oh0[nt][0] = Elem<T>::to_acc(e0[0]); oh0[nt][1] = Elem<T>::to_acc(e0[1]); oh0[nt][2] = oh0[nt][3] = 0.f;
                oh1[nt][2] = Elem<T>::to_acc(e1[0]); oh1[nt][3] = Elem<T>::to_acc(e1[1]); oh1[nt][0] = oh1[nt][1] = 0.f;
            }
        }
        __syncwarp();   // every lane has read its Q rows (they may now hold the output) and the list is written

        // ---- union blocks referenced by this warp's 8 tokens, two per step --------------------------------------
        for (int k = 0; k < n_act; k += 2) {
            const uint32_t two = lds_u16(s_act + k);
            const bool has1 = k + 1 < n_act;
            int pj[2];
            pj[0] = int(two & 0xffu);
            pj[1] = has1 ? int(two >> 8) : pj[0];
            bool far[2][2], anyfar = false;
            float2 bias[2][2];   // [block][column] -> (head 0, head 1)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const uint2 pu = lds_u64(s_pu + (pj[j] * 8 + q4 * 2) * 4);
                const bool mem = ((mk >> pj[j]) & 1ull) != 0ull && (j == 0 || has1);
                const uint32_t i0 = cl_win_index(pu.x, pib, mem, far[j][0]), i1 = cl_win_index(pu.y, pib, mem, far[j][1]);
                bias[j][0] = lds_f2(s_win + i0 * 8u);
                bias[j][1] = lds_f2(s_win + i1 * 8u);
                anyfar = anyfar || far[j][0] || far[j][1];
            }
            if (__any_sync(0xffffffffu, anyfar)) {   // rare: an offset outside the window, or a padded row
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
            float s0[2][4], s1[2][4];
#pragma unroll
            for (int j = 0; j < 2; ++j) cl_scores<T, C>(s0[j], s1[j], qa, s_kv + pj[j] * G::BLK, lane);
            // logits in the log2 domain: head 0 in s0[j][0..1], head 1 in s1[j][2..3]
            float lm[2] = {-INFINITY, -INFINITY};
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    s0[j][cc] = fmaf(s0[j][cc], p.scale_log2, bias[j][cc].x);
                    s1[j][2 + cc] = fmaf(s1[j][2 + cc], p.scale_log2, bias[j][cc].y);
                    lm[0] = fmaxf(lm[0], s0[j][cc]);
                    lm[1] = fmaxf(lm[1], s1[j][2 + cc]);
                }
            const bool need = lm[0] > m[0] + 8.f || lm[1] > m[1] + 8.f;
            if (__any_sync(0xffffffffu, need)) {
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    float v = lm[hh];
                    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
                    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
                    const float mn = fmaxf(m[hh], v);
                    const float f = f_ex2(m[hh] - mn);
                    m[hh] = mn;
                    ol[2 * hh] *= f; ol[2 * hh + 1] *= f;
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) {
                        if (hh == 0) { oh0[nt][0] *= f; oh0[nt][1] *= f; }
                        else { oh1[nt][2] *= f; oh1[nt][3] *= f; }
                    }
                }
            }
            uint32_t pa[4];
            pa[0] = f_pack2<T>(f_ex2(s0[0][0] - m[0]), f_ex2(s0[0][1] - m[0]));
            pa[1] = f_pack2<T>(f_ex2(s1[0][2] - m[1]), f_ex2(s1[0][3] - m[1]));
            pa[2] = f_pack2<T>(f_ex2(s0[1][0] - m[0]), f_ex2(s0[1][1] - m[0]));
            pa[3] = f_pack2<T>(f_ex2(s1[1][2] - m[1]), f_ex2(s1[1][3] - m[1]));
            f_mma<T>(ol, pa, onesb, onesb);   // row sums of the rounded probabilities (column 0)
            cl_accum_rows<T, C, true>(oh0, oh1, pa, s_kv + pj[(lane >> 3) & 1] * G::BLK, lane);
        }

        // ---- epilogue: O / l through the warp's Q rows (same swizzle), 16-byte stores ----------------------------
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const float l = __shfl_sync(0xffffffffu, ol[2 * hh], lane & ~3);
            const float inv = __fdividef(1.f, l);
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const int j = hh * (C / 8) + nt;
                const uint32_t a = cl_tile_off<C>(tok, j);
                const float x0 = hh == 0 ? oh0[nt][0] : oh1[nt][2], x1 = hh == 0 ? oh0[nt][1] : oh1[nt][3];
                *reinterpret_cast<uint32_t *>(gen + G::OFF_Q + a + q4 * 4) = f_pack2<T>(x0 * inv, x1 * inv);
            }
            if (q4 == 0 && live)
                p.lse[(int64_t(b) * p.H + h0 + hh) * p.N + t0 + tok] = (m[hh] + __log2f(l)) * 0.6931471805599453f;
        }
        __syncwarp();
#pragma unroll
        for (int e = lane; e < 8 * (G::QB / 16); e += 32) {
            const int row = warp * 8 + e / (G::QB / 16), ch = e % (G::QB / 16);
            const uint32_t a = cl_tile_off<C>(row, ch);
            if (row < ti)
                *reinterpret_cast<uint4 *>(out_tile + int64_t(row) * p.out.sN + ch * 8) = *reinterpret_cast<const uint4 *>(gen + G::OFF_Q + a);
        }
        }
        fence_proxy_async();   // the staging writes above precede the next item's async-proxy writes to the same rows
        __syncthreads();       // all warps are done with this item's rows before the next copies land
        }
    }
}

template <typename T, int C>
static int launch_cl_fwd(const CUtensorMap &tm_kv, const CUtensorMap &tm_q, const ClFwdParams &p, cudaStream_t st) {
    using G = ClGeom<C>;
    const int smem = G::total(p.ucap);
    if (smem > 227 * 1024) { set_error("fused_cl_fwd: union of %d blocks does not fit shared memory", p.ucap); return AFFB200_E_ARG; }
    auto kern = fused_cl_fwd_kernel<T, C>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("fused_cl_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return AFFB200_E_CUDA; }
    const int64_t items = int64_t(p.H / 2) * p.B * p.n_tiles;
    if (items >= (int64_t(1) << 31)) { set_error("fused_cl_fwd: too many tiles"); return AFFB200_E_ARG; }
    const int grid = int(items < 2 * kNumSMs ? items : 2 * kNumSMs);
    kern<<<grid, kClThreads, smem, st>>>(tm_kv, tm_q, p);
    return check_launch("fused_cl_fwd");
}

}  // namespace affb200

using namespace affb200;

extern "C" int affb200_fused_cl_fwd(const affb200_view *q, const affb200_view *k, const affb200_view *v,
                                    const affb200_view *out, const float *table, const float *blank_k,
                                    const float *blank_v, const affb200_cl_plan *plan, float *lse, int32_t *guard,
                                    int64_t B, int64_t H, int64_t N, int64_t c, int64_t rel_pos_width,
                                    int64_t table_width, float scale, int dtype, void *stream) {
    if (!q || !k || !v || !out || !plan || !table || !blank_k || !blank_v || !lse || !guard) {
        set_error("fused_cl_fwd: null pointer");
        return AFFB200_E_ARG;
    }
    if (B <= 0 || N <= 0 || H <= 0 || H % 2 != 0 || (c != 16 && c != 32) || (dtype != AFFB200_F16 && dtype != AFFB200_BF16)) {
        set_error("fused_cl_fwd: needs an even head count, head dim 16 or 32 and f16/bf16 (H=%lld c=%lld dtype=%d)",
                  (long long)H, (long long)c, dtype);
        return AFFB200_E_ARG;
    }
    const int64_t es = 2;
    if (reinterpret_cast<uintptr_t>(v->p) != reinterpret_cast<uintptr_t>(k->p) + uintptr_t(c * es) || k->sH != 2 * c ||
        v->sH != 2 * c || k->sN != v->sN || k->sB != v->sB || q->sH != c || out->sH != c) {
        set_error("fused_cl_fwd: k/v must be the [H][2][c] halves of one kv row and q/out rows [H][c]");
        return AFFB200_E_ARG;
    }
    if (!aligned16(q->p) || !aligned16(k->p) || !aligned16(out->p) || (q->sN * es) % 16 || (k->sN * es) % 16 ||
        (out->sN * es) % 16 || (q->sB * es) % 16 || (k->sB * es) % 16 || (out->sB * es) % 16) {
        set_error("fused_cl_fwd: rows must be 16-byte aligned");
        return AFFB200_E_ARG;
    }
    if (plan->ucap < 1 || plan->ucap > AFFB200_CL_UMAX || !plan->q_hdr || !plan->q_ucl || !plan->q_mask || !plan->pos16) {
        set_error("fused_cl_fwd: incomplete plan");
        return AFFB200_E_ARG;
    }
    ClFwdParams p;
    p.out = ClView{out->p, out->sB, out->sN, out->sH};
    p.table = table; p.blank_k = blank_k; p.blank_v = blank_v;
    p.hdr = plan->q_hdr; p.ucl = plan->q_ucl; p.qmask = reinterpret_cast<const unsigned long long *>(plan->q_mask);
    p.pos16 = plan->pos16; p.lse = lse; p.guard = guard;
    p.hdr2 = plan->q_hdr2; p.ucl2 = plan->q_hdr2 ? plan->q_ucl2 : nullptr;
    if (plan->q_hdr2 && !plan->q_ucl2) { set_error("fused_cl_fwd: q_hdr2 without q_ucl2"); return AFFB200_E_ARG; }
    p.B = int(B); p.H = int(H); p.N = int(N); p.n_tiles = int((N + kClT - 1) / kClT); p.ucap = plan->ucap;
    p.w = int(rel_pos_width); p.tw = int(table_width);
    p.scale_log2 = scale * kLog2e;
    CUtensorMap tm_kv, tm_q;
    int rc = cl_make_tensor_map(&tm_kv, k->p, dtype, 2 * H * c, N, B, k->sN, k->sB, 64, 8, false);
    if (rc) return rc;
    rc = cl_make_tensor_map(&tm_q, q->p, dtype, H * c, N, B, q->sN, q->sB, int(2 * c), kClT, c == 16);
    if (rc) return rc;
    auto st = (cudaStream_t)stream;
    if (dtype == AFFB200_BF16)
        return c == 32 ? launch_cl_fwd<__nv_bfloat16, 32>(tm_kv, tm_q, p, st) : launch_cl_fwd<__nv_bfloat16, 16>(tm_kv, tm_q, p, st);
    return c == 32 ? launch_cl_fwd<__half, 32>(tm_kv, tm_q, p, st) : launch_cl_fwd<__half, 16>(tm_kv, tm_q, p, st);
}
