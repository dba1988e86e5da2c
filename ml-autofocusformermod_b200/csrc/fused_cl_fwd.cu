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

template <int C, int HG> struct ClGeom {
    static_assert(HG == 2, "rows 0-7 / 8-15 of the m16 tiles are the two heads of a group");
    static constexpr int KVB = HG * 2 * C * 2;   // bytes of one row's K|V for the head group
    static constexpr int KVS = KVB + 16;         // padded row stride
    static constexpr int QB = HG * C * 2;
    static constexpr int QS = QB + 16;
    static constexpr int HB = 2 * C * 2;         // bytes between heads inside a K|V row
    static constexpr int OFF_UCL = 16;
    static constexpr int OFF_ACT = 144;          // per-warp list of referenced union blocks (8 x 64 bytes)
    static constexpr int OFF_WIN = 768;
    static constexpr int OFF_POS = OFF_WIN + ((kClWin * kClWin + 1) * HG * 4 + 127) / 128 * 128;
    __host__ __device__ static constexpr int off_q(int ucap) { return OFF_POS + (ucap + 1) * 32; }
    __host__ __device__ static constexpr int off_kv(int ucap) { return off_q(ucap) + kClT * QS; }
    __host__ __device__ static constexpr int total(int ucap) { return off_kv(ucap) + (ucap + 1) * 8 * KVS; }
};

struct ClView {
    const void *p;
    int64_t sB, sN, sH;    // element strides
};

struct ClFwdParams {
    ClView q, kv, out;     // kv: K of head h at p + b*sB + i*sN + h*sH, V follows at +C
    const float *table, *blank_k, *blank_v;
    const int32_t *hdr;
    const uint16_t *ucl;
    const unsigned long long *qmask;
    const uint32_t *pos16;
    float *lse;
    int32_t *guard;        // incremented when a tile had to be skipped (speculated plans)
    int B, H, N, n_tiles, ucap, w, tw;
    float scale_log2;
};

template <typename T> __device__ __forceinline__ uint32_t cl_ones2();
template <> __device__ __forceinline__ uint32_t cl_ones2<__nv_bfloat16>() { return 0x3f803f80u; }
template <> __device__ __forceinline__ uint32_t cl_ones2<__half>() { return 0x3c003c00u; }

// d += A_lo·B with only rows 0-7 of A (registers a0, a2) / only rows 8-15 (a1, a3); the other half is zero
template <typename T> __device__ __forceinline__ void cl_mma_lo(float (&d)[4], uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
    const uint32_t a[4] = {a0, 0u, a2, 0u};
    f_mma<T>(d, a, b0, b1);
}
template <typename T> __device__ __forceinline__ void cl_mma_hi(float (&d)[4], uint32_t a1, uint32_t a3, uint32_t b0, uint32_t b1) {
    const uint32_t a[4] = {0u, a1, 0u, a3};
    f_mma<T>(d, a, b0, b1);
}

// S (16 x 8: rows 0-7 = head 0, rows 8-15 = head 1 of the warp's 8 tokens) for one staged block at `kb`
template <typename T, int C, int KVS, int HB>
__device__ __forceinline__ void cl_scores(float (&s)[4], const uint32_t (&qa)[C / 16][4], uint32_t kb, int lane) {
    s[0] = s[1] = s[2] = s[3] = 0.f;
    if constexpr (C == 32) {
        uint32_t kf[4];
        cl_ldsm_x4(kf, kb + (lane & 7) * KVS + (lane >> 3) * 16);
        cl_mma_lo<T>(s, qa[0][0], qa[0][2], kf[0], kf[1]);
        cl_mma_lo<T>(s, qa[1][0], qa[1][2], kf[2], kf[3]);
        cl_ldsm_x4(kf, kb + (lane & 7) * KVS + HB + (lane >> 3) * 16);
        cl_mma_hi<T>(s, qa[0][1], qa[0][3], kf[0], kf[1]);
        cl_mma_hi<T>(s, qa[1][1], qa[1][3], kf[2], kf[3]);
    } else {
        uint32_t kf[4];   // (head 0: chunks 0,1), (head 1: chunks 0,1)
        cl_ldsm_x4(kf, kb + (lane & 7) * KVS + (lane >> 4) * HB + ((lane >> 3) & 1) * 16);
        cl_mma_lo<T>(s, qa[0][0], qa[0][2], kf[0], kf[1]);
        cl_mma_hi<T>(s, qa[0][1], qa[0][3], kf[2], kf[3]);
    }
}

template <typename T, int C, int HG>
__global__ void __launch_bounds__(kClThreads, 2)
fused_cl_fwd_kernel(const ClFwdParams p) {
    using G = ClGeom<C, HG>;
    constexpr int NT = C / 8;     // output n-tiles per head
    constexpr int KS = C / 16;    // k-steps of the S contraction per head
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, q4 = lane & 3;
    const uint32_t sm = dsm_addr();
    const uint32_t bar = sm, s_act = sm + G::OFF_ACT + warp * 64, s_win = sm + G::OFF_WIN, s_pos = sm + G::OFF_POS;
    const uint32_t s_q = sm + G::off_q(p.ucap), s_kv = sm + G::off_kv(p.ucap);
    unsigned char *gen = affb200_dsm;
    const int ucap = p.ucap;
    const uint32_t items = uint32_t(p.H / HG) * p.B * p.n_tiles;
    const T *qp = static_cast<const T *>(p.q.p), *kvp = static_cast<const T *>(p.kv.p);
    T *outp = static_cast<T *>(const_cast<void *>(p.out.p));
    const int nblk = (p.N + 7) >> 3;
    const uint32_t onesb = g == 0 ? cl_ones2<T>() : 0u;   // B fragment of a column of ones (column 0)

    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    // rows 1..7 of the blank pseudo-block stay zero for the whole kernel
    for (int e = tid; e < 8 * G::KVS / 16; e += kClThreads)
        *reinterpret_cast<uint4 *>(gen + G::off_kv(ucap) + ucap * 8 * G::KVS + e * 16) = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();

    uint32_t phase = 0;
    int cur_hg = -1;

    for (uint32_t item = blockIdx.x; item < items; item += gridDim.x) {
        const int tile = int(item % uint32_t(p.n_tiles));
        const uint32_t ib = item / uint32_t(p.n_tiles);
        const int b = int(ib % uint32_t(p.B)), hg = int(ib / uint32_t(p.B));
        const int h0 = hg * HG;
        const int t0 = tile * kClT, ti = min(kClT, p.N - t0);
        const int hdr = p.hdr[b * p.n_tiles + tile];
        const int U = hdr & AFFB200_PLAN_COUNT_MASK;
        const bool skip = (hdr & (AFFB200_PLAN_OVERFLOW | AFFB200_CL_BAD)) != 0 || U > ucap;
        T *out_tile = outp + int64_t(b) * p.out.sB + int64_t(t0) * p.out.sN + h0 * p.out.sH;
        if (skip) {   // speculated plan that did not hold: defined output, guard tripped
            for (int e = tid; e < ti * (G::QB / 16); e += kClThreads) {
                const int row = e / (G::QB / 16), ch = e % (G::QB / 16);
                *reinterpret_cast<uint4 *>(out_tile + int64_t(row) * p.out.sN + ch * 8) = make_uint4(0u, 0u, 0u, 0u);
            }
            for (int e = tid; e < ti * HG; e += kClThreads)
                p.lse[(int64_t(b) * p.H + h0 + e / ti) * p.N + t0 + e % ti] = 0.f;
            if (tid == 0 && hg == 0) atomicAdd(p.guard, 1);
            continue;
        }
        if (hg != cur_hg) {   // head group changed: table window and blank row of this group
            cur_hg = hg;
            cl_fill_window<HG>(reinterpret_cast<float *>(gen + G::OFF_WIN), p.table, p.H, h0, p.w, p.tw);
            for (int e = tid; e < HG * 2 * C; e += kClThreads) {
                const int hh = e / (2 * C), r = e % (2 * C);
                const float v = r < C ? __ldg(p.blank_k + (h0 + hh) * C + r) : __ldg(p.blank_v + (h0 + hh) * C + r - C);
                reinterpret_cast<T *>(gen + G::off_kv(ucap) + ucap * 8 * G::KVS)[hh * 2 * C + r] = Elem<T>::from_acc(v);
            }
        }
        const uint16_t *ucl_g = p.ucl + (int64_t(b) * p.n_tiles + tile) * kClUMax;
        if (tid < U) reinterpret_cast<uint16_t *>(gen + G::OFF_UCL)[tid] = ucl_g[tid];
        {
            // rows of the union that exist: only the image's last block can be partial, and the union is ascending.
            // Its missing rows are zeroed (they are masked, but 0 x garbage must stay 0 in P·V).
            const int last = ucl_g[U - 1];
            const int missing = last == nblk - 1 ? nblk * 8 - p.N : 0;
            for (int e = tid; e < missing * (G::KVS / 16); e += kClThreads)
                *reinterpret_cast<uint4 *>(gen + G::off_kv(ucap) + (U * 8 - missing) * G::KVS + e * 16) = make_uint4(0u, 0u, 0u, 0u);
            if (tid == 0)
                mbar_expect_tx(bar, uint32_t(U * 8 - missing) * G::KVB + uint32_t(ti) * G::QB + uint32_t(U) * 32u);
        }
        fence_proxy_async();
        __syncthreads();
        // ---- stage: one bulk copy per row ------------------------------------------------------------------
        {
            const T *kv_img = kvp + int64_t(b) * p.kv.sB + h0 * p.kv.sH;
            for (int e = tid; e < U * 8; e += kClThreads) {
                const int blk = reinterpret_cast<const uint16_t *>(gen + G::OFF_UCL)[e >> 3];
                const int row = blk * 8 + (e & 7);
                if (row < p.N) bulk_g2s(s_kv + e * G::KVS, kv_img + int64_t(row) * p.kv.sN, G::KVB, bar);
            }
            if (tid < ti)
                bulk_g2s(s_q + tid * G::QS, qp + int64_t(b) * p.q.sB + int64_t(t0 + tid) * p.q.sN + h0 * p.q.sH, G::QB, bar);
            if (tid >= 64 && tid - 64 < U) {
                const int blk = reinterpret_cast<const uint16_t *>(gen + G::OFF_UCL)[tid - 64];
                bulk_g2s(s_pos + (tid - 64) * 32, p.pos16 + int64_t(b) * nblk * 8 + blk * 8, 32u, bar);
            }
        }
        // ---- per-token state and the warp's list of referenced blocks, while the copies fly --------------------
        const int tok = warp * 8 + g;
        const bool live = tok < ti;
        const unsigned long long mk = live ? p.qmask[int64_t(b) * p.N + t0 + tok] : 0ull;
        const uint32_t pib = (live ? p.pos16[int64_t(b) * nblk * 8 + t0 + tok] : 0u) - kClPosBias;
        int n_act;
        {
            const uint32_t alo = __reduce_or_sync(0xffffffffu, uint32_t(mk)), ahi = __reduce_or_sync(0xffffffffu, uint32_t(mk >> 32));
            const uint32_t lt = (1u << lane) - 1u;
            const int clo = __popc(alo);
            if ((alo >> lane) & 1u) gen[G::OFF_ACT + warp * 64 + __popc(alo & lt)] = uint8_t(lane);
            if ((ahi >> lane) & 1u) gen[G::OFF_ACT + warp * 64 + clo + __popc(ahi & lt)] = uint8_t(32 + lane);
            n_act = clo + __popc(ahi);
        }
        mbar_wait(bar, phase);
        phase ^= 1u;

        // ---- Q fragments (a0/a2: head 0, a1/a3: head 1), blank token ------------------------------------------
        uint32_t qa[KS][4];
#pragma unroll
        for (int ks = 0; ks < KS; ++ks)
            cl_ldsm_x4(qa[ks], s_q + (warp * 8 + (lane & 7)) * G::QS + ((lane >> 3) & 1) * C * 2 + ks * 32 + (lane >> 4) * 16);
        float o[NT][4], ol[4], m[2];
        {
            float s[4];
            cl_scores<T, C, G::KVS, G::HB>(s, qa, s_kv + ucap * 8 * G::KVS, lane);
            m[0] = __shfl_sync(0xffffffffu, s[0], lane & ~3) * p.scale_log2;
            m[1] = __shfl_sync(0xffffffffu, s[2], lane & ~3) * p.scale_log2;
            ol[0] = ol[2] = q4 == 0 ? 1.f : 0.f;
            ol[1] = ol[3] = 0.f;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const float2 bv = __ldg(reinterpret_cast<const float2 *>(p.blank_v + (h0 + hh) * C + nt * 8 + q4 * 2));
                    // the staged blank_v is rounded to T like every other V row; use the same values
                    o[nt][2 * hh] = Elem<T>::to_acc(Elem<T>::from_acc(bv.x));
                    o[nt][2 * hh + 1] = Elem<T>::to_acc(Elem<T>::from_acc(bv.y));
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
                const uint2 pu = lds_u64(s_pos + (pj[j] * 8 + q4 * 2) * 4);
                const bool mem = ((mk >> pj[j]) & 1ull) != 0ull && (j == 0 || has1);
                const uint32_t i0 = cl_win_index(pu.x, pib, mem, far[j][0]), i1 = cl_win_index(pu.y, pib, mem, far[j][1]);
                bias[j][0] = lds_f2(s_win + i0 * 8u);
                bias[j][1] = lds_f2(s_win + i1 * 8u);
                anyfar = anyfar || far[j][0] || far[j][1];
            }
            if (__any_sync(0xffffffffu, anyfar)) {   // rare: an offset outside the window, or a padded row
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const uint2 pu = lds_u64(s_pos + (pj[j] * 8 + q4 * 2) * 4);
#pragma unroll
                    for (int cc = 0; cc < 2; ++cc)
                        if (far[j][cc]) {
                            bias[j][cc].x = cl_bias_far(p.table, p.H, h0, p.w, p.tw, cc ? pu.y : pu.x, pib);
                            bias[j][cc].y = cl_bias_far(p.table, p.H, h0 + 1, p.w, p.tw, cc ? pu.y : pu.x, pib);
                        }
                }
            }
            float s[2][4];
#pragma unroll
            for (int j = 0; j < 2; ++j) cl_scores<T, C, G::KVS, G::HB>(s[j], qa, s_kv + pj[j] * 8 * G::KVS, lane);
            // logits in the log2 domain: s[j][2*head + column]
            float lm[2] = {-INFINITY, -INFINITY};
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    s[j][cc] = fmaf(s[j][cc], p.scale_log2, bias[j][cc].x);
                    s[j][2 + cc] = fmaf(s[j][2 + cc], p.scale_log2, bias[j][cc].y);
                    lm[0] = fmaxf(lm[0], s[j][cc]);
                    lm[1] = fmaxf(lm[1], s[j][2 + cc]);
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
                    for (int nt = 0; nt < NT; ++nt) { o[nt][2 * hh] *= f; o[nt][2 * hh + 1] *= f; }
                }
            }
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    s[j][cc] = f_ex2(s[j][cc] - m[0]);
                    s[j][2 + cc] = f_ex2(s[j][2 + cc] - m[1]);
                }
            uint32_t pa[4];
            pa[0] = f_pack2<T>(s[0][0], s[0][1]); pa[1] = f_pack2<T>(s[0][2], s[0][3]);
            pa[2] = f_pack2<T>(s[1][0], s[1][1]); pa[3] = f_pack2<T>(s[1][2], s[1][3]);
            f_mma<T>(ol, pa, onesb, onesb);   // row sums of the rounded probabilities (column 0)
            const uint32_t vb = s_kv + (pj[(lane >> 3) & 1] * 8 + (lane & 7)) * G::KVS + C * 2 + (lane >> 4) * 16;
#pragma unroll
            for (int np = 0; np < C / 16; ++np) {
                uint32_t vf[4];
                cl_ldsm_x4_t(vf, vb + np * 32);
                cl_mma_lo<T>(o[2 * np], pa[0], pa[2], vf[0], vf[1]);
                cl_mma_lo<T>(o[2 * np + 1], pa[0], pa[2], vf[2], vf[3]);
                cl_ldsm_x4_t(vf, vb + G::HB + np * 32);
                cl_mma_hi<T>(o[2 * np], pa[1], pa[3], vf[0], vf[1]);
                cl_mma_hi<T>(o[2 * np + 1], pa[1], pa[3], vf[2], vf[3]);
            }
        }

        // ---- epilogue ------------------------------------------------------------------------------------------
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const float l = __shfl_sync(0xffffffffu, ol[2 * hh], lane & ~3);
            const float inv = __fdividef(1.f, l);
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
                *reinterpret_cast<uint32_t *>(gen + G::off_q(ucap) + tok * G::QS + (hh * C + nt * 8 + q4 * 2) * 2) =
                    f_pack2<T>(o[nt][2 * hh] * inv, o[nt][2 * hh + 1] * inv);
            if (q4 == 0 && live)
                p.lse[(int64_t(b) * p.H + h0 + hh) * p.N + t0 + tok] = (m[hh] + __log2f(l)) * 0.6931471805599453f;
        }
        __syncwarp();
#pragma unroll
        for (int e = lane; e < 8 * (G::QB / 16); e += 32) {
            const int row = warp * 8 + e / (G::QB / 16), ch = e % (G::QB / 16);
            if (row < ti)
                *reinterpret_cast<uint4 *>(out_tile + int64_t(row) * p.out.sN + ch * 8) =
                    *reinterpret_cast<const uint4 *>(gen + G::off_q(ucap) + row * G::QS + ch * 16);
        }
        __syncthreads();   // all warps are done with this item's rows before the next copies land
    }
}

template <typename T, int C, int HG>
static int launch_cl_fwd(const ClFwdParams &p, cudaStream_t st) {
    using G = ClGeom<C, HG>;
    const int smem = G::total(p.ucap);
    if (smem > 227 * 1024) { set_error("fused_cl_fwd: union of %d blocks does not fit shared memory", p.ucap); return AFFB200_E_ARG; }
    auto kern = fused_cl_fwd_kernel<T, C, HG>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("fused_cl_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return AFFB200_E_CUDA; }
    const int64_t items = int64_t(p.H / HG) * p.B * p.n_tiles;
    if (items >= (int64_t(1) << 31)) { set_error("fused_cl_fwd: too many tiles"); return AFFB200_E_ARG; }
    const int grid = int(items < 2 * kNumSMs ? items : 2 * kNumSMs);
    kern<<<grid, kClThreads, smem, st>>>(p);
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
    p.q = ClView{q->p, q->sB, q->sN, q->sH};
    p.kv = ClView{k->p, k->sB, k->sN, k->sH};
    p.out = ClView{out->p, out->sB, out->sN, out->sH};
    p.table = table; p.blank_k = blank_k; p.blank_v = blank_v;
    p.hdr = plan->q_hdr; p.ucl = plan->q_ucl; p.qmask = reinterpret_cast<const unsigned long long *>(plan->q_mask);
    p.pos16 = plan->pos16; p.lse = lse; p.guard = guard;
    p.B = int(B); p.H = int(H); p.N = int(N); p.n_tiles = int((N + kClT - 1) / kClT); p.ucap = plan->ucap;
    p.w = int(rel_pos_width); p.tw = int(table_width);
    p.scale_log2 = scale * kLog2e;
    auto st = (cudaStream_t)stream;
    if (dtype == AFFB200_BF16)
        return c == 32 ? launch_cl_fwd<__nv_bfloat16, 32, 2>(p, st) : launch_cl_fwd<__nv_bfloat16, 16, 2>(p, st);
    return c == 32 ? launch_cl_fwd<__half, 32, 2>(p, st) : launch_cl_fwd<__half, 16, 2>(p, st);
}
