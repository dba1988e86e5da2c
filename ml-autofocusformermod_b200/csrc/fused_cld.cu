// fused_cld.cu — block-structured neighbourhood attention for images of at most 128 tokens (sm_100a): DENSE masked
// attention over the image with every operand resident on chip.  (Stage 3 of the 224-pixel AFF configs: 125 tokens,
// 18 of the 25 blocks of AFF-Tiny-1/5; reference models/aff_transformer.py:120-155 between the Linears.)
//
// With clusters of 8 consecutive rows an image of n <= 128 tokens has at most 16 key blocks, so a token's
// neighbourhood is a 16-bit membership word and one CTA per (image, head) can hold the image's K|V (16 TMA boxes of
// 8 rows x 128 bytes, 128-byte swizzle), the packed positions, and the head's relative-position table (the central
// (2R+1)^2 rows in the storage type, <= 32 KB) in shared memory.  A warp owns 16 query tokens: S = Q·K^T per key block
// in mma.sync accumulators, bias = table[pos_u - pos_i] by one 2-byte shared-memory load, membership by a bit test,
// online softmax in registers, P fed back as the A operand of P·V — no S/P tile, no per-slot gather, no plan beyond
// the block ids.  The blank token is key row 0 of a 17th pseudo-block.
//
// Backward (same CTA shape, two phases):
//   A  per 16 query tokens: S, P = exp2(S - lse), dP = dO·V^T, dS = P∘(dP - delta), dQ = scale·dS·K in registers;
//      P and dS go to two swizzled [128][128] tiles in shared memory and dS also to ds_work (ascending-block slot
//      order, what affb200_pe_feat_grad / affb200_table_grad_u16 consume with plan->q_pe);
//   B  per 16 key rows: dV = P^T·dO and dK = scale·dS^T·Q accumulate in registers over the token tiles whose union
//      holds the rows (transposed ldmatrix of the two tiles) — no atomics, fixed order.  Blank-token gradients per CTA.
#include "fused_cl.cuh"

namespace affb200 {

constexpr int kCdThreads = 256;      // 8 warps x 16 query tokens
constexpr int kCdN = 128;            // tokens per image, at most
constexpr int kCdC = 32;             // head dim
constexpr int kCdBlank = 16;         // index of the blank pseudo-block
constexpr int kCdTabMax = 16384;     // table entries (2 bytes each) that fit: (2R+1)^2 <= 16384 -> R <= 63

struct CdParams {
    ClView q, out, dout, dq, dk, dv;     // [B][N][H][c] element strides (dk / dv: the halves of a kv row)
    const void *table_t;                 // [H][tab_stride] storage type, entries (dy + w) * tw + (dx + w)
    const float *blank_k, *blank_v;      // [H][c] fp32
    const uint16_t *blk;                 // [B][N][NB] block ids (kClAbsent: fully padded group)
    const uint32_t *pos16;               // [B][npad]
    float *lse;                          // [B][H][N]
    void *ds;                            // [B][H][N][M] storage type
    float *blank_part;                   // [B][H][2][c]
    int B, H, N, NB, npad, w, tw, tab_stride;
    float scale_log2, scale;
};

// shared-memory layout (bytes)
constexpr int CD_BAR = 0;
constexpr int CD_UNI = 16;                        // per-warp union words [8] u32
constexpr int CD_PI = 64;                         // packed positions [128] u32
constexpr int CD_MEM = CD_PI + 512;               // membership words [128] u16 (+ pad)
constexpr int CD_PB = CD_MEM + 256;               // backward: P_blank, dS_blank [2][128] fp32
constexpr int CD_KV = 2048;                       // 17 boxes of 1024 bytes
static_assert(CD_PB + 2 * 128 * 4 <= CD_KV, "blank scratch overlaps the K|V boxes");
constexpr int CD_TAB = CD_KV + 17 * 1024;         // table, <= 32 KB
__host__ __device__ constexpr int cd_tab_bytes(int tw) { return ((tw * tw + 7) / 8) * 16; }
// backward only, after the table: P tile, dS tile (256-byte rows, swizzled).  The Q and dO tiles of phase B (64-byte
// rows, 64-byte swizzle) take the place of the K|V boxes, which are dead after phase A.
__host__ __device__ constexpr int cd_off_p(int tw) { return (CD_TAB + cd_tab_bytes(tw) + 127) / 128 * 128; }

template <typename T> __device__ __forceinline__ float cd_to_f(uint32_t bits16);
template <> __device__ __forceinline__ float cd_to_f<__nv_bfloat16>(uint32_t b) { return __uint_as_float(b << 16); }
template <> __device__ __forceinline__ float cd_to_f<__half>(uint32_t b) { return __half2float(__ushort_as_half((unsigned short)b)); }

// bias (already times log2 e) of (token with biased position pib, key with position pu)
template <typename T>
__device__ __forceinline__ float cd_bias(uint32_t s_tab, uint32_t pu, uint32_t pib, int tw, uint32_t last) {
    const uint32_t d = pu - pib;
    const uint32_t idx = min((d >> 16) * uint32_t(tw) + (d & 0xffffu), last);   // clamp: padded rows, tripped guards
    return cd_to_f<T>(lds_u16(s_tab + idx * 2u)) * kLog2e;
}

// swizzled offsets: 64-byte rows (Q / dO tiles), 256-byte rows (P / dS tiles), 16-byte chunk j of row r
__device__ __forceinline__ uint32_t cd_sw64(int r, int j) { return uint32_t(r * 64 + ((j ^ ((r >> 1) & 3)) << 4)); }
__device__ __forceinline__ uint32_t cd_sw256(int r, int j) { return uint32_t(r * 256 + ((j ^ (r & 7)) << 4)); }

// common prologue: barrier, TMA boxes of the image's K|V for head h, table, positions; membership words; blank block.
// Returns after everything has landed (all threads).
template <typename T>
__device__ __forceinline__ void cd_stage(const CUtensorMap *tm_kv, const CdParams &p, int b, int h, uint32_t sm) {
    unsigned char *gen = affb200_dsm;
    const int tid = threadIdx.x;
    const int nblk = (p.N + 7) >> 3;
    const uint32_t bar = sm + CD_BAR;
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    const int pos_n = min(kCdN, p.npad);
    const uint32_t tab_bytes = uint32_t(cd_tab_bytes(p.tw));
    if (tid == 0) {
        mbar_expect_tx(bar, uint32_t(nblk) * 1024u + tab_bytes + uint32_t(pos_n) * 4u);
        bulk_g2s(sm + CD_TAB, static_cast<const unsigned char *>(p.table_t) + int64_t(h) * p.tab_stride * 2, tab_bytes, bar);
        bulk_g2s(sm + CD_PI, p.pos16 + int64_t(b) * p.npad, uint32_t(pos_n) * 4u, bar);
    }
    if (tid < nblk) tma_load_3d(sm + CD_KV + tid * 1024, tm_kv, h * 2 * kCdC, tid * 8, b, bar);
    // meanwhile: membership words from the block ids, invalid positions beyond the copied range, the blank block
    if (tid < kCdN) {
        uint32_t mk = 0;
        if (tid < p.N) {
            const uint16_t *bp = p.blk + (int64_t(b) * p.N + tid) * p.NB;
            for (int j = 0; j < p.NB; ++j) {
                const uint32_t id = __ldg(bp + j);
                if (id < 16u) mk |= 1u << id;
            }
        }
        reinterpret_cast<uint16_t *>(gen + CD_MEM)[tid] = uint16_t(mk);
        if (tid >= pos_n) reinterpret_cast<uint32_t *>(gen + CD_PI)[tid] = kClPosInvalid;
    }
    for (int e = tid; e < 1024 / 4; e += kCdThreads) reinterpret_cast<uint32_t *>(gen + CD_KV + kCdBlank * 1024)[e] = 0u;
    // boxes past the image's last block are never read as keys, but phase B of the backward reads whole key pairs
    for (int e = tid + nblk * 256; e < 16 * 256; e += kCdThreads) reinterpret_cast<uint32_t *>(gen + CD_KV)[e] = 0u;
    __syncthreads();
    if (tid < 2 * kCdC) {   // row 0 of the blank block: K chunks 0-3, V chunks 4-7 (row 0 is stored unswizzled)
        const float v = tid < kCdC ? __ldg(p.blank_k + h * kCdC + tid) : __ldg(p.blank_v + h * kCdC + tid - kCdC);
        reinterpret_cast<T *>(gen + CD_KV + kCdBlank * 1024)[tid] = Elem<T>::from_acc(v);
    }
    mbar_wait(bar, 0);
    __syncthreads();
}

// A fragments (16 rows x 32 channels) of rows r0 + g, r0 + g + 8 straight from global memory; rows >= n read as zero
template <typename T>
__device__ __forceinline__ void cd_load_a_global(uint32_t (&a)[2][4], const ClView &v, int b, int h, int r0, int n, int lane) {
    const int g = lane >> 2, q4 = lane & 3;
    const T *base = static_cast<const T *>(v.p) + int64_t(b) * v.sB + int64_t(h) * v.sH;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const int row = r0 + g + half * 8;
        const uint32_t *rp = reinterpret_cast<const uint32_t *>(base + int64_t(row) * v.sN);
#pragma unroll
        for (int ks = 0; ks < 2; ++ks) {
            a[ks][half] = row < n ? __ldg(rp + ks * 8 + q4) : 0u;
            a[ks][half + 2] = row < n ? __ldg(rp + ks * 8 + 4 + q4) : 0u;
        }
    }
}

// S (or dP) n-tile of one key block: acc = A · X_blk^T, X = K (chunks 0-3) or V (chunks 4-7) of the block's 8 rows
template <typename T, bool IS_V>
__device__ __forceinline__ void cd_dot_block(float (&acc)[4], const uint32_t (&a)[2][4], uint32_t blk_addr, int lane) {
    uint32_t f[4];
    cl_ldsm_x4(f, blk_addr + cl_sw(lane & 7, (IS_V ? 4 : 0) + (lane >> 3)));
    acc[0] = acc[1] = acc[2] = acc[3] = 0.f;
    f_mma<T>(acc, a[0], f[0], f[1]);
    f_mma<T>(acc, a[1], f[2], f[3]);
}

// acc[nt] += P · X for the 16 rows of two blocks (k = their rows), X = V (chunks 4-7) or K (chunks 0-3), through
// ldmatrix.trans; blk0 / blk1 = shared addresses of the two blocks
template <typename T, bool IS_V>
__device__ __forceinline__ void cd_accum_pair(float (&acc)[4][4], const uint32_t (&pa)[4], uint32_t blk0, uint32_t blk1, int lane) {
    const uint32_t mine = ((lane >> 3) & 1) ? blk1 : blk0;
    uint32_t f[4];
#pragma unroll
    for (int np = 0; np < 2; ++np) {
        cl_ldsm_x4_t(f, mine + cl_sw(lane & 7, (IS_V ? 4 : 0) + np * 2 + (lane >> 4)));
        f_mma<T>(acc[2 * np], pa, f[0], f[1]);
        f_mma<T>(acc[2 * np + 1], pa, f[2], f[3]);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kCdThreads)
fused_cld_fwd_kernel(const __grid_constant__ CUtensorMap tm_kv, const CdParams p) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, q4 = lane & 3;
    const int b = blockIdx.x / p.H, h = blockIdx.x % p.H;
    const uint32_t sm = dsm_addr();
    cd_stage<T>(&tm_kv, p, b, h, sm);
    const int r0 = warp * 16;
    if (r0 >= p.N) return;
    const uint32_t s_kv = sm + CD_KV, s_tab = sm + CD_TAB, s_pi = sm + CD_PI;
    const uint32_t last = uint32_t(p.tw * p.tw - 1);
    const uint32_t wbias = (uint32_t(p.w) << 16) | uint32_t(p.w);

    uint32_t qa[2][4];
    cd_load_a_global<T>(qa, p.q, b, h, r0, p.N, lane);
    const int rowA = r0 + g, rowB = rowA + 8;
    const uint32_t mkA = lds_u16(sm + CD_MEM + rowA * 2), mkB = lds_u16(sm + CD_MEM + rowB * 2);
    const uint32_t pibA = lds_u32(s_pi + rowA * 4) - wbias, pibB = lds_u32(s_pi + rowB * 4) - wbias;
    const uint32_t uni = __reduce_or_sync(0xffffffffu, mkA | mkB);
    const uint32_t onesb = g == 0 ? cl_ones2<T>() : 0u;

    float o[4][4], ol[4], m[2];
    {   // blank token first: finite running maxima from the start
        float s[4];
        cd_dot_block<T, false>(s, qa, s_kv + kCdBlank * 1024, lane);
        m[0] = __shfl_sync(0xffffffffu, s[0], lane & ~3) * p.scale_log2;
        m[1] = __shfl_sync(0xffffffffu, s[2], lane & ~3) * p.scale_log2;
        ol[0] = ol[2] = q4 == 0 ? 1.f : 0.f;     // p_blank = exp2(0) = 1 in column 0
        ol[1] = ol[3] = 0.f;
        // O starts as 1 * blank_v (as staged: rounded to T)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
            const uint32_t v = lds_u32(s_kv + kCdBlank * 1024 + 64 + nt * 16 + q4 * 4);
            const T *e = reinterpret_cast<const T *>(&v);
            o[nt][0] = o[nt][2] = Elem<T>::to_acc(e[0]);
            o[nt][1] = o[nt][3] = Elem<T>::to_acc(e[1]);
        }
    }
    for (int jp = 0; jp < 8; ++jp) {
        if (((uni >> (2 * jp)) & 3u) == 0u) continue;
        float s[2][4];
        float lm[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int blk = 2 * jp + j;
            cd_dot_block<T, false>(s[j], qa, s_kv + blk * 1024, lane);
            const uint2 pu = lds_u64(s_pi + (blk * 8 + q4 * 2) * 4);
            const bool inA = (mkA >> blk) & 1u, inB = (mkB >> blk) & 1u;
            const bool v0 = pu.x != kClPosInvalid, v1 = pu.y != kClPosInvalid;
            s[j][0] = inA && v0 ? fmaf(s[j][0], p.scale_log2, cd_bias<T>(s_tab, pu.x, pibA, p.tw, last)) : -INFINITY;
            s[j][1] = inA && v1 ? fmaf(s[j][1], p.scale_log2, cd_bias<T>(s_tab, pu.y, pibA, p.tw, last)) : -INFINITY;
            s[j][2] = inB && v0 ? fmaf(s[j][2], p.scale_log2, cd_bias<T>(s_tab, pu.x, pibB, p.tw, last)) : -INFINITY;
            s[j][3] = inB && v1 ? fmaf(s[j][3], p.scale_log2, cd_bias<T>(s_tab, pu.y, pibB, p.tw, last)) : -INFINITY;
            lm[0] = fmaxf(lm[0], fmaxf(s[j][0], s[j][1]));
            lm[1] = fmaxf(lm[1], fmaxf(s[j][2], s[j][3]));
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
                for (int nt = 0; nt < 4; ++nt) { o[nt][2 * hh] *= f; o[nt][2 * hh + 1] *= f; }
            }
        }
        uint32_t pa[4];
        pa[0] = f_pack2<T>(f_ex2(s[0][0] - m[0]), f_ex2(s[0][1] - m[0]));
        pa[1] = f_pack2<T>(f_ex2(s[0][2] - m[1]), f_ex2(s[0][3] - m[1]));
        pa[2] = f_pack2<T>(f_ex2(s[1][0] - m[0]), f_ex2(s[1][1] - m[0]));
        pa[3] = f_pack2<T>(f_ex2(s[1][2] - m[1]), f_ex2(s[1][3] - m[1]));
        f_mma<T>(ol, pa, onesb, onesb);      // row sums of the rounded probabilities (column 0)
        cd_accum_pair<T, true>(o, pa, s_kv + (2 * jp) * 1024, s_kv + (2 * jp + 1) * 1024, lane);
    }
    T *outp = static_cast<T *>(const_cast<void *>(p.out.p)) + int64_t(b) * p.out.sB + int64_t(h) * p.out.sH;
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
        const int row = hh ? rowB : rowA;
        const float l = __shfl_sync(0xffffffffu, ol[2 * hh], lane & ~3);
        const float inv = __fdividef(1.f, l);
        if (row < p.N) {
            uint32_t *rp = reinterpret_cast<uint32_t *>(outp + int64_t(row) * p.out.sN);
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) rp[nt * 4 + q4] = f_pack2<T>(o[nt][2 * hh] * inv, o[nt][2 * hh + 1] * inv);
            if (q4 == 0) p.lse[(int64_t(b) * p.H + h) * p.N + row] = (m[hh] + __log2f(l)) * 0.6931471805599453f;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// backward
// ---------------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kCdThreads)
fused_cld_bwd_kernel(const __grid_constant__ CUtensorMap tm_kv, const CdParams p) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, q4 = lane & 3;
    const int b = blockIdx.x / p.H, h = blockIdx.x % p.H;
    const uint32_t sm = dsm_addr();
    unsigned char *gen = affb200_dsm;
    const int off_q = CD_KV, off_do = CD_KV + kCdN * 64, off_p = cd_off_p(p.tw), off_ds = off_p + kCdN * 256;
    // (the P / dS tiles need no clearing: phase B reads a (token tile, key pair) only if phase A wrote all of it — both
    //  decide by the same union word)
    for (int e = tid; e < 2 * kCdN; e += kCdThreads) reinterpret_cast<float *>(gen + CD_PB)[e] = 0.f;
    if (tid < 8) reinterpret_cast<uint32_t *>(gen + CD_UNI)[tid] = 0u;
    cd_stage<T>(&tm_kv, p, b, h, sm);
    const uint32_t s_kv = sm + CD_KV, s_tab = sm + CD_TAB, s_pi = sm + CD_PI;
    const uint32_t last = uint32_t(p.tw * p.tw - 1);
    const uint32_t wbias = (uint32_t(p.w) << 16) | uint32_t(p.w);
    const int r0 = warp * 16;
    const int M = p.NB * 8;

    // ---- phase A: the warp's 16 query tokens ---------------------------------------------------------------------
    if (r0 < p.N) {
        uint32_t qa[2][4], ga[2][4];
        cd_load_a_global<T>(qa, p.q, b, h, r0, p.N, lane);
        cd_load_a_global<T>(ga, p.dout, b, h, r0, p.N, lane);
        const int rowA = r0 + g, rowB = rowA + 8;
        const bool okA = rowA < p.N, okB = rowB < p.N;
        const uint32_t mkA = lds_u16(sm + CD_MEM + rowA * 2), mkB = lds_u16(sm + CD_MEM + rowB * 2);
        const uint32_t pibA = lds_u32(s_pi + rowA * 4) - wbias, pibB = lds_u32(s_pi + rowB * 4) - wbias;
        const uint32_t uni = __reduce_or_sync(0xffffffffu, mkA | mkB);
        if (lane == 0) reinterpret_cast<uint32_t *>(gen + CD_UNI)[warp] = uni;
        // delta = rowsum(dO ∘ O), lse (log2 domain)
        float delta[2], lse2[2];
        {
            const T *ob = static_cast<const T *>(p.out.p) + int64_t(b) * p.out.sB + int64_t(h) * p.out.sH;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const int row = hh ? rowB : rowA;
                float acc = 0.f;
                if (row < p.N) {
                    const uint32_t *rp = reinterpret_cast<const uint32_t *>(ob + int64_t(row) * p.out.sN);
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks)
#pragma unroll
                        for (int c2 = 0; c2 < 2; ++c2) {
                            const uint32_t ov = __ldg(rp + ks * 8 + c2 * 4 + q4), gv = ga[ks][hh + 2 * c2];
                            const T *oe = reinterpret_cast<const T *>(&ov), *ge = reinterpret_cast<const T *>(&gv);
                            acc = fmaf(Elem<T>::to_acc(oe[0]), Elem<T>::to_acc(ge[0]), acc);
                            acc = fmaf(Elem<T>::to_acc(oe[1]), Elem<T>::to_acc(ge[1]), acc);
                        }
                }
                acc += __shfl_xor_sync(0xffffffffu, acc, 1);
                acc += __shfl_xor_sync(0xffffffffu, acc, 2);
                delta[hh] = acc;
                lse2[hh] = row < p.N ? __ldg(p.lse + (int64_t(b) * p.H + h) * p.N + row) * kLog2e : 0.f;
            }
        }
        T *dsb = static_cast<T *>(p.ds) + (int64_t(b) * p.H + h) * p.N * M;
        float dq[4][4];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) dq[nt][0] = dq[nt][1] = dq[nt][2] = dq[nt][3] = 0.f;
        {   // blank token
            float s[4], dp[4];
            cd_dot_block<T, false>(s, qa, s_kv + kCdBlank * 1024, lane);
            cd_dot_block<T, true>(dp, ga, s_kv + kCdBlank * 1024, lane);
            const float pA = q4 == 0 ? f_ex2(s[0] * p.scale_log2 - lse2[0]) : 0.f;
            const float pB = q4 == 0 ? f_ex2(s[2] * p.scale_log2 - lse2[1]) : 0.f;
            const float dA = pA * (dp[0] - delta[0]), dB = pB * (dp[2] - delta[1]);
            if (q4 == 0) {
                float *pb = reinterpret_cast<float *>(gen + CD_PB);
                if (okA) { pb[rowA] = pA; pb[kCdN + rowA] = dA; }
                if (okB) { pb[rowB] = pB; pb[kCdN + rowB] = dB; }
            }
            uint32_t da[4] = {f_pack2<T>(dA, 0.f), f_pack2<T>(dB, 0.f), 0u, 0u};
            cd_accum_pair<T, false>(dq, da, s_kv + kCdBlank * 1024, s_kv + kCdBlank * 1024, lane);
        }
        for (int jp = 0; jp < 8; ++jp) {
            if (((uni >> (2 * jp)) & 3u) == 0u) continue;
            uint32_t da[4];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int blk = 2 * jp + j;
                float s[4], dp[4];
                cd_dot_block<T, false>(s, qa, s_kv + blk * 1024, lane);
                cd_dot_block<T, true>(dp, ga, s_kv + blk * 1024, lane);
                const uint2 pu = lds_u64(s_pi + (blk * 8 + q4 * 2) * 4);
                const bool inA = (mkA >> blk) & 1u, inB = (mkB >> blk) & 1u;
                const bool v0 = pu.x != kClPosInvalid, v1 = pu.y != kClPosInvalid;
                float pr[4];
                pr[0] = inA && v0 ? f_ex2(fmaf(s[0], p.scale_log2, cd_bias<T>(s_tab, pu.x, pibA, p.tw, last)) - lse2[0]) : 0.f;
                pr[1] = inA && v1 ? f_ex2(fmaf(s[1], p.scale_log2, cd_bias<T>(s_tab, pu.y, pibA, p.tw, last)) - lse2[0]) : 0.f;
                pr[2] = inB && v0 ? f_ex2(fmaf(s[2], p.scale_log2, cd_bias<T>(s_tab, pu.x, pibB, p.tw, last)) - lse2[1]) : 0.f;
                pr[3] = inB && v1 ? f_ex2(fmaf(s[3], p.scale_log2, cd_bias<T>(s_tab, pu.y, pibB, p.tw, last)) - lse2[1]) : 0.f;
                const float d0 = pr[0] * (dp[0] - delta[0]), d1 = pr[1] * (dp[1] - delta[0]);
                const float d2 = pr[2] * (dp[2] - delta[1]), d3 = pr[3] * (dp[3] - delta[1]);
                const uint32_t pwA = f_pack2<T>(pr[0], pr[1]), pwB = f_pack2<T>(pr[2], pr[3]);
                da[2 * j] = f_pack2<T>(d0, d1);
                da[2 * j + 1] = f_pack2<T>(d2, d3);
                *reinterpret_cast<uint32_t *>(gen + off_p + cd_sw256(rowA, blk) + q4 * 4) = pwA;
                *reinterpret_cast<uint32_t *>(gen + off_p + cd_sw256(rowB, blk) + q4 * 4) = pwB;
                *reinterpret_cast<uint32_t *>(gen + off_ds + cd_sw256(rowA, blk) + q4 * 4) = da[2 * j];
                *reinterpret_cast<uint32_t *>(gen + off_ds + cd_sw256(rowB, blk) + q4 * 4) = da[2 * j + 1];
            }
            cd_accum_pair<T, false>(dq, da, s_kv + (2 * jp) * 1024, s_kv + (2 * jp + 1) * 1024, lane);
        }
        // dS for the bias gradient, in ascending-block slot order: the warp copies its 16 rows' member chunks out of the
        // dS tile, 16 bytes per (token, block); slots behind the token's real blocks (fully padded groups) are zero
        __syncwarp();
        for (int e = lane; e < 16 * p.NB; e += 32) {
            const int row = r0 + e / p.NB, s8 = e % p.NB;
            if (row >= p.N) continue;
            uint32_t mk = lds_u16(sm + CD_MEM + row * 2);
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (s8 < __popc(mk)) {
                for (int k = 0; k < s8; ++k) mk &= mk - 1u;          // drop the s8 lowest set bits
                v = *reinterpret_cast<const uint4 *>(gen + off_ds + cd_sw256(row, __ffs(mk) - 1));
            }
            *reinterpret_cast<uint4 *>(dsb + int64_t(row) * M + s8 * 8) = v;
        }
        T *dqb = static_cast<T *>(const_cast<void *>(p.dq.p)) + int64_t(b) * p.dq.sB + int64_t(h) * p.dq.sH;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const int row = hh ? rowB : rowA;
            if (row < p.N) {
                uint32_t *rp = reinterpret_cast<uint32_t *>(dqb + int64_t(row) * p.dq.sN);
#pragma unroll
                for (int nt = 0; nt < 4; ++nt) rp[nt * 4 + q4] = f_pack2<T>(dq[nt][2 * hh] * p.scale, dq[nt][2 * hh + 1] * p.scale);
            }
        }
    }
    __syncthreads();
    {   // Q and dO tiles (rows >= n zero) over the K|V boxes
        const T *qb = static_cast<const T *>(p.q.p) + int64_t(b) * p.q.sB + int64_t(h) * p.q.sH;
        const T *gb = static_cast<const T *>(p.dout.p) + int64_t(b) * p.dout.sB + int64_t(h) * p.dout.sH;
        for (int e = tid; e < kCdN * 4; e += kCdThreads) {
            const int row = e >> 2, ch = e & 3;
            uint4 vq = make_uint4(0u, 0u, 0u, 0u), vg = vq;
            if (row < p.N) {
                vq = __ldg(reinterpret_cast<const uint4 *>(qb + int64_t(row) * p.q.sN) + ch);
                vg = __ldg(reinterpret_cast<const uint4 *>(gb + int64_t(row) * p.dout.sN) + ch);
            }
            *reinterpret_cast<uint4 *>(gen + off_q + cd_sw64(row, ch)) = vq;
            *reinterpret_cast<uint4 *>(gen + off_do + cd_sw64(row, ch)) = vg;
        }
    }
    __syncthreads();

    // ---- phase B: the warp's 16 key rows (blocks 2w, 2w + 1) -----------------------------------------------------
    const int k0 = warp * 16;
    if (k0 < p.N) {
        float dk[4][4], dv[4][4];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) dk[nt][e] = dv[nt][e] = 0.f;
        const uint32_t mine = 3u << (2 * warp);
        for (int tt = 0; tt < 8; ++tt) {
            if ((reinterpret_cast<const uint32_t *>(gen + CD_UNI)[tt] & mine) == 0u) continue;
            const int tok = tt * 16 + (lane & 7) + ((lane >> 4) & 1) * 8;      // row of the tile this lane addresses
            const int chunk = 2 * warp + ((lane >> 3) & 1);
            uint32_t pa[4], sa[4], f[4];
            cl_ldsm_x4_t(pa, sm + off_p + cd_sw256(tok, chunk));
            cl_ldsm_x4_t(sa, sm + off_ds + cd_sw256(tok, chunk));
            const int trow = tt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;     // k rows of the B operand
#pragma unroll
            for (int np = 0; np < 2; ++np) {
                cl_ldsm_x4_t(f, sm + off_do + cd_sw64(trow, np * 2 + (lane >> 4)));
                f_mma<T>(dv[2 * np], pa, f[0], f[1]);
                f_mma<T>(dv[2 * np + 1], pa, f[2], f[3]);
                cl_ldsm_x4_t(f, sm + off_q + cd_sw64(trow, np * 2 + (lane >> 4)));
                f_mma<T>(dk[2 * np], sa, f[0], f[1]);
                f_mma<T>(dk[2 * np + 1], sa, f[2], f[3]);
            }
        }
        T *dkb = static_cast<T *>(const_cast<void *>(p.dk.p)) + int64_t(b) * p.dk.sB + int64_t(h) * p.dk.sH;
        T *dvb = static_cast<T *>(const_cast<void *>(p.dv.p)) + int64_t(b) * p.dv.sB + int64_t(h) * p.dv.sH;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const int row = k0 + g + hh * 8;
            if (row < p.N) {
                uint32_t *kp = reinterpret_cast<uint32_t *>(dkb + int64_t(row) * p.dk.sN);
                uint32_t *vp = reinterpret_cast<uint32_t *>(dvb + int64_t(row) * p.dv.sN);
#pragma unroll
                for (int nt = 0; nt < 4; ++nt) {
                    kp[nt * 4 + q4] = f_pack2<T>(dk[nt][2 * hh] * p.scale, dk[nt][2 * hh + 1] * p.scale);
                    vp[nt * 4 + q4] = f_pack2<T>(dv[nt][2 * hh], dv[nt][2 * hh + 1]);
                }
            }
        }
    }
    // ---- blank-token gradients: d blank_k = scale * sum_tok dS_blank[tok] Q[tok], d blank_v = sum_tok P_blank[tok] dO[tok]
    {
        const float *pb = reinterpret_cast<const float *>(gen + CD_PB);
        __syncthreads();                                                     // phase B is done with the P tile
        float *acc = reinterpret_cast<float *>(gen + off_p);                 // [8 warps][2][32] over the P tile
        const int ch = lane;       // one channel per lane, 16 tokens per warp
        float ak = 0.f, av = 0.f;
        for (int t = warp * 16; t < min(p.N, warp * 16 + 16); ++t) {
            const T qv = *reinterpret_cast<const T *>(gen + off_q + cd_sw64(t, ch >> 3) + (ch & 7) * 2);
            const T gv = *reinterpret_cast<const T *>(gen + off_do + cd_sw64(t, ch >> 3) + (ch & 7) * 2);
            ak = fmaf(pb[kCdN + t], Elem<T>::to_acc(qv), ak);
            av = fmaf(pb[t], Elem<T>::to_acc(gv), av);
        }
        acc[warp * 64 + ch] = ak;
        acc[warp * 64 + 32 + ch] = av;
        __syncthreads();
        if (tid < 64) {     // fixed-order sum over the 8 warps
            float v = 0.f;
#pragma unroll
            for (int wv = 0; wv < 8; ++wv) v += acc[wv * 64 + tid];
            p.blank_part[((int64_t(b) * p.H + h) * 2 + (tid >> 5)) * kCdC + (tid & 31)] = v * (tid < 32 ? p.scale : 1.f);
        }
    }
}

static int cd_check(const char *what, const affb200_view *q, const affb200_view *k, const affb200_view *v,
                    const affb200_view *out, int64_t B, int64_t H, int64_t N, int64_t M, int64_t c, int64_t w, int64_t tw,
                    int64_t tab_stride, int64_t npad, int dtype) {
    if (B <= 0 || H <= 0 || N <= 0 || N > kCdN || M <= 0 || M % 8 != 0 || M > 128 || c != kCdC ||
        (dtype != AFFB200_F16 && dtype != AFFB200_BF16) || w < 0 || tw != 2 * w + 1 || tw * tw > kCdTabMax ||
        tab_stride < tw * tw || tab_stride % 8 != 0 || npad < N || npad % 64 != 0 || B * H >= (int64_t(1) << 31)) {
        set_error("%s: needs N <= 128, head dim 32, f16/bf16, M a multiple of 8, a table of (2w+1)^2 <= 16384 rows with a "
                  "row stride that is a multiple of 8 (B=%lld H=%lld N=%lld M=%lld c=%lld w=%lld tw=%lld)", what,
                  (long long)B, (long long)H, (long long)N, (long long)M, (long long)c, (long long)w, (long long)tw);
        return AFFB200_E_ARG;
    }
    if (reinterpret_cast<uintptr_t>(v->p) != reinterpret_cast<uintptr_t>(k->p) + uintptr_t(c * 2) || k->sH != 2 * c ||
        v->sH != 2 * c || k->sN != v->sN || k->sB != v->sB) {
        set_error("%s: k / v must be the [H][2][c] halves of one kv row", what);
        return AFFB200_E_ARG;
    }
    if (!aligned16(q->p) || !aligned16(k->p) || !aligned16(out->p) || (q->sN * 2) % 16 || (k->sN * 2) % 16 ||
        (out->sN * 2) % 16 || (q->sB * 2) % 16 || (k->sB * 2) % 16 || (out->sB * 2) % 16 || (q->sH * 2) % 16 || (out->sH * 2) % 16) {
        set_error("%s: rows must be 16-byte aligned", what);
        return AFFB200_E_ARG;
    }
    return AFFB200_OK;
}

}  // namespace affb200

using namespace affb200;

extern "C" int affb200_fused_cld_fwd(const affb200_view *q, const affb200_view *k, const affb200_view *v,
                                     const affb200_view *out, const void *table_t, int64_t tab_stride,
                                     const float *blank_k, const float *blank_v, const uint16_t *blk,
                                     const uint32_t *pos16, int64_t npad, float *lse, int64_t B, int64_t H, int64_t N,
                                     int64_t M, int64_t c, int64_t rel_pos_width, int64_t table_width, float scale,
                                     int dtype, void *stream) {
    if (!q || !k || !v || !out || !table_t || !blank_k || !blank_v || !blk || !pos16 || !lse) {
        set_error("fused_cld_fwd: null pointer");
        return AFFB200_E_ARG;
    }
    if (int e = cd_check("fused_cld_fwd", q, k, v, out, B, H, N, M, c, rel_pos_width, table_width, tab_stride, npad, dtype)) return e;
    CdParams p{};
    p.q = ClView{q->p, q->sB, q->sN, q->sH};
    p.out = ClView{out->p, out->sB, out->sN, out->sH};
    p.table_t = table_t; p.blank_k = blank_k; p.blank_v = blank_v; p.blk = blk; p.pos16 = pos16; p.lse = lse;
    p.B = int(B); p.H = int(H); p.N = int(N); p.NB = int(M / 8); p.npad = int(npad);
    p.w = int(rel_pos_width); p.tw = int(table_width); p.tab_stride = int(tab_stride);
    p.scale = scale; p.scale_log2 = scale * kLog2e;
    CUtensorMap tm_kv;
    if (int rc = cl_make_tensor_map(&tm_kv, k->p, dtype, 2 * H * c, N, B, k->sN, k->sB, 64, 8, false)) return rc;
    const int smem = CD_TAB + cd_tab_bytes(p.tw);
    auto st = (cudaStream_t)stream;
    auto launch = [&](auto kern) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) { set_error("fused_cld_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return int(AFFB200_E_CUDA); }
        kern<<<unsigned(B * H), kCdThreads, smem, st>>>(tm_kv, p);
        return check_launch("fused_cld_fwd");
    };
    return dtype == AFFB200_BF16 ? launch(fused_cld_fwd_kernel<__nv_bfloat16>) : launch(fused_cld_fwd_kernel<__half>);
}

extern "C" int affb200_fused_cld_bwd(const affb200_view *q, const affb200_view *k, const affb200_view *v,
                                     const affb200_view *out, const affb200_view *dout, const affb200_view *dq,
                                     const affb200_view *dk, const affb200_view *dv, const void *table_t,
                                     int64_t tab_stride, const float *blank_k, const float *blank_v,
                                     const uint16_t *blk, const uint32_t *pos16, int64_t npad, const float *lse,
                                     void *ds_work, float *blank_part, int64_t B, int64_t H, int64_t N, int64_t M,
                                     int64_t c, int64_t rel_pos_width, int64_t table_width, float scale, int dtype,
                                     void *stream) {
    if (!q || !k || !v || !out || !dout || !dq || !dk || !dv || !table_t || !blank_k || !blank_v || !blk || !pos16 || !lse ||
        !ds_work || !blank_part) {
        set_error("fused_cld_bwd: null pointer");
        return AFFB200_E_ARG;
    }
    if (int e = cd_check("fused_cld_bwd", q, k, v, out, B, H, N, M, c, rel_pos_width, table_width, tab_stride, npad, dtype)) return e;
    if (!aligned16(dout->p) || !aligned16(dq->p) || !aligned16(dk->p) || (dout->sN * 2) % 16 || (dq->sN * 2) % 16 ||
        (dk->sN * 2) % 16 || (dout->sB * 2) % 16 || (dq->sB * 2) % 16 || (dk->sB * 2) % 16 || (dout->sH * 2) % 16 ||
        (dq->sH * 2) % 16 || (dk->sH * 2) % 16 || (dv->sH * 2) % 16 || dv->sN != dk->sN || dv->sB != dk->sB) {
        set_error("fused_cld_bwd: gradient rows must be 16-byte aligned");
        return AFFB200_E_ARG;
    }
    CdParams p{};
    p.q = ClView{q->p, q->sB, q->sN, q->sH};
    p.out = ClView{out->p, out->sB, out->sN, out->sH};
    p.dout = ClView{dout->p, dout->sB, dout->sN, dout->sH};
    p.dq = ClView{dq->p, dq->sB, dq->sN, dq->sH};
    p.dk = ClView{dk->p, dk->sB, dk->sN, dk->sH};
    p.dv = ClView{dv->p, dv->sB, dv->sN, dv->sH};
    p.table_t = table_t; p.blank_k = blank_k; p.blank_v = blank_v; p.blk = blk; p.pos16 = pos16;
    p.lse = const_cast<float *>(lse); p.ds = ds_work; p.blank_part = blank_part;
    p.B = int(B); p.H = int(H); p.N = int(N); p.NB = int(M / 8); p.npad = int(npad);
    p.w = int(rel_pos_width); p.tw = int(table_width); p.tab_stride = int(tab_stride);
    p.scale = scale; p.scale_log2 = scale * kLog2e;
    CUtensorMap tm_kv;
    if (int rc = cl_make_tensor_map(&tm_kv, k->p, dtype, 2 * H * c, N, B, k->sN, k->sB, 64, 8, false)) return rc;
    const int smem = cd_off_p(p.tw) + 2 * kCdN * 256;
    auto st = (cudaStream_t)stream;
    auto launch = [&](auto kern) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) { set_error("fused_cld_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return int(AFFB200_E_CUDA); }
        kern<<<unsigned(B * H), kCdThreads, smem, st>>>(tm_kv, p);
        return check_launch("fused_cld_bwd");
    };
    return dtype == AFFB200_BF16 ? launch(fused_cld_bwd_kernel<__nv_bfloat16>) : launch(fused_cld_bwd_kernel<__half>);
}
