// fused_cl.cuh — shared pieces of the block-structured fused neighbourhood attention (fused_cl_fwd.cu,
// fused_cl_bwd.cu) and its plan builders (fused_cl_plan.cu).
//
// Structure exploited (reference models/point_utils.py:203-206, models/aff_transformer.py:476-480): after
// space_filling_cluster the tokens of a cluster are stored contiguously and a neighbourhood is the member lists of
// the nnc nearest clusters, so nbhd_idx[b,i,8q..8q+7] are 8 CONSECUTIVE rows starting at a multiple of 8 (clusters of
// 8, 16, 24 ... rows are 1, 2, 3 ... such blocks); only the padded tail of the last cluster is irregular (rows >= N,
// masked).  A neighbourhood is therefore M/8 block ids, and a tile of 64 consecutive tokens references a union of
// ~25-40 blocks (measured on the AFF pipelines).  The kernels treat a tile as DENSE MASKED attention over that union:
//   * every union row is staged once per tile with ONE bulk-async copy (cp.async.bulk -> UBLKCP, completion on an
//     mbarrier -> SYNCS) of the row's K and V for a group of heads (256 B), instead of 16-byte cp.async pieces;
//   * S = Q·K_U^T is produced block by block in mma.sync accumulator registers (an n-tile of 8 columns IS one block),
//     masked by the row's 64-bit membership word over the tile's union, biased with table[pos_u - pos_i] from a
//     shared-memory window of the relative-position table (no pe_idx stream), exponentiated in registers and fed
//     straight back as the A operand of P·V_U — no S/P round trip through shared memory, no per-slot gather;
//   * a warp owns 8 tokens (one cluster of the space-filling order) and BOTH heads of a head group: rows 0-7 of its
//     m16 tiles are head 0, rows 8-15 head 1 of the same tokens (the operand halves of the other head are zero), so
//     membership and the bias index are computed once per (token, row) for both heads, and a warp visits only the
//     union blocks its 8 tokens reference (~9 of the tile's ~30; each token uses 6).
#pragma once
#include "fused_attn.cuh"
#include <cuda.h>

namespace affb200 {

constexpr int kClT = 64;        // tokens per query tile / key rows per key tile
constexpr int kClThreads = 256; // 8 warps x (8 tokens x 2 heads)
constexpr int kClUMax = AFFB200_CL_UMAX;   // union blocks per tile (membership is a 64-bit word)
constexpr int kClWin = 32;      // side of the shared-memory window of the relative-position table
constexpr int kClWinHalf = 16;  // offsets -16 .. 15 are served from the window, the rest from the global table
constexpr uint32_t kClPosBias = (uint32_t(kClWinHalf) << 16) | uint32_t(kClWinHalf);
constexpr uint32_t kClPosInvalid = 0xffffffffu;   // packed position of a padded row (>= N)
constexpr int kClAbsent = 0xffff;                 // block id of a fully padded group of 8 slots

// ---- mbarrier + bulk async copy (PTX ISA 8.x; SASS: SYNCS.*, UBLKCP) --------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
// global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned; completes `bytes` on the mbarrier
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// tiled (TMA) copy of one box of a 3-D tensor: global -> shared, full box bytes completed on the mbarrier, out-of-range
// elements zero-filled (SASS: UTMALDG)
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, int c0, int c1, int c2, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// Host: tensor map of a [B][N][W] (W contiguous) 2-byte tensor with element strides sB, sN, box {box_w, box_rows, 1},
// 16-byte chunks XOR-swizzled inside 128-byte (or 64-byte) lines.  cuTensorMapEncodeTiled is resolved through the
// runtime (no link-time dependency on libcuda).
inline int cl_make_tensor_map(CUtensorMap *map, const void *base, int dtype, int64_t W, int64_t N, int64_t B, int64_t sN,
                              int64_t sB, int box_w, int box_rows, bool swizzle64) {
    typedef CUresult (*Fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                           const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                           CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static Fn fn = [] {
        void *pfn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &pfn, cudaEnableDefault, &q) != cudaSuccess) pfn = nullptr;
        return reinterpret_cast<Fn>(pfn);
    }();
    if (!fn) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return AFFB200_E_CUDA; }

    const cuuint64_t dims[3] = {cuuint64_t(W), cuuint64_t(N), cuuint64_t(B)};
    const cuuint64_t strides[2] = {cuuint64_t(sN) * 2u, cuuint64_t(B > 1 ? sB : sN * N) * 2u};
    const cuuint32_t box[3] = {cuuint32_t(box_w), cuuint32_t(box_rows), 1u};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    auto encode = [&]() {
        return fn(map, dtype == AFFB200_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3,
                  const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    };
    CUresult r = encode();
    if (r == CUDA_ERROR_INVALID_CONTEXT) {
        // a thread that only ever inherited its device (autograd's worker) has no driver context bound: setting the
        // current device again binds the primary context to this thread (and is legal under stream capture)
        int dev = 0;
        if (cudaGetDevice(&dev) == cudaSuccess && cudaSetDevice(dev) == cudaSuccess) r = encode();
    }
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d)", int(r)); return AFFB200_E_CUDA; }
    return AFFB200_OK;
}

__device__ __forceinline__ void cl_ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void cl_ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void cl_ldsm_x2(uint32_t (&r)[2], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t addr) {
    uint16_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint2 lds_u64(uint32_t addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ float2 lds_f2(uint32_t addr) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ float lds_f1(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

// Relative-position bias of (token i, union row u) for a group of HG heads.
//   d = pos_u - (pos_i - (16,16)) as packed (y << 16 | x) integers; when both offsets lie in [-16, 16) the two fields of
//   d are the window coordinates and the entry comes from shared memory ([32*32 + 1][HG] floats, the last entry -inf:
//   where non-members are pointed); otherwise (rare: far neighbours of sparse stages, padded rows) cl_bias_far
//   rebuilds the offsets and reads the global table [P][H].  Values are pre-multiplied by log2(e).
struct ClBiasCtx {
    uint32_t win;          // shared-memory address of the window
    const float *table;    // [P][H] fp32
    int H, h0, w, tw;      // heads, first head of the group, rel_pos_width, table_width
};
constexpr uint32_t kClWinSentinel = kClWin * kClWin;   // index of the -inf entry

__device__ __forceinline__ uint32_t cl_win_index(uint32_t pu, uint32_t pi_biased, bool member, bool &far) {
    const uint32_t d = pu - pi_biased;
    far = member && (d & 0xffe0ffe0u) != 0u;
    return member ? (((d >> 11) | (d & 31u)) & 1023u) : kClWinSentinel;
}

// a padded row (pos_u == kClPosInvalid) yields -inf
static __device__ __noinline__ float cl_bias_far(const float *__restrict__ table, int H, int h, int w, int tw, uint32_t pu,
                                          uint32_t pi_biased) {
    if (pu == kClPosInvalid) return -INFINITY;
    const uint32_t pi = pi_biased + kClPosBias;
    // clamped into the table: a replayed launch whose speculated table radius did not hold (the plan's guard reports
    // it and the result is discarded) must still not read outside the table
    const int dx = max(-w, min(w, int(pu & 0xffffu) - int(pi & 0xffffu)));
    const int dy = max(-w, min(w, int(pu >> 16) - int(pi >> 16)));
    const int pe = (dy + w) * tw + dx + w;
    return __ldg(table + int64_t(pe) * H + h) * kLog2e;
}

// fill the window for head group h0..h0+HG-1 (all threads of the CTA; caller synchronises)
template <int HG>
__device__ __forceinline__ void cl_fill_window(float *win, const float *__restrict__ table, int H, int h0, int w, int tw) {
    for (int e = threadIdx.x; e < kClWin * kClWin; e += blockDim.x) {
        const int dy = (e >> 5) - kClWinHalf, dx = (e & 31) - kClWinHalf;
        const bool ok = dy >= -w && dy <= w && dx >= -w && dx <= w;
        const int64_t pe = int64_t(dy + w) * tw + dx + w;
#pragma unroll
        for (int hh = 0; hh < HG; ++hh) win[e * HG + hh] = ok ? __ldg(table + pe * H + h0 + hh) * kLog2e : 0.f;
    }
    if (threadIdx.x < HG) win[kClWin * kClWin * HG + threadIdx.x] = -INFINITY;
}

// Shared-memory layout.  K|V blocks and the Q tile are written by TMA with the 128-byte (Q at head dim 16: 64-byte)
// XOR swizzle: the 16-byte chunk j of row r of an 8-row x 128-byte tile sits at r*128 + ((j ^ r) << 4), so the eight
// row addresses of an ldmatrix fall into eight different bank groups without padding.
template <int C> struct ClGeom {
    static constexpr int KVB = 8 * C;            // bytes of one row's K|V for a group of 2 heads: [K0 V0 K1 V1]
    static constexpr int TPB = KVB / 128;        // 1024-byte tiles per block (head dim 32: one per head)
    static constexpr int BLK = 8 * KVB;          // bytes of a staged block
    static constexpr int QB = 4 * C;             // bytes of a token's Q (or output) row for the group
    static constexpr int OFF_ACT = 128;          // per-warp list of referenced union blocks (8 x 64 bytes)
    static constexpr int OFF_MK = 640;           // membership words of the tile's tokens [64] u64
    static constexpr int OFF_PI = 1152;          // packed positions of the tile's tokens [64] u32
    static constexpr int OFF_PU = 1408;          // packed positions of the union rows [64 blocks][8] u32
    static constexpr int OFF_WIN = 3456;         // bias window [32*32 + 1][2] fp32
    static constexpr int OFF_Q = 12288;          // Q tile / output staging [64][QB]
    static constexpr int OFF_KV = OFF_Q + kClT * 128;
    __host__ __device__ static constexpr int total(int ucap) { return OFF_KV + (ucap + 1) * BLK; }
    static_assert(OFF_WIN + (kClWin * kClWin + 1) * 8 <= OFF_Q, "window overlaps the Q tile");
};

struct ClView {
    const void *p;
    int64_t sB, sN, sH;    // element strides
};


template <typename T> __device__ __forceinline__ uint32_t cl_ones2();
template <> __device__ __forceinline__ uint32_t cl_ones2<__nv_bfloat16>() { return 0x3f803f80u; }
template <> __device__ __forceinline__ uint32_t cl_ones2<__half>() { return 0x3c003c00u; }

// swizzled byte offset of chunk j of row r inside a 1024-byte tile
__device__ __forceinline__ uint32_t cl_sw(int r, int j) { return uint32_t(r * 128 + ((j ^ r) << 4)); }

// Products of the warp's 16 rows (rows 0-7: its 8 tokens under head 0, rows 8-15: the same tokens under head 1) with
// the rows of one staged block: s0 = A·X_h0^T is meaningful in its head-0 rows (registers 0, 1), s1 = A·X_h1^T in its
// head-1 rows (registers 2, 3); the other halves are ignored products.  X = K (IS_V false) or V (true).
template <typename T, int C, bool IS_V = false>
__device__ __forceinline__ void cl_scores(float (&s0)[4], float (&s1)[4], const uint32_t (&qa)[C / 16][4], uint32_t blk_addr,
                                          int lane) {
    const int r = lane & 7;
    s0[0] = s0[1] = s0[2] = s0[3] = 0.f;
    s1[0] = s1[1] = s1[2] = s1[3] = 0.f;
    uint32_t kf[4];
    if constexpr (C == 32) {
        cl_ldsm_x4(kf, blk_addr + cl_sw(r, (IS_V ? 4 : 0) + (lane >> 3)));
        f_mma<T>(s0, qa[0], kf[0], kf[1]);
        f_mma<T>(s0, qa[1], kf[2], kf[3]);
        cl_ldsm_x4(kf, blk_addr + 1024 + cl_sw(r, (IS_V ? 4 : 0) + (lane >> 3)));
        f_mma<T>(s1, qa[0], kf[0], kf[1]);
        f_mma<T>(s1, qa[1], kf[2], kf[3]);
    } else {
        cl_ldsm_x4(kf, blk_addr + cl_sw(r, (lane >> 4) * 4 + (IS_V ? 2 : 0) + ((lane >> 3) & 1)));   // head 0 | head 1
        f_mma<T>(s0, qa[0], kf[0], kf[1]);
        f_mma<T>(s1, qa[0], kf[2], kf[3]);
    }
}

// acc0[nt] += P·X_h0, acc1[nt] += P·X_h1 for the 16 rows of two staged blocks (k = the blocks' rows), X = V or K through
// ldmatrix.trans; `blk_lane` = address of the block this lane addresses (lanes 8-15 and 24-31: the second block).
// acc0 is meaningful in its head-0 rows (registers 0, 1), acc1 in its head-1 rows (registers 2, 3).
template <typename T, int C, bool IS_V>
__device__ __forceinline__ void cl_accum_rows(float (&acc0)[C / 8][4], float (&acc1)[C / 8][4], const uint32_t (&pa)[4],
                                              uint32_t blk_lane, int lane) {
    const int r = lane & 7;
    uint32_t vf[4];
    if constexpr (C == 32) {
#pragma unroll
        for (int np = 0; np < 2; ++np) {
            cl_ldsm_x4_t(vf, blk_lane + cl_sw(r, (IS_V ? 4 : 0) + np * 2 + (lane >> 4)));
            f_mma<T>(acc0[2 * np], pa, vf[0], vf[1]);
            f_mma<T>(acc0[2 * np + 1], pa, vf[2], vf[3]);
            cl_ldsm_x4_t(vf, blk_lane + 1024 + cl_sw(r, (IS_V ? 4 : 0) + np * 2 + (lane >> 4)));
            f_mma<T>(acc1[2 * np], pa, vf[0], vf[1]);
            f_mma<T>(acc1[2 * np + 1], pa, vf[2], vf[3]);
        }
    } else {
        cl_ldsm_x4_t(vf, blk_lane + cl_sw(r, (IS_V ? 2 : 0) + (lane >> 4)));
        f_mma<T>(acc0[0], pa, vf[0], vf[1]);
        f_mma<T>(acc0[1], pa, vf[2], vf[3]);
        cl_ldsm_x4_t(vf, blk_lane + cl_sw(r, 4 + (IS_V ? 2 : 0) + (lane >> 4)));
        f_mma<T>(acc1[0], pa, vf[0], vf[1]);
        f_mma<T>(acc1[1], pa, vf[2], vf[3]);
    }
}

// A fragments of the warp's 8 token rows of a swizzled [64][2 heads x C] tile: registers 0, 2 = head 0 (rows 0-7 of the
// m16 tile), 1, 3 = head 1 (rows 8-15)
template <int C>
__device__ __forceinline__ void cl_load_afrag(uint32_t (&qa)[C / 16][4], uint32_t tile_addr, int warp, int lane) {
    const int r8 = lane & 7, row = warp * 8 + r8, i = lane >> 3;
#pragma unroll
    for (int ks = 0; ks < C / 16; ++ks) {
        if constexpr (C == 32) cl_ldsm_x4(qa[ks], tile_addr + row * 128 + ((((i & 1) * 4 + 2 * ks + (i >> 1)) ^ r8) << 4));
        else cl_ldsm_x4(qa[ks], tile_addr + row * 64 + ((((i & 1) * 2 + (i >> 1)) ^ ((row >> 1) & 3)) << 4));
    }
}
// byte offset of chunk j of token row `tok` in such a tile
template <int C> __device__ __forceinline__ uint32_t cl_tile_off(int tok, int j) {
    return C == 32 ? uint32_t(tok * 128 + ((j ^ (tok & 7)) << 4)) : uint32_t(tok * 64 + ((j ^ ((tok >> 1) & 3)) << 4));
}

}  // namespace affb200
