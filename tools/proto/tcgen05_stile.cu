// tcgen05_stile.cu — measured prototype asked for by the round-1 review: the score tile of dense masked attention over a
// cluster union, S = Q·K^T for 128 tokens x 128 keys x head dim 32 (the stage-3 tile of AFF-Tiny-1/5: one image-head),
// produced (a) by tcgen05.mma into TMEM (one elected thread, M128 N128 K16 x 2, operands in TMA-written 64-byte-swizzled
// shared-memory tiles through UMMA descriptors) and read back lane-per-row with tcgen05.ld for the softmax statistics,
// and (b) by mma.sync m16n8k16 from ldmatrix on THE SAME staged tiles with the statistics in registers (what
// csrc/fused_cld.cu does).  Both write the row maximum and the row sum of exp2(s - max) of every row; the program
// checks them against a CPU evaluation and times both kernels over the same items with CUDA events.
// Stand-alone:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o tcgen05_stile tcgen05_stile.cu
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int TM = 128, TN = 128, TK = 32;       // tokens, keys, head dim
constexpr int TILE_BYTES = TM * TK * 2;          // 8 KB per operand tile

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
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
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
// 16-byte chunk j of row r of a [rows][64-byte] tile written by TMA with the 64-byte swizzle
__device__ __forceinline__ uint32_t sw64(int r, int j) { return uint32_t(r * 64 + ((j ^ ((r >> 1) & 3)) << 4)); }

// ---- (a) tcgen05 ---------------------------------------------------------------------------------------------------
// shared-memory matrix descriptor, K-major operand, 64-byte swizzle: rows of 64 bytes, 8-row groups 512 bytes apart
__device__ __forceinline__ uint64_t umma_desc_k_sw64(uint32_t saddr) {
    uint64_t d = 0;
    d |= uint64_t((saddr & 0x3ffffu) >> 4);          // start address, 16-byte units
    d |= uint64_t(1) << 16;                          // leading byte offset: unused for swizzled K-major (1)
    d |= uint64_t(512 >> 4) << 32;                   // stride byte offset between 8-row groups
    d |= uint64_t(1) << 46;                          // descriptor version (sm_100)
    d |= uint64_t(4) << 61;                          // layout type: SWIZZLE_64B
    return d;
}
// instruction descriptor: D fp32, A/B bf16, both K-major, N = 128, M = 128
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | (uint32_t(TN >> 3) << 17) | (uint32_t(TM >> 4) << 24);

__global__ void __launch_bounds__(128)
stile_tcgen05_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk, float scale_log2,
                     float *__restrict__ row_max, float *__restrict__ row_sum) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) uint64_t bars[2];
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, item = blockIdx.x;
    const uint32_t sq = smem_u32(smem), sk = sq + TILE_BYTES;
    const uint32_t bar_ld = smem_u32(&bars[0]), bar_mma = smem_u32(&bars[1]);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(128u));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    if (tid == 0) {
        mbar_init(bar_ld, 1);
        mbar_init(bar_mma, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    if (tid == 0) {
        mbar_expect_tx(bar_ld, 2 * TILE_BYTES);
        tma_load_2d(sq, &tmq, 0, item * TM, bar_ld);
        tma_load_2d(sk, &tmk, 0, item * TN, bar_ld);
    }
    mbar_wait(bar_ld, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (tid == 0) {
        const uint64_t da = umma_desc_k_sw64(sq), db = umma_desc_k_sw64(sk);
#pragma unroll
        for (int k = 0; k < TK / 16; ++k) {
            const uint32_t acc = k > 0 ? 1u : 0u;
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                ::"r"(tmem), "l"(da + uint64_t(k * 2)), "l"(db + uint64_t(k * 2)), "r"(kIdesc), "r"(acc) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_mma) : "memory");
    }
    mbar_wait(bar_mma, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // lane-per-row statistics: thread = row 32*warp + lane, 4 loads of 32 columns, two passes (max, then sum)
    const uint32_t taddr = tmem + (uint32_t(warp * 32) << 16);
    float mx = -INFINITY, sum = 0.f;
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
#pragma unroll
        for (int c0 = 0; c0 < TN; c0 += 32) {
            uint32_t r[32];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                         "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                         "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                           "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                           "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                           "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                         : "r"(taddr + uint32_t(c0)));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const float v = __uint_as_float(r[j]) * scale_log2;
                if (pass == 0) mx = fmaxf(mx, v);
                else sum += exp2f(v - mx);
            }
        }
    }
    row_max[int64_t(item) * TM + tid] = mx;
    row_sum[int64_t(item) * TM + tid] = sum;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u));
}

// ---- (b) mma.sync on the same staged tiles ---------------------------------------------------------------------------
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(256)
stile_mma_sync_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk, float scale_log2,
                      float *__restrict__ row_max, float *__restrict__ row_sum) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) uint64_t bars[1];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, item = blockIdx.x;
    const uint32_t sq = smem_u32(smem), sk = sq + TILE_BYTES, bar_ld = smem_u32(&bars[0]);
    if (tid == 0) {
        mbar_init(bar_ld, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
        mbar_expect_tx(bar_ld, 2 * TILE_BYTES);
        tma_load_2d(sq, &tmq, 0, item * TM, bar_ld);
        tma_load_2d(sk, &tmk, 0, item * TN, bar_ld);
    }
    mbar_wait(bar_ld, 0);
    // warp: 16 rows; A fragments of both k-steps
    uint32_t qa[2][4];
    {
        const int row = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, cj = lane >> 4;
#pragma unroll
        for (int ks = 0; ks < 2; ++ks) ldsm_x4(qa[ks], sq + sw64(row, ks * 2 + cj));
    }
    float s[16][4];
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nb = 0; nb < 16; ++nb) {
        uint32_t kf[4];
        ldsm_x4(kf, sk + sw64(nb * 8 + (lane & 7), lane >> 3));
        s[nb][0] = s[nb][1] = s[nb][2] = s[nb][3] = 0.f;
        mma_bf16(s[nb], qa[0], kf[0], kf[1]);
        mma_bf16(s[nb], qa[1], kf[2], kf[3]);
#pragma unroll
        for (int e = 0; e < 4; ++e) s[nb][e] *= scale_log2;
        mx[0] = fmaxf(mx[0], fmaxf(s[nb][0], s[nb][1]));
        mx[1] = fmaxf(mx[1], fmaxf(s[nb][2], s[nb][3]));
    }
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
        mx[hh] = fmaxf(mx[hh], __shfl_xor_sync(0xffffffffu, mx[hh], 1));
        mx[hh] = fmaxf(mx[hh], __shfl_xor_sync(0xffffffffu, mx[hh], 2));
    }
    float sum[2] = {0.f, 0.f};
#pragma unroll
    for (int nb = 0; nb < 16; ++nb) {
        sum[0] += exp2f(s[nb][0] - mx[0]) + exp2f(s[nb][1] - mx[0]);
        sum[1] += exp2f(s[nb][2] - mx[1]) + exp2f(s[nb][3] - mx[1]);
    }
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
        sum[hh] += __shfl_xor_sync(0xffffffffu, sum[hh], 1);
        sum[hh] += __shfl_xor_sync(0xffffffffu, sum[hh], 2);
    }
    if ((lane & 3) == 0) {
        const int64_t r0 = int64_t(item) * TM + warp * 16 + g;
        row_max[r0] = mx[0]; row_sum[r0] = sum[0];
        row_max[r0 + 8] = mx[1]; row_sum[r0 + 8] = sum[1];
    }
}

// ---- host ------------------------------------------------------------------------------------------------------------
static bool make_map(CUtensorMap *map, const void *base, int64_t rows) {
    typedef CUresult (*Fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                           const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                           CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void *pfn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &pfn, cudaEnableDefault, &q) != cudaSuccess || !pfn) return false;
    const cuuint64_t dims[2] = {cuuint64_t(TK), cuuint64_t(rows)};
    const cuuint64_t strides[1] = {cuuint64_t(TK) * 2u};
    const cuuint32_t box[2] = {cuuint32_t(TK), cuuint32_t(TM)};
    const cuuint32_t estr[2] = {1u, 1u};
    return reinterpret_cast<Fn>(pfn)(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
                                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int main(int argc, char **argv) {
    const int items = argc > 1 ? atoi(argv[1]) : 2048;      // image-heads (stage 3 of AFF-Tiny-1/5 at batch 256: 2048)
    const int iters = argc > 2 ? atoi(argv[2]) : 50;
    const size_t n = size_t(items) * TM * TK;
    std::vector<__nv_bfloat16> hq(n), hk(n);
    uint32_t seed = 12345u;
    auto rnd = [&]() { seed = seed * 1664525u + 1013904223u; return float(int(seed >> 9) % 2001 - 1000) / 1000.f; };
    for (size_t i = 0; i < n; ++i) { hq[i] = __float2bfloat16(rnd()); hk[i] = __float2bfloat16(rnd()); }
    __nv_bfloat16 *dq, *dk;
    float *mx[2], *sm[2];
    CK(cudaMalloc(&dq, n * 2)); CK(cudaMalloc(&dk, n * 2));
    for (int v = 0; v < 2; ++v) { CK(cudaMalloc(&mx[v], size_t(items) * TM * 4)); CK(cudaMalloc(&sm[v], size_t(items) * TM * 4)); }
    CK(cudaMemcpy(dq, hq.data(), n * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dk, hk.data(), n * 2, cudaMemcpyHostToDevice));
    CUtensorMap tmq, tmk;
    if (!make_map(&tmq, dq, int64_t(items) * TM) || !make_map(&tmk, dk, int64_t(items) * TN)) { printf("tensor map failed\n"); return 1; }
    const float scale_log2 = (1.f / sqrtf(float(TK))) * 1.4426950408889634f;
    const int smem = 2 * TILE_BYTES;
    auto run = [&](int v) {
        if (v == 0) stile_tcgen05_kernel<<<items, 128, smem>>>(tmq, tmk, scale_log2, mx[0], sm[0]);
        else stile_mma_sync_kernel<<<items, 256, smem>>>(tmq, tmk, scale_log2, mx[1], sm[1]);
    };
    float ms[2];
    for (int v = 0; v < 2; ++v) {
        for (int i = 0; i < 3; ++i) run(v);
        CK(cudaDeviceSynchronize());
        cudaEvent_t a, b;
        CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
        CK(cudaEventRecord(a));
        for (int i = 0; i < iters; ++i) run(v);
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        CK(cudaEventElapsedTime(&ms[v], a, b));
        ms[v] /= iters;
    }
    // check both against the CPU on a few items
    std::vector<float> hm[2], hs[2];
    for (int v = 0; v < 2; ++v) {
        hm[v].resize(size_t(items) * TM); hs[v].resize(size_t(items) * TM);
        CK(cudaMemcpy(hm[v].data(), mx[v], hm[v].size() * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(hs[v].data(), sm[v], hs[v].size() * 4, cudaMemcpyDeviceToHost));
    }
    double err[2] = {0, 0};
    const int check_items[3] = {0, items / 2, items - 1};
    for (int ci = 0; ci < 3; ++ci) {
        const int it = check_items[ci];
        for (int r = 0; r < TM; ++r) {
            double m = -1e30, srow[TN];
            for (int c = 0; c < TN; ++c) {
                double acc = 0;
                for (int k = 0; k < TK; ++k)
                    acc += double(__bfloat162float(hq[(size_t(it) * TM + r) * TK + k])) * double(__bfloat162float(hk[(size_t(it) * TN + c) * TK + k]));
                srow[c] = acc * scale_log2;
                m = fmax(m, srow[c]);
            }
            double s = 0;
            for (int c = 0; c < TN; ++c) s += exp2(srow[c] - m);
            for (int v = 0; v < 2; ++v) {
                err[v] = fmax(err[v], fabs(hm[v][size_t(it) * TM + r] - m));
                err[v] = fmax(err[v], fabs(hs[v][size_t(it) * TM + r] - s) / s);
            }
        }
    }
    const double flop = 2.0 * items * TM * TN * TK;
    printf("{\"items\": %d, \"tile\": \"128x128x32 bf16\", \"tcgen05_us\": %.2f, \"mma_sync_us\": %.2f, \"tcgen05_max_err\": %.3g, "
           "\"mma_sync_max_err\": %.3g, \"tcgen05_tflops\": %.1f, \"mma_sync_tflops\": %.1f, \"operand_gbs_tcgen05\": %.0f, "
           "\"operand_gbs_mma_sync\": %.0f}\n",
           items, ms[0] * 1e3, ms[1] * 1e3, err[0], err[1], flop / ms[0] / 1e9, flop / ms[1] / 1e9,
           2.0 * n * 2 / ms[0] / 1e6, 2.0 * n * 2 / ms[1] / 1e6);
    return (err[0] < 1e-3 && err[1] < 1e-3) ? 0 : 2;
}
