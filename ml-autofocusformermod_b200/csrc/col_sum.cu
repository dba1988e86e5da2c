// col_sum.cu — column sums of a row-major [R][C] matrix in fp32: the bias gradient of every Linear around the
// attention (q, kv, proj, fc1, fc2, the merge Linear; reference aff_transformer.py:85-88,96-100,285), which autograd
// computes with a generic strided reduction (≈ 10 % of the bf16 training step at 224²).  One pass at 16-byte loads:
// a CTA owns a block of 256 columns (32 lanes x 8) or the whole row when C is smaller, walks a slab of rows, and
// leaves fp32 partials [slab][C]; a second kernel adds the slabs in a fixed order (no atomics: bit-reproducible).
#include "common.cuh"
#include <algorithm>
#include <atomic>

namespace affb200 {

constexpr int kCT = 256;

template <typename T> __device__ __forceinline__ void cs_load8(const T *p, float (&o)[8]) {
    if constexpr (sizeof(T) == 4) {
        const float4 a = __ldg(reinterpret_cast<const float4 *>(p)), b = __ldg(reinterpret_cast<const float4 *>(p) + 1);
        o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
    } else {
        const uint4 r = __ldg(reinterpret_cast<const uint4 *>(p));
        const T *e = reinterpret_cast<const T *>(&r);
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = Elem<T>::to_acc(e[k]);
    }
}

// grid (slabs, column blocks).  LPR lanes cover the 8*LPR columns of the block; kCT/LPR rows per pass.
template <typename T, int LPR>
__global__ void __launch_bounds__(kCT)
col_sum_kernel(const T *__restrict__ x, float *__restrict__ part, int64_t R, int C, int64_t rows_per_slab) {
    constexpr int RPC = kCT / LPR;
    __shared__ float red[kCT / 32][LPR * 8];
    const int l = threadIdx.x % LPR, rl = threadIdx.x / LPR, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c0 = (blockIdx.y * LPR + l) * 8;
    const bool col_ok = c0 < C;
    const int64_t r_lo = int64_t(blockIdx.x) * rows_per_slab, r_hi = min(R, r_lo + rows_per_slab);
    float acc[8], acc2[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = acc2[e] = 0.f;
    int64_t r = r_lo + rl;
    for (; r + RPC < r_hi; r += 2 * RPC) {  // two independent loads in flight
        float a[8], b[8];
        if (col_ok) {
            cs_load8<T>(x + r * C + c0, a);
            cs_load8<T>(x + (r + RPC) * C + c0, b);
#pragma unroll
            for (int e = 0; e < 8; ++e) { acc[e] += a[e]; acc2[e] += b[e]; }
        }
    }
    if (r < r_hi && col_ok) {
        float a[8];
        cs_load8<T>(x + r * C + c0, a);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += a[e];
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        float a = acc[e] + acc2[e];
#pragma unroll
        for (int o = LPR; o < 32; o <<= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (lane < LPR) red[warp][l * 8 + e] = a;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < LPR * 8; c += kCT) {
        float a = 0.f;
#pragma unroll
        for (int w = 0; w < kCT / 32; ++w) a += red[w][c];
        const int col = blockIdx.y * LPR * 8 + c;
        if (col < C) part[int64_t(blockIdx.x) * C + col] = a;
    }
}

// out[c] = sum over slabs: a block owns 32 columns, 32 threads per column stride over the slabs (coalesced 128-byte
// rows of partials), then a fixed-order tree over the 32 partial sums
__global__ void __launch_bounds__(1024)
col_sum_finish_kernel(const float *__restrict__ part, int n_slabs, int C, float *__restrict__ out, int accumulate) {
    __shared__ float red[32][33];
    const int cl = threadIdx.x & 31, sub = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + cl;
    float a = 0.f;
    if (c < C)
        for (int s = sub; s < n_slabs; s += 32) a += part[int64_t(s) * C + c];
    red[sub][cl] = a;
    __syncthreads();
    if (sub == 0 && c < C) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 32; ++k) t += red[k][cl];
        out[c] = accumulate ? out[c] + t : t;
    }
}

int launch_col_sum_finish(const float *part, int n_slabs, int C, float *out, int accumulate, cudaStream_t st) {
    col_sum_finish_kernel<<<unsigned((C + 31) / 32), 1024, 0, st>>>(part, n_slabs, C, out, accumulate);
    return check_launch("col_sum_finish");
}

static void col_sum_geometry(int64_t R, int64_t C, int *lpr, int *n_cblk, int *n_slabs, int64_t *rows_per_slab) {
    const int nch = int(C / 8);
    *lpr = nch <= 8 ? 8 : (nch <= 16 ? 16 : 32);
    *n_cblk = (nch + *lpr - 1) / *lpr;
    const int rpc = kCT / *lpr;
    int64_t slabs = std::max<int64_t>(1, (int64_t(kNumSMs) * 4) / *n_cblk);
    slabs = std::min<int64_t>(slabs, ceil_div64(R, int64_t(rpc) * 4));
    slabs = std::max<int64_t>(slabs, 1);
    *rows_per_slab = ceil_div64(ceil_div64(R, slabs), rpc) * rpc;
    *n_slabs = int(ceil_div64(R, *rows_per_slab));
}

}  // namespace affb200

using namespace affb200;

extern "C" int64_t affb200_col_sum_work_floats(int64_t R, int64_t C) {
    if (R <= 0 || C <= 0 || C % 8 != 0) return 0;
    int lpr, n_cblk, n_slabs;
    int64_t rps;
    col_sum_geometry(R, C, &lpr, &n_cblk, &n_slabs, &rps);
    return int64_t(n_slabs) * C;
}

extern "C" int affb200_col_sum(const void *x, float *out, float *work, int64_t R, int64_t C, int dtype, int accumulate,
                               void *stream) {
    if (R < 0 || C <= 0 || C % 8 != 0 || C >= (int64_t(1) << 24)) { set_error("col_sum: C must be a positive multiple of 8"); return AFFB200_E_ARG; }
    auto st = (cudaStream_t)stream;
    if (!out) { set_error("col_sum: null pointer"); return AFFB200_E_ARG; }
    if (R == 0) {
        if (accumulate) return AFFB200_OK;
        cudaMemsetAsync(out, 0, size_t(C) * 4, st);
        return check_launch("col_sum memset", 0);
    }
    if (!x || !work || !aligned16(x)) { set_error("col_sum: null or misaligned pointer"); return AFFB200_E_ARG; }
    int lpr, n_cblk, n_slabs;
    int64_t rps;
    col_sum_geometry(R, C, &lpr, &n_cblk, &n_slabs, &rps);
    const dim3 grid{unsigned(n_slabs), unsigned(n_cblk), 1u};
    int e = dispatch_dtype(dtype, [&](auto tag) -> int {
        using T = decltype(tag);
        if constexpr (sizeof(T) == 8) {
            set_error("col_sum: f32 / f16 / bf16 only");
            return AFFB200_E_ARG;
        } else {
            if (lpr == 8) col_sum_kernel<T, 8><<<grid, kCT, 0, st>>>((const T *)x, work, R, int(C), rps);
            else if (lpr == 16) col_sum_kernel<T, 16><<<grid, kCT, 0, st>>>((const T *)x, work, R, int(C), rps);
            else col_sum_kernel<T, 32><<<grid, kCT, 0, st>>>((const T *)x, work, R, int(C), rps);
            return check_launch("col_sum");
        }
    });
    if (e) return e;
    return launch_col_sum_finish(work, n_slabs, int(C), out, accumulate, st);
}

// ---- allocation-free variant --------------------------------------------------------------------------------
// affb200_col_sum_static: the same two kernels with the slab partials in a file-scope device array instead of a
// caller-provided workspace, for callers that cannot allocate (CUDA graph capture, serving loops).  No counterpart
// in the reference.  The array (1 MiB) is part of the module image: every GPU that loads the library holds it until
// the process exits.  In host code `g_static_part` names a host-side stand-in, not device memory, so its device
// address is looked up once per device with cudaGetSymbolAddress while that device is current.
namespace affb200 {

constexpr int64_t kStaticWorkFloats = AFFB200_COL_SUM_STATIC_MAX_FLOATS;
static __device__ __align__(256) float g_static_part[kStaticWorkFloats];

constexpr int kMaxDevices = 64;
static std::atomic<float *> g_static_ptr[kMaxDevices];

static float *static_part_address() {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        set_error("col_sum_static: cudaGetDevice failed or device ordinal %d unsupported (%s)", dev,
                  cudaGetErrorString(e));
        return nullptr;
    }
    float *ptr = g_static_ptr[dev].load(std::memory_order_acquire);
    if (ptr) return ptr;
    void *sym = nullptr;
    e = cudaGetSymbolAddress(&sym, g_static_part);
    if (e != cudaSuccess || !sym) {
        set_error("col_sum_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        return nullptr;
    }
    g_static_ptr[dev].store(static_cast<float *>(sym), std::memory_order_release);
    return static_cast<float *>(sym);
}

}  // namespace affb200

extern "C" int64_t affb200_col_sum_static_max_floats(void) { return kStaticWorkFloats; }

extern "C" int affb200_col_sum_static(const void *x, float *out, int64_t R, int64_t C, int dtype, int accumulate,
                                      void *stream) {
    if (R <= 0 || C <= 0 || C % 8 != 0 || C >= (int64_t(1) << 24)) {
        set_error("col_sum_static: R must be positive and C a positive multiple of 8");
        return AFFB200_E_ARG;
    }
    if (!x || !out || !aligned16(x)) { set_error("col_sum_static: null or misaligned pointer"); return AFFB200_E_ARG; }
    int lpr, n_cblk, n_slabs;
    int64_t rps;
    col_sum_geometry(R, C, &lpr, &n_cblk, &n_slabs, &rps);
    const int64_t need = int64_t(n_slabs) * C;
    if (need > kStaticWorkFloats) {
        set_error("col_sum_static: [%lld][%lld] needs %lld floats of scratch, the built-in array holds %lld; use "
                  "affb200_col_sum with a caller workspace", (long long)R, (long long)C, (long long)need,
                  (long long)kStaticWorkFloats);
        return AFFB200_E_WORKSPACE;
    }
    if (dtype != AFFB200_F32 && dtype != AFFB200_F16 && dtype != AFFB200_BF16) {
        set_error("col_sum_static: f32 / f16 / bf16 only");
        return AFFB200_E_ARG;
    }
    float *part = static_part_address();
    if (!part) return AFFB200_E_CUDA;
    auto st = (cudaStream_t)stream;
    const dim3 grid{unsigned(n_slabs), unsigned(n_cblk), 1u};
    // every partial [slab][col] that the finish kernel reads is written by the first kernel of this call: nothing in
    // the shared array has to start at zero, whatever an earlier call left there
    int e = dispatch_dtype(dtype, [&](auto tag) -> int {
        using T = decltype(tag);
        if constexpr (sizeof(T) == 8) {
            return AFFB200_E_ARG;
        } else {
            if (lpr == 8) col_sum_kernel<T, 8><<<grid, kCT, 0, st>>>((const T *)x, part, R, int(C), rps);
            else if (lpr == 16) col_sum_kernel<T, 16><<<grid, kCT, 0, st>>>((const T *)x, part, R, int(C), rps);
            else col_sum_kernel<T, 32><<<grid, kCT, 0, st>>>((const T *)x, part, R, int(C), rps);
            return check_launch("col_sum_static");
        }
    });
    if (e) return e;
    return launch_col_sum_finish(part, n_slabs, int(C), out, accumulate, st);
}
