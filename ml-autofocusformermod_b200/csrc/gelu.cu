// gelu.cu — the exact (erf) GELU of the Mlp between fc1 and fc2 (reference models/aff_transformer.py:35-51: nn.GELU)
// for 2-byte activations.  On B200 the stock element-wise kernels are bound by the fp32 erf/exp arithmetic (about 30
// issue slots per element forward, 45 backward), not by HBM.  A bf16/f16 input has only 65 536 distinct values, so
// both passes become table lookups in shared memory:
//   forward   y  = F[bits(x)]                      F: 65 536 entries of T, generated with the stock formula, so the
//                                                  result is bit-identical to computing erf per element
//   backward  dx = dy · G[bits(|x|)]  (x >= 0),  dy · (1 − G[bits(|x|)])  (x < 0)
//                                                  G: 32 768 fp32 entries of Phi(x) + x·phi(x); the identity
//                                                  G(−x) = 1 − G(x) halves the table to the 128 KB one SM can hold
// and optionally the column sums of dx (the bias gradient of fc1) are accumulated in the same pass.
// One CTA of 1024 threads per SM; the 128 KB table is pulled from L2 once per CTA.
#include "common.cuh"
#include <algorithm>
#include <type_traits>

namespace affb200 {

constexpr int kGT = 1024;
constexpr int kGeluTabBytes = 128 * 1024;

int launch_col_sum_finish(const float *part, int n_slabs, int C, float *out, int accumulate, cudaStream_t st);

template <typename T> __device__ __forceinline__ T g_from_bits(uint32_t b) {
    const uint16_t s = uint16_t(b);
    return *reinterpret_cast<const T *>(&s);
}
template <typename T> __device__ __forceinline__ uint32_t g_bits(T v) { return *reinterpret_cast<const uint16_t *>(&v); }

// stock formulas (ATen GeluCUDAKernelImpl / GeluBackwardCUDAKernelImpl, approximate='none'), fp32 arithmetic
template <typename T>
__global__ void __launch_bounds__(256)
gelu_tables_kernel(uint16_t *__restrict__ fwd_tab, float *__restrict__ bwd_tab) {
    const uint32_t code = blockIdx.x * 256 + threadIdx.x;  // < 65536
    const float x = Elem<T>::to_acc(g_from_bits<T>(code));
    constexpr float kAlpha = 0.70710678118654752440f;
    fwd_tab[code] = uint16_t(g_bits<T>(Elem<T>::from_acc(x * 0.5f * (1.f + erff(x * kAlpha)))));
    if (code < 32768u) {
        constexpr float kBeta = 1.12837916709551257390f * 0.70710678118654752440f * 0.5f;
        const float cdf = 0.5f * (1.f + erff(x * kAlpha));
        const float pdf = expf(-0.5f * x * x) * kBeta;
        bwd_tab[code] = cdf + x * pdf;
    }
}

template <typename T>
__global__ void __launch_bounds__(kGT, 1)
gelu_fwd_kernel(const T *__restrict__ x, T *__restrict__ y, int64_t n8, const uint16_t *__restrict__ tab_g) {
    extern __shared__ __align__(16) uint16_t tab[];
    for (int i = threadIdx.x; i < kGeluTabBytes / 16; i += kGT)
        reinterpret_cast<uint4 *>(tab)[i] = __ldg(reinterpret_cast<const uint4 *>(tab_g) + i);
    __syncthreads();
    const int64_t stride = int64_t(gridDim.x) * kGT;
#pragma unroll 2
    for (int64_t g = int64_t(blockIdx.x) * kGT + threadIdx.x; g < n8; g += stride) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(x) + g);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) o[k] = uint32_t(tab[w[k] & 0xffffu]) | (uint32_t(tab[w[k] >> 16]) << 16);
        reinterpret_cast<uint4 *>(y)[g] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

// dx = dy · GELU'(x); COLS: also part[blockIdx.x][C] = column sums of the (rounded) dx over this CTA's rows.
// Thread -> (row lane, 8-column chunk): chunks = C/8 consecutive threads cover one row.
template <typename T, bool COLS>
__global__ void __launch_bounds__(kGT, 1)
gelu_bwd_kernel(const T *__restrict__ dy, const T *__restrict__ x, T *__restrict__ dx, int64_t R, int C,
                const float *__restrict__ tab_g, float *__restrict__ part) {
    extern __shared__ __align__(16) float tabf[];
    for (int i = threadIdx.x; i < kGeluTabBytes / 16; i += kGT)
        reinterpret_cast<uint4 *>(tabf)[i] = __ldg(reinterpret_cast<const uint4 *>(tab_g) + i);
    __syncthreads();
    const int chunks = C >> 3;
    const int rpp = COLS ? kGT / chunks : 0;           // rows per pass of the CTA
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    auto one = [&](int64_t g) {
        const uint4 xv = __ldg(reinterpret_cast<const uint4 *>(x) + g);
        const uint4 dv = __ldg(reinterpret_cast<const uint4 *>(dy) + g);
        const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w}, dw[4] = {dv.x, dv.y, dv.z, dv.w};
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t x0 = xw[k] & 0xffffu, x1 = xw[k] >> 16;
            float g0 = tabf[x0 & 0x7fffu], g1 = tabf[x1 & 0x7fffu];
            if (x0 & 0x8000u) g0 = 1.f - g0;
            if (x1 & 0x8000u) g1 = 1.f - g1;
            const T r0 = Elem<T>::from_acc(Elem<T>::to_acc(g_from_bits<T>(dw[k] & 0xffffu)) * g0);
            const T r1 = Elem<T>::from_acc(Elem<T>::to_acc(g_from_bits<T>(dw[k] >> 16)) * g1);
            if (COLS) { acc[2 * k] += Elem<T>::to_acc(r0); acc[2 * k + 1] += Elem<T>::to_acc(r1); }
            o[k] = g_bits<T>(r0) | (g_bits<T>(r1) << 16);
        }
        reinterpret_cast<uint4 *>(dx)[g] = make_uint4(o[0], o[1], o[2], o[3]);
    };
    if constexpr (COLS) {
        const int rl = threadIdx.x / chunks, ch = threadIdx.x - rl * chunks;
        if (rl < rpp) {
            const int64_t stride = int64_t(gridDim.x) * rpp;
#pragma unroll 2
            for (int64_t r = int64_t(blockIdx.x) * rpp + rl; r < R; r += stride) one(r * chunks + ch);
        }
        __syncthreads();                       // the table is dead: its storage takes the per-row-lane partials
        float *red = tabf;                     // [rpp][C]
        if (rl < rpp) {
#pragma unroll
            for (int e = 0; e < 8; ++e) red[rl * C + ch * 8 + e] = acc[e];
        }
        __syncthreads();
        for (int c = threadIdx.x; c < C; c += kGT) {
            float a = 0.f;
            for (int k = 0; k < rpp; ++k) a += red[k * C + c];
            part[int64_t(blockIdx.x) * C + c] = a;
        }
    } else {
        const int64_t n8 = R * chunks, stride = int64_t(gridDim.x) * kGT;
#pragma unroll 2
        for (int64_t g = int64_t(blockIdx.x) * kGT + threadIdx.x; g < n8; g += stride) one(g);
    }
}

static int gelu_grid(int64_t n8) { return int(std::max<int64_t>(1, std::min<int64_t>(kNumSMs, (n8 + kGT - 1) / kGT))); }

}  // namespace affb200

using namespace affb200;

// fwd_tab: 65536 x 2 bytes, bwd_tab: 32768 x 4 bytes (both 128 KB, 16-byte aligned), filled for `dtype` (f16 / bf16)
extern "C" int affb200_gelu_tables(int dtype, void *fwd_tab, float *bwd_tab, void *stream) {
    if (!fwd_tab || !bwd_tab || !aligned16(fwd_tab) || !aligned16(bwd_tab)) { set_error("gelu_tables: null or misaligned pointer"); return AFFB200_E_ARG; }
    auto st = (cudaStream_t)stream;
    if (dtype == AFFB200_BF16) gelu_tables_kernel<__nv_bfloat16><<<256, 256, 0, st>>>((uint16_t *)fwd_tab, bwd_tab);
    else if (dtype == AFFB200_F16) gelu_tables_kernel<__half><<<256, 256, 0, st>>>((uint16_t *)fwd_tab, bwd_tab);
    else { set_error("gelu_tables: f16 / bf16 only"); return AFFB200_E_ARG; }
    return check_launch("gelu_tables");
}

extern "C" int affb200_gelu_fwd(const void *x, void *y, int64_t n, int dtype, const void *fwd_tab, void *stream) {
    if (n < 0 || n % 8 != 0) { set_error("gelu_fwd: element count must be a multiple of 8"); return AFFB200_E_ARG; }
    if (n == 0) return AFFB200_OK;
    if (!x || !y || !fwd_tab || !aligned16(x) || !aligned16(y) || !aligned16(fwd_tab)) { set_error("gelu_fwd: null or misaligned pointer"); return AFFB200_E_ARG; }
    if (dtype != AFFB200_BF16 && dtype != AFFB200_F16) { set_error("gelu_fwd: f16 / bf16 only"); return AFFB200_E_ARG; }
    auto st = (cudaStream_t)stream;
    auto kern = gelu_fwd_kernel<__nv_bfloat16>;   // a pure bit-pattern lookup: one instantiation serves both types
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kGeluTabBytes) != cudaSuccess)
        return check_launch("cudaFuncSetAttribute", 0);
    kern<<<gelu_grid(n / 8), kGT, kGeluTabBytes, st>>>((const __nv_bfloat16 *)x, (__nv_bfloat16 *)y, n / 8,
                                                       (const uint16_t *)fwd_tab);
    return check_launch("gelu_fwd");
}

// dx = dy·GELU'(x) over a row-major [R][C] matrix; col_sums != NULL: col_sums[c] (+)= sum_r dx[r][c] (fp32; needs
// C <= 8192 and work = affb200_gelu_bwd_work_floats(C) floats)
extern "C" int64_t affb200_gelu_bwd_work_floats(int64_t C) { return C > 0 ? int64_t(kNumSMs) * C : 0; }

extern "C" int affb200_gelu_bwd(const void *dy, const void *x, void *dx, int64_t R, int64_t C, int dtype,
                                const float *bwd_tab, float *col_sums, int accumulate, float *work, void *stream) {
    if (R < 0 || C <= 0 || C % 8 != 0 || C > 8192) { set_error("gelu_bwd: C must be a multiple of 8, at most 8192"); return AFFB200_E_ARG; }
    if (R == 0) return AFFB200_OK;
    if (!dy || !x || !dx || !bwd_tab || !aligned16(dy) || !aligned16(x) || !aligned16(dx) || !aligned16(bwd_tab) ||
        (col_sums && !work)) {
        set_error("gelu_bwd: null or misaligned pointer");
        return AFFB200_E_ARG;
    }
    auto st = (cudaStream_t)stream;
    const int64_t n8 = R * (C / 8);
    return dispatch_dtype(dtype, [&](auto tag) -> int {
        using T0 = decltype(tag);
        if constexpr (std::is_same<T0, __half>::value || std::is_same<T0, __nv_bfloat16>::value) {
            using T = T0;
            if (col_sums) {
                auto kern = gelu_bwd_kernel<T, true>;
                if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kGeluTabBytes) != cudaSuccess)
                    return check_launch("cudaFuncSetAttribute", 0);
                const int rpp = kGT / int(C / 8);
                const int grid = int(std::max<int64_t>(1, std::min<int64_t>(kNumSMs, (R + rpp - 1) / rpp)));
                kern<<<grid, kGT, kGeluTabBytes, st>>>((const T *)dy, (const T *)x, (T *)dx, R, int(C), bwd_tab, work);
                if (int e = check_launch("gelu_bwd")) return e;
                return launch_col_sum_finish(work, grid, int(C), col_sums, accumulate, st);
            }
            auto kern = gelu_bwd_kernel<T, false>;
            if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kGeluTabBytes) != cudaSuccess)
                return check_launch("cudaFuncSetAttribute", 0);
            kern<<<gelu_grid(n8), kGT, kGeluTabBytes, st>>>((const T *)dy, (const T *)x, (T *)dx, R, int(C), bwd_tab, nullptr);
            return check_launch("gelu_bwd");
        } else {
            set_error("gelu_bwd: f16 / bf16 only");
            return AFFB200_E_ARG;
        }
    });
}
