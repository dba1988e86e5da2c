// batch_norm.cu — training-mode BatchNorm2d followed by the exact (erf) GELU over channels-last activations: the
// `bn` + `act1` pair of the stem that feeds the off-grid stages (reference models/aff_transformer.py:524-527, 541).
// The tensor is the largest activation of the network ([B, 112, 112, C/2]); the stock channels-last kernels walk it
// with a handful of CTAs per channel group.  Here every pass streams rows with 16-byte loads from all SMs:
//
//   forward   bn_stats (per-CTA Welford partials) -> bn_stats_finish (mean, 1/std, running statistics)
//             -> bn_apply (normalise, affine, GELU; one read, one write)
//   backward  bn_bwd_reduce (sum g, sum g·(x−mean), g = dy·GELU'(z) recomputed from x) -> bn_bwd_finish
//             -> bn_bwd_elemt (dx)
//
// Numerics follow the stock operators step by step so that the fused op rounds where they round: z = BN(x) is
// rounded to the storage type before the GELU, g is rounded to the storage type before the BatchNorm backward.
// Partials are combined in a fixed order (no atomics: bit-reproducible).
//
// Mapping: LPR = C/8 lanes share a row, lane l owns channels [8l, 8l+8); a CTA pass covers 256/LPR rows.
#include "common.cuh"
#include <algorithm>
#include <type_traits>

namespace affb200 {

constexpr int kBT = 256;

struct Wf3 {
    float n, mean, m2;
};
// Chan et al. pairwise combination of (count, mean, centred second moment)
__device__ __forceinline__ Wf3 wf_merge(Wf3 a, Wf3 b) {
    const float n = a.n + b.n;
    if (n == 0.f) return a;
    const float d = b.mean - a.mean, f = b.n / n;
    return Wf3{n, a.mean + d * f, a.m2 + b.m2 + d * d * a.n * f};
}

__device__ __forceinline__ float bn_gelu(float x) {
    return x * 0.5f * (1.f + erff(x * 0.70710678118654752440f));
}
__device__ __forceinline__ float bn_gelu_grad(float x, float dy) {
    constexpr float kBeta = 1.12837916709551257390f * 0.70710678118654752440f * 0.5f;
    const float cdf = 0.5f * (1.f + erff(x * 0.70710678118654752440f));
    const float pdf = expf(-0.5f * x * x) * kBeta;
    return dy * (cdf + x * pdf);
}
template <typename T> __device__ __forceinline__ float bn_round(float x) { return Elem<T>::to_acc(Elem<T>::from_acc(x)); }

// per-CTA partial statistics: part[blk][3][C] = (count, mean, M2) of the CTA's rows, per channel
template <typename T, int LPR>
__global__ void __launch_bounds__(kBT)
bn_stats_kernel(const T *__restrict__ x, float *__restrict__ part, int64_t R) {
    constexpr int RPC = kBT / LPR, C = LPR * 8;
    __shared__ float red[8][3][C];
    const int l = threadIdx.x % LPR, rl = threadIdx.x / LPR, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t stride = int64_t(gridDim.x) * RPC;
    int64_t r = int64_t(blockIdx.x) * RPC + rl;
    // sums of (x − K) with K = the thread's first row: no cancellation between the mean and the spread
    float K[8], s[8], ss[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) K[e] = s[e] = ss[e] = 0.f;
    if (r < R) load8f<T>(x + r * C + l * 8, K);
    float cnt = 0.f;
#pragma unroll 4
    for (; r < R; r += stride) {
        float v[8];
        load8f<T>(x + r * C + l * 8, v);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float d = v[e] - K[e];
            s[e] += d;
            ss[e] = fmaf(d, d, ss[e]);
        }
        cnt += 1.f;
    }
    Wf3 w[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const float m = cnt > 0.f ? s[e] / cnt : 0.f;
        w[e] = Wf3{cnt, K[e] + m, cnt > 0.f ? fmaxf(ss[e] - s[e] * m, 0.f) : 0.f};
    }
    // rows of the warp that share the channel chunk, then the 8 warps
#pragma unroll
    for (int o = LPR; o < 32; o <<= 1) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            Wf3 b;
            b.n = __shfl_xor_sync(0xffffffffu, w[e].n, o);
            b.mean = __shfl_xor_sync(0xffffffffu, w[e].mean, o);
            b.m2 = __shfl_xor_sync(0xffffffffu, w[e].m2, o);
            // both lanes of a pair must produce the same result: combine in lane order
            w[e] = (lane & o) ? wf_merge(b, w[e]) : wf_merge(w[e], b);
        }
    }
    if (lane < LPR) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            red[warp][0][l * 8 + e] = w[e].n;
            red[warp][1][l * 8 + e] = w[e].mean;
            red[warp][2][l * 8 + e] = w[e].m2;
        }
    }
    __syncthreads();
    if (threadIdx.x < C) {
        const int c = threadIdx.x;
        Wf3 a{red[0][0][c], red[0][1][c], red[0][2][c]};
#pragma unroll
        for (int k = 1; k < 8; ++k) a = wf_merge(a, Wf3{red[k][0][c], red[k][1][c], red[k][2][c]});
        float *o = part + int64_t(blockIdx.x) * 3 * C;
        o[c] = a.n; o[C + c] = a.mean; o[2 * C + c] = a.m2;
    }
}

// one warp per channel: lane s combines partials s, s+32, ..., then the lanes combine; also the running statistics
// (momentum update with the unbiased variance, torch.nn.BatchNorm2d semantics)
__global__ void __launch_bounds__(1024)
bn_stats_finish_kernel(const float *__restrict__ part, int n_blk, int C, float eps, float momentum,
                       float *__restrict__ mean, float *__restrict__ rstd, float *__restrict__ running_mean,
                       float *__restrict__ running_var) {
    const int c = blockIdx.x * 32 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (c >= C) return;
    Wf3 a{0.f, 0.f, 0.f};
    for (int k = lane; k < n_blk; k += 32) {
        const float *p = part + int64_t(k) * 3 * C;
        a = wf_merge(a, Wf3{p[c], p[C + c], p[2 * C + c]});
    }
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        Wf3 b;
        b.n = __shfl_xor_sync(0xffffffffu, a.n, o);
        b.mean = __shfl_xor_sync(0xffffffffu, a.mean, o);
        b.m2 = __shfl_xor_sync(0xffffffffu, a.m2, o);
        a = (lane & o) ? wf_merge(b, a) : wf_merge(a, b);
    }
    if (lane == 0) {
        const float var = a.m2 / a.n;
        mean[c] = a.mean;
        rstd[c] = rsqrtf(var + eps);
        if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * a.mean;
        if (running_var) running_var[c] = (1.f - momentum) * running_var[c] + momentum * (a.m2 / fmaxf(a.n - 1.f, 1.f));
    }
}

template <typename T, int LPR, bool GELU>
__global__ void __launch_bounds__(kBT)
bn_apply_kernel(const T *__restrict__ x, const float *__restrict__ gamma, const float *__restrict__ beta,
                const float *__restrict__ mean, const float *__restrict__ rstd, T *__restrict__ y, int64_t R) {
    constexpr int RPC = kBT / LPR, C = LPR * 8;
    const int l = threadIdx.x % LPR, rl = threadIdx.x / LPR;
    float w[8], b[8], m[8], inv[8];
    load8f<float>(gamma + l * 8, w);
    load8f<float>(beta + l * 8, b);
    load8f<float>(mean + l * 8, m);
    load8f<float>(rstd + l * 8, inv);
    const int64_t stride = int64_t(gridDim.x) * RPC;
#pragma unroll 4
    for (int64_t r = int64_t(blockIdx.x) * RPC + rl; r < R; r += stride) {
        float v[8];
        load8f<T>(x + r * C + l * 8, v);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float z = w[e] * (v[e] - m[e]) * inv[e] + b[e];
            v[e] = GELU ? bn_gelu(bn_round<T>(z)) : z;
        }
        store8f<T>(y + r * C + l * 8, v);
    }
}

// g = dy (· GELU'(z), z = BN(x) rounded to T), rounded to T: what the stock GELU backward hands to the BatchNorm backward
// TAB (2-byte T): GELU'(z) from the 32 768-entry fp32 half table of gelu.cu in shared memory (z has 65 536 bit patterns;
// G(-z) = 1 - G(z)) instead of erf + exp per element, which bound both backward kernels.
template <typename T, bool GELU, bool TAB = false>
__device__ __forceinline__ float bn_g(float dy, float xm, float w, float inv, float b, const float *tabf = nullptr) {
    if constexpr (!GELU) {
        return dy;
    } else if constexpr (TAB) {
        const T z = Elem<T>::from_acc(w * xm * inv + b);
        const uint32_t bits = *reinterpret_cast<const uint16_t *>(&z);
        float g = tabf[bits & 0x7fffu];
        if (bits & 0x8000u) g = 1.f - g;
        return bn_round<T>(dy * g);
    } else {
        return bn_round<T>(bn_gelu_grad(bn_round<T>(w * xm * inv + b), dy));
    }
}

constexpr int kBnTabBytes = 128 * 1024;
constexpr int kBnTabThreads = 1024;
__device__ __forceinline__ void bn_load_table(float *tabf, const float *tab_g, int nthreads) {
    for (int i = threadIdx.x; i < kBnTabBytes / 16; i += nthreads)
        reinterpret_cast<uint4 *>(tabf)[i] = __ldg(reinterpret_cast<const uint4 *>(tab_g) + i);
    __syncthreads();
}

// per-CTA partial sums: part[blk][2][C] = (sum g, sum g·(x − mean))
template <typename T, int LPR, bool GELU, int BT = kBT, bool TAB = false>
__global__ void __launch_bounds__(BT)
bn_bwd_reduce_kernel(const T *__restrict__ dy, const T *__restrict__ x, const float *__restrict__ gamma,
                     const float *__restrict__ beta, const float *__restrict__ mean, const float *__restrict__ rstd,
                     float *__restrict__ part, int64_t R, const float *__restrict__ tab_g = nullptr) {
    constexpr int RPC = BT / LPR, C = LPR * 8;
    __shared__ float red[BT / 32][2][C];
    extern __shared__ __align__(16) float bn_tabf[];
    if constexpr (TAB) bn_load_table(bn_tabf, tab_g, BT);
    const int l = threadIdx.x % LPR, rl = threadIdx.x / LPR, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float w[8], b[8], m[8], inv[8], a1[8], a2[8];
    load8f<float>(gamma + l * 8, w);
    load8f<float>(beta + l * 8, b);
    load8f<float>(mean + l * 8, m);
    load8f<float>(rstd + l * 8, inv);
#pragma unroll
    for (int e = 0; e < 8; ++e) a1[e] = a2[e] = 0.f;
    const int64_t stride = int64_t(gridDim.x) * RPC;
#pragma unroll 2
    for (int64_t r = int64_t(blockIdx.x) * RPC + rl; r < R; r += stride) {
        float v[8], d[8];
        load8f<T>(x + r * C + l * 8, v);
        load8f<T>(dy + r * C + l * 8, d);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float xm = v[e] - m[e];
            const float g = bn_g<T, GELU, TAB>(d[e], xm, w[e], inv[e], b[e], bn_tabf);
            a1[e] += g;
            a2[e] = fmaf(g, xm, a2[e]);
        }
    }
#pragma unroll
    for (int o = LPR; o < 32; o <<= 1) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            a1[e] += __shfl_xor_sync(0xffffffffu, a1[e], o);
            a2[e] += __shfl_xor_sync(0xffffffffu, a2[e], o);
        }
    }
    if (lane < LPR) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            red[warp][0][l * 8 + e] = a1[e];
            red[warp][1][l * 8 + e] = a2[e];
        }
    }
    __syncthreads();
    if (threadIdx.x < 2 * C) {
        const int which = threadIdx.x / C, c = threadIdx.x % C;
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < BT / 32; ++k) acc += red[k][which][c];
        part[(int64_t(blockIdx.x) * 2 + which) * C + c] = acc;
    }
}

// sums[0][c] = sum g, sums[1][c] = sum g·(x − mean);  dbeta = sums[0], dgamma = sums[1]·rstd
__global__ void __launch_bounds__(1024)
bn_bwd_finish_kernel(const float *__restrict__ part, int n_blk, int C, const float *__restrict__ rstd,
                     float *__restrict__ sums, float *__restrict__ dgamma, float *__restrict__ dbeta) {
    const int c = blockIdx.x * 32 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (c >= C) return;
    float a1 = 0.f, a2 = 0.f;
    for (int k = lane; k < n_blk; k += 32) {
        a1 += part[(int64_t(k) * 2 + 0) * C + c];
        a2 += part[(int64_t(k) * 2 + 1) * C + c];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a1 += __shfl_xor_sync(0xffffffffu, a1, o);
        a2 += __shfl_xor_sync(0xffffffffu, a2, o);
    }
    if (lane == 0) {
        sums[c] = a1;
        sums[C + c] = a2;
        if (dbeta) dbeta[c] = a1;
        if (dgamma) dgamma[c] = a2 * rstd[c];
    }
}

// dx = (g − mean(g) − (x − mean)·rstd²·mean(g·(x − mean)))·rstd·gamma
template <typename T, int LPR, bool GELU, int BT = kBT, bool TAB = false>
__global__ void __launch_bounds__(BT)
bn_bwd_elemt_kernel(const T *__restrict__ dy, const T *__restrict__ x, const float *__restrict__ gamma,
                    const float *__restrict__ beta, const float *__restrict__ mean, const float *__restrict__ rstd,
                    const float *__restrict__ sums, T *__restrict__ dx, int64_t R, const float *__restrict__ tab_g = nullptr) {
    constexpr int RPC = BT / LPR, C = LPR * 8;
    extern __shared__ __align__(16) float bn_tabf[];
    if constexpr (TAB) bn_load_table(bn_tabf, tab_g, BT);
    const int l = threadIdx.x % LPR, rl = threadIdx.x / LPR;
    float w[8], b[8], m[8], inv[8], mdy[8], f1[8], f2[8];
    load8f<float>(gamma + l * 8, w);
    load8f<float>(beta + l * 8, b);
    load8f<float>(mean + l * 8, m);
    load8f<float>(rstd + l * 8, inv);
    load8f<float>(sums + l * 8, mdy);
    load8f<float>(sums + C + l * 8, f1);
    const float norm = 1.f / float(R);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        mdy[e] *= norm;
        f1[e] = inv[e] * inv[e] * f1[e] * norm;
        f2[e] = w[e] * inv[e];
    }
    const int64_t stride = int64_t(gridDim.x) * RPC;
#pragma unroll 2
    for (int64_t r = int64_t(blockIdx.x) * RPC + rl; r < R; r += stride) {
        float v[8], d[8];
        load8f<T>(x + r * C + l * 8, v);
        load8f<T>(dy + r * C + l * 8, d);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float xm = v[e] - m[e];
            const float g = bn_g<T, GELU, TAB>(d[e], xm, w[e], inv[e], b[e], bn_tabf);
            v[e] = (g - mdy[e] - xm * f1[e]) * f2[e];
        }
        store8f<T>(dx + r * C + l * 8, v);
    }
}

static int bn_grid(int64_t R, int lpr) {
    const int64_t rpc = kBT / lpr;
    return int(std::min<int64_t>((R + rpc - 1) / rpc, int64_t(kNumSMs) * 8));
}

static bool bn_shape_ok(const char *what, int64_t R, int64_t C, int dtype) {
    if (R <= 0 || (C != 8 && C != 16 && C != 32 && C != 64 && C != 128)) {
        set_error("%s: need R > 0 and C in {8,16,32,64,128} (got R=%lld C=%lld)", what, (long long)R, (long long)C);
        return false;
    }
    if (dtype != AFFB200_F32 && dtype != AFFB200_F16 && dtype != AFFB200_BF16) {
        set_error("%s: f32 / f16 / bf16 only", what);
        return false;
    }
    return true;
}

}  // namespace affb200

using namespace affb200;

// floats of scratch for either direction: the per-CTA partials plus the two per-channel sums of the backward
extern "C" int64_t affb200_batch_norm_work_floats(int64_t R, int64_t C) {
    if (R <= 0 || C <= 0 || C % 8) return 0;
    return int64_t(bn_grid(R, int(C / 8))) * 3 * C + 2 * C;
}

#define AFF_BN_LPR(CALL)                                                        \
    switch (int(C / 8)) {                                                       \
        case 1: CALL(1); break;                                                 \
        case 2: CALL(2); break;                                                 \
        case 4: CALL(4); break;                                                 \
        case 8: CALL(8); break;                                                 \
        default: CALL(16); break;                                               \
    }

extern "C" int affb200_batch_norm_fwd(const void *x, const float *gamma, const float *beta, float *running_mean,
                                      float *running_var, float momentum, float eps, int gelu, void *y, float *mean,
                                      float *rstd, float *work, int64_t R, int64_t C, int dtype, void *stream) {
    if (!bn_shape_ok("batch_norm_fwd", R, C, dtype)) return AFFB200_E_ARG;
    if (!x || !gamma || !beta || !y || !mean || !rstd || !work || !aligned16(x) || !aligned16(y) || !aligned16(gamma) ||
        !aligned16(beta) || !aligned16(mean) || !aligned16(rstd)) {
        set_error("batch_norm_fwd: null or misaligned pointer");
        return AFFB200_E_ARG;
    }
    auto st = (cudaStream_t)stream;
    const int grid = bn_grid(R, int(C / 8));
    return dispatch_dtype(dtype, [&](auto tag) -> int {
        using T = decltype(tag);
        if constexpr (std::is_same<T, double>::value) {
            return AFFB200_E_ARG;
        } else {
#define AFF_BN_STATS(L_) bn_stats_kernel<T, L_><<<grid, kBT, 0, st>>>((const T *)x, work, R)
            AFF_BN_LPR(AFF_BN_STATS)
#undef AFF_BN_STATS
            bn_stats_finish_kernel<<<unsigned((C + 31) / 32), 1024, 0, st>>>(work, grid, int(C), eps, momentum, mean, rstd,
                                                                             running_mean, running_var);
#define AFF_BN_APPLY(L_)                                                                                          \
    do {                                                                                                          \
        if (gelu) bn_apply_kernel<T, L_, true><<<grid, kBT, 0, st>>>((const T *)x, gamma, beta, mean, rstd, (T *)y, R); \
        else bn_apply_kernel<T, L_, false><<<grid, kBT, 0, st>>>((const T *)x, gamma, beta, mean, rstd, (T *)y, R);     \
    } while (0)
            AFF_BN_LPR(AFF_BN_APPLY)
#undef AFF_BN_APPLY
            return check_launch("batch_norm_fwd", 3);
        }
    });
}

// y = act(gamma * (x - mean) * rstd + beta) with GIVEN per-channel mean / rstd: evaluation-mode BatchNorm (running
// statistics) as one pass, GELU included.
extern "C" int affb200_batch_norm_apply(const void *x, const float *gamma, const float *beta, const float *mean,
                                        const float *rstd, int gelu, void *y, int64_t R, int64_t C, int dtype,
                                        void *stream) {
    if (!bn_shape_ok("batch_norm_apply", R, C, dtype)) return AFFB200_E_ARG;
    if (!x || !gamma || !beta || !mean || !rstd || !y || !aligned16(x) || !aligned16(y) || !aligned16(gamma) ||
        !aligned16(beta) || !aligned16(mean) || !aligned16(rstd)) {
        set_error("batch_norm_apply: null or misaligned pointer");
        return AFFB200_E_ARG;
    }
    auto st = (cudaStream_t)stream;
    const int grid = bn_grid(R, int(C / 8));
    return dispatch_dtype(dtype, [&](auto tag) -> int {
        using T = decltype(tag);
        if constexpr (std::is_same<T, double>::value) {
            return AFFB200_E_ARG;
        } else {
#define AFF_BN_APPLY2(L_)                                                                                         \
    do {                                                                                                          \
        if (gelu) bn_apply_kernel<T, L_, true><<<grid, kBT, 0, st>>>((const T *)x, gamma, beta, mean, rstd, (T *)y, R); \
        else bn_apply_kernel<T, L_, false><<<grid, kBT, 0, st>>>((const T *)x, gamma, beta, mean, rstd, (T *)y, R);     \
    } while (0)
            AFF_BN_LPR(AFF_BN_APPLY2)
#undef AFF_BN_APPLY2
            return check_launch("batch_norm_apply");
        }
    });
}

extern "C" int affb200_batch_norm_bwd(const void *dy, const void *x, const float *gamma, const float *beta,
                                      const float *mean, const float *rstd, int gelu, void *dx, float *dgamma,
                                      float *dbeta, float *work, int64_t R, int64_t C, int dtype,
                                      const float *gelu_bwd_tab, void *stream) {
    if (!bn_shape_ok("batch_norm_bwd", R, C, dtype)) return AFFB200_E_ARG;
    if (gelu_bwd_tab && !aligned16(gelu_bwd_tab)) { set_error("batch_norm_bwd: misaligned table"); return AFFB200_E_ARG; }
    if (!dy || !x || !gamma || !beta || !mean || !rstd || !dx || !work || !aligned16(dy) || !aligned16(x) ||
        !aligned16(dx) || !aligned16(gamma) || !aligned16(beta) || !aligned16(mean) || !aligned16(rstd) || !aligned16(work)) {
        set_error("batch_norm_bwd: null or misaligned pointer");
        return AFFB200_E_ARG;
    }
    auto st = (cudaStream_t)stream;
    const int grid = bn_grid(R, int(C / 8));
    float *sums = work, *part = work + 2 * C;
    return dispatch_dtype(dtype, [&](auto tag) -> int {
        using T = decltype(tag);
        if constexpr (std::is_same<T, double>::value) {
            return AFFB200_E_ARG;
        } else {
            if constexpr (sizeof(T) == 2) {
                if (gelu && gelu_bwd_tab) {   // one 1024-thread CTA per SM around the 128 KB derivative table
                    const int rpc = kBnTabThreads / int(C / 8);
                    const int tg = int(std::min<int64_t>((R + rpc - 1) / rpc, int64_t(kNumSMs)));
#define AFF_BN_TAB(L_)                                                                                                  \
    do {                                                                                                                \
        auto kr = bn_bwd_reduce_kernel<T, L_, true, kBnTabThreads, true>;                                               \
        auto ke = bn_bwd_elemt_kernel<T, L_, true, kBnTabThreads, true>;                                                \
        if (cudaFuncSetAttribute(kr, cudaFuncAttributeMaxDynamicSharedMemorySize, kBnTabBytes) != cudaSuccess ||        \
            cudaFuncSetAttribute(ke, cudaFuncAttributeMaxDynamicSharedMemorySize, kBnTabBytes) != cudaSuccess)          \
            return check_launch("cudaFuncSetAttribute", 0);                                                             \
        kr<<<tg, kBnTabThreads, kBnTabBytes, st>>>((const T *)dy, (const T *)x, gamma, beta, mean, rstd, part, R, gelu_bwd_tab); \
        bn_bwd_finish_kernel<<<unsigned((C + 31) / 32), 1024, 0, st>>>(part, tg, int(C), rstd, sums, dgamma, dbeta);   \
        ke<<<tg, kBnTabThreads, kBnTabBytes, st>>>((const T *)dy, (const T *)x, gamma, beta, mean, rstd, sums, (T *)dx, R, gelu_bwd_tab); \
    } while (0)
                    AFF_BN_LPR(AFF_BN_TAB)
#undef AFF_BN_TAB
                    return check_launch("batch_norm_bwd", 3);
                }
            }
#define AFF_BN_RED(L_)                                                                                                  \
    do {                                                                                                                \
        if (gelu) bn_bwd_reduce_kernel<T, L_, true><<<grid, kBT, 0, st>>>((const T *)dy, (const T *)x, gamma, beta, mean, rstd, part, R); \
        else bn_bwd_reduce_kernel<T, L_, false><<<grid, kBT, 0, st>>>((const T *)dy, (const T *)x, gamma, beta, mean, rstd, part, R);     \
    } while (0)
            AFF_BN_LPR(AFF_BN_RED)
#undef AFF_BN_RED
            bn_bwd_finish_kernel<<<unsigned((C + 31) / 32), 1024, 0, st>>>(part, grid, int(C), rstd, sums, dgamma, dbeta);
#define AFF_BN_EL(L_)                                                                                                   \
    do {                                                                                                                \
        if (gelu) bn_bwd_elemt_kernel<T, L_, true><<<grid, kBT, 0, st>>>((const T *)dy, (const T *)x, gamma, beta, mean, rstd, sums, (T *)dx, R); \
        else bn_bwd_elemt_kernel<T, L_, false><<<grid, kBT, 0, st>>>((const T *)dy, (const T *)x, gamma, beta, mean, rstd, sums, (T *)dx, R);     \
    } while (0)
            AFF_BN_LPR(AFF_BN_EL)
#undef AFF_BN_EL
            return check_launch("batch_norm_bwd", 3);
        }
    });
}
