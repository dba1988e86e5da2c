// layer_norm.cu — LayerNorm over the channel dim of token rows, the op on either side of the off-grid attention
// (norm1 -> q/kv Linear, norm2 -> Mlp, the norm of ClusterMerging; reference models/aff_transformer.py:199-200,
// 223-229, 284, 362).  Under autocast the reference runs nn.LayerNorm in fp32: the bf16 residual stream is copied to
// fp32, normalised (fp32 out) and copied back to bf16 by the next Linear — 20 bytes per element for a 4-byte job.
// Here one pass reads the row in its storage type, keeps it in registers, does the statistics in fp32 (two-pass:
// mean, then centred second moment) and writes the result directly in the consumer's type; the backward does dx and
// the per-CTA partial sums of dgamma/dbeta in one pass, a second small kernel adds the partials in a fixed order
// (no atomics: bit-reproducible).
//
// Mapping: LPR lanes share a row (8, 16 or 32), lane l owns the 8-element chunks l, l+LPR, ... (NCH of them).
#include "common.cuh"
#include <algorithm>

namespace affb200 {

constexpr int kLT = 256;

template <typename T> __device__ __forceinline__ void ln_load8(const T *p, float (&o)[8]) {
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
template <typename T> __device__ __forceinline__ void ln_store8(T *p, const float (&v)[8]) {
    if constexpr (sizeof(T) == 4) {
        reinterpret_cast<float4 *>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
        reinterpret_cast<float4 *>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
    } else {
        uint4 r;
        T *e = reinterpret_cast<T *>(&r);
#pragma unroll
        for (int k = 0; k < 8; ++k) e[k] = Elem<T>::from_acc(v[k]);
        *reinterpret_cast<uint4 *>(p) = r;
    }
}
template <int LPR> __device__ __forceinline__ float ln_group_sum(float v) {
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <typename TI, typename TO, int LPR, int NCH>
__global__ void __launch_bounds__(kLT)
ln_fwd_kernel(const TI *__restrict__ x, const float *__restrict__ gamma, const float *__restrict__ beta,
              TO *__restrict__ y, float *__restrict__ mean, float *__restrict__ rstd, int64_t R, int C, float eps) {
    constexpr int RPC = kLT / LPR;  // rows per CTA pass
    const int l = threadIdx.x % LPR, rl = threadIdx.x / LPR;
    const int nch = C >> 3;
    const float invC = 1.f / float(C);
    for (int64_t r0 = int64_t(blockIdx.x) * RPC; r0 < R; r0 += int64_t(gridDim.x) * RPC) {
        const int64_t r = r0 + rl;
        const bool live = r < R;
        float v[NCH][8];
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            const int ch = l + k * LPR;
#pragma unroll
            for (int e = 0; e < 8; ++e) v[k][e] = 0.f;
            if (live && ch < nch) ln_load8<TI>(x + r * C + ch * 8, v[k]);
#pragma unroll
            for (int e = 0; e < 8; ++e) s += v[k][e];
        }
        const float mu = ln_group_sum<LPR>(s) * invC;
        float sq = 0.f;
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            const int ch = l + k * LPR;
            if (ch < nch) {
#pragma unroll
                for (int e = 0; e < 8; ++e) { const float d = v[k][e] - mu; sq += d * d; }
            }
        }
        const float rs = rsqrtf(ln_group_sum<LPR>(sq) * invC + eps);
        if (live) {
#pragma unroll
            for (int k = 0; k < NCH; ++k) {
                const int ch = l + k * LPR;
                if (ch < nch) {
                    float g[8], bt[8], o[8];
                    ln_load8<float>(gamma + ch * 8, g);
                    ln_load8<float>(beta + ch * 8, bt);
#pragma unroll
                    for (int e = 0; e < 8; ++e) o[e] = (v[k][e] - mu) * rs * g[e] + bt[e];
                    ln_store8<TO>(y + r * C + ch * 8, o);
                }
            }
            if (l == 0) { mean[r] = mu; rstd[r] = rs; }
        }
    }
}

// dx and per-CTA partials part[blk][2][C] = (sum dy*xhat, sum dy) over the CTA's rows
template <typename TI, typename TG, int LPR, int NCH>
__global__ void __launch_bounds__(kLT)
ln_bwd_kernel(const TG *__restrict__ dy, const TI *__restrict__ x, const float *__restrict__ gamma,
              const float *__restrict__ mean, const float *__restrict__ rstd, const TI *__restrict__ dres,
              TI *__restrict__ dx, float *__restrict__ part, int64_t R, int C) {
    constexpr int RPC = kLT / LPR;
    extern __shared__ float red[];  // [8 warps][C]
    const int l = threadIdx.x % LPR, rl = threadIdx.x / LPR, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nch = C >> 3;
    const float invC = 1.f / float(C);
    float ag[NCH][8], ab[NCH][8], gm[NCH][8];
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
        const int ch = l + k * LPR;
#pragma unroll
        for (int e = 0; e < 8; ++e) ag[k][e] = ab[k][e] = gm[k][e] = 0.f;
        if (ch < nch) ln_load8<float>(gamma + ch * 8, gm[k]);
    }
    for (int64_t r0 = int64_t(blockIdx.x) * RPC; r0 < R; r0 += int64_t(gridDim.x) * RPC) {
        const int64_t r = r0 + rl;
        const bool live = r < R;
        const float mu = live ? __ldg(mean + r) : 0.f, rs = live ? __ldg(rstd + r) : 0.f;
        float xh[NCH][8], g[NCH][8];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            const int ch = l + k * LPR;
#pragma unroll
            for (int e = 0; e < 8; ++e) xh[k][e] = g[k][e] = 0.f;
            if (live && ch < nch) {
                float xv[8], dv[8];
                ln_load8<TI>(x + r * C + ch * 8, xv);
                ln_load8<TG>(dy + r * C + ch * 8, dv);
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    xh[k][e] = (xv[e] - mu) * rs;
                    ag[k][e] += dv[e] * xh[k][e];
                    ab[k][e] += dv[e];
                    g[k][e] = dv[e] * gm[k][e];
                    s1 += g[k][e];
                    s2 += g[k][e] * xh[k][e];
                }
            }
        }
        s1 = ln_group_sum<LPR>(s1) * invC;
        s2 = ln_group_sum<LPR>(s2) * invC;
        if (live) {
#pragma unroll
            for (int k = 0; k < NCH; ++k) {
                const int ch = l + k * LPR;
                if (ch < nch) {
                    float o[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) o[e] = rs * (g[k][e] - s1 - xh[k][e] * s2);
                    if (dres) {  // gradient arriving over the residual branch that forked off x
                        float rv[8];
                        ln_load8<TI>(dres + r * C + ch * 8, rv);
#pragma unroll
                        for (int e = 0; e < 8; ++e) o[e] += rv[e];
                    }
                    ln_store8<TI>(dx + r * C + ch * 8, o);
                }
            }
        }
    }
    // rows of one warp that share a lane position, then the 8 warps through shared memory, dgamma then dbeta
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            const int ch = l + k * LPR;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                float a = pass == 0 ? ag[k][e] : ab[k][e];
#pragma unroll
                for (int o = LPR; o < 32; o <<= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
                if (lane < LPR && ch < nch) red[warp * C + ch * 8 + e] = a;
            }
        }
        __syncthreads();
        for (int c = threadIdx.x; c < C; c += kLT) {
            float a = 0.f;
#pragma unroll
            for (int w = 0; w < kLT / 32; ++w) a += red[w * C + c];
            part[(int64_t(blockIdx.x) * 2 + pass) * C + c] = a;
        }
        __syncthreads();
    }
}

// dgamma / dbeta = sum of the CTA partials: a block owns 32 columns, 32 threads per column stride over the CTAs
__global__ void __launch_bounds__(1024)
ln_bwd_finish_kernel(const float *__restrict__ part, int nblk, int C, float *__restrict__ dgamma,
                     float *__restrict__ dbeta, int accumulate) {
    __shared__ float red[32][33];
    const int cl = threadIdx.x & 31, sub = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + cl;
    const int pass = blockIdx.y;
    float a = 0.f;
    if (c < C)
        for (int b = sub; b < nblk; b += 32) a += part[(int64_t(b) * 2 + pass) * C + c];
    red[sub][cl] = a;
    __syncthreads();
    if (sub == 0 && c < C) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 32; ++k) t += red[k][cl];
        float *o = (pass == 0 ? dgamma : dbeta) + c;
        *o = accumulate ? *o + t : t;
    }
}

static int ln_grid(int64_t R, int rpc, int per_sm) {
    return int(std::max<int64_t>(1, std::min<int64_t>(ceil_div64(R, rpc), int64_t(kNumSMs) * per_sm)));
}

template <typename TI, typename TO>
static int ln_fwd_launch(const void *x, const float *gamma, const float *beta, void *y, float *mean, float *rstd,
                         int64_t R, int C, float eps, cudaStream_t st) {
    const int nch = C / 8;
#define AFF_LNF(LPR_, NCH_)                                                                                       \
    ln_fwd_kernel<TI, TO, LPR_, NCH_><<<ln_grid(R, kLT / LPR_, 8), kLT, 0, st>>>((const TI *)x, gamma, beta, (TO *)y, \
                                                                                  mean, rstd, R, C, eps)
    if (nch <= 8) AFF_LNF(8, 1);
    else if (nch <= 16) AFF_LNF(16, 1);
    else if (nch <= 32) AFF_LNF(32, 1);
    else if (nch <= 64) AFF_LNF(32, 2);
    else if (nch <= 128) AFF_LNF(32, 4);
    else AFF_LNF(32, 8);
#undef AFF_LNF
    return check_launch("layer_norm_fwd");
}

template <typename TI, typename TG>
static int ln_bwd_launch(const void *dy, const void *x, const float *gamma, const float *mean, const float *rstd,
                         const void *dres, void *dx, float *part, int nblk, int64_t R, int C, cudaStream_t st) {
    const int nch = C / 8;
    const size_t smem = size_t(kLT / 32) * C * 4;
#define AFF_LNB(LPR_, NCH_)                                                                                      \
    ln_bwd_kernel<TI, TG, LPR_, NCH_><<<nblk, kLT, smem, st>>>((const TG *)dy, (const TI *)x, gamma, mean, rstd,  \
                                                               (const TI *)dres, (TI *)dx, part, R, C)
    if (nch <= 8) AFF_LNB(8, 1);
    else if (nch <= 16) AFF_LNB(16, 1);
    else if (nch <= 32) AFF_LNB(32, 1);
    else if (nch <= 64) AFF_LNB(32, 2);
    else AFF_LNB(32, 4);
#undef AFF_LNB
    return check_launch("layer_norm_bwd");
}

template <typename F> static int ln_dispatch3(int dtype, F &&f) {
    switch (dtype) {
        case AFFB200_F32: return f(float{});
        case AFFB200_F16: return f(__half{});
        case AFFB200_BF16: return f(__nv_bfloat16{});
        default: set_error("layer_norm: unsupported dtype enum %d", dtype); return AFFB200_E_ARG;
    }
}

}  // namespace affb200

using namespace affb200;

extern "C" int affb200_layer_norm_fwd(const void *x, const float *gamma, const float *beta, void *y, float *mean,
                                      float *rstd, int64_t R, int64_t C, float eps, int in_dtype, int out_dtype,
                                      void *stream) {
    if (R < 0 || C <= 0 || C % 8 != 0 || C > 2048) { set_error("layer_norm_fwd: C must be a multiple of 8, <= 2048"); return AFFB200_E_ARG; }
    if (R == 0) return AFFB200_OK;
    if (!x || !gamma || !beta || !y || !mean || !rstd || !aligned16(x) || !aligned16(y) || !aligned16(gamma) || !aligned16(beta)) {
        set_error("layer_norm_fwd: null or misaligned pointer");
        return AFFB200_E_ARG;
    }
    auto st = (cudaStream_t)stream;
    return ln_dispatch3(in_dtype, [&](auto ti) {
        using TI = decltype(ti);
        return ln_dispatch3(out_dtype, [&](auto to) {
            using TO = decltype(to);
            return ln_fwd_launch<TI, TO>(x, gamma, beta, y, mean, rstd, R, int(C), eps, st);
        });
    });
}

extern "C" int64_t affb200_layer_norm_bwd_work_floats(int64_t R, int64_t C) {
    if (R <= 0 || C <= 0) return 0;
    return int64_t(kNumSMs) * 4 * 2 * C;
}

extern "C" int affb200_layer_norm_bwd(const void *dy, const void *x, const float *gamma, const float *mean,
                                      const float *rstd, const void *dres, void *dx, float *dgamma, float *dbeta,
                                      float *work, int64_t R, int64_t C, int x_dtype, int dy_dtype, int accumulate,
                                      void *stream) {
    if (R < 0 || C <= 0 || C % 8 != 0 || C > 1024) { set_error("layer_norm_bwd: C must be a multiple of 8, <= 1024"); return AFFB200_E_ARG; }
    auto st = (cudaStream_t)stream;
    if (!dgamma || !dbeta) { set_error("layer_norm_bwd: null pointer"); return AFFB200_E_ARG; }
    if (R == 0) {
        if (accumulate) return AFFB200_OK;
        cudaMemsetAsync(dgamma, 0, size_t(C) * 4, st);
        cudaMemsetAsync(dbeta, 0, size_t(C) * 4, st);
        return check_launch("layer_norm_bwd memset", 0);
    }
    if (!dy || !x || !gamma || !mean || !rstd || !dx || !work || !aligned16(dy) || !aligned16(x) || !aligned16(dx) ||
        !aligned16(gamma) || !aligned16(dres)) {
        set_error("layer_norm_bwd: null or misaligned pointer");
        return AFFB200_E_ARG;
    }
    const int nch = int(C / 8);
    const int lpr = nch <= 8 ? 8 : (nch <= 16 ? 16 : 32);
    const int nblk = ln_grid(R, kLT / lpr, 4);
    int e = ln_dispatch3(x_dtype, [&](auto ti) {
        using TI = decltype(ti);
        return ln_dispatch3(dy_dtype, [&](auto tg) {
            using TG = decltype(tg);
            return ln_bwd_launch<TI, TG>(dy, x, gamma, mean, rstd, dres, dx, work, nblk, R, int(C), st);
        });
    });
    if (e) return e;
    ln_bwd_finish_kernel<<<dim3(unsigned((C + 31) / 32), 2), 1024, 0, st>>>(work, nblk, int(C), dgamma, dbeta, accumulate);
    return check_launch("layer_norm_bwd_finish");
}
