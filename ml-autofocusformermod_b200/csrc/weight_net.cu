// weight_net.cu — the PointConv weight table of ClusterMerging: weights_table = GELU(LayerNorm_4(Linear_{5->4}(pre_table)))
// (reference models/aff_transformer.py:262-264,343: `self.weight_net(pre_table)`, an nn.Sequential evaluated on the
// (2w+1)^2 x 5 relative-position feature table in every merge).  The stock chain is a GEMM with N=4, a LayerNorm over
// 4 columns and a GELU over 12 321 x 4 elements forward, and backward a LayerNorm gamma/beta reduction that runs in
// one column block (≈ 60 µs per merge for 49 284 elements) plus two more skinny GEMMs.  Here one thread owns a table
// row: 5 inputs, 4 outputs, everything in registers; the backward recomputes the row and leaves the 32 parameter
// gradient sums (dW 4x5, db 4, dgamma 4, dbeta 4) as per-CTA partials for the fixed-order column-sum finish.
//
// Arithmetic follows what autocast does to the three modules: the Linear in the compute dtype T (inputs, weight and
// bias rounded to T, fp32 accumulate, result rounded to T), LayerNorm and GELU in fp32; the gradient that reaches the
// Linear is rounded to T (autograd casts it to the Linear's output dtype).  T = float: no rounding anywhere.
#include "common.cuh"
#include <algorithm>

namespace affb200 {

int launch_col_sum_finish(const float *part, int n_slabs, int C, float *out, int accumulate, cudaStream_t st);

constexpr int kWnT = 128;     // threads per CTA
constexpr int kWnSums = 32;   // dW[4][5], db[4], dgamma[4], dbeta[4]

template <typename T> __device__ __forceinline__ float wn_round(float x) {
    if constexpr (sizeof(T) == 4) return x;
    else return Elem<T>::to_acc(Elem<T>::from_acc(x));
}

struct WnParams {
    float w[4][5], b[4], gamma[4], beta[4];
};

template <typename T>
__device__ __forceinline__ void wn_load_params(const float *w, const float *b, const float *gamma, const float *beta,
                                               WnParams &p) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int k = 0; k < 5; ++k) p.w[j][k] = wn_round<T>(__ldg(w + j * 5 + k));
        p.b[j] = wn_round<T>(__ldg(b + j));
        p.gamma[j] = __ldg(gamma + j);
        p.beta[j] = __ldg(beta + j);
    }
}

// one table row: x[5] (rounded to T) -> xhat[4], rstd, z[4] = LayerNorm output
template <typename T>
__device__ __forceinline__ void wn_row(const float *feat_row, const WnParams &p, float eps, float (&x)[5],
                                       float (&xhat)[4], float &rstd, float (&z)[4]) {
#pragma unroll
    for (int k = 0; k < 5; ++k) x[k] = wn_round<T>(__ldg(feat_row + k));
    float h[4], mean = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        float a = p.b[j];
#pragma unroll
        for (int k = 0; k < 5; ++k) a = fmaf(x[k], p.w[j][k], a);
        h[j] = wn_round<T>(a);
        mean += h[j];
    }
    mean *= 0.25f;
    float var = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        h[j] -= mean;
        var = fmaf(h[j], h[j], var);
    }
    rstd = rsqrtf(var * 0.25f + eps);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        xhat[j] = h[j] * rstd;
        z[j] = fmaf(xhat[j], p.gamma[j], p.beta[j]);
    }
}

template <typename T>
__global__ void __launch_bounds__(kWnT)
weight_net_fwd_kernel(const float *__restrict__ feat, const float *__restrict__ w, const float *__restrict__ b,
                      const float *__restrict__ gamma, const float *__restrict__ beta, float eps, int P,
                      float *__restrict__ out) {
    WnParams p;
    wn_load_params<T>(w, b, gamma, beta, p);
    for (int r = blockIdx.x * kWnT + threadIdx.x; r < P; r += gridDim.x * kWnT) {
        float x[5], xhat[4], z[4], rstd;
        wn_row<T>(feat + int64_t(r) * 5, p, eps, x, xhat, rstd, z);
        float4 o;
        o.x = z[0] * 0.5f * (1.f + erff(z[0] * 0.70710678118654752440f));
        o.y = z[1] * 0.5f * (1.f + erff(z[1] * 0.70710678118654752440f));
        o.z = z[2] * 0.5f * (1.f + erff(z[2] * 0.70710678118654752440f));
        o.w = z[3] * 0.5f * (1.f + erff(z[3] * 0.70710678118654752440f));
        reinterpret_cast<float4 *>(out)[r] = o;
    }
}

template <typename T>
__global__ void __launch_bounds__(kWnT)
weight_net_bwd_kernel(const float *__restrict__ feat, const float *__restrict__ w, const float *__restrict__ b,
                      const float *__restrict__ gamma, const float *__restrict__ beta, float eps, int P,
                      const float *__restrict__ d_out, float *__restrict__ part) {
    __shared__ float red[kWnT / 32][kWnSums];
    WnParams p;
    wn_load_params<T>(w, b, gamma, beta, p);
    float s[kWnSums];
#pragma unroll
    for (int i = 0; i < kWnSums; ++i) s[i] = 0.f;
    for (int r = blockIdx.x * kWnT + threadIdx.x; r < P; r += gridDim.x * kWnT) {
        float x[5], xhat[4], z[4], rstd;
        wn_row<T>(feat + int64_t(r) * 5, p, eps, x, xhat, rstd, z);
        const float4 go = __ldg(reinterpret_cast<const float4 *>(d_out) + r);
        const float g[4] = {go.x, go.y, go.z, go.w};
        float dxh[4], m1 = 0.f, m2 = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            constexpr float kBeta = 1.12837916709551257390f * 0.70710678118654752440f * 0.5f;
            const float cdf = 0.5f * (1.f + erff(z[j] * 0.70710678118654752440f));
            const float pdf = expf(-0.5f * z[j] * z[j]) * kBeta;
            const float dz = g[j] * (cdf + z[j] * pdf);
            s[24 + j] = fmaf(dz, xhat[j], s[24 + j]);     // dgamma
            s[28 + j] += dz;                              // dbeta
            dxh[j] = dz * p.gamma[j];
            m1 += dxh[j];
            m2 = fmaf(dxh[j], xhat[j], m2);
        }
        m1 *= 0.25f;
        m2 *= 0.25f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float dh = wn_round<T>(rstd * (dxh[j] - m1 - xhat[j] * m2));
            s[20 + j] += dh;                              // db
#pragma unroll
            for (int k = 0; k < 5; ++k) s[j * 5 + k] = fmaf(dh, x[k], s[j * 5 + k]);   // dW
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < kWnSums; ++i) {
        float a = s[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (lane == 0) red[warp][i] = a;
    }
    __syncthreads();
    if (threadIdx.x < kWnSums) {
        float a = 0.f;
#pragma unroll
        for (int k = 0; k < kWnT / 32; ++k) a += red[k][threadIdx.x];
        part[blockIdx.x * kWnSums + threadIdx.x] = a;
    }
}

static int wn_grid(int64_t P) { return int(std::max<int64_t>(1, std::min<int64_t>((P + kWnT - 1) / kWnT, kNumSMs))); }

static int wn_check(const char *what, int64_t P, int dtype) {
    if (P < 0 || P >= (int64_t(1) << 31)) { set_error("%s: unsupported table size", what); return AFFB200_E_ARG; }
    if (dtype != AFFB200_F32 && dtype != AFFB200_F16 && dtype != AFFB200_BF16) {
        set_error("%s: compute dtype must be f32 / f16 / bf16", what);
        return AFFB200_E_ARG;
    }
    return AFFB200_OK;
}

}  // namespace affb200

using namespace affb200;

extern "C" int64_t affb200_weight_net_work_floats(int64_t P) { return P > 0 ? int64_t(wn_grid(P)) * kWnSums : 0; }

extern "C" int affb200_weight_net_fwd(const float *feat, const float *w, const float *b, const float *gamma,
                                      const float *beta, float eps, int64_t P, int compute_dtype, float *out,
                                      void *stream) {
    if (int e = wn_check("weight_net_fwd", P, compute_dtype)) return e;
    if (P == 0) return AFFB200_OK;
    if (!feat || !w || !b || !gamma || !beta || !out || !aligned16(out)) {
        set_error("weight_net_fwd: null or misaligned pointer");
        return AFFB200_E_ARG;
    }
    auto st = (cudaStream_t)stream;
    const int grid = wn_grid(P);
    if (compute_dtype == AFFB200_BF16)
        weight_net_fwd_kernel<__nv_bfloat16><<<grid, kWnT, 0, st>>>(feat, w, b, gamma, beta, eps, int(P), out);
    else if (compute_dtype == AFFB200_F16)
        weight_net_fwd_kernel<__half><<<grid, kWnT, 0, st>>>(feat, w, b, gamma, beta, eps, int(P), out);
    else
        weight_net_fwd_kernel<float><<<grid, kWnT, 0, st>>>(feat, w, b, gamma, beta, eps, int(P), out);
    return check_launch("weight_net_fwd");
}

// grads[32] = dW[4][5], db[4], dgamma[4], dbeta[4] (overwritten); work = affb200_weight_net_work_floats(P) floats
extern "C" int affb200_weight_net_bwd(const float *feat, const float *w, const float *b, const float *gamma,
                                      const float *beta, float eps, int64_t P, int compute_dtype, const float *d_out,
                                      float *grads, float *work, void *stream) {
    if (int e = wn_check("weight_net_bwd", P, compute_dtype)) return e;
    auto st = (cudaStream_t)stream;
    if (!grads) { set_error("weight_net_bwd: null pointer"); return AFFB200_E_ARG; }
    if (P == 0) {
        cudaMemsetAsync(grads, 0, kWnSums * 4, st);
        return check_launch("weight_net_bwd memset", 0);
    }
    if (!feat || !w || !b || !gamma || !beta || !d_out || !work || !aligned16(d_out)) {
        set_error("weight_net_bwd: null or misaligned pointer");
        return AFFB200_E_ARG;
    }
    const int grid = wn_grid(P);
    if (compute_dtype == AFFB200_BF16)
        weight_net_bwd_kernel<__nv_bfloat16><<<grid, kWnT, 0, st>>>(feat, w, b, gamma, beta, eps, int(P), d_out, work);
    else if (compute_dtype == AFFB200_F16)
        weight_net_bwd_kernel<__half><<<grid, kWnT, 0, st>>>(feat, w, b, gamma, beta, eps, int(P), d_out, work);
    else
        weight_net_bwd_kernel<float><<<grid, kWnT, 0, st>>>(feat, w, b, gamma, beta, eps, int(P), d_out, work);
    if (int e = check_launch("weight_net_bwd")) return e;
    return launch_col_sum_finish(work, grid, kWnSums, grads, 0, st);
}
