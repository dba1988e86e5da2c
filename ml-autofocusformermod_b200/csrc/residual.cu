// residual.cu — residual connection with stochastic depth as one pass:  out = a + x * s[b]   (a optional)
// Reference: `feat = shortcut + self.drop_path(feat)` (models/aff_transformer.py:224-229; timm DropPath scales the
// kept samples of the batch by 1/(1-p) and zeroes the dropped ones).  Stock PyTorch runs a broadcast multiply
// (scalar-indexed, not vectorised) and an add; the backward another broadcast multiply.  Here the per-sample factor
// s[b] * s_scale (fp32: the 0/1 keep draw times 1/(1-p)) is applied inside the residual add with 16-byte accesses; the backward of the x branch is
// the same kernel without `a`.
#include "common.cuh"
#include <algorithm>

namespace affb200 {

template <typename T>
__global__ void __launch_bounds__(256)
scale_add_kernel(const T *__restrict__ a, const T *__restrict__ x, const float *__restrict__ s, float s_scale,
                 T *__restrict__ out, int64_t per_sample_vec, int64_t total_vec) {
    constexpr int V = 16 / sizeof(T);
    for (int64_t i = int64_t(blockIdx.x) * 256 + threadIdx.x; i < total_vec; i += int64_t(gridDim.x) * 256) {
        const float sc = __ldg(s + i / per_sample_vec) * s_scale;
        float xv[V], av[V];
        load_vec<T, V>(x + i * V, xv);
        if (a) {
            load_vec<T, V>(a + i * V, av);
#pragma unroll
            for (int k = 0; k < V; ++k)  // product rounded to T, then the add: no fused multiply-add, as the two-op reference
                xv[k] = __fadd_rn(Elem<T>::to_acc(Elem<T>::from_acc(__fmul_rn(xv[k], sc))), av[k]);
        } else {
#pragma unroll
            for (int k = 0; k < V; ++k) xv[k] *= sc;
        }
        store_vec<T, V>(out + i * V, xv);
    }
}

// y[r][c] = x[r][c] + bias[c] over channels-last rows, the sum rounded to T (bias rounded to T first: the conv bias
// as autocast hands it to the stock add).  C/V consecutive threads cover a row; the bias chunk stays in registers.
template <typename T>
__global__ void __launch_bounds__(256)
bias_rows_kernel(const T *__restrict__ x, const float *__restrict__ bias, T *__restrict__ y, int chunks, int64_t total_vec) {
    constexpr int V = 16 / sizeof(T);
    const int64_t stride = int64_t(gridDim.x) * 256;     // host makes it a multiple of `chunks`
    int64_t i = int64_t(blockIdx.x) * 256 + threadIdx.x;
    const int ch = int(i % chunks);
    float bv[V];
#pragma unroll
    for (int k = 0; k < V; ++k) bv[k] = Elem<T>::to_acc(Elem<T>::from_acc(__ldg(bias + ch * V + k)));
    for (; i < total_vec; i += stride) {
        float xv[V];
        load_vec<T, V>(x + i * V, xv);
#pragma unroll
        for (int k = 0; k < V; ++k) xv[k] += bv[k];
        store_vec<T, V>(y + i * V, xv);
    }
}

}  // namespace affb200

using namespace affb200;

extern "C" int affb200_bias_rows(const void *x, const float *bias, void *y, int64_t R, int64_t C, int dtype, void *stream) {
    if (R < 0 || C <= 0) { set_error("bias_rows: bad size"); return AFFB200_E_ARG; }
    if (R == 0) return AFFB200_OK;
    if (!x || !bias || !y || !aligned16(x) || !aligned16(y)) { set_error("bias_rows: null or misaligned pointer"); return AFFB200_E_ARG; }
    return dispatch_dtype(dtype, [&](auto tag) -> int {
        using T = decltype(tag);
        if constexpr (sizeof(T) == 8) {
            set_error("bias_rows: f32 / f16 / bf16 only");
            return AFFB200_E_ARG;
        } else {
            constexpr int V = 16 / sizeof(T);
            if (C % V != 0 || C / V > 256 || 256 % (C / V) != 0) {
                set_error("bias_rows: C / %d must divide 256", V);
                return AFFB200_E_ARG;
            }
            const int chunks = int(C / V);
            const int64_t tv = R * chunks;
            const unsigned grid = unsigned(std::min<int64_t>((tv + 255) / 256, int64_t(kNumSMs) * 16));
            bias_rows_kernel<T><<<grid, 256, 0, (cudaStream_t)stream>>>((const T *)x, bias, (T *)y, chunks, tv);
            return check_launch("bias_rows");
        }
    });
}

extern "C" int affb200_scale_add(const void *a, const void *x, const float *s, float s_scale, void *out, int64_t B,
                                 int64_t per_sample, int dtype, void *stream) {
    if (B < 0 || per_sample < 0) { set_error("scale_add: negative size"); return AFFB200_E_ARG; }
    if (B * per_sample == 0) return AFFB200_OK;
    if (!x || !s || !out) { set_error("scale_add: null pointer"); return AFFB200_E_ARG; }
    return dispatch_dtype(dtype, [&](auto tag) -> int {
        using T = decltype(tag);
        if constexpr (sizeof(T) == 8) {
            set_error("scale_add: f32 / f16 / bf16 only");
            return AFFB200_E_ARG;
        } else {
            constexpr int V = 16 / sizeof(T);
            if (per_sample % V != 0 || !aligned16(x) || !aligned16(out) || (a && !aligned16(a))) {
                set_error("scale_add: per-sample size must be a multiple of %d elements and pointers 16-byte aligned", V);
                return AFFB200_E_ARG;
            }
            const int64_t pv = per_sample / V, tv = B * pv;
            const unsigned grid = unsigned(std::min<int64_t>((tv + 255) / 256, int64_t(kNumSMs) * 16));
            scale_add_kernel<T><<<grid, 256, 0, (cudaStream_t)stream>>>((const T *)a, (const T *)x, s, s_scale, (T *)out, pv, tv);
            return check_launch("scale_add");
        }
    });
}
