// conv_stem.cu — the first convolution of the patch embedding: 3 -> Cout channels, 3x3, stride 2, padding 1, on
// channels-last images (reference models/aff_transformer.py:524,541: `self.proj1 = nn.Conv2d(in_chans, embed_dim // 2,
// 3, 2, 1)`).  The library kernel chosen for this shape (K = 27, three input channels padded to a tensor-core
// friendly count by an extra pass) runs at a tenth of the bandwidth bound: 433 µs for 77 MB in / 205 MB out at batch
// 256.  Here a CTA owns four output rows of one image: the nine input rows it needs are staged once in shared memory
// as fp32 (with the left padding pixel), the 27 x Cout weights likewise, and a thread produces 4 rows x 16 channels of
// one output column: per tap one broadcast weight load serves the four rows.  fp32 accumulate, one rounding.
#include "common.cuh"
#include <algorithm>

namespace affb200 {

constexpr int kCsRows = 4;     // output rows per CTA
constexpr int kCsCh = 16;      // output channels per thread

template <typename T, int MAXT>
__global__ void __launch_bounds__(MAXT)
conv_stem_fwd_kernel(const T *__restrict__ x, const float *__restrict__ w, T *__restrict__ y, int H, int W, int Ho,
                     int Wo, int Cout) {
    extern __shared__ __align__(16) float cs_smem[];
    const int row_f = (W + 1) * 3;                     // floats per staged input row (pixel -1 first)
    float *ws = cs_smem;                               // [27][Cout]  (k = (kh*3 + kw)*3 + ci)
    float *in_s = cs_smem + 27 * Cout;                 // [2*kCsRows + 1][row_f]
    const int b = blockIdx.y, oy0 = blockIdx.x * kCsRows;
    const int nthr = blockDim.x;
    // weights: OIHW fp32 master [Cout][3][3][3] -> [k][o], rounded to T as the autocast copy is
    for (int i = threadIdx.x; i < 27 * Cout; i += nthr) {
        const int o = i / 27, r = i - o * 27;          // r = ci*9 + kh*3 + kw
        const int ci = r / 9, kk = r - ci * 9;
        ws[(kk * 3 + ci) * Cout + o] = Elem<T>::to_acc(Elem<T>::from_acc(__ldg(w + i)));
    }
    // input rows 2*oy0 - 1 .. 2*oy0 + 2*kCsRows - 1
    const int n_rows = 2 * kCsRows + 1, vec_per_row = W * 3 * int(sizeof(T)) / 16, V = 16 / int(sizeof(T));
    for (int i = threadIdx.x; i < n_rows * (vec_per_row + 1); i += nthr) {
        const int r = i / (vec_per_row + 1), v = i - r * (vec_per_row + 1);
        const int iy = 2 * oy0 - 1 + r;
        float *dst = in_s + r * row_f;
        if (v == vec_per_row) {                        // the padding pixel at ix = -1
            dst[0] = dst[1] = dst[2] = 0.f;
            continue;
        }
        float vals[16 / sizeof(T)];
        if (iy >= 0 && iy < H) {
            load_vec<T, 16 / sizeof(T)>(x + ((int64_t(b) * H + iy) * W) * 3 + int64_t(v) * V, vals);
        } else {
#pragma unroll
            for (int k = 0; k < V; ++k) vals[k] = 0.f;
        }
#pragma unroll
        for (int k = 0; k < V; ++k) dst[3 + v * V + k] = vals[k];
    }
    __syncthreads();
    const int hg = Cout / kCsCh;
    const int ox = threadIdx.x / hg, h = threadIdx.x - ox * hg;
    if (ox >= Wo) return;
    float acc[kCsRows][kCsCh];
#pragma unroll
    for (int r = 0; r < kCsRows; ++r)
#pragma unroll
        for (int j = 0; j < kCsCh; ++j) acc[r][j] = 0.f;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
#pragma unroll
            for (int ci = 0; ci < 3; ++ci) {
                const int k = (kh * 3 + kw) * 3 + ci;
                float wv[kCsCh];
#pragma unroll
                for (int q = 0; q < kCsCh / 4; ++q) {
                    const float4 t = *reinterpret_cast<const float4 *>(ws + k * Cout + h * kCsCh + q * 4);
                    wv[q * 4] = t.x; wv[q * 4 + 1] = t.y; wv[q * 4 + 2] = t.z; wv[q * 4 + 3] = t.w;
                }
#pragma unroll
                for (int r = 0; r < kCsRows; ++r) {
                    const float v = in_s[(2 * r + kh) * row_f + (2 * ox + kw) * 3 + ci];
#pragma unroll
                    for (int j = 0; j < kCsCh; ++j) acc[r][j] = fmaf(v, wv[j], acc[r][j]);
                }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < kCsRows; ++r) {
        const int oy = oy0 + r;
        if (oy >= Ho) break;
        T *dst = y + ((int64_t(b) * Ho + oy) * Wo + ox) * Cout + h * kCsCh;
        constexpr int V2 = 16 / sizeof(T);
#pragma unroll
        for (int q = 0; q < kCsCh / V2; ++q) {
            float o[V2];
#pragma unroll
            for (int k = 0; k < V2; ++k) o[k] = acc[r][q * V2 + k];
            store_vec<T, V2>(dst + q * V2, o);
        }
    }
}

}  // namespace affb200

using namespace affb200;

// y[B][Ho][Wo][Cout] = conv3x3/s2/p1(x[B][H][W][3], w[Cout][3][3][3]) without bias; x / y of `dtype` (f16/bf16),
// w fp32 (rounded to dtype inside); H, W even, Cout a multiple of 16, (W/2) * (Cout/16) <= 256 threads (128 registers each).
extern "C" int affb200_conv_stem_fwd(const void *x, const float *w, void *y, int64_t B, int64_t H, int64_t W,
                                     int64_t Cout, int dtype, void *stream) {
    if (B < 0 || H <= 0 || W <= 0 || H % 2 || W % 2 || Cout <= 0 || Cout % 16 || (W / 2) * (Cout / 16) > 256 ||
        B > 65535 || (W * 3 * 2) % 16 != 0) {
        set_error("conv_stem_fwd: needs even H, W with W*6 a multiple of 16, Cout a multiple of 16, (W/2)*(Cout/16) <= 256");
        return AFFB200_E_ARG;
    }
    if (dtype != AFFB200_BF16 && dtype != AFFB200_F16) { set_error("conv_stem_fwd: f16 / bf16 only"); return AFFB200_E_ARG; }
    if (B == 0) return AFFB200_OK;
    if (!x || !w || !y || !aligned16(x) || !aligned16(y)) { set_error("conv_stem_fwd: null or misaligned pointer"); return AFFB200_E_ARG; }
    const int Ho = int(H / 2), Wo = int(W / 2);
    const size_t smem = (size_t(27) * Cout + size_t(2 * kCsRows + 1) * (W + 1) * 3) * 4;
    if (smem > 200 * 1024) { set_error("conv_stem_fwd: image too wide"); return AFFB200_E_ARG; }
    const dim3 grid{unsigned((Ho + kCsRows - 1) / kCsRows), unsigned(B), 1u};
    const int threads = ((Wo * int(Cout / 16)) + 31) / 32 * 32;
    auto st = (cudaStream_t)stream;
    auto launch = [&](auto kern, auto tag) -> int {
        using T = decltype(tag);
        if (smem > 48 * 1024 && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)) != cudaSuccess)
            return check_launch("cudaFuncSetAttribute", 0);
        kern<<<grid, threads, smem, st>>>((const T *)x, w, (T *)y, int(H), int(W), Ho, Wo, int(Cout));
        return AFFB200_OK;
    };
    int e;
    if (dtype == AFFB200_BF16) e = launch(conv_stem_fwd_kernel<__nv_bfloat16, 256>, __nv_bfloat16{});
    else e = launch(conv_stem_fwd_kernel<__half, 256>, __half{});
    if (e) return e;
    return check_launch("conv_stem_fwd");
}
