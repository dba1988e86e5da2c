// common.cuh — shared device helpers for libaffb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/affb200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libaffb200 targets sm_100a (B200) only"
#endif

// The dynamic shared memory of every kernel, declared at global scope so that inline PTX can name it: `mov.u32 r,
// affb200_dsm` is the array's address in the shared window, a link-time constant.  Going through
// __cvta_generic_to_shared(pointer) instead makes ptxas form the generic address first (S2R SR_CgaCtaId + LEA) and,
// under register pressure, rebuild it in front of every store of an unrolled loop.
extern __shared__ __align__(1024) unsigned char affb200_dsm[];

namespace affb200 {

__device__ __forceinline__ uint32_t dsm_addr() {
    uint32_t a;
    asm("mov.u32 %0, affb200_dsm;" : "=r"(a));
    return a;
}

void set_error(const char *fmt, ...);
int check_launch(const char *what, int n_kernels = 1);  // also counts kernels enqueued (affb200_launch_count)
void count_launches(int n);

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs
constexpr int kUMaxDense = AFFB200_PLAN_UMAX;

// ---- element traits: storage type T, accumulator type, 16-byte vector width ------------------------------
template <typename T> struct Elem;
template <> struct Elem<float> {
    using Acc = float;
    static constexpr int VEC = 4;
    __device__ __forceinline__ static float to_acc(float x) { return x; }
    __device__ __forceinline__ static float from_acc(float x) { return x; }
};
template <> struct Elem<double> {
    using Acc = double;
    static constexpr int VEC = 2;
    __device__ __forceinline__ static double to_acc(double x) { return x; }
    __device__ __forceinline__ static double from_acc(double x) { return x; }
};
template <> struct Elem<__half> {
    using Acc = float;
    static constexpr int VEC = 8;
    __device__ __forceinline__ static float to_acc(__half x) { return __half2float(x); }
    __device__ __forceinline__ static __half from_acc(float x) { return __float2half_rn(x); }
};
template <> struct Elem<__nv_bfloat16> {
    using Acc = float;
    static constexpr int VEC = 8;
    __device__ __forceinline__ static float to_acc(__nv_bfloat16 x) { return __bfloat162float(x); }
    __device__ __forceinline__ static __nv_bfloat16 from_acc(float x) { return __float2bfloat16_rn(x); }
};

// ---- W-wide loads/stores; W == Elem<T>::VEC is one 128-bit access, W == 1 is the scalar fallback -----------
template <typename T, int W> struct Pack {
    typename Elem<T>::Acc v[W];
};

template <typename T, int W>
__device__ __forceinline__ void load_vec(const T *__restrict__ p, typename Elem<T>::Acc (&out)[W]) {
    if constexpr (W == 1) {
        out[0] = Elem<T>::to_acc(__ldg(p));
    } else {
        static_assert(W * sizeof(T) == 16, "vector path is 128-bit");
        uint4 raw = __ldg(reinterpret_cast<const uint4 *>(p));
        const T *e = reinterpret_cast<const T *>(&raw);
#pragma unroll
        for (int t = 0; t < W; ++t) out[t] = Elem<T>::to_acc(e[t]);
    }
}

template <typename T, int W>
__device__ __forceinline__ void store_vec(T *__restrict__ p, const typename Elem<T>::Acc (&in)[W]) {
    if constexpr (W == 1) {
        p[0] = Elem<T>::from_acc(in[0]);
    } else {
        static_assert(W * sizeof(T) == 16, "vector path is 128-bit");
        uint4 raw;
        T *e = reinterpret_cast<T *>(&raw);
#pragma unroll
        for (int t = 0; t < W; ++t) e[t] = Elem<T>::from_acc(in[t]);
        *reinterpret_cast<uint4 *>(p) = raw;
    }
}

// 8 consecutive elements (16-byte aligned for 2-byte types, 32-byte for fp32) <-> fp32 registers
template <typename T> __device__ __forceinline__ void load8f(const T *p, float (&o)[8]) {
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
template <typename T> __device__ __forceinline__ void store8f(T *p, const float (&v)[8]) {
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

// streaming (evict-first) 128-bit store for write-once outputs
__device__ __forceinline__ void st_cs_u4(void *p, uint4 v) {
    asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- division of a non-negative int32 by a runtime-constant divisor (host precomputes the magic) ---------
struct FastDiv {
    uint32_t d, mul, shr;
    __host__ FastDiv() : d(1), mul(0), shr(0) {}
    __host__ explicit FastDiv(uint32_t div) : d(div) {
        // round-up method, exact for all n < 2^31 (Granlund-Montgomery)
        if (div == 1) { mul = 0; shr = 0; return; }
        uint32_t l = 0;
        while ((1u << l) < div) ++l;  // ceil(log2 d)
        uint64_t m = ((uint64_t(1) << (31 + l)) + div - 1) / div;  // ceil(2^(31+l)/d) fits in 32 bits
        mul = uint32_t(m);
        shr = l;
    }
    __device__ __forceinline__ uint32_t div(uint32_t n) const {
        if (d == 1) return n;
        return __umulhi(n << 1, mul) >> shr;  // umulhi(2n, mul) = (n*mul) >> 31, then >> l
    }
};

__host__ __device__ __forceinline__ int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

template <typename F> inline int dispatch_dtype(int dtype, F &&f) {
    switch (dtype) {
        case AFFB200_F32: return f(float{});
        case AFFB200_F16: return f(__half{});
        case AFFB200_BF16: return f(__nv_bfloat16{});
        case AFFB200_F64: return f(double{});
        default: set_error("unsupported dtype enum %d", dtype); return AFFB200_E_ARG;
    }
}

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// dense-tile launchers (clusten_dense.cu), instantiated for __half and __nv_bfloat16
template <typename T>
int launch_gather_dot_dense(const T *x, const T *y, const affb200_nbhd_plan *pl, T *out, int64_t B, int64_t H,
                            int64_t Nq, int64_t Nk, int64_t M, int64_t c, cudaStream_t st);
template <typename T>
int launch_gather_wsum_dense(const T *w, const T *y, const affb200_nbhd_plan *pl, T *out, int64_t B, int64_t H,
                             int64_t Nq, int64_t Nk, int64_t M, int64_t c, cudaStream_t st);
template <typename T>
int launch_scatter_wsum_dense(const T *w, const T *x, const affb200_nbhd_plan *pl, T *dy, int64_t B, int64_t H,
                              int64_t Nq, int64_t Nk, int64_t M, int64_t c, cudaStream_t st);

// x / dy addressed as [b][h][row] with element strides {sB, sH, sN}; dy = alpha * sum(...)
template <typename T>
int launch_scatter_wsum_dense_strided(const T *w, const T *x, const affb200_nbhd_plan *pl, T *dy, int64_t B, int64_t H,
                                      int64_t Nq, int64_t Nk, int64_t M, int64_t c, const int64_t *x_str,
                                      const int64_t *y_str, float alpha, cudaStream_t st);
// d_table[p][h] from d_vals [B][H][N*M] indexed by packed uint16 bias indices (attn_glue.cu)
int launch_table_grad_u16(const void *d_vals, const uint16_t *pe, int64_t B, int64_t H, int64_t NM, int64_t P,
                          float *d_table, int dtype, cudaStream_t st);

}  // namespace affb200
