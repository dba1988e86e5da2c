// attn_glue.cu — the relative-position bias of ClusterAttention as one in-place pass and its table gradient.
//
// Reference (models/aff_transformer.py:126-137): pe_table = Linear(5->H)(pre_table)  [P x H];
//   pos_embed = pe_table.gather(pe_idx) -> [B,N,M,H] -> permute -> [B,H,N,M];  attn = attn + pos_embed;
//   attn = attn + (1 - mask) * (-100).
// Forward here: attn[b,h,i,j] += table[pe_idx[b,i,j], h] (+ -100 where mask == 0), in place on the QK output, one
// read and one write of attn, pe_idx/mask read once for all heads.
// Backward: d_attn passes through; d_table[p,h] = sum over {(b,i,j): pe_idx = p} of d_attn[b,h,i,j].  The reference
// gets this from autograd's scatter-add / index_put (one global atomic or a sort per element).  Here each CTA keeps a
// private fp32 copy of one head's table column in shared memory (P floats, 49 KB at 224x224), accumulates its share
// of the (b,i,j) range with shared-memory atomics, and flushes only the entries it touched — neighbourhoods are
// local, so a CTA touches a few hundred distinct offsets — with one global atomic each.
#include "common.cuh"
#include <algorithm>

namespace affb200 {

constexpr int kBT = 256;

template <typename T>
__global__ void __launch_bounds__(kBT)
attn_bias_add_kernel(T *__restrict__ attn, const float *__restrict__ table_f32, const int64_t *__restrict__ pe_idx,
                     const int64_t *__restrict__ mask, int H, int64_t NM, int64_t total /* B*N*M */) {
    const int64_t t = int64_t(blockIdx.x) * kBT + threadIdx.x;
    if (t >= total) return;
    const int64_t b = t / NM;
    const int64_t r = t - b * NM;
    const int64_t p = __ldg(pe_idx + t);
    const float mterm = (mask && __ldg(mask + t) == 0) ? -100.f : 0.f;
    for (int h = 0; h < H; ++h) {
        T *a = attn + (b * H + h) * NM + r;
        // same rounding sequence as the reference's two adds in the tensor dtype
        float v = Elem<T>::to_acc(*a);
        v = Elem<T>::to_acc(Elem<T>::from_acc(v + __ldg(table_f32 + p * H + h)));
        if (mask) v = Elem<T>::to_acc(Elem<T>::from_acc(v + mterm));
        *a = Elem<T>::from_acc(v);
    }
}

// grid (chunks, H); dynamic smem: P floats + P/32 touched-bitmap words
template <typename T>
__global__ void __launch_bounds__(kBT)
attn_bias_bwd_kernel(const T *__restrict__ d_attn, const int64_t *__restrict__ pe_idx, int H, int64_t NM,
                     int64_t total, int P, int channel_last, float *__restrict__ d_table /* [P][H], pre-zeroed */) {
    extern __shared__ float hist[];
    uint32_t *touched = reinterpret_cast<uint32_t *>(hist + P);
    const int nw = (P + 31) >> 5;
    const int h = blockIdx.y;
    for (int p = threadIdx.x; p < P; p += kBT) hist[p] = 0.f;
    for (int w = threadIdx.x; w < nw; w += kBT) touched[w] = 0u;
    __syncthreads();
    const int64_t per = (total + gridDim.x - 1) / gridDim.x;
    const int64_t t0 = int64_t(blockIdx.x) * per, t1 = min(total, t0 + per);
    for (int64_t t = t0 + threadIdx.x; t < t1; t += kBT) {
        const int64_t b = t / NM;
        const int64_t r = t - b * NM;
        const int p = int(__ldg(pe_idx + t));
        // [B][H][N*M] (attention logits) or [B][N*M][H] (PointConv weights, aff_transformer.py:349)
        const float g = Elem<T>::to_acc(__ldg(d_attn + (channel_last ? t * H + h : (b * H + h) * NM + r)));
        atomicAdd(hist + p, g);
        atomicOr(touched + (p >> 5), 1u << (p & 31));
    }
    __syncthreads();
    for (int w = threadIdx.x; w < nw; w += kBT) {
        uint32_t bits = touched[w];
        while (bits) {
            const int bit = __ffs(bits) - 1;
            bits &= bits - 1;
            const int p = (w << 5) + bit;
            atomicAdd(d_table + int64_t(p) * H + h, hist[p]);
        }
    }
}

// fallback for tables that do not fit in shared memory (MixRes: 1023^2 entries): direct global atomics
template <typename T>
__global__ void __launch_bounds__(kBT)
attn_bias_bwd_global_kernel(const T *__restrict__ d_attn, const int64_t *__restrict__ pe_idx, int H, int64_t NM,
                            int64_t total, int channel_last, float *__restrict__ d_table) {
    const int64_t t = int64_t(blockIdx.x) * kBT + threadIdx.x;
    if (t >= total) return;
    const int64_t b = t / NM;
    const int64_t r = t - b * NM;
    const int64_t p = __ldg(pe_idx + t);
    for (int h = 0; h < H; ++h)
        atomicAdd(d_table + p * H + h,
                  Elem<T>::to_acc(__ldg(d_attn + (channel_last ? t * H + h : (b * H + h) * NM + r))));
}

// same histogram scheme, indices taken from the fused-attention plan's packed uint16 form (bit 15 = masked slot)
constexpr int kTG = 512;
template <typename T>
__global__ void __launch_bounds__(kTG)
table_grad_u16_kernel(const T *__restrict__ d_vals, const uint16_t *__restrict__ pe, int H, int64_t NM, int64_t total,
                      int P, float *__restrict__ d_table) {
    extern __shared__ float hist[];
    const int h = blockIdx.y;
    for (int p = threadIdx.x; p < P; p += kTG) hist[p] = 0.f;
    __syncthreads();
    // chunk boundaries are multiples of 8 entries so that every thread handles 8 consecutive entries with one
    // 16-byte load of the indices and one of the values (NM is a multiple of 8 whenever M is)
    const int64_t groups = (total + 7) / 8;
    const int64_t per = (groups + gridDim.x - 1) / gridDim.x;
    const int64_t g0 = int64_t(blockIdx.x) * per, g1 = min(groups, g0 + per);
    const bool vec = (NM % 8 == 0) && sizeof(T) == 2;
    for (int64_t g = g0 + threadIdx.x; g < g1; g += kTG) {
        const int64_t t = g * 8;
        if (vec) {
            const int64_t b = t / NM;
            const int64_t r = t - b * NM;
            const uint4 pv = __ldg(reinterpret_cast<const uint4 *>(pe + t));
            const uint4 dv = __ldg(reinterpret_cast<const uint4 *>(d_vals + (b * H + h) * NM + r));
            const uint16_t *p8 = reinterpret_cast<const uint16_t *>(&pv);
            const T *d8 = reinterpret_cast<const T *>(&dv);
#pragma unroll
            for (int k = 0; k < 8; ++k) atomicAdd(hist + (p8[k] & 0x7fffu), Elem<T>::to_acc(d8[k]));
        } else {
            for (int k = 0; k < 8 && t + k < total; ++k) {
                const int64_t b = (t + k) / NM;
                const int64_t r = (t + k) - b * NM;
                atomicAdd(hist + (__ldg(pe + t + k) & 0x7fffu), Elem<T>::to_acc(__ldg(d_vals + (b * H + h) * NM + r)));
            }
        }
    }
    __syncthreads();
    for (int p = threadIdx.x; p < P; p += kTG) {
        const float v = hist[p];
        if (v != 0.f) atomicAdd(d_table + int64_t(p) * H + h, v);
    }
}

// The bias table of ClusterAttention is itself a Linear of a fixed feature table, table = feat·W^T + b
// (reference aff_transformer.py:78,125: pos_embed(pre_table), 5 features per relative offset), and nothing but W and
// b receive its gradient.  So instead of the [P][H] histogram d_table (one shared-memory atomic per slot and head),
// the gradient goes straight to the 6 numbers per head
//     G[h][f] = sum over slots of d_vals[slot][h] · feat8[pe[slot]][f]         f < 8
// (feat8 rows padded to 8 floats, a column of ones gives db), accumulated in registers: dW = G[:, :F], db = G[:, F].
// grid (chunks, head groups of HG); part[chunk][H][8]; a second kernel adds the chunks in a fixed order.
template <typename T, int HG>
__global__ void __launch_bounds__(256)
pe_feat_grad_kernel(const T *__restrict__ d_vals, const uint16_t *__restrict__ pe, const float *__restrict__ feat8, int H,
                    int64_t NM, int64_t total, float *__restrict__ part) {
    __shared__ float red[8][HG][8];
    const int h0 = blockIdx.y * HG, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float acc[HG][6];
#pragma unroll
    for (int hh = 0; hh < HG; ++hh)
#pragma unroll
        for (int f = 0; f < 6; ++f) acc[hh][f] = 0.f;
    const int64_t groups = total / 8;   // NM % 8 == 0 (checked by the launcher)
    const int64_t per = (groups + gridDim.x - 1) / gridDim.x;
    const int64_t g0 = int64_t(blockIdx.x) * per, g1 = min(groups, g0 + per);
    // (image, offset in the image) of this thread's group, advanced without a 64-bit division per iteration
    int64_t b = 0, r = 0;
    if (g0 + threadIdx.x < g1) {
        b = ((g0 + threadIdx.x) * 8) / NM;
        r = (g0 + threadIdx.x) * 8 - b * NM;
    }
    for (int64_t g = g0 + threadIdx.x; g < g1; g += 256, r += 256 * 8) {
        while (r >= NM) { r -= NM; ++b; }
        const int64_t t = g * 8;
        const uint4 pv = __ldg(reinterpret_cast<const uint4 *>(pe + t));
        uint4 dv[HG];
#pragma unroll
        for (int hh = 0; hh < HG; ++hh) {
            dv[hh] = make_uint4(0u, 0u, 0u, 0u);
            if (h0 + hh < H) dv[hh] = __ldg(reinterpret_cast<const uint4 *>(d_vals + (b * H + h0 + hh) * NM + r));
        }
        const uint16_t *p8 = reinterpret_cast<const uint16_t *>(&pv);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float *fr = feat8 + size_t(p8[k] & 0x7fffu) * 8;
            const float4 fa = __ldg(reinterpret_cast<const float4 *>(fr));
            const float2 fb = __ldg(reinterpret_cast<const float2 *>(fr + 4));
#pragma unroll
            for (int hh = 0; hh < HG; ++hh) {
                const float d = Elem<T>::to_acc(reinterpret_cast<const T *>(&dv[hh])[k]);
                acc[hh][0] = fmaf(d, fa.x, acc[hh][0]); acc[hh][1] = fmaf(d, fa.y, acc[hh][1]);
                acc[hh][2] = fmaf(d, fa.z, acc[hh][2]); acc[hh][3] = fmaf(d, fa.w, acc[hh][3]);
                acc[hh][4] = fmaf(d, fb.x, acc[hh][4]); acc[hh][5] = fmaf(d, fb.y, acc[hh][5]);
            }
        }
    }
#pragma unroll
    for (int hh = 0; hh < HG; ++hh)
#pragma unroll
        for (int f = 0; f < 6; ++f) {
            float a = acc[hh][f];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
            if (lane == 0) red[warp][hh][f] = a;
        }
    __syncthreads();
    if (threadIdx.x < HG * 8) {
        const int hh = threadIdx.x / 8, f = threadIdx.x % 8;
        float a = 0.f;
        if (f < 6)
#pragma unroll
            for (int w = 0; w < 8; ++w) a += red[w][hh][f];
        if (h0 + hh < H) part[(int64_t(blockIdx.x) * H + h0 + hh) * 8 + f] = a;
    }
}

// out[e] = sum over chunks, in a fixed order: a block owns 32 outputs, 32 threads per output stride over the chunks
// (a single thread walking the ~600 chunks is a 27 us chain of dependent loads), then a fixed-order sum of the 32
__global__ void __launch_bounds__(1024)
pe_feat_grad_finish_kernel(const float *__restrict__ part, int n_chunks, int H8, float *__restrict__ out) {
    __shared__ float red[32][33];
    const int el = threadIdx.x & 31, sub = threadIdx.x >> 5;
    const int e = blockIdx.x * 32 + el;
    float a = 0.f;
    if (e < H8)
        for (int c = sub; c < n_chunks; c += 32) a += part[int64_t(c) * H8 + e];
    red[sub][el] = a;
    __syncthreads();
    if (sub == 0 && e < H8) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 32; ++k) t += red[k][el];
        out[e] = t;
    }
}

// out[t][:] = table[idx[t]][:] for rows of ROWB bytes (4, 8 or 16): one thread per gathered row.  The PointConv weights
// of ClusterMerging are rows of 4 floats (reference aff_transformer.py:347-349); the stock indexing kernel spends a
// whole CTA on each 16-byte row.
template <typename V>
__global__ void __launch_bounds__(256)
table_gather_kernel(const V *__restrict__ table, const int64_t *__restrict__ idx, V *__restrict__ out, int64_t total) {
    const int64_t t = int64_t(blockIdx.x) * 256 + threadIdx.x;
    if (t < total) out[t] = __ldg(table + __ldg(idx + t));
}
template <typename T>
__global__ void __launch_bounds__(256)
table_gather_generic_kernel(const T *__restrict__ table, const int64_t *__restrict__ idx, T *__restrict__ out,
                            int64_t total, int IC) {
    const int64_t e = int64_t(blockIdx.x) * 256 + threadIdx.x;
    if (e >= total * IC) return;
    const int64_t t = e / IC;
    out[e] = __ldg(table + __ldg(idx + t) * IC + (e - t * IC));
}

int launch_table_grad_u16(const void *d_vals, const uint16_t *pe, int64_t B, int64_t H, int64_t NM, int64_t P,
                          float *d_table, int dtype, cudaStream_t st) {
    if (cudaMemsetAsync(d_table, 0, size_t(P * H) * 4, st) != cudaSuccess) return check_launch("table_grad memset", 0);
    const int64_t total = B * NM;
    if (total == 0) return AFFB200_OK;
    const size_t smem = size_t(P) * 4;  // P <= 32768 -> <= 128 KB
    return dispatch_dtype(dtype, [&](auto tag) {
        using T = decltype(tag);
        auto kern = table_grad_u16_kernel<T>;
        if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        const int per_sm = smem > 100 * 1024 ? 1 : 2;
        int chunks = int(std::max<int64_t>(1, (int64_t(kNumSMs) * per_sm) / H));
        chunks = int(std::min<int64_t>(chunks, (total + 4095) / 4096));
        kern<<<dim3(unsigned(chunks), unsigned(H)), kTG, smem, st>>>((const T *)d_vals, pe, int(H), NM, total, int(P),
                                                                     d_table);
        return check_launch("table_grad_u16");
    });
}

}  // namespace affb200

using namespace affb200;

extern "C" int affb200_attn_bias_add(void *attn, const float *table, const int64_t *pe_idx, const int64_t *mask,
                                     int64_t B, int64_t H, int64_t N, int64_t M, int dtype, void *stream) {
    if (B < 0 || H < 0 || N < 0 || M < 0 || H >= (1 << 16)) { set_error("attn_bias_add: bad dimensions"); return AFFB200_E_ARG; }
    const int64_t total = B * N * M;
    if (total * H == 0) return AFFB200_OK;
    if (!attn || !table || !pe_idx) { set_error("attn_bias_add: null pointer"); return AFFB200_E_ARG; }
    return dispatch_dtype(dtype, [&](auto tag) {
        using T = decltype(tag);
        attn_bias_add_kernel<T><<<unsigned((total + kBT - 1) / kBT), kBT, 0, (cudaStream_t)stream>>>(
            (T *)attn, table, pe_idx, mask, int(H), N * M, total);
        return check_launch("attn_bias_add");
    });
}

extern "C" int affb200_attn_bias_bwd(const void *d_attn, const int64_t *pe_idx, int64_t B, int64_t H, int64_t N,
                                     int64_t M, int64_t P, int channel_last, float *d_table, int dtype, void *stream) {
    if (B < 0 || H < 0 || N < 0 || M < 0 || P < 0 || H >= 65536 || P >= (int64_t(1) << 31)) {
        set_error("attn_bias_bwd: bad dimensions");
        return AFFB200_E_ARG;
    }
    auto st = (cudaStream_t)stream;
    if (P * H == 0) return AFFB200_OK;
    if (!d_table) { set_error("attn_bias_bwd: null pointer"); return AFFB200_E_ARG; }
    if (cudaMemsetAsync(d_table, 0, size_t(P * H) * 4, st) != cudaSuccess) return check_launch("attn_bias_bwd memset", 0);
    const int64_t total = B * N * M;
    if (total == 0) return AFFB200_OK;
    if (!d_attn || !pe_idx) { set_error("attn_bias_bwd: null pointer"); return AFFB200_E_ARG; }
    const size_t smem = size_t(P) * 4 + size_t((P + 31) / 32) * 4;
    return dispatch_dtype(dtype, [&](auto tag) {
        using T = decltype(tag);
        if (smem <= 200 * 1024) {
            auto kern = attn_bias_bwd_kernel<T>;
            if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
            const int per_sm = smem > 100 * 1024 ? 1 : 2;
            int chunks = int(std::max<int64_t>(1, (int64_t(kNumSMs) * per_sm) / H));
            chunks = int(std::min<int64_t>(chunks, (total + 4095) / 4096));
            kern<<<dim3(unsigned(chunks), unsigned(H)), kBT, smem, st>>>((const T *)d_attn, pe_idx, int(H), N * M,
                                                                         total, int(P), channel_last, d_table);
        } else {
            attn_bias_bwd_global_kernel<T><<<unsigned((total + kBT - 1) / kBT), kBT, 0, st>>>(
                (const T *)d_attn, pe_idx, int(H), N * M, total, channel_last, d_table);
        }
        return check_launch("attn_bias_bwd");
    });
}

extern "C" int affb200_table_gather(const void *table, const int64_t *idx, void *out, int64_t total, int64_t IC,
                                    int elem_bytes, void *stream) {
    if (total < 0 || IC <= 0 || (elem_bytes != 2 && elem_bytes != 4 && elem_bytes != 8)) {
        set_error("table_gather: bad dimensions");
        return AFFB200_E_ARG;
    }
    if (total == 0) return AFFB200_OK;
    if (!table || !idx || !out) { set_error("table_gather: null pointer"); return AFFB200_E_ARG; }
    auto st = (cudaStream_t)stream;
    const int64_t rowb = IC * elem_bytes;
    const unsigned grid = unsigned((total + 255) / 256);
    const uintptr_t al = reinterpret_cast<uintptr_t>(table) | reinterpret_cast<uintptr_t>(out);
    if (rowb == 16 && (al & 15) == 0)
        table_gather_kernel<uint4><<<grid, 256, 0, st>>>((const uint4 *)table, idx, (uint4 *)out, total);
    else if (rowb == 8 && (al & 7) == 0)
        table_gather_kernel<uint2><<<grid, 256, 0, st>>>((const uint2 *)table, idx, (uint2 *)out, total);
    else if (rowb == 4 && (al & 3) == 0)
        table_gather_kernel<uint32_t><<<grid, 256, 0, st>>>((const uint32_t *)table, idx, (uint32_t *)out, total);
    else {
        const unsigned g2 = unsigned((total * IC + 255) / 256);
        if (elem_bytes == 2)
            table_gather_generic_kernel<uint16_t><<<g2, 256, 0, st>>>((const uint16_t *)table, idx, (uint16_t *)out, total, int(IC));
        else if (elem_bytes == 4)
            table_gather_generic_kernel<uint32_t><<<g2, 256, 0, st>>>((const uint32_t *)table, idx, (uint32_t *)out, total, int(IC));
        else
            table_gather_generic_kernel<uint64_t><<<g2, 256, 0, st>>>((const uint64_t *)table, idx, (uint64_t *)out, total, int(IC));
    }
    return check_launch("table_gather");
}

static int pe_feat_chunks(int64_t total) {
    return int(std::max<int64_t>(1, std::min<int64_t>(int64_t(kNumSMs) * 4, total / (8 * 256 * 2))));
}

extern "C" int64_t affb200_pe_feat_grad_work_floats(int64_t B, int64_t H, int64_t NM) {
    if (B <= 0 || H <= 0 || NM <= 0) return 0;
    return int64_t(pe_feat_chunks(B * NM)) * H * 8;
}

extern "C" int affb200_pe_feat_grad(const void *d_vals, const uint16_t *q_pe, const float *feat8, int64_t B, int64_t H,
                                    int64_t NM, float *out, float *work, int dtype, void *stream) {
    if (B <= 0 || H <= 0 || NM <= 0 || NM % 8 != 0 || H > 65535) {
        set_error("pe_feat_grad: needs B, H > 0 and N*M a positive multiple of 8");
        return AFFB200_E_ARG;
    }
    if (!d_vals || !q_pe || !feat8 || !out || !work || !aligned16(d_vals) || !aligned16(q_pe) || !aligned16(feat8)) {
        set_error("pe_feat_grad: null or misaligned pointer");
        return AFFB200_E_ARG;
    }
    auto st = (cudaStream_t)stream;
    const int64_t total = B * NM;
    const int chunks = pe_feat_chunks(total);
    const int HG = H <= 2 ? 2 : 4;   // heads per CTA: two-head stages do not pay for the FMAs of two absent heads
    const dim3 grid{unsigned(chunks), unsigned((H + HG - 1) / HG)};
    int e = dispatch_dtype(dtype, [&](auto tag) -> int {
        using T = decltype(tag);
        if constexpr (sizeof(T) != 2) {
            set_error("pe_feat_grad: f16 / bf16 only");
            return AFFB200_E_ARG;
        } else {
            if (HG == 2)
                pe_feat_grad_kernel<T, 2><<<grid, 256, 0, st>>>((const T *)d_vals, q_pe, feat8, int(H), NM, total, work);
            else
                pe_feat_grad_kernel<T, 4><<<grid, 256, 0, st>>>((const T *)d_vals, q_pe, feat8, int(H), NM, total, work);
            return check_launch("pe_feat_grad");
        }
    });
    if (e) return e;
    pe_feat_grad_finish_kernel<<<unsigned((H * 8 + 31) / 32), 1024, 0, st>>>(work, chunks, int(H * 8), out);
    return check_launch("pe_feat_grad_finish");
}

// d_table[p][h] = sum of d_vals[b][h][slot] over the slots whose packed bias index is p (masked slots included: their
// d_vals are zero); d_vals [B][H][NM], pe [B][NM] packed as by affb200_fused_attn_pack_pe.  d_table fully overwritten.
extern "C" int affb200_table_grad_u16(const void *d_vals, const uint16_t *pe, int64_t B, int64_t H, int64_t NM, int64_t P,
                                      float *d_table, int dtype, void *stream) {
    if (!d_vals || !pe || !d_table || B <= 0 || H <= 0 || NM <= 0 || P <= 0 || P > 32768) {
        affb200::set_error("table_grad_u16: bad arguments");
        return AFFB200_E_ARG;
    }
    return affb200::launch_table_grad_u16(d_vals, pe, B, H, NM, P, d_table, dtype, (cudaStream_t)stream);
}
