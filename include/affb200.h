/*
 * affb200.h — C ABI of libaffb200.so: B200 (sm_100a) kernels for AutoFocusFormer's off-grid attention path.
 *
 * This is the drop-in boundary below the reference's three pybind modules
 *   clustenqk_cuda / clustenav_cuda / clustenwf_cuda   (reference clusten/src/clusten{qk,av,wf}_cuda.cpp:48-51)
 * and below the library kernels its Python glue calls for index building
 *   (pykeops argKmin: models/point_utils.py:49-57; torch.sort/topk/gather: models/point_utils.py:148-206,
 *    models/aff_transformer.py:295-341,476-487).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to a dense row-major ("contiguous") array unless marked host;
 *   - inputs are borrowed for the duration of the enqueue, outputs are caller-allocated and fully overwritten
 *     (no pre-zeroing required);
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*); calls never synchronise the host;
 *   - return value 0 = enqueued; negative = AFFB200_E_* and affb200_last_error() holds a thread-local message;
 *   - re-entrant and thread-agnostic (forward runs on the caller's thread, backward on autograd's thread);
 *   - like the reference (no bounds checks, clustenqk_cuda_kernel.cu:38), out-of-range nbhd_idx is undefined.
 *
 * Element types: accumulate in fp32 for f32/f16/bf16 (the reference accumulates in the storage type and has no
 * bf16 dispatch: clustenqk_cuda_kernel.cu:78), in fp64 for f64.
 */
#ifndef AFFB200_H
#define AFFB200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AFFB200_ABI_VERSION 1

typedef enum {
    AFFB200_F32 = 0,
    AFFB200_F16 = 1,
    AFFB200_BF16 = 2,
    AFFB200_F64 = 3
} affb200_dtype;

enum {
    AFFB200_OK = 0,
    AFFB200_E_ARG = -1,      /* bad dimension / null pointer / unsupported dtype */
    AFFB200_E_CUDA = -2,     /* a CUDA runtime call or kernel launch failed */
    AFFB200_E_WORKSPACE = -3 /* caller-provided workspace too small */
};

int affb200_abi_version(void);
const char *affb200_last_error(void);
/* Compile-time target of the kernels in this library, e.g. "sm_100a". */
const char *affb200_target_arch(void);

/* ------------------------------------------------------------------------------------------------------------
 * Transposed neighbour index ("who references key row r"), used by the scatter halves of the backward passes
 * instead of the reference's atomics (clustenqk_cuda_kernel.cu:125, clustenav_cuda_kernel.cu:121,
 * clustenwf_cuda_kernel.cu:129).  CSR per batch element:
 *   off[b][0..Nk]      int32, off[b][r]..off[b][r+1] = range of entries referencing key row r
 *   ent[b][0..Nq*M)    int32, entry e = i*M + j  (query token i, neighbour slot j), ascending within a row
 * `work` is scratch of affb200_csr_work_bytes() bytes.  nbhd_idx: [B][Nq][M] int64.
 * ---------------------------------------------------------------------------------------------------------- */
size_t affb200_csr_off_bytes(int64_t B, int64_t Nk);
size_t affb200_csr_ent_bytes(int64_t B, int64_t Nq, int64_t M);
size_t affb200_csr_work_bytes(int64_t B, int64_t Nk);
int affb200_csr_build(const int64_t *nbhd_idx, int64_t B, int64_t Nq, int64_t Nk, int64_t M,
                      int32_t *off, int32_t *ent, void *work, int sorted, void *stream);

/* ------------------------------------------------------------------------------------------------------------
 * QK: replaces clustenqk_cuda.forward / .backward (clustenqk_cuda.cpp:25-45, clustenqk_cuda_kernel.cu:18-178).
 *   attn[b,h,i,j] = sum_c q[b,h,i,c] * k[b,h,idx[b,i,j],c]
 *   q [B][H][Nq][c], k [B][H][Nk][c] (NOT pre-transposed: clusten.py:30's permute is not needed),
 *   idx [B][Nq][M] int64, attn [B][H][Nq][M].
 * Backward: d_q[b,h,i,c] = sum_j k[idx]*d_attn ; d_k[b,h,r,c] = sum_{(i,j): idx[b,i,j]=r} q[b,h,i,c]*d_attn[b,h,i,j]
 * ---------------------------------------------------------------------------------------------------------- */
int affb200_qk_fwd(const void *q, const void *k, const int64_t *idx, void *attn,
                   int64_t B, int64_t H, int64_t Nq, int64_t Nk, int64_t M, int64_t c,
                   int dtype, void *stream);
int affb200_qk_bwd(const void *d_attn, const void *q, const void *k, const int64_t *idx,
                   const int32_t *csr_off, const int32_t *csr_ent,
                   void *d_q, void *d_k,
                   int64_t B, int64_t H, int64_t Nq, int64_t Nk, int64_t M, int64_t c,
                   int dtype, void *stream);

/* ------------------------------------------------------------------------------------------------------------
 * AV: replaces clustenav_cuda.forward / .backward (clustenav_cuda.cpp:25-45, clustenav_cuda_kernel.cu:18-218).
 *   out[b,h,i,c] = sum_j attn[b,h,i,j] * v[b,h,idx[b,i,j],c]
 * Backward: d_attn[b,h,i,j] = sum_c v[idx]*d_out ; d_v[b,h,r,c] = sum_{(i,j)->r} attn[b,h,i,j]*d_out[b,h,i,c]
 * ---------------------------------------------------------------------------------------------------------- */
int affb200_av_fwd(const void *attn, const void *v, const int64_t *idx, void *out,
                   int64_t B, int64_t H, int64_t Nq, int64_t Nk, int64_t M, int64_t c,
                   int dtype, void *stream);
int affb200_av_bwd(const void *d_out, const void *attn, const void *v, const int64_t *idx,
                   const int32_t *csr_off, const int32_t *csr_ent,
                   void *d_attn, void *d_v,
                   int64_t B, int64_t H, int64_t Nq, int64_t Nk, int64_t M, int64_t c,
                   int dtype, void *stream);

/* ------------------------------------------------------------------------------------------------------------
 * WF: replaces clustenwf_cuda.forward / .backward (clustenwf_cuda.cpp:25-45, clustenwf_cuda_kernel.cu:18-229).
 *   out[b,i,ic,c] = sum_j w[b,i,j,ic] * f[b,idx[b,i,j],c]
 *   w [B][Nq][M][IC], f [B][Nk][C], idx [B][Nq][M] int64, out [B][Nq][IC][C].
 * Backward: d_w[b,i,j,ic] = sum_c f[idx]*d_out[b,i,ic,c] ; d_f[b,r,c] = sum_{(i,j)->r} sum_ic d_out[b,i,ic,c]*w[b,i,j,ic]
 * ---------------------------------------------------------------------------------------------------------- */
int affb200_wf_fwd(const void *w, const void *f, const int64_t *idx, void *out,
                   int64_t B, int64_t Nq, int64_t Nk, int64_t M, int64_t IC, int64_t C,
                   int dtype, void *stream);
int affb200_wf_bwd(const void *d_out, const void *w, const void *f, const int64_t *idx,
                   const int32_t *csr_off, const int32_t *csr_ent,
                   void *d_w, void *d_f,
                   int64_t B, int64_t Nq, int64_t Nk, int64_t M, int64_t IC, int64_t C,
                   int dtype, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* AFFB200_H */
