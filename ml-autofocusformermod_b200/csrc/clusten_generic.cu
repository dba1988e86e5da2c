// clusten_generic.cu — index-agnostic CLUSTEN kernels (any nbhd_idx, duplicates allowed, Nq != Nk allowed).
//
// The eight reference kernels (clusten/src/clusten{qk,av,wf}_cuda_kernel.cu) reduce to three primitives:
//   gather-dot   out[z,i,j] = sum_c x[z,i,c] * y[z,idx[b,i,j],c]          QK fwd, AV bwd (d_attn), WF bwd (d_w)
//   gather-wsum  out[z,i,c] = sum_j w[z,i,j] * y[z,idx[b,i,j],c]          AV fwd, QK bwd (d_q),    WF fwd
//   csr-wsum     dy[z,r,c]  = sum_{e=(i,j): idx[b,i,j]=r} w[z,e]*x[z,i,c]  QK bwd (d_k), AV bwd (d_v), WF bwd (d_f)
// csr-wsum replaces the reference's per-element atomics with a gather over the transposed index (csr_build.cu):
// one fp32 accumulation per output element, one write, no read-modify-write traffic, run-to-run deterministic.
//
// Memory behaviour: every neighbour row is fetched with 128-bit loads (one row = c*sizeof(T) contiguous bytes);
// rows shared between the tokens of a CTA's tile are served by L1 (a tile is TI consecutive tokens, which the
// space-filling order makes spatially adjacent, so their neighbourhoods overlap).
#include "common.cuh"

namespace affb200 {

constexpr int kThreads = 256;

// ---------------------------------------------------------------------------------------------------------
// gather-dot.  grid (ceil(Nq/TI), B); each thread owns (token, slot) pairs of the tile and loops over heads,
// so nbhd_idx is read exactly once (coalesced int64) and reused for all H heads.  x rows of the tile are staged
// in shared memory as accumulator-type values.
// ---------------------------------------------------------------------------------------------------------
template <typename T, int W, int CT>
__global__ void __launch_bounds__(kThreads)
gather_dot_kernel(const T *__restrict__ x, const T *__restrict__ y, const int64_t *__restrict__ idx,
                  T *__restrict__ out, int H, int Nq, int Nk, int M, int c_rt, int TI) {
    using Acc = typename Elem<T>::Acc;
    const int c = CT > 0 ? CT : c_rt;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Acc *x_s = reinterpret_cast<Acc *>(smem_raw);  // [H][TI][c]

    const int b = blockIdx.y;
    const int i0 = blockIdx.x * TI;
    const int ti = min(TI, Nq - i0);
    const int tid = threadIdx.x;

    for (int h = 0; h < H; ++h) {
        const T *xg = x + ((int64_t(b) * H + h) * Nq + i0) * c;
        Acc *xs = x_s + int64_t(h) * TI * c;
        for (int t = tid; t < ti * c; t += kThreads) xs[t] = Elem<T>::to_acc(xg[t]);
    }
    __syncthreads();

    const int64_t *idx_g = idx + (int64_t(b) * Nq + i0) * M;
    for (int p = tid; p < ti * M; p += kThreads) {
        const int il = p / M;
        const int r = int(idx_g[p]);
        for (int h = 0; h < H; ++h) {
            const T *yrow = y + ((int64_t(b) * H + h) * Nk + r) * c;
            const Acc *xr = x_s + (int64_t(h) * TI + il) * c;
            Acc acc = Acc(0);
            if constexpr (CT > 0) {
                Acc yv[CT / W][W];
#pragma unroll
                for (int v = 0; v < CT / W; ++v) load_vec<T, W>(yrow + v * W, yv[v]);
#pragma unroll
                for (int v = 0; v < CT / W; ++v)
#pragma unroll
                    for (int t = 0; t < W; ++t) acc += xr[v * W + t] * yv[v][t];
            } else {
                for (int cc = 0; cc < c; cc += W) {
                    Acc yv[W];
                    load_vec<T, W>(yrow + cc, yv);
#pragma unroll
                    for (int t = 0; t < W; ++t) acc += xr[cc + t] * yv[t];
                }
            }
            out[((int64_t(b) * H + h) * Nq + i0) * M + p] = Elem<T>::from_acc(acc);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// gather-dot, lane-group form (c a compile-time multiple of the 16-byte vector width W, LPR = c/W <= 16 lanes).
// The kernel above gives a thread a whole neighbour row: one warp load then touches 32 different 128-byte lines and
// the L1 tag stage (one line per clock) is the bound (fp32 c=32: 8 loads x 32 lines per 32 outputs — measured 2.3 ms
// at the AFF-Tiny stage-1 shape, the same as the reference kernel).  Here LPR lanes share a row, so a warp load
// covers 32/LPR whole rows (4 lines for fp32 c=32); a group walks LPR consecutive (token, slot) pairs, keeps the LPR
// partial dot products in registers and reduces them with a transposing butterfly (LPR-1 shuffles for LPR sums,
// lane l ends up with pair l), so the 32 results of a warp leave as one coalesced store.
// SAME: M % LPR == 0, all pairs of a group belong to one token -> its x chunk is read once per head.
// ---------------------------------------------------------------------------------------------------------
template <typename Acc, int N>
__device__ __forceinline__ Acc transpose_reduce(Acc (&v)[N], int l) {
#pragma unroll
    for (int s = N / 2; s >= 1; s >>= 1) {
        const bool up = (l & s) != 0;
#pragma unroll
        for (int i = 0; i < s; ++i) {
            const Acc keep = up ? v[i + s] : v[i];
            const Acc send = up ? v[i] : v[i + s];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
        }
    }
    return v[0];
}

template <typename T, int CT, bool SAME>
__global__ void __launch_bounds__(kThreads)
gather_dot_group_kernel(const T *__restrict__ x, const T *__restrict__ y, const int64_t *__restrict__ idx,
                        T *__restrict__ out, int H, int Nq, int Nk, int M, int TI) {
    using Acc = typename Elem<T>::Acc;
    constexpr int W = Elem<T>::VEC;
    constexpr int LPR = CT / W;
    static_assert(LPR >= 1 && LPR <= 16 && (LPR & (LPR - 1)) == 0 && CT % W == 0, "lane-group shape");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Acc *x_s = reinterpret_cast<Acc *>(smem_raw);  // [H][TI][CT]

    const int b = blockIdx.y;
    const int i0 = blockIdx.x * TI;
    const int ti = min(TI, Nq - i0);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int l = lane % LPR, g0 = lane - l;  // lane in group, first lane of the group

    for (int h = 0; h < H; ++h) {
        const T *xg = x + ((int64_t(b) * H + h) * Nq + i0) * CT;
        Acc *xs = x_s + int64_t(h) * TI * CT;
        for (int t = tid; t < ti * CT; t += kThreads) xs[t] = Elem<T>::to_acc(xg[t]);
    }
    __syncthreads();

    const int P = ti * M;
    const int64_t *idx_g = idx + (int64_t(b) * Nq + i0) * M;
    for (int base = warp * 32; base < P; base += (kThreads / 32) * 32) {
        const int p = base + lane;
        const int rl = p < P ? int(idx_g[p]) : 0;
        const int p0 = base + g0;                  // first pair of this group
        const int il0 = min(p0, P - 1) / M;
        for (int h = 0; h < H; ++h) {
            const T *yz = y + (int64_t(b) * H + h) * Nk * CT + l * W;
            const Acc *xz = x_s + int64_t(h) * TI * CT + l * W;
            Acc xv[W];
            if constexpr (SAME) {
#pragma unroll
                for (int t = 0; t < W; ++t) xv[t] = xz[il0 * CT + t];
            }
            Acc part[LPR];
#pragma unroll
            for (int r = 0; r < LPR; ++r) {
                const int row = __shfl_sync(0xffffffffu, rl, g0 + r);
                Acc yv[W];
                load_vec<T, W>(yz + int64_t(row) * CT, yv);
                if constexpr (!SAME) {
                    const int il = min(p0 + r, P - 1) / M;
#pragma unroll
                    for (int t = 0; t < W; ++t) xv[t] = xz[il * CT + t];
                }
                Acc a = Acc(0);
#pragma unroll
                for (int t = 0; t < W; ++t) a += xv[t] * yv[t];
                part[r] = a;
            }
            const Acc sum = transpose_reduce<Acc, LPR>(part, l);
            if (p < P) out[((int64_t(b) * H + h) * Nq + i0) * M + p] = Elem<T>::from_acc(sum);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// gather-wsum.  grid (ceil(Nq/TI), B); LPR = c/W lanes cooperate on one output row, each owning W channels.
// nbhd_idx of the tile is narrowed to int32 in shared memory once and reused by all heads.
// WW = vector width of the weight loads (W when M % W == 0 and rows are 16-byte aligned, else 1).
// ---------------------------------------------------------------------------------------------------------
template <typename T, int W, int WW>
__global__ void __launch_bounds__(kThreads)
gather_wsum_kernel(const T *__restrict__ w, const T *__restrict__ y, const int64_t *__restrict__ idx,
                   T *__restrict__ out, int H, int Nq, int Nk, int M, int c, int TI) {
    using Acc = typename Elem<T>::Acc;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int *idx_s = reinterpret_cast<int *>(smem_raw);  // [TI][MS]
    const int MS = M | 1;                            // odd row stride: conflict-free across lane groups

    const int b = blockIdx.y;
    const int i0 = blockIdx.x * TI;
    const int ti = min(TI, Nq - i0);
    const int tid = threadIdx.x;

    const int64_t *idx_g = idx + (int64_t(b) * Nq + i0) * M;
    for (int p = tid; p < ti * M; p += kThreads) {
        const int il = p / M;
        idx_s[il * MS + (p - il * M)] = int(idx_g[p]);
    }
    __syncthreads();

    const int LPR = c / W;
    const int items = ti * H * LPR;
    for (int it = tid; it < items; it += kThreads) {
        const int part = it % LPR;
        const int t2 = it / LPR;
        const int il = t2 % ti;
        const int h = t2 / ti;
        const int64_t z = int64_t(b) * H + h;
        const T *wrow = w + (z * Nq + i0 + il) * M;
        const T *ybase = y + z * Nk * c + part * W;
        const int *ir = idx_s + il * MS;
        Acc acc[W];
#pragma unroll
        for (int t = 0; t < W; ++t) acc[t] = Acc(0);
        for (int j0 = 0; j0 < M; j0 += WW) {
            Acc wv[WW];
            load_vec<T, WW>(wrow + j0, wv);
            Acc yv[WW][W];
#pragma unroll
            for (int jj = 0; jj < WW; ++jj) load_vec<T, W>(ybase + int64_t(ir[j0 + jj]) * c, yv[jj]);
#pragma unroll
            for (int jj = 0; jj < WW; ++jj)
#pragma unroll
                for (int t = 0; t < W; ++t) acc[t] += wv[jj] * yv[jj][t];
        }
        store_vec<T, W>(out + (z * Nq + i0 + il) * c + part * W, acc);
    }
}

// ---------------------------------------------------------------------------------------------------------
// csr-wsum.  One lane group (LPR lanes) per output row (z, r); walks the row's reference list.
// ---------------------------------------------------------------------------------------------------------
template <typename T, int W>
__global__ void __launch_bounds__(kThreads)
csr_wsum_kernel(const T *__restrict__ w, const T *__restrict__ x, const int32_t *__restrict__ off,
                const int32_t *__restrict__ ent, T *__restrict__ dy, int H, int Nq, int Nk, int M, int c,
                FastDiv divM, int64_t n_groups) {
    using Acc = typename Elem<T>::Acc;
    const int LPR = c / W;
    const int64_t g = int64_t(blockIdx.x) * kThreads + threadIdx.x;
    const int64_t grp = g / LPR;
    if (grp >= n_groups) return;
    const int part = int(g - grp * LPR);
    const int64_t z = grp / Nk;
    const int r = int(grp - z * Nk);
    const int b = int(z / H);
    const int32_t *o = off + int64_t(b) * (Nk + 1) + r;
    const int beg = o[0], end = o[1];
    const int32_t *e_p = ent + int64_t(b) * Nq * M;
    const T *wz = w + z * int64_t(Nq) * M;
    const T *xz = x + z * int64_t(Nq) * c + part * W;
    Acc acc[W];
#pragma unroll
    for (int t = 0; t < W; ++t) acc[t] = Acc(0);
    int p = beg;
    for (; p + 4 <= end; p += 4) {
        int e[4];
        Acc wv[4], xv[4][W];
#pragma unroll
        for (int u = 0; u < 4; ++u) e[u] = __ldg(e_p + p + u);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            wv[u] = Elem<T>::to_acc(__ldg(wz + e[u]));
            load_vec<T, W>(xz + int64_t(divM.div(uint32_t(e[u]))) * c, xv[u]);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int t = 0; t < W; ++t) acc[t] += wv[u] * xv[u][t];
    }
    for (; p < end; ++p) {
        const int e = __ldg(e_p + p);
        const Acc wv = Elem<T>::to_acc(__ldg(wz + e));
        Acc xv[W];
        load_vec<T, W>(xz + int64_t(divM.div(uint32_t(e))) * c, xv);
#pragma unroll
        for (int t = 0; t < W; ++t) acc[t] += wv * xv[t];
    }
    store_vec<T, W>(dy + grp * c + part * W, acc);
}

// ---------------------------------------------------------------------------------------------------------
// WF forward: out[b,i,ic,c] = sum_j w[b,i,j,ic] * f[b,idx[b,i,j],c].  Each gathered row is read ONCE and feeds
// all ICT inner channels (the reference re-gathers it per inner channel: clustenwf_cuda_kernel.cu:41-47).
// ICT = 4 (model value, aff_transformer.py:267) or 1 (any IC; grid.y = IC).
// ---------------------------------------------------------------------------------------------------------
template <typename T, int N> __device__ __forceinline__ void load_small(const T *p, typename Elem<T>::Acc (&o)[N]) {
    if constexpr (N == 4 && sizeof(T) == 2) {
        uint2 raw = __ldg(reinterpret_cast<const uint2 *>(p));
        const T *e = reinterpret_cast<const T *>(&raw);
#pragma unroll
        for (int t = 0; t < 4; ++t) o[t] = Elem<T>::to_acc(e[t]);
    } else if constexpr (N == 4 && sizeof(T) == 4) {
        uint4 raw = __ldg(reinterpret_cast<const uint4 *>(p));
        const T *e = reinterpret_cast<const T *>(&raw);
#pragma unroll
        for (int t = 0; t < 4; ++t) o[t] = Elem<T>::to_acc(e[t]);
    } else {
#pragma unroll
        for (int t = 0; t < N; ++t) o[t] = Elem<T>::to_acc(__ldg(p + t));
    }
}

template <typename T, int W, int ICT>
__global__ void __launch_bounds__(kThreads)
wf_fwd_kernel(const T *__restrict__ w, const T *__restrict__ f, const int64_t *__restrict__ idx,
              T *__restrict__ out, int Nq, int Nk, int M, int IC, int C, int64_t n_groups) {
    using Acc = typename Elem<T>::Acc;
    const int LPR = C / W;
    const int64_t g = int64_t(blockIdx.x) * kThreads + threadIdx.x;
    const int64_t grp = g / LPR;  // (b, i)
    if (grp >= n_groups) return;
    const int part = int(g - grp * LPR);
    const int ic0 = blockIdx.y * ICT;
    const int64_t b = grp / Nq;
    const int64_t *ir = idx + grp * M;
    const T *wrow = w + grp * M * IC + ic0;
    const T *fb = f + b * Nk * C + part * W;
    Acc acc[ICT][W];
#pragma unroll
    for (int q = 0; q < ICT; ++q)
#pragma unroll
        for (int t = 0; t < W; ++t) acc[q][t] = Acc(0);
#pragma unroll 4
    for (int j = 0; j < M; ++j) {
        const int64_t r = __ldg(ir + j);
        Acc fv[W], wv[ICT];
        load_vec<T, W>(fb + r * C, fv);
        load_small<T, ICT>(wrow + int64_t(j) * IC, wv);
#pragma unroll
        for (int q = 0; q < ICT; ++q)
#pragma unroll
            for (int t = 0; t < W; ++t) acc[q][t] += wv[q] * fv[t];
    }
#pragma unroll
    for (int q = 0; q < ICT; ++q) store_vec<T, W>(out + (grp * IC + ic0 + q) * C + part * W, acc[q]);
}

// WF backward, weights half: d_w[b,i,j,ic] = sum_c f[b,idx,c] * d_out[b,i,ic,c].  G lanes (power of two) per
// (i,j) pair stride over the C channels with 128-bit loads, then butterfly-reduce.
template <typename T, int W, int ICT>
__global__ void __launch_bounds__(kThreads)
wf_bwd_w_kernel(const T *__restrict__ d_out, const T *__restrict__ f, const int64_t *__restrict__ idx,
                T *__restrict__ d_w, int Nq, int Nk, int M, int IC, int C, int G, FastDiv divM,
                int64_t n_pairs) {
    using Acc = typename Elem<T>::Acc;
    const int64_t g = int64_t(blockIdx.x) * kThreads + threadIdx.x;
    const int64_t pair = g / G;
    const int lane = int(g - pair * G);
    const bool valid = pair < n_pairs;
    const int ic0 = blockIdx.y * ICT;
    Acc acc[ICT];
#pragma unroll
    for (int q = 0; q < ICT; ++q) acc[q] = Acc(0);
    if (valid) {
        const int64_t bi = pair / M;  // b*Nq + i
        const int64_t b = bi / Nq;
        const int64_t r = __ldg(idx + pair);
        const T *frow = f + (b * Nk + r) * C;
        const T *drow = d_out + (bi * IC + ic0) * C;
        for (int cc = lane * W; cc < C; cc += G * W) {
            Acc fv[W];
            load_vec<T, W>(frow + cc, fv);
#pragma unroll
            for (int q = 0; q < ICT; ++q) {
                Acc dv[W];
                load_vec<T, W>(drow + int64_t(q) * C + cc, dv);
#pragma unroll
                for (int t = 0; t < W; ++t) acc[q] += fv[t] * dv[t];
            }
        }
    }
    for (int o = G >> 1; o > 0; o >>= 1)
#pragma unroll
        for (int q = 0; q < ICT; ++q) acc[q] += __shfl_xor_sync(0xffffffffu, acc[q], o);
    if (valid && lane == 0) {
#pragma unroll
        for (int q = 0; q < ICT; ++q) d_w[pair * IC + ic0 + q] = Elem<T>::from_acc(acc[q]);
    }
}

// N partial sums per lane, L lanes (powers of two, N >= L): after log2(L) exchange steps lane l of the group holds the
// complete sums [l*N/L, (l+1)*N/L) in v[0 .. N/L).  N-N/L shuffles in all, against N*log2(L) for N butterflies.
template <typename Acc, int N, int L>
__device__ __forceinline__ void transpose_reduce_block(Acc (&v)[N], int l) {
    static_assert(N >= L && (N & (N - 1)) == 0 && (L & (L - 1)) == 0, "powers of two");
#pragma unroll
    for (int bit = L / 2, half = N / 2; bit >= 1; bit >>= 1, half >>= 1) {
        const bool up = (l & bit) != 0;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            const Acc keep = up ? v[i + half] : v[i];
            const Acc send = up ? v[i] : v[i + half];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
        }
    }
}

// WF backward, weights half, grouped form (IC == 4, C = G*NCH*W): G lanes own one token's d_out rows (4 x C, converted
// once, in registers) and walk 8 of its slots; the 8 x 4 partial sums per lane meet in one transposing reduction, so
// lane l of a group ends with the four inner-channel sums of slot l/.. and the group's 32 results leave as contiguous
// stores.  grid (groups of the image, 1, B): no 64-bit divisions (the pair-per-lane-group kernel above spent most of
// its 320 instructions per pair on two of them and on re-converting the d_out rows for every slot).
template <typename T, int G, int NCH>
__global__ void __launch_bounds__(kThreads)
wf_bwd_w_group_kernel(const T *__restrict__ d_out, const T *__restrict__ f, const int64_t *__restrict__ idx,
                      T *__restrict__ d_w, int Nq, int Nk, int M, FastDiv div_gpt, int gpt, int n_groups) {
    using Acc = typename Elem<T>::Acc;
    constexpr int W = Elem<T>::VEC, ICT = 4, PP = 8, C = G * NCH * W;
    const int b = blockIdx.z;
    const int gthread = blockIdx.x * kThreads + threadIdx.x;
    const int grp = gthread / G, l = gthread % G;
    const bool live = grp < n_groups;
    const int gq = live ? grp : n_groups - 1;           // dead lanes of the last warp compute on a valid group
    const int tok = int(div_gpt.div(uint32_t(gq)));
    const int j0 = (gq - tok * gpt) * PP;
    const int64_t bt = int64_t(b) * Nq + tok;
    Acc dv[ICT][NCH][W];
#pragma unroll
    for (int q = 0; q < ICT; ++q)
#pragma unroll
        for (int n = 0; n < NCH; ++n) load_vec<T, W>(d_out + (bt * ICT + q) * C + (n * G + l) * W, dv[q][n]);
    const int64_t *ir = idx + bt * M;
    const T *fb = f + int64_t(b) * Nk * C;
    Acc acc[PP * ICT];
#pragma unroll
    for (int p = 0; p < PP; ++p) {
        const int j = min(j0 + p, M - 1);
        const int64_t r = __ldg(ir + j);
        Acc a[ICT] = {Acc(0), Acc(0), Acc(0), Acc(0)};
#pragma unroll
        for (int n = 0; n < NCH; ++n) {
            Acc fv[W];
            load_vec<T, W>(fb + r * C + (n * G + l) * W, fv);
#pragma unroll
            for (int q = 0; q < ICT; ++q)
#pragma unroll
                for (int t = 0; t < W; ++t) a[q] += fv[t] * dv[q][n][t];
        }
#pragma unroll
        for (int q = 0; q < ICT; ++q) acc[p * ICT + q] = a[q];
    }
    transpose_reduce_block<Acc, PP * ICT, G>(acc, l);
    // lane l holds results [l*R, (l+1)*R) of the group's 32, R = 32/G; result k = slot j0 + k/4, inner channel k%4
    constexpr int R = PP * ICT / G;
    if (live) {
        T *o = d_w + (bt * M + j0) * ICT;
#pragma unroll
        for (int k = 0; k < R; ++k) {
            const int e = l * R + k;
            if (j0 + e / ICT < M) o[e] = Elem<T>::from_acc(acc[k]);
        }
    }
}

template <typename T, int G, int NCH>
static int launch_wf_bwd_w_group(const T *d_out, const T *f, const int64_t *idx, T *d_w, int64_t B, int64_t Nq,
                                 int64_t Nk, int64_t M, cudaStream_t st) {
    const int gpt = int((M + 7) / 8);
    const int64_t n_groups = Nq * gpt;
    dim3 grid(unsigned(ceil_div64(n_groups * G, kThreads)), 1u, unsigned(B));
    wf_bwd_w_group_kernel<T, G, NCH><<<grid, kThreads, 0, st>>>(d_out, f, idx, d_w, int(Nq), int(Nk), int(M),
                                                               FastDiv{uint32_t(gpt)}, gpt, int(n_groups));
    return check_launch("wf_bwd_w_group");
}

// WF backward, feature half: d_f[b,r,c] = sum_{e=(i,j)->r} sum_ic d_out[b,i,ic,c] * w[b,i,j,ic]  (CSR walk).
template <typename T, int W, int ICT>
__global__ void __launch_bounds__(kThreads)
wf_bwd_f_kernel(const T *__restrict__ d_out, const T *__restrict__ w, const int32_t *__restrict__ off,
                const int32_t *__restrict__ ent, T *__restrict__ d_f, int Nq, int Nk, int M, int IC, int C,
                FastDiv divM, int64_t n_groups) {
    using Acc = typename Elem<T>::Acc;
    const int LPR = C / W;
    const int64_t g = int64_t(blockIdx.x) * kThreads + threadIdx.x;
    const int64_t grp = g / LPR;  // (b, r)
    if (grp >= n_groups) return;
    const int part = int(g - grp * LPR);
    const int64_t b = grp / Nk;
    const int r = int(grp - b * Nk);
    const int32_t *o = off + b * (Nk + 1) + r;
    const int beg = o[0], end = o[1];
    const int32_t *e_p = ent + b * Nq * M;
    const T *wb = w + b * Nq * M * IC;
    const T *db = d_out + b * Nq * IC * C + part * W;
    Acc acc[W];
#pragma unroll
    for (int t = 0; t < W; ++t) acc[t] = Acc(0);
#pragma unroll 2
    for (int p = beg; p < end; ++p) {
        const int e = __ldg(e_p + p);
        const int64_t i = divM.div(uint32_t(e));
        if constexpr (ICT > 0) {
            Acc wv[ICT];
            load_small<T, ICT>(wb + int64_t(e) * IC, wv);
            Acc dv[ICT][W];
#pragma unroll
            for (int q = 0; q < ICT; ++q) load_vec<T, W>(db + (i * IC + q) * C, dv[q]);
#pragma unroll
            for (int q = 0; q < ICT; ++q)
#pragma unroll
                for (int t = 0; t < W; ++t) acc[t] += wv[q] * dv[q][t];
        } else {
            for (int q = 0; q < IC; ++q) {
                const Acc wv = Elem<T>::to_acc(__ldg(wb + int64_t(e) * IC + q));
                Acc dv[W];
                load_vec<T, W>(db + (i * IC + q) * C, dv);
#pragma unroll
                for (int t = 0; t < W; ++t) acc[t] += wv * dv[t];
            }
        }
    }
    store_vec<T, W>(d_f + grp * C + part * W, acc);
}

// =========================================================================================================
// host-side launchers
// =========================================================================================================
static bool fits_i32(int64_t v) { return v >= 0 && v < (int64_t(1) << 31); }

static int check_dims(const char *op, int64_t B, int64_t H, int64_t Nq, int64_t Nk, int64_t M, int64_t c) {
    if (B < 0 || H < 0 || Nq < 0 || Nk < 0 || M < 0 || c < 0) {
        set_error("%s: negative dimension", op);
        return AFFB200_E_ARG;
    }
    if (!fits_i32(Nq * M) || !fits_i32(Nk) || !fits_i32(c) || !fits_i32(H) || B > 65535 * int64_t(1024)) {
        set_error("%s: dimension too large (Nq*M, Nk, c, H must fit int32)", op);
        return AFFB200_E_ARG;
    }
    return AFFB200_OK;
}

// tokens per tile so that the dynamic shared memory of a tile stays within `budget` bytes
static int pick_tile(int64_t per_token_bytes, int64_t Nq, int budget = 40 * 1024, int cap = 32) {
    int64_t ti = per_token_bytes > 0 ? budget / per_token_bytes : cap;
    if (ti > cap) ti = cap;
    if (ti > Nq) ti = Nq;
    if (ti < 1) ti = 1;
    return int(ti);
}

template <typename T, int CT>
static int launch_gather_dot_group(const T *x, const T *y, const int64_t *idx, T *out, int64_t B, int64_t H,
                                   int64_t Nq, int64_t Nk, int64_t M, int TI, size_t smem, dim3 grid,
                                   cudaStream_t st) {
    constexpr int LPR = CT / Elem<T>::VEC;
    if constexpr (LPR > 16) {
        return 1;
    } else {
        auto go = [&](auto kern) {
            if (smem > 48 * 1024)
                cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
            kern<<<grid, kThreads, smem, st>>>(x, y, idx, out, int(H), int(Nq), int(Nk), int(M), TI);
        };
        if (M % LPR == 0) go(gather_dot_group_kernel<T, CT, true>);
        else go(gather_dot_group_kernel<T, CT, false>);
        return check_launch("gather_dot_group");
    }
}

template <typename T>
int launch_gather_dot(const T *x, const T *y, const int64_t *idx, T *out, int64_t B, int64_t H, int64_t Nq,
                      int64_t Nk, int64_t M, int64_t c, cudaStream_t st) {
    using Acc = typename Elem<T>::Acc;
    if (B * H * Nq * M == 0) return AFFB200_OK;
    if (c == 0) return cudaMemsetAsync(out, 0, size_t(B * H * Nq * M) * sizeof(T), st) == cudaSuccess
                           ? AFFB200_OK : check_launch("memset");
    constexpr int W = Elem<T>::VEC;
    const int TI = pick_tile(H * c * int64_t(sizeof(Acc)), Nq, 96 * 1024);
    const size_t smem = size_t(H) * TI * c * sizeof(Acc);
    dim3 grid(unsigned(ceil_div64(Nq, TI)), unsigned(B));
    const bool vec = (c % W == 0) && aligned16(y);
    if (vec) {
        int e = 1;  // 1: shape not covered by the lane-group kernel
        if (c == 32) e = launch_gather_dot_group<T, 32>(x, y, idx, out, B, H, Nq, Nk, M, TI, smem, grid, st);
        else if (c == 16) e = launch_gather_dot_group<T, 16>(x, y, idx, out, B, H, Nq, Nk, M, TI, smem, grid, st);
        else if (c == 64) e = launch_gather_dot_group<T, 64>(x, y, idx, out, B, H, Nq, Nk, M, TI, smem, grid, st);
        if (e <= 0) return e;
    }
#define AFF_GD(WW_, CT_)                                                                                       \
    do {                                                                                                       \
        auto kern = gather_dot_kernel<T, WW_, CT_>;                                                            \
        if (smem > 48 * 1024)                                                                                  \
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));                \
        kern<<<grid, kThreads, smem, st>>>(x, y, idx, out, int(H), int(Nq), int(Nk), int(M), int(c), TI);      \
    } while (0)
    if (vec && c == 32) AFF_GD(W, 32);
    else if (vec && c == 16) AFF_GD(W, 16);
    else if (vec && c == 64) AFF_GD(W, 64);
    else if (vec) AFF_GD(W, 0);
    else AFF_GD(1, 0);
#undef AFF_GD
    return check_launch("gather_dot");
}

template <typename T>
int launch_gather_wsum(const T *w, const T *y, const int64_t *idx, T *out, int64_t B, int64_t H, int64_t Nq,
                       int64_t Nk, int64_t M, int64_t c, cudaStream_t st) {
    if (B * H * Nq * c == 0) return AFFB200_OK;
    if (M == 0) return cudaMemsetAsync(out, 0, size_t(B * H * Nq * c) * sizeof(T), st) == cudaSuccess
                           ? AFFB200_OK : check_launch("memset");
    constexpr int W = Elem<T>::VEC;
    const int MS = int(M) | 1;
    const int TI = pick_tile(int64_t(MS) * 4, Nq);
    const size_t smem = size_t(TI) * MS * sizeof(int);
    dim3 grid(unsigned(ceil_div64(Nq, TI)), unsigned(B));
    const bool vec = (c % W == 0) && aligned16(y) && aligned16(out);
    const bool wvec = (M % W == 0) && aligned16(w);
    if (vec && wvec)
        gather_wsum_kernel<T, W, W><<<grid, kThreads, smem, st>>>(w, y, idx, out, int(H), int(Nq), int(Nk), int(M), int(c), TI);
    else if (vec)
        gather_wsum_kernel<T, W, 1><<<grid, kThreads, smem, st>>>(w, y, idx, out, int(H), int(Nq), int(Nk), int(M), int(c), TI);
    else
        gather_wsum_kernel<T, 1, 1><<<grid, kThreads, smem, st>>>(w, y, idx, out, int(H), int(Nq), int(Nk), int(M), int(c), TI);
    return check_launch("gather_wsum");
}

template <typename T>
int launch_csr_wsum(const T *w, const T *x, const int32_t *off, const int32_t *ent, T *dy, int64_t B, int64_t H,
                    int64_t Nq, int64_t Nk, int64_t M, int64_t c, cudaStream_t st) {
    if (B * H * Nk * c == 0) return AFFB200_OK;
    if (Nq * M == 0) return cudaMemsetAsync(dy, 0, size_t(B * H * Nk * c) * sizeof(T), st) == cudaSuccess
                                ? AFFB200_OK : check_launch("memset");
    constexpr int W = Elem<T>::VEC;
    const bool vec = (c % W == 0) && aligned16(x) && aligned16(dy);
    const int64_t n_groups = B * H * Nk;
    const FastDiv divM{uint32_t(M)};
    if (vec) {
        const int64_t threads = n_groups * (c / W);
        csr_wsum_kernel<T, W><<<unsigned(ceil_div64(threads, kThreads)), kThreads, 0, st>>>(
            w, x, off, ent, dy, int(H), int(Nq), int(Nk), int(M), int(c), divM, n_groups);
    } else {
        const int64_t threads = n_groups * c;
        csr_wsum_kernel<T, 1><<<unsigned(ceil_div64(threads, kThreads)), kThreads, 0, st>>>(
            w, x, off, ent, dy, int(H), int(Nq), int(Nk), int(M), int(c), divM, n_groups);
    }
    return check_launch("csr_wsum");
}

template <typename T>
int launch_wf_fwd(const T *w, const T *f, const int64_t *idx, T *out, int64_t B, int64_t Nq, int64_t Nk,
                  int64_t M, int64_t IC, int64_t C, cudaStream_t st) {
    if (B * Nq * IC * C == 0) return AFFB200_OK;
    if (M == 0) return cudaMemsetAsync(out, 0, size_t(B * Nq * IC * C) * sizeof(T), st) == cudaSuccess
                           ? AFFB200_OK : check_launch("memset");
    constexpr int W = Elem<T>::VEC;
    const bool vec = (C % W == 0) && aligned16(f) && aligned16(out);
    const bool ic4 = (IC == 4) && aligned16(w);
    const int64_t n_groups = B * Nq;
    const int64_t threads = n_groups * (vec ? C / W : C);
    dim3 grid(unsigned(ceil_div64(threads, kThreads)), ic4 ? 1u : unsigned(IC));
#define AFF_WF(W_, I_) wf_fwd_kernel<T, W_, I_><<<grid, kThreads, 0, st>>>(w, f, idx, out, int(Nq), int(Nk), int(M), int(IC), int(C), n_groups)
    if (vec && ic4) AFF_WF(W, 4);
    else if (vec) AFF_WF(W, 1);
    else if (ic4) AFF_WF(1, 4);
    else AFF_WF(1, 1);
#undef AFF_WF
    return check_launch("wf_fwd");
}

template <typename T>
int launch_wf_bwd_w(const T *d_out, const T *f, const int64_t *idx, T *d_w, int64_t B, int64_t Nq, int64_t Nk,
                    int64_t M, int64_t IC, int64_t C, cudaStream_t st) {
    if (B * Nq * M * IC == 0) return AFFB200_OK;
    if (C == 0) return cudaMemsetAsync(d_w, 0, size_t(B * Nq * M * IC) * sizeof(T), st) == cudaSuccess
                           ? AFFB200_OK : check_launch("memset");
    constexpr int W = Elem<T>::VEC;
    const bool vec = (C % W == 0) && aligned16(f) && aligned16(d_out);
    const int64_t lpr = vec ? C / W : C;
    int G = 1;
    while (G * 2 <= 32 && G * 2 <= lpr) G *= 2;
    const bool ic4 = (IC == 4);
    if (vec && ic4 && B <= 65535 && Nq * ((M + 7) / 8) * 32 < (int64_t(1) << 31)) {   // grouped kernel: C = G*NCH*W
        const int64_t lanes = C / W;
#define AFF_WG(G_, N_) return launch_wf_bwd_w_group<T, G_, N_>(d_out, f, idx, d_w, B, Nq, Nk, M, st)
        if (lanes == 4) AFF_WG(4, 1);
        if (lanes == 8) AFF_WG(8, 1);
        if (lanes == 16) AFF_WG(16, 1);
        if (lanes == 32) AFF_WG(32, 1);
        if (lanes == 64) AFF_WG(32, 2);
        if (lanes == 128 && sizeof(T) <= 4) AFF_WG(32, 4);
#undef AFF_WG
    }
    const int64_t n_pairs = B * Nq * M;
    dim3 grid(unsigned(ceil_div64(n_pairs * G, kThreads)), ic4 ? 1u : unsigned(IC));
    const FastDiv divM{uint32_t(M)};
#define AFF_WB(W_, I_) wf_bwd_w_kernel<T, W_, I_><<<grid, kThreads, 0, st>>>(d_out, f, idx, d_w, int(Nq), int(Nk), int(M), int(IC), int(C), G, divM, n_pairs)
    if (vec && ic4) AFF_WB(W, 4);
    else if (vec) AFF_WB(W, 1);
    else if (ic4) AFF_WB(1, 4);
    else AFF_WB(1, 1);
#undef AFF_WB
    return check_launch("wf_bwd_w");
}

template <typename T>
int launch_wf_bwd_f(const T *d_out, const T *w, const int32_t *off, const int32_t *ent, T *d_f, int64_t B,
                    int64_t Nq, int64_t Nk, int64_t M, int64_t IC, int64_t C, cudaStream_t st) {
    if (B * Nk * C == 0) return AFFB200_OK;
    if (Nq * M * IC == 0) return cudaMemsetAsync(d_f, 0, size_t(B * Nk * C) * sizeof(T), st) == cudaSuccess
                                     ? AFFB200_OK : check_launch("memset");
    constexpr int W = Elem<T>::VEC;
    const bool vec = (C % W == 0) && aligned16(d_out) && aligned16(d_f);
    const bool ic4 = (IC == 4) && aligned16(w);
    const int64_t n_groups = B * Nk;
    const int64_t threads = n_groups * (vec ? C / W : C);
    const unsigned grid = unsigned(ceil_div64(threads, kThreads));
    const FastDiv divM{uint32_t(M)};
#define AFF_WFB(W_, I_) wf_bwd_f_kernel<T, W_, I_><<<grid, kThreads, 0, st>>>(d_out, w, off, ent, d_f, int(Nq), int(Nk), int(M), int(IC), int(C), divM, n_groups)
    if (vec && ic4) AFF_WFB(W, 4);
    else if (vec) AFF_WFB(W, 0);
    else if (ic4) AFF_WFB(1, 4);
    else AFF_WFB(1, 0);
#undef AFF_WFB
    return check_launch("wf_bwd_f");
}

}  // namespace affb200

// =========================================================================================================
// C ABI
// =========================================================================================================
using namespace affb200;

#define AFF_REQUIRE(cond, msg)                         \
    do {                                               \
        if (!(cond)) {                                 \
            set_error("%s: %s", __func__, msg);        \
            return AFFB200_E_ARG;                      \
        }                                              \
    } while (0)

template <typename T> struct IsDenseType { static constexpr bool value = false; };
template <> struct IsDenseType<__half> { static constexpr bool value = true; };
template <> struct IsDenseType<__nv_bfloat16> { static constexpr bool value = true; };

static bool dense_shape_ok(int64_t M, int64_t c, const void *a, const void *b, const void *o) {
    return (c == 16 || c == 32) && M >= 1 && M <= AFFB200_PLAN_MAX_M && aligned16(a) && aligned16(b) && aligned16(o);
}
static bool has_qplan(const affb200_nbhd_plan *p) { return p && p->tq > 0 && p->q_hdr && p->q_rows && p->q_loc; }
static bool has_kplan(const affb200_nbhd_plan *p) { return p && p->tk > 0 && p->k_hdr && p->k_utok && p->k_rloc; }
static bool has_csr(const affb200_nbhd_plan *p) { return p && p->csr_off && p->csr_ent; }

// out[z,i,j] = sum_c x[z,i,c]*y[z,idx,c]
template <typename T>
static int run_gather_dot(const T *x, const T *y, const int64_t *idx, const affb200_nbhd_plan *plan, T *out,
                          int64_t B, int64_t H, int64_t Nq, int64_t Nk, int64_t M, int64_t c, cudaStream_t st) {
    if constexpr (IsDenseType<T>::value) {
        if (has_qplan(plan) && dense_shape_ok(M, c, x, y, out) && B * H * Nq > 0)
            return launch_gather_dot_dense<T>(x, y, plan, out, B, H, Nq, Nk, M, c, st);
    }
    return launch_gather_dot<T>(x, y, idx, out, B, H, Nq, Nk, M, c, st);
}
// out[z,i,c] = sum_j w[z,i,j]*y[z,idx,c]
template <typename T>
static int run_gather_wsum(const T *w, const T *y, const int64_t *idx, const affb200_nbhd_plan *plan, T *out,
                           int64_t B, int64_t H, int64_t Nq, int64_t Nk, int64_t M, int64_t c, cudaStream_t st) {
    if constexpr (IsDenseType<T>::value) {
        if (has_qplan(plan) && dense_shape_ok(M, c, w, y, out) && B * H * Nq > 0)
            return launch_gather_wsum_dense<T>(w, y, plan, out, B, H, Nq, Nk, M, c, st);
    }
    return launch_gather_wsum<T>(w, y, idx, out, B, H, Nq, Nk, M, c, st);
}
// dy[z,r,c] = sum_{(i,j)->r} w[z,i,j]*x[z,i,c]
template <typename T>
static int run_scatter_wsum(const char *op, const T *w, const T *x, const affb200_nbhd_plan *plan, T *dy, int64_t B,
                            int64_t H, int64_t Nq, int64_t Nk, int64_t M, int64_t c, cudaStream_t st) {
    if constexpr (IsDenseType<T>::value) {
        if (has_kplan(plan) && dense_shape_ok(M, c, w, x, dy) && B * H * Nk > 0 && Nq > 0)
            return launch_scatter_wsum_dense<T>(w, x, plan, dy, B, H, Nq, Nk, M, c, st);
    }
    if (!has_csr(plan) && B * Nq * M > 0) {
        set_error("%s: the plan holds neither a usable key-tile plan nor the transposed index (csr_off/csr_ent)", op);
        return AFFB200_E_ARG;
    }
    return launch_csr_wsum<T>(w, x, plan ? plan->csr_off : nullptr, plan ? plan->csr_ent : nullptr, dy, B, H, Nq, Nk,
                              M, c, st);
}

extern "C" int affb200_qk_fwd(const void *q, const void *k, const int64_t *idx, const affb200_nbhd_plan *plan,
                              void *attn, int64_t B, int64_t H, int64_t Nq, int64_t Nk, int64_t M, int64_t c,
                              int dtype, void *stream) {
    if (int e = check_dims("qk_fwd", B, H, Nq, Nk, M, c)) return e;
    AFF_REQUIRE(B * H * Nq * M == 0 || (attn && idx && (c == 0 || (q && k))), "null pointer");
    return dispatch_dtype(dtype, [&](auto tag) {
        using T = decltype(tag);
        return run_gather_dot<T>((const T *)q, (const T *)k, idx, plan, (T *)attn, B, H, Nq, Nk, M, c,
                                 (cudaStream_t)stream);
    });
}

extern "C" int affb200_qk_bwd(const void *d_attn, const void *q, const void *k, const int64_t *idx,
                              const affb200_nbhd_plan *plan, void *d_q, void *d_k, int64_t B,
                              int64_t H, int64_t Nq, int64_t Nk, int64_t M, int64_t c, int dtype, void *stream) {
    if (int e = check_dims("qk_bwd", B, H, Nq, Nk, M, c)) return e;
    AFF_REQUIRE(B * H * Nq * M * c == 0 || (d_attn && q && k && idx), "null pointer");
    AFF_REQUIRE(B * H * Nq * c == 0 || d_q, "null pointer");
    AFF_REQUIRE(B * H * Nk * c == 0 || d_k, "null pointer");
    return dispatch_dtype(dtype, [&](auto tag) {
        using T = decltype(tag);
        auto st = (cudaStream_t)stream;
        if (int e = run_gather_wsum<T>((const T *)d_attn, (const T *)k, idx, plan, (T *)d_q, B, H, Nq, Nk, M, c, st))
            return e;
        return run_scatter_wsum<T>("qk_bwd", (const T *)d_attn, (const T *)q, plan, (T *)d_k, B, H, Nq, Nk, M, c, st);
    });
}

extern "C" int affb200_av_fwd(const void *attn, const void *v, const int64_t *idx, const affb200_nbhd_plan *plan,
                              void *out, int64_t B, int64_t H, int64_t Nq, int64_t Nk, int64_t M, int64_t c,
                              int dtype, void *stream) {
    if (int e = check_dims("av_fwd", B, H, Nq, Nk, M, c)) return e;
    AFF_REQUIRE(B * H * Nq * c == 0 || (out && (M == 0 || (attn && v && idx))), "null pointer");
    return dispatch_dtype(dtype, [&](auto tag) {
        using T = decltype(tag);
        return run_gather_wsum<T>((const T *)attn, (const T *)v, idx, plan, (T *)out, B, H, Nq, Nk, M, c,
                                  (cudaStream_t)stream);
    });
}

extern "C" int affb200_av_bwd(const void *d_out, const void *attn, const void *v, const int64_t *idx,
                              const affb200_nbhd_plan *plan, void *d_attn, void *d_v, int64_t B,
                              int64_t H, int64_t Nq, int64_t Nk, int64_t M, int64_t c, int dtype, void *stream) {
    if (int e = check_dims("av_bwd", B, H, Nq, Nk, M, c)) return e;
    AFF_REQUIRE(B * H * Nq * M * c == 0 || (d_out && attn && v && idx && d_attn), "null pointer");
    AFF_REQUIRE(B * H * Nk * c == 0 || d_v, "null pointer");
    return dispatch_dtype(dtype, [&](auto tag) {
        using T = decltype(tag);
        auto st = (cudaStream_t)stream;
        if (int e = run_gather_dot<T>((const T *)d_out, (const T *)v, idx, plan, (T *)d_attn, B, H, Nq, Nk, M, c, st))
            return e;
        return run_scatter_wsum<T>("av_bwd", (const T *)attn, (const T *)d_out, plan, (T *)d_v, B, H, Nq, Nk, M, c, st);
    });
}

extern "C" int affb200_wf_fwd(const void *w, const void *f, const int64_t *idx, const affb200_nbhd_plan *plan,
                              void *out, int64_t B, int64_t Nq, int64_t Nk, int64_t M, int64_t IC, int64_t C,
                              int dtype, void *stream) {
    (void)plan;
    if (int e = check_dims("wf_fwd", B, IC, Nq, Nk, M, C)) return e;
    AFF_REQUIRE(B * Nq * IC * C == 0 || (out && (M == 0 || (w && f && idx))), "null pointer");
    return dispatch_dtype(dtype, [&](auto tag) {
        using T = decltype(tag);
        return launch_wf_fwd<T>((const T *)w, (const T *)f, idx, (T *)out, B, Nq, Nk, M, IC, C, (cudaStream_t)stream);
    });
}

extern "C" int affb200_wf_bwd(const void *d_out, const void *w, const void *f, const int64_t *idx,
                              const affb200_nbhd_plan *plan, void *d_w, void *d_f, int64_t B,
                              int64_t Nq, int64_t Nk, int64_t M, int64_t IC, int64_t C, int dtype, void *stream) {
    if (int e = check_dims("wf_bwd", B, IC, Nq, Nk, M, C)) return e;
    AFF_REQUIRE(B * Nq * M * IC * C == 0 || (d_out && w && f && idx && d_w), "null pointer");
    AFF_REQUIRE(B * Nk * C == 0 || d_f, "null pointer");
    AFF_REQUIRE(B * Nq * M == 0 || has_csr(plan), "transposed index (csr_off/csr_ent) is required");
    return dispatch_dtype(dtype, [&](auto tag) {
        using T = decltype(tag);
        auto st = (cudaStream_t)stream;
        if (int e = launch_wf_bwd_w<T>((const T *)d_out, (const T *)f, idx, (T *)d_w, B, Nq, Nk, M, IC, C, st))
            return e;
        return launch_wf_bwd_f<T>((const T *)d_out, (const T *)w, plan ? plan->csr_off : nullptr,
                                  plan ? plan->csr_ent : nullptr, (T *)d_f, B, Nq, Nk, M, IC, C, st);
    });
}
