// index_build.cu — index builders of the off-grid attention path (sm_100a):
//   balanced space-filling clustering   (reference models/point_utils.py:61-208, default variant)
//   stable fp32 argsort                 (torch.sort call sites: point_utils.py:159, mixres_neighbour.py:735)
//   brute-force k nearest neighbours    (pykeops argKmin / Kmin_argKmin: point_utils.py:49-57)
//   neighbourhood / mask / pe_idx build (models/aff_transformer.py:476-487; mixres_neighbour.py:518-528)
// Integer outputs are bit-exact with the reference's: every fp32 operation is issued in the reference's order with
// explicit round-to-nearest intrinsics (no FMA contraction), and orderings follow the rules of the library kernels
// the reference calls (stable ascending sort; kNN ties -> lowest index).
//
// Sorting works on 64-bit composites (key bits, token index) — unique, so the result is the stable sort by
// construction: tiles of 2048 are sorted in shared memory and merged by rank (see tile_sort_kernel below).  The
// slices are short (N <= 30 720 tokens per image); a radix sort's multi-pass global traffic and fixed launch chain
// (torch.sort: ~70 us at every N here) is what this avoids.
#include "common.cuh"
#include <algorithm>
#include <cmath>

namespace affb200 {

constexpr int kIT = 256;  // threads per CTA of the index kernels

__device__ __forceinline__ uint32_t f32_sortable(float x) {
    const uint32_t u = __float_as_uint(x);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// ---------------------------------------------------------------------------------------------------------
// Space-filling key, part 1: patch rank and distance ratio per token, running maximum of the ratio.
// Geometry (host-computed exactly as point_utils.py:89-97): npw x nph patches of size (plw, plh) in fp32.
// Patches are scanned row by row with alternating direction (:124-133), so rank r <-> (y = r / npw,
// x = y odd ? npw-1-r%npw : r%npw); patch centre = ((g*pl) + pl/2) - 0.5 (:138).
// ---------------------------------------------------------------------------------------------------------
struct SfcGeom {
    int npw, nph, nump;
    float plw, plh;
};

__device__ __forceinline__ void sfc_centre(const SfcGeom &g, int r, float &cx, float &cy) {
    const int y = r / g.npw, xr = r - y * g.npw;
    const int x = (y & 1) ? g.npw - 1 - xr : xr;
    cx = __fsub_rn(__fadd_rn(__fmul_rn(float(x), g.plw), __fdiv_rn(g.plw, 2.0f)), 0.5f);
    cy = __fsub_rn(__fadd_rn(__fmul_rn(float(y), g.plh), __fdiv_rn(g.plh, 2.0f)), 0.5f);
}

__device__ __forceinline__ void sfc_prev_next(const SfcGeom &g, int r, float &px, float &py, float &nx, float &ny) {
    float c0x, c0y, c1x, c1y;
    if (r > 0) {
        sfc_centre(g, r - 1, px, py);
    } else {  // prev[0] = c[0] - (c[1] - c[0])   (:143)
        sfc_centre(g, 0, c0x, c0y);
        sfc_centre(g, 1, c1x, c1y);
        px = __fsub_rn(c0x, __fsub_rn(c1x, c0x));
        py = __fsub_rn(c0y, __fsub_rn(c1y, c0y));
    }
    if (r < g.nump - 1) {
        sfc_centre(g, r + 1, nx, ny);
    } else {  // next[last] = c[last] + (c[last] - c[last-1])   (:146)
        sfc_centre(g, g.nump - 1, c0x, c0y);
        sfc_centre(g, g.nump - 2, c1x, c1y);
        nx = __fadd_rn(c0x, __fsub_rn(c0x, c1x));
        ny = __fadd_rn(c0y, __fsub_rn(c0y, c1y));
    }
}

__global__ void __launch_bounds__(kIT)
sfc_ratio_kernel(const float *__restrict__ pos, int64_t total, SfcGeom g, int32_t *__restrict__ rank,
                 float *__restrict__ ratio, uint32_t *__restrict__ max_bits) {
    const int64_t t = int64_t(blockIdx.x) * kIT + threadIdx.x;
    float r = 0.f;
    if (t < total) {
        const float2 p = __ldg(reinterpret_cast<const float2 *>(pos) + t);
        const float cx = floorf(__fdiv_rn(p.x, g.plw)), cy = floorf(__fdiv_rn(p.y, g.plh));
        const long long patch = (long long)__fadd_rn(cx, __fmul_rn(cy, float(g.npw)));  // (:149-150)
        const int y = int(patch / g.npw), x = int(patch - (long long)y * g.npw);
        const int rk = y * g.npw + ((y & 1) ? g.npw - 1 - x : x);
        float px, py, nx, ny;
        sfc_prev_next(g, rk, px, py, nx, ny);
        const float dxp = __fsub_rn(p.x, px), dyp = __fsub_rn(p.y, py);
        const float dxn = __fsub_rn(p.x, nx), dyn = __fsub_rn(p.y, ny);
        const float dp = __fadd_rn(__fmul_rn(dxp, dxp), __fmul_rn(dyp, dyp));
        const float dn = __fadd_rn(__fmul_rn(dxn, dxn), __fmul_rn(dyn, dyn));
        r = __fdiv_rn(dp, __fadd_rn(dn, 1e-5f));  // (:156)
        rank[t] = rk;
        ratio[t] = r;
    }
    // ratio >= 0, so the uint bit pattern orders like the float; one atomic per warp
    uint32_t m = __float_as_uint(r);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m) atomicMax(max_bits, m);
}

// part 2: key = rank * (max + 1) + ratio (:158) as a 64-bit composite (key bits, token index)
__global__ void __launch_bounds__(kIT)
sfc_key_kernel(const int32_t *__restrict__ rank, const float *__restrict__ ratio, const uint32_t *__restrict__ max_bits,
               int64_t total, int N, uint64_t *__restrict__ comp) {
    const int64_t t = int64_t(blockIdx.x) * kIT + threadIdx.x;
    if (t >= total) return;
    const float scale = __fadd_rn(__uint_as_float(*max_bits), 1.0f);
    const float key = __fadd_rn(__fmul_rn(float(rank[t]), scale), ratio[t]);
    comp[t] = (uint64_t(f32_sortable(key)) << 32) | uint32_t(t % N);
}

// ---------------------------------------------------------------------------------------------------------
// Peano / Hilbert curve keys (point_utils.py:211-282 / :285-349): per level the cell of the token inside the
// current square (floor(rem / cur_h), then rem %= cur_h, fp32 like the reference's tensor ops), mapped through the
// curve's digit table under the orientation accumulated from the coarser levels; the key is the fp32 sum
// orders[L-1] + sum_i orders[i] * base^(L-i-1) in the reference's order of additions.
// Peano (:236-268): digit table [[2,3,8],[1,4,7],[0,5,6]] indexed x + 3y; a level flips x of the finer levels
// when its y is 1 and flips y when its x is 1 (x parity for the widened first level, :229-233).
// Hilbert (:303-340): rotation state r; digit = rot_res[r mod 4][x + 2y]; r += rotate_dict[r mod 3][x + 2y].
// ---------------------------------------------------------------------------------------------------------
constexpr int kCurveMaxLevels = 16;
struct CurveGeom {
    int type;      // 1 peano, 2 hilbert
    int levels;
    int first_w;   // widened first level (peano, h != w), 0 when absent
    float cur_h[kCurveMaxLevels];
    float weight[kCurveMaxLevels];   // base^(L-i-1) as fp32
};

__device__ __forceinline__ int pymod(int a, int n) { const int r = a % n; return r < 0 ? r + n : r; }

__global__ void __launch_bounds__(kIT)
curve_key_kernel(const float *__restrict__ pos, int64_t total, CurveGeom g, float *__restrict__ key) {
    const int64_t t = int64_t(blockIdx.x) * kIT + threadIdx.x;
    if (t >= total) return;
    const float2 p = __ldg(reinterpret_cast<const float2 *>(pos) + t);
    float rx = p.x, ry = p.y;
    float ord[kCurveMaxLevels];
    int invx = 1, invy = 1, rot = 0;
    const int pe_tab[9] = {2, 3, 8, 1, 4, 7, 0, 5, 6};
    const int hil_rot[12] = {-1, 1, 0, 0, 0, -1, 0, 1, 1, 0, -1, 0};
    const int hil_res[16] = {0, 3, 1, 2, 2, 3, 1, 0, 2, 1, 3, 0, 0, 1, 3, 2};
#pragma unroll
    for (int i = 0; i < kCurveMaxLevels; ++i) {
        if (i < g.levels) {
            const float ch = g.cur_h[i];
            const int lx = int(floorf(__fdiv_rn(rx, ch))), ly = int(floorf(__fdiv_rn(ry, ch)));
            rx = fmodf(rx, ch);
            ry = fmodf(ry, ch);
            if (g.type == 1) {
                const int ix = invx > 0 ? lx : 2 - lx, iy = invy > 0 ? ly : 2 - ly;
                int o;
                if (i == 0 && g.first_w) {
                    const int c6 = ix % 6;
                    const int base = c6 < 3 ? pe_tab[min(max(iy, 0), 2) * 3 + c6] : pe_tab[(2 - min(max(iy, 0), 2)) * 3 + c6 - 3];
                    o = base + (ix / 3) * 9;
                    invy *= (lx & 1) ? -1 : 1;
                } else {
                    o = pe_tab[min(max(iy, 0), 2) * 3 + min(max(ix, 0), 2)];
                    invy *= lx == 1 ? -1 : 1;
                }
                invx *= ly == 1 ? -1 : 1;
                ord[i] = float(o);
            } else {
                const int cell = min(max(lx + ly * 2, 0), 3);
                ord[i] = float(hil_res[pymod(rot, 4) * 4 + cell]);
                rot += hil_rot[pymod(rot, 3) * 4 + cell];
            }
        } else {
            ord[i] = 0.f;
        }
    }
    float k = 0.f;
#pragma unroll
    for (int i = 0; i < kCurveMaxLevels; ++i)
        if (i == g.levels - 1) k = ord[i];
#pragma unroll
    for (int i = 0; i < kCurveMaxLevels; ++i)
        if (i < g.levels - 1) k = __fadd_rn(k, __fmul_rn(ord[i], g.weight[i]));
    key[t] = k;
}

// Clustering key with the anchor order given as tables (Peano / Hilbert anchors, point_utils.py:104-122,135-156):
// grid_rank[cell] = position of the anchor cell in the curve, prev/next[rank] = centres of the neighbouring anchors.
__global__ void __launch_bounds__(kIT)
sfc_ratio_table_kernel(const float *__restrict__ pos, int64_t total, int npw, int nump, float plw, float plh,
                       const int32_t *__restrict__ grid_rank, const float2 *__restrict__ prev,
                       const float2 *__restrict__ next, int32_t *__restrict__ rank, float *__restrict__ ratio,
                       uint32_t *__restrict__ max_bits) {
    const int64_t t = int64_t(blockIdx.x) * kIT + threadIdx.x;
    float r = 0.f;
    if (t < total) {
        const float2 p = __ldg(reinterpret_cast<const float2 *>(pos) + t);
        const float cx = floorf(__fdiv_rn(p.x, plw)), cy = floorf(__fdiv_rn(p.y, plh));
        long long cell = (long long)__fadd_rn(cx, __fmul_rn(cy, float(npw)));
        cell = cell < 0 ? 0 : (cell >= nump ? nump - 1 : cell);
        const int rk = __ldg(grid_rank + cell);
        const float2 pv = __ldg(prev + rk), nx = __ldg(next + rk);
        const float dxp = __fsub_rn(p.x, pv.x), dyp = __fsub_rn(p.y, pv.y);
        const float dxn = __fsub_rn(p.x, nx.x), dyn = __fsub_rn(p.y, nx.y);
        const float dp = __fadd_rn(__fmul_rn(dxp, dxp), __fmul_rn(dyp, dyp));
        const float dn = __fadd_rn(__fmul_rn(dxn, dxn), __fmul_rn(dyn, dyn));
        r = __fdiv_rn(dp, __fadd_rn(dn, 1e-5f));
        rank[t] = rk;
        ratio[t] = r;
    }
    uint32_t m = __float_as_uint(r);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m) atomicMax(max_bits, m);
}

__global__ void __launch_bounds__(kIT)
f32_key_kernel(const float *__restrict__ x, int64_t total, int N, int descending, uint64_t *__restrict__ comp) {
    const int64_t t = int64_t(blockIdx.x) * kIT + threadIdx.x;
    if (t >= total) return;
    uint32_t k = f32_sortable(__ldg(x + t));
    if (descending) k = ~k;
    comp[t] = (uint64_t(k) << 32) | uint32_t(t % N);
}

// ---------------------------------------------------------------------------------------------------------
// Sort of 64-bit composites (key bits << 32 | token index: unique, so any comparison sort is the stable sort).
//   1. tile_sort_kernel: every CTA sorts one tile of 2048 composites in shared memory (bitonic network).  A slice
//      that fits one tile is finished here: order / pos_sorted are written directly.
//   2. tile_rank_kernel: rank(e) = position in its own sorted tile + sum over the other tiles of lower_bound(tile, e),
//      a binary search in the tile staged in shared memory.  grid (T, B, JS): CTA (ti, b, js) ranks the elements of
//      tile ti against the js-th slice of the tiles; partial ranks meet in one atomicAdd per element when JS > 1.
//   3. rank_scatter_kernel: order[b][rank] = i.
// Work per slice is N log(2048) + N (N / 2048) log(2048) compares instead of the N^2 of a plain rank count (30 720
// tokens: 5.4 M instead of 944 M), still without multi-pass global traffic; measured in profiles/r02_sort.md.
// ---------------------------------------------------------------------------------------------------------
constexpr int kSortTile = 2048;     // composites per shared-memory tile (16 KB)
constexpr int kSortThreads = 1024;
constexpr uint64_t kSortPad = ~0ull;   // above every composite (token index < N <= 2^31)

__global__ void __launch_bounds__(kSortThreads)
tile_sort_kernel(const uint64_t *__restrict__ comp, int N, uint64_t *__restrict__ sorted, int64_t *__restrict__ order,
                 const float *__restrict__ pos, float *__restrict__ pos_sorted) {
    __shared__ uint64_t s[kSortTile];
    const int tid = threadIdx.x, b = blockIdx.y, base = blockIdx.x * kSortTile;
    const int nt = min(kSortTile, N - base);
    int n2 = 64;
    while (n2 < nt) n2 <<= 1;
    const uint64_t *cb = comp + int64_t(b) * N + base;
    for (int t = tid; t < n2; t += kSortThreads) s[t] = t < nt ? cb[t] : kSortPad;
    __syncthreads();
    for (int k = 2; k <= n2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < (n2 >> 1); t += kSortThreads) {
                const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                const int l = i | j;
                const uint64_t x = s[i], y = s[l];
                if ((x > y) == ((i & k) == 0)) { s[i] = y; s[l] = x; }
            }
            __syncthreads();
        }
    }
    if (gridDim.x == 1) {     // the whole slice: finished
        for (int t = tid; t < nt; t += kSortThreads) {
            const int i = int(uint32_t(s[t]));
            const int64_t dst = int64_t(b) * N + t;
            order[dst] = i;
            if (pos) reinterpret_cast<float2 *>(pos_sorted)[dst] = __ldg(reinterpret_cast<const float2 *>(pos) + int64_t(b) * N + i);
        }
    } else {
        uint64_t *out = sorted + (int64_t(b) * gridDim.x + blockIdx.x) * kSortTile;
        for (int t = tid; t < kSortTile; t += kSortThreads) out[t] = t < n2 ? s[t] : kSortPad;
    }
}

__global__ void __launch_bounds__(kSortThreads)
tile_rank_kernel(const uint64_t *__restrict__ sorted, int N, int T, int tiles_per_slice, int32_t *__restrict__ rank_out) {
    __shared__ uint64_t s[kSortTile];
    const int tid = threadIdx.x, ti = blockIdx.x, b = blockIdx.y;
    const uint64_t *sb = sorted + int64_t(b) * T * kSortTile;
    const uint64_t e0 = sb[int64_t(ti) * kSortTile + tid], e1 = sb[int64_t(ti) * kSortTile + kSortThreads + tid];
    const int j0 = blockIdx.z * tiles_per_slice, j1 = min(T, j0 + tiles_per_slice);
    int c0 = 0, c1 = 0;
    for (int tj = j0; tj < j1; ++tj) {
        if (tj == ti) { c0 += tid; c1 += kSortThreads + tid; continue; }
        __syncthreads();
        for (int t = tid; t < kSortTile; t += kSortThreads) s[t] = sb[int64_t(tj) * kSortTile + t];
        __syncthreads();
        int p0 = 0, p1 = 0;     // lower_bound over 2048 entries (pads compare above everything)
#pragma unroll
        for (int step = kSortTile >> 1; step > 0; step >>= 1) {
            p0 += s[p0 + step - 1] < e0 ? step : 0;
            p1 += s[p1 + step - 1] < e1 ? step : 0;
        }
        c0 += p0 + (s[p0] < e0 ? 1 : 0);
        c1 += p1 + (s[p1] < e1 ? 1 : 0);
    }
    if (e0 != kSortPad) {
        int32_t *r = rank_out + int64_t(b) * N + int(uint32_t(e0));
        if (gridDim.z == 1) *r = c0; else atomicAdd(r, c0);
    }
    if (e1 != kSortPad) {
        int32_t *r = rank_out + int64_t(b) * N + int(uint32_t(e1));
        if (gridDim.z == 1) *r = c1; else atomicAdd(r, c1);
    }
}

// order[b][rank[b][i]] = i ; optionally pos_sorted[b][rank] = pos[b][i]
__global__ void __launch_bounds__(kIT)
rank_scatter_kernel(const int32_t *__restrict__ rank, int64_t total, int N, int64_t *__restrict__ order,
                    const float *__restrict__ pos, float *__restrict__ pos_sorted) {
    const int64_t t = int64_t(blockIdx.x) * kIT + threadIdx.x;
    if (t >= total) return;
    const int64_t b = t / N;
    const int i = int(t - b * N);
    const int64_t dst = b * N + rank[t];
    order[dst] = i;
    if (pos) reinterpret_cast<float2 *>(pos_sorted)[dst] = __ldg(reinterpret_cast<const float2 *>(pos) + t);
}

// cluster centres: mean of the m (or fewer, last cluster) sorted positions of each cluster (:183-192)
__global__ void __launch_bounds__(kIT)
cluster_mean_kernel(const float *__restrict__ pos_sorted, int64_t total, int N, int K, int m, float *__restrict__ mean) {
    const int64_t t = int64_t(blockIdx.x) * kIT + threadIdx.x;
    if (t >= total) return;
    const int64_t b = t / K;
    const int c = int(t - b * K);
    const int i0 = c * m, cnt = min(m, N - i0);
    const float2 *p = reinterpret_cast<const float2 *>(pos_sorted) + b * N + i0;
    float sx = 0.f, sy = 0.f;
    for (int q = 0; q < cnt; ++q) {
        const float2 v = __ldg(p + q);
        sx = __fadd_rn(sx, v.x);
        sy = __fadd_rn(sy, v.y);
    }
    reinterpret_cast<float2 *>(mean)[t] = make_float2(__fdiv_rn(sx, float(cnt)), __fdiv_rn(sy, float(cnt)));
}

// ---------------------------------------------------------------------------------------------------------
// k nearest neighbours, brute force.  One thread per query keeps its K best (distance, index) pairs sorted
// ascending; the database is streamed through shared memory in index order and a candidate is inserted only on a
// strict '<' against the current entries behind it, so equal distances keep the lower index first — the rule of
// KeOps' ArgKMin reduction.  d = sqrt(dx*dx + dy*dy) in fp32 as the reference's LazyTensor formula
// ((q-x)**2).sum(-1)**0.5 (point_utils.py:52); comparisons are made on d (the rounded square root), not on d².
// ---------------------------------------------------------------------------------------------------------
constexpr int kKnnTile = 2048;

template <int K>
__global__ void __launch_bounds__(kIT)
knn2d_kernel(const float *__restrict__ query, const float *__restrict__ db, int Nq, int Nd, int k_rt,
             int64_t *__restrict__ idx_out, float *__restrict__ dist_out) {
    __shared__ float2 tile[kKnnTile];
    const int b = blockIdx.y;
    const int i = blockIdx.x * kIT + threadIdx.x;
    const bool live = i < Nq;
    float2 q = make_float2(0.f, 0.f);
    if (live) q = __ldg(reinterpret_cast<const float2 *>(query) + int64_t(b) * Nq + i);
    float bd[K];
    int bi[K];
#pragma unroll
    for (int t = 0; t < K; ++t) { bd[t] = __int_as_float(0x7f800000); bi[t] = -1; }
    const float2 *dbb = reinterpret_cast<const float2 *>(db) + int64_t(b) * Nd;
    for (int base = 0; base < Nd; base += kKnnTile) {
        const int nt = min(kKnnTile, Nd - base);
        __syncthreads();
        for (int t = threadIdx.x; t < nt; t += kIT) tile[t] = __ldg(dbb + base + t);
        __syncthreads();
        if (!live) continue;
        for (int t = 0; t < nt; ++t) {
            const float2 x = tile[t];
            const float dx = __fsub_rn(q.x, x.x), dy = __fsub_rn(q.y, x.y);
            const float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
            if (d < bd[K - 1] || bi[K - 1] < 0) {
                // insertion keeping ascending order; '<' keeps earlier (lower-index) entries ahead on ties
                float cd = d;
                int ci = base + t;
                bool shifting = false;  // once the candidate has found its slot, everything behind it moves down
#pragma unroll
                for (int s = 0; s < K; ++s) {
                    const bool take = shifting || (bi[s] < 0) || (cd < bd[s]);
                    const float td = bd[s];
                    const int tix = bi[s];
                    if (take) { bd[s] = cd; bi[s] = ci; cd = td; ci = tix; shifting = true; }
                }
            }
        }
    }
    if (live) {
        int64_t *o = idx_out + (int64_t(b) * Nq + i) * k_rt;
        for (int t = 0; t < k_rt; ++t) o[t] = bi[t];
        if (dist_out) {
            float *od = dist_out + (int64_t(b) * Nq + i) * k_rt;
            for (int t = 0; t < k_rt; ++t) od[t] = bd[t];
        }
    }
}

// Nearest database point in L1 distance, first index on ties: `argmin(cdist(query, db, p=1), dim=2)` — the reference's
// find_pos_org_order (maskfiner_up_down.py:150-154), which materialises the Nq x Nd distance matrix (N up to 16 384).
// Same shared-memory tiling of the database as knn2d; the scan stops improving only on a strictly smaller distance,
// so the lowest index wins ties as torch.argmin does.
__global__ void __launch_bounds__(kIT)
nearest_l1_kernel(const float *__restrict__ query, const float *__restrict__ db, int Nq, int Nd,
                  int64_t *__restrict__ idx_out) {
    __shared__ float2 tile[kKnnTile];
    const int b = blockIdx.y;
    const int i = blockIdx.x * kIT + threadIdx.x;
    const bool live = i < Nq;
    float2 q = make_float2(0.f, 0.f);
    if (live) q = __ldg(reinterpret_cast<const float2 *>(query) + int64_t(b) * Nq + i);
    float best = __int_as_float(0x7f800000);
    int bi = 0;
    const float2 *dbb = reinterpret_cast<const float2 *>(db) + int64_t(b) * Nd;
    for (int base = 0; base < Nd; base += kKnnTile) {
        const int nt = min(kKnnTile, Nd - base);
        __syncthreads();
        for (int t = threadIdx.x; t < nt; t += kIT) tile[t] = __ldg(dbb + base + t);
        __syncthreads();
        if (!live) continue;
        for (int t = 0; t < nt; ++t) {
            const float2 x = tile[t];
            const float d = __fadd_rn(fabsf(__fsub_rn(q.x, x.x)), fabsf(__fsub_rn(q.y, x.y)));
            if (d < best) { best = d; bi = base + t; }
        }
    }
    if (live) idx_out[int64_t(b) * Nq + i] = bi;
}

// ---------------------------------------------------------------------------------------------------------
// Neighbourhood build (aff_transformer.py:478-487): for token i and its nnc nearest clusters,
//   nbhd_idx[i, q*m + r] = member(nearest[i,q]*m + r),  member(t) = t < N ? t : 0          (point_utils.py:203-206)
//   mask[i, q*m + r]     = (nearest[i,q]*m + r) < N                                         (:186-189)
//   rel = pos[nbhd] - (pos[i] - rel_pos_width) ; pe_idx = long(rel_y * table_width + rel_x) (:483-487)
// One thread per (token, slot); all three int64 outputs are written coalesced.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kIT)
nbhd_build_kernel(const float *__restrict__ pos, const int64_t *__restrict__ nearest, int64_t total, int N, int m,
                  int nnc, float rel_pos_width, float table_width, int clamp, int64_t *__restrict__ nbhd,
                  int64_t *__restrict__ mask, int64_t *__restrict__ pe_idx) {
    const int64_t t = int64_t(blockIdx.x) * kIT + threadIdx.x;
    if (t >= total) return;
    const int M = m * nnc;
    const int64_t bi = t / M;  // b*N + i
    const int slot = int(t - bi * M);
    const int q = slot / m, r = slot - q * m;
    const int64_t b = bi / N;
    const int64_t raw = __ldg(nearest + bi * nnc + q) * m + r;
    const bool valid = raw < N;
    const int64_t row = valid ? raw : 0;
    nbhd[t] = row;
    if (mask) mask[t] = valid ? 1 : 0;
    const float2 pi = __ldg(reinterpret_cast<const float2 *>(pos) + bi);
    const float2 pn = __ldg(reinterpret_cast<const float2 *>(pos) + b * N + row);
    float rx = __fsub_rn(pn.x, __fsub_rn(pi.x, rel_pos_width));
    float ry = __fsub_rn(pn.y, __fsub_rn(pi.y, rel_pos_width));
    if (clamp) {  // mixres_neighbour.py:527
        rx = fminf(fmaxf(rx, 0.f), table_width - 1.f);
        ry = fminf(fmaxf(ry, 0.f), table_width - 1.f);
    }
    pe_idx[t] = (int64_t)__fadd_rn(__fmul_rn(ry, table_width), rx);
}


// ---------------------------------------------------------------------------------------------------------
// Merge-centre selection (aff_transformer.py:295-326).  One CTA per image.
//   grid prior   stride 2: (pos % 2).sum == 0 ; else ada = 2^(ceil(log2 d2nd) + 1), (long(pos) % long(ada)).sum == 0
//   final        = grid + lp*alpha (product rounded in lp's dtype) + reserve_mask * (-100)
//   sampled      = torch.topk(final, sample, sorted=False) in its CUDA order: indices with final > v_k ascending,
//                  then indices with final == v_k ascending (v_k = the sample-th largest value)
//   reserved     = indices with (pos % 2*stride).sum == 0, ascending, appended after the sampled ones
// status[b] bit 0: reserved count != reserve_num (the reference's reshape would raise); bit 1: ada stride 0.
// ---------------------------------------------------------------------------------------------------------
constexpr int kSelT = 1024;

__device__ __forceinline__ int block_scan_1024(int v, int *warp_buf, int &total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int s = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, s, d);
        if (lane >= d) s += t;
    }
    __syncthreads();
    if (lane == 31) warp_buf[wid] = s;
    __syncthreads();
    if (wid == 0) {
        int ws = warp_buf[lane];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, ws, d);
            if (lane >= d) ws += t;
        }
        warp_buf[lane] = ws;
    }
    __syncthreads();
    total = warp_buf[31];
    return s - v + (wid > 0 ? warp_buf[wid - 1] : 0);
}

template <typename T>
__global__ void __launch_bounds__(kSelT)
merge_select_kernel(const float *__restrict__ pos, const T *__restrict__ lp, const float *__restrict__ d2nd /* [B][N][2] self-kNN distances */, int N,
                    int stride, float alpha, int sample, int reserve_num, int reserve_on,
                    float *__restrict__ final_out, int64_t *__restrict__ idx_out, int32_t *__restrict__ status) {
    extern __shared__ float fin[];  // [N]
    __shared__ int warp_buf[32];
    __shared__ int hist_s[256];
    __shared__ uint32_t sel_s[2];
    __shared__ int flags_s;
    const int b = blockIdx.x, tid = threadIdx.x;
    const float2 *pb = reinterpret_cast<const float2 *>(pos) + int64_t(b) * N;
    if (tid == 0) flags_s = 0;
    __syncthreads();
    for (int i = tid; i < N; i += kSelT) {
        const float2 p = __ldg(pb + i);
        float grid;
        if (stride == 2) {
            grid = (__fadd_rn(fmodf(p.x, 2.f), fmodf(p.y, 2.f)) == 0.f) ? 1.f : 0.f;
        } else {
            const float ada = exp2f(__fadd_rn(ceilf(log2f(__ldg(d2nd + (int64_t(b) * N + i) * 2 + 1))), 1.f));
            const long long al = (long long)ada;
            if (al == 0) { grid = 0.f; atomicOr(&flags_s, 2); }
            else grid = (((long long)p.x % al) + ((long long)p.y % al) == 0) ? 1.f : 0.f;
        }
        float f = grid;
        if (lp) {
            const float scaled = Elem<T>::to_acc(Elem<T>::from_acc(__fmul_rn(Elem<T>::to_acc(lp[int64_t(b) * N + i]), alpha)));
            f = __fadd_rn(f, scaled);
        }
        if (reserve_on) {
            const float s2 = float(stride * 2);
            const float rm = (__fadd_rn(fmodf(p.x, s2), fmodf(p.y, s2)) == 0.f) ? 1.f : 0.f;
            f = __fadd_rn(f, __fmul_rn(rm, -100.f));
        }
        fin[i] = f;
        final_out[int64_t(b) * N + i] = f;
    }
    __syncthreads();
    int64_t *out = idx_out + int64_t(b) * (sample + (reserve_on ? reserve_num : 0));
    // v_k: the element whose stable descending rank is sample-1
    if (sample > 0) {
        // block radix select over the order-preserving bit patterns: 4 passes of 8 bits from the top, each a shared
        // 256-bin histogram of the elements that still match the prefix (the O(N^2) rank count it replaces took
        // 0.6 ms at N = 3136)
        uint32_t prefix = 0u, mask = 0u;
        int need = sample;
        for (int shift = 24; shift >= 0; shift -= 8) {
            if (tid < 256) hist_s[tid] = 0;
            __syncthreads();
            for (int i = tid; i < N; i += kSelT) {
                const uint32_t key = f32_sortable(fin[i]);
                if ((key & mask) == prefix) atomicAdd(&hist_s[(key >> shift) & 255u], 1);
            }
            __syncthreads();
            if (tid < 32) {  // lane l owns bins 255-8l .. 248-8l, largest digit first
                int loc[8], s = 0;
#pragma unroll
                for (int k = 0; k < 8; ++k) { loc[k] = hist_s[255 - 8 * tid - k]; s += loc[k]; }
                int incl = s;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int t = __shfl_up_sync(0xffffffffu, incl, d);
                    if (tid >= d) incl += t;
                }
                int c = incl - s;
                if (c < need && need <= incl) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        if (c < need && need <= c + loc[k]) {
                            sel_s[0] = prefix | (uint32_t(255 - 8 * tid - k) << shift);
                            sel_s[1] = uint32_t(need - c);
                        }
                        c += loc[k];
                    }
                }
            }
            __syncthreads();
            prefix = sel_s[0];
            need = int(sel_s[1]);
            mask |= 0xffu << shift;
        }
        const float vk = __uint_as_float((prefix & 0x80000000u) ? (prefix ^ 0x80000000u) : ~prefix);
        int base = 0;
        for (int i0 = 0; i0 < N; i0 += kSelT) {  // pass 1: strictly greater, ascending index
            const int i = i0 + tid;
            const int has = (i < N && fin[i] > vk) ? 1 : 0;
            int tot;
            const int ex = block_scan_1024(has, warp_buf, tot);
            if (has) out[base + ex] = i;
            base += tot;
        }
        for (int i0 = 0; i0 < N && base < sample; i0 += kSelT) {  // pass 2: ties with v_k, ascending index
            const int i = i0 + tid;
            const int has = (i < N && fin[i] == vk) ? 1 : 0;
            int tot;
            const int ex = block_scan_1024(has, warp_buf, tot);
            if (has && base + ex < sample) out[base + ex] = i;
            base += tot;
        }
    }
    if (reserve_on) {
        const float s2 = float(stride * 2);
        int base = 0;
        for (int i0 = 0; i0 < N; i0 += kSelT) {
            const int i = i0 + tid;
            int has = 0;
            if (i < N) {
                const float2 p = __ldg(pb + i);
                has = (__fadd_rn(fmodf(p.x, s2), fmodf(p.y, s2)) == 0.f) ? 1 : 0;
            }
            int tot;
            const int ex = block_scan_1024(has, warp_buf, tot);
            if (has && base + ex < reserve_num) out[sample + base + ex] = i;
            base += tot;
        }
        if (tid == 0 && base != reserve_num) atomicOr(&flags_s, 1);
    }
    __syncthreads();
    if (tid == 0) status[b] = flags_s;
}

// gather rows of an int64 [B][N][M] tensor at idx [B][K] -> [B][K][M]  (aff_transformer.py:335-338), 16-byte accesses
__global__ void __launch_bounds__(kIT)
gather_rows_i64_kernel(const int64_t *__restrict__ src, const int64_t *__restrict__ idx, int64_t total, int N, int K,
                       int M2, int64_t *__restrict__ dst) {
    const int64_t t = int64_t(blockIdx.x) * kIT + threadIdx.x;  // over B*K*M2 (M2 = M/2 longlong2 per row)
    if (t >= total) return;
    const int64_t bk = t / M2;
    const int c = int(t - bk * M2);
    const int64_t b = bk / K;
    const int64_t row = __ldg(idx + bk);
    reinterpret_cast<longlong2 *>(dst)[t] = __ldg(reinterpret_cast<const longlong2 *>(src) + (b * N + row) * M2 + c);
}

}  // namespace affb200

// =========================================================================================================
// C ABI
// =========================================================================================================
using namespace affb200;

static inline size_t al16(size_t x) { return (x + 15) & ~size_t(15); }
static inline unsigned blocks_for(int64_t total) { return unsigned((total + kIT - 1) / kIT); }

// workspace of the sort: composites u64[B*N] | ranks i32[B*N] | sorted tiles u64[B*T*2048] (slices longer than a tile)
static size_t sort_work_bytes(int64_t B, int64_t N) {
    const int64_t T = (N + kSortTile - 1) / kSortTile;
    return al16(size_t(B * N) * 8) + al16(size_t(B * N) * 4) + (T > 1 ? size_t(B * T) * kSortTile * 8 : 0);
}

// comp -> order (and optionally pos_sorted); comp, rank and the sorted tiles live in `work` (rank follows comp)
static int rank_sort(uint64_t *comp, int32_t *rank, int64_t B, int64_t N, int64_t *order, const float *pos,
                     float *pos_sorted, cudaStream_t st) {
    const int T = int((N + kSortTile - 1) / kSortTile);
    uint64_t *sorted = reinterpret_cast<uint64_t *>(reinterpret_cast<char *>(rank) + al16(size_t(B * N) * 4));
    tile_sort_kernel<<<dim3(T, unsigned(B)), kSortThreads, 0, st>>>(comp, int(N), sorted, order, pos, pos_sorted);
    if (T == 1) return check_launch("tile_sort", 1);
    const int64_t ctas = int64_t(T) * B;
    int js = 1;
    if (ctas < 2 * kNumSMs) js = int(std::min<int64_t>((2 * kNumSMs + ctas - 1) / ctas, T));
    const int per = (T + js - 1) / js;
    js = (T + per - 1) / per;
    if (js > 1 && cudaMemsetAsync(rank, 0, size_t(B * N) * 4, st) != cudaSuccess) return check_launch("rank memset", 0);
    tile_rank_kernel<<<dim3(T, unsigned(B), js), kSortThreads, 0, st>>>(sorted, int(N), T, per, rank);
    rank_scatter_kernel<<<blocks_for(B * N), kIT, 0, st>>>(rank, B * N, int(N), order, pos, pos_sorted);
    return check_launch("rank_sort", 3);
}

static bool idx_dims_ok(const char *what, int64_t B, int64_t N) {
    if (B < 0 || N < 0 || B > 65535 || N >= (int64_t(1) << 31) || B * N >= (int64_t(1) << 40)) {
        set_error("%s: unsupported dimensions (0 <= B <= 65535, N < 2^31)", what);
        return false;
    }
    return true;
}

extern "C" size_t affb200_sfc_work_bytes(int64_t B, int64_t N) {
    return al16(size_t(B * N) * 4) * 2 + 16 + sort_work_bytes(B, N);
}

extern "C" int affb200_sfc_cluster(const float *pos, int64_t B, int64_t N, int64_t m, int64_t h, int64_t w,
                                   int64_t *ranking, float *pos_sorted, float *mean, void *work, void *stream) {
    if (!idx_dims_ok("sfc_cluster", B, N)) return AFFB200_E_ARG;
    if (m <= 0 || h <= 0 || w <= 0) { set_error("sfc_cluster: m, h, w must be positive"); return AFFB200_E_ARG; }
    if (B * N == 0) return AFFB200_OK;
    if (!pos || !ranking || !pos_sorted || !mean || !work) { set_error("sfc_cluster: null pointer"); return AFFB200_E_ARG; }
    // anchor-grid geometry in double, then fp32 patch lengths, as point_utils.py:89-97,137
    const int64_t K = (N + m - 1) / m;
    const double patch_len = std::sqrt(double(h) * double(w) / double(K));
    const int nph = int(std::nearbyint(double(h) / patch_len));  // Python round(): half to even, as nearbyint
    const int npw = int(std::nearbyint(double(w) / patch_len));
    if (nph < 1 || npw < 1 || int64_t(nph) * npw < 3) {
        set_error("sfc_cluster: anchor grid %dx%d too small (the reference needs >= 3 patches)", nph, npw);
        return AFFB200_E_ARG;
    }
    SfcGeom g;
    g.npw = npw; g.nph = nph; g.nump = nph * npw;
    g.plw = float(double(w) / npw); g.plh = float(double(h) / nph);
    auto st = (cudaStream_t)stream;
    char *wp = reinterpret_cast<char *>(work);
    int32_t *rank = reinterpret_cast<int32_t *>(wp); wp += al16(size_t(B * N) * 4);
    float *ratio = reinterpret_cast<float *>(wp);    wp += al16(size_t(B * N) * 4);
    uint32_t *max_bits = reinterpret_cast<uint32_t *>(wp); wp += 16;
    uint64_t *comp = reinterpret_cast<uint64_t *>(wp); wp += al16(size_t(B * N) * 8);
    int32_t *prank = reinterpret_cast<int32_t *>(wp);
    if (cudaMemsetAsync(max_bits, 0, 16, st) != cudaSuccess) return check_launch("sfc memset", 0);
    sfc_ratio_kernel<<<blocks_for(B * N), kIT, 0, st>>>(pos, B * N, g, rank, ratio, max_bits);
    sfc_key_kernel<<<blocks_for(B * N), kIT, 0, st>>>(rank, ratio, max_bits, B * N, int(N), comp);
    if (int e = check_launch("sfc_key", 2)) return e;
    if (int e = rank_sort(comp, prank, B, N, ranking, pos, pos_sorted, st)) return e;
    cluster_mean_kernel<<<blocks_for(B * K), kIT, 0, st>>>(pos_sorted, B * K, int(N), int(K), int(m), mean);
    return check_launch("cluster_mean");
}

// Python's math.log(x, base) is log(x)/log(base) in doubles, round() is half-to-even like nearbyint
static int curve_geom(int64_t h, int64_t w, int curve, CurveGeom &g) {
    const double base = curve == 1 ? 3.0 : 2.0;
    g.type = curve;
    g.levels = int(std::ceil(std::log(double(h)) / std::log(base)));
    if (g.levels < 1) { set_error("curve_key: h too short"); return AFFB200_E_ARG; }
    if (g.levels > kCurveMaxLevels) { set_error("curve_key: more than %d levels", kCurveMaxLevels); return AFFB200_E_ARG; }
    g.first_w = 0;
    if (h != w) {
        const int fw = int(std::nearbyint(base * (double(w) / double(h))));
        if (fw != int(base)) g.first_w = fw;
    }
    if (curve == 2 && g.first_w) {
        set_error("curve_key: the Hilbert order needs round(2*w/h) == 2 (the reference indexes a 4-entry table with "
                  "x + y*first_w, point_utils.py:322-338)");
        return AFFB200_E_ARG;
    }
    if (curve == 1 && g.first_w < 0) { set_error("curve_key: bad aspect ratio"); return AFFB200_E_ARG; }
    double last = double(h);
    for (int i = 0; i < kCurveMaxLevels; ++i) { g.cur_h[i] = 1.f; g.weight[i] = 0.f; }
    for (int i = 0; i < g.levels; ++i) {
        last = last / base;
        g.cur_h[i] = float(last);
        g.weight[i] = float(std::pow(base * base, double(g.levels - i - 1)));
    }
    return AFFB200_OK;
}

extern "C" int affb200_curve_key(const float *pos, int64_t B, int64_t N, int64_t h, int64_t w, int curve, float *key,
                                 void *stream) {
    if (!idx_dims_ok("curve_key", B, N)) return AFFB200_E_ARG;
    if (curve != 1 && curve != 2) { set_error("curve_key: curve must be 1 (peano) or 2 (hilbert)"); return AFFB200_E_ARG; }
    if (h <= 0 || w <= 0) { set_error("curve_key: h, w must be positive"); return AFFB200_E_ARG; }
    if (B * N == 0) return AFFB200_OK;
    if (!pos || !key) { set_error("curve_key: null pointer"); return AFFB200_E_ARG; }
    CurveGeom g;
    if (int e = curve_geom(h, w, curve, g)) return e;
    curve_key_kernel<<<blocks_for(B * N), kIT, 0, (cudaStream_t)stream>>>(pos, B * N, g, key);
    return check_launch("curve_key");
}

extern "C" int affb200_sfc_cluster_tables(const float *pos, int64_t B, int64_t N, int64_t m, int64_t npw, int64_t nump,
                                          float plw, float plh, const int32_t *grid_rank, const float *prev,
                                          const float *next, int64_t *ranking, float *pos_sorted, float *mean,
                                          void *work, void *stream) {
    if (!idx_dims_ok("sfc_cluster_tables", B, N)) return AFFB200_E_ARG;
    if (m <= 0 || npw <= 0 || nump <= 0 || nump >= (int64_t(1) << 31)) {
        set_error("sfc_cluster_tables: m, npw, nump must be positive");
        return AFFB200_E_ARG;
    }
    if (B * N == 0) return AFFB200_OK;
    if (!pos || !ranking || !pos_sorted || !mean || !work || !grid_rank || !prev || !next) {
        set_error("sfc_cluster_tables: null pointer");
        return AFFB200_E_ARG;
    }
    const int64_t K = (N + m - 1) / m;
    auto st = (cudaStream_t)stream;
    char *wp = reinterpret_cast<char *>(work);
    int32_t *rank = reinterpret_cast<int32_t *>(wp); wp += al16(size_t(B * N) * 4);
    float *ratio = reinterpret_cast<float *>(wp);    wp += al16(size_t(B * N) * 4);
    uint32_t *max_bits = reinterpret_cast<uint32_t *>(wp); wp += 16;
    uint64_t *comp = reinterpret_cast<uint64_t *>(wp); wp += al16(size_t(B * N) * 8);
    int32_t *prank = reinterpret_cast<int32_t *>(wp);
    if (cudaMemsetAsync(max_bits, 0, 16, st) != cudaSuccess) return check_launch("sfc memset", 0);
    sfc_ratio_table_kernel<<<blocks_for(B * N), kIT, 0, st>>>(pos, B * N, int(npw), int(nump), plw, plh, grid_rank,
                                                              reinterpret_cast<const float2 *>(prev),
                                                              reinterpret_cast<const float2 *>(next), rank, ratio,
                                                              max_bits);
    sfc_key_kernel<<<blocks_for(B * N), kIT, 0, st>>>(rank, ratio, max_bits, B * N, int(N), comp);
    if (int e = check_launch("sfc_key", 2)) return e;
    if (int e = rank_sort(comp, prank, B, N, ranking, pos, pos_sorted, st)) return e;
    cluster_mean_kernel<<<blocks_for(B * K), kIT, 0, st>>>(pos_sorted, B * K, int(N), int(K), int(m), mean);
    return check_launch("cluster_mean");
}

extern "C" size_t affb200_argsort_work_bytes(int64_t B, int64_t N) { return sort_work_bytes(B, N); }

extern "C" int affb200_argsort_f32(const float *x, int64_t B, int64_t N, int descending, int64_t *order, void *work,
                                   void *stream) {
    if (!idx_dims_ok("argsort_f32", B, N)) return AFFB200_E_ARG;
    if (B * N == 0) return AFFB200_OK;
    if (!x || !order || !work) { set_error("argsort_f32: null pointer"); return AFFB200_E_ARG; }
    auto st = (cudaStream_t)stream;
    uint64_t *comp = reinterpret_cast<uint64_t *>(work);
    int32_t *prank = reinterpret_cast<int32_t *>(reinterpret_cast<char *>(work) + al16(size_t(B * N) * 8));
    f32_key_kernel<<<blocks_for(B * N), kIT, 0, st>>>(x, B * N, int(N), descending, comp);
    if (int e = check_launch("argsort key")) return e;
    return rank_sort(comp, prank, B, N, order, nullptr, nullptr, st);
}

extern "C" int affb200_knn2d(const float *query, const float *db, int64_t B, int64_t Nq, int64_t Nd, int64_t k,
                             int64_t *idx, float *dist, void *stream) {
    if (!idx_dims_ok("knn2d", B, Nq) || !idx_dims_ok("knn2d", B, Nd)) return AFFB200_E_ARG;
    if (k < 0 || k > 32 || (k > Nd && B * Nq > 0)) {
        set_error("knn2d: k must be in [0, min(32, Nd)]");
        return AFFB200_E_ARG;
    }
    if (B * Nq * k == 0) return AFFB200_OK;
    if (!query || !db || !idx) { set_error("knn2d: null pointer"); return AFFB200_E_ARG; }
    auto st = (cudaStream_t)stream;
    dim3 grid(unsigned((Nq + kIT - 1) / kIT), unsigned(B));
#define AFF_KNN(K_) knn2d_kernel<K_><<<grid, kIT, 0, st>>>(query, db, int(Nq), int(Nd), int(k), idx, dist)
    if (k <= 2) AFF_KNN(2);
    else if (k <= 6) AFF_KNN(6);
    else if (k <= 8) AFF_KNN(8);
    else if (k <= 16) AFF_KNN(16);
    else AFF_KNN(32);
#undef AFF_KNN
    return check_launch("knn2d");
}

extern "C" int affb200_nearest_l1(const float *query, const float *db, int64_t B, int64_t Nq, int64_t Nd, int64_t *idx,
                                  void *stream) {
    if (!idx_dims_ok("nearest_l1", B, Nq) || Nd <= 0 || Nd >= (int64_t(1) << 31)) {
        if (Nd <= 0 || Nd >= (int64_t(1) << 31)) set_error("nearest_l1: the database must hold 1..2^31-1 points");
        return AFFB200_E_ARG;
    }
    if (B * Nq == 0) return AFFB200_OK;
    if (!query || !db || !idx) { set_error("nearest_l1: null pointer"); return AFFB200_E_ARG; }
    dim3 grid(unsigned((Nq + kIT - 1) / kIT), unsigned(B));
    nearest_l1_kernel<<<grid, kIT, 0, (cudaStream_t)stream>>>(query, db, int(Nq), int(Nd), idx);
    return check_launch("nearest_l1");
}

extern "C" int affb200_nbhd_build(const float *pos, const int64_t *nearest, int64_t B, int64_t N, int64_t m,
                                  int64_t nnc, int64_t rel_pos_width, int64_t table_width, int clamp,
                                  int64_t *nbhd_idx, int64_t *mask, int64_t *pe_idx, void *stream) {
    if (!idx_dims_ok("nbhd_build", B, N)) return AFFB200_E_ARG;
    if (m <= 0 || nnc < 0 || N * m * nnc >= (int64_t(1) << 31)) { set_error("nbhd_build: bad m / nnc"); return AFFB200_E_ARG; }
    const int64_t total = B * N * m * nnc;
    if (total == 0) return AFFB200_OK;
    if (!pos || !nearest || !nbhd_idx || !pe_idx) { set_error("nbhd_build: null pointer"); return AFFB200_E_ARG; }
    nbhd_build_kernel<<<blocks_for(total), kIT, 0, (cudaStream_t)stream>>>(
        pos, nearest, total, int(N), int(m), int(nnc), float(rel_pos_width), float(table_width), clamp, nbhd_idx, mask,
        pe_idx);
    return check_launch("nbhd_build");
}

extern "C" size_t affb200_merge_select_work_bytes(int64_t B, int64_t N) {
    return al16(size_t(B * N) * 2 * 8) + al16(size_t(B * N) * 2 * 4);  // self-kNN (k=2) indices and distances
}

extern "C" int affb200_merge_select(const float *pos, const void *lp, int lp_dtype, int64_t B, int64_t N,
                                    int64_t stride, double alpha, int64_t sample, int64_t reserve_num, int reserve_on,
                                    float *final_prob, int64_t *idx_out, int32_t *status, void *work, void *stream) {
    if (!idx_dims_ok("merge_select", B, N)) return AFFB200_E_ARG;
    if (stride < 1 || sample < 0 || reserve_num < 0 || sample > N) { set_error("merge_select: bad stride / counts"); return AFFB200_E_ARG; }
    if (B == 0) return AFFB200_OK;
    if (N * 4 > 200 * 1024) { set_error("merge_select: N > 51200 tokens per image is not supported"); return AFFB200_E_ARG; }
    if (!pos || !final_prob || !idx_out || !status || (stride != 2 && !work)) { set_error("merge_select: null pointer"); return AFFB200_E_ARG; }
    auto st = (cudaStream_t)stream;
    float *d2nd = nullptr;
    if (stride != 2) {  // distance to the nearest other token (aff_transformer.py:298-299), laid out [B][N]
        if (N < 2) { set_error("merge_select: the adaptive grid prior needs N >= 2"); return AFFB200_E_ARG; }
        int64_t *ki = reinterpret_cast<int64_t *>(work);
        float *kd = reinterpret_cast<float *>(reinterpret_cast<char *>(work) + al16(size_t(B * N) * 16));
        if (int e = affb200_knn2d(pos, pos, B, N, N, 2, ki, kd, stream)) return e;
        // compact column 1 in place: d2nd[t] = kd[2t+1]  (strided read inside the kernel instead of a copy)
        d2nd = kd;
    }
    const size_t smem = size_t(N) * 4;
    return dispatch_dtype(lp ? lp_dtype : AFFB200_F32, [&](auto tag) {
        using T = decltype(tag);
        auto kern = merge_select_kernel<T>;
        if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        kern<<<unsigned(B), kSelT, smem, st>>>(pos, (const T *)lp, d2nd, int(N), int(stride), float(alpha), int(sample),
                                               int(reserve_num), reserve_on, final_prob, idx_out, status);
        return check_launch("merge_select");
    });
}

extern "C" int affb200_gather_rows_i64(const int64_t *src, const int64_t *idx, int64_t B, int64_t N, int64_t K,
                                       int64_t M, int64_t *dst, void *stream) {
    if (!idx_dims_ok("gather_rows_i64", B, N) || K < 0 || M < 0 || (M & 1)) {
        if (M & 1) set_error("gather_rows_i64: M must be even");
        return AFFB200_E_ARG;
    }
    const int64_t total = B * K * (M / 2);
    if (total == 0) return AFFB200_OK;
    if (!src || !idx || !dst) { set_error("gather_rows_i64: null pointer"); return AFFB200_E_ARG; }
    gather_rows_i64_kernel<<<blocks_for(total), kIT, 0, (cudaStream_t)stream>>>(src, idx, total, int(N), int(K),
                                                                                int(M / 2), dst);
    return check_launch("gather_rows_i64");
}
