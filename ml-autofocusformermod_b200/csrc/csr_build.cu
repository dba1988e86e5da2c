// csr_build.cu — transposed neighbour index (per batch element CSR: key row -> referencing (token, slot) entries).
// Built once per nbhd_idx tensor and reused by every backward call of the stage (nbhd_idx is constant across
// the 2..18 blocks of a stage: reference models/aff_transformer.py:489-495).
#include "common.cuh"

namespace affb200 {

constexpr int kCsrThreads = 256;
constexpr int kWarpSortMax = 1024;  // rows up to this length are rank-sorted by one warp in shared memory

// off[b][r+1] += 1 for every entry referencing r   (off pre-zeroed)
__global__ void __launch_bounds__(kCsrThreads)
csr_count_kernel(const int64_t *__restrict__ idx, int32_t *__restrict__ off, int64_t NqM, int Nk, int64_t total) {
    const int64_t g = int64_t(blockIdx.x) * kCsrThreads + threadIdx.x;
    if (g >= total) return;
    const int64_t b = g / NqM;
    const int64_t r = __ldg(idx + g);
    atomicAdd(off + b * (Nk + 1) + r + 1, 1);
}

// per batch element: inclusive scan of off[b][1..Nk] in place; cur[b][r] = off[b][r]
__global__ void __launch_bounds__(1024)
csr_scan_kernel(int32_t *__restrict__ off, int32_t *__restrict__ cur, int Nk) {
    __shared__ int warp_sum[32];
    __shared__ int carry_s;
    int32_t *o = off + int64_t(blockIdx.x) * (Nk + 1);
    int32_t *c = cur + int64_t(blockIdx.x) * Nk;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < Nk; base += 1024) {
        const int r = base + tid;
        const int v = r < Nk ? o[r + 1] : 0;
        int s = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, s, d);
            if (lane >= d) s += t;
        }
        if (lane == 31) warp_sum[wid] = s;
        __syncthreads();
        if (wid == 0) {
            int ws = warp_sum[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, ws, d);
                if (lane >= d) ws += t;
            }
            warp_sum[lane] = ws;
        }
        __syncthreads();
        const int carry = carry_s;
        const int incl = carry + s + (wid > 0 ? warp_sum[wid - 1] : 0);
        if (r < Nk) {
            o[r + 1] = incl;
            c[r] = incl - v;
        }
        __syncthreads();
        if (tid == 1023) carry_s = incl;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kCsrThreads)
csr_fill_kernel(const int64_t *__restrict__ idx, int32_t *__restrict__ cur, int32_t *__restrict__ ent,
                int64_t NqM, int Nk, int64_t total) {
    const int64_t g = int64_t(blockIdx.x) * kCsrThreads + threadIdx.x;
    if (g >= total) return;
    const int64_t b = g / NqM;
    const int e = int(g - b * NqM);
    const int64_t r = __ldg(idx + g);
    const int p = atomicAdd(cur + b * Nk + r, 1);
    ent[b * NqM + p] = e;
}

// Make entry order canonical (ascending e) so the backward sums are run-to-run deterministic.  One warp per
// row: entries are distinct, so rank = number of smaller entries.  Rows longer than kWarpSortMax go to a
// worklist handled by csr_sort_long_kernel.
__global__ void __launch_bounds__(kCsrThreads)
csr_sort_rows_kernel(const int32_t *__restrict__ off, int32_t *__restrict__ ent, int64_t NqM, int Nk,
                     int64_t n_rows, int32_t *__restrict__ long_count, int64_t *__restrict__ long_rows) {
    __shared__ int buf[kCsrThreads / 32][kWarpSortMax];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row = int64_t(blockIdx.x) * (kCsrThreads / 32) + wid;
    if (row >= n_rows) return;
    const int64_t b = row / Nk;
    const int r = int(row - b * Nk);
    const int32_t *o = off + b * (Nk + 1) + r;
    const int beg = o[0], len = o[1] - o[0];
    if (len <= 1) return;
    if (len > kWarpSortMax) {
        if (lane == 0) long_rows[atomicAdd(long_count, 1)] = row;
        return;
    }
    int32_t *seg = ent + b * NqM + beg;
    int *s = buf[wid];
    for (int t = lane; t < len; t += 32) s[t] = seg[t];
    __syncwarp();
    for (int t = lane; t < len; t += 32) {
        const int v = s[t];
        int rank = 0;
        for (int u = 0; u < len; ++u) rank += (s[u] < v);
        seg[rank] = v;
    }
}

// bitonic sort of long rows in global memory, one CTA per worklist item (grid-stride over the worklist)
__global__ void __launch_bounds__(1024)
csr_sort_long_kernel(const int32_t *__restrict__ off, int32_t *__restrict__ ent, int64_t NqM, int Nk,
                     const int32_t *__restrict__ long_count, const int64_t *__restrict__ long_rows) {
    const int n_long = *long_count;
    for (int item = blockIdx.x; item < n_long; item += gridDim.x) {
        const int64_t row = long_rows[item];
        const int64_t b = row / Nk;
        const int r = int(row - b * Nk);
        const int32_t *o = off + b * (Nk + 1) + r;
        const int len = o[1] - o[0];
        int32_t *seg = ent + b * NqM + o[0];
        int n2 = 1;
        while (n2 < len) n2 <<= 1;
        // all-ascending bitonic network (mirror step, then halving steps): with virtual +inf padding beyond
        // len no comparator ever needs to move a padding element, so pairs touching u >= len are skipped.
        for (int k = 2; k <= n2; k <<= 1) {
            for (int j = k; j > 1; j >>= 1) {
                const int xmask = (j == k) ? (k - 1) : ((j >> 1));
                for (int t = threadIdx.x; t < n2; t += blockDim.x) {
                    const int u = t ^ xmask;
                    if (u > t && u < len) {
                        const int a = seg[t], c = seg[u];
                        if (a > c) { seg[t] = c; seg[u] = a; }
                    }
                }
                __syncthreads();
            }
        }
        __syncthreads();
    }
}

}  // namespace affb200

using namespace affb200;

extern "C" size_t affb200_csr_off_bytes(int64_t B, int64_t Nk) { return size_t(B) * size_t(Nk + 1) * 4; }
extern "C" size_t affb200_csr_ent_bytes(int64_t B, int64_t Nq, int64_t M) { return size_t(B) * size_t(Nq) * size_t(M) * 4; }
// cursors [B][Nk] int32 + long-row counter (16 B) + long-row worklist [B*Nk] int64
extern "C" size_t affb200_csr_work_bytes(int64_t B, int64_t Nk) {
    size_t cur = (size_t(B) * size_t(Nk) * 4 + 15) & ~size_t(15);
    return cur + 16 + size_t(B) * size_t(Nk) * 8;
}

extern "C" int affb200_csr_build(const int64_t *nbhd_idx, int64_t B, int64_t Nq, int64_t Nk, int64_t M,
                                 int32_t *off, int32_t *ent, void *work, int sorted, void *stream) {
    if (B < 0 || Nq < 0 || Nk < 0 || M < 0 || Nq * M >= (int64_t(1) << 31) || Nk >= (int64_t(1) << 31) - 1) {
        set_error("csr_build: bad dimensions");
        return AFFB200_E_ARG;
    }
    if (B == 0) return AFFB200_OK;
    if (!off || (!ent && Nq * M > 0) || (!work && Nk > 0) || (!nbhd_idx && Nq * M > 0)) {
        set_error("csr_build: null pointer");
        return AFFB200_E_ARG;
    }
    auto st = (cudaStream_t)stream;
    const int64_t NqM = Nq * M, total = B * NqM;
    if (cudaMemsetAsync(off, 0, affb200_csr_off_bytes(B, Nk), st) != cudaSuccess) return check_launch("csr memset");
    if (total == 0 || Nk == 0) return AFFB200_OK;
    int32_t *cur = reinterpret_cast<int32_t *>(work);
    const size_t cur_bytes = (size_t(B) * size_t(Nk) * 4 + 15) & ~size_t(15);
    int32_t *long_count = reinterpret_cast<int32_t *>(reinterpret_cast<char *>(work) + cur_bytes);
    int64_t *long_rows = reinterpret_cast<int64_t *>(reinterpret_cast<char *>(work) + cur_bytes + 16);
    const unsigned gblocks = unsigned(ceil_div64(total, kCsrThreads));
    csr_count_kernel<<<gblocks, kCsrThreads, 0, st>>>(nbhd_idx, off, NqM, int(Nk), total);
    csr_scan_kernel<<<unsigned(B), 1024, 0, st>>>(off, cur, int(Nk));
    csr_fill_kernel<<<gblocks, kCsrThreads, 0, st>>>(nbhd_idx, cur, ent, NqM, int(Nk), total);
    if (sorted) {
        if (cudaMemsetAsync(long_count, 0, 16, st) != cudaSuccess) return check_launch("csr memset");
        const int64_t n_rows = B * Nk;
        csr_sort_rows_kernel<<<unsigned(ceil_div64(n_rows, kCsrThreads / 32)), kCsrThreads, 0, st>>>(
            off, ent, NqM, int(Nk), n_rows, long_count, long_rows);
        csr_sort_long_kernel<<<kNumSMs, 1024, 0, st>>>(off, ent, NqM, int(Nk), long_count, long_rows);
    }
    return check_launch("csr_build", sorted ? 5 : 3);
}
