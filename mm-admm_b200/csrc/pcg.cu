// General sparse path of the ADMM x-update: preconditioned conjugate gradients on a fixed CSR matrix, FP64, sm_100a.
//
// The reference solves  t x = vec  with Eigen::ConjugateGradient + Jacobi preconditioner (src/MeshIntegrator.cpp:45-55,
// 137-138,159-160); its one genuinely non-diagonal solve (Method 2) goes through lib/LASolver: level-of-fill ILU
// (ILU_class.cpp:98-444), triangular solves (:447-527), CG / CGSTAB accelerators (accel_class.cpp:280-491), matmult (:521-549),
// dot (:496-519).  On the ADMM matrix itself t is numerically diagonal and the product path uses the fused
// scatter/divide kernel k_xupdate; the kernels here are the general form the north star asks for, verified against LASolver
// built from the reference's own sources (oracle/_ref) and against scipy:
//
//   k_spmv_dot<LANES>   y = A x with LANES threads per row (warp-shuffle row reduction) fused with the partial sums of x.y
//   k_tri_level<LOWER>  one level of the level-scheduled ILU(0) triangular solves (rows of a level are independent)
//   k_tri_cluster<LOWER> ALL levels of one triangular solve in ONE launch: a thread-block cluster of 8 x 256 threads walks the
//                       levels with a hardware cluster barrier between them instead of a kernel boundary (natural-order ILU(0)
//                       on a structured 2D mesh has thousands of levels of a few hundred rows: launch latency was the cost)
//   k_cg_update         x += alpha p, r -= alpha q, z = r / diag (Jacobi) fused with the partial sums of r.z and r.r
//   k_dot2              partial sums of a.b and c.c                      (ILU path: after the triangular solves)
//   k_scalars           fixed-order final reduction + the CG scalars (alpha, beta) kept on the device
//   k_xpay              p = z + beta p
//
// All reductions are two-stage with a fixed summation order (deterministic, like the rest of the library).  The matrix does
// not change between steps, so ILU(0) is factored once at creation (host, row by row in CSR order -- IKJ variant on the
// matrix's own sparsity pattern) together with the level schedule of both triangular factors.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>

#include "../../include/mmadmm.h"

namespace {

constexpr int TPB = 256;
inline int cdiv(size_t a, size_t b) { return (int)((a + b - 1) / b); }
// MMADMM_DEBUG_SETUP=1: host-side phase times of the set-up paths on stderr
struct SetupTimer {
    bool on; timespec t0; const char *what;
    explicit SetupTimer(const char *w) : on(getenv("MMADMM_DEBUG_SETUP") != nullptr), what(w) { if (on) clock_gettime(CLOCK_MONOTONIC, &t0); }
    void lap(const char *phase) {
        if (!on) return;
        timespec t1; clock_gettime(CLOCK_MONOTONIC, &t1);
        fprintf(stderr, "[mmadmm setup] %s: %s %.1f ms\n", what, phase, (t1.tv_sec - t0.tv_sec) * 1e3 + (t1.tv_nsec - t0.tv_nsec) * 1e-6);
        t0 = t1;
    }
};

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// block sum valid in thread 0 (blockDim.x = TPB)
__device__ __forceinline__ double block_sum_d(double v, double *sh) {
    v = warp_sum_d(v);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[wid] = v;
    __syncthreads();
    double r = 0.0;
    if (wid == 0) {
        r = (lane < (int)(blockDim.x >> 5)) ? sh[lane] : 0.0;
        r = warp_sum_d(r);
    }
    return r;
}

// y = A x, LANES threads per row; part[blockIdx] = sum over the block's rows of x_i * y_i (used as p.Ap) when part != null
template <int LANES>
__global__ void __launch_bounds__(TPB) k_spmv_dot(int n, const int *__restrict__ rowptr, const int *__restrict__ col,
                                                   const double *__restrict__ val, const double *__restrict__ x,
                                                   double *__restrict__ y, double *part) {
    __shared__ double sh[32];
    const int row = (blockIdx.x * TPB + threadIdx.x) / LANES;
    const int sub = threadIdx.x % LANES;
    double acc = 0.0;
    if (row < n) {
        const int k1 = rowptr[row + 1];
        for (int k = rowptr[row] + sub; k < k1; k += LANES) acc += val[k] * __ldg(x + col[k]);
    }
#pragma unroll
    for (int o = LANES / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    double d = 0.0;
    if (row < n && sub == 0) {
        y[row] = acc;
        d = acc * x[row];
    }
    if (part) {
        const double s = block_sum_d(d, sh);
        if (threadIdx.x == 0) part[blockIdx.x] = s;
    }
}

// SELL-C-sigma form of the same product (C = 32: one warp per slice of 32 rows, sigma = 4096: rows are sorted
// by length inside windows of 4096 rows so that a slice pads little).  A slice stores its entries column-major
// ([entry j of the row][32 rows]): every load of values / column indices is one coalesced 256 / 128-byte line, one thread owns a
// row and accumulates in ascending column order (the order of LASolver's matmult, accel_class.cpp:521-549).  Padding: value 0,
// column 0.
__global__ void __launch_bounds__(TPB) k_spmv_sell(int nSlices, const int *__restrict__ slicePtr, const int *__restrict__ scol,
                                                    const double *__restrict__ sval, const int *__restrict__ perm,
                                                    const double *__restrict__ x, double *__restrict__ y, double *part) {
    __shared__ double sh[32];
    const int gt = blockIdx.x * TPB + threadIdx.x;
    const int slice = gt >> 5, lane = gt & 31;
    double d = 0.0;
    if (slice < nSlices) {
        const int row = perm[slice * 32 + lane];
        const int p0 = slicePtr[slice], w = (slicePtr[slice + 1] - p0) >> 5;
        double acc = 0.0;
        for (int j = 0; j < w; j++) {
            const int k = p0 + j * 32 + lane;
            acc += sval[k] * __ldg(x + scol[k]);
        }
        if (row >= 0) {
            y[row] = acc;
            d = acc * x[row];
        }
    }
    if (part) {
        const double s = block_sum_d(d, sh);
        if (threadIdx.x == 0) part[blockIdx.x] = s;
    }
}

// One level of a triangular solve with the ILU(0) factors stored in A's pattern (unit lower, upper with diagonal):
// rows[lo..hi) are mutually independent.  LOWER: y_i = b_i - sum_{j<i} L_ij y_j ; UPPER: x_i = (y_i - sum_{j>i} U_ij x_j) / U_ii.
template <bool LOWER>
__global__ void __launch_bounds__(TPB) k_tri_level(int lo, int hi, const int *__restrict__ rows, const int *__restrict__ rowptr,
                                                    const int *__restrict__ col, const int *__restrict__ diagPos,
                                                    const double *__restrict__ lu, const double *__restrict__ rhs, double *out) {
    const int t = lo + blockIdx.x * TPB + threadIdx.x;
    if (t >= hi) return;
    const int i = rows[t];
    const int dp = diagPos[i];
    double s = rhs[i];
    if (LOWER) {
        for (int k = rowptr[i]; k < dp; k++) s -= lu[k] * out[col[k]];
        out[i] = s;
    } else {
        for (int k = dp + 1; k < rowptr[i + 1]; k++) s -= lu[k] * out[col[k]];
        out[i] = s / lu[dp];
    }
}

// All levels [l0, l1) of one triangular solve in one launch by ONE thread-block cluster (TC_CLUSTER CTAs x TC_THREADS threads).
// Rows of a level are spread over the cluster's threads; levels are separated by the hardware cluster barrier, split into
// its two halves: after a thread has stored its row's result it ARRIVES (release), then fetches the NEXT level's row --
// extent (kbe: row extents stored in level order, so they are indexed by t, not through rows[t]), right-hand side and up to
// TC_CACHE (column, factor) pairs into registers -- and only then WAITS (acquire).  None of that data is written by this
// kernel, so issuing those loads inside the barrier is safe, and their latency overlaps the barrier's.  After the wait a row
// costs one round trip to L2 for its `out` operands (ld.global.cg: values written by other CTAs, never a stale L1 line),
// the FMA chain in ascending k (same order, same bits as k_tri_level) and the store.
constexpr int TC_THREADS = 256, TC_CLUSTER = 8, TC_CACHE = 12;
__device__ __forceinline__ void cluster_arrive_release() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait_acquire() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
template <bool LOWER>
__global__ void __cluster_dims__(TC_CLUSTER, 1, 1) __launch_bounds__(TC_THREADS)
k_tri_cluster(int l0, int l1, const int *__restrict__ lvlOff, const int *__restrict__ rows, const int2 *__restrict__ kbe,
              const int *__restrict__ col, const double *__restrict__ lu, const double *__restrict__ rhs, double *out) {
    const int tid = blockIdx.x * TC_THREADS + threadIdx.x;
    constexpr int NTHR = TC_THREADS * TC_CLUSTER;
    int lo = lvlOff[l0], hi = lvlOff[l0 + 1];
    int i = 0, kb = 0, ke = 0, c[TC_CACHE];
    double s = 0.0, dv = 1.0, a[TC_CACHE];
    // registers <- this thread's first row of the level starting at lo
    auto fetch = [&](int t) {
        const int2 e = kbe[t];
        i = rows[t];
        kb = e.x; ke = e.y;
        s = rhs[i];
        if (!LOWER) dv = lu[kb - 1];   // U rows start right after the diagonal
#pragma unroll
        for (int j = 0; j < TC_CACHE; j++)
            if (kb + j < ke) { c[j] = col[kb + j]; a[j] = lu[kb + j]; }
    };
    if (lo + tid < hi) fetch(lo + tid);
    for (int l = l0; l < l1; l++) {
        if (lo + tid < hi) {
            double o[TC_CACHE];
#pragma unroll
            for (int j = 0; j < TC_CACHE; j++) o[j] = (kb + j < ke) ? __ldcg(out + c[j]) : 0.0;
#pragma unroll
            for (int j = 0; j < TC_CACHE; j++)
                if (kb + j < ke) s -= a[j] * o[j];
            for (int k = kb + TC_CACHE; k < ke; k++) s -= lu[k] * __ldcg(out + col[k]);
            out[i] = LOWER ? s : s / dv;
        }
        for (int t = lo + tid + NTHR; t < hi; t += NTHR) {   // levels wider than the cluster: further passes, no register cache
            const int2 e = kbe[t];
            const int r = rows[t];
            double sr = rhs[r];
            for (int k = e.x; k < e.y; k++) sr -= lu[k] * __ldcg(out + col[k]);
            out[r] = LOWER ? sr : sr / lu[e.x - 1];
        }
        cluster_arrive_release();
        if (l + 1 < l1) {
            lo = hi; hi = lvlOff[l + 2];
            if (lo + tid < hi) fetch(lo + tid);
        }
        cluster_wait_acquire();
    }
}

// ILU(0) numeric factor on A's own pattern, one L-level per launch (rows of a level need only rows of earlier levels): IKJ order
// exactly as LASolver's level-0 factor (ILU_class.cpp:300-444) and the numpy restatement run it on the host -- for k < i in row i
// ascending: l_ik = a_ik / u_kk, then a_ij -= l_ik u_kj for the j > k of pattern(i), found by merging the two sorted rows.  The
// multiply and the subtract are rounded separately (__dmul_rn / __dsub_rn: host compilers do not contract them), so the factor
// is bit-identical to the host one.  flag: 1 = zero pivot.
__global__ void __launch_bounds__(TPB) k_ilu_level(int lo, int hi, const int *__restrict__ rows, const int *__restrict__ rowptr,
                                                    const int *__restrict__ col, const int *__restrict__ diagPos, double *lu, int *flag) {
    const int t = lo + blockIdx.x * TPB + threadIdx.x;
    if (t >= hi) return;
    const int i = rows[t];
    const int dp = diagPos[i], re = rowptr[i + 1];
    for (int k = rowptr[i]; k < dp; k++) {
        const int c = col[k];
        const int dc = diagPos[c];
        const double piv = lu[dc];
        if (piv == 0.0) { *flag = 1; return; }
        const double lik = lu[k] / piv;
        lu[k] = lik;
        int p = k + 1;
        const int ce = rowptr[c + 1];
        for (int kk = dc + 1; kk < ce && p < re; kk++) {
            const int cj = col[kk];
            while (p < re && col[p] < cj) p++;
            if (p < re && col[p] == cj) { lu[p] = __dsub_rn(lu[p], __dmul_rn(lik, lu[kk])); p++; }
        }
    }
    if (lu[dp] == 0.0) *flag = 1;
}
// The same factorisation with a WARP per row (rows of at most 64 stored entries): the row lives in registers, two entries per
// lane; for every L entry k in turn the pivot row's U part is read coalesced and matched against the row by shuffles, so a row
// costs two dependent round trips per L entry instead of two per (L entry, U entry) pair.  Every stored entry still receives its
// updates in ascending k with separately rounded multiply and subtract: same bits as k_ilu_level and as the host factor.
__global__ void __launch_bounds__(TPB) k_ilu_level_warp(int lo, int hi, const int *__restrict__ rows, const int *__restrict__ rowptr,
                                                         const int *__restrict__ col, const int *__restrict__ diagPos, double *lu, int *flag) {
    const int lane = threadIdx.x & 31;
    const int t = lo + (blockIdx.x * TPB + threadIdx.x) / 32;
    if (t >= hi) return;
    const int i = rows[t], r0 = rowptr[i], len = rowptr[i + 1] - r0, dpl = diagPos[i] - r0;
    int c0 = lane < len ? col[r0 + lane] : 0x7fffffff, c1 = lane + 32 < len ? col[r0 + lane + 32] : 0x7fffffff;
    double v0 = lane < len ? lu[r0 + lane] : 0.0, v1 = lane + 32 < len ? lu[r0 + lane + 32] : 0.0;
    for (int k = 0; k < dpl; k++) {
        const int ck = __shfl_sync(0xffffffffu, k < 32 ? c0 : c1, k & 31);
        const double vk = __shfl_sync(0xffffffffu, k < 32 ? v0 : v1, k & 31);
        const int dc = diagPos[ck], ce = rowptr[ck + 1];
        const double piv = lu[dc];
        if (piv == 0.0) { if (lane == 0) *flag = 1; return; }
        const double lik = vk / piv;
        if (lane == (k & 31)) { if (k < 32) v0 = lik; else v1 = lik; }
        for (int base = dc + 1; base < ce; base += 32) {
            const int kk = base + lane;
            const int cu = kk < ce ? col[kk] : -1;
            const double vu = kk < ce ? lu[kk] : 0.0;
            const int cnt = min(32, ce - base);
            for (int j = 0; j < cnt; j++) {
                const int cj = __shfl_sync(0xffffffffu, cu, j);
                const double vj = __shfl_sync(0xffffffffu, vu, j);
                if (c0 == cj) v0 = __dsub_rn(v0, __dmul_rn(lik, vj));
                if (c1 == cj) v1 = __dsub_rn(v1, __dmul_rn(lik, vj));
            }
        }
    }
    if (lane < len) lu[r0 + lane] = v0;
    if (lane + 32 < len) lu[r0 + lane + 32] = v1;
    const double dv = __shfl_sync(0xffffffffu, dpl < 32 ? v0 : v1, dpl & 31);
    if (dv == 0.0 && lane == 0) *flag = 1;
}
__global__ void k_inv_diag(int n, const int *__restrict__ diagPos, const double *__restrict__ val, double *invDiag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) invDiag[i] = 1.0 / val[diagPos[i]];   // Eigen's DiagonalPreconditioner stores 1/diag
}

// ---- block rows (option "row_block" = BS) ----
// A GROUP is BS consecutive rows [base, base + BS) solved by ONE thread in natural order (ascending for L, descending for U): the
// dependencies inside a group are honoured by the sequence, so only dependencies on OTHER groups define levels.  On the Method 2
// Jacobian (D x D blocks per vertex pair) BS = D halves (2D) / thirds (3D) the number of levels, BS = 2D shortens the wavefront
// once more.  Every row is still the same FMA chain in ascending k; an operand of the thread's own group comes from a register
// instead of memory: bit-identical to the row-per-thread kernels (tests/test_zz_gpu_method2.py).
template <int BS> __device__ __forceinline__ double pick_own(const double (&own)[BS], int g) {
    double v = own[0];
#pragma unroll
    for (int q = 1; q < BS; q++) v = (g == q) ? own[q] : v;
    return v;
}
template <bool LOWER, int BS>
__device__ __forceinline__ void tri_group_plain(int n, int base, const int *__restrict__ rowptr, const int *__restrict__ diagPos,
                                                const int *__restrict__ col, const double *__restrict__ lu,
                                                const double *__restrict__ rhs, double *out) {
    double own[BS];
#pragma unroll
    for (int j = 0; j < BS; j++) own[j] = 0.0;
#pragma unroll
    for (int jj = 0; jj < BS; jj++) {
        const int j = LOWER ? jj : BS - 1 - jj;
        const int i = base + j;
        if (i >= n) continue;   // the last group may be short
        const int dp = diagPos[i];
        const int kb = LOWER ? rowptr[i] : dp + 1, ke = LOWER ? dp : rowptr[i + 1];
        double s = rhs[i];
        for (int k = kb; k < ke; k++) {
            const int c = col[k], g = c - base;
            const double v = (g >= 0 && g < BS) ? pick_own<BS>(own, g) : __ldcg(out + c);
            s -= lu[k] * v;
        }
        own[j] = LOWER ? s : s / lu[dp];
        out[i] = own[j];
    }
}
// one level of groups, full-width launch
template <bool LOWER, int BS>
__global__ void __launch_bounds__(TPB) k_tri_level_blk(int n, int lo, int hi, const int *__restrict__ groups, const int *__restrict__ rowptr,
                                                        const int *__restrict__ col, const int *__restrict__ diagPos,
                                                        const double *__restrict__ lu, const double *__restrict__ rhs, double *out) {
    const int t = lo + blockIdx.x * TPB + threadIdx.x;
    if (t >= hi) return;
    tri_group_plain<LOWER, BS>(n, groups[t], rowptr, diagPos, col, lu, rhs, out);
}
// k_tri_cluster on groups: the first CACHE (column, factor) pairs of every row of the thread's group are fetched inside the split
// cluster barrier; after the wait ALL their operands that live in other groups are requested at once (one L2 round trip per
// level, not one per row), then the BS chains run in sequence.
template <bool LOWER, int BS, int CACHE>
__global__ void __cluster_dims__(TC_CLUSTER, 1, 1) __launch_bounds__(TC_THREADS)
k_tri_cluster_blk(int n, int l0, int l1, const int *__restrict__ lvlOff, const int *__restrict__ groups, const int2 *__restrict__ kbe,
                  const int *__restrict__ rowptr, const int *__restrict__ diagPos, const int *__restrict__ col,
                  const double *__restrict__ lu, const double *__restrict__ rhs, double *out) {
    const int tid = blockIdx.x * TC_THREADS + threadIdx.x;
    constexpr int NTHR = TC_THREADS * TC_CLUSTER;
    int lo = lvlOff[l0], hi = lvlOff[l0 + 1];
    int base = 0, kb[BS], ke[BS], c[BS][CACHE];
    double s[BS], dv[BS], a[BS][CACHE];
#pragma unroll
    for (int j = 0; j < BS; j++) { kb[j] = ke[j] = 0; s[j] = 0.0; dv[j] = 1.0; }
    auto fetch = [&](int t) {
        base = groups[t];
#pragma unroll
        for (int j = 0; j < BS; j++) {
            const int2 e = kbe[(size_t)t * BS + j];   // rows past n (short last group): empty extent
            kb[j] = e.x; ke[j] = e.y;
            if (base + j < n) {
                s[j] = rhs[base + j];
                if (!LOWER) dv[j] = lu[e.x - 1];   // U rows start right after the diagonal
            }
#pragma unroll
            for (int q = 0; q < CACHE; q++)
                if (e.x + q < e.y) { c[j][q] = col[e.x + q]; a[j][q] = lu[e.x + q]; }
        }
    };
    if (lo + tid < hi) fetch(lo + tid);
    for (int l = l0; l < l1; l++) {
        if (lo + tid < hi) {
            double o[BS][CACHE], own[BS];
#pragma unroll
            for (int j = 0; j < BS; j++) {
                own[j] = 0.0;
#pragma unroll
                for (int q = 0; q < CACHE; q++) {
                    const int g = c[j][q] - base;
                    o[j][q] = (kb[j] + q < ke[j] && !(g >= 0 && g < BS)) ? __ldcg(out + c[j][q]) : 0.0;
                }
            }
#pragma unroll
            for (int jj = 0; jj < BS; jj++) {
                const int j = LOWER ? jj : BS - 1 - jj;
                double sj = s[j];
#pragma unroll
                for (int q = 0; q < CACHE; q++)
                    if (kb[j] + q < ke[j]) {
                        const int g = c[j][q] - base;
                        const double v = (g >= 0 && g < BS) ? pick_own<BS>(own, g) : o[j][q];
                        sj -= a[j][q] * v;
                    }
                for (int k = kb[j] + CACHE; k < ke[j]; k++) {
                    const int cc = col[k], g = cc - base;
                    const double v = (g >= 0 && g < BS) ? pick_own<BS>(own, g) : __ldcg(out + cc);
                    sj -= lu[k] * v;
                }
                own[j] = LOWER ? sj : sj / dv[j];
                if (base + j < n) out[base + j] = own[j];
            }
        }
        for (int t = lo + tid + NTHR; t < hi; t += NTHR)   // levels wider than the cluster: further passes, no register cache
            tri_group_plain<LOWER, BS>(n, groups[t], rowptr, diagPos, col, lu, rhs, out);
        cluster_arrive_release();
        if (l + 1 < l1) {
            lo = hi; hi = lvlOff[l + 2];
            if (lo + tid < hi) fetch(lo + tid);
        }
        cluster_wait_acquire();
    }
}

// ---- level-ELL copy of the factor ("tri_layout" = 1) ----
// The cluster kernels above read every row where CSR put it: along a wavefront the rows of one level lie ~1 MB apart (one DRAM
// page and one TLB entry per row, 4- and 8-byte loads that share no sector between the threads of a warp) and the extent of a row
// must arrive before its entries can be requested.  Here every level that fits one pass of the cluster (<= 2048 groups) owns a
// CONTIGUOUS block: plane (j, q) = entry q of row j of the groups, in the level's group order, so thread t reads element t of each
// plane -- coalesced, one page, no dependent load.  U blocks carry the diagonal as plane 0.  meta[l] = {first group, last group + 1,
// block offset or -1 (level not in this layout), planes per row}.  Pads: column -1.  The arithmetic is untouched (same chains,
// ascending entries): bit-identical to the CSR kernels.
__global__ void k_lell_fill(size_t slots, const int *__restrict__ src, const double *__restrict__ lu, double *val) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < slots) { const int k = src[i]; val[i] = k >= 0 ? lu[k] : 1.0; }
}
template <bool LOWER, int BS, int CACHE>
__global__ void __cluster_dims__(TC_CLUSTER, 1, 1) __launch_bounds__(TC_THREADS)
k_tri_cluster_lell(int n, int l0, int l1, const int4 *__restrict__ meta, const int *__restrict__ groups, const int *__restrict__ colT,
                   const double *__restrict__ valT, const double *__restrict__ rhs, double *out
#ifdef MMADMM_TRI_PROF
                   , int pmode   // timing experiments only (results are wrong for pmode != 0): 1 no `out` gathers, 2 no factor loads, 3 barrier only, 4 no fences
#endif
) {
#ifndef MMADMM_TRI_PROF
    constexpr int pmode = 0;
#endif
    const int tid = blockIdx.x * TC_THREADS + threadIdx.x;
    constexpr int P0 = LOWER ? 0 : 1;
    int4 cur = meta[l0];
    int4 nxt = (l0 + 1 < l1) ? meta[l0 + 1] : cur;
    int base = 0, baseN = 0, c[BS][CACHE];
    double s[BS], dv[BS], a[BS][CACHE];
#pragma unroll
    for (int j = 0; j < BS; j++) { s[j] = 0.0; dv[j] = 1.0; }
    // registers <- this thread's group of level mt: right-hand sides, diagonals, the first CACHE planes of every row
    auto fetch = [&](const int4 &mt, int b) {
        const int w = mt.y - mt.x, MP = mt.w + P0;
        const size_t o0 = (size_t)mt.z + tid;
#pragma unroll
        for (int j = 0; j < BS; j++) {
            s[j] = (b + j < n) ? rhs[b + j] : 0.0;
            if (!LOWER) dv[j] = valT[o0 + (size_t)(j * MP) * w];
#pragma unroll
            for (int q = 0; q < CACHE; q++) {
                c[j][q] = -1;
                if (q < mt.w && pmode < 2) {
                    const size_t o = o0 + (size_t)(j * MP + P0 + q) * w;
                    c[j][q] = colT[o];
                    a[j][q] = valT[o];
                }
            }
        }
    };
    if (tid < cur.y - cur.x) { base = groups[cur.x + tid]; fetch(cur, base); }
    if (l0 + 1 < l1 && tid < nxt.y - nxt.x) baseN = groups[nxt.x + tid];
    for (int l = l0; l < l1; l++) {
        const int w = cur.y - cur.x, MP = cur.w + P0;
        if (tid < w) {
            double o[BS][CACHE], own[BS];
#pragma unroll
            for (int j = 0; j < BS; j++) {
                own[j] = 0.0;
#pragma unroll
                for (int q = 0; q < CACHE; q++) {
                    const int cc = c[j][q], g = cc - base;
                    o[j][q] = (cc >= 0 && !(g >= 0 && g < BS) && pmode != 1) ? __ldcg(out + cc) : 0.0;
                }
            }
#pragma unroll
            for (int jj = 0; jj < BS; jj++) {
                const int j = LOWER ? jj : BS - 1 - jj;
                double sj = s[j];
#pragma unroll
                for (int q = 0; q < CACHE; q++) {
                    const int cc = c[j][q], g = cc - base;
                    if (cc >= 0) {
                        const double v = (g >= 0 && g < BS) ? pick_own<BS>(own, g) : o[j][q];
                        sj -= a[j][q] * v;
                    }
                }
                for (int q = CACHE; q < cur.w; q++) {   // rows longer than the register cache: the remaining planes, still coalesced
                    const size_t off = (size_t)cur.z + tid + (size_t)(j * MP + P0 + q) * w;
                    const int cc = colT[off], g = cc - base;
                    if (cc >= 0) {
                        const double v = (g >= 0 && g < BS) ? pick_own<BS>(own, g) : __ldcg(out + cc);
                        sj -= valT[off] * v;
                    }
                }
                own[j] = LOWER ? sj : sj / dv[j];
                if (base + j < n && pmode != 3) out[base + j] = own[j];
            }
        }
        if (pmode == 4) asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); else cluster_arrive_release();
        if (l + 1 < l1) {
            cur = nxt; base = baseN;
            if (tid < cur.y - cur.x && pmode != 3) fetch(cur, base);
            if (l + 2 < l1) {
                nxt = meta[l + 2];
                if (tid < nxt.y - nxt.x) baseN = groups[nxt.x + tid];
            }
        }
        if (pmode == 4) asm volatile("barrier.cluster.wait.aligned;" ::: "memory"); else cluster_wait_acquire();
    }
}

// ---- dataflow ("sync-free") triangular solve: "tri_mode" = 2 ----
// No barrier between levels at all.  The result vector is pre-filled with a sentinel bit pattern; a row's 8-byte result IS its
// ready flag: a consumer polls its operands with gpu-scope relaxed loads until none is the sentinel, then runs the same FMA chain
// in ascending k and publishes its own result with one relaxed store (no fence: there is no separate flag to order against).
// Work is cut into CHUNKS of 32 groups in level order, every level padded to a whole chunk, so the 32 lanes of a warp never
// depend on each other (a lane spinning on a lane of its own warp would wait forever at the reconvergence point).  Warps of a
// persistent grid CLAIM chunks in order from an atomic counter: a chunk only depends on chunks claimed earlier by warps that
// are running, so the wavefront cannot deadlock whatever the residency of the grid.  Factor entries come from a chunk-ELL copy
// (plane (j, q) of a chunk = 32 contiguous entries: coalesced, loaded before the operands are polled), U chunks carry the
// diagonal as plane 0.  A level costs one store-to-visible plus one L2 round trip instead of a cluster barrier with its
// release/acquire fences (measured: 1.8-2.5 us of a 3.4-5.3 us level), and all SMs stream the factor.
// Polling is bounded (trap after ~1 s): a protocol bug is a CUDA error, never a hung GPU.
constexpr unsigned long long TRI_SENTINEL = 0xFFF8DEADBEEF0001ull;   // a NaN payload no computation produces
__device__ __forceinline__ double ld_relaxed_gpu(const double *p) {
    double v;
    asm volatile("ld.relaxed.gpu.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_gpu(double *p, double v) { asm volatile("st.relaxed.gpu.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory"); }
__global__ void k_tri_sf_prepare(int n, double *out, unsigned *counter) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __longlong_as_double((long long)TRI_SENTINEL);
    if (i == 0) *counter = 0u;
}
// one operand, polled on its own (rows longer than the register cache)
__device__ __forceinline__ double tri_sf_wait(const double *p) {
    for (int spins = 0;; spins++) {
        const double v = ld_relaxed_gpu(p);
        if ((unsigned long long)__double_as_longlong(v) != TRI_SENTINEL) return v;
        if (spins > 4) __nanosleep(64);
        if (spins > (1 << 24)) __trap();
    }
}
#ifdef MMADMM_TRI_PROF
// timing experiments: per chunk {claimed, factor planes in registers, proxy operand seen, all operands seen, results stored} (ns)
__device__ unsigned long long g_sf_prof[5 * 65536];
__device__ int g_sf_spins[65536];   // lane 0: polls issued in the per-lane wait of the chunk
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define SF_STAMP(k) do { if (lane == 0 && chunk >= profLo && chunk < profLo + 65536) g_sf_prof[5 * (chunk - profLo) + (k)] = gtimer(); } while (0)
#else
#define SF_STAMP(k) do { } while (0)
#endif
template <bool LOWER, int BS, int CACHE>
__global__ void __launch_bounds__(TC_THREADS) k_tri_syncfree(int n, int nChunks, const longlong2 *__restrict__ meta, const int *__restrict__ items,
                                                            const int *__restrict__ colT, const double *__restrict__ valT,
                                                            const double *__restrict__ rhs, double *out, unsigned *counter, int proxy
#ifdef MMADMM_TRI_PROF
                                                            , int profLo
#endif
) {
    static_assert(BS * CACHE <= 32, "operand mask is one word");
    const int lane = threadIdx.x & 31;
    constexpr int P0 = LOWER ? 0 : 1;
    int chunk = 0;
    if (lane == 0) chunk = (int)atomicAdd(counter, 1u);
    chunk = __shfl_sync(0xffffffffu, chunk, 0);
    while (chunk < nChunks) {
        int next = 0;
        if (lane == 0) next = (int)atomicAdd(counter, 1u);   // claimed now, needed after this chunk: the round trip is hidden
        SF_STAMP(0);
        const longlong2 mt = meta[chunk];
        const int m = (int)mt.y, MP = m + P0;
        const int base = items[(size_t)chunk * 32 + lane];   // -1: padding lane (its planes hold column -1)
        const size_t o0 = (size_t)mt.x + lane;
        int c[BS][CACHE];
        double a[BS][CACHE], s[BS], dv[BS], o[BS][CACHE], own[BS];
        unsigned need = 0u;
#pragma unroll
        for (int j = 0; j < BS; j++) {
            own[j] = 0.0;
            s[j] = (base >= 0 && base + j < n) ? rhs[base + j] : 0.0;
            dv[j] = LOWER ? 1.0 : valT[o0 + (size_t)(j * MP) * 32];
#pragma unroll
            for (int q = 0; q < CACHE; q++) {
                c[j][q] = -1;
                if (q < m) {
                    const size_t off = o0 + (size_t)(j * MP + P0 + q) * 32;
                    c[j][q] = colT[off];
                    a[j][q] = valT[off];
                }
                const int g = c[j][q] - base;
                o[j][q] = 0.0;
                if (c[j][q] >= 0 && !(g >= 0 && g < BS)) need |= 1u << (j * CACHE + q);
            }
        }
        // Operands of other groups: per round trip every lane requests ALL it still misses and only then looks at the results
        // (requests tested one by one, or lanes of a warp waiting in different branches, serialise the round trips); what
        // arrived stays in its register, the rest is requested again.  "tri_proxy" 1 puts a cheaper first stage in front for
        // warps that run far ahead of the wavefront: ONE lane polls ONE operand (the nearest row: highest column of an L row,
        // lowest of a U row) until it has arrived.
        SF_STAMP(1);
        if (proxy) {
            const unsigned waiting = __ballot_sync(0xffffffffu, need != 0u);
            if (waiting) {
                if (lane == __ffs(waiting) - 1) {
                    const int b = LOWER ? 31 - __clz(need) : __ffs(need) - 1;
                    const double *p = out;
#pragma unroll
                    for (int j = 0; j < BS; j++)
#pragma unroll
                        for (int q = 0; q < CACHE; q++)
                            if (b == j * CACHE + q) p = out + c[j][q];
                    for (int spins = 0;; spins++) {
                        if ((unsigned long long)__double_as_longlong(ld_relaxed_gpu(p)) != TRI_SENTINEL) break;
                        if (spins > (1 << 24)) __trap();
                    }
                }
                __syncwarp();
            }
        }
        SF_STAMP(2);
        for (int spins = 0; need; spins++) {
#pragma unroll
            for (int j = 0; j < BS; j++)
#pragma unroll
                for (int q = 0; q < CACHE; q++)
                    if (need & (1u << (j * CACHE + q))) o[j][q] = ld_relaxed_gpu(out + c[j][q]);
#pragma unroll
            for (int j = 0; j < BS; j++)
#pragma unroll
                for (int q = 0; q < CACHE; q++)
                    if ((need & (1u << (j * CACHE + q))) && (unsigned long long)__double_as_longlong(o[j][q]) != TRI_SENTINEL) need &= ~(1u << (j * CACHE + q));
            if (spins > (1 << 24)) __trap();
        }
        __syncwarp();
        SF_STAMP(3);
#pragma unroll
        for (int jj = 0; jj < BS; jj++) {
            const int j = LOWER ? jj : BS - 1 - jj;
            double sj = s[j];
#pragma unroll
            for (int q = 0; q < CACHE; q++) {
                const int cc = c[j][q], g = cc - base;
                if (cc >= 0) {
                    const double v = (g >= 0 && g < BS) ? pick_own<BS>(own, g) : o[j][q];
                    sj -= a[j][q] * v;
                }
            }
            for (int q = CACHE; q < m; q++) {
                const size_t off = o0 + (size_t)(j * MP + P0 + q) * 32;
                const int cc = colT[off], g = cc - base;
                if (cc >= 0) {
                    const double v = (g >= 0 && g < BS) ? pick_own<BS>(own, g) : tri_sf_wait(out + cc);
                    sj -= valT[off] * v;
                }
            }
            own[j] = LOWER ? sj : sj / dv[j];
            if (base >= 0 && base + j < n) st_relaxed_gpu(out + base + j, own[j]);
        }
        SF_STAMP(4);
        chunk = __shfl_sync(0xffffffffu, next, 0);
    }
}

// ---- dataflow solve on BS x BS blocks ----
// On a matrix made of dense BS x BS blocks (the Method 2 Jacobian: one block per pair of adjacent vertices) the BS rows of a group
// read the SAME operands: the generic kernel above requests each of them once per row, 8 bytes at a time, and what a hop of the
// wavefront costs is exactly that -- scattered strong loads, 32 sectors per warp instruction (measured: a level = 2.6 us with
// 14 operand loads per lane, 4.3 us with 28).  Here a lane holds its group's neighbour BLOCKS: one operand request per
// neighbour group (a 16-byte load for BS = 2), shared by the BS rows.  Operands that were computed long ago (the cell-centre
// rows of the upper factor) are picked up when the chunk is claimed, off the critical path.  Row arithmetic in CSR order:
// L row = neighbour blocks ascending, then the group's own columns; U row = own columns, then neighbour blocks: same bits.
template <int BS> __device__ __forceinline__ void ld_group_raw(const double *p, double (&v)[BS]) {
    if (BS == 2) {
        asm volatile("ld.relaxed.gpu.global.v2.f64 {%0, %1}, [%2];" : "=d"(v[0]), "=d"(v[1]) : "l"(p) : "memory");
    } else {
#pragma unroll
        for (int k = 0; k < BS; k++) v[k] = ld_relaxed_gpu(p + k);
    }
}
template <int BS> __device__ __forceinline__ bool group_ready(const double (&v)[BS]) {
    bool ok = true;
#pragma unroll
    for (int k = 0; k < BS; k++) ok = ok && ((unsigned long long)__double_as_longlong(v[k]) != TRI_SENTINEL);
    return ok;
}
template <bool LOWER, int BS, int NBC>
__global__ void __launch_bounds__(TC_THREADS, 1) k_tri_sf_block(int n, int nChunks, const longlong2 *__restrict__ meta, const int *__restrict__ items,
                                                            const int *__restrict__ nbrT, const double *__restrict__ valT,
                                                            const double *__restrict__ rhs, double *out, unsigned *counter, int proxy
#ifdef MMADMM_TRI_PROF
                                                            , int profLo, int pmode   // pmode (timing only, results wrong): 1 no proxy phase, 2 no FP chain, 4 no early batch
#endif
) {
#ifndef MMADMM_TRI_PROF
    constexpr int pmode = 0;
#endif
    const int lane = threadIdx.x & 31;
    constexpr int BB = BS * BS;
    int chunk = 0;
    if (lane == 0) chunk = (int)atomicAdd(counter, 1u);
    chunk = __shfl_sync(0xffffffffu, chunk, 0);
    while (chunk < nChunks) {
        int next = 0;
        if (lane == 0) next = (int)atomicAdd(counter, 1u);
        SF_STAMP(0);
        const longlong2 mt = meta[chunk];                    // x: first value slot, y: first index slot << 8 | neighbour planes
        const int m = (int)(mt.y & 255);
        const size_t oV = (size_t)mt.x + lane, oI = (size_t)(mt.y >> 8) + lane;
        const int base = items[(size_t)chunk * 32 + lane];   // first row of the lane's group; -1: padding lane
        double d[BB], a[NBC][BB], o[NBC][BS], s[BS], own[BS];
        int nb[NBC];
        unsigned need = 0u;
#pragma unroll
        for (int k = 0; k < BB; k++) d[k] = valT[oV + (size_t)k * 32];
#pragma unroll
        for (int r = 0; r < BS; r++) { s[r] = base >= 0 ? rhs[base + r] : 0.0; own[r] = 0.0; }
#pragma unroll
        for (int q = 0; q < NBC; q++) {
            nb[q] = q < m ? nbrT[oI + (size_t)q * 32] : -1;
            if (nb[q] >= 0) {
                need |= 1u << q;
#pragma unroll
                for (int k = 0; k < BB; k++) a[q][k] = valT[oV + (size_t)(BB * (1 + q) + k) * 32];
            }
        }
        // operands that have been there for a long time: one batch now, while nobody waits for this chunk
        // (all requests are issued before any result is looked at: one round trip for the batch, not one per neighbour)
        {   // (straight into the operand registers: a group that is not complete yet is simply requested again)
#pragma unroll
            for (int q = 0; q < NBC; q++)
                if ((need & (1u << q)) && !(pmode & 4)) ld_group_raw<BS>(out + nb[q], o[q]);
#pragma unroll
            for (int q = 0; q < NBC; q++)
                if ((need & (1u << q)) && !(pmode & 4) && group_ready<BS>(o[q])) need &= ~(1u << q);
        }
        // optional ("tri_proxy" 1): a far-ahead warp waits through ONE lane polling ONE neighbour (see k_tri_syncfree).  Off by
        // default here: with one 16-byte request per lane per round trip the polls of all waiting warps together stay far
        // below what L2 serves, and the extra detection step cost 0.9 us per level (12.7 vs 16.5 ms per application).
        SF_STAMP(1);
        {
            const unsigned waiting = (!proxy || (pmode & 1)) ? 0u : __ballot_sync(0xffffffffu, need != 0u);
            if (waiting) {
                if (lane == __ffs(waiting) - 1) {
                    const int b = LOWER ? 31 - __clz(need) : __ffs(need) - 1;
                    const double *p = out;
#pragma unroll
                    for (int q = 0; q < NBC; q++)
                        if (b == q) p = out + nb[q];
                    for (int spins = 0;; spins++) {
                        if ((unsigned long long)__double_as_longlong(ld_relaxed_gpu(p)) != TRI_SENTINEL) break;
                        if (spins > (1 << 24)) __trap();
                    }
                }
                __syncwarp();
            }
        }
        // then every lane waits for its own neighbours: per round trip ALL still-missing ones are requested before any result is
        // looked at (a request followed by its own test, neighbour after neighbour, serialises the round trips -- and so do
        // lanes that wait for different neighbour slots: measured 0.51 us per poll instead of the 0.19 us of one relaxed load)
        SF_STAMP(2);
        for (int spins = 0; need; spins++) {
#pragma unroll
            for (int q = 0; q < NBC; q++)
                if (need & (1u << q)) ld_group_raw<BS>(out + nb[q], o[q]);
#pragma unroll
            for (int q = 0; q < NBC; q++)
                if ((need & (1u << q)) && group_ready<BS>(o[q])) need &= ~(1u << q);
            if (spins > (1 << 24)) __trap();
#ifdef MMADMM_TRI_PROF
            if (lane == 0 && chunk >= profLo && chunk < profLo + 65536) g_sf_spins[chunk - profLo] = spins + 1;
#endif
        }
        __syncwarp();
        SF_STAMP(3);
#pragma unroll
        for (int rr = 0; rr < BS; rr++) {
            const int r = LOWER ? rr : BS - 1 - rr;
            double sj = s[r];
            if (!LOWER) {
#pragma unroll
                for (int cidx = 0; cidx < BS; cidx++)
                    if (cidx > r) sj -= d[r * BS + cidx] * own[cidx];
            }
#pragma unroll
            for (int q = 0; q < NBC; q++)
                if (nb[q] >= 0 && !(pmode & 2)) {
#pragma unroll
                    for (int cidx = 0; cidx < BS; cidx++) sj -= a[q][r * BS + cidx] * o[q][cidx];
                }
            for (int q = NBC; q < m; q++) {   // more neighbours than the register cache holds
                const int nq = nbrT[oI + (size_t)q * 32];
                if (nq >= 0) {
#pragma unroll
                    for (int cidx = 0; cidx < BS; cidx++) sj -= valT[oV + (size_t)(BB * (1 + q) + r * BS + cidx) * 32] * tri_sf_wait(out + nq + cidx);
                }
            }
            if (LOWER) {
#pragma unroll
                for (int cidx = 0; cidx < BS; cidx++)
                    if (cidx < r) sj -= d[r * BS + cidx] * own[cidx];
            }
            own[r] = (LOWER || (pmode & 2)) ? sj : sj / d[r * BS + r];
        }
        if (base >= 0) {
            if (BS == 2) asm volatile("st.relaxed.gpu.global.v2.f64 [%0], {%1, %2};" ::"l"(out + base), "d"(own[0]), "d"(own[1]) : "memory");
            else {
#pragma unroll
                for (int r = 0; r < BS; r++) st_relaxed_gpu(out + base + r, own[r]);
            }
        }
        SF_STAMP(4);
        chunk = __shfl_sync(0xffffffffu, next, 0);
    }
}

// Index planes of the block form of the chunk-major factor copy, built on the device (the host only lays out chunks): for the
// group of (chunk, lane) the neighbour planes (first column of every neighbour block on this side) and, per value slot, the
// position of its entry in the CSR factor (what k_lell_fill gathers from).  Pads keep the -1 of the memset.
__global__ void k_chell_block_fill(int nChunks, int bs, int lower, const longlong2 *__restrict__ meta, const int *__restrict__ items,
                                   const int *__restrict__ rp, const int *__restrict__ cl, int *nbr, int *src) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nChunks * 32) return;
    const int c = t >> 5, r = t & 31, b = items[t];
    if (b < 0) return;
    const longlong2 mt = meta[c];
    const size_t oV = (size_t)mt.x + r, oI = (size_t)(mt.y >> 8) + r;
    const int BB = bs * bs, r0 = rp[b], nblk = (rp[b + 1] - r0) / bs;
    int p = 0;
    while (cl[r0 + p * bs] != b) p++;
    const int cnt = lower ? p : nblk - p - 1, first = lower ? 0 : p + 1;
    for (int rr = 0; rr < bs; rr++)
        for (int cc = 0; cc < bs; cc++) src[oV + (size_t)(rr * bs + cc) * 32] = rp[b + rr] + p * bs + cc;
    for (int q = 0; q < cnt; q++) {
        nbr[oI + (size_t)q * 32] = cl[r0 + (first + q) * bs];
        for (int rr = 0; rr < bs; rr++)
            for (int cc = 0; cc < bs; cc++) src[oV + (size_t)(BB * (1 + q) + rr * bs + cc) * 32] = rp[b + rr] + (first + q) * bs + cc;
    }
}

// scal layout (device doubles): 0 rz, 1 pAp, 2 alpha, 3 beta, 4 rr, 5 rzNew, 6 bb
enum { S_RZ = 0, S_PAP = 1, S_ALPHA = 2, S_BETA = 3, S_RR = 4, S_RZNEW = 5, S_BB = 6, S_COUNT = 8 };

// x += alpha p ; r -= alpha q ; JACOBI: z = r * invDiag ; partial sums of r.z (JACOBI) and r.r
template <bool JACOBI>
__global__ void __launch_bounds__(TPB) k_cg_update(int n, const double *scal, const double *__restrict__ p,
                                                    const double *__restrict__ q, const double *__restrict__ invDiag, double *x,
                                                    double *r, double *z, double *partRZ, double *partRR) {
    __shared__ double sh[32];
    const int i = blockIdx.x * TPB + threadIdx.x;
    const double alpha = scal[S_ALPHA];
    double rz = 0.0, rr = 0.0;
    if (i < n) {
        x[i] += alpha * p[i];
        const double ri = r[i] - alpha * q[i];
        r[i] = ri;
        rr = ri * ri;
        if (JACOBI) {
            const double zi = ri * invDiag[i];
            z[i] = zi;
            rz = ri * zi;
        }
    }
    const double s1 = block_sum_d(rz, sh);
    const double s2 = block_sum_d(rr, sh);
    if (threadIdx.x == 0) { partRZ[blockIdx.x] = s1; partRR[blockIdx.x] = s2; }
}

// partial sums of a.b and c.c
__global__ void __launch_bounds__(TPB) k_dot2(int n, const double *__restrict__ a, const double *__restrict__ b,
                                               const double *__restrict__ c, double *partAB, double *partCC) {
    __shared__ double sh[32];
    const int i = blockIdx.x * TPB + threadIdx.x;
    double ab = 0.0, cc = 0.0;
    if (i < n) { ab = a[i] * b[i]; cc = c[i] * c[i]; }
    const double s1 = block_sum_d(ab, sh);
    const double s2 = block_sum_d(cc, sh);
    if (threadIdx.x == 0) { partAB[blockIdx.x] = s1; partCC[blockIdx.x] = s2; }
}

__global__ void k_jacobi(int n, const double *__restrict__ r, const double *__restrict__ invDiag, double *z) {
    const int i = blockIdx.x * TPB + threadIdx.x;
    if (i < n) z[i] = r[i] * invDiag[i];
}

// Final (fixed-order) reduction of up to two partial arrays by one block, then the CG scalar algebra.
//   mode 0: scal[S_PAP] = sum(p0); alpha = rz / pAp
//   mode 1: scal[S_RZNEW] = sum(p0); scal[S_RR] = sum(p1); beta = rzNew / rz; rz = rzNew
//   mode 2: initial: scal[S_RZ] = sum(p0); scal[S_RR] = sum(p1)
//   mode 3: scal[S_BB] = sum(p1)
__global__ void __launch_bounds__(1024) k_scalars(int nPart, const double *p0, const double *p1, double *scal, int mode) {
    __shared__ double sh[32];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < nPart; i += blockDim.x) {   // strided: every thread's chain has a fixed order
        if (p0) a += p0[i];
        if (p1) b += p1[i];
    }
    const double sa = block_sum_d(a, sh);
    const double sb = block_sum_d(b, sh);
    if (threadIdx.x == 0) {
        if (mode == 0) { scal[S_PAP] = sa; scal[S_ALPHA] = scal[S_RZ] / sa; }
        else if (mode == 1) { scal[S_RZNEW] = sa; scal[S_RR] = sb; scal[S_BETA] = sa / scal[S_RZ]; scal[S_RZ] = sa; }
        else if (mode == 2) { scal[S_RZ] = sa; scal[S_RR] = sb; }
        else { scal[S_BB] = sb; }
    }
}

__global__ void __launch_bounds__(TPB) k_xpay(int n, const double *scal, const double *__restrict__ z, double *p) {
    const int i = blockIdx.x * TPB + threadIdx.x;
    if (i < n) p[i] = z[i] + scal[S_BETA] * p[i];
}

// ---- BiCGSTAB vector kernels (LASolver scaler_cgstab, lib/LASolver/accel_class.cpp:280-399); scalars come from the host ----
// p = r + beta (p - omega v)
__global__ void __launch_bounds__(TPB) k_bi_p(int n, double beta, double omega, const double *__restrict__ r,
                                               const double *__restrict__ v, double *p) {
    const int i = blockIdx.x * TPB + threadIdx.x;
    if (i < n) p[i] = r[i] + beta * (p[i] - omega * v[i]);
}
// s = r - alpha v
__global__ void __launch_bounds__(TPB) k_bi_s(int n, double alpha, const double *__restrict__ r, const double *__restrict__ v, double *s) {
    const int i = blockIdx.x * TPB + threadIdx.x;
    if (i < n) s[i] = r[i] - alpha * v[i];
}
// x += alpha vbar + omega zvec ; r = s - omega t ; partial sums of r.r
__global__ void __launch_bounds__(TPB) k_bi_xr(int n, double alpha, double omega, const double *__restrict__ vbar,
                                                const double *__restrict__ zvec, const double *__restrict__ s,
                                                const double *__restrict__ t, double *x, double *r, double *partRR) {
    __shared__ double sh[32];
    const int i = blockIdx.x * TPB + threadIdx.x;
    double rr = 0.0;
    if (i < n) {
        const double step = alpha * vbar[i] + omega * zvec[i];
        x[i] = x[i] + step;
        const double ri = s[i] - omega * t[i];
        r[i] = ri;
        rr = ri * ri;
    }
    const double s1 = block_sum_d(rr, sh);
    if (threadIdx.x == 0) partRR[blockIdx.x] = s1;
}

// FP64 pipe peak: 8 independent DFMA chains per thread, no memory traffic
__global__ void __launch_bounds__(256) k_dfma_peak(int iters, double seed, double *out) {
    double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; i++) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    const double s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 123.456) out[0] = s;   // never true: keeps the chains alive
}

}  // namespace

struct mmadmm_pcg {
    int n = 0, nnz = 0, precond = 0, device = 0, lanes = 1;
    cudaStream_t stream = nullptr;
    int *rowptr = nullptr, *col = nullptr, *diagPos = nullptr;
    double *val = nullptr, *lu = nullptr, *invDiag = nullptr;
    int *rowsL = nullptr, *rowsU = nullptr;       // rows ordered by level
    int *lvlL_d = nullptr, *lvlU_d = nullptr;     // level offsets on the device (k_tri_cluster)
    int2 *kbeL = nullptr, *kbeU = nullptr;        // off-diagonal extent [kb, ke) of the rows in level order (k_tri_cluster)
    int triMode = 1;                              // 1: one cluster launch per triangular solve; 0: one launch per level
    int *flag_d = nullptr;                        // zero-pivot flag of the device ILU(0) factor
    int rowBlock = 1;                             // "row_block": rows per group of the triangular solves (1 = row per thread)
    int *grpL = nullptr, *grpU = nullptr;         // base rows of the groups in level order
    int *lvlGL_d = nullptr, *lvlGU_d = nullptr;   // group-level offsets (device)
    int2 *kbeGL = nullptr, *kbeGU = nullptr;      // [group in level order][row of the group]: off-diagonal extent
    std::vector<int> lvlGL, lvlGU;                // group-level offsets (host)
    int triLayout = 0;                            // "tri_layout": 1 = level-ELL copy of the factor for the cluster kernels
    struct Chell {                                // chunk-ELL copy of the factor for the dataflow solve ("tri_mode" 2)
        longlong2 *meta_d = nullptr; int *items = nullptr, *col = nullptr, *src = nullptr; double *val = nullptr;
        size_t slots = 0; int nChunks = 0;
        bool blocked = false;                     // BS x BS block form (k_tri_sf_block): col = neighbour planes, val = block planes
    } chellL, chellU;
    unsigned *sfCounter = nullptr;
    int sfGridX2 = 2;                             // dataflow grid = sfGrid * sfGridX2 / 2 CTAs ("tri_sf_ctas_per_sm_x2")
    int sfGrid = 0;                               // CTAs of the dataflow kernel (SM count x resident CTAs per SM)
    int sfProxy = 0;                              // "tri_proxy": block dataflow kernel waits through one proxy lane per warp first
    int sfBlocks = 1;                             // "tri_blocks": use the BS x BS block form of the dataflow solve where the pattern allows
    struct Lell {
        int4 *meta_d = nullptr; int *col = nullptr, *src = nullptr; double *val = nullptr; size_t slots = 0;
        std::vector<int4> meta;
    } lellL, lellU;
    std::vector<int> diagPos_h;
    int iluThreadKernel = 0;                      // "ilu_thread_kernel" 1: thread-per-row factor kernel even where the warp form fits (tests)
    int maxRowLen = 0;                            // longest row (stored entries): <= 64 -> warp-per-row ILU(0) factor kernel
    std::vector<int> rowsL_h;                     // rows in lower-level order (host), for the deferred part of the row schedule
    bool rowSchedule = false;                     // ensure_row_schedule has run
    bool valHostStale = false;                    // val_h older than the device values (after mmadmm_pcg_refactor_device)
    int spmvFormat = 0;                           // 0: CSR, sub-warp per row; 1: SELL-32-4096 (built on first use)
    int nSlices = 0;
    long long sellEntries = 0;
    int *slicePtr = nullptr, *scol = nullptr, *perm = nullptr;
    double *sval = nullptr;
    std::vector<int> rowptr_h, col_h;             // host copy of the pattern (SELL build)
    std::vector<double> val_h;
    std::vector<int> lvlL, lvlU;                  // level offsets (host)
    double *r = nullptr, *z = nullptr, *p = nullptr, *q = nullptr, *x = nullptr, *b = nullptr, *tmp = nullptr;
    double *part0 = nullptr, *part1 = nullptr, *scal = nullptr, *scal_h = nullptr;
    double *bi[4] = {nullptr, nullptr, nullptr, nullptr};   // BiCGSTAB work vectors (res0, svec, zvec, tvec), allocated on first use
    int nblk = 0;
    long long launches = 0;
    std::string err;
};

#define PCK(call)                                                                                   \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                            \
            return MMADMM_ECUDA;                                                                    \
        }                                                                                           \
    } while (0)

namespace {

void launch_spmv(mmadmm_pcg *h, const double *x, double *y, double *part) {
    const int n = h->n;
    if (h->spmvFormat == 1) {
        k_spmv_sell<<<cdiv((size_t)h->nSlices * 32, TPB), TPB, 0, h->stream>>>(h->nSlices, h->slicePtr, h->scol, h->sval, h->perm, x, y, part);
        h->launches++;
        return;
    }
    const size_t threads = (size_t)n * h->lanes;
    const int g = cdiv(threads, TPB);
    switch (h->lanes) {
        case 1: k_spmv_dot<1><<<g, TPB, 0, h->stream>>>(n, h->rowptr, h->col, h->val, x, y, part); break;
        case 2: k_spmv_dot<2><<<g, TPB, 0, h->stream>>>(n, h->rowptr, h->col, h->val, x, y, part); break;
        case 4: k_spmv_dot<4><<<g, TPB, 0, h->stream>>>(n, h->rowptr, h->col, h->val, x, y, part); break;
        case 8: k_spmv_dot<8><<<g, TPB, 0, h->stream>>>(n, h->rowptr, h->col, h->val, x, y, part); break;
        case 16: k_spmv_dot<16><<<g, TPB, 0, h->stream>>>(n, h->rowptr, h->col, h->val, x, y, part); break;
        default: k_spmv_dot<32><<<g, TPB, 0, h->stream>>>(n, h->rowptr, h->col, h->val, x, y, part); break;
    }
    h->launches++;
}
int spmv_parts(const mmadmm_pcg *h) {
    return h->spmvFormat == 1 ? cdiv((size_t)h->nSlices * 32, TPB) : cdiv((size_t)h->n * h->lanes, TPB);
}
// CSR -> SELL-32-4096 on the host (once), uploaded next to the CSR arrays
int build_sell(mmadmm_pcg *h) {
    if (h->slicePtr) return MMADMM_OK;
    if (h->valHostStale) {
        PCK(cudaSetDevice(h->device));
        PCK(cudaStreamSynchronize(h->stream));
        h->val_h.resize((size_t)h->nnz);
        PCK(cudaMemcpy(h->val_h.data(), h->val, sizeof(double) * (size_t)h->nnz, cudaMemcpyDeviceToHost));
        h->valHostStale = false;
    }
    const int n = h->n, C = 32, SIGMA = 4096;
    const std::vector<int> &rp = h->rowptr_h, &cl = h->col_h;
    const std::vector<double> &vl = h->val_h;
    const int nSlices = (n + C - 1) / C;
    std::vector<int> perm((size_t)nSlices * C, -1);
    for (int i = 0; i < n; i++) perm[i] = i;
    for (int w0 = 0; w0 < n; w0 += SIGMA) {
        const int w1 = std::min(n, w0 + SIGMA);
        std::stable_sort(perm.begin() + w0, perm.begin() + w1, [&](int a, int b) { return rp[a + 1] - rp[a] > rp[b + 1] - rp[b]; });
    }
    std::vector<int> sp(nSlices + 1, 0);
    for (int s = 0; s < nSlices; s++) {
        int w = 0;
        for (int r = 0; r < C; r++) { const int row = perm[(size_t)s * C + r]; if (row >= 0) w = std::max(w, rp[row + 1] - rp[row]); }
        sp[s + 1] = sp[s] + w * C;
    }
    std::vector<int> scol((size_t)sp[nSlices], 0);
    std::vector<double> sval((size_t)sp[nSlices], 0.0);
    for (int s = 0; s < nSlices; s++)
        for (int r = 0; r < C; r++) {
            const int row = perm[(size_t)s * C + r];
            if (row < 0) continue;
            for (int k = rp[row], j = 0; k < rp[row + 1]; k++, j++) {
                scol[(size_t)sp[s] + (size_t)j * C + r] = cl[k];
                sval[(size_t)sp[s] + (size_t)j * C + r] = vl[k];
            }
        }
    PCK(cudaSetDevice(h->device));
    PCK(cudaMalloc(&h->slicePtr, sizeof(int) * sp.size()));
    PCK(cudaMalloc(&h->perm, sizeof(int) * perm.size()));
    PCK(cudaMalloc(&h->scol, sizeof(int) * std::max<size_t>(scol.size(), 1)));
    PCK(cudaMalloc(&h->sval, sizeof(double) * std::max<size_t>(sval.size(), 1)));
    PCK(cudaMemcpy(h->slicePtr, sp.data(), sizeof(int) * sp.size(), cudaMemcpyHostToDevice));
    PCK(cudaMemcpy(h->perm, perm.data(), sizeof(int) * perm.size(), cudaMemcpyHostToDevice));
    PCK(cudaMemcpy(h->scol, scol.data(), sizeof(int) * scol.size(), cudaMemcpyHostToDevice));
    PCK(cudaMemcpy(h->sval, sval.data(), sizeof(double) * sval.size(), cudaMemcpyHostToDevice));
    PCK(cudaDeviceSynchronize());   // as in build_groups: the uploads above must have landed before h->stream reads them
    h->nSlices = nSlices;
    h->sellEntries = sp[nSlices];
    return MMADMM_OK;
}

int lell_refresh(mmadmm_pcg *h);
// lu <- ILU(0) of the values in h->val (device), by the L-level schedule; h->flag_d is the zero-pivot flag
int ilu_factor_device(mmadmm_pcg *h) {
    PCK(cudaMemcpyAsync(h->lu, h->val, sizeof(double) * (size_t)h->nnz, cudaMemcpyDeviceToDevice, h->stream));
    PCK(cudaMemsetAsync(h->flag_d, 0, sizeof(int), h->stream));
    const int nlev = (int)h->lvlL.size() - 1;
    for (int l = 0; l < nlev; l++) {   // (level 0 rows have no entries left of the diagonal: only their pivot is checked)
        const int lo = h->lvlL[l], hi = h->lvlL[l + 1];
        if (h->maxRowLen <= 64 && !h->iluThreadKernel) k_ilu_level_warp<<<cdiv((size_t)(hi - lo) * 32, TPB), TPB, 0, h->stream>>>(lo, hi, h->rowsL, h->rowptr, h->col, h->diagPos, h->lu, h->flag_d);
        else k_ilu_level<<<cdiv(hi - lo, TPB), TPB, 0, h->stream>>>(lo, hi, h->rowsL, h->rowptr, h->col, h->diagPos, h->lu, h->flag_d);
        h->launches++;
    }
    int flag = 0;
    PCK(cudaMemcpyAsync(&flag, h->flag_d, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    PCK(cudaStreamSynchronize(h->stream));
    PCK(cudaGetLastError());
    if (flag) { h->err = "zero pivot in ILU(0)"; return MMADMM_ESINGULAR; }
    return lell_refresh(h);
}

// The part of the row-per-thread schedule that only tri_mode 0 / 1 with row_block 1 read (mmadmm_pcg_create builds just what
// the factorisation needs): upper levels and rows in upper-level order, row extents in level order, level offsets on the device.
int ensure_row_schedule(mmadmm_pcg *h) {
    if (h->rowSchedule || h->precond != 1) return MMADMM_OK;
    const int n = h->n;
    const std::vector<int> &rp = h->rowptr_h, &cl = h->col_h, &dg = h->diagPos_h, &rowsL = h->rowsL_h;
    PCK(cudaSetDevice(h->device));
    std::vector<int> lev(n, 0), rowsU(n);
    int nu = 0;
    for (int i = n - 1; i >= 0; i--) {
        int l = 0;
        for (int k = dg[i] + 1; k < rp[i + 1]; k++) l = std::max(l, lev[cl[k]] + 1);
        lev[i] = l; nu = std::max(nu, l + 1);
    }
    h->lvlU.assign(nu + 1, 0);
    for (int i = 0; i < n; i++) h->lvlU[lev[i] + 1]++;
    for (int l = 0; l < nu; l++) h->lvlU[l + 1] += h->lvlU[l];
    {
        std::vector<int> cur(h->lvlU.begin(), h->lvlU.end() - 1);
        for (int i = 0; i < n; i++) rowsU[cur[lev[i]]++] = i;
    }
    std::vector<int2> eL(n), eU(n);
    for (int t = 0; t < n; t++) {
        eL[t] = make_int2(rp[rowsL[t]], dg[rowsL[t]]);
        eU[t] = make_int2(dg[rowsU[t]] + 1, rp[rowsU[t] + 1]);
    }
    auto up = [&](void **dst, const void *src, size_t bytes) -> cudaError_t {
        cudaError_t e = cudaMalloc(dst, std::max<size_t>(bytes, 8));
        return e != cudaSuccess ? e : cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice);
    };
    PCK(up((void **)&h->rowsU, rowsU.data(), sizeof(int) * (size_t)n));
    PCK(up((void **)&h->kbeL, eL.data(), sizeof(int2) * (size_t)n));
    PCK(up((void **)&h->kbeU, eU.data(), sizeof(int2) * (size_t)n));
    PCK(up((void **)&h->lvlL_d, h->lvlL.data(), sizeof(int) * h->lvlL.size()));
    PCK(up((void **)&h->lvlU_d, h->lvlU.data(), sizeof(int) * h->lvlU.size()));
    PCK(cudaDeviceSynchronize());   // pageable uploads must have landed before h->stream (non-blocking) reads them
    h->rowSchedule = true;
    return MMADMM_OK;
}

// ---- host-side planning of the group schedules (no device needed: mmadmm_pcg_plan_levels, tests/test_pcg_plan.py) ----
// Level of every group [g*bs, min(n, (g+1)*bs)) of the lower (upper) solve: 1 + the highest level among the OTHER groups that hold
// a column left (right) of the group in one of its rows; rows of a group are solved in sequence by one thread.  Returns the
// number of levels.
int plan_group_levels(int n, int bs, bool lower, const int *rp, const int *cl, const int *dg, std::vector<int> &lev) {
    const int ng = (n + bs - 1) / bs;
    lev.assign(ng, 0);
    int nlev = 0;
    if (lower) {
        for (int g = 0; g < ng; g++) {
            int l = 0;
            for (int i = g * bs; i < std::min(n, (g + 1) * bs); i++)
                for (int k = rp[i]; k < dg[i]; k++)
                    if (cl[k] < g * bs) l = std::max(l, lev[cl[k] / bs] + 1);
            lev[g] = l; nlev = std::max(nlev, l + 1);
        }
    } else {
        for (int g = ng - 1; g >= 0; g--) {
            int l = 0;
            for (int i = g * bs; i < std::min(n, (g + 1) * bs); i++)
                for (int k = dg[i] + 1; k < rp[i + 1]; k++)
                    if (cl[k] >= (g + 1) * bs) l = std::max(l, lev[cl[k] / bs] + 1);
            lev[g] = l; nlev = std::max(nlev, l + 1);
        }
    }
    return nlev;
}
// Block form of the dataflow solve (k_tri_sf_block) is possible when every group is bs full rows whose columns are the same
// aligned runs of bs (dense bs x bs blocks, the group's own block among them).
bool plan_blockable(int n, int bs, const int *rp, const int *cl) {
    if ((bs != 2 && bs != 3) || n % bs) return false;
    const int ng = n / bs;
    for (int g = 0; g < ng; g++) {
        const int i0 = g * bs, len0 = rp[i0 + 1] - rp[i0];
        if (len0 % bs) return false;
        bool own = false;
        for (int k = rp[i0]; k < rp[i0 + 1]; k += bs) {
            if (cl[k] % bs) return false;
            for (int c = 1; c < bs; c++)
                if (cl[k + c] != cl[k] + c) return false;
            own = own || cl[k] == i0;
        }
        if (!own) return false;
        for (int r = 1; r < bs; r++) {
            if (rp[i0 + r + 1] - rp[i0 + r] != len0) return false;
            for (int k = 0; k < len0; k++)
                if (cl[rp[i0 + r] + k] != cl[rp[i0] + k]) return false;
        }
    }
    return true;
}

void free_groups(mmadmm_pcg *h) {
    void *ptrs[6] = {h->grpL, h->grpU, h->lvlGL_d, h->lvlGU_d, h->kbeGL, h->kbeGU};
    for (void *q : ptrs) {
        const cudaError_t e = cudaFree(q);
        if (e != cudaSuccess) h->err = std::string("free_groups: cudaFree: ") + cudaGetErrorString(e);
    }
    h->grpL = h->grpU = h->lvlGL_d = h->lvlGU_d = nullptr;
    h->kbeGL = h->kbeGU = nullptr;
    h->lvlGL.clear(); h->lvlGU.clear();
    for (mmadmm_pcg::Chell *L : {&h->chellL, &h->chellU}) {
        cudaFree(L->meta_d); cudaFree(L->items); cudaFree(L->col); cudaFree(L->src); cudaFree(L->val);
        L->meta_d = nullptr; L->items = L->col = L->src = nullptr; L->val = nullptr; L->slots = 0; L->nChunks = 0; L->blocked = false;
    }
    for (mmadmm_pcg::Lell *L : {&h->lellL, &h->lellU}) {
        cudaFree(L->meta_d); cudaFree(L->col); cudaFree(L->src); cudaFree(L->val);
        L->meta_d = nullptr; L->col = L->src = nullptr; L->val = nullptr; L->slots = 0; L->meta.clear();
    }
}
// factor values -> the level-ELL copies (after every numeric factorisation)
int lell_refresh(mmadmm_pcg *h) {
    for (mmadmm_pcg::Lell *L : {&h->lellL, &h->lellU})
        if (L->slots) {
            k_lell_fill<<<cdiv(L->slots, TPB), TPB, 0, h->stream>>>(L->slots, L->src, h->lu, L->val);
            h->launches++;
        }
    for (mmadmm_pcg::Chell *L : {&h->chellL, &h->chellU})
        if (L->slots) {
            k_lell_fill<<<cdiv(L->slots, TPB), TPB, 0, h->stream>>>(L->slots, L->src, h->lu, L->val);
            h->launches++;
        }
    PCK(cudaGetLastError());
    return MMADMM_OK;
}
// Level schedule of the groups [g*bs, (g+1)*bs): level(g) = 1 + max level of the OTHER groups its rows depend on.
int build_groups(mmadmm_pcg *h, int bs) {
    free_groups(h);
    h->rowBlock = 1;
    if (bs == 1 && !h->triLayout && h->triMode != 2) return MMADMM_OK;
    const int n = h->n;
    if (h->precond != 1 || (bs != 1 && bs != 2 && bs != 3 && bs != 4 && bs != 6)) {
        h->err = "row_block: 1, 2, 3, 4 or 6 rows per group, on an ILU(0) handle";
        return MMADMM_EINVAL;
    }
    const int ng = (n + bs - 1) / bs;   // the last group may be short
    const std::vector<int> &rp = h->rowptr_h, &cl = h->col_h, &dg = h->diagPos_h;
    std::vector<int> lev(ng, 0);
    auto bucket = [&](int nlev, std::vector<int> &grp, std::vector<int> &off) {
        off.assign(nlev + 1, 0);
        for (int g = 0; g < ng; g++) off[lev[g] + 1]++;
        for (int l = 0; l < nlev; l++) off[l + 1] += off[l];
        grp.resize(ng);
        std::vector<int> cur(off.begin(), off.end() - 1);
        for (int g = 0; g < ng; g++) grp[cur[lev[g]]++] = g * bs;
    };
    auto upload = [&](const std::vector<int> &grp, const std::vector<int> &off, bool lower, int **grp_d, int **off_d, int2 **kbe_d) -> cudaError_t {
        std::vector<int2> e((size_t)ng * bs);
        for (int t = 0; t < ng; t++)
            for (int j = 0; j < bs; j++) {
                const int i = grp[t] + j;
                e[(size_t)t * bs + j] = i >= n ? make_int2(0, 0) : lower ? make_int2(rp[i], dg[i]) : make_int2(dg[i] + 1, rp[i + 1]);
            }
        cudaError_t er;
        if ((er = cudaMalloc(grp_d, sizeof(int) * (size_t)ng)) != cudaSuccess) return er;
        if ((er = cudaMalloc(off_d, sizeof(int) * off.size())) != cudaSuccess) return er;
        if ((er = cudaMalloc(kbe_d, sizeof(int2) * e.size())) != cudaSuccess) return er;
        if ((er = cudaMemcpy(*grp_d, grp.data(), sizeof(int) * (size_t)ng, cudaMemcpyHostToDevice)) != cudaSuccess) return er;
        if ((er = cudaMemcpy(*off_d, off.data(), sizeof(int) * off.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return er;
        return cudaMemcpy(*kbe_d, e.data(), sizeof(int2) * e.size(), cudaMemcpyHostToDevice);
    };
    // level-ELL block of every level of at most one cluster pass (see k_tri_cluster_lell)
    auto lell = [&](const std::vector<int> &grp, const std::vector<int> &off, bool lower, mmadmm_pcg::Lell &L) -> cudaError_t {
        const int nlev = (int)off.size() - 1, P0 = lower ? 0 : 1, maxw = TC_THREADS * TC_CLUSTER;
        auto len = [&](int i) { return lower ? dg[i] - rp[i] : rp[i + 1] - dg[i] - 1; };
        L.meta.assign(nlev, make_int4(0, 0, -1, 0));
        size_t slots = 0;
        for (int l = 0; l < nlev; l++) {
            const int lo = off[l], hi = off[l + 1], w = hi - lo;
            L.meta[l] = make_int4(lo, hi, -1, 0);
            if (w > maxw) continue;
            int m = 0;
            for (int t = lo; t < hi; t++)
                for (int j = 0; j < bs; j++)
                    if (grp[t] + j < n) m = std::max(m, len(grp[t] + j));
            const size_t need = (size_t)bs * (m + P0) * w;
            if (slots + need >= (size_t)1 << 31) continue;   // offsets are ints: the rest stays in CSR
            L.meta[l].z = (int)slots; L.meta[l].w = m;
            slots += need;
        }
        std::vector<int> colv(slots, -1), srcv(slots, -1);
        for (int l = 0; l < nlev; l++) {
            const int4 mt = L.meta[l];
            if (mt.z < 0) continue;
            const int w = mt.y - mt.x, MP = mt.w + P0;
            for (int t = mt.x; t < mt.y; t++)
                for (int j = 0; j < bs; j++) {
                    const int i = grp[t] + j;
                    if (i >= n) continue;
                    const size_t o0 = (size_t)mt.z + (t - mt.x);
                    if (!lower) { colv[o0 + (size_t)(j * MP) * w] = -2; srcv[o0 + (size_t)(j * MP) * w] = dg[i]; }
                    const int kb = lower ? rp[i] : dg[i] + 1;
                    for (int q = 0; q < len(i); q++) {
                        const size_t o = o0 + (size_t)(j * MP + P0 + q) * w;
                        colv[o] = cl[kb + q]; srcv[o] = kb + q;
                    }
                }
        }
        L.slots = slots;
        cudaError_t er;
        if ((er = cudaMalloc(&L.meta_d, sizeof(int4) * std::max<size_t>(L.meta.size(), 1))) != cudaSuccess) return er;
        if ((er = cudaMemcpy(L.meta_d, L.meta.data(), sizeof(int4) * L.meta.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return er;
        if (!slots) return cudaSuccess;
        if ((er = cudaMalloc(&L.col, sizeof(int) * slots)) != cudaSuccess) return er;
        if ((er = cudaMalloc(&L.src, sizeof(int) * slots)) != cudaSuccess) return er;
        if ((er = cudaMalloc(&L.val, sizeof(double) * slots)) != cudaSuccess) return er;
        if ((er = cudaMemcpy(L.col, colv.data(), sizeof(int) * slots, cudaMemcpyHostToDevice)) != cudaSuccess) return er;
        return cudaMemcpy(L.src, srcv.data(), sizeof(int) * slots, cudaMemcpyHostToDevice);
    };
    // chunk-ELL copy for the dataflow solve: chunks of 32 groups in level order, every level padded to whole chunks
    auto chell = [&](const std::vector<int> &grp, const std::vector<int> &off, bool lower, mmadmm_pcg::Chell &L) -> cudaError_t {
        const int nlev = (int)off.size() - 1, P0 = lower ? 0 : 1;
        auto len = [&](int i) { return lower ? dg[i] - rp[i] : rp[i + 1] - dg[i] - 1; };
        std::vector<int> items;
        items.reserve((size_t)ng + 32 * (size_t)nlev);
        for (int l = 0; l < nlev; l++) {
            for (int t = off[l]; t < off[l + 1]; t++) items.push_back(grp[t]);
            while (items.size() % 32) items.push_back(-1);
        }
        const int nChunks = (int)(items.size() / 32);
        std::vector<longlong2> meta(nChunks);
        size_t slots = 0;
        for (int c = 0; c < nChunks; c++) {
            int m = 0;
            for (int r = 0; r < 32; r++) {
                const int b = items[(size_t)c * 32 + r];
                if (b < 0) continue;
                for (int j = 0; j < bs; j++)
                    if (b + j < n) m = std::max(m, len(b + j));
            }
            meta[c] = make_longlong2((long long)slots, m);
            slots += (size_t)bs * (m + P0) * 32;
        }
        std::vector<int> colv(slots, -1), srcv(slots, -1);
#pragma omp parallel for schedule(static)
        for (int c = 0; c < nChunks; c++) {   // chunks own disjoint slots
            const int MP = (int)meta[c].y + P0;
            for (int r = 0; r < 32; r++) {
                const int b = items[(size_t)c * 32 + r];
                if (b < 0) continue;
                const size_t o0 = (size_t)meta[c].x + r;
                for (int j = 0; j < bs; j++) {
                    const int i = b + j;
                    if (i >= n) continue;
                    if (!lower) { colv[o0 + (size_t)(j * MP) * 32] = -2; srcv[o0 + (size_t)(j * MP) * 32] = dg[i]; }
                    const int kb = lower ? rp[i] : dg[i] + 1;
                    for (int q = 0; q < len(i); q++) {
                        const size_t o = o0 + (size_t)(j * MP + P0 + q) * 32;
                        colv[o] = cl[kb + q]; srcv[o] = kb + q;
                    }
                }
            }
        }
        L.slots = slots; L.nChunks = nChunks;
        cudaError_t er;
        if ((er = cudaMalloc(&L.meta_d, sizeof(longlong2) * std::max<size_t>(meta.size(), 1))) != cudaSuccess) return er;
        if ((er = cudaMalloc(&L.items, sizeof(int) * std::max<size_t>(items.size(), 1))) != cudaSuccess) return er;
        if ((er = cudaMalloc(&L.col, sizeof(int) * std::max<size_t>(slots, 1))) != cudaSuccess) return er;
        if ((er = cudaMalloc(&L.src, sizeof(int) * std::max<size_t>(slots, 1))) != cudaSuccess) return er;
        if ((er = cudaMalloc(&L.val, sizeof(double) * std::max<size_t>(slots, 1))) != cudaSuccess) return er;
        if ((er = cudaMemcpy(L.meta_d, meta.data(), sizeof(longlong2) * meta.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return er;
        if ((er = cudaMemcpy(L.items, items.data(), sizeof(int) * items.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return er;
        if ((er = cudaMemcpy(L.col, colv.data(), sizeof(int) * slots, cudaMemcpyHostToDevice)) != cudaSuccess) return er;
        return cudaMemcpy(L.src, srcv.data(), sizeof(int) * slots, cudaMemcpyHostToDevice);
    };
    auto chell_block = [&](const std::vector<int> &grp, const std::vector<int> &off, bool lower, mmadmm_pcg::Chell &L) -> cudaError_t {
        const int nlev = (int)off.size() - 1, BB = bs * bs;
        std::vector<int> items;
        items.reserve((size_t)ng + 32 * (size_t)nlev);
        for (int l = 0; l < nlev; l++) {
            for (int t = off[l]; t < off[l + 1]; t++) items.push_back(grp[t]);
            while (items.size() % 32) items.push_back(-1);
        }
        const int nChunks = (int)(items.size() / 32);
        // position of the group's own block in its row, number of neighbour blocks on this side
        auto ownPos = [&](int b) { int p = 0; while (cl[rp[b] + p * bs] != b) p++; return p; };
        auto nNbr = [&](int b) { const int p = ownPos(b); return lower ? p : (rp[b + 1] - rp[b]) / bs - p - 1; };
        std::vector<longlong2> meta(nChunks);
        size_t slotsV = 0, slotsI = 0;
        for (int c = 0; c < nChunks; c++) {
            int m = 0;
            for (int r = 0; r < 32; r++) { const int b = items[(size_t)c * 32 + r]; if (b >= 0) m = std::max(m, nNbr(b)); }
            if (m > 255) return cudaErrorInvalidValue;
            meta[c] = make_longlong2((long long)slotsV, (long long)((slotsI << 8) | (size_t)m));
            slotsV += (size_t)BB * (1 + m) * 32;
            slotsI += (size_t)m * 32;
        }
        L.slots = slotsV; L.nChunks = nChunks; L.blocked = true;
        const size_t nI = std::max<size_t>(slotsI, 1), nV = std::max<size_t>(slotsV, 1);
        cudaError_t er;
        if ((er = cudaMalloc(&L.meta_d, sizeof(longlong2) * std::max<size_t>(meta.size(), 1))) != cudaSuccess) return er;
        if ((er = cudaMalloc(&L.items, sizeof(int) * std::max<size_t>(items.size(), 1))) != cudaSuccess) return er;
        if ((er = cudaMalloc(&L.col, sizeof(int) * nI)) != cudaSuccess) return er;
        if ((er = cudaMalloc(&L.src, sizeof(int) * nV)) != cudaSuccess) return er;
        if ((er = cudaMalloc(&L.val, sizeof(double) * nV)) != cudaSuccess) return er;
        if ((er = cudaMemcpy(L.meta_d, meta.data(), sizeof(longlong2) * meta.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return er;
        if ((er = cudaMemcpy(L.items, items.data(), sizeof(int) * items.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return er;
        if ((er = cudaDeviceSynchronize()) != cudaSuccess) return er;   // pageable uploads landed before h->stream reads them
        // the index planes (0.2 GB per factor at C2 size) are written by the device, not filled on the host and uploaded
        if ((er = cudaMemsetAsync(L.col, 0xFF, sizeof(int) * nI, h->stream)) != cudaSuccess) return er;
        if ((er = cudaMemsetAsync(L.src, 0xFF, sizeof(int) * nV, h->stream)) != cudaSuccess) return er;
        k_chell_block_fill<<<cdiv((size_t)nChunks * 32, TPB), TPB, 0, h->stream>>>(nChunks, bs, lower ? 1 : 0, L.meta_d, L.items, h->rowptr, h->col, L.col, L.src);
        h->launches++;
        return cudaGetLastError();
    };
    SetupTimer tg("build_groups");
    const bool useBlocks = h->triMode == 2 && h->sfBlocks && plan_blockable(n, bs, rp.data(), cl.data());
    tg.lap("block-structure check");
    PCK(cudaSetDevice(h->device));
    if (!h->sfCounter) PCK(cudaMalloc(&h->sfCounter, sizeof(unsigned)));
    std::vector<int> grp;
    const int nl = plan_group_levels(n, bs, true, rp.data(), cl.data(), dg.data(), lev);
    tg.lap("lower group levels");
    bucket(nl, grp, h->lvlGL);
    PCK(upload(grp, h->lvlGL, true, &h->grpL, &h->lvlGL_d, &h->kbeGL));
    if (h->triLayout && h->triMode != 2) PCK(lell(grp, h->lvlGL, true, h->lellL));
    if (h->triMode == 2) PCK(useBlocks ? chell_block(grp, h->lvlGL, true, h->chellL) : chell(grp, h->lvlGL, true, h->chellL));
    tg.lap("lower schedule + factor copy layout + uploads");
    const int nu = plan_group_levels(n, bs, false, rp.data(), cl.data(), dg.data(), lev);
    tg.lap("upper group levels");
    bucket(nu, grp, h->lvlGU);
    PCK(upload(grp, h->lvlGU, false, &h->grpU, &h->lvlGU_d, &h->kbeGU));
    if (h->triLayout && h->triMode != 2) PCK(lell(grp, h->lvlGU, false, h->lellU));
    if (h->triMode == 2) PCK(useBlocks ? chell_block(grp, h->lvlGU, false, h->chellU) : chell(grp, h->lvlGU, false, h->chellU));
    tg.lap("upper schedule + factor copy layout + uploads");
    h->rowBlock = bs;
    // cudaMemcpy from pageable memory may return before the DMA has landed, and h->stream does not synchronise with the legacy
    // stream: without this the fill kernel below could read a half-uploaded index array (seen as a rare, per-process-stable mismatch)
    PCK(cudaDeviceSynchronize());
    return lell_refresh(h);
}

// One triangular solve.  tri_mode 1: runs of consecutive levels that each fit a few passes of the cluster's 2048 threads go
// out as ONE k_tri_cluster launch (natural-order ILU(0) on a structured 2D mesh: thousands of levels of a few hundred rows);
// wider levels (e.g. all cell-centre rows of the upper factor, plane-sized levels in 3D) keep one full-width launch each.
template <bool LOWER>
void tri_solve(mmadmm_pcg *h, const std::vector<int> &lvl, const int *lvl_d, const int *rows, const int2 *kbe, const double *rhs,
               double *out) {
    const int nlev = (int)lvl.size() - 1;
    const int wide = 4 * TC_THREADS * TC_CLUSTER;
    auto level = [&](int l) {
        const int lo = lvl[l], hi = lvl[l + 1];
        k_tri_level<LOWER><<<cdiv(hi - lo, TPB), TPB, 0, h->stream>>>(lo, hi, rows, h->rowptr, h->col, h->diagPos, h->lu, rhs, out);
        h->launches++;
    };
    int l = 0;
    while (l < nlev) {
        int e = l;
        if (h->triMode == 1 && lvl_d && kbe)
            while (e < nlev && lvl[e + 1] - lvl[e] <= wide) e++;
        if (e - l >= 4) {
            k_tri_cluster<LOWER><<<TC_CLUSTER, TC_THREADS, 0, h->stream>>>(l, e, lvl_d, rows, kbe, h->col, h->lu, rhs, out);
            h->launches++;
            l = e;
        } else {
            level(l);
            l++;
        }
    }
}

// Same on groups of BS rows ("row_block").
template <bool LOWER, int BS, int CACHE>
void tri_solve_blk(mmadmm_pcg *h, const std::vector<int> &lvl, const int *lvl_d, const int *groups, const int2 *kbe, const double *rhs,
                   double *out) {
    const int nlev = (int)lvl.size() - 1;
    const int wide = 4 * TC_THREADS * TC_CLUSTER;
    const mmadmm_pcg::Chell &C = LOWER ? h->chellL : h->chellU;
    if (h->triMode == 2 && C.nChunks > 0) {   // the whole solve: sentinel fill + one persistent launch
        k_tri_sf_prepare<<<cdiv(h->n, TPB), TPB, 0, h->stream>>>(h->n, out, h->sfCounter);
        const int grid = std::min(std::max(1, h->sfGrid * h->sfGridX2 / 2), cdiv(C.nChunks, TC_THREADS / 32));   // chunks are claimed dynamically: CTAs that are not resident at once are harmless
        if (C.blocked) {
#ifdef MMADMM_TRI_PROF
            static const int profLoB = getenv("MMADMM_SF_PROF_LO") ? atoi(getenv("MMADMM_SF_PROF_LO")) : 0;
            static const int profU = getenv("MMADMM_SF_PROF_UPPER") ? 1 : 0;
            const int pl = (LOWER != (profU == 1)) ? profLoB : (1 << 30);
            static const int pmodeB = getenv("MMADMM_TRI_PROF_MODE") ? atoi(getenv("MMADMM_TRI_PROF_MODE")) : 0;
            if (BS == 2) k_tri_sf_block<LOWER, 2, 8><<<grid, TC_THREADS, 0, h->stream>>>(h->n, C.nChunks, C.meta_d, C.items, C.col, C.val, rhs, out, h->sfCounter, h->sfProxy, pl, pmodeB);
            else k_tri_sf_block<LOWER, 3, 8><<<grid, TC_THREADS, 0, h->stream>>>(h->n, C.nChunks, C.meta_d, C.items, C.col, C.val, rhs, out, h->sfCounter, h->sfProxy, pl, pmodeB);
#else
            if (BS == 2) k_tri_sf_block<LOWER, 2, 8><<<grid, TC_THREADS, 0, h->stream>>>(h->n, C.nChunks, C.meta_d, C.items, C.col, C.val, rhs, out, h->sfCounter, h->sfProxy);
            else k_tri_sf_block<LOWER, 3, 8><<<grid, TC_THREADS, 0, h->stream>>>(h->n, C.nChunks, C.meta_d, C.items, C.col, C.val, rhs, out, h->sfCounter, h->sfProxy);
#endif
            h->launches += 2;
            return;
        }
#ifdef MMADMM_TRI_PROF
        static const int profLo = getenv("MMADMM_SF_PROF_LO") ? atoi(getenv("MMADMM_SF_PROF_LO")) : 0;
        k_tri_syncfree<LOWER, BS, (BS * CACHE <= 32 ? CACHE : 32 / BS)><<<grid, TC_THREADS, 0, h->stream>>>(h->n, C.nChunks, C.meta_d, C.items, C.col, C.val, rhs, out,
                                                                                                        h->sfCounter, h->sfProxy, LOWER ? profLo : (1 << 30));
#else
        k_tri_syncfree<LOWER, BS, (BS * CACHE <= 32 ? CACHE : 32 / BS)><<<grid, TC_THREADS, 0, h->stream>>>(h->n, C.nChunks, C.meta_d, C.items, C.col, C.val, rhs, out,
                                                                                                        h->sfCounter, h->sfProxy);
#endif
        h->launches += 2;
        return;
    }
    const mmadmm_pcg::Lell &L = LOWER ? h->lellL : h->lellU;
    const bool useLell = h->triMode == 1 && h->triLayout == 1 && (int)L.meta.size() == nlev;
    int l = 0;
    while (l < nlev) {
        int e = l;
        if (useLell) {
            while (e < nlev && L.meta[e].z >= 0) e++;
            if (e - l >= 4) {
#ifdef MMADMM_TRI_PROF
                static const int pmode = getenv("MMADMM_TRI_PROF_MODE") ? atoi(getenv("MMADMM_TRI_PROF_MODE")) : 0;
                k_tri_cluster_lell<LOWER, BS, CACHE><<<TC_CLUSTER, TC_THREADS, 0, h->stream>>>(h->n, l, e, L.meta_d, groups, L.col, L.val, rhs, out, pmode);
#else
                k_tri_cluster_lell<LOWER, BS, CACHE><<<TC_CLUSTER, TC_THREADS, 0, h->stream>>>(h->n, l, e, L.meta_d, groups, L.col, L.val, rhs, out);
#endif
                h->launches++;
                l = e;
                continue;
            }
            e = l;
        }
        if (h->triMode == 1)
            while (e < nlev && lvl[e + 1] - lvl[e] <= wide) e++;
        if (e - l >= 4 && !useLell) {
            k_tri_cluster_blk<LOWER, BS, CACHE><<<TC_CLUSTER, TC_THREADS, 0, h->stream>>>(h->n, l, e, lvl_d, groups, kbe, h->rowptr, h->diagPos,
                                                                                          h->col, h->lu, rhs, out);
            h->launches++;
            l = e;
        } else {
            const int lo = lvl[l], hi = lvl[l + 1];
            k_tri_level_blk<LOWER, BS><<<cdiv(hi - lo, TPB), TPB, 0, h->stream>>>(h->n, lo, hi, groups, h->rowptr, h->col, h->diagPos, h->lu, rhs, out);
            h->launches++;
            l++;
        }
    }
}
template <int BS, int CACHE> void apply_ilu_blk(mmadmm_pcg *h, const double *r, double *z) {
    tri_solve_blk<true, BS, CACHE>(h, h->lvlGL, h->lvlGL_d, h->grpL, h->kbeGL, r, h->tmp);
    tri_solve_blk<false, BS, CACHE>(h, h->lvlGU, h->lvlGU_d, h->grpU, h->kbeGU, h->tmp, z);
}

// z = (LU)^{-1} r by level-scheduled forward and backward substitution
int apply_ilu(mmadmm_pcg *h, const double *r, double *z) {
    switch (h->grpL ? h->rowBlock : 0) {
        case 1: apply_ilu_blk<1, 14>(h, r, z); return MMADMM_OK;
        case 2: apply_ilu_blk<2, 14>(h, r, z); return MMADMM_OK;
        case 3: apply_ilu_blk<3, 10>(h, r, z); return MMADMM_OK;
        case 4: apply_ilu_blk<4, 8>(h, r, z); return MMADMM_OK;
        case 6: apply_ilu_blk<6, 5>(h, r, z); return MMADMM_OK;
        default: break;
    }
    { int rc = ensure_row_schedule(h); if (rc) return rc; }
    tri_solve<true>(h, h->lvlL, h->lvlL_d, h->rowsL, h->kbeL, r, h->tmp);
    tri_solve<false>(h, h->lvlU, h->lvlU_d, h->rowsU, h->kbeU, h->tmp, z);
    return MMADMM_OK;
}

}  // namespace

extern "C" {

int mmadmm_pcg_plan_levels(int n, const int *rowptr, const int *col, int bs, int *levL, int *levU, int *info3) {
    if (n <= 0 || !rowptr || !col || (bs != 1 && bs != 2 && bs != 3 && bs != 4 && bs != 6)) return MMADMM_EINVAL;
    std::vector<int> dg(n, -1);
    for (int i = 0; i < n; i++) {
        for (int k = rowptr[i]; k < rowptr[i + 1]; k++) {
            if (col[k] < 0 || col[k] >= n || (k > rowptr[i] && col[k] <= col[k - 1])) return MMADMM_EINVAL;
            if (col[k] == i) dg[i] = k;
        }
        if (dg[i] < 0) return MMADMM_EINVAL;
    }
    std::vector<int> lev;
    const int nl = plan_group_levels(n, bs, true, rowptr, col, dg.data(), lev);
    if (levL) std::copy(lev.begin(), lev.end(), levL);
    const int nu = plan_group_levels(n, bs, false, rowptr, col, dg.data(), lev);
    if (levU) std::copy(lev.begin(), lev.end(), levU);
    if (info3) { info3[0] = nl; info3[1] = nu; info3[2] = plan_blockable(n, bs, rowptr, col) ? 1 : 0; }
    return MMADMM_OK;
}

int mmadmm_pcg_create(int n, const int *rowptr, const int *col, const double *val, int precond, int device, mmadmm_pcg **out) {
    if (!out) return MMADMM_EINVAL;
    *out = nullptr;
    if (n <= 0 || !rowptr || !col || !val || precond < 0 || precond > 1) return MMADMM_EINVAL;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return MMADMM_ENODEVICE;
    mmadmm_pcg *h = new mmadmm_pcg();
    *out = h;
    SetupTimer tm("pcg_create");
    h->n = n; h->nnz = rowptr[n]; h->precond = precond; h->device = device;
    const int nnz = h->nnz;
    h->rowptr_h.assign(rowptr, rowptr + n + 1); h->col_h.assign(col, col + nnz);
    h->valHostStale = true;   // the host copy of the values (SELL build only) is fetched from the device when needed
    // pattern checks: sorted columns, diagonal present
    std::vector<int> diagPos(n, -1);
    for (int i = 0; i < n; i++) {
        for (int k = rowptr[i]; k < rowptr[i + 1]; k++) {
            if (col[k] < 0 || col[k] >= n || (k > rowptr[i] && col[k] <= col[k - 1])) {
                h->err = "CSR rows must have strictly ascending column indices in range";
                return MMADMM_EINVAL;
            }
            if (col[k] == i) diagPos[i] = k;
        }
        if (diagPos[i] < 0) { h->err = "every row needs a stored diagonal entry"; return MMADMM_EINVAL; }
        h->maxRowLen = std::max(h->maxRowLen, rowptr[i + 1] - rowptr[i]);
    }
    tm.lap("host copies + pattern check");
    const double avg = (double)nnz / n;
    h->lanes = avg <= 3 ? 1 : avg <= 6 ? 2 : avg <= 12 ? 4 : avg <= 24 ? 8 : avg <= 48 ? 16 : 32;
    PCK(cudaSetDevice(device));
    PCK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    auto up = [&](void **dst, const void *src, size_t bytes) -> cudaError_t {
        cudaError_t e = cudaMalloc(dst, std::max<size_t>(bytes, 8));
        if (e != cudaSuccess) return e;
        return src ? cudaMemcpyAsync(*dst, src, bytes, cudaMemcpyHostToDevice, h->stream) : cudaSuccess;
    };
    PCK(up((void **)&h->rowptr, rowptr, sizeof(int) * (size_t)(n + 1)));
    PCK(up((void **)&h->col, col, sizeof(int) * (size_t)nnz));
    PCK(up((void **)&h->val, val, sizeof(double) * (size_t)nnz));
    PCK(up((void **)&h->diagPos, diagPos.data(), sizeof(int) * (size_t)n));
    h->diagPos_h = diagPos;
    std::vector<double> invDiag(n);
    for (int i = 0; i < n; i++) invDiag[i] = 1.0 / val[diagPos[i]];   // Eigen's DiagonalPreconditioner stores 1/diag
    PCK(up((void **)&h->invDiag, invDiag.data(), sizeof(double) * (size_t)n));
    std::vector<int> rowsL, rowsU;
    if (precond == 1) {
        tm.lap("uploads queued");
        // level schedule: level(i) = 1 + max level of the rows it depends on
        std::vector<int> lev(n, 0);
        int nl = 0;
        for (int i = 0; i < n; i++) {
            int l = 0;
            for (int k = rowptr[i]; k < diagPos[i]; k++) l = std::max(l, lev[col[k]] + 1);
            lev[i] = l; nl = std::max(nl, l + 1);
        }
        auto bucket = [&](int nlev, std::vector<int> &rows, std::vector<int> &off) {
            off.assign(nlev + 1, 0);
            for (int i = 0; i < n; i++) off[lev[i] + 1]++;
            for (int l = 0; l < nlev; l++) off[l + 1] += off[l];
            rows.resize(n);
            std::vector<int> cur(off.begin(), off.end() - 1);
            for (int i = 0; i < n; i++) rows[cur[lev[i]]++] = i;
        };
        bucket(nl, rowsL, h->lvlL);
        PCK(up((void **)&h->lu, nullptr, sizeof(double) * (size_t)nnz));
        PCK(up((void **)&h->flag_d, nullptr, sizeof(int)));
        PCK(up((void **)&h->rowsL, rowsL.data(), sizeof(int) * (size_t)n));
        // the rest of the row-per-thread schedule (upper levels, row extents in level order, level offsets on the device) is only
        // read by tri_mode 0 / 1 with row_block 1: built on first use (ensure_row_schedule)
        tm.lap("level schedules + uploads");
        PCK(cudaStreamSynchronize(h->stream));   // rowsL is a local
        h->rowsL_h.swap(rowsL);
        { int rcf = ilu_factor_device(h); if (rcf) return rcf; }   // the numeric factor runs on the device, level by level
        tm.lap("device ILU(0) factor");
    }
    const size_t vb = sizeof(double) * (size_t)n;
    PCK(up((void **)&h->r, nullptr, vb)); PCK(up((void **)&h->z, nullptr, vb)); PCK(up((void **)&h->p, nullptr, vb));
    PCK(up((void **)&h->q, nullptr, vb)); PCK(up((void **)&h->x, nullptr, vb)); PCK(up((void **)&h->b, nullptr, vb));
    PCK(up((void **)&h->tmp, nullptr, vb));
    h->nblk = std::max(std::max(cdiv(n, TPB), spmv_parts(h)), cdiv((size_t)((n + 31) / 32) * 32, TPB));
    PCK(up((void **)&h->part0, nullptr, sizeof(double) * (size_t)h->nblk));
    PCK(up((void **)&h->part1, nullptr, sizeof(double) * (size_t)h->nblk));
    PCK(up((void **)&h->scal, nullptr, sizeof(double) * S_COUNT));
    PCK(cudaMallocHost(&h->scal_h, sizeof(double) * S_COUNT));
    PCK(cudaStreamSynchronize(h->stream));
    return MMADMM_OK;
}

int mmadmm_pcg_set_option(mmadmm_pcg *h, const char *name, int value) {
    if (!h || !name) return MMADMM_EINVAL;
    if (!std::strcmp(name, "tri_mode")) {
        if (value < 0 || value > 2) return MMADMM_EINVAL;
        const bool rebuild = (value == 2) != (h->triMode == 2) && h->precond == 1;
        h->triMode = value;
        if (rebuild) {   // the dataflow solve has its own copy of the factor (chunk-ELL)
            PCK(cudaSetDevice(h->device));
            PCK(cudaStreamSynchronize(h->stream));
            if (!h->sfGrid) {
                cudaDeviceProp pr;
                PCK(cudaGetDeviceProperties(&pr, h->device));
                h->sfGrid = pr.multiProcessorCount;   // 1 CTA of 256 threads per SM (register-bound)
            }
            int rc = build_groups(h, h->rowBlock);
            if (rc) return rc;
            PCK(cudaStreamSynchronize(h->stream));
            PCK(cudaGetLastError());
        }
        return MMADMM_OK;
    }
    if (!std::strcmp(name, "ilu_thread_kernel")) { h->iluThreadKernel = value ? 1 : 0; return MMADMM_OK; }
    if (!std::strcmp(name, "tri_proxy")) { h->sfProxy = value ? 1 : 0; return MMADMM_OK; }
    if (!std::strcmp(name, "tri_sf_ctas_per_sm_x2")) { if (value < 1 || value > 8) return MMADMM_EINVAL; h->sfGridX2 = value; return MMADMM_OK; }
    if (!std::strcmp(name, "tri_blocks")) {   // 1 (default): block form of the dataflow solve where the pattern is made of BS x BS blocks
        h->sfBlocks = value ? 1 : 0;
        if (h->triMode == 2 && h->precond == 1) {
            PCK(cudaSetDevice(h->device));
            PCK(cudaStreamSynchronize(h->stream));
            int rc = build_groups(h, h->rowBlock);
            if (rc) return rc;
            PCK(cudaStreamSynchronize(h->stream));
            PCK(cudaGetLastError());
        }
        return MMADMM_OK;
    }
    if (!std::strcmp(name, "tri_schedule")) {   // row_block | tri_layout << 8 in one build
        if (h->precond != 1) { h->err = "tri_schedule needs an ILU(0) handle"; return MMADMM_EINVAL; }
        PCK(cudaSetDevice(h->device));
        PCK(cudaStreamSynchronize(h->stream));
        h->triLayout = (value >> 8) & 1;
        if ((value >> 12) & 3) h->triMode = ((value >> 12) & 3) - 1;   // bits 12-13: tri_mode + 1 (0 = keep)
        if (h->triMode == 2 && !h->sfGrid) {
            cudaDeviceProp pr;
            PCK(cudaGetDeviceProperties(&pr, h->device));
            h->sfGrid = pr.multiProcessorCount;
        }
        int rc = build_groups(h, value & 0xff);
        if (rc) return rc;
        PCK(cudaStreamSynchronize(h->stream));
        PCK(cudaGetLastError());
        return MMADMM_OK;
    }
    if (!std::strcmp(name, "tri_layout")) {
        if (value != 0 && value != 1) return MMADMM_EINVAL;
        if (h->precond != 1) { h->err = "tri_layout needs an ILU(0) handle"; return MMADMM_EINVAL; }
        PCK(cudaSetDevice(h->device));
        PCK(cudaStreamSynchronize(h->stream));
        h->triLayout = value;
        int rc = build_groups(h, h->rowBlock);
        if (rc) return rc;
        PCK(cudaStreamSynchronize(h->stream));
        PCK(cudaGetLastError());
        return MMADMM_OK;
    }
    if (!std::strcmp(name, "row_block")) {
        PCK(cudaSetDevice(h->device));
        PCK(cudaStreamSynchronize(h->stream));
        { cudaError_t e0 = cudaGetLastError(); if (e0 != cudaSuccess) { h->err = std::string("row_block: error pending on entry: ") + cudaGetErrorString(e0); return MMADMM_ECUDA; } }
        int rc = build_groups(h, value);
        if (rc) return rc;
        { cudaError_t e1 = cudaGetLastError(); if (e1 != cudaSuccess) { h->err = std::string("row_block: error left by build_groups: ") + cudaGetErrorString(e1) + " / " + h->err; return MMADMM_ECUDA; } }
        return MMADMM_OK;
    }
    if (!std::strcmp(name, "spmv_format")) {   // 0: CSR (sub-warp per row), 1: SELL-32-4096
        if (value != 0 && value != 1) return MMADMM_EINVAL;
        if (value == 1) { int rc = build_sell(h); if (rc) return rc; PCK(cudaGetLastError()); }
        h->spmvFormat = value;
        return MMADMM_OK;
    }
    if (!std::strcmp(name, "sell_padding_permille")) return h->nnz ? (int)((h->sellEntries - h->nnz) * 1000 / h->nnz) : 0;   // query
    h->err = std::string("unknown option ") + name;
    return MMADMM_EINVAL;
}

int mmadmm_pcg_info(mmadmm_pcg *h, int *info4) {
    if (!h || !info4) return MMADMM_EINVAL;
    info4[0] = h->lanes;
    if (!h->grpL) { int rc = ensure_row_schedule(h); if (rc) return rc; }   // the upper row levels are built on first use
    const std::vector<int> &lL = h->grpL ? h->lvlGL : h->lvlL, &lU = h->grpL ? h->lvlGU : h->lvlU;
    info4[1] = (int)lL.size() > 0 ? (int)lL.size() - 1 : 0;
    info4[2] = (int)lU.size() > 0 ? (int)lU.size() - 1 : 0;
    info4[3] = (int)std::min<long long>(h->launches, 2147483647LL);
    return MMADMM_OK;
}

int mmadmm_pcg_spmv(mmadmm_pcg *h, const double *x, double *y) {
    if (!h || !x || !y) return MMADMM_EINVAL;
    PCK(cudaSetDevice(h->device));
    PCK(cudaMemcpyAsync(h->p, x, sizeof(double) * (size_t)h->n, cudaMemcpyHostToDevice, h->stream));
    launch_spmv(h, h->p, h->q, nullptr);
    PCK(cudaMemcpyAsync(y, h->q, sizeof(double) * (size_t)h->n, cudaMemcpyDeviceToHost, h->stream));
    PCK(cudaStreamSynchronize(h->stream));
    PCK(cudaGetLastError());
    return MMADMM_OK;
}

int mmadmm_pcg_precond(mmadmm_pcg *h, const double *r, double *z) {
    if (!h || !r || !z) return MMADMM_EINVAL;
    PCK(cudaSetDevice(h->device));
    PCK(cudaMemcpyAsync(h->r, r, sizeof(double) * (size_t)h->n, cudaMemcpyHostToDevice, h->stream));
    if (h->precond == 1) { int rca = apply_ilu(h, h->r, h->z); if (rca) return rca; }
    else { k_jacobi<<<cdiv(h->n, TPB), TPB, 0, h->stream>>>(h->n, h->r, h->invDiag, h->z); h->launches++; }
    PCK(cudaMemcpyAsync(z, h->z, sizeof(double) * (size_t)h->n, cudaMemcpyDeviceToHost, h->stream));
    PCK(cudaStreamSynchronize(h->stream));
    PCK(cudaGetLastError());
    return MMADMM_OK;
}

int mmadmm_pcg_get_ilu(mmadmm_pcg *h, double *lu) {
    if (!h || !lu || h->precond != 1) return MMADMM_EINVAL;
    PCK(cudaMemcpy(lu, h->lu, sizeof(double) * (size_t)h->nnz, cudaMemcpyDeviceToHost));
    return MMADMM_OK;
}

// Device-resident solve: b_dev -> x_dev (both n doubles on h's device), zero initial guess like Eigen's solve().
// Stops when ||r||_2 <= tol ||b||_2 (Eigen::ConjugateGradient's criterion) or after maxit iterations.
int mmadmm_pcg_refactor_device(mmadmm_pcg *h, const double *val_dev) {
    if (!h || !val_dev) return MMADMM_EINVAL;
    PCK(cudaSetDevice(h->device));
    PCK(cudaMemcpyAsync(h->val, val_dev, sizeof(double) * (size_t)h->nnz, cudaMemcpyDeviceToDevice, h->stream));
    k_inv_diag<<<cdiv(h->n, TPB), TPB, 0, h->stream>>>(h->n, h->diagPos, h->val, h->invDiag);
    h->launches++;
    if (h->slicePtr) {   // the SELL copy holds values too: rebuild it from the new ones (host-side build, rare)
        h->valHostStale = true;   // build_sell fetches the new values
        PCK(cudaStreamSynchronize(h->stream));
        cudaFree(h->slicePtr); cudaFree(h->scol); cudaFree(h->sval); cudaFree(h->perm);
        h->slicePtr = h->scol = h->perm = nullptr; h->sval = nullptr;
        int rc = build_sell(h);
        if (rc) return rc;
    } else {
        h->valHostStale = true;
    }
    if (h->precond == 1) return ilu_factor_device(h);
    PCK(cudaStreamSynchronize(h->stream));
    return MMADMM_OK;
}

int mmadmm_pcg_solve_device(mmadmm_pcg *h, const double *b_dev, double *x_dev, double tol, int maxit, int *iters, double *relres) {
    if (!h || !b_dev || !x_dev) return MMADMM_EINVAL;
    PCK(cudaSetDevice(h->device));
    const int n = h->n, g = cdiv(n, TPB);
    cudaStream_t st = h->stream;
    const size_t vb = sizeof(double) * (size_t)n;
    PCK(cudaMemsetAsync(x_dev, 0, vb, st));
    PCK(cudaMemcpyAsync(h->r, b_dev, vb, cudaMemcpyDeviceToDevice, st));   // r = b - A*0
    // z = M^{-1} r ; rz, rr(=bb)
    if (h->precond == 1) { int rca = apply_ilu(h, h->r, h->z); if (rca) return rca; }
    else k_jacobi<<<g, TPB, 0, st>>>(n, h->r, h->invDiag, h->z);
    k_dot2<<<g, TPB, 0, st>>>(n, h->r, h->z, h->r, h->part0, h->part1);
    k_scalars<<<1, 1024, 0, st>>>(g, h->part0, h->part1, h->scal, 2);
    k_scalars<<<1, 1024, 0, st>>>(g, nullptr, h->part1, h->scal, 3);
    PCK(cudaMemcpyAsync(h->p, h->z, vb, cudaMemcpyDeviceToDevice, st));
    PCK(cudaMemcpyAsync(h->scal_h, h->scal, sizeof(double) * S_COUNT, cudaMemcpyDeviceToHost, st));
    PCK(cudaStreamSynchronize(st));
    h->launches += 4;
    const double bb = h->scal_h[S_BB];
    int it = 0;
    double rr = h->scal_h[S_RR];
    if (bb == 0.0) { if (iters) *iters = 0; if (relres) *relres = 0.0; return MMADMM_OK; }
    const double thresh = tol * tol * bb;
    while (rr > thresh && it < maxit) {
        launch_spmv(h, h->p, h->q, h->part0);                                   // q = A p, partials of p.q
        k_scalars<<<1, 1024, 0, st>>>(spmv_parts(h), h->part0, nullptr, h->scal, 0);   // alpha = rz / pAp
        if (h->precond == 1) {
            k_cg_update<false><<<g, TPB, 0, st>>>(n, h->scal, h->p, h->q, h->invDiag, x_dev, h->r, h->z, h->part0, h->part1);
            { int rca = apply_ilu(h, h->r, h->z); if (rca) return rca; }
            k_dot2<<<g, TPB, 0, st>>>(n, h->r, h->z, h->r, h->part0, h->part1);
            h->launches++;
        } else {
            k_cg_update<true><<<g, TPB, 0, st>>>(n, h->scal, h->p, h->q, h->invDiag, x_dev, h->r, h->z, h->part0, h->part1);
        }
        k_scalars<<<1, 1024, 0, st>>>(g, h->part0, h->part1, h->scal, 1);       // beta, rz, rr
        k_xpay<<<g, TPB, 0, st>>>(n, h->scal, h->z, h->p);
        PCK(cudaMemcpyAsync(h->scal_h, h->scal, sizeof(double) * S_COUNT, cudaMemcpyDeviceToHost, st));
        PCK(cudaStreamSynchronize(st));
        h->launches += 4;
        rr = h->scal_h[S_RR];
        it++;
        if (!(rr == rr)) { h->err = "non-finite residual in PCG (matrix not SPD?)"; return MMADMM_ESINGULAR; }
    }
    PCK(cudaGetLastError());
    if (iters) *iters = it;
    if (relres) *relres = std::sqrt(rr / bb);
    return MMADMM_OK;
}

// BiCGSTAB as the reference's LASolver runs it for Method 2 (lib/LASolver/MatrixIter.cpp:693-800 driver, accel_class.cpp:229-273
// start values, :280-399 one iteration; configured at src/Mesh.cpp:266-304: iaccel = 0, new_rhat = 0, toler = 0): right-
// preconditioned by the ILU(0) factors, rhat = r0 = b (zero initial guess), alpha = rholst = omega = 1 and p = A vbar = 0 at
// the start, `tiny` = 1e-300 guards in beta and omega, stop when ||r|| / ||b|| < resid_reduc.  *iters = iterations done, or
// -1 when nitmax iterations did not converge (LASolver's convention).  The four dot products per iteration are reduced on
// the device in a fixed order and read back (8-byte copies): the scalar recurrences run on the host.
int mmadmm_pcg_bicgstab_device(mmadmm_pcg *h, const double *b_dev, double *x_dev, double resid_reduc, int nitmax, int *iters,
                               double *relres) {
    if (!h || !b_dev || !x_dev) return MMADMM_EINVAL;
    if (h->precond != 1) { h->err = "BiCGSTAB needs the ILU(0) preconditioner (precond = 1)"; return MMADMM_EINVAL; }
    PCK(cudaSetDevice(h->device));
    const int n = h->n, g = cdiv(n, TPB);
    cudaStream_t st = h->stream;
    const size_t vb = sizeof(double) * (size_t)n;
    for (int k = 0; k < 4; k++)
        if (!h->bi[k]) PCK(cudaMalloc(&h->bi[k], vb));
    double *res = h->r, *res0 = h->bi[0], *pvec = h->p, *vbar = h->z, *avbar = h->q, *svec = h->bi[1], *zvec = h->bi[2], *tvec = h->bi[3];
    // two dot products a.b and c.c -> host
    auto dots = [&](const double *a, const double *bb, const double *c, double *ab, double *cc) -> cudaError_t {
        k_dot2<<<g, TPB, 0, st>>>(n, a, bb, c, h->part0, h->part1);
        k_scalars<<<1, 1024, 0, st>>>(g, h->part0, h->part1, h->scal, 2);
        h->launches += 2;
        cudaError_t e = cudaMemcpyAsync(h->scal_h, h->scal, sizeof(double) * S_COUNT, cudaMemcpyDeviceToHost, st);
        if (e != cudaSuccess) return e;
        e = cudaStreamSynchronize(st);
        *ab = h->scal_h[S_RZ]; *cc = h->scal_h[S_RR];
        return e;
    };
    PCK(cudaMemsetAsync(x_dev, 0, vb, st));
    PCK(cudaMemcpyAsync(res, b_dev, vb, cudaMemcpyDeviceToDevice, st));
    PCK(cudaMemcpyAsync(res0, b_dev, vb, cudaMemcpyDeviceToDevice, st));
    PCK(cudaMemsetAsync(pvec, 0, vb, st));
    PCK(cudaMemsetAsync(avbar, 0, vb, st));
    double rho = 0.0, bb2 = 0.0;
    PCK(dots(res0, res, res, &rho, &bb2));
    const double rmsi = std::sqrt(bb2);
    if (rmsi == 0.0) { if (iters) *iters = 0; if (relres) *relres = 0.0; return MMADMM_OK; }
    const double tiny = 1.e-300;
    double alpha = 1.0, rholst = 1.0, omega = 1.0, rms = rmsi;
    int nitr = 0;
    bool conv = false;
    for (int it = 1; it <= nitmax; it++) {
        nitr++;
        if (it > 1) { double dummy; PCK(dots(res0, res, res, &rho, &dummy)); }
        double beta = rho / (rholst + tiny);
        beta *= (alpha / (omega + tiny));
        rholst = rho;
        k_bi_p<<<g, TPB, 0, st>>>(n, beta, omega, res, avbar, pvec);
        { int rca = apply_ilu(h, pvec, vbar); if (rca) return rca; }
        launch_spmv(h, vbar, avbar, nullptr);
        double r0av, dummy2;
        PCK(dots(res0, avbar, avbar, &r0av, &dummy2));
        alpha = rho / r0av;
        k_bi_s<<<g, TPB, 0, st>>>(n, alpha, res, avbar, svec);
        { int rca = apply_ilu(h, svec, zvec); if (rca) return rca; }
        launch_spmv(h, zvec, tvec, nullptr);
        double ts, tt;
        PCK(dots(tvec, svec, tvec, &ts, &tt));
        omega = ts / (tt + tiny);
        k_bi_xr<<<g, TPB, 0, st>>>(n, alpha, omega, vbar, zvec, svec, tvec, x_dev, res, h->part1);
        k_scalars<<<1, 1024, 0, st>>>(g, nullptr, h->part1, h->scal, 3);
        h->launches += 4;
        PCK(cudaMemcpyAsync(h->scal_h, h->scal, sizeof(double) * S_COUNT, cudaMemcpyDeviceToHost, st));
        PCK(cudaStreamSynchronize(st));
        rms = std::sqrt(h->scal_h[S_BB]);
        if (!(rms == rms)) { h->err = "non-finite residual in BiCGSTAB (breakdown)"; return MMADMM_ESINGULAR; }
        if (rms / rmsi < resid_reduc) { conv = true; break; }
    }
    PCK(cudaGetLastError());
    if (iters) *iters = conv ? nitr : -1;
    if (relres) *relres = rms / rmsi;
    return MMADMM_OK;
}

// host-buffer form of the above
int mmadmm_pcg_bicgstab(mmadmm_pcg *h, const double *b, double *x, double resid_reduc, int nitmax, int *iters, double *relres) {
    if (!h || !b || !x) return MMADMM_EINVAL;
    PCK(cudaSetDevice(h->device));
    PCK(cudaMemcpyAsync(h->b, b, sizeof(double) * (size_t)h->n, cudaMemcpyHostToDevice, h->stream));
    int rc = mmadmm_pcg_bicgstab_device(h, h->b, h->x, resid_reduc, nitmax, iters, relres);
    if (rc) return rc;
    PCK(cudaMemcpyAsync(x, h->x, sizeof(double) * (size_t)h->n, cudaMemcpyDeviceToHost, h->stream));
    PCK(cudaStreamSynchronize(h->stream));
    return MMADMM_OK;
}

int mmadmm_pcg_solve(mmadmm_pcg *h, const double *b, double *x, double tol, int maxit, int *iters, double *relres) {
    if (!h || !b || !x) return MMADMM_EINVAL;
    PCK(cudaSetDevice(h->device));
    PCK(cudaMemcpyAsync(h->b, b, sizeof(double) * (size_t)h->n, cudaMemcpyHostToDevice, h->stream));
    int rc = mmadmm_pcg_solve_device(h, h->b, h->x, tol, maxit, iters, relres);
    if (rc) return rc;
    PCK(cudaMemcpyAsync(x, h->x, sizeof(double) * (size_t)h->n, cudaMemcpyDeviceToHost, h->stream));
    PCK(cudaStreamSynchronize(h->stream));
    return MMADMM_OK;
}

// average device time (ms) of one SpMV over `reps` back-to-back launches (CUDA events on the solver's stream)
int mmadmm_pcg_time_spmv(mmadmm_pcg *h, int reps, double *ms) {
    if (!h || reps <= 0 || !ms) return MMADMM_EINVAL;
    PCK(cudaSetDevice(h->device));
    cudaEvent_t e0, e1;
    PCK(cudaEventCreate(&e0)); PCK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; i++) launch_spmv(h, h->p, h->q, h->part0);
    PCK(cudaEventRecord(e0, h->stream));
    for (int i = 0; i < reps; i++) launch_spmv(h, h->p, h->q, h->part0);
    PCK(cudaEventRecord(e1, h->stream));
    PCK(cudaEventSynchronize(e1));
    float t = 0;
    PCK(cudaEventElapsedTime(&t, e0, e1));
    *ms = t / reps;
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return MMADMM_OK;
}

// Measured FP64 FMA throughput of the device (TFLOP/s, 2 flops per DFMA): the denominator of the prox kernel's FP64-pipe
// figure (MEASURED_PEAKS.json has no FP64 entry; SURVEY.md 8d asks for a DFMA microbenchmark on the box).
int mmadmm_pcg_time_precond(mmadmm_pcg *h, int reps, double *ms) {
    if (!h || !ms || reps <= 0) return MMADMM_EINVAL;
    PCK(cudaSetDevice(h->device));
    cudaEvent_t e0, e1;
    PCK(cudaEventCreate(&e0)); PCK(cudaEventCreate(&e1));
    PCK(cudaMemsetAsync(h->r, 0, sizeof(double) * (size_t)h->n, h->stream));
    auto apply = [&]() -> int {
        if (h->precond == 1) return apply_ilu(h, h->r, h->z);
        k_jacobi<<<cdiv(h->n, TPB), TPB, 0, h->stream>>>(h->n, h->r, h->invDiag, h->z);
        h->launches++;
        return MMADMM_OK;
    };
    { int rca = apply(); if (rca) return rca; }
    PCK(cudaEventRecord(e0, h->stream));
    for (int k = 0; k < reps; k++) { int rca = apply(); if (rca) return rca; }
    PCK(cudaEventRecord(e1, h->stream));
    PCK(cudaStreamSynchronize(h->stream));
    PCK(cudaGetLastError());
    float t = 0.f;
    PCK(cudaEventElapsedTime(&t, e0, e1));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *ms = (double)t / reps;
    return MMADMM_OK;
}

#ifdef MMADMM_TRI_PROF
// timing experiments: group-level offsets of the lower (which = 0) / upper (1) schedule; returns the number of entries
int mmadmm_pcg_debug_levels(mmadmm_pcg *h, int which, int *out, int cap) {
    const std::vector<int> &v = which ? h->lvlGU : h->lvlGL;
    for (int i = 0; i < (int)v.size() && i < cap; i++) out[i] = v[i];
    return (int)v.size();
}
// timing experiments: per-chunk time stamps of the last LOWER dataflow solve + the chunk's first item's level is the caller's business
int mmadmm_pcg_debug_sf_spins(int *out65536) { return cudaMemcpyFromSymbol(out65536, g_sf_spins, sizeof(int) * 65536) == cudaSuccess ? 0 : 2; }
int mmadmm_pcg_debug_sf_prof(unsigned long long *out5x65536) {
    return cudaMemcpyFromSymbol(out5x65536, g_sf_prof, sizeof(unsigned long long) * 5 * 65536) == cudaSuccess ? 0 : 2;
}
#endif
int mmadmm_measure_fp64_peak(int device, double *tflops) {
    if (!tflops) return MMADMM_EINVAL;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return MMADMM_ENODEVICE;
    if (cudaSetDevice(device) != cudaSuccess) return MMADMM_ECUDA;
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    double *out = nullptr;
    if (cudaMalloc(&out, 8) != cudaSuccess) return MMADMM_ECUDA;
    const int iters = 1 << 13, grid = sms * 8, tpb = 256;   // ~1 ms per launch
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_dfma_peak<<<grid, tpb>>>(iters, 1.0, out);
    double best = 0.0;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        k_dfma_peak<<<grid, tpb>>>(iters, 1.0 + rep, out);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return MMADMM_ECUDA; }
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double tf = 2.0 * 8.0 * (double)iters * grid * tpb / (ms * 1e-3) / 1e12;
        best = tf > best ? tf : best;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(out);
    *tflops = best;
    return MMADMM_OK;
}

const char *mmadmm_pcg_last_error(mmadmm_pcg *h) { return h ? h->err.c_str() : ""; }

void mmadmm_pcg_destroy(mmadmm_pcg *h) {
    if (!h) return;
    cudaSetDevice(h->device);
    cudaFree(h->rowptr); cudaFree(h->col); cudaFree(h->diagPos); cudaFree(h->val); cudaFree(h->lu); cudaFree(h->invDiag);
    cudaFree(h->rowsL); cudaFree(h->rowsU); cudaFree(h->lvlL_d); cudaFree(h->lvlU_d); cudaFree(h->kbeL); cudaFree(h->kbeU);
    free_groups(h);
    cudaFree(h->flag_d); cudaFree(h->sfCounter);
    cudaFree(h->r); cudaFree(h->z); cudaFree(h->p); cudaFree(h->q); cudaFree(h->x); cudaFree(h->b); cudaFree(h->tmp);
    cudaFree(h->part0); cudaFree(h->part1); cudaFree(h->scal);
    for (int k = 0; k < 4; k++) cudaFree(h->bi[k]);
    cudaFree(h->slicePtr); cudaFree(h->scol); cudaFree(h->sval); cudaFree(h->perm);
    if (h->scal_h) cudaFreeHost(h->scal_h);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

}  // extern "C"
