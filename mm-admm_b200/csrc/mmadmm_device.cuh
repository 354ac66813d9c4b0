// Device-side FP64 math of the MM-ADMM hot path for sm_100a.
//
//   block_grad<D>      AdaptationFunctional<D>::blockGrad      (reference src/AdaptationFunctional.cpp:102-287)
//   eval_monitor<D>    MeshInterpolator<D>::evalMonitorOnGrid  (reference src/MeshInterpolator.cpp:286-342)
//   find_cell          utils::findLimInfMeshPoint              (reference src/MeshUtils.h:45-54)
//
// Everything is fixed-size and fully unrolled so it lives in registers; the pow() calls of the reference
// (exponents 1.5, 0.5, -0.5, 2.25, 1.25 only, because p = 3/2 and theta = 1/3 are hard-coded there,
// AdaptationFunctional.cpp:210-211) are sqrt chains.  No tensor cores: the matrices are 2x2 / 3x3.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mmadmm {

// Element state (z, u, inverse Hessians) is stored TILE-MAJOR: [tile of 32 elements][row][32], so everything one warp
// needs of a tile is one contiguous block (z, u: n*256 bytes; B: n*n*256 bytes) -- one TMA bulk copy each way, whole
// DRAM pages, and all rows of an element sit within a few KB for the vertex-centric gathers of the x-update.
constexpr int TILE = 32;
__host__ __device__ __forceinline__ size_t tm_index(size_t e, int row, int rows) {
    return ((e >> 5) * (size_t)rows + (size_t)row) * TILE + (e & 31);
}

// z and u (n = D(D+1) doubles per element, vertex-major inside an element): [tile of 32 elements][vertex v][32][D] -- a tile is
// one contiguous n*256-byte block (one bulk copy), and the D components of a slot (element, vertex) are adjacent, so the
// vertex-centric gathers of the x-update fetch a slot with one 16-byte load (2D) / from one or two sectors (3D) instead of D
// loads from D different 256-byte rows.  k = D*v + l.
template <int D> __host__ __device__ __forceinline__ size_t zu_index(size_t e, int k) {
    return (((e >> 5) * (D + 1) + (size_t)(k / D)) * TILE + (e & 31)) * D + (k % D);
}
// offset of component k inside a tile image, relative to the lane's base (lane * D)
template <int D> __host__ __device__ constexpr int zu_off(int k) { return (k / D) * TILE * D + (k % D); }

// Inverse Hessians (n x n per element).  2D: tile-major like z / u ([tile of 32][n*n entries][32]).  3D: sub-tiles of 8 elements,
// [sub-tile][144 entries][8], entry (row, col) at position row*12 + (col ^ ((row/3) & 1)): the 3D prox kernel gives every element
// four lanes (one per vertex) that read different rows (or columns) of the same sub-tile at the same time, and this pairwise swap
// of the columns in the rows of odd vertices spreads those four groups over both halves of the shared-memory banks (a sub-tile is
// one contiguous 9216-byte block: one TMA bulk copy each way).
constexpr int SUB3 = 8;
__host__ __device__ __forceinline__ int b3_pos(int row, int col) { return row * 12 + (col ^ ((row / 3) & 1)); }
template <int D> __host__ __device__ __forceinline__ size_t b_index(size_t e, int row, int col) {
    if (D == 2) return tm_index(e, row * 6 + col, 36);
    return ((e >> 3) * 144 + (size_t)b3_pos(row, col)) * SUB3 + (e & 7);
}

enum NodeTypeDev { NT_BOUNDARY_FREE = 0, NT_BOUNDARY_FIXED = 1, NT_INTERIOR = 2 };  // src/NodeType.h:4-8
enum StatusBits { ST_INVERTED = 1, ST_SINGULAR_HESS = 2, ST_BFGS_MAX = 4, ST_NONFINITE = 8, ST_PEER_TIMEOUT = 16 };

// Monitor grid (read-only after init).  Rows are grid points, D*D doubles each (3D rows are padded to 10 doubles so
// that every row is five aligned 16-byte loads), row index zInd*(nx+1)*(ny+1) + yInd*(nx+1) + xInd exactly as the
// reference's lookup reads it.
template <int D> struct GridRow { static constexpr int stride = (D == 2) ? 4 : 10; };
struct MonGrid {
    const double *vals;
    const double *iso;           // non-null when every grid tensor is m*I (checked bitwise at init): the (g+1)^D scalars m.
                                 // Interpolating m and forming m*I is bit-identical to interpolating the D*D entries and
                                 // reads D*D (2D) / 10 (3D) times fewer bytes per lookup.
    const double *ax, *ay, *az;  // axis coordinates, nx+1 / ny+1 / nz+1 entries (utils::linspace values)
    int nx, ny, nz;
    double o[3];                 // axis origin a[0]
    double rh[3];                // 1 / (a[1] - a[0]): the cell guess multiplies instead of dividing (the bilinear
                                 // interpolant is continuous across cells, so a one-ulp different guess on a grid line
                                 // changes the result at rounding level only)
};

template <int D> struct SMat {
    double a[D * D];
    __device__ __forceinline__ double &operator()(int i, int j) { return a[i * D + j]; }
    __device__ __forceinline__ double operator()(int i, int j) const { return a[i * D + j]; }
};

template <int D> __device__ __forceinline__ double sdet(const SMat<D> &m);
template <> __device__ __forceinline__ double sdet<2>(const SMat<2> &m) { return m(0, 0) * m(1, 1) - m(1, 0) * m(0, 1); }
template <> __device__ __forceinline__ double sdet<3>(const SMat<3> &m) {
    return m(0, 0) * (m(1, 1) * m(2, 2) - m(1, 2) * m(2, 1)) - m(0, 1) * (m(1, 0) * m(2, 2) - m(1, 2) * m(2, 0)) +
           m(0, 2) * (m(1, 0) * m(2, 1) - m(1, 1) * m(2, 0));
}
template <int D> __device__ __forceinline__ SMat<D> sinv(const SMat<D> &m);
template <> __device__ __forceinline__ SMat<2> sinv<2>(const SMat<2> &m) {
    SMat<2> r;
    double invdet = 1.0 / sdet<2>(m);
    r(0, 0) = m(1, 1) * invdet;
    r(1, 0) = -m(1, 0) * invdet;
    r(0, 1) = -m(0, 1) * invdet;
    r(1, 1) = m(0, 0) * invdet;
    return r;
}
template <> __device__ __forceinline__ SMat<3> sinv<3>(const SMat<3> &m) {
    SMat<3> r;
    // cofactor form (what a fixed-size 3x3 inverse does): r(j,i) = cof(i,j)/det
    double c00 = m(1, 1) * m(2, 2) - m(1, 2) * m(2, 1);
    double c10 = m(2, 1) * m(0, 2) - m(2, 2) * m(0, 1);
    double c20 = m(0, 1) * m(1, 2) - m(0, 2) * m(1, 1);
    double invdet = 1.0 / (c00 * m(0, 0) + c10 * m(1, 0) + c20 * m(2, 0));
    r(0, 0) = c00 * invdet; r(0, 1) = c10 * invdet; r(0, 2) = c20 * invdet;
    r(1, 0) = (m(1, 2) * m(2, 0) - m(1, 0) * m(2, 2)) * invdet;
    r(1, 1) = (m(2, 2) * m(0, 0) - m(2, 0) * m(0, 2)) * invdet;
    r(1, 2) = (m(0, 2) * m(1, 0) - m(0, 0) * m(1, 2)) * invdet;
    r(2, 0) = (m(1, 0) * m(2, 1) - m(1, 1) * m(2, 0)) * invdet;
    r(2, 1) = (m(2, 0) * m(0, 1) - m(2, 1) * m(0, 0)) * invdet;
    r(2, 2) = (m(0, 0) * m(1, 1) - m(0, 1) * m(1, 0)) * invdet;
    return r;
}
// m^{-1} / s with a single division (adjugate times 1/(det*s))
template <int D> __device__ __forceinline__ SMat<D> sinv_scaled(const SMat<D> &m, double s);
template <> __device__ __forceinline__ SMat<2> sinv_scaled<2>(const SMat<2> &m, double s) {
    SMat<2> r;
    double invdet = 1.0 / (sdet<2>(m) * s);
    r(0, 0) = m(1, 1) * invdet;
    r(1, 0) = -m(1, 0) * invdet;
    r(0, 1) = -m(0, 1) * invdet;
    r(1, 1) = m(0, 0) * invdet;
    return r;
}
template <> __device__ __forceinline__ SMat<3> sinv_scaled<3>(const SMat<3> &m, double s) {
    SMat<3> r;
    double c00 = m(1, 1) * m(2, 2) - m(1, 2) * m(2, 1);
    double c10 = m(2, 1) * m(0, 2) - m(2, 2) * m(0, 1);
    double c20 = m(0, 1) * m(1, 2) - m(0, 2) * m(1, 1);
    double invdet = 1.0 / ((c00 * m(0, 0) + c10 * m(1, 0) + c20 * m(2, 0)) * s);
    r(0, 0) = c00 * invdet; r(0, 1) = c10 * invdet; r(0, 2) = c20 * invdet;
    r(1, 0) = (m(1, 2) * m(2, 0) - m(1, 0) * m(2, 2)) * invdet;
    r(1, 1) = (m(2, 2) * m(0, 0) - m(2, 0) * m(0, 2)) * invdet;
    r(1, 2) = (m(0, 2) * m(1, 0) - m(0, 0) * m(1, 2)) * invdet;
    r(2, 0) = (m(1, 0) * m(2, 1) - m(1, 1) * m(2, 0)) * invdet;
    r(2, 1) = (m(2, 0) * m(0, 1) - m(2, 1) * m(0, 0)) * invdet;
    r(2, 2) = (m(0, 0) * m(1, 1) - m(0, 1) * m(1, 0)) * invdet;
    return r;
}
template <int D> __device__ __forceinline__ SMat<D> smul(const SMat<D> &A, const SMat<D> &B) {
    SMat<D> C;
#pragma unroll
    for (int i = 0; i < D; i++)
#pragma unroll
        for (int j = 0; j < D; j++) {
            double s = A(i, 0) * B(0, j);
#pragma unroll
            for (int k = 1; k < D; k++) s += A(i, k) * B(k, j);
            C(i, j) = s;
        }
    return C;
}
// A * B^T
template <int D> __device__ __forceinline__ SMat<D> smul_nt(const SMat<D> &A, const SMat<D> &B) {
    SMat<D> C;
#pragma unroll
    for (int i = 0; i < D; i++)
#pragma unroll
        for (int j = 0; j < D; j++) {
            double s = A(i, 0) * B(j, 0);
#pragma unroll
            for (int k = 1; k < D; k++) s += A(i, k) * B(j, k);
            C(i, j) = s;
        }
    return C;
}
// A^T * B
template <int D> __device__ __forceinline__ SMat<D> smul_tn(const SMat<D> &A, const SMat<D> &B) {
    SMat<D> C;
#pragma unroll
    for (int i = 0; i < D; i++)
#pragma unroll
        for (int j = 0; j < D; j++) {
            double s = A(0, i) * B(0, j);
#pragma unroll
            for (int k = 1; k < D; k++) s += A(k, i) * B(k, j);
            C(i, j) = s;
        }
    return C;
}

// t^(d*p/2) and t^(d*p/2 - 1) with p = 3/2:  D=2 -> t^1.5, t^0.5 ; D=3 -> t^2.25, t^1.25
template <int D> __device__ __forceinline__ void pow_pair(double t, double &p_full, double &p_m1) {
    if (D == 2) {
        double s = sqrt(t);
        p_m1 = s;
        p_full = t * s;
    } else {
        double q = sqrt(sqrt(t));
        p_m1 = t * q;
        p_full = t * p_m1;
    }
}

// utils::findLimInfMeshPoint: uint32 wrap for negatives => clamp to the last cell (size-2 = nIntervals-1)
__device__ __forceinline__ int find_cell(double w, double a0, double rh, int nIntervals) {
    uint32_t guess = (uint32_t)(int)((w - a0) * rh);
    if (guess > (uint32_t)(nIntervals - 1)) guess = (uint32_t)(nIntervals - 1);
    return (int)guess;
}

template <int D> __device__ __forceinline__ void eval_monitor(const MonGrid &g, const double *pnt, SMat<D> &mVal);

template <> __device__ __forceinline__ void eval_monitor<2>(const MonGrid &g, const double *pnt, SMat<2> &mVal) {
    int xInd = find_cell(pnt[0], g.o[0], g.rh[0], g.nx);
    int yInd = find_cell(pnt[1], g.o[1], g.rh[1], g.ny);
    double x0 = __ldg(g.ax + xInd), x1 = __ldg(g.ax + xInd + 1);
    double y0 = __ldg(g.ay + yInd), y1 = __ldg(g.ay + yInd + 1);
    double x = pnt[0], y = pnt[1];
    double norm = (1 / ((x1 - x0) * (y1 - y0)));
    double c0 = norm * (x1 - x) * (y1 - y);
    double c1 = norm * (x - x0) * (y1 - y);
    double c2 = norm * (x1 - x) * (y - y0);
    double c3 = norm * (x - x0) * (y - y0);
    // two adjacent grid points are 64 contiguous bytes (2 x 4 doubles): four 16-byte loads per grid row
    const double2 *r0 = reinterpret_cast<const double2 *>(g.vals) + 2 * ((size_t)yInd * (g.nx + 1) + xInd);
    const double2 *r1 = r0 + 2 * (size_t)(g.nx + 1);
    double2 g00a = __ldg(r0), g00b = __ldg(r0 + 1), g10a = __ldg(r0 + 2), g10b = __ldg(r0 + 3);
    double2 g01a = __ldg(r1), g01b = __ldg(r1 + 1), g11a = __ldg(r1 + 2), g11b = __ldg(r1 + 3);
    mVal.a[0] = c0 * g00a.x + c1 * g10a.x + c2 * g01a.x + c3 * g11a.x;
    mVal.a[1] = c0 * g00a.y + c1 * g10a.y + c2 * g01a.y + c3 * g11a.y;
    mVal.a[2] = c0 * g00b.x + c1 * g10b.x + c2 * g01b.x + c3 * g11b.x;
    mVal.a[3] = c0 * g00b.y + c1 * g10b.y + c2 * g01b.y + c3 * g11b.y;
}

template <> __device__ __forceinline__ void eval_monitor<3>(const MonGrid &g, const double *pnt, SMat<3> &mVal) {
    int xInd = find_cell(pnt[0], g.o[0], g.rh[0], g.nx);
    int yInd = find_cell(pnt[1], g.o[1], g.rh[1], g.ny);
    int zInd = find_cell(pnt[2], g.o[2], g.rh[2], g.nz);
    double x0 = __ldg(g.ax + xInd), x1 = __ldg(g.ax + xInd + 1);
    double y0 = __ldg(g.ay + yInd), y1 = __ldg(g.ay + yInd + 1);
    double z0 = __ldg(g.az + zInd), z1 = __ldg(g.az + zInd + 1);
    // utils::triLinearInterpolation (src/MeshUtils.h:65-80)
    double xd = (pnt[0] - x0) / (x1 - x0);
    double yd = (pnt[1] - y0) / (y1 - y0);
    double zd = (pnt[2] - z0) / (z1 - z0);
    double c[8];
    c[0] = (1 - xd) * (1 - yd) * (1 - zd);
    c[1] = xd * (1 - yd) * (1 - zd);
    c[2] = (1 - xd) * yd * (1 - zd);
    c[3] = xd * yd * (1 - zd);
    c[4] = (1 - xd) * (1 - yd) * zd;
    c[5] = xd * (1 - yd) * zd;
    c[6] = (1 - xd) * yd * zd;
    c[7] = xd * yd * zd;
    const size_t sl = (size_t)(g.nx + 1) * (g.ny + 1);
    const size_t base = (size_t)zInd * sl + (size_t)yInd * (g.nx + 1) + xInd;
#pragma unroll
    for (int k = 0; k < 9; k++) mVal.a[k] = 0.0;
#pragma unroll
    for (int cc = 0; cc < 8; cc++) {
        size_t row = base + ((cc & 1) ? 1 : 0) + ((cc & 2) ? (size_t)(g.nx + 1) : 0) + ((cc & 4) ? sl : 0);
        const double2 *r = reinterpret_cast<const double2 *>(g.vals + row * GridRow<3>::stride);
        const double2 v0 = __ldg(r), v1 = __ldg(r + 1), v2 = __ldg(r + 2), v3 = __ldg(r + 3), v4 = __ldg(r + 4);
        mVal.a[0] += c[cc] * v0.x; mVal.a[1] += c[cc] * v0.y; mVal.a[2] += c[cc] * v1.x; mVal.a[3] += c[cc] * v1.y;
        mVal.a[4] += c[cc] * v2.x; mVal.a[5] += c[cc] * v2.y; mVal.a[6] += c[cc] * v3.x; mVal.a[7] += c[cc] * v3.y;
        mVal.a[8] += c[cc] * v4.x;
    }
}

// Isotropic grids (every tensor m*I): the scalar m by the same bi/tri-linear expressions, term order included, so that
// m*I equals what eval_monitor<D> returns bit for bit.
template <int D> __device__ __forceinline__ double eval_monitor_iso(const MonGrid &g, const double *pnt);
template <> __device__ __forceinline__ double eval_monitor_iso<2>(const MonGrid &g, const double *pnt) {
    int xInd = find_cell(pnt[0], g.o[0], g.rh[0], g.nx);
    int yInd = find_cell(pnt[1], g.o[1], g.rh[1], g.ny);
    double x0 = __ldg(g.ax + xInd), x1 = __ldg(g.ax + xInd + 1);
    double y0 = __ldg(g.ay + yInd), y1 = __ldg(g.ay + yInd + 1);
    double x = pnt[0], y = pnt[1];
    double norm = (1 / ((x1 - x0) * (y1 - y0)));
    double c0 = norm * (x1 - x) * (y1 - y);
    double c1 = norm * (x - x0) * (y1 - y);
    double c2 = norm * (x1 - x) * (y - y0);
    double c3 = norm * (x - x0) * (y - y0);
    const double *r0 = g.iso + (size_t)yInd * (g.nx + 1) + xInd;
    const double *r1 = r0 + (g.nx + 1);
    const double m00 = __ldg(r0), m10 = __ldg(r0 + 1), m01 = __ldg(r1), m11 = __ldg(r1 + 1);
    return c0 * m00 + c1 * m10 + c2 * m01 + c3 * m11;
}
template <> __device__ __forceinline__ double eval_monitor_iso<3>(const MonGrid &g, const double *pnt) {
    int xInd = find_cell(pnt[0], g.o[0], g.rh[0], g.nx);
    int yInd = find_cell(pnt[1], g.o[1], g.rh[1], g.ny);
    int zInd = find_cell(pnt[2], g.o[2], g.rh[2], g.nz);
    double x0 = __ldg(g.ax + xInd), x1 = __ldg(g.ax + xInd + 1);
    double y0 = __ldg(g.ay + yInd), y1 = __ldg(g.ay + yInd + 1);
    double z0 = __ldg(g.az + zInd), z1 = __ldg(g.az + zInd + 1);
    double xd = (pnt[0] - x0) / (x1 - x0);
    double yd = (pnt[1] - y0) / (y1 - y0);
    double zd = (pnt[2] - z0) / (z1 - z0);
    const size_t sl = (size_t)(g.nx + 1) * (g.ny + 1);
    const double *b = g.iso + (size_t)zInd * sl + (size_t)yInd * (g.nx + 1) + xInd;
    const double v0 = __ldg(b), v1 = __ldg(b + 1), v2 = __ldg(b + (g.nx + 1)), v3 = __ldg(b + (g.nx + 1) + 1);
    const double v4 = __ldg(b + sl), v5 = __ldg(b + sl + 1), v6 = __ldg(b + sl + (g.nx + 1)), v7 = __ldg(b + sl + (g.nx + 1) + 1);
    double m = 0.0;
    m += ((1 - xd) * (1 - yd) * (1 - zd)) * v0;
    m += (xd * (1 - yd) * (1 - zd)) * v1;
    m += ((1 - xd) * yd * (1 - zd)) * v2;
    m += (xd * yd * (1 - zd)) * v3;
    m += ((1 - xd) * (1 - yd) * zd) * v4;
    m += (xd * (1 - yd) * zd) * v5;
    m += ((1 - xd) * yd * zd) * v6;
    m += (xd * yd * zd) * v7;
    return m;
}
// run-time selection (unit tests / ABI lookups only; the hot kernels are compiled per path)
template <int D> __device__ __forceinline__ void eval_monitor_any(const MonGrid &g, const double *pnt, SMat<D> &mVal) {
    if (g.iso) {
        const double m = eval_monitor_iso<D>(g, pnt);
#pragma unroll
        for (int k = 0; k < D * D; k++) mVal.a[k] = (k % (D + 1) == 0) ? m : 0.0;
    } else {
        eval_monitor<D>(g, pnt, mVal);
    }
}

// L2 prefetch of the grid rows eval_monitor(pnt) will read
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
template <int D, bool ISO = false> __device__ __forceinline__ void prefetch_monitor(const MonGrid &g, const double *pnt) {
    int xInd = find_cell(pnt[0], g.o[0], g.rh[0], g.nx);
    int yInd = find_cell(pnt[1], g.o[1], g.rh[1], g.ny);
    if (ISO) {
        const size_t sl = (size_t)(g.nx + 1) * (g.ny + 1);
        const double *r0 = g.iso + (size_t)yInd * (g.nx + 1) + xInd + ((D == 3) ? (size_t)find_cell(pnt[2], g.o[2], g.rh[2], g.nz) * sl : 0);
        prefetch_l2(r0); prefetch_l2(r0 + 1);
        prefetch_l2(r0 + (g.nx + 1)); prefetch_l2(r0 + (g.nx + 1) + 1);
        if (D == 3) { prefetch_l2(r0 + sl); prefetch_l2(r0 + sl + 1); prefetch_l2(r0 + sl + (g.nx + 1)); prefetch_l2(r0 + sl + (g.nx + 1) + 1); }
        return;
    }
    if (D == 2) {
        const double *r0 = g.vals + 4 * ((size_t)yInd * (g.nx + 1) + xInd);
        prefetch_l2(r0);
        prefetch_l2(r0 + 7);
        prefetch_l2(r0 + 4 * (size_t)(g.nx + 1));
        prefetch_l2(r0 + 4 * (size_t)(g.nx + 1) + 7);
    } else {
        int zInd = find_cell(pnt[2], g.o[2], g.rh[2], g.nz);
        const size_t sl = (size_t)(g.nx + 1) * (g.ny + 1);
        const double *r0 = g.vals + GridRow<3>::stride * ((size_t)zInd * sl + (size_t)yInd * (g.nx + 1) + xInd);
#pragma unroll
        for (int cc = 0; cc < 4; cc++) {
            const double *r = r0 + GridRow<3>::stride * (((cc & 1) ? (size_t)(g.nx + 1) : 0) + ((cc & 2) ? sl : 0));
            prefetch_l2(r);
            prefetch_l2(r + 2 * GridRow<3>::stride - 1);
        }
    }
}

// Boundary sliding (OPTION, off by default): Mesh::projectOntoBoundary -> projection2D / projection3D (src/Mesh.cpp:118-241) on the
// face lists of Mesh::buildFaceList (:61-112).  All three call sites are commented out in the reference (:636-642, :975-984,
// :1020-1026); option "boundary_projection" re-enables the two on the ADMM path.  Face geometry comes from `vp`, the node
// positions at the end of the previous outer step (the reference's *Vp).
struct FaceArgs {
    const int *ptr = nullptr, *idx = nullptr;    // vertex -> boundary faces (ascending face id, the reference's std::set order)
    const int *list = nullptr;                   // nFaces x D vertex ids
    const double *vp = nullptr;                  // P x D
    int mask = 0;                                // bit 0: after every element's BFGS solve in prox; bit 1: in updateAfterStep
};
__device__ __forceinline__ int sgn_d(double v) { return (0.0 < v) - (v < 0.0); }
template <int D> __device__ __noinline__ void project_onto_boundary(const FaceArgs &fa, int nodeId, double *nodeIn);
template <> __device__ __noinline__ void project_onto_boundary<2>(const FaceArgs &fa, int nodeId, double *nodeIn) {
    const double node[2] = {nodeIn[0], nodeIn[1]};
    double best[2] = {node[0], node[1]};
    double minDist = INFINITY;
    for (int q = fa.ptr[nodeId]; q < fa.ptr[nodeId + 1]; q++) {
        const int *pid = fa.list + (size_t)fa.idx[q] * 2;
        const double x1[2] = {fa.vp[(size_t)pid[0] * 2], fa.vp[(size_t)pid[0] * 2 + 1]};
        const double x2[2] = {fa.vp[(size_t)pid[1] * 2], fa.vp[(size_t)pid[1] * 2 + 1]};
        const double u[2] = {x2[0] - x1[0], x2[1] - x1[1]};
        const double w[2] = {node[0] - x1[0], node[1] - x1[1]};
        const double c = __ddiv_rn(__dadd_rn(__dmul_rn(u[0], w[0]), __dmul_rn(u[1], w[1])), __dadd_rn(__dmul_rn(u[0], u[0]), __dmul_rn(u[1], u[1])));
        const double proj[2] = {c * u[0], c * u[1]};
        const double d0 = proj[0] - w[0], d1 = proj[1] - w[1];
        double dist = sqrt(__dadd_rn(__dmul_rn(d0, d0), __dmul_rn(d1, d1)));
        const double t = __ddiv_rn(sqrt(__dadd_rn(__dmul_rn(proj[0], proj[0]), __dmul_rn(proj[1], proj[1]))),
                                   sqrt(__dadd_rn(__dmul_rn(u[0], u[0]), __dmul_rn(u[1], u[1]))));
        const bool same = sgn_d(u[0]) == sgn_d(proj[0]) && sgn_d(u[1]) == sgn_d(proj[1]);
        if (dist < minDist && same && t > 0.0 && t < 1.0) {
            minDist = dist;
            best[0] = __dadd_rn(__dmul_rn(1.0 - t, x1[0]), __dmul_rn(t, x2[0]));
            best[1] = __dadd_rn(__dmul_rn(1.0 - t, x1[1]), __dmul_rn(t, x2[1]));
        } else if (!same) {
            const double e0 = x1[0] - node[0], e1 = x1[1] - node[1];
            dist = sqrt(__dadd_rn(__dmul_rn(e0, e0), __dmul_rn(e1, e1)));
            if (dist < minDist) { minDist = dist; best[0] = x1[0]; best[1] = x1[1]; }
        } else if (t > 1.0) {
            const double e0 = x2[0] - node[0], e1 = x2[1] - node[1];
            dist = sqrt(__dadd_rn(__dmul_rn(e0, e0), __dmul_rn(e1, e1)));
            if (dist < minDist) { minDist = dist; best[0] = x2[0]; best[1] = x2[1]; }
        }
    }
    nodeIn[0] = best[0]; nodeIn[1] = best[1];
}
template <> __device__ __noinline__ void project_onto_boundary<3>(const FaceArgs &fa, int nodeId, double *nodeIn) {
    const double node[3] = {nodeIn[0], nodeIn[1], nodeIn[2]};
    double best[3] = {0.0, 0.0, 0.0};
    bool projWorked = false;
    const double CHECK_EPS = 1e-10;
    double minDist = INFINITY;
    // explicit non-fused arithmetic: which face wins (dist < minDist, b >= 1e-10) is a comparison of nearly equal numbers
    auto cross = [](const double *a, const double *b, double *c) {
        c[0] = __dsub_rn(__dmul_rn(a[1], b[2]), __dmul_rn(a[2], b[1]));
        c[1] = __dsub_rn(__dmul_rn(a[2], b[0]), __dmul_rn(a[0], b[2]));
        c[2] = __dsub_rn(__dmul_rn(a[0], b[1]), __dmul_rn(a[1], b[0]));
    };
    auto dot = [](const double *a, const double *b) { return __dadd_rn(__dadd_rn(__dmul_rn(a[0], b[0]), __dmul_rn(a[1], b[1])), __dmul_rn(a[2], b[2])); };
    for (int qi = fa.ptr[nodeId]; qi < fa.ptr[nodeId + 1]; qi++) {
        const int *pid = fa.list + (size_t)fa.idx[qi] * 3;
        const double *p0 = fa.vp + (size_t)pid[0] * 3, *p1 = fa.vp + (size_t)pid[1] * 3, *p2 = fa.vp + (size_t)pid[2] * 3;
        double u[3], v[3], w[3], n[3], uw[3], wv[3], b[3];
#pragma unroll
        for (int l = 0; l < 3; l++) { u[l] = p1[l] - p0[l]; v[l] = p2[l] - p0[l]; w[l] = node[l] - p0[l]; }
        cross(u, v, n);
        const double temp = __ddiv_rn(1.0, dot(n, n));
        cross(u, w, uw);
        cross(w, v, wv);
        b[2] = __dmul_rn(dot(uw, n), temp);
        b[1] = __dmul_rn(dot(wv, n), temp);
        b[0] = __dsub_rn(__dsub_rn(1.0, b[1]), b[2]);
        double np[3];
#pragma unroll
        for (int l = 0; l < 3; l++) np[l] = __dadd_rn(__dadd_rn(__dmul_rn(b[0], p0[l]), __dmul_rn(b[1], p1[l])), __dmul_rn(b[2], p2[l]));
        const double f0 = np[0] - node[0], f1 = np[1] - node[1], f2 = np[2] - node[2];
        const double dist = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(f0, f0), __dmul_rn(f1, f1)), __dmul_rn(f2, f2)));
        if (dist < minDist && b[0] >= CHECK_EPS && b[1] >= CHECK_EPS && b[2] >= CHECK_EPS) {
            minDist = dist;
            best[0] = np[0]; best[1] = np[1]; best[2] = np[2];
            projWorked = true;
        }
    }
    if (projWorked) { nodeIn[0] = best[0]; nodeIn[1] = best[1]; nodeIn[2] = best[2]; }
}

// Constants of the functional that the reference recomputes per call (AdaptationFunctional.cpp:176-201,
// 219-220): computed once on the host with the reference's own expressions (libm pow) and passed by value.
template <int D> struct FuncConsts {
    double Ehat[D * D];  // reference simplex scaled to volume 1/N (used when !compMesh)
    double dpow;         // pow(d, d*p/2)
    double w2;           // w*w (ADMM penalty)
};

// Energy G*|K| and (optionally) the unregularised element gradient.  z: n = D(D+1) doubles, vertex-major.
// Returns absK*G.  `inverted` is set when det(E) <= 0 (the reference asserts there, :174).
// Everything of blockGrad after the D+1 monitor lookups: mPre[i] (general) or mIso[i] (ISO: tensor i is mIso[i]*I) are the
// monitor values at the D+1 vertices.  Split out so that a kernel may obtain the lookups elsewhere (the 3D prox kernel
// evaluates one vertex per warp and exchanges them through shared memory).
template <int D, bool GRAD, bool ISO>
__device__ __forceinline__ double block_grad_core(const double *z, const double *xi, bool compMesh, const SMat<D> *mPre,
                                                  const double *mIso, const FuncConsts<D> &fc, double *grad, bool &inverted) {
    constexpr int n = D * (D + 1);
    const double dFact = (D == 2) ? 2.0 : 6.0;
    const double d = (double)D;
    const double p = 1.5;
    const double theta = 1.0 / 3.0;
    SMat<D> M;
#pragma unroll
    for (int k = 0; k < D * D; k++) M.a[k] = 0.0;
#pragma unroll
    for (int i = 0; i < D + 1; i++) {
        if (ISO) {
#pragma unroll
            for (int k = 0; k < D; k++) M.a[k * (D + 1)] += mIso[i];
        } else {
#pragma unroll
            for (int k = 0; k < D * D; k++) M.a[k] += mPre[i].a[k];
        }
    }
    SMat<D> E, Ehat;
#pragma unroll
    for (int j = 0; j < D; j++)
#pragma unroll
        for (int l = 0; l < D; l++) {
            E(l, j) = z[D * (j + 1) + l] - z[l];
            Ehat(l, j) = compMesh ? (xi[D * (j + 1) + l] - xi[l]) : fc.Ehat[l * D + j];
        }
    double Edet = sdet<D>(E);
    if (!(Edet > 0)) inverted = true;

    SMat<D> Einv = sinv<D>(E);
    SMat<D> FJ = smul<D>(Ehat, Einv);
    double detFJ = sdet<D>(FJ);

    // Monitor-dependent pieces.  General: Minv = M^{-1}/(D+1) (sic, :157), MinvJt = Minv FJ^T, tr(FJ Minv FJ^T), det.
    // ISO: M = ms I, so Minv = mu I with mu = 1/(ms (D+1)), tr = mu |FJ|_F^2, det(Minv) = mu^D -- the same quantities
    // without the 3x3 inverse and four of the 3x3 products (rounding-level differences only).
    SMat<D> Minv, MinvJt;
    double mu = 0.0, fro = 0.0, trJMJt, detM;
    if (ISO) {
        mu = 1.0 / (M.a[0] * ((double)D + 1));
#pragma unroll
        for (int k = 0; k < D * D; k++) fro = (k == 0) ? FJ.a[0] * FJ.a[0] : fro + FJ.a[k] * FJ.a[k];
        trJMJt = mu * fro;
        detM = rsqrt((D == 2) ? mu * mu : mu * mu * mu);
    } else {
        Minv = sinv_scaled<D>(M, (double)D + 1);
        MinvJt = smul_nt<D>(Minv, FJ);  // Minv * FJ^T
        SMat<D> JMJt = smul<D>(FJ, MinvJt);
        trJMJt = JMJt(0, 0);
#pragma unroll
        for (int i = 1; i < D; i++) trJMJt += JMJt(i, i);
        detM = rsqrt(sdet<D>(Minv));   // sqrt(1/det(Minv))
    }

    double trp, trpm1;
    pow_pair<D>(trJMJt, trp, trpm1);
    double q = detFJ / detM;
    double G = theta * detM * trp + (1.0 - 2.0 * theta) * fc.dpow * detM * (q * sqrt(q));
    double absK = fabs(Edet / dFact);
    if (!GRAD) return absK * G;

    double sJ = d * p * theta * detM * trpm1;
    double rsDetM = rsqrt(detM);        // detM^(1-p)
    double sDetFJ = sqrt(detFJ);        // detFJ^(p-1)
    double dGddet = p * (1.0 - 2.0 * theta) * fc.dpow * rsDetM * sDetFJ;
    const double s1 = -0.5 * theta * d * p * detM * trpm1;
    const double s2 = (0.5 * theta * detM * trp + ((0.5 - theta) * (1.0 - p) * fc.dpow) * rsDetM * (detFJ * sDetFJ));
    double basisComb[D];
#pragma unroll
    for (int l = 0; l < D; l++) basisComb[l] = 0.0;
    SMat<D> T;     // Einv * dGdJ * FJ with dGdJ = sJ Minv FJ^T
    if (ISO) {
        // dGdM = s1 mu^2 FJ^T FJ + s2 mu I and mPre[j+1] - mPre[0] = (m_{j+1} - m_0) I: only tr(dGdM) is needed
        const double trdGdM = s1 * (mu * mu) * fro + (double)D * (s2 * mu);
#pragma unroll
        for (int j = 0; j < D; j++) {
            const double tr = trdGdM * (mIso[j + 1] - mIso[0]);
#pragma unroll
            for (int l = 0; l < D; l++) basisComb[l] += Einv(j, l) * tr;
        }
        SMat<D> ES = smul<D>(Einv, smul_tn<D>(FJ, FJ));
        const double sJm = sJ * mu;
#pragma unroll
        for (int k = 0; k < D * D; k++) T.a[k] = sJm * ES.a[k];
    } else {
        SMat<D> dGdJ;
#pragma unroll
        for (int k = 0; k < D * D; k++) dGdJ.a[k] = sJ * MinvJt.a[k];
        // Minv^T * FJ^T * FJ * Minv; reference order ((Minv^T FJ^T) FJ) Minv, reassociated (rounding-level) to reuse FJ*Minv
        SMat<D> A = smul_tn<D>(Minv, smul_tn<D>(FJ, smul<D>(FJ, Minv)));
        SMat<D> dGdM;
#pragma unroll
        for (int k = 0; k < D * D; k++) dGdM.a[k] = s1 * A.a[k] + s2 * Minv.a[k];
#pragma unroll
        for (int j = 0; j < D; j++) {
            // trace(dGdM * (mPre[j+1] - mPre[0]))
            double tr = 0.0;
#pragma unroll
            for (int a = 0; a < D; a++)
#pragma unroll
                for (int b = 0; b < D; b++) tr += dGdM(a, b) * (mPre[j + 1](b, a) - mPre[0](b, a));
#pragma unroll
            for (int l = 0; l < D; l++) basisComb[l] += Einv(j, l) * tr;
        }
        T = smul<D>(smul<D>(Einv, dGdJ), FJ);
    }
    double c1 = (-G + dGddet * detFJ);
    SMat<D> vLoc;
#pragma unroll
    for (int r = 0; r < D; r++)
#pragma unroll
        for (int l = 0; l < D; l++) vLoc(r, l) = (c1 * Einv(r, l) + T(r, l)) - basisComb[l] * (1.0 / ((double)D + 1.0));
    double g0[D];
#pragma unroll
    for (int l = 0; l < D; l++) {
        double s = 0.0;
#pragma unroll
        for (int r = 0; r < D; r++) s += vLoc(r, l);
        g0[l] = s + basisComb[l];
    }
#pragma unroll
    for (int l = 0; l < D; l++) grad[l] = g0[l] * absK;
#pragma unroll
    for (int nn = 1; nn < D + 1; nn++)
#pragma unroll
        for (int l = 0; l < D; l++) grad[D * nn + l] = (-vLoc(nn - 1, l)) * absK;
    (void)n;
    return absK * G;
}

// Isotropic blockGrad for ONE vertex q of the element (k_prox_vw: warp q owns vertex q).  Same quantities as
// block_grad_core<D, true, true>, but the gradient block of vertex q needs only one row vector a of Einv:
//   q >= 1: grad_q = -vLoc(q-1, :) |K|,  vLoc(r, :) = c1 Einv(r, :) + sJ mu Einv(r, :) (FJ^T FJ) - bc/(D+1)       (a = Einv(q-1, :))
//   q == 0: grad_0 = (sum_r vLoc(r, :) + bc) |K|                                                            (a = sum_r Einv(r, :))
// so the 3x3 products Einv*(FJ^T FJ), the scaled T and the full vLoc are replaced by one vector-matrix product.  For q >= 1 the
// arithmetic is the reference's operation for operation; for q == 0 the row sum is taken before the product (rounding level).
template <int D>
__device__ __forceinline__ double block_grad_own_iso(const double *z, const double *xi, bool compMesh, const double *mIso,
                                                     const FuncConsts<D> &fc, int q, double *g, bool &inverted) {
    const double dFact = (D == 2) ? 2.0 : 6.0;
    const double d = (double)D;
    const double p = 1.5;
    const double theta = 1.0 / 3.0;
    double ms = mIso[0];
#pragma unroll
    for (int i = 1; i < D + 1; i++) ms += mIso[i];
    SMat<D> E, Ehat;
#pragma unroll
    for (int j = 0; j < D; j++)
#pragma unroll
        for (int l = 0; l < D; l++) {
            E(l, j) = z[D * (j + 1) + l] - z[l];
            Ehat(l, j) = compMesh ? (xi[D * (j + 1) + l] - xi[l]) : fc.Ehat[l * D + j];
        }
    const double Edet = sdet<D>(E);
    if (!(Edet > 0)) inverted = true;
    const SMat<D> Einv = sinv<D>(E);
    const SMat<D> FJ = smul<D>(Ehat, Einv);
    const double detFJ = sdet<D>(FJ);
    const double mu = 1.0 / (ms * ((double)D + 1));
    double fro = 0.0;
#pragma unroll
    for (int k = 0; k < D * D; k++) fro = (k == 0) ? FJ.a[0] * FJ.a[0] : fro + FJ.a[k] * FJ.a[k];
    const double trJMJt = mu * fro;
    const double detM = rsqrt((D == 2) ? mu * mu : mu * mu * mu);
    double trp, trpm1;
    pow_pair<D>(trJMJt, trp, trpm1);
    const double qq = detFJ / detM;
    const double G = theta * detM * trp + (1.0 - 2.0 * theta) * fc.dpow * detM * (qq * sqrt(qq));
    const double absK = fabs(Edet / dFact);
    const double sJ = d * p * theta * detM * trpm1;
    const double rsDetM = rsqrt(detM);
    const double sDetFJ = sqrt(detFJ);
    const double dGddet = p * (1.0 - 2.0 * theta) * fc.dpow * rsDetM * sDetFJ;
    const double s1 = -0.5 * theta * d * p * detM * trpm1;
    const double s2 = (0.5 * theta * detM * trp + ((0.5 - theta) * (1.0 - p) * fc.dpow) * rsDetM * (detFJ * sDetFJ));
    const double trdGdM = s1 * (mu * mu) * fro + (double)D * (s2 * mu);
    double bc[D];
#pragma unroll
    for (int l = 0; l < D; l++) bc[l] = 0.0;
#pragma unroll
    for (int j = 0; j < D; j++) {
        const double tr = trdGdM * (mIso[j + 1] - mIso[0]);
#pragma unroll
        for (int l = 0; l < D; l++) bc[l] += Einv(j, l) * tr;
    }
    double a[D];
#pragma unroll
    for (int k = 0; k < D; k++) {
        double sum = Einv(0, k);
        double row = Einv(0, k);
#pragma unroll
        for (int r = 1; r < D; r++) { sum += Einv(r, k); row = (q == r + 1) ? Einv(r, k) : row; }
        a[k] = (q == 0) ? sum : row;
    }
    const SMat<D> S = smul_tn<D>(FJ, FJ);
    const double sJm = sJ * mu;
    const double c1 = (-G + dGddet * detFJ);
    const double inv = 1.0 / ((double)D + 1.0);
#pragma unroll
    for (int l = 0; l < D; l++) {
        double es = a[0] * S(0, l);
#pragma unroll
        for (int k = 1; k < D; k++) es += a[k] * S(k, l);
        const double v = (c1 * a[l] + sJm * es);
        g[l] = ((q == 0) ? ((v - d * (bc[l] * inv)) + bc[l]) : -(v - bc[l] * inv)) * absK;
    }
    return absK * G;
}

// Energy G*|K| and (optionally) the unregularised element gradient.  z: n = D(D+1) doubles, vertex-major.
// Returns absK*G.  `inverted` is set when det(E) <= 0 (the reference asserts there, :174).
// ISO (compile-time): every grid tensor is m*I, so the D+1 vertex tensors are D+1 scalars (4 instead of 36 live doubles
// in 3D) and M = (sum m_i) I; the arithmetic downstream is the general one on that M.
template <int D, bool GRAD, bool ISO = false>
__device__ __forceinline__ double block_grad(const double *z, const double *xi, bool compMesh, const MonGrid &mg,
                                             const FuncConsts<D> &fc, double *grad, bool &inverted) {
    SMat<D> mPre[ISO ? 1 : D + 1];
    double mIso[ISO ? D + 1 : 1];
#pragma unroll
    for (int i = 0; i < D + 1; i++) {
        if (ISO) mIso[i] = eval_monitor_iso<D>(mg, z + i * D);
        else eval_monitor<D>(mg, z + i * D, mPre[i]);
    }
    return block_grad_core<D, GRAD, ISO>(z, xi, compMesh, mPre, mIso, fc, grad, inverted);
}

// Regularised + masked gradient as Mesh::computeBlockGrad(.., regularize=true) returns it
// (src/Mesh.cpp:755-772 on top of AdaptationFunctional.cpp:279-282).  Returns the UNregularised energy (Igt).
template <int D, bool ISO = false>
__device__ __forceinline__ double prox_grad(const double *z, const double *dxpu, const double *xi, bool compMesh,
                                            const MonGrid &mg, const FuncConsts<D> &fc, unsigned fixedBits, double *grad,
                                            bool &inverted) {
    constexpr int n = D * (D + 1);
    double Igt = block_grad<D, true, ISO>(z, xi, compMesh, mg, fc, grad, inverted);
#pragma unroll
    for (int k = 0; k < n; k++) grad[k] += fc.w2 * (-dxpu[k] + z[k]);
#pragma unroll
    for (int v = 0; v < D + 1; v++)
        if (fixedBits & (1u << v)) {
#pragma unroll
            for (int l = 0; l < D; l++) grad[D * v + l] = 0.0;
        }
    return Igt;
}

// Same as prox_grad, with DXpU read from shared memory (tile image in the zu_index layout, dxs = tile + lane*D) after the gradient is formed,
// so that it is not live in registers across block_grad.
template <int D, bool ISO = false>
__device__ __forceinline__ double prox_grad_s(const double *z, const double *dxs, const double *xi, bool compMesh,
                                              const MonGrid &mg, const FuncConsts<D> &fc, unsigned fixedBits, double *grad,
                                              bool &inverted) {
    constexpr int n = D * (D + 1);
    double Igt = block_grad<D, true, ISO>(z, xi, compMesh, mg, fc, grad, inverted);
#pragma unroll
    for (int k = 0; k < n; k++) grad[k] += fc.w2 * (-dxs[zu_off<D>(k)] + z[k]);
#pragma unroll
    for (int v = 0; v < D + 1; v++)
        if (fixedBits & (1u << v)) {
#pragma unroll
            for (int l = 0; l < D; l++) grad[D * v + l] = 0.0;
        }
    return Igt;
}

// ------------------------------------------------------------------------------------------------
// TMA (bulk async copy) + mbarrier helpers: the inverse-Hessian tile of a CTA is staged through shared memory
// with cp.async.bulk (SASS UBLKCP) so that 79 % (2D) / 88 % (3D) of the bytes of the prox kernel never pass
// through the register file or the LSU, and the load overlaps the first gradient evaluation.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, unsigned parity) {
    unsigned ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a protocol error must surface as a CUDA error, never as a hung GPU.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
    for (int i = 0; i < (1 << 26); i++)
        if (mbar_try_wait(bar, parity)) return;
    __trap();
}
// global -> shared, completion counted in bytes on `bar`; bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, unsigned bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// shared -> global (bulk async group)
__device__ __forceinline__ void bulk_s2g(void *dst_gmem, const void *src_smem, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// 2D tensor-map TMA: one instruction moves a whole [planes][32 elements] tile (SASS UTMALDG / UTMASTG)
__device__ __forceinline__ void tma_load_2d(void *dst_smem, const void *tmap, int c0, int c1, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tma_store_2d(const void *tmap, int c0, int c1, const void *src_smem) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];" ::"l"(tmap), "r"(c0), "r"(c1),
                 "r"(smem_u32(src_smem))
                 : "memory");
}
// per-thread asynchronous gather global -> shared (LDGSTS): no registers are held while the data is in flight
__device__ __forceinline__ void cp_async16(void *dst_smem, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(void *dst_smem, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(void *dst_smem, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
// make generic-proxy shared-memory writes visible to the async proxy (before a bulk store reads them)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------
// NVLink peer-memory exchange: flags are monotonically increasing epochs written with system scope after the data; readers
// wait on them with a TIME bound (ranks enter a step at independent host times: a peer whose host is busy for a while must
// not cost its neighbours their CUDA context).  A wait that runs out sets ST_PEER_TIMEOUT in the status word and returns;
// every later wait of the same run returns at once, and the host reports MMADMM_EPEER on every rank.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void flag_release(unsigned long long *flag, unsigned long long v) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ bool flag_wait(const unsigned long long *flag, unsigned long long v, unsigned long long timeoutNs, unsigned *status) {
    unsigned long long cur;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(cur) : "l"(flag) : "memory");
    if (cur >= v) return true;
    if (*(volatile unsigned *)status & ST_PEER_TIMEOUT) return false;
    const unsigned long long t0 = global_timer_ns();
    for (unsigned spin = 1;; spin++) {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(cur) : "l"(flag) : "memory");
        if (cur >= v) return true;
        __nanosleep(spin < 64 ? 32 : 256);
        if ((spin & 255u) == 0 && global_timer_ns() - t0 > timeoutNs) break;
    }
    atomicOr(status, ST_PEER_TIMEOUT);
    return false;
}
// What a kernel needs to store its partial sums of the shared vertices straight into the neighbours' mailboxes and, as the
// last CTA of the grid to finish, to publish the epoch (fused form of the halo push).  list == nullptr: nothing to push.
struct HaloPush {
    const int *ptr = nullptr;                    // [nShared+1] offsets into nbr / pos
    const int *nbr = nullptr;                    // neighbour index q of the entry
    const int *pos = nullptr;                    // position inside that neighbour's receive segment for this rank
    double *const *peerRecv = nullptr;           // [nNbr] receive segment (parity 0) in the neighbour's mailbox
    const size_t *peerStride = nullptr;          // [nNbr] doubles between the two parity halves there
    unsigned long long *const *peerFlag = nullptr;   // [nNbr] the neighbour's flag slot for this rank
    unsigned *ticket = nullptr;                  // CTA completion counter (self-resetting)
    int nNbr = 0;
    unsigned long long epoch = 0;
};
template <int D> __device__ __forceinline__ void halo_push_store(const HaloPush &hp, int sh, const double *s) {
    for (int k = hp.ptr[sh]; k < hp.ptr[sh + 1]; k++) {
        const int q = hp.nbr[k];
        double *dst = hp.peerRecv[q] + ((hp.epoch & 1ull) ? hp.peerStride[q] : 0) + (size_t)hp.pos[k] * D;
#pragma unroll
        for (int l = 0; l < D; l++) dst[l] = s[l];
    }
}
// Call with the whole CTA, after every thread's halo_push_store: the last CTA of the grid publishes the epoch to all neighbours.
__device__ __forceinline__ void halo_push_publish(const HaloPush &hp) {
    __shared__ unsigned lastFlag;
    __threadfence_system();      // this thread's peer stores are visible system-wide before its CTA takes a ticket
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned t = atomicAdd(hp.ticket, 1u);
        lastFlag = (t == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (lastFlag) {
        __threadfence();
        if ((int)threadIdx.x < hp.nNbr) flag_release(hp.peerFlag[threadIdx.x], hp.epoch);
        if (threadIdx.x == 0) *hp.ticket = 0u;
    }
}

// ------------------------------------------------------------------------------------------------
// reductions
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// Sum over the block; result valid in thread 0.  blockDim.x multiple of 32, <= 1024.
__device__ __forceinline__ double block_sum(double v, double *smem /* >= 32 doubles */) {
    v = warp_sum(v);
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) smem[wid] = v;
    __syncthreads();
    double r = 0.0;
    if (wid == 0) {
        r = (lane < (int)(blockDim.x >> 5)) ? smem[lane] : 0.0;
        r = warp_sum(r);
    }
    return r;
}

}  // namespace mmadmm
