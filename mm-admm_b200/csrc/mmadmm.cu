// libmmadmm_b200: hand-written FP64 sm_100a kernels + C ABI for the MM-ADMM implicit step.
// See include/mmadmm.h for the boundary and DESIGN.md for the data layout and the roofline of each kernel.
//
// Hot path per ADMM iteration (reference src/MeshIntegrator.cpp:144-172):
//   k_prox      K1+K2+K3a  DXpU = D x + u (gather), per-element BFGS prox (Mesh::bfgsOptSimplex,
//                          src/Mesh.cpp:777-872), u = DXpU - z, ||z - zPrev||^2 and sum(Ih) partials
//   k_xupdate   K3b+K4     vec = tau*xBar + dt^2 WD^T(w (z-u)) by vertex->element incidence gather (deterministic),
//                          x = vec / diag(t)  (t is numerically diagonal)
//               K5         ||D x - z||^2 partials in the same pass;  k_finalize reduces the three scalars in a fixed order
// Element state is SoA over elements (component-major planes of length E) so every warp access is
// one contiguous 256-byte line.
#include <cuda_runtime.h>
#include <cub/cub.cuh>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include <dlfcn.h>
#include <pthread.h>
#include <time.h>

#include "../../include/mmadmm.h"
#include "mmadmm_device.cuh"
#include "partition.h"

using namespace mmadmm;

#define CK(call)                                                                                      \
    do {                                                                                              \
        cudaError_t _e = (call);                                                                      \
        if (_e != cudaSuccess) {                                                                      \
            char _b[512];                                                                             \
            snprintf(_b, sizeof(_b), "%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
            ctx->err = _b;                                                                            \
            return MMADMM_ECUDA;                                                                      \
        }                                                                                             \
    } while (0)

namespace {

constexpr int TPB = 128;       // element kernels: threads per block
constexpr int TPB_V = 256;     // vertex kernels
constexpr int MAXV = 4;        // D+1 <= 4
enum ProxFlags { PF_PREFETCH_GRID = 1, PF_ISO = 2 };   // PF_ISO: scalar monitor grid (isotropic monitors), compile-time path   // template switches of the persistent prox kernel
constexpr int ES_ALIGN = 128;  // element planes are padded to full prox tiles (TMA bulk copies need whole, aligned tiles)

template <int D> struct ElemArgs {
    int E, P;
    size_t Es;                   // plane stride of the SoA element state (E rounded up to a multiple of ES_ALIGN)
    const double *zsrc;          // non-null: z enters the sweep as D*zsrc (gathered), the z planes are not read
    int uZero;                   // u == 0 on entry (before the first step is taken): the u planes are not read
    const int *F[MAXV];          // connectivity planes F[v][e]
    const uint8_t *fixedBits;    // bit v: vertex v of the element is BOUNDARY_FIXED
    const uint8_t *nonIntBits;   // bit v: vertex v is not INTERIOR
    const double *x;             // P x D interleaved
    const double *Vc;            // P x D or null
    double *z, *u, *B;           // SoA planes of length E: z[n], u[n], B[n*n]
    MonGrid mg;
    FuncConsts<D> fc;
    int compMesh;
    double tol;
    double *partIh, *partDual;   // per-block partial sums
    unsigned long long *hist;    // 52 bins (may be null)
    unsigned *status;
    FaceArgs fa;                 // boundary sliding option (fa.mask & 1: project the FREE vertices' z after the BFGS solve)
};

// the D components of slot (element e, local vertex v) of a z / u array (zu_index layout: adjacent doubles)
template <int D> __device__ __forceinline__ void load_slot(const double *zu, size_t e, int v, double *out) {
    const double *p = zu + zu_index<D>(e, D * v);
    if (D == 2) {
        const double2 t = __ldg(reinterpret_cast<const double2 *>(p));
        out[0] = t.x; out[1] = t.y;
    } else {
        out[0] = __ldg(p); out[1] = __ldg(p + 1); out[2] = __ldg(p + 2);
    }
}
template <int D> __device__ __forceinline__ void gather_vertex(const double *x, int v, double *out) {
    if (D == 2) {
        double2 t = __ldg(reinterpret_cast<const double2 *>(x) + v);
        out[0] = t.x; out[1] = t.y;
    } else {
        const double *p = x + (size_t)3 * v;
        out[0] = __ldg(p); out[1] = __ldg(p + 1); out[2] = __ldg(p + 2);
    }
}

// ------------------------------------------------------------------------------------------------
// init kernels
// ------------------------------------------------------------------------------------------------
// reOrientElements (src/Mesh.cpp:243-260) + transpose of F to planes + per-element mask bits
template <int D>
__global__ void k_setup_F(int E, const int *Frow, int *F0, int *F1, int *F2, int *F3, const double *x,
                          const uint8_t *mask, uint8_t *fixedBits, uint8_t *nonIntBits) {
    int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E) return;
    int f[D + 1];
#pragma unroll
    for (int v = 0; v < D + 1; v++) f[v] = Frow[(size_t)e * (D + 1) + v];
    double p0[D];
    gather_vertex<D>(x, f[0], p0);
    SMat<D> Em;
#pragma unroll
    for (int j = 0; j < D; j++) {
        double pj[D];
        gather_vertex<D>(x, f[j + 1], pj);
#pragma unroll
        for (int l = 0; l < D; l++) Em(l, j) = pj[l] - p0[l];
    }
    if (sdet<D>(Em) < 0) { int t = f[2]; f[2] = f[1]; f[1] = t; }
    F0[e] = f[0]; F1[e] = f[1]; F2[e] = f[2];
    if (D == 3) F3[e] = f[D];
    unsigned fb = 0, nb = 0;
#pragma unroll
    for (int v = 0; v < D + 1; v++) {
        int m = mask[f[v]];
        if (m == NT_BOUNDARY_FIXED) fb |= 1u << v;
        if (m != NT_INTERIOR) nb |= 1u << v;
    }
    fixedBits[e] = (uint8_t)fb;
    nonIntBits[e] = (uint8_t)nb;
}

// keys/vals for the vertex->(element,local vertex) incidence; slot id = e*4 + v
template <int D> __global__ void k_inc_pairs(int E, const int *F0, const int *F1, const int *F2, const int *F3, int *keys, int *vals) {
    int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E) return;
    const int *F[4] = {F0, F1, F2, F3};
#pragma unroll
    for (int v = 0; v < D + 1; v++) {
        keys[(size_t)e * (D + 1) + v] = F[v][e];
        vals[(size_t)e * (D + 1) + v] = e * 4 + v;
    }
}
template <int D> __global__ void k_fill_identity(size_t Es, double *B) {
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= Es) return;
    for (int k = 0; k < D * (D + 1); k++) B[b_index<D>(e, k, k)] = 1.0;
}
__global__ void k_scatter_owned(int P, int D, const int *verts, const uint8_t *owner, const double *x, double *xg) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P || !owner[p]) return;
    for (int l = 0; l < D; l++) xg[(size_t)verts[p] * D + l] = x[(size_t)p * D + l];
}
// key = first incident element of the vertex (E for an isolated vertex), val = vertex
__global__ void k_vorder_keys(int P, int E, const int *incPtr, const int *incIdx, int *keys, int *vals) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    keys[p] = (incPtr[p + 1] > incPtr[p]) ? (incIdx[incPtr[p]] >> 2) : E;
    vals[p] = p;
}
__global__ void k_status_to_double(const unsigned *st, double *out) { *out = (double)*st; }
__global__ void k_count_keys(size_t n, const int *keys, int *cnt) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) atomicAdd(&cnt[keys[i]], 1);
}
// diag(t) = tau + dt^2 * sum_{valence} (w*w)*1.0     (src/MeshIntegrator.cpp:45-51)
__global__ void k_tdiag(int P, const int *incPtr, const int *valence, double m, double dtsq, double w, double *tdiag) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    double s = 0.0;
    int cnt = valence ? valence[p] : (incPtr[p + 1] - incPtr[p]);   // distributed: the GLOBAL valence
    for (int k = 0; k < cnt; k++) s += (w * w) * 1.0;
    tdiag[p] = m + dtsq * s;
}

// ---- monitor grid: exact 1-NN of every grid point among the mesh vertices (uniform bins), then Jacobi smoothing
struct BinGeom { int nb[3]; double lo[3], inv[3], minbw; int maxr; };

template <int D> __global__ void k_bin_ids(int P, const double *x, BinGeom bg, int *bin, int *ids) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    int c[3] = {0, 0, 0};
#pragma unroll
    for (int l = 0; l < D; l++) {
        int k = (int)((x[(size_t)p * D + l] - bg.lo[l]) * bg.inv[l]);
        c[l] = min(max(k, 0), bg.nb[l] - 1);
    }
    bin[p] = (c[2] * bg.nb[1] + c[1]) * bg.nb[0] + c[0];
    ids[p] = p;
}

// one thread per grid point; rings of bins outwards until the best candidate is provably nearest.
// Distances use explicit non-fused arithmetic so that exact ties are exact on every platform; ties -> lowest index.
template <int D>
__global__ void k_nn_grid(MonGrid mg, double *gridOut, const double *x, const double *monVertex, BinGeom bg,
                          const int *binStart, const int *items) {
    const int nx = mg.nx, ny = mg.ny, nz = (D == 3) ? mg.nz : 0;
    size_t total = (size_t)(nx + 1) * (ny + 1) * (nz + 1);
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    int i = (int)(t % (nx + 1));
    int j = (int)((t / (nx + 1)) % (ny + 1));
    int k3 = (int)(t / ((size_t)(nx + 1) * (ny + 1)));
    double q[3] = {mg.ax[i], mg.ay[j], (D == 3) ? mg.az[k3] : 0.0};
    int c[3] = {0, 0, 0};
#pragma unroll
    for (int l = 0; l < D; l++) {
        int k = (int)((q[l] - bg.lo[l]) * bg.inv[l]);
        c[l] = min(max(k, 0), bg.nb[l] - 1);
    }
    int best = -1;
    double bestd = INFINITY;
    for (int r = 0; r <= bg.maxr; r++) {
        int z0 = (D == 3) ? c[2] - r : 0, z1 = (D == 3) ? c[2] + r : 0;
        for (int kz = z0; kz <= z1; kz++) {
            if (kz < 0 || kz >= bg.nb[2]) continue;
            for (int ky = c[1] - r; ky <= c[1] + r; ky++) {
                if (ky < 0 || ky >= bg.nb[1]) continue;
                bool edgeRow = (abs(ky - c[1]) == r) || (D == 3 && abs(kz - c[2]) == r);
                int step = edgeRow ? 1 : max(2 * r, 1);  // interior rows of the shell: only the two end bins
                for (int kx = c[0] - r; kx <= c[0] + r; kx += step) {
                    if (kx < 0 || kx >= bg.nb[0]) continue;
                    size_t b = ((size_t)kz * bg.nb[1] + ky) * bg.nb[0] + kx;
                    for (int s = binStart[b]; s < binStart[b + 1]; s++) {
                        int p = items[s];
                        double d2 = 0.0;
#pragma unroll
                        for (int l = 0; l < D; l++) {
                            double df = __dsub_rn(q[l], x[(size_t)p * D + l]);
                            d2 = __dadd_rn(d2, __dmul_rn(df, df));
                        }
                        if (d2 < bestd || (d2 == bestd && p < best)) { bestd = d2; best = p; }
                    }
                }
            }
        }
        if (best >= 0) {
            double reach = r * bg.minbw;
            if (bestd < reach * reach) break;
        }
    }
    // 2D: row j*(nx+1)+i.  3D (sic, reference :198,234): row (nx+1)(ny+1)k + i*(nx+1) + j, i.e. transposed in x<->y
    size_t row = (D == 2) ? ((size_t)j * (nx + 1) + i) : ((size_t)(nx + 1) * (ny + 1) * k3 + (size_t)i * (nx + 1) + j);
    constexpr int DD = D * D, RS = GridRow<D>::stride;
#pragma unroll
    for (int k = 0; k < DD; k++) gridOut[row * RS + k] = monVertex[(size_t)best * DD + k];
}

// smoothMonitorGrid (src/MeshInterpolator.cpp:365-404): Jacobi sweep from `src` into `dst`, interior nodes only
template <int D> __global__ void k_smooth(int nx, int ny, int nz, const double *src, double *dst) {
    constexpr int DD = D * D, RS = GridRow<D>::stride;
    size_t rows = (size_t)(nx + 1) * (ny + 1) * ((D == 3) ? (nz + 1) : 1);
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= rows * DD) return;
    size_t row = t / DD;
    int k = (int)(t % DD);
    int i = (int)(row % (nx + 1));
    int j = (int)((row / (nx + 1)) % (ny + 1));
    int k3 = (int)(row / ((size_t)(nx + 1) * (ny + 1)));
    bool interior = (i >= 1 && i < nx && j >= 1 && j < ny) && (D == 2 || (k3 >= 1 && k3 < nz));
    if (!interior) return;
    size_t sx = RS, sy = (size_t)(nx + 1) * RS, sz = (size_t)(nx + 1) * (ny + 1) * RS;
    const double *c = src + row * RS + k;
    double v;
    // explicit non-fused arithmetic: the grid is input data for everything downstream, keep it bit-reproducible
    if (D == 2) {
        v = __dmul_rn(0.6, c[0]);
        v = __dadd_rn(v, __dmul_rn(0.1, c[sx]));
        v = __dadd_rn(v, __dmul_rn(0.1, *(c - sx)));
        v = __dadd_rn(v, __dmul_rn(0.1, c[sy]));
        v = __dadd_rn(v, __dmul_rn(0.1, *(c - sy)));
    } else {
        const double h = 0.4 / 6.0;
        v = __dmul_rn(0.6, c[0]);
        v = __dadd_rn(v, __dmul_rn(h, c[sx]));
        v = __dadd_rn(v, __dmul_rn(h, *(c - sx)));
        v = __dadd_rn(v, __dmul_rn(h, c[sy]));
        v = __dadd_rn(v, __dmul_rn(h, *(c - sy)));
        v = __dadd_rn(v, __dmul_rn(h, c[sz]));
        v = __dadd_rn(v, __dmul_rn(h, *(c - sz)));
    }
    dst[row * RS + k] = v;
}

// ------------------------------------------------------------------------------------------------
// element kernels
// ------------------------------------------------------------------------------------------------
template <int D> __device__ __forceinline__ void load_elem_ids(const ElemArgs<D> &a, int e, int *vid) {
#pragma unroll
    for (int v = 0; v < D + 1; v++) vid[v] = __ldg(a.F[v] + e);
}
template <int D> __device__ __forceinline__ void load_xi(const ElemArgs<D> &a, const int *vid, double *xi) {
    if (a.compMesh) {
#pragma unroll
        for (int v = 0; v < D + 1; v++) gather_vertex<D>(a.Vc, vid[v], xi + D * v);
    }
}

// K8: Mesh::computeEnergy (src/Mesh.cpp:496-530); coordinates gathered from `x`
template <int D> __global__ void __launch_bounds__(TPB) k_energy(ElemArgs<D> a, double *part) {
    constexpr int n = D * (D + 1);
    __shared__ double red[32];
    int e = blockIdx.x * TPB + threadIdx.x;
    double en = 0.0;
    if (e < a.E) {
        int vid[D + 1];
        load_elem_ids<D>(a, e, vid);
        double z[n], xi[n], g[n];
#pragma unroll
        for (int v = 0; v < D + 1; v++) gather_vertex<D>(a.x, vid[v], z + D * v);
        load_xi<D>(a, vid, xi);
        bool inv = false;
        en = block_grad<D, false>(z, xi, a.compMesh, a.mg, a.fc, g, inv);
        if (inv) atomicOr(a.status, ST_INVERTED);
    }
    double s = block_sum(en, red);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
}

// K6a: element part of Mesh::eulerGrad (src/Mesh.cpp:582-624): masked, un-regularised element gradients -> scratch planes.
// `bits` selects the slices that are dropped: fixedBits for eulerGrad (computeBlockGrad zeroes FIXED slices), nonIntBits for
// eulerStepMod (src/Mesh.cpp:533-579: only INTERIOR vertices collect).  ENERGY: also the block sums of the element energies.
template <int D, bool ENERGY> __global__ void __launch_bounds__(TPB) k_euler_elem(ElemArgs<D> a, double *gs, const uint8_t *bits, double *part) {
    constexpr int n = D * (D + 1);
    __shared__ double red[32];
    int e = blockIdx.x * TPB + threadIdx.x;
    double en = 0.0;
    if (e < a.E) {
        int vid[D + 1];
        load_elem_ids<D>(a, e, vid);
        double z[n], xi[n], g[n];
#pragma unroll
        for (int v = 0; v < D + 1; v++) gather_vertex<D>(a.x, vid[v], z + D * v);
        load_xi<D>(a, vid, xi);
        bool inv = false;
        en = block_grad<D, true>(z, xi, a.compMesh, a.mg, a.fc, g, inv);
        if (inv) atomicOr(a.status, ST_INVERTED);
        unsigned fb = bits[e];
#pragma unroll
        for (int v = 0; v < D + 1; v++)
#pragma unroll
            for (int l = 0; l < D; l++) gs[(size_t)(D * v + l) * a.Es + e] = (fb & (1u << v)) ? 0.0 : g[D * v + l];
    }
    if (ENERGY) {
        double s = block_sum(en, red);
        if (threadIdx.x == 0) part[blockIdx.x] = s;
    }
}

// DXpU = D x + u and the sweep's entry z for element e (src/MeshIntegrator.cpp:120-127,146)
template <int D>
__device__ __forceinline__ void load_z_dx(const ElemArgs<D> &a, int e, const int *vid, double *z, double *dx) {
    constexpr int n = D * (D + 1);
#pragma unroll
    for (int v = 0; v < D + 1; v++) gather_vertex<D>(a.x, vid[v], dx + D * v);
    if (!a.uZero) {
#pragma unroll
        for (int k = 0; k < n; k++) dx[k] += a.u[zu_index<D>(e, k)];
    }
    if (a.zsrc) {
#pragma unroll
        for (int v = 0; v < D + 1; v++) gather_vertex<D>(a.zsrc, vid[v], z + D * v);
    } else {
#pragma unroll
        for (int k = 0; k < n; k++) z[k] = a.z[zu_index<D>(e, k)];
    }
}

// first prox sweep only (hessComputed == false): finite-difference Hessian of the regularised, masked gradient,
// unit diagonal on non-interior vertices, general inverse by partial-pivot LU (src/Mesh.cpp:793-817).
template <int D> __global__ void __launch_bounds__(TPB) k_hess_init(ElemArgs<D> a) {
    constexpr int n = D * (D + 1);
    int e = blockIdx.x * TPB + threadIdx.x;
    if (e >= a.E) return;
    int vid[D + 1];
    load_elem_ids<D>(a, e, vid);
    unsigned fb = a.fixedBits[e], nb = a.nonIntBits[e];
    double z[n], dx[n], xi[n], Gk[n], G1[n];
    load_z_dx<D>(a, e, vid, z, dx);
    load_xi<D>(a, vid, xi);
    bool inv = false;
    prox_grad<D>(z, dx, xi, a.compMesh, a.mg, a.fc, fb, Gk, inv);
    const double h = 2.0 * 1.4901161193847656e-08;  // 2*sqrt(DBL_EPSILON)
    double H[n * n];
    for (int i = 0; i < n; i++) {
        double zp[n];
#pragma unroll
        for (int k = 0; k < n; k++) zp[k] = (k == i) ? (z[k] + h) : z[k];
        prox_grad<D>(zp, dx, xi, a.compMesh, a.mg, a.fc, fb, G1, inv);
#pragma unroll
        for (int r = 0; r < n; r++) H[r * n + i] = (G1[r] - Gk[r]) / h;
    }
    if (inv) atomicOr(a.status, ST_INVERTED);
    for (int v = 0; v < D + 1; v++)
        if (nb & (1u << v))
            for (int l = 0; l < D; l++) H[(D * v + l) * n + (D * v + l)] = 1.0;
    // partial-pivot LU in place
    int perm[n];
    for (int i = 0; i < n; i++) perm[i] = i;
    bool singular = false;
    for (int k = 0; k < n; k++) {
        int piv = k;
        double best = fabs(H[k * n + k]);
        for (int i = k + 1; i < n; i++) {
            double v = fabs(H[i * n + k]);
            if (v > best) { best = v; piv = i; }
        }
        if (best == 0.0) { singular = true; break; }
        if (piv != k) {
            for (int j = 0; j < n; j++) { double t = H[k * n + j]; H[k * n + j] = H[piv * n + j]; H[piv * n + j] = t; }
            int t = perm[k]; perm[k] = perm[piv]; perm[piv] = t;
        }
        for (int i = k + 1; i < n; i++) H[i * n + k] /= H[k * n + k];
        for (int i = k + 1; i < n; i++) {
            double l = H[i * n + k];
            for (int j = k + 1; j < n; j++) H[i * n + j] -= l * H[k * n + j];
        }
    }
    if (singular) { atomicOr(a.status, ST_SINGULAR_HESS); return; }
    for (int c = 0; c < n; c++) {
        double y[n];
        for (int i = 0; i < n; i++) {
            double s = (perm[i] == c) ? 1.0 : 0.0;
            for (int j = 0; j < i; j++) s -= H[i * n + j] * y[j];
            y[i] = s;
        }
        for (int i = n - 1; i >= 0; i--) {
            double s = y[i];
            for (int j = i + 1; j < n; j++) s -= H[i * n + j] * y[j];
            y[i] = s / H[i * n + i];
        }
        for (int i = 0; i < n; i++) a.B[b_index<D>(e, i, c)] = y[i];
    }
}

// K1+K2+K3a: the per-element BFGS prox (Mesh::prox -> bfgsOptSimplex, src/Mesh.cpp:930-994, 777-872), one thread per
// element.  The CTA's tile of inverse Hessians (n*n planes x TPB elements) is brought into shared memory by TMA
// bulk copies while the threads gather their vertices and evaluate the entry gradient; the three passes over
// Bkinv (pk = -B g; B y and y^T B; rank-2 update) run out of shared memory ([plane][thread]: conflict-free), and the
// updated tile goes back with TMA bulk stores.  Also: DXpU = D x + u (K1), u = DXpU - z (K3a), sum(Ih) and
// ||z - zPrev||^2 partials (K5).
template <int D, int TPB_, int MINB, bool COMP, bool ISO>
__global__ void __launch_bounds__(TPB_, MINB) k_prox(ElemArgs<D> a) {
    constexpr int n = D * (D + 1);
    constexpr int NN = n * n;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *Bs = reinterpret_cast<double *>(smem_raw);                 // [TPB_/32 tiles][NN][32]
    uint64_t *bar = reinterpret_cast<uint64_t *>(Bs + (size_t)NN * TPB_);
    double *red = reinterpret_cast<double *>(bar + 2);                 // 32 doubles
    int *shist = reinterpret_cast<int *>(red + 32);                    // 52 ints
    const int tid = threadIdx.x;
    const size_t e0 = (size_t)blockIdx.x * TPB_;
    if (tid == 0) mbar_init(bar, 1);
    if (a.hist && tid < 52) shist[tid] = 0;
    __syncthreads();
    if (tid == 0) mbar_expect_tx(bar, (unsigned)(NN * TPB_ * sizeof(double)));
    if (tid < TPB_ / TILE)   // one contiguous block per 32-element tile
        bulk_g2s(Bs + (size_t)tid * NN * TILE, a.B + (e0 / TILE + tid) * NN * TILE, (unsigned)(NN * TILE * sizeof(double)), bar);
    const int e = (int)e0 + tid;
    double ih = 0.0, dual2 = 0.0;
    const bool active = e < a.E;
    double z[n], dx[n], xi[COMP ? n : 1], Gk[n];
    unsigned fb = 0;
    bool inv = false;
    if (active) {
        int vid[D + 1];
        load_elem_ids<D>(a, e, vid);
        fb = a.fixedBits[e];
        load_z_dx<D>(a, e, vid, z, dx);
        if (COMP) {
#pragma unroll
            for (int v = 0; v < D + 1; v++) gather_vertex<D>(a.Vc, vid[v], xi + D * v);
        }
        ih = prox_grad<D, ISO>(z, dx, xi, COMP, a.mg, a.fc, fb, Gk, inv);   // Ihsave: energy at entry
    }
    mbar_wait(bar, 0);
    if (active) {
        double *Bt = Bs + (size_t)(tid >> 5) * NN * TILE + (tid & 31);
#define BS(k) Bt[(k) * TILE]
        double dz[n];
#pragma unroll
        for (int k = 0; k < n; k++) dz[k] = 0.0;
        int iter;
        for (iter = 0; iter < 50; iter++) {
            double pk[n], G1[n];
#pragma unroll
            for (int r = 0; r < n; r++) {
                double s = BS(r * n) * Gk[0];
#pragma unroll
                for (int c = 1; c < n; c++) s += BS(r * n + c) * Gk[c];
                pk[r] = -s;
            }
            if (iter == 0) {
#pragma unroll
                for (int k = 0; k < n; k++) { dz[k] = z[k]; z[k] += pk[k]; }   // dz holds zPrev until the end
            } else {
#pragma unroll
                for (int k = 0; k < n; k++) z[k] += pk[k];
            }
            prox_grad<D, ISO>(z, dx, xi, COMP, a.mg, a.fc, fb, G1, inv);
            double Ix = 0.0, yk[n];
#pragma unroll
            for (int k = 0; k < n; k++) { Ix += fabs(G1[k]); yk[k] = G1[k] - Gk[k]; Gk[k] = G1[k]; }
            double c2 = pk[0] * yk[0];
#pragma unroll
            for (int k = 1; k < n; k++) c2 += pk[k] * yk[k];
            // one pass over B for both B*yk (rows) and yk^T*B (columns); accumulation order as in the two-pass form
            double By[n], yB[n];
#pragma unroll
            for (int r = 0; r < n; r++) {
                double s = 0.0;
#pragma unroll
                for (int c = 0; c < n; c++) {
                    const double b = BS(r * n + c);
                    s = (c == 0) ? b * yk[0] : s + b * yk[c];
                    yB[c] = (r == 0) ? yk[0] * b : yB[c] + yk[r] * b;
                }
                By[r] = s;
            }
            double yBy = yk[0] * By[0];
#pragma unroll
            for (int k = 1; k < n; k++) yBy += yk[k] * By[k];
            const double ic2 = 1.0 / c2;
            const double c1 = (c2 + yBy) * (ic2 * ic2);
#pragma unroll
            for (int r = 0; r < n; r++) {
                const double a1 = c1 * pk[r], a3 = pk[r] * ic2, a2 = By[r] * ic2;
#pragma unroll
                for (int c = 0; c < n; c++) BS(r * n + c) += (a1 * pk[c] - a2 * pk[c]) - a3 * yB[c];
            }
            if (Ix < a.tol) break;
        }
#undef BS
        if (inv) atomicOr(a.status, ST_INVERTED);
        if (iter == 50) atomicOr(a.status, ST_BFGS_MAX);
        if (a.hist) atomicAdd(&shist[(iter == 50) ? 51 : iter + 1], 1);
        if (a.fa.mask & 1) {
            const unsigned fr = (unsigned)a.nonIntBits[e] & ~fb;
#pragma unroll
            for (int v = 0; v < D + 1; v++)
                if (fr & (1u << v)) {
                    double pt[D];
#pragma unroll
                    for (int l = 0; l < D; l++) pt[l] = z[D * v + l];
                    project_onto_boundary<D>(a.fa, __ldg(a.F[v] + e), pt);
#pragma unroll
                    for (int l = 0; l < D; l++) z[D * v + l] = pt[l];
                }
        }
#pragma unroll
        for (int k = 0; k < n; k++) {
            a.z[zu_index<D>(e, k)] = z[k];
            a.u[zu_index<D>(e, k)] = dx[k] - z[k];    // u = DXpU - z
            const double d = z[k] - dz[k];
            dual2 += d * d;
        }
    }
    // updated tile back to HBM: generic-proxy writes -> async proxy, then bulk stores
    fence_async_smem();
    __syncthreads();
    if (tid < TPB_ / TILE) {
        bulk_s2g(a.B + (e0 / TILE + tid) * NN * TILE, Bs + (size_t)tid * NN * TILE, (unsigned)(NN * TILE * sizeof(double)));
        bulk_commit();
    }
    double s1 = block_sum(ih, red);
    double s2 = block_sum(dual2, red);
    if (tid == 0) { a.partIh[blockIdx.x] = s1; a.partDual[blockIdx.x] = s2; }
    if (a.hist) {
        __syncthreads();
        if (tid < 52 && shist[tid]) atomicAdd(&a.hist[tid], (unsigned long long)shist[tid]);
    }
    if (tid < TPB_ / TILE) bulk_wait_read0();   // shared memory must stay allocated until the bulk stores have read it
}

// The 3D prox kernel (D = 3: n = 12, a 12 x 12 inverse Hessian = 1152 bytes per tet).  Thread-per-element cannot work here -- the
// eight 12-vectors of the BFGS update plus blockGrad's 3x3 temporaries need ~300 registers -- so every element gets FOUR lanes,
// one per vertex: a warp owns a sub-tile of 8 elements, lane = el + 8 q (el = element of the sub-tile, q = vertex).  A thread
// keeps its vertex's D components of every BFGS vector and works on rows 3q..3q+2 of its element's inverse Hessian, which sits
// in the warp's shared memory as [entry][8] (one contiguous 9216-byte block per sub-tile in HBM: one TMA bulk copy each way,
// SASS UBLKCP; layout b3_pos, mmadmm_device.cuh).  What the other three vertices hold arrives by warp shuffles.  There is no CTA
// barrier and no shared-memory exchange buffer: warps are independent, persistent, and software-pipelined like the 2D kernel --
// the own-vertex inputs of the next sub-tile (x gather, u, z, mask byte) are fetched by per-thread cp.async while this one is
// solved, the connectivity two sub-tiles ahead, and the sub-tile's inverse Hessians land while the entry gradient is evaluated.
// The part of blockGrad after the monitor lookups is evaluated by all four lanes of an element (it needs all of z); each lane
// forms only its own vertex's block of the gradient (block_grad_own_iso).
// (Round 1's k_prox_vw gave a CTA of four vertex WARPS one 32-element tile and paid four CTA barriers per BFGS iteration:
// 3.81 ms on C4; this form: see DESIGN.md 3.2.)
// KEEP: the thread's own rows of B (3 x 12) stay in registers from pass 1 to pass 3 (and across BFGS iterations): 108 instead
// of 180 shared-memory accesses per thread and iteration, at 72 more registers (A/B variants, "prox_variant" 3-5).
template <bool COMP, bool ISO, int WPB, int MINB, bool KEEP = false>
__global__ void __launch_bounds__(WPB * 32, MINB) k_prox_q(const __grid_constant__ ElemArgs<3> a) {
    constexpr int D = 3, n = 12, NN = 144, NV = 4, DD = 9;
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int WARP_DOUBLES = NN * SUB3 + (COMP ? 4 : 3) * D * 32;           // B sub-tile | next x | next u | next z | (next Vc)
    constexpr int WARP_BYTES = WARP_DOUBLES * 8 + 2 * 32 * 4 + 16;                // + connectivity slot, mask-word slot, mbarrier
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int el = lane & 7, q = lane >> 3;
    unsigned char *wbase = smem_raw + (size_t)warp * ((WARP_BYTES + 127) / 128 * 128);
    double *Bs = reinterpret_cast<double *>(wbase);           // [144][8]
    double *Xn = Bs + NN * SUB3;                               // [3][32]  thread-private slots
    double *Un = Xn + D * 32;
    double *Zn = Un + D * 32;
    double *Cn = Zn + D * 32;                                  // COMP only
    int *Vn = reinterpret_cast<int *>(Bs + WARP_DOUBLES);      // [32] own vertex id of the sub-tile after next
    int *Fn = Vn + 32;                                         // [32] the word of fixedBits that holds the element's byte
    uint64_t *barB = reinterpret_cast<uint64_t *>(Fn + 32);
    const int gw = blockIdx.x * WPB + warp, GW = gridDim.x * WPB;
    const int nSub = (a.E + SUB3 - 1) / SUB3;
    if (lane == 0) mbar_init(barB, 1);
    __syncwarp();

    // own rows of the element's B, all columns: entry (3q+rl, c) at ((3q+rl)*12 + (c ^ (q&1)))*8 + el
    const int qodd = q & 1;
    const double *rowE = Bs + (size_t)(3 * q) * 12 * SUB3 + el + (qodd ? SUB3 : 0);    // even c
    const double *rowO = Bs + (size_t)(3 * q) * 12 * SUB3 + el - (qodd ? SUB3 : 0);    // odd c
#define BROW(rl, c) ((((c) & 1) ? rowO : rowE)[((rl) * 12 + (c)) * SUB3])
    // own columns, all rows: entry (r, 3q+cl) at (r*12 + ((3q+cl) ^ ((r/3)&1)))*8 + el
    const double *colE = Bs + (size_t)(3 * q) * SUB3 + el;                              // rows of even vertices; + cl*8
    int coff[D];
#pragma unroll
    for (int cl = 0; cl < D; cl++) coff[cl] = ((3 * q + cl) ^ 1) * SUB3 + el;            // rows of odd vertices
#define BCOL(r, cl) (((((r) / 3) & 1) ? (Bs + coff[cl]) : (colE + (cl) * SUB3))[(r) * 12 * SUB3])

    auto prefetch = [&](int t, int vid) {   // own-vertex inputs of sub-tile t (planes and F are padded to whole tiles: always in bounds)
        const int e = t * SUB3 + el;
#pragma unroll
        for (int l = 0; l < D; l++) {
            const int sl = l * 32 + lane;
            cp_async8(Xn + sl, a.x + (size_t)D * vid + l);
            if (!a.uZero) cp_async8(Un + sl, a.u + zu_index<D>(e, D * q + l));
            if (a.zsrc) cp_async8(Zn + sl, a.zsrc + (size_t)D * vid + l);
            else cp_async8(Zn + sl, a.z + zu_index<D>(e, D * q + l));
            if (COMP) cp_async8(Cn + sl, a.Vc + (size_t)D * vid + l);
        }
        cp_async4(Fn + lane, reinterpret_cast<const int *>(a.fixedBits) + (e >> 2));
    };
    auto prefetch_vid = [&](int t) {
        if (t < nSub) cp_async4(Vn + lane, a.F[q] + (size_t)t * SUB3 + el);
    };
    // the D components every vertex of the element holds -> all n of them (vertex-major), by shuffles
    auto all_of = [&](const double *own, double *all) {
#pragma unroll
        for (int v = 0; v < NV; v++)
#pragma unroll
            for (int l = 0; l < D; l++) all[D * v + l] = __shfl_sync(FULL, own[l], el + 8 * v);
    };
    auto sum4 = [&](double part) {   // sum over the element's vertices in the order v = 0..3 (identical on its four lanes)
        double sacc = __shfl_sync(FULL, part, el);
#pragma unroll
        for (int v = 1; v < NV; v++) sacc += __shfl_sync(FULL, part, el + 8 * v);
        return sacc;
    };

    int t = gw;
    if (t < nSub) {
        prefetch(t, __ldg(a.F[q] + (size_t)t * SUB3 + el));
        prefetch_vid(t + GW);
    }
    cp_async_commit();
    double ihAcc = 0.0, dualAcc = 0.0;
    for (int it = 0; t < nSub; t += GW, it++) {
        const int e = t * SUB3 + el;
        const bool active = e < a.E;
        if (lane == 0) {   // this sub-tile's inverse Hessians, once the previous store has drained out of shared memory
            bulk_wait_read0();
            mbar_expect_tx(barB, (unsigned)(NN * SUB3 * sizeof(double)));
            bulk_g2s(Bs, a.B + (size_t)t * NN * SUB3, (unsigned)(NN * SUB3 * sizeof(double)), barB);
        }
        cp_async_wait_all();
        double zo[D], dxo[D], zprev[D], Gk[D], xi[COMP ? n : 1];
#pragma unroll
        for (int l = 0; l < D; l++) {
            const int sl = l * 32 + lane;
            zo[l] = Zn[sl];
            dxo[l] = a.uZero ? Xn[sl] : Xn[sl] + Un[sl];
            zprev[l] = zo[l];
        }
        if (COMP) {
            double co[D];
#pragma unroll
            for (int l = 0; l < D; l++) co[l] = Cn[l * 32 + lane];
            all_of(co, xi);
        }
        const bool ownFixed = ((unsigned)Fn[lane] >> (8 * (e & 3) + q)) & 1u;
        const int vidNext = Vn[lane];
        {   // next sub-tile's inputs (the slots just read are this thread's own), connectivity of the one after
            const int tn = t + GW;
            if (tn < nSub) {
                prefetch(tn, vidNext);
                prefetch_vid(tn + GW);
            }
            cp_async_commit();
        }
        bool inv = false;
        // regularised, masked gradient of the own vertex at z = (zo of the four lanes) (Mesh::computeBlockGrad, regularize = true);
        // returns the un-regularised element energy.  Executed by all lanes (shuffles inside).
        auto own_gradient = [&](double *g) -> double {
            double z[n];
            all_of(zo, z);
            double en;
            if constexpr (ISO) {
                const double mOwn = eval_monitor_iso<D>(a.mg, zo);
                double mIso[NV];
#pragma unroll
                for (int v = 0; v < NV; v++) mIso[v] = __shfl_sync(FULL, mOwn, el + 8 * v);
                en = block_grad_own_iso<D>(z, xi, COMP, mIso, a.fc, q, g, inv);
            } else {
                SMat<D> mOwn, mPre[NV];
                eval_monitor<D>(a.mg, zo, mOwn);
#pragma unroll
                for (int v = 0; v < NV; v++)
#pragma unroll
                    for (int k = 0; k < DD; k++) mPre[v].a[k] = __shfl_sync(FULL, mOwn.a[k], el + 8 * v);
                double grad[n], mDummy[1] = {0.0};
                en = block_grad_core<D, true, false>(z, xi, COMP, mPre, mDummy, a.fc, grad, inv);
#pragma unroll
                for (int l = 0; l < D; l++) g[l] = 0.0;
#pragma unroll
                for (int v = 0; v < NV; v++)
                    if (v == q) {
#pragma unroll
                        for (int l = 0; l < D; l++) g[l] = grad[D * v + l];
                    }
            }
#pragma unroll
            for (int l = 0; l < D; l++) g[l] = ownFixed ? 0.0 : g[l] + a.fc.w2 * (-dxo[l] + zo[l]);
            return en;
        };
        const double ih = own_gradient(Gk);     // Ihsave: energy at entry (identical on the element's four lanes)
        mbar_wait(barB, (unsigned)(it & 1));
        bool done = !active;
        int iters = 0;
        double Br[KEEP ? D : 1][KEEP ? n : 1];
        for (int iter = 0; iter < 50; iter++) {
            double pk[D], yk[D], By[D], yB[D], G1[D];
            {   // pass 1: pk = -B Gk (own rows), full step
                double Gall[n];
                all_of(Gk, Gall);
                if (KEEP && iter == 0) {
#pragma unroll
                    for (int rl = 0; rl < D; rl++)
#pragma unroll
                        for (int c = 0; c < n; c++) Br[KEEP ? rl : 0][KEEP ? c : 0] = BROW(rl, c);
                }
#pragma unroll
                for (int rl = 0; rl < D; rl++) {
                    double sacc = (KEEP ? Br[KEEP ? rl : 0][0] : BROW(rl, 0)) * Gall[0];
#pragma unroll
                    for (int c = 1; c < n; c++) sacc += (KEEP ? Br[KEEP ? rl : 0][KEEP ? c : 0] : BROW(rl, c)) * Gall[c];
                    pk[rl] = -sacc;
                }
            }
            if (!done) {
#pragma unroll
                for (int l = 0; l < D; l++) zo[l] += pk[l];
            }
            own_gradient(G1);
            double ixp = 0.0, c2p = 0.0;
#pragma unroll
            for (int l = 0; l < D; l++) {
                ixp += fabs(G1[l]);
                yk[l] = G1[l] - Gk[l];
                c2p = (l == 0) ? pk[0] * yk[0] : c2p + pk[l] * yk[l];
                if (!done) Gk[l] = G1[l];
            }
            {   // pass 2: B yk (own rows), yk^T B (own columns)
                double Yall[n];
                all_of(yk, Yall);
#pragma unroll
                for (int rl = 0; rl < D; rl++) {
                    double sacc = (KEEP ? Br[KEEP ? rl : 0][0] : BROW(rl, 0)) * Yall[0];
#pragma unroll
                    for (int c = 1; c < n; c++) sacc += (KEEP ? Br[KEEP ? rl : 0][KEEP ? c : 0] : BROW(rl, c)) * Yall[c];
                    By[rl] = sacc;
                }
#pragma unroll
                for (int cl = 0; cl < D; cl++) {
                    double sacc = Yall[0] * BCOL(0, cl);
#pragma unroll
                    for (int r = 1; r < n; r++) sacc += Yall[r] * BCOL(r, cl);
                    yB[cl] = sacc;
                }
            }
            double yByp = yk[0] * By[0];
#pragma unroll
            for (int l = 1; l < D; l++) yByp += yk[l] * By[l];
            const double Ix = sum4(ixp), c2 = sum4(c2p), yBy = sum4(yByp);
            const bool conv = Ix < a.tol;
            {   // pass 3: rank-2 update of the own rows (src/Mesh.cpp:846-856)
                double Pall[n], YBall[n];
                all_of(pk, Pall);
                all_of(yB, YBall);
                if (!done) {
                    const double ic2 = 1.0 / c2;
                    const double c1 = (c2 + yBy) * (ic2 * ic2);
#pragma unroll
                    for (int rl = 0; rl < D; rl++) {
                        const double a1 = c1 * pk[rl], a3 = pk[rl] * ic2, a2 = By[rl] * ic2;
#pragma unroll
                        for (int c = 0; c < n; c++) {
                            double *bp = const_cast<double *>(&BROW(rl, c));
                            if (KEEP) {
                                Br[KEEP ? rl : 0][KEEP ? c : 0] += (a1 * Pall[c] - a2 * Pall[c]) - a3 * YBall[c];
                                *bp = Br[KEEP ? rl : 0][KEEP ? c : 0];
                            } else {
                                *bp += (a1 * Pall[c] - a2 * Pall[c]) - a3 * YBall[c];
                            }
                        }
                    }
                    iters = iter + 1;
                    if (conv) done = true;
                }
            }
            if (__all_sync(FULL, done)) break;
        }
        int bin = -1;
        if (active) {
            if (inv) atomicOr(a.status, ST_INVERTED);
            if (!done) atomicOr(a.status, ST_BFGS_MAX);
            if (q == 0) { bin = done ? iters : 51; ihAcc += ih; }
            if ((a.fa.mask & 1) && !ownFixed && (((unsigned)a.nonIntBits[e] >> q) & 1u)) {   // own vertex BOUNDARY_FREE
                double pt[D];
#pragma unroll
                for (int l = 0; l < D; l++) pt[l] = zo[l];
                project_onto_boundary<D>(a.fa, __ldg(a.F[q] + e), pt);
#pragma unroll
                for (int l = 0; l < D; l++) zo[l] = pt[l];
            }
#pragma unroll
            for (int l = 0; l < D; l++) {
                a.z[zu_index<D>(e, D * q + l)] = zo[l];
                a.u[zu_index<D>(e, D * q + l)] = dxo[l] - zo[l];    // u = DXpU - z
                const double dd = zo[l] - zprev[l];
                dualAcc += dd * dd;
            }
        }
        if (a.hist) {   // one atomic per distinct iteration count in the warp
            const unsigned peers = __match_any_sync(FULL, bin);
            if (bin >= 0 && lane == __ffs(peers) - 1) atomicAdd(&a.hist[bin], (unsigned long long)__popc(peers));
        }
        // updated sub-tile back to HBM: generic-proxy writes -> async proxy, then one bulk store
        fence_async_smem();
        __syncwarp();
        if (lane == 0) {
            bulk_s2g(a.B + (size_t)t * NN * SUB3, Bs, (unsigned)(NN * SUB3 * sizeof(double)));
            bulk_commit();
        }
    }
#undef BROW
#undef BCOL
    cp_async_wait_all();
    if (lane == 0) bulk_wait_read0();   // shared memory must stay allocated until the last store has read it
    ihAcc = warp_sum(ihAcc);
    dualAcc = warp_sum(dualAcc);
    if (lane == 0) { a.partIh[gw] = ihAcc; a.partDual[gw] = dualAcc; }
}

// Persistent, software-pipelined form of the warp-tile prox kernel (the default for D = 2).  One warp walks tiles
// t = gw, gw + GW, ... of 32 elements.  While tile t is being solved, everything tile t+1 needs except its inverse
// Hessians is already on its way into the other half of a double buffer: vertex gathers (x, zsrc, Vc) by per-thread
// cp.async (LDGSTS, no registers held), the u / z planes by one tensor-map TMA each; the connectivity of tile t+2 is
// loaded into three registers.  The inverse-Hessian tile itself is single-buffered: its TMA load is issued as soon as
// the previous tile's TMA store has drained and is covered by the entry-gradient evaluation.
template <int D> __device__ __forceinline__ void gather_async(double *dst /* [D+1][32][D] */, const double *src, const int *vid, int lane) {
#pragma unroll
    for (int v = 0; v < D + 1; v++) {
        double *d = dst + ((size_t)v * 32 + lane) * D;
        const double *s = src + (size_t)D * vid[v];
        if (D == 2) cp_async16(d, s);
        else { cp_async8(d, s); cp_async8(d + 1, s + 1); cp_async8(d + 2, s + 2); }
    }
}

template <int D, int WPB, int MINB, bool COMP, int FLAGS>
__global__ void __launch_bounds__(WPB * 32, MINB) k_prox_p(const __grid_constant__ ElemArgs<D> a) {
    constexpr int n = D * (D + 1);
    constexpr int NN = n * n;
    constexpr int SLOTS = COMP ? 4 : 3;                 // per buffer: xg | u->DXpU | z (planes or gathered) | xi
    constexpr int BUF_DOUBLES = SLOTS * n * 32;
    constexpr int WARP_DOUBLES = NN * 32 + 2 * BUF_DOUBLES;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *Bs = reinterpret_cast<double *>(smem_raw) + (size_t)warp * WARP_DOUBLES;   // [NN][32]
    double *bufs = Bs + NN * 32;
    uint64_t *bars = reinterpret_cast<uint64_t *>(reinterpret_cast<double *>(smem_raw) + (size_t)WPB * WARP_DOUBLES) + 4 * warp;
    uint64_t *barB = &bars[2];
    const int gw = blockIdx.x * WPB + warp, GW = gridDim.x * WPB;
    const int nTiles = (a.E + 31) >> 5;
    const bool haveU = !a.uZero, haveZ = (a.zsrc == nullptr);
    const unsigned smallBytes = (unsigned)(((haveU ? n : 0) + (haveZ ? n : 0)) * 32 * sizeof(double));
    if (lane == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_init(barB, 1); }
    __syncwarp();

    // connectivity of a tile -> shared memory ([D+1][32] ints), asynchronously; the F planes are padded to Es
    auto issue_ids = [&](int tile, int *dst) {
#pragma unroll
        for (int v = 0; v < D + 1; v++) cp_async4(dst + v * 32 + lane, a.F[v] + (size_t)tile * 32 + lane);
    };
    auto issue_small = [&](int tile, int buf, const int *vid) {
        double *b = bufs + (size_t)buf * BUF_DOUBLES;
        gather_async<D>(b, a.x, vid, lane);
        if (!haveZ) gather_async<D>(b + 2 * n * 32, a.zsrc, vid, lane);
        if (COMP) gather_async<D>(b + 3 * n * 32, a.Vc, vid, lane);
        if (lane == 0 && smallBytes) {
            mbar_expect_tx(&bars[buf], smallBytes);
            if (haveU) bulk_g2s(b + n * 32, a.u + (size_t)tile * n * TILE, (unsigned)(n * TILE * sizeof(double)), &bars[buf]);
            if (haveZ) bulk_g2s(b + 2 * n * 32, a.z + (size_t)tile * n * TILE, (unsigned)(n * TILE * sizeof(double)), &bars[buf]);
        }
    };

    double ihAcc = 0.0, dualAcc = 0.0;
    int t = gw, tn = gw + GW;
    if (t < nTiles) {
        int vid[D + 1];
        const int e = t * 32 + lane;
#pragma unroll
        for (int v = 0; v < D + 1; v++) vid[v] = __ldg(a.F[v] + e);     // padded planes: always in bounds
        issue_small(t, 0, vid);
        if (tn < nTiles) issue_ids(tn, reinterpret_cast<int *>(bufs + BUF_DOUBLES));   // -> xg slot of buffer 1
        cp_async_commit();
    }
    for (int it = 0; t < nTiles; t = tn, tn += GW, it++) {
        const int buf = it & 1;
        double *b = bufs + (size_t)buf * BUF_DOUBLES;
        double *bn = bufs + (size_t)(buf ^ 1) * BUF_DOUBLES;
        const int e = t * 32 + lane;
        const bool active = e < a.E;
        const unsigned fb = active ? (unsigned)a.fixedBits[e] : 0u;   // consumed after the first gradient evaluation
        cp_async_wait_all();                                            // this tile's gathers and the next tile's connectivity
        if (smallBytes) mbar_wait(&bars[buf], (unsigned)((it >> 1) & 1));
        // every per-tile operand sits in shared memory as a tile image in the zu_index layout ([vertex][32][D]); X = base + lane*D
        double *xg = b + lane * D, *dxs = b + n * 32 + lane * D, *zs = b + 2 * n * 32 + lane * D, *xis = b + 3 * n * 32 + lane * D;
        double z[n], xi[COMP ? n : 1], Gk[n];
        bool inv = false;
        double ih = 0.0;
        // DXpU = D x + u -> dxs;  z at entry -> z[] (zs keeps it: zPrev of the dual residual)
#pragma unroll
        for (int k = 0; k < n; k++) {
            const double xv = xg[zu_off<D>(k)];
            dxs[zu_off<D>(k)] = haveU ? xv + dxs[zu_off<D>(k)] : xv;
            z[k] = zs[zu_off<D>(k)];
        }
        if (COMP) {
#pragma unroll
            for (int v = 0; v < D + 1; v++)
#pragma unroll
                for (int l = 0; l < D; l++) xi[D * v + l] = xis[(size_t)v * 32 * D + l];
        }
        // everything but B for the next tile; connectivity for the one after (into this buffer's spent xg slot)
        if (tn < nTiles) {
            int vid[D + 1];
#pragma unroll
            for (int v = 0; v < D + 1; v++) vid[v] = reinterpret_cast<const int *>(bn)[v * 32 + lane];
            __syncwarp();   // every lane has its ids (and has consumed xg) before the slots are overwritten
            issue_small(tn, buf ^ 1, vid);
            if (tn + GW < nTiles) issue_ids(tn + GW, reinterpret_cast<int *>(b));
        }
        cp_async_commit();
        if (lane == 0) {
            // this tile's inverse Hessians, once the previous tile's store has drained out of shared memory
            bulk_wait_read0();
            mbar_expect_tx(barB, (unsigned)(NN * 32 * sizeof(double)));
            bulk_g2s(Bs, a.B + (size_t)t * NN * TILE, (unsigned)(NN * TILE * sizeof(double)), barB);
        }
        if (active) ih = prox_grad_s<D, (FLAGS & PF_ISO) != 0>(z, dxs, xi, COMP, a.mg, a.fc, fb, Gk, inv);   // Ihsave: energy at entry
        mbar_wait(barB, (unsigned)(it & 1));
        int bin = -1;
        if (active) {
            double *Bt = Bs + lane;
#define BS(k) Bt[(k) * 32]
            int iter;
            for (iter = 0; iter < 50; iter++) {
                double pk[n], G1[n];
#pragma unroll
                for (int r = 0; r < n; r++) {
                    double s = BS(r * n) * Gk[0];
#pragma unroll
                    for (int c = 1; c < n; c++) s += BS(r * n + c) * Gk[c];
                    pk[r] = -s;
                    z[r] += pk[r];
                }
                prox_grad_s<D, (FLAGS & PF_ISO) != 0>(z, dxs, xi, COMP, a.mg, a.fc, fb, G1, inv);
                double Ix = 0.0, yk[n];
#pragma unroll
                for (int k = 0; k < n; k++) { Ix += fabs(G1[k]); yk[k] = G1[k] - Gk[k]; Gk[k] = G1[k]; }
                double c2 = pk[0] * yk[0];
#pragma unroll
                for (int k = 1; k < n; k++) c2 += pk[k] * yk[k];
                double By[n], yB[n];
#pragma unroll
                for (int r = 0; r < n; r++) {
                    double s = 0.0;
#pragma unroll
                    for (int c = 0; c < n; c++) {
                        const double bb = BS(r * n + c);
                        s = (c == 0) ? bb * yk[0] : s + bb * yk[c];
                        yB[c] = (r == 0) ? yk[0] * bb : yB[c] + yk[r] * bb;
                    }
                    By[r] = s;
                }
                double yBy = yk[0] * By[0];
#pragma unroll
                for (int k = 1; k < n; k++) yBy += yk[k] * By[k];
                const double ic2 = 1.0 / c2;
                const double c1 = (c2 + yBy) * (ic2 * ic2);
#pragma unroll
                for (int r = 0; r < n; r++) {
                    const double a1 = c1 * pk[r], a3 = pk[r] * ic2, a2 = By[r] * ic2;
#pragma unroll
                    for (int c = 0; c < n; c++) BS(r * n + c) += (a1 * pk[c] - a2 * pk[c]) - a3 * yB[c];
                }
                if (Ix < a.tol) break;
            }
#undef BS
            if (inv) atomicOr(a.status, ST_INVERTED);
            if (iter == 50) atomicOr(a.status, ST_BFGS_MAX);
            bin = (iter == 50) ? 51 : iter + 1;
            if (a.fa.mask & 1) {   // "Moving boundary points" (src/Mesh.cpp:975-984): rare, boundary elements only
                const unsigned fr = (unsigned)a.nonIntBits[e] & ~fb;
                if (fr) {
#pragma unroll
                    for (int v = 0; v < D + 1; v++)
                        if (fr & (1u << v)) {
                            double pt[D];     // only this copy has its address taken: z stays in registers
#pragma unroll
                            for (int l = 0; l < D; l++) pt[l] = z[D * v + l];
                            project_onto_boundary<D>(a.fa, __ldg(a.F[v] + e), pt);
#pragma unroll
                            for (int l = 0; l < D; l++) z[D * v + l] = pt[l];
                        }
                }
            }
            double dual2 = 0.0;
#pragma unroll
            for (int k = 0; k < n; k++) {
                a.z[(size_t)t * n * TILE + lane * D + zu_off<D>(k)] = z[k];
                a.u[(size_t)t * n * TILE + lane * D + zu_off<D>(k)] = dxs[zu_off<D>(k)] - z[k];    // u = DXpU - z
                const double d = z[k] - zs[zu_off<D>(k)];
                dual2 += d * d;
            }
            ihAcc += ih;
            dualAcc += dual2;
        }
        if (a.hist) {   // one atomic per distinct iteration count in the warp
            const unsigned peers = __match_any_sync(0xffffffffu, bin);
            if (bin >= 0 && lane == __ffs(peers) - 1) atomicAdd(&a.hist[bin], (unsigned long long)__popc(peers));
        }
        // updated tile back to HBM: generic-proxy writes -> async proxy, then one tensor-map store (TMA stores beat
        // per-thread STG of the 36 planes by 25 % here: profiles/r01_b_summary.md)
        fence_async_smem();
        __syncwarp();
        if (lane == 0) {
            bulk_s2g(a.B + (size_t)t * NN * TILE, Bs, (unsigned)(NN * TILE * sizeof(double)));
            bulk_commit();
        }
        if (tn < nTiles && (FLAGS & PF_PREFETCH_GRID)) {
            // the next tile's vertices have landed by now: pull the monitor-grid rows its entry gradient will read into L2
            cp_async_wait_all();
            if (haveZ) mbar_wait(&bars[buf ^ 1], (unsigned)(((it + 1) >> 1) & 1));
            const double *zn = bn + 2 * n * 32;
#pragma unroll
            for (int v = 0; v < D + 1; v++) {
                double pnt[D];
#pragma unroll
                for (int l = 0; l < D; l++) pnt[l] = zn[((size_t)v * 32 + lane) * D + l];
                prefetch_monitor<D, (FLAGS & PF_ISO) != 0>(a.mg, pnt);
            }
        }
    }
    cp_async_wait_all();
    if (lane == 0) bulk_wait_read0();   // shared memory must stay allocated until the last store has read it
    ihAcc = warp_sum(ihAcc);
    dualAcc = warp_sum(dualAcc);
    if (lane == 0) { a.partIh[gw] = ihAcc; a.partDual[gw] = dualAcc; }
}

// z = D * src (gather), src is a P x D vertex vector
template <int D> __global__ void __launch_bounds__(TPB) k_applyD(ElemArgs<D> a, const double *src) {
    int e = blockIdx.x * TPB + threadIdx.x;
    if (e >= a.E) return;
    int vid[D + 1];
    load_elem_ids<D>(a, e, vid);
#pragma unroll
    for (int v = 0; v < D + 1; v++) {
        double p[D];
        gather_vertex<D>(src, vid[v], p);
#pragma unroll
        for (int l = 0; l < D; l++) a.z[zu_index<D>(e, D * v + l)] = p[l];
    }
}

// fixed-order final reduction of up to 3 partial arrays (lengths n0,n1,n2) -> out[0..2]
// `out` may be mapped host memory (single-GPU contexts: the host reads the scalars after a stream synchronise, no copy is
// queued); then the status word travels the same way.
__global__ void __launch_bounds__(1024) k_finalize(int n0, const double *p0, int n1, const double *p1, int n2, const double *p2,
                                                   double *out, const unsigned *statusIn = nullptr, unsigned *statusOut = nullptr) {
    __shared__ double red[32];
    const double *ps[3] = {p0, p1, p2};
    const int ns[3] = {n0, n1, n2};
    for (int q = 0; q < 3; q++) {
        if (!ps[q]) continue;
        double s = 0.0;
        for (int i = threadIdx.x; i < ns[q]; i += blockDim.x) s += ps[q][i];
        double r = block_sum(s, red);
        if (threadIdx.x == 0) out[q] = r;
        __syncthreads();
    }
    if (statusOut && threadIdx.x == 0) *statusOut = *statusIn;
}

// ------------------------------------------------------------------------------------------------
// vertex kernels
// ------------------------------------------------------------------------------------------------
// x of the first x-solve of a step at one vertex component (see k_xfirst): one definition for k_xfirst and for the speculative
// form inside k_xupdate, so that both round identically
__device__ __forceinline__ double xfirst_value(double m, double dtsq, double w, double val, double xb, double zs, double us, double td) {
    // explicit roundings (no contraction, like the oracle's -ffp-contract=off): identical bits in both kernels
    const double sv = w * (w * __dsub_rn(__dmul_rn(val, zs), us));
    return __dadd_rn(__dmul_rn(m, xb), __dmul_rn(dtsq, sv)) / td;
}
// What a k_xupdate launch needs to ALSO be the last kernel of a single-GPU step:
//  * next: the next outer step's predictor + first x-solve (k_xfirst<EXTRAP>) computed speculatively from this update's
//    result (x_new, usum+ and xPrev are all at hand in the thread): xBarNext = 2 x_new - xPrev, xNext = first solve.  If the
//    host then ends the step (the usual case: one ADMM iteration per step), the next step starts by swapping buffers
//    instead of launching k_xfirst (0.04 ms and 227 MB on C2 against 100 MB more here);
//  * fin: the fixed-order final reduction of the residual partials (k_finalize) by the last CTA to finish.
struct StepTail {
    const double *xPrev = nullptr;
    double *xNext = nullptr, *xBarNext = nullptr;
    int nProx = 0;                       // fin: number of prox partials; 0 = no fused reduction
    const double *partIh = nullptr, *partDual = nullptr;
    double *out = nullptr;               // 3 scalars (mapped host memory)
    const unsigned *statusIn = nullptr;
    unsigned *statusOut = nullptr, *ticket = nullptr;
};

// K3b+K4(+K5): x = (m*xBar + dt^2 * sum_{inc, ascending element} w*(w*(z-u))) / tdiag   (src/MeshIntegrator.cpp:137-138,159-160)
//   ZSRC : z == D*zsrc, so every incident slot of vertex p holds zsrc[p] (first solve of a step; z planes not read)
//   UZERO: u == 0 (before the first step; u planes not read)
//   PRIMAL: also accumulate ||D x - z||^2 = sum_p sum_{slots of p} |x_p - z_slot|^2 from the z values already in
//           registers (src/MeshIntegrator.cpp:162) -> per-block partials
//   UREC : the dual planes are not read either.  usum_p = sum of u over the slots of p is carried per vertex: the prox sweep
//          set u+ = (x_p + u) - z+ in every slot, so usum+ = usum + val_p*x_p - zsum_p, and the right-hand side is
//          w(w(zsum - usum+)).  Vertices shared with another rank keep the slot-by-slot form (their sums are partial).
//          The exact form (UREC = false) also stores usum, which re-anchors the recurrence ("usum_period" option).
template <int D, bool ZSRC, bool UZERO, bool PRIMAL, bool UREC, int CH = 4, int MINB = 1>
__global__ void __launch_bounds__(TPB_V, MINB) k_xupdate(int P, size_t Es, const int *incPtr, const int *incIdx, const double *z,
                                                   const double *u, const double *zsrc, const double *xBar,
                                                   const double *tdiag, double m, double dtsq, double w, double *x,
                                                   double *partPrimal, const int *sharedSlot, double *sPart, const int *vorder,
                                                   double *usum, const HaloPush hp, const int skipShared, const StepTail tail) {
    static_assert(!UREC || (PRIMAL && !ZSRC && !UZERO), "the recurrence belongs to the update that follows a prox sweep");
    __shared__ double red[32];
    __shared__ int lastCta;
    const int tix = blockIdx.x * TPB_V + threadIdx.x;
    double prim = 0.0;
    // skipShared: the vertices shared with other ranks were handled by an earlier launch over the shared list (their partial
    // sums are already on their way to the neighbours while this launch does the bulk of the vertices)
    if (tix < P && !(skipShared && sharedSlot[vorder[tix]] >= 0)) {
        const int p = vorder[tix];   // vertices in the order of their first incident element: neighbouring threads
                                     // read neighbouring element slots in every plane (2.4x less DRAM traffic)
        double s[D], zs[D], xn[D], us[D];
#pragma unroll
        for (int l = 0; l < D; l++) { s[l] = 0.0; us[l] = 0.0; zs[l] = ZSRC ? zsrc[(size_t)D * p + l] : 0.0; }
        const int k0 = incPtr[p], k1 = incPtr[p + 1];
        const int sh = sharedSlot ? sharedSlot[p] : -1;
        const bool rec = UREC && sh < 0;
        // UREC: the primal share sum_s |x_new - z_s|^2 is formed in the SAME pass as zsum, shifted by the vertex's previous
        // position xo (known before the pass; |z_s - xo| and |x_new - xo| are step-sized, so nothing large cancels):
        //   sum_s |d - e_s|^2 = val |d|^2 - 2 d . S1 + S2,  d = x_new - xo, e_s = z_s - xo, S1 = sum e_s, S2 = sum |e_s|^2.
        // The second pass over the slots (half of the kernel's L1 requests) disappears; x itself is unchanged bit for bit.
        double xo[D], S1[D], S2 = 0.0, usn[D];
#pragma unroll
        for (int l = 0; l < D; l++) { xo[l] = UREC ? x[(size_t)D * p + l] : 0.0; S1[l] = 0.0; usn[l] = 0.0; }
        if (UREC && sh >= 0) {   // shared with another rank: slot-by-slot form, plain loop (few vertices)
            for (int k = k0; k < k1; k++) {
                const int slot = __ldg(incIdx + k);
                const size_t e = (size_t)(slot >> 2);
                const int v = slot & 3;
                double zq[D], uq[D];
                load_slot<D>(z, e, v, zq);
                load_slot<D>(u, e, v, uq);
#pragma unroll
                for (int l = 0; l < D; l++) s[l] += w * (w * (zq[l] - uq[l]));
            }
        } else {
            // slots in chunks of CH: all index loads, then all value loads, then the (ordered) accumulation
            for (int kb = k0; kb < k1; kb += CH) {
                int slot[CH];
                double zv[CH][D], uv[UREC ? 1 : CH][D];
#pragma unroll
                for (int j = 0; j < CH; j++) slot[j] = (kb + j < k1) ? __ldg(incIdx + kb + j) : -1;
#pragma unroll
                for (int j = 0; j < CH; j++) {
                    const size_t e = (size_t)(slot[j] >> 2);
                    const int v = slot[j] & 3;
                    if (ZSRC || slot[j] < 0) {
#pragma unroll
                        for (int l = 0; l < D; l++) zv[j][l] = zs[l];
                    } else {
                        load_slot<D>(z, e, v, zv[j]);
                    }
                    if (!UREC) {
                        if (UZERO || slot[j] < 0) {
#pragma unroll
                            for (int l = 0; l < D; l++) uv[j][l] = 0.0;
                        } else {
                            load_slot<D>(u, e, v, uv[j]);
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < CH; j++)
                    if (slot[j] >= 0) {
#pragma unroll
                        for (int l = 0; l < D; l++) {
                            if (UREC) {
                                s[l] += zv[j][l];   // zsum
                                const double el = zv[j][l] - xo[l];
                                S1[l] += el;
                                S2 += el * el;
                            } else { s[l] += w * (w * (zv[j][l] - uv[j][l])); us[l] += uv[j][l]; }
                        }
                    }
            }
        }
        if (PRIMAL && sh < 0) {
#pragma unroll
            for (int l = 0; l < D; l++) {
                if (rec) {
                    const double un = (usum[(size_t)D * p + l] + (double)(k1 - k0) * xo[l]) - s[l];
                    usum[(size_t)D * p + l] = un;
                    usn[l] = un;
                    s[l] = w * (w * (s[l] - un));
                } else {
                    usum[(size_t)D * p + l] = us[l];
                }
            }
        }
        if (sh >= 0) {   // shared with another rank: only the partial sum is known here (finished by k_xfinish)
#pragma unroll
            for (int l = 0; l < D; l++) sPart[(size_t)sh * D + l] = s[l];
            if (hp.ptr) halo_push_store<D>(hp, sh, s);   // ... and goes straight into the sharers' mailboxes over NVLink
        }
        double td = tdiag[p];
#pragma unroll
        for (int l = 0; l < D; l++) {
            xn[l] = ((m * xBar[(size_t)D * p + l]) + dtsq * s[l]) / td;
            if (sh < 0) x[(size_t)D * p + l] = xn[l];
        }
        if (PRIMAL && rec) {
            double acc = S2;
#pragma unroll
            for (int l = 0; l < D; l++) {
                const double dl = xn[l] - xo[l];
                acc += dl * ((double)(k1 - k0) * dl - 2.0 * S1[l]);
            }
            prim = acc > 0.0 ? acc : 0.0;
        }
        if (UREC && rec && tail.xNext) {   // the next step's predictor and first x-solve, speculatively (see StepTail)
            const double val = (double)(k1 - k0);
#pragma unroll
            for (int l = 0; l < D; l++) {
                const size_t i = (size_t)D * p + l;
                const double xb = 2 * xn[l] - tail.xPrev[i];
                tail.xBarNext[i] = xb;
                tail.xNext[i] = xfirst_value(m, dtsq, w, val, xb, xb, usn[l], td);
            }
        }
        if (PRIMAL && !rec && sh < 0) {   // second pass over the (L1-resident) slots: differences, not expanded squares
            for (int k = k0; k < k1; k++) {
                int slot = __ldg(incIdx + k);
                size_t e = (size_t)(slot >> 2);
                int v = slot & 3;
                double zv[D];
                if (ZSRC) {
#pragma unroll
                    for (int l = 0; l < D; l++) zv[l] = zs[l];
                } else {
                    load_slot<D>(z, e, v, zv);
                }
#pragma unroll
                for (int l = 0; l < D; l++) {
                    double df = xn[l] - zv[l];
                    prim += df * df;
                }
            }
        }
    }
    if (PRIMAL) {
        double r = block_sum(prim, red);
        if (threadIdx.x == 0) partPrimal[blockIdx.x] = r;
    }
    if (hp.ptr) halo_push_publish(hp);
    if (UREC && tail.nProx > 0) {   // the CTA that finishes last reduces the step's three scalars in k_finalize's fixed order
        if (threadIdx.x == 0) {
            __threadfence();
            const unsigned t = atomicAdd(tail.ticket, 1u);
            lastCta = (t == gridDim.x - 1);
            if (lastCta) *tail.ticket = 0;   // ready for the next launch (stream order)
        }
        __syncthreads();
        if (!lastCta) return;
        __threadfence();
        const double *ps[3] = {tail.partIh, partPrimal, tail.partDual};
        const int ns[3] = {tail.nProx, (int)gridDim.x, tail.nProx};
#pragma unroll
        for (int q = 0; q < 3; q++) {
            double sq = 0.0;
            for (int i = threadIdx.x; i < ns[q]; i += TPB_V) sq += __ldcg(ps[q] + i);
            const double r = block_sum(sq, red);
            if (threadIdx.x == 0) tail.out[q] = r;
            __syncthreads();
        }
        if (threadIdx.x == 0) *tail.statusOut = *tail.statusIn;
    }
}

// Distributed x-update, shared vertices: x from the exchanged total, and their share of ||D x - z||^2 over the LOCAL slots
template <int D, bool ZSRC, bool PRIMAL>
__global__ void __launch_bounds__(TPB_V) k_xfinish(int nShared, const int *sharedIdx, const double *sTot, size_t Es, const int *incPtr,
                                                   const int *incIdx, const double *z, const double *zsrc, const double *xBar,
                                                   const double *tdiag, double m, double dtsq, double *x, double *partPrimal) {
    __shared__ double red[32];
    const int i = blockIdx.x * TPB_V + threadIdx.x;
    double prim = 0.0;
    if (i < nShared) {
        const int p = sharedIdx[i];
        const double td = tdiag[p];
        double xn[D];
#pragma unroll
        for (int l = 0; l < D; l++) {
            xn[l] = ((m * xBar[(size_t)D * p + l]) + dtsq * sTot[(size_t)i * D + l]) / td;
            x[(size_t)D * p + l] = xn[l];
        }
        if (PRIMAL) {
            for (int k = incPtr[p]; k < incPtr[p + 1]; k++) {
                const int slot = __ldg(incIdx + k);
                const size_t e = (size_t)(slot >> 2);
                const int v = slot & 3;
#pragma unroll
                for (int l = 0; l < D; l++) {
                    const double zv = ZSRC ? zsrc[(size_t)D * p + l] : __ldg(z + zu_index<D>(e, D * v + l));
                    const double df = xn[l] - zv;
                    prim += df * df;
                }
            }
        }
    }
    if (PRIMAL) {
        double r = block_sum(prim, red);
        if (threadIdx.x == 0) partPrimal[blockIdx.x] = r;
    }
}
// Distributed predictor, shared vertices: xBar = x - (dt/tau) * (exchanged gradient total)
template <int D>
__global__ void k_predict_finish(int nShared, const int *sharedIdx, const double *sTot, double dtOverTau, const double *x,
                                 double *xBar, double *xPrev) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nShared) return;
    const int p = sharedIdx[i];
#pragma unroll
    for (int l = 0; l < D; l++) {
        const double xv = x[(size_t)D * p + l];
        xBar[(size_t)D * p + l] = xv - dtOverTau * sTot[(size_t)i * D + l];
        xPrev[(size_t)D * p + l] = xv;
    }
}

// K6b: grad = sum of element gradients (ascending element order, as the reference's serial loop accumulates);
// xBar = x - (dt/tau)*grad ; xPrev = x     (src/Mesh.cpp:649-674, src/MeshIntegrator.cpp:119)
template <int D>
__global__ void __launch_bounds__(TPB_V) k_predict_grad(int P, size_t Es, const int *incPtr, const int *incIdx, const double *gs,
                                                        double dtOverTau, const double *x, double *xBar, double *xPrev,
                                                        const int *sharedSlot, double *sPart, const int *vorder) {
    const int tix = blockIdx.x * TPB_V + threadIdx.x;
    if (tix >= P) return;
    const int p = vorder[tix];
    double s[D];
#pragma unroll
    for (int l = 0; l < D; l++) s[l] = 0.0;
    int k0 = incPtr[p], k1 = incPtr[p + 1];
    for (int k = k0; k < k1; k++) {
        int slot = __ldg(incIdx + k);
        size_t e = (size_t)(slot >> 2);
        int v = slot & 3;
#pragma unroll
        for (int l = 0; l < D; l++) s[l] += __ldg(gs + (size_t)(D * v + l) * Es + e);
    }
    const int sh = sharedSlot ? sharedSlot[p] : -1;
    if (sh >= 0) {
#pragma unroll
        for (int l = 0; l < D; l++) sPart[(size_t)sh * D + l] = s[l];
        return;
    }
#pragma unroll
    for (int l = 0; l < D; l++) {
        double xv = x[(size_t)D * p + l];
        xBar[(size_t)D * p + l] = xv - dtOverTau * s[l];
        xPrev[(size_t)D * p + l] = xv;
    }
}
// K7: xBar = 2x - xPrev ; xPrev = x
__global__ void __launch_bounds__(TPB_V) k_predict_extrap(size_t N, const double *x, double *xBar, double *xPrev) {
    size_t i = (size_t)blockIdx.x * TPB_V + threadIdx.x;
    if (i >= N) return;
    double xv = x[i];
    xBar[i] = 2 * xv - xPrev[i];
    xPrev[i] = xv;
}

// First x-solve of a step (src/MeshIntegrator.cpp:137-138) without touching element state: z = D zsrc there, so
// sum_slots w(w(z - u)) = w(w(val_p*zsrc_p - usum_p)) with the per-vertex dual sums that k_xupdate keeps.  One thread per
// component, natural vertex order (everything coalesced).  EXTRAP: fused with K7 (xBar = 2x - xPrev, xPrev = x; zsrc = xBar).
// Vertices shared with another rank only get the predictor; their x comes from k_xupdate<ZSRC> over the shared list.
template <int D, bool EXTRAP>
__global__ void __launch_bounds__(TPB_V) k_xfirst(size_t N, const int *incPtr, const double *usum, const double *zsrc, double *xBar,
                                                  double *xPrev, const double *tdiag, double m, double dtsq, double w, double *x,
                                                  const int *sharedSlot, const int skipShared) {
    const size_t i = (size_t)blockIdx.x * TPB_V + threadIdx.x;
    if (i >= N) return;
    const int p = (int)(i / D);
    if (skipShared && sharedSlot[p] >= 0) return;   // predictor of the shared vertices already done (k_predict_shared)
    double xb, zs;
    if (EXTRAP) {
        const double xv = x[i];
        xb = 2 * xv - xPrev[i];
        xBar[i] = xb;
        xPrev[i] = xv;
        zs = xb;
    } else {
        xb = xBar[i];
        zs = zsrc[i];
    }
    if (sharedSlot && sharedSlot[p] >= 0) return;
    const double val = (double)(incPtr[p + 1] - incPtr[p]);
    x[i] = xfirst_value(m, dtsq, w, val, xb, zs, usum[i], tdiag[p]);
}

// K7 for the vertices shared with other ranks only (so that their partial sums can be formed and pushed before the bulk pass)
template <int D> __global__ void k_predict_shared(int nShared, const int *sharedIdx, const double *x, double *xBar, double *xPrev) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nShared) return;
    const int p = sharedIdx[i];
#pragma unroll
    for (int l = 0; l < D; l++) {
        const double xv = x[(size_t)D * p + l];
        xBar[(size_t)D * p + l] = 2 * xv - xPrev[(size_t)D * p + l];
        xPrev[(size_t)D * p + l] = xv;
    }
}

// Isotropy of the monitor grid: flag stays 1 iff every row is m*I bitwise (equal diagonal, zero off-diagonal); the scalars
// go to `iso`.  All analytic monitors of the reference's driver except MEx2 are of this form (Experiments/TestMonitors).
template <int D> __global__ void k_iso_extract(size_t rows, const double *grid, double *iso, int *flag) {
    size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    const double *g = grid + r * GridRow<D>::stride;
    const double m = g[0];
    bool ok = true;
#pragma unroll
    for (int a = 0; a < D; a++)
#pragma unroll
        for (int b = 0; b < D; b++) ok = ok && (a == b ? (g[a * D + b] == m) : (g[a * D + b] == 0.0));
    iso[r] = m;
    if (!ok) *flag = 0;
}

// updateAfterStep with the projection of the BOUNDARY_FREE nodes enabled (src/Mesh.cpp:1020-1026): x_p <- projectOntoBoundary(p, x_p)
template <int D> __global__ void k_project_x(int P, const uint8_t *mask, FaceArgs fa, double *x) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P || mask[p] != NT_BOUNDARY_FREE) return;
    double pt[D];
#pragma unroll
    for (int l = 0; l < D; l++) pt[l] = x[(size_t)D * p + l];
    project_onto_boundary<D>(fa, p, pt);
#pragma unroll
    for (int l = 0; l < D; l++) x[(size_t)D * p + l] = pt[l];
}
__global__ void k_copy(size_t n, const double *src, double *dst) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
}
// x (P x D interleaved) -> column-major P x D (Eigen's default storage of the caller's Vp)
__global__ void k_x_to_colmajor(int P, int D, const double *x, double *xT) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    for (int l = 0; l < D; l++) xT[(size_t)l * P + p] = x[(size_t)D * p + l];
}

// ---- Method 2 (backward Euler) kernels ----
// K9: Mesh::buildEulerJac + FSubJac (src/Mesh.cpp:1111-1135, 1161-1261).  One thread per vertex p owns the D rows D p .. D p + D-1
// of the block-CSR Jacobian (vertex adjacency adjPtr/adjIdx, ascending, self included; row q of p starts at
// D D adjPtr[p] + q D deg(p), entry (t, c) = column D adjIdx[t] + c) and walks p's incident elements in ascending order as the
// reference's std::set does: Gk = blockGrad at Vp (a.x points at the Vp snapshot), then per direction i a forward difference
// with h = 10 sqrt(eps) that fills ROW i with d grad / d x_{p,i} (the reference stores the perturbed direction as the row).
// BOUNDARY_FIXED vertices: identity block, and only from the elements in which p is local vertex 0 (reference quirk, :1210-1218).
// Finally a *= dt/tau and the diagonal gets + 1.
template <int D>
__global__ void __launch_bounds__(128) k_m2_jac(ElemArgs<D> a, const uint8_t *mask, const int *incPtr, const int *incIdx,
                                                const int *adjPtr, const int *adjIdx, double dtOverTau, double *val) {
    constexpr int n = D * (D + 1);
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.P) return;
    const int a0 = adjPtr[p], deg = adjPtr[p + 1] - a0;
    double *rowBase = val + (size_t)D * D * a0;
    const int rowLen = D * deg;
    for (int k = 0; k < D * rowLen; k++) rowBase[k] = 0.0;
    const double h = 10.0 * 1.4901161193847656e-08;   // 10 * sqrt(2^-52)
    const bool fixed = mask[p] == NT_BOUNDARY_FIXED;
    bool inv = false;
    for (int k = incPtr[p]; k < incPtr[p + 1]; k++) {
        const int slot = incIdx[k];
        const int e = slot >> 2, off = slot & 3;
        int vid[D + 1], pos[D + 1];
        load_elem_ids<D>(a, e, vid);
#pragma unroll
        for (int j = 0; j < D + 1; j++) {
            int lo = 0, hi = deg - 1;
            while (lo < hi) {   // adjIdx[a0 .. a0+deg) ascending and contains vid[j]
                const int mid = (lo + hi) >> 1;
                if (adjIdx[a0 + mid] < vid[j]) lo = mid + 1; else hi = mid;
            }
            pos[j] = lo;
        }
        if (fixed) {
            if (off == 0)
                for (int r = 0; r < D; r++) rowBase[r * rowLen + D * pos[0] + r] += 1.0;
            continue;
        }
        double xl[n], xi[n], Gk[n], G1[n];
#pragma unroll
        for (int v = 0; v < D + 1; v++) gather_vertex<D>(a.x, vid[v], xl + D * v);
        load_xi<D>(a, vid, xi);
        block_grad<D, true>(xl, xi, a.compMesh, a.mg, a.fc, Gk, inv);
        for (int i = 0; i < D; i++) {
            double keep = 0.0;
#pragma unroll
            for (int q = 0; q < n; q++)
                if (q == D * off + i) { keep = xl[q]; xl[q] = keep + h; }
            block_grad<D, true>(xl, xi, a.compMesh, a.mg, a.fc, G1, inv);
#pragma unroll
            for (int q = 0; q < n; q++)
                if (q == D * off + i) xl[q] = keep;
#pragma unroll
            for (int j = 0; j < D + 1; j++)
#pragma unroll
                for (int c = 0; c < D; c++) rowBase[i * rowLen + D * pos[j] + c] += (G1[D * j + c] - Gk[D * j + c]) / h;
        }
    }
    if (inv) atomicOr(a.status, ST_INVERTED);
    for (int q = 0; q < D; q++)
        for (int t = 0; t < deg; t++)
            for (int c = 0; c < D; c++) {
                double v = rowBase[q * rowLen + D * t + c] * dtOverTau;
                if (adjIdx[a0 + t] == p && c == q) v += 1.0;
                rowBase[q * rowLen + D * t + c] = v;
            }
}

// K10: vertex side of Mesh::backwardsEulerStep (src/Mesh.cpp:1263-1341).  s = sum of the masked element gradients of the vertex
// (ascending element order).  MODE 0, the explicit-Euler guess: x -= (dt/tau) s.  MODE 1, the Newton residual:
// F = (dt/tau) s + (x - xn), rhs = -F (the right-hand side of J dx = -F), block partial sums of |F| (the reference's lpNorm<1>).
template <int D, int MODE>
__global__ void __launch_bounds__(TPB_V) k_m2_gather(int P, size_t Es, const int *incPtr, const int *incIdx, const double *gs,
                                                     double dtOverTau, double *x, const double *xn, double *rhs, double *part) {
    __shared__ double red[32];
    const int p = blockIdx.x * TPB_V + threadIdx.x;
    double l1 = 0.0;
    if (p < P) {
        double s[D];
#pragma unroll
        for (int l = 0; l < D; l++) s[l] = 0.0;
        for (int k = incPtr[p]; k < incPtr[p + 1]; k++) {
            const int slot = __ldg(incIdx + k);
            const size_t e = (size_t)(slot >> 2);
            const int v = slot & 3;
#pragma unroll
            for (int l = 0; l < D; l++) s[l] += __ldg(gs + (size_t)(D * v + l) * Es + e);
        }
#pragma unroll
        for (int l = 0; l < D; l++) {
            const size_t i = (size_t)D * p + l;
            if (MODE == 0) {
                x[i] -= dtOverTau * s[l];
            } else {
                const double f = s[l] * dtOverTau + (x[i] - xn[i]);
                rhs[i] = -f;
                l1 += fabs(f);
            }
        }
    }
    if (MODE == 1) {
        const double b = block_sum(l1, red);
        if (threadIdx.x == 0) part[blockIdx.x] = b;
    }
}

__global__ void k_m2_axpy(size_t N, const double *dx, double *x) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) x[i] += dx[i];
}

// ---- unit-parity helpers ----
template <int D> __global__ void k_eval_monitor(MonGrid mg, int npts, const double *pts, double *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npts) return;
    double p[D];
#pragma unroll
    for (int l = 0; l < D; l++) p[l] = pts[(size_t)i * D + l];
    SMat<D> m;
    eval_monitor_any<D>(mg, p, m);
#pragma unroll
    for (int k = 0; k < D * D; k++) out[(size_t)i * D * D + k] = m.a[k];
}
template <int D>
__global__ void k_block_grad_dbg(ElemArgs<D> a, int e0, int ne, const double *zs, const double *dxpu, int computeGrad,
                                 int regularize, int applyMask, double *energy, double *grad, double *igt) {
    constexpr int n = D * (D + 1);
    int e = blockIdx.x * blockDim.x + threadIdx.x;   // row of the caller's arrays; mesh element e0 + e
    if (e >= ne) return;
    int vid[D + 1];
    load_elem_ids<D>(a, e0 + e, vid);
    double z[n], xi[n], g[n], dx[n];
#pragma unroll
    for (int k = 0; k < n; k++) { z[k] = zs[(size_t)e * n + k]; dx[k] = regularize ? dxpu[(size_t)e * n + k] : 0.0; g[k] = 0.0; }
    load_xi<D>(a, vid, xi);
    bool inv = false;
    double en = computeGrad ? block_grad<D, true>(z, xi, a.compMesh, a.mg, a.fc, g, inv)
                            : block_grad<D, false>(z, xi, a.compMesh, a.mg, a.fc, g, inv);
    double I = en;
    if (regularize) {
        double sq = 0.0;
#pragma unroll
        for (int k = 0; k < n; k++) { double t = dx[k] - z[k]; sq += t * t; g[k] += a.fc.w2 * (-dx[k] + z[k]); }
        en += 0.5 * a.fc.w2 * sq;
    }
    if (applyMask) {
        unsigned fb = a.fixedBits[e0 + e];
#pragma unroll
        for (int v = 0; v < D + 1; v++)
            if (fb & (1u << v))
#pragma unroll
                for (int l = 0; l < D; l++) g[D * v + l] = 0.0;
    }
    energy[e] = en;
    if (igt) igt[e] = I;
    if (grad && computeGrad)
#pragma unroll
        for (int k = 0; k < n; k++) grad[(size_t)e * n + k] = g[k];
    if (inv) atomicOr(a.status, ST_INVERTED);
}

}  // namespace

// =====================================================================================================
// context
// =====================================================================================================
struct mmadmm_ctx {
    int D = 0, n = 0, P = 0, E = 0, device = 0;
    size_t Es = 0;               // element-plane stride
    int proxVariant = 0, xupVariant = 0;
    // ---- distributed (element-block partition, one process per GPU) ----
    bool dist = false;
    int rank = 0, nranks = 1, Pglobal = 0, Eglobal = 0;
    void *comm = nullptr;                       // ncclComm_t
    // NVLink peer-memory exchange (CUDA IPC): every rank owns a mailbox that its neighbours store into directly
    bool p2p = false;
    unsigned char *mailbox = nullptr;           // this rank's mailbox (device memory, exported through cudaIpcMemHandle)
    std::vector<void *> peerMap;                // opened mailboxes of the other ranks (null for self / unopened)
    double **peerRecv_d = nullptr;              // [nNbr] neighbour's receive segment for this rank (parity 0)
    size_t *peerRecvStride_d = nullptr;         // [nNbr] doubles between the two parity halves of that neighbour's receive area
    unsigned long long **peerHaloFlag_d = nullptr;   // [nNbr] neighbour's flag slot for this rank
    double **peerScal_d = nullptr;              // [nranks] every rank's scalar table row for this rank (parity 0)
    unsigned long long **peerScalFlag_d = nullptr;   // [nranks]
    unsigned long long *haloFlags_d = nullptr, *scalFlags_d = nullptr;   // inside the own mailbox
    double *mbScal_d = nullptr, *mbRecv_d = nullptr;                      // inside the own mailbox
    int *nbrPtr_d = nullptr;
    unsigned long long haloEpoch = 0, scalEpoch = 0;
    unsigned long long peerTimeoutNs = 300ull * 1000000000ull;   // bound of every wait on a peer's flag (option "peer_timeout_s")
    int *pushPtr_d = nullptr, *pushNbr_d = nullptr, *pushPos_d = nullptr;   // per shared vertex: (neighbour, position) it is pushed to
    unsigned *pushTicket_d = nullptr;
    int nShared = 0, nNbrItems = 0, nblkSh = 0;
    std::vector<int> nbrRank, nbrPtr, nbrListHost;   // host copies of the plan's neighbour segments
    int *sharedIdx = nullptr, *sharedSlot = nullptr, *nbrList = nullptr, *contribPtr = nullptr, *contrib = nullptr, *verts = nullptr;
    uint8_t *owner = nullptr;
    double *xloc_h = nullptr;
    std::vector<int> verts_h;
    std::vector<std::pair<int, int>> runs;   // (first local vertex, length) of consecutive global ids
    double *sPart = nullptr, *sTot = nullptr, *sendBuf = nullptr, *recvBuf = nullptr, *scalAll_d = nullptr, *scalAll_h = nullptr, *scalAll_hd = nullptr;
    int nProxPart = 0;           // number of per-block (or per-warp) partial sums the last prox launch wrote
    int numSMs = 148;
    bool smemAttrSet[16] = {false};
    bool compMesh = false, gradUse = false;
    cudaStream_t stream = nullptr;
    int *F[4] = {nullptr, nullptr, nullptr, nullptr};
    uint8_t *mask = nullptr, *fixedBits = nullptr, *nonIntBits = nullptr;
    double *x = nullptr, *xPrev = nullptr, *xBar = nullptr, *Vc = nullptr, *tdiag = nullptr;
    // single-GPU step tail (StepTail): the next step's xBar and first x-solve written by the last k_xupdate<UREC> of a step;
    // specValid: they belong to the state x / xPrev / usum / dt / rho / tau as it is now (anything that touches those clears it)
    double *xSpec = nullptr, *xBarSpec = nullptr;
    bool specValid = false;
    unsigned *finTicket = nullptr;
    // option "step_tail": bit 0 speculative first solve, bit 1 final reduction in k_xupdate.  Measured (scripts/gpu_r2_r.sh, C2
    // ms/step): 0: 0.770, 1: 0.759, 2: 0.776, 3: 0.771 -- the ticket + fences of the fused reduction cost the 4,100 CTAs of
    // k_xupdate more (+0.019 ms) than the k_finalize launch they replace (0.011 ms), so only bit 0 is on by default
    int tailMode = 1;
    long long specUsed = 0;
    double *z = nullptr, *u = nullptr, *B = nullptr, *gs = nullptr;
    int *incPtr = nullptr, *incIdx = nullptr, *vorder = nullptr, *valence_d = nullptr;
    double *grid = nullptr, *ax = nullptr, *ay = nullptr, *az = nullptr;
    double *xT = nullptr;               // column-major staging of x for mmadmm_get_x_colmajor
    double *gridIso = nullptr;          // scalar monitor grid (isotropic monitors only)
    bool gridIsIso = false, isoOn = true, l2Persist = true;
    int nx = 0, ny = 0, nz = 0;
    size_t gridRows = 0;
    double rho = 0, w = 0, tau = 0, m = 0, dt = 0;
    double Ehat[9] = {0}, dpow = 0;
    bool stepTaken = false, hessComputed = false;
    int stepsTaken = 0;
    int projMask = 0;         // option "boundary_projection" (FaceArgs::mask)
    int *facePtr_d = nullptr, *faceIdx_d = nullptr, *faceList_d = nullptr;
    int nFaces = 0;
    double *usum = nullptr;   // per-vertex sum of the dual over the vertex's slots (k_xupdate / k_xfirst)
    int usumPeriod = 64;      // every usumPeriod-th x-update after a prox sweep re-gathers usum exactly; 0: slot-by-slot x-updates only
    int usumAge = 0;
    int nblkE = 0, nblkV = 0, nblkProx = 0;
    double *partIh = nullptr, *partDual = nullptr, *partPrimal = nullptr, *scal_d = nullptr, *scal_h = nullptr, *scal_hd = nullptr;
    unsigned *status_d = nullptr, *status_h = nullptr, *status_hd = nullptr;   // *_hd: device-side address of the mapped host buffer
    unsigned long long *hist_d = nullptr;
    bool statsOn = true, timingOn = false;
    long long admmIters = 0, proxCalls = 0, launches = 0;
    double phaseMs[5] = {0, 0, 0, 0, 0};
    long long phaseCnt[5] = {0, 0, 0, 0, 0};
    std::vector<cudaEvent_t> evPool;
    std::vector<int> evPhase;
    size_t evUsed = 0;
    std::string err;
    // Method 2 (backward Euler) state, built on the first mmadmm_backward_euler_step
    struct M2 {
        bool init = false, stepTaken = false, jacValid = false;
        int N = 0, nnz = 0, nblk = 0, vpStamp = -1;
        double jacDt = 0.0;
        int *adjPtr_d = nullptr, *adjIdx_d = nullptr;
        std::vector<int> rowptr_h, col_h;
        std::vector<double> val_h;
        double *val_d = nullptr, *vp0 = nullptr, *xn = nullptr, *rhs = nullptr, *dx = nullptr, *part = nullptr;
        mmadmm_pcg *solver = nullptr;
        long long newtonIters = 0, linIters = 0, jacBuilds = 0;
        int triMode = 2;   // mmadmm_pcg "tri_mode" of the Method 2 solver (option "m2_tri_mode"): 2 = dataflow solve
        int rowBlock = -1; // mmadmm_pcg "row_block" of the Method 2 solver (option "m2_row_block"); -1 = D (one vertex per thread)
        int triLayout = 1; // mmadmm_pcg "tri_layout" of the Method 2 solver (option "m2_tri_layout"): 1 = level-ELL copy of the factor
    } m2;

    // checkpoint of the whole integrator state (mmadmm_checkpoint / mmadmm_restore)
    struct Ckpt {
        bool valid = false, stepTaken = false, hessComputed = false;
        int stepsTaken = 0, usumAge = 0;
        long long admmIters = 0, proxCalls = 0;
        double *x = nullptr, *xPrev = nullptr, *xBar = nullptr, *z = nullptr, *u = nullptr, *B = nullptr, *usum = nullptr;
        unsigned long long *hist = nullptr;
    } ck;
    // write-back of the node positions into the caller's host array after every step (Mesh::updateAfterStep, src/Mesh.cpp:1030-1035)
    struct Writeback {
        int mode = 0;                 // 0 off, 1 strict (synchronous, the reference's semantics), 2 deferred (overlapped, fenced)
        int colMajor = 0;             // 1: caller's array is column-major P x D (Eigen::MatrixXd Vp)
        double *host = nullptr;       // caller-owned; GLOBAL P x D on a distributed context (this rank writes its own rows)
        double *stage[2] = {nullptr, nullptr};
        cudaStream_t copyStream = nullptr;
        cudaEvent_t evSnap[2] = {nullptr, nullptr}, evCopied[2] = {nullptr, nullptr};
        bool inFlight[2] = {false, false};
        int next = 0;
        long long copies = 0;
        // MMADMM_DEBUG_WB=1 / MMADMM_WB_SNAP=ce (experiments)
        bool debug = getenv("MMADMM_DEBUG_WB") != nullptr;
        bool snapByCE = getenv("MMADMM_WB_SNAP") && !strcmp(getenv("MMADMM_WB_SNAP"), "ce");
        cudaEvent_t evCopyStart[2] = {nullptr, nullptr};
        long long dbgNs[7] = {0, 0, 0, 0, 0, 0, 0}, dbgMax[7] = {0, 0, 0, 0, 0, 0, 0};
        double dbgCopyMs = 0, dbgCopyMax = 0;
        long long dbgCopyN = 0, dbgSyncNs[3] = {0, 0, 0}, dbgSyncN[3] = {0, 0, 0};
    } wb;

    double gridO[3] = {0, 0, 0}, gridRh[3] = {0, 0, 0};
    double locMin[3] = {0, 0, 0}, locMax[3] = {-1, -1, -1};   // bounding box of this rank's vertices (create_impl)
    MonGrid monGrid() const {
        MonGrid mg;
        mg.vals = grid; mg.iso = (isoOn && gridIsIso) ? gridIso : nullptr; mg.ax = ax; mg.ay = ay; mg.az = az; mg.nx = nx; mg.ny = ny; mg.nz = nz;
        for (int l = 0; l < 3; l++) { mg.o[l] = gridO[l]; mg.rh[l] = gridRh[l]; }
        return mg;
    }
    template <int D_> ElemArgs<D_> elemArgs(double tol) const {
        ElemArgs<D_> a;
        a.E = E; a.P = P; a.Es = Es; a.zsrc = nullptr; a.uZero = 0;
        for (int v = 0; v < MAXV; v++) a.F[v] = F[v];
        a.fixedBits = fixedBits; a.nonIntBits = nonIntBits;
        a.x = x; a.Vc = Vc; a.z = z; a.u = u; a.B = B;
        a.mg = monGrid();
        for (int k = 0; k < D_ * D_; k++) a.fc.Ehat[k] = Ehat[k];
        a.fc.dpow = dpow; a.fc.w2 = w * w;
        a.compMesh = compMesh ? 1 : 0;
        a.tol = tol;
        a.partIh = partIh; a.partDual = partDual;
        a.hist = statsOn ? hist_d : nullptr;
        a.status = status_d;
        a.fa.ptr = facePtr_d; a.fa.idx = faceIdx_d; a.fa.list = faceList_d; a.fa.vp = xPrev; a.fa.mask = projMask;
        return a;
    }
    void phaseBegin(int ph) {
        if (!timingOn) return;
        if (evUsed + 2 > evPool.size()) {
            for (int i = 0; i < 64; i++) { cudaEvent_t e; cudaEventCreate(&e); evPool.push_back(e); evPhase.push_back(0); }
        }
        evPhase[evUsed] = ph;
        cudaEventRecord(evPool[evUsed], stream);
    }
    void phaseEnd() {
        if (!timingOn) return;
        cudaEventRecord(evPool[evUsed + 1], stream);
        evUsed += 2;
    }
    // Elapsed times are read when somebody asks (mmadmm_get_timing) or when the pool is large, not after every step: the
    // synchronise + query per phase would sit in the host gap between two steps.
    void phaseResolve() {
        if (evUsed == 0) return;
        for (size_t i = 0; i + 1 < evUsed; i += 2) {
            float ms = 0;
            cudaEventSynchronize(evPool[i + 1]);
            cudaEventElapsedTime(&ms, evPool[i], evPool[i + 1]);
            phaseMs[evPhase[i]] += ms;
            phaseCnt[evPhase[i]]++;
        }
        evUsed = 0;
    }
};

namespace {

inline int cdiv(size_t a, size_t b) { return (int)((a + b - 1) / b); }

// ---- NCCL, loaded at run time (single-GPU use of this library needs no NCCL at all) ----
struct NcclApi {
    bool ok = false;
    int (*getUniqueId)(void *) = nullptr;
    void *commInitRank = nullptr;
    int (*commDestroy)(void *) = nullptr;
    int (*groupStart)() = nullptr;
    int (*groupEnd)() = nullptr;
    int (*send)(const void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    int (*recv)(void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    int (*allGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
    int (*allReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    const char *(*getErrorString)(int) = nullptr;
};
struct NcclIdBlob { char b[128]; };
constexpr int NCCL_FLOAT64 = 8, NCCL_SUM = 0;   // ncclDataType_t / ncclRedOp_t values (nccl.h)

NcclApi &nccl() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api;
    tried = true;
    // MMADMM_NCCL_LIB (a path) first, then the sonames; a libnccl the process has already loaded (e.g. torch's) is found by soname
    const char *cands[] = {getenv("MMADMM_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *c : cands) {
        if (!c || !c[0]) continue;
        h = dlopen(c, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) return api;
    api.getUniqueId = (int (*)(void *))dlsym(h, "ncclGetUniqueId");
    api.commInitRank = dlsym(h, "ncclCommInitRank");
    api.commDestroy = (int (*)(void *))dlsym(h, "ncclCommDestroy");
    api.groupStart = (int (*)())dlsym(h, "ncclGroupStart");
    api.groupEnd = (int (*)())dlsym(h, "ncclGroupEnd");
    api.send = (int (*)(const void *, size_t, int, int, void *, cudaStream_t))dlsym(h, "ncclSend");
    api.recv = (int (*)(void *, size_t, int, int, void *, cudaStream_t))dlsym(h, "ncclRecv");
    api.allGather = (int (*)(const void *, void *, size_t, int, void *, cudaStream_t))dlsym(h, "ncclAllGather");
    api.allReduce = (int (*)(const void *, void *, size_t, int, int, void *, cudaStream_t))dlsym(h, "ncclAllReduce");
    api.getErrorString = (const char *(*)(int))dlsym(h, "ncclGetErrorString");
    api.ok = api.getUniqueId && api.commInitRank && api.commDestroy && api.groupStart && api.groupEnd && api.send && api.recv &&
             api.allGather && api.allReduce;
    return api;
}

#define NK(call)                                                                                       \
    do {                                                                                               \
        int _r = (call);                                                                               \
        if (_r != 0) {                                                                                 \
            char _b[512];                                                                              \
            snprintf(_b, sizeof(_b), "%s:%d: %s -> NCCL error %d (%s)", __FILE__, __LINE__, #call, _r, \
                     nccl().getErrorString ? nccl().getErrorString(_r) : "?");                         \
            ctx->err = _b;                                                                             \
            return MMADMM_ENCCL;                                                                       \
        }                                                                                              \
    } while (0)

// ---- shared-vertex partial-sum exchange (see partition.h) ----
template <int D> __global__ void k_halo_pack(int nItems, const int *nbrList, const double *sPart, double *sendBuf) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nItems) return;
    const int i = nbrList[j];
#pragma unroll
    for (int l = 0; l < D; l++) sendBuf[(size_t)j * D + l] = sPart[(size_t)i * D + l];
}
// total = sum of the sharers' partials in ascending rank order (identical on every sharer)
template <int D>
__global__ void k_halo_combine(int nShared, const int *contribPtr, const int *contrib, const double *sPart, const double *recvBuf,
                               double *sTot) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nShared) return;
    double s[D];
#pragma unroll
    for (int l = 0; l < D; l++) s[l] = 0.0;
    for (int k = contribPtr[i]; k < contribPtr[i + 1]; k++) {
        const int c = contrib[k];
        const double *src = (c < 0) ? (sPart + (size_t)i * D) : (recvBuf + (size_t)c * D);
#pragma unroll
        for (int l = 0; l < D; l++) s[l] += src[l];
    }
#pragma unroll
    for (int l = 0; l < D; l++) sTot[(size_t)i * D + l] = s[l];
}


// ---- NVLink peer-memory exchange ----------------------------------------------------------------------------------
// Flags are monotonically increasing epochs written with system scope after the data; readers spin (bounded) on them.
// One block per neighbour: store this rank's partials of the vertices shared with neighbour q straight into q's mailbox
// (receive half `epoch & 1`), then publish the epoch in q's flag slot for this rank.
template <int D>
__global__ void __launch_bounds__(1024) k_halo_push(const int *nbrPtr, const int *nbrList, const double *sPart, double *const *peerRecv,
                                                    const size_t *peerStride, unsigned long long *const *peerFlag,
                                                    unsigned long long epoch) {
    const int q = blockIdx.x;
    const int j0 = nbrPtr[q], j1 = nbrPtr[q + 1];
    double *dst = peerRecv[q] + ((epoch & 1ull) ? peerStride[q] : 0);
    for (int j = j0 + threadIdx.x; j < j1; j += blockDim.x) {
        const int i = nbrList[j];
#pragma unroll
        for (int l = 0; l < D; l++) dst[(size_t)(j - j0) * D + l] = sPart[(size_t)i * D + l];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) flag_release(peerFlag[q], epoch);
}
// total = sum of the sharers' partials in ascending rank order, reading the partials the neighbours pushed
template <int D>
__device__ __forceinline__ void halo_wait_all(int nNbr, const unsigned long long *flags, unsigned long long epoch,
                                              unsigned long long timeoutNs, unsigned *status) {
    for (int q = threadIdx.x; q < nNbr; q += blockDim.x) flag_wait(flags + q, epoch, timeoutNs, status);
    __syncthreads();
}
template <int D>
__device__ __forceinline__ void halo_total(int i, const int *contribPtr, const int *contrib, const double *sPart, const double *recvBuf,
                                           double *s) {
#pragma unroll
    for (int l = 0; l < D; l++) s[l] = 0.0;
    for (int k = contribPtr[i]; k < contribPtr[i + 1]; k++) {
        const int c = contrib[k];
#pragma unroll
        for (int l = 0; l < D; l++) s[l] += (c < 0) ? sPart[(size_t)i * D + l] : __ldcg(recvBuf + (size_t)c * D + l);
    }
}
template <int D>
__global__ void k_halo_combine_p2p(int nShared, int nNbr, const int *contribPtr, const int *contrib, const double *sPart,
                                   const double *recvBase, size_t recvStride, const unsigned long long *flags,
                                   unsigned long long epoch, double *sTot, unsigned long long timeoutNs, unsigned *status) {
    halo_wait_all<D>(nNbr, flags, epoch, timeoutNs, status);
    const double *recvBuf = recvBase + ((epoch & 1ull) ? recvStride : 0);
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nShared) return;
    double s[D];
    halo_total<D>(i, contribPtr, contrib, sPart, recvBuf, s);
#pragma unroll
    for (int l = 0; l < D; l++) sTot[(size_t)i * D + l] = s[l];
}
// Fused k_halo_combine_p2p + k_xfinish: wait for the neighbours' epoch, sum the sharers' partials in ascending rank order
// (identical on every sharer), finish x of the shared vertices and their share of ||D x - z||^2 over the LOCAL slots.
template <int D, bool ZSRC, bool PRIMAL>
__global__ void __launch_bounds__(TPB_V) k_xfinish_p2p(int nShared, int nNbr, const int *sharedIdx, const int *contribPtr, const int *contrib,
                                                       const double *sPart, const double *recvBase, size_t recvStride,
                                                       const unsigned long long *flags, unsigned long long epoch,
                                                       unsigned long long timeoutNs, unsigned *status, const int *incPtr,
                                                       const int *incIdx, const double *z, const double *zsrc, const double *xBar,
                                                       const double *tdiag, double m, double dtsq, double *x, double *partPrimal) {
    __shared__ double red[32];
    halo_wait_all<D>(nNbr, flags, epoch, timeoutNs, status);
    const double *recvBuf = recvBase + ((epoch & 1ull) ? recvStride : 0);
    const int i = blockIdx.x * TPB_V + threadIdx.x;
    double prim = 0.0;
    if (i < nShared) {
        double s[D], xn[D];
        halo_total<D>(i, contribPtr, contrib, sPart, recvBuf, s);
        const int p = sharedIdx[i];
        const double td = tdiag[p];
#pragma unroll
        for (int l = 0; l < D; l++) {
            xn[l] = ((m * xBar[(size_t)D * p + l]) + dtsq * s[l]) / td;
            x[(size_t)D * p + l] = xn[l];
        }
        if (PRIMAL) {
            for (int k = incPtr[p]; k < incPtr[p + 1]; k++) {
                const int slot = __ldg(incIdx + k);
                const size_t e = (size_t)(slot >> 2);
                const int v = slot & 3;
#pragma unroll
                for (int l = 0; l < D; l++) {
                    const double zv = ZSRC ? zsrc[(size_t)D * p + l] : __ldg(z + zu_index<D>(e, D * v + l));
                    const double df = xn[l] - zv;
                    prim += df * df;
                }
            }
        }
    }
    if (PRIMAL) {
        double r = block_sum(prim, red);
        if (threadIdx.x == 0) partPrimal[blockIdx.x] = r;
    }
}
// Fused k_finalize + all-gather of the per-rank scalars through the mailboxes + sum in rank order (one CTA): the fixed-order
// reduction of up to 3 partial arrays into scal[off..], then thread r stores this rank's row {scal[0..3], status} into rank r's
// table and publishes the epoch, waits for rank r's row, and thread 0 adds the rows in ascending rank order -- identical on every
// rank, so every rank takes the same break decision (src/MeshIntegrator.cpp:162-169) -- into mapped host memory.
__global__ void __launch_bounds__(1024) k_finalize_dist(int n0, const double *p0, int n1, const double *p1, int n2, const double *p2,
                                                        double *scal, int off, unsigned *status, int nranks, double *const *peerScal,
                                                        unsigned long long *const *peerFlag, const double *mbScal,
                                                        const unsigned long long *flags, unsigned long long epoch,
                                                        unsigned long long timeoutNs, double *scalHost, unsigned *statusHost) {
    __shared__ double red[32];
    __shared__ double rows[32 * 5];
    const double *ps[3] = {p0, p1, p2};
    const int ns[3] = {n0, n1, n2};
    for (int q = 0; q < 3; q++) {
        if (!ps[q]) continue;
        double s = 0.0;
        for (int i = threadIdx.x; i < ns[q]; i += blockDim.x) s += ps[q][i];
        double r = block_sum(s, red);
        if (threadIdx.x == 0) scal[off + q] = r;
        __syncthreads();
    }
    const int r = threadIdx.x;
    if (r < nranks) {
        double *dst = peerScal[r] + ((epoch & 1ull) ? (size_t)nranks * 8 : 0);
#pragma unroll
        for (int k = 0; k < 4; k++) dst[k] = scal[k];
        dst[4] = (double)(*(volatile unsigned *)status & ~(unsigned)ST_PEER_TIMEOUT);
        flag_release(peerFlag[r], epoch);
        const bool ok = flag_wait(flags + r, epoch, timeoutNs, status);
        const double *src = mbScal + ((epoch & 1ull) ? (size_t)nranks * 8 : 0) + (size_t)r * 8;
#pragma unroll
        for (int k = 0; k < 5; k++) rows[5 * r + k] = ok ? __ldcg(src + k) : 0.0;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned st = *(volatile unsigned *)status & ST_PEER_TIMEOUT;
        double acc[4] = {0.0, 0.0, 0.0, 0.0};
        for (int q = 0; q < nranks; q++) {
#pragma unroll
            for (int k = 0; k < 4; k++) acc[k] += rows[5 * q + k];
            st |= (unsigned)rows[5 * q + 4];
        }
#pragma unroll
        for (int k = 0; k < 4; k++) scalHost[k] = acc[k];
        *statusHost = st;
        __threadfence_system();
    }
}

// sPart (partials of the shared vertices) -> sTot (totals): NVLink peer stores (default) or NCCL point-to-point
template <int D> int halo_exchange(mmadmm_ctx *ctx) {
    if (!ctx->dist || ctx->nShared == 0) return MMADMM_OK;
    cudaStream_t st = ctx->stream;
    if (ctx->p2p) {
        const unsigned long long ep = ++ctx->haloEpoch;
        const int nNbr = (int)ctx->nbrRank.size();
        k_halo_push<D><<<nNbr, 1024, 0, st>>>(ctx->nbrPtr_d, ctx->nbrList, ctx->sPart, ctx->peerRecv_d, ctx->peerRecvStride_d,
                                              ctx->peerHaloFlag_d, ep);
        k_halo_combine_p2p<D><<<cdiv(ctx->nShared, 256), 256, 0, st>>>(ctx->nShared, nNbr, ctx->contribPtr, ctx->contrib, ctx->sPart,
                                                                      ctx->mbRecv_d, (size_t)std::max(ctx->nNbrItems, 1) * D,
                                                                      ctx->haloFlags_d, ep, ctx->sTot, ctx->peerTimeoutNs, ctx->status_d);
        ctx->launches += 2;
        return MMADMM_OK;
    }
    k_halo_pack<D><<<cdiv(ctx->nNbrItems, 256), 256, 0, st>>>(ctx->nNbrItems, ctx->nbrList, ctx->sPart, ctx->sendBuf);
    NcclApi &N = nccl();
    NK(N.groupStart());
    for (size_t q = 0; q < ctx->nbrRank.size(); q++) {
        const size_t off = (size_t)ctx->nbrPtr[q] * D, cnt = (size_t)(ctx->nbrPtr[q + 1] - ctx->nbrPtr[q]) * D;
        NK(N.send(ctx->sendBuf + off, cnt, NCCL_FLOAT64, ctx->nbrRank[q], ctx->comm, st));
        NK(N.recv(ctx->recvBuf + off, cnt, NCCL_FLOAT64, ctx->nbrRank[q], ctx->comm, st));
    }
    NK(N.groupEnd());
    k_halo_combine<D><<<cdiv(ctx->nShared, 256), 256, 0, st>>>(ctx->nShared, ctx->contribPtr, ctx->contrib, ctx->sPart, ctx->recvBuf,
                                                               ctx->sTot);
    ctx->launches += 2;
    return MMADMM_OK;
}

// Fixed-order reduction of up to three partial arrays into scal[off .. off+2], then the scalars (+ the status word) where the
// host reads them after one stream synchronise: single GPU -- k_finalize writes mapped host memory; NVLink peer path --
// k_finalize_dist also exchanges the per-rank rows through the mailboxes and adds them in rank order on the device; NCCL
// fallback -- ncclAllGather + host sum.  The break decision of the ADMM loop is identical on all ranks
// (src/MeshIntegrator.cpp:162-169).
int finalize_scalars(mmadmm_ctx *ctx, int n0, const double *p0, int n1, const double *p1, int n2, const double *p2, int off) {
    cudaStream_t st = ctx->stream;
    if (!ctx->dist) {
        k_finalize<<<1, 1024, 0, st>>>(n0, p0, n1, p1, n2, p2, ctx->scal_hd + off, ctx->status_d, ctx->status_hd);
        ctx->launches += 1;
        if (ctx->wb.debug) {
            timespec t0, t1;
            clock_gettime(CLOCK_MONOTONIC, &t0);
            CK(cudaStreamSynchronize(st));
            clock_gettime(CLOCK_MONOTONIC, &t1);
            ctx->wb.dbgSyncNs[ctx->wb.mode] += (t1.tv_sec - t0.tv_sec) * 1000000000ll + (t1.tv_nsec - t0.tv_nsec);
            ctx->wb.dbgSyncN[ctx->wb.mode]++;
            return MMADMM_OK;
        }
        CK(cudaStreamSynchronize(st));
        return MMADMM_OK;
    }
    if (ctx->p2p) {
        const unsigned long long ep = ++ctx->scalEpoch;
        k_finalize_dist<<<1, 1024, 0, st>>>(n0, p0, n1, p1, n2, p2, ctx->scal_d, off, ctx->status_d, ctx->nranks, ctx->peerScal_d,
                                            ctx->peerScalFlag_d, ctx->mbScal_d, ctx->scalFlags_d, ep, ctx->peerTimeoutNs, ctx->scal_hd,
                                            ctx->status_hd);
        ctx->launches += 1;
        CK(cudaStreamSynchronize(st));
        return MMADMM_OK;
    }
    k_finalize<<<1, 1024, 0, st>>>(n0, p0, n1, p1, n2, p2, ctx->scal_d + off, nullptr, nullptr);
    k_status_to_double<<<1, 1, 0, st>>>(ctx->status_d, ctx->scal_d + 4);
    NK(nccl().allGather(ctx->scal_d, ctx->scalAll_d, 5, NCCL_FLOAT64, ctx->comm, st));
    CK(cudaMemcpyAsync(ctx->scalAll_h, ctx->scalAll_d, sizeof(double) * 5 * ctx->nranks, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    unsigned stat = 0;
    for (int k = 0; k < 4; k++) ctx->scal_h[k] = 0.0;
    for (int r = 0; r < ctx->nranks; r++) {
        for (int k = 0; k < 4; k++) ctx->scal_h[k] += ctx->scalAll_h[5 * r + k];
        stat |= (unsigned)ctx->scalAll_h[5 * r + 4];
    }
    *ctx->status_h = stat;
    ctx->launches += 3;
    return MMADMM_OK;
}


// NVLink peer-memory exchange set-up (collective over the communicator).  Every rank exports a mailbox through CUDA IPC;
// the handles and the neighbour tables travel once through ncclAllGather; afterwards halo sums and the scalar all-gather
// are plain stores into the neighbours' mailboxes + epoch flags (k_halo_push / k_scal_push), no NCCL call per iteration.
// Falls back to the NCCL path on every rank if any rank cannot take part (MMADMM_NO_P2P=1, too many neighbours, IPC
// refused by the platform).
constexpr int P2P_MAXNBR = 16, P2P_MAXRANKS = 32, P2P_TAB = 48;
constexpr size_t MB_OFF_HALOFLAG = 0, MB_OFF_SCALFLAG = 128, MB_OFF_SCAL = 384;
inline size_t mb_off_recv(int nranks) { return MB_OFF_SCAL + (size_t)2 * nranks * 64; }

int setup_p2p(mmadmm_ctx *ctx, int D) {
    cudaStream_t st = ctx->stream;
    NcclApi &N = nccl();
    const int nranks = ctx->nranks, me = ctx->rank, nNbr = (int)ctx->nbrRank.size();
    const char *env = getenv("MMADMM_NO_P2P");
    bool can = !(env && env[0] == '1') && nNbr <= P2P_MAXNBR && nranks <= P2P_MAXRANKS;
    const size_t recvHalf = (size_t)std::max(ctx->nNbrItems, 1) * D;
    const size_t mbBytes = mb_off_recv(nranks) + sizeof(double) * 2 * recvHalf;
    cudaIpcMemHandle_t hnd;
    memset(&hnd, 0, sizeof(hnd));
    if (can) {
        if (cudaMalloc(&ctx->mailbox, mbBytes) != cudaSuccess || cudaMemset(ctx->mailbox, 0, mbBytes) != cudaSuccess ||
            cudaIpcGetMemHandle(&hnd, ctx->mailbox) != cudaSuccess) {
            cudaGetLastError();
            can = false;
        }
    }
    CK(cudaDeviceSynchronize());
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    std::vector<double> tab(P2P_TAB, 0.0), all((size_t)P2P_TAB * nranks, 0.0);
    tab[0] = can ? 1.0 : 0.0; tab[1] = nNbr; tab[2] = ctx->nNbrItems;
    for (int q = 0; q < nNbr && q < P2P_MAXNBR; q++) { tab[3 + q] = ctx->nbrRank[q]; tab[19 + q] = ctx->nbrPtr[q]; }
    memcpy(&tab[40], &hnd, 64);
    double *tab_d = nullptr, *all_d = nullptr;
    CK(cudaMalloc(&tab_d, sizeof(double) * P2P_TAB));
    CK(cudaMalloc(&all_d, sizeof(double) * P2P_TAB * nranks));
    auto allgather = [&]() -> int {
        CK(cudaMemcpyAsync(tab_d, tab.data(), sizeof(double) * P2P_TAB, cudaMemcpyHostToDevice, st));
        NK(N.allGather(tab_d, all_d, P2P_TAB, NCCL_FLOAT64, ctx->comm, st));
        CK(cudaMemcpyAsync(all.data(), all_d, sizeof(double) * P2P_TAB * nranks, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        return MMADMM_OK;
    };
    int rc = allgather();
    if (rc) return rc;
    bool every = true;
    for (int r = 0; r < nranks; r++) every = every && all[(size_t)r * P2P_TAB] == 1.0;
    ctx->peerMap.assign(nranks, nullptr);
    bool opened = every;
    if (every) {
        for (int r = 0; r < nranks && opened; r++) {
            if (r == me) { ctx->peerMap[r] = ctx->mailbox; continue; }
            cudaIpcMemHandle_t h;
            memcpy(&h, &all[(size_t)r * P2P_TAB + 40], 64);
            if (cudaIpcOpenMemHandle(&ctx->peerMap[r], h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                cudaGetLastError();
                ctx->peerMap[r] = nullptr;
                opened = false;
            }
        }
    }
    // consensus on the outcome (second, tiny collective: reuse slot 0 of the table)
    std::vector<double> keep(all);
    tab[0] = opened ? 1.0 : 0.0;
    rc = allgather();
    if (rc) return rc;
    bool ok = true;
    for (int r = 0; r < nranks; r++) ok = ok && all[(size_t)r * P2P_TAB] == 1.0;
    cudaFree(tab_d); cudaFree(all_d);
    if (!ok) {
        for (int r = 0; r < nranks; r++)
            if (r != me && ctx->peerMap[r]) cudaIpcCloseMemHandle(ctx->peerMap[r]);
        ctx->peerMap.assign(nranks, nullptr);
        if (ctx->mailbox) { cudaFree(ctx->mailbox); ctx->mailbox = nullptr; }
        ctx->p2p = false;
        return MMADMM_OK;
    }
    all = keep;
    std::vector<double *> peerRecv(std::max(nNbr, 1), nullptr), peerScal(nranks, nullptr);
    std::vector<size_t> peerStride(std::max(nNbr, 1), 0);
    std::vector<unsigned long long *> peerHF(std::max(nNbr, 1), nullptr), peerSF(nranks, nullptr);
    for (int q = 0; q < nNbr; q++) {
        const int r = ctx->nbrRank[q];
        const double *tr = &all[(size_t)r * P2P_TAB];
        int qq = -1;
        for (int k = 0; k < (int)tr[1]; k++) if ((int)tr[3 + k] == me) qq = k;
        if (qq < 0) { ctx->err = "asymmetric neighbour lists in the partition plan"; return MMADMM_EINVAL; }
        unsigned char *base = (unsigned char *)ctx->peerMap[r];
        peerRecv[q] = (double *)(base + mb_off_recv(nranks)) + (size_t)tr[19 + qq] * D;
        peerStride[q] = (size_t)std::max((int)tr[2], 1) * D;
        peerHF[q] = (unsigned long long *)(base + MB_OFF_HALOFLAG) + qq;
    }
    for (int r = 0; r < nranks; r++) {
        unsigned char *base = (unsigned char *)ctx->peerMap[r];
        peerScal[r] = (double *)(base + MB_OFF_SCAL) + (size_t)me * 8;
        peerSF[r] = (unsigned long long *)(base + MB_OFF_SCALFLAG) + me;
    }
    auto up = [&](void **dst, const void *src, size_t bytes) -> cudaError_t {
        cudaError_t e = cudaMalloc(dst, std::max<size_t>(bytes, 8));
        if (e != cudaSuccess) return e;
        return cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice);
    };
    CK(up((void **)&ctx->peerRecv_d, peerRecv.data(), sizeof(double *) * peerRecv.size()));
    CK(up((void **)&ctx->peerRecvStride_d, peerStride.data(), sizeof(size_t) * peerStride.size()));
    CK(up((void **)&ctx->peerHaloFlag_d, peerHF.data(), sizeof(void *) * peerHF.size()));
    CK(up((void **)&ctx->peerScal_d, peerScal.data(), sizeof(double *) * nranks));
    CK(up((void **)&ctx->peerScalFlag_d, peerSF.data(), sizeof(void *) * nranks));
    CK(up((void **)&ctx->nbrPtr_d, ctx->nbrPtr.data(), sizeof(int) * ctx->nbrPtr.size()));
    {   // per shared vertex: the (neighbour, position in that neighbour's receive segment) pairs its partial sum is stored to
        // by the kernel that computes it (fused halo push of k_xupdate)
        std::vector<int> pptr((size_t)ctx->nShared + 1, 0), pnbr(ctx->nbrListHost.size()), ppos(ctx->nbrListHost.size());
        for (int i : ctx->nbrListHost) pptr[(size_t)i + 1]++;
        for (int i = 0; i < ctx->nShared; i++) pptr[(size_t)i + 1] += pptr[i];
        std::vector<int> cur(pptr.begin(), pptr.end() - 1);
        for (int q = 0; q < nNbr; q++)
            for (int j = ctx->nbrPtr[q]; j < ctx->nbrPtr[q + 1]; j++) {
                const int i = ctx->nbrListHost[j];
                pnbr[cur[i]] = q; ppos[cur[i]] = j - ctx->nbrPtr[q];
                cur[i]++;
            }
        CK(up((void **)&ctx->pushPtr_d, pptr.data(), sizeof(int) * pptr.size()));
        CK(up((void **)&ctx->pushNbr_d, pnbr.data(), sizeof(int) * pnbr.size()));
        CK(up((void **)&ctx->pushPos_d, ppos.data(), sizeof(int) * ppos.size()));
        CK(cudaMalloc(&ctx->pushTicket_d, sizeof(unsigned)));
        CK(cudaMemset(ctx->pushTicket_d, 0, sizeof(unsigned)));
    }
    ctx->haloFlags_d = (unsigned long long *)(ctx->mailbox + MB_OFF_HALOFLAG);
    ctx->scalFlags_d = (unsigned long long *)(ctx->mailbox + MB_OFF_SCALFLAG);
    ctx->mbScal_d = (double *)(ctx->mailbox + MB_OFF_SCAL);
    ctx->mbRecv_d = (double *)(ctx->mailbox + mb_off_recv(nranks));
    ctx->p2p = true;
    return MMADMM_OK;
}

// CSR by key: sorted vals (stable => ascending input order inside a key) and ptr[nkeys+1]
int build_csr(mmadmm_ctx *ctx, size_t nItems, int nKeys, int *keys, int *vals, int **ptrOut, int **valsOut) {
    cudaStream_t st = ctx->stream;
    int *cnt = nullptr, *ptr = nullptr, *keysOut = nullptr, *valsSorted = nullptr;
    CK(cudaMalloc(&cnt, sizeof(int) * ((size_t)nKeys + 1)));
    CK(cudaMalloc(&ptr, sizeof(int) * ((size_t)nKeys + 1)));
    CK(cudaMalloc(&keysOut, sizeof(int) * nItems));
    CK(cudaMalloc(&valsSorted, sizeof(int) * nItems));
    CK(cudaMemsetAsync(cnt, 0, sizeof(int) * ((size_t)nKeys + 1), st));
    k_count_keys<<<cdiv(nItems, 256), 256, 0, st>>>(nItems, keys, cnt);
    void *tmp = nullptr;
    size_t tmpBytes = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmpBytes, cnt, ptr, nKeys + 1, st));
    CK(cudaMalloc(&tmp, tmpBytes));
    CK(cub::DeviceScan::ExclusiveSum(tmp, tmpBytes, cnt, ptr, nKeys + 1, st));
    CK(cudaFree(tmp));
    int endBit = 1;
    while ((1ll << endBit) < (long long)nKeys) endBit++;
    tmp = nullptr; tmpBytes = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, keys, keysOut, vals, valsSorted, (int)nItems, 0, endBit, st));
    CK(cudaMalloc(&tmp, tmpBytes));
    CK(cub::DeviceRadixSort::SortPairs(tmp, tmpBytes, keys, keysOut, vals, valsSorted, (int)nItems, 0, endBit, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaFree(tmp));
    CK(cudaFree(cnt));
    CK(cudaFree(keysOut));
    *ptrOut = ptr;
    *valsOut = valsSorted;
    ctx->launches += 4;
    return MMADMM_OK;
}

// VpHost / monVertexHost: the P vertices the grid is built from (the GLOBAL mesh on a distributed context)
template <int D> int build_monitor_grid(mmadmm_ctx *ctx, int P, const double *VpHost, const double *monVertexHost) {
    cudaStream_t st = ctx->stream;
    constexpr int DD = D * D;
    double *xg = nullptr;
    CK(cudaMalloc(&xg, sizeof(double) * (size_t)P * D));
    CK(cudaMemcpyAsync(xg, VpHost, sizeof(double) * (size_t)P * D, cudaMemcpyHostToDevice, st));
    // MeshInterpolator::updateMesh (src/MeshInterpolator.cpp:67-130): grid size, bounding box, axes (host, init only)
    ctx->nx = (int)std::pow((double)((size_t)P * D), 1.0 / D);
    ctx->ny = ctx->nx;
    ctx->nz = (D == 2) ? 1 : ctx->nx;
    double mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int p = 0; p < P; p++)
        for (int l = 0; l < D; l++) {
            double v = VpHost[(size_t)p * D + l];
            mn[l] = (v < mn[l]) ? v : mn[l];
            mx[l] = (v > mx[l]) ? v : mx[l];
        }
    auto linspace = [](double xa, double xb, int ns, std::vector<double> &x) {  // src/MeshUtils.h:24-29
        x.resize(ns + 1);
        for (int i = 0; i < ns + 1; i++) x[i] = xa + ((double)i) * (xb - xa) / ns;
    };
    std::vector<double> hx, hy, hz;
    linspace(mn[0], mx[0], ctx->nx, hx);
    linspace(mn[1], mx[1], ctx->ny, hy);
    if (D == 3) linspace(mn[2], mx[2], ctx->nz, hz); else hz.assign(2, 0.0);
    CK(cudaMalloc(&ctx->ax, sizeof(double) * hx.size()));
    CK(cudaMalloc(&ctx->ay, sizeof(double) * hy.size()));
    CK(cudaMalloc(&ctx->az, sizeof(double) * hz.size()));
    CK(cudaMemcpyAsync(ctx->ax, hx.data(), sizeof(double) * hx.size(), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(ctx->ay, hy.data(), sizeof(double) * hy.size(), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(ctx->az, hz.data(), sizeof(double) * hz.size(), cudaMemcpyHostToDevice, st));
    ctx->gridRows = (size_t)(ctx->nx + 1) * (ctx->ny + 1) * ((D == 3) ? (ctx->nz + 1) : 1);
    double *gridTmp = nullptr, *monV = nullptr;
    constexpr int RS = GridRow<D>::stride;
    CK(cudaMalloc(&ctx->grid, sizeof(double) * ctx->gridRows * RS));
    CK(cudaMalloc(&gridTmp, sizeof(double) * ctx->gridRows * RS));
    CK(cudaMemsetAsync(ctx->grid, 0, sizeof(double) * ctx->gridRows * RS, st));
    CK(cudaMalloc(&monV, sizeof(double) * (size_t)P * DD));
    CK(cudaMemcpyAsync(monV, monVertexHost, sizeof(double) * (size_t)P * DD, cudaMemcpyHostToDevice, st));

    // uniform bins over the vertex bounding box (about 2 vertices per bin)
    BinGeom bg;
    int per = std::max(1, (int)std::floor(std::pow((double)P / 2.0, 1.0 / D)));
    double minbw = INFINITY;
    for (int l = 0; l < 3; l++) { bg.nb[l] = 1; bg.lo[l] = 0; bg.inv[l] = 0; }
    for (int l = 0; l < D; l++) {
        bg.nb[l] = per; bg.lo[l] = mn[l];
        double ext = mx[l] - mn[l];
        bg.inv[l] = ext > 0 ? per / ext : 0.0;
        if (ext > 0) minbw = std::min(minbw, ext / per);
    }
    // the stop test uses a slightly shrunken bin width so that rounding in the bin assignment can never
    // let a closer vertex hide in an unvisited ring
    bg.minbw = std::isfinite(minbw) ? minbw * (1.0 - 1e-9) : 0.0;
    bg.maxr = per;
    size_t nbins = (size_t)bg.nb[0] * bg.nb[1] * bg.nb[2];
    int *bin = nullptr, *ids = nullptr, *binStart = nullptr, *items = nullptr;
    CK(cudaMalloc(&bin, sizeof(int) * (size_t)P));
    CK(cudaMalloc(&ids, sizeof(int) * (size_t)P));
    k_bin_ids<D><<<cdiv(P, 256), 256, 0, st>>>(P, xg, bg, bin, ids);
    int rc = build_csr(ctx, (size_t)P, (int)nbins, bin, ids, &binStart, &items);
    if (rc) return rc;
    ctx->gridO[0] = hx[0]; ctx->gridO[1] = hy[0]; ctx->gridO[2] = hz[0];
    ctx->gridRh[0] = 1.0 / (hx[1] - hx[0]); ctx->gridRh[1] = 1.0 / (hy[1] - hy[0]);
    ctx->gridRh[2] = (D == 3) ? 1.0 / (hz[1] - hz[0]) : 0.0;
    MonGrid mg = ctx->monGrid();
    k_nn_grid<D><<<cdiv(ctx->gridRows, 128), 128, 0, st>>>(mg, ctx->grid, xg, monV, bg, binStart, items);
    int nSmooth = (D == 2) ? 5 : 2;  // src/MeshInterpolator.cpp:246-252
    for (int it = 0; it < nSmooth; it++) {
        CK(cudaMemcpyAsync(gridTmp, ctx->grid, sizeof(double) * ctx->gridRows * RS, cudaMemcpyDeviceToDevice, st));
        k_smooth<D><<<cdiv(ctx->gridRows * DD, 256), 256, 0, st>>>(ctx->nx, ctx->ny, ctx->nz, gridTmp, ctx->grid);
    }
    {   // isotropic fast path of the lookups
        int *flag = nullptr, one = 1;
        CK(cudaMalloc(&flag, sizeof(int)));
        CK(cudaMalloc(&ctx->gridIso, sizeof(double) * ctx->gridRows));
        CK(cudaMemcpyAsync(flag, &one, sizeof(int), cudaMemcpyHostToDevice, st));
        k_iso_extract<D><<<cdiv(ctx->gridRows, 256), 256, 0, st>>>(ctx->gridRows, ctx->grid, ctx->gridIso, flag);
        CK(cudaMemcpyAsync(&one, flag, sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        cudaFree(flag);
        ctx->gridIsIso = one != 0;
        if (!ctx->gridIsIso) { cudaFree(ctx->gridIso); ctx->gridIso = nullptr; }
        // Keep the scalar grid resident in L2 (persisting access-policy window on the library's stream): the prox kernels
        // stream gigabytes of inverse Hessians through L2 per launch, which would otherwise keep evicting the 25-34 MB the
        // monitor lookups hit.  Best effort: ignored where the device grants no persisting L2.
        // Measured: 3D (8 values in 4 lines per lookup, no prefetch one tile ahead) 4.44 -> 4.12 ms on C4; 2D loses (0.584 -> 0.647 ms
        // on C2: the carve-out shrinks the L2 left for the vertex vectors and the 2D kernel already prefetches its lookups), so
        // the window is set for D = 3 only.
        if (D == 3 && ctx->gridIsIso && ctx->l2Persist && !getenv("MMADMM_NO_L2PERSIST")) {
            int maxPersist = 0, maxWin = 0;
            cudaDeviceGetAttribute(&maxPersist, cudaDevAttrMaxPersistingL2CacheSize, ctx->device);
            cudaDeviceGetAttribute(&maxWin, cudaDevAttrMaxAccessPolicyWindowSize, ctx->device);
            // On a distributed context every rank holds the whole grid (built from the global vertex set) but looks up only the
            // slab its own vertices live in: element blocks of the box mesh cut along z, the slowest grid index, so that slab is
            // one contiguous range of planes.  The window covers those planes (+2 on either side: nodes move) instead of a
            // fraction of everything -- at N = 8 on the C4 ladder the grid is 199 MB, the slab 25 MB.
            const size_t sl = (size_t)(ctx->nx + 1) * (ctx->ny + 1);
            long z0 = 0, z1 = ctx->nz;
            if (ctx->dist && ctx->locMax[2] >= ctx->locMin[2]) {
                z0 = std::max(0l, (long)std::floor((ctx->locMin[2] - ctx->gridO[2]) * ctx->gridRh[2]) - 2);
                z1 = std::min((long)ctx->nz, (long)std::floor((ctx->locMax[2] - ctx->gridO[2]) * ctx->gridRh[2]) + 3);
            }
            const size_t bytes = sizeof(double) * sl * (size_t)(z1 - z0 + 1);
            if (maxPersist > 0 && maxWin > 0) {
                const size_t lim = std::min<size_t>((size_t)maxPersist, bytes);
                cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, lim);
                cudaStreamAttrValue av;
                memset(&av, 0, sizeof(av));
                av.accessPolicyWindow.base_ptr = ctx->gridIso + sl * (size_t)z0;
                av.accessPolicyWindow.num_bytes = std::min<size_t>(bytes, (size_t)maxWin);
                av.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)lim / (double)av.accessPolicyWindow.num_bytes);
                av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                if (cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &av) != cudaSuccess) cudaGetLastError();
            }
        }
    }
    ctx->launches += 3 + nSmooth;
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    cudaFree(gridTmp); cudaFree(monV); cudaFree(bin); cudaFree(ids); cudaFree(binStart); cudaFree(items); cudaFree(xg);
    return MMADMM_OK;
}

// Quantities that differ from the local mesh on a distributed context (null / equal to the local ones otherwise)
struct InitGlobals {
    int Eglobal;                 // element count in the reference-simplex scaling (AdaptationFunctional.cpp:190-200)
    const int *valence;          // global valence of the local vertices (diagonal of t), or null
    int Pgrid;                   // vertices the monitor grid is built from
    const double *VpGrid, *monGrid;
};

template <int D>
int create_impl(mmadmm_ctx *ctx, const double *Vp, const double *Vc, const int *F, const int *mask, const InitGlobals &G) {
    constexpr int n = D * (D + 1);
    const int P = ctx->P, E = ctx->E;
    cudaStream_t st = ctx->stream;
    ctx->Es = ((size_t)E + ES_ALIGN - 1) / ES_ALIGN * ES_ALIGN;
    const size_t NV = (size_t)P * D, NZ = ctx->Es * n;
    CK(cudaMalloc(&ctx->x, sizeof(double) * NV));
    CK(cudaMalloc(&ctx->xPrev, sizeof(double) * NV));
    CK(cudaMalloc(&ctx->xBar, sizeof(double) * NV));
    CK(cudaMalloc(&ctx->xSpec, sizeof(double) * NV));
    CK(cudaMalloc(&ctx->xBarSpec, sizeof(double) * NV));
    CK(cudaMalloc(&ctx->finTicket, sizeof(unsigned)));
    CK(cudaMemsetAsync(ctx->finTicket, 0, sizeof(unsigned), st));
    CK(cudaMalloc(&ctx->tdiag, sizeof(double) * (size_t)P));
    CK(cudaMemcpyAsync(ctx->x, Vp, sizeof(double) * NV, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(ctx->xPrev, ctx->x, sizeof(double) * NV, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(ctx->xBar, ctx->x, sizeof(double) * NV, cudaMemcpyDeviceToDevice, st));
    if (Vc) {
        CK(cudaMalloc(&ctx->Vc, sizeof(double) * NV));
        CK(cudaMemcpyAsync(ctx->Vc, Vc, sizeof(double) * NV, cudaMemcpyHostToDevice, st));
    }
    std::vector<uint8_t> m8(P);
    for (int p = 0; p < P; p++) m8[p] = (uint8_t)mask[p];
    CK(cudaMalloc(&ctx->mask, (size_t)P));
    CK(cudaMemcpyAsync(ctx->mask, m8.data(), (size_t)P, cudaMemcpyHostToDevice, st));
    int *Frow = nullptr;
    CK(cudaMalloc(&Frow, sizeof(int) * (size_t)E * (D + 1)));
    CK(cudaMemcpyAsync(Frow, F, sizeof(int) * (size_t)E * (D + 1), cudaMemcpyHostToDevice, st));
    for (int v = 0; v < D + 1; v++) {   // padded to whole tiles (asynchronous tile loads read the padding)
        CK(cudaMalloc(&ctx->F[v], sizeof(int) * ctx->Es));
        CK(cudaMemsetAsync(ctx->F[v], 0, sizeof(int) * ctx->Es, st));
    }
    CK(cudaMalloc(&ctx->fixedBits, ctx->Es));   // whole tiles: k_prox_vw fetches the aligned word around an element's byte
    CK(cudaMemsetAsync(ctx->fixedBits, 0, ctx->Es, st));
    CK(cudaMalloc(&ctx->nonIntBits, (size_t)E));
    k_setup_F<D><<<cdiv(E, 256), 256, 0, st>>>(E, Frow, ctx->F[0], ctx->F[1], ctx->F[2], ctx->F[3], ctx->x, ctx->mask,
                                               ctx->fixedBits, ctx->nonIntBits);
    // incidence
    int *keys = nullptr, *vals = nullptr;
    CK(cudaMalloc(&keys, sizeof(int) * (size_t)E * (D + 1)));
    CK(cudaMalloc(&vals, sizeof(int) * (size_t)E * (D + 1)));
    k_inc_pairs<D><<<cdiv(E, 256), 256, 0, st>>>(E, ctx->F[0], ctx->F[1], ctx->F[2], ctx->F[3], keys, vals);
    int rc = build_csr(ctx, (size_t)E * (D + 1), P, keys, vals, &ctx->incPtr, &ctx->incIdx);
    if (rc) return rc;
    cudaFree(keys); cudaFree(vals); cudaFree(Frow);
    ctx->launches += 2;
    {   // vertex processing order of the vertex kernels
        int *vk = nullptr, *vv = nullptr, *vptr = nullptr;
        CK(cudaMalloc(&vk, sizeof(int) * (size_t)P));
        CK(cudaMalloc(&vv, sizeof(int) * (size_t)P));
        k_vorder_keys<<<cdiv(P, 256), 256, 0, st>>>(P, E, ctx->incPtr, ctx->incIdx, vk, vv);
        ctx->launches++;
        rc = build_csr(ctx, (size_t)P, E + 1, vk, vv, &vptr, &ctx->vorder);
        if (rc) return rc;
        cudaFree(vk); cudaFree(vv); cudaFree(vptr);
    }

    // functional constants, with the reference's own expressions (AdaptationFunctional.cpp:176-201, 219-220)
    {
        double Eh[3][3] = {{0}};
        double dFact = (D == 2) ? 2.0 : 6.0;
        if (D == 2) {
            Eh[0][0] = 1.0; Eh[1][0] = 0.0; Eh[0][1] = 1.0 / 2.0; Eh[1][1] = std::sqrt(3) / 2.0;
        } else {
            Eh[0][0] = -2.0; Eh[1][0] = 0.0;  Eh[2][0] = -2.0;
            Eh[0][1] = 0.0;  Eh[1][1] = -2.0; Eh[2][1] = -2.0;
            Eh[0][2] = -2.0; Eh[1][2] = -2.0; Eh[2][2] = 0.0;
        }
        double det;
        if (D == 2) det = Eh[0][0] * Eh[1][1] - Eh[1][0] * Eh[0][1];
        else
            det = Eh[0][0] * (Eh[1][1] * Eh[2][2] - Eh[1][2] * Eh[2][1]) - Eh[0][1] * (Eh[1][0] * Eh[2][2] - Eh[1][2] * Eh[2][0]) +
                  Eh[0][2] * (Eh[1][0] * Eh[2][1] - Eh[1][1] * Eh[2][0]);
        double s1 = std::pow((dFact / std::fabs(det)), 1.0 / ((double)D));
        double s2 = (double)std::pow(G.Eglobal, 1.0 / D);
        for (int a = 0; a < D; a++)
            for (int b = 0; b < D; b++) { double v = Eh[a][b]; v *= s1; v /= s2; ctx->Ehat[a * D + b] = v; }
        ctx->dpow = std::pow((double)D, (double)D * 1.5 / 2.0);
    }
    ctx->w = 0.5 * std::sqrt(ctx->rho);  // src/Mesh.cpp:451 (the JSON "w" is ignored by the reference)
    ctx->m = ctx->tau;
    if (G.valence) {   // kept: mmadmm_set_dt rebuilds the diagonal
        CK(cudaMalloc(&ctx->valence_d, sizeof(int) * (size_t)P));
        CK(cudaMemcpyAsync(ctx->valence_d, G.valence, sizeof(int) * (size_t)P, cudaMemcpyHostToDevice, st));
    }
    k_tdiag<<<cdiv(P, 256), 256, 0, st>>>(P, ctx->incPtr, ctx->valence_d, ctx->m, ctx->dt * ctx->dt, ctx->w, ctx->tdiag);
    CK(cudaStreamSynchronize(st));
    ctx->launches++;

    for (int l = 0; l < D; l++) {
        double lo = INFINITY, hi = -INFINITY;
        for (int pp = 0; pp < P; pp++) { const double v = Vp[(size_t)pp * D + l]; lo = v < lo ? v : lo; hi = v > hi ? v : hi; }
        ctx->locMin[l] = lo; ctx->locMax[l] = hi;
    }
    rc = build_monitor_grid<D>(ctx, G.Pgrid, G.VpGrid, G.monGrid);
    if (rc) return rc;

    // element state: z = D x, u = 0, hessInvs = I
    CK(cudaMalloc(&ctx->z, sizeof(double) * NZ));
    CK(cudaMalloc(&ctx->u, sizeof(double) * NZ));
    CK(cudaMalloc(&ctx->B, sizeof(double) * NZ * n));
    CK(cudaMalloc(&ctx->gs, sizeof(double) * NZ));
    CK(cudaMemsetAsync(ctx->z, 0, sizeof(double) * NZ, st));
    CK(cudaMemsetAsync(ctx->u, 0, sizeof(double) * NZ, st));
    CK(cudaMalloc(&ctx->usum, sizeof(double) * (size_t)P * D));
    CK(cudaMemsetAsync(ctx->usum, 0, sizeof(double) * (size_t)P * D, st));
    CK(cudaMemsetAsync(ctx->B, 0, sizeof(double) * NZ * n, st));
    k_fill_identity<D><<<cdiv(ctx->Es, 256), 256, 0, st>>>(ctx->Es, ctx->B);   // hessInvs = I (src/Mesh.cpp:456-464)
    ctx->launches++;
    CK(cudaDeviceGetAttribute(&ctx->numSMs, cudaDevAttrMultiProcessorCount, ctx->device));
    ctx->nblkE = cdiv(E, TPB);
    ctx->nblkV = cdiv(P, TPB_V);
    ctx->nblkProx = std::max(cdiv(E, 32), 148 * 64);   // upper bound over the prox tile sizes / persistent warps
    CK(cudaMalloc(&ctx->partIh, sizeof(double) * ctx->nblkProx));
    CK(cudaMalloc(&ctx->partDual, sizeof(double) * ctx->nblkProx));
    CK(cudaMalloc(&ctx->partPrimal, sizeof(double) * (size_t)(std::max(ctx->nblkE, ctx->nblkV) + 3 * ctx->nblkV + 1)));
    CK(cudaMalloc(&ctx->scal_d, sizeof(double) * 8));
    CK(cudaMemsetAsync(ctx->scal_d, 0, sizeof(double) * 8, st));
    CK(cudaHostAlloc(&ctx->scal_h, sizeof(double) * 8, cudaHostAllocMapped));
    CK(cudaHostGetDevicePointer(&ctx->scal_hd, ctx->scal_h, 0));
    std::memset(ctx->scal_h, 0, sizeof(double) * 8);
    CK(cudaMalloc(&ctx->status_d, sizeof(unsigned)));
    CK(cudaHostAlloc(&ctx->status_h, sizeof(unsigned), cudaHostAllocMapped));
    CK(cudaHostGetDevicePointer(&ctx->status_hd, ctx->status_h, 0));
    *ctx->status_h = 0;
    CK(cudaMemsetAsync(ctx->status_d, 0, sizeof(unsigned), st));
    CK(cudaMalloc(&ctx->hist_d, sizeof(unsigned long long) * 52));
    CK(cudaMemsetAsync(ctx->hist_d, 0, sizeof(unsigned long long) * 52, st));
    ElemArgs<D> a = ctx->elemArgs<D>(0.0);
    k_applyD<D><<<ctx->nblkE, TPB, 0, st>>>(a, ctx->x);
    ctx->launches++;
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return MMADMM_OK;
}

// Mesh::buildFaceList (src/Mesh.cpp:61-112) for the boundary-sliding option: a simplex with exactly D non-INTERIOR vertices
// contributes one boundary face (those vertices in local order); per vertex the faces it belongs to in ascending face id.
int build_face_lists(mmadmm_ctx *ctx) {
    if (ctx->facePtr_d) return MMADMM_OK;
    const int D = ctx->D, V = D + 1, P = ctx->P, E = ctx->E;
    std::vector<int> Fh((size_t)E * V), plane(E);
    std::vector<uint8_t> mask(P);
    CK(cudaStreamSynchronize(ctx->stream));
    for (int v = 0; v < V; v++) {
        CK(cudaMemcpy(plane.data(), ctx->F[v], sizeof(int) * (size_t)E, cudaMemcpyDeviceToHost));
        for (int e = 0; e < E; e++) Fh[(size_t)e * V + v] = plane[e];
    }
    CK(cudaMemcpy(mask.data(), ctx->mask, (size_t)P, cudaMemcpyDeviceToHost));
    std::vector<int> faces;
    for (int e = 0; e < E; e++) {
        int sum = 0;
        for (int j = 0; j < V; j++) sum += mask[Fh[(size_t)e * V + j]] != NT_INTERIOR;
        if (sum == D)
            for (int j = 0; j < V; j++)
                if (mask[Fh[(size_t)e * V + j]] != NT_INTERIOR) faces.push_back(Fh[(size_t)e * V + j]);
    }
    const int nF = (int)(faces.size() / D);
    std::vector<int> ptr((size_t)P + 1, 0), idx;
    for (int pass = 0; pass < 2; pass++) {
        std::vector<int> fill;
        if (pass == 1) {
            for (int p = 0; p < P; p++) ptr[p + 1] += ptr[p];
            idx.resize(ptr[P]);
            fill.assign(ptr.begin(), ptr.end() - 1);
        }
        for (int f = 0; f < nF; f++)
            for (int j = 0; j < D; j++) {
                bool dup = false;
                for (int jj = 0; jj < j; jj++) dup = dup || faces[(size_t)f * D + jj] == faces[(size_t)f * D + j];
                if (dup) continue;
                if (pass == 0) ptr[faces[(size_t)f * D + j] + 1]++;
                else idx[fill[faces[(size_t)f * D + j]]++] = f;
            }
    }
    CK(cudaMalloc(&ctx->facePtr_d, sizeof(int) * ptr.size()));
    CK(cudaMalloc(&ctx->faceIdx_d, sizeof(int) * std::max<size_t>(idx.size(), 1)));
    CK(cudaMalloc(&ctx->faceList_d, sizeof(int) * std::max<size_t>(faces.size(), 1)));
    CK(cudaMemcpy(ctx->facePtr_d, ptr.data(), sizeof(int) * ptr.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->faceIdx_d, idx.data(), sizeof(int) * idx.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->faceList_d, faces.data(), sizeof(int) * faces.size(), cudaMemcpyHostToDevice));
    ctx->nFaces = nF;
    return MMADMM_OK;
}

int check_status(mmadmm_ctx *ctx) {
    unsigned s = *ctx->status_h;
    if (s & ST_PEER_TIMEOUT) {
        ctx->err = "a peer rank did not reach the exchange within peer_timeout_s: it died, or its host stalled for longer than that";
        return MMADMM_EPEER;
    }
    if (s & ST_INVERTED) { ctx->err = "inverted element (det E <= 0): the reference asserts here (AdaptationFunctional.cpp:174)"; return MMADMM_EINVERTED; }
    if (s & ST_SINGULAR_HESS) { ctx->err = "singular finite-difference Hessian"; return MMADMM_ESINGULAR; }
    return MMADMM_OK;
}


// ---- write-back of the node positions (Mesh::updateAfterStep, src/Mesh.cpp:1030-1035) -------------------------------------
// After every mmadmm_step the final x is snapshotted device-to-device into one of two staging buffers (transposed there when
// the caller's array is column-major) and copied to the caller's page-locked array on a separate copy stream, so the PCIe
// transfer of step s overlaps the kernels of step s+1.  mode 1 (strict) fences at once: the reference's semantics.
int ensure_local_runs(mmadmm_ctx *ctx) {
    if (!ctx->dist || !ctx->verts_h.empty()) return MMADMM_OK;
    ctx->verts_h.resize(ctx->P);   // runs of consecutive global ids (a strip / slab partition has two or three)
    CK(cudaMemcpy(ctx->verts_h.data(), ctx->verts, sizeof(int) * (size_t)ctx->P, cudaMemcpyDeviceToHost));
    for (int l = 0; l < ctx->P;) {
        int m = l + 1;
        while (m < ctx->P && ctx->verts_h[m] == ctx->verts_h[m - 1] + 1) m++;
        ctx->runs.push_back({l, m - l});
        l = m;
    }
    return MMADMM_OK;
}
int wb_fence(mmadmm_ctx *ctx) {
    mmadmm_ctx::Writeback &w = ctx->wb;
    for (int b = 0; b < 2; b++)
        if (w.inFlight[b]) {
            CK(cudaEventSynchronize(w.evCopied[b]));
            w.inFlight[b] = false;
        }
    return MMADMM_OK;
}
// staging buffers, copy stream and events: allocated when a write-back target is set (mmadmm_set_writeback), never inside a step
int wb_setup(mmadmm_ctx *ctx) {
    mmadmm_ctx::Writeback &w = ctx->wb;
    if (w.copyStream) return MMADMM_OK;
    const size_t bytes = sizeof(double) * (size_t)ctx->P * ctx->D;
    CK(cudaStreamCreateWithFlags(&w.copyStream, cudaStreamNonBlocking));
    for (int b = 0; b < 2; b++) {
        CK(cudaMalloc(&w.stage[b], bytes));
        CK(cudaEventCreateWithFlags(&w.evSnap[b], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&w.evCopied[b], w.debug ? cudaEventDefault : cudaEventDisableTiming));
        if (w.debug) CK(cudaEventCreate(&w.evCopyStart[b]));
    }
    return MMADMM_OK;
}
int wb_issue(mmadmm_ctx *ctx) {
    mmadmm_ctx::Writeback &w = ctx->wb;
    const int D = ctx->D, P = ctx->P;
    // MMADMM_DEBUG_WB=1: host time of every queueing call and device time of every copy, reported at destroy
    auto now = []() { timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec * 1000000000ll + t.tv_nsec; };
    const bool dbg = w.debug;
    long long tPrev = dbg ? now() : 0;
    auto lap = [&](int k) {
        if (!dbg) return;
        const long long t = now();
        w.dbgNs[k] += t - tPrev;
        if (t - tPrev > w.dbgMax[k]) w.dbgMax[k] = t - tPrev;
        tPrev = t;
    };
    const size_t bytes = sizeof(double) * (size_t)P * D;
    {
        int rcs = wb_setup(ctx);
        if (rcs) return rcs;
        lap(0);
    }
    const int b = w.next;
    w.next ^= 1;
    if (w.inFlight[b]) {
        if (dbg && w.copies >= 2) {   // duration of the copy that used this stage two steps ago
            float ms = 0;
            if (cudaEventSynchronize(w.evCopied[b]) == cudaSuccess && cudaEventElapsedTime(&ms, w.evCopyStart[b], w.evCopied[b]) == cudaSuccess) {
                w.dbgCopyMs += ms; w.dbgCopyN++;
                if (ms > w.dbgCopyMax) w.dbgCopyMax = ms;
            }
        }
        CK(cudaStreamWaitEvent(ctx->stream, w.evCopied[b], 0));   // two steps old: the stage is free again
    }
    lap(1);
    // snapshot by a kernel, not a copy-engine memcpy: the stream stays on the compute engine between two steps
    if (w.snapByCE && !w.colMajor) CK(cudaMemcpyAsync(w.stage[b], ctx->x, bytes, cudaMemcpyDeviceToDevice, ctx->stream));
    else if (w.colMajor) k_x_to_colmajor<<<cdiv(P, 256), 256, 0, ctx->stream>>>(P, D, ctx->x, w.stage[b]);
    else k_copy<<<cdiv((size_t)P * D, 256), 256, 0, ctx->stream>>>((size_t)P * D, ctx->x, w.stage[b]);
    ctx->launches++;
    lap(2);
    CK(cudaEventRecord(w.evSnap[b], ctx->stream));
    CK(cudaStreamWaitEvent(w.copyStream, w.evSnap[b], 0));
    if (dbg) CK(cudaEventRecord(w.evCopyStart[b], w.copyStream));
    lap(3);
    if (!ctx->dist) {
        CK(cudaMemcpyAsync(w.host, w.stage[b], bytes, cudaMemcpyDeviceToHost, w.copyStream));
    } else {   // this rank's vertices into their rows of the caller's GLOBAL array, one copy per run of consecutive ids
        int rc = ensure_local_runs(ctx);
        if (rc) return rc;
        for (auto &r : ctx->runs) {
            const size_t g0 = (size_t)ctx->verts_h[r.first];
            if (w.colMajor) {
                for (int l = 0; l < D; l++)
                    CK(cudaMemcpyAsync(w.host + (size_t)l * ctx->Pglobal + g0, w.stage[b] + (size_t)l * P + r.first,
                                       sizeof(double) * (size_t)r.second, cudaMemcpyDeviceToHost, w.copyStream));
            } else {
                CK(cudaMemcpyAsync(w.host + g0 * D, w.stage[b] + (size_t)r.first * D, sizeof(double) * (size_t)r.second * D,
                                   cudaMemcpyDeviceToHost, w.copyStream));
            }
        }
    }
    lap(4);
    CK(cudaEventRecord(w.evCopied[b], w.copyStream));
    w.inFlight[b] = true;
    w.copies++;
    lap(5);
    if (w.mode == 1) { int rc = wb_fence(ctx); lap(6); return rc; }
    return MMADMM_OK;
}

template <int D, int TPB_> constexpr size_t prox_smem_bytes() {
    return sizeof(double) * (size_t)(D * (D + 1)) * (D * (D + 1)) * TPB_ + 16 + sizeof(double) * 32 + sizeof(int) * 52;
}

template <int D, int TPB_, int MINB> int launch_prox_t(mmadmm_ctx *ctx, const ElemArgs<D> &a, int slot) {
    constexpr size_t smem = prox_smem_bytes<D, TPB_>();
    const bool iso = a.mg.iso != nullptr;
    auto kern = ctx->compMesh ? (iso ? k_prox<D, TPB_, MINB, true, true> : k_prox<D, TPB_, MINB, true, false>)
                              : (iso ? k_prox<D, TPB_, MINB, false, true> : k_prox<D, TPB_, MINB, false, false>);
    if (!ctx->smemAttrSet[slot + (iso ? 8 : 0)]) {
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ctx->smemAttrSet[slot + (iso ? 8 : 0)] = true;
    }
    kern<<<cdiv(ctx->E, TPB_), TPB_, smem, ctx->stream>>>(a);
    ctx->nProxPart = cdiv(ctx->E, TPB_);
    ctx->launches++;
    return MMADMM_OK;
}

// Tile shape / occupancy variants of the prox kernel ("prox_variant" option; 0 is the tuned default)
template <int D> int launch_prox(mmadmm_ctx *ctx, const ElemArgs<D> &a);
template <int D, int WPB, int MINB, int FLAGS> int launch_prox_p(mmadmm_ctx *ctx, const ElemArgs<D> &a, int slot) {
    constexpr int n = D * (D + 1);
    const size_t slots = ctx->compMesh ? 4 : 3;
    const size_t smem = sizeof(double) * (size_t)WPB * (n * n * 32 + 2 * slots * n * 32) + 32 * WPB;
    const bool iso = a.mg.iso != nullptr;
    auto kern = ctx->compMesh ? (iso ? k_prox_p<D, WPB, MINB, true, FLAGS | PF_ISO> : k_prox_p<D, WPB, MINB, true, FLAGS>)
                              : (iso ? k_prox_p<D, WPB, MINB, false, FLAGS | PF_ISO> : k_prox_p<D, WPB, MINB, false, FLAGS>);
    if (!ctx->smemAttrSet[slot + (iso ? 8 : 0)]) {
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ctx->smemAttrSet[slot + (iso ? 8 : 0)] = true;
    }
    const int grid = std::min(ctx->numSMs * MINB, cdiv(ctx->E, WPB * 32));
    kern<<<grid, WPB * 32, smem, ctx->stream>>>(a);
    ctx->nProxPart = grid * WPB;
    ctx->launches++;
    return MMADMM_OK;
}
template <> int launch_prox<2>(mmadmm_ctx *ctx, const ElemArgs<2> &a) {
    switch (ctx->proxVariant) {
        case 1: return launch_prox_p<2, 2, 6, 0>(ctx, a, 1);                  // no monitor-grid prefetch
        case 2: return launch_prox_p<2, 4, 3, PF_PREFETCH_GRID>(ctx, a, 2);   // 4-warp CTAs
        case 3: return launch_prox_p<2, 6, 2, PF_PREFETCH_GRID>(ctx, a, 3);   // 6-warp CTAs
        case 4: return launch_prox_t<2, 128, 4>(ctx, a, 4);                   // CTA-tile form (non-persistent)
        default: return launch_prox_p<2, 2, 6, PF_PREFETCH_GRID>(ctx, a, 0);
    }
}
template <int WPB, int MINB, bool KEEP = false> int launch_prox_q(mmadmm_ctx *ctx, const ElemArgs<3> &a, int slot) {
    const bool iso = a.mg.iso != nullptr;
    const size_t warpBytes = ((size_t)(144 * SUB3 + (ctx->compMesh ? 4 : 3) * 3 * 32) * 8 + 2 * 32 * 4 + 16 + 127) / 128 * 128;
    const size_t smem = warpBytes * WPB;
    auto kern = ctx->compMesh ? (iso ? k_prox_q<true, true, WPB, MINB, KEEP> : k_prox_q<true, false, WPB, MINB, KEEP>)
                              : (iso ? k_prox_q<false, true, WPB, MINB, KEEP> : k_prox_q<false, false, WPB, MINB, KEEP>);
    if (!ctx->smemAttrSet[slot + (iso ? 8 : 0)]) {
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ctx->smemAttrSet[slot + (iso ? 8 : 0)] = true;
    }
    const int grid = std::min(ctx->numSMs * MINB, cdiv(ctx->E, WPB * SUB3));    // persistent: each warp walks sub-tiles with this stride
    kern<<<grid, WPB * 32, smem, ctx->stream>>>(a);
    ctx->nProxPart = grid * WPB;
    ctx->launches++;
    return MMADMM_OK;
}
template <> int launch_prox<3>(mmadmm_ctx *ctx, const ElemArgs<3> &a) {
    switch (ctx->proxVariant) {
        case 1: return launch_prox_q<2, 6>(ctx, a, 1);         // 2-warp CTAs
        case 2: return launch_prox_q<4, 4>(ctx, a, 2);         // 16 warps per SM at 128 registers
        case 3: return launch_prox_q<4, 2, true>(ctx, a, 3);   // own B rows in registers, 8 warps per SM (255 registers)
        case 4: return launch_prox_q<2, 5, true>(ctx, a, 4);   // own B rows in registers, 10 warps per SM (204 registers)
        case 5: return launch_prox_q<4, 3, true>(ctx, a, 5);   // own B rows in registers, 12 warps per SM (168 registers)
        default: return launch_prox_q<4, 3>(ctx, a, 0);        // 12 warps per SM
    }
}

// x-update (+ primal residual partials).  Distributed: local kernel -> shared-vertex partials -> exchange -> k_xfinish.
//   sharedOnly: the vertices that are not shared were already solved by k_xfirst
// The x-update in three pieces.  On a distributed context with the NVLink mailbox exchange they run in the order
//   xupdate_shared (partial sums of the shared vertices -> sPart AND straight into the neighbours' mailboxes)
//   bulk pass over all other vertices            <- the halo is in flight meanwhile
//   xupdate_finish (wait for the neighbours' epoch, rank-ordered sum, x and primal share of the shared vertices)
// so the latency of the exchange and the skew between the ranks hide behind the bulk pass.
template <int D, bool ZSRC, bool UZERO, bool PRIMAL, bool UREC>
int xupdate_shared(mmadmm_ctx *ctx, const double *zsrc, bool push, unsigned long long &ep) {
    const double dtsq = ctx->dt * ctx->dt;
    HaloPush hp;
    if (push) {
        ep = ++ctx->haloEpoch;
        hp.ptr = ctx->pushPtr_d; hp.nbr = ctx->pushNbr_d; hp.pos = ctx->pushPos_d;
        hp.peerRecv = ctx->peerRecv_d; hp.peerStride = ctx->peerRecvStride_d; hp.peerFlag = ctx->peerHaloFlag_d;
        hp.ticket = ctx->pushTicket_d; hp.nNbr = (int)ctx->nbrRank.size(); hp.epoch = ep;
    }
    // primal partials of this launch are all zero (shared vertices leave theirs to xupdate_finish): scratch region
    k_xupdate<D, ZSRC, UZERO, PRIMAL, UREC><<<ctx->nblkSh, TPB_V, 0, ctx->stream>>>(
        ctx->nShared, ctx->Es, ctx->incPtr, ctx->incIdx, ctx->z, ctx->u, zsrc, ctx->xBar, ctx->tdiag, ctx->m, dtsq, ctx->w, ctx->x,
        ctx->partPrimal + ctx->nblkV + ctx->nblkSh, ctx->sharedSlot, ctx->sPart, ctx->sharedIdx, ctx->usum, hp, 0, StepTail());
    ctx->launches++;
    return MMADMM_OK;
}
template <int D, bool ZSRC, bool UZERO, bool PRIMAL, bool UREC>
int xupdate_bulk(mmadmm_ctx *ctx, const double *zsrc, bool skipShared, const StepTail &tail = StepTail()) {
    const double dtsq = ctx->dt * ctx->dt;
    auto kern = k_xupdate<D, ZSRC, UZERO, PRIMAL, UREC>;
    if constexpr (UREC) {   // slots per chunk / resident CTAs of the recurrence form ("xup_variant", A/B runs; 0 = tuned default)
        switch (ctx->xupVariant) {
            case 1: kern = k_xupdate<D, ZSRC, UZERO, PRIMAL, UREC, 4, 4>; break;
            case 2: kern = k_xupdate<D, ZSRC, UZERO, PRIMAL, UREC, 2, 6>; break;
            case 3: kern = k_xupdate<D, ZSRC, UZERO, PRIMAL, UREC, 8, 2>; break;
            case 4: kern = k_xupdate<D, ZSRC, UZERO, PRIMAL, UREC, 4, 2>; break;
            case 5: kern = k_xupdate<D, ZSRC, UZERO, PRIMAL, UREC, 2, 8>; break;
            case 6: kern = k_xupdate<D, ZSRC, UZERO, PRIMAL, UREC, 8, 1>; break;
            case 7: kern = k_xupdate<D, ZSRC, UZERO, PRIMAL, UREC, 2, 5>; break;
            case 8: kern = k_xupdate<D, ZSRC, UZERO, PRIMAL, UREC, 4, 5>; break;
            // measured with the single-pass form (scripts/gpu_r2_q.sh): 2D 2 slots per chunk at 6 CTAs per SM (40 registers, no
            // spills) 0.143 ms against 0.147 (2,8) / 0.145 (4,4) / 0.166 (8,2); 3D 4 slots at 4 CTAs 0.256 against 0.280 (8,1) / 0.295 (2,6)
            default: kern = (D == 2) ? k_xupdate<D, ZSRC, UZERO, PRIMAL, UREC, 2, 6> : k_xupdate<D, ZSRC, UZERO, PRIMAL, UREC, 4, 4>; break;
        }
    }
    kern<<<ctx->nblkV, TPB_V, 0, ctx->stream>>>(
        ctx->P, ctx->Es, ctx->incPtr, ctx->incIdx, ctx->z, ctx->u, zsrc, ctx->xBar, ctx->tdiag, ctx->m, dtsq, ctx->w, ctx->x,
        ctx->partPrimal, ctx->dist ? ctx->sharedSlot : nullptr, ctx->sPart, ctx->vorder, ctx->usum, HaloPush(), skipShared ? 1 : 0, tail);
    ctx->launches++;
    return MMADMM_OK;
}
template <int D, bool ZSRC, bool PRIMAL>
int xupdate_finish(mmadmm_ctx *ctx, const double *zsrc, unsigned long long ep) {
    const double dtsq = ctx->dt * ctx->dt;
    if (ctx->p2p) {
        k_xfinish_p2p<D, ZSRC, PRIMAL><<<ctx->nblkSh, TPB_V, 0, ctx->stream>>>(
            ctx->nShared, (int)ctx->nbrRank.size(), ctx->sharedIdx, ctx->contribPtr, ctx->contrib, ctx->sPart, ctx->mbRecv_d,
            (size_t)std::max(ctx->nNbrItems, 1) * D, ctx->haloFlags_d, ep, ctx->peerTimeoutNs, ctx->status_d, ctx->incPtr, ctx->incIdx,
            ctx->z, zsrc, ctx->xBar, ctx->tdiag, ctx->m, dtsq, ctx->x, ctx->partPrimal + ctx->nblkV);
        ctx->launches++;
        return MMADMM_OK;
    }
    int rc = halo_exchange<D>(ctx);
    if (rc) return rc;
    k_xfinish<D, ZSRC, PRIMAL><<<ctx->nblkSh, TPB_V, 0, ctx->stream>>>(ctx->nShared, ctx->sharedIdx, ctx->sTot, ctx->Es, ctx->incPtr,
                                                                      ctx->incIdx, ctx->z, zsrc, ctx->xBar, ctx->tdiag, ctx->m, dtsq,
                                                                      ctx->x, ctx->partPrimal + ctx->nblkV);
    ctx->launches++;
    return MMADMM_OK;
}
// x-update (+ primal residual partials).  sharedOnly: the other vertices were (or, with `mid`, are being) solved by k_xfirst;
// `mid` is launched between the push of the shared vertices' partials and the wait for the neighbours'.
struct NoMid { void operator()() const {} };
template <int D, bool ZSRC, bool UZERO, bool PRIMAL, bool UREC = false, class Mid = NoMid>
int launch_xupdate_t(mmadmm_ctx *ctx, const double *zsrc, bool sharedOnly = false, Mid mid = Mid()) {
    const bool exch = ctx->dist && ctx->nShared > 0;
    unsigned long long ep = 0;
    int rc = MMADMM_OK;
    if (exch && ctx->p2p) {
        rc = xupdate_shared<D, ZSRC, UZERO, PRIMAL, UREC>(ctx, zsrc, true, ep);
        if (rc) return rc;
        if (!sharedOnly) rc = xupdate_bulk<D, ZSRC, UZERO, PRIMAL, UREC>(ctx, zsrc, true);
        if (rc) return rc;
        mid();
        return xupdate_finish<D, ZSRC, PRIMAL>(ctx, zsrc, ep);
    }
    mid();
    if (!sharedOnly) rc = xupdate_bulk<D, ZSRC, UZERO, PRIMAL, UREC>(ctx, zsrc, false);   // shared vertices: partial sums into sPart
    else if (exch) rc = xupdate_shared<D, ZSRC, UZERO, PRIMAL, UREC>(ctx, zsrc, false, ep);
    if (rc || !exch) return rc;
    return xupdate_finish<D, ZSRC, PRIMAL>(ctx, zsrc, ep);
}

template <int D> int step_impl(mmadmm_ctx *ctx, int nIters, double tol, double *IhOut, int *itersOut, double *primalOut, double *dualOut) {
    cudaStream_t st = ctx->stream;
    const int P = ctx->P;
    const size_t NV = (size_t)P * D;
    const bool fast = ctx->usumPeriod > 0;   // x-updates on the per-vertex dual sums (k_xfirst, k_xupdate<UREC>)
    // Single GPU: the previous step's last k_xupdate already wrote this step's xBar and first x-solve (StepTail).  Taking them
    // over is a rotation of buffers: xPrev <- x, x <- xSpec, xBar <- xBarSpec -- exactly what k_xfirst<EXTRAP> would store.
    const bool tailOk = fast && !ctx->dist && !(ctx->projMask & 2);
    const bool specUse = ctx->specValid && tailOk && !ctx->gradUse && ctx->stepsTaken > 2;
    ctx->specValid = false;
    if (specUse) {
        double *t = ctx->xPrev;
        ctx->xPrev = ctx->x; ctx->x = ctx->xSpec; ctx->xSpec = t;
        std::swap(ctx->xBar, ctx->xBarSpec);
        ctx->specUsed++;
    }
    ElemArgs<D> a = ctx->elemArgs<D>(tol / 100);   // BFGS tolerance tol/100, 50 iterations (src/Mesh.cpp:968)
    const int gE = ctx->nblkE, gV = ctx->nblkV;
    const bool ovl = ctx->dist && ctx->nShared > 0 && ctx->p2p;   // exchange overlapped with the bulk vertex passes
    bool extrapFused = false;

    // predictX (src/Mesh.cpp:649-674) fused with xPrev = x (src/MeshIntegrator.cpp:119)
    ctx->phaseBegin(0);
    if (ctx->gradUse || ctx->stepsTaken <= 2) {
        k_euler_elem<D, false><<<gE, TPB, 0, st>>>(a, ctx->gs, a.fixedBits, nullptr);
        k_predict_grad<D><<<gV, TPB_V, 0, st>>>(P, ctx->Es, ctx->incPtr, ctx->incIdx, ctx->gs, ctx->dt / ctx->tau, ctx->x, ctx->xBar,
                                                ctx->xPrev, ctx->dist ? ctx->sharedSlot : nullptr, ctx->sPart, ctx->vorder);
        ctx->launches += 2;
        if (ctx->dist && ctx->nShared > 0) {   // assembled gradient of the shared vertices
            int rcx = halo_exchange<D>(ctx);
            if (rcx) return rcx;
            k_predict_finish<D><<<cdiv(ctx->nShared, 256), 256, 0, st>>>(ctx->nShared, ctx->sharedIdx, ctx->sTot, ctx->dt / ctx->tau, ctx->x,
                                                                         ctx->xBar, ctx->xPrev);
            ctx->launches++;
        }
    } else if (fast) {   // predictor and first x-solve in one vertex-local pass (k_xfirst<EXTRAP>), launched in the next section
        extrapFused = true;
        if (ovl) {       // the shared vertices' predictor first: their partial sums are pushed before the bulk pass
            k_predict_shared<D><<<cdiv(ctx->nShared, 256), 256, 0, st>>>(ctx->nShared, ctx->sharedIdx, ctx->x, ctx->xBar, ctx->xPrev);
            ctx->launches += 1;
        }
    } else {
        k_predict_extrap<<<cdiv(NV, TPB_V), TPB_V, 0, st>>>(NV, ctx->x, ctx->xBar, ctx->xPrev);
        ctx->launches += 1;
    }
    ctx->phaseEnd();
    // x = xBar; z = D x (z = D xPrev on the very first step); u = 0 until a step has been taken (:120-127).
    // z is not materialised: the first x-solve and the first prox sweep gather it from the vertex vector `zsrc`.
    const double *zsrc = (ctx->stepsTaken == 0) ? ctx->xPrev : ctx->xBar;
    const bool uZero = !ctx->stepTaken;
    // first x-solve (:137-138)
    ctx->phaseBegin(2);
    {
        auto xfirst = [&]() {   // all vertices that are not shared with another rank; no element state is read
            if (!fast || specUse) return;
            const int *ss = ctx->dist ? ctx->sharedSlot : nullptr;
            if (extrapFused)
                k_xfirst<D, true><<<cdiv(NV, TPB_V), TPB_V, 0, st>>>(NV, ctx->incPtr, ctx->usum, nullptr, ctx->xBar, ctx->xPrev, ctx->tdiag,
                                                                     ctx->m, ctx->dt * ctx->dt, ctx->w, ctx->x, ss, ovl ? 1 : 0);
            else
                k_xfirst<D, false><<<cdiv(NV, TPB_V), TPB_V, 0, st>>>(NV, ctx->incPtr, ctx->usum, zsrc, ctx->xBar, ctx->xPrev, ctx->tdiag,
                                                                      ctx->m, ctx->dt * ctx->dt, ctx->w, ctx->x, ss, 0);
            ctx->launches += 1;
        };
        int rcx = uZero ? launch_xupdate_t<D, true, true, false, false>(ctx, zsrc, fast, xfirst)
                        : launch_xupdate_t<D, true, false, false, false>(ctx, zsrc, fast, xfirst);
        if (rcx) return rcx;
    }
    ctx->phaseEnd();

    double Ihstart = 0, primal = 0, dual = 0;
    int i;
    bool specWritten = false;
    for (i = 0; i < nIters; i++) {
        a.zsrc = (i == 0) ? zsrc : nullptr;
        a.uZero = (i == 0 && uZero) ? 1 : 0;
        if (!ctx->hessComputed) {
            ctx->phaseBegin(4);
            k_hess_init<D><<<gE, TPB, 0, st>>>(a);
            ctx->phaseEnd();
            ctx->launches++;
        }
        ctx->phaseBegin(1);
        int rcl = launch_prox<D>(ctx, a);
        if (rcl) return rcl;
        ctx->phaseEnd();
        ctx->hessComputed = true;
        ctx->stepTaken = true;
        ctx->phaseBegin(2);
        bool fusedFin = false;
        {
            const bool rec = fast && ctx->usumAge + 1 < ctx->usumPeriod;
            ctx->usumAge = rec ? ctx->usumAge + 1 : 0;
            int rcx = MMADMM_OK;
            specWritten = false;
            if (rec && tailOk && ctx->tailMode) {   // the step's last kernel (see StepTail)
                StepTail tl;
                if (ctx->tailMode & 1) { tl.xPrev = ctx->xPrev; tl.xNext = ctx->xSpec; tl.xBarNext = ctx->xBarSpec; specWritten = true; }
                if (ctx->tailMode & 2) {
                    tl.nProx = ctx->nProxPart; tl.partIh = ctx->partIh; tl.partDual = ctx->partDual; tl.out = ctx->scal_hd;
                    tl.statusIn = ctx->status_d; tl.statusOut = ctx->status_hd; tl.ticket = ctx->finTicket;
                    fusedFin = true;
                }
                rcx = xupdate_bulk<D, false, false, true, true>(ctx, nullptr, false, tl);
            } else {
                rcx = rec ? launch_xupdate_t<D, false, false, true, true>(ctx, nullptr) : launch_xupdate_t<D, false, false, true>(ctx, nullptr);
            }
            if (rcx) return rcx;
        }
        ctx->phaseEnd();
        ctx->phaseBegin(3);
        const int nbp = ctx->nProxPart;
        int rc = MMADMM_OK;
        if (fusedFin) {   // the scalars and the status word are in mapped host memory when the stream has drained
            CK(cudaStreamSynchronize(st));
        } else {          // synchronises; sums over ranks on a distributed context
            rc = finalize_scalars(ctx, nbp, ctx->partIh, gV + ((ctx->dist && ctx->nShared > 0) ? ctx->nblkSh : 0), ctx->partPrimal, nbp,
                                  ctx->partDual, 0);
        }
        ctx->phaseEnd();
        if (rc) return rc;
        ctx->admmIters++;
        ctx->proxCalls += (long long)ctx->E;
        rc = check_status(ctx);
        if (rc) return rc;
        if (i == 0) Ihstart = ctx->scal_h[0];
        primal = std::sqrt(ctx->scal_h[1]);
        dual = std::sqrt(ctx->scal_h[2]);
        if (primal < tol && dual < tol) break;
    }
    if (ctx->evUsed >= 4096) ctx->phaseResolve();
    if (ctx->projMask & 2) {   // updateAfterStep's projection of the FREE nodes (option), face geometry from the old Vp = xPrev
        k_project_x<D><<<cdiv(P, 256), 256, 0, st>>>(P, ctx->mask, a.fa, ctx->x);
        ctx->launches++;
    }
    CK(cudaGetLastError());
    ctx->stepsTaken++;
    ctx->specValid = specWritten;
    if (ctx->wb.mode) {   // updateAfterStep: Vp <- x
        int rcw = wb_issue(ctx);
        if (rcw) return rcw;
    }
    if (IhOut) *IhOut = Ihstart;
    if (itersOut) *itersOut = (i == nIters) ? nIters : i + 1;
    if (primalOut) *primalOut = primal;
    if (dualOut) *dualOut = dual;
    return MMADMM_OK;
}

template <int D> int energy_impl(mmadmm_ctx *ctx, double *I) {
    cudaStream_t st = ctx->stream;
    ElemArgs<D> a = ctx->elemArgs<D>(0.0);
    k_energy<D><<<ctx->nblkE, TPB, 0, st>>>(a, ctx->partPrimal);
    ctx->launches += 1;
    int rcg = finalize_scalars(ctx, ctx->nblkE, ctx->partPrimal, 0, nullptr, 0, nullptr, 3);   // -> scal[3]
    if (rcg) return rcg;
    CK(cudaGetLastError());
    *I = ctx->scal_h[3];
    return check_status(ctx);
}

// Method 1: MeshIntegrator::eulerStep (src/MeshIntegrator.cpp:87-95) = Mesh::eulerStepMod (src/Mesh.cpp:533-579) + x -= (dt/tau) grad.
// Returns the energy at entry.  K6a with the INTERIOR-only mask, then K6b writing into x itself.
template <int D> int euler_impl(mmadmm_ctx *ctx, double *IhOut) {
    cudaStream_t st = ctx->stream;
    ElemArgs<D> a = ctx->elemArgs<D>(0.0);
    k_euler_elem<D, true><<<ctx->nblkE, TPB, 0, st>>>(a, ctx->gs, a.nonIntBits, ctx->partPrimal);
    k_predict_grad<D><<<ctx->nblkV, TPB_V, 0, st>>>(ctx->P, ctx->Es, ctx->incPtr, ctx->incIdx, ctx->gs, ctx->dt / ctx->tau, ctx->x, ctx->x,
                                                    ctx->xPrev, ctx->dist ? ctx->sharedSlot : nullptr, ctx->sPart, ctx->vorder);
    ctx->launches += 2;
    if (ctx->dist && ctx->nShared > 0) {
        int rcx = halo_exchange<D>(ctx);
        if (rcx) return rcx;
        k_predict_finish<D><<<cdiv(ctx->nShared, 256), 256, 0, st>>>(ctx->nShared, ctx->sharedIdx, ctx->sTot, ctx->dt / ctx->tau, ctx->x,
                                                                     ctx->x, ctx->xPrev);
        ctx->launches++;
    }
    int rcg = finalize_scalars(ctx, ctx->nblkE, ctx->partPrimal, 0, nullptr, 0, nullptr, 3);
    if (rcg) return rcg;
    CK(cudaGetLastError());
    if (IhOut) *IhOut = ctx->scal_h[3];
    return check_status(ctx);
}

// ---- Method 2: MeshIntegrator::backwardsEulerStep (src/MeshIntegrator.cpp:68-76) -> Mesh::backwardsEulerStep (src/Mesh.cpp:1263-1341) ----
int m2_init(mmadmm_ctx *ctx) {
    mmadmm_ctx::M2 &m = ctx->m2;
    const int D = ctx->D, V = D + 1, P = ctx->P, E = ctx->E;
    // vertex adjacency = the structure Mesh::buildMatrix gives LASolver (src/Mesh.cpp:305-340): every vertex couples to all
    // vertices it shares a simplex with (itself included); D x D dense blocks
    std::vector<int> Fh((size_t)E * V), plane(E);
    CK(cudaStreamSynchronize(ctx->stream));
    for (int v = 0; v < V; v++) {
        CK(cudaMemcpy(plane.data(), ctx->F[v], sizeof(int) * (size_t)E, cudaMemcpyDeviceToHost));
        for (int e = 0; e < E; e++) Fh[(size_t)e * V + v] = plane[e];
    }
    std::vector<size_t> ptr((size_t)P + 1, 0);
    for (size_t k = 0; k < Fh.size(); k++) ptr[Fh[k] + 1] += V;
    for (int p = 0; p < P; p++) ptr[p + 1] += ptr[p];
    std::vector<int> nb(ptr[P]);
    {
        std::vector<size_t> cur(ptr.begin(), ptr.end() - 1);
        for (int e = 0; e < E; e++)
            for (int i = 0; i < V; i++)
                for (int j = 0; j < V; j++) nb[cur[Fh[(size_t)e * V + i]]++] = Fh[(size_t)e * V + j];
    }
    std::vector<int> adjPtr(P + 1, 0), adjIdx;
#pragma omp parallel for schedule(static)
    for (int p = 0; p < P; p++) {   // sorted, de-duplicated in place; the count goes to adjPtr[p + 1]
        auto b = nb.begin() + ptr[p], e2 = nb.begin() + ptr[p + 1];
        std::sort(b, e2);
        const int cnt = (int)(std::unique(b, e2) - b);
        adjPtr[p + 1] = cnt ? cnt : 1;   // a vertex no element uses still needs its diagonal
    }
    for (int p = 0; p < P; p++) adjPtr[p + 1] += adjPtr[p];
    adjIdx.resize((size_t)adjPtr[P]);
#pragma omp parallel for schedule(static)
    for (int p = 0; p < P; p++) {
        const int cnt = adjPtr[p + 1] - adjPtr[p];
        if (ptr[p + 1] == ptr[p]) adjIdx[adjPtr[p]] = p;
        else std::copy(nb.begin() + ptr[p], nb.begin() + ptr[p] + cnt, adjIdx.begin() + adjPtr[p]);
    }
    const size_t nnz = (size_t)D * D * adjIdx.size();
    if (nnz > (size_t)std::numeric_limits<int>::max()) { ctx->err = "Method 2: Jacobian has more than 2^31 entries"; return MMADMM_EINVAL; }
    m.N = P * D; m.nnz = (int)nnz;
    m.rowptr_h.resize((size_t)m.N + 1); m.col_h.resize(nnz); m.val_h.resize(nnz);
#pragma omp parallel for schedule(static)
    for (int p = 0; p < P; p++) {
        const int a0 = adjPtr[p], deg = adjPtr[p + 1] - a0;
        for (int q = 0; q < D; q++) {
            const size_t r0 = (size_t)D * D * a0 + (size_t)q * D * deg;
            m.rowptr_h[(size_t)D * p + q] = (int)r0;
            for (int t = 0; t < deg; t++)
                for (int c = 0; c < D; c++) m.col_h[r0 + (size_t)D * t + c] = D * adjIdx[a0 + t] + c;
        }
    }
    m.rowptr_h[m.N] = (int)nnz;
    CK(cudaMalloc(&m.adjPtr_d, sizeof(int) * (size_t)(P + 1)));
    CK(cudaMalloc(&m.adjIdx_d, sizeof(int) * adjIdx.size()));
    CK(cudaMemcpy(m.adjPtr_d, adjPtr.data(), sizeof(int) * (size_t)(P + 1), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(m.adjIdx_d, adjIdx.data(), sizeof(int) * adjIdx.size(), cudaMemcpyHostToDevice));
    const size_t vb = sizeof(double) * (size_t)m.N;
    CK(cudaMalloc(&m.val_d, sizeof(double) * nnz));
    CK(cudaMalloc(&m.vp0, vb)); CK(cudaMalloc(&m.xn, vb)); CK(cudaMalloc(&m.rhs, vb)); CK(cudaMalloc(&m.dx, vb));
    m.nblk = cdiv(P, TPB_V);
    CK(cudaMalloc(&m.part, sizeof(double) * (size_t)m.nblk));
    m.init = true;
    return MMADMM_OK;
}

// Jacobian values on the device, then the ILU(0) factor + level schedules (mmadmm_pcg_create): jac->sfac + the numeric factor
template <int D> int m2_build_jac(mmadmm_ctx *ctx, double dt) {
    mmadmm_ctx::M2 &m = ctx->m2;
    if (m.jacValid && m.jacDt == dt) return MMADMM_OK;   // Vp and dt unchanged: a rebuild would reproduce the same values
    cudaStream_t st = ctx->stream;
    ElemArgs<D> a = ctx->elemArgs<D>(0.0);
    a.x = m.vp0;
    k_m2_jac<D><<<cdiv(ctx->P, 128), 128, 0, st>>>(a, ctx->mask, ctx->incPtr, ctx->incIdx, m.adjPtr_d, m.adjIdx_d, dt / ctx->tau, m.val_d);
    ctx->launches++;
    if (m.solver) {
        // a rebuild: same pattern, level schedules and buffers; numeric ILU(0) re-factorisation on the device (jac->factor)
        CK(cudaStreamSynchronize(st));
        CK(cudaGetLastError());
        int rcr = mmadmm_pcg_refactor_device(m.solver, m.val_d);
        if (rcr) { ctx->err = std::string("Method 2 Jacobian: ") + mmadmm_pcg_last_error(m.solver); return rcr; }
        m.jacValid = true; m.jacDt = dt; m.jacBuilds++;
        return MMADMM_OK;
    }
    CK(cudaMemcpyAsync(m.val_h.data(), m.val_d, sizeof(double) * (size_t)m.nnz, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    int rc = mmadmm_pcg_create(m.N, m.rowptr_h.data(), m.col_h.data(), m.val_h.data(), 1, ctx->device, &m.solver);
    if (rc) {
        ctx->err = std::string("Method 2 Jacobian: ") + (m.solver ? mmadmm_pcg_last_error(m.solver) : "solver allocation failed");
        if (m.solver) { mmadmm_pcg_destroy(m.solver); m.solver = nullptr; }
        return rc;
    }
    rc = mmadmm_pcg_set_option(m.solver, "tri_schedule", (m.rowBlock < 0 ? D : m.rowBlock) | (m.triLayout << 8) | ((m.triMode + 1) << 12));
    if (rc) { ctx->err = std::string("Method 2 tri_schedule: ") + mmadmm_pcg_last_error(m.solver); return rc; }
    m.jacValid = true; m.jacDt = dt; m.jacBuilds++;
    return MMADMM_OK;
}

template <int D> int beuler_impl(mmadmm_ctx *ctx, double dt, double tol, double *IhOut, int *newtonOut, int *linOut) {
    if (ctx->dist) { ctx->err = "Method 2 (backward Euler) runs on single-GPU contexts only"; return MMADMM_EINVAL; }
    mmadmm_ctx::M2 &m = ctx->m2;
    cudaStream_t st = ctx->stream;
    if (!m.init) { int rc = m2_init(ctx); if (rc) return rc; }
    const size_t vb = sizeof(double) * (size_t)m.N;
    if (m.vpStamp != ctx->stepsTaken) {
        // Vp as the reference holds it: the positions at construction or after the last Method-0 step (updateAfterStep);
        // Methods 1 and 2 never refresh it, so the snapshot is taken when this method is first entered after either
        CK(cudaMemcpyAsync(m.vp0, ctx->x, vb, cudaMemcpyDeviceToDevice, st));
        m.vpStamp = ctx->stepsTaken;
        m.jacValid = false;
    }
    const double s = dt / ctx->tau;
    const double SAFETY_FAC = 1.0 / 10.0;
    const int MAX_ITERS = 1000;
    ElemArgs<D> a = ctx->elemArgs<D>(0.0);
    CK(cudaMemcpyAsync(m.xn, ctx->x, vb, cudaMemcpyDeviceToDevice, st));
    // initial guess: one explicit Euler step
    k_euler_elem<D, true><<<ctx->nblkE, TPB, 0, st>>>(a, ctx->gs, a.nonIntBits, ctx->partPrimal);
    k_m2_gather<D, 0><<<m.nblk, TPB_V, 0, st>>>(ctx->P, ctx->Es, ctx->incPtr, ctx->incIdx, ctx->gs, s, ctx->x, m.xn, m.rhs, m.part);
    ctx->launches += 2;
    if (!m.stepTaken) { int rc = m2_build_jac<D>(ctx, dt); if (rc) return rc; }
    int nIter = 0, lin = 0;
    double g1Prev = std::numeric_limits<double>::infinity(), Ih = 0.0;
    do {
        k_euler_elem<D, true><<<ctx->nblkE, TPB, 0, st>>>(a, ctx->gs, a.nonIntBits, ctx->partPrimal);
        k_m2_gather<D, 1><<<m.nblk, TPB_V, 0, st>>>(ctx->P, ctx->Es, ctx->incPtr, ctx->incIdx, ctx->gs, s, ctx->x, m.xn, m.rhs, m.part);
        ctx->launches += 2;
        int rcg = finalize_scalars(ctx, m.nblk, m.part, ctx->nblkE, ctx->partPrimal, 0, nullptr, 2);
        if (rcg) return rcg;
        CK(cudaGetLastError());
        const double g1 = ctx->scal_h[2];
        Ih = ctx->scal_h[3];
        int rcs = check_status(ctx);
        if (rcs) return rcs;
        if (!(g1 == g1)) { ctx->err = "Method 2: non-finite Newton residual"; return MMADMM_ESINGULAR; }
        if (g1 < SAFETY_FAC * tol) break;
        if (!m.stepTaken || std::fabs(g1 - g1Prev) / g1 < 0.25) {
            int rc = m2_build_jac<D>(ctx, dt);
            if (rc) return rc;
            m.stepTaken = true;
        }
        int cgIter = 0;
        int rc = mmadmm_pcg_bicgstab_device(m.solver, m.rhs, m.dx, 1.e-6, 10000, &cgIter, nullptr);
        if (rc) { ctx->err = std::string("Method 2 linear solve: ") + mmadmm_pcg_last_error(m.solver); return rc; }
        if (cgIter <= 0) {   // the reference asserts cgIter > 0 (src/Mesh.cpp:1329); LASolver returns -1 when not converged
            ctx->err = "Method 2: BiCGSTAB did not converge (the reference asserts cgIter > 0, src/Mesh.cpp:1329)";
            return MMADMM_ESINGULAR;
        }
        lin += cgIter;
        k_m2_axpy<<<cdiv((size_t)m.N, 256), 256, 0, st>>>((size_t)m.N, m.dx, ctx->x);
        ctx->launches++;
        nIter++;
        g1Prev = g1;
    } while (nIter < MAX_ITERS);
    CK(cudaStreamSynchronize(st));
    m.newtonIters += nIter; m.linIters += lin;
    if (IhOut) *IhOut = Ih;
    if (newtonOut) *newtonOut = nIter;
    if (linOut) *linOut = lin;
    return MMADMM_OK;
}

}  // namespace

// =====================================================================================================
// C ABI
// =====================================================================================================
extern "C" {

const char *mmadmm_version(void) { return "mmadmm_b200 0.1 (sm_100a, fp64)"; }

int mmadmm_create(int D, int P, int E, const double *Vp, const double *Vc, const int *F, const int *mask,
                  const double *monVertex, double rho, double tau, double dt, int gradUse, int device, mmadmm_ctx **out) {
    if (!out) return MMADMM_EINVAL;
    *out = nullptr;
    if ((D != 2 && D != 3) || P <= 0 || E <= 0 || !Vp || !F || !mask || !monVertex) return MMADMM_EINVAL;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return MMADMM_ENODEVICE;
    mmadmm_ctx *ctx = new mmadmm_ctx();
    *out = ctx;  // returned even on failure so that mmadmm_last_error can be read; caller destroys it
    ctx->D = D; ctx->n = D * (D + 1); ctx->P = P; ctx->E = E; ctx->device = device;
    ctx->compMesh = (Vc != nullptr); ctx->gradUse = gradUse != 0;
    ctx->rho = rho; ctx->tau = tau; ctx->dt = dt;
    CK(cudaSetDevice(device));
    CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    for (size_t p = 0; p < (size_t)E * (D + 1); p++)
        if (F[p] < 0 || F[p] >= P) { ctx->err = "connectivity index out of range"; return MMADMM_EINVAL; }
    InitGlobals G{E, nullptr, P, Vp, monVertex};
    int rc = (D == 2) ? create_impl<2>(ctx, Vp, Vc, F, mask, G) : create_impl<3>(ctx, Vp, Vc, F, mask, G);
    return rc;
}

int mmadmm_comm_unique_id(void *id128) {
    if (!id128) return MMADMM_EINVAL;
    NcclApi &N = nccl();
    if (!N.ok) return MMADMM_ENCCL;
    NcclIdBlob id;
    memset(&id, 0, sizeof(id));
    if (N.getUniqueId(&id) != 0) return MMADMM_ENCCL;
    memcpy(id128, &id, sizeof(id));
    return MMADMM_OK;
}

int mmadmm_create_dist(int D, int P, int E, const double *Vp, const double *Vc, const int *F, const int *mask,
                       const double *monVertex, double rho, double tau, double dt, int gradUse, int device, int rank, int nranks,
                       const void *id128, mmadmm_ctx **out) {
    if (!out) return MMADMM_EINVAL;
    *out = nullptr;
    if ((D != 2 && D != 3) || P <= 0 || E <= 0 || !Vp || !F || !mask || !monVertex || nranks < 1 || rank < 0 || rank >= nranks)
        return MMADMM_EINVAL;
    if (nranks == 1) return mmadmm_create(D, P, E, Vp, Vc, F, mask, monVertex, rho, tau, dt, gradUse, device, out);
    if (!id128) return MMADMM_EINVAL;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return MMADMM_ENODEVICE;
    mmadmm_ctx *ctx = new mmadmm_ctx();
    *out = ctx;
    NcclApi &N = nccl();
    if (!N.ok) { ctx->err = "libnccl.so.2 could not be loaded"; return MMADMM_ENCCL; }
    PartitionPlan plan;
    if (!build_partition_plan(D, P, E, F, nranks, rank, plan)) { ctx->err = "bad mesh or partition arguments"; return MMADMM_EINVAL; }
    const int Pl = (int)plan.verts.size(), El = plan.e1 - plan.e0;
    if (El <= 0) { ctx->err = "empty element block: more ranks than 128-element tiles"; return MMADMM_EINVAL; }
    ctx->D = D; ctx->n = D * (D + 1); ctx->P = Pl; ctx->E = El; ctx->device = device;
    ctx->Pglobal = P; ctx->Eglobal = E; ctx->rank = rank; ctx->nranks = nranks; ctx->dist = true;
    ctx->compMesh = (Vc != nullptr); ctx->gradUse = gradUse != 0;
    ctx->rho = rho; ctx->tau = tau; ctx->dt = dt;
    CK(cudaSetDevice(device));
    CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    cudaStream_t st = ctx->stream;
    {   // communicator: ncclCommInitRank(&comm, nranks, id (by value), rank)
        typedef int (*InitFn)(void **, int, NcclIdBlob, int);
        NcclIdBlob id;
        memcpy(&id, id128, sizeof(id));
        NK(((InitFn)N.commInitRank)(&ctx->comm, nranks, id, rank));
    }
    // local mesh
    std::vector<double> VpL((size_t)Pl * D), VcL;
    std::vector<int> maskL((size_t)Pl);
    if (Vc) VcL.resize((size_t)Pl * D);
    for (int l = 0; l < Pl; l++) {
        const int g = plan.verts[l];
        for (int k = 0; k < D; k++) {
            VpL[(size_t)l * D + k] = Vp[(size_t)g * D + k];
            if (Vc) VcL[(size_t)l * D + k] = Vc[(size_t)g * D + k];
        }
        maskL[l] = mask[g];
    }
    // halo structures
    ctx->nShared = (int)plan.shared.size();
    ctx->nNbrItems = (int)plan.nbrList.size();
    ctx->nbrRank = plan.nbrRank;
    ctx->nbrPtr = plan.nbrPtr;
    ctx->nbrListHost = plan.nbrList;
    auto upload = [&](int **dst, const std::vector<int> &v) -> cudaError_t {
        cudaError_t e = cudaMalloc(dst, sizeof(int) * std::max<size_t>(v.size(), 1));
        if (e != cudaSuccess) return e;
        return cudaMemcpyAsync(*dst, v.data(), sizeof(int) * v.size(), cudaMemcpyHostToDevice, st);
    };
    std::vector<int> slot((size_t)Pl, -1);
    for (int i = 0; i < ctx->nShared; i++) slot[plan.shared[i]] = i;
    CK(upload(&ctx->sharedIdx, plan.shared));
    CK(upload(&ctx->sharedSlot, slot));
    CK(upload(&ctx->nbrList, plan.nbrList));
    CK(upload(&ctx->contribPtr, plan.contribPtr));
    CK(upload(&ctx->contrib, plan.contrib));
    CK(upload(&ctx->verts, plan.verts));
    CK(cudaMalloc(&ctx->owner, std::max<size_t>(Pl, 1)));
    CK(cudaMemcpyAsync(ctx->owner, plan.owner.data(), (size_t)Pl, cudaMemcpyHostToDevice, st));
    const size_t shD = (size_t)std::max(ctx->nShared, 1) * D, nbD = (size_t)std::max(ctx->nNbrItems, 1) * D;
    CK(cudaMalloc(&ctx->sPart, sizeof(double) * shD));
    CK(cudaMalloc(&ctx->sTot, sizeof(double) * shD));
    CK(cudaMalloc(&ctx->sendBuf, sizeof(double) * nbD));
    CK(cudaMalloc(&ctx->recvBuf, sizeof(double) * nbD));
    CK(cudaMalloc(&ctx->scalAll_d, sizeof(double) * 5 * nranks));
    CK(cudaHostAlloc(&ctx->scalAll_h, sizeof(double) * 5 * nranks, cudaHostAllocMapped));
    CK(cudaHostGetDevicePointer(&ctx->scalAll_hd, ctx->scalAll_h, 0));
    ctx->nblkSh = cdiv(std::max(ctx->nShared, 1), TPB_V);
    if (const char *pt = getenv("MMADMM_PEER_TIMEOUT_S")) {
        const long v = atol(pt);
        if (v > 0) ctx->peerTimeoutNs = (unsigned long long)v * 1000000000ull;
    }
    CK(cudaStreamSynchronize(st));
    {
        int rcp = setup_p2p(ctx, D);
        if (rcp) return rcp;
    }
    InitGlobals G{E, plan.valence.data(), P, Vp, monVertex};
    int rc = (D == 2) ? create_impl<2>(ctx, VpL.data(), Vc ? VcL.data() : nullptr, plan.Floc.data(), maskL.data(), G)
                      : create_impl<3>(ctx, VpL.data(), Vc ? VcL.data() : nullptr, plan.Floc.data(), maskL.data(), G);
    return rc;
}

int mmadmm_local_sizes(mmadmm_ctx *ctx, int *Ploc, int *Eloc) {
    if (!ctx) return MMADMM_EINVAL;
    if (Ploc) *Ploc = ctx->P;
    if (Eloc) *Eloc = ctx->E;
    return MMADMM_OK;
}

int mmadmm_set_dt(mmadmm_ctx *ctx, double dt) {
    if (!ctx || !(dt >= 0)) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    ctx->dt = dt;
    ctx->specValid = false;
    k_tdiag<<<cdiv(ctx->P, 256), 256, 0, ctx->stream>>>(ctx->P, ctx->incPtr, ctx->valence_d, ctx->m, dt * dt, ctx->w, ctx->tdiag);
    ctx->launches++;
    CK(cudaStreamSynchronize(ctx->stream));
    return MMADMM_OK;
}

int mmadmm_step(mmadmm_ctx *ctx, int nIters, double tol, double *Ih, int *itersTaken, double *primal, double *dual) {
    if (!ctx) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    return (ctx->D == 2) ? step_impl<2>(ctx, nIters, tol, Ih, itersTaken, primal, dual)
                         : step_impl<3>(ctx, nIters, tol, Ih, itersTaken, primal, dual);
}

int mmadmm_euler_step(mmadmm_ctx *ctx, double *Ih) {
    if (!ctx) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    ctx->specValid = false;
    return (ctx->D == 2) ? euler_impl<2>(ctx, Ih) : euler_impl<3>(ctx, Ih);
}

int mmadmm_backward_euler_step(mmadmm_ctx *ctx, double dt, double tol, double *Ih, int *newtonIters, int *linIters) {
    if (!ctx || !(tol > 0.0)) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    if (!(dt > 0.0)) dt = ctx->dt;
    ctx->specValid = false;
    return (ctx->D == 2) ? beuler_impl<2>(ctx, dt, tol, Ih, newtonIters, linIters) : beuler_impl<3>(ctx, dt, tol, Ih, newtonIters, linIters);
}

int mmadmm_backward_euler_jacobian(mmadmm_ctx *ctx, int *sizes2, int *rowptr, int *col, double *val) {
    if (!ctx) return MMADMM_EINVAL;
    if (ctx->dist) { ctx->err = "Method 2 (backward Euler) runs on single-GPU contexts only"; return MMADMM_EINVAL; }
    CK(cudaSetDevice(ctx->device));
    mmadmm_ctx::M2 &m = ctx->m2;
    if (!m.init) { int rc = m2_init(ctx); if (rc) return rc; }
    if (val && !m.jacValid) {
        if (m.vpStamp != ctx->stepsTaken) {
            CK(cudaMemcpyAsync(m.vp0, ctx->x, sizeof(double) * (size_t)m.N, cudaMemcpyDeviceToDevice, ctx->stream));
            m.vpStamp = ctx->stepsTaken;
        }
        int rc = (ctx->D == 2) ? m2_build_jac<2>(ctx, ctx->dt) : m2_build_jac<3>(ctx, ctx->dt);
        if (rc) return rc;
    }
    if (sizes2) { sizes2[0] = m.N; sizes2[1] = m.nnz; }
    if (rowptr) std::memcpy(rowptr, m.rowptr_h.data(), sizeof(int) * ((size_t)m.N + 1));
    if (col) std::memcpy(col, m.col_h.data(), sizeof(int) * (size_t)m.nnz);
    if (val) CK(cudaMemcpy(val, m.val_d, sizeof(double) * (size_t)m.nnz, cudaMemcpyDeviceToHost));   // val_h is only the first build's copy
    return MMADMM_OK;
}

int mmadmm_energy(mmadmm_ctx *ctx, double *I) {
    if (!ctx || !I) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    return (ctx->D == 2) ? energy_impl<2>(ctx, I) : energy_impl<3>(ctx, I);
}

// ------------------------------------------------------------------------------------------------
// Allocation-free energy evaluation (no counterpart in the reference; see include/mmadmm.h).
// k_energy's per-block partial sums go into ONE file-scope device array instead of a cudaMalloc'ed buffer.  Every GPU that loads
// the library holds the array (1 MiB + 256 B) until exit; every call on a device shares it, so calls that overlap on the GPU race.
// Nothing in it has to start at zero: a call overwrites partials [0, nblk) and reads back exactly those.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_energy_scratch[MMADMM_NOALLOC_MAX_BLOCKS + 32];

// In host code `g_energy_scratch` is only a host-side stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static double *energy_scratch_ptr(int device, cudaError_t *err) {
    static double *ptrs[64] = {};
    static pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
    *err = cudaSuccess;
    if (device < 0 || device >= 64) { *err = cudaErrorInvalidDevice; return nullptr; }
    pthread_mutex_lock(&mu);
    double *p = ptrs[device];
    if (!p) {
        int cur = -1;
        *err = cudaGetDevice(&cur);
        if (*err == cudaSuccess && cur != device) *err = cudaSetDevice(device);
        if (*err == cudaSuccess) *err = cudaGetSymbolAddress((void **)&p, g_energy_scratch);
        if (*err == cudaSuccess) ptrs[device] = p; else p = nullptr;
    }
    pthread_mutex_unlock(&mu);
    return p;
}

int mmadmm_energy_noalloc(mmadmm_ctx *ctx, const double *x_dev, double *I) {
    if (!ctx || !I) return MMADMM_EINVAL;
    if (ctx->dist) { ctx->err = "mmadmm_energy_noalloc: single-GPU contexts only"; return MMADMM_EINVAL; }
    if (ctx->nblkE > MMADMM_NOALLOC_MAX_BLOCKS) {
        char b[256];
        snprintf(b, sizeof(b), "mmadmm_energy_noalloc: %d elements need %d partial sums, the static scratch holds %d (at most %lld elements)",
                 ctx->E, ctx->nblkE, MMADMM_NOALLOC_MAX_BLOCKS, (long long)MMADMM_NOALLOC_MAX_BLOCKS * TPB);
        ctx->err = b;
        return MMADMM_EINVAL;
    }
    CK(cudaSetDevice(ctx->device));
    cudaError_t se;
    double *scratch = energy_scratch_ptr(ctx->device, &se);
    CK(se);
    cudaStream_t st = ctx->stream;
    if (ctx->D == 2) {
        ElemArgs<2> a = ctx->elemArgs<2>(0.0);
        if (x_dev) a.x = x_dev;
        k_energy<2><<<ctx->nblkE, TPB, 0, st>>>(a, scratch);
    } else {
        ElemArgs<3> a = ctx->elemArgs<3>(0.0);
        if (x_dev) a.x = x_dev;
        k_energy<3><<<ctx->nblkE, TPB, 0, st>>>(a, scratch);
    }
    // same fixed-order final reduction as mmadmm_energy; the scalar and the status word land in the context's mapped host words
    k_finalize<<<1, 1024, 0, st>>>(ctx->nblkE, scratch, 0, nullptr, 0, nullptr, ctx->scal_hd + 3, ctx->status_d, ctx->status_hd);
    ctx->launches += 2;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));
    *I = ctx->scal_h[3];
    return check_status(ctx);
}

int mmadmm_get_x(mmadmm_ctx *ctx, double *x) {
    if (!ctx || !x) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    { int rcf = wb_fence(ctx); if (rcf) return rcf; }
    if (!ctx->dist) {
        CK(cudaMemcpyAsync(x, ctx->x, sizeof(double) * (size_t)ctx->P * ctx->D, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        return MMADMM_OK;
    }
    // every rank contributes the vertices it owns (lowest rank holding them); zeros elsewhere; sum == assembly
    const size_t NG = (size_t)ctx->Pglobal * ctx->D;
    double *xg = nullptr;
    CK(cudaMalloc(&xg, sizeof(double) * NG));
    CK(cudaMemsetAsync(xg, 0, sizeof(double) * NG, ctx->stream));
    k_scatter_owned<<<cdiv((size_t)ctx->P, 256), 256, 0, ctx->stream>>>(ctx->P, ctx->D, ctx->verts, ctx->owner, ctx->x, xg);
    ctx->launches++;
    NK(nccl().allReduce(xg, xg, NG, NCCL_FLOAT64, NCCL_SUM, ctx->comm, ctx->stream));
    CK(cudaMemcpyAsync(x, xg, sizeof(double) * NG, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaFree(xg));
    return MMADMM_OK;
}
int mmadmm_get_x_colmajor(mmadmm_ctx *ctx, double *VpColMajor) {
    if (!ctx || !VpColMajor || ctx->dist) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    { int rcf = wb_fence(ctx); if (rcf) return rcf; }
    const size_t bytes = sizeof(double) * (size_t)ctx->P * ctx->D;
    if (!ctx->xT) CK(cudaMalloc(&ctx->xT, bytes));
    k_x_to_colmajor<<<cdiv(ctx->P, 256), 256, 0, ctx->stream>>>(ctx->P, ctx->D, ctx->x, ctx->xT);
    ctx->launches++;
    CK(cudaMemcpyAsync(VpColMajor, ctx->xT, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return MMADMM_OK;
}

int mmadmm_update_monitor(mmadmm_ctx *ctx, const double *monVertex) {
    if (!ctx || !monVertex) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    const int Pg = ctx->dist ? ctx->Pglobal : ctx->P;
    std::vector<double> xh((size_t)Pg * ctx->D);
    int rc = mmadmm_get_x(ctx, xh.data());     // current node positions (assembled over the ranks on a distributed context)
    if (rc) return rc;
    CK(cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->grid); cudaFree(ctx->gridIso); cudaFree(ctx->ax); cudaFree(ctx->ay); cudaFree(ctx->az);
    ctx->grid = ctx->gridIso = ctx->ax = ctx->ay = ctx->az = nullptr;
    ctx->gridIsIso = false;
    return (ctx->D == 2) ? build_monitor_grid<2>(ctx, Pg, xh.data(), monVertex) : build_monitor_grid<3>(ctx, Pg, xh.data(), monVertex);
}

int mmadmm_set_writeback(mmadmm_ctx *ctx, double *VpHost, int colMajor, int mode) {
    if (!ctx || mode < 0 || mode > 2 || (mode != 0 && !VpHost)) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    int rc = wb_fence(ctx);
    if (rc) return rc;
    if (mode != 0 && ctx->dist) {
        rc = ensure_local_runs(ctx);
        if (rc) return rc;
        if (ctx->runs.size() > 256) { ctx->err = "write-back: this rank's vertices form more than 256 runs of consecutive global ids"; return MMADMM_EINVAL; }
    }
    ctx->wb.mode = mode; ctx->wb.colMajor = colMajor ? 1 : 0; ctx->wb.host = VpHost;
    if (mode != 0) return wb_setup(ctx);
    return MMADMM_OK;
}
int mmadmm_writeback_fence(mmadmm_ctx *ctx) {
    if (!ctx) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    return wb_fence(ctx);
}

int mmadmm_checkpoint(mmadmm_ctx *ctx) {
    if (!ctx) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    mmadmm_ctx::Ckpt &c = ctx->ck;
    const size_t NV = sizeof(double) * (size_t)ctx->P * ctx->D, NZ = sizeof(double) * ctx->Es * ctx->n;
    if (!c.x) {
        CK(cudaMalloc(&c.x, NV)); CK(cudaMalloc(&c.xPrev, NV)); CK(cudaMalloc(&c.xBar, NV)); CK(cudaMalloc(&c.usum, NV));
        CK(cudaMalloc(&c.z, NZ)); CK(cudaMalloc(&c.u, NZ)); CK(cudaMalloc(&c.B, NZ * ctx->n));
        CK(cudaMalloc(&c.hist, sizeof(unsigned long long) * 52));
    }
    cudaStream_t st = ctx->stream;
    CK(cudaMemcpyAsync(c.x, ctx->x, NV, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(c.xPrev, ctx->xPrev, NV, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(c.xBar, ctx->xBar, NV, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(c.usum, ctx->usum, NV, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(c.z, ctx->z, NZ, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(c.u, ctx->u, NZ, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(c.B, ctx->B, NZ * ctx->n, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(c.hist, ctx->hist_d, sizeof(unsigned long long) * 52, cudaMemcpyDeviceToDevice, st));
    CK(cudaStreamSynchronize(st));
    c.stepTaken = ctx->stepTaken; c.hessComputed = ctx->hessComputed; c.stepsTaken = ctx->stepsTaken; c.usumAge = ctx->usumAge;
    c.admmIters = ctx->admmIters; c.proxCalls = ctx->proxCalls;
    c.valid = true;
    return MMADMM_OK;
}
int mmadmm_restore(mmadmm_ctx *ctx) {
    if (!ctx) return MMADMM_EINVAL;
    mmadmm_ctx::Ckpt &c = ctx->ck;
    if (!c.valid) { ctx->err = "mmadmm_restore: no checkpoint has been taken"; return MMADMM_EINVAL; }
    ctx->specValid = false;
    CK(cudaSetDevice(ctx->device));
    int rc = wb_fence(ctx);
    if (rc) return rc;
    const size_t NV = sizeof(double) * (size_t)ctx->P * ctx->D, NZ = sizeof(double) * ctx->Es * ctx->n;
    cudaStream_t st = ctx->stream;
    CK(cudaMemcpyAsync(ctx->x, c.x, NV, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(ctx->xPrev, c.xPrev, NV, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(ctx->xBar, c.xBar, NV, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(ctx->usum, c.usum, NV, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(ctx->z, c.z, NZ, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(ctx->u, c.u, NZ, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(ctx->B, c.B, NZ * ctx->n, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(ctx->hist_d, c.hist, sizeof(unsigned long long) * 52, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemsetAsync(ctx->status_d, 0, sizeof(unsigned), st));
    CK(cudaStreamSynchronize(st));
    *ctx->status_h = 0;
    ctx->stepTaken = c.stepTaken; ctx->hessComputed = c.hessComputed; ctx->stepsTaken = c.stepsTaken; ctx->usumAge = c.usumAge;
    ctx->admmIters = c.admmIters; ctx->proxCalls = c.proxCalls;
    ctx->m2.vpStamp = -1; ctx->m2.jacValid = false; ctx->m2.stepTaken = false;
    return MMADMM_OK;
}

int mmadmm_host_register(void *p, unsigned long bytes) {
    if (!p || !bytes) return MMADMM_EINVAL;
    if (cudaHostRegister(p, bytes, cudaHostRegisterDefault) != cudaSuccess) { cudaGetLastError(); return MMADMM_ECUDA; }
    return MMADMM_OK;
}
int mmadmm_host_unregister(void *p) {
    if (!p) return MMADMM_EINVAL;
    if (cudaHostUnregister(p) != cudaSuccess) { cudaGetLastError(); return MMADMM_ECUDA; }
    return MMADMM_OK;
}

int mmadmm_get_x_local(mmadmm_ctx *ctx, double *xGlobal) {
    if (!ctx || !xGlobal) return MMADMM_EINVAL;
    if (!ctx->dist) return mmadmm_get_x(ctx, xGlobal);
    const int D = ctx->D;
    { int rcf = wb_fence(ctx); if (rcf) return rcf; }
    {
        int rcr = ensure_local_runs(ctx);
        if (rcr) return rcr;
    }
    if (ctx->runs.size() <= 64) {   // straight into the caller's array, one copy per run
        for (auto &r : ctx->runs)
            CK(cudaMemcpyAsync(xGlobal + (size_t)ctx->verts_h[r.first] * D, ctx->x + (size_t)r.first * D,
                               sizeof(double) * (size_t)r.second * D, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        return MMADMM_OK;
    }
    if (!ctx->xloc_h) CK(cudaMallocHost(&ctx->xloc_h, sizeof(double) * (size_t)ctx->P * D));
    CK(cudaMemcpyAsync(ctx->xloc_h, ctx->x, sizeof(double) * (size_t)ctx->P * D, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int l = 0; l < ctx->P; l++)
        for (int k = 0; k < D; k++) xGlobal[(size_t)ctx->verts_h[l] * D + k] = ctx->xloc_h[(size_t)l * D + k];
    return MMADMM_OK;
}
int mmadmm_get_x_device(mmadmm_ctx *ctx, double *x_dev) {
    if (!ctx || !x_dev) return MMADMM_EINVAL;
    CK(cudaMemcpyAsync(x_dev, ctx->x, sizeof(double) * (size_t)ctx->P * ctx->D, cudaMemcpyDeviceToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return MMADMM_OK;
}

int mmadmm_get_F(mmadmm_ctx *ctx, int *F) {
    if (!ctx || !F) return MMADMM_EINVAL;
    const int V = ctx->D + 1;
    std::vector<int> plane(ctx->E);
    for (int v = 0; v < V; v++) {
        CK(cudaMemcpy(plane.data(), ctx->F[v], sizeof(int) * (size_t)ctx->E, cudaMemcpyDeviceToHost));
        for (int e = 0; e < ctx->E; e++) F[(size_t)e * V + v] = plane[e];
    }
    return MMADMM_OK;
}

int mmadmm_grid_dims(mmadmm_ctx *ctx, int *dims) {
    if (!ctx || !dims) return MMADMM_EINVAL;
    dims[0] = ctx->nx; dims[1] = ctx->ny; dims[2] = ctx->nz;
    return MMADMM_OK;
}

int mmadmm_get_state(mmadmm_ctx *ctx, int which, double *out) {
    if (!ctx || !out) return MMADMM_EINVAL;
    CK(cudaStreamSynchronize(ctx->stream));
    const size_t E = ctx->E;
    const int n = ctx->n;
    if (which == MMADMM_STATE_Z || which == MMADMM_STATE_U || which == MMADMM_STATE_HESSINV) {
        int planes = (which == MMADMM_STATE_HESSINV) ? n * n : n;
        const double *src = (which == MMADMM_STATE_Z) ? ctx->z : (which == MMADMM_STATE_U) ? ctx->u : ctx->B;
        std::vector<double> tmp(ctx->Es * planes);   // tile-major on the device -> element-major for the caller
        CK(cudaMemcpy(tmp.data(), src, sizeof(double) * tmp.size(), cudaMemcpyDeviceToHost));
        for (size_t e = 0; e < E; e++)
            for (int k = 0; k < planes; k++) {
                size_t src_i;
                if (which == MMADMM_STATE_HESSINV) src_i = (ctx->D == 2) ? b_index<2>(e, k / n, k % n) : b_index<3>(e, k / n, k % n);
                else src_i = (ctx->D == 2) ? zu_index<2>(e, k) : zu_index<3>(e, k);
                out[e * planes + k] = tmp[src_i];
            }
        return MMADMM_OK;
    }
    if (which == MMADMM_STATE_XBAR || which == MMADMM_STATE_XPREV) {
        CK(cudaMemcpy(out, which == MMADMM_STATE_XBAR ? ctx->xBar : ctx->xPrev, sizeof(double) * (size_t)ctx->P * ctx->D, cudaMemcpyDeviceToHost));
        return MMADMM_OK;
    }
    if (which == MMADMM_STATE_MONGRID) {   // rows unpadded to D*D values
        const int DD = ctx->D * ctx->D, RS = (ctx->D == 2) ? GridRow<2>::stride : GridRow<3>::stride;
        std::vector<double> tmp(ctx->gridRows * RS);
        CK(cudaMemcpy(tmp.data(), ctx->grid, sizeof(double) * ctx->gridRows * RS, cudaMemcpyDeviceToHost));
        for (size_t r = 0; r < ctx->gridRows; r++)
            for (int k = 0; k < DD; k++) out[r * DD + k] = tmp[r * RS + k];
        return MMADMM_OK;
    }
    return MMADMM_EINVAL;
}

int mmadmm_eval_monitor(mmadmm_ctx *ctx, const double *pts, int npts, double *out) {
    if (!ctx || !pts || !out || npts <= 0) return MMADMM_EINVAL;
    const int D = ctx->D;
    double *dp = nullptr, *dout = nullptr;
    CK(cudaMalloc(&dp, sizeof(double) * (size_t)npts * D));
    CK(cudaMalloc(&dout, sizeof(double) * (size_t)npts * D * D));
    CK(cudaMemcpy(dp, pts, sizeof(double) * (size_t)npts * D, cudaMemcpyHostToDevice));
    MonGrid mg = ctx->monGrid();
    if (D == 2) k_eval_monitor<2><<<cdiv(npts, 128), 128, 0, ctx->stream>>>(mg, npts, dp, dout);
    else k_eval_monitor<3><<<cdiv(npts, 128), 128, 0, ctx->stream>>>(mg, npts, dp, dout);
    ctx->launches++;
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaGetLastError());
    CK(cudaMemcpy(out, dout, sizeof(double) * (size_t)npts * D * D, cudaMemcpyDeviceToHost));
    cudaFree(dp); cudaFree(dout);
    return MMADMM_OK;
}

int mmadmm_block_grad(mmadmm_ctx *ctx, int ne, const double *zs, const double *dxpu, int computeGrad, int regularize,
                      int applyMask, double *energy, double *grad, double *igt) {
    return mmadmm_block_grad_at(ctx, 0, ne, zs, dxpu, computeGrad, regularize, applyMask, energy, grad, igt);
}

int mmadmm_block_grad_at(mmadmm_ctx *ctx, int e0, int ne, const double *zs, const double *dxpu, int computeGrad, int regularize,
                         int applyMask, double *energy, double *grad, double *igt) {
    if (!ctx || !zs || !energy || ne <= 0 || e0 < 0 || e0 > ctx->E - ne || (regularize && !dxpu)) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    const int n = ctx->n;
    double *dz = nullptr, *dd = nullptr, *de = nullptr, *dg = nullptr, *di = nullptr;
    CK(cudaMalloc(&dz, sizeof(double) * (size_t)ne * n));
    CK(cudaMalloc(&dd, sizeof(double) * (size_t)ne * n));
    CK(cudaMalloc(&de, sizeof(double) * (size_t)ne));
    CK(cudaMalloc(&dg, sizeof(double) * (size_t)ne * n));
    CK(cudaMalloc(&di, sizeof(double) * (size_t)ne));
    CK(cudaMemcpy(dz, zs, sizeof(double) * (size_t)ne * n, cudaMemcpyHostToDevice));
    if (regularize) CK(cudaMemcpy(dd, dxpu, sizeof(double) * (size_t)ne * n, cudaMemcpyHostToDevice));
    CK(cudaMemsetAsync(ctx->status_d, 0, sizeof(unsigned), ctx->stream));
    if (ctx->D == 2) {
        ElemArgs<2> a = ctx->elemArgs<2>(0.0);
        k_block_grad_dbg<2><<<cdiv(ne, 128), 128, 0, ctx->stream>>>(a, e0, ne, dz, dd, computeGrad, regularize, applyMask, de, dg, di);
    } else {
        ElemArgs<3> a = ctx->elemArgs<3>(0.0);
        k_block_grad_dbg<3><<<cdiv(ne, 128), 128, 0, ctx->stream>>>(a, e0, ne, dz, dd, computeGrad, regularize, applyMask, de, dg, di);
    }
    ctx->launches++;
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaGetLastError());
    CK(cudaMemcpy(energy, de, sizeof(double) * (size_t)ne, cudaMemcpyDeviceToHost));
    if (grad) CK(cudaMemcpy(grad, dg, sizeof(double) * (size_t)ne * n, cudaMemcpyDeviceToHost));
    if (igt) CK(cudaMemcpy(igt, di, sizeof(double) * (size_t)ne, cudaMemcpyDeviceToHost));
    CK(cudaMemsetAsync(ctx->status_d, 0, sizeof(unsigned), ctx->stream));  // probing inverted simplices is legal here
    cudaFree(dz); cudaFree(dd); cudaFree(de); cudaFree(dg); cudaFree(di);
    return MMADMM_OK;
}

int mmadmm_get_stats(mmadmm_ctx *ctx, long long *hist52, long long *proxCalls, long long *admmIters, long long *kernelLaunches) {
    if (!ctx) return MMADMM_EINVAL;
    if (hist52) {
        unsigned long long h[52];
        CK(cudaStreamSynchronize(ctx->stream));
        CK(cudaMemcpy(h, ctx->hist_d, sizeof(h), cudaMemcpyDeviceToHost));
        for (int k = 0; k < 52; k++) hist52[k] = (long long)h[k];
    }
    if (proxCalls) *proxCalls = ctx->proxCalls;
    if (admmIters) *admmIters = ctx->admmIters;
    if (kernelLaunches) *kernelLaunches = ctx->launches;
    return MMADMM_OK;
}

int mmadmm_get_timing(mmadmm_ctx *ctx, double *ms5, long long *counts5, int reset) {
    if (!ctx) return MMADMM_EINVAL;
    CK(cudaSetDevice(ctx->device));
    ctx->phaseResolve();
    for (int i = 0; i < 5; i++) {
        if (ms5) ms5[i] = ctx->phaseMs[i];
        if (counts5) counts5[i] = ctx->phaseCnt[i];
        if (reset) { ctx->phaseMs[i] = 0; ctx->phaseCnt[i] = 0; }
    }
    return MMADMM_OK;
}

int mmadmm_set_option(mmadmm_ctx *ctx, const char *name, int value) {
    if (!ctx || !name) return MMADMM_EINVAL;
    if (!strcmp(name, "spec_used")) return (int)std::min<long long>(ctx->specUsed, 1 << 30);   // query: steps started from the tail
    ctx->specValid = false;
    if (!strcmp(name, "step_tail")) {   // single-GPU step tail (StepTail): bit 0 speculative first solve, bit 1 fused final reduction
        if (value < 0 || value > 3) return MMADMM_EINVAL;
        ctx->tailMode = value;
        return MMADMM_OK;
    }
    if (!strcmp(name, "timing")) {
        ctx->timingOn = value != 0;
        if (ctx->timingOn) {   // event pool up front: creating events is not something to do between two steps
            CK(cudaSetDevice(ctx->device));
            while (ctx->evPool.size() < 1024) { cudaEvent_t e; CK(cudaEventCreate(&e)); ctx->evPool.push_back(e); ctx->evPhase.push_back(0); }
        }
        return MMADMM_OK;
    }
    if (!strcmp(name, "stats")) { ctx->statsOn = value != 0; return MMADMM_OK; }
    if (!strcmp(name, "m2_tri_mode")) {   // Method 2 triangular solves: 2 = dataflow (one persistent launch), 1 = cluster launches, 0 = one launch per level
        if (value < 0 || value > 2) return MMADMM_EINVAL;
        ctx->m2.triMode = value;
        if (ctx->m2.solver) {
            int rc = mmadmm_pcg_set_option(ctx->m2.solver, "tri_mode", ctx->m2.triMode);
            if (rc) { ctx->err = std::string("m2_tri_mode: ") + mmadmm_pcg_last_error(ctx->m2.solver); return rc; }
        }
        return MMADMM_OK;
    }
    if (!strcmp(name, "m2_row_block") || !strcmp(name, "m2_tri_layout")) {
        // rows per thread of the Method 2 triangular solves: 1, D (default, -1), 2 D; level-ELL copy of the factor: 1 (default) / 0
        if (name[3] == 'r') ctx->m2.rowBlock = value; else ctx->m2.triLayout = value ? 1 : 0;
        if (ctx->m2.solver) {
            int rc = mmadmm_pcg_set_option(ctx->m2.solver, "tri_schedule", (ctx->m2.rowBlock < 0 ? ctx->D : ctx->m2.rowBlock) | (ctx->m2.triLayout << 8) | ((ctx->m2.triMode + 1) << 12));
            if (rc) { ctx->err = std::string(name) + ": " + mmadmm_pcg_last_error(ctx->m2.solver); return rc; }
        }
        return MMADMM_OK;
    }
    if (!strcmp(name, "prox_variant")) { ctx->proxVariant = value; return MMADMM_OK; }
    if (!strcmp(name, "xup_variant")) { ctx->xupVariant = value; return MMADMM_OK; }
    if (!strcmp(name, "usum_period")) {
        if (value < 0) return MMADMM_EINVAL;
        ctx->usumPeriod = value; ctx->usumAge = 0;
        return MMADMM_OK;
    }
    if (!strcmp(name, "boundary_projection")) {   // real boundary sliding: see include/mmadmm.h
        if (value < 0 || value > 3) return MMADMM_EINVAL;
        if (value && ctx->dist) { ctx->err = "boundary_projection is available on single-GPU contexts only"; return MMADMM_EINVAL; }
        if (value) {
            CK(cudaSetDevice(ctx->device));
            int rcf = build_face_lists(ctx);
            if (rcf) return rcf;
        }
        ctx->projMask = value;
        return MMADMM_OK;
    }
    if (!strcmp(name, "boundary_faces")) return ctx->nFaces;   // query: number of boundary faces (after boundary_projection was set)
    if (!strcmp(name, "peer_timeout_s")) {   // bound (seconds) of every wait on a peer rank's flag; the wait ends with MMADMM_EPEER
        if (value <= 0) return MMADMM_EINVAL;
        ctx->peerTimeoutNs = (unsigned long long)value * 1000000000ull;
        return MMADMM_OK;
    }
    if (!strcmp(name, "iso_grid")) { ctx->isoOn = value != 0; return MMADMM_OK; }   // 0: always interpolate the full tensors
    if (!strcmp(name, "p2p_active")) return ctx->p2p ? MMADMM_OK : MMADMM_EINVAL;   // query: NVLink peer-memory exchange in use
    if (!strcmp(name, "iso_grid_active")) return (ctx->isoOn && ctx->gridIsIso) ? MMADMM_OK : MMADMM_EINVAL;   // query
    return MMADMM_EINVAL;
}

void *mmadmm_stream(mmadmm_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

void mmadmm_destroy(mmadmm_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    wb_fence(ctx);
    if (ctx->wb.debug && ctx->wb.copies) {
        const mmadmm_ctx::Writeback &w = ctx->wb;
        static const char *nm[7] = {"setup", "wait-stage", "snapshot", "events", "memcpyAsync", "record", "fence(strict)"};
        fprintf(stderr, "mmadmm: write-back: %lld copies; copy duration on the copy stream: avg %.3f ms, max %.3f ms over %lld\n", w.copies,
                w.dbgCopyN ? w.dbgCopyMs / w.dbgCopyN : 0.0, w.dbgCopyMax, w.dbgCopyN);
        for (int m = 0; m < 3; m++)
            if (w.dbgSyncN[m]) fprintf(stderr, "mmadmm:   host wait in the per-iteration stream synchronise, write-back mode %d: avg %.1f us over %lld\n", m,
                                       w.dbgSyncNs[m] * 1e-3 / w.dbgSyncN[m], w.dbgSyncN[m]);
        for (int k = 0; k < 7; k++)
            fprintf(stderr, "mmadmm:   host %-14s avg %8.1f us  max %8.1f us\n", nm[k], w.dbgNs[k] * 1e-3 / w.copies, w.dbgMax[k] * 1e-3);
    }
    for (int b = 0; b < 2; b++) {
        cudaFree(ctx->wb.stage[b]);
        if (ctx->wb.evSnap[b]) cudaEventDestroy(ctx->wb.evSnap[b]);
        if (ctx->wb.evCopied[b]) cudaEventDestroy(ctx->wb.evCopied[b]);
    }
    if (ctx->wb.copyStream) cudaStreamDestroy(ctx->wb.copyStream);
    cudaFree(ctx->ck.x); cudaFree(ctx->ck.xPrev); cudaFree(ctx->ck.xBar); cudaFree(ctx->ck.usum); cudaFree(ctx->ck.z); cudaFree(ctx->ck.u);
    cudaFree(ctx->ck.B); cudaFree(ctx->ck.hist);
    for (int v = 0; v < 4; v++) cudaFree(ctx->F[v]);
    cudaFree(ctx->mask); cudaFree(ctx->fixedBits); cudaFree(ctx->nonIntBits);
    cudaFree(ctx->xSpec); cudaFree(ctx->xBarSpec); cudaFree(ctx->finTicket);
    cudaFree(ctx->x); cudaFree(ctx->xPrev); cudaFree(ctx->xBar); cudaFree(ctx->Vc); cudaFree(ctx->tdiag);
    cudaFree(ctx->z); cudaFree(ctx->u); cudaFree(ctx->usum); cudaFree(ctx->B); cudaFree(ctx->gs);
    cudaFree(ctx->incPtr); cudaFree(ctx->incIdx); cudaFree(ctx->vorder); cudaFree(ctx->valence_d);
    for (size_t r = 0; r < ctx->peerMap.size(); r++)
        if ((int)r != ctx->rank && ctx->peerMap[r]) cudaIpcCloseMemHandle(ctx->peerMap[r]);
    cudaFree(ctx->mailbox); cudaFree(ctx->peerRecv_d); cudaFree(ctx->peerRecvStride_d); cudaFree(ctx->peerHaloFlag_d);
    cudaFree(ctx->peerScal_d); cudaFree(ctx->peerScalFlag_d); cudaFree(ctx->nbrPtr_d);
    cudaFree(ctx->facePtr_d); cudaFree(ctx->faceIdx_d); cudaFree(ctx->faceList_d);
    cudaFree(ctx->pushPtr_d); cudaFree(ctx->pushNbr_d); cudaFree(ctx->pushPos_d); cudaFree(ctx->pushTicket_d);
    cudaFree(ctx->grid); cudaFree(ctx->gridIso); cudaFree(ctx->xT); cudaFree(ctx->ax); cudaFree(ctx->ay); cudaFree(ctx->az);
    cudaFree(ctx->partIh); cudaFree(ctx->partDual); cudaFree(ctx->partPrimal); cudaFree(ctx->scal_d);
    if (ctx->scal_h) cudaFreeHost(ctx->scal_h);
    cudaFree(ctx->status_d);
    if (ctx->status_h) cudaFreeHost(ctx->status_h);
    cudaFree(ctx->hist_d);
    if (ctx->m2.solver) { mmadmm_pcg_destroy(ctx->m2.solver); cudaSetDevice(ctx->device); }
    cudaFree(ctx->m2.adjPtr_d); cudaFree(ctx->m2.adjIdx_d); cudaFree(ctx->m2.val_d); cudaFree(ctx->m2.vp0); cudaFree(ctx->m2.xn);
    cudaFree(ctx->m2.rhs); cudaFree(ctx->m2.dx); cudaFree(ctx->m2.part);
    cudaFree(ctx->sharedIdx); cudaFree(ctx->sharedSlot); cudaFree(ctx->nbrList); cudaFree(ctx->contribPtr); cudaFree(ctx->contrib);
    cudaFree(ctx->verts); cudaFree(ctx->owner); cudaFree(ctx->sPart); cudaFree(ctx->sTot); cudaFree(ctx->sendBuf); cudaFree(ctx->recvBuf);
    cudaFree(ctx->scalAll_d);
    if (ctx->scalAll_h) cudaFreeHost(ctx->scalAll_h);
    if (ctx->xloc_h) cudaFreeHost(ctx->xloc_h);
    if (ctx->comm && nccl().ok) nccl().commDestroy(ctx->comm);
    for (auto e : ctx->evPool) cudaEventDestroy(e);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *mmadmm_last_error(mmadmm_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }

}  // extern "C"
