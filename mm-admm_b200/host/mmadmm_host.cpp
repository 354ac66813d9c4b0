// Host-side helpers of the B200 MM-ADMM package (product code, plain C++17, no CUDA):
// the mesh generators and analytic monitor functions that the reference's driver evaluates once at
// start-up (main.cpp:632-733, src/MeshUtils.h:82-335, Experiments/TestMonitors/MEx*.h).  They are exposed
// with C linkage so that the Python harness (tests, bench.py) can build the same inputs the C++ host
// classes in this directory build.  Nothing here is on the per-step hot path.
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <limits>

#include "MonitorFunctions.h"

extern "C" {

// MonitorFunction<D>::evaluateAtVertices for monitor slot `kind` of main.cpp's Mvals / Mvals3D tables
// (main.cpp:836-863).  X: P x D row-major, out: P x D*D row-major flattened tensors.
int mmadmm_host_monitor_eval(int D, int kind, int P, const double *X, double *out) {
    if (D == 2) {
        MonitorFunction<2> *m = mmadmm_host::makeMonitor2(kind);
        if (!m) return 1;
        m->evaluateAtVertices(X, P, out);
        delete m;
    } else if (D == 3) {
        MonitorFunction<3> *m = mmadmm_host::makeMonitor3(kind);
        if (!m) return 1;
        m->evaluateAtVertices(X, P, out);
        delete m;
    } else {
        return 1;
    }
    return 0;
}

// utils::generateUniformRectMesh<D>: 4 triangles / 12 tetrahedra per cell around a cell-centre vertex.
// Sizes: P = (nx+1)(ny+1)[(nz+1)] + nx*ny[*nz];  E = 4 nx ny | 12 nx ny nz.  bType = NodeType of boundary nodes.
// Quirks kept from the reference: box bounds are truncated to int; the 2D mask uses i/(ny+1) and the 3D mask
// i/(nx+1), i%(ny+1) (equivalent when nx == ny).
int mmadmm_host_uniform_rect_mesh(int D, int nx, int ny, int nz, double xa_, double xb_, double ya_, double yb_,
                                  double za_, double zb_, int bType, double *Vp, int *F, int *mask) {
    if ((D != 2 && D != 3) || nx <= 0 || ny <= 0 || (D == 3 && nz <= 0)) return 1;
    const int INTERIOR = 2, FIXED = 1;
    const int xa = (int)xa_, xb = (int)xb_, ya = (int)ya_, yb = (int)yb_;
    const int za = (D == 3) ? (int)za_ : 0, zb = (D == 3) ? (int)zb_ : 0;
    const double hx = (xb - xa) / ((double)nx), hy = (yb - ya) / ((double)ny);
    const double hz = (D == 3) ? (zb - za) / ((double)nz) : 0;
    size_t o = 0;
    if (D == 2) {
        const int row = nx + 1, nLat = (nx + 1) * (ny + 1);
        for (int j = 0; j <= ny; j++)
            for (int i = 0; i <= nx; i++, o++) { Vp[2 * o] = xa + hx * i; Vp[2 * o + 1] = ya + hy * j; }
        for (int j = 0; j < ny; j++)
            for (int i = 0; i < nx; i++, o++) { Vp[2 * o] = xa + hx * i + hx / 2.0; Vp[2 * o + 1] = ya + hy * j + hy / 2.0; }
        size_t t = 0;
        for (int j = 0; j < ny; j++)
            for (int i = 0; i < nx; i++) {
                const int sw = i + j * row, se = sw + 1, nw = sw + row, ne = nw + 1, c = nLat + i + j * nx;
                const int tri[4][3] = {{sw, c, nw}, {c, ne, nw}, {c, ne, se}, {sw, se, c}};  // left, top, right, bottom
                for (int q = 0; q < 4; q++, t++)
                    for (int r = 0; r < 3; r++) F[3 * t + r] = tri[q][r];
            }
        for (int p = 0; p < nLat + nx * ny; p++) mask[p] = INTERIOR;
        for (int p = 0; p < nLat; p++) {
            const int iOff = p % (nx + 1), jOff = p / (ny + 1);
            const bool onX = (iOff == 0 || iOff == nx), onY = (jOff == 0 || jOff == ny);
            mask[p] = (onX || onY) ? bType : INTERIOR;
            if (onX && onY) mask[p] = FIXED;
        }
        return 0;
    }
    const int row = nx + 1, slab = (nx + 1) * (ny + 1), nLat = slab * (nz + 1);
    for (int k = 0; k <= nz; k++)
        for (int j = 0; j <= ny; j++)
            for (int i = 0; i <= nx; i++, o++) { Vp[3 * o] = xa + hx * i; Vp[3 * o + 1] = ya + hy * j; Vp[3 * o + 2] = za + hz * k; }
    for (int k = 0; k < nz; k++)
        for (int j = 0; j < ny; j++)
            for (int i = 0; i < nx; i++, o++) {
                Vp[3 * o] = xa + hx * i + hx / 2.0; Vp[3 * o + 1] = ya + hy * j + hy / 2.0; Vp[3 * o + 2] = za + hz * k + hz / 2.0;
            }
    size_t t = 0;
    for (int k = 0; k < nz; k++)
        for (int j = 0; j < ny; j++)
            for (int i = 0; i < nx; i++) {
                // the 8 cell corners c[dz][dy][dx]
                int c[2][2][2];
                for (int dz = 0; dz < 2; dz++)
                    for (int dy = 0; dy < 2; dy++)
                        for (int dx = 0; dx < 2; dx++) c[dz][dy][dx] = (i + dx) + (j + dy) * row + (k + dz) * slab;
                const int mid = nLat + i + j * nx + k * (nx * ny);
                // two tets per face, in the reference's order: bottom, top, left, right, back, front
                const int tet[12][3] = {
                    {c[0][0][0], c[0][0][1], c[0][1][1]}, {c[0][0][0], c[0][1][0], c[0][1][1]},
                    {c[1][0][0], c[1][0][1], c[1][1][1]}, {c[1][0][0], c[1][1][0], c[1][1][1]},
                    {c[0][0][0], c[0][1][0], c[1][1][0]}, {c[0][0][0], c[1][0][0], c[1][1][0]},
                    {c[0][0][1], c[0][1][1], c[1][1][1]}, {c[0][0][1], c[1][0][1], c[1][1][1]},
                    {c[0][0][0], c[0][0][1], c[1][0][0]}, {c[0][0][1], c[1][0][1], c[1][0][0]},
                    {c[0][1][0], c[0][1][1], c[1][1][0]}, {c[0][1][1], c[1][1][1], c[1][1][0]}};
                for (int q = 0; q < 12; q++, t++) {
                    for (int r = 0; r < 3; r++) F[4 * t + r] = tet[q][r];
                    F[4 * t + 3] = mid;
                }
            }
    for (int p = 0; p < nLat + nx * ny * nz; p++) mask[p] = INTERIOR;
    for (int k = 0; k <= nz; k++)
        for (int p = 0; p < slab; p++) {
            const int iOff = p / (nx + 1), jOff = p % (ny + 1);
            const bool onI = (iOff == 0 || iOff == nx), onJ = (jOff == 0 || jOff == ny), onK = (k == 0 || k == nz);
            const int id = k * slab + p;
            if (onI || onJ || onK) mask[id] = bType;
            if ((onI && onJ) || (onI && onK) || (onJ && onK)) mask[id] = FIXED;  // edges and corners of the box are fixed
        }
    return 0;
}

// Disc of radius R around (cx, cy) as N concentric hexagonal rings: ring k has 6k points at radius k R / N, so arc and ring
// spacing stay within 5 % of each other and every triangle is close to equilateral.  P = 1 + 3N(N+1), E = 6 N^2; elements
// are emitted ring by ring (spatially coherent, so contiguous element blocks are annuli).  Outer-ring nodes get bType.
// This is the scalable stand-in for the reference's BaseCircle fixtures (Experiments/Results/BaseCircle/CircleEx*.txt,
// unit discs of unknown origin): utils::meshFromLevelSetFun cannot provide it, the reference algorithm inverts an element
// of its meshes in the first step (tests/test_gpu_dropin.py).  Used by BASELINE config 3 (bench.py --workload c3).
int mmadmm_host_disc_mesh(int N, double R, double cx, double cy, int bType, double *Vp, int *F, int *mask) {
    if (N <= 0 || !(R > 0) || !Vp || !F || !mask) return 1;
    const double PI = 3.14159265358979323846;
    auto start = [](long long k) { return 1 + 3 * k * (k - 1); };   // first point of ring k >= 1
    Vp[0] = cx; Vp[1] = cy; mask[0] = 2;
    size_t t = 0;
    for (int k = 1; k <= N; k++) {
        const long long s0 = start(k), s1 = (k > 1) ? start(k - 1) : 0;
        const int nk = 6 * k, nk1 = 6 * (k - 1);
        const double r = R * k / N;
        for (int j = 0; j < nk; j++) {
            const double th = 2.0 * PI * j / nk;
            Vp[2 * (s0 + j)] = cx + r * std::cos(th);
            Vp[2 * (s0 + j) + 1] = cy + r * std::sin(th);
            mask[s0 + j] = (k == N) ? bType : 2;
        }
        for (int sec = 0; sec < 6; sec++)
            for (int i = 0; i < k; i++) {
                const int o0 = (int)(s0 + (sec * k + i) % nk), o1 = (int)(s0 + (sec * k + i + 1) % nk);
                if (k == 1) { F[3 * t] = 0; F[3 * t + 1] = o0; F[3 * t + 2] = o1; t++; continue; }
                const int ia = (int)(s1 + (sec * (k - 1) + i) % nk1);
                F[3 * t] = ia; F[3 * t + 1] = o0; F[3 * t + 2] = o1; t++;
                if (i < k - 1) {
                    const int ib = (int)(s1 + (sec * (k - 1) + i + 1) % nk1);
                    F[3 * t] = ia; F[3 * t + 1] = o1; F[3 * t + 2] = ib; t++;
                }
            }
    }
    return 0;
}

}  // extern "C"
