// =====================================================================================
// mmadmm_oracle.cpp -- CPU ORACLE.  TEST INFRASTRUCTURE ONLY.
//
// A plain C++17 (no Eigen, no nanoflann) restatement of the reference's implicit MMPDE
// step (connortannahill/MM-ADMM @ 895f4e3), function for function, with the reference
// file:line each function follows.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load this library.  The product path
// (mm-admm_b200/) never links, imports or calls it.
//
// Pinning: the reference itself cannot be compiled in this container (Eigen, nanoflann
// and nlohmann/json are un-fetched submodules).  The oracle is pinned to
// the reference's own committed energy traces Experiments/Results/{Monitor210,Monitor310,
// 3DMonitor210,...}/Ih0.txt at their 6 printed digits (tests/test_oracle_kat.py); beyond
// 6 digits parity with the reference is unpinned.
//
// Third-party arithmetic restated here (absent from /root/reference):
//   Eigen @ f0f1d79   fixed-size determinant()/inverse() for 2x2, 3x3 (cofactor forms),
//                     inverse() for 6x6 / 12x12 (PartialPivLU + solve(Identity)),
//                     ConjugateGradient<Lower|Upper, DiagonalPreconditioner> on the
//                     (numerically diagonal) ADMM matrix == exact diagonal solve.
//   nanoflann @ 6fb746e  exact 1-NN (L2_Simple_Adaptor); ties -> lowest vertex index.
// Build: see oracle/Makefile  (g++ -O3 -ffp-contract=off -fopenmp; the reference builds
// with -O3 -msse2, i.e. without FMA contraction).
// =====================================================================================
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>
#include <algorithm>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

enum { BOUNDARY_FREE = 0, BOUNDARY_FIXED = 1, INTERIOR = 2 };  // src/NodeType.h:4-8

// ----------------------------------------------------------------------------------
// Small dense algebra in Eigen's conventions (row-major D x D arrays here; only values matter)
// ----------------------------------------------------------------------------------
template <int D> struct Mat { double a[D][D]; };

inline double det2(const double m[2][2]) { return m[0][0] * m[1][1] - m[1][0] * m[0][1]; }
inline double det3h(const double m[3][3], int a, int b, int c) {
    return m[0][a] * (m[1][b] * m[2][c] - m[1][c] * m[2][b]);
}
inline double det3(const double m[3][3]) {
    return det3h(m, 0, 1, 2) - det3h(m, 1, 0, 2) + det3h(m, 2, 0, 1);
}
template <int D> inline double det(const Mat<D> &m);
template <> inline double det<2>(const Mat<2> &m) { return det2(m.a); }
template <> inline double det<3>(const Mat<3> &m) { return det3(m.a); }

template <int D> inline Mat<D> inverse(const Mat<D> &m);
template <> inline Mat<2> inverse<2>(const Mat<2> &m) {
    Mat<2> r;
    double invdet = 1.0 / det2(m.a);
    r.a[0][0] = m.a[1][1] * invdet;
    r.a[1][0] = -m.a[1][0] * invdet;
    r.a[0][1] = -m.a[0][1] * invdet;
    r.a[1][1] = m.a[0][0] * invdet;
    return r;
}
inline double cof3(const double m[3][3], int i, int j) {
    int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return m[i1][j1] * m[i2][j2] - m[i1][j2] * m[i2][j1];
}
template <> inline Mat<3> inverse<3>(const Mat<3> &m) {
    Mat<3> r;
    double c0 = cof3(m.a, 0, 0), c1 = cof3(m.a, 1, 0), c2 = cof3(m.a, 2, 0);
    double d = c0 * m.a[0][0] + c1 * m.a[1][0] + c2 * m.a[2][0];
    double invdet = 1.0 / d;
    r.a[0][0] = c0 * invdet; r.a[0][1] = c1 * invdet; r.a[0][2] = c2 * invdet;
    r.a[1][0] = cof3(m.a, 0, 1) * invdet; r.a[1][1] = cof3(m.a, 1, 1) * invdet; r.a[1][2] = cof3(m.a, 2, 1) * invdet;
    r.a[2][0] = cof3(m.a, 0, 2) * invdet; r.a[2][1] = cof3(m.a, 1, 2) * invdet; r.a[2][2] = cof3(m.a, 2, 2) * invdet;
    return r;
}
template <int D> inline Mat<D> mul(const Mat<D> &A, const Mat<D> &B) {
    Mat<D> C;
    for (int i = 0; i < D; i++)
        for (int j = 0; j < D; j++) {
            double s = A.a[i][0] * B.a[0][j];
            for (int k = 1; k < D; k++) s += A.a[i][k] * B.a[k][j];
            C.a[i][j] = s;
        }
    return C;
}
template <int D> inline Mat<D> transpose(const Mat<D> &A) {
    Mat<D> C;
    for (int i = 0; i < D; i++) for (int j = 0; j < D; j++) C.a[i][j] = A.a[j][i];
    return C;
}
template <int D> inline double trace(const Mat<D> &A) {
    double s = A.a[0][0];
    for (int i = 1; i < D; i++) s += A.a[i][i];
    return s;
}

// General n x n inverse: partial-pivot LU, then solve against the identity
// (what Eigen's inverse() does for dynamic/large fixed sizes: PartialPivLU::inverse()).
static bool luInverse(int n, const double *Ain, double *Ainv) {
    std::vector<double> lu(Ain, Ain + n * n);
    std::vector<int> perm(n);
    for (int i = 0; i < n; i++) perm[i] = i;
    for (int k = 0; k < n; k++) {
        int piv = k; double best = std::fabs(lu[k * n + k]);
        for (int i = k + 1; i < n; i++) {
            double v = std::fabs(lu[i * n + k]);
            if (v > best) { best = v; piv = i; }
        }
        if (best == 0.0) return false;
        if (piv != k) {
            for (int j = 0; j < n; j++) std::swap(lu[k * n + j], lu[piv * n + j]);
            std::swap(perm[k], perm[piv]);
        }
        for (int i = k + 1; i < n; i++) lu[i * n + k] /= lu[k * n + k];
        for (int i = k + 1; i < n; i++) {
            double l = lu[i * n + k];
            for (int j = k + 1; j < n; j++) lu[i * n + j] -= l * lu[k * n + j];
        }
    }
    for (int c = 0; c < n; c++) {
        std::vector<double> y(n);
        for (int i = 0; i < n; i++) {
            double s = (perm[i] == c) ? 1.0 : 0.0;
            for (int j = 0; j < i; j++) s -= lu[i * n + j] * y[j];
            y[i] = s;
        }
        for (int i = n - 1; i >= 0; i--) {
            double s = y[i];
            for (int j = i + 1; j < n; j++) s -= lu[i * n + j] * y[j];
            y[i] = s / lu[i * n + i];
        }
        for (int i = 0; i < n; i++) Ainv[i * n + c] = y[i];
    }
    return true;
}

// ----------------------------------------------------------------------------------
// utils:: (src/MeshUtils.h)
// ----------------------------------------------------------------------------------
// src/MeshUtils.h:24-29
static void linspace(double xa, double xb, int ns, std::vector<double> &x) {
    x.resize(ns + 1);
    for (int i = 0; i < ns + 1; i++) x[i] = xa + ((double)i) * (xb - xa) / ns;
}
// src/MeshUtils.h:45-54  (uint32_t guess: the "< 0" branch is dead, negatives wrap to the last cell)
static inline int findLimInfMeshPoint(double w, const std::vector<double> &w_mesh) {
    uint32_t guess = (int)((w - w_mesh[0]) / (w_mesh[1] - w_mesh[0]));
    if (guess > w_mesh.size() - 2) guess = (uint32_t)(w_mesh.size() - 2);
    return (int)guess;
}
// src/MeshUtils.h:65-80
static inline void triLinearInterpolation(double x, double y, double z, const double xMesh[2],
                                          const double yMesh[2], const double zMesh[2], double coefs[8]) {
    double xd = (x - xMesh[0]) / (xMesh[1] - xMesh[0]);
    double yd = (y - yMesh[0]) / (yMesh[1] - yMesh[0]);
    double zd = (z - zMesh[0]) / (zMesh[1] - zMesh[0]);
    coefs[0] = (1 - xd) * (1 - yd) * (1 - zd);
    coefs[1] = xd * (1 - yd) * (1 - zd);
    coefs[2] = (1 - xd) * yd * (1 - zd);
    coefs[3] = xd * yd * (1 - zd);
    coefs[4] = (1 - xd) * (1 - yd) * zd;
    coefs[5] = xd * (1 - yd) * zd;
    coefs[6] = (1 - xd) * yd * zd;
    coefs[7] = xd * yd * zd;
}

// ----------------------------------------------------------------------------------
// The oracle state: Mesh<D> + MeshInterpolator<D> + MeshIntegrator<D> rolled into one
// ----------------------------------------------------------------------------------
struct Oracle {
    int D = 0, n = 0, P = 0, E = 0;
    bool compMesh = false, gradUse = false;
    std::vector<double> Vp, Vc;  // P x D row-major
    std::vector<int> F;          // E x (D+1) row-major (re-oriented)
    std::vector<int> mask;       // P
    // MeshInterpolator
    int nx = 0, ny = 0, nz = 0;
    std::vector<double> gx, gy, gz;
    std::vector<double> monVals;      // P x D^2
    std::vector<double> monGridVals;  // (nx+1)(ny+1)(nz+1) x D^2
    // Mesh
    double rho = 0, w = 0, tau = 0, m = 0;
    bool stepTaken = false, hessComputed = false;
    std::vector<double> hessInvs;  // E x n x n, row-major inside an element
    std::vector<double> Ih;        // E
    std::vector<double> grad;      // P*D (eulerGrad)
    // MeshIntegrator
    double dt = 0;
    int stepsTaken = 0;
    std::vector<double> x, xPrev, xBar, z, zPrev, uBar, DXpU, vec, tdiag;
    // vertex -> (element, local vertex) incidence in ascending element order (the order in
    // which Eigen's column-major SpMV WD_T * v accumulates into a row)
    std::vector<int> incPtr, incIdx;
    // statistics
    long long bfgsHist[52];   // histogram of BFGS iterations per element-prox (index 51 = hit max)
    long long proxCalls = 0;  // element-prox calls
    int lastAdmmIters = 0;
    double lastPrimal = 0, lastDual = 0;
    int nthreads = 1;
    int error = 0;  // 1 = inverted element seen (reference would assert), 2 = singular FD Hessian
    // Boundary sliding (src/Mesh.cpp:61-241).  The reference builds the face lists at init (buildFaceList, :420) but all three call
    // sites of projectOntoBoundary are commented out (:636-642 old eulerStep, :975-984 prox, :1020-1026 updateAfterStep); projMask
    // re-enables them as an OPTION: bit 0 = after every element's BFGS solve in prox, bit 1 = in updateAfterStep.  Default 0 = what
    // the reference executes.  No reference result exists with it on: new behaviour, parity GPU vs this oracle only.
    int projMask = 0;
    std::vector<int> faceList;             // nFaces x D vertex ids
    std::vector<int> facePtr, faceIdx;     // faceConnects: vertex -> faces, ascending face id (std::set<int> order)
};

// src/MeshInterpolator.cpp:286-342
template <int D>
static inline void evalMonitorOnGrid(const Oracle &o, const double *pnt, Mat<D> &mVal) {
    int xInd = findLimInfMeshPoint(pnt[0], o.gx);
    int yInd = findLimInfMeshPoint(pnt[1], o.gy);
    const int nx = o.nx, ny = o.ny;
    const double *g = o.monGridVals.data();
    constexpr int DD = D * D;
    if (D == 2) {
        double xMesh[2] = {o.gx[xInd], o.gx[xInd + 1]};
        double yMesh[2] = {o.gy[yInd], o.gy[yInd + 1]};
        double x = pnt[0], y = pnt[1];
        double norm = (1 / ((xMesh[1] - xMesh[0]) * (yMesh[1] - yMesh[0])));
        double coefs[4] = {norm * (xMesh[1] - x) * (yMesh[1] - y), norm * (x - xMesh[0]) * (yMesh[1] - y),
                           norm * (xMesh[1] - x) * (y - yMesh[0]), norm * (x - xMesh[0]) * (y - yMesh[0])};
        for (int k = 0; k < DD; k++) {
            mVal.a[k / D][k % D] = coefs[0] * g[(size_t)(yInd * (nx + 1) + xInd) * DD + k] +
                                   coefs[1] * g[(size_t)(yInd * (nx + 1) + xInd + 1) * DD + k] +
                                   coefs[2] * g[(size_t)((yInd + 1) * (nx + 1) + xInd) * DD + k] +
                                   coefs[3] * g[(size_t)((yInd + 1) * (nx + 1) + xInd + 1) * DD + k];
        }
    } else {
        int zInd = findLimInfMeshPoint(pnt[2], o.gz);
        double coefs[8] = {0.0};
        double xMesh[2] = {o.gx[xInd], o.gx[xInd + 1]};
        double yMesh[2] = {o.gy[yInd], o.gy[yInd + 1]};
        double zMesh[2] = {o.gz[zInd], o.gz[zInd + 1]};
        triLinearInterpolation(pnt[0], pnt[1], pnt[2], xMesh, yMesh, zMesh, coefs);
        double mFlat[DD];
        for (int k = 0; k < DD; k++) mFlat[k] = 0.0;
        size_t sl = (size_t)(nx + 1) * (ny + 1);
        size_t rows[8] = {zInd * sl + (size_t)yInd * (nx + 1) + xInd,
                          zInd * sl + (size_t)yInd * (nx + 1) + xInd + 1,
                          zInd * sl + (size_t)(yInd + 1) * (nx + 1) + xInd,
                          zInd * sl + (size_t)(yInd + 1) * (nx + 1) + xInd + 1,
                          (zInd + 1) * sl + (size_t)yInd * (nx + 1) + xInd,
                          (zInd + 1) * sl + (size_t)yInd * (nx + 1) + xInd + 1,
                          (zInd + 1) * sl + (size_t)(yInd + 1) * (nx + 1) + xInd,
                          (zInd + 1) * sl + (size_t)(yInd + 1) * (nx + 1) + xInd + 1};
        for (int c = 0; c < 8; c++)
            for (int k = 0; k < DD; k++) mFlat[k] += coefs[c] * g[rows[c] * DD + k];
        for (int k = 0; k < DD; k++) mVal.a[k / D][k % D] = mFlat[k];
    }
}

// src/AdaptationFunctional.cpp:102-287
template <int D>
static double blockGrad(Oracle &o, int zId, const double *z, const double *xi, double *grad,
                        bool computeGrad, bool regularize, double &Igt) {
    constexpr int n = D * (D + 1);
    double Ih = 0.0, detFJ;
    Mat<D> E, FJ, Ehat, M, vLoc, Einv, dGdJ, dGdM;
    double basisComb[D], gradSimplex[D], dGdX[D];
    const double dFact = (D == 2) ? 2.0 : 6.0;
    double absK;

    // Interpolate the monitor function (:143-153)
    Mat<D> mPre[D + 1];
    for (int i = 0; i < D; i++) for (int j = 0; j < D; j++) M.a[i][j] = 0.0;
    for (int i = 0; i < D + 1; i++) {
        Mat<D> mTemp;
        evalMonitorOnGrid<D>(o, z + i * D, mTemp);
        mPre[i] = mTemp;
        for (int a = 0; a < D; a++) for (int b = 0; b < D; b++) M.a[a][b] += mTemp.a[a][b];
    }
    // Minv = M^{-1} / (D+1)  (:157, sic)
    Mat<D> Minv = inverse<D>(M);
    for (int a = 0; a < D; a++) for (int b = 0; b < D; b++) Minv.a[a][b] /= ((double)D + 1);

    double G, dGddet;
    // Edge matrices (:162-169)
    for (int j = 0; j < D; j++) {
        int nn = j + 1;
        for (int l = 0; l < D; l++) {
            E.a[l][j] = z[D * nn + l] - z[l];
            if (o.compMesh) Ehat.a[l][j] = xi[D * nn + l] - xi[l];
        }
    }
    double Edet = det<D>(E);
    if (!(Edet > 0)) o.error = 1;  // reference: assert(Edet > 0) (:174)

    if (!o.compMesh) {  // (:176-201)
        if (D == 2) {
            Ehat.a[0][0] = 1.0; Ehat.a[1][0] = 0.0;
            Ehat.a[0][1] = 1.0 / 2.0; Ehat.a[1][1] = std::sqrt(3) / 2.0;
        } else {
            Ehat.a[0][0] = -2.0; Ehat.a[1][0] = 0.0;  Ehat.a[2][0] = -2.0;
            Ehat.a[0][1] = 0.0;  Ehat.a[1][1] = -2.0; Ehat.a[2][1] = -2.0;
            Ehat.a[0][2] = -2.0; Ehat.a[1][2] = -2.0; Ehat.a[2][2] = 0.0;
        }
        double s1 = std::pow((dFact / std::fabs(det<D>(Ehat))), 1.0 / ((double)D));
        for (int a = 0; a < D; a++) for (int b = 0; b < D; b++) Ehat.a[a][b] *= s1;
        double s2 = (double)std::pow(o.E, 1.0 / D);
        for (int a = 0; a < D; a++) for (int b = 0; b < D; b++) Ehat.a[a][b] /= s2;
    }

    Einv = inverse<D>(E);
    FJ = mul<D>(Ehat, Einv);
    detFJ = det<D>(FJ);

    double d = (double)D;
    const double p = 1.5;
    const double theta = 1.0 / 3.0;

    Mat<D> FJt = transpose<D>(FJ);
    Mat<D> MinvJt = mul<D>(Minv, FJt);
    Mat<D> JMJt = mul<D>(FJ, MinvJt);
    double trJMJt = trace<D>(JMJt);
    double detM = std::sqrt(1.0 / det<D>(Minv));

    G = theta * detM * std::pow(trJMJt, d * p / 2.0) +
        (1.0 - 2.0 * theta) * std::pow(d, d * p / 2.0) * detM * std::pow(detFJ / detM, p);

    absK = std::fabs(Edet / dFact);

    if (!computeGrad) {  // (:224-230)
        if (regularize) {
            double sq = 0.0;
            for (int k = 0; k < n; k++) { double t = o.DXpU[(size_t)n * zId + k] - z[k]; sq += t * t; }
            return absK * G + 0.5 * o.w * o.w * sq;
        }
        return absK * G;
    }

    // (:232-236)
    {
        double s = d * p * theta * detM * std::pow(trJMJt, d * p / 2.0 - 1);
        for (int a = 0; a < D; a++) for (int b = 0; b < D; b++) dGdJ.a[a][b] = s * MinvJt.a[a][b];
    }
    dGddet = p * (1.0 - 2.0 * theta) * std::pow(d, (d * p) / 2.0) * std::pow(detM, 1.0 - p) * std::pow(detFJ, p - 1);
    {
        double s1 = -0.5 * theta * d * p * detM * std::pow(trJMJt, d * p / 2.0 - 1);
        Mat<D> Mt = transpose<D>(Minv);
        Mat<D> T = mul<D>(mul<D>(mul<D>(Mt, FJt), FJ), Minv);
        double s2 = (0.5 * theta * detM * std::pow(trJMJt, d * p / 2.0) +
                     ((0.5 - theta) * (1.0 - p) * std::pow(d, d * p / 2.0)) * std::pow(detM, 1 - p) * std::pow(detFJ, p));
        for (int a = 0; a < D; a++)
            for (int b = 0; b < D; b++) dGdM.a[a][b] = s1 * T.a[a][b] + s2 * Minv.a[a][b];
    }
    for (int l = 0; l < D; l++) { dGdX[l] = 0.0; basisComb[l] = 0.0; }

    // (:239-244)
    for (int j = 0; j < D; j++) {
        int nn = j + 1;
        Mat<D> dm;
        for (int a = 0; a < D; a++) for (int b = 0; b < D; b++) dm.a[a][b] = mPre[nn].a[a][b] - mPre[0].a[a][b];
        double tr = trace<D>(mul<D>(dGdM, dm));
        for (int l = 0; l < D; l++) basisComb[l] += Einv.a[j][l] * tr;
    }

    // (:246-250)
    double c1 = (-G + dGddet * detFJ);
    {
        Mat<D> T = mul<D>(mul<D>(Einv, dGdJ), FJ);
        for (int a = 0; a < D; a++) for (int b = 0; b < D; b++) vLoc.a[a][b] = c1 * Einv.a[a][b] + T.a[a][b];
    }
    for (int r = 0; r < D; r++)
        for (int l = 0; l < D; l++) vLoc.a[r][l] -= (basisComb[l]) / ((double)D + 1.0);

    // (:253-271)
    for (int l = 0; l < D; l++) gradSimplex[l] = 0.0;
    for (int r = 0; r < D; r++) for (int l = 0; l < D; l++) gradSimplex[l] += vLoc.a[r][l];
    for (int l = 0; l < D; l++) gradSimplex[l] += basisComb[l] + dGdX[l];
    for (int l = 0; l < D; l++) grad[l] = gradSimplex[l];
    for (int nn = 1; nn < D + 1; nn++) for (int l = 0; l < D; l++) grad[D * nn + l] = -vLoc.a[nn - 1][l];
    for (int k = 0; k < n; k++) grad[k] *= absK;

    Ih = absK * G;
    Igt = Ih;

    if (regularize) {  // (:279-282)
        double sq = 0.0;
        for (int k = 0; k < n; k++) { double t = o.DXpU[(size_t)n * zId + k] - z[k]; sq += t * t; }
        Ih += 0.5 * o.w * o.w * sq;
        for (int k = 0; k < n; k++) grad[k] += o.w * o.w * (-o.DXpU[(size_t)n * zId + k] + z[k]);
    }
    return Ih;
}

// src/Mesh.cpp:755-772
template <int D>
static double computeBlockGrad(Oracle &o, int zId, const double *z, const double *xi, double *grad,
                               bool computeGrad, bool regularize, double &Igt) {
    double Ix = blockGrad<D>(o, zId, z, xi, grad, computeGrad, regularize, Igt);
    for (int i = 0; i < D + 1; i++) {
        int pntId = o.F[(size_t)zId * (D + 1) + i];
        if (o.mask[pntId] == BOUNDARY_FIXED)
            for (int l = 0; l < D; l++) grad[D * i + l] = 0.0;
    }
    return Ix;
}

// src/Mesh.cpp:61-112 buildFaceList: a simplex with exactly D non-INTERIOR vertices contributes one boundary face (its non-interior
// vertices in local order); faceConnects = per vertex the set of faces it belongs to.
template <int D>
static void buildFaceList(Oracle &o) {
    o.faceList.clear();
    for (int i = 0; i < o.E; i++) {
        int sum = 0;
        for (int j = 0; j < D + 1; j++) sum += (o.mask[o.F[(size_t)i * (D + 1) + j]] != INTERIOR) ? 1 : 0;
        if (sum == D)
            for (int j = 0; j < D + 1; j++)
                if (o.mask[o.F[(size_t)i * (D + 1) + j]] != INTERIOR) o.faceList.push_back(o.F[(size_t)i * (D + 1) + j]);
    }
    const int nF = (int)(o.faceList.size() / D);
    o.facePtr.assign(o.P + 1, 0);
    for (int f = 0; f < nF; f++)
        for (int j = 0; j < D; j++) {
            bool dup = false;   // a set: a face lists a vertex once
            for (int jj = 0; jj < j; jj++) dup = dup || o.faceList[(size_t)f * D + jj] == o.faceList[(size_t)f * D + j];
            if (!dup) o.facePtr[o.faceList[(size_t)f * D + j] + 1]++;
        }
    for (int p = 0; p < o.P; p++) o.facePtr[p + 1] += o.facePtr[p];
    o.faceIdx.resize(o.facePtr[o.P]);
    std::vector<int> fill(o.facePtr.begin(), o.facePtr.end() - 1);
    for (int f = 0; f < nF; f++)
        for (int j = 0; j < D; j++) {
            bool dup = false;
            for (int jj = 0; jj < j; jj++) dup = dup || o.faceList[(size_t)f * D + jj] == o.faceList[(size_t)f * D + j];
            if (!dup) o.faceIdx[fill[o.faceList[(size_t)f * D + j]]++] = f;
        }
}
static inline int sgn(double val) { return (0.0 < val) - (val < 0.0); }   // src/Mesh.cpp:114-116
// src/Mesh.cpp:118-177 projection2D: closest point on the boundary edges that meet at the node, edge geometry from Vp
static void projection2D(const Oracle &o, int nodeId, double *nodeIn) {
    const double node[2] = {nodeIn[0], nodeIn[1]};
    double best[2] = {node[0], node[1]};
    double minDist = INFINITY;
    for (int q = o.facePtr[nodeId]; q < o.facePtr[nodeId + 1]; q++) {
        const int *pid = &o.faceList[(size_t)o.faceIdx[q] * 2];
        const double x1[2] = {o.Vp[(size_t)pid[0] * 2], o.Vp[(size_t)pid[0] * 2 + 1]};
        const double x2[2] = {o.Vp[(size_t)pid[1] * 2], o.Vp[(size_t)pid[1] * 2 + 1]};
        const double u[2] = {x2[0] - x1[0], x2[1] - x1[1]};
        const double w[2] = {node[0] - x1[0], node[1] - x1[1]};
        const double c = (u[0] * w[0] + u[1] * w[1]) / (u[0] * u[0] + u[1] * u[1]);
        const double proj[2] = {c * u[0], c * u[1]};
        double dist = std::sqrt((proj[0] - w[0]) * (proj[0] - w[0]) + (proj[1] - w[1]) * (proj[1] - w[1]));
        const double t = std::sqrt(proj[0] * proj[0] + proj[1] * proj[1]) / std::sqrt(u[0] * u[0] + u[1] * u[1]);
        const bool same = sgn(u[0]) == sgn(proj[0]) && sgn(u[1]) == sgn(proj[1]);
        if (dist < minDist && same && t > 0.0 && t < 1.0) {
            minDist = dist;
            best[0] = (1.0 - t) * x1[0] + t * x2[0];
            best[1] = (1.0 - t) * x1[1] + t * x2[1];
        } else if (!same) {
            dist = std::sqrt((x1[0] - node[0]) * (x1[0] - node[0]) + (x1[1] - node[1]) * (x1[1] - node[1]));
            if (dist < minDist) { minDist = dist; best[0] = x1[0]; best[1] = x1[1]; }
        } else if (t > 1.0) {
            dist = std::sqrt((x2[0] - node[0]) * (x2[0] - node[0]) + (x2[1] - node[1]) * (x2[1] - node[1]));
            if (dist < minDist) { minDist = dist; best[0] = x2[0]; best[1] = x2[1]; }
        }
    }
    nodeIn[0] = best[0]; nodeIn[1] = best[1];
}
// src/Mesh.cpp:179-231 projection3D: barycentric projection onto the boundary triangles that meet at the node; the node stays
// where it is when no projection lands inside a triangle (all b >= 1e-10)
static void projection3D(const Oracle &o, int nodeId, double *nodeIn) {
    const double node[3] = {nodeIn[0], nodeIn[1], nodeIn[2]};
    double best[3] = {0, 0, 0};
    bool projWorked = false;
    const double CHECK_EPS = 1e-10;
    double minDist = INFINITY;
    auto cross = [](const double *a, const double *b, double *c) {
        c[0] = a[1] * b[2] - a[2] * b[1]; c[1] = a[2] * b[0] - a[0] * b[2]; c[2] = a[0] * b[1] - a[1] * b[0];
    };
    auto dot = [](const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; };
    for (int qi = o.facePtr[nodeId]; qi < o.facePtr[nodeId + 1]; qi++) {
        const int *pid = &o.faceList[(size_t)o.faceIdx[qi] * 3];
        const double *p0 = &o.Vp[(size_t)pid[0] * 3], *p1 = &o.Vp[(size_t)pid[1] * 3], *p2 = &o.Vp[(size_t)pid[2] * 3];
        double u[3], v[3], w[3], n[3], uw[3], wv[3], b[3];
        for (int l = 0; l < 3; l++) { u[l] = p1[l] - p0[l]; v[l] = p2[l] - p0[l]; w[l] = node[l] - p0[l]; }
        cross(u, v, n);
        const double temp = 1.0 / dot(n, n);
        cross(u, w, uw);
        cross(w, v, wv);
        b[2] = dot(uw, n) * temp;
        b[1] = dot(wv, n) * temp;
        b[0] = 1.0 - b[1] - b[2];
        double np[3];
        for (int l = 0; l < 3; l++) np[l] = b[0] * p0[l] + b[1] * p1[l] + b[2] * p2[l];
        const double dist = std::sqrt((np[0] - node[0]) * (np[0] - node[0]) + (np[1] - node[1]) * (np[1] - node[1]) +
                                      (np[2] - node[2]) * (np[2] - node[2]));
        if (dist < minDist && b[0] >= CHECK_EPS && b[1] >= CHECK_EPS && b[2] >= CHECK_EPS) {
            minDist = dist;
            best[0] = np[0]; best[1] = np[1]; best[2] = np[2];
            projWorked = true;
        }
    }
    if (projWorked) { nodeIn[0] = best[0]; nodeIn[1] = best[1]; nodeIn[2] = best[2]; }
}
template <int D> static inline void projectOntoBoundary(const Oracle &o, int nodeId, double *node) {   // src/Mesh.cpp:233-241
    if (D == 2) projection2D(o, nodeId, node); else projection3D(o, nodeId, node);
}

// src/Mesh.cpp:777-872
template <int D>
static double bfgsOptSimplex(Oracle &o, int zId, double *z, const double *xi, int nIter, double tol, int &itersOut) {
    constexpr int n = D * (D + 1);
    double h = 2.0 * std::sqrt(std::numeric_limits<double>::epsilon());
    double zPurt[n], Gkp1[n], Gk[n];
    double Igt;
    computeBlockGrad<D>(o, zId, z, xi, Gk, true, true, Igt);
    double Ihsave = Igt;
    double *Hst = &o.hessInvs[(size_t)zId * n * n];
    if (!o.hessComputed) {
        double H[n * n];
        for (int k = 0; k < n; k++) zPurt[k] = z[k];
        for (int i = 0; i < n; i++) {
            zPurt[i] += h;
            computeBlockGrad<D>(o, zId, zPurt, xi, Gkp1, true, true, Igt);
            for (int r = 0; r < n; r++) H[r * n + i] = (Gkp1[r] - Gk[r]) / h;
            zPurt[i] = z[i];
        }
        for (int nn = 0; nn < D + 1; nn++) {
            int pntId = o.F[(size_t)zId * (D + 1) + nn];
            if (o.mask[pntId] != INTERIOR)
                for (int mm = 0; mm < D; mm++) H[(D * nn + mm) * n + (D * nn + mm)] = 1.0;
        }
        if (!luInverse(n, H, Hst)) o.error = 2;
    }

    double Bkinv[n * n];
    std::memcpy(Bkinv, Hst, sizeof(Bkinv));
    double pk[n], yk[n];
    double c1, c2;
    int iter;
    for (iter = 0; iter < nIter; iter++) {
        // pk = -Bkinv * Gk
        for (int r = 0; r < n; r++) {
            double s = Bkinv[r * n] * Gk[0];
            for (int c = 1; c < n; c++) s += Bkinv[r * n + c] * Gk[c];
            pk[r] = -s;
        }
        for (int k = 0; k < n; k++) z[k] += pk[k];
        computeBlockGrad<D>(o, zId, z, xi, Gkp1, true, true, Igt);
        double Ix = 0;
        for (int i = 0; i < n; i++) Ix += std::fabs(Gkp1[i]);
        for (int k = 0; k < n; k++) yk[k] = Gkp1[k] - Gk[k];

        // c2 = pk.yk ; c1 = (pk.yk + yk.(Bkinv yk)) / c2^2
        double pkyk = pk[0] * yk[0];
        for (int k = 1; k < n; k++) pkyk += pk[k] * yk[k];
        c2 = pkyk;
        double By[n], yB[n];
        for (int r = 0; r < n; r++) {
            double s = Bkinv[r * n] * yk[0];
            for (int c = 1; c < n; c++) s += Bkinv[r * n + c] * yk[c];
            By[r] = s;
        }
        for (int c = 0; c < n; c++) {
            double s = yk[0] * Bkinv[c];
            for (int r = 1; r < n; r++) s += yk[r] * Bkinv[r * n + c];
            yB[c] = s;
        }
        double yBy = yk[0] * By[0];
        for (int k = 1; k < n; k++) yBy += yk[k] * By[k];
        c1 = (pkyk + yBy) / (std::pow(c2, 2.0));
        // Bkinv += c1 pk pk^T - Bkinv (yk pk^T)/c2 - pk (yk^T Bkinv)/c2  (whole RHS from the old Bkinv)
        for (int r = 0; r < n; r++)
            for (int c = 0; c < n; c++)
                Bkinv[r * n + c] += (c1 * (pk[r] * pk[c]) - (By[r] * pk[c]) / c2) - (pk[r] * yB[c]) / c2;

        for (int k = 0; k < n; k++) Gk[k] = Gkp1[k];
        if (Ix < tol) break;
    }
    itersOut = (iter == nIter) ? nIter + 1 : iter + 1;
    std::memcpy(Hst, Bkinv, sizeof(Bkinv));
    return Ihsave;
}

// src/Mesh.cpp:930-994
template <int D>
static double prox(Oracle &o, double tol) {
    constexpr int n = D * (D + 1);
    const int E = o.E;
    std::vector<long long> hist((size_t)o.nthreads * 52, 0);
#pragma omp parallel for num_threads(o.nthreads) schedule(static)
    for (int i = 0; i < E; i++) {
        double z_i[n], xi_i[n];
        if (o.compMesh)
            for (int nn = 0; nn < D + 1; nn++)
                for (int l = 0; l < D; l++) xi_i[nn * D + l] = o.Vc[(size_t)o.F[(size_t)i * (D + 1) + nn] * D + l];
        for (int k = 0; k < n; k++) z_i[k] = o.z[(size_t)n * i + k];
        int its = 0;
        o.Ih[i] = bfgsOptSimplex<D>(o, i, z_i, xi_i, 50, tol / 100, its);
        if (o.projMask & 1)   // "Moving boundary points" (src/Mesh.cpp:975-984, commented out in the reference)
            for (int nn = 0; nn < D + 1; nn++) {
                const int v = o.F[(size_t)i * (D + 1) + nn];
                if (o.mask[v] == BOUNDARY_FREE) projectOntoBoundary<D>(o, v, z_i + nn * D);
            }
        for (int l = 0; l < n; l++) o.z[(size_t)n * i + l] = z_i[l];
        int tid = 0;
#ifdef _OPENMP
        tid = omp_get_thread_num();
#endif
        hist[(size_t)tid * 52 + its]++;
    }
    for (int t = 0; t < o.nthreads; t++) for (int k = 0; k < 52; k++) o.bfgsHist[k] += hist[(size_t)t * 52 + k];
    o.proxCalls += E;
    o.hessComputed = true;
    double s = 0.0;  // Ih->sum()
    for (int i = 0; i < E; i++) s += o.Ih[i];
    return s;
}

// src/Mesh.cpp:582-624 (serial in the reference; element order fixes the accumulation order)
template <int D>
static double eulerGrad(Oracle &o, const std::vector<double> &x, std::vector<double> &grad) {
    constexpr int n = D * (D + 1);
    std::fill(grad.begin(), grad.end(), 0.0);
    double Ihorig = 0.0, Igt;
    for (int i = 0; i < o.E; i++) {
        double x_i[n], xi_i[n], gradSimp[n];
        for (int nn = 0; nn < D + 1; nn++) {
            int v = o.F[(size_t)i * (D + 1) + nn];
            for (int l = 0; l < D; l++) {
                if (o.compMesh) xi_i[nn * D + l] = o.Vc[(size_t)v * D + l];
                x_i[nn * D + l] = x[(size_t)D * v + l];
            }
        }
        Ihorig += computeBlockGrad<D>(o, i, x_i, xi_i, gradSimp, true, false, Igt);
        for (int nn = 0; nn < D + 1; nn++) {
            int off = o.F[(size_t)i * (D + 1) + nn];
            for (int l = 0; l < D; l++) grad[(size_t)D * off + l] += gradSimp[D * nn + l];
        }
    }
    return Ihorig;
}

// Method 1 (explicit Euler).  src/Mesh.cpp:533-579 eulerStepMod: blockGrad directly (no FIXED zeroing), but only INTERIOR
// vertices collect gradient, so FIXED and FREE boundary vertices both stay put; src/MeshIntegrator.cpp:87-95 eulerStep:
// x -= (dt/tau) grad, returns the energy at entry.  Vp is NOT refreshed (no updateAfterStep on this path).
template <int D>
static double eulerStepMod(Oracle &o, const std::vector<double> &x, std::vector<double> &grad) {
    constexpr int n = D * (D + 1);
    std::fill(grad.begin(), grad.end(), 0.0);
    double Ihorig = 0.0, Igt;
    for (int i = 0; i < o.E; i++) {
        double x_i[n], xi_i[n], gradSimp[n];
        for (int nn = 0; nn < D + 1; nn++) {
            int v = o.F[(size_t)i * (D + 1) + nn];
            for (int l = 0; l < D; l++) {
                if (o.compMesh) xi_i[nn * D + l] = o.Vc[(size_t)v * D + l];
                x_i[nn * D + l] = x[(size_t)D * v + l];
            }
        }
        Ihorig += blockGrad<D>(o, i, x_i, xi_i, gradSimp, true, false, Igt);
        for (int nn = 0; nn < D + 1; nn++) {
            int off = o.F[(size_t)i * (D + 1) + nn];
            if (o.mask[off] != INTERIOR) continue;
            for (int l = 0; l < D; l++) grad[(size_t)D * off + l] += gradSimp[D * nn + l];
        }
    }
    return Ihorig;
}
template <int D>
static double eulerStep(Oracle &o) {
    double Ih = eulerStepMod<D>(o, o.x, o.grad);
    for (size_t i = 0; i < o.x.size(); i++) o.x[i] -= (o.dt / o.tau) * o.grad[i];
    return Ih;
}

// Method 2 (backward Euler), Jacobian of the residual.  src/Mesh.cpp:1111-1135 buildEulerJac + :1161-1261 FSubJac on
// the CSR pattern (ia, ja) that buildMatrix (src/Mesh.cpp:305-337) gives LASolver: rows D*v+p couple to all D columns of
// every vertex sharing a simplex with v.  Quirks kept: the finite differences (h = 10 sqrt(eps), forward) are taken at
// *Vp, which this method never refreshes (no updateAfterStep on the path), not at the iterate x; the perturbed direction
// i fills ROW i of `derivs` (d grad / d x_{off,i}, i.e. the Hessian column used as a row); for a BOUNDARY_FIXED vertex
// derivs is the identity only when the vertex is local vertex 0 of the simplex (r == c with c in [D off, D off + D));
// values are summed in ascending simplex order, then scaled by dt/tau, then 1 is added on the diagonal.
template <int D>
static void buildEulerJac(Oracle &o, const int *ia, const int *ja, double *a) {
    constexpr int n = D * (D + 1);
    const double h = 10.0 * std::sqrt(std::numeric_limits<double>::epsilon());
    const int N = o.P * D;
    for (int k = 0; k < ia[N]; k++) a[k] = 0.0;
    for (int pntId = 0; pntId < o.P; pntId++) {
        for (int q = o.incPtr[pntId]; q < o.incPtr[pntId + 1]; q++) {
            const int sId = o.incIdx[q] / (D + 1);  // ascending element order == std::set iteration order
            double xiLoc[n], xLoc[n], Gk[n], Gkp1[n], xPurt[n], derivs[D][n], Igt;
            int off = 0;
            for (int nn = 0; nn < D + 1; nn++) {
                int v = o.F[(size_t)sId * (D + 1) + nn];
                if (v == pntId) off = nn;
                for (int l = 0; l < D; l++) {
                    if (o.compMesh) xiLoc[nn * D + l] = o.Vc[(size_t)v * D + l];
                    xLoc[nn * D + l] = o.Vp[(size_t)v * D + l];
                }
            }
            blockGrad<D>(o, sId, xLoc, xiLoc, Gk, true, false, Igt);
            for (int c = 0; c < n; c++) xPurt[c] = xLoc[c];
            if (o.mask[pntId] == BOUNDARY_FIXED) {
                for (int r = 0; r < D; r++)
                    for (int c = 0; c < n; c++) derivs[r][c] = 0.0;
                for (int r = 0; r < D; r++)
                    for (int c = D * off; c < D * (off + 1); c++)
                        if (r == c) derivs[r][c] = 1.0;
            } else {
                for (int i = 0; i < D; i++) {
                    xPurt[D * off + i] += h;
                    blockGrad<D>(o, sId, xPurt, xiLoc, Gkp1, true, false, Igt);
                    for (int c = 0; c < n; c++) derivs[i][c] = (Gkp1[c] - Gk[c]) / h;
                    xPurt[D * off + i] = xLoc[D * off + i];
                }
            }
            for (int p = 0; p < D; p++) {
                const int row = D * pntId + p;
                for (int k = ia[row]; k < ia[row + 1]; k++) {
                    const int colIndex = ja[k] / D, colOff = ja[k] % D;
                    for (int j = 0; j < D + 1; j++)
                        if (o.F[(size_t)sId * (D + 1) + j] == colIndex) a[k] += derivs[p][D * j + colOff];
                }
            }
        }
    }
    for (int r = 0; r < N; r++)
        for (int k = ia[r]; k < ia[r + 1]; k++) {
            a[k] *= (o.dt / o.tau);
            if (ja[k] == r) a[k] += 1.0;
        }
}

// src/Mesh.cpp:496-530 (reads *Vp, ignores its argument)
template <int D>
static double computeEnergy(Oracle &o) {
    constexpr int n = D * (D + 1);
    double Ih = 0.0, Igt;
    for (int i = 0; i < o.E; i++) {
        double x_i[n], xi_i[n], g[n];
        for (int nn = 0; nn < D + 1; nn++) {
            int v = o.F[(size_t)i * (D + 1) + nn];
            for (int l = 0; l < D; l++) {
                if (o.compMesh) xi_i[nn * D + l] = o.Vc[(size_t)v * D + l];
                x_i[nn * D + l] = o.Vp[(size_t)v * D + l];
            }
        }
        Ih += computeBlockGrad<D>(o, i, x_i, xi_i, g, false, false, Igt);
    }
    return Ih;
}

// src/Mesh.cpp:649-674
template <int D>
static void predictX(Oracle &o) {
    const size_t N = o.x.size();
    if (o.gradUse || o.stepsTaken <= 2) {
        eulerGrad<D>(o, o.x, o.grad);
        for (size_t i = 0; i < N; i++) o.xBar[i] = o.x[i] - (o.dt / o.tau) * o.grad[i];
    } else {
        for (size_t i = 0; i < N; i++) o.xBar[i] = 2 * o.x[i] - o.xPrev[i];
    }
}

// z = Dmat * x   (src/Mesh.cpp:703-753: D is a 0/1 selection matrix)
template <int D>
static void applyD(const Oracle &o, const std::vector<double> &x, std::vector<double> &out) {
    constexpr int n = D * (D + 1);
    for (int e = 0; e < o.E; e++)
        for (int v = 0; v < D + 1; v++) {
            int p = o.F[(size_t)e * (D + 1) + v];
            for (int l = 0; l < D; l++) out[(size_t)n * e + D * v + l] = x[(size_t)D * p + l];
        }
}

// vec = m*xBar + dt^2 * (WD_T * (w*(z-uBar))) ;  x = cg.solve(vec)     (src/MeshIntegrator.cpp:137-138,159-160)
// t = M + dt^2 WD^T WD is numerically diagonal => Jacobi-preconditioned CG == diagonal solve.
template <int D>
static void xUpdate(Oracle &o) {
    constexpr int n = D * (D + 1);
    const double dtsq = o.dt * o.dt;
    for (int p = 0; p < o.P; p++) {
        for (int l = 0; l < D; l++) {
            double s = 0.0;
            for (int k = o.incPtr[p]; k < o.incPtr[p + 1]; k++) {
                int e = o.incIdx[k] / (D + 1), v = o.incIdx[k] % (D + 1);
                size_t idx = (size_t)n * e + D * v + l;
                s += o.w * (o.w * (o.z[idx] - o.uBar[idx]));
            }
            size_t r = (size_t)D * p + l;
            o.vec[r] = (o.m * o.xBar[r]) + dtsq * s;
            o.x[r] = o.vec[r] / o.tdiag[r];
        }
    }
}

// src/MeshIntegrator.cpp:101-191
template <int D>
static double step(Oracle &o, int nIters, double tol) {
    constexpr int n = D * (D + 1);
    const size_t NZ = (size_t)o.E * n;
    predictX<D>(o);
    o.xPrev = o.x;
    o.x = o.xBar;
    applyD<D>(o, o.x, o.z);
    if (!o.stepTaken) std::fill(o.uBar.begin(), o.uBar.end(), 0.0);
    if (o.stepsTaken == 0) applyD<D>(o, o.xPrev, o.z);

    double Ihstart = 0;
    xUpdate<D>(o);

    int i;
    double IhCur = 0, primal = 0, duel = 0;
    for (i = 0; i < nIters; i++) {
        applyD<D>(o, o.x, o.DXpU);
        for (size_t k = 0; k < NZ; k++) o.DXpU[k] += o.uBar[k];
        o.zPrev = o.z;
        IhCur = prox<D>(o, tol);
        if (i == 0) Ihstart = IhCur;
        o.stepTaken = true;
        for (size_t k = 0; k < NZ; k++) o.uBar[k] = o.DXpU[k] - o.z[k];
        xUpdate<D>(o);
        // primal = ||D x - z||, duel = ||z - zPrev||
        double sp = 0.0, sd = 0.0;
        for (int e = 0; e < o.E; e++)
            for (int v = 0; v < D + 1; v++) {
                int p = o.F[(size_t)e * (D + 1) + v];
                for (int l = 0; l < D; l++) {
                    size_t idx = (size_t)n * e + D * v + l;
                    double a = o.x[(size_t)D * p + l] - o.z[idx];
                    double b = o.z[idx] - o.zPrev[idx];
                    sp += a * a; sd += b * b;
                }
            }
        primal = std::sqrt(sp); duel = std::sqrt(sd);
        if (primal < tol && duel < tol) break;
    }
    o.lastAdmmIters = (i == nIters) ? nIters : i + 1;
    o.lastPrimal = primal; o.lastDual = duel;
    // updateAfterStep: Vp <- x  (src/Mesh.cpp:1016-1036); the projection of the FREE nodes there (:1020-1026) is commented out
    if (o.projMask & 2)
        for (int p = 0; p < o.P; p++)
            if (o.mask[p] == BOUNDARY_FREE) projectOntoBoundary<D>(o, p, &o.x[(size_t)D * p]);
    o.Vp = o.x;
    o.stepsTaken++;
    return Ihstart;
}

// src/Mesh.cpp:243-260
template <int D>
static void reOrientElements(Oracle &o) {
    for (int i = 0; i < o.E; i++) {
        Mat<D> Em;
        int *f = &o.F[(size_t)i * (D + 1)];
        for (int j = 0; j < D; j++)
            for (int l = 0; l < D; l++) Em.a[l][j] = o.Vp[(size_t)f[j + 1] * D + l] - o.Vp[(size_t)f[0] * D + l];
        if (det<D>(Em) < 0) std::swap(f[1], f[2]);
    }
}

// exact 1-NN: brute force over uniform bins, lowest index on ties (nanoflann knnSearch(k=1) restated)
struct NNBins {
    int D; int nb[3]; double lo[3], inv[3];
    std::vector<int> start, items;
};
static void buildBins(const Oracle &o, NNBins &b) {
    const int D = o.D;
    b.D = D;
    double hi[3] = {0, 0, 0};
    for (int l = 0; l < 3; l++) { b.lo[l] = 0; b.nb[l] = 1; b.inv[l] = 0; }
    for (int l = 0; l < D; l++) {
        double mn = INFINITY, mx = -INFINITY;
        for (int p = 0; p < o.P; p++) { double v = o.Vp[(size_t)p * D + l]; mn = std::min(mn, v); mx = std::max(mx, v); }
        b.lo[l] = mn; hi[l] = mx;
    }
    int per = std::max(1, (int)std::floor(std::pow((double)o.P / 2.0, 1.0 / D)));
    for (int l = 0; l < D; l++) { b.nb[l] = per; double ext = hi[l] - b.lo[l]; b.inv[l] = ext > 0 ? per / ext : 0.0; }
    size_t nbins = (size_t)b.nb[0] * b.nb[1] * b.nb[2];
    b.start.assign(nbins + 1, 0);
    std::vector<int> binOf(o.P);
    for (int p = 0; p < o.P; p++) {
        int c[3] = {0, 0, 0};
        for (int l = 0; l < D; l++) {
            int k = (int)((o.Vp[(size_t)p * D + l] - b.lo[l]) * b.inv[l]);
            c[l] = std::min(std::max(k, 0), b.nb[l] - 1);
        }
        binOf[p] = (c[2] * b.nb[1] + c[1]) * b.nb[0] + c[0];
        b.start[binOf[p] + 1]++;
    }
    for (size_t i = 0; i < nbins; i++) b.start[i + 1] += b.start[i];
    b.items.resize(o.P);
    std::vector<int> fill(b.start.begin(), b.start.end() - 1);
    for (int p = 0; p < o.P; p++) b.items[fill[binOf[p]]++] = p;
}
static int nearestVertex(const Oracle &o, const NNBins &b, const double *q) {
    const int D = o.D;
    int c[3] = {0, 0, 0};
    double bw[3] = {INFINITY, INFINITY, INFINITY};
    for (int l = 0; l < D; l++) {
        int k = (int)((q[l] - b.lo[l]) * b.inv[l]);
        c[l] = std::min(std::max(k, 0), b.nb[l] - 1);
        if (b.inv[l] > 0) bw[l] = 1.0 / b.inv[l];
    }
    double minbw = std::min(bw[0], std::min(bw[1], bw[2]));
    int best = -1; double bestd = INFINITY;
    int maxr = std::max(b.nb[0], std::max(b.nb[1], b.nb[2]));
    for (int r = 0; r <= maxr; r++) {
        // visit the shell of bins at Chebyshev distance r
        int z0 = (D == 3) ? c[2] - r : 0, z1 = (D == 3) ? c[2] + r : 0;
        for (int kz = z0; kz <= z1; kz++) {
            if (kz < 0 || kz >= b.nb[2]) continue;
            for (int ky = c[1] - r; ky <= c[1] + r; ky++) {
                if (ky < 0 || ky >= b.nb[1]) continue;
                for (int kx = c[0] - r; kx <= c[0] + r; kx++) {
                    if (kx < 0 || kx >= b.nb[0]) continue;
                    int cheb = std::max(std::abs(kx - c[0]), std::max(std::abs(ky - c[1]), (D == 3) ? std::abs(kz - c[2]) : 0));
                    if (cheb != r) continue;
                    size_t bin = ((size_t)kz * b.nb[1] + ky) * b.nb[0] + kx;
                    for (int t = b.start[bin]; t < b.start[bin + 1]; t++) {
                        int p = b.items[t];
                        double d2 = 0.0;
                        for (int l = 0; l < D; l++) { double df = q[l] - o.Vp[(size_t)p * D + l]; d2 += df * df; }
                        if (d2 < bestd || (d2 == bestd && p < best)) { bestd = d2; best = p; }
                    }
                }
            }
        }
        // every unvisited bin is at least r*minbw away from the query (query lies inside bin c, or
        // outside the bbox on the far side of it) -> safe to stop once best is strictly closer.
        if (best >= 0 && std::isfinite(minbw)) {
            double reach = r * minbw;
            if (bestd < reach * reach) break;
        }
    }
    return best;
}

// src/MeshInterpolator.cpp:67-130 (updateMesh), 165-241 (nearestNeighGridMap), 365-404 (smoothMonitorGrid)
static void buildMonitorGrid(Oracle &o) {
    const int D = o.D, DD = D * D;
    o.nx = (int)std::pow((double)((size_t)o.P * D), 1.0 / D);
    o.ny = o.nx;
    o.nz = (D == 2) ? 1 : o.nx;
    size_t rows = (size_t)(o.nx + 1) * (o.ny + 1) * (o.nz + 1);
    o.monGridVals.assign(rows * DD, 0.0);
    double mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int p = 0; p < o.P; p++)
        for (int l = 0; l < D; l++) {
            double v = o.Vp[(size_t)p * D + l];
            mn[l] = (v < mn[l]) ? v : mn[l];
            mx[l] = (v > mx[l]) ? v : mx[l];
        }
    linspace(mn[0], mx[0], o.nx, o.gx);
    linspace(mn[1], mx[1], o.ny, o.gy);
    if (D == 3) linspace(mn[2], mx[2], o.nz, o.gz);

    NNBins bins;
    buildBins(o, bins);
    const int nx = o.nx, ny = o.ny, nz = o.nz;
    if (D == 2) {
#pragma omp parallel for num_threads(o.nthreads) schedule(static)
        for (int i = 0; i < nx + 1; i++)
            for (int j = 0; j < ny + 1; j++) {
                double q[2] = {o.gx[i], o.gy[j]};
                int v = nearestVertex(o, bins, q);
                for (int k = 0; k < DD; k++) o.monGridVals[((size_t)j * (nx + 1) + i) * DD + k] = o.monVals[(size_t)v * DD + k];
            }
    } else {
#pragma omp parallel for num_threads(o.nthreads) schedule(static)
        for (int k3 = 0; k3 < nz + 1; k3++)
            for (int i = 0; i < nx + 1; i++)
                for (int j = 0; j < ny + 1; j++) {
                    double q[3] = {o.gx[i], o.gy[j], o.gz[k3]};
                    int v = nearestVertex(o, bins, q);
                    // sic: (x_i, y_j) stored at i*(nx+1)+j -- transposed w.r.t. the lookup (:198,234 vs :329-336)
                    size_t row = (size_t)(nx + 1) * (ny + 1) * k3 + (size_t)i * (nx + 1) + j;
                    for (int k = 0; k < DD; k++) o.monGridVals[row * DD + k] = o.monVals[(size_t)v * DD + k];
                }
    }
    // smoothing (Jacobi on a copy; interior nodes only)
    int nSmooth = (D == 2) ? 5 : 2;
    std::vector<double> tmp;
    for (int iter = 0; iter < nSmooth; iter++) {
        tmp = o.monGridVals;
        if (D == 2) {
            for (int i = 1; i < nx; i++)
                for (int j = 1; j < ny; j++)
                    for (int k = 0; k < DD; k++) {
                        double v = 0.6 * tmp[((size_t)j * (nx + 1) + i) * DD + k];
                        v += 0.1 * tmp[((size_t)j * (nx + 1) + i + 1) * DD + k];
                        v += 0.1 * tmp[((size_t)j * (nx + 1) + i - 1) * DD + k];
                        v += 0.1 * tmp[((size_t)(j + 1) * (nx + 1) + i) * DD + k];
                        v += 0.1 * tmp[((size_t)(j - 1) * (nx + 1) + i) * DD + k];
                        o.monGridVals[((size_t)j * (nx + 1) + i) * DD + k] = v;
                    }
        } else {
            double h = 0.4 / 6.0;
            size_t sl = (size_t)(nx + 1) * (ny + 1);
            for (int k3 = 1; k3 < nz; k3++)
                for (int i = 1; i < nx; i++)
                    for (int j = 1; j < ny; j++) {
                        size_t c = sl * k3 + (size_t)j * (nx + 1) + i;
                        for (int k = 0; k < DD; k++) {
                            double v = 0.6 * tmp[c * DD + k] + h * tmp[(c + 1) * DD + k] + h * tmp[(c - 1) * DD + k] +
                                       h * tmp[(c + (nx + 1)) * DD + k] + h * tmp[(c - (nx + 1)) * DD + k] +
                                       h * tmp[(c + sl) * DD + k] + h * tmp[(c - sl) * DD + k];
                            o.monGridVals[c * DD + k] = v;
                        }
                    }
        }
    }
}

// ----------------------------------------------------------------------------------
// Monitor functions  (Experiments/TestMonitors/MEx*.h; main.cpp:836-863 index tables)
// ----------------------------------------------------------------------------------
static double mex5_u(double x, double y) {  // MEx5.h:10-14
    double r = std::sqrt(std::pow(x - 0.7, 2.0) + std::pow(y - 0.5, 2.0));
    double theta = std::atan((y - 0.5) / (x - 0.7));
    return 1.0 + 9.0 / (1.0 + 100.0 * r * r * std::pow(std::cos(theta - 20.0 * r * r), 2.0));
}
static double mex53d_u(double x, double y, double z) {  // MEx53D.h:10-16
    double r = std::sqrt(std::pow(x - 0.7, 2.0) + std::pow(y - 0.5, 2.0) + std::pow(z - 0.5, 2));
    double theta = std::atan((y - 0.5) / (x - 0.7));
    double psi = std::atan((z - 0.5) / (x - 0.7));
    return 1.0 + 9.0 / (1.0 + 100.0 * r * r * std::pow(std::cos(theta + psi - 20.0 * r * r), 2.0));
}
// Evaluate monitor `kind` (position in main.cpp's Mvals / Mvals3D tables) at x; M is D x D row-major.
static void monitorEval(int D, int kind, const double *x, double *M) {
    const int DD = D * D;
    for (int k = 0; k < DD; k++) M[k] = 0.0;
    double s = 1.0;
    const double PI = 3.141592653589793238462643383;
    const double h = 2.0 * std::sqrt(std::numeric_limits<double>::epsilon());
    if (D == 2) {
        switch (kind) {
        case 0: s = 1.0; break;  // MEx0
        case 1: {                // MEx1
            double sq = (x[0] - 0.5) * (x[0] - 0.5) + (x[1] - 0.5) * (x[1] - 0.5);
            s = (1 + 20.0 / (1 + 20.0 * (sq)));
        } break;
        case 2: {  // MEx2 (anisotropic)
            double lam1 = 1 + (1.0 / std::cosh(50 * (x[0] + x[1] - 1.0) * (x[0] + x[1] - 1.0)));
            double lam2 = 1.0 / lam1;
            double v[2] = {(1.0 / std::sqrt(2.0)) * 1.0, (1.0 / std::sqrt(2.0)) * 1.0};
            double vo[2] = {(1.0 / std::sqrt(2.0)) * 1.0, -(1.0 / std::sqrt(2.0)) * 1.0};
            for (int a = 0; a < 2; a++)
                for (int b = 0; b < 2; b++) M[a * 2 + b] = (lam1 * v[a]) * v[b] + (lam2 * vo[a]) * vo[b];
            return;
        }
        case 3:  // MEx3
            s = std::sqrt(0.01 / (2.0 + std::cos(8.0 * PI * std::sqrt(std::pow(x[0] - 0.5, 2) + std::pow(x[1] - 0.5, 2)))));
            break;
        case 4: {  // MEx4
            double eps = 0.01;
            double g0 = ((1.0 / (1.0 + std::exp((x[0] + h + x[1] - 1) / (2.0 * eps)))) -
                         (1.0 / (1.0 + std::exp((x[0] - h + x[1] - 1) / (2.0 * eps))))) / (2.0 * h);
            double g1 = ((1.0 / (1.0 + std::exp((x[0] + x[1] + h - 1) / (2.0 * eps)))) -
                         (1.0 / (1.0 + std::exp((x[0] + x[1] - h - 1) / (2.0 * eps))))) / (2.0 * h);
            double nrm = std::sqrt(g0 * g0 + g1 * g1);
            s = std::pow(1 + std::pow(nrm, 2.0), 1.0 / 4.0);
        } break;
        case 5: {  // MEx5
            double g0 = (mex5_u(x[0] + h, x[1]) - mex5_u(x[0] - h, x[1])) / (2.0 * h);
            double g1 = (mex5_u(x[0], x[1] + h) - mex5_u(x[0], x[1] - h)) / (2.0 * h);
            double nrm = std::sqrt(g0 * g0 + g1 * g1);
            s = std::pow(1 + std::pow(nrm, 2.0), 1.0 / 4.0);
        } break;
        default: s = 1.0;
        }
    } else {
        switch (kind) {
        case 0: case 4: s = 1.0; break;  // MEx0<3> (slot 4 is M03D again, main.cpp:858)
        case 1: {                        // MEx13D
            double sq = (x[0] - 0.5) * (x[0] - 0.5) + (x[1] - 0.5) * (x[1] - 0.5) + (x[2] - 0.5) * (x[2] - 0.5);
            s = (1 + 20.0 / (1 + 20.0 * (sq)));
        } break;
        case 2: case 3:  // MEx23D, MEx33D (identical bodies)
            s = std::pow(0.01 / (2.0 + std::cos(8.0 * PI * std::sqrt(std::pow(x[0] - 0.5, 2) + std::pow(x[1] - 0.5, 2) +
                                                                 std::pow(x[2] - 0.5, 2)))), 1.0 / 2.0);
            break;
        case 5: {  // MEx53D (sic: grad(1) assigned twice, MEx53D.h:20-22)
            double g0 = (mex53d_u(x[0] + h, x[1], x[2]) - mex53d_u(x[0] - h, x[1], x[2])) / (2.0 * h);
            double g1 = (mex53d_u(x[0], x[1], x[2] + h) - mex53d_u(x[0], x[1], x[2] - h)) / (2.0 * h);
            double nrm = std::sqrt(g0 * g0 + g1 * g1);
            s = std::pow(1 + std::pow(nrm, 2.0), 1.0 / 4.0);
        } break;
        default: s = 1.0;
        }
    }
    for (int a = 0; a < D; a++) M[a * D + a] = 1.0 * s;
}

template <int D>
static void initIntegrator(Oracle &o) {
    constexpr int n = D * (D + 1);
    // MeshIntegrator ctor (src/MeshIntegrator.cpp:15-62)
    o.x = o.Vp; o.xPrev = o.Vp; o.xBar = o.Vp;
    o.z.assign((size_t)o.E * n, 0.0);
    applyD<D>(o, o.x, o.z);
    o.zPrev = o.z;
    o.uBar.assign(o.z.size(), 0.0);
    o.DXpU = o.z;
    o.vec.assign(o.x.size(), 0.0);
    o.grad.assign(o.x.size(), 0.0);
    // incidence, ascending element order
    o.incPtr.assign(o.P + 1, 0);
    for (size_t k = 0; k < o.F.size(); k++) o.incPtr[o.F[k] + 1]++;
    for (int p = 0; p < o.P; p++) o.incPtr[p + 1] += o.incPtr[p];
    o.incIdx.resize(o.F.size());
    std::vector<int> fill(o.incPtr.begin(), o.incPtr.end() - 1);
    for (size_t k = 0; k < o.F.size(); k++) o.incIdx[fill[o.F[k]]++] = (int)k;
    // t = M + dt^2 * WD^T W D : diagonal tau + dt^2 * sum_{valence} (w*w)*1.0
    o.tdiag.assign(o.x.size(), 0.0);
    const double dtsq = o.dt * o.dt;
    for (int p = 0; p < o.P; p++) {
        double s = 0.0;
        for (int k = o.incPtr[p]; k < o.incPtr[p + 1]; k++) s += (o.w * o.w) * 1.0;
        for (int l = 0; l < D; l++) o.tdiag[(size_t)D * p + l] = o.m + dtsq * s;
    }
}

}  // namespace

// ======================================================================================
// C entry points (ctypes / bench)
// ======================================================================================
extern "C" {

// Mesh<D>::meshInit (src/Mesh.cpp:384-472) + MeshIntegrator ctor.  F is copied and re-oriented.
void *orc_create(int D, int P, int E, const double *Vp, const double *Vc, const int *F, const int *mask,
                 const double *monVertex, double rho, double tau, double dt, int gradUse, int nthreads) {
    Oracle *o = new Oracle();
    o->D = D; o->n = D * (D + 1); o->P = P; o->E = E;
    o->Vp.assign(Vp, Vp + (size_t)P * D);
    o->compMesh = (Vc != nullptr);
    if (Vc) o->Vc.assign(Vc, Vc + (size_t)P * D);
    o->F.assign(F, F + (size_t)E * (D + 1));
    o->mask.assign(mask, mask + P);
    o->monVals.assign(monVertex, monVertex + (size_t)P * D * D);
    o->gradUse = gradUse != 0;
    o->nthreads = nthreads > 0 ? nthreads : 1;
    o->rho = rho; o->tau = tau; o->m = tau; o->dt = dt;
    std::memset(o->bfgsHist, 0, sizeof(o->bfgsHist));
    if (D == 2) reOrientElements<2>(*o); else reOrientElements<3>(*o);
    buildMonitorGrid(*o);
    o->w = 0.5 * std::sqrt(rho);  // src/Mesh.cpp:451 (JSON "w" ignored)
    o->hessInvs.assign((size_t)E * o->n * o->n, 0.0);
    for (int e = 0; e < E; e++)
        for (int k = 0; k < o->n; k++) o->hessInvs[(size_t)e * o->n * o->n + k * o->n + k] = 1.0;
    o->Ih.assign(E, 0.0);
    if (D == 2) initIntegrator<2>(*o); else initIntegrator<3>(*o);
    return o;
}
void orc_destroy(void *h) { delete (Oracle *)h; }
// Deep copy of the whole state (bench.py: every timed epoch restarts from the same post-warm-up snapshot instead of
// repeating the warm-up steps) and the OpenMP thread count of the prox loop (src/Mesh.cpp:429,437 omp_set_num_threads).
void *orc_clone(void *h) { return new Oracle(*(Oracle *)h); }
// Boundary sliding option (see Oracle::projMask); builds the face lists (Mesh::buildFaceList) on first use.
void orc_set_projection(void *h, int mask) {
    Oracle &o = *(Oracle *)h;
    if (mask && o.facePtr.empty()) { if (o.D == 2) buildFaceList<2>(o); else buildFaceList<3>(o); }
    o.projMask = mask;
}
// What the commented-out body of Mesh::setUp would do (src/Mesh.cpp:1006-1014: mapEvaluator->updateMesh(Vp, F);
// mapEvaluator->interpolateMonitor(*Mon)): rebuild the monitor interpolation grid from the CURRENT node positions Vp with the
// monitor evaluated there.  (In the reference interpolateMonitor can run only once -- it deletes monGridTemp,
// src/MeshInterpolator.cpp:243-259 -- so this is a "next" row of SURVEY 8f, not something the reference executes.)
void orc_update_monitor(void *h, const double *monVertex) {
    Oracle &o = *(Oracle *)h;
    o.monVals.assign(monVertex, monVertex + (size_t)o.P * o.D * o.D);
    buildMonitorGrid(o);
}
int orc_face_count(void *h) { Oracle &o = *(Oracle *)h; return (int)(o.faceList.size() / o.D); }
void orc_project(void *h, int nodeId, double *node) {
    Oracle &o = *(Oracle *)h;
    if (o.facePtr.empty()) { if (o.D == 2) buildFaceList<2>(o); else buildFaceList<3>(o); }
    if (o.D == 2) projection2D(o, nodeId, node); else projection3D(o, nodeId, node);
}
void orc_set_threads(void *h, int nthreads) { ((Oracle *)h)->nthreads = nthreads > 0 ? nthreads : 1; }

int orc_step(void *h, int nIters, double tol, double *Ih, int *iters, double *primal, double *dual) {
    Oracle &o = *(Oracle *)h;
    double v = (o.D == 2) ? step<2>(o, nIters, tol) : step<3>(o, nIters, tol);
    if (Ih) *Ih = v;
    if (iters) *iters = o.lastAdmmIters;
    if (primal) *primal = o.lastPrimal;
    if (dual) *dual = o.lastDual;
    return o.error;
}
int orc_euler_step(void *h, double *Ih) {
    Oracle &o = *(Oracle *)h;
    double v = (o.D == 2) ? eulerStep<2>(o) : eulerStep<3>(o);
    if (Ih) *Ih = v;
    return o.error;
}
// Method 2 primitives (the Newton loop of src/Mesh.cpp:1263-1341 is driven from tests/oracle_lib.py, with the linear
// solves done by the reference's own LASolver in oracle/_ref or by the restated ILU(0)/CGSTAB of tests/pcg_oracle.py).
int orc_euler_step_mod(void *h, const double *x, double *grad, double *Ih) {
    Oracle &o = *(Oracle *)h;
    std::vector<double> xv(x, x + (size_t)o.P * o.D), g((size_t)o.P * o.D);
    double v = (o.D == 2) ? eulerStepMod<2>(o, xv, g) : eulerStepMod<3>(o, xv, g);
    std::memcpy(grad, g.data(), g.size() * sizeof(double));
    if (Ih) *Ih = v;
    return o.error;
}
int orc_build_euler_jac(void *h, const int *ia, const int *ja, double *a) {
    Oracle &o = *(Oracle *)h;
    if (o.D == 2) buildEulerJac<2>(o, ia, ja, a); else buildEulerJac<3>(o, ia, ja, a);
    return o.error;
}
void orc_set_x(void *h, const double *x) { Oracle &o = *(Oracle *)h; std::memcpy(o.x.data(), x, o.x.size() * sizeof(double)); }
double orc_energy(void *h) {
    Oracle &o = *(Oracle *)h;
    return (o.D == 2) ? computeEnergy<2>(o) : computeEnergy<3>(o);
}
void orc_get_x(void *h, double *out) { Oracle &o = *(Oracle *)h; std::memcpy(out, o.x.data(), o.x.size() * sizeof(double)); }
void orc_get_z(void *h, double *out) { Oracle &o = *(Oracle *)h; std::memcpy(out, o.z.data(), o.z.size() * sizeof(double)); }
void orc_get_u(void *h, double *out) { Oracle &o = *(Oracle *)h; std::memcpy(out, o.uBar.data(), o.uBar.size() * sizeof(double)); }
void orc_get_hess(void *h, double *out) { Oracle &o = *(Oracle *)h; std::memcpy(out, o.hessInvs.data(), o.hessInvs.size() * sizeof(double)); }
void orc_get_F(void *h, int *out) { Oracle &o = *(Oracle *)h; std::memcpy(out, o.F.data(), o.F.size() * sizeof(int)); }
void orc_grid_dims(void *h, int *dims) { Oracle &o = *(Oracle *)h; dims[0] = o.nx; dims[1] = o.ny; dims[2] = o.nz; }
void orc_get_grid(void *h, double *out) { Oracle &o = *(Oracle *)h; std::memcpy(out, o.monGridVals.data(), o.monGridVals.size() * sizeof(double)); }
void orc_get_axes(void *h, double *ax, double *ay, double *az) {
    Oracle &o = *(Oracle *)h;
    std::memcpy(ax, o.gx.data(), o.gx.size() * sizeof(double));
    std::memcpy(ay, o.gy.data(), o.gy.size() * sizeof(double));
    if (o.D == 3 && az) std::memcpy(az, o.gz.data(), o.gz.size() * sizeof(double));
}
void orc_get_stats(void *h, long long *hist52, long long *proxCalls) {
    Oracle &o = *(Oracle *)h;
    std::memcpy(hist52, o.bfgsHist, sizeof(o.bfgsHist));
    *proxCalls = o.proxCalls;
}
double orc_w(void *h) { return ((Oracle *)h)->w; }

// evalMonitorOnGrid at npts points (row-major npts x D) -> npts x D^2
void orc_eval_monitor(void *h, const double *pts, int npts, double *out) {
    Oracle &o = *(Oracle *)h;
    for (int i = 0; i < npts; i++) {
        if (o.D == 2) { Mat<2> m; evalMonitorOnGrid<2>(o, pts + 2 * i, m); std::memcpy(out + 4 * (size_t)i, m.a, 4 * sizeof(double)); }
        else { Mat<3> m; evalMonitorOnGrid<3>(o, pts + 3 * i, m); std::memcpy(out + 9 * (size_t)i, m.a, 9 * sizeof(double)); }
    }
}

// blockGrad on ne stand-alone simplices (element ids 0..ne-1 index F/mask/DXpU of the handle):
// zs: ne x n, dxpu: ne x n (used when regularize), out energy[ne], grad[ne x n], igt[ne]
void orc_block_grad(void *h, int ne, const double *zs, const double *dxpu, int computeGrad, int regularize,
                    int applyMask, double *energy, double *grad, double *igt) {
    Oracle &o = *(Oracle *)h;
    const int n = o.n;
    std::vector<double> saved;
    if (regularize) { saved = o.DXpU; std::memcpy(o.DXpU.data(), dxpu, (size_t)ne * n * sizeof(double)); }
    for (int e = 0; e < ne; e++) {
        double xi[12];
        if (o.compMesh)
            for (int v = 0; v < o.D + 1; v++)
                for (int l = 0; l < o.D; l++) xi[v * o.D + l] = o.Vc[(size_t)o.F[(size_t)e * (o.D + 1) + v] * o.D + l];
        double g[12] = {0}, I = 0, en;
        if (o.D == 2)
            en = applyMask ? computeBlockGrad<2>(o, e, zs + (size_t)e * n, xi, g, computeGrad, regularize, I)
                           : blockGrad<2>(o, e, zs + (size_t)e * n, xi, g, computeGrad, regularize, I);
        else
            en = applyMask ? computeBlockGrad<3>(o, e, zs + (size_t)e * n, xi, g, computeGrad, regularize, I)
                           : blockGrad<3>(o, e, zs + (size_t)e * n, xi, g, computeGrad, regularize, I);
        energy[e] = en;
        if (igt) igt[e] = I;
        if (grad) std::memcpy(grad + (size_t)e * n, g, n * sizeof(double));
    }
    if (regularize) o.DXpU = saved;
}

// Monitor table: out[P x D^2] = monitor `kind` at the rows of X (MonitorFunction::evaluateAtVertices,
// src/MonitorFunction.cpp:15-32)
void orc_monitor_eval(int D, int kind, int P, const double *X, double *out) {
    for (int p = 0; p < P; p++) monitorEval(D, kind, X + (size_t)p * D, out + (size_t)p * D * D);
}

// n x n inverse as used for the FD Hessian (src/Mesh.cpp:816)
int orc_lu_inverse(int n, const double *A, double *Ainv) { return luInverse(n, A, Ainv) ? 0 : 1; }

// utils::generateUniformRectMesh<D> (src/MeshUtils.h:82-335).  xa..zb are truncated to int as in
// the reference (:90-96).  Caller allocates Vp[(nPnts) x D], F[nElem x (D+1)], mask[nPnts].
void orc_uniform_rect_mesh(int D, int nx, int ny, int nz, double xa_, double xb_, double ya_, double yb_,
                           double za_, double zb_, int bType, double *Vp, int *F, int *mask) {
    int xa = (int)xa_, xb = (int)xb_, ya = (int)ya_, yb = (int)yb_;
    int za = (D == 3) ? (int)za_ : 0, zb = (D == 3) ? (int)zb_ : 0;
    double hx = (xb - xa) / ((double)nx), hy = (yb - ya) / ((double)ny);
    double hz = (D == 3) ? (zb - za) / ((double)nz) : 0;
    size_t off = 0;
    if (D == 2) {
        for (int j = 0; j <= ny; j++) for (int i = 0; i <= nx; i++) { Vp[off * 2] = xa + hx * i; Vp[off * 2 + 1] = ya + hy * j; off++; }
        for (int j = 0; j < ny; j++) for (int i = 0; i < nx; i++) { Vp[off * 2] = xa + hx * i + hx / 2.0; Vp[off * 2 + 1] = ya + hy * j + hy / 2.0; off++; }
        int stride = (nx + 1) * (ny + 1);
        off = 0;
        for (int j = 0; j < ny; j++)
            for (int i = 0; i < nx; i++) {
                int a = i + j * (nx + 1), b = i + 1 + j * (nx + 1), c = i + (j + 1) * (nx + 1), d = i + 1 + (j + 1) * (nx + 1);
                int mid = stride + i + j * nx;
                int t[4][3] = {{a, mid, c}, {mid, d, c}, {mid, d, b}, {a, b, mid}};
                for (int q = 0; q < 4; q++) { for (int r = 0; r < 3; r++) F[off * 3 + r] = t[q][r]; off++; }
            }
        int nP = (nx + 1) * (ny + 1) + nx * ny;
        for (int i = 0; i < nP; i++) mask[i] = INTERIOR;
        for (int i = 0; i < (nx + 1) * (ny + 1); i++) {
            int iOff = i % (nx + 1), jOff = i / (ny + 1);  // sic (:162-163)
            bool bnd = (iOff == 0) || (iOff == nx) || (jOff == 0) || (jOff == ny);
            mask[i] = bnd ? bType : INTERIOR;
            if ((iOff == 0 && jOff == 0) || (iOff == nx && jOff == 0) || (iOff == 0 && jOff == ny) || (iOff == nx && jOff == ny))
                mask[i] = BOUNDARY_FIXED;
        }
    } else {
        for (int k = 0; k <= nz; k++) for (int j = 0; j <= ny; j++) for (int i = 0; i <= nx; i++) {
            Vp[off * 3] = xa + hx * i; Vp[off * 3 + 1] = ya + hy * j; Vp[off * 3 + 2] = za + hz * k; off++; }
        for (int k = 0; k < nz; k++) for (int j = 0; j < ny; j++) for (int i = 0; i < nx; i++) {
            Vp[off * 3] = xa + hx * i + hx / 2.0; Vp[off * 3 + 1] = ya + hy * j + hy / 2.0; Vp[off * 3 + 2] = za + hz * k + hz / 2.0; off++; }
        int stride = (nx + 1) * (ny + 1) * (nz + 1);
        off = 0;
        auto V = [&](int i, int j, int k) { return i + j * (nx + 1) + k * (nx + 1) * (ny + 1); };
        for (int k = 0; k < nz; k++)
            for (int j = 0; j < ny; j++)
                for (int i = 0; i < nx; i++) {
                    int mid = stride + i + j * nx + k * (nx * ny);
                    int t[12][3] = {
                        {V(i, j, k), V(i + 1, j, k), V(i + 1, j + 1, k)},              // bot
                        {V(i, j, k), V(i, j + 1, k), V(i + 1, j + 1, k)},
                        {V(i, j, k + 1), V(i + 1, j, k + 1), V(i + 1, j + 1, k + 1)},  // top
                        {V(i, j, k + 1), V(i, j + 1, k + 1), V(i + 1, j + 1, k + 1)},
                        {V(i, j, k), V(i, j + 1, k), V(i, j + 1, k + 1)},              // left
                        {V(i, j, k), V(i, j, k + 1), V(i, j + 1, k + 1)},
                        {V(i + 1, j, k), V(i + 1, j + 1, k), V(i + 1, j + 1, k + 1)},  // right
                        {V(i + 1, j, k), V(i + 1, j, k + 1), V(i + 1, j + 1, k + 1)},
                        {V(i, j, k), V(i + 1, j, k), V(i, j, k + 1)},                  // back
                        {V(i + 1, j, k), V(i + 1, j, k + 1), V(i, j, k + 1)},
                        {V(i, j + 1, k), V(i + 1, j + 1, k), V(i, j + 1, k + 1)},      // front
                        {V(i + 1, j + 1, k), V(i + 1, j + 1, k + 1), V(i, j + 1, k + 1)}};
                    for (int q = 0; q < 12; q++) { for (int r = 0; r < 3; r++) F[off * 4 + r] = t[q][r]; F[off * 4 + 3] = mid; off++; }
                }
        int nP = (nx + 1) * (ny + 1) * (nz + 1) + nx * ny * nz;
        for (int i = 0; i < nP; i++) mask[i] = INTERIOR;
        for (int k = 0; k < nz + 1; k++)
            for (int i = 0; i < (nx + 1) * (ny + 1); i++) {
                int iOff = i / (nx + 1), jOff = i % (ny + 1);  // sic (:297-298)
                bool bnd = (iOff == 0) || (iOff == nx) || (jOff == 0) || (jOff == ny) || (k == 0) || (k == nz);
                int o2 = k * (nx + 1) * (ny + 1) + i;
                if (bnd) mask[o2] = bType;
                bool corner = (iOff == 0 && jOff == 0) || (iOff == nx && jOff == 0) || (iOff == 0 && jOff == ny) ||
                              (iOff == nx && jOff == ny) || (iOff == 0 && k == 0) || (iOff == nx && k == 0) ||
                              (iOff == 0 && k == nz) || (iOff == nx && k == nz) || (k == 0 && jOff == 0) ||
                              (k == nz && jOff == 0) || (k == 0 && jOff == ny) || (k == nz && jOff == ny);
                if (corner) mask[o2] = BOUNDARY_FIXED;
            }
    }
}

}  // extern "C"
