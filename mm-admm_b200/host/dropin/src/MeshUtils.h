// utils:: mesh generators and readers of the MM-ADMM driver layer, with the reference's signatures
// (src/MeshUtils.h) so that main.cpp compiles against this header unchanged.  Everything here runs once, on the
// host, before the first step.  The results (points, simplices, masks) are value-identical to the reference's for
// the generators it can actually run; the implementations are linear-time where the reference is quadratic
// (removeRow-in-a-loop and the relabelling triple loop of meshFromLevelSetFun).
#ifndef MESH_UTILS_H
#define MESH_UTILS_H

#include <Eigen/Dense>
#include "MeshIO.h"
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <fstream>
#include <functional>
#include <iostream>
#include <limits>
#include <map>
#include <set>
#include <sstream>
#include <string>
#include <tuple>
#include <unordered_map>
#include <vector>

#include "Mesh.h"
#include "NodeType.h"
#include "MeshIO.h"

using namespace std;

namespace utils {

// [xa, xb] split into ns intervals (ns + 1 points); src/MeshUtils.h:24-29
inline void linspace(double xa, double xb, int ns, vector<double> &x) {
    x.resize(ns + 1);
    for (int i = 0; i <= ns; i++) x[i] = xa + ((double)i) * (xb - xa) / ns;
}

inline void computeCentroid2D(int simplexId, Eigen::MatrixXd *X, Eigen::MatrixXi *F, Eigen::Vector<double, 2> &pnt) {
    for (int l = 0; l < 2; l++)
        pnt[l] = 1.0 / 3.0 * ((*X)((*F)(simplexId, 0), l) + (*X)((*F)(simplexId, 1), l) + (*X)((*F)(simplexId, 2), l));
}

// Index of the uniform-axis cell holding w (src/MeshUtils.h:45-54).  The reference stores the guess in a uint32_t,
// so a negative guess wraps to a huge value and is clamped to the LAST cell, not the first; kept.
inline int findLimInfMeshPoint(double w, vector<double> &w_mesh) {
    uint32_t guess = (int)((w - w_mesh.at(0)) / (w_mesh.at(1) - w_mesh.at(0)));
    if (guess > w_mesh.size() - 2) guess = w_mesh.size() - 2;
    return guess;
}

inline void biLinearInterpolation(double x, double y, double xMesh[2], double yMesh[2], double *coefs) {
    const double s = (1 / ((xMesh[1] - xMesh[0]) * (yMesh[1] - yMesh[0])));
    coefs[0] = s * (xMesh[1] - x) * (yMesh[1] - y);
    coefs[1] = s * (x - xMesh[0]) * (yMesh[1] - y);
    coefs[2] = s * (xMesh[1] - x) * (y - yMesh[0]);
    coefs[3] = s * (x - xMesh[0]) * (y - yMesh[0]);
}

inline void triLinearInterpolation(double x, double y, double z, double xMesh[2], double yMesh[2], double zMesh[2],
                                   double coefs[8]) {
    const double xd = (x - xMesh[0]) / (xMesh[1] - xMesh[0]);
    const double yd = (y - yMesh[0]) / (yMesh[1] - yMesh[0]);
    const double zd = (z - zMesh[0]) / (zMesh[1] - zMesh[0]);
    for (int c = 0; c < 8; c++)
        coefs[c] = ((c & 1) ? xd : 1 - xd) * ((c & 2) ? yd : 1 - yd) * ((c & 4) ? zd : 1 - zd);
}

// Uniform box mesh: every cell is split around its centre into 4 triangles / 12 tetrahedra (src/MeshUtils.h:82-335).
// Points: lattice nodes (x fastest) followed by the cell centres.  Quirks kept: the box bounds are truncated to int;
// the 2D mask tests i/(ny+1), the 3D mask i/(nx+1) and i%(ny+1); box edges and corners are always FIXED.
template <int D>
inline void generateUniformRectMesh(unordered_map<string,double> params, Eigen::MatrixXd *Vc,
        Eigen::MatrixXi *F, vector<NodeType> *boundaryMask, NodeType bType) {
    const int nx = (int)params["nx"], ny = (int)params["ny"], nz = (D == 3) ? (int)params["nz"] : 0;
    const int xa = params["xa"], xb = params["xb"], ya = params["ya"], yb = params["yb"];
    const int za = (D == 3) ? params["za"] : 0, zb = (D == 3) ? params["zb"] : 0;
    const double hx = (xb - xa) / ((double)nx), hy = (yb - ya) / ((double)ny);
    const double hz = (D == 3) ? (zb - za) / ((double)nz) : 0;
    Eigen::MatrixXd &V = *Vc;
    Eigen::MatrixXi &T = *F;
    vector<NodeType> &mask = *boundaryMask;
    int o = 0;
    if (D == 2) {
        const int row = nx + 1, nLat = (nx + 1) * (ny + 1);
        for (int j = 0; j <= ny; j++)
            for (int i = 0; i <= nx; i++, o++) { V(o, 0) = xa + hx * i; V(o, 1) = ya + hy * j; }
        for (int j = 0; j < ny; j++)
            for (int i = 0; i < nx; i++, o++) { V(o, 0) = xa + hx * i + hx / 2.0; V(o, 1) = ya + hy * j + hy / 2.0; }
        int t = 0;
        for (int j = 0; j < ny; j++)
            for (int i = 0; i < nx; i++) {
                const int sw = i + j * row, se = sw + 1, nw = sw + row, ne = nw + 1, c = nLat + i + j * nx;
                const int tri[4][3] = {{sw, c, nw}, {c, ne, nw}, {c, ne, se}, {sw, se, c}};   // left, top, right, bottom
                for (int q = 0; q < 4; q++, t++)
                    for (int r = 0; r < 3; r++) T(t, r) = tri[q][r];
            }
        for (int p = 0; p < nLat + nx * ny; p++) mask.at(p) = NodeType::INTERIOR;
        for (int p = 0; p < nLat; p++) {
            const int iOff = p % (nx + 1), jOff = p / (ny + 1);
            const bool onX = (iOff == 0 || iOff == nx), onY = (jOff == 0 || jOff == ny);
            mask.at(p) = (onX || onY) ? bType : NodeType::INTERIOR;
            if (onX && onY) mask.at(p) = NodeType::BOUNDARY_FIXED;
        }
        return;
    }
    const int row = nx + 1, slab = (nx + 1) * (ny + 1), nLat = slab * (nz + 1);
    for (int k = 0; k <= nz; k++)
        for (int j = 0; j <= ny; j++)
            for (int i = 0; i <= nx; i++, o++) { V(o, 0) = xa + hx * i; V(o, 1) = ya + hy * j; V(o, 2) = za + hz * k; }
    for (int k = 0; k < nz; k++)
        for (int j = 0; j < ny; j++)
            for (int i = 0; i < nx; i++, o++) {
                V(o, 0) = xa + hx * i + hx / 2.0; V(o, 1) = ya + hy * j + hy / 2.0; V(o, 2) = za + hz * k + hz / 2.0;
            }
    int t = 0;
    for (int k = 0; k < nz; k++)
        for (int j = 0; j < ny; j++)
            for (int i = 0; i < nx; i++) {
                int c[2][2][2];   // cell corners c[dz][dy][dx]
                for (int dz = 0; dz < 2; dz++)
                    for (int dy = 0; dy < 2; dy++)
                        for (int dx = 0; dx < 2; dx++) c[dz][dy][dx] = (i + dx) + (j + dy) * row + (k + dz) * slab;
                const int mid = nLat + i + j * nx + k * (nx * ny);
                // two tetrahedra per face, faces in the reference's order: bottom, top, left, right, back, front
                const int tet[12][3] = {
                    {c[0][0][0], c[0][0][1], c[0][1][1]}, {c[0][0][0], c[0][1][0], c[0][1][1]},
                    {c[1][0][0], c[1][0][1], c[1][1][1]}, {c[1][0][0], c[1][1][0], c[1][1][1]},
                    {c[0][0][0], c[0][1][0], c[1][1][0]}, {c[0][0][0], c[1][0][0], c[1][1][0]},
                    {c[0][0][1], c[0][1][1], c[1][1][1]}, {c[0][0][1], c[1][0][1], c[1][1][1]},
                    {c[0][0][0], c[0][0][1], c[1][0][0]}, {c[0][0][1], c[1][0][1], c[1][0][0]},
                    {c[0][1][0], c[0][1][1], c[1][1][0]}, {c[0][1][1], c[1][1][1], c[1][1][0]}};
                for (int q = 0; q < 12; q++, t++) {
                    for (int r = 0; r < 3; r++) T(t, r) = tet[q][r];
                    T(t, 3) = mid;
                }
            }
    for (int p = 0; p < nLat + nx * ny * nz; p++) mask.at(p) = NodeType::INTERIOR;
    for (int k = 0; k <= nz; k++)
        for (int p = 0; p < slab; p++) {
            const int iOff = p / (nx + 1), jOff = p % (ny + 1);
            const bool onI = (iOff == 0 || iOff == nx), onJ = (jOff == 0 || jOff == ny), onK = (k == 0 || k == nz);
            const int id = k * slab + p;
            if (onI || onJ || onK) mask.at(id) = bType;
            if ((onI && onJ) || (onI && onK) || (onJ && onK)) mask.at(id) = NodeType::BOUNDARY_FIXED;
        }
}

// src/MeshUtils.h:338-346
inline void removeRow(Eigen::MatrixXi &matrix, unsigned int rowToRemove) {
    const int rows = matrix.rows() - 1, cols = matrix.cols();
    Eigen::MatrixXi out(rows, cols);
    for (int j = 0; j < cols; j++)
        for (int i = 0; i < rows; i++) out(i, j) = matrix(i < (int)rowToRemove ? i : i + 1, j);
    matrix = out;
}

// all rows whose flag is set, in one pass (what a removeRow loop over descending ids produces)
inline void removeRows(Eigen::MatrixXi &matrix, const vector<char> &drop) {
    int keep = 0;
    for (int i = 0; i < matrix.rows(); i++) keep += !drop[i];
    Eigen::MatrixXi out(keep, matrix.cols());
    for (int i = 0, o = 0; i < matrix.rows(); i++) {
        if (drop[i]) continue;
        for (int j = 0; j < matrix.cols(); j++) out(o, j) = matrix(i, j);
        o++;
    }
    matrix = out;
}

inline double phiSdf(double x, double y, std::function<double(double,double)> phiFun) { return phiFun(x, y); }
inline double phiSdf(double x, double y, double z, std::function<double(double,double,double)> phiFun) { return phiFun(x, y, z); }

// src/MeshUtils.h:369-386: the 2D projection hard-codes the unit normal of a circle centred at (0.5, 0.5)
inline void interpolateBoundaryLocation(Eigen::Vector<double, 2> &pnt, std::function<double(double, double)> phiFun,
        std::vector<int> &nVals, std::vector<std::tuple<double, double>> &bb) {
    (void)nVals; (void)bb;
    const double xv = pnt[0] - 0.5, yv = pnt[1] - 0.5;
    Eigen::Vector<double, 2> g;
    g[0] = xv / sqrt(xv * xv + yv * yv);
    g[1] = yv / sqrt(xv * xv + yv * yv);
    pnt = pnt - phiFun(pnt[0], pnt[1]) * g;
}
// src/MeshUtils.h:388-402: central-difference normal
inline void interpolateBoundaryLocation(Eigen::Vector<double, 3> &pnt, std::function<double(double, double, double)> phiFun,
        std::vector<int> &nVals, std::vector<std::tuple<double, double>> &bb) {
    (void)nVals; (void)bb;
    const double h = 2.0 * sqrt(std::numeric_limits<double>::epsilon());
    Eigen::Vector<double, 3> g;
    g[0] = (phiFun(pnt[0] + h, pnt[1], pnt[2]) - phiFun(pnt[0] - h, pnt[1], pnt[2])) / (2.0 * h);
    g[1] = (phiFun(pnt[0], pnt[1] + h, pnt[2]) - phiFun(pnt[0], pnt[1] - h, pnt[2])) / (2.0 * h);
    g[2] = (phiFun(pnt[0], pnt[1], pnt[2] + h) - phiFun(pnt[0], pnt[1], pnt[2] - h)) / (2.0 * h);
    g.normalize();
    pnt = pnt - phiFun(pnt[0], pnt[1], pnt[2]) * g;
}

// Mesh of {phi < 0} cut out of the uniform box mesh (src/MeshUtils.h:404-538): simplices whose vertices are all
// outside (phi > -1e-12) are dropped, outside vertices of the remaining simplices are projected onto phi = 0, unused
// points are removed and the rest renumbered in ascending order.  Value-identical to the reference, quirks included:
//  * the boundary mask is NOT compacted: bType marks are left at the OLD point numbers, and the final pass marks
//    FIXED at the NEW numbers wherever |phi| < 1e-12 -- so the mask keeps the box mesh's length;
//  * Vc is resized but never filled by the reference (it reads uninitialised memory); here Vc = the compacted
//    un-projected box coordinates, which is what the code evidently intended and is only read when CompMesh is on.
inline void meshFromLevelSetFun(std::function<double(double, double)> phiFun, std::vector<int> &nVals,
        std::vector<std::tuple<double, double>> &bb, Eigen::MatrixXd *Vc, Eigen::MatrixXd *Vp, Eigen::MatrixXi *F,
        std::vector<NodeType> *boundaryMask, NodeType bType) {
    const double EPS = 1e-12;
    const int D = 2, nx = nVals.at(0), ny = nVals.at(1);
    std::unordered_map<std::string,double> params;
    params["nx"] = nx; params["ny"] = ny;
    params["xa"] = std::get<0>(bb.at(0)); params["xb"] = std::get<1>(bb.at(0));
    params["ya"] = std::get<0>(bb.at(1)); params["yb"] = std::get<1>(bb.at(1));
    const int nBox = (nx + 1) * (ny + 1) + nx * ny;
    Vp->resize(nBox, D);
    F->resize(4 * nx * ny, D + 1);
    boundaryMask->resize(nBox);
    generateUniformRectMesh<D>(params, Vp, F, boundaryMask, bType);
    for (auto &mk : *boundaryMask) mk = NodeType::INTERIOR;
    Eigen::MatrixXd box(*Vp);

    std::vector<double> phiAt(nBox);
    for (int p = 0; p < nBox; p++) phiAt[p] = phiFun((*Vp)(p, 0), (*Vp)(p, 1));
    std::vector<char> drop(F->rows(), 0);
    for (int s = 0; s < F->rows(); s++)
        drop[s] = phiAt[(*F)(s, 0)] > -EPS && phiAt[(*F)(s, 1)] > -EPS && phiAt[(*F)(s, 2)] > -EPS;
    removeRows(*F, drop);

    std::vector<char> used(nBox, 0);
    for (int i = 0; i < F->rows(); i++) for (int j = 0; j < F->cols(); j++) used[(*F)(i, j)] = 1;
    Eigen::Vector<double, D> xPnt;
    for (int p = 0; p < nBox; p++) {
        if (!used[p]) continue;
        xPnt = (*Vp)(p, Eigen::placeholders::all);
        const double phi = phiAt[p];
        if (std::abs(phi) < EPS || phi > 0) {
            interpolateBoundaryLocation(xPnt, phiFun, nVals, bb);
            boundaryMask->at(p) = bType;
        }
        (*Vp)(p, Eigen::placeholders::all) = xPnt;
    }
    std::vector<int> newId(nBox, -1);
    int nUsed = 0;
    for (int p = 0; p < nBox; p++) if (used[p]) newId[p] = nUsed++;
    Eigen::MatrixXd VpNew(nUsed, D), VcNew(nUsed, D);
    for (int p = 0; p < nBox; p++) {
        if (!used[p]) continue;
        for (int l = 0; l < D; l++) { VpNew(newId[p], l) = (*Vp)(p, l); VcNew(newId[p], l) = box(p, l); }
    }
    for (int i = 0; i < F->rows(); i++) for (int j = 0; j < F->cols(); j++) (*F)(i, j) = newId[(*F)(i, j)];
    *Vp = VpNew;
    *Vc = VcNew;
    for (int p = 0; p < Vp->rows(); p++)
        if (std::abs(phiFun((*Vp)(p, 0), (*Vp)(p, 1))) < EPS) boundaryMask->at(p) = NodeType::BOUNDARY_FIXED;
}

// 3D counterpart (src/MeshUtils.h:540-667).  The reference version cannot be reproduced: it renumbers the points in
// DESCENDING order while relabelling F in ascending order (labels get mapped twice), and finally deletes the
// caller's Vc / Vp and re-points its own local copies of the pointers, leaving the caller with dangling pointers.
// This version does what the 2D one does (ascending renumbering, results written into the caller's matrices).
inline void meshFromLevelSetFun(std::function<double(double, double, double)> phiFun, std::vector<int> &nVals,
        std::vector<std::tuple<double, double>> &bb, Eigen::MatrixXd *Vc, Eigen::MatrixXd *Vp, Eigen::MatrixXi *F,
        std::vector<NodeType> *boundaryMask, NodeType bType) {
    const double EPS = 1e-12;
    const int D = 3, nx = nVals.at(0), ny = nVals.at(1), nz = nVals.at(2);
    std::unordered_map<std::string,double> params;
    params["nx"] = nx; params["ny"] = ny; params["nz"] = nz;
    params["xa"] = std::get<0>(bb.at(0)); params["xb"] = std::get<1>(bb.at(0));
    params["ya"] = std::get<0>(bb.at(1)); params["yb"] = std::get<1>(bb.at(1));
    params["za"] = std::get<0>(bb.at(2)); params["zb"] = std::get<1>(bb.at(2));
    const int nBox = (nx + 1) * (ny + 1) * (nz + 1) + nx * ny * nz;
    Vp->resize(nBox, D);
    F->resize(12 * nx * ny * nz, D + 1);
    boundaryMask->resize(nBox);
    generateUniformRectMesh<3>(params, Vp, F, boundaryMask, bType);
    for (auto &mk : *boundaryMask) mk = NodeType::INTERIOR;
    Eigen::MatrixXd box(*Vp);

    std::vector<double> phiAt(nBox);
    for (int p = 0; p < nBox; p++) phiAt[p] = phiFun((*Vp)(p, 0), (*Vp)(p, 1), (*Vp)(p, 2));
    std::vector<char> drop(F->rows(), 0);
    for (int s = 0; s < F->rows(); s++) {
        bool out = true;
        for (int v = 0; v < 4; v++) out = out && phiAt[(*F)(s, v)] > -EPS;
        drop[s] = out;
    }
    removeRows(*F, drop);
    std::vector<char> used(nBox, 0);
    for (int i = 0; i < F->rows(); i++) for (int j = 0; j < F->cols(); j++) used[(*F)(i, j)] = 1;
    std::vector<int> newId(nBox, -1);
    int nUsed = 0;
    for (int p = 0; p < nBox; p++) if (used[p]) newId[p] = nUsed++;
    Eigen::MatrixXd VpNew(nUsed, D), VcNew(nUsed, D);
    std::vector<NodeType> maskNew(nUsed, NodeType::INTERIOR);
    Eigen::Vector<double, 3> xPnt;
    for (int p = 0; p < nBox; p++) {
        if (!used[p]) continue;
        xPnt = (*Vp)(p, Eigen::placeholders::all);
        if (phiAt[p] > -EPS) {
            interpolateBoundaryLocation(xPnt, phiFun, nVals, bb);
            maskNew[newId[p]] = bType;
        }
        for (int l = 0; l < D; l++) { VpNew(newId[p], l) = xPnt[l]; VcNew(newId[p], l) = box(p, l); }
    }
    for (int i = 0; i < F->rows(); i++) for (int j = 0; j < F->cols(); j++) (*F)(i, j) = newId[(*F)(i, j)];
    *Vp = VpNew;
    *Vc = VcNew;
    *boundaryMask = maskNew;
}

// Comma-separated simplices and points, one mask value per line (src/MeshUtils.h:669-733).  The mask reader of the
// reference pushes one extra (repeated) entry at end of file; kept, the Mesh only reads the first P entries.
// utils::readTriangles (src/MeshUtils.h:669-733).  Whole-file scans instead of a stringstream per line (MeshIO.h); each of the
// three files may also be the binary table Mesh::output* writes with MMADMM_BINARY_IO=1.  The mask keeps the reference's quirk:
// its `while (fin) { fin >> tmp; push_back(tmp); }` loop pushes one entry more than the file holds (the failed extraction
// at end of file stores 0 since C++11), so the mask has P + 1 entries.
inline void readTriangles(int D, const char *triFileName, const char *pntFileName, const char *maskFileName,
        Eigen::MatrixXi &F, Eigen::MatrixXd &Vp, vector<NodeType> &boundaryMask) {
    vector<double> pntData;
    vector<int> triData, maskData;
    if (!meshio::read_ints(triFileName, triData)) { cerr << "cannot open " << triFileName << endl; assert(false); exit(1); }
    if (!meshio::read_doubles(pntFileName, pntData)) { cerr << "cannot open " << pntFileName << endl; exit(1); }
    boundaryMask.clear();
    if (meshio::read_ints(maskFileName, maskData)) {
        for (size_t i = 0; i < maskData.size(); i++) boundaryMask.push_back((NodeType)maskData[i]);
        boundaryMask.push_back((NodeType)0);
    } else {
        boundaryMask.push_back((NodeType)0);     // unreadable file: the reference's loop still pushes one (zero) entry
    }
    F.resize(triData.size() / (D + 1), D + 1);
    Vp.resize(pntData.size() / D, D);
    size_t off = 0;
    for (int i = 0; i < F.rows(); i++) for (int j = 0; j < F.cols(); j++) F(i, j) = triData.at(off++);
    off = 0;
    for (int i = 0; i < Vp.rows(); i++) for (int j = 0; j < Vp.cols(); j++) Vp(i, j) = pntData.at(off++);
}

}  // namespace utils
#endif
