// Mesh<D>: host-side mirror of the reference class (src/Mesh.h:16-127) over the C ABI of libmmadmm_b200.so.
//
// Same constructors, same public data the driver touches (Vc, Vp, F, boundaryMask as NON-OWNING pointers to the
// caller's objects, src/Mesh.cpp:399-414), same side effects: the caller's F is re-oriented in place at
// construction (reOrientElements, src/Mesh.cpp:243-260) and the caller's Vp is overwritten after every step
// (updateAfterStep, src/Mesh.cpp:1030-1035).  All mesh state used by the integrator lives on the GPU inside an
// opaque mmadmm_ctx; there is no CPU implementation behind this class -- without a CUDA device the constructor
// reports the error and aborts, which is what the reference does on a failed assert.
//
// What the reference's Mesh does on the host and this one does not: the D/W/M sparse matrices, face lists and
// simplex maps are never formed (the ADMM system matrix is diagonal, incidence replaces D).
// Method 1 (explicit Euler) runs on the device through mmadmm_euler_step, Method 2 (backward Euler, the reference's
// LASolver path) through mmadmm_backward_euler_step (FD Jacobian kernel + ILU(0)/BiCGSTAB kernels).
#ifndef MESH_H
#define MESH_H

#include <Eigen/Dense>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <map>
#include <set>
#include <vector>

#include "AdaptationFunctional.h"
#include "MonitorFunction.h"
#include "NodeType.h"
#include "mmadmm.h"

using namespace std;

template <int D=-1>
class Mesh {
public:
    double m;
    int integrationMode;
    int gradUse;

    Mesh(Eigen::MatrixXd &Xc, Eigen::MatrixXd &Xp, Eigen::MatrixXi &F, vector<NodeType> &boundaryMask,
            MonitorFunction<D> *M, int numThreads, double rho, double w, double tau, int integrationMode, bool gradUse) {
        this->integrationMode = integrationMode;
        this->compMesh = true;
        meshInit(Xc, Xp, F, boundaryMask, M, numThreads, rho, w, tau, gradUse);
    }
    Mesh(Eigen::MatrixXd &Xp, Eigen::MatrixXi &F, vector<NodeType> &boundaryMask,
            MonitorFunction<D> *M, int numThreads, double rho, double w, double tau, int integrationMode, bool gradUse) {
        this->integrationMode = integrationMode;
        this->compMesh = false;
        meshInit(Xp, Xp, F, boundaryMask, M, numThreads, rho, w, tau, gradUse);
    }
    ~Mesh() {
        if (vpRegistered) mmadmm_host_unregister(Vp->data());
        if (ctx) mmadmm_destroy(ctx);
        delete monitorEvals;
    }
    Mesh(const Mesh &) = delete;
    Mesh &operator=(const Mesh &) = delete;

    // src/Mesh.cpp:384-472.  The JSON "w" argument is ignored like in the reference (w = 0.5*sqrt(rho), :451);
    // numThreads has no meaning on the GPU and is kept for signature compatibility.
    void meshInit(Eigen::MatrixXd &Xc, Eigen::MatrixXd &Xp, Eigen::MatrixXi &F, vector<NodeType> &boundaryMask,
            MonitorFunction<D> *Mon, int numThreads, double rho, double w, double tau, bool gradUse) {
        (void)numThreads; (void)w;
        this->gradUse = gradUse;
        this->Vc = compMesh ? &Xc : nullptr;
        this->Vp = &Xp;
        this->F = &F;
        this->boundaryMask = &boundaryMask;
        this->Mon = Mon;
        this->tau = tau;
        this->m = tau;
        this->rho = rho;
        this->w = 0.5 * sqrt(rho);
        this->nPnts = Xp.rows();
        // integrationMode: 0 ADMM (step), 1 explicit Euler (eulerStep), anything else backward Euler (main.cpp:182-191);
        // the device context serves all three, the Method 2 state is built on the first backwardsEulerStep
        const int P = Xp.rows(), E = F.rows();
        if ((int)boundaryMask.size() < P) {
            cerr << "mm-admm_b200: boundary mask has " << boundaryMask.size() << " entries for " << P << " points" << endl;
            exit(1);
        }
        // the monitor tensor at the vertices: the only host-side evaluation of the user's functor
        monitorEvals = new Eigen::MatrixXd(P, D*D);
        Mon->evaluateAtVertices(Xp, F, *monitorEvals);

        // C ABI layouts: row-major interleaved
        vector<double> vp((size_t)P * D), vc, mon((size_t)P * D * D);
        vector<int> f((size_t)E * (D + 1)), mask((size_t)P);
        for (int p = 0; p < P; p++) {
            for (int l = 0; l < D; l++) vp[(size_t)p * D + l] = Xp(p, l);
            for (int k = 0; k < D*D; k++) mon[(size_t)p * D * D + k] = (*monitorEvals)(p, k);
            mask[p] = (int)boundaryMask[p];
        }
        if (compMesh) {
            vc.resize((size_t)P * D);
            for (int p = 0; p < P; p++) for (int l = 0; l < D; l++) vc[(size_t)p * D + l] = Xc(p, l);
        }
        for (int e = 0; e < E; e++) for (int v = 0; v <= D; v++) f[(size_t)e * (D + 1) + v] = F(e, v);

        const char *dev = getenv("MMADMM_DEVICE");
        check(mmadmm_create(D, P, E, vp.data(), compMesh ? vc.data() : nullptr, f.data(), mask.data(), mon.data(), rho, tau,
                            /*dt, set by MeshIntegrator*/ 0.0, gradUse ? 1 : 0, dev ? atoi(dev) : 0, &ctx), "mmadmm_create");
        // the reference permutes the caller's F (positive orientation)
        check(mmadmm_get_F(ctx, f.data()), "mmadmm_get_F");
        for (int e = 0; e < E; e++) for (int v = 0; v <= D; v++) F(e, v) = f[(size_t)e * (D + 1) + v];
        // page-lock the caller's Vp: updateAfterStep then copies straight into it at PCIe speed
        vpRegistered = mmadmm_host_register(Xp.data(), sizeof(double) * (unsigned long)P * D) == MMADMM_OK;
    }

    // evalMonitorOnGrid at one point (src/MeshInterpolator.cpp:286-342): the GPU-resident interpolation
    void evalMonitorAtPoint(Eigen::Vector<double,D> &x, Eigen::Matrix<double,D,D> &mVal) {
        double pt[D], out[D*D];
        for (int l = 0; l < D; l++) pt[l] = x(l);
        check(mmadmm_eval_monitor(ctx, pt, 1, out), "mmadmm_eval_monitor");
        for (int k = 0; k < D*D; k++) mVal(k / D, k % D) = out[k];
    }

    void setUp() {}   // src/Mesh.cpp:1006-1014: body commented out in the reference
    int getNPnts() { return nPnts; }

    // src/Mesh.cpp:996-1004: interleaved copy of Vp
    void copyX(Eigen::VectorXd &tar) {
        for (int p = 0; p < Vp->rows(); p++) for (int l = 0; l < D; l++) tar(p * D + l) = (*Vp)(p, l);
    }
    // src/Mesh.cpp:496-530: the functional over the current physical mesh (the argument is ignored there too)
    double computeEnergy(Eigen::VectorXd &x) {
        (void)x;
        double I = 0;
        check(mmadmm_energy(ctx, &I), "mmadmm_energy");
        return I;
    }
    // src/Mesh.cpp:1016-1036: Vp <- x.  x lives on the device; the arguments are kept for the signature.
    void updateAfterStep(double dt, Eigen::VectorXd &xPrev, Eigen::VectorXd &x) {
        (void)dt; (void)xPrev; (void)x;
        syncVp();
    }
    // Vp is column-major (Eigen's default): the transposition runs on the device
    void syncVp() { check(mmadmm_get_x_colmajor(ctx, Vp->data()), "mmadmm_get_x_colmajor"); }

    // result writers, formats of src/Mesh.cpp:1049-1095
    void outputSimplices(const char *fname) {
        std::ofstream out(fname);
        for (int i = 0; i < F->rows(); i++) {
            for (int j = 0; j < D; j++) out << (*F)(i, j) << ", ";
            out << (*F)(i, D) << endl;
        }
    }
    void outputPoints(const char *fname) {
        std::ofstream out(fname);
        for (int i = 0; i < Vp->rows(); i++) {
            for (int j = 0; j < D-1; j++) out << (*Vp)(i, j) << ", ";
            out << (*Vp)(i, D-1) << endl;
        }
    }
    void outputBoundaryNodes(const char *fname) {
        std::ofstream out(fname);
        for (int i = 0; i < Vp->rows(); i++) {
            if (boundaryMask->at(i) == INTERIOR) continue;
            for (int j = 0; j < D-1; j++) out << (*Vp)(i, j) << ", ";
            out << (*Vp)(i, D-1) << endl;
        }
    }

    // status code of the C ABI -> the reference's behaviour on a failed assert / exit(1): message + abort
    void check(int rc, const char *what) {
        if (rc == MMADMM_OK) return;
        static const char *names[] = {"OK", "EINVAL", "ECUDA", "EINVERTED", "ESINGULAR", "ENCCL", "ENODEVICE"};
        cerr << "mm-admm_b200: " << what << " failed: " << (rc >= 0 && rc <= 6 ? names[rc] : "?")
             << (ctx && mmadmm_last_error(ctx) ? string(" -- ") + mmadmm_last_error(ctx) : string()) << endl;
        if (rc == MMADMM_ENODEVICE) cerr << "mm-admm_b200: no CUDA device; this build has no CPU path" << endl;
        abort();
    }

    mmadmm_ctx *ctx = nullptr;
    int nSteps = 0;
    Eigen::MatrixXd *Vc = nullptr;
    Eigen::MatrixXd *Vp = nullptr;
    Eigen::MatrixXi *F = nullptr;
    vector<NodeType> *boundaryMask = nullptr;
    Eigen::MatrixXd *monitorEvals = nullptr;
    MonitorFunction<D> *Mon = nullptr;
    bool stepTaken = false;
    bool hessComputed = false;
    bool compMesh = true;
    double tau = 0;
    double rho = 0;
    double w = 0;
    int nPnts = 0;
    double Ihcur = 0, Ihprev = 0;

private:
    bool vpRegistered = false;
};

#endif
