// MeshIntegrator<D>: host-side mirror of the reference class (src/MeshIntegrator.h:12-50).  step() is one call into
// the CUDA library (mmadmm_step: predictor, x-updates, per-simplex BFGS prox sweeps, dual update, residual test --
// src/MeshIntegrator.cpp:101-191), followed by the reference's updateAfterStep (the caller's Vp is overwritten).
#ifndef SOLVER_H
#define SOLVER_H

#include <Eigen/Dense>
#include <cmath>
#include <ctime>
#include <string>
#include <unordered_map>

#include "Mesh.h"

using namespace std;

template <int D>
class MeshIntegrator {
public:
    MeshIntegrator(double dt, Mesh<D> &a) {
        this->a = &a;
        this->dt = dt;
        this->dtPrev = dt;
        stepsTaken = 0;
        proxTime = cgTime = multTime = predTime = 0;
        a.broadcast(Mesh<D>::OP_SETDT, 0, dt);
        a.check(mmadmm_set_dt(a.ctx, dt), "mmadmm_set_dt");   // diagonal of t = M + dt^2 (WD)^T(WD), :45-51
        a.check(mmadmm_set_option(a.ctx, "timing", 1), "mmadmm_set_option");
    }
    ~MeshIntegrator() {}

    // One implicit MMPDE step by ADMM.  Returns the energy at entry to the first prox sweep, like the reference.
    double step(int nIters, double tol) {
        double Ih = 0, primal = 0, dual = 0;
        int iters = 0;
        a->broadcast(Mesh<D>::OP_STEP, nIters, tol);
        a->check(mmadmm_step(a->ctx, nIters, tol, &Ih, &iters, &primal, &dual), "mmadmm_step");
        lastIters = iters; lastPrimal = primal; lastDual = dual;
        a->stepTaken = true;
        a->hessComputed = true;
        a->updateAfterStep(dt, xPrevUnused, xUnused);
        // the reference accumulates CPU clock() seconds of the prox loop and of predictX; here: device time of
        // the prox kernels and of the predictor kernels
        double ms[5]; long long cnt[5];
        if (mmadmm_get_timing(a->ctx, ms, cnt, 1) == MMADMM_OK) { predTime += ms[0] * 1e-3; proxTime += (ms[1] + ms[4]) * 1e-3; cgTime += ms[2] * 1e-3; }
        energyCur = Ih;
        stepsTaken++;
        a->nSteps = stepsTaken;
        return Ih;
    }
    // Method 1 (src/MeshIntegrator.cpp:87-95).  As in the reference there is no updateAfterStep on this path: Vp follows at done().
    double eulerStep(double tol) {
        (void)tol;
        double Ih = 0;
        a->broadcast(Mesh<D>::OP_EULER);
        a->check(mmadmm_euler_step(a->ctx, &Ih), "mmadmm_euler_step");
        return Ih;
    }
    // Method 2 (src/MeshIntegrator.cpp:68-76 -> Mesh::backwardsEulerStep, src/Mesh.cpp:1263-1341).  No updateAfterStep here either.
    double backwardsEulerStep(double dt, double tol) {
        double Ih = 0;
        a->check(mmadmm_backward_euler_step(a->ctx, dt, tol, &Ih, nullptr, nullptr), "mmadmm_backward_euler_step");
        return Ih;
    }
    double getEnergy() { return a->computeEnergy(xUnused); }
    void done() { a->syncVp(); }

    // x / z writers (src/MeshIntegrator.cpp:217-243); x is read back from the device
    void outputX(const char *fname) {
        std::vector<double> x((size_t)a->getNPnts() * D);
        a->broadcast(Mesh<D>::OP_GETX);
        a->check(mmadmm_get_x(a->ctx, x.data()), "mmadmm_get_x");
        std::ofstream out(fname);
        for (int i = 0; i < a->getNPnts(); i++) {
            for (int j = 0; j < D-1; j++) out << x[(size_t)i*D+j] << ", ";
            out << x[(size_t)i*D+(D-1)] << endl;
        }
    }
    void outputZ(const char *fname) {
        const int n = D*(D+1), E = a->F->rows();
        if (a->distributed) { cerr << "mm-admm_b200: outputZ is not available on a multi-GPU run (element state stays on its rank)" << endl; return; }
        std::vector<double> z((size_t)E * n);
        a->check(mmadmm_get_state(a->ctx, MMADMM_STATE_Z, z.data()), "mmadmm_get_state");
        std::ofstream out(fname);
        for (size_t i = 0; i < z.size() / D; i++) {
            for (int j = 0; j < D-1; j++) out << z[i*D+j] << ", ";
            out << z[i*D+(D-1)] << endl;
        }
    }

    double proxTime;
    double multTime;
    double cgTime;
    double predTime;

    Mesh<D> *a;
    double dt;
    double dtPrev;
    double energyCur = INFINITY;
    int stepsTaken;
    int lastIters = 0;
    double lastPrimal = 0, lastDual = 0;

private:
    Eigen::VectorXd xUnused, xPrevUnused;
};

#endif
