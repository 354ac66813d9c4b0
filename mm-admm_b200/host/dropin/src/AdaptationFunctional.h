// AdaptationFunctional<D> / HuangFunctional<D>: interface types of the reference (src/AdaptationFunctional.h:8-54,
// src/HuangFunctional.h:11-39) kept for source compatibility, virtual G / dGdJ / dGddet / dGdM / dGdX included (HuangFunctional.h
// holds their Huang-functional bodies, restated from src/HuangFunctional.cpp:35-90).  In the reference those virtuals are
// never called -- blockGrad inlines Huang's functional with p = 3/2, theta = 1/3 (src/AdaptationFunctional.cpp:210-211)
// -- and the same arithmetic is the device function block_grad<D> in mm-admm_b200/csrc/mmadmm_device.cuh.
// blockGrad here evaluates element zId of a live Mesh through the C ABI (mmadmm_block_grad_at).
#ifndef ADAPTATION_FUNCTIONAL_H
#define ADAPTATION_FUNCTIONAL_H

#include <Eigen/Dense>
#include "MonitorFunction.h"
#include "mmadmm.h"

template <int D=-1>
class AdaptationFunctional {
public:
    bool boundaryNode = false;
    int N = 0;
    bool compMesh = false;
    explicit AdaptationFunctional(mmadmm_ctx *ctx_) : ctx(ctx_) {}
    virtual ~AdaptationFunctional() {}
    // energy of simplex zId at coordinates z (+ optional ADMM penalty w^2/2 |z - DXpU_i|^2 when regularize), its
    // gradient, and Igt = the un-regularised energy.  xi (computational coordinates) is taken from the context.
    double blockGrad(int zId, Eigen::Vector<double, D*(D+1)> &z, Eigen::Vector<double, D*(D+1)> &dxpu,
                     Eigen::Vector<double, D*(D+1)> &grad, bool computeGrad, bool regularize, double &Igt) {
        double e = 0, g[D*(D+1)], zz[D*(D+1)], dd[D*(D+1)];
        for (int k = 0; k < D*(D+1); k++) { zz[k] = z(k); dd[k] = dxpu(k); }
        int rc = mmadmm_block_grad_at(ctx, zId, 1, zz, regularize ? dd : nullptr, computeGrad, regularize, 0, &e, g, &Igt);
        if (rc != MMADMM_OK) return NAN;
        if (computeGrad) for (int k = 0; k < D*(D+1); k++) grad(k) = g[k];
        return e;
    }
protected:
    mmadmm_ctx *ctx;
    // the functional's pieces as the reference declares them (src/AdaptationFunctional.h:19-31); M is the INVERSE monitor
    // tensor in every implementation the reference ships
    virtual double G(Eigen::Matrix<double,D,D> &J, double det, Eigen::Matrix<double,D,D> &M, Eigen::Vector<double,D> &x) = 0;
    virtual void dGdJ(Eigen::Matrix<double,D,D> &J, double det, Eigen::Matrix<double,D,D> &M, Eigen::Vector<double,D> &x,
                      Eigen::Matrix<double,D,D> &out) = 0;
    virtual double dGddet(Eigen::Matrix<double,D,D> &J, double det, Eigen::Matrix<double,D,D> &M, Eigen::Vector<double,D> &x) = 0;
    virtual void dGdM(Eigen::Matrix<double,D,D> &J, double det, Eigen::Matrix<double,D,D> &M, Eigen::Vector<double,D> &x,
                      Eigen::Matrix<double,D,D> &out) = 0;
    virtual void dGdX(Eigen::Matrix<double,D,D> &J, double det, Eigen::Matrix<double,D,D> &M, Eigen::Vector<double,D> &x,
                      Eigen::Vector<double,D> &out) = 0;
};

#include "HuangFunctional.h"

#endif
