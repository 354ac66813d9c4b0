// HuangFunctional<D>: the reference's subclass (src/HuangFunctional.h:11-39) with its virtual pieces of Huang's functional
//   G(J, det J, M^-1) = theta sqrt(det M) tr(J M^-1 J^T)^{dp/2} + (1 - 2 theta) d^{dp/2} sqrt(det M) (det J / sqrt(det M))^p
// restated on the host from src/HuangFunctional.cpp:35-90, quirks kept: G, dGdJ and dGdM hard-code p = 3/2, theta = 1/3
// whatever the constructor was given (the reference constructs it with theta = p = 0, src/Mesh.cpp:468-470); dGddet uses
// theta = 1/2 -- so it returns 0 -- and the MEMBER p; dGdX is zero.  Nothing on the ADMM path calls these: the
// device function block_grad<D> carries the arithmetic that blockGrad inlines.  tests/dropin_host_check.cpp checks them
// against a finite-difference derivative of G.
#ifndef HUANG_FUNCTIONAL_H
#define HUANG_FUNCTIONAL_H
#include <cmath>
#include "AdaptationFunctional.h"

template <int D>
class HuangFunctional : public AdaptationFunctional<D> {
public:
    explicit HuangFunctional(mmadmm_ctx *ctx_, double w_ = 0.0, double theta_ = 0.0, double p_ = 0.0)
        : AdaptationFunctional<D>(ctx_), theta(theta_), p(p_), w(w_) {}
    // public forwarding, for callers that hold a HuangFunctional (the reference keeps them protected)
    double evalG(Eigen::Matrix<double,D,D> &J, double det, Eigen::Matrix<double,D,D> &Minv, Eigen::Vector<double,D> &x) { return G(J, det, Minv, x); }
    void evaldGdJ(Eigen::Matrix<double,D,D> &J, double det, Eigen::Matrix<double,D,D> &Minv, Eigen::Vector<double,D> &x,
                  Eigen::Matrix<double,D,D> &out) { dGdJ(J, det, Minv, x, out); }
    double evaldGddet(Eigen::Matrix<double,D,D> &J, double det, Eigen::Matrix<double,D,D> &Minv, Eigen::Vector<double,D> &x) { return dGddet(J, det, Minv, x); }
    void evaldGdM(Eigen::Matrix<double,D,D> &J, double det, Eigen::Matrix<double,D,D> &Minv, Eigen::Vector<double,D> &x,
                  Eigen::Matrix<double,D,D> &out) { dGdM(J, det, Minv, x, out); }
    void evaldGdX(Eigen::Matrix<double,D,D> &J, double det, Eigen::Matrix<double,D,D> &Minv, Eigen::Vector<double,D> &x,
                  Eigen::Vector<double,D> &out) { dGdX(J, det, Minv, x, out); }
protected:
    double theta;
    double p;
    double w;
    static double trJMJt(Eigen::Matrix<double,D,D> &J, Eigen::Matrix<double,D,D> &Minv) {
        Eigen::Matrix<double,D,D> Jt(J.transpose());
        Eigen::Matrix<double,D,D> A(J * Minv * Jt);
        double t = 0;
        for (int i = 0; i < D; i++) t += A(i, i);
        return t;
    }
    double G(Eigen::Matrix<double,D,D> &J, double detJ, Eigen::Matrix<double,D,D> &Minv, Eigen::Vector<double,D> &x) override {
        (void)x;
        double d = (double)D, p = 1.5, theta = 1.0 / 3.0;
        double detM = sqrt(1.0 / Minv.determinant());
        return theta * detM * pow(trJMJt(J, Minv), d * p / 2.0) + (1.0 - 2.0 * theta) * pow(d, d * p / 2.0) * detM * pow(detJ / detM, p);
    }
    void dGdJ(Eigen::Matrix<double,D,D> &J, double detJ, Eigen::Matrix<double,D,D> &Minv, Eigen::Vector<double,D> &x,
              Eigen::Matrix<double,D,D> &out) override {
        (void)x; (void)detJ;
        double d = (double)D, p = 1.5, theta = 1.0 / 3.0;
        double detM = sqrt(1.0 / Minv.determinant());
        Eigen::Matrix<double,D,D> Jt(J.transpose());
        out = Minv * Jt;
        out *= d * p * theta * detM * pow(trJMJt(J, Minv), d * p / 2.0 - 1);
    }
    double dGddet(Eigen::Matrix<double,D,D> &J, double detJ, Eigen::Matrix<double,D,D> &Minv, Eigen::Vector<double,D> &x) override {
        (void)J; (void)x;
        double d = (double)D;
        double theta = 1.0 / 2.0;     // sic (src/HuangFunctional.cpp:64): the factor (1 - 2 theta) vanishes
        double detM = sqrt(1.0 / Minv.determinant());
        return p * (1.0 - 2.0 * theta) * pow(d, (d * p) / 2.0) * pow(detM, 1.0 - p) * pow(detJ, p - 1);
    }
    void dGdM(Eigen::Matrix<double,D,D> &J, double detJ, Eigen::Matrix<double,D,D> &Minv, Eigen::Vector<double,D> &x,
              Eigen::Matrix<double,D,D> &out) override {
        (void)x;
        double d = (double)D, p = 1.5, theta = 1.0 / 3.0;
        double detM = sqrt(1.0 / Minv.determinant());
        Eigen::Matrix<double,D,D> Jt(J.transpose()), Mt(Minv.transpose());
        double tr = trJMJt(J, Minv);
        Eigen::Matrix<double,D,D> A(Mt * Jt * J * Minv);
        A *= -0.5 * theta * d * p * detM * pow(tr, d * p / 2.0 - 1);
        Eigen::Matrix<double,D,D> B(Minv);
        B *= 0.5 * theta * detM * pow(tr, d * p / 2.0) + ((0.5 - theta) * (1.0 - p) * pow(d, d * p / 2.0)) * pow(detM, 1 - p) * pow(detJ, p);
        out = A + B;
    }
    void dGdX(Eigen::Matrix<double,D,D> &J, double detJ, Eigen::Matrix<double,D,D> &Minv, Eigen::Vector<double,D> &x,
              Eigen::Vector<double,D> &out) override {
        (void)J; (void)detJ; (void)Minv; (void)x;
        out.setZero();
    }
};

#endif
