// The analytic test monitors of the reference (Experiments/TestMonitors/MEx*.h), against this package's
// MonitorFunction<D> interface.  Slot numbering follows main.cpp:836-863:
//   2D: 0 MEx0 identity, 1 MEx1 bump, 2 MEx2 anisotropic, 3 MEx3 rings, 4 MEx4 shock layer on x+y=1, 5 MEx5 spiral
//   3D: 0 MEx0, 1 MEx13D, 2 MEx23D, 3 MEx33D, 4 MEx0 (again), 5 MEx53D
#ifndef MMADMM_HOST_MONITOR_FUNCTIONS_H
#define MMADMM_HOST_MONITOR_FUNCTIONS_H
#include <cmath>
#include <limits>

#include "MonitorFunction.h"

namespace mmadmm_host {

template <int D> inline void setScaledIdentity(double *M, double s) {
    for (int k = 0; k < D * D; k++) M[k] = 0.0;
    for (int a = 0; a < D; a++) M[a * D + a] = 1.0 * s;
}
inline double fdStep() { return 2.0 * std::sqrt(std::numeric_limits<double>::epsilon()); }

template <int D> class MEx0 : public MonitorFunction<D> {  // MEx0.h
public:
    void operator()(const double *, double *M) override { setScaledIdentity<D>(M, 1.0); }
};

template <int D> class MEx1 : public MonitorFunction<D> {  // MEx1.h / MEx13D.h: 1 + mu1/(1 + mu2 |x - c|^2)
public:
    void operator()(const double *x, double *M) override {
        const double mu_1 = 20, mu_2 = 20;
        double sq = 0.0;
        for (int l = 0; l < D; l++) sq += (x[l] - 0.5) * (x[l] - 0.5);
        setScaledIdentity<D>(M, (1 + mu_1 / (1 + mu_2 * (sq))));
    }
};

class MEx2 : public MonitorFunction<2> {  // MEx2.h: anisotropic, aligned with x + y = 1
public:
    void operator()(const double *x, double *M) override {
        const double t = x[0] + x[1] - 1.0;
        const double lam1 = 1 + (1.0 / std::cosh(50 * t * t));
        const double lam2 = 1.0 / lam1;
        const double r = (1.0 / std::sqrt(2.0)) * 1.0;
        const double v[2] = {r, r}, vo[2] = {r, -r};
        for (int a = 0; a < 2; a++)
            for (int b = 0; b < 2; b++) M[a * 2 + b] = (lam1 * v[a]) * v[b] + (lam2 * vo[a]) * vo[b];
    }
};

template <int D> class MEx3 : public MonitorFunction<D> {  // MEx3.h / MEx23D.h / MEx33D.h: concentric rings
public:
    void operator()(const double *x, double *M) override {
        const double PI = 3.141592653589793238462643383;
        double r2 = 0.0;
        for (int l = 0; l < D; l++) r2 += std::pow(x[l] - 0.5, 2);
        const double a = 0.01 / (2.0 + std::cos(8.0 * PI * std::sqrt(r2)));
        setScaledIdentity<D>(M, (D == 2) ? std::sqrt(a) : std::pow(a, 1.0 / 2.0));
    }
};

class MEx4 : public MonitorFunction<2> {  // MEx4.h: arc-length monitor of a shock layer on x + y = 1
    static double u(double s) { return 1.0 / (1.0 + std::exp(s / (2.0 * 0.01))); }
public:
    void operator()(const double *x, double *M) override {
        const double h = fdStep();
        const double g0 = (u(x[0] + h + x[1] - 1) - u(x[0] - h + x[1] - 1)) / (2.0 * h);
        const double g1 = (u(x[0] + x[1] + h - 1) - u(x[0] + x[1] - h - 1)) / (2.0 * h);
        const double nrm = std::sqrt(g0 * g0 + g1 * g1);
        setScaledIdentity<2>(M, std::pow(1 + std::pow(nrm, 2.0), 1.0 / 4.0));
    }
};

class MEx5 : public MonitorFunction<2> {  // MEx5.h: spiral front
    static double u(double x, double y) {
        const double r = std::sqrt(std::pow(x - 0.7, 2.0) + std::pow(y - 0.5, 2.0));
        const double theta = std::atan((y - 0.5) / (x - 0.7));
        return 1.0 + 9.0 / (1.0 + 100.0 * r * r * std::pow(std::cos(theta - 20.0 * r * r), 2.0));
    }
public:
    void operator()(const double *x, double *M) override {
        const double h = fdStep();
        const double g0 = (u(x[0] + h, x[1]) - u(x[0] - h, x[1])) / (2.0 * h);
        const double g1 = (u(x[0], x[1] + h) - u(x[0], x[1] - h)) / (2.0 * h);
        const double nrm = std::sqrt(g0 * g0 + g1 * g1);
        setScaledIdentity<2>(M, std::pow(1 + std::pow(nrm, 2.0), 1.0 / 4.0));
    }
};

class MEx53D : public MonitorFunction<3> {  // MEx53D.h.  As in the reference, d/dy never reaches the norm
    static double u(double x, double y, double z) {  // (grad(1) is assigned twice there, MEx53D.h:20-22).
        const double r = std::sqrt(std::pow(x - 0.7, 2.0) + std::pow(y - 0.5, 2.0) + std::pow(z - 0.5, 2));
        const double theta = std::atan((y - 0.5) / (x - 0.7));
        const double psi = std::atan((z - 0.5) / (x - 0.7));
        return 1.0 + 9.0 / (1.0 + 100.0 * r * r * std::pow(std::cos(theta + psi - 20.0 * r * r), 2.0));
    }
public:
    void operator()(const double *x, double *M) override {
        const double h = fdStep();
        const double g0 = (u(x[0] + h, x[1], x[2]) - u(x[0] - h, x[1], x[2])) / (2.0 * h);
        const double g1 = (u(x[0], x[1], x[2] + h) - u(x[0], x[1], x[2] - h)) / (2.0 * h);
        const double nrm = std::sqrt(g0 * g0 + g1 * g1);
        setScaledIdentity<3>(M, std::pow(1 + std::pow(nrm, 2.0), 1.0 / 4.0));
    }
};

inline MonitorFunction<2> *makeMonitor2(int kind) {
    switch (kind) {
    case 0: return new MEx0<2>();
    case 1: return new MEx1<2>();
    case 2: return new MEx2();
    case 3: return new MEx3<2>();
    case 4: return new MEx4();
    case 5: return new MEx5();
    }
    return nullptr;
}
inline MonitorFunction<3> *makeMonitor3(int kind) {
    switch (kind) {
    case 0: case 4: return new MEx0<3>();
    case 1: return new MEx1<3>();
    case 2: case 3: return new MEx3<3>();
    case 5: return new MEx53D();
    }
    return nullptr;
}

}  // namespace mmadmm_host
#endif
