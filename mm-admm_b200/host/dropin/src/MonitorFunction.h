// MonitorFunction<D>: the user's metric tensor M(x).  Same interface as the reference (src/MonitorFunction.h:10-20),
// so the monitor plug-ins in Experiments/TestMonitors/MEx*.h compile against it unchanged.  The tensor is
// evaluated on the host exactly once per vertex (evaluateAtVertices, src/MonitorFunction.cpp:15-32); from then on
// the GPU works from the interpolation grid built out of that table (mmadmm_create).
#ifndef MONITOR_FUNCTION_H
#define MONITOR_FUNCTION_H

#include <Eigen/Dense>
#include <vector>

using namespace std;   // the reference header does this and main.cpp relies on it

template <int D>
class MonitorFunction {
public:
    virtual void operator()(Eigen::Vector<double,D> &x, Eigen::Matrix<double,D,D> &M) = 0;
    virtual ~MonitorFunction() {}
    // Row v of M receives the row-major flattened tensor at vertex v.  F is unused, as in the reference.
    void evaluateAtVertices(Eigen::MatrixXd &X, Eigen::MatrixXi &F, Eigen::MatrixXd &M) {
        (void)F;
        Eigen::Matrix<double,D,D> mon;
        Eigen::Vector<double,D> x;
        for (int v = 0; v < X.rows(); v++) {
            mon.setZero();
            x = X.row(v);
            (*this)(x, mon);
            for (int k = 0; k < D*D; k++) M(v, k) = mon(k / D, k % D);
        }
    }
    // Flat row-major table (P x D*D) for the C ABI (monVertex of mmadmm_create).
    void evaluateAtVertices(Eigen::MatrixXd &X, std::vector<double> &table) {
        table.assign((size_t)X.rows() * D * D, 0.0);
        Eigen::Matrix<double,D,D> mon;
        Eigen::Vector<double,D> x;
        for (int v = 0; v < X.rows(); v++) {
            mon.setZero();
            x = X.row(v);
            (*this)(x, mon);
            for (int k = 0; k < D*D; k++) table[(size_t)v * D * D + k] = mon(k / D, k % D);
        }
    }
    void evaluateAtPoint(Eigen::MatrixXd &X, Eigen::MatrixXi &F, Eigen::MatrixXd &M) { evaluateAtVertices(X, F, M); }
};

#endif
