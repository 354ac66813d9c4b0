// MonitorFunction<D>: the user-supplied metric tensor M(x) that drives the mesh adaptation.
// Mirrors the reference interface src/MonitorFunction.h:10-20 (virtual operator()(x, M), evaluateAtVertices)
// without Eigen: x is D doubles, M is D x D row-major.  The tensor is only ever evaluated on the host, once,
// at the mesh vertices (src/MonitorFunction.cpp:15-32); the GPU then works from the interpolation grid.
#ifndef MMADMM_HOST_MONITOR_FUNCTION_H
#define MMADMM_HOST_MONITOR_FUNCTION_H
#include <cstddef>

template <int D> class MonitorFunction {
public:
    virtual void operator()(const double *x, double *M) = 0;
    virtual ~MonitorFunction() {}
    // Row p of `out` receives the row-major flattened tensor at vertex p (X: P x D row-major).
    void evaluateAtVertices(const double *X, int P, double *out) {
        for (int p = 0; p < P; p++) {
            double *M = out + (size_t)p * D * D;
            for (int k = 0; k < D * D; k++) M[k] = 0.0;
            (*this)(X + (size_t)p * D, M);
        }
    }
};
#endif
