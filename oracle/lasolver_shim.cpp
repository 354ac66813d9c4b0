// Test infrastructure: a C entry point into the reference's own sparse solver (lib/LASolver), compiled together with
// the reference's sources where they lie under /root/reference into oracle/_ref/liblasolver_ref.so (oracle/Makefile).
// Only tests/ may load it.  It drives LASolver the way the reference's Method 2 does (src/Mesh.cpp:266-304, 1284-1323:
// natural ordering, level-of-fill ILU, no scaling) with the accelerator selectable (iaccel = -1: conjugate gradients).
#include <cstring>
#include <new>

#include "SparseItObj.h"

using namespace SparseItObj;

extern "C" int lasolver_solve(int n, const int *ia, const int *ja, const double *a, const double *b, int iaccel, int level,
                              double resid_reduc, int nitmax, double *x, int *nitr) {
    try {
        MatrixIter m(n, ia, ja);
        for (int k = 0; k < ia[n]; k++) m.aValue(k) = a[k];
        for (int i = 0; i < n; i++) m.bValue(i) = b[i];
        double *tol = new double[n];
        for (int i = 0; i < n; i++) tol[i] = 0.0;       // only the residual-reduction criterion stops the iteration
        m.set_toler(tol);
        ParamIter p;
        p.order = 0; p.level = level; p.drop_ilu = 0; p.iscal = 0; p.nitmax = nitmax; p.ipiv = 0;
        p.resid_reduc = resid_reduc; p.info = 0; p.drop_tol = 1e-3; p.new_rhat = 0; p.iaccel = iaccel; p.north = 10;
        m.sfac(p);
        int it = 0;
        m.solve(p, x, it);
        *nitr = it;
        delete[] tol;
        return 0;
    } catch (...) {
        return 1;
    }
}
