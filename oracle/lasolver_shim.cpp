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

// A persistent LASolver matrix driven the way the reference's Method 2 (backward Euler) drives it: the structure through
// MatrixStruc::set_entry in the caller's order + pack() (src/Mesh.cpp:305-340), ParamIter exactly as buildMatrix sets it
// (src/Mesh.cpp:266-304: natural order, ILU level 0, no scaling, CGSTAB, resid_reduc 1e-6, nitmax 10000), sfac ONCE
// (src/Mesh.cpp:1284-1286), then per Newton iteration aValue(k) = ..., bValue(i) = ..., set_toler(0), solve
// (src/Mesh.cpp:1311-1323; every solve re-factors numerically because aValue() marks the factors stale).
struct LasHandle {
    MatrixIter *m = nullptr;
    ParamIter p;
    int n = 0;
};

extern "C" void *las_open(int n, long long nent, const int *rows, const int *cols) {
    try {
        MatrixStruc s(n, 0);
        for (long long k = 0; k < nent; k++) s.set_entry(rows[k], cols[k]);
        s.pack();
        LasHandle *h = new LasHandle;
        h->m = new MatrixIter(s);
        h->n = n;
        ParamIter &p = h->p;
        p.order = 0; p.level = 0; p.drop_ilu = 0; p.iscal = 0; p.nitmax = 10000; p.ipiv = 0;
        p.resid_reduc = 1.e-6; p.info = 0; p.drop_tol = 1.e-3; p.new_rhat = 0; p.iaccel = 0; p.north = 10;
        return h;
    } catch (...) {
        return nullptr;
    }
}
extern "C" int las_nnz(void *hv) { LasHandle *h = (LasHandle *)hv; return h->m->rowEndPlusOne(h->n - 1); }
extern "C" void las_pattern(void *hv, int *ia, int *ja) {
    LasHandle *h = (LasHandle *)hv;
    for (int r = 0; r < h->n; r++) {
        ia[r] = h->m->rowBegin(r);
        for (int k = h->m->rowBegin(r); k < h->m->rowEndPlusOne(r); k++) ja[k] = h->m->getColIndex(k);
    }
    ia[h->n] = h->m->rowEndPlusOne(h->n - 1);
}
extern "C" int las_sfac(void *hv) {
    LasHandle *h = (LasHandle *)hv;
    try { h->m->sfac(h->p); return 0; } catch (...) { return 1; }
}
extern "C" int las_solve(void *hv, const double *a, const double *b, double *x, int *nitr) {
    LasHandle *h = (LasHandle *)hv;
    try {
        const int nnz = las_nnz(hv);
        if (a) for (int k = 0; k < nnz; k++) h->m->aValue(k) = a[k];
        for (int i = 0; i < h->n; i++) h->m->bValue(i) = b[i];
        double *tol = new double[h->n];
        for (int i = 0; i < h->n; i++) tol[i] = 0.0;
        h->m->set_toler(tol);
        delete[] tol;
        int it = 0;
        h->m->solve(h->p, x, it);
        *nitr = it;
        return 0;
    } catch (...) {
        return 1;
    }
}
extern "C" void las_close(void *hv) {
    LasHandle *h = (LasHandle *)hv;
    if (h) { delete h->m; delete h; }
}
