// Mesh<D>: host-side mirror of the reference class (src/Mesh.h:16-127) over the C ABI of libmmadmm_b200.so.
//
// Same constructors, same public data the driver touches (Vc, Vp, F, boundaryMask as NON-OWNING pointers to the
// caller's objects, src/Mesh.cpp:399-414), same side effects: the caller's F is re-oriented in place at
// construction (reOrientElements, src/Mesh.cpp:243-260) and the caller's Vp is overwritten after every step
// (updateAfterStep, src/Mesh.cpp:1030-1035).  All mesh state used by the integrator lives on the GPU inside an
// opaque mmadmm_ctx; there is no CPU implementation behind this class -- without a CUDA device the constructor
// reports the error and aborts, which is what the reference does on a failed assert.
//
// Multi-GPU (the reference's only parallelism knob is numThreads -> omp_set_num_threads, src/Mesh.cpp:429,437):
// with MMADMM_GPUS=N in the environment (N >= 2; "threads" = take the driver's numThreads argument) the constructor forks N-1
// worker processes BEFORE the first CUDA call -- one process per GPU, as the library's element-block partition wants it.
// Every process builds its rank of the distributed context (mmadmm_create_dist) from the same global mesh; the workers then
// serve the collective calls (step / energy / get_x / ...) that this process, rank 0, issues through pipes, and never return
// into the driver.  main.cpp runs unchanged and unaware: one process writes the result files.  done() assembles the global
// mesh; between steps only rank 0's vertices are refreshed in the caller's Vp.  Method 2 is single-GPU only.
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
#include <cstring>
#include <string>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

#include "AdaptationFunctional.h"
#include "MonitorFunction.h"
#include "NodeType.h"
#include "MeshIO.h"
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
        if (ctx) mmadmm_writeback_fence(ctx);
        if (vpRegistered) mmadmm_host_unregister(Vp->data());
        broadcast(OP_EXIT);
        if (ctx) mmadmm_destroy(ctx);
        for (size_t r = 0; r < workers.size(); r++) {
            close(workers[r].fd);
            int st = 0;
            waitpid(workers[r].pid, &st, 0);
        }
        delete monitorEvals;
    }
    Mesh(const Mesh &) = delete;
    Mesh &operator=(const Mesh &) = delete;

    // src/Mesh.cpp:384-472.  The JSON "w" argument is ignored like in the reference (w = 0.5*sqrt(rho), :451);
    // numThreads has no meaning on the GPU and is kept for signature compatibility.
    void meshInit(Eigen::MatrixXd &Xc, Eigen::MatrixXd &Xp, Eigen::MatrixXi &F, vector<NodeType> &boundaryMask,
            MonitorFunction<D> *Mon, int numThreads, double rho, double w, double tau, bool gradUse) {
        (void)w;
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
        const int dev0 = dev ? atoi(dev) : 0;
        int nGpus = 1;
        if (const char *g = getenv("MMADMM_GPUS")) nGpus = !strcmp(g, "threads") ? numThreads : atoi(g);
        if (nGpus < 1) nGpus = 1;
        if (nGpus > 1 && integrationMode != 0 && integrationMode != 1) {
            cerr << "mm-admm_b200: Method 2 (backward Euler) runs on one GPU; ignoring MMADMM_GPUS=" << nGpus << endl;
            nGpus = 1;
        }
        if (nGpus == 1) {
            check(mmadmm_create(D, P, E, vp.data(), compMesh ? vc.data() : nullptr, f.data(), mask.data(), mon.data(), rho, tau,
                                /*dt, set by MeshIntegrator*/ 0.0, gradUse ? 1 : 0, dev0, &ctx), "mmadmm_create");
            // the reference permutes the caller's F (positive orientation)
            check(mmadmm_get_F(ctx, f.data()), "mmadmm_get_F");
            for (int e = 0; e < E; e++) for (int v = 0; v <= D; v++) F(e, v) = f[(size_t)e * (D + 1) + v];
        } else {
            // one process per GPU: fork the workers before this process has made its first CUDA call
            cout.flush(); cerr.flush(); fflush(nullptr);
            int idPipe[32][2];
            if (nGpus > 32) nGpus = 32;
            for (int r = 1; r < nGpus; r++) {
                int cmd[2];
                if (pipe(cmd) != 0 || pipe(idPipe[r]) != 0) { perror("mm-admm_b200: pipe"); exit(1); }
                pid_t pid = fork();
                if (pid < 0) { perror("mm-admm_b200: fork"); exit(1); }
                if (pid == 0) {   // worker: rank r, never returns into the driver
                    close(cmd[1]); close(idPipe[r][1]);
                    for (size_t q = 0; q < workers.size(); q++) close(workers[q].fd);
                    char id[128];
                    if (read(idPipe[r][0], id, 128) != 128) _exit(3);
                    mmadmm_ctx *c = nullptr;
                    int rc = mmadmm_create_dist(D, P, E, vp.data(), compMesh ? vc.data() : nullptr, f.data(), mask.data(), mon.data(),
                                                rho, tau, 0.0, gradUse ? 1 : 0, dev0 + r, r, nGpus, id, &c);
                    if (rc != MMADMM_OK) {
                        fprintf(stderr, "mm-admm_b200: rank %d: mmadmm_create_dist failed (%d): %s\n", r, rc, c ? mmadmm_last_error(c) : "");
                        _exit(4);
                    }
                    serve(c, cmd[0], r, P);
                    _exit(0);
                }
                close(cmd[0]); close(idPipe[r][0]);
                workers.push_back({pid, cmd[1]});
            }
            char id[128];
            check(mmadmm_comm_unique_id(id), "mmadmm_comm_unique_id");
            for (int r = 1; r < nGpus; r++) {
                if (write(idPipe[r][1], id, 128) != 128) { perror("mm-admm_b200: write"); exit(1); }
                close(idPipe[r][1]);
            }
            check(mmadmm_create_dist(D, P, E, vp.data(), compMesh ? vc.data() : nullptr, f.data(), mask.data(), mon.data(), rho, tau, 0.0,
                                     gradUse ? 1 : 0, dev0, 0, nGpus, id, &ctx), "mmadmm_create_dist");
            // re-orientation of the caller's F: the library keeps only this rank's block, so restate src/Mesh.cpp:243-260 here
            for (int e = 0; e < E; e++) {
                Eigen::Matrix<double,D,D> Em;
                for (int j = 0; j < D; j++) for (int l = 0; l < D; l++) Em(l, j) = Xp(F(e, j + 1), l) - Xp(F(e, 0), l);
                if (Em.determinant() < 0) { int t = F(e, 2); F(e, 2) = F(e, 1); F(e, 1) = t; }
            }
            distributed = true;
        }
        // page-lock the caller's Vp and let every step write it back (Mesh::updateAfterStep, src/Mesh.cpp:1030-1035): the copy of
        // step s overlaps step s+1 and is fenced in done(), outputPoints(), computeEnergy() (include/mmadmm.h,
        // mmadmm_set_writeback).  MMADMM_WRITEBACK=strict: the copy has landed when step() returns; =off: done() only.
        vpRegistered = mmadmm_host_register(Xp.data(), sizeof(double) * (unsigned long)P * D) == MMADMM_OK;
        int mode = 2;
        if (const char *wbm = getenv("MMADMM_WRITEBACK")) mode = !strcmp(wbm, "strict") ? 1 : (!strcmp(wbm, "off") ? 0 : 2);
        check(mmadmm_set_writeback(ctx, Xp.data(), /*column-major*/ 1, mode), "mmadmm_set_writeback");
    }

    // ---- multi-GPU plumbing: rank 0 (this process) tells the workers which collective call comes next ----
    enum Op { OP_EXIT = 0, OP_STEP = 1, OP_ENERGY = 2, OP_EULER = 3, OP_GETX = 4, OP_SETDT = 5 };
    struct Cmd { int op; int nIters; double a, b; };
    struct Worker { pid_t pid; int fd; };
    void broadcast(int op, int nIters = 0, double a = 0, double b = 0) {
        Cmd c{op, nIters, a, b};
        for (size_t r = 0; r < workers.size(); r++)
            if (write(workers[r].fd, &c, sizeof(c)) != (ssize_t)sizeof(c)) { cerr << "mm-admm_b200: worker " << r + 1 << " is gone" << endl; abort(); }
    }
    static void serve(mmadmm_ctx *c, int fd, int rank, int P) {
        std::vector<double> xg;
        Cmd cmd;
        for (;;) {
            ssize_t got = read(fd, &cmd, sizeof(cmd));
            if (got != (ssize_t)sizeof(cmd) || cmd.op == OP_EXIT) break;   // EOF: rank 0 died or finished
            int rc = MMADMM_OK;
            double v = 0;
            switch (cmd.op) {
                case OP_STEP: rc = mmadmm_step(c, cmd.nIters, cmd.a, &v, nullptr, nullptr, nullptr); break;
                case OP_ENERGY: rc = mmadmm_energy(c, &v); break;
                case OP_EULER: rc = mmadmm_euler_step(c, &v); break;
                case OP_SETDT: rc = mmadmm_set_dt(c, cmd.a); break;
                case OP_GETX: xg.resize((size_t)P * D); rc = mmadmm_get_x(c, xg.data()); break;
                default: rc = MMADMM_EINVAL;
            }
            if (rc != MMADMM_OK) {
                fprintf(stderr, "mm-admm_b200: rank %d: call %d failed (%d): %s\n", rank, cmd.op, rc, mmadmm_last_error(c));
                _exit(5);
            }
        }
        mmadmm_destroy(c);
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
        check(mmadmm_writeback_fence(ctx), "mmadmm_writeback_fence");
        broadcast(OP_ENERGY);
        check(mmadmm_energy(ctx, &I), "mmadmm_energy");
        return I;
    }
    // src/Mesh.cpp:1016-1036: Vp <- x.  After a step the library has already queued the write-back (mmadmm_set_writeback in
    // meshInit); the arguments are kept for the signature.
    void updateAfterStep(double dt, Eigen::VectorXd &xPrev, Eigen::VectorXd &x) {
        (void)dt; (void)xPrev; (void)x;
    }
    // Vp <- x now, whole mesh (done(); also after Method 1 / 2 steps, which do not write back).  Vp is column-major (Eigen's
    // default): the transposition runs on the device; a distributed context assembles the global vector first.
    void syncVp() {
        if (!distributed) { check(mmadmm_get_x_colmajor(ctx, Vp->data()), "mmadmm_get_x_colmajor"); return; }
        std::vector<double> xg((size_t)Vp->rows() * D);
        broadcast(OP_GETX);
        check(mmadmm_get_x(ctx, xg.data()), "mmadmm_get_x");
        for (int p = 0; p < Vp->rows(); p++) for (int l = 0; l < D; l++) (*Vp)(p, l) = xg[(size_t)p * D + l];
    }
    void fenceVp() { check(mmadmm_writeback_fence(ctx), "mmadmm_writeback_fence"); }

    // result writers, formats of src/Mesh.cpp:1049-1095 (byte-identical text; formatted by all host threads, MeshIO.h);
    // MMADMM_BINARY_IO=1 also writes the exact binary twin "<fname>.bin"
    void outputSimplices(const char *fname) {
        std::function<int(long, int)> v = [this](long i, int j) { return (*F)(i, j); };
        if (!meshio::write_text<int>(fname, F->rows(), D + 1, v)) cerr << "mm-admm_b200: cannot write " << fname << endl;
        if (meshio::binary_twin()) meshio::write_bin<int>((string(fname) + ".bin").c_str(), F->rows(), D + 1, v);
    }
    void outputPoints(const char *fname) {
        fenceVp();
        std::function<double(long, int)> v = [this](long i, int j) { return (*Vp)(i, j); };
        if (!meshio::write_text<double>(fname, Vp->rows(), D, v)) cerr << "mm-admm_b200: cannot write " << fname << endl;
        if (meshio::binary_twin()) meshio::write_bin<double>((string(fname) + ".bin").c_str(), Vp->rows(), D, v);
    }
    void outputBoundaryNodes(const char *fname) {
        fenceVp();
        std::function<double(long, int)> v = [this](long i, int j) { return (*Vp)(i, j); };
        std::function<bool(long)> keep = [this](long i) { return boundaryMask->at(i) != INTERIOR; };
        if (!meshio::write_text<double>(fname, Vp->rows(), D, v, keep)) cerr << "mm-admm_b200: cannot write " << fname << endl;
    }

    // status code of the C ABI -> the reference's behaviour on a failed assert / exit(1): message + abort
    void check(int rc, const char *what) {
        if (rc == MMADMM_OK) return;
        static const char *names[] = {"OK", "EINVAL", "ECUDA", "EINVERTED", "ESINGULAR", "ENCCL", "ENODEVICE", "EPEER"};
        cerr << "mm-admm_b200: " << what << " failed: " << (rc >= 0 && rc <= 7 ? names[rc] : "?")
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
    bool distributed = false;
    std::vector<Worker> workers;

private:
    bool vpRegistered = false;
};

#endif
