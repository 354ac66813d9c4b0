/* mmadmm.h -- C ABI of the B200-native MM-ADMM implicit-step library (libmmadmm_b200.so).
 *
 * This is the drop-in boundary for the hot path of connortannahill/MM-ADMM @ 895f4e3:
 * the reference's Mesh<D> / MeshIntegrator<D> host classes (re-implemented in
 * mm-admm_b200/host/) call these entry points instead of Eigen/OpenMP.  Plain pointers and
 * sizes only; all arrays are caller-owned, row-major / interleaved, FP64 or int32, and are
 * copied at the call.  Every function returns 0 on success or an MMADMM_E* code; no C++
 * exception crosses this boundary.  A context is NOT thread-safe; one host thread drives it.
 *
 * Layouts:  Vp, Vc, x        P x D   row-major  (x[D*v + l])         reference: interleaved VectorXd x
 *           F                E x (D+1) row-major                     reference: Eigen::MatrixXi F (values)
 *           mask             P  ints, NodeType values 0/1/2          reference: src/NodeType.h:4-8
 *           monVertex        P x D*D row-major flattened tensors     reference: MonitorFunction::evaluateAtVertices
 *           z, u             E x n   (n = D*(D+1), vertex-major inside an element)
 *           hessInv          E x n x n row-major
 */
#ifndef MMADMM_H
#define MMADMM_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mmadmm_ctx mmadmm_ctx;

enum {
    MMADMM_OK = 0,
    MMADMM_EINVAL = 1,      /* bad argument */
    MMADMM_ECUDA = 2,       /* CUDA runtime error (see mmadmm_last_error) */
    MMADMM_EINVERTED = 3,   /* an element with det(E) <= 0 was evaluated; reference: assert(Edet > 0),
                               src/AdaptationFunctional.cpp:174 */
    MMADMM_ESINGULAR = 4,   /* singular finite-difference Hessian (reference: Eigen inverse() -> inf/nan) */
    MMADMM_ENCCL = 5,       /* NCCL error */
    MMADMM_ENODEVICE = 6,   /* no CUDA device: there is no CPU fallback */
    MMADMM_EPEER = 7        /* distributed context: a peer rank did not reach an exchange within "peer_timeout_s" (default 300 s;
                               env MMADMM_PEER_TIMEOUT_S).  Ranks enter a step at independent host times: waiting is normal, the
                               bound only catches a rank that died.  Reported on every surviving rank; the CUDA context stays usable. */
};

/* State selectors for mmadmm_get_state (inspection / parity tests). */
enum { MMADMM_STATE_Z = 0, MMADMM_STATE_U = 1, MMADMM_STATE_HESSINV = 2, MMADMM_STATE_XBAR = 3,
       MMADMM_STATE_XPREV = 4, MMADMM_STATE_MONGRID = 5 };

/* Replaces Mesh<D>::Mesh / meshInit (src/Mesh.h:22-25, src/Mesh.cpp:384-472) + MeshIntegrator<D>::MeshIntegrator
 * (src/MeshIntegrator.cpp:15-62): re-orients F (reOrientElements, src/Mesh.cpp:243-260), builds the monitor
 * interpolation grid on the device (MeshInterpolator::updateMesh / nearestNeighGridMap / smoothMonitorGrid,
 * src/MeshInterpolator.cpp:67-130,165-241,365-404), w = 0.5*sqrt(rho) (src/Mesh.cpp:451), hessInvs = I, x = xPrev =
 * xBar = Vp, z = D x, u = 0, and the diagonal of t = M + dt^2 (WD)^T(WD).  Vc may be NULL (CompMesh false).
 * `device` is the CUDA device ordinal. */
int mmadmm_create(int D, int P, int E, const double *Vp, const double *Vc, const int *F, const int *mask,
                  const double *monVertex, double rho, double tau, double dt, int gradUse, int device,
                  mmadmm_ctx **out);

/* Replaces MeshIntegrator<D>::step(nIters, tol) (src/MeshIntegrator.cpp:101-191), including Mesh::predictX,
 * Mesh::prox -> bfgsOptSimplex, the x-update solves and the residual test.  Ih = energy returned by the
 * reference (sum of the un-regularised element energies at entry to the first prox sweep). */
int mmadmm_step(mmadmm_ctx *ctx, int nIters, double tol, double *Ih, int *itersTaken, double *primal, double *dual);

/* Replaces MeshIntegrator<D>::eulerStep (Method 1, src/MeshIntegrator.cpp:87-95): grad = Mesh::eulerStepMod
 * (src/Mesh.cpp:533-579: element gradients without regularisation, collected by INTERIOR vertices only), then
 * x -= (dt/tau) grad.  Ih = energy at entry.  Like the reference, this path does not go through updateAfterStep:
 * the caller reads the positions with mmadmm_get_x when it wants them. */
int mmadmm_euler_step(mmadmm_ctx *ctx, double *Ih);

/* Method 2: one MeshIntegrator::backwardsEulerStep(dt, tol) (src/MeshIntegrator.cpp:68-76 -> Mesh::backwardsEulerStep,
 * src/Mesh.cpp:1263-1341): Newton iteration on F(x) = x - xn + (dt/tau) grad I(x) = 0 from the explicit-Euler guess, gradient
 * by eulerStepMod (INTERIOR vertices only), Jacobian I + (dt/tau) * forward differences of the element gradients (FSubJac,
 * src/Mesh.cpp:1161-1261; h = 10 sqrt(eps); taken at the vertex positions Vp of construction / the last Method-0 step, which
 * this method never refreshes; rebuilt only while no step has been taken or when the 1-norm of F stalls), linear solves by
 * ILU(0) + BiCGSTAB (mmadmm_pcg_bicgstab_device; residual reduction 1e-6), stop at ||F||_1 < tol / 10 or after 1000 Newton
 * iterations.  Ih = energy at the last evaluated iterate (the value the reference returns); newtonIters / linIters (sum of
 * the inner iterations) may be NULL.  dt is the step the reference's driver passes (main.cpp:190; it enters as dt/tau in the
 * guess, the residual and the Jacobian); dt <= 0 uses the context's dt.  Single-GPU contexts only (MMADMM_EINVAL otherwise). */
int mmadmm_backward_euler_step(mmadmm_ctx *ctx, double dt, double tol, double *Ih, int *newtonIters, int *linIters);
/* The Method 2 Jacobian in CSR form (after the first backward-Euler step or mmadmm_backward_euler_jacobian with val != NULL
 * forcing a build): sizes2 = {rows, nnz}; rowptr[rows+1], col[nnz], val[nnz] may each be NULL. */
int mmadmm_backward_euler_jacobian(mmadmm_ctx *ctx, int *sizes2, int *rowptr, int *col, double *val);

/* The time step belongs to MeshIntegrator<D>, not to Mesh<D> (src/MeshIntegrator.cpp:15-17,45-51): a host Mesh<D> may
 * create the context before any integrator exists (dt = 0) and the integrator's constructor then sets dt, which
 * rebuilds the diagonal of t = M + dt^2 (WD)^T(WD).  Must be called before the first mmadmm_step. */
int mmadmm_set_dt(mmadmm_ctx *ctx, double dt);

/* What the commented-out body of Mesh<D>::setUp would do every step (src/Mesh.cpp:1006-1014: mapEvaluator->updateMesh(Vp, F);
 * mapEvaluator->interpolateMonitor(*Mon)): rebuild the monitor interpolation grid (bounding box, axes, exact 1-NN vertex -> grid
 * map, Jacobi smoothing; src/MeshInterpolator.cpp:67-130,165-259,365-404) on the device from the CURRENT node positions, with the
 * monitor tensors monVertex (P x D*D; GLOBAL P on a distributed context, same array on every rank: collective call) that the
 * caller evaluated at those positions (mmadmm_get_x).  For moving-front problems; the reference itself never re-interpolates
 *  and the default path does not call this. */
int mmadmm_update_monitor(mmadmm_ctx *ctx, const double *monVertex);

/* Replaces MeshIntegrator<D>::getEnergy -> Mesh::computeEnergy (src/Mesh.cpp:496-530). */
int mmadmm_energy(mmadmm_ctx *ctx, double *I);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same energy as mmadmm_energy (same
 * kernels, same fixed-order reduction, bit-identical result), but the per-block partial sums go into a file-scope
 * `static __device__` array of MMADMM_NOALLOC_MAX_BLOCKS doubles (1 MiB; every GPU that loads the library holds it until
 * exit) and the call makes no cudaMalloc / cudaFree / allocating library call.
 *   x_dev : node positions to evaluate at, DEVICE pointer, P x D row-major (as mmadmm_get_x_device writes them); NULL = the
 *           context's current positions.  I : host pointer, receives the energy.
 * Limits: single-GPU contexts only; at most MMADMM_NOALLOC_MAX_BLOCKS * 128 = 16,777,216 elements (larger meshes ->
 * MMADMM_EINVAL with a message in mmadmm_last_error).  All calls on one GPU share the one array, which keeps the last
 * call's partial sums (nothing in it has to start at zero: a call overwrites what it reads): calls that overlap on the GPU
 * (two contexts / two host threads on the same device) race -- serialise them. */
#define MMADMM_NOALLOC_MAX_BLOCKS 131072
int mmadmm_energy_noalloc(mmadmm_ctx *ctx, const double *x_dev, double *I);

/* Replaces Mesh::updateAfterStep / copyX (src/Mesh.cpp:996-1004,1016-1036): current node positions, P x D.
 * On a distributed context this is a collective call that assembles the GLOBAL vector on every rank. */
int mmadmm_get_x(mmadmm_ctx *ctx, double *x);
/* Same, but into a caller-provided *device* pointer (no host copy). */
int mmadmm_get_x_device(mmadmm_ctx *ctx, double *x_dev);

/* Same into a COLUMN-major P x D host array -- the storage of the caller's Eigen::MatrixXd Vp that Mesh::updateAfterStep
 * writes (src/Mesh.cpp:1030-1035): the transposition runs on the device, the copy goes straight into the caller's matrix.
 * Single-GPU contexts only. */
int mmadmm_get_x_colmajor(mmadmm_ctx *ctx, double *VpColMajor);
/* Page-lock / unlock a caller-owned host range (cudaHostRegister) so that the per-step write-back of Vp runs at PCIe speed.
 * Optional: failure only means the copy is staged by the driver. */
int mmadmm_host_register(void *p, unsigned long bytes);
int mmadmm_host_unregister(void *p);

/* Replaces the per-step Mesh::updateAfterStep (src/Mesh.cpp:1016-1036: Vp <- x after every MeshIntegrator::step) without a
 * call per step: once a target is set, every successful mmadmm_step ends by writing the node positions into `VpHost`
 * (row-major interleaved P x D, or column-major P x D = the storage of the caller's Eigen::MatrixXd when colMajor != 0; on a
 * distributed context the GLOBAL array, of which this rank writes the rows of its own vertices).
 *   mode 1 (strict):   the copy has landed when mmadmm_step returns -- the reference's semantics.
 *   mode 2 (deferred): x is snapshotted on the device at the end of the step and copied on a separate copy stream while the
 *                      next step runs.  CONTRACT: the caller must not read VpHost until mmadmm_writeback_fence (or
 *                      mmadmm_get_x*, mmadmm_restore, mmadmm_destroy, which fence) has returned, and must keep VpHost alive
 *                      until then.  The reference's driver never reads Vp between steps (main.cpp:180-215), only after
 *                      MeshIntegrator::done(); the host classes fence in done(), outputPoints() and getEnergy().
 *   mode 0: off (default).  VpHost should be page-locked (mmadmm_host_register) for the copy to be asynchronous at all.
 * Methods 1 and 2 do not write back (the reference does not call updateAfterStep there either). */
int mmadmm_set_writeback(mmadmm_ctx *ctx, double *VpHost, int colMajor, int mode);
int mmadmm_writeback_fence(mmadmm_ctx *ctx);

/* Checkpoint / restore of the whole integrator state on the device (x, xPrev, xBar, z, u, inverse Hessians, dual sums, step
 * counters, statistics): restore returns the context to exactly the state of the last checkpoint (bit-identical continuation),
 * also after a run has broken down (status word cleared).  The reference has no counterpart ("Checkpoint / resume:
 * none"); bench.py restarts its timed epochs from one post-warm-up checkpoint.  Costs one extra copy of the state in HBM.
 * On a distributed context every rank must call it at the same point of the run. */
int mmadmm_checkpoint(mmadmm_ctx *ctx);
int mmadmm_restore(mmadmm_ctx *ctx);

/* The re-oriented connectivity (the reference permutes the caller's F in place). */
int mmadmm_get_F(mmadmm_ctx *ctx, int *F);

/* Monitor grid geometry: dims = {nx, ny, nz} intervals; axes have nx+1 / ny+1 / nz+1 entries. */
int mmadmm_grid_dims(mmadmm_ctx *ctx, int *dims);

/* Copy a piece of integrator state to the host in the layouts above (parity tests / debugging). */
int mmadmm_get_state(mmadmm_ctx *ctx, int which, double *out);

/* evalMonitorOnGrid at npts host points (npts x D) -> npts x D*D   (src/MeshInterpolator.cpp:286-342). */
int mmadmm_eval_monitor(mmadmm_ctx *ctx, const double *pts, int npts, double *out);

/* Mesh::computeBlockGrad on the context's first `ne` elements with caller-supplied coordinates
 * (src/Mesh.cpp:755-772 -> src/AdaptationFunctional.cpp:102-287).  zs, dxpu: ne x n. grad/igt may be NULL. */
int mmadmm_block_grad(mmadmm_ctx *ctx, int ne, const double *zs, const double *dxpu, int computeGrad, int regularize,
                      int applyMask, double *energy, double *grad, double *igt);
/* Same on mesh elements e0 .. e0+ne-1 (zId of the reference's blockGrad); rows of zs/dxpu/energy/grad/igt are 0..ne-1. */
int mmadmm_block_grad_at(mmadmm_ctx *ctx, int e0, int ne, const double *zs, const double *dxpu, int computeGrad,
                         int regularize, int applyMask, double *energy, double *grad, double *igt);

/* Statistics since creation: hist52[k] = number of element-prox solves that ran k BFGS iterations
 * (k = 51: hit the 50-iteration cap, the reference prints "MAX IN BFGS"); proxCalls = element-prox solves;
 * admmIters = ADMM iterations executed; kernelLaunches = kernels launched by this library. */
int mmadmm_get_stats(mmadmm_ctx *ctx, long long *hist52, long long *proxCalls, long long *admmIters,
                     long long *kernelLaunches);

/* Device time (ms, CUDA events on the library's stream) accumulated per phase since the last reset:
 * ms[0]=predict, ms[1]=prox kernel, ms[2]=x-update, ms[3]=residual+reduce, ms[4]=hessian init.
 * Enabled with mmadmm_set_option(ctx, "timing", 1); counts[i] = launches timed. */
int mmadmm_get_timing(mmadmm_ctx *ctx, double *ms5, long long *counts5, int reset);

/* Options: "timing" (0/1), "stats" (0/1, default 1), "prox_variant" (kernel tile/occupancy variant, debugging),
 * "iso_grid" (0/1, default 1): when every tensor of the monitor grid is m*I bitwise, look m up in a scalar grid
 * (bit-identical results, D*D..10x fewer bytes per lookup); "iso_grid_active" is a query: returns 0 iff that path is on;
 * "usum_period" (default 64): the x-updates work on per-vertex sums of the dual variable that follow a recurrence
 *  and are re-gathered exactly every usum_period-th update; 0 = every x-update gathers z - u slot by slot,
 * the literal form of src/MeshIntegrator.cpp:137-138,159-160 (results agree to rounding, tests hold 1e-12);
 * "p2p_active" is a query: returns 0 iff a distributed context exchanges through NVLink peer stores;
 * "peer_timeout_s" (default 300): see MMADMM_EPEER;
 * "boundary_projection" (default 0): REAL BOUNDARY SLIDING, new behaviour.  The reference carries Mesh::projectOntoBoundary ->
 * projection2D / projection3D and the face lists of buildFaceList (src/Mesh.cpp:61-241), but all three call sites are commented
 * out (:636-642, :975-984, :1020-1026), so what it executes for BOUNDARY_FREE nodes is "gradient live, Hessian diagonal 1, no
 * projection" -- the default here too.  Bit 0 re-enables the call in Mesh::prox (the z slots of the FREE vertices of an element are
 * projected onto the boundary faces meeting at that vertex, after the element's BFGS solve); bit 1 the call in updateAfterStep
 * (x of every FREE node, before Vp <- x).  Face geometry comes from the node positions at the end of the previous outer step
 * (the reference's *Vp).  No reference result exists with it enabled: parity is GPU vs the CPU oracle only (unpinned beyond it).
 * Single-GPU contexts. */
int mmadmm_set_option(mmadmm_ctx *ctx, const char *name, int value);

/* The CUDA stream every kernel of this context is launched on (cudaStream_t as void*). */
void *mmadmm_stream(mmadmm_ctx *ctx);

/* ---- multi-GPU: element-block partition, one process per GPU (the reference has no counterpart, its only
 * parallelism is the OpenMP loop over simplices at src/Mesh.cpp:945-948) ---- */

/* Host-side partition plan of `rank` out of `nranks` computed from the GLOBAL connectivity (no device needed).
 * sizes6 = {e0, e1, nLocalVertices, nSharedVertices, nNeighbours, lenNeighbourLists}. */
typedef struct mmadmm_plan mmadmm_plan;
enum { MMADMM_PLAN_VERTS = 0,       /* local vertex -> global vertex, ascending            [nLocalVertices] */
       MMADMM_PLAN_FLOC = 1,        /* owned elements in local vertex ids                  [(e1-e0)*(D+1)]  */
       MMADMM_PLAN_VALENCE = 2,     /* GLOBAL valence of every local vertex                [nLocalVertices] */
       MMADMM_PLAN_SHARED = 3,      /* local ids of vertices shared with other ranks       [nSharedVertices] */
       MMADMM_PLAN_NBR_RANK = 4,    /* neighbour ranks, ascending                          [nNeighbours]    */
       MMADMM_PLAN_NBR_PTR = 5,     /* segment offsets into NBR_LIST                       [nNeighbours+1]  */
       MMADMM_PLAN_NBR_LIST = 6,    /* per neighbour: indices into SHARED, ascending global vertex */
       MMADMM_PLAN_CONTRIB_PTR = 7, /* per shared vertex: offsets into CONTRIB             [nSharedVertices+1] */
       MMADMM_PLAN_CONTRIB = 8,     /* contributions in ascending rank order: receive-buffer position, or -1 = own */
       MMADMM_PLAN_OWNER = 9 };     /* 1 if this rank is the lowest rank holding the vertex [nLocalVertices] */
int mmadmm_plan_create(int D, int P, int E, const int *F, int nranks, int rank, mmadmm_plan **out);
int mmadmm_plan_sizes(const mmadmm_plan *plan, int *sizes6);
int mmadmm_plan_get(const mmadmm_plan *plan, int which, int *out);
void mmadmm_plan_destroy(mmadmm_plan *plan);

/* 128-byte NCCL unique id for the communicator of a distributed context (call on rank 0, broadcast it by any means). */
int mmadmm_comm_unique_id(void *id128);

/* Distributed form of mmadmm_create: every rank passes the same GLOBAL mesh and its (rank, nranks, id).  The rank keeps
 * the element block of its plan and private copies of the vertices those elements touch; per ADMM iteration the
 * shared-vertex partial sums of D^T(w(z-u)) are exchanged with ncclSend/ncclRecv and the three residual scalars with
 * one ncclAllGather.  mmadmm_step / mmadmm_energy are collective calls and return identical values on every rank. */
int mmadmm_create_dist(int D, int P, int E, const double *Vp, const double *Vc, const int *F, const int *mask,
                       const double *monVertex, double rho, double tau, double dt, int gradUse, int device, int rank,
                       int nranks, const void *id128, mmadmm_ctx **out);

/* Number of local vertices / elements of a context (P, E for a single-GPU context). */
int mmadmm_local_sizes(mmadmm_ctx *ctx, int *Ploc, int *Eloc);
/* Local vertices' positions scattered into a caller array of GLOBAL size P x D (other rows untouched; no collective). */
int mmadmm_get_x_local(mmadmm_ctx *ctx, double *xGlobal);

/* ---- general sparse path of the x-update: preconditioned CG on a fixed CSR matrix (mm-admm_b200/csrc/pcg.cu) ----
 * Replaces Eigen::ConjugateGradient<SparseMatrix, Lower|Upper> with its default Jacobi preconditioner
 * (src/MeshIntegrator.cpp:54-55,138,160: compute(t), solve(vec); zero initial guess, stop at ||r|| <= tol ||b||) and, with
 * precond = 1, the level-0 ILU + CG path of lib/LASolver (ILU_class.cpp:98-527, accel_class.cpp:402-491) that the
 * reference's Method 2 uses.  On the ADMM matrix itself t is diagonal and mmadmm_step does not call this.
 * CSR: rowptr[n+1], strictly ascending col[] per row, stored diagonal required.  precond: 0 = Jacobi, 1 = ILU(0). */
typedef struct mmadmm_pcg mmadmm_pcg;
int mmadmm_pcg_create(int n, const int *rowptr, const int *col, const double *val, int precond, int device, mmadmm_pcg **out);
/* x = A^{-1} b from host arrays; iters = CG iterations taken, relres = ||r|| / ||b|| at exit. */
int mmadmm_pcg_solve(mmadmm_pcg *h, const double *b, double *x, double tol, int maxit, int *iters, double *relres);
/* Same with device pointers on the solver's device (no host copies). */
int mmadmm_pcg_solve_device(mmadmm_pcg *h, const double *b_dev, double *x_dev, double tol, int maxit, int *iters, double *relres);
/* BiCGSTAB exactly as LASolver runs it for the reference's Method 2 (lib/LASolver/accel_class.cpp:229-399 scaler_cgstab driven
 * by MatrixIter::solve, MatrixIter.cpp:635-800, with the ParamIter of src/Mesh.cpp:266-304: iaccel = 0, ILU level 0, natural
 * order, no scaling, zero initial guess, toler = 0): stops when ||r|| / ||b|| < resid_reduc.  Needs precond = 1.
 * iters = iterations done, -1 if nitmax iterations did not converge (LASolver's convention). */
int mmadmm_pcg_bicgstab(mmadmm_pcg *h, const double *b, double *x, double resid_reduc, int nitmax, int *iters, double *relres);
int mmadmm_pcg_bicgstab_device(mmadmm_pcg *h, const double *b_dev, double *x_dev, double resid_reduc, int nitmax, int *iters,
                               double *relres);
/* y = A x (host arrays): the SpMV kernel alone (reference: matmult, accel_class.cpp:521-549). */
int mmadmm_pcg_spmv(mmadmm_pcg *h, const double *x, double *y);
/* z = M^{-1} r: Jacobi scaling or the level-scheduled ILU(0) triangular solves (reference: scaler_ILU::solve, ILU_class.cpp:447-527). */
int mmadmm_pcg_precond(mmadmm_pcg *h, const double *r, double *z);
/* ILU(0) factor values in A's CSR pattern (unit lower below the diagonal, U on and above it). */
int mmadmm_pcg_get_ilu(mmadmm_pcg *h, double *lu);
/* New values on the SAME pattern, from a DEVICE array in the handle's CSR order: replaces LASolver's numeric re-factorisation
 * (jac->factor after set_entry, lib/LASolver/ILU_class.cpp:300-444, as Mesh::buildEulerJac runs it at every Jacobian rebuild,
 * src/Mesh.cpp:1111-1261).  Pattern, level schedules and buffers are kept; the Jacobi diagonal and (precond = 1) the ILU(0)
 * factor are recomputed on the device, the factor level by level in LASolver's IKJ order with separately rounded multiply and
 * subtract: bit-identical to a factor computed on the host.  MMADMM_ESINGULAR on a zero pivot. */
int mmadmm_pcg_refactor_device(mmadmm_pcg *h, const double *val_dev);
/* "spmv_format": 0 (default) = CSR with a sub-warp per row; 1 = SELL-C-sigma (C = 32 rows per slice = one warp, sigma = 4096-row
 * sorting window, column-major slices: every value / index load is one coalesced line; a thread owns a row and accumulates in
 * ascending column order like LASolver's matmult, accel_class.cpp:521-549), built from the CSR arrays on first use.  Both serve
 * every solver entry point.  "sell_padding_permille" is a query: padding entries per 1000 stored entries.
 * "tri_mode": how the ILU(0) triangular solves (LASolver's scaler_ILU::solve, ILU_class.cpp:447-527) run.  0 = one launch per
 * level.  1 (default of a bare handle) = every run of consecutive levels of at most 8192 rows is ONE launch of a thread-block
 * cluster that walks the levels with cluster barriers (wider levels keep a launch each).  2 (what the Method 2 context uses;
 * option "m2_tri_mode") = DATAFLOW: one persistent launch per solve, no barrier at all -- the result vector is pre-filled with
 * a sentinel, a row's 8-byte result is its own ready flag, consumers poll their operands through L2, warps claim 32-group
 * chunks in level order from an atomic counter; the factor is read from a chunk-major copy (coalesced), in BS x BS block
 * form when the pattern is made of dense blocks and row_block = BS ("tri_blocks" = 0 forces the generic form).
 * "tri_layout": 1 = the cluster kernels of tri_mode 1 read a level-major (ELL) copy of the factor instead of CSR.
 * "tri_schedule" = row_block | tri_layout << 8 | (tri_mode + 1) << 12 sets the three at once (one rebuild of the schedules;
 * bits 12-13 zero = keep tri_mode).  "tri_proxy" 1: the dataflow kernels let a waiting warp poll through one lane first (default 0).  "tri_sf_ctas_per_sm_x2" (1..8, default 2): CTAs of the dataflow kernels = SMs x value / 2.
 * "ilu_thread_kernel" 1: the ILU(0) factor kernel with a thread per row even where the warp-per-row form fits (rows of at
 * most 64 entries); same bits (tests).
 * The copies cost memory: chunk-major factor copy = (values 8 B + source index 4 B) per stored entry + padding, once per
 * triangular factor (C2-size Method 2 Jacobian, 58.7 M entries: about 1.3 GB next to the 0.7 GB of CSR values and indices).
 * All modes and layouts run every row's FMA chain in ascending column order: results are bit-identical.
 * "row_block": BS = 1 (default), 2, 3, 4 or 6 (the last group may be short): the triangular solves schedule GROUPS of BS consecutive rows, one
 * thread solving its group's rows in natural order with the operands of its own group in registers, so only dependencies
 * between groups make levels: on a Jacobian with D x D vertex blocks BS = D halves (2D) / thirds (3D) the level count (the
 * Method 2 context sets BS = D; option "m2_row_block").  Same FMA chains in the same order: bit-identical to BS = 1. */
int mmadmm_pcg_set_option(mmadmm_pcg *h, const char *name, int value);
/* Host-only (no device): the level schedule that "row_block" = bs gives the triangular solves of an ILU(0) factor on this pattern
 * (LASolver solves row by row, ILU_class.cpp:447-527; the levels are this library's parallel order of the same arithmetic).
 * levL / levU (may be NULL): level of every group [g bs, (g+1) bs) of the lower / upper solve, ceil(n / bs) entries each;
 * info3 = {lower levels, upper levels, 1 if the pattern is made of dense bs x bs blocks (block form of the dataflow solve)}. */
int mmadmm_pcg_plan_levels(int n, const int *rowptr, const int *col, int bs, int *levL, int *levU, int *info3);
/* info4 = {threads per row of the SpMV kernel, lower levels, upper levels (of groups when row_block > 1), kernels launched so far}. */
int mmadmm_pcg_info(mmadmm_pcg *h, int *info4);
/* Average device time (ms, CUDA events) of one SpMV + dot launch over `reps` launches. */
int mmadmm_pcg_time_spmv(mmadmm_pcg *h, int reps, double *ms);
/* Average device time (ms, CUDA events) of one preconditioner application z = M^{-1} r (Jacobi scaling, or the two ILU(0)
 * triangular solves in the current "tri_mode" / "row_block") over `reps` applications. */
int mmadmm_pcg_time_precond(mmadmm_pcg *h, int reps, double *ms);
const char *mmadmm_pcg_last_error(mmadmm_pcg *h);
/* Measured FP64 FMA peak of the device in TFLOP/s (DFMA microbenchmark, 8 independent chains per thread). */
int mmadmm_measure_fp64_peak(int device, double *tflops);
void mmadmm_pcg_destroy(mmadmm_pcg *h);

void mmadmm_destroy(mmadmm_ctx *ctx);
const char *mmadmm_last_error(mmadmm_ctx *ctx);
const char *mmadmm_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MMADMM_H */
