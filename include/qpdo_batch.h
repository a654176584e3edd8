/* qpdo_batch.h -- ADDITIVE extension of the qpdo.h C ABI: many independent small QPs at once.
 *
 * The reference has no batch entry (SURVEY.md 8(b) "additive extension"): its user loops over
 * qpdo_setup / qpdo_solve / qpdo_cleanup per problem (src/qpdo.c:49,304,591).  This header is
 * what such a loop binds instead when all problems share one sparsity pattern (MPC-style,
 * BASELINE.json configs[4]): every instance runs the SAME algorithm as qpdo_solve
 * (src/qpdo.c:304-476 with iteration.c, newton.c, linesearch.c, termination.c, scaling.c), one
 * CTA per QP with the whole solver state in shared memory.  qpdo.h itself is untouched.
 *
 * Plain pointers and sizes only.  Every function returns 0 on success or a negative code;
 * qpdo_dev_last_error() (qpdo_cuda.h) has the text.  There is no CPU fallback.
 */
#ifndef QPDO_B200_QPDO_BATCH_H
#define QPDO_B200_QPDO_BATCH_H

#include "qpdo.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct qpdo_batch qpdo_batch;

/* per-QP result summary: what ranks gather over NCCL; fields as QPDOInfo
 * (reference include/types.h:53-72) */
typedef struct {
    int    status_val;          /* reference include/constants.h:17-26 */
    int    iterations;          /* loop passes incl. outer updates, src/qpdo.c:455 */
    int    oterations;
    int    n_active;            /* final size of the active set, src/newton.c:96-107 */
    double objective;
    double res_prim_norm;
    double res_dual_norm;
} qpdo_batch_summary;           /* 40 bytes */

/* Shared structure: Q = lower triangle (n x n) and A (m x n), 32-bit CSC patterns with sorted
 * rows, no duplicates.  `capacity` = most instances one solve call will carry.  `settings`
 * follows qpdo_set_default_settings / validate_settings (src/validate.c:34-170); verbose,
 * print_interval and max_time are ignored.  Runs the fill-reducing symbolic analysis once
 * (replaces the per-refactor cholmod_analyze, src/cholmod_interface.c:10).  device < 0: current. */
int  qpdo_batch_create(qpdo_batch **out, int n, int m, const int *Qp, const int *Qi,
                       const int *Ap, const int *Ai, int capacity,
                       const QPDOSettings *settings, int device);
void qpdo_batch_destroy(qpdo_batch *b);

/* Solve `count` <= capacity instances.  HOST buffers, instance-major:
 *   Qx [count][nnzQ]  Ax [count][nnzA]  q [count][n]  l, u [count][m]  (bounds clipped to +-1e20)
 *   c  [count] or NULL (constant terms, QPDOData.c)
 *   x0 [count][n], y0 [count][m] or NULL: qpdo_warm_start's arguments (src/qpdo.c:217-299)
 * Outputs (any may be NULL): x [count][n], y [count][m] = QPDOSolution; summary [count];
 *   active_set [count][m] bytes = QPDOCholmod.active_set.
 * Copies host->device, runs the batch kernel, copies back; synchronous. */
int  qpdo_batch_solve(qpdo_batch *b, int count, const double *Qx, const double *Ax,
                      const double *q, const double *l, const double *u, const double *c,
                      const double *x0, const double *y0,
                      double *x, double *y, qpdo_batch_summary *summary, unsigned char *active_set);

/* Same with DEVICE buffers already resident in HBM (same layouts); asynchronous on the
 * batch's stream unless `sync` != 0.  *kernel_ms (if non-NULL, implies sync) receives the
 * CUDA-event time of the batch kernel alone. */
int  qpdo_batch_solve_device(qpdo_batch *b, int count, const double *dQx, const double *dAx,
                             const double *dq, const double *dl, const double *du, const double *dc,
                             const double *dx0, const double *dy0,
                             double *dx, double *dy, qpdo_batch_summary *dsummary,
                             unsigned char *dactive, int sync, float *kernel_ms);

/* One process, `ngpu` GPUs of one box: batches[g] = a batch created on device g with the same pattern and settings.
 * Instance i goes to GPU floor(i * ngpu / count) (contiguous blocks); every block runs through its
 * GPU's chunked copy / kernel pipeline, all GPUs concurrently; buffers and results exactly as qpdo_batch_solve.
 * There is no inter-GPU traffic (the summaries of all blocks land in the caller's one array); a multi-process job
 * gathers them with one all-gather instead (bench.py). */
int  qpdo_batch_solve_multi(qpdo_batch *const *batches, int ngpu, int count, const double *Qx, const double *Ax,
                            const double *q, const double *l, const double *u, const double *c,
                            const double *x0, const double *y0,
                            double *x, double *y, qpdo_batch_summary *summary, unsigned char *active_set);

/* Work-queue order.  on != 0: every solve that writes summaries also leaves, on the device, the order "longest
 * instance first" by the iteration counts it just produced; the next solve of the SAME number of instances pulls its
 * queue in that order (a receding-horizon loop re-solves similar problems; the ragged tail of a launch is set by
 * where the longest instances start).  Results do not depend on the order.  Default: off (natural order). */
int  qpdo_batch_set_adaptive_order(qpdo_batch *b, int on);

/* Convenience for callers holding the reference's own structs: `count` QPDOData problems
 * (include/types.h:81-90) that share one pattern; returns -4 if they do not.  Creates a batch,
 * solves, destroys.  x / y: arrays of `count` host pointers (QPDOSolution.x / .y) or NULL. */
int  qpdo_batch_solve_data(const QPDOData *const *data, int count, const QPDOSettings *settings,
                           double *const *x, double *const *y, qpdo_batch_summary *summary, int device);

/* diagnostics: per-phase clock64 accounting inside the batch kernel.  count == 0 switches it on for
 * the following solves; count > 0 returns the mean SM cycles per instance of the last solve in
 * cycles8[0..7] = { setup + scaling, residuals + outer updates, Newton preparation, K assembly,
 * factorization + forward solve, backward solve, direction SpMVs, linesearch + iterate update }. */
int  qpdo_batch_timing(qpdo_batch *b, int count, double *cycles8);

/* diagnostics: shared memory per CTA, CTAs launched, nonzeros of L under the batch ordering */
int  qpdo_batch_info(const qpdo_batch *b, int *smem_bytes, int *ctas, int *lnz, int *threads);

#ifdef __cplusplus
}
#endif
#endif
