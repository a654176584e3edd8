/* qpdo_io.h -- problem input for the drop-in library.
 *
 * The reference has no problem reader: its users build QPDOData in MATLAB (interfaces/mex/qpdo.m) or
 * by hand in C.  This additive entry reads a QP in free-format QPS (the Maros-Meszaros format:
 * ROWS / COLUMNS / RHS / RANGES / BOUNDS / QUADOBJ or QMATRIX) straight into the reference's own
 * QPDOData (include/types.h:81-90), ready for qpdo_setup:
 *
 *     min 1/2 x'Qx + q'x + c   s.t.  l <= A x <= u
 *
 * QPDO has no separate variable bounds, so every variable whose QPS bound is not (-inf, +inf) gets
 * an identity row appended below the m0 constraint rows (what qpdo.m's callers do); infinite
 * bounds are +-QPDO_INFTY = 1e20 (reference include/constants.h:38).  Q is returned as its lower
 * triangle, CSC, stype = -1 (what the mex passes, qpdo_mex.c:150); A is CSC with sorted rows.
 * Host-only code: no CUDA call is made.
 */
#ifndef QPDO_IO_H
#define QPDO_IO_H

#include "qpdo.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Returns 0 and a newly allocated problem in *out, or a negative code (-1 cannot open, -2 syntax /
 * unknown name, with the offending line number in *err_line when err_line != NULL, -3 allocation).
 * m_rows (may be NULL) receives m0, the number of QPS constraint rows; rows m0 .. m-1 are the
 * variable bounds, in column order. */
int qpdo_read_qps(const char *path, QPDOData **out, c_int *m_rows, c_int *err_line);

/* Frees a problem returned by qpdo_read_qps / qpdo_read_mtx_problem (NULL-safe). */
void qpdo_free_data(QPDOData *data);

/* MatrixMarket (the format of CHOLMOD's reader, reference SuiteSparse/CHOLMOD/Check/cholmod_read.c;
 * QPDO itself never calls it).  Reads "matrix coordinate|array real|integer|pattern
 * general|symmetric|skew-symmetric" into CSC with sorted rows and summed duplicates: symmetric files
 * give the lower triangle with stype = -1 (what qpdo_setup wants for Q), everything else stype = 0
 * (skew-symmetric is expanded).  Same return codes as qpdo_read_qps. */
int qpdo_read_mtx(const char *path, cholmod_sparse **out, c_int *err_line);
void qpdo_free_sparse(cholmod_sparse *S);

/* A vector stored as an n x 1 or 1 x n MatrixMarket matrix (array or coordinate); *out is malloc'ed. */
int qpdo_read_mtx_vector(const char *path, c_float **out, c_int *len, c_int *err_line);

/* A whole problem from MatrixMarket files.  A_path is required (m x n, general).  Q_path NULL = no
 * quadratic term; a Q tagged general is reduced to its lower triangle.  q_path / l_path / u_path
 * NULL = zeros / -QPDO_INFTY / +QPDO_INFTY; bounds beyond +-1e20 (or "inf") are clipped to +-QPDO_INFTY
 * as qpdo.m:138-139 does.  -4 = the files' dimensions do not agree. */
int qpdo_read_mtx_problem(const char *Q_path, const char *A_path, const char *q_path, const char *l_path,
                          const char *u_path, QPDOData **out, c_int *err_line);

#ifdef __cplusplus
}
#endif
#endif
