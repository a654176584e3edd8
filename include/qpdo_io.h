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

/* Frees a problem returned by qpdo_read_qps (NULL-safe). */
void qpdo_free_data(QPDOData *data);

#ifdef __cplusplus
}
#endif
#endif
