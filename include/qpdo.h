/* qpdo.h -- public C API of the B200-native QPDO Newton hot path.
 *
 * Drop-in boundary for the reference's solver API (reference include/qpdo.h:14-56): the same
 * eight entry points with the same signatures, and layout-identical public structs (reference
 * include/types.h:14-224) in the DLONG + PROFILING configuration, which is the only build recipe
 * the reference ships (interfaces/mex/qpdo_make.m:300,331): c_float = double, c_int = long.
 *
 * A caller compiled against the reference headers (which pull in cholmod.h) links against this
 * library unchanged.  A caller that has no CHOLMOD headers includes this file instead: the three
 * CHOLMOD structs that appear in the public structs are declared below as layout-compatible
 * shims (SuiteSparse/CHOLMOD/Include/cholmod_core.h:1214-1263 sparse, :1894-1905 dense); the
 * factor type stays opaque because nothing in the solver API dereferences it.
 */
#ifndef QPDO_B200_QPDO_H
#define QPDO_B200_QPDO_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef double c_float;   /* reference global_opts.h:41 (Real)            */
typedef long   c_int;     /* reference global_opts.h:15 (SuiteSparse_long) */

/* ------------------------------------------------------------------ CHOLMOD layout shims -- */
#ifndef CHOLMOD_H
typedef struct cholmod_sparse_struct {
    size_t nrow, ncol, nzmax;
    void  *p;        /* column pointers, ncol+1 (c_int when itype == CHOLMOD_LONG)   */
    void  *i;        /* row indices                                                   */
    void  *nz;       /* per-column counts when !packed                                */
    void  *x;        /* values (double)                                               */
    void  *z;
    int    stype;    /* 0 unsymmetric, >0 upper stored, <0 lower stored               */
    int    itype;    /* CHOLMOD_INT 0 / CHOLMOD_INTLONG 1 / CHOLMOD_LONG 2            */
    int    xtype;    /* CHOLMOD_PATTERN 0 / CHOLMOD_REAL 1                            */
    int    dtype;    /* CHOLMOD_DOUBLE 0                                              */
    int    sorted;
    int    packed;
} cholmod_sparse;

typedef struct cholmod_dense_struct {
    size_t nrow, ncol, nzmax, d;
    void  *x;
    void  *z;
    int    xtype, dtype;
} cholmod_dense;

typedef struct cholmod_factor_struct cholmod_factor;   /* opaque here */
#define QPDO_CHOLMOD_SHIM 1
#endif

#define QPDO_CHOLMOD_INT   0
#define QPDO_CHOLMOD_LONG  2
#define QPDO_CHOLMOD_REAL  1

/* ------------------------------------------------------------------ status codes ---------- */
/* reference include/constants.h:17-26 */
#define QPDO_SOLVED                  (1)
#define QPDO_DUAL_TERMINATED         (2)
#define QPDO_NON_CVX                 (-1)
#define QPDO_PRIMAL_DUAL_INFEASIBLE  (-2)
#define QPDO_PRIMAL_INFEASIBLE       (-3)
#define QPDO_DUAL_INFEASIBLE         (-4)
#define QPDO_MAX_ITER_REACHED        (-5)
#define QPDO_MAX_TIME_REACHED        (-6)
#define QPDO_UNSOLVED                (-10)
#define QPDO_ERROR                   (-99)

#ifndef QPDO_NULL
#define QPDO_NULL 0
#endif
#ifndef QPDO_INFTY
#define QPDO_INFTY ((c_float)1E+20)     /* reference constants.h:38 */
#endif

/* hidden constants of the reference that change iterates (constants.h:62,69) */
#define QPDO_MIN_SCALING      (1e-9)
#define QPDO_MAX_RANK_UPDATE  100

/* ------------------------------------------------------------------ public structs -------- */
/* reference types.h:14-17 (linesearch scratch element; kept for layout only) */
typedef struct array_element { c_float x; size_t i; } array_element;

/* reference types.h:27-30 */
typedef struct { c_float *x; c_float *y; } QPDOSolution;

typedef struct QPDO_TIMER QPDOTimer;

/* reference types.h:40-47 */
typedef struct {
    c_float *D, *Dinv, *E, *Einv;
    c_float  c, cinv;
} QPDOScaling;

/* reference types.h:53-72 (PROFILING layout) */
typedef struct {
    c_int   iterations;
    c_int   oterations;
    char    status[32];
    c_int   status_val;
    c_float res_prim_norm, res_dual_norm, res_prim_in_norm, res_dual_in_norm;
    c_float objective;
    c_float setup_time, solve_time, run_time;
} QPDOInfo;

/* reference types.h:81-90 */
typedef struct {
    size_t          n, m;
    cholmod_sparse *Q;     /* n x n, CSC, stype != 0 (mex passes the lower triangle, -1) */
    cholmod_sparse *A;     /* m x n, CSC                                                */
    c_float        *q;
    c_float         c;
    c_float        *l, *u;
} QPDOData;

/* reference types.h:96-116 */
typedef struct {
    c_float max_time;
    c_int   max_iter;
    c_int   inner_max_iter;
    c_float eps_abs;
    c_float eps_abs_in;
    c_float eps_prim_inf;
    c_float eps_dual_inf;
    c_float rho;
    c_float theta;
    c_float delta;
    c_float mu_min;
    c_int   proximal;
    c_float sigma_init;
    c_float sigma_upd;
    c_float sigma_min;
    c_int   scaling;
    c_int   verbose;
    c_int   print_interval;
    c_int   reset_newton_iter;
} QPDOSettings;

/* reference types.h:121-142.  In this library LD points at the opaque device context
 * (struct qpdo_dev), LD_Q is unused, the cholmod_dense members are host shims over the same
 * buffers as work->dx etc., At_scale / At_sqrt_mu are NULL (they live on the device). */
typedef struct {
    cholmod_factor *LD;
    cholmod_factor *LD_Q;
    cholmod_dense  *E_temp, *D_temp, *linsys_rhs, *dx, *dy, *Adx, *Qdx, *Atdy;
    c_int   reset_newton;
    c_int  *active_set;
    c_int  *active_set_old;
    c_int   n_active_set;
    c_int  *enter;
    c_int   n_enter;
    c_int  *leave;
    c_int   n_leave;
    cholmod_dense  *At_scale;
    cholmod_sparse *At_sqrt_mu;
} QPDOCholmod;

/* reference types.h:147-224.  Host members that callers read after qpdo_solve (x, y, dx, dy,
 * solution, info, chol->active_set; see interfaces/mex/qpdo_mex.c:237-280) are filled from the
 * device at the end of every solve; the remaining vectors are allocated for layout/ownership
 * compatibility and mirror device state only where stated in DESIGN.md. */
typedef struct {
    QPDOData *data;
    c_float *x, *y, *Ax, *Qx, *Aty;
    c_int    initialized;
    c_float *temp_m, *temp_n, *temp_2m;
    c_float *mu, *sqrt_mu;
    c_float  sqrt_mu_min, sqrt_delta;
    c_int    n_mu_changed;
    c_float  sigma;
    c_int    sigma_mined;
    c_float  norm_q;
    c_float *xbar, *ybar, *dx, *dy;
    c_float  tau;
    c_float *Qdx, *Adx, *Atdy;
    c_float *w, *z, *df, *res_prim, *res_dual, *res_prim_old, *res_prim_in, *res_dual_in;
    c_float *linsys_rhs;
    c_float  res_prim_norm_old, res_dual_norm_old;
    c_float  ls_eta, ls_beta;
    c_float *ls_delta, *ls_alpha;
    array_element *ls_taus;
    c_int   *ls_idx_L, *ls_idx_P, *ls_idx_J;
    c_float  eps_prim, eps_dual, eps_prim_in, eps_dual_in, eps_in;
    c_float *D_temp, *E_temp;
    QPDOCholmod  *chol;
    QPDOSettings *settings;
    QPDOScaling  *scaling;
    QPDOSolution *solution;
    QPDOInfo     *info;
    QPDOTimer    *timer;
} QPDOWorkspace;

/* ------------------------------------------------------------------ the eight entry points - */
/* replaces reference src/qpdo.c:24-44 */
void qpdo_set_default_settings(QPDOSettings *settings);
/* replaces src/qpdo.c:49-212; returns QPDO_NULL on invalid data / settings / no usable GPU */
QPDOWorkspace *qpdo_setup(const QPDOData *data, const QPDOSettings *settings);
/* replaces src/qpdo.c:217-299; either pointer may be NULL (zeros) */
void qpdo_warm_start(QPDOWorkspace *work, c_float *x_warm_start, c_float *y_warm_start);
/* replaces src/qpdo.c:304-476 */
void qpdo_solve(QPDOWorkspace *work);
/* replaces src/qpdo.c:481-517 */
void qpdo_update_settings(QPDOWorkspace *work, const QPDOSettings *settings);
/* replaces src/qpdo.c:522-544; either pointer may be NULL (unchanged) */
void qpdo_update_bounds(QPDOWorkspace *work, const c_float *bmin, const c_float *bmax);
/* replaces src/qpdo.c:549-586 */
void qpdo_update_q(QPDOWorkspace *work, const c_float *q);
/* replaces src/qpdo.c:591-689; NULL-safe */
void qpdo_cleanup(QPDOWorkspace *work);

/* helpers some callers link (reference include/util.h:14,26; the mex gateway includes util.h) */
QPDOSettings *copy_settings(const QPDOSettings *settings);
void update_status(QPDOInfo *info, c_int status_val);

#ifdef __cplusplus
}
#endif
#endif /* QPDO_B200_QPDO_H */
