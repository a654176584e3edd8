/* symbolic.h -- host-side symbolic analysis of the regularised Newton matrix
 *
 *     K = Q + sigma*I + A' diag(omega) A          (n x n, SPD)
 *
 * on the fixed SUPERSET pattern pattern(Q) U pattern(A'A) (every constraint row present).  It runs
 * once per qpdo_setup and replaces the analyze-on-every-refactor of the reference
 * (src/cholmod_interface.c:8-30 -> cholmod_analyze): fill-reducing minimum-degree ordering,
 * elimination tree, postorder, column counts, relaxed supernodes, multifrontal assembly maps and
 * the level schedule the GPU numeric factorization (factor.cu) and solves (trisolve.cu) follow.
 * Written from scratch; CHOLMOD/AMD sources are not used.
 */
#ifndef QPDO_B200_SYMBOLIC_H
#define QPDO_B200_SYMBOLIC_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct qpdo_symbolic {
    int      n, m;
    int      nsuper;       /* number of supernodes (= fronts)                                 */
    int      nlevel;       /* height of the supernodal assembly tree                           */
    int     *perm;         /* [n]  perm[new] = old                                              */
    int     *pinv;         /* [n]  pinv[old] = new                                              */
    int     *sn_start;     /* [nsuper+1] first (permuted) column of each supernode             */
    int     *sn_nr;        /* [nsuper]   number of rows strictly below the pivot block         */
    int64_t *sn_rowptr;    /* [nsuper+1] offsets into rows[] / rel[]                           */
    int     *rows;         /* sorted permuted row indices below the pivot block                */
    int     *rel;          /* rows[] expressed as positions in the PARENT front (0..Np-1)     */
    int     *sn_parent;    /* [nsuper]   parent supernode or -1                                */
    int     *child_ptr;    /* [nsuper+1]                                                        */
    int     *child_idx;    /* children in ascending order (fixed summation order)             */
    int64_t *Lp;           /* [nsuper+1] offset of each L panel (N x nc, column major, ld = N) */
    int64_t *Up;           /* [nsuper+1] offset of each update matrix (nr x nr, ld = nr)       */
    int     *sn_level;     /* [nsuper]   height above the leaves                               */
    int     *level_ptr;    /* [nlevel+1]                                                        */
    int     *level_sn;     /* supernodes ordered by level, then by decreasing front order      */
    int     *col2sn;       /* [n] supernode of each permuted column                             */
    /* entries of the lower triangle of P K P' (CSC, sorted) and where they land in Lx */
    int64_t  nnzK;
    int     *Kp;           /* [n+1]                                                             */
    int     *Ki;           /* [nnzK] permuted row index                                         */
    int64_t *Kdst;         /* [nnzK] offset into Lx                                             */
    int     *Kq;           /* [nnzK] index into the (lower CSC) values of Q, or -1             */
    int64_t *Ktp;          /* [nnzK+1] range of contributing (row, a_ir*a_ic) products         */
    int64_t  ntriple;
    int     *Kt_row;       /* [ntriple] constraint row i                                        */
    int     *Kt_a1;        /* [ntriple] index into CSC values of A of a_ir                      */
    int     *Kt_a2;        /* [ntriple] index into CSC values of A of a_ic                      */
    /* statistics */
    int64_t  lnz;          /* sum of column counts (incl. diagonal), excluding relaxation zeros */
    int64_t  lnz_stored;   /* doubles in Lx (panels, incl. unused upper part of pivot blocks)  */
    int64_t  usize;        /* doubles in the update-matrix arena                                */
    double   flops;        /* sum_j colcount_j^2 (CHOLMOD's definition, rowcolcounts.c:516-528) */
    int      max_front;    /* largest front order N = nc + nr                                  */
    int      max_nc;
} qpdo_symbolic;

/* Inputs are clean int32 CSC with sorted, duplicate-free row indices: Q lower triangle (n x n),
 * A (m x n).  ordering: 0 = natural, 1 = minimum degree (default), 2 = nested dissection (level structures, minimum-degree leaves).  Returns NULL on failure and
 * writes a message into err (if non-NULL, at least 256 bytes). */
qpdo_symbolic *qpdo_symbolic_analyze(int n, int m, const int *Qp, const int *Qi,
                                     const int *Ap, const int *Ai, int ordering, char *err);
void qpdo_symbolic_free(qpdo_symbolic *S);

/* fill-reducing ordering alone (exposed for tests): graph in CSC-like adjacency form without
 * self loops, symmetric; perm_out[new] = old */
int qpdo_order_mindeg(int n, const int64_t *gp, const int *gi, int *perm_out);
/* nested dissection by BFS level structures with minimum-degree leaves of at most `leaf` vertices */
int qpdo_order_nd(int n, const int64_t *gp, const int *gi, int *perm_out, int leaf);

#ifdef __cplusplus
}
#endif
#endif
