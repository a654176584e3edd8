/* qpdo_io.c -- free-format QPS reader into the reference's QPDOData (include/qpdo_io.h).
 *
 * Host-only C.  The reference has no reader; the output is exactly what a C caller
 * of qpdo_setup would build by hand (reference src/qpdo.c:49-110 deep-copies it): Q lower CSC with
 * stype -1, A CSC with sorted row indices, l <= A x <= u with variable bounds as identity rows.
 */
#define _POSIX_C_SOURCE 200809L
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "qpdo_io.h"

/* ------------------------------------------------------------------ name table (open addressing) */
typedef struct { char **name; long *slot; long count, cap, nslot; } names;

static unsigned long hash_str(const char *s)
{
    unsigned long h = 1469598103934665603UL;
    for (; *s; s++) { h ^= (unsigned char)*s; h *= 1099511628211UL; }
    return h;
}
static int names_grow(names *t)
{
    long nslot = t->nslot ? 2 * t->nslot : 1024;
    long *slot = malloc(sizeof(long) * nslot);
    if (!slot) return -1;
    for (long k = 0; k < nslot; k++) slot[k] = -1;
    for (long k = 0; k < t->count; k++) {
        unsigned long h = hash_str(t->name[k]) & (nslot - 1);
        while (slot[h] >= 0) h = (h + 1) & (nslot - 1);
        slot[h] = k;
    }
    free(t->slot);
    t->slot = slot; t->nslot = nslot;
    return 0;
}
static long names_find(const names *t, const char *s)
{
    if (!t->nslot) return -1;
    unsigned long h = hash_str(s) & (t->nslot - 1);
    while (t->slot[h] >= 0) {
        if (strcmp(t->name[t->slot[h]], s) == 0) return t->slot[h];
        h = (h + 1) & (t->nslot - 1);
    }
    return -1;
}
static long names_add(names *t, const char *s)        /* index of s, inserted when new; -2 on allocation failure */
{
    long k = names_find(t, s);
    if (k >= 0) return k;
    if (2 * (t->count + 1) > t->nslot && names_grow(t)) return -2;
    if (t->count == t->cap) {
        long cap = t->cap ? 2 * t->cap : 256;
        char **nm = realloc(t->name, sizeof(char *) * cap);
        if (!nm) return -2;
        t->name = nm; t->cap = cap;
    }
    if (!(t->name[t->count] = strdup(s))) return -2;
    unsigned long h = hash_str(s) & (t->nslot - 1);
    while (t->slot[h] >= 0) h = (h + 1) & (t->nslot - 1);
    t->slot[h] = t->count;
    return t->count++;
}
static void names_free(names *t)
{
    for (long k = 0; k < t->count; k++) free(t->name[k]);
    free(t->name); free(t->slot);
    memset(t, 0, sizeof(*t));
}

/* ------------------------------------------------------------------ triplets -> CSC */
typedef struct { long *r, *c; double *v; long count, cap; } triplets;

static int trip_push(triplets *t, long r, long c, double v)
{
    if (t->count == t->cap) {
        long cap = t->cap ? 2 * t->cap : 1024;
        long *nr = realloc(t->r, sizeof(long) * cap);   if (!nr) return -1; t->r = nr;
        long *nc = realloc(t->c, sizeof(long) * cap);   if (!nc) return -1; t->c = nc;
        double *nv = realloc(t->v, sizeof(double) * cap); if (!nv) return -1; t->v = nv;
        t->cap = cap;
    }
    t->r[t->count] = r; t->c[t->count] = c; t->v[t->count] = v; t->count++;
    return 0;
}
static void trip_free(triplets *t) { free(t->r); free(t->c); free(t->v); memset(t, 0, sizeof(*t)); }

/* CSC with sorted rows, duplicates summed (what cholmod_triplet_to_sparse does) */
static cholmod_sparse *to_csc(const triplets *t, long nrow, long ncol, int stype)
{
    cholmod_sparse *S = calloc(1, sizeof(*S));
    long *p = calloc((size_t)ncol + 1, sizeof(long));
    long *ri = malloc(sizeof(long) * (size_t)(t->count ? t->count : 1));
    double *x = malloc(sizeof(double) * (size_t)(t->count ? t->count : 1));
    long *next = malloc(sizeof(long) * (size_t)(ncol ? ncol : 1));
    if (!S || !p || !ri || !x || !next) { free(S); free(p); free(ri); free(x); free(next); return NULL; }
    for (long k = 0; k < t->count; k++) p[t->c[k] + 1]++;
    for (long j = 0; j < ncol; j++) { p[j + 1] += p[j]; next[j] = p[j]; }
    for (long k = 0; k < t->count; k++) { long d = next[t->c[k]]++; ri[d] = t->r[k]; x[d] = t->v[k]; }
    long out = 0;
    for (long j = 0; j < ncol; j++) {
        const long b = p[j], e = p[j + 1];
        for (long a = b + 1; a < e; a++) {              /* insertion sort by row: columns are short */
            long r = ri[a]; double v = x[a]; long q = a - 1;
            while (q >= b && ri[q] > r) { ri[q + 1] = ri[q]; x[q + 1] = x[q]; q--; }
            ri[q + 1] = r; x[q + 1] = v;
        }
        const long start = out;
        for (long a = b; a < e; a++) {
            if (out > start && ri[out - 1] == ri[a]) x[out - 1] += x[a];
            else { ri[out] = ri[a]; x[out] = x[a]; out++; }
        }
        p[j] = start;
    }
    p[ncol] = out;
    free(next);
    S->nrow = (size_t)nrow; S->ncol = (size_t)ncol; S->nzmax = (size_t)(out ? out : 1);
    S->p = p; S->i = ri; S->x = x; S->nz = NULL; S->z = NULL;
    S->stype = stype; S->itype = 2; S->xtype = 1; S->dtype = 0; S->sorted = 1; S->packed = 1;
    return S;
}
static void free_sparse(cholmod_sparse *S) { if (S) { free(S->p); free(S->i); free(S->x); free(S); } }

void qpdo_free_data(QPDOData *d)
{
    if (!d) return;
    free_sparse(d->Q); free_sparse(d->A);
    free(d->q); free(d->l); free(d->u);
    free(d);
}

/* ------------------------------------------------------------------ the reader */
enum { S_NONE, S_ROWS, S_COLUMNS, S_RHS, S_RANGES, S_BOUNDS, S_QUADOBJ, S_QMATRIX, S_OBJSENSE, S_SKIP };

static int split(char *line, char **tok, int maxtok)
{
    int n = 0;
    char *save = NULL;
    for (char *t = strtok_r(line, " \t\r\n", &save); t && n < maxtok; t = strtok_r(NULL, " \t\r\n", &save)) tok[n++] = t;
    return n;
}

int qpdo_read_qps(const char *path, QPDOData **out, c_int *m_rows, c_int *err_line)
{
    if (out) *out = NULL;
    if (err_line) *err_line = 0;
    FILE *fp = path ? fopen(path, "r") : NULL;
    if (!fp || !out) { if (fp) fclose(fp); return -1; }

    names rows = {0}, cols = {0};
    triplets At = {0}, Qt = {0};
    char *rtype = NULL;           /* per row: 'N' 'E' 'L' 'G' */
    long *rcon = NULL;            /* per row: constraint index or -1 */
    double *rhs = NULL, *rng = NULL, *lo = NULL, *up = NULL;
    char *has_rng = NULL;
    long rcap = 0, ccap = 0, m0 = 0, objrow = -1, lineno = 0;
    double cconst = 0.0;
    int section = S_NONE, rc = 0;
    char *line = NULL;
    size_t lcap = 0;

#define FAIL(code) do { rc = (code); goto done; } while (0)
    while (getline(&line, &lcap, fp) >= 0) {
        lineno++;
        if (line[0] == '*' || line[0] == '\n' || line[0] == '\r' || line[0] == '\0') continue;
        const int header = !(line[0] == ' ' || line[0] == '\t');
        char *tok[8];
        const int nt = split(line, tok, 8);
        if (nt == 0) continue;
        if (header) {
            if (!strcmp(tok[0], "NAME")) section = S_NONE;
            else if (!strcmp(tok[0], "ROWS")) section = S_ROWS;
            else if (!strcmp(tok[0], "COLUMNS")) section = S_COLUMNS;
            else if (!strcmp(tok[0], "RHS")) section = S_RHS;
            else if (!strcmp(tok[0], "RANGES")) section = S_RANGES;
            else if (!strcmp(tok[0], "BOUNDS")) section = S_BOUNDS;
            else if (!strcmp(tok[0], "QUADOBJ")) section = S_QUADOBJ;
            else if (!strcmp(tok[0], "QMATRIX")) section = S_QMATRIX;
            else if (!strcmp(tok[0], "ENDATA")) break;
            else if (!strcmp(tok[0], "OBJSENSE")) {
                /* "OBJSENSE MAX" on the header line, or MAX / MIN on the next data line: QPDO minimises; a maximisation
                 * file is rejected with its line number instead of being solved as a minimisation */
                section = S_OBJSENSE;
                if (nt > 1 && (!strcmp(tok[1], "MAX") || !strcmp(tok[1], "MAXIMIZE"))) FAIL(-2);
            }
            else section = S_SKIP;                    /* OBJSENSE_MAXIMIZE-style extensions of other dialects etc.: skipped */
            continue;
        }
        switch (section) {
        case S_OBJSENSE:
            if (!strcmp(tok[0], "MAX") || !strcmp(tok[0], "MAXIMIZE")) FAIL(-2);
            break;
        case S_ROWS: {
            if (nt < 2) FAIL(-2);
            const long before = rows.count, r = names_add(&rows, tok[1]);
            if (r == -2) FAIL(-3);
            if (r < before) FAIL(-2);                 /* duplicate row name */
            if (rows.count > rcap) {
                long cap = rcap ? 2 * rcap : 256;
                char *a = realloc(rtype, (size_t)cap); if (!a) FAIL(-3); rtype = a;
                long *b = realloc(rcon, sizeof(long) * cap); if (!b) FAIL(-3); rcon = b;
                rcap = cap;
            }
            const char ty = tok[0][0];
            if (ty != 'N' && ty != 'E' && ty != 'L' && ty != 'G') FAIL(-2);
            rtype[r] = ty;
            if (ty == 'N') { rcon[r] = -1; if (objrow < 0) objrow = r; }
            else rcon[r] = m0++;
            break;
        }
        case S_COLUMNS: {
            if (nt >= 3 && !strcmp(tok[1], "'MARKER'")) break;
            if (nt != 3 && nt != 5) FAIL(-2);
            const long ncol_before = cols.count, j = names_add(&cols, tok[0]);
            if (j == -2) FAIL(-3);
            if (cols.count > ccap) {
                long cap = ccap ? 2 * ccap : 256;
                double *b = realloc(lo, sizeof(double) * cap); if (!b) FAIL(-3); lo = b;
                double *c = realloc(up, sizeof(double) * cap); if (!c) FAIL(-3); up = c;
                ccap = cap;
            }
            if (j >= ncol_before) { lo[j] = 0.0; up[j] = QPDO_INFTY; }       /* QPS default bounds [0, +inf) */
            for (int k = 1; k + 1 < nt; k += 2) {
                const long r = names_find(&rows, tok[k]);
                if (r < 0) FAIL(-2);
                const double v = strtod(tok[k + 1], NULL);
                if (rcon[r] >= 0) { if (trip_push(&At, rcon[r], j, v)) FAIL(-3); }
                else if (r == objrow) { if (trip_push(&Qt, -1, j, v)) FAIL(-3); }   /* linear cost, parked with row -1 */
            }
            break;
        }
        case S_RHS: case S_RANGES: {
            if (!rhs) {
                rhs = calloc((size_t)(m0 ? m0 : 1), sizeof(double)); rng = calloc((size_t)(m0 ? m0 : 1), sizeof(double));
                has_rng = calloc((size_t)(m0 ? m0 : 1), 1);
                if (!rhs || !rng || !has_rng) FAIL(-3);
            }
            const int first = (nt % 2) ? 1 : 0;       /* odd token count: the set name leads */
            if (nt < 2) FAIL(-2);
            for (int k = first; k + 1 < nt; k += 2) {
                const long r = names_find(&rows, tok[k]);
                if (r < 0) FAIL(-2);
                const double v = strtod(tok[k + 1], NULL);
                if (rcon[r] < 0) { if (section == S_RHS && r == objrow) cconst = -v; continue; }
                if (section == S_RHS) rhs[rcon[r]] = v; else { rng[rcon[r]] = v; has_rng[rcon[r]] = 1; }
            }
            break;
        }
        case S_BOUNDS: {
            /* type [set] column [value] */
            const char *ty = tok[0];
            const int noval = !strcmp(ty, "FR") || !strcmp(ty, "MI") || !strcmp(ty, "PL") || !strcmp(ty, "BV");
            const char *cname; double v = 0.0;
            if (noval) { if (nt == 2) cname = tok[1]; else if (nt >= 3) cname = tok[2]; else FAIL(-2); }
            else { if (nt == 3) { cname = tok[1]; v = strtod(tok[2], NULL); } else if (nt >= 4) { cname = tok[2]; v = strtod(tok[3], NULL); } else FAIL(-2); }
            const long j = names_find(&cols, cname);
            if (j < 0) FAIL(-2);
            if (!strcmp(ty, "UP") || !strcmp(ty, "UI")) { up[j] = v; if (v < 0.0 && lo[j] == 0.0) lo[j] = -QPDO_INFTY; }
            else if (!strcmp(ty, "LO") || !strcmp(ty, "LI")) lo[j] = v;
            else if (!strcmp(ty, "FX")) { lo[j] = v; up[j] = v; }
            else if (!strcmp(ty, "FR")) { lo[j] = -QPDO_INFTY; up[j] = QPDO_INFTY; }
            else if (!strcmp(ty, "MI")) lo[j] = -QPDO_INFTY;
            else if (!strcmp(ty, "PL")) up[j] = QPDO_INFTY;
            else if (!strcmp(ty, "BV")) { lo[j] = 0.0; up[j] = 1.0; }
            else FAIL(-2);
            break;
        }
        case S_QUADOBJ: case S_QMATRIX: {
            if (nt != 3) FAIL(-2);
            const long a = names_find(&cols, tok[0]), b = names_find(&cols, tok[1]);
            if (a < 0 || b < 0) FAIL(-2);
            const double v = strtod(tok[2], NULL);
            if (section == S_QMATRIX && a < b) break;             /* both triangles listed: keep the lower one */
            if (trip_push(&Qt, a > b ? a : b, a > b ? b : a, v)) FAIL(-3);
            break;
        }
        default: break;
        }
    }
    goto build;

build: {
        const long n = cols.count;
        double *q = calloc((size_t)(n ? n : 1), sizeof(double));
        if (!q) FAIL(-3);
        /* split the parked cost entries (row -1) from the quadratic ones */
        triplets Qq = {0};
        for (long k = 0; k < Qt.count; k++) {
            if (Qt.r[k] < 0) q[Qt.c[k]] += Qt.v[k];
            else if (trip_push(&Qq, Qt.r[k], Qt.c[k], Qt.v[k])) { free(q); trip_free(&Qq); FAIL(-3); }
        }
        if (!rhs) { rhs = calloc((size_t)(m0 ? m0 : 1), sizeof(double)); rng = calloc((size_t)(m0 ? m0 : 1), sizeof(double)); has_rng = calloc((size_t)(m0 ? m0 : 1), 1); }
        if (!rhs || !rng || !has_rng) { free(q); trip_free(&Qq); FAIL(-3); }
        /* variable bounds -> identity rows */
        long nb = 0;
        for (long j = 0; j < n; j++) if (lo[j] > -QPDO_INFTY || up[j] < QPDO_INFTY) nb++;
        const long m = m0 + nb;
        double *l = malloc(sizeof(double) * (size_t)(m ? m : 1)), *u = malloc(sizeof(double) * (size_t)(m ? m : 1));
        if (!l || !u) { free(q); free(l); free(u); trip_free(&Qq); FAIL(-3); }
        for (long r = 0; r < rows.count; r++) {
            const long i = rcon[r];
            if (i < 0) continue;
            const double b = rhs[i], R = rng[i];
            switch (rtype[r]) {
            case 'E': l[i] = b; u[i] = b; if (has_rng[i]) { if (R < 0) l[i] = b + R; else u[i] = b + R; } break;
            case 'L': u[i] = b; l[i] = has_rng[i] ? b - fabs(R) : -QPDO_INFTY; break;
            default:  l[i] = b; u[i] = has_rng[i] ? b + fabs(R) : QPDO_INFTY; break;       /* 'G' */
            }
        }
        long i = m0;
        int bad = 0;
        for (long j = 0; j < n && !bad; j++)
            if (lo[j] > -QPDO_INFTY || up[j] < QPDO_INFTY) {
                bad = trip_push(&At, i, j, 1.0);
                l[i] = lo[j] > -QPDO_INFTY ? lo[j] : -QPDO_INFTY; u[i] = up[j] < QPDO_INFTY ? up[j] : QPDO_INFTY; i++;
            }
        QPDOData *d = bad ? NULL : calloc(1, sizeof(*d));
        if (!d) { free(q); free(l); free(u); trip_free(&Qq); FAIL(-3); }
        d->n = (size_t)n; d->m = (size_t)m; d->q = q; d->c = cconst; d->l = l; d->u = u;
        d->Q = to_csc(&Qq, n, n, -1);
        d->A = to_csc(&At, m, n, 0);
        trip_free(&Qq);
        if (!d->Q || !d->A) { qpdo_free_data(d); FAIL(-3); }
        *out = d;
        if (m_rows) *m_rows = (c_int)m0;
    }
done:
    if (rc == -2 && err_line) *err_line = (c_int)lineno;
    free(line); fclose(fp);
    names_free(&rows); names_free(&cols); trip_free(&At); trip_free(&Qt);
    free(rtype); free(rcon); free(rhs); free(rng); free(has_rng); free(lo); free(up);
    return rc;
#undef FAIL
}

/* ------------------------------------------------------------------ MatrixMarket
 * The format CHOLMOD's own reader takes (reference SuiteSparse/CHOLMOD/Check/cholmod_read.c, not used
 * by QPDO itself).  Written from the format definition: banner "%%MatrixMarket matrix <coordinate|array>
 * <real|integer|pattern> <general|symmetric|skew-symmetric>", '%' comments, a size line, then 1-based
 * entries (coordinate) or column-major values (array). */
enum { MM_GENERAL, MM_SYMMETRIC, MM_SKEW };

static void lower_str(char *s) { for (; *s; s++) if (*s >= 'A' && *s <= 'Z') *s = (char)(*s - 'A' + 'a'); }

/* Reads the banner and the size line.  coordinate: dims = nrow ncol nnz; array: dims = nrow ncol. */
static int mm_header(FILE *fp, char **line, size_t *cap, long *lineno, int *is_array, int *is_pattern, int *sym, long dims[3])
{
    if (getline(line, cap, fp) < 0) return -2;
    ++*lineno;
    char *tok[8];
    const int nt = split(*line, tok, 8);
    if (nt < 5) return -2;
    for (int k = 0; k < 5; k++) lower_str(tok[k]);
    if (strcmp(tok[0], "%%matrixmarket") || strcmp(tok[1], "matrix")) return -2;
    if (!strcmp(tok[2], "coordinate")) *is_array = 0; else if (!strcmp(tok[2], "array")) *is_array = 1; else return -2;
    if (!strcmp(tok[3], "pattern")) *is_pattern = 1;
    else if (!strcmp(tok[3], "real") || !strcmp(tok[3], "integer") || !strcmp(tok[3], "double")) *is_pattern = 0;
    else return -2;                                                                  /* complex: not a QP */
    if (!strcmp(tok[4], "general")) *sym = MM_GENERAL; else if (!strcmp(tok[4], "symmetric")) *sym = MM_SYMMETRIC;
    else if (!strcmp(tok[4], "skew-symmetric")) *sym = MM_SKEW; else return -2;
    if (*is_array && *is_pattern) return -2;
    for (;;) {
        if (getline(line, cap, fp) < 0) return -2;
        ++*lineno;
        const char *s = *line + strspn(*line, " \t\r\n");
        if (*s == '%' || *s == '\0') continue;
        const int want = *is_array ? 2 : 3;
        char *end;
        for (int k = 0; k < want; k++) {
            dims[k] = strtol(s, &end, 10);
            if (end == s || dims[k] < 0) return -2;
            s = end;
        }
        if (s[strspn(s, " \t\r\n")] != '\0') return -2;
        return 0;
    }
}

int qpdo_read_mtx(const char *path, cholmod_sparse **out, c_int *err_line)
{
    if (out) *out = NULL;
    if (err_line) *err_line = 0;
    FILE *fp = path ? fopen(path, "r") : NULL;
    if (!fp || !out) { if (fp) fclose(fp); return -1; }
    char *line = NULL;
    size_t cap = 0;
    long lineno = 0, dims[3] = {0, 0, 0};
    int is_array = 0, is_pattern = 0, sym = MM_GENERAL;
    triplets T = {0};
    int rc = mm_header(fp, &line, &cap, &lineno, &is_array, &is_pattern, &sym, dims);
    const long nrow = dims[0], ncol = dims[1];
    if (rc == 0 && sym != MM_GENERAL && nrow != ncol) rc = -2;
    if (rc == 0 && !is_array) {
        long seen = 0;
        while (seen < dims[2]) {
            if (getline(&line, &cap, fp) < 0) { rc = -2; break; }
            lineno++;
            const char *s = line + strspn(line, " \t\r\n");
            if (*s == '%' || *s == '\0') continue;
            char *end;
            long i = strtol(s, &end, 10);  if (end == s) { rc = -2; break; }  s = end;
            long j = strtol(s, &end, 10);  if (end == s) { rc = -2; break; }  s = end;
            double v = 1.0;
            if (!is_pattern) { v = strtod(s, &end); if (end == s) { rc = -2; break; } }
            if (i < 1 || i > nrow || j < 1 || j > ncol) { rc = -2; break; }
            i--; j--;
            if (sym != MM_GENERAL && i < j) { long t = i; i = j; j = t; if (sym == MM_SKEW) v = -v; }   /* keep the lower triangle */
            if (sym == MM_SKEW && i == j) { rc = -2; break; }
            if (trip_push(&T, i, j, v)) { rc = -3; break; }
            if (sym == MM_SKEW && trip_push(&T, j, i, -v)) { rc = -3; break; }
            seen++;
        }
    } else if (rc == 0) {                                /* array: column-major; symmetric files hold the lower triangle */
        for (long j = 0; j < ncol && rc == 0; j++)
            for (long i = (sym == MM_GENERAL ? 0 : (sym == MM_SKEW ? j + 1 : j)); i < nrow && rc == 0; ) {
                if (getline(&line, &cap, fp) < 0) { rc = -2; break; }
                lineno++;
                const char *s = line + strspn(line, " \t\r\n");
                if (*s == '%' || *s == '\0') continue;
                char *end;
                for (;;) {                               /* several values on one line are accepted */
                    const double v = strtod(s, &end);
                    if (end == s) break;
                    s = end;
                    if (i >= nrow) { rc = -2; break; }
                    if (v != 0.0 && trip_push(&T, i, j, v)) { rc = -3; break; }
                    if (sym == MM_SKEW && v != 0.0 && trip_push(&T, j, i, -v)) { rc = -3; break; }
                    i++;
                }
                if (rc == 0 && s[strspn(s, " \t\r\n")] != '\0') rc = -2;
            }
    }
    if (rc == 0) {
        *out = to_csc(&T, nrow, ncol, sym == MM_SYMMETRIC ? -1 : 0);
        if (!*out) rc = -3;
    }
    if (rc == -2 && err_line) *err_line = (c_int)lineno;
    trip_free(&T); free(line); fclose(fp);
    return rc;
}

void qpdo_free_sparse(cholmod_sparse *S) { free_sparse(S); }

int qpdo_read_mtx_vector(const char *path, c_float **out, c_int *len, c_int *err_line)
{
    if (out) *out = NULL;
    if (len) *len = 0;
    cholmod_sparse *S = NULL;
    const int rc = out && len ? qpdo_read_mtx(path, &S, err_line) : -1;
    if (rc) return rc;
    if (S->ncol != 1 && S->nrow != 1) { free_sparse(S); return -2; }
    const long n = (long)(S->ncol == 1 ? S->nrow : S->ncol);
    double *v = calloc((size_t)(n ? n : 1), sizeof(double));
    if (!v) { free_sparse(S); return -3; }
    const long *p = S->p, *ri = S->i;
    const double *x = S->x;
    for (long j = 0; j < (long)S->ncol; j++)
        for (long a = p[j]; a < p[j + 1]; a++) v[S->ncol == 1 ? ri[a] : j] = x[a];
    free_sparse(S);
    *out = v; *len = (c_int)n;
    return 0;
}

/* Keeps the lower triangle of a general square matrix (a Q written without the symmetric tag). */
static cholmod_sparse *lower_of(const cholmod_sparse *S)
{
    triplets T = {0};
    const long *p = S->p, *ri = S->i;
    const double *x = S->x;
    for (long j = 0; j < (long)S->ncol; j++)
        for (long a = p[j]; a < p[j + 1]; a++)
            if (ri[a] >= j && trip_push(&T, ri[a], j, x[a])) { trip_free(&T); return NULL; }
    cholmod_sparse *L = to_csc(&T, (long)S->nrow, (long)S->ncol, -1);
    trip_free(&T);
    return L;
}

int qpdo_read_mtx_problem(const char *Q_path, const char *A_path, const char *q_path, const char *l_path,
                          const char *u_path, QPDOData **out, c_int *err_line)
{
    if (out) *out = NULL;
    if (!out || !A_path) return -1;
    QPDOData *d = calloc(1, sizeof(*d));
    if (!d) return -3;
    int rc = qpdo_read_mtx(A_path, &d->A, err_line);
    if (rc == 0 && d->A->stype != 0) rc = -2;            /* A is a general matrix */
    const long m = rc ? 0 : (long)d->A->nrow, n = rc ? 0 : (long)d->A->ncol;
    if (rc == 0 && Q_path) {
        rc = qpdo_read_mtx(Q_path, &d->Q, err_line);
        if (rc == 0 && ((long)d->Q->nrow != n || (long)d->Q->ncol != n)) rc = -4;
        if (rc == 0 && d->Q->stype == 0) {
            cholmod_sparse *L = lower_of(d->Q);
            free_sparse(d->Q); d->Q = L;
            if (!L) rc = -3;
        }
    } else if (rc == 0) {                                /* no Q: an LP, the empty lower triangle */
        triplets none = {0};
        d->Q = to_csc(&none, n, n, -1);
        if (!d->Q) rc = -3;
    }
    const char *vp[3] = { q_path, l_path, u_path };
    double **dst[3] = { &d->q, &d->l, &d->u };
    const double fill[3] = { 0.0, -QPDO_INFTY, QPDO_INFTY };
    for (int k = 0; k < 3 && rc == 0; k++) {
        const long want = k == 0 ? n : m;
        if (vp[k]) {
            c_int len = 0;
            rc = qpdo_read_mtx_vector(vp[k], dst[k], &len, err_line);
            if (rc == 0 && (long)len != want) rc = -4;
        } else {
            double *v = malloc(sizeof(double) * (size_t)(want ? want : 1));
            if (!v) { rc = -3; break; }
            for (long a = 0; a < want; a++) v[a] = fill[k];
            *dst[k] = v;
        }
    }
    if (rc == 0)
        for (long i = 0; i < m; i++) {                   /* +-inf spelled as huge numbers or "inf" */
            if (!(d->l[i] > -QPDO_INFTY)) d->l[i] = -QPDO_INFTY;
            if (!(d->u[i] < QPDO_INFTY)) d->u[i] = QPDO_INFTY;
        }
    if (rc) { qpdo_free_data(d); return rc; }
    d->n = (size_t)n; d->m = (size_t)m; d->c = 0.0;
    *out = d;
    return 0;
}
