/*
 * oracle/simplex_oracle.c -- TEST INFRASTRUCTURE ONLY (CPU oracle, never shipped).
 *
 * Scalar C restatement of the reference's two-phase slack-form simplex
 * (/root/reference/week-2/diet/simplex.py).  It works on the CONDENSED tableau --
 * the (n+1) x (1+ncols) doubles that can ever be non-zero --
 * instead of the reference's nvars x nvars list-of-lists with None holes, and it
 * reproduces, bit for bit, every rounding, every snap-to-zero, the CPython-set
 * entering order, the degenerate-skip ratio test and the three crash classes.
 *
 * Build with -ffp-contract=off: the reference rounds the multiply and the add
 * separately (simplex.py:188).
 *
 * Function -> reference map
 *   snap()                 fp_ops.iszero + "= 0.0"        simplex.py:29-34, 191-193, 336, 344, 415
 *   do_pivot()             SlackForm.pivot                simplex.py:348-385
 *                          make_pivot_constraint          simplex.py:325-345
 *                          Constraint.substitute / sexp   simplex.py:185-214
 *   pick_entering()        pick_entering_idx              simplex.py:390-392 (+140-149)
 *   pick_leaving()         pick_leaving_idx / min_index   simplex.py:394-436, 157-161
 *   run_loop()             SlackForm.solve                simplex.py:442-491
 *   lp_oracle_simplex()    Simplex.solve                  simplex.py:586-630
 *                          find_basic_feasible            simplex.py:514-584
 *                          make_auxiliary_form            simplex.py:632-660
 *
 * rule = LP_RULE_BLAND is the "corrected" mode of SURVEY 8(f) rank 3 -- NOT reference behaviour,
 * no reference to pin it to ("parity unpinned" for this mode; tests check it against scipy
 * HiGHS instead).  Same tableau arithmetic (two roundings, snap), different decisions:
 *   - entering = LOWEST VARIABLE ID with c[j] > tol (Bland) instead of CPython-set order;
 *   - ratio test keeps zero ratios (the reference skips them, Appendix A.8), ties -> lowest id;
 *   - phase dispatch on the true minimum of b (zeros are zeros); min b >= 0 starts phase 2,
 *     so the AttributeError class (:626) is gone; None-tracking (TypeError :409) is off;
 *   - x_aux still basic (at level 0) after phase 1 is pivoted out on the lowest-id non-basic
 *     variable with a non-zero coefficient in its row instead of raising IndexError (:541);
 *     if the row is all zero it is a redundant constraint: it stays, inert, and no column is dropped.
 * rule = LP_RULE_DANTZIG is LP_RULE_BLAND with largest-coefficient pricing: entering = the column
 * with the largest c[j] > tol (ties -> lowest id); once more than LP_DEGENERATE_STREAK pivots in a
 * row (within one run of the loop) had a zero leaving ratio, Bland's entering choice is used until
 * a pivot with a non-zero ratio happens.
 * rule = LP_RULE_DEVEX is LP_RULE_DANTZIG priced by Devex reference-framework weights (Forrest & Goldfarb's
 * approximation of steepest edge): every non-basic position k carries a weight w[k]; entering = the column
 * with the largest (c*c)/w among c > tol (ties -> lowest id); a pivot with entering position p, new pivot row
 * nr[] (make_pivot_constraint, :325-345) and wq = w[p] sets w[k] = max(w[k], (nr[k]*nr[k])*wq) for k != p and
 * w[p] = max((nr[p]*nr[p])*wq, 1) (position p now holds the leaving variable).  Element-wise, no reduction:
 * the same bits whatever kernel, thread mapping or GPU count serves the LP.  Weights are reset to 1 (a new
 * reference framework) where a pivot loop is entered: before the forced phase-1 pivot, before an x_aux
 * drive-out pivot, and before phase 2.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "lp_oracle.h"

#define NONE_INITIAL (-1)

typedef struct {
    int n, m;
    int ncols;   /* non-basic positions currently alive (m, or m+1 during phase 1) */
    int stride;  /* doubles per tableau row = 1 + (m+1) */
    double *T;   /* rows 0..n-1: basic rows, row n: objective; col 0: constant */
    int *rowvar; /* [n]    variable id that owns basic row i */
    int *colvar; /* [m+1]  variable id at non-basic position k */
    int *varpos; /* [n+m+1] row index if basic, column position if non-basic */
    char *isbasic;
    /* None bookkeeping -- only to reproduce the TypeError of simplex.py:409.
     * In the reference each basic row's set of None entries is: all slack columns
     * (row never rewritten), or exactly {entering id of the last pivot that rewrote
     * the row} (substitute sets return_row[e+1] = None, simplex.py:209, 343); after
     * phase 1 the clean-up at simplex.py:551 adds every then-basic column. */
    int *none_col;   /* [n] NONE_INITIAL or a variable id */
    char *untouched; /* [n] row not rewritten since the phase-1 -> phase-2 clean-up */
    char *basis0;    /* [n+m+1] basic at the clean-up (simplex.py:551) */
    pyset_t dep, indep;
    double tol;
    int npiv, maxpiv;
    int rule;
    int streak;     /* LP_RULE_DANTZIG: degenerate pivots in a row in the current loop */
    int last_zero;  /* the ratio of the leaving row picked last was 0 */
    int32_t *log;
    int logcap;
    double *newrow;
    double *wts; /* LP_RULE_DEVEX: reference-framework weight of non-basic position k, [m+1] */
    int threads; /* 1 = scalar; otherwise the rows of one pivot are swept by this many host threads (0 = all cores) */
} lp_t;

static void devex_reset(lp_t *s)
{
    if (s->rule == LP_RULE_DEVEX)
        for (int k = 0; k <= s->m; ++k) s->wts[k] = 1.0;
}

/* test instrumentation: x_aux drive-out pivots performed since the last reset (not thread-safe) */
static int g_driveouts = 0;
int lp_oracle_driveouts(int reset)
{
    const int v = g_driveouts;
    if (reset) g_driveouts = 0;
    return v;
}

static inline double snap(double v, double tol) { return fabs(v) <= tol ? 0.0 : v; }

static void log_pivot(lp_t *s, int e, int l)
{
    if (s->log && s->npiv < s->logcap) {
        s->log[2 * s->npiv] = e;
        s->log[2 * s->npiv + 1] = l;
    }
    s->npiv++;
}

/* Constraint.substitute / sexp (simplex.py:185-214) applied to rows [lo, hi) of the tableau for the pivot
 * (entering position p, leaving row r, entering id e); newrow = the scaled pivot row. */
static void sweep_rows(lp_t *s, int p, int r, int e, int W, int lo, int hi)
{
    const int n = s->n, S = s->stride;
    const double tol = s->tol;
    const double *nr = s->newrow;
    for (int i = lo; i < hi; ++i) {
        if (i == r) continue;
        double *Ti = s->T + (size_t)i * S;
        const double f = Ti[1 + p];
        if (fabs(f) <= tol) { /* "not coeff or iszero(coeff)": row returned unchanged, :203-205 */
            Ti[1 + p] = 0.0;  /* A[i][l] stays None */
            continue;
        }
        for (int k = 0; k < W; ++k) {
            const double a = (k == 1 + p) ? 0.0 : Ti[k];
            const double prod = f * nr[k]; /* rounded multiply ... */
            Ti[k] = snap(a + prod, tol);   /* ... then rounded add, then snap (sexp, :185-194) */
        }
        if (i < n) {
            s->none_col[i] = e;
            s->untouched[i] = 0;
        }
    }
}

#define SWEEP_ROWS 32 /* rows per work item of the threaded sweep */
typedef struct {
    lp_t *s;
    int p, r, e, W;
} sweep_job_t;

static void sweep_job(void *arg, int item)
{
    sweep_job_t *j = (sweep_job_t *)arg;
    const int lo = item * SWEEP_ROWS;
    const int hi = lo + SWEEP_ROWS < j->s->n + 1 ? lo + SWEEP_ROWS : j->s->n + 1;
    sweep_rows(j->s, j->p, j->r, j->e, j->W, lo, hi);
}

/* SlackForm.pivot(entering e, leaving l), simplex.py:348-385 */
static void do_pivot(lp_t *s, int e, int l)
{
    const int n = s->n, S = s->stride, W = 1 + s->ncols;
    const double tol = s->tol;
    const int p = s->varpos[e]; /* column position of the entering variable */
    const int r = s->varpos[l]; /* row of the leaving variable */
    double *Tr = s->T + (size_t)r * S;
    double *nr = s->newrow;

    /* make_pivot_constraint, simplex.py:325-345 */
    const double piv = Tr[1 + p];
    double pinv = -1.0 * (1.0 / piv);
    pinv = snap(pinv, tol);
    for (int k = 0; k < W; ++k) nr[k] = snap(pinv * Tr[k], tol);
    nr[1 + p] = snap(-1.0 * pinv, tol); /* position p now belongs to l: coefficient 1/pivot */
    if (s->rule == LP_RULE_DEVEX) {
        const double wq = s->wts[p];
        for (int k = 0; k < s->ncols; ++k) {
            const double a = nr[1 + k];
            const double cand = (a * a) * wq;
            if (k == p) s->wts[k] = cand > 1.0 ? cand : 1.0;
            else if (cand > s->wts[k]) s->wts[k] = cand;
        }
    }

    /* substitute into every other basic row and into the objective (row n).  Rows are independent of each
     * other (each reads only itself and nr[]), so sweeping them on several host threads gives the same
     * bits as the scalar loop whatever the thread count (tests/test_oracle_big.py asserts it). */
    if (s->threads != 1 && (size_t)(n + 1) * (size_t)W >= ((size_t)1 << 16)) {
        sweep_job_t job = {s, p, r, e, W};
        lp_oracle_parallel_for((n + 1 + SWEEP_ROWS - 1) / SWEEP_ROWS, s->threads, sweep_job, &job);
    } else {
        sweep_rows(s, p, r, e, W, 0, n + 1);
    }
    memcpy(Tr, nr, sizeof(double) * (size_t)W); /* store(entering_idx, row), :352 */
    s->none_col[r] = e;
    s->untouched[r] = 0;

    s->rowvar[r] = e;
    s->colvar[p] = l;
    s->varpos[e] = r;
    s->varpos[l] = p;
    s->isbasic[e] = 1;
    s->isbasic[l] = 0;

    /* set_replace(dependent, rem=l, add=e); set_replace(independent, rem=e, add=l), :369-381 */
    pyset_remove(&s->dep, l);
    pyset_add(&s->dep, e);
    pyset_remove(&s->indep, e);
    pyset_add(&s->indep, l);
    log_pivot(s, e, l);
}

/* first j, in iteration order of the `independent` set, with c[j] > tol (simplex.py:390-392) */
static int pick_entering(const lp_t *s)
{
    const double *c = s->T + (size_t)s->n * s->stride;
    if (s->rule == LP_RULE_DEVEX && s->streak <= LP_DEGENERATE_STREAK) {
        int best = -1;
        double bestd = 0.0;
        for (int k = 0; k < s->ncols; ++k) {
            const double v = c[1 + k];
            if (!(v > s->tol)) continue;
            const double d = (v * v) / s->wts[k];
            if (best < 0 || d > bestd || (d == bestd && s->colvar[k] < best)) {
                best = s->colvar[k];
                bestd = d;
            }
        }
        return best;
    }
    if (s->rule == LP_RULE_DANTZIG && s->streak <= LP_DEGENERATE_STREAK) {
        int best = -1;
        double bestv = 0.0;
        for (int k = 0; k < s->ncols; ++k) {
            const double v = c[1 + k];
            if (v > s->tol && (best < 0 || v > bestv || (v == bestv && s->colvar[k] < best))) {
                best = s->colvar[k];
                bestv = v;
            }
        }
        return best;
    }
    if (s->rule != LP_RULE_REFERENCE) {
        int best = -1;
        for (int k = 0; k < s->ncols; ++k)
            if (c[1 + k] > s->tol && (best < 0 || s->colvar[k] < best)) best = s->colvar[k];
        return best;
    }
    for (int i = 0; i < s->indep.size; ++i) {
        const int v = s->indep.table[i];
        if (v >= 0 && c[1 + s->varpos[v]] > s->tol) return v;
    }
    return -1;
}

/* pick_leaving_idx, simplex.py:394-436.  Returns variable id, -1 = unbounded, -2 = TypeError. */
static int pick_leaving(lp_t *s, int e)
{
    const int n = s->n, S = s->stride, p = s->varpos[e];
    const int bland = (s->rule != LP_RULE_REFERENCE); /* the corrected ratio test (both corrected rules) */
    for (int i = 0; i < n && !bland; ++i) { /* -1 * None at :409 */
        if ((s->none_col[i] == NONE_INITIAL && e < n) || s->none_col[i] == e ||
            (s->untouched[i] && s->basis0[e]))
            return -2;
    }
    int best = -1;
    double bestr = 0.0;
    for (int i = 0; i < n; ++i) {
        const double *Ti = s->T + (size_t)i * S;
        const double coeff = -1.0 * Ti[1 + p];
        if (!(coeff > s->tol)) continue;                 /* ispositive(coeff), :413 */
        const double ratio = snap(Ti[0] / coeff, s->tol); /* :414-415 */
        if (ratio == 0.0 && !bland) continue;            /* falsy -> key +inf in min_index, :161 */
        const int v = s->rowvar[i];
        /* min() over ascending variable id keeps the first minimum */
        if (best < 0 || ratio < bestr || (ratio == bestr && v < best)) {
            best = v;
            bestr = ratio;
        }
    }
    s->last_zero = (best >= 0 && bestr == 0.0);
    return best;
}

/* SlackForm.solve loop, simplex.py:442-466.  Returns 0 when no positive coefficient is left. */
static int run_loop(lp_t *s)
{
    s->streak = 0;
    for (;;) {
        const int e = pick_entering(s);
        if (e < 0) return 0;
        const int l = pick_leaving(s, e);
        if (l == -2) return LP_STATUS_REF_TYPE_ERROR;
        if (l < 0) return LP_STATUS_INFINITY;
        s->streak = s->last_zero ? s->streak + 1 : 0;
        if (s->npiv >= s->maxpiv) return LP_STATUS_ITERATION_LIMIT;
        do_pivot(s, e, l);
    }
}

int lp_oracle_simplex(int n, int m, const double *A, const double *b, const double *c, double tol,
                      int max_pivots, double *x, double *objective, int32_t *npivots,
                      int32_t *pivot_log, int log_cap)
{
    return lp_oracle_simplex_rule(n, m, A, b, c, tol, max_pivots, LP_RULE_REFERENCE, x, objective, npivots,
                                  pivot_log, log_cap);
}

int lp_oracle_simplex_rule(int n, int m, const double *A, const double *b, const double *c, double tol,
                           int max_pivots, int rule, double *x, double *objective, int32_t *npivots,
                           int32_t *pivot_log, int log_cap)
{
    return lp_oracle_simplex_state(n, m, A, b, c, tol, max_pivots, rule, 1, x, objective, npivots, pivot_log,
                                   log_cap, NULL, NULL);
}

/* The same solve with (a) the rows of every pivot swept by `threads` host threads (1 = scalar, 0 = all
 * cores; bit-identical for any count, rows are independent) so that a config-5-size tableau
 * (16384 x 32768, 8.6 GB moved per pivot) is checkable in seconds per pivot, and (b) an export of the
 * tableau the solve stopped on, indexed by VARIABLE id so that it does not depend on how an engine lays
 * its columns out (both nullable, nv = n + m + 1):
 *   obj_by_var[v]    objective-row coefficient of v if v is non-basic, else 0;  obj_by_var[nv] = the
 *                    objective row's constant (the value v of simplex.py:480)
 *   const_by_var[v]  constant-column entry of v's row if v is basic, else 0
 * With max_pivots = K this is the state after the first K pivots (status LP_STATUS_ITERATION_LIMIT). */
int lp_oracle_simplex_state(int n, int m, const double *A, const double *b, const double *c, double tol,
                            int max_pivots, int rule, int threads, double *x, double *objective,
                            int32_t *npivots, int32_t *pivot_log, int log_cap, double *obj_by_var,
                            double *const_by_var)
{
    lp_t s;
    memset(&s, 0, sizeof(s));
    s.threads = threads;
    const int nv = n + m + 1;
    int status = LP_STATUS_BOUNDED;
    s.n = n;
    s.m = m;
    s.stride = m + 2;
    s.tol = tol;
    s.maxpiv = max_pivots;
    s.rule = rule;
    s.log = pivot_log;
    s.logcap = log_cap;
    s.T = (double *)calloc((size_t)(n + 1) * (size_t)s.stride, sizeof(double));
    s.newrow = (double *)calloc((size_t)s.stride, sizeof(double));
    s.rowvar = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    s.colvar = (int *)malloc(sizeof(int) * (size_t)(m + 1));
    s.varpos = (int *)malloc(sizeof(int) * (size_t)nv);
    s.isbasic = (char *)calloc((size_t)nv, 1);
    s.wts = (double *)calloc((size_t)(m + 2), sizeof(double));
    s.none_col = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    s.untouched = (char *)calloc((size_t)(n > 0 ? n : 1), 1);
    s.basis0 = (char *)calloc((size_t)nv, 1);
    for (int j = 0; j < m; ++j) x[j] = 0.0;
    *objective = 0.0;

    /* min_constant, idx = min_index(b): zero keys count as +inf, first minimum wins (:591, :157-161) */
    int idx = 0;
    {
        double bestkey = INFINITY;
        for (int i = 0; i < n; ++i) {
            const double key = (b[i] != 0.0 || rule != LP_RULE_REFERENCE) ? b[i] : INFINITY;
            if (key < bestkey) {
                bestkey = key;
                idx = i;
            }
        }
    }
    const double min_c = b[idx];
    const int phase1 = (min_c < 0.0);
    if (!phase1 && !(min_c > tol) && rule == LP_RULE_REFERENCE) { /* self.assignment undefined -> AttributeError, :626 */
        status = LP_STATUS_REF_ATTRIBUTE_ERROR;
        goto done_noset;
    }

    /* SlackForm.__init__, simplex.py:230-302 (plus the x_aux column of :632-660 in phase 1) */
    s.ncols = phase1 ? m + 1 : m;
    for (int i = 0; i < n; ++i) {
        double *Ti = s.T + (size_t)i * s.stride;
        Ti[0] = b[i];
        for (int j = 0; j < m; ++j) Ti[1 + j] = -A[(size_t)i * m + j];
        if (phase1) Ti[1 + m] = 1.0; /* -(-1) */
        s.rowvar[i] = i;
        s.varpos[i] = i;
        s.isbasic[i] = 1;
        s.none_col[i] = NONE_INITIAL;
    }
    for (int k = 0; k < s.ncols; ++k) {
        s.colvar[k] = n + k;
        s.varpos[n + k] = k;
    }
    {
        double *obj = s.T + (size_t)n * s.stride;
        obj[0] = 0.0;
        if (phase1) {
            for (int j = 0; j < m; ++j) obj[1 + j] = 0.0;
            obj[1 + m] = -1.0;
        } else {
            for (int j = 0; j < m; ++j) obj[1 + j] = c[j];
        }
    }
    pyset_init_range(&s.dep, 0, n);
    pyset_init_range(&s.indep, n, n + s.ncols);

    if (phase1) {
        const int aux = n + m;
        if (s.npiv >= s.maxpiv) { status = LP_STATUS_ITERATION_LIMIT; goto done; }
        devex_reset(&s);
        do_pivot(&s, aux, idx); /* aux_sf.pivot(n + m, min_idx), :524 */
        status = run_loop(&s);  /* :526 */
        if (status != 0) goto done;
        if (!(fabs(s.T[(size_t)n * s.stride]) <= tol)) { /* not iszero(opt) -> InfeasibleError, :528-530 */
            status = LP_STATUS_NO_SOLUTION;
            goto done;
        }
        if (s.isbasic[aux] && rule != LP_RULE_REFERENCE) { /* drive x_aux out of the basis */
            const double *Tr = s.T + (size_t)s.varpos[aux] * s.stride;
            int ev = -1;
            for (int k = 0; k < s.ncols; ++k)
                if (fabs(Tr[1 + k]) > tol && (ev < 0 || s.colvar[k] < ev)) ev = s.colvar[k];
            if (ev >= 0) {
                if (s.npiv >= s.maxpiv) { status = LP_STATUS_ITERATION_LIMIT; goto done; }
                devex_reset(&s);
                do_pivot(&s, ev, aux);
                ++g_driveouts;
                status = run_loop(&s); /* the objective is -x_aux again: ends at once */
                if (status != 0) goto done;
            }
        } else if (s.isbasic[aux]) { /* new_obj.get_coefficient(n+m) -> IndexError, :541 -> :183 */
            status = LP_STATUS_REF_INDEX_ERROR;
            goto done;
        }
        /* rebuild the original objective by substituting the basic rows in
         * iteration order of the `dependent` set, :537-541 */
        double *o = (double *)calloc((size_t)nv, sizeof(double));
        double o0 = 0.0;
        for (int j = 0; j < m; ++j) o[n + j] = c[j];
        for (int t = 0; t < s.dep.size; ++t) {
            const int v = s.dep.table[t];
            if (v < 0) continue;
            const double f = o[v];
            if (fabs(f) <= tol) continue;
            const double *Tr = s.T + (size_t)s.varpos[v] * s.stride;
            o0 = snap(o0 + f * Tr[0], tol);
            for (int k = 0; k < s.ncols; ++k) {
                const int w = s.colvar[k];
                if (w == aux) continue; /* zip() truncates the x_aux column away, :190 */
                const double prod = f * Tr[1 + k];
                o[w] = snap(o[w] + prod, tol);
            }
            for (int i = 0; i < n; ++i) { /* basic columns: a + f*0, then snapped */
                const int w = s.rowvar[i];
                if (w != v) o[w] = snap(o[w], tol);
            }
            o[v] = 0.0; /* return_row[idx+1] = None, :209 */
        }
        /* clean-up, :547-575 */
        for (int v = 0; v < nv; ++v) s.basis0[v] = s.isbasic[v];
        for (int i = 0; i < n; ++i) s.untouched[i] = 1;
        if (!s.isbasic[aux]) {
            const int kaux = s.varpos[aux], last = s.ncols - 1;
            if (kaux != last) {
                for (int i = 0; i <= n; ++i)
                    s.T[(size_t)i * s.stride + 1 + kaux] = s.T[(size_t)i * s.stride + 1 + last];
                s.colvar[kaux] = s.colvar[last];
                s.varpos[s.colvar[kaux]] = kaux;
            }
            s.ncols = m;
        }
        {
            double *obj = s.T + (size_t)n * s.stride;
            obj[0] = o0;
            for (int k = 0; k < s.ncols; ++k) obj[1 + k] = o[s.colvar[k]];
        }
        free(o);
        if (!s.isbasic[aux]) pyset_remove(&s.indep, aux); /* :574 -- the same set objects live on into phase 2 */
    }

    devex_reset(&s);
    status = run_loop(&s); /* :599 / :619 */
    if (status == 0) {
        for (int j = 0; j < m; ++j) /* assignment, :474-478, 603 */
            x[j] = s.isbasic[n + j] ? s.T[(size_t)s.varpos[n + j] * s.stride] : 0.0;
        *objective = s.T[(size_t)n * s.stride]; /* opt_value = v, :480 */
    }
done:
    if (obj_by_var) {
        const double *obj = s.T + (size_t)n * s.stride;
        for (int v = 0; v <= nv; ++v) obj_by_var[v] = 0.0;
        for (int k = 0; k < s.ncols; ++k) obj_by_var[s.colvar[k]] = obj[1 + k];
        obj_by_var[nv] = obj[0];
    }
    if (const_by_var) {
        for (int v = 0; v < nv; ++v) const_by_var[v] = 0.0;
        for (int i = 0; i < n; ++i) const_by_var[s.rowvar[i]] = s.T[(size_t)i * s.stride];
    }
    pyset_free(&s.dep);
    pyset_free(&s.indep);
    goto release;
done_noset:
    if (obj_by_var) for (int v = 0; v <= nv; ++v) obj_by_var[v] = 0.0;
    if (const_by_var) for (int v = 0; v < nv; ++v) const_by_var[v] = 0.0;
release:
    *npivots = s.npiv;
    free(s.T); free(s.newrow); free(s.rowvar); free(s.colvar); free(s.varpos);
    free(s.wts);
    free(s.isbasic); free(s.none_col); free(s.untouched); free(s.basis0);
    return status;
}

/* Simplex.verify_bounds, simplex.py:663-671: sum = 0; sum += row[i] * ansx[i] left to right (product and sum
 * rounded separately); assert(sum <= b[r] + 1e-3) -- the reference ignores its tolerance argument, the caller
 * passes 1e-3 to reproduce it.  Returns the first row whose assertion fires, -1 when the answer passes. */
int lp_oracle_verify_bounds(int n, int m, const double *A, const double *b, const double *x, double tolerance)
{
    for (int r = 0; r < n; ++r) {
        double sum = 0.0;
        for (int i = 0; i < m; ++i) {
            const double prod = A[(size_t)r * m + i] * x[i];
            sum = sum + prod;
        }
        if (!(sum <= b[r] + tolerance)) return r;
    }
    return -1;
}

/* ---- batch driver for the timed CPU baseline: pthreads pulling problems off a shared counter ---- */
typedef struct {
    int B, n, m;
    const double *A, *b, *c;
    double tol;
    int max_pivots, rule;
    int32_t *status;
    double *x, *objective;
    int32_t *npivots;
} simplex_job_t;

static void simplex_job(void *arg, int p)
{
    simplex_job_t *j = (simplex_job_t *)arg;
    const int n = j->n, m = j->m;
    j->status[p] = lp_oracle_simplex_rule(n, m, j->A + (size_t)p * n * m, j->b + (size_t)p * n,
                                          j->c + (size_t)p * m, j->tol, j->max_pivots, j->rule,
                                          j->x + (size_t)p * m, j->objective + p, j->npivots + p, NULL, 0);
}

int lp_oracle_simplex_batch(int B, int n, int m, const double *A, const double *b, const double *c,
                            double tol, int max_pivots, int32_t *status, double *x, double *objective,
                            int32_t *npivots, int threads)
{
    return lp_oracle_simplex_batch_rule(B, n, m, A, b, c, tol, max_pivots, LP_RULE_REFERENCE, status, x,
                                        objective, npivots, threads);
}

int lp_oracle_simplex_batch_rule(int B, int n, int m, const double *A, const double *b, const double *c,
                                 double tol, int max_pivots, int rule, int32_t *status, double *x,
                                 double *objective, int32_t *npivots, int threads)
{
    simplex_job_t j = {B, n, m, A, b, c, tol, max_pivots, rule, status, x, objective, npivots};
    return lp_oracle_parallel_for(B, threads, simplex_job, &j);
}
